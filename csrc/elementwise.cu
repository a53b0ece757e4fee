// Fused elementwise kernels: Keras-form Adam / Adagrad / SGD over the flat parameter space (K12/K13 of
// SURVEY.md §2.5) and the MNIST MinMax + reshape + one-hot preprocessing pass (K14).
// Update rules follow Keras 2 as named by the reference ('adam' /root/reference/ddl_mnist_aztk.py:195,
// 'adagrad' /root/reference/ddl_nyiso_aztk.py:207); the step counter lives in device memory so a captured
// CUDA graph replays the right bias correction.
#include "common.cuh"
#include "shadow.cuh"
#include "../include/ddls_b200.h"

namespace ddls {

constexpr int EW_THREADS = 256;

static inline int ew_grid(size_t n_vec) {
  size_t b = (n_vec + EW_THREADS - 1) / EW_THREADS;
  const size_t cap = 148 * 8;
  return static_cast<int>(b < 1 ? 1 : (b > cap ? cap : b));
}

__global__ void __launch_bounds__(EW_THREADS) refresh_shadows_kernel(const float4* __restrict__ p, size_t n4, ShadowSet sh) {
  pdl_wait();
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
    if (4 * i + 4 <= sh.lo || 4 * i >= sh.hi) continue;
    const float4 pi = p[i];
    const float v[4] = {pi.x, pi.y, pi.z, pi.w};
    write_shadows(sh, 4 * i, v);
  }
}

// step[0] = completed steps, step[1] = arrival ticket.  Every block reads step[0] on entry; the last
// block to finish publishes step[0]+1, so no block can observe the new value mid-kernel.
// `advance` = 0: a range call that is not the last of its training step (per-bucket updates launched while backward is
// still running) leaves the step count alone; every call still consumes and resets the ticket.
DDLS_DEVINL void publish_step(int* step, int t, int advance) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    unsigned ticket = atomicAdd(reinterpret_cast<unsigned*>(step + 1), 1u);
    if (ticket == gridDim.x - 1) {
      if (advance) step[0] = t;
      step[1] = 0;
      __threadfence();
    }
  }
}

__global__ void __launch_bounds__(EW_THREADS) adam_kernel(float4* __restrict__ p, float4* __restrict__ g,
                                                          float4* __restrict__ m, float4* __restrict__ v, int* step,
                                                          size_t n4, float lr, float b1, float b2, float eps,
                                                          float gs, int zero_grads, ShadowSet sh, size_t elem0, int advance) {
  pdl_wait();
  const int t = *reinterpret_cast<volatile int*>(step) + 1;
  __shared__ float s_lr_t;                 // two powf per thread cost more than the update itself: one thread computes
  if (threadIdx.x == 0) s_lr_t = lr * sqrtf(1.0f - powf(b2, (float)t)) / (1.0f - powf(b1, (float)t));
  __syncthreads();
  const float lr_t = s_lr_t;
  const float ob1 = 1.0f - b1, ob2 = 1.0f - b2;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
    float4 gi = g[i], mi = m[i], vi = v[i], pi = p[i];
    gi.x *= gs; gi.y *= gs; gi.z *= gs; gi.w *= gs;
    mi.x = b1 * mi.x + ob1 * gi.x; mi.y = b1 * mi.y + ob1 * gi.y; mi.z = b1 * mi.z + ob1 * gi.z; mi.w = b1 * mi.w + ob1 * gi.w;
    vi.x = b2 * vi.x + ob2 * gi.x * gi.x; vi.y = b2 * vi.y + ob2 * gi.y * gi.y;
    vi.z = b2 * vi.z + ob2 * gi.z * gi.z; vi.w = b2 * vi.w + ob2 * gi.w * gi.w;
    pi.x -= lr_t * mi.x / (sqrtf(vi.x) + eps); pi.y -= lr_t * mi.y / (sqrtf(vi.y) + eps);
    pi.z -= lr_t * mi.z / (sqrtf(vi.z) + eps); pi.w -= lr_t * mi.w / (sqrtf(vi.w) + eps);
    m[i] = mi; v[i] = vi; p[i] = pi;
    if (zero_grads) g[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    const float pv[4] = {pi.x, pi.y, pi.z, pi.w};
    write_shadows(sh, elem0 + 4 * i, pv);
  }
  publish_step(step, t, advance);
}

__global__ void __launch_bounds__(EW_THREADS) adagrad_kernel(float4* __restrict__ p, float4* __restrict__ g,
                                                             float4* __restrict__ a, int* step, size_t n4, float lr,
                                                             float eps, float gs, int zero_grads, ShadowSet sh, size_t elem0, int advance) {
  pdl_wait();
  const int t = *reinterpret_cast<volatile int*>(step) + 1;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
    float4 gi = g[i], ai = a[i], pi = p[i];
    gi.x *= gs; gi.y *= gs; gi.z *= gs; gi.w *= gs;
    ai.x += gi.x * gi.x; ai.y += gi.y * gi.y; ai.z += gi.z * gi.z; ai.w += gi.w * gi.w;
    pi.x -= lr * gi.x / (sqrtf(ai.x) + eps); pi.y -= lr * gi.y / (sqrtf(ai.y) + eps);
    pi.z -= lr * gi.z / (sqrtf(ai.z) + eps); pi.w -= lr * gi.w / (sqrtf(ai.w) + eps);
    a[i] = ai; p[i] = pi;
    if (zero_grads) g[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    const float pv[4] = {pi.x, pi.y, pi.z, pi.w};
    write_shadows(sh, elem0 + 4 * i, pv);
  }
  publish_step(step, t, advance);
}

__global__ void __launch_bounds__(EW_THREADS) sgd_kernel(float4* __restrict__ p, float4* __restrict__ g, int* step,
                                                         size_t n4, float lr, float gs, int zero_grads, ShadowSet sh, size_t elem0,
                                                         int advance) {
  pdl_wait();
  const int t = *reinterpret_cast<volatile int*>(step) + 1;
  const float a = lr * gs;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
    float4 gi = g[i], pi = p[i];
    pi.x -= a * gi.x; pi.y -= a * gi.y; pi.z -= a * gi.z; pi.w -= a * gi.w;
    p[i] = pi;
    if (zero_grads) g[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    const float pv[4] = {pi.x, pi.y, pi.z, pi.w};
    write_shadows(sh, elem0 + 4 * i, pv);
  }
  publish_step(step, t, advance);
}

__global__ void __launch_bounds__(EW_THREADS) mnist_preprocess_kernel(const uint8_t* __restrict__ pix,
                                                                      const int64_t* __restrict__ labels,
                                                                      float* __restrict__ x, float* __restrict__ onehot,
                                                                      int n, int npix, int classes, float o_min,
                                                                      float scale, float n_min) {
  pdl_wait();
  const size_t total4 = (size_t)n * npix / 4;
  const uchar4* pix4 = reinterpret_cast<const uchar4*>(pix);
  float4* x4 = reinterpret_cast<float4*>(x);
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total4; i += stride) {
    uchar4 u = pix4[i];
    x4[i] = make_float4(((float)u.x - o_min) * scale + n_min, ((float)u.y - o_min) * scale + n_min,
                        ((float)u.z - o_min) * scale + n_min, ((float)u.w - o_min) * scale + n_min);
  }
  if (onehot != nullptr) {
    const size_t tot = (size_t)n * classes;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < tot; i += stride) {
      const int r = (int)(i / classes), c = (int)(i % classes);
      onehot[i] = (labels[r] == c) ? 1.0f : 0.0f;
    }
  }
}

static int check_flat(const void* a, size_t n, const char* who) {
  if ((reinterpret_cast<uintptr_t>(a) & 15) != 0 || (n & 3) != 0) {
    set_last_error("%s: flat buffers must be 16-byte aligned with a length multiple of 4 (got n=%zu)", who, n);
    return 1;
  }
  return 0;
}

}  // namespace ddls

using namespace ddls;

extern "C" int ddls_adam_step_range(float* params, float* grads, float* m, float* v, int* step, size_t n, size_t elem0,
                                    int advance, float lr, float beta1, float beta2, float eps, float grad_scale,
                                    int zero_grads, const DdlsShadowSpec* shadows, int n_shadows, void* stream) {
  if (n == 0) return 0;
  if (check_flat(params, n, "ddls_adam_step") || check_flat(grads, n, "ddls_adam_step") || (elem0 & 3)) return 1;
  DDLS_CUDA_OK(launch_k(adam_kernel, dim3(ew_grid(n / 4)), dim3(EW_THREADS), 0, static_cast<cudaStream_t>(stream),
      reinterpret_cast<float4*>(params), reinterpret_cast<float4*>(grads), reinterpret_cast<float4*>(m),
      reinterpret_cast<float4*>(v), step, n / 4, lr, beta1, beta2, eps, grad_scale, zero_grads,
      make_shadows(shadows, n_shadows), elem0, advance));
  count_launch(1);
  return check_cuda(cudaGetLastError(), "adam_kernel launch");
}

extern "C" int ddls_adam_step(float* params, float* grads, float* m, float* v, int* step, size_t n, float lr,
                              float beta1, float beta2, float eps, float grad_scale, int zero_grads,
                              const DdlsShadowSpec* shadows, int n_shadows, void* stream) {
  return ddls_adam_step_range(params, grads, m, v, step, n, 0, 1, lr, beta1, beta2, eps, grad_scale, zero_grads, shadows,
                              n_shadows, stream);
}

extern "C" int ddls_adagrad_step_range(float* params, float* grads, float* acc, int* step, size_t n, size_t elem0,
                                       int advance, float lr, float eps, float grad_scale, int zero_grads,
                                       const DdlsShadowSpec* shadows, int n_shadows, void* stream) {
  if (n == 0) return 0;
  if (check_flat(params, n, "ddls_adagrad_step") || check_flat(grads, n, "ddls_adagrad_step") || (elem0 & 3)) return 1;
  DDLS_CUDA_OK(launch_k(adagrad_kernel, dim3(ew_grid(n / 4)), dim3(EW_THREADS), 0, static_cast<cudaStream_t>(stream),
      reinterpret_cast<float4*>(params), reinterpret_cast<float4*>(grads), reinterpret_cast<float4*>(acc), step, n / 4,
      lr, eps, grad_scale, zero_grads, make_shadows(shadows, n_shadows), elem0, advance));
  count_launch(1);
  return check_cuda(cudaGetLastError(), "adagrad_kernel launch");
}

extern "C" int ddls_adagrad_step(float* params, float* grads, float* acc, int* step, size_t n, float lr, float eps,
                                 float grad_scale, int zero_grads, const DdlsShadowSpec* shadows, int n_shadows,
                                 void* stream) {
  return ddls_adagrad_step_range(params, grads, acc, step, n, 0, 1, lr, eps, grad_scale, zero_grads, shadows, n_shadows, stream);
}

extern "C" int ddls_sgd_step_range(float* params, float* grads, int* step, size_t n, size_t elem0, int advance, float lr,
                                   float grad_scale, int zero_grads, const DdlsShadowSpec* shadows, int n_shadows,
                                   void* stream) {
  if (n == 0) return 0;
  if (check_flat(params, n, "ddls_sgd_step") || check_flat(grads, n, "ddls_sgd_step") || (elem0 & 3)) return 1;
  DDLS_CUDA_OK(launch_k(sgd_kernel, dim3(ew_grid(n / 4)), dim3(EW_THREADS), 0, static_cast<cudaStream_t>(stream),
      reinterpret_cast<float4*>(params), reinterpret_cast<float4*>(grads), step, n / 4, lr, grad_scale, zero_grads,
      make_shadows(shadows, n_shadows), elem0, advance));
  count_launch(1);
  return check_cuda(cudaGetLastError(), "sgd_kernel launch");
}

extern "C" int ddls_sgd_step(float* params, float* grads, int* step, size_t n, float lr, float grad_scale,
                             int zero_grads, const DdlsShadowSpec* shadows, int n_shadows, void* stream) {
  return ddls_sgd_step_range(params, grads, step, n, 0, 1, lr, grad_scale, zero_grads, shadows, n_shadows, stream);
}

namespace ddls {
// Minibatch gather from a device-resident dataset (the B200 form of "the partition is cached in executor memory",
// /root/reference/ddl_mnist_aztk.py:156-161): rows perm[(k*B + r) mod N] of X and Y are copied into the step's static
// input buffers, k being a DEVICE-side batch cursor that the last block advances — so the gather is one node of the
// captured training-step graph and a replay needs no host-side input copy at all.  Rows are copied in 4-byte words
// (16-byte vectors when the row size allows).
template <typename VecT>
__global__ void __launch_bounds__(EW_THREADS) gather_batch_kernel(const VecT* __restrict__ X, const VecT* __restrict__ Y,
                                                                  VecT* __restrict__ x, VecT* __restrict__ y,
                                                                  const int* __restrict__ perm, unsigned* cursor,
                                                                  int B, long long N, int xw, int yw) {
  pdl_wait();
  const unsigned long long k = cursor[0];
  const size_t total_x = (size_t)B * xw, total = total_x + (size_t)B * yw;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const bool is_x = i < total_x;
    const size_t j = is_x ? i : i - total_x;
    const int w = is_x ? xw : yw;
    const int r = (int)(j / w), c = (int)(j - (size_t)r * w);
    long long row = (long long)((k * (unsigned long long)B + r) % (unsigned long long)N);
    if (perm != nullptr) row = perm[row];
    if (is_x) x[j] = X[(size_t)row * xw + c]; else y[j] = Y[(size_t)row * yw + c];
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    if (atomicAdd(cursor + 1, 1u) == gridDim.x - 1) { cursor[1] = 0; cursor[0] = (unsigned)(k + 1); __threadfence(); }
  }
}
}  // namespace ddls

extern "C" int ddls_gather_batch(const void* X, const void* Y, void* x, void* y, const int* perm, unsigned* cursor, int B,
                                 long long N, long long row_x_bytes, long long row_y_bytes, void* stream) {
  using namespace ddls;
  if (B <= 0 || N <= 0) { set_last_error("ddls_gather_batch: empty batch or dataset"); return 1; }
  if (row_x_bytes % 4 != 0 || row_y_bytes % 4 != 0 || row_x_bytes <= 0 || row_y_bytes <= 0) {
    set_last_error("ddls_gather_batch: row sizes must be positive multiples of 4 bytes (got %lld, %lld)", row_x_bytes, row_y_bytes);
    return 1;
  }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const bool v16 = row_x_bytes % 16 == 0 && row_y_bytes % 16 == 0 &&
                   ((reinterpret_cast<uintptr_t>(X) | reinterpret_cast<uintptr_t>(Y) | reinterpret_cast<uintptr_t>(x) |
                     reinterpret_cast<uintptr_t>(y)) & 15) == 0;
  if (v16) {
    const int xw = (int)(row_x_bytes / 16), yw = (int)(row_y_bytes / 16);
    int blocks = ew_grid((size_t)B * (xw + yw));
    if (blocks > 64) blocks = 64;
    launch_k(gather_batch_kernel<uint4>, dim3(blocks), dim3(EW_THREADS), 0, st, static_cast<const uint4*>(X), static_cast<const uint4*>(Y),
                                                               static_cast<uint4*>(x), static_cast<uint4*>(y), perm, cursor, B, N, xw, yw);
  } else {
    const int xw = (int)(row_x_bytes / 4), yw = (int)(row_y_bytes / 4);
    int blocks = ew_grid((size_t)B * (xw + yw));
    if (blocks > 64) blocks = 64;
    launch_k(gather_batch_kernel<uint32_t>, dim3(blocks), dim3(EW_THREADS), 0, st, static_cast<const uint32_t*>(X), static_cast<const uint32_t*>(Y),
                                                                  static_cast<uint32_t*>(x), static_cast<uint32_t*>(y), perm, cursor, B, N, xw, yw);
  }
  count_launch(1);
  return check_cuda(cudaGetLastError(), "gather_batch_kernel launch");
}

extern "C" int ddls_mnist_preprocess(const uint8_t* pixels, const int64_t* labels, float* x, float* onehot, int n,
                                     int npix, int classes, float o_min, float scale, float n_min, void* stream) {
  if (n <= 0) return 0;
  if (((size_t)n * npix) % 4 != 0) { set_last_error("ddls_mnist_preprocess: n*npix must be a multiple of 4"); return 1; }
  launch_k(mnist_preprocess_kernel, dim3(ew_grid((size_t)n * npix / 4)), dim3(EW_THREADS), 0, static_cast<cudaStream_t>(stream), 
      pixels, labels, x, onehot, n, npix, classes, o_min, scale, n_min);
  count_launch(1);
  return check_cuda(cudaGetLastError(), "mnist_preprocess_kernel launch");
}

extern "C" int ddls_refresh_shadows(const float* params, size_t n, const DdlsShadowSpec* shadows, int n_shadows,
                                    void* stream) {
  if (n == 0 || n_shadows <= 0) return 0;
  if (check_flat(params, n, "ddls_refresh_shadows")) return 1;
  launch_k(refresh_shadows_kernel, dim3(ew_grid(n / 4)), dim3(EW_THREADS), 0, static_cast<cudaStream_t>(stream), 
      reinterpret_cast<const float4*>(params), n / 4, make_shadows(shadows, n_shadows));
  count_launch(1);
  return check_cuda(cudaGetLastError(), "refresh_shadows_kernel launch");
}
