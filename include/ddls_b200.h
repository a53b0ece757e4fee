/* C ABI of libddls_b200.so — hand-written sm_100a kernels and the peer-memory runtime.
 *
 * Every function returns 0 on success and a non-zero code on failure; ddls_last_error() then
 * returns a thread-local message.  Pointers are raw device pointers unless noted; `stream` is a
 * cudaStream_t passed as void* (NULL = legacy default stream).  Citations name the place in the
 * reference whose behaviour the kernel reproduces.
 */
#ifndef DDLS_B200_H_
#define DDLS_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

const char* ddls_last_error(void);
int ddls_version(void);
/* Number of kernels this library has launched in this process (bench.py reports it). */
unsigned long long ddls_launch_count(void);
/* Programmatic dependent launch between the kernels of a training step (on by default, DDLS_PDL=0 at start-up or
 * ddls_set_pdl(0) switch it off): the next kernel's prologue overlaps the tail of its predecessor. */
int ddls_set_pdl(int on);

/* ---- generic tcgen05 GEMM ------------------------------------------------------------------ */
#define DDLS_GEMM_RELU 1u
#define DDLS_GEMM_BIAS 2u
#define DDLS_GEMM_OUT_BF16 4u
#define DDLS_GEMM_ATOMIC_ADD 8u   /* internal */
#define DDLS_GEMM_A_F32 16u       /* A is fp32 in memory (converted to bf16 while staging) */
#define DDLS_GEMM_B_F32 32u       /* B is fp32 in memory */
#define DDLS_GEMM_SPLITK_WS 64u   /* internal */
#define DDLS_GEMM_ACCUMULATE 128u /* C += result (fp32 C) */
#define DDLS_GEMM_FORCE_V1 256u   /* use the register-staged mainloop even when the cp.async one applies */
#define DDLS_GEMM_NO_TMA 512u     /* do not route large problems to the TMA / warp-specialised kernel */
/* C[M,N] = op(A)·op(B)ᵀ (+bias, ReLU, cast).  a_mn_major=0: A is [M,K] row-major; 1: A is [K,M]
 * row-major.  Same for B over N.  Operands are bf16, or fp32 with the *_F32 flags.  split_k>1 writes
 * fp32 partial tiles to `workspace` (split_k*M*N*4 bytes of scratch) and a second kernel sums them
 * and applies the epilogue — no atomics.  (Keras Dense / matmul, /root/reference/ddl_mnist_aztk.py:189-191.) */
int ddls_gemm_bf16(const void* A, const void* B, void* C, const float* bias, int M, int N, int K, int lda, int ldb,
                   int ldc, int a_mn_major, int b_mn_major, int split_k, unsigned flags, void* workspace,
                   size_t workspace_bytes, void* stream);
/* The persistent TMA-fed, warp-specialised kernel (csrc/tma_gemm.cu) directly: bf16 operands, 16-byte aligned
 * bases, lda/ldb/N multiples of 8, no split-K.  ddls_gemm_bf16 routes to it on its own for M >= 256, split_k == 1. */
int ddls_gemm_tma(const void* A, const void* B, void* C, const float* bias, int M, int N, int K, int lda, int ldb,
                  int ldc, int a_mn_major, int b_mn_major, unsigned flags, void* stream);

/* ---- fused optimizers over a flat fp32 parameter space --------------------------------------
 * Keras-2 update rules named by the reference ('adam' ddl_mnist_aztk.py:195, ddl_nyiso_aztk.py:255;
 * 'adagrad' ddl_nyiso_aztk.py:207).  `step` is a device int32 counter the kernel increments, so a
 * captured CUDA graph replays correctly.  grad_scale multiplies the gradient first (1/N, 1/passes).
 * If zero_grads != 0 the gradient buffer is cleared after use (saves a memset launch). */
#define DDLS_MAX_SHADOWS 4
/* A bf16 copy of the parameter matrix at flat element offset `begin` ([rows, cols] row-major) kept at dst with
 * row stride ld (>= cols, multiple of 8; padding columns stay zero).  The optimizer kernels refresh it in the
 * same pass that updates the fp32 master ("fused cast"). */
typedef struct {
  unsigned long long begin;
  unsigned rows, cols, ld, kind;
  void* dst;
} DdlsShadowSpec;
int ddls_adam_step(float* params, float* grads, float* m, float* v, int* step, size_t n, float lr, float beta1,
                   float beta2, float eps, float grad_scale, int zero_grads, const DdlsShadowSpec* shadows,
                   int n_shadows, void* stream);
int ddls_adagrad_step(float* params, float* grads, float* acc, int* step, size_t n, float lr, float eps,
                      float grad_scale, int zero_grads, const DdlsShadowSpec* shadows, int n_shadows, void* stream);
int ddls_sgd_step(float* params, float* grads, int* step, size_t n, float lr, float grad_scale, int zero_grads,
                  const DdlsShadowSpec* shadows, int n_shadows, void* stream);
/* Range forms: params / grads / slots point at flat element `elem0` of the parameter space (shadow specs use flat indices)
 * and cover n elements; `advance` = 0 leaves the step count alone (every range call of a training step but the last).
 * Used to update each gradient bucket as soon as it is final, on a side stream, while backward is still running. */
int ddls_adam_step_range(float* params, float* grads, float* m, float* v, int* step, size_t n, size_t elem0, int advance,
                         float lr, float beta1, float beta2, float eps, float grad_scale, int zero_grads,
                         const DdlsShadowSpec* shadows, int n_shadows, void* stream);
int ddls_adagrad_step_range(float* params, float* grads, float* acc, int* step, size_t n, size_t elem0, int advance,
                            float lr, float eps, float grad_scale, int zero_grads, const DdlsShadowSpec* shadows,
                            int n_shadows, void* stream);
int ddls_sgd_step_range(float* params, float* grads, int* step, size_t n, size_t elem0, int advance, float lr,
                        float grad_scale, int zero_grads, const DdlsShadowSpec* shadows, int n_shadows, void* stream);
int ddls_refresh_shadows(const float* params, size_t n, const DdlsShadowSpec* shadows, int n_shadows, void* stream);

/* ---- on-device preprocessing (MinMax + reshape + one-hot, ddl_mnist_aztk.py:126-144) ---------- */
/* pixels uint8 [n,784] -> x fp32 [n,28,28,1] = (p - o_min)*scale + n_min ; labels int64 [n] -> onehot fp32 [n,classes] */
int ddls_mnist_preprocess(const uint8_t* pixels, const int64_t* labels, float* x, float* onehot, int n, int npix,
                          int classes, float o_min, float scale, float n_min, void* stream);

/* Minibatch gather from a device-resident dataset: rows perm[(k*B + r) mod N] (perm may be NULL) of X [N, row_x_bytes] and
 * Y [N, row_y_bytes] -> x [B, ...], y [B, ...]; k = cursor[0] is a device-side batch counter the kernel advances
 * (cursor: uint32[2], zero-initialised), so the gather can be the first node of a captured training-step graph. */
int ddls_gather_batch(const void* X, const void* Y, void* x, void* y, const int* perm, unsigned* cursor, int B, long long N,
                      long long row_x_bytes, long long row_y_bytes, void* stream);

/* ---- persistent LSTM(128) / GRU(128) recurrences, tcgen05 (csrc/rnn128.cu; Keras LSTM / GRU of
 * /root/reference/ddl_nyiso_aztk.py:201-203, 249-251; x [B,T] fp32, gate order i,f,c,o / z,r,h) ------------------
 * ddls_rnn128_pack: recurrent [128, G*128], w_ih [G*128], bias [G*128] (fp32, Keras layout) -> bf16 UMMA
 * image [18][G*16][8][8] (16 hidden k-groups, 1 k-group of input-kernel/bias hi-lo rows, 1 zero k-group);
 * pulled into shared memory by the kernels below with bulk TMA copies. */
int ddls_rnn128_pack(const float* recurrent, const float* w_ih, const float* bias, void* image, int gates, void* stream);
/* B8 = ceil(B/8).  hs: [T][B8][128] x 16 B (h_t, 8 batch rows per vector = K-major UMMA image);
 * gates: [T][B8][4][128] x 16 B; cs: [T][B8][128] x 32 B; dz: [T][B8][512] x 16 B (K-major UMMA image). */
int ddls_lstm128_forward_v2(const float* x, const void* image, float* out, void* hs, void* gates, void* cs, int B, int T,
                            int return_sequences, int save_for_backward, int tile, void* stream);
int ddls_lstm128_backward_v2(const float* dout, const float* x, const void* image, const void* gates, const void* cs,
                             void* dz, float* d_wih, float* d_bias, int B, int T, int return_sequences, int tile,
                             void* stream);
/* GRU: hs, rh: [T][B8][128] x 16 B images (h_t, r_t*h_{t-1}); zrh: [T][B8][3][128] x 16 B (z, r, hh);
 * hprev: [T][B8][128] x 32 B (fp32 h_{t-1}); da: [T][B8][384] x 16 B image of d(az), d(ar), d(ah). */
int ddls_gru128_forward_v2(const float* x, const void* image, float* out, void* hs, void* rh, void* zrh, void* hprev,
                           int B, int T, int return_sequences, int save_for_backward, int tile, void* stream);
int ddls_gru128_backward_v2(const float* dout, const float* x, const void* image, const void* zrh, const void* hprev,
                            void* da, float* d_wih, float* d_bias, int B, int T, int return_sequences, int tile,
                            void* stream);
/* C[128, n] (ldc) = sum_g A_img[g]^T · B_img[g][b_noff : b_noff+n]; images as above, kgroups % 8 == 0, n % 64 == 0. */
int ddls_rnn128_wgrad(const void* a_img, const void* b_img, float* C, int kgroups, int n, int b_ntotal, int b_noff,
                      int ldc, void* stream);

/* ---- large-hidden LSTM (csrc/lstm_big.cu; BASELINE.json config 4: hidden 1024, seq 168) ------------------
 * 4H-wide tensors use the permuted column order n = (unit/32)*128 + gate*32 + unit%32.  H % 128 == 0. */
int ddls_lstm_big_perm_cast(const float* src, void* dst_bf16, int rows, int H, void* stream);   /* [rows,4H] Keras -> perm bf16 */
int ddls_lstm_big_perm_vec(const float* src, float* dst, int H, void* stream);                 /* [4H] Keras -> perm */
int ddls_lstm_big_unperm(const float* src, float* dst, int rows, int H, int accumulate, void* stream);
/* out[keras col] += sum_r w[r]*X[r][perm col] (w NULL: column sums); X bf16 [R,4H]; caller zeroes out. */
int ddls_lstm_big_colsum(const void* X, const float* w, float* out_plain, float* out_weighted, int R, int H, void* stream);
/* Second-generation step kernels on the TMA mainloop (csrc/lstm_big_tma.cu).  zx1: input projection of a first layer
 * with input dimension 1, zx[t][b][:] = x[b][t]*wih_perm + bias_perm (bf16).  step_fwd2 runs time step t of one layer
 * over whole-sequence buffers: hs [T+1][B][H] bf16 (hs[0] = initial state), cs [T+1][B][H] fp32, zx [T][B][4H] bf16
 * with the bias folded in, gates [T][B][4H] bf16 (or NULL); it writes hs[t+1], cs[t+1], gates[t]. */
int ddls_lstm_big_zx1(const float* x, const float* wih_perm, const float* bias_perm, void* zx, int B, int T, int H,
                      void* stream);
int ddls_lstm_big_step_fwd2(const void* hs, const void* u_perm, const void* zx, float* cs, void* gates, float* h_out_f32,
                            int B, int T, int H, int t, void* stream);
/* BPTT step t on whole-sequence buffers (dz [T][B][4H] bf16, gates, cs as above, dc [B][H] fp32 scratch): K = 4H is
 * split over a 4-CTA cluster whose partial accumulators are exchanged through distributed shared memory. */
int ddls_lstm_big_step_bwd2(void* dz, const void* u_perm, const void* dout_bf16, const float* dout_f32, const void* gates,
                            const float* cs, float* dc, int B, int T, int H, int t, void* stream);
/* Whole-sequence persistent kernels (csrc/lstm_big_seq.cu): ONE cooperative launch per layer (per group of batch tiles
 * that fits the SMs) runs all T steps with c / dc in registers and a per-batch-tile step barrier.  `sync` is a device
 * array of at least 32*ceil(B/128) uint32 (scratch, zeroed by the call; one 128-byte line per batch tile).  dout: [T][B][H] bf16 gradient from the layer
 * above (or NULL), dout_last: [B][H] fp32 gradient of the final hidden state (or NULL).
 * ddls_lstm_big_seq_tiles_per_launch returns how many 128-row batch tiles one launch holds (0 = layer too wide: use the
 * per-step kernels). */
int ddls_lstm_big_seq_tiles_per_launch(int H, int backward);
int ddls_lstm_big_seq_fwd(void* hs, const void* u_perm, const void* zx, float* cs, void* gates, float* h_out_f32, void* sync,
                          int B, int T, int H, void* stream);
int ddls_lstm_big_seq_bwd(void* dz, const void* u_perm, const void* dout, const float* dout_last, const void* gates,
                          const float* cs, void* sync, int B, int T, int H, void* stream);


/* ---- fused output heads (forward + backward in one launch) -------------------------------------- */
/* Dense(H->O, linear) + Keras mean_squared_error (ddl_nyiso_aztk.py:203,207).  y == NULL: pred only.
 * One CTA; every output (pred, loss[1], dh [B,H], dwd [H,O], dbd [O]) is overwritten. */
int ddls_dense_mse(const float* h, const float* wd, const float* bd, const float* y, float* pred, float* loss,
                   float* dh, float* dwd, float* dbd, int B, int H, int O, void* stream);
/* Dense(F->C) + softmax + categorical_crossentropy (ddl_mnist_aztk.py:191-196).  onehot == NULL:
 * probabilities only.  B/8 CTAs; dw [F,C], db [C], dprev_bias [F] are ACCUMULATED (caller zeroes; they are
 * slices of the flat gradient buffer).  loss[1], probs, da are overwritten.  da: fp32 [B,F] and/or bf16
 * [B,ldz] (zero padded).  relu_input folds the previous layer's ReLU in and emits its bias gradient.
 * scratch: (B/8 + 2) floats, zero-initialised ONCE by the caller (holds a self-resetting arrival ticket
 * and the per-CTA loss partials). */
int ddls_dense_softmax_xent(const float* a, const float* w, const float* bias, const float* onehot, float* probs,
                            float* loss, float* da_f32, void* da_bf16, int ldz, float* dw, float* db,
                            float* dprev_bias, float* scratch, int B, int F, int C, int relu_input, void* stream);

/* ---- MNIST CNN trunk (/root/reference/ddl_mnist_aztk.py:180-190), NHWC, Keras kernel layouts --------
 * conv1+ReLU: x [B,28,28,1] fp32 -> a1 [B,26,26,32] bf16. */
int ddls_cnn_conv1_fwd(const float* x, const float* w1, const float* b1, void* a1, int B, void* stream);
/* conv2+bias+ReLU+maxpool 2x2 (csrc/cnn_tma.cu: persistent CTAs, one 4-D TMA box per filter tap with 64-byte swizzle,
 * tcgen05 implicit GEMM, pool + arg-max in the epilogue): a1 -> pooled [B,12,12,32] bf16 + arg-max idx u8. */
int ddls_cnn_conv2_pool_fwd(const void* a1, const float* w2, const float* b2, void* pooled, void* pidx, int B,
                            void* stream);
/* dz = da*(a>0) as bf16 [B,ldz] (zero padded) and db[f] += sum_b dz  (caller zeroes db). */
int ddls_relu_bwd_bias(const float* da, const float* a, void* dz, float* db, int B, int F, int ldz, void* stream);
/* max-pool + ReLU backward: dflat [B,12,12,32] bf16 -> dz2 [B,24,24,32] bf16; db2[c] += sum (caller zeroes). */
int ddls_cnn_unpool_bwd(const void* dflat, const void* pooled, const void* pidx, void* dz2, float* db2, int B,
                        void* stream);
/* conv2 weight gradient dw2 [3,3,32,32] fp32 on the same TMA boxes (MN-major operands, K = pixels); one partial dW per
 * persistent CTA in `scratch` (ddls_cnn_conv2_wgrad_scratch_bytes(B) lets every CTA run; at least 36,864 bytes). */
size_t ddls_cnn_conv2_wgrad_scratch_bytes(int B);
int ddls_cnn_conv2_wgrad(const void* a1, const void* dz2, float* dw2, void* scratch, size_t scratch_bytes, int B,
                         void* stream);
/* conv2 data gradient fused with conv1's ReLU mask: dz1 [B,26,26,32] bf16 (zero padding = TMA out-of-bounds fill). */
int ddls_cnn_conv2_dgrad(const void* dz2, const float* w2, const void* a1, void* dz1, int B, void* stream);
/* conv1 gradients: dw1_db1 = [3*3*1*32 kernel gradient | 32 bias gradient] fp32; scratch B*320*4 bytes. */
int ddls_cnn_conv1_wgrad(const float* x, const void* dz1, float* dw1_db1, void* scratch, size_t scratch_bytes, int B,
                         void* stream);
/* Same result as ddls_cnn_conv1_wgrad for callers that cannot allocate GPU memory (fixed memory budgets,
 * loops that must not hit cudaMalloc's device-wide synchronisation, bare C callers): the scratch is a
 * 1.25 MiB array built into the library, held by every GPU that loads it until exit.  B must be
 * <= DDLS_NOALLOC_CONV1_MAX_BATCH, larger batches are rejected with an error.  The first call on a device
 * resolves the array's address (cudaGetSymbolAddress) — make it before any stream capture on that device.
 * Calls share the array: calls that overlap on the GPU race; it needs no clearing between calls.
 * No counterpart in the reference (which has no native code and no notion of GPU scratch). */
#define DDLS_NOALLOC_CONV1_MAX_BATCH 1024
int ddls_cnn_conv1_wgrad_noalloc(const float* x, const void* dz1, float* dw1_db1, int B, void* stream);

/* ---- symmetric peer-memory heap (CUDA VMM + optional NVSwitch multicast) -------------------------------
 * Replaces dist-keras' pickle-over-TCP parameter exchange.  File descriptors returned
 * here are passed between ranks by the caller (SCM_RIGHTS over a Unix socket). */
int ddls_multicast_supported(int device);
int ddls_symm_granularity(int device, int world, size_t* out);
int ddls_symm_create(int device, size_t size, unsigned long long* handle_out, void** ptr_out, int* fd_out);
int ddls_symm_import(int device, int fd, size_t size, unsigned long long* handle_out, void** ptr_out);
int ddls_symm_unmap(void* ptr, size_t size, unsigned long long handle);
int ddls_mc_create(int world, size_t size, unsigned long long* mc_handle_out, int* fd_out);
int ddls_mc_import(int fd, unsigned long long* mc_handle_out);
int ddls_mc_add_device(unsigned long long mc_handle, int device);
int ddls_mc_bind_and_map(unsigned long long mc_handle, unsigned long long mem_handle, size_t size, int device,
                         void** mc_ptr_out);

/* ---- peer-memory collectives (one kernel each; offsets are symmetric offsets into the heaps) ---------- */
#define DDLS_MAX_RANKS 8
#define DDLS_COMM_MAX_BLOCKS 128
#define DDLS_COMM_CONTROL_BYTES 8192 /* flags [128][8] u32 at offset 0, epoch counter at 4096 */
typedef struct {
  void* base[DDLS_MAX_RANKS]; /* this process' mappings of every rank's heap */
  void* mc_base;              /* multicast mapping or NULL */
  int world;
  int rank;
  int emulate;                /* 1: all ranks live on this device; one cooperative launch covers them */
  unsigned timeout_ms;        /* flag-wait timeout (0 = 20 s); on expiry the kernel prints and traps */
  uint64_t flags_off;
  uint64_t seq_off;
} DdlsCommDesc;
enum { DDLS_AR_ONESHOT = 0, DDLS_AR_TWOSHOT = 1, DDLS_AR_NVLS = 2, DDLS_AR_ONESHOT_LL = 3 };
/* DDLS_AR_ONESHOT_LL: flag-in-data packets (8 bytes = 4 payload + epoch), no flag round and no system fence; its staging
 * region must be its own, zero-initialised, 2 * world * n * 8 bytes (fp32 wire) and never used by the other algorithms. */
/* out = scale * sum_ranks(in) over fp32 [n].  stage_off / stage_bytes: a scratch region in each heap, used as two
 * parity halves of stage_bytes/2 (each must hold world*n wire elements; the halves are fixed so that calls of different
 * sizes never overlap).  wire_bf16: values cross NVLink as bf16 (fp32 accumulate).  blocks = 0 picks by size. */
#define DDLS_WIRE_BF16 1   /* wire_flags: values cross NVLink as bf16, fp32 accumulate */
#define DDLS_LL_MC_PUSH 2 /* wire_flags: flag-in-data packets are pushed with one store through the multicast mapping */
int ddls_comm_allreduce(const DdlsCommDesc* d, int algo, uint64_t in_off, uint64_t out_off, uint64_t stage_off,
                        uint64_t stage_bytes, uint64_t n, float scale, int wire_flags, int blocks, void* stream);
/* Flag-in-data allreduce of the gradient range at grad_off FUSED with the optimizer update (kind 1 Keras Adam: slot1 = m,
 * slot2 = v; 2 Adagrad: slot1 = accumulator; 3 SGD): the thread that finishes summing a gradient vector updates the
 * matching parameter vector; `scale` already contains 1/N and the optimizer's gradient scale; step = the optimizer's
 * device counter {steps, ticket}.  With zero_grads the gradient range is cleared, else it holds the reduced gradient.
 * n_opt <= n: the optimizer owns the first n_opt elements; the rest of the message (metric slots) is only reduced. */
int ddls_comm_allreduce_opt(const DdlsCommDesc* d, uint64_t grad_off, uint64_t stage_off, uint64_t stage_bytes, uint64_t n,
                            uint64_t n_opt, float scale, int wire_flags, int kind, float* params, float* slot1, float* slot2, int* step,
                            float lr, float b1, float b2, float eps, int zero_grads, int advance_step, void* stream);
/* Sharded data-parallel step of ONE gradient bucket (csrc/comm.cu, dp_step_kernel): reduce-scatter of the bucket's
 * gradients (multimem.ld_reduce, or peer loads without multicast), the Keras optimizer on this rank's 1/world slice
 * (rank r owns vectors [r*ceil(nvec/world), ...) of the bucket; its slots at slot*_off in ITS heap are the only valid
 * ones), and an all-gather of the new parameters by multicast store; then every rank clears its gradient bucket and
 * refreshes its bf16 shadow weights locally.  All offsets are symmetric heap offsets of the bucket's first element;
 * elem0 is that element's index in the flat parameter space (shadow specs are in flat indices).  advance_step: only
 * the last bucket kernel of a training step publishes step+1.  Replaces dist-keras' commit -> pull -> set_weights
 * cycle (/root/reference/ddl_mnist_aztk.py:216-219 via SURVEY.md E5/E6) for synchronous training. */
int ddls_comm_dp_step(const DdlsCommDesc* d, uint64_t grad_off, uint64_t param_off, uint64_t slot1_off, uint64_t slot2_off,
                      uint64_t n, uint64_t n_opt, uint64_t elem0, float scale, int kind, int* step, float lr, float b1,
                      float b2, float eps, int zero_grads, int advance_step, const DdlsShadowSpec* shadows, int n_shadows,
                      int blocks, void* stream);
/* Debug aid: later ddls_comm_dp_step launches write 5 globaltimer stamps per block (entry, after barrier A, slice done,
 * after barrier B, end) into buf (device memory, 5 * blocks uint64); NULL switches it off. */
int ddls_comm_dp_step_trace(unsigned long long* buf);
int ddls_comm_broadcast(const DdlsCommDesc* d, uint64_t off, uint64_t bytes, int root, int use_multicast, int blocks,
                        void* stream);
int ddls_comm_allgather(const DdlsCommDesc* d, uint64_t in_off, uint64_t out_off, uint64_t chunk_bytes, int blocks,
                        void* stream);
int ddls_comm_barrier(const DdlsCommDesc* d, void* stream);
/* ADAG/DynSGD commit+pull against a center variable replicated at center_off in every heap
 * (dist-keras ADAGWorker/ADAGParameterServer, /root/reference/ddl_mnist_aztk.py:216-219 via SURVEY.md E5/E6). */
int ddls_comm_adag_commit(const DdlsCommDesc* d, float* w, float* pulled, uint64_t center_off, uint64_t updates_off,
                          uint64_t n, float scale, int staleness_aware, unsigned* last_update, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DDLS_B200_H_ */
