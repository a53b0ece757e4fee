"""Per-step vs fixed cost of the v2 recurrent forward kernels: time vs T (CUDA events, warm)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributeddeeplearningspark_b200.ops import recurrent as R
dev = torch.device('cuda', 0)
H = 128
def timeit(fn, n=50):
    for _ in range(5): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e3
for cell, G in (('lstm', 4), ('gru', 3)):
    k = torch.randn(1, G * H, device=dev) * 0.1; r = torch.randn(H, G * H, device=dev) * 0.05; b = torch.zeros(G * H, device=dev)
    img = R.pack_image(r, k, b, G)
    for B in (32, 64):
        for T in (1, 5, 25, 50, 100):
            x2 = torch.rand(B, T, device=dev); out = torch.empty(B, H, device=dev)
            if cell == 'lstm':
                hs, gates, cs, dz = R.lstm_v2_buffers(B, T, dev)
                f_save = lambda: R.lstm_v2_forward(x2, img, out, hs, gates, cs, B, T, False, True)
                f_nosave = lambda: R.lstm_v2_forward(x2, img, out, None, None, None, B, T, False, False)
            else:
                hs, rh, zrh, hp, da = R.gru_v2_buffers(B, T, dev)
                f_save = lambda: R.gru_v2_forward(x2, img, out, hs, rh, zrh, hp, B, T, False, True)
                f_nosave = lambda: R.gru_v2_forward(x2, img, out, None, None, None, None, B, T, False, False)
            print(f'{cell} B={B} T={T:3d}: save {timeit(f_save):8.2f} us   nosave {timeit(f_nosave):8.2f} us', flush=True)
