"""Per-step timeline of the persistent stacked-LSTM forward kernel (clock64 stamps of CTA (0,0)).
Slots: 0 barrier passed, 1 last h block issued, 2 first K-block landed, 3 last MMA issued, 4 accumulator complete,
5 stores issued, 6 fences done, 7 arrival posted."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import distributeddeeplearningspark_b200 as ddls
from distributeddeeplearningspark_b200 import _native as N
from distributeddeeplearningspark_b200.ops import lstm_big

B = int(sys.argv[1]) if len(sys.argv) > 1 else 512
T, H = 168, 1024
dev = torch.device('cuda', 0)
m = ddls.models.stacked_lstm(seed=0).to(dev)
layers = m.layers[:-1]
bufs = lstm_big.StackedLSTMBuffers(layers, B, T, dev)
x2 = torch.rand(B, T, device=dev)
st = N.stream_ptr(dev)
lstm_big.prepare_weights(layers, bufs, st)
if len(sys.argv) > 2:
    print('fence mode', sys.argv[2]); N.lib().ddls_lstm_big_seq_set_fence_mode(int(sys.argv[2]))
lstm_big.forward(layers, bufs, x2, st, save=True)
torch.cuda.synchronize()
NCTA = ((B + 127) // 128) * (4 * H // 128)
trace = torch.zeros(T * 8 + T * 2 * NCTA, dtype=torch.int64, device=dev)
N.lib().ddls_lstm_big_seq_set_trace.argtypes = [ctypes.c_void_p]
N.lib().ddls_lstm_big_seq_set_trace(trace.data_ptr())
lstm_big.forward(layers[:1], bufs, x2, st, save=True)      # first layer only
torch.cuda.synchronize()
N.lib().ddls_lstm_big_seq_set_trace(None)
tr = trace[:T * 8].view(T, 8).cpu().double()
ga = trace[T * 8:].view(T, 2, NCTA).cpu().double() / 1e3   # us
names = ['barrier', 'A issued', 'kb0 landed', 'mma issued', 'acc done', 'h stores', 'fences', 'arrive']
clk = 1.965e3  # cycles per us at max clock
print('--- forward')
print('step period (us):', float((tr[60:160, 0] - tr[59:159, 0]).mean() / clk))
for a, b in [(0, 2), (2, 3), (3, 4), (4, 5), (5, 6), (6, 7), (0, 1)]:
    d = (tr[50:160, b] - tr[50:160, a]) / clk
    print(f'{names[a]:>11} -> {names[b]:<11}: mean {float(d.mean()):6.2f} us  min {float(d.min()):6.2f}  max {float(d.max()):6.2f}')
d = (tr[51:160, 0] - tr[50:159, 7]) / clk
print(f'     arrive -> next barrier: mean {float(d.mean()):6.2f} us  min {float(d.min()):6.2f}  max {float(d.max()):6.2f}')

mt = (B + 127) // 128
arr = ga[60:160, 1]                                  # arrival time of every CTA, steps 60..159
rel = arr - arr.min(dim=1, keepdim=True).values
print('arrival spread over all CTAs (us): mean max %.2f, mean of means %.2f' % (float(rel.max(dim=1).values.mean()), float(rel.mean())))
late = rel.mean(dim=0)
order = torch.argsort(late, descending=True)
print('latest CTAs (cta = n_tile*%d + m_tile): ' % mt, [(int(c), round(float(late[c]), 2)) for c in order[:8]])
print('earliest CTAs: ', [(int(c), round(float(late[c]), 2)) for c in order[-4:]])
bar = ga[61:160, 0] - ga[60:159, 1].max(dim=1, keepdim=True).values   # last arrival -> each CTA passes the barrier
print('last arrival -> barrier passed (us): mean %.2f  max-CTA mean %.2f' % (float(bar.mean()), float(bar.max(dim=1).values.mean())))
work = (ga[60:160, 1] - ga[60:160, 0]).mean(dim=0)       # barrier passed -> arrival, per CTA
print('per-CTA work time (us): min %.2f  median %.2f  max %.2f' % (float(work.min()), float(work.median()), float(work.max())))
print('work by n_tile (mean over m):', [round(float(work.view(-1, mt)[n].mean()), 2) for n in range(work.numel() // mt)])
passd = (ga[61:160, 0] - ga[60:159, 1].max(dim=1, keepdim=True).values).mean(dim=0)
print('barrier pass delay by n_tile:', [round(float(passd.view(-1, mt)[n].mean()), 2) for n in range(work.numel() // mt)])
# backward, top layer
print('--- backward')
trace.zero_()
N.lib().ddls_lstm_big_seq_set_trace_bwd.argtypes = [ctypes.c_void_p]
N.lib().ddls_lstm_big_seq_set_trace_bwd(trace.data_ptr())
dh = torch.randn(B, H, device=dev)
i = 1
N.call('ddls_lstm_big_seq_bwd', bufs.dz[i].data_ptr(), bufs.u_perm[i].data_ptr(), 0, dh.data_ptr(), bufs.gates[i].data_ptr(),
       bufs.cs[i].data_ptr(), bufs.sync.data_ptr(), B, T, H, st)
torch.cuda.synchronize()
N.lib().ddls_lstm_big_seq_set_trace_bwd(None)
tr = trace[:T * 8].view(T, 8).cpu().double()
names = ['barrier', 'kb0 landed', 'mma issued', 'acc seen', 'exchanged', 'dz stores', 'fences', 'arrive']
print('step period (us):', float((tr[60:160, 0] - tr[59:159, 0]).mean() / clk))
for a, b in [(0, 1), (1, 2), (2, 3), (3, 4), (4, 5), (5, 6), (6, 7)]:
    d = (tr[50:160, b] - tr[50:160, a]) / clk
    print(f'{names[a]:>11} -> {names[b]:<11}: mean {float(d.mean()):6.2f} us  min {float(d.min()):6.2f}  max {float(d.max()):6.2f}')
d = (tr[51:160, 0] - tr[50:159, 7]) / clk
print(f'     arrive -> next barrier: mean {float(d.mean()):6.2f} us  min {float(d.min()):6.2f}  max {float(d.max()):6.2f}')
