import os, sys
os.environ['CUDA_LAUNCH_BLOCKING'] = '1'
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributeddeeplearningspark_b200.models import layers as L
from distributeddeeplearningspark_b200.ops import recurrent as R
dev = torch.device('cuda', 0)
torch.manual_seed(1)
k = (torch.rand(1, 512, device=dev) - 0.5) * 0.4
r = torch.empty(128, 512); L.orthogonal_(r); r = r.to(dev)
b = (torch.rand(512, device=dev) - 0.5) * 0.2
x = torch.rand(32, 25, 1, device=dev)
ref = L.lstm_reference(x, k, r, b)
try:
    out = R.lstm_layer(x, k, r, b)
    torch.cuda.synchronize()
    print('max abs err', float((out - ref).abs().max()), 'ref max', float(ref.abs().max()))
except Exception as e:
    print('ERROR', type(e).__name__, str(e)[:300])
