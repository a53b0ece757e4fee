import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributeddeeplearningspark_b200.models import mnist_cnn
from distributeddeeplearningspark_b200.ops import cnn, fused
dev = torch.device('cuda', 0)
m = mnist_cnn(seed=0).to(dev)
with torch.no_grad():
    for l in m.layers:
        if hasattr(l, 'bias') and l.bias is not None:
            l.bias.uniform_(-0.05, 0.05)
for B in (128, 5):
    torch.manual_seed(B)
    x = torch.rand(B, 28, 28, 1, device=dev)
    a = x
    for l in m.layers[:8]:
        a = l(a)
    w = torch.randn_like(a)
    (a * w).sum().backward()
    ref = [p.grad.clone() for _, p in m.named_keras_weights()[:6]]
    for _, p in m.named_keras_weights():
        p.grad = None
    plan = fused._plan(m)
    out = cnn.trunk(plan, x)
    print('B', B, 'fwd rel', float((out - a).abs().max() / a.abs().max()), 'l2', float((out - a).norm() / a.norm()))
    (out * w).sum().backward()
    for n, (_, p), r in zip(['c1k', 'c1b', 'c2k', 'c2b', 'd1k', 'd1b'], m.named_keras_weights()[:6], ref):
        g = p.grad
        print(n, 'maxrel', float((g - r).abs().max() / r.abs().max()), 'l2', float((g - r).norm() / r.norm()))
    for _, p in m.named_keras_weights():
        p.grad = None
