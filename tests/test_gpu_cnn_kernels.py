"""Each CNN kernel in isolation against a plain fp32 PyTorch computation on the same (bf16-rounded) inputs."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

from distributeddeeplearningspark_b200 import _native as N
from distributeddeeplearningspark_b200.ops import cnn as C     # registers prototypes
from distributeddeeplearningspark_b200.ops.gemm import workspace


def _rel(a, b):
    return float((a.float() - b.float()).abs().max() / (b.float().abs().max() + 1e-12))


def _conv_ref(x_nhwc, w_hwio, b=None):
    y = F.conv2d(x_nhwc.permute(0, 3, 1, 2), w_hwio.permute(3, 2, 0, 1), b)
    return y.permute(0, 2, 3, 1).contiguous()


@pytest.fixture(scope='module')
def dev():
    return torch.device('cuda', 0)


@pytest.mark.parametrize('B', [3, 128])
def test_conv1_fwd(dev, B):
    torch.manual_seed(0)
    x = torch.rand(B, 28, 28, 1, device=dev)
    w = torch.randn(3, 3, 1, 32, device=dev) * 0.3
    b = torch.randn(32, device=dev) * 0.1
    a1 = torch.empty(B, 26, 26, 32, dtype=torch.bfloat16, device=dev)
    N.call('ddls_cnn_conv1_fwd', x.data_ptr(), w.data_ptr(), b.data_ptr(), a1.data_ptr(), B, N.stream_ptr(dev))
    ref = torch.relu(_conv_ref(x, w, b))
    assert _rel(a1, ref) < 6e-3


@pytest.mark.parametrize('entry', ['ddls_cnn_conv2_pool_fwd'])
@pytest.mark.parametrize('B', [3, 128])
def test_conv2_pool_fwd(dev, B, entry):
    torch.manual_seed(1)
    a1 = torch.relu(torch.randn(B, 26, 26, 32, device=dev)).bfloat16()
    w = (torch.randn(3, 3, 32, 32, device=dev) * 0.08)
    b = torch.randn(32, device=dev) * 0.1
    pooled = torch.empty(B, 12, 12, 32, dtype=torch.bfloat16, device=dev)
    pidx = torch.empty(B, 12, 12, 32, dtype=torch.uint8, device=dev)
    N.call(entry, a1.data_ptr(), w.data_ptr(), b.data_ptr(), pooled.data_ptr(), pidx.data_ptr(), B,
           N.stream_ptr(dev))
    z = _conv_ref(a1.float(), w.bfloat16().float(), b)                       # [B,24,24,32]
    zr = z.reshape(B, 12, 2, 12, 2, 32).permute(0, 1, 3, 2, 4, 5).reshape(B, 12, 12, 4, 32)
    ref_val, ref_idx = zr.max(dim=3)
    assert _rel(pooled, torch.relu(ref_val)) < 1e-2
    # arg-max must agree except where the two best window values are within rounding distance
    top2 = zr.topk(2, dim=3).values
    clear = (top2[:, :, :, 0] - top2[:, :, :, 1]) > 1e-2
    assert bool((pidx.long()[clear] == ref_idx[clear]).all())
    assert int(pidx.max()) <= 3


def test_unpool_bwd(dev):
    torch.manual_seed(2)
    B = 9
    dflat = torch.randn(B, 12, 12, 32, device=dev).bfloat16()
    pooled = torch.relu(torch.randn(B, 12, 12, 32, device=dev)).bfloat16()
    pidx = torch.randint(0, 4, (B, 12, 12, 32), device=dev, dtype=torch.uint8)
    dz2 = torch.empty(B, 24, 24, 32, dtype=torch.bfloat16, device=dev)
    db2 = torch.zeros(32, device=dev)
    N.call('ddls_cnn_unpool_bwd', dflat.data_ptr(), pooled.data_ptr(), pidx.data_ptr(), dz2.data_ptr(), db2.data_ptr(), B,
           N.stream_ptr(dev))
    g = torch.where(pooled.float() > 0, dflat.float(), torch.zeros((), device=dev))
    onehot = F.one_hot(pidx.long(), 4).permute(0, 1, 2, 4, 3).float()          # [B,12,12,4,32]
    ref = (onehot * g.unsqueeze(3)).reshape(B, 12, 12, 2, 2, 32).permute(0, 1, 3, 2, 4, 5).reshape(B, 24, 24, 32)
    torch.testing.assert_close(dz2.float(), ref)
    torch.testing.assert_close(db2, g.sum(dim=(0, 1, 2)), rtol=1e-4, atol=1e-3)


@pytest.mark.parametrize('entry', ['ddls_cnn_conv2_wgrad'])
@pytest.mark.parametrize('B', [2, 128])
def test_conv2_wgrad(dev, B, entry):
    torch.manual_seed(3)
    a1 = torch.relu(torch.randn(B, 26, 26, 32, device=dev)).bfloat16()
    dz2 = (torch.randn(B, 24, 24, 32, device=dev) * (torch.rand(B, 24, 24, 32, device=dev) < 0.25)).bfloat16()
    dw = torch.empty(3, 3, 32, 32, device=dev)
    ws = workspace(dev, N.lib().ddls_cnn_conv2_wgrad_scratch_bytes(B))
    N.call(entry, a1.data_ptr(), dz2.data_ptr(), dw.data_ptr(), ws.data_ptr(), ws.numel(), B,
           N.stream_ptr(dev))
    w = torch.zeros(3, 3, 32, 32, device=dev, requires_grad=True)
    (_conv_ref(a1.float(), w) * dz2.float()).sum().backward()
    assert _rel(dw, w.grad) < 5e-3


@pytest.mark.parametrize('entry', ['ddls_cnn_conv2_dgrad'])
@pytest.mark.parametrize('B', [2, 128])
def test_conv2_dgrad(dev, B, entry):
    torch.manual_seed(4)
    a1 = torch.relu(torch.randn(B, 26, 26, 32, device=dev)).bfloat16()
    dz2 = torch.randn(B, 24, 24, 32, device=dev).bfloat16()
    w = torch.randn(3, 3, 32, 32, device=dev) * 0.1
    dz1 = torch.empty(B, 26, 26, 32, dtype=torch.bfloat16, device=dev)
    N.call(entry, dz2.data_ptr(), w.data_ptr(), a1.data_ptr(), dz1.data_ptr(), B, N.stream_ptr(dev))
    xin = a1.float().clone().requires_grad_(True)
    (_conv_ref(xin, w.bfloat16().float()) * dz2.float()).sum().backward()
    ref = xin.grad * (a1.float() > 0)
    assert _rel(dz1, ref) < 1e-2


@pytest.mark.parametrize('B', [2, 128])
def test_conv1_wgrad(dev, B):
    torch.manual_seed(5)
    x = torch.rand(B, 28, 28, 1, device=dev)
    dz1 = torch.randn(B, 26, 26, 32, device=dev).bfloat16()
    out = torch.empty(320, device=dev)
    ws = workspace(dev, B * 320 * 4)
    N.call('ddls_cnn_conv1_wgrad', x.data_ptr(), dz1.data_ptr(), out.data_ptr(), ws.data_ptr(), ws.numel(), B,
           N.stream_ptr(dev))
    w = torch.zeros(3, 3, 1, 32, device=dev, requires_grad=True)
    b = torch.zeros(32, device=dev, requires_grad=True)
    (_conv_ref(x, w, b) * dz1.float()).sum().backward()
    assert _rel(out[:288].view(3, 3, 1, 32), w.grad) < 2e-3
    assert _rel(out[288:], b.grad) < 2e-3


def test_relu_bwd_bias(dev):
    torch.manual_seed(6)
    B, Fd = 128, 225
    da = torch.randn(B, Fd, device=dev)
    a = torch.relu(torch.randn(B, Fd, device=dev))
    dz = torch.empty(B, 232, dtype=torch.bfloat16, device=dev)
    db = torch.zeros(Fd, device=dev)
    N.call('ddls_relu_bwd_bias', da.data_ptr(), a.data_ptr(), dz.data_ptr(), db.data_ptr(), B, Fd, 232, N.stream_ptr(dev))
    ref = da * (a > 0)
    assert _rel(dz[:, :Fd], ref) < 5e-3 and float(dz[:, Fd:].float().abs().max()) == 0.0
    torch.testing.assert_close(db, ref.sum(0), rtol=1e-4, atol=1e-4)
