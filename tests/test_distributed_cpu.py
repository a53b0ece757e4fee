"""BASELINE.json config 1: MNIST CNN, 2 ranks, local barrier backend + gloo (plumbing, no GPU)."""
import numpy as np
import pytest
import torch

from distributeddeeplearningspark_b200 import spark
from distributeddeeplearningspark_b200.parallel import plan_buckets
from distributeddeeplearningspark_b200.spark import LocalBarrierBackend

ENV = {'DDLS_FORCE_CPU': '1', 'OMP_NUM_THREADS': '2'}


def _sync_train(steps, compression, passes):
    import distributeddeeplearningspark_b200 as ddls
    from distributeddeeplearningspark_b200.data import mnist_like_arrays
    from distributeddeeplearningspark_b200.trainers import categorical_crossentropy
    torch.set_num_threads(2)
    ddls.init()
    r, n = ddls.rank(), ddls.size()
    model = ddls.models.mnist_cnn(seed=100 + r)            # different init per rank on purpose
    opt = ddls.optim.KerasAdam(model.parameters())
    ddls.broadcast_parameters(opt, root_rank=0)
    after_bcast = ddls.allgather(opt.flat.flat_params[:4096].clone()).reshape(n, -1)
    dopt = ddls.DistributedOptimizer(opt, compression=ddls.Compression(compression), backward_passes_per_step=passes,
                                     bucket_bytes=1 << 20)
    px, lab = mnist_like_arrays(16 * steps * passes, seed=r)
    x = torch.tensor(px, dtype=torch.float32).reshape(-1, 28, 28, 1) / 250.0
    y = torch.eye(10)[torch.tensor(lab)]
    losses = []
    for s in range(steps * passes):
        xb, yb = x[s * 16:(s + 1) * 16], y[s * 16:(s + 1) * 16]
        if s % passes == 0:
            dopt.zero_grad()
        loss = categorical_crossentropy(model(xb), yb)
        loss.backward()
        dopt.step()
        losses.append(float(loss.detach()))
    allw = ddls.allgather(opt.flat.flat_params.clone()).reshape(n, -1)
    return {'rank': r, 'size': n, 'bcast_equal': bool((after_bcast[0] == after_bcast[1]).all()),
            'weights_equal': bool((allw[0] == allw[1]).all()), 'losses': losses,
            'n_buckets': len(dopt.buckets), 'w0': allw[0][:8].tolist()}


@pytest.mark.parametrize('compression,passes', [('none', 1), ('bf16', 1), ('none', 2)])
def test_mnist_cnn_two_ranks_gloo(compression, passes):
    out = spark.run(_sync_train, args=(3, compression, passes), num_proc=2, backend=LocalBarrierBackend(timeout=300),
                    env=ENV)
    assert [o['rank'] for o in out] == [0, 1] and all(o['size'] == 2 for o in out)
    assert all(o['bcast_equal'] for o in out)          # identical weights after broadcast
    assert all(o['weights_equal'] for o in out)        # identical weights after N synchronous steps
    assert out[0]['n_buckets'] >= 2                    # 1 MB buckets split the 4.2 MB model
    assert all(np.isfinite(o['losses']).all() for o in out)


def _metric_slots_job():
    import distributeddeeplearningspark_b200 as ddls
    ddls.init()
    r, n = ddls.rank(), ddls.size()
    outs = {}
    for slots in (0, 3):
        torch.manual_seed(11)
        ps = [torch.nn.Parameter(torch.randn(50)), torch.nn.Parameter(torch.randn(7, 3))]   # 50 -> padded to 52
        opt = ddls.optim.KerasAdagrad(ps, metric_slots=slots)
        dopt = ddls.DistributedOptimizer(opt)
        assert (dopt.metrics is None) == (slots == 0)
        for k in range(3):
            torch.manual_seed(100 * k + r)
            opt.flat.flat_grads[:opt.flat.numel].copy_(torch.randn(opt.flat.numel))
            if slots:
                dopt.metrics[0] = float(r + 1 + k)
                dopt.metrics[1] = 10.0 * (r + 1)
            dopt.step()
            if slots:   # averaged like the gradients, not consumed or cleared by the optimizer
                assert abs(float(dopt.metrics[0]) - (k + (n + 1) / 2.0)) < 1e-6, float(dopt.metrics[0])
                assert abs(float(dopt.metrics[1]) - 10.0 * (n + 1) / 2.0) < 1e-5
            opt.zero_grad()
        outs[slots] = opt.flat.flat_params.clone()
        assert opt.flat.flat_params.numel() == 76 and opt.flat.flat_grads.numel() == 76 + (4 if slots else 0)
    return {'rank': r, 'same': bool(torch.equal(outs[0], outs[3])), 'buckets_hi': dopt.buckets[-1].hi}


def test_metric_slots_ride_in_the_top_gradient_bucket():
    """K20: metric slots in the tail of the gradient buffer are reduced with the last bucket and do not disturb the
    parameter update."""
    out = spark.run(_metric_slots_job, num_proc=2, backend=LocalBarrierBackend(timeout=300), env=ENV)
    assert all(o['same'] for o in out)
    assert max(o['buckets_hi'] for o in out) == 80


def test_bucket_plan_is_static_and_tiles_the_flat_space():
    segs, off = [], 0
    for n in (288, 32, 9216, 32, 1036800, 228, 2250, 12):   # MNIST CNN tensors, 4-aligned
        segs.append((off, n)); off += (n + 3) // 4 * 4
    plan = plan_buckets(segs, 256 * 1024)
    assert plan == plan_buckets(segs, 256 * 1024)
    covered = sorted((lo, hi) for lo, hi, _ in plan)
    assert covered[0][0] == 0 and covered[-1][1] == off
    for (a, b), (c, d) in zip(covered[:-1], covered[1:]):
        assert b == c
    assert plan[0][2][0] == len(segs) - 1              # first bucket holds the last layer (backward order)
    assert sorted(i for _, _, idxs in plan for i in idxs) == list(range(len(segs)))


def _collectives():
    import distributeddeeplearningspark_b200 as ddls
    ddls.init()
    r = ddls.rank()
    t = torch.full((5,), float(r + 1))
    avg = ddls.allreduce(t)
    tot = ddls.allreduce(t, op=ddls.Sum)
    g = ddls.allgather(torch.tensor([[r, r * 10]]))
    b = ddls.broadcast(torch.tensor([r + 7.0]), root_rank=1)
    objs = ddls.allgather_object({'r': r})
    return avg.tolist(), tot.tolist(), g.tolist(), b.tolist(), objs


def test_collective_functions_gloo():
    out = spark.run(_collectives, num_proc=2, backend=LocalBarrierBackend(timeout=180), env=ENV)
    for avg, tot, g, b, objs in out:
        assert avg == [1.5] * 5 and tot == [3.0] * 5
        assert g == [[0, 0], [1, 10]] and b == [8.0] and objs == [{'r': 0}, {'r': 1}]


def test_allreduce_algorithm_selection_follows_the_measured_crossovers(monkeypatch):
    """select_algo / ll_max_bytes (host logic, no GPU needed): flag-in-data while (ranks-1) x bytes <= 800 KB, then NVLS;
    two-shot at two ranks from 2 MB; fenced one-shot / two-shot without multicast (profiles/sweep_n{2,4,8}.json)."""
    from distributeddeeplearningspark_b200.parallel import peer
    for k in ('DDLS_ALLREDUCE_ALGO', 'DDLS_LL_INBOUND_BYTES', 'DDLS_LL_MAX_BYTES', 'DDLS_ONESHOT_MAX_BYTES'):
        monkeypatch.delenv(k, raising=False)

    class Heap(peer._HeapBase):
        def __init__(self, size):
            self.size = size
            self._init_heap(1 << 30)
            self._init_ll_stage(size)

    KB, MB = 1 << 10, 1 << 20
    h2, h4, h8 = Heap(2), Heap(4), Heap(8)
    assert h2.ll_max_bytes(2) == 512 * KB and h4.ll_max_bytes(4) == 800 * KB // 3 and h8.ll_max_bytes(8) == 800 * KB // 7
    lstm = 66692 * 4                                   # the NYISO LSTM gradient bucket
    assert h2.select_algo(lstm, True, True) == peer.AR_ONESHOT_LL
    assert h4.select_algo(lstm, True, True) == peer.AR_ONESHOT_LL
    assert h8.select_algo(lstm, True, True) == peer.AR_NVLS
    assert h8.select_algo(64 * KB, True, True) == peer.AR_ONESHOT_LL
    assert h8.select_algo(4 * MB, True, True) == peer.AR_NVLS and h4.select_algo(64 * MB, True, True) == peer.AR_NVLS
    assert h2.select_algo(1 * MB, True, True) == peer.AR_NVLS and h2.select_algo(4 * MB, True, True) == peer.AR_TWOSHOT
    assert h8.select_algo(1 * MB, True, False) == peer.AR_ONESHOT and h8.select_algo(4 * MB, True, False) == peer.AR_TWOSHOT
    assert h8.select_algo(4 * MB, False, True) == peer.AR_ONESHOT          # output not symmetric: push into staging only
    monkeypatch.setenv('DDLS_ALLREDUCE_ALGO', 'twoshot')
    assert h8.select_algo(4 * KB, True, True) == peer.AR_TWOSHOT


def test_shard_bounds_match_the_kernels_slice_partition():
    """Rank r owns vectors [r*ceil(nvec/N), ...) of a bucket (csrc/comm.cu dp_step_kernel): the host-side bounds used to
    gather the sharded optimizer slots tile the bucket exactly, whatever the bucket size."""
    from distributeddeeplearningspark_b200.parallel.distributed_optimizer import DistributedOptimizer, _Bucket

    class Fake:
        shard_bounds = DistributedOptimizer.shard_bounds

    for size in (2, 3, 8):
        for lo, hi, numel in ((0, 4, 4), (128, 128 + 50008, 128 + 50000), (0, 1037252, 1037252), (64, 64 + 36, 64 + 36)):
            f = Fake(); f.size = size
            f.flat = type('F', (), {'numel': numel})()
            f.comm = type('C', (), {'rank': 0})()
            b = _Bucket(lo, hi, [])
            spans = [f.shard_bounds(b, r) for r in range(size)]
            assert spans[0][0] == lo
            for (a0, a1), (b0, b1) in zip(spans[:-1], spans[1:]):
                assert a1 == min(b0, numel) or (a1 == numel and b0 >= numel)   # contiguous up to the optimizer's range
            assert max(s[1] for s in spans) == min(hi, numel) and all(s[0] % 4 == 0 for s in spans)
