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
