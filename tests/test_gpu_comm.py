"""Peer-memory collectives.  Single-GPU tests emulate the ranks inside ONE cooperative launch
(never as separate launches that wait on each other); multi-GPU tests run one process per GPU."""
import pytest
import torch

pytestmark = pytest.mark.gpu

from distributeddeeplearningspark_b200.parallel.peer import AR_ONESHOT, AR_ONESHOT_LL, AR_TWOSHOT, EmulatedPeerGroup


@pytest.mark.parametrize('world', [2, 4, 8])
@pytest.mark.parametrize('algo', [AR_ONESHOT, AR_TWOSHOT])
@pytest.mark.parametrize('n,blocks', [(4, 1), (1000, 1), (66688, 0), (1048576 + 12, 0), (300000, 7)])
def test_emulated_allreduce_matches_fp64_sum(world, algo, n, blocks):
    dev = torch.device('cuda', 0)
    g = EmulatedPeerGroup(world, dev, heap_bytes=(2 * world + 2) * max(n, 1 << 20) * 4 + (16 << 20))
    off, bufs = g.alloc(n)
    torch.manual_seed(n + world)
    src = [torch.randn(n, device=dev) for _ in range(world)]
    for b, s in zip(bufs, src):
        b.copy_(s)
    g.allreduce(off, n, algo, scale=1.0 / world, blocks=blocks)
    torch.cuda.synchronize()
    ref = (torch.stack(src).double().sum(0) / world).float()
    for r in range(world):
        torch.testing.assert_close(bufs[r], ref, rtol=1e-5, atol=1e-6)
    for r in range(1, world):                       # rank-ordered summation => bit-identical replicas
        assert torch.equal(bufs[0], bufs[r])


@pytest.mark.parametrize('world', [2, 8])
@pytest.mark.parametrize('n,blocks', [(4, 1), (1000, 0), (65536, 0), (50000, 3)])
@pytest.mark.parametrize('wire_bf16', [False, True])
def test_emulated_flag_in_data_allreduce(world, n, blocks, wire_bf16):
    """One-shot with 8-byte {payload, epoch} packets: no flag round, no system fence."""
    dev = torch.device('cuda', 0)
    g = EmulatedPeerGroup(world, dev)
    off, bufs = g.alloc(n)
    for it in range(6):                             # both parities, increasing epochs over the same slots
        torch.manual_seed(n + world + it)
        src = [torch.randn(n, device=dev) for _ in range(world)]
        for b, s in zip(bufs, src):
            b.copy_(s)
        g.allreduce(off, n, AR_ONESHOT_LL, scale=0.5, wire_bf16=wire_bf16, blocks=blocks)
        torch.cuda.synchronize()
        wire = [s.bfloat16().float() if wire_bf16 else s for s in src]
        ref = (torch.stack(wire).double().sum(0) * 0.5).float()
        torch.testing.assert_close(bufs[world - 1], ref, rtol=1e-5, atol=1e-5)
        for r in range(1, world):
            assert torch.equal(bufs[0], bufs[r])


def test_emulated_allreduce_mixed_sizes_keep_parity_halves_apart():
    """A small call followed by a larger one must not share staging bytes (a fast rank may already push the second
    while a slow one still reads the first): the parity halves are fixed, not derived from the message size."""
    dev = torch.device('cuda', 0)
    world = 4
    g = EmulatedPeerGroup(world, dev)
    offs = [g.alloc(n) for n in (1024, 300000, 4096, 200000)]
    for it in range(20):
        for (off, bufs), n in zip(offs, (1024, 300000, 4096, 200000)):
            for r, b in enumerate(bufs):
                b.fill_(float(r + 1 + it))
            g.allreduce(off, n, AR_ONESHOT if it % 2 else AR_TWOSHOT, scale=1.0)
        torch.cuda.synchronize()
        want = float(sum(r + 1 + it for r in range(world)))
        for off, bufs in offs:
            assert float(bufs[0].min()) == want and float(bufs[world - 1].max()) == want


def test_emulated_allreduce_bf16_wire_and_back_to_back_calls():
    dev = torch.device('cuda', 0)
    world, n = 4, 50000
    g = EmulatedPeerGroup(world, dev)
    off, bufs = g.alloc(n)
    off2, bufs2 = g.alloc(4096)
    torch.manual_seed(0)
    for it in range(60):                            # alternating sizes/algorithms reuse flags and staging parity
        src = [torch.randn(n, device=dev) for _ in range(world)]
        for b, s in zip(bufs, src):
            b.copy_(s)
        algo = AR_ONESHOT if it % 3 else AR_TWOSHOT
        g.allreduce(off, n, algo, scale=1.0, wire_bf16=(it % 2 == 0))
        small = [torch.full((4096,), float(r + it), device=dev) for r in range(world)]
        for b, s in zip(bufs2, small):
            b.copy_(s)
        g.allreduce(off2, 4096, AR_ONESHOT, scale=1.0)
        torch.cuda.synchronize()
        if it % 2 == 0:
            ref = torch.stack([s.bfloat16().float() for s in src]).sum(0)
            tol = dict(rtol=1e-5, atol=1e-5)
        else:
            ref = torch.stack(src).sum(0)
            tol = dict(rtol=1e-5, atol=1e-5)
        torch.testing.assert_close(bufs[1], ref, **tol)
        assert torch.equal(bufs[0], bufs[3])
        assert float(bufs2[2][0]) == sum(r + it for r in range(world))


def test_emulated_flag_protocol_stress_100k_back_to_back_calls():
    """Race / reuse stress of the flag protocols: 10^5 back-to-back small allreduces with no host
    synchronisation in between, alternating two buffers and the three small-message algorithms, so every flag word,
    both staging parities and both packet halves are reused ~10^4 times; results are checked every 10,000 calls."""
    dev = torch.device('cuda', 0)
    world, n = 4, 1024
    g = EmulatedPeerGroup(world, dev)
    sets = [g.alloc(n) for _ in range(2)]
    src = [torch.full((n,), float(r + 1), device=dev) for r in range(world)]
    want = float(sum(range(1, world + 1)))
    algos = (AR_ONESHOT_LL, AR_ONESHOT, AR_TWOSHOT)
    for it in range(100000):
        off, bufs = sets[it & 1]
        for b, s_ in zip(bufs, src):
            b.copy_(s_, non_blocking=True)
        g.allreduce(off, n, algos[it % 3], scale=1.0)
        if it % 10000 == 9999:
            torch.cuda.synchronize()
            for b in bufs:
                assert float(b.min()) == want == float(b.max()), (it, float(b.min()), float(b.max()))


@pytest.mark.parametrize('world', [2, 8])
def test_emulated_broadcast_every_root_and_allgather(world):
    dev = torch.device('cuda', 0)
    g = EmulatedPeerGroup(world, dev)
    n = 70000
    off, bufs = g.alloc(n)
    for root in range(world):
        for r, b in enumerate(bufs):
            b.fill_(float(r + 1 + 10 * root))
        g.broadcast(off, n * 4, root)
        torch.cuda.synchronize()
        for b in bufs:
            assert float(b.min()) == float(b.max()) == float(root + 1 + 10 * root)
    chunk = 4096
    in_off, ins = g.alloc(chunk // 4)
    out_off, outs = g.alloc(chunk // 4 * world)
    for r, b in enumerate(ins):
        b.fill_(float(r))
    g.allgather(in_off, out_off, chunk)
    g.barrier()
    torch.cuda.synchronize()
    ref = torch.arange(world, device=dev, dtype=torch.float32).repeat_interleave(chunk // 4)
    for o in outs:
        assert torch.equal(o, ref)


@pytest.mark.parametrize('world', [2, 4, 8])
@pytest.mark.parametrize('kind', ['adam', 'adagrad', 'sgd'])
def test_emulated_sharded_dp_step_matches_allreduce_then_optimizer(world, kind):
    """ddls_comm_dp_step (reduce-scatter -> optimizer on the owner's slice -> parameter all-gather, one kernel) against
    a plain fp32 restatement: average the ranks' gradients, apply the Keras update."""
    from distributeddeeplearningspark_b200.ops.optimizer import make_specs
    dev = torch.device('cuda', 0)
    g = EmulatedPeerGroup(world, dev)
    n_opt, tail, elem0 = 50000, 8, 128
    n = n_opt + tail
    goff, grads = g.alloc(n)
    poff, params = g.alloc(n_opt)
    s1off, s1 = g.alloc(n_opt)
    s2off, s2 = g.alloc(n_opt)
    step = torch.zeros(2, dtype=torch.int32, device=dev)
    rows, cols, ld, mbeg = 37, 225, 232, 1000
    shadow = torch.zeros((rows, ld), dtype=torch.bfloat16, device=dev)
    specs, nspecs = make_specs([(elem0 + mbeg, rows, cols, ld, shadow)])
    torch.manual_seed(world)
    p_ref = torch.randn(n_opt, device=dev)
    for p in params:
        p.copy_(p_ref)
    m_ref, v_ref = torch.zeros_like(p_ref), torch.zeros_like(p_ref)
    lr, b1, b2, eps = (1e-3, 0.9, 0.999, 1e-7) if kind == 'adam' else (0.01, 0.0, 0.0, 1e-7)
    k = {'adam': 1, 'adagrad': 2, 'sgd': 3}[kind]
    for t in range(1, 4):
        src = [torch.randn(n, device=dev) for _ in range(world)]
        for b, s_ in zip(grads, src):
            b.copy_(s_)
        g.dp_step(goff, poff, s1off, s2off, n, n_opt, elem0, 1.0 / world, k, step, lr, b1, b2, eps, zero_grads=True,
                  advance=True, specs=specs, nspecs=nspecs, blocks=(3 if t == 2 else 0))
        torch.cuda.synchronize()
        gm = (torch.stack(src).double().sum(0) / world).float()
        ga = gm[:n_opt]
        if kind == 'adam':
            m_ref = b1 * m_ref + (1 - b1) * ga
            v_ref = b2 * v_ref + (1 - b2) * ga * ga
            lr_t = lr * (1 - b2 ** t) ** 0.5 / (1 - b1 ** t)
            p_ref = p_ref - lr_t * m_ref / (v_ref.sqrt() + eps)
        elif kind == 'adagrad':
            m_ref = m_ref + ga * ga
            p_ref = p_ref - lr * ga / (m_ref.sqrt() + eps)
        else:
            p_ref = p_ref - lr * ga
        assert int(step[0]) == t
        for r in range(world):
            torch.testing.assert_close(params[r], p_ref, rtol=1e-5, atol=1e-6)
            assert torch.equal(params[r], params[0])                      # replicas bit-identical
            assert float(grads[r][:n_opt].abs().max()) == 0.0             # cleared on every rank
            torch.testing.assert_close(grads[r][n_opt:], gm[n_opt:], rtol=1e-5, atol=1e-6)   # metric tail: reduced, kept
        # slots are valid on the owner of each slice
        nvec = n // 4
        per = (nvec + world - 1) // world
        for r in range(world):
            lo, hi = min(nvec, r * per) * 4, min(min(nvec, (r + 1) * per) * 4, n_opt)
            if kind != 'sgd':
                torch.testing.assert_close(s1[r][lo:hi], m_ref[lo:hi], rtol=1e-4, atol=1e-6)
            if kind == 'adam':
                torch.testing.assert_close(s2[r][lo:hi], v_ref[lo:hi], rtol=1e-4, atol=1e-7)
        want = params[0][mbeg - 0:mbeg + rows * cols].view(rows, cols).bfloat16()
        assert torch.equal(shadow[:, :cols], want) and float(shadow[:, cols:].float().abs().max()) == 0.0


def test_emulated_sharded_dp_step_two_buckets_share_one_step_counter():
    """Two bucket kernels per training step: only the last one publishes step+1, both see the same Adam t."""
    dev = torch.device('cuda', 0)
    world = 4
    g = EmulatedPeerGroup(world, dev)
    n = 4096
    offs = [g.alloc(n) for _ in range(4)]         # grads, params, m, v
    step = torch.zeros(2, dtype=torch.int32, device=dev)
    for t in range(1, 4):
        for b in offs[0][1]:
            b.fill_(1.0)
        half = n // 2
        for lo, adv in ((half, False), (0, True)):
            o = [x[0] + 4 * lo for x in offs]
            g.dp_step(o[0], o[1], o[2], o[3], half, half, lo, 1.0 / world, 1, step, 1e-3, 0.9, 0.999, 1e-7, advance=adv)
        torch.cuda.synchronize()
        assert int(step[0]) == t
        p = offs[1][1][0]
        assert float((p[:half] - p[half:]).abs().max()) == 0.0          # both halves took the same bias correction


# --------------------------------------------------------------------------------------------------------
# real multi-process, multi-GPU
# --------------------------------------------------------------------------------------------------------
def _multi_rank_checks():
    import distributeddeeplearningspark_b200 as ddls
    from distributeddeeplearningspark_b200.parallel.peer import AR_NVLS, AR_ONESHOT, AR_ONESHOT_LL, AR_TWOSHOT, PeerMemoryComm
    ddls.init()
    r, n = ddls.rank(), ddls.size()
    comm = ddls.runtime.communicator()
    assert isinstance(comm, PeerMemoryComm), type(comm)
    dev = ddls.device()
    res = {'rank': r, 'mc': bool(comm.mc_base)}
    algos = [AR_ONESHOT, AR_TWOSHOT] + ([AR_NVLS] if comm.mc_base else [])
    for numel in (4, 66688, 1048576, 12 * 1048576):
        t = comm.alloc(numel)
        for algo in algos + ([AR_ONESHOT_LL] if numel * 4 <= 256 * 1024 else []):
            if algo in (AR_ONESHOT, AR_TWOSHOT) and 2 * n * numel * 4 > (256 << 20):
                continue                                # push staging (2 x ranks x message) would not fit the default heap
            torch.manual_seed(100 + r)
            t.copy_(torch.randn(numel, device=dev))
            mine = t.clone()
            comm.allreduce_(t, ddls.Sum, algo=algo)
            gathered = comm.allgather(mine)          # this library's own all-gather (chunked through the heap)
            ref = gathered.double().sum(0).float()
            err = float((t - ref).abs().max())
            assert err < 1e-4 * max(1.0, float(ref.abs().max())), (numel, algo, err)
            chk = comm.allgather(t)
            assert all(torch.equal(chk[0], c) for c in chk), ('replicas differ', numel, algo)
    # flag-in-data packets pushed through the multicast mapping (one store reaches every peer): many back-to-back calls
    # over both parities, checked against the per-peer push
    if comm.mc_base:
        t = comm.alloc(16384 + 4)
        bad = 0
        for it in range(300):
            torch.manual_seed(1000 + 17 * it + r)
            src = torch.randn(t.numel(), device=dev)
            comm._ll_mc_push = True
            t.copy_(src); comm.allreduce_(t, ddls.Sum, algo=AR_ONESHOT_LL); a = t.clone()
            comm._ll_mc_push = False
            t.copy_(src); comm.allreduce_(t, ddls.Sum, algo=AR_ONESHOT_LL)
            bad += int(not torch.equal(a, t))
        res['ll_mc_push_mismatches'] = bad
        comm._ll_mc_push = False
    # bf16 on the wire through the switch (NVLS): contributions rounded to bf16, fp32 accumulation, bf16 result
    if comm.mc_base:
        for numel in (1048576, 3 * 1048576 + 8):
            t = comm.alloc(numel)
            vals = []
            for rr in range(n):
                torch.manual_seed(4000 + rr)
                vals.append(torch.randn(numel, device=dev))
            t.copy_(vals[r])
            comm.allreduce_(t, ddls.Average, algo=AR_NVLS, wire_bf16=True)
            ref = (torch.stack([v.bfloat16().float() for v in vals]).sum(0) / n)
            err = float((t - ref).abs().max())
            assert err < 2e-2 * max(1.0, float(ref.abs().max())), ('nvls bf16', numel, err)
            assert torch.equal(t, t.bfloat16().float())       # the result really is a bf16 value
    # non-heap tensor path + average + odd size
    x = torch.full((1001,), float(r + 1), device=dev)
    ddls.allreduce_(x)
    assert abs(float(x[0]) - (n + 1) / 2.0) < 1e-6
    # broadcast from every root (heap and non-heap tensors)
    for root in range(n):
        b = comm.alloc(4096); b.fill_(float(r))
        comm.broadcast_(b, root)
        assert float(b.min()) == float(b.max()) == float(root)
        y = torch.full((777,), float(r), device=dev)
        ddls.broadcast_(y, root)
        assert float(y.max()) == float(root) == float(y.min())
    g = ddls.allgather(torch.tensor([[float(r), 2.0 * r]], device=dev))
    assert g.tolist() == [[float(i), 2.0 * i] for i in range(n)]
    # asynchronous ADAG center in multicast memory: center = init + sum over ranks/commits of scale*(W - pulled)
    from distributeddeeplearningspark_b200.trainers import MulticastCenter
    wflat = torch.full((4096,), 1.0, device=dev)
    center = MulticastCenter(comm, wflat, staleness_aware=False)
    for k in range(3):
        wflat += float(r + 1)                       # "local training" moved the weights by (r+1)
        center.commit_pull(wflat, 0.5)              # commit 0.5*(r+1) to every replica, then pull
    center.finalize()
    expect = 1.0 + 3 * 0.5 * sum(range(1, n + 1))
    assert abs(float(center.center.min()) - expect) < 1e-4 and abs(float(center.center.max()) - expect) < 1e-4
    assert center.total_updates() == 3 * n
    # synchronous training: identical weights on every rank after a few steps
    model = ddls.models.nyiso_lstm(seed=r).to(dev)
    opt = ddls.optim.KerasAdam(model.parameters(), comm=comm)
    ddls.broadcast_parameters(opt, 0)
    dopt = ddls.DistributedOptimizer(opt)
    from distributeddeeplearningspark_b200.trainers import StepEngine
    eng = StepEngine(model, optimizer=dopt, loss='mean_squared_error', device=dev, use_cuda_graph=True)
    torch.manual_seed(r)
    for _ in range(8):
        eng.step(torch.rand(32, 25, 1, device=dev), torch.rand(32, 1, device=dev))
    w = opt.flat.flat_params.clone()
    ws = comm.allgather(w)
    res['weights_equal'] = all(torch.equal(ws[0], v) for v in ws)
    res['graph'] = eng._graph is not None
    res['fused'] = bool(dopt._fused)
    # fused allreduce + optimizer kernel == allreduce kernel followed by the optimizer kernel (same gradients fed to both;
    # end-to-end training runs are not compared because the recurrent kernels' atomics make tiny gradients differ
    # from run to run and Adagrad's g/(sqrt(a)+eps) amplifies that)
    import os
    finals = {}
    for fused in ('1', '0'):
        os.environ['DDLS_FUSED_OPT'] = fused
        for kind in ('adam', 'adagrad', 'sgd'):
            torch.manual_seed(5)
            ps = [torch.nn.Parameter(torch.randn(1000, device=dev)), torch.nn.Parameter(torch.randn(37, 5, device=dev))]
            slots = 4 if kind == 'adam' else 0            # K20: metric slots ride in the (single) bucket
            o2 = ddls.optim.get_optimizer(kind, ps, comm=comm, metric_slots=slots)
            o2.zero_grads_in_step = True
            d2 = ddls.DistributedOptimizer(o2)
            assert bool(d2._fused) == (fused == '1')
            for k in range(5):
                torch.manual_seed(1000 * k + r)
                o2.flat.flat_grads[:o2.flat.numel].copy_(torch.randn(o2.flat.numel, device=dev))
                if slots:
                    d2.metrics[0] = float(r + k)
                d2.step()
                if slots:
                    assert abs(float(d2.metrics[0]) - (k + (n - 1) / 2.0)) < 1e-5, (fused, float(d2.metrics[0]))
            torch.cuda.synchronize()
            assert int(o2.step_count.item()) == 5 and float(o2.flat.flat_grads[:o2.flat.numel].abs().max()) == 0.0
            finals[(fused, kind)] = o2.flat.flat_params.clone()
    os.environ.pop('DDLS_FUSED_OPT')
    res['fused_err'] = {k: float((finals[('1', k)] - finals[('0', k)]).abs().max()) for k in ('adam', 'adagrad', 'sgd')}
    res['fused_matches'] = all(v < 1e-6 for v in res['fused_err'].values())
    torch.cuda.synchronize()
    res.update(_sharded_step_checks(ddls, comm, dev, r, n))
    return res


def _sharded_step_checks(ddls, comm, dev, r, n):
    """Multi-bucket fused data parallelism: the big bucket runs the sharded kernel (multimem reduce-scatter -> optimizer on
    the owner's slice -> multicast parameter all-gather), the small one the flag-in-data kernel; compared with the unfused
    path (allreduce kernels + one optimizer kernel) on identical gradients."""
    import os
    res = {}
    finals, slots = {}, {}
    for fused in ('1', '0'):
        os.environ['DDLS_FUSED_OPT'] = fused
        for kind in ('adam', 'adagrad'):
            torch.manual_seed(7)
            ps = [torch.nn.Parameter(torch.randn(3000, device=dev)), torch.nn.Parameter(torch.randn(37, 5, device=dev)),
                  torch.nn.Parameter(torch.randn(2600, 225, device=dev))]      # 585,000 elements: its own >2 MB bucket
            o2 = ddls.optim.get_optimizer(kind, ps, comm=comm, metric_slots=4)
            o2.zero_grads_in_step = True
            shadow = o2.add_shadow(ps[2], 232)
            d2 = ddls.DistributedOptimizer(o2)
            assert len(d2.buckets) == 2 and bool(d2._fused) == (fused == '1')
            for k in range(4):
                torch.manual_seed(1000 * k + r)
                o2.flat.flat_grads[:o2.flat.numel].copy_(torch.randn(o2.flat.numel, device=dev))
                d2.metrics[1] = float(r + k)
                d2.step()
                assert abs(float(d2.metrics[1]) - (k + (n - 1) / 2.0)) < 1e-5
            torch.cuda.synchronize()
            if fused == '1':
                assert [d2._bucket_mode(b) for b in d2.buckets][0] == 'dp', [d2._bucket_mode(b) for b in d2.buckets]
            assert int(o2.step_count.item()) == 4 and float(o2.flat.flat_grads[:o2.flat.numel].abs().max()) == 0.0
            assert torch.equal(shadow[:, :225], ps[2].detach().bfloat16())
            sd = d2.state_dict()                        # collective: gathers the sharded slots
            finals[(fused, kind)] = o2.flat.flat_params.clone()
            slots[(fused, kind)] = {k_: v_.to(dev) for k_, v_ in sd['slots'].items()}
            ws = comm.allgather(finals[(fused, kind)])
            assert all(torch.equal(ws[0], w_) for w_ in ws), 'replicas differ after the sharded step'
    os.environ.pop('DDLS_FUSED_OPT')
    res['sharded_err'] = {k: float((finals[('1', k)] - finals[('0', k)]).abs().max()) for k in ('adam', 'adagrad')}
    res['sharded_slot_err'] = {k: max(float((slots[('1', k)][s_] - slots[('0', k)][s_]).abs().max()) for s_ in slots[('1', k)])
                               for k in ('adam', 'adagrad')}
    res['sharded_matches'] = all(v < 1e-6 for v in res['sharded_err'].values()) and \
        all(v < 1e-6 for v in res['sharded_slot_err'].values())
    # the MNIST CNN end to end: two buckets (sharded + flag-in-data), CUDA graph, weights identical on every rank
    from distributeddeeplearningspark_b200.trainers import StepEngine
    model = ddls.models.mnist_cnn(seed=r).to(dev)
    opt = ddls.optim.KerasAdam(model.parameters(), comm=comm)
    ddls.broadcast_parameters(opt, 0)
    dopt = ddls.DistributedOptimizer(opt)
    eng = StepEngine(model, optimizer=dopt, loss='categorical_crossentropy', device=dev, use_cuda_graph=True)
    torch.manual_seed(100 + r)
    losses = []
    x = torch.rand(128, 28, 28, 1, device=dev)
    y = torch.nn.functional.one_hot(torch.randint(0, 10, (128,), device=dev), 10).float()
    for _ in range(12):
        losses.append(float(eng.step(x, y)))
    w = opt.flat.flat_params.clone()
    ws = comm.allgather(w)
    res['cnn_weights_equal'] = all(torch.equal(ws[0], v) for v in ws)
    res['cnn_modes'] = [dopt._bucket_mode(b) for b in dopt.buckets]
    res['cnn_loss'] = (losses[0], losses[-1])
    res['cnn_steps'] = int(opt.step_count.item())
    res['cnn_shadow_ok'] = bool(torch.equal(eng._direct.w1s[:, :225], model.layers_list[6].kernel.detach().bfloat16()))
    # checkpoint on CUDA ranks with sharded optimizer slots: rank 0 saves (the slots are gathered first), a fresh model +
    # optimizer on every rank resumes (load on rank 0, broadcast), and one more step of both runs lands on the same weights
    import tempfile
    from distributeddeeplearningspark_b200.ckpt import resume, save_checkpoint
    path = ddls.broadcast_object(os.path.join(tempfile.mkdtemp(prefix='ddls_ckpt_'), 'cnn') if r == 0 else None, 0)
    save_checkpoint(path, model, dopt, step=12)
    ddls.barrier(); torch.cuda.synchronize()
    model2 = ddls.models.mnist_cnn(seed=50 + r).to(dev)
    opt2 = ddls.optim.KerasAdam(model2.parameters(), comm=comm)
    dopt2 = ddls.DistributedOptimizer(opt2)
    eng2 = StepEngine(model2, optimizer=dopt2, loss='categorical_crossentropy', device=dev, use_cuda_graph=False)
    eng2.step(x, y)                                   # builds the fused step and its bf16 shadows from the wrong weights
    meta = resume(path, model2, dopt2)
    res['resume_step'] = int(meta.get('step') or -1), int(opt2.step_count.item())
    la, lb = float(eng.step(x, y)), float(eng2.step(x, y))
    torch.cuda.synchronize()
    res['resume_loss'] = (la, lb)
    res['resume_err'] = float((opt.flat.flat_params - opt2.flat.flat_params).abs().max())
    res['resume_ok'] = abs(la - lb) < 1e-4 * max(1.0, abs(la)) and res['resume_err'] < 5e-5
    return res


@pytest.mark.multigpu
def test_peer_memory_comm_multi_gpu():
    from distributeddeeplearningspark_b200 import spark
    from distributeddeeplearningspark_b200.spark import LocalBarrierBackend
    n = min(torch.cuda.device_count(), 8)
    out = spark.run(_multi_rank_checks, num_proc=n, backend=LocalBarrierBackend(timeout=600))
    assert [o['rank'] for o in out] == list(range(n))
    assert all(o['weights_equal'] for o in out) and all(o['graph'] for o in out)
    assert all(o['fused'] for o in out) and all(o['fused_matches'] for o in out), [o.get('fused_err') for o in out]
    assert all(o.get('ll_mc_push_mismatches', 0) == 0 for o in out), [o.get('ll_mc_push_mismatches') for o in out]
    assert all(o['sharded_matches'] for o in out), [(o.get('sharded_err'), o.get('sharded_slot_err')) for o in out]
    assert all(o['cnn_weights_equal'] and o['cnn_shadow_ok'] for o in out)
    assert all(o['cnn_modes'][0] == 'dp' and o['cnn_steps'] == 12 for o in out), [(o['cnn_modes'], o['cnn_steps']) for o in out]
    assert all(o['cnn_loss'][1] < o['cnn_loss'][0] for o in out), [o['cnn_loss'] for o in out]
    assert all(o['resume_ok'] and o['resume_step'] == (12, 12) for o in out), [(o['resume_step'], o['resume_loss'], o['resume_err']) for o in out]
    import json, os
    os.makedirs('gpurun_out', exist_ok=True)
    with open('gpurun_out/multigpu_checks.json', 'w') as f:
        json.dump(out, f, indent=1, default=str)


def _stalled_peer():
    """Rank 1 never joins the collective: rank 0's kernel must give up after DDLS_COMM_TIMEOUT_MS, not spin for ever."""
    import time
    import distributeddeeplearningspark_b200 as ddls
    ddls.init()
    comm = ddls.runtime.communicator()
    t = comm.alloc(1024)
    if ddls.rank() == 1:
        time.sleep(40)                                   # alive (the launcher sees no dead process) but silent
        return 'late'
    t0 = time.time()
    comm.allreduce_(t, ddls.Sum, algo=AR_ONESHOT)
    torch.cuda.synchronize()                             # the trapped kernel surfaces here as a CUDA error
    return time.time() - t0


@pytest.mark.multigpu
def test_silent_peer_trips_the_flag_timeout_and_surfaces_as_stage_error():
    """Failure detection on the device side: a flag wait that exceeds DDLS_COMM_TIMEOUT_MS prints rank / peer /
    epoch and traps; the launcher turns the resulting CUDA error of that rank into a BarrierStageError and tears the
    stalled peer down."""
    import time
    from distributeddeeplearningspark_b200 import spark
    from distributeddeeplearningspark_b200.spark import BarrierStageError, LocalBarrierBackend
    t0 = time.time()
    with pytest.raises(BarrierStageError, match='rank 0 failed'):
        spark.run(_stalled_peer, num_proc=2, backend=LocalBarrierBackend(timeout=120), env={'DDLS_COMM_TIMEOUT_MS': '2000'})
    assert time.time() - t0 < 35, 'the timeout did not fire before the stalled peer woke up'
