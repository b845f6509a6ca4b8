"""Fused peer-memory collectives over NVLink: needs >= 2 GPUs (gpurun --gpus 2)."""
import pytest
import torch

from tests.dist_utils import run_distributed

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]


def _reduce_worker(rank, world):
    import torch.distributed as dist
    from ddl25spring_b200.parallel.symm_impl import SymmetricArena, shard_bounds
    a = SymmetricArena()
    n, G = 1_200_000, 3
    rows = a.alloc("rows", G * n * 4).view(G, n)
    res = a.alloc("res", n * 4)
    g = torch.Generator(device="cuda").manual_seed(100 + rank)
    rows.copy_(torch.randn(G, n, generator=g, device="cuda"))
    allrows = [torch.empty(G, n, device="cuda") for _ in range(world)]
    dist.all_gather(allrows, rows.contiguous())
    U = torch.stack([allrows[k % world][k // world] for k in range(G * world)])
    K = G * world
    ptrs = [a.peer_ptrs("rows")[k % world] + (k // world) * n * 4 for k in range(K)]
    w = [(k + 1) / sum(range(1, K + 1)) for k in range(K)]
    d0, d1 = shard_bounds(n, world, rank)
    out = {}
    for rep in range(3):                      # repeated use: epochs / counters must keep working
        a.reduce(0, ptrs, w, a.peer_ptrs("res"), d0, d1)
    torch.cuda.synchronize(); a.check()
    exp = (U.double() * torch.tensor(w, device="cuda", dtype=torch.float64)[:, None]).sum(0).float()
    out["wsum_err"] = float((res - exp).abs().max())
    a.reduce(1, ptrs, w, a.peer_ptrs("res"), d0, d1, rank_lo=(K - 1) // 2 if K % 2 else K // 2 - 1, rank_hi=(K - 1) // 2 + 1 if K % 2 else K // 2 + 1)
    torch.cuda.synchronize(); a.check()
    s = U.sort(0).values
    med = s[(K - 1) // 2] if K % 2 else 0.5 * (s[K // 2 - 1] + s[K // 2])
    out["median_err"] = float((res - med).abs().max())
    # timing: device time of the fused all-reduce-like op, max over ranks
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.barrier(); torch.cuda.synchronize()
    e0.record()
    for _ in range(10):
        a.reduce(0, ptrs, w, a.peer_ptrs("res"), d0, d1)
    e1.record(); torch.cuda.synchronize(); a.check()
    t = torch.tensor([e0.elapsed_time(e1) / 10], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    out["fused_ms"] = float(t)
    # NCCL baseline for the same math: local weighted partial sum + all_reduce
    part = torch.empty(n, device="cuda")
    wl = torch.tensor([w[k] for k in range(K) if k % world == rank], device="cuda")
    torch.cuda.synchronize(); e0.record()
    for _ in range(10):
        torch.mv(rows.t(), wl, out=part)
        dist.all_reduce(part)
    e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / 10], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    out["nccl_ms"] = float(t)
    return out


@pytest.mark.timeout(300)
def test_symm_reduce_matches_oracle():
    world = min(torch.cuda.device_count(), 8)
    outs = run_distributed(_reduce_worker, world, timeout=240, backend="nccl")
    for o in outs:
        assert o["wsum_err"] < 1e-5 and o["median_err"] < 1e-6, o
    print("fused vs nccl ms:", outs[0]["fused_ms"], outs[0]["nccl_ms"])


def _dp_worker(rank, world, optimizer, graph, nvls):
    """DataParallel with the fused peer-memory reducer (reduce-scatter + optimizer on the owned shard + all-gather of the new
    parameters in one kernel per bucket; with graph=True captured as branches of the step graph) against the NCCL all-reduce +
    whole-arena optimizer kernel."""
    import copy
    import torch.nn.functional as F
    from ddl25spring_b200.parallel.dp import DataParallel
    from ddl25spring_b200.parallel.symm_impl import SymmetricArena, SymmetricGradReducer
    torch.manual_seed(0)
    m = torch.nn.Sequential(torch.nn.Linear(64, 512), torch.nn.ReLU(), torch.nn.Linear(512, 512), torch.nn.ReLU(),
                            torch.nn.Linear(512, 10)).cuda()
    m2 = copy.deepcopy(m)
    arena = SymmetricArena()
    n = sum(p.numel() for p in m.parameters()) + 1024
    try:
        red = SymmetricGradReducer(arena, n, use_nvls=nvls)
    except Exception as e:
        if nvls:
            return ("unsupported", repr(e))
        raise
    kw = dict(bucket_mb=0.25, optimizer=optimizer, lr=0.05 if optimizer == "sgd" else 1e-3, weight_decay=0.01,
              momentum=0.9 if optimizer == "sgd" else 0.0)
    dp = DataParallel(m, reducer=red, graph=graph, **kw)
    ref = DataParallel(m2, **kw)                                              # NCCL all_reduce path
    assert dp.fused_step and not ref.fused_step
    g = torch.Generator().manual_seed(rank)
    x, y = torch.randn(32, 64, generator=g).cuda(), torch.randint(0, 10, (32,), generator=g).cuda()
    lf = lambda mod, a, b: F.cross_entropy(mod(a), b)
    l1 = [float(dp.train_step(lf, x, y)) for _ in range(5)]
    l2 = [float(ref.train_step(lf, x, y)) for _ in range(5)]
    torch.cuda.synchronize(); arena.check()
    werr = float((dp.arena.flat - ref.arena.flat).abs().max())
    gleft = float(dp.arena.grad.abs().max())                                 # the fused kernels re-zero the gradient arena
    return ("ok", werr, len(dp.buckets), max(abs(a - b) for a, b in zip(l1, l2)), gleft)


@pytest.mark.timeout(400)
@pytest.mark.parametrize("optimizer,graph,nvls", [("sgd", False, False), ("adam", False, False), ("adamw", True, False),
                                                  ("sgd", True, False), ("sgd", True, True)])
def test_dp_with_fused_reducer_matches_nccl(optimizer, graph, nvls):
    world = min(torch.cuda.device_count(), 8)
    outs = run_distributed(_dp_worker, world, (optimizer, graph, nvls), timeout=300, backend="nccl")
    if outs[0][0] == "unsupported":
        pytest.skip("NVLS multicast not available: " + outs[0][1])
    for _, err, nb, lerr, gleft in outs:
        assert err < 2e-5 and nb > 1 and lerr < 1e-4 and gleft == 0.0, (err, nb, lerr, gleft)
    assert len({o[1] for o in outs}) >= 1


def _wa_worker(rank, world):
    import copy
    import torch.nn.functional as F
    from ddl25spring_b200.parallel.dp import WeightAveraging
    from ddl25spring_b200.parallel.symm_impl import SymmetricArena, SymmetricWeightAverager
    torch.manual_seed(0)
    m = torch.nn.Sequential(torch.nn.Linear(32, 256), torch.nn.ReLU(), torch.nn.Linear(256, 10)).cuda()
    m2 = copy.deepcopy(m)
    arena = SymmetricArena()
    n = sum(p.numel() for p in m.parameters()) + 512
    wa = WeightAveraging(m, lr=0.05, averager=SymmetricWeightAverager(arena, n))
    ref = WeightAveraging(m2, lr=0.05)
    g = torch.Generator().manual_seed(rank)
    x, y = torch.randn(16, 32, generator=g).cuda(), torch.randint(0, 10, (16,), generator=g).cuda()
    for _ in range(3):
        for w in (wa, ref):
            F.cross_entropy(w.module(x), y).backward()
            w.step()
    torch.cuda.synchronize(); arena.check()
    return float((wa.arena.flat - ref.arena.flat).abs().max()), wa.arena.flat.clone().cpu()


@pytest.mark.timeout(300)
def test_weight_averaging_fused_matches_nccl():
    world = min(torch.cuda.device_count(), 8)
    outs = run_distributed(_wa_worker, world, timeout=240, backend="nccl")
    for err, _ in outs:
        assert err < 1e-6, err
    torch.testing.assert_close(outs[0][1], outs[-1][1], rtol=0, atol=0)       # replicas identical after averaging


def _timeout_worker(rank, world):
    """A rank that never launches its kernel: the others must time out, raise, and publish nothing."""
    from ddl25spring_b200.parallel.symm_impl import SymmetricArena, shard_bounds
    a = SymmetricArena()
    n = 4096
    rows = a.alloc("rows", n * 4)
    res = a.alloc("res", n * 4)
    rows.fill_(1.0 + rank)
    res.fill_(-7.0)
    torch.cuda.synchronize()
    import torch.distributed as dist
    dist.barrier()
    raised = False
    if rank != 0:                                   # rank 0 "crashed": it never joins the collective
        d0, d1 = shard_bounds(n, world, rank)
        a.reduce(0, a.peer_ptrs("rows"), [1.0] * world, a.peer_ptrs("res"), d0, d1)
        torch.cuda.synchronize()
        try:
            a.check()
        except RuntimeError:
            raised = True
    dist.barrier()
    torch.cuda.synchronize()
    return raised, float(res.min()), float(res.max())


@pytest.mark.timeout(300)
def test_missing_peer_times_out_loudly_and_publishes_nothing():
    world = min(torch.cuda.device_count(), 8)
    outs = run_distributed(_timeout_worker, world, timeout=240, backend="nccl")
    assert all(o[0] for o in outs[1:]) and not outs[0][0]
    for _, lo, hi in outs:
        assert lo == -7.0 and hi == -7.0          # nobody's result buffer was touched


def _fedavg_worker(rank, world, agg_name):
    from ddl25spring_b200.core.data import SyntheticMNIST, split
    from ddl25spring_b200.fl.aggregators import make_aggregator
    from ddl25spring_b200.fl.packed_server import PackedFedAvgServer
    from ddl25spring_b200.parallel.symm import make_collective
    ds = SyntheticMNIST(n_train=2400, n_test=400, seed=2, noise=0.8)
    subsets = split(ds.train, 12, True, 3)
    kw = dict(n_byzantine=1) if agg_name == "krum" else {}
    outs = {}
    for name, coll in (("fused", make_collective(two_phase_select=(agg_name == "median2"))), ("nccl", None)):
        srv = PackedFedAvgServer(0.05, 50, subsets, 0.5, 1, 3, test_set=ds.test, device="cuda", p_drop1=0.0, p_drop2=0.0,
                                 aggregator=make_aggregator(agg_name.rstrip("2"), **kw), collective=coll)
        res = srv.run(2)
        torch.cuda.synchronize()
        outs[name] = (srv.arena.flat.clone().cpu(), res.test_accuracy)
    return outs


@pytest.mark.timeout(400)
@pytest.mark.parametrize("agg_name", ["mean", "median", "median2", "krum"])
def test_packed_fedavg_fused_equals_nccl_path(agg_name):
    world = min(torch.cuda.device_count(), 8)
    outs = run_distributed(_fedavg_worker, world, (agg_name,), timeout=300, backend="nccl")
    for o in outs:
        torch.testing.assert_close(o["fused"][0], o["nccl"][0], rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(outs[0]["fused"][0], outs[-1]["fused"][0], rtol=0, atol=0)


def _vfl_worker(rank, world):
    from ddl25spring_b200.core.data import SyntheticHeart
    from ddl25spring_b200.parallel import vfl
    from ddl25spring_b200.parallel.symm_impl import FusedVFLGather, SymmetricArena
    tab = SyntheticHeart().encoded()
    parts = vfl.balanced_partition(tab.columns, world)
    idx = tab.column_index(parts[rank])
    x_local, y = tab.x[:256, idx].cuda(), tab.y[:256].cuda()
    dims = [2 * len(p) for p in parts]
    arena = SymmetricArena()
    out = {}
    for name in ("fused", "allgather"):
        hook = FusedVFLGather(arena, 256, dims, 128) if name == "fused" else None
        net = vfl.DistributedVFL(len(idx), dims, act_on_logits=False, seed=3, gather_linear=hook).cuda()
        net.bottom.dropout.p = 0.0
        opt = torch.optim.SGD(net.parameters(), lr=0.1)
        losses = []
        for _ in range(4):
            opt.zero_grad()
            loss = torch.nn.functional.cross_entropy(net(x_local), y)
            loss.backward()
            opt.step()
            losses.append(float(loss))
        torch.cuda.synchronize(); arena.check()
        out[name] = (losses, net.top_model.fc1.weight.detach().cpu().clone(), net.bottom.fc1.weight.detach().cpu().clone())
    return out


@pytest.mark.timeout(300)
def test_vfl_fused_gather_gemm_equals_allgather_path():
    world = min(torch.cuda.device_count(), 8)
    outs = run_distributed(_vfl_worker, world, timeout=240, backend="nccl")
    for o in outs:
        assert o["fused"][0] == pytest.approx(o["allgather"][0], rel=2e-4)
        torch.testing.assert_close(o["fused"][1], o["allgather"][1], rtol=1e-3, atol=1e-5)
        torch.testing.assert_close(o["fused"][2], o["allgather"][2], rtol=1e-3, atol=1e-5)


def _generic_worker(rank, world, agg_name, with_bn):
    import torch.nn.functional as Fn
    from ddl25spring_b200.core.data import SyntheticMNIST, split
    from ddl25spring_b200.fl.aggregators import make_aggregator
    from ddl25spring_b200.fl.attacks import SignFlip
    from ddl25spring_b200.fl.server import FedAvgServer
    from ddl25spring_b200.parallel.symm_impl import SymmetricAggregator
    import numpy as np

    def model_fn():
        layers = [torch.nn.Flatten(), torch.nn.Linear(784, 64)]
        if with_bn:
            layers.append(torch.nn.BatchNorm1d(64))
        return torch.nn.Sequential(*layers, torch.nn.ReLU(), torch.nn.Linear(64, 10))
    ds = SyntheticMNIST(n_train=1600, n_test=400, seed=2, noise=0.8)
    subsets = split(ds.train, 8, True, 3)
    mal = np.zeros(8, dtype=bool); mal[1] = True
    kw = dict(n_byzantine=1, multi=4) if agg_name == "krum" else {}
    outs = {}
    coll = None
    for name in ("fused", "nccl"):
        srv = FedAvgServer(0.05, 50, subsets, 1.0, 1, 3, model_fn=model_fn, test_set=ds.test, device="cuda", include_buffers=with_bn,
                           loss_fn=lambda o, t: Fn.cross_entropy(o, t), aggregator=make_aggregator(agg_name, **kw),
                           attack=SignFlip(4.0), malicious=mal)
        if name == "fused":
            coll = SymmetricAggregator(max_rows_per_rank=4, row_elems=srv.arena.size)
            srv.collective = coll
        res = srv.run(2)
        torch.cuda.synchronize()
        outs[name] = (srv.arena.flat.clone().cpu(), res.test_accuracy)
    return outs


@pytest.mark.timeout(400)
@pytest.mark.parametrize("agg_name,with_bn", [("mean", False), ("median", True), ("krum", True)])
def test_generic_server_fused_equals_nccl_path(agg_name, with_bn):
    world = min(torch.cuda.device_count(), 4)
    outs = run_distributed(_generic_worker, world, (agg_name, with_bn), timeout=300, backend="nccl")
    for o in outs:
        assert torch.isfinite(o["fused"][0]).all()
        torch.testing.assert_close(o["fused"][0], o["nccl"][0], rtol=1e-4, atol=1e-5)


def _rowgather_worker(rank, world):
    import torch.distributed as dist
    from ddl25spring_b200.ops.gemm import gram_rowgather_ptrs
    from ddl25spring_b200.parallel.symm_impl import SymmetricArena, shard_bounds
    a = SymmetricArena()
    G, n = 8, 200_000
    rows = a.alloc("rows", G * n * 4).view(G, n)
    g = torch.Generator(device="cuda").manual_seed(7 + rank)
    rows.copy_(0.5 + 0.01 * torch.randn(G, n, generator=g, device="cuda"))      # large common offset: the hard case for TF32
    allrows = [torch.empty(G, n, device="cuda") for _ in range(world)]
    dist.all_gather(allrows, rows.contiguous())
    U = torch.cat(allrows).double()                                            # p-major row order
    d0, d1 = shard_bounds(n, world, rank, align=32)
    out = torch.empty(G * world, G * world, device="cuda")
    a.barrier()
    gram_rowgather_ptrs(a.peer_ptrs("rows"), G, n, d0, d1, out, precise=True)
    a.barrier()
    torch.cuda.synchronize(); a.check()
    exp = U[:, d0:d1] @ U[:, d0:d1].T
    return float((out.double() - exp).abs().max() / exp.abs().max())


@pytest.mark.timeout(300)
def test_fused_allgather_gram_over_peers():
    world = min(torch.cuda.device_count(), 8)
    errs = run_distributed(_rowgather_worker, world, timeout=240, backend="nccl")
    assert max(errs) < 5e-5, errs


def _krum_fused_gather_worker(rank, world):
    from ddl25spring_b200.core.data import SyntheticMNIST, split
    from ddl25spring_b200.fl.aggregators import Krum
    from ddl25spring_b200.fl.packed_server import PackedFedAvgServer
    from ddl25spring_b200.parallel.symm import make_collective
    K = 8 * world
    ds = SyntheticMNIST(n_train=K * 100, n_test=200, seed=2, noise=0.8)
    subsets = split(ds.train, K, True, 3)
    sel = {}
    coll = make_collective()
    for name, fg in (("fused_gather", True), ("delta_path", False)):
        agg = Krum(n_byzantine=2, multi=K - 4, fused_gather=fg)
        srv = PackedFedAvgServer(0.05, 50, subsets, 1.0, 1, 3, test_set=ds.test, device="cuda", p_drop1=0.0, p_drop2=0.0,
                                 aggregator=agg, collective=coll)
        srv.run(1)
        torch.cuda.synchronize()
        sel[name] = (sorted(agg.last_selected.cpu().tolist()), srv.arena.flat.clone().cpu())
    return sel


@pytest.mark.timeout(400)
def test_krum_fused_allgather_gram_selects_like_delta_path():
    world = min(torch.cuda.device_count(), 8)
    outs = run_distributed(_krum_fused_gather_worker, world, timeout=300, backend="nccl")
    for o in outs:
        assert o["fused_gather"][0] == o["delta_path"][0]
        torch.testing.assert_close(o["fused_gather"][1], o["delta_path"][1], rtol=1e-5, atol=1e-6)


def _nvls_worker(rank, world):
    import torch.distributed as dist
    from ddl25spring_b200.parallel.nvls import NvlsAllReduce
    from ddl25spring_b200.parallel.symm_impl import SymmetricArena
    arena = SymmetricArena()
    n = 11_181_696
    nv = NvlsAllReduce.create(n, arena)
    if nv is None:
        return {"supported": False}
    g = torch.Generator(device="cuda").manual_seed(rank)
    x = torch.randn(n, generator=g, device="cuda")
    ref = x.clone()
    dist.all_reduce(ref)
    nv.buf[:n].copy_(x)
    nv.all_reduce_(0, n)
    torch.cuda.synchronize(); arena.check()
    err = float((nv.buf[:n] - ref).abs().max())
    out = {"supported": True, "err": err}
    for name, fn in (("nvls", lambda: nv.all_reduce_(0, n)), ("nccl", lambda: dist.all_reduce(ref))):
        for _ in range(3):
            fn()
        torch.cuda.synchronize(); dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            fn()
        e1.record(); torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / 10], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        out[name + "_ms"] = float(t)
    return out


@pytest.mark.timeout(300)
def test_nvls_allreduce_matches_nccl():
    world = min(torch.cuda.device_count(), 8)
    outs = run_distributed(_nvls_worker, world, timeout=240, backend="nccl")
    if not outs[0]["supported"]:
        pytest.skip("multicast (NVLS) not available on this box")
    for o in outs:
        assert o["err"] < 1e-3, o
    print("NVLS all-reduce 44.7 MB: ours %.3f ms, NCCL %.3f ms" % (outs[0]["nvls_ms"], outs[0]["nccl_ms"]))


def _pp_nccl_worker(rank, world, schedule, n_micro, graph=False):
    """Two-stage pipeline over NCCL p2p with the stage modules on the hand-written kernels."""
    from ddl25spring_b200.models.llama import LLama, causalLLMLoss
    from ddl25spring_b200.ops.nn import use_kernels
    from ddl25spring_b200.parallel import pp
    V, D, T = 512, 96, 64
    torch.manual_seed(0)
    model = LLama(None, V, D, 6, "cuda", 4, T)
    tokens = torch.randint(0, V, (n_micro, T), generator=torch.Generator().manual_seed(1)).cuda()
    mesh = pp.Mesh(world, 1)
    stage_mod = use_kernels(model.split_stages(world)[mesh.stage_id])
    st = pp.PipelineStage(stage_mod, mesh, (1, T, D), schedule=schedule, n_micro=n_micro,
                          loss_fn=lambda out, tgt: causalLLMLoss(out, tgt, V), device="cuda", graph=graph)
    mbs = list(tokens.split(1))
    st.zero_grad()
    losses = st.run(inputs=mbs, targets=mbs)
    if graph:                                           # replay once more: the graphs must be reusable across iterations
        st.zero_grad()
        losses = st.run(inputs=mbs, targets=mbs)
    torch.cuda.synchronize()
    # oracle on this rank: the unsplit model in plain PyTorch
    torch.manual_seed(0)
    ref = LLama(None, V, D, 6, "cuda", 4, T)
    loss = sum(causalLLMLoss(ref(t), t, V) for t in tokens.split(1)) / n_micro
    loss.backward()
    ref_stage = dict(ref.split_stages(world)[mesh.stage_id].named_parameters())
    worst = 0.0
    for n, p in stage_mod.named_parameters():
        g, r = p.grad, ref_stage[n].grad
        worst = max(worst, float((g - r).norm() / r.norm().clamp_min(1e-12)))
    return {"worst_rel": worst, "n_losses": len(losses), "stage": mesh.stage_id}


@pytest.mark.timeout(300)
@pytest.mark.parametrize("schedule,n_micro,graph", [("gpipe", 3, False), ("1f1b", 4, False), ("1f1b", 4, True)])
def test_pipeline_over_nccl_on_kernels_matches_unsplit_model(schedule, n_micro, graph):
    outs = run_distributed(_pp_nccl_worker, 2, (schedule, n_micro, graph), timeout=240, backend="nccl")
    for o in outs:
        assert o["worst_rel"] < 5e-3, o
    assert [o for o in outs if o["stage"] == 1][0]["n_losses"] == n_micro


def _pp_conv_worker(rank, world, which):
    """2-stage model-parallel split of the conv models (BASELINE config 5) over NCCL p2p, stages on the NHWC tcgen05 stack;
    gradients must equal the unsplit model's."""
    import torch.nn.functional as Fn
    from ddl25spring_b200.parallel import pp
    torch.manual_seed(0)
    n_micro, mb = 2, 16
    if which == "resnet":
        from ddl25spring_b200.models.resnet_nhwc import resnet18_nhwc, split_resnet18_nhwc_two_stage
        full = resnet18_nhwc().cuda()
        torch.manual_seed(0)
        ref = resnet18_nhwc().cuda()
        stages = split_resnet18_nhwc_two_stage(full)
        ref_stages = split_resnet18_nhwc_two_stage(ref)
        act_shape = (mb, 4, 4, 128)
        g = torch.Generator().manual_seed(1)
        xs = [torch.randn(mb, 3, 32, 32, generator=g).cuda() for _ in range(n_micro)]
        ys = [torch.randint(0, 10, (mb,), generator=g).cuda() for _ in range(n_micro)]
        loss_fn = lambda out, tgt: Fn.cross_entropy(out, tgt)
        fwd_full = lambda x: ref(x)
    else:
        from ddl25spring_b200.models.dcgan_nhwc import DiscriminatorNHWC
        full = DiscriminatorNHWC(16).cuda()
        torch.manual_seed(0)
        ref = DiscriminatorNHWC(16).cuda()
        stages, ref_stages = full.stages(), ref.stages()
        g = torch.Generator().manual_seed(1)
        xs = [(torch.rand(mb, 3, 64, 64, generator=g) * 2 - 1).cuda() for _ in range(n_micro)]
        ys = [torch.randint(0, 2, (mb,), generator=g).float().cuda() for _ in range(n_micro)]
        with torch.no_grad():
            act_shape = tuple(stages[0](xs[0]).shape)
        loss_fn = lambda out, tgt: Fn.binary_cross_entropy_with_logits(out.view(-1), tgt)
        fwd_full = lambda x: ref(x)
    mesh = pp.Mesh(2, 1)
    st = pp.PipelineStage(stages[mesh.stage_id], mesh, act_shape, schedule="1f1b", n_micro=n_micro, loss_fn=loss_fn, device="cuda")
    st.zero_grad()
    losses = st.run(inputs=xs, targets=ys)
    torch.cuda.synchronize()
    # oracle on this rank: the unsplit model, micro-batch by micro-batch (BatchNorm statistics are per micro-batch in both)
    for x, y in zip(xs, ys):
        (loss_fn(fwd_full(x), y) / n_micro).backward()
    worst = 0.0
    refp = dict(ref_stages[mesh.stage_id].named_parameters())
    for n, p in stages[mesh.stage_id].named_parameters():
        r = refp[n].grad
        worst = max(worst, float((p.grad - r).norm() / r.norm().clamp_min(1e-12)))
    return {"worst_rel": worst, "stage": mesh.stage_id, "n_losses": len(losses)}


@pytest.mark.timeout(300)
@pytest.mark.parametrize("which", ["resnet", "dcgan_d"])
def test_conv_model_two_stage_split_over_nccl_matches_unsplit(which):
    outs = run_distributed(_pp_conv_worker, 2, (which,), timeout=240, backend="nccl")
    for o in outs:
        assert o["worst_rel"] < 2e-3, o
    assert [o for o in outs if o["stage"] == 1][0]["n_losses"] == 2


def _pp_dp_worker(rank, world, n_stages, schedule):
    """PP x DP hybrid on GPUs (the configuration the reference deadlocks on, lab/homework-1.ipynb cell 48): n_stages pipeline
    stages x world/n_stages pipelines; after allreduce_grads every replica of a stage holds the mean gradient over pipelines,
    equal to the unsplit model's gradient on the union of the pipelines' batches."""
    from ddl25spring_b200.models.llama import LLama, causalLLMLoss
    from ddl25spring_b200.ops.nn import use_kernels
    from ddl25spring_b200.parallel import pp
    V, D, T, n_micro = 512, 96, 64, 2
    n_pipe = world // n_stages
    mesh = pp.Mesh(n_stages, n_pipe)
    torch.manual_seed(0)
    model = LLama(None, V, D, 6, "cuda", 2 * n_stages, T)
    stage_mod = use_kernels(model.split_stages(n_stages)[mesh.stage_id])
    st = pp.PipelineStage(stage_mod, mesh, (1, T, D), schedule=schedule, n_micro=n_micro,
                          loss_fn=lambda out, tgt: causalLLMLoss(out, tgt, V), device="cuda")
    all_tokens = torch.randint(0, V, (n_pipe, n_micro, T), generator=torch.Generator().manual_seed(1)).cuda()
    mine = list(all_tokens[mesh.pipeline_id].split(1))
    st.zero_grad()
    for _ in range(2):                                  # two iterations: communicators / schedules are reusable
        st.zero_grad()
        st.run(inputs=mine, targets=mine)
        st.allreduce_grads()
    torch.cuda.synchronize()
    torch.manual_seed(0)
    ref = LLama(None, V, D, 6, "cuda", 2 * n_stages, T)
    loss = sum(causalLLMLoss(ref(t), t, V) for p in range(n_pipe) for t in all_tokens[p].split(1)) / (n_micro * n_pipe)
    loss.backward()
    ref_stage = dict(ref.split_stages(n_stages)[mesh.stage_id].named_parameters())
    worst = 0.0
    for n, p in stage_mod.named_parameters():
        r = ref_stage[n].grad
        worst = max(worst, float((p.grad - r).norm() / r.norm().clamp_min(1e-12)))
    return {"worst_rel": worst, "stage": mesh.stage_id, "pipeline": mesh.pipeline_id}


@pytest.mark.timeout(400)
@pytest.mark.parametrize("schedule", ["gpipe", "1f1b"])
def test_pp_dp_hybrid_over_nccl(schedule):
    n = torch.cuda.device_count()
    if n < 4:
        pytest.skip("PP x DP needs >= 4 GPUs (2 stages x 2 pipelines)")
    world = 8 if n >= 8 else 4
    outs = run_distributed(_pp_dp_worker, world, (2, schedule), timeout=300, backend="nccl")
    for o in outs:
        assert o["worst_rel"] < 5e-3, o
    assert {o["pipeline"] for o in outs} == set(range(world // 2))


def _fed_dcgan_worker(rank, world):
    from ddl25spring_b200.gen.fed_dcgan import FederatedDCGAN
    from ddl25spring_b200.parallel.symm_impl import SymmetricArena
    g = torch.Generator().manual_seed(0)
    clients = [torch.rand(24, 3, 64, 64, generator=g) * 2 - 1 for _ in range(2 * world)]
    out = {}
    for name in ("fused", "nccl"):
        fed = FederatedDCGAN(clients, 1.0, local_steps=1, batch_size=8, seed=3, device="cuda", ngf=16, ndf=16,
                             collective=SymmetricArena() if name == "fused" else None)
        fed.run(1)
        torch.cuda.synchronize()
        first = (fed.glob.ga.flat.clone().cpu(), fed.glob.da.flat.clone().cpu())
        fed.run(1)
        torch.cuda.synchronize()
        out[name] = first + (fed.glob.ga.flat.clone().cpu(), fed.glob.da.flat.clone().cpu())
    return out


@pytest.mark.timeout(400)
def test_federated_dcgan_fused_aggregation_equals_nccl():
    world = min(torch.cuda.device_count(), 8)
    outs = run_distributed(_fed_dcgan_worker, world, timeout=300, backend="nccl")
    for o in outs:
        # round 1: same local training, the two aggregation paths differ only in summation order
        torch.testing.assert_close(o["fused"][0], o["nccl"][0], rtol=1e-5, atol=2e-6)
        torch.testing.assert_close(o["fused"][1], o["nccl"][1], rtol=1e-5, atol=2e-6)
        # round 2: Adam (lr 2e-4) turns last-bit differences of the averaged model into update differences of up to ~lr
        torch.testing.assert_close(o["fused"][2], o["nccl"][2], rtol=1e-3, atol=1e-3)
        torch.testing.assert_close(o["fused"][3], o["nccl"][3], rtol=1e-3, atol=1e-3)
    torch.testing.assert_close(outs[0]["fused"][2], outs[-1]["fused"][2], rtol=0, atol=0)      # replicas stay identical
