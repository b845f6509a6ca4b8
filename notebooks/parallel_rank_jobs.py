"""Workers of notebooks/data_and_pipeline_parallel.ipynb: spawned ranks import their entry function from a module on disk
(functions defined in a notebook cell cannot be pickled into a spawned process).  Counterparts of the rank programs
lab/tutorial_1b/DP/gradient_aggr/intro_DP_GA.py, weight_aggr/intro_DP_WA.py and PP/1F1B/intro_PP_1F1B_MB.py / _MP.py."""

VOCAB, DMODEL, LAYERS, HEADS, CTX = 256, 32, 4, 2, 32


def _device(rank, world):
    import torch
    return f"cuda:{rank}" if torch.cuda.is_available() and torch.cuda.device_count() >= world else "cpu"


def dp_rank(rank, world, mode, iters):
    """Data parallel by gradient ("grad") or weight ("weight") aggregation; returns (losses, parameter checksum)."""
    import torch

    from ddl25spring_b200.core.data import SyntheticTokens
    from ddl25spring_b200.models.llama import LLama, causalLLMLoss
    from ddl25spring_b200.parallel.dp import DataParallel, WeightAveraging
    dev = _device(rank, world)
    torch.manual_seed(0)                                     # the labs rely on identical seeds for identical replicas
    net = LLama(None, VOCAB, DMODEL, HEADS, dev, LAYERS, CTX)
    loss_fn = lambda m, t: causalLLMLoss(m(t), t, VOCAB)
    ds = iter(SyntheticTokens(VOCAB, 2, CTX, skip=rank * 5000))     # every rank reads its own slice of the stream
    losses = []
    if mode == "grad":
        dp = DataParallel(net, optimizer="adam", lr=3e-3)
        for _ in range(iters):
            losses.append(float(dp.train_step(loss_fn, next(ds).to(dev))))
    else:
        wa = WeightAveraging(net, torch.optim.Adam(net.parameters(), lr=3e-3))
        for _ in range(iters):
            loss = loss_fn(net, next(ds).to(dev))
            loss.backward()
            wa.step()
            losses.append(float(loss))
    print(f"rank {rank}/{world} on {dev}: {mode} aggregation, losses {[round(l, 3) for l in losses]}", flush=True)
    return losses, float(sum(p.double().sum() for p in net.parameters()))


def pp_rank(rank, world, schedule, n_micro, pipelines, iters):
    """One stage of ``pipelines`` pipelines of ``world // pipelines`` stages; the last stage of each pipeline returns its losses."""
    import torch

    from ddl25spring_b200.core.data import SyntheticTokens
    from ddl25spring_b200.models.llama import LLama, causalLLMLoss
    from ddl25spring_b200.parallel.pp import Mesh, PipelineStage
    dev = _device(rank, world)
    mesh = Mesh(world // pipelines, pipelines)
    torch.manual_seed(0)
    full = LLama(None, VOCAB, DMODEL, HEADS, dev, LAYERS, CTX)
    stage = full.split_stages(mesh.n_stages)[mesh.stage_id]
    st = PipelineStage(stage, mesh, (1, CTX, DMODEL), schedule=schedule, n_micro=n_micro,
                       loss_fn=lambda o, t: causalLLMLoss(o, t, VOCAB), device=dev)
    opt = torch.optim.Adam(stage.parameters(), lr=3e-3)
    ds = iter(SyntheticTokens(VOCAB, n_micro, CTX, skip=mesh.pipeline_id * 5000))
    out = []
    for _ in range(iters):
        opt.zero_grad()
        mbs = list(next(ds).split(1))
        losses = st.run(inputs=mbs, targets=mbs)
        st.allreduce_grads()                                 # across pipelines (no-op with one pipeline)
        opt.step()
        if st.is_last:
            out.append(float(sum(losses)) / n_micro)
    print(f"rank {rank}: pipeline {mesh.pipeline_id} stage {mesh.stage_id} done {out}", flush=True)
    return out if st.is_last else None
