"""Pipeline schedules (as data) and the stage executor over gloo."""
import pytest
import torch

from ddl25spring_b200.parallel import pp
from tests.dist_utils import run_distributed

VOCAB, D, H, L, T = 64, 24, 2, 3, 8


@pytest.mark.parametrize("name", ["gpipe", "1f1b", "sequential"])
@pytest.mark.parametrize("stages", [1, 2, 3, 4])
@pytest.mark.parametrize("micro", [1, 2, 3, 8])
def test_schedules_are_deadlock_free(name, stages, micro):
    if name == "sequential":
        micro = 1
    assert pp.check_schedule(name, stages, micro) >= 2 * micro


def test_1f1b_bounds_activations_in_flight():
    for stages in (2, 3, 4):
        for s in range(stages):
            ops, live, peak = pp.one_f_one_b(stages, s, 8), 0, 0
            for op in ops:
                live += 1 if op.kind == "F" else -1
                peak = max(peak, live)
            assert peak <= stages - s, (stages, s, peak)         # GPipe would hold all 8


def _make_model():
    from ddl25spring_b200.models.llama import LLama
    torch.manual_seed(0)
    return LLama(None, VOCAB, D, H, "cpu", L, T)


def _batch(n):
    g = torch.Generator().manual_seed(1)
    return torch.randint(0, VOCAB, (n, T), generator=g)


def _pp_worker(rank, world, schedule, n_micro, n_pipelines):
    from ddl25spring_b200.models.llama import causalLLMLoss
    n_stages = world // n_pipelines
    mesh = pp.Mesh(n_stages, n_pipelines)
    model = _make_model()
    stage_mod = model.split_stages(n_stages)[mesh.stage_id]
    tokens = _batch(n_micro * n_pipelines)                       # every pipeline gets its own slice
    mine = tokens[mesh.pipeline_id * n_micro:(mesh.pipeline_id + 1) * n_micro]
    mbs = list(mine.split(1))
    st = pp.PipelineStage(stage_mod, mesh, (1, T, D), schedule=schedule, n_micro=n_micro,
                          loss_fn=lambda out, tgt: causalLLMLoss(out, tgt, VOCAB), device="cpu")
    for p in stage_mod.parameters():
        p.grad = None
    losses = st.run(inputs=mbs, targets=mbs)
    st.allreduce_grads()
    return {n: p.grad.clone() for n, p in stage_mod.named_parameters() if p.grad is not None}, losses, mesh.stage_id


@pytest.mark.timeout(180)
@pytest.mark.parametrize("schedule,n_micro", [("gpipe", 3), ("1f1b", 4), ("sequential", 1)])
def test_pipeline_grads_equal_unsplit_model(schedule, n_micro):
    from ddl25spring_b200.models.llama import causalLLMLoss
    outs = run_distributed(_pp_worker, 3, (schedule, n_micro, 1), timeout=150)
    model = _make_model()
    tokens = _batch(n_micro)
    loss = sum(causalLLMLoss(model(t), t, VOCAB) for t in tokens.split(1)) / n_micro
    loss.backward()
    ref_stages = model.split_stages(3)
    for grads, losses, sid in outs:
        ref = dict(ref_stages[sid].named_parameters())
        assert grads, sid
        for n, g in grads.items():
            torch.testing.assert_close(g, ref[n].grad, rtol=1e-4, atol=1e-6)
    last = [o for o in outs if o[2] == 2][0]
    assert len(last[1]) == n_micro


@pytest.mark.timeout(240)
def test_pp_x_dp_hybrid_completes_and_averages():
    """2 pipelines x 3 stages = the 6-rank mesh on which the reference deadlocks (homework-1 cell 48)."""
    from ddl25spring_b200.models.llama import causalLLMLoss
    n_micro = 2
    outs = run_distributed(_pp_worker, 6, ("1f1b", n_micro, 2), timeout=200)
    model = _make_model()
    tokens = _batch(2 * n_micro)
    loss = sum(causalLLMLoss(model(t), t, VOCAB) for t in tokens.split(1)) / (2 * n_micro)
    loss.backward()
    ref_stages = model.split_stages(3)
    for grads, _, sid in outs:
        ref = dict(ref_stages[sid].named_parameters())
        for n, g in grads.items():
            torch.testing.assert_close(g, ref[n].grad, rtol=1e-4, atol=1e-6)


def _two_stage_worker(rank, world, which):
    import torch.nn.functional as Fn
    mesh = pp.Mesh(2, 1)
    torch.manual_seed(0)
    if which == "resnet":
        from ddl25spring_b200.models.resnet import resnet18, split_resnet18_two_stage
        full = resnet18().eval()                          # eval: BN uses running stats, so micro-batching is exact
        stages = split_resnet18_two_stage(full)
        x = torch.randn(4, 3, 32, 32, generator=torch.Generator().manual_seed(1))
        y = torch.randint(0, 10, (4,), generator=torch.Generator().manual_seed(2))
        act_shape, loss_fn = (2, 128, 4, 4), (lambda o, t: Fn.cross_entropy(o, t))
    else:
        from ddl25spring_b200.models.dcgan import Generator
        full = Generator(nz=16, ngf=8).eval()
        stages = full.stages(3)
        x = torch.randn(4, 16, 1, 1, generator=torch.Generator().manual_seed(1))
        y = torch.zeros(4, 3, 64, 64)
        act_shape, loss_fn = (2, 16, 16, 16), (lambda o, t: Fn.mse_loss(o, t))
    st = pp.PipelineStage(stages[rank], mesh, act_shape, schedule="1f1b", n_micro=2, loss_fn=loss_fn, device="cpu")
    st.run(inputs=list(x.split(2)), targets=list(y.split(2)))
    mine = {n: p.grad.clone() for n, p in stages[rank].named_parameters() if p.grad is not None}
    for p in full.parameters():
        p.grad = None
    out = full(x)
    (sum(loss_fn(o, t) for o, t in zip(out.split(2), y.split(2))) / 2).backward()
    ref = {n: p.grad.clone() for n, p in stages[rank].named_parameters() if p.grad is not None}
    return max(float((mine[n] - ref[n]).abs().max()) for n in mine), len(mine)


@pytest.mark.timeout(200)
@pytest.mark.parametrize("which", ["resnet", "dcgan"])
def test_two_stage_model_parallel_split(which):
    """BASELINE.json config 5: 2-stage model-parallel split (ResNet-18 / DCGAN generator) == unsplit model."""
    outs = run_distributed(_two_stage_worker, 2, (which,), timeout=150)
    for err, n in outs:
        assert n > 0 and err < 1e-4, (err, n)


@pytest.mark.parametrize("name", ["gpipe", "1f1b"])
@pytest.mark.parametrize("stages", [2, 3, 4, 5])
@pytest.mark.parametrize("micro", [1, 2, 3, 4, 8])
def test_executor_p2p_order_is_deadlock_free_under_nccl_stream_semantics(name, stages, micro):
    """NCCL runs a rank's p2p operations in issue order on one stream per communicator (rendezvous, no tags).  With activations
    and gradients on one communicator the executor's pre-posted receives hang (observed on 2 B200s); with one communicator per
    direction - what Mesh creates under NCCL - every schedule completes."""
    assert pp.check_stream_ordered_p2p(name, stages, micro, per_direction=True) > 0
    with pytest.raises(RuntimeError, match="deadlock"):
        pp.check_stream_ordered_p2p(name, stages, micro, per_direction=False)
