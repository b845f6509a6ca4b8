"""Packed MnistCnn engine (hand-written grouped kernels) vs. the plain PyTorch model in fp32."""
import pytest
import torch
import torch.nn.functional as F

from ddl25spring_b200.core.arena import FlatArena
from ddl25spring_b200.models.mnist_cnn import MnistCnn

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _model_and_flat(seed=0, p=0.0):
    torch.manual_seed(seed)
    m = MnistCnn(p_drop1=p, p_drop2=p).to(DEV)
    arena = FlatArena.from_module(m)
    assert arena.size == 1_200_000
    return m, arena


def test_arena_layout_matches_engine_offsets(cuda_lib):
    from ddl25spring_b200.fl.packed import OFF
    m, arena = _model_and_flat()
    for s in arena.specs:
        assert OFF[s.name] == s.offset, s


@pytest.mark.parametrize("n", [7, 100, 300])
def test_eval_forward_matches_torch(cuda_lib, n):
    from ddl25spring_b200.fl.packed import PackedMnistEngine
    old = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    try:
        m, arena = _model_and_flat(1)
        m.eval()
        x = torch.randn(n, 1, 28, 28, device=DEV)
        y = torch.randint(0, 10, (n,), device=DEV)
        with torch.no_grad():
            exp = m.fc2(F.relu(m.fc1(m.features(x))))
        eng = PackedMnistEngine(1, 128, train=False)
        correct = torch.zeros((), dtype=torch.int32, device=DEV)
        got = eng.eval_logits(arena.flat, x, y, correct)
        err = float((got - exp).abs().max() / exp.abs().max())
        assert err < 5e-3, err
        assert abs(int(correct) - int((exp.argmax(1) == y).sum())) <= max(1, n // 50)
    finally:
        torch.backends.cudnn.allow_tf32 = old


def _grads(m, x, y):
    m.train()
    m.zero_grad()
    F.nll_loss(m(x), y).backward()
    return [p.grad.detach().clone() for p in m.parameters()]


@pytest.mark.parametrize("G,B,bszs", [(2, 32, [32, 20]), (3, 100, [100, 100, 37])])
def test_train_step_matches_autograd(cuda_lib, G, B, bszs):
    """Gradients implied by one engine step vs torch autograd in fp32.  conv2 runs in TF32 here exactly
    as the reference's cuDNN convs do by default, so the yardstick for the conv-dependent tensors is
    the error cuDNN-TF32 itself makes against fp32 on the same batch (ReLU-boundary flips dominate)."""
    from ddl25spring_b200.fl.packed import PackedMnistEngine
    old = torch.backends.cudnn.allow_tf32
    try:
        lr = 0.05
        models, arenas = zip(*[_model_and_flat(10 + g) for g in range(G)])
        W = torch.stack([a.flat for a in arenas]).contiguous()
        X = torch.randn(G, B, 784, device=DEV)
        Y = torch.randint(0, 10, (G, B), device=DEV)
        perm = torch.stack([torch.randperm(B, device=DEV) for _ in range(G)]).int().contiguous()
        eng = PackedMnistEngine(G, B, p_drop1=0.0, p_drop2=0.0)
        eng.prep_weights(W)
        W0 = W.clone()
        eng.train_step(W, X, B * 784, Y, B, perm, B, 0, bszs, lr, seed=1)
        torch.cuda.synchronize()
        for g in range(G):
            x = X[g][perm[g, :bszs[g]].long()].view(-1, 1, 28, 28)
            y = Y[g][perm[g, :bszs[g]].long()]
            torch.backends.cudnn.allow_tf32 = False
            g32 = _grads(models[g], x, y)
            torch.backends.cudnn.allow_tf32 = True
            gtf = _grads(models[g], x, y)
            got_flat = (W0[g] - W[g]) / lr
            for spec, a, b in zip(arenas[g].specs, g32, gtf):
                got = got_flat[spec.offset:spec.offset + spec.numel].view(spec.shape)
                mx = a.abs().max().clamp_min(1e-8)
                err = float((got - a).abs().max() / mx)
                yard = float((b - a).abs().max() / mx)
                assert err < max(2.0 * yard, 2e-3), (g, spec.name, err, yard)
            # the loss the head kernel reports is the mean NLL of the client's minibatch
            torch.backends.cudnn.allow_tf32 = False
    finally:
        torch.backends.cudnn.allow_tf32 = old


def test_local_train_learns(cuda_lib):
    from ddl25spring_b200.core.data import SyntheticMNIST
    from ddl25spring_b200.fl.packed import PackedMnistEngine
    ds = SyntheticMNIST(n_train=1600, n_test=500, seed=1, noise=0.7)
    G, n = 4, 400
    X = ds.train.images.view(G, n, 784).to(DEV).contiguous()
    Y = ds.train.targets.view(G, n).to(DEV).contiguous()
    m, arena = _model_and_flat(3)
    W = arena.flat[None, :].repeat(G, 1).contiguous()
    epochs = 3
    perm = torch.stack([torch.cat([torch.randperm(n) for _ in range(epochs)]) for _ in range(G)]).int().to(DEV)
    eng = PackedMnistEngine(G, 50)
    steps = eng.local_train(W, X, Y, perm, [n] * G, 50, epochs, 0.05, seed=7)
    assert steps == epochs * 8
    ev = PackedMnistEngine(1, 512, train=False)
    xt, yt = ds.test.images.to(DEV), ds.test.targets.to(DEV)
    c0 = torch.zeros((), dtype=torch.int32, device=DEV)
    ev.eval_logits(arena.flat, xt, yt, c0, want_logits=False)
    c1 = torch.zeros((), dtype=torch.int32, device=DEV)
    ev.eval_logits(W.mean(0).contiguous(), xt, yt, c1, want_logits=False)
    assert int(c1) > int(c0) + 150, (int(c0), int(c1))
    assert torch.isfinite(W).all()


def test_u8_transport_is_bit_identical_to_resident_floats():
    """End-to-end mode ships raw uint8 pixels and normalises on the device: same bits as the host transform, same model."""
    import torch
    from ddl25spring_b200 import ops
    from ddl25spring_b200.core.data import MNIST_MEAN, MNIST_STD, SyntheticMNIST, split
    from ddl25spring_b200.fl.packed_server import PackedFedAvgServer
    u8 = torch.randint(0, 256, (1000, 784), dtype=torch.uint8)
    host = u8.float().div_(255.0).sub_(MNIST_MEAN).div_(MNIST_STD)
    assert torch.equal(ops.u8_normalize(u8.cuda(), MNIST_MEAN, MNIST_STD).cpu(), host)
    ds = SyntheticMNIST(n_train=1600, n_test=256, seed=0)
    flats = []
    for on_dev in (True, False):
        srv = PackedFedAvgServer(0.05, 50, split(ds.train, 8, True, 1), 0.5, 1, 3, test_set=ds.test, device="cuda",
                                 data_on_device=on_dev, prefetch_next=on_dev)
        srv.run(2)
        if not on_dev:
            assert srv.X_u8 is not None
            assert srv.h2d_bytes_last_round == 4 * 200 * 784 + 4 * 200 * 8 + 4 * 200 * 4      # u8 pixels + int64 labels + int32 perms
        flats.append(srv.arena.flat.detach().clone())
    assert torch.equal(flats[0], flats[1])


def test_packed_server_resumes_from_checkpoint(cuda_lib, tmp_path):
    """PackedFedAvgServer.run(checkpointer=...): a run stopped after 2 rounds and continued by a new server object equals
    the uninterrupted 4-round run bit for bit (the sampling RNG is saved as it was before the next round's prefetch draw)."""
    from ddl25spring_b200.core.checkpoint import RoundCheckpointer
    from ddl25spring_b200.core.data import SyntheticMNIST, split
    from ddl25spring_b200.fl.packed_server import PackedFedAvgServer
    ds = SyntheticMNIST(n_train=2000, n_test=500, seed=4, noise=0.8)
    subsets = split(ds.train, 10, True, 3)

    def make():
        return PackedFedAvgServer(0.05, 50, subsets, 0.4, 1, 7, test_set=ds.test, device="cuda")

    full = make()
    ref = full.run(4)
    ck = RoundCheckpointer(str(tmp_path), every=1)
    a = make()
    a.run(2, checkpointer=ck)
    del a
    b = make()
    res = b.run(4, checkpointer=ck)
    torch.cuda.synchronize()
    assert res.test_accuracy == ref.test_accuracy
    assert torch.equal(b.arena.flat, full.arena.flat)
