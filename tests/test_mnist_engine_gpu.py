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
