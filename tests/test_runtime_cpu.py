"""CPU behaviour of the GPU-runtime features: graph capture is a no-op off-GPU, device-side helpers have host fallbacks."""
import torch
import torch.nn.functional as F

from ddl25spring_b200 import ops
from ddl25spring_b200.core.data import MNIST_MEAN, MNIST_STD
from ddl25spring_b200.utils.graphs import GraphedStep, graphed_train_step


def test_graphed_step_is_a_plain_call_on_cpu():
    calls = []

    def fn(a, b):
        calls.append(1)
        return a + b

    g = GraphedStep(fn, (torch.ones(3), torch.ones(3)))
    assert not g.enabled and g.graph is None
    assert torch.equal(g(torch.ones(3), 2 * torch.ones(3)), 3 * torch.ones(3))
    assert len(calls) == 1                                   # no warm-up runs on CPU


def test_graphed_train_step_trains_on_cpu():
    torch.manual_seed(0)
    model = torch.nn.Linear(4, 2)
    opt = torch.optim.SGD(model.parameters(), lr=0.1)
    x, y = torch.randn(8, 4), torch.randint(0, 2, (8,))
    step = graphed_train_step(model, F.cross_entropy, opt, x, y)
    l0 = float(step(x, y))
    for _ in range(20):
        l1 = float(step(x, y))
    assert l1 < l0


def test_dataparallel_and_fl_graph_flags_fall_back_on_cpu():
    from ddl25spring_b200.fl.client import ModelSlot
    from ddl25spring_b200.models.mnist_cnn import MnistCnn
    from ddl25spring_b200.parallel.dp import DataParallel
    dp = DataParallel(torch.nn.Linear(4, 4), optimizer="sgd", lr=0.1, graph=True)
    assert dp.graph is False
    loss = dp.train_step(lambda m, x: m(x).square().mean(), torch.randn(2, 4))
    assert loss == loss
    assert ModelSlot(MnistCnn, "cpu", graph=True).graph is False


def test_u8_normalize_and_adam_tensor_step_host_paths():
    u8 = torch.randint(0, 256, (5, 784), dtype=torch.uint8)
    assert torch.equal(ops.u8_normalize(u8, MNIST_MEAN, MNIST_STD), u8.float().div_(255.0).sub_(MNIST_MEAN).div_(MNIST_STD))
    p1, p2 = torch.ones(8), torch.ones(8)
    g = torch.full((8,), 0.5)
    m1, v1, m2, v2 = (torch.zeros(8) for _ in range(4))
    ops.adam_step(p1, g, m1, v1, 3, 1e-2)
    ops.adam_step(p2, g, m2, v2, torch.tensor(3, dtype=torch.int32), 1e-2)
    assert torch.equal(p1, p2)


def test_gate_up_weights_are_stacked_and_stay_stacked_in_a_flat_arena():
    """use_kernels lays w_gate / w_up out back to back (one GEMM over [Wg; Wu] on the GPU); values, Parameters and state_dict
    keys are untouched, a FlatArena keeps the pair adjacent (also in the gradient arena), and CPU forward is unchanged."""
    import torch
    from ddl25spring_b200.core.arena import FlatArena
    from ddl25spring_b200.models.llama import LLama
    from ddl25spring_b200.ops.nn import _direct_grad_slot, _stacked, use_kernels
    torch.manual_seed(0)
    net = LLama(None, 64, 32, 2, "cpu", 2, 8)
    tok = torch.randint(0, 64, (2, 8))
    before = {k: v.clone() for k, v in net.state_dict().items()}
    y0 = net(tok)
    use_kernels(net)
    mlps = [m for m in net.modules() if hasattr(m, "w_gate")]
    assert mlps and all(_stacked(m.w_gate.weight.data, m.w_up.weight.data) for m in mlps)
    assert all(torch.equal(before[k], v) for k, v in net.state_dict().items())
    arena = FlatArena.from_module(net, with_grad=True)
    for m in mlps:
        assert _stacked(m.w_gate.weight.data, m.w_up.weight.data) and _stacked(m.w_gate.weight.grad, m.w_up.weight.grad)
        assert m.w_gate.weight._ddl_direct_grad
        with torch.no_grad():
            assert _direct_grad_slot(m.w_gate.weight) is m.w_gate.weight.grad      # backward runs with grad mode off
        assert _direct_grad_slot(m.w_gate.weight) is None                          # ... never while a graph is being built
    torch.testing.assert_close(net(tok), y0)
    net(tok).sum().backward()
    assert float(arena.grad.abs().sum()) > 0
