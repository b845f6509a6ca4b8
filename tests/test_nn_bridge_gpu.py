"""Autograd bridges (Linear on tcgen05, fused xent, RMSNorm) vs torch autograd in fp32."""
import pytest
import torch
import torch.nn.functional as F

from ddl25spring_b200.ops import nn as knn

pytestmark = pytest.mark.gpu
DEV = "cuda"


def test_kernel_linear_fwd_bwd(cuda_lib):
    torch.manual_seed(0)
    x = torch.randn(6, 50, 96, device=DEV, requires_grad=True)
    lin = torch.nn.Linear(96, 160).to(DEV)
    y0 = lin(x); y0.square().sum().backward()
    g0 = (x.grad.clone(), lin.weight.grad.clone(), lin.bias.grad.clone())
    x.grad = None; lin.zero_grad()
    y1 = knn.kernel_linear(x, lin.weight, lin.bias); y1.square().sum().backward()
    torch.testing.assert_close(y1, y0, rtol=1e-4, atol=1e-4)
    for a, b in zip((x.grad, lin.weight.grad, lin.bias.grad), g0):
        torch.testing.assert_close(a, b, rtol=2e-4, atol=2e-3)


def test_kernel_cross_entropy_and_rmsnorm(cuda_lib):
    torch.manual_seed(1)
    lg = torch.randn(64, 1000, device=DEV, requires_grad=True)
    t = torch.randint(0, 1000, (64,), device=DEV)
    t[::7] = -100
    l0 = F.cross_entropy(lg, t, ignore_index=-100); l0.backward(); g0 = lg.grad.clone(); lg.grad = None
    l1 = knn.kernel_cross_entropy(lg, t); l1.backward()
    torch.testing.assert_close(l1, l0, rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(lg.grad, g0, rtol=1e-4, atol=1e-7)
    x = torch.randn(8, 33, 288, device=DEV, requires_grad=True)
    w = torch.randn(288, device=DEV, requires_grad=True)
    ref = x * torch.rsqrt(x.pow(2).mean(-1, keepdim=True) + 1e-5) * w
    ref.sin().sum().backward(); gx, gw = x.grad.clone(), w.grad.clone(); x.grad = None; w.grad = None
    out = knn.kernel_rms_norm(x, w); out.sin().sum().backward()
    torch.testing.assert_close(out, ref, rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(x.grad, gx, rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(w.grad, gw, rtol=1e-4, atol=1e-4)


def test_llama_on_kernels_matches_torch(cuda_lib):
    from ddl25spring_b200.models.llama import LLama, causalLLMLoss
    import copy
    torch.manual_seed(0)
    a = LLama(None, 512, 96, 6, DEV, 2, 64)
    b = knn.use_kernels(copy.deepcopy(a))
    tok = torch.randint(0, 512, (3, 64), device=DEV)
    la = causalLLMLoss(a(tok), tok, 512); la.backward()
    lb = knn.kernel_causal_lm_loss(b(tok), tok); lb.backward()
    torch.testing.assert_close(lb, la, rtol=1e-4, atol=1e-4)
    for (n, p), q in zip(a.named_parameters(), b.parameters()):
        torch.testing.assert_close(q.grad, p.grad, rtol=5e-3, atol=5e-5, msg=n)
