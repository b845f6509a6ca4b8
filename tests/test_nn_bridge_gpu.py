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


@pytest.mark.parametrize("shape", [(64, 48), (32, 16, 8, 8), (5, 3, 4, 4)])
def test_kernel_batchnorm_matches_torch(cuda_lib, shape):
    torch.manual_seed(0)
    C = shape[1]
    Mod = torch.nn.BatchNorm1d if len(shape) == 2 else torch.nn.BatchNorm2d
    a, b = Mod(C).to(DEV), Mod(C).to(DEV)
    with torch.no_grad():
        a.weight.uniform_(0.5, 1.5); a.bias.normal_()
        b.load_state_dict(a.state_dict())
    for step in range(2):
        x = torch.randn(*shape, device=DEV) * 2 + 1
        xa, xb = x.clone().requires_grad_(), x.clone().requires_grad_()
        ya = a(xa); (ya * ya.detach().sign()).sum().backward()
        yb = knn.kernel_batch_norm(xb, b); (yb * yb.detach().sign()).sum().backward()
        torch.testing.assert_close(yb, ya, rtol=1e-4, atol=1e-5)
        torch.testing.assert_close(xb.grad, xa.grad, rtol=1e-3, atol=1e-5)
        torch.testing.assert_close(b.weight.grad, a.weight.grad, rtol=1e-3, atol=1e-4)
        torch.testing.assert_close(b.bias.grad, a.bias.grad, rtol=1e-3, atol=1e-4)
        torch.testing.assert_close(b.running_mean, a.running_mean, rtol=1e-4, atol=1e-6)
        torch.testing.assert_close(b.running_var, a.running_var, rtol=1e-4, atol=1e-6)
    a.eval(); b.eval()
    x = torch.randn(*shape, device=DEV)
    with torch.no_grad():
        torch.testing.assert_close(knn.kernel_batch_norm(x, b), a(x), rtol=1e-4, atol=1e-5)


@pytest.mark.parametrize("B,H,T,D", [(3, 6, 256, 48), (2, 2, 100, 32), (1, 4, 300, 64), (2, 4, 128, 64), (1, 2, 64, 16), (1, 3, 512, 48)])
def test_kernel_causal_attention(cuda_lib, B, H, T, D):
    """T % 32 == 0: grouped tcgen05 GEMMs (3xTF32) + row-wise softmax kernels; other T: the CUDA-core flash kernels."""
    torch.manual_seed(0)
    q, k, v = (torch.randn(B, H, T, D, device=DEV, requires_grad=True) for _ in range(3))
    ref = F.scaled_dot_product_attention(q.double(), k.double(), v.double(), is_causal=True)
    w = torch.randn_like(ref)
    (ref * w).sum().backward()
    g = [t.grad.clone() for t in (q, k, v)]
    for t in (q, k, v):
        t.grad = None
    out = knn.kernel_causal_attention(q, k, v)
    (out * w.float()).sum().backward()
    torch.testing.assert_close(out.double(), ref, rtol=1e-4, atol=1e-5)
    for t, gg in zip((q, k, v), g):
        torch.testing.assert_close(t.grad.double(), gg.double(), rtol=1e-3, atol=2e-5)


def test_rope_then_tensor_core_attention_chain(cuda_lib):
    """The fused QKV split + RoPE kernel hands 32-column-padded q / k / v to the attention GEMMs as views: the chain equals
    the PyTorch statement, forward and gradient of the packed qkv."""
    from ddl25spring_b200.models.llama import apply_rope, rope_tables
    torch.manual_seed(0)
    B, T, H, hd = 2, 128, 6, 48
    cos, sin = rope_tables(hd, 256, device=DEV)
    qkv = torch.randn(B, T, 3 * H * hd, device=DEV, requires_grad=True)
    q, k, v = knn.kernel_rope_qkv(qkv, cos, sin, H)
    assert q.shape == (B, H, T, hd) and q.stride(2) == 64
    o = knn.kernel_causal_attention(q, k, v)
    w = torch.randn(B, H, T, hd, device=DEV)
    (g1,) = torch.autograd.grad((o * w).sum(), qkv)
    qr, kr, vr = qkv.view(B, T, 3, H, hd).permute(2, 0, 3, 1, 4)
    orf = F.scaled_dot_product_attention(apply_rope(qr, cos, sin), apply_rope(kr, cos, sin), vr, is_causal=True)
    (g2,) = torch.autograd.grad((orf * w).sum(), qkv)
    torch.testing.assert_close(o, orf, rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(g1, g2, rtol=1e-3, atol=2e-5)


@pytest.mark.gpu
def test_kernel_rope_qkv_and_swiglu_match_torch(cuda_lib):
    from ddl25spring_b200.models.llama import apply_rope, rope_tables
    from ddl25spring_b200.ops.nn import kernel_rope_qkv, kernel_swiglu
    torch.manual_seed(0)
    B, T, H, hd = 3, 100, 6, 48
    cos, sin = rope_tables(hd, 256, device="cuda")
    qkv = torch.randn(B, T, 3 * H * hd, device="cuda", requires_grad=True)
    q, k, v = kernel_rope_qkv(qkv, cos, sin, H)
    qr, kr, vr = qkv.view(B, T, 3, H, hd).permute(2, 0, 3, 1, 4)
    qr, kr = apply_rope(qr, cos, sin), apply_rope(kr, cos, sin)
    for a, b in ((q, qr), (k, kr), (v, vr)):
        assert torch.allclose(a, b, atol=1e-6)
    gq, gk, gv = torch.randn_like(q), torch.randn_like(k), torch.randn_like(v)
    g1 = torch.autograd.grad((q * gq).sum() + (k * gk).sum() + (v * gv).sum(), qkv)[0]
    g2 = torch.autograd.grad((qr * gq).sum() + (kr * gk).sum() + (vr * gv).sum(), qkv)[0]
    assert torch.allclose(g1, g2, atol=1e-5)
    g = torch.randn(5, 77, 768, device="cuda", requires_grad=True)
    u = torch.randn(5, 77, 768, device="cuda", requires_grad=True)
    h, hr = kernel_swiglu(g, u), F.silu(g) * u
    assert torch.allclose(h, hr, atol=1e-5, rtol=1e-5)
    dh = torch.randn_like(h)
    for a, b in zip(torch.autograd.grad(h, (g, u), dh), torch.autograd.grad(hr, (g, u), dh)):
        assert torch.allclose(a, b, atol=1e-5, rtol=1e-4)


@pytest.mark.parametrize("M,K,N,bias", [(64, 30, 64, True), (37, 60, 2, True), (256, 13, 26, True), (100, 48, 31, False), (8, 256, 2, True),
                                        (1025, 31, 48, True)])
def test_kernel_linear_ragged_widths(cuda_lib, M, K, N, bias):
    """The tabular models' layer widths (30 -> 64, 256 -> 2, party slabs of 13 / 26 columns ...): forward, input gradient,
    weight gradient and bias gradient all come from the tcgen05 GEMM (padded, bias-augmented operands), vs fp32 autograd."""
    torch.manual_seed(M + K + N)
    x = torch.randn(M, K, device=DEV, requires_grad=True)
    lin = torch.nn.Linear(K, N, bias=bias).to(DEV)
    y0 = lin(x); (y0.sin()).sum().backward()
    g0 = [x.grad.clone(), lin.weight.grad.clone()] + ([lin.bias.grad.clone()] if bias else [])
    x.grad = None; lin.zero_grad()
    from ddl25spring_b200 import ops
    ops.reset_launch_count()
    y1 = knn.kernel_linear(x, lin.weight, lin.bias); (y1.sin()).sum().backward()
    assert ops.launch_count() >= 6                      # pad2d x2 + gemm, pad2d + 2 gemms: all ours
    torch.testing.assert_close(y1, y0, rtol=1e-4, atol=1e-4)
    g1 = [x.grad, lin.weight.grad] + ([lin.bias.grad] if bias else [])
    for a, b in zip(g1, g0):
        torch.testing.assert_close(a, b, rtol=3e-4, atol=3e-4)


def test_tabular_models_train_on_kernels_like_torch(cuda_lib):
    """HeartDiseaseNN / Autoencoder / VFLNetwork with use_kernels + the fused arena optimizer follow the ATen + torch.optim
    trajectories (same seeds, dropout off)."""
    import copy
    from ddl25spring_b200.models.tabular import Autoencoder, BottomModel, HeartDiseaseNN
    from ddl25spring_b200.ops.optim import ArenaOptimizer
    from ddl25spring_b200.parallel import vfl
    torch.manual_seed(0)
    x = torch.randn(256, 30, device=DEV)
    y = torch.randint(0, 2, (256,), device=DEV)
    a = HeartDiseaseNN(30, p_drop=0.0).to(DEV)
    b = knn.use_kernels(copy.deepcopy(a))
    oa, ob = torch.optim.AdamW(a.parameters()), ArenaOptimizer(b, "adamw")
    for _ in range(5):
        for net, opt, lf in ((a, oa, F.cross_entropy), (b, ob, knn.kernel_cross_entropy)):
            opt.zero_grad()
            loss = lf(net(x), y)
            loss.backward()
            opt.step()
    for p, q in zip(a.parameters(), b.parameters()):
        torch.testing.assert_close(q, p, rtol=2e-3, atol=2e-4)
    # VAE: Linear + BatchNorm1d + fused loss + fused Adam
    from ddl25spring_b200.gen.vae import fused_vae_loss
    data = torch.randn(192, 31, device=DEV)
    va = Autoencoder(31, 48, 32, 16).to(DEV)
    vb = knn.use_kernels(copy.deepcopy(va))
    oa, ob = torch.optim.Adam(va.parameters(), lr=1e-3), ArenaOptimizer(vb, "adam", lr=1e-3)
    la = lb = None
    for i in range(4):
        for net, opt in ((va, oa), (vb, ob)):
            torch.manual_seed(100 + i)                      # same reparameterisation noise
            opt.zero_grad()
            rec, mu, lv = net(data)
            loss = fused_vae_loss(rec, data, mu, lv)
            loss.backward()
            opt.step()
            if net is va:
                la = float(loss)
            else:
                lb = float(loss)
    assert lb == pytest.approx(la, rel=2e-3)


def test_vfl_network_runs_on_kernels(cuda_lib):
    from ddl25spring_b200 import ops
    from ddl25spring_b200.core.data import SyntheticHeart
    from ddl25spring_b200.parallel import vfl
    tab = SyntheticHeart().encoded()
    parts = vfl.contiguous_raw_partition(tab, 4)
    torch.manual_seed(0)
    net = vfl.build_vfl(tab, parts).to(DEV)
    ops.reset_launch_count()
    hist = net.train_with_settings(6, 64, 4, parts, tab.x[:640], tab.y[:640])
    assert net._on_kernels and ops.launch_count() > 100
    assert hist[-1] < hist[0]
    acc, loss = net.test(tab.x[640:840], tab.y[640:840])
    assert float(acc) > 0.5


def test_kernel_bce_logits_matches_torch(cuda_lib):
    torch.manual_seed(0)
    for target in (0.0, 1.0):
        x = (3 * torch.randn(64, 1, 1, 1, device=DEV)).requires_grad_()
        ref = F.binary_cross_entropy_with_logits(x, torch.full_like(x, target))
        (g0,) = torch.autograd.grad(ref, x)
        out = knn.kernel_bce_logits(x, target)
        (g1,) = torch.autograd.grad(out * 2.0, x)
        torch.testing.assert_close(out, ref, rtol=1e-5, atol=1e-6)
        torch.testing.assert_close(g1, 2.0 * g0, rtol=1e-5, atol=1e-7)


def test_second_graph_capture_survives_a_dead_cycle_owning_the_first_graph(cuda_lib):
    """A finished model's step graph kept alive only by a reference cycle must not be collected in the middle of the next
    capture (destroying a CUDA graph there invalidates it): GraphedStep collects first and holds the collector off."""
    import gc
    from ddl25spring_b200.utils.graphs import GraphedStep

    class Holder:
        pass

    def make():
        lin = torch.nn.Linear(64, 64).to(DEV)
        h = Holder()

        def step(x):
            h.n = getattr(h, "n", 0) + 1                 # closure -> holder -> graphed step -> closure: a cycle
            y = knn.kernel_linear(x, lin.weight, lin.bias)
            junk = [[] for _ in range(2000)]             # enough container allocations to trigger automatic collections
            for j in junk:
                j.append(junk)
            return y.sum()

        h.g = GraphedStep(step, (torch.randn(32, 64, device=DEV),))
        return h

    gc.collect()
    h1 = make()
    del h1                                               # dead cycle, first graph still alive inside it
    h2 = make()
    out = h2.g(torch.ones(32, 64, device=DEV))
    assert torch.isfinite(out).item()
    assert gc.isenabled()


def test_stacked_gate_up_swiglu_matches_torch(cuda_lib):
    """One GEMM against [Wg; Wu] + stacked SwiGLU kernel, forward and all three gradients, vs fp32 torch autograd."""
    from ddl25spring_b200.models.llama import GatedMLP
    torch.manual_seed(3)
    mlp = GatedMLP(96, 160).to(DEV)
    x = torch.randn(5, 40, 96, device=DEV, requires_grad=True)
    y0 = mlp(x); y0.square().sum().backward()
    g0 = [x.grad.clone()] + [p.grad.clone() for p in mlp.parameters()]
    x.grad = None; mlp.zero_grad()
    knn.use_kernels(mlp)
    assert mlp.gate_up_fn is not None and knn._stacked(mlp.w_gate.weight, mlp.w_up.weight)
    y1 = mlp(x); y1.square().sum().backward()
    torch.testing.assert_close(y1, y0, rtol=1e-4, atol=1e-4)
    for a, b in zip([x.grad] + [p.grad for p in mlp.parameters()], g0):
        torch.testing.assert_close(a, b, rtol=2e-4, atol=2e-3)
    # weights that are NOT back to back in memory take the two-GEMM path and give the same answer
    mlp.w_up.weight.data = mlp.w_up.weight.data.clone()
    assert not knn._stacked(mlp.w_gate.weight, mlp.w_up.weight)
    torch.testing.assert_close(mlp(x), y1, rtol=1e-5, atol=1e-5)


def test_weight_gradients_accumulate_straight_into_arena_slots(cuda_lib):
    """Parameters bound to a flat gradient arena get their weight gradient from the GEMM epilogue (C = A B + C): same values as
    autograd's own accumulation over two backward passes, post-accumulate-grad hooks run once per parameter and pass."""
    from ddl25spring_b200.core.arena import FlatArena
    from ddl25spring_b200.models.llama import LLama
    torch.manual_seed(5)
    V = 512
    ref = LLama(None, V, 96, 2, DEV, 2, 64)
    net = LLama(None, V, 96, 2, DEV, 2, 64)
    net.load_state_dict(ref.state_dict())
    knn.use_kernels(net)
    arena = FlatArena.from_module(net, with_grad=True)
    fired = {}
    for name, p in net.named_parameters():
        assert p._ddl_direct_grad
        p.register_post_accumulate_grad_hook((lambda n: (lambda _p: fired.__setitem__(n, fired.get(n, 0) + 1)))(name))
    tok = torch.randint(0, V, (2, 64), device=DEV)
    for _ in range(2):                                   # gradients of two passes add up in both models
        knn.kernel_causal_lm_loss(net(tok), tok).backward()
        F.cross_entropy(ref(tok)[:, :-1].reshape(-1, V), tok[:, 1:].reshape(-1)).backward()
    names = [n for n, _ in net.named_parameters()]
    assert {n: fired.get(n, 0) for n in names} == {n: 2 for n in names}
    for (name, p), q in zip(net.named_parameters(), ref.parameters()):
        assert p.grad.data_ptr() >= arena.grad.data_ptr() and p.grad.data_ptr() < arena.grad.data_ptr() + arena.grad.numel() * 4, name
        torch.testing.assert_close(p.grad, q.grad, rtol=2e-3, atol=2e-4, msg=name)
