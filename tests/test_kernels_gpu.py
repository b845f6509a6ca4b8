"""Numerics of every hand-written sm_100a kernel against a plain PyTorch fp32/fp64 oracle."""
import math

import pytest
import torch

from ddl25spring_b200 import ops
from ddl25spring_b200.ops import reference as ref

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _updates(K, D, seed=0, scale=1.0):
    g = torch.Generator(device="cpu").manual_seed(seed)
    return (scale * torch.randn(K, D, generator=g)).to(DEV)


# ------------------------------------------------------------------ K1
@pytest.mark.parametrize("K,D", [(1, 128), (3, 1024), (10, 1_199_872 + 128), (64, 40_000), (130, 4096)])
def test_weighted_sum(cuda_lib, K, D):
    u = _updates(K, D)
    w = torch.rand(K, dtype=torch.float64)
    w = (w / w.sum()).float()
    out = ops.weighted_sum(u, w)
    exp = ref.weighted_sum(u.double().cpu(), w.double()).float()
    torch.testing.assert_close(out.cpu(), exp, rtol=1e-5, atol=1e-5)


def test_weighted_sum_fedsgd_epilogue(cuda_lib):
    K, D, lr = 7, 8192, 0.05
    g = _updates(K, D)
    p = torch.randn(D, device=DEV)
    w = torch.full((K,), 1.0 / K)
    out = ops.weighted_sum(g, w, base=p, alpha=1.0, beta=-lr)
    exp = p.cpu() - lr * ref.weighted_sum(g.cpu(), w)
    torch.testing.assert_close(out.cpu(), exp, rtol=1e-5, atol=1e-6)


# ------------------------------------------------------------------ K4 / K5
@pytest.mark.parametrize("K", [1, 2, 3, 5, 8, 13, 16, 32, 33, 64, 100])
def test_coordinate_median(cuda_lib, K):
    u = _updates(K, 5000 + K, seed=K)
    for lower in (False, True):
        out = ops.coordinate_median(u, lower=lower)
        exp = ref.coordinate_median(u.cpu(), lower=lower)
        torch.testing.assert_close(out.cpu(), exp, rtol=0, atol=1e-6)
    if K % 2 == 1:
        torch.testing.assert_close(ops.coordinate_median(u).cpu(), u.cpu().median(dim=0).values)


@pytest.mark.parametrize("K,trim", [(5, 1), (10, 2), (16, 3), (64, 12), (64, 0), (21, 10)])
def test_trimmed_mean(cuda_lib, K, trim):
    u = _updates(K, 7777, seed=K + trim)
    out = ops.trimmed_mean(u, trim)
    exp = ref.trimmed_mean(u.double().cpu(), trim).float()
    torch.testing.assert_close(out.cpu(), exp, rtol=1e-5, atol=1e-6)


# ------------------------------------------------------------------ K6
def test_attack_perturb(cuda_lib):
    D = 4096
    u, g = torch.randn(D, device=DEV), torch.randn(D, device=DEV)
    out = ops.attack_perturb(u, g, scale=3.0)
    torch.testing.assert_close(out, ref.sign_flip(u, g, 3.0))
    out = ops.attack_perturb(u, None, scale=1.0)
    torch.testing.assert_close(out, -u)
    noisy = ops.attack_perturb(u, g, scale=1.0, noise=0.5, seed=3)
    d = (noisy - ref.sign_flip(u, g, 1.0))
    assert d.abs().max() <= 0.5 + 1e-6 and d.std() > 0.2


# ------------------------------------------------------------------ K15
def test_sgd_step(cuda_lib):
    n = 1_199_872 + 128
    p, g = torch.randn(n, device=DEV), torch.randn(n, device=DEV)
    p_ref, buf_ref = p.clone().cpu(), torch.zeros(n)
    buf = torch.zeros(n, device=DEV)
    for _ in range(3):
        ops.sgd_step(p, g, lr=0.01, weight_decay=1e-3, momentum=0.9, momentum_buf=buf)
        ref.sgd_step(p_ref, g.cpu(), 0.01, 1e-3, 0.9, buf_ref)
    torch.testing.assert_close(p.cpu(), p_ref, rtol=1e-5, atol=1e-6)
    q = torch.randn(n, device=DEV)
    q_ref = q.clone()
    ops.sgd_step(q, g, lr=0.1)
    torch.testing.assert_close(q, q_ref - 0.1 * g)


@pytest.mark.parametrize("decoupled", [False, True])
def test_adam_step_matches_torch_optim(cuda_lib, decoupled):
    n = 4096
    p0, grads = torch.randn(n), [torch.randn(n) for _ in range(5)]
    tp = torch.nn.Parameter(p0.clone())
    opt = (torch.optim.AdamW if decoupled else torch.optim.Adam)([tp], lr=8e-4, weight_decay=0.01)
    p = p0.clone().to(DEV)
    m, v = torch.zeros(n, device=DEV), torch.zeros(n, device=DEV)
    for i, g in enumerate(grads, start=1):
        tp.grad = g.clone()
        opt.step()
        ops.adam_step(p, g.to(DEV), m, v, i, lr=8e-4, weight_decay=0.01, decoupled=decoupled)
    torch.testing.assert_close(p.cpu(), tp.detach(), rtol=2e-5, atol=2e-6)


# ------------------------------------------------------------------ K13 / K14 / K17
@pytest.mark.parametrize("N,C", [(100, 10), (64, 2), (37, 1000), (8, 32000)])
def test_softmax_xent_index(cuda_lib, N, C):
    x = torch.randn(N, C, device=DEV) * 3
    t = torch.randint(0, C, (N,), device=DEV)
    loss, grad = ops.softmax_xent(x, t)
    xr = x.detach().cpu().double().requires_grad_()
    lr_ = torch.nn.functional.cross_entropy(xr, t.cpu())
    lr_.backward()
    torch.testing.assert_close(loss.cpu().double(), lr_.detach(), rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(grad.cpu().double(), xr.grad, rtol=1e-4, atol=1e-6)


def test_softmax_xent_soft_targets(cuda_lib):
    N, C = 64, 2
    x = torch.randn(N, C, device=DEV)
    t = torch.nn.functional.one_hot(torch.randint(0, C, (N,)), C).float().to(DEV)
    loss, grad = ops.softmax_xent(x, t)
    xr = x.detach().cpu().requires_grad_()
    l2 = torch.nn.CrossEntropyLoss()(xr, t.cpu())
    l2.backward()
    torch.testing.assert_close(loss.cpu(), l2.detach(), rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(grad.cpu(), xr.grad, rtol=1e-4, atol=1e-6)


def test_eval_top1(cuda_lib):
    x = torch.randn(10_000, 10, device=DEV)
    t = torch.randint(0, 10, (10_000,), device=DEV)
    assert int(ops.eval_top1(x, t)) == int((x.argmax(1) == t).sum())


def test_vae_loss(cuda_lib):
    xr, x = torch.randn(64, 31, device=DEV), torch.randn(64, 31, device=DEV)
    mu, lv = torch.randn(64, 16, device=DEV), torch.randn(64, 16, device=DEV) * 0.3
    loss, (gx, gm, gl) = ops.vae_loss(xr, x, mu, lv, want_grad=True)
    a, b, c = (t.detach().cpu().double().requires_grad_() for t in (xr, mu, lv))
    l2 = ref.vae_reparam_loss(a, x.cpu().double(), b, c)
    l2.backward()
    torch.testing.assert_close(loss.cpu().double(), l2.detach(), rtol=1e-5, atol=1e-3)
    torch.testing.assert_close(gx.cpu().double(), a.grad, rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(gm.cpu().double(), b.grad, rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(gl.cpu().double(), c.grad, rtol=1e-4, atol=1e-6)


# ------------------------------------------------------------------ K3 select
@pytest.mark.parametrize("K,f,multi", [(10, 2, 1), (64, 12, 1), (64, 12, 20), (7, 1, 3)])
def test_krum_select_from_exact_gram(cuda_lib, K, f, multi):
    u = _updates(K, 3000, seed=K)
    u[:f] += 5.0  # make the first f rows outliers
    G = (u.double() @ u.double().T).float()
    sel, w, scores = ops.krum_select(u, f, multi, gram_matrix=G)
    exp_scores = ref.krum_scores(u.cpu(), f)
    torch.testing.assert_close(scores.cpu().double(), exp_scores, rtol=2e-3, atol=1e-2)
    exp_sel = torch.argsort(exp_scores, stable=True)[:multi]
    assert set(sel.cpu().tolist()) == set(exp_sel.tolist())
    assert not (set(sel.cpu().tolist()) & set(range(f)))
    assert abs(float(w.sum()) - 1.0) < 1e-6


def test_krum_never_selects_non_finite_rows(cuda_lib):
    """A Byzantine row with NaN / Inf entries (or one scaled up until its Gram entries overflow fp32) must get an
    infinite score: the CPU oracle rejects it (NaN sorts last) and so must the GPU kernel."""
    K, f = 12, 3
    u = _updates(K, 4096, seed=5)
    u[0, 7] = float("nan")
    u[1, 9] = float("inf")
    u[2] *= 1e25                                     # ||u||^2 overflows fp32 -> inf in the Gram matrix
    from ddl25spring_b200.ops.gemm import gram_tcgen05
    G = gram_tcgen05(u)
    sel, w, scores = ops.krum_select(u, f, 2, gram_matrix=G)
    assert not (set(sel.cpu().tolist()) & {0, 1, 2}), sel
    assert torch.isinf(scores[:3]).all() and torch.isfinite(scores[3:]).all(), scores
    agg = ops.weighted_sum(u, w)
    assert torch.isfinite(agg).all()
    # honest rows: same choice as the fp64 oracle restricted to finite rows
    exp = ref.krum_scores(u.cpu(), f)
    order = [i for i in torch.argsort(exp, stable=True).tolist() if i >= 3][:2]
    assert set(sel.cpu().tolist()) == set(order)
