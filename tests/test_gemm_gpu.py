"""tcgen05 GEMM (K10) and Gram (K3) against fp64 oracles."""
import pytest
import torch

from ddl25spring_b200.ops import gemm as G

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _rel_err(got, exp):
    return float((got.double() - exp).abs().max() / exp.abs().max().clamp_min(1e-12))


@pytest.mark.parametrize("M,N,K", [(128, 128, 32), (128, 128, 256), (100, 10, 128), (300, 200, 96), (64, 64, 4096),
                                   (800, 128, 9216)])
@pytest.mark.parametrize("precise", [False, True])
def test_gemm_tn(cuda_lib, M, N, K, precise):
    a, b = torch.randn(M, K, device=DEV), torch.randn(N, K, device=DEV)
    bias = torch.randn(N, device=DEV)
    got = G.gemm(a, b, bias=bias, act="relu", precise=precise)
    exp = torch.relu(a.double() @ b.double().T + bias.double())
    err = _rel_err(got, exp)
    assert err < (1e-5 if precise else 3e-3), err


@pytest.mark.parametrize("precise", [False, True])
def test_gemm_mn_major_operands(cuda_lib, precise):
    Bsz, nin, nout = 100, 256, 128
    x, w, dy = (torch.randn(Bsz, nin, device=DEV), torch.randn(nout, nin, device=DEV),
                torch.randn(Bsz, nout, device=DEV))
    tol = 1e-5 if precise else 3e-3
    dx = G.linear_dgrad(dy, w, precise=precise)
    assert _rel_err(dx, dy.double() @ w.double()) < tol
    dw = G.linear_wgrad(dy, x, precise=precise)
    assert _rel_err(dw, dy.double().T @ x.double()) < tol


@pytest.mark.parametrize("fixup", [False, True])
def test_gemm_grouped_and_splitk(cuda_lib, fixup, monkeypatch):
    monkeypatch.setattr(G, "USE_FIXUP", fixup)      # separate reduce kernel / in-kernel cooperative fix-up
    Gn, M, N, K = 8, 100, 128, 1024
    a, b = torch.randn(Gn, M, K, device=DEV), torch.randn(Gn, N, K, device=DEV)
    bias = torch.randn(Gn, N, device=DEV)
    exp = torch.baddbmm(bias.double()[:, None, :], a.double(), b.double().transpose(1, 2))
    got = G.gemm(a, b, bias=bias, precise=True)
    assert _rel_err(got, exp) < 1e-5
    got2 = G.gemm(a, b, bias=bias, precise=True, n_split=4)
    assert _rel_err(got2, exp) < 1e-5
    got3 = G.gemm(a, b, bias=bias, precise=True, n_split=4)
    assert torch.equal(got2, got3), "split-K must be deterministic"


def test_gemm_bf16(cuda_lib):
    M, N, K = 256, 128, 512
    a, b = torch.randn(M, K, device=DEV).bfloat16(), torch.randn(N, K, device=DEV).bfloat16()
    got = G.gemm(a, b)
    assert _rel_err(got, a.double() @ b.double().T) < 1e-5   # exact products of bf16 inputs, fp32 accumulate


@pytest.mark.parametrize("K,D", [(64, 1_199_872 + 128), (10, 50_000), (128, 300_000)])
def test_gram(cuda_lib, K, D):
    u = 0.01 * torch.randn(K, D, device=DEV)
    exp = u.double() @ u.double().T
    got = G.gram_tcgen05(u)
    assert _rel_err(got, exp) < 3e-3
    gotp = G.gram_tcgen05(u, precise=True)
    assert _rel_err(gotp, exp) < 5e-5
    assert torch.equal(G.gram_tcgen05(u), got), "Gram must be deterministic"


@pytest.mark.parametrize("K,D", [(3, 4096), (8, 1_000_000 + 32 * 7), (16, 250_016), (33, 64_000), (64, 320_000), (5, 32), (8, 96)])
def test_gram_row_packed(cuda_lib, K, D):
    """K <= 64 rows, D % 32 == 0: the row-packed SYRK kernel (128 / pow2(K) column slices share one MMA tile), plain TF32,
    3xTF32 and 3xTF32 with the centre subtracted in shared memory (Gram of the deltas), strided rows, vs fp64."""
    torch.manual_seed(K + D)
    buf = torch.randn(K, D + 64, device=DEV)                 # strided rows: ld = D + 64
    u = buf[:, :D]
    u.mul_(0.01)
    exp = u.double() @ u.double().T
    got = G.gram_tcgen05(u)
    assert _rel_err(got, exp) < 3e-3
    assert torch.equal(G.gram_tcgen05(u), got), "Gram must be deterministic"
    assert _rel_err(G.gram_tcgen05(u, precise=True), exp) < 5e-5
    c = 0.5 + 0.1 * torch.randn(D, device=DEV)               # large common offset: raw-weight Gram would lose the deltas
    w = (u + c[None, :]).contiguous()
    gotc = G.gram_tcgen05(w, center=c)
    d = (w.double() - c.double()[None, :])
    assert _rel_err(gotc, d @ d.T) < 5e-5
    out = torch.full((K, K), 7.0, device=DEV)
    assert G.gram_tcgen05(u, out=out, n_split=1) is out and _rel_err(out, exp) < 3e-3


@pytest.mark.parametrize("dims", [[14, 16, 14, 16], [64, 32], [16, 8, 12], [7, 4, 6, 13, 40]])
def test_gather_gemm_local_parts(cuda_lib, dims):
    """K7 with all parts in local memory: act(cat(parts) @ W.T + b)."""
    M, N = 200, 128
    ld = (max(dims) + 3) // 4 * 4
    parts = [torch.randn(M, ld, device=DEV) for _ in dims]
    for p_, d_ in zip(parts, dims):
        p_[:, d_:] = 0          # contract: padding columns of a part are zero
    W, b = torch.randn(N, sum(dims), device=DEV), torch.randn(N, device=DEV)
    got = G.gather_linear_ptrs([p.data_ptr() for p in parts], dims, [ld] * len(dims), M, W, b, act="leaky_relu")
    cat = torch.cat([p[:, :d] for p, d in zip(parts, dims)], 1).double()
    exp = torch.nn.functional.leaky_relu(cat @ W.double().T + b.double(), 0.01)
    assert _rel_err(got, exp) < 1e-5


@pytest.mark.parametrize("precise", [False, True])
@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, False), (True, True)])
def test_persistent_gemm_many_tiles(cuda_lib, precise, a_mn, b_mn):
    """Grids of >= 2 x 148 tiles run the persistent kernel (one CTA per SM walks the tiles, accumulator double-buffered in
    TMEM): every operand layout, a ragged last tile in both dimensions, bias + ReLU, and accumulation into C."""
    torch.manual_seed(11)
    M, N, K = 128 * 9 + 32, 128 * 40 + 96, 288            # 10 x 41 = 410 tiles
    a = torch.randn((K, M) if a_mn else (M, K), device=DEV)
    b = torch.randn((K, N) if b_mn else (N, K), device=DEV)
    bias = torch.randn(N, device=DEV)
    A = a.double().T if a_mn else a.double()
    B = b.double() if b_mn else b.double().T
    tol = 1e-5 if precise else 3e-3
    exp = torch.relu(A @ B + bias.double())
    got = G.gemm(a, b, a_mn=a_mn, b_mn=b_mn, bias=bias, act="relu", precise=precise)
    assert _rel_err(got, exp) < tol
    c0 = torch.randn(M, N, device=DEV)
    out = c0.clone()
    G.gemm(a, b, a_mn=a_mn, b_mn=b_mn, precise=precise, out=out, alpha=0.5, beta=1.0)
    assert _rel_err(out, 0.5 * (A @ B) + c0.double()) < tol
    again = G.gemm(a, b, a_mn=a_mn, b_mn=b_mn, bias=bias, act="relu", precise=precise)
    assert torch.equal(got, again)
