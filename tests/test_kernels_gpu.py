"""Kernel-level GPU tests: every C-ABI entry point against a plain torch fp32 statement of the same op.

These are floating-point kernels, so the checker here is torch fp32 on the same device (the model-level
parity tests against oracle/ are in test_parity_gpu.py).  Tolerances are written at each comparison:
fp32 paths 1e-5 relative (north_star), bf16 paths 2e-2 relative to the tensor's scale.
"""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def nat():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from obst import native
    native.load()
    return native


def rel_err(a, b):
    a = a.double()
    b = b.double()
    return float((a - b).norm() / (b.norm() + 1e-30))


def max_rel(a, b):
    a = a.double()
    b = b.double()
    return float((a - b).abs().max() / (b.abs().max() + 1e-30))


DEV = "cuda"


def randn(*shape, dtype=torch.bfloat16, scale=1.0, seed=0):
    g = torch.Generator(device="cpu").manual_seed(seed)
    return (torch.randn(*shape, generator=g) * scale).to(dtype).to(DEV)


# ---------------------------------------------------------------------------------------------------
# GEMM: both back ends, every operand layout the training step uses
# ---------------------------------------------------------------------------------------------------
BACKENDS = [1, 2]


@pytest.mark.parametrize("backend", BACKENDS)
@pytest.mark.parametrize("M,N,K", [(256, 512, 512), (200, 64, 96), (128, 256, 32), (384, 8192, 512)])
def test_gemm_kmajor_kmajor(nat, backend, M, N, K):
    a = randn(M, K, seed=1)
    b = randn(N, K, seed=2)
    d = torch.empty(M, N, dtype=torch.bfloat16, device=DEV)
    nat.gemm(nat.operand(a, K, 1), nat.operand(b, K, 1), d, M, N, K, d_ld=N, backend=backend)
    ref = a.float() @ b.float().t()
    assert max_rel(d, ref) < 1e-2  # bf16 output rounding (2^-8) dominates


@pytest.mark.parametrize("backend", BACKENDS)
@pytest.mark.parametrize("epi", [0, 1, 2, 3, 4])
def test_gemm_b_mnmajor_epilogues(nat, backend, epi):
    M, N, K = 256, 512, 256
    a = randn(M, K, seed=3)
    b = randn(K, N, seed=4)  # [K, N]: N contiguous = MN-major B (a weight read in place by the forward pass)
    r = randn(M, N, seed=5, scale=4.0)
    d = r.clone()
    nat.gemm(nat.operand(a, K, 1), nat.operand(b, 1, N), d, M, N, K, d_ld=N, R=d if epi >= 2 else None, epi=epi,
             backend=backend)
    acc = a.float() @ b.float()
    ref = {0: acc, 1: acc.relu(), 2: r.float() + acc, 3: r.float() - acc, 4: acc * (r.float() > 0)}[epi]
    assert max_rel(d, ref) < 1e-2


@pytest.mark.parametrize("backend", BACKENDS)
@pytest.mark.parametrize("M,N,K,splits", [(512, 512, 4096, 0), (64, 4096, 2048, 3), (4096, 512, 1024, 1)])
def test_gemm_wgrad_mnmajor_accumulate(nat, backend, M, N, K, splits):
    a = randn(K, M, seed=6)  # activations [T, f]: reduction index is the row
    b = randn(K, N, seed=7)
    d0 = randn(M, N, dtype=torch.float32, seed=8)
    d = d0.clone()
    nat.gemm(nat.operand(a, 1, M), nat.operand(b, 1, N), d, M, N, K, d_ld=N, accumulate=True, splits=splits,
             backend=backend)
    ref = d0 + a.float().t() @ b.float()
    assert max_rel(d, ref) < 1e-4  # fp32 accumulate of exact bf16 products; only summation order differs


@pytest.mark.parametrize("backend", BACKENDS)
def test_gemm_grouped_heads(nat, backend):
    T, H, J, K = 512, 8, 256, 128
    n = randn(T, H, J, seed=9)
    w = randn(H, J, K, seed=10, scale=0.1)
    x1 = randn(T, H, K, seed=11)
    y = x1.clone()
    # y[t,h,:] = x1[t,h,:] + n[t,h,:] @ w[h]   (W_out, residual add fused, in place)
    nat.gemm(nat.operand(n, H * J, 1, s_glo=J), nat.operand(w, 1, K, s_glo=J * K), y, T, K, J, GLO=H, d_ld=H * K,
             d_glo=K, R=y, epi=2, backend=backend)
    ref = x1.float() + torch.einsum("thj,hjk->thk", n.float(), w.float())
    assert max_rel(y, ref) < 1e-2


@pytest.mark.parametrize("backend", BACKENDS)
@pytest.mark.parametrize("S,K", [(256, 64), (512, 128)])
def test_gemm_token_mix_forward_and_dx(nat, backend, S, K):
    Bn, H = 3, 4
    F = H * K
    mix = torch.tril(randn(H, S, S, seed=12, scale=0.05).float()).to(torch.bfloat16)
    x = randn(Bn, S, H, K, seed=13)
    out = torch.empty_like(x)
    # out[b,s,h,k] = sum_{t<=s} mix[h,s,t] x[b,t,h,k]
    nat.gemm(nat.operand(mix, S, 1, s_glo=S * S), nat.operand(x, 1, F, s_glo=K, s_ghi=S * F), out, S, K, S, GLO=H,
             GHI=Bn, d_ld=F, d_glo=K, d_ghi=S * F, tri=1, backend=backend)
    ref = torch.einsum("hst,bthk->bshk", mix.float(), x.float())
    assert max_rel(out, ref) < 1e-2
    # dx[b,t,h,k] = sum_{s>=t} mix[h,s,t] dout[b,s,h,k]   (A = mix^T, read MN-major)
    dx = torch.empty_like(x)
    nat.gemm(nat.operand(mix, 1, S, s_glo=S * S), nat.operand(x, 1, F, s_glo=K, s_ghi=S * F), dx, S, K, S, GLO=H,
             GHI=Bn, d_ld=F, d_glo=K, d_ghi=S * F, tri=2, backend=backend)
    ref = torch.einsum("hst,bshk->bthk", mix.float(), x.float())
    assert max_rel(dx, ref) < 1e-2


@pytest.mark.parametrize("backend", BACKENDS)
def test_gemm_token_mix_weight_grad(nat, backend):
    Bn, H, S, K = 5, 4, 256, 64
    F = H * K
    do = randn(Bn, S, H, K, seed=14)
    x = randn(Bn, S, H, K, seed=15)
    dm = torch.zeros(H, S, S, dtype=torch.float32, device=DEV)
    # dM[h,s,t] = [t<=s] sum_{b,k} do[b,s,h,k] x[b,t,h,k]  (outer reduction index = b)
    nat.gemm(nat.operand(do, F, 1, s_ko=S * F, s_glo=K), nat.operand(x, F, 1, s_ko=S * F, s_glo=K), dm, S, S, K,
             KO=Bn, GLO=H, d_ld=S, d_glo=S * S, tri=3, accumulate=True, backend=backend)
    ref = torch.tril(torch.einsum("bshk,bthk->hst", do.float(), x.float()))
    assert max_rel(dm, ref) < 1e-4
    assert float(torch.triu(dm, diagonal=1).abs().max()) == 0.0  # masked entries never receive gradient


def test_gemm_fp32_simt(nat):
    M, N, K = 96, 72, 50
    a = randn(M, K, dtype=torch.float32, seed=16)
    b = randn(K, N, dtype=torch.float32, seed=17)
    d = torch.empty(M, N, dtype=torch.float32, device=DEV)
    nat.gemm(nat.operand(a, K, 1), nat.operand(b, 1, N), d, M, N, K, d_ld=N, dtype_in=torch.float32)
    ref = (a.double() @ b.double())
    assert max_rel(d, ref) < 1e-5


def test_gemm_tc_matches_simt_bitwise_shape_sweep(nat):
    """auto backend must pick tcgen05 for bf16 and agree with the SIMT kernel on identical inputs."""
    for (M, N, K) in [(128, 128, 64), (1024, 1024, 1024), (640, 320, 448)]:
        a = randn(M, K, seed=M)
        b = randn(N, K, seed=N + 1)
        d1 = torch.empty(M, N, dtype=torch.float32, device=DEV)
        d2 = torch.empty(M, N, dtype=torch.float32, device=DEV)
        nat.gemm(nat.operand(a, K, 1), nat.operand(b, K, 1), d1, M, N, K, d_ld=N, backend=1)
        nat.gemm(nat.operand(a, K, 1), nat.operand(b, K, 1), d2, M, N, K, d_ld=N, backend=0)
        assert max_rel(d2, d1) < 1e-5


# ---------------------------------------------------------------------------------------------------
# group norm / activations / residual / embedding / cross-entropy / casts
# ---------------------------------------------------------------------------------------------------
def gelu_ref(x):
    return 0.5 * x * (1 + torch.tanh(0.044715 * x ** 3 + math.sqrt(2 / math.pi) * x))


def gn_ref(x, gamma, beta, act):
    xh = x - x.mean(-1, keepdim=True)
    y = xh * torch.rsqrt((xh * xh).mean(-1, keepdim=True) + 1e-5) * gamma + beta
    return gelu_ref(y) if act else y


@pytest.mark.parametrize("dtype,tol", [(torch.float32, 1e-5), (torch.bfloat16, 2e-2)])
@pytest.mark.parametrize("width", [32, 64, 512, 1024])
@pytest.mark.parametrize("act", [0, 1])
@pytest.mark.parametrize("saved_stats", [False, True])
def test_gn_fwd_bwd(nat, dtype, tol, width, act, saved_stats):
    T, H = 300, 8
    x = randn(T, H, width, dtype=dtype, seed=20, scale=2.0)
    gamma = (1 + 0.02 * torch.randn(H, width)).to(dtype).to(DEV)
    beta = (0.02 * torch.randn(H, width)).to(dtype).to(DEV)
    dy = randn(T, H, width, dtype=dtype, seed=21)
    y = torch.empty_like(x)
    stats = torch.full((T * H, 2), float("nan"), device=DEV) if saved_stats else None
    nat.gn_fwd(x, gamma, beta, y, H, width, act, stats=stats)
    if saved_stats:  # (mean, rstd) per row, for the backward kernel
        xd = x.double().view(T * H, width)
        assert rel_err(stats[:, 0], xd.mean(-1)) < 1e-5 + 1e-6
        assert rel_err(stats[:, 1], (xd.var(-1, unbiased=False) + 1e-5).rsqrt()) < 1e-5
    xr = x.double().requires_grad_(True)
    gr = gamma.double().requires_grad_(True)
    br = beta.double().requires_grad_(True)
    yr = gn_ref(xr, gr, br, act)
    assert rel_err(y, yr.detach()) < tol
    yr.backward(dy.double())
    dx0 = randn(T, H, width, dtype=dtype, seed=22)
    dx = dx0.clone()
    dg = torch.zeros(H, width, device=DEV)
    db = torch.zeros(H, width, device=DEV)
    nat.gn_bwd(dy, x, gamma, beta, dx, dg, db, H, width, act=act, dx_accumulate=True, stats=stats)
    assert rel_err(dx, dx0.double() + xr.grad) < tol
    assert rel_err(dg, gr.grad) < tol
    assert rel_err(db, br.grad) < tol


# The kernels the benchmark runs: from 4096 tokens on, bf16 rows of 512 / 1024 elements take the streaming ("ring")
# kernels of gn_fwd_ring.cu / gn_bwd_ring.cu, which the small case above never reaches.  Every template instantiation
# (width x fused gelu x saved statistics x dx accumulation x ReLU mask) against the torch float64 statement; 8200
# tokens is not a multiple of the 128-token block, so the ragged tail is covered too.
@pytest.mark.parametrize("T", [4096, 8200])
@pytest.mark.parametrize("width", [256, 512, 1024])
@pytest.mark.parametrize("act", [0, 1])
def test_gn_fwd_streaming_kernels(nat, T, width, act):
    H = 8
    x = randn(T, H, width, seed=60, scale=2.0) + 0.5
    gamma = (1 + 0.02 * torch.randn(H, width)).to(torch.bfloat16).to(DEV)
    beta = (0.02 * torch.randn(H, width)).to(torch.bfloat16).to(DEV)
    yr = gn_ref(x.double(), gamma.double(), beta.double(), act)
    xd = x.double().view(T * H, width)
    for with_stats in (False, True):
        y = torch.full_like(x, float("nan"))
        stats = torch.full((T * H, 2), float("nan"), device=DEV) if with_stats else None
        before = nat.path_count(nat.PATH_GN_FWD_RING)
        nat.gn_fwd(x, gamma, beta, y, H, width, act, stats=stats)
        assert nat.path_count(nat.PATH_GN_FWD_RING) == before + 1, "the streaming forward kernel did not run"
        assert rel_err(y, yr) < 2e-2 and bool(torch.isfinite(y.float()).all())
        assert max_rel(y, yr) < 2e-2  # no row may be off: a wrong row would hide in the L2 norm of 8200 x 8 rows
        if with_stats:
            assert rel_err(stats[:, 0], xd.mean(-1)) < 1e-5
            assert rel_err(stats[:, 1], (xd.var(-1, unbiased=False) + 1e-5).rsqrt()) < 1e-5
    # no scale / no shift
    y = torch.empty_like(x)
    nat.gn_fwd(x, None, None, y, H, width, act)
    assert rel_err(y, gn_ref(x.double(), 1.0, 0.0, act)) < 2e-2


@pytest.mark.parametrize("T", [4096, 8200])
@pytest.mark.parametrize("width,relu_mask", [(256, False), (512, False), (1024, False), (1024, True), (512, True)])
@pytest.mark.parametrize("act", [0, 1])
@pytest.mark.parametrize("dx_accumulate", [False, True])
def test_gn_bwd_streaming_kernels(nat, T, width, relu_mask, act, dx_accumulate):
    H = 8
    x = randn(T, H, width, seed=61, scale=2.0) + 0.5
    if relu_mask:
        x = x.relu()  # the W_mid path: the normalised tensor is a ReLU output
    gamma = (1 + 0.02 * torch.randn(H, width)).to(torch.bfloat16).to(DEV)
    beta = (0.02 * torch.randn(H, width)).to(torch.bfloat16).to(DEV)
    dy = randn(T, H, width, seed=62)
    y = torch.empty_like(x)
    stats = torch.empty(T * H, 2, device=DEV)
    nat.gn_fwd(x, gamma, beta, y, H, width, act, stats=stats)
    xr = x.double().requires_grad_(True)
    gr = gamma.double().requires_grad_(True)
    br = beta.double().requires_grad_(True)
    gn_ref(xr, gr, br, act).backward(dy.double())
    dx_ref = xr.grad * (x > 0) if relu_mask else xr.grad
    dx0 = randn(T, H, width, seed=63)
    for saved in (True, False):
        dx = dx0.clone() if dx_accumulate else torch.full_like(x, float("nan"))
        dg = torch.zeros(H, width, device=DEV)
        db = torch.zeros(H, width, device=DEV)
        before = nat.path_count(nat.PATH_GN_BWD_RING)
        nat.gn_bwd(dy, x, gamma, beta, dx, dg, db, H, width, act=act, relu_mask=relu_mask, dx_accumulate=dx_accumulate,
                   stats=stats if saved else None)
        if saved:  # the training step always passes the forward's statistics: that is the streaming kernel's case
            assert nat.path_count(nat.PATH_GN_BWD_RING) == before + 1, "the streaming backward kernel did not run"
        want = dx0.double() + dx_ref if dx_accumulate else dx_ref
        assert rel_err(dx, want) < 2e-2 and bool(torch.isfinite(dx.float()).all())
        assert max_rel(dx, want) < 4e-2
        assert rel_err(dg, gr.grad) < 2e-2
        assert rel_err(db, br.grad) < 2e-2


def test_gn_bwd_relu_mask(nat):
    T, H, W = 64, 8, 64
    x = randn(T, H, W, dtype=torch.float32, seed=23).relu()
    dy = randn(T, H, W, dtype=torch.float32, seed=24)
    dx = torch.empty_like(x)
    nat.gn_bwd(dy, x, None, None, dx, None, None, H, W, relu_mask=True)
    xr = x.double().requires_grad_(True)
    gn_ref(xr, 1.0, 0.0, 0).backward(dy.double())
    assert rel_err(dx, xr.grad * (x > 0)) < 1e-5


@pytest.mark.parametrize("dtype,tol", [(torch.float32, 1e-5), (torch.bfloat16, 2e-2)])
@pytest.mark.parametrize("kind", [0, 1])
def test_activations(nat, dtype, tol, kind):
    n = 8 * 1000 + 3
    x = randn(n, dtype=dtype, seed=25, scale=2.0)
    dy = randn(n, dtype=dtype, seed=26)
    y = torch.empty_like(x)
    dx = torch.empty_like(x)
    nat.act_fwd(kind, x, y)
    nat.act_bwd(kind, x, dy, dx)
    xr = x.double().requires_grad_(True)
    yr = xr.relu() if kind == 0 else gelu_ref(xr)
    yr.backward(dy.double())
    assert rel_err(y, yr.detach()) < tol
    assert rel_err(dx, xr.grad) < tol


@pytest.mark.parametrize("dtype,tol", [(torch.float32, 1e-5), (torch.bfloat16, 2e-2)])
@pytest.mark.parametrize("name", ["sigmoid", "tanh", "lecun_tanh", "silu", "mish", "mtf_mish", "softsign", "exp"])
def test_other_registered_activations(nat, name, dtype, tol):
    """Every other entry of ACTIVATIONS (src/model/activation.py:190-200) against the oracle, including the
    reference's own non-analytic backward of silu and mish."""
    from oracle import obst_oracle as orc
    from obst.model.activation import ACTIVATION_KINDS
    kind = ACTIVATION_KINDS[name]
    x = randn(4099, dtype=dtype, seed=50, scale=2.0)
    dy = randn(4099, dtype=dtype, seed=51)
    y, dx = torch.empty_like(x), torch.empty_like(x)
    nat.act_fwd(kind, x, y)
    nat.act_bwd(kind, x, dy, dx)
    xn, gn = x.double().cpu().numpy(), dy.double().cpu().numpy()
    assert rel_err(y, torch.from_numpy(orc.ACT_FWD[name](xn)).to(DEV)) < tol
    assert rel_err(dx, torch.from_numpy(orc.ACT_BWD[name](xn) * gn).to(DEV)) < tol


@pytest.mark.parametrize("dtype,tol", [(torch.float32, 1e-5), (torch.bfloat16, 2e-2)])
def test_glu_gate(nat, dtype, tol):
    a, b, dy = (randn(3001, dtype=dtype, seed=s) for s in (52, 53, 54))
    out, da, db = torch.empty_like(a), torch.empty_like(a), torch.empty_like(a)
    nat.gate_fwd(a, b, out)
    nat.gate_bwd(a, b, dy, da, db)
    ad, bd = a.double().requires_grad_(True), b.double().requires_grad_(True)
    ref = ad * torch.sigmoid(bd)
    ref.backward(dy.double())
    assert rel_err(out, ref.detach()) < tol and rel_err(da, ad.grad) < tol and rel_err(db, bd.grad) < tol


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_axpby(nat, dtype):
    a = randn(4099, dtype=dtype, seed=27)
    b = randn(4099, dtype=dtype, seed=28)
    out = torch.empty_like(a)
    nat.axpby(a, b, -1.0, out)
    assert torch.equal(out, (a.float() - b.float()).to(dtype))
    nat.axpby(a, b, 1.0, a)
    assert torch.equal(a, (out.float() + 2 * b.float()).to(dtype)) or rel_err(a, out.float() + 2 * b.float()) < 1e-2


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_embedding_gather_scatter(nat, dtype):
    V, W, T = 256, 64, 5000
    table = randn(V, W, dtype=dtype, seed=29)
    idx = torch.randint(0, V, (T,), dtype=torch.int32, device=DEV)
    out = torch.empty(T, W, dtype=dtype, device=DEV)
    nat.embed_gather(idx, table, out)
    assert torch.equal(out, table[idx.long()])
    dout = randn(T, W, dtype=dtype, seed=30)
    dt = torch.zeros(V, W, device=DEV)
    nat.embed_scatter_add(idx, dout, dt)
    ref = torch.zeros(V, W, device=DEV, dtype=torch.float64).index_add_(0, idx.long(), dout.double())
    assert rel_err(dt, ref) < 1e-5


@pytest.mark.parametrize("dtype,tol", [(torch.float32, 1e-5), (torch.bfloat16, 2e-2)])
def test_xent(nat, dtype, tol):
    T, V = 3000, 256
    logits = randn(T, V, dtype=dtype, seed=31, scale=3.0)
    tgt = torch.randint(0, V, (T,), dtype=torch.int32, device=DEV)
    dl = torch.empty_like(logits)
    sums = torch.zeros(3, dtype=torch.float64, device=DEV)
    n_global = 2 * T  # the normaliser is the GLOBAL token count, not this shard's
    nat.xent_fwd_bwd(logits, tgt, dl, sums, 1.0 / n_global, 1e-4)
    lr = logits.double().requires_grad_(True)
    logz = torch.logsumexp(lr, -1)
    nll = (logz - lr.gather(1, tgt.long()[:, None])[:, 0]).sum()
    z2 = (logz ** 2).sum()
    loss = nll / n_global + 1e-4 * z2 / n_global
    loss.backward()
    nll, z2 = nll.detach(), z2.detach()
    assert abs(float(sums[0]) - float(nll)) / float(nll) < 1e-5
    assert abs(float(sums[1]) - float(z2)) / float(z2) < 1e-5
    assert float(sums[2]) == float((logits.float().argmax(-1) == tgt).sum())
    assert rel_err(dl, lr.grad) < tol
    # in place, as the model calls it (dlogits IS logits): same sums, same gradient
    inplace = logits.clone()
    sums2 = torch.zeros(3, dtype=torch.float64, device=DEV)
    nat.xent_fwd_bwd(inplace, tgt, inplace, sums2, 1.0 / n_global, 1e-4)
    assert torch.equal(inplace, dl)
    assert torch.allclose(sums2, sums, rtol=1e-12, atol=0)


def test_cast_param_tril(nat):
    H, S = 3, 64
    src = randn(H, S, S, dtype=torch.float32, seed=32)
    dst = torch.empty(H, S, S, dtype=torch.bfloat16, device=DEV)
    nat.cast_param(src, dst, tril_s=S)
    assert torch.equal(dst, torch.tril(src).to(torch.bfloat16))
    dst2 = torch.empty(H, S, S, dtype=torch.float32, device=DEV)
    nat.cast_param(src, dst2)
    assert torch.equal(dst2, src)


@pytest.mark.parametrize("shape,dtype", [((3, 64, 96), torch.bfloat16), ((1, 33, 7), torch.float32),
                                         ((8, 512, 512), torch.bfloat16), ((2, 1, 65), torch.bfloat16)])
def test_transpose_param(nat, shape, dtype):
    src = randn(*shape, dtype=dtype, seed=33)
    dst = torch.full((shape[0], shape[2], shape[1]), 7.0, dtype=dtype, device=DEV)
    nat.transpose_param(src, dst)
    assert torch.equal(dst, src.transpose(1, 2).contiguous())  # a copy: bit-exact
    with pytest.raises(nat.NativeError):
        nat.transpose_param(src, src)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("vocab", [256, 50, 1000])
def test_sample_step(nat, dtype, vocab):
    """obst_sample_step against the defining lines (src/run/inference.py:88-95) on random logits: greedy, with
    supplied uniforms (bit-identical choice), position 0, and the built-in generator's statistics."""
    B, S = 5, 12
    logits = randn(B, S, vocab, dtype=dtype, seed=40, scale=2.0)
    tok0 = torch.randint(0, vocab, (B, S), dtype=torch.int32, device=DEV)
    lf = logits.float()
    for pos in (1, 7, S - 1):
        tok = tok0.clone()
        nat.sample_step(logits, tok, pos, 0.0)
        exp = tok0.clone()
        exp[:, pos] = lf[:, pos - 1].argmax(-1).int()
        assert torch.equal(tok, exp)
        u = torch.rand(B, vocab, device=DEV)
        tok = tok0.clone()
        nat.sample_step(logits, tok, pos, 1.5, uniform=u)
        uc = u.clamp(1e-9, 0.99999994)
        exp[:, pos] = (lf[:, pos - 1] + torch.log(-torch.log(uc)) * -1.5).argmax(-1).int()
        assert torch.equal(tok, exp)
    tok = tok0.clone()
    nat.sample_step(logits, tok, 0, 0.7)
    assert torch.equal(tok[:, 1:], tok0[:, 1:]) and int(tok[:, 0].abs().max()) == 0
    # built-in generator: at a huge temperature the choice is uniform over the vocabulary, and seeds differ
    toks = torch.zeros(4096, 2, dtype=torch.int32, device=DEV)
    big = torch.zeros(4096, 2, 8, dtype=dtype, device=DEV)
    nat.sample_step(big, toks, 1, 1e3, seed=1)
    counts = torch.bincount(toks[:, 1].long(), minlength=8).float()
    assert float((counts - 512).abs().max()) < 120          # ~5.6 sigma of Binomial(4096, 1/8)
    toks2 = torch.zeros(4096, 2, dtype=torch.int32, device=DEV)
    nat.sample_step(big, toks2, 1, 1e3, seed=2)
    assert not torch.equal(toks, toks2)
    with pytest.raises(nat.NativeError):
        nat.sample_step(logits, tok0.clone(), S, 0.0)


# ---------------------------------------------------------------------------------------------------
# edge cases: empty inputs, ragged sizes, the largest mixing map of the reference configs
# ---------------------------------------------------------------------------------------------------
def test_empty_inputs_are_no_ops(nat):
    e = torch.empty(0, dtype=torch.bfloat16, device=DEV)
    before = nat.launch_count()
    nat.gn_fwd(e.view(0, 8, 64), None, None, e.view(0, 8, 64), 8, 64)
    nat.gn_bwd(e.view(0, 8, 64), e.view(0, 8, 64), None, None, e.view(0, 8, 64), None, None, 8, 64)
    nat.act_fwd(0, e, e)
    nat.axpby(e, e, 1.0, e)
    nat.xent_fwd_bwd(e.view(0, 256), torch.empty(0, dtype=torch.int32, device=DEV), e.view(0, 256),
                     torch.zeros(3, dtype=torch.float64, device=DEV), 1.0, 1e-4)
    nat.embed_gather(torch.empty(0, dtype=torch.int32, device=DEV), torch.zeros(4, 8, dtype=torch.bfloat16, device=DEV),
                     e.view(0, 8))
    assert nat.launch_count() == before  # nothing was launched, nothing failed


def test_gemm_rejects_bad_descriptors(nat):
    a = randn(128, 64)
    d = torch.empty(128, 128, dtype=torch.bfloat16, device=DEV)
    with pytest.raises(nat.NativeError, match="epilogue 2 needs R"):
        nat.gemm(nat.operand(a, 64, 1), nat.operand(a, 64, 1), d, 128, 128, 64, d_ld=128, epi=2)
    with pytest.raises(nat.NativeError, match="accumulate needs fp32"):
        nat.gemm(nat.operand(a, 64, 1), nat.operand(a, 64, 1), d, 128, 128, 64, d_ld=128, accumulate=True)
    with pytest.raises(nat.NativeError, match="tri 3 needs accumulate"):
        nat.gemm(nat.operand(a, 64, 1), nat.operand(a, 64, 1), d, 128, 128, 64, d_ld=128, tri=3)
    with pytest.raises(nat.NativeError):  # tcgen05 backend cannot address a row stride of 5 elements
        nat.gemm(nat.operand(a, 5, 1), nat.operand(a, 64, 1), d, 16, 128, 4, d_ld=128, backend=2)


@pytest.mark.parametrize("M,N,K", [(1000, 520, 136), (257, 72, 1000), (129, 8, 64)])
def test_gemm_ragged_sizes_tc(nat, M, N, K):
    a = randn(M, K, seed=40)
    b = randn(N, K, seed=41)
    d = torch.full((M + 3, N), 7.0, dtype=torch.bfloat16, device=DEV)  # 3 guard rows must stay untouched
    nat.gemm(nat.operand(a, K, 1), nat.operand(b, K, 1), d, M, N, K, d_ld=N, backend=2)
    assert max_rel(d[:M], a.float() @ b.float().t()) < 1e-2
    assert float((d[M:].float() - 7.0).abs().max()) == 0.0


def test_token_mix_at_the_longest_context_of_the_reference_configs(nat):
    """configs/32ctx_mixer.json: sequence 2048, 256 features per head."""
    Bn, H, S, K = 2, 2, 2048, 256
    F = H * K
    mix = torch.tril(randn(H, S, S, seed=42, scale=0.02).float()).to(torch.bfloat16)
    x = randn(Bn, S, H, K, seed=43)
    out = torch.empty_like(x)
    nat.gemm(nat.operand(mix, S, 1, s_glo=S * S), nat.operand(x, 1, F, s_glo=K, s_ghi=S * F), out, S, K, S, GLO=H,
             GHI=Bn, d_ld=F, d_glo=K, d_ghi=S * F, tri=1)
    assert max_rel(out, torch.einsum("hst,bthk->bshk", mix.float(), x.float())) < 1e-2
    dm = torch.zeros(H, S, S, device=DEV)
    nat.gemm(nat.operand(out, F, 1, s_ko=S * F, s_glo=K), nat.operand(x, F, 1, s_ko=S * F, s_glo=K), dm, S, S, K,
             KO=Bn, GLO=H, d_ld=S, d_glo=S * S, tri=3, accumulate=True)
    ref = torch.tril(torch.einsum("bshk,bthk->hst", out.float(), x.float()))
    assert max_rel(dm, ref) < 1e-4


# ---------------------------------------------------------------------------------------------------
# The group norm between W_mid and W_out folded into the GEMMs around it, forward AND backward
# (obst_gemm epilogues 5..11 + the small kernels of csrc/fused_norm.cu), against float64 autograd of
#   h = relu(h1 @ W_mid);  n = group_norm(h) * gamma + beta;  y = R +/- n @ W_out (per head)
# (src/model/basic.py:47-57,122-126; src/model/normalization.py:22-34).
# ---------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("T,G,Kc,Np,sign", [(512, 8, 1024, 512, -1.0), (384, 4, 512, 256, 1.0), (300, 2, 256, 256, -1.0)])
def test_folded_group_norm_forward_and_backward_chain(nat, T, G, Kc, Np, sign):
    bf = torch.bfloat16
    K1 = 256
    h1 = randn(T, K1, seed=1)
    w_mid = randn(K1, G * Kc, scale=K1 ** -0.5, seed=2)
    gamma = (1.0 + 0.3 * randn(G, Kc, dtype=torch.float32, seed=3)).to(bf)
    beta = (0.2 * randn(G, Kc, dtype=torch.float32, seed=4)).to(bf)
    w_out = randn(G, Kc, Np, scale=Kc ** -0.5, seed=5)
    res = randn(T, G, Np, seed=6)
    dy = randn(T, G, Np, seed=7)
    # ---- device chain -------------------------------------------------------------------------------
    part = torch.empty((T, G * Kc // 128, 2), dtype=torch.float32, device=DEV)
    h = torch.empty((T, G * Kc), dtype=bf, device=DEV)
    nat.gemm(nat.operand(h1, K1, 1), nat.operand(w_mid, 1, G * Kc), h, T, G * Kc, K1, d_ld=G * Kc, epi=7, backend=2,
             aux=(part,))
    stats = torch.empty((T * G, 2), dtype=torch.float32, device=DEV)
    nat.norm_stats_finalize(part, stats, T, G, Kc // 128, Kc)
    wg_t = torch.empty((G, Np, Kc), dtype=bf, device=DEV)
    cvec = torch.empty((G, Np), dtype=torch.float32, device=DEV)
    dvec = torch.empty((G, Np), dtype=torch.float32, device=DEV)
    gamma_f = torch.empty((G, Kc), dtype=torch.float32, device=DEV)
    wg = torch.empty((G, Kc, Np), dtype=bf, device=DEV)
    nat.fold_norm_weight(nat.fold_items([(w_out, gamma, beta, wg_t, cvec, dvec, gamma_f, wg)]), 1, G, Kc, Np)
    assert torch.equal(wg, wg_t.transpose(1, 2))
    y_plain = torch.empty((T, G, Np), dtype=bf, device=DEV)
    nat.gemm(nat.operand(h, G * Kc, 1, s_glo=Kc), nat.operand(wg_t, Kc, 1, s_glo=Kc * Np), y_plain, T, Np, Kc, GLO=G,
             d_ld=G * Np, d_glo=Np, R=res, epi=5 if sign > 0 else 6, backend=2, aux=(stats, cvec, dvec))
    y = torch.empty((T, G, Np), dtype=bf, device=DEV)
    nparts = 2 * ((Np + 255) // 256)
    dots = torch.full((T, G, nparts, 2), float("nan"), dtype=torch.float32, device=DEV)
    nat.gemm(nat.operand(h, G * Kc, 1, s_glo=Kc), nat.operand(wg_t, Kc, 1, s_glo=Kc * Np), y, T, Np, Kc, GLO=G,
             d_ld=G * Np, d_glo=Np, R=res, epi=9 if sign > 0 else 10, backend=2, aux=(stats, cvec, dvec, dy, dots))
    assert torch.equal(y, y_plain)  # the dot products ride along: the output itself is that of epilogue 5 / 6
    dyr = torch.empty_like(dy)
    v = torch.zeros((2, G, Np), dtype=torch.float32, device=DEV)
    nat.rowscale_colsum(dy, stats, dyr, v[0], v[1], T, G, Np)
    dw = torch.zeros((G, Kc, Np), dtype=torch.float32, device=DEV)
    dgamma = torch.zeros((G, Kc), dtype=torch.float32, device=DEV)
    dbeta = torch.zeros((G, Kc), dtype=torch.float32, device=DEV)
    nat.gemm(nat.operand(h, 1, G * Kc, s_glo=Kc), nat.operand(dyr, 1, G * Np, s_glo=Np), dw, Kc, Np, T, GLO=G,
             d_ld=Np, d_glo=Kc * Np, accumulate=True, epi=8, backend=2, aux=(gamma_f, w_out, dgamma))
    nat.norm_affine_grad(w_out, v[0], v[1], dgamma, dbeta, G, Kc, Np)
    nat.norm_wgrad_rank1(dw, gamma, beta, v[0], v[1], G, Kc, Np)
    coeffs = torch.empty((T, G, 4), dtype=torch.float32, device=DEV)
    nat.norm_bwd_coeffs(dots, stats, coeffs, T, G, nparts, Kc)
    dh = torch.empty_like(h)
    nat.gemm(nat.operand(dy, G * Np, 1, s_glo=Np), nat.operand(wg, Np, 1, s_glo=Kc * Np), dh, T, Kc, Np, GLO=G,
             d_ld=G * Kc, d_glo=Kc, R=h, epi=11, backend=2, aux=(coeffs,))
    torch.cuda.synchronize()
    assert torch.isfinite(dots).all()  # every partial slot was written
    # ---- float64 autograd on the same (bf16-valued) inputs; the ReLU output is taken as stored -----------------
    pre = (h1.double() @ w_mid.double()).requires_grad_(True)
    g64 = gamma.double().requires_grad_(True)
    b64 = beta.double().requires_grad_(True)
    w64 = w_out.double().requires_grad_(True)
    hr = torch.relu(pre).view(T, G, Kc)
    mu = hr.mean(-1, keepdim=True)
    var = hr.var(-1, unbiased=False, keepdim=True)
    n = (hr - mu) * torch.rsqrt(var + 1e-5) * g64 + b64
    f = torch.einsum("tgk,gkn->tgn", n, w64)
    y_ref = res.double() + sign * f
    f.backward(dy.double())
    assert rel_err(h.view(T, G, Kc), hr.detach()) < 1e-2
    assert rel_err(y, y_ref.detach()) < 2e-2                      # bf16 tier (north_star): 2e-2
    assert rel_err(dh.view(T, G * Kc), pre.grad) < 2e-2
    assert rel_err(dw, w64.grad) < 2e-2
    assert rel_err(dgamma, g64.grad) < 2e-2
    assert rel_err(dbeta, b64.grad) < 2e-2


def test_folded_norm_epilogues_reject_what_they_cannot_run(nat):
    """Epilogues 9 / 10 / 11 exist in the CTA-pair tensor-core kernel only and need their auxiliary tensors: anything
    else is refused with an error, never run wrongly."""
    bf = torch.bfloat16
    G, Kc, Np = 2, 256, 256
    stats = torch.zeros(128 * G, 2, device=DEV)
    cvec = torch.zeros(G, Np, device=DEV)
    wg_t = torch.zeros(G, Np, Kc, dtype=bf, device=DEV)

    def fwd(T, epi, aux, backend=2):
        h = torch.zeros(T, G * Kc, dtype=bf, device=DEV)
        res = torch.zeros(T, G, Np, dtype=bf, device=DEV)
        y = torch.empty(T, G, Np, dtype=bf, device=DEV)
        nat.gemm(nat.operand(h, G * Kc, 1, s_glo=Kc), nat.operand(wg_t, Kc, 1, s_glo=Kc * Np), y, T, Np, Kc, GLO=G,
                 d_ld=G * Np, d_glo=Np, R=res, epi=epi, backend=backend, aux=aux)

    dy = torch.zeros(128, G, Np, dtype=bf, device=DEV)
    dots = torch.zeros(128, G, 2, 2, device=DEV)
    with pytest.raises(nat.NativeError, match="CTA-pair"):      # one 128-row tile: the 1-CTA kernel would run
        fwd(128, 10, (stats, cvec, cvec, dy, dots))
    with pytest.raises(nat.NativeError, match="aux3"):          # the second tile is missing
        fwd(512, 10, (torch.zeros(512 * G, 2, device=DEV), cvec, cvec))
    with pytest.raises(nat.NativeError, match="SIMT"):          # no fused-norm epilogue in the fp32-FMA kernel
        fwd(512, 10, (torch.zeros(512 * G, 2, device=DEV), cvec, cvec, torch.zeros(512, G, Np, dtype=bf, device=DEV),
                      torch.zeros(512, G, 2, 2, device=DEV)), backend=1)
    with pytest.raises(nat.NativeError, match="aux0"):
        fwd(512, 11, ())
    with pytest.raises(nat.NativeError):
        nat.norm_bwd_coeffs(torch.zeros(8, device=DEV), torch.zeros(8, device=DEV), torch.zeros(9, device=DEV)[1:], 1, 1, 4, 256)


@pytest.mark.parametrize("T,F,z", [(1000, 512, 1e-4), (4096, 4096, 1e-4), (300, 256, 0.0)])
def test_lm_head_with_the_cross_entropy_in_its_epilogue(nat, T, F, z):
    """obst_gemm epilogue 12: LM head -> softmax cross-entropy + z-loss + accuracy as one kernel
    (src/model/__init__.py:144-145 -> src/mtf_wrapper.py:64-71, accuracy :185-187).  Against the two-kernel path
    (same bf16-rounded logits: the sums agree to float rounding, the hit count exactly, the gradient to an ulp of
    bf16) and against float64 autograd; row tails, out-of-range targets (clamped like the separate kernel does)."""
    bf, V = torch.bfloat16, 256
    x = randn(T, F, seed=71, scale=1.0)
    w = randn(F, V, seed=72, scale=3.0 / F ** 0.5)
    tgt = torch.randint(0, V, (T,), dtype=torch.int32, device=DEV, generator=torch.Generator(DEV).manual_seed(5))
    tgt[3], tgt[T - 1] = -7, V + 40
    n_global = 3 * T
    # two kernels
    logits = torch.empty(T, V, dtype=bf, device=DEV)
    nat.gemm(nat.operand(x, F, 1), nat.operand(w, 1, V), logits, T, V, F, d_ld=V, backend=2)
    dl = torch.empty_like(logits)
    sums = torch.zeros(3, dtype=torch.float64, device=DEV)
    nat.xent_fwd_bwd(logits, tgt, dl, sums, 1.0 / n_global, z)
    # one kernel, twice into the same accumulator (it adds)
    fused = torch.full((T, V), float("nan"), dtype=bf, device=DEV)
    fsums = torch.zeros(3, dtype=torch.float64, device=DEV)
    for _ in range(2):
        nat.gemm(nat.operand(x, F, 1), nat.operand(w, 1, V), fused, T, V, F, d_ld=V, backend=2, epi=12,
                 aux=(tgt, None, None, None, fsums), scalars=(1.0 / n_global, z))
    assert torch.allclose(fsums[:2] / 2, sums[:2], rtol=2e-6, atol=0), (fsums, sums)
    assert float(fsums[2]) == 2 * float(sums[2])
    assert torch.isfinite(fused.float()).all()
    same = (fused == dl).float().mean().item()
    assert same > 0.98, same
    assert torch.allclose(fused.float(), dl.float(), rtol=8e-3, atol=1e-12)
    # float64 autograd on the stored logits
    lr = logits.double().requires_grad_(True)
    logz = torch.logsumexp(lr, -1)
    clamped = tgt.long().clamp(0, V - 1)
    nll = (logz - lr.gather(1, clamped[:, None])[:, 0]).sum()
    ((nll + z * (logz ** 2).sum()) / n_global).backward()
    assert abs(float(fsums[0]) / 2 - float(nll.detach())) / float(nll.detach()) < 1e-5
    assert rel_err(fused, lr.grad) < 2e-2


def test_lm_head_loss_epilogue_rejects_what_it_cannot_run(nat):
    bf = torch.bfloat16
    sums = torch.zeros(3, dtype=torch.float64, device=DEV)

    def run(T, V, aux, backend=2):
        x = torch.zeros(T, 64, dtype=bf, device=DEV)
        w = torch.zeros(64, V, dtype=bf, device=DEV)
        nat.gemm(nat.operand(x, 64, 1), nat.operand(w, 1, V), torch.empty(T, V, dtype=bf, device=DEV), T, V, 64,
                 d_ld=V, backend=backend, epi=12, aux=aux, scalars=(1.0, 0.0))

    tgt = torch.zeros(512, dtype=torch.int32, device=DEV)
    with pytest.raises(nat.NativeError, match="N == 256"):
        run(512, 128, (tgt, None, None, None, sums))
    with pytest.raises(nat.NativeError, match="aux4|aux0"):
        run(512, 256, (tgt,))
    with pytest.raises(nat.NativeError, match="CTA-pair"):
        run(128, 256, (tgt, None, None, None, sums))
    with pytest.raises(nat.NativeError, match="SIMT"):
        run(512, 256, (tgt, None, None, None, sums), backend=1)
