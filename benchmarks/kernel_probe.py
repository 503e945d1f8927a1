"""Times the individual kernels at 32big_mixer shapes (CUDA events, L2-cold inputs) -- a development probe."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "homebrewnlp-mtf_b200"))
from obst import native as nat  # noqa: E402

DEV = "cuda"


def timeit(fn, iters=10, warm=3):
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=DEV)
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    tot = 0.0
    for _ in range(iters):
        flush.zero_()
        s = torch.cuda.Event(enable_timing=True)
        e = torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize()
        tot += s.elapsed_time(e)
    return tot / iters


def main():
    if os.environ.get("PROBE_LIB"):  # A/B of two builds on the same box
        nat.LIB_PATH = os.environ["PROBE_LIB"]
    nat.load()
    T = int(os.environ.get("PROBE_T", 65536))
    F, I, H, K, K2, S = 4096, 512, 8, 512, 1024, 512
    bf = torch.bfloat16
    x = torch.randn(T, F, device=DEV, dtype=bf)
    win = torch.randn(F, I, device=DEV, dtype=bf)
    h1 = torch.empty(T, I, device=DEV, dtype=bf)
    wmid = torch.randn(I, H * K2, device=DEV, dtype=bf)
    h2 = torch.empty(T, H * K2, device=DEV, dtype=bf)
    wout = torch.randn(H, K2, K, device=DEV, dtype=bf)
    y = torch.randn(T, F, device=DEV, dtype=bf)
    rows = []

    def rep(name, ms, flops=None, bytes_=None):
        s = f"{name:34s} {ms:8.3f} ms"
        if flops:
            s += f"  {flops / ms / 1e9:8.1f} TFLOP/s"
        if bytes_:
            s += f"  {bytes_ / ms / 1e6:8.1f} GB/s"
        print(s, flush=True)

    only = os.environ.get("PROBE_ONLY", "")  # "gn": only the bandwidth kernels (short enough to run under ncu)
    dh2 = torch.empty(T, H * K2, device=DEV, dtype=bf)
    out = torch.empty(T, F, device=DEV, dtype=bf)
    for be, tag in ((2, "tc"),) if only in ("", "gemm") else ():
        ms = timeit(lambda: nat.gemm(nat.operand(x, F, 1), nat.operand(win, 1, I), h1, T, I, F, d_ld=I, epi=1, backend=be))
        rep(f"W_in fwd [{tag}] T x512 x4096", ms, 2 * T * I * F)
        ms = timeit(lambda: nat.gemm(nat.operand(h1, I, 1), nat.operand(wmid, 1, H * K2), h2, T, H * K2, I, d_ld=H * K2, epi=1, backend=be))
        rep(f"W_mid fwd [{tag}] T x8192 x512", ms, 2 * T * I * H * K2)
        ms = timeit(lambda: nat.gemm(nat.operand(h2, H * K2, 1, s_glo=K2), nat.operand(wout, 1, K, s_glo=K2 * K), y, T, K, K2, GLO=H, d_ld=F, d_glo=K, R=y, epi=2, backend=be))
        rep(f"W_out fwd grouped [{tag}]", ms, 2 * T * H * K2 * K)
        # the same forward GEMMs with the weight stored transposed (B K-major instead of MN-major)
        win_t = win.t().contiguous()
        ms = timeit(lambda: nat.gemm(nat.operand(x, F, 1), nat.operand(win_t, F, 1), h1, T, I, F, d_ld=I, epi=1, backend=be))
        rep(f"W_in fwd, W^T stored [{tag}]", ms, 2 * T * I * F)
        wmid_t = wmid.t().contiguous()
        ms = timeit(lambda: nat.gemm(nat.operand(h1, I, 1), nat.operand(wmid_t, I, 1), h2, T, H * K2, I, d_ld=H * K2, epi=1, backend=be))
        rep(f"W_mid fwd, W^T stored [{tag}]", ms, 2 * T * I * H * K2)
        wout_t = wout.transpose(1, 2).contiguous()
        ms = timeit(lambda: nat.gemm(nat.operand(h2, H * K2, 1, s_glo=K2), nat.operand(wout_t, K2, 1, s_glo=K2 * K), y, T, K, K2, GLO=H, d_ld=F, d_glo=K, R=y, epi=2, backend=be))
        rep(f"W_out fwd grouped, W^T stored [{tag}]", ms, 2 * T * H * K2 * K)
        # dgrads
        ms = timeit(lambda: nat.gemm(nat.operand(y, F, 1, s_glo=K), nat.operand(wout, K, 1, s_glo=K2 * K), dh2, T, K2, K, GLO=H, d_ld=H * K2, d_glo=K2, backend=be))
        rep(f"W_out dgrad grouped [{tag}]", ms, 2 * T * H * K2 * K)
        ms = timeit(lambda: nat.gemm(nat.operand(h2, H * K2, 1), nat.operand(wmid, H * K2, 1), h1, T, I, H * K2, d_ld=I, R=h1, epi=4, backend=be))
        rep(f"W_mid dgrad [{tag}]", ms, 2 * T * I * H * K2)
        ms = timeit(lambda: nat.gemm(nat.operand(h1, I, 1), nat.operand(win, I, 1), x, T, F, I, d_ld=F, backend=be))
        rep(f"W_in dgrad [{tag}]", ms, 2 * T * I * F)
        # wgrads
        gw = torch.zeros(F, I, device=DEV)
        ms = timeit(lambda: nat.gemm(nat.operand(x, 1, F), nat.operand(h1, 1, I), gw, F, I, T, d_ld=I, accumulate=True, backend=be))
        rep(f"W_in wgrad [{tag}]", ms, 2 * T * I * F)
        gm = torch.zeros(I, H * K2, device=DEV)
        ms = timeit(lambda: nat.gemm(nat.operand(h1, 1, I), nat.operand(h2, 1, H * K2), gm, I, H * K2, T, d_ld=H * K2, accumulate=True, backend=be))
        rep(f"W_mid wgrad [{tag}]", ms, 2 * T * I * H * K2)
        go = torch.zeros(H, K2, K, device=DEV)
        ms = timeit(lambda: nat.gemm(nat.operand(h2, 1, H * K2, s_glo=K2), nat.operand(y, 1, F, s_glo=K), go, K2, K, T, GLO=H, d_ld=K, d_glo=K2 * K, accumulate=True, backend=be))
        rep(f"W_out wgrad grouped [{tag}]", ms, 2 * T * H * K2 * K)
        # token mixing
        Bn = T // S
        mix = torch.tril(torch.randn(H, S, S, device=DEV) * 0.01).to(bf)
        ms = timeit(lambda: nat.gemm(nat.operand(mix, S, 1, s_glo=S * S), nat.operand(x, 1, F, s_glo=K, s_ghi=S * F), out, S, K, S, GLO=H, GHI=Bn, d_ld=F, d_glo=K, d_ghi=S * F, tri=1, backend=be))
        rep(f"mix fwd [{tag}] (causal flops)", ms, 2 * T * F * (S + 1) / 2)
        ms = timeit(lambda: nat.gemm(nat.operand(mix, 1, S, s_glo=S * S), nat.operand(x, 1, F, s_glo=K, s_ghi=S * F), out, S, K, S, GLO=H, GHI=Bn, d_ld=F, d_glo=K, d_ghi=S * F, tri=2, backend=be))
        rep(f"mix dx [{tag}] (causal flops)", ms, 2 * T * F * (S + 1) / 2)
        dm = torch.zeros(H, S, S, device=DEV)
        ms = timeit(lambda: nat.gemm(nat.operand(y, F, 1, s_ko=S * F, s_glo=K), nat.operand(x, F, 1, s_ko=S * F, s_glo=K), dm, S, S, K, KO=Bn, GLO=H, d_ld=S, d_glo=S * S, tri=3, accumulate=True, backend=be))
        rep(f"mix dM [{tag}] (causal flops)", ms, 2 * T * F * (S + 1) / 2)
        # layout experiment: the same products with the operands head-major, [batch, heads, seq, features]
        # (every (batch, head) slab contiguous) instead of the stream's [batch, seq, heads, features]
        HM = dict(s_glo=S * K, s_ghi=H * S * K)
        ms = timeit(lambda: nat.gemm(nat.operand(mix, S, 1, s_glo=S * S), nat.operand(x, 1, K, **HM), out, S, K, S, GLO=H, GHI=Bn, d_ld=K, d_glo=S * K, d_ghi=H * S * K, tri=1, backend=be))
        rep(f"mix fwd, head-major in+out [{tag}]", ms, 2 * T * F * (S + 1) / 2)
        ms = timeit(lambda: nat.gemm(nat.operand(mix, S, 1, s_glo=S * S), nat.operand(x, 1, K, **HM), out, S, K, S, GLO=H, GHI=Bn, d_ld=F, d_glo=K, d_ghi=S * F, tri=1, backend=be))
        rep(f"mix fwd, head-major in [{tag}]", ms, 2 * T * F * (S + 1) / 2)
        ms = timeit(lambda: nat.gemm(nat.operand(mix, S, 1, s_glo=S * S), nat.operand(x, 1, F, s_glo=K, s_ghi=S * F), out, S, K, S, GLO=H, GHI=Bn, d_ld=K, d_glo=S * K, d_ghi=H * S * K, tri=1, backend=be))
        rep(f"mix fwd, head-major out [{tag}]", ms, 2 * T * F * (S + 1) / 2)
        ms = timeit(lambda: nat.gemm(nat.operand(mix, 1, S, s_glo=S * S), nat.operand(x, 1, K, **HM), out, S, K, S, GLO=H, GHI=Bn, d_ld=K, d_glo=S * K, d_ghi=H * S * K, tri=2, backend=be))
        rep(f"mix dx, head-major in+out [{tag}]", ms, 2 * T * F * (S + 1) / 2)
        ms = timeit(lambda: nat.gemm(nat.operand(y, K, 1, s_ko=H * S * K, s_glo=S * K), nat.operand(x, K, 1, s_ko=H * S * K, s_glo=S * K), dm, S, S, K, KO=Bn, GLO=H, d_ld=S, d_glo=S * S, tri=3, accumulate=True, backend=be))
        rep(f"mix dM, head-major operands [{tag}]", ms, 2 * T * F * (S + 1) / 2)
    if only == "gemm":
        return
    # cuBLAS yardstick on this box
    a = torch.randn(8192, 8192, device=DEV, dtype=bf)
    b = torch.randn(8192, 8192, device=DEV, dtype=bf)
    ms = timeit(lambda: torch.matmul(a, b))
    rep("torch.matmul 8192^3 (cuBLAS)", ms, 2 * 8192 ** 3)
    ms = timeit(lambda: torch.matmul(x, win))
    rep("torch.matmul W_in shape (cuBLAS)", ms, 2 * T * I * F)
    ms = timeit(lambda: torch.matmul(h1, wmid, out=h2))
    rep("torch.matmul W_mid fwd shape (cuBLAS)", ms, 2 * T * I * H * K2)
    ms = timeit(lambda: torch.matmul(h2, wmid.t(), out=h1))
    rep("torch.matmul W_mid dgrad shape (cuBLAS)", ms, 2 * T * I * H * K2)
    # bandwidth kernels
    gam = torch.ones(H, K, device=DEV, dtype=bf)
    bet = torch.zeros(H, K, device=DEV, dtype=bf)
    ms = timeit(lambda: nat.gn_fwd(x, gam, bet, y, H, K, 0))
    rep("gn_fwd [T,8,512]", ms, bytes_=2 * T * F * 2)
    dg = torch.zeros(H, K, device=DEV)
    db = torch.zeros(H, K, device=DEV)
    ms = timeit(lambda: nat.gn_bwd(y, x, gam, bet, out, dg, db, H, K))
    rep("gn_bwd [T,8,512]", ms, bytes_=3 * T * F * 2)
    st = torch.empty(T * H, 2, device=DEV)
    nat.gn_fwd(x, gam, bet, c_tmp := torch.empty_like(x), H, K, 0, stats=st)
    ms = timeit(lambda: nat.gn_fwd(x, gam, bet, c_tmp, H, K, 0, stats=st))
    rep("gn_fwd +stats [T,8,512]", ms, bytes_=2 * T * F * 2)
    ms = timeit(lambda: nat.gn_bwd(y, x, gam, bet, out, dg, db, H, K, stats=st))
    rep("gn_bwd saved stats [T,8,512]", ms, bytes_=3 * T * F * 2)
    ms = timeit(lambda: nat.gn_bwd(y, x, gam, bet, out, dg, db, H, K, 1, stats=st))
    rep("gn_bwd + gelu saved stats", ms, bytes_=3 * T * F * 2)
    ms = timeit(lambda: nat.gn_bwd(y, x, gam, bet, out, dg, db, H, K, 0, False, True, stats=st))
    rep("gn_bwd dx+= saved stats", ms, bytes_=4 * T * F * 2)
    ms = timeit(lambda: nat.gn_fwd(x, gam, bet, y, H, K, 1))
    rep("gn_fwd + gelu [T,8,512]", ms, bytes_=2 * T * F * 2)
    ms = timeit(lambda: nat.gn_bwd(y, x, gam, bet, out, dg, db, H, K, 1))
    rep("gn_bwd + gelu [T,8,512]", ms, bytes_=3 * T * F * 2)
    ms = timeit(lambda: nat.gn_bwd(y, x, gam, bet, out, dg, db, H, K, 0, False, True))
    rep("gn_bwd dx+= [T,8,512]", ms, bytes_=4 * T * F * 2)
    gam2 = torch.ones(H, K2, device=DEV, dtype=bf)
    bet2 = torch.zeros(H, K2, device=DEV, dtype=bf)
    n2 = torch.empty_like(h2)
    h2.normal_()
    ms = timeit(lambda: nat.gn_fwd(h2, gam2, bet2, n2, H, K2, 0))
    rep("gn_fwd [T,8,1024]", ms, bytes_=2 * T * H * K2 * 2)
    dg2 = torch.zeros(H, K2, device=DEV)
    db2 = torch.zeros(H, K2, device=DEV)
    ms = timeit(lambda: nat.gn_bwd(dh2, h2, gam2, bet2, n2, dg2, db2, H, K2, 0, True))
    rep("gn_bwd relu-mask [T,8,1024]", ms, bytes_=3 * T * H * K2 * 2)
    st2 = torch.empty(T * H, 2, device=DEV)
    nat.gn_fwd(h2, gam2, bet2, n2, H, K2, 0, stats=st2)
    ms = timeit(lambda: nat.gn_bwd(dh2, h2, gam2, bet2, n2, dg2, db2, H, K2, 0, True, stats=st2))
    rep("gn_bwd relu-mask saved stats [T,8,1024]", ms, bytes_=3 * T * H * K2 * 2)
    ms = timeit(lambda: nat.gn_bwd(dh2, h2, gam2, bet2, n2, dg2, db2, H, K2, 0, False, False, stats=st2))
    rep("gn_bwd saved stats [T,8,1024]", ms, bytes_=3 * T * H * K2 * 2)
    # 256-wide heads (32ctx_mixer / 32mixer_group): [T, 16, 256] has the bytes of [T, 8, 512]
    x4 = x.view(T, 16, 256)
    gam4 = torch.ones(16, 256, device=DEV, dtype=bf)
    bet4 = torch.zeros(16, 256, device=DEV, dtype=bf)
    st4 = torch.empty(T * 16, 2, device=DEV)
    dg4 = torch.zeros(16, 256, device=DEV)
    db4 = torch.zeros(16, 256, device=DEV)
    ms = timeit(lambda: nat.gn_fwd(x4, gam4, bet4, c_tmp, 16, 256, 0, stats=st4))
    rep("gn_fwd +stats [T,16,256]", ms, bytes_=2 * T * F * 2)
    ms = timeit(lambda: nat.gn_bwd(y, x4, gam4, bet4, out, dg4, db4, 16, 256, stats=st4))
    rep("gn_bwd saved stats [T,16,256]", ms, bytes_=3 * T * F * 2)
    ms = timeit(lambda: nat.gn_bwd(y, x4, gam4, bet4, out, dg4, db4, 16, 256, 1, False, True, stats=st4))
    rep("gn_bwd + gelu dx+= saved stats [T,16,256]", ms, bytes_=4 * T * F * 2)
    ms = timeit(lambda: nat.axpby(x, y, 1.0, out))
    rep("axpby [T,4096]", ms, bytes_=3 * T * F * 2)
    c = torch.empty_like(x)
    ms = timeit(lambda: c.copy_(x))
    rep("torch copy_ [T,4096]", ms, bytes_=2 * T * F * 2)


if __name__ == "__main__":
    main()
