"""Times the GEMMs of the folded group norm (forward epilogues 6 / 10, dgrad epilogues 0 / 11, wgrad epilogue 8 with
and without the row dots) and the separate norm backward they replace, at 32big_mixer's W_out shape -- a development
probe (CUDA events, L2 flushed between iterations)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "homebrewnlp-mtf_b200"))
from obst import native as nat  # noqa: E402

DEV = "cuda"


def timeit(fn, iters=10):
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=DEV)
    for _ in range(2):
        fn()
    tot = 0.0
    for _ in range(iters):
        flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize()
        tot += s.elapsed_time(e)
    return tot / iters


def main():
    nat.load()
    T = int(os.environ.get("PROBE_T", 131072))
    G, Kc, Np = 8, 1024, 512
    bf = torch.bfloat16
    h = torch.randn(T, G * Kc, device=DEV, dtype=bf).relu_()
    wg_t = torch.randn(G, Np, Kc, device=DEV, dtype=bf) * 0.03
    w = torch.randn(G, Kc, Np, device=DEV, dtype=bf) * 0.03
    res = torch.randn(T, G, Np, device=DEV, dtype=bf)
    dy = torch.randn(T, G, Np, device=DEV, dtype=bf)
    y = torch.empty(T, G, Np, device=DEV, dtype=bf)
    stats = torch.rand(T * G, 2, device=DEV) + 0.5
    cvec = torch.randn(G, Np, device=DEV)
    dvec = torch.randn(G, Np, device=DEV)
    gamma_f = torch.rand(G, Kc, device=DEV) + 0.5
    gamma = gamma_f.to(bf)
    dots = torch.empty(T, G, 4, 2, device=DEV)
    coeffs = torch.randn(T, G, 4, device=DEV)
    dn = torch.empty(T, G * Kc, device=DEV, dtype=bf)
    dh = torch.empty(T, G * Kc, device=DEV, dtype=bf)
    dw = torch.zeros(G, Kc, Np, device=DEV)
    dgamma = torch.zeros(G, Kc, device=DEV)
    dbeta = torch.zeros(G, Kc, device=DEV)
    fl = 2.0 * T * G * Kc * Np

    def fwd(epi, aux):
        nat.gemm(nat.operand(h, G * Kc, 1, s_glo=Kc), nat.operand(wg_t, Kc, 1, s_glo=Kc * Np), y, T, Np, Kc, GLO=G,
                 d_ld=G * Np, d_glo=Np, R=res, epi=epi, backend=2, aux=aux)

    def dgrad(epi, out, R, aux):
        nat.gemm(nat.operand(dy, G * Np, 1, s_glo=Np), nat.operand(w, Np, 1, s_glo=Kc * Np), out, T, Kc, Np, GLO=G,
                 d_ld=G * Kc, d_glo=Kc, R=R, epi=epi, backend=2, aux=aux)

    def wgrad(aux):
        nat.gemm(nat.operand(h, 1, G * Kc, s_glo=Kc), nat.operand(dy, 1, G * Np, s_glo=Np), dw, Kc, Np, T, GLO=G,
                 d_ld=Np, d_glo=Kc * Np, accumulate=True, epi=8, backend=2, aux=aux)

    rows = [
        ("W_out fwd, epilogue 6", lambda: fwd(6, (stats, cvec, dvec))),
        ("W_out fwd, epilogue 10 (+ dots)", lambda: fwd(10, (stats, cvec, dvec, dy, dots))),
        ("W_out dgrad, epilogue 0", lambda: dgrad(0, dn, None, ())),
        ("W_out dgrad, epilogue 4 (relu mask)", lambda: dgrad(4, dn, h, ())),
        ("W_out dgrad, epilogue 11", lambda: dgrad(11, dh, h, (coeffs,))),
        ("W_out wgrad, epilogue 8", lambda: wgrad((gamma_f,))),
        ("W_out wgrad, epilogue 8 + dots", lambda: wgrad((gamma_f, w, dgamma))),
        ("gn_bwd relu-mask saved stats", lambda: nat.gn_bwd(dn, h, gamma, gamma, dh, dgamma, dbeta, G, Kc,
                                                           relu_mask=True, stats=stats)),
    ]
    only = os.environ.get("PROBE_ONLY", "")  # substring of a row's name: run just those (short enough for ncu)
    iters = int(os.environ.get("PROBE_ITERS", 10))
    for name, fn in rows:
        if only and only not in name:
            continue
        ms = timeit(fn, iters)
        print(f"{name:40s} {ms:8.3f} ms  {fl / ms / 1e9:8.1f} TFLOP/s", flush=True)


if __name__ == "__main__":
    main()
