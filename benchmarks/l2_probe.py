"""Development probe: does a micro-batch small enough for the activations to stay in the 126 MB L2 pay off?
Times the forward kernel chain of one depth (both sub-blocks) per token at several micro-batch sizes, replayed from
a CUDA graph (so the host launch cost does not hide the answer)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "homebrewnlp-mtf_b200"))
from obst import native as nat  # noqa: E402

nat.load()
bf = torch.bfloat16
DEV = "cuda"
F, I, H, K, K2, S = 4096, 512, 8, 512, 1024, 512
win = torch.randn(F, I, device=DEV, dtype=bf) * 0.02
wmid = torch.randn(I, H * K2, device=DEV, dtype=bf) * 0.02
wout = torch.randn(H, K2, K, device=DEV, dtype=bf) * 0.02
mixa = torch.tril(torch.randn(H, S, S, device=DEV) * 0.01).to(bf)
gam = torch.ones(H, K, device=DEV, dtype=bf)
bet = torch.zeros(H, K, device=DEV, dtype=bf)
gam2 = torch.ones(H, K2, device=DEV, dtype=bf)
bet2 = torch.zeros(H, K2, device=DEV, dtype=bf)

for T in (4096, 8192, 16384, 32768, 131072):
    Bn = T // S
    x1 = torch.randn(T, F, device=DEV, dtype=bf)
    x2 = torch.randn(T, F, device=DEV, dtype=bf)
    a = torch.empty(T, F, device=DEV, dtype=bf)
    h1 = torch.empty(T, I, device=DEV, dtype=bf)
    h2 = torch.empty(T, H * K2, device=DEV, dtype=bf)
    n = torch.empty(T, H * K2, device=DEV, dtype=bf)
    m1 = torch.empty(T, F, device=DEV, dtype=bf)
    c = torch.empty(T, F, device=DEV, dtype=bf)

    def chain():
        # block_config[0]
        nat.gn_fwd(x2, gam, bet, a, H, K, 0)
        nat.gemm(nat.operand(a, F, 1), nat.operand(win, 1, I), h1, T, I, F, d_ld=I, epi=1)
        nat.gemm(nat.operand(h1, I, 1), nat.operand(wmid, 1, H * K2), h2, T, H * K2, I, d_ld=H * K2, epi=1)
        nat.gn_fwd(h2, gam2, bet2, n, H, K2, 0)
        nat.gemm(nat.operand(n, H * K2, 1, s_glo=K2), nat.operand(wout, 1, K, s_glo=K2 * K), x1, T, K, K2, GLO=H,
                 d_ld=F, d_glo=K, R=x1, epi=2)
        # block_config[1] (streams swapped)
        nat.gn_fwd(x1, gam, bet, a, H, K, 0)
        nat.gemm(nat.operand(mixa, S, 1, s_glo=S * S), nat.operand(a, 1, F, s_glo=K, s_ghi=S * F), m1, S, K, S, GLO=H,
                 GHI=Bn, d_ld=F, d_glo=K, d_ghi=S * F, tri=1)
        nat.gn_fwd(m1, gam, bet, c, H, K, 1)
        nat.gemm(nat.operand(mixa, S, 1, s_glo=S * S), nat.operand(c, 1, F, s_glo=K, s_ghi=S * F), x2, S, K, S, GLO=H,
                 GHI=Bn, d_ld=F, d_glo=K, d_ghi=S * F, R=x2, epi=2, tri=1)

    reps = max(1, 131072 // T)  # same number of tokens per measurement
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        chain()
        chain()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            for _ in range(reps):
                chain()
    torch.cuda.synchronize()
    for _ in range(3):
        g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    print(f"T={T:7d} ({reps:3d} chains/graph): {ms:8.3f} ms per 131072 tokens -> {ms * 1e6 / (reps * T):7.2f} ns/token", flush=True)
