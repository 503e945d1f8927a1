"""Development probe: cost of MN-major operands in the tcgen05 GEMM (same M,N,K; 4 layout combinations)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "homebrewnlp-mtf_b200"))
sys.path.insert(0, ROOT)
from benchmarks.kernel_probe import timeit  # noqa: E402
from obst import native as nat  # noqa: E402

nat.load()
bf = torch.bfloat16
for (M, N, K) in ((65536, 512, 8192), (65536, 512, 4096), (4096, 512, 65536), (512, 8192, 65536)):
    a_k = torch.randn(M, K, device="cuda", dtype=bf)
    b_k = torch.randn(N, K, device="cuda", dtype=bf)
    a_m = a_k.t().contiguous()  # [K, M]
    b_m = b_k.t().contiguous()  # [K, N]
    d = torch.empty(M, N, device="cuda", dtype=bf)
    for name, A, B in (("A K-major, B K-major", nat.operand(a_k, K, 1), nat.operand(b_k, K, 1)),
                       ("A K-major, B MN-major", nat.operand(a_k, K, 1), nat.operand(b_m, 1, N)),
                       ("A MN-major, B K-major", nat.operand(a_m, 1, M), nat.operand(b_k, K, 1)),
                       ("A MN-major, B MN-major", nat.operand(a_m, 1, M), nat.operand(b_m, 1, N))):
        ms = timeit(lambda: nat.gemm(A, B, d, M, N, K, d_ld=N, backend=2))
        print(f"M={M} N={N} K={K}  {name:24s} {ms:7.3f} ms  {2 * M * N * K / ms / 1e9:7.1f} TFLOP/s", flush=True)
