"""Development probe: is the GEMM main loop bound by bytes staged per SM?  Same M, K; N = 256 / 128 tiles."""
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
M, K = 65536, 8192
for N in (512, 256, 128):
    a = torch.randn(M, K, device="cuda", dtype=bf)
    b = torch.randn(N, K, device="cuda", dtype=bf)
    d = torch.empty(M, N, device="cuda", dtype=bf)
    ms = timeit(lambda: nat.gemm(nat.operand(a, K, 1), nat.operand(b, K, 1), d, M, N, K, d_ld=N, backend=2))
    print(f"M={M} N={N} K={K}: {ms:7.3f} ms {2 * M * N * K / ms / 1e9:7.1f} TFLOP/s", flush=True)
    ms = timeit(lambda: torch.matmul(a, b.t()))
    print(f"   cuBLAS          : {ms:7.3f} ms {2 * M * N * K / ms / 1e9:7.1f} TFLOP/s", flush=True)
