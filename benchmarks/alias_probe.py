"""Does the relative placement of a GEMM's input and output buffers matter?  The token-mixing product reads
x[b, :, h, cols] and writes out[b, :, h, cols]: with both tensors at the same offset inside equally aligned
allocations every CTA's loads and stores fall on the same DRAM channels / banks.  Times the causal mix with the
output buffer shifted by various byte offsets (development probe)."""
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
    H, K, S = 8, 512, 512
    F = H * K
    Bn = T // S
    bf = torch.bfloat16
    x = torch.randn(T, F, device=DEV, dtype=bf)
    mix = torch.tril(torch.randn(H, S, S, device=DEV) * 0.01).to(bf)
    big = torch.empty(T * F + (64 << 20), device=DEV, dtype=bf)
    print("x %% 2MiB = %d, big %% 2MiB = %d" % (x.data_ptr() % (2 << 20), big.data_ptr() % (2 << 20)))
    for off_bytes in (0, 128, 512, 1024, 2048, 4096, 8192, 16384, 65536, 1 << 20, (1 << 20) + 4096, 3 << 20, (32 << 20) + 8192):
        out = big[off_bytes // 2: off_bytes // 2 + T * F].view(T, F)
        ms = timeit(lambda: nat.gemm(nat.operand(mix, S, 1, s_glo=S * S), nat.operand(x, 1, F, s_glo=K, s_ghi=S * F), out,
                                     S, K, S, GLO=H, GHI=Bn, d_ld=F, d_glo=K, d_ghi=S * F, tri=1, backend=2))
        print(f"mix fwd, out shifted by {off_bytes:>9d} B: {ms:.3f} ms", flush=True)
    # layout experiment, alternating so that clock drift hits both alike: stream layout [b, s, h, f] against head-major
    # [b, h, s, f] for the B operand and / or the output
    out = big[:T * F].view(T, F)
    SL = dict(s_glo=K, s_ghi=S * F)
    HM = dict(s_glo=S * K, s_ghi=H * S * K)
    variants = {
        "stream in, stream out": (nat.operand(x, 1, F, **SL), dict(d_ld=F, d_glo=K, d_ghi=S * F)),
        "head-major in, head-major out": (nat.operand(x, 1, K, **HM), dict(d_ld=K, d_glo=S * K, d_ghi=H * S * K)),
        "head-major in, stream out": (nat.operand(x, 1, K, **HM), dict(d_ld=F, d_glo=K, d_ghi=S * F)),
        "stream in, head-major out": (nat.operand(x, 1, F, **SL), dict(d_ld=K, d_glo=S * K, d_ghi=H * S * K)),
    }
    for rnd in range(3):
        for name, (b_op, dk) in variants.items():
            ms = timeit(lambda: nat.gemm(nat.operand(mix, S, 1, s_glo=S * S), b_op, out, S, K, S, GLO=H, GHI=Bn, tri=1,
                                         backend=2, **dk), 6)
            print(f"round {rnd} mix fwd, {name}: {ms:.3f} ms", flush=True)
    # the same for a bandwidth kernel: group norm forward
    g = torch.ones(H, K, device=DEV, dtype=bf)
    for off_bytes in (0, 4096, 1 << 20, (1 << 20) + 4096):
        out = big[off_bytes // 2: off_bytes // 2 + T * F].view(T, F)
        ms = timeit(lambda: nat.gn_fwd(x, g, g, out, H, K))
        print(f"gn_fwd, out shifted by {off_bytes:>9d} B: {ms:.3f} ms", flush=True)


if __name__ == "__main__":
    main()
