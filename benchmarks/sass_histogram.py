"""SASS opcode histogram of the built objects (cuobjdump -sass): the tracked evidence that the GEMM is tcgen05 / TMEM /
TMA code and that the bandwidth kernels use cp.async rings.  Runs here (no GPU needed):

    python benchmarks/sass_histogram.py > profiles/r2_sass_histogram.md
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BUILD = os.path.join(ROOT, "homebrewnlp-mtf_b200", "csrc", "build")
# SASS mnemonic prefixes that only exist for the Blackwell tensor / TMA paths (B200_PROFILING.md, "What proves ...")
MARKERS = ["UTCHMMA", "UTCQMMA", "UTCBAR", "UTCCP", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP", "UTMAPF", "SYNCS",
           "LDGSTS", "HMMA", "HGMMA"]


def histogram(obj):
    out = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True, check=True).stdout
    per_kernel = collections.OrderedDict()
    cur = None
    for line in out.splitlines():
        m = re.match(r"\s+Function : (\S+)", line)
        if m:
            name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
            cur = per_kernel.setdefault(name.split("(")[0], collections.Counter())
            continue
        m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\w+\s+)?([A-Z][\w.]*)", line)
        if m and cur is not None:
            cur[m.group(1)] += 1
    return per_kernel


def main():
    objs = sys.argv[1:] or ["gemm_tc.o", "optimizer_ring.o", "gn_fwd_ring.o", "gn_bwd_ring.o"]
    print("# SASS opcode histograms (cuobjdump -sass of homebrewnlp-mtf_b200/csrc/build/*.o, sm_100a)\n")
    for o in objs:
        per_kernel = histogram(os.path.join(BUILD, o))
        print(f"## {o}\n")
        for name, cnt in per_kernel.items():
            total = sum(cnt.values())
            marks = {k: sum(v for op, v in cnt.items() if op.startswith(k)) for k in MARKERS}
            marks = {k: v for k, v in marks.items() if v}
            if o != "gemm_tc.o" and not marks:
                continue
            print(f"* `{name}`: {total} instructions; Blackwell / async-copy mnemonics: "
                  + (", ".join(f"`{k}*` x {v}" for k, v in marks.items()) or "none"))
            if o == "gemm_tc.o":
                full = sorted((op for op in cnt if any(op.startswith(k) for k in MARKERS)), key=lambda x: -cnt[x])
                print("  * exact forms: " + ", ".join(f"`{op}` x {cnt[op]}" for op in full))
                top = ", ".join(f"`{op}` {v}" for op, v in cnt.most_common(12))
                print(f"  * most frequent: {top}")
        print()


if __name__ == "__main__":
    main()
