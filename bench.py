#!/usr/bin/env python
"""bench.py -- train tokens/sec of the OBST mixer training step (BASELINE.json's metric).

    python bench.py --gpus N --steps K --warmup W            (N > 1: launched under torchrun, one rank per GPU)
    python bench.py --impl reference ...                      (the CPU restatement timed on the host cores)

One "step" = one full training step (forward, reversible backward with recompute, gradient allreduce,
fused optimizer update) of configs/32big_mixer.json's shape on synthetic tokens with random-init weights.
The GLOBAL batch (1024 x 512 tokens) is fixed: N ranks take 1024/N rows each ("scaling": "strong").
Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "homebrewnlp-mtf_b200"))

METRIC = "train tokens/sec (32big_mixer shape)"
UNIT = "tokens/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="32big_mixer",
                    choices=["32big_mixer", "32ctx_mixer", "32mixer_group", "tiny", "optimizer"],
                    help="'optimizer' = BASELINE.json configs[4]: the fused optimizer update alone over the 32big_mixer "
                         "parameter set (HBM roofline)")
    ap.add_argument("--micro-batch", type=int, default=0, help="rows per micro-batch (0 = auto)")
    ap.add_argument("--depth", type=int, default=0, help="override depth (debug only: makes the number INVALID)")
    ap.add_argument("--cpu-rows", type=int, default=1, help="sequences per step of the CPU baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-fold-norm", action="store_true",
                    help="A/B switch: run the group norm between W_mid and W_out as its own kernel instead of folding "
                         "it into the two GEMMs")
    ap.add_argument("--no-fold-norm-bwd", action="store_true",
                    help="A/B switch: keep the mid norm's backward as a separate kernel (dn written and re-read)")
    ap.add_argument("--no-fuse-loss", action="store_true",
                    help="A/B switch: LM head and softmax cross-entropy as two kernels (logits written and re-read)")
    ap.add_argument("--store-activations", action="store_true",
                    help="extension: keep activations instead of the reference's reversible recompute (reported "
                         "separately; the default run is the reference algorithm)")
    return ap.parse_args()


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            d = json.load(f)
        return {"hbm_gbs": d.get("hbm_gbs", 6650.0), "tf_burst": d.get("bf16_tflops", 1590.0),
                "tf_sustained": d.get("bf16_tflops_sustained", d.get("bf16_tflops", 1590.0)), "source": "measured"}
    return {"hbm_gbs": 6650.0, "tf_burst": 1590.0, "tf_sustained": 1400.0, "source": "fallback"}


class ClockSampler(threading.Thread):
    """nvidia-smi clocks + throttle reasons DURING the timed region (B200_PROFILING.md's clocks line)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index = index
        self.stop_flag = threading.Event()
        self.samples = []

    def run(self):
        while not self.stop_flag.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                      "-i", str(self.index)], capture_output=True, text=True, timeout=5).stdout
                parts = [p.strip() for p in out.strip().split(",")]
                if len(parts) >= 7:
                    self.samples.append(parts)
            except Exception:
                pass
            self.stop_flag.wait(0.2)

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unavailable"]}
        sm = [float(s[0]) for s in self.samples if s[0].replace(".", "").isdigit()]
        mx = [float(s[1]) for s in self.samples if s[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(s[3 + i].lower().startswith("active") for s in self.samples)]
        pw = [float(s[2]) for s in self.samples if s[2].replace(".", "").isdigit()]
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "reasons": reasons, "samples": len(self.samples)}


# ------------------------------------------------------------------------------------------------------------
# CPU baseline / reference arm: the oracle (a port: the reference needs TF 2.7 + mesh-tensorflow, not installable)
# ------------------------------------------------------------------------------------------------------------
def pin_host_threads() -> int:
    """Use every host core for the CPU arm, and say how many.  torch.distributed.run exports OMP_NUM_THREADS=1 to its
    workers, which silently halves-or-worse a NumPy/BLAS baseline launched under it (the round-1 records showed the
    CPU arm dropping from 56 to 26 tokens/s at N >= 2); the limit is therefore set explicitly, before and after the
    BLAS library is loaded."""
    cores = os.cpu_count() or 1
    for var in ("OMP_NUM_THREADS", "OPENBLAS_NUM_THREADS", "MKL_NUM_THREADS"):
        os.environ[var] = str(cores)
    try:
        import threadpoolctl
        import numpy  # noqa: F401  (loads the BLAS the limit applies to)
        threadpoolctl.threadpool_limits(limits=cores)
        used = max([cores] + [int(p.get("num_threads", 0)) for p in threadpoolctl.threadpool_info()])
    except Exception:
        used = cores
    return min(used, cores) if used else cores


def cpu_step_rate(workload: str, rows: int, steps: int, warmup: int, depth: int = 0):
    """The CPU restatement (oracle/, kind "port") on the host cores, bounded: forward + reversible backward are
    timed on `rows` sequences at FULL depth (per-token work), the optimizer update once over the FULL parameter
    set (per-step work); tokens/s of the full step is composed from the two:
        T_step = (tokens_per_step / sample_tokens) * t_fwd_bwd + t_optimizer."""
    import numpy as np
    from oracle import obst_oracle as orc
    from obst.configs import mixer_config
    cfg = mixer_config(workload)
    if depth:
        cfg["depth"] = depth
    full_tokens = cfg["train_batch_size"] * cfg["sequence_length"]
    cfg["train_batch_size"] = rows
    sh = orc.Shapes(cfg)
    cores = pin_host_threads()
    params = orc.init_params_fast(sh, seed=0)
    state = orc.init_opt_state(params)
    tx, ty = orc.synthetic_tokens(sh)
    tokens = rows * sh.sequence_length
    times = []
    grads = None
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        _, _, st = orc.forward(sh, params, tx, ty, dtype=np.float32, n_global=full_tokens)
        grads = orc.backward(sh, st)
        times.append(time.perf_counter() - t0)
        del st
    t_fb = sum(times[warmup:]) / len(times[warmup:])
    t0 = time.perf_counter()
    orc.optimizer_step(sh, params, grads, state, orc.learning_rate(sh, 2048))
    t_opt = time.perf_counter() - t0
    t_step = full_tokens / tokens * t_fb + t_opt
    value = full_tokens / t_step
    sample = (f"{workload} (depth {sh.depth}, all {orc.count_params(sh)} params): fwd + reversible bwd (with recompute) "
              f"timed on {rows} sequence(s) = {tokens} tokens, mean of {len(times) - warmup} after {warmup} warm-up: "
              f"{t_fb:.2f} s; optimizer update over the full parameter set once: {t_opt:.2f} s; composed full step "
              f"({full_tokens} tokens) = {t_step:.0f} s; numpy fp32 + BLAS on {cores} host cores")
    return value, cores, sample, t_fb


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    pin_host_threads()
    if args.workload == "optimizer":
        return run_reference_optimizer(args)
    value, cores, sample, sec = cpu_step_rate(args.workload, args.cpu_rows, args.steps, args.warmup, args.depth)
    line = {"impl": "reference", "metric": METRIC.replace("32big_mixer", args.workload), "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"{args.workload} (bounded sample per step: fwd+bwd of {args.cpu_rows} sequence(s) at "
                                   f"full depth; value = composed full-step tokens/s, see cpu_baseline.sample)"},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0,
            "note": "reference needs tensorflow==2.7.0 + mesh-tensorflow (not installable: py3.12, no network); "
                    "this arm times oracle/obst_oracle.py, the CPU restatement of the same step"}
    print(json.dumps(line), flush=True)


OPT_METRIC = "fused optimizer update, 32big_mixer parameter set: algorithmic GB/s (14 B/param)"


def cpu_optimizer_rate(steps: int, warmup: int):
    """The oracle's optimizer update over the full 32big_mixer parameter set on the host cores: GB/s on the same
    14 B/param algorithmic basis (it is NumPy fp32, op by op: the reference's own formulation of the update)."""
    import numpy as np
    from oracle import obst_oracle as orc
    from obst.configs import mixer_config
    sh = orc.Shapes(mixer_config("32big_mixer"))
    cores = pin_host_threads()
    params = orc.init_params_fast(sh, seed=0)
    state = orc.init_opt_state(params)
    rng = np.random.default_rng(0)
    grads = {k: orc.round_bf16((rng.standard_normal(v.shape) * 1e-3).astype(np.float32)) for k, v in params.items()}
    n = sum(v.size for v in params.values())
    times = []
    for _ in range(warmup + steps):
        t0 = time.perf_counter()
        orc.optimizer_step(sh, params, grads, state, 0.005)
        times.append(time.perf_counter() - t0)
    sec = sum(times[warmup:]) / max(1, len(times[warmup:]))
    return 14.0 * n / sec / 1e9, cores, sec, n


def run_reference_optimizer(args):
    steps, warmup = min(args.steps, 3), min(args.warmup, 1)
    value, cores, sec, n = cpu_optimizer_rate(steps, warmup)
    sample = (f"oracle optimizer_step (NumPy fp32, op by op) over all {n} parameters of 32big_mixer, mean of {steps} "
              f"after {warmup} warm-up: {sec:.2f} s per update on {cores} host cores")
    line = {"impl": "reference", "metric": OPT_METRIC, "value": value, "unit": "GB/s", "n_gpus": args.gpus,
            "steps": steps, "warmup": warmup, "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "optimizer update over the 32big_mixer parameter set (357 tensors)"},
            "cpu_baseline": {"value": value, "unit": "GB/s", "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def run_b200_optimizer(args):
    """BASELINE.json configs[4]: the fused update alone.  Gradients (bf16, ~N(0, 1e-3): SURVEY 8(d)) are resident in
    the hand-over buffer; one step = what FusedOptimizer.step() runs in a training step after the hand-over: norms
    (sum g^2 over the bf16 gradients, weight norms carried over), update, accumulator finalize, transposed-shadow
    refresh.  L2 is flushed between timed iterations (the state is 4.8 GB, but the tail of one update would
    otherwise serve the head of the next)."""
    import torch
    from obst import native
    from obst.configs import mixer_config
    from obst.dataclass import ModelParameter
    from obst.optimizer import _ensure_state
    from obst.run.train import get_train_model
    if int(os.environ.get("WORLD_SIZE", "1")) > 1 and int(os.environ.get("RANK", "0")) != 0:
        return  # the update is replicated, not sharded: one rank measures it ("replicas only")
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
    native.load()
    import contextlib
    import io
    cfg = mixer_config("32big_mixer", train_batch_size=1)
    cfg["device"] = f"cuda:{torch.cuda.current_device()}"
    with contextlib.redirect_stdout(io.StringIO()):
        params = ModelParameter(cfg)
    S = cfg["sequence_length"]
    tok = torch.randint(0, 256, (1, S + 1), dtype=torch.int32, device="cuda")
    # one real step on a single sequence creates every variable, shadow and the optimizer state
    get_train_model(params, None, None, None, tok[:, :-1, None].contiguous(), tok[:, 1:, None].contiguous(), None,
                    None, None, 0)
    st = _ensure_state(params)
    n = st.n_params
    assert st.grad_lp, "the 32big_mixer dtypes hand bf16 gradients to the optimizer"
    host_g = (torch.randn(st.flat_grad_lp.numel(), dtype=torch.float32) * 1e-3).to(torch.bfloat16).pin_memory()
    st.flat_grad_lp.copy_(host_g)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    def prepare():
        st.zero_grad()                       # clears the norm sums and the hand-over bookkeeping
        st.finalized.update(st.item_ranges)  # gradients are already in the hand-over buffer

    def timed(iters, refresh_transposed, e2e=False):
        tot = 0.0
        for _ in range(iters):
            prepare()
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            if e2e:
                st.flat_grad_lp.copy_(host_g, non_blocking=True)
            st.step(0.005, refresh_transposed)
            if e2e:
                probe = st.norms[:2].cpu()  # noqa: F841  (device -> host read of a result of the step)
            e1.record()
            torch.cuda.synchronize()
            tot += e0.elapsed_time(e1)
        return tot / iters

    timed(max(args.warmup, 3), True)
    sampler = ClockSampler(torch.cuda.current_device())
    sampler.start()
    native.launch_count_reset()
    ring0 = native.path_count(native.PATH_OPT_RING)
    iters = max(args.steps, 20)
    ms = timed(iters, True)
    launches = native.launch_count()
    ring = native.path_count(native.PATH_OPT_RING) - ring0
    ms_no_t = timed(iters, False)
    ms_e2e = timed(3, True, e2e=True)
    # the update kernel alone (the dominant kernel): CUDA events around its launch
    lib = native.load()
    orig = lib.obst_opt_update
    ev = []

    def wrapped(*a):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        rc = orig(*a)
        e1.record()
        ev.append((e0, e1))
        return rc

    lib.obst_opt_update = wrapped
    for _ in range(10):
        prepare()
        flush.zero_()
        st.step(0.005, False)
    torch.cuda.synchronize()
    lib.obst_opt_update = orig
    ms_update = sum(a.elapsed_time(b) for a, b in ev) / len(ev)
    sampler.stop_flag.set()
    sampler.join(timeout=2)
    pk = peaks()
    alg_bytes = 14.0 * n
    value = alg_bytes / (ms * 1e-3) / 1e9
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "optimizer_traffic.json")
    if os.path.exists(tpath):
        with open(tpath) as f:
            traffic = json.load(f).get("dram_bytes_per_launch")
    achieved = alg_bytes / (ms_update * 1e-3) / 1e9
    line = {"metric": OPT_METRIC, "value": value, "unit": "GB/s", "n_gpus": 1, "steps": iters,
            "warmup": max(args.warmup, 3), "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"fused optimizer update over the 32big_mixer parameter set: {len(st.variables)} "
                                   f"tensors, {n} parameters; fp32 masters, bf16 momentum / SM3 slots, bf16 gradients "
                                   f"~N(0, 1e-3), bf16 shadows + transposed shadows",
                       "l2": "flushed between timed iterations (256 MiB write)",
                       "step": "obst_opt_norms (sum g^2; weight norms carried) + obst_opt_update + obst_opt_finalize + "
                               "obst_transpose_batched",
                       "bytes_per_param": {"algorithmic": 14, "moved": "norms 2 + update 16 (g 2, p 4+4, m 2+2, "
                                                                       "shadow 2) + transposed shadows ~4"},
                       "ms_without_transposed_refresh": ms_no_t, "ms_update_kernel": ms_update,
                       "streaming_kernel_launches": ring},
            "clocks": sampler.summary(),
            "e2e": {"value": alg_bytes / (ms_e2e * 1e-3) / 1e9, "unit": "GB/s", "ms_per_step": ms_e2e,
                    "h2d_bytes_per_step": host_g.numel() * 2, "d2h_bytes_per_step": 16,
                    "note": "gradients copied from pinned host memory every step: PCIe-bound, reported for the contract"},
            "gpu_launches": launches,
            "roofline": {"bound": "hbm", "kernel": "obst::opt_update_ring_kernel", "achieved": achieved,
                         "peak": pk["hbm_gbs"], "unit": "GB/s", "frac": achieved / pk["hbm_gbs"], "traffic": traffic,
                         "peak_source": f"{pk['source']} (hbm_gbs)",
                         "whole_update_achieved": value, "whole_update_frac": value / pk["hbm_gbs"],
                         "algorithmic_bytes_per_launch": alg_bytes}}
    if not args.no_cpu_baseline:
        v, cores, sec, _ = cpu_optimizer_rate(1, 0)
        line["cpu_baseline"] = {"value": v, "unit": "GB/s", "cores": cores, "kind": "port",
                                "sample": f"oracle optimizer_step over the same {n} parameters once: {sec:.2f} s"}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------------------
# B200 arm
# ------------------------------------------------------------------------------------------------------------
def run_b200(args):
    if args.workload == "optimizer":
        return run_b200_optimizer(args)
    import torch
    import torch.distributed as dist
    from obst import native
    from obst.configs import mixer_config
    from obst.dataclass import ModelParameter
    from obst.parallel import GradReducer, shard_rows
    from obst.run.train import get_train_model, read_scalars

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: there is no CPU fallback for the product path")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    native.load()

    cfg = mixer_config(args.workload)
    if args.depth:
        cfg["depth"] = args.depth
    gb = cfg["train_batch_size"]
    seq = cfg["sequence_length"]
    r0, r1 = shard_rows(gb, rank, world)
    rows = r1 - r0
    mb = args.micro_batch or max(1, min(rows, (32768 if args.store_activations else 131072) // seq))
    cfg["revnet_store_activations"] = bool(args.store_activations)
    cfg["fold_mid_norm"] = not args.no_fold_norm
    cfg["fold_mid_norm_backward"] = not args.no_fold_norm_bwd
    cfg["fuse_lm_head_loss"] = not args.no_fuse_loss
    cfg["micro_batch_size"] = mb
    cfg["device"] = f"cuda:{local_rank}"
    import contextlib
    import io
    with contextlib.redirect_stdout(io.StringIO()) if rank else contextlib.nullcontext():
        params = ModelParameter(cfg)
    if world > 1:
        params.grad_reducer = GradReducer(params)
    # global_step starts at 0 exactly like a fresh reference run: linear warm-up, lr = 0.01 * step / 4096

    # synthetic tokens: x = w[:, :S], y = w[:, 1:S+1]  (src/inputs.py:543-551), pinned host memory
    g = torch.Generator().manual_seed(134567)
    w = torch.randint(0, cfg["vocab_size"], (gb, seq + 1), generator=g, dtype=torch.int32)[r0:r1]
    host_x = w[:, :-1].contiguous().unsqueeze(-1).pin_memory()
    host_y = w[:, 1:].contiguous().unsqueeze(-1).pin_memory()
    dev_x = host_x.cuda(non_blocking=True)
    dev_y = host_y.cuda(non_blocking=True)

    def step_resident():
        return get_train_model(params, None, None, None, dev_x, dev_y, None, None, None, params.manual_global_step)

    def step_e2e():
        x = host_x.cuda(non_blocking=True)
        y = host_y.cuda(non_blocking=True)
        out = get_train_model(params, None, None, None, x, y, None, None, None, params.manual_global_step)
        return read_scalars(params, out[3])

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms: float) -> float:
        if world == 1:
            return ms
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    for _ in range(max(args.warmup, 1)):
        out = step_resident()
    scal = read_scalars(params, out[3])
    barrier()

    # ---- timed region 1: inputs resident in HBM, device timing -------------------------------------------
    sampler = ClockSampler(local_rank)
    sampler.start()
    native.launch_count_reset()
    native.gemm_profile(True)
    barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(args.steps):
        out = step_resident()
    ev1.record()
    barrier()
    ms_total = max_over_ranks(ev0.elapsed_time(ev1))
    launches = native.launch_count()
    gemm_ms, gemm_flops, gemm_n = native.gemm_profile(False)
    # ---- timed region 2: end to end through the public API with HOST buffers ----------------------------
    barrier()
    ev2, ev3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev2.record()
    for _ in range(args.steps):
        scal = step_e2e()
    ev3.record()
    barrier()
    e2e_ms_total = max_over_ranks(ev2.elapsed_time(ev3))
    sampler.stop_flag.set()
    sampler.join(timeout=2)

    tokens_per_step = gb * seq
    ms_per_step = ms_total / args.steps
    value = tokens_per_step / (ms_per_step * 1e-3)
    e2e_value = tokens_per_step / (e2e_ms_total / args.steps * 1e-3)
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    from obst.analysis import count_params, flops_per_token_fwd
    flops_tok = 3.0 * flops_per_token_fwd(params)
    pk = peaks()
    # dominant kernel: gemm_tc_kernel (every contraction of the step).  achieved = algorithmic FLOPs (3 x forward,
    # causal-aware, recompute NOT credited) that this rank's launches carry / summed launch durations (CUDA events
    # around every launch on the launching stream, inside timed region 1).
    rank_tokens = rows * seq * args.steps
    achieved = flops_tok * rank_tokens / (gemm_ms * 1e-3) / 1e12 if gemm_ms > 0 else None
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "gemm_traffic.json")
    if os.path.exists(tpath):
        with open(tpath) as f:
            traffic = json.load(f).get("dram_bytes_per_launch")
    roofline = {"bound": "tensor", "kernel": "obst::gemm_tc_kernel", "achieved": achieved, "peak": pk["tf_sustained"],
                "unit": "TFLOP/s", "frac": (achieved / pk["tf_sustained"]) if achieved else None, "traffic": traffic,
                "peak_source": f"{pk['source']} (bf16_tflops_sustained: kernel timed inside a long step)",
                "launches": gemm_n, "kernel_ms_per_step": gemm_ms / args.steps,
                "kernel_share_of_step": gemm_ms / ms_total if ms_total else None,
                "executed_tflops": gemm_flops / (gemm_ms * 1e-3) / 1e12 if gemm_ms > 0 else None,
                "whole_step_frac_of_peak": flops_tok * tokens_per_step / world / (ms_per_step * 1e-3) / 1e12 /
                pk["tf_sustained"]}
    line = {"metric": METRIC.replace("32big_mixer", args.workload), "value": value, "unit": UNIT, "n_gpus": world,
            "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "bf16" if cfg["calculation_dtype"] == "bfloat16" else "f32",
            "data": "synthetic",
            "config": {"workload": f"{args.workload}: depth {params.depth}, global batch {gb} x seq {seq} "
                                   f"({tokens_per_step} tokens/step), {count_params(params)} params, random init",
                       "parallelism": f"dp{world}", "micro_batch_rows": mb, "rows_per_rank": rows,
                       "l2": "inputs larger than L2 (every activation tensor is >= 512 MiB; weights 1.4 GB)",
                       "loss": scal["loss"], "accuracy": scal["accuracy"]},
            "clocks": sampler.summary(),
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": 2 * rows * seq * 4 * world,
                    "d2h_bytes_per_step": 24 * world, "ms_per_step": e2e_ms_total / args.steps},
            "gpu_launches": launches, "roofline": roofline}
    if args.depth:
        line["INVALID"] = "depth override"
    if args.store_activations:
        line["config"]["extension"] = "revnet_store_activations (no recompute / reconstruction in backward)"
    if world == 1 and not args.no_cpu_baseline:
        v, cores, sample, _ = cpu_step_rate(args.workload, args.cpu_rows, 1, 0, args.depth)
        line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


class _JsonOnlyStdout:
    """The contract is ONE JSON line on stdout.  Everything else -- the reference-style config warnings, torch
    chatter, and lines native libraries write straight to file descriptor 1 (NCCL prints its version there on some
    boxes) -- goes to stderr: fd 1 is pointed at stderr for the whole run and the JSON line is written to a
    duplicate of the original descriptor."""

    def __init__(self):
        sys.stdout.flush()
        self.fd = os.dup(1)
        os.dup2(2, 1)
        self.json_open = False

    def write(self, text):
        if text.lstrip().startswith("{\""):
            self.json_open = True
            os.write(self.fd, text.encode())
        elif self.json_open and text.strip() == "":
            os.write(self.fd, text.encode())  # the newline that print() sends after the JSON text
            self.json_open = False
        else:
            self.json_open = False
            sys.stderr.write(text)

    def flush(self):
        sys.stderr.flush()


if __name__ == "__main__":
    a = parse()
    sys.stdout = _JsonOnlyStdout()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)
