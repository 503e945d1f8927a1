"""Autoregressive sampling throughput: the reference's loop -- one full forward pass of the
token window per generated position (src/run/inference.py:76-95) -- on the 32big_mixer shape, CUDA events."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "homebrewnlp-mtf_b200"))
from obst import native  # noqa: E402
from obst.configs import mixer_config  # noqa: E402
from obst.dataclass import ModelParameter  # noqa: E402
from obst.run.inference import autoregressive_model  # noqa: E402


def main():
    workload = sys.argv[1] if len(sys.argv) > 1 else "32big_mixer"
    rows = int(sys.argv[2]) if len(sys.argv) > 2 else 256
    steps = int(sys.argv[3]) if len(sys.argv) > 3 else 8
    native.load()
    cfg = mixer_config(workload, train_batch_size=rows)
    params = ModelParameter(cfg)
    S = cfg["sequence_length"]
    tok = torch.randint(0, cfg["vocab_size"], (rows, S, 1), dtype=torch.int32, device="cuda")
    autoregressive_model(params, token_x_input=tok, initial_pos=S - 2, sampling_temperature=0.7, seed=1)  # warm-up
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    native.launch_count_reset()
    e0.record()
    out, _ = autoregressive_model(params, token_x_input=tok, initial_pos=S - steps, sampling_temperature=0.7, seed=1)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    print(json.dumps({"workload": f"{workload}: sampling, {rows} sequences x context {S}, {steps} positions, "
                                  f"temperature 0.7", "ms_per_position": ms / steps,
                      "generated_tokens_per_s": rows * steps / (ms * 1e-3),
                      "window_tokens_per_s_forward": rows * S * steps / (ms * 1e-3),
                      "gpu_launches": int(native.launch_count()),
                      "changed_tokens": int((out[:, S - steps:] != tok[:, S - steps:]).sum())}))


if __name__ == "__main__":
    main()
