#!/usr/bin/env python
"""Generates tests/golden/tiny_step.npz from oracle/obst_oracle.py (float64 forward/backward, fp32 optimizer).

The reference holds no golden vectors for this path and cannot be run here (TF 2.7 + mesh-tensorflow), so these
fixtures pin the ORACLE against silent edits, not the oracle against the reference ("parity unpinned").
Run from the repo root:  python tests/golden/make_golden.py"""
import copy
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import obst_oracle as orc  # noqa: E402


def main():
    cfg = copy.deepcopy(orc.TINY_CONFIG)
    cfg.update(depth=2, train_batch_size=2, sequence_length=64)
    sh = orc.Shapes(cfg)
    p = orc.init_params(sh, seed=0)
    tx, ty = orc.synthetic_tokens(sh)
    trace = {}
    loss, acc, state = orc.forward(sh, p, tx, ty, dtype=np.float64, trace=trace)
    grads = orc.backward(sh, state)
    st = orc.init_opt_state(p)
    p2 = {k: v.copy() for k, v in p.items()}
    orc.optimizer_step(sh, p2, {k: v.astype(np.float32) for k, v in grads.items()}, st, orc.learning_rate(sh, 2048))
    out = {"loss": loss, "accuracy": acc, "lr_2048": orc.learning_rate(sh, 2048),
           "logits_head": trace["logits"][:4, :8], "x2_last_head": trace[f"x2_{sh.depth - 1}_1"][0, :2, 0, :8]}
    for k in sorted(grads):
        out[f"gnorm/{k}"] = np.linalg.norm(grads[k])
        out[f"pnew_sum/{k}"] = p2[k].astype(np.float64).sum()
        out[f"mom_norm/{k}"] = np.linalg.norm(st[k]["momentum"])
        out[f"dim0_sum/{k}"] = st[k]["dim0"].astype(np.float64).sum()
    np.savez(os.path.join(os.path.dirname(os.path.abspath(__file__)), "tiny_step.npz"), **out)
    print("wrote tiny_step.npz:", len(out), "entries; loss", loss)


if __name__ == "__main__":
    main()
