"""Per-variable comparison of the product's gradients with the reference-run goldens (development helper)."""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "homebrewnlp-mtf_b200"))
from tests.test_reference_models_gpu import GOLDEN, _nested, rel  # noqa: E402


def main():
    from obst.dataclass import ModelParameter, gpt_mode
    from obst.run.train import get_train_model, read_scalars
    for case in sys.argv[1:]:
        with np.load(GOLDEN) as f:
            rec = _nested(f, case)
        cfg = json.loads(str(rec["config"]))
        created = [str(n) for n in rec["created"]]
        params = ModelParameter(gpt_mode(dict(cfg, gemm_backend=1)))
        params.import_variables = [rec["init"][n] for n in created]
        params.global_step = 1500
        x = torch.from_numpy(rec["tok_x"]).to(params.device)
        y = torch.from_numpy(rec["tok_y"]).to(params.device)
        step = rec["steps"]["0"]
        twin = None
        if "bf16" in case:
            with np.load(GOLDEN) as f:
                twin = _nested(f, case.replace("bf16", "f32") + "_twin")["steps"]["0"]["grads"]
        out = get_train_model(params, None, None, None, x, y, None, None, None, params.manual_global_step)
        print(case, "loss", read_scalars(params, out[3])["loss"], float(step["loss"]))
        for n in created:
            var = params.variable_cache[n]
            g = step["grads"].get(n)
            mine = None if var.grad is None else var.grad.detach().float().cpu().numpy()
            if g is None or mine is None:
                print(f"  {n}: ref {None if g is None else np.linalg.norm(g):} mine {None if mine is None else np.linalg.norm(mine)}")
                continue
            extra = ""
            if twin is not None:
                t = twin[n]
                extra = f" | f32 truth {np.linalg.norm(t):.3e}: ref-bf16 off {rel(g, t):.2e}, mine off {rel(mine.reshape(t.shape), t):.2e}"
            print(f"  {n.replace('gpt0v/body0v/', '')[-70:]:70s} |ref| {np.linalg.norm(g):.3e} |mine| {np.linalg.norm(mine):.3e} rel {rel(mine.reshape(g.shape), g):.2e}{extra}")


if __name__ == "__main__":
    main()
