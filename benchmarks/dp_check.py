"""Run under torchrun (N ranks, NCCL): one data-parallel step of the tiny fp32 config must give every rank the
same parameters as a single-process step on the full batch (gradients are SUMMED with the global loss normaliser,
adaptive-clip norms are taken after the allreduce)."""
import copy
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "homebrewnlp-mtf_b200"))
from oracle import obst_oracle as orc  # noqa: E402  (checker only: seeded weights + tokens)
from tests.helpers import make_params, to_np  # noqa: E402
from obst.parallel import GradReducer, shard_rows  # noqa: E402
from obst.run.train import get_train_model, read_scalars  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    for dtype, tol in (("float32", 1e-5), ("bfloat16", 2e-2)):
        cfg = copy.deepcopy(orc.TINY_CONFIG)
        cfg.update(depth=2, train_batch_size=8, sequence_length=128, features_per_head=64, calculation_dtype=dtype,
                   optimizer_slice_dtype=dtype, device=f"cuda:{local}")
        sh = orc.Shapes(cfg)
        op = orc.init_params(sh, seed=0)
        tx, ty = orc.synthetic_tokens(sh)
        r0, r1 = shard_rows(8, rank, world)
        # ---- data parallel
        params = make_params(cfg, op, micro_batch_size=2)
        params.grad_reducer = GradReducer(params)
        params.global_step = 2048
        out = get_train_model(params, None, None, None, torch.from_numpy(tx[r0:r1]).cuda(),
                              torch.from_numpy(ty[r0:r1]).cuda(), None, None, None, 0)
        dp_scal = read_scalars(params, out[3])
        dp = [to_np(v.data) for v in params.variable_cache.values()]
        # ---- single process, full batch
        single = make_params(cfg, op)
        single.global_step = 2048
        out = get_train_model(single, None, None, None, torch.from_numpy(tx).cuda(), torch.from_numpy(ty).cuda(), None,
                              None, None, 0)
        sp_scal = read_scalars(single, out[3])
        sp = [to_np(v.data) for v in single.variable_cache.values()]
        worst = max(float(np.linalg.norm(a - b) / (np.linalg.norm(b) + 1e-30)) for a, b in zip(dp, sp))
        assert abs(dp_scal["loss"] - sp_scal["loss"]) / sp_scal["loss"] < tol, (dp_scal, sp_scal)
        assert worst < (1e-5 if dtype == "float32" else 5e-3), worst
        # every rank holds identical parameters
        flat = torch.cat([v.data.reshape(-1) for v in params.variable_cache.values()])
        ref = flat.clone()
        dist.broadcast(ref, 0)
        assert torch.equal(flat, ref)
        if rank == 0:
            print(f"dp_check[{dtype}] world={world}: loss dp {dp_scal['loss']:.6f} single {sp_scal['loss']:.6f}, "
                  f"worst param rel diff {worst:.2e}, allreduced {params.grad_reducer.bytes_reduced} B -- OK",
                  flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
