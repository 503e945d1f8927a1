"""The N>1 host logic on CPU: world_size 2, gloo.  Each rank holds a flat fp32 gradient buffer laid out like the
optimizer's (depth buckets first, shared/input/output bucket last); GradReducer must SUM every bucket exactly
once whether buckets were launched early (on_depth_done during the last backward) or by finish()."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


class _State:
    def __init__(self, rank):
        g = torch.Generator().manual_seed(100 + rank)
        self.flat_grad = torch.randn(1000, generator=g)
        self.depth_ranges = {2: [0, 300], 1: [300, 600], 0: [600, 900], -1: [900, 1000]}


class _Params:
    pass


def _worker(rank, world, port, early, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, os.path.join(root, "homebrewnlp-mtf_b200"))
    from obst.parallel import GradReducer
    params = _Params()
    params.optimizer_state = _State(rank)
    params.loss_sums = torch.tensor([1.0 + rank, 2.0, 3.0], dtype=torch.float64)
    red = GradReducer(params)
    assert red.world == world and red.rank == rank
    # a non-final micro-batch: nothing may be reduced
    red.begin_backward(last=False)
    red.on_depth_done(2)
    before = params.optimizer_state.flat_grad.clone()
    red.begin_backward(last=True)
    for d in early:
        red.on_depth_done(d)
    red.finish()
    ok = torch.equal(before, _State(rank).flat_grad)
    norms = torch.tensor([1.0, 2.0 + rank * 2.0 ** -50], dtype=torch.float64)  # last-bit disagreement between ranks
    red.agree(norms)
    torch.save({"grad": params.optimizer_state.flat_grad, "sums": params.loss_sums, "ok": ok, "norms": norms,
                "bytes": red.bytes_reduced}, os.path.join(out, f"r{rank}.pt"))
    dist.destroy_process_group()


@pytest.mark.parametrize("early", [(), (2, 1, 0), (2,)])
def test_grad_reducer_world2_gloo(tmp_path, early):
    port = _free_port()
    mp.spawn(_worker, args=(2, port, early, str(tmp_path)), nprocs=2, join=True)
    r0 = torch.load(tmp_path / "r0.pt")
    r1 = torch.load(tmp_path / "r1.pt")
    expect = _State(0).flat_grad + _State(1).flat_grad
    assert r0["ok"] and r1["ok"]
    assert torch.allclose(r0["grad"], expect) and torch.equal(r0["grad"], r1["grad"])  # summed once, identical
    assert r0["sums"].tolist() == [3.0, 4.0, 6.0]
    assert r0["norms"].tolist() == [1.0, 2.0] and torch.equal(r0["norms"], r1["norms"])  # rank 0's copy everywhere
    assert r0["bytes"] == 1000 * 4 + 3 * 8
