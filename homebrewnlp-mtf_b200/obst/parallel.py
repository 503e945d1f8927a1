"""Data parallelism: the reference's `batch:b` mesh axis (src/dataclass.py:247-252) as one process per GPU.

Every rank holds all parameters and optimizer state, takes train_batch_size/world rows of the batch, and the
weight gradients are SUM-allreduced (NCCL over NVLink; gloo on CPU for tests) in depth buckets, in the dtype the
graph hands the optimizer (bf16 when calculation_dtype is bfloat16: what an MTF einsum over the batch mesh axis
reduces in).  During the last
backward pass a depth's bucket is launched on a side stream as soon as that depth's two reversible sub-blocks
have been back-propagated, so the transfer overlaps the rest of backward; the shared token-mixing maps and the
input/output variables (final only at the very end) form the last bucket together with the loss scalars.
The loss normaliser is the GLOBAL token count, so gradients are summed, never averaged."""
import typing

import torch
import torch.distributed as dist


class GradReducer:
    def __init__(self, params, group=None):
        self.params = params
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.last = False
        self.launched: typing.Set[int] = set()
        self.handles: list = []
        self.comm_stream: typing.Optional[torch.cuda.Stream] = None
        self.bytes_reduced = 0

    # -- called by get_optimizer / the reversible backward ----------------------------------------------
    def begin_backward(self, last: bool):
        self.last = last
        self.launched = set()
        self.handles = []

    def _flat(self) -> torch.Tensor:
        """The buffer that is reduced: the gradients in the dtype the graph hands the optimizer (activation dtype,
        src/optimizer/__init__.py:180) -- bf16 for the 32*_mixer configs, which halves the NVLink bytes."""
        state = self.params.optimizer_state
        fn = getattr(state, "reduce_buffer", None)
        return fn() if fn is not None else state.flat_grad

    def _hand_over(self, key: int):
        """Round the bucket's fp32 accumulation into the hand-over buffer (no-op when that IS the fp32 buffer)."""
        fn = getattr(self.params.optimizer_state, "hand_over", None)
        if fn is not None:
            fn(key, with_norms=False)

    def _reduce(self, t: torch.Tensor):
        if self.world == 1 or t.numel() == 0:
            return
        self.bytes_reduced += t.numel() * t.element_size()
        if t.is_cuda:
            if self.comm_stream is None:
                self.comm_stream = torch.cuda.Stream(device=t.device)
            self.comm_stream.wait_stream(torch.cuda.current_stream(t.device))
            with torch.cuda.stream(self.comm_stream):
                dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
        else:
            dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)

    def agree(self, t: torch.Tensor):
        """Make every rank use rank 0's copy of a few reduction results (per-tensor norm sums).  Each rank accumulates
        them with float64 atomics in its own order; a last-bit difference in one rank's clip scale would let that
        replica's weights drift for the rest of the run, and nothing downstream would notice."""
        if self.world == 1 or t.numel() == 0:
            return
        src = dist.get_global_rank(self.group, 0) if self.group is not None else 0
        dist.broadcast(t, src=src, group=self.group)

    def on_depth_done(self, depth: int):
        """Gradients of every variable owned by `depth` are final (both sub-blocks back-propagated)."""
        if not self.last or self.world == 1:
            return
        rng = self.params.optimizer_state.depth_ranges.get(depth)
        if rng is None or depth in self.launched:
            return
        self.launched.add(depth)
        self._hand_over(depth)
        self._reduce(self._flat()[rng[0]:rng[1]])

    def finish(self):
        """Reduce whatever has not been launched yet (always: the shared / input / output bucket and the loss
        scalars), then make the compute stream wait for the side stream."""
        if self.world == 1:
            return
        state = self.params.optimizer_state
        for key, rng in state.depth_ranges.items():
            if key not in self.launched:
                self.launched.add(key)
                self._hand_over(key)
                self._reduce(self._flat()[rng[0]:rng[1]])
        extra = getattr(state, "extra_reduce_buffers", None)
        for buf in (extra() if extra is not None else []):
            self._reduce(buf)
        sums = getattr(self.params, "loss_sums", None)
        if sums is not None:
            self._reduce(sums)
        if self.comm_stream is not None:
            torch.cuda.current_stream().wait_stream(self.comm_stream)


def shard_rows(total_rows: int, rank: int, world: int) -> typing.Tuple[int, int]:
    """[start, stop) of the batch rows owned by `rank`; the batch must divide evenly (as MTF requires of `b`)."""
    if total_rows % world:
        raise ValueError(f"train_batch_size {total_rows} is not divisible by the data-parallel size {world}")
    per = total_rows // world
    return rank * per, (rank + 1) * per
