"""obst_opt_step_noalloc: the fused optimizer update for a host that cannot allocate device memory.

The regular entry points (obst_opt_norms / obst_opt_update / obst_opt_finalize) need a tensor table, a work list,
norm sums and SM3 staging buffers that the caller puts into device memory.  Extension code that runs inside a
process whose framework has claimed the whole GPU, or under a runtime that forbids driver allocation calls, has no
way to create those; obst_opt_step_noalloc keeps them in a fixed block inside the library image instead.  This
test drives both paths through the C-ABI on the same tensor and checks that they agree, twice in a row (the second
call reuses the block the first one left dirty, and starts from non-zero optimizer state).

Collected last on purpose (file name): a `-x` run has then already exercised everything else."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def nat():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from obst import native
    native.load()
    return native


def _reference_step(nat, p, shadow, g, m, accs, dims, decay, tril_s, state_dtype, hyper):
    """The existing path: caller-allocated table / work list / norms / staging."""
    dev = p.device
    cols = dims[-1]
    rows = p.numel() // cols
    t = nat.OptTensor()
    t.p, t.shadow, t.g, t.m = p.data_ptr(), shadow.data_ptr(), g.data_ptr(), m.data_ptr()
    stage = [torch.zeros(d, dtype=torch.float32, device=dev) for d in dims]
    for i, d in enumerate(dims):
        t.acc[i] = accs[i].data_ptr()
        t.acc_new[i] = stage[i].data_ptr()
        t.dims[i] = d
    t.numel, t.ndim, t.decay, t.tril_s = p.numel(), len(dims), decay, tril_s
    table = torch.frombuffer(bytearray(bytes(t)), dtype=torch.uint8).to(dev)
    rpi = min(2048, max(1, 32768 // cols))
    r0 = list(range(0, rows, rpi))
    it = torch.zeros(len(r0), dtype=torch.int32, device=dev)
    ir0 = torch.tensor(r0, dtype=torch.int32, device=dev)
    irs = torch.tensor([min(rpi, rows - r) for r in r0], dtype=torch.int32, device=dev)
    norms = torch.zeros(2, dtype=torch.float64, device=dev)
    lib, st = nat.load(), nat.stream_ptr()
    sdt = nat.dtype_code(state_dtype)
    clip, beta, nesterov, lr, wd = hyper
    nat.check(lib.obst_opt_norms(table.data_ptr(), 1, it.data_ptr(), ir0.data_ptr(), irs.data_ptr(), len(r0),
                                 nat.BF16, 1, norms.data_ptr(), 0, None, 0, 1, st), "obst_opt_norms")
    nat.check(lib.obst_opt_update(table.data_ptr(), 1, it.data_ptr(), ir0.data_ptr(), irs.data_ptr(), len(r0), cols,
                                  int(cols % 4 == 0), nat.BF16, sdt, 1, norms.data_ptr(), None, clip, beta, nesterov,
                                  lr, wd, 0, st), "obst_opt_update")
    nat.check(lib.obst_opt_finalize(table.data_ptr(), 1, sdt, st), "obst_opt_finalize")
    torch.cuda.synchronize()


def _noalloc_step(nat, p, shadow, g, m, accs, dims, decay, tril_s, state_dtype, hyper):
    lib = nat.load()
    acc_arr = (ctypes.c_void_p * 4)(*[a.data_ptr() for a in accs], *([None] * (4 - len(accs))))
    dim_arr = (ctypes.c_int32 * 4)(*dims, *([0] * (4 - len(dims))))
    clip, beta, nesterov, lr, wd = hyper
    # driver-level view of the device's memory: unlike torch's own counters it also sees a cudaMalloc made inside the
    # library (the first call on a device may still trigger lazy module loading, so callers warm the entry point up
    # once before they measure)
    torch.cuda.synchronize()
    free_before, _ = torch.cuda.mem_get_info()
    rc = lib.obst_opt_step_noalloc(p.data_ptr(), shadow.data_ptr(), g.data_ptr(), m.data_ptr(), acc_arr, dim_arr,
                                   len(dims), decay, tril_s, nat.BF16, nat.dtype_code(state_dtype), 1, clip, beta,
                                   nesterov, lr, wd, nat.stream_ptr())
    nat.check(rc, "obst_opt_step_noalloc")
    torch.cuda.synchronize()
    free_after, _ = torch.cuda.mem_get_info()
    return free_before - free_after


# (2, 12288): a last dim the first version accepted on paper and could not launch (shared memory); wide rows are now
# split into 512-column chunks
@pytest.mark.parametrize("dims,state_dtype,tril_s", [((8, 96, 128), torch.bfloat16, 0), ((4, 64, 64), torch.float32, 64),
                                                     ((256, 6), torch.bfloat16, 0), ((2, 12288), torch.bfloat16, 0)])
def test_noalloc_update_matches_the_table_driven_update(nat, dims, state_dtype, tril_s):
    g0 = torch.Generator().manual_seed(7)
    dev = "cuda"
    hyper = (0.003, 0.9, 1, 0.005, 1e-4)

    def fresh():
        gen = torch.Generator().manual_seed(11)
        p = (torch.randn(dims, generator=gen) * 0.05).to(dev)
        shadow = torch.zeros(dims, dtype=torch.bfloat16, device=dev)
        m = torch.zeros(dims, dtype=state_dtype, device=dev)
        accs = [torch.zeros(d, dtype=state_dtype, device=dev) for d in dims]
        return p, shadow, m, accs

    grads = [(torch.randn(dims, generator=g0) * 1e-3).to(dev) for _ in range(3)]
    a, b = fresh(), fresh()
    # the oracle's op-by-op statement of the same chain, on the same tensor ("win_0": a weight that decays)
    import types

    import numpy as np

    from oracle import obst_oracle as orc
    sh = types.SimpleNamespace(calculation_dtype="bfloat16",
                               optimizer_slice_dtype="bfloat16" if state_dtype == torch.bfloat16 else "float32",
                               optimizer="adaptive_clip:0.003-sm3-momentum:0.9:1:1-learning_rate", weight_decay=hyper[4])
    op = {"win_0": b[0].cpu().numpy().copy()}
    ost = orc.init_opt_state(op)
    warm = fresh()  # first call on this device: symbol look-up and lazy module loading happen here
    _noalloc_step(nat, warm[0], warm[1], grads[2].clone(), warm[2], warm[3], list(dims), 1, tril_s, state_dtype, hyper)
    for step in range(2):  # two calls in a row: the second starts from non-zero state and a used scratch block
        _reference_step(nat, a[0], a[1], grads[step].clone(), a[2], a[3], list(dims), 1, tril_s, state_dtype, hyper)
        grew = _noalloc_step(nat, b[0], b[1], grads[step].clone(), b[2], b[3], list(dims), 1, tril_s, state_dtype,
                             hyper)
        assert grew <= 0, f"device memory in use grew by {grew} bytes across obst_opt_step_noalloc"
        # same kernels, same arithmetic; only the order of the fp64 norm atomics may differ -> 1e-6 relative
        for x, y in ((a[0], b[0]), (a[1], b[1]), (a[2], b[2]), *zip(a[3], b[3])):
            x, y = x.double(), y.double()
            assert float((x - y).norm() / (x.norm() + 1e-30)) < 1e-6
        assert float(b[2].float().abs().max()) > 0
        # ... and against the oracle: master weight, momentum and accumulators bit for bit
        orc.optimizer_step(sh, op, {"win_0": orc.round_bf16(grads[step].cpu().numpy())}, ost, hyper[3])
        assert np.array_equal(b[0].cpu().numpy(), op["win_0"])
        assert np.array_equal(b[2].float().cpu().numpy(), ost["win_0"]["momentum"])
        for i, acc in enumerate(b[3]):
            assert np.array_equal(acc.float().cpu().numpy(), ost["win_0"][f"dim{i}"])


def test_noalloc_rejects_what_does_not_fit(nat):
    lib = nat.load()
    dev = "cuda"
    p = torch.zeros(8, device=dev)
    acc_arr = (ctypes.c_void_p * 4)(p.data_ptr(), p.data_ptr(), None, None)
    too_wide = (ctypes.c_int32 * 4)(2, 20000, 0, 0)
    rc = lib.obst_opt_step_noalloc(p.data_ptr(), None, p.data_ptr(), p.data_ptr(), acc_arr, too_wide, 2, 0, 0, nat.BF16,
                                   nat.F32, 0, 0.003, 0.9, 1, 0.0, 0.0, nat.stream_ptr())
    assert rc == -1 and b"staging holds" in lib.obst_last_error()
    # staging is large enough (4096 + 4096 + 64 entries), the work list is not: 16.7 M rows of 8 = 8192 items
    too_many = (ctypes.c_int32 * 4)(4096, 4096, 8, 0)
    acc2 = (ctypes.c_void_p * 4)(p.data_ptr(), p.data_ptr(), p.data_ptr(), None)
    rc = lib.obst_opt_step_noalloc(p.data_ptr(), None, p.data_ptr(), p.data_ptr(), acc2, too_many, 3, 0, 0, nat.BF16,
                                   nat.F32, 0, 0.003, 0.9, 1, 0.0, 0.0, nat.stream_ptr())
    assert rc == -1 and b"work items" in lib.obst_last_error()
    odd = torch.zeros(9, device=dev)[1:]  # 4-byte aligned only
    rc = lib.obst_opt_step_noalloc(odd.data_ptr(), None, p.data_ptr(), p.data_ptr(), acc2, (ctypes.c_int32 * 4)(2, 4),
                                   2, 0, 0, nat.BF16, nat.F32, 0, 0.003, 0.9, 1, 0.0, 0.0, nat.stream_ptr())
    assert rc == -1 and b"16-byte aligned" in lib.obst_last_error()
    bad_rank = (ctypes.c_int32 * 4)(1, 1, 1, 1)
    rc = lib.obst_opt_step_noalloc(p.data_ptr(), None, p.data_ptr(), p.data_ptr(), acc2, bad_rank, 5, 0, 0, nat.BF16,
                                   nat.F32, 0, 0.003, 0.9, 1, 0.0, 0.0, nat.stream_ptr())
    assert rc == -1 and b"ndim" in lib.obst_last_error()
