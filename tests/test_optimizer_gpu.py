"""The fused optimizer update (obst_opt_norms / obst_grad_finalize / obst_opt_update / obst_opt_finalize) against the
oracle's op-by-op fp32 statement of the reference chain
    adaptive_clip:0.003 - sm3 - momentum:0.9:1:1 - learning_rate  (+ weight decay, assign_sub)
src/optimizer/optimizers.py:60-84,114-128, src/optimizer/__init__.py:31-66 -- at kernel level, through the C-ABI, in
every dtype combination the model uses: the benchmark's (bf16 slots, bf16 hand-over gradients, bf16 shadows, carried
weight norms from step 2 on) and the fp32 one of the tiny config.

Bar: the kernels evaluate the same fp32 op sequence with correctly rounded operations, so momentum, SM3 accumulators
and master weights must be BIT-IDENTICAL to the oracle (torch.equal on the raw values), shadows must be the bf16
rounding of the new master with the causal mask applied, over three consecutive steps."""
import ctypes
import os
import types

import numpy as np
import pytest
import torch

from oracle import obst_oracle as orc

pytestmark = pytest.mark.gpu

CHAIN = "adaptive_clip:0.003-sm3-momentum:0.9:1:1-learning_rate"
DEV = "cuda"


@pytest.fixture(scope="module")
def nat():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from obst import native
    native.load()
    return native


class Bank:
    """A table of tensors laid out like FusedOptimizer does it, driven through the raw entry points."""

    def __init__(self, nat, shapes, state_dtype, act_bf16, g_store_bf16, tril, decay, item_elems=32768):
        self.nat = nat
        self.shapes = shapes
        self.state_dtype = state_dtype
        self.act_bf16 = act_bf16
        self.g_store_bf16 = g_store_bf16
        gen = torch.Generator().manual_seed(5)
        self.p = [(torch.randn(s, generator=gen) * 0.05).to(DEV) for s in shapes]
        adt = torch.bfloat16 if act_bf16 else torch.float32
        self.shadow = [torch.zeros(s, dtype=adt, device=DEV) for s in shapes]
        self.m = [torch.zeros(s, dtype=state_dtype, device=DEV) for s in shapes]
        self.acc = [[torch.zeros(d, dtype=state_dtype, device=DEV) for d in s] for s in shapes]
        self.acc_new = [[torch.zeros(d, dtype=torch.float32, device=DEV) for d in s] for s in shapes]
        self.g_acc = [torch.zeros(s, dtype=torch.float32, device=DEV) for s in shapes]
        self.g_lp = [torch.zeros(s, dtype=torch.bfloat16, device=DEV) for s in shapes] if g_store_bf16 else None
        n = len(shapes)
        table = (nat.OptTensor * n)()
        items = []
        self.max_cols, self.vec4 = 1, True
        for ti, s in enumerate(shapes):
            t = table[ti]
            t.p, t.shadow, t.m = self.p[ti].data_ptr(), self.shadow[ti].data_ptr(), self.m[ti].data_ptr()
            if g_store_bf16:
                t.g, t.g_acc, t.g_bf16 = self.g_lp[ti].data_ptr(), self.g_acc[ti].data_ptr(), 1
            else:
                t.g, t.g_acc, t.g_bf16 = self.g_acc[ti].data_ptr(), None, 0
            for i, d in enumerate(s):
                t.acc[i], t.acc_new[i], t.dims[i] = self.acc[ti][i].data_ptr(), self.acc_new[ti][i].data_ptr(), d
            t.numel, t.ndim, t.decay, t.tril_s = int(np.prod(s)), len(s), int(decay[ti]), int(tril[ti])
            cols = s[-1]
            rows = t.numel // cols
            self.max_cols = max(self.max_cols, cols)
            self.vec4 &= cols % 4 == 0
            rpi = min(2048, max(1, item_elems // min(cols, 512)))
            for r0 in range(0, rows, rpi):
                items.append((ti, r0, min(rpi, rows - r0)))
        self.table = torch.frombuffer(bytearray(bytes(table)), dtype=torch.uint8).to(DEV)
        self.n = n
        self.n_items = len(items)
        self.it = torch.tensor([i[0] for i in items], dtype=torch.int32, device=DEV)
        self.ir0 = torch.tensor([i[1] for i in items], dtype=torch.int32, device=DEV)
        self.irs = torch.tensor([i[2] for i in items], dtype=torch.int32, device=DEV)
        self.norms = torch.zeros(2 * n, dtype=torch.float64, device=DEV)
        self.wnext = torch.zeros(n, dtype=torch.float64, device=DEV)
        self.carry = False

    def step(self, grads, hyper):
        nat, lib, st = self.nat, self.nat.load(), self.nat.stream_ptr()
        clip, beta, nesterov, lr, wd = hyper
        for dst, g in zip(self.g_acc, grads):
            dst.copy_(g)
        act = nat.BF16 if self.act_bf16 else nat.F32
        sdt = nat.dtype_code(self.state_dtype)
        args = (self.table.data_ptr(), self.n, self.it.data_ptr(), self.ir0.data_ptr(), self.irs.data_ptr(), self.n_items)
        self.norms.zero_()
        flags = int(self.carry)
        if self.g_store_bf16:  # hand-over + sum g^2 in one pass, as on a single rank
            nat.check(lib.obst_grad_finalize(*args, self.norms.data_ptr(), 1, st), "obst_grad_finalize")
            flags |= 6
            grad_bf16 = 2
        else:
            grad_bf16 = int(self.act_bf16)
        nat.check(lib.obst_opt_norms(*args, act, grad_bf16 & 1, self.norms.data_ptr(), flags, self.wnext.data_ptr(), 0,
                                     self.n, st), "obst_opt_norms")
        nat.check(lib.obst_opt_update(*args, self.max_cols, int(self.vec4), act, sdt, grad_bf16, self.norms.data_ptr(),
                                      self.wnext.data_ptr(), clip, beta, nesterov, lr, wd, 0, st), "obst_opt_update")
        nat.check(lib.obst_opt_finalize(self.table.data_ptr(), self.n, sdt, st), "obst_opt_finalize")
        torch.cuda.synchronize()
        self.carry = clip > 0


def oracle_shapes(act_bf16, state_bf16, wd):
    return types.SimpleNamespace(calculation_dtype="bfloat16" if act_bf16 else "float32",
                                 optimizer_slice_dtype="bfloat16" if state_bf16 else "float32", optimizer=CHAIN,
                                 weight_decay=wd)


# 32big_mixer's tensor kinds at reduced sizes: W_in-like, W_mid-like (1024-wide rows: two column chunks), a causal
# mixing map, norm scale (one short item), the rank-4 input linear, a 256-wide LM head, a last dim of 6 (generic kernel)
SHAPES = [(8, 96, 128), (64, 8, 1024), (4, 64, 64), (8, 512), (1, 16, 8, 512), (8, 64, 1, 256), (40, 6), (2, 20000)]
TRIL = [0, 0, 64, 0, 0, 0, 0, 0]
DECAY = [1, 1, 0, 0, 0, 0, 1, 0]
NAMES = ["win_0", "wmid_0", "mix_a", "g0_0", "w0", "wo", "win_1", "embed"]  # oracle decay rule: win_/wmid_/wout_


@pytest.mark.parametrize("act_bf16,state_dtype,g_store_bf16,odd", [
    (True, torch.bfloat16, True, False),     # the benchmark's combination (configs/32big_mixer.json:60-63)
    (True, torch.bfloat16, False, False),    # bf16 activations, fp32 gradient buffer rounded on read
    (True, torch.float32, True, False),
    (False, torch.float32, False, False),    # the tiny fp32 config
    # a last dim that is not a multiple of 4 sends the whole launch to the generic kernel
    (True, torch.bfloat16, True, True),
    (True, torch.bfloat16, False, True),
    (False, torch.float32, False, True),
])
def test_fused_update_is_bit_identical_to_the_op_by_op_chain(nat, act_bf16, state_dtype, g_store_bf16, odd):
    shapes = SHAPES if odd else [s for s in SHAPES if s[-1] % 4 == 0]
    keep = [SHAPES.index(s) for s in shapes]
    tril, decay, names = [TRIL[i] for i in keep], [DECAY[i] for i in keep], [NAMES[i] for i in keep]
    bank = Bank(nat, shapes, state_dtype, act_bf16, g_store_bf16, tril, decay)
    sh = oracle_shapes(act_bf16, state_dtype == torch.bfloat16, 1e-4)
    op = {n: p.cpu().numpy().copy() for n, p in zip(names, bank.p)}
    ost = orc.init_opt_state(op)
    gen = torch.Generator().manual_seed(9)
    hyper_of = lambda step: (0.003, 0.9, 1, 0.005 + 0.001 * step, 1e-4)  # noqa: E731
    for step in range(3):  # step 0: norms of the weights are read; steps 1-2: carried from the previous update
        grads = [torch.randn(s, generator=gen) * (1e-3 if i % 2 else 3e-2) for i, s in enumerate(shapes)]
        bank.step([g.to(DEV) for g in grads], hyper_of(step))
        handed = {n: (orc.round_bf16(g.numpy()) if act_bf16 else g.numpy().astype(np.float32))
                  for n, g in zip(names, grads)}
        orc.optimizer_step(sh, op, handed, ost, hyper_of(step)[3])
        for ti, n in enumerate(names):
            ref_p = torch.from_numpy(op[n])
            assert torch.equal(bank.p[ti].cpu(), ref_p), (step, n, float((bank.p[ti].cpu() - ref_p).abs().max()))
            ref_m = torch.from_numpy(ost[n]["momentum"]).to(state_dtype)
            assert torch.equal(bank.m[ti].cpu(), ref_m), (step, n, "momentum")
            for i in range(len(shapes[ti])):
                ref_a = torch.from_numpy(ost[n][f"dim{i}"]).to(state_dtype)
                assert torch.equal(bank.acc[ti][i].cpu(), ref_a), (step, n, f"dim{i}")
                assert float(bank.acc_new[ti][i].abs().max()) == 0.0  # staging left clean for the next step
            seen = torch.from_numpy(orc.round_bf16(op[n]) if act_bf16 else op[n])
            if tril[ti]:
                seen = torch.tril(seen)
            assert torch.equal(bank.shadow[ti].float().cpu(), seen), (step, n, "shadow")
        if g_store_bf16:  # the hand-over buffer holds the rounded gradient
            for ti, n in enumerate(names):
                assert torch.equal(bank.g_lp[ti].float().cpu(), torch.from_numpy(handed[n])), (step, n)


def test_macro_batching_mode_moves_the_state_but_not_the_master(nat):
    """preview bit 0 (the reference's non-final macro-batches, src/run/train.py:63-69): momentum and accumulators
    are updated exactly as by a committed update, the master stays, and the shadow receives
    cast(cast(p) - cast(update)) -- the value the next macro-batch's graph reads."""
    shapes = [(8, 96, 128), (4, 64, 64)]
    a = Bank(nat, shapes, torch.bfloat16, True, True, [0, 64], [1, 0])
    b = Bank(nat, shapes, torch.bfloat16, True, True, [0, 64], [1, 0])
    gen = torch.Generator().manual_seed(3)
    grads = [(torch.randn(s, generator=gen) * 1e-2).to(DEV) for s in shapes]
    hyper = (0.003, 0.9, 1, 0.01, 1e-4)
    p0 = [t.clone() for t in b.p]
    b.step(grads, hyper)  # the committed update: reference for the state and for the size of the update
    lib, st = nat.load(), nat.stream_ptr()
    for dst, g in zip(a.g_acc, grads):
        dst.copy_(g)
    args = (a.table.data_ptr(), a.n, a.it.data_ptr(), a.ir0.data_ptr(), a.irs.data_ptr(), a.n_items)
    nat.check(lib.obst_grad_finalize(*args, a.norms.data_ptr(), 1, st), "obst_grad_finalize")
    nat.check(lib.obst_opt_norms(*args, nat.BF16, 0, a.norms.data_ptr(), 6, a.wnext.data_ptr(), 0, a.n, st), "norms")
    nat.check(lib.obst_opt_update(*args, a.max_cols, 1, nat.BF16, nat.BF16, 2, a.norms.data_ptr(), None, *hyper, 1, st),
              "obst_opt_update")
    nat.check(lib.obst_opt_finalize(a.table.data_ptr(), a.n, nat.BF16, st), "obst_opt_finalize")
    torch.cuda.synchronize()
    for i in range(len(shapes)):
        assert torch.equal(a.p[i], p0[i])                       # master untouched
        assert torch.equal(a.m[i], b.m[i])                      # state as after a committed update
        assert all(torch.equal(x, y) for x, y in zip(a.acc[i], b.acc[i]))
        step = (p0[i] - b.p[i])                                 # the update the committed path applied (fp32)
        want = (p0[i].bfloat16().float() - step.bfloat16().float()).bfloat16()
        if i == 1:
            want = torch.tril(want)
        # p0 - b.p re-derives the update with one extra rounding: allow the last bf16 bit
        assert float((a.shadow[i].float() - want.float()).abs().max()) <= float(want.float().abs().max()) * 2 ** -7


def test_transpose_batched(nat):
    shapes = [(3, 64, 96), (1, 33, 7), (8, 512, 128), (2, 1, 65)]
    src = [torch.randn(s).to(torch.bfloat16).to(DEV) for s in shapes]
    dst = [torch.full((s[0], s[2], s[1]), 7.0, dtype=torch.bfloat16, device=DEV) for s in shapes]
    items = (nat.TransposeItem * len(shapes))()
    tiles = 0
    for it, s, a, b in zip(items, shapes, src, dst):
        it.src, it.dst, it.batch, it.rows, it.cols, it.tile0 = a.data_ptr(), b.data_ptr(), s[0], s[1], s[2], tiles
        tiles += s[0] * (-(-s[1] // 64)) * (-(-s[2] // 64))
    table = torch.frombuffer(bytearray(bytes(items)), dtype=torch.uint8).to(DEV)
    nat.check(nat.load().obst_transpose_batched(nat.BF16, table.data_ptr(), len(shapes), tiles, nat.stream_ptr()),
              "obst_transpose_batched")
    for a, b in zip(src, dst):
        assert torch.equal(b, a.transpose(1, 2).contiguous())


@pytest.mark.parametrize("case,act_bf16,state_dtype,g_store_bf16", [
    ("hot_path", False, torch.float32, False),
    ("hot_path_bf16", True, torch.bfloat16, True),
    ("hot_path_bf16", True, torch.bfloat16, False),
])
def test_fused_update_against_the_reference_run(nat, case, act_bf16, state_dtype, g_store_bf16):
    """The CUDA update against what the REFERENCE'S OWN get_optimizer produced (tests/golden/reference_optimizer.npz:
    /root/reference's src/optimizer run on NumPy stand-ins, see tests/golden/make_reference_vectors.py), three
    consecutive steps on the recorded gradients: master weights to 2e-6, momentum and SM3 accumulators BIT FOR BIT in
    the state dtype (fp32 slots: 1e-6, the two sides sum the clip norms in different orders)."""
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reference_optimizer.npz")
    with np.load(path) as f:
        keys = [k for k in f.files if k.startswith(case + "|")]
        rec = {k[len(case) + 1:]: f[k] for k in keys}
    names = [str(n) for n in rec["names"]]
    use = [i for i, n in enumerate(names) if rec[f"init|{n}"].ndim >= 1]          # the rank-0 variable takes adam
    shapes = [tuple(rec[f"init|{names[i]}"].shape) for i in use]
    decay = [1 if "bottleneck" in names[i] else 0 for i in use]                     # src/optimizer/__init__.py:49-57
    bank = Bank(nat, shapes, state_dtype, act_bf16, g_store_bf16, [0] * len(use), decay)
    for p, i in zip(bank.p, use):
        p.copy_(torch.from_numpy(rec[f"init|{names[i]}"]))
    opt_key = "adaptive_clip_0.003-sm3-momentum_0.9_1_1-learning_rate"
    for step in range(3):
        grads = [torch.from_numpy(rec[f"steps|{step}|grads|0|{i}"]).to(DEV) for i in use]
        if act_bf16:
            grads = [g.to(torch.bfloat16).float() for g in grads]
        bank.step(grads, (0.003, 0.9, 1, float(rec[f"steps|{step}|lr"]), 1e-4))
        for k, i in enumerate(use):
            n = names[i]
            want = torch.from_numpy(rec[f"steps|{step}|state|{n}"]).to(DEV)
            err = float((bank.p[k] - want).norm() / want.norm())
            assert err < 2e-6, (step, n, err)
            m_want = torch.from_numpy(rec[f"steps|{step}|state|momentum{i}v/{n}/{opt_key}/momentum"]).to(DEV)
            if state_dtype == torch.bfloat16:
                assert torch.equal(bank.m[k], m_want.to(torch.bfloat16)), (step, n, "momentum")
            else:
                assert float((bank.m[k] - m_want).abs().max() / m_want.abs().max()) < 1e-6, (step, n)
            for d in range(len(shapes[k])):
                a_want = torch.from_numpy(rec[f"steps|{step}|state|sm3{i}v/{n}/{opt_key}/dim{d}"]).to(DEV)
                if state_dtype == torch.bfloat16:
                    assert torch.equal(bank.acc[k][d], a_want.to(torch.bfloat16)), (step, n, d)
                else:
                    assert float((bank.acc[k][d] - a_want).abs().max() / a_want.abs().max()) < 1e-6, (step, n, d)
