"""ctypes binding of libobst_b200.so (the C-ABI declared in include/obst_b200.h).

PyTorch is only the plumbing here: it owns device memory and the CUDA stream; every kernel on the
hot path is one of the entry points below.  There is no CPU fallback: if the shared library is
missing or a call fails, this module raises.
"""
import ctypes
import os
import typing

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(os.path.dirname(_HERE), "libobst_b200.so")

F32, BF16 = 0, 1

c_void_p = ctypes.c_void_p
c_int = ctypes.c_int
c_i32 = ctypes.c_int32
c_i64 = ctypes.c_int64
c_float = ctypes.c_float


class Operand(ctypes.Structure):
    _fields_ = [("ptr", c_void_p), ("s_mn", c_i64), ("s_k", c_i64), ("s_ko", c_i64), ("s_glo", c_i64),
                ("s_ghi", c_i64)]


class GemmDesc(ctypes.Structure):
    _fields_ = [("dtype_in", c_i32), ("dtype_out", c_i32), ("M", c_i32), ("N", c_i32), ("KI", c_i32), ("KO", c_i32),
                ("GLO", c_i32), ("GHI", c_i32), ("A", Operand), ("B", Operand), ("D", c_void_p), ("R", c_void_p),
                ("d_ld", c_i64), ("d_glo", c_i64), ("d_ghi", c_i64), ("tri", c_i32), ("epi", c_i32),
                ("accumulate", c_i32), ("splits", c_i32), ("aux0", c_void_p), ("aux1", c_void_p), ("aux2", c_void_p),
                ("aux3", c_void_p), ("aux4", c_void_p), ("scalar0", c_float), ("scalar1", c_float)]


class OptTensor(ctypes.Structure):
    _fields_ = [("p", c_void_p), ("shadow", c_void_p), ("g", c_void_p), ("m", c_void_p), ("acc", c_void_p * 4),
                ("acc_new", c_void_p * 4), ("g_acc", c_void_p), ("numel", c_i64), ("dims", c_i32 * 4), ("ndim", c_i32),
                ("decay", c_i32), ("tril_s", c_i32), ("g_bf16", c_i32)]


class FoldItem(ctypes.Structure):  # obst_fold_item
    _fields_ = [("w", c_void_p), ("gamma", c_void_p), ("beta", c_void_p), ("wg_t", c_void_p), ("cvec", c_void_p),
                ("dvec", c_void_p), ("gamma_f", c_void_p), ("wg", c_void_p)]


class TransposeItem(ctypes.Structure):  # obst_transpose_item
    _fields_ = [("src", c_void_p), ("dst", c_void_p), ("batch", c_i32), ("rows", c_i32), ("cols", c_i32),
                ("tile0", c_i32)]


class OptStageArgs(ctypes.Structure):  # obst_opt_stage_args
    _fields_ = [("stage", c_i32), ("dtype_act", c_i32), ("dtype_state", c_i32), ("flags", c_i32),
                ("f", c_float * 4), ("grad_base", c_void_p), ("work", c_void_p), ("work2", c_void_p),
                ("state_a", c_void_p), ("state_b", c_void_p), ("scalars_in", c_void_p), ("scalars_out", c_void_p),
                ("sums", c_void_p), ("sums2", c_void_p)]


# the OBST_STAGE_* enum of include/obst_b200.h
(STAGE_REDUCE, STAGE_ROUND_BF16, STAGE_COPY, STAGE_ADAPTIVE_CLIP, STAGE_L2NORM_CLIP, STAGE_GLOBAL_L2NORM_CLIP,
 STAGE_VALUE_CLIP, STAGE_GRAD_CENTRALISE, STAGE_WEIGHT_CENTRALISE, STAGE_LEARNING_RATE, STAGE_MOMENTUM, STAGE_ADAM,
 STAGE_NOVOGRAD, STAGE_GRAFT_COMBINE, STAGE_APPLY) = range(15)

# name -> (restype, argtypes); must list every symbol include/obst_b200.h declares (tests check this)
SIGNATURES: typing.Dict[str, typing.Tuple[typing.Any, typing.List[typing.Any]]] = {
    "obst_version": (c_int, []),
    "obst_crc32c": (ctypes.c_uint32, [c_void_p, c_i64, ctypes.c_uint32]),
    "obst_crc32c_portable": (ctypes.c_uint32, [c_void_p, c_i64, ctypes.c_uint32]),
    "obst_last_error": (ctypes.c_char_p, []),
    "obst_launch_count": (c_i64, []),
    "obst_launch_count_reset": (None, []),
    "obst_path_count": (c_i64, [c_i32]),
    "obst_gemm": (c_int, [ctypes.POINTER(GemmDesc), c_int, c_void_p]),
    "obst_gn_fwd": (c_int, [c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_i64, c_i32, c_i32, c_i32,
                            c_void_p]),
    "obst_gn_bwd": (c_int, [c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                            c_i64, c_i32, c_i32, c_i32, c_i32, c_i32, c_void_p]),
    "obst_norm_stats_finalize": (c_int, [c_void_p, c_void_p, c_i64, c_i32, c_i32, c_i32, c_void_p]),
    "obst_fold_norm_weight": (c_int, [c_void_p, c_i32, c_i32, c_i32, c_i32, c_void_p]),
    "obst_rowscale_colsum": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_i64, c_i32, c_i32, c_void_p]),
    "obst_norm_wgrad_rank1": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_i32, c_i32, c_i32, c_void_p]),
    "obst_norm_bwd_coeffs": (c_int, [c_void_p, c_void_p, c_void_p, c_i64, c_i32, c_i32, c_i32, c_void_p]),
    "obst_norm_affine_grad": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_i32, c_i32, c_i32, c_void_p]),
    "obst_act_fwd": (c_int, [c_int, c_i32, c_void_p, c_void_p, c_i64, c_void_p]),
    "obst_act_bwd": (c_int, [c_int, c_i32, c_void_p, c_void_p, c_void_p, c_i64, c_void_p]),
    "obst_gate_fwd": (c_int, [c_int, c_void_p, c_void_p, c_void_p, c_i64, c_void_p]),
    "obst_gate_bwd": (c_int, [c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_i64, c_void_p]),
    "obst_axpby": (c_int, [c_int, c_void_p, c_void_p, c_float, c_void_p, c_i64, c_void_p]),
    "obst_sum_heads": (c_int, [c_int, c_void_p, c_void_p, c_i64, c_i32, c_i32, c_void_p]),
    "obst_broadcast_heads": (c_int, [c_int, c_void_p, c_void_p, c_i64, c_i32, c_i32, c_void_p]),
    "obst_swap_axes": (c_int, [c_int, c_void_p, c_void_p, c_i64, c_i32, c_i32, c_i32, c_void_p]),
    "obst_mul": (c_int, [c_int, c_void_p, c_void_p, c_void_p, c_i64, c_void_p]),
    "obst_scale": (c_int, [c_int, c_void_p, c_void_p, c_void_p, c_i64, c_void_p]),
    "obst_dot_accumulate": (c_int, [c_int, c_void_p, c_void_p, c_void_p, c_i64, c_void_p]),
    "obst_dropout": (c_int, [c_int, c_void_p, c_void_p, c_i64, c_float, ctypes.c_uint64, c_void_p]),
    "obst_attn_softmax_fwd": (c_int, [c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_i64, c_i32,
                                      c_i32, c_i32, c_float, c_i32, c_void_p]),
    "obst_attn_softmax_bwd": (c_int, [c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                      c_void_p, c_i64, c_i32, c_i32, c_i32, c_float, c_i32, c_void_p]),
    "obst_bias_add": (c_int, [c_int, c_void_p, c_void_p, c_void_p, c_i64, c_i64, c_void_p]),
    "obst_colsum_accumulate": (c_int, [c_int, c_void_p, c_void_p, c_i64, c_i64, c_void_p]),
    "obst_cumsum": (c_int, [c_int, c_void_p, c_void_p, c_i64, c_i32, c_i64, c_i32, c_i32, c_void_p]),
    "obst_embed_gather": (c_int, [c_int, c_void_p, c_void_p, c_void_p, c_i64, c_i32, c_i32, c_void_p]),
    "obst_embed_scatter_add": (c_int, [c_int, c_void_p, c_void_p, c_void_p, c_i64, c_i32, c_i32, c_void_p]),
    "obst_xent_fwd_bwd": (c_int, [c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_i64, c_i32, c_float, c_float,
                                  c_void_p]),
    "obst_cast_param": (c_int, [c_int, c_void_p, c_void_p, c_i64, c_i32, c_void_p]),
    "obst_transpose_param": (c_int, [c_int, c_void_p, c_void_p, c_i64, c_i32, c_i32, c_void_p]),
    "obst_opt_norms": (c_int, [c_void_p, c_i32, c_void_p, c_void_p, c_void_p, c_i32, c_i32, c_i32, c_void_p,
                               c_i32, c_void_p, c_i32, c_i32, c_void_p]),
    "obst_opt_update": (c_int, [c_void_p, c_i32, c_void_p, c_void_p, c_void_p, c_i32, c_i32, c_i32, c_i32, c_i32,
                                c_i32, c_void_p, c_void_p, c_float, c_float, c_i32, c_float, c_float, c_i32,
                                c_void_p]),
    "obst_grad_finalize": (c_int, [c_void_p, c_i32, c_void_p, c_void_p, c_void_p, c_i32, c_void_p, c_i32, c_void_p]),
    "obst_transpose_batched": (c_int, [c_int, c_void_p, c_i32, c_i32, c_void_p]),
    "obst_opt_finalize": (c_int, [c_void_p, c_i32, c_i32, c_void_p]),
    "obst_sample_step": (c_int, [c_int, c_void_p, c_void_p, c_void_p, c_i32, c_i32, c_i32, c_i32, c_float,
                                 ctypes.c_uint64, c_void_p]),
    "obst_opt_stage": (c_int, [c_void_p, c_i32, c_void_p, c_void_p, c_void_p, c_i32, ctypes.POINTER(OptStageArgs),
                               c_void_p]),
    "obst_opt_sm3_stage": (c_int, [c_void_p, c_i32, c_void_p, c_void_p, c_void_p, c_i32, c_i32, c_i32, c_i32,
                                   c_void_p]),
    "obst_opt_step_noalloc": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, ctypes.POINTER(c_void_p),
                                      ctypes.POINTER(c_i32), c_i32, c_i32, c_i32, c_i32, c_i32, c_i32, c_float,
                                      c_float, c_i32, c_float, c_float, c_void_p]),
}

_lib = None


class NativeError(RuntimeError):
    pass


def load() -> ctypes.CDLL:
    """Load the shared library (once).  Raises if it has not been built: there is no fallback path."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise NativeError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                          f"(or `make -C homebrewnlp-mtf_b200/csrc`) first; there is no CPU fallback")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def crc32c(data, crc: int = 0) -> int:
    """CRC-32C of a bytes-like object or a C-contiguous numpy array / CPU tensor (host memory), continuing from `crc`."""
    if isinstance(data, torch.Tensor):
        if data.is_cuda or not data.is_contiguous():
            raise NativeError("crc32c: a contiguous CPU tensor is expected")
        return load().obst_crc32c(data.data_ptr(), data.numel() * data.element_size(), crc)
    if hasattr(data, "ctypes"):  # numpy
        if not data.flags["C_CONTIGUOUS"]:
            raise NativeError("crc32c: a C-contiguous array is expected")
        return load().obst_crc32c(data.ctypes.data, data.nbytes, crc)
    data = bytes(data) if not isinstance(data, bytes) else data
    return load().obst_crc32c(data, len(data), crc)


def check(rc: int, what: str = ""):
    if rc != 0:
        msg = load().obst_last_error().decode("utf-8", "replace")
        raise NativeError(f"{what} failed with code {rc}: {msg}")


def dtype_code(dtype: torch.dtype) -> int:
    if dtype == torch.float32:
        return F32
    if dtype == torch.bfloat16:
        return BF16
    raise NativeError(f"unsupported dtype {dtype}")


def stream_ptr() -> int:
    return torch.cuda.current_stream().cuda_stream


def _p(t: typing.Optional[torch.Tensor]):
    return None if t is None else t.data_ptr()


def _cuda(*tensors):
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise NativeError("obst kernels need CUDA tensors (no CPU fallback)")


# ---- thin wrappers -----------------------------------------------------------------------------------------

def launch_count() -> int:
    return int(load().obst_launch_count())


def launch_count_reset():
    load().obst_launch_count_reset()


# the OBST_PATH_* enum of include/obst_b200.h
PATH_GEMM_TC, PATH_GEMM_SIMT, PATH_GN_FWD_RING, PATH_GN_BWD_RING, PATH_OPT_RING = range(5)


def path_count(which: int) -> int:
    """Calls routed to a specialised kernel so far (obst_path_count)."""
    return int(load().obst_path_count(which))


def operand(t: torch.Tensor, s_mn: int, s_k: int, s_ko: int = 0, s_glo: int = 0, s_ghi: int = 0,
            offset: int = 0) -> Operand:
    return Operand(t.data_ptr() + offset * t.element_size(), s_mn, s_k, s_ko, s_glo, s_ghi)


_PROFILE = None  # list of (start_event, end_event, flops) while gemm_profile(True)


def gemm_profile(on: bool):
    """bench.py's live roofline measurement: CUDA events around every obst_gemm launch on the launching stream.
    gemm_profile(True) starts collecting; gemm_profile(False) synchronises and returns
    (summed launch duration in ms, executed useful FLOPs, number of launches)."""
    global _PROFILE
    if on:
        _PROFILE = []
        return None
    rec, _PROFILE = _PROFILE, None
    if not rec:
        return 0.0, 0.0, 0
    torch.cuda.synchronize()
    ms = sum(s.elapsed_time(e) for s, e, _ in rec)
    return ms, float(sum(f for _, _, f in rec)), len(rec)


def gemm(A: Operand, B: Operand, D: torch.Tensor, M: int, N: int, KI: int, *, KO: int = 1, GLO: int = 1, GHI: int = 1,
         d_ld: int, d_glo: int = 0, d_ghi: int = 0, d_offset: int = 0, R: typing.Optional[torch.Tensor] = None,
         r_offset: typing.Optional[int] = None, tri: int = 0, epi: int = 0, accumulate: bool = False, splits: int = 0,
         dtype_in: torch.dtype = torch.bfloat16, backend: int = 0, aux=(), scalars=(0.0, 0.0)):
    """D[g,m,n] = epi(sum_k A[g,m,k] B[g,n,k]) -- see include/obst_b200.h for the full semantics."""
    _cuda(D, R)
    desc = GemmDesc()
    desc.dtype_in = dtype_code(dtype_in)
    desc.dtype_out = dtype_code(D.dtype)
    desc.M, desc.N, desc.KI, desc.KO, desc.GLO, desc.GHI = M, N, KI, KO, GLO, GHI
    desc.A, desc.B = A, B
    desc.D = D.data_ptr() + d_offset * D.element_size()
    if R is not None:
        ro = d_offset if r_offset is None else r_offset
        desc.R = R.data_ptr() + ro * R.element_size()
    else:
        desc.R = None
    desc.d_ld, desc.d_glo, desc.d_ghi = d_ld, d_glo, d_ghi
    desc.tri, desc.epi, desc.accumulate, desc.splits = tri, epi, int(accumulate), splits
    desc.scalar0, desc.scalar1 = scalars
    for i, a in enumerate(aux):  # epilogues 5..11: row statistics / folded-norm vectors / row scales / second tiles
        if a is None:
            continue
        _cuda(a)
        setattr(desc, f"aux{i}", a.data_ptr())
    if _PROFILE is not None:
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record()
        check(load().obst_gemm(ctypes.byref(desc), backend, stream_ptr()), "obst_gemm")
        ev1.record()
        useful = (M + 1) / (2.0 * M) if tri in (1, 2, 3) else 1.0  # causal half of a square [M, M] map
        _PROFILE.append((ev0, ev1, 2.0 * M * N * KI * KO * GLO * GHI * useful))
        return
    check(load().obst_gemm(ctypes.byref(desc), backend, stream_ptr()), "obst_gemm")


def gn_fwd(x, gamma, beta, y, heads: int, width: int, act: int = 0, stats=None):
    """stats: optional fp32 [rows, 2] tensor that receives (mean, rstd) per normalised row for gn_bwd."""
    _cuda(x, y)
    rows = x.numel() // width
    if stats is not None and (stats.dtype != torch.float32 or stats.numel() != 2 * rows or not stats.is_cuda):
        raise NativeError("gn_fwd: stats must be a CUDA fp32 tensor of 2 * rows elements")
    check(load().obst_gn_fwd(dtype_code(x.dtype), _p(x), _p(gamma), _p(beta), _p(y), _p(stats), rows, heads, width,
                             act, stream_ptr()), "obst_gn_fwd")


def gn_bwd(dy, x, gamma, beta, dx, dgamma, dbeta, heads: int, width: int, act: int = 0, relu_mask: bool = False,
           dx_accumulate: bool = False, stats=None):
    _cuda(dy, x, dx)
    rows = x.numel() // width
    if stats is not None and (stats.dtype != torch.float32 or stats.numel() != 2 * rows or not stats.is_cuda):
        raise NativeError("gn_bwd: stats must be a CUDA fp32 tensor of 2 * rows elements")
    check(load().obst_gn_bwd(dtype_code(x.dtype), _p(dy), _p(x), _p(gamma), _p(beta), _p(dx), _p(dgamma), _p(dbeta),
                             _p(stats), rows, heads, width, act, int(relu_mask), int(dx_accumulate), stream_ptr()),
          "obst_gn_bwd")


def act_fwd(kind: int, x, y):
    _cuda(x, y)
    check(load().obst_act_fwd(dtype_code(x.dtype), kind, _p(x), _p(y), x.numel(), stream_ptr()), "obst_act_fwd")


def act_bwd(kind: int, x, dy, dx):
    _cuda(x, dy, dx)
    check(load().obst_act_bwd(dtype_code(x.dtype), kind, _p(x), _p(dy), _p(dx), x.numel(), stream_ptr()),
          "obst_act_bwd")


def gate_fwd(a, b, out):
    _cuda(a, b, out)
    check(load().obst_gate_fwd(dtype_code(a.dtype), _p(a), _p(b), _p(out), a.numel(), stream_ptr()), "obst_gate_fwd")


def gate_bwd(a, b, dy, da, db):
    _cuda(a, b, dy, da, db)
    check(load().obst_gate_bwd(dtype_code(a.dtype), _p(a), _p(b), _p(dy), _p(da), _p(db), a.numel(), stream_ptr()),
          "obst_gate_bwd")


def axpby(a, b, sign: float, out):
    _cuda(a, b, out)
    check(load().obst_axpby(dtype_code(a.dtype), _p(a), _p(b), sign, _p(out), a.numel(), stream_ptr()), "obst_axpby")


def sum_heads(x, out, rows: int, heads: int, width: int):
    _cuda(x, out)
    check(load().obst_sum_heads(dtype_code(x.dtype), _p(x), _p(out), rows, heads, width, stream_ptr()), "obst_sum_heads")


def broadcast_heads(dy, dx, rows: int, heads: int, width: int):
    _cuda(dy, dx)
    check(load().obst_broadcast_heads(dtype_code(dy.dtype), _p(dy), _p(dx), rows, heads, width, stream_ptr()),
          "obst_broadcast_heads")


def swap_axes(x, out, outer: int, a: int, mid: int, b: int):
    _cuda(x, out)
    check(load().obst_swap_axes(dtype_code(x.dtype), _p(x), _p(out), outer, a, mid, b, stream_ptr()), "obst_swap_axes")


def mul(a, b, out):
    _cuda(a, b, out)
    check(load().obst_mul(dtype_code(a.dtype), _p(a), _p(b), _p(out), a.numel(), stream_ptr()), "obst_mul")


def scale(x, scalar, out):
    _cuda(x, scalar, out)
    check(load().obst_scale(dtype_code(x.dtype), _p(x), _p(scalar), _p(out), x.numel(), stream_ptr()), "obst_scale")


def dot_accumulate(a, b, out):
    """out[0] (fp32) += sum(a * b)"""
    _cuda(a, b, out)
    if out.dtype != torch.float32:
        raise NativeError("dot_accumulate: the accumulator must be fp32")
    check(load().obst_dot_accumulate(dtype_code(a.dtype), _p(a), _p(b), _p(out), a.numel(), stream_ptr()),
          "obst_dot_accumulate")


def dropout(x, out, keep: float, seed: int):
    _cuda(x, out)
    check(load().obst_dropout(dtype_code(x.dtype), _p(x), _p(out), x.numel(), float(keep), int(seed) & (2 ** 64 - 1),
                              stream_ptr()), "obst_dropout")


def norm_stats_finalize(partials, stats, rows: int, heads: int, per_head: int, width: int):
    _cuda(partials, stats)
    check(load().obst_norm_stats_finalize(_p(partials), _p(stats), rows, heads, per_head, width, stream_ptr()),
          "obst_norm_stats_finalize")


def fold_items(entries) -> torch.Tensor:
    """Device table of obst_fold_item from (w, gamma, beta, wg_t, cvec, dvec, gamma_f[, wg]) tensor tuples."""
    items = (FoldItem * len(entries))()
    for it, entry in zip(items, entries):
        w, gamma, beta, wg_t, cvec, dvec, gamma_f = entry[:7]
        it.wg = _p(entry[7]) if len(entry) > 7 else None
        _cuda(w, wg_t, cvec, dvec, gamma_f)
        if w.dtype != torch.bfloat16:
            raise NativeError("fold_norm_weight: bf16 weights only")
        it.w, it.gamma, it.beta, it.wg_t = _p(w), _p(gamma), _p(beta), _p(wg_t)
        it.cvec, it.dvec, it.gamma_f = _p(cvec), _p(dvec), _p(gamma_f)
    return torch.frombuffer(bytearray(bytes(items)), dtype=torch.uint8).to(entries[0][0].device)


def fold_norm_weight(table: torch.Tensor, nitems: int, heads: int, kin: int, kout: int):
    check(load().obst_fold_norm_weight(_p(table), nitems, heads, kin, kout, stream_ptr()), "obst_fold_norm_weight")


def rowscale_colsum(dy, stats, dyr, v1, v2, tokens: int, heads: int, width: int):
    _cuda(dy, stats, dyr, v1, v2)
    check(load().obst_rowscale_colsum(_p(dy), _p(stats), _p(dyr), _p(v1), _p(v2), tokens, heads, width, stream_ptr()),
          "obst_rowscale_colsum")


def norm_bwd_coeffs(partials, stats, coeffs, rows: int, heads: int, nparts: int, width: int):
    """(a, b, c) per (row, head) of the folded norm's backward from epilogue 9 / 10's partial dot products."""
    _cuda(partials, stats, coeffs)
    check(load().obst_norm_bwd_coeffs(_p(partials), _p(stats), _p(coeffs), rows, heads, nparts, width, stream_ptr()),
          "obst_norm_bwd_coeffs")


def norm_affine_grad(w, v1, v2, dgamma, dbeta, heads: int, kin: int, kout: int):
    """dgamma -= W v1, dbeta += W v2 per head (the halves of the norm's scale / shift gradients that need no dn)."""
    _cuda(w, v1, v2)
    check(load().obst_norm_affine_grad(_p(w), _p(v1), _p(v2), _p(dgamma), _p(dbeta), heads, kin, kout, stream_ptr()),
          "obst_norm_affine_grad")


def norm_wgrad_rank1(dw, gamma, beta, v1, v2, heads: int, kin: int, kout: int):
    _cuda(dw, v1, v2)
    check(load().obst_norm_wgrad_rank1(_p(dw), _p(gamma), _p(beta), _p(v1), _p(v2), heads, kin, kout, stream_ptr()),
          "obst_norm_wgrad_rank1")


def attn_softmax_fwd(logits, bias, add_map, mul_map, soft, probs, batch, seq, heads, keys, scale, masked):
    _cuda(logits, soft)
    check(load().obst_attn_softmax_fwd(dtype_code(logits.dtype), _p(logits), _p(bias), _p(add_map), _p(mul_map), _p(soft),
                                       _p(probs), batch, seq, heads, keys, float(scale), int(masked), stream_ptr()),
          "obst_attn_softmax_fwd")


def attn_softmax_bwd(dprobs, soft, add_map, mul_map, dlogits, dbias, dadd, dmul, batch, seq, heads, keys, scale,
                     masked):
    _cuda(dprobs, soft, dlogits)
    check(load().obst_attn_softmax_bwd(dtype_code(soft.dtype), _p(dprobs), _p(soft), _p(add_map), _p(mul_map),
                                       _p(dlogits), _p(dbias), _p(dadd), _p(dmul), batch, seq, heads, keys, float(scale),
                                       int(masked), stream_ptr()), "obst_attn_softmax_bwd")


def bias_add(x, bias, out):
    _cuda(x, bias, out)
    n = bias.numel()
    check(load().obst_bias_add(dtype_code(x.dtype), _p(x), _p(bias), _p(out), x.numel() // n, n, stream_ptr()),
          "obst_bias_add")


def colsum_accumulate(x, out):
    _cuda(x, out)
    n = out.numel()
    check(load().obst_colsum_accumulate(dtype_code(x.dtype), _p(x), _p(out), x.numel() // n, n, stream_ptr()),
          "obst_colsum_accumulate")


def cumsum(x, out, outer: int, seq: int, inner: int, reverse: bool = False, scale: int = 0):
    """Running sum along the middle axis of [outer, seq, inner] (obst_cumsum)."""
    _cuda(x, out)
    if x.data_ptr() == out.data_ptr():
        raise NativeError("cumsum: in-place operation is not supported")
    check(load().obst_cumsum(dtype_code(x.dtype), _p(x), _p(out), outer, seq, inner, int(reverse), scale, stream_ptr()),
          "obst_cumsum")


def embed_gather(idx, table, out):
    _cuda(idx, table, out)
    vocab, width = table.shape
    check(load().obst_embed_gather(dtype_code(table.dtype), _p(idx), _p(table), _p(out), idx.numel(), vocab, width,
                                   stream_ptr()), "obst_embed_gather")


def embed_scatter_add(idx, dout, dtable):
    _cuda(idx, dout, dtable)
    vocab, width = dtable.shape
    check(load().obst_embed_scatter_add(dtype_code(dout.dtype), _p(idx), _p(dout), _p(dtable), idx.numel(), vocab,
                                        width, stream_ptr()), "obst_embed_scatter_add")


def xent_fwd_bwd(logits, targets, dlogits, sums, inv_n: float, z_loss: float):
    _cuda(logits, targets, dlogits, sums)
    vocab = logits.shape[-1]
    tokens = logits.numel() // vocab
    check(load().obst_xent_fwd_bwd(dtype_code(logits.dtype), _p(logits), _p(targets), _p(dlogits), _p(sums), tokens,
                                   vocab, inv_n, z_loss, stream_ptr()), "obst_xent_fwd_bwd")


def cast_param(src, dst, tril_s: int = 0):
    _cuda(src, dst)
    check(load().obst_cast_param(dtype_code(dst.dtype), _p(src), _p(dst), src.numel(), tril_s, stream_ptr()),
          "obst_cast_param")


def sample_step(logits, tokens, position: int, temperature: float = 0.0, seed: int = 0, uniform=None):
    """tokens[b, position] <- Gumbel-argmax of logits[b, position - 1] (obst_sample_step)."""
    _cuda(logits, tokens)
    if tokens.dtype != torch.int32 or not tokens.is_contiguous() or not logits.is_contiguous():
        raise NativeError("sample_step: tokens must be contiguous int32, logits contiguous")
    vocab = logits.shape[-1]
    batch = tokens.shape[0]
    seq = tokens.numel() // max(batch, 1)
    if logits.numel() != batch * seq * vocab:
        raise NativeError(f"sample_step: logits {tuple(logits.shape)} do not match tokens {tuple(tokens.shape)}")
    check(load().obst_sample_step(dtype_code(logits.dtype), _p(logits), _p(tokens), _p(uniform), batch, seq, vocab,
                                  int(position), float(temperature), int(seed) & (2 ** 64 - 1), stream_ptr()),
          "obst_sample_step")


def transpose_param(src, dst):
    """dst[..., c, r] = src[..., r, c] (contiguous tensors, leading dims are the batch)."""
    _cuda(src, dst)
    rows, cols = src.shape[-2], src.shape[-1]
    batch = src.numel() // max(1, rows * cols)
    check(load().obst_transpose_param(dtype_code(src.dtype), _p(src), _p(dst), batch, rows, cols, stream_ptr()),
          "obst_transpose_param")
