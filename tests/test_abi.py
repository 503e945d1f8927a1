"""The C-ABI library loads and exports every symbol include/obst_b200.h declares; the ctypes mirror of its
structs has the C compiler's layout.  No compute calls (no GPU needed)."""
import ctypes
import os
import re
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "obst_b200.h")


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(obst_[a-z0-9_]+)\s*\(", src)))


@pytest.fixture(scope="module")
def lib():
    import __graft_entry__ as g
    from obst import native
    if not os.path.exists(native.LIB_PATH):
        g.build()
    return native.load()


def test_header_symbols_are_exported_and_bound(lib):
    from obst import native
    names = declared_symbols()
    assert len(names) >= 17
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/obst_b200.h but not exported"
        assert n in native.SIGNATURES, f"{n} has no ctypes signature in obst/native.py"
    assert sorted(native.SIGNATURES) == names


def test_version_error_text_and_launch_counter(lib):
    assert lib.obst_version() == 100
    assert isinstance(lib.obst_last_error(), bytes)
    lib.obst_launch_count_reset()
    assert lib.obst_launch_count() == 0


def test_argument_validation_needs_no_gpu(lib):
    """Entry points reject bad arguments before touching the device (error behaviour of the boundary)."""
    from obst import native
    assert lib.obst_gemm(None, 0, None) == -1
    assert b"null descriptor" in lib.obst_last_error()
    d = native.GemmDesc()
    d.M = d.N = d.KI = d.KO = d.GLO = d.GHI = 0
    assert lib.obst_gemm(ctypes.byref(d), 0, None) == -1
    assert lib.obst_xent_fwd_bwd(0, None, None, None, None, 4, 4096, 1.0, 0.0, None) == -1
    assert b"vocab" in lib.obst_last_error()
    assert lib.obst_act_fwd(0, 77, None, None, 8, None) == -1
    assert lib.obst_gn_fwd(1, None, None, None, None, None, 8, 8, 100, 0, None) == -1  # width not a multiple of 8
    from obst import native
    a = native.OptStageArgs()
    assert lib.obst_opt_stage(None, 1, None, None, None, 1, ctypes.byref(a), None) == -1
    assert b"null argument" in lib.obst_last_error()
    one = ctypes.c_int32(0)
    ptr = ctypes.addressof(one)
    a.stage = 99
    assert lib.obst_opt_stage(ptr, 1, ptr, ptr, ptr, 1, ctypes.byref(a), None) == -1 and b"bad stage" in lib.obst_last_error()
    a.stage = native.STAGE_ADAM
    assert lib.obst_opt_stage(ptr, 1, ptr, ptr, ptr, 1, ctypes.byref(a), None) == -1   # no buffers
    assert lib.obst_opt_sm3_stage(ptr, 1, ptr, ptr, ptr, 1, 512, 1, 7, None) == -1 and b"state dtype" in lib.obst_last_error()
    assert lib.obst_opt_update(ptr, 1, ptr, ptr, ptr, 1, 512, 1, 1, 7, 0, ptr, None, 0.003, 0.9, 1, 0.01, 0.0, 0,
                               None) == -1 and b"state dtype" in lib.obst_last_error()
    assert lib.obst_grad_finalize(ptr, 1, ptr, ptr, ptr, 1, None, 1, None) == -1 and b"norms" in lib.obst_last_error()
    assert lib.obst_transpose_batched(1, None, 3, 5, None) == -1 and b"null item" in lib.obst_last_error()
    assert lib.obst_transpose_batched(1, None, 0, 0, None) == 0


def test_struct_layout_matches_the_c_compiler(tmp_path):
    from obst import native
    c = tmp_path / "sz.c"
    c.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "obst_b200.h"\nint main(){'
                 'printf("%zu %zu %zu %zu %zu %zu %zu %zu %zu %zu %zu %zu %zu\\n", sizeof(obst_operand), '
                 'sizeof(obst_gemm_desc), '
                 'sizeof(obst_opt_tensor), offsetof(obst_gemm_desc, D), offsetof(obst_opt_tensor, numel), '
                 'offsetof(obst_opt_tensor, tril_s), sizeof(obst_opt_stage_args), '
                 'offsetof(obst_opt_stage_args, grad_base), offsetof(obst_opt_stage_args, sums2), '
                 'offsetof(obst_opt_tensor, g_acc), offsetof(obst_opt_tensor, g_bf16), sizeof(obst_transpose_item), '
                 'offsetof(obst_transpose_item, tile0)); return 0;}')
    exe = tmp_path / "sz"
    subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), str(c), "-o", str(exe)], check=True)
    out = subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.split()
    got = [ctypes.sizeof(native.Operand), ctypes.sizeof(native.GemmDesc), ctypes.sizeof(native.OptTensor),
           native.GemmDesc.D.offset, native.OptTensor.numel.offset, native.OptTensor.tril_s.offset,
           ctypes.sizeof(native.OptStageArgs), native.OptStageArgs.grad_base.offset,
           native.OptStageArgs.sums2.offset, native.OptTensor.g_acc.offset, native.OptTensor.g_bf16.offset,
           ctypes.sizeof(native.TransposeItem), native.TransposeItem.tile0.offset]
    assert [int(v) for v in out] == got


def test_product_never_imports_the_oracle():
    """oracle/ is test infrastructure: nothing under homebrewnlp-mtf_b200/ may import or execute it."""
    pkg = os.path.join(ROOT, "homebrewnlp-mtf_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in text.replace("# oracle", ""), os.path.join(dirpath, f)


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    from obst import native
    monkeypatch.setattr(native, "_lib", None)
    monkeypatch.setattr(native, "LIB_PATH", str(tmp_path / "nope.so"))
    with pytest.raises(native.NativeError):
        native.load()
