"""The ctypes mirror of the C ABI must match include/tadnn.h, and argument validation in the native
library must refuse bad calls with a message *before* touching the device (so these run without a GPU)."""
import ctypes as C
import os
import re
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "tadnn.h")


def _lib():
    from tadnn import _native
    if not _native.available():
        pytest.skip("libtadnn.so is not built (python build_native.py)")
    return _native.lib()


def _c_layout(tmp_path, struct, fields):
    """Compile a probe against the header with the host compiler and return {field: offset, '': sizeof}."""
    cc = shutil.which("gcc") or shutil.which("cc")
    if cc is None:
        pytest.skip("no host C compiler")
    lines = ['#include <stdio.h>', '#include <stddef.h>', f'#include "{HEADER}"', 'int main(void) {',
             f'  printf(" %zu\\n", sizeof({struct}));']
    lines += [f'  printf("{f} %zu\\n", offsetof({struct}, {f}));' for f in fields]
    lines += ['  return 0;', '}']
    src = tmp_path / f"probe_{struct}.c"
    src.write_text("\n".join(lines))
    exe = tmp_path / f"probe_{struct}"
    subprocess.run([cc, "-std=c99", "-o", str(exe), str(src)], check=True, capture_output=True)
    out = subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout
    res = {}
    for ln in out.splitlines():
        name, _, val = ln.rpartition(" ")
        res[name.strip()] = int(val)
    return res


def test_gemm_args_layout_matches_header(tmp_path):
    from tadnn._native import GemmArgs
    names = [f[0] for f in GemmArgs._fields_]
    c = _c_layout(tmp_path, "tadnn_gemm_args", names)
    assert c[""] == C.sizeof(GemmArgs)
    for n in names:
        assert getattr(GemmArgs, n).offset == c[n], n


def test_head_args_layout_matches_header(tmp_path):
    from tadnn.ops.cnn import HeadArgs
    names = [f[0] for f in HeadArgs._fields_]
    c = _c_layout(tmp_path, "tadnn_head_args", names)
    assert c[""] == C.sizeof(HeadArgs)
    for n in names:
        assert getattr(HeadArgs, n).offset == c[n], n


def test_every_declared_entry_point_is_exported():
    lib = _lib()
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    declared = set(re.findall(r"\b(tadnn_[a-z0-9_]+)\s*\(", text))
    assert len(declared) > 50
    missing = [n for n in sorted(declared) if not hasattr(lib, n)]
    assert not missing, missing


def test_exported_entry_points_are_declared():
    _lib()
    from tadnn import _native
    nm = shutil.which("nm")
    if nm is None:
        pytest.skip("no nm")
    out = subprocess.run([nm, "-D", "--defined-only", _native.LIB_PATH], check=True, capture_output=True, text=True).stdout
    exported = {ln.split()[-1] for ln in out.splitlines() if ln.split() and ln.split()[-1].startswith("tadnn_")}
    text = open(HEADER).read()
    undeclared = [n for n in sorted(exported) if not re.search(r"\b" + n + r"\s*\(", text)]
    assert not undeclared, undeclared


def test_smalln_validation_messages():
    lib = _lib()
    buf = (C.c_char * 256)()  # stands in for device memory: validation must fail before any dereference
    p = C.cast(buf, C.c_void_p)
    rc = lib.tadnn_linear_smalln_fwd(p, 64, p, 64, None, p, 17, 4, 17, 64, None)
    assert rc != 0 and b"outside 1..16" in lib.tadnn_last_error()
    rc = lib.tadnn_linear_smalln_fwd(p, 64, p, 64, None, p, 16, 4, 10, 60, None)
    assert rc != 0 and b"multiple of 8" in lib.tadnn_last_error()
    rc = lib.tadnn_linear_smalln_dgrad(p, 16, p, 60, None, 0, p, 64, 4, 10, 64, None)
    assert rc != 0 and b"16-byte aligned rows" in lib.tadnn_last_error()
    rc = lib.tadnn_linear_smalln_wgrad(p, 8, p, 64, p, 64, None, 4, 10, 64, 0, None)
    assert rc != 0 and b"bad arguments" in lib.tadnn_last_error()  # gradient pitch 8 < N = 10


def test_preprocess_validation_messages():
    lib = _lib()
    buf = (C.c_char * 256)()
    p = C.cast(buf, C.c_void_p)
    taps = (C.c_float * 7)(*([1.0] * 7))
    args = lambda n_taps, H, W, su: (p, p, 0, C.c_longlong(1), H, W, taps, n_taps, C.c_float(0), C.c_float(su),
                                     C.c_float(0), C.c_float(1), C.c_float(1e-4), C.c_float(1e-4), None)
    rc = lib.tadnn_preprocess_yuv_lcn(*args(6, 32, 32, 1.0))
    assert rc != 0 and b"odd tap count" in lib.tadnn_last_error()
    rc = lib.tadnn_preprocess_yuv_lcn(*args(7, 128, 128, 1.0))
    assert rc != 0 and b"5120-pixel" in lib.tadnn_last_error()
    rc = lib.tadnn_preprocess_yuv_lcn(*args(7, 32, 32, 0.0))
    assert rc != 0 and b"zero standard deviation" in lib.tadnn_last_error()


def test_launch_counter_is_untouched_by_refused_calls():
    lib = _lib()
    from tadnn import _native
    before = _native.launch_count()
    buf = (C.c_char * 64)()
    lib.tadnn_linear_smalln_fwd(C.cast(buf, C.c_void_p), 64, C.cast(buf, C.c_void_p), 64, None, C.cast(buf, C.c_void_p), 17, 4, 17, 64, None)
    assert _native.launch_count() == before
