"""CPU: the C-ABI library loads and exports every symbol include/gsr.h declares (no compute calls)."""
import ctypes
import os
import re

import pytest

from helpers import ROOT


def _header_decls():
    src = open(os.path.join(ROOT, "include", "gsr.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return re.findall(r"GSR_API\s+([\w\s\*]+?)\s*(gsr_\w+)\s*\(([^;]*?)\)\s*;", src, flags=re.S)


def test_header_declares_entry_points():
    names = [n for _, n, _ in _header_decls()]
    for must in ["gsr_forward_prepare", "gsr_forward_render", "gsr_backward", "gsr_mark_visible", "gsr_debug_dump",
                 "gsr_geometry_bytes", "gsr_image_bytes", "gsr_binning_bytes", "gsr_last_error", "gsr_version"]:
        assert must in names


def test_library_exports_every_declared_symbol():
    from diff_gaussian_rasterization import _cabi
    lib = ctypes.CDLL(_cabi.LIB_PATH)
    for _, name, _ in _header_decls():
        assert hasattr(lib, name), f"{name} declared in gsr.h but not exported"


def test_ctypes_signatures_match_header():
    from diff_gaussian_rasterization import _cabi

    def ctype_of(a):
        a = a.strip()
        if a in ("void", ""):
            return None
        if "*" in a:
            return ctypes.c_void_p
        t = " ".join(a.split()[:-1])
        return {"int": ctypes.c_int, "float": ctypes.c_float, "int64_t": ctypes.c_int64, "size_t": ctypes.c_size_t,
                "uint32_t": ctypes.c_uint32}[t]

    decls = _header_decls()
    assert {n for _, n, _ in decls} == set(_cabi.SIGNATURES), "binding and header list different entry points"
    for _, name, args in decls:
        want = [c for c in (ctype_of(a) for a in args.split(",")) if c is not None]
        assert want == _cabi.SIGNATURES[name][1], name


def test_size_queries_need_no_gpu():
    from diff_gaussian_rasterization import _cabi
    lib = _cabi.lib
    assert lib.gsr_version() >= 1000
    g1, g2 = lib.gsr_geometry_bytes(1000), lib.gsr_geometry_bytes(2000)
    assert 0 < g1 < g2
    assert lib.gsr_image_bytes(640, 360) >= 640 * 360 * 4 + 40 * 23 * 8
    assert lib.gsr_binning_bytes(1000, 640, 360) >= 16 * 1000
    assert lib.gsr_profile_count() == 9      # stages with a timer (the tile ranges come out of the tile sort)
    assert lib.gsr_profile_name(0) == b"preprocess"
    # the moment buffer of the model step: 2 x (11 per-Gaussian floats + 3 dense dc floats + 3 M SH floats); gsr_parallel
    # re-lays it out for the padded count of the shared multi-GPU step and must agree with the library on its size
    import gsr_parallel
    for P, M in [(0, 16), (1, 1), (4999, 16), (1036800, 16), (12, 4)]:
        assert lib.gsr_gaussian_adam_state_floats(P, M) == 2 * (14 + 3 * M) * P == gsr_parallel._state_floats(P, M)


def test_header_is_plain_c(tmp_path):
    """include/gsr.h is the boundary a C host binds: it must compile as C99 and link against the library by name."""
    import shutil
    import subprocess
    if shutil.which("gcc") is None:
        pytest.skip("gcc not available")
    src = tmp_path / "use_gsr.c"
    src.write_text('#include "gsr.h"\n#include <stdio.h>\n'
                   "int main(void) { printf(\"%d %zu\\n\", gsr_version(), gsr_geometry_bytes(1000)); return 0; }\n")
    inc = os.path.join(ROOT, "include")
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-pedantic", "-I", inc, "-fsyntax-only", str(src)], check=True)
    lib = os.path.join(ROOT, "cf-3dgs_b200", "libgsr_b200.so")
    if os.path.exists(lib):
        exe = tmp_path / "use_gsr"
        subprocess.run(["gcc", "-std=c99", "-I", inc, str(src), "-o", str(exe), lib, "-Wl,-rpath," + os.path.dirname(lib)],
                       check=True)
        out = subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.split()
        assert int(out[0]) >= 1000 and int(out[1]) > 0
