"""CPU: the C-ABI library loads and exports exactly what include/srb200.h declares;
struct layouts of the ctypes/numpy mirrors match the header; no compute without a GPU."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "srb200.h")


def _declared():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(srb_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_what_the_binding_lists():
    from softwarerenderer_b200 import _native as N
    assert _declared() == sorted(N.EXPORTS)


def test_library_exports_every_declared_symbol():
    from softwarerenderer_b200 import _native as N
    L = N.lib()
    for name in _declared():
        assert hasattr(L, name), f"{name} is declared in srb200.h but not exported by libsrb200.so"
    assert L.srb_version() >= 100


def test_exported_symbols_are_plain_c():
    from softwarerenderer_b200 import _native as N
    out = subprocess.run(["nm", "-D", "--defined-only", N.LIB_PATH], capture_output=True, text=True).stdout
    syms = {ln.split()[-1] for ln in out.splitlines() if " T " in ln}
    assert set(_declared()) <= syms
    # nothing from torch / no C++ types leak through the boundary
    assert not [s for s in syms if "torch" in s.lower() or "at::" in s]


def test_header_compiles_as_c_and_struct_sizes_match(tmp_path):
    from softwarerenderer_b200 import _native as N
    src = tmp_path / "sz.c"
    src.write_text('#include <stdio.h>\n#include "srb200.h"\nint main(void){printf("%zu %zu %zu\\n",'
                   'sizeof(SrbUniforms),sizeof(SrbDrawCmd),sizeof(SrbStats));return 0;}\n')
    exe = tmp_path / "sz"
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)])
    u, d, s = (int(x) for x in subprocess.check_output([str(exe)]).split())
    assert u == C.sizeof(N.SrbUniforms)
    assert d == N.DRAW_CMD_DTYPE.itemsize == 248
    assert s == C.sizeof(N.SrbStats)


def test_oracle_uniforms_layout_matches_product_uniforms():
    from oracle import pyoracle as O
    from softwarerenderer_b200 import _native as N
    assert [(n, C.sizeof(t)) for n, t in O.OrcUniforms._fields_] == [(n, C.sizeof(t)) for n, t in N.SrbUniforms._fields_]


def test_no_cpu_fallback_without_a_device():
    """Without a CUDA device srb_create must FAIL (SRB_ERR_CUDA), never fall back to the CPU."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    import softwarerenderer_b200 as srb
    with pytest.raises(srb.SrbError) as e:
        srb.MainWindow(64, 64)
    assert e.value.code == -2 and "no CPU fallback" in str(e.value)


def test_argument_errors_do_not_need_a_device():
    from softwarerenderer_b200 import _native as N
    L = N.lib()
    ctx = C.c_void_p()
    assert L.srb_create(0, 0, 10, C.byref(ctx)) == N.SRB_ERR_ARG           # ArgumentException analogue (Rasterizer.cs:71-74)
    assert b"positive non-zero" in L.srb_last_error(None)
    assert L.srb_create(0, 70000, 10, C.byref(ctx)) == N.SRB_ERR_ARG
    assert L.srb_destroy(None) == N.SRB_ERR_ARG
    assert L.srb_end_frame(None) == N.SRB_ERR_ARG


def test_product_does_not_link_or_import_the_oracle():
    from softwarerenderer_b200 import _native as N
    out = subprocess.run(["ldd", N.LIB_PATH], capture_output=True, text=True).stdout
    assert "oracle" not in out
    pkg = os.path.join(ROOT, "softwarerenderer_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dp, f)).read()
                assert "pyoracle" not in text and "liboracle" not in text and "import oracle" not in text, f


def test_csharp_binding_covers_the_header():
    """csharp/Native.cs (the P/Invoke stub a maintainer adds) binds every declared symbol."""
    text = open(os.path.join(ROOT, "csharp", "Native.cs")).read()
    bound = set(re.findall(r"extern\s+\w+\s+(srb_[a-z0-9_]+)\s*\(", text))
    assert bound == set(_declared())
