"""CPU-side checks of the drop-in boundary: the C-ABI library loads, exports every symbol that
include/canvas_b200.h and include/framework.h declare, keeps the reference's struct layouts, and
fails loudly (no CPU fallback) when no GPU is present."""
import ctypes as C
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    import __graft_entry__ as g
    if not os.path.exists(os.path.join(ROOT, "canvas_b200", "libcanvas_b200.so")):
        g.build()
    from canvas_b200 import _abi
    return _abi.load()


def declared_functions(header):
    text = open(os.path.join(ROOT, "include", header)).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    names = set()
    for m in re.finditer(r"^\s*(?:const\s+)?[A-Za-z_][\w\s\*]*?\b(\w+)\s*\([^;{]*\)\s*;", text, flags=re.M):
        line = m.group(0)
        if "typedef" in line or re.search(r"\bstatic\b", line) or "(*" in line:      # the keyword, not a name like cv_..._static_f32
            continue
        names.add(m.group(1))
    return names


def test_every_cv_symbol_is_exported(lib):
    names = declared_functions("canvas_b200.h")
    assert len(names) > 40
    missing = [n for n in sorted(names) if not hasattr(lib, n)]
    assert not missing, "declared in include/canvas_b200.h but not exported: %s" % missing


def test_binding_covers_header(lib):
    from canvas_b200 import _abi
    assert set(_abi.exported_names()) == declared_functions("canvas_b200.h")


def test_struct_layouts_match_c():
    src = r'''
    #include "canvas_b200.h"
    #include <stdio.h>
    #include <stddef.h>
    int main(void) {
        printf("%zu %zu %zu %zu %zu %zu %zu %zu %zu\n", sizeof(rgba_f16), sizeof(box2i), sizeof(rgba_frame_f16), sizeof(rgba_frame_f32),
               sizeof(cv_coded), sizeof(cv_layer), sizeof(cv_source), sizeof(cv_frame_job), offsetof(cv_frame_job, out));
        return 0;
    }'''
    exe = "/tmp/cv_layout_probe"
    subprocess.run(["gcc", "-x", "c", "-", "-I", os.path.join(ROOT, "include"), "-o", exe], input=src.encode(), check=True)
    sizes = [int(v) for v in subprocess.run([exe], capture_output=True, check=True).stdout.split()]
    from canvas_b200 import _abi as A
    assert sizes == [8, 16, 40, 40, C.sizeof(A.Coded), C.sizeof(A.Layer), C.sizeof(A.Source), C.sizeof(A.FrameJob), A.FrameJob.out.offset]


def test_no_cpu_fallback_without_gpu(lib):
    if lib.cv_device_count() > 0:
        pytest.skip("a GPU is present")
    assert lib.cv_init(0) < 0
    assert b"no CPU fallback" in lib.cv_last_error()
    from canvas_b200 import _abi as A
    frame = A.FrameF16()
    frame.current_window = A.Box2i(0, 0, 5, 5)
    coded = A.Coded()
    assert lib.cv_convert_ycbcr8(C.byref(frame), C.byref(coded), None) < 0
    assert frame.current_window.empty          # the reference's error convention: empty window


def test_host_tables_match_oracle(lib):
    import numpy as np
    import oraclelib as O
    for which in range(4):
        want = np.zeros(65536, np.uint16)
        O.oracle().orc_build_lut(which, O.ptr(want))
        got = np.ctypeslib.as_array(lib.cv_host_lut(which), shape=(65536,))
        assert np.array_equal(got, want)
    want = np.zeros(65536, np.uint8)
    O.oracle().orc_build_gamma45(O.ptr(want))
    assert np.array_equal(np.ctypeslib.as_array(lib.cv_host_gamma45_ramp(), shape=(65536,)), want)


def test_nothing_but_the_headers_is_exported(lib):
    """The drop-in exports include/*.h and nothing else: no kernel stubs, no cvi_* internals, no CUDA runtime
    (canvas_b200/csrc/exports.map)."""
    out = subprocess.run(["nm", "-D", "--defined-only", os.path.join(ROOT, "canvas_b200", "libcanvas_b200.so")],
                         capture_output=True, text=True, check=True).stdout
    exported = {line.split()[-1] for line in out.splitlines() if line.strip()}
    declared = declared_functions("canvas_b200.h") | declared_functions("framework.h") | {
        "init_half", "half_convert_from_float", "half_convert_to_float", "half_convert_from_float_fast",
        "half_convert_to_float_fast", "half_lookup"}
    extra = sorted(exported - declared)
    assert not extra, "exported but not declared in include/*.h: %s" % extra
    missing = sorted(declared - exported)
    assert not missing, "declared in include/*.h but not exported: %s" % missing
