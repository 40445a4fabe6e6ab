"""The reference's own C API (include/framework.h, include/half.h) as exported by libcanvas_b200.so.

CPU part: symbols, workspace bookkeeping, FIR taps, coded images, time helpers, plain copies.
GPU part (-m gpu): the host-pointer entry points against the oracle, bit for bit, including a
workspace built from C-callback sources exactly like the reference's tests build theirs.
"""
import ctypes as C
import os
import random

import numpy as np
import pytest

import oraclelib as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

GET_FRAME = C.CFUNCTYPE(None, C.c_void_p, C.c_int, C.POINTER(O.Frame16))
GET_FRAME_32 = C.CFUNCTYPE(None, C.c_void_p, C.c_int, C.POINTER(O.Frame32))


class SourceFuncs(C.Structure):
    """video_frame_source_funcs (framework.h:185-194)"""
    _fields_ = [("flags", C.c_int), ("get_frame", C.c_void_p), ("get_frame_32", C.c_void_p), ("get_frame_gl", C.c_void_p)]


class VideoSource(C.Structure):
    _fields_ = [("obj", C.c_void_p), ("funcs", C.POINTER(SourceFuncs))]


class FirFilter(C.Structure):
    _fields_ = [("coeff", C.POINTER(C.c_float)), ("width", C.c_int), ("center", C.c_int)]


class Rational(C.Structure):
    _fields_ = [("n", C.c_int32), ("d", C.c_uint32)]


@pytest.fixture(scope="module")
def lib():
    from canvas_b200 import _abi
    L = _abi.load()
    L.workspace_create.restype = C.c_void_p
    L.workspace_add_item.restype = C.c_void_p
    L.workspace_add_item.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_int64, C.c_void_p]
    L.workspace_get_item.restype = C.c_void_p
    L.workspace_get_item.argtypes = [C.c_void_p, C.c_int]
    L.workspace_get_length.argtypes = [C.c_void_p]
    L.workspace_remove_item.argtypes = [C.c_void_p]
    L.workspace_free.argtypes = [C.c_void_p]
    L.workspace_get_item_pos.argtypes = [C.c_void_p, C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
    L.workspace_get_item_offset.restype = C.c_int64
    L.workspace_get_item_offset.argtypes = [C.c_void_p]
    L.workspace_get_item_tag.restype = C.c_void_p
    L.workspace_get_item_tag.argtypes = [C.c_void_p]
    L.workspace_update_item.argtypes = [C.c_void_p] + [C.POINTER(C.c_int64)] * 4 + [C.POINTER(C.c_void_p)] * 2
    L.workspace_as_video_source.argtypes = [C.c_void_p, C.POINTER(VideoSource)]
    L.video_get_frame_f16.argtypes = [C.POINTER(VideoSource), C.c_int, C.POINTER(O.Frame16)]
    L.video_get_frame_f32.argtypes = [C.POINTER(VideoSource), C.c_int, C.POINTER(O.Frame32)]
    L.video_copy_frame_f16.argtypes = [C.POINTER(O.Frame16), C.POINTER(O.Frame16)]
    L.video_copy_frame_alpha_f32.argtypes = [C.POINTER(O.Frame32), C.POINTER(O.Frame32), C.c_float]
    L.video_mix_cross_f32.argtypes = [C.POINTER(O.Frame32), C.POINTER(O.Frame32), C.POINTER(O.Frame32), C.c_float]
    L.video_mix_over_f32.argtypes = [C.POINTER(O.Frame32), C.POINTER(O.Frame32), C.c_float]
    L.video_mix_cross_f32_pull.argtypes = [C.POINTER(O.Frame32), C.POINTER(VideoSource), C.c_int, C.POINTER(VideoSource), C.c_int, C.c_float]
    L.filter_createTriangle.argtypes = [C.c_float, C.c_float, C.POINTER(FirFilter)]
    L.filter_free.argtypes = [C.POINTER(FirFilter)]
    L.coded_image_alloc0.restype = C.POINTER(O.RefCodedImage)
    L.coded_image_alloc0.argtypes = [C.POINTER(C.c_int), C.POINTER(C.c_int), C.c_int]
    L.get_frame_time.restype = C.c_int64
    L.get_frame_time.argtypes = [C.POINTER(Rational), C.c_int]
    L.get_time_frame.argtypes = [C.POINTER(Rational), C.c_int64]
    L.video_get_gamma45_ramp.restype = C.POINTER(C.c_uint8)
    L.video_reconstruct_dv.argtypes = [C.POINTER(O.Frame16), C.POINTER(O.RefCodedImage)]
    L.video_subsample_dv.restype = C.POINTER(O.RefCodedImage)
    L.video_subsample_dv.argtypes = [C.POINTER(O.Frame16)]
    for name in ("video_transfer_rec709_to_linear_scene", "video_transfer_rec709_to_linear_display",
                 "video_transfer_linear_to_rec709", "video_transfer_linear_to_sRGB"):
        getattr(L, name).argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
    return L


FRAMEWORK_SYMBOLS = """init_half half_convert_from_float half_convert_to_float half_convert_from_float_fast
half_convert_to_float_fast half_lookup get_frame_time get_time_frame gettime video_get_frame_f16 video_get_frame_f32
video_copy_frame_f16 video_copy_frame_alpha_f32 video_mix_cross_f32_pull video_mix_cross_f32 video_mix_over_f32
video_scale_bilinear_f32 video_scale_bilinear_f32_pull video_color_rgb_to_xyz_sdtv video_color_xyz_to_srgb
video_transfer_rec709_to_linear_scene video_transfer_rec709_to_linear_display video_transfer_linear_to_rec709
video_transfer_linear_to_sRGB video_get_gamma45_ramp coded_image_alloc coded_image_alloc0 video_reconstruct_dv
video_subsample_dv workspace_create workspace_get_length workspace_add_item workspace_get_item workspace_remove_item
workspace_as_video_source workspace_free workspace_get_item_pos workspace_get_item_offset workspace_set_item_offset
workspace_get_item_source workspace_set_item_source workspace_get_item_tag workspace_set_item_tag workspace_update_item
filter_createTriangle filter_free""".split()


def test_framework_symbols_are_exported(lib):
    missing = [n for n in FRAMEWORK_SYMBOLS if not hasattr(lib, n)]
    assert not missing, missing


def test_framework_header_declares_exactly_these():
    from test_abi import declared_functions
    declared = declared_functions("framework.h") | declared_functions("half.h")
    assert declared <= set(FRAMEWORK_SYMBOLS), declared - set(FRAMEWORK_SYMBOLS)
    assert len(declared) >= 35


def test_triangle_taps_match_oracle(lib):
    for sub, offset in [(4.0, 0.0), (2.0, 0.0), (0.25, 0.0), (0.5, 0.0), (1.0, 0.0), (3.0, 0.25), (0.2, -0.5), (2.5, 0.5)]:
        want = (C.c_float * 64)()
        w, c = C.c_int(), C.c_int()
        assert O.oracle().orc_triangle(sub, offset, want, 64, C.byref(w), C.byref(c)) == 0
        f = FirFilter()
        lib.filter_createTriangle(sub, offset, C.byref(f))
        assert (f.width, f.center) == (w.value, c.value)
        assert [f.coeff[i] for i in range(f.width)] == [want[i] for i in range(w.value)]
        lib.filter_free(C.byref(f))
    # caller-supplied array that is too small: untouched coefficients, center -1, width = need
    small = (C.c_float * 2)(7.0, 7.0)
    f = FirFilter(C.cast(small, C.POINTER(C.c_float)), 2, 0)
    lib.filter_createTriangle(4.0, 0.0, C.byref(f))
    assert (f.width, f.center, small[0]) == (7, -1, 7.0)


def test_time_helpers(lib):
    for n, d in [(24000, 1001), (30000, 1001), (25, 1), (60, 1)]:
        rate = Rational(n, d)
        for frame in [0, 1, 2, 1000, 86399, -5]:
            t = lib.get_frame_time(C.byref(rate), frame)
            assert t == (frame * 1000000000 * d) // n + 1 if frame >= 0 else True
            if frame >= 0:
                assert lib.get_time_frame(C.byref(rate), t) == frame


def test_coded_image_alloc0(lib):
    strides = (C.c_int * 3)(720, 180, 0)
    lines = (C.c_int * 3)(480, 480, 480)
    im = lib.coded_image_alloc0(strides, lines, 3)
    assert im.contents.stride[0] == 720 and im.contents.line_count[1] == 480
    assert im.contents.data[2] is None and im.contents.data[3] is None
    plane = np.ctypeslib.as_array(C.cast(im.contents.data[0], C.POINTER(C.c_uint8)), shape=(480 * 720,))
    assert not plane.any()
    C.CFUNCTYPE(None, C.c_void_p)(im.contents.free_func)(im)


def test_workspace_bookkeeping(lib):
    ws = lib.workspace_create()
    tags = {}
    rng = random.Random(3)
    items = []
    for i in range(200):
        x, length, z, offset = rng.randrange(-50, 50), rng.randrange(1, 40), rng.randrange(-5, 5), rng.randrange(0, 9)
        it = lib.workspace_add_item(ws, None, x, length, offset, z, C.c_void_p(i + 1))
        items.append(it)
        tags[i + 1] = [x, length, z, offset]
    assert lib.workspace_get_length(ws) == 200

    # random updates / removals, as tests/process/video/VideoWorkspace.py does
    for _ in range(300):
        it = rng.choice(items)
        tag = lib.workspace_get_item_tag(it)
        x, z = C.c_int64(rng.randrange(-50, 50)), C.c_int64(rng.randrange(-5, 5))
        lib.workspace_update_item(it, C.byref(x), None, C.byref(z), None, None, None)
        tags[tag][0], tags[tag][2] = x.value, z.value
    for it in items[:50]:
        del tags[lib.workspace_get_item_tag(it)]
        lib.workspace_remove_item(it)
    assert lib.workspace_get_length(ws) == 150

    # workspace_get_item walks the items ordered by (x, z), like the reference's leftsort
    seen = []
    for i in range(150):
        it = lib.workspace_get_item(ws, i)
        x, w, z = C.c_int64(), C.c_int64(), C.c_int64()
        lib.workspace_get_item_pos(it, C.byref(x), C.byref(w), C.byref(z))
        tag = lib.workspace_get_item_tag(it)
        assert [x.value, w.value, z.value, lib.workspace_get_item_offset(it)] == tags[tag]
        seen.append((x.value, z.value))
    assert seen == sorted(seen)
    lib.workspace_free(ws)


def test_copy_frame_f16_is_a_plain_copy(lib):
    src = O.synth_frame((-4, -4, 60, 40), 8, (2, 2, 55, 20))
    want, got = O.HostFrame16((0, 0, 47, 31)), O.HostFrame16((0, 0, 47, 31))
    O.oracle().orc_copy_frame_f16(C.byref(want.c), C.byref(src.c))
    lib.video_copy_frame_f16(C.byref(got.c), C.byref(src.c))
    assert got.current.tuple() == want.current.tuple()
    assert np.array_equal(got.array, want.array)


def test_gamma45_ramp_pointer(lib):
    want = np.zeros(65536, np.uint8)
    O.oracle().orc_build_gamma45(O.ptr(want))
    assert np.array_equal(np.ctypeslib.as_array(lib.video_get_gamma45_ramp(), shape=(65536,)), want)


# ------------------------------------------------------------------------------------------ GPU

class FrameSource:
    """An f16 frame as a video_source, like RgbaFrameF16.c:27-38 (get_frame = video_copy_frame_f16)."""

    def __init__(self, lib, frame, f32_only=False):
        self.lib, self.frame = lib, frame

        def get16(obj, index, out):
            lib.video_copy_frame_f16(out, C.byref(self.frame.c))

        def get32(obj, index, out):
            # an f32-only source: widen on the host with the oracle (test scaffolding, not product)
            tmp = O.HostFrame16(O.Box(*out.contents.full.tuple()).tuple())
            O.oracle().orc_copy_frame_f16(C.byref(tmp.c), C.byref(self.frame.c))
            wide = O.Frame32(out.contents.data, out.contents.full, out.contents.current)
            O.oracle().orc_frame_f16_to_f32(C.byref(wide), C.byref(tmp.c))
            out.contents.current = wide.current

        self.cb16, self.cb32 = GET_FRAME(get16), GET_FRAME_32(get32)
        self.funcs = SourceFuncs(0, None if f32_only else C.cast(self.cb16, C.c_void_p), C.cast(self.cb32, C.c_void_p) if f32_only else None, None)
        self.source = VideoSource(None, C.pointer(self.funcs))


@pytest.mark.gpu
@pytest.mark.parametrize("f32_only", [(), (1,), (0, 2)])
def test_workspace_pulled_through_video_get_frame_f16(lib, f32_only):
    full = (0, 0, 95, 53)
    frames = [O.synth_frame(full, 40 + i, cw) for i, cw in enumerate([(0, 0, 95, 53), (10, 5, 60, 40), (31, 20, 95, 53), (3, 2, 40, 50)])]
    sources = [FrameSource(lib, f, i in f32_only) for i, f in enumerate(frames)]
    ws = lib.workspace_create()
    z = [3, 1, 2, 0]
    for s, zi in zip(sources, z):
        lib.workspace_add_item(ws, C.addressof(s.source), 10, 5, 0, zi, None)

    vs = VideoSource()
    lib.workspace_as_video_source(ws, C.byref(vs))

    order = [frames[i] for i in np.argsort(z)]
    want = O.HostFrame16(full)
    arr = (O.Frame16 * 4)(*[f.c for f in order])
    O.oracle().orc_composite_f16(C.byref(want.c), arr, 4)

    got = O.HostFrame16(full)
    lib.video_get_frame_f16(C.byref(vs), 12, C.byref(got.c))
    assert got.current.tuple() == want.current.tuple()
    assert np.array_equal(got.array, want.array)

    lib.video_get_frame_f16(C.byref(vs), 15, C.byref(got.c))       # x + length: nothing shows
    assert got.current.empty
    lib.video_get_frame_f16(C.byref(vs), 9, C.byref(got.c))
    assert got.current.empty
    lib.workspace_free(ws)


@pytest.mark.gpu
@pytest.mark.parametrize("mix", [1.0, 0.4, 0.0])
def test_host_mixes_match_oracle(lib, mix):
    from test_gpu_parity_misc import synth_f32
    full = (0, 0, 47, 31)
    for windows in [((0, 0, 47, 31), (0, 0, 47, 31)), ((0, 0, 30, 31), (10, 4, 47, 20)), ((0, 0, -1, -1), (3, 3, 30, 30))]:
        out, b = synth_f32(full, 1, windows[0]), synth_f32(full, 2, windows[1])
        want = out.copy()
        O.oracle().orc_mix_over_f32(C.byref(want.c), C.byref(b.c), mix)
        lib.video_mix_over_f32(C.byref(out.c), C.byref(b.c), mix)
        assert out.current.tuple() == want.current.tuple() or (out.current.empty and want.current.empty)
        if not want.current.empty:
            assert np.array_equal(out.window_view().view(np.uint32), want.window_view().view(np.uint32))

        a = synth_f32(full, 3, windows[0])
        want = O.HostFrame32(full)
        got = O.HostFrame32(full)
        O.oracle().orc_mix_cross_f32(C.byref(want.c), C.byref(a.c), C.byref(b.c), mix)
        lib.video_mix_cross_f32(C.byref(got.c), C.byref(a.c), C.byref(b.c), mix)
        assert got.current.tuple() == want.current.tuple() or (got.current.empty and want.current.empty)
        if not want.current.empty:
            assert np.array_equal(got.window_view().view(np.uint32), want.window_view().view(np.uint32))


@pytest.mark.gpu
def test_half_function_pointers_and_transfers(lib):
    lib.init_half()
    rng = np.random.default_rng(1)
    floats = rng.integers(0, 2 ** 32, size=4099, dtype=np.uint64).astype(np.uint32).view(np.float32)
    want = np.zeros(floats.shape, np.uint16)
    O.oracle().orc_f2h_n(O.ptr(want), O.ptr(floats), floats.size)
    fn = C.cast(C.c_void_p.in_dll(lib, "half_convert_from_float").value, C.CFUNCTYPE(None, C.c_void_p, C.c_void_p, C.c_int))
    got = np.zeros(floats.shape, np.uint16)
    fn(O.ptr(got), O.ptr(floats), floats.size)
    assert np.array_equal(got, want)

    halves = np.arange(65536, dtype=np.uint16)
    table = np.zeros(65536, np.uint16)
    O.oracle().orc_build_lut(O.LUT_LINEAR_TO_SRGB, O.ptr(table))
    out = np.zeros(65536, np.uint16)
    lib.video_transfer_linear_to_sRGB(O.ptr(out), O.ptr(halves), 65536)
    assert np.array_equal(out, table)


@pytest.mark.gpu
def test_dv_entry_points_match_oracle(lib):
    pic = O.synth_coded(720, 480, 9, 4, O.LAYOUT_PLANAR, O.MATRIX_709, (0, -1), 0, True)
    full = (0, -1, 719, 478)
    want = O.HostFrame16(full)
    O.oracle().orc_reconstruct(C.byref(want.c), C.byref(pic.c))

    image = O.RefCodedImage()
    for i, p in enumerate(pic.planes):
        image.data[i], image.stride[i], image.line_count[i] = p.ctypes.data, p.shape[1], p.shape[0]
    got = O.HostFrame16(full)
    lib.video_reconstruct_dv(C.byref(got.c), C.byref(image))
    assert got.current.tuple() == want.current.tuple()
    assert np.array_equal(got.array, want.array)

    # and back: f16 -> 4:1:1 planes, bit-exact, input frame transformed in place
    frame = O.synth_frame(full, 17)
    want_frame = frame.copy()
    want_pic = O.CodedPicture(720, 480, 4, O.LAYOUT_PLANAR, O.MATRIX_709, (0, -1))
    O.oracle().orc_subsample(C.byref(want_frame.c), C.byref(want_pic.c))
    planar = lib.video_subsample_dv(C.byref(frame.c))
    for i, (w, h) in enumerate([(720, 480), (180, 480), (180, 480)]):
        plane = np.ctypeslib.as_array(C.cast(planar.contents.data[i], C.POINTER(C.c_uint8)), shape=(h, w))
        assert np.array_equal(plane, want_pic.planes[i])
    assert np.array_equal(frame.array, want_frame.array)
    C.CFUNCTYPE(None, C.c_void_p)(planar.contents.free_func)(planar)
