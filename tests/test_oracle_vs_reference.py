"""Pins the oracle (oracle/oracle.c) against the reference's OWN C code compiled unmodified
(oracle/_ref/libcanvasref.so, built by `make -C oracle ref` from /root/reference).

CPU-only.  Skipped when the reference library has not been built (it is built in the
build container and travels to the GPU box as a prebuilt .so; /root/reference itself is
never read at test time).
"""
import ctypes as C

import numpy as np
import pytest

import oraclelib as ol
from oraclelib import Box, HostFrame16, HostFrame32, ptr

pytestmark = pytest.mark.skipif(not ol.have_reference(), reason="oracle/_ref/libcanvasref.so not built")


def test_half_convert_from_float_all_exponents():
    """half.c:47-51: table-driven truncation == the oracle's bit-op restatement."""
    O, R = ol.oracle(), ol.reference()
    rng = np.random.default_rng(1)
    bits = rng.integers(0, 2 ** 32, size=4_000_000, dtype=np.uint64).astype(np.uint32)
    # add every exponent boundary and the special values
    edges = []
    for e in range(256):
        for m in (0, 1, 0x1FFF, 0x2000, 0x3FFFFF, 0x400000, 0x7FFFFF):
            for s in (0, 1):
                edges.append((s << 31) | (e << 23) | m)
    bits = np.concatenate([bits, np.array(edges, dtype=np.uint32)])
    f = bits.view(np.float32)
    want = ol.ref_half_from_float(R, f)
    got = np.zeros(f.shape, np.uint16)
    O.orc_f2h_n(ptr(got), ptr(f), f.size)
    assert np.array_equal(got, want)


def test_half_convert_to_float_exhaustive():
    """half.c:31-37: all 65536 inputs."""
    O, R = ol.oracle(), ol.reference()
    h = np.arange(65536, dtype=np.uint16)
    want = ol.ref_half_to_float(R, h)
    got = np.zeros(65536, np.float32)
    O.orc_h2f_n(ptr(got), ptr(h), h.size)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))


@pytest.mark.parametrize("which,name", [
    (ol.LUT_709_TO_LINEAR_SCENE, "video_transfer_rec709_to_linear_scene"),
    (ol.LUT_709_TO_LINEAR_DISPLAY, "video_transfer_rec709_to_linear_display"),
    (ol.LUT_LINEAR_TO_709, "video_transfer_linear_to_rec709"),
    (ol.LUT_LINEAR_TO_SRGB, "video_transfer_linear_to_sRGB"),
])
def test_transfer_tables(which, name):
    """gammatab.c:83-250: every one of the 65536 table entries."""
    O, R = ol.oracle(), ol.reference()
    h = np.arange(65536, dtype=np.uint16)
    want = np.zeros(65536, np.uint16)
    getattr(R, name)(ptr(want), ptr(h), 65536)
    got = np.zeros(65536, np.uint16)
    O.orc_build_lut(which, ptr(got))
    assert np.array_equal(got, want)


def test_gamma45_ramp():
    """gammatab.c:8-38."""
    O, R = ol.oracle(), ol.reference()
    want = np.ctypeslib.as_array(R.video_get_gamma45_ramp(), shape=(65536,))
    got = np.zeros(65536, np.uint8)
    O.orc_build_gamma45(ptr(got))
    assert np.array_equal(got, want)


@pytest.mark.parametrize("sub,offset", [(4.0, 0.0), (2.0, 0.0), (0.25, 0.0), (0.5, 0.0), (1.0, 0.0), (3.0, 0.25), (0.3, -0.5)])
def test_triangle_filter(sub, offset):
    """filter.c:24-76."""
    O, R = ol.oracle(), ol.reference()

    class Fir(C.Structure):
        _fields_ = [("coeff", C.POINTER(C.c_float)), ("width", C.c_int), ("center", C.c_int)]

    fir = Fir()
    R.filter_createTriangle.argtypes = [C.c_float, C.c_float, C.POINTER(Fir)]
    R.filter_createTriangle(sub, offset, C.byref(fir))
    want = np.array([fir.coeff[i] for i in range(fir.width)], np.float32)

    got = np.zeros(64, np.float32)
    w, c = C.c_int(), C.c_int()
    assert O.orc_triangle(sub, offset, ptr(got), 64, C.byref(w), C.byref(c)) == 0
    assert (w.value, c.value) == (fir.width, fir.center)
    assert np.array_equal(got[:w.value].view(np.uint32), want.view(np.uint32))


def _dv_picture(seed, full_range=False):
    return ol.synth_coded(720, 480, seed, sub_x=4, pic=(0, -1), stride_pad=(seed % 3) * 8, full_range=full_range)


def _as_ref_coded(pic):
    im = ol.RefCodedImage()
    for i, p in enumerate(pic.planes):
        im.data[i], im.stride[i], im.line_count[i] = p.ctypes.data, p.shape[1], p.shape[0]
    return im


@pytest.mark.parametrize("full", [(0, -1, 719, 478), (0, 0, 719, 479), (-16, -8, 300, 200), (100, 50, 800, 600), (5, 7, 6, 7), (0, 500, 100, 600)])
@pytest.mark.parametrize("seed,full_range", [(1, False), (2, True)])
def test_reconstruct_dv(full, seed, full_range):
    """video_reconstruct.c:50-137 at its own hard-coded parameters, several caller windows.
    (A caller window wholly to the right/left of the picture makes the reference call
    rgba_f32_to_f16 with a negative count, :128, and run off the row: it segfaults, so that
    case has no reference answer; the oracle returns an empty current_window.)"""
    O, R = ol.oracle(), ol.reference()
    pic = _dv_picture(seed, full_range)
    want, got = HostFrame16(full), HostFrame16(full)
    want.array[:] = 0x1234
    got.array[:] = 0x1234
    im = _as_ref_coded(pic)
    R.video_reconstruct_dv(C.byref(want.c), C.byref(im))
    O.orc_reconstruct(C.byref(got.c), C.byref(pic.c))
    assert got.current.tuple() == want.current.tuple()
    if not want.current.empty:
        assert np.array_equal(got.array, want.array)


def test_reconstruct_dv_survey_known_answer():
    """pattern Y=16+((x+y)%220), Cb=16+((3x+y)%225), Cr=16+((5x+2y)%225)."""
    O = ol.oracle()
    pic = ol.CodedPicture(720, 480, sub_x=4, pic=(0, -1))
    yy, xx = np.mgrid[0:480, 0:720]
    cy, cx = np.mgrid[0:480, 0:180]
    pic.set_samples((16 + ((xx + yy) % 220)).astype(np.uint8), (16 + ((3 * cx + cy) % 225)).astype(np.uint8),
                    (16 + ((5 * cx + 2 * cy) % 225)).astype(np.uint8))
    got = HostFrame16((0, -1, 719, 478))
    O.orc_reconstruct(C.byref(got.c), C.byref(pic.c))
    assert got.current.tuple() == (0, -1, 719, 478)
    assert ol.fnv1a(got.array) == 0x7CCDB6B9
    assert [hex(v) for v in got.array[0, 0]] == ["0xb198", "0x2fd0", "0xb298", "0x3c00"]
    assert [hex(v) for v in got.array[0, 3]] == ["0xb150", "0x2ff9", "0xb25d", "0x3c00"]


@pytest.mark.parametrize("current", [(0, -1, 719, 478), (3, 5, 700, 400), (-5, -5, 900, 900), (101, 7, 333, 9)])
@pytest.mark.parametrize("seed", [3, 4])
def test_subsample_dv(current, seed):
    """video_subsample.c:99-187 at its own parameters: planes bit-exact, input mutated the same way."""
    O, R = ol.oracle(), ol.reference()
    full = (-8, -8, 1000, 1000) if current[0] < 0 or current[2] > 719 else (0, -1, 719, 478)
    src = ol.synth_frame(full, seed, current)
    # push some values out of [0,1] (super-whites and negatives) so truncation/wrap paths are hit
    src.array[::7, ::5, 0] = 0x3E00      # 1.5
    src.array[::11, ::3, 2] = 0xB400     # -0.25
    a, b = src.copy(), src.copy()

    planes = R.video_subsample_dv(C.byref(a.c)).contents
    want = [np.ctypeslib.as_array(C.cast(planes.data[i], C.POINTER(C.c_uint8)), shape=(planes.line_count[i], planes.stride[i])).copy()
            for i in range(3)]

    pic = ol.CodedPicture(720, 480, sub_x=4, pic=(0, -1))
    O.orc_subsample(C.byref(b.c), C.byref(pic.c))
    for i in range(3):
        assert np.array_equal(pic.planes[i], want[i]), f"plane {i}"
    assert np.array_equal(a.array, b.array)


def _rand_frame32(full, current, seed):
    rng = np.random.default_rng(seed)
    f = HostFrame32(full, current)
    h = rng.integers(0, 0x3C01, size=f.array.shape, dtype=np.uint16)        # halves in [0, 1]
    h[..., 3] = rng.choice(np.array([0, 0x3400, 0x3800, 0x3A00, 0x3C00, 0x2E66], np.uint16), size=h.shape[:2])
    ol.oracle().orc_h2f_n(ptr(f.array), ptr(h), h.size)
    return f


FULL = (-4, -2, 59, 37)
# Windows for which the reference's min.x-vs-min.y comparison (video_mix.c:137,265) and its
# disjoint-window gap handling (:161-169) are neutral, so intended == reference semantics.
NEUTRAL_PAIRS = [
    ((-4, -2, 59, 37), (-4, -2, 59, 37)),
    ((-4, -2, 59, 37), (10, 5, 40, 30)),
    ((-4, -2, 59, 37), (10, 5, 80, 60)),      # b sticks out of full on the right/bottom
    ((0, 0, 30, 20), (0, 10, 50, 37)),         # equal min.x, overlap
    ((0, 0, 30, 20), (0, 21, 30, 37)),         # equal x-range, vertically adjacent (no gap)
    ((5, 5, 40, 30), (5, 5, 40, 30)),
]


@pytest.mark.parametrize("awin,bwin", NEUTRAL_PAIRS)
@pytest.mark.parametrize("mix", [0.25, 0.0, 1.0, 0.6180339, 1.5, -0.5])
def test_mix_over_f32(awin, bwin, mix):
    """video_mix.c:237-370."""
    O, R = ol.oracle(), ol.reference()
    bwin_c = Box(*bwin)
    inner = (max(bwin[0], FULL[0]), max(bwin[1], FULL[1]), min(bwin[2], FULL[2]), min(bwin[3], FULL[3]))
    base = _rand_frame32(FULL, awin, 10)
    top = _rand_frame32(FULL, inner, 11)
    want, got = base.copy(), base.copy()
    R.video_mix_over_f32(C.byref(want.c), C.byref(top.c), mix)
    O.orc_mix_over_f32(C.byref(got.c), C.byref(top.c), mix)
    assert got.current.tuple() == want.current.tuple()
    assert np.array_equal(got.window_view().view(np.uint32), want.window_view().view(np.uint32))
    del bwin_c


@pytest.mark.parametrize("awin,bwin", NEUTRAL_PAIRS + [((0, 0, -1, -1), (3, 3, 20, 20)), ((3, 3, 20, 20), (0, 0, -1, -1))])
@pytest.mark.parametrize("mix", [0.25, 0.0, 1.0, 0.6180339, 2.0])
def test_mix_cross_f32(awin, bwin, mix):
    """video_mix.c:107-235.  The left-strip source pick at :137 compares a.min.x with b.min.y, so the
    pairs here keep that comparison equal to the intended a.min.x < b.min.x."""
    O, R = ol.oracle(), ol.reference()
    clip = lambda w: (max(w[0], FULL[0]), max(w[1], FULL[1]), min(w[2], FULL[2]), min(w[3], FULL[3]))
    a = _rand_frame32(FULL, clip(awin), 20)
    b = _rand_frame32(FULL, clip(bwin), 21)
    want, got = HostFrame32(FULL), HostFrame32(FULL)
    R.video_mix_cross_f32(C.byref(want.c), C.byref(a.c), C.byref(b.c), mix)
    O.orc_mix_cross_f32(C.byref(got.c), C.byref(a.c), C.byref(b.c), mix)
    assert got.current.tuple() == want.current.tuple()
    assert np.array_equal(got.window_view().view(np.uint32), want.window_view().view(np.uint32))


def test_reference_window_bugs_are_where_we_say():
    """Documents the two reference bugs: the oracle's intended semantics differ
    from the compiled reference exactly on the pixels the reference leaves stale / mis-picks."""
    O, R = ol.oracle(), ol.reference()
    full = (0, 0, 39, 29)
    # left strip: out cw (5,0)-(20,20), b cw (2,10)-(25,15): row 12, x=2..4 should take b
    base = _rand_frame32(full, (5, 0, 20, 20), 30)
    top = _rand_frame32(full, (2, 10, 25, 15), 31)
    want, got = base.copy(), base.copy()
    R.video_mix_over_f32(C.byref(want.c), C.byref(top.c), 1.0)
    O.orc_mix_over_f32(C.byref(got.c), C.byref(top.c), 1.0)
    assert got.current.tuple() == want.current.tuple() == (2, 0, 25, 20)
    diff = np.any(got.array.view(np.uint32) != want.array.view(np.uint32), axis=2)
    ys, xs = np.nonzero(diff)
    assert set(xs.tolist()) <= {2, 3, 4} and set(ys.tolist()) <= set(range(10, 16))
    assert np.array_equal(got.array[12, 2:5], top.array[12, 2:5])


@pytest.mark.parametrize("count", [1, 2, 4, 8])
def test_workspace_composite_f16(count):
    """workspace.c:494-550 via video_get_frame_f16 (main.c:34-77), layers given out of z order."""
    O, R = ol.oracle(), ol.reference()
    full = (0, 0, 63, 47)
    wins = [(0, 0, 63, 47), (8, 6, 50, 40), (20, 10, 90, 30), (4, 20, 40, 60), (0, 0, 63, 47), (30, 5, 33, 44), (1, 1, 62, 46), (16, 16, 47, 31)]
    layers = [ol.synth_frame((-2, -3, 95, 63), 40 + i, wins[i]) for i in range(count)]
    # bottom -> top order for the oracle; the reference sorts by z itself
    order = list(range(count))
    z = [10 * i + 3 for i in order]
    perm = list(reversed(order))
    arr = (ol.Frame16 * count)(*[layers[i].c for i in perm])
    zs = (C.c_int64 * count)(*[z[i] for i in perm])
    want, got = HostFrame16(full), HostFrame16(full)
    R.ref_workspace_composite_f16(arr, zs, count, C.byref(want.c))
    arr2 = (ol.Frame16 * count)(*[layers[i].c for i in order])
    O.orc_composite_f16(C.byref(got.c), arr2, count)
    assert got.current.tuple() == want.current.tuple()
    assert np.array_equal(got.window_view(), want.window_view())


def test_workspace_composite_empty():
    O, R = ol.oracle(), ol.reference()
    want, got = HostFrame16((0, 0, 7, 7)), HostFrame16((0, 0, 7, 7))
    R.ref_workspace_composite_f16(None, None, 0, C.byref(want.c))
    O.orc_composite_f16(C.byref(got.c), None, 0)
    assert want.current.empty and got.current.empty


@pytest.mark.parametrize("mix", [0.0, 0.25, 0.5, 1.0, 0.999])
def test_cross_pull_f16(mix):
    """video_mix_cross_f32_pull (video_mix.c:46-71) pulled as f16."""
    O, R = ol.oracle(), ol.reference()
    full = (0, 0, 63, 47)
    a = ol.synth_frame(full, 50)
    b = ol.synth_frame(full, 51, (0, 8, 63, 40))
    want, got = HostFrame16(full), HostFrame16(full)
    R.ref_cross_pull_f16(C.byref(a.c), C.byref(b.c), mix, C.byref(want.c))
    O.orc_cross_f16(C.byref(got.c), C.byref(a.c), C.byref(b.c), mix)
    assert got.current.tuple() == want.current.tuple()
    assert np.array_equal(got.window_view(), want.window_view())


def test_encode_u8():
    """widget_gl.c:290-307 with the reference's sRGB table."""
    O, R = ol.oracle(), ol.reference()
    full = (-3, -1, 60, 30)
    src = ol.synth_frame(full, 60, (0, 0, 55, 30))
    src.array[3, 3] = (0x7C00, 0xFC00, 0x7E00, 0x0001)
    ramp = np.zeros(65536, np.uint8)
    O.orc_build_intent_ramp(1.25, ptr(ramp))
    lut = np.zeros(65536, np.uint16)
    O.orc_build_lut(ol.LUT_LINEAR_TO_SRGB, ptr(lut))
    a, b = src.copy(), src.copy()
    want = np.zeros((full[3] - full[1] + 1, full[2] - full[0] + 1, 4), np.uint8)
    got = want.copy()
    R.ref_encode_u8(C.byref(a.c), ptr(ramp), ptr(want))
    O.orc_encode_u8(C.byref(b.c), ptr(lut), ptr(ramp), ptr(got))
    assert np.array_equal(got, want)
    assert np.array_equal(a.array, b.array)


# ------------------------------------------------------------------ scaler (video_scale.c:231-285)

SCALE_CASES = [
    # (fx, fy, target_point, source_point, source window, target full)
    (2.0, 2.0, (0.0, 0.0), (0.0, 0.0), (0, 0, 39, 23), (0, 0, 79, 47)),
    (0.5, 0.5, (0.0, 0.0), (0.0, 0.0), (0, 0, 79, 47), (0, 0, 39, 23)),
    (1.0, 3.0, (2.0, 1.5), (2.0, -0.25), (3, 2, 50, 20), (-4, -6, 60, 70)),        # vertical only, fractional points
    (0.37, 1.0, (5.0, 0.0), (11.5, 0.0), (-10, 0, 90, 12), (0, -3, 44, 20)),       # horizontal only, downsample
    (1.7, 0.6, (10.0, 10.0), (20.25, 15.5), (0, 0, 63, 63), (-20, -20, 100, 60)),  # mixed: vertical (smaller) first
    (0.25, 3.5, (0.0, 0.0), (0.0, 0.0), (8, 4, 71, 19), (0, 0, 30, 80)),           # horizontal (smaller) first
    (1.0, 1.0, (3.0, 4.0), (3.0, 4.0), (2, 2, 20, 20), (0, 0, 15, 30)),            # identity: plain windowed copy
    (1.0, 1.0, (3.5, 4.0), (3.0, 4.0), (2, 2, 20, 20), (0, 0, 25, 30)),            # factor 1 with a sub-pixel shift
    (4.0, 4.0, (0.0, 0.0), (0.0, 0.0), (5, 5, 9, 9), (100, 100, 120, 120)),        # target window misses the image
    (1.0 / 16.0, 1.0 / 3.0, (0.0, 0.0), (0.0, 0.0), (0, 0, 255, 59), (0, 0, 20, 25)),
]


@pytest.mark.parametrize("case", SCALE_CASES)
def test_scale_bilinear_f32(case):
    """video_scale_bilinear_f32: separable triangle filter, taps from filter_createTriangle, both pass orders."""
    fx, fy, tp, sp, swin, tfull = case
    O, R = ol.oracle(), ol.reference()
    rng = np.random.default_rng(int(fx * 1000 + fy * 10))
    sfull = (swin[0] - 3, swin[1] - 2, swin[2] + 4, swin[3] + 1)
    src = HostFrame32(sfull, swin, rng.uniform(-0.5, 1.5, size=(sfull[3] - sfull[1] + 1, sfull[2] - sfull[0] + 1, 4)).astype(np.float32))

    want = HostFrame32(tfull)
    want.array[...] = 7.0                       # the scaler zero-fills the whole target before it accumulates
    R.video_scale_bilinear_f32(C.byref(want.c), ol.V2f(*tp), C.byref(src.c), ol.V2f(*sp), ol.V2f(fx, fy))
    got = HostFrame32(tfull)
    got.array[...] = 9.0
    O.orc_scale_bilinear_f32(C.byref(got.c), tp[0], tp[1], C.byref(src.c), sp[0], sp[1], fx, fy)

    if want.current.empty:
        assert got.current.empty
    else:
        assert got.current.tuple() == want.current.tuple()
    if fx == 1.0 and fy == 1.0 and tp == sp:
        w, g = want.window_view(), got.window_view()        # a plain copy: pixels outside the window are untouched
    else:
        w, g = want.array, got.array
    assert np.array_equal(g.view(np.uint32), w.view(np.uint32))
