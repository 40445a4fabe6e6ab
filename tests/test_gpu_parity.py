"""GPU parity: the sm_100a path (through the C-ABI, libcanvas_b200.so) against the CPU oracle
(oracle/liboracle.so, itself pinned bit-for-bit against the reference's own C code).

Bar: bit-exact for 8-bit / integer outputs and window boxes; f16 RGBA is ASSERTED bit-exact too
(the contract in BASELINE.json is <= 1 half ulp, max abs error <= 2^-10 on [0,1]; the ulp report
is printed on failure).
"""
import ctypes as C
import os

import numpy as np
import pytest

import oraclelib as O

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev():
    from canvas_b200 import device
    device.init(0)
    return device


def half_ulp_distance(a, b):
    """Distance in representable halves between raw half bit patterns (sign-magnitude aware)."""
    def key(x):
        x = x.astype(np.int32)
        return np.where(x & 0x8000, -(x & 0x7FFF), x & 0x7FFF)
    return np.abs(key(a) - key(b))


def assert_frames_equal(gpu_array, gpu_window, host_frame, what):
    gx0, gy0, gx1, gy1 = tuple(gpu_window)
    if host_frame.current.empty:
        # "empty <=> max < min" (framework.h:80-82); the reference does not canonicalise empty boxes
        assert gx1 < gx0 or gy1 < gy0, "%s: window %r, oracle says empty" % (what, tuple(gpu_window))
        return
    assert tuple(gpu_window) == host_frame.current.tuple(), "%s: window %r != oracle %r" % (what, tuple(gpu_window), host_frame.current)
    b = host_frame.current
    f = host_frame.full
    ys = slice(b.min_y - f.min_y, b.max_y - f.min_y + 1)
    xs = slice(b.min_x - f.min_x, b.max_x - f.min_x + 1)
    got, want = gpu_array[ys, xs], host_frame.array[ys, xs]
    if not np.array_equal(got, want):
        d = half_ulp_distance(got, want)
        bad = np.argwhere(d > 0)
        y, x, c = bad[0]
        raise AssertionError("%s: %d of %d halves differ, max %d ulp; first at (x=%d,y=%d,c=%d): gpu %04x oracle %04x"
                             % (what, len(bad), got.size, d.max(), x + b.min_x, y + b.min_y, c, got[y, x, c], want[y, x, c]))


def to_device(dev, hf):
    return dev.DeviceFrame16.from_numpy(hf.array, hf.full.tuple(), hf.current.tuple())


def poisoned(dev, full):
    f = dev.DeviceFrame16(full)
    f.fill(0x5A)
    return f


# ------------------------------------------------------------------ composite

COMPOSITE_CASES = {
    # name: (full, [layer (full, current)])
    "two_full": ((0, 0, 63, 47), [((0, 0, 63, 47), (0, 0, 63, 47))] * 2),
    "eight_full": ((0, 0, 127, 63), [((0, 0, 127, 63), (0, 0, 127, 63))] * 8),
    "single": ((0, 0, 31, 15), [((0, 0, 31, 15), (2, 3, 20, 9))]),
    "partial_windows": ((0, 0, 95, 63), [((0, 0, 95, 63), (0, 0, 95, 63)), ((0, 0, 95, 63), (10, 5, 60, 40)),
                                         ((0, 0, 95, 63), (31, 20, 95, 63)), ((0, 0, 95, 63), (3, 50, 40, 60))]),
    "odd_origins": ((-3, -1, 60, 40), [((-3, -1, 60, 40), (-3, -1, 60, 40)), ((-7, -5, 71, 51), (-5, 0, 33, 17)),
                                       ((1, 2, 50, 30), (1, 2, 50, 30))]),
    "disjoint": ((0, 0, 63, 63), [((0, 0, 63, 63), (0, 0, 9, 9)), ((0, 0, 63, 63), (20, 20, 29, 29)),
                                  ((0, 0, 63, 63), (5, 5, 24, 24))]),
    "outside_full": ((0, 0, 39, 29), [((0, 0, 39, 29), (0, 0, 39, 29)), ((-20, -10, 59, 49), (-20, -10, 10, 12)),
                                      ((30, 20, 90, 70), (30, 20, 90, 70)), ((100, 100, 120, 120), (100, 100, 120, 120))]),
    "empty_bottom": ((0, 0, 31, 31), [((0, 0, 31, 31), (0, 0, -1, -1)), ((0, 0, 31, 31), (4, 4, 20, 20)),
                                      ((0, 0, 31, 31), (8, 1, 30, 12))]),
    "odd_width": ((0, 0, 44, 20), [((0, 0, 44, 20), (0, 0, 44, 20)), ((0, 0, 44, 20), (7, 3, 33, 18))]),
    "sixteen": ((0, 0, 63, 31), [((0, 0, 63, 31), (i, i % 5, 63 - i, 31 - (i % 3))) for i in range(16)]),
    # rows of several 192-pixel (3 pairs per lane) / 128-pixel segments: window edges inside, on and next to segment
    # boundaries; five ring slots take the 3-pair window-classifying kernels, nine the 2-pair ones
    "wide_windows_5": ((0, 0, 1001, 11), [((0, 0, 1001, 11), (0, 0, 1001, 11)), ((0, 0, 1001, 11), (191, 1, 576, 9)),
                                          ((0, 0, 1001, 11), (192, 0, 383, 11)), ((-6, -2, 1100, 14), (385, 3, 1001, 7)),
                                          ((0, 0, 1001, 11), (7, 2, 130, 10))]),
    "wide_windows_9": ((-2, 0, 899, 8), [((-2, 0, 899, 8), (-2, 0, 899, 8))] +
                                        [((-2, 0, 899, 8), (-2 + 97 * i, i % 3, 899 - 61 * i, 8 - (i % 2))) for i in range(1, 9)]),
}


@pytest.mark.parametrize("name", sorted(COMPOSITE_CASES))
def test_composite_over_matches_oracle(dev, name):
    full, specs = COMPOSITE_CASES[name]
    layers = [O.synth_frame(lf, 100 + i, cw) for i, (lf, cw) in enumerate(specs)]

    want = O.HostFrame16(full)
    want.array[:] = 0x5A5A
    arr = (O.Frame16 * len(layers))(*[l.c for l in layers])
    O.oracle().orc_composite_f16(C.byref(want.c), arr, len(layers))

    dlayers = [to_device(dev, l) for l in layers]
    out = poisoned(dev, full)
    dev.composite_over(out, dlayers)
    assert_frames_equal(out.download(), out.current.tuple(), want, "composite/" + name)

    # pixels outside the result window must not have been written
    got = out.download()
    mask = np.ones(got.shape[:2], bool)
    if not want.current.empty:
        b, f = want.current, want.full
        mask[b.min_y - f.min_y: b.max_y - f.min_y + 1, b.min_x - f.min_x: b.max_x - f.min_x + 1] = False
    assert np.all(got[mask] == 0x5A5A), "composite/%s wrote outside current_window" % name


def test_composite_empty_inputs(dev):
    out = poisoned(dev, (0, 0, 15, 15))
    dev.composite_over(out, [])
    assert out.current.empty
    layer = to_device(dev, O.synth_frame((0, 0, 15, 15), 3, (0, 0, -1, -1)))
    dev.composite_over(out, [layer, layer])
    assert out.current.empty


def test_composite_deep_stack_chains_without_extra_rounding(dev):
    full = (0, 0, 47, 31)
    layers = [O.synth_frame(full, 500 + i, (i % 7, i % 5, 47 - (i % 4), 31 - (i % 6))) for i in range(37)]
    want = O.HostFrame16(full)
    arr = (O.Frame16 * len(layers))(*[l.c for l in layers])
    O.oracle().orc_composite_f16(C.byref(want.c), arr, len(layers))
    out = poisoned(dev, full)
    dev.composite_over(out, [to_device(dev, l) for l in layers])
    assert_frames_equal(out.download(), out.current.tuple(), want, "composite/deep")


# ------------------------------------------------------------------ crossfade

@pytest.mark.parametrize("mix", [0.0, 1.0, 0.25, 0.5, 0.999, 1e-3, -0.5, 1.5, float("nan")])
@pytest.mark.parametrize("windows", [((0, 0, 63, 35), (0, 0, 63, 35)), ((0, 0, 40, 35), (10, 4, 63, 20)),
                                     ((0, 0, 9, 9), (20, 20, 29, 29)), ((0, 0, -1, -1), (3, 3, 30, 30)),
                                     ((5, 0, 20, 20), (2, 10, 25, 15))])
def test_cross_matches_oracle(dev, mix, windows):
    full = (0, 0, 63, 35)
    a = O.synth_frame(full, 11, windows[0])
    b = O.synth_frame(full, 12, windows[1])
    want = O.HostFrame16(full)
    O.oracle().orc_cross_f16(C.byref(want.c), C.byref(a.c), C.byref(b.c), mix)

    out = poisoned(dev, full)
    dev.mix_cross(out, to_device(dev, a), to_device(dev, b), mix)
    assert_frames_equal(out.download(), out.current.tuple(), want, "cross mix=%r windows=%r" % (mix, windows))


TINY_MIXES = [1e-20, 1e-38, 1.1754944e-38, 1e-42, 5.0e-8, 1.0 - 5.96e-8]      # ... FLT_MIN, a subnormal, just under 2^-24, 1 - 2^-24


@pytest.mark.parametrize("mix", TINY_MIXES)
def test_tiny_mixes_keep_the_ieee_division(dev, mix):
    """A positive mix below 2^-24 makes alpha * mix smaller than anything the shared-reciprocal division of the fast path
    is exact for: such jobs must take the IEEE-division path.  The crossfade (video_mix.c:192-206) with the
    weight on either side, and a layer whose alpha is scaled by it (video_mix.c:323-337)."""
    full = (0, 0, 63, 35)
    a, b = O.synth_frame(full, 21, full), O.synth_frame(full, 22, (8, 2, 60, 30))
    for m in (mix, 1.0 - mix if mix < 0.5 else mix):
        want = O.HostFrame16(full)
        O.oracle().orc_cross_f16(C.byref(want.c), C.byref(a.c), C.byref(b.c), m)
        out = poisoned(dev, full)
        dev.mix_cross(out, to_device(dev, a), to_device(dev, b), m)
        assert_frames_equal(out.download(), out.current.tuple(), want, "cross mix=%r" % m)

    # a layer over a bottom with its alpha scaled by the tiny mix: the f32 frame operations are the oracle here
    bottom32, top32, acc = O.HostFrame32(full), O.HostFrame32(full), O.HostFrame32(full)
    O.oracle().orc_frame_f16_to_f32(C.byref(bottom32.c), C.byref(a.c))
    O.oracle().orc_frame_f16_to_f32(C.byref(top32.c), C.byref(b.c))
    O.oracle().orc_copy_frame_alpha_f32(C.byref(acc.c), C.byref(bottom32.c), 1.0)
    O.oracle().orc_mix_over_f32(C.byref(acc.c), C.byref(top32.c), mix)
    want = O.HostFrame16(full)
    O.oracle().orc_frame_f32_to_f16(C.byref(want.c), C.byref(acc.c))
    out = poisoned(dev, full)
    dev.composite_over(out, [to_device(dev, a), (to_device(dev, b), mix)])
    dev.sync()
    assert_frames_equal(out.download(), out.current.tuple(), want, "layer mix=%r" % mix)


# ------------------------------------------------------------------ conversion

def device_coded(dev, pic, zero=False):
    return dev.DeviceCoded(pic.planes, pic.width, pic.height, pic.sub_x, pic.layout, pic.matrix, pic.pic, zero=zero)


CONVERT_CASES = [
    # width, height, sub_x, layout, matrix, pic, full, stride_pad, full_range
    (720, 480, 4, O.LAYOUT_PLANAR, O.MATRIX_709, (0, -1), (0, -1, 719, 478), 0, False),      # the reference's own DV case
    (720, 480, 4, O.LAYOUT_PLANAR, O.MATRIX_709, (0, -1), (-10, -10, 400, 200), 16, True),
    (96, 40, 2, O.LAYOUT_PLANAR, O.MATRIX_709, (0, 0), (0, 0, 95, 39), 0, True),
    (96, 40, 2, O.LAYOUT_UYVY, O.MATRIX_709, (0, 0), (0, 0, 95, 39), 0, True),
    (96, 40, 2, O.LAYOUT_YUY2, O.MATRIX_601, (0, 0), (0, 0, 95, 39), 8, True),
    (97, 33, 2, O.LAYOUT_PLANAR, O.MATRIX_601, (3, -2), (0, 0, 120, 40), 3, True),           # odd width, odd offset
    (97, 33, 2, O.LAYOUT_UYVY, O.MATRIX_709, (-5, 1), (0, 0, 63, 63), 0, True),
    (50, 20, 1, O.LAYOUT_PLANAR, O.MATRIX_709, (1, 1), (0, 0, 63, 31), 0, True),             # 4:4:4
    (101, 17, 4, O.LAYOUT_PLANAR, O.MATRIX_601, (2, 0), (0, 0, 127, 31), 1, True),
    (64, 16, 2, O.LAYOUT_PLANAR, O.MATRIX_709, (200, 200), (0, 0, 63, 31), 0, False),        # picture outside the frame
]


@pytest.mark.parametrize("case", CONVERT_CASES, ids=lambda c: "%dx%d_s%d_l%d_m%d" % c[:5])
def test_convert_matches_oracle(dev, case):
    w, h, sub, layout, matrix, pic, full, pad, full_range = case
    picture = O.synth_coded(w, h, 7, sub, layout, matrix, pic, pad, full_range)
    want = O.HostFrame16(full)
    O.oracle().orc_reconstruct(C.byref(want.c), C.byref(picture.c))

    out = poisoned(dev, full)
    dev.convert_ycbcr8(out, device_coded(dev, picture))
    assert_frames_equal(out.download(), out.current.tuple(), want, "convert %r" % (case,))


def test_convert_all_byte_values(dev):
    """Every (Y', C) byte pair appears: 256x256 picture with Y = x, Cb = y, Cr = 255 - y (4:4:4)."""
    pic = O.CodedPicture(256, 256, 1, O.LAYOUT_PLANAR, O.MATRIX_709)
    yy, xx = np.mgrid[0:256, 0:256].astype(np.uint8)
    pic.set_samples(xx, yy, (255 - yy).astype(np.uint8))
    want = O.HostFrame16((0, 0, 255, 255))
    O.oracle().orc_reconstruct(C.byref(want.c), C.byref(pic.c))
    out = poisoned(dev, (0, 0, 255, 255))
    dev.convert_ycbcr8(out, device_coded(dev, pic))
    assert_frames_equal(out.download(), out.current.tuple(), want, "convert/all-bytes")


# ------------------------------------------------------------------ fused timeline frame

@pytest.mark.parametrize("mix", [0.0, 0.3, 1.0])
@pytest.mark.parametrize("layout", [O.LAYOUT_PLANAR, O.LAYOUT_UYVY])
def test_timeline_frame_matches_oracle(dev, mix, layout):
    full = (0, 0, 127, 71)
    pa = O.synth_coded(128, 72, 21, 2, layout, O.MATRIX_709, (0, 0), 0, True)
    pb = O.synth_coded(120, 60, 22, 2, layout, O.MATRIX_601, (4, 6), 0, True)
    overlays = [O.synth_frame(full, 31, (10, 10, 90, 50)), O.synth_frame((-8, -8, 140, 80), 32, (-8, 30, 140, 80)),
                O.synth_frame(full, 33, (100, 0, 127, 71))]

    want = O.HostFrame16(full)
    arr = (O.Frame16 * len(overlays))(*[o.c for o in overlays])
    O.oracle().orc_timeline_frame_f16(C.byref(want.c), C.byref(pa.c), C.byref(pb.c), mix, arr, len(overlays))

    out = poisoned(dev, full)
    job = dev.make_job(out=out, a=device_coded(dev, pa), b=device_coded(dev, pb), mix_b=mix,
                       layers=[to_device(dev, o) for o in overlays])
    dev.render_frames([job])
    dev.sync()
    assert_frames_equal(out.download(), out.current.tuple(), want, "timeline mix=%r layout=%d" % (mix, layout))


@pytest.mark.parametrize("matrix", [O.MATRIX_709, O.MATRIX_601])
@pytest.mark.parametrize("layout", [O.LAYOUT_PLANAR, O.LAYOUT_UYVY, O.LAYOUT_YUY2])
def test_corner_bytes_through_the_layered_kernels(dev, matrix, layout):
    """Only bytes 0, 1, 254, 255 in every plane: the extremes of the converted range (past 2.0 and below -1.0), where the
    64 KiB compact conversion table of the layered kernels ends.  Uniform and windowed forms, plus the bare conversion."""
    w, h = 256, 24
    full = (0, 0, w - 1, h - 1)
    rng = np.random.default_rng(5 + matrix * 3 + layout)
    pick = np.array([0, 1, 254, 255], np.uint8)
    pic = O.CodedPicture(w, h, 2, layout, matrix, (0, 0))
    pic.set_samples(pick[rng.integers(0, 4, (h, w))], pick[rng.integers(0, 4, (h, w // 2))], pick[rng.integers(0, 4, (h, w // 2))])
    dpic = device_coded(dev, pic)

    for boxes in ([full, full], [(20, 2, 201, 20), (0, 0, 99, 23)]):
        overlays = [O.synth_frame(full, 61 + i, b) for i, b in enumerate(boxes)]
        want = O.HostFrame16(full)
        arr = (O.Frame16 * 2)(*[o.c for o in overlays])
        O.oracle().orc_timeline_frame_f16(C.byref(want.c), C.byref(pic.c), None, 0.0, arr, 2)
        out = poisoned(dev, full)
        dev.render_frames([dev.make_job(out=out, a=dpic, layers=[to_device(dev, o) for o in overlays])])
        dev.sync()
        assert_frames_equal(out.download(), out.current.tuple(), want, "corner bytes under %r" % (boxes,))

    want = O.HostFrame16(full)
    O.oracle().orc_reconstruct(C.byref(want.c), C.byref(pic.c))
    out = poisoned(dev, full)
    dev.convert_ycbcr8(out, dpic)
    dev.sync()
    assert_frames_equal(out.download(), out.current.tuple(), want, "corner bytes, bare conversion")
    values = want.array[..., :3].view(np.float16)
    assert values.max() > 1.0 and values.min() < 0.0


def test_render_frames_batch_equals_single(dev):
    """A batch (grid-walked jobs from device memory) gives the same pixels as one job per launch."""
    full = (0, 0, 95, 53)
    pics = [O.synth_coded(96, 54, 40 + i, 2, O.LAYOUT_UYVY, O.MATRIX_709, (0, 0), 0, True) for i in range(3)]
    overlays = [O.synth_frame(full, 60 + i, (i * 3, i, 95 - i, 53 - 2 * i)) for i in range(4)]
    dpics = [device_coded(dev, p) for p in pics]
    dover = [to_device(dev, o) for o in overlays]

    jobs, outs = [], []
    for f in range(9):
        out = poisoned(dev, full)
        outs.append(out)
        if f % 3 == 0:
            jobs.append(dev.make_job(out=out, layers=[dover[f % 4], dover[(f + 1) % 4]]))
        elif f % 3 == 1:
            jobs.append(dev.make_job(out=out, a=dpics[f % 3], b=dpics[(f + 1) % 3], mix_b=f / 9.0, layers=[dover[f % 4]]))
        else:
            jobs.append(dev.make_job(out=out, a=dover[0], b=dover[1], mix_b=0.5, layers=[dover[2]]))
    dev.render_frames(jobs)
    dev.sync()
    batch = [(o.download(), o.current.tuple()) for o in outs]

    for f, job in enumerate(jobs):
        outs[f].fill(0x5A)
        dev.render_frames([job])
        dev.sync()
        single = outs[f].download()
        assert batch[f][1] == outs[f].current.tuple()
        assert np.array_equal(batch[f][0], single), "job %d differs between batch and single launch" % f


# ------------------------------------------------------------------ values outside the fast path's range

def assert_equal_up_to_nan_payload(got, want, what):
    """Bit-exact, except that any NaN matches any NaN (x86 propagates payloads, the GPU canonicalises)."""
    nan_g = ((got & 0x7C00) == 0x7C00) & ((got & 0x03FF) != 0)
    nan_w = ((want & 0x7C00) == 0x7C00) & ((want & 0x03FF) != 0)
    same = (got == want) | (nan_g & nan_w)
    if not same.all():
        bad = np.argwhere(~same)
        y, x, c = bad[0]
        raise AssertionError("%s: %d halves differ; first at (%d,%d,%d): gpu %04x oracle %04x"
                             % (what, len(bad), x, y, c, got[y, x, c], want[y, x, c]))


@pytest.mark.parametrize("kind", ["negative_and_zero_signs", "alpha_above_one", "inf_and_huge", "any_bits"])
def test_composite_unusual_values_take_the_exact_path(dev, kind):
    full = (0, 0, 127, 31)
    rng = np.random.default_rng(hash(kind) % 1000)
    layers = []
    for i in range(5):
        f = O.synth_frame(full, 900 + i)
        a = f.array
        pick = rng.random(a.shape[:2]) < 0.3
        if kind == "negative_and_zero_signs":
            a[pick, 0] |= 0x8000                                  # negative red
            a[rng.random(a.shape[:2]) < 0.2, 1] = 0x8000          # -0.0 green
            a[rng.random(a.shape[:2]) < 0.2, 3] = 0x3C00
        elif kind == "alpha_above_one":
            a[pick, 3] = rng.integers(0x3C01, 0x4400, size=int(pick.sum()), dtype=np.uint16)
            a[rng.random(a.shape[:2]) < 0.1, 3] = 0xB800          # alpha -0.5
        elif kind == "inf_and_huge":
            a[pick, 2] = 0x7C00
            a[rng.random(a.shape[:2]) < 0.1, 0] = 0x7BFF
            a[rng.random(a.shape[:2]) < 0.1, 3] = 0x7C00
        else:
            a[pick] = rng.integers(0, 65536, size=(int(pick.sum()), 4), dtype=np.uint16)
        layers.append(f)

    want = O.HostFrame16(full)
    arr = (O.Frame16 * len(layers))(*[l.c for l in layers])
    O.oracle().orc_composite_f16(C.byref(want.c), arr, len(layers))
    out = poisoned(dev, full)
    dev.composite_over(out, [to_device(dev, l) for l in layers])
    assert out.current.tuple() == want.current.tuple()
    assert_equal_up_to_nan_payload(out.download(), want.array, "composite/" + kind)


def test_composite_large_random_stack_bit_exact(dev):
    """8 full 1080p layers: ~16.6 M divisions per layer through the shared-reciprocal fast path."""
    full = (0, 0, 1919, 1079)
    layers = [O.synth_frame(full, 700 + i) for i in range(8)]
    want = O.HostFrame16(full)
    arr = (O.Frame16 * 8)(*[l.c for l in layers])
    O.oracle().orc_composite_f16(C.byref(want.c), arr, 8)
    out = poisoned(dev, full)
    dev.composite_over(out, [to_device(dev, l) for l in layers])
    assert_frames_equal(out.download(), out.current.tuple(), want, "composite/1080p x8")


# ------------------------------------------------------------------ same-signature batches (persistent k_seg)

@pytest.mark.parametrize("shape", ["convert", "cross_u8", "timeline", "layers_u8", "windows", "timeline_windows"])
def test_uniform_batch_matches_oracle(dev, shape):
    """Runs of jobs with one signature go through ONE persistent launch (k_seg BATCH); every frame is
    checked against the oracle, and the launch count proves the batch form was taken."""
    w, h = 200, 37                      # 100 pairs per row: a short last segment in every row
    full = (0, 0, w - 1, h - 1)
    orc = O.oracle()
    n = 7
    pics_a = [O.synth_coded(w, h, 500 + i, 2, O.LAYOUT_UYVY, O.MATRIX_709, (0, 0), 0, True) for i in range(n)]
    pics_b = [O.synth_coded(w, h, 600 + i, 2, O.LAYOUT_UYVY, O.MATRIX_709, (0, 0), 0, True) for i in range(n)]
    frames = [O.synth_frame(full, 700 + i, full) for i in range(n + 3)]
    # the same partial windows in every job (one geometry), odd edges included: the classifying mode, batched
    boxes = [(0, 0, w - 1, h - 1), (13, 3, 150, 30), (64, 0, 199, 20), (1, 17, 77, 36)]
    windowed = [[O.synth_frame(full, 900 + 10 * i + k, boxes[k]) for k in range(4)] for i in range(n)] if "windows" in shape else []
    dev_w = [[to_device(dev, f) for f in row] for row in windowed]
    dev_a, dev_b = [device_coded(dev, p) for p in pics_a], [device_coded(dev, p) for p in pics_b]
    dev_f = [to_device(dev, f) for f in frames]
    table = dev.display_table(1.25)
    srgb, ramp = np.zeros(65536, np.uint16), np.zeros(65536, np.uint8)
    orc.orc_build_lut(O.LUT_LINEAR_TO_SRGB, O.ptr(srgb))
    orc.orc_build_intent_ramp(1.25, O.ptr(ramp))

    jobs, outs, u8s, wants, want_u8 = [], [], [], [], []
    for i in range(n):
        want = O.HostFrame16(full)
        out = poisoned(dev, full)
        u8 = None
        if shape == "convert":
            orc.orc_reconstruct(C.byref(want.c), C.byref(pics_a[i].c))
            job = dev.make_job(out=out, a=dev_a[i])
        elif shape == "cross_u8":
            mix = 0.125 + 0.1 * i
            orc.orc_cross_f16(C.byref(want.c), C.byref(frames[i].c), C.byref(frames[i + 1].c), mix)
            u8 = dev.DeviceBuffer(w * h * 4)
            job = dev.make_job(out=out, a=dev_f[i], b=dev_f[i + 1], mix_b=mix, out_u8=u8, u8_table=table)
        elif shape == "timeline":
            mix = 0.05 + 0.13 * i
            layers = [frames[i + 1], frames[i + 2], frames[i + 3]]
            arr = (O.Frame16 * 3)(*[l.c for l in layers])
            orc.orc_timeline_frame_f16(C.byref(want.c), C.byref(pics_a[i].c), C.byref(pics_b[i].c), mix, arr, 3)
            job = dev.make_job(out=out, a=dev_a[i], b=dev_b[i], mix_b=mix, layers=[dev_f[i + 1], dev_f[i + 2], dev_f[i + 3]])
        elif shape == "windows":
            arr = (O.Frame16 * 4)(*[l.c for l in windowed[i]])
            orc.orc_composite_f16(C.byref(want.c), arr, 4)
            job = dev.make_job(out=out, layers=dev_w[i])
        elif shape == "timeline_windows":
            mix = 0.05 + 0.13 * i
            arr = (O.Frame16 * 3)(*[l.c for l in windowed[i][1:]])
            orc.orc_timeline_frame_f16(C.byref(want.c), C.byref(pics_a[i].c), C.byref(pics_b[i].c), mix, arr, 3)
            job = dev.make_job(out=out, a=dev_a[i], b=dev_b[i], mix_b=mix, layers=dev_w[i][1:])
        else:
            layers = [frames[i], frames[i + 1], frames[i + 2]]
            arr = (O.Frame16 * 3)(*[l.c for l in layers])
            orc.orc_composite_f16(C.byref(want.c), arr, 3)
            u8 = dev.DeviceBuffer(w * h * 4)
            job = dev.make_job(out=out, layers=[dev_f[i], dev_f[i + 1], dev_f[i + 2]], out_u8=u8, u8_table=table)
        if u8 is not None:
            copy = want.copy()
            target = np.zeros((h, w, 4), np.uint8)
            orc.orc_encode_u8(C.byref(copy.c), O.ptr(srgb), O.ptr(ramp), O.ptr(target))
            want_u8.append(target)
        jobs.append(job); outs.append(out); u8s.append(u8); wants.append(want)

    before = dev.lib().cv_kernel_launches()
    dev.render_frames(jobs)
    dev.sync()
    if not os.environ.get("CV_NO_SEG_BATCH") and os.environ.get("CV_SEG_BATCH_GENERAL", "1") != "0":
        assert dev.lib().cv_kernel_launches() - before == 1, "the run was not taken as one batch launch"
    for i in range(n):
        assert_frames_equal(outs[i].download(), outs[i].current.tuple(), wants[i], "%s batch job %d" % (shape, i))
        if u8s[i] is not None:
            assert np.array_equal(u8s[i].download(np.uint8, (h, w, 4)), want_u8[i]), "%s batch job %d: 8-bit output" % (shape, i)


def test_large_mixed_batch_in_chunks(dev):
    """cv_render_frames builds and launches a call in chunks of 1024 jobs; inside a chunk jobs are grouped by kind.
    2 600 jobs of three kinds interleaved (a 4:2:2 conversion, a two-layer stack with a partial window, a crossfade),
    each with its own output: every output against the oracle; then a bad job late in the batch makes the call fail."""
    from canvas_b200 import _abi as A
    full = (0, 0, 63, 7)
    orc = O.oracle()
    pic = O.synth_coded(64, 8, 5, 2, O.LAYOUT_UYVY, O.MATRIX_709, (0, 0))
    dpic = device_coded(dev, pic)
    la, lb = O.synth_frame(full, 11, full), O.synth_frame(full, 12, (6, 1, 50, 6))
    dla, dlb = to_device(dev, la), to_device(dev, lb)

    want_convert = O.HostFrame16(full)
    orc.orc_reconstruct(C.byref(want_convert.c), C.byref(pic.c))
    want_stack = O.HostFrame16(full)
    orc.orc_composite_f16(C.byref(want_stack.c), (O.Frame16 * 2)(la.c, lb.c), 2)
    want_cross = O.HostFrame16(full)
    orc.orc_cross_f16(C.byref(want_cross.c), C.byref(la.c), C.byref(lb.c), 0.375)
    wants = [want_convert, want_stack, want_cross]

    count = 2600
    outs = [dev.DeviceFrame16(full) for _ in range(count)]
    jobs = []
    for i, out in enumerate(outs):
        kind = (i * 7 + i // 5) % 3
        if kind == 0:
            jobs.append(dev.make_job(out=out, a=dpic))
        elif kind == 1:
            jobs.append(dev.make_job(out=out, layers=[dla, dlb]))
        else:
            jobs.append(dev.make_job(out=out, a=dla, b=dlb, mix_b=0.375))
    dev.render_frames(jobs)
    dev.sync()
    for i in range(0, count, 37):           # a spread of them (downloading all 2 600 would only cost time)
        kind = (i * 7 + i // 5) % 3
        assert_frames_equal(outs[i].download(), outs[i].current.tuple(), wants[kind], "job %d kind %d" % (i, kind))
    for i in (1023, 1024, 1025, 2047, 2048, 2599):
        kind = (i * 7 + i // 5) % 3
        assert_frames_equal(outs[i].download(), outs[i].current.tuple(), wants[kind], "job %d kind %d" % (i, kind))

    bad = dev.make_job(out=outs[0], a=dpic)
    bad.layer_count = 99
    arr = (dev.FrameJob * 1500)(*(jobs[:1400] + [bad] + jobs[1401:1500]))
    assert dev.lib().cv_render_frames(arr, 1500, None) != 0
    dev.sync()
