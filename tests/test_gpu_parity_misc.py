"""GPU parity, part 2: half pack/unpack, transfer tables, encode side, 8-bit outputs, f32 frame ops."""
import ctypes as C

import numpy as np
import pytest

import oraclelib as O
from test_gpu_parity import assert_frames_equal, to_device, poisoned, device_coded, dev  # noqa: F401

pytestmark = pytest.mark.gpu


def test_half_from_float_bit_exact(dev):
    rng = np.random.default_rng(5)
    bits = rng.integers(0, 2 ** 32, size=1 << 20, dtype=np.uint64).astype(np.uint32)
    special = np.array([0, 0x80000000, 0x7F800000, 0xFF800000, 0x7FC00000, 0x7F800001, 0x7F802000, 0x477FE000, 0x477FFFFF,
                        0x47800000, 0x33000000, 0x32FFFFFF, 0x38800000, 0x387FFFFF, 0x3F800000, 0x3F801FFF], np.uint32)
    bits = np.concatenate([special, bits, special[:3]])         # odd tail exercises the scalar remainder
    floats = bits.view(np.float32)
    want = np.zeros(floats.shape, np.uint16)
    O.oracle().orc_f2h_n(O.ptr(want), O.ptr(floats), floats.size)

    src = dev.DeviceBuffer.from_numpy(floats)
    dst = dev.DeviceBuffer(2 * floats.size)
    dev.check(dev.lib().cv_half_from_float(dst.ptr, src.ptr, floats.size, None))
    got = dst.download(np.uint16, floats.shape)
    assert np.array_equal(got, want)


def test_half_to_float_exact_for_all_65536(dev):
    halves = np.arange(65536, dtype=np.uint16)
    want = np.zeros(65536, np.float32)
    O.oracle().orc_h2f_n(O.ptr(want), O.ptr(halves), 65536)
    src = dev.DeviceBuffer.from_numpy(halves)
    dst = dev.DeviceBuffer(4 * 65536)
    dev.check(dev.lib().cv_half_to_float(dst.ptr, src.ptr, 65536, None))
    got = dst.download(np.uint32, (65536,))
    assert np.array_equal(got, want.view(np.uint32))


@pytest.mark.parametrize("which", [0, 1, 2, 3])
def test_transfer_tables_all_inputs(dev, which):
    table = np.zeros(65536, np.uint16)
    O.oracle().orc_build_lut(which, O.ptr(table))
    host = np.ctypeslib.as_array(dev.lib().cv_host_lut(which), shape=(65536,))
    assert np.array_equal(host, table), "host-built table %d differs from the oracle's" % which

    halves = np.concatenate([np.arange(65536, dtype=np.uint16), np.arange(65535, 0, -7, dtype=np.uint16)])
    buf = dev.DeviceBuffer.from_numpy(halves)
    dev.check(dev.lib().cv_transfer(which, buf.ptr, buf.ptr, halves.size, None))       # in place, like the reference allows
    got = buf.download(np.uint16, halves.shape)
    assert np.array_equal(got, table[halves])


def test_gamma45_and_intent_ramps(dev):
    want = np.zeros(65536, np.uint8)
    O.oracle().orc_build_gamma45(O.ptr(want))
    got = np.ctypeslib.as_array(dev.lib().cv_host_gamma45_ramp(), shape=(65536,))
    assert np.array_equal(got, want)
    for intent in (1.25, 1.0, 0.8):
        O.oracle().orc_build_intent_ramp(intent, O.ptr(want))
        mine = np.zeros(65536, np.uint8)
        dev.check(dev.lib().cv_build_intent_ramp(intent, O.ptr(mine)))
        assert np.array_equal(mine, want)


@pytest.mark.parametrize("write_back", [0, 1])
@pytest.mark.parametrize("current", [(0, 0, 79, 44), (3, 2, 60, 40), (0, 0, -1, -1)])
def test_encode_u8_matches_oracle(dev, write_back, current):
    full = (0, 0, 79, 44)
    frame = O.synth_frame(full, 77, current)
    srgb = np.zeros(65536, np.uint16)
    ramp = np.zeros(65536, np.uint8)
    O.oracle().orc_build_lut(O.LUT_LINEAR_TO_SRGB, O.ptr(srgb))
    O.oracle().orc_build_intent_ramp(1.25, O.ptr(ramp))
    want_frame = frame.copy()
    want_u8 = np.full((45, 80, 4), 0x5A, np.uint8)
    O.oracle().orc_encode_u8(C.byref(want_frame.c), O.ptr(srgb), O.ptr(ramp), O.ptr(want_u8))

    dframe = to_device(dev, frame)
    target = dev.DeviceBuffer.from_numpy(np.full((45, 80, 4), 0x5A, np.uint8))
    table = dev.display_table(1.25)
    dev.check(dev.lib().cv_encode_u8(C.byref(dframe.c), table, target.ptr, write_back, None))
    got_u8 = target.download(np.uint8, (45, 80, 4))
    assert np.array_equal(got_u8, want_u8)
    got_frame = dframe.download()
    assert np.array_equal(got_frame, want_frame.array if write_back else frame.array)


def test_to_argb32_matches_oracle(dev):
    full = (-2, -3, 60, 30)
    frame = O.synth_frame(full, 78, (1, 0, 50, 22))
    ramp45 = np.zeros(65536, np.uint8)
    O.oracle().orc_build_gamma45(O.ptr(ramp45))
    want = np.zeros((23, 50), np.uint32)
    O.oracle().orc_to_argb32(C.byref(frame.c), O.ptr(ramp45), O.ptr(want))
    dframe = to_device(dev, frame)
    out = dev.DeviceBuffer(want.nbytes)
    dev.check(dev.lib().cv_to_argb32(C.byref(dframe.c), out.ptr, None))
    assert np.array_equal(out.download(np.uint32, want.shape), want)


SUBSAMPLE_CASES = [
    # width, height, sub, layout, matrix, pic, full, current
    (720, 480, 4, O.LAYOUT_PLANAR, O.MATRIX_709, (0, -1), (0, -1, 719, 478), (0, -1, 719, 478)),     # DV, like DVSubsampleFilter.c:51-66
    (720, 480, 4, O.LAYOUT_PLANAR, O.MATRIX_709, (0, -1), (0, -1, 719, 478), (13, 5, 600, 300)),
    (96, 40, 2, O.LAYOUT_PLANAR, O.MATRIX_709, (0, 0), (0, 0, 95, 39), (0, 0, 95, 39)),
    (96, 40, 2, O.LAYOUT_UYVY, O.MATRIX_601, (0, 0), (-4, -4, 99, 43), (1, 1, 90, 35)),
    (96, 40, 2, O.LAYOUT_YUY2, O.MATRIX_709, (0, 0), (0, 0, 95, 39), (0, 0, 95, 39)),
    (50, 20, 1, O.LAYOUT_PLANAR, O.MATRIX_709, (2, 1), (0, 0, 63, 31), (0, 0, 63, 31)),
    (97, 21, 2, O.LAYOUT_PLANAR, O.MATRIX_601, (0, 0), (0, 0, 96, 20), (0, 0, 96, 20)),
]


@pytest.mark.parametrize("case", SUBSAMPLE_CASES, ids=lambda c: "%dx%d_s%d_l%d_m%d" % c[:5])
def test_subsample_matches_oracle_bit_exact(dev, case):
    w, h, sub, layout, matrix, pic, full, current = case
    frame = O.synth_frame(full, 91, current)
    want_frame = frame.copy()
    want_pic = O.CodedPicture(w, h, sub, layout, matrix, pic)
    O.oracle().orc_subsample(C.byref(want_frame.c), C.byref(want_pic.c))

    dframe = to_device(dev, frame)
    blank = O.CodedPicture(w, h, sub, layout, matrix, pic)
    dpic = device_coded(dev, blank, zero=True)
    dev.subsample_ycbcr8(dframe, dpic)
    dev.sync()
    for got, want in zip(dpic.download_planes(), want_pic.planes):
        assert np.array_equal(got, want)
    # the input frame is transformed in place over the covered window (video_subsample.c:147)
    assert np.array_equal(dframe.download(), want_frame.array)


# ------------------------------------------------------------------ f32 frame ops

def synth_f32(full, seed, current):
    h = O.synth_frame(full, seed, current)
    f = O.HostFrame32(full, current)
    O.oracle().orc_h2f_n(O.ptr(f.array), O.ptr(h.array), h.array.size)
    return f


def to_device32(dev, f):
    return dev.DeviceFrame32.from_numpy(f.array, f.full.tuple(), f.current.tuple())


def assert_f32_equal(dframe, want, what):
    assert dframe.current.tuple() == want.current.tuple(), what
    if want.current.empty:
        return
    b, f = want.current, want.full
    ys = slice(b.min_y - f.min_y, b.max_y - f.min_y + 1)
    xs = slice(b.min_x - f.min_x, b.max_x - f.min_x + 1)
    got = dframe.download()[ys, xs].view(np.uint32)
    exp = want.array[ys, xs].view(np.uint32)
    assert np.array_equal(got, exp), "%s: %d f32 values differ" % (what, np.count_nonzero(got != exp))


WINDOW_PAIRS = [((0, 0, 47, 31), (0, 0, 47, 31)), ((0, 0, 30, 31), (10, 4, 47, 20)), ((0, 0, 9, 9), (20, 20, 29, 29)),
                ((0, 0, -1, -1), (3, 3, 30, 30)), ((3, 3, 30, 30), (0, 0, -1, -1))]


@pytest.mark.parametrize("mix", [1.0, 0.4, 0.0])
@pytest.mark.parametrize("windows", WINDOW_PAIRS)
def test_mix_over_f32(dev, mix, windows):
    full = (0, 0, 47, 31)
    out = synth_f32(full, 1, windows[0])
    b = synth_f32(full, 2, windows[1])
    want = out.copy()
    O.oracle().orc_mix_over_f32(C.byref(want.c), C.byref(b.c), mix)
    dout, db = to_device32(dev, out), to_device32(dev, b)
    dev.check(dev.lib().cv_mix_over_f32(C.byref(dout.c), C.byref(db.c), mix, None))
    assert_f32_equal(dout, want, "over_f32 mix=%r %r" % (mix, windows))


@pytest.mark.parametrize("mix", [0.0, 0.4, 1.0])
@pytest.mark.parametrize("windows", WINDOW_PAIRS)
def test_mix_cross_f32(dev, mix, windows):
    full = (0, 0, 47, 31)
    a = synth_f32(full, 3, windows[0])
    b = synth_f32(full, 4, windows[1])
    want = O.HostFrame32(full)
    O.oracle().orc_mix_cross_f32(C.byref(want.c), C.byref(a.c), C.byref(b.c), mix)
    da, db, dout = to_device32(dev, a), to_device32(dev, b), dev.DeviceFrame32(full)
    dev.check(dev.lib().cv_mix_cross_f32(C.byref(dout.c), C.byref(da.c), C.byref(db.c), mix, None))
    assert_f32_equal(dout, want, "cross_f32 mix=%r %r" % (mix, windows))


@pytest.mark.parametrize("alpha", [1.0, 0.5, 0.0])
def test_copy_frame_alpha_f32_and_hops(dev, alpha):
    full = (0, 0, 47, 31)
    src = synth_f32((-4, -4, 60, 40), 5, (2, 2, 55, 20))
    want = O.HostFrame32(full)
    O.oracle().orc_copy_frame_alpha_f32(C.byref(want.c), C.byref(src.c), alpha)
    dsrc, dout = to_device32(dev, src), dev.DeviceFrame32(full)
    dev.check(dev.lib().cv_copy_frame_alpha_f32(C.byref(dout.c), C.byref(dsrc.c), alpha, None))
    assert_f32_equal(dout, want, "copy_alpha %r" % alpha)

    # precision hops (main.c:56-70, 126-139)
    h = O.synth_frame(full, 6, (1, 2, 40, 30))
    dh = to_device(dev, h)
    d32 = dev.DeviceFrame32(full)
    dev.check(dev.lib().cv_frame_f16_to_f32(C.byref(d32.c), C.byref(dh.c), None))
    back = poisoned(dev, full)
    dev.check(dev.lib().cv_frame_f32_to_f16(C.byref(back.c), C.byref(d32.c), None))
    assert_frames_equal(back.download(), back.current.tuple(), h, "f16->f32->f16")


def test_copy_frame_f16(dev):
    src = O.synth_frame((-4, -4, 60, 40), 8, (2, 2, 55, 20))
    want = O.HostFrame16((0, 0, 47, 31))
    O.oracle().orc_copy_frame_f16(C.byref(want.c), C.byref(src.c))
    out = poisoned(dev, (0, 0, 47, 31))
    dsrc = to_device(dev, src)
    dev.check(dev.lib().cv_copy_frame_f16(C.byref(out.c), C.byref(dsrc.c), None))
    dev.sync()
    assert_frames_equal(out.download(), out.current.tuple(), want, "copy_frame_f16")


def test_encode_u8_and_argb32_at_1080p(dev):
    """frames large enough that every thread of k_encode_u8 / k_to_argb32 walks several pixels per iteration (the
    four-pixel unroll and its tail), with an inset window: 8-bit outputs and the written-back frame bit for bit"""
    full = (0, 0, 1919, 1079)
    current = (3, 2, 1900, 1071)
    frame = O.synth_frame(full, 79, current)
    srgb = np.zeros(65536, np.uint16)
    ramp = np.zeros(65536, np.uint8)
    O.oracle().orc_build_lut(O.LUT_LINEAR_TO_SRGB, O.ptr(srgb))
    O.oracle().orc_build_intent_ramp(1.25, O.ptr(ramp))
    for write_back in (0, 1):
        want_frame = frame.copy()
        want_u8 = np.full((1080, 1920, 4), 0x5A, np.uint8)
        O.oracle().orc_encode_u8(C.byref(want_frame.c), O.ptr(srgb), O.ptr(ramp), O.ptr(want_u8))
        dframe = to_device(dev, frame)
        target = dev.DeviceBuffer.from_numpy(np.full((1080, 1920, 4), 0x5A, np.uint8))
        dev.check(dev.lib().cv_encode_u8(C.byref(dframe.c), dev.display_table(1.25), target.ptr, write_back, None))
        assert np.array_equal(target.download(np.uint8, (1080, 1920, 4)), want_u8)
        assert np.array_equal(dframe.download(), want_frame.array if write_back else frame.array)

    ramp45 = np.zeros(65536, np.uint8)
    O.oracle().orc_build_gamma45(O.ptr(ramp45))
    rows, cols = current[3] - current[1] + 1, current[2] - current[0] + 1
    want = np.zeros((rows, cols), np.uint32)
    O.oracle().orc_to_argb32(C.byref(frame.c), O.ptr(ramp45), O.ptr(want))
    dframe = to_device(dev, frame)
    out = dev.DeviceBuffer(want.nbytes)
    dev.check(dev.lib().cv_to_argb32(C.byref(dframe.c), out.ptr, None))
    assert np.array_equal(out.download(np.uint32, want.shape), want)


def test_half_from_float_at_every_truncation_boundary(dev):
    """every half value as a float, the float just above it and the float just below the next half, both signs, finite
    halves only (so that whole quads take the hardware-conversion path of k_half_from_float), plus the same mixed with
    values that force the bit-exact integer path: truncation, subnormals and the boundaries bit for bit"""
    halves = np.arange(0x7C00, dtype=np.uint16)                      # finite, positive
    exact = np.zeros(halves.size, np.float32)
    O.oracle().orc_h2f_n(O.ptr(exact), O.ptr(halves), halves.size)
    bits = exact.view(np.uint32)
    nxt = np.concatenate([bits[1:], np.array([0x47800000], np.uint32)])           # the next half as a float (65536 after the last)
    patterns = np.concatenate([bits, bits + 1, nxt - 1])
    patterns = np.concatenate([patterns, patterns | 0x80000000]).astype(np.uint32)
    mixed = patterns.copy()
    mixed[::7] = 0x7FC01234                                          # a NaN with a payload in every second quad or so
    mixed[3::11] = 0x4B000000
    for values in (patterns, mixed):
        floats = values.view(np.float32)
        want = np.zeros(floats.shape, np.uint16)
        O.oracle().orc_f2h_n(O.ptr(want), O.ptr(floats), floats.size)
        src = dev.DeviceBuffer.from_numpy(floats)
        dst = dev.DeviceBuffer(2 * floats.size)
        dev.check(dev.lib().cv_half_from_float(dst.ptr, src.ptr, floats.size, None))
        assert np.array_equal(dst.download(np.uint16, floats.shape), want)
