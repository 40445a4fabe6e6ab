"""GPU parity for SURVEY.md section 8(f) ranks 2-3: cv_color_* (color.c:100-160, oracle pinned against the compiled
reference in test_color_mpeg2.py), cv_crop_f16 (video_filter.c:79-95) and cv_subsample_mpeg2
(video_subsample.c:212-290; GLSL-only in the reference, parity unpinned: the oracle is the shader restated in C).
Everything here is compared bit for bit."""
import ctypes as C

import numpy as np
import pytest

import oraclelib as O
from test_color_mpeg2 import COLOR_WINDOWS, special_halves, mpeg2_picture, solid_frame

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev():
    from canvas_b200 import device
    device.init(0)
    return device


@pytest.mark.parametrize("which", ["rgb_to_xyz_sdtv", "xyz_to_srgb"])
@pytest.mark.parametrize("full,current", COLOR_WINDOWS + [((0, 0, 1919, 1079), (1, 0, 1918, 1079))])
def test_color_device_matches_oracle(dev, which, full, current):
    src = special_halves(O.synth_frame(full, 23, current))
    want = src.copy()
    getattr(O.oracle(), "orc_color_" + which)(C.byref(want.c))
    frame = dev.DeviceFrame16.from_numpy(src.array, src.full.tuple(), src.current.tuple())
    dev.check(getattr(dev.lib(), "cv_color_" + which)(C.byref(frame.c), None), "cv_color_" + which)
    got = frame.download()
    assert frame.current.tuple() == want.current.tuple()
    assert np.array_equal(got, want.array)          # pixels outside the window included: they must be untouched


@pytest.mark.parametrize("which", ["rgb_to_xyz_sdtv", "xyz_to_srgb"])
def test_color_host_entry_point(dev, which):
    """the reference's exported symbol, host frame in place"""
    lib = dev.lib()
    fn = getattr(lib, "video_color_" + which)
    fn.argtypes, fn.restype = [C.POINTER(O.Frame16)], None
    src = O.synth_frame((-3, -2, 60, 40), 5, (0, 0, 50, 33))
    want, got = src.copy(), src.copy()
    getattr(O.oracle(), "orc_color_" + which)(C.byref(want.c))
    fn(C.byref(got.c))
    assert got.current.tuple() == want.current.tuple()
    assert np.array_equal(got.array, want.array)


@pytest.mark.parametrize("rect", [(5, 3, 20, 10), (-10, -10, 100, 100), (40, 40, 50, 50), (2, 1, 2, 14), (0, 5, 31, 6)])
def test_crop_matches_oracle(dev, rect):
    from canvas_b200 import _abi as A
    src = O.synth_frame((0, 0, 31, 15), 3, (2, 1, 29, 14))
    want = src.copy()
    O.oracle().orc_crop_f16(C.byref(want.c), C.byref(O.Box(*rect)))
    frame = dev.DeviceFrame16.from_numpy(src.array, src.full.tuple(), src.current.tuple())
    dev.check(dev.lib().cv_crop_f16(C.byref(frame.c), C.byref(A.Box2i(*rect)), None), "cv_crop_f16")
    assert frame.current.tuple() == want.current.tuple()
    assert np.array_equal(frame.download(), want.array)


MPEG2_CASES = [
    # (width, height, pic, frame full, frame current, matrix)
    (32, 16, (0, 0), (0, 0, 31, 15), None, O.MATRIX_601),
    (16, 12, (2, 2), (-4, -2, 27, 17), (2, 2, 19, 11), O.MATRIX_601),
    (64, 48, (-3, 5), (-8, -8, 80, 60), (0, 0, 50, 40), O.MATRIX_709),
    (720, 480, (0, 0), (0, 0, 719, 479), (0, 0, 719, 479), O.MATRIX_601),          # the reference's own call
    (720, 480, (0, 0), (0, 0, 719, 479), (10, 7, 700, 470), O.MATRIX_601),
    (1920, 1080, (0, 0), (0, 0, 1919, 1079), None, O.MATRIX_709),
    (2, 2, (0, 0), (0, 0, 1, 1), None, O.MATRIX_601),
]


def run_mpeg2(dev, frame_host, width, height, pic, matrix):
    want = mpeg2_picture(width, height, pic, matrix)
    O.oracle().orc_subsample_mpeg2(C.byref(frame_host.c), C.byref(want.c))
    frame = dev.DeviceFrame16.from_numpy(frame_host.array, frame_host.full.tuple(), frame_host.current.tuple())
    planes = [np.full(p.shape, 0xEE, np.uint8) for p in want.planes]
    coded = dev.DeviceCoded(planes, width, height, sub_x=2, matrix=matrix, pic=pic)
    dev.subsample_mpeg2(frame, coded)
    got = coded.download_planes()
    assert np.array_equal(frame.download(), frame_host.array)                       # the source is read-only here
    return got, want.planes


@pytest.mark.parametrize("case", MPEG2_CASES)
def test_mpeg2_device_matches_oracle(dev, case):
    width, height, pic, full, current, matrix = case
    src = O.synth_frame(full, 41, current)
    if width <= 720:
        special_halves(src)
    got, want = run_mpeg2(dev, src, width, height, pic, matrix)
    for name, g, w in zip("Y Cb Cr".split(), got, want):
        assert np.array_equal(g, w), (case, name, int((g != w).sum()))


def test_mpeg2_empty_window_is_black(dev):
    src = O.synth_frame((0, 0, 31, 15), 2, (0, 0, -1, -1))
    got, want = run_mpeg2(dev, src, 32, 16, (0, 0), O.MATRIX_601)
    assert np.all(got[0] == 16) and np.all(got[1] == 128) and np.all(got[2] == 128)
    for g, w in zip(got, want):
        assert np.array_equal(g, w)


def test_mpeg2_4k_solid_property(dev):
    """full size, size-independent property: a solid frame gives flat planes equal to the small-case oracle value"""
    rgb = (0.25, 0.5, 0.125)
    small, frame = mpeg2_picture(8, 8), solid_frame((0, 0, 7, 7), rgb)
    O.oracle().orc_subsample_mpeg2(C.byref(frame.c), C.byref(small.c))
    big = solid_frame((0, 0, 3839, 2159), rgb)
    dframe = dev.DeviceFrame16.from_numpy(big.array, big.full.tuple())
    planes = [np.zeros((2160, 3840), np.uint8), np.zeros((1080, 1920), np.uint8), np.zeros((1080, 1920), np.uint8)]
    coded = dev.DeviceCoded(planes, 3840, 2160, sub_x=2, matrix=O.MATRIX_601)
    dev.subsample_mpeg2(dframe, coded)
    for g, w in zip(coded.download_planes(), small.planes):
        assert np.all(g == w[0, 0])


def test_mpeg2_rejects_bad_pictures(dev):
    from canvas_b200 import CanvasError
    frame = dev.DeviceFrame16((0, 0, 15, 15))
    for width, height in ((15, 16), (16, 15), (0, 16)):
        planes = [np.zeros((16, 16), np.uint8), np.zeros((8, 8), np.uint8), np.zeros((8, 8), np.uint8)]
        coded = dev.DeviceCoded(planes, width, height, sub_x=2)
        with pytest.raises(CanvasError):
            dev.subsample_mpeg2(frame, coded)


def test_mpeg2_subsample_filter_type(dev):
    """process.MPEG2SubsampleFilter over a solid colour source and over a frame (MPEG2SubsampleFilter.c:45-58)"""
    from canvas_b200 import process
    src = O.synth_frame((0, 0, 719, 479), 77, (4, 2, 715, 470))
    want = mpeg2_picture(720, 480, (0, 0), O.MATRIX_601)
    O.oracle().orc_subsample_mpeg2(C.byref(src.c), C.byref(want.c))
    image = process.MPEG2SubsampleFilter(process.RgbaFrameF16((0, 0, 719, 479), (4, 2, 715, 470), src.array)).get_coded_image(0)
    assert image.stride == [720, 360, 360] and image.line_count == [480, 240, 240]
    for data, plane in zip(image.data, want.planes):
        assert data == plane.tobytes()
