"""SURVEY.md section 8(f) ranks 2 and 3, CPU side: the 3x3 colour transforms (color.c:100-160 -- the oracle is pinned bit for bit
against the compiled reference), the crop shader (video_filter.c:79-95) and the MPEG-2 4:2:0 subsample shaders
(video_subsample.c:212-290) -- both GLSL-only in the reference, so their restatements are checked for the properties
the shader text implies (parity unpinned)."""
import ctypes as C

import numpy as np
import pytest

import oraclelib as ol



def f2h(values):
    a = np.ascontiguousarray(values, np.float32)
    out = np.zeros(a.shape, np.uint16)
    ol.oracle().orc_f2h_n(ol.ptr(out), ol.ptr(a), a.size)
    return out


COLOR_WINDOWS = [((0, 0, 63, 31), None), ((-5, -3, 40, 20), (-2, 0, 33, 17)), ((0, 0, 15, 7), (4, 2, 4, 2))]


def special_halves(frame):
    """sprinkles infinities, NaNs, subnormals, negative zero and large values over a frame's array"""
    flat = frame.array.reshape(-1)
    specials = np.array([0x7C00, 0xFC00, 0x0001, 0x83FF, 0x8000, 0x7BFF, 0xFBFF, 0x3C00, 0x0000, 0x7800, 0xF800], np.uint16)
    spots = np.arange(specials.size) * (flat.size // specials.size)
    flat[spots] = specials
    return frame


@pytest.mark.skipif(not ol.have_reference(), reason="oracle/_ref/libcanvasref.so not built")
@pytest.mark.parametrize("which", ["video_color_rgb_to_xyz_sdtv", "video_color_xyz_to_srgb"])
@pytest.mark.parametrize("full,current", COLOR_WINDOWS)
def test_color_oracle_matches_reference(which, full, current):
    O, R = ol.oracle(), ol.reference()
    src = special_halves(ol.synth_frame(full, 11, current))
    want, got = src.copy(), src.copy()
    getattr(R, which)(C.byref(want.c))
    getattr(O, which.replace("video_", "orc_"))(C.byref(got.c))
    assert got.current.tuple() == want.current.tuple()
    assert np.array_equal(got.array, want.array)
    assert not np.array_equal(got.window_view(), src.window_view())


def mpeg2_picture(width, height, pic=(0, 0), matrix=ol.MATRIX_601):
    p = ol.CodedPicture(width, height, 2, ol.LAYOUT_PLANAR, matrix, pic)
    # 4:2:0: the chroma planes have half the rows as well
    p.planes[1] = np.zeros((height // 2, width // 2), np.uint8)
    p.planes[2] = np.zeros((height // 2, width // 2), np.uint8)
    for i in (1, 2):
        p.c.plane[i] = p.planes[i].ctypes.data
        p.c.stride[i] = p.planes[i].shape[1]
    return p


def solid_frame(full, rgb, current=None):
    f = ol.HostFrame16(full, current)
    f.array[..., :3] = f2h(np.asarray(rgb, np.float32))
    f.array[..., 3] = 0x3C00
    return f


def test_mpeg2_solid_colours_are_flat_and_studio_range():
    """a solid frame has flat planes: black -> (16, 128, 128), white -> (235, 128, 128) (:237, :284)"""
    O = ol.oracle()
    for rgb, expect in (((0, 0, 0), (16, 128, 128)), ((1, 1, 1), (235, 128, 128))):
        pic, frame = mpeg2_picture(32, 16), solid_frame((0, 0, 31, 15), rgb)
        O.orc_subsample_mpeg2(C.byref(frame.c), C.byref(pic.c))
        for plane, value in zip(pic.planes, expect):
            assert np.all(plane == value), (rgb, value, np.unique(plane))
    # a saturated linear red: Y' = .299 -> 16 + .299*219 = 81.48, Pb = -.168736 -> 90.2, Pr = .5 -> 240
    pic, frame = mpeg2_picture(32, 16), solid_frame((0, 0, 31, 15), (1, 0, 0))
    O.orc_subsample_mpeg2(C.byref(frame.c), C.byref(pic.c))
    assert [int(p[3, 3]) for p in pic.planes] == [81, 90, 240]


def test_mpeg2_chroma_rows_follow_the_interlace_pattern():
    """chroma row j draws 12/16 from luma row 2j (j even) or 2j+1 (j odd) and 4/16 from row 2j+2 / 2j-1 (:265-281)"""
    O = ol.oracle()
    full = (0, 0, 15, 15)
    for lit_row in range(16):
        f = solid_frame(full, (0, 0, 0))
        f.array[lit_row, :, 0] = f2h(np.float32(1.0))         # one red row
        pic = mpeg2_picture(16, 16)
        O.orc_subsample_mpeg2(C.byref(f.c), C.byref(pic.c))
        touched = {j for j in range(8) if pic.planes[2][j, 4] != 128}
        expect = set()
        for j in range(8):
            near, far = (2 * j + 1, 2 * j - 1) if j & 1 else (2 * j, 2 * j + 2)
            if lit_row in (near, min(max(far, 0), 15)):
                expect.add(j)
        assert touched == expect, (lit_row, touched, expect)
        # the luma plane shows exactly that row
        assert set(np.nonzero(pic.planes[0][:, 5] != 16)[0]) == {lit_row}


def test_mpeg2_window_and_offset():
    """pixels outside current_window read as zero (black after the transfer function); pic_x/pic_y shift the picture"""
    O = ol.oracle()
    full, current = (-4, -2, 27, 17), (2, 2, 19, 11)
    f = solid_frame(full, (1, 1, 1), current)
    pic = mpeg2_picture(16, 12, pic=(0, 0))
    O.orc_subsample_mpeg2(C.byref(f.c), C.byref(pic.c))
    luma = pic.planes[0]
    assert np.all(luma[2:12, 2:16] == 235) and np.all(luma[:2, :] == 16) and np.all(luma[:, :2] == 16)
    shifted = mpeg2_picture(16, 12, pic=(2, 2))
    O.orc_subsample_mpeg2(C.byref(f.c), C.byref(shifted.c))
    assert np.all(shifted.planes[0][:10, :] == 235) and np.all(shifted.planes[0][10:, :] == 16)
    assert np.all(shifted.planes[1][:4, :] == 128)


def test_crop_oracle():
    O = ol.oracle()
    f = ol.synth_frame((0, 0, 31, 15), 3, (2, 1, 29, 14))
    before = f.array.copy()
    O.orc_crop_f16(C.byref(f.c), C.byref(ol.Box(5, 3, 20, 10)))
    assert f.current.tuple() == (2, 1, 29, 14)
    assert np.array_equal(f.array[3:11, 5:21], before[3:11, 5:21])
    inside = np.zeros((16, 32), bool)
    inside[3:11, 5:21] = True
    window = np.zeros((16, 32), bool)
    window[1:15, 2:30] = True
    assert np.all(f.array[window & ~inside] == 0)
    assert np.array_equal(f.array[~window], before[~window])        # outside the current window: untouched
