"""Host proof of the division-free studio-range decode used by the fast 4:2:2 conversion
(canvas_b200/csrc/cv_kernels.cu: studio_decode2): q0 = a*r, rem = fma(-d, q0, a), q = fma(rem, r, q0)
with r = RN(1/d) must equal the IEEE quotient a/d of video_reconstruct.c:31-39 for every byte."""
import ctypes as C
import ctypes.util

import numpy as np

libm = C.CDLL(ctypes.util.find_library("m"))
libm.fmaf.restype = C.c_float
libm.fmaf.argtypes = [C.c_float, C.c_float, C.c_float]


def decode(byte, bias, divisor):
    f32 = np.float32
    r = f32(1.0) / f32(divisor)
    a = f32(byte) - f32(bias)
    q0 = f32(a * r)
    rem = libm.fmaf(float(-f32(divisor)), float(q0), float(a))
    return f32(libm.fmaf(rem, float(r), float(q0)))


def test_luma_and_chroma_decode_equal_ieee_division_for_all_bytes():
    for byte in range(256):
        want_y = np.float32(np.float32(byte) - np.float32(16.0)) / np.float32(219.0)
        want_c = np.float32(np.float32(byte) - np.float32(128.0)) / np.float32(224.0)
        assert decode(byte, 16.0, 219.0).tobytes() == np.float32(want_y).tobytes(), byte
        assert decode(byte, 128.0, 224.0).tobytes() == np.float32(want_c).tobytes(), byte


def test_byte_to_float_trick_is_exact():
    for byte in range(256):
        bits = np.uint32(0x4B000000 | byte)
        assert np.frombuffer(bits.tobytes(), np.float32)[0] == np.float32(8388608.0 + byte)


def test_converted_values_stay_inside_the_compact_table():
    """The 64 KiB conversion table staged by the layered kernels (cv_runtime.c: g_host_lut709_compact) holds the
    positive halves below 0x4080 (2.25) and the negative halves of magnitude below 0x3F80 (1.875).  Every operation
    of the conversion is monotone in each input, so its extremes are taken at corner bytes: walk all of them, both
    matrices, even and odd pixel."""
    import itertools
    f32 = np.float32
    matrices = {"709": (1.5748, -0.187324, -0.468124, 1.8556), "601": (1.402, -0.344136, -0.714136, 1.772)}
    lo, hi = 0.0, 0.0
    for (m2, m4, m5, m7) in matrices.values():
        m2, m4, m5, m7 = f32(m2), f32(m4), f32(m5), f32(m7)
        for Y, cb0, cr0, cb1, cr1 in itertools.product((0, 255), repeat=5):
            y = f32(f32(Y) - f32(16)) / f32(219)
            c = [f32(f32(v) - f32(128)) / f32(224) for v in (cb0, cr0, cb1, cr1)]
            for cb, cr in ((c[0], c[1]), (f32(c[0] * f32(.5)) + f32(c[2] * f32(.5)), f32(c[1] * f32(.5)) + f32(c[3] * f32(.5)))):
                values = (y + f32(cr * m2), f32(y + f32(cb * m4)) + f32(cr * m5), y + f32(cb * m7))
                lo, hi = min(lo, *map(float, values)), max(hi, *map(float, values))
    assert hi < 2.25 and lo > -1.875, (lo, hi)
    assert 2.1 < hi and lo < -1.1          # the bound is not vacuous: the corners really reach past 2.0 and -1.0
