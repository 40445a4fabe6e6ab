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
