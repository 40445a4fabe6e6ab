"""ctypes bindings for the CHECKERS: oracle/liboracle.so (our CPU restatement) and,
when present, oracle/_ref/libcanvasref.so (the reference's own C compiled unmodified).

Test infrastructure only -- the product (canvas_b200/) never imports this.
"""
import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_SO = os.path.join(ROOT, "oracle", "liboracle.so")
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libcanvasref.so")
REF_FAST_SO = os.path.join(ROOT, "oracle", "_ref", "libcanvasref_fast.so")


class Box(C.Structure):
    """box2i (include/framework.h:62-64): inclusive min/max corners."""
    _fields_ = [("min_x", C.c_int32), ("min_y", C.c_int32), ("max_x", C.c_int32), ("max_y", C.c_int32)]

    def __init__(self, min_x=0, min_y=0, max_x=-1, max_y=-1):
        super().__init__(min_x, min_y, max_x, max_y)

    def tuple(self):
        return (self.min_x, self.min_y, self.max_x, self.max_y)

    @property
    def width(self):
        return max(0, self.max_x - self.min_x + 1)

    @property
    def height(self):
        return max(0, self.max_y - self.min_y + 1)

    @property
    def empty(self):
        return self.max_x < self.min_x or self.max_y < self.min_y

    def __repr__(self):
        return f"Box{self.tuple()}"


class Frame16(C.Structure):
    """rgba_frame_f16 (include/framework.h:167-171)."""
    _fields_ = [("data", C.c_void_p), ("full", Box), ("current", Box)]


class Frame32(C.Structure):
    """rgba_frame_f32 (include/framework.h:173-177)."""
    _fields_ = [("data", C.c_void_p), ("full", Box), ("current", Box)]


class Coded(C.Structure):
    """orc_coded (oracle/oracle.h)."""
    _fields_ = [("plane", C.c_void_p * 3), ("stride", C.c_int32 * 3), ("layout", C.c_int32),
                ("width", C.c_int32), ("height", C.c_int32), ("sub_x", C.c_int32), ("matrix", C.c_int32),
                ("pic_x", C.c_int32), ("pic_y", C.c_int32)]


class RefCodedImage(C.Structure):
    """coded_image (include/framework.h:470-476)."""
    _fields_ = [("data", C.c_void_p * 4), ("stride", C.c_int * 4), ("line_count", C.c_int * 4), ("free_func", C.c_void_p)]


LAYOUT_PLANAR, LAYOUT_UYVY, LAYOUT_YUY2 = 0, 1, 2
MATRIX_709, MATRIX_601 = 0, 1
LUT_709_TO_LINEAR_SCENE, LUT_709_TO_LINEAR_DISPLAY, LUT_LINEAR_TO_709, LUT_LINEAR_TO_SRGB = 0, 1, 2, 3


def ptr(a):
    return a.ctypes.data_as(C.c_void_p)


class HostFrame16:
    """A numpy-backed f16 RGBA frame: array shape (h, w, 4) uint16 (raw half bits)."""

    def __init__(self, full, current=None, array=None):
        self.full = Box(*full)
        h, w = self.full.height, self.full.width
        self.array = np.zeros((h, w, 4), dtype=np.uint16) if array is None else np.ascontiguousarray(array, dtype=np.uint16)
        assert self.array.shape == (h, w, 4)
        self.c = Frame16(self.array.ctypes.data, self.full, Box(*(current if current is not None else full)))

    @property
    def current(self):
        return self.c.current

    def window_view(self, box=None):
        box = box or self.c.current
        if box.empty:
            return self.array[0:0, 0:0]
        return self.array[box.min_y - self.full.min_y: box.max_y - self.full.min_y + 1,
                          box.min_x - self.full.min_x: box.max_x - self.full.min_x + 1]

    def copy(self):
        return HostFrame16(self.full.tuple(), self.c.current.tuple(), self.array.copy())


class HostFrame32:
    def __init__(self, full, current=None, array=None):
        self.full = Box(*full)
        h, w = self.full.height, self.full.width
        self.array = np.zeros((h, w, 4), dtype=np.float32) if array is None else np.ascontiguousarray(array, dtype=np.float32)
        assert self.array.shape == (h, w, 4)
        self.c = Frame32(self.array.ctypes.data, self.full, Box(*(current if current is not None else full)))

    @property
    def current(self):
        return self.c.current

    def window_view(self, box=None):
        box = box or self.c.current
        if box.empty:
            return self.array[0:0, 0:0]
        return self.array[box.min_y - self.full.min_y: box.max_y - self.full.min_y + 1,
                          box.min_x - self.full.min_x: box.max_x - self.full.min_x + 1]

    def copy(self):
        return HostFrame32(self.full.tuple(), self.c.current.tuple(), self.array.copy())


class CodedPicture:
    """8-bit Y'CbCr picture, planar or packed, numpy-backed."""

    def __init__(self, width, height, sub_x=2, layout=LAYOUT_PLANAR, matrix=MATRIX_709, pic=(0, 0), stride_pad=0):
        self.width, self.height, self.sub_x, self.layout, self.matrix, self.pic = width, height, sub_x, layout, matrix, pic
        cw = (width + sub_x - 1) // sub_x
        if layout == LAYOUT_PLANAR:
            self.planes = [np.zeros((height, width + stride_pad), np.uint8),
                           np.zeros((height, cw + stride_pad), np.uint8),
                           np.zeros((height, cw + stride_pad), np.uint8)]
        else:
            assert sub_x == 2
            self.planes = [np.zeros((height, 4 * cw + stride_pad), np.uint8)]
        self.c = Coded()
        for i, p in enumerate(self.planes):
            self.c.plane[i] = p.ctypes.data
            self.c.stride[i] = p.shape[1]
        self.c.layout, self.c.width, self.c.height = layout, width, height
        self.c.sub_x, self.c.matrix, self.c.pic_x, self.c.pic_y = sub_x, matrix, pic[0], pic[1]

    def set_samples(self, y, cb, cr):
        """y: (h, w) ; cb, cr: (h, ceil(w/sub))."""
        h, w = self.height, self.width
        cw = cb.shape[1]
        if self.layout == LAYOUT_PLANAR:
            self.planes[0][:, :w] = y
            self.planes[1][:, :cw] = cb
            self.planes[2][:, :cw] = cr
        else:
            p = self.planes[0]
            ypad = np.zeros((h, 2 * cw), np.uint8)
            ypad[:, :w] = y
            if self.layout == LAYOUT_UYVY:
                p[:, 0:4 * cw:4], p[:, 1:4 * cw:4], p[:, 2:4 * cw:4], p[:, 3:4 * cw:4] = cb, ypad[:, 0::2], cr, ypad[:, 1::2]
            else:
                p[:, 0:4 * cw:4], p[:, 1:4 * cw:4], p[:, 2:4 * cw:4], p[:, 3:4 * cw:4] = ypad[:, 0::2], cb, ypad[:, 1::2], cr

    def get_samples(self):
        h, w = self.height, self.width
        cw = (w + self.sub_x - 1) // self.sub_x
        if self.layout == LAYOUT_PLANAR:
            return self.planes[0][:, :w].copy(), self.planes[1][:, :cw].copy(), self.planes[2][:, :cw].copy()
        p = self.planes[0]
        y = np.zeros((h, 2 * cw), np.uint8)
        if self.layout == LAYOUT_UYVY:
            cb, y[:, 0::2], cr, y[:, 1::2] = p[:, 0:4 * cw:4], p[:, 1:4 * cw:4], p[:, 2:4 * cw:4], p[:, 3:4 * cw:4]
        else:
            y[:, 0::2], cb, y[:, 1::2], cr = p[:, 0:4 * cw:4], p[:, 1:4 * cw:4], p[:, 2:4 * cw:4], p[:, 3:4 * cw:4]
        return y[:, :w].copy(), cb.copy(), cr.copy()


_oracle = None
_ref = None


def oracle():
    """liboracle.so with argtypes set.  Raises if it has not been built (`make -C oracle`)."""
    global _oracle
    if _oracle is None:
        L = C.CDLL(ORACLE_SO)
        L.orc_f2h.argtypes, L.orc_f2h.restype = [C.c_float], C.c_uint16
        L.orc_h2f.argtypes, L.orc_h2f.restype = [C.c_uint16], C.c_float
        L.orc_f2h_n.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
        L.orc_h2f_n.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
        L.orc_build_lut.argtypes = [C.c_int, C.c_void_p]
        L.orc_build_gamma45.argtypes = [C.c_void_p]
        L.orc_build_intent_ramp.argtypes = [C.c_float, C.c_void_p]
        L.orc_triangle.argtypes = [C.c_float, C.c_float, C.c_void_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.orc_reconstruct.argtypes = [C.POINTER(Frame16), C.POINTER(Coded)]
        L.orc_subsample.argtypes = [C.POINTER(Frame16), C.POINTER(Coded)]
        L.orc_copy_frame_f16.argtypes = [C.POINTER(Frame16), C.POINTER(Frame16)]
        L.orc_copy_frame_alpha_f32.argtypes = [C.POINTER(Frame32), C.POINTER(Frame32), C.c_float]
        L.orc_mix_cross_f32.argtypes = [C.POINTER(Frame32), C.POINTER(Frame32), C.POINTER(Frame32), C.c_float]
        L.orc_mix_over_f32.argtypes = [C.POINTER(Frame32), C.POINTER(Frame32), C.c_float]
        L.orc_scale_bilinear_f32.argtypes = [C.POINTER(Frame32), C.c_float, C.c_float, C.POINTER(Frame32), C.c_float, C.c_float,
                                             C.c_float, C.c_float]
        L.orc_subsample_mpeg2.argtypes = [C.POINTER(Frame16), C.POINTER(Coded)]
        L.orc_color_rgb_to_xyz_sdtv.argtypes = [C.POINTER(Frame16)]
        L.orc_color_xyz_to_srgb.argtypes = [C.POINTER(Frame16)]
        L.orc_crop_f16.argtypes = [C.POINTER(Frame16), C.POINTER(Box)]
        L.orc_frame_f16_to_f32.argtypes = [C.POINTER(Frame32), C.POINTER(Frame16)]
        L.orc_frame_f32_to_f16.argtypes = [C.POINTER(Frame16), C.POINTER(Frame32)]
        L.orc_composite_f16.argtypes = [C.POINTER(Frame16), C.POINTER(Frame16), C.c_int]
        L.orc_cross_f16.argtypes = [C.POINTER(Frame16), C.POINTER(Frame16), C.POINTER(Frame16), C.c_float]
        L.orc_timeline_frame_f16.argtypes = [C.POINTER(Frame16), C.POINTER(Coded), C.POINTER(Coded), C.c_float,
                                             C.POINTER(Frame16), C.c_int]
        L.orc_encode_u8.argtypes = [C.POINTER(Frame16), C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_to_argb32.argtypes = [C.POINTER(Frame16), C.c_void_p, C.c_void_p]
        L.orc_hash32.argtypes, L.orc_hash32.restype = [C.c_uint32, C.c_uint32], C.c_uint32
        L.orc_fill_frame.argtypes = [C.POINTER(Frame16), C.c_uint32]
        L.orc_fill_bytes.argtypes = [C.c_void_p, C.c_size_t, C.c_uint32, C.c_int, C.c_int]
        L.orc_fnv1a.argtypes, L.orc_fnv1a.restype = [C.c_void_p, C.c_size_t], C.c_uint32
        _oracle = L
    return _oracle


def have_reference():
    return os.path.exists(REF_SO)


def reference(fast=False):
    """The reference's own C (compiled from /root/reference by `make -C oracle ref`)."""
    global _ref
    if fast:
        L = C.CDLL(REF_FAST_SO)
        _set_ref_types(L)
        L.ref_init()
        return L
    if _ref is None:
        L = C.CDLL(REF_SO)
        _set_ref_types(L)
        L.ref_init()
        _ref = L
    return _ref


class V2f(C.Structure):
    """v2f (include/framework.h): passed BY VALUE to video_scale_bilinear_f32."""
    _fields_ = [("x", C.c_float), ("y", C.c_float)]


def _set_ref_types(L):
    L.video_scale_bilinear_f32.argtypes = [C.POINTER(Frame32), V2f, C.POINTER(Frame32), V2f, V2f]
    L.video_scale_bilinear_f32.restype = None
    for name in ("video_color_rgb_to_xyz_sdtv", "video_color_xyz_to_srgb"):
        getattr(L, name).argtypes, getattr(L, name).restype = [C.POINTER(Frame16)], None
    L.video_get_gamma45_ramp.restype = C.POINTER(C.c_uint8)
    for name in ("video_transfer_rec709_to_linear_scene", "video_transfer_rec709_to_linear_display",
                 "video_transfer_linear_to_rec709", "video_transfer_linear_to_sRGB"):
        getattr(L, name).argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
    L.video_reconstruct_dv.argtypes = [C.POINTER(Frame16), C.POINTER(RefCodedImage)]
    L.video_subsample_dv.argtypes, L.video_subsample_dv.restype = [C.POINTER(Frame16)], C.POINTER(RefCodedImage)
    L.video_copy_frame_f16.argtypes = [C.POINTER(Frame16), C.POINTER(Frame16)]
    L.video_copy_frame_alpha_f32.argtypes = [C.POINTER(Frame32), C.POINTER(Frame32), C.c_float]
    L.video_mix_cross_f32.argtypes = [C.POINTER(Frame32), C.POINTER(Frame32), C.POINTER(Frame32), C.c_float]
    L.video_mix_over_f32.argtypes = [C.POINTER(Frame32), C.POINTER(Frame32), C.c_float]
    L.ref_workspace_composite_f16.argtypes = [C.POINTER(Frame16), C.POINTER(C.c_int64), C.c_int, C.POINTER(Frame16)]
    L.ref_cross_pull_f16.argtypes = [C.POINTER(Frame16), C.POINTER(Frame16), C.c_float, C.POINTER(Frame16)]
    L.ref_encode_u8.argtypes = [C.POINTER(Frame16), C.c_void_p, C.c_void_p]
    L.ref_timeline_create.restype = C.c_void_p
    L.ref_timeline_create.argtypes = [C.POINTER(Frame16), C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int64),
                                      C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.c_int]
    L.ref_timeline_get_frame_f16.argtypes = [C.c_void_p, C.c_int, C.POINTER(Frame16)]
    L.ref_timeline_free.argtypes = [C.c_void_p]
    L.ref_time_timeline.restype = C.c_int64
    L.ref_time_timeline.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.POINTER(Frame16), C.c_int]
    L.ref_now_ns.restype = C.c_int64


def ref_half_from_float(L, arr):
    """Call through the reference's exported function POINTER half_convert_from_float (half.c:87-105)."""
    fn = C.cast(C.c_void_p.in_dll(L, "half_convert_from_float").value, C.CFUNCTYPE(None, C.c_void_p, C.c_void_p, C.c_int))
    arr = np.ascontiguousarray(arr, np.float32)
    out = np.zeros(arr.shape, np.uint16)
    fn(ptr(out), ptr(arr), arr.size)
    return out


def ref_half_to_float(L, arr):
    fn = C.cast(C.c_void_p.in_dll(L, "half_convert_to_float").value, C.CFUNCTYPE(None, C.c_void_p, C.c_void_p, C.c_int))
    arr = np.ascontiguousarray(arr, np.uint16)
    out = np.zeros(arr.shape, np.float32)
    fn(ptr(out), ptr(arr), arr.size)
    return out


def fnv1a(data):
    a = np.ascontiguousarray(data)
    return oracle().orc_fnv1a(ptr(a), a.nbytes)


def synth_frame(full, seed, current=None):
    """Deterministic synthetic f16 frame (orc_fill_frame); current window defaults to full."""
    f = HostFrame16(full, current)
    oracle().orc_fill_frame(C.byref(f.c), seed)
    return f


def synth_bytes(shape, seed, lo, hi):
    a = np.zeros(shape, np.uint8)
    oracle().orc_fill_bytes(ptr(a), a.size, seed, lo, hi)
    return a


def synth_coded(width, height, seed, sub_x=2, layout=LAYOUT_PLANAR, matrix=MATRIX_709, pic=(0, 0), stride_pad=0,
                full_range=False):
    pic_ = CodedPicture(width, height, sub_x, layout, matrix, pic, stride_pad)
    cw = (width + sub_x - 1) // sub_x
    ylo, yhi, clo, chi = (0, 255, 0, 255) if full_range else (16, 235, 16, 240)
    pic_.set_samples(synth_bytes((height, width), seed * 3 + 1, ylo, yhi),
                     synth_bytes((height, cw), seed * 3 + 2, clo, chi),
                     synth_bytes((height, cw), seed * 3 + 3, clo, chi))
    return pic_
