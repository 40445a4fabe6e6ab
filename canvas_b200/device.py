"""Host-side helpers over the C-ABI: device-resident frames and coded pictures, numpy in/out.

Pixels never get computed here: every operation is a call into libcanvas_b200.so.  numpy is
used only to hold host copies (raw half bits as uint16, shape (h, w, 4)).
"""
import ctypes as C

import numpy as np

from . import _abi
from ._abi import (Box2i, FrameF16, FrameF32, Coded, Layer, Source, FrameJob, CanvasError, check,
                   CV_MEM_DEVICE, CV_LAYOUT_PLANAR, CV_LAYOUT_UYVY, CV_LAYOUT_YUY2, CV_MATRIX_709, CV_MATRIX_601,
                   CV_SRC_NONE, CV_SRC_F16, CV_SRC_CODED, CV_SRC_F32, CV_MAX_LAYERS, LayerOps, CV_OP_WEAVE, CV_OP_GAIN_OFFSET)

_initialised = set()


def init(device=0):
    """cv_init(device); raises CanvasError if there is no B200 (no CPU fallback exists)."""
    lib = _abi.load()
    check(lib.cv_init(device), "cv_init(%d)" % device)
    _initialised.add(device)
    return lib


def lib():
    L = _abi.load()
    if not _initialised:
        init(0)
    return L


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


def sync(stream=None):
    check(lib().cv_stream_sync(stream), "cv_stream_sync")


def _quiesce():
    """cv_free's contract (include/canvas_b200.h): the work that used a block has finished before the block goes back
    to the pool.  These Python objects do not know which streams touched them (and __del__ runs whenever the collector
    likes), so they wait for the whole device; code that frees on its own stream uses cv_free_async."""
    lib().cv_device_sync()


class DeviceFrame16:
    """An rgba_frame_f16 whose data pointer lives in HBM (cv_frame_alloc_f16)."""

    def __init__(self, full, current=None):
        self.c = FrameF16()
        box = Box2i(*full)
        check(lib().cv_frame_alloc_f16(C.byref(self.c), C.byref(box), CV_MEM_DEVICE), "cv_frame_alloc_f16")
        if current is not None:
            self.c.current_window = Box2i(*current)

    @classmethod
    def from_numpy(cls, array, full, current=None, stream=None):
        f = cls(full, current if current is not None else full)
        f.upload(array, stream)
        return f

    @property
    def full(self):
        return self.c.full_window

    @property
    def current(self):
        return self.c.current_window

    def upload(self, array, stream=None):
        a = np.ascontiguousarray(array, dtype=np.uint16)
        assert a.shape == (self.full.height, self.full.width, 4), (a.shape, self.full)
        check(lib().cv_upload(self.c.data, _ptr(a), a.nbytes, stream), "cv_upload")
        sync(stream)

    def download(self, stream=None):
        a = np.empty((self.full.height, self.full.width, 4), dtype=np.uint16)
        if a.nbytes:
            check(lib().cv_download(_ptr(a), self.c.data, a.nbytes, stream), "cv_download")
            sync(stream)
        return a

    def fill(self, byte=0, stream=None):
        n = self.full.height * self.full.width * 8
        if n:
            check(lib().cv_memset(self.c.data, byte, n, stream), "cv_memset")

    def free(self):
        if self.c.data:
            _quiesce()
            lib().cv_frame_free_f16(C.byref(self.c), CV_MEM_DEVICE)

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class DeviceFrame32:
    """An rgba_frame_f32 in HBM."""

    def __init__(self, full, current=None):
        self.c = FrameF32()
        box = Box2i(*full)
        check(lib().cv_frame_alloc_f32(C.byref(self.c), C.byref(box), CV_MEM_DEVICE), "cv_frame_alloc_f32")
        if current is not None:
            self.c.current_window = Box2i(*current)

    @classmethod
    def from_numpy(cls, array, full, current=None, stream=None):
        f = cls(full, current if current is not None else full)
        a = np.ascontiguousarray(array, dtype=np.float32)
        assert a.shape == (f.full.height, f.full.width, 4)
        check(lib().cv_upload(f.c.data, _ptr(a), a.nbytes, stream), "cv_upload")
        sync(stream)
        return f

    @property
    def full(self):
        return self.c.full_window

    @property
    def current(self):
        return self.c.current_window

    def download(self, stream=None):
        a = np.empty((self.full.height, self.full.width, 4), dtype=np.float32)
        if a.nbytes:
            check(lib().cv_download(_ptr(a), self.c.data, a.nbytes, stream), "cv_download")
            sync(stream)
        return a

    def free(self):
        if self.c.data:
            _quiesce()
            lib().cv_frame_free_f32(C.byref(self.c), CV_MEM_DEVICE)

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class DeviceBuffer:
    """A raw device allocation (cv_malloc)."""

    def __init__(self, nbytes):
        self.ptr = C.c_void_p()
        self.nbytes = int(nbytes)
        check(lib().cv_malloc(C.byref(self.ptr), self.nbytes, CV_MEM_DEVICE), "cv_malloc")

    @classmethod
    def from_numpy(cls, array, stream=None):
        a = np.ascontiguousarray(array)
        b = cls(max(1, a.nbytes))
        if a.nbytes:
            check(lib().cv_upload(b.ptr, _ptr(a), a.nbytes, stream), "cv_upload")
            sync(stream)
        return b

    def download(self, dtype, shape, stream=None):
        a = np.empty(shape, dtype=dtype)
        if a.nbytes:
            check(lib().cv_download(_ptr(a), self.ptr, a.nbytes, stream), "cv_download")
            sync(stream)
        return a

    def zero(self, stream=None):
        check(lib().cv_memset(self.ptr, 0, self.nbytes, stream), "cv_memset")

    def free(self):
        if self.ptr:
            _quiesce()
            lib().cv_free(self.ptr, CV_MEM_DEVICE)
            self.ptr = C.c_void_p()

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class DeviceCoded:
    """An 8-bit Y'CbCr picture in HBM (cv_coded).  `planes` are host arrays of shape (h, stride)."""

    def __init__(self, planes, width, height, sub_x=2, layout=CV_LAYOUT_PLANAR, matrix=CV_MATRIX_709, pic=(0, 0), zero=False):
        self.buffers = []
        self.shapes = [p.shape for p in planes]
        self.c = Coded()
        for i, p in enumerate(planes):
            p = np.ascontiguousarray(p, dtype=np.uint8)
            b = DeviceBuffer.from_numpy(p)
            if zero:
                b.zero()
            self.buffers.append(b)
            self.c.plane[i] = b.ptr.value
            self.c.stride[i] = p.shape[1]
        self.c.layout, self.c.width, self.c.height = layout, width, height
        self.c.sub_x, self.c.matrix, self.c.pic_x, self.c.pic_y = sub_x, matrix, pic[0], pic[1]

    def download_planes(self):
        sync()
        return [b.download(np.uint8, s) for b, s in zip(self.buffers, self.shapes)]


def _source(obj):
    s = Source()
    if obj is None:
        s.kind = CV_SRC_NONE
    elif isinstance(obj, DeviceFrame16):
        s.kind, s.f16 = CV_SRC_F16, obj.c
    elif isinstance(obj, DeviceFrame32):
        s.kind, s.f32 = CV_SRC_F32, obj.c
    elif isinstance(obj, DeviceCoded):
        s.kind, s.coded = CV_SRC_CODED, obj.c
    else:
        raise TypeError("not a device source: %r" % (obj,))
    return s


def make_job(out=None, a=None, b=None, mix_b=0.0, layers=(), out_f32=None, out_u8=None, u8_table=None, full=None):
    """Builds a cv_frame_job.  layers: DeviceFrame16, (DeviceFrame16, mix) or (DeviceFrame16, mix, ops) bottom -> top;
    ops = dict(gain=, offset=, weave=DeviceFrame16 supplying the even rows) or None: the layer's fused filters (cv_layer_ops)."""
    job = FrameJob()
    job.a, job.b, job.mix_b = _source(a), _source(b), mix_b
    if len(layers) > CV_MAX_LAYERS:
        raise ValueError("a frame job fuses at most %d layers" % CV_MAX_LAYERS)
    job.layer_count = len(layers)
    keep_ops = []
    for i, entry in enumerate(layers):
        frame, mix, ops = (tuple(entry) + (None,))[:3] if isinstance(entry, tuple) else (entry, 1.0, None)
        job.layers[i].frame, job.layers[i].mix = frame.c, mix
        if ops:
            o = LayerOps()
            if "gain" in ops:
                o.flags |= CV_OP_GAIN_OFFSET
                o.gain, o.offset = ops["gain"], ops.get("offset", 0.0)
            if ops.get("weave") is not None:
                o.flags |= CV_OP_WEAVE
                o.even_rows = ops["weave"].c
            job.layers[i].ops = C.pointer(o)
            keep_ops.append((o, ops))
    if out is not None:
        job.out = C.pointer(out.c)
    if out_f32 is not None:
        job.out_f32 = C.pointer(out_f32.c)
    if out_u8 is not None:
        job.out_u8 = out_u8.ptr if isinstance(out_u8, DeviceBuffer) else out_u8
        job.u8_table = u8_table.ptr if isinstance(u8_table, DeviceBuffer) else u8_table
    if full is not None:
        job.full_window = Box2i(*full)
    job._keep = (out, out_f32, a, b, layers, out_u8, u8_table, keep_ops)
    return job


def render_frames(jobs, stream=None):
    """cv_render_frames on a list of jobs (one launch per kernel shape)."""
    arr = (FrameJob * len(jobs))(*jobs)
    check(lib().cv_render_frames(arr, len(jobs), stream), "cv_render_frames")
    # ctypes copied the structs: write the result windows back into the callers' jobs
    for src, dst in zip(arr, jobs):
        dst.current_window = src.current_window
    return arr


def composite_over(out, layers, stream=None):
    arr = (Layer * max(1, len(layers)))()
    for i, entry in enumerate(layers):
        frame, mix = entry if isinstance(entry, tuple) else (entry, 1.0)
        arr[i].frame, arr[i].mix = frame.c, mix
    check(lib().cv_composite_over_f16(C.byref(out.c), arr, len(layers), stream), "cv_composite_over_f16")


def mix_cross(out, a, b, mix_b, stream=None):
    check(lib().cv_mix_cross_f16(C.byref(out.c), C.byref(a.c), C.byref(b.c), mix_b, stream), "cv_mix_cross_f16")


def convert_ycbcr8(out, coded, stream=None):
    check(lib().cv_convert_ycbcr8(C.byref(out.c), C.byref(coded.c), stream), "cv_convert_ycbcr8")


def subsample_ycbcr8(frame, coded, stream=None):
    check(lib().cv_subsample_ycbcr8(C.byref(frame.c), C.byref(coded.c), stream), "cv_subsample_ycbcr8")


def subsample_mpeg2(frame, coded, stream=None):
    check(lib().cv_subsample_mpeg2(C.byref(frame.c), C.byref(coded.c), stream), "cv_subsample_mpeg2")


def display_table(rendering_intent=1.25):
    p = C.c_void_p()
    check(lib().cv_upload_display_table(rendering_intent, C.byref(p)), "cv_upload_display_table")
    return p


class Timer:
    """CUDA-event timing on the stream the kernels are launched on."""

    def __init__(self, stream=None):
        self.stream = stream
        self.start, self.stop = C.c_void_p(), C.c_void_p()
        check(lib().cv_event_create(C.byref(self.start)), "cv_event_create")
        check(lib().cv_event_create(C.byref(self.stop)), "cv_event_create")

    def begin(self):
        check(lib().cv_event_record(self.start, self.stream), "cv_event_record")

    def end(self):
        check(lib().cv_event_record(self.stop, self.stream), "cv_event_record")

    def elapsed_ms(self):
        check(lib().cv_event_sync(self.stop), "cv_event_sync")
        ms = C.c_float()
        check(lib().cv_event_elapsed_ms(self.start, self.stop, C.byref(ms)), "cv_event_elapsed_ms")
        return ms.value
