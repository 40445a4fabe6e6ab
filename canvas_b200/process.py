"""The fluggo.media.process type surface on the B200 path.

Same names, constructor arguments, attributes and error behaviour as the reference's CPython
extension (/root/reference/src/process/*.c), so its tests read the same with
`from canvas_b200 import process`.  What differs is where frames live and who computes them:
a source does not fill a caller's host buffer, it DESCRIBES its frame (a device-resident f16 / f32
frame, an 8-bit coded picture, a crossfade of two descriptions, a stack of layers); the description of
a whole graph -- DVReconstructionFilter -> VideoMixFilter -> VideoWorkspace -- is then evaluated by ONE
fused launch of libcanvas_b200.so (cv_render_frames), with the reference's intermediate roundings
(f16 after a conversion, f32 everywhere else, one truncation when an f16 frame is asked for).
No pixel is computed on the CPU: numpy only carries host copies of results.

Reference files followed: VideoSource.c:26-36, RgbaFrameF16.c:100-260, RgbaFrameF32.c,
SolidColorVideoSource.c:31-103, EmptyVideoSource.c, VideoPassThroughFilter.c:44-102,
VideoSequence.c:87-122,160-300, VideoMixFilter.c:39-75, VideoWorkspace.c:205-262,406-470,
DVReconstructionFilter.c:34-75, DVSubsampleFilter.c:32-75, CodedImageSource.c:50-274,
basicframefuncs.c:75-330, AnimationFunc.c:23-446, main.c:150-268.
"""
import bisect
import ctypes as C
import threading
import time

import numpy as np

from . import _abi, _cbases, device as dev
from ._abi import Box2i, check
from .basetypes import box2i, rgba, v2i

# The C base types of the CPython extension (fluggo.media.process, canvas_b200/csrc/py/py_process.c): they carry the
# capsules C consumers bind to.  Plain Python bases when the extension has not been built.
if _cbases.VideoSource is None:
    _cbases.load()
_VideoSourceBase = _cbases.VideoSource or object
_CodedImageSourceBase = _cbases.CodedImageSource or object
_FrameFunctionBase = _cbases.FrameFunction or object

POINT_HOLD, POINT_LINEAR = 0, 1                  # AnimationFunc.c:23-24

_INT_MIN, _INT_MAX = -2 ** 31, 2 ** 31 - 1


def _box(obj):
    """box2i / 4-sequence / ((x, y), (x, y)) -> box2i"""
    if isinstance(obj, box2i):
        return obj
    seq = tuple(obj)
    if len(seq) == 4:
        return box2i(*seq)
    if len(seq) == 2:
        return box2i(seq[0], seq[1])
    raise TypeError("expected a box2i")


def _lround(v):
    """C lround: halfway cases away from zero (basicframefuncs.c:316-326)."""
    return int(np.floor(abs(v) + 0.5) * (1 if v >= 0 else -1))


def _clip(value, lo, hi):
    return min(max(value, lo), hi)


# ----------------------------------------------------------------------------- frame functions

class FrameFunction(_FrameFunctionBase):
    """Base of the frame functions: get_values(frames) -> list of 4-tuples (FrameFunction.c / pyframework.h:69-74).
    Derived from the extension's C type, so every frame function carries the `_frame_function_funcs` capsule."""

    def _values(self, frame):
        raise NotImplementedError

    def get_values(self, frames):
        if isinstance(frames, (int, float)):
            return [tuple(self._values(float(frames)))]
        return [tuple(self._values(float(f))) for f in frames]


class FrameFuncPassThroughFilter(FrameFunction):
    """FrameFuncPassThroughFilter(source, offset=0.0): the source's values at frame + offset; a constant source
    (number, tuple, box) answers with the constant (FrameFuncPassThroughFilter.c:33-75)."""

    def __init__(self, source, offset=0.0):
        self.offset = float(offset)
        self._holder = _Holder(source)
        self._source = source

    def source(self):
        return self._source if self._holder.func is not None else None

    def set_source(self, source):
        self._holder, self._source = _Holder(source), source

    def _values(self, frame):
        return self._holder.values(frame + self.offset) if self._holder.func is not None else tuple(self._holder.constant)


class LinearFrameFunc(FrameFunction):
    def __init__(self, a, b):
        self.a, self.b = float(np.float32(a)), float(np.float32(b))

    def _values(self, frame):
        return (float(np.float32(frame * self.a + self.b)), 0.0, 0.0, 0.0)        # basicframefuncs.c:83-91: float outputs


class LerpFunc(FrameFunction):
    """LerpFunc(start, end, length): start + frame * (end - start) / length per component (basicframefuncs.c:157-165)."""

    def __init__(self, start, end, length):
        self.length = float(length)
        if self.length <= 0.0:
            raise Exception("length must be greater than 1.")
        pad = lambda seq: [float(np.float32(v)) for v in list(seq)[:4]] + [0.0] * (4 - min(4, len(list(seq))))
        self.start, self.end = pad(start), pad(end)

    def _values(self, frame):
        # the reference subtracts in float (box2f members), then divides and adds in double
        return tuple(frame * float(np.float32(e) - np.float32(s)) / self.length + s for s, e in zip(self.start, self.end))


def _parse_value(value):
    if isinstance(value, tuple):
        if not value:
            raise ValueError("An empty tuple was passed.")
        if len(value) > 4:
            raise ValueError("One of the tuples passed has more than four entries (%d)." % len(value))
        return [float(v) for v in value] + [0.0] * (4 - len(value))
    return [float(value), 0.0, 0.0, 0.0]


class AnimationPoint:
    def __init__(self, type, frame, value):
        self.type = int(type)
        self._frame = float(frame)
        self.values = _parse_value(value)
        self._owner = None

    @property
    def frame(self):
        return self._frame

    @frame.setter
    def frame(self, value):
        self._frame = float(value)
        if self._owner is not None:
            self._owner._resort()

    @property
    def value(self):
        return tuple(self.values)

    @value.setter
    def value(self, v):
        self.values = _parse_value(v)


class AnimationFunc(FrameFunction):
    """Piecewise hold / linear function of the frame (AnimationFunc.c:410-460)."""

    def __init__(self):
        self._points = []
        self._lock = threading.RLock()

    def add(self, *args):
        point = args[0] if len(args) == 1 and isinstance(args[0], AnimationPoint) else AnimationPoint(*args)
        with self._lock:
            point._owner = self
            keys = [p._frame for p in self._points]
            self._points.insert(bisect.bisect_right(keys, point._frame), point)
        return point

    def remove(self, point):
        with self._lock:
            self._points.remove(point)
            point._owner = None

    def _resort(self):
        with self._lock:
            self._points.sort(key=lambda p: p._frame)

    def __len__(self):
        return len(self._points)

    def __getitem__(self, i):
        return self._points[i]

    def _values(self, frame):
        with self._lock:
            pts = self._points
            if not pts:
                return (0.0, 0.0, 0.0, 0.0)
            i = bisect.bisect_right([p._frame for p in pts], frame) - 1
            if i < 0:
                return tuple(pts[0].values)
            left = pts[i]
            right = pts[i + 1] if i + 1 < len(pts) else None
            if right is None or left.type == POINT_HOLD:
                return tuple(left.values)
            if left.type == POINT_LINEAR:
                distance = right._frame - left._frame
                return tuple((r * (frame - left._frame) + l * (right._frame - frame)) / distance
                             for l, r in zip(left.values, right.values))
            return (0.0, 0.0, 0.0, 0.0)


class _Holder:
    """FrameFunctionHolder (basicframefuncs.c:180-275): a constant (box, tuple, number) or a frame function."""

    def __init__(self, source, default=(0.0, 0.0, 0.0, 0.0)):
        self.func, self.constant = None, list(default)
        if source is None:
            return
        if isinstance(source, FrameFunction) or hasattr(source, "get_values"):
            self.func = source
        elif hasattr(source, "_frame_function_funcs") and _cbases.cmodule is not None:
            self.func = _ForeignFrameFunction(source)           # another extension's function, through its capsule
        elif isinstance(source, box2i):
            self.constant = [float(v) for v in source.flat()]
        elif isinstance(source, tuple):
            self.constant = _parse_value(source)
        else:
            try:
                self.constant = [float(source), 0.0, 0.0, 0.0]
            except (TypeError, ValueError):
                raise Exception("The source didn't have an acceptable _frame_function_funcs attribute.")

    def values(self, frame):
        return self.func.get_values(float(frame))[0] if self.func is not None else tuple(self.constant)

    def f32(self, frame):
        return float(np.float32(self.values(frame)[0]))

    def box(self, frame):
        v = self.values(frame)
        return box2i(_lround(v[0]), _lround(v[1]), _lround(v[2]), _lround(v[3]))

    def rgba_f32(self, frame):
        return [float(np.float32(c)) for c in self.values(frame)]


class _ForeignFrameFunction:
    """A frame function that only offers the C interface (a `_frame_function_funcs` capsule, pyframework.h:69-93)."""

    def __init__(self, obj):
        self.obj = obj

    def get_values(self, frames):
        if isinstance(frames, (int, float)):
            frames = [frames]
        return _cbases.cmodule.framefunc_values(self.obj, [float(f) for f in frames])


# ----------------------------------------------------------------------------- descriptions

class _Desc:
    """What a source contributes at one frame: kind in {'f16', 'f32', 'coded', 'cross', 'stack'} (None = nothing).
    An 'f16' leaf may carry `ops`: per-pixel filters still to be applied to it (gain/offset, the field weave) -- evaluated
    inside the launch that consumes the frame when it ends up as a workspace layer (cv_layer_ops), else by _plain()."""
    __slots__ = ("kind", "frame", "a", "b", "mix", "layers", "ops")

    def __init__(self, kind, frame=None, a=None, b=None, mix=0.0, layers=(), ops=None):
        self.kind, self.frame, self.a, self.b, self.mix, self.layers, self.ops = kind, frame, a, b, mix, list(layers), ops

    def layer(self):
        """this f16 leaf as an entry of make_job(layers=...)"""
        return (self.frame, 1.0, self.ops) if self.ops else self.frame


def _fuse_filters():
    """CV_PY_NO_FUSED_FILTERS=1 keeps the per-pixel filters as separate passes (for A/B tests)"""
    import os
    return os.environ.get("CV_PY_NO_FUSED_FILTERS", "0") in ("", "0")


def _flat(window):
    return _box(window).flat()


def _leaf(desc, full):
    """A description usable as a bottom source of one job: leaves stay, composites are rendered to f32 first
    (the reference pulls nested filters through get_frame_32, main.c:106-145: no rounding in between)."""
    if desc is None or (desc.kind in ("f16", "f32", "coded") and not desc.ops):
        return desc
    out = dev.DeviceFrame32(full)
    _run(desc, full, out_f32=out)
    return _Desc("f32", out) if not out.current.empty else None


def _run(desc, full, out=None, out_f32=None):
    """Evaluates `desc` into a device frame with as few launches as its shape needs."""
    target = out if out is not None else out_f32
    if desc is None:
        target.c.current_window = Box2i()
        return
    if desc.kind == "stack":
        bottom, layers = desc.a, desc.layers
        # fused when everything above the bottom is a plain f16 frame (cv_frame_job's shape)
        if all(l is not None and l.kind == "f16" for l in layers) and len(layers) <= _abi.CV_MAX_LAYERS:
            entries = [l.layer() for l in layers]
            if bottom is not None and bottom.kind == "f16" and bottom.ops and len(entries) < _abi.CV_MAX_LAYERS:
                # a filtered bottom frame: the first layer of a layers-only job (copied with its filters applied)
                entries, bottom = [bottom.layer()] + entries, None
            bottom = bottom if (bottom is None or bottom.kind != "stack") else _leaf(bottom, full)
            a, b, mix = (bottom.a, bottom.b, bottom.mix) if bottom is not None and bottom.kind == "cross" else (bottom, None, 0.0)
            a, b = _leaf(a, full), _leaf(b, full)
            job = dev.make_job(out=out, out_f32=out_f32, a=a.frame if a else None, b=b.frame if b else None, mix_b=mix, layers=entries)
            if bottom is not None and bottom.kind == "cross" and (a is None or b is None):
                # video_mix_cross_f32 with one side empty is copy_frame_alpha of the other (video_mix.c:111-118):
                # keep both slots so the library applies the same rule
                job = _cross_job(out, out_f32, a, b, mix, full, entries)
            dev.render_frames([job])
            target.c.current_window = job.current_window
            return
        # general stack: f32 accumulator, one "over" per layer (workspace.c:530-547)
        acc = dev.DeviceFrame32(full)
        _run(bottom, full, out_f32=acc)
        for layer in layers:
            if layer is None:
                continue
            if acc.current.empty:
                _run(layer, full, out_f32=acc)
                continue
            temp = dev.DeviceFrame32(full)
            _run(layer, full, out_f32=temp)
            check(dev.lib().cv_mix_over_f32(C.byref(acc.c), C.byref(temp.c), 1.0, None), "cv_mix_over_f32")
        _run(_Desc("f32", acc) if not acc.current.empty else None, full, out=out, out_f32=out_f32)
        return
    if desc.kind == "cross":
        a, b = _leaf(desc.a, full), _leaf(desc.b, full)
        job = _cross_job(out, out_f32, a, b, desc.mix, full, [])
        dev.render_frames([job])
        target.c.current_window = job.current_window
        return
    if desc.ops:       # a filtered frame on its own: a one-layer job, the filters run inside it
        job = dev.make_job(out=out, out_f32=out_f32, layers=[desc.layer()])
    else:
        job = dev.make_job(out=out, out_f32=out_f32, a=desc.frame)
    dev.render_frames([job])
    target.c.current_window = job.current_window


def _cross_job(out, out_f32, a, b, mix, full, layers):
    """VideoMixFilter: video_mix_cross_f32_pull (video_mix.c:46-71).  An absent side is an empty frame."""
    def side(d):
        if d is not None:
            return d.frame
        empty = dev.DeviceFrame16(full, current=(0, 0, -1, -1))
        return empty
    return dev.make_job(out=out, out_f32=out_f32, a=side(a), b=side(b), mix_b=mix, layers=layers)


# ----------------------------------------------------------------------------- frames

def _describe_of(source, frame_index, full):
    """The description of ANY video source: one of this module's (its own _describe), or a foreign object that only
    offers the C interface -- a `_video_frame_source_funcs` capsule (pyframework.h:52) -- which is pulled through its
    vtable into host memory (py_video_take_source + video_get_frame_f16, GIL released) and sent to the device."""
    if isinstance(source, VideoSource):
        return source._describe(frame_index, full)
    if _cbases.cmodule is None or not hasattr(source, "_video_frame_source_funcs"):
        raise Exception("The source didn't have an acceptable _video_frame_source_funcs attribute.")
    pixels, cw = _cbases.cmodule.pull_video_f16(source, int(frame_index), tuple(full))
    if cw[2] < cw[0] or cw[3] < cw[1]:
        return None
    box = _box(full)
    data = np.frombuffer(pixels, np.uint16).reshape(box.height, box.width, 4)
    return _Desc("f16", dev.DeviceFrame16.from_numpy(data, box.flat(), cw))


class VideoSource(_VideoSourceBase):
    """Base type: get_frame_f16 / get_frame_f32 (VideoSource.c:26-36).  Derived from the extension's C type, so every
    source carries the `_video_frame_source_funcs` capsule; _pull_f16 / _pull_f32 are what its vtable calls."""

    def _describe(self, frame_index, full):
        raise NotImplementedError

    def _pull_f16(self, frame_index, window):
        frame = self.get_frame_f16(frame_index, window)
        return frame.data, frame.current_window.flat()

    def _pull_f32(self, frame_index, window):
        frame = self.get_frame_f32(frame_index, window)
        return frame.data, frame.current_window.flat()

    def get_frame_f16(self, frame_index, data_window, force_gl=False):
        full = _flat(data_window)
        out = dev.DeviceFrame16(full)
        _run(self._describe(int(frame_index), full), full, out=out)
        dev.sync()
        return RgbaFrameF16._from_device(out)

    def get_frame_f32(self, frame_index, data_window, force_gl=False):
        full = _flat(data_window)
        out = dev.DeviceFrame32(full)
        _run(self._describe(int(frame_index), full), full, out_f32=out)
        dev.sync()
        return RgbaFrameF32._from_device(out)


class _HostFrame(VideoSource):
    def __init__(self, full_window, current_window=None):
        self.full_window = _box(full_window)
        self.current_window = _box(current_window) if current_window is not None else box2i()

    def _index(self, x, y):
        if not (self.current_window.min.x <= x <= self.current_window.max.x and self.current_window.min.y <= y <= self.current_window.max.y):
            return None                                     # RgbaFrameF16.c:106-109: outside the defined window
        return (y - self.full_window.min.y, x - self.full_window.min.x)


class RgbaFrameF16(_HostFrame):
    """A host copy of an f16 RGBA frame (raw half bits, numpy uint16 (h, w, 4)); also a VideoSource (RgbaFrameF16.c:27-60)."""

    def __init__(self, full_window, current_window=None, data=None):
        super().__init__(full_window, current_window)
        shape = (self.full_window.height, self.full_window.width, 4)
        self.data = np.zeros(shape, np.uint16) if data is None else np.ascontiguousarray(data, np.uint16).reshape(shape)

    @classmethod
    def _from_device(cls, frame):
        f = cls(frame.full.tuple(), frame.current.tuple())
        if frame.full.width and frame.full.height:
            f.data = frame.download()
        return f

    def pixel(self, x, y):
        at = self._index(x, y)
        if at is None:
            return None
        return rgba(*[float(v) for v in self.data[at].view(np.float16)])

    def to_argb32_bytes(self):
        """Premultiplied ARGB32 of current_window through the gamma-0.45 ramp (RgbaFrameF16.c:114-149), on the GPU."""
        cw = self.current_window
        if cw.empty:
            return b""
        frame = dev.DeviceFrame16.from_numpy(self.data, self.full_window.flat(), cw.flat())
        out = dev.DeviceBuffer(cw.width * cw.height * 4)
        check(dev.lib().cv_to_argb32(C.byref(frame.c), out.ptr, None), "cv_to_argb32")
        return out.download(np.uint32, (cw.height, cw.width)).tobytes()

    def _describe(self, frame_index, full):
        if self.current_window.empty:
            return None
        return _Desc("f16", dev.DeviceFrame16.from_numpy(self.data, self.full_window.flat(), self.current_window.flat()))


class RgbaFrameF32(_HostFrame):
    def __init__(self, full_window, current_window=None, data=None):
        super().__init__(full_window, current_window)
        shape = (self.full_window.height, self.full_window.width, 4)
        self.data = np.zeros(shape, np.float32) if data is None else np.ascontiguousarray(data, np.float32).reshape(shape)

    @classmethod
    def _from_device(cls, frame):
        f = cls(frame.full.tuple(), frame.current.tuple())
        if frame.full.width and frame.full.height:
            f.data = frame.download()
        return f

    def pixel(self, x, y):
        at = self._index(x, y)
        if at is None:
            return None
        return rgba(*[float(v) for v in self.data[at]])

    def _describe(self, frame_index, full):
        if self.current_window.empty:
            return None
        return _Desc("f32", dev.DeviceFrame32.from_numpy(self.data, self.full_window.flat(), self.current_window.flat()))


# ----------------------------------------------------------------------------- sources and filters

class EmptyVideoSource(VideoSource):
    def _describe(self, frame_index, full):
        return None


class SolidColorVideoSource(VideoSource):
    """SolidColorVideoSource(color[, window]); both may be frame functions (SolidColorVideoSource.c:31-103)."""

    def __init__(self, color, window=None):
        self._color = _Holder(color)
        self._window = _Holder(window, default=(_INT_MIN, _INT_MIN, _INT_MAX, _INT_MAX))

    def _describe(self, frame_index, full):
        w = self._window.box(frame_index)
        window = Box2i(_clip(w.min.x, _INT_MIN, _INT_MAX), _clip(w.min.y, _INT_MIN, _INT_MAX),
                       _clip(w.max.x, _INT_MIN, _INT_MAX), _clip(w.max.y, _INT_MIN, _INT_MAX))
        color = (C.c_float * 4)(*self._color.rgba_f32(frame_index))
        # native in both precisions: the f32 colour stays exact for f32 consumers, f16 consumers get its truncation
        frame = dev.DeviceFrame32(full)
        check(dev.lib().cv_solid_frame_f32(C.byref(frame.c), C.byref(window), color, None), "cv_solid_frame_f32")
        return None if frame.current.empty else _Desc("f32", frame)


class VideoPassThroughFilter(VideoSource):
    """source shifted by `offset`, empty outside [start_frame, end_frame) (VideoPassThroughFilter.c:44-102)."""

    def __init__(self, source, offset=0, start_frame=None, end_frame=None):
        self._lock = threading.RLock()
        self._source = source
        self.offset, self.start_frame, self.end_frame = int(offset), start_frame, end_frame

    def source(self):
        return self._source

    def set_source(self, source):
        with self._lock:
            self._source = source

    def _describe(self, frame_index, full):
        with self._lock:
            if self.start_frame is not None and frame_index < self.start_frame:
                return None
            if self.end_frame is not None and frame_index >= self.end_frame:
                return None
            if self._source is None:
                return None
            return _describe_of(self._source, frame_index + self.offset, full)


class VideoSequence(VideoSource):
    """A list of (source, offset, length) played one after the other (VideoSequence.c:87-122, 160-300)."""

    def __init__(self):
        self._elements = []
        self._lock = threading.RLock()

    def __len__(self):
        return len(self._elements)

    def __getitem__(self, i):
        return self._elements[i]

    def __setitem__(self, i, value):
        with self._lock:
            self._elements[i] = self._parse(value)

    @staticmethod
    def _parse(value):
        source, offset, length = value
        if int(length) < 1:
            raise ValueError("Length cannot be less than one.")
        return (source, int(offset), int(length))

    def insert(self, i, value):
        with self._lock:
            self._elements.insert(i, self._parse(value))

    def append(self, value):
        with self._lock:
            self._elements.append(self._parse(value))

    def get_start_frame(self, i):
        return sum(e[2] for e in self._elements[:i])

    def _describe(self, frame_index, full):
        with self._lock:
            if frame_index < 0:
                return None
            start = 0
            for source, offset, length in self._elements:
                if start <= frame_index < start + length:
                    return _describe_of(source, frame_index - start + offset, full) if source is not None else None
                start += length
            return None                                             # before frame 0 or past the end (VideoSequence.c:59-83)


class VideoMixFilter(VideoSource):
    """VideoMixFilter(src_a, src_b, mix_b): crossfade, mix_b a number or a frame function (VideoMixFilter.c:39-75)."""

    def __init__(self, src_a, src_b, mix_b):
        self.src_a, self.src_b = src_a, src_b
        self._mix = _Holder(mix_b)

    def _describe(self, frame_index, full):
        mix_b = _clip(self._mix.f32(frame_index), 0.0, 1.0)
        # video_mix_cross_f32_pull short-circuits the ends (video_mix.c:50-55)
        if mix_b == 0.0:
            return _describe_of(self.src_a, frame_index, full)
        if mix_b == 1.0:
            return _describe_of(self.src_b, frame_index, full)
        return _Desc("cross", a=_describe_of(self.src_a, frame_index, full), b=_describe_of(self.src_b, frame_index, full), mix=mix_b)


class VideoWorkspaceItem:
    def __init__(self, workspace, source, offset, x, length, z, tag):
        self._workspace = workspace
        self.x, self.length, self.z, self.offset, self.source, self.tag = int(x), int(length), int(z), int(offset), source, tag
        self._serial = workspace._next_serial()

    def update(self, **kw):
        """update(x=, length=, z=, offset=, source=, tag=) (VideoWorkspace.c:229-262)."""
        for key in kw:
            if key not in ("x", "length", "z", "offset", "source", "tag"):
                raise TypeError("'%s' is an invalid keyword argument for this function" % key)
        with self._workspace._lock:
            for key, value in kw.items():
                setattr(self, key, value if key in ("source", "tag") else int(value))


class VideoWorkspace(VideoSource):
    """Items (source, offset, x, length, z, tag) composited bottom-up by z (VideoWorkspace.c:406-470, workspace.c:494-550)."""

    def __init__(self):
        self._items = []
        self._lock = threading.RLock()
        self._serial = 0

    def _next_serial(self):
        self._serial += 1
        return self._serial

    def add(self, source, offset=0, x=0, length=0, z=0, tag=None):
        with self._lock:
            item = VideoWorkspaceItem(self, source, offset, x, length, z, tag)
            self._items.append(item)
            return item

    def remove(self, item):
        with self._lock:
            if item not in self._items:
                raise ValueError("item is not in this workspace")
            self._items.remove(item)

    def __len__(self):
        return len(self._items)

    def __getitem__(self, i):
        with self._lock:
            return sorted(self._items, key=lambda it: (it.x, it.z, it._serial))[i]       # the reference's left-sorted order

    def _describe(self, frame_index, full):
        with self._lock:
            showing = [it for it in self._items if it.x <= frame_index < it.x + it.length]
            showing.sort(key=lambda it: (it.z, -it._serial))       # bottom first; later-added lower among equals
            descs = [_describe_of(it.source, frame_index - it.x + it.offset, full) if it.source is not None else None for it in showing]
        descs = [d for d in descs if d is not None]
        if not descs:
            return None
        if len(descs) == 1:
            return descs[0]
        return _Desc("stack", a=descs[0], layers=descs[1:])


class VideoScaler(VideoSource):
    """VideoScaler(source, target_point, source_point, scale_factors, source_rect): bilinear (triangle) resampling that
    maps source_point onto target_point; every argument may be a frame function (VideoScaler.c:33-78, video_scale.c:287-319)."""

    def __init__(self, source, target_point, source_point, scale_factors, source_rect):
        self.source = source
        self._target_point, self._source_point = _Holder(target_point), _Holder(source_point)
        self._factors, self._rect = _Holder(scale_factors), _Holder(source_rect)

    def _describe(self, frame_index, full):
        if self.source is None:
            return None
        v2 = lambda holder: tuple(float(np.float32(v)) for v in holder.values(frame_index)[:2])
        tp, sp, fac = v2(self._target_point), v2(self._source_point), v2(self._factors)
        rect = self._rect.box(frame_index)
        if fac[0] == 0.0 or fac[1] == 0.0:
            return None
        if fac == (1.0, 1.0) and tp == sp:
            return _describe_of(self.source, frame_index, full)
        # the part of the source the target can see, computed with C's (int) truncation (video_scale.c:301-307)
        f32, trunc = np.float32, lambda v: int(np.trunc(v))
        want = box2i(trunc(f32(sp[0]) - (f32(tp[0]) - f32(full[0])) / f32(fac[0])) - 1,
                     trunc(f32(sp[1]) - (f32(tp[1]) - f32(full[1])) / f32(fac[1])) - 1,
                     trunc(f32(sp[0]) + (f32(full[2]) - f32(tp[0])) / f32(fac[0])) + 1,
                     trunc(f32(sp[1]) + (f32(full[3]) - f32(tp[1])) / f32(fac[1])) + 1)
        fetch = (max(want.min.x, rect.min.x), max(want.min.y, rect.min.y), min(want.max.x, rect.max.x), min(want.max.y, rect.max.y))
        if fetch[2] < fetch[0] or fetch[3] < fetch[1]:
            return None
        src = dev.DeviceFrame32(fetch)
        _run(_describe_of(self.source, frame_index, fetch), fetch, out_f32=src)
        out = dev.DeviceFrame32(full)
        V = _abi.V2f
        check(dev.lib().cv_scale_bilinear_f32(C.byref(out.c), V(*tp), C.byref(src.c), V(*sp), V(*fac), None), "cv_scale_bilinear_f32")
        return None if out.current.empty else _Desc("f32", out)


class VideoGainOffsetFilter(VideoSource):
    """VideoGainOffsetFilter(source, gain, offset): rgb * gain + offset, alpha untouched (VideoGainOffsetFilter.c:66-82;
    GL-only in the reference -- here an f32 pass on the device)."""

    def __init__(self, source, gain, offset):
        self.source = source
        self._gain, self._offset = _Holder(gain), _Holder(offset)

    def _describe(self, frame_index, full):
        if self.source is None:
            return None
        inner = _describe_of(self.source, frame_index, full)
        if inner is not None and inner.kind == "f16" and not (inner.ops and "gain" in inner.ops) and _fuse_filters():
            # an f16 frame: the filter rides on it and is evaluated in f32 inside the launch that consumes the frame
            # (cv_layer_ops; the same bits as the separate pass below, tests/test_gpu_layer_ops.py)
            ops = dict(inner.ops or {}, gain=self._gain.f32(frame_index), offset=self._offset.f32(frame_index))
            return _Desc("f16", inner.frame, ops=ops)
        frame = dev.DeviceFrame32(full)
        _run(inner, full, out_f32=frame)
        if frame.current.empty:
            return None
        check(dev.lib().cv_gain_offset_f32(C.byref(frame.c), self._gain.f32(frame_index), self._offset.f32(frame_index), None),
              "cv_gain_offset_f32")
        return _Desc("f32", frame)


class Pulldown23RemovalFilter(VideoSource):
    """Pulldown23RemovalFilter(source, offset): 30 -> 24 fps; three of four output frames are whole source frames, the
    fourth weaves the odd rows of one with the even rows of the next (Pulldown23RemovalFilter.c:46-105)."""

    def __init__(self, source, offset):
        self.source, self.offset = source, int(offset)

    def _describe(self, frame_index, full):
        if self.source is None:
            return None
        phase = (frame_index + 3) & 3 if self.offset == 4 else (frame_index + self.offset) & 3
        base = ((frame_index + self.offset) >> 2) * 5 - self.offset
        if phase != 2:
            return _describe_of(self.source, base + {0: 0, 1: 1, 3: 4}[phase], full)
        first = _describe_of(self.source, base + 2, full)
        if first is not None and first.kind == "f16" and not first.ops and _fuse_filters():
            # both fields are plain f16 frames: the weave becomes a row-parity select on the layer's pointer inside the
            # consuming launch -- when the second frame defines every pixel the first one contributes
            second = _describe_of(self.source, base + 3, full)
            if second is not None and second.kind == "f16" and not second.ops:
                box = _box(full)
                clip = lambda w: (max(w[0], box.min.x), max(w[1], box.min.y), min(w[2], box.max.x), min(w[3], box.max.y))
                a, b = clip(first.frame.current.tuple()), clip(second.frame.current.tuple())
                if a[2] < a[0] or a[3] < a[1]:
                    return None
                if b[0] <= a[0] and b[1] <= a[1] and b[2] >= a[2] and b[3] >= a[3]:
                    return _Desc("f16", first.frame, ops={"weave": second.frame})
        frame = dev.DeviceFrame16(full)
        _run(first, full, out=frame)
        if frame.current.empty:
            return None
        window = frame.current.tuple()
        other = dev.DeviceFrame16(window)
        _run(_describe_of(self.source, base + 3, window), window, out=other)
        # the reference copies those rows whatever the second frame defined; rows it did not define are left as they are
        check(dev.lib().cv_weave_fields_f16(C.byref(frame.c), C.byref(other.c), None), "cv_weave_fields_f16")
        return _Desc("f16", frame)


# ----------------------------------------------------------------------------- coded images (DV)

class CodedImage:
    """CodedImage(data, stride, line_count): up to four planes of bytes (CodedImageSource.c:252-274)."""

    def __init__(self, data, stride, line_count):
        self.data, self.stride, self.line_count = list(data), list(stride), list(line_count)


class CodedImageSource(_CodedImageSourceBase):
    """Base of the coded image sources; carries the `_coded_image_source_funcs` capsule (pyframework.h:132), whose
    getFrame calls _pull_coded."""

    def get_coded_image(self, frame):
        raise NotImplementedError

    def _pull_coded(self, frame):
        image = self.get_coded_image(frame)
        if image is None:
            return None
        return [(bytes(d) if d is not None else b"", s, n) for d, s, n in zip(image.data, image.stride, image.line_count) if d is not None]


def _coded_image_of(source, frame):
    """source.get_coded_image(frame), or -- for a foreign source -- coded_image_source.getFrame through its capsule."""
    if hasattr(source, "get_coded_image"):
        return source.get_coded_image(frame)
    if _cbases.cmodule is None or not hasattr(source, "_coded_image_source_funcs"):
        raise Exception("The source didn't have an acceptable _coded_image_source_funcs attribute.")
    planes = _cbases.cmodule.pull_coded(source, int(frame))
    if planes is None:
        return None
    return CodedImage([p[0] for p in planes], [p[1] for p in planes], [p[2] for p in planes])


_DV_WINDOW = (0, -1, 719, 478)                  # video_reconstruct.c:69-76, video_subsample.c:112-117


class DVReconstructionFilter(VideoSource):
    """Planar 4:1:1 Rec.709 720x480 -> linear f16 RGBA at picture offset (0, -1) (DVReconstructionFilter.c:34-75)."""

    def __init__(self, source):
        self.source = source

    def _describe(self, frame_index, full):
        image = _coded_image_of(self.source, frame_index) if self.source is not None else None
        if image is None:
            return None
        widths = (720, 180, 180)
        planes = []
        for p in range(3):
            if image.data[p] is None:
                return None
            plane = np.frombuffer(bytes(image.data[p]), np.uint8)[: image.stride[p] * image.line_count[p]]
            planes.append(plane.reshape(image.line_count[p], image.stride[p]))
            if image.stride[p] < widths[p] or image.line_count[p] < 480:
                return None
        coded = dev.DeviceCoded(planes, 720, 480, sub_x=4, layout=_abi.CV_LAYOUT_PLANAR, matrix=_abi.CV_MATRIX_709, pic=(0, -1))
        return _Desc("coded", coded)


class DVSubsampleFilter(CodedImageSource):
    """Linear f16 RGBA -> planar 4:1:1 Rec.709 studio-range bytes (DVSubsampleFilter.c:32-75, video_subsample.c:99-187)."""

    def __init__(self, source):
        self.source = source

    def get_coded_image(self, frame):
        out = dev.DeviceFrame16(_DV_WINDOW)
        _run(_describe_of(self.source, int(frame), _DV_WINDOW), _DV_WINDOW, out=out)
        planes = [np.zeros((480, w), np.uint8) for w in (720, 180, 180)]
        coded = dev.DeviceCoded(planes, 720, 480, sub_x=4, layout=_abi.CV_LAYOUT_PLANAR, matrix=_abi.CV_MATRIX_709, pic=(0, -1), zero=True)
        dev.subsample_ycbcr8(out, coded)
        got = coded.download_planes()
        return CodedImage([p.tobytes() for p in got], [720, 180, 180], [480, 480, 480])


class MPEG2SubsampleFilter(CodedImageSource):
    """Linear f16 RGBA -> planar 8-bit 4:2:0, Rec.601, 720 x 480, interlace-aware chroma (MPEG2SubsampleFilter.c:45-58,
    video_subsample.c:212-290: GLSL in the reference, a CUDA kernel here)."""

    _WINDOW = (0, 0, 719, 479)

    def __init__(self, source):
        self.source = source

    def get_coded_image(self, frame):
        out = dev.DeviceFrame16(self._WINDOW)
        _run(_describe_of(self.source, int(frame), self._WINDOW), self._WINDOW, out=out)
        planes = [np.zeros((h, w), np.uint8) for w, h in ((720, 480), (360, 240), (360, 240))]
        coded = dev.DeviceCoded(planes, 720, 480, sub_x=2, layout=_abi.CV_LAYOUT_PLANAR, matrix=_abi.CV_MATRIX_601, pic=(0, 0), zero=True)
        dev.subsample_mpeg2(out, coded)
        got = coded.download_planes()
        return CodedImage([p.tobytes() for p in got], [720, 360, 360], [480, 240, 240])


class VideoPullQueueItem:
    """The handle enqueue() returns; cancel() stops the callback (and the render, if it has not started)."""

    def __init__(self):
        self.active = True
        self._future = None

    def cancel(self):
        self.active = False
        if self._future is not None:
            self._future.cancel()


class VideoPullQueue:
    """VideoPullQueue(): frames rendered ahead of need on two worker threads, `callback(frame_index, frame, user_data)`
    when one is ready (VideoPullQueue.c:100-183).  The reference hands the callback to the glib main loop; there is no
    main loop here, so it runs on the worker thread that rendered the frame."""

    def __init__(self, device=None):
        from concurrent.futures import ThreadPoolExecutor
        self._device = dev.lib().cv_get_device() if device is None else int(device)
        self._pool = ThreadPoolExecutor(max_workers=2, thread_name_prefix="VideoPullQueue")

    def _work(self, item, source, frame_index, window, callback, user_data):
        if not item.active:
            return
        check(dev.lib().cv_set_device(self._device), "cv_set_device")       # the library's device selection is per thread
        frame = source.get_frame_f16(frame_index, window)
        if item.active:
            callback(frame_index, frame, user_data)

    def enqueue(self, source, frame_index, window, callback, user_data):
        item = VideoPullQueueItem()
        item._future = self._pool.submit(self._work, item, source, int(frame_index), window, callback, user_data)
        return item

    def close(self):
        self._pool.shutdown(wait=True)


# ----------------------------------------------------------------------------- module functions

class _Rational(C.Structure):
    _fields_ = [("n", C.c_int32), ("d", C.c_uint32)]


def _rational(rate):
    return _Rational(int(rate.numerator), int(rate.denominator))


def get_frame_time(rate, frame):
    """Time in nanoseconds of `frame` at the Fraction frame rate (main.c:252-258)."""
    fn = _abi.load().get_frame_time
    fn.restype, fn.argtypes = C.c_int64, [C.POINTER(_Rational), C.c_int]
    return fn(C.byref(_rational(rate)), int(frame))


def get_time_frame(rate, time_ns):
    fn = _abi.load().get_time_frame
    fn.restype, fn.argtypes = C.c_int, [C.POINTER(_Rational), C.c_int64]
    return fn(C.byref(_rational(rate)), int(time_ns))


def time_get_frame(source, min_frame, max_frame, data_window=(0, 0, 1, 1)):
    """Renders min_frame..max_frame into a device frame and returns the nanoseconds it took (main.c:150-200)."""
    full = _flat(data_window)
    out = dev.DeviceFrame16(full)
    dev.sync()
    start = time.perf_counter_ns()
    for frame in range(int(min_frame), int(max_frame) + 1):
        _run(_describe_of(source, frame, full), full, out=out)
    dev.sync()
    return time.perf_counter_ns() - start


# ----------------------------------------------------------------------------- fluggo.media.process
# The extension module re-exports this module's public names (the reference's surface); done here, at the end, so that
# it works whichever of the two modules was imported first.
if _cbases.cmodule is not None:
    for _name, _value in list(globals().items()):
        if not _name.startswith("_") and _name not in ("np", "C", "dev", "bisect", "threading", "time"):
            setattr(_cbases.cmodule, _name, _value)
