"""ctypes binding of libcanvas_b200.so (include/canvas_b200.h + include/framework.h).

This is the reference-side binding a maintainer would add: plain structs
and pointers, no torch types.  The library is loaded from inside the package directory; if it
is missing or no B200 can be initialised every entry point raises -- there is no CPU fallback.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libcanvas_b200.so")

CV_MEM_DEVICE, CV_MEM_PINNED = 0, 1
CV_LUT_709_TO_LINEAR_SCENE, CV_LUT_709_TO_LINEAR_DISPLAY, CV_LUT_LINEAR_TO_709, CV_LUT_LINEAR_TO_SRGB = 0, 1, 2, 3
CV_LAYOUT_PLANAR, CV_LAYOUT_UYVY, CV_LAYOUT_YUY2 = 0, 1, 2
CV_MATRIX_709, CV_MATRIX_601 = 0, 1
CV_SRC_NONE, CV_SRC_F16, CV_SRC_CODED, CV_SRC_F32 = 0, 1, 2, 3
CV_MAX_LAYERS = 16
CV_TIMELINE_STATIC = 1


class CanvasError(RuntimeError):
    pass


class V2i(C.Structure):
    _fields_ = [("x", C.c_int32), ("y", C.c_int32)]


class Box2i(C.Structure):
    """box2i (include/framework.h): inclusive corners, empty <=> max < min."""
    _fields_ = [("min", V2i), ("max", V2i)]

    def __init__(self, min_x=0, min_y=0, max_x=-1, max_y=-1):
        super().__init__(V2i(min_x, min_y), V2i(max_x, max_y))

    def tuple(self):
        return (self.min.x, self.min.y, self.max.x, self.max.y)

    @property
    def width(self):
        return max(0, self.max.x - self.min.x + 1)

    @property
    def height(self):
        return max(0, self.max.y - self.min.y + 1)

    @property
    def empty(self):
        return self.max.x < self.min.x or self.max.y < self.min.y

    def __repr__(self):
        return "Box2i%r" % (self.tuple(),)


class FrameF16(C.Structure):
    """rgba_frame_f16"""
    _fields_ = [("data", C.c_void_p), ("full_window", Box2i), ("current_window", Box2i)]


class FrameF32(C.Structure):
    """rgba_frame_f32"""
    _fields_ = [("data", C.c_void_p), ("full_window", Box2i), ("current_window", Box2i)]


class V2f(C.Structure):
    """v2f: passed by value"""
    _fields_ = [("x", C.c_float), ("y", C.c_float)]


class Coded(C.Structure):
    """cv_coded"""
    _fields_ = [("plane", C.c_void_p * 3), ("stride", C.c_int32 * 3), ("layout", C.c_int32),
                ("width", C.c_int32), ("height", C.c_int32), ("sub_x", C.c_int32), ("matrix", C.c_int32),
                ("pic_x", C.c_int32), ("pic_y", C.c_int32)]


CV_OP_WEAVE, CV_OP_COLOR, CV_OP_GAIN_OFFSET, CV_OP_CROP = 1, 2, 4, 8
CV_COLOR_RGB_TO_XYZ_SDTV, CV_COLOR_XYZ_TO_SRGB = 1, 2


class LayerOps(C.Structure):
    """cv_layer_ops: per-pixel filters fused into a layer"""
    _fields_ = [("flags", C.c_int32), ("color", C.c_int32), ("gain", C.c_float), ("offset", C.c_float), ("crop", Box2i),
                ("even_rows", FrameF16)]


class Layer(C.Structure):
    """cv_layer"""
    _fields_ = [("frame", FrameF16), ("mix", C.c_float), ("reserved", C.c_int32), ("ops", C.POINTER(LayerOps))]


class Source(C.Structure):
    """cv_source"""
    _fields_ = [("kind", C.c_int32), ("reserved", C.c_int32), ("f16", FrameF16), ("f32", FrameF32), ("coded", Coded)]


class FrameJob(C.Structure):
    """cv_frame_job"""
    _fields_ = [("a", Source), ("b", Source), ("mix_b", C.c_float), ("layer_count", C.c_int32),
                ("layers", Layer * CV_MAX_LAYERS), ("out", C.POINTER(FrameF16)), ("out_f32", C.POINTER(FrameF32)),
                ("out_u8", C.c_void_p), ("u8_table", C.c_void_p), ("full_window", Box2i), ("current_window", Box2i)]


class Clip(C.Structure):
    """cv_clip"""
    _fields_ = [("frames", C.POINTER(Source)), ("count", C.c_int32), ("reserved", C.c_int32)]


class CodedImageSourceFuncs(C.Structure):
    """coded_image_source_funcs (include/framework.h)"""
    _fields_ = [("flags", C.c_int), ("getFrame", C.c_void_p)]


class CodedImageSource(C.Structure):
    """coded_image_source"""
    _fields_ = [("obj", C.c_void_p), ("funcs", C.POINTER(CodedImageSourceFuncs))]


class Pull(C.Structure):
    """cv_pull: pictures pulled from the host while a timeline renders"""
    _fields_ = [("coded", C.POINTER(CodedImageSource)), ("format", Coded), ("video", C.c_void_p), ("full_window", Box2i)]


class TimelineItem(C.Structure):
    """cv_timeline_item"""
    _fields_ = [("x", C.c_int64), ("length", C.c_int64), ("z", C.c_int64), ("offset", C.c_int64),
                ("clips", C.POINTER(Clip)), ("clips_b", C.POINTER(Clip)), ("mix_first", C.c_float), ("mix_last", C.c_float),
                ("ramp_start", C.c_int64), ("ramp_length", C.c_int64), ("pull", C.POINTER(Pull)), ("pull_b", C.POINTER(Pull))]


class TimelineTarget(C.Structure):
    """cv_timeline_target"""
    _fields_ = [("out", C.POINTER(FrameF16)), ("out_u8", C.POINTER(C.c_void_p)), ("u8_table", C.c_void_p),
                ("ring", C.c_int32), ("reserved", C.c_int32)]


TIMELINE_SINK = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_int, C.c_int64, C.POINTER(FrameJob), C.c_void_p)


class TimelineRun(C.Structure):
    """cv_timeline_run"""
    _fields_ = [("items", C.POINTER(TimelineItem)), ("item_count", C.c_int32), ("n_gpus", C.c_int32),
                ("devices", C.POINTER(C.c_int32)), ("first", C.c_int64), ("last", C.c_int64), ("chunk", C.c_int32),
                ("flags", C.c_int32), ("targets", C.POINTER(TimelineTarget)), ("sink", TIMELINE_SINK), ("user", C.c_void_p)]


class TimelineStats(C.Structure):
    """cv_timeline_stats"""
    _fields_ = [("device", C.c_int32), ("reserved", C.c_int32), ("frames", C.c_int64), ("launches", C.c_uint64),
                ("device_ms", C.c_double), ("wall_ms", C.c_double), ("h2d_bytes", C.c_uint64)]


_PROTOTYPES = {
    # name: (restype, argtypes)
    "cv_device_count": (C.c_int, []),
    "cv_init": (C.c_int, [C.c_int]),
    "cv_set_device": (C.c_int, [C.c_int]),
    "cv_get_device": (C.c_int, []),
    "cv_shutdown": (None, []),
    "cv_last_error": (C.c_char_p, []),
    "cv_version": (C.c_char_p, []),
    "cv_kernel_launches": (C.c_uint64, []),
    "cv_stream_create": (C.c_int, [C.POINTER(C.c_void_p)]),
    "cv_stream_destroy": (C.c_int, [C.c_void_p]),
    "cv_stream_sync": (C.c_int, [C.c_void_p]),
    "cv_device_sync": (C.c_int, []),
    "cv_event_create": (C.c_int, [C.POINTER(C.c_void_p)]),
    "cv_event_destroy": (C.c_int, [C.c_void_p]),
    "cv_event_record": (C.c_int, [C.c_void_p, C.c_void_p]),
    "cv_event_sync": (C.c_int, [C.c_void_p]),
    "cv_event_elapsed_ms": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(C.c_float)]),
    "cv_malloc": (C.c_int, [C.POINTER(C.c_void_p), C.c_size_t, C.c_int]),
    "cv_free": (C.c_int, [C.c_void_p, C.c_int]),
    "cv_free_async": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p]),
    "cv_memset": (C.c_int, [C.c_void_p, C.c_int, C.c_size_t, C.c_void_p]),
    "cv_upload": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    "cv_download": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    "cv_copy_d2d": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    "cv_flush_l2": (C.c_int, [C.c_void_p]),
    "cv_frame_alloc_f16": (C.c_int, [C.POINTER(FrameF16), C.POINTER(Box2i), C.c_int]),
    "cv_frame_free_f16": (C.c_int, [C.POINTER(FrameF16), C.c_int]),
    "cv_frame_alloc_f32": (C.c_int, [C.POINTER(FrameF32), C.POINTER(Box2i), C.c_int]),
    "cv_frame_free_f32": (C.c_int, [C.POINTER(FrameF32), C.c_int]),
    "cv_frame_upload_f16": (C.c_int, [C.POINTER(FrameF16), C.POINTER(FrameF16), C.c_void_p]),
    "cv_frame_download_f16": (C.c_int, [C.POINTER(FrameF16), C.POINTER(FrameF16), C.c_void_p]),
    "cv_frame_upload_f32": (C.c_int, [C.POINTER(FrameF32), C.POINTER(FrameF32), C.c_void_p]),
    "cv_frame_download_f32": (C.c_int, [C.POINTER(FrameF32), C.POINTER(FrameF32), C.c_void_p]),
    "cv_host_lut": (C.POINTER(C.c_uint16), [C.c_int]),
    "cv_host_gamma45_ramp": (C.POINTER(C.c_uint8), []),
    "cv_build_intent_ramp": (C.c_int, [C.c_float, C.c_void_p]),
    "cv_half_from_float": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    "cv_half_to_float": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    "cv_half_lookup": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    "cv_transfer": (C.c_int, [C.c_int, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    "cv_convert_ycbcr8": (C.c_int, [C.POINTER(FrameF16), C.POINTER(Coded), C.c_void_p]),
    "cv_subsample_ycbcr8": (C.c_int, [C.POINTER(FrameF16), C.POINTER(Coded), C.c_void_p]),
    "cv_copy_frame_f16": (C.c_int, [C.POINTER(FrameF16), C.POINTER(FrameF16), C.c_void_p]),
    "cv_mix_cross_f16": (C.c_int, [C.POINTER(FrameF16), C.POINTER(FrameF16), C.POINTER(FrameF16), C.c_float, C.c_void_p]),
    "cv_composite_over_f16": (C.c_int, [C.POINTER(FrameF16), C.POINTER(Layer), C.c_int, C.c_void_p]),
    "cv_copy_frame_alpha_f32": (C.c_int, [C.POINTER(FrameF32), C.POINTER(FrameF32), C.c_float, C.c_void_p]),
    "cv_mix_cross_f32": (C.c_int, [C.POINTER(FrameF32), C.POINTER(FrameF32), C.POINTER(FrameF32), C.c_float, C.c_void_p]),
    "cv_mix_over_f32": (C.c_int, [C.POINTER(FrameF32), C.POINTER(FrameF32), C.c_float, C.c_void_p]),
    "cv_frame_f16_to_f32": (C.c_int, [C.POINTER(FrameF32), C.POINTER(FrameF16), C.c_void_p]),
    "cv_frame_f32_to_f16": (C.c_int, [C.POINTER(FrameF16), C.POINTER(FrameF32), C.c_void_p]),
    "cv_solid_frame_f16": (C.c_int, [C.POINTER(FrameF16), C.POINTER(Box2i), C.POINTER(C.c_float), C.c_void_p]),
    "cv_solid_frame_f32": (C.c_int, [C.POINTER(FrameF32), C.POINTER(Box2i), C.POINTER(C.c_float), C.c_void_p]),
    "cv_scale_bilinear_f32": (C.c_int, [C.POINTER(FrameF32), V2f, C.POINTER(FrameF32), V2f, V2f, C.c_void_p]),
    "cv_scale_bilinear_static_f32": (C.c_int, [C.POINTER(FrameF32), V2f, C.POINTER(FrameF32), V2f, V2f, C.c_void_p]),
    "cv_gain_offset_f32": (C.c_int, [C.POINTER(FrameF32), C.c_float, C.c_float, C.c_void_p]),
    "cv_weave_fields_f16": (C.c_int, [C.POINTER(FrameF16), C.POINTER(FrameF16), C.c_void_p]),
    "cv_color_rgb_to_xyz_sdtv": (C.c_int, [C.POINTER(FrameF16), C.c_void_p]),
    "cv_color_xyz_to_srgb": (C.c_int, [C.POINTER(FrameF16), C.c_void_p]),
    "cv_crop_f16": (C.c_int, [C.POINTER(FrameF16), C.POINTER(Box2i), C.c_void_p]),
    "cv_subsample_mpeg2": (C.c_int, [C.POINTER(FrameF16), C.POINTER(Coded), C.c_void_p]),
    "cv_upload_display_table": (C.c_int, [C.c_float, C.POINTER(C.c_void_p)]),
    "cv_encode_u8": (C.c_int, [C.POINTER(FrameF16), C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]),
    "cv_to_argb32": (C.c_int, [C.POINTER(FrameF16), C.c_void_p, C.c_void_p]),
    "cv_render_frames": (C.c_int, [C.POINTER(FrameJob), C.c_int, C.c_void_p]),
    "cv_timeline_showing": (C.c_int, [C.POINTER(TimelineItem), C.c_int, C.c_int64, C.POINTER(C.c_int), C.c_int]),
    "cv_timeline_job": (C.c_int, [C.POINTER(TimelineItem), C.c_int, C.c_int, C.c_int64, C.POINTER(FrameJob)]),
    "cv_timeline_device_of": (C.c_int, [C.c_int64, C.c_int64, C.c_int, C.c_int]),
    "cv_timeline_frames_of": (C.c_int64, [C.c_int, C.c_int64, C.c_int64, C.c_int, C.c_int]),
    "cv_timeline_render": (C.c_int, [C.POINTER(TimelineRun), C.POINTER(TimelineStats)]),
}

_lib = None


def load():
    """Loads the library (once) and sets the prototypes.  Raises CanvasError when it is missing."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise CanvasError("canvas_b200: %s is missing -- run `python -c 'import __graft_entry__ as g; g.build()'` "
                              "(there is no CPU fallback)" % LIB_PATH)
        lib = C.CDLL(LIB_PATH)
        for name, (restype, argtypes) in _PROTOTYPES.items():
            fn = getattr(lib, name)
            fn.restype, fn.argtypes = restype, argtypes
        _lib = lib
    return _lib


def exported_names():
    return sorted(_PROTOTYPES)


def check(rc, what="canvas_b200 call"):
    if rc != 0:
        message = load().cv_last_error().decode("utf-8", "replace")
        raise CanvasError("%s failed (%d): %s" % (what, rc, message))
