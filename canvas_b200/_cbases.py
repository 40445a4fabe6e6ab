"""Where the C base types of the CPython extension `fluggo.media.process` (canvas_b200/csrc/py/py_process.c) are handed
to canvas_b200.process, whichever of the two gets imported first.

The extension defines VideoSource / CodedImageSource / FrameFunction in C -- the types that carry the
`_video_frame_source_funcs` / `_coded_image_source_funcs` / `_frame_function_funcs` capsules C consumers bind to
(/root/reference/include/pyframework.h:48-132) -- and stores them here from its PyInit_process; canvas_b200.process
derives its classes from them.  Without the built extension the classes fall back to plain Python bases: everything
renders the same, only the capsule protocol is missing."""
import importlib
import os
import sys

VideoSource = CodedImageSource = FrameFunction = None
cmodule = None
loading = False


def load():
    """Imports the extension (its init fills in the names above).  Safe to call again, and from inside the extension's
    own initialisation."""
    global loading
    if cmodule is not None or loading:
        return cmodule
    loading = True
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    if root not in sys.path:
        sys.path.insert(0, root)
    try:
        importlib.import_module("fluggo.media.process")
    except ImportError:
        pass
    finally:
        loading = False
    return cmodule
