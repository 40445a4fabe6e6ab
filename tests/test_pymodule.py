"""The CPython extension module fluggo.media.process (canvas_b200/csrc/py/py_process.c): the capsule protocol of the
reference's extension (/root/reference/include/pyframework.h:48-132, src/process/main.c:30-66, 359-417), and the
reference's OWN test files run unmodified against it.

The reference's tests live under /root/reference/tests/process (build container) or, on the GPU box, in the git-ignored
copy __graft_entry__.build() leaves under oracle/_ref/tests (like the compiled reference itself); CANVAS_REFERENCE_TESTS
overrides both.  They are loaded from there at run time -- never copied into the repository."""
import importlib.util
import os
import subprocess
import sys
import sysconfig
import unittest

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def reference_tests_dir():
    for path in (os.environ.get("CANVAS_REFERENCE_TESTS"), "/root/reference/tests", os.path.join(ROOT, "oracle", "_ref", "tests")):
        if path and os.path.isfile(os.path.join(path, "process", "frame_func.py")):
            return path
    return None


def run_reference_test_file(relative):
    """Loads one of the reference's test files as it is and runs its unittest cases."""
    base = reference_tests_dir()
    if base is None:
        pytest.skip("the reference's tests are not available here")
    path = os.path.join(base, relative)
    spec = importlib.util.spec_from_file_location("reference_test_" + os.path.basename(relative)[:-3], path)
    module = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(module)
    suite = unittest.defaultTestLoader.loadTestsFromModule(module)
    assert suite.countTestCases() > 0
    result = unittest.TestResult()
    suite.run(result)
    problems = result.errors + result.failures
    assert not problems, "\n".join(text for _, text in problems)
    return result.testsRun


@pytest.fixture(scope="module")
def process():
    from fluggo.media import process
    return process


def test_module_is_a_c_extension_with_the_reference_surface(process):
    assert process.__file__.endswith(sysconfig.get_config_var("EXT_SUFFIX"))
    for name in ("VideoSource", "RgbaFrameF16", "RgbaFrameF32", "SolidColorVideoSource", "EmptyVideoSource", "VideoPassThroughFilter",
                 "VideoSequence", "VideoMixFilter", "VideoWorkspace", "DVReconstructionFilter", "DVSubsampleFilter", "CodedImageSource",
                 "CodedImage", "LerpFunc", "LinearFrameFunc", "AnimationFunc", "AnimationPoint", "FrameFuncPassThroughFilter", "VideoScaler",
                 "VideoPullQueue", "VideoGainOffsetFilter", "Pulldown23RemovalFilter", "MPEG2SubsampleFilter", "POINT_HOLD", "POINT_LINEAR",
                 "time_get_frame", "get_frame_time", "get_time_frame", "NS_PER_SEC"):
        assert hasattr(process, name), name
    from canvas_b200 import process as implementation
    assert process.VideoWorkspace is implementation.VideoWorkspace
    # every source type is an instance of the extension's C base type
    c_base = [t for t in process.VideoWorkspace.__mro__ if t.__module__ == "fluggo.media.process"]
    assert c_base and c_base[0].__name__ == "VideoSource"


def test_capsules_have_the_reference_names(process):
    import ctypes as C
    C.pythonapi.PyCapsule_IsValid.argtypes = [C.py_object, C.c_char_p]
    C.pythonapi.PyCapsule_IsValid.restype = C.c_int
    source = process.SolidColorVideoSource((1.0, 0.5, 0.25, 1.0))
    func = process.LerpFunc((0.0,), (1.0,), 10)

    class Pictures(process.CodedImageSource):
        def get_coded_image(self, frame):
            return None

    for obj, name in ((source, "_video_frame_source_funcs"), (process.VideoWorkspace(), "_video_frame_source_funcs"),
                      (func, "_frame_function_funcs"), (process.AnimationFunc(), "_frame_function_funcs"),
                      (Pictures(), "_coded_image_source_funcs")):
        capsule = getattr(obj, name)
        assert C.pythonapi.PyCapsule_IsValid(capsule, name.encode()) == 1, name


def test_frame_function_through_its_capsule(process):
    func = process.AnimationFunc()
    func.add(process.AnimationPoint(process.POINT_HOLD, 0.0, 4.0))
    func.add(process.AnimationPoint(process.POINT_LINEAR, 1.0, 2.0))
    func.add(process.AnimationPoint(process.POINT_LINEAR, 2.0, 6.0))
    frames = [-0.5, 0.0, 0.75, 1.0, 1.25, 1.3, 2.0, 2.5]
    assert process.framefunc_values(func, frames) == [tuple(v) for v in func.get_values(frames)]
    with pytest.raises(Exception, match="_frame_function_funcs"):
        process.framefunc_values(object(), [0.0])


def test_coded_image_source_through_its_capsule(process):
    planes = [np.arange(6 * 4, dtype=np.uint8).reshape(4, 6), np.arange(3 * 4, dtype=np.uint8).reshape(4, 3) + 100]

    class Pictures(process.CodedImageSource):
        def get_coded_image(self, frame):
            if frame < 0:
                return None
            return process.CodedImage([(p + frame).astype(np.uint8).tobytes() for p in planes], [6, 3], [4, 4])

    source = Pictures()
    got = process.pull_coded(source, 5)
    assert [(g[1], g[2]) for g in got] == [(6, 4), (3, 4)]
    assert got[0][0] == (planes[0] + 5).astype(np.uint8).tobytes() and got[1][0] == (planes[1] + 5).astype(np.uint8).tobytes()
    assert process.pull_coded(source, -1) is None


def test_take_source_rejects_objects_without_a_capsule(process):
    with pytest.raises(Exception, match="_video_frame_source_funcs"):
        process.pull_video_f16(object(), 0, (0, 0, 1, 1))


def test_reference_frame_func_tests_run_unmodified():
    """/root/reference/tests/process/frame_func.py: LerpFunc and AnimationFunc (host logic: no GPU needed)"""
    assert run_reference_test_file(os.path.join("process", "frame_func.py")) == 3


# ----------------------------------------------------------------------------- GPU

@pytest.mark.gpu
@pytest.mark.parametrize("relative,cases", [(os.path.join("process", "video", "RgbaFrameF16.py"), 1),
                                            (os.path.join("process", "video", "SolidColorVideoSource.py"), 4),
                                            (os.path.join("process", "video", "VideoWorkspace.py"), 1)])
def test_reference_video_tests_run_unmodified(relative, cases):
    assert run_reference_test_file(relative) == cases


@pytest.mark.gpu
def test_c_consumer_pulls_our_sources_through_the_capsule(process):
    """pull_video_f16 is a C consumer: py_video_take_source( obj ) + video_get_frame_f16 with the GIL released, the way an
    encoder or the pull queue holds a source (src/process/main.c:30-66; VideoPullQueue.c:102)."""
    from canvas_b200.basetypes import box2i
    solid = process.SolidColorVideoSource((0.25, 0.5, 0.75, 1.0), box2i(1, 1, 5, 3))
    over = process.SolidColorVideoSource((1.0, 0.0, 0.0, 0.5), box2i(3, 0, 7, 4))
    workspace = process.VideoWorkspace()
    workspace.add(source=solid, x=0, length=10, z=0, offset=0)
    workspace.add(source=over, x=0, length=10, z=1, offset=0)
    window = (0, 0, 7, 4)
    for source in (solid, workspace):
        want = source.get_frame_f16(2, window)
        pixels, current = process.pull_video_f16(source, 2, window)
        assert current == want.current_window.flat()
        got = np.frombuffer(pixels, np.uint16).reshape(5, 8, 4)
        x0, y0, x1, y1 = current
        assert np.array_equal(got[y0:y1 + 1, x0:x1 + 1], want.data[y0:y1 + 1, x0:x1 + 1])
        wide, current32 = process.pull_video_f32(source, 2, window)
        assert current32 == current
        want32 = source.get_frame_f32(2, window)
        assert np.array_equal(np.frombuffer(wide, np.float32).reshape(5, 8, 4)[y0:y1 + 1, x0:x1 + 1], want32.data[y0:y1 + 1, x0:x1 + 1])


@pytest.fixture(scope="module")
def foreign(tmp_path_factory):
    """another extension's source type: only the capsule, none of our classes (tests/cext/foreign_source.c)"""
    out = tmp_path_factory.mktemp("cext")
    so = os.path.join(str(out), "foreign_source" + sysconfig.get_config_var("EXT_SUFFIX"))
    subprocess.run(["gcc", "-O1", "-shared", "-fPIC", "-I", sysconfig.get_paths()["include"], "-I", os.path.join(ROOT, "include"),
                    os.path.join(ROOT, "tests", "cext", "foreign_source.c"), "-o", so], check=True)
    sys.path.insert(0, str(out))
    import foreign_source
    return foreign_source


@pytest.mark.gpu
def test_foreign_sources_are_pulled_through_their_capsule(process, foreign):
    """A source from ANOTHER extension joins our graphs: VideoMixFilter / VideoWorkspace pull it through its vtable
    (py_video_take_source), upload what it produced and composite on the device -- checked against the oracle."""
    import ctypes as C
    import oraclelib as O
    full = (0, 0, 63, 23)
    gradient = foreign.ForeignGradient(4, 2, 50, 20)
    assert not isinstance(gradient, process.VideoSource)

    def host_frame(frame_index):
        f = O.HostFrame16(full, (4, 2, 50, 20))
        for y in range(2, 21):
            for x in range(4, 51):
                f.array[y, x] = [(x * 131 + y * 71 + frame_index * 17 + c * 977) % 0x3C00 for c in range(3)] + [0x3800 + ((x + y) & 1) * 0x0400]
        return f

    ours = O.synth_frame(full, 77, (0, 0, 40, 23))
    ours_source = process.RgbaFrameF16(full, ours.current.tuple(), ours.array)

    # a crossfade of the foreign source and one of ours
    got = process.VideoMixFilter(gradient, ours_source, 0.25).get_frame_f16(3, full)
    want = O.HostFrame16(full)
    theirs = host_frame(3)
    O.oracle().orc_cross_f16(C.byref(want.c), C.byref(theirs.c), C.byref(ours.c), 0.25)
    assert got.current_window.flat() == want.current.tuple()
    x0, y0, x1, y1 = want.current.tuple()            # pixels outside current_window are undefined (framework.rst:27-30)
    assert np.array_equal(got.data[y0:y1 + 1, x0:x1 + 1], want.array[y0:y1 + 1, x0:x1 + 1])

    # ... and a workspace with the foreign source on top, read at its own frame index (offset)
    workspace = process.VideoWorkspace()
    workspace.add(source=ours_source, x=0, length=100, z=0, offset=0)
    workspace.add(source=gradient, x=0, length=100, z=5, offset=7)
    got = workspace.get_frame_f16(2, full)
    top = host_frame(9)                             # (kept alive: the ctypes struct only borrows its pixels)
    layers = (O.Frame16 * 2)(ours.c, top.c)
    O.oracle().orc_composite_f16(C.byref(want.c), layers, 2)
    assert got.current_window.flat() == want.current.tuple()
    x0, y0, x1, y1 = want.current.tuple()
    bad = np.argwhere(np.any(got.data[y0:y1 + 1, x0:x1 + 1] != want.array[y0:y1 + 1, x0:x1 + 1], axis=2))
    assert len(bad) == 0, "first differing pixels (row, column): %s" % bad[:8].tolist()
    assert gradient.calls >= 2
