"""The fluggo.media.process type surface (canvas_b200.process).

Host logic (frame functions, workspace bookkeeping, sequences) runs on the CPU; everything that
produces pixels is marked gpu.  The first GPU tests restate the reference's own tests
(tests/process/video/RgbaFrameF16.py, SolidColorVideoSource.py with force_gl=False,
VideoWorkspace.py, tests/process/frame_func.py) against this package; the later ones compare whole
graphs with the oracle bit for bit."""
import ctypes as C
import random
from fractions import Fraction

import numpy as np
import pytest

import oraclelib as O
from canvas_b200 import process
from canvas_b200.basetypes import box2i, rgba, v2i


def f2h(values):
    a = np.asarray(values, np.float32)
    out = np.zeros(a.shape, np.uint16)
    O.oracle().orc_f2h_n(O.ptr(out), O.ptr(a), a.size)
    return out


def valid_mask(frame):
    """Boolean (h, w) mask of the pixels inside the frame's current window."""
    m = np.zeros(frame.array.shape[:2], bool)
    c, f = frame.current, frame.full
    if not c.empty:
        m[c.min_y - f.min_y: c.max_y - f.min_y + 1, c.min_x - f.min_x: c.max_x - f.min_x + 1] = True
    return m


def almost(a, b, places=6):
    assert len(a) == len(b)
    for x, y in zip(a, b):
        assert round(abs(x - y), places) == 0, (a, b)


# ---- tests/process/frame_func.py --------------------------------------------------------------

def test_lerp_func():
    func = process.LerpFunc((1.0, 2.0, 3.0, 4.0), (-1.0, -2.0, -3.0, -4.0), 4)
    almost((1.0, 2.0, 3.0, 4.0), func.get_values(0)[0])
    almost((0.5, 1.0, 1.5, 2.0), func.get_values(1)[0])
    almost((0.0, 0.0, 0.0, 0.0), func.get_values(2)[0])
    almost((-0.5, -1.0, -1.5, -2.0), func.get_values(3)[0])
    almost((-1.0, -2.0, -3.0, -4.0), func.get_values(4)[0])
    want = [(-1.0, -2.0, -3.0, -4.0), (0.5, 1.0, 1.5, 2.0), (0.0, 0.0, 0.0, 0.0), (1.0, 2.0, 3.0, 4.0), (-0.5, -1.0, -1.5, -2.0)]
    for a, b in zip(want, func.get_values([4, 1, 2, 0, 3])):
        almost(a, b)
    with pytest.raises(Exception):
        process.LerpFunc((0,), (1,), 0)


def test_frame_func_pass_through_filter():
    """FrameFuncPassThroughFilter.c:33-75: values of the source at frame + offset; constants pass as they are"""
    lerp = process.LerpFunc((0.0, 10.0, 0.0, 1.0), (4.0, 14.0, -4.0, 1.0), 4)
    shifted = process.FrameFuncPassThroughFilter(lerp, 1.5)
    assert shifted.source() is lerp
    for frame in (0, 1, 2.25):
        almost(shifted.get_values(frame)[0], lerp.get_values(frame + 1.5)[0])
    almost(shifted.get_values([0, 2])[1], lerp.get_values(3.5)[0])
    constant = process.FrameFuncPassThroughFilter((1.0, 2.0, 3.0, 4.0), 7.0)
    assert constant.source() is None
    almost(constant.get_values(5)[0], (1.0, 2.0, 3.0, 4.0))
    constant.set_source(lerp)
    almost(constant.get_values(0)[0], lerp.get_values(7.0)[0])
    # a pass-through is itself a frame function: it can drive a source's colour
    assert isinstance(shifted, process.FrameFunction)


def test_animation_func_basic_and_move():
    func = process.AnimationFunc()
    func.add(process.AnimationPoint(process.POINT_HOLD, 0.0, 4.0))
    func.add(process.AnimationPoint(process.POINT_LINEAR, 1.0, 2.0))
    func.add(process.AnimationPoint(process.POINT_LINEAR, 2.0, 6.0))
    table = {-0.5: 4.0, 0.0: 4.0, 0.25: 4.0, 0.5: 4.0, 0.75: 4.0, 1.0: 2.0, 1.25: 3.0, 1.5: 4.0, 1.75: 5.0, 2.0: 6.0, 2.5: 6.0}
    for t, v in table.items():
        assert abs(func.get_values(t)[0][0] - v) < 1e-7
    keys = list(table)
    random.Random(3).shuffle(keys)
    for t in keys:
        assert abs(func.get_values(t)[0][0] - table[t]) < 1e-7

    func = process.AnimationFunc()
    pt1 = func.add(process.POINT_HOLD, 0.0, 4.0)
    pt2 = func.add(process.POINT_LINEAR, 2.0, 6.0)
    pt3 = func.add(process.POINT_LINEAR, 1.0, 2.0)
    assert (func[0], func[1], func[2]) == (pt1, pt3, pt2)
    pt3.frame = 3.0
    assert (func[0], func[1], func[2]) == (pt1, pt2, pt3)


def test_basetypes():
    assert box2i((0, 0), (2, 2)) == box2i(0, 0, 2, 2)
    assert box2i(0, 0, 2, 2).size == v2i(3, 3) and not box2i(0, 0, 2, 2).empty
    assert box2i().empty and not box2i()
    assert rgba(1, 2, 3, 4).a == 4


def test_workspace_bookkeeping_without_gpu():
    ws = process.VideoWorkspace()
    a = ws.add(source=None, x=10, length=5, z=2, offset=1)
    b = ws.add(source=None, x=3, length=5, z=7)
    c = ws.add(source=None, x=10, length=1, z=-1, tag="c")
    assert len(ws) == 3 and [ws[i] for i in range(3)] == [b, c, a]
    a.update(x=0, z=3)
    assert ws[0] is a and (a.x, a.z, a.length, a.offset) == (0, 3, 5, 1)
    with pytest.raises(TypeError):
        a.update(bogus=1)
    ws.remove(b)
    assert len(ws) == 2
    with pytest.raises(ValueError):
        ws.remove(b)
    assert ws._describe(500, (0, 0, 1, 1)) is None          # nothing shows: no GPU needed


def test_sequence_bookkeeping_without_gpu():
    seq = process.VideoSequence()
    seq.append((None, 0, 10))
    seq.append((None, 5, 3))
    seq.insert(1, (None, 1, 7))
    assert len(seq) == 3 and [seq.get_start_frame(i) for i in range(3)] == [0, 10, 17]
    with pytest.raises(ValueError):
        seq.append((None, 0, 0))
    assert seq._describe(-1, (0, 0, 1, 1)) is None and seq._describe(20, (0, 0, 1, 1)) is None


def test_time_helpers():
    rate = Fraction(30000, 1001)
    for frame in (0, 1, 29, 30, 1000):
        t = process.get_frame_time(rate, frame)
        assert process.get_time_frame(rate, t) == frame
    assert process.get_frame_time(Fraction(24, 1), 24) == 1000000001      # main.c:25 adds one nanosecond


# ---- the reference's own video tests, on the GPU ----------------------------------------------

@pytest.mark.gpu
def test_reference_rgba_frame_f16_solid():
    color = (1.0, 0.5, 0.333333, 0.2)
    solid = process.SolidColorVideoSource(color, box2i((0, 0), (2, 2)))
    frame = solid.get_frame_f16(0, box2i((0, 0), (3, 3)))
    assert frame.current_window == box2i(0, 0, 2, 2)
    assert frame.full_window == box2i(0, 0, 3, 3)
    almost(frame.pixel(0, 0), color, 3)
    frame2 = frame.get_frame_f16(0, box2i(-1, -1, 1, 1))
    assert frame2.current_window == box2i(0, 0, 1, 1)
    assert frame2.full_window == box2i(-1, -1, 1, 1)
    almost(frame2.pixel(0, 0), color, 3)
    assert frame2.pixel(-1, -1) is None
    # exactly the truncated halves, everywhere in the window
    want = f2h(color)
    assert np.array_equal(frame.data[0:3, 0:3], np.broadcast_to(want, (3, 3, 4)))


@pytest.mark.gpu
def test_reference_solid_color_video_source():
    color = (1.0, 0.5, 0.333333, 0.2)
    solid = process.SolidColorVideoSource(color)
    frame = solid.get_frame_f32(0, box2i(0, 0, 3, 3), force_gl=False)
    assert frame.current_window == box2i(0, 0, 3, 3)
    almost(frame.pixel(0, 0), color)

    solid = process.SolidColorVideoSource(color, box2i(0, 0, 2, 2))
    frame = solid.get_frame_f32(0, box2i(0, 0, 3, 3))
    assert frame.current_window == box2i(0, 0, 2, 2)
    frame2 = frame.get_frame_f32(0, box2i(-1, -1, 1, 1))
    assert frame2.current_window == box2i(0, 0, 1, 1)
    almost(frame2.pixel(0, 0), color)

    solid = process.SolidColorVideoSource(process.LerpFunc((0.5, 0.25, 2.0, 1.0), (-0.5, -0.25, -2.0, 0.0), 2))
    almost(solid.get_frame_f32(0, box2i(0, 0, 0, 0)).pixel(0, 0), rgba(0.5, 0.25, 2.0, 1.0))
    almost(solid.get_frame_f32(1, box2i(0, 0, 0, 0)).pixel(0, 0), rgba(0.0, 0.0, 0.0, 0.5))
    almost(solid.get_frame_f32(2, box2i(0, 0, 0, 0)).pixel(0, 0), rgba(-0.5, -0.25, -2.0, 0.0))

    solid = process.SolidColorVideoSource(rgba(0.0, 0.0, 1.0, 1.0), process.LerpFunc((-2, -2, 2, 2), (-4, -4, 0, 6), 2))
    assert solid.get_frame_f32(0, box2i(-5, -5, 5, 6)).current_window == box2i(-2, -2, 2, 2)
    assert solid.get_frame_f32(1, box2i(-5, -5, 5, 6)).current_window == box2i(-3, -3, 1, 4)
    assert solid.get_frame_f32(2, box2i(-5, -5, 5, 6)).current_window == box2i(-4, -4, 0, 6)


@pytest.mark.gpu
def test_reference_video_workspace_random():
    rnd = random.Random(11)
    red = process.SolidColorVideoSource(process.LerpFunc((0, 0, 0, 1), (100, 0, 0, 1), 100))
    green = process.SolidColorVideoSource(process.LerpFunc((0, 0, 0, 1), (0, 100, 0, 1), 100))
    blue = process.SolidColorVideoSource(process.LerpFunc((0, 0, 0, 1), (0, 0, 100, 1), 100))
    workspace = process.VideoWorkspace()

    def getcolor(source, frame):
        return source.get_frame_f32(frame, box2i(0, 0, 0, 0)).pixel(0, 0)

    def top_colour(frame):
        showing = [it for it in workspace._items if it.x <= frame < it.x + it.length]
        if not showing:
            return None
        top = max(showing, key=lambda it: (it.z, -it._serial))
        return getcolor(top.source, frame - top.x + top.offset)

    for i in range(600):
        action = rnd.randint(1, 7)
        items = [workspace[k] for k in range(len(workspace))]
        if action == 1 and items:
            rnd.choice(items).update(x=rnd.randint(0, 1000))
        elif action == 2 and items:
            rnd.choice(items).update(z=rnd.randint(-10, 10))
        elif action == 3 and items:
            rnd.choice(items).update(length=rnd.randint(1, 100))
        elif action == 4 and items:
            rnd.choice(items).update(offset=rnd.randint(-20, 20))
        elif action == 5 and items:
            workspace.remove(rnd.choice(items))
        elif action == 6:
            for _ in range(3):
                frame = rnd.randint(-100, 1100)
                got = getcolor(workspace, frame)
                want = top_colour(frame)
                # every source is opaque (alpha 1): the top item's colour shows through unchanged
                if want is None:
                    assert got is None
                else:
                    almost(got, want, 4)
        else:
            workspace.add(source=rnd.choice((red, green, blue)), x=rnd.randint(0, 1000), z=rnd.randint(-10, 10),
                          length=rnd.randint(1, 100), offset=rnd.randint(-20, 20))


# ---- whole graphs against the oracle ----------------------------------------------------------

class _Pictures(process.CodedImageSource):
    """A CodedImageSource over synthetic 720x480 4:1:1 planes (what a DV decoder would hand over)."""

    def __init__(self, seeds):
        self.pics = [O.synth_coded(720, 480, s, 4, O.LAYOUT_PLANAR, O.MATRIX_709, pic=(0, -1)) for s in seeds]

    def get_coded_image(self, frame):
        p = self.pics[frame % len(self.pics)]
        return process.CodedImage([pl.tobytes() for pl in p.planes], [pl.shape[1] for pl in p.planes], [pl.shape[0] for pl in p.planes])


def _host(frame16):
    return O.HostFrame16(frame16.full_window.flat(), frame16.current_window.flat(), frame16.data)


@pytest.mark.gpu
def test_dv_graph_matches_oracle():
    """DVReconstructionFilter x2 -> VideoMixFilter(AnimationFunc) -> VideoWorkspace with two f16 overlays, frame by frame."""
    full = (0, -1, 719, 478)
    pa, pb = _Pictures([1, 2, 3]), _Pictures([7, 8])
    mix = process.AnimationFunc()
    mix.add(process.POINT_LINEAR, 0.0, 0.0)
    mix.add(process.POINT_LINEAR, 8.0, 1.0)
    fade = process.VideoMixFilter(process.DVReconstructionFilter(pa), process.DVReconstructionFilter(pb), mix)
    over = [O.synth_frame(full, 50 + i, cw) for i, cw in enumerate([(100, 50, 500, 300), (301, 100, 719, 478)])]
    overlays = [process.RgbaFrameF16(o.full.tuple(), o.current.tuple(), o.array) for o in over]

    ws = process.VideoWorkspace()
    ws.add(source=fade, x=0, length=9, z=0, offset=0)
    ws.add(source=overlays[0], x=2, length=5, z=1)
    ws.add(source=overlays[1], x=0, length=9, z=2)

    orc = O.oracle()
    for f in range(9):
        got = ws.get_frame_f16(f, box2i(*full))
        layers = ([over[0]] if 2 <= f < 7 else []) + [over[1]]
        arr = (O.Frame16 * len(layers))(*[l.c for l in layers])
        want = O.HostFrame16(full)
        m = float(np.float32(f / 8.0))
        orc.orc_timeline_frame_f16(C.byref(want.c), C.byref(pa.pics[f % 3].c), C.byref(pb.pics[f % 2].c), m, arr, len(layers))
        assert got.current_window.flat() == want.current.tuple()
        assert np.array_equal(got.data, want.array), "frame %d" % f


@pytest.mark.gpu
def test_subsample_filter_round_trip_matches_oracle():
    full = (0, -1, 719, 478)
    src = O.synth_frame(full, 77, full)
    frame = process.RgbaFrameF16(src.full.tuple(), src.current.tuple(), src.array)
    image = process.DVSubsampleFilter(frame).get_coded_image(0)

    want = O.synth_coded(720, 480, 1, 4, O.LAYOUT_PLANAR, O.MATRIX_709, pic=(0, -1))
    for p in want.planes:
        p[...] = 0
    copy = O.HostFrame16(full, full, src.array.copy())
    O.oracle().orc_subsample(C.byref(copy.c), C.byref(want.c))
    for got, plane in zip(image.data, want.planes):
        assert np.array_equal(np.frombuffer(got, np.uint8).reshape(plane.shape), plane)


@pytest.mark.gpu
def test_sequence_passthrough_and_nested_mix():
    full = (0, 0, 63, 31)
    frames = [O.synth_frame(full, 90 + i, (2 * i, i, 63 - i, 31)) for i in range(3)]
    srcs = [process.RgbaFrameF16(f.full.tuple(), f.current.tuple(), f.array) for f in frames]

    seq = process.VideoSequence()
    seq.append((srcs[0], 0, 2))
    seq.append((srcs[1], 0, 3))
    assert np.array_equal(seq.get_frame_f16(1, box2i(*full)).data[valid_mask(frames[0])], frames[0].array[valid_mask(frames[0])])
    assert seq.get_frame_f16(4, box2i(*full)).current_window.flat() == frames[1].current.tuple()
    assert seq.get_frame_f16(5, box2i(*full)).current_window.empty

    pt = process.VideoPassThroughFilter(srcs[2], offset=3, start_frame=1, end_frame=4)
    assert pt.get_frame_f16(0, box2i(*full)).current_window.empty
    assert pt.get_frame_f16(4, box2i(*full)).current_window.empty
    assert pt.get_frame_f16(2, box2i(*full)).current_window.flat() == frames[2].current.tuple()

    # a crossfade whose side is itself a crossfade: the inner result stays f32 (no rounding in between)
    inner = process.VideoMixFilter(srcs[0], srcs[1], 0.25)
    outer = process.VideoMixFilter(inner, srcs[2], 0.5)
    got = outer.get_frame_f16(0, box2i(*full))
    orc = O.oracle()
    a32, b32, c32, t32 = (O.HostFrame32(full) for _ in range(4))
    orc.orc_frame_f16_to_f32(C.byref(a32.c), C.byref(frames[0].c))
    orc.orc_frame_f16_to_f32(C.byref(b32.c), C.byref(frames[1].c))
    orc.orc_frame_f16_to_f32(C.byref(c32.c), C.byref(frames[2].c))
    orc.orc_mix_cross_f32(C.byref(t32.c), C.byref(a32.c), C.byref(b32.c), 0.25)
    out32 = O.HostFrame32(full)
    orc.orc_mix_cross_f32(C.byref(out32.c), C.byref(t32.c), C.byref(c32.c), 0.5)
    want = O.HostFrame16(full)
    orc.orc_frame_f32_to_f16(C.byref(want.c), C.byref(out32.c))
    assert got.current_window.flat() == want.current.tuple()
    mask = valid_mask(want)
    assert np.array_equal(got.data[mask], want.array[mask])


@pytest.mark.gpu
def test_to_argb32_bytes_and_time_get_frame():
    full = (0, 0, 31, 15)
    f = O.synth_frame(full, 5, (3, 2, 30, 15))
    frame = process.RgbaFrameF16(f.full.tuple(), f.current.tuple(), f.array)
    want = np.zeros((14, 28), np.uint32)
    ramp45 = np.zeros(65536, np.uint8)
    O.oracle().orc_build_gamma45(O.ptr(ramp45))
    O.oracle().orc_to_argb32(C.byref(f.c), O.ptr(ramp45), O.ptr(want))
    assert frame.to_argb32_bytes() == want.tobytes()
    assert process.time_get_frame(frame, 0, 5, box2i(*full)) > 0


@pytest.mark.gpu
def test_video_pull_queue():
    """VideoPullQueue.c:100-183: frames rendered on worker threads, callback(frame_index, frame, user_data); cancel()"""
    import threading
    solid = process.SolidColorVideoSource(process.LerpFunc((0.0, 0.0, 0.0, 1.0), (8.0, 4.0, 2.0, 1.0), 8), box2i(0, 0, 15, 7))
    queue = process.VideoPullQueue()
    got, done = {}, threading.Event()

    def callback(frame_index, frame, user_data):
        got[frame_index] = (frame, user_data)
        if len(got) == 8:
            done.set()

    items = [queue.enqueue(solid, i, box2i(-2, -2, 20, 10), callback, "tag%d" % i) for i in range(8)]
    assert done.wait(60), sorted(got)
    queue.close()
    for i in range(8):
        frame, tag = got[i]
        assert tag == "tag%d" % i and frame.current_window == box2i(0, 0, 15, 7)
        almost(frame.pixel(3, 3), rgba(float(i), i / 2.0, i / 4.0, 1.0))
        assert frame.pixel(-1, -1) is None

    # a cancelled item never calls back
    queue = process.VideoPullQueue()
    gate = threading.Event()

    class Slow(process.VideoSource):
        def _describe(self, frame_index, full):
            gate.wait(30)
            return solid._describe(frame_index, full)

    calls = []
    slow = Slow()
    first = [queue.enqueue(slow, i, box2i(0, 0, 3, 3), lambda *a: calls.append(a[0]), None) for i in range(2)]    # both workers busy
    late = queue.enqueue(slow, 9, box2i(0, 0, 3, 3), lambda *a: calls.append(a[0]), None)
    late.cancel()
    first[1].cancel()
    gate.set()
    queue.close()
    assert calls == [0]
