"""The framework.h drop-in path with HOST frames, the way a C client of the reference uses it
(canvas_b200/csrc/client/host_client.c -> libcanvas_client.so): host f16 frames as video sources, a workspace built with
workspace_add_item, frames pulled with video_get_frame_f16 -- from one thread and from several at once (the pull queue's
contract, /root/reference/src/process/VideoPullQueue.c:99-110: get_frame is called concurrently from worker threads).
Results are compared with the oracle bit for bit; the item query is compared with a brute-force walk over 3000 items
(/root/reference/src/cprocess/workspace.c:243-307, 494-550)."""
import ctypes as C
import os
import random

import numpy as np
import pytest

import oraclelib as O

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def libs():
    from canvas_b200 import _abi as A
    from canvas_b200 import device
    device.init(0)
    client = C.CDLL(os.path.join(ROOT, "canvas_b200", "libcanvas_client.so"))
    client.client_timeline_create.restype = C.c_void_p
    client.client_timeline_create.argtypes = [C.POINTER(A.FrameF16), C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int64),
                                              C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.c_int]
    client.client_time_timeline.restype = C.c_int64
    client.client_time_timeline.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.POINTER(A.FrameF16)]
    client.client_timeline_get_frame_f16.argtypes = [C.c_void_p, C.c_int, C.POINTER(A.FrameF16)]
    client.client_timeline_free.argtypes = [C.c_void_p]
    return A, client


def build_timeline(A, client, full, clips, specs):
    """clips: list of lists of HostFrame16; specs: (x, length, offset, z) per clip."""
    flat = [f for clip in clips for f in clip]
    frames = (A.FrameF16 * len(flat))()
    for i, f in enumerate(flat):
        frames[i].data, frames[i].full_window, frames[i].current_window = f.array.ctypes.data, A.Box2i(*f.full.tuple()), A.Box2i(*f.current.tuple())
    n = len(clips)
    firsts, at = [], 0
    for clip in clips:
        firsts.append(at)
        at += len(clip)
    ints, longs = C.c_int * n, C.c_int64 * n
    tl = client.client_timeline_create(frames, ints(*firsts), ints(*[len(c) for c in clips]), longs(*[s[0] for s in specs]),
                                       longs(*[s[1] for s in specs]), longs(*[s[2] for s in specs]), longs(*[s[3] for s in specs]), n)
    return tl, (frames, flat)


def oracle_frame(full, clips, specs, frame):
    showing = sorted([(s[3], i) for i, s in enumerate(specs) if s[0] <= frame < s[0] + s[1]])
    layers = [clips[i][(frame - specs[i][0] + specs[i][2]) % len(clips[i])] for _, i in showing]
    out = O.HostFrame16(full)
    arr = (O.Frame16 * max(1, len(layers)))(*[l.c for l in layers])
    O.oracle().orc_composite_f16(C.byref(out.c), arr, len(layers))
    return out


def test_pull_through_workspace_matches_oracle(libs):
    A, client = libs
    full = (0, 0, 319, 95)
    windows = [full, (20, 10, 300, 80), (101, 0, 319, 60), (0, 40, 150, 95)]
    clips = [[O.synth_frame(full, 100 * k + i, windows[k]) for i in range(3)] for k in range(4)]
    specs = [(0, 50, 0, 0), (2, 10, 1, 1), (0, 50, 5, 2), (4, 3, 0, 3)]
    tl, keep = build_timeline(A, client, full, clips, specs)
    for frame in (0, 1, 2, 4, 6, 7, 13, 49, 50):
        got = np.zeros((96, 320, 4), np.uint16)
        out = A.FrameF16()
        out.data, out.full_window = got.ctypes.data, A.Box2i(*full)
        client.client_timeline_get_frame_f16(tl, frame, C.byref(out))
        want = oracle_frame(full, clips, specs, frame)
        assert out.current_window.tuple() == want.current.tuple(), frame
        if not want.current.empty:
            x0, y0, x1, y1 = want.current.tuple()
            assert np.array_equal(got[y0:y1 + 1, x0:x1 + 1], want.array[y0:y1 + 1, x0:x1 + 1]), frame
    client.client_timeline_free(tl)


def test_concurrent_pulls_from_four_threads(libs):
    """Four threads pull 64 frames of one workspace concurrently, each on its own stream and staging slabs; every
    timeline frame shows the same pictures here (one-frame clips), so every thread's last result must be the oracle's."""
    A, client = libs
    full = (0, 0, 639, 179)
    windows = [full, (64, 16, 600, 170), (0, 0, 400, 100)]
    clips = [[O.synth_frame(full, 7 + k, windows[k])] for k in range(3)]
    specs = [(0, 1 << 30, 0, k) for k in range(3)]
    tl, keep = build_timeline(A, client, full, clips, specs)
    threads = 4
    outs_np = [np.zeros((180, 640, 4), np.uint16) for _ in range(threads)]
    outs = (A.FrameF16 * threads)()
    for t in range(threads):
        outs[t].data, outs[t].full_window = outs_np[t].ctypes.data, A.Box2i(*full)
    assert client.client_time_timeline(tl, 0, 64, threads, outs) > 0
    want = oracle_frame(full, clips, specs, 0)
    for t in range(threads):
        assert outs[t].current_window.tuple() == want.current.tuple()
        assert np.array_equal(outs_np[t], want.array), "thread %d" % t
    client.client_timeline_free(tl)


def test_item_query_over_three_thousand_items(libs):
    """workspace_add_item / update / remove over 3000 items, then the items the compositor pulls at a frame (order and
    source frame index) against a brute-force walk: x <= f < x + length, ascending z, index f - x + offset."""
    A, _ = libs
    lib = A.load()
    lib.workspace_create.restype = C.c_void_p
    lib.workspace_add_item.restype = C.c_void_p
    lib.workspace_add_item.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_int64, C.c_void_p]
    lib.workspace_remove_item.argtypes = [C.c_void_p]
    lib.workspace_update_item.argtypes = [C.c_void_p] + [C.c_void_p] * 6
    lib.workspace_as_video_source.argtypes = [C.c_void_p, C.c_void_p]
    lib.workspace_free.argtypes = [C.c_void_p]
    lib.video_get_frame_f16.argtypes = [C.c_void_p, C.c_int, C.POINTER(A.FrameF16)]

    class Funcs(C.Structure):
        _fields_ = [("flags", C.c_int), ("get_frame", C.c_void_p), ("get_frame_32", C.c_void_p), ("get_frame_gl", C.c_void_p)]

    class Source(C.Structure):
        _fields_ = [("obj", C.c_void_p), ("funcs", C.POINTER(Funcs))]

    pulled = []
    GET = C.CFUNCTYPE(None, C.c_void_p, C.c_int, C.POINTER(A.FrameF16))

    @GET
    def get_frame(obj, frame_index, frame):
        pulled.append((obj, frame_index))
        frame.contents.current_window = A.Box2i()           # nothing to show: the walk is what is being tested

    funcs = Funcs(0, C.cast(get_frame, C.c_void_p), None, None)
    rng = random.Random(5)
    ws = lib.workspace_create()
    items, sources = {}, []
    for i in range(3000):
        src = Source(i + 1, C.pointer(funcs))
        sources.append(src)
        x, length, offset, z = rng.randrange(0, 5000), rng.randrange(1, 400), rng.randrange(0, 50), rng.randrange(-5, 2000000)
        items[i] = [lib.workspace_add_item(ws, C.addressof(src), x, length, offset, z, None), x, length, offset, z]
    for i in rng.sample(sorted(items), 300):                # move some, stretch some, drop some
        handle, x, length, offset, z = items[i]
        nx, nl = C.c_int64(rng.randrange(0, 5000)), C.c_int64(rng.randrange(1, 900))
        lib.workspace_update_item(handle, C.addressof(nx), C.addressof(nl), None, None, None, None)
        items[i][1], items[i][2] = nx.value, nl.value
    for i in rng.sample(sorted(items), 200):
        lib.workspace_remove_item(items.pop(i)[0])

    wsrc = Source()
    lib.workspace_as_video_source(ws, C.addressof(wsrc))
    out = A.FrameF16()
    data = np.zeros((4, 4, 4), np.uint16)
    out.data, out.full_window = data.ctypes.data, A.Box2i(0, 0, 3, 3)
    for frame in (0, 17, 1234, 2500, 4999, 5200, 6000):
        del pulled[:]
        lib.video_get_frame_f16(C.addressof(wsrc), frame, C.byref(out))
        want = sorted((z, i) for i, (_, x, length, offset, z) in items.items() if x <= frame < x + length)
        distinct = len({z for z, _ in want}) == len(want)
        if distinct:
            assert pulled == [(i + 1, frame - items[i][1] + items[i][3]) for _, i in want], frame
        else:
            assert sorted(pulled) == sorted((i + 1, frame - items[i][1] + items[i][3]) for _, i in want), frame
    lib.workspace_free(ws)
