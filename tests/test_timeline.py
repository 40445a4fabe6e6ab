"""The frame-parallel timeline scheduler (cv_timeline_*, include/canvas_b200.h).

CPU part: the stateless item query and the per-frame job follow workspace.c:494-550
(x <= f < x + length, ascending z, source frame f - x + offset), the block-cyclic partition covers
every frame exactly once, and the N>1 plumbing of bench.py (barrier, max/sum over ranks) works over
gloo with two processes.  GPU part: cv_timeline_render against the oracle, frame by frame."""
import ctypes as C
import os
import subprocess
import sys

import numpy as np
import pytest

import oraclelib as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    import __graft_entry__ as g
    if not os.path.exists(os.path.join(ROOT, "canvas_b200", "libcanvas_b200.so")):
        g.build()
    from canvas_b200 import _abi
    return _abi.load()


def test_struct_layouts_match_c():
    src = r'''
    #include "canvas_b200.h"
    #include <stdio.h>
    #include <stddef.h>
    int main(void) {
        printf("%zu %zu %zu %zu %zu %zu %zu\n", sizeof(cv_clip), sizeof(cv_timeline_item), sizeof(cv_timeline_target),
               sizeof(cv_timeline_run), sizeof(cv_timeline_stats), offsetof(cv_timeline_run, targets), offsetof(cv_timeline_item, ramp_start));
        return 0;
    }'''
    exe = "/tmp/cv_timeline_layout_probe"
    subprocess.run(["gcc", "-x", "c", "-", "-I", os.path.join(ROOT, "include"), "-o", exe], input=src.encode(), check=True)
    sizes = [int(v) for v in subprocess.run([exe], capture_output=True, check=True).stdout.split()]
    from canvas_b200 import _abi as A
    assert sizes == [C.sizeof(A.Clip), C.sizeof(A.TimelineItem), C.sizeof(A.TimelineTarget), C.sizeof(A.TimelineRun),
                     C.sizeof(A.TimelineStats), A.TimelineRun.targets.offset, A.TimelineItem.ramp_start.offset]


@pytest.mark.parametrize("first,last,chunk,n", [(0, 999, 16, 8), (5, 5, 8, 4), (-7, 120, 3, 2), (0, 30, 32, 4), (10, 9, 8, 2), (0, 1000, 1, 1)])
def test_partition_covers_every_frame_once(lib, first, last, chunk, n):
    owners = [lib.cv_timeline_device_of(f, first, chunk, n) for f in range(first, last + 1)]
    assert all(0 <= o < n for o in owners)
    counts = [owners.count(g) for g in range(n)]
    assert counts == [lib.cv_timeline_frames_of(g, first, last, chunk, n) for g in range(n)]
    assert sum(counts) == max(0, last - first + 1)
    # block-cyclic: consecutive frames of one chunk share a GPU, neighbouring chunks do not (n > 1)
    for i, f in enumerate(range(first, last + 1)):
        assert owners[i] == ((f - first) // chunk) % n
    assert max(counts) - min(counts) <= chunk


def make_items(A, specs, n_slots=1):
    """specs: (x, length, z, offset, clip_len, kind, has_b).  Sources are fakes: only kinds and markers matter here."""
    keep = []
    items = (A.TimelineItem * len(specs))()
    for i, (x, length, z, offset, clip_len, kind, has_b) in enumerate(specs):
        def clips(tag):
            arr = (A.Clip * n_slots)()
            for s in range(n_slots):
                frames = (A.Source * clip_len)()
                for k in range(clip_len):
                    frames[k].kind = kind
                    frames[k].reserved = tag * 1000000 + s * 10000 + k           # marker: which clip / slot / ring entry
                    frames[k].f16.full_window = A.Box2i(0, 0, 3, 3)
                    frames[k].f16.current_window = A.Box2i(i, s, 3, k % 4)
                arr[s].frames, arr[s].count = frames, clip_len
                keep.append(frames)
            keep.append(arr)
            return arr
        items[i].x, items[i].length, items[i].z, items[i].offset = x, length, z, offset
        items[i].clips = clips(2 * i)
        if has_b:
            items[i].clips_b = clips(2 * i + 1)
            items[i].mix_first, items[i].mix_last, items[i].ramp_start, items[i].ramp_length = 0.0, 1.0, offset, length - 1
    return items, keep


def showing_reference(specs, frame):
    """workspace.c:504-547 restated: active items, bottom first (ascending z; later-added lower among equals)."""
    live = [i for i, s in enumerate(specs) if s[0] <= frame < s[0] + s[1]]
    return sorted(live, key=lambda i: (specs[i][2], -i))


def test_showing_matches_workspace_semantics(lib):
    from canvas_b200 import _abi as A
    rng = np.random.default_rng(5)
    specs = [(int(rng.integers(-20, 60)), int(rng.integers(1, 50)), int(rng.integers(-3, 4)), int(rng.integers(-5, 30)), 4, A.CV_SRC_F16, False)
             for _ in range(40)]
    items, keep = make_items(A, specs)
    idx = (C.c_int * 64)()
    for frame in range(-30, 120):
        n = lib.cv_timeline_showing(items, len(specs), frame, idx, 64)
        assert list(idx[:n]) == showing_reference(specs, frame)


def test_job_follows_item_walk(lib):
    from canvas_b200 import _abi as A
    # bottom: a crossfade of two coded clips; above: three f16 overlays, one of them only part of the time
    specs = [(0, 100, 0, 7, 5, A.CV_SRC_CODED, True),
             (0, 100, 1, 0, 3, A.CV_SRC_F16, False),
             (10, 20, 3, 2, 4, A.CV_SRC_F16, False),
             (0, 100, 2, -1, 6, A.CV_SRC_F16, False)]
    items, keep = make_items(A, specs, n_slots=2)
    for slot in (0, 1):
        for frame in (0, 9, 10, 17, 29, 30, 99):
            job = A.FrameJob()
            n = lib.cv_timeline_job(items, len(specs), slot, frame, C.byref(job))
            order = showing_reference(specs, frame)
            assert n == len(order)
            src_frame = frame - 0 + 7
            assert job.a.kind == A.CV_SRC_CODED and job.a.reserved == 0 * 1000000 + slot * 10000 + src_frame % 5
            assert job.b.kind == A.CV_SRC_CODED and job.b.reserved == 1 * 1000000 + slot * 10000 + src_frame % 5
            t = (src_frame - 7) / 99.0
            want_mix = np.float32(0.0 * (1.0 - t) + 1.0 * t) if 0.0 < t < 1.0 else np.float32(min(max(t, 0.0), 1.0))
            assert job.mix_b == want_mix
            assert job.layer_count == len(order) - 1
            for k, i in enumerate(order[1:]):
                x, length, z, offset, clip_len, _, _ = specs[i]
                ring = (frame - x + offset) % clip_len
                got = job.layers[k].frame.current_window.tuple()
                assert got == (i, slot, 3, ring % 4), (frame, k, i, got)
                assert job.layers[k].mix == 1.0
    # nothing showing -> empty job
    job = A.FrameJob()
    assert lib.cv_timeline_job(items, len(specs), 0, 500, C.byref(job)) == 0
    assert job.a.kind == A.CV_SRC_NONE and job.layer_count == 0


def test_job_rejects_unsupported_stacks(lib):
    from canvas_b200 import _abi as A
    specs = [(0, 10, 0, 0, 2, A.CV_SRC_F16, False), (0, 10, 1, 0, 2, A.CV_SRC_CODED, False)]
    items, keep = make_items(A, specs)
    job = A.FrameJob()
    assert lib.cv_timeline_job(items, 2, 0, 3, C.byref(job)) == -5          # CV_ERR_UNSUPPORTED: coded above the bottom
    assert b"bottom" in lib.cv_last_error()


GLOO_WORKER = r'''
import os, sys
sys.path.insert(0, %(root)r)
import bench
from canvas_b200 import _abi
lib = _abi.load()
dist, rank, world = bench.dist_setup(world_size_hint := int(os.environ["WORLD_SIZE"]))
assert world == 2 and dist.get_backend() == "gloo"
first, last, chunk = 0, 999, 16
mine = [f for f in range(first, last + 1) if lib.cv_timeline_device_of(f, first, chunk, world) == rank]
assert len(mine) == lib.cv_timeline_frames_of(rank, first, last, chunk, world)
dist.barrier()
total = bench.dist_sum(dist, float(len(mine)))
slowest = bench.dist_max(dist, 10.0 + rank)
assert total == 1000.0 and slowest == 11.0, (total, slowest)
dist.barrier()
dist.destroy_process_group()
print("rank %%d ok %%d frames" %% (rank, len(mine)))
'''


def test_two_rank_plumbing_over_gloo(tmp_path):
    """The N>1 path of bench.py: one process per GPU, frames sharded by index, barrier + max/sum over ranks."""
    script = tmp_path / "worker.py"
    script.write_text(GLOO_WORKER % {"root": ROOT})
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT="29517", CUDA_VISIBLE_DEVICES="")
    proc = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                           "--master-addr", "127.0.0.1", "--master-port", "29517", str(script)],
                          env=env, capture_output=True, text=True, timeout=300)
    assert proc.returncode == 0, proc.stdout + proc.stderr
    assert "rank 0 ok 504 frames" in proc.stdout and "rank 1 ok 496 frames" in proc.stdout


# ----------------------------------------------------------------------------- GPU

def _client():
    from canvas_b200 import _abi as A
    client = C.CDLL(os.path.join(ROOT, "canvas_b200", "libcanvas_client.so"))
    client.client_coded_clip_create.restype = C.c_void_p
    client.client_coded_clip_create.argtypes = [C.POINTER(C.c_void_p), C.c_int, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    client.client_coded_clip_source.restype = C.POINTER(A.CodedImageSource)
    client.client_coded_clip_source.argtypes = [C.c_void_p]
    client.client_coded_clip_pulls.restype = C.c_long
    client.client_coded_clip_pulls.argtypes = [C.c_void_p]
    client.client_coded_clip_free.argtypes = [C.c_void_p]
    client.client_video_clip_create.restype = C.c_void_p
    client.client_video_clip_create.argtypes = [C.POINTER(A.FrameF16), C.c_int]
    client.client_video_clip_free.argtypes = [C.c_void_p]
    return client


def _timeline_on_gpus(lib, n_gpus, ring=2, same_windows=False, static=False, pull=False):
    """static: the fixed block-cyclic deal (frame -> GPU known in advance); otherwise chunks are stolen.
    pull: the two 4:2:2 clips come from the host through coded_image_source.getFrame (a: ordinary memory, staged through
    the worker's pinned ring; b: pinned memory, copied from directly) and the first overlay through a video_source."""
    from canvas_b200 import _abi as A
    from canvas_b200 import device as dev
    w, h = 192, 64
    full = (0, 0, w - 1, h - 1)
    frames_total, chunk = 23, 3

    # host-side clips (the oracle sees these), then one device copy per GPU
    pics_a = [O.synth_coded(w, h, 100 + i, 2, O.LAYOUT_UYVY, O.MATRIX_709) for i in range(3)]
    pics_b = [O.synth_coded(w, h, 200 + i, 2, O.LAYOUT_YUY2, O.MATRIX_601) for i in range(2)]
    over1 = [O.synth_frame(full, 300 + i, (16, 8, 150, 50)) for i in range(2)]
    over2 = [O.synth_frame(full, 400 + i, (40 + (0 if same_windows else 2 * i), 0, w - 1, h - 1)) for i in range(4)]

    keep, clips = [], {}
    pulls, client, coded_clips = {}, None, {}
    if pull:
        client = _client()
        dev.init(0)
        for name, pics in (("a", pics_a), ("b", pics_b)):
            planes = (C.c_void_p * len(pics))()
            for k, pic in enumerate(pics):
                src = pic.planes[0]
                if name == "b":                     # pinned: the scheduler copies straight out of it
                    mem = C.c_void_p()
                    dev.check(lib.cv_malloc(C.byref(mem), src.nbytes, A.CV_MEM_PINNED), "cv_malloc(pinned)")
                    C.memmove(mem, src.ctypes.data, src.nbytes)
                    planes[k] = mem.value
                else:
                    planes[k] = src.ctypes.data
            stride, lines = (C.c_int * 1)(pics[0].planes[0].shape[1]), (C.c_int * 1)(h)
            clip = client.client_coded_clip_create(planes, len(pics), 1, stride, lines)
            coded_clips[name] = clip
            p = A.Pull()
            p.coded = client.client_coded_clip_source(clip)
            p.format.layout, p.format.width, p.format.height = pics[0].layout, w, h
            p.format.sub_x, p.format.matrix = 2, pics[0].matrix
            pulls[name] = p
            keep += [planes, stride, lines]
        frames16 = (A.FrameF16 * len(over1))()
        for k, hst in enumerate(over1):
            frames16[k].data, frames16[k].full_window = hst.array.ctypes.data, A.Box2i(*hst.full.tuple())
            frames16[k].current_window = A.Box2i(*hst.current.tuple())
        p = A.Pull()
        p.video = client.client_video_clip_create(frames16, len(over1))
        p.full_window = A.Box2i(*full)
        pulls["o1"] = p
        keep.append(frames16)
    for name, hosts in (("a", pics_a), ("b", pics_b), ("o1", over1), ("o2", over2)):
        if name in pulls:
            continue
        arr = (A.Clip * n_gpus)()
        for g in range(n_gpus):
            dev.init(g)
            srcs = (A.Source * len(hosts))()
            for k, hst in enumerate(hosts):
                if name in ("a", "b"):
                    d = dev.DeviceCoded(hst.planes, hst.width, hst.height, hst.sub_x, hst.layout, hst.matrix, hst.pic)
                    srcs[k].kind, srcs[k].coded = A.CV_SRC_CODED, d.c
                else:
                    d = dev.DeviceFrame16.from_numpy(hst.array, hst.full.tuple(), hst.current.tuple())
                    srcs[k].kind, srcs[k].f16 = A.CV_SRC_F16, d.c
                keep.append(d)
            arr[g].frames, arr[g].count = srcs, len(hosts)
            keep.append(srcs)
        clips[name] = arr

    items = (A.TimelineItem * 3)()
    items[0].x, items[0].length, items[0].z, items[0].offset = 0, frames_total, 0, 1
    if pull:
        items[0].pull, items[0].pull_b = C.pointer(pulls["a"]), C.pointer(pulls["b"])
    else:
        items[0].clips, items[0].clips_b = clips["a"], clips["b"]
    items[0].mix_first, items[0].mix_last, items[0].ramp_start, items[0].ramp_length = 0.0, 1.0, 1, frames_total - 1
    items[1].x, items[1].length, items[1].z, items[1].offset = 4, 12, 5, 0
    if pull:
        items[1].pull = C.pointer(pulls["o1"])
    else:
        items[1].clips = clips["o1"]
    items[2].x, items[2].length, items[2].z, items[2].offset = 0, frames_total, 9, 3
    items[2].clips = clips["o2"]

    targets = (A.TimelineTarget * n_gpus)()
    outs = []
    for g in range(n_gpus):
        dev.init(g)
        frames = (A.FrameF16 * ring)()
        for r in range(ring):
            f = dev.DeviceFrame16(full)
            keep.append(f)
            frames[r] = f.c
        targets[g].out, targets[g].ring = frames, ring
        outs.append(frames)

    results = {}
    host_bufs = {f: np.zeros((h, w, 4), np.uint16) for f in range(frames_total)}

    @A.TIMELINE_SINK
    def sink(user, slot, frame, job, stream):
        out = job.contents.out.contents
        results[frame] = (slot, out.current_window.tuple())
        return lib.cv_download(host_bufs[frame].ctypes.data_as(C.c_void_p), out.data, host_bufs[frame].nbytes, stream)

    run = A.TimelineRun()
    run.items, run.item_count, run.n_gpus = items, 3, n_gpus
    run.first, run.last, run.chunk = 0, frames_total - 1, chunk
    run.targets, run.sink = targets, sink
    run.flags = A.CV_TIMELINE_STATIC if static else 0
    stats = (A.TimelineStats * n_gpus)()
    before = lib.cv_kernel_launches()
    dev.check(lib.cv_timeline_render(C.byref(run), stats), "cv_timeline_render")
    # one fused launch per frame, or fewer: consecutive frames of one shape on a GPU share a persistent launch
    assert 0 < lib.cv_kernel_launches() - before <= frames_total
    if same_windows and ring >= 4 and n_gpus == 1:
        # frames 0 | 1-3 | 4-15 | 16-21 | 22 have one shape each (mix 0 and 1 short-circuit the crossfade): groups of `ring`
        # frames, at most one extra launch per shape change
        assert lib.cv_kernel_launches() - before <= (frames_total + ring - 1) // ring + 4
    assert sum(s.frames for s in stats) == frames_total
    if static:
        assert [s.frames for s in stats] == [lib.cv_timeline_frames_of(g, 0, frames_total - 1, chunk, n_gpus) for g in range(n_gpus)]
    assert all(s.device_ms > 0 for s in stats if s.frames)
    assert sorted(results) == list(range(frames_total))            # every frame exactly once, whoever rendered it
    if pull:
        # frames 0 and 22 short-circuit the crossfade (mix 0 / 1): one picture pulled there, two everywhere else
        assert client.client_coded_clip_pulls(coded_clips["a"]) + client.client_coded_clip_pulls(coded_clips["b"]) >= 2 * frames_total - 2
        assert sum(s.h2d_bytes for s in stats) >= (2 * frames_total - 2) * w * h * 2

    orc = O.oracle()
    for f in range(frames_total):
        slot, cw = results[f]
        if static:
            assert slot == lib.cv_timeline_device_of(f, 0, chunk, n_gpus)
        src = f + 1
        t = (src - 1) / float(frames_total - 1)
        mix = float(np.float32(t)) if 0.0 < t < 1.0 else min(max(t, 0.0), 1.0)
        layers = []
        if 4 <= f < 16:
            layers.append(over1[(f - 4) % 2])
        layers.append(over2[(f + 3) % 4])
        arr = (O.Frame16 * len(layers))(*[l.c for l in layers])
        want = O.HostFrame16(full)
        # video_mix_cross_pull short-circuits mix 0 / 1 (video_mix.c:50-55); the oracle entry takes care of it
        orc.orc_timeline_frame_f16(C.byref(want.c), C.byref(pics_a[src % 3].c), C.byref(pics_b[src % 2].c), mix, arr, len(layers))
        assert cw == want.current.tuple()
        assert np.array_equal(host_bufs[f], want.array), "frame %d differs from the oracle" % f
    dev.init(0)


@pytest.mark.gpu
def test_timeline_render_one_gpu(lib):
    _timeline_on_gpus(lib, 1)
    _timeline_on_gpus(lib, 1, static=True)


@pytest.mark.gpu
def test_timeline_render_pulls_sources_from_the_host(lib):
    """coded_image_source.getFrame and video_source pulls through the worker's pinned ring and copy stream"""
    _timeline_on_gpus(lib, 1, ring=4, pull=True)
    _timeline_on_gpus(lib, 1, ring=2, pull=True, static=True)


@pytest.mark.gpu
def test_timeline_render_groups_frames_of_one_shape(lib):
    """with a deep output ring the worker submits groups of frames; frames of one shape share a persistent launch"""
    _timeline_on_gpus(lib, 1, ring=6, same_windows=True)


@pytest.mark.gpu
def test_timeline_render_all_gpus(lib):
    n = lib.cv_device_count()
    if n < 2:
        pytest.skip("one GPU visible")
    _timeline_on_gpus(lib, min(n, 8))
    _timeline_on_gpus(lib, min(n, 8), static=True)
    _timeline_on_gpus(lib, min(n, 8), ring=5, same_windows=True)
    _timeline_on_gpus(lib, min(n, 8), ring=4, pull=True)
