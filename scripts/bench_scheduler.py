#!/usr/bin/env python
"""BASELINE.json configs[4] through the C scheduler itself: a 3840x2160, 1000-frame synthetic timeline (two 4:2:2 clips
crossfading under a linear ramp + three f16 overlay clips with partial windows) rendered by ONE cv_timeline_render() call
that partitions the frames over the GPUs of this box (one host thread + one stream per GPU, no collective).

    python scripts/bench_scheduler.py --gpus N [--frames 1000] [--chunk 16] [--ring 16] [--sink none|u8]

Prints one JSON line: frames/s from the call's wall clock, the slowest GPU's device time, launches, per-GPU shares.
`--sink u8` also writes the 8-bit display encode of every frame and copies it to pinned host memory from the sink
(the end-to-end flavour: 33 MB per frame device->host).  This is a measurement of the scheduler, not the contract bench
(bench.py); the oracle is not involved (tests/test_timeline.py checks the same call frame by frame against it)."""
import argparse
import ctypes as C
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--frames", type=int, default=1000)
    ap.add_argument("--chunk", type=int, default=16)
    ap.add_argument("--ring", type=int, default=16)
    ap.add_argument("--clip", type=int, default=8, help="distinct source frames per clip and GPU (indexed f mod clip)")
    ap.add_argument("--sink", default="none", choices=["none", "u8"])
    ap.add_argument("--size", default="4k", choices=["4k", "1080p"])
    args = ap.parse_args()

    from canvas_b200 import _abi as A
    from canvas_b200 import device as dev
    lib = A.load()
    n = args.gpus
    if lib.cv_device_count() < n:
        raise SystemExit("only %d GPU(s) visible" % lib.cv_device_count())
    w, h = (3840, 2160) if args.size == "4k" else (1920, 1080)
    full = (0, 0, w - 1, h - 1)
    windows = [(w // 8, h // 8, w - 1, h - 1), (0, 0, 3 * w // 4 - 1, h - 1), (w // 4, 0, w - 1, 3 * h // 4 - 1)]
    rng = np.random.default_rng(7)
    base_uyvy = rng.integers(16, 236, size=(h, w * 2), dtype=np.uint8)
    base_layer = np.empty((h, w, 4), np.uint16)
    base_layer[..., :3] = rng.uniform(0.0, 1.0, size=(h, w, 3)).astype(np.float16).view(np.uint16)
    base_layer[..., 3] = rng.choice(np.array([0.0, 0.25, 0.5, 0.75, 1.0], np.float16), size=(h, w)).view(np.uint16)

    keep, clips = [], {}
    for name in ("a", "b", "o0", "o1", "o2"):
        arr = (A.Clip * n)()
        for g in range(n):
            dev.init(g)
            srcs = (A.Source * args.clip)()
            for k in range(args.clip):
                if name in ("a", "b"):
                    d = dev.DeviceCoded([np.roll(base_uyvy, 4 * (k + 1) + (64 if name == "b" else 0), axis=1)], w, h, 2,
                                        A.CV_LAYOUT_UYVY, A.CV_MATRIX_709)
                    srcs[k].kind, srcs[k].coded = A.CV_SRC_CODED, d.c
                else:
                    d = dev.DeviceFrame16.from_numpy(np.roll(base_layer, 17 * k + 1, axis=1), full, windows[int(name[1])])
                    srcs[k].kind, srcs[k].f16 = A.CV_SRC_F16, d.c
                keep.append(d)
            arr[g].frames, arr[g].count = srcs, args.clip
            keep.append(srcs)
        clips[name] = arr

    items = (A.TimelineItem * 4)()
    items[0].x, items[0].length, items[0].z, items[0].offset = 0, args.frames, 0, 0
    items[0].clips, items[0].clips_b = clips["a"], clips["b"]
    items[0].mix_first, items[0].mix_last, items[0].ramp_start, items[0].ramp_length = 0.0, 1.0, -1, args.frames + 1
    for i in range(3):
        it = items[1 + i]
        it.x, it.length, it.z, it.offset, it.clips = 0, args.frames, 1 + i, i, clips["o%d" % i]

    targets = (A.TimelineTarget * n)()
    pinned = []
    for g in range(n):
        dev.init(g)
        frames = (A.FrameF16 * args.ring)()
        for r in range(args.ring):
            f = dev.DeviceFrame16(full)
            keep.append(f)
            frames[r] = f.c
        targets[g].out, targets[g].ring = frames, args.ring
        if args.sink == "u8":
            table = dev.display_table(1.25)
            bufs = (C.c_void_p * args.ring)()
            for r in range(args.ring):
                b = dev.DeviceBuffer(w * h * 4)
                keep.append(b)
                bufs[r] = b.ptr.value
            targets[g].out_u8 = C.cast(bufs, C.POINTER(C.c_void_p))
            targets[g].u8_table = table.ptr if hasattr(table, 'ptr') else table
            host = C.c_void_p()
            dev.check(lib.cv_malloc(C.byref(host), w * h * 4 * args.ring, A.CV_MEM_PINNED), "cv_malloc(pinned)")
            pinned.append(host)
            keep += [table, bufs]
    dev.init(0)

    counters = [0] * n

    @A.TIMELINE_SINK
    def sink(user, slot, frame, job, stream):
        k = counters[slot] % args.ring
        counters[slot] += 1
        return lib.cv_download(C.c_void_p(pinned[slot].value + k * w * h * 4), job.contents.out_u8, w * h * 4, stream)

    run = A.TimelineRun()
    run.items, run.item_count, run.n_gpus = items, 4, n
    run.first, run.last, run.chunk = 0, args.frames - 1, args.chunk
    run.targets = targets
    if args.sink == "u8":
        run.sink = sink
    stats = (A.TimelineStats * n)()

    warm = A.TimelineRun()
    C.memmove(C.byref(warm), C.byref(run), C.sizeof(run))
    warm.last = min(args.frames, 4 * args.chunk * n) - 1
    dev.check(lib.cv_timeline_render(C.byref(warm), stats), "cv_timeline_render (warm-up)")

    before = lib.cv_kernel_launches()
    t0 = time.perf_counter()
    dev.check(lib.cv_timeline_render(C.byref(run), stats), "cv_timeline_render")
    wall = time.perf_counter() - t0
    device_ms = max(s.device_ms for s in stats)
    coverage = sum((b[2] - b[0] + 1) * (b[3] - b[1] + 1) for b in windows)
    bytes_per_frame = 2 * w * h * 2 + 8 * coverage + 8 * w * h + (4 * w * h if args.sink == "u8" else 0)
    print(json.dumps({
        "metric": "timeline frames/sec through cv_timeline_render", "value": args.frames / wall, "unit": "frames/s",
        "n_gpus": n, "frames": args.frames, "chunk": args.chunk, "ring": args.ring, "sink": args.sink, "size": args.size,
        "wall_ms": wall * 1e3, "slowest_gpu_device_ms": device_ms, "device_fps": args.frames / (device_ms * 1e-3),
        "gpu_launches": int(lib.cv_kernel_launches() - before), "frames_per_gpu": [int(s.frames) for s in stats],
        "algorithmic_gb_per_s": bytes_per_frame * args.frames / wall / 1e9}))


if __name__ == "__main__":
    main()
