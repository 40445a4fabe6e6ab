"""Host time of cv_render_frames() per job against the device time of the same batch (is a workload host-bound?)."""
import ctypes as C
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench

import argparse
args = argparse.Namespace(gpus=1, no_numa=False, no_batch=False)
B = bench.Bench(args)
dev, lib = B.dev, B.lib
for name in sys.argv[1:] or ["1080p_convert", "1080p_cross_u8", "4k_convert"]:
    sets = B.build(name, pinned=False)
    n = len(sets)
    per_step = max(n, (bench.WORKLOADS[name][3] // n) * n)
    jobs = (dev.FrameJob * per_step)(*[sets[f % n].job for f in range(per_step)])
    for _ in range(3):
        dev.check(lib.cv_render_frames(jobs, per_step, B.stream), "render")
    dev.sync(B.stream)
    t = dev.Timer(B.stream)
    t.begin()
    t0 = time.perf_counter()
    reps = 6
    for _ in range(reps):
        dev.check(lib.cv_render_frames(jobs, per_step, B.stream), "render")
    host = time.perf_counter() - t0
    t.end()
    ms = t.elapsed_ms()
    print("%-22s %6d jobs/call  host %.2f us/job  device %.2f us/job" % (name, per_step, host * 1e6 / (reps * per_step), ms * 1e3 / (reps * per_step)))
    for s in sets:
        s.free()

# the host's own cost per job: the same call on frames so small that the device is never the limit
import numpy as np
w, h, count = 128, 16, 8192
full = (0, 0, w - 1, h - 1)
pics = [dev.DeviceCoded([np.full((h, 2 * w), 128, np.uint8)], w, h, 2, bench.dev_abi().CV_LAYOUT_UYVY if hasattr(bench, "dev_abi") else dev._abi.CV_LAYOUT_UYVY, dev._abi.CV_MATRIX_709) for _ in range(4)]
outs = [dev.DeviceFrame16(full) for _ in range(4)]
small = [dev.make_job(out=outs[i % 4], a=pics[i % 4]) for i in range(count)]
arr = (dev.FrameJob * count)(*small)
for _ in range(2):
    dev.check(lib.cv_render_frames(arr, count, B.stream), "render")
dev.sync(B.stream)
t0 = time.perf_counter()
for _ in range(4):
    dev.check(lib.cv_render_frames(arr, count, B.stream), "render")
host = time.perf_counter() - t0
dev.sync(B.stream)
total = time.perf_counter() - t0
print("tiny frames (%dx%d conversions): host %.2f us/job to enqueue, %.2f us/job until done" % (w, h, host * 1e6 / (4 * count), total * 1e6 / (4 * count)))

# the same for windowed layer stacks (k_rect: the host plans rectangles of constant coverage)
w, h, count = 256, 32, 4096
full = (0, 0, w - 1, h - 1)
wins = [full, (16, 4, 200, 28), (64, 0, 255, 20), (0, 10, 127, 31)]
layers = [[dev.DeviceFrame16(full, current=win) for win in wins] for _ in range(4)]
for group in layers:
    for f in group:
        f.fill(0x3C)
outs = [dev.DeviceFrame16(full) for _ in range(4)]
small = [dev.make_job(out=outs[i % 4], layers=layers[i % 4]) for i in range(count)]
arr = (dev.FrameJob * count)(*small)
for _ in range(2):
    dev.check(lib.cv_render_frames(arr, count, B.stream), "render")
dev.sync(B.stream)
t0 = time.perf_counter()
for _ in range(4):
    dev.check(lib.cv_render_frames(arr, count, B.stream), "render")
host = time.perf_counter() - t0
dev.sync(B.stream)
total = time.perf_counter() - t0
print("tiny frames (%dx%d, 4 windowed layers): host %.2f us/job to enqueue, %.2f us/job until done" % (w, h, host * 1e6 / (4 * count), total * 1e6 / (4 * count)))
