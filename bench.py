#!/usr/bin/env python
"""bench.py -- frames/sec of the per-pixel frame path on B200, with its HBM roofline and the
reference CPU path timed beside it.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME]

Default workload (BASELINE.json: the configuration the metric is quoted on):
    4k_8layer_composite   3840x2160, 8 full-window f16 RGBA layers composited "over" -> f16 out,
                          72 algorithmic bytes per pixel = 597.2 MB per frame.
Other workloads (explicit --workload only; parity-test configurations, not the bench line):
    4k_composite_convert  bottom layer is an 8-bit 4:2:2 picture converted in-kernel + 7 f16 layers (66 B/px)
    4k_convert            4:2:2 Rec.709 -> linear f16 RGBA (10 B/px)
    1080p_convert         same at 1920x1080 (BASELINE.json configs[0]); 1080p_convert_planar: Y, Cb, Cr planes
    1080p_cross_u8        two f16 sources crossfaded, display-encoded to 8-bit (20 B/px)
    4k_timeline           2 x (4:2:2 -> f16) -> crossfade -> + 3 overlays -> f16 (36 B/px at full coverage)
    1080p_4layer_windows  4-layer workspace composite, layers 1-3 with partial windows (BASELINE.json configs[2];
                          bytes = 8 x (sum of layer window areas + output area) = 56.1 MB per frame)

A "step" renders 16 (4K) or 120 (1080p) frames with ONE cv_render_frames() call -- frames of one shape become one
persistent kernel launch; the plain 8-layer composite launches once per frame -- from a ring of input sets that
together exceed L2 several times.  `value` is device-resident throughput (CUDA events on the launching stream);
`e2e` moves every frame's inputs host->device from pinned memory and the result device->host
inside the timed region.  Multi-GPU: frames are sharded by index over ranks (one process per
GPU, no data-path collective; torch.distributed only provides the barrier and the max over ranks).
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

FRAMES_PER_STEP = 16
INPUT_SETS = 3          # at least; small frames get more so that the ring of sets stays larger than L2
RING_BYTES = 400e6      # >= 3 x the 126 MB L2


# dram__bytes_read.sum + dram__bytes_write.sum per FRAME, from the committed `ncu --set full` captures
# (profiles/r1_final.md; the windows workload from profiles/r1_general.md)
NCU_TRAFFIC_PER_FRAME = {
    "4k_8layer_composite": 588e6, "4k_timeline": 4881e6 / 16, "4k_convert": 1268e6 / 16, "4k_composite_convert": 8893e6 / 16,
    "1080p_convert": 2460e6 / 120, "1080p_convert_planar": 2512e6 / 120, "1080p_cross_u8": 4958e6 / 120,
    "1080p_4layer_windows": 6412e6 / 120,
}


def frames_per_step(workload):
    """4K: 16 frames a step; 1080p: 120 (BASELINE.json configs[0] is quoted on 120 frames), so that a step is a few ms"""
    return FRAMES_PER_STEP if workload.startswith("4k") else 120


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


WORKLOADS = {
    # name: (width, height, bytes per pixel, description)
    "4k_8layer_composite": (3840, 2160, 72, "3840x2160 8-layer f16 alpha-over workspace composite -> f16"),
    "4k_composite_convert": (3840, 2160, 66, "3840x2160 4:2:2 bottom converted in-kernel + 7 f16 layers -> f16"),
    "4k_convert": (3840, 2160, 10, "3840x2160 8-bit 4:2:2 Rec.709 -> linear f16 RGBA"),
    "1080p_convert": (1920, 1080, 10, "1920x1080 8-bit 4:2:2 Rec.709 -> linear f16 RGBA"),
    "1080p_convert_planar": (1920, 1080, 10, "1920x1080 8-bit planar 4:2:2 Rec.709 (Y, Cb, Cr planes) -> linear f16 RGBA"),
    "1080p_cross_u8": (1920, 1080, 20, "1920x1080 two-source f16 crossfade -> 8-bit display encode"),
    "4k_timeline": (3840, 2160, 36, "3840x2160 2x(4:2:2->f16) crossfade + 3 f16 overlays -> f16"),
    "1080p_4layer_windows": (1920, 1080, None, "1920x1080 4-layer alpha-over workspace composite, partial layer windows -> f16"),
}


# ----------------------------------------------------------------------------- synthetic inputs

# layer windows of 1080p_4layer_windows (layer 0 covers the frame); inclusive corners
WINDOWS_1080P = [(0, 0, 1919, 1079), (200, 100, 1719, 979), (480, 0, 1919, 700), (0, 540, 959, 1079)]


def workload_bytes(name):
    """Algorithmic bytes per frame: every input pixel read once, every output pixel written once."""
    w, h, per_px, _ = WORKLOADS[name]
    if per_px is not None:
        return w * h * per_px
    area = lambda b: (b[2] - b[0] + 1) * (b[3] - b[1] + 1)
    return 8 * (sum(area(b) for b in WINDOWS_1080P) + w * h)


def synth_layer(rng, h, w):
    """Raw half bits: rgb uniform halves in [0,1], alpha from {0,1/4,1/2,3/4,1,1} or a random fraction."""
    a = np.empty((h, w, 4), np.uint16)
    a[..., :3] = rng.integers(0, 0x3C01, size=(h, w, 3), dtype=np.uint16)
    choice = rng.integers(0, 8, size=(h, w), dtype=np.uint8)
    table = np.array([0x0000, 0x3400, 0x3800, 0x3A00, 0x3C00, 0x3C00, 0, 0], np.uint16)
    alpha = table[choice]
    frac = rng.integers(0x2000, 0x3C00, size=(h, w), dtype=np.uint16)
    a[..., 3] = np.where(choice >= 6, frac, alpha)
    return a


PLANAR = False
CONTENT = "noise"        # --content: "noise" = independent uniform bytes (the default; the worst case for the table gathers),
                         #            "video" = smooth gradients + 2 LSB of noise, the spatial coherence of camera pictures


def synth_uyvy(rng, h, w):
    if CONTENT == "noise":
        return rng.integers(16, 236, size=(h, w * 2), dtype=np.uint8)
    y, x = np.mgrid[0:h, 0:w * 2].astype(np.float32)
    phase = rng.uniform(0, 6.28, 3)
    p = 126 + 60 * np.sin(x / (w * 0.31) + phase[0]) + 40 * np.sin(y / (h * 0.23) + phase[1]) + 8 * np.sin((x + y) / 97.0 + phase[2])
    p += rng.integers(-2, 3, size=p.shape)
    return np.clip(p, 16, 235).astype(np.uint8)


class ClockSampler(threading.Thread):
    """Samples nvidia-smi clocks / throttle reasons while the timed region runs."""

    FIELDS = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu):
        super().__init__(daemon=True)
        self.gpu, self.samples, self.stop_flag = gpu, [], threading.Event()

    def run(self):
        while not self.stop_flag.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.FIELDS,
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                parts = [p.strip() for p in out.strip().split(",")]
                if len(parts) >= 6:
                    self.samples.append(parts)
            except Exception:
                pass
            self.stop_flag.wait(0.2)

    def finish(self):
        self.stop_flag.set()
        self.join(timeout=6)
        sm = [float(s[0]) for s in self.samples if s[0].replace(".", "").isdigit()]
        mx = [float(s[1]) for s in self.samples if s[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for s in self.samples for n, v in zip(names, s[2:6]) if v.lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(self.samples)}


# ----------------------------------------------------------------------------- our arm

class PinnedArray:
    """A numpy view over cudaMallocHost memory."""

    def __init__(self, dev, shape, dtype):
        self.dev = dev
        self.nbytes = int(np.prod(shape)) * np.dtype(dtype).itemsize
        self.ptr = C.c_void_p()
        dev.check(dev.lib().cv_malloc(C.byref(self.ptr), self.nbytes, 1), "cv_malloc(pinned)")
        buf = (C.c_uint8 * self.nbytes).from_address(self.ptr.value)
        self.array = np.frombuffer(buf, dtype=dtype).reshape(shape)


class InputSet:
    """Device-resident inputs + output of one frame, with pinned host mirrors for the e2e leg."""

    def __init__(self, dev, name, rng, base):
        w, h, _, _ = WORKLOADS[name]
        full = (0, 0, w - 1, h - 1)
        self.dev, self.full = dev, full
        self.layers, self.coded, self.host, self.uploads = [], [], [], []
        self.out = dev.DeviceFrame16(full)
        self.out_u8 = None
        self.d2h_bytes = w * h * 8

        def add_layer(i, window=None):
            arr = np.roll(base["layer"][i % len(base["layer"])], 17 * len(self.layers) + 1, axis=1)
            frame = dev.DeviceFrame16.from_numpy(arr, full, window)
            pinned = PinnedArray(dev, arr.shape, np.uint16)
            pinned.array[...] = arr
            self.layers.append(frame)
            self.uploads.append((frame.c.data, pinned))

        def add_coded(i, planar=False):
            planar = planar or PLANAR           # --planar: every coded source as three planes (the reference's coded_image layout)
            arr = np.roll(base["uyvy"][i % len(base["uyvy"])], 4 * (len(self.coded) + 1), axis=1)
            if planar:          # the same samples as three planes: UYVY = Cb Y0 Cr Y1
                planes = [np.ascontiguousarray(arr[:, 1::2]), np.ascontiguousarray(arr[:, 0::4]), np.ascontiguousarray(arr[:, 2::4])]
                pic = dev.DeviceCoded(planes, w, h, 2, dev._abi.CV_LAYOUT_PLANAR, dev.CV_MATRIX_709)
            else:
                planes = [arr]
                pic = dev.DeviceCoded(planes, w, h, 2, dev.CV_LAYOUT_UYVY, dev.CV_MATRIX_709)
            self.coded.append(pic)
            for buffer, plane in zip(pic.buffers, planes):
                pinned = PinnedArray(dev, plane.shape, np.uint8)
                pinned.array[...] = plane
                self.uploads.append((buffer.ptr, pinned))

        if name == "4k_8layer_composite":
            for i in range(8):
                add_layer(i)
            self.job = dev.make_job(out=self.out, layers=self.layers)
        elif name == "4k_composite_convert":
            add_coded(0)
            for i in range(7):
                add_layer(i)
            self.job = dev.make_job(out=self.out, a=self.coded[0], layers=self.layers)
        elif name in ("4k_convert", "1080p_convert", "1080p_convert_planar"):
            add_coded(0, planar=name.endswith("planar"))
            self.job = dev.make_job(out=self.out, a=self.coded[0])
        elif name == "1080p_cross_u8":
            add_layer(0)
            add_layer(1)
            self.out_u8 = dev.DeviceBuffer(w * h * 4)
            self.table = base["table"]
            self.job = dev.make_job(a=self.layers[0], b=self.layers[1], mix_b=0.375, out_u8=self.out_u8,
                                    u8_table=self.table, full=full)
            self.d2h_bytes = w * h * 4
        elif name == "4k_timeline":
            add_coded(0)
            add_coded(1)
            for i in range(3):
                add_layer(i)
            self.job = dev.make_job(out=self.out, a=self.coded[0], b=self.coded[1], mix_b=0.375, layers=self.layers)
        elif name == "1080p_4layer_windows":
            for i, window in enumerate(WINDOWS_1080P):
                add_layer(i, window)
            self.job = dev.make_job(out=self.out, layers=self.layers)
        else:
            raise SystemExit("unknown workload %r" % name)

        self.h2d_bytes = sum(p.nbytes for _, p in self.uploads)
        self.result = PinnedArray(dev, (self.d2h_bytes,), np.uint8)

    def upload_inputs(self, stream):
        for dst, pinned in self.uploads:
            self.dev.check(self.dev.lib().cv_upload(dst, pinned.ptr, pinned.nbytes, stream), "cv_upload")

    def download_result(self, stream):
        src = self.out_u8.ptr if self.out_u8 is not None else self.out.c.data
        self.dev.check(self.dev.lib().cv_download(self.result.ptr, src, self.d2h_bytes, stream), "cv_download")


def make_base(dev, name, seed):
    """A few distinct random source images; input sets use shifted copies (distinct memory, cheap to make)."""
    w, h, _, _ = WORKLOADS[name]
    rng = np.random.default_rng(seed)
    base = {"layer": [synth_layer(rng, h, w) for _ in range(2)], "uyvy": [synth_uyvy(rng, h, w) for _ in range(2)]}
    if name == "1080p_cross_u8":
        base["table"] = dev.display_table(1.25)
    return base


def dist_setup(n_gpus):
    """torch.distributed for the barrier / max-over-ranks only (no data-path collective)."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world <= 1:
        return None, 0, 1
    import torch
    import torch.distributed as dist
    rank, local = int(os.environ["RANK"]), int(os.environ.get("LOCAL_RANK", os.environ["RANK"]))
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    use_nccl = torch.cuda.is_available()
    if use_nccl:
        torch.cuda.set_device(local)
    dist.init_process_group("nccl" if use_nccl else "gloo", rank=rank, world_size=world)
    return dist, rank, world


def dist_max(dist, value):
    if dist is None:
        return value
    import torch
    t = torch.tensor([value], dtype=torch.float64, device="cuda" if dist.get_backend() == "nccl" else "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def dist_sum(dist, value):
    if dist is None:
        return value
    import torch
    t = torch.tensor([value], dtype=torch.float64, device="cuda" if dist.get_backend() == "nccl" else "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())


def run_ours(args):
    from canvas_b200 import device as dev
    for k in ("CV_LAYOUT_UYVY", "CV_MATRIX_709"):
        setattr(dev, k, getattr(dev._abi, k))

    dist, rank, world = dist_setup(args.gpus)
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dev.init(local)
    lib = dev.lib()

    name = args.workload
    w, h, _, description = WORKLOADS[name]
    frame_bytes = workload_bytes(name)
    batch = not args.no_batch      # one cv_render_frames() call per step (the library launches one kernel per frame)
    FRAMES_PER_STEP = frames_per_step(name)

    base = make_base(dev, name, 1234 + rank)
    n_sets = max(INPUT_SETS, int(np.ceil(RING_BYTES / frame_bytes)))
    sets = [InputSet(dev, name, None, base) for _ in range(n_sets)]
    stream = C.c_void_p()
    dev.check(lib.cv_stream_create(C.byref(stream)), "cv_stream_create")

    def step(first_frame):
        if batch:
            jobs = [sets[(first_frame + f) % n_sets].job for f in range(FRAMES_PER_STEP)]
            dev.render_frames(jobs, stream)
        else:
            for f in range(FRAMES_PER_STEP):
                dev.render_frames([sets[(first_frame + f) % n_sets].job], stream)

    def barrier():
        dev.check(lib.cv_device_sync(), "cv_device_sync")
        if dist is not None:
            dist.barrier()
        dev.check(lib.cv_device_sync(), "cv_device_sync")

    # ---- device-resident throughput ------------------------------------------------------------
    frame = 0
    for _ in range(args.warmup):
        step(frame)
        frame += FRAMES_PER_STEP
    barrier()

    sampler = ClockSampler(local)
    sampler.start()
    launches_before = lib.cv_kernel_launches()
    timer = dev.Timer(stream)
    timer.begin()
    for _ in range(args.steps):
        step(frame)
        frame += FRAMES_PER_STEP
    timer.end()
    ms = timer.elapsed_ms()
    barrier()
    launches = lib.cv_kernel_launches() - launches_before
    clocks = sampler.finish()

    ms_max = dist_max(dist, ms)
    frames_rank = args.steps * FRAMES_PER_STEP
    frames_total = dist_sum(dist, float(frames_rank))
    value = frames_total / (ms_max * 1e-3)

    # ---- end to end: pinned host inputs -> HBM -> kernel -> pinned host result ------------------
    streams = []
    for _ in range(min(n_sets, 4)):
        s = C.c_void_p()
        dev.check(lib.cv_stream_create(C.byref(s)), "cv_stream_create")
        streams.append(s)

    def e2e_frames(count, first):
        for f in range(count):
            i = (first + f) % n_sets
            st = streams[i % len(streams)]
            sets[i].upload_inputs(st)
            dev.render_frames([sets[i].job], st)
            sets[i].download_result(st)

    e2e_steps = max(1, min(args.steps, 4))
    e2e_frames(n_sets, 0)
    barrier()
    t0 = time.perf_counter()
    e2e_frames(e2e_steps * FRAMES_PER_STEP, 0)
    for s in streams:
        dev.check(lib.cv_stream_sync(s), "cv_stream_sync")
    e2e_s = time.perf_counter() - t0
    barrier()
    e2e_s = dist_max(dist, e2e_s)
    e2e_value = dist_sum(dist, float(e2e_steps * FRAMES_PER_STEP)) / e2e_s

    peak, peak_source = load_peaks()
    kernel_ms = ms / max(1, launches)
    frames_per_launch = frames_rank / max(1, launches)
    achieved = frame_bytes * frames_per_launch / (kernel_ms * 1e-3) / 1e9

    line = {
        "metric": "f16 composite frames/sec (%s)" % name, "value": value, "unit": "frames/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_max / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f16 pixels, f32 arithmetic", "data": ("synthetic" if CONTENT == "noise" else "synthetic (%s-like coded sources)" % CONTENT) + (", planar 4:2:2" if PLANAR else ""),
        "config": {"workload": name, "description": description, "width": w, "height": h,
                   "frames_per_step": FRAMES_PER_STEP, "input_sets": n_sets,
                   "l2": "inputs larger than L2 (%.0f MB per frame, %d rotating sets)" % (frame_bytes / 1e6, n_sets),
                   "sharding": "frames by index, one process per GPU, no collective"},
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": (NCU_TRAFFIC_PER_FRAME[name] * frames_per_launch if name in NCU_TRAFFIC_PER_FRAME else None),
                     "traffic_source": "profiles/r1_final.md (ncu --set full, per frame x frames per launch)",
                     "peak_source": peak_source, "algorithmic_bytes_per_frame": frame_bytes,
                     "kernel_ms": kernel_ms},
        "e2e": {"value": e2e_value, "unit": "frames/s", "h2d_bytes_per_step": sets[0].h2d_bytes * FRAMES_PER_STEP,
                "d2h_bytes_per_step": sets[0].d2h_bytes * FRAMES_PER_STEP, "steps": e2e_steps},
        "gpu_launches": int(launches), "clocks": clocks,
    }

    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline(name, seconds=20.0)

    if rank == 0:
        print(json.dumps(line))
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


# ----------------------------------------------------------------------------- CPU baseline / reference arm

def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def cpu_threads(name):
    """All usable host cores, bounded by memory (each in-flight 4K frame needs ~0.5 GB of f32 temporaries)."""
    w, h, _, _ = WORKLOADS[name]
    per_thread = w * h * (8 + 16 + 16 + 8 + 8) * 1.2
    try:
        import psutil
        budget = psutil.virtual_memory().available * 0.5
    except Exception:
        budget = 16e9
    return int(max(1, min(host_cores(), budget // per_thread)))


def cpu_run(name, frames, threads):
    """Renders `frames` frames of workload `name` on the CPU, `threads` frames in flight.
    4k_8layer_composite runs the REFERENCE's own C code (oracle/_ref/libcanvasref_fast.so, built with the
    reference's flags: workspace.c -> video_mix_over_f32 through video_get_frame_f16); the workloads whose
    sizes/layouts the reference hard-codes out (4:2:2, 1080p/4K conversion) run the oracle port.
    Returns (seconds, kind)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oraclelib as O
    w, h, _, _ = WORKLOADS[name]
    full = (0, 0, w - 1, h - 1)
    rng = np.random.default_rng(99)

    if name == "4k_8layer_composite" and os.path.exists(O.REF_FAST_SO):
        L = O.reference(fast=True)
        layers = [O.HostFrame16(full, full, synth_layer(rng, h, w)) for _ in range(8)]
        frames_c = (O.Frame16 * 8)(*[l.c for l in layers])
        starts = (C.c_int * 8)(*range(8))
        counts = (C.c_int * 8)(*([1] * 8))
        i64 = C.c_int64 * 8
        tl = L.ref_timeline_create(frames_c, starts, counts, i64(*([0] * 8)), i64(*([1 << 30] * 8)), i64(*([0] * 8)), i64(*range(8)), 8)
        outs = [O.HostFrame16(full) for _ in range(threads)]
        outs_c = (O.Frame16 * threads)(*[o.c for o in outs])
        L.ref_time_timeline(tl, 0, min(threads, frames), threads, outs_c, threads)        # warm: page in the buffers
        ns = L.ref_time_timeline(tl, 0, frames, threads, outs_c, threads)
        L.ref_timeline_free(tl)
        return ns * 1e-9, "reference"

    # oracle port, frame-parallel through a thread pool (ctypes releases the GIL)
    from concurrent.futures import ThreadPoolExecutor
    orc = O.oracle()
    layers = [O.HostFrame16(full, full, synth_layer(rng, h, w)) for _ in range(8 if "8layer" in name else (7 if "composite" in name else (4 if "4layer" in name else 3)))]
    windowed = [O.HostFrame16(full, win, l.array) for l, win in zip(layers, WINDOWS_1080P)] if name == "1080p_4layer_windows" else []
    pics = [O.synth_coded(w, h, 5 + i, 2, O.LAYOUT_UYVY, O.MATRIX_709) for i in range(2)]
    srgb, ramp = np.zeros(65536, np.uint16), np.zeros(65536, np.uint8)
    orc.orc_build_lut(O.LUT_LINEAR_TO_SRGB, O.ptr(srgb))
    orc.orc_build_intent_ramp(1.25, O.ptr(ramp))
    outs = [O.HostFrame16(full) for _ in range(threads)]
    u8s = [np.zeros((h, w, 4), np.uint8) for _ in range(threads)]

    def one(slot):
        out = outs[slot]
        if name == "4k_8layer_composite":
            arr = (O.Frame16 * 8)(*[l.c for l in layers])
            orc.orc_composite_f16(C.byref(out.c), arr, 8)
        elif name == "4k_composite_convert":
            arr = (O.Frame16 * 7)(*[l.c for l in layers[:7]])
            orc.orc_timeline_frame_f16(C.byref(out.c), C.byref(pics[0].c), None, 0.0, arr, 7)
        elif name in ("4k_convert", "1080p_convert", "1080p_convert_planar"):
            orc.orc_reconstruct(C.byref(out.c), C.byref(pics[0].c))
        elif name == "1080p_cross_u8":
            orc.orc_cross_f16(C.byref(out.c), C.byref(layers[0].c), C.byref(layers[1].c), 0.375)
            orc.orc_encode_u8(C.byref(out.c), O.ptr(srgb), O.ptr(ramp), O.ptr(u8s[slot]))
        elif name == "4k_timeline":
            arr = (O.Frame16 * 3)(*[l.c for l in layers[:3]])
            orc.orc_timeline_frame_f16(C.byref(out.c), C.byref(pics[0].c), C.byref(pics[1].c), 0.375, arr, 3)
        elif name == "1080p_4layer_windows":
            arr = (O.Frame16 * 4)(*[l.c for l in windowed])
            orc.orc_composite_f16(C.byref(out.c), arr, 4)

    def worker(slot):
        n = frames // threads + (1 if slot < frames % threads else 0)
        for _ in range(n):
            one(slot)

    with ThreadPoolExecutor(threads) as pool:
        list(pool.map(one, range(threads)))             # warm-up: one frame per thread
        t0 = time.perf_counter()
        list(pool.map(worker, range(threads)))
        seconds = time.perf_counter() - t0
    return seconds, "port"


def cpu_baseline(name, seconds=20.0):
    """Bounded sample of the same workload on the host cores (about `seconds` of CPU work)."""
    threads = cpu_threads(name)
    t, kind = cpu_run(name, threads, threads)                   # calibration: one frame per thread
    frames = int(max(threads, min(64 * threads, threads * round(seconds / max(t, 1e-3)))))
    t, kind = cpu_run(name, frames, threads)
    return {"value": frames / t, "unit": "frames/s", "cores": threads, "kind": kind,
            "sample": "%d frames of %s, %d frames in flight (one per thread), %.1f s" % (frames, name, threads, t)}


def run_reference(args):
    """--impl reference: the reference's own CPU path (or the oracle port where it has none) on the host cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    name = args.workload
    w, h, _, description = WORKLOADS[name]
    threads = cpu_threads(name)
    t1, kind = cpu_run(name, threads, threads)
    per_step = int(max(threads, min(16 * threads, threads * round(8.0 / max(t1, 1e-3)))))   # ~8 s per step
    for _ in range(max(0, min(args.warmup, 1))):
        cpu_run(name, threads, threads)
    total_t = 0.0
    steps = max(1, min(args.steps, 3))
    for _ in range(steps):
        t, kind = cpu_run(name, per_step, threads)
        total_t += t
    value = steps * per_step / total_t
    line = {
        "impl": "reference", "metric": "f16 composite frames/sec (%s)" % name, "value": value, "unit": "frames/s",
        "n_gpus": int(os.environ.get("WORLD_SIZE", "1")), "steps": steps, "warmup": min(args.warmup, 1),
        "ms_per_step": total_t / steps * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f16 pixels, f32 arithmetic", "data": ("synthetic" if CONTENT == "noise" else "synthetic (%s-like coded sources)" % CONTENT) + (", planar 4:2:2" if PLANAR else ""),
        "config": {"workload": name, "description": description, "width": w, "height": h,
                   "frames_per_step": per_step, "note": "CPU reference arm: steps bounded to %d x %d frames" % (steps, per_step)},
        "cpu_baseline": {"value": value, "unit": "frames/s", "cores": threads, "kind": kind,
                         "sample": "%d steps x %d frames of %s, %d frames in flight" % (steps, per_step, name, threads)},
        "e2e": {"value": value, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="4k_8layer_composite", choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--planar", action="store_true", help="4:2:2 sources as three planes instead of packed UYVY")
    ap.add_argument("--content", default="noise", choices=["noise", "video"],
                    help="coded sources: independent random bytes (default) or smooth picture-like content")
    ap.add_argument("--no-batch", action="store_true", help="one launch per frame even for microsecond kernels")
    args = ap.parse_args()
    global CONTENT, PLANAR
    CONTENT, PLANAR = args.content, args.planar
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
