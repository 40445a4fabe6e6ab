#!/usr/bin/env python
"""bench.py -- frames/sec of the per-pixel frame path on B200, with its HBM roofline and the
reference CPU path timed beside it.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME]

Headline workload (BASELINE.json: the configuration the metric is quoted on):
    4k_8layer_composite   3840x2160, 8 full-window f16 RGBA layers composited "over" -> f16 out,
                          72 algorithmic bytes per pixel = 597.2 MB per frame.
The other BASELINE.json configurations are measured in the same run (each about 0.3 s of device time) and
reported in the line's `other_workloads` array; `--workload NAME` makes any of them the headline instead:
    4k_composite_convert  bottom layer is an 8-bit 4:2:2 picture converted in-kernel + 7 f16 layers (66 B/px)
    4k_convert            4:2:2 Rec.709 -> linear f16 RGBA (10 B/px)
    1080p_convert         same at 1920x1080 (BASELINE.json configs[0]); 1080p_convert_planar: Y, Cb, Cr planes
    1080p_cross_u8        two f16 sources crossfaded, display-encoded to 8-bit (20 B/px; configs[1])
    1080p_4layer_windows  4-layer workspace composite, layers 1-3 with partial windows (configs[2];
                          bytes = 8 x (sum of layer window areas + output area) = 56.1 MB per frame)
    4k_timeline           2 x (4:2:2 -> f16) -> crossfade -> + 3 full-window overlays -> f16 (36 B/px)
    4k_timeline_windows   the same with PARTIAL overlay windows (configs[4]'s frame; 28.6 B/px)
    4k_4layer_filters     4 full-window layers, three of them with per-pixel filters fused into the launch (40 B/px)

A "step" renders enough frames to last about 50 ms (512 frames at 4K for the headline) with ONE
cv_render_frames() call -- frames of one shape become persistent kernel launches of up to 256 frames; the plain
8-layer composite launches once per frame -- from a ring of input sets that together exceed L2 several times.
`value` is device-resident throughput (CUDA events on the launching stream); `e2e` moves every frame's inputs
host->device from pinned memory and the result device->host inside the timed region.  After the timed region the
outputs the timed launches wrote are compared with the oracle (`parity_checked`, `max_ulp`).
Multi-GPU: frames are sharded by index over ranks (one process per GPU, no data-path collective;
torch.distributed only provides the barrier and the max over ranks); `scheduler` is the same timeline rendered by
ONE cv_timeline_render() call over all GPUs (rank 0, while the other ranks idle).
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

INPUT_SETS = 3          # at least; small frames get more so that the ring of sets stays larger than L2
RING_BYTES = 400e6      # >= 3 x the 126 MB L2
STEP_MS = 50.0          # a step lasts about this long, so that 20 steps are a second of sustained load

WORKLOADS = {
    # name: (width, height, bytes per pixel, frames per step (about STEP_MS of device time), description)
    "4k_8layer_composite": (3840, 2160, 72, 512, "3840x2160 8-layer f16 alpha-over workspace composite -> f16"),
    "4k_composite_convert": (3840, 2160, 66, 448, "3840x2160 4:2:2 bottom converted in-kernel + 7 f16 layers -> f16"),
    "4k_convert": (3840, 2160, 10, 3072, "3840x2160 8-bit 4:2:2 Rec.709 -> linear f16 RGBA"),
    "1080p_convert": (1920, 1080, 10, 12000, "1920x1080 8-bit 4:2:2 Rec.709 -> linear f16 RGBA"),
    "1080p_convert_planar": (1920, 1080, 10, 9600, "1920x1080 8-bit planar 4:2:2 Rec.709 (Y, Cb, Cr planes) -> linear f16 RGBA"),
    "1080p_cross_u8": (1920, 1080, 20, 6000, "1920x1080 two-source f16 crossfade -> 8-bit display encode"),
    "4k_timeline": (3840, 2160, 36, 640, "3840x2160 2x(4:2:2->f16) crossfade + 3 f16 overlays -> f16"),
    "4k_timeline_windows": (3840, 2160, None, 512, "3840x2160 2x(4:2:2->f16) crossfade + 3 f16 overlays with partial windows -> f16"),
    "1080p_4layer_windows": (1920, 1080, None, 3840, "1920x1080 4-layer alpha-over workspace composite, partial layer windows -> f16"),
    "4k_4layer_filters": (3840, 2160, 40, 512, "3840x2160 4-layer composite, layers 1-3 with fused per-pixel filters (colour transform / gain+offset / crop) -> f16"),
}
# what the driver-run line reports next to the headline: BASELINE.json configs[0..4] and their variants
OTHER_WORKLOADS = ["1080p_convert", "1080p_convert_planar", "1080p_cross_u8", "1080p_4layer_windows", "4k_composite_convert",
                   "4k_convert", "4k_timeline", "4k_timeline_windows", "4k_4layer_filters"]

# the filters fused into layers 1..3 of 4k_4layer_filters (cv_layer_ops): flags, colour transform, gain, offset, crop
LAYER_FILTERS = [None, (2, 1, 0.0, 0.0, None), (4, 0, 1.25, -0.03125, None), (4 | 8, 0, 0.75, 0.0625, (400, 200, 3400, 1900))]

# layer windows of 1080p_4layer_windows (layer 0 covers the frame); inclusive corners
WINDOWS_1080P = [(0, 0, 1919, 1079), (200, 100, 1719, 979), (480, 0, 1919, 700), (0, 540, 959, 1079)]


def timeline_windows(w, h):
    """Overlay windows of the configs[4] timeline frame (the same boxes scripts/bench_scheduler.py uses)."""
    return [(w // 8, h // 8, w - 1, h - 1), (0, 0, 3 * w // 4 - 1, h - 1), (w // 4, 0, w - 1, 3 * h // 4 - 1)]


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def load_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum per FRAME by workload, written by scripts/ncu_summary.py from
    the committed `ncu --set full` captures (profiles/traffic.json)."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            return json.load(f)
    except Exception:
        return {}


def workload_bytes(name):
    """Algorithmic bytes per frame: every input pixel read once, every output pixel written once."""
    w, h, per_px = WORKLOADS[name][:3]
    if per_px is not None:
        return w * h * per_px
    area = lambda b: (b[2] - b[0] + 1) * (b[3] - b[1] + 1)
    if name == "4k_timeline_windows":
        return 4 * w * h + 8 * (sum(area(b) for b in timeline_windows(w, h)) + w * h)
    return 8 * (sum(area(b) for b in WINDOWS_1080P) + w * h)


# ----------------------------------------------------------------------------- synthetic inputs

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
            self.stop_flag.wait(0.1)

    def finish(self):
        self.stop_flag.set()
        self.join(timeout=6)
        sm = [float(s[0]) for s in self.samples if s[0].replace(".", "").isdigit()]
        mx = [float(s[1]) for s in self.samples if s[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for s in self.samples for n, v in zip(names, s[2:6]) if v.lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(self.samples)}


# ----------------------------------------------------------------------------- NUMA placement

def gpu_numa_cpus(local_gpu):
    """The CPUs of the NUMA node the GPU's PCIe root hangs off (None when the box does not say)."""
    try:
        out = subprocess.run(["nvidia-smi", "-i", str(local_gpu), "--query-gpu=pci.bus_id", "--format=csv,noheader"],
                             capture_output=True, text=True, timeout=10).stdout.strip().lower()
        if not out:
            return None
        bus = out[-12:] if len(out) > 12 else out          # 00000000:1B:00.0 -> 0000:1b:00.0
        node = int(open("/sys/bus/pci/devices/%s/numa_node" % bus).read())
        if node < 0:
            return None
        cpus = set()
        for part in open("/sys/devices/system/node/node%d/cpulist" % node).read().strip().split(","):
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        usable = cpus & set(os.sched_getaffinity(0))
        return (node, sorted(usable)) if usable else None
    except Exception:
        return None


def pin_to_gpu_node(local_gpu):
    """Runs this process on the GPU's NUMA node, so that pinned staging memory is first-touched there."""
    found = gpu_numa_cpus(local_gpu)
    if not found:
        return None
    node, cpus = found
    try:
        os.sched_setaffinity(0, cpus)
        return {"node": node, "cpus": len(cpus)}
    except Exception:
        return None


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

    def free(self):
        if self.ptr:
            self.array = None
            self.dev.lib().cv_free(self.ptr, 1)
            self.ptr = C.c_void_p()


class InputSet:
    """Device-resident inputs + output of one frame, pinned host mirrors for the e2e leg and the host copies the
    oracle check reads (`host_layers`: (array, window); `host_coded`: (UYVY rows, planar?))."""

    def __init__(self, dev, name, base, pinned=True):
        w, h = WORKLOADS[name][:2]
        full = (0, 0, w - 1, h - 1)
        self.dev, self.full, self.name = dev, full, name
        self.layers, self.coded, self.uploads = [], [], []
        self.host_layers, self.host_coded = [], []
        self.out = dev.DeviceFrame16(full)
        self.out_u8 = None
        self.mix_b = 0.375
        self.d2h_bytes = w * h * 8

        def mirror(dst, arr, dtype):
            if pinned:
                p = PinnedArray(dev, arr.shape, dtype)
                p.array[...] = arr
                self.uploads.append((dst, p))

        def add_layer(i, window=None):
            arr = np.roll(base["layer"][i % len(base["layer"])], 17 * len(self.layers) + 1, axis=1)
            frame = dev.DeviceFrame16.from_numpy(arr, full, window)
            self.layers.append(frame)
            self.host_layers.append((arr, window or full))
            mirror(frame.c.data, arr, np.uint16)

        def add_coded(i, planar=False):
            planar = planar or PLANAR           # --planar: every coded source as three planes (the reference's coded_image layout)
            arr = np.roll(base["uyvy"][i % len(base["uyvy"])], 4 * (len(self.coded) + 1), axis=1)
            if planar:          # the same samples as three planes: UYVY = Cb Y0 Cr Y1
                planes = [np.ascontiguousarray(arr[:, 1::2]), np.ascontiguousarray(arr[:, 0::4]), np.ascontiguousarray(arr[:, 2::4])]
                pic = dev.DeviceCoded(planes, w, h, 2, dev._abi.CV_LAYOUT_PLANAR, dev._abi.CV_MATRIX_709)
            else:
                planes = [arr]
                pic = dev.DeviceCoded(planes, w, h, 2, dev._abi.CV_LAYOUT_UYVY, dev._abi.CV_MATRIX_709)
            self.coded.append(pic)
            self.host_coded.append((arr, planar))
            for buffer, plane in zip(pic.buffers, planes):
                mirror(buffer.ptr, plane, np.uint8)

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
            self.job = dev.make_job(a=self.layers[0], b=self.layers[1], mix_b=self.mix_b, out_u8=self.out_u8,
                                    u8_table=self.table, full=full)
            self.d2h_bytes = w * h * 4
        elif name in ("4k_timeline", "4k_timeline_windows"):
            add_coded(0)
            add_coded(1)
            windows = timeline_windows(w, h) if name.endswith("windows") else [None] * 3
            for i in range(3):
                add_layer(i, windows[i])
            self.job = dev.make_job(out=self.out, a=self.coded[0], b=self.coded[1], mix_b=self.mix_b, layers=self.layers)
        elif name == "1080p_4layer_windows":
            for i, window in enumerate(WINDOWS_1080P):
                add_layer(i, window)
            self.job = dev.make_job(out=self.out, layers=self.layers)
        elif name == "4k_4layer_filters":
            for i in range(4):
                add_layer(i)
            self.job = dev.make_job(out=self.out, layers=self.layers)
            self.layer_ops = self.attach_filters(self.job)
        else:
            raise SystemExit("unknown workload %r" % name)

        self.h2d_bytes = sum(p.nbytes for _, p in self.uploads)
        self.result = PinnedArray(dev, (self.d2h_bytes,), np.uint8) if pinned else None

    def attach_filters(self, job):
        """cv_layer.ops of the 4k_4layer_filters workload"""
        A = self.dev._abi
        ops = (A.LayerOps * len(LAYER_FILTERS))()
        for i, spec in enumerate(LAYER_FILTERS):
            if spec is None:
                continue
            ops[i].flags, ops[i].color, ops[i].gain, ops[i].offset = spec[:4]
            if spec[4]:
                ops[i].crop = A.Box2i(*spec[4])
            job.layers[i].ops = C.pointer(ops[i])
        return ops

    def job_into(self, out):
        """The same frame program writing into another output frame (so that every frame of a batch can be checked)."""
        d = self.dev
        if self.name == "4k_4layer_filters":
            job = d.make_job(out=out, layers=self.layers)
            job._ops = self.attach_filters(job)
            return job
        if self.name == "1080p_cross_u8":
            return d.make_job(a=self.layers[0], b=self.layers[1], mix_b=self.mix_b, out_u8=out, u8_table=self.table, full=self.full)
        return d.make_job(out=out, a=self.coded[0] if self.coded else None, b=self.coded[1] if len(self.coded) > 1 else None,
                          mix_b=self.mix_b if len(self.coded) > 1 else 0.0, layers=self.layers)

    def upload_inputs(self, stream):
        for dst, pinned in self.uploads:
            self.dev.check(self.dev.lib().cv_upload(dst, pinned.ptr, pinned.nbytes, stream), "cv_upload")

    def download_result(self, stream):
        src = self.out_u8.ptr if self.out_u8 is not None else self.out.c.data
        self.dev.check(self.dev.lib().cv_download(self.result.ptr, src, self.d2h_bytes, stream), "cv_download")

    def output_array(self, out=None):
        """The frame the last launch wrote: (h, w, 4) uint16 halves, or uint8 for the 8-bit workload."""
        w, h = WORKLOADS[self.name][:2]
        if self.out_u8 is not None:
            return (out or self.out_u8).download(np.uint8, (h, w, 4))
        return (out or self.out).download()

    def free(self):
        for _, p in self.uploads:
            p.free()
        if self.result is not None:
            self.result.free()
        for f in self.layers + [self.out]:
            f.free()
        for c in self.coded:
            for b in c.buffers:
                b.free()
        if self.out_u8 is not None:
            self.out_u8.free()
        self.uploads, self.layers, self.coded = [], [], []


def make_base(dev, name, seed):
    """A few distinct random source images; input sets use shifted copies (distinct memory, cheap to make)."""
    w, h = WORKLOADS[name][:2]
    rng = np.random.default_rng(seed)
    base = {"layer": [synth_layer(rng, h, w) for _ in range(2)], "uyvy": [synth_uyvy(rng, h, w) for _ in range(2)]}
    if name == "1080p_cross_u8":
        base["table"] = dev.display_table(1.25)
    return base


def count_sets(name):
    return max(INPUT_SETS, int(np.ceil(RING_BYTES / workload_bytes(name))))


# ----------------------------------------------------------------------------- the checker (oracle; never timed on our arm)

def expected_output(inputs):
    """What the oracle (oracle/liboracle.so, pinned bit for bit against the compiled reference) produces for the frame
    program of `inputs`: (array, current_window tuple).  Test infrastructure: used after the timed region only."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oraclelib as O
    orc = O.oracle()
    name, full = inputs.name, inputs.full
    w, h = WORKLOADS[name][:2]
    out = O.HostFrame16(full)

    def coded(arr, planar):
        pic = O.CodedPicture(w, h, 2, O.LAYOUT_PLANAR if planar else O.LAYOUT_UYVY, O.MATRIX_709)
        pic.set_samples(arr[:, 1::2], arr[:, 0::4], arr[:, 2::4])       # UYVY = Cb Y0 Cr Y1
        return pic

    layers = [O.HostFrame16(full, win, arr) for arr, win in inputs.host_layers]
    pics = [coded(arr, planar) for arr, planar in inputs.host_coded]
    arr = (O.Frame16 * max(1, len(layers)))(*[l.c for l in layers])

    if name == "1080p_cross_u8":
        orc.orc_cross_f16(C.byref(out.c), C.byref(layers[0].c), C.byref(layers[1].c), inputs.mix_b)
        srgb, ramp = np.zeros(65536, np.uint16), np.zeros(65536, np.uint8)
        orc.orc_build_lut(O.LUT_LINEAR_TO_SRGB, O.ptr(srgb))
        orc.orc_build_intent_ramp(1.25, O.ptr(ramp))
        u8 = np.zeros((h, w, 4), np.uint8)
        orc.orc_encode_u8(C.byref(out.c), O.ptr(srgb), O.ptr(ramp), O.ptr(u8))
        return u8, out.current.tuple()
    if name == "4k_4layer_filters":
        class OrcOps(C.Structure):
            _fields_ = [("flags", C.c_int32), ("color", C.c_int32), ("gain", C.c_float), ("offset", C.c_float), ("crop", O.Box),
                        ("even_rows", O.Frame16)]
        ops = (OrcOps * len(layers))()
        for i, spec in enumerate(LAYER_FILTERS):
            if spec is not None:
                ops[i].flags, ops[i].color, ops[i].gain, ops[i].offset = spec[:4]
                if spec[4]:
                    ops[i].crop = O.Box(*spec[4])
        orc.orc_composite_ops_f16.argtypes = [C.POINTER(O.Frame16), C.POINTER(O.Frame16), C.c_void_p, C.c_int]
        orc.orc_composite_ops_f16(C.byref(out.c), arr, C.addressof(ops), len(layers))
    elif not pics:
        orc.orc_composite_f16(C.byref(out.c), arr, len(layers))
    elif not layers and len(pics) == 1:
        orc.orc_reconstruct(C.byref(out.c), C.byref(pics[0].c))
    else:
        orc.orc_timeline_frame_f16(C.byref(out.c), C.byref(pics[0].c), C.byref(pics[1].c) if len(pics) > 1 else None,
                                   inputs.mix_b if len(pics) > 1 else 0.0, arr, len(layers))
    return out.array, out.current.tuple()


def max_ulp(got, want):
    """Largest distance between two arrays of raw half bits (or bytes), in units in the last place."""
    if got.dtype == np.uint8:
        return int(np.abs(got.astype(np.int32) - want.astype(np.int32)).max()) if got.size else 0
    if np.array_equal(got, want):
        return 0
    def ordered(a):
        a = a.astype(np.int32)
        return np.where(a & 0x8000, -(a & 0x7FFF), a)
    return int(np.abs(ordered(got) - ordered(want)).max())


def check_sets(sets, which):
    """Compares the outputs the last launches wrote for sets[i], i in `which`, with the oracle."""
    worst, checked = 0, 0
    for i in which:
        want, window = expected_output(sets[i])
        got = sets[i].output_array()
        x0, y0, x1, y1 = window
        worst = max(worst, max_ulp(got[y0:y1 + 1, x0:x1 + 1], want[y0:y1 + 1, x0:x1 + 1]))
        checked += 1
    return {"parity_checked": True, "max_ulp": worst, "frames_checked": checked}


# ----------------------------------------------------------------------------- measurement

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


def dist_reduce(dist, value, op):
    if dist is None:
        return value
    import torch
    t = torch.tensor([value], dtype=torch.float64, device="cuda" if dist.get_backend() == "nccl" else "cpu")
    dist.all_reduce(t, op=getattr(dist.ReduceOp, op))
    return float(t.item())


def dist_max(dist, value):
    return dist_reduce(dist, value, "MAX")


def dist_sum(dist, value):
    return dist_reduce(dist, value, "SUM")


class Bench:
    """One device, one launching stream, the barrier of the job."""

    def __init__(self, args):
        from canvas_b200 import device as dev
        self.dev, self.args = dev, args
        self.dist, self.rank, self.world = dist_setup(args.gpus)
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        self.numa = pin_to_gpu_node(self.local) if not args.no_numa else None
        dev.init(self.local)
        self.lib = dev.lib()
        self.stream = C.c_void_p()
        dev.check(self.lib.cv_stream_create(C.byref(self.stream)), "cv_stream_create")
        self.peak, self.peak_source = load_peaks()
        self.traffic = load_traffic()

    def host_signal(self, key):
        if self.dist is not None:
            self.dist.distributed_c10d._get_default_store().set(key, "1")

    def host_wait(self, key):
        """blocks on the rendezvous (TCP) store: no GPU work while waiting"""
        if self.dist is not None:
            import datetime
            self.dist.distributed_c10d._get_default_store().wait([key], datetime.timedelta(seconds=600))

    def barrier(self):
        self.dev.check(self.lib.cv_device_sync(), "cv_device_sync")
        if self.dist is not None:
            self.dist.barrier()
        self.dev.check(self.lib.cv_device_sync(), "cv_device_sync")

    def build(self, name, pinned):
        base = make_base(self.dev, name, 1234 + self.rank)
        n_sets = count_sets(name)
        return [InputSet(self.dev, name, base, pinned=pinned) for _ in range(n_sets)]

    def measure(self, name, sets, steps, warmup, frames_per_step=None, sample_clocks=True):
        """Device-resident throughput of `steps` steps; returns the pieces of a bench line."""
        dev, lib = self.dev, self.lib
        n_sets = len(sets)
        per_step = frames_per_step or WORKLOADS[name][3]
        per_step = max(n_sets, (per_step // n_sets) * n_sets)           # every step submits the same job array
        jobs = (dev.FrameJob * per_step)(*[sets[f % n_sets].job for f in range(per_step)])

        def step():
            if self.args.no_batch:
                for f in range(per_step):
                    dev.check(lib.cv_render_frames(C.byref(jobs, f * C.sizeof(dev.FrameJob)), 1, self.stream), "cv_render_frames")
            else:
                dev.check(lib.cv_render_frames(jobs, per_step, self.stream), "cv_render_frames")

        for _ in range(warmup):
            step()
        self.barrier()

        sampler = ClockSampler(self.local) if sample_clocks else None
        if sampler:
            sampler.start()
        launches_before = lib.cv_kernel_launches()
        timer = dev.Timer(self.stream)
        timer.begin()
        for _ in range(steps):
            step()
        timer.end()
        ms = timer.elapsed_ms()
        self.barrier()
        launches = lib.cv_kernel_launches() - launches_before
        clocks = sampler.finish() if sampler else None

        ms_max = dist_max(self.dist, ms)
        frames_rank = steps * per_step
        frames_total = dist_sum(self.dist, float(frames_rank))
        frame_bytes = workload_bytes(name)
        kernel_ms = ms / max(1, launches)
        frames_per_launch = frames_rank / max(1, launches)
        achieved = frame_bytes * frames_per_launch / (kernel_ms * 1e-3) / 1e9
        traffic = self.traffic.get(name)
        return {
            "value": frames_total / (ms_max * 1e-3), "ms_per_step": ms_max / steps, "frames_per_step": per_step,
            "launches": int(launches), "clocks": clocks,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": self.peak, "unit": "GB/s", "frac": achieved / self.peak,
                         "traffic": (traffic["dram_bytes_per_frame"] * frames_per_launch if traffic else None),
                         "traffic_source": (traffic["source"] if traffic else None),
                         "peak_source": self.peak_source, "algorithmic_bytes_per_frame": frame_bytes,
                         "kernel_ms": kernel_ms, "frames_per_launch": frames_per_launch},
        }

    def e2e(self, sets, frames):
        """Pinned host inputs -> HBM -> kernel -> pinned host result, every frame, several streams deep."""
        dev, lib = self.dev, self.lib
        n_sets = len(sets)
        streams = []
        for _ in range(min(n_sets, 4)):
            s = C.c_void_p()
            dev.check(lib.cv_stream_create(C.byref(s)), "cv_stream_create")
            streams.append(s)

        def run(count):
            for f in range(count):
                i = f % n_sets
                st = streams[i % len(streams)]
                sets[i].upload_inputs(st)
                dev.render_frames([sets[i].job], st)
                sets[i].download_result(st)
            for s in streams:
                dev.check(lib.cv_stream_sync(s), "cv_stream_sync")

        run(n_sets)
        self.barrier()
        t0 = time.perf_counter()
        run(frames)
        seconds = time.perf_counter() - t0
        self.barrier()
        for s in streams:
            lib.cv_stream_destroy(s)
        seconds = dist_max(self.dist, seconds)
        return dist_sum(self.dist, float(frames)) / seconds


def plugin_e2e(B, name, sets, check=True):
    """The same workload through the REFERENCE's interface with host buffers (include/framework.h): the layers of each
    input set are host f16 frames wrapped as video sources, placed on a workspace with workspace_add_item, and the
    composite is pulled with video_get_frame_f16(&workspace_source, frame, &host_frame) from several threads -- the
    calls canvas_b200/csrc/client/host_client.c makes, and the calls the reference arm makes against the reference's own C
    (oracle/ref_harness.c).  Frames live in ordinary (pageable) host memory, as the reference's PyMem_Malloc frames do."""
    from canvas_b200 import _abi as A
    w, h = WORKLOADS[name][:2]
    full = A.Box2i(0, 0, w - 1, h - 1)
    client = C.CDLL(os.path.join(ROOT, "canvas_b200", "libcanvas_client.so"))
    client.client_timeline_create.restype = C.c_void_p
    client.client_timeline_create.argtypes = [C.POINTER(A.FrameF16), C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int64),
                                              C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.c_int]
    client.client_time_timeline.restype = C.c_int64
    client.client_time_timeline.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.POINTER(A.FrameF16)]
    client.client_timeline_get_frame_f16.argtypes = [C.c_void_p, C.c_int, C.POINTER(A.FrameF16)]
    client.client_timeline_free.argtypes = [C.c_void_p]

    n_layers, n_sets = len(sets[0].host_layers), len(sets)
    keep, frames = [], (A.FrameF16 * (n_layers * n_sets))()
    for k in range(n_layers):               # clip k = layer k of every input set: timeline frame f shows set f mod n_sets
        for i, inputs in enumerate(sets):
            arr, window = inputs.host_layers[k]
            arr = np.ascontiguousarray(arr)
            keep.append(arr)
            f = frames[k * n_sets + i]
            f.data, f.full_window, f.current_window = arr.ctypes.data, full, A.Box2i(*window)
    ints, longs = C.c_int * n_layers, C.c_int64 * n_layers
    tl = client.client_timeline_create(frames, ints(*[k * n_sets for k in range(n_layers)]), ints(*([n_sets] * n_layers)),
                                       longs(*([0] * n_layers)), longs(*([1 << 30] * n_layers)), longs(*([0] * n_layers)),
                                       longs(*range(n_layers)), n_layers)

    threads = max(1, min(16, host_cores()))
    outs_np = [np.zeros((h, w, 4), np.uint16) for _ in range(threads)]
    outs = (A.FrameF16 * threads)()
    for t in range(threads):
        outs[t].data, outs[t].full_window = outs_np[t].ctypes.data, full

    client.client_time_timeline(tl, 0, threads, threads, outs)            # warm-up: contexts, slabs, pool
    B.barrier()
    count = 8 * threads if name.startswith("4k") else 64 * threads
    ns = client.client_time_timeline(tl, 0, count, threads, outs)
    B.barrier()
    seconds = dist_max(B.dist, ns * 1e-9)
    result = {"value": dist_sum(B.dist, float(count)) / seconds, "unit": "frames/s", "threads": threads, "frames_timed": count,
              "call": "video_get_frame_f16(workspace_as_video_source(...), frame, &host_frame) over %d host-frame sources per frame "
                      "(include/framework.h; pageable host memory in and out)" % n_layers}
    # what the host itself can copy: each output frame costs n_layers + 1 whole-frame memcpys on the CPU (every source
    # writes its frame into the caller's buffer; the result is copied into the caller's frame)
    if B.rank == 0:
        client.client_host_copy_rate.restype = C.c_double
        client.client_host_copy_rate.argtypes = [C.c_int, C.c_size_t, C.c_int]
        rate = client.client_host_copy_rate(threads, w * h * 8, 6)
        result["host_copy_gbs"] = rate / 1e9
        result["host_copy_bound_fps"] = rate / ((n_layers + 1) * w * h * 8) if rate else None
    if check:
        one = A.FrameF16()
        got = np.zeros((h, w, 4), np.uint16)
        one.data, one.full_window = got.ctypes.data, full
        client.client_timeline_get_frame_f16(tl, 0, C.byref(one))
        want, window = expected_output(sets[0])
        x0, y0, x1, y1 = window
        result["parity_checked"] = True
        result["max_ulp"] = max_ulp(got[y0:y1 + 1, x0:x1 + 1], want[y0:y1 + 1, x0:x1 + 1]) if one.current_window.tuple() == window else -1
    client.client_timeline_free(tl)
    return result


def measure_small_kernels(B, seconds=0.15):
    """The encode side, the colour transform, the scaler and the flat half conversions as single kernels: frames/s over
    rotating device buffers larger than L2, achieved GB/s against their algorithmic bytes."""
    dev, lib = B.dev, B.lib
    A = dev._abi
    rng = np.random.default_rng(3)
    results = []

    def timed(name, kernel, note, bytes_per_frame, calls, sets):
        for i in range(3 * sets):
            calls(i % sets)
        B.barrier()
        reps = sets
        while True:
            timer = dev.Timer(B.stream)
            timer.begin()
            for i in range(reps):
                calls(i % sets)
            timer.end()
            ms = timer.elapsed_ms()
            if ms >= seconds * 1e3 or reps >= 1 << 16:
                break
            reps = int(reps * max(2.0, min(16.0, 1.2 * seconds * 1e3 / max(ms, 1e-3))))
        gbs = bytes_per_frame * reps / (ms * 1e-3) / 1e9
        results.append({"op": name, "kernel": kernel, "note": note, "value": reps / (ms * 1e-3), "unit": "frames/s",
                        "algorithmic_bytes_per_frame": bytes_per_frame, "achieved_gbs": gbs, "roofline_frac": gbs / B.peak})

    w, h = 3840, 2160
    full = (0, 0, w - 1, h - 1)
    base = synth_layer(rng, h, w)
    sets = 5                                                    # 5 x 66 MB f16 frames: 2.6 x L2
    frames = [dev.DeviceFrame16.from_numpy(np.roll(base, 13 * i, axis=1), full, full) for i in range(sets)]

    # f16 -> packed 8-bit 4:2:2 (video_subsample.c:99-187 generalised), table written back in place: 8 + 8 + 2 B/px
    pics = [dev.DeviceCoded([np.zeros((h, 2 * w), np.uint8)], w, h, 2, A.CV_LAYOUT_UYVY, A.CV_MATRIX_709) for _ in range(sets)]
    timed("4k_subsample_422", "k_subsample<1>", "f16 -> UYVY, linear->Rec.709 table applied in place", w * h * 18,
          lambda i: dev.check(lib.cv_subsample_ycbcr8(C.byref(frames[i].c), C.byref(pics[i].c), B.stream), "cv_subsample_ycbcr8"), sets)
    # f16 -> planar 4:2:0 (the MPEG-2 shaders, video_subsample.c:212-290): 8 + 1.5 B/px
    planar = [dev.DeviceCoded([np.zeros((h, w), np.uint8), np.zeros((h // 2, w // 2), np.uint8), np.zeros((h // 2, w // 2), np.uint8)],
                              w, h, 2, A.CV_LAYOUT_PLANAR, A.CV_MATRIX_601) for _ in range(sets)]
    timed("4k_subsample_mpeg2_420", "k_subsample_mpeg2", "f16 -> planar 4:2:0, interlace-aware 6-tap chroma", int(w * h * 9.5),
          lambda i: dev.check(lib.cv_subsample_mpeg2(C.byref(frames[i].c), C.byref(planar[i].c), B.stream), "cv_subsample_mpeg2"), sets)
    # display encode with write-back (widget_gl.c:290-307): 8 + 8 + 4 B/px
    table = dev.display_table(1.25)
    u8 = [dev.DeviceBuffer(w * h * 4) for _ in range(sets)]
    timed("4k_encode_u8_writeback", "k_encode_u8", "f16 -> 8-bit display RGBA, sRGB halves written back", w * h * 20,
          lambda i: dev.check(lib.cv_encode_u8(C.byref(frames[i].c), table, u8[i].ptr, 1, B.stream), "cv_encode_u8"), sets)
    # colour transform in place (color.c:100-131): 8 + 8 B/px
    timed("4k_color_rgb_to_xyz", "k_color", "f16 in place: table, 3x3 in f32, truncating pack", w * h * 16,
          lambda i: dev.check(lib.cv_color_rgb_to_xyz_sdtv(C.byref(frames[i].c), B.stream), "cv_color_rgb_to_xyz_sdtv"), sets)
    # premultiplied ARGB32 (RgbaFrameF16.c:114-149): 8 + 4 B/px
    timed("4k_to_argb32", "k_to_argb32", "f16 -> premultiplied ARGB32 through the gamma-0.45 ramp", w * h * 12,
          lambda i: dev.check(lib.cv_to_argb32(C.byref(frames[i].c), u8[i].ptr, B.stream), "cv_to_argb32"), sets)
    for f in frames + u8:
        f.free()
    for p in pics + planar:
        for b in p.buffers:
            b.free()

    # flat half <-> float (half.c:31-51): 4 + 2 B per value, 2^26 values per call
    count = 1 << 26
    floats = [dev.DeviceBuffer(count * 4) for _ in range(3)]
    halves = [dev.DeviceBuffer(count * 2) for _ in range(3)]
    for b in floats:
        lib.cv_memset(b.ptr, 0x3C, count * 4, B.stream)
    timed("half_from_float_64M", "k_half_from_float", "truncating pack, subnormals kept", count * 6,
          lambda i: dev.check(lib.cv_half_from_float(halves[i].ptr, floats[i].ptr, count, B.stream), "cv_half_from_float"), 3)
    timed("half_to_float_64M", "k_half_to_float", "exact widening", count * 6,
          lambda i: dev.check(lib.cv_half_to_float(floats[i].ptr, halves[i].ptr, count, B.stream), "cv_half_to_float"), 3)
    for b in floats + halves:
        b.free()

    # the bilinear scaler (video_scale.c:231-319): 1080p f32 -> 4K f32, both axes (factor 2).  Enlargement, because for
    # factors below one the reference sizes its intermediate frame with `* factor` (:253-256) and only that share of the
    # target is computed; algorithmic = source + target, the intermediate frame stays on chip
    src_full, dst_full = (0, 0, w // 2 - 1, h // 2 - 1), (0, 0, w - 1, h - 1)
    f32 = np.ascontiguousarray(base.view(np.float16).astype(np.float32)[:h // 2, :w // 2])
    sources = [dev.DeviceFrame32.from_numpy(np.roll(f32, 7 * i, axis=1), src_full, src_full) for i in range(3)]
    targets = [dev.DeviceFrame32(dst_full) for _ in range(3)]
    zero, twice = A.V2f(0.0, 0.0), A.V2f(2.0, 2.0)
    timed("1080p_to_4k_scale_f32", "k_scale_fused<0>", "triangle-filter resampling, both passes in one launch (intermediate frame in shared memory, tap tables cached)",
          (w * h + (w // 2) * (h // 2)) * 16,
          lambda i: dev.check(lib.cv_scale_bilinear_f32(C.byref(targets[i].c), zero, C.byref(sources[i].c), zero, twice, B.stream), "cv_scale_bilinear_f32"), 3)
    for f in sources + targets:
        f.free()
    return results


def scheduler_bench(B, n_gpus, frames=1000, chunk=16, ring=32, clip=4):
    """BASELINE.json configs[4] through the C scheduler: a 3840x2160, `frames`-frame timeline -- two 4:2:2 clips crossfading
    under a linear ramp + three f16 overlay clips with PARTIAL windows -- rendered by ONE cv_timeline_render() call that
    hands chunks of frames to the GPUs of this box (one worker thread, a compute and a copy stream per GPU, no collective).
      resident   every clip resident on every GPU, f16 output left in HBM (the device-side capability of the scheduler);
      pulled_u8  end to end: the 4:2:2 pictures are PULLED from pinned host memory through coded_image_source.getFrame
                 (4 B/px host->device per frame), overlays resident (static graphics), the 8-bit display encode of every
                 frame copied to pinned host memory from the sink (4 B/px device->host);
      pulled_422 the same input, the result subsampled to 8-bit 4:2:2 on the device (cv_subsample_ycbcr8 from the sink)
                 and copied back: 2 B/px device->host, what an encoder takes."""
    from canvas_b200 import _abi as A
    dev, lib = B.dev, B.lib
    w, h = 3840, 2160
    full = (0, 0, w - 1, h - 1)
    windows = timeline_windows(w, h)
    rng = np.random.default_rng(7)
    base_uyvy = rng.integers(16, 236, size=(h, w * 2), dtype=np.uint8)
    base_layer = synth_layer(rng, h, w)

    client = C.CDLL(os.path.join(ROOT, "canvas_b200", "libcanvas_client.so"))
    client.client_coded_clip_create.restype = C.c_void_p
    client.client_coded_clip_create.argtypes = [C.POINTER(C.c_void_p), C.c_int, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    client.client_coded_clip_source.restype = C.POINTER(A.CodedImageSource)
    client.client_coded_clip_source.argtypes = [C.c_void_p]

    keep, clips, pulls = [], {}, {}
    for name in ("a", "b", "o0", "o1", "o2"):
        arr = (A.Clip * n_gpus)()
        for g in range(n_gpus):
            dev.init(g)
            srcs = (A.Source * clip)()
            for k in range(clip):
                if name in ("a", "b"):
                    d = dev.DeviceCoded([np.roll(base_uyvy, 4 * (k + 1) + (64 if name == "b" else 0), axis=1)], w, h, 2,
                                        A.CV_LAYOUT_UYVY, A.CV_MATRIX_709)
                    srcs[k].kind, srcs[k].coded = A.CV_SRC_CODED, d.c
                else:
                    d = dev.DeviceFrame16.from_numpy(np.roll(base_layer, 17 * k + 1, axis=1), full, windows[int(name[1])])
                    srcs[k].kind, srcs[k].f16 = A.CV_SRC_F16, d.c
                keep.append(d)
            arr[g].frames, arr[g].count = srcs, clip
            keep.append(srcs)
        clips[name] = arr
    dev.init(B.local)
    for name in ("a", "b"):                 # the same pictures in pinned host memory, behind coded_image_source.getFrame
        planes = (C.c_void_p * clip)()
        for k in range(clip):
            pic = np.roll(base_uyvy, 4 * (k + 1) + (64 if name == "b" else 0), axis=1)
            pinned = PinnedArray(dev, pic.shape, np.uint8)
            pinned.array[...] = pic
            planes[k] = pinned.ptr.value
            keep.append(pinned)
        stride, lines = (C.c_int * 1)(2 * w), (C.c_int * 1)(h)
        p = A.Pull()
        p.coded = client.client_coded_clip_source(client.client_coded_clip_create(planes, clip, 1, stride, lines))
        p.format.layout, p.format.width, p.format.height, p.format.sub_x, p.format.matrix = A.CV_LAYOUT_UYVY, w, h, 2, A.CV_MATRIX_709
        pulls[name] = p
        keep += [planes, stride, lines]

    def make_items(pulled):
        items = (A.TimelineItem * 4)()
        items[0].x, items[0].length, items[0].z, items[0].offset = 0, frames, 0, 0
        if pulled:
            items[0].pull, items[0].pull_b = C.pointer(pulls["a"]), C.pointer(pulls["b"])
        else:
            items[0].clips, items[0].clips_b = clips["a"], clips["b"]
        items[0].mix_first, items[0].mix_last, items[0].ramp_start, items[0].ramp_length = 0.0, 1.0, -1, frames + 1
        for i in range(3):
            it = items[1 + i]
            it.x, it.length, it.z, it.offset, it.clips = 0, frames, 1 + i, i, clips["o%d" % i]
        return items

    targets, pinned_out, u8_rings = (A.TimelineTarget * n_gpus)(), [], []
    for g in range(n_gpus):
        dev.init(g)
        outs = (A.FrameF16 * ring)()
        bufs = (C.c_void_p * ring)()
        for r in range(ring):
            f = dev.DeviceFrame16(full)
            b = dev.DeviceBuffer(w * h * 4)
            keep += [f, b]
            outs[r], bufs[r] = f.c, b.ptr.value
        targets[g].out, targets[g].ring = outs, ring
        u8_rings.append(bufs)
        table = dev.display_table(1.25)
        targets[g].u8_table = table
        host = PinnedArray(dev, (ring, h, w, 4), np.uint8)
        pinned_out.append(host)
        keep += [outs, bufs, table, host]
    dev.init(B.local)

    counters = [0] * n_gpus

    @A.TIMELINE_SINK
    def sink(user, slot, frame, job, stream):
        k = counters[slot] % ring
        counters[slot] += 1
        return lib.cv_download(C.c_void_p(pinned_out[slot].ptr.value + k * w * h * 4), job.contents.out_u8, w * h * 4, stream)

    # 4:2:2 8-bit out (what an encoder takes, X264VideoEncoder.c:295-312): cv_subsample_ycbcr8 from the sink, 2 B/px back
    coded_out = []
    for g in range(n_gpus):
        dev.init(g)
        coded_out.append([dev.DeviceCoded([np.zeros((h, 2 * w), np.uint8)], w, h, 2, A.CV_LAYOUT_UYVY, A.CV_MATRIX_709) for _ in range(ring)])
    dev.init(B.local)
    keep.append(coded_out)

    @A.TIMELINE_SINK
    def sink_422(user, slot, frame, job, stream):
        k = counters[slot] % ring
        counters[slot] += 1
        pic = coded_out[slot][k]
        rc = lib.cv_subsample_ycbcr8(job.contents.out, C.byref(pic.c), stream)
        return rc or lib.cv_download(C.c_void_p(pinned_out[slot].ptr.value + k * w * h * 4), pic.buffers[0].ptr, w * h * 2, stream)

    coverage = sum((b[2] - b[0] + 1) * (b[3] - b[1] + 1) for b in windows)
    result = {"n_gpus": n_gpus, "frames": frames, "chunk": chunk, "ring": ring, "workload": "4k_timeline_windows (BASELINE.json configs[4])"}
    modes = ("resident", "pulled_u8", "pulled_422") if not os.environ.get("CV_BENCH_SCHEDULER_MODES") else tuple(os.environ["CV_BENCH_SCHEDULER_MODES"].split(","))
    for mode in modes:                      # also: "pulled" (4:2:2 in, f16 left in HBM), "resident_u8" (8-bit out only)
        items = make_items(mode.startswith("pulled"))
        run = A.TimelineRun()
        run.items, run.item_count, run.n_gpus = items, 4, n_gpus
        run.first, run.last, run.chunk, run.targets = 0, frames - 1, chunk, targets
        for g in range(n_gpus):             # the 8-bit targets exist only in the end-to-end modes
            targets[g].out_u8 = C.cast(u8_rings[g], C.POINTER(C.c_void_p)) if mode.endswith("_u8") else None
        if mode.endswith("_u8"):
            run.sink = sink
        elif mode.endswith("_422"):
            run.sink = sink_422
        stats = (A.TimelineStats * n_gpus)()
        warm = A.TimelineRun()
        C.memmove(C.byref(warm), C.byref(run), C.sizeof(run))
        warm.last = min(frames, 4 * chunk * n_gpus) - 1
        dev.check(lib.cv_timeline_render(C.byref(warm), stats), "cv_timeline_render (warm-up)")
        before = lib.cv_kernel_launches()
        t0 = time.perf_counter()
        dev.check(lib.cv_timeline_render(C.byref(run), stats), "cv_timeline_render")
        wall = time.perf_counter() - t0
        device_ms = max(s.device_ms for s in stats)
        bytes_per_frame = 4 * w * h + 8 * coverage + 8 * w * h + (4 * w * h if mode.endswith("_u8") else 0) + (18 * w * h if mode.endswith("_422") else 0)
        result[mode] = {"value": frames / wall, "unit": "frames/s", "wall_ms": wall * 1e3, "slowest_gpu_device_ms": device_ms,
                        "device_fps": frames / (device_ms * 1e-3), "gpu_launches": int(lib.cv_kernel_launches() - before),
                        "frames_per_gpu": [int(s.frames) for s in stats], "algorithmic_gb_per_s": bytes_per_frame * frames / wall / 1e9,
                        "h2d_bytes_per_frame": (sum(int(s.h2d_bytes) for s in stats) / frames), "d2h_bytes_per_frame": (4 * w * h if mode.endswith("_u8") else (2 * w * h if mode.endswith("_422") else 0))}
    for k in keep:
        if hasattr(k, "free"):
            try:
                k.free()
            except Exception:
                pass
    return result


def run_ours(args):
    B = Bench(args)
    dev, name = B.dev, args.workload
    w, h, _, _, description = WORKLOADS[name]

    sets = B.build(name, pinned=True)
    m = B.measure(name, sets, args.steps, args.warmup)
    # the outputs the timed launches wrote, against the oracle: the first and the last input set of the ring
    parity = check_sets(sets, sorted({0, len(sets) - 1})) if not args.no_check else {"parity_checked": False}

    e2e_frames = max(len(sets), min(args.steps * m["frames_per_step"], 96 if name.startswith("4k") else 960))
    e2e_value = B.e2e(sets, e2e_frames)
    h2d, d2h = sets[0].h2d_bytes, sets[0].d2h_bytes

    line = {
        "metric": "f16 composite frames/sec (%s)" % name, "value": m["value"], "unit": "frames/s",
        "n_gpus": B.world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": m["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f16 pixels, f32 arithmetic", "data": ("synthetic" if CONTENT == "noise" else "synthetic (%s-like coded sources)" % CONTENT) + (", planar 4:2:2" if PLANAR else ""),
        "config": {"workload": name, "description": description, "width": w, "height": h,
                   "frames_per_step": m["frames_per_step"], "input_sets": len(sets),
                   "l2": "inputs larger than L2 (%.0f MB per frame, %d rotating sets)" % (workload_bytes(name) / 1e6, len(sets)),
                   "sharding": "frames by index, one process per GPU, no collective", "numa": B.numa},
        "roofline": m["roofline"],
        "e2e": {"value": e2e_value, "unit": "frames/s", "h2d_bytes_per_step": h2d * m["frames_per_step"],
                "d2h_bytes_per_step": d2h * m["frames_per_step"], "frames_timed": e2e_frames,
                "h2d_bytes_per_frame": h2d, "d2h_bytes_per_frame": d2h},
        "gpu_launches": m["launches"], "clocks": m["clocks"],
    }
    line.update(parity)
    if sets[0].host_layers and not sets[0].host_coded and sets[0].out_u8 is None and not args.no_plugin:
        line["e2e_plugin"] = plugin_e2e(B, name, sets, check=not args.no_check)
    for s in sets:
        s.free()
    sets = None

    # ---- the other BASELINE.json configurations, about 0.3 s of device time each ------------------------------------
    if not args.no_others:
        others = []
        for other in OTHER_WORKLOADS:
            if other == name:
                continue
            osets = B.build(other, pinned=False)
            om = B.measure(other, osets, steps=6, warmup=3, sample_clocks=False)
            entry = {"workload": other, "value": om["value"], "unit": "frames/s", "ms_per_step": om["ms_per_step"],
                     "frames_per_step": om["frames_per_step"], "steps": 6, "gpu_launches": om["launches"],
                     "roofline_frac": om["roofline"]["frac"], "achieved_gbs": om["roofline"]["achieved"],
                     "algorithmic_bytes_per_frame": om["roofline"]["algorithmic_bytes_per_frame"],
                     "traffic": om["roofline"]["traffic"]}
            if not args.no_check:
                entry.update(check_sets(osets, [0]))
            others.append(entry)
            for s in osets:
                s.free()
        line["other_workloads"] = others
    if not args.no_others or args.only_kernels:
        line["other_kernels"] = measure_small_kernels(B)

    # ---- the frame-parallel scheduler itself: ONE cv_timeline_render() call over all the GPUs of the job -------------
    if not args.no_scheduler:
        B.barrier()
        # The other ranks' GPUs are driven by rank 0's worker threads now.  They must idle on the HOST: an NCCL barrier
        # would park a spinning kernel on every waiting GPU, so they wait on a key of the rendezvous store instead.
        if B.rank == 0:
            n_sched = max(B.world, args.gpus if B.dist is None else B.world)
            n_sched = min(n_sched, B.lib.cv_device_count())
            try:
                line["scheduler"] = scheduler_bench(B, n_sched, frames=1000)
            finally:
                B.host_signal("scheduler_done")
        else:
            B.host_wait("scheduler_done")
        B.barrier()

    if B.rank == 0 and B.world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline(name, seconds=20.0)

    if B.rank == 0:
        print(json.dumps(line))
    if B.dist is not None:
        B.dist.barrier()
        B.dist.destroy_process_group()


# ----------------------------------------------------------------------------- CPU baseline / reference arm

REF_FAST_FLAGS = "-O3 -march=x86-64-v3 -mtune=generic -fno-math-errno -fno-signed-zeros (the reference's SConstruct:77,83 with x86-64-v3 for native, so that the build travels)"


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def cpu_threads(name):
    """All usable host cores, bounded by memory (each in-flight 4K frame needs ~0.5 GB of f32 temporaries)."""
    w, h = WORKLOADS[name][:2]
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
    w, h = WORKLOADS[name][:2]
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
    n_layers = {"4k_8layer_composite": 8, "4k_composite_convert": 7, "1080p_4layer_windows": 4, "4k_4layer_filters": 4}.get(name, 3)
    layers = [O.HostFrame16(full, full, synth_layer(rng, h, w)) for _ in range(n_layers)]
    windows = WINDOWS_1080P if name == "1080p_4layer_windows" else ([full] + timeline_windows(w, h) if name == "4k_timeline_windows" else [])
    windowed = [O.HostFrame16(full, win, l.array) for l, win in zip(layers, windows)]
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
        elif name == "4k_timeline_windows":
            arr = (O.Frame16 * 3)(*[l.c for l in windowed[1:4]])
            orc.orc_timeline_frame_f16(C.byref(out.c), C.byref(pics[0].c), C.byref(pics[1].c), 0.375, arr, 3)
        elif name == "1080p_4layer_windows":
            arr = (O.Frame16 * 4)(*[l.c for l in windowed])
            orc.orc_composite_f16(C.byref(out.c), arr, 4)
        elif name == "4k_4layer_filters":
            arr = (O.Frame16 * 4)(*[l.c for l in layers])
            orc.orc_composite_f16(C.byref(out.c), arr, 4)          # the stack without its filters (a lower bound of the CPU cost)

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
    """Bounded sample of the same workload on the host cores (about `seconds` of CPU work), all cores and -- what the
    reference itself does: one frame on one thread (src/process/VideoPullQueue.c:110, main.c:120-167) -- one thread."""
    threads = cpu_threads(name)
    t0, kind = cpu_run(name, threads, threads)                  # calibration: one frame per thread
    frames = int(max(threads, min(64 * threads, threads * round(seconds * 0.75 / max(t0, 1e-3)))))
    t, kind = cpu_run(name, frames, threads)
    single = int(max(2, min(64, round(seconds * 0.25 / max(t0, 1e-3)))))       # a frame alone takes at most t0
    t1, _ = cpu_run(name, single, 1)
    return {"value": frames / t, "unit": "frames/s", "cores": threads, "kind": kind,
            "sample": "%d frames of %s, %d frames in flight (one per thread), %.1f s" % (frames, name, threads, t),
            "one_thread": {"value": single / t1, "unit": "frames/s", "cores": 1,
                           "sample": "%d frames on one thread, %.1f s" % (single, t1)},
            "flags": REF_FAST_FLAGS if kind == "reference" else "oracle/Makefile: -O2 -ffp-contract=off"}


def run_reference(args):
    """--impl reference: the reference's own CPU path (or the oracle port where it has none) on the host cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    name = args.workload
    w, h, _, _, description = WORKLOADS[name]
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
                         "sample": "%d steps x %d frames of %s, %d frames in flight" % (steps, per_step, name, threads),
                         "flags": REF_FAST_FLAGS if kind == "reference" else "oracle/Makefile: -O2 -ffp-contract=off"},
        "e2e": {"value": value, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="4k_8layer_composite", choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-others", action="store_true", help="skip the other_workloads array")
    ap.add_argument("--only-kernels", action="store_true", help="with --no-others: still measure the other_kernels array")
    ap.add_argument("--no-check", action="store_true", help="skip the oracle comparison after the timed region")
    ap.add_argument("--no-numa", action="store_true", help="do not pin the process to the GPU's NUMA node")
    ap.add_argument("--no-plugin", action="store_true", help="skip e2e_plugin (the framework.h pull with host frames)")
    ap.add_argument("--no-scheduler", action="store_true", help="skip the cv_timeline_render() measurement over all GPUs")
    ap.add_argument("--planar", action="store_true", help="4:2:2 sources as three planes instead of packed UYVY")
    ap.add_argument("--content", default="noise", choices=["noise", "video"],
                    help="coded sources: independent random bytes (default) or smooth picture-like content")
    ap.add_argument("--no-batch", action="store_true", help="one cv_render_frames() call per frame")
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
