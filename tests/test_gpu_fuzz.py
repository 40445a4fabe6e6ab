"""Randomised parity of whole frame programs (cv_render_frames) against the oracle: random sizes and origins (even and odd:
the row-segment kernels and the per-item fallback), 4:2:2 bottoms in every layout with random picture offsets, crossfades,
0-6 overlays with random windows, single launches and same-geometry batches.  Bit for bit.
CV_FUZZ_CASES scales the number of cases (default 40, seeds fixed)."""
import ctypes as C
import os

import numpy as np
import pytest

import oraclelib as O
from test_gpu_parity import assert_frames_equal, device_coded, poisoned, to_device

pytestmark = pytest.mark.gpu

CASES = int(os.environ.get("CV_FUZZ_CASES", "40"))


@pytest.fixture(scope="module")
def dev():
    from canvas_b200 import device
    device.init(0)
    return device


def random_box(rng, full, allow_outside=True):
    x0, y0, x1, y1 = full
    w, h = x1 - x0 + 1, y1 - y0 + 1
    if rng.random() < 0.35:
        return full
    a = int(rng.integers(x0 - (8 if allow_outside else 0), x1))
    b = int(rng.integers(a, x1 + (9 if allow_outside else 1)))
    c = int(rng.integers(y0 - (3 if allow_outside else 0), y1 + 1))
    d = int(rng.integers(c, y1 + (4 if allow_outside else 1)))
    return (a, c, b, d)


def clip(box, full):
    return (max(box[0], full[0]), max(box[1], full[1]), min(box[2], full[2]), min(box[3], full[3]))


@pytest.mark.parametrize("seed", range(CASES))
def test_random_frame_program(dev, seed):
    rng = np.random.default_rng(1000 + seed)
    orc = O.oracle()
    w = int(rng.choice([66, 128, 130, 200, 257, 384, 640, 700]))
    h = int(rng.integers(3, 28))
    ox = int(rng.choice([0, 0, 0, -6, 10, 3, -1]))       # odd origins leave the row-segment kernels
    oy = int(rng.integers(-3, 4))
    full = (ox, oy, ox + w - 1, oy + h - 1)
    n_jobs = int(rng.choice([1, 1, 2, 5]))
    bottom = rng.choice(["none", "single", "cross"], p=[0.3, 0.35, 0.35])
    n_layers = int(rng.integers(0 if bottom != "none" else 1, 7))
    mixes = [float(rng.choice([0.0, 1.0, 0.25, rng.random()])) for _ in range(n_jobs)]

    # geometry shared by the jobs of the batch
    def coded_shape():
        layout = int(rng.choice([O.LAYOUT_PLANAR, O.LAYOUT_UYVY, O.LAYOUT_YUY2]))
        matrix = int(rng.choice([O.MATRIX_709, O.MATRIX_601]))
        if rng.random() < 0.6:
            pw, ph, pic = w, h, (ox, oy)
        else:
            pw = int(rng.integers(8, w + 20)) & ~1
            ph = int(rng.integers(2, h + 4))
            pic = (ox + 2 * int(rng.integers(-4, 8)) + (1 if rng.random() < 0.15 else 0), oy + int(rng.integers(-2, 3)))
        return pw, ph, layout, matrix, pic

    shape_a, shape_b = coded_shape(), coded_shape()
    layer_fulls, layer_windows = [], []
    for _ in range(n_layers):
        lf = full if rng.random() < 0.7 else (full[0] - 2 * int(rng.integers(0, 5)), full[1] - int(rng.integers(0, 3)),
                                              full[2] + 2 * int(rng.integers(0, 5)), full[3] + int(rng.integers(0, 3)))
        box = random_box(rng, full)
        layer_fulls.append(lf)
        layer_windows.append(clip(box, lf))

    jobs, outs, wants, keep = [], [], [], []
    for j in range(n_jobs):
        pa = pb = None
        if bottom != "none":
            pw, ph, layout, matrix, pic = shape_a
            pa = O.synth_coded(pw, ph, 10 * seed + j, 2, layout, matrix, pic, 0, True)
        if bottom == "cross":
            pw, ph, layout, matrix, pic = shape_b
            pb = O.synth_coded(pw, ph, 10 * seed + j + 5, 2, layout, matrix, pic, 0, True)
        layers = []
        for k in range(n_layers):
            win = layer_windows[k]
            if win[2] < win[0] or win[3] < win[1]:
                win = (0, 0, -1, -1)
            layers.append(O.synth_frame(layer_fulls[k], 100 * seed + 10 * j + k, win))
        want = O.HostFrame16(full)
        arr = (O.Frame16 * max(1, n_layers))(*[l.c for l in layers]) if n_layers else None
        orc.orc_timeline_frame_f16(C.byref(want.c), C.byref(pa.c) if pa else None, C.byref(pb.c) if pb else None,
                                   mixes[j], arr, n_layers)
        out = poisoned(dev, full)
        da = device_coded(dev, pa) if pa else None
        db = device_coded(dev, pb) if pb else None
        dl = [to_device(dev, l) for l in layers]
        jobs.append(dev.make_job(out=out, a=da, b=db, mix_b=mixes[j] if pb else 0.0, layers=dl))
        outs.append(out); wants.append(want); keep.append((pa, pb, layers, da, db, dl))

    dev.render_frames(jobs)
    dev.sync()
    what = "seed %d: %dx%d at (%d,%d), bottom %s, %d layers, %d job(s)" % (seed, w, h, ox, oy, bottom, n_layers, n_jobs)
    for j in range(n_jobs):
        assert_frames_equal(outs[j].download(), outs[j].current.tuple(), wants[j], "%s, job %d" % (what, j))


@pytest.mark.parametrize("seed", range(max(8, CASES // 4)))
def test_random_full_window_stack(dev, seed):
    """the TMA-ring composite (every layer covers the frame): random sizes up to a few thousand pixels wide, so that the
    warps' row cursors wrap in every way; optional 8-bit output sends the same stack through the row-segment kernel"""
    rng = np.random.default_rng(5000 + seed)
    orc = O.oracle()
    w = 2 * int(rng.integers(17, 2100)) if rng.random() < 0.5 else 2 * int(rng.integers(17, 200))
    h = int(rng.integers(1, 40)) if w > 600 else int(rng.integers(1, 300))
    ox, oy = 2 * int(rng.integers(-3, 4)), int(rng.integers(-3, 4))
    full = (ox, oy, ox + w - 1, oy + h - 1)
    n = int(rng.integers(1, 9))
    layers = [O.synth_frame(full, 7000 + 16 * seed + k) for k in range(n)]
    want = O.HostFrame16(full)
    arr = (O.Frame16 * n)(*[l.c for l in layers])
    orc.orc_composite_f16(C.byref(want.c), arr, n)
    out = poisoned(dev, full)
    dl = [to_device(dev, l) for l in layers]
    u8 = table = None
    if rng.random() < 0.3:
        u8, table = dev.DeviceBuffer(w * h * 4), dev.display_table(1.25)
    dev.render_frames([dev.make_job(out=out, layers=dl, out_u8=u8, u8_table=table)])
    dev.sync()
    assert_frames_equal(out.download(), out.current.tuple(), want, "seed %d: %d layers %dx%d" % (seed, n, w, h))
    if u8 is not None:
        srgb, ramp = np.zeros(65536, np.uint16), np.zeros(65536, np.uint8)
        orc.orc_build_lut(O.LUT_LINEAR_TO_SRGB, O.ptr(srgb))
        orc.orc_build_intent_ramp(1.25, O.ptr(ramp))
        target = np.zeros((h, w, 4), np.uint8)
        orc.orc_encode_u8(C.byref(want.c), O.ptr(srgb), O.ptr(ramp), O.ptr(target))
        assert np.array_equal(u8.download(np.uint8, (h, w, 4)), target), "seed %d: 8-bit output" % seed
