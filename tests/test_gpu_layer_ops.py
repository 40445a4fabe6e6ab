"""Per-pixel filters fused into workspace layers (cv_layer_ops): the field weave of Pulldown23RemovalFilter
(/root/reference/src/process/Pulldown23RemovalFilter.c:82-104), the colour transforms (src/cprocess/color.c:100-160),
gain/offset (video_filter.c:27-39) and crop (video_filter.c:79-95) evaluated in registers on a layer's pixels inside the
composite launch.  Three ways to the same bits:
  fused     one cv_render_frames() launch with cv_layer.ops set,
  separate  cv_weave_fields_f16 / cv_color_* / cv_gain_offset_f32 / cv_crop as their own passes, then the f32 stack,
  oracle    oracle/oracle.c:orc_composite_ops_f16 (each filter as its own pass over the oracle's pinned functions).
"""
import ctypes as C

import numpy as np
import pytest

import oraclelib as O

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gpu():
    from canvas_b200 import _abi as A
    from canvas_b200 import device as dev
    dev.init(0)
    O.oracle().orc_composite_ops_f16.argtypes = [C.POINTER(O.Frame16), C.POINTER(O.Frame16), C.c_void_p, C.c_int]
    O.oracle().orc_weave_fields_f16.argtypes = [C.POINTER(O.Frame16), C.POINTER(O.Frame16)]
    return A, dev


class OrcOps(C.Structure):
    _fields_ = [("flags", C.c_int32), ("color", C.c_int32), ("gain", C.c_float), ("offset", C.c_float), ("crop", O.Box),
                ("even_rows", O.Frame16)]


def sane_frame(full, seed, window):
    """synthetic halves without Inf/NaN in the colours (the colour tables map them to themselves; fine, but noisy)"""
    return O.synth_frame(full, seed, window)


def run_case(gpu, full, specs, seeds=100):
    """specs: per layer (window, ops dict or None); ops dict keys: weave(bool), color, gain, offset, crop"""
    A, dev = gpu
    hosts, evens, orc_ops, keep = [], [], (OrcOps * len(specs))(), []
    layers = (A.Layer * len(specs))()
    ops_arr = (A.LayerOps * len(specs))()
    for i, (window, ops) in enumerate(specs):
        host = sane_frame(full, seeds + i, window)
        hosts.append(host)
        d = dev.DeviceFrame16.from_numpy(host.array, full, window)
        keep.append(d)
        layers[i].frame, layers[i].mix = d.c, 1.0
        if not ops:
            continue
        flags = 0
        if ops.get("weave"):
            even = sane_frame(full, seeds + 50 + i, ops.get("even_window", window))
            evens.append(even)
            de = dev.DeviceFrame16.from_numpy(even.array, full, even.current.tuple())
            keep.append(de)
            ops_arr[i].even_rows, orc_ops[i].even_rows = de.c, even.c
            flags |= A.CV_OP_WEAVE
        if ops.get("color"):
            ops_arr[i].color = orc_ops[i].color = ops["color"]
            flags |= A.CV_OP_COLOR
        if "gain" in ops:
            ops_arr[i].gain = orc_ops[i].gain = ops["gain"]
            ops_arr[i].offset = orc_ops[i].offset = ops.get("offset", 0.0)
            flags |= A.CV_OP_GAIN_OFFSET
        if "crop" in ops:
            ops_arr[i].crop, orc_ops[i].crop = A.Box2i(*ops["crop"]), O.Box(*ops["crop"])
            flags |= A.CV_OP_CROP
        ops_arr[i].flags = orc_ops[i].flags = flags
        layers[i].ops = C.pointer(ops_arr[i])

    # oracle
    want = O.HostFrame16(full)
    arr = (O.Frame16 * len(specs))(*[h.c for h in hosts])
    O.oracle().orc_composite_ops_f16(C.byref(want.c), arr, C.addressof(orc_ops), len(specs))

    # fused: one launch
    out = dev.DeviceFrame16(full)
    job = A.FrameJob()
    job.layer_count = len(specs)
    for i in range(len(specs)):
        job.layers[i] = layers[i]
    job.out = C.pointer(out.c)
    before = dev.lib().cv_kernel_launches()
    dev.render_frames([job])
    dev.sync()
    launches = dev.lib().cv_kernel_launches() - before
    assert out.current.tuple() == want.current.tuple()
    x0, y0, x1, y1 = want.current.tuple()
    got = out.download()
    ys, xs = slice(y0 - full[1], y1 - full[1] + 1), slice(x0 - full[0], x1 - full[0] + 1)       # array indices of the window
    assert got[ys, xs].size == (y1 - y0 + 1) * (x1 - x0 + 1) * 4
    assert np.array_equal(got[ys, xs], want.array[ys, xs]), "fused layer filters differ from the oracle"
    return launches, want, (hosts, evens)


COLOR_TO_XYZ, COLOR_TO_SRGB = 1, 2


def test_each_filter_alone_and_all_together(gpu):
    full = (0, 0, 255, 47)
    for ops in ({"gain": 1.25, "offset": -0.0625}, {"crop": (40, 5, 200, 40)}, {"color": COLOR_TO_XYZ}, {"color": COLOR_TO_SRGB},
                {"weave": True}, {"weave": True, "color": COLOR_TO_XYZ, "gain": 0.75, "offset": 0.125, "crop": (10, 3, 250, 44)}):
        launches, _, _ = run_case(gpu, full, [(full, None), (full, ops), (full, None)])
        assert launches == 1, ops          # evaluated inside the composite launch (k_rect), no extra pass


def test_filters_on_windowed_layers_with_odd_edges(gpu):
    """partial windows (rectangles of constant coverage), odd window edges (ragged pairs -> the per-pixel path), a
    filtered bottom layer, a crop rectangle that leaves the layer's window, a weave frame larger than the layer"""
    full = (-4, -3, 299, 60)
    specs = [((-4, -3, 299, 60), {"gain": 2.0, "offset": 0.5}),
             ((21, 4, 250, 50), {"color": COLOR_TO_XYZ, "crop": (0, 10, 400, 30)}),
             ((100, -3, 299, 33), {"weave": True, "even_window": (96, -3, 299, 40), "gain": 0.5, "offset": 0.0}),
             ((-4, 20, 155, 60), None)]
    launches, _, _ = run_case(gpu, full, specs, seeds=300)
    assert launches == 1


def test_fused_equals_separate_passes(gpu):
    """the same chain as separate device passes over each layer -- weave, colour, gain/offset on the widened frame, crop --
    then stacked in f32: identical bits"""
    A, dev = gpu
    full = (0, 0, 191, 31)
    window = (8, 2, 180, 29)
    ops = {"weave": True, "color": COLOR_TO_SRGB, "gain": 1.5, "offset": -0.25, "crop": (20, 4, 150, 20)}
    _, want, (hosts, evens) = run_case(gpu, full, [(full, None), (window, ops)], seeds=500)

    lib = dev.lib()
    bottom = dev.DeviceFrame16.from_numpy(hosts[0].array, full, full)
    top = dev.DeviceFrame16.from_numpy(hosts[1].array, full, window)
    even = dev.DeviceFrame16.from_numpy(evens[0].array, full, window)
    dev.check(lib.cv_weave_fields_f16(C.byref(top.c), C.byref(even.c), None), "weave")
    dev.check(lib.cv_color_xyz_to_srgb(C.byref(top.c), None), "colour")
    top32 = dev.DeviceFrame32(full)
    dev.check(lib.cv_frame_f16_to_f32(C.byref(top32.c), C.byref(top.c), None), "widen")
    dev.check(lib.cv_gain_offset_f32(C.byref(top32.c), 1.5, -0.25, None), "gain/offset")
    # crop on the f32 frame: zero fill outside the rectangle (cv_crop_f16 is the f16 form of the same shader)
    data = top32.download()
    x0, y0, x1, y1 = window
    mask = np.zeros(data.shape[:2], bool)
    mask[4:21, 20:151] = True
    inside = np.zeros(data.shape[:2], bool)
    inside[y0:y1 + 1, x0:x1 + 1] = True
    data[inside & ~mask] = 0.0
    top32 = dev.DeviceFrame32.from_numpy(data, full, window)
    acc = dev.DeviceFrame32(full)
    dev.check(lib.cv_frame_f16_to_f32(C.byref(acc.c), C.byref(bottom.c), None), "widen")
    dev.check(lib.cv_mix_over_f32(C.byref(acc.c), C.byref(top32.c), 1.0, None), "over")
    out = dev.DeviceFrame16(full)
    dev.check(lib.cv_frame_f32_to_f16(C.byref(out.c), C.byref(acc.c), None), "pack")
    dev.sync()
    assert np.array_equal(out.download(), want.array)


def test_weave_matches_restated_reference(gpu):
    """cv_weave_fields_f16 against the restatement of Pulldown23RemovalFilter.c:82-104, bit for bit"""
    A, dev = gpu
    full = (0, -2, 99, 37)
    for window, other in (((0, -2, 99, 37), (0, -2, 99, 37)), ((0, 1, 80, 30), (0, -2, 99, 37)), ((0, 0, 99, 37), (0, 4, 60, 20))):
        frame, even = O.synth_frame(full, 1, window), O.synth_frame(full, 2, other)
        want = frame.copy()
        O.oracle().orc_weave_fields_f16(C.byref(want.c), C.byref(even.c))
        d, e = dev.DeviceFrame16.from_numpy(frame.array, full, window), dev.DeviceFrame16.from_numpy(even.array, full, other)
        dev.check(dev.lib().cv_weave_fields_f16(C.byref(d.c), C.byref(e.c), None), "cv_weave_fields_f16")
        dev.sync()
        assert np.array_equal(d.download(), want.array)


def test_batch_of_filtered_frames_is_one_launch(gpu):
    A, dev = gpu
    full = (0, 0, 127, 23)
    specs = [(full, None), ((16, 2, 120, 20), {"gain": 1.125, "offset": 0.03125, "crop": (30, 0, 90, 23)})]
    jobs, outs, wants, keep = [], [], [], []
    for f in range(5):
        hosts = [O.synth_frame(full, 900 + 10 * f + i, w) for i, (w, _) in enumerate(specs)]
        devs = [dev.DeviceFrame16.from_numpy(h.array, full, h.current.tuple()) for h in hosts]
        ops = A.LayerOps()
        ops.flags, ops.gain, ops.offset, ops.crop = A.CV_OP_GAIN_OFFSET | A.CV_OP_CROP, 1.125, 0.03125, A.Box2i(30, 0, 90, 23)
        out = dev.DeviceFrame16(full)
        job = A.FrameJob()
        job.layer_count = 2
        job.layers[0].frame, job.layers[0].mix = devs[0].c, 1.0
        job.layers[1].frame, job.layers[1].mix, job.layers[1].ops = devs[1].c, 1.0, C.pointer(ops)
        job.out = C.pointer(out.c)
        orc = (OrcOps * 2)()
        orc[1].flags, orc[1].gain, orc[1].offset, orc[1].crop = 12, 1.125, 0.03125, O.Box(30, 0, 90, 23)
        want = O.HostFrame16(full)
        O.oracle().orc_composite_ops_f16(C.byref(want.c), (O.Frame16 * 2)(*[h.c for h in hosts]), C.addressof(orc), 2)
        jobs.append(job); outs.append(out); wants.append(want); keep += [hosts, devs, ops]
    before = dev.lib().cv_kernel_launches()
    dev.render_frames(jobs)
    dev.sync()
    assert dev.lib().cv_kernel_launches() - before == 1
    for out, want in zip(outs, wants):
        assert np.array_equal(out.download(), want.array)


def test_python_filters_ride_on_workspace_layers(gpu, monkeypatch):
    """fluggo.media.process level: VideoGainOffsetFilter and the weaving frame of Pulldown23RemovalFilter over f16 frames
    inside a VideoWorkspace are evaluated inside the ONE composite launch (cv_layer_ops); the result has the bits of the
    same graph with the filters as separate passes (CV_PY_NO_FUSED_FILTERS=1), as f16 and as f32, and on their own."""
    A, dev = gpu
    from canvas_b200 import process
    from canvas_b200.basetypes import box2i
    full = (0, -1, 95, 46)
    for weave_fits in (True, False):
        check_python_filter_graph(dev, monkeypatch, full, weave_fits)


def check_python_filter_graph(dev, monkeypatch, full, weave_fits):
    """weave_fits: the frame that supplies the even rows (clip frame 3) defines every pixel of the woven one (clip frame 2),
    so the weave can ride on the layer; otherwise that filter stays a separate pass (and the results still agree)"""
    from canvas_b200 import process
    from canvas_b200.basetypes import box2i
    partial = (8, 3, 90, 40)
    frames = [O.synth_frame(full, 1200 + i, partial if i % 3 == (2 if weave_fits else 0) else full) for i in range(10)]

    class Clip(process.VideoSource):
        def _describe(self, frame_index, window):
            f = frames[frame_index % len(frames)]
            return process.RgbaFrameF16(f.full.tuple(), f.current.tuple(), f.array)._describe(frame_index, window)

    def graph():
        ws = process.VideoWorkspace()
        bottom = O.synth_frame(full, 77, full)
        ws.add(source=process.RgbaFrameF16(full, full, bottom.array), x=0, length=100, z=0)
        ws.add(source=process.VideoGainOffsetFilter(Clip(), 1.5, -0.125), x=0, length=100, z=1)
        # output frame 2 with offset 0 is the weaving phase (frames 2 and 3 of the clip); gain/offset on top of the weave
        ws.add(source=process.VideoGainOffsetFilter(process.Pulldown23RemovalFilter(Clip(), 0), 0.75, 0.0625), x=0, length=100, z=2)
        return ws

    results = {}
    for fused in (True, False):
        monkeypatch.setenv("CV_PY_NO_FUSED_FILTERS", "0" if fused else "1")
        ws = graph()
        before = dev.lib().cv_kernel_launches()
        f16 = ws.get_frame_f16(2, box2i(*full))
        launches = dev.lib().cv_kernel_launches() - before
        f32 = ws.get_frame_f32(2, box2i(*full))
        alone = process.VideoGainOffsetFilter(process.Pulldown23RemovalFilter(Clip(), 0), 0.75, 0.0625).get_frame_f16(2, box2i(*full))
        results[fused] = (f16, f32, alone, launches)
    (a16, a32, aalone, fused_launches), (b16, b32, balone, separate_launches) = results[True], results[False]
    assert separate_launches > 3 and (fused_launches == 1 if weave_fits else fused_launches > 1)
    for x, y in ((a16, b16), (a32, b32), (aalone, balone)):
        assert x.current_window.flat() == y.current_window.flat()
        x0, y0, x1, y1 = x.current_window.flat()
        xs, ys = slice(x0 - full[0], x1 - full[0] + 1), slice(y0 - full[1], y1 - full[1] + 1)
        assert np.array_equal(np.asarray(x.data)[ys, xs].view(np.uint8), np.asarray(y.data)[ys, xs].view(np.uint8))


def test_random_layer_filters(gpu):
    """random stacks (1-6 layers, random windows with even and odd edges, some outside the frame) with random filter
    combinations per layer; fused launch against the oracle.  CV_OPS_FUZZ_CASES scales it (default 40)."""
    import os
    rng = np.random.default_rng(77)
    for case in range(int(os.environ.get("CV_OPS_FUZZ_CASES", "40"))):
        w, h = int(rng.choice([66, 130, 200, 257, 384])), int(rng.integers(4, 40))
        ox, oy = int(rng.choice([0, 0, -6, 3, 10])), int(rng.integers(-3, 4))
        full = (ox, oy, ox + w - 1, oy + h - 1)

        def box():
            if rng.random() < 0.3:
                return full
            a = int(rng.integers(ox - 4, ox + w - 1)); b = int(rng.integers(a, ox + w + 4))
            c = int(rng.integers(oy - 2, oy + h)); d = int(rng.integers(c, oy + h + 2))
            return (a, c, b, d)

        specs = []
        for _ in range(int(rng.integers(1, 7))):
            clip = lambda b: (max(b[0], full[0]), max(b[1], full[1]), min(b[2], full[2]), min(b[3], full[3]))
            window = clip(box())                # these frames share the output's full_window
            if window[2] < window[0] or window[3] < window[1]:
                window = full
            ops = {}
            if rng.random() < 0.3:              # the frame that supplies the even rows defines at least the layer's window
                ops["weave"] = True
                ops["even_window"] = clip((window[0] - int(rng.integers(0, 5)), window[1] - int(rng.integers(0, 4)),
                                           window[2] + int(rng.integers(0, 5)), window[3] + int(rng.integers(0, 4))))
            if rng.random() < 0.35:
                ops["color"] = int(rng.choice([COLOR_TO_XYZ, COLOR_TO_SRGB]))
            if rng.random() < 0.5:
                ops["gain"], ops["offset"] = float(np.float32(rng.uniform(-0.5, 2.0))), float(np.float32(rng.uniform(-0.25, 0.25)))
            if rng.random() < 0.35:
                ops["crop"] = box()
            specs.append((window, ops or None))
        run_case(gpu, full, specs, seeds=2000 + 20 * case)
