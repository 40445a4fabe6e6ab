"""GPU parity of the next rows of SURVEY.md section 8(f): the bilinear scaler (video_scale.c:231-319, oracle pinned
against the compiled reference in test_oracle_vs_reference.py), the 2:3 pulldown field weave
(Pulldown23RemovalFilter.c:46-105) and gain/offset (video_filter.c:27-39, GL-only in the reference: parity unpinned,
checked against its arithmetic restated in numpy f32)."""
import ctypes as C

import numpy as np
import pytest

import oraclelib as O
from test_oracle_vs_reference import SCALE_CASES

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev():
    from canvas_b200 import device
    device.init(0)
    return device


def source_frame(case, seed=None):
    fx, fy, tp, sp, swin, tfull = case
    rng = np.random.default_rng(int(fx * 1000 + fy * 10) if seed is None else seed)
    sfull = (swin[0] - 3, swin[1] - 2, swin[2] + 4, swin[3] + 1)
    data = rng.uniform(-0.5, 1.5, size=(sfull[3] - sfull[1] + 1, sfull[2] - sfull[0] + 1, 4)).astype(np.float32)
    return O.HostFrame32(sfull, swin, data)


def check_scaled(got_array, got_window, want, identity, what):
    if want.current.empty:
        assert got_window[2] < got_window[0] or got_window[3] < got_window[1], what
    else:
        assert tuple(got_window) == want.current.tuple(), what
    if identity:
        b, f = want.current, want.full
        ys, xs = slice(b.min_y - f.min_y, b.max_y - f.min_y + 1), slice(b.min_x - f.min_x, b.max_x - f.min_x + 1)
        assert np.array_equal(got_array[ys, xs].view(np.uint32), want.array[ys, xs].view(np.uint32)), what
    else:
        assert np.array_equal(got_array.view(np.uint32), want.array.view(np.uint32)), what      # zero fill included


@pytest.mark.parametrize("case", SCALE_CASES)
def test_scale_device_matches_oracle(dev, case):
    from canvas_b200 import _abi as A
    fx, fy, tp, sp, swin, tfull = case
    src = source_frame(case)
    want = O.HostFrame32(tfull)
    O.oracle().orc_scale_bilinear_f32(C.byref(want.c), tp[0], tp[1], C.byref(src.c), sp[0], sp[1], fx, fy)

    dsrc = dev.DeviceFrame32.from_numpy(src.array, src.full.tuple(), src.current.tuple())
    out = dev.DeviceFrame32(tfull)
    dev.check(dev.lib().cv_memset(out.c.data, 0x5A, out.full.width * out.full.height * 16, None), "cv_memset")
    dev.check(dev.lib().cv_scale_bilinear_f32(C.byref(out.c), A.V2f(*tp), C.byref(dsrc.c), A.V2f(*sp), A.V2f(fx, fy), None), "cv_scale_bilinear_f32")
    dev.sync()
    check_scaled(out.download(), out.current.tuple(), want, fx == 1.0 and fy == 1.0 and tp == sp, "device scale %r" % (case,))


@pytest.mark.parametrize("case", SCALE_CASES[:6])
def test_scale_host_entry_point_matches_oracle(dev, case):
    """framework.h's video_scale_bilinear_f32 with HOST frames (the reference's own signature, v2f by value)."""
    from canvas_b200 import _abi as A
    lib = dev.lib()
    lib.video_scale_bilinear_f32.argtypes = [C.POINTER(O.Frame32), A.V2f, C.POINTER(O.Frame32), A.V2f, A.V2f]
    lib.video_scale_bilinear_f32.restype = None
    fx, fy, tp, sp, swin, tfull = case
    src = source_frame(case)
    want = O.HostFrame32(tfull)
    O.oracle().orc_scale_bilinear_f32(C.byref(want.c), tp[0], tp[1], C.byref(src.c), sp[0], sp[1], fx, fy)
    got = O.HostFrame32(tfull)
    got.array[...] = 3.0
    lib.video_scale_bilinear_f32(C.byref(got.c), A.V2f(*tp), C.byref(src.c), A.V2f(*sp), A.V2f(fx, fy))
    check_scaled(got.array, got.current.tuple(), want, False, "host scale %r" % (case,))


def test_cached_tap_tables_hits_and_replacement(dev):
    """cv_scale.c keeps the uploaded tap tables of the last 12 geometries per device: 20 geometries in rotation, three
    rounds (misses, replacements, then -- for the last 12 -- hits), different source pixels every round, each result
    against the oracle; then the same geometry on two streams at once."""
    from canvas_b200 import _abi as A
    cases = []
    for i in range(20):
        fx, fy = (0.4 + 0.13 * i, 2.2 - 0.09 * i) if i % 2 else (1.9 - 0.07 * i, 0.45 + 0.11 * i)
        cases.append((fx, fy, (1.0 + i, 2.0), (0.5 * i, 1.25), (i, 2, 60 + i, 40), (-3, -2, 90, 70 - i)))
    for rounds in range(3):
        for index, case in enumerate(cases):
            fx, fy, tp, sp, swin, tfull = case
            src = source_frame(case, seed=1000 * rounds + index)
            want = O.HostFrame32(tfull)
            O.oracle().orc_scale_bilinear_f32(C.byref(want.c), tp[0], tp[1], C.byref(src.c), sp[0], sp[1], fx, fy)
            dsrc = dev.DeviceFrame32.from_numpy(src.array, src.full.tuple(), src.current.tuple())
            out = dev.DeviceFrame32(tfull)
            dev.check(dev.lib().cv_memset(out.c.data, 0x33, out.full.width * out.full.height * 16, None), "cv_memset")
            dev.check(dev.lib().cv_scale_bilinear_f32(C.byref(out.c), A.V2f(*tp), C.byref(dsrc.c), A.V2f(*sp), A.V2f(fx, fy), None), "scale")
            dev.sync()
            check_scaled(out.download(), out.current.tuple(), want, False, "round %d %r" % (rounds, case))

    # a table uploaded on one stream and used from another at once: the second stream waits for the upload
    case = (1.3, 0.7, (2.0, 3.0), (1.5, 0.25), (0, 0, 700, 500), (0, 0, 900, 360))
    fx, fy, tp, sp, swin, tfull = case
    src = source_frame(case, seed=5)
    want = O.HostFrame32(tfull)
    O.oracle().orc_scale_bilinear_f32(C.byref(want.c), tp[0], tp[1], C.byref(src.c), sp[0], sp[1], fx, fy)
    dsrc = dev.DeviceFrame32.from_numpy(src.array, src.full.tuple(), src.current.tuple())
    streams = [C.c_void_p(), C.c_void_p()]
    outs = [dev.DeviceFrame32(tfull), dev.DeviceFrame32(tfull)]
    for st in streams:
        dev.check(dev.lib().cv_stream_create(C.byref(st)), "cv_stream_create")
    for st, out in zip(streams, outs):
        dev.check(dev.lib().cv_scale_bilinear_f32(C.byref(out.c), A.V2f(*tp), C.byref(dsrc.c), A.V2f(*sp), A.V2f(fx, fy), st), "scale")
    for st, out in zip(streams, outs):
        dev.sync(st)
        check_scaled(out.download(), out.current.tuple(), want, False, "two streams")
        dev.lib().cv_stream_destroy(st)


def test_fused_scaler_equals_two_pass_on_random_geometries(dev):
    """k_scale_fused (one launch, intermediate frame in shared memory) against the two-pass form -- which
    cv_scale_bilinear_static_f32 still is -- and against the oracle: 60 random geometries (CV_SCALE_FUZZ_CASES scales it) (factors 0.15 .. 5 per axis,
    fractional points, source windows inside larger frames, negative origins, targets that the image only partly covers)"""
    from canvas_b200 import _abi as A
    rng = np.random.default_rng(20260)
    fused_launches = 0
    import os
    count = int(os.environ.get("CV_SCALE_FUZZ_CASES", "60"))
    for index in range(count):
        fx, fy = (float(np.float32(np.exp(rng.uniform(np.log(0.15), np.log(5.0))))) for _ in range(2))
        sw, sh = int(rng.integers(8, 400)), int(rng.integers(6, 240))
        sx0, sy0 = int(rng.integers(-50, 50)), int(rng.integers(-50, 50))
        swin = (sx0, sy0, sx0 + sw - 1, sy0 + sh - 1)
        sp = (float(np.float32(sx0 + rng.uniform(0, sw))), float(np.float32(sy0 + rng.uniform(0, sh))))
        tw, th = int(rng.integers(8, 500)), int(rng.integers(6, 300))
        tx0, ty0 = int(rng.integers(-60, 60)), int(rng.integers(-60, 60))
        tfull = (tx0, ty0, tx0 + tw - 1, ty0 + th - 1)
        tp = (float(np.float32(tx0 + rng.uniform(0, tw))), float(np.float32(ty0 + rng.uniform(0, th))))
        case = (fx, fy, tp, sp, swin, tfull)
        src = source_frame(case, seed=7000 + index)
        want = O.HostFrame32(tfull)
        O.oracle().orc_scale_bilinear_f32(C.byref(want.c), tp[0], tp[1], C.byref(src.c), sp[0], sp[1], fx, fy)
        dsrc = dev.DeviceFrame32.from_numpy(src.array, src.full.tuple(), src.current.tuple())
        results = []
        for entry in ("cv_scale_bilinear_f32", "cv_scale_bilinear_static_f32"):
            out = dev.DeviceFrame32(tfull)
            dev.check(dev.lib().cv_memset(out.c.data, 0x5A, out.full.width * out.full.height * 16, None), "cv_memset")
            before = dev.lib().cv_kernel_launches()
            dev.check(getattr(dev.lib(), entry)(C.byref(out.c), A.V2f(*tp), C.byref(dsrc.c), A.V2f(*sp), A.V2f(fx, fy), None), entry)
            dev.sync()
            if entry == "cv_scale_bilinear_f32" and dev.lib().cv_kernel_launches() - before == 1:
                fused_launches += 1
            check_scaled(out.download(), out.current.tuple(), want, False, "%s %r" % (entry, case))
            results.append((out.current.tuple(), out.download()))
        assert np.array_equal(results[0][1].view(np.uint32), results[1][1].view(np.uint32)), case
    assert fused_launches >= count // 4 or os.environ.get("CV_NO_SCALE_FUSED", "0") not in ("", "0")          # the others: strong reductions whose tiles do not fit, empty results


def test_scaler_from_four_threads_with_more_geometries_than_the_cache_holds(dev):
    """four host threads, a stream each, 16 geometries in different orders (the cache of uploaded tap tables holds 12 per
    device: hits, misses and replacements interleave); every result against the oracle"""
    import threading
    from canvas_b200 import _abi as A
    cases, wants, sources = [], [], []
    for i in range(16):
        fx, fy = (0.5 + 0.2 * i, 2.6 - 0.12 * i)
        case = (fx, fy, (2.0 + i, 1.0), (0.25 * i, 3.5), (i, 1, 90 + i, 60), (-2, -3, 120, 80 + i))
        src = source_frame(case, seed=300 + i)
        want = O.HostFrame32(case[5])
        O.oracle().orc_scale_bilinear_f32(C.byref(want.c), case[2][0], case[2][1], C.byref(src.c), case[3][0], case[3][1], fx, fy)
        cases.append(case); wants.append(want)
        sources.append(dev.DeviceFrame32.from_numpy(src.array, src.full.tuple(), src.current.tuple()))
    failures = []

    def worker(seed):
        try:
            rng = np.random.default_rng(seed)
            dev.check(dev.lib().cv_set_device(0), "cv_set_device")          # the device is a per-thread choice, like CUDA's
            stream = C.c_void_p()
            dev.check(dev.lib().cv_stream_create(C.byref(stream)), "cv_stream_create")
            for _ in range(3):
                for i in rng.permutation(16):
                    fx, fy, tp, sp, swin, tfull = cases[i]
                    out = dev.DeviceFrame32(tfull)
                    dev.check(dev.lib().cv_scale_bilinear_f32(C.byref(out.c), A.V2f(*tp), C.byref(sources[i].c), A.V2f(*sp), A.V2f(fx, fy), stream), "scale")
                    got = out.download(stream)
                    check_scaled(got, out.current.tuple(), wants[i], False, "thread %d case %d" % (seed, i))
            dev.lib().cv_stream_destroy(stream)
        except BaseException as e:          # noqa: BLE001 -- reported by the main thread
            failures.append(repr(e))

    threads = [threading.Thread(target=worker, args=(s,)) for s in range(4)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not failures, failures[:2]


def test_scale_1080p_to_4k_and_back_properties(dev):
    """Full-size properties: a constant image stays constant in the interior under 2x enlargement (the triangle taps
    of every target pixel sum to one) and under 1/2 reduction (taps normalised, filter.c:69-75).  The reduction is run
    one axis at a time: the two-axis form sizes its intermediate frame with `* factor` (video_scale.c:253-256, :270-273),
    which for factors below one keeps only the first `factor` share of the target -- kept as the reference has it, and
    covered bit for bit by the mixed-factor cases above."""
    from canvas_b200 import _abi as A
    w, h = 1920, 1080
    small = (0, 0, w - 1, h - 1)
    big = (0, 0, 2 * w - 1, 2 * h - 1)
    flat = np.full((h, w, 4), 0.625, np.float32)
    dsrc = dev.DeviceFrame32.from_numpy(flat, small, small)
    up = dev.DeviceFrame32(big)
    dev.check(dev.lib().cv_scale_bilinear_f32(C.byref(up.c), A.V2f(0, 0), C.byref(dsrc.c), A.V2f(0, 0), A.V2f(2.0, 2.0), None), "scale up")
    dev.sync()
    a = up.download()
    assert up.current.tuple()[0] <= 0 and up.current.tuple()[2] >= 2 * w - 2
    assert np.all(a[2:-2, 2:-2] == np.float32(0.625))

    half_w = dev.DeviceFrame32((0, 0, w - 1, 2 * h - 1))
    dev.check(dev.lib().cv_scale_bilinear_f32(C.byref(half_w.c), A.V2f(0, 0), C.byref(up.c), A.V2f(0, 0), A.V2f(0.5, 1.0), None), "scale down x")
    down = dev.DeviceFrame32(small)
    dev.check(dev.lib().cv_scale_bilinear_f32(C.byref(down.c), A.V2f(0, 0), C.byref(half_w.c), A.V2f(0, 0), A.V2f(1.0, 0.5), None), "scale down y")
    dev.sync()
    assert down.current.tuple() == small
    b = down.download()
    assert np.allclose(b[2:-2, 2:-2], 0.625, rtol=0, atol=2e-7)


def test_video_scaler_type_matches_oracle(dev):
    from canvas_b200 import process
    from canvas_b200.basetypes import box2i
    case = SCALE_CASES[4]
    fx, fy, tp, sp, swin, tfull = case
    src = source_frame(case)
    source = process.RgbaFrameF32(src.full.tuple(), src.current.tuple(), src.array)
    rect = box2i(-1000, -1000, 1000, 1000)
    got = process.VideoScaler(source, tp, sp, (fx, fy), rect).get_frame_f32(0, box2i(*tfull))

    # the pull form first fetches the visible part of the source (video_scale.c:301-311), then scales
    f32 = np.float32
    fetch = (int(np.trunc(f32(sp[0]) - (f32(tp[0]) - f32(tfull[0])) / f32(fx))) - 1, int(np.trunc(f32(sp[1]) - (f32(tp[1]) - f32(tfull[1])) / f32(fy))) - 1,
             int(np.trunc(f32(sp[0]) + (f32(tfull[2]) - f32(tp[0])) / f32(fx))) + 1, int(np.trunc(f32(sp[1]) + (f32(tfull[3]) - f32(tp[1])) / f32(fy))) + 1)
    part = O.HostFrame32(fetch)
    O.oracle().orc_copy_frame_alpha_f32(C.byref(part.c), C.byref(src.c), 1.0)
    want = O.HostFrame32(tfull)
    O.oracle().orc_scale_bilinear_f32(C.byref(want.c), tp[0], tp[1], C.byref(part.c), sp[0], sp[1], fx, fy)
    assert got.current_window.flat() == want.current.tuple()
    b, f = want.current, want.full
    ys, xs = slice(b.min_y - f.min_y, b.max_y - f.min_y + 1), slice(b.min_x - f.min_x, b.max_x - f.min_x + 1)
    assert np.array_equal(got.data[ys, xs].view(np.uint32), want.array[ys, xs].view(np.uint32))


def test_pulldown_removal_weaves_fields(dev):
    from canvas_b200 import process
    from canvas_b200.basetypes import box2i
    full = (0, -1, 47, 22)
    frames = [O.synth_frame(full, 900 + i, full) for i in range(12)]

    class Clip(process.VideoSource):
        def _describe(self, frame_index, window):
            f = frames[frame_index % len(frames)]
            return process.RgbaFrameF16(f.full.tuple(), f.current.tuple(), f.array)._describe(frame_index, window)

    for offset in range(5):
        filt = process.Pulldown23RemovalFilter(Clip(), offset)
        for out_frame in range(1 if offset else 0, 8):
            phase = (out_frame + 3) & 3 if offset == 4 else (out_frame + offset) & 3
            base = ((out_frame + offset) >> 2) * 5 - offset
            got = filt.get_frame_f16(out_frame, box2i(*full))
            if phase != 2:
                want = frames[(base + {0: 0, 1: 1, 3: 4}[phase]) % 12].array
            else:
                want = frames[(base + 2) % 12].array.copy()
                other = frames[(base + 3) % 12].array
                ys = [y for y in range(full[1], full[3] + 1) if y % 2 == 0]
                for y in ys:
                    want[y - full[1]] = other[y - full[1]]
            assert got.current_window.flat() == full
            assert np.array_equal(got.data, want), (offset, out_frame)


def test_gain_offset_arithmetic(dev):
    from canvas_b200 import process
    from canvas_b200.basetypes import box2i
    full = (0, 0, 63, 31)
    rng = np.random.default_rng(4)
    data = rng.uniform(-1, 2, size=(32, 64, 4)).astype(np.float32)
    src = process.RgbaFrameF32(full, (3, 2, 60, 29), data)
    gain, offset = 1.75, -0.125
    got = process.VideoGainOffsetFilter(src, gain, offset).get_frame_f32(0, box2i(*full))
    assert got.current_window.flat() == (3, 2, 60, 29)
    want = data.copy()
    want[..., :3] = want[..., :3] * np.float32(gain) + np.float32(offset)       # mul, then add: two roundings
    assert np.array_equal(got.data[2:30, 3:61].view(np.uint32), want[2:30, 3:61].view(np.uint32))


def test_scaler_on_a_second_gpu(dev):
    """the tap-table cache and the static scratch are per device: the same geometry on GPU 1 after GPU 0, both entry
    points, against the oracle (skipped on a one-GPU box)"""
    import threading
    from canvas_b200 import _abi as A
    lib = dev.lib()
    if lib.cv_device_count() < 2:
        pytest.skip("needs two GPUs")
    case = (1.5, 1.25, (2.0, 1.0), (0.5, 3.5), (3, 1, 200, 120), (-2, -3, 320, 170))
    fx, fy, tp, sp, swin, tfull = case
    src = source_frame(case, seed=99)
    want = O.HostFrame32(tfull)
    O.oracle().orc_scale_bilinear_f32(C.byref(want.c), tp[0], tp[1], C.byref(src.c), sp[0], sp[1], fx, fy)
    failures = []

    def on_device(index):
        try:
            dev.check(lib.cv_init(index), "cv_init")            # selects the device for this thread
            dsrc = dev.DeviceFrame32.from_numpy(src.array, src.full.tuple(), src.current.tuple())
            for entry in ("cv_scale_bilinear_f32", "cv_scale_bilinear_static_f32", "cv_scale_bilinear_f32"):
                out = dev.DeviceFrame32(tfull)
                dev.check(getattr(lib, entry)(C.byref(out.c), A.V2f(*tp), C.byref(dsrc.c), A.V2f(*sp), A.V2f(fx, fy), None), entry)
                dev.sync()
                check_scaled(out.download(), out.current.tuple(), want, False, "%s on GPU %d" % (entry, index))
                out.free()
            dsrc.free()
        except BaseException as e:          # noqa: BLE001
            failures.append("GPU %d: %r" % (index, e))

    for index in (0, 1, 0):
        t = threading.Thread(target=on_device, args=(index,))
        t.start()
        t.join()
    assert not failures, failures
