"""cv_scale_bilinear_static_f32 (canvas_b200/csrc/cv_static.cu): the two-axis scaler with its scratch memory -- tap tables
and the intermediate frame -- in a static device array instead of pooled allocations.  No counterpart in the reference;
the check is against the existing entry point cv_scale_bilinear_f32
through the C-ABI, bit for bit (the tolerance every scaler test here uses: none), on calls in a row that share the array.

Named test_zz_* so that pytest collects it last: a failure here must not hide the other tests under `pytest -x`.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

CV_ERR_UNSUPPORTED = -5

# (fx, fy, target_point, source_point, source window, target full)
CASES = [
    (1.5, 1.5, (0.0, 0.0), (0.0, 0.0), (0, 0, 1279, 719), (0, 0, 1919, 1079)),        # 720p -> 1080p, both passes, large
    (2.0, 2.0, (0.0, 0.0), (0.0, 0.0), (0, 0, 39, 23), (0, 0, 79, 47)),
    (0.5, 0.5, (0.0, 0.0), (0.0, 0.0), (0, 0, 79, 47), (0, 0, 39, 23)),
    (1.0, 3.0, (2.0, 1.5), (2.0, -0.25), (3, 2, 50, 20), (-4, -6, 60, 70)),           # vertical only
    (0.37, 1.0, (5.0, 0.0), (11.5, 0.0), (-10, 0, 90, 12), (0, -3, 44, 20)),          # horizontal only
    (1.7, 0.6, (10.0, 10.0), (20.25, 15.5), (0, 0, 63, 63), (-20, -20, 100, 60)),     # vertical (smaller factor) first
    (0.25, 3.5, (0.0, 0.0), (0.0, 0.0), (8, 4, 71, 19), (0, 0, 30, 80)),              # horizontal first
    (1.0, 1.0, (3.0, 4.0), (3.0, 4.0), (2, 2, 20, 20), (0, 0, 15, 30)),               # identity: windowed copy
    (1.0, 1.0, (3.5, 4.0), (3.0, 4.0), (2, 2, 20, 20), (0, 0, 25, 30)),               # factor 1 with a sub-pixel shift
    (4.0, 4.0, (0.0, 0.0), (0.0, 0.0), (5, 5, 9, 9), (100, 100, 120, 120)),           # target misses the image
    (1.0 / 16.0, 1.0 / 3.0, (0.0, 0.0), (0.0, 0.0), (0, 0, 255, 59), (0, 0, 20, 25)),
]


@pytest.fixture(scope="module")
def gpu():
    from canvas_b200 import _abi as A
    from canvas_b200 import device as dev
    dev.init(0)
    return A, dev


def source_of(dev, case, seed):
    fx, fy, tp, sp, swin, tfull = case
    rng = np.random.default_rng(seed)
    sfull = (swin[0] - 3, swin[1] - 2, swin[2] + 4, swin[3] + 1)
    data = rng.uniform(-0.5, 1.5, size=(sfull[3] - sfull[1] + 1, sfull[2] - sfull[0] + 1, 4)).astype(np.float32)
    return dev.DeviceFrame32.from_numpy(data, sfull, swin)


def run(gpu, entry, case, src, fill):
    """one call of `entry` into a fresh target prefilled with `fill` bytes; returns (rc, window, pixels)"""
    A, dev = gpu
    fx, fy, tp, sp, swin, tfull = case
    out = dev.DeviceFrame32(tfull)
    dev.check(dev.lib().cv_memset(out.c.data, fill, out.full.width * out.full.height * 16, None), "cv_memset")
    rc = getattr(dev.lib(), entry)(C.byref(out.c), A.V2f(*tp), C.byref(src.c), A.V2f(*sp), A.V2f(fx, fy), None)
    dev.sync()
    w = out.current
    return rc, (w.min.x, w.min.y, w.max.x, w.max.y), out.download()


def same(a, b):
    (rc_a, win_a, px_a), (rc_b, win_b, px_b) = a, b
    empty_a, empty_b = win_a[2] < win_a[0] or win_a[3] < win_a[1], win_b[2] < win_b[0] or win_b[3] < win_b[1]
    return rc_a == rc_b == 0 and (win_a == win_b or (empty_a and empty_b)) and np.array_equal(px_a.view(np.uint32), px_b.view(np.uint32))


def test_static_scaler_equals_the_allocating_one_on_calls_in_a_row(gpu):
    """every case twice in a row through the static entry point, interleaved with the other cases so that each call finds
    the previous call's tables and intermediate frame in the array: bits and windows equal cv_scale_bilinear_f32's"""
    A, dev = gpu
    for index, case in enumerate(CASES):
        src = source_of(dev, case, 4000 + index)
        want = run(gpu, "cv_scale_bilinear_f32", case, src, 0x5A)
        first = run(gpu, "cv_scale_bilinear_static_f32", case, src, 0x5A)
        second = run(gpu, "cv_scale_bilinear_static_f32", case, src, 0x5A)
        assert same(first, want), "first static call differs: %r (%s)" % (case, dev.lib().cv_last_error())
        assert same(second, want), "second static call differs: %r" % (case,)

    # large, then small, then large again: stale contents of a bigger call must not reach a smaller one and vice versa
    big, small = CASES[0], CASES[5]
    src_big, src_small = source_of(dev, big, 1), source_of(dev, small, 2)
    want_big, want_small = run(gpu, "cv_scale_bilinear_f32", big, src_big, 0xA5), run(gpu, "cv_scale_bilinear_f32", small, src_small, 0xA5)
    for case, src, want in ((big, src_big, want_big), (small, src_small, want_small), (big, src_big, want_big), (small, src_small, want_small)):
        assert same(run(gpu, "cv_scale_bilinear_static_f32", case, src, 0xA5), want), case


def test_static_scaler_launches_its_own_kernel_and_allocates_nothing(gpu):
    A, dev = gpu
    import torch
    case = CASES[0]
    src = source_of(dev, case, 77)
    run(gpu, "cv_scale_bilinear_static_f32", case, src, 0)          # set-up (symbol address) happens on the first call
    fx, fy, tp, sp, swin, tfull = case
    out = dev.DeviceFrame32(tfull)
    dev.sync()
    free_before = torch.cuda.mem_get_info(0)[0]
    launches_before = dev.lib().cv_kernel_launches()
    for _ in range(2):
        dev.check(dev.lib().cv_scale_bilinear_static_f32(C.byref(out.c), A.V2f(*tp), C.byref(src.c), A.V2f(*sp), A.V2f(fx, fy), None),
                  "cv_scale_bilinear_static_f32")
    dev.sync()
    assert dev.lib().cv_kernel_launches() - launches_before == 4          # two passes per call
    assert torch.cuda.mem_get_info(0)[0] == free_before                   # no device memory taken or returned


def test_sizes_beyond_the_static_scratch_are_rejected(gpu):
    A, dev = gpu
    # 4K target from a 4K-wide source: the intermediate frame is 3840 x 1080 = 4.1 Mi pixels > CV_STATIC_MAX_TEMP_PIXELS
    case = (1.0 + 1.0 / 64.0, 2.0, (0.0, 0.0), (0.0, 0.0), (0, 0, 3839, 1079), (0, 0, 3839, 2159))
    fx, fy, tp, sp, swin, tfull = case
    src = dev.DeviceFrame32(swin, swin)
    out = dev.DeviceFrame32(tfull)
    launches_before = dev.lib().cv_kernel_launches()
    rc = dev.lib().cv_scale_bilinear_static_f32(C.byref(out.c), A.V2f(*tp), C.byref(src.c), A.V2f(*sp), A.V2f(fx, fy), None)
    assert rc == CV_ERR_UNSUPPORTED
    message = dev.lib().cv_last_error().decode()
    assert "static scratch" in message and "pixels" in message, message
    assert dev.lib().cv_kernel_launches() == launches_before
    w = out.current
    assert w.max.x < w.min.x or w.max.y < w.min.y

    # more target lines on one axis than the tables hold
    wide = dev.DeviceFrame32((0, 0, 16999, 1))
    narrow = dev.DeviceFrame32((0, 0, 8499, 1), (0, 0, 8499, 1))
    dev.check(dev.lib().cv_memset(narrow.c.data, 0, 8500 * 2 * 16, None), "cv_memset")
    rc = dev.lib().cv_scale_bilinear_static_f32(C.byref(wide.c), A.V2f(0, 0), C.byref(narrow.c), A.V2f(0, 0), A.V2f(2.0, 1.0), None)
    assert rc == CV_ERR_UNSUPPORTED and "target lines" in dev.lib().cv_last_error().decode()
    # ... while the allocating entry point takes both
    dev.check(dev.lib().cv_scale_bilinear_f32(C.byref(wide.c), A.V2f(0, 0), C.byref(narrow.c), A.V2f(0, 0), A.V2f(2.0, 1.0), None), "scale")
    dev.sync()
