"""Finds the first failing case of tests/test_gpu_layer_ops.py::test_random_layer_filters and describes the difference."""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import oraclelib as O
import test_gpu_layer_ops as T
from canvas_b200 import _abi as A
from canvas_b200 import device as dev

dev.init(0)
O.oracle().orc_composite_ops_f16.argtypes = [C.POINTER(O.Frame16), C.POINTER(O.Frame16), C.c_void_p, C.c_int]
O.oracle().orc_weave_fields_f16.argtypes = [C.POINTER(O.Frame16), C.POINTER(O.Frame16)]
gpu = (A, dev)

# replay the generator of the test
rng = np.random.default_rng(77)
COLOR_TO_XYZ, COLOR_TO_SRGB = 1, 2
limit = int(os.environ.get("CV_OPS_FUZZ_CASES", "40"))
only = int(sys.argv[1]) if len(sys.argv) > 1 else None
for case in range(limit):
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
        window = clip(box())
        if window[2] < window[0] or window[3] < window[1]:
            window = full
        ops = {}
        if rng.random() < 0.3:
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
    if only is not None and case != only:
        continue
    try:
        T.run_case(gpu, full, specs, seeds=2000 + 20 * case)
    except AssertionError as e:
        print("case", case, "full", full)
        for s in specs:
            print("   ", s)
        print(str(e)[:200])
        # redo by hand to locate the difference
        hosts = []
        import inspect
        src = inspect.getsource(T.run_case)
        # run_case again but capturing arrays: patch np.array_equal
        captured = {}
        real = np.array_equal
        def spy(a, b):
            captured["got"], captured["want"] = a.copy(), b.copy()
            return real(a, b)
        np.array_equal = spy
        made = []
        real_sane = T.sane_frame
        def rec(full_, seed_, window_):
            f = real_sane(full_, seed_, window_)
            made.append((seed_, window_, f))
            return f
        T.sane_frame = rec
        try:
            T.run_case(gpu, full, specs, seeds=2000 + 20 * case)
        except AssertionError:
            pass
        np.array_equal = real
        if "got" in captured:
            g, wv = captured["got"], captured["want"]
            bad = np.any(g != wv, axis=2)
            ys, xs = np.nonzero(bad)
            print("window-relative mismatches:", bad.sum(), "of", bad.size, "rows", ys.min(), ys.max(), "cols", xs.min(), xs.max())
            for y, x in list(zip(ys, xs))[:8]:
                print("   at", (x, y), "got", g[y, x], "want", wv[y, x])
            # the inputs at those places (frame coordinates = window-relative + the result window's corner; here the window
            # is the union of the layer windows)
            wx0 = min(sp[0][0] for sp in specs); wy0 = min(sp[0][1] for sp in specs)
            for seed_, window_, f in made:
                print("   input seed", seed_, "window", window_, "full", f.full.tuple())
                for y, x in list(zip(ys, xs))[:8]:
                    fx, fy = x + wx0 - f.full.min_x, y + wy0 - f.full.min_y
                    print("      frame px (%d,%d):" % (x + wx0, y + wy0), f.array[fy, fx])
        break
else:
    print("no failing case in", limit)
