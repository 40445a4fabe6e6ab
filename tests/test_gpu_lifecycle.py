"""cv_shutdown() followed by cv_init(): everything the library caches per device -- pooled blocks, the pinned pool, the
scaler's uploaded tap tables, per-thread staging slabs -- is either released or survives intact, and the second life
produces the bits of the first.  Run in a child process: the other test modules keep device frames alive across tests."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

SCRIPT = r"""
import ctypes as C, sys
import numpy as np
sys.path.insert(0, %(root)r)
sys.path.insert(0, %(tests)r)
from canvas_b200 import _abi as A
from canvas_b200 import device as dev

def one_life():
    dev.init(0)
    lib = dev.lib()
    rng = np.random.default_rng(5)
    # the scaler twice (a miss, then a hit of the tap-table cache), fused and single-axis geometries
    src = rng.uniform(0, 1, size=(90, 160, 4)).astype(np.float32)
    dsrc = dev.DeviceFrame32.from_numpy(src, (0, 0, 159, 89), (0, 0, 159, 89))
    results = []
    for factors in ((1.5, 1.25), (1.0, 2.0)):
        for _ in range(2):
            out = dev.DeviceFrame32((0, 0, 239, 119))
            dev.check(lib.cv_scale_bilinear_f32(C.byref(out.c), A.V2f(0, 0), C.byref(dsrc.c), A.V2f(0, 0), A.V2f(*factors), None), "scale")
            dev.sync()
            results.append(out.download().copy())
            out.free()
    # a layer stack through the device ABI
    full = (0, 0, 127, 47)
    layers = [dev.DeviceFrame16.from_numpy(rng.integers(0, 0x3C00, size=(48, 128, 4), dtype=np.uint16), full, full) for _ in range(3)]
    out = dev.DeviceFrame16(full)
    dev.render_frames([dev.make_job(out=out, layers=layers)])
    dev.sync()
    results.append(out.download().copy())
    for f in layers + [out, dsrc]:
        f.free()
    dev.sync()
    lib.cv_shutdown()
    dev._initialised.clear()
    return results

first = one_life()
second = one_life()
third = one_life()
assert len(first) == len(second) == len(third) == 5
for a, b, c in zip(first, second, third):
    assert np.array_equal(a.view(np.uint8), b.view(np.uint8)) and np.array_equal(a.view(np.uint8), c.view(np.uint8))
# after the last shutdown every call reports the missing initialisation instead of touching freed state
rc = dev._abi.load().cv_device_sync()
assert rc == -1, rc
print("lifecycle ok")
"""


def test_shutdown_and_reinit_in_a_child_process():
    code = SCRIPT % {"root": ROOT, "tests": os.path.join(ROOT, "tests")}
    run = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=600)
    assert run.returncode == 0 and "lifecycle ok" in run.stdout, (run.stdout[-2000:], run.stderr[-4000:])
