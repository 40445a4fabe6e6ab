"""The MEASURED configurations against the oracle, at their real sizes and in their real launch forms.

Every bench.py workload is built exactly as bench.py builds it (its own InputSet / make_base), then
  * ONE frame goes through a single cv_render_frames() call (the single-launch form: k_stack_tma at 3840x2160, the
    non-batch k_seg instantiations), and
  * one full step's worth of frames of the BATCH form -- 16 x 4K or 120 x 1080p jobs in ONE cv_render_frames() call,
    i.e. the persistent BATCH=1 launches over count x segments with the job ring -- writes into a distinct output per
    frame, and EVERY frame of it is compared,
bit for bit with oracle/liboracle.so (which tests/test_oracle_vs_reference.py pins against the compiled reference).
Reference semantics: /root/reference/src/cprocess/workspace.c:494-550, video_mix.c:107-370, video_reconstruct.c:50-137,
src/cprocess/widget_gl.c:290-307.
"""
import ctypes as C
from concurrent.futures import ThreadPoolExecutor

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

import bench


@pytest.fixture(scope="module")
def dev():
    from canvas_b200 import device
    device.init(0)
    return device


DISTINCT = 3        # distinct input sets per workload (the oracle renders each once; 4K frames cost it about a second)


def batch_frames(name):
    return 16 if name.startswith("4k") else 120


def window_equal(got, want, window):
    x0, y0, x1, y1 = window
    return np.array_equal(got[y0:y1 + 1, x0:x1 + 1], want[y0:y1 + 1, x0:x1 + 1])


@pytest.mark.parametrize("name", sorted(bench.WORKLOADS))
def test_bench_workload_matches_oracle_at_full_size(dev, name):
    w, h = bench.WORKLOADS[name][:2]
    base = bench.make_base(dev, name, 4321)
    sets = [bench.InputSet(dev, name, base, pinned=False) for _ in range(DISTINCT)]
    with ThreadPoolExecutor(DISTINCT) as pool:          # ctypes releases the GIL: the oracle renders the sets side by side
        expected = list(pool.map(bench.expected_output, sets))

    # ---- the single-launch form ----
    launches = dev.lib().cv_kernel_launches()
    dev.render_frames([sets[0].job])
    dev.sync()
    assert dev.lib().cv_kernel_launches() - launches == 1
    want, window = expected[0]
    if sets[0].out_u8 is None:
        assert sets[0].out.current.tuple() == window
    assert window == (0, 0, w - 1, h - 1)
    assert window_equal(sets[0].output_array(), want, window), "%s: single launch differs from the oracle" % name

    # ---- one step of the batch form, a distinct output per frame ----
    count = batch_frames(name)
    full = (0, 0, w - 1, h - 1)
    if sets[0].out_u8 is not None:
        outs = [dev.DeviceBuffer(w * h * 4) for _ in range(count)]
    else:
        outs = [dev.DeviceFrame16(full) for _ in range(count)]
    jobs = [sets[f % DISTINCT].job_into(outs[f]) for f in range(count)]
    launches = dev.lib().cv_kernel_launches()
    dev.render_frames(jobs)
    dev.sync()
    used = dev.lib().cv_kernel_launches() - launches
    if name == "4k_8layer_composite":
        assert used == count            # the TMA composite launches once per frame
    else:
        assert used == 1, "%s: %d launches for one step" % (name, used)

    for f in range(count):
        want, window = expected[f % DISTINCT]
        got = sets[0].output_array(outs[f])
        assert window_equal(got, want, window), "%s: frame %d of the batch differs from the oracle" % (name, f)
        outs[f].free()
    for s in sets:
        s.free()


def test_bench_in_line_check_reports_zero_ulp(dev):
    """bench.py's own post-run check (what puts parity_checked / max_ulp into the bench line)."""
    name = "1080p_4layer_windows"
    base = bench.make_base(dev, name, 99)
    sets = [bench.InputSet(dev, name, base, pinned=False) for _ in range(2)]
    dev.render_frames([s.job for s in sets])
    dev.sync()
    result = bench.check_sets(sets, [0, 1])
    assert result == {"parity_checked": True, "max_ulp": 0, "frames_checked": 2}
    # ... and it does see a difference when there is one
    a = sets[0].output_array()
    b = a.copy()
    b[10, 10, 0] ^= 1
    assert bench.max_ulp(a, b) == 1
    for s in sets:
        s.free()
