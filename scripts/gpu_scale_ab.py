"""Times cv_scale_bilinear_f32 (device frames) for a few shapes; run once as is and once with CV_NO_SCALE_FUSED=1 for the
two-pass form.  Prints frames/s and GB/s of algorithmic bytes (source window + target)."""
import ctypes as C
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from canvas_b200 import _abi as A
from canvas_b200 import device as dev

dev.init(0)
lib = dev.lib()
CASES = [("4k->1080p", (3840, 2160), (1920, 1080), (0.5, 0.5)),
         ("1080p->4k", (1920, 1080), (3840, 2160), (2.0, 2.0)),
         ("4k->720p", (3840, 2160), (1280, 720), (1.0 / 3.0, 1.0 / 3.0)),
         ("1080p->720p x-first", (1920, 1080), (1280, 720), (0.6666, 0.66667)),
         ("4k anamorphic", (3840, 2160), (2880, 2160 * 2 // 3), (0.75, 2.0 / 3.0))]
rng = np.random.default_rng(0)
if len(sys.argv) > 1:
    CASES = [c for c in CASES if c[0] == sys.argv[1]]
for name, (sw, sh), (tw, th), (fx, fy) in CASES:
    sfull, tfull = (0, 0, sw - 1, sh - 1), (0, 0, tw - 1, th - 1)
    data = rng.uniform(0, 1, size=(sh, sw, 4)).astype(np.float32)
    sources = [dev.DeviceFrame32.from_numpy(np.roll(data, 5 * i, axis=1), sfull, sfull) for i in range(3)]
    targets = [dev.DeviceFrame32(tfull) for _ in range(3)]
    z, f = A.V2f(0.0, 0.0), A.V2f(fx, fy)
    def call(i):
        dev.check(lib.cv_scale_bilinear_f32(C.byref(targets[i % 3].c), z, C.byref(sources[i % 3].c), z, f, None), "scale")
    for i in range(6):
        call(i)
    dev.sync()
    before = lib.cv_kernel_launches()
    t = dev.Timer()
    n = 60
    t.begin()
    for i in range(n):
        call(i)
    t.end()
    ms = t.elapsed_ms() / n
    bytes_ = (sw * sh + tw * th) * 16          # nominal; for factors < 1 the reference computes only a share (video_scale.c:253-256)
    print("%-22s %8.4f ms  %8.0f fps  %7.0f GB/s  launches/call %.1f" % (name, ms, 1000.0 / ms, bytes_ / ms / 1e6, (lib.cv_kernel_launches() - before) / n))
    for x in sources + targets:
        x.free()
