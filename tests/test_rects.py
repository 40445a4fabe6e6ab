"""The host-side rectangle decomposition of windowed frame programs (canvas_b200/csrc/cv_rects.c) against a
brute-force per-pixel evaluation of the coverage rule of the exact path -- tests/rect_selftest.c, plain C, no GPU.
(The semantics being decomposed: /root/reference/src/cprocess/video_mix.c:123-235, 251-368 and workspace.c:530-547.)"""
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_rectangle_plans_match_per_pixel_coverage(tmp_path):
    exe = str(tmp_path / "rect_selftest")
    subprocess.run(["gcc", "-O1", "-Wall", "-I", os.path.join(ROOT, "include"), "-I", os.path.join(ROOT, "canvas_b200", "csrc"),
                    "-I", "/usr/local/cuda/include", os.path.join(ROOT, "tests", "rect_selftest.c"),
                    os.path.join(ROOT, "canvas_b200", "csrc", "cv_rects.c"), "-o", exe], check=True)
    for seed in (1, 2, 3):
        out = subprocess.run([exe, "4000", str(seed)], capture_output=True, text=True)
        assert out.returncode == 0 and out.stdout.startswith("ok"), out.stdout + out.stderr
