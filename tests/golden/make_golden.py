"""Generates tests/golden/reference_vectors.npz from the REFERENCE's own C code
(oracle/_ref/libcanvasref.so = /root/reference/src/cprocess compiled unmodified, `make -C oracle ref`).

Run in the build container only (needs /root/reference):   python tests/golden/make_golden.py
The vectors are small on purpose; tests/test_golden.py checks the oracle against them everywhere
(also on the GPU box, where /root/reference does not exist), and the GPU parity tests check the
kernels against the oracle.
"""
import ctypes as C
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import oraclelib as O  # noqa: E402


# (fx, fy, target_point, source_point, source window, target full window)
SCALE_GOLDEN = [
    (2.0, 2.0, (0.0, 0.0), (0.0, 0.0), (0, 0, 39, 23), (0, 0, 79, 47)),
    (1.0, 3.0, (2.0, 1.5), (2.0, -0.25), (3, 2, 50, 20), (-4, -6, 60, 70)),
    (0.37, 1.0, (5.0, 0.0), (11.5, 0.0), (-10, 0, 90, 12), (0, -3, 44, 20)),
    (1.7, 0.6, (10.0, 10.0), (20.25, 15.5), (0, 0, 63, 63), (-20, -20, 100, 60)),
    (0.25, 3.5, (0.0, 0.0), (0.0, 0.0), (8, 4, 71, 19), (0, 0, 30, 80)),
]


def main():
    L = O.reference()
    out = {}

    # transfer tables + ramps, hashed (FNV-1a 32) and sampled
    halves = np.arange(65536, dtype=np.uint16)
    for name in ("video_transfer_rec709_to_linear_scene", "video_transfer_rec709_to_linear_display",
                 "video_transfer_linear_to_rec709", "video_transfer_linear_to_sRGB"):
        table = np.zeros(65536, np.uint16)
        getattr(L, name)(O.ptr(table), O.ptr(halves), 65536)
        out[name + "_fnv"] = np.uint32(O.fnv1a(table))
        out[name + "_every_97th"] = table[::97].copy()
    ramp = np.ctypeslib.as_array(L.video_get_gamma45_ramp(), shape=(65536,)).copy()
    out["gamma45_fnv"] = np.uint32(O.fnv1a(ramp))
    out["gamma45_every_97th"] = ramp[::97].copy()

    # half conversion: structured + random bit patterns
    rng = np.random.default_rng(2024)
    bits = np.concatenate([rng.integers(0, 2 ** 32, size=20000, dtype=np.uint64).astype(np.uint32),
                           np.array([0, 0x80000000, 0x7F800000, 0x7F800001, 0x7FC00000, 0x477FE000, 0x477FFFFF, 0x47800000,
                                     0x33000000, 0x32FFFFFF, 0x38800000, 0x387FFFFF, 0x3F801FFF], np.uint32)])
    out["f2h_in_bits"] = bits
    out["f2h_out"] = O.ref_half_from_float(L, bits.view(np.float32))
    out["h2f_out_bits"] = O.ref_half_to_float(L, halves).view(np.uint32)[::13].copy()

    # DV reconstruction of a deterministic picture, cropped window
    pic = O.synth_coded(720, 480, 5, 4, O.LAYOUT_PLANAR, O.MATRIX_709, (0, -1), 0, True)
    image = O.RefCodedImage()
    for i, p in enumerate(pic.planes):
        image.data[i], image.stride[i], image.line_count[i] = p.ctypes.data, p.shape[1], p.shape[0]
    frame = O.HostFrame16((600, 400, 719, 478))
    L.video_reconstruct_dv(C.byref(frame.c), C.byref(image))
    out["dv_reconstruct_window"] = np.array(frame.current.tuple(), np.int32)
    out["dv_reconstruct_pixels"] = frame.array.copy()
    whole = O.HostFrame16((0, -1, 719, 478))
    L.video_reconstruct_dv(C.byref(whole.c), C.byref(image))
    out["dv_reconstruct_full_fnv"] = np.uint32(O.fnv1a(whole.array))

    # DV subsample of a deterministic frame
    src = O.synth_frame((0, -1, 719, 478), 6)
    planar = L.video_subsample_dv(C.byref(src.c))
    for i, (w, h) in enumerate([(720, 480), (180, 480), (180, 480)]):
        plane = np.ctypeslib.as_array(C.cast(planar.contents.data[i], C.POINTER(C.c_uint8)), shape=(h, w))
        out["dv_subsample_plane%d_fnv" % i] = np.uint32(O.fnv1a(plane))
        out["dv_subsample_plane%d_rows" % i] = plane[100:104].copy()
    out["dv_subsample_frame_after_fnv"] = np.uint32(O.fnv1a(src.array))

    # f32 mixes on bug-neutral windows
    def f32(seed, full, cw):
        h = O.synth_frame(full, seed, cw)
        f = O.HostFrame32(full, cw)
        f.array[...] = O.ref_half_to_float(L, h.array)
        return f

    full = (0, 0, 63, 35)
    for tag, wa, wb in [("full", full, full), ("inset", full, (8, 6, 40, 30))]:
        for mix in (1.0, 0.375):
            o, b = f32(1, full, wa), f32(2, full, wb)
            L.video_mix_over_f32(C.byref(o.c), C.byref(b.c), mix)
            out["over_%s_%s" % (tag, mix)] = o.array.view(np.uint32).copy()
            a, b, r = f32(3, full, wa), f32(4, full, wb), O.HostFrame32(full)
            L.video_mix_cross_f32(C.byref(r.c), C.byref(a.c), C.byref(b.c), mix)
            if tag == "full":
                out["cross_%s_%s" % (tag, mix)] = r.array.view(np.uint32).copy()

    # a 4-layer workspace pulled as f16 (workspace.c + main.c)
    layers = [O.synth_frame(full, 20 + i, cw) for i, cw in enumerate([full, (8, 6, 40, 30), (8, 2, 63, 20), (8, 10, 50, 35)])]
    arr = (O.Frame16 * 4)(*[l.c for l in layers])
    z = (C.c_int64 * 4)(0, 1, 2, 3)
    res = O.HostFrame16(full)
    L.ref_workspace_composite_f16(arr, z, 4, C.byref(res.c))
    out["workspace_window"] = np.array(res.current.tuple(), np.int32)
    out["workspace_pixels"] = res.array.copy()

    # the scaler (video_scale.c:231-285): window + FNV of the whole target for a fixed list of cases
    for i, (fx, fy, tp, sp, swin, tfull) in enumerate(SCALE_GOLDEN):
        sfull = (swin[0] - 3, swin[1] - 2, swin[2] + 4, swin[3] + 1)
        data = np.random.default_rng(700 + i).uniform(-0.5, 1.5, size=(sfull[3] - sfull[1] + 1, sfull[2] - sfull[0] + 1, 4)).astype(np.float32)
        src = O.HostFrame32(sfull, swin, data)
        dst = O.HostFrame32(tfull)
        L.video_scale_bilinear_f32(C.byref(dst.c), O.V2f(*tp), C.byref(src.c), O.V2f(*sp), O.V2f(fx, fy))
        out["scale_%d_window" % i] = np.array(dst.current.tuple(), np.int64)
        out["scale_%d_fnv" % i] = np.uint32(O.fnv1a(dst.array))
        out["scale_%d_row" % i] = dst.array[dst.array.shape[0] // 2].view(np.uint32).copy()

    # color.c:100-160 on a seeded frame with a partial window
    for which in ("rgb_to_xyz_sdtv", "xyz_to_srgb"):
        frame = O.synth_frame((-4, -2, 91, 53), 31, (0, 0, 80, 47))
        getattr(L, "video_color_" + which)(C.byref(frame.c))
        out["color_%s_fnv" % which] = np.uint32(O.fnv1a(frame.array))
        out["color_%s_row" % which] = frame.array[20].copy()

    np.savez_compressed(os.path.join(HERE, "reference_vectors.npz"), **out)
    print("wrote %d arrays" % len(out))


if __name__ == "__main__":
    main()
