/*
 * oracle/oracle.h -- TEST INFRASTRUCTURE ONLY.
 *
 * Plain-C CPU restatement of the per-pixel frame pipeline of fluggo/Canvas
 * (src/cprocess).  It exists to CHECK the CUDA product path; nothing under
 * canvas_b200/ may include, link or call it.  Only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs use it.
 *
 * Parity pin: every function here is validated bit-for-bit against the reference's
 * own C code compiled unmodified (oracle/_ref/libcanvasref.so, `make ref`) by
 * tests/test_oracle_vs_reference.py in the build container, and against the golden
 * vectors under tests/golden/ (generated from that library by
 * tests/golden/make_golden.py) everywhere else.
 *
 * Where the reference only has a hard-coded case (720x480 planar 4:1:1 Rec.709,
 * video_reconstruct.c:52,79) the functions here take the size / subsampling /
 * matrix / picture offset as parameters and reproduce the reference exactly at its
 * own parameter values.
 *
 * Window semantics of mix/over are the INTENDED ones (see DESIGN.md "reference
 * bugs"): inside the union box a pixel covered by neither input is zero and a pixel
 * covered by one input takes that input, whatever the relation of the windows'
 * min.x / min.y (the reference compares min.x with min.y at video_mix.c:137,265 and
 * leaves stale pixels for disjoint windows, :161-169).
 */
#ifndef CANVAS_ORACLE_H
#define CANVAS_ORACLE_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Same layout as the reference's box2i / rgba_frame_* (include/framework.h:46-71,155-183). */
typedef struct { int32_t min_x, min_y, max_x, max_y; } orc_box;
typedef struct { uint16_t r, g, b, a; } orc_px16;
typedef struct { float r, g, b, a; } orc_px32;
typedef struct { uint8_t r, g, b, a; } orc_px8;
typedef struct { orc_px16 *data; orc_box full, current; } orc_frame16;
typedef struct { orc_px32 *data; orc_box full, current; } orc_frame32;

enum { ORC_LUT_709_TO_LINEAR_SCENE = 0, ORC_LUT_709_TO_LINEAR_DISPLAY = 1,
       ORC_LUT_LINEAR_TO_709 = 2, ORC_LUT_LINEAR_TO_SRGB = 3 };
enum { ORC_LAYOUT_PLANAR = 0, ORC_LAYOUT_UYVY = 1, ORC_LAYOUT_YUY2 = 2 };
enum { ORC_MATRIX_709 = 0, ORC_MATRIX_601 = 1 };

/* Coded picture description (generalises coded_image, framework.h:470-476). */
typedef struct {
    uint8_t *plane[3];      /* planar: Y, Cb, Cr; packed: plane[0] only */
    int32_t stride[3];
    int32_t layout;         /* ORC_LAYOUT_* */
    int32_t width, height;  /* luma size */
    int32_t sub_x;          /* horizontal chroma subsampling: 1, 2 or 4 (co-sited left) */
    int32_t matrix;         /* ORC_MATRIX_* */
    int32_t pic_x, pic_y;   /* picOffset: frame coordinate of luma sample (0,0) */
} orc_coded;

/* half.c:31-51 + genhalf.py (table semantics restated as bit operations) */
uint16_t orc_f2h( float value );
float orc_h2f( uint16_t value );
void orc_f2h_n( uint16_t *out, const float *in, size_t count );
void orc_h2f_n( float *out, const uint16_t *in, size_t count );

/* gammatab.c:8-38, 48-66, 83-250; widget_gl.c:949-969 */
void orc_build_lut( int which, uint16_t *table /*[65536]*/ );
void orc_build_gamma45( uint8_t *ramp /*[65536]*/ );
void orc_build_intent_ramp( float rendering_intent, uint8_t *ramp /*[65536]*/ );

/* filter.c:24-76.  Returns 0, or -1 if cap is too small (width then holds the need). */
int orc_triangle( float sub, float offset, float *coeff, int cap, int *width, int *center );

/* video_reconstruct.c:50-137 (parameterised) */
void orc_reconstruct( orc_frame16 *frame, const orc_coded *image );
/* video_subsample.c:99-187 (parameterised; planes must be zero-initialised by the
 * caller like coded_image_alloc0; mutates frame in place exactly like :147) */
void orc_subsample( orc_frame16 *frame, const orc_coded *image );

/* video_subsample.c:212-290 (GLSL in the reference -- PARITY UNPINNED: no CPU twin, no GL here): the MPEG-2 4:2:0
 * shaders restated in C.  Choices the shader leaves to the GL implementation are fixed as: pow() = libm powf on the
 * exact value of the half texel, dot products summed left to right, sin(y*3.14156) taken as its intended +-1,
 * texels outside current_window are zero, coordinates clamp to the picture edge, a normalised byte is
 * lrintf(clamp(v,0,1)*255) (NaN -> 0).  Planes: Y width x height, Cb/Cr width/2 x height/2.  frame is not modified. */
void orc_subsample_mpeg2( const orc_frame16 *frame, const orc_coded *image );

/* color.c:100-131, :134-160, in place over current_window */
void orc_color_rgb_to_xyz_sdtv( orc_frame16 *frame );
void orc_color_xyz_to_srgb( orc_frame16 *frame );

/* video_filter.c:79-95 (GLSL only, parity unpinned): pixels of current_window outside rect become zero */
void orc_crop_f16( orc_frame16 *frame, const orc_box *rect );

/* video_mix.c:27-44, 73-105, 107-235, 237-370 */
void orc_copy_frame_f16( orc_frame16 *out, const orc_frame16 *in );
void orc_copy_frame_alpha_f32( orc_frame32 *out, const orc_frame32 *in, float alpha );
void orc_mix_cross_f32( orc_frame32 *out, const orc_frame32 *a, const orc_frame32 *b, float mix_b );
void orc_mix_over_f32( orc_frame32 *out, const orc_frame32 *b, float mix_b );

/* main.c:34-77 / 106-145 conversions between precisions over current_window */
/* video_scale.c:231-285 (target_point, source_point, factors passed as scalars) */
void orc_scale_bilinear_f32( orc_frame32 *target, float tpx, float tpy, const orc_frame32 *source, float spx, float spy, float fx, float fy );
void orc_frame_f16_to_f32( orc_frame32 *out, const orc_frame16 *in );
void orc_frame_f32_to_f16( orc_frame16 *out, const orc_frame32 *in );

/* workspace.c:494-550 pulled through main.c:34-77: layers bottom->top, each copied into
 * out->full (video_copy_frame_f16), converted to f32, composited with mix 1, packed. */
void orc_composite_f16( orc_frame16 *out, const orc_frame16 *layers, int count );

/* src/process/Pulldown23RemovalFilter.c:82-104: the even rows of frame's current window come from `even_rows`
 * (intended form: copied from the window's left edge; the reference starts at the temporary's x = 0, equal for min.x == 0) */
void orc_weave_fields_f16( orc_frame16 *frame, const orc_frame16 *even_rows );

/* The compositor with a filter chain in front of each layer -- weave, colour transform (f16), gain/offset (f32, GLSL
 * restated: PARITY UNPINNED for that step), crop (zeros; GLSL restated) -- each applied as its own pass, then
 * workspace.c:530-547.  ops may be NULL; ops[i].flags == 0 leaves layer i alone. */
enum { ORC_OP_WEAVE = 1, ORC_OP_COLOR = 2, ORC_OP_GAIN_OFFSET = 4, ORC_OP_CROP = 8 };
enum { ORC_COLOR_RGB_TO_XYZ_SDTV = 1, ORC_COLOR_XYZ_TO_SRGB = 2 };
typedef struct { int32_t flags, color; float gain, offset; orc_box crop; orc_frame16 even_rows; } orc_layer_ops;
void orc_composite_ops_f16( orc_frame16 *out, const orc_frame16 *layers, const orc_layer_ops *ops, int count );

/* VideoMixFilter path: video_mix.c:46-71 pulled as f16 (main.c:44-72). */
void orc_cross_f16( orc_frame16 *out, const orc_frame16 *a, const orc_frame16 *b, float mix_b );

/* Timeline frame (BASELINE.json configs[4]): bottom = crossfade of two converted coded
 * pictures (either may be NULL -> other one only / none), then overlays over it. */
void orc_timeline_frame_f16( orc_frame16 *out, const orc_coded *a, const orc_coded *b, float mix_b,
    const orc_frame16 *overlays, int overlay_count );

/* widget_gl.c:290-307: sRGB transfer in place on current window, then ramp per channel */
void orc_encode_u8( orc_frame16 *frame, const uint16_t *lut_srgb, const uint8_t *ramp, orc_px8 *target );
/* RgbaFrameF16.c:114-149: premultiplied ARGB32 of the current window, tightly packed */
void orc_to_argb32( const orc_frame16 *frame, const uint8_t *ramp45, uint32_t *out );

/* Deterministic synthetic inputs (same integer hash as the device generator). */
uint32_t orc_hash32( uint32_t seed, uint32_t index );
void orc_fill_frame( orc_frame16 *frame, uint32_t seed );    /* fills full window, sets nothing else */
void orc_fill_bytes( uint8_t *data, size_t count, uint32_t seed, int lo, int hi );

uint32_t orc_fnv1a( const void *data, size_t count );

#ifdef __cplusplus
}
#endif

#endif
