/*
 * include/canvas_b200.h -- the thin C-ABI over the sm_100a kernels (libcanvas_b200.so).
 *
 * Plain C: pointers, sizes and PODs only.  No C++/torch types cross this boundary.
 * A frame that lives in device memory is described by the SAME struct the
 * reference uses on the host (rgba_frame_f16, include/framework.h) -- only the
 * data pointer is a device pointer obtained from cv_frame_alloc_f16().
 *
 * Every function returns 0 or a negative cv_err.  Functions that produce a frame
 * ALSO follow the reference's error convention: on failure the frame's
 * current_window is set empty (reference src/cprocess/main.c:36-39).
 * current_window of an output is pure box arithmetic and is written on the host
 * before the call returns; pixels are valid once the stream has been synchronised.
 *
 * There is no CPU fallback anywhere behind this header: if the CUDA runtime or a
 * B200 is missing, cv_init() fails and every other call returns CV_ERR_NOT_INITIALISED.
 *
 * Reference interface each entry replaces (paths relative to /root/reference):
 *   cv_half_*            src/cprocess/half.c:31-51,82-85 (half_convert_*, half_lookup)
 *   cv_transfer          src/cprocess/gammatab.c:83-250 (video_transfer_*)
 *   cv_convert_ycbcr8    src/cprocess/video_reconstruct.c:50-137 (video_reconstruct_dv)
 *   cv_subsample_ycbcr8  src/cprocess/video_subsample.c:99-187 (video_subsample_dv)
 *   cv_subsample_mpeg2   src/cprocess/video_subsample.c:212-290 (video_subsample_mpeg2_gl's shaders)
 *   cv_color_*           src/cprocess/color.c:100-160 (video_color_rgb_to_xyz_sdtv, video_color_xyz_to_srgb)
 *   cv_copy_frame_f16    src/cprocess/video_mix.c:27-44
 *   cv_mix_cross_f16     src/cprocess/video_mix.c:46-71,107-235 pulled as f16 (main.c:44-72)
 *   cv_composite_over_f16 src/cprocess/workspace.c:494-550 + video_mix.c:237-370
 *   cv_mix_*_f32         src/cprocess/video_mix.c:73-105,107-235,237-370 on f32 frames
 *   cv_encode_u8         src/cprocess/widget_gl.c:290-307,949-969
 *   cv_to_argb32         src/process/RgbaFrameF16.c:114-149
 *   cv_render_frames     the fused form of DVReconstructionFilter -> VideoMixFilter ->
 *                        VideoWorkspace pulled through video_get_frame_f16
 *   cv_timeline_*        src/cprocess/workspace.c:243-307,494-550 as a stateless,
 *                        frame-parallel multi-GPU scheduler
 */
#ifndef CANVAS_B200_H
#define CANVAS_B200_H

#include "framework.h"

#if defined(__cplusplus)
extern "C" {
#endif

typedef enum {
    CV_OK = 0,
    CV_ERR_NOT_INITIALISED = -1,    /* cv_init not called, or it failed */
    CV_ERR_CUDA = -2,               /* a CUDA runtime call failed; see cv_last_error() */
    CV_ERR_ARGUMENT = -3,
    CV_ERR_MEMORY = -4,
    CV_ERR_UNSUPPORTED = -5,        /* e.g. sub_x not in {1,2,4}, packed layout with sub_x != 2 */
    CV_ERR_NO_DEVICE = -6
} cv_err;

typedef void *cv_stream;            /* a cudaStream_t; NULL = the library's default stream of the current device */
typedef struct cv_event_s *cv_event;

enum { CV_MEM_DEVICE = 0, CV_MEM_PINNED = 1 };

enum { CV_LUT_709_TO_LINEAR_SCENE = 0, CV_LUT_709_TO_LINEAR_DISPLAY = 1,
       CV_LUT_LINEAR_TO_709 = 2, CV_LUT_LINEAR_TO_SRGB = 3, CV_LUT_COUNT = 4 };
enum { CV_LAYOUT_PLANAR = 0, CV_LAYOUT_UYVY = 1, CV_LAYOUT_YUY2 = 2 };
enum { CV_MATRIX_709 = 0, CV_MATRIX_601 = 1 };

enum { CV_MAX_LAYERS = 16 };        /* over-layers fused into one launch; more are chained through an f32 frame */

/* ------------------------------------------------------------------ runtime */

int cv_device_count( void );
/* Initialises `device` (creates its default stream, builds the transfer tables on the
 * host with libm powf exactly like gammatab.c:83-250 and uploads them).  Idempotent
 * per device; makes `device` current for the calling thread. */
int cv_init( int device );
int cv_set_device( int device );
int cv_get_device( void );
void cv_shutdown( void );
const char *cv_last_error( void );          /* thread-local message of the last failure */
const char *cv_version( void );
/* Number of kernels this library has launched on any device since load (all threads). */
uint64_t cv_kernel_launches( void );

int cv_stream_create( cv_stream *stream );
int cv_stream_destroy( cv_stream stream );
int cv_stream_sync( cv_stream stream );
int cv_device_sync( void );

int cv_event_create( cv_event *event );
int cv_event_destroy( cv_event event );
int cv_event_record( cv_event event, cv_stream stream );
int cv_event_sync( cv_event event );
int cv_event_elapsed_ms( cv_event start, cv_event stop, float *ms );

/* ------------------------------------------------------------------- memory */

int cv_malloc( void **ptr, size_t bytes, int where /* CV_MEM_* */ );
/* cv_free returns a device block to the pool at once: the caller must have synchronised every stream that used it.
 * cv_free_async orders the free behind the work already queued on `stream` (the block is not handed out again before
 * that work has finished) -- what a thread that renders on its own stream uses. */
int cv_free( void *ptr, int where );
int cv_free_async( void *ptr, int where, cv_stream stream );
int cv_memset( void *dev, int byte, size_t bytes, cv_stream stream );
int cv_upload( void *dev, const void *host, size_t bytes, cv_stream stream );      /* async when host is pinned */
int cv_download( void *host, const void *dev, size_t bytes, cv_stream stream );
int cv_copy_d2d( void *dst, const void *src, size_t bytes, cv_stream stream );
/* Writes `bytes` of scratch (> L2) so the next timed launch starts from a cold L2. */
int cv_flush_l2( cv_stream stream );

/* The half-float frame allocator (replaces PyMem_Malloc in RgbaFrameF16.c:205 and the
 * g_slice temporaries of main.c:50, video_mix.c:62, workspace.c:538): 256-byte aligned,
 * pooled per device so steady-state rendering never calls cudaMalloc. */
int cv_frame_alloc_f16( rgba_frame_f16 *frame, const box2i *full_window, int where );
int cv_frame_free_f16( rgba_frame_f16 *frame, int where );
int cv_frame_alloc_f32( rgba_frame_f32 *frame, const box2i *full_window, int where );
int cv_frame_free_f32( rgba_frame_f32 *frame, int where );
/* Copies the pixels of full_window (same full_window on both sides) and current_window. */
int cv_frame_upload_f16( rgba_frame_f16 *dev, const rgba_frame_f16 *host, cv_stream stream );
int cv_frame_download_f16( rgba_frame_f16 *host, const rgba_frame_f16 *dev, cv_stream stream );
int cv_frame_upload_f32( rgba_frame_f32 *dev, const rgba_frame_f32 *host, cv_stream stream );
int cv_frame_download_f32( rgba_frame_f32 *host, const rgba_frame_f32 *dev, cv_stream stream );

/* Host copies of the tables the kernels gather from (built once per process). */
const half *cv_host_lut( int which /* CV_LUT_* */ );
const uint8_t *cv_host_gamma45_ramp( void );
/* ramp[i] = lrint(clamp(powf(h2f(i), intent) * 255, 0, 255)), widget_gl.c:949-969. */
int cv_build_intent_ramp( float rendering_intent, uint8_t *ramp /* [65536] host */ );

/* ------------------------------------------------ half pack / unpack, tables */

int cv_half_from_float( half *out, const float *in, size_t count, cv_stream stream );      /* device pointers */
int cv_half_to_float( float *out, const half *in, size_t count, cv_stream stream );
int cv_half_lookup( const half *table /* device, [65536] */, half *out, const half *in, size_t count, cv_stream stream );
int cv_transfer( int which /* CV_LUT_* */, half *out, const half *in, size_t count, cv_stream stream );

/* ---------------------------------------------------------- pixel operations */

/*
 * An 8-bit Y'CbCr picture in device memory (generalises coded_image,
 * framework.h:470-476, whose only reference consumer hard-codes 720x480 planar 4:1:1
 * Rec.709 at picture offset (0,-1), video_reconstruct.c:52,79).  Chroma is subsampled
 * horizontally by sub_x and co-sited with the left luma sample.  Packed layouts are
 * 4:2:2 by definition: UYVY = Cb Y0 Cr Y1, YUY2 = Y0 Cb Y1 Cr, plane[0] only.
 */
typedef struct {
    uint8_t *plane[3];          /* planar: Y, Cb, Cr */
    int32_t stride[3];          /* bytes per row */
    int32_t layout;             /* CV_LAYOUT_* */
    int32_t width, height;      /* luma samples */
    int32_t sub_x;              /* 1, 2 or 4 */
    int32_t matrix;             /* CV_MATRIX_* */
    int32_t pic_x, pic_y;       /* frame coordinate of luma sample (0,0) */
} cv_coded;

/* Y'CbCr -> linear f16 RGBA: studio-range decode, triangle chroma reconstruction, 3x3
 * matrix in f32, truncating half pack, rec709->linear-scene table on R, G, B and A.
 * frame->current_window = picture box intersected with frame->full_window. */
int cv_convert_ycbcr8( rgba_frame_f16 *frame, const cv_coded *image, cv_stream stream );

/* Linear f16 RGBA -> Y'CbCr (planes must be zeroed by the caller, like coded_image_alloc0).
 * Like the reference (video_subsample.c:147) the linear->Rec.709 table is applied IN PLACE
 * to all four channels of the window of `frame` that the picture covers. */
int cv_subsample_ycbcr8( rgba_frame_f16 *frame, const cv_coded *image, cv_stream stream );

int cv_copy_frame_f16( rgba_frame_f16 *out, const rgba_frame_f16 *in, cv_stream stream );

/* VideoMixFilter pulled as f16: mix_b is clamped to [0,1]; 0 and 1 fetch one source only. */
int cv_mix_cross_f16( rgba_frame_f16 *out, const rgba_frame_f16 *a, const rgba_frame_f16 *b, float mix_b, cv_stream stream );

/* One stacked layer: `frame.current_window` places it, `mix` scales its alpha (the
 * workspace uses 1).  frame.data may be NULL only when current_window is empty. */
/*
 * Per-pixel filters FUSED into a layer: what the editor chains in front of a workspace item (all GL-only in the
 * reference except the colour transforms and the field weave) is evaluated in registers on the layer's pixels as they
 * are read, before "over" -- no intermediate frame, no extra HBM traffic.  Applied in this order:
 *   CV_OP_WEAVE        the even rows come from `even_rows` instead (Pulldown23RemovalFilter.c:82-104: field 1 of the next
 *                      frame); even_rows.current_window must cover the layer's window, pixels are untouched bits;
 *   CV_OP_COLOR        video_color_rgb_to_xyz_sdtv / video_color_xyz_to_srgb (color.c:100-160): [table,] exact widening,
 *                      3x3 matrix in f32, truncating pack to half [, table] -- the same roundings as the separate pass;
 *   CV_OP_GAIN_OFFSET  rgb * gain + offset in f32, alpha untouched (video_filter.c:27-39); the result stays f32;
 *   CV_OP_CROP         pixels of the layer outside `crop` become (0, 0, 0, 0), its window stays (video_filter.c:79-95).
 * The result is bit-identical to running cv_weave_fields_f16 / cv_color_* / cv_gain_offset_f32 / cv_crop_f16 as separate
 * passes and stacking the outcome (tests/test_gpu_layer_ops.py).
 */
enum { CV_OP_WEAVE = 1, CV_OP_COLOR = 2, CV_OP_GAIN_OFFSET = 4, CV_OP_CROP = 8 };
enum { CV_COLOR_RGB_TO_XYZ_SDTV = 1, CV_COLOR_XYZ_TO_SRGB = 2 };

typedef struct {
    int32_t flags;              /* CV_OP_* */
    int32_t color;              /* CV_COLOR_* */
    float gain, offset;
    box2i crop;
    rgba_frame_f16 even_rows;   /* device frame, pair-aligned like the layer's */
} cv_layer_ops;

typedef struct {
    rgba_frame_f16 frame;
    float mix;
    int32_t reserved;
    const cv_layer_ops *ops;    /* NULL: the layer as it is */
} cv_layer;

/* Workspace compositor: layers[0] is fetched into out, each further layer goes "over" it,
 * all in f32 registers, one truncating pack at the end.  Any count >= 0; count == 0 or all
 * layers empty gives an empty current_window.  More than CV_MAX_LAYERS layers are chained
 * through a pooled f32 frame (still one rounding, at the end). */
int cv_composite_over_f16( rgba_frame_f16 *out, const cv_layer *layers, int count, cv_stream stream );

/* The f32 frame operations of video_mix.c on device frames (in-place on out where the
 * reference is in-place). */
int cv_copy_frame_alpha_f32( rgba_frame_f32 *out, const rgba_frame_f32 *in, float alpha, cv_stream stream );
int cv_mix_cross_f32( rgba_frame_f32 *out, const rgba_frame_f32 *a, const rgba_frame_f32 *b, float mix_b, cv_stream stream );
int cv_mix_over_f32( rgba_frame_f32 *out, const rgba_frame_f32 *b, float mix_b, cv_stream stream );
/* main.c:56-70 / 126-139: precision hops over in->current_window. */
int cv_frame_f16_to_f32( rgba_frame_f32 *out, const rgba_frame_f16 *in, cv_stream stream );
int cv_frame_f32_to_f16( rgba_frame_f16 *out, const rgba_frame_f32 *in, cv_stream stream );

/* SolidColorVideoSource (src/process/SolidColorVideoSource.c:52-103): current_window = window
 * intersected with full_window, every pixel of it = color (f16: truncated to half once). */
int cv_solid_frame_f16( rgba_frame_f16 *frame, const box2i *window, const float color[4], cv_stream stream );
int cv_solid_frame_f32( rgba_frame_f32 *frame, const box2i *window, const float color[4], cv_stream stream );

/* video_scale_bilinear_f32 (src/cprocess/video_scale.c:231-285) on device frames: separable triangle-filter
 * resampling that maps source_point onto target_point with the given factors (> 1 enlarges).  The whole
 * target full_window is written (zero where no tap reaches); current_window as the reference computes it. */
int cv_scale_bilinear_f32( rgba_frame_f32 *target, v2f target_point, const rgba_frame_f32 *source, v2f source_point, v2f factors, cv_stream stream );

/* cv_scale_bilinear_f32 for callers that may not allocate (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * Same arguments, same result bits and same current_window as cv_scale_bilinear_f32, but the call allocates nothing
 * on the device or the host: the tap tables and the intermediate frame between the two passes live in one static
 * device array of CV_STATIC_SCRATCH_BYTES that every GPU loading the library holds until exit (cv_static.cu).
 * Limits, rejected with CV_ERR_UNSUPPORTED before anything is launched: an intermediate frame of at most
 * CV_STATIC_MAX_TEMP_PIXELS pixels (1920x1080 fits), at most CV_STATIC_MAX_LINES target lines and
 * CV_STATIC_MAX_TAPS filter taps per axis.  All calls on a device share the array, which keeps the last call's
 * contents (everything read from it is written first by the same call, nothing needs clearing); calls that overlap
 * on the GPU -- two streams, two threads -- RACE: issue one at a time, stream-ordered after the previous one. */
#define CV_STATIC_SCRATCH_BYTES     37880320u           /* 32 MiB intermediate frame + 2 x (start | src | coef) tap tables */
#define CV_STATIC_MAX_TEMP_PIXELS   (2u << 20)
#define CV_STATIC_MAX_LINES         16384u
#define CV_STATIC_MAX_TAPS          (256u << 10)
int cv_scale_bilinear_static_f32( rgba_frame_f32 *target, v2f target_point, const rgba_frame_f32 *source, v2f source_point, v2f factors, cv_stream stream );

/* video_filter_gain_offset_gl (src/cprocess/video_filter.c:27-72; a GLSL shader in the reference, no CPU twin):
 * rgb = rgb * gain + offset over current_window, alpha untouched, in place. */
int cv_gain_offset_f32( rgba_frame_f32 *frame, float gain, float offset, cv_stream stream );
/* The field weave of Pulldown23RemovalFilter (src/process/Pulldown23RemovalFilter.c:83-101): the even rows of
 * frame's current window are taken from `even_rows`. */
int cv_weave_fields_f16( rgba_frame_f16 *frame, const rgba_frame_f16 *even_rows, cv_stream stream );

/* video_color_rgb_to_xyz_sdtv / video_color_xyz_to_srgb (src/cprocess/color.c:100-131, :134-160), in place over
 * current_window: [rec709->linear-scene table on all four channels,] 3x3 matrix in f32 on exact-widened halves,
 * truncating pack [, linear->sRGB table on all four channels]. */
int cv_color_rgb_to_xyz_sdtv( rgba_frame_f16 *frame, cv_stream stream );
int cv_color_xyz_to_srgb( rgba_frame_f16 *frame, cv_stream stream );
/* video_filter_crop_gl (src/cprocess/video_filter.c:79-95; GLSL only in the reference): pixels of current_window
 * outside crop_rect become (0,0,0,0); the window itself is unchanged. */
int cv_crop_f16( rgba_frame_f16 *frame, const box2i *crop_rect, cv_stream stream );
/* video_subsample_mpeg2_gl (src/cprocess/video_subsample.c:212-290,314-520; GLSL only in the reference): linear f16
 * RGBA -> planar 8-bit 4:2:0 (Y width x height, Cb/Cr width/2 x height/2), interlace-aware 6-tap chroma, the frame
 * is NOT modified.  image: CV_LAYOUT_PLANAR, even width/height, pic_x/pic_y = frame coordinate of luma (0,0);
 * sub_x is ignored.  The reference call is width 720, height 480, CV_MATRIX_601, picture at (0,0). */
int cv_subsample_mpeg2( const rgba_frame_f16 *frame, const cv_coded *image, cv_stream stream );

/* Display encode: target[px] = ramp[ sRGB[ frame[px] ] ] per channel (alpha too) over
 * frame->current_window; target is full_window-sized.  `u8_table` is a device copy of the
 * COMPOSED table from cv_upload_display_table().  With write_back != 0 the sRGB-encoded
 * halves are also stored back into `frame`, as the widget does to its private copy. */
int cv_upload_display_table( float rendering_intent, uint8_t **table /* device, [65536] */ );
int cv_encode_u8( rgba_frame_f16 *frame, const uint8_t *u8_table, rgba_u8 *target, int write_back, cv_stream stream );
/* Premultiplied ARGB32 of current_window, tightly packed (RgbaFrameF16.to_argb32_bytes). */
int cv_to_argb32( const rgba_frame_f16 *frame, uint32_t *out, cv_stream stream );

/* ------------------------------------------------------- fused frame programs */

enum { CV_SRC_NONE = 0, CV_SRC_F16 = 1, CV_SRC_CODED = 2, CV_SRC_F32 = 3 };

typedef struct {
    int32_t kind;               /* CV_SRC_* */
    int32_t reserved;
    rgba_frame_f16 f16;         /* CV_SRC_F16 */
    rgba_frame_f32 f32;         /* CV_SRC_F32 */
    cv_coded coded;             /* CV_SRC_CODED: converted in-kernel exactly like cv_convert_ycbcr8 */
} cv_source;

/*
 * One output frame of a filter graph of the shape the editor builds
 * (fluggo/editor/graph/video.py:84-159): an optional bottom -- source `a` alone, or the
 * crossfade of `a` and `b` -- under `layer_count` over-layers.  Everything is evaluated in
 * one kernel: each input pixel is read once, the output written once, with the exact
 * intermediate roundings of the unfused reference chain (f16 pack + table after a
 * conversion, f32 everywhere else, one truncating pack at the end).
 * Outputs: out (f16), and/or out_f32, and/or out_u8 through u8_table.
 */
typedef struct {
    cv_source a, b;
    float mix_b;
    int32_t layer_count;
    cv_layer layers[CV_MAX_LAYERS];
    rgba_frame_f16 *out;        /* may be NULL when only out_f32 / out_u8 is wanted (then `window` carries the geometry) */
    rgba_frame_f32 *out_f32;
    rgba_u8 *out_u8;            /* device, full_window-sized, or NULL */
    const uint8_t *u8_table;    /* device composed display table, needed with out_u8 */
    box2i full_window;          /* used when out == NULL && out_f32 == NULL */
    box2i current_window;       /* result window, written by cv_render_frames */
} cv_frame_job;

/* Renders `count` independent jobs with as few launches as their shapes allow
 * (jobs of one shape share a launch: grid.y = job). */
int cv_render_frames( cv_frame_job *jobs, int count, cv_stream stream );

/* ------------------------------------------- frame-parallel multi-GPU timeline */

/*
 * The reference renders a timeline one frame at a time on one thread
 * (workspace_get_frame_f32, src/cprocess/workspace.c:494-550).  A frame depends only on its
 * index, so here a range of frames is partitioned over the GPUs of one box with no collective:
 * chunks of consecutive frames go to whichever GPU is free next (or, with CV_TIMELINE_STATIC, to
 * gpu slot = ((frame - first) / chunk) mod n_gpus), one host thread, a compute stream and a copy stream per GPU.
 *
 * A clip is a device-resident ring of sources (one ring PER GPU: frames are produced where their
 * inputs live); source frame i of a clip is frames[i mod count].
 */
typedef struct {
    const cv_source *frames;
    int32_t count;
    int32_t reserved;
} cv_clip;

/* One workspace item (workspace_add_item, framework.h:573; workspace.c:309-345): shown at frames
 * x <= f < x + length, stacked by ascending z, reading its clip at f - x + offset.  With clips_b the
 * item is the crossfade VideoMixFilter(clips, clips_b, mix_b) where mix_b ramps linearly from
 * mix_first at source frame ramp_start to mix_last at ramp_start + ramp_length (a two-point linear
 * AnimationFunc, src/process/AnimationFunc.c:439-446: interpolated in double, then (float)).
 * Only the bottom item showing at a frame may be a crossfade or a coded / f32 source. */
typedef struct {
    int64_t x, length, z, offset;
    const cv_clip *clips;           /* [n_gpus] */
    const cv_clip *clips_b;         /* [n_gpus] or NULL */
    float mix_first, mix_last;
    int64_t ramp_start, ramp_length;
    const struct cv_pull_s *pull;   /* when set, `clips` is ignored: the item's pictures are pulled from the host (below) */
    const struct cv_pull_s *pull_b; /* likewise for the crossfade's second source */
} cv_timeline_item;

/*
 * Pictures PULLED from the host while the timeline renders, through the reference's own source interfaces:
 *   coded   coded_image *getFrame( obj, source_frame, quality_hint 0 ) (include/framework.h:513-523; what
 *           DVReconstructionFilter_get_frame calls per frame, src/process/DVReconstructionFilter.c:53-71).  The picture is
 *           released with its own free_func once it has been copied.  `format` says what the planes hold (layout, width,
 *           height, sub_x, matrix, pic_x / pic_y; its plane pointers are ignored, strides come from the coded_image);
 *   video   a video_source pulled with video_get_frame_f16( source, source_frame, frame over full_window )
 *           (src/cprocess/workspace.c:530-547 pulls every item this way).
 * Exactly one of the two is set.  getFrame / get_frame are called on the GPU's worker thread (several GPUs call
 * concurrently, as the reference's pull queue does).  Each worker stages what it pulls through its own ring of pinned
 * buffers and a dedicated copy stream -- one cudaMemcpyAsync per plane, the kernels waiting on an event; memory that is
 * already pinned (cv_malloc( ..., CV_MEM_PINNED ), cudaHostRegister) is copied from directly.
 */
typedef struct cv_pull_s {
    coded_image_source *coded;
    cv_coded format;
    video_source *video;
    box2i full_window;
} cv_pull;

/* Indices (into items) of the items showing at `frame`, bottom first; returns how many show. */
int cv_timeline_showing( const cv_timeline_item *items, int count, int64_t frame, int *indices, int capacity );
/* Fills job->a, b, mix_b, layers, layer_count for `frame` on GPU slot `slot` (outputs are left to
 * the caller); returns the number of items showing (0 = empty frame) or a negative cv_err.  Pure
 * host logic: usable without a GPU. */
int cv_timeline_job( const cv_timeline_item *items, int count, int slot, int64_t frame, cv_frame_job *job );
/* The partition: which GPU slot renders `frame`, and how many frames of [first, last] a slot gets. */
int cv_timeline_device_of( int64_t frame, int64_t first, int chunk, int n_gpus );
int64_t cv_timeline_frames_of( int slot, int64_t first, int64_t last, int chunk, int n_gpus );

/* Per-GPU outputs: a ring of device frames (and optionally 8-bit targets) reused round-robin; a deeper ring (8-16)
 * lets more frames share a launch. */
typedef struct {
    rgba_frame_f16 *out;            /* [ring] device frames from cv_frame_alloc_f16 on that GPU */
    rgba_u8 **out_u8;               /* [ring] device buffers or NULL */
    const uint8_t *u8_table;        /* that GPU's composed display table (cv_upload_display_table) */
    int32_t ring;
    int32_t reserved;
} cv_timeline_target;

/* Called on the GPU's own host thread after frame `frame` was enqueued on `stream` (frames are enqueued in groups
 * of up to min(ring, 16): frames of one shape share one persistent launch; the sink is called once per frame when
 * its group has been enqueued); it may enqueue copies out of job->out / job->out_u8 on the same stream (the ring
 * slot is not reused before they finish).  Frames of one GPU arrive in ascending order; different GPUs call
 * concurrently. */
typedef int (*cv_timeline_sink)( void *user, int slot, int64_t frame, const cv_frame_job *job, cv_stream stream );

typedef struct {
    const cv_timeline_item *items;
    int32_t item_count;
    int32_t n_gpus;
    const int32_t *devices;         /* [n_gpus] CUDA ordinals, or NULL for 0..n_gpus-1 */
    int64_t first, last;            /* inclusive */
    int32_t chunk;                  /* consecutive frames a GPU takes at a time (8..32 keeps input rings warm) */
    int32_t flags;                  /* CV_TIMELINE_* */
    const cv_timeline_target *targets;  /* [n_gpus] */
    cv_timeline_sink sink;          /* may be NULL */
    void *user;
} cv_timeline_run;

typedef struct {
    int32_t device;
    int32_t reserved;
    int64_t frames;
    uint64_t launches;
    double device_ms;               /* CUDA events around this GPU's share, on its stream */
    double wall_ms;
    uint64_t h2d_bytes;             /* what this GPU's worker pulled from the host (cv_pull items) */
} cv_timeline_stats;

/* By default a GPU that finishes its chunk takes the next undistributed one (a shared atomic counter: no GPU is left
 * with a tail of frames while the others idle).  CV_TIMELINE_STATIC keeps the fixed block-cyclic deal of
 * cv_timeline_device_of() -- frame -> GPU is then known in advance. */
enum { CV_TIMELINE_STATIC = 1 };

/* Renders frames first..last; blocks until every GPU has finished.  stats: [n_gpus] or NULL. */
int cv_timeline_render( const cv_timeline_run *run, cv_timeline_stats *stats );

#if defined(__cplusplus)
}
#endif

#endif
