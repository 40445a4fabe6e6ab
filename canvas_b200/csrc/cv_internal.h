/*
 * canvas_b200/csrc/cv_internal.h -- shared between the runtime, the host-side job
 * builder and the kernels.  Not installed; the public surface is the headers under include/.
 */
#ifndef CV_INTERNAL_H
#define CV_INTERNAL_H

#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>
#include <pthread.h>
#include "canvas_b200.h"

#define CV_MAX_DEVICES 16

/* ---- device-side job description (what one frame program needs, pre-digested) ---- */

typedef struct DevBox { int32_t x0, y0, x1, y1; } DevBox;      /* inclusive, empty <=> x1 < x0 || y1 < y0 */

typedef struct DevSrc {
    int32_t kind;               /* CV_SRC_* */
    int32_t pitch;              /* pixels per row of the f16/f32 frame */
    int32_t ox, oy;             /* full_window.min of the f16/f32 frame */
    const void *data;           /* f16 or f32 pixels */
    const void *origin;         /* f16: data - (oy * pitch + ox) pixels */
    const void *lo, *hi;        /* f16: the 16-byte aligned byte range [lo, hi) inside the frame's allocation */
    DevBox cw;                  /* valid pixels, already clipped to the output's full_window */
    /* CV_SRC_CODED */
    const uint8_t *plane[3];
    int32_t stride[3];
    int32_t layout, width, height, sub_shift, matrix, pic_x, pic_y;
    int32_t last_chroma;        /* (width - 1) / sub_x: the last chroma sample that exists */
} DevSrc;

typedef struct DevLayer {
    const void *data;           /* f16 pixels */
    const void *origin;         /* data - (oy * pitch + ox) pixels: the address pixel (0,0) would have */
    const void *lo, *hi;        /* the 16-byte aligned byte range [lo, hi) inside the frame's allocation (bulk copies stay inside) */
    int32_t pitch, ox, oy;
    float mix;
    DevBox cw;                  /* clipped to the output's full_window; never empty (empty layers are dropped) */
    DevBox under;               /* box already defined below this layer (running union); may be empty */
    /* fused per-pixel filters (cv_layer_ops), evaluated on the layer's pixels as they are read */
    int32_t op_flags, op_color; /* CV_OP_*, CV_COLOR_* */
    float gain, offset;
    DevBox crop;
    const void *weave_origin;   /* CV_OP_WEAVE: where pixel (0,0) of the frame that supplies the even rows would be */
    int32_t weave_pitch, op_pad;
} DevLayer;

enum { DEV_BOTTOM_NONE = 0, DEV_BOTTOM_SINGLE = 1, DEV_BOTTOM_CROSS = 2 };

typedef struct DevJob {
    DevBox region;              /* pixels to write == the result's current_window */
    int32_t full_x, full_y;     /* output full_window.min */
    int32_t out_pitch;
    int32_t bottom_mode;        /* DEV_BOTTOM_* */
    void *out16;                /* rgba_f16 or NULL */
    void *out32;                /* rgba_f32 or NULL */
    void *out8;                 /* rgba_u8 or NULL */
    const uint8_t *u8tab;
    float mix_a, mix_b;         /* crossfade weights (mix_a = 1 - mix_b computed once in f32) */
    float single_alpha;         /* alpha scale of a single bottom source (1 unless copy_frame_alpha) */
    int32_t n_layers;
    int32_t flags;              /* JOB_FLAG_* (cv_kernels.cu) */
    const uint16_t *op_lut_709, *op_lut_srgb;   /* the colour transforms' tables (rec709->linear scene, linear->sRGB), 65536 halves each */
    float one;                  /* 1.0f, opaque to the compiler (see add2_unfused in cv_kernels.cu) */
    /* work decomposition: items of PX pixels, item (row, i) covers x = item_x0 + i*PX .. */
    int32_t item_x0;
    uint32_t items_per_row, rows, total_items, div_magic;
    DevSrc a, b;
    DevLayer layer[CV_MAX_LAYERS];
} DevJob;

/* ---- argument blocks of the small kernels ---- */

typedef struct EncodeArgs {
    uint2 *frame; uint32_t *target; const uint16_t *srgb; const uint8_t *ramp;
    int32_t pitch, ox, oy; DevBox win; uint32_t width, total, magic; int32_t write_back;
} EncodeArgs;

typedef struct ArgbArgs {
    const uint2 *frame; uint32_t *out; const uint8_t *ramp45;
    int32_t pitch, ox, oy; DevBox win; uint32_t width, total, magic;
} ArgbArgs;

typedef struct SubsampleArgs {
    uint2 *frame; int32_t pitch, ox, oy;
    uint8_t *plane[3]; int32_t stride[3];
    int32_t layout, sub_shift, matrix, pic_x, pic_y;
    int32_t wmin, wmax, row0, rows;         /* plane-local window */
    int32_t tx0; uint32_t tx_count, total, magic;
    const uint16_t *lut;                    /* linear -> Rec.709 */
} SubsampleArgs;

typedef struct LookupWindowArgs {
    uint2 *frame; int32_t pitch, ox, oy; DevBox win; uint32_t width, total, magic; const uint16_t *lut;
} LookupWindowArgs;

typedef struct OverF32Args {
    float4 *out; const float4 *b;
    int32_t out_pitch, out_ox, out_oy, b_pitch, b_ox, b_oy;
    DevBox out_cw, b_cw, region; float mix; uint32_t width, total, magic;
} OverF32Args;

typedef struct FillArgs {
    void *data; int32_t pitch, ox, oy, bytes; DevBox win; uint32_t width, total, magic; uint32_t word[4];
} FillArgs;

typedef struct ScaleArgs {
    float4 *target; int32_t t_pitch, t_ox, t_oy; uint32_t t_width, total, magic;
    const float4 *source; int32_t s_pitch, s_ox, s_oy;
    int32_t vertical, cross_lo, cross_hi;
    const int32_t *start, *src; const float *coef;       /* per target line: taps start[t]..start[t+1]-1 */
} ScaleArgs;

/* both passes in one launch (cv_scale_fused.cu): a block = a TX x TY tile of the target, its slice of the temporary in
 * shared memory.  Pass 1 fills the temporary (lines1 lines starting at coordinate origin1 along its axis, cross range
 * cross1 on the other), pass 2 reads it. */
enum { CV_SCALE_FUSED_TX = 64, CV_SCALE_FUSED_TY = 16, CV_SCALE_FUSED_TY_SHORT = 8,      /* tile rows: enlargements / the rest */
       CV_SCALE_FUSED_MAX_TEMP = 3072,      /* pixels of the temporary per tile: 48 KB of shared memory (measured: with 96 KB and
                                             * 16-row tiles a 3:1 reduction runs at 0.029 ms against 0.023 ms for the two passes;
                                             * with 8-row tiles it fits 48 KB and runs at 0.019 ms) */
       CV_SCALE_FUSED_MAX_TAPS = 1024 };    /* taps of one pass that a tile's lines may have (staged in shared memory) */
typedef struct FusedScaleArgs {
    float4 *target; int32_t t_pitch, t_ox, t_oy; uint32_t t_width, t_height;
    const float4 *source; int32_t s_pitch, s_ox, s_oy;
    int32_t x_first;                                     /* 1: horizontal pass first */
    int32_t origin1, lines1, cross1_lo, cross1_hi;
    const int32_t *start1, *src1; const float *coef1;
    int32_t cross2_lo, cross2_hi;
    const int32_t *start2, *src2; const float *coef2;   /* indexed by target line along the second axis */
    uint32_t temp_pixels;                                /* largest slice of the temporary any tile needs */
    float one;                                           /* 1.0f, opaque to the compiler (see tap_accumulate) */
    uint32_t cap1, cap2;                                 /* most taps of a pass that any tile's lines have, rounded up to 4 */
    int32_t tile_rows;                                   /* CV_SCALE_FUSED_TY or CV_SCALE_FUSED_TY_SHORT */
} FusedScaleArgs;

typedef struct ColorArgs {
    uint2 *frame; int32_t pitch, ox, oy; DevBox win; uint32_t width, total, magic;
    const uint16_t *pre, *post;                             /* tables applied to all four channels before / after, or NULL */
    float m[9];                                             /* out.x = (r*m[0] + g*m[1]) + b*m[2], ... */
    float one;                                              /* 1.0f, opaque to the compiler (see add2_unfused in cv_device.cuh) */
} ColorArgs;

typedef struct Mpeg2Args {
    const uint2 *frame; int32_t pitch, ox, oy; DevBox cw;   /* f16 source, valid pixels (others read as zero) */
    int32_t pic_x, pic_y, width, height, matrix;            /* picture placement in frame coordinates, even width/height */
    uint8_t *plane[3]; int32_t stride[3];
    const float *to709;                                     /* half bits -> f32 linear->Rec.709, 65536 entries */
    uint32_t luma_total, luma_magic, chroma_total, chroma_width, chroma_magic;
} Mpeg2Args;

enum { JOB_FLAG_SIMPLE = 1,     /* every layer covers the whole region, mix == 1, no bottom: no window tests */
       JOB_FLAG_EXACT = 2,      /* f32 frames involved: IEEE-division path only */
       JOB_FLAG_SEG = 4,        /* every f16 input is pair-aligned with the output: the row-segment kernel (k_seg) applies */
       JOB_FLAG_OPS = 8 };      /* some layer carries fused per-pixel filters: k_rect (or the per-pixel path) renders it */

/* Kernel shape selectors */
enum { KB_LAYERS = 0,           /* no bottom sources: layer[0] is fetched, the rest go over it */
       KB_GENERAL = 1,          /* bottom from a/b (f16 or f32 sources), no table needed */
       KB_CODED = 2,            /* bottom from a/b with at least one coded source: 128 KiB table in SMEM */
       KB_COUNT = 3 };

/* ---- windowed jobs as rectangles of constant coverage (cv_rects.c -> k_rect) ---- */

enum { CV_RECT_MAX_BANDS = 40, CV_RECT_MAX_PIECES = 2048 };
enum { RECT_FIRST_COPY = 1,     /* nothing lies below the first covering layer: it is fetched as it is */
       RECT_RAGGED = 2 };       /* a one-pair column split by a window edge: per-item code */

/* what one warp renders at a time: `pairs` pixel pairs of one row starting at pair column pair0 (relative to item_x0) */
typedef struct RectPiece {
    uint32_t pair0;
    uint16_t pairs;
    uint8_t bottom;             /* bit 0: source a covers it, bit 1: source b */
    uint8_t flags;              /* RECT_* */
    uint32_t mask;              /* bit k: layer k covers it */
    uint32_t reserved;
} RectPiece;

/* rows y0 .. y0 + rows - 1 share one list of `ppr` pieces per row (entries first_entry ..).  The kernel's unit of work
 * is a piece over a GROUP of up to rows_per_unit consecutive rows of its band (decoded once, then walked row by row);
 * units of a band are numbered group-major, `ustart` = units of the bands above in one frame. */
typedef struct RectBand { int32_t y0; uint32_t rows, ppr, magic, first_entry, ustart; } RectBand;

typedef struct RectPlan {
    int32_t n_bands;
    uint32_t n_entries, rows_per_unit, units_per_frame, frame_magic;
    uint64_t row_pieces_per_frame;          /* sum over bands of rows x ppr (statistics, load balance) */
    RectBand band[CV_RECT_MAX_BANDS];
} RectPlan;

/* ---- per-device state -------------------------------------------------------------- */

typedef struct CvDevice {
    int ready;
    int index;
    int sm_count;
    size_t l2_bytes;
    cudaStream_t stream;                /* library default stream */
    half *lut[CV_LUT_COUNT];            /* device copies, 65536 entries each */
    half *lut709_compact;               /* rec709->linear restricted to the halves a conversion can produce (cv_runtime.c) */
    uint8_t *gamma45;                   /* device copy */
    float *to709_f32;                   /* linear -> Rec.709 as f32 per half bit pattern (MPEG-2 subsample) */
    void *flush_buf; size_t flush_bytes;
    DevJob *job_ring_dev;               /* device ring for batched job descriptors */
    DevJob *job_ring_host;              /* pinned mirror */
    uint32_t job_ring_size, job_ring_pos;
    cudaEvent_t *job_ring_events;       /* one per ring segment */
    RectPiece *piece_ring_dev, *piece_ring_host;    /* CV_RECT_MAX_PIECES entries per job-ring segment (k_rect's piece lists) */
    pthread_mutex_t lock;               /* guards the job ring and the pool */
} CvDevice;

#ifdef __cplusplus
extern "C" {
#endif

CvDevice *cvi_device( void );                       /* current device state or NULL (sets last error) */
int cvi_fail( int code, const char *fmt, ... );     /* records the message, returns code */
int cvi_cuda( cudaError_t e, const char *what );    /* 0 or CV_ERR_CUDA */
cudaStream_t cvi_stream( cv_stream s );
void cvi_count_launch( unsigned n );

/* pooled device allocations (cv_memory.c).  cvi_pool_free marks a block reusable at once (the caller has synchronised
 * the work that used it); cvi_pool_free_async records an event behind the work queued on `stream` and the block is
 * handed out again only after that event has completed. */
int cvi_pool_alloc( void **ptr, size_t bytes );
int cvi_pool_free( void *ptr );
int cvi_pool_free_async( void *ptr, cudaStream_t stream );
/* pooled pinned host memory for staging (never returned to the OS before cv_shutdown) */
int cvi_pinned_alloc( void **ptr, size_t bytes );
void cvi_pinned_free( void *ptr );
void cvi_pinned_release_all( void );

/* per-thread context of the framework.h entry points (cv_host_ctx.c): own stream, pinned staging slabs */
enum { CV_HOST_SLABS = 4 };
typedef struct CvHostCtx {
    int device;
    int next;
    cudaStream_t stream;
    struct { void *ptr; size_t bytes; cudaEvent_t done; int busy; } slab[CV_HOST_SLABS];
} CvHostCtx;
CvHostCtx *cvi_host_ctx( void );                                    /* NULL (error recorded) when no B200 can be initialised */
void *cvi_host_slab( CvHostCtx *ctx, size_t bytes, int *index );    /* pinned staging memory; waits for the slab's previous copy */
void cvi_host_slab_done( CvHostCtx *ctx, int index );               /* the slab's contents are in flight on ctx->stream */
void cvi_pool_release_all( int device );
void cvi_scale_cache_release( int device );                         /* cv_scale.c: cached tap tables go before the pool does */

/* kernels (cv_kernels.cu) */
int cvi_launch_jobs( int shape, int px /*1|2*/, const DevJob *host_jobs, int count, cudaStream_t stream );
int cvi_launch_half_from_float( half *out, const float *in, size_t count, cudaStream_t stream );
int cvi_launch_half_to_float( float *out, const half *in, size_t count, cudaStream_t stream );
int cvi_launch_lookup( const half *table, half *out, const half *in, size_t count, cudaStream_t stream );
int cvi_launch_flush( void *buf, size_t bytes, cudaStream_t stream );
int cvi_launch_encode_u8( const EncodeArgs *args, cudaStream_t stream );
int cvi_launch_to_argb32( const ArgbArgs *args, cudaStream_t stream );
int cvi_launch_subsample( const SubsampleArgs *args, cudaStream_t stream );
int cvi_launch_lookup_window( const LookupWindowArgs *args, cudaStream_t stream );
int cvi_launch_over_f32( const OverF32Args *args, cudaStream_t stream );
int cvi_launch_fill( const FillArgs *args, cudaStream_t stream );
int cvi_launch_scale_axis( const ScaleArgs *args, cudaStream_t stream );
int cvi_launch_scale_fused( const FusedScaleArgs *args, cudaStream_t stream );     /* cv_scale_fused.cu */
int cvi_launch_gain_offset( const FillArgs *args, cudaStream_t stream );
int cvi_launch_color( const ColorArgs *args, cudaStream_t stream );
int cvi_launch_subsample_mpeg2( const Mpeg2Args *args, cudaStream_t stream );
int cvi_kernels_init( void );                       /* per-device function attributes */

/* rectangles of constant coverage (cv_rects.c): 0 and a filled plan, or -1 when the job is not one the rectangle
 * kernel takes (f32 frames, layer mixes other than 1, 4:1:1 sources, too many pieces) */
int cvi_rect_applies( const DevJob *job );
int cvi_rect_plan( const DevJob *job, int ppl, int rows_per_unit, RectPlan *plan, RectPiece *pieces, int capacity );
/* cv_rect.cu: the rectangle-walking kernel (CV_NO_RECT=1 disables it: windowed jobs then use k_seg's classifying mode) */
int cvi_rect_enabled( void );
int cvi_rect_ppl( const DevJob *job );          /* pixel pairs per lane and piece the kernel will be launched with */
int cvi_rect_rows( const DevJob *job, int ppl, int count );     /* rows per unit of work for a launch over `count` frames */
int cvi_launch_rect( int shape, int ppl, const DevJob *job, const DevJob *dev_jobs, int count, const RectPlan *plan,
                     const RectPiece *dev_pieces, cudaStream_t stream );
uint32_t cvi_div_magic( uint32_t d );

#ifdef __cplusplus
}
#endif

#endif
