/*
 * include/framework.h -- canvas-b200 re-issue of the frame types and the video
 * source / filter C API of the per-pixel path.
 *
 * It replaces the video part of /root/reference/include/framework.h (geometry
 * :46-137, pixel/frame types :155-183, source vtable :185-213, free functions
 * :225-253, coded images :470-528, workspace :571-587, FIR filters :618-650).
 * Layouts, names and argument meaning are the reference's, so a C caller or the
 * fluggo.media.process extension compiles against it unchanged; what differs:
 *
 *   - no <glib.h>, no <GL/glew.h>: gint/gpointer/GFreeFunc are spelled as the C
 *     types they alias, and the GL slot of the vtable keeps its offset through
 *     a uint32_t texture name but is never filled (no graphics stack here);
 *   - audio, codec packets, presentation clocks and the display widget are out
 *     of scope of this library and are not declared;
 *   - every function declared here is implemented on the GPU (sm_100a kernels
 *     behind include/canvas_b200.h).  Frames passed through THIS header hold
 *     HOST pointers, exactly as in the reference; they are staged through
 *     pinned memory.  Device-resident frames use the cv_* interface.
 */
#ifndef CANVAS_B200_FRAMEWORK_H
#define CANVAS_B200_FRAMEWORK_H

#include <stddef.h>
#include <stdint.h>
#include <inttypes.h>
#include <stdbool.h>
#include "half.h"

#if !defined(EXPORT)
#define EXPORT __attribute__((visibility("default")))
#endif

#define NS_PER_SEC INT64_C(1000000000)

#if defined(__cplusplus)
extern "C" {
#endif

/* ------------------------------------------------------------------ geometry */

typedef struct { int32_t n; uint32_t d; } rational;
typedef struct { int32_t x, y; } v2i;
typedef struct { v2i min, max; } box2i;     /* inclusive corners; empty <=> max < min */
typedef struct { float x, y; } v2f;
typedef struct { v2f min, max; } box2f;

static inline void v2i_add( v2i *r, const v2i *a, const v2i *b ) { r->x = a->x + b->x; r->y = a->y + b->y; }
static inline void v2i_subtract( v2i *r, const v2i *a, const v2i *b ) { r->x = a->x - b->x; r->y = a->y - b->y; }

#if !defined(__CUDACC__)     /* nvcc already supplies min/max(int, int) with this meaning */
static inline int min( int a, int b ) { return b > a ? a : b; }
static inline int max( int a, int b ) { return b < a ? a : b; }
#endif
static inline int clamp( int v, int lo, int hi ) { return min( max( v, lo ), hi ); }

/* NaN handling matters for clampf (a NaN mix value clamps to the LOW bound). */
static inline float minf( float a, float b ) { return a < b ? a : b; }
static inline float maxf( float a, float b ) { return a > b ? a : b; }
static inline float clampf( float v, float lo, float hi ) { return minf( maxf( v, lo ), hi ); }

static inline void box2i_set( box2i *b, int min_x, int min_y, int max_x, int max_y ) {
    b->min.x = min_x; b->min.y = min_y; b->max.x = max_x; b->max.y = max_y;
}
static inline void box2i_set_empty( box2i *b ) { box2i_set( b, 0, 0, -1, -1 ); }
static inline bool box2i_is_empty( const box2i *b ) { return b->max.x < b->min.x || b->max.y < b->min.y; }

static inline void box2i_intersect( box2i *r, const box2i *p, const box2i *q ) {
    box2i_set( r, max( p->min.x, q->min.x ), max( p->min.y, q->min.y ),
                  min( p->max.x, q->max.x ), min( p->max.y, q->max.y ) );
}
static inline void box2i_union( box2i *r, const box2i *p, const box2i *q ) {
    box2i_set( r, min( p->min.x, q->min.x ), min( p->min.y, q->min.y ),
                  max( p->max.x, q->max.x ), max( p->max.y, q->max.y ) );
}
/* Turns an inverted range into the gap it encloses (used by the 9-region walk). */
static inline void box2i_normalize( box2i *b ) {
    if( b->min.x > b->max.x ) { int32_t t = b->min.x - 1; b->min.x = b->max.x + 1; b->max.x = t; }
    if( b->min.y > b->max.y ) { int32_t t = b->min.y - 1; b->min.y = b->max.y + 1; b->max.y = t; }
}
static inline void box2i_get_size( const box2i *b, v2i *size ) {
    size->x = b->max.x < b->min.x ? 0 : b->max.x - b->min.x + 1;
    size->y = b->max.y < b->min.y ? 0 : b->max.y - b->min.y + 1;
}

int64_t get_frame_time( const rational *frame_rate, int frame );
int get_time_frame( const rational *frame_rate, int64_t time );
int64_t gettime( void );

/* -------------------------------------------------------------------- frames */

typedef struct { half r, g, b, a; } rgba_f16;       /* 8 bytes, straight (non-premultiplied) alpha */
typedef struct { uint8_t r, g, b, a; } rgba_u8;
typedef struct { float r, g, b, a; } rgba_f32;

/*
 * data holds the pixels of full_window, row-major, no padding: the pixel (x, y)
 * is data[(y - full.min.y) * width(full) + (x - full.min.x)].  The CALLER owns
 * data and sets full_window; a source writes inside full_window only and sets
 * current_window to the box of valid pixels (empty = nothing / failure).
 */
typedef struct { rgba_f16 *data; box2i full_window, current_window; } rgba_frame_f16;
typedef struct { rgba_f32 *data; box2i full_window, current_window; } rgba_frame_f32;
typedef struct { uint32_t texture; box2i full_window, current_window; } rgba_frame_gl;     /* never produced here */

static inline rgba_f16 *video_get_pixel_f16( rgba_frame_f16 *f, int x, int y ) {
    const int pitch = f->full_window.max.x - f->full_window.min.x + 1;
    return f->data + ((y - f->full_window.min.y) * pitch + (x - f->full_window.min.x));
}
static inline rgba_f32 *video_get_pixel_f32( rgba_frame_f32 *f, int x, int y ) {
    const int pitch = f->full_window.max.x - f->full_window.min.x + 1;
    return f->data + ((y - f->full_window.min.y) * pitch + (x - f->full_window.min.x));
}

typedef void (*video_get_frame_func)( void *self, int frame_index, rgba_frame_f16 *frame );
typedef void (*video_get_frame_32_func)( void *self, int frame_index, rgba_frame_f32 *frame );
typedef void (*video_get_frame_gl_func)( void *self, int frame_index, rgba_frame_gl *frame );

/* Any subset of the slots may be NULL; video_get_frame_* converts between precisions. */
typedef struct {
    int flags;                              /* reserved, zero */
    video_get_frame_func get_frame;
    video_get_frame_32_func get_frame_32;
    video_get_frame_gl_func get_frame_gl;   /* left NULL by every source of this library */
} video_frame_source_funcs;

typedef struct { void *obj; video_frame_source_funcs *funcs; } video_source;

static inline void rgba_f32_to_f16( rgba_f16 *out, const rgba_f32 *in, int count ) {
    half_convert_from_float( &out->r, &in->r, 4 * count );
}
static inline void rgba_f16_to_f32( rgba_f32 *out, const rgba_f16 *in, int count ) {
    half_convert_to_float( &out->r, &in->r, 4 * count );
}

void video_get_frame_f16( video_source *source, int frame_index, rgba_frame_f16 *frame );
void video_get_frame_f32( video_source *source, int frame_index, rgba_frame_f32 *frame );

void video_copy_frame_f16( rgba_frame_f16 *out, rgba_frame_f16 *in );
void video_copy_frame_alpha_f32( rgba_frame_f32 *out, rgba_frame_f32 *in, float alpha );
void video_mix_cross_f32_pull( rgba_frame_f32 *out, video_source *a, int frame_a, video_source *b, int frame_b, float mix_b );
void video_mix_cross_f32( rgba_frame_f32 *out, rgba_frame_f32 *a, rgba_frame_f32 *b, float mix_b );
void video_mix_over_f32( rgba_frame_f32 *out, rgba_frame_f32 *b, float mix_b );
/* video_scale.c:231-319 (reference framework.h:246-247) */
void video_scale_bilinear_f32( rgba_frame_f32 *target, v2f target_point, rgba_frame_f32 *source, v2f source_point, v2f factors );
void video_scale_bilinear_f32_pull( rgba_frame_f32 *target, v2f target_point, video_source *source, int frame, box2i *source_rect, v2f source_point, v2f factors );
/* color.c:100-160: exported by the reference's library (EXPORT) without a declaration in its framework.h */
void video_color_rgb_to_xyz_sdtv( rgba_frame_f16 *frame );
void video_color_xyz_to_srgb( rgba_frame_f16 *frame );

/* 65536-entry transfer tables applied as gathers; out may alias in. */
void video_transfer_rec709_to_linear_scene( half *out, const half *in, size_t count );
void video_transfer_rec709_to_linear_display( half *out, const half *in, size_t count );
void video_transfer_linear_to_rec709( half *out, const half *in, size_t count );
void video_transfer_linear_to_sRGB( half *out, const half *in, size_t count );
const uint8_t *video_get_gamma45_ramp( void );

/* -------------------------------------------------------------- coded images */

#define CODED_IMAGE_MAX_PLANES 4

typedef struct {
    void *data[CODED_IMAGE_MAX_PLANES];
    int stride[CODED_IMAGE_MAX_PLANES];
    int line_count[CODED_IMAGE_MAX_PLANES];
    void (*free_func)( void *image );
} coded_image;

coded_image *coded_image_alloc( const int *strides, const int *line_counts, int count );
coded_image *coded_image_alloc0( const int *strides, const int *line_counts, int count );

typedef coded_image *(*coded_image_getFrameFunc)( void *self, int frame, int quality_hint );
typedef struct { int flags; coded_image_getFrameFunc getFrame; } coded_image_source_funcs;
typedef struct { void *obj; coded_image_source_funcs *funcs; } coded_image_source;

/* 720x480 planar 4:1:1 Rec.709 picture at frame offset (0,-1), as in the reference. */
void video_reconstruct_dv( rgba_frame_f16 *frame, coded_image *planar );
coded_image *video_subsample_dv( rgba_frame_f16 *frame );

/* ----------------------------------------------------------------- workspace */

typedef struct workspace_t_tag workspace_t;
typedef struct workspace_item_t_tag workspace_item_t;

workspace_t *workspace_create( void );
int workspace_get_length( workspace_t *workspace );
workspace_item_t *workspace_add_item( workspace_t *self, void *source, int64_t x, int64_t width, int64_t offset, int64_t z, void *tag );
workspace_item_t *workspace_get_item( workspace_t *self, int index );
void workspace_remove_item( workspace_item_t *item );
void workspace_as_video_source( workspace_t *workspace, video_source *source );
void workspace_free( workspace_t *workspace );
void workspace_get_item_pos( workspace_item_t *item, int64_t *x, int64_t *width, int64_t *z );
int64_t workspace_get_item_offset( workspace_item_t *item );
void workspace_set_item_offset( workspace_item_t *item, int64_t offset );
void *workspace_get_item_source( workspace_item_t *item );
void workspace_set_item_source( workspace_item_t *item, void *source );
void *workspace_get_item_tag( workspace_item_t *item );
void workspace_set_item_tag( workspace_item_t *item, void *tag );
void workspace_update_item( workspace_item_t *item, int64_t *x, int64_t *width, int64_t *z, int64_t *offset, void **source, void **tag );

/* --------------------------------------------------------------- FIR filters */

typedef struct {
    float *coeff;       /* taps */
    int width;          /* number of taps */
    int center;         /* index of the centre tap */
} fir_filter;

/*
 * Triangle taps for 1:sub up-sampling (sub > 1, not normalised) or sub:1
 * down-sampling (sub < 1, normalised to unit sum).  With coeff == NULL the array is
 * allocated; otherwise width is the capacity on entry, and when it is too small
 * coeff is left alone, center becomes -1 and width the capacity needed.
 */
void filter_createTriangle( float sub, float offset, fir_filter *filter );
void filter_free( fir_filter *filter );

#if defined(__cplusplus)
}
#endif

#endif
