/*
 * canvas_b200/csrc/cv_scale.c -- the separable triangle-filter scaler (cv_scale_bilinear_f32).
 *
 * Follows /root/reference/src/cprocess/video_scale.c:24-319.  The reference scatters (upsampling) or
 * gathers (downsampling) one line at a time on the CPU, rebuilding the FIR taps for every line with
 * filter_createTriangle (filter.c:24-76).  Here the HOST turns each axis into a tap table -- for every
 * target line the source lines that reach it, in ascending order, with their coefficients, computed
 * by the same filter_createTriangle in f32 -- and ONE kernel launch per axis gathers every target
 * pixel from it (k_scale_axis): 0 + in*c + in*c ... in the reference's order, products and sums
 * rounded separately, so the f32 result has the reference's bits.  The whole target is written
 * (pixels no tap reaches are zero, like video_fill_zero_f32, :24-31).
 */
#define _GNU_SOURCE
#include "cv_internal.h"

#include <limits.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

enum { SCALE_MAX_TAPS = 256 };

typedef struct AxisPlan {
    int t_lo, t_hi;             /* target lines */
    int used_lo, used_hi;       /* target lines some tap reached (INT_MAX / INT_MIN when none) */
    int32_t *start;             /* [lines + 1] into src / coef */
    int32_t *src;
    float *coef;
    int total;
} AxisPlan;

static void plan_free( AxisPlan *p ) {
    free( p->start ); free( p->src ); free( p->coef );
    memset( p, 0, sizeof(*p) );
}

/* Visits every (target line, source line, coefficient) of one axis in the reference's order:
 * upsampling walks the source lines (video_scale.c:58-88), downsampling the target lines (:90-121). */
typedef void (*tap_visitor)( void *closure, int t, int s, float c );

static int walk_axis( float tmin, int t_lo, int t_hi, float smin, int s_lo, int s_hi, float factor, tap_visitor visit, void *closure ) {
    float buffer[SCALE_MAX_TAPS];
    const int up = factor > 1.0f;
    const int outer_lo = up ? s_lo : t_lo, outer_hi = up ? s_hi : t_hi;

    for( int outer = outer_lo; outer <= outer_hi; outer++ ) {
        const float where = up ? ((float) outer - smin) * factor + tmin : ((float) outer - tmin) / factor + smin;
        const int base = (int) floorf( where );

        fir_filter filter = { .coeff = buffer, .width = SCALE_MAX_TAPS, .center = 0 };
        filter_createTriangle( factor, where - (float) base, &filter );
        if( filter.center < 0 && filter.width > SCALE_MAX_TAPS )
            return cvi_fail( CV_ERR_UNSUPPORTED, "cv_scale_bilinear_f32: factor %g needs %d taps (at most %d)", factor, filter.width, SCALE_MAX_TAPS );

        for( int tap = 0; tap < filter.width; tap++ ) {
            const int inner = base - filter.center + tap;
            const int t = up ? inner : outer, s = up ? outer : inner;
            if( t < t_lo || t > t_hi || s < s_lo || s > s_hi )
                continue;
            visit( closure, t, s, buffer[tap] );
        }
    }
    return CV_OK;
}

static void count_tap( void *closure, int t, int s, float c ) {
    AxisPlan *p = (AxisPlan *) closure;
    (void) s; (void) c;
    p->start[t - p->t_lo + 1]++;
    if( t < p->used_lo ) p->used_lo = t;
    if( t > p->used_hi ) p->used_hi = t;
}

typedef struct Filler { AxisPlan *plan; int32_t *cursor; } Filler;

static void fill_tap( void *closure, int t, int s, float c ) {
    Filler *f = (Filler *) closure;
    const int at = f->cursor[t - f->plan->t_lo]++;
    f->plan->src[at] = s;
    f->plan->coef[at] = c;
}

static int plan_axis( AxisPlan *p, float tmin, int t_lo, int t_hi, float smin, int s_lo, int s_hi, float factor ) {
    memset( p, 0, sizeof(*p) );
    p->t_lo = t_lo; p->t_hi = t_hi;
    p->used_lo = INT_MAX; p->used_hi = INT_MIN;
    const int lines = t_hi >= t_lo ? t_hi - t_lo + 1 : 0;

    p->start = (int32_t *) calloc( (size_t) lines + 2, sizeof(int32_t) );
    if( !p->start ) return cvi_fail( CV_ERR_MEMORY, "cv_scale_bilinear_f32: out of host memory" );

    int rc = walk_axis( tmin, t_lo, t_hi, smin, s_lo, s_hi, factor, count_tap, p );
    if( rc ) { plan_free( p ); return rc; }

    for( int i = 0; i < lines; i++ )
        p->start[i + 1] += p->start[i];
    p->total = lines ? p->start[lines] : 0;

    p->src = (int32_t *) malloc( sizeof(int32_t) * (size_t)( p->total + 1 ) );
    p->coef = (float *) malloc( sizeof(float) * (size_t)( p->total + 1 ) );
    int32_t *cursor = (int32_t *) malloc( sizeof(int32_t) * (size_t)( lines + 1 ) );
    if( !p->src || !p->coef || !cursor ) { free( cursor ); plan_free( p ); return cvi_fail( CV_ERR_MEMORY, "cv_scale_bilinear_f32: out of host memory" ); }
    memcpy( cursor, p->start, sizeof(int32_t) * (size_t) lines );

    Filler filler = { p, cursor };
    rc = walk_axis( tmin, t_lo, t_hi, smin, s_lo, s_hi, factor, fill_tap, &filler );
    free( cursor );
    if( rc ) plan_free( p );
    return rc;
}

static int imax2( int a, int b ) { return a > b ? a : b; }
static int imin2( int a, int b ) { return a < b ? a : b; }

/* one axis: target (whole full_window written) from source->current_window */
static int scale_axis( rgba_frame_f32 *target, float tmin, const rgba_frame_f32 *source, float smin, float factor, int vertical, cudaStream_t stream ) {
    const box2i s = source->current_window, t = target->full_window;
    const int tw = t.max.x - t.min.x + 1, th = t.max.y - t.min.y + 1;

    if( tw <= 0 || th <= 0 ) {
        box2i_set_empty( &target->current_window );
        return CV_OK;
    }

    if( factor == 1.0f && tmin == smin ) {
        /* video_scale.c:44-47: zero fill, then a plain windowed copy */
        int rc = cvi_cuda( cudaMemsetAsync( target->data, 0, sizeof(rgba_f32) * (size_t) tw * (size_t) th, stream ), "cudaMemsetAsync" );
        if( rc ) return rc;
        return cv_copy_frame_alpha_f32( target, source, 1.0f, stream );
    }

    const int cross_lo = vertical ? imax2( s.min.x, t.min.x ) : imax2( s.min.y, t.min.y );
    const int cross_hi = vertical ? imin2( s.max.x, t.max.x ) : imin2( s.max.y, t.max.y );

    AxisPlan plan;
    int rc = vertical ? plan_axis( &plan, tmin, t.min.y, t.max.y, smin, s.min.y, s.max.y, factor )
                      : plan_axis( &plan, tmin, t.min.x, t.max.x, smin, s.min.x, s.max.x, factor );
    if( rc ) { box2i_set_empty( &target->current_window ); return rc; }

    /* the table travels through one pooled device block: [start | src | coef] */
    const int lines = vertical ? th : tw;
    const size_t n_start = (size_t) lines + 1, n_taps = (size_t) plan.total + 1;
    const size_t bytes = sizeof(int32_t) * (n_start + n_taps) + sizeof(float) * n_taps;
    void *dev_table = NULL, *host_table = malloc( bytes );
    if( !host_table ) { plan_free( &plan ); return cvi_fail( CV_ERR_MEMORY, "cv_scale_bilinear_f32: out of host memory" ); }
    memcpy( host_table, plan.start, sizeof(int32_t) * n_start );
    memcpy( (int32_t *) host_table + n_start, plan.src, sizeof(int32_t) * n_taps );
    memcpy( (int32_t *) host_table + n_start + n_taps, plan.coef, sizeof(float) * n_taps );

    rc = cvi_pool_alloc( &dev_table, bytes );
    if( !rc ) rc = cvi_cuda( cudaMemcpyAsync( dev_table, host_table, bytes, cudaMemcpyHostToDevice, stream ), "cudaMemcpyAsync(taps)" );
    /* the copy reads pageable memory: it has been staged by the time the call returns */

    if( !rc ) {
        ScaleArgs A;
        memset( &A, 0, sizeof(A) );
        A.target = (float4 *) target->data; A.t_pitch = tw; A.t_ox = t.min.x; A.t_oy = t.min.y;
        A.t_width = (uint32_t) tw; A.total = (uint32_t) tw * (uint32_t) th; A.magic = cvi_div_magic( (uint32_t) tw );
        A.source = (const float4 *) source->data;
        A.s_pitch = source->full_window.max.x - source->full_window.min.x + 1;
        A.s_ox = source->full_window.min.x; A.s_oy = source->full_window.min.y;
        A.vertical = vertical; A.cross_lo = cross_lo; A.cross_hi = cross_hi;
        A.start = (const int32_t *) dev_table;
        A.src = A.start + n_start;
        A.coef = (const float *)( A.src + n_taps );
        rc = cvi_launch_scale_axis( &A, stream );
    }

    if( dev_table )             /* the block returns to the pool once the kernel that reads it has run (stream-ordered) */
        cvi_pool_free_async( dev_table, stream );
    free( host_table );

    if( rc ) box2i_set_empty( &target->current_window );
    else if( vertical ) box2i_set( &target->current_window, cross_lo, plan.used_lo, cross_hi, plan.used_hi );
    else if( cross_lo <= cross_hi ) box2i_set( &target->current_window, plan.used_lo, cross_lo, plan.used_hi, cross_hi );
    else box2i_set( &target->current_window, INT_MAX, cross_lo, INT_MIN, cross_hi );      /* :139-140 untouched by an empty row loop */

    plan_free( &plan );
    return rc;
}

int cv_scale_bilinear_f32( rgba_frame_f32 *target, v2f target_point, const rgba_frame_f32 *source, v2f source_point, v2f factors, cv_stream stream ) {
    if( !target || !source ) return cvi_fail( CV_ERR_ARGUMENT, "cv_scale_bilinear_f32: NULL argument" );
    if( !cvi_device() ) { box2i_set_empty( &target->current_window ); return CV_ERR_NOT_INITIALISED; }
    if( !(factors.x > 0.0f) || !(factors.y > 0.0f) ) {
        box2i_set_empty( &target->current_window );
        return cvi_fail( CV_ERR_ARGUMENT, "cv_scale_bilinear_f32: factors must be positive" );
    }
    cudaStream_t s = cvi_stream( stream );

    /* video_scale.c:231-245: an identity axis is skipped */
    if( factors.x == 1.0f && target_point.x == source_point.x ) {
        if( factors.y == 1.0f && target_point.y == source_point.y )
            return cv_copy_frame_alpha_f32( target, source, 1.0f, stream );
        return scale_axis( target, target_point.y, source, source_point.y, factors.y, 1, s );
    }
    if( factors.y == 1.0f && target_point.y == source_point.y )
        return scale_axis( target, target_point.x, source, source_point.x, factors.x, 0, s );

    /* :247-284: the axis with the smaller factor first, through a temporary whose window uses C's (int) truncation */
    const box2i tf = target->full_window, sc = source->current_window;
    rgba_frame_f32 temp;
    box2i want;
    const int x_first = factors.x < factors.y;

    if( x_first )
        box2i_set( &want, (int)( source_point.x - (target_point.x - (float) tf.min.x) * factors.x ), sc.min.y,
                          (int)( source_point.x + ((float) tf.max.x - target_point.x) * factors.x ), sc.max.y );
    else
        box2i_set( &want, sc.min.x, (int)( source_point.y - (target_point.y - (float) tf.min.y) * factors.y ),
                          sc.max.x, (int)( source_point.y + ((float) tf.max.y - target_point.y) * factors.y ) );
    box2i_intersect( &want, &want, &tf );

    if( box2i_is_empty( &want ) ) {
        /* nothing of the intermediate frame falls into the target: the result is a zeroed, empty frame */
        const int tw = tf.max.x - tf.min.x + 1, th = tf.max.y - tf.min.y + 1;
        box2i_set_empty( &target->current_window );
        if( tw > 0 && th > 0 )
            return cvi_cuda( cudaMemsetAsync( target->data, 0, sizeof(rgba_f32) * (size_t) tw * (size_t) th, s ), "cudaMemsetAsync" );
        return CV_OK;
    }

    int rc = cv_frame_alloc_f32( &temp, &want, CV_MEM_DEVICE );
    if( rc ) { box2i_set_empty( &target->current_window ); return rc; }

    if( x_first ) {
        rc = scale_axis( &temp, target_point.x, source, source_point.x, factors.x, 0, s );
        if( !rc ) rc = scale_axis( target, target_point.y, &temp, source_point.y, factors.y, 1, s );
    }
    else {
        rc = scale_axis( &temp, target_point.y, source, source_point.y, factors.y, 1, s );
        if( !rc ) rc = scale_axis( target, target_point.x, &temp, source_point.x, factors.x, 0, s );
    }
    cvi_pool_free_async( temp.data, s );             /* behind the second pass, without waiting for it */
    if( rc ) box2i_set_empty( &target->current_window );
    return rc;
}
