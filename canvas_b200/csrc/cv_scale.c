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
#include <pthread.h>
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

/* ------------------------------------------------------------- cache of uploaded tap tables */

/*
 * A video scales every frame of a clip by the same geometry, so the tap tables -- a pure function of the points, the
 * factors, the source's current window and the target's full window -- are planned and uploaded once and kept in
 * HBM per device (a few hundred KB each), least recently used of SCALE_CACHE_WAYS replaced.  A hit costs a key
 * compare and the launch; a miss plans on the host and uploads on the caller's stream (other streams wait on the
 * entry's `ready` event).  An entry's table may be in use by launches already queued on any stream, so replacing one
 * synchronises the device first (rare: more than SCALE_CACHE_WAYS geometries in rotation).  The lock is held from
 * lookup to launch so that no entry is replaced between the two.  CV_NO_SCALE_CACHE=1 turns hits off (for timing
 * the miss path); cv_shutdown() drops everything (cvi_scale_cache_release).
 */
enum { SCALE_CACHE_WAYS = 12, SCALE_KIND_AXIS = 1, SCALE_KIND_FUSED = 2 };
enum { ENTRY_FREE = 0, ENTRY_READY = 1, ENTRY_NOT_APPLICABLE = 2 };

typedef struct ScaleKey {
    int32_t kind, order;            /* order: `vertical` for one axis, `x_first` for both */
    float tp[2], sp[2], factor[2];
    box2i source_cw, target_full;
} ScaleKey;

typedef struct ScaleEntry {
    int state;
    uint64_t stamp;
    ScaleKey key;
    void *table;                    /* pooled device block holding every array the args point into */
    cudaEvent_t ready;              /* recorded behind the upload */
    ScaleArgs axis;                 /* templates: frame pointers, pitches and origins are filled in per call */
    FusedScaleArgs fused;
    box2i window;                   /* the target's current_window after the launch */
} ScaleEntry;

static ScaleEntry g_scale_cache[CV_MAX_DEVICES][SCALE_CACHE_WAYS];
static pthread_mutex_t g_scale_lock = PTHREAD_MUTEX_INITIALIZER;
static uint64_t g_scale_stamp;

static int cache_enabled( void ) {
    static int state = -1;
    if( state < 0 ) { const char *e = getenv( "CV_NO_SCALE_CACHE" ); state = !( e && *e && *e != '0' ); }
    return state;
}

static void entry_drop( ScaleEntry *e, int synchronise ) {
    if( e->state == ENTRY_READY ) {
        if( synchronise ) cudaDeviceSynchronize();       /* launches queued on any stream may still read the table */
        if( e->table ) cvi_pool_free( e->table );
        if( e->ready ) cudaEventDestroy( e->ready );
    }
    memset( e, 0, sizeof(*e) );
}

/* cv_shutdown(), with the device current and idle, before the pool goes */
void cvi_scale_cache_release( int device ) {
    if( device < 0 || device >= CV_MAX_DEVICES ) return;
    pthread_mutex_lock( &g_scale_lock );
    for( int i = 0; i < SCALE_CACHE_WAYS; i++ )
        entry_drop( &g_scale_cache[device][i], 0 );
    pthread_mutex_unlock( &g_scale_lock );
}

/* with g_scale_lock held: the entry of `key` (state READY / NOT_APPLICABLE) or a cleared slot to fill (state FREE) */
static ScaleEntry *cache_find( int device, const ScaleKey *key ) {
    ScaleEntry *set = g_scale_cache[device], *victim = NULL;
    for( int i = 0; i < SCALE_CACHE_WAYS && cache_enabled(); i++ )
        if( set[i].state != ENTRY_FREE && !memcmp( &set[i].key, key, sizeof(*key) ) ) {
            set[i].stamp = ++g_scale_stamp;
            return &set[i];
        }
    for( int i = 0; i < SCALE_CACHE_WAYS && !victim; i++ )
        if( set[i].state == ENTRY_FREE ) victim = &set[i];
    for( int i = 0; i < SCALE_CACHE_WAYS && ( !victim || victim->state != ENTRY_FREE ); i++ )
        if( !victim || set[i].stamp < victim->stamp ) victim = &set[i];
    entry_drop( victim, 1 );
    victim->key = *key;
    victim->stamp = ++g_scale_stamp;
    return victim;
}

/* host arrays -> one pooled device block on `stream`; records the entry's ready event */
static int entry_upload( ScaleEntry *e, const int32_t *host_words, size_t words, cudaStream_t stream ) {
    int rc = cvi_pool_alloc( &e->table, sizeof(int32_t) * words );
    if( !rc ) rc = cvi_cuda( cudaMemcpyAsync( e->table, host_words, sizeof(int32_t) * words, cudaMemcpyHostToDevice, stream ), "cudaMemcpyAsync(taps)" );
    /* the copy reads pageable memory: it has been staged by the time the call returns */
    if( !rc ) rc = cvi_cuda( cudaEventCreateWithFlags( &e->ready, cudaEventDisableTiming ), "cudaEventCreate" );
    if( !rc ) rc = cvi_cuda( cudaEventRecord( e->ready, stream ), "cudaEventRecord" );
    if( rc ) {
        if( e->table ) cvi_pool_free_async( e->table, stream );
        if( e->ready ) cudaEventDestroy( e->ready );
        e->table = NULL; e->ready = NULL;
    }
    return rc;
}

static int imax2( int a, int b ) { return a > b ? a : b; }
static int imin2( int a, int b ) { return a < b ? a : b; }

/* current_window of an axis pass (video_scale.c:123-140, 224-228) */
static void window_after( box2i *w, const AxisPlan *plan, int vertical, int cross_lo, int cross_hi ) {
    if( vertical ) box2i_set( w, cross_lo, plan->used_lo, cross_hi, plan->used_hi );
    else if( cross_lo <= cross_hi ) box2i_set( w, plan->used_lo, cross_lo, plan->used_hi, cross_hi );
    else box2i_set( w, INT_MAX, cross_lo, INT_MIN, cross_hi );      /* :139-140 untouched by an empty row loop */
}

static void key_init( ScaleKey *key, int kind, int order, const box2i *source_cw, const box2i *target_full ) {
    memset( key, 0, sizeof(*key) );
    key->kind = kind; key->order = order;
    key->source_cw = *source_cw; key->target_full = *target_full;
}

/* one axis: target (whole full_window written) from source->current_window */
static int scale_axis( rgba_frame_f32 *target, float tmin, const rgba_frame_f32 *source, float smin, float factor, int vertical, cudaStream_t stream ) {
    const box2i s = source->current_window, t = target->full_window;
    const int tw = t.max.x - t.min.x + 1, th = t.max.y - t.min.y + 1;
    CvDevice *dev = cvi_device();
    if( !dev ) { box2i_set_empty( &target->current_window ); return CV_ERR_NOT_INITIALISED; }

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

    ScaleKey key;
    key_init( &key, SCALE_KIND_AXIS, vertical, &s, &t );
    key.tp[0] = tmin; key.sp[0] = smin; key.factor[0] = factor;

    pthread_mutex_lock( &g_scale_lock );
    ScaleEntry *e = cache_find( dev->index, &key );
    int rc = CV_OK;

    if( e->state == ENTRY_FREE ) {
        const int cross_lo = vertical ? imax2( s.min.x, t.min.x ) : imax2( s.min.y, t.min.y );
        const int cross_hi = vertical ? imin2( s.max.x, t.max.x ) : imin2( s.max.y, t.max.y );
        AxisPlan plan;
        rc = vertical ? plan_axis( &plan, tmin, t.min.y, t.max.y, smin, s.min.y, s.max.y, factor )
                      : plan_axis( &plan, tmin, t.min.x, t.max.x, smin, s.min.x, s.max.x, factor );
        if( !rc ) {
            /* the table travels as one block: [start | src | coef] */
            const int lines = vertical ? th : tw;
            const size_t n_start = (size_t) lines + 1, n_taps = (size_t) plan.total + 1;
            const size_t words = n_start + 2 * n_taps;
            int32_t *host_table = (int32_t *) malloc( sizeof(int32_t) * words );
            if( !host_table ) rc = cvi_fail( CV_ERR_MEMORY, "cv_scale_bilinear_f32: out of host memory" );
            else {
                memcpy( host_table, plan.start, sizeof(int32_t) * n_start );
                memcpy( host_table + n_start, plan.src, sizeof(int32_t) * n_taps );
                memcpy( host_table + n_start + n_taps, plan.coef, sizeof(float) * n_taps );
                rc = entry_upload( e, host_table, words, stream );
                free( host_table );
            }
            if( !rc ) {
                ScaleArgs *A = &e->axis;
                A->t_pitch = tw; A->t_ox = t.min.x; A->t_oy = t.min.y;
                A->t_width = (uint32_t) tw; A->total = (uint32_t) tw * (uint32_t) th; A->magic = cvi_div_magic( (uint32_t) tw );
                A->vertical = vertical; A->cross_lo = cross_lo; A->cross_hi = cross_hi;
                A->start = (const int32_t *) e->table;
                A->src = A->start + n_start;
                A->coef = (const float *)( A->src + n_taps );
                window_after( &e->window, &plan, vertical, cross_lo, cross_hi );
                e->state = ENTRY_READY;
            }
            plan_free( &plan );
        }
    }

    if( !rc ) rc = cvi_cuda( cudaStreamWaitEvent( stream, e->ready, 0 ), "cudaStreamWaitEvent(taps)" );
    if( !rc ) {
        ScaleArgs A = e->axis;
        A.target = (float4 *) target->data;
        A.source = (const float4 *) source->data;
        A.s_pitch = source->full_window.max.x - source->full_window.min.x + 1;
        A.s_ox = source->full_window.min.x; A.s_oy = source->full_window.min.y;
        rc = cvi_launch_scale_axis( &A, stream );
    }
    if( rc ) box2i_set_empty( &target->current_window );
    else target->current_window = e->window;
    pthread_mutex_unlock( &g_scale_lock );
    return rc;
}

/* ------------------------------------------------------------- both passes in one launch (cv_scale_fused.cu) */

/* The largest slice of the temporary a tile of `per_tile` consecutive second-pass lines reads, in lines of the
 * temporary; -1 when a tile's taps do not lie between its first and its last one (the kernel's assumption). */
static int widest_slice( const AxisPlan *p, int lines, int per_tile ) {
    int widest = 0;
    for( int t0 = 0; t0 < lines; t0 += per_tile ) {
        const int t1 = t0 + per_tile < lines ? t0 + per_tile : lines;
        const int first = p->start[t0], last = p->start[t1];
        if( first >= last ) continue;
        int lo = INT_MAX, hi = INT_MIN;
        for( int i = first; i < last; i++ ) {
            if( p->src[i] < lo ) lo = p->src[i];
            if( p->src[i] > hi ) hi = p->src[i];
        }
        if( lo != p->src[first] || hi != p->src[last - 1] ) return -1;
        if( hi - lo + 1 > widest ) widest = hi - lo + 1;
    }
    return widest;
}

/* the most taps any tile's lines have: tiles of `per_tile` lines, the first one starting at line `first_line` of the
 * plan (negative when the tile grid starts outside the plan's lines) */
static int most_taps( const AxisPlan *p, int lines, int first_line, int tiled_lines, int per_tile ) {
    int most = 0;
    for( int t0 = 0; t0 < tiled_lines; t0 += per_tile ) {
        int a = first_line + t0, b = first_line + ( t0 + per_tile < tiled_lines ? t0 + per_tile : tiled_lines );
        if( a < 0 ) a = 0;
        if( b > lines ) b = lines;
        if( b > a && p->start[b] - p->start[a] > most ) most = p->start[b] - p->start[a];
    }
    return most;
}

static int fused_enabled( void ) {
    static int state = -1;
    if( state < 0 ) { const char *e = getenv( "CV_NO_SCALE_FUSED" ); state = !( e && *e && *e != '0' ); }
    return state;
}

/* plans both passes into the entry: 1 = ready, 0 = not applicable, < 0 = error */
static int fused_build( ScaleEntry *e, const box2i *s, const box2i *t, v2f target_point, v2f source_point, v2f factors,
                        const box2i *want, int x_first, cudaStream_t stream ) {
    const int tw = t->max.x - t->min.x + 1, th = t->max.y - t->min.y + 1;

    /* pass 1: source -> temporary (full window `want`); pass 2: temporary -> target */
    const int vertical1 = !x_first, vertical2 = x_first;
    const float tmin1 = x_first ? target_point.x : target_point.y, smin1 = x_first ? source_point.x : source_point.y;
    const float tmin2 = x_first ? target_point.y : target_point.x, smin2 = x_first ? source_point.y : source_point.x;
    const float factor1 = x_first ? factors.x : factors.y, factor2 = x_first ? factors.y : factors.x;
    if( ( factor1 == 1.0f && tmin1 == smin1 ) || ( factor2 == 1.0f && tmin2 == smin2 ) ) return 0;      /* a copy pass */

    const int cross1_lo = vertical1 ? imax2( s->min.x, want->min.x ) : imax2( s->min.y, want->min.y );
    const int cross1_hi = vertical1 ? imin2( s->max.x, want->max.x ) : imin2( s->max.y, want->max.y );
    AxisPlan plan1, plan2;
    if( vertical1 ? plan_axis( &plan1, tmin1, want->min.y, want->max.y, smin1, s->min.y, s->max.y, factor1 )
                  : plan_axis( &plan1, tmin1, want->min.x, want->max.x, smin1, s->min.x, s->max.x, factor1 ) )
        return 0;                               /* the two-pass path reports it */
    box2i temp_cw;
    window_after( &temp_cw, &plan1, vertical1, cross1_lo, cross1_hi );
    if( box2i_is_empty( &temp_cw ) ) { plan_free( &plan1 ); return 0; }

    const int cross2_lo = vertical2 ? imax2( temp_cw.min.x, t->min.x ) : imax2( temp_cw.min.y, t->min.y );
    const int cross2_hi = vertical2 ? imin2( temp_cw.max.x, t->max.x ) : imin2( temp_cw.max.y, t->max.y );
    if( vertical2 ? plan_axis( &plan2, tmin2, t->min.y, t->max.y, smin2, temp_cw.min.y, temp_cw.max.y, factor2 )
                  : plan_axis( &plan2, tmin2, t->min.x, t->max.x, smin2, temp_cw.min.x, temp_cw.max.x, factor2 ) ) {
        plan_free( &plan1 );
        return 0;
    }

    const int lines1 = vertical1 ? want->max.y - want->min.y + 1 : want->max.x - want->min.x + 1;
    const int lines2 = vertical2 ? th : tw;
    /* tile height: 16 rows when both axes enlarge (1080p -> 4K: 0.050 ms against 0.060 ms with 8), 8 rows otherwise
     * (4K -> 1080p 0.021 against 0.025 ms; a 3:1 reduction fits only with 8 and then beats the two passes, 0.019 against
     * 0.023 ms) */
    const int tile_rows = ( factors.x > 1.0f && factors.y > 1.0f ) ? CV_SCALE_FUSED_TY : CV_SCALE_FUSED_TY_SHORT;
    const int slice = widest_slice( &plan2, lines2, vertical2 ? tile_rows : CV_SCALE_FUSED_TX );
    const long long temp_pixels = (long long) slice * ( vertical2 ? CV_SCALE_FUSED_TX : tile_rows );
    /* pass 1 is indexed by the tile's lines along the OTHER axis (rows of the tile for a vertical first pass) */
    const int taps1 = vertical1 ? most_taps( &plan1, lines1, t->min.y - want->min.y, th, tile_rows )
                                : most_taps( &plan1, lines1, t->min.x - want->min.x, tw, CV_SCALE_FUSED_TX );
    const int taps2 = most_taps( &plan2, lines2, 0, lines2, vertical2 ? tile_rows : CV_SCALE_FUSED_TX );
    if( slice < 0 || temp_pixels > CV_SCALE_FUSED_MAX_TEMP || taps1 > CV_SCALE_FUSED_MAX_TAPS || taps2 > CV_SCALE_FUSED_MAX_TAPS ) {
        plan_free( &plan1 ); plan_free( &plan2 );
        return 0;
    }

    /* one block: [start1 | src1 | coef1 | start2 | src2 | coef2] */
    const size_t n_start1 = (size_t) lines1 + 1, n_taps1 = (size_t) plan1.total + 1;
    const size_t n_start2 = (size_t) lines2 + 1, n_taps2 = (size_t) plan2.total + 1;
    const size_t words = n_start1 + 2 * n_taps1 + n_start2 + 2 * n_taps2;
    int32_t *host_table = (int32_t *) malloc( sizeof(int32_t) * words );
    int rc = host_table ? CV_OK : cvi_fail( CV_ERR_MEMORY, "cv_scale_bilinear_f32: out of host memory" );
    if( !rc ) {
        int32_t *h = host_table;
        memcpy( h, plan1.start, sizeof(int32_t) * n_start1 ); h += n_start1;
        memcpy( h, plan1.src, sizeof(int32_t) * n_taps1 ); h += n_taps1;
        memcpy( h, plan1.coef, sizeof(float) * n_taps1 ); h += n_taps1;
        memcpy( h, plan2.start, sizeof(int32_t) * n_start2 ); h += n_start2;
        memcpy( h, plan2.src, sizeof(int32_t) * n_taps2 ); h += n_taps2;
        memcpy( h, plan2.coef, sizeof(float) * n_taps2 );
        rc = entry_upload( e, host_table, words, stream );
        free( host_table );
    }
    if( !rc ) {
        FusedScaleArgs *A = &e->fused;
        A->t_pitch = tw; A->t_ox = t->min.x; A->t_oy = t->min.y;
        A->t_width = (uint32_t) tw; A->t_height = (uint32_t) th;
        A->x_first = x_first;
        A->origin1 = vertical1 ? want->min.y : want->min.x; A->lines1 = lines1;
        A->cross1_lo = cross1_lo; A->cross1_hi = cross1_hi;
        A->cross2_lo = cross2_lo; A->cross2_hi = cross2_hi;
        const int32_t *d = (const int32_t *) e->table;
        A->start1 = d; d += n_start1;
        A->src1 = d; d += n_taps1;
        A->coef1 = (const float *) d; d += n_taps1;
        A->start2 = d; d += n_start2;
        A->src2 = d; d += n_taps2;
        A->coef2 = (const float *) d;
        A->temp_pixels = (uint32_t)( temp_pixels > 0 ? temp_pixels : 1 );
        A->one = 1.0f;
        A->tile_rows = tile_rows;
        A->cap1 = (uint32_t)( ( taps1 + 3 ) & ~3 ); A->cap2 = (uint32_t)( ( taps2 + 3 ) & ~3 );
        window_after( &e->window, &plan2, vertical2, cross2_lo, cross2_hi );
    }
    plan_free( &plan1 ); plan_free( &plan2 );
    return rc ? rc : 1;
}

/* 1 = rendered, 0 = not applicable (the caller takes the two-pass path), < 0 = error */
static int scale_fused( rgba_frame_f32 *target, v2f target_point, const rgba_frame_f32 *source, v2f source_point, v2f factors,
                        const box2i *want, int x_first, cudaStream_t stream ) {
    CvDevice *dev = cvi_device();
    const box2i s = source->current_window, t = target->full_window;
    if( !dev || !fused_enabled() || t.max.x < t.min.x || t.max.y < t.min.y ) return 0;

    ScaleKey key;
    key_init( &key, SCALE_KIND_FUSED, x_first, &s, &t );
    key.tp[0] = target_point.x; key.tp[1] = target_point.y;
    key.sp[0] = source_point.x; key.sp[1] = source_point.y;
    key.factor[0] = factors.x; key.factor[1] = factors.y;

    pthread_mutex_lock( &g_scale_lock );
    ScaleEntry *e = cache_find( dev->index, &key );
    int rc = 1;
    if( e->state == ENTRY_FREE ) {
        rc = fused_build( e, &s, &t, target_point, source_point, factors, want, x_first, stream );
        if( rc >= 0 ) e->state = rc ? ENTRY_READY : ENTRY_NOT_APPLICABLE;
    }
    else if( e->state == ENTRY_NOT_APPLICABLE ) rc = 0;

    if( rc > 0 ) {
        int err = cvi_cuda( cudaStreamWaitEvent( stream, e->ready, 0 ), "cudaStreamWaitEvent(taps)" );
        if( !err ) {
            FusedScaleArgs A = e->fused;
            A.target = (float4 *) target->data;
            A.source = (const float4 *) source->data;
            A.s_pitch = source->full_window.max.x - source->full_window.min.x + 1;
            A.s_ox = source->full_window.min.x; A.s_oy = source->full_window.min.y;
            err = cvi_launch_scale_fused( &A, stream );
        }
        if( err ) { box2i_set_empty( &target->current_window ); rc = err; }
        else target->current_window = e->window;
    }
    else if( rc < 0 ) box2i_set_empty( &target->current_window );
    pthread_mutex_unlock( &g_scale_lock );
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

    int rc = scale_fused( target, target_point, source, source_point, factors, &want, x_first, s );
    if( rc ) return rc < 0 ? rc : CV_OK;

    rc = cv_frame_alloc_f32( &temp, &want, CV_MEM_DEVICE );
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
