/*
 * canvas_b200/csrc/cv_static.cu -- cv_scale_bilinear_static_f32: the two-axis bilinear scaler
 * (cv_scale_bilinear_f32, cv_scale.c) for callers that may not allocate.
 *
 * No counterpart in the reference.  cv_scale_bilinear_f32 takes
 * its scratch memory -- the per-axis tap tables and the intermediate frame between the two passes -- from the
 * device pool, which can reach cudaMalloc.  Code under CUDA graph capture, tight serving loops and harnesses
 * that forbid allocation cannot call that.  This entry point allocates NOTHING, on the device or on the host:
 * its scratch is one file-scope `static __device__` array of CV_STATIC_SCRATCH_BYTES that every GPU loading
 * the library holds until exit, and the host-side tap tables are built in file-scope arrays as well.
 *
 * Consequences the caller has to know (also in include/canvas_b200.h):
 *   - sizes beyond the stated maxima are rejected with CV_ERR_UNSUPPORTED, nothing is launched;
 *   - every call on a device shares the one array, which keeps the previous call's contents.  Nothing in it has
 *     to start at zero: both tap tables are written completely (host -> device copy on the call's stream)
 *     before the kernel that reads them, and the first pass writes the WHOLE intermediate frame (zero where
 *     no tap reaches) before the second reads it -- so no clearing is needed or done;
 *   - calls that overlap on the GPU (different streams, or different host threads) race on the array and on
 *     the host-side tables: one call at a time per process, stream-ordered after the previous one.
 *
 * The arithmetic is k_scale_axis's (cv_kernels.cu): 0 + in*c + in*c ... over a host-built table of the taps
 * of filter_createTriangle, products and sums rounded separately (-fmad=false), so results are bit-equal to
 * cv_scale_bilinear_f32's.
 */
#include "cv_internal.h"

#include <limits.h>
#include <math.h>
#include <string.h>
#include <pthread.h>

/* ------------------------------------------------------------- the static scratch and its layout */

/* CV_STATIC_MAX_TEMP_PIXELS (2 Mi pixels of 16 bytes), CV_STATIC_MAX_LINES and CV_STATIC_MAX_TAPS per axis: include/canvas_b200.h */
#define CV_STATIC_FILTER_TAPS       256                 /* taps of one line (as cv_scale.c's SCALE_MAX_TAPS) */

#define CV_STATIC_TEMP_BYTES        ( (size_t) CV_STATIC_MAX_TEMP_PIXELS * 16u )
#define CV_STATIC_START_BYTES       ( ( (size_t) CV_STATIC_MAX_LINES + 64u ) * 4u )         /* start[lines + 1], padded to 256 */
#define CV_STATIC_TAP_BYTES         ( (size_t) CV_STATIC_MAX_TAPS * 4u )
#define CV_STATIC_TABLE_BYTES       ( CV_STATIC_START_BYTES + 2u * CV_STATIC_TAP_BYTES )   /* start | src | coef */
#define CV_STATIC_SCRATCH_TOTAL     ( CV_STATIC_TEMP_BYTES + 2u * CV_STATIC_TABLE_BYTES )

/* Device memory; in HOST code this name is only a stand-in for cudaGetSymbolAddress, never an address. */
static __device__ __align__(256) unsigned char g_static_scratch[CV_STATIC_SCRATCH_TOTAL];

static_assert( CV_STATIC_SCRATCH_TOTAL == CV_STATIC_SCRATCH_BYTES, "include/canvas_b200.h states the size of the static scratch" );
static_assert( CV_STATIC_START_BYTES % 256u == 0 && CV_STATIC_TAP_BYTES % 256u == 0, "table parts stay 256-byte aligned" );

/* host-side tables of one axis (file scope: no malloc; shared by all calls like the device array) */
typedef struct HostTable {
    int32_t start[CV_STATIC_MAX_LINES + 2];
    int32_t cursor[CV_STATIC_MAX_LINES + 2];
    int32_t src[CV_STATIC_MAX_TAPS];
    float coef[CV_STATIC_MAX_TAPS];
    int total, used_lo, used_hi;
} HostTable;

static HostTable g_host_table[2];

/* per device: the array's device address, fetched once with that device current */
enum { CV_STATIC_MAX_DEVICES = 64 };
static unsigned char *g_scratch_address[CV_STATIC_MAX_DEVICES];
static pthread_mutex_t g_scratch_lock = PTHREAD_MUTEX_INITIALIZER;

static int scratch_for_current_device( unsigned char **out ) {
    int device = -1;
    int rc = cvi_cuda( cudaGetDevice( &device ), "cudaGetDevice" );
    if( rc ) return rc;
    if( device < 0 || device >= CV_STATIC_MAX_DEVICES )
        return cvi_fail( CV_ERR_UNSUPPORTED, "cv_scale_bilinear_static_f32: device ordinal %d beyond %d", device, CV_STATIC_MAX_DEVICES );

    pthread_mutex_lock( &g_scratch_lock );
    if( !g_scratch_address[device] ) {
        void *address = NULL;
        rc = cvi_cuda( cudaGetSymbolAddress( &address, g_static_scratch ), "cudaGetSymbolAddress(g_static_scratch)" );
        if( !rc && !address ) rc = cvi_fail( CV_ERR_CUDA, "cv_scale_bilinear_static_f32: the static scratch has no device address" );
        if( !rc ) g_scratch_address[device] = (unsigned char *) address;
    }
    *out = g_scratch_address[device];
    pthread_mutex_unlock( &g_scratch_lock );
    return rc;
}

/* ------------------------------------------------------------- kernel: one axis, taps from the static array */

typedef struct StaticScaleArgs {
    float4 *target; int32_t t_pitch, t_ox, t_oy; uint32_t t_width, t_height;
    const float4 *source; int32_t s_pitch, s_ox, s_oy;
    int32_t vertical, cross_lo, cross_hi;
    const int32_t *start, *src; const float *coef;       /* inside g_static_scratch; taps of line t: start[t]..start[t+1]-1 */
} StaticScaleArgs;

/*
 * The gather of k_scale_axis (cv_kernels.cu; video_scale.c:33-228): a thread per target pixel walks the taps
 * of its line in ascending source order, acc = acc + in*c with multiply and add rounded separately.  Threads
 * along x (128-bit loads and stores, coalesced for the vertical pass), rows strided over blockIdx.y.
 */
__global__ void __launch_bounds__( 256 ) k_scale_axis_static( const __grid_constant__ StaticScaleArgs A ) {
    const uint32_t col = blockIdx.x * blockDim.x + threadIdx.x;
    if( col >= A.t_width ) return;
    const int x = A.t_ox + (int) col;

    for( uint32_t row = blockIdx.y; row < A.t_height; row += gridDim.y ) {
        const int y = A.t_oy + (int) row;
        const int line = A.vertical ? (int) row : (int) col, cross = A.vertical ? x : y;
        float4 acc = make_float4( 0.0f, 0.0f, 0.0f, 0.0f );

        if( cross >= A.cross_lo && cross <= A.cross_hi ) {
            const int first = __ldg( A.start + line ), last = __ldg( A.start + line + 1 );
            for( int tap = first; tap < last; tap++ ) {
                const int s = __ldg( A.src + tap );
                const float c = __ldg( A.coef + tap );
                const size_t index = A.vertical ? (size_t)(s - A.s_oy) * (size_t) A.s_pitch + (size_t)(x - A.s_ox)
                                                : (size_t)(y - A.s_oy) * (size_t) A.s_pitch + (size_t)(s - A.s_ox);
                const float4 in = A.source[index];        /* plain load: the intermediate frame is written by the pass before */
                acc.x = acc.x + in.x * c;
                acc.y = acc.y + in.y * c;
                acc.z = acc.z + in.z * c;
                acc.w = acc.w + in.w * c;
            }
        }
        A.target[(size_t) row * (size_t) A.t_pitch + col] = acc;
    }
}

/* ------------------------------------------------------------- host: tap tables without malloc */

/* Every (target line, source line, coefficient) of one axis in the reference's order: upsampling walks the
 * source lines (video_scale.c:58-88), downsampling the target lines (:90-121).  pass 0 counts, pass 1 fills. */
static int walk_axis_static( HostTable *T, int pass, float tmin, int t_lo, int t_hi, float smin, int s_lo, int s_hi, float factor ) {
    float buffer[CV_STATIC_FILTER_TAPS];
    const int up = factor > 1.0f;
    const int outer_lo = up ? s_lo : t_lo, outer_hi = up ? s_hi : t_hi;

    for( int outer = outer_lo; outer <= outer_hi; outer++ ) {
        const float where = up ? ((float) outer - smin) * factor + tmin : ((float) outer - tmin) / factor + smin;
        const int base = (int) floorf( where );

        fir_filter filter;
        memset( &filter, 0, sizeof(filter) );
        filter.coeff = buffer; filter.width = CV_STATIC_FILTER_TAPS; filter.center = 0;
        filter_createTriangle( factor, where - (float) base, &filter );          /* given a buffer it does not allocate */
        if( filter.center < 0 && filter.width > CV_STATIC_FILTER_TAPS )
            return cvi_fail( CV_ERR_UNSUPPORTED, "cv_scale_bilinear_static_f32: factor %g needs %d taps per line (at most %d)",
                             (double) factor, filter.width, CV_STATIC_FILTER_TAPS );

        for( int tap = 0; tap < filter.width; tap++ ) {
            const int inner = base - filter.center + tap;
            const int t = up ? inner : outer, s = up ? outer : inner;
            if( t < t_lo || t > t_hi || s < s_lo || s > s_hi )
                continue;
            if( pass == 0 ) {
                T->start[t - t_lo + 1]++;
                if( t < T->used_lo ) T->used_lo = t;
                if( t > T->used_hi ) T->used_hi = t;
            }
            else {
                const int at = T->cursor[t - t_lo]++;
                T->src[at] = s;
                T->coef[at] = buffer[tap];
            }
        }
    }
    return CV_OK;
}

static int plan_axis_static( HostTable *T, float tmin, int t_lo, int t_hi, float smin, int s_lo, int s_hi, float factor ) {
    const int lines = t_hi >= t_lo ? t_hi - t_lo + 1 : 0;
    if( (unsigned) lines > CV_STATIC_MAX_LINES )
        return cvi_fail( CV_ERR_UNSUPPORTED, "cv_scale_bilinear_static_f32: %d target lines on one axis, the static scratch holds %u",
                         lines, CV_STATIC_MAX_LINES );
    T->used_lo = INT_MAX; T->used_hi = INT_MIN; T->total = 0;
    memset( T->start, 0, sizeof(int32_t) * (size_t)( lines + 2 ) );

    int rc = walk_axis_static( T, 0, tmin, t_lo, t_hi, smin, s_lo, s_hi, factor );
    if( rc ) return rc;

    long long sum = 0;
    for( int i = 0; i < lines; i++ ) {
        sum += T->start[i + 1];
        if( sum > (long long) CV_STATIC_MAX_TAPS )
            return cvi_fail( CV_ERR_UNSUPPORTED, "cv_scale_bilinear_static_f32: more than %u taps on one axis, beyond the static scratch",
                             CV_STATIC_MAX_TAPS );
        T->start[i + 1] = (int32_t) sum;
    }
    T->total = (int) sum;
    memcpy( T->cursor, T->start, sizeof(int32_t) * (size_t)( lines + 1 ) );
    return walk_axis_static( T, 1, tmin, t_lo, t_hi, smin, s_lo, s_hi, factor );
}

/* ------------------------------------------------------------- host: one axis, then both */

static int imax2( int a, int b ) { return a > b ? a : b; }
static int imin2( int a, int b ) { return a < b ? a : b; }

/* copy_frame with alpha 1 (video_mix.c:40-78 as cv_copy_frame_alpha_f32 does it for alpha == 1): the window both
 * frames share is copied, the rest of the target is left alone */
static int copy_window_static( rgba_frame_f32 *target, const rgba_frame_f32 *source, cudaStream_t stream ) {
    if( target->data == source->data ) return CV_OK;          /* video_mix.c:76-78 */
    box2i inner;
    box2i_intersect( &inner, &source->current_window, &target->full_window );
    target->current_window = inner;
    if( box2i_is_empty( &inner ) ) { box2i_set_empty( &target->current_window ); return CV_OK; }

    const size_t t_pitch = (size_t)( target->full_window.max.x - target->full_window.min.x + 1 );
    const size_t s_pitch = (size_t)( source->full_window.max.x - source->full_window.min.x + 1 );
    const rgba_f32 *from = source->data + (size_t)( inner.min.y - source->full_window.min.y ) * s_pitch + (size_t)( inner.min.x - source->full_window.min.x );
    rgba_f32 *to = target->data + (size_t)( inner.min.y - target->full_window.min.y ) * t_pitch + (size_t)( inner.min.x - target->full_window.min.x );
    return cvi_cuda( cudaMemcpy2DAsync( to, t_pitch * sizeof(rgba_f32), from, s_pitch * sizeof(rgba_f32),
                                        (size_t)( inner.max.x - inner.min.x + 1 ) * sizeof(rgba_f32), (size_t)( inner.max.y - inner.min.y + 1 ),
                                        cudaMemcpyDeviceToDevice, stream ), "cudaMemcpy2DAsync(static scale copy)" );
}

/* one axis: the whole target full_window is written from source->current_window; `slot` picks the host table and
 * the table region of the static array (the two passes of one call use different ones) */
static int scale_axis_static( rgba_frame_f32 *target, float tmin, const rgba_frame_f32 *source, float smin, float factor, int vertical,
                              unsigned char *scratch, int slot, cudaStream_t stream ) {
    const box2i s = source->current_window, t = target->full_window;
    const int tw = t.max.x - t.min.x + 1, th = t.max.y - t.min.y + 1;

    if( tw <= 0 || th <= 0 ) { box2i_set_empty( &target->current_window ); return CV_OK; }

    if( factor == 1.0f && tmin == smin ) {            /* video_scale.c:44-47: zero fill, then a plain windowed copy */
        int rc = cvi_cuda( cudaMemsetAsync( target->data, 0, sizeof(rgba_f32) * (size_t) tw * (size_t) th, stream ), "cudaMemsetAsync" );
        if( rc ) { box2i_set_empty( &target->current_window ); return rc; }
        return copy_window_static( target, source, stream );
    }

    const int cross_lo = vertical ? imax2( s.min.x, t.min.x ) : imax2( s.min.y, t.min.y );
    const int cross_hi = vertical ? imin2( s.max.x, t.max.x ) : imin2( s.max.y, t.max.y );

    HostTable *T = &g_host_table[slot];
    int rc = vertical ? plan_axis_static( T, tmin, t.min.y, t.max.y, smin, s.min.y, s.max.y, factor )
                      : plan_axis_static( T, tmin, t.min.x, t.max.x, smin, s.min.x, s.max.x, factor );
    if( rc ) { box2i_set_empty( &target->current_window ); return rc; }

    const int lines = vertical ? th : tw;
    unsigned char *table = scratch + CV_STATIC_TEMP_BYTES + (size_t) slot * CV_STATIC_TABLE_BYTES;
    int32_t *d_start = (int32_t *) table;
    int32_t *d_src = (int32_t *)( table + CV_STATIC_START_BYTES );
    float *d_coef = (float *)( table + CV_STATIC_START_BYTES + CV_STATIC_TAP_BYTES );

    /* pageable source: staged by the time each call returns, so the host table is free for the next call */
    rc = cvi_cuda( cudaMemcpyAsync( d_start, T->start, sizeof(int32_t) * (size_t)( lines + 1 ), cudaMemcpyHostToDevice, stream ), "cudaMemcpyAsync(static taps)" );
    if( !rc && T->total ) {
        rc = cvi_cuda( cudaMemcpyAsync( d_src, T->src, sizeof(int32_t) * (size_t) T->total, cudaMemcpyHostToDevice, stream ), "cudaMemcpyAsync(static taps)" );
        if( !rc ) rc = cvi_cuda( cudaMemcpyAsync( d_coef, T->coef, sizeof(float) * (size_t) T->total, cudaMemcpyHostToDevice, stream ), "cudaMemcpyAsync(static taps)" );
    }

    if( !rc ) {
        StaticScaleArgs A;
        memset( &A, 0, sizeof(A) );
        A.target = (float4 *) target->data; A.t_pitch = tw; A.t_ox = t.min.x; A.t_oy = t.min.y;
        A.t_width = (uint32_t) tw; A.t_height = (uint32_t) th;
        A.source = (const float4 *) source->data;
        A.s_pitch = source->full_window.max.x - source->full_window.min.x + 1;
        A.s_ox = source->full_window.min.x; A.s_oy = source->full_window.min.y;
        A.vertical = vertical; A.cross_lo = cross_lo; A.cross_hi = cross_hi;
        A.start = d_start; A.src = d_src; A.coef = d_coef;

        const unsigned bx = ( (unsigned) tw + 255u ) / 256u;
        unsigned by = (unsigned) th;
        const unsigned wave = 148u * 8u;                          /* rows beyond a few waves are strided */
        if( bx * by > 4u * wave ) by = imax2( 1, (int)( 4u * wave / bx ) );
        k_scale_axis_static<<<dim3( bx, by ), 256, 0, stream>>>( A );
        cvi_count_launch( 1 );
        rc = cvi_cuda( cudaGetLastError(), "k_scale_axis_static launch" );
    }

    if( rc ) box2i_set_empty( &target->current_window );
    else if( vertical ) box2i_set( &target->current_window, cross_lo, T->used_lo, cross_hi, T->used_hi );
    else if( cross_lo <= cross_hi ) box2i_set( &target->current_window, T->used_lo, cross_lo, T->used_hi, cross_hi );
    else box2i_set( &target->current_window, INT_MAX, cross_lo, INT_MIN, cross_hi );      /* video_scale.c:139-140 untouched by an empty row loop */
    return rc;
}

/* ------------------------------------------------------------- the entry point */

extern "C" int cv_scale_bilinear_static_f32( rgba_frame_f32 *target, v2f target_point, const rgba_frame_f32 *source, v2f source_point,
                                             v2f factors, cv_stream stream ) {
    if( !target || !source ) return cvi_fail( CV_ERR_ARGUMENT, "cv_scale_bilinear_static_f32: NULL argument" );
    if( !cvi_device() ) { box2i_set_empty( &target->current_window ); return CV_ERR_NOT_INITIALISED; }
    if( !(factors.x > 0.0f) || !(factors.y > 0.0f) ) {
        box2i_set_empty( &target->current_window );
        return cvi_fail( CV_ERR_ARGUMENT, "cv_scale_bilinear_static_f32: factors must be positive" );
    }
    cudaStream_t s = cvi_stream( stream );

    unsigned char *scratch = NULL;
    int rc = scratch_for_current_device( &scratch );
    if( rc ) { box2i_set_empty( &target->current_window ); return rc; }

    /* video_scale.c:231-245: an identity axis is skipped */
    if( factors.x == 1.0f && target_point.x == source_point.x ) {
        if( factors.y == 1.0f && target_point.y == source_point.y )
            return copy_window_static( target, source, s );
        return scale_axis_static( target, target_point.y, source, source_point.y, factors.y, 1, scratch, 0, s );
    }
    if( factors.y == 1.0f && target_point.y == source_point.y )
        return scale_axis_static( target, target_point.x, source, source_point.x, factors.x, 0, scratch, 0, s );

    /* :247-284: the axis with the smaller factor first, through an intermediate frame whose window uses C's (int)
     * truncation -- here the head of the static array instead of a pooled block */
    const box2i tf = target->full_window, sc = source->current_window;
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

    const unsigned long long temp_pixels = (unsigned long long)( want.max.x - want.min.x + 1 ) * (unsigned long long)( want.max.y - want.min.y + 1 );
    if( temp_pixels > CV_STATIC_MAX_TEMP_PIXELS ) {
        box2i_set_empty( &target->current_window );
        return cvi_fail( CV_ERR_UNSUPPORTED, "cv_scale_bilinear_static_f32: the intermediate frame needs %llu pixels, the static scratch holds %u "
                         "(use cv_scale_bilinear_f32, which allocates)", temp_pixels, CV_STATIC_MAX_TEMP_PIXELS );
    }

    rgba_frame_f32 temp;
    memset( &temp, 0, sizeof(temp) );
    temp.data = (rgba_f32 *) scratch;                /* 256-byte aligned head of the array */
    temp.full_window = want;
    temp.current_window = want;

    if( x_first ) {
        rc = scale_axis_static( &temp, target_point.x, source, source_point.x, factors.x, 0, scratch, 0, s );
        if( !rc ) rc = scale_axis_static( target, target_point.y, &temp, source_point.y, factors.y, 1, scratch, 1, s );
    }
    else {
        rc = scale_axis_static( &temp, target_point.y, source, source_point.y, factors.y, 1, scratch, 0, s );
        if( !rc ) rc = scale_axis_static( target, target_point.x, &temp, source_point.x, factors.x, 0, scratch, 1, s );
    }
    if( rc ) box2i_set_empty( &target->current_window );
    return rc;
}
