/*
 * canvas_b200/csrc/cv_runtime.c -- device bring-up, streams/events, memory and the host-built
 * transfer tables behind include/canvas_b200.h.  Plain C over the CUDA runtime API.
 *
 * Tables follow /root/reference/src/cprocess/gammatab.c:8-38,48-66,83-250 and
 * widget_gl.c:949-969: they are computed HERE on the host with libm powf and uploaded, never
 * recomputed on the device, so the gathers return the reference's bits.
 */
#define _GNU_SOURCE
#include "cv_internal.h"

#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

static CvDevice g_devices[CV_MAX_DEVICES];
static pthread_mutex_t g_init_lock = PTHREAD_MUTEX_INITIALIZER;
static __thread char t_error[512];
static __thread int t_device = -1;
static uint64_t g_launches;

static half *g_host_lut[CV_LUT_COUNT];
static uint8_t *g_host_gamma45;
static float *g_host_to709_f32;
static half *g_host_lut709_compact;
static pthread_once_t g_tables_once = PTHREAD_ONCE_INIT;

/* ------------------------------------------------------------------ errors */

int cvi_fail( int code, const char *fmt, ... ) {
    va_list ap;
    va_start( ap, fmt );
    vsnprintf( t_error, sizeof(t_error), fmt, ap );
    va_end( ap );
    return code;
}

int cvi_cuda( cudaError_t e, const char *what ) {
    if( e == cudaSuccess )
        return CV_OK;
    return cvi_fail( CV_ERR_CUDA, "%s: %s", what, cudaGetErrorString( e ) );
}

const char *cv_last_error( void ) { return t_error; }
const char *cv_version( void ) { return "canvas-b200 0.1 (sm_100a)"; }

void cvi_count_launch( unsigned n ) { __atomic_fetch_add( &g_launches, (uint64_t) n, __ATOMIC_RELAXED ); }
uint64_t cv_kernel_launches( void ) { return __atomic_load_n( &g_launches, __ATOMIC_RELAXED ); }

/* ------------------------------------------------- half helpers (host side) */

/* half.c:47-51 / genhalf.py:25-57 as bit operations: truncation, subnormals kept */
static uint16_t host_f2h( float value ) {
    union { float f; uint32_t u; } v = { .f = value };
    const uint32_t sign = (v.u >> 16) & 0x8000u, mant = v.u & 0x007FFFFFu;
    const int e = (int)((v.u >> 23) & 0xFF) - 127;

    if( e < -24 ) return (uint16_t) sign;
    if( e < -14 ) return (uint16_t)(sign | ((0x0400u >> (-14 - e)) + (mant >> (-e - 1))));
    if( e <= 15 ) return (uint16_t)(sign | (((uint32_t)(e + 15) << 10) + (mant >> 13)));
    if( e < 128 ) return (uint16_t)(sign | 0x7C00u);
    return (uint16_t)(sign | (0x7C00u + (mant >> 13)));
}

/* half.c:31-37: exact */
static float host_h2f( uint16_t h ) {
    union { float f; uint32_t u; } v;
    const uint32_t sign = ((uint32_t) h & 0x8000u) << 16;
    uint32_t e = (h >> 10) & 0x1F, m = h & 0x3FF;

    if( e == 0 ) {
        if( m == 0 ) { v.u = sign; return v.f; }
        int shift = 0;
        while( !(m & 0x400) ) { m <<= 1; shift++; }
        v.u = sign | ((uint32_t)(113 - shift) << 23) | ((m & 0x3FF) << 13);
    }
    else if( e == 31 ) v.u = sign | 0x7F800000u | (m << 13);
    else v.u = sign | ((e + 112) << 23) | (m << 13);
    return v.f;
}

/* ------------------------------------------------------------------ tables */

static float tf_709_to_linear( float v ) {           /* gammatab.c:48-55 */
    if( v < 4.5f * 0.018f ) return v / 4.5f;
    return powf( (v + 0.099f) / 1.099f, 1.0f / 0.45f );
}
static float tf_linear_to_709( float v ) {           /* gammatab.c:58-66 */
    if( v < 0.018f ) return v * 4.5f;
    return 1.099f * powf( v, 0.45f ) - 0.099f;
}
static float tf_linear_to_srgb( float v ) {          /* gammatab.c:201-211 */
    const float a = 0.055;
    if( v <= 0.0031308f ) return v * 12.92f;
    return (1.0f + a) * powf( v, 1.0f / 2.4f ) - a;
}

static void build_tables( void ) {
    for( int which = 0; which < CV_LUT_COUNT; which++ ) {
        half *t = (half *) malloc( sizeof(half) * HALF_COUNT );
        for( int i = 0; i < HALF_COUNT; i++ ) {
            float v = host_h2f( (uint16_t) i );
            switch( which ) {
                case CV_LUT_709_TO_LINEAR_SCENE:   v = tf_709_to_linear( v ); break;
                case CV_LUT_709_TO_LINEAR_DISPLAY: v = (v < 0.0f) ? 0.0f : powf( v, 2.5f ); break;     /* gammatab.c:145-150 */
                case CV_LUT_LINEAR_TO_709:         v = tf_linear_to_709( v ); break;
                default:                           v = tf_linear_to_srgb( v ); break;
            }
            t[i] = host_f2h( v );
        }
        g_host_lut[which] = t;
    }

    /* The rec709->linear table restricted to what a Y'CbCr conversion can produce.  For any byte input and either
     * matrix the converted value lies in (-1.14, 2.15) (extremes: B = y + 1.8556 cb at Y = 255/0, Cb = 255/0;
     * tests/test_decode_tables.py walks the corners), so 0x4080 positive halves [0, 2.25) and 0x3F80 negative
     * halves (-1.875, -0] suffice: 32768 entries.  Layout: entry h for a positive half, entry 0x7FFF - m for the
     * negative half 0x8000 + m -- the index is h ^ 0xFFFF for negatives, one XOR with the replicated sign. */
    g_host_lut709_compact = (half *) malloc( sizeof(half) * 32768 );
    for( int i = 0; i < 32768; i++ )
        g_host_lut709_compact[i] = g_host_lut[CV_LUT_709_TO_LINEAR_SCENE][i < 0x4080 ? i : 0x8000 + (0x7FFF - i)];

    /* video_subsample.c:227-232, 258-263: the MPEG-2 shader's linear_to_rec709() on a half texel, kept in f32 */
    g_host_to709_f32 = (float *) malloc( sizeof(float) * HALF_COUNT );
    for( int i = 0; i < HALF_COUNT; i++ )
        g_host_to709_f32[i] = tf_linear_to_709( host_h2f( (uint16_t) i ) );

    /* gammatab.c:8-38: truncating (uint8_t) of the clamped value; NaN clamps to 0 */
    g_host_gamma45 = (uint8_t *) malloc( HALF_COUNT );
    for( int i = 0; i < HALF_COUNT; i++ ) {
        const float v = clampf( powf( host_h2f( (uint16_t) i ), 0.45f ) * 255.0f, 0.0f, 255.0f );
        g_host_gamma45[i] = (uint8_t)(int) v;
    }
}

const half *cv_host_lut( int which ) {
    if( which < 0 || which >= CV_LUT_COUNT ) return NULL;
    pthread_once( &g_tables_once, build_tables );
    return g_host_lut[which];
}

const uint8_t *cv_host_gamma45_ramp( void ) {
    pthread_once( &g_tables_once, build_tables );
    return g_host_gamma45;
}

int cv_build_intent_ramp( float rendering_intent, uint8_t *ramp ) {
    if( !ramp ) return cvi_fail( CV_ERR_ARGUMENT, "cv_build_intent_ramp: NULL ramp" );
    for( int i = 0; i < HALF_COUNT; i++ )
        ramp[i] = (uint8_t) lrintf( clampf( powf( host_h2f( (uint16_t) i ), rendering_intent ) * 255.0f, 0.0f, 255.0f ) );
    return CV_OK;
}

/* ------------------------------------------------------------------ devices */

int cv_device_count( void ) {
    int n = 0;
    if( cudaGetDeviceCount( &n ) != cudaSuccess ) { cudaGetLastError(); return 0; }
    return n > CV_MAX_DEVICES ? CV_MAX_DEVICES : n;
}

CvDevice *cvi_device( void ) {
    if( t_device < 0 || !g_devices[t_device].ready ) {
        cvi_fail( CV_ERR_NOT_INITIALISED, "canvas_b200: cv_init() has not succeeded on this thread's device (no CPU fallback exists)" );
        return NULL;
    }
    return &g_devices[t_device];
}

cudaStream_t cvi_stream( cv_stream s ) {
    if( s ) return (cudaStream_t) s;
    CvDevice *dev = cvi_device();
    return dev ? dev->stream : 0;
}

int cv_init( int device ) {
    const int count = cv_device_count();
    if( count <= 0 )
        return cvi_fail( CV_ERR_NO_DEVICE, "cv_init: no CUDA device (the sm_100a kernels have no CPU fallback)" );
    if( device < 0 || device >= count )
        return cvi_fail( CV_ERR_ARGUMENT, "cv_init: device %d out of range (0..%d)", device, count - 1 );

    int rc = cvi_cuda( cudaSetDevice( device ), "cudaSetDevice" );
    if( rc ) return rc;

    pthread_mutex_lock( &g_init_lock );
    CvDevice *dev = &g_devices[device];

    if( !dev->ready ) {
        struct cudaDeviceProp prop;
        rc = cvi_cuda( cudaGetDeviceProperties( &prop, device ), "cudaGetDeviceProperties" );
        if( !rc && prop.major != 10 )
            rc = cvi_fail( CV_ERR_NO_DEVICE, "cv_init: device %d is sm_%d%d; this library carries sm_100a code only", device, prop.major, prop.minor );

        if( !rc ) {
            memset( dev, 0, sizeof(*dev) );
            dev->index = device;
            dev->sm_count = prop.multiProcessorCount;
            dev->l2_bytes = (size_t) prop.l2CacheSize;
            pthread_mutex_init( &dev->lock, NULL );
            rc = cvi_cuda( cudaStreamCreateWithFlags( &dev->stream, cudaStreamNonBlocking ), "cudaStreamCreate" );
        }

        pthread_once( &g_tables_once, build_tables );

        for( int i = 0; !rc && i < CV_LUT_COUNT; i++ ) {
            rc = cvi_cuda( cudaMalloc( (void **) &dev->lut[i], sizeof(half) * HALF_COUNT ), "cudaMalloc(table)" );
            if( !rc ) rc = cvi_cuda( cudaMemcpy( dev->lut[i], g_host_lut[i], sizeof(half) * HALF_COUNT, cudaMemcpyHostToDevice ), "cudaMemcpy(table)" );
        }
        if( !rc ) rc = cvi_cuda( cudaMalloc( (void **) &dev->gamma45, HALF_COUNT ), "cudaMalloc(gamma45)" );
        if( !rc ) rc = cvi_cuda( cudaMemcpy( dev->gamma45, g_host_gamma45, HALF_COUNT, cudaMemcpyHostToDevice ), "cudaMemcpy(gamma45)" );
        if( !rc ) rc = cvi_cuda( cudaMalloc( (void **) &dev->lut709_compact, sizeof(half) * 32768 ), "cudaMalloc(compact table)" );
        if( !rc ) rc = cvi_cuda( cudaMemcpy( dev->lut709_compact, g_host_lut709_compact, sizeof(half) * 32768, cudaMemcpyHostToDevice ), "cudaMemcpy(compact table)" );
        if( !rc ) rc = cvi_cuda( cudaMalloc( (void **) &dev->to709_f32, sizeof(float) * HALF_COUNT ), "cudaMalloc(to709_f32)" );
        if( !rc ) rc = cvi_cuda( cudaMemcpy( dev->to709_f32, g_host_to709_f32, sizeof(float) * HALF_COUNT, cudaMemcpyHostToDevice ), "cudaMemcpy(to709_f32)" );

        if( !rc ) {
            dev->ready = 1;
            t_device = device;
            rc = cvi_kernels_init();
            if( rc ) dev->ready = 0;
        }
    }

    pthread_mutex_unlock( &g_init_lock );
    if( !rc ) t_device = device;
    return rc;
}

int cv_set_device( int device ) {
    if( device < 0 || device >= CV_MAX_DEVICES || !g_devices[device].ready )
        return cvi_fail( CV_ERR_NOT_INITIALISED, "cv_set_device: device %d has not been initialised with cv_init", device );
    int rc = cvi_cuda( cudaSetDevice( device ), "cudaSetDevice" );
    if( !rc ) t_device = device;
    return rc;
}

int cv_get_device( void ) { return t_device; }

void cv_shutdown( void ) {
    pthread_mutex_lock( &g_init_lock );
    for( int d = 0; d < CV_MAX_DEVICES; d++ ) {
        CvDevice *dev = &g_devices[d];
        if( !dev->ready ) continue;
        cudaSetDevice( d );
        cudaDeviceSynchronize();
        cvi_scale_cache_release( d );
        cvi_pool_release_all( d );
        for( int i = 0; i < CV_LUT_COUNT; i++ ) cudaFree( dev->lut[i] );
        cudaFree( dev->gamma45 );
        cudaFree( dev->to709_f32 );
        cudaFree( dev->lut709_compact );
        cudaFree( dev->flush_buf );
        cudaFree( dev->job_ring_dev );
        cudaFreeHost( dev->job_ring_host );
        cudaFree( dev->piece_ring_dev );
        cudaFreeHost( dev->piece_ring_host );
        if( dev->job_ring_events ) {
            for( int i = 0; i < 8; i++ ) cudaEventDestroy( dev->job_ring_events[i] );
            free( dev->job_ring_events );
        }
        cudaStreamDestroy( dev->stream );
        dev->ready = 0;
    }
    cvi_pinned_release_all();
    t_device = -1;
    pthread_mutex_unlock( &g_init_lock );
}

/* ------------------------------------------------------- streams and events */

int cv_stream_create( cv_stream *stream ) {
    if( !cvi_device() ) return CV_ERR_NOT_INITIALISED;
    cudaStream_t s;
    int rc = cvi_cuda( cudaStreamCreateWithFlags( &s, cudaStreamNonBlocking ), "cudaStreamCreate" );
    if( !rc ) *stream = (cv_stream) s;
    return rc;
}
int cv_stream_destroy( cv_stream stream ) { return cvi_cuda( cudaStreamDestroy( (cudaStream_t) stream ), "cudaStreamDestroy" ); }
int cv_stream_sync( cv_stream stream ) {
    if( !cvi_device() ) return CV_ERR_NOT_INITIALISED;
    return cvi_cuda( cudaStreamSynchronize( cvi_stream( stream ) ), "cudaStreamSynchronize" );
}
int cv_device_sync( void ) {
    if( !cvi_device() ) return CV_ERR_NOT_INITIALISED;
    return cvi_cuda( cudaDeviceSynchronize(), "cudaDeviceSynchronize" );
}

int cv_event_create( cv_event *event ) {
    if( !cvi_device() ) return CV_ERR_NOT_INITIALISED;
    cudaEvent_t e;
    int rc = cvi_cuda( cudaEventCreate( &e ), "cudaEventCreate" );
    if( !rc ) *event = (cv_event) e;
    return rc;
}
int cv_event_destroy( cv_event event ) { return cvi_cuda( cudaEventDestroy( (cudaEvent_t) event ), "cudaEventDestroy" ); }
int cv_event_record( cv_event event, cv_stream stream ) {
    if( !cvi_device() ) return CV_ERR_NOT_INITIALISED;
    return cvi_cuda( cudaEventRecord( (cudaEvent_t) event, cvi_stream( stream ) ), "cudaEventRecord" );
}
int cv_event_sync( cv_event event ) { return cvi_cuda( cudaEventSynchronize( (cudaEvent_t) event ), "cudaEventSynchronize" ); }
int cv_event_elapsed_ms( cv_event start, cv_event stop, float *ms ) {
    return cvi_cuda( cudaEventElapsedTime( ms, (cudaEvent_t) start, (cudaEvent_t) stop ), "cudaEventElapsedTime" );
}
