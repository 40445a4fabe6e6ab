/*
 * canvas_b200/csrc/cv_host_framework.c -- the reference's own C API (include/framework.h,
 * include/half.h) implemented on top of the device C-ABI.  This is the drop-in boundary for C
 * callers and for the fluggo.media.process extension: same symbols, same argument meaning, same
 * "empty current_window on failure" convention, but every pixel is produced by the sm_100a
 * kernels.  Frames passed here hold HOST pointers (as in the reference); they are moved to HBM,
 * processed and moved back on the CALLING THREAD's own stream (cv_host_ctx.c), so concurrent callers never
 * queue behind each other; frames this file pulls from a source are produced straight into pinned staging
 * slabs (true asynchronous DMA), frames handed in by the caller are copied from wherever they live (the
 * CUDA driver stages pageable memory itself).  Callers that can keep frames resident use include/canvas_b200.h.
 *
 * Replaces (paths relative to /root/reference/src/cprocess):
 *   half.c:87-105 (function-pointer binding), main.c:24-145 (time helpers, video_get_frame_*),
 *   video_mix.c:27-105 (copies) :46-71,107-370 (mixes), gammatab.c:8-38,83-250 (transfer entry
 *   points), filter.c:24-76,150-153, video_reconstruct.c:50-137, video_subsample.c:24-73,99-187.
 */
#define _GNU_SOURCE
#include "cv_internal.h"

#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

/* ------------------------------------------------------------------ bring-up */

/* the calling thread's stream (and through it the device), or NULL when no B200 can be used */
static CvHostCtx *host_ctx( void ) {
    CvHostCtx *ctx = cvi_host_ctx();
    if( !ctx ) {
        static int warned;
        if( !warned ) {
            warned = 1;
            fprintf( stderr, "canvas_b200: %s\n", cv_last_error() );
        }
    }
    return ctx;
}

static int host_ready( void ) { return host_ctx() != NULL; }
static cv_stream host_stream( void ) { CvHostCtx *ctx = cvi_host_ctx(); return ctx ? (cv_stream) ctx->stream : NULL; }

/* device scratch for one host<->device round trip */
typedef struct Scratch { void *dev; size_t bytes; } Scratch;

static int scratch_up( Scratch *s, const void *host, size_t bytes ) {
    s->dev = NULL;
    s->bytes = bytes;
    int rc = cv_malloc( &s->dev, bytes ? bytes : 1, CV_MEM_DEVICE );
    if( !rc && host && bytes )
        rc = cv_upload( s->dev, host, bytes, host_stream() );
    return rc;
}

static int scratch_down( Scratch *s, void *host ) {
    int rc = CV_OK;
    if( host && s->bytes )
        rc = cv_download( host, s->dev, s->bytes, host_stream() );
    if( !rc )
        rc = cv_stream_sync( host_stream() );
    return rc;
}

static void scratch_free( Scratch *s ) {
    if( s->dev ) {
        cv_free_async( s->dev, CV_MEM_DEVICE, host_stream() );
        s->dev = NULL;
    }
}

static size_t window_pixels( const box2i *w ) {
    v2i size;
    box2i_get_size( w, &size );
    return (size_t) size.x * (size_t) size.y;
}

/* ------------------------------------------------------------------ half.h */

static void gpu_half_from_float( half *out, const float *in, int count ) {
    if( count <= 0 || !host_ready() ) return;
    Scratch src = { 0 }, dst = { 0 };
    if( !scratch_up( &src, in, sizeof(float) * (size_t) count ) && !scratch_up( &dst, NULL, sizeof(half) * (size_t) count ) &&
            !cv_half_from_float( (half *) dst.dev, (const float *) src.dev, (size_t) count, host_stream() ) )
        scratch_down( &dst, out );
    scratch_free( &src );
    scratch_free( &dst );
}

static void gpu_half_to_float( float *out, const half *in, int count ) {
    if( count <= 0 || !host_ready() ) return;
    Scratch src = { 0 }, dst = { 0 };
    if( !scratch_up( &src, in, sizeof(half) * (size_t) count ) && !scratch_up( &dst, NULL, sizeof(float) * (size_t) count ) &&
            !cv_half_to_float( (float *) dst.dev, (const half *) src.dev, (size_t) count, host_stream() ) )
        scratch_down( &dst, out );
    scratch_free( &src );
    scratch_free( &dst );
}

static void gpu_half_lookup( const half *table, half *out, const half *in, int count ) {
    if( count <= 0 || !host_ready() ) return;
    Scratch tab = { 0 }, buf = { 0 };
    if( !scratch_up( &tab, table, sizeof(half) * HALF_COUNT ) && !scratch_up( &buf, in, sizeof(half) * (size_t) count ) &&
            !cv_half_lookup( (const half *) tab.dev, (half *) buf.dev, (const half *) buf.dev, (size_t) count, host_stream() ) )
        scratch_down( &buf, out );
    scratch_free( &tab );
    scratch_free( &buf );
}

void (*half_convert_from_float)( half *, const float *, int ) = gpu_half_from_float;
void (*half_convert_to_float)( float *, const half *, int ) = gpu_half_to_float;
void (*half_convert_from_float_fast)( half *, const float *, int ) = gpu_half_from_float;
void (*half_convert_to_float_fast)( float *, const half *, int ) = gpu_half_to_float;
void (*half_lookup)( const half *, half *, const half *, int ) = gpu_half_lookup;

void init_half( void ) {
    if( !host_ready() ) {
        fprintf( stderr, "canvas_b200: init_half(): no usable B200 and no CPU fallback exists -- aborting\n" );
        abort();
    }
    half_convert_from_float = half_convert_from_float_fast = gpu_half_from_float;
    half_convert_to_float = half_convert_to_float_fast = gpu_half_to_float;
    half_lookup = gpu_half_lookup;
}

/* ------------------------------------------------------- transfer functions */

static void transfer( int which, half *out, const half *in, size_t count ) {
    if( !count || !host_ready() ) return;
    Scratch buf = { 0 };
    if( !scratch_up( &buf, in, sizeof(half) * count ) && !cv_transfer( which, (half *) buf.dev, (const half *) buf.dev, count, host_stream() ) )
        scratch_down( &buf, out );
    scratch_free( &buf );
}

void video_transfer_rec709_to_linear_scene( half *out, const half *in, size_t count )   { transfer( CV_LUT_709_TO_LINEAR_SCENE, out, in, count ); }
void video_transfer_rec709_to_linear_display( half *out, const half *in, size_t count ) { transfer( CV_LUT_709_TO_LINEAR_DISPLAY, out, in, count ); }
void video_transfer_linear_to_rec709( half *out, const half *in, size_t count )         { transfer( CV_LUT_LINEAR_TO_709, out, in, count ); }
void video_transfer_linear_to_sRGB( half *out, const half *in, size_t count )           { transfer( CV_LUT_LINEAR_TO_SRGB, out, in, count ); }

const uint8_t *video_get_gamma45_ramp( void ) { return cv_host_gamma45_ramp(); }

/* ------------------------------------------------------------------ time */

int64_t get_frame_time( const rational *frame_rate, int frame ) {
    /* main.c:22-25: nanosecond at which `frame` starts, plus one so that get_time_frame maps it back */
    const int64_t ticks = (int64_t) frame * NS_PER_SEC * (int64_t) frame_rate->d;
    return ticks / (int64_t) frame_rate->n + 1;
}

int get_time_frame( const rational *frame_rate, int64_t time ) {
    return (int)( (time * (int64_t) frame_rate->n) / (NS_PER_SEC * (int64_t) frame_rate->d) );
}

int64_t gettime( void ) {
    struct timespec now;
    clock_gettime( CLOCK_MONOTONIC, &now );
    return (int64_t) now.tv_sec * NS_PER_SEC + now.tv_nsec;
}

/* ------------------------------------------------------------------ FIR taps */

void filter_createTriangle( float sub, float offset, fir_filter *filter ) {
    /* filter.c:24-76: taps of 1 - |t| sampled at the integer positions strictly inside
     * (offset - width, offset + width); normalised to unit sum when down-sampling */
    const float width = (sub < 1.0f) ? 1.0f / sub : sub;
    float lo = ceilf( offset - width ), hi = floorf( offset + width );
    if( lo == offset - width ) lo += 1.0f;
    if( hi == offset + width ) hi -= 1.0f;

    const int taps = (int) hi - (int) lo + 1;

    if( filter->coeff ) {
        if( filter->width < taps ) {
            filter->width = taps;
            filter->center = -1;
            return;
        }
    }
    else {
        filter->coeff = (float *) malloc( sizeof(float) * (size_t)(taps > 0 ? taps : 1) );
    }

    filter->width = taps;
    filter->center = -(int) lo;

    float sum = 0.0f;
    for( int i = 0; i < taps; i++ ) {
        filter->coeff[i] = 1.0f - fabsf( (1.0f / width) * ((float)(i - filter->center) - offset) );
        sum += filter->coeff[i];
    }

    if( sub < 1.0f && sum != 0.0f )
        for( int i = 0; i < taps; i++ )
            filter->coeff[i] /= sum;
}

void filter_free( fir_filter *filter ) {
    free( filter->coeff );
    filter->coeff = NULL;
}

/* ------------------------------------------------------------------ coded images */

static void coded_image_release( void *image ) { free( image ); }

static coded_image *coded_image_new( const int *strides, const int *line_counts, int count, int zero ) {
    if( count < 0 || count > CODED_IMAGE_MAX_PLANES ) return NULL;

    /* one block: the struct, then the planes back to back (video_subsample.c:24-73) */
    size_t total = sizeof(coded_image);
    for( int i = 0; i < count; i++ )
        total += (size_t) strides[i] * (size_t) line_counts[i];

    coded_image *image = (coded_image *)( zero ? calloc( 1, total ) : malloc( total ) );
    if( !image ) return NULL;
    if( !zero ) memset( image, 0, sizeof(coded_image) );

    uint8_t *cursor = (uint8_t *)( image + 1 );
    for( int i = 0; i < count; i++ ) {
        const size_t bytes = (size_t) strides[i] * (size_t) line_counts[i];
        image->data[i] = bytes ? cursor : NULL;
        image->stride[i] = strides[i];
        image->line_count[i] = line_counts[i];
        cursor += bytes;
    }
    image->free_func = coded_image_release;
    return image;
}

coded_image *coded_image_alloc( const int *strides, const int *line_counts, int count )  { return coded_image_new( strides, line_counts, count, 0 ); }
coded_image *coded_image_alloc0( const int *strides, const int *line_counts, int count ) { return coded_image_new( strides, line_counts, count, 1 ); }

/* ------------------------------------------------- host frame <-> device frame */

/* Rows of `win` between a host frame and a device frame that share full_window. */
static int copy_window( void *dst, const void *src, const box2i *full, const box2i *win, size_t px, enum cudaMemcpyKind kind ) {
    if( box2i_is_empty( win ) ) return CV_OK;
    v2i size, wsize;
    box2i_get_size( full, &size );
    box2i_get_size( win, &wsize );
    const size_t pitch = (size_t) size.x * px;
    const size_t offset = ((size_t)(win->min.y - full->min.y) * (size_t) size.x + (size_t)(win->min.x - full->min.x)) * px;
    return cvi_cuda( cudaMemcpy2DAsync( (char *) dst + offset, pitch, (const char *) src + offset, pitch,
        (size_t) wsize.x * px, (size_t) wsize.y, kind, cvi_stream( host_stream() ) ), "cudaMemcpy2DAsync(window)" );
}

/* A device twin of a host f16/f32 frame: same windows, pooled HBM, the valid pixels uploaded. */
static int twin_f16( rgba_frame_f16 *dev, const rgba_frame_f16 *host, int upload ) {
    int rc = cv_frame_alloc_f16( dev, &host->full_window, CV_MEM_DEVICE );
    if( rc ) return rc;
    box2i_intersect( &dev->current_window, &host->current_window, &host->full_window );
    if( upload )
        rc = copy_window( dev->data, host->data, &host->full_window, &dev->current_window, sizeof(rgba_f16), cudaMemcpyHostToDevice );
    return rc;
}

static int twin_f32( rgba_frame_f32 *dev, const rgba_frame_f32 *host, int upload ) {
    int rc = cv_frame_alloc_f32( dev, &host->full_window, CV_MEM_DEVICE );
    if( rc ) return rc;
    box2i_intersect( &dev->current_window, &host->current_window, &host->full_window );
    if( upload )
        rc = copy_window( dev->data, host->data, &host->full_window, &dev->current_window, sizeof(rgba_f32), cudaMemcpyHostToDevice );
    return rc;
}

static void release_f16( rgba_frame_f16 *dev ) {
    if( dev->data ) { cv_free_async( dev->data, CV_MEM_DEVICE, host_stream() ); dev->data = NULL; }
}
static void release_f32( rgba_frame_f32 *dev ) {
    if( dev->data ) { cv_free_async( dev->data, CV_MEM_DEVICE, host_stream() ); dev->data = NULL; }
}

/* result window + its pixels back into the caller's frame */
static int fetch_back_f16( rgba_frame_f16 *host, const rgba_frame_f16 *dev ) {
    host->current_window = dev->current_window;
    int rc = copy_window( host->data, dev->data, &host->full_window, &dev->current_window, sizeof(rgba_f16), cudaMemcpyDeviceToHost );
    if( !rc ) rc = cv_stream_sync( host_stream() );
    return rc;
}
static int fetch_back_f32( rgba_frame_f32 *host, const rgba_frame_f32 *dev ) {
    host->current_window = dev->current_window;
    int rc = copy_window( host->data, dev->data, &host->full_window, &dev->current_window, sizeof(rgba_f32), cudaMemcpyDeviceToHost );
    if( !rc ) rc = cv_stream_sync( host_stream() );
    return rc;
}

/* ---------------------------------------------------------- video_get_frame_* */

void video_get_frame_f16( video_source *source, int frame_index, rgba_frame_f16 *frame ) {
    if( !source || !source->funcs ) {
        box2i_set_empty( &frame->current_window );
        return;
    }

    if( source->funcs->get_frame ) {
        source->funcs->get_frame( source->obj, frame_index, frame );
        return;
    }

    if( !source->funcs->get_frame_32 || !host_ready() ) {
        /* only a GL slot (no graphics stack here), or no device: nothing can be produced */
        box2i_set_empty( &frame->current_window );
        return;
    }

    /* main.c:44-72: pull as f32 into a temporary of the same window (a pinned staging slab), then pack (truncating) */
    CvHostCtx *ctx = host_ctx();
    int slab = -1;
    rgba_frame_f32 temp;
    temp.data = (rgba_f32 *) cvi_host_slab( ctx, sizeof(rgba_f32) * (window_pixels( &frame->full_window ) + 1), &slab );
    temp.full_window = frame->full_window;
    temp.current_window = frame->full_window;
    if( !temp.data ) { box2i_set_empty( &frame->current_window ); return; }

    source->funcs->get_frame_32( source->obj, frame_index, &temp );

    rgba_frame_f32 dev32 = { 0 };
    rgba_frame_f16 dev16 = { 0 };
    int rc = CV_OK;

    if( !box2i_is_empty( &temp.current_window ) ) {
        rc = twin_f32( &dev32, &temp, 1 );
        cvi_host_slab_done( ctx, slab );
        if( !rc ) rc = cv_frame_alloc_f16( &dev16, &frame->full_window, CV_MEM_DEVICE );
        if( !rc ) rc = cv_frame_f32_to_f16( &dev16, &dev32, host_stream() );
        if( !rc ) rc = fetch_back_f16( frame, &dev16 );
    }

    if( rc || box2i_is_empty( &temp.current_window ) )
        frame->current_window = temp.current_window;
    if( rc )
        box2i_set_empty( &frame->current_window );

    release_f32( &dev32 );
    release_f16( &dev16 );
}

void video_get_frame_f32( video_source *source, int frame_index, rgba_frame_f32 *frame ) {
    if( !source || !source->funcs ) {
        box2i_set_empty( &frame->current_window );
        return;
    }

    if( source->funcs->get_frame_32 ) {
        source->funcs->get_frame_32( source->obj, frame_index, frame );
        return;
    }

    if( !source->funcs->get_frame || !host_ready() ) {
        box2i_set_empty( &frame->current_window );
        return;
    }

    /* main.c:116-141: pull as f16 (into a pinned staging slab), widen exactly */
    CvHostCtx *ctx = host_ctx();
    int slab = -1;
    rgba_frame_f16 temp;
    temp.data = (rgba_f16 *) cvi_host_slab( ctx, sizeof(rgba_f16) * (window_pixels( &frame->full_window ) + 1), &slab );
    temp.full_window = frame->full_window;
    temp.current_window = frame->full_window;
    if( !temp.data ) { box2i_set_empty( &frame->current_window ); return; }

    source->funcs->get_frame( source->obj, frame_index, &temp );

    rgba_frame_f16 dev16 = { 0 };
    rgba_frame_f32 dev32 = { 0 };
    int rc = CV_OK;

    if( !box2i_is_empty( &temp.current_window ) ) {
        rc = twin_f16( &dev16, &temp, 1 );
        cvi_host_slab_done( ctx, slab );
        if( !rc ) rc = cv_frame_alloc_f32( &dev32, &frame->full_window, CV_MEM_DEVICE );
        if( !rc ) rc = cv_frame_f16_to_f32( &dev32, &dev16, host_stream() );
        if( !rc ) rc = fetch_back_f32( frame, &dev32 );
    }

    if( rc || box2i_is_empty( &temp.current_window ) )
        frame->current_window = temp.current_window;
    if( rc )
        box2i_set_empty( &frame->current_window );

    release_f16( &dev16 );
    release_f32( &dev32 );
}

/* ------------------------------------------------------------------ copies */

void video_copy_frame_f16( rgba_frame_f16 *out, rgba_frame_f16 *in ) {
    /* video_mix.c:27-44: pure data movement between two host frames (no arithmetic to offload) */
    box2i inner;
    box2i_intersect( &inner, &out->full_window, &in->current_window );
    out->current_window = inner;
    if( box2i_is_empty( &inner ) ) return;

    v2i size;
    box2i_get_size( &inner, &size );
    for( int y = inner.min.y; y <= inner.max.y; y++ )
        memmove( video_get_pixel_f16( out, inner.min.x, y ), video_get_pixel_f16( in, inner.min.x, y ), sizeof(rgba_f16) * (size_t) size.x );
}

void video_copy_frame_alpha_f32( rgba_frame_f32 *out, rgba_frame_f32 *in, float alpha ) {
    alpha = clampf( alpha, 0.0f, 1.0f );

    if( out == in && alpha == 1.0f ) return;
    if( alpha == 0.0f ) { box2i_set_empty( &out->current_window ); return; }

    if( alpha == 1.0f ) {
        box2i inner;
        box2i_intersect( &inner, &out->full_window, &in->current_window );
        out->current_window = inner;
        if( box2i_is_empty( &inner ) ) return;
        v2i size;
        box2i_get_size( &inner, &size );
        for( int y = inner.min.y; y <= inner.max.y; y++ )
            memmove( video_get_pixel_f32( out, inner.min.x, y ), video_get_pixel_f32( in, inner.min.x, y ), sizeof(rgba_f32) * (size_t) size.x );
        return;
    }

    if( !host_ready() ) { box2i_set_empty( &out->current_window ); return; }

    rgba_frame_f32 din = { 0 }, dout = { 0 };
    int rc = twin_f32( &din, in, 1 );
    if( !rc ) rc = cv_frame_alloc_f32( &dout, &out->full_window, CV_MEM_DEVICE );
    if( !rc ) rc = cv_copy_frame_alpha_f32( &dout, &din, alpha, host_stream() );
    if( !rc ) rc = fetch_back_f32( out, &dout );
    if( rc ) box2i_set_empty( &out->current_window );
    release_f32( &din );
    release_f32( &dout );
}

/* ------------------------------------------------------------------ mixes */

void video_mix_cross_f32( rgba_frame_f32 *out, rgba_frame_f32 *a, rgba_frame_f32 *b, float mix_b ) {
    if( !host_ready() ) { box2i_set_empty( &out->current_window ); return; }

    rgba_frame_f32 da = { 0 }, db = { 0 }, dout = { 0 };
    int rc = twin_f32( &da, a, 1 );
    if( !rc ) rc = twin_f32( &db, b, 1 );
    if( !rc ) rc = cv_frame_alloc_f32( &dout, &out->full_window, CV_MEM_DEVICE );
    if( !rc ) rc = cv_mix_cross_f32( &dout, &da, &db, mix_b, host_stream() );
    if( !rc ) rc = fetch_back_f32( out, &dout );
    if( rc ) box2i_set_empty( &out->current_window );
    release_f32( &da );
    release_f32( &db );
    release_f32( &dout );
}

void video_mix_over_f32( rgba_frame_f32 *out, rgba_frame_f32 *b, float mix_b ) {
    if( !host_ready() ) { box2i_set_empty( &out->current_window ); return; }

    rgba_frame_f32 dout = { 0 }, db = { 0 };
    int rc = twin_f32( &dout, out, 1 );
    if( !rc ) rc = twin_f32( &db, b, 1 );
    if( !rc ) rc = cv_mix_over_f32( &dout, &db, mix_b, host_stream() );
    if( !rc ) rc = fetch_back_f32( out, &dout );
    if( rc ) box2i_set_empty( &out->current_window );
    release_f32( &dout );
    release_f32( &db );
}

void video_mix_cross_f32_pull( rgba_frame_f32 *out, video_source *a, int frame_a, video_source *b, int frame_b, float mix_b ) {
    /* video_mix.c:46-71 */
    mix_b = clampf( mix_b, 0.0f, 1.0f );

    if( mix_b == 0.0f ) { video_get_frame_f32( a, frame_a, out ); return; }
    if( mix_b == 1.0f ) { video_get_frame_f32( b, frame_b, out ); return; }

    rgba_frame_f32 temp;
    temp.data = (rgba_f32 *) malloc( sizeof(rgba_f32) * (window_pixels( &out->full_window ) + 1) );
    temp.full_window = out->full_window;
    if( !temp.data ) { box2i_set_empty( &out->current_window ); return; }

    video_get_frame_f32( a, frame_a, out );
    video_get_frame_f32( b, frame_b, &temp );
    video_mix_cross_f32( out, out, &temp, mix_b );
    free( temp.data );
}

/* ------------------------------------------------------------------ scaler */

void video_scale_bilinear_f32( rgba_frame_f32 *target, v2f target_point, rgba_frame_f32 *source, v2f source_point, v2f factors ) {
    /* video_scale.c:231-285; the scaler zero-fills the whole target, so all of full_window comes back */
    if( !host_ready() ) { box2i_set_empty( &target->current_window ); return; }

    rgba_frame_f32 dt = { 0 }, ds = { 0 };
    int rc = twin_f32( &dt, target, 0 );
    if( !rc ) rc = twin_f32( &ds, source, 1 );
    if( !rc ) rc = cv_scale_bilinear_f32( &dt, target_point, &ds, source_point, factors, host_stream() );
    if( !rc ) {
        const int identity = factors.x == 1.0f && factors.y == 1.0f && target_point.x == source_point.x && target_point.y == source_point.y;
        target->current_window = dt.current_window;
        rc = copy_window( target->data, dt.data, &target->full_window, identity ? &dt.current_window : &target->full_window,
                          sizeof(rgba_f32), cudaMemcpyDeviceToHost );
        if( !rc ) rc = cv_stream_sync( host_stream() );
    }
    if( rc ) box2i_set_empty( &target->current_window );
    release_f32( &dt );
    release_f32( &ds );
}

void video_scale_bilinear_f32_pull( rgba_frame_f32 *target, v2f target_point, video_source *source, int frame, box2i *source_rect, v2f source_point, v2f factors ) {
    /* video_scale.c:287-319: fetch just the part of the source the target can see, then scale */
    if( factors.x == 0.0f || factors.y == 0.0f ) {
        box2i_set_empty( &target->current_window );
        return;
    }
    if( factors.x == 1.0f && factors.y == 1.0f && target_point.x == source_point.x && target_point.y == source_point.y ) {
        video_get_frame_f32( source, frame, target );
        return;
    }

    rgba_frame_f32 temp;
    const box2i tf = target->full_window;
    box2i_set( &temp.full_window,
        (int)( source_point.x - (target_point.x - (float) tf.min.x) / factors.x ) - 1,
        (int)( source_point.y - (target_point.y - (float) tf.min.y) / factors.y ) - 1,
        (int)( source_point.x + ((float) tf.max.x - target_point.x) / factors.x ) + 1,
        (int)( source_point.y + ((float) tf.max.y - target_point.y) / factors.y ) + 1 );
    box2i_intersect( &temp.full_window, &temp.full_window, source_rect );

    temp.data = (rgba_f32 *) malloc( sizeof(rgba_f32) * (window_pixels( &temp.full_window ) + 1) );
    if( !temp.data ) { box2i_set_empty( &target->current_window ); return; }

    video_get_frame_f32( source, frame, &temp );
    video_scale_bilinear_f32( target, target_point, &temp, source_point, factors );
    free( temp.data );
}

/* ------------------------------------------------------------------ colour transforms */

static void color_in_place( rgba_frame_f16 *frame, int (*transform)( rgba_frame_f16 *, cv_stream ) ) {
    /* color.c:100-160 work in place on current_window and have no way to report failure */
    if( !frame || box2i_is_empty( &frame->current_window ) || !host_ready() ) return;
    rgba_frame_f16 dev = { 0 };
    const box2i keep = frame->current_window;
    int rc = twin_f16( &dev, frame, 1 );
    if( !rc ) rc = transform( &dev, host_stream() );
    if( !rc ) rc = fetch_back_f16( frame, &dev );
    frame->current_window = keep;
    release_f16( &dev );
}

void video_color_rgb_to_xyz_sdtv( rgba_frame_f16 *frame ) { color_in_place( frame, cv_color_rgb_to_xyz_sdtv ); }
void video_color_xyz_to_srgb( rgba_frame_f16 *frame )     { color_in_place( frame, cv_color_xyz_to_srgb ); }

/* ------------------------------------------------------------------ DV conversion */

enum { DV_WIDTH = 720, DV_HEIGHT = 480, DV_SUB_X = 4, DV_PIC_X = 0, DV_PIC_Y = -1 };

void video_reconstruct_dv( rgba_frame_f16 *frame, coded_image *planar ) {
    /* video_reconstruct.c:50-137: 720x480 planar 4:1:1, Rec.709, picture at (0,-1) */
    if( !planar || !host_ready() ) { box2i_set_empty( &frame->current_window ); return; }

    Scratch plane[3] = { { 0 } };
    cv_coded image;
    memset( &image, 0, sizeof(image) );
    int rc = CV_OK;

    for( int i = 0; i < 3 && !rc; i++ ) {
        const int lines = planar->line_count[i] < DV_HEIGHT ? planar->line_count[i] : DV_HEIGHT;
        rc = scratch_up( &plane[i], planar->data[i], (size_t) planar->stride[i] * (size_t)(lines > 0 ? lines : 0) );
        image.plane[i] = (uint8_t *) plane[i].dev;
        image.stride[i] = planar->stride[i];
    }
    image.layout = CV_LAYOUT_PLANAR; image.width = DV_WIDTH; image.height = DV_HEIGHT;
    image.sub_x = DV_SUB_X; image.matrix = CV_MATRIX_709; image.pic_x = DV_PIC_X; image.pic_y = DV_PIC_Y;

    rgba_frame_f16 dev = { 0 };
    if( !rc ) rc = cv_frame_alloc_f16( &dev, &frame->full_window, CV_MEM_DEVICE );
    if( !rc ) rc = cv_convert_ycbcr8( &dev, &image, host_stream() );
    if( !rc ) rc = fetch_back_f16( frame, &dev );
    if( rc ) box2i_set_empty( &frame->current_window );

    release_f16( &dev );
    for( int i = 0; i < 3; i++ ) scratch_free( &plane[i] );
}

coded_image *video_subsample_dv( rgba_frame_f16 *frame ) {
    /* video_subsample.c:99-187; like the reference, the linear->Rec.709 table is also applied in
     * place to the covered window of `frame` (:147) */
    const int strides[3] = { DV_WIDTH, DV_WIDTH / DV_SUB_X, DV_WIDTH / DV_SUB_X };
    const int lines[3] = { DV_HEIGHT, DV_HEIGHT, DV_HEIGHT };
    coded_image *planar = coded_image_alloc0( strides, lines, 3 );
    if( !planar || !host_ready() ) return planar;

    Scratch plane[3] = { { 0 } };
    cv_coded image;
    memset( &image, 0, sizeof(image) );
    int rc = CV_OK;

    for( int i = 0; i < 3 && !rc; i++ ) {
        rc = scratch_up( &plane[i], planar->data[i], (size_t) strides[i] * lines[i] );      /* zeros */
        image.plane[i] = (uint8_t *) plane[i].dev;
        image.stride[i] = strides[i];
    }
    image.layout = CV_LAYOUT_PLANAR; image.width = DV_WIDTH; image.height = DV_HEIGHT;
    image.sub_x = DV_SUB_X; image.matrix = CV_MATRIX_709; image.pic_x = DV_PIC_X; image.pic_y = DV_PIC_Y;

    rgba_frame_f16 dev = { 0 };
    if( !rc ) rc = twin_f16( &dev, frame, 1 );
    if( !rc ) rc = cv_subsample_ycbcr8( &dev, &image, host_stream() );
    for( int i = 0; i < 3 && !rc; i++ )
        rc = cv_download( planar->data[i], plane[i].dev, (size_t) strides[i] * lines[i], host_stream() );
    if( !rc ) {
        const box2i keep = frame->current_window;
        rc = fetch_back_f16( frame, &dev );
        frame->current_window = keep;
    }

    release_f16( &dev );
    for( int i = 0; i < 3; i++ ) scratch_free( &plane[i] );
    return planar;
}
