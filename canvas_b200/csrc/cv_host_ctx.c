/*
 * canvas_b200/csrc/cv_host_ctx.c -- the per-thread context of the framework.h entry points.
 *
 * The reference's threading contract (/root/reference/src/process/VideoPullQueue.c:99-110, widget_gl.c:280):
 * get_frame may be called concurrently from any number of threads, none of them special.  Every calling thread
 * therefore gets its OWN stream and its own ring of pinned staging slabs on first use, and nothing on this path waits
 * for another thread's work: a pull from the widget thread and the pull queue's two threads overlap their copies and
 * kernels on the device instead of queueing behind one shared stream.
 *
 *   - the slabs (pinned blocks from the pool of cv_memory.c, returned to it when the thread ends) are where pulled sources write their pixels (video_get_frame_f16(source, ..., &frame_in_slab)), so the
 *     host->device copy that follows is a true asynchronous DMA out of pinned memory and the thread goes on pulling the
 *     next layer while it runs; a slab is reused only after the event recorded behind its last copy has completed;
 *   - device frames come from the per-device pool and go back to it stream-ordered (cvi_pool_free_async), so a block
 *     is never handed to another stream while this thread's kernels may still touch it.
 */
#define _GNU_SOURCE
#include "cv_internal.h"

#include <stdlib.h>
#include <string.h>

static pthread_key_t g_ctx_key;
static pthread_once_t g_ctx_once = PTHREAD_ONCE_INIT;

static void ctx_destroy( void *p ) {
    CvHostCtx *ctx = (CvHostCtx *) p;
    if( !ctx ) return;
    if( cudaSetDevice( ctx->device ) == cudaSuccess ) {
        if( ctx->stream ) cudaStreamSynchronize( ctx->stream );
        for( int i = 0; i < CV_HOST_SLABS; i++ ) {
            if( ctx->slab[i].ptr ) cvi_pinned_free( ctx->slab[i].ptr );       /* back to the pool: the next thread's first pull finds it */
            if( ctx->slab[i].done ) cudaEventDestroy( ctx->slab[i].done );
        }
        if( ctx->stream ) cudaStreamDestroy( ctx->stream );
    }
    cudaGetLastError();
    free( ctx );
}

static void ctx_key_create( void ) { pthread_key_create( &g_ctx_key, ctx_destroy ); }

CvHostCtx *cvi_host_ctx( void ) {
    pthread_once( &g_ctx_once, ctx_key_create );
    CvHostCtx *ctx = (CvHostCtx *) pthread_getspecific( g_ctx_key );
    if( ctx ) {
        if( cv_get_device() != ctx->device && cv_set_device( ctx->device ) ) return NULL;
        return ctx;
    }

    if( cv_get_device() < 0 ) {
        const char *env = getenv( "CANVAS_B200_DEVICE" );
        if( cv_init( env ? atoi( env ) : 0 ) != CV_OK )
            return NULL;
    }

    ctx = (CvHostCtx *) calloc( 1, sizeof(*ctx) );
    if( !ctx ) { cvi_fail( CV_ERR_MEMORY, "out of host memory" ); return NULL; }
    ctx->device = cv_get_device();
    if( cvi_cuda( cudaStreamCreateWithFlags( &ctx->stream, cudaStreamNonBlocking ), "cudaStreamCreate(host context)" ) ) {
        free( ctx );
        return NULL;
    }
    pthread_setspecific( g_ctx_key, ctx );
    return ctx;
}

void *cvi_host_slab( CvHostCtx *ctx, size_t bytes, int *index ) {
    const int i = ctx->next;
    ctx->next = (ctx->next + 1) % CV_HOST_SLABS;
    if( index ) *index = i;

    if( ctx->slab[i].done && ctx->slab[i].busy ) {
        cudaEventSynchronize( ctx->slab[i].done );         /* the copy that last used this slab */
        ctx->slab[i].busy = 0;
    }
    if( ctx->slab[i].bytes < bytes ) {
        if( ctx->slab[i].ptr ) cvi_pinned_free( ctx->slab[i].ptr );
        ctx->slab[i].ptr = NULL;
        ctx->slab[i].bytes = 0;
        /* from the pool of pinned blocks (cv_memory.c), which outlives the thread: pinning 66 MB costs milliseconds and
         * serialises in the driver, and callers with short-lived threads would otherwise pay it on every first pull.
         * Grown in 1 MiB steps so that frames of one size share slabs */
        const size_t want = (bytes + ((size_t) 1 << 20) - 1) & ~(((size_t) 1 << 20) - 1);
        if( cvi_pinned_alloc( &ctx->slab[i].ptr, want ) ) {
            ctx->slab[i].ptr = NULL;
            return NULL;
        }
        ctx->slab[i].bytes = want;
    }
    if( !ctx->slab[i].done &&
            cvi_cuda( cudaEventCreateWithFlags( &ctx->slab[i].done, cudaEventDisableTiming ), "cudaEventCreate(staging slab)" ) )
        return NULL;
    return ctx->slab[i].ptr;
}

void cvi_host_slab_done( CvHostCtx *ctx, int index ) {
    if( index < 0 || index >= CV_HOST_SLABS || !ctx->slab[index].done ) return;
    if( cudaEventRecord( ctx->slab[index].done, ctx->stream ) == cudaSuccess )
        ctx->slab[index].busy = 1;
}
