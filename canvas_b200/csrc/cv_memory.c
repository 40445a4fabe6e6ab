/*
 * canvas_b200/csrc/cv_memory.c -- device / pinned memory, the pooled frame allocator and the
 * host<->device frame copies of include/canvas_b200.h.
 *
 * The reference allocates every frame with PyMem_Malloc (src/process/RgbaFrameF16.c:205) and
 * every temporary with g_slice_alloc per call (src/cprocess/main.c:50,122, video_mix.c:62,
 * workspace.c:538).  Here frames live in HBM (or pinned host memory for staging) and come from
 * a per-device pool, so steady-state rendering performs no cudaMalloc/cudaFree at all.
 */
#include "cv_internal.h"

#include <stdlib.h>
#include <string.h>

typedef struct PoolBlock { void *ptr; size_t bytes; int in_use; int pending; cudaEvent_t event; } PoolBlock;

typedef struct Pool {
    PoolBlock *blocks;
    int count, capacity;
} Pool;

static Pool g_pools[CV_MAX_DEVICES];

#define POOL_GRANULE ((size_t) 1 << 16)        /* sizes are rounded to 64 KiB so equal-sized frames share blocks */

int cvi_pool_alloc( void **ptr, size_t bytes ) {
    CvDevice *dev = cvi_device();
    if( !dev ) return CV_ERR_NOT_INITIALISED;
    if( bytes == 0 ) bytes = 1;
    bytes = (bytes + POOL_GRANULE - 1) & ~(POOL_GRANULE - 1);

    Pool *pool = &g_pools[dev->index];
    pthread_mutex_lock( &dev->lock );

    /* the smallest fitting free block whose last user has finished; a block still waiting for its stream-ordered free
     * (cvi_pool_free_async) is taken only when nothing else fits, and then waited for */
    int best = -1, best_pending = -1;
    for( int i = 0; i < pool->count; i++ ) {
        PoolBlock *b = &pool->blocks[i];
        if( b->in_use || b->bytes < bytes || b->bytes > bytes + bytes / 4 )
            continue;
        if( b->pending && cudaEventQuery( b->event ) == cudaSuccess )
            b->pending = 0;
        if( !b->pending ) {
            if( best < 0 || b->bytes < pool->blocks[best].bytes ) best = i;
        }
        else if( best_pending < 0 || b->bytes < pool->blocks[best_pending].bytes )
            best_pending = i;
    }
    cudaGetLastError();             /* cudaEventQuery reports cudaErrorNotReady through the sticky-free error slot */

    if( best < 0 && best_pending >= 0 ) {
        cudaEventSynchronize( pool->blocks[best_pending].event );
        pool->blocks[best_pending].pending = 0;
        best = best_pending;
    }

    if( best >= 0 ) {
        pool->blocks[best].in_use = 1;
        *ptr = pool->blocks[best].ptr;
        pthread_mutex_unlock( &dev->lock );
        return CV_OK;
    }

    void *fresh = NULL;
    cudaError_t e = cudaMalloc( &fresh, bytes );

    if( e != cudaSuccess ) {
        /* give cached blocks back and retry once */
        cudaGetLastError();
        for( int i = 0; i < pool->count; ) {
            if( !pool->blocks[i].in_use ) {
                if( pool->blocks[i].pending ) cudaEventSynchronize( pool->blocks[i].event );
                if( pool->blocks[i].event ) cudaEventDestroy( pool->blocks[i].event );
                cudaFree( pool->blocks[i].ptr );
                pool->blocks[i] = pool->blocks[--pool->count];
            }
            else i++;
        }
        e = cudaMalloc( &fresh, bytes );
    }

    if( e != cudaSuccess ) {
        pthread_mutex_unlock( &dev->lock );
        cudaGetLastError();
        return cvi_fail( CV_ERR_MEMORY, "cv pool: cudaMalloc(%zu) failed: %s", bytes, cudaGetErrorString( e ) );
    }

    if( pool->count == pool->capacity ) {
        pool->capacity = pool->capacity ? pool->capacity * 2 : 64;
        pool->blocks = (PoolBlock *) realloc( pool->blocks, sizeof(PoolBlock) * pool->capacity );
    }
    pool->blocks[pool->count].ptr = fresh;
    pool->blocks[pool->count].bytes = bytes;
    pool->blocks[pool->count].in_use = 1;
    pool->blocks[pool->count].pending = 0;
    pool->blocks[pool->count].event = NULL;
    pool->count++;
    *ptr = fresh;
    pthread_mutex_unlock( &dev->lock );
    return CV_OK;
}

int cvi_pool_free( void *ptr ) {
    if( !ptr ) return CV_OK;
    CvDevice *dev = cvi_device();
    if( !dev ) return CV_ERR_NOT_INITIALISED;
    Pool *pool = &g_pools[dev->index];
    int rc = CV_ERR_ARGUMENT;

    pthread_mutex_lock( &dev->lock );
    for( int i = 0; i < pool->count; i++ ) {
        if( pool->blocks[i].ptr == ptr ) {
            pool->blocks[i].in_use = 0;
            rc = CV_OK;
            break;
        }
    }
    pthread_mutex_unlock( &dev->lock );

    if( rc ) return cvi_fail( rc, "cv pool: %p was not allocated by this device's pool", ptr );
    return rc;
}

int cvi_pool_free_async( void *ptr, cudaStream_t stream ) {
    if( !ptr ) return CV_OK;
    CvDevice *dev = cvi_device();
    if( !dev ) return CV_ERR_NOT_INITIALISED;
    Pool *pool = &g_pools[dev->index];
    int rc = CV_ERR_ARGUMENT;

    pthread_mutex_lock( &dev->lock );
    for( int i = 0; i < pool->count; i++ ) {
        PoolBlock *b = &pool->blocks[i];
        if( b->ptr != ptr ) continue;
        rc = CV_OK;
        if( !b->event && cudaEventCreateWithFlags( &b->event, cudaEventDisableTiming ) != cudaSuccess )
            b->event = NULL;
        if( b->event && cudaEventRecord( b->event, stream ) == cudaSuccess )
            b->pending = 1;
        else {
            cudaGetLastError();
            cudaStreamSynchronize( stream );        /* no event: order the free by waiting */
            b->pending = 0;
        }
        b->in_use = 0;
        break;
    }
    pthread_mutex_unlock( &dev->lock );

    if( rc ) return cvi_fail( rc, "cv pool: %p was not allocated by this device's pool", ptr );
    return rc;
}

void cvi_pool_release_all( int device ) {
    Pool *pool = &g_pools[device];
    for( int i = 0; i < pool->count; i++ ) {
        if( pool->blocks[i].event ) cudaEventDestroy( pool->blocks[i].event );
        cudaFree( pool->blocks[i].ptr );
    }
    free( pool->blocks );
    memset( pool, 0, sizeof(*pool) );
}

/* ------------------------------------------------------------ pooled pinned staging memory */

/* cudaMallocHost costs milliseconds per call: the staging buffers of the timeline workers are kept in a process-wide
 * pool (same size classes as the device pool) and handed out again by the next cv_timeline_render() call. */
typedef struct PinnedBlock { void *ptr; size_t bytes; int in_use; } PinnedBlock;
static PinnedBlock *g_pinned;
static int g_pinned_count, g_pinned_capacity;
static pthread_mutex_t g_pinned_lock = PTHREAD_MUTEX_INITIALIZER;

int cvi_pinned_alloc( void **ptr, size_t bytes ) {
    if( bytes == 0 ) bytes = 1;
    bytes = (bytes + POOL_GRANULE - 1) & ~(POOL_GRANULE - 1);
    pthread_mutex_lock( &g_pinned_lock );
    int best = -1;
    for( int i = 0; i < g_pinned_count; i++ )
        if( !g_pinned[i].in_use && g_pinned[i].bytes >= bytes && g_pinned[i].bytes <= bytes + bytes / 4 &&
                (best < 0 || g_pinned[i].bytes < g_pinned[best].bytes) )
            best = i;
    if( best >= 0 ) {
        g_pinned[best].in_use = 1;
        *ptr = g_pinned[best].ptr;
        pthread_mutex_unlock( &g_pinned_lock );
        return CV_OK;
    }
    pthread_mutex_unlock( &g_pinned_lock );

    void *fresh = NULL;
    int rc = cvi_cuda( cudaMallocHost( &fresh, bytes ), "cudaMallocHost(staging)" );
    if( rc ) return rc;
    pthread_mutex_lock( &g_pinned_lock );
    if( g_pinned_count == g_pinned_capacity ) {
        g_pinned_capacity = g_pinned_capacity ? g_pinned_capacity * 2 : 64;
        g_pinned = (PinnedBlock *) realloc( g_pinned, sizeof(PinnedBlock) * (size_t) g_pinned_capacity );
    }
    g_pinned[g_pinned_count].ptr = fresh; g_pinned[g_pinned_count].bytes = bytes; g_pinned[g_pinned_count].in_use = 1;
    g_pinned_count++;
    pthread_mutex_unlock( &g_pinned_lock );
    *ptr = fresh;
    return CV_OK;
}

void cvi_pinned_free( void *ptr ) {
    if( !ptr ) return;
    pthread_mutex_lock( &g_pinned_lock );
    for( int i = 0; i < g_pinned_count; i++ )
        if( g_pinned[i].ptr == ptr ) { g_pinned[i].in_use = 0; break; }
    pthread_mutex_unlock( &g_pinned_lock );
}

void cvi_pinned_release_all( void ) {
    /* blocks in use belong to live per-thread contexts (their staging slabs): they stay listed and are pooled again
     * when their thread ends */
    pthread_mutex_lock( &g_pinned_lock );
    int kept = 0;
    for( int i = 0; i < g_pinned_count; i++ ) {
        if( g_pinned[i].in_use ) g_pinned[kept++] = g_pinned[i];
        else cudaFreeHost( g_pinned[i].ptr );
    }
    g_pinned_count = kept;
    if( !kept ) {
        free( g_pinned );
        g_pinned = NULL;
        g_pinned_capacity = 0;
    }
    pthread_mutex_unlock( &g_pinned_lock );
}

/* ------------------------------------------------------------ raw memory */

int cv_malloc( void **ptr, size_t bytes, int where ) {
    if( !ptr ) return cvi_fail( CV_ERR_ARGUMENT, "cv_malloc: NULL result pointer" );
    if( !cvi_device() ) return CV_ERR_NOT_INITIALISED;
    if( where == CV_MEM_PINNED )
        return cvi_cuda( cudaMallocHost( ptr, bytes ? bytes : 1 ), "cudaMallocHost" );
    return cvi_pool_alloc( ptr, bytes );
}

int cv_free( void *ptr, int where ) {
    if( !ptr ) return CV_OK;
    if( where == CV_MEM_PINNED )
        return cvi_cuda( cudaFreeHost( ptr ), "cudaFreeHost" );
    return cvi_pool_free( ptr );
}

int cv_free_async( void *ptr, int where, cv_stream stream ) {
    if( !ptr ) return CV_OK;
    if( where == CV_MEM_PINNED ) {
        int rc = cv_stream_sync( stream );
        return rc ? rc : cvi_cuda( cudaFreeHost( ptr ), "cudaFreeHost" );
    }
    if( !cvi_device() ) return CV_ERR_NOT_INITIALISED;
    return cvi_pool_free_async( ptr, cvi_stream( stream ) );
}

int cv_memset( void *dev, int byte, size_t bytes, cv_stream stream ) {
    if( !cvi_device() ) return CV_ERR_NOT_INITIALISED;
    return cvi_cuda( cudaMemsetAsync( dev, byte, bytes, cvi_stream( stream ) ), "cudaMemsetAsync" );
}

int cv_upload( void *dev, const void *host, size_t bytes, cv_stream stream ) {
    if( !cvi_device() ) return CV_ERR_NOT_INITIALISED;
    return cvi_cuda( cudaMemcpyAsync( dev, host, bytes, cudaMemcpyHostToDevice, cvi_stream( stream ) ), "cudaMemcpyAsync(H2D)" );
}

int cv_download( void *host, const void *dev, size_t bytes, cv_stream stream ) {
    if( !cvi_device() ) return CV_ERR_NOT_INITIALISED;
    return cvi_cuda( cudaMemcpyAsync( host, dev, bytes, cudaMemcpyDeviceToHost, cvi_stream( stream ) ), "cudaMemcpyAsync(D2H)" );
}

int cv_copy_d2d( void *dst, const void *src, size_t bytes, cv_stream stream ) {
    if( !cvi_device() ) return CV_ERR_NOT_INITIALISED;
    return cvi_cuda( cudaMemcpyAsync( dst, src, bytes, cudaMemcpyDeviceToDevice, cvi_stream( stream ) ), "cudaMemcpyAsync(D2D)" );
}

int cv_flush_l2( cv_stream stream ) {
    CvDevice *dev = cvi_device();
    if( !dev ) return CV_ERR_NOT_INITIALISED;

    if( !dev->flush_buf ) {
        /* 2x L2 (B200: 2 x 126 MB) so that every line of the previous working set is displaced */
        dev->flush_bytes = dev->l2_bytes ? dev->l2_bytes * 2 : ((size_t) 256 << 20);
        int rc = cvi_cuda( cudaMalloc( &dev->flush_buf, dev->flush_bytes ), "cudaMalloc(flush buffer)" );
        if( rc ) { dev->flush_buf = NULL; return rc; }
        cudaMemset( dev->flush_buf, 0, dev->flush_bytes );
    }
    return cvi_launch_flush( dev->flush_buf, dev->flush_bytes, cvi_stream( stream ) );
}

/* --------------------------------------------------------------- frames */

static size_t window_pixels( const box2i *w ) {
    v2i size;
    box2i_get_size( w, &size );
    return (size_t) size.x * (size_t) size.y;
}

static int frame_alloc( void **data, const box2i *full_window, size_t pixel_bytes, int where ) {
    if( !data || !full_window ) return cvi_fail( CV_ERR_ARGUMENT, "cv_frame_alloc: NULL argument" );
    return cv_malloc( data, window_pixels( full_window ) * pixel_bytes, where );
}

int cv_frame_alloc_f16( rgba_frame_f16 *frame, const box2i *full_window, int where ) {
    if( !frame ) return cvi_fail( CV_ERR_ARGUMENT, "cv_frame_alloc_f16: NULL frame" );
    void *data = NULL;
    int rc = frame_alloc( &data, full_window, sizeof(rgba_f16), where );
    frame->data = (rgba_f16 *) data;
    if( full_window ) frame->full_window = *full_window;
    box2i_set_empty( &frame->current_window );
    return rc;
}

int cv_frame_alloc_f32( rgba_frame_f32 *frame, const box2i *full_window, int where ) {
    if( !frame ) return cvi_fail( CV_ERR_ARGUMENT, "cv_frame_alloc_f32: NULL frame" );
    void *data = NULL;
    int rc = frame_alloc( &data, full_window, sizeof(rgba_f32), where );
    frame->data = (rgba_f32 *) data;
    if( full_window ) frame->full_window = *full_window;
    box2i_set_empty( &frame->current_window );
    return rc;
}

int cv_frame_free_f16( rgba_frame_f16 *frame, int where ) {
    if( !frame ) return CV_OK;
    int rc = cv_free( frame->data, where );
    frame->data = NULL;
    return rc;
}

int cv_frame_free_f32( rgba_frame_f32 *frame, int where ) {
    if( !frame ) return CV_OK;
    int rc = cv_free( frame->data, where );
    frame->data = NULL;
    return rc;
}

static int same_window( const box2i *a, const box2i *b ) {
    return a->min.x == b->min.x && a->min.y == b->min.y && a->max.x == b->max.x && a->max.y == b->max.y;
}

int cv_frame_upload_f16( rgba_frame_f16 *dev, const rgba_frame_f16 *host, cv_stream stream ) {
    if( !dev || !host || !same_window( &dev->full_window, &host->full_window ) )
        return cvi_fail( CV_ERR_ARGUMENT, "cv_frame_upload_f16: frames must share full_window" );
    dev->current_window = host->current_window;
    return cv_upload( dev->data, host->data, window_pixels( &host->full_window ) * sizeof(rgba_f16), stream );
}

int cv_frame_download_f16( rgba_frame_f16 *host, const rgba_frame_f16 *dev, cv_stream stream ) {
    if( !dev || !host || !same_window( &dev->full_window, &host->full_window ) )
        return cvi_fail( CV_ERR_ARGUMENT, "cv_frame_download_f16: frames must share full_window" );
    host->current_window = dev->current_window;
    return cv_download( host->data, dev->data, window_pixels( &host->full_window ) * sizeof(rgba_f16), stream );
}

int cv_frame_upload_f32( rgba_frame_f32 *dev, const rgba_frame_f32 *host, cv_stream stream ) {
    if( !dev || !host || !same_window( &dev->full_window, &host->full_window ) )
        return cvi_fail( CV_ERR_ARGUMENT, "cv_frame_upload_f32: frames must share full_window" );
    dev->current_window = host->current_window;
    return cv_upload( dev->data, host->data, window_pixels( &host->full_window ) * sizeof(rgba_f32), stream );
}

int cv_frame_download_f32( rgba_frame_f32 *host, const rgba_frame_f32 *dev, cv_stream stream ) {
    if( !dev || !host || !same_window( &dev->full_window, &host->full_window ) )
        return cvi_fail( CV_ERR_ARGUMENT, "cv_frame_download_f32: frames must share full_window" );
    host->current_window = dev->current_window;
    return cv_download( host->data, dev->data, window_pixels( &host->full_window ) * sizeof(rgba_f32), stream );
}
