/*
 * canvas_b200/csrc/cv_host_workspace.c -- the workspace (timeline) API of include/framework.h.
 *
 * Replaces /root/reference/src/cprocess/workspace.c: item bookkeeping (:309-492) and the video
 * compositor workspace_get_frame_f32 (:494-550).  Differences in mechanism, not in results:
 *
 *   - the reference keeps two sorted GSequences plus a moving "composite list" that is updated
 *     under a mutex as the playhead moves (:243-307), a sequential-access optimisation.  Here the
 *     set of items showing at a frame is a STATELESS query (x <= frame < x + length, ordered by z)
 *     over an array sorted by x with a running maximum of the items' ends: a binary search finds the
 *     last item that starts at or before the frame, and the walk back from it stops as soon as no
 *     earlier item can still be showing -- O(log n + items near the frame), so any number of threads /
 *     GPUs can render different frames of one 10 000-item workspace at once;
 *   - layers are accumulated in HBM on the CALLING THREAD's own stream (cv_host_ctx.c): each source is
 *     pulled with the reference's own vtable (video_get_frame_f16/_f32) straight into a pinned staging
 *     slab, its valid rows go up as an asynchronous copy while the next source is being pulled, and the
 *     whole stack is composited by ONE fused launch when every source is f16-native (the f32 "over"
 *     kernel otherwise); one download.  Nothing here waits for another thread's work.
 *
 * Order of equal-z items: the later-added item is the lower one (what the reference build used by
 * the oracle does; the reference leaves it to GSequence's insertion order, so it is unpinned).
 */
#define _GNU_SOURCE
#include "cv_internal.h"

#include <stdlib.h>
#include <string.h>

struct workspace_item_t_tag {
    workspace_t *workspace;
    int64_t x, length, z, offset;
    void *source, *tag;
    uint64_t serial;            /* insertion order */
};

struct workspace_t_tag {
    pthread_mutex_t lock;
    workspace_item_t **items;   /* kept sorted by (x, z) like the reference's leftsort (workspace_get_item order) */
    int64_t *max_end;           /* max_end[i] = the largest x + length among items[0..i] */
    int count, capacity;
    uint64_t next_serial;
};

static int by_position( const workspace_item_t *a, const workspace_item_t *b ) {
    if( a->x != b->x ) return a->x < b->x ? -1 : 1;
    if( a->z != b->z ) return a->z < b->z ? -1 : 1;
    return a->serial < b->serial ? -1 : (a->serial > b->serial);
}

/* bottom first: ascending z; among equals the later-added item is lower */
static int by_depth( const void *pa, const void *pb ) {
    const workspace_item_t *a = *(workspace_item_t *const *) pa, *b = *(workspace_item_t *const *) pb;
    if( a->z != b->z ) return a->z < b->z ? -1 : 1;
    return a->serial > b->serial ? -1 : (a->serial < b->serial);
}

static void refresh_ends( workspace_t *ws, int from ) {
    for( int i = from < 0 ? 0 : from; i < ws->count; i++ ) {
        const int64_t end = ws->items[i]->x + ws->items[i]->length;
        ws->max_end[i] = (i > 0 && ws->max_end[i - 1] > end) ? ws->max_end[i - 1] : end;
    }
}

/* index of the first item that sorts after `item` (binary search) */
static int upper_bound( const workspace_t *ws, const workspace_item_t *item ) {
    int lo = 0, hi = ws->count;
    while( lo < hi ) {
        const int mid = lo + (hi - lo) / 2;
        if( by_position( ws->items[mid], item ) <= 0 ) lo = mid + 1; else hi = mid;
    }
    return lo;
}

static void insert_sorted( workspace_t *ws, workspace_item_t *item ) {
    if( ws->count == ws->capacity ) {
        ws->capacity = ws->capacity ? ws->capacity * 2 : 16;
        ws->items = (workspace_item_t **) realloc( ws->items, sizeof(*ws->items) * (size_t) ws->capacity );
        ws->max_end = (int64_t *) realloc( ws->max_end, sizeof(*ws->max_end) * (size_t) ws->capacity );
    }
    const int at = upper_bound( ws, item );
    memmove( &ws->items[at + 1], &ws->items[at], sizeof(*ws->items) * (size_t)(ws->count - at) );
    ws->items[at] = item;
    ws->count++;
    refresh_ends( ws, at );
}

static void detach( workspace_t *ws, workspace_item_t *item ) {
    /* the item sits just before its own upper bound */
    int at = upper_bound( ws, item ) - 1;
    if( at < 0 || ws->items[at] != item ) {
        for( at = 0; at < ws->count && ws->items[at] != item; at++ ) ;
        if( at == ws->count ) return;
    }
    memmove( &ws->items[at], &ws->items[at + 1], sizeof(*ws->items) * (size_t)(ws->count - at - 1) );
    ws->count--;
    refresh_ends( ws, at );
}

workspace_t *workspace_create( void ) {
    workspace_t *ws = (workspace_t *) calloc( 1, sizeof(*ws) );
    if( ws ) pthread_mutex_init( &ws->lock, NULL );
    return ws;
}

int workspace_get_length( workspace_t *ws ) {
    pthread_mutex_lock( &ws->lock );
    const int n = ws->count;
    pthread_mutex_unlock( &ws->lock );
    return n;
}

workspace_item_t *workspace_add_item( workspace_t *ws, void *source, int64_t x, int64_t width, int64_t offset, int64_t z, void *tag ) {
    workspace_item_t *item = (workspace_item_t *) calloc( 1, sizeof(*item) );
    if( !item ) return NULL;
    item->workspace = ws;
    item->x = x; item->length = width; item->offset = offset; item->z = z;
    item->source = source; item->tag = tag;

    pthread_mutex_lock( &ws->lock );
    item->serial = ws->next_serial++;
    insert_sorted( ws, item );
    pthread_mutex_unlock( &ws->lock );
    return item;
}

workspace_item_t *workspace_get_item( workspace_t *ws, int index ) {
    pthread_mutex_lock( &ws->lock );
    workspace_item_t *item = (index >= 0 && index < ws->count) ? ws->items[index] : NULL;
    pthread_mutex_unlock( &ws->lock );
    return item;
}

void workspace_remove_item( workspace_item_t *item ) {
    workspace_t *ws = item->workspace;
    pthread_mutex_lock( &ws->lock );
    detach( ws, item );
    pthread_mutex_unlock( &ws->lock );
    free( item );
}

void workspace_free( workspace_t *ws ) {
    for( int i = 0; i < ws->count; i++ )
        free( ws->items[i] );
    free( ws->items );
    free( ws->max_end );
    pthread_mutex_destroy( &ws->lock );
    free( ws );
}

void workspace_get_item_pos( workspace_item_t *item, int64_t *x, int64_t *width, int64_t *z ) {
    if( x ) *x = item->x;
    if( width ) *width = item->length;
    if( z ) *z = item->z;
}

int64_t workspace_get_item_offset( workspace_item_t *item ) { return item->offset; }
void workspace_set_item_offset( workspace_item_t *item, int64_t offset ) { item->offset = offset; }
void *workspace_get_item_source( workspace_item_t *item ) { return item->source; }
void workspace_set_item_source( workspace_item_t *item, void *source ) { item->source = source; }
void *workspace_get_item_tag( workspace_item_t *item ) { return item->tag; }
void workspace_set_item_tag( workspace_item_t *item, void *tag ) { item->tag = tag; }

void workspace_update_item( workspace_item_t *item, int64_t *x, int64_t *width, int64_t *z, int64_t *offset, void **source, void **tag ) {
    workspace_t *ws = item->workspace;
    pthread_mutex_lock( &ws->lock );
    const int moves = (x && *x != item->x) || (z && *z != item->z);
    if( moves ) detach( ws, item );
    if( x ) item->x = *x;
    if( width ) item->length = *width;
    if( z ) item->z = *z;
    if( offset ) item->offset = *offset;
    if( source ) item->source = *source;
    if( tag ) item->tag = *tag;
    if( moves ) insert_sorted( ws, item );
    else if( width ) refresh_ends( ws, 0 );
    pthread_mutex_unlock( &ws->lock );
}

/* ------------------------------------------------------------------ rendering */

typedef struct Showing { video_source *source; int frame; } Showing;

/* The items showing at `frame`, bottom first (workspace.c:510-520, 530, 542). */
static int showing_at( workspace_t *ws, int frame, Showing **out ) {
    pthread_mutex_lock( &ws->lock );

    /* the last item with x <= frame, then back while an earlier item may still reach the frame */
    int lo = 0, hi = ws->count;
    while( lo < hi ) {
        const int mid = lo + (hi - lo) / 2;
        if( ws->items[mid]->x <= (int64_t) frame ) lo = mid + 1; else hi = mid;
    }

    workspace_item_t **hit = (workspace_item_t **) malloc( sizeof(*hit) * (size_t)(lo + 1) );
    int n = 0;
    for( int i = lo - 1; i >= 0 && ws->max_end[i] > (int64_t) frame; i-- ) {
        workspace_item_t *it = ws->items[i];
        if( (int64_t) frame < it->x + it->length )
            hit[n++] = it;
    }
    qsort( hit, (size_t) n, sizeof(*hit), by_depth );

    Showing *list = (Showing *) malloc( sizeof(Showing) * (size_t)(n + 1) );
    for( int i = 0; i < n; i++ ) {
        list[i].source = (video_source *) hit[i]->source;
        list[i].frame = (int)( (int64_t) frame - hit[i]->x + hit[i]->offset );
    }
    pthread_mutex_unlock( &ws->lock );
    free( hit );
    *out = list;
    return n;
}

static size_t window_pixels( const box2i *w ) {
    v2i size;
    box2i_get_size( w, &size );
    return (size_t) size.x * (size_t) size.y;
}

static int all_f16_native( const Showing *list, int n ) {
    for( int i = 0; i < n; i++ )
        if( !list[i].source || !list[i].source->funcs || !list[i].source->funcs->get_frame )
            return 0;
    return 1;
}

/* rows of `win` between a host buffer and a device buffer that both hold full_window, on the thread's stream */
static int copy_rows( void *dst, const void *src, const box2i *full, const box2i *win, size_t px, enum cudaMemcpyKind kind, cudaStream_t stream ) {
    if( box2i_is_empty( win ) ) return CV_OK;
    v2i size, wsize;
    box2i_get_size( full, &size );
    box2i_get_size( win, &wsize );
    const size_t pitch = (size_t) size.x * px;
    const size_t offset = ((size_t)(win->min.y - full->min.y) * (size_t) size.x + (size_t)(win->min.x - full->min.x)) * px;
    if( wsize.x == size.x )         /* whole rows: one contiguous copy */
        return cvi_cuda( cudaMemcpyAsync( (char *) dst + offset, (const char *) src + offset, pitch * (size_t) wsize.y, kind, stream ), "cudaMemcpyAsync(rows)" );
    return cvi_cuda( cudaMemcpy2DAsync( (char *) dst + offset, pitch, (const char *) src + offset, pitch,
        (size_t) wsize.x * px, (size_t) wsize.y, kind, stream ), "cudaMemcpy2DAsync(window)" );
}

static void copy_window_to_caller( void *dst, const void *src, const box2i *full, const box2i *win, size_t px ) {
    if( box2i_is_empty( win ) ) return;
    v2i size, wsize;
    box2i_get_size( full, &size );
    box2i_get_size( win, &wsize );
    for( int y = win->min.y; y <= win->max.y; y++ ) {
        const size_t offset = ((size_t)(y - full->min.y) * (size_t) size.x + (size_t)(win->min.x - full->min.x)) * px;
        memcpy( (char *) dst + offset, (const char *) src + offset, (size_t) wsize.x * px );
    }
}

/* Every showing source is f16-native: pull each into a pinned slab, send its valid rows up asynchronously while the
 * next one is pulled, ONE fused composite launch, one download. */
static int render_fused_f16( CvHostCtx *ctx, const Showing *list, int n, rgba_frame_f16 *frame ) {
    const size_t bytes = sizeof(rgba_f16) * window_pixels( &frame->full_window );
    cv_layer *layers = (cv_layer *) calloc( (size_t) n, sizeof(cv_layer) );
    int rc = layers ? CV_OK : cvi_fail( CV_ERR_MEMORY, "out of host memory" );

    for( int i = 0; i < n && !rc; i++ ) {
        int slab = -1;
        rgba_frame_f16 host;
        host.data = (rgba_f16 *) cvi_host_slab( ctx, bytes, &slab );
        host.full_window = frame->full_window;
        host.current_window = frame->full_window;
        if( !host.data ) { rc = CV_ERR_MEMORY; break; }

        video_get_frame_f16( list[i].source, list[i].frame, &host );

        rc = cv_frame_alloc_f16( &layers[i].frame, &frame->full_window, CV_MEM_DEVICE );
        box2i_intersect( &layers[i].frame.current_window, &host.current_window, &frame->full_window );
        if( !rc ) rc = copy_rows( layers[i].frame.data, host.data, &frame->full_window, &layers[i].frame.current_window,
                                  sizeof(rgba_f16), cudaMemcpyHostToDevice, ctx->stream );
        cvi_host_slab_done( ctx, slab );
        layers[i].mix = 1.0f;
    }

    rgba_frame_f16 out = { 0 };
    if( !rc ) rc = cv_frame_alloc_f16( &out, &frame->full_window, CV_MEM_DEVICE );
    if( !rc ) rc = cv_composite_over_f16( &out, layers, n, (cv_stream) ctx->stream );

    if( !rc ) {
        int slab = -1;
        void *staging = cvi_host_slab( ctx, bytes, &slab );
        if( !staging ) rc = CV_ERR_MEMORY;
        if( !rc ) rc = copy_rows( staging, out.data, &frame->full_window, &out.current_window, sizeof(rgba_f16), cudaMemcpyDeviceToHost, ctx->stream );
        if( !rc ) rc = cvi_cuda( cudaStreamSynchronize( ctx->stream ), "cudaStreamSynchronize" );
        if( !rc ) {
            /* only the result window is written into the caller's frame */
            frame->current_window = out.current_window;
            copy_window_to_caller( frame->data, staging, &frame->full_window, &out.current_window, sizeof(rgba_f16) );
        }
    }

    for( int i = 0; i < n; i++ )
        if( layers && layers[i].frame.data ) { cvi_pool_free_async( layers[i].frame.data, ctx->stream ); layers[i].frame.data = NULL; }
    if( out.data ) cvi_pool_free_async( out.data, ctx->stream );
    free( layers );
    return rc;
}

/* The general case (some source is f32-only): accumulate in an f32 frame in HBM, two layer buffers in turn so that the
 * upload of one layer overlaps the "over" of the previous one. */
static int render_accumulate_f32( CvHostCtx *ctx, const Showing *list, int n, rgba_frame_f32 *frame, int into_caller, rgba_frame_f32 *keep_on_device ) {
    const size_t bytes = sizeof(rgba_f32) * window_pixels( &frame->full_window );
    rgba_frame_f32 acc = { 0 }, layer[2] = { { 0 }, { 0 } };
    int rc = cv_frame_alloc_f32( &acc, &frame->full_window, CV_MEM_DEVICE );
    if( !rc && n > 1 ) rc = cv_frame_alloc_f32( &layer[0], &frame->full_window, CV_MEM_DEVICE );
    if( !rc && n > 2 ) rc = cv_frame_alloc_f32( &layer[1], &frame->full_window, CV_MEM_DEVICE );

    for( int i = 0; i < n && !rc; i++ ) {
        int slab = -1;
        rgba_frame_f32 host;
        host.data = (rgba_f32 *) cvi_host_slab( ctx, bytes, &slab );
        host.full_window = frame->full_window;
        host.current_window = frame->full_window;
        if( !host.data ) { rc = CV_ERR_MEMORY; break; }

        video_get_frame_f32( list[i].source, list[i].frame, &host );

        rgba_frame_f32 *target = (i == 0) ? &acc : &layer[(i - 1) & 1];
        box2i_intersect( &target->current_window, &host.current_window, &frame->full_window );
        rc = copy_rows( target->data, host.data, &frame->full_window, &target->current_window, sizeof(rgba_f32), cudaMemcpyHostToDevice, ctx->stream );
        cvi_host_slab_done( ctx, slab );
        if( !rc && i > 0 ) rc = cv_mix_over_f32( &acc, target, 1.0f, (cv_stream) ctx->stream );
    }

    if( !rc && into_caller ) {
        int slab = -1;
        void *staging = cvi_host_slab( ctx, bytes, &slab );
        if( !staging ) rc = CV_ERR_MEMORY;
        if( !rc ) rc = copy_rows( staging, acc.data, &frame->full_window, &acc.current_window, sizeof(rgba_f32), cudaMemcpyDeviceToHost, ctx->stream );
        if( !rc ) rc = cvi_cuda( cudaStreamSynchronize( ctx->stream ), "cudaStreamSynchronize" );
        if( !rc ) copy_window_to_caller( frame->data, staging, &frame->full_window, &acc.current_window, sizeof(rgba_f32) );
    }
    if( !rc ) frame->current_window = acc.current_window;

    if( layer[0].data ) cvi_pool_free_async( layer[0].data, ctx->stream );
    if( layer[1].data ) cvi_pool_free_async( layer[1].data, ctx->stream );
    if( !rc && keep_on_device ) *keep_on_device = acc;
    else if( acc.data ) cvi_pool_free_async( acc.data, ctx->stream );
    return rc;
}

static void workspace_get_frame_f32( void *self, int frame_index, rgba_frame_f32 *frame ) {
    Showing *list = NULL;
    const int n = showing_at( (workspace_t *) self, frame_index, &list );
    CvHostCtx *ctx = n ? cvi_host_ctx() : NULL;

    if( n == 0 || !ctx ) {                          /* workspace.c:504-508 */
        box2i_set_empty( &frame->current_window );
        free( list );
        return;
    }

    if( render_accumulate_f32( ctx, list, n, frame, 1, NULL ) )
        box2i_set_empty( &frame->current_window );
    free( list );
}

static void workspace_get_frame_f16( void *self, int frame_index, rgba_frame_f16 *frame ) {
    Showing *list = NULL;
    const int n = showing_at( (workspace_t *) self, frame_index, &list );
    CvHostCtx *ctx = n ? cvi_host_ctx() : NULL;

    if( n == 0 || !ctx ) {
        box2i_set_empty( &frame->current_window );
        free( list );
        return;
    }

    int rc;
    if( all_f16_native( list, n ) ) {
        rc = render_fused_f16( ctx, list, n, frame );
    }
    else {
        /* as the reference's dispatcher would: composite in f32, pack once at the end (main.c:44-72) -- the f32 result
         * stays in HBM and only the packed halves travel back */
        rgba_frame_f32 wide = { 0 }, acc = { 0 };
        rgba_frame_f16 dev16 = { 0 };
        wide.full_window = frame->full_window;
        rc = render_accumulate_f32( ctx, list, n, &wide, 0, &acc );

        if( !rc && !box2i_is_empty( &acc.current_window ) ) {
            const size_t bytes = sizeof(rgba_f16) * window_pixels( &frame->full_window );
            int slab = -1;
            void *staging = cvi_host_slab( ctx, bytes, &slab );
            if( !staging ) rc = CV_ERR_MEMORY;
            if( !rc ) rc = cv_frame_alloc_f16( &dev16, &frame->full_window, CV_MEM_DEVICE );
            if( !rc ) rc = cv_frame_f32_to_f16( &dev16, &acc, (cv_stream) ctx->stream );
            if( !rc ) rc = copy_rows( staging, dev16.data, &frame->full_window, &dev16.current_window, sizeof(rgba_f16), cudaMemcpyDeviceToHost, ctx->stream );
            if( !rc ) rc = cvi_cuda( cudaStreamSynchronize( ctx->stream ), "cudaStreamSynchronize" );
            if( !rc ) {
                copy_window_to_caller( frame->data, staging, &frame->full_window, &dev16.current_window, sizeof(rgba_f16) );
                frame->current_window = dev16.current_window;
            }
        }
        else if( !rc ) {
            frame->current_window = acc.current_window;
        }

        if( acc.data ) cvi_pool_free_async( acc.data, ctx->stream );
        if( dev16.data ) cvi_pool_free_async( dev16.data, ctx->stream );
    }

    if( rc )
        box2i_set_empty( &frame->current_window );
    free( list );
}

static video_frame_source_funcs workspace_video_funcs = {
    .flags = 0,
    .get_frame = workspace_get_frame_f16,
    .get_frame_32 = workspace_get_frame_f32,
    .get_frame_gl = NULL,
};

void workspace_as_video_source( workspace_t *workspace, video_source *source ) {
    source->obj = workspace;
    source->funcs = &workspace_video_funcs;
}
