/*
 * canvas_b200/csrc/cv_host_workspace.c -- the workspace (timeline) API of include/framework.h.
 *
 * Replaces /root/reference/src/cprocess/workspace.c: item bookkeeping (:309-492) and the video
 * compositor workspace_get_frame_f32 (:494-550).  Differences in mechanism, not in results:
 *
 *   - the reference keeps two sorted GSequences plus a moving "composite list" that is updated
 *     under a mutex as the playhead moves (:243-307), a sequential-access optimisation.  Here the
 *     set of items showing at a frame is a STATELESS query (x <= frame < x + length, ordered by z),
 *     so any number of threads / GPUs can render different frames of one workspace at once;
 *   - layers are accumulated in HBM: each source is pulled with the reference's own vtable
 *     (video_get_frame_f16/_f32), uploaded once, and the whole stack is composited by ONE fused
 *     launch when every source is f16-native, or by the f32 "over" kernel otherwise; one download.
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

static void insert_sorted( workspace_t *ws, workspace_item_t *item ) {
    if( ws->count == ws->capacity ) {
        ws->capacity = ws->capacity ? ws->capacity * 2 : 16;
        ws->items = (workspace_item_t **) realloc( ws->items, sizeof(*ws->items) * (size_t) ws->capacity );
    }
    int at = ws->count;
    while( at > 0 && by_position( item, ws->items[at - 1] ) < 0 ) {
        ws->items[at] = ws->items[at - 1];
        at--;
    }
    ws->items[at] = item;
    ws->count++;
}

static void detach( workspace_t *ws, workspace_item_t *item ) {
    for( int i = 0; i < ws->count; i++ ) {
        if( ws->items[i] == item ) {
            memmove( &ws->items[i], &ws->items[i + 1], sizeof(*ws->items) * (size_t)(ws->count - i - 1) );
            ws->count--;
            return;
        }
    }
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
    pthread_mutex_unlock( &ws->lock );
}

/* ------------------------------------------------------------------ rendering */

typedef struct Showing { video_source *source; int frame; } Showing;

/* The items showing at `frame`, bottom first (workspace.c:510-520, 530, 542). */
static int showing_at( workspace_t *ws, int frame, Showing **out ) {
    pthread_mutex_lock( &ws->lock );
    workspace_item_t **hit = (workspace_item_t **) malloc( sizeof(*hit) * (size_t)(ws->count + 1) );
    int n = 0;
    for( int i = 0; i < ws->count; i++ ) {
        workspace_item_t *it = ws->items[i];
        if( it->x <= (int64_t) frame && (int64_t) frame < it->x + it->length )
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

/* Every showing source is f16-native: pull, upload the valid rows, ONE fused composite launch. */
static int render_fused_f16( const Showing *list, int n, rgba_frame_f16 *frame ) {
    cv_layer *layers = (cv_layer *) calloc( (size_t) n, sizeof(cv_layer) );
    rgba_frame_f16 host;
    host.data = (rgba_f16 *) malloc( sizeof(rgba_f16) * (window_pixels( &frame->full_window ) + 1) );
    host.full_window = frame->full_window;
    int rc = (layers && host.data) ? CV_OK : CV_ERR_MEMORY;

    for( int i = 0; i < n && !rc; i++ ) {
        host.current_window = frame->full_window;
        video_get_frame_f16( list[i].source, list[i].frame, &host );
        rc = cv_frame_alloc_f16( &layers[i].frame, &frame->full_window, CV_MEM_DEVICE );
        if( !rc ) rc = cv_frame_upload_f16( &layers[i].frame, &host, NULL );
        if( !rc ) rc = cv_stream_sync( NULL );          /* `host` is reused for the next layer */
        layers[i].mix = 1.0f;
    }

    rgba_frame_f16 out = { 0 };
    if( !rc ) rc = cv_frame_alloc_f16( &out, &frame->full_window, CV_MEM_DEVICE );
    if( !rc ) rc = cv_composite_over_f16( &out, layers, n, NULL );
    if( !rc ) rc = cv_frame_download_f16( &host, &out, NULL );
    if( !rc ) rc = cv_stream_sync( NULL );

    if( !rc ) {
        /* only the result window is written into the caller's frame */
        frame->current_window = out.current_window;
        if( !box2i_is_empty( &out.current_window ) ) {
            v2i size;
            box2i_get_size( &out.current_window, &size );
            for( int y = out.current_window.min.y; y <= out.current_window.max.y; y++ )
                memcpy( video_get_pixel_f16( frame, out.current_window.min.x, y ), video_get_pixel_f16( &host, out.current_window.min.x, y ),
                    sizeof(rgba_f16) * (size_t) size.x );
        }
    }

    cv_stream_sync( NULL );
    for( int i = 0; i < n; i++ )
        if( layers && layers[i].frame.data ) cv_frame_free_f16( &layers[i].frame, CV_MEM_DEVICE );
    if( out.data ) cv_frame_free_f16( &out, CV_MEM_DEVICE );
    free( layers );
    free( host.data );
    return rc;
}

/* The general case (some source is f32-only): accumulate in an f32 frame in HBM. */
static int render_accumulate_f32( const Showing *list, int n, rgba_frame_f32 *frame ) {
    rgba_frame_f32 host, acc = { 0 }, layer = { 0 };
    host.data = (rgba_f32 *) malloc( sizeof(rgba_f32) * (window_pixels( &frame->full_window ) + 1) );
    host.full_window = frame->full_window;
    int rc = host.data ? CV_OK : CV_ERR_MEMORY;

    if( !rc ) rc = cv_frame_alloc_f32( &acc, &frame->full_window, CV_MEM_DEVICE );
    if( !rc ) rc = cv_frame_alloc_f32( &layer, &frame->full_window, CV_MEM_DEVICE );

    for( int i = 0; i < n && !rc; i++ ) {
        host.current_window = frame->full_window;
        video_get_frame_f32( list[i].source, list[i].frame, &host );
        rgba_frame_f32 *target = (i == 0) ? &acc : &layer;
        rc = cv_frame_upload_f32( target, &host, NULL );
        if( !rc ) box2i_intersect( &target->current_window, &host.current_window, &frame->full_window );
        if( !rc && i > 0 ) rc = cv_mix_over_f32( &acc, &layer, 1.0f, NULL );
        if( !rc ) rc = cv_stream_sync( NULL );
    }

    if( !rc ) rc = cv_frame_download_f32( &host, &acc, NULL );
    if( !rc ) rc = cv_stream_sync( NULL );

    if( !rc ) {
        frame->current_window = acc.current_window;
        if( !box2i_is_empty( &acc.current_window ) ) {
            v2i size;
            box2i_get_size( &acc.current_window, &size );
            for( int y = acc.current_window.min.y; y <= acc.current_window.max.y; y++ )
                memcpy( video_get_pixel_f32( frame, acc.current_window.min.x, y ), video_get_pixel_f32( &host, acc.current_window.min.x, y ),
                    sizeof(rgba_f32) * (size_t) size.x );
        }
    }

    if( acc.data ) cv_frame_free_f32( &acc, CV_MEM_DEVICE );
    if( layer.data ) cv_frame_free_f32( &layer, CV_MEM_DEVICE );
    free( host.data );
    return rc;
}

static int device_ready( void ) {
    if( cv_get_device() >= 0 ) return 1;
    const char *env = getenv( "CANVAS_B200_DEVICE" );
    return cv_init( env ? atoi( env ) : 0 ) == CV_OK;
}

static void workspace_get_frame_f32( void *self, int frame_index, rgba_frame_f32 *frame ) {
    Showing *list = NULL;
    const int n = showing_at( (workspace_t *) self, frame_index, &list );

    if( n == 0 || !device_ready() ) {               /* workspace.c:504-508 */
        box2i_set_empty( &frame->current_window );
        free( list );
        return;
    }

    if( render_accumulate_f32( list, n, frame ) )
        box2i_set_empty( &frame->current_window );
    free( list );
}

static void workspace_get_frame_f16( void *self, int frame_index, rgba_frame_f16 *frame ) {
    Showing *list = NULL;
    const int n = showing_at( (workspace_t *) self, frame_index, &list );

    if( n == 0 || !device_ready() ) {
        box2i_set_empty( &frame->current_window );
        free( list );
        return;
    }

    int rc;
    if( all_f16_native( list, n ) ) {
        rc = render_fused_f16( list, n, frame );
    }
    else {
        /* as the reference's dispatcher would: composite in f32, pack once at the end (main.c:44-72) */
        rgba_frame_f32 wide, dev32 = { 0 };
        rgba_frame_f16 dev16 = { 0 };
        wide.data = (rgba_f32 *) malloc( sizeof(rgba_f32) * (window_pixels( &frame->full_window ) + 1) );
        wide.full_window = frame->full_window;
        rc = wide.data ? render_accumulate_f32( list, n, &wide ) : CV_ERR_MEMORY;

        if( !rc && !box2i_is_empty( &wide.current_window ) ) {
            rgba_frame_f16 host;
            host.data = (rgba_f16 *) malloc( sizeof(rgba_f16) * (window_pixels( &frame->full_window ) + 1) );
            host.full_window = frame->full_window;
            rc = cv_frame_alloc_f32( &dev32, &frame->full_window, CV_MEM_DEVICE );
            if( !rc ) rc = cv_frame_upload_f32( &dev32, &wide, NULL );
            if( !rc ) rc = cv_frame_alloc_f16( &dev16, &frame->full_window, CV_MEM_DEVICE );
            if( !rc ) rc = cv_frame_f32_to_f16( &dev16, &dev32, NULL );
            if( !rc && host.data ) rc = cv_frame_download_f16( &host, &dev16, NULL );
            if( !rc ) rc = cv_stream_sync( NULL );
            if( !rc && host.data ) {
                v2i size;
                box2i_get_size( &dev16.current_window, &size );
                for( int y = dev16.current_window.min.y; y <= dev16.current_window.max.y; y++ )
                    memcpy( video_get_pixel_f16( frame, dev16.current_window.min.x, y ), video_get_pixel_f16( &host, dev16.current_window.min.x, y ),
                        sizeof(rgba_f16) * (size_t) size.x );
                frame->current_window = dev16.current_window;
            }
            free( host.data );
        }
        else if( !rc ) {
            frame->current_window = wide.current_window;
        }

        cv_stream_sync( NULL );
        if( dev32.data ) cv_frame_free_f32( &dev32, CV_MEM_DEVICE );
        if( dev16.data ) cv_frame_free_f16( &dev16, CV_MEM_DEVICE );
        free( wide.data );
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
