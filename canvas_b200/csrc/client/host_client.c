/*
 * canvas_b200/csrc/client/host_client.c -- a C CLIENT of the drop-in boundary (builds canvas_b200/libcanvas_client.so,
 * linked against libcanvas_b200.so, include/framework.h only).
 *
 * It does what a host application of the reference does with src/cprocess: keeps f16 frames in HOST memory, wraps them
 * as video sources behind the video_frame_source_funcs vtable (the way src/process/RgbaFrameF16.c:27-38 does: get_frame =
 * video_copy_frame_f16 out of the stored frame), places them on a workspace with workspace_add_item
 * (framework.h:573) and pulls composited frames with video_get_frame_f16( &workspace_source, frame, &host_frame )
 * (src/cprocess/main.c:34) from several threads at once (the pull queue's contract, src/process/VideoPullQueue.c:99-110).
 *
 * bench.py's `e2e_plugin` times exactly this against the library, and the reference arm does the same calls against
 * the reference's own C (oracle/ref_harness.c) -- same call on both arms, host buffers in, host buffer out.
 * Nothing here knows about CUDA.
 */
#define _GNU_SOURCE
#include "framework.h"

#include <pthread.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

/* ---- host frames as sources ---------------------------------------------------------------- */

typedef struct {
    rgba_frame_f16 *frames;     /* source frame i is frames[i mod count] */
    int count;
} clip_source;

static void clip_get_frame( void *self, int frame_index, rgba_frame_f16 *frame ) {
    const clip_source *clip = (const clip_source *) self;
    int i = frame_index % clip->count;
    if( i < 0 ) i += clip->count;
    video_copy_frame_f16( frame, &clip->frames[i] );
}

static video_frame_source_funcs clip_funcs = { .flags = 0, .get_frame = clip_get_frame, .get_frame_32 = NULL, .get_frame_gl = NULL };

typedef struct {
    workspace_t *workspace;
    video_source as_source;
    video_source *sources;
    clip_source *clips;
    int count;
} client_timeline;

/* item i shows clip i = frames[first[i] .. first[i] + clip_count[i]) at timeline frames [x, x + length), stacked by z */
EXPORT client_timeline *client_timeline_create( rgba_frame_f16 *frames, const int *first, const int *clip_count,
        const int64_t *x, const int64_t *length, const int64_t *offset, const int64_t *z, int count ) {
    client_timeline *tl = (client_timeline *) calloc( 1, sizeof(*tl) );
    if( !tl ) return NULL;
    tl->workspace = workspace_create();
    tl->sources = (video_source *) calloc( (size_t)( count > 0 ? count : 1 ), sizeof(video_source) );
    tl->clips = (clip_source *) calloc( (size_t)( count > 0 ? count : 1 ), sizeof(clip_source) );
    tl->count = count;

    for( int i = 0; i < count; i++ ) {
        tl->clips[i].frames = frames + first[i];
        tl->clips[i].count = clip_count[i];
        tl->sources[i].obj = &tl->clips[i];
        tl->sources[i].funcs = &clip_funcs;
        workspace_add_item( tl->workspace, &tl->sources[i], x[i], length[i], offset[i], z[i], NULL );
    }
    workspace_as_video_source( tl->workspace, &tl->as_source );
    return tl;
}

EXPORT void client_timeline_get_frame_f16( client_timeline *tl, int frame_index, rgba_frame_f16 *out ) {
    video_get_frame_f16( &tl->as_source, frame_index, out );
}

EXPORT void client_timeline_free( client_timeline *tl ) {
    if( !tl ) return;
    workspace_free( tl->workspace );
    free( tl->sources );
    free( tl->clips );
    free( tl );
}

/* ---- 8-bit pictures as a coded_image_source -------------------------------------------------- */

/*
 * A ring of 8-bit pictures in host memory behind the reference's coded_image_source interface (framework.h:513-523):
 * getFrame( frame ) hands out a coded_image whose planes POINT INTO the ring (a decoder handing out its frame buffer;
 * src/process/CodedImageSource.c:118-223 copies instead because its pictures come from Python objects) and whose
 * free_func releases only the small descriptor.  The frame-parallel scheduler (cv_timeline_render) pulls these.
 */
typedef struct {
    uint8_t *const *planes;     /* [count * plane_count] */
    int count, plane_count;
    int stride[CODED_IMAGE_MAX_PLANES], lines[CODED_IMAGE_MAX_PLANES];
    coded_image_source as_source;
    coded_image_source_funcs funcs;
    long pulls;                 /* how many pictures were handed out (tests) */
} coded_clip;

static void coded_descriptor_free( void *image ) { free( image ); }

static coded_image *coded_clip_get_frame( void *self, int frame, int quality_hint ) {
    coded_clip *clip = (coded_clip *) self;
    (void) quality_hint;
    int i = frame % clip->count;
    if( i < 0 ) i += clip->count;
    coded_image *image = (coded_image *) calloc( 1, sizeof(coded_image) );
    if( !image ) return NULL;
    for( int p = 0; p < clip->plane_count; p++ ) {
        image->data[p] = clip->planes[i * clip->plane_count + p];
        image->stride[p] = clip->stride[p];
        image->line_count[p] = clip->lines[p];
    }
    image->free_func = coded_descriptor_free;
    __sync_fetch_and_add( &clip->pulls, 1 );
    return image;
}

EXPORT coded_clip *client_coded_clip_create( uint8_t *const *planes, int count, int plane_count, const int *stride, const int *lines ) {
    coded_clip *clip = (coded_clip *) calloc( 1, sizeof(*clip) );
    if( !clip ) return NULL;
    clip->planes = planes; clip->count = count; clip->plane_count = plane_count;
    for( int p = 0; p < plane_count && p < CODED_IMAGE_MAX_PLANES; p++ ) { clip->stride[p] = stride[p]; clip->lines[p] = lines[p]; }
    clip->funcs.flags = 0;
    clip->funcs.getFrame = coded_clip_get_frame;
    clip->as_source.obj = clip;
    clip->as_source.funcs = &clip->funcs;
    return clip;
}

EXPORT coded_image_source *client_coded_clip_source( coded_clip *clip ) { return &clip->as_source; }
EXPORT long client_coded_clip_pulls( coded_clip *clip ) { return clip->pulls; }
EXPORT void client_coded_clip_free( coded_clip *clip ) { free( clip ); }

/* a clip of host f16 frames as a plain video_source (for cv_pull.video) */
EXPORT video_source *client_video_clip_create( rgba_frame_f16 *frames, int count ) {
    struct both { video_source source; clip_source clip; } *b = (struct both *) calloc( 1, sizeof(*b) );
    if( !b ) return NULL;
    b->clip.frames = frames; b->clip.count = count;
    b->source.obj = &b->clip; b->source.funcs = &clip_funcs;
    return &b->source;
}
EXPORT void client_video_clip_free( video_source *source ) { free( source ); }

/* ---- several threads pulling frames --------------------------------------------------------- */

typedef struct {
    client_timeline *tl;
    rgba_frame_f16 *outs;       /* one output frame per thread */
    int first_frame, count;
    int next;                   /* next frame to hand out */
    pthread_mutex_t lock;
} pull_run;

typedef struct { pull_run *run; int slot; } pull_worker;

static void *pull_thread( void *arg ) {
    pull_worker *w = (pull_worker *) arg;
    pull_run *run = w->run;
    for( ;; ) {
        pthread_mutex_lock( &run->lock );
        const int job = run->next < run->count ? run->next++ : -1;
        pthread_mutex_unlock( &run->lock );
        if( job < 0 ) return NULL;
        video_get_frame_f16( &run->tl->as_source, run->first_frame + job, &run->outs[w->slot] );
    }
}

static int64_t now_ns( void ) {
    struct timespec ts;
    clock_gettime( CLOCK_MONOTONIC, &ts );
    return (int64_t) ts.tv_sec * NS_PER_SEC + ts.tv_nsec;
}

/* Pulls frames [first_frame, first_frame + count) on `threads` threads (thread t writes outs[t]); returns elapsed ns. */
EXPORT int64_t client_time_timeline( client_timeline *tl, int first_frame, int count, int threads, rgba_frame_f16 *outs ) {
    pull_run run = { tl, outs, first_frame, count, 0, PTHREAD_MUTEX_INITIALIZER };
    pthread_t *ids = (pthread_t *) calloc( (size_t) threads, sizeof(pthread_t) );
    pull_worker *workers = (pull_worker *) calloc( (size_t) threads, sizeof(pull_worker) );
    const int64_t begin = now_ns();
    for( int t = 0; t < threads; t++ ) {
        workers[t].run = &run;
        workers[t].slot = t;
        pthread_create( &ids[t], NULL, pull_thread, &workers[t] );
    }
    for( int t = 0; t < threads; t++ )
        pthread_join( ids[t], NULL );
    const int64_t elapsed = now_ns() - begin;
    free( ids );
    free( workers );
    return elapsed;
}

/* ------------------------------------------------------------- host copy rate (a diagnostic for bench.py)
 *
 * Every frame a host-frame source hands over is a memcpy of the whole frame into the caller's buffer: the plug-in
 * path of a workspace over N such sources moves N + 1 frames through the CPU per output frame, whatever the GPU
 * does.  This measures what `threads` threads copying `bytes`-sized pageable buffers at once achieve on this host,
 * so that the e2e_plugin figure can be read against it. */
typedef struct copy_worker { unsigned char *from, *to; size_t bytes; int repeats; } copy_worker;

static void *copy_thread( void *arg ) {
    copy_worker *w = (copy_worker *) arg;
    for( int i = 0; i < w->repeats; i++ )
        memcpy( w->to, w->from, w->bytes );
    return NULL;
}

/* returns the copy rate in bytes copied per second over all threads (0 on allocation failure) */
EXPORT double client_host_copy_rate( int threads, size_t bytes, int repeats ) {
    pthread_t *ids = (pthread_t *) calloc( (size_t) threads, sizeof(pthread_t) );
    copy_worker *workers = (copy_worker *) calloc( (size_t) threads, sizeof(copy_worker) );
    int ok = ids && workers;
    for( int t = 0; t < threads && ok; t++ ) {
        workers[t].from = (unsigned char *) malloc( bytes );
        workers[t].to = (unsigned char *) malloc( bytes );
        workers[t].bytes = bytes; workers[t].repeats = repeats;
        ok = workers[t].from && workers[t].to;
        if( ok ) { memset( workers[t].from, t + 1, bytes ); memset( workers[t].to, 0, bytes ); }     /* pages touched */
    }
    double rate = 0.0;
    if( ok ) {
        const int64_t begin = now_ns();
        for( int t = 0; t < threads; t++ ) pthread_create( &ids[t], NULL, copy_thread, &workers[t] );
        for( int t = 0; t < threads; t++ ) pthread_join( ids[t], NULL );
        const int64_t elapsed = now_ns() - begin;
        rate = (double) bytes * repeats * threads / ( (double) elapsed * 1e-9 );
    }
    for( int t = 0; t < threads && workers; t++ ) { free( workers[t].from ); free( workers[t].to ); }
    free( ids ); free( workers );
    return rate;
}
