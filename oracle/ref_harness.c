/*
 * oracle/ref_harness.c -- TEST INFRASTRUCTURE ONLY (linked into oracle/_ref/libcanvasref.so).
 *
 * Thin drivers AROUND the unmodified reference functions so tests and the CPU
 * baseline can call them with plain pointers: nothing here re-implements pixel
 * arithmetic.  Every pixel produced through this file comes from the reference's
 * own code compiled from /root/reference/src/cprocess.
 *
 *  - frame-as-source vtable: same as src/process/RgbaFrameF16.c:27-38
 *    (get_frame = video_copy_frame_f16 from the stored frame).
 *  - ref_workspace_*: build a reference workspace (src/cprocess/workspace.c) from
 *    f16 frames and pull one composited frame through video_get_frame_f16
 *    (src/cprocess/main.c:34-77 -> workspace_get_frame_f32, workspace.c:494-550).
 *  - ref_cross_pull_f16: VideoMixFilter's path (src/process/VideoMixFilter.c:62-67
 *    -> video_mix_cross_f32_pull, video_mix.c:46-71) pulled as f16.
 *  - ref_time_*: pthread frame-parallel timing loops (one frame per thread at a
 *    time; the reference itself renders a frame on one thread).
 */
#define _GNU_SOURCE
#include "framework.h"
#include <pthread.h>
#include <time.h>

/* ---- frame as source ----------------------------------------------------- */
static void frame_source_get_frame( void *self, int frame_index, rgba_frame_f16 *frame ) {
    (void) frame_index;
    video_copy_frame_f16( frame, (rgba_frame_f16 *) self );
}

static video_frame_source_funcs frame_source_funcs = {
    .get_frame = frame_source_get_frame,
};

/* A ring of frames as a source: frame_index selects ring[frame_index mod count]. */
typedef struct {
    rgba_frame_f16 *frames;
    int count;
} ring_source;

static void ring_source_get_frame( void *self, int frame_index, rgba_frame_f16 *frame ) {
    ring_source *ring = (ring_source *) self;
    int i = frame_index % ring->count;
    if( i < 0 )
        i += ring->count;
    video_copy_frame_f16( frame, &ring->frames[i] );
}

static video_frame_source_funcs ring_source_funcs = {
    .get_frame = ring_source_get_frame,
};

EXPORT void ref_init( void ) {
    init_half();
}

EXPORT int64_t ref_now_ns( void ) {
    struct timespec ts;
    clock_gettime( CLOCK_MONOTONIC, &ts );
    return (int64_t) ts.tv_sec * INT64_C(1000000000) + ts.tv_nsec;
}

/* ---- workspace composite -------------------------------------------------- */
/*
 * layers[i] is an f16 frame (data/full_window/current_window set by the caller),
 * added with x=0, length=1<<30, offset=0, z=z[i].  Pulls frame 0 as f16 into out.
 */
EXPORT void ref_workspace_composite_f16( rgba_frame_f16 *layers, const int64_t *z, int count, rgba_frame_f16 *out ) {
    workspace_t *ws = workspace_create();
    video_source *sources = calloc( count ? count : 1, sizeof(video_source) );

    for( int i = 0; i < count; i++ ) {
        sources[i].obj = &layers[i];
        sources[i].funcs = &frame_source_funcs;
        workspace_add_item( ws, &sources[i], 0, INT64_C(1) << 30, 0, z[i], NULL );
    }

    video_source wsrc;
    workspace_as_video_source( ws, &wsrc );
    video_get_frame_f16( &wsrc, 0, out );

    workspace_free( ws );
    free( sources );
}

/*
 * General timeline form: item i shows ring i (ring_counts[i] frames starting at
 * rings[ring_starts[i]]) at timeline frames [x, x+length), source frame index
 * frame - x + offset (workspace.c:530,542).
 */
typedef struct {
    workspace_t *ws;
    video_source wsrc;
    video_source *sources;
    ring_source *rings;
    int count;
} ref_timeline;

EXPORT ref_timeline *ref_timeline_create( rgba_frame_f16 *frames, const int *ring_starts, const int *ring_counts,
        const int64_t *x, const int64_t *length, const int64_t *offset, const int64_t *z, int count ) {
    ref_timeline *tl = calloc( 1, sizeof(ref_timeline) );
    tl->ws = workspace_create();
    tl->sources = calloc( count ? count : 1, sizeof(video_source) );
    tl->rings = calloc( count ? count : 1, sizeof(ring_source) );
    tl->count = count;

    for( int i = 0; i < count; i++ ) {
        tl->rings[i].frames = frames + ring_starts[i];
        tl->rings[i].count = ring_counts[i];
        tl->sources[i].obj = &tl->rings[i];
        tl->sources[i].funcs = &ring_source_funcs;
        workspace_add_item( tl->ws, &tl->sources[i], x[i], length[i], offset[i], z[i], NULL );
    }

    workspace_as_video_source( tl->ws, &tl->wsrc );
    return tl;
}

EXPORT void ref_timeline_get_frame_f16( ref_timeline *tl, int frame_index, rgba_frame_f16 *out ) {
    video_get_frame_f16( &tl->wsrc, frame_index, out );
}

EXPORT void ref_timeline_free( ref_timeline *tl ) {
    workspace_free( tl->ws );
    free( tl->sources );
    free( tl->rings );
    free( tl );
}

/* ---- crossfade pulled as f16 ---------------------------------------------- */
typedef struct {
    video_source a, b;
    float mix_b;
} cross_filter;

static void cross_filter_get_frame_32( void *self, int frame_index, rgba_frame_f32 *frame ) {
    cross_filter *f = (cross_filter *) self;
    video_mix_cross_f32_pull( frame, &f->a, frame_index, &f->b, frame_index, f->mix_b );
}

static video_frame_source_funcs cross_filter_funcs = {
    .get_frame_32 = cross_filter_get_frame_32,
};

EXPORT void ref_cross_pull_f16( rgba_frame_f16 *a, rgba_frame_f16 *b, float mix_b, rgba_frame_f16 *out ) {
    cross_filter f = { { a, &frame_source_funcs }, { b, &frame_source_funcs }, mix_b };
    video_source src = { &f, &cross_filter_funcs };
    video_get_frame_f16( &src, 0, out );
}

/* ---- 8-bit encodes (loops lifted around the reference's LUT functions) ----- */
/*
 * widget soft path, src/cprocess/widget_gl.c:290-307: sRGB transfer in place on the
 * current window (all four channels), then ramp[] per channel.  The ramp is built by
 * the caller's restatement of widget_gl.c:949-969 (that function is welded to the
 * widget object and cannot be called without GL); the transfer is the reference's.
 */
EXPORT void ref_encode_u8( rgba_frame_f16 *frame, const uint8_t *ramp, rgba_u8 *target ) {
    v2i size;
    box2i_get_size( &frame->full_window, &size );

    for( int y = frame->current_window.min.y; y <= frame->current_window.max.y; y++ ) {
        rgba_u8 *target_row = target + (y - frame->full_window.min.y) * size.x - frame->full_window.min.x;
        rgba_f16 *source_row = video_get_pixel_f16( frame, 0, y );

        video_transfer_linear_to_sRGB( &source_row[frame->current_window.min.x].r,
            &source_row[frame->current_window.min.x].r,
            (frame->current_window.max.x - frame->current_window.min.x + 1) * 4 );

        for( int x = frame->current_window.min.x; x <= frame->current_window.max.x; x++ ) {
            target_row[x].r = ramp[source_row[x].r];
            target_row[x].g = ramp[source_row[x].g];
            target_row[x].b = ramp[source_row[x].b];
            target_row[x].a = ramp[source_row[x].a];
        }
    }
}

/* ---- frame-parallel timing ------------------------------------------------- */
typedef void (*ref_job_func)( void *closure, int job );

typedef struct {
    ref_job_func func;
    void *closure;
    int next, count;
    pthread_mutex_t mutex;
} job_pool;

static void *job_thread( void *arg ) {
    job_pool *pool = (job_pool *) arg;
    for( ;; ) {
        pthread_mutex_lock( &pool->mutex );
        int job = pool->next < pool->count ? pool->next++ : -1;
        pthread_mutex_unlock( &pool->mutex );
        if( job < 0 )
            return NULL;
        pool->func( pool->closure, job );
    }
}

static int64_t run_jobs( ref_job_func func, void *closure, int count, int threads ) {
    job_pool pool = { func, closure, 0, count, PTHREAD_MUTEX_INITIALIZER };
    pthread_t *tids = calloc( threads, sizeof(pthread_t) );
    int64_t start = ref_now_ns();
    for( int i = 0; i < threads; i++ )
        pthread_create( &tids[i], NULL, job_thread, &pool );
    for( int i = 0; i < threads; i++ )
        pthread_join( tids[i], NULL );
    int64_t end = ref_now_ns();
    free( tids );
    return end - start;
}

typedef struct {
    ref_timeline *tl;
    rgba_frame_f16 *outs;       /* one per thread-slot: outs[job % out_count] must not collide -> out_count >= threads */
    int first_frame;
    int out_count;
    pthread_mutex_t *locks;
} timeline_job;

static void timeline_job_func( void *closure, int job ) {
    timeline_job *tj = (timeline_job *) closure;
    int slot = job % tj->out_count;
    pthread_mutex_lock( &tj->locks[slot] );
    video_get_frame_f16( &tj->tl->wsrc, tj->first_frame + job, &tj->outs[slot] );
    pthread_mutex_unlock( &tj->locks[slot] );
}

/* Renders frames [first, first+count) on `threads` threads; returns elapsed ns. */
EXPORT int64_t ref_time_timeline( ref_timeline *tl, int first_frame, int count, int threads, rgba_frame_f16 *outs, int out_count ) {
    timeline_job tj = { tl, outs, first_frame, out_count, calloc( out_count, sizeof(pthread_mutex_t) ) };
    for( int i = 0; i < out_count; i++ )
        pthread_mutex_init( &tj.locks[i], NULL );
    int64_t ns = run_jobs( timeline_job_func, &tj, count, threads );
    free( tj.locks );
    return ns;
}
