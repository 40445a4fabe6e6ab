/*
 * canvas_b200/csrc/cv_timeline.c -- the frame-parallel multi-GPU scheduler (cv_timeline_*).
 *
 * What it replaces: the reference renders a timeline by calling workspace_get_frame_f32 once per
 * frame on one thread (/root/reference/src/cprocess/workspace.c:494-550), helped only by the
 * sequential "composite list" cache under a mutex (:243-307) and a 2-thread pull queue
 * (src/process/VideoPullQueue.c:102-110).  Frames never interact -- a frame is a pure function of
 * its index (docs/sphinx/framework.rst:16-18; workspace.c:530,542 is index arithmetic) -- so here
 *
 *   - cv_timeline_job() is the STATELESS form of the compositor's item walk: items showing at
 *     frame f are those with x <= f < x + length, ordered bottom-first by ascending z, each reading
 *     its source at f - x + offset (workspace.c:504-547); an item may be a crossfade of two clips
 *     with a linear mix ramp (VideoMixFilter + AnimationFunc, src/process/AnimationFunc.c:439-446);
 *   - cv_timeline_render() partitions [first, last] block-cyclically over n GPUs
 *     (gpu = ((f - first) / chunk) mod n), one host thread and one stream per GPU, no collective
 *     and no shared state except the caller's sink.  Every frame is one fused launch
 *     (cv_render_frames) on device-resident inputs.
 *
 * Plain C + pthreads; CUDA is reached only through the cv_* ABI.
 */
#define _GNU_SOURCE
#include "cv_internal.h"

#include <stdlib.h>
#include <string.h>
#include <time.h>

int cv_timeline_device_of( int64_t frame, int64_t first, int chunk, int n_gpus ) {
    if( n_gpus < 1 || chunk < 1 || frame < first ) return -1;
    return (int)( ((frame - first) / chunk) % n_gpus );
}

int64_t cv_timeline_frames_of( int slot, int64_t first, int64_t last, int chunk, int n_gpus ) {
    if( n_gpus < 1 || chunk < 1 || slot < 0 || slot >= n_gpus || last < first ) return 0;
    const int64_t total = last - first + 1, round = (int64_t) chunk * n_gpus;
    const int64_t whole = total / round, rest = total % round;
    int64_t mine = whole * chunk;
    const int64_t begin = (int64_t) slot * chunk;
    if( rest > begin ) mine += (rest - begin < chunk) ? rest - begin : chunk;
    return mine;
}

/* AnimationFunc.c:439-446 for a two-point linear segment: (right * (f - left_frame) + left * (right_frame - f)) / distance
 * in double -- the reference's own expression, so that the (float) below rounds the same double -- then (float) */
static float ramp_value( const cv_timeline_item *it, int64_t source_frame ) {
    if( it->ramp_length <= 0 ) return it->mix_first;
    if( source_frame <= it->ramp_start ) return it->mix_first;
    if( source_frame >= it->ramp_start + it->ramp_length ) return it->mix_last;
    const double f = (double) source_frame, left = (double) it->ramp_start, right = (double)( it->ramp_start + it->ramp_length );
    return (float)( ((double) it->mix_last * (f - left) + (double) it->mix_first * (right - f)) / (right - left) );
}

static const cv_source *clip_frame( const cv_clip *clip, int64_t source_frame ) {
    if( !clip || clip->count <= 0 || !clip->frames ) return NULL;
    int64_t i = source_frame % clip->count;
    if( i < 0 ) i += clip->count;
    return &clip->frames[i];
}

typedef struct Showing { const cv_timeline_item *item; int order; } Showing;

/* bottom first: ascending z; among equals the later-added item is lower (cv_host_workspace.c: by_depth) */
static int by_depth( const void *pa, const void *pb ) {
    const Showing *a = (const Showing *) pa, *b = (const Showing *) pb;
    if( a->item->z != b->item->z ) return a->item->z < b->item->z ? -1 : 1;
    return a->order > b->order ? -1 : (a->order < b->order);
}

int cv_timeline_showing( const cv_timeline_item *items, int count, int64_t frame, int *indices, int capacity ) {
    if( count < 0 || (count && !items) ) return cvi_fail( CV_ERR_ARGUMENT, "cv_timeline_showing: bad items" );
    Showing *show = (Showing *) malloc( sizeof(Showing) * (size_t)( count ? count : 1 ) );
    if( !show ) return cvi_fail( CV_ERR_MEMORY, "cv_timeline_showing: out of host memory" );

    int n = 0;
    for( int i = 0; i < count; i++ ) {
        if( items[i].x <= frame && frame < items[i].x + items[i].length ) {         /* workspace.c:243-307 as a query */
            show[n].item = &items[i];
            show[n].order = i;
            n++;
        }
    }
    qsort( show, (size_t) n, sizeof(Showing), by_depth );
    for( int i = 0; i < n && i < capacity; i++ )
        indices[i] = show[i].order;
    free( show );
    return n;
}

int cv_timeline_job( const cv_timeline_item *items, int count, int slot, int64_t frame, cv_frame_job *job ) {
    if( !job || slot < 0 ) return cvi_fail( CV_ERR_ARGUMENT, "cv_timeline_job: bad argument" );

    int order[CV_MAX_LAYERS + 1];
    const int n = cv_timeline_showing( items, count, frame, order, CV_MAX_LAYERS + 1 );
    if( n < 0 ) return n;
    if( n > CV_MAX_LAYERS + 1 )
        return cvi_fail( CV_ERR_UNSUPPORTED, "cv_timeline_job: %d items show at frame %lld (at most %d)", n, (long long) frame, CV_MAX_LAYERS + 1 );

    job->a.kind = CV_SRC_NONE;
    job->b.kind = CV_SRC_NONE;
    job->mix_b = 0.0f;
    job->layer_count = 0;

    for( int k = 0; k < n; k++ ) {
        const cv_timeline_item *it = &items[order[k]];
        const int64_t source_frame = frame - it->x + it->offset;                    /* workspace.c:530,542 */
        const cv_source *sa = clip_frame( it->clips ? &it->clips[slot] : NULL, source_frame );
        const cv_source *sb = it->clips_b ? clip_frame( &it->clips_b[slot], source_frame ) : NULL;
        if( !sa ) return cvi_fail( CV_ERR_ARGUMENT, "cv_timeline_job: item %d has no clip for device slot %d", order[k], slot );

        const int simple_f16 = !sb && sa->kind == CV_SRC_F16;

        if( k == 0 && !simple_f16 ) {
            /* the bottom of the stack may be any source, or the crossfade of two (VideoMixFilter) */
            job->a = *sa;
            if( sb ) { job->b = *sb; job->mix_b = ramp_value( it, source_frame ); }
        }
        else if( simple_f16 ) {
            if( job->layer_count >= CV_MAX_LAYERS )
                return cvi_fail( CV_ERR_UNSUPPORTED, "cv_timeline_job: more than %d f16 layers at frame %lld", CV_MAX_LAYERS, (long long) frame );
            cv_layer *L = &job->layers[job->layer_count++];
            L->frame = sa->f16;
            L->mix = 1.0f;                                                          /* workspace.c:547: over with mix 1.0 */
            L->reserved = 0;
        }
        else
            return cvi_fail( CV_ERR_UNSUPPORTED, "cv_timeline_job: item %d above the bottom must be a plain f16 clip", order[k] );
    }
    return n;
}

/* ------------------------------------------------------------------ the scheduler */

typedef struct Worker {
    const cv_timeline_run *run;
    int slot;
    pthread_t thread;
    int rc;
    char error[256];
    cv_timeline_stats stats;
} Worker;

enum { CV_TIMELINE_GROUP = 16 };        /* frames per cv_render_frames() call of a GPU's worker */

static double now_ms( void ) {
    struct timespec ts;
    clock_gettime( CLOCK_MONOTONIC, &ts );
    return ts.tv_sec * 1e3 + ts.tv_nsec * 1e-6;
}

static void *worker_main( void *arg ) {
    Worker *w = (Worker *) arg;
    const cv_timeline_run *run = w->run;
    const int device = run->devices ? run->devices[w->slot] : w->slot;
    const cv_timeline_target *target = &run->targets[w->slot];
    cv_stream stream = NULL;
    cv_event begin = NULL, end = NULL, *slot_done = NULL;
    cv_frame_job *jobs = NULL;
    int64_t *frames = NULL;
    int rc;

    w->stats.device = device;
    if( (rc = cv_init( device )) ) goto fail;
    if( (rc = cv_stream_create( &stream )) ) goto fail;
    if( (rc = cv_event_create( &begin )) || (rc = cv_event_create( &end )) ) goto fail;
    if( target->ring < 1 || !target->out ) { rc = cvi_fail( CV_ERR_ARGUMENT, "cv_timeline_render: device slot %d has no output ring", w->slot ); goto fail; }

    slot_done = (cv_event *) calloc( (size_t) target->ring, sizeof(cv_event) );
    if( !slot_done ) { rc = cvi_fail( CV_ERR_MEMORY, "cv_timeline_render: out of host memory" ); goto fail; }
    for( int i = 0; i < target->ring; i++ )
        if( (rc = cv_event_create( &slot_done[i] )) ) goto fail;

    const double t0 = now_ms();
    const uint64_t launches0 = cv_kernel_launches();
    if( (rc = cv_event_record( begin, stream )) ) goto fail;

    /* Frames are submitted in groups of up to `group` (one cv_render_frames() call: frames of one shape become ONE
     * persistent launch, tables staged once); a group never holds more frames than the output ring has slots. */
    int group = target->ring < CV_TIMELINE_GROUP ? target->ring : CV_TIMELINE_GROUP;
    jobs = (cv_frame_job *) calloc( (size_t) group, sizeof(cv_frame_job) );
    frames = (int64_t *) calloc( (size_t) group, sizeof(int64_t) );
    if( !jobs || !frames ) { rc = cvi_fail( CV_ERR_MEMORY, "cv_timeline_render: out of host memory" ); goto fail; }

    int64_t produced = 0;
    int64_t f = run->first;
    while( f <= run->last ) {
        int n = 0;
        for( ; f <= run->last && n < group; f++ ) {
            if( cv_timeline_device_of( f, run->first, run->chunk, run->n_gpus ) != w->slot )
                continue;

            const int ring_slot = (int)( (produced + n) % target->ring );
            if( produced + n >= target->ring )              /* the sink's copies out of this slot have finished */
                if( (rc = cv_event_sync( slot_done[ring_slot] )) ) goto fail;

            cv_frame_job *job = &jobs[n];
            memset( job, 0, sizeof(*job) );
            rc = cv_timeline_job( run->items, run->item_count, w->slot, f, job );
            if( rc < 0 ) goto fail;

            job->out = &target->out[ring_slot];
            if( target->out_u8 ) {
                job->out_u8 = target->out_u8[ring_slot];
                job->u8_table = target->u8_table;
            }
            frames[n++] = f;
        }
        if( n == 0 )
            break;
        if( (rc = cv_render_frames( jobs, n, stream )) ) goto fail;

        for( int i = 0; i < n; i++ ) {
            if( run->sink ) {
                rc = run->sink( run->user, w->slot, frames[i], &jobs[i], stream );
                if( rc ) { cvi_fail( rc, "cv_timeline_render: the sink failed at frame %lld", (long long) frames[i] ); goto fail; }
            }
            if( (rc = cv_event_record( slot_done[(int)( (produced + i) % target->ring )], stream )) ) goto fail;
        }
        produced += n;
    }

    if( (rc = cv_event_record( end, stream )) ) goto fail;
    if( (rc = cv_stream_sync( stream )) ) goto fail;
    w->stats.frames = produced;
    w->stats.wall_ms = now_ms() - t0;
    w->stats.launches = cv_kernel_launches() - launches0;      /* all threads' launches in that interval */
    {
        float ms = 0.0f;
        if( (rc = cv_event_elapsed_ms( begin, end, &ms )) ) goto fail;
        w->stats.device_ms = ms;
    }
    rc = CV_OK;

fail:
    if( rc ) {
        strncpy( w->error, cv_last_error(), sizeof(w->error) - 1 );
        if( stream ) cv_stream_sync( stream );
    }
    if( slot_done ) {
        for( int i = 0; i < target->ring; i++ )
            if( slot_done[i] ) cv_event_destroy( slot_done[i] );
        free( slot_done );
    }
    free( jobs );
    free( frames );
    if( begin ) cv_event_destroy( begin );
    if( end ) cv_event_destroy( end );
    if( stream ) cv_stream_destroy( stream );
    w->rc = rc;
    return NULL;
}

int cv_timeline_render( const cv_timeline_run *run, cv_timeline_stats *stats ) {
    if( !run || run->n_gpus < 1 || run->n_gpus > CV_MAX_DEVICES || run->chunk < 1 || !run->targets )
        return cvi_fail( CV_ERR_ARGUMENT, "cv_timeline_render: bad run description" );
    if( run->item_count < 0 || (run->item_count && !run->items) )
        return cvi_fail( CV_ERR_ARGUMENT, "cv_timeline_render: bad items" );
    if( run->last < run->first )
        return CV_OK;

    Worker *workers = (Worker *) calloc( (size_t) run->n_gpus, sizeof(Worker) );
    if( !workers ) return cvi_fail( CV_ERR_MEMORY, "cv_timeline_render: out of host memory" );

    int started = 0, rc = CV_OK;
    for( int g = 0; g < run->n_gpus; g++ ) {
        workers[g].run = run;
        workers[g].slot = g;
        if( pthread_create( &workers[g].thread, NULL, worker_main, &workers[g] ) ) {
            rc = cvi_fail( CV_ERR_MEMORY, "cv_timeline_render: pthread_create failed" );
            break;
        }
        started++;
    }
    for( int g = 0; g < started; g++ ) {
        pthread_join( workers[g].thread, NULL );
        if( workers[g].rc && !rc )
            rc = cvi_fail( workers[g].rc, "device slot %d: %s", g, workers[g].error );
        if( stats ) stats[g] = workers[g].stats;
    }
    free( workers );
    return rc;
}
