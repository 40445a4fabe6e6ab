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
 *   - cv_timeline_render() hands chunks of consecutive frames to n GPUs, one host thread per GPU, no
 *     collective.  By default a GPU that finishes a chunk takes the next one from a shared atomic counter
 *     (no GPU is left with a long tail; CV_TIMELINE_STATIC keeps the block-cyclic deal
 *     gpu = ((f - first) / chunk) mod n).  Every frame is one fused launch (cv_render_frames);
 *   - an item's pictures are either resident on every GPU (cv_clip) or PULLED from the host while the
 *     timeline renders (cv_pull: the reference's coded_image_source.getFrame, framework.h:513-523 /
 *     DVReconstructionFilter.c:53-71, or a video_source pulled as f16, workspace.c:530-547): each GPU's
 *     worker stages them through its own ring of pinned buffers and a dedicated COPY stream, one
 *     cudaMemcpyAsync per plane, the compute stream waiting on an event -- so the upload of frame f+1
 *     overlaps the kernel of frame f.  Worker threads and their pinned rings are bound to the NUMA node of
 *     the GPU they drive.
 *
 * Plain C + pthreads; CUDA is reached only through the cv_* ABI.
 */
#define _GNU_SOURCE
#include "cv_internal.h"

#include <stdlib.h>
#include <string.h>
#include <time.h>
#include <sched.h>
#include <stdio.h>
#include <stdatomic.h>

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

        /* a pulled item is described by its format only: the scheduler fills the pointers once the picture is in HBM */
        cv_source pulled_a, pulled_b;
        const cv_source *sa = NULL, *sb = NULL;
        if( it->pull ) {
            memset( &pulled_a, 0, sizeof(pulled_a) );
            if( it->pull->coded ) { pulled_a.kind = CV_SRC_CODED; pulled_a.coded = it->pull->format; memset( pulled_a.coded.plane, 0, sizeof(pulled_a.coded.plane) ); }
            else { pulled_a.kind = CV_SRC_F16; pulled_a.f16.full_window = it->pull->full_window; box2i_set_empty( &pulled_a.f16.current_window ); }
            sa = &pulled_a;
        }
        else
            sa = clip_frame( it->clips ? &it->clips[slot] : NULL, source_frame );
        if( it->pull_b ) {
            memset( &pulled_b, 0, sizeof(pulled_b) );
            if( it->pull_b->coded ) { pulled_b.kind = CV_SRC_CODED; pulled_b.coded = it->pull_b->format; memset( pulled_b.coded.plane, 0, sizeof(pulled_b.coded.plane) ); }
            else { pulled_b.kind = CV_SRC_F16; pulled_b.f16.full_window = it->pull_b->full_window; box2i_set_empty( &pulled_b.f16.current_window ); }
            sb = &pulled_b;
        }
        else if( it->clips_b )
            sb = clip_frame( &it->clips_b[slot], source_frame );
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

typedef struct Shared {
    const cv_timeline_run *run;
    atomic_llong next_chunk;            /* chunk stealing: the next undistributed chunk */
    atomic_int failed;
} Shared;

typedef struct Worker {
    Shared *shared;
    int slot;
    pthread_t thread;
    int rc;
    char error[256];
    cv_timeline_stats stats;
} Worker;

enum { CV_TIMELINE_GROUP = 16,          /* frames per cv_render_frames() call of a GPU's worker */
       CV_PULLS_MAX = CV_MAX_LAYERS + 2 };

static double now_ms( void ) {
    struct timespec ts;
    clock_gettime( CLOCK_MONOTONIC, &ts );
    return ts.tv_sec * 1e3 + ts.tv_nsec * 1e-6;
}

/* Runs the calling thread on the CPUs of the NUMA node the GPU's PCIe root belongs to, so that the pinned staging
 * memory it allocates next is local to the copy engine that reads it.  Silently does nothing where the box does not
 * say (single-node machines, containers without /sys). */
static void bind_to_device_node( int device ) {
    char bus[32] = { 0 }, path[128], line[4096];
    if( cudaDeviceGetPCIBusId( bus, (int) sizeof(bus), device ) != cudaSuccess ) { cudaGetLastError(); return; }
    for( char *c = bus; *c; c++ ) if( *c >= 'A' && *c <= 'Z' ) *c = (char)( *c - 'A' + 'a' );
    snprintf( path, sizeof(path), "/sys/bus/pci/devices/%s/numa_node", bus );
    FILE *f = fopen( path, "r" );
    int node = -1;
    if( f ) { if( fscanf( f, "%d", &node ) != 1 ) node = -1; fclose( f ); }
    if( node < 0 ) return;
    snprintf( path, sizeof(path), "/sys/devices/system/node/node%d/cpulist", node );
    f = fopen( path, "r" );
    if( !f ) return;
    if( !fgets( line, sizeof(line), f ) ) { fclose( f ); return; }
    fclose( f );

    cpu_set_t allowed, want;
    CPU_ZERO( &want );
    if( sched_getaffinity( 0, sizeof(allowed), &allowed ) ) return;
    int any = 0;
    for( char *tok = strtok( line, ",\n" ); tok; tok = strtok( NULL, ",\n" ) ) {
        int lo = 0, hi = 0;
        const int got = sscanf( tok, "%d-%d", &lo, &hi );
        if( got < 1 ) continue;
        if( got == 1 ) hi = lo;
        for( int c = lo; c <= hi && c < CPU_SETSIZE; c++ )
            if( CPU_ISSET( c, &allowed ) ) { CPU_SET( c, &want ); any = 1; }
    }
    if( any ) sched_setaffinity( 0, sizeof(want), &want );
}

/* one pulled picture of one ring entry: pinned host buffer, device buffer, and -- when the source handed out pinned
 * memory itself, which is then copied from directly -- the image to release once the copy has been consumed */
typedef struct Stage { void *pinned, *dev; size_t bytes; coded_image *held; } Stage;
typedef struct Entry { Stage stage[CV_PULLS_MAX]; } Entry;

/* device buffer (and, when the picture has to be staged, pinned host buffer) of at least `bytes`; both come from pools,
 * so a second cv_timeline_render() call finds the blocks of the first */
static int stage_reserve( Stage *s, size_t bytes, int need_pinned ) {
    int rc = CV_OK;
    if( s->bytes < bytes ) {
        if( s->pinned ) cvi_pinned_free( s->pinned );
        if( s->dev ) cv_free( s->dev, CV_MEM_DEVICE );
        s->pinned = s->dev = NULL; s->bytes = 0;
        rc = cv_malloc( &s->dev, bytes, CV_MEM_DEVICE );
        if( !rc ) s->bytes = bytes;
    }
    if( !rc && need_pinned && !s->pinned )
        rc = cvi_pinned_alloc( &s->pinned, s->bytes );
    return rc;
}

static void stage_release_held( Stage *s ) {
    if( s->held ) {
        if( s->held->free_func ) s->held->free_func( s->held );
        s->held = NULL;
    }
}

static int host_memory_is_pinned( const void *p ) {
    struct cudaPointerAttributes attr;
    if( cudaPointerGetAttributes( &attr, p ) != cudaSuccess ) { cudaGetLastError(); return 0; }
    return attr.type == cudaMemoryTypeHost;
}

/* coded_image_source.getFrame -> HBM: one cudaMemcpyAsync per plane on the copy stream (CodedImageSource.c:118-223 is
 * the producer side of this call, DVReconstructionFilter.c:53-71 its consumer in the reference) */
static int pull_coded( const cv_pull *pull, int64_t source_frame, Stage *st, cv_stream copy, cv_coded *out, uint64_t *bytes_moved ) {
    coded_image *image = pull->coded->funcs->getFrame( pull->coded->obj, (int) source_frame, 0 );
    if( !image ) return cvi_fail( CV_ERR_ARGUMENT, "cv_timeline_render: a coded image source returned no picture for frame %lld", (long long) source_frame );

    const int planes = pull->format.layout == CV_LAYOUT_PLANAR ? 3 : 1;
    size_t size[3] = { 0, 0, 0 }, total = 0;
    for( int i = 0; i < planes; i++ ) {
        int lines = image->line_count[i] < pull->format.height ? image->line_count[i] : pull->format.height;
        if( lines < 0 || !image->data[i] ) lines = 0;
        size[i] = ((size_t) image->stride[i] * (size_t) lines + 255) & ~(size_t) 255;
        total += size[i];
    }
    const int direct = total && host_memory_is_pinned( image->data[0] );        /* pinned already: copied from where it lies */
    int rc = stage_reserve( st, total ? total : 256, !direct );

    *out = pull->format;
    size_t at = 0;
    for( int i = 0; i < planes && !rc; i++ ) {
        const size_t bytes = (size_t) image->stride[i] * (size_t)( size[i] ? (image->line_count[i] < pull->format.height ? image->line_count[i] : pull->format.height) : 0 );
        const void *from = image->data[i];
        if( !direct && bytes ) {
            memcpy( (char *) st->pinned + at, image->data[i], bytes );
            from = (char *) st->pinned + at;
        }
        if( bytes ) rc = cv_upload( (char *) st->dev + at, from, bytes, copy );
        out->plane[i] = (uint8_t *) st->dev + at;
        out->stride[i] = image->stride[i];
        *bytes_moved += bytes;
        at += size[i];
    }
    for( int i = planes; i < 3; i++ ) { out->plane[i] = NULL; out->stride[i] = 0; }

    if( direct && !rc ) st->held = image;               /* released when this ring entry comes round again */
    else if( image->free_func ) image->free_func( image );
    return rc;
}

/* a video_source pulled as f16 into pinned memory, its valid rows sent up (workspace.c:530-547 pulls every item this way) */
static int pull_video( const cv_pull *pull, int64_t source_frame, Stage *st, cv_stream copy, rgba_frame_f16 *out, uint64_t *bytes_moved ) {
    v2i size;
    box2i_get_size( &pull->full_window, &size );
    const size_t pitch = (size_t) size.x * sizeof(rgba_f16), bytes = pitch * (size_t) size.y;
    int rc = stage_reserve( st, bytes ? bytes : 256, 1 );
    if( rc ) return rc;

    rgba_frame_f16 host;
    host.data = (rgba_f16 *) st->pinned;
    host.full_window = pull->full_window;
    host.current_window = pull->full_window;
    video_get_frame_f16( pull->video, (int) source_frame, &host );

    out->data = (rgba_f16 *) st->dev;
    out->full_window = pull->full_window;
    box2i_intersect( &out->current_window, &host.current_window, &pull->full_window );
    if( box2i_is_empty( &out->current_window ) ) { box2i_set_empty( &out->current_window ); return CV_OK; }

    const box2i *cw = &out->current_window;
    const size_t offset = (size_t)( cw->min.y - pull->full_window.min.y ) * pitch;       /* whole rows of the window */
    const size_t rows = (size_t)( cw->max.y - cw->min.y + 1 );
    *bytes_moved += rows * pitch;
    return cv_upload( (char *) st->dev + offset, (char *) st->pinned + offset, rows * pitch, copy );
}

static void *worker_main( void *arg ) {
    Worker *w = (Worker *) arg;
    Shared *shared = w->shared;
    const cv_timeline_run *run = shared->run;
    const int device = run->devices ? run->devices[w->slot] : w->slot;
    const cv_timeline_target *target = &run->targets[w->slot];
    cv_stream stream = NULL, copy = NULL;
    cv_event begin = NULL, end = NULL, uploaded = NULL, *slot_done = NULL;
    cv_frame_job *jobs = NULL;
    int64_t *frames = NULL;
    Entry *entries = NULL;
    int rc, any_pull = 0;

    w->stats.device = device;
    bind_to_device_node( device );
    if( (rc = cv_init( device )) ) goto fail;
    if( (rc = cv_stream_create( &stream )) || (rc = cv_stream_create( &copy )) ) goto fail;
    if( (rc = cv_event_create( &begin )) || (rc = cv_event_create( &end )) || (rc = cv_event_create( &uploaded )) ) goto fail;
    if( target->ring < 1 || !target->out ) { rc = cvi_fail( CV_ERR_ARGUMENT, "cv_timeline_render: device slot %d has no output ring", w->slot ); goto fail; }

    for( int i = 0; i < run->item_count; i++ )
        if( run->items[i].pull || run->items[i].pull_b ) any_pull = 1;

    slot_done = (cv_event *) calloc( (size_t) target->ring, sizeof(cv_event) );
    entries = (Entry *) calloc( (size_t) target->ring, sizeof(Entry) );
    if( !slot_done || !entries ) { rc = cvi_fail( CV_ERR_MEMORY, "cv_timeline_render: out of host memory" ); goto fail; }
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

    const int64_t total_chunks = (run->last - run->first) / run->chunk + 1;
    const int fixed_deal = (run->flags & CV_TIMELINE_STATIC) != 0;
    int64_t produced = 0, my_chunk = w->slot;
    int64_t next = 0, chunk_last = -1;          /* what is left of the chunk this worker holds */

    for( ;; ) {
        /* a group: up to `group` frames, continuing into the worker's next chunk when the one it holds runs out */
        int n = 0;
        while( n < group ) {
            if( next > chunk_last ) {
                /* the next chunk of consecutive frames: stolen from the shared counter, or this slot's turn of the fixed deal */
                int64_t chunk_index;
                if( fixed_deal ) { chunk_index = my_chunk; my_chunk += run->n_gpus; }
                else chunk_index = atomic_fetch_add( &shared->next_chunk, 1 );
                if( chunk_index >= total_chunks || atomic_load( &shared->failed ) )
                    break;
                next = run->first + chunk_index * run->chunk;
                chunk_last = (next + run->chunk - 1 < run->last) ? next + run->chunk - 1 : run->last;
            }
            const int64_t f = next++;

            const int ring_slot = (int)( (produced + n) % target->ring );
            if( produced + n >= target->ring )              /* the kernels and the sink's copies that used this slot have finished */
                if( (rc = cv_event_sync( slot_done[ring_slot] )) ) goto fail;

            cv_frame_job *job = &jobs[n];
            memset( job, 0, sizeof(*job) );
            rc = cv_timeline_job( run->items, run->item_count, w->slot, f, job );
            if( rc < 0 ) goto fail;

            if( any_pull ) {
                /* the same walk as cv_timeline_job: fill in what has to come from the host */
                int order[CV_MAX_LAYERS + 1];
                const int showing = cv_timeline_showing( run->items, run->item_count, f, order, CV_MAX_LAYERS + 1 );
                Entry *entry = &entries[ring_slot];
                int layer = 0, stage = 0;
                for( int k = 0; k < showing; k++ ) {
                    const cv_timeline_item *it = &run->items[order[k]];
                    const int64_t source_frame = f - it->x + it->offset;
                    const int bottom = (k == 0) && (job->a.kind != CV_SRC_NONE);
                    const cv_pull *pulls[2] = { it->pull, it->pull_b };
                    for( int side = 0; side < 2; side++ ) {
                        const cv_pull *p = pulls[side];
                        if( !p ) continue;
                        Stage *st = &entry->stage[stage++];
                        stage_release_held( st );
                        if( p->coded ) {
                            cv_source *dst = side ? &job->b : &job->a;
                            rc = pull_coded( p, source_frame, st, copy, &dst->coded, &w->stats.h2d_bytes );
                        }
                        else if( bottom ) {
                            cv_source *dst = side ? &job->b : &job->a;
                            rc = pull_video( p, source_frame, st, copy, &dst->f16, &w->stats.h2d_bytes );
                        }
                        else
                            rc = pull_video( p, source_frame, st, copy, &job->layers[layer].frame, &w->stats.h2d_bytes );
                        if( rc ) goto fail;
                    }
                    if( !bottom ) layer++;
                }
            }

            job->out = &target->out[ring_slot];
            if( target->out_u8 ) {
                job->out_u8 = target->out_u8[ring_slot];
                job->u8_table = target->u8_table;
            }
            frames[n++] = f;
        }
        if( n == 0 )
            break;

        if( any_pull ) {            /* the kernels of this group start when its uploads have landed */
            if( (rc = cv_event_record( uploaded, copy )) ) goto fail;
            if( (rc = cvi_cuda( cudaStreamWaitEvent( (cudaStream_t) stream, (cudaEvent_t) uploaded, 0 ), "cudaStreamWaitEvent" )) ) goto fail;
        }
        if( (rc = cv_render_frames( jobs, n, stream )) ) goto fail;

        for( int i = 0; i < n; i++ ) {
            if( run->sink ) {
                rc = run->sink( run->user, w->slot, frames[i], &jobs[i], stream );
                if( rc ) {
                    char why[160];
                    strncpy( why, cv_last_error(), sizeof(why) - 1 );
                    why[sizeof(why) - 1] = 0;
                    cvi_fail( rc, "cv_timeline_render: the sink failed at frame %lld (%s)", (long long) frames[i], why );
                    goto fail;
                }
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
        atomic_store( &shared->failed, 1 );
        strncpy( w->error, cv_last_error(), sizeof(w->error) - 1 );
        if( stream ) cv_stream_sync( stream );
    }
    if( copy ) cv_stream_sync( copy );
    if( entries ) {
        for( int i = 0; i < target->ring; i++ )
            for( int j = 0; j < CV_PULLS_MAX; j++ ) {
                Stage *st = &entries[i].stage[j];
                stage_release_held( st );
                if( st->pinned ) cvi_pinned_free( st->pinned );
                if( st->dev ) cv_free( st->dev, CV_MEM_DEVICE );
            }
        free( entries );
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
    if( uploaded ) cv_event_destroy( uploaded );
    if( stream ) cv_stream_destroy( stream );
    if( copy ) cv_stream_destroy( copy );
    w->rc = rc;
    return NULL;
}

int cv_timeline_render( const cv_timeline_run *run, cv_timeline_stats *stats ) {
    if( !run || run->n_gpus < 1 || run->n_gpus > CV_MAX_DEVICES || run->chunk < 1 || !run->targets )
        return cvi_fail( CV_ERR_ARGUMENT, "cv_timeline_render: bad run description" );
    if( run->item_count < 0 || (run->item_count && !run->items) )
        return cvi_fail( CV_ERR_ARGUMENT, "cv_timeline_render: bad items" );
    for( int i = 0; i < run->item_count; i++ ) {
        const cv_pull *pulls[2] = { run->items[i].pull, run->items[i].pull_b };
        for( int side = 0; side < 2; side++ ) {
            const cv_pull *p = pulls[side];
            if( !p ) continue;
            if( !p->coded == !p->video )
                return cvi_fail( CV_ERR_ARGUMENT, "cv_timeline_render: item %d: a cv_pull names exactly one of coded / video", i );
            if( p->coded && (!p->coded->funcs || !p->coded->funcs->getFrame) )
                return cvi_fail( CV_ERR_ARGUMENT, "cv_timeline_render: item %d: coded image source without getFrame", i );
        }
    }
    if( run->last < run->first )
        return CV_OK;

    Worker *workers = (Worker *) calloc( (size_t) run->n_gpus, sizeof(Worker) );
    if( !workers ) return cvi_fail( CV_ERR_MEMORY, "cv_timeline_render: out of host memory" );
    Shared shared;
    shared.run = run;
    atomic_init( &shared.next_chunk, 0 );
    atomic_init( &shared.failed, 0 );

    int started = 0, rc = CV_OK;
    for( int g = 0; g < run->n_gpus; g++ ) {
        workers[g].shared = &shared;
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
