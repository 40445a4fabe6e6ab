/*
 * canvas_b200/csrc/cv_api.c -- host side of the pixel operations of include/canvas_b200.h:
 * window arithmetic, short-circuits and the translation of a frame program (cv_frame_job)
 * into the pre-digested DevJob the kernels consume.  Plain C; no pixel is touched on the CPU.
 *
 * Window rules follow /root/reference/src/cprocess/video_mix.c (copy :27-44, :73-105, cross
 * :107-235 with the pull short-circuits :46-71, over :237-370), workspace.c:494-550 and
 * video_reconstruct.c:69-76, with the INTENDED semantics where the reference's 9-region walk
 * is buggy (min.x compared with min.y at video_mix.c:137,265; stale pixels for disjoint windows
 * :161-169): inside the union box, a pixel covered by nobody is zero.
 */
#include "cv_internal.h"

#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------ box helpers */

static DevBox to_dev( const box2i *b ) {
    DevBox d = { b->min.x, b->min.y, b->max.x, b->max.y };
    return d;
}
static box2i from_dev( const DevBox *d ) {
    box2i b;
    box2i_set( &b, d->x0, d->y0, d->x1, d->y1 );
    return b;
}
static int dev_empty( const DevBox *b ) { return b->x1 < b->x0 || b->y1 < b->y0; }
static DevBox dev_empty_box( void ) { DevBox d = { 0, 0, -1, -1 }; return d; }
static DevBox dev_intersect( const DevBox *a, const DevBox *b ) {
    DevBox d = { max( a->x0, b->x0 ), max( a->y0, b->y0 ), min( a->x1, b->x1 ), min( a->y1, b->y1 ) };
    if( dev_empty( &d ) ) d = dev_empty_box();
    return d;
}
static DevBox dev_union( const DevBox *a, const DevBox *b ) {
    if( dev_empty( a ) ) return *b;
    if( dev_empty( b ) ) return *a;
    DevBox d = { min( a->x0, b->x0 ), min( a->y0, b->y0 ), max( a->x1, b->x1 ), max( a->y1, b->y1 ) };
    return d;
}
static int box_width( const box2i *b ) { return b->max.x < b->min.x ? 0 : b->max.x - b->min.x + 1; }
static int same_window( const box2i *a, const box2i *b ) {
    return a->min.x == b->min.x && a->min.y == b->min.y && a->max.x == b->max.x && a->max.y == b->max.y;
}

/* ------------------------------------------------------------- job builder */

static void f16_bounds( const void *data, const box2i *full, const void **lo, const void **hi );

typedef struct BuildOptions {
    int pull_short_circuit;     /* crossfade: mix 0 / 1 fetch one source only (video_mix.c:50-55) */
    float single_alpha;         /* alpha scale when the bottom is one source (copy_frame_alpha) */
} BuildOptions;

static int fill_source( DevSrc *d, const cv_source *s, const DevBox *full ) {
    memset( d, 0, sizeof(*d) );
    d->kind = s->kind;
    d->cw = dev_empty_box();

    switch( s->kind ) {
        case CV_SRC_NONE:
            return CV_OK;

        case CV_SRC_F16: {
            DevBox cw = to_dev( &s->f16.current_window );
            d->cw = dev_intersect( &cw, full );
            d->data = s->f16.data;
            d->pitch = box_width( &s->f16.full_window );
            d->ox = s->f16.full_window.min.x;
            d->oy = s->f16.full_window.min.y;
            d->origin = (const char *) d->data - ((int64_t) d->oy * d->pitch + d->ox) * (int64_t) sizeof(rgba_f16);
            f16_bounds( d->data, &s->f16.full_window, &d->lo, &d->hi );
            if( !dev_empty( &d->cw ) && (!d->data || ((uintptr_t) d->data & 7)) )
                return cvi_fail( CV_ERR_ARGUMENT, "f16 source: data is NULL or not 8-byte aligned" );
            return CV_OK;
        }

        case CV_SRC_F32: {
            DevBox cw = to_dev( &s->f32.current_window );
            d->cw = dev_intersect( &cw, full );
            d->data = s->f32.data;
            d->pitch = box_width( &s->f32.full_window );
            d->ox = s->f32.full_window.min.x;
            d->oy = s->f32.full_window.min.y;
            if( !dev_empty( &d->cw ) && (!d->data || ((uintptr_t) d->data & 15)) )
                return cvi_fail( CV_ERR_ARGUMENT, "f32 source: data is NULL or not 16-byte aligned" );
            return CV_OK;
        }

        case CV_SRC_CODED: {
            const cv_coded *c = &s->coded;
            if( c->sub_x != 1 && c->sub_x != 2 && c->sub_x != 4 )
                return cvi_fail( CV_ERR_UNSUPPORTED, "coded source: sub_x %d (1, 2 or 4 supported)", c->sub_x );
            if( c->matrix != CV_MATRIX_709 && c->matrix != CV_MATRIX_601 )
                return cvi_fail( CV_ERR_UNSUPPORTED, "coded source: unknown matrix %d", c->matrix );
            if( c->width <= 0 || c->height <= 0 || !c->plane[0] )
                return cvi_fail( CV_ERR_ARGUMENT, "coded source: empty picture" );

            if( c->layout == CV_LAYOUT_PLANAR ) {
                if( !c->plane[1] || !c->plane[2] )
                    return cvi_fail( CV_ERR_ARGUMENT, "coded source: planar layout needs three planes" );
            }
            else if( c->layout == CV_LAYOUT_UYVY || c->layout == CV_LAYOUT_YUY2 ) {
                if( c->sub_x != 2 )
                    return cvi_fail( CV_ERR_UNSUPPORTED, "coded source: packed layouts are 4:2:2 (sub_x 2)" );
                if( ((uintptr_t) c->plane[0] & 3) || (c->stride[0] & 3) )
                    return cvi_fail( CV_ERR_ARGUMENT, "coded source: packed rows must be 4-byte aligned" );
            }
            else
                return cvi_fail( CV_ERR_UNSUPPORTED, "coded source: unknown layout %d", c->layout );

            /* video_reconstruct.c:69-76 */
            DevBox pic = { c->pic_x, c->pic_y, c->pic_x + c->width - 1, c->pic_y + c->height - 1 };
            d->cw = dev_intersect( &pic, full );
            for( int i = 0; i < 3; i++ ) { d->plane[i] = c->plane[i]; d->stride[i] = c->stride[i]; }
            d->layout = c->layout; d->width = c->width; d->height = c->height;
            d->sub_shift = c->sub_x == 4 ? 2 : (c->sub_x == 2 ? 1 : 0);
            d->matrix = c->matrix; d->pic_x = c->pic_x; d->pic_y = c->pic_y;
            d->last_chroma = (c->width - 1) / c->sub_x;
            return CV_OK;
        }
    }
    return cvi_fail( CV_ERR_ARGUMENT, "unknown source kind %d", s->kind );
}

static int f16_pair_ok( const void *data, int pitch, int ox, int full_x ) {
    return ((uintptr_t) data & 15) == 0 && (pitch & 1) == 0 && (((int64_t) ox - full_x) & 1) == 0;
}

/* the 16-byte aligned part of a frame's allocation: bulk copies of whole pixel pairs must stay inside it */
static void f16_bounds( const void *data, const box2i *full, const void **lo, const void **hi ) {
    const int w = box_width( full ), h = full->max.y < full->min.y ? 0 : full->max.y - full->min.y + 1;
    const uintptr_t begin = (uintptr_t) data, end = begin + (uintptr_t) w * (uintptr_t) h * sizeof(rgba_f16);
    *lo = (const void *)( (begin + 15) & ~(uintptr_t) 15 );
    *hi = (const void *)( end & ~(uintptr_t) 15 );
}

/*
 * Fills `dj` from `job`; returns the kernel shape in *shape and the pixels per item in *px.
 * dj->total_items == 0 means "nothing to launch" (empty result).  Writes the result window into
 * job->current_window and the output frame structs.
 */
static int build_job( cv_frame_job *job, const BuildOptions *opt, DevJob *dj, int *shape, int *px ) {
    memset( dj, 0, sizeof(*dj) );

    box2i full;
    if( job->out ) full = job->out->full_window;
    else if( job->out_f32 ) full = job->out_f32->full_window;
    else full = job->full_window;

    if( job->out && job->out_f32 && !same_window( &job->out->full_window, &job->out_f32->full_window ) )
        return cvi_fail( CV_ERR_ARGUMENT, "frame job: out and out_f32 must share full_window" );
    if( job->out_u8 && !job->u8_table )
        return cvi_fail( CV_ERR_ARGUMENT, "frame job: out_u8 needs u8_table" );
    if( job->layer_count < 0 || job->layer_count > CV_MAX_LAYERS )
        return cvi_fail( CV_ERR_ARGUMENT, "frame job: layer_count %d (0..%d)", job->layer_count, CV_MAX_LAYERS );

    box2i_set_empty( &job->current_window );
    if( job->out ) box2i_set_empty( &job->out->current_window );
    if( job->out_f32 ) box2i_set_empty( &job->out_f32->current_window );

    const DevBox dfull = to_dev( &full );
    if( dev_empty( &dfull ) )
        return CV_OK;

    dj->full_x = full.min.x;
    dj->full_y = full.min.y;
    dj->out_pitch = box_width( &full );
    dj->out16 = job->out ? job->out->data : NULL;
    dj->out32 = job->out_f32 ? job->out_f32->data : NULL;
    dj->out8 = job->out_u8;
    dj->u8tab = job->u8_table;
    dj->single_alpha = 1.0f;
    dj->one = 1.0f;

    if( (job->out && !dj->out16) || (job->out_f32 && !dj->out32) )
        return cvi_fail( CV_ERR_ARGUMENT, "frame job: output frame without data" );
    if( ((uintptr_t) dj->out16 & 7) || ((uintptr_t) dj->out32 & 15) || ((uintptr_t) dj->out8 & 3) )
        return cvi_fail( CV_ERR_ARGUMENT, "frame job: misaligned output pointer" );

    /* ---- bottom ---- */
    int rc;
    DevSrc sa, sb;
    if( (rc = fill_source( &sa, &job->a, &dfull )) ) return rc;
    if( (rc = fill_source( &sb, &job->b, &dfull )) ) return rc;

    DevBox under = dev_empty_box();
    dj->bottom_mode = DEV_BOTTOM_NONE;

    if( sa.kind != CV_SRC_NONE && sb.kind != CV_SRC_NONE ) {
        const float mix_b = clampf( job->mix_b, 0.0f, 1.0f );
        const float mix_a = 1.0f - mix_b;

        if( opt->pull_short_circuit && mix_b == 0.0f ) {
            dj->a = sa; dj->bottom_mode = DEV_BOTTOM_SINGLE;
        }
        else if( opt->pull_short_circuit && mix_b == 1.0f ) {
            dj->a = sb; dj->bottom_mode = DEV_BOTTOM_SINGLE;
        }
        else if( dev_empty( &sa.cw ) ) {         /* video_mix.c:111-114 -> copy_frame_alpha(b, mix_b) */
            dj->a = sb; dj->bottom_mode = DEV_BOTTOM_SINGLE; dj->single_alpha = mix_b;
        }
        else if( dev_empty( &sb.cw ) ) {         /* video_mix.c:115-118 */
            dj->a = sa; dj->bottom_mode = DEV_BOTTOM_SINGLE; dj->single_alpha = mix_a;
        }
        else {
            dj->a = sa; dj->b = sb; dj->bottom_mode = DEV_BOTTOM_CROSS;
            dj->mix_a = mix_a; dj->mix_b = mix_b;
            under = dev_union( &sa.cw, &sb.cw );
        }
    }
    else if( sa.kind != CV_SRC_NONE || sb.kind != CV_SRC_NONE ) {
        dj->a = (sa.kind != CV_SRC_NONE) ? sa : sb;
        dj->bottom_mode = DEV_BOTTOM_SINGLE;
        dj->single_alpha = clampf( opt->single_alpha, 0.0f, 1.0f );
    }

    if( dj->bottom_mode == DEV_BOTTOM_SINGLE ) {
        /* video_copy_frame_alpha_f32: alpha 0 or an empty window leaves nothing (video_mix.c:80-93) */
        if( dj->single_alpha == 0.0f || dev_empty( &dj->a.cw ) )
            dj->bottom_mode = DEV_BOTTOM_NONE;
        else
            under = dj->a.cw;
    }

    int coded = 0;
    if( dj->bottom_mode != DEV_BOTTOM_NONE )
        coded = dj->a.kind == CV_SRC_CODED || (dj->bottom_mode == DEV_BOTTOM_CROSS && dj->b.kind == CV_SRC_CODED);

    /* ---- layers (workspace.c:530-547 / video_mix.c:243-249) ---- */
    int pair_ok = dj->out16 ? f16_pair_ok( dj->out16, dj->out_pitch, dj->full_x, dj->full_x ) : ((dj->out_pitch & 1) == 0);
    int n = 0;

    for( int i = 0; i < job->layer_count; i++ ) {
        const cv_layer *L = &job->layers[i];
        const float mix = clampf( L->mix, 0.0f, 1.0f );
        DevBox cw = to_dev( &L->frame.current_window );
        cw = dev_intersect( &cw, &dfull );

        if( dev_empty( &cw ) || mix == 0.0f )
            continue;
        if( !L->frame.data || ((uintptr_t) L->frame.data & 7) )
            return cvi_fail( CV_ERR_ARGUMENT, "layer %d: data is NULL or not 8-byte aligned", i );

        DevLayer *d = &dj->layer[n++];
        d->data = L->frame.data;
        d->pitch = box_width( &L->frame.full_window );
        d->ox = L->frame.full_window.min.x;
        d->oy = L->frame.full_window.min.y;
        d->origin = (const char *) d->data - ((int64_t) d->oy * d->pitch + d->ox) * (int64_t) sizeof(rgba_f16);
        f16_bounds( d->data, &L->frame.full_window, &d->lo, &d->hi );
        d->mix = mix;
        d->cw = cw;
        d->under = under;
        under = dev_union( &under, &cw );
        pair_ok = pair_ok && f16_pair_ok( d->data, d->pitch, d->ox, dj->full_x );

        /* fused per-pixel filters of this layer */
        if( L->ops && L->ops->flags ) {
            const cv_layer_ops *ops = L->ops;
            if( ops->flags & ~(CV_OP_WEAVE | CV_OP_COLOR | CV_OP_GAIN_OFFSET | CV_OP_CROP) )
                return cvi_fail( CV_ERR_ARGUMENT, "layer %d: unknown filter flags %#x", i, ops->flags );
            d->op_flags = ops->flags;
            if( ops->flags & CV_OP_COLOR ) {
                if( ops->color != CV_COLOR_RGB_TO_XYZ_SDTV && ops->color != CV_COLOR_XYZ_TO_SRGB )
                    return cvi_fail( CV_ERR_ARGUMENT, "layer %d: unknown colour transform %d", i, ops->color );
                d->op_color = ops->color;
            }
            d->gain = ops->gain; d->offset = ops->offset;
            d->crop = to_dev( &ops->crop );
            if( ops->flags & CV_OP_WEAVE ) {
                const rgba_frame_f16 *e = &ops->even_rows;
                DevBox ecw = to_dev( &e->current_window );
                if( !e->data || ((uintptr_t) e->data & 7) )
                    return cvi_fail( CV_ERR_ARGUMENT, "layer %d: even_rows frame without data", i );
                if( ecw.x0 > cw.x0 || ecw.y0 > cw.y0 || ecw.x1 < cw.x1 || ecw.y1 < cw.y1 )
                    return cvi_fail( CV_ERR_UNSUPPORTED, "layer %d: even_rows must cover the layer's window (weave the frames with cv_weave_fields_f16 instead)", i );
                d->weave_pitch = box_width( &e->full_window );
                d->weave_origin = (const char *) e->data - ((int64_t) e->full_window.min.y * d->weave_pitch + e->full_window.min.x) * (int64_t) sizeof(rgba_f16);
                pair_ok = pair_ok && f16_pair_ok( e->data, d->weave_pitch, e->full_window.min.x, dj->full_x );
            }
            dj->flags |= JOB_FLAG_OPS;
        }
    }
    if( dj->flags & JOB_FLAG_OPS ) {
        CvDevice *dev = cvi_device();
        if( !dev ) return CV_ERR_NOT_INITIALISED;
        dj->op_lut_709 = (const uint16_t *) dev->lut[CV_LUT_709_TO_LINEAR_SCENE];
        dj->op_lut_srgb = (const uint16_t *) dev->lut[CV_LUT_LINEAR_TO_SRGB];
    }
    dj->n_layers = n;

    /* ---- result window and work decomposition ---- */
    dj->region = under;
    if( dev_empty( &under ) )
        return CV_OK;

    job->current_window = from_dev( &under );
    if( job->out ) job->out->current_window = job->current_window;
    if( job->out_f32 ) job->out_f32->current_window = job->current_window;

    *px = pair_ok ? 2 : 1;
    *shape = coded ? KB_CODED : (dj->bottom_mode != DEV_BOTTOM_NONE ? KB_GENERAL : KB_LAYERS);

    /* f32 frames keep the IEEE-division path (their results are visible at full f32 precision) */
    if( dj->out32 || (dj->bottom_mode != DEV_BOTTOM_NONE && (dj->a.kind == CV_SRC_F32 ||
            (dj->bottom_mode == DEV_BOTTOM_CROSS && dj->b.kind == CV_SRC_F32))) )
        dj->flags |= JOB_FLAG_EXACT;

    /* The shared-reciprocal division of the fast path assumes that a non-zero alpha is >= 2^-48 (a half's smallest
     * alpha 2^-24 squared).  A positive mix below 2^-24 breaks that bound (alpha = a * mix may fall under the
     * reciprocal's floor): such jobs keep the IEEE-division path. */
    {
        const float floor_mix = 5.9604645e-8f;      /* 2^-24 */
        int tiny = 0;
        if( dj->bottom_mode == DEV_BOTTOM_CROSS )
            tiny = (dj->mix_a > 0.0f && dj->mix_a < floor_mix) || (dj->mix_b > 0.0f && dj->mix_b < floor_mix);
        else if( dj->bottom_mode == DEV_BOTTOM_SINGLE )
            tiny = dj->single_alpha > 0.0f && dj->single_alpha < floor_mix;
        for( int i = 0; i < n; i++ )
            tiny = tiny || dj->layer[i].mix < floor_mix;         /* layers with mix 0 were dropped above */
        if( tiny )
            dj->flags |= JOB_FLAG_EXACT;
    }

    /* the common workspace case: full-coverage layers, no window tests needed in the kernel */
    if( *shape == KB_LAYERS && *px == 2 && !(dj->flags & JOB_FLAG_OPS) && ((under.x0 - dj->full_x) & 1) == 0 && ((under.x1 - under.x0 + 1) & 1) == 0 ) {
        int simple = 1;
        for( int i = 0; i < n; i++ ) {
            const DevBox *c = &dj->layer[i].cw;
            if( c->x0 != under.x0 || c->y0 != under.y0 || c->x1 != under.x1 || c->y1 != under.y1 || dj->layer[i].mix != 1.0f )
                simple = 0;
        }
        if( simple )
            dj->flags |= JOB_FLAG_SIMPLE;
    }

    /* the row-segment kernel also moves f16 bottom sources as aligned pixel pairs */
    if( *px == 2 && !(dj->flags & JOB_FLAG_EXACT) ) {
        int seg = 1;
        if( dj->bottom_mode != DEV_BOTTOM_NONE && dj->a.kind == CV_SRC_F16 )
            seg = seg && f16_pair_ok( dj->a.data, dj->a.pitch, dj->a.ox, dj->full_x );
        if( dj->bottom_mode == DEV_BOTTOM_CROSS && dj->b.kind == CV_SRC_F16 )
            seg = seg && f16_pair_ok( dj->b.data, dj->b.pitch, dj->b.ox, dj->full_x );
        if( dj->out8 && !dj->out16 && (dj->out_pitch & 1) )
            seg = 0;
        if( seg )
            dj->flags |= JOB_FLAG_SEG;
    }

    if( *px == 2 ) {
        dj->item_x0 = dj->full_x + ((under.x0 - dj->full_x) & ~1);
        dj->items_per_row = (uint32_t)( (under.x1 - dj->item_x0) / 2 + 1 );
    }
    else {
        dj->item_x0 = under.x0;
        dj->items_per_row = (uint32_t)( under.x1 - under.x0 + 1 );
    }
    dj->rows = (uint32_t)( under.y1 - under.y0 + 1 );

    const uint64_t total = (uint64_t) dj->items_per_row * dj->rows;
    if( total > 0xFFFFFFFFu )
        return cvi_fail( CV_ERR_UNSUPPORTED, "frame job: window too large" );
    dj->total_items = (uint32_t) total;
    dj->div_magic = cvi_div_magic( dj->items_per_row );
    return CV_OK;
}

/* ------------------------------------------------------------ frame programs */

static int render_jobs( cv_frame_job *jobs, int count, const BuildOptions *opt, cv_stream stream ) {
    if( !cvi_device() ) {
        for( int i = 0; i < count; i++ ) {
            box2i_set_empty( &jobs[i].current_window );
            if( jobs[i].out ) box2i_set_empty( &jobs[i].out->current_window );
            if( jobs[i].out_f32 ) box2i_set_empty( &jobs[i].out_f32->current_window );
        }
        return CV_ERR_NOT_INITIALISED;
    }
    if( count <= 0 ) return CV_OK;
    if( !jobs ) return cvi_fail( CV_ERR_ARGUMENT, "cv_render_frames: NULL jobs" );

    cudaStream_t s = cvi_stream( stream );

    if( count == 1 ) {
        DevJob dj;
        int shape = 0, px = 1;
        int rc = build_job( &jobs[0], opt, &dj, &shape, &px );
        if( rc || dj.total_items == 0 ) return rc;
        return cvi_launch_jobs( shape, px, &dj, 1, s );
    }

    /* Built and launched in chunks of RENDER_CHUNK jobs: the device starts on the first chunk while the host builds the
     * next, and the two scratch arrays stay small enough (3 MB) for the allocator to reuse them without fresh pages -- a
     * 12 000-job call used to build 36 MB of descriptors before its first launch, which made small frames host-bound.
     * Inside a chunk jobs are grouped by (shape, px): they are independent, so their order inside the stream is free. */
    enum { RENDER_CHUNK = 1024 };            /* a multiple of the job ring's segment (256 jobs = one launch) */
    const int chunk = count < RENDER_CHUNK ? count : RENDER_CHUNK;
    DevJob *built = (DevJob *) malloc( sizeof(DevJob) * (size_t) chunk );
    DevJob *group = NULL;
    int *key = (int *) malloc( sizeof(int) * (size_t) chunk );
    if( !built || !key ) {
        free( built ); free( key );
        return cvi_fail( CV_ERR_MEMORY, "cv_render_frames: out of host memory" );
    }

    int rc = CV_OK;
    for( int base = 0; base < count && !rc; base += chunk ) {
        const int m = count - base < chunk ? count - base : chunk;
        int first_key = -2, mixed = 0;
        for( int i = 0; i < m && !rc; i++ ) {
            int shape = 0, px = 1;
            rc = build_job( &jobs[base + i], opt, &built[i], &shape, &px );
            key[i] = built[i].total_items ? shape * 2 + (px == 2) : -1;
            if( first_key == -2 ) first_key = key[i];
            else if( key[i] != first_key ) mixed = 1;
        }
        if( rc ) break;

        if( !mixed ) {                          /* the usual case: one kind of job, launched from where it was built */
            if( first_key >= 0 )
                rc = cvi_launch_jobs( first_key / 2, (first_key & 1) ? 2 : 1, built, m, s );
            continue;
        }
        if( !group && !( group = (DevJob *) malloc( sizeof(DevJob) * (size_t) chunk ) ) ) {
            rc = cvi_fail( CV_ERR_MEMORY, "cv_render_frames: out of host memory" );
            break;
        }
        for( int k = 0; k < KB_COUNT * 2 && !rc; k++ ) {
            int n = 0;
            for( int i = 0; i < m; i++ )
                if( key[i] == k ) group[n++] = built[i];
            if( n )
                rc = cvi_launch_jobs( k / 2, (k & 1) ? 2 : 1, group, n, s );
        }
    }

    free( built ); free( group ); free( key );
    return rc;
}

int cv_render_frames( cv_frame_job *jobs, int count, cv_stream stream ) {
    const BuildOptions opt = { 1, 1.0f };
    return render_jobs( jobs, count, &opt, stream );
}

int cv_convert_ycbcr8( rgba_frame_f16 *frame, const cv_coded *image, cv_stream stream ) {
    if( !frame || !image ) return cvi_fail( CV_ERR_ARGUMENT, "cv_convert_ycbcr8: NULL argument" );
    cv_frame_job job;
    memset( &job, 0, sizeof(job) );
    job.a.kind = CV_SRC_CODED;
    job.a.coded = *image;
    job.out = frame;
    return cv_render_frames( &job, 1, stream );
}

int cv_mix_cross_f16( rgba_frame_f16 *out, const rgba_frame_f16 *a, const rgba_frame_f16 *b, float mix_b, cv_stream stream ) {
    if( !out || !a || !b ) return cvi_fail( CV_ERR_ARGUMENT, "cv_mix_cross_f16: NULL argument" );
    cv_frame_job job;
    memset( &job, 0, sizeof(job) );
    job.a.kind = CV_SRC_F16; job.a.f16 = *a;
    job.b.kind = CV_SRC_F16; job.b.f16 = *b;
    job.mix_b = mix_b;
    job.out = out;
    return cv_render_frames( &job, 1, stream );
}

int cv_composite_over_f16( rgba_frame_f16 *out, const cv_layer *layers, int count, cv_stream stream ) {
    if( !out || (count > 0 && !layers) ) return cvi_fail( CV_ERR_ARGUMENT, "cv_composite_over_f16: NULL argument" );
    if( count < 0 ) return cvi_fail( CV_ERR_ARGUMENT, "cv_composite_over_f16: negative count" );

    cv_frame_job job;
    memset( &job, 0, sizeof(job) );

    if( count <= CV_MAX_LAYERS ) {
        job.layer_count = count;
        if( count ) memcpy( job.layers, layers, sizeof(cv_layer) * (size_t) count );
        job.out = out;
        return cv_render_frames( &job, 1, stream );
    }

    /* deep stacks: CV_MAX_LAYERS at a time through a pooled f32 accumulator (no extra rounding) */
    rgba_frame_f32 acc;
    int rc = cv_frame_alloc_f32( &acc, &out->full_window, CV_MEM_DEVICE );
    if( rc ) { box2i_set_empty( &out->current_window ); return rc; }

    int done = 0;
    while( done < count && !rc ) {
        const int n = (count - done > CV_MAX_LAYERS) ? CV_MAX_LAYERS : count - done;
        const int last = (done + n == count);
        memset( &job, 0, sizeof(job) );
        if( done ) { job.a.kind = CV_SRC_F32; job.a.f32 = acc; }
        job.layer_count = n;
        memcpy( job.layers, layers + done, sizeof(cv_layer) * (size_t) n );
        if( last ) job.out = out; else job.out_f32 = &acc;
        rc = cv_render_frames( &job, 1, stream );
        done += n;
    }

    /* the accumulator goes back to the pool; later users are ordered behind us on `stream` only,
     * so wait for the chain before another stream may pick the block up */
    if( !rc ) rc = cv_stream_sync( stream );
    cv_frame_free_f32( &acc, CV_MEM_DEVICE );
    if( rc ) box2i_set_empty( &out->current_window );
    return rc;
}

int cv_copy_frame_f16( rgba_frame_f16 *out, const rgba_frame_f16 *in, cv_stream stream ) {
    if( !out || !in ) return cvi_fail( CV_ERR_ARGUMENT, "cv_copy_frame_f16: NULL argument" );
    if( !cvi_device() ) { box2i_set_empty( &out->current_window ); return CV_ERR_NOT_INITIALISED; }

    /* video_mix.c:27-44: a strided row copy, done by the copy engine (bit-preserving) */
    box2i inner;
    box2i_intersect( &inner, &out->full_window, &in->current_window );
    out->current_window = inner;
    if( box2i_is_empty( &inner ) ) {
        box2i_set_empty( &out->current_window );
        return CV_OK;
    }

    const size_t out_pitch = (size_t) box_width( &out->full_window ) * sizeof(rgba_f16);
    const size_t in_pitch = (size_t) box_width( &in->full_window ) * sizeof(rgba_f16);
    const rgba_f16 *src = in->data + (size_t)(inner.min.y - in->full_window.min.y) * box_width( &in->full_window ) + (inner.min.x - in->full_window.min.x);
    rgba_f16 *dst = out->data + (size_t)(inner.min.y - out->full_window.min.y) * box_width( &out->full_window ) + (inner.min.x - out->full_window.min.x);

    return cvi_cuda( cudaMemcpy2DAsync( dst, out_pitch, src, in_pitch, (size_t) box_width( &inner ) * sizeof(rgba_f16),
        (size_t)(inner.max.y - inner.min.y + 1), cudaMemcpyDeviceToDevice, cvi_stream( stream ) ), "cudaMemcpy2DAsync(copy_frame)" );
}

/* ------------------------------------------------------------- f32 frame ops */

int cv_copy_frame_alpha_f32( rgba_frame_f32 *out, const rgba_frame_f32 *in, float alpha, cv_stream stream ) {
    if( !out || !in ) return cvi_fail( CV_ERR_ARGUMENT, "cv_copy_frame_alpha_f32: NULL argument" );
    alpha = clampf( alpha, 0.0f, 1.0f );

    if( out->data == in->data && alpha == 1.0f )      /* video_mix.c:76-78 */
        return cvi_device() ? CV_OK : CV_ERR_NOT_INITIALISED;

    cv_frame_job job;
    memset( &job, 0, sizeof(job) );
    job.a.kind = CV_SRC_F32; job.a.f32 = *in;
    job.out_f32 = out;
    const BuildOptions opt = { 0, alpha };
    return render_jobs( &job, 1, &opt, stream );
}

int cv_mix_cross_f32( rgba_frame_f32 *out, const rgba_frame_f32 *a, const rgba_frame_f32 *b, float mix_b, cv_stream stream ) {
    if( !out || !a || !b ) return cvi_fail( CV_ERR_ARGUMENT, "cv_mix_cross_f32: NULL argument" );
    cv_frame_job job;
    memset( &job, 0, sizeof(job) );
    job.a.kind = CV_SRC_F32; job.a.f32 = *a;
    job.b.kind = CV_SRC_F32; job.b.f32 = *b;
    job.mix_b = mix_b;
    job.out_f32 = out;
    const BuildOptions opt = { 0, 1.0f };        /* video_mix_cross_f32 itself has no 0/1 short-circuit */
    return render_jobs( &job, 1, &opt, stream );
}

int cv_mix_over_f32( rgba_frame_f32 *out, const rgba_frame_f32 *b, float mix_b, cv_stream stream ) {
    if( !out || !b ) return cvi_fail( CV_ERR_ARGUMENT, "cv_mix_over_f32: NULL argument" );
    if( !cvi_device() ) return CV_ERR_NOT_INITIALISED;
    mix_b = clampf( mix_b, 0.0f, 1.0f );

    if( box2i_is_empty( &out->current_window ) )          /* video_mix.c:243-246 */
        return cv_copy_frame_alpha_f32( out, b, mix_b, stream );
    if( box2i_is_empty( &b->current_window ) || mix_b == 0.0f )
        return CV_OK;

    const DevBox full = to_dev( &out->full_window );
    DevBox ocw = to_dev( &out->current_window ), bcw = to_dev( &b->current_window );
    ocw = dev_intersect( &ocw, &full );
    bcw = dev_intersect( &bcw, &full );
    const DevBox region = dev_union( &ocw, &bcw );
    if( dev_empty( &region ) ) return CV_OK;

    OverF32Args args;
    memset( &args, 0, sizeof(args) );
    args.out = (float4 *) out->data; args.b = (const float4 *) b->data;
    args.out_pitch = box_width( &out->full_window ); args.out_ox = out->full_window.min.x; args.out_oy = out->full_window.min.y;
    args.b_pitch = box_width( &b->full_window ); args.b_ox = b->full_window.min.x; args.b_oy = b->full_window.min.y;
    args.out_cw = ocw; args.b_cw = bcw; args.region = region; args.mix = mix_b;
    args.width = (uint32_t)( region.x1 - region.x0 + 1 );
    args.total = args.width * (uint32_t)( region.y1 - region.y0 + 1 );
    args.magic = cvi_div_magic( args.width );

    out->current_window = from_dev( &region );
    return cvi_launch_over_f32( &args, cvi_stream( stream ) );
}

int cv_frame_f16_to_f32( rgba_frame_f32 *out, const rgba_frame_f16 *in, cv_stream stream ) {
    if( !out || !in ) return cvi_fail( CV_ERR_ARGUMENT, "cv_frame_f16_to_f32: NULL argument" );
    cv_frame_job job;
    memset( &job, 0, sizeof(job) );
    job.a.kind = CV_SRC_F16; job.a.f16 = *in;
    job.out_f32 = out;
    return cv_render_frames( &job, 1, stream );
}

int cv_frame_f32_to_f16( rgba_frame_f16 *out, const rgba_frame_f32 *in, cv_stream stream ) {
    if( !out || !in ) return cvi_fail( CV_ERR_ARGUMENT, "cv_frame_f32_to_f16: NULL argument" );
    cv_frame_job job;
    memset( &job, 0, sizeof(job) );
    job.a.kind = CV_SRC_F32; job.a.f32 = *in;
    job.out = out;
    return cv_render_frames( &job, 1, stream );
}

/* ------------------------------------------------------------- solid colour */

/* half.c:47-51 as bit operations (truncation, subnormals kept): four values on the host */
static uint16_t f2h_bits( float value ) {
    union { float f; uint32_t u; } v = { .f = value };
    const uint32_t sign = (v.u >> 16) & 0x8000u, mant = v.u & 0x007FFFFFu;
    const int e = (int)((v.u >> 23) & 0xFF) - 127;
    if( e < -24 ) return (uint16_t) sign;
    if( e < -14 ) return (uint16_t)(sign | ((0x0400u >> (-14 - e)) + (mant >> (-e - 1))));
    if( e <= 15 ) return (uint16_t)(sign | (((uint32_t)(e + 15) << 10) + (mant >> 13)));
    if( e < 128 ) return (uint16_t)(sign | 0x7C00u);
    return (uint16_t)(sign | (0x7C00u + (mant >> 13)));
}

static int solid_fill( void *data, const box2i *full, box2i *current, const box2i *window, int bytes, const uint32_t word[4], cv_stream stream ) {
    if( !cvi_device() ) { box2i_set_empty( current ); return CV_ERR_NOT_INITIALISED; }
    box2i_intersect( current, window, full );
    if( box2i_is_empty( current ) )
        return CV_OK;
    if( !data ) { box2i_set_empty( current ); return cvi_fail( CV_ERR_ARGUMENT, "cv_solid_frame: frame without data" ); }

    FillArgs A;
    memset( &A, 0, sizeof(A) );
    A.data = data; A.pitch = box_width( full ); A.ox = full->min.x; A.oy = full->min.y; A.bytes = bytes;
    A.win = to_dev( current );
    A.width = (uint32_t) box_width( current );
    A.total = A.width * (uint32_t)( current->max.y - current->min.y + 1 );
    A.magic = cvi_div_magic( A.width );
    memcpy( A.word, word, sizeof(A.word) );
    int rc = cvi_launch_fill( &A, cvi_stream( stream ) );
    if( rc ) box2i_set_empty( current );
    return rc;
}

int cv_solid_frame_f16( rgba_frame_f16 *frame, const box2i *window, const float color[4], cv_stream stream ) {
    if( !frame || !window || !color ) return cvi_fail( CV_ERR_ARGUMENT, "cv_solid_frame_f16: NULL argument" );
    /* SolidColorVideoSource.c:67-68: the f32 colour is truncated to half once, then replicated */
    const uint32_t word[4] = { (uint32_t) f2h_bits( color[0] ) | ((uint32_t) f2h_bits( color[1] ) << 16),
                               (uint32_t) f2h_bits( color[2] ) | ((uint32_t) f2h_bits( color[3] ) << 16), 0, 0 };
    return solid_fill( frame->data, &frame->full_window, &frame->current_window, window, 8, word, stream );
}

int cv_solid_frame_f32( rgba_frame_f32 *frame, const box2i *window, const float color[4], cv_stream stream ) {
    if( !frame || !window || !color ) return cvi_fail( CV_ERR_ARGUMENT, "cv_solid_frame_f32: NULL argument" );
    uint32_t word[4];
    memcpy( word, color, sizeof(word) );
    return solid_fill( frame->data, &frame->full_window, &frame->current_window, window, 16, word, stream );
}

/* ------------------------------------------------- gain / offset, field weave */

int cv_gain_offset_f32( rgba_frame_f32 *frame, float gain, float offset, cv_stream stream ) {
    if( !frame ) return cvi_fail( CV_ERR_ARGUMENT, "cv_gain_offset_f32: NULL frame" );
    if( !cvi_device() ) return CV_ERR_NOT_INITIALISED;
    box2i cw;
    box2i_intersect( &cw, &frame->current_window, &frame->full_window );
    if( box2i_is_empty( &cw ) ) return CV_OK;
    if( !frame->data ) return cvi_fail( CV_ERR_ARGUMENT, "cv_gain_offset_f32: frame without data" );

    FillArgs A;
    memset( &A, 0, sizeof(A) );
    A.data = frame->data; A.pitch = box_width( &frame->full_window ); A.ox = frame->full_window.min.x; A.oy = frame->full_window.min.y;
    A.bytes = 16; A.win = to_dev( &cw );
    A.width = (uint32_t) box_width( &cw );
    A.total = A.width * (uint32_t)( cw.max.y - cw.min.y + 1 );
    A.magic = cvi_div_magic( A.width );
    memcpy( &A.word[0], &gain, 4 );
    memcpy( &A.word[1], &offset, 4 );
    return cvi_launch_gain_offset( &A, cvi_stream( stream ) );
}

int cv_weave_fields_f16( rgba_frame_f16 *frame, const rgba_frame_f16 *even_rows, cv_stream stream ) {
    /* Pulldown23RemovalFilter.c:83-101: the even rows of frame's current window are replaced by the same rows
     * of `even_rows` (the copy engine does the strided row copy; pixels are untouched bits) */
    if( !frame || !even_rows ) return cvi_fail( CV_ERR_ARGUMENT, "cv_weave_fields_f16: NULL argument" );
    if( !cvi_device() ) return CV_ERR_NOT_INITIALISED;
    box2i cw;
    box2i_intersect( &cw, &frame->current_window, &even_rows->current_window );
    box2i_intersect( &cw, &cw, &frame->full_window );
    if( box2i_is_empty( &cw ) ) return CV_OK;

    const int first = (cw.min.y + 1) & ~1;
    if( first > cw.max.y ) return CV_OK;
    const size_t rows = (size_t)( (cw.max.y - first) / 2 + 1 );
    const size_t dpitch = (size_t) box_width( &frame->full_window ) * sizeof(rgba_f16);
    const size_t spitch = (size_t) box_width( &even_rows->full_window ) * sizeof(rgba_f16);
    const char *src = (const char *) even_rows->data + (size_t)( first - even_rows->full_window.min.y ) * spitch
                    + (size_t)( cw.min.x - even_rows->full_window.min.x ) * sizeof(rgba_f16);
    char *dst = (char *) frame->data + (size_t)( first - frame->full_window.min.y ) * dpitch
              + (size_t)( cw.min.x - frame->full_window.min.x ) * sizeof(rgba_f16);
    return cvi_cuda( cudaMemcpy2DAsync( dst, 2 * dpitch, src, 2 * spitch, (size_t) box_width( &cw ) * sizeof(rgba_f16), rows,
                                        cudaMemcpyDeviceToDevice, cvi_stream( stream ) ), "cudaMemcpy2DAsync(weave)" );
}

/* ------------------------------------------------- colour transforms, crop */

static int color_transform( rgba_frame_f16 *frame, const float m[9], int pre, int post, cv_stream stream, const char *who ) {
    if( !frame ) return cvi_fail( CV_ERR_ARGUMENT, "%s: NULL frame", who );
    CvDevice *dev = cvi_device();
    if( !dev ) return CV_ERR_NOT_INITIALISED;
    box2i cw;
    box2i_intersect( &cw, &frame->current_window, &frame->full_window );
    if( box2i_is_empty( &cw ) ) return CV_OK;
    if( !frame->data ) return cvi_fail( CV_ERR_ARGUMENT, "%s: frame without data", who );

    ColorArgs A;
    memset( &A, 0, sizeof(A) );
    A.frame = (uint2 *) frame->data; A.pitch = box_width( &frame->full_window );
    A.ox = frame->full_window.min.x; A.oy = frame->full_window.min.y;
    A.win = to_dev( &cw );
    A.width = (uint32_t) box_width( &cw );
    A.total = A.width * (uint32_t)( cw.max.y - cw.min.y + 1 );
    A.magic = cvi_div_magic( A.width );
    A.pre = pre >= 0 ? (const uint16_t *) dev->lut[pre] : NULL;
    A.post = post >= 0 ? (const uint16_t *) dev->lut[post] : NULL;
    memcpy( A.m, m, sizeof(A.m) );
    A.one = 1.0f;
    return cvi_launch_color( &A, cvi_stream( stream ) );
}

int cv_color_rgb_to_xyz_sdtv( rgba_frame_f16 *frame, cv_stream stream ) {
    /* color.c:111-114: columns r, g, b of the SMPTE-C RGB -> XYZ matrix, written here by output row */
    static const float m[9] = { 0.3936f, 0.3652f, 0.1916f,   0.2124f, 0.7010f, 0.0865f,   0.0187f, 0.1119f, 0.9582f };
    return color_transform( frame, m, CV_LUT_709_TO_LINEAR_SCENE, -1, stream, "cv_color_rgb_to_xyz_sdtv" );
}

int cv_color_xyz_to_srgb( rgba_frame_f16 *frame, cv_stream stream ) {
    /* color.c:139-142 */
    static const float m[9] = { 3.2410f, -1.5374f, -0.4986f,   -0.9692f, 1.8760f, 0.0416f,   0.0556f, -0.2040f, 1.0570f };
    return color_transform( frame, m, -1, CV_LUT_LINEAR_TO_SRGB, stream, "cv_color_xyz_to_srgb" );
}

int cv_crop_f16( rgba_frame_f16 *frame, const box2i *crop_rect, cv_stream stream ) {
    /* video_filter.c:79-95: pixels of the frame outside crop_rect become (0,0,0,0); the frame keeps its window.
     * Up to four zero fills: the rows above and below the rectangle, the strips left and right of it. */
    if( !frame || !crop_rect ) return cvi_fail( CV_ERR_ARGUMENT, "cv_crop_f16: NULL argument" );
    if( !cvi_device() ) return CV_ERR_NOT_INITIALISED;
    box2i cw, keep;
    box2i_intersect( &cw, &frame->current_window, &frame->full_window );
    if( box2i_is_empty( &cw ) ) return CV_OK;
    if( !frame->data ) return cvi_fail( CV_ERR_ARGUMENT, "cv_crop_f16: frame without data" );
    box2i_intersect( &keep, &cw, crop_rect );

    box2i strips[4];
    int count = 0;
    if( box2i_is_empty( &keep ) )
        strips[count++] = cw;
    else {
        if( keep.min.y > cw.min.y ) strips[count++] = (box2i) { { cw.min.x, cw.min.y }, { cw.max.x, keep.min.y - 1 } };
        if( keep.max.y < cw.max.y ) strips[count++] = (box2i) { { cw.min.x, keep.max.y + 1 }, { cw.max.x, cw.max.y } };
        if( keep.min.x > cw.min.x ) strips[count++] = (box2i) { { cw.min.x, keep.min.y }, { keep.min.x - 1, keep.max.y } };
        if( keep.max.x < cw.max.x ) strips[count++] = (box2i) { { keep.max.x + 1, keep.min.y }, { cw.max.x, keep.max.y } };
    }

    for( int i = 0; i < count; i++ ) {
        FillArgs A;
        memset( &A, 0, sizeof(A) );
        A.data = frame->data; A.pitch = box_width( &frame->full_window );
        A.ox = frame->full_window.min.x; A.oy = frame->full_window.min.y; A.bytes = 8;
        A.win = to_dev( &strips[i] );
        A.width = (uint32_t) box_width( &strips[i] );
        A.total = A.width * (uint32_t)( strips[i].max.y - strips[i].min.y + 1 );
        A.magic = cvi_div_magic( A.width );
        int rc = cvi_launch_fill( &A, cvi_stream( stream ) );
        if( rc ) return rc;
    }
    return CV_OK;
}

/* ------------------------------------------------- MPEG-2 4:2:0 subsample */

int cv_subsample_mpeg2( const rgba_frame_f16 *frame, const cv_coded *image, cv_stream stream ) {
    if( !frame || !image ) return cvi_fail( CV_ERR_ARGUMENT, "cv_subsample_mpeg2: NULL argument" );
    CvDevice *dev = cvi_device();
    if( !dev ) return CV_ERR_NOT_INITIALISED;
    if( image->layout != CV_LAYOUT_PLANAR )
        return cvi_fail( CV_ERR_UNSUPPORTED, "cv_subsample_mpeg2: 4:2:0 is planar" );
    if( image->width <= 0 || image->height <= 0 || (image->width & 1) || (image->height & 1) )
        return cvi_fail( CV_ERR_ARGUMENT, "cv_subsample_mpeg2: picture %d x %d (both must be even and positive)", image->width, image->height );
    if( image->matrix != CV_MATRIX_709 && image->matrix != CV_MATRIX_601 )
        return cvi_fail( CV_ERR_ARGUMENT, "cv_subsample_mpeg2: matrix %d", image->matrix );
    if( !image->plane[0] || !image->plane[1] || !image->plane[2] )
        return cvi_fail( CV_ERR_ARGUMENT, "cv_subsample_mpeg2: NULL plane" );
    if( image->stride[0] < image->width || image->stride[1] < image->width / 2 || image->stride[2] < image->width / 2 )
        return cvi_fail( CV_ERR_ARGUMENT, "cv_subsample_mpeg2: stride shorter than the row" );

    box2i cw;
    box2i_intersect( &cw, &frame->current_window, &frame->full_window );
    if( !box2i_is_empty( &cw ) && !frame->data )
        return cvi_fail( CV_ERR_ARGUMENT, "cv_subsample_mpeg2: frame without data" );

    Mpeg2Args A;
    memset( &A, 0, sizeof(A) );
    A.frame = (const uint2 *) frame->data; A.pitch = box_width( &frame->full_window );
    A.ox = frame->full_window.min.x; A.oy = frame->full_window.min.y;
    A.cw = to_dev( &cw );
    A.pic_x = image->pic_x; A.pic_y = image->pic_y; A.width = image->width; A.height = image->height; A.matrix = image->matrix;
    for( int i = 0; i < 3; i++ ) { A.plane[i] = image->plane[i]; A.stride[i] = image->stride[i]; }
    A.to709 = dev->to709_f32;
    A.luma_total = (uint32_t) image->width * (uint32_t) image->height;
    A.luma_magic = cvi_div_magic( (uint32_t) image->width );
    A.chroma_width = (uint32_t) image->width / 2;
    A.chroma_total = A.chroma_width * ((uint32_t) image->height / 2);
    A.chroma_magic = cvi_div_magic( A.chroma_width );
    return cvi_launch_subsample_mpeg2( &A, cvi_stream( stream ) );
}

/* --------------------------------------------------- half arrays and tables */

int cv_half_from_float( half *out, const float *in, size_t count, cv_stream stream ) {
    if( !cvi_device() ) return CV_ERR_NOT_INITIALISED;
    if( count && (!out || !in) ) return cvi_fail( CV_ERR_ARGUMENT, "cv_half_from_float: NULL argument" );
    return cvi_launch_half_from_float( out, in, count, cvi_stream( stream ) );
}

int cv_half_to_float( float *out, const half *in, size_t count, cv_stream stream ) {
    if( !cvi_device() ) return CV_ERR_NOT_INITIALISED;
    if( count && (!out || !in) ) return cvi_fail( CV_ERR_ARGUMENT, "cv_half_to_float: NULL argument" );
    return cvi_launch_half_to_float( out, in, count, cvi_stream( stream ) );
}

int cv_half_lookup( const half *table, half *out, const half *in, size_t count, cv_stream stream ) {
    if( !cvi_device() ) return CV_ERR_NOT_INITIALISED;
    if( count && (!table || !out || !in) ) return cvi_fail( CV_ERR_ARGUMENT, "cv_half_lookup: NULL argument" );
    return cvi_launch_lookup( table, out, in, count, cvi_stream( stream ) );
}

int cv_transfer( int which, half *out, const half *in, size_t count, cv_stream stream ) {
    CvDevice *dev = cvi_device();
    if( !dev ) return CV_ERR_NOT_INITIALISED;
    if( which < 0 || which >= CV_LUT_COUNT ) return cvi_fail( CV_ERR_ARGUMENT, "cv_transfer: unknown table %d", which );
    return cv_half_lookup( dev->lut[which], out, in, count, stream );
}

/* ------------------------------------------------------------ 8-bit encodes */

int cv_upload_display_table( float rendering_intent, uint8_t **table ) {
    CvDevice *dev = cvi_device();
    if( !dev ) return CV_ERR_NOT_INITIALISED;
    if( !table ) return cvi_fail( CV_ERR_ARGUMENT, "cv_upload_display_table: NULL result" );

    uint8_t *ramp = (uint8_t *) malloc( HALF_COUNT ), *composed = (uint8_t *) malloc( HALF_COUNT );
    if( !ramp || !composed ) { free( ramp ); free( composed ); return cvi_fail( CV_ERR_MEMORY, "out of host memory" ); }

    /* widget_gl.c:297-306: sRGB transfer first, then the rendering-intent ramp */
    cv_build_intent_ramp( rendering_intent, ramp );
    const half *srgb = cv_host_lut( CV_LUT_LINEAR_TO_SRGB );
    for( int i = 0; i < HALF_COUNT; i++ )
        composed[i] = ramp[srgb[i]];

    void *devptr = NULL;
    int rc = cvi_pool_alloc( &devptr, HALF_COUNT );
    if( !rc ) rc = cvi_cuda( cudaMemcpy( devptr, composed, HALF_COUNT, cudaMemcpyHostToDevice ), "cudaMemcpy(display table)" );
    free( ramp ); free( composed );
    *table = (uint8_t *) devptr;
    return rc;
}

int cv_encode_u8( rgba_frame_f16 *frame, const uint8_t *u8_table, rgba_u8 *target, int write_back, cv_stream stream ) {
    CvDevice *dev = cvi_device();
    if( !dev ) return CV_ERR_NOT_INITIALISED;
    if( !frame || !u8_table || !target ) return cvi_fail( CV_ERR_ARGUMENT, "cv_encode_u8: NULL argument" );

    box2i win;
    box2i_intersect( &win, &frame->current_window, &frame->full_window );
    if( box2i_is_empty( &win ) ) return CV_OK;

    if( !write_back ) {
        /* one pass through the frame program: source = frame, output = u8 only */
        cv_frame_job job;
        memset( &job, 0, sizeof(job) );
        job.a.kind = CV_SRC_F16; job.a.f16 = *frame;
        job.out_u8 = target; job.u8_table = u8_table;
        job.full_window = frame->full_window;
        return cv_render_frames( &job, 1, stream );
    }

    /* with write-back the sRGB-encoded halves also replace the frame's pixels (widget_gl.c:297-299) */
    EncodeArgs args;
    memset( &args, 0, sizeof(args) );
    args.frame = (uint2 *) frame->data; args.target = (uint32_t *) target;
    args.srgb = dev->lut[CV_LUT_LINEAR_TO_SRGB]; args.ramp = u8_table;
    args.pitch = box_width( &frame->full_window ); args.ox = frame->full_window.min.x; args.oy = frame->full_window.min.y;
    args.win = to_dev( &win );
    args.width = (uint32_t) box_width( &win );
    args.total = args.width * (uint32_t)( win.max.y - win.min.y + 1 );
    args.magic = cvi_div_magic( args.width );
    args.write_back = 1;
    return cvi_launch_encode_u8( &args, cvi_stream( stream ) );
}

int cv_to_argb32( const rgba_frame_f16 *frame, uint32_t *out, cv_stream stream ) {
    CvDevice *dev = cvi_device();
    if( !dev ) return CV_ERR_NOT_INITIALISED;
    if( !frame || !out ) return cvi_fail( CV_ERR_ARGUMENT, "cv_to_argb32: NULL argument" );
    if( box2i_is_empty( &frame->current_window ) ) return CV_OK;

    ArgbArgs args;
    memset( &args, 0, sizeof(args) );
    args.frame = (const uint2 *) frame->data; args.out = out; args.ramp45 = dev->gamma45;
    args.pitch = box_width( &frame->full_window ); args.ox = frame->full_window.min.x; args.oy = frame->full_window.min.y;
    args.win = to_dev( &frame->current_window );
    args.width = (uint32_t) box_width( &frame->current_window );
    args.total = args.width * (uint32_t)( frame->current_window.max.y - frame->current_window.min.y + 1 );
    args.magic = cvi_div_magic( args.width );
    return cvi_launch_to_argb32( &args, cvi_stream( stream ) );
}

/* ------------------------------------------------------- RGB -> Y'CbCr 8-bit */

int cv_subsample_ycbcr8( rgba_frame_f16 *frame, const cv_coded *image, cv_stream stream ) {
    CvDevice *dev = cvi_device();
    if( !dev ) return CV_ERR_NOT_INITIALISED;
    if( !frame || !image ) return cvi_fail( CV_ERR_ARGUMENT, "cv_subsample_ycbcr8: NULL argument" );
    if( image->sub_x != 1 && image->sub_x != 2 && image->sub_x != 4 )
        return cvi_fail( CV_ERR_UNSUPPORTED, "cv_subsample_ycbcr8: sub_x %d", image->sub_x );
    if( image->layout != CV_LAYOUT_PLANAR && image->sub_x != 2 )
        return cvi_fail( CV_ERR_UNSUPPORTED, "cv_subsample_ycbcr8: packed layouts are 4:2:2" );
    if( image->matrix != CV_MATRIX_709 && image->matrix != CV_MATRIX_601 )
        return cvi_fail( CV_ERR_UNSUPPORTED, "cv_subsample_ycbcr8: unknown matrix" );
    if( image->layout != CV_LAYOUT_PLANAR && (((uintptr_t) image->plane[0] & 3) || (image->stride[0] & 3)) )
        return cvi_fail( CV_ERR_ARGUMENT, "cv_subsample_ycbcr8: packed rows must be 4-byte aligned" );

    /* video_subsample.c:122-128 */
    box2i pic, win;
    box2i_set( &pic, image->pic_x, image->pic_y, image->pic_x + image->width - 1, image->pic_y + image->height - 1 );
    box2i_intersect( &win, &pic, &frame->current_window );
    if( box2i_is_empty( &win ) ) return CV_OK;

    const int shift = image->sub_x == 4 ? 2 : (image->sub_x == 2 ? 1 : 0);

    SubsampleArgs args;
    memset( &args, 0, sizeof(args) );
    args.frame = (uint2 *) frame->data;
    args.pitch = box_width( &frame->full_window ); args.ox = frame->full_window.min.x; args.oy = frame->full_window.min.y;
    for( int i = 0; i < 3; i++ ) { args.plane[i] = image->plane[i]; args.stride[i] = image->stride[i]; }
    args.layout = image->layout; args.sub_shift = shift; args.matrix = image->matrix;
    args.pic_x = image->pic_x; args.pic_y = image->pic_y;
    args.wmin = win.min.x - image->pic_x; args.wmax = win.max.x - image->pic_x;
    args.row0 = win.min.y - image->pic_y; args.rows = win.max.y - win.min.y + 1;
    args.tx0 = args.wmin >> shift;
    args.tx_count = (uint32_t)( (args.wmax >> shift) - args.tx0 + 1 );
    args.total = args.tx_count * (uint32_t) args.rows;
    args.magic = cvi_div_magic( args.tx_count );
    args.lut = dev->lut[CV_LUT_LINEAR_TO_709];

    /* one pass: coded bytes out, and -- :147 -- the table applied in place to the window of the INPUT frame by the
     * thread that owns each pixel (neighbours get each other's values through shuffles, never by re-reading) */
    return cvi_launch_subsample( &args, cvi_stream( stream ) );
}
