/*
 * canvas_b200/csrc/cv_rects.c -- host-side decomposition of a WINDOWED frame program into rectangles of
 * constant coverage, for the rectangle-walking kernel (k_rect, cv_kernels.cu).
 *
 * The reference walks nine regions per mix (above / below / left / right / inside of the two windows,
 * /root/reference/src/cprocess/video_mix.c:123-235 for the crossfade, :251-368 for "over") and does so once per layer
 * (workspace.c:530-547).  The fused frame program sees all windows at once: cutting the result window at every
 * window edge gives a grid of cells in which the set of inputs that cover a pixel is constant, so a cell needs no
 * window test at all -- it is a small "every input covers everything" job.  This file builds that grid:
 *
 *   - columns are cut on the PAIR grid (a pixel pair = 16 bytes is the unit the kernels move); a window edge that
 *     falls inside a pair makes that pair a one-pair-wide "ragged" column, rendered by the per-item code;
 *   - rows are cut at every window's top and bottom; consecutive rows with the same column structure form a BAND;
 *   - the cells of a band with equal coverage are merged into runs, and a run is split into PIECES of at most
 *     32 * ppl pairs (what one warp renders of a row at a time), balanced so that no piece is a short tail;
 *   - the kernel's unit of work is a piece over a GROUP of consecutive rows of its band (rows_per_unit): everything
 *     about a unit -- which inputs, where -- is decoded once and its rows are then walked by adding pitches.
 *
 * A piece records which inputs cover it: the bottom class (none / a / b / crossfade of both), a bit per layer, and
 * whether the first layer that covers it is fetched as it is (nothing below it: video_copy_frame_alpha_f32,
 * video_mix.c:243-246) or goes "over" zeros (it lies inside the union box of what is below, where uncovered pixels
 * are zero -- the intended semantics of video_mix.c:161-169).
 *
 * Pure host arithmetic on boxes: no pixel is touched here.
 */
#include "cv_internal.h"

#include <stdlib.h>
#include <string.h>

typedef struct Cut { int32_t at; } Cut;

static int cmp_int( const void *a, const void *b ) {
    const int32_t x = *(const int32_t *) a, y = *(const int32_t *) b;
    return x < y ? -1 : (x > y);
}

static int sort_unique( int32_t *v, int n ) {
    qsort( v, (size_t) n, sizeof(int32_t), cmp_int );
    int m = 0;
    for( int i = 0; i < n; i++ )
        if( m == 0 || v[m - 1] != v[i] ) v[m++] = v[i];
    return m;
}

static int box_is_empty( const DevBox *b ) { return b->x1 < b->x0 || b->y1 < b->y0; }

/* pixels [x0, x1] x rows [y0, y1] all inside b */
static int box_covers( const DevBox *b, int x0, int x1, int y0, int y1 ) {
    return !box_is_empty( b ) && b->x0 <= x0 && x1 <= b->x1 && b->y0 <= y0 && y1 <= b->y1;
}

typedef struct Cell { uint32_t mask; uint8_t bottom, flags; } Cell;

static int same_cell( const Cell *a, const Cell *b ) { return a->mask == b->mask && a->bottom == b->bottom && a->flags == b->flags; }

/* a coded source the pair form applies to (what source_uniform() in cv_kernels.cu asks, minus the coverage) */
static int coded_pairable( const DevSrc *s, const DevJob *job ) {
    if( s->kind != CV_SRC_CODED ) return 1;
    if( s->sub_shift != 1 || ((job->item_x0 - s->pic_x) & 1) ) return 0;
    return s->layout != CV_LAYOUT_PLANAR || ((((uintptr_t) s->plane[0] | (uintptr_t) s->stride[0]) & 1) == 0);
}

int cvi_rect_applies( const DevJob *job ) {
    if( !(job->flags & JOB_FLAG_SEG) || (job->flags & JOB_FLAG_EXACT) || job->out32 || !(job->out16 || job->out8) )
        return 0;
    if( job->bottom_mode != DEV_BOTTOM_NONE && !coded_pairable( &job->a, job ) ) return 0;
    if( job->bottom_mode == DEV_BOTTOM_CROSS && !coded_pairable( &job->b, job ) ) return 0;
    if( job->bottom_mode != DEV_BOTTOM_NONE && job->a.kind == CV_SRC_F32 ) return 0;
    if( job->bottom_mode == DEV_BOTTOM_CROSS && job->b.kind == CV_SRC_F32 ) return 0;
    for( int k = 0; k < job->n_layers; k++ )
        if( job->layer[k].mix != 1.0f ) return 0;
    return 1;
}

/* adds the pair-grid cuts of the pixel interval [x0, x1] (inclusive, inside the region); pairs split by an edge
 * become ragged columns */
static void add_x_cuts( const DevJob *job, int x0, int x1, int32_t *cuts, int *n, int32_t *ragged, int *n_ragged ) {
    const int lo = x0 - job->item_x0, hi = x1 - job->item_x0;
    if( lo & 1 ) { ragged[(*n_ragged)++] = lo >> 1; cuts[(*n)++] = lo >> 1; cuts[(*n)++] = (lo >> 1) + 1; }
    else cuts[(*n)++] = lo >> 1;
    if( hi & 1 ) cuts[(*n)++] = (hi >> 1) + 1;
    else { ragged[(*n_ragged)++] = hi >> 1; cuts[(*n)++] = hi >> 1; cuts[(*n)++] = (hi >> 1) + 1; }
}

int cvi_rect_plan( const DevJob *job, int ppl, int rows_per_unit, RectPlan *plan, RectPiece *pieces, int capacity ) {
    enum { MAX_BOXES = CV_MAX_LAYERS + 2, MAX_CUTS = 4 * (MAX_BOXES + 1) + 2 };
    int32_t xcut[MAX_CUTS], ycut[MAX_CUTS], ragged[2 * (MAX_BOXES + 1)];
    int nx = 0, ny = 0, n_ragged = 0;
    const DevBox *R = &job->region;
    const int n = job->n_layers;

    memset( plan, 0, sizeof(*plan) );
    if( !cvi_rect_applies( job ) || box_is_empty( R ) || rows_per_unit < 1 ) return -1;

    /* the region itself, then every input's window (already clipped to the output's full window; clip to the region) */
    add_x_cuts( job, R->x0, R->x1, xcut, &nx, ragged, &n_ragged );
    ycut[ny++] = R->y0; ycut[ny++] = R->y1 + 1;

    const DevBox *boxes[MAX_BOXES];
    int n_boxes = 0;
    if( job->bottom_mode != DEV_BOTTOM_NONE ) boxes[n_boxes++] = &job->a.cw;
    if( job->bottom_mode == DEV_BOTTOM_CROSS ) boxes[n_boxes++] = &job->b.cw;
    for( int k = 0; k < n; k++ ) boxes[n_boxes++] = &job->layer[k].cw;

    for( int i = 0; i < n_boxes; i++ ) {
        const DevBox *b = boxes[i];
        if( box_is_empty( b ) ) continue;
        const int x0 = b->x0 > R->x0 ? b->x0 : R->x0, x1 = b->x1 < R->x1 ? b->x1 : R->x1;
        const int y0 = b->y0 > R->y0 ? b->y0 : R->y0, y1 = b->y1 < R->y1 ? b->y1 : R->y1;
        if( x1 < x0 || y1 < y0 ) continue;
        add_x_cuts( job, x0, x1, xcut, &nx, ragged, &n_ragged );
        ycut[ny++] = y0; ycut[ny++] = y1 + 1;
    }
    nx = sort_unique( xcut, nx );
    ny = sort_unique( ycut, ny );
    n_ragged = sort_unique( ragged, n_ragged );

    const uint32_t piece_pairs = 32u * (uint32_t) ppl;
    int n_entries = 0, n_bands = 0;
    Cell cells[MAX_CUTS];

    for( int yi = 0; yi + 1 < ny; yi++ ) {
        const int y0 = ycut[yi], y1 = ycut[yi + 1] - 1;

        /* coverage of every column of this row range */
        for( int xi = 0; xi + 1 < nx; xi++ ) {
            const int c0 = xcut[xi], c1 = xcut[xi + 1];                     /* pair columns [c0, c1) */
            const int px0 = job->item_x0 + 2 * c0, px1 = job->item_x0 + 2 * c1 - 1;
            Cell *c = &cells[xi];
            c->mask = 0; c->bottom = 0; c->flags = 0;

            for( int r = 0; r < n_ragged; r++ )
                if( ragged[r] == c0 && c1 == c0 + 1 ) c->flags = RECT_RAGGED;
            if( c->flags & RECT_RAGGED ) continue;                           /* the per-item code needs no coverage */

            if( job->bottom_mode != DEV_BOTTOM_NONE && box_covers( &job->a.cw, px0, px1, y0, y1 ) ) c->bottom |= 1;
            if( job->bottom_mode == DEV_BOTTOM_CROSS && box_covers( &job->b.cw, px0, px1, y0, y1 ) ) c->bottom |= 2;
            int first = 1;
            for( int k = 0; k < n; k++ ) {
                if( !box_covers( &job->layer[k].cw, px0, px1, y0, y1 ) ) continue;
                c->mask |= 1u << k;
                if( first && !c->bottom && !box_covers( &job->layer[k].under, px0, px1, y0, y1 ) )
                    c->flags |= RECT_FIRST_COPY;
                first = 0;
            }
        }

        /* runs of equal coverage -> pieces */
        const int band_first = n_entries;
        for( int xi = 0; xi + 1 < nx; ) {
            int xj = xi + 1;
            while( xj + 1 < nx && same_cell( &cells[xi], &cells[xj] ) ) xj++;
            const uint32_t c0 = (uint32_t) xcut[xi], width = (uint32_t)( xcut[xj] - xcut[xi] );
            const uint32_t count = (width + piece_pairs - 1) / piece_pairs;
            const uint32_t base = width / count, extra = width % count;
            uint32_t at = c0;
            for( uint32_t i = 0; i < count; i++ ) {
                if( n_entries >= capacity ) return -1;
                RectPiece *p = &pieces[n_entries++];
                p->pair0 = at;
                p->pairs = (uint16_t)( base + (i < extra ? 1u : 0u) );
                p->bottom = cells[xi].bottom;
                p->flags = cells[xi].flags;
                p->mask = cells[xi].mask;
                p->reserved = 0;
                at += p->pairs;
            }
            xi = xj;
        }

        /* the same column structure as the band above: extend it */
        const int ppr = n_entries - band_first;
        if( n_bands > 0 ) {
            RectBand *prev = &plan->band[n_bands - 1];
            if( prev->ppr == (uint32_t) ppr && prev->y0 + (int32_t) prev->rows == y0 &&
                    memcmp( &pieces[prev->first_entry], &pieces[band_first], sizeof(RectPiece) * (size_t) ppr ) == 0 ) {
                prev->rows += (uint32_t)( y1 - y0 + 1 );
                n_entries = band_first;
                continue;
            }
        }
        if( n_bands >= CV_RECT_MAX_BANDS ) return -1;
        RectBand *b = &plan->band[n_bands++];
        b->y0 = y0;
        b->rows = (uint32_t)( y1 - y0 + 1 );
        b->ppr = (uint32_t) ppr;
        b->magic = cvi_div_magic( (uint32_t) ppr );
        b->first_entry = (uint32_t) band_first;
    }

    /* units: a piece over a group of rows_per_unit rows of its band */
    uint64_t units = 0, row_pieces = 0;
    for( int i = 0; i < n_bands; i++ ) {
        RectBand *b = &plan->band[i];
        b->ustart = (uint32_t) units;
        units += (uint64_t)( (b->rows + (uint32_t) rows_per_unit - 1) / (uint32_t) rows_per_unit ) * b->ppr;
        row_pieces += (uint64_t) b->rows * b->ppr;
        if( units > 0x7FFFFFFFu ) return -1;
    }

    plan->n_bands = n_bands;
    plan->n_entries = (uint32_t) n_entries;
    plan->rows_per_unit = (uint32_t) rows_per_unit;
    plan->units_per_frame = (uint32_t) units;
    plan->frame_magic = cvi_div_magic( (uint32_t) units );
    plan->row_pieces_per_frame = row_pieces;
    return units ? 0 : -1;
}
