/*
 * tests/rect_selftest.c -- CPU check of the rectangle decomposition (canvas_b200/csrc/cv_rects.c), compiled and run by
 * tests/test_rects.py.  Random window sets are turned into a DevJob the way build_job() (cv_api.c) does it, the plan is
 * built, and every pixel pair of the result window is compared with the per-pixel rule of the exact path
 * (pixel_exact(), cv_device.cuh): which sources and layers cover it, and whether the first covering layer lies inside
 * the running union box of what is below it.  Pieces must tile each row of their band exactly once.
 */
#include "cv_internal.h"

#include <stdio.h>
#include <stdlib.h>
#include <string.h>

uint32_t cvi_div_magic( uint32_t d ) { return d <= 1 ? 0xFFFFFFFFu : (uint32_t)( (UINT64_C(1) << 32) / d ); }

static uint32_t rng_state;
static uint32_t rnd( void ) { rng_state = rng_state * 1664525u + 1013904223u; return rng_state >> 8; }
static int rnd_in( int lo, int hi ) { return lo + (int)( rnd() % (uint32_t)( hi - lo + 1 ) ); }

static int empty( const DevBox *b ) { return b->x1 < b->x0 || b->y1 < b->y0; }
static int inside( const DevBox *b, int x, int y ) { return x >= b->x0 && x <= b->x1 && y >= b->y0 && y <= b->y1; }
static DevBox unite( const DevBox *a, const DevBox *b ) {
    if( empty( a ) ) return *b;
    if( empty( b ) ) return *a;
    DevBox d = { a->x0 < b->x0 ? a->x0 : b->x0, a->y0 < b->y0 ? a->y0 : b->y0, a->x1 > b->x1 ? a->x1 : b->x1, a->y1 > b->y1 ? a->y1 : b->y1 };
    return d;
}
static DevBox random_box( int w, int h ) {
    DevBox b;
    b.x0 = rnd_in( 0, w - 1 ); b.x1 = rnd_in( b.x0, w - 1 );
    b.y0 = rnd_in( 0, h - 1 ); b.y1 = rnd_in( b.y0, h - 1 );
    if( rnd() % 4 == 0 ) { b.x0 = 0; b.x1 = w - 1; }
    if( rnd() % 4 == 0 ) { b.y0 = 0; b.y1 = h - 1; }
    return b;
}

/* returns 0, or a positive code naming the first broken property */
static int one_case( int ppl, int verbose ) {
    const int w = rnd_in( 2, 400 ), h = rnd_in( 1, 24 ), full_x = rnd_in( -5, 5 );
    static DevJob job;
    memset( &job, 0, sizeof(job) );
    job.flags = JOB_FLAG_SEG;
    job.out16 = (void *) 16;
    job.full_x = full_x;
    job.bottom_mode = (int)( rnd() % 3 );
    job.n_layers = rnd_in( job.bottom_mode ? 0 : 1, 6 );

    DevBox under = { 0, 0, -1, -1 };
    if( job.bottom_mode != DEV_BOTTOM_NONE ) {
        job.a.kind = CV_SRC_F16; job.a.cw = random_box( w, h ); job.a.cw.x0 += full_x; job.a.cw.x1 += full_x;
        under = job.a.cw;
    }
    if( job.bottom_mode == DEV_BOTTOM_CROSS ) {
        job.b.kind = CV_SRC_F16; job.b.cw = random_box( w, h ); job.b.cw.x0 += full_x; job.b.cw.x1 += full_x;
        under = unite( &job.a.cw, &job.b.cw );
    }
    for( int k = 0; k < job.n_layers; k++ ) {
        job.layer[k].mix = 1.0f;
        job.layer[k].cw = random_box( w, h ); job.layer[k].cw.x0 += full_x; job.layer[k].cw.x1 += full_x;
        job.layer[k].under = under;
        under = unite( &under, &job.layer[k].cw );
    }
    job.region = under;
    job.item_x0 = full_x + ((under.x0 - full_x) & ~1);
    job.items_per_row = (uint32_t)( (under.x1 - job.item_x0) / 2 + 1 );
    job.rows = (uint32_t)( under.y1 - under.y0 + 1 );

    static RectPlan plan;
    static RectPiece pieces[CV_RECT_MAX_PIECES];
    const int rpu = rnd_in( 1, 9 );
    if( cvi_rect_plan( &job, ppl, rpu, &plan, pieces, CV_RECT_MAX_PIECES ) )
        return 1;

    uint32_t start = 0;
    uint64_t row_pieces = 0;
    int y = under.y0;
    for( int b = 0; b < plan.n_bands; b++ ) {
        const RectBand *B = &plan.band[b];
        if( B->y0 != y || B->ustart != start || B->ppr == 0 || B->magic != cvi_div_magic( B->ppr ) ) return 2;
        uint32_t at = 0;
        for( uint32_t i = 0; i < B->ppr; i++ ) {
            const RectPiece *p = &pieces[B->first_entry + i];
            if( p->pair0 != at || p->pairs == 0 || p->pairs > 32 * ppl ) return 3;
            at += p->pairs;

            for( int row = B->y0; row < B->y0 + (int) B->rows; row++ ) {
                for( uint32_t pair = p->pair0; pair < p->pair0 + p->pairs; pair++ ) {
                    const int x = job.item_x0 + 2 * (int) pair;
                    /* what the exact path would do at the two pixels */
                    int bottom[2] = { 0, 0 }, first_copy[2] = { 0, 0 }, in_region[2];
                    uint32_t mask[2] = { 0, 0 };
                    for( int s = 0; s < 2; s++ ) {
                        in_region[s] = inside( &job.region, x + s, row );
                        if( job.bottom_mode != DEV_BOTTOM_NONE && inside( &job.a.cw, x + s, row ) ) bottom[s] |= 1;
                        if( job.bottom_mode == DEV_BOTTOM_CROSS && inside( &job.b.cw, x + s, row ) ) bottom[s] |= 2;
                        int seen = 0;
                        for( int k = 0; k < job.n_layers; k++ ) {
                            if( !inside( &job.layer[k].cw, x + s, row ) ) continue;
                            mask[s] |= 1u << k;
                            if( !seen && !bottom[s] && !inside( &job.layer[k].under, x + s, row ) ) first_copy[s] = 1;
                            seen = 1;
                        }
                    }
                    const int uniform = in_region[0] && in_region[1] && bottom[0] == bottom[1] && mask[0] == mask[1] && first_copy[0] == first_copy[1];
                    if( p->flags & RECT_RAGGED )
                        continue;                               /* per-item code: always right */
                    if( !uniform ) {
                        /* a pair that needs the per-item code must be ragged -- unless both pixels agree anyway */
                        if( verbose ) fprintf( stderr, "pair %u row %d not uniform but piece not ragged\n", pair, row );
                        return 5;
                    }
                    if( p->bottom != bottom[0] || p->mask != mask[0] ) return 6;
                    if( ((p->flags & RECT_FIRST_COPY) != 0) != first_copy[0] && mask[0] ) return 7;
                }
            }
        }
        if( at != job.items_per_row ) return 8;
        start += B->ppr * ((B->rows + (uint32_t) rpu - 1) / (uint32_t) rpu);
        row_pieces += (uint64_t) B->ppr * B->rows;
        y += (int) B->rows;
    }
    if( y != under.y1 + 1 || start != plan.units_per_frame || row_pieces != plan.row_pieces_per_frame || plan.rows_per_unit != (uint32_t) rpu ) return 9;
    return 0;
}

int main( int argc, char **argv ) {
    const int cases = argc > 1 ? atoi( argv[1] ) : 2000;
    rng_state = argc > 2 ? (uint32_t) atoi( argv[2] ) : 12345u;
    for( int i = 0; i < cases; i++ ) {
        const uint32_t saved = rng_state;
        const int ppl = 2 + (int)( rnd() % 2 );
        const int rc = one_case( ppl, 0 );
        if( rc ) {
            printf( "case %d (state %u, ppl %d): property %d broken\n", i, saved, ppl, rc );
            return 1;
        }
    }
    printf( "ok %d cases\n", cases );
    return 0;
}
