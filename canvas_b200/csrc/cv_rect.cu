/*
 * canvas_b200/csrc/cv_rect.cu -- k_rect: frame programs whose inputs have partial windows, rendered as rectangles of
 * constant coverage (the decomposition is host arithmetic on boxes, cv_rects.c), and its launcher.
 * Compiled like cv_kernels.cu: sm_100a only, -fmad=false.
 */
#include "cv_device.cuh"

/* ------------------------------------------------ windowed frame programs as rectangles of constant coverage */

/*
 * k_rect: a frame program whose inputs have PARTIAL windows, rendered without a single window test.  The host cuts
 * the result window at every window edge (cv_rects.c) into bands of rows and, per band, pieces of at most 32*PPL
 * pixel pairs inside which the set of covering inputs is constant.  A warp's unit of work is one piece over a GROUP
 * of consecutive rows: which ring slots to fill, which layers to run, where the sources and the output start is
 * decoded ONCE per unit (one unit ahead, so that its loads are never waited for); the rows of the unit are then walked
 * by adding pitches -- per row a warp arms one mbarrier, issues one bulk copy (UBLKCP) per covering f16 input,
 * loads the 4:2:2 groups of the next row into registers, and composites the row that has arrived:
 *   - absent inputs cost nothing (no copy is issued, the layer is skipped by a warp-uniform test on the piece's mask);
 *   - the first covering layer of a piece with nothing below is fetched as it is (RECT_FIRST_COPY), otherwise the
 *     stack starts from zeros -- the "pixels covered by nobody are zero" rule of the union box;
 *   - a one-pair column split by an odd window edge (RECT_RAGGED) goes through the per-item code, out of line.
 * It replaces the window-classifying mode of k_seg (per-segment classification, 168 registers, 11 warps per SM) for
 * every job whose layer mixes are 1 -- what the workspace compositor produces (workspace.c:547).
 */
struct RectArgs {
    int32_t n_bottom, slots;                /* f16 bottom sources staged in ring slots 0..n_bottom-1, then one slot per layer */
    uint32_t stage_bytes, lut_bytes, u8_bytes;
    uint32_t all_units;                     /* frames x plan.units_per_frame */
};

#define CV_RECT_PARK_BYTES 576u          /* per warp: 64 bytes lane-uniform + 32 x 16 bytes per lane (the unit waiting to be rendered) */

#ifndef CV_RECT_MAX_THREADS
#define CV_RECT_MAX_THREADS 384       /* 168 registers: two units and the 4:2:2 groups of two rows stay in registers (512 threads spilled them) */
#endif
/* The instantiations with fused layer filters are bounded for 512 threads instead (128 registers, 16 to 96 bytes of
 * spills): measured on the same box, 4K 4 layers with filters 7 620 -> 8 190 fps, while the plain ones lose at 512
 * (1080p 4 layers 90.7 k -> 63.0 k at three pairs per lane, 90.1 k at two; the 4K timeline 11.8 k -> 10.7 k). */
#define CV_RECT_OPS_MAX_THREADS 512

template<int SHAPE, int PPL, bool OPS>
__global__ void __launch_bounds__( OPS ? CV_RECT_OPS_MAX_THREADS : CV_RECT_MAX_THREADS, 1 )
k_rect( const __grid_constant__ DevJob J0, const DevJob *__restrict__ jobs, const __grid_constant__ RectPlan plan,
        const RectPiece *__restrict__ pieces, const __grid_constant__ RectArgs A,
        const uint16_t *__restrict__ table /* staged: the compact conversion table */, const uint16_t *__restrict__ full_table ) {
    extern __shared__ __align__(128) unsigned char smem[];
    constexpr uint32_t SEG_BYTES = 32u * PPL * 16u;
    constexpr bool COMPACT = SHAPE == KB_CODED;

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, warps = blockDim.x >> 5;
    const int n = J0.n_layers, nb = A.n_bottom;
    const float2 one = make_float2( J0.one, J0.one );
    const uint32_t table_bytes = A.lut_bytes + A.u8_bytes;

    const uint16_t *lut = reinterpret_cast<const uint16_t *>( smem );
    const uint16_t *slow_lut = COMPACT ? full_table : lut;
    const uint8_t *u8s = smem + A.lut_bytes;
    uint64_t *table_bar = reinterpret_cast<uint64_t *>( smem + table_bytes );
    uint64_t *bars = table_bar + 2 + warp * 2;
    unsigned char *ring = smem + table_bytes + (((size_t)( warps + 1 ) * 16 + 127) & ~(size_t) 127) + (size_t) warp * 2 * A.stage_bytes;

    if( table_bytes && threadIdx.x == 0 ) {
        mbar_init( table_bar, 1 );
        asm volatile( "fence.mbarrier_init.release.cluster;" ::: "memory" );
        mbar_expect_tx( table_bar, table_bytes );
        for( uint32_t off = 0; off < A.lut_bytes; off += 32768u )
            bulk_g2s( smem + off, reinterpret_cast<const char *>( table ) + off, 32768u, table_bar );
        for( uint32_t off = 0; off < A.u8_bytes; off += 32768u )
            bulk_g2s( smem + A.lut_bytes + off, reinterpret_cast<const char *>( J0.u8tab ) + off, 32768u, table_bar );
    }
    if( lane == 0 ) {
        mbar_init( bars + 0, 1 );
        mbar_init( bars + 1, 1 );
        asm volatile( "fence.mbarrier_init.release.cluster;" ::: "memory" );
    }
    __syncwarp();

    const bool has_bottom = SHAPE != KB_LAYERS && J0.bottom_mode != DEV_BOTTOM_NONE;
    const bool cross_job = has_bottom && J0.bottom_mode == DEV_BOTTOM_CROSS;
    const bool a_coded = SHAPE == KB_CODED && has_bottom && J0.a.kind == CV_SRC_CODED;
    const bool b_coded = SHAPE == KB_CODED && cross_job && J0.b.kind == CV_SRC_CODED;
    const bool a_f16 = has_bottom && J0.a.kind == CV_SRC_F16, b_f16 = cross_job && J0.b.kind == CV_SRC_F16;
    const int slot_b = a_f16 ? 1 : 0;
    const bool f16_only = J0.out16 && !J0.out8;
    const uint32_t out_row_bytes = (uint32_t) J0.out_pitch * 8u;

    const uint32_t first = blockIdx.x * warps + warp, stride = gridDim.x * warps;
    const uint32_t mine = first < A.all_units ? (A.all_units - first - 1) / stride + 1 : 0;

    /* a unit: one piece over up to plan.rows_per_unit rows; `src` / `dst16` are this LANE's addresses.  The unit being
     * rendered lives in registers; the NEXT one is decoded a whole unit ahead and waits in shared memory (one record per
     * warp: the lane-uniform part once, the lanes' source addresses next to it), so that looking ahead costs no registers
     * in the row loop. */
    struct Unit {
        const DevJob *job;
        char *dst16;                    /* this lane's first output pair */
        int y, x0;
        uint32_t rows, pairs, mask, bottom, flags;
        uint32_t row_bytes;             /* what one row's copies bring into the ring (the mbarrier's transaction count) */
        float mix_a, mix_b;
        const char *src;                /* this lane's ring slot: where its next row comes from (NULL: not covered) */
        uint32_t src_pitch;
    };
    struct ParkedCommon { const DevJob *job; char *dst16; int y, x0; uint32_t rows, pairs, mask, bottom, flags, row_bytes; float mix_a, mix_b; };
    struct ParkedLane { const char *src; uint32_t src_pitch, pad; };
    static_assert( sizeof(ParkedCommon) == 56 && sizeof(ParkedLane) == 16, "parked unit layout" );
    unsigned char *const park_base = smem + table_bytes + (((size_t)( warps + 1 ) * 16 + 127) & ~(size_t) 127) + (size_t) warps * 2 * A.stage_bytes
                                   + (size_t) warp * CV_RECT_PARK_BYTES;
    ParkedCommon *const park_common = reinterpret_cast<ParkedCommon *>( park_base );
    ParkedLane *const park_lane = reinterpret_cast<ParkedLane *>( park_base + 64 ) + lane;

    auto park = [&]( const Unit &U ) {
        if( lane == 0 ) {
            ParkedCommon c;
            c.job = U.job; c.dst16 = U.dst16; c.y = U.y; c.x0 = U.x0; c.rows = U.rows; c.pairs = U.pairs; c.mask = U.mask;
            c.bottom = U.bottom; c.flags = U.flags; c.row_bytes = U.row_bytes; c.mix_a = U.mix_a; c.mix_b = U.mix_b;
            *park_common = c;
        }
        park_lane->src = U.src; park_lane->src_pitch = U.src_pitch;
        __syncwarp();
    };
    auto unpark = [&]( Unit &U ) {
        const ParkedCommon c = *park_common;
        U.job = c.job; U.dst16 = c.dst16 + lane * 16; U.y = c.y; U.x0 = c.x0; U.rows = c.rows; U.pairs = c.pairs; U.mask = c.mask;
        U.bottom = c.bottom; U.flags = c.flags; U.row_bytes = c.row_bytes; U.mix_a = c.mix_a; U.mix_b = c.mix_b;
        U.src = park_lane->src; U.src_pitch = park_lane->src_pitch;
    };

    auto decode = [&]( uint32_t index, Unit &U ) {
        uint32_t frame = 0, r = index;
        if( jobs ) fast_divmod( index, plan.units_per_frame, plan.frame_magic, frame, r );
        int band = 0;
        while( band + 1 < plan.n_bands && r >= plan.band[band + 1].ustart ) band++;
        const RectBand &B = plan.band[band];
        uint32_t group, pc;
        fast_divmod( r - B.ustart, B.ppr, B.magic, group, pc );
        const RectPiece e = pieces[B.first_entry + pc];
        const uint32_t row0 = group * plan.rows_per_unit;

        U.job = jobs ? jobs + frame : &J0;
        const DevJob &J = *U.job;
        U.y = B.y0 + (int) row0;
        U.rows = min( plan.rows_per_unit, B.rows - row0 );
        U.x0 = J0.item_x0 + 2 * (int) e.pair0;
        U.pairs = e.pairs; U.mask = e.mask; U.bottom = e.bottom; U.flags = e.flags;
        U.mix_a = J.mix_a; U.mix_b = J.mix_b;
        U.dst16 = static_cast<char *>( J.out16 ) + ((int64_t)( U.y - J0.full_y ) * J0.out_pitch + ( U.x0 - J0.full_x )) * 8;   /* + lane * 16 when unparked */

        const bool ragged = (e.flags & RECT_RAGGED) != 0;
        U.src = NULL; U.src_pitch = 0;
        if( lane < A.slots && !ragged ) {
            bool present; const void *origin; int pitch;
            if( lane < nb ) {
                const bool is_a = (lane == 0 && a_f16);
                present = is_a ? (e.bottom & 1u) != 0 : (e.bottom & 2u) != 0;
                origin = is_a ? J.a.origin : J.b.origin; pitch = is_a ? J0.a.pitch : J0.b.pitch;
            }
            else {
                present = (e.mask >> (lane - nb)) & 1u;
                origin = J.layer[lane - nb].origin; pitch = J0.layer[lane - nb].pitch;
            }
            if( present ) {
                U.src = static_cast<const char *>( origin ) + ((int64_t) U.y * pitch + U.x0) * 8;
                U.src_pitch = (uint32_t) pitch * 8u;
            }
        }
        uint32_t copies = ragged ? 0u : (uint32_t) __popc( e.mask );
        if( !ragged && a_f16 && (e.bottom & 1u) ) copies++;
        if( !ragged && b_f16 && (e.bottom & 2u) ) copies++;
        U.row_bytes = (uint32_t) e.pairs * 16u * copies;
    };

    /* starts everything row `y` of unit U needs: its bulk copies into ring stage `stage`, its 4:2:2 groups into registers */
    auto issue_row = [&]( Unit &U, int y, int stage, CodedRegs<PPL> &ga, CodedRegs<PPL> &gb ) {
        if( SHAPE == KB_CODED && !(U.flags & RECT_RAGGED) ) {
            SegPos P;
            P.y = y; P.x0 = U.x0; P.pairs = U.pairs;
            if( a_coded && (U.bottom & 1u) ) coded_load<PPL>( J0.a, U.job->a.plane, P, lane, ga );
            if( b_coded && (U.bottom & 2u) ) coded_load<PPL>( J0.b, U.job->b.plane, P, lane, gb );
        }
        if( !A.slots )
            return;
        if( lane == 0 )
            mbar_expect_tx( bars + stage, U.row_bytes );
        __syncwarp();           /* every lane has finished reading this stage's previous contents */
        if( U.src ) {
            const char *from = U.src;
            if constexpr( OPS ) {       /* CV_OP_WEAVE: this layer's even rows come from its other frame (Pulldown23RemovalFilter.c:93-99) */
                if( lane >= nb && (J0.layer[lane - nb].op_flags & CV_OP_WEAVE) && !(y & 1) )
                    from = static_cast<const char *>( U.job->layer[lane - nb].weave_origin ) + ((int64_t) y * J0.layer[lane - nb].weave_pitch + U.x0) * 8;
            }
            bulk_g2s( ring + (size_t) stage * A.stage_bytes + (uint32_t) lane * SEG_BYTES, from, U.pairs * 16u, bars + stage );
            U.src += U.src_pitch;
        }
    };

    Unit cur;
    CodedRegs<PPL> ca, cb, na, nbg;
    #pragma unroll
    for( int q = 0; q < PPL; q++ ) ca.w[q] = cb.w[q] = na.w[q] = nbg.w[q] = 0x80808080u;
    ca.extra = cb.extra = na.extra = nbg.extra = 0x80808080u;

    if( mine ) {
        Unit head;
        decode( first, head );
        issue_row( head, head.y, 0, na, nbg );
        park( head );
    }

    if( table_bytes ) {
        __syncthreads();            /* table_bar initialised before anyone polls it */
        mbar_wait( table_bar, 0 );
    }

    CodedConst const_a, const_b;
    if( SHAPE == KB_CODED ) {
        const_a = coded_const( J0.a.layout, J0.a.matrix, lut );
        const_b = coded_const( J0.b.layout, J0.b.matrix, lut );
    }

    uint32_t it = 0;                /* rows rendered so far: ring stage and barrier phase */
    for( uint32_t u = 0; u < mine; u++ ) {
        unpark( cur );
        __syncwarp();                   /* everyone has taken the parked unit before the next one is written over it */
        const bool more = u + 1 < mine;
        if( more ) {
            Unit ahead;
            decode( first + (u + 1) * stride, ahead );
            park( ahead );
        }

        const DevJob &J = *cur.job;
        char *dst16 = cur.dst16;
        const bool whole = cur.pairs == 32u * PPL && f16_only;

        for( uint32_t r = 0; r < cur.rows; r++, it++, dst16 += out_row_bytes ) {
            const int s = (int)( it & 1u );
            const int y = cur.y + (int) r;
            ca = na; cb = nbg;
            if( r + 1 < cur.rows ) issue_row( cur, y + 1, s ^ 1, na, nbg );
            else if( more ) {           /* the first row of the next unit: its record, with the advanced source address written back */
                Unit ahead;
                unpark( ahead );
                issue_row( ahead, ahead.y, s ^ 1, na, nbg );
                park_lane->src = ahead.src;
            }

            if( A.slots )
                mbar_wait( bars + s, (it >> 1) & 1u );

            if( cur.flags & RECT_RAGGED ) {
                const uint32_t item0 = (uint32_t)( y - J0.region.y0 ) * J0.items_per_row + (uint32_t)( (cur.x0 - J0.item_x0) >> 1 );
                seg_ragged_call<SHAPE, PPL>( J, item0, cur.pairs, (uint32_t) lane, slow_lut );
                continue;
            }

            const unsigned char *mine_px = ring + (size_t) s * A.stage_bytes + (uint32_t) lane * 16u;

            P2 acc[PPL];
            Guard guard[PPL];
            #pragma unroll
            for( int q = 0; q < PPL; q++ )
                acc[q].r = acc[q].g = acc[q].b = acc[q].a = make_float2( 0.0f, 0.0f );

            /* ---- bottom: nothing, one source (alpha scaled by its weight), or the crossfade of both ---- */
            if( SHAPE != KB_LAYERS && cur.bottom ) {
                if( cur.bottom == 3u ) {
                    #pragma unroll
                    for( int q = 0; q < PPL; q++ ) {
                        P2 va, vb;
                        if( a_coded ) va = unpack_pair( coded_halves_from_regs<PPL, COMPACT>( const_a, ca, q, lane, lut, one ) );
                        else va = unpack_pair( lds_u4( mine_px + q * 512 ) );
                        if( b_coded ) vb = unpack_pair( coded_halves_from_regs<PPL, COMPACT>( const_b, cb, q, lane, lut, one ) );
                        else vb = unpack_pair( lds_u4( mine_px + slot_b * SEG_BYTES + q * 512 ) );
                        guard[q].see( va );
                        guard[q].see( vb );
                        acc[q] = cross_pair( va, vb, cur.mix_a, cur.mix_b, one );
                    }
                }
                else {
                    const bool from_a = cur.bottom == 1u;
                    const float weight = cross_job ? (from_a ? cur.mix_a : cur.mix_b) : J0.single_alpha;
                    #pragma unroll
                    for( int q = 0; q < PPL; q++ ) {
                        P2 v;
                        if( from_a ) {
                            if( a_coded ) v = unpack_pair( coded_halves_from_regs<PPL, COMPACT>( const_a, ca, q, lane, lut, one ) );
                            else v = unpack_pair( lds_u4( mine_px + q * 512 ) );
                        }
                        else {
                            if( b_coded ) v = unpack_pair( coded_halves_from_regs<PPL, COMPACT>( const_b, cb, q, lane, lut, one ) );
                            else v = unpack_pair( lds_u4( mine_px + slot_b * SEG_BYTES + q * 512 ) );
                        }
                        guard[q].see( v );
                        v.a = __fmul2_rn( v.a, make_float2( weight, weight ) );
                        acc[q] = v;
                    }
                }
            }

            /* ---- the layers that cover this piece ---- */
            bool fetch = (cur.flags & RECT_FIRST_COPY) != 0;
            uint32_t todo = cur.mask;
            const unsigned char *slot = mine_px + (size_t) nb * SEG_BYTES;
            for( int k = 0; k < n && todo; k++, slot += SEG_BYTES, todo >>= 1 ) {
                if( !(todo & 1u) )
                    continue;
                P2 b[PPL];
                if( OPS && J0.layer[k].op_flags ) {
                    /* the layer's fused filters, on its pixels as they leave the ring (cv_layer_ops) */
                    #pragma unroll
                    for( int q = 0; q < PPL; q++ ) {
                        const uint4 raw = lds_u4( slot + q * 512 );
                        const int x = cur.x0 + 2 * ( q * 32 + lane );
                        b[q] = layer_ops_pair( J0, J0.layer[k], raw, x, y );
                    }
                }
                else {
                    #pragma unroll
                    for( int q = 0; q < PPL; q++ )
                        b[q] = unpack_pair( lds_u4( slot + q * 512 ) );
                }
                if( fetch ) {           /* nothing below: the layer as it is (mix 1) */
                    #pragma unroll
                    for( int q = 0; q < PPL; q++ ) {
                        guard[q].see( b[q] );
                        acc[q] = b[q];
                    }
                    fetch = false;
                }
                else {
                    #pragma unroll
                    for( int q = 0; q < PPL; q++ ) {
                        guard[q].see( b[q] );
                        over_pair( acc[q], b[q], 1.0f, one );
                    }
                }
            }

            /* ---- store ---- */
            bool any_failed = false;
            #pragma unroll
            for( int q = 0; q < PPL; q++ )
                any_failed = any_failed || guard[q].failed();

            if( whole && !any_failed ) {
                #pragma unroll
                for( int q = 0; q < PPL; q++ ) {
                    const uint2 h0 = pack_px_guarded( acc[q].r.x, acc[q].g.x, acc[q].b.x, acc[q].a.x );
                    const uint2 h1 = pack_px_guarded( acc[q].r.y, acc[q].g.y, acc[q].b.y, acc[q].a.y );
                    st_stream_u4( dst16 + q * 512, make_uint4( h0.x, h0.y, h1.x, h1.y ) );
                }
                continue;
            }

            SegOut out;
            out.out16 = J.out16; out.out8 = J.out8;
            #pragma unroll
            for( int q = 0; q < PPL; q++ ) {
                const uint32_t pair = (uint32_t) q * 32u + (uint32_t) lane;
                if( pair >= cur.pairs )
                    continue;
                const int x = cur.x0 + (int) pair * 2;
                uint2 h0, h1;
                if( guard[q].failed() ) {
                    h0 = pack_px( pixel_exact<SHAPE>( J, x, y, slow_lut ) );
                    h1 = pack_px( pixel_exact<SHAPE>( J, x + 1, y, slow_lut ) );
                }
                else {
                    h0 = pack_px_guarded( acc[q].r.x, acc[q].g.x, acc[q].b.x, acc[q].a.x );
                    h1 = pack_px_guarded( acc[q].r.y, acc[q].g.y, acc[q].b.y, acc[q].a.y );
                }
                seg_store_halves( J0, out, u8s, x, y, true, true, h0, h1 );
            }
        }
    }
}


/* ------------------------------------------------------------------ launcher */

static int env_int( const char *name, int fallback ) {
    const char *s = getenv( name );
    return s ? atoi( s ) : fallback;
}

/* windowed jobs whose layer mixes are 1: the rectangle walk (k_rect) instead of the window-classifying mode of k_seg */
extern "C" int cvi_rect_enabled( void ) {
    static int disabled = -1;
    if( disabled < 0 ) disabled = env_int( "CV_NO_RECT", 0 );
    return !disabled;
}

extern "C" int cvi_rect_ppl( const DevJob *job ) {
    static int forced = -1;
    if( forced < 0 ) forced = env_int( "CV_RECT_PPL", 0 );
    if( job->flags & JOB_FLAG_OPS ) return 2;
    if( forced == 2 || forced == 3 ) return forced;
    int slots = job->n_layers;
    if( job->bottom_mode != DEV_BOTTOM_NONE && job->a.kind == CV_SRC_F16 ) slots++;
    if( job->bottom_mode == DEV_BOTTOM_CROSS && job->b.kind == CV_SRC_F16 ) slots++;
    return slots <= 6 ? 3 : 2;
}

/* Rows per unit: as many as keep every warp of the machine busy with several units (a unit is decoded once, its rows
 * cost only pitch additions), at most 8.  `row_pieces` = rows x pieces per row x frames of the launch. */
extern "C" int cvi_rect_rows( const DevJob *job, int ppl, int count ) {
    static int forced = -1;
    if( forced < 0 ) forced = env_int( "CV_RECT_ROWS", 0 );
    if( forced >= 1 && forced <= 64 ) return forced;
    CvDevice *dev = cvi_device();
    const uint64_t warps = (uint64_t)( dev ? dev->sm_count : 148 ) * (CV_RECT_MAX_THREADS / 32);
    const uint64_t per_row = (job->items_per_row + 32u * (uint32_t) ppl - 1) / (32u * (uint32_t) ppl);
    const uint64_t row_pieces = per_row * job->rows * (uint64_t)( count > 0 ? count : 1 );
    uint64_t rows = row_pieces / (warps * 6);
    return rows < 1 ? 1 : (rows > 8 ? 8 : (int) rows);
}

template<int SHAPE, int PPL, bool OPS>
static int launch_rect_ppl( CvDevice *dev, const DevJob *job, const DevJob *dev_jobs, int count, const RectPlan *plan,
                            const RectPiece *dev_pieces, cudaStream_t stream ) {
    static int attr_done[CV_MAX_DEVICES];
    const size_t limit = 227 * 1024 - 1024;
    if( !attr_done[dev->index] ) {
        int rc = cvi_cuda( cudaFuncSetAttribute( k_rect<SHAPE, PPL, OPS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) limit ), "cudaFuncSetAttribute(k_rect)" );
        if( rc ) return rc;
        attr_done[dev->index] = 1;
    }

    RectArgs A;
    memset( &A, 0, sizeof(A) );
    if( job->bottom_mode != DEV_BOTTOM_NONE && job->a.kind == CV_SRC_F16 ) A.n_bottom++;
    if( job->bottom_mode == DEV_BOTTOM_CROSS && job->b.kind == CV_SRC_F16 ) A.n_bottom++;
    A.slots = A.n_bottom + job->n_layers;
    A.lut_bytes = SHAPE == KB_CODED ? 65536u : 0u;         /* the compact conversion table (cv_runtime.c) */
    A.u8_bytes = job->out8 ? 65536u : 0u;
    A.stage_bytes = (uint32_t) A.slots * 32u * PPL * 16u;
    if( (uint64_t) plan->units_per_frame * (uint64_t) count > 0x7FFFFFFFu )
        return cvi_fail( CV_ERR_UNSUPPORTED, "k_rect: batch too large" );
    A.all_units = plan->units_per_frame * (uint32_t) count;

    const size_t tables = (size_t) A.lut_bytes + A.u8_bytes;
    static int max_warps = 0, small_warps = 0;          /* per instantiation */
    if( !max_warps ) {
        constexpr int most = ( OPS ? CV_RECT_OPS_MAX_THREADS : CV_RECT_MAX_THREADS ) / 32;
        max_warps = env_int( "CV_RECT_WARPS", most );
        if( max_warps < 1 || max_warps > most ) max_warps = most;
        small_warps = env_int( "CV_RECT_SMALL_WARPS", 4 );
        if( small_warps < 1 || small_warps > max_warps ) small_warps = max_warps < 4 ? max_warps : 4;
    }
    int warps = tables ? max_warps : small_warps;
    size_t smem;
    for( ;; ) {
        smem = tables + (((size_t)( warps + 1 ) * 16 + 127) & ~(size_t) 127) + (size_t) warps * 2 * A.stage_bytes + (size_t) warps * CV_RECT_PARK_BYTES;
        if( smem <= limit || warps == 1 ) break;
        warps--;
    }
    if( smem > limit )
        return cvi_fail( CV_ERR_UNSUPPORTED, "k_rect: %d ring slots do not fit shared memory", A.slots );
    const int threads = warps * 32;

    static __thread int cached_key = -1, cached_blocks = 0;
    const int key = ((dev->index * 64 + A.slots) * 4 + (int)(tables / 65536)) * 32 + warps;
    if( cached_key != key ) {
        int per_sm = 0;
        if( cudaOccupancyMaxActiveBlocksPerMultiprocessor( &per_sm, k_rect<SHAPE, PPL, OPS>, threads, smem ) != cudaSuccess || per_sm < 1 ) {
            cudaGetLastError();
            per_sm = 1;
        }
        const int cap = env_int( "CV_RECT_CTAS_PER_SM", 0 );
        if( cap > 0 && cap < per_sm ) per_sm = cap;
        cached_blocks = per_sm * dev->sm_count;
        cached_key = key;
    }
    uint64_t blocks = ((uint64_t) A.all_units + warps - 1) / warps;
    if( blocks > (uint64_t) cached_blocks ) blocks = (uint64_t) cached_blocks;

    const uint16_t *full_table = reinterpret_cast<const uint16_t *>( dev->lut[CV_LUT_709_TO_LINEAR_SCENE] );
    k_rect<SHAPE, PPL, OPS><<<(unsigned) blocks, threads, smem, stream>>>( *job, count > 1 ? dev_jobs : NULL, *plan, dev_pieces, A,
        reinterpret_cast<const uint16_t *>( dev->lut709_compact ), full_table );
    cvi_count_launch( 1 );
    return cvi_cuda( cudaGetLastError(), "k_rect launch" );
}

template<int SHAPE>
static int launch_rect( CvDevice *dev, int ppl, const DevJob *job, const DevJob *dev_jobs, int count, const RectPlan *plan,
                        const RectPiece *dev_pieces, cudaStream_t stream ) {
    if( job->flags & JOB_FLAG_OPS )     /* layers with fused filters: the two-pair instantiation that evaluates them */
        return launch_rect_ppl<SHAPE, 2, true>( dev, job, dev_jobs, count, plan, dev_pieces, stream );
    if( ppl == 3 ) return launch_rect_ppl<SHAPE, 3, false>( dev, job, dev_jobs, count, plan, dev_pieces, stream );
    return launch_rect_ppl<SHAPE, 2, false>( dev, job, dev_jobs, count, plan, dev_pieces, stream );
}

extern "C" int cvi_launch_rect( int shape, int ppl, const DevJob *job, const DevJob *dev_jobs, int count, const RectPlan *plan,
                                const RectPiece *dev_pieces, cudaStream_t stream ) {
    CvDevice *dev = cvi_device();
    if( !dev ) return CV_ERR_NOT_INITIALISED;
    switch( shape ) {
        case KB_LAYERS:  return launch_rect<KB_LAYERS>( dev, ppl, job, dev_jobs, count, plan, dev_pieces, stream );
        case KB_GENERAL: return launch_rect<KB_GENERAL>( dev, ppl, job, dev_jobs, count, plan, dev_pieces, stream );
        case KB_CODED:   return launch_rect<KB_CODED>( dev, ppl, job, dev_jobs, count, plan, dev_pieces, stream );
    }
    return cvi_fail( CV_ERR_ARGUMENT, "cvi_launch_rect: unknown shape %d", shape );
}
