/*
 * canvas_b200/csrc/cv_scale_fused.cu -- both passes of video_scale_bilinear_f32 in ONE kernel.
 *
 * /root/reference/src/cprocess/video_scale.c:247-284 resamples the axis with the smaller factor into a temporary
 * frame and the other axis from that temporary.  As two launches (k_scale_axis, cv_kernels.cu) the temporary costs
 * a write and a read of HBM that the algorithm does not need.  Here a block produces a TX x TY tile of the target (TY = 16
 * rows for enlargements, whose slices of the temporary are small, 8 otherwise -- measured per geometry, cv_scale.c):
 *   1. the tile's slices of both tap tables go to shared memory (two rounds of independent loads instead of a
 *      chain of dependent ones per pixel),
 *   2. exactly the pixels of the temporary that the tile's second-pass taps touch are evaluated into shared memory --
 *      f32, the same 0 + in*c + in*c ... in the same tap order, multiply and add rounded separately (-fmad=false),
 *      so each one has the bits the two-pass form stores; a warp per row of the slice, lanes along x (coalesced),
 *   3. the tile is gathered from there: a thread owns a column and TY/4 rows.
 * Pixels of the temporary that neighbouring tiles share are evaluated by both (the overlap comes out of L2).
 *
 * The host (cv_scale.c) plans both axes as before and checks, per tile, that the slice of the temporary and both
 * table slices fit (CV_SCALE_FUSED_MAX_TEMP / _MAX_TAPS) and that a tile's second-pass taps lie between its first
 * and its last one; otherwise it keeps the two-pass path.
 */
#include "cv_internal.h"

#define FUSED_THREADS 256
#define FUSED_WARPS ( FUSED_THREADS / 32 )

/* behind the temporary in dynamic shared memory: start1[TX + 1 .. 68] start2[68] src1[cap1] coef1[cap1] src2[cap2]
 * coef2[cap2], cap = the most taps any tile has in that pass (a multiple of 4, from the host) */
#define FUSED_START_WORDS 68
struct FusedTables {
    int32_t *start1, *start2, *src1, *src2;
    float *coef1, *coef2;
    __device__ __forceinline__ FusedTables( void *base, uint32_t cap1, uint32_t cap2 ) {
        int32_t *w = reinterpret_cast<int32_t *>( base );
        start1 = w; start2 = w + FUSED_START_WORDS;
        src1 = w + 2 * FUSED_START_WORDS; coef1 = reinterpret_cast<float *>( src1 + cap1 );
        src2 = src1 + 2 * cap1; coef2 = reinterpret_cast<float *>( src2 + cap2 );
    }
};
static size_t fused_table_bytes( uint32_t cap1, uint32_t cap2 ) { return sizeof(int32_t) * ( 2 * FUSED_START_WORDS + 2 * (size_t) cap1 + 2 * (size_t) cap2 ); }

/* acc = acc + in*k per channel, product and sum rounded separately, two channels per instruction (FMUL2, then the
 * sum as fma(p, ONE, acc) with ONE = 1.0f from the kernel arguments: ptxas would contract a packed mul + add into one
 * FFMA2 even under -fmad=false, see add2_unfused in cv_device.cuh) */
struct Acc { float2 lo, hi; };
__device__ __forceinline__ void tap_accumulate( Acc &acc, const float4 in, const float k, const float2 one ) {
    const float2 kk = make_float2( k, k );
    acc.lo = __ffma2_rn( __fmul2_rn( make_float2( in.x, in.y ), kk ), one, acc.lo );
    acc.hi = __ffma2_rn( __fmul2_rn( make_float2( in.z, in.w ), kk ), one, acc.hi );
}
__device__ __forceinline__ float4 acc_pixel( const Acc &acc ) { return make_float4( acc.lo.x, acc.lo.y, acc.hi.x, acc.hi.y ); }
__device__ __forceinline__ Acc acc_zero() { Acc a; a.lo = make_float2( 0.0f, 0.0f ); a.hi = make_float2( 0.0f, 0.0f ); return a; }

/* first pass = vertical (X_FIRST == 0): temporary( xs, yt ) over source columns and target rows
 * first pass = horizontal (X_FIRST == 1): temporary( xt, ys ) over target columns and source rows */
template<int X_FIRST, int TY>
__global__ void __launch_bounds__( FUSED_THREADS, 8 ) k_scale_fused( const __grid_constant__ FusedScaleArgs A ) {
    extern __shared__ float4 temp[];
    constexpr int TX = CV_SCALE_FUSED_TX, ROWS = TY / ( FUSED_THREADS / TX );
    const FusedTables T( temp + A.temp_pixels, A.cap1, A.cap2 );

    const int c0 = (int) blockIdx.x * TX, r0 = (int) blockIdx.y * TY;
    const int nc = min( TX, (int) A.t_width - c0 ), nr = min( TY, (int) A.t_height - r0 );
    const int tid = (int) threadIdx.x, warp = tid >> 5, lane = tid & 31;

    /* ---- 1. table slices.  Pass 2 is indexed by the tile's lines along its axis; pass 1 by the lines of the
     *         temporary along ITS axis, which are the tile's lines along the other axis, shifted by origin1 */
    const int line2_0 = X_FIRST ? r0 : c0, n2 = X_FIRST ? nr : nc;
    const int want1_0 = ( X_FIRST ? A.t_ox + c0 : A.t_oy + r0 ) - A.origin1, want1_n = X_FIRST ? nc : nr;
    const int line1_0 = max( want1_0, 0 ), n1 = max( min( want1_0 + want1_n, A.lines1 ) - line1_0, 0 );

    if( tid <= n2 ) T.start2[tid] = __ldg( A.start2 + line2_0 + tid );
    else if( n1 > 0 && tid >= 128 && tid - 128 <= n1 ) T.start1[tid - 128] = __ldg( A.start1 + line1_0 + tid - 128 );
    __syncthreads();
    const int base1 = n1 ? T.start1[0] : 0, count1 = n1 ? T.start1[n1] - base1 : 0;
    const int base2 = T.start2[0], count2 = T.start2[n2] - base2;
#pragma unroll 1
    for( int i = tid; i < count1; i += FUSED_THREADS ) { T.src1[i] = __ldg( A.src1 + base1 + i ); T.coef1[i] = __ldg( A.coef1 + base1 + i ); }
#pragma unroll 1
    for( int i = tid; i < count2; i += FUSED_THREADS ) { T.src2[i] = __ldg( A.src2 + base2 + i ); T.coef2[i] = __ldg( A.coef2 + base2 + i ); }
    __syncthreads();
    const float2 one = make_float2( A.one, A.one );

    /* the slice of the temporary: from the tile's first second-pass tap to its last (checked by the host) */
    const int lo = count2 > 0 ? T.src2[0] : 0;
    const int span = count2 > 0 ? T.src2[count2 - 1] - lo + 1 : 0;

    /* ---- 2. the temporary */
    if( !X_FIRST ) {
        /* rows = the tile's target rows (a warp each), columns lo .. lo + span - 1 of the source */
        /* the slice as one run of nr * span pixels over all threads (a narrow slice -- enlargement -- would leave most
         * lanes of a warp-per-row walk idle); (r, c) advance by the block's stride without dividing again */
        if( span > 0 ) {
            const int step_r = FUSED_THREADS / span, step_c = FUSED_THREADS - step_r * span;
            int r = tid / span, c = tid - r * span;
            const float4 *origin = A.source - ( (ptrdiff_t) A.s_ox + (ptrdiff_t) A.s_oy * A.s_pitch );
#pragma unroll 1
            for( ; r < nr; r += step_r, c += step_c ) {
                if( c >= span ) { c -= span; if( ++r >= nr ) break; }
                const int line = want1_0 + r, xs = lo + c;
                Acc acc = acc_zero();
                if( line >= 0 && line < A.lines1 && xs >= A.cross1_lo && xs <= A.cross1_hi ) {
                    const int first = T.start1[line - line1_0] - base1, last = T.start1[line - line1_0 + 1] - base1;
                    const float4 *column = origin + xs;
#pragma unroll 2
                    for( int tap = first; tap < last; tap++ )
                        tap_accumulate( acc, __ldg( column + (ptrdiff_t) T.src1[tap] * A.s_pitch ), T.coef1[tap], one );
                }
                temp[r * span + c] = acc_pixel( acc );
            }
        }
    }
    else {
        /* columns = the tile's target columns (lanes), rows lo .. lo + span - 1 of the source (a warp each) */
        for( int r = warp; r < span; r += FUSED_WARPS ) {
            const int ys = lo + r;
            const bool inside = ys >= A.cross1_lo && ys <= A.cross1_hi;
            const float4 *row = A.source + ( (ptrdiff_t)( ys - A.s_oy ) * A.s_pitch - A.s_ox );
            for( int c = lane; c < nc; c += 32 ) {
                const int line = want1_0 + c;
                Acc acc = acc_zero();
                if( inside && line >= 0 && line < A.lines1 ) {
                    const int first = T.start1[line - line1_0] - base1, last = T.start1[line - line1_0 + 1] - base1;
#pragma unroll 2
                    for( int tap = first; tap < last; tap++ )
                        tap_accumulate( acc, __ldg( row + T.src1[tap] ), T.coef1[tap], one );
                }
                temp[r * TX + c] = acc_pixel( acc );
            }
        }
    }
    __syncthreads();

    /* ---- 3. the tile out of shared memory: a thread owns column c and rows rr, rr + 4, ...; everything is written
     *         (zero where no tap reaches) */
    const int c = tid % TX, rr = tid / TX;
    if( c >= nc ) return;
    const int x = A.t_ox + c0 + c;
    Acc acc[ROWS];
#pragma unroll
    for( int k = 0; k < ROWS; k++ ) acc[k] = acc_zero();
    float4 *out = A.target + ( (size_t)( r0 + rr ) * (size_t) A.t_pitch + (size_t)( c0 + c ) );
    const size_t out_step = (size_t) A.t_pitch * ( FUSED_THREADS / TX );

    if( !X_FIRST ) {
        /* taps belong to the column: each one is read once and applied to all of the thread's rows (rows beyond the
         * tile read a pixel of the slice that exists -- row rr -- and are not stored) */
        const int first = T.start2[c] - base2, last = T.start2[c + 1] - base2;
        const float4 *mine = temp + rr * span - lo;
        const int row_step = span * ( FUSED_THREADS / TX );
#pragma unroll 1
        for( int tap = first; tap < last; tap++ ) {
            const float4 *p = mine + T.src2[tap];
            const float k2 = T.coef2[tap];
#pragma unroll
            for( int k = 0; k < ROWS; k++ )
                tap_accumulate( acc[k], p[rr + k * ( FUSED_THREADS / TX ) < nr ? k * row_step : 0], k2, one );
        }
#pragma unroll
        for( int k = 0; k < ROWS; k++ ) {
            const int r = rr + k * ( FUSED_THREADS / TX ), y = A.t_oy + r0 + r;
            if( r >= nr ) continue;
            const bool inside = y >= A.cross2_lo && y <= A.cross2_hi;
            out[k * out_step] = inside ? acc_pixel( acc[k] ) : make_float4( 0.0f, 0.0f, 0.0f, 0.0f );
        }
    }
    else {
        const bool inside = x >= A.cross2_lo && x <= A.cross2_hi;
        const float4 *mine = temp + c - lo * TX;
#pragma unroll
        for( int k = 0; k < ROWS; k++ ) {
            const int r = rr + k * ( FUSED_THREADS / TX );
            if( r >= nr ) continue;
            if( inside ) {
                const int first = T.start2[r] - base2, last = T.start2[r + 1] - base2;
#pragma unroll 2
                for( int tap = first; tap < last; tap++ )
                    tap_accumulate( acc[k], mine[T.src2[tap] * TX], T.coef2[tap], one );
            }
            out[k * out_step] = acc_pixel( acc[k] );
        }
    }
}

extern "C" int cvi_launch_scale_fused( const FusedScaleArgs *args, cudaStream_t stream ) {
    CvDevice *dev = cvi_device();
    if( !dev ) return CV_ERR_NOT_INITIALISED;
    if( !args->t_width || !args->t_height ) return CV_OK;
    static_assert( CV_SCALE_FUSED_TX >= CV_SCALE_FUSED_TY && CV_SCALE_FUSED_TX + 1 <= 128 && FUSED_THREADS % CV_SCALE_FUSED_TX == 0
                   && CV_SCALE_FUSED_TY % ( FUSED_THREADS / CV_SCALE_FUSED_TX ) == 0 && CV_SCALE_FUSED_TY_SHORT % ( FUSED_THREADS / CV_SCALE_FUSED_TX ) == 0, "tile shape" );
    if( args->temp_pixels > CV_SCALE_FUSED_MAX_TEMP || args->cap1 > CV_SCALE_FUSED_MAX_TAPS || args->cap2 > CV_SCALE_FUSED_MAX_TAPS || ( args->cap1 | args->cap2 ) % 4
            || ( args->tile_rows != CV_SCALE_FUSED_TY && args->tile_rows != CV_SCALE_FUSED_TY_SHORT ) )
        return cvi_fail( CV_ERR_ARGUMENT, "k_scale_fused: %u pixels of the temporary per tile (at most %u), %d rows", args->temp_pixels, (unsigned) CV_SCALE_FUSED_MAX_TEMP, args->tile_rows );

    static int configured[CV_MAX_DEVICES];          /* more than 48 KB of dynamic shared memory is opt-in, per device */
    const size_t most = (size_t) CV_SCALE_FUSED_MAX_TEMP * sizeof(float4) + fused_table_bytes( CV_SCALE_FUSED_MAX_TAPS, CV_SCALE_FUSED_MAX_TAPS );
    if( !configured[dev->index] ) {
        int rc = cvi_cuda( cudaFuncSetAttribute( k_scale_fused<0, CV_SCALE_FUSED_TY>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) most ), "cudaFuncSetAttribute(k_scale_fused)" );
        if( !rc ) rc = cvi_cuda( cudaFuncSetAttribute( k_scale_fused<1, CV_SCALE_FUSED_TY>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) most ), "cudaFuncSetAttribute(k_scale_fused)" );
        if( !rc ) rc = cvi_cuda( cudaFuncSetAttribute( k_scale_fused<0, CV_SCALE_FUSED_TY_SHORT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) most ), "cudaFuncSetAttribute(k_scale_fused)" );
        if( !rc ) rc = cvi_cuda( cudaFuncSetAttribute( k_scale_fused<1, CV_SCALE_FUSED_TY_SHORT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) most ), "cudaFuncSetAttribute(k_scale_fused)" );
        if( rc ) return rc;
        configured[dev->index] = 1;
    }

    const dim3 grid( ( args->t_width + CV_SCALE_FUSED_TX - 1 ) / CV_SCALE_FUSED_TX, ( args->t_height + args->tile_rows - 1 ) / args->tile_rows );
    const size_t shared = (size_t) args->temp_pixels * sizeof(float4) + fused_table_bytes( args->cap1, args->cap2 );
    if( args->tile_rows == CV_SCALE_FUSED_TY ) {
        if( args->x_first ) k_scale_fused<1, CV_SCALE_FUSED_TY><<<grid, FUSED_THREADS, shared, stream>>>( *args );
        else k_scale_fused<0, CV_SCALE_FUSED_TY><<<grid, FUSED_THREADS, shared, stream>>>( *args );
    }
    else {
        if( args->x_first ) k_scale_fused<1, CV_SCALE_FUSED_TY_SHORT><<<grid, FUSED_THREADS, shared, stream>>>( *args );
        else k_scale_fused<0, CV_SCALE_FUSED_TY_SHORT><<<grid, FUSED_THREADS, shared, stream>>>( *args );
    }
    cvi_count_launch( 1 );
    return cvi_cuda( cudaGetLastError(), "k_scale_fused launch" );
}
