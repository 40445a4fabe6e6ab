/*
 * oracle/oracle.c -- TEST INFRASTRUCTURE ONLY.  See oracle.h for the rules.
 *
 * CPU restatement of the reference algorithm for the per-pixel path.  Written from
 * the behaviour of the reference files cited per function (paths relative to
 * /root/reference); compiled with -O2 -ffp-contract=off so that every f32
 * intermediate is rounded exactly once per C operation, which is what the CUDA
 * kernels reproduce with __fmul_rn/__fadd_rn/__fdiv_rn.
 */
#include "oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <stdint.h>

/* ---- small box helpers (framework.h:80-137) -------------------------------- */
static inline int imin( int a, int b ) { return a < b ? a : b; }
static inline int imax( int a, int b ) { return a > b ? a : b; }
static inline int box_empty( const orc_box *b ) { return b->max_x < b->min_x || b->max_y < b->min_y; }
static inline void box_set_empty( orc_box *b ) { b->min_x = 0; b->min_y = 0; b->max_x = -1; b->max_y = -1; }
static inline orc_box box_intersect( const orc_box *a, const orc_box *b ) {
    orc_box r = { imax( a->min_x, b->min_x ), imax( a->min_y, b->min_y ), imin( a->max_x, b->max_x ), imin( a->max_y, b->max_y ) };
    return r;
}
static inline orc_box box_union( const orc_box *a, const orc_box *b ) {
    orc_box r = { imin( a->min_x, b->min_x ), imin( a->min_y, b->min_y ), imax( a->max_x, b->max_x ), imax( a->max_y, b->max_y ) };
    return r;
}
static inline int box_contains( const orc_box *b, int x, int y ) {
    return x >= b->min_x && x <= b->max_x && y >= b->min_y && y <= b->max_y;
}
static inline int box_width( const orc_box *b ) { return b->max_x < b->min_x ? 0 : b->max_x - b->min_x + 1; }
static inline int box_height( const orc_box *b ) { return b->max_y < b->min_y ? 0 : b->max_y - b->min_y + 1; }
static inline size_t box_area( const orc_box *b ) { return (size_t) box_width( b ) * (size_t) box_height( b ); }

static inline orc_px16 *px16( const orc_frame16 *f, int x, int y ) {
    return &f->data[(size_t)(y - f->full.min_y) * (size_t)(f->full.max_x - f->full.min_x + 1) + (x - f->full.min_x)];
}
static inline orc_px32 *px32( const orc_frame32 *f, int x, int y ) {
    return &f->data[(size_t)(y - f->full.min_y) * (size_t)(f->full.max_x - f->full.min_x + 1) + (x - f->full.min_x)];
}

static inline float clampf( float v, float lo, float hi ) {
    /* framework.h:127-137: minf(maxf(v, lo), hi) with `a < b ? a : b` forms (NaN -> lo) */
    float m = v > lo ? v : lo;
    return m < hi ? m : hi;
}

/* ---- half <-> float --------------------------------------------------------
 * half.c:47-51 with genhalf.py:25-57: out = base[e9] + (mantissa >> shift[e9]).
 * That is truncation (round toward zero) with subnormals kept; finite values of
 * magnitude >= 65536 become infinity; NaN payloads are truncated to their top ten
 * bits (and so may become infinity).
 */
uint16_t orc_f2h( float value ) {
    union { float f; uint32_t i; } u = { .f = value };
    uint32_t sign = (u.i >> 16) & 0x8000u;
    int e = (int)((u.i >> 23) & 0xFF) - 127;
    uint32_t mant = u.i & 0x007FFFFFu;

    if( e < -24 )
        return (uint16_t) sign;
    if( e < -14 )
        return (uint16_t)(sign | ((0x0400u >> (-14 - e)) + (mant >> (-e - 1))));
    if( e <= 15 )
        return (uint16_t)(sign | (((uint32_t)(e + 15) << 10) + (mant >> 13)));
    if( e < 128 )
        return (uint16_t)(sign | 0x7C00u);
    return (uint16_t)(sign | (0x7C00u + (mant >> 13)));
}

/* half.c:31-37 with genhalf.py:59-94: exact widening, subnormals normalised. */
float orc_h2f( uint16_t value ) {
    union { float f; uint32_t i; } u;
    uint32_t sign = ((uint32_t) value & 0x8000u) << 16;
    uint32_t e = (value >> 10) & 0x1F;
    uint32_t m = value & 0x3FF;

    if( e == 0 ) {
        if( m == 0 ) {
            u.i = sign;
        }
        else {
            int shift = 0;
            while( !(m & 0x400) ) {
                m <<= 1;
                shift++;
            }
            u.i = sign | ((uint32_t)(113 - shift) << 23) | ((m & 0x3FF) << 13);
        }
    }
    else if( e == 31 ) {
        u.i = sign | 0x7F800000u | (m << 13);
    }
    else {
        u.i = sign | ((e + 112) << 23) | (m << 13);
    }

    return u.f;
}

void orc_f2h_n( uint16_t *out, const float *in, size_t count ) {
    for( size_t i = 0; i < count; i++ )
        out[i] = orc_f2h( in[i] );
}

void orc_h2f_n( float *out, const uint16_t *in, size_t count ) {
    for( size_t i = 0; i < count; i++ )
        out[i] = orc_h2f( in[i] );
}

/* ---- transfer tables --------------------------------------------------------- */
/* gammatab.c:48-55 */
static float rec709_to_linear( float in ) {
    const float transition = 4.5f * 0.018f;
    if( in < transition )
        return in / 4.5f;
    return powf( (in + 0.099f) / 1.099f, 1.0f / 0.45f );
}

/* gammatab.c:58-66 */
static float linear_to_rec709( float in ) {
    const float transition = 0.018f;
    if( in < transition )
        return in * 4.5f;
    return 1.099f * powf( in, 0.45f ) - 0.099f;
}

/* gammatab.c:201-211 (note: `a` is a float initialised from the double 0.055) */
static float linear_to_srgb( float in ) {
    const float transition = 0.0031308f;
    const float a = 0.055;
    if( in <= transition )
        return in * 12.92f;
    return (1.0f + a) * powf( in, 1.0f / 2.4f ) - a;
}

void orc_build_lut( int which, uint16_t *table ) {
    for( int i = 0; i < 65536; i++ ) {
        float f = orc_h2f( (uint16_t) i );
        switch( which ) {
            case ORC_LUT_709_TO_LINEAR_SCENE:   f = rec709_to_linear( f ); break;
            case ORC_LUT_709_TO_LINEAR_DISPLAY: f = (f < 0.0f) ? 0.0f : powf( f, 2.5f ); break;  /* gammatab.c:145-150 */
            case ORC_LUT_LINEAR_TO_709:         f = linear_to_rec709( f ); break;
            default:                            f = linear_to_srgb( f ); break;
        }
        table[i] = orc_f2h( f );
    }
}

/* C's (uint8_t) of an out-of-range float is undefined; gcc/x86 converts through a
 * 32-bit cvttss2si and keeps the low byte.  clampf() already bounds these callers. */
static inline uint8_t float_to_u8_wrap( float f ) {
    if( !(fabsf( f ) < 2147483648.0f) )
        return 0;       /* x86 "integer indefinite" 0x80000000 -> low byte 0 */
    return (uint8_t)(int32_t) f;
}

/* gammatab.c:8-38 */
void orc_build_gamma45( uint8_t *ramp ) {
    for( int i = 0; i < 65536; i++ )
        ramp[i] = float_to_u8_wrap( clampf( powf( orc_h2f( (uint16_t) i ), 0.45f ) * 255.0f, 0.0f, 255.0f ) );
}

/* widget_gl.c:949-969 */
void orc_build_intent_ramp( float rendering_intent, uint8_t *ramp ) {
    for( int i = 0; i < 65536; i++ )
        ramp[i] = (uint8_t) lrintf( clampf( powf( orc_h2f( (uint16_t) i ), rendering_intent ) * 255.0f, 0.0f, 255.0f ) );
}

/* ---- FIR triangle: filter.c:24-76 -------------------------------------------- */
int orc_triangle( float sub, float offset, float *coeff, int cap, int *width_out, int *center_out ) {
    const int down = sub < 1.0f;
    const float width = down ? (1.0f / sub) : sub;

    float left_edge = ceilf( offset - width );
    float right_edge = floorf( offset + width );

    if( left_edge == offset - width )
        left_edge++;
    if( right_edge == offset + width )
        right_edge--;

    const int full_width = (int) right_edge - (int) left_edge + 1;
    *width_out = full_width;

    if( cap < full_width ) {
        *center_out = -1;
        return -1;
    }

    const int center = -(int) left_edge;
    *center_out = center;

    float sum = 0.0f;
    for( int i = 0; i < full_width; i++ ) {
        coeff[i] = 1.0f - fabsf( (1.0f / width) * ((i - center) - offset) );
        sum += coeff[i];
    }

    if( sub < 1.0f && sum != 0.0f ) {
        for( int i = 0; i < full_width; i++ )
            coeff[i] /= sum;
    }

    return 0;
}

/* ---- coded picture access ------------------------------------------------------ */
static inline uint8_t coded_luma( const orc_coded *im, int x, int row ) {
    switch( im->layout ) {
        case ORC_LAYOUT_UYVY: return im->plane[0][(size_t) row * im->stride[0] + 2 * x + 1];
        case ORC_LAYOUT_YUY2: return im->plane[0][(size_t) row * im->stride[0] + 2 * x];
        default:              return im->plane[0][(size_t) row * im->stride[0] + x];
    }
}

/* Packed layouts are 4:2:2 by definition: byte groups Cb Y0 Cr Y1 (UYVY) / Y0 Cb Y1 Cr (YUY2). */
static inline uint8_t coded_chroma( const orc_coded *im, int which /*0 cb, 1 cr*/, int cx, int row ) {
    switch( im->layout ) {
        case ORC_LAYOUT_UYVY: return im->plane[0][(size_t) row * im->stride[0] + 4 * cx + (which ? 2 : 0)];
        case ORC_LAYOUT_YUY2: return im->plane[0][(size_t) row * im->stride[0] + 4 * cx + (which ? 3 : 1)];
        default:              return im->plane[1 + which][(size_t) row * im->stride[1 + which] + cx];
    }
}

static inline uint8_t *coded_luma_ptr( const orc_coded *im, int x, int row ) {
    switch( im->layout ) {
        case ORC_LAYOUT_UYVY: return &im->plane[0][(size_t) row * im->stride[0] + 2 * x + 1];
        case ORC_LAYOUT_YUY2: return &im->plane[0][(size_t) row * im->stride[0] + 2 * x];
        default:              return &im->plane[0][(size_t) row * im->stride[0] + x];
    }
}

static inline uint8_t *coded_chroma_ptr( const orc_coded *im, int which, int cx, int row ) {
    switch( im->layout ) {
        case ORC_LAYOUT_UYVY: return &im->plane[0][(size_t) row * im->stride[0] + 4 * cx + (which ? 2 : 0)];
        case ORC_LAYOUT_YUY2: return &im->plane[0][(size_t) row * im->stride[0] + 4 * cx + (which ? 3 : 1)];
        default:              return &im->plane[1 + which][(size_t) row * im->stride[1 + which] + cx];
    }
}

/* video_reconstruct.c:54-66 */
static const float ycbcr_to_rgb[2][3][3] = {
    { { 1.0f,  0.0f,       1.5748f },       /* Rec. 709 */
      { 1.0f, -0.187324f, -0.468124f },
      { 1.0f,  1.8556f,    0.0f } },
    { { 1.0f,  0.0f,       1.402f },        /* Rec. 601 (the commented-out table :56-58) */
      { 1.0f, -0.344136f, -0.714136f },
      { 1.0f,  1.772f,     0.0f } },
};

/* video_subsample.c:104-108; Rec. 601 from the MPEG-2 shader's matrix :406-410 is not a
 * CPU path in the reference: the 601 forward table here is Poynton's, documented in DESIGN.md. */
static const float rgb_to_ycbcr[2][3][3] = {
    { {  0.2126f,    0.7152f,    0.0722f   },
      { -0.114572f, -0.385428f,  0.5f      },
      {  0.5f,      -0.454153f, -0.045847f } },
    { {  0.299f,     0.587f,     0.114f    },
      { -0.168736f, -0.331264f,  0.5f      },
      {  0.5f,      -0.418688f, -0.081312f } },
};

/* ---- video_reconstruct_dv, parameterised: video_reconstruct.c:50-137 ---------- */
void orc_reconstruct( orc_frame16 *frame, const orc_coded *im ) {
    static uint16_t *lut = NULL;
    if( !lut ) {
        uint16_t *t = malloc( sizeof(uint16_t) * 65536 );
        orc_build_lut( ORC_LUT_709_TO_LINEAR_SCENE, t );
        lut = t;
    }

    const int W = im->width, H = im->height, sub = im->sub_x;
    const float (*m)[3] = ycbcr_to_rgb[im->matrix];

    /* :69-76 */
    frame->current.min_x = imax( im->pic_x, frame->full.min_x );
    frame->current.min_y = imax( im->pic_y, frame->full.min_y );
    frame->current.max_x = imin( W + im->pic_x - 1, frame->full.max_x );
    frame->current.max_y = imin( H + im->pic_y - 1, frame->full.max_y );

    if( box_empty( &frame->current ) )
        return;

    float coeff[16];
    int fw, center;
    orc_triangle( (float) sub, 0.0f, coeff, 16, &fw, &center );

    float *temp_cb = malloc( sizeof(float) * W ), *temp_cr = malloc( sizeof(float) * W );
    const int lo = frame->current.min_x - im->pic_x, hi = frame->current.max_x - im->pic_x;    /* plane-local */

    for( int row = frame->current.min_y - im->pic_y; row <= frame->current.max_y - im->pic_y; row++ ) {
        memset( temp_cb, 0, sizeof(float) * W );
        memset( temp_cr, 0, sizeof(float) * W );

        /* :98-109 scatter-accumulate in ascending chroma x */
        const int endx = (W - 1) / sub;

        for( int x = 0; x <= endx; x++ ) {
            float cb = (coded_chroma( im, 0, x, row ) - 128.0f) / 224.0f;
            float cr = (coded_chroma( im, 1, x, row ) - 128.0f) / 224.0f;

            for( int i = imax( lo, x * sub - center ); i <= imin( hi, x * sub + (fw - center - 1) ); i++ ) {
                temp_cb[i] += cb * coeff[i - x * sub + center];
                temp_cr[i] += cr * coeff[i - x * sub + center];
            }
        }

        /* :111-131 */
        for( int i = lo; i <= hi; i++ ) {
            float y = (coded_luma( im, i, row ) - 16.0f) / 219.0f;
            float rgba[4];

            rgba[0] = y * m[0][0] + temp_cb[i] * m[0][1] + temp_cr[i] * m[0][2];
            rgba[1] = y * m[1][0] + temp_cb[i] * m[1][1] + temp_cr[i] * m[1][2];
            rgba[2] = y * m[2][0] + temp_cb[i] * m[2][1] + temp_cr[i] * m[2][2];
            rgba[3] = 1.0f;

            orc_px16 *out = px16( frame, i + im->pic_x, row + im->pic_y );
            out->r = lut[orc_f2h( rgba[0] )];
            out->g = lut[orc_f2h( rgba[1] )];
            out->b = lut[orc_f2h( rgba[2] )];
            out->a = lut[orc_f2h( rgba[3] )];
        }
    }

    free( temp_cb );
    free( temp_cr );
}

/* ---- video_subsample_dv, parameterised: video_subsample.c:99-187 --------------- */
void orc_subsample( orc_frame16 *frame, const orc_coded *im ) {
    static uint16_t *lut = NULL;
    if( !lut ) {
        uint16_t *t = malloc( sizeof(uint16_t) * 65536 );
        orc_build_lut( ORC_LUT_LINEAR_TO_709, t );
        lut = t;
    }

    const int W = im->width, H = im->height, sub = im->sub_x;
    const float (*m)[3] = rgb_to_ycbcr[im->matrix];

    /* :122-128 */
    orc_box window = {
        imax( im->pic_x, frame->current.min_x ), imax( im->pic_y, frame->current.min_y ),
        imin( W + im->pic_x - 1, frame->current.max_x ), imin( H + im->pic_y - 1, frame->current.max_y ) };

    if( box_empty( &window ) )
        return;

    const int window_width = window.max_x - window.min_x + 1;

    float coeff[16];
    int fw, center;
    orc_triangle( 1.0f / (float) sub, 0.0f, coeff, 16, &fw, &center );

    float *temp_cb = malloc( sizeof(float) * window_width ), *temp_cr = malloc( sizeof(float) * window_width );

    for( int row = window.min_y - im->pic_y; row <= window.max_y - im->pic_y; row++ ) {
        orc_px16 *in = px16( frame, window.min_x, row + im->pic_y );

        for( int x = 0; x < window_width; x++ ) {
            /* :147 in-place transfer on all four channels of the INPUT frame */
            in[x].r = lut[in[x].r];
            in[x].g = lut[in[x].g];
            in[x].b = lut[in[x].b];
            in[x].a = lut[in[x].a];

            float r = orc_h2f( in[x].r ), g = orc_h2f( in[x].g ), b = orc_h2f( in[x].b );

            float y = (r * m[0][0] + g * m[0][1] + b * m[0][2]) * 219.0f + 16.0f;
            temp_cb[x] = r * m[1][0] + g * m[1][1] + b * m[1][2];
            temp_cr[x] = r * m[2][0] + g * m[2][1] + b * m[2][2];

            /* plane-local luma x = frame x - pic_x (the reference has pic_x == 0, :164) */
            *coded_luma_ptr( im, x + window.min_x - im->pic_x, row ) = float_to_u8_wrap( y );
        }

        /* :167-179; window coordinates are plane-local after removing pic_x */
        const int wmin = window.min_x - im->pic_x, wmax = window.max_x - im->pic_x;

        for( int tx = wmin / sub; tx <= wmax / sub; tx++ ) {
            float cb = 0.0f, cr = 0.0f;

            for( int sx = imax( wmin, tx * sub - center ); sx <= imin( wmax, tx * sub + (fw - center - 1) ); sx++ ) {
                cb += temp_cb[sx - wmin] * coeff[sx - tx * sub + center];
                cr += temp_cr[sx - wmin] * coeff[sx - tx * sub + center];
            }

            *coded_chroma_ptr( im, 0, tx, row ) = float_to_u8_wrap( cb * 224.0f + 128.0f );
            *coded_chroma_ptr( im, 1, tx, row ) = float_to_u8_wrap( cr * 224.0f + 128.0f );
        }
    }

    free( temp_cb );
    free( temp_cr );
}

/* ---- MPEG-2 4:2:0 shaders restated: video_subsample.c:212-290 -------------------- */
typedef struct { float y, pb, pr; } ypbpr;

static ypbpr mpeg2_texel( const orc_frame16 *frame, const orc_coded *im, const float *to709, int px, int py ) {
    /* GL_CLAMP_TO_EDGE on the picture, then the frame pixel the texel maps to (:404-405 input_offset) */
    px = imin( imax( px, 0 ), im->width - 1 ) + im->pic_x;
    py = imin( imax( py, 0 ), im->height - 1 ) + im->pic_y;
    float r = 0.0f, g = 0.0f, b = 0.0f;
    if( box_contains( &frame->current, px, py ) && box_contains( &frame->full, px, py ) ) {
        const orc_px16 *p = px16( frame, px, py );
        r = to709[p->r]; g = to709[p->g]; b = to709[p->b];      /* linear_to_rec709(), :227-232 */
    }
    const float (*m)[3] = rgb_to_ycbcr[im->matrix];             /* color * rgb2yuv, :235, :276-281 */
    ypbpr out = {
        (r * m[0][0] + g * m[0][1]) + b * m[0][2],
        (r * m[1][0] + g * m[1][1]) + b * m[1][2],
        (r * m[2][0] + g * m[2][1]) + b * m[2][2] };
    return out;
}

static uint8_t unorm8( float v ) {
    if( v != v ) return 0;
    return (uint8_t) lrintf( clampf( v, 0.0f, 1.0f ) * 255.0f );
}

void orc_subsample_mpeg2( const orc_frame16 *frame, const orc_coded *im ) {
    static float *to709 = NULL;
    if( !to709 ) {
        float *t = malloc( sizeof(float) * 65536 );
        for( int i = 0; i < 65536; i++ )
            t[i] = linear_to_rec709( orc_h2f( (uint16_t) i ) );
        to709 = t;
    }

    const float ys = 219.0f / 255.0f, yo = 16.0f / 255.0f, cs = 224.0f / 255.0f, co = 128.0f / 255.0f;

    /* luma shader, :234-239 */
    for( int y = 0; y < im->height; y++ )
        for( int x = 0; x < im->width; x++ )
            im->plane[0][(size_t) y * im->stride[0] + x] = unorm8( mpeg2_texel( frame, im, to709, x, y ).y * ys + yo );

    /* chroma shader, :265-286: luma_coord = 2 * chroma coordinate; near/far rows by the parity of the chroma row */
    static const float weight[6] = { 3.0f / 16.0f, 6.0f / 16.0f, 3.0f / 16.0f, 1.0f / 16.0f, 2.0f / 16.0f, 1.0f / 16.0f };

    for( int j = 0; j < im->height / 2; j++ ) {
        const int odd = j & 1;
        const int near_row = 2 * j + (odd ? 1 : 0), far_row = 2 * j + (odd ? -1 : 2);

        for( int i = 0; i < im->width / 2; i++ ) {
            float cb = 0.0f, cr = 0.0f;

            for( int t = 0; t < 6; t++ ) {
                const ypbpr v = mpeg2_texel( frame, im, to709, 2 * i + (t % 3) - 1, t < 3 ? near_row : far_row );
                if( t == 0 ) { cb = weight[0] * v.pb; cr = weight[0] * v.pr; }
                else { cb = cb + weight[t] * v.pb; cr = cr + weight[t] * v.pr; }
            }

            im->plane[1][(size_t) j * im->stride[1] + i] = unorm8( cb * cs + co );
            im->plane[2][(size_t) j * im->stride[2] + i] = unorm8( cr * cs + co );
        }
    }
}

/* ---- color.c:100-160 -------------------------------------------------------------- */
static void color_matrix( orc_frame16 *frame, const float a[3], const float b[3], const float c[3], int pre, int post ) {
    static uint16_t *lut[4] = { NULL, NULL, NULL, NULL };
    const int need[2] = { pre, post };
    for( int i = 0; i < 2; i++ ) {
        if( need[i] >= 0 && !lut[need[i]] ) {
            uint16_t *t = malloc( sizeof(uint16_t) * 65536 );
            orc_build_lut( need[i], t );
            lut[need[i]] = t;
        }
    }

    const orc_box box = box_intersect( &frame->current, &frame->full );

    for( int y = box.min_y; y <= box.max_y; y++ ) {
        for( int x = box.min_x; x <= box.max_x; x++ ) {
            orc_px16 *h = px16( frame, x, y );
            if( pre >= 0 ) { h->r = lut[pre][h->r]; h->g = lut[pre][h->g]; h->b = lut[pre][h->b]; h->a = lut[pre][h->a]; }

            /* mult_mat_xyz (:35-43): a, b, c are the matrix COLUMNS */
            const float r = orc_h2f( h->r ), g = orc_h2f( h->g ), bl = orc_h2f( h->b ), al = orc_h2f( h->a );
            h->r = orc_f2h( r * a[0] + g * b[0] + bl * c[0] );
            h->g = orc_f2h( r * a[1] + g * b[1] + bl * c[1] );
            h->b = orc_f2h( r * a[2] + g * b[2] + bl * c[2] );
            h->a = orc_f2h( al );

            if( post >= 0 ) { h->r = lut[post][h->r]; h->g = lut[post][h->g]; h->b = lut[post][h->b]; h->a = lut[post][h->a]; }
        }
    }
}

void orc_color_rgb_to_xyz_sdtv( orc_frame16 *frame ) {
    static const float r[3] = { 0.3936f, 0.2124f, 0.0187f }, g[3] = { 0.3652f, 0.7010f, 0.1119f }, b[3] = { 0.1916f, 0.0865f, 0.9582f };
    color_matrix( frame, r, g, b, ORC_LUT_709_TO_LINEAR_SCENE, -1 );
}

void orc_color_xyz_to_srgb( orc_frame16 *frame ) {
    static const float r[3] = { 3.2410f, -0.9692f, 0.0556f }, g[3] = { -1.5374f, 1.8760f, -0.2040f }, b[3] = { -0.4986f, 0.0416f, 1.0570f };
    color_matrix( frame, r, g, b, -1, ORC_LUT_LINEAR_TO_SRGB );
}

/* ---- video_filter.c:79-95 (crop shader) -------------------------------------------- */
void orc_crop_f16( orc_frame16 *frame, const orc_box *rect ) {
    const orc_box box = box_intersect( &frame->current, &frame->full );
    for( int y = box.min_y; y <= box.max_y; y++ )
        for( int x = box.min_x; x <= box.max_x; x++ )
            if( !box_contains( rect, x, y ) )
                memset( px16( frame, x, y ), 0, sizeof(orc_px16) );
}

/* ---- copies: video_mix.c:27-44, 73-105 ----------------------------------------- */
void orc_copy_frame_f16( orc_frame16 *out, const orc_frame16 *in ) {
    orc_box inner = box_intersect( &out->full, &in->current );
    out->current = inner;

    if( box_empty( &inner ) )
        return;

    for( int y = inner.min_y; y <= inner.max_y; y++ )
        memcpy( px16( out, inner.min_x, y ), px16( in, inner.min_x, y ), sizeof(orc_px16) * box_width( &inner ) );
}

void orc_copy_frame_alpha_f32( orc_frame32 *out, const orc_frame32 *in, float alpha ) {
    alpha = clampf( alpha, 0.0f, 1.0f );

    if( out == in && alpha == 1.0f )
        return;

    if( alpha == 0.0f ) {
        box_set_empty( &out->current );
        return;
    }

    orc_box inner = box_intersect( &out->full, &in->current );
    out->current = inner;

    if( box_empty( &inner ) )
        return;

    for( int y = inner.min_y; y <= inner.max_y; y++ ) {
        orc_px32 *row_out = px32( out, inner.min_x, y );
        memmove( row_out, px32( in, inner.min_x, y ), sizeof(orc_px32) * box_width( &inner ) );

        if( alpha != 1.0f ) {
            for( int x = 0; x < box_width( &inner ); x++ )
                row_out[x].a *= alpha;
        }
    }
}

/* ---- bilinear scaler: video_scale.c:24-285 -------------------------------------- */

static orc_frame32 alloc32( const orc_box *full );

/*
 * One axis of the separable triangle-filter scaler (video_scale.c:33-126 vertical, :129-228
 * horizontal), written once over "lines" (rows for the vertical pass, columns for the horizontal
 * one).  Upsampling walks the SOURCE lines in ascending order and spreads each over the target
 * lines its triangle reaches; downsampling walks the TARGET lines and gathers the source lines
 * under its triangle, also ascending -- so every target pixel is  0 + in*c + in*c ...  with the
 * source index increasing, products and sums rounded separately in f32.
 */
enum { ORC_SCALE_MAX_TAPS = 256 };

static void scale_axis( orc_frame32 *target, float tmin, const orc_frame32 *source, float smin, float factor, int vertical ) {
    const orc_box srect = source->current, trect = target->full;

    /* the other axis: where source and target overlap (:36-37, :136-137) */
    const int cross_lo = vertical ? imax( srect.min_x, trect.min_x ) : imax( srect.min_y, trect.min_y );
    const int cross_hi = vertical ? imin( srect.max_x, trect.max_x ) : imin( srect.max_y, trect.max_y );
    const int s_lo = vertical ? srect.min_y : srect.min_x, s_hi = vertical ? srect.max_y : srect.max_x;
    const int t_lo = vertical ? trect.min_y : trect.min_x, t_hi = vertical ? trect.max_y : trect.max_x;

    memset( target->data, 0, box_area( &trect ) * sizeof(orc_px32) );      /* video_fill_zero_f32, :24-31 */

    if( factor == 1.0f && tmin == smin ) {                                  /* :44-47 */
        orc_copy_frame_alpha_f32( target, source, 1.0f );
        return;
    }

    int used_lo = INT32_MAX, used_hi = INT32_MIN;
    float coeff[ORC_SCALE_MAX_TAPS];

    const int upsample = factor > 1.0f;
    const int outer_lo = upsample ? s_lo : t_lo, outer_hi = upsample ? s_hi : t_hi;

    for( int outer = outer_lo; outer <= outer_hi; outer++ ) {
        /* where this line falls on the other grid (:62, :94) */
        const float where = upsample ? (outer - smin) * factor + tmin : (outer - tmin) / factor + smin;
        const int base = (int) floorf( where );
        int width = 0, center = 0;
        if( orc_triangle( factor, where - base, coeff, ORC_SCALE_MAX_TAPS, &width, &center ) )
            continue;

        for( int tap = 0; tap < width; tap++ ) {
            const int inner = base - center + tap;
            const int t = upsample ? inner : outer, sidx = upsample ? outer : inner;

            if( upsample ? (t < t_lo || t > t_hi) : (sidx < s_lo || sidx > s_hi) )
                continue;

            const float c = coeff[tap];
            for( int k = cross_lo; k <= cross_hi; k++ ) {
                const orc_px32 *in = vertical ? px32( source, k, sidx ) : px32( source, sidx, k );
                orc_px32 *out = vertical ? px32( target, k, t ) : px32( target, t, k );
                out->r += in->r * c;
                out->g += in->g * c;
                out->b += in->b * c;
                out->a += in->a * c;
            }

            /* the horizontal pass only learns its extent inside the row loop (:185-186, :216-217) */
            if( vertical || cross_lo <= cross_hi ) {
                used_lo = imin( used_lo, t );
                used_hi = imax( used_hi, t );
            }
        }
    }

    if( vertical ) { target->current.min_x = cross_lo; target->current.max_x = cross_hi; target->current.min_y = used_lo; target->current.max_y = used_hi; }
    else           { target->current.min_x = used_lo; target->current.max_x = used_hi; target->current.min_y = cross_lo; target->current.max_y = cross_hi; }
}

/* video_scale_bilinear_f32, video_scale.c:231-285: one pass when the other axis is the identity, else the
 * axis with the smaller factor first through a temporary whose window is computed with C's (int) truncation */
void orc_scale_bilinear_f32( orc_frame32 *target, float tpx, float tpy, const orc_frame32 *source, float spx, float spy, float fx, float fy ) {
    if( fx == 1.0f && tpx == spx ) {
        if( fy == 1.0f && tpy == spy ) {
            orc_copy_frame_alpha_f32( target, source, 1.0f );
            return;
        }
        scale_axis( target, tpy, source, spy, fy, 1 );
        return;
    }
    if( fy == 1.0f && tpy == spy ) {
        scale_axis( target, tpx, source, spx, fx, 0 );
        return;
    }

    orc_frame32 temp;
    if( fx < fy ) {
        orc_box want = { (int)( spx - (tpx - target->full.min_x) * fx ), source->current.min_y,
                         (int)( spx + (target->full.max_x - tpx) * fx ), source->current.max_y };
        want = box_intersect( &want, &target->full );
        temp = alloc32( &want );
        scale_axis( &temp, tpx, source, spx, fx, 0 );
        scale_axis( target, tpy, &temp, spy, fy, 1 );
    }
    else {
        orc_box want = { source->current.min_x, (int)( spy - (tpy - target->full.min_y) * fy ),
                         source->current.max_x, (int)( spy + (target->full.max_y - tpy) * fy ) };
        want = box_intersect( &want, &target->full );
        temp = alloc32( &want );
        scale_axis( &temp, tpy, source, spy, fy, 1 );
        scale_axis( target, tpx, &temp, spx, fx, 0 );
    }
    free( temp.data );
}

/* ---- crossfade: video_mix.c:107-235 (intended window semantics) ------------------ */
void orc_mix_cross_f32( orc_frame32 *out, const orc_frame32 *a, const orc_frame32 *b, float mix_b ) {
    mix_b = clampf( mix_b, 0.0f, 1.0f );
    const float mix_a = 1.0f - mix_b;

    if( box_empty( &a->current ) ) {
        orc_copy_frame_alpha_f32( out, b, mix_b );
        return;
    }
    else if( box_empty( &b->current ) ) {
        orc_copy_frame_alpha_f32( out, a, mix_a );
        return;
    }

    orc_box outer = box_union( &a->current, &b->current );
    outer = box_intersect( &outer, &out->full );

    const orc_px32 zero = { 0.0f, 0.0f, 0.0f, 0.0f };

    for( int y = outer.min_y; y <= outer.max_y; y++ ) {
        for( int x = outer.min_x; x <= outer.max_x; x++ ) {
            const int in_a = box_contains( &a->current, x, y ), in_b = box_contains( &b->current, x, y );
            orc_px32 *o = px32( out, x, y );

            if( in_a && in_b ) {
                /* :192-206 */
                const orc_px32 pa = *px32( a, x, y ), pb = *px32( b, x, y );
                float alpha_a = pa.a * mix_a;
                float alpha_b = pb.a * mix_b;
                float alpha = alpha_a + alpha_b;

                if( alpha != 0.0f ) {
                    o->r = (pa.r * alpha_a + pb.r * alpha_b) / alpha;
                    o->g = (pa.g * alpha_a + pb.g * alpha_b) / alpha;
                    o->b = (pa.b * alpha_a + pb.b * alpha_b) / alpha;
                    o->a = alpha;
                }
                else {
                    *o = zero;
                }
            }
            else if( in_a ) {
                orc_px32 p = *px32( a, x, y );
                p.a *= mix_a;
                *o = p;
            }
            else if( in_b ) {
                orc_px32 p = *px32( b, x, y );
                p.a *= mix_b;
                *o = p;
            }
            else {
                *o = zero;
            }
        }
    }

    out->current = outer;
}

/* ---- alpha-over: video_mix.c:237-370 (intended window semantics) ------------------ */
void orc_mix_over_f32( orc_frame32 *out, const orc_frame32 *b, float mix_b ) {
    mix_b = clampf( mix_b, 0.0f, 1.0f );

    if( box_empty( &out->current ) ) {
        orc_copy_frame_alpha_f32( out, b, mix_b );
        return;
    }
    else if( box_empty( &b->current ) || mix_b == 0.0f ) {
        return;
    }

    orc_box outer = box_union( &out->current, &b->current );
    outer = box_intersect( &outer, &out->full );

    const orc_px32 zero = { 0.0f, 0.0f, 0.0f, 0.0f };

    for( int y = outer.min_y; y <= outer.max_y; y++ ) {
        for( int x = outer.min_x; x <= outer.max_x; x++ ) {
            const int in_o = box_contains( &out->current, x, y ), in_b = box_contains( &b->current, x, y );
            orc_px32 *o = px32( out, x, y );

            if( in_o && in_b ) {
                /* :323-337 */
                const orc_px32 pb = *px32( b, x, y );
                float alpha_b = pb.a * mix_b;
                float alpha_a = o->a * (1.0f - pb.a * mix_b);
                float alpha = alpha_a + alpha_b;

                if( alpha != 0.0f ) {
                    o->r = (o->r * alpha_a + pb.r * alpha_b) / alpha;
                    o->g = (o->g * alpha_a + pb.g * alpha_b) / alpha;
                    o->b = (o->b * alpha_a + pb.b * alpha_b) / alpha;
                    o->a = alpha;
                }
                else {
                    *o = zero;
                }
            }
            else if( in_o ) {
                /* untouched */
            }
            else if( in_b ) {
                orc_px32 p = *px32( b, x, y );
                p.a *= mix_b;
                *o = p;
            }
            else {
                *o = zero;
            }
        }
    }

    out->current = outer;
}

/* ---- precision hops: main.c:56-70, 126-139 ------------------------------------------ */
void orc_frame_f16_to_f32( orc_frame32 *out, const orc_frame16 *in ) {
    out->current = in->current;
    if( box_empty( &in->current ) )
        return;
    for( int y = in->current.min_y; y <= in->current.max_y; y++ )
        orc_h2f_n( &px32( out, in->current.min_x, y )->r, &px16( in, in->current.min_x, y )->r, 4 * (size_t) box_width( &in->current ) );
}

void orc_frame_f32_to_f16( orc_frame16 *out, const orc_frame32 *in ) {
    out->current = in->current;
    if( box_empty( &in->current ) )
        return;
    for( int y = in->current.min_y; y <= in->current.max_y; y++ )
        orc_f2h_n( &px16( out, in->current.min_x, y )->r, &px32( in, in->current.min_x, y )->r, 4 * (size_t) box_width( &in->current ) );
}

static orc_frame32 alloc32( const orc_box *full ) {
    orc_frame32 f;
    f.data = malloc( sizeof(orc_px32) * (box_area( full ) ? box_area( full ) : 1) );
    f.full = *full;
    box_set_empty( &f.current );
    return f;
}

static orc_frame16 alloc16( const orc_box *full ) {
    orc_frame16 f;
    f.data = malloc( sizeof(orc_px16) * (box_area( full ) ? box_area( full ) : 1) );
    f.full = *full;
    box_set_empty( &f.current );
    return f;
}

/* video_get_frame_f32 on an f16 frame-source (main.c:116-141 + RgbaFrameF16.c:27-30) */
static void fetch_f16_as_f32( orc_frame32 *out, const orc_frame16 *source ) {
    orc_frame16 temp = alloc16( &out->full );
    orc_copy_frame_f16( &temp, source );
    orc_frame_f16_to_f32( out, &temp );
    free( temp.data );
}

/* ---- workspace compositor: workspace.c:494-550 through main.c:34-77 ------------------ */
void orc_composite_f16( orc_frame16 *out, const orc_frame16 *layers, int count ) {
    if( count <= 0 ) {
        box_set_empty( &out->current );     /* workspace.c:504-508 */
        return;
    }

    orc_frame32 acc = alloc32( &out->full );
    fetch_f16_as_f32( &acc, &layers[0] );   /* :530 */

    if( count > 1 ) {
        orc_frame32 temp = alloc32( &out->full );

        for( int i = 1; i < count; i++ ) {  /* :541-544 */
            fetch_f16_as_f32( &temp, &layers[i] );
            orc_mix_over_f32( &acc, &temp, 1.0f );
        }

        free( temp.data );
    }

    orc_frame_f32_to_f16( out, &acc );
    free( acc.data );
}

/* ---- 2:3 pulldown field weave: src/process/Pulldown23RemovalFilter.c:82-104 -------------------
 * The even rows (field 1) of frame's current window are replaced by the same rows of `even_rows`.  The reference
 * copies them out of a temporary it pulled over frame->current_window, starting at that temporary's pixel x = 0
 * (video_get_pixel_f16( &tempFrame, 0, i ), :97-99) -- which is the window's left edge only when min.x == 0; the
 * intended copy (from min.x) is what is restated here, and what the reference does for every window with min.x == 0. */
void orc_weave_fields_f16( orc_frame16 *frame, const orc_frame16 *even_rows ) {
    orc_box box = box_intersect( &frame->current, &even_rows->current );
    box = box_intersect( &box, &frame->full );
    if( box_empty( &box ) )
        return;
    for( int y = (box.min_y + 1) & ~1; y <= box.max_y; y += 2 )
        memcpy( px16( frame, box.min_x, y ), px16( even_rows, box.min_x, y ), sizeof(orc_px16) * (size_t)( box.max_x - box.min_x + 1 ) );
}

/* ---- the workspace compositor with per-pixel filters in front of its layers -------------------
 * What the editor builds when an item's source is a chain Pulldown23RemovalFilter -> colour transform ->
 * VideoGainOffsetFilter -> crop: every filter is applied to the layer as its own pass (the functions above and below,
 * in that order), the filtered layers are stacked by workspace.c:530-547.  Precision of the chain: the weave and the
 * colour transform produce f16 frames, gain / offset works on the f32 frame the compositor pulls (video_filter.c:27-39
 * on an f32 texture) and its result stays f32. */
void orc_composite_ops_f16( orc_frame16 *out, const orc_frame16 *layers, const orc_layer_ops *ops, int count ) {
    if( count <= 0 ) {
        box_set_empty( &out->current );
        return;
    }
    orc_frame32 acc = alloc32( &out->full ), temp = alloc32( &out->full );

    for( int i = 0; i < count; i++ ) {
        orc_frame16 filtered = alloc16( &out->full );
        orc_copy_frame_f16( &filtered, &layers[i] );
        const orc_layer_ops *o = ops ? &ops[i] : NULL;

        if( o && (o->flags & ORC_OP_WEAVE) ) orc_weave_fields_f16( &filtered, &o->even_rows );
        if( o && (o->flags & ORC_OP_COLOR) ) {
            if( o->color == ORC_COLOR_RGB_TO_XYZ_SDTV ) orc_color_rgb_to_xyz_sdtv( &filtered );
            else orc_color_xyz_to_srgb( &filtered );
        }
        orc_frame32 *target = (i == 0) ? &acc : &temp;
        orc_frame_f16_to_f32( target, &filtered );
        if( o && (o->flags & ORC_OP_GAIN_OFFSET) ) {
            const orc_box box = box_intersect( &target->current, &target->full );
            for( int y = box.min_y; y <= box.max_y; y++ )
                for( int x = box.min_x; x <= box.max_x; x++ ) {
                    orc_px32 *p = px32( target, x, y );
                    p->r = p->r * o->gain + o->offset; p->g = p->g * o->gain + o->offset; p->b = p->b * o->gain + o->offset;
                }
        }
        if( o && (o->flags & ORC_OP_CROP) ) {
            const orc_box box = box_intersect( &target->current, &target->full );
            for( int y = box.min_y; y <= box.max_y; y++ )
                for( int x = box.min_x; x <= box.max_x; x++ )
                    if( !box_contains( &o->crop, x, y ) )
                        memset( px32( target, x, y ), 0, sizeof(orc_px32) );
        }
        if( i > 0 ) orc_mix_over_f32( &acc, &temp, 1.0f );
        free( filtered.data );
    }

    orc_frame_f32_to_f16( out, &acc );
    free( acc.data );
    free( temp.data );
}

/* ---- crossfade pulled as f16: video_mix.c:46-71 via main.c:44-72 ---------------------- */
void orc_cross_f16( orc_frame16 *out, const orc_frame16 *a, const orc_frame16 *b, float mix_b ) {
    orc_frame32 acc = alloc32( &out->full );
    mix_b = clampf( mix_b, 0.0f, 1.0f );

    if( mix_b == 0.0f ) {
        fetch_f16_as_f32( &acc, a );
    }
    else if( mix_b == 1.0f ) {
        fetch_f16_as_f32( &acc, b );
    }
    else {
        orc_frame32 temp = alloc32( &out->full );
        fetch_f16_as_f32( &acc, a );
        fetch_f16_as_f32( &temp, b );
        orc_mix_cross_f32( &acc, &acc, &temp, mix_b );
        free( temp.data );
    }

    orc_frame_f32_to_f16( out, &acc );
    free( acc.data );
}

/* ---- timeline frame (configs[4]) ------------------------------------------------------- */
static void fetch_coded_as_f32( orc_frame32 *out, const orc_coded *im ) {
    /* DVReconstructionFilter.c:53-71 (f16-only source) pulled as f32: main.c:116-141 */
    orc_frame16 temp = alloc16( &out->full );
    orc_reconstruct( &temp, im );
    orc_frame_f16_to_f32( out, &temp );
    free( temp.data );
}

void orc_timeline_frame_f16( orc_frame16 *out, const orc_coded *a, const orc_coded *b, float mix_b,
        const orc_frame16 *overlays, int overlay_count ) {
    orc_frame32 acc = alloc32( &out->full );
    int have_bottom = 0;

    if( a && b ) {
        /* video_mix_cross_f32_pull, video_mix.c:46-71 */
        mix_b = clampf( mix_b, 0.0f, 1.0f );

        if( mix_b == 0.0f ) {
            fetch_coded_as_f32( &acc, a );
        }
        else if( mix_b == 1.0f ) {
            fetch_coded_as_f32( &acc, b );
        }
        else {
            orc_frame32 temp = alloc32( &out->full );
            fetch_coded_as_f32( &acc, a );
            fetch_coded_as_f32( &temp, b );
            orc_mix_cross_f32( &acc, &acc, &temp, mix_b );
            free( temp.data );
        }
        have_bottom = 1;
    }
    else if( a || b ) {
        fetch_coded_as_f32( &acc, a ? a : b );
        have_bottom = 1;
    }

    orc_frame32 temp = alloc32( &out->full );

    for( int i = 0; i < overlay_count; i++ ) {
        if( !have_bottom ) {
            fetch_f16_as_f32( &acc, &overlays[i] );
            have_bottom = 1;
            continue;
        }
        fetch_f16_as_f32( &temp, &overlays[i] );
        orc_mix_over_f32( &acc, &temp, 1.0f );
    }

    free( temp.data );

    if( !have_bottom )
        box_set_empty( &acc.current );

    orc_frame_f32_to_f16( out, &acc );
    free( acc.data );
}

/* ---- 8-bit encodes --------------------------------------------------------------------- */
/* widget_gl.c:290-307 */
void orc_encode_u8( orc_frame16 *frame, const uint16_t *lut_srgb, const uint8_t *ramp, orc_px8 *target ) {
    const int width = box_width( &frame->full );

    if( box_empty( &frame->current ) )
        return;

    for( int y = frame->current.min_y; y <= frame->current.max_y; y++ ) {
        for( int x = frame->current.min_x; x <= frame->current.max_x; x++ ) {
            orc_px16 *p = px16( frame, x, y );
            orc_px8 *t = &target[(size_t)(y - frame->full.min_y) * width + (x - frame->full.min_x)];

            p->r = lut_srgb[p->r];
            p->g = lut_srgb[p->g];
            p->b = lut_srgb[p->b];
            p->a = lut_srgb[p->a];

            t->r = ramp[p->r];
            t->g = ramp[p->g];
            t->b = ramp[p->b];
            t->a = ramp[p->a];
        }
    }
}

/* RgbaFrameF16.c:114-149 */
void orc_to_argb32( const orc_frame16 *frame, const uint8_t *ramp45, uint32_t *out ) {
    if( box_empty( &frame->current ) )
        return;

    const int w = box_width( &frame->current );

    for( int y = frame->current.min_y; y <= frame->current.max_y; y++ ) {
        const orc_px16 *row = px16( frame, frame->current.min_x, y );

        for( int x = 0; x < w; x++ ) {
            uint32_t a = ramp45[row[x].a];

            out[(size_t)(y - frame->current.min_y) * w + x] =
                (a << 24) |
                (((ramp45[row[x].r] * a >> 8) & 0xFF) << 16) |
                (((ramp45[row[x].g] * a >> 8) & 0xFF) << 8) |
                ((ramp45[row[x].b] * a >> 8) & 0xFF);
        }
    }
}

/* ---- synthetic inputs -------------------------------------------------------------------- */
uint32_t orc_hash32( uint32_t seed, uint32_t index ) {
    uint32_t h = index * 0x9E3779B1u + seed * 0x85EBCA77u + 0x165667B1u;
    h ^= h >> 15;
    h *= 0x2C1B3C6Du;
    h ^= h >> 12;
    h *= 0x297A2D39u;
    h ^= h >> 15;
    return h;
}

/*
 * Pixel i of the full window (row-major): u0 = hash(seed, 2i), u1 = hash(seed, 2i+1).
 * r,g,b = f2h( 16-bit field * 2^-16 ) in [0,1); alpha picked by (u1 >> 16) & 7:
 * 0 -> 0, 1 -> 1/4, 2 -> 1/2, 3 -> 3/4, 4,5 -> 1, 6,7 -> 13-bit fraction.
 */
void orc_fill_frame( orc_frame16 *frame, uint32_t seed ) {
    const size_t count = box_area( &frame->full );

    for( size_t i = 0; i < count; i++ ) {
        uint32_t u0 = orc_hash32( seed, (uint32_t)(2 * i) ), u1 = orc_hash32( seed, (uint32_t)(2 * i + 1) );
        orc_px16 *p = &frame->data[i];

        p->r = orc_f2h( (float)(u0 & 0xFFFF) * (1.0f / 65536.0f) );
        p->g = orc_f2h( (float)(u0 >> 16) * (1.0f / 65536.0f) );
        p->b = orc_f2h( (float)(u1 & 0xFFFF) * (1.0f / 65536.0f) );

        uint32_t sel = (u1 >> 16) & 7;
        float alpha;

        switch( sel ) {
            case 0: alpha = 0.0f; break;
            case 1: alpha = 0.25f; break;
            case 2: alpha = 0.5f; break;
            case 3: alpha = 0.75f; break;
            case 4: case 5: alpha = 1.0f; break;
            default: alpha = (float)(u1 >> 19) * (1.0f / 8192.0f); break;
        }

        p->a = orc_f2h( alpha );
    }
}

/* byte i = lo + hash(seed, i) % (hi - lo + 1) */
void orc_fill_bytes( uint8_t *data, size_t count, uint32_t seed, int lo, int hi ) {
    const uint32_t span = (uint32_t)(hi - lo + 1);
    for( size_t i = 0; i < count; i++ )
        data[i] = (uint8_t)(lo + (int)(orc_hash32( seed, (uint32_t) i ) % span));
}

uint32_t orc_fnv1a( const void *data, size_t count ) {
    const uint8_t *p = (const uint8_t *) data;
    uint32_t h = 0x811C9DC5u;
    for( size_t i = 0; i < count; i++ )
        h = (h ^ p[i]) * 0x01000193u;
    return h;
}
