/*
 * include/half.h -- canvas-b200 re-issue of the half-float entry points.
 *
 * Replaces /root/reference/include/half.h:26-35 for the per-pixel path.  The five
 * entry points stay global FUNCTION POINTERS with the same names and argument
 * meaning, so C code written against the reference header links unchanged; here
 * init_half() binds them to launchers of sm_100a kernels (host buffers are
 * staged through pinned memory), never to CPU loops.
 *
 * Semantics kept bit-for-bit (reference src/cprocess/half.c:31-51, genhalf.py):
 *   - float -> half TRUNCATES (round toward zero), keeps subnormals, maps finite
 *     magnitudes >= 65536 to infinity and keeps the top ten NaN payload bits;
 *   - half -> float is exact;
 *   - half_lookup is out[i] = table[in[i]] over a 65536-entry table.
 * The "_fast" pair is bound to the same kernels (the reference's fast variants
 * only relax the handling of subnormals/NaN; being exact is allowed).
 */
#ifndef CANVAS_B200_HALF_H
#define CANVAS_B200_HALF_H

#include <stdint.h>

#if defined(__cplusplus)
extern "C" {
#endif

typedef uint16_t half;

enum { HALF_COUNT = 65536 };

/* Binds the pointers below.  Fails loudly (abort with a message) when no CUDA
 * device can be initialised: there is no CPU fallback. */
void init_half( void );

extern void (*half_convert_from_float)( half *out, const float *in, int count );
extern void (*half_convert_to_float)( float *out, const half *in, int count );
extern void (*half_convert_from_float_fast)( half *out, const float *in, int count );
extern void (*half_convert_to_float_fast)( float *out, const half *in, int count );
extern void (*half_lookup)( const half *table, half *out, const half *in, int count );

#if defined(__cplusplus)
}
#endif

#endif
