// sa_device.cuh — device-side arithmetic shared by the kernels of libsagpu (sm_100a).
//
// Parity contract: the reference computes in double precision with an APPROXIMATE, non-associative logAdd
// (impl/pairwiseAligner.c:298-318).  Posteriors are only reproducible to 1e-4 if the same cubic pieces, the same
// range tests, the same 7.5 cut-off AND the same accumulation order are used, so everything here
// is FP64 with explicit operation order; the library is compiled with -fmad=false and FMA is used only where
// written out as fma().
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#define SA_NEG_INF (__longlong_as_double(0xfff0000000000000LL))

// The reference's cubic coefficients are FLOAT literals promoted to double (impl/pairwiseAligner.c:305-311): the
// float bit pattern is selected per range (one 32-bit word per coefficient) and widened exactly with F2F.
#define SA_FB(f) ((unsigned) __float_as_int(f))

// logAdd (impl/pairwiseAligner.c:314-318): max when either operand is -inf or |x-y| >= 7.5, else
// min + lookup(|x-y|) with lookup = the reference's four cubics for log(1 + e^x), Horner without contraction.
// Branch-free, integer tests on the high word wherever that is exact.
__device__ __forceinline__ double sa_logAdd(double x, double y) {
    const double diff = __dsub_rn(x, y);
    const int dh = __double2hiint(diff);
    const bool xlt = dh < 0;                       // x < y  (diff == +0 when equal; NaN only when both are -inf)
    const double hi = xlt ? y : x;
    const double lo = xlt ? x : y;
    const int ah = dh & 0x7fffffff;                // high word of |diff|
    // |diff| >= 7.5, |diff| == inf (one operand -inf) and NaN (both -inf) all have ah >= hi(7.5) = 0x401E0000
    const bool far = ah >= 0x401E0000;
    const double d = __hiloint2double(far ? 0 : ah, far ? 0 : __double2loint(diff));   // clamped: cubic result unused when far
    // range index as three nested 0/1 flags; the coefficient bit patterns are picked with integer multiply-adds
    // (wrap-around arithmetic is exact), which keeps the selection off the branch unit
    const unsigned i1 = d > 1.0, i2 = d > 2.5, i3 = d > 4.5;
#define SA_PICK(a0, a1, a2, a3)                                                                                   \
    ((double) __int_as_float((int) (SA_FB(a0) + i1 * (SA_FB(a1) - SA_FB(a0)) + i2 * (SA_FB(a2) - SA_FB(a1)) +      \
                                    i3 * (SA_FB(a3) - SA_FB(a2)))))
    const double c3 = SA_PICK(-0.009350833524763f, -0.014532321752540f, -0.004605031767994f, -0.000458661602210f);
    const double c2 = SA_PICK(0.130659527668286f, 0.139942324101744f, 0.063427417320019f, 0.009695946122598f);
    const double c1 = SA_PICK(0.498799810682272f, 0.495635523139337f, 0.695956496475118f, 0.930734667215156f);
    const double c0 = SA_PICK(0.693203116424741f, 0.692140569840976f, 0.514272634594009f, 0.168037164329057f);
#undef SA_PICK
    double r = __dmul_rn(c3, d);
    r = __dmul_rn(__dadd_rn(r, c2), d);
    r = __dmul_rn(__dadd_rn(r, c1), d);
    r = __dadd_rn(__dadd_rn(r, c0), lo);
    return far ? hi : r;
}
