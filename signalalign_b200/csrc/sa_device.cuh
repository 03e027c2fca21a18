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

// The reference's four cubics for log(1 + e^x) on [0,1], (1,2.5], (2.5,4.5], (4.5,7.5): FLOAT literals promoted
// to double (impl/pairwiseAligner.c:305-311).  They live in a shared-memory table with one row (c3 c2 c1 c0) per value of
// n = ceil(2 |x - y|) = 0..15: the range tests d > 1, d > 2.5, d > 4.5 are exactly n >= 3, n >= 6, n >= 10 (2 d is exact
// and the thresholds are half-integers), so one DFMA replaces three DSETP and the predicate arithmetic, and a logAdd
// fetches its coefficients with two conflict-free LDS.128.  Row 16 is all zeros: the "far" case (|x - y| >= 7.5, or an
// operand -inf) evaluates 0 + max(x, y) through the same instructions.  Every kernel calls sa_logAdd_init() once per CTA.
__shared__ __align__(16) double sa_lt[68];
__shared__ __align__(16) float sa_ltf[68];      // single precision, linear coefficient minus one (sa_logAddF, the FP32 band kernel)

__device__ __forceinline__ void sa_logAdd_init() {
    const float c[16] = {-0.009350833524763f, 0.130659527668286f, 0.498799810682272f, 0.693203116424741f,
                         -0.014532321752540f, 0.139942324101744f, 0.495635523139337f, 0.692140569840976f,
                         -0.004605031767994f, 0.063427417320019f, 0.695956496475118f, 0.514272634594009f,
                         -0.000458661602210f, 0.009695946122598f, 0.930734667215156f, 0.168037164329057f};
    for (int i = threadIdx.x; i < 68; i += blockDim.x) {
        const int n = i >> 2, range = (n >= 3) + (n >= 6) + (n >= 10), k = range * 4 + (i & 3);
        float v = c[0];
#pragma unroll
        for (int q = 1; q < 16; q++) v = k == q ? c[q] : v;
        sa_lt[i] = n >= 16 ? 0.0 : (double) v;
        sa_ltf[i] = n >= 16 ? 0.0f : ((i & 3) == 2 ? v - 1.0f : v);
    }
    __syncthreads();
}

// logAdd (impl/pairwiseAligner.c:314-318): max when either operand is -inf or |x-y| >= 7.5, else
// min + lookup(|x-y|), Horner without contraction.  Branch-free, 19 instructions:
//   * far = |diff| >= 7.5, +inf (one operand -inf) or NaN (both -inf), decided on the HIGH WORD of |diff| as an integer
//     (7.5 = 0x401E0000 00000000: the low word cannot matter);
//   * ONE two-word select picks the operand the polynomial is added to: the larger one when far (the zero row then adds
//     +0, which is exact), the smaller one otherwise;
//   * the polynomial's argument is |diff| with its high word clamped to that of 8.0, so that it stays finite in the far case
//     (0 * inf would be NaN); for |diff| < 7.5 the clamp is the identity.
__device__ __forceinline__ double sa_logAdd(double x, double y) {
    const double diff = __dsub_rn(x, y);
    const int dh = __double2hiint(diff);
    const unsigned ah = (unsigned) dh & 0x7fffffffu;
    const bool far = ah >= 0x401E0000u;
    const bool xlt = dh < 0;                       // x < y  (diff == +0 when equal)
    const double base = (far == xlt) ? y : x;      // far: max(x, y); else min(x, y)
    const double d = __hiloint2double((int) umin(ah, 0x40200000u), __double2loint(diff));
    // n = ceil(2 d): 2 d + 2^52 rounded UP leaves it in the low mantissa word (one DFMA on the FP64 pipe)
    unsigned n = (unsigned) __double2loint(__fma_ru(d, 2.0, 4503599627370496.0));
    n = far ? 16u : n;
    const double2 ca = *reinterpret_cast<const double2 *>(sa_lt + 4 * n);
    const double2 cb = *reinterpret_cast<const double2 *>(sa_lt + 4 * n + 2);
    double r = __dmul_rn(ca.x, d);
    r = __dmul_rn(__dadd_rn(r, ca.y), d);
    r = __dmul_rn(__dadd_rn(r, cb.x), d);
    return __dadd_rn(__dadd_rn(r, cb.y), base);
}

// The same logAdd in single precision for the FP32 band kernel (sa_wide32.cuh).  The reference returns min + P(d) with
// d = |x - y| and P one of its four cubics for log(1 + e^d); since min = max - d this is max + (P(d) - d), and P(d) - d is
// the same cubic with 1 taken off its linear coefficient (the table above stores it that way).  That form needs neither
// the "far" test nor a select between the operands: d is clamped to 8 on its bit pattern (which also absorbs +-inf and the
// NaN of -inf - -inf), ceil(2 d) = 16 from 7.5 up, and row 16 is zeros, so a far pair evaluates max + 0.  The range index
// comes out of one FFMA.RU (2 d + 2^23 rounded up leaves it in the low mantissa bits, and shifted left by four it IS the
// byte offset of the row once 2^23's own bits are folded into the base address), the four coefficients arrive in one
// LDS.128, the cubic is three FFMA: 11 instructions.  Thresholds are the reference's except at the single float d == 7.5f
// (the cubic's 5.5e-4 instead of 0), which stands for true differences on both sides of 7.5 in equal measure.
// (sa_logAddKey: the folded base address.  The kernel parks it in shared memory and reads it back once per diagonal — a
// value that comes out of a load is one the assembler cannot split into "base" and "constant" again at every use)
__device__ __forceinline__ unsigned sa_logAddKey() { return (unsigned) __cvta_generic_to_shared(sa_ltf) - 0xB0000000u; }
__device__ __forceinline__ float sa_logAddF(float x, float y, unsigned key) {
    const float diff = __fsub_rn(x, y);
    const float d = __uint_as_float(umin(__float_as_uint(diff) & 0x7fffffffu, 0x41000000u));
    const unsigned nb = __float_as_uint(__fmaf_ru(d, 2.0f, 8388608.0f));           // 0x4B000000 + ceil(2 d)
    const unsigned addr = (nb << 4) + key;
    float4 c;
    asm("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(c.x), "=f"(c.y), "=f"(c.z), "=f"(c.w) : "r"(addr));
    const float r = fmaf(fmaf(fmaf(c.x, d, c.y), d, c.z), d, c.w);
    return __fadd_rn(fmaxf(x, y), r);
}

// exp() for the expectation sums (cell_signal_updateExpectations, impl/pairwiseAligner.c:914-944): every band cell adds
// seven exp(f + b + eP + tP - total) terms, so the library exp() (about 35 instructions with its special-case branches)
// was 45 % of the expectations kernel.  Here: n = rint(a log2 e) by the 1.5 * 2^52 shift, Cody-Waite reduction with a
// two-word ln 2, degree-10 Taylor polynomial on |r| <= 0.347 (truncation 2.2e-13 relative), 2^n added into the exponent
// field.  Results below the normal range (a < -708) are flushed to zero: each is < 1e-307 against sums >= 1e-3
// (the parity bar on the sums is 1e-6 relative).  NaN, +inf and a > 700 take the library exp().
__device__ __noinline__ double sa_expSlow(double a) { return exp(a); }
__device__ __forceinline__ double sa_expSum(double a) {
    if (!(a <= 700.0)) return sa_expSlow(a);
    const double t = fma(a, 1.4426950408889634, 6755399441055744.0);
    const int n = __double2loint(t);
    const double nf = __dsub_rn(t, 6755399441055744.0);
    double r = fma(nf, -6.93147180369123816490e-01, a);
    r = fma(nf, -1.90821492927058770002e-10, r);
    double p = 2.7557319223985893e-07;            // 1/10!
    p = fma(p, r, 2.7557319223985888e-06);        // 1/9!
    p = fma(p, r, 2.4801587301587302e-05);        // 1/8!
    p = fma(p, r, 1.9841269841269841e-04);        // 1/7!
    p = fma(p, r, 1.3888888888888889e-03);        // 1/6!
    p = fma(p, r, 8.3333333333333332e-03);        // 1/5!
    p = fma(p, r, 4.1666666666666664e-02);        // 1/4!
    p = fma(p, r, 1.6666666666666666e-01);        // 1/3!
    p = fma(p, r, 0.5);
    p = fma(p, r, 1.0);
    p = fma(p, r, 1.0);
    const double v = __hiloint2double(__double2hiint(p) + (n << 20), __double2loint(p));
    return a < -708.0 ? 0.0 : v;
}
