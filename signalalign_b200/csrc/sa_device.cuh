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

// log(1 + e^x) on [0, 7.5): the reference's four cubics; coefficients are FLOAT literals promoted to double
// (impl/pairwiseAligner.c:305-311), evaluated by Horner without contraction.
__device__ __forceinline__ double sa_lookup(double x) {
    double c3, c2, c1, c0;
    if (x <= 1.0) {
        c3 = (double) -0.009350833524763f; c2 = (double) 0.130659527668286f;
        c1 = (double) 0.498799810682272f;  c0 = (double) 0.693203116424741f;
    } else if (x <= 2.5) {
        c3 = (double) -0.014532321752540f; c2 = (double) 0.139942324101744f;
        c1 = (double) 0.495635523139337f;  c0 = (double) 0.692140569840976f;
    } else if (x <= 4.5) {
        c3 = (double) -0.004605031767994f; c2 = (double) 0.063427417320019f;
        c1 = (double) 0.695956496475118f;  c0 = (double) 0.514272634594009f;
    } else {
        c3 = (double) -0.000458661602210f; c2 = (double) 0.009695946122598f;
        c1 = (double) 0.930734667215156f;  c0 = (double) 0.168037164329057f;
    }
    return __dadd_rn(__dmul_rn(__dadd_rn(__dmul_rn(__dadd_rn(__dmul_rn(c3, x), c2), x), c1), x), c0);
}

// logAdd (impl/pairwiseAligner.c:314-318): max when either is -inf or |x-y| >= 7.5, else min + lookup(|x-y|)
__device__ __forceinline__ double sa_logAdd(double x, double y) {
    double hi = x < y ? y : x;
    double lo = x < y ? x : y;
    double d = __dsub_rn(hi, lo);
    if (lo == SA_NEG_INF || d >= 7.5) return hi;
    return __dadd_rn(sa_lookup(d), lo);
}
