// sa_variants.cuh — variant calls of aligned pairs and their per-position marginals on the device.
//
// Reference: writePosteriorProbsVC (impl/signalMachine.c:161-232) writes one TSV row (event, reference position, called
// base, posterior "%f") per ambiguity character X (AMBIG_BASE, inc/pairwiseAligner.h:32) of the reference kmer of every
// aligned pair; MarginalizeVariants.get_data (src/signalalign/variantCaller.py:37-89) reads the rows back, sums the
// posterior per (position, base), normalises over the bases of a position and averages the positions of a read.
//
// Both steps collapse once the coordinate algebra of the writer is carried out.  With o the offset of an X inside the
// pair's kmer and t = x + o its index in the aligned sequence:
//   * forward template / backward complement: reported position = (x + referenceSequenceOffset) + o = t + offset;
//   * otherwise the writer reverse-complements the kmer (X at q = k-1-o), reports
//     (refLength - k) - (x + refLength - offset) + q = offset - 1 - t, and reads the base at (k-1) - q = o;
// so in every case the row is (position(t), pathKmer[o], p): a pair contributes p to mass[t][base at o] for every X it
// covers.  One CTA per read; the masses are accumulated with shared / global atomics (the reference's own summation
// order is Python's set order, i.e. unspecified: results agree to rounding, 1e-13 relative in the tests).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>
#include "signalalign_b200.h"

struct VarRead {
    int64_t pairOff, nPairs;      // into VarParams.pairs
    int64_t rankOff, targetLen;   // into VarParams.xRank: rank of t among the X positions of the target, -1 elsewhere
    int64_t massOff;              // into VarParams.mass: nX x V doubles
    int32_t nX, pad;
};

struct VarParams {
    const VarRead *reads;
    const sa_pair_t *pairs;
    const int32_t *xRank;
    double *mass;
    int32_t digitToVariant[256];  // alphabet digit -> column (index in the sorted variants), -1 if the base is not a variant
    int32_t A, k, V, roundLikeTsv;
    int32_t *err;                 // 1: a called base outside `variants`, 2: pair outside the target
};

// the six decimals a "%f" field keeps, in millionths: exact decimal rounding of the double p / 1e7, ties (only possible
// where p / 1e7 is a dyadic rational) to even, as glibc's printf does
__device__ __forceinline__ int32_t var_round6(int32_t p) {
    const int32_t q = p / 10, digit = p - 10 * q;
    if (digit != 5) return q + (digit > 5);
    const double v = __ddiv_rn((double) p, 1e7);
    const double e = fma(v, 1e7, -(double) p);          // sign of (stored double - exact decimal)
    if (e > 0) return q + 1;
    if (e < 0) return q;
    return q + (q & 1);
}

template <int NT>
__global__ void __launch_bounds__(NT) sa_variant_mass_kernel(const __grid_constant__ VarParams P) {
    const VarRead R = P.reads[blockIdx.x];
    const sa_pair_t *pairs = P.pairs + R.pairOff;
    const int32_t *rank = P.xRank + R.rankOff;
    double *mass = P.mass + R.massOff;
    int64_t pw[16];
    pw[0] = 1;
    for (int i = 1; i < 16; i++) pw[i] = pw[i - 1] * P.A;
    for (int64_t i = threadIdx.x; i < R.nPairs; i += NT) {
        const sa_pair_t a = pairs[i];
        if (a.x < 0 || (int64_t) a.x + P.k > R.targetLen) { atomicMax(P.err, 2); continue; }
        double pv = 0.0;
        bool have = false;
        for (int o = 0; o < P.k; o++) {
            const int32_t r = rank[a.x + o];
            if (r < 0) continue;
            if (!have) {
                pv = P.roundLikeTsv ? __ddiv_rn((double) var_round6(a.p), 1e6) : __ddiv_rn((double) a.p, 1e7);
                have = true;
            }
            const int digit = (int) ((a.kmerIdx / pw[P.k - 1 - o]) % P.A);
            const int32_t col = P.digitToVariant[digit];
            if (col < 0) { atomicMax(P.err, 1); continue; }
            atomicAdd(mass + (int64_t) r * P.V + col, pv);
        }
    }
}
