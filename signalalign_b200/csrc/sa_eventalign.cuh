// sa_eventalign.cuh — adaptive banded event -> kmer alignment on the device.
//
// Reference: adaptive_banded_simple_event_align2 (impl/eventAligner.c:899-1235): the Suzuki-Kasahara adaptive band as
// nanopolish uses it.  Band b holds 100 cells of the anti-diagonal (event + 1) + (kmer + 1) = b; its lower-left corner moves
// DOWN (next event) or RIGHT (next kmer) from band to band, decided by comparing the two end cells of the previous band
// (:1006-1018).  Cells are a Viterbi maximum over step (diagonal, band b-2), stay (up, band b-1) and skip (left, band b-1)
// with the ..._MeanOnly emission (impl/stateMachine.c:557-605), scores rounded to float exactly where the reference rounds
// them (:1087-1098); trim states along kmer -1 (:1044-1053); then the best end cell in the last kmer's column (:1136-1148),
// the backtrack with the reference's QC figures (:1157-1223).
//
// One CTA per read, thread o = cell o of the band; the last three bands live in shared memory as DOUBLES (the trim cells
// and the start cell are not float-rounded in the reference, and the move rule compares them as doubles); the trace bytes,
// the band corners and the last kmer's column go to global memory for the backtrack, which one thread walks.  The band
// sweep is sequential by construction (the corner of band b depends on the scores of band b-1); the batch supplies the
// parallelism.  Every operation keeps the reference's order (the library is built with -fmad=false, divisions IEEE), the
// per-kmer log(sigma) and the transition logs come from the host's libm, so the output is bit-identical.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

struct EaRead {
    int64_t kmerOff, evOff, traceOff, bandOff, pairOff;     // into EaParams.kmerId / ev (and lastCol) / trace / llE, llK / pairs
    int32_t nKmers, nEvents;
    double scale, shift, var, logInvVar;
    double lpSkip, lpStay, lpStep, lpTrim, eventsPerKmer;
};

struct EaParams {
    const EaRead *reads;
    const int32_t *kmerId;
    const float *ev;
    const double *table;      // K x 5 match table as loaded (level mean, level sd, ...)
    const double *c1;         // K: log_inv_sqrt_2pi - log(level sd), evaluated on the host
    uint8_t *trace;           // per read nBands x 100
    int32_t *llE, *llK;       // per read nBands: lower-left corner of every band
    double *lastCol;          // per read nEvents: score of cell (event, last kmer), -inf where not filled
    int32_t *pairs;           // per read 2 x (nEvents + nKmers): (kmer, event), written from the END backwards
    double *qc;               // per read 4: aligned pairs, sum of emissions, max gap, spanned
};

constexpr int kEaBand = 100;

// emissions_signal_strawManGetKmerEventMatchProbWithDescaling_MeanOnly, operation order kept:
//   x = (event + var * mu - scale * mu - shift) / var;  a = (x - mu) / sigma;
//   log(1 / var) + ((log_inv_sqrt_2pi - log(sigma)) + (-0.5 * a * a))
__device__ __forceinline__ double ea_emission(const EaParams &P, const EaRead &R, int32_t kid, double ev) {
    const double mu = P.table[(int64_t) kid * 5], sigma = P.table[(int64_t) kid * 5 + 1];
    const double x = __ddiv_rn(__dsub_rn(__dsub_rn(__dadd_rn(ev, __dmul_rn(R.var, mu)), __dmul_rn(R.scale, mu)), R.shift), R.var);
    const double a = __ddiv_rn(__dsub_rn(x, mu), sigma);
    return __dadd_rn(R.logInvVar, __dadd_rn(P.c1[kid], __dmul_rn(__dmul_rn(-0.5, a), a)));
}

__global__ void __launch_bounds__(128) sa_event_align_kernel(const __grid_constant__ EaParams P) {
    __shared__ double ring[3][kEaBand];
    __shared__ float redS[128];
    __shared__ int32_t redE[128];
    const EaRead R = P.reads[blockIdx.x];
    const int o = threadIdx.x;
    const int32_t nK = R.nKmers, nE = R.nEvents;
    const int32_t *kid = P.kmerId + R.kmerOff;
    const float *ev = P.ev + R.evOff;
    uint8_t *trace = P.trace + R.traceOff;
    int32_t *llE = P.llE + R.bandOff, *llK = P.llK + R.bandOff;
    double *lastCol = P.lastCol + R.evOff;
    const int32_t nBands = (nE + 1) + (nK + 1);
    const double NEG = -INFINITY;
    for (int32_t e = o; e < nE; e += 128) lastCol[e] = NEG;
    // bands 0 and 1 (:984-999)
    int32_t e2 = kEaBand / 2 - 1, k2 = -1 - kEaBand / 2;      // corner of band b - 2
    int32_t e1 = e2 + 1, k1 = k2;                             // corner of band b - 1 (move down)
    if (o < kEaBand) {
        ring[0][o] = (o == -1 - k2) ? 0.0 : NEG;
        ring[1][o] = (o == e1 - 0) ? R.lpTrim : NEG;
        trace[o] = 0;
        trace[kEaBand + o] = (o == e1 - 0) ? 1 : 0;
    }
    if (o == 0) { llE[0] = e2; llK[0] = k2; llE[1] = e1; llK[1] = k1; }
    __syncthreads();
    for (int32_t b = 2; b < nBands; b++) {
        const double *prev = ring[(b - 1) % 3], *prev2 = ring[(b - 2) % 3];
        double *cur = ring[b % 3];
        const double ll = prev[0], ur = prev[kEaBand - 1];
        bool right;
        if (ll == NEG && ur == NEG) right = (b & 1) == 1;
        else right = ll < ur;                                  // Suzuki's rule
        const int32_t e0 = right ? e1 : e1 + 1, k0 = right ? k1 + 1 : k1;
        if (o == 0) { llE[b] = e0; llK[b] = k0; }
        if (o < kEaBand) {
            double val = NEG;
            uint8_t from = 0;
            const int32_t eIdx = e0 - o, kIdx = k0 + o;
            if (kIdx == -1) {                                  // trim state
                if (eIdx >= 0 && eIdx < nE) { val = __dmul_rn(R.lpTrim, (double) (eIdx + 1)); from = 1; }
            }
            int32_t lo = max(max(0 - k0, e0 - (nE - 1)), 0);
            int32_t hi = min(min(nK - k0, e0 + 1), kEaBand);
            if (o >= lo && o < hi) {
                const int32_t oUp = e1 - (eIdx - 1), oLeft = (kIdx - 1) - k1, oDiag = (kIdx - 1) - k2;
                const float up = (oUp >= 0 && oUp < kEaBand) ? (float) prev[oUp] : -INFINITY;
                const float left = (oLeft >= 0 && oLeft < kEaBand) ? (float) prev[oLeft] : -INFINITY;
                const float diag = (oDiag >= 0 && oDiag < kEaBand) ? (float) prev2[oDiag] : -INFINITY;
                const double lp = ea_emission(P, R, kid[kIdx], (double) ev[eIdx]);
                const float sD = (float) __dadd_rn(__dadd_rn((double) diag, R.lpStep), lp);
                const float sU = (float) __dadd_rn(__dadd_rn((double) up, R.lpStay), lp);
                const float sL = (float) __dadd_rn((double) left, R.lpSkip);
                float mx = sD;
                from = 0;
                mx = sU > mx ? sU : mx;
                from = mx == sU ? 1 : from;
                mx = sL > mx ? sL : mx;
                from = mx == sL ? 2 : from;
                val = (double) mx;
                if (kIdx == nK - 1) lastCol[eIdx] = val;
            }
            cur[o] = val;
            trace[(int64_t) b * kEaBand + o] = from;
        }
        e2 = e1; k2 = k1; e1 = e0; k1 = k0;
        __syncthreads();
    }
    // best end cell in the last kmer's column (:1136-1148): the FIRST event with the strictly largest float score
    {
        float best = -INFINITY;
        int32_t bestE = 0x7fffffff;
        for (int32_t e = o; e < nE; e += 128) {
            const int32_t bi = (e + 1) + nK;                   // band of (e, nK - 1)
            const int32_t off = llE[bi] - e;
            if (off >= 0 && off < kEaBand) {
                const float s = (float) __dadd_rn(lastCol[e], __dmul_rn((double) (nE - e), R.lpTrim));
                if (s > best) { best = s; bestE = e; }
            }
        }
        redS[o] = best; redE[o] = bestE;
        __syncthreads();
        for (int st = 64; st > 0; st >>= 1) {
            if (o < st) {
                const float s2 = redS[o + st];
                const int32_t x2 = redE[o + st];
                if (s2 > redS[o] || (s2 == redS[o] && x2 < redE[o])) { redS[o] = s2; redE[o] = x2; }
            }
            __syncthreads();
        }
    }
    if (o != 0) return;
    // backtrack (:1157-1196)
    int32_t eCur = (redS[0] > -INFINITY) ? redE[0] : 0, kCur = nK - 1;
    int32_t *pairs = P.pairs + R.pairOff;
    const int64_t cap = (int64_t) nE + nK;
    int64_t n = 0;
    double sumEm = 0.0;
    int32_t gap = 0, maxGap = 0, firstK = -1, lastK = -1;
    while (kCur >= 0 && eCur >= 0 && n < cap) {
        pairs[2 * (cap - 1 - n)] = kCur;
        pairs[2 * (cap - 1 - n) + 1] = eCur;
        if (n == 0) lastK = kCur;
        firstK = kCur;
        n++;
        sumEm = __dadd_rn(sumEm, ea_emission(P, R, kid[kCur], (double) ev[eCur]));
        const int32_t bi = (eCur + 1) + (kCur + 1);
        const int32_t off = llE[bi] - eCur;
        if (off < 0 || off >= kEaBand) break;                   // (left the band: the reference reads outside its row here)
        const uint8_t from = trace[(int64_t) bi * kEaBand + off];
        if (from == 0) { kCur--; eCur--; gap = 0; }
        else if (from == 1) { eCur--; gap = 0; }
        else { kCur--; gap++; maxGap = max(gap, maxGap); }
    }
    double *qc = P.qc + (int64_t) blockIdx.x * 4;
    qc[0] = (double) n; qc[1] = sumEm; qc[2] = (double) maxGap; qc[3] = (firstK == 0 && lastK == nK - 1) ? 1.0 : 0.0;
}
