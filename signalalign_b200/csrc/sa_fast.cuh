// sa_fast.cuh — what the band kernels (sa_cta.cuh: register-resident, windows up to 256 columns; sa_wide.cuh: shared-
// memory resident, any width) share: the per-item / per-launch parameter blocks, the on-the-fly band iterator and the
// geometry pass that classifies every work item before the sweeps are launched.
//
// Eligibility for the band kernels (decided by sa_geometry_kernel + host): un-ambiguous reference (one path per cell),
// Gaussian MeanOnly or HDP emission, uniform X-gap emission, GAP_SWITCH_TO_X == -inf (always true for models loaded from
// a .model file).  Everything else runs in the general kernel (sa_general.cuh).
//
// Arithmetic is FP64 with the reference's cubic logAdd (sa_device.cuh); accumulation order per cell is the
// reference's (see sa_general.cuh header), so the integer outputs agree with the general kernel / the oracle.
#pragma once
#include "sa_device.cuh"
#include "sa_internal.h"

struct FItem {
    int64_t evOff;      // event means of the region
    int64_t kmerOff;    // int32 kmer index per reference position (lX entries)
    int64_t anchorOff;  // int32 (x, y) pairs, region-relative
    int64_t tbOff;      // traceback diagonals T_0 < T_1 < ... (int32), filled by sa_geometry_kernel
    int64_t pairOff, pairCap;
    int32_t nAnchors, nTbCap;
    int32_t lX, lY;
    int32_t flags;      // bit0 ragged left, bit1 ragged right
    int32_t pad;
    double scale, shift, invVar, logInvVar;
    double var;
    int64_t fullOff;    // split mode: offset (doubles) of the item's full-length row buffer in FParams.scratch
    int64_t assignOff, assignCap;   // HDP expectations: the item's slot in FParams.assignK / assignE
};

struct FBlock {          // split mode: one traceback block of an item
    int32_t item;
    int32_t d;           // the traceback starts on this diagonal
    int32_t tracedBackTo, tracedBackFrom;
    int64_t pairOff, pairCap;
};

struct FGeo {            // per item, written by sa_geometry_kernel
    int64_t cells;
    int32_t maxWidth, nTb, maxRows;
    int32_t maxSpan;     // max over d of the columns the diagonals d-2, d-1, d touch (hull of their bands)
    int32_t monotone;    // band never moves backwards in x
    int32_t pad;
    int64_t maxBlockDoubles;   // wide kernel: doubles of forward rows one traceback block keeps alive (compact rows)
};

struct FParams {
    const FItem *items;
    int32_t nItems;
    int32_t *counter;
    const int32_t *order;    // items in processing order (heaviest first), nOrder entries
    int32_t nOrder;
    const double *evMean;
    const int32_t *kmerIdx;
    const int32_t *anchors;
    int32_t *tb;
    FGeo *geo;
    const double *ft;        // K * 5: mu, 1/sdMatch, -0.9189.. - log sdMatch, 1/sdGapY, -0.9189.. - log sdGapY
    double trans[9];
    double eX;               // uniform X-gap emission (log 0.1)
    double threshold;
    double logThrGate;       // log(threshold) - 1e-6: below it exp() cannot reach the threshold (-inf if threshold <= 0)
    int32_t minDiags, traceBack, expansion;
    int32_t rowsCap;         // rows per ring
    double *scratch;         // per warp: rowsCap * rowStride doubles
    sa_pair_t *pairs;
    int64_t *pairCount;
    double *expOut;          // expectations: 9 transition sums (from * 3 + to) + likelihood per item
    // HDP emission (band kernel <.., HD = true>): uniform grid, posterior predictive and spline slopes per row, kmer -> row
    const double *grid, *pp, *slopes;
    const int32_t *kmerToRow;
    int32_t G;
    // split mode
    const FBlock *blocks;
    int32_t nBlocks;
    int64_t *blockPairCount;
    // HDP expectations: (kmer position, event) of every match term >= threshold (reference continuousHmm.c, the
    // assignments of cell_signal_updateExpectationsAndAssignments), in the reference's order
    int32_t *assignK, *assignE;
    int64_t *assignCount;
    double gridDx, gridInvDx; // grid[1] - grid[0] and its reciprocal
    // wide kernel (sa_wide.cuh): window slots, per-CTA row ring (scratch + blockIdx * wRingDoubles) and header ring
    int32_t wNSW, wRowsCap;
    int64_t wRingDoubles;
    int4 *wHdr;
    double *wFold;            // per-CTA 2 * wNSW doubles when the fold terms do not fit shared memory, else nullptr
    int32_t wPlanes;          // planes of a full stored row (3, HDP 4); geometry pass
    int32_t wAllFull;         // expectations: every row keeps all planes
    int32_t wGeoAlways;       // size the wide kernel's row ring for every item, not only those wider than 254 columns
    double *wRings;           // per-CTA 6 * wNSW doubles: emission rings of the window classes too large for shared memory
    double *wOff;             // FP32 wide kernel: per-CTA forward offsets, one per 16 diagonals
    int32_t wOffCap;
};

// ---------------------------------------------------------------------------------------------- band iterator
// band_construct (reference impl/pairwiseAligner.c:195-246) evaluated on the fly from the anchors, one diagonal
// per call; all lanes run it redundantly (uniform values, broadcast loads).
struct BandIt {
    const int32_t *anch;
    int32_t nA, ai, lX, lY, e;
    int32_t xay, pxay, pxmy, nxay, nxmy, xL, yL, xU, yU;
    __device__ __forceinline__ void init(const int32_t *a, int32_t n, int32_t lx, int32_t ly, int32_t ex) {
        anch = a; nA = n; lX = lx; lY = ly; e = ex; ai = 0;
        xay = pxay = pxmy = nxay = nxmy = xL = yL = xU = yU = 0;
    }
    static __device__ __forceinline__ int32_t clampc(int32_t z, int32_t l) { return z < 0 ? 0 : (z > l ? l : z); }
    // band of diagonal `xay` as the inclusive column range [xlo, xhi]; then advances to xay + 1
    __device__ __forceinline__ void next(int32_t &xlo, int32_t &xhi) {
        int32_t l = xL - yL, r = xU - yU;
        if ((xay + l) & 1) l++;   // band_avoidOffByOne: both bounds move UP to the diagonal's parity
        if ((xay + r) & 1) r++;
        int32_t t;
        if ((t = xL - (xay + l) / 2) > 0) l += 2 * t;
        if ((t = (xay - l) / 2 - yL) > 0) l += 2 * t;
        if ((t = (xay + r) / 2 - xU) > 0) r -= 2 * t;
        if ((t = yU - (xay - r) / 2) > 0) r -= 2 * t;
        xlo = (xay + l) / 2;
        xhi = (xay + r) / 2;
        if (nxay == xay) {
            pxay = nxay; pxmy = nxmy;
            int32_t x = lX, y = lY;
            if (ai < nA) { x = anch[2 * ai] + 1; y = anch[2 * ai + 1] + 1; ai++; }
            nxay = x + y; nxmy = x - y;
            xL = clampc((pxay + pxmy - e) / 2, lX);
            yL = clampc((nxay - nxmy + e) / 2, lY);
            xU = clampc((nxay + nxmy + e) / 2, lX);
            yU = clampc((pxay - pxmy - e) / 2, lY);
        }
        xay++;
    }
    // position the iterator ON diagonal d0 >= 1, as if next() had been called for 0 .. d0 - 1: diagonal d lies between
    // the last anchor whose diagonal (x + 1) + (y + 1) is < d and the first one >= d (or the end point (lX, lY))
    __device__ __forceinline__ void seek(int32_t d0) {
        int32_t lo = 0, hi = nA;
        while (lo < hi) {
            const int32_t mid = (lo + hi) >> 1;
            if (anch[2 * mid] + anch[2 * mid + 1] + 2 >= d0) hi = mid; else lo = mid + 1;
        }
        int32_t px = 0, py = 0, x = lX, y = lY;
        if (lo > 0) { px = anch[2 * (lo - 1)] + 1; py = anch[2 * (lo - 1) + 1] + 1; }
        ai = lo;
        if (ai < nA) { x = anch[2 * ai] + 1; y = anch[2 * ai + 1] + 1; ai++; }
        pxay = px + py; pxmy = px - py;
        nxay = x + y; nxmy = x - y;
        xL = clampc((pxay + pxmy - e) / 2, lX);
        yL = clampc((nxay - nxmy + e) / 2, lY);
        xU = clampc((nxay + nxmy + e) / 2, lX);
        yU = clampc((pxay - pxmy - e) / 2, lY);
        xay = d0;
    }
};

// One WARP per item: records what the host needs to size buffers and pick the kernel class, and the traceback
// diagonals the band kernels follow (getPosteriorProbsWithBanding's traceback condition, reference
// impl/pairwiseAligner.c:1493-1497, depends on geometry only).  The diagonals 0 .. lX+lY are cut into 32 contiguous
// ranges, one per lane; a lane enters its range with BandIt::seek and walks it incrementally.  The traceback list is
// periodic when no diagonal is wider than 2 * expansion + 1 (always, for banded nanopore reads); otherwise lane 0
// walks the band once more for it.
// (32-thread CTAs held to 32 registers: the pass of the NEXT batch runs next to the persistent band kernel of the one in
// flight, in the 1024 registers that kernel leaves free in every SM sub-partition.)
__global__ void __maxnreg__(40) sa_geometry_kernel(const FParams P) {
    const int it = blockIdx.x, lane = threadIdx.x;
    if (it >= P.nItems) return;
    const FItem &I = P.items[it];
    const int32_t nDiag = I.lX + I.lY;
    FGeo g;
    g.cells = 0; g.maxWidth = 0; g.nTb = 0; g.maxRows = 1; g.maxSpan = 0; g.monotone = 1; g.pad = 0; g.maxBlockDoubles = 0;
    if (nDiag == 0) {
        if (lane == 0) P.geo[it] = g;
        return;
    }
    const int32_t *anch = P.anchors + 2 * I.anchorOff;
    {
        const int32_t total = nDiag + 1, len = (total + 31) / 32;
        const int32_t d0 = lane * len, d1 = min(d0 + len, total);
        if (d0 < d1) {
            BandIt b;
            b.init(anch, I.nAnchors, I.lX, I.lY, P.expansion);
            const int32_t ds = d0 >= 2 ? d0 - 2 : 0;   // two warm-up diagonals: the predecessors of d0
            if (ds >= 1) b.seek(ds);
            int32_t pxlo = 0, pxhi = 0, ppxlo = 0, ppxhi = 0;
            for (int32_t d = ds; d < d1; d++) {
                int32_t xlo, xhi;
                b.next(xlo, xhi);
                if (d >= d0) {
                    const int32_t w = xhi - xlo + 1;
                    g.cells += w;
                    g.maxWidth = max(g.maxWidth, w);
                    if (d > 0 && (xlo < pxlo || xhi < pxhi)) g.monotone = 0;
                    int32_t lo = xlo, hi = xhi;
                    if (d >= 1) { lo = min(lo, pxlo); hi = max(hi, pxhi); }
                    if (d >= 2) { lo = min(lo, ppxlo); hi = max(hi, ppxhi); }
                    g.maxSpan = max(g.maxSpan, hi - lo + 1);
                }
                ppxlo = pxlo; ppxhi = pxhi; pxlo = xlo; pxhi = xhi;
            }
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        g.cells += __shfl_xor_sync(0xffffffffu, g.cells, o);
        g.maxWidth = max(g.maxWidth, __shfl_xor_sync(0xffffffffu, g.maxWidth, o));
        g.maxSpan = max(g.maxSpan, __shfl_xor_sync(0xffffffffu, g.maxSpan, o));
        g.monotone &= __shfl_xor_sync(0xffffffffu, g.monotone, o);
    }
    int32_t *tb = P.tb + I.tbOff;
    if (lane == 0) {
    int32_t tracedBackTo = 0;
    if (g.maxWidth <= P.expansion * 2 + 1) {
        for (;;) {   // the first diagonal >= tracedBackTo + minDiags, or the last one
            const int32_t d = min(tracedBackTo + P.minDiags, nDiag);
            const bool atEnd = d == nDiag;
            g.maxRows = max(g.maxRows, d - tracedBackTo + 1);
            if (g.nTb < I.nTbCap) tb[g.nTb] = d;
            g.nTb++;
            if (atEnd) break;
            tracedBackTo = d - (P.traceBack + 1);
        }
    } else {
        BandIt b;
        b.init(anch, I.nAnchors, I.lX, I.lY, P.expansion);
        for (int32_t d = 0; d <= nDiag; d++) {
            int32_t xlo, xhi;
            b.next(xlo, xhi);
            if (d == 0) continue;
            const bool atEnd = d == nDiag;
            const bool tbp = d >= tracedBackTo + P.minDiags && xhi - xlo + 1 <= P.expansion * 2 + 1;
            if (atEnd || tbp) {
                g.maxRows = max(g.maxRows, d - tracedBackTo + 1);
                if (g.nTb < I.nTbCap) tb[g.nTb] = d;
                g.nTb++;
                tracedBackTo = d - (atEnd ? 0 : P.traceBack + 1);
            }
        }
    }
    if (g.nTb > I.nTbCap) g.monotone = 0;  // cannot happen (capacity is nDiag / (minDiags - traceBack - 1) + 2); be safe
    }
    // Wide kernel: the doubles of compact forward rows each traceback block keeps alive — rows tracedBackTo .. T, the match
    // plane always, all planes where the traceback reads them: above tracedBackFrom, on every 10th posterior diagonal and
    // its predecessor, and the rows carried over from the previous block.  One block per lane, exact.
    g.nTb = __shfl_sync(0xffffffffu, g.nTb, 0);
    g.maxRows = __shfl_sync(0xffffffffu, g.maxRows, 0);
    g.monotone = __shfl_sync(0xffffffffu, g.monotone, 0);
    if ((g.maxSpan > 254 || P.wGeoAlways || !g.monotone) && g.nTb <= I.nTbCap) {
        __syncwarp();                          // lane 0's traceback list is visible to the warp
        int64_t need = 0;
        for (int32_t k = lane; k < g.nTb; k += 32) {
            const int32_t T = tb[k], prevT = k > 0 ? tb[k - 1] : 0;
            const int32_t to = k > 0 ? prevT - (P.traceBack + 1) : 0;
            const int32_t from = T - (T == nDiag ? 0 : P.traceBack + 1);
            BandIt b;
            b.init(anch, I.nAnchors, I.lX, I.lY, P.expansion);
            if (to >= 1) b.seek(to);
            int64_t sum = 0;
            for (int32_t d = to; d <= T; d++) {
                int32_t xlo, xhi;
                b.next(xlo, xhi);
                const int32_t m = (from - d) % 10;
                const bool full = P.wAllFull || d > from || m <= 1 || (k > 0 && d <= prevT) || d == 0;
                sum += (int64_t) (xhi - xlo + 1) * (full ? P.wPlanes : P.wPlanes - 2);
            }
            need = max(need, sum);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) need = max(need, __shfl_xor_sync(0xffffffffu, need, o));
        g.maxBlockDoubles = need;
    }
    if (lane == 0) P.geo[it] = g;
}
