// sa_fast.cuh — the FAST band kernel: one WARP per work item, live diagonals in registers, neighbour cells by
// warp shuffle, forward rows spilled to an HBM ring for the blocked traceback.
//
// Eligibility (decided by sa_geometry_kernel + host): un-ambiguous reference (one path per cell), Gaussian
// MeanOnly emission, uniform X-gap emission, GAP_SWITCH_TO_X == -inf (always true for models loaded from a
// .model file), band never wider than 32*C - 4 cells and never moving backwards in x.  Everything else runs in
// the general kernel (sa_general.cuh).
//
// Layout.  A warp owns NS = 32*C "slots"; slot s = x mod NS holds matrix column x for as long as x is inside the
// sliding window [wlo, wlo+NS).  Lane l owns slots l*C .. l*C+C-1.  Along a diagonal d the cell of column x is
// (x, y = d - x), so with x fixed per slot:
//      lower  (x-1, y  ) on d-1  -> slot s-1        middle (x-1, y-1) on d-2 -> slot s-1
//      upper  (x,   y-1) on d-1  -> slot s (same register)
// i.e. one neighbour exchange per step (the top cell of lane l-1), the per-column kmer constants stay in registers
// until the column leaves the band, and the event values stream through the slots systolically
// (ev(d)[x] = ev(d-1)[x-1]).  The backward pass is the mirror image (neighbours in slot s+1).
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
    int32_t maxSpan;     // max over d of xhi(d) - xlo(d-2) + 1: the columns three consecutive diagonals touch
    int32_t monotone;    // band never moves backwards in x
    int32_t pad;
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
__global__ void __maxnreg__(32) sa_geometry_kernel(const FParams P) {
    const int it = blockIdx.x, lane = threadIdx.x;
    if (it >= P.nItems) return;
    const FItem &I = P.items[it];
    const int32_t nDiag = I.lX + I.lY;
    FGeo g;
    g.cells = 0; g.maxWidth = 0; g.nTb = 0; g.maxRows = 1; g.maxSpan = 0; g.monotone = 1; g.pad = 0;
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
            int32_t pxlo = 0, pxhi = 0, ppxlo = 0;
            for (int32_t d = ds; d < d1; d++) {
                int32_t xlo, xhi;
                b.next(xlo, xhi);
                if (d >= d0) {
                    const int32_t w = xhi - xlo + 1;
                    g.cells += w;
                    g.maxWidth = max(g.maxWidth, w);
                    if (d > 0 && (xlo < pxlo || xhi < pxhi)) g.monotone = 0;
                    g.maxSpan = max(g.maxSpan, xhi - (d >= 2 ? ppxlo : (d >= 1 ? pxlo : xlo)) + 1);
                }
                ppxlo = pxlo; pxlo = xlo; pxhi = xhi;
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
    if (lane != 0) return;
    int32_t *tb = P.tb + I.tbOff;
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
    P.geo[it] = g;
}

// ---------------------------------------------------------------------------------------------- fast kernel
template <int C>
struct FastK {
    static constexpr int NS = 32 * C;          // slots per warp
    static constexpr int ROW = 3 * NS + 2;     // doubles per stored forward row: 2-double header + M, X, Y planes

    struct Consts {  // per slot: emission constants of the slot's current column
        double km[C];   // scale * mu + shift
        double iM[C];   // 1 / (var * sdMatch)
        double cM[C];   // log(1/var) - 0.9189.. - log sdMatch
        double iY[C];
        double cY[C];
    };

    static __device__ __forceinline__ double shflD(double v, int srcLane) {
        return __shfl_sync(0xffffffffu, v, srcLane);
    }

    // column currently held by slot s for a window starting at wlo
    static __device__ __forceinline__ int32_t slotX(int32_t s, int32_t wlo) { return wlo + ((s - wlo) & (NS - 1)); }

    static __device__ __forceinline__ void loadConsts(const FParams &P, const FItem &I, int32_t x, double &km,
                                                      double &iM, double &cM, double &iY, double &cY) {
        // column 0 is the NULL kmer (every emission -inf); columns 1..lX read the table
        if (x >= 1 && x <= I.lX) {
            const double *t = P.ft + (int64_t) P.kmerIdx[I.kmerOff + x - 1] * 5;
            km = fma(I.scale, t[0], I.shift);
            iM = t[1] * I.invVar;
            cM = t[2] + I.logInvVar;
            iY = t[3] * I.invVar;
            cY = t[4] + I.logInvVar;
        } else {
            km = 0.0; iM = 0.0; iY = 0.0; cM = SA_NEG_INF; cY = SA_NEG_INF;
        }
    }
    static __device__ __forceinline__ double loadEv(const FParams &P, const FItem &I, int32_t y) {
        return (y >= 1 && y <= I.lY) ? P.evMean[I.evOff + y - 1] : 0.0;
    }

    // (re)load every slot for a window starting at wlo, events for diagonal d
    static __device__ __forceinline__ void fillWindow(const FParams &P, const FItem &I, int lane, int32_t wlo, int32_t d,
                                                      Consts &k, double (&ev)[C]) {
#pragma unroll
        for (int j = 0; j < C; j++) {
            int32_t x = slotX(lane * C + j, wlo);
            loadConsts(P, I, x, k.km[j], k.iM[j], k.cM[j], k.iY[j], k.cY[j]);
            ev[j] = loadEv(P, I, d - x);
        }
    }

    // forward: advance the window so that it covers [.., xhi]; slots that wrap get a new column
    static __device__ __forceinline__ void slideUp(const FParams &P, const FItem &I, int lane, int32_t &wlo, int32_t xhi,
                                                   int32_t d, Consts &k, double (&ev)[C], double (&a)[3][C],
                                                   double (&b)[3][C]) {
        while (xhi > wlo + NS - C) {   // keep the top C-1 columns of the window inactive (column order == lane order)
            const int32_t s = wlo & (NS - 1), xn = wlo + NS;
            double nkm, niM, ncM, niY, ncY;
            loadConsts(P, I, xn, nkm, niM, ncM, niY, ncY);   // uniform address: one broadcast load per value
            const double nev = loadEv(P, I, d - xn);
#pragma unroll
            for (int j = 0; j < C; j++) {   // select form (no conditional stores: keeps the arrays in registers)
                const bool hit = lane * C + j == s;
                k.km[j] = hit ? nkm : k.km[j]; k.iM[j] = hit ? niM : k.iM[j]; k.cM[j] = hit ? ncM : k.cM[j];
                k.iY[j] = hit ? niY : k.iY[j]; k.cY[j] = hit ? ncY : k.cY[j];
                ev[j] = hit ? nev : ev[j];
#pragma unroll
                for (int t = 0; t < 3; t++) {
                    a[t][j] = hit ? SA_NEG_INF : a[t][j];
                    b[t][j] = hit ? SA_NEG_INF : b[t][j];
                }
            }
            wlo++;
        }
    }

    static __device__ __forceinline__ void emissions(const Consts &k, int j, double ev, double &eM, double &eY) {
        double d0 = ev - k.km[j];
        double aM = d0 * k.iM[j], aY = d0 * k.iY[j];
        eM = fma(-0.5 * aM, aM, k.cM[j]);
        eY = fma(-0.5 * aY, aY, k.cY[j]);
    }

    // f(d) from f1 = f(d-1) and f2 = f(d-2); the result overwrites f2.  nb1 / nb2: top cell of lane-1 on d-1 / d-2.
    static __device__ __forceinline__ void forwardStep(const FParams &P, int lane, int32_t wlo, int32_t xlo, int32_t xhi,
                                                       const Consts &k, const double (&ev)[C], double (&f1)[3][C],
                                                       double (&f2)[3][C], double (&nb1)[3], const double (&nb2)[3]) {
        const double *T = P.trans;
        const int src = (lane + 31) & 31;
        nb1[0] = shflD(f1[0][C - 1], src); nb1[1] = shflD(f1[1][C - 1], src); nb1[2] = shflD(f1[2][C - 1], src);
        const double cOX = P.eX + T[T_GAP_OPEN_X], cEX = P.eX + T[T_GAP_EXTEND_X];
#pragma unroll
        for (int j = C - 1; j >= 0; j--) {
            double lM = j > 0 ? f1[0][j - 1] : nb1[0], lX = j > 0 ? f1[1][j - 1] : nb1[1];
            double mM = j > 0 ? f2[0][j - 1] : nb2[0], mX = j > 0 ? f2[1][j - 1] : nb2[1],
                   mY = j > 0 ? f2[2][j - 1] : nb2[2];
            double uM = f1[0][j], uY = f1[2][j];
            double eM, eY;
            emissions(k, j, ev[j], eM, eY);
            // stateMachine3_cellCalculate order: first contribution lands on -inf (logAdd(-inf, v) == v exactly)
            double X = sa_logAdd(lM + cOX, lX + cEX);
            double M = sa_logAdd(sa_logAdd(mM + (eM + T[T_MATCH_CONTINUE]), mX + (eM + T[T_MATCH_FROM_GAP_X])),
                                 mY + (eM + T[T_MATCH_FROM_GAP_Y]));
            double Y = sa_logAdd(uM + (eY + T[T_GAP_OPEN_Y]), uY + (eY + T[T_GAP_EXTEND_Y]));
            int32_t x = slotX(lane * C + j, wlo);
            bool act = x >= xlo && x <= xhi;
            f2[0][j] = act ? M : SA_NEG_INF;
            f2[1][j] = act ? X : SA_NEG_INF;
            f2[2][j] = act ? Y : SA_NEG_INF;
        }
    }

    static __device__ __forceinline__ void storeRow(double *row, int lane, int32_t xlo, int32_t xhi, bool full,
                                                    const double (&f)[3][C]) {
        if (lane == 0) {
            int2 h = make_int2(xlo, xhi);
            *reinterpret_cast<int2 *>(row) = h;
        }
        double *m = row + 2 + lane * C;
#pragma unroll
        for (int j = 0; j < C; j += 2) {
            *reinterpret_cast<double2 *>(m + j) = make_double2(f[0][j], f[0][j + 1]);
            if (full) {
                *reinterpret_cast<double2 *>(m + NS + j) = make_double2(f[1][j], f[1][j + 1]);
                *reinterpret_cast<double2 *>(m + 2 * NS + j) = make_double2(f[2][j], f[2][j + 1]);
            }
        }
    }

    // events move one slot up per forward diagonal; the window's lowest column reads its event from memory
    static __device__ __forceinline__ void shiftEvUp(const FParams &P, const FItem &I, int lane, int32_t wlo, int32_t d,
                                                     double (&ev)[C]) {
        double in = shflD(ev[C - 1], (lane + 31) & 31);
#pragma unroll
        for (int j = C - 1; j > 0; j--) ev[j] = ev[j - 1];
        ev[0] = in;
        const int32_t s = wlo & (NS - 1);
        const double fresh = loadEv(P, I, d - wlo);
#pragma unroll
        for (int j = 0; j < C; j++) ev[j] = (lane * C + j == s) ? fresh : ev[j];
    }
    static __device__ __forceinline__ void shiftEvDown(const FParams &P, const FItem &I, int lane, int32_t wlo, int32_t d,
                                                       double (&ev)[C]) {
        double in = shflD(ev[0], (lane + 1) & 31);
#pragma unroll
        for (int j = 0; j < C - 1; j++) ev[j] = ev[j + 1];
        ev[C - 1] = in;
        const int32_t xt = wlo + NS - 1, s = xt & (NS - 1);
        const double fresh = loadEv(P, I, d - xt);
#pragma unroll
        for (int j = 0; j < C; j++) ev[j] = (lane * C + j == s) ? fresh : ev[j];
    }

    // backward: lower the window so that it covers [xlo, ..]
    static __device__ __forceinline__ void slideDown(const FParams &P, const FItem &I, int lane, int32_t &wlo, int32_t xlo,
                                                     int32_t d, Consts &k, double (&ev)[C], double (&pa)[2][C],
                                                     double (&pb)[2][C], double (&q)[3][C]) {
        while (xlo < wlo) {
            wlo--;
            const int32_t s = wlo & (NS - 1);
            double nkm, niM, ncM, niY, ncY;
            loadConsts(P, I, wlo, nkm, niM, ncM, niY, ncY);
            const double nev = loadEv(P, I, d - wlo);
#pragma unroll
            for (int j = 0; j < C; j++) {
                const bool hit = lane * C + j == s;
                k.km[j] = hit ? nkm : k.km[j]; k.iM[j] = hit ? niM : k.iM[j]; k.cM[j] = hit ? ncM : k.cM[j];
                k.iY[j] = hit ? niY : k.iY[j]; k.cY[j] = hit ? ncY : k.cY[j];
                ev[j] = hit ? nev : ev[j];
                pa[0][j] = hit ? SA_NEG_INF : pa[0][j]; pb[0][j] = hit ? SA_NEG_INF : pb[0][j];
                pa[1][j] = hit ? 0.0 : pa[1][j]; pb[1][j] = hit ? 0.0 : pb[1][j];
                q[0][j] = hit ? SA_NEG_INF : q[0][j]; q[1][j] = hit ? SA_NEG_INF : q[1][j];
                q[2][j] = hit ? 0.0 : q[2][j];
            }
        }
    }

    // b(d) by gathering (see sa_general.cuh for the order argument).
    //   p1 = {M, eM} of d+1, p2 = {M, eM} of d+2 (overwritten with d), q = {X, Y, eY} of d+1 (overwritten with d)
    static __device__ __forceinline__ void backwardStep(const FParams &P, int lane, int32_t wlo, int32_t xlo, int32_t xhi,
                                                        const Consts &k, const double (&ev)[C], double (&p1)[2][C],
                                                        double (&p2)[2][C], double (&q)[3][C]) {
        const double *T = P.trans;
        const int src = (lane + 1) & 31;
        // neighbours in slot s+1: bottom cell of lane+1
        double nA_M = shflD(p2[0][0], src), nA_e = shflD(p2[1][0], src), nL_X = shflD(q[0][0], src);
        const double cOX = P.eX + T[T_GAP_OPEN_X], cEX = P.eX + T[T_GAP_EXTEND_X];
#pragma unroll
        for (int j = 0; j < C; j++) {
            double aM = j < C - 1 ? p2[0][j + 1] : nA_M, aE = j < C - 1 ? p2[1][j + 1] : nA_e;
            double lX = j < C - 1 ? q[0][j + 1] : nL_X;
            double uY = q[1][j], uE = q[2][j];
            double M = sa_logAdd(sa_logAdd(aM + (aE + T[T_MATCH_CONTINUE]), uY + (uE + T[T_GAP_OPEN_Y])), lX + cOX);
            double X = sa_logAdd(aM + (aE + T[T_MATCH_FROM_GAP_X]), lX + cEX);
            double Y = sa_logAdd(aM + (aE + T[T_MATCH_FROM_GAP_Y]), uY + (uE + T[T_GAP_EXTEND_Y]));
            double eM, eY;
            emissions(k, j, ev[j], eM, eY);
            int32_t x = slotX(lane * C + j, wlo);
            bool act = x >= xlo && x <= xhi;
            p2[0][j] = act ? M : SA_NEG_INF;
            p2[1][j] = eM;
            q[0][j] = act ? X : SA_NEG_INF;
            q[1][j] = act ? Y : SA_NEG_INF;
            q[2][j] = eY;
        }
    }

    // sequential logAdd fold of the per-slot values v[] in column order (= ascending xmy), skipping the cells that
    // cannot change the accumulator: logAdd(acc, v) == acc whenever v <= acc - 7.5, and acc >= every value folded
    // so far, so v <= prefixMax - 7.6 is an exact no-op.
    static __device__ __forceinline__ double orderedFold(const double (&v)[C], int lane, int32_t wlo) {
        const int l0 = (wlo & (NS - 1)) / C;   // lane holding the window's first column
        const int rot = (lane - l0) & 31;      // this lane's position in column order
        // the window start may sit mid-lane: slots below it (j < j0) in lane l0 hold the window's LAST columns,
        // which are never active (band width <= NS - 4), i.e. -inf, so plain lane-major order is column order.
        double mx = v[0];
#pragma unroll
        for (int j = 1; j < C; j++) mx = fmax(mx, v[j]);
        // exclusive prefix max over lanes in rotated order
        double run = mx;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            double t = shflD(run, (lane - o) & 31);
            if (rot >= o) run = fmax(run, t);
        }
        double excl = shflD(run, (lane + 31) & 31);
        if (rot == 0) excl = SA_NEG_INF;
        unsigned cand[C];
        double pm = excl;
#pragma unroll
        for (int j = 0; j < C; j++) {
            bool c = v[j] > pm - 7.6;   // also false for v == -inf
            cand[j] = __ballot_sync(0xffffffffu, c);
            pm = fmax(pm, v[j]);
        }
        unsigned any = 0;
#pragma unroll
        for (int j = 0; j < C; j++) any |= cand[j];
        // rotate so that bit r corresponds to rotated position r
        unsigned anyR = l0 ? ((any >> l0) | (any << (32 - l0))) : any;
        double acc = SA_NEG_INF;
        while (anyR) {
            int r = __ffs(anyR) - 1;
            anyR &= anyR - 1;
            int ln = (r + l0) & 31;
#pragma unroll
            for (int j = 0; j < C; j++) {
                double t = shflD(v[j], ln);
                if ((cand[j] >> ln) & 1u) acc = sa_logAdd(acc, t);
            }
        }
        return acc;
    }

    // rows d (-> cur) and d-1 (-> prv) back into registers after a traceback; nbPrev = top cell of lane-1 on d-1
    // (the step d+1 re-fetches the neighbour of f(d) itself)
    static __device__ __forceinline__ void restoreForward(const FParams &P, const double *ring, int lane, int32_t d,
                                                          double (&cur)[3][C], double (&prv)[3][C], double (&nbPrev)[3]) {
        const double *r0 = ring + (int64_t) (d % P.rowsCap) * ROW + 2 + lane * C;
        const double *r1 = ring + (int64_t) ((d - 1) % P.rowsCap) * ROW + 2 + lane * C;
#pragma unroll
        for (int s3 = 0; s3 < 3; s3++)
#pragma unroll
            for (int j = 0; j < C; j++) {
                cur[s3][j] = r0[s3 * NS + j];
                prv[s3][j] = r1[s3 * NS + j];
            }
        const int src = (lane + 31) & 31;
        nbPrev[0] = shflD(prv[0][C - 1], src);
        nbPrev[1] = shflD(prv[1][C - 1], src);
        nbPrev[2] = shflD(prv[2][C - 1], src);
    }

    // total probability (every 10th call) + posterior match probabilities of diagonal d2.
    //   cur = {M, eM} of d2, nxt = {M, eM} of d2+1, q = {X, Y, eY} of d2
    static __device__ __forceinline__ void posteriorDiag(const FParams &P, const FItem &I, const double *ring, int lane,
                                                         int32_t d2, int32_t bw, int32_t bxlo, int32_t bxhi, bool newTotal,
                                                         const double (&cur)[2][C], const double (&nxt)[2][C],
                                                         const double (&q)[3][C], double &total, int64_t &nPairs) {
        const double *T = P.trans;
        const double *row = ring + (int64_t) (d2 % P.rowsCap) * ROW + 2;
        double fM[C];
#pragma unroll
        for (int j = 0; j < C; j += 2) {
            double2 t = *reinterpret_cast<const double2 *>(row + lane * C + j);
            fM[j] = t.x; fM[j + 1] = t.y;
        }
        if (newTotal) {
            // diagonalCalculationTotalProbability (reference pairwiseAligner.c:1335-1353)
            double v[C];
#pragma unroll
            for (int j = 0; j < C; j++) {
                double fX = row[NS + lane * C + j], fY = row[2 * NS + lane * C + j];
                double t = fM[j] + cur[0][j];
                t = sa_logAdd(t, fX + q[0][j]);
                t = sa_logAdd(t, fY + q[1][j]);
                v[j] = t;
            }
            double tot = orderedFold(v, lane, bw);
            // matches jumping over d2: middle-only forward step from f(d2-1) into the cells of d2+1
            const double *rowm = ring + (int64_t) ((d2 - 1) % P.rowsCap) * ROW + 2;
#pragma unroll
            for (int j = 0; j < C; j++) {
                int sp = (lane * C + j + NS - 1) & (NS - 1);   // slot of column x-1
                double gM = rowm[sp], gX = rowm[NS + sp], gY = rowm[2 * NS + sp];
                double eM = nxt[1][j];
                double m = sa_logAdd(sa_logAdd(gM + (eM + T[T_MATCH_CONTINUE]), gX + (eM + T[T_MATCH_FROM_GAP_X])),
                                     gY + (eM + T[T_MATCH_FROM_GAP_Y]));
                v[j] = m + nxt[0][j];
            }
            double tot2 = orderedFold(v, lane, bw);
            total = sa_logAdd(tot, tot2);
        }
        // diagonalCalculationPosteriorMatchProbs (reference pairwiseAligner.c:1355-1421)
        unsigned bal[C];
        double pp[C];
#pragma unroll
        for (int j = 0; j < C; j++) {
            int32_t x = slotX(lane * C + j, bw);
            int32_t y = d2 - x;
            pp[j] = exp((fM[j] + cur[0][j]) - total);
            bool hit = x >= bxlo && x <= bxhi && x > 0 && y > 0 && pp[j] >= P.threshold;
            bal[j] = __ballot_sync(0xffffffffu, hit);
        }
        unsigned any = 0;
#pragma unroll
        for (int j = 0; j < C; j++) any |= bal[j];
        if (any) {
            const int l0 = (bw & (NS - 1)) / C;
            const int rot = (lane - l0) & 31;
            unsigned below = rot ? (0xffffffffu >> (32 - rot)) : 0u;                // rotated positions < rot
            unsigned belowL = l0 ? ((below << l0) | (below >> (32 - l0))) : below;  // back to lane bits
            int before = 0, totalHits = 0;
#pragma unroll
            for (int j = 0; j < C; j++) {
                before += __popc(bal[j] & belowL);
                totalHits += __popc(bal[j]);
            }
#pragma unroll
            for (int j = 0; j < C; j++) {
                if ((bal[j] >> lane) & 1u) {
                    int64_t w = nPairs + before;
                    if (w < I.pairCap) {
                        double p1 = pp[j] > 1.0 ? 1.0 : pp[j];
                        int32_t x = slotX(lane * C + j, bw);
                        sa_pair_t o;
                        o.p = (int32_t) floor(p1 * 10000000.0);
                        o.x = x - 1; o.y = d2 - x - 1;
                        o.kmerIdx = P.kmerIdx[I.kmerOff + x - 1];
                        P.pairs[I.pairOff + w] = o;
                    }
                    before++;
                }
            }
            nPairs += totalHits;
        }
    }

    static __device__ void run(const FParams &P, const FItem &I, int it, double *ring, int lane) {
        const double *T = P.trans;
        const int32_t nDiag = I.lX + I.lY;
        int64_t nPairs = 0;
        if (nDiag == 0) {
            if (lane == 0) P.pairCount[it] = 0;
            return;
        }
        const int32_t *tbList = P.tb + I.tbOff;
        const int32_t nTb = P.geo[it].nTb;
        BandIt band;
        band.init(P.anchors + 2 * I.anchorOff, I.nAnchors, I.lX, I.lY, P.expansion);
        Consts k;
        double ev[C];
        double fa[3][C], fb[3][C], nbA[3], nbB[3];
        int32_t wlo = 0;
        // diagonal 0: the single cell (0, 0) with the start / ragged-start state (reference stateMachine.c:1134-1145)
        int32_t xlo, xhi;
        band.next(xlo, xhi);
        fillWindow(P, I, lane, wlo, 0, k, ev);
        {
            bool ragged = I.flags & 1;
#pragma unroll
            for (int j = 0; j < C; j++) {
                bool c0 = lane * C + j == 0;
                fa[0][j] = (c0 && !ragged) ? 0.0 : SA_NEG_INF;
                fa[1][j] = (c0 && ragged) ? 0.0 : SA_NEG_INF;
                fa[2][j] = (c0 && ragged) ? 0.0 : SA_NEG_INF;
                fb[0][j] = fb[1][j] = fb[2][j] = SA_NEG_INF;
            }
            nbA[0] = nbA[1] = nbA[2] = SA_NEG_INF;
            nbB[0] = nbB[1] = nbB[2] = SA_NEG_INF;
        }
        storeRow(ring, lane, xlo, xhi, true, fa);  // row of diagonal 0
        int32_t tracedBackTo = 0;
        int tbI = 0;
        int32_t nextT = nTb > 0 ? tbList[0] : nDiag;
        // which rows need the X / Y planes: the traceback recomputes the total probability on every 10th posterior
        // diagonal counted down from tracedBackFrom, from f(d) and f(d-1) (pairwiseAligner.c:1538, 1335-1353); the
        // rows above tracedBackFrom are carried into the next block whose schedule differs, so they are kept whole.
        int32_t curFrom = nextT - (nextT == nDiag ? 0 : P.traceBack + 1);
        for (int32_t d = 1; d <= nDiag; d++) {
            band.next(xlo, xhi);
            shiftEvUp(P, I, lane, wlo, d, ev);
            // f(d-1) lives in fa when d is odd, in fb when d is even
            if (d & 1) {
                slideUp(P, I, lane, wlo, xhi, d, k, ev, fa, fb);
                forwardStep(P, lane, wlo, xlo, xhi, k, ev, fa, fb, nbA, nbB);
            } else {
                slideUp(P, I, lane, wlo, xhi, d, k, ev, fb, fa);
                forwardStep(P, lane, wlo, xlo, xhi, k, ev, fb, fa, nbB, nbA);
            }
            {
                int32_t back = curFrom - d;
                bool full = back < 0 || (back % 10) <= 1;
                double *row = ring + (int64_t) (d % P.rowsCap) * ROW;
                if (d & 1) storeRow(row, lane, xlo, xhi, full, fb);
                else storeRow(row, lane, xlo, xhi, full, fa);
            }
            if (d != nextT) continue;
            // ------------------------------------------------------------------ traceback from diagonal d
            __syncwarp();
            const bool atEnd = d == nDiag;
            const int32_t tracedBackFrom = curFrom;
            {
                // the forward registers (fa, fb, nbA, nbB) are dead from here: rows d and d-1 are in the ring with all
                // three planes and are reloaded after the traceback; k / ev are reused for the backward sweep and
                // refilled afterwards, which keeps the live register set of the two sweeps disjoint.
                double pA[2][C], pB[2][C], q[3][C];
                Consts &kb = k;
                double (&evb)[C] = ev;
                int32_t bw = wlo;
                // end / ragged-end state on diagonal d (reference stateMachine.c:1147-1173)
                const bool rag = atEnd && (I.flags & 2);
                const double e0 = rag ? (T[T_GAP_OPEN_X] + T[T_GAP_OPEN_Y]) / 2.0 : T[T_MATCH_CONTINUE];
                const double e1 = rag ? T[T_GAP_EXTEND_X] : T[T_MATCH_FROM_GAP_X];
                const double e2 = rag ? T[T_GAP_EXTEND_Y] : T[T_MATCH_FROM_GAP_Y];
#pragma unroll
                for (int j = 0; j < C; j++) {
                    int32_t x = slotX(lane * C + j, bw);
                    bool act = x >= xlo && x <= xhi;
                    double eM, eY;
                    emissions(kb, j, evb[j], eM, eY);
                    pA[0][j] = act ? e0 : SA_NEG_INF; pA[1][j] = eM;
                    q[0][j] = act ? e1 : SA_NEG_INF; q[1][j] = act ? e2 : SA_NEG_INF; q[2][j] = eY;
                    pB[0][j] = SA_NEG_INF; pB[1][j] = 0.0;
                }
                double total = SA_NEG_INF;
                int32_t nPost = 0;
                int32_t bxlo = xlo, bxhi = xhi;
                // pA holds diagonal d2 when (d - d2) is even, pB when odd
                for (int32_t d2 = d; d2 > tracedBackTo; d2--) {
                    const bool inA = ((d - d2) & 1) == 0;
                    if (d2 < d) {
                        const double *row = ring + (int64_t) (d2 % P.rowsCap) * ROW;
                        int2 h = *reinterpret_cast<const int2 *>(row);
                        bxlo = h.x; bxhi = h.y;
                        shiftEvDown(P, I, lane, bw, d2, evb);
                        slideDown(P, I, lane, bw, bxlo, d2, kb, evb, pA, pB, q);
                        if (inA) backwardStep(P, lane, bw, bxlo, bxhi, kb, evb, pB, pA, q);
                        else backwardStep(P, lane, bw, bxlo, bxhi, kb, evb, pA, pB, q);
                    }
                    if (d2 > tracedBackFrom) continue;
                    const bool newTotal = nPost++ % 10 == 0;
                    if (inA) posteriorDiag(P, I, ring, lane, d2, bw, bxlo, bxhi, newTotal, pA, pB, q, total, nPairs);
                    else posteriorDiag(P, I, ring, lane, d2, bw, bxlo, bxhi, newTotal, pB, pA, q, total, nPairs);
                }
            }
            tracedBackTo = tracedBackFrom;
            tbI++;
            nextT = tbI < nTb ? tbList[tbI] : nDiag;
            curFrom = nextT - (nextT == nDiag ? 0 : P.traceBack + 1);
            if (atEnd) break;
            // restore the forward sweep: f(d) / f(d-1) from the ring, constants and events for the window at d
            if (d & 1) restoreForward(P, ring, lane, d, fb, fa, nbA);
            else restoreForward(P, ring, lane, d, fa, fb, nbB);
            fillWindow(P, I, lane, wlo, d, k, ev);
        }
        if (lane == 0) P.pairCount[it] = nPairs;
    }
};

template <int C>
__global__ void __launch_bounds__(128) sa_fast_kernel(const FParams P) {
    sa_logAdd_init();
    const int lane = threadIdx.x & 31;
    const int warp = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    double *ring = P.scratch + (int64_t) warp * P.rowsCap * FastK<C>::ROW;
    for (;;) {
        int n = 0;
        if (lane == 0) n = atomicAdd(P.counter, 1);
        n = __shfl_sync(0xffffffffu, n, 0);
        if (n >= P.nOrder) break;
        int it = P.order[n];
        FastK<C>::run(P, P.items[it], it, ring, lane);
        __syncwarp();
    }
}
