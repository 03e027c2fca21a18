// sa_general.cuh — the GENERAL band kernel: one CTA per work item, any band width, any number of paths per
// cell (ambiguous alphabets), every emission kind, alignment or expectations.
//
// It is the catch-all behind the fast register-resident kernel (sa_fast.cuh): items whose band does not fit
// the fast kernel's lane layout, or whose reference has ambiguity codes (multi-path HDCells), run here.
//
// Reference semantics implemented (paths relative to the signalAlign tree):
//   driver / blocked traceback ........ impl/pairwiseAligner.c:1450-1590
//   forward cell ........................ impl/stateMachine.c:1306-1437 via doTransitionForward (pairwiseAligner.c:852)
//   backward cell ....................... the reference SCATTERS (pairwiseAligner.c:866); here each target cell
//                                         GATHERS its contributions in exactly the order the scatter delivers them
//                                         (middle from xay+2, then upper-of (xay+1, xmy-1), then lower-of
//                                         (xay+1, xmy+1)), which keeps the non-associative logAdd bit-compatible
//   total probability ................... impl/pairwiseAligner.c:1335-1353 (sequential fold, same order)
//   posterior / expectations callbacks .. impl/pairwiseAligner.c:1355-1443, 914-968
#pragma once
#include "sa_device.cuh"
#include "sa_internal.h"

struct GItem {
    int64_t evOff;        // first event of the region in the event arrays
    int64_t pathOffOff;   // lX + 2 entries in pathOff (region-relative prefix of path counts; column 0 = NULL kmer)
    int64_t pathKmerOff;  // base of the region's concrete kmer indices
    int64_t diagOff;      // lX + lY + 1 entries in the band arrays
    int64_t pairOff, pairCap;
    int64_t assignOff, assignCap;
    int32_t lX, lY;
    int32_t flags;        // bit0 ragged left, bit1 ragged right
    int32_t pad;
    double scale, shift, var, invVar, logInvVar;
};

struct GParams {
    const GItem *items;      // indexed by global item id
    const int32_t *order;    // processing order (heaviest first): nItems global ids
    int32_t nItems;
    int32_t *counter;
    const int32_t *pathOff;
    const int32_t *pathKmer;
    const int32_t *pcCol;    // column of every path-cell (same indexing as pathKmer): spares the search for it (sa_mp_kernel)
    int32_t mpCap;           // sa_mp_kernel: path-cells per diagonal its shared-memory working set holds (0: it stays in global memory)
    const double *evMean, *evNoise, *evLogNoise;
    const int32_t *bandL, *bandW;
    const int64_t *bandOff;  // prefix of path-cells over the diagonals of the item
    // model
    const double *kt;     // K * SA_KT_STRIDE
    const double *gapX;   // K
    double trans[9];
    int32_t kind, A;
    int64_t suffixMod;    // A^(k-1)
    // floor(n / A) and floor(n / suffixMod) for kmer indices n < 2^31 as a multiply and a shift (round-up magic numbers:
    // M = ceil(2^sh / D), sh = 31 + ceil(log2 D)); path_checkLegal runs for every (path, neighbour path) pair of every cell
    uint32_t divA_M, divS_M;
    int32_t divA_sh, divS_sh;
    int64_t G;
    const double *grid, *pp, *slopes;
    const int32_t *kmerToRow;
    // alignment parameters
    double threshold;
    double logThrGate;    // log(threshold) - 1e-6: a cell whose f + b - total is below it cannot reach the threshold (no exp)
    int64_t minDiags, traceBack, expansion;
    // per-CTA scratch slab: [ring*3][3*maxDiagPC*3][2*maxWidth]
    double *scratch;
    int64_t slabDoubles, ringSize /* power of two, path-cells */, maxDiagPC, maxWidth;
    // outputs
    sa_pair_t *pairs;
    int64_t *pairCount;    // per item: number of pairs the item produces (may exceed pairCap -> host retries)
    double *expOut;        // per item: 9 transitions + likelihood
    int32_t *assignK, *assignE;
    int64_t *assignCount;
};

struct GCtx {
    const GParams *P;
    const GItem *I;
    const int32_t *pathOff;   // item-local
    const int32_t *pathKmer;
    const int32_t *pcCol;     // item-local (sa_mp_kernel)
    const int32_t *bandL, *bandW;
    const int64_t *bandOff;
    double *F, *B, *vals;
    double *E;                // sa_mp_kernel: match / gap-Y emission of every path-cell of the forward ring (2 per path-cell)
    int64_t ringMask;
    int64_t bStride;          // path-cells per diagonal buffer of B (maxDiagPC, or sa_mp_kernel's shared-memory capacity)
};

// (A non-inlined single copy of this function was tried against the instruction-cache stalls ncu shows: 5208 instead of
// 6776 SASS instructions, but 122 registers instead of 96 and no gain — 2.44 against 2.53 GCUPS on the config-2 shape.)
__device__ __forceinline__ double g_emit(const GCtx &c, int32_t ki, int64_t y, bool match) {
    const GParams &P = *c.P;
    if (ki < 0) return SA_NEG_INF;  // NULL kmer of matrix column 0
    double mean = y > 0 ? P.evMean[c.I->evOff + y - 1] : SA_NEG_INF;  // NULLEVENT, pairwiseAligner.c:325
    const double *t = P.kt + (int64_t) ki * SA_KT_STRIDE + ((match || P.kind == SA_EMIT_HDP) ? 0 : 6);
    double mu = t[0];
    if (P.kind == SA_EMIT_HDP) {  // stateMachine.c:527-554 -> hdp.c:2588-2612 -> hdp_math_utils.c:471-495
        double q = __ddiv_rn(__dsub_rn(__dsub_rn(__dadd_rn(mean, __dmul_rn(c.I->var, mu)), __dmul_rn(c.I->scale, mu)),
                                       c.I->shift), c.I->var);
        int64_t n = P.G;
        const double *x = P.grid;
        const double *yv = P.pp + (int64_t) P.kmerToRow[ki] * n;
        const double *s = P.slopes + (int64_t) P.kmerToRow[ki] * n;
        double v;
        if (q <= x[0]) v = __dsub_rn(yv[0], __dmul_rn(s[0], __dsub_rn(x[0], q)));
        else if (q >= x[n - 1]) v = __dadd_rn(yv[n - 1], __dmul_rn(s[n - 1], __dsub_rn(q, x[n - 1])));
        else {
            double dx = __dsub_rn(x[1], x[0]);
            int64_t il = (int64_t) __ddiv_rn(__dsub_rn(q, x[0]), dx), ir = il + 1;
            double dy = __dsub_rn(yv[ir], yv[il]);
            double a = __dsub_rn(__dmul_rn(s[il], dx), dy), b = __dsub_rn(dy, __dmul_rn(s[ir], dx));
            double tl = __ddiv_rn(__dsub_rn(q, x[il]), dx), tr = __dsub_rn(1.0, tl);
            v = __dadd_rn(__dadd_rn(__dmul_rn(tr, yv[il]), __dmul_rn(tl, yv[ir])),
                          __dmul_rn(__dmul_rn(tl, tr), __dadd_rn(__dmul_rn(a, tr), __dmul_rn(b, tl))));
        }
        double dens = v > 0.0 ? v : 0.0;
        return log(__dmul_rn(c.I->invVar, dens));
    }
    double sd = t[1], lsd = t[2];
    double xq = mean;
    if (P.kind != SA_EMIT_GAUSS_NOISE_SCALED)  // descale, stateMachine.c:344-348
        xq = __ddiv_rn(__dsub_rn(__dsub_rn(__dadd_rn(mean, __dmul_rn(c.I->var, mu)), __dmul_rn(c.I->scale, mu)),
                                 c.I->shift), c.I->var);
    double lp;
    if (sd == 0.0) lp = SA_NEG_INF;  // logGaussPdf, stateMachine.c:296-306
    else {
        double a = __ddiv_rn(__dsub_rn(xq, mu), sd);
        lp = __dadd_rn(__dsub_rn(-0.91893853320467267, lsd), __dmul_rn(__dmul_rn(-0.5, a), a));
    }
    if (P.kind == SA_EMIT_GAUSS_MEANONLY_DESCALE) return __dadd_rn(c.I->logInvVar, lp);
    // inverse-Gaussian noise term, stateMachine.c:285-294
    double noise = y > 0 ? P.evNoise[c.I->evOff + y - 1] : 0.0;
    double lnoise = y > 0 ? P.evLogNoise[c.I->evOff + y - 1] : SA_NEG_INF;
    if (P.kind == SA_EMIT_GAUSS_NOISE_DESCALE && noise == 0.0) noise = 0.000000001;
    double nm = t[3], nl = t[4], lnl = t[5];
    double a2 = __ddiv_rn(__dsub_rn(noise, nm), nm);
    double ln = __ddiv_rn(__dsub_rn(__dsub_rn(__dsub_rn(lnl, 1.8378770664093453), __dmul_rn(3.0, lnoise)),
                                    __ddiv_rn(__dmul_rn(__dmul_rn(nl, a2), a2), noise)), 2.0);
    return __dadd_rn(lp, ln);
}

__device__ __forceinline__ double g_gapX(const GCtx &c, int32_t ki) {
    if (c.P->kind == SA_EMIT_HDP) return -2.3025850929940455;  // literal in stateMachine3HDP_cellCalculate (:1394)
    return ki < 0 ? SA_NEG_INF : c.P->gapX[ki];                // emissions_kmer_getGapProb (:208-225)
}
__device__ __forceinline__ bool g_legal(const GCtx &c, int32_t kq, int32_t kp) {  // path_checkLegal (:595-621)
    if (kq < 0 || kp < 0) return true;
    const uint32_t pre = (uint32_t) (((uint64_t) (uint32_t) kp * c.P->divA_M) >> c.P->divA_sh);   // kp / A
    const uint32_t hi = (uint32_t) (((uint64_t) (uint32_t) kq * c.P->divS_M) >> c.P->divS_sh);    // kq / suffixMod
    return (uint32_t) kq - hi * (uint32_t) c.P->suffixMod == pre;                                  // kq % suffixMod
}

// geometry of diagonal d of the current item
struct GDiag {
    int32_t L, W;
    int64_t off;
    int32_t x0, y0;
};
__device__ __forceinline__ GDiag g_diag(const GCtx &c, int64_t d) {
    GDiag g;
    g.L = c.bandL[d]; g.W = c.bandW[d]; g.off = c.bandOff[d];
    g.x0 = (int32_t) ((d + g.L) / 2);
    g.y0 = (int32_t) ((d - g.L) / 2);
    return g;
}
// path-cell index (within diagonal) of the cell with coordinate xmy; -1 if outside the band
__device__ __forceinline__ int64_t g_cellPC(const GCtx &c, const GDiag &g, int64_t xmy, int32_t *x) {
    if (xmy < g.L || xmy > (int64_t) g.L + 2 * (int64_t) (g.W - 1)) return -1;
    int32_t ci = (int32_t) ((xmy - g.L) / 2);
    *x = g.x0 + ci;
    return c.pathOff[*x] - c.pathOff[g.x0];
}
__device__ __forceinline__ double *g_F(const GCtx &c, const GDiag &g, int64_t pc) {
    return c.F + ((g.off + pc) & c.ringMask) * 3;
}
__device__ __forceinline__ double *g_B(const GCtx &c, int64_t d, int64_t pc) {
    return c.B + ((d % 3) * c.bStride + pc) * 3;
}

// Work is dealt out per PATH-CELL, not per cell: thread j of a diagonal takes the j-th (cell, path) pair, so a warp that
// holds a multi-path column does not serialise that column's paths while its other lanes wait (the forward and backward
// values of a path depend only on the neighbour cells, never on the sibling paths of the same cell).  The cell of
// path-cell j is found by bisection over the column prefix; with `extra` paths beyond one per column in the diagonal the
// bracket is extra + 1 columns wide (0 steps where the window holds no ambiguity code).
__device__ __forceinline__ int32_t g_locate(const GCtx &c, const GDiag &g, int32_t base, int32_t nPC, int32_t j) {
    const int32_t extra = nPC - g.W;
    int32_t lo = j - extra > 0 ? j - extra : 0, hi = j < g.W - 1 ? j : g.W - 1;
    while (lo < hi) {   // largest ci whose first path-cell is <= j
        const int32_t mid = (lo + hi + 1) >> 1;
        if (c.pathOff[g.x0 + mid] - base <= j) lo = mid;
        else hi = mid - 1;
    }
    return lo;
}

__device__ void g_forwardDiag(const GCtx &c, int64_t d) {
    const double *T = c.P->trans;
    const bool swX = T[T_GAP_SWITCH_TO_X] > SA_NEG_INF || T[T_GAP_SWITCH_TO_X] != T[T_GAP_SWITCH_TO_X];
    GDiag g = g_diag(c, d), g1 = g_diag(c, d - 1), g2 = g1;
    bool has2 = d >= 2;
    if (has2) g2 = g_diag(c, d - 2);
    const int32_t pcBase = c.pathOff[g.x0], nPC = c.pathOff[g.x0 + g.W] - pcBase;
    for (int32_t j = threadIdx.x; j < nPC; j += blockDim.x) {
        const int32_t ci = g_locate(c, g, pcBase, nPC, j);
        int64_t xmy = (int64_t) g.L + 2 * ci;
        int32_t x = g.x0 + ci, y = g.y0 - ci, xl = 0, xm = 0, xu = 0;
        int64_t pc = c.pathOff[x] - pcBase;
        int32_t nP = c.pathOff[x + 1] - c.pathOff[x];
        int64_t pl = g_cellPC(c, g1, xmy - 1, &xl), pu = g_cellPC(c, g1, xmy + 1, &xu);
        int64_t pm = has2 ? g_cellPC(c, g2, xmy, &xm) : -1;
        const int32_t *kC = c.pathKmer + c.pathOff[x];
        {
            const int32_t p = j - (int32_t) pc;
            int32_t kp = kC[p];
            double M = SA_NEG_INF, X = SA_NEG_INF, Y = SA_NEG_INF;
            if (pl >= 0) {
                int32_t nQ = c.pathOff[xl + 1] - c.pathOff[xl];
                const int32_t *kQ = c.pathKmer + c.pathOff[xl];
                double eP = g_gapX(c, kp);
                for (int32_t q = 0; q < nQ; q++)
                    if (g_legal(c, kQ[q], kp)) {
                        const double *f = g_F(c, g1, pl + q);
                        X = sa_logAdd(X, __dadd_rn(f[0], __dadd_rn(eP, T[T_GAP_OPEN_X])));
                        X = sa_logAdd(X, __dadd_rn(f[1], __dadd_rn(eP, T[T_GAP_EXTEND_X])));
                        // GAP_SWITCH_TO_X is -inf in every model loaded from a .model file: logAdd(X, -inf) == X exactly
                        if (swX) X = sa_logAdd(X, __dadd_rn(f[2], __dadd_rn(eP, T[T_GAP_SWITCH_TO_X])));
                    }
            }
            if (pm >= 0) {
                int32_t nQ = c.pathOff[xm + 1] - c.pathOff[xm];
                const int32_t *kQ = c.pathKmer + c.pathOff[xm];
                double eP = g_emit(c, kp, y, true);
                for (int32_t q = 0; q < nQ; q++)
                    if (g_legal(c, kQ[q], kp)) {
                        const double *f = g_F(c, g2, pm + q);
                        M = sa_logAdd(M, __dadd_rn(f[0], __dadd_rn(eP, T[T_MATCH_CONTINUE])));
                        M = sa_logAdd(M, __dadd_rn(f[1], __dadd_rn(eP, T[T_MATCH_FROM_GAP_X])));
                        M = sa_logAdd(M, __dadd_rn(f[2], __dadd_rn(eP, T[T_MATCH_FROM_GAP_Y])));
                    }
            }
            if (pu >= 0) {
                double eP = g_emit(c, kp, y, false);
                for (int32_t q = 0; q < nP; q++)
                    if (kC[q] == kp) {  // same column -> same path list; stString_eq on the kmers
                        const double *f = g_F(c, g1, pu + q);
                        Y = sa_logAdd(Y, __dadd_rn(f[0], __dadd_rn(eP, T[T_GAP_OPEN_Y])));
                        Y = sa_logAdd(Y, __dadd_rn(f[2], __dadd_rn(eP, T[T_GAP_EXTEND_Y])));
                    }
            }
            double *o = g_F(c, g, pc + p);
            o[0] = M; o[1] = X; o[2] = Y;
        }
    }
}

// b[d] for every cell by gathering from b[d+1] / b[d+2] (which exist up to the traceback start diagonal `top`)
__device__ void g_backwardDiag(const GCtx &c, int64_t d, int64_t top) {
    const double *T = c.P->trans;
    const bool swX = T[T_GAP_SWITCH_TO_X] > SA_NEG_INF || T[T_GAP_SWITCH_TO_X] != T[T_GAP_SWITCH_TO_X];
    GDiag g = g_diag(c, d), g1 = g, g2 = g;
    bool has1 = d + 1 <= top, has2 = d + 2 <= top;
    if (has1) g1 = g_diag(c, d + 1);
    if (has2) g2 = g_diag(c, d + 2);
    const int32_t pcBase = c.pathOff[g.x0], nPC = c.pathOff[g.x0 + g.W] - pcBase;
    for (int32_t j = threadIdx.x; j < nPC; j += blockDim.x) {   // one path-cell per thread (see g_locate)
        const int32_t ci = g_locate(c, g, pcBase, nPC, j);
        int64_t xmy = (int64_t) g.L + 2 * ci;
        int32_t x = g.x0 + ci, y = g.y0 - ci, xa = 0, xu = 0, xl = 0;
        int64_t pc = c.pathOff[x] - pcBase;
        int32_t nQ = c.pathOff[x + 1] - c.pathOff[x];
        int64_t pa = has2 ? g_cellPC(c, g2, xmy, &xa) : -1;       // (x+1, y+1): this cell is its "middle"
        int64_t pu = has1 ? g_cellPC(c, g1, xmy - 1, &xu) : -1;   // (x,   y+1): this cell is its "upper"
        int64_t pl = has1 ? g_cellPC(c, g1, xmy + 1, &xl) : -1;   // (x+1, y  ): this cell is its "lower"
        const int32_t *kZ = c.pathKmer + c.pathOff[x];
        {
            const int32_t q = j - (int32_t) pc;
            int32_t kq = kZ[q];
            double M = SA_NEG_INF, X = SA_NEG_INF, Y = SA_NEG_INF;
            if (pa >= 0) {
                int32_t nP = c.pathOff[xa + 1] - c.pathOff[xa];
                const int32_t *kP = c.pathKmer + c.pathOff[xa];
                for (int32_t p = 0; p < nP; p++)
                    if (g_legal(c, kq, kP[p])) {
                        double eP = g_emit(c, kP[p], y + 1, true);
                        double a = g_B(c, d + 2, pa + p)[0];
                        M = sa_logAdd(M, __dadd_rn(a, __dadd_rn(eP, T[T_MATCH_CONTINUE])));
                        X = sa_logAdd(X, __dadd_rn(a, __dadd_rn(eP, T[T_MATCH_FROM_GAP_X])));
                        Y = sa_logAdd(Y, __dadd_rn(a, __dadd_rn(eP, T[T_MATCH_FROM_GAP_Y])));
                    }
            }
            if (pu >= 0) {
                double eP = g_emit(c, kq, y + 1, false);
                for (int32_t p = 0; p < nQ; p++)
                    if (kZ[p] == kq) {
                        double a = g_B(c, d + 1, pu + p)[2];
                        M = sa_logAdd(M, __dadd_rn(a, __dadd_rn(eP, T[T_GAP_OPEN_Y])));
                        Y = sa_logAdd(Y, __dadd_rn(a, __dadd_rn(eP, T[T_GAP_EXTEND_Y])));
                    }
            }
            if (pl >= 0) {
                int32_t nP = c.pathOff[xl + 1] - c.pathOff[xl];
                const int32_t *kP = c.pathKmer + c.pathOff[xl];
                for (int32_t p = 0; p < nP; p++)
                    if (g_legal(c, kq, kP[p])) {
                        double eP = g_gapX(c, kP[p]);
                        double a = g_B(c, d + 1, pl + p)[1];
                        M = sa_logAdd(M, __dadd_rn(a, __dadd_rn(eP, T[T_GAP_OPEN_X])));
                        X = sa_logAdd(X, __dadd_rn(a, __dadd_rn(eP, T[T_GAP_EXTEND_X])));
                        if (swX) Y = sa_logAdd(Y, __dadd_rn(a, __dadd_rn(eP, T[T_GAP_SWITCH_TO_X])));
                    }
            }
            double *o = g_B(c, d, pc + q);
            o[0] = M; o[1] = X; o[2] = Y;
        }
    }
}

// diagonalCalculationTotalProbability (pairwiseAligner.c:1335-1353); result broadcast through *sTotal (shared)
__device__ void g_totalProbability(const GCtx &c, int64_t d, int64_t top, double *sTotal) {
    const double *T = c.P->trans;
    GDiag g = g_diag(c, d);
    // per-cell totals of f[d] . b[d]
    for (int32_t ci = threadIdx.x; ci < g.W; ci += blockDim.x) {
        int32_t x = g.x0 + ci;
        int64_t pc = c.pathOff[x] - c.pathOff[g.x0];
        int32_t nP = c.pathOff[x + 1] - c.pathOff[x];
        const int32_t *kC = c.pathKmer + c.pathOff[x];
        double cell = SA_NEG_INF;
        for (int32_t p = 0; p < nP; p++)
            for (int32_t q = 0; q < nP; q++)
                if (kC[p] == kC[q]) {
                    const double *f = g_F(c, g, pc + p), *b = g_B(c, d, pc + q);
                    double t = __dadd_rn(f[0], b[0]);
                    t = sa_logAdd(t, __dadd_rn(f[1], b[1]));
                    t = sa_logAdd(t, __dadd_rn(f[2], b[2]));
                    cell = sa_logAdd(cell, t);
                }
        c.vals[ci] = cell;
    }
    // matches jumping over d: forward middle-only step from f[d-1] into the shape of b[d+1], dotted with b[d+1]
    bool hasMatch = d + 1 <= top && d - 1 >= 0;
    GDiag gb = g, gf = g;
    if (hasMatch) {
        gb = g_diag(c, d + 1);
        gf = g_diag(c, d - 1);
        for (int32_t ci = threadIdx.x; ci < gb.W; ci += blockDim.x) {
            int64_t xmy = (int64_t) gb.L + 2 * ci;
            int32_t x = gb.x0 + ci, y = gb.y0 - ci, xm = 0;
            int64_t pc = c.pathOff[x] - c.pathOff[gb.x0];
            int32_t nP = c.pathOff[x + 1] - c.pathOff[x];
            const int32_t *kC = c.pathKmer + c.pathOff[x];
            int64_t pm = g_cellPC(c, gf, xmy, &xm);
            double cell = SA_NEG_INF;
            for (int32_t p = 0; p < nP; p++) {
                double M = SA_NEG_INF;
                if (pm >= 0) {
                    int32_t nQ = c.pathOff[xm + 1] - c.pathOff[xm];
                    const int32_t *kQ = c.pathKmer + c.pathOff[xm];
                    double eP = g_emit(c, kC[p], y, true);
                    for (int32_t q = 0; q < nQ; q++)
                        if (g_legal(c, kQ[q], kC[p])) {
                            const double *f = g_F(c, gf, pm + q);
                            M = sa_logAdd(M, __dadd_rn(f[0], __dadd_rn(eP, T[T_MATCH_CONTINUE])));
                            M = sa_logAdd(M, __dadd_rn(f[1], __dadd_rn(eP, T[T_MATCH_FROM_GAP_X])));
                            M = sa_logAdd(M, __dadd_rn(f[2], __dadd_rn(eP, T[T_MATCH_FROM_GAP_Y])));
                        }
                }
                for (int32_t q = 0; q < nP; q++)
                    if (kC[p] == kC[q]) {
                        const double *b = g_B(c, d + 1, pc + q);
                        // cell_dotProduct with X = Y = -inf on the match diagonal: the two extra logAdds return t
                        double t = __dadd_rn(M, b[0]);
                        cell = sa_logAdd(cell, t);
                    }
            }
            c.vals[c.P->maxWidth + ci] = cell;
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        // The reference's sequential folds, both chains in flight.  In a band of at most 256 cells a value 7.6 or more
        // below the running maximum is an exact no-op (the accumulator stays within 0.1 of that maximum: each logAdd
        // returns at least max - 5e-4, and |x - y| >= 7.5 returns the larger operand unchanged), so it is skipped
        // without touching the dependent chain — as the band kernel's fold does (sa_cta.cuh, orderedFold2).
        // (any width: the accumulator never falls below the running maximum — logAdd returns at least its larger operand —,
        // which is the rule the wide kernel's fold applies to windows of up to 4096 cells with byte-identical results)
        const bool skipOk = true;
        const int32_t nB = hasMatch ? gb.W : 0, nMax = g.W > nB ? g.W : nB;
        const double *vb = c.vals + c.P->maxWidth;
        double tot = SA_NEG_INF, t2 = SA_NEG_INF, mx = SA_NEG_INF, mx2 = SA_NEG_INF;
        for (int32_t ci = 0; ci < nMax; ci++) {
            if (ci < g.W) {
                const double v = c.vals[ci];
                if (!(skipOk && v <= mx - 7.6)) tot = sa_logAdd(tot, v);
                mx = v > mx ? v : mx;
            }
            if (ci < nB) {
                const double v = vb[ci];
                if (!(skipOk && v <= mx2 - 7.6)) t2 = sa_logAdd(t2, v);
                mx2 = v > mx2 ? v : mx2;
            }
        }
        if (hasMatch) tot = sa_logAdd(tot, t2);
        *sTotal = tot;
    }
    __syncthreads();
}

// exclusive block scan of one int per thread; returns the thread's offset, *total = block sum
// (sWarp holds two sets of per-warp sums used alternately — `phase` counts the calls, identically in every thread —, so
// that one barrier per call is enough: a set is rewritten two calls, i.e. at least one barrier, after its last reader)
__device__ __forceinline__ int g_blockScan(int v, int *sWarpBase, int *total, int &phase) {
    int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nW = (blockDim.x + 31) >> 5;
    int inc = v;
    for (int o = 1; o < 32; o <<= 1) {
        int t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    int *sWarp = sWarpBase + 8 * (phase & 1);
    phase++;
    if (lane == 31) sWarp[w] = inc;
    __syncthreads();
    int base = 0, tot = 0;
    for (int i = 0; i < nW; i++) {
        if (i < w) base += sWarp[i];
        tot += sWarp[i];
    }
    *total = tot;
    return base + inc - v;
}

// diagonalCalculationPosteriorMatchProbs (pairwiseAligner.c:1355-1421), pairs written in the reference's order
__device__ void g_posteriorDiag(const GCtx &c, int64_t d, double total, int64_t *nPairs, int *sWarp, int &phase) {
    GDiag g = g_diag(c, d);
    for (int32_t base = 0; base < g.W; base += blockDim.x) {
        int32_t ci = base + threadIdx.x;
        int cnt = 0;
        int32_t x = 0, y = 0, nP = 0;
        int64_t pc = 0;
        if (ci < g.W) {
            x = g.x0 + ci; y = g.y0 - ci;
            if (x > 0 && y > 0) {
                pc = c.pathOff[x] - c.pathOff[g.x0];
                nP = c.pathOff[x + 1] - c.pathOff[x];
                const int32_t *kC = c.pathKmer + c.pathOff[x];
                for (int32_t fp = 0; fp < nP; fp++)
                    for (int32_t bp = 0; bp < nP; bp++)
                        if (kC[fp] == kC[bp]) {
                            double v = __dsub_rn(__dadd_rn(g_F(c, g, pc + fp)[0], g_B(c, d, pc + bp)[0]), total);
                            if (v >= c.P->logThrGate && exp(v) >= c.P->threshold) cnt++;
                        }
            }
        }
        int tot;
        int off = g_blockScan(cnt, sWarp, &tot, phase);
        if (cnt > 0) {
            const int32_t *kC = c.pathKmer + c.pathOff[x];
            int64_t w = *nPairs + off;
            for (int32_t fp = 0; fp < nP; fp++)
                for (int32_t bp = 0; bp < nP; bp++)
                    if (kC[fp] == kC[bp]) {
                        double v = __dsub_rn(__dadd_rn(g_F(c, g, pc + fp)[0], g_B(c, d, pc + bp)[0]), total);
                        double pp = v >= c.P->logThrGate ? exp(v) : 0.0;
                        if (v >= c.P->logThrGate && pp >= c.P->threshold) {
                            if (pp > 1.0) pp = 1.0;
                            if (w < c.I->pairCap) {
                                sa_pair_t o;
                                o.p = (int32_t) floor(__dmul_rn(pp, 10000000.0));
                                o.x = x - 1; o.y = y - 1; o.kmerIdx = kC[fp];
                                c.P->pairs[c.I->pairOff + w] = o;
                            }
                            w++;
                        }
                    }
        }
        *nPairs += tot;
    }
}

// diagonalCalculation_Expectations (pairwiseAligner.c:1423-1443): current = b[d], neighbours from f[d-1], f[d-2].
// `floorDiag` = tracedBackTo of the running block: forward diagonals below it have been freed by the reference
// (pairwiseAligner.c:1556-1569), so their cells are NULL to the callback.
__device__ void g_expectationsDiag(const GCtx &c, int64_t d, double total, int64_t floorDiag, double *acc9,
                                   int64_t *nAssign, int *sWarp, int &phase) {
    const double *T = c.P->trans;
    const bool hdp = c.P->kind == SA_EMIT_HDP;
    GDiag g = g_diag(c, d), g1 = g, g2 = g;
    bool has1 = d - 1 >= 0 && d - 1 >= floorDiag, has2 = d - 2 >= 0 && d - 2 >= floorDiag;
    if (has1) g1 = g_diag(c, d - 1);
    if (has2) g2 = g_diag(c, d - 2);
    for (int32_t base = 0; base < g.W; base += blockDim.x) {
        int32_t ci = base + threadIdx.x;
        int cnt = 0;
        for (int pass = 0; pass < (hdp ? 2 : 1); pass++) {
            int64_t w = 0;
            if (pass == 1) {
                int tot;
                int off = g_blockScan(cnt, sWarp, &tot, phase);
                w = *nAssign + off;
                *nAssign += tot;
            }
            if (ci >= g.W) continue;
            int64_t xmy = (int64_t) g.L + 2 * ci;
            int32_t x = g.x0 + ci, y = g.y0 - ci, xl = 0, xm = 0, xu = 0;
            int64_t pc = c.pathOff[x] - c.pathOff[g.x0];
            int32_t nP = c.pathOff[x + 1] - c.pathOff[x];
            int64_t pl = has1 ? g_cellPC(c, g1, xmy - 1, &xl) : -1, pu = has1 ? g_cellPC(c, g1, xmy + 1, &xu) : -1;
            int64_t pm = has2 ? g_cellPC(c, g2, xmy, &xm) : -1;
            const int32_t *kC = c.pathKmer + c.pathOff[x];
            for (int32_t p = 0; p < nP; p++) {
                int32_t kp = kC[p];
                const double *b = g_B(c, d, pc + p);
                if (pass == 0 && pl >= 0) {
                    int32_t nQ = c.pathOff[xl + 1] - c.pathOff[xl];
                    const int32_t *kQ = c.pathKmer + c.pathOff[xl];
                    double eP = g_gapX(c, kp);
                    for (int32_t q = 0; q < nQ; q++)
                        if (g_legal(c, kQ[q], kp)) {
                            const double *f = g_F(c, g1, pl + q);
                            acc9[0 * 3 + 1] += exp(__dsub_rn(__dadd_rn(__dadd_rn(f[0], b[1]), __dadd_rn(eP, T[T_GAP_OPEN_X])), total));
                            acc9[1 * 3 + 1] += exp(__dsub_rn(__dadd_rn(__dadd_rn(f[1], b[1]), __dadd_rn(eP, T[T_GAP_EXTEND_X])), total));
                            acc9[2 * 3 + 1] += exp(__dsub_rn(__dadd_rn(__dadd_rn(f[2], b[1]), __dadd_rn(eP, T[T_GAP_SWITCH_TO_X])), total));
                        }
                }
                if (pm >= 0) {
                    int32_t nQ = c.pathOff[xm + 1] - c.pathOff[xm];
                    const int32_t *kQ = c.pathKmer + c.pathOff[xm];
                    double eP = g_emit(c, kp, y, true);
                    for (int32_t q = 0; q < nQ; q++)
                        if (g_legal(c, kQ[q], kp)) {
                            const double *f = g_F(c, g2, pm + q);
                            const int tix[3] = {T_MATCH_CONTINUE, T_MATCH_FROM_GAP_X, T_MATCH_FROM_GAP_Y};
                            for (int s = 0; s < 3; s++) {
                                double pr = exp(__dsub_rn(__dadd_rn(__dadd_rn(f[s], b[0]), __dadd_rn(eP, T[tix[s]])), total));
                                if (pass == 0) {
                                    acc9[s * 3 + 0] += pr;
                                    if (hdp && pr >= c.P->threshold) cnt++;
                                } else if (pr >= c.P->threshold) {
                                    if (w < c.I->assignCap) {
                                        // kmer pointer = sX->get(x-1) (index -1 clamps to 0, :497-499); event = y-1
                                        c.P->assignK[c.I->assignOff + w] = x - 1 >= 0 ? x - 1 : 0;
                                        c.P->assignE[c.I->assignOff + w] = y - 1;
                                    }
                                    w++;
                                }
                            }
                        }
                }
                if (pass == 0 && pu >= 0) {
                    double eP = g_emit(c, kp, y, false);
                    for (int32_t q = 0; q < nP; q++)
                        if (kC[q] == kp) {
                            const double *f = g_F(c, g1, pu + q);
                            acc9[0 * 3 + 2] += exp(__dsub_rn(__dadd_rn(__dadd_rn(f[0], b[2]), __dadd_rn(eP, T[T_GAP_OPEN_Y])), total));
                            acc9[2 * 3 + 2] += exp(__dsub_rn(__dadd_rn(__dadd_rn(f[2], b[2]), __dadd_rn(eP, T[T_GAP_EXTEND_Y])), total));
                        }
                }
            }
        }
    }
}

#ifdef SA_GEN_MINB   // A/B knob: resident CTAs per SM the register allocation is held to (default: ptxas' own choice, 96)
#define SA_GEN_BOUNDS __launch_bounds__(128, SA_GEN_MINB)
#else
#define SA_GEN_BOUNDS __launch_bounds__(128)
#endif
template <bool EXPECT>
__global__ void SA_GEN_BOUNDS sa_general_kernel(const GParams P) {
    __shared__ int sItem;
    __shared__ double sTotal;
    __shared__ int sWarp[16];
    __shared__ double sAcc[8][9];
    sa_logAdd_init();
    int scanPhase = 0;                    // g_blockScan's call count (which of its two sets of warp sums is next)
    for (;;) {
        __syncthreads();
        if (threadIdx.x == 0) sItem = atomicAdd(P.counter, 1);
        __syncthreads();
        if (sItem >= P.nItems) break;
        const int it = P.order[sItem];
        const GItem &I = P.items[it];
        GCtx c;
        c.P = &P; c.I = &I;
        c.pathOff = P.pathOff + I.pathOffOff;
        c.pathKmer = P.pathKmer + I.pathKmerOff;
        c.bandL = P.bandL + I.diagOff; c.bandW = P.bandW + I.diagOff; c.bandOff = P.bandOff + I.diagOff;
        c.F = P.scratch + (int64_t) blockIdx.x * P.slabDoubles;
        c.B = c.F + P.ringSize * 3;
        c.vals = c.B + 3 * P.maxDiagPC * 3;
        c.ringMask = P.ringSize - 1;
        c.bStride = P.maxDiagPC;
        const int64_t nDiag = (int64_t) I.lX + I.lY;
        int64_t nPairs = 0, nAssign = 0;
        double acc9[9];
        for (int i = 0; i < 9; i++) acc9[i] = 0.0;
        double likelihood = 0.0;
        if (nDiag > 0) {
            // diagonal 0: start / ragged start state (stateMachine.c:1134-1145)
            {
                GDiag g0 = g_diag(c, 0);
                bool ragged = I.flags & 1;
                for (int64_t j = threadIdx.x; j < g0.W; j += blockDim.x) {
                    double *o = g_F(c, g0, j);
                    o[0] = ragged ? SA_NEG_INF : 0.0;
                    o[1] = ragged ? 0.0 : SA_NEG_INF;
                    o[2] = ragged ? 0.0 : SA_NEG_INF;
                }
            }
            int64_t tracedBackTo = 0;
            for (int64_t d = 1; d <= nDiag; d++) {
                __syncthreads();
                g_forwardDiag(c, d);
                bool atEnd = d == nDiag;
                bool tbPoint = d >= tracedBackTo + P.minDiags && (int64_t) c.bandW[d] <= P.expansion * 2 + 1;
                if (!(atEnd || tbPoint)) continue;
                __syncthreads();
                // end / ragged end state on the traceback start diagonal (stateMachine.c:1147-1173)
                {
                    GDiag gt = g_diag(c, d);
                    int32_t x0 = gt.x0;
                    int64_t nPC = c.pathOff[x0 + gt.W] - c.pathOff[x0];
                    bool rag = atEnd && (I.flags & 2);
                    double e0 = rag ? __ddiv_rn(__dadd_rn(P.trans[T_GAP_OPEN_X], P.trans[T_GAP_OPEN_Y]), 2.0) : P.trans[T_MATCH_CONTINUE];
                    double e1 = rag ? P.trans[T_GAP_EXTEND_X] : P.trans[T_MATCH_FROM_GAP_X];
                    double e2 = rag ? P.trans[T_GAP_EXTEND_Y] : P.trans[T_MATCH_FROM_GAP_Y];
                    for (int64_t j = threadIdx.x; j < nPC; j += blockDim.x) {
                        double *o = g_B(c, d, j);
                        o[0] = e0; o[1] = e1; o[2] = e2;
                    }
                }
                int64_t tracedBackFrom = d - (atEnd ? 0 : P.traceBack + 1);
                double total = SA_NEG_INF;
                int64_t nPost = 0;
                // One barrier interval per diagonal: b[d2] is complete at the top; the callback of d2 (reads f and
                // b[d2], b[d2+1]) and the backward step INTO d2-1 (reads b[d2], b[d2+1]; writes the third slot of the
                // 3-row backward ring) do not depend on each other.
                for (int64_t d2 = d; d2 > tracedBackTo; d2--) {
                    __syncthreads();
                    const bool post = d2 <= tracedBackFrom;
                    if (post && nPost++ % 10 == 0) {
                        g_totalProbability(c, d2, d, &sTotal);
                        total = sTotal;
                    }
                    if (d2 - 1 > tracedBackTo) g_backwardDiag(c, d2 - 1, d);
                    if (post) {
                        if (EXPECT) {
                            likelihood += total;
                            g_expectationsDiag(c, d2, total, tracedBackTo, acc9, &nAssign, sWarp, scanPhase);
                        } else {
                            g_posteriorDiag(c, d2, total, &nPairs, sWarp, scanPhase);
                        }
                    }
                }
                tracedBackTo = tracedBackFrom;
            }
        }
        if (EXPECT) {
            // block reduction of the 9 transition sums
            int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
            for (int i = 0; i < 9; i++) {
                double v = acc9[i];
                for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
                if (lane == 0) sAcc[w][i] = v;
            }
            __syncthreads();
            if (threadIdx.x == 0) {
                int nW = (blockDim.x + 31) >> 5;
                for (int i = 0; i < 9; i++) {
                    double v = 0.0;
                    for (int j = 0; j < nW; j++) v += sAcc[j][i];
                    P.expOut[(int64_t) it * 10 + i] = v;
                }
                P.expOut[(int64_t) it * 10 + 9] = likelihood;
                if (P.assignCount) P.assignCount[it] = nAssign;
            }
        } else if (threadIdx.x == 0) {
            P.pairCount[it] = nPairs;
        }
    }
}


// =====================================================================================================================
// sa_mp_kernel — the alignment pass of the general kernel with its WORKING SET out of global memory.
//
// Same inputs, same arithmetic, same order of every logAdd as sa_general_kernel<false> (the results are byte-identical);
// what changes is where the sweep's neighbours live and how a thread finds its cell:
//   * the three most recent diagonals of the sweep in flight (f(d-1), f(d-2) and the one being written; b(d+1), b(d+2)
//     and the one being written during a traceback) are ONE set of three buffers indexed by path-cell within the
//     diagonal — in shared memory when the widest diagonal of the batch fits (GParams.mpCap), else in the per-CTA
//     global slab the general kernel uses for its backward ring.  The forward ring in global memory is still written on
//     every diagonal (the traceback reads f from it) but no longer read by the forward sweep; after a traceback the two
//     newest forward diagonals are copied back from it;
//   * thread j of a diagonal owns path-cell j; its column comes from the per-path-cell array pcCol (one load) instead of
//     a bisection over the column prefix; neighbours are located by column (32-bit), not by x - y arithmetic in 64 bits;
//   * band descriptors of d, d-1, d-2 are carried from step to step instead of being re-read three times per step;
//   * 256 threads per CTA.
// Expectations (and the HDP assignment lists) stay on sa_general_kernel<true>.
struct MDiag {
    int32_t x0, W, pc0, nPC;     // first column, cells, item-local index of its first path-cell, path-cells
    int64_t off;                 // path-cells of the item before this diagonal (forward ring position)
};
__device__ __forceinline__ MDiag m_diag(const GCtx &c, int64_t d) {
    MDiag g;
    const int32_t L = c.bandL[d];
    g.W = c.bandW[d]; g.off = c.bandOff[d];
    g.x0 = (int32_t) ((d + L) / 2);
    g.pc0 = c.pathOff[g.x0];
    g.nPC = c.pathOff[g.x0 + g.W] - g.pc0;
    return g;
}
__device__ __forceinline__ double *m_W(const GCtx &c, int64_t d, int32_t pc) {      // working set, as g_B addresses it
    return c.B + ((int64_t) (d % 3) * c.bStride + pc) * 3;
}
__device__ __forceinline__ double *m_F(const GCtx &c, const MDiag &g, int32_t pc) {
    return c.F + ((g.off + pc) & c.ringMask) * 3;
}
__device__ __forceinline__ double *m_E(const GCtx &c, const MDiag &g, int32_t pc) {
    return c.E + ((g.off + pc) & c.ringMask) * 2;
}
// path-cell (within diagonal g) of column x's first path, -1 when the column is outside the band of g
__device__ __forceinline__ int32_t m_colPC(const GCtx &c, const MDiag &g, int32_t x) {
    return (x >= g.x0 && x < g.x0 + g.W) ? c.pathOff[x] - g.pc0 : -1;
}

__device__ __forceinline__ void m_forwardDiag(const GCtx &c, int64_t d, const MDiag &g, const MDiag &g1, const MDiag &g2, bool has2) {
    const double *T = c.P->trans;
    const bool swX = T[T_GAP_SWITCH_TO_X] > SA_NEG_INF || T[T_GAP_SWITCH_TO_X] != T[T_GAP_SWITCH_TO_X];
    for (int32_t j = threadIdx.x; j < g.nPC; j += blockDim.x) {
        const int32_t x = c.pcCol[g.pc0 + j], y = (int32_t) d - x;
        const int32_t colPC = c.pathOff[x] - g.pc0, nP = c.pathOff[x + 1] - c.pathOff[x];
        const int32_t pl = m_colPC(c, g1, x - 1), pu = m_colPC(c, g1, x), pm = has2 ? m_colPC(c, g2, x - 1) : -1;
        const int32_t *kC = c.pathKmer + c.pathOff[x];
        const int32_t kp = kC[j - colPC];
        // both emissions of the path-cell, kept beside the forward ring: the backward sweep reads them back instead of
        // evaluating the two divisions again (same kmer, same event: the same bits)
        const double eMv = g_emit(c, kp, y, true), eYv = g_emit(c, kp, y, false);
        {
            double *e = m_E(c, g, j);
            e[0] = eMv; e[1] = eYv;
        }
        double M = SA_NEG_INF, X = SA_NEG_INF, Y = SA_NEG_INF;
        if (pl >= 0) {
            const int32_t nQ = c.pathOff[x] - c.pathOff[x - 1];
            const int32_t *kQ = c.pathKmer + c.pathOff[x - 1];
            const double eP = g_gapX(c, kp);
            for (int32_t q = 0; q < nQ; q++)
                if (g_legal(c, kQ[q], kp)) {
                    const double *f = m_W(c, d - 1, pl + q);
                    X = sa_logAdd(X, __dadd_rn(f[0], __dadd_rn(eP, T[T_GAP_OPEN_X])));
                    X = sa_logAdd(X, __dadd_rn(f[1], __dadd_rn(eP, T[T_GAP_EXTEND_X])));
                    if (swX) X = sa_logAdd(X, __dadd_rn(f[2], __dadd_rn(eP, T[T_GAP_SWITCH_TO_X])));
                }
        }
        if (pm >= 0) {
            const int32_t nQ = c.pathOff[x] - c.pathOff[x - 1];
            const int32_t *kQ = c.pathKmer + c.pathOff[x - 1];
            const double eP = eMv;
            for (int32_t q = 0; q < nQ; q++)
                if (g_legal(c, kQ[q], kp)) {
                    const double *f = m_W(c, d - 2, pm + q);
                    M = sa_logAdd(M, __dadd_rn(f[0], __dadd_rn(eP, T[T_MATCH_CONTINUE])));
                    M = sa_logAdd(M, __dadd_rn(f[1], __dadd_rn(eP, T[T_MATCH_FROM_GAP_X])));
                    M = sa_logAdd(M, __dadd_rn(f[2], __dadd_rn(eP, T[T_MATCH_FROM_GAP_Y])));
                }
        }
        if (pu >= 0) {
            const double eP = eYv;
            for (int32_t q = 0; q < nP; q++)
                if (kC[q] == kp) {
                    const double *f = m_W(c, d - 1, pu + q);
                    Y = sa_logAdd(Y, __dadd_rn(f[0], __dadd_rn(eP, T[T_GAP_OPEN_Y])));
                    Y = sa_logAdd(Y, __dadd_rn(f[2], __dadd_rn(eP, T[T_GAP_EXTEND_Y])));
                }
        }
        double *o = m_W(c, d, j);
        o[0] = M; o[1] = X; o[2] = Y;
        double *r = m_F(c, g, j);
        r[0] = M; r[1] = X; r[2] = Y;
    }
}

// b(d) by gathering from b(d+1) / b(d+2) in the working set (g_backwardDiag's order)
__device__ __forceinline__ void m_backwardDiag(const GCtx &c, int64_t d, const MDiag &g, const MDiag &g1, const MDiag &g2,
                                               bool has1, bool has2) {
    const double *T = c.P->trans;
    const bool swX = T[T_GAP_SWITCH_TO_X] > SA_NEG_INF || T[T_GAP_SWITCH_TO_X] != T[T_GAP_SWITCH_TO_X];
    for (int32_t j = threadIdx.x; j < g.nPC; j += blockDim.x) {
        const int32_t x = c.pcCol[g.pc0 + j];
        const int32_t colPC = c.pathOff[x] - g.pc0, nQ = c.pathOff[x + 1] - c.pathOff[x];
        const int32_t pa = has2 ? m_colPC(c, g2, x + 1) : -1;      // (x+1, y+1): this cell is its "middle"
        const int32_t pu = has1 ? m_colPC(c, g1, x) : -1;          // (x,   y+1): this cell is its "upper"
        const int32_t pl = has1 ? m_colPC(c, g1, x + 1) : -1;      // (x+1, y  ): this cell is its "lower"
        const int32_t *kZ = c.pathKmer + c.pathOff[x];
        const int32_t kq = kZ[j - colPC];
        double M = SA_NEG_INF, X = SA_NEG_INF, Y = SA_NEG_INF;
        if (pa >= 0) {
            const int32_t nP = c.pathOff[x + 2] - c.pathOff[x + 1];
            const int32_t *kP = c.pathKmer + c.pathOff[x + 1];
            for (int32_t p = 0; p < nP; p++)
                if (g_legal(c, kq, kP[p])) {
                    const double eP = m_E(c, g2, pa + p)[0];          // g_emit(c, kP[p], y + 1, true), as the forward sweep left it
                    const double a = m_W(c, d + 2, pa + p)[0];
                    M = sa_logAdd(M, __dadd_rn(a, __dadd_rn(eP, T[T_MATCH_CONTINUE])));
                    X = sa_logAdd(X, __dadd_rn(a, __dadd_rn(eP, T[T_MATCH_FROM_GAP_X])));
                    Y = sa_logAdd(Y, __dadd_rn(a, __dadd_rn(eP, T[T_MATCH_FROM_GAP_Y])));
                }
        }
        if (pu >= 0) {
            for (int32_t p = 0; p < nQ; p++)
                if (kZ[p] == kq) {
                    const double eP = m_E(c, g1, pu + p)[1];          // g_emit(c, kq, y + 1, false): same kmer, same event
                    const double a = m_W(c, d + 1, pu + p)[2];
                    M = sa_logAdd(M, __dadd_rn(a, __dadd_rn(eP, T[T_GAP_OPEN_Y])));
                    Y = sa_logAdd(Y, __dadd_rn(a, __dadd_rn(eP, T[T_GAP_EXTEND_Y])));
                }
        }
        if (pl >= 0) {
            const int32_t nP = c.pathOff[x + 2] - c.pathOff[x + 1];
            const int32_t *kP = c.pathKmer + c.pathOff[x + 1];
            for (int32_t p = 0; p < nP; p++)
                if (g_legal(c, kq, kP[p])) {
                    const double eP = g_gapX(c, kP[p]);
                    const double a = m_W(c, d + 1, pl + p)[1];
                    M = sa_logAdd(M, __dadd_rn(a, __dadd_rn(eP, T[T_GAP_OPEN_X])));
                    X = sa_logAdd(X, __dadd_rn(a, __dadd_rn(eP, T[T_GAP_EXTEND_X])));
                    if (swX) Y = sa_logAdd(Y, __dadd_rn(a, __dadd_rn(eP, T[T_GAP_SWITCH_TO_X])));
                }
        }
        double *o = m_W(c, d, j);
        o[0] = M; o[1] = X; o[2] = Y;
    }
}

// Launch: 256 threads, dynamic shared memory = 3 * mpCap * 3 doubles (0 when the working set stays in global memory)
__global__ void __launch_bounds__(256) sa_mp_kernel(const GParams P) {
    extern __shared__ __align__(16) double sa_mpsm[];
    __shared__ int sItem;
    __shared__ double sTotal;
    __shared__ int sWarp[16];
    sa_logAdd_init();
    int scanPhase = 0;                    // g_blockScan's call count (which of its two sets of warp sums is next)
    for (;;) {
        __syncthreads();
        if (threadIdx.x == 0) sItem = atomicAdd(P.counter, 1);
        __syncthreads();
        if (sItem >= P.nItems) break;
        const int it = P.order[sItem];
        const GItem &I = P.items[it];
        GCtx c;
        c.P = &P; c.I = &I;
        c.pathOff = P.pathOff + I.pathOffOff;
        c.pathKmer = P.pathKmer + I.pathKmerOff;
        c.pcCol = P.pcCol + I.pathKmerOff;
        c.bandL = P.bandL + I.diagOff; c.bandW = P.bandW + I.diagOff; c.bandOff = P.bandOff + I.diagOff;
        c.F = P.scratch + (int64_t) blockIdx.x * P.slabDoubles;
        double *slabB = c.F + P.ringSize * 3;
        c.B = P.mpCap > 0 ? sa_mpsm : slabB;
        c.bStride = P.mpCap > 0 ? P.mpCap : P.maxDiagPC;
        c.vals = slabB + 3 * P.maxDiagPC * 3;
        c.E = c.vals + 2 * P.maxWidth;
        c.ringMask = P.ringSize - 1;
        const int64_t nDiag = (int64_t) I.lX + I.lY;
        int64_t nPairs = 0;
        if (nDiag > 0) {
            MDiag g1 = m_diag(c, 0), g2 = g1;       // d - 1, d - 2
            {
                const bool ragged = I.flags & 1;
                for (int32_t j = threadIdx.x; j < g1.nPC; j += blockDim.x) {
                    double *o = m_W(c, 0, j), *r = m_F(c, g1, j);
                    o[0] = r[0] = ragged ? SA_NEG_INF : 0.0;
                    o[1] = r[1] = ragged ? 0.0 : SA_NEG_INF;
                    o[2] = r[2] = ragged ? 0.0 : SA_NEG_INF;
                }
            }
            int64_t tracedBackTo = 0;
            for (int64_t d = 1; d <= nDiag; d++) {
                __syncthreads();
                const MDiag g = m_diag(c, d);
                m_forwardDiag(c, d, g, g1, g2, d >= 2);
                const bool atEnd = d == nDiag;
                const bool tbPoint = d >= tracedBackTo + P.minDiags && (int64_t) g.W <= P.expansion * 2 + 1;
                if (!(atEnd || tbPoint)) { g2 = g1; g1 = g; continue; }
                __syncthreads();
                {   // end / ragged end state on the traceback start diagonal (stateMachine.c:1147-1173)
                    const bool rag = atEnd && (I.flags & 2);
                    const double e0 = rag ? __ddiv_rn(__dadd_rn(P.trans[T_GAP_OPEN_X], P.trans[T_GAP_OPEN_Y]), 2.0) : P.trans[T_MATCH_CONTINUE];
                    const double e1 = rag ? P.trans[T_GAP_EXTEND_X] : P.trans[T_MATCH_FROM_GAP_X];
                    const double e2 = rag ? P.trans[T_GAP_EXTEND_Y] : P.trans[T_MATCH_FROM_GAP_Y];
                    for (int32_t j = threadIdx.x; j < g.nPC; j += blockDim.x) {
                        double *o = m_W(c, d, j);
                        o[0] = e0; o[1] = e1; o[2] = e2;
                    }
                }
                const int64_t tracedBackFrom = d - (atEnd ? 0 : P.traceBack + 1);
                double total = SA_NEG_INF;
                int64_t nPost = 0;
                MDiag b0 = g, b1 = g, b2 = g;       // d2 - 1, d2, d2 + 1 while the loop runs (b1 = current callback diagonal)
                for (int64_t d2 = d; d2 > tracedBackTo; d2--) {
                    __syncthreads();
                    const bool post = d2 <= tracedBackFrom;
                    if (post && nPost++ % 10 == 0) {
                        g_totalProbability(c, d2, d, &sTotal);
                        total = sTotal;
                    }
                    if (d2 - 1 > tracedBackTo) {
                        b0 = m_diag(c, d2 - 1);
                        m_backwardDiag(c, d2 - 1, b0, b1, b2, true, d2 + 1 <= d);
                    }
                    if (post) g_posteriorDiag(c, d2, total, &nPairs, sWarp, scanPhase);
                    b2 = b1; b1 = b0;
                }
                tracedBackTo = tracedBackFrom;
                if (!atEnd) {       // the forward sweep goes on: f(d), f(d - 1) back into the working set
                    __syncthreads();
                    for (int32_t j = threadIdx.x; j < g.nPC; j += blockDim.x) {
                        const double *r = m_F(c, g, j);
                        double *o = m_W(c, d, j);
                        o[0] = r[0]; o[1] = r[1]; o[2] = r[2];
                    }
                    for (int32_t j = threadIdx.x; j < g1.nPC; j += blockDim.x) {
                        const double *r = m_F(c, g1, j);
                        double *o = m_W(c, d - 1, j);
                        o[0] = r[0]; o[1] = r[1]; o[2] = r[2];
                    }
                }
                g2 = g1; g1 = g;
            }
        }
        if (threadIdx.x == 0) P.pairCount[it] = nPairs;
    }
}
