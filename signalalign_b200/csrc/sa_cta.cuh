// sa_cta.cuh — the CTA band kernel: one CTA of W warps per work item, C (1 or 2) adjacent band cells per thread.
//
// Same eligibility and arithmetic as the warp kernel it replaces (un-ambiguous reference, Gaussian MeanOnly emission,
// FP64 cubic logAdd in the reference's accumulation order), different mapping:
//
//   * thread t owns slots C t .. C t + C - 1 of NS = 32 W C slots; slot = column mod NS inside a sliding window
//     [wlo, wlo + NS), so along a diagonal the `upper` neighbour (x, y-1) is the thread's own register, `lower` /
//     `middle` (x-1, .) of the thread's upper cell are its own lower cell, and those of the lower cell come from thread
//     t-1 through a double-buffered shared-memory mailbox (3 doubles per thread), one split barrier per diagonal.
//     Two cells per thread halve everything that is per thread and per diagonal (mailbox, barrier, band and ring
//     bookkeeping) and double the independent logAdd chains in flight.
//   * per-thread state is a dozen doubles (f(d-1), f(d-2) of one column), so registers allow ~5 CTAs of 128 threads per
//     SM (20 warps) instead of the 8 warps the 4-cells-per-lane layout reached: the FP64 dependency chains of the
//     logAdd are hidden by warp interleaving instead of by compiler-scheduled ILP.
//   * everything a step needs besides the DP state comes from shared-memory rings filled cooperatively and coalesced:
//     per-column emission constants (5 doubles, scale/shift/var folded in), event means, and the band of the next NS
//     diagonals (band_construct in closed form per diagonal from the staged anchors).  No global load sits on the
//     forward critical path.
//   * forward rows go to the same HBM ring as before (row = [xlo, xhi | M plane | X plane | Y plane]) for the blocked
//     traceback; the backward sweep mirrors the forward one (neighbours from thread t+1).
//   * total probability: the per-cell terms go through shared memory and warp 0 folds them in column order (the fold
//     is sequential and non-associative), skipping exact no-ops; the result is broadcast through shared memory.
//   * posterior pairs: exp() only for cells that can reach the threshold; hits are counted per warp (ballot) and written
//     one barrier later in the reference's order.
#pragma once
#include "sa_fast.cuh"

// Split CTA barrier (mbarrier in shared memory): a thread ARRIVES right after posting its mailbox values and WAITS only
// when it needs its neighbour's, so the row store, the loop bookkeeping and the next cell's emissions overlap the wait
// for the slowest warp.  arrive = release, wait = acquire (CTA scope), so the mailbox writes are visible after the wait.
__device__ __forceinline__ void sa_mbInit(unsigned long long *bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"((unsigned) __cvta_generic_to_shared(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void sa_mbArrive(unsigned long long *bar) {
    asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.shared::cta.b64 st, [%0];\n\t}" ::"r"((unsigned) __cvta_generic_to_shared(bar)) : "memory");
}
__device__ __forceinline__ void sa_mbWait(unsigned long long *bar, unsigned &phase) {
    unsigned ok;
    const unsigned addr = (unsigned) __cvta_generic_to_shared(bar);
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(addr), "r"(phase) : "memory");
    } while (!ok);
    phase ^= 1u;
}

// EX = false: aligned pairs (diagonalCalculationPosteriorMatchProbs); EX = true: transition expectations + likelihood
// (diagonalCalculation_Expectations, reference impl/pairwiseAligner.c:1423-1443) on the same sweeps.
// HD = true: HDP emission (stateMachine3HDP: the same kmer density for the match and the gap-Y state) instead of the
// Gaussian tables; per column the ring then holds var * mu, scale * mu and the kmer's row of the spline tables.
template <int W, int C, bool EX, bool HD>
struct CtaK {
    static_assert(C == 1 || C == 2, "one or two cells per thread");
    static constexpr int NT = 32 * W;        // threads
    static constexpr int NS = NT * C;        // slots
    static constexpr int NL = NS / 32;       // slots per lane when ONE warp walks all of them (the ordered fold)
    static constexpr int NH = W * C;         // hit-mask words per diagonal: [warp][cell]
    static constexpr int ROW = (HD ? 4 : 3) * NS + 2;   // doubles per stored forward row (HDP: + the emission plane)
    static constexpr int RN = 2 * NS;        // entries of the column / event rings
    // Expectation terms below e^-50 = 2e-22 are not evaluated: a read adds ~1e7 terms to sums of order 1..1e4 that are
    // compared at 1e-6 relative (structural zeros: 1e-9 absolute).  The test is per WARP (__any_sync): on most diagonals
    // the alignment path crosses only one of the CTA's warps, and the other skips all fourteen exp() of its two cells.
    static constexpr double kExpZero = -50.0;

    struct Smem {
        double cst[5][RN];       // km, iM, cM, iY, cY per column (index = column & (RN - 1))
        double evs[RN];          // event mean per y (index = y & (RN - 1)); 0 outside [1, lY]
        double nbx[2][3][NT];    // neighbour mailbox: every thread's 3 newest values, double buffered by diagonal parity
        double foldv[2][NS];     // total-probability terms by slot
        double sTotal;           // ... and their ordered fold (warp 0 computes, everybody reads)
        int2 bandTab[NT];        // [xlo, xhi] of forward diagonals bd0 .. bd0 + NT - 1
        int32_t aX[NT + 1], aY[NT + 1], aD[NT + 1];   // staged anchors (+1 coordinates) and their diagonals
        alignas(16) unsigned hitM[2][NH < 4 ? 4 : NH];   // posterior hits per warp and cell (ballot), double buffered
        unsigned hitA[2][3][(EX && HD) ? NH : 1];        // HDP expectations: match terms >= threshold, per from-state
        int32_t sItem, sAi;
        unsigned long long mbar; // split barrier of the diagonal steps
    };

    static __device__ __forceinline__ int32_t slotX(int32_t s, int32_t wlo) { return wlo + ((s - wlo) & (NS - 1)); }

    // ------------------------------------------------------------------------------------------ ring refills
    // constants of columns [c0, c0 + RN): column 0 is the NULL kmer, columns outside [1, lX] never become active
    static __device__ __forceinline__ void fillConsts(const FParams &P, const FItem &I, Smem &S, int t, int32_t c0) {
        for (int i = t; i < RN; i += NT) {
            const int32_t x = c0 + i, s = x & (RN - 1);
            double km = 0.0, iM = 0.0, iY = 0.0, cM = SA_NEG_INF, cY = SA_NEG_INF;
            if (HD) {
                // (km, iM, cM) = (var * mu, scale * mu, row of the kmer in the spline tables or -1 for the NULL kmer)
                long long row = -1;
                if (x >= 1 && x <= I.lX) {
                    const int32_t k = P.kmerIdx[I.kmerOff + x - 1];
                    const double mu = P.ft[(int64_t) k * 5];
                    km = I.var * mu;
                    iM = I.scale * mu;
                    row = P.kmerToRow[k];
                }
                cM = __longlong_as_double(row);
            } else if (x >= 1 && x <= I.lX) {
                const double *ft = P.ft + (int64_t) P.kmerIdx[I.kmerOff + x - 1] * 5;
                km = fma(I.scale, ft[0], I.shift);
                iM = ft[1] * I.invVar;
                cM = ft[2] + I.logInvVar;
                iY = ft[3] * I.invVar;
                cY = ft[4] + I.logInvVar;
            }
            S.cst[0][s] = km; S.cst[1][s] = iM; S.cst[2][s] = cM; S.cst[3][s] = iY; S.cst[4][s] = cY;
        }
    }
    static __device__ __forceinline__ void fillEvents(const FParams &P, const FItem &I, Smem &S, int t, int32_t y0) {
        for (int i = t; i < RN; i += NT) {
            const int32_t y = y0 + i;
            S.evs[y & (RN - 1)] = (y >= 1 && y <= I.lY) ? P.evMean[I.evOff + y - 1] : 0.0;
        }
    }
    // make the rings cover the window [wlo, wlo + NS) on diagonal d; uniform decisions, barriers only when refilling
    static __device__ __forceinline__ void ensureRings(const FParams &P, const FItem &I, Smem &S, int t, int32_t wlo,
                                                       int32_t d, int32_t &cLo, int32_t &eLo) {
        const int32_t ymin = d - (wlo + NS - 1);
        const bool needC = (unsigned) (wlo - cLo) > (unsigned) (RN - NS);    // wlo < cLo || wlo + NS > cLo + RN
        const bool needE = (unsigned) (ymin - eLo) > (unsigned) (RN - NS);
        if (needC || needE) {
            __syncthreads();   // nobody still reads the old contents
            if (needC) { cLo = wlo - NS / 2; fillConsts(P, I, S, t, cLo); }
            if (needE) { eLo = ymin - NS / 2; fillEvents(P, I, S, t, eLo); }
            __syncthreads();
        }
    }

    // band_construct (reference impl/pairwiseAligner.c:195-246) in closed form for diagonals d0 .. d0 + NT - 1:
    // between two consecutive anchors the band only depends on the diagonal, and the anchor interval of diagonal d is
    // (number of anchors whose diagonal x + y + 2 is < d).  Anchors are staged in shared memory, then every thread
    // evaluates its own diagonal.
    static __device__ __forceinline__ void fillBand(const FParams &P, const FItem &I, Smem &S, int t, int32_t d0) {
        __syncthreads();
        const int32_t base = S.sAi;            // anchors with diagonal < d0 - 1 (lower bound of every count below)
        const int32_t *anch = P.anchors + 2 * I.anchorOff;
        for (int j = t; j <= NT; j += NT) {
            const int32_t a = base - 1 + j;
            int32_t X = 0, Y = 0, D = -0x7fffffff;
            if (a >= I.nAnchors) { X = I.lX; Y = I.lY; D = 0x7fffffff; }
            else if (a >= 0) { X = anch[2 * a] + 1; Y = anch[2 * a + 1] + 1; D = X + Y; }
            S.aX[j] = X; S.aY[j] = Y; S.aD[j] = D;
        }
        __syncthreads();
        const int32_t d = d0 + t, nDiag = I.lX + I.lY;
        // c = number of staged real anchors (entries 1..NT) with diagonal < d
        int lo = 0, hi = NT;                   // invariant: entries 1..lo are < d, entries hi+1.. are >= d
        while (lo < hi) {
            const int mid = (lo + hi + 1) >> 1;
            if (S.aD[mid] < d) lo = mid; else hi = mid - 1;
        }
        const int c = lo < NT ? lo : NT - 1;   // (c + 1 <= NT always holds: anchor diagonals grow by >= 2)
        const int32_t pX = S.aX[c], pY = S.aY[c], nX = S.aX[c + 1], nY = S.aY[c + 1];
        const int32_t pxay = pX + pY, pxmy = pX - pY, nxay = nX + nY, nxmy = nX - nY, e = P.expansion;
        const int32_t xL = BandIt::clampc((pxay + pxmy - e) / 2, I.lX), yL = BandIt::clampc((nxay - nxmy + e) / 2, I.lY);
        const int32_t xU = BandIt::clampc((nxay + nxmy + e) / 2, I.lX), yU = BandIt::clampc((pxay - pxmy - e) / 2, I.lY);
        int32_t l = xL - yL, r = xU - yU, q;
        if ((d + l) & 1) l++;
        if ((d + r) & 1) r++;
        if ((q = xL - (d + l) / 2) > 0) l += 2 * q;
        if ((q = (d - l) / 2 - yL) > 0) l += 2 * q;
        if ((q = (d + r) / 2 - xU) > 0) r -= 2 * q;
        if ((q = yU - (d - r) / 2) > 0) r -= 2 * q;
        S.bandTab[t] = d <= nDiag ? make_int2((d + l) / 2, (d + r) / 2) : make_int2(0, 0);
        if (t == NT - 1) S.sAi = base + c;
        __syncthreads();
    }

    // ------------------------------------------------------------------------------------------ neighbour exchange
    // every thread posts the three values its neighbours need; one barrier later the neighbour reads them.  (Plain
    // shared-memory traffic instead of shuffles: no lane-dependent branches at the warp boundaries, 3 STS + 3 LDS.)
    static __device__ __forceinline__ void publish(Smem &S, int buf, int t, const double (&v)[3]) {
#pragma unroll
        for (int i = 0; i < 3; i++) S.nbx[buf][i][t] = v[i];
    }
    static __device__ __forceinline__ void fetchUp(const Smem &S, int buf, int t, double (&o)[3]) {   // from slot t - 1
        const int src = (t + NT - 1) & (NT - 1);
#pragma unroll
        for (int i = 0; i < 3; i++) o[i] = S.nbx[buf][i][src];
    }
    static __device__ __forceinline__ void fetchDown(const Smem &S, int buf, int t, double (&o)[3]) {  // from slot t + 1
        const int src = (t + 1) & (NT - 1);
#pragma unroll
        for (int i = 0; i < 3; i++) o[i] = S.nbx[buf][i][src];
    }

    // get_nanopore_kmer_density (reference stateMachine.c:527-554 -> hdp.c:2588-2612 -> hdp_math_utils.c:471-495):
    // cubic Hermite spline of the kmer's posterior predictive on the uniform grid, linear beyond its ends
    static __device__ __forceinline__ double hdpLogDensity(const FParams &P, const FItem &I, long long row, double q) {
        const int32_t n = P.G;
        const double *yv = P.pp + row * n, *sv = P.slopes + row * n;
        const double x0 = P.grid[0], xn = P.grid[n - 1];
        double v;
        if (q <= x0) v = yv[0] - sv[0] * (x0 - q);
        else if (q >= xn) v = yv[n - 1] + sv[n - 1] * (q - xn);
        else {
            // (reciprocal multiplies instead of the reference's three divisions: the knot index can only differ for q
            // within an ulp of a knot, where both intervals give the knot's value; results move by ~1e-16 relative)
            const double dx = P.gridDx, rdx = P.gridInvDx;
            int32_t il = (int32_t) ((q - x0) * rdx);
            il = il > n - 2 ? n - 2 : il;
            const int32_t ir = il + 1;
            const double yl = yv[il], yr = yv[ir], dy = yr - yl;
            const double a = sv[il] * dx - dy, b = dy - sv[ir] * dx;
            const double tl = (q - P.grid[il]) * rdx, tr = 1.0 - tl;
            v = (tr * yl + tl * yr) + (tl * tr) * (a * tr + b * tl);
        }
        const double dens = v > 0.0 ? v : 0.0;
        return log(I.invVar * dens);
    }

    static __device__ __forceinline__ void emissions(const FParams &P, const FItem &I, const Smem &S, int32_t x, int32_t y,
                                                     double &eM, double &eY) {
        const int cs = x & (RN - 1);
        if (HD) {
            const long long row = __double_as_longlong(S.cst[2][cs]);
            // de-scaled event mean (stateMachine.c:344-348): (mean + var mu - scale mu - shift) / var
            const double q = (((S.evs[y & (RN - 1)] + S.cst[0][cs]) - S.cst[1][cs]) - I.shift) * I.invVar;
            double e = SA_NEG_INF;
            if (row >= 0) e = hdpLogDensity(P, I, row, q);
            eM = e; eY = e;
            return;
        }
        const double d0 = S.evs[y & (RN - 1)] - S.cst[0][cs];
        const double aM = d0 * S.cst[1][cs], aY = d0 * S.cst[3][cs];
        eM = fma(-0.5 * aM, aM, S.cst[2][cs]);
        eY = fma(-0.5 * aY, aY, S.cst[4][cs]);
    }

    // The two sequential logAdd folds of S.foldv[0][*] and S.foldv[1][*] in column order, run by ONE warp.  The fold is
    // a dependent chain of logAdds (non-associative: the order is the reference's), so the work is to make the chain
    // short, exactly:
    //   * no-op cells: logAdd(acc, v) == acc whenever v <= acc - 7.5, and acc >= every value folded so far (the cubic
    //     is positive on [0, 7.5)), so v <= prefixMax - 7.6 never changes the accumulator;
    //   * reset cells: the cubic overshoots log(1 + e^-d) by < 1.1e-4, hence acc <= prefixMax + ln(NS) + NS * 1.1e-4
    //     < prefixMax + 5.6, and a cell with v >= prefixMax + 13.2 makes |acc - v| >= 7.5: the result is v itself,
    //     whatever came before.  The chain starts at the LAST such cell (the first finite cell is one).
    // On a diagonal the terms rise by tens of nats per cell towards the alignment path and fall off behind it, so a
    // chain is a handful of cells instead of every cell left of the path.  Candidates are read straight from shared
    // memory (uniform address), the two folds advance in lockstep: two independent dependency chains in flight.
    static __device__ __forceinline__ void orderedFold2(const Smem &S, int lane, int32_t wlo, double &out0, double &out1) {
        const int s0 = wlo & (NS - 1);
        unsigned m[2][NL];
#pragma unroll
        for (int f = 0; f < 2; f++) {
            double carry = SA_NEG_INF;      // max over the previous 32-cell blocks
            int lastReset = 0;
#pragma unroll
            for (int k = 0; k < NL; k++) {
                const double v = S.foldv[f][(s0 + k * 32 + lane) & (NS - 1)];
                double run = v;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const double q = __shfl_up_sync(0xffffffffu, run, o);
                    if (lane >= o) run = fmax(run, q);
                }
                double pm = __shfl_up_sync(0xffffffffu, run, 1);
                if (lane == 0) pm = SA_NEG_INF;
                pm = fmax(pm, carry);                       // exclusive prefix max in cell order
                const unsigned mc = __ballot_sync(0xffffffffu, v > pm - 7.6);     // false for -inf
                const unsigned mr = __ballot_sync(0xffffffffu, v >= pm + 13.2);
                m[f][k] = mc;
                if (mr) lastReset = k * 32 + 31 - __clz(mr);
                carry = fmax(carry, __shfl_sync(0xffffffffu, run, 31));
            }
#pragma unroll
            for (int k = 0; k < NL; k++) {
                if (k * 32 + 31 < lastReset) m[f][k] = 0;
                else if (k * 32 <= lastReset) m[f][k] &= 0xffffffffu << (lastReset - k * 32);
            }
        }
        double acc0 = SA_NEG_INF, acc1 = SA_NEG_INF;
#pragma unroll
        for (int k = 0; k < NL; k++) {
            unsigned a = m[0][k], b = m[1][k];
            while (a | b) {                                  // uniform
                if (a) {
                    const int i = __ffs(a) - 1;
                    a &= a - 1;
                    acc0 = sa_logAdd(acc0, S.foldv[0][(s0 + k * 32 + i) & (NS - 1)]);
                }
                if (b) {
                    const int i = __ffs(b) - 1;
                    b &= b - 1;
                    acc1 = sa_logAdd(acc1, S.foldv[1][(s0 + k * 32 + i) & (NS - 1)]);
                }
            }
        }
        out0 = acc0; out1 = acc1;
    }

    // f(d) of this thread's cell from the own column on d-1 (f1), and the left neighbour on d-1 (n1) and d-2 (n2)
    static __device__ __forceinline__ void forwardCell(const FParams &P, const Smem &S, int32_t x, int32_t d, bool act,
                                                       double cOX, double cEX, double eM, double eY,
                                                       const double (&f1)[3], const double (&n1)[3],
                                                       const double (&n2)[3], double (&o)[3]) {
        const double *T = P.trans;
        // stateMachine3_cellCalculate order; the first contribution lands on -inf (logAdd(-inf, v) == v exactly)
        const double X = sa_logAdd(n1[0] + cOX, n1[1] + cEX);
        const double M = sa_logAdd(sa_logAdd(n2[0] + (eM + T[T_MATCH_CONTINUE]), n2[1] + (eM + T[T_MATCH_FROM_GAP_X])),
                                   n2[2] + (eM + T[T_MATCH_FROM_GAP_Y]));
        const double Y = sa_logAdd(f1[0] + (eY + T[T_GAP_OPEN_Y]), f1[2] + (eY + T[T_GAP_EXTEND_Y]));
        o[0] = act ? M : SA_NEG_INF;
        o[1] = act ? X : SA_NEG_INF;
        o[2] = act ? Y : SA_NEG_INF;
    }

    static __device__ __forceinline__ void storeRow(double *row, int t, int32_t xlo, int32_t xhi, bool full,
                                                    const double (&f)[C][3], const double (&e)[C]) {
        if (t == 0) *reinterpret_cast<int2 *>(row) = make_int2(xlo, xhi);
        if (HD) {                              // the cell's log density: the traceback reads it back instead of
            double *q = row + 2 + 3 * NS + t * C;   // evaluating the spline and the log a second time
            if (C == 2) *reinterpret_cast<double2 *>(q) = make_double2(e[0], e[C - 1]);
            else q[0] = e[0];
        }
#pragma unroll
        for (int pl = 0; pl < 3; pl++) {
            if (pl > 0 && !full) break;
            double *q = row + 2 + pl * NS + t * C;       // (16-byte aligned for C == 2: ROW and NS are even)
            if (C == 2) *reinterpret_cast<double2 *>(q) = make_double2(f[0][pl], f[C - 1][pl]);
            else q[0] = f[0][pl];
        }
    }
    static __device__ __forceinline__ void loadRow(const double *row, int pl, int t, double (&o)[C]) {
        const double *q = row + 2 + pl * NS + t * C;
        if (C == 2) {
            const double2 v = *reinterpret_cast<const double2 *>(q);
            o[0] = v.x; o[C - 1] = v.y;
        } else o[0] = q[0];
    }

    // number of posterior hits in slots [0, k) from the ballots m[warp * C + cell]; slot = (32 warp + lane) C + cell
    static __device__ __forceinline__ int hitsBelow(const unsigned (&m)[NH], int k) {
        int n = 0;
#pragma unroll
        for (int w = 0; w < W; w++) {
            const int lo = w * 32 * C;
            if (k >= lo + 32 * C) {
#pragma unroll
                for (int j = 0; j < C; j++) n += __popc(m[w * C + j]);
            } else if (k > lo) {
                const int r = k - lo, l = r / C, jj = r % C;    // l <= 31
                const unsigned below = (1u << l) - 1u;
#pragma unroll
                for (int j = 0; j < C; j++) {
                    n += __popc(m[w * C + j] & below);
                    if (j < jj) n += (m[w * C + j] >> l) & 1u;
                }
            }
        }
        return n;
    }

    // blocked traceback from diagonal d down to tracedBackTo + 1 (reference impl/pairwiseAligner.c:1500-1570):
    // backward sweep mirrored on the forward one, posterior pairs for diagonals <= tracedBackFrom
    static __device__ __forceinline__ void traceback(const FParams &P, const FItem &I, Smem &S, const double *ring, int t,
                                                     int lane, int warp, int32_t d, int32_t rowI, int32_t xlo, int32_t xhi,
                                                     int32_t bw, int32_t &cLo, int32_t &eLo, int32_t tracedBackTo,
                                                     int32_t tracedBackFrom, bool atEnd, double cOX, double cEX,
                                                     unsigned &phase, int64_t &nPairs, double (&acc)[7], double &lik,
                                                     bool &yxNaN, int32_t cap) {
        const double *T = P.trans;
        // end / ragged-end state on diagonal d (reference stateMachine.c:1147-1173)
        const bool rag = atEnd && (I.flags & 2);
        const double e0 = rag ? (T[T_GAP_OPEN_X] + T[T_GAP_OPEN_Y]) / 2.0 : T[T_MATCH_CONTINUE];
        const double e1 = rag ? T[T_GAP_EXTEND_X] : T[T_MATCH_FROM_GAP_X];
        const double e2 = rag ? T[T_GAP_EXTEND_Y] : T[T_MATCH_FROM_GAP_Y];
        // own columns: diagonal d2 (cur, q) and d2 + 1 (nxt)
        int32_t myx[C];
        double curM[C], curE[C], qX[C], qY[C], qE[C], nxtM[C], nxtE[C];
        double n1M = SA_NEG_INF, n1E = 0.0;    // (M, eM) one diagonal up of the column right of the thread's last cell
#pragma unroll
        for (int j = 0; j < C; j++) {
            myx[j] = slotX(t * C + j, bw);
            const bool act = myx[j] >= xlo && myx[j] <= xhi;
            double eM, eY;
            if (HD) {
                double ev[C];
                loadRow(ring + (int64_t) rowI * ROW, 3, t, ev);
                eM = eY = ev[j];
            } else emissions(P, I, S, myx[j], d - myx[j], eM, eY);
            curM[j] = act ? e0 : SA_NEG_INF; curE[j] = eM;
            qX[j] = act ? e1 : SA_NEG_INF; qY[j] = act ? e2 : SA_NEG_INF; qE[j] = eY;
            nxtM[j] = SA_NEG_INF; nxtE[j] = 0.0;
        }
        sa_mbWait(&S.mbar, phase);             // everybody has posted (and finished reading for) forward step d
        double total = SA_NEG_INF;
        int32_t postMod = 0, bxlo = xlo, bxhi = xhi;   // postMod: posterior diagonals since the last total, mod 10
        // posterior hits of the previous diagonal, written after the next barrier (needs every warp's ballot)
        bool pendValid = false, pendHit[C];
        int32_t pendP[C], pendX[C], pendD = 0, pendS0 = 0;
#pragma unroll
        for (int j = 0; j < C; j++) { pendHit[j] = false; pendP[j] = 0; pendX[j] = 0; }
        auto flush = [&]() {
            unsigned m[NH];
            int tot = 0;
            {
                const uint4 *hp = reinterpret_cast<const uint4 *>(S.hitM[pendD & 1]);
#pragma unroll
                for (int w4 = 0; w4 < (NH + 3) / 4; w4++) {
                    const uint4 q = hp[w4];
                    const unsigned qq[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
                    for (int w = 0; w < 4; w++)
                        if (w4 * 4 + w < NH) { m[w4 * 4 + w] = qq[w]; tot += __popc(qq[w]); }
                }
            }
            bool anyHit = false;
#pragma unroll
            for (int j = 0; j < C; j++) anyHit |= pendHit[j];
            if (anyHit) {
                const int a = hitsBelow(m, pendS0);
#pragma unroll
                for (int j = 0; j < C; j++) {
                    if (!pendHit[j]) continue;
                    const int sl = t * C + j, b = hitsBelow(m, sl);
                    const int before = pendS0 <= sl ? b - a : tot - a + b;
                    const int64_t w = nPairs + before;
                    if (w < I.pairCap) {
                        sa_pair_t o;
                        o.p = pendP[j]; o.x = pendX[j] - 1; o.y = pendD - pendX[j] - 1;
                        o.kmerIdx = P.kmerIdx[I.kmerOff + pendX[j] - 1];
                        P.pairs[I.pairOff + w] = o;
                    }
                }
            }
            nPairs += tot;
            pendValid = false;
        };
        // HDP expectations: the assignments of the previous diagonal (cell order, then from-state M, X, Y), written after
        // the next barrier like the pairs; nPairs counts them
        unsigned pendA[C];
#pragma unroll
        for (int j = 0; j < C; j++) pendA[j] = 0u;
        auto flushA = [&]() {
            unsigned m[3][NH];
            int tot[3], all = 0;
#pragma unroll
            for (int s3 = 0; s3 < 3; s3++) {
                tot[s3] = 0;
#pragma unroll
                for (int w = 0; w < NH; w++) { m[s3][w] = S.hitA[pendD & 1][s3][w]; tot[s3] += __popc(m[s3][w]); }
                all += tot[s3];
            }
#pragma unroll
            for (int j = 0; j < C; j++) {
                if (!pendA[j]) continue;
                const int sl = t * C + j;
                int before = 0;
#pragma unroll
                for (int s3 = 0; s3 < 3; s3++) {
                    const int a = hitsBelow(m[s3], pendS0), b = hitsBelow(m[s3], sl);
                    before += pendS0 <= sl ? b - a : tot[s3] - a + b;
                }
                int64_t w = nPairs + before;
#pragma unroll
                for (int s3 = 0; s3 < 3; s3++) {
                    if (!((pendA[j] >> s3) & 1u)) continue;
                    if (w < I.assignCap) {
                        // kmer position x - 1 (clamped at 0 like the reference's index), event y - 1
                        P.assignK[I.assignOff + w] = pendX[j] - 1 >= 0 ? pendX[j] - 1 : 0;
                        P.assignE[I.assignOff + w] = pendD - pendX[j] - 1;
                    }
                    w++;
                }
            }
            nPairs += all;
            pendValid = false;
        };
        int2 hdrNext = make_int2(0, 0);   // band of d2 - 1 (the expectations also need it on the block's lowest diagonal)
        if (EX || d - 1 > tracedBackTo) hdrNext = *reinterpret_cast<const int2 *>(ring + (int64_t) (rowI == 0 ? cap - 1 : rowI - 1) * ROW);
        rowI = rowI + 1 == cap ? 0 : rowI + 1;
        for (int32_t d2 = d; d2 > tracedBackTo; d2--) {
            rowI = rowI == 0 ? cap - 1 : rowI - 1;             // ring row of d2
            const int32_t rowP = rowI == 0 ? cap - 1 : rowI - 1;   // ... of d2 - 1
            const double *row = ring + (int64_t) rowI * ROW;
            const bool post = d2 <= tracedBackFrom;
            if (EX && d2 - 3 >= tracedBackTo) {
                // the expectations read f(d2 - 1) and f(d2 - 2) of every cell; f(d2 - 2) is first touched on this step
                // and comes from HBM (a traceback block's rows exceed L2): start the next step's row now
                int32_t r3 = rowI - 3;
                r3 = r3 < 0 ? r3 + cap : r3;
                const double *q = ring + (int64_t) r3 * ROW + 2 + t * C;
#pragma unroll
                for (int pl = 0; pl < 3; pl++) asm volatile("prefetch.global.L1 [%0];" ::"l"(q + pl * NS));
            }
            double fM[C];
#pragma unroll
            for (int j = 0; j < C; j++) fM[j] = 0.0;
            if (post && !EX) loadRow(row, 0, t, fM);
            if (d2 < d) {
                bxlo = hdrNext.x; bxhi = hdrNext.y;
                if (EX || d2 - 1 > tracedBackTo)
                    hdrNext = *reinterpret_cast<const int2 *>(ring + (int64_t) rowP * ROW);
                const bool slid = bxlo < bw;
                if (slid) bw = bxlo;           // lower the window; recycled slots restart at -inf
                ensureRings(P, I, S, t, bw, d2, cLo, eLo);
                if (slid) {                    // (uniform branch: the per-slot checks run on slide steps only)
#pragma unroll
                    for (int j = 0; j < C; j++) {
                        const int32_t x = slotX(t * C + j, bw);
                        if (x != myx[j]) {
                            myx[j] = x;
                            curM[j] = SA_NEG_INF; curE[j] = 0.0; nxtM[j] = SA_NEG_INF; nxtE[j] = 0.0;
                            qX[j] = SA_NEG_INF; qY[j] = SA_NEG_INF; qE[j] = 0.0;
                        }
                    }
                }
                double eM[C], eY[C];
                if (HD) {
                    loadRow(row, 3, t, eM);
#pragma unroll
                    for (int j = 0; j < C; j++) eY[j] = eM[j];
                } else {
#pragma unroll
                    for (int j = 0; j < C; j++) emissions(P, I, S, myx[j], d2 - myx[j], eM[j], eY[j]);
                }
                sa_mbWait(&S.mbar, phase);     // mailbox and hit masks of diagonal d2 + 1 are visible
                if (pendValid) {
                    if (EX && HD) flushA(); else flush();
                }
                double nb[3];
                fetchDown(S, (d2 + 1) & 1, t, nb);
                // ascending j: cell j reads the values cell j + 1 held BEFORE this step
#pragma unroll
                for (int j = 0; j < C; j++) {
                    const bool last = j == C - 1;
                    const double aM = last ? n1M : nxtM[last ? j : j + 1], aE = last ? n1E : nxtE[last ? j : j + 1];
                    const double lX = last ? nb[2] : qX[last ? j : j + 1];
                    // gather in the order the reference's scatter delivers (see sa_general.cuh)
                    const double M = sa_logAdd(sa_logAdd(aM + (aE + T[T_MATCH_CONTINUE]), qY[j] + (qE[j] + T[T_GAP_OPEN_Y])), lX + cOX);
                    const double X = sa_logAdd(aM + (aE + T[T_MATCH_FROM_GAP_X]), lX + cEX);
                    const double Y = sa_logAdd(aM + (aE + T[T_MATCH_FROM_GAP_Y]), qY[j] + (qE[j] + T[T_GAP_EXTEND_Y]));
                    const bool act = myx[j] >= bxlo && myx[j] <= bxhi;
                    nxtM[j] = curM[j]; nxtE[j] = curE[j];
                    curM[j] = act ? M : SA_NEG_INF; curE[j] = eM[j];
                    qX[j] = act ? X : SA_NEG_INF; qY[j] = act ? Y : SA_NEG_INF; qE[j] = eY[j];
                }
                n1M = nb[0]; n1E = nb[1];
            }
            {
                const double pub[3] = {curM[0], curE[0], qX[0]};   // the left neighbour thread reads the FIRST cell
                publish(S, d2 & 1, t, pub);
            }
            if (!post) { sa_mbArrive(&S.mbar); continue; }
            const bool newTotal = postMod == 0;
            postMod = postMod == 9 ? 0 : postMod + 1;
            if (newTotal) {
                // diagonalCalculationTotalProbability (reference pairwiseAligner.c:1335-1353)
                double fX[C], fY[C];
                if (EX) loadRow(row, 0, t, fM);
                loadRow(row, 1, t, fX);
                loadRow(row, 2, t, fY);
                const double *rowm = ring + (int64_t) rowP * ROW + 2;
#pragma unroll
                for (int j = 0; j < C; j++) {
                    const int sl = t * C + j;
                    double v = fM[j] + curM[j];
                    v = sa_logAdd(v, fX[j] + qX[j]);
                    v = sa_logAdd(v, fY[j] + qY[j]);
                    // matches jumping over d2: middle-only forward step from f(d2 - 1) into the cells of d2 + 1
                    const int sp = (sl + NS - 1) & (NS - 1);
                    const double gM = rowm[sp], gX = rowm[NS + sp], gY = rowm[2 * NS + sp];
                    const double m = sa_logAdd(sa_logAdd(gM + (nxtE[j] + T[T_MATCH_CONTINUE]), gX + (nxtE[j] + T[T_MATCH_FROM_GAP_X])),
                                               gY + (nxtE[j] + T[T_MATCH_FROM_GAP_Y]));
                    S.foldv[0][sl] = v;
                    S.foldv[1][sl] = m + nxtM[j];
                }
                __syncthreads();
                if (warp == 0) {               // the fold is sequential: one warp does it, the others keep their issue slots
                    double tot, tot2;
                    orderedFold2(S, lane, bw, tot, tot2);
                    if (lane == 0) S.sTotal = sa_logAdd(tot, tot2);
                }
                __syncthreads();
                total = S.sTotal;
            }
            if (EX) {
                // diagonalCalculation_Expectations: current = b(d2), neighbours from f(d2 - 1) and f(d2 - 2); a neighbour
                // counts when its cell lies inside the band of its diagonal (and the diagonal was not freed:
                // d2 - 2 >= tracedBackTo).  Every term is exp(((f + b) + (eP + tP)) - total), as the reference adds it.
                if (t == 0) lik += total;
                const bool has2 = d2 - 2 >= tracedBackTo;
                const int32_t rowPP = rowP == 0 ? cap - 1 : rowP - 1;
                int2 h2 = make_int2(1, 0);
                if (has2) h2 = *reinterpret_cast<const int2 *>(ring + (int64_t) rowPP * ROW);
                const double *r1 = ring + (int64_t) rowP * ROW + 2, *r2 = ring + (int64_t) rowPP * ROW + 2;
                const int2 h1 = hdrNext;
#pragma unroll
                for (int j = 0; j < C; j++) {
                    const int32_t x = myx[j];
                    const bool here = x >= bxlo && x <= bxhi;
                    const bool lower = here && x - 1 >= h1.x && x - 1 <= h1.y;
                    const bool upper = here && x >= h1.x && x <= h1.y;
                    const bool middle = here && x - 1 >= h2.x && x - 1 <= h2.y;
                    const int sl = x & (NS - 1), sp = (x - 1) & (NS - 1);
                    if (__any_sync(0xffffffffu, lower)) {
                        const double f0 = r1[sp], f1 = r1[NS + sp];
                        const double a0 = ((f0 + qX[j]) + cOX) - total, a1 = ((f1 + qX[j]) + cEX) - total;
                        // terms below kExpZero (and -inf) are skipped, NaN is not
                        if (__any_sync(0xffffffffu, lower && !(a0 < kExpZero))) acc[1] += lower ? sa_expSum(a0) : 0.0;
                        if (__any_sync(0xffffffffu, lower && !(a1 < kExpZero))) acc[3] += lower ? sa_expSum(a1) : 0.0;
                        yxNaN |= lower && !(total > SA_NEG_INF);   // exp(-inf - total): NaN exactly when total is -inf
                    }
                    double pr[3] = {0.0, 0.0, 0.0};   // the three match terms (HDP: also the assignment test)
                    if (__any_sync(0xffffffffu, middle)) {
                        const double f0 = r2[sp], f1 = r2[NS + sp], f2 = r2[2 * NS + sp];
                        const double a0 = ((f0 + curM[j]) + (curE[j] + T[T_MATCH_CONTINUE])) - total;
                        const double a1 = ((f1 + curM[j]) + (curE[j] + T[T_MATCH_FROM_GAP_X])) - total;
                        const double a2 = ((f2 + curM[j]) + (curE[j] + T[T_MATCH_FROM_GAP_Y])) - total;
                        if (__any_sync(0xffffffffu, middle && !(a0 < kExpZero))) { pr[0] = middle ? sa_expSum(a0) : 0.0; acc[0] += pr[0]; }
                        if (__any_sync(0xffffffffu, middle && !(a1 < kExpZero))) { pr[1] = middle ? sa_expSum(a1) : 0.0; acc[2] += pr[1]; }
                        if (__any_sync(0xffffffffu, middle && !(a2 < kExpZero))) { pr[2] = middle ? sa_expSum(a2) : 0.0; acc[4] += pr[2]; }
                    }
                    if (HD) {
                        pendA[j] = 0u;
#pragma unroll
                        for (int s3 = 0; s3 < 3; s3++) {
                            const bool hs = middle && pr[s3] >= P.threshold;
                            S.hitA[d2 & 1][s3][warp * C + j] = __ballot_sync(0xffffffffu, hs);
                            pendA[j] |= hs ? (1u << s3) : 0u;
                        }
                        pendX[j] = x;
                    }
                    if (__any_sync(0xffffffffu, upper)) {
                        const double f0 = r1[sl], f2 = r1[2 * NS + sl];
                        const double a0 = ((f0 + qY[j]) + (qE[j] + T[T_GAP_OPEN_Y])) - total;
                        const double a2 = ((f2 + qY[j]) + (qE[j] + T[T_GAP_EXTEND_Y])) - total;
                        if (__any_sync(0xffffffffu, upper && !(a0 < kExpZero))) acc[5] += upper ? sa_expSum(a0) : 0.0;
                        if (__any_sync(0xffffffffu, upper && !(a2 < kExpZero))) acc[6] += upper ? sa_expSum(a2) : 0.0;
                    }
                }
                if (HD) { pendValid = true; pendD = d2; pendS0 = bw & (NS - 1); }
                sa_mbArrive(&S.mbar);
                continue;
            }
            // diagonalCalculationPosteriorMatchProbs (reference pairwiseAligner.c:1355-1421)
#pragma unroll
            for (int j = 0; j < C; j++) {
                const int32_t x = myx[j];
                const double arg = (fM[j] + curM[j]) - total;
                const bool cand = x >= bxlo && x <= bxhi && x > 0 && d2 - x > 0 && arg >= P.logThrGate;
                bool hit = false;
                if (__any_sync(0xffffffffu, cand)) {
                    const double pp = exp(arg);
                    hit = cand && pp >= P.threshold;
                    pendP[j] = (int32_t) floor((pp > 1.0 ? 1.0 : pp) * 10000000.0);
                }
                S.hitM[d2 & 1][warp * C + j] = __ballot_sync(0xffffffffu, hit);   // every lane stores the same word
                pendHit[j] = hit; pendX[j] = x;
            }
            pendValid = true; pendD = d2; pendS0 = bw & (NS - 1);
            sa_mbArrive(&S.mbar);
        }
        sa_mbWait(&S.mbar, phase);             // the last diagonal's hit masks
        if (pendValid) {
            if (EX && HD) flushA(); else flush();
        }
    }

    // FWD = true: the forward sweep only, every row kept (ring of lX + lY + 1 rows = no wrap); the traceback blocks of
    // the item are then independent work items of sa_cta_block_kernel (split mode for reads that dominate a batch)
    template <bool FWD>
    static __device__ void run(const FParams &P, const FItem &I, int it, double *ring, Smem &S, unsigned &phase) {
        const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
        const double *T = P.trans;
        const int32_t nDiag = I.lX + I.lY;
        int64_t nPairs = 0;
        // expectations: MM, MX, XM, XX, YM, MY, YY sums of this thread's cells; likelihood (thread 0)
        double acc[7] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0}, lik = 0.0;
        bool yxNaN = false;
        if (nDiag == 0) {
            if (t == 0) {
                P.pairCount[it] = 0;
                if (EX && HD) P.assignCount[it] = 0;
                if (EX)
                    for (int i = 0; i < 10; i++) P.expOut[(int64_t) it * 10 + i] = 0.0;
            }
            return;
        }
        const int32_t *tbList = P.tb + I.tbOff;
        const int32_t nTb = P.geo[it].nTb;
        __syncthreads();                       // previous item finished with shared memory
        if (t == 0) S.sAi = 0;
        int32_t wlo = 0, cLo = 0, eLo = 0, bd0 = 1;
        cLo = wlo - NS / 2; fillConsts(P, I, S, t, cLo);
        eLo = (0 - (wlo + NS - 1)) - NS / 2; fillEvents(P, I, S, t, eLo);
        fillBand(P, I, S, t, bd0);             // (barriers inside)
        // diagonal 0: the cell (0, 0) with the start / ragged-start state (reference stateMachine.c:1134-1145)
        // f(d-1) / f(d-2) of the own cells by parity, and the same two of the column left of the thread's first cell
        double fa[C][3], fb[C][3], nbA[3], nbB[3];
        int32_t myx[C];
        {
            const bool ragged = I.flags & 1;
#pragma unroll
            for (int j = 0; j < C; j++) {
                const bool c0 = t == 0 && j == 0;
                myx[j] = slotX(t * C + j, wlo);
                fa[j][0] = (c0 && !ragged) ? 0.0 : SA_NEG_INF;
                fa[j][1] = (c0 && ragged) ? 0.0 : SA_NEG_INF;
                fa[j][2] = (c0 && ragged) ? 0.0 : SA_NEG_INF;
                fb[j][0] = fb[j][1] = fb[j][2] = SA_NEG_INF;
            }
            nbA[0] = nbA[1] = nbA[2] = SA_NEG_INF;
            nbB[0] = nbB[1] = nbB[2] = SA_NEG_INF;
        }
        {
            const double e0[C] = {};
            storeRow(ring, t, 0, 0, true, fa, e0);
        }
        publish(S, 0, t, fa[C - 1]);           // the right neighbour thread reads the LAST cell
        sa_mbArrive(&S.mbar);
        int32_t tracedBackTo = 0;
        int tbI = 0;
        int32_t nextT = nTb > 0 ? tbList[0] : nDiag;
        int32_t curFrom = nextT - (nextT == nDiag ? 0 : P.traceBack + 1);
        const double cOX = P.eX + T[T_GAP_OPEN_X], cEX = P.eX + T[T_GAP_EXTEND_X];
        const int32_t cap = FWD ? nDiag + 1 : P.rowsCap;
        int32_t rowI = 0;                       // ring row of diagonal d (= d mod cap, kept incrementally)
        int32_t backMod = (curFrom - 1) % 10;   // (curFrom - d) mod 10 while it is >= 0
        for (int32_t d = 1; d <= nDiag; d++) {
            rowI = rowI + 1 == cap ? 0 : rowI + 1;
            if (d >= bd0 + NT) { bd0 = d; fillBand(P, I, S, t, bd0); }
            const int2 bnd = S.bandTab[d - bd0];
            const int32_t xlo = bnd.x, xhi = bnd.y;
            const bool slid = xhi > wlo + NS - 1;
            if (slid) wlo = xhi - (NS - 1);    // slide the window up; recycled slots restart at -inf
            ensureRings(P, I, S, t, wlo, d, cLo, eLo);
            if (slid) {                        // (uniform branch: the per-slot checks run on slide steps only)
#pragma unroll
                for (int j = 0; j < C; j++) {
                    const int32_t x = slotX(t * C + j, wlo);
                    if (x != myx[j]) {
                        myx[j] = x;
                        fa[j][0] = fa[j][1] = fa[j][2] = SA_NEG_INF;
                        fb[j][0] = fb[j][1] = fb[j][2] = SA_NEG_INF;
                    }
                }
            }
            bool act[C];
            double eM[C], eY[C];
#pragma unroll
            for (int j = 0; j < C; j++) {
                const int32_t x = myx[j];
                act[j] = x >= xlo && x <= xhi;
                emissions(P, I, S, x, d - x, eM[j], eY[j]);
            }
            const bool full = EX || curFrom - d < 0 || backMod <= 1;   // the expectations read all three planes of every row
            backMod = backMod == 0 ? 9 : backMod - 1;
            double *row = ring + (int64_t) rowI * ROW;
            sa_mbWait(&S.mbar, phase);         // the mailbox of step d - 1 is visible
            // f(d-1) lives in fa when d is odd, in fb when d is even; the result overwrites f(d-2).  Descending j: cell j
            // reads f(d-2) of cell j - 1 before that cell overwrites it.
            if (d & 1) {
                fetchUp(S, (d - 1) & 1, t, nbA);
#pragma unroll
                for (int j = C - 1; j >= 0; j--) {
                    if (j > 0) forwardCell(P, S, myx[j], d, act[j], cOX, cEX, eM[j], eY[j], fa[j], fa[j > 0 ? j - 1 : 0], fb[j > 0 ? j - 1 : 0], fb[j]);
                    else forwardCell(P, S, myx[j], d, act[j], cOX, cEX, eM[j], eY[j], fa[j], nbA, nbB, fb[j]);
                }
                publish(S, d & 1, t, fb[C - 1]);
                sa_mbArrive(&S.mbar);
                storeRow(row, t, xlo, xhi, full, fb, eM);
            } else {
                fetchUp(S, (d - 1) & 1, t, nbB);
#pragma unroll
                for (int j = C - 1; j >= 0; j--) {
                    if (j > 0) forwardCell(P, S, myx[j], d, act[j], cOX, cEX, eM[j], eY[j], fb[j], fb[j > 0 ? j - 1 : 0], fa[j > 0 ? j - 1 : 0], fa[j]);
                    else forwardCell(P, S, myx[j], d, act[j], cOX, cEX, eM[j], eY[j], fb[j], nbB, nbA, fa[j]);
                }
                publish(S, d & 1, t, fa[C - 1]);
                sa_mbArrive(&S.mbar);
                storeRow(row, t, xlo, xhi, full, fa, eM);
            }
            if (d != nextT) continue;
            // ------------------------------------------------------------------ traceback from diagonal d
            const bool atEnd = d == nDiag;
            const int32_t tracedBackFrom = curFrom;
            if (!FWD)
                traceback(P, I, S, ring, t, lane, warp, d, rowI, xlo, xhi, wlo, cLo, eLo, tracedBackTo, tracedBackFrom,
                          atEnd, cOX, cEX, phase, nPairs, acc, lik, yxNaN, cap);
            tracedBackTo = tracedBackFrom;
            tbI++;
            nextT = tbI < nTb ? tbList[tbI] : nDiag;
            curFrom = nextT - (nextT == nDiag ? 0 : P.traceBack + 1);
            backMod = (curFrom - (d + 1)) % 10;
            if (atEnd) break;
            if (FWD) continue;                 // nothing disturbed the sweep
            // restore the forward sweep: f(d), f(d-1) from the ring (all three planes were kept above tracedBackFrom)
            __syncthreads();
            {
                const double *r0 = ring + (int64_t) rowI * ROW;
                const double *r1 = ring + (int64_t) (rowI == 0 ? cap - 1 : rowI - 1) * ROW;
                double cur[C][3], prv[C][3], nb[3];
#pragma unroll
                for (int s3 = 0; s3 < 3; s3++) {
                    double a[C], b[C];
                    loadRow(r0, s3, t, a);
                    loadRow(r1, s3, t, b);
#pragma unroll
                    for (int j = 0; j < C; j++) { cur[j][s3] = a[j]; prv[j][s3] = b[j]; }
                }
                publish(S, (d - 1) & 1, t, prv[C - 1]);
                __syncthreads();
                fetchUp(S, (d - 1) & 1, t, nb);
                publish(S, d & 1, t, cur[C - 1]);
                sa_mbArrive(&S.mbar);
#pragma unroll
                for (int s3 = 0; s3 < 3; s3++) {
#pragma unroll
                    for (int j = 0; j < C; j++) {
                        if (d & 1) { fb[j][s3] = cur[j][s3]; fa[j][s3] = prv[j][s3]; }
                        else { fa[j][s3] = cur[j][s3]; fb[j][s3] = prv[j][s3]; }
                    }
                    if (d & 1) nbA[s3] = nb[s3]; else nbB[s3] = nb[s3];
                }
#pragma unroll
                for (int j = 0; j < C; j++) myx[j] = slotX(t * C + j, wlo);
            }
        }
        if (EX) {
            // CTA reduction of the seven sums (summation order differs from the reference's cell order: ~1e-13 relative)
            __syncthreads();
            double *red = &S.foldv[0][0];
#pragma unroll
            for (int i = 0; i < 7; i++) {
                double v = acc[i];
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
                if (lane == 0) red[warp * 8 + i] = v;
            }
            const unsigned nanMask = __ballot_sync(0xffffffffu, yxNaN);
            if (lane == 0) red[warp * 8 + 7] = nanMask ? 1.0 : 0.0;
            __syncthreads();
            if (t == 0) {
                double v[8];
                for (int i = 0; i < 8; i++) {
                    v[i] = 0.0;
                    for (int w = 0; w < W; w++) v[i] += red[w * 8 + i];
                }
                double *o = P.expOut + (int64_t) it * 10;   // from * 3 + to, states M, X, Y
                o[0] = v[0]; o[1] = v[1]; o[2] = v[5];
                o[3] = v[2]; o[4] = v[3]; o[5] = 0.0;
                o[6] = v[4]; o[7] = v[7] != 0.0 ? __longlong_as_double(0x7ff8000000000000LL) : 0.0; o[8] = v[6];
                o[9] = lik;
            }
        }
        if (FWD) sa_mbWait(&S.mbar, phase);    // consume the last step's arrival: the next item starts with a clean phase
        if (t == 0 && !FWD) {
            if (EX && HD) { P.pairCount[it] = 0; P.assignCount[it] = nPairs; }
            else P.pairCount[it] = nPairs;
        }
    }

    // One traceback block of an item whose forward rows are all in `ring` (row index = diagonal): the body of the
    // reference's traceback point (impl/pairwiseAligner.c:1500-1570) as an independent work item.
    static __device__ void runBlock(const FParams &P, const FBlock &B, int blk, Smem &S, unsigned &phase) {
        const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
        const double *T = P.trans;
        FItem I = P.items[B.item];
        I.pairOff = B.pairOff; I.pairCap = B.pairCap;
        const int32_t nDiag = I.lX + I.lY, d = B.d;
        double *ring = P.scratch + I.fullOff;
        const int2 hdr = *reinterpret_cast<const int2 *>(ring + (int64_t) d * ROW);
        const int32_t xlo = hdr.x, xhi = hdr.y;
        __syncthreads();                       // previous work item finished with shared memory
        // any window [bw, bw + NS) over the band of d works: slots are columns mod NS, order is by column
        int32_t bw = xlo, cLo = bw - NS / 2, eLo = (d - (bw + NS - 1)) - NS / 2;
        fillConsts(P, I, S, t, cLo);
        fillEvents(P, I, S, t, eLo);
        __syncthreads();
        sa_mbArrive(&S.mbar);                  // stands for the forward step the traceback waits for
        int64_t nPairs = 0;
        double acc[7] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0}, lik = 0.0;
        bool yxNaN = false;
        const double cOX = P.eX + T[T_GAP_OPEN_X], cEX = P.eX + T[T_GAP_EXTEND_X];
        traceback(P, I, S, ring, t, lane, warp, d, d, xlo, xhi, bw, cLo, eLo, B.tracedBackTo, B.tracedBackFrom,
                  d == nDiag, cOX, cEX, phase, nPairs, acc, lik, yxNaN, nDiag + 1);
        if (t == 0) P.blockPairCount[blk] = nPairs;
    }
};

// resident CTAs per SM the register allocation is held to
template <int W, int C, bool EX, bool HD> struct CtaOcc {
#ifndef SA_HD_OCC
#define SA_HD_OCC 7
#endif
    static constexpr int kMinBlocks = (HD && !EX) ? (W == 1 ? 2 * SA_HD_OCC : W == 2 ? SA_HD_OCC : SA_HD_OCC / 2)
                                    : EX ? (W == 1 ? 12 : W == 2 ? 6 : 3)
                                         : C == 1 ? (W == 1 ? 16 : W == 2 ? 10 : W == 4 ? 5 : 2)
                                                  : (W == 1 ? 16 : W == 2 ? 8 : 4);
};

template <int W, int C, bool EX, bool HD>
__global__ void __launch_bounds__(32 * W, CtaOcc<W, C, EX, HD>::kMinBlocks) sa_cta_kernel(const FParams P) {
    __shared__ typename CtaK<W, C, EX, HD>::Smem S;
    if (threadIdx.x == 0) sa_mbInit(&S.mbar, 32 * W);
    sa_logAdd_init();                          // (ends with a barrier: the mbarrier is initialised for everybody)
    unsigned phase = 0;
    double *ring = P.scratch + (int64_t) blockIdx.x * P.rowsCap * CtaK<W, C, EX, HD>::ROW;
    for (;;) {
        __syncthreads();
        if (threadIdx.x == 0) S.sItem = atomicAdd(P.counter, 1);
        __syncthreads();
        const int n = S.sItem;
        if (n >= P.nOrder) break;
        const int it = P.order[n];
        CtaK<W, C, EX, HD>::template run<false>(P, P.items[it], it, ring, S, phase);
    }
}

// Split mode, phase 1: forward sweeps of the listed items into their full-length row buffers (P.scratch + fullOff).
template <int W, int C, bool HD>
__global__ void __launch_bounds__(32 * W, CtaOcc<W, C, false, HD>::kMinBlocks) sa_cta_forward_kernel(const FParams P) {
    __shared__ typename CtaK<W, C, false, HD>::Smem S;
    if (threadIdx.x == 0) sa_mbInit(&S.mbar, 32 * W);
    sa_logAdd_init();
    unsigned phase = 0;
    for (;;) {
        __syncthreads();
        if (threadIdx.x == 0) S.sItem = atomicAdd(P.counter, 1);
        __syncthreads();
        const int n = S.sItem;
        if (n >= P.nOrder) break;
        const int it = P.order[n];
        CtaK<W, C, false, HD>::template run<true>(P, P.items[it], it, P.scratch + P.items[it].fullOff, S, phase);
    }
}

// Split mode, phase 2: the traceback blocks of those items, each an independent work item (P.blocks, heaviest first).
template <int W, int C, bool HD>
__global__ void __launch_bounds__(32 * W, CtaOcc<W, C, false, HD>::kMinBlocks) sa_cta_block_kernel(const FParams P) {
    __shared__ typename CtaK<W, C, false, HD>::Smem S;
    if (threadIdx.x == 0) sa_mbInit(&S.mbar, 32 * W);
    sa_logAdd_init();
    unsigned phase = 0;
    for (;;) {
        __syncthreads();
        if (threadIdx.x == 0) S.sItem = atomicAdd(P.counter, 1);
        __syncthreads();
        const int n = S.sItem;
        if (n >= P.nBlocks) break;
        CtaK<W, C, false, HD>::runBlock(P, P.blocks[n], n, S, phase);
    }
}
