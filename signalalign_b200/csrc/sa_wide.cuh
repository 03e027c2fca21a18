// sa_wide.cuh — the WIDE band kernel: one CTA per work item, ANY band width up to the window NSW (a launch parameter:
// 128 .. 4096 slots), DP state resident in shared memory, cells of a diagonal dealt out to the threads round-robin.
//
// Why a second band kernel.  The register-resident kernel (sa_cta.cuh) ties a matrix column to a thread, so its window
// is 32 W C <= 256 columns and a thread idles whenever its column is outside the band.  Guide alignments trimmed the way
// signalMachine trims them (constraintDiagonalTrim = 14 from both ends of every match run, reference
// impl/pairwiseAligner.c:1624-1658, default :2029) leave sparse anchors: between two anchor runs the band is the whole
// rectangle they span (band_construct :195-246), hundreds to thousands of cells per diagonal, and narrow (expansion + 1)
// only next to the anchors.  Here:
//
//   * the forward state f(d-1), f(d-2) of every live cell sits in shared memory, three planes (M, X, Y) x two buffers
//     (diagonal parity).  A cell is addressed by its ANTI-DIAGONAL OFFSET: slot(d, x) = (x - ceil(d / 2) + H) mod NSW,
//     i.e. floor(xmy / 2) — the reference's own DpDiagonal index (dpDiagonal_getCell :1114-1120).  The `middle` neighbour
//     (x-1, y-1) on d-2 then has the SAME slot in the SAME buffer as the cell being computed, and it is read by that cell
//     only: the update is in place, two buffers suffice (48 B per slot), and no thread writes what another still reads.
//     `lower` (x-1, y) and `upper` (x, y-1) on d-1 are the slots j - 1 + (d & 1) and j + (d & 1) of the other buffer.
//     Slots outside the band of their diagonal hold -inf (kept so by clearing the few slots the band leaves behind), so
//     band edges need no tests in the cell code.
//   * cell i of diagonal d (column xlo(d) + i) goes to thread i mod NT: every thread is busy whatever the width, one
//     barrier per diagonal.  Per-column emission constants, event means and the band table live in shared-memory rings
//     refilled cooperatively, exactly as in sa_cta.cuh.
//   * forward rows go to a per-CTA HBM ring in COMPACT form (w(d) cells per plane, not the window size): the match plane
//     always, the X / Y planes only where the traceback reads them (the rows a total-probability refresh touches and the
//     rows kept for the next block); a header ring holds (xlo, xhi, offset) per row.
//   * the blocked traceback mirrors the forward sweep in gather form (reference accumulation order, see sa_general.cuh);
//     b(d2) overwrites b(d2 + 2) in place the same way.  Total probability: per-cell terms to shared memory, every warp
//     classifies its own 32-cell chunks (exact no-op and reset cells from the exclusive prefix maximum), warp 0 runs the
//     remaining short logAdd chain in the reference's order.  Posterior hits are recorded as one ballot word per chunk and
//     written one barrier later, in the reference's order, by the CTA's last warp.
//
// Arithmetic, accumulation order and outputs are those of sa_cta.cuh / sa_general.cuh: FP64, the reference's cubic
// logAdd, pairs bit-identical to the other kernels.
#pragma once
#include "sa_fast.cuh"

// EX: transition expectations + likelihood (diagonalCalculation_Expectations, reference impl/pairwiseAligner.c:1423-1443);
// HD: HDP emission (stateMachine3HDP_cellCalculate: one kmer density for the match and the gap-Y state);
// GR: the emission rings and the fold terms live in a per-CTA global-memory scratch (L1 / L2 resident) instead of shared
//     memory — the 4096-slot window, whose DP state alone takes 192 KB.
// NSWT / NTT: the window (slots) and the threads per CTA as compile-time constants (0: run-time values, for the helpers
// sa_wide32.cuh borrows) — masks, plane offsets and shared-memory addresses of the hot loops become immediates.
template <bool EX, bool HD, bool GR, int NSWT = 0, int NTT = 0>
struct WideK {
    // an opaque move: the value stays in a register instead of being rebuilt from special registers / kernel parameters
    template <typename T> static __device__ __forceinline__ T *pin(T *p) {
        T *q;
        asm volatile("mov.u64 %0, %1;" : "=l"(q) : "l"(p));
        return q;
    }
    static __device__ __forceinline__ int pinnedTid() {
        int t;
        asm volatile("mov.u32 %0, %%tid.x;" : "=r"(t));
        return t;
    }
    static constexpr int BT = 256;             // diagonals per descriptor table
    static constexpr double kExpZero = -50.0;  // expectation terms below e^-50 are not evaluated (see sa_cta.cuh)
    static constexpr int PL = HD ? 4 : 3;      // planes of a full stored row (HDP: + the cell's log density)

    // CTA-uniform view of the dynamic shared memory (all derived from one __shared__ base: LDS / STS addressing).
    // The per-slot arrays are ARRAYS OF STRUCTS (3 doubles of state, 5 doubles of constants per slot): one address per
    // slot, the planes at immediate offsets; strides of 24 B and 40 B keep 64-bit accesses of a warp conflict-free.
    struct Sm {
        double *st;        // [2][NSW][3]   DP state by diagonal parity, slot, (M, X, Y)
        double *cst;       // [NSW][5]      km, iM, cM, iY, cY per column (index = column & mask)
        double *evs;       // [NSW]         event mean per y (index = y & mask)
        double *foldv;     // [2][NSW]      total-probability terms by cell index (shared or global memory)
        float2 *cmax;      // [2][NSW/32]   per 32-cell chunk: max of the terms rounded down / up to float
        unsigned *candM;   // [2][NSW/32]   cells that can change the fold's accumulator
        unsigned *resetM;  // [2][NSW/32]   cells that replace it
        unsigned *hitW;    // [2][3][NSW/32] posterior hits per chunk, double buffered by diagonal parity (EX && HD: per from-state)
        int4 *desc;        // [BT]          per diagonal of the table: xlo, xhi, row offset (doubles), planes stored
        int4 *descB;       // [BT + 2]      the same for the traceback: entry q = header of diagonal (table top) - q
        int32_t *aX, *aY, *aD;   // [BT + 1] staged anchors
        double *sTotal;
        int32_t *sInt;     // [0] work item, [1] anchor cursor, [2] row-ring write offset after the table
        int32_t NSW, mask, nCh;  // nCh = NSW / 32
    };

    static __host__ __device__ size_t smemBytes(int NSW, bool foldInSmem) {
        size_t d = (size_t) 6 * NSW + (GR ? 0 : 5 * NSW + NSW + (foldInSmem ? 2 * NSW : 0)) + 2;   // doubles (+ sTotal, pad)
        size_t b = d * 8;
        b += (size_t) (NSW / 32) * (2 * 8 + 2 * 4 + 2 * 4 + 6 * 4);   // cmax, candM, resetM, hitW
        b += (size_t) (2 * BT + 2) * 16 + 3 * (BT + 1) * 4 + 16 + 16;
        return b;
    }
    static __device__ __forceinline__ void carve(Sm &S, double *base, int NSWarg, double *foldGlobal, double *ringsGlobal) {
        const int NSW = NSWT ? NSWT : NSWarg;
        S.NSW = NSW; S.mask = NSW - 1; S.nCh = NSW / 32;
        double *p = base;
        S.st = p; p += 6 * NSW;
        if (GR) { S.cst = pin(ringsGlobal); S.evs = S.cst + 5 * NSW; S.foldv = foldGlobal; }
        else {
            S.cst = p; p += 5 * NSW;
            S.evs = p; p += NSW;
            if (foldGlobal) S.foldv = foldGlobal; else { S.foldv = p; p += 2 * NSW; }
        }
        S.sTotal = p; p += 2;
        S.desc = reinterpret_cast<int4 *>(p); p += 2 * BT;          // (16-byte aligned: every count above is even)
        S.descB = reinterpret_cast<int4 *>(p); p += 2 * (BT + 2);
        S.cmax = reinterpret_cast<float2 *>(p); p += 2 * S.nCh;
        unsigned *u = reinterpret_cast<unsigned *>(p);
        S.candM = u; u += 2 * S.nCh;
        S.resetM = u; u += 2 * S.nCh;
        S.hitW = u; u += 6 * S.nCh;
        int32_t *q = reinterpret_cast<int32_t *>(u);
        S.aX = q; q += BT + 1;
        S.aY = q; q += BT + 1;
        S.aD = q; q += BT + 1;
        S.sInt = q;
    }

    // per-CTA forward-row storage: a ring of doubles (rows never straddle its end) and a ring of row headers
    struct Rows {
        double *ring;
        int4 *hdr;          // (xlo, xhi, offset in doubles, planes)
        int32_t cap;        // header entries (> rows of a traceback block + BT)
        int32_t ringDoubles;
    };

    // ------------------------------------------------------------------------------------------ rings
    // emission constants of column x into its ring slot (column 0 is the NULL kmer; columns outside [1, lX] are never
    // active, their entries only have to be harmless numbers)
    static __device__ __forceinline__ void fillConst1(const FParams &P, const FItem &I, const Sm &S, int32_t x) {
        double km = 0.0, iM = 0.0, iY = 0.0, cM = SA_NEG_INF, cY = SA_NEG_INF;
        if (HD) {
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
        double *c = S.cst + (x & S.mask) * 5;
        c[0] = km; c[1] = iM; c[2] = cM; c[3] = iY; c[4] = cY;
    }
    // Make the rings hold columns [cl, cl + span] and events [el, el + span] (span < NSW - 2 * kSlack).  cLo / eLo: first
    // index the ring currently holds (it holds NSW consecutive ones).  fwd: the sweep moves towards larger indices, so a
    // refill puts the slack ahead.  Uniform decisions; barriers only when refilling; only indices not yet present are loaded.
    static constexpr int kSlack = 8;
    static __device__ __forceinline__ void ensureRings(const FParams &P, const FItem &I, const Sm &S, int t, int NT,
                                                       int32_t cl, int32_t el, int32_t span, bool fwd, int32_t &cLo, int32_t &eLo) {
        const unsigned room = (unsigned) (S.NSW - 1 - span);            // cl - cLo may be 0 .. room
        const bool needC = (unsigned) (cl - cLo) > room, needE = (unsigned) (el - eLo) > room;
        if (!(needC || needE)) return;
        __syncthreads();                       // nobody still reads the entries that are replaced
        if (needC) {
            const int32_t nLo = fwd ? cl - kSlack : cl + span + kSlack - (S.NSW - 1);
            // new range [nLo, nLo + NSW); entries of the overlap with the old range stay
            for (int32_t k = t; k < S.NSW; k += NT) {
                const int32_t x = nLo + k;
                if (x < cLo || x >= cLo + S.NSW) fillConst1(P, I, S, x);
            }
            cLo = nLo;
        }
        if (needE) {
            const int32_t nLo = fwd ? el - kSlack : el + span + kSlack - (S.NSW - 1);
            for (int32_t k = t; k < S.NSW; k += NT) {
                const int32_t y = nLo + k;
                if (y < eLo || y >= eLo + S.NSW)
                    S.evs[y & S.mask] = (y >= 1 && y <= I.lY) ? P.evMean[I.evOff + y - 1] : 0.0;
            }
            eLo = nLo;
        }
        __syncthreads();
    }

    // Descriptor table of diagonals d0 .. d0 + BT - 1: the band (band_construct in closed form from the staged anchors,
    // see CtaK::fillBand), the planes the row keeps (all of them where the traceback reads the X / Y states: above
    // tracedBackFrom of the diagonal's block, on every 10th posterior diagonal and its predecessor) and the row's offset
    // in the CTA's row ring (prefix sum by warp 0; the table starts at the ring's origin when it would straddle its end).
    // The headers go to the global header ring in one coalesced write.  wr: ring write offset, rowI0: header slot of d0.
    static __device__ __forceinline__ void fillTable(const FParams &P, const FItem &I, const Sm &S, const Rows &RW, int t, int NT,
                                                     int32_t d0, int32_t rowI0, int tbI, int32_t nTb, const int32_t *tbList,
                                                     int32_t &wr) {
        __syncthreads();
        const int32_t base = S.sInt[1];        // anchors with diagonal < d0 - 1
        const int32_t *anch = P.anchors + 2 * I.anchorOff;
        for (int j = t; j <= BT; j += NT) {
            const int32_t a = base - 1 + j;
            int32_t X = 0, Y = 0, D = -0x7fffffff;
            if (a >= I.nAnchors) { X = I.lX; Y = I.lY; D = 0x7fffffff; }
            else if (a >= 0) { X = anch[2 * a] + 1; Y = anch[2 * a + 1] + 1; D = X + Y; }
            S.aX[j] = X; S.aY[j] = Y; S.aD[j] = D;
        }
        __syncthreads();
        const int32_t nDiag = I.lX + I.lY;
        int32_t cLast = 0;
        for (int q = t; q < BT; q += NT) {
            const int32_t d = d0 + q;
            int lo = 0, hi = BT;               // invariant: entries 1..lo are < d, entries hi+1.. are >= d
            while (lo < hi) {
                const int mid = (lo + hi + 1) >> 1;
                if (S.aD[mid] < d) lo = mid; else hi = mid - 1;
            }
            const int c = lo < BT ? lo : BT - 1;
            const int32_t pX = S.aX[c], pY = S.aY[c], nX = S.aX[c + 1], nY = S.aY[c + 1];
            const int32_t pxay = pX + pY, pxmy = pX - pY, nxay = nX + nY, nxmy = nX - nY, e = P.expansion;
            const int32_t xL = BandIt::clampc((pxay + pxmy - e) / 2, I.lX), yL = BandIt::clampc((nxay - nxmy + e) / 2, I.lY);
            const int32_t xU = BandIt::clampc((nxay + nxmy + e) / 2, I.lX), yU = BandIt::clampc((pxay - pxmy - e) / 2, I.lY);
            int32_t l = xL - yL, r = xU - yU, z;
            if ((d + l) & 1) l++;
            if ((d + r) & 1) r++;
            if ((z = xL - (d + l) / 2) > 0) l += 2 * z;
            if ((z = (d - l) / 2 - yL) > 0) l += 2 * z;
            if ((z = (d + r) / 2 - xU) > 0) r -= 2 * z;
            if ((z = yU - (d - r) / 2) > 0) r -= 2 * z;
            int4 D4 = make_int4(0, -1, 0, 0);
            if (d <= nDiag) {
                int k = tbI;                   // the block of d ends at the first traceback diagonal >= d
                while (k < nTb && tbList[k] < d) k++;
                const int32_t T = k < nTb ? tbList[k] : nDiag;
                const int32_t from = T - (T == nDiag ? 0 : P.traceBack + 1);
                const bool full = EX || d > from || (from - d) % 10 <= 1;
                D4 = make_int4((d + l) / 2, (d + r) / 2, 0, full ? PL : PL - 2);
            }
            S.desc[q] = D4;
            if (q == BT - 1) cLast = c;
        }
        if (((BT - 1) % NT) == t) S.sInt[1] = base + cLast;   // the thread that evaluated the table's last diagonal
        __syncthreads();
        if (t < 32) {                          // row offsets: exclusive prefix of planes * width, 8 entries per lane
            int32_t sz[BT / 32], sum = 0;
#pragma unroll
            for (int k = 0; k < BT / 32; k++) {
                const int4 D4 = S.desc[t * (BT / 32) + k];
                sz[k] = D4.w * (D4.y - D4.x + 1);
                sum += sz[k];
            }
            int32_t inc = sum;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int32_t v = __shfl_up_sync(0xffffffffu, inc, o);
                if (t >= o) inc += v;
            }
            const int32_t total = __shfl_sync(0xffffffffu, inc, 31);
            const int32_t start = wr + total > RW.ringDoubles ? 0 : wr;
            int32_t off = start + inc - sum;
#pragma unroll
            for (int k = 0; k < BT / 32; k++) {
                S.desc[t * (BT / 32) + k].z = off;
                off += sz[k];
            }
            if (t == 0) S.sInt[2] = start + total;
        }
        __syncthreads();
        wr = S.sInt[2];
        for (int q = t; q < BT; q += NT) {
            if (d0 + q > nDiag) break;
            int32_t r = rowI0 + q;
            if (r >= RW.cap) r -= RW.cap;
            RW.hdr[r] = S.desc[q];
        }
    }

    // ------------------------------------------------------------------------------------------ emissions
    static __device__ __forceinline__ double hdpLogDensity(const FParams &P, const FItem &I, long long row, double q) {
        const int32_t n = P.G;
        const double *yv = P.pp + row * n, *sv = P.slopes + row * n;
        const double x0 = P.grid[0], xn = P.grid[n - 1];
        double v;
        if (q <= x0) v = yv[0] - sv[0] * (x0 - q);
        else if (q >= xn) v = yv[n - 1] + sv[n - 1] * (q - xn);
        else {
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
    // match and gap-Y emission of cell (x, y) from the rings
    // ring read: a shared-memory load, or (rings in global memory) ld.global.ca said explicitly — the pinned base pointer is
    // opaque to the compiler, which would otherwise emit generic loads
    static __device__ __forceinline__ double ldR(const double *p) { return GR ? __ldca(p) : *p; }
    static __device__ __forceinline__ void emissions(const FParams &P, const FItem &I, const Sm &S, int32_t x, int32_t y,
                                                     double &eM, double &eY) {
        const double *c = S.cst + (x & S.mask) * 5;
        const double ev = ldR(S.evs + (y & S.mask));
        if (HD) {
            const long long row = __double_as_longlong(ldR(c + 2));
            const double q = (((ev + ldR(c)) - ldR(c + 1)) - I.shift) * I.invVar;
            double e = SA_NEG_INF;
            if (row >= 0) e = hdpLogDensity(P, I, row, q);
            eM = e; eY = e;
            return;
        }
        const double d0 = ev - ldR(c);
        const double aM = d0 * ldR(c + 1), aY = d0 * ldR(c + 3);
        eM = fma(-0.5 * aM, aM, ldR(c + 2));
        eY = fma(-0.5 * aY, aY, ldR(c + 4));
    }
    static __device__ __forceinline__ double emissionM(const FParams &P, const FItem &I, const Sm &S, int32_t x, int32_t y) {
        const double *c = S.cst + (x & S.mask) * 5;
        const double ev = ldR(S.evs + (y & S.mask));
        if (HD) {
            const long long row = __double_as_longlong(ldR(c + 2));
            const double q = (((ev + ldR(c)) - ldR(c + 1)) - I.shift) * I.invVar;
            return row >= 0 ? hdpLogDensity(P, I, row, q) : SA_NEG_INF;
        }
        const double aM = (ev - ldR(c)) * ldR(c + 1);
        return fma(-0.5 * aM, aM, ldR(c + 2));
    }
    static __device__ __forceinline__ double emissionY(const FParams &P, const FItem &I, const Sm &S, int32_t x, int32_t y) {
        if (HD) return emissionM(P, I, S, x, y);
        const double *c = S.cst + (x & S.mask) * 5;
        const double aY = (ldR(S.evs + (y & S.mask)) - ldR(c)) * ldR(c + 3);
        return fma(-0.5 * aY, aY, ldR(c + 4));
    }

    // ------------------------------------------------------------------------------------------ state helpers
    // jb(d): slot of column x on diagonal d is (x + jb(d)) & mask
    static __device__ __forceinline__ int32_t jbase(int32_t d, int32_t H) { return H - ((d + 1) >> 1); }

    // both buffers in one pass: buffer (dA & 1) <- valA on the band [loA, hiA] of dA, buffer (dB & 1) <- rows of dB (or
    // -inf everywhere when rowB == nullptr); everything else -inf.  valA: constant triple (traceback start) or a row.
    static __device__ __forceinline__ void loadState(const Sm &S, int t, int NT, int32_t H, int32_t dA, int32_t loA, int32_t hiA,
                                                     const double *rowA, const double (&cA)[3], int32_t dB, int32_t loB,
                                                     int32_t hiB, const double *rowB) {
        const int32_t jA = jbase(dA, H), jB = jbase(dB, H);
        double *bufA = S.st + (dA & 1) * 3 * S.NSW, *bufB = S.st + (dB & 1) * 3 * S.NSW;
        const int32_t wA = hiA - loA + 1, wB = hiB - loB + 1;
        for (int32_t s = t; s < S.NSW; s += NT) {
            {
                const int32_t i = (s - jA - loA) & S.mask;   // cell index of the column that owns slot s on dA
                double v0 = SA_NEG_INF, v1 = SA_NEG_INF, v2 = SA_NEG_INF;
                if (i < wA) {
                    if (rowA) { v0 = rowA[i]; v1 = rowA[wA + i]; v2 = rowA[2 * wA + i]; }
                    else { v0 = cA[0]; v1 = cA[1]; v2 = cA[2]; }
                }
                bufA[3 * s] = v0; bufA[3 * s + 1] = v1; bufA[3 * s + 2] = v2;
            }
            {
                const int32_t i = (s - jB - loB) & S.mask;
                double v0 = SA_NEG_INF, v1 = SA_NEG_INF, v2 = SA_NEG_INF;
                if (rowB && i < wB) { v0 = rowB[i]; v1 = rowB[wB + i]; v2 = rowB[2 * wB + i]; }
                bufB[3 * s] = v0; bufB[3 * s + 1] = v1; bufB[3 * s + 2] = v2;
            }
        }
    }

    // ------------------------------------------------------------------------------------------ total probability
    // chunk maximum of a fold term, rounded down (.x) and up (.y) to float: thresholds below keep 0.1 of slack, the two
    // roundings make the classification safe at any magnitude
    static __device__ __forceinline__ void chunkMax(const Sm &S, int f, int c, int lane, double v) {
        float lo = __double2float_rd(v), hi = __double2float_ru(v);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            lo = fmaxf(lo, __shfl_xor_sync(0xffffffffu, lo, o));
            hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, o));
        }
        if (lane == 0) S.cmax[f * S.nCh + c] = make_float2(lo, hi);
    }
    // Every warp classifies its chunks of fold f (n cells): with pm = exclusive prefix maximum in cell order,
    //   candidate  v >  pm - 7.6 : anything lower is an exact no-op (logAdd returns the accumulator, which is >= pm);
    //   reset      v >= pm + R   : the accumulator is < pm + ln(n) + n * 1.1e-4 (the cubic overshoots log(1 + e^-d) by less
    //                              than 1.1e-4), so |acc - v| >= 7.5 and the result is v itself whatever came before;
    //                              R = 7.5 + ln(4096) + 0.46 + slack = 16.5 covers every window size.
    static __device__ __forceinline__ void classify(const Sm &S, int f, int32_t n, int lane, int warp, int NW) {
        const int32_t nChunks = (n + 31) >> 5;
        for (int32_t c = warp; c < nChunks; c += NW) {
            const int32_t i = c * 32 + lane;
            const double v = i < n ? S.foldv[f * S.NSW + i] : SA_NEG_INF;
            float lo = -INFINITY, hi = -INFINITY;     // carry: maximum over the chunks before c
            for (int32_t k = lane; k < c; k += 32) {
                const float2 m = S.cmax[f * S.nCh + k];
                lo = fmaxf(lo, m.x); hi = fmaxf(hi, m.y);
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                lo = fmaxf(lo, __shfl_xor_sync(0xffffffffu, lo, o));
                hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, o));
            }
            float rl = __double2float_rd(v), rh = __double2float_ru(v);   // inclusive scan in cell order
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const float ql = __shfl_up_sync(0xffffffffu, rl, o), qh = __shfl_up_sync(0xffffffffu, rh, o);
                if (lane >= o) { rl = fmaxf(rl, ql); rh = fmaxf(rh, qh); }
            }
            float pl = __shfl_up_sync(0xffffffffu, rl, 1), ph = __shfl_up_sync(0xffffffffu, rh, 1);
            if (lane == 0) { pl = -INFINITY; ph = -INFINITY; }
            pl = fmaxf(pl, lo); ph = fmaxf(ph, hi);
            const unsigned mc = __ballot_sync(0xffffffffu, v > (double) pl - 7.6);    // false for -inf
            const unsigned mr = __ballot_sync(0xffffffffu, v >= (double) ph + 16.5);
            if (lane == 0) { S.candM[f * S.nCh + c] = mc; S.resetM[f * S.nCh + c] = mr; }
        }
    }
    // warp 0: the reference's two sequential folds from their last reset cells, both chains in flight
    static __device__ __forceinline__ double foldChains(const Sm &S, int32_t n0, int32_t n1, int lane) {
        int32_t R[2];
#pragma unroll
        for (int f = 0; f < 2; f++) {
            const int32_t nChunks = ((f == 0 ? n0 : n1) + 31) >> 5;
            int32_t best = 0;
            for (int32_t c = lane; c < nChunks; c += 32) {
                const unsigned m = S.resetM[f * S.nCh + c];
                if (m) best = max(best, c * 32 + 31 - __clz(m));
            }
            R[f] = __reduce_max_sync(0xffffffffu, best);
        }
        const int32_t nc0 = (n0 + 31) >> 5, nc1 = (n1 + 31) >> 5;
        int32_t c0 = R[0] >> 5, c1 = R[1] >> 5;
        unsigned m0 = c0 < nc0 ? (S.candM[c0] & (0xffffffffu << (R[0] & 31))) : 0u;
        unsigned m1 = c1 < nc1 ? (S.candM[S.nCh + c1] & (0xffffffffu << (R[1] & 31))) : 0u;
        double acc0 = SA_NEG_INF, acc1 = SA_NEG_INF;
        for (;;) {                              // (uniform: every lane runs the same chain)
            while (!m0 && ++c0 < nc0) m0 = S.candM[c0];
            while (!m1 && ++c1 < nc1) m1 = S.candM[S.nCh + c1];
            if (!(m0 | m1)) break;
            if (m0) {
                const int i = __ffs(m0) - 1;
                m0 &= m0 - 1;
                acc0 = sa_logAdd(acc0, S.foldv[c0 * 32 + i]);
            }
            if (m1) {
                const int i = __ffs(m1) - 1;
                m1 &= m1 - 1;
                acc1 = sa_logAdd(acc1, S.foldv[S.NSW + c1 * 32 + i]);
            }
        }
        return sa_logAdd(acc0, acc1);
    }

    // ------------------------------------------------------------------------------------------ traceback
    // blocked traceback from diagonal d down to tracedBackTo + 1 (reference impl/pairwiseAligner.c:1500-1570)
    static __device__ __forceinline__ void traceback(const FParams &P, const FItem &I, const Sm &S, const Rows &RW, int t, int NT, int32_t H,
                                     int32_t d, int32_t rowI, int32_t &cLo, int32_t &eLo, int32_t tracedBackTo,
                                     int32_t tracedBackFrom, bool atEnd, double cOX, double cEX, int64_t &nPairs,
                                     double (&acc)[7], double &lik, bool &yxNaN) {
        const int lane = t & 31, warp = t >> 5, NW = NT >> 5;
        const double *T = P.trans;
        const int32_t cap = RW.cap;
        // end / ragged-end state on diagonal d (reference stateMachine.c:1147-1173)
        const bool rag = atEnd && (I.flags & 2);
        const double endS[3] = {rag ? (T[T_GAP_OPEN_X] + T[T_GAP_OPEN_Y]) / 2.0 : T[T_MATCH_CONTINUE],
                                rag ? T[T_GAP_EXTEND_X] : T[T_MATCH_FROM_GAP_X], rag ? T[T_GAP_EXTEND_Y] : T[T_MATCH_FROM_GAP_Y]};
        // row headers of the next BT steps from the global header ring into shared memory (entry q = diagonal top - q; two
        // more for the total-probability / expectation reads of d2 - 1 and d2 - 2): one LDS.128 per step instead of global loads
        auto fillBack = [&](int32_t rowTop) {
            for (int32_t q = t; q < BT + 2; q += NT) {
                int32_t r = rowTop - q;
                if (r < 0) r += cap;
                S.descB[q] = RW.hdr[r];
            }
        };
        int32_t tTop = d;
        {
            const int4 hTop = RW.hdr[rowI];
            __syncthreads();                    // the forward step's reads of the state are over
            loadState(S, t, NT, H, d, hTop.x, hTop.y, nullptr, endS, d - 1, 0, -1, nullptr);
            fillBack(rowI);
            __syncthreads();
        }
        double total = SA_NEG_INF;
        int32_t postMod = 0;                    // posterior diagonals since the last total, mod 10
        int4 hUp = make_int4(0, -1, 0, 0);      // header of d2 + 1
        int4 hUp2 = make_int4(0, -1, 0, 0);     // header of d2 + 2
        int4 hCur = make_int4(0, -1, 0, 0);
        // pending posterior hits of the previous diagonal (hit words in S.hitW[pendD & 1])
        bool pend = false;
        int32_t pendD = 0, pendXlo = 0, pendW = 0;
        int32_t pendOff = 0;
        double pendTotal = 0.0;
        auto flush = [&]() {
            // the CTA's last warp: hit words in chunk order; lane k owns the words k, k + 32, ...
            if (warp == NW - 1) {
                const int32_t nWords = (pendW + 31) >> 5;
                {                               // most diagonals have no hit at all
                    unsigned any = 0u;
                    for (int32_t c = lane; c < nWords; c += 32) any |= S.hitW[(pendD & 1) * 3 * S.nCh + c];
                    if (!__any_sync(0xffffffffu, any != 0u)) { pend = false; return; }
                }
                const double *rowM = RW.ring + pendOff;
                const int32_t jb = jbase(pendD, H);
                const double *bM = S.st + (pendD & 1) * 3 * S.NSW;   // (array of structs: M at 3 * slot)
                int64_t base = nPairs;
                for (int32_t c0 = 0; c0 < nWords; c0 += 32) {
                    const int32_t c = c0 + lane;
                    unsigned wv = c < nWords ? S.hitW[(pendD & 1) * 3 * S.nCh + c] : 0u;
                    const int cnt = __popc(wv);
                    int inc = cnt;
#pragma unroll
                    for (int o = 1; o < 32; o <<= 1) {
                        const int q = __shfl_up_sync(0xffffffffu, inc, o);
                        if (lane >= o) inc += q;
                    }
                    int64_t w = base + (inc - cnt);
                    while (wv) {
                        const int bit = __ffs(wv) - 1;
                        wv &= wv - 1;
                        const int32_t i = c * 32 + bit, x = pendXlo + i;
                        const double pp = exp((rowM[i] + bM[3 * ((x + jb) & S.mask)]) - pendTotal);
                        if (w < I.pairCap) {
                            sa_pair_t o;
                            o.p = (int32_t) floor((pp > 1.0 ? 1.0 : pp) * 10000000.0);
                            o.x = x - 1; o.y = pendD - x - 1;
                            o.kmerIdx = P.kmerIdx[I.kmerOff + x - 1];
                            P.pairs[I.pairOff + w] = o;
                        }
                        w++;
                    }
                    base += __shfl_sync(0xffffffffu, inc, 31);
                }
                nPairs = base;
            }
            pend = false;
        };
        // HDP expectations: assignments of the previous diagonal, cell order then from-state M, X, Y
        auto flushA = [&]() {
            if (warp == NW - 1) {
                const int32_t nWords = (pendW + 31) >> 5;
                int64_t base = nPairs;
                for (int32_t c0 = 0; c0 < nWords; c0 += 32) {
                    const int32_t c = c0 + lane;
                    unsigned wv[3];
                    int cnt = 0;
#pragma unroll
                    for (int s3 = 0; s3 < 3; s3++) {
                        wv[s3] = c < nWords ? S.hitW[((pendD & 1) * 3 + s3) * S.nCh + c] : 0u;
                        cnt += __popc(wv[s3]);
                    }
                    int inc = cnt;
#pragma unroll
                    for (int o = 1; o < 32; o <<= 1) {
                        const int q = __shfl_up_sync(0xffffffffu, inc, o);
                        if (lane >= o) inc += q;
                    }
                    int64_t w = base + (inc - cnt);
                    unsigned any = wv[0] | wv[1] | wv[2];
                    while (any) {
                        const int bit = __ffs(any) - 1;
                        any &= any - 1;
                        const int32_t x = pendXlo + c * 32 + bit;
#pragma unroll
                        for (int s3 = 0; s3 < 3; s3++) {
                            if (!((wv[s3] >> bit) & 1u)) continue;
                            if (w < I.assignCap) {
                                P.assignK[I.assignOff + w] = x - 1 >= 0 ? x - 1 : 0;
                                P.assignE[I.assignOff + w] = pendD - x - 1;
                            }
                            w++;
                        }
                    }
                    base += __shfl_sync(0xffffffffu, inc, 31);
                }
                nPairs = base;
            }
            pend = false;
        };
        for (int32_t d2 = d; d2 > tracedBackTo; d2--) {
            int32_t q = tTop - d2;
            if (q == BT) {                      // (the previous step's barrier separates the table's last readers from the refill)
                fillBack(rowI);
                tTop = d2; q = 0;
                __syncthreads();
            }
            hCur = S.descB[q];
            const int32_t bxlo = hCur.x, bxhi = hCur.y, w = bxhi - bxlo + 1, par = d2 & 1, jb = jbase(d2, H);
            const bool post = d2 <= tracedBackFrom;
            double *cur = S.st + par * 3 * S.NSW;                 // b(d2 + 2) -> b(d2)
            const double *prv = S.st + (par ^ 1) * 3 * S.NSW;     // b(d2 + 1)
            const double *row = RW.ring + hCur.z;
            {
                // what this step evaluates emissions for: cells (x + 1, y + 1), (x, y + 1), (x, y) of the band of d2 and the
                // cells of d2 + 1 (total probability)
                const bool up = hUp.y >= hUp.x;
                const int32_t cl = up ? min(bxlo, hUp.x) : bxlo, ch = up ? max(bxhi + 1, hUp.y) : bxhi + 1;
                const int32_t el = up ? min(d2 - bxhi, d2 + 1 - hUp.y) : d2 - bxhi;
                const int32_t eh = up ? max(d2 - bxlo + 1, d2 + 1 - hUp.x) : d2 - bxlo + 1;
                ensureRings(P, I, S, t, NT, cl, el, max(ch - cl, eh - el), false, cLo, eLo);
            }
            // posterior diagonals since the last total-probability refresh, mod 10 (reference :1538)
            const bool newTotal = post && postMod == 0;
            if (post) postMod = postMod == 9 ? 0 : postMod + 1;
            // the posterior of a cell needs nothing but its own b(d2) and the running total: evaluated in the same pass
            // unless this diagonal refreshes the total (or starts the traceback, where b(d2) is the end state)
            const bool inlinePost = !EX && post && !newTotal;
            if (d2 < d) {
                const int32_t nChunks = (w + 31) >> 5;
                for (int32_t c = warp; c < nChunks; c += NW) {
                    const int32_t i = c * 32 + lane;
                    bool cand = false;
                    double arg = 0.0;
                    // the stored forward value comes from L2 or HBM: asked for before the cell's arithmetic, used after it
                    const double fM = (inlinePost && i < w) ? __ldcs(row + i) : 0.0;
                    if (i < w) {
                        const int32_t x = bxlo + i, y = d2 - x;
                        const int32_t j = (x + jb) & S.mask, sq = (j - 1 + par) & S.mask, sl = (sq + 1) & S.mask;
                        // a = (x + 1, y + 1) on d2 + 2 (same slot), q = (x, y + 1) and l = (x + 1, y) on d2 + 1
                        const double aM = cur[3 * j], qY = prv[3 * sq + 2], lX = prv[3 * sl + 1];
                        double aE, qE;
                        if (HD) {               // the forward sweep kept every cell's log density
                            aE = qE = 0.0;
                            if (hUp2.y >= hUp2.x && x + 1 >= hUp2.x && x + 1 <= hUp2.y)
                                aE = (RW.ring + hUp2.z)[(hUp2.w - 1) * (hUp2.y - hUp2.x + 1) + (x + 1 - hUp2.x)];
                            if (x >= hUp.x && x <= hUp.y)
                                qE = (RW.ring + hUp.z)[(hUp.w - 1) * (hUp.y - hUp.x + 1) + (x - hUp.x)];
                        } else {
                            aE = emissionM(P, I, S, x + 1, y + 1);
                            qE = emissionY(P, I, S, x, y + 1);
                        }
                        // gather in the order the reference's scatter delivers (see sa_general.cuh)
                        const double M = sa_logAdd(sa_logAdd(aM + (aE + T[T_MATCH_CONTINUE]), qY + (qE + T[T_GAP_OPEN_Y])), lX + cOX);
                        const double X = sa_logAdd(aM + (aE + T[T_MATCH_FROM_GAP_X]), lX + cEX);
                        const double Y = sa_logAdd(aM + (aE + T[T_MATCH_FROM_GAP_Y]), qY + (qE + T[T_GAP_EXTEND_Y]));
                        cur[3 * j] = M; cur[3 * j + 1] = X; cur[3 * j + 2] = Y;
                        if (inlinePost) {
                            arg = (fM + M) - total;
                            cand = x > 0 && y > 0 && arg >= P.logThrGate;
                        }
                    }
                    if (inlinePost) {
                        bool hit = false;
                        if (__any_sync(0xffffffffu, cand)) hit = cand && exp(arg) >= P.threshold;
                        const unsigned m = __ballot_sync(0xffffffffu, hit);
                        if (lane == 0) S.hitW[par * 3 * S.nCh + c] = m;
                    }
                }
                // slots of d2 + 2 whose cell (x', .) has no successor (x' - 1, .) on d2: back to -inf
                if (hUp2.y >= hUp2.x) {
                    const int32_t jb2 = jb - 1;   // jbase(d2 + 2)
                    const int32_t nLow = max(0, min(hUp2.y, bxlo) - hUp2.x + 1);          // x' <= bxlo
                    const int32_t hiStart = max(hUp2.x, bxhi + 2);                         // x' >= bxhi + 2
                    const int32_t nHigh = max(0, hUp2.y - hiStart + 1);
                    for (int32_t k = NT - 1 - t; k < nLow + nHigh; k += NT) {
                        const int32_t xq = k < nLow ? hUp2.x + k : hiStart + (k - nLow);
                        const int32_t s = (xq + jb2) & S.mask;
                        cur[3 * s] = SA_NEG_INF; cur[3 * s + 1] = SA_NEG_INF; cur[3 * s + 2] = SA_NEG_INF;
                    }
                }
            }
            if (pend) {
                if (EX && HD) flushA(); else flush();
            }
            if (post) {
                if (newTotal) {
                    // diagonalCalculationTotalProbability (reference pairwiseAligner.c:1335-1353)
                    const int32_t nChunks = (w + 31) >> 5;
                    for (int32_t c = warp; c < nChunks; c += NW) {
                        const int32_t i = c * 32 + lane;
                        double v = SA_NEG_INF;
                        if (i < w) {
                            const int32_t j = (bxlo + i + jb) & S.mask;
                            v = row[i] + cur[3 * j];
                            v = sa_logAdd(v, row[w + i] + cur[3 * j + 1]);
                            v = sa_logAdd(v, row[2 * w + i] + cur[3 * j + 2]);
                            S.foldv[i] = v;
                        }
                        chunkMax(S, 0, c, lane, v);
                    }
                    // matches jumping over d2: middle-only forward step from f(d2 - 1) into the cells of d2 + 1
                    const int4 hDn = S.descB[q + 1];     // header of d2 - 1 (row tracedBackTo still exists)
                    const int32_t w1 = hUp.y - hUp.x + 1, wD = hDn.y - hDn.x + 1;
                    const int32_t n1 = w1 > 0 ? w1 : 0;
                    {
                        const double *rowD = RW.ring + hDn.z;
                        const int32_t jb1 = jbase(d2 + 1, H);
                        const int32_t nChunks1 = (n1 + 31) >> 5;
                        for (int32_t c = warp; c < nChunks1; c += NW) {
                            const int32_t i = c * 32 + lane;
                            double v = SA_NEG_INF;
                            if (i < n1) {
                                const int32_t x = hUp.x + i, y = d2 + 1 - x, iD = x - 1 - hDn.x;
                                double gM = SA_NEG_INF, gX = SA_NEG_INF, gY = SA_NEG_INF;
                                if (iD >= 0 && iD < wD) { gM = rowD[iD]; gX = rowD[wD + iD]; gY = rowD[2 * wD + iD]; }
                                double e;
                                if (HD) e = (RW.ring + hUp.z)[(hUp.w - 1) * w1 + i];
                                else e = emissionM(P, I, S, x, y);
                                const double m = sa_logAdd(sa_logAdd(gM + (e + T[T_MATCH_CONTINUE]), gX + (e + T[T_MATCH_FROM_GAP_X])),
                                                           gY + (e + T[T_MATCH_FROM_GAP_Y]));
                                v = m + prv[3 * ((x + jb1) & S.mask)];
                                S.foldv[S.NSW + i] = v;
                            }
                            chunkMax(S, 1, c, lane, v);
                        }
                    }
                    __syncthreads();
                    classify(S, 0, w, lane, warp, NW);
                    classify(S, 1, n1, lane, warp, NW);
                    __syncthreads();
                    if (warp == 0) {
                        const double tot = foldChains(S, w, n1, lane);
                        if (lane == 0) *S.sTotal = tot;
                    }
                    __syncthreads();
                    total = *S.sTotal;
                }
                if (EX) {
                    // diagonalCalculation_Expectations: current = b(d2), neighbours from f(d2 - 1) and f(d2 - 2); a neighbour
                    // counts when its cell lies inside the band of its diagonal and the diagonal was not freed
                    // (d2 - 2 >= tracedBackTo).  Every term is exp(((f + b) + (eP + tP)) - total), as the reference adds it.
                    if (t == 0) lik += total;
                    const bool has2 = d2 - 2 >= tracedBackTo;
                    const int4 hDn = S.descB[q + 1];
                    int4 h2 = make_int4(1, 0, 0, 0);
                    if (has2) h2 = S.descB[q + 2];
                    const int32_t w1 = hDn.y - hDn.x + 1, w2 = h2.y - h2.x + 1;
                    const double *r1 = RW.ring + hDn.z, *r2 = RW.ring + h2.z;
                    const int32_t nChunks = (w + 31) >> 5;
                    for (int32_t c = warp; c < nChunks; c += NW) {
                        const int32_t i = c * 32 + lane;
                        const bool here = i < w;
                        const int32_t x = bxlo + i, y = d2 - x, j = (x + jb) & S.mask;
                        const bool lower = here && x - 1 >= hDn.x && x - 1 <= hDn.y;
                        const bool upper = here && x >= hDn.x && x <= hDn.y;
                        const bool middle = here && x - 1 >= h2.x && x - 1 <= h2.y;
                        double bM = 0.0, bX = 0.0, bY = 0.0, eM = 0.0, eY = 0.0;
                        if (here) {
                            bM = cur[3 * j]; bX = cur[3 * j + 1]; bY = cur[3 * j + 2];
                            if (HD) eM = eY = row[3 * w + i];
                            else emissions(P, I, S, x, y, eM, eY);
                        }
                        if (__any_sync(0xffffffffu, lower)) {
                            const int32_t i1 = lower ? x - 1 - hDn.x : 0;
                            const double f0 = lower ? r1[i1] : 0.0, f1 = lower ? r1[w1 + i1] : 0.0;
                            const double a0 = ((f0 + bX) + cOX) - total, a1 = ((f1 + bX) + cEX) - total;
                            if (__any_sync(0xffffffffu, lower && !(a0 < kExpZero))) acc[1] += lower ? sa_expSum(a0) : 0.0;
                            if (__any_sync(0xffffffffu, lower && !(a1 < kExpZero))) acc[3] += lower ? sa_expSum(a1) : 0.0;
                            yxNaN |= lower && !(total > SA_NEG_INF);
                        }
                        double pr[3] = {0.0, 0.0, 0.0};
                        if (__any_sync(0xffffffffu, middle)) {
                            const int32_t i2 = middle ? x - 1 - h2.x : 0;
                            const double f0 = middle ? r2[i2] : 0.0, f1 = middle ? r2[w2 + i2] : 0.0, f2 = middle ? r2[2 * w2 + i2] : 0.0;
                            const double a0 = ((f0 + bM) + (eM + T[T_MATCH_CONTINUE])) - total;
                            const double a1 = ((f1 + bM) + (eM + T[T_MATCH_FROM_GAP_X])) - total;
                            const double a2 = ((f2 + bM) + (eM + T[T_MATCH_FROM_GAP_Y])) - total;
                            if (__any_sync(0xffffffffu, middle && !(a0 < kExpZero))) { pr[0] = middle ? sa_expSum(a0) : 0.0; acc[0] += pr[0]; }
                            if (__any_sync(0xffffffffu, middle && !(a1 < kExpZero))) { pr[1] = middle ? sa_expSum(a1) : 0.0; acc[2] += pr[1]; }
                            if (__any_sync(0xffffffffu, middle && !(a2 < kExpZero))) { pr[2] = middle ? sa_expSum(a2) : 0.0; acc[4] += pr[2]; }
                        }
                        if (HD) {
#pragma unroll
                            for (int s3 = 0; s3 < 3; s3++) {
                                const unsigned m = __ballot_sync(0xffffffffu, middle && pr[s3] >= P.threshold);
                                if (lane == 0) S.hitW[(par * 3 + s3) * S.nCh + c] = m;
                            }
                        }
                        if (__any_sync(0xffffffffu, upper)) {
                            const int32_t i1 = upper ? x - hDn.x : 0;
                            const double f0 = upper ? r1[i1] : 0.0, f2 = upper ? r1[2 * w1 + i1] : 0.0;
                            const double a0 = ((f0 + bY) + (eY + T[T_GAP_OPEN_Y])) - total;
                            const double a2 = ((f2 + bY) + (eY + T[T_GAP_EXTEND_Y])) - total;
                            if (__any_sync(0xffffffffu, upper && !(a0 < kExpZero))) acc[5] += upper ? sa_expSum(a0) : 0.0;
                            if (__any_sync(0xffffffffu, upper && !(a2 < kExpZero))) acc[6] += upper ? sa_expSum(a2) : 0.0;
                        }
                    }
                    if (HD) { pend = true; pendD = d2; pendXlo = bxlo; pendW = w; }
                } else {
                    // diagonalCalculationPosteriorMatchProbs (reference pairwiseAligner.c:1355-1421)
                    const int32_t nChunks = inlinePost && d2 < d ? 0 : (w + 31) >> 5;
                    for (int32_t c = warp; c < nChunks; c += NW) {
                        const int32_t i = c * 32 + lane, x = bxlo + i;
                        bool cand = false;
                        double arg = 0.0;
                        if (i < w) {
                            arg = (row[i] + cur[3 * ((x + jb) & S.mask)]) - total;
                            cand = x > 0 && d2 - x > 0 && arg >= P.logThrGate;
                        }
                        bool hit = false;
                        if (__any_sync(0xffffffffu, cand)) hit = cand && exp(arg) >= P.threshold;
                        const unsigned m = __ballot_sync(0xffffffffu, hit);
                        if (lane == 0) S.hitW[par * 3 * S.nCh + c] = m;
                    }
                    pend = true; pendD = d2; pendXlo = bxlo; pendW = w; pendOff = hCur.z; pendTotal = total;
                }
            }
            __syncthreads();                    // b(d2) and the hit words are complete
            hUp2 = hUp; hUp = hCur;
            rowI = rowI == 0 ? cap - 1 : rowI - 1;
        }
        if (pend) {
            if (EX && HD) flushA(); else flush();
        }
    }

    // ------------------------------------------------------------------------------------------ one work item
    static __device__ __forceinline__ void run(const FParams &P, const FItem &I, int it, const Sm &S, const Rows &RWin) {
        const int NT = NTT ? NTT : (int) blockDim.x, NW = NT >> 5;
        const int t = pinnedTid(), lane = t & 31, warp = t >> 5;
        Rows RW = RWin;
        RW.ring = pin(RWin.ring);
        const double *T = P.trans;
        const int32_t nDiag = I.lX + I.lY;
        int64_t nPairs = 0;
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
        const int32_t H = I.lY + 2;             // keeps x - ceil(d / 2) + H positive
        __syncthreads();                        // the previous item is done with shared memory
        if (t == 0) S.sInt[1] = 0;
        int32_t cLo = -kSlack, eLo = -kSlack, bd0 = 1;
        for (int32_t k = t; k < S.NSW; k += NT) {
            fillConst1(P, I, S, cLo + k);
            const int32_t y = eLo + k;
            S.evs[y & S.mask] = (y >= 1 && y <= I.lY) ? P.evMean[I.evOff + y - 1] : 0.0;
        }
        // diagonal 0: the cell (0, 0) with the start / ragged-start state (reference stateMachine.c:1134-1145)
        {
            const bool ragged = I.flags & 1;
            const double s0[3] = {ragged ? SA_NEG_INF : 0.0, ragged ? 0.0 : SA_NEG_INF, ragged ? 0.0 : SA_NEG_INF};
            loadState(S, t, NT, H, 0, 0, 0, nullptr, s0, -1, 0, -1, nullptr);
            if (t == 0) {
                RW.ring[0] = s0[0]; RW.ring[1] = s0[1]; RW.ring[2] = s0[2];
                if (HD) RW.ring[3] = 0.0;
                RW.hdr[0] = make_int4(0, 0, 0, PL);
            }
        }
        int32_t tracedBackTo = 0;
        int tbI = 0;
        int32_t nextT = nTb > 0 ? tbList[0] : nDiag;
        int32_t curFrom = nextT - (nextT == nDiag ? 0 : P.traceBack + 1);
        const double cOX = P.eX + T[T_GAP_OPEN_X], cEX = P.eX + T[T_GAP_EXTEND_X];
        int32_t rowI = 0;                       // header-ring row of diagonal d
        int32_t wr = PL;                        // next free double of the row ring
        fillTable(P, I, S, RW, t, NT, bd0, 1, tbI, nTb, tbList, wr);
        __syncthreads();
        int32_t xlo1 = 0, xhi1 = 0, xlo2 = 0, xhi2 = -1;   // bands of d - 1 and d - 2
        for (int32_t d = 1; d <= nDiag; d++) {
            rowI = rowI + 1 == RW.cap ? 0 : rowI + 1;
            if (d >= bd0 + BT) {
                bd0 = d;
                fillTable(P, I, S, RW, t, NT, bd0, rowI, tbI, nTb, tbList, wr);
                __syncthreads();
            }
            const int4 D4 = S.desc[d - bd0];
            const int32_t xlo = D4.x, xhi = D4.y, w = xhi - xlo + 1, par = d & 1, jb = jbase(d, H);
            ensureRings(P, I, S, t, NT, xlo, d - xhi, w - 1, true, cLo, eLo);
            const bool full = D4.w == PL;
            double *row = RW.ring + D4.z;
            double *cur = S.st + par * 3 * S.NSW;                       // f(d - 2) -> f(d)
            const double *prv = S.st + (par ^ 1) * 3 * S.NSW;           // f(d - 1)
            const int32_t nChunks = (w + 31) >> 5;
            for (int32_t c = warp; c < nChunks; c += NW) {
                const int32_t i = c * 32 + lane;
                if (i < w) {
                    const int32_t x = xlo + i, y = d - x;
                    const int32_t j = (x + jb) & S.mask, sl = (j - 1 + par) & S.mask, su = (sl + 1) & S.mask;
                    double eM, eY;
                    emissions(P, I, S, x, y, eM, eY);
                    double *pj = cur + 3 * j;
                    const double *pl = prv + 3 * sl, *pu = prv + 3 * su;
                    const double n1M = pl[0], n1X = pl[1];
                    const double f1M = pu[0], f1Y = pu[2];
                    const double n2M = pj[0], n2X = pj[1], n2Y = pj[2];
                    // stateMachine3_cellCalculate order; the first contribution lands on -inf (logAdd(-inf, v) == v exactly)
                    const double X = sa_logAdd(n1M + cOX, n1X + cEX);
                    const double M = sa_logAdd(sa_logAdd(n2M + (eM + T[T_MATCH_CONTINUE]), n2X + (eM + T[T_MATCH_FROM_GAP_X])),
                                               n2Y + (eM + T[T_MATCH_FROM_GAP_Y]));
                    const double Y = sa_logAdd(f1M + (eY + T[T_GAP_OPEN_Y]), f1Y + (eY + T[T_GAP_EXTEND_Y]));
                    pj[0] = M; pj[1] = X; pj[2] = Y;
                    __stcs(row + i, M);                 // (streaming: the rows must not push the constant rings out of L1)
                    if (full) { __stcs(row + w + i, X); __stcs(row + 2 * w + i, Y); }
                    if (HD) __stcs(row + (D4.w - 1) * w + i, eM);
                }
            }
            // slots of d - 2 whose cell (x', .) has no successor (x' + 1, .) on d: back to -inf
            if (xhi2 >= xlo2) {
                const int32_t jb2 = jb + 1;     // jbase(d - 2)
                const int32_t nLow = max(0, min(xhi2, xlo - 2) - xlo2 + 1);       // x' <= xlo - 2
                const int32_t hiStart = max(xlo2, xhi);                            // x' >= xhi
                const int32_t nHigh = max(0, xhi2 - hiStart + 1);
                for (int32_t k = NT - 1 - t; k < nLow + nHigh; k += NT) {
                    const int32_t xq = k < nLow ? xlo2 + k : hiStart + (k - nLow);
                    double *ps = cur + 3 * ((xq + jb2) & S.mask);
                    ps[0] = SA_NEG_INF; ps[1] = SA_NEG_INF; ps[2] = SA_NEG_INF;
                }
            }
            xlo2 = xlo1; xhi2 = xhi1; xlo1 = xlo; xhi1 = xhi;
            __syncthreads();                    // f(d) is complete
            if (d != nextT) continue;
            // ------------------------------------------------------------------ traceback from diagonal d
            const bool atEnd = d == nDiag;
            const int32_t tracedBackFrom = curFrom;
            traceback(P, I, S, RW, t, NT, H, d, rowI, cLo, eLo, tracedBackTo, tracedBackFrom, atEnd, cOX, cEX, nPairs, acc, lik, yxNaN);
            tracedBackTo = tracedBackFrom;
            tbI++;
            nextT = tbI < nTb ? tbList[tbI] : nDiag;
            curFrom = nextT - (nextT == nDiag ? 0 : P.traceBack + 1);
            if (atEnd) break;
            // restore the forward sweep: f(d), f(d - 1) from the ring (all planes were kept above tracedBackFrom)
            __syncthreads();
            {
                const int4 h0 = RW.hdr[rowI], h1 = RW.hdr[rowI == 0 ? RW.cap - 1 : rowI - 1];
                const double none[3] = {SA_NEG_INF, SA_NEG_INF, SA_NEG_INF};
                loadState(S, t, NT, H, d, h0.x, h0.y, RW.ring + h0.z, none, d - 1, h1.x, h1.y, RW.ring + h1.z);
                xlo1 = h0.x; xhi1 = h0.y; xlo2 = h1.x; xhi2 = h1.y;
            }
            __syncthreads();
        }
        if (EX) {
            __syncthreads();
            double *red = S.foldv;              // (NW <= 32 warps x 8 doubles <= 2 NSW)
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
                    for (int w = 0; w < NW; w++) v[i] += red[w * 8 + i];
                }
                double *o = P.expOut + (int64_t) it * 10;   // from * 3 + to, states M, X, Y
                o[0] = v[0]; o[1] = v[1]; o[2] = v[5];
                o[3] = v[2]; o[4] = v[3]; o[5] = 0.0;
                o[6] = v[4]; o[7] = v[7] != 0.0 ? __longlong_as_double(0x7ff8000000000000LL) : 0.0; o[8] = v[6];
                o[9] = lik;
            }
        }
        // the pair count lives in the flushing warp
        if (t == NT - 1) {
            if (EX && HD) { P.pairCount[it] = 0; P.assignCount[it] = nPairs; }
            else P.pairCount[it] = nPairs;
        }
    }
};

// Persistent grid, items pulled heaviest-first from an atomic queue.  Launch: blockDim = NT, dynamic shared memory =
// WideK::smemBytes(NSW, P.wFold == nullptr).  One instantiation per window class (sa_gpu.cu: SA_W64_CLASSES).
template <bool EX, bool HD, bool GR, int NSW, int NT, int MINB>
__global__ void __launch_bounds__(NT, MINB) sa_wide_kernel(const __grid_constant__ FParams P) {
    extern __shared__ __align__(16) double sa_wsm[];
    using W = WideK<EX, HD, GR, NSW, NT>;
    typename W::Sm S;
    W::carve(S, sa_wsm, NSW, P.wFold ? P.wFold + (int64_t) blockIdx.x * 2 * NSW : nullptr,
             GR ? P.wRings + (int64_t) blockIdx.x * 6 * NSW : nullptr);
    sa_logAdd_init();
    typename W::Rows RW;
    RW.ring = P.scratch + (int64_t) blockIdx.x * P.wRingDoubles;
    RW.hdr = P.wHdr + (int64_t) blockIdx.x * P.wRowsCap;
    RW.cap = P.wRowsCap;
    RW.ringDoubles = (int32_t) P.wRingDoubles;
    for (;;) {
        __syncthreads();
        if (threadIdx.x == 0) S.sInt[0] = atomicAdd(P.counter, 1);
        __syncthreads();
        const int n = S.sInt[0];
        if (n >= P.nOrder) break;
        const int it = P.order[n];
        W::run(P, P.items[it], it, S, RW);
    }
}
