// sa_estimate.cuh — per-read scale / shift / var / drift estimation on the device.
//
// Reference: signalUtils_estimateNanoporeParams (impl/signalMachineUtils.c:186-225), which signalMachine runs before every
// alignment and every expectations pass (impl/signalMachine.c:778, 855):
//   nanopore_getOneDAssignmentsFromRead   impl/nanopore.c:601-631   kmer i of the base-called strand <-> event map[i], taken
//                                                                   when map[i] exceeds every event index taken so far
//   nanopore_compute_mean_scale_params    impl/nanopore.c:756-826   weighted least squares, 3 unknowns (shift, scale, drift)
//                                                                   + var = sqrt(mean weighted squared residual)
//   nanopore_compute_noise_scale_params   impl/nanopore.c:878-946   the same in 2 unknowns on the event sd
//   nanopore_lineq_solve                  impl/nanopore.c:691-748   Gaussian elimination (pivot quirk kept)
//   nanopore_adjustEventsForDriftP        impl/nanopore.c:633-638   mean -= time * drift for every event of the strand
//
// One CTA per strand.  The results are BIT-IDENTICAL to the reference's: every one of its sums is a sequential chain of
// double additions in list order, and floating-point addition is not associative, so the chains are kept — but only the
// chains.  Per chunk of NT rows the whole CTA evaluates, in parallel, the selection (a block-wide exclusive prefix maximum
// of the event map), the kmer index, the table look-ups and the 14 (pass 1) / 2 (pass 2) products of every row; rows that
// are not selected contribute an exact +0.0.  Then one thread per accumulator adds its NT products in row order.  The
// library is compiled with -fmad=false, divisions and square roots are IEEE (nvcc defaults), so every product is rounded
// as the reference's x86-64 build rounds it.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

struct EstStrand {
    int64_t readOff, mapOff, evOff;   // offsets into EstParams.reads / maps / (mean, sd, time)
    int32_t readLength, nEvents;
};

struct EstParams {
    const EstStrand *strands;
    const char *reads;
    const int32_t *maps;              // event index per base of the strand
    double *mean;                     // event means (adjusted in place when adjust != 0)
    const double *sd, *time;
    const double *table;              // K x 5 match table as loaded from the .model file
    int32_t code[256];                // character -> index in the sorted alphabet, -1 outside
    int32_t A, k;
    int32_t adjust;
    double *out;                      // 8 per strand: scale, shift, var, shift_sd, scale_sd, var_sd, drift, #assignments
    int32_t *err;                     // 1: character outside the alphabet, 2: no assignments / singular system, 3: map out of range
};

// nanopore_lineq_solve (impl/nanopore.c:691-748), n = 2 or 3; false when the reference would exit ("not invertible")
__device__ inline bool est_lineq(const double *A, const double *b, double *x, int n) {
    const double MACHEP = 1.11022302462515654042E-16;
    double aux[9];
    for (int i = 0; i < n; i++) {
        x[i] = b[i];
        for (int j = 0; j < n; j++) aux[i * n + j] = A[i * n + j];
    }
    for (int i = 0; i < n; i++) {
        if (fabs(aux[i * n + i]) < MACHEP) {
            int swap = i + 1;
            if (swap >= n) return false;
            while (aux[swap * n + i] < MACHEP) {
                swap++;
                if (swap >= n) return false;
            }
            for (int j = 0; j < n; j++) {
                const double t = aux[i * n + j];
                aux[i * n + j] = aux[swap * n + j];
                aux[swap * n + j] = t;
            }
            x[i] = x[swap];
            x[swap] = x[i];
        }
        double factor = __ddiv_rn(1.0, aux[i * n + i]);
        x[i] = __dmul_rn(x[i], factor);
        for (int j = 0; j < n; j++) aux[i * n + j] = __dmul_rn(aux[i * n + j], factor);
        for (int sub = i + 1; sub < n; sub++) {
            factor = aux[sub * n + i];
            x[sub] = __dsub_rn(x[sub], __dmul_rn(factor, x[i]));
            for (int j = 0; j < n; j++) aux[sub * n + j] = __dsub_rn(aux[sub * n + j], __dmul_rn(factor, aux[i * n + j]));
        }
    }
    for (int i = n - 1; i >= 0; i--)
        for (int sub = i - 1; sub >= 0; sub--) {
            const double factor = aux[sub * n + i];
            x[sub] = __dsub_rn(x[sub], __dmul_rn(factor, x[i]));
            for (int j = 0; j < n; j++) aux[sub * n + j] = __dsub_rn(aux[sub * n + j], __dmul_rn(factor, aux[i * n + j]));
        }
    return true;
}

template <int NT>
__global__ void __launch_bounds__(NT) sa_estimate_kernel(const __grid_constant__ EstParams P) {
    constexpr int NA = 14, NW = NT / 32;
    __shared__ double prod[NA][NT];
    __shared__ double sAcc[NA + 2];
    __shared__ double sBeta[5];
    __shared__ int32_t sWarpMax[NW];
    __shared__ int32_t sCarry, sCount, sBad;
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const EstStrand S = P.strands[blockIdx.x];
    const char *read = P.reads + S.readOff;
    const int32_t *map = P.maps + S.mapOff;
    double *mean = P.mean + S.evOff;
    const double *sd = P.sd + S.evOff, *tm = P.time + S.evOff;
    double *out = P.out + (int64_t) blockIdx.x * 8;
    const int32_t rows = S.readLength - (P.k - 1);
    if (t == 0) { sBad = 0; sCount = 0; }
    if (t < 5) sBeta[t] = 0.0;
    __syncthreads();
    for (int pass = 0; pass < 2; pass++) {
        const int na = pass == 0 ? NA : 2;
        double acc = 0.0;                       // accumulator of thread t < na
        if (t == 0) sCarry = -1;
        __syncthreads();
        for (int32_t c0 = 0; c0 < rows; c0 += NT) {
            const int32_t i = c0 + t;
            const bool valid = i < rows;
            const int32_t e = valid ? map[i] : -1;
            // exclusive prefix maximum of the event index in row order (-1 before the first row)
            int32_t run = e;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int32_t q = __shfl_up_sync(0xffffffffu, run, o);
                if (lane >= o) run = max(run, q);
            }
            if (lane == 31) sWarpMax[warp] = run;
            __syncthreads();
            int32_t pm = __shfl_up_sync(0xffffffffu, run, 1);
            if (lane == 0) pm = -1;
            pm = max(pm, sCarry);
            for (int w = 0; w < warp; w++) pm = max(pm, sWarpMax[w]);
            const bool sel = valid && e > pm;
            double v[NA];
#pragma unroll
            for (int a = 0; a < NA; a++) v[a] = 0.0;
            if (sel) {
                int64_t id = 0;
                bool bad = false;
                for (int q = 0; q < P.k; q++) {
                    const int32_t cc = P.code[(unsigned char) read[i + q]];
                    bad |= cc < 0;
                    id = id * P.A + (cc < 0 ? 0 : cc);
                }
                if (bad) atomicMax(&sBad, 1);
                if (e >= S.nEvents) { atomicMax(&sBad, 3); }
                else if (!bad) {
                    const double *row = P.table + id * 5;
                    const double level_mean = row[0], level_sd = row[1], noise_mean = row[2], noise_sd = row[3];
                    const double event = mean[e], noise = sd[e], time = tm[e];
                    if (pass == 0) {
                        const double inv_var = __ddiv_rn(1.0, __dmul_rn(level_sd, level_sd));
                        const double scaled_mean = __dmul_rn(level_mean, inv_var), scaled_time = __dmul_rn(time, inv_var);
                        v[0] = inv_var; v[1] = scaled_mean; v[2] = scaled_time;
                        v[3] = __dmul_rn(scaled_mean, level_mean); v[4] = __dmul_rn(scaled_mean, time);
                        v[5] = __dmul_rn(scaled_time, time);
                        v[6] = __dmul_rn(inv_var, event); v[7] = __dmul_rn(scaled_mean, event); v[8] = __dmul_rn(scaled_time, event);
                        const double inv_nv = __ddiv_rn(1.0, __dmul_rn(noise_sd, noise_sd));
                        const double scaled_nm = __dmul_rn(noise_mean, inv_nv);
                        v[9] = inv_nv; v[10] = scaled_nm; v[11] = __dmul_rn(scaled_nm, noise_mean);
                        v[12] = __dmul_rn(inv_nv, noise); v[13] = __dmul_rn(scaled_nm, noise);
                    } else {
                        const double predicted = __dadd_rn(__dadd_rn(sBeta[0], __dmul_rn(sBeta[1], level_mean)), __dmul_rn(sBeta[2], time));
                        const double residual = __dsub_rn(event, predicted);
                        v[0] = __ddiv_rn(__dmul_rn(residual, residual), __dmul_rn(level_sd, level_sd));
                        const double predicted_n = __dadd_rn(sBeta[3], __dmul_rn(sBeta[4], noise_mean));
                        const double residual_n = __dsub_rn(noise, predicted_n);
                        v[1] = __ddiv_rn(__dmul_rn(residual_n, residual_n), __dmul_rn(noise_sd, noise_sd));
                    }
                }
            }
#pragma unroll
            for (int a = 0; a < NA; a++)
                if (a < na) prod[a][t] = v[a];
            if (pass == 0) {
                const int nSel = __syncthreads_count(sel);      // (also the barrier that publishes prod)
                if (t == 0) sCount += nSel;
            } else __syncthreads();
            if (t < na) {
                const int lim = min(NT, rows - c0);
                for (int j = 0; j < lim; j++) acc = __dadd_rn(acc, prod[t][j]);     // the reference's chain, row order
            }
            if (t == NT - 1) sCarry = max(sCarry, max(pm, e));   // (thread NT-1's inclusive maximum covers the chunk)
            __syncthreads();
        }
        if (t < na) sAcc[t] = acc;
        __syncthreads();
        if (t == 0) {
            const int32_t n = sCount;
            if (pass == 0) {
                bool ok = n > 0 && sBad == 0;
                if (ok) {
                    const double XWX[9] = {sAcc[0], sAcc[1], sAcc[2], sAcc[1], sAcc[3], sAcc[4], sAcc[2], sAcc[4], sAcc[5]};
                    const double XWy[3] = {sAcc[6], sAcc[7], sAcc[8]};
                    const double NWX[4] = {sAcc[9], sAcc[10], sAcc[10], sAcc[11]};
                    const double NWy[2] = {sAcc[12], sAcc[13]};
                    double b3[3], b2[2];
                    ok = est_lineq(XWX, XWy, b3, 3) && est_lineq(NWX, NWy, b2, 2);
                    sBeta[0] = b3[0]; sBeta[1] = b3[1]; sBeta[2] = b3[2]; sBeta[3] = b2[0]; sBeta[4] = b2[1];
                }
                if (!ok && sBad == 0) sBad = 2;
            } else {
                out[0] = sBeta[1]; out[1] = sBeta[0]; out[6] = sBeta[2];           // scale, shift, drift
                out[2] = sqrt(__ddiv_rn(sAcc[0], (double) n));                     // var
                out[3] = sBeta[3]; out[4] = sBeta[4];                              // shift_sd, scale_sd
                out[5] = sqrt(__ddiv_rn(sAcc[1], (double) n));                     // var_sd
                out[7] = (double) n;
                if (sBad) atomicMax(P.err, sBad);
            }
        }
        __syncthreads();
        if (sBad) break;
    }
    if (sBad) {
        if (t == 0) atomicMax(P.err, sBad);
        return;
    }
    if (P.adjust) {
        const double drift = sBeta[2];
        for (int32_t i = t; i < S.nEvents; i += NT) mean[i] = __dsub_rn(mean[i], __dmul_rn(tm[i], drift));
    }
}
