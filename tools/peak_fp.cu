// peak_fp.cu — measured issue peaks of the FP64 / FP32 / integer pipes the band kernels live on (B200, sm_100a).
//
// MEASURED_PEAKS.json only carries HBM GB/s and bf16 TFLOP/s; this path is bound by the FP64 pipe (bit-exact parity
// with the reference's double arithmetic), so the roofline denominator is measured here: lane-instructions per second
// of back-to-back independent DFMA / DADD / DMUL / FFMA / IMAD chains with every SM full.
//
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/peak_fp tools/peak_fp.cu && tools/peak_fp
#include <cstdio>
#include <cuda_runtime.h>

template <int OP>
__global__ void __launch_bounds__(256) pipe_kernel(double *out, int iters, double seed) {
    constexpr int N = 8;  // independent chains per thread
    double a[N];
    float f[N];
    int q[N];
#pragma unroll
    for (int i = 0; i < N; i++) {
        a[i] = seed + threadIdx.x * 1e-9 + i;
        f[i] = (float) a[i];
        q[i] = threadIdx.x + i;
    }
    const double m = 1.0000001, c = 1e-9;
    const float mf = 1.0000001f, cf = 1e-9f;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int r = 0; r < 8; r++) {
#pragma unroll
            for (int i = 0; i < N; i++) {
                if (OP == 0) a[i] = fma(a[i], m, c);
                if (OP == 1) a[i] = __dadd_rn(a[i], c);
                if (OP == 2) a[i] = __dmul_rn(a[i], m);
                if (OP == 3) f[i] = fmaf(f[i], mf, cf);
                if (OP == 4) q[i] = q[i] * 3 + it;
            }
        }
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < N; i++) s += a[i] + f[i] + q[i];
    if (s == 123.456) out[0] = s;
}

template <int OP>
static double run(const char *name, int nSM) {
    double *d;
    cudaMalloc(&d, 8);
    const int iters = 4096, blocks = nSM * 8;
    pipe_kernel<OP><<<blocks, 256>>>(d, 64, 1.0);
    cudaDeviceSynchronize();
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    float best = 1e30f;
    for (int rep = 0; rep < 5; rep++) {
        cudaEventRecord(e0);
        pipe_kernel<OP><<<blocks, 256>>>(d, iters, 1.0);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    double ops = (double) blocks * 256 * iters * 8 * 8;
    double rate = ops / (best * 1e-3);
    printf("{\"pipe\": \"%s\", \"lane_instr_per_s\": %.4e, \"per_sm_per_clk_at_1965MHz\": %.2f, \"ms\": %.3f}\n", name, rate,
           rate / nSM / 1.965e9, best);
    cudaFree(d);
    return rate;
}

int main() {
    int nSM = 148;
    cudaDeviceGetAttribute(&nSM, cudaDevAttrMultiProcessorCount, 0);
    printf("{\"sms\": %d}\n", nSM);
    run<0>("dfma", nSM);
    run<1>("dadd", nSM);
    run<2>("dmul", nSM);
    run<3>("ffma", nSM);
    run<4>("imad", nSM);
    return 0;
}
