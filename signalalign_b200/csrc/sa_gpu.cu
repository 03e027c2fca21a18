// sa_gpu.cu — device side of libsagpu: model upload, batch packing, kernel launches, result assembly.
//
// A batch = many getAlignedPairsUsingAnchors / getExpectationsUsingAnchors calls (one per read) executed together.
// Host work here is integer-only preparation that the reference also does on the host (split points, kmer
// indices / ambiguity paths, band geometry); all floating-point DP work runs in the kernels.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <numeric>
#include <vector>
#include <cuda_runtime.h>
#include "sa_internal.h"
#include "sa_general.cuh"
#include "sa_fast.cuh"
#include "sa_cta.cuh"
#include "sa_wide.cuh"
#include "sa_wide32.cuh"
#include "sa_estimate.cuh"
#include "sa_variants.cuh"
#include "sa_eventalign.cuh"

#define CK(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess) {                                                                   \
            sa_set_error("CUDA error %s at %s:%d: %s", #call, __FILE__, __LINE__, cudaGetErrorString(e_)); \
            return SA_ERR_CUDA;                                                                    \
        }                                                                                          \
    } while (0)

// ------------------------------------------------------------------------------------------------ device model
struct DeviceModel {
    double *kt = nullptr, *ft = nullptr, *gapX = nullptr, *grid = nullptr, *pp = nullptr, *slopes = nullptr;
    int32_t *kmerToRow = nullptr;
    double *raw = nullptr;     // K x 5 match table as loaded (per-read parameter estimation, sa_estimate.cuh); uploaded on first use
};

void sa_model_drop_device(const sa_model *m) {
    if (!m->dev) return;
    cudaFree(m->dev->kt); cudaFree(m->dev->ft); cudaFree(m->dev->gapX); cudaFree(m->dev->grid); cudaFree(m->dev->pp);
    cudaFree(m->dev->slopes); cudaFree(m->dev->kmerToRow); cudaFree(m->dev->raw);
    delete m->dev;
    m->dev = nullptr;
    m->devDevice = -1;
}

static int uploadModel(const sa_model *m) {
    int dev;
    CK(cudaGetDevice(&dev));
    if (m->dev && m->devDevice == dev) return SA_OK;
    sa_model_drop_device(m);
    DeviceModel *d = new DeviceModel;
    std::vector<double> kt((size_t) m->K * SA_KT_STRIDE);
    for (int64_t i = 0; i < m->K; i++)
        for (int t = 0; t < 2; t++) {
            const double *src = (t == 0 ? m->match.data() : m->gapY.data()) + i * 5;
            double *dst = kt.data() + i * SA_KT_STRIDE + t * 6;
            dst[0] = src[0];
            dst[1] = src[1];
            dst[2] = log(src[1]);  // l_sigma of emissions_signal_logGaussPdf, evaluated by the host libm like the reference
            dst[3] = src[2];
            dst[4] = src[4];
            dst[5] = log(src[4]);
        }
    CK(cudaMalloc(&d->kt, kt.size() * sizeof(double)));
    CK(cudaMemcpy(d->kt, kt.data(), kt.size() * sizeof(double), cudaMemcpyHostToDevice));
    // fast-kernel table: mu, 1/sd, -0.5*log(2*pi) - log(sd) for the match and gap-Y tables (sd == 0 -> emission -inf)
    std::vector<double> ft((size_t) m->K * 5);
    for (int64_t i = 0; i < m->K; i++) {
        double sdM = m->match[i * 5 + 1], sdY = m->gapY[i * 5 + 1];
        ft[i * 5 + 0] = m->match[i * 5];
        ft[i * 5 + 1] = sdM == 0.0 ? 0.0 : 1.0 / sdM;
        ft[i * 5 + 2] = sdM == 0.0 ? -INFINITY : -0.91893853320467267 - log(sdM);
        ft[i * 5 + 3] = sdY == 0.0 ? 0.0 : 1.0 / sdY;
        ft[i * 5 + 4] = sdY == 0.0 ? -INFINITY : -0.91893853320467267 - log(sdY);
    }
    CK(cudaMalloc(&d->ft, ft.size() * sizeof(double)));
    CK(cudaMemcpy(d->ft, ft.data(), ft.size() * sizeof(double), cudaMemcpyHostToDevice));
    CK(cudaMalloc(&d->gapX, m->K * sizeof(double)));
    CK(cudaMemcpy(d->gapX, m->gapX.data(), m->K * sizeof(double), cudaMemcpyHostToDevice));
    if (m->G > 0) {
        CK(cudaMalloc(&d->grid, m->G * sizeof(double)));
        CK(cudaMemcpy(d->grid, m->grid.data(), m->G * sizeof(double), cudaMemcpyHostToDevice));
        CK(cudaMalloc(&d->pp, m->postPred.size() * sizeof(double)));
        CK(cudaMemcpy(d->pp, m->postPred.data(), m->postPred.size() * sizeof(double), cudaMemcpyHostToDevice));
        CK(cudaMalloc(&d->slopes, m->slopes.size() * sizeof(double)));
        CK(cudaMemcpy(d->slopes, m->slopes.data(), m->slopes.size() * sizeof(double), cudaMemcpyHostToDevice));
        CK(cudaMalloc(&d->kmerToRow, m->K * sizeof(int32_t)));
        CK(cudaMemcpy(d->kmerToRow, m->kmerToRow.data(), m->K * sizeof(int32_t), cudaMemcpyHostToDevice));
    }
    m->dev = d;
    m->devDevice = dev;
    return SA_OK;
}

// ------------------------------------------------------------------------------------------------ device init
extern "C" int sa_gpu_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
    return n;
}

extern "C" int sa_gpu_init(int device) {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
        sa_set_error("sa_gpu_init: no CUDA device available (%s); libsagpu has no CPU fallback",
                     e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
        return SA_ERR_CUDA;
    }
    if (device < 0 || device >= n) {
        sa_set_error("sa_gpu_init: device %d out of range (have %d)", device, n);
        return SA_ERR_CUDA;
    }
    CK(cudaSetDevice(device));
    CK(cudaFree(0));
    return SA_OK;
}

// ------------------------------------------------------------------------------------------------ batch
#include <chrono>
struct PhaseTimer {  // SA_TIMING=1 prints host-side phase times to stderr
    bool on;
    const char *what;
    std::chrono::steady_clock::time_point t0;
    explicit PhaseTimer(const char *w) : on(getenv("SA_TIMING") != nullptr), what(w), t0(std::chrono::steady_clock::now()) {}
    void lap(const char *name) {
        if (!on) return;
        auto t1 = std::chrono::steady_clock::now();
        fprintf(stderr, "[sa timing] %s / %s: %.1f ms\n", what, name, std::chrono::duration<double, std::milli>(t1 - t0).count());
        t0 = t1;
    }
};

// page-locked staging memory, cached for the life of the process (pinning is far slower than the copies it serves)
#include <mutex>
struct PinnedSlot { void *p = nullptr; size_t cap = 0; bool busy = false; };
static std::mutex g_pinMu;
static std::vector<PinnedSlot> g_pins;
static void *pinAcquire(size_t bytes) {
    std::lock_guard<std::mutex> lk(g_pinMu);
    int best = -1;
    size_t total = 0;
    for (int i = 0; i < (int) g_pins.size(); i++) {
        total += g_pins[i].cap;
        if (g_pins[i].busy || !g_pins[i].p) continue;
        if (g_pins[i].cap >= bytes && (best < 0 || g_pins[i].cap < g_pins[best].cap)) best = i;
    }
    if (best < 0) {
        // nothing idle fits: add a slab (never re-size an idle one — it is somebody's steady-state size and replacing it
        // makes every call pay cudaHostAlloc again); idle slabs are only given back once the pool passes 32 GB
        size_t cap = bytes + bytes / 4 + 4096;
        if (total + cap > ((size_t) 32 << 30))
            for (auto &sl : g_pins)
                if (!sl.busy && sl.p) { cudaFreeHost(sl.p); sl.p = nullptr; sl.cap = 0; }
        int slot = -1;
        for (int i = 0; i < (int) g_pins.size(); i++)
            if (!g_pins[i].p) { slot = i; break; }
        if (slot < 0) { g_pins.push_back(PinnedSlot()); slot = (int) g_pins.size() - 1; }
        PinnedSlot &sl = g_pins[slot];
        if (cudaHostAlloc(&sl.p, cap, cudaHostAllocDefault) != cudaSuccess) { sl.p = nullptr; sl.cap = 0; return nullptr; }
        sl.cap = cap;
        sl.busy = false;
        best = slot;
    }
    g_pins[best].busy = true;
    return g_pins[best].p;
}
static bool pinRelease(void *p) {
    if (!p) return false;
    std::lock_guard<std::mutex> lk(g_pinMu);
    for (auto &sl : g_pins)
        if (sl.p == p) { sl.busy = false; return true; }
    return false;
}
// sa_result_free (sa_host.cpp): the pairs array of a result is a slab of this pool, D2H'ed into directly
bool sa_pinned_release(void *p) { return pinRelease(p); }

// Teardown for a long-lived host process (signalMachine itself just exits): waits for the device, gives the idle
// page-locked slabs and the cached device pool memory back.  Results still held by the caller keep their slabs; models
// and batches are freed by their own calls.  The library stays usable afterwards (caches refill on the next call).
extern "C" int sa_gpu_shutdown(void) {
    if (cudaDeviceSynchronize() != cudaSuccess) {
        sa_set_error("sa_gpu_shutdown: %s", cudaGetErrorString(cudaGetLastError()));
        return SA_ERR_CUDA;
    }
    {
        std::lock_guard<std::mutex> lk(g_pinMu);
        for (auto &sl : g_pins)
            if (!sl.busy && sl.p) { cudaFreeHost(sl.p); sl.p = nullptr; sl.cap = 0; }
    }
    int dev = 0;
    cudaMemPool_t pool;
    if (cudaGetDevice(&dev) == cudaSuccess && cudaDeviceGetDefaultMemPool(&pool, dev) == cudaSuccess) cudaMemPoolTrimTo(pool, 0);
    return SA_OK;
}
template <class T>
struct PinVec {  // the few vector operations the packer needs, over pinned memory
    T *p = nullptr;
    size_t n = 0;
    bool resize(size_t count) {
        pinRelease(p);
        p = nullptr;
        n = count;
        if (count == 0) return true;
        p = (T *) pinAcquire(count * sizeof(T));
        return p != nullptr;
    }
    T *data() { return p; }
    const T *data() const { return p; }
    size_t size() const { return n; }
    bool empty() const { return n == 0; }
    T &operator[](size_t i) { return p[i]; }
    const T &operator[](size_t i) const { return p[i]; }
    ~PinVec() { pinRelease(p); }
};

// Device buffers come from the device's default memory pool in stream order (cudaMallocAsync / cudaFreeAsync) with the
// pool's release threshold lifted, so that a batch neither pays cudaMalloc for its multi-GB slots every call nor stalls
// the other batches in flight with the device-wide synchronisation of cudaFree.
static int poolInit() {
    static bool done[64] = {false};
    int dev = 0;
    CK(cudaGetDevice(&dev));
    if (dev < 64 && done[dev]) return SA_OK;
    cudaMemPool_t pool;
    CK(cudaDeviceGetDefaultMemPool(&pool, dev));
    uint64_t keep = UINT64_MAX;
    CK(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep));
    if (dev < 64) done[dev] = true;
    return SA_OK;
}

template <class T>
struct DevBuf {
    T *p = nullptr;
    size_t n = 0;
    cudaStream_t s = nullptr;   // the stream the buffer was allocated on (and is freed on)
    int alloc(size_t count, cudaStream_t st) {
        release();
        n = count;
        s = st;
        if (count == 0) return SA_OK;
        cudaError_t e = cudaMallocAsync((void **) &p, count * sizeof(T), st);
        if (e == cudaErrorMemoryAllocation) {
            // the pool keeps what earlier batches freed (release threshold lifted); blocks of other sizes fragment it.  Let the
            // pending frees land, hand the unused blocks back to the driver and ask once more.
            (void) cudaGetLastError();
            int dev = 0;
            cudaMemPool_t pool;
            if (cudaDeviceSynchronize() == cudaSuccess && cudaGetDevice(&dev) == cudaSuccess &&
                cudaDeviceGetDefaultMemPool(&pool, dev) == cudaSuccess)
                cudaMemPoolTrimTo(pool, 0);
            e = cudaMallocAsync((void **) &p, count * sizeof(T), st);
        }
        if (e != cudaSuccess) {
            p = nullptr;
            sa_set_error("CUDA error cudaMallocAsync(%zu bytes) at %s:%d: %s", count * sizeof(T), __FILE__, __LINE__, cudaGetErrorString(e));
            (void) cudaGetLastError();
            return SA_ERR_CUDA;
        }
        return SA_OK;
    }
    int upload(const std::vector<T> &v, cudaStream_t s, int64_t *bytes) {
        int rc = alloc(v.size(), s);
        if (rc) return rc;
        if (!v.empty()) {
            CK(cudaMemcpyAsync(p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice, s));
            if (bytes) *bytes += (int64_t) (v.size() * sizeof(T));
        }
        return SA_OK;
    }
    int upload(const PinVec<T> &v, cudaStream_t s, int64_t *bytes) {
        int rc = alloc(v.size(), s);
        if (rc) return rc;
        if (!v.empty()) {
            CK(cudaMemcpyAsync(p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice, s));
            if (bytes) *bytes += (int64_t) (v.size() * sizeof(T));
        }
        return SA_OK;
    }
    void release() {
        if (p) cudaFreeAsync(p, s);
        p = nullptr;
        n = 0;
    }
    ~DevBuf() { release(); }
};

// wide band kernel classes: window slots, threads per CTA (sa_wide.cuh)
static const int kWideClasses = 5;
static const int kWideNSW[kWideClasses] = {256, 512, 1024, 2048, 4096};
static const size_t kSmemCap = 227 * 1024 - 1024;   // opt-in shared memory per CTA minus the kernels' static part (logAdd table)

// one work item = one sub-region of one read after getSplitPoints
struct ItemRec {
    int64_t read, x1, y1, lX, lY;
    int32_t region;
    int32_t flags;            // bit0 ragged left, bit1 ragged right
    int64_t evOff, pathOffOff, pathKmerOff, anchorOff, nAnchors;
    int64_t maxP;
    int64_t pairOff = 0, pairCap = 0, assignOff = 0, assignCap = 0;
    int64_t cells = 0;        // band cells (integer band logic, same as the reference's)
    int64_t work = 0;         // path-cells
    int kernel = 0;           // 0 general; 1, 2, 4, 8 = warps of the register-resident band kernel; 100 + c = wide class c
    double scale, shift, var;
};

struct sa_batch {
    const sa_model *model = nullptr;
    sa_params_t params;
    bool expectations = false;
    int64_t nReads = 0;
    struct StreamOwner {         // declared before every DevBuf: destroyed after their stream-ordered frees were queued
        cudaStream_t s = nullptr;
        ~StreamOwner() { if (s) cudaStreamDestroy(s); }
    } streamOwner, auxOwner, latOwner, wideOwner[5];
    cudaEvent_t evWideJoin[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
    cudaStream_t stream = nullptr;  // the band kernels
    // everything else (H2D, geometry pass, compaction, D2H) on a HIGH-priority stream: when several batches are in
    // flight, these small kernels must overtake the not yet resident CTAs of another batch's persistent band kernel
    cudaStream_t aux = nullptr;
    cudaEvent_t evPrepared = nullptr;
    // latency class: the few reads whose own sweep is long against the whole batch run on the one-cell-per-thread
    // mapping (<4, 1>: half the instructions per thread and diagonal) next to the throughput kernel, on their own stream
    cudaStream_t lat = nullptr;
    cudaEvent_t evFork = nullptr, evJoin = nullptr, evJoinSplit = nullptr;
    std::vector<int32_t> latOrder;
    int latBlocks = 0;
    int32_t latRows = 1;
    // split mode (the longest of those, while their full-length row buffers fit a memory budget): the forward sweep
    // runs alone on the one-cell mapping keeping every row; the traceback blocks then are independent work items
    std::vector<int32_t> splitOrder;
    std::vector<FBlock> blocks;          // block b of the list; blockItemFirst[i] .. : blocks of split item i in order
    std::vector<int64_t> blockCountHost;
    int blockGrid = 0;
    cudaEvent_t ev[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    std::vector<ItemRec> items;  // global item id = index
    // shared inputs
    DevBuf<double> dEvMean, dEvNoise, dEvLogNoise;
    DevBuf<int32_t> dPathOff, dPathKmer, dPcCol, dAnchors;
    int mpThreads = 0;         // > 0: the alignment pass of the general items runs on sa_mp_kernel
    size_t mpSmem = 0;
    // outputs (indexed by global item id)
    DevBuf<sa_pair_t> dPairs;
    DevBuf<int64_t> dPairCount, dAssignCount;
    DevBuf<double> dExpOut;
    DevBuf<int32_t> dAssignK, dAssignE;
    // general kernel
    std::vector<GItem> gItems;      // indexed by global item id (entries of fast items unused)
    std::vector<int32_t> gOrder;
    DevBuf<GItem> dGItems;
    DevBuf<int32_t> dGOrder, dBandL, dBandW, dCounters;
    DevBuf<int64_t> dBandOff;
    DevBuf<double> dGScratch;
    GParams gp;
    int gCta = 0;
    // fast kernel
    std::vector<FItem> fItems;      // indexed by global item id
    std::vector<int32_t> fOrder[4];   // CTA kernel classes W = 1, 2, 4, 8
    DevBuf<FItem> dFItems;
    DevBuf<int32_t> dFOrder[4], dTb;
    DevBuf<FGeo> dGeo;
    DevBuf<double> dFScratch, dLatScratch, dSplitRows;
    DevBuf<int32_t> dLatOrder, dSplitOrder;
    DevBuf<FBlock> dBlocks;
    DevBuf<int64_t> dBlockCount;
    DevBuf<sa_pair_t> dSplitPairs;
    FParams fp;
    int fBlocks[4] = {0, 0, 0, 0};
    int32_t fRows[4] = {1, 1, 1, 1};
    // wide kernel (sa_wide.cuh), classes by window size
    std::vector<int32_t> wOrder[kWideClasses];
    DevBuf<int32_t> dWOrder[kWideClasses];
    DevBuf<double> dWRing[kWideClasses], dWFold[kWideClasses], dWRings[kWideClasses], dWOff[kWideClasses];
    bool fp32 = false;                  // the wide classes run on sa_wide32_kernel (Gaussian alignment, SA_PRECISION_FP32)
    int32_t wOffCap[kWideClasses] = {0, 0, 0, 0, 0};
    DevBuf<int4> dWHdr[kWideClasses];
    int wBlocks[kWideClasses] = {0, 0, 0, 0, 0};
    int32_t wRows[kWideClasses] = {1, 1, 1, 1, 1};
    int64_t wRingDoubles[kWideClasses] = {0, 0, 0, 0, 0};
    bool oneCell = false;      // SA_CTA_C1: the 128-slot class with one cell per thread (A/B runs only)
    int64_t bandCells = 0, pathCells = 0, events = 0;
    int64_t h2dBytes = 0, d2hBytes = 0;
    float lastMs[3] = {0, 0, 0};
    int64_t lastLaunches = 0;
    std::vector<int64_t> pairCountHost, assignCountHost;
    ~sa_batch() {
        if (stream) cudaStreamSynchronize(stream);   // the buffers are freed in the order of the aux stream
        for (auto &e : ev)
            if (e) cudaEventDestroy(e);
        if (evPrepared) cudaEventDestroy(evPrepared);
        if (evFork) cudaEventDestroy(evFork);
        if (evJoin) cudaEventDestroy(evJoin);
        if (evJoinSplit) cudaEventDestroy(evJoinSplit);
        for (auto &e : evWideJoin)
            if (e) cudaEventDestroy(e);
    }
};

static int64_t nextPow2(int64_t v) {
    int64_t p = 1;
    while (p < v) p <<= 1;
    return p;
}

// Models the band kernel handles: the production Gaussian emission, or the HDP emission for alignment runs
static bool modelFastEligible(const sa_model *m, bool expectations) {
    const bool hdp = m->kind == SA_EMIT_HDP;
    if (m->kind != SA_EMIT_GAUSS_MEANONLY_DESCALE && !hdp) return false;
    if (hdp && m->G < 2) return false;
    if (!(m->trans[T_GAP_SWITCH_TO_X] == -INFINITY)) return false;
    // the HDP machine's X-gap emission is the literal log(0.1) (stateMachine.c:1394), whatever the table says
    for (int64_t i = 0; i < m->K; i++) {
        if (!hdp && m->gapX[i] != m->gapX[0]) return false;
        if (!hdp && m->gapY[i * 5] != m->match[i * 5]) return false;  // the fast kernel shares the de-scaled mean
    }
    return true;
}

// CTA band kernel classes by slot count NS = 32, 64, 128, 256 (an item goes to the narrowest class its band fits):
// <warps, cells per thread> = <1,1>, <1,2>, <2,2>, <4,2>
template <bool EX, bool HD>
static int ctaOccupancyT(int c, bool oneCell, int *bps) {
    if (c == 0) CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(bps, sa_cta_kernel<1, 1, EX, HD>, 32, 0));
    if (c == 1) CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(bps, sa_cta_kernel<1, 2, EX, HD>, 32, 0));
    if (c == 2 && !oneCell) CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(bps, sa_cta_kernel<2, 2, EX, HD>, 64, 0));
    if (c == 2 && oneCell) CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(bps, sa_cta_kernel<4, 1, false, false>, 128, 0));
    if (c == 3) CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(bps, sa_cta_kernel<4, 2, EX, HD>, 128, 0));
    return SA_OK;
}
template <bool EX, bool HD>
static void ctaLaunchT(int c, bool oneCell, int blocks, cudaStream_t st, const FParams &f) {
    if (c == 0) sa_cta_kernel<1, 1, EX, HD><<<blocks, 32, 0, st>>>(f);
    else if (c == 1) sa_cta_kernel<1, 2, EX, HD><<<blocks, 32, 0, st>>>(f);
    else if (c == 2 && !oneCell) sa_cta_kernel<2, 2, EX, HD><<<blocks, 64, 0, st>>>(f);
    else if (c == 2) sa_cta_kernel<4, 1, false, false><<<blocks, 128, 0, st>>>(f);
    else sa_cta_kernel<4, 2, EX, HD><<<blocks, 128, 0, st>>>(f);
}
// ex: the expectations variant of the kernels (getExpectationsUsingAnchors) instead of the aligned-pairs one;
// hdp: the HDP emission instead of the Gaussian tables (with ex: the kmer / event assignment lists as well)
static int ctaOccupancy(int c, bool oneCell, bool ex, bool hdp, int *bps) {
    if (hdp && ex) return ctaOccupancyT<true, true>(c, false, bps);
    if (hdp) return ctaOccupancyT<false, true>(c, false, bps);
    return ex ? ctaOccupancyT<true, false>(c, false, bps) : ctaOccupancyT<false, false>(c, oneCell, bps);
}
static void ctaLaunch(int c, bool oneCell, bool ex, bool hdp, int blocks, cudaStream_t st, const FParams &f) {
    if (hdp && ex) ctaLaunchT<true, true>(c, false, blocks, st, f);
    else if (hdp) ctaLaunchT<false, true>(c, false, blocks, st, f);
    else if (ex) ctaLaunchT<true, false>(c, false, blocks, st, f);
    else ctaLaunchT<false, false>(c, oneCell, blocks, st, f);
}

// wide kernel: dynamic shared memory of a window class, occupancy and launch by (expectations, HDP) variant
// (gr: emission rings and fold terms in global memory — the 4096-slot class, whose DP state alone fills shared memory)
static size_t wideSmemBytes(bool gr, int NSW, bool foldInSmem) {
    return gr ? WideK<false, false, true>::smemBytes(NSW, false) : WideK<false, false, false>::smemBytes(NSW, foldInSmem);
}
// One instantiation of the FP64 wide kernel per window class and (expectations, HDP) variant: <rings in global memory,
// window, threads, CTAs per SM the registers are held to>; window and thread count are compile-time constants there.
#define SA_W64_CLASSES(X) X(0, false, 256, 128, 5) X(1, false, 512, 256, 3) X(2, false, 1024, 512, 2) X(3, true, 2048, 512, 2) X(4, true, 4096, 1024, 1)
template <bool EX, bool HD>
static int wideOccupancyT(int c, size_t smem, int *bps) {
    static bool raised[kWideClasses] = {false, false, false, false, false};
#define X(C, GR, NSW, NT, MB)                                                                                                      \
    if (c == C) {                                                                                                                 \
        if (smem > 48 * 1024 && !raised[C]) {                                                                                     \
            CK(cudaFuncSetAttribute(sa_wide_kernel<EX, HD, GR, NSW, NT, MB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) kSmemCap)); \
            raised[C] = true;                                                                                                     \
        }                                                                                                                         \
        CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(bps, sa_wide_kernel<EX, HD, GR, NSW, NT, MB>, NT, smem));                \
    }
    SA_W64_CLASSES(X)
#undef X
    return SA_OK;
}
static int wideOccupancy(int c, bool hdp, bool ex, size_t smem, int *bps) {
    if (hdp && ex) return wideOccupancyT<true, true>(c, smem, bps);
    if (hdp) return wideOccupancyT<false, true>(c, smem, bps);
    return ex ? wideOccupancyT<true, false>(c, smem, bps) : wideOccupancyT<false, false>(c, smem, bps);
}
template <bool EX, bool HD>
static void wideLaunchT(int c, int blocks, size_t smem, cudaStream_t st, const FParams &f) {
#define X(C, GR, NSW, NT, MB) if (c == C) sa_wide_kernel<EX, HD, GR, NSW, NT, MB><<<blocks, NT, smem, st>>>(f);
    SA_W64_CLASSES(X)
#undef X
}
static void wideLaunch(int c, bool hdp, bool ex, int blocks, size_t smem, cudaStream_t st, const FParams &f) {
    if (hdp && ex) wideLaunchT<true, true>(c, blocks, smem, st, f);
    else if (hdp) wideLaunchT<false, true>(c, blocks, smem, st, f);
    else if (ex) wideLaunchT<true, false>(c, blocks, smem, st, f);
    else wideLaunchT<false, false>(c, blocks, smem, st, f);
}
// ---- precision of the wide band kernel for Gaussian alignment runs: FP64 (default, bit-faithful) or FP32 state relative to
// per-diagonal offsets (sa_wide32.cuh; within the contract's 1e-4).  SA_PRECISION=fp32 in the environment sets the default.
static int g_precision = -1;
static int currentPrecision() {
    if (g_precision < 0) {
        const char *e = getenv("SA_PRECISION");
        g_precision = (e && (!strcmp(e, "fp32") || !strcmp(e, "FP32") || !strcmp(e, "f32"))) ? SA_PRECISION_FP32 : SA_PRECISION_FP64;
    }
    return g_precision;
}
extern "C" int sa_gpu_set_precision(int precision) {
    if (precision != SA_PRECISION_FP64 && precision != SA_PRECISION_FP32) {
        sa_set_error("sa_gpu_set_precision: unknown precision %d", precision);
        return SA_ERR_INPUT;
    }
    g_precision = precision;
    return SA_OK;
}
extern "C" int sa_gpu_get_precision(void) { return currentPrecision(); }

// does the class keep its emission rings in global memory?  From the 2048-slot window on: with the rings in shared memory
// one CTA of that class fills an SM (1024 threads, half of them idle on a typical diagonal); with only the DP state
// there, two CTAs of 512 threads fit and overlap each other's barriers (+15 % on the contract pattern).
static bool wideGlobalRings(int c) {
#define X(C, GR, NSW, NT, MB) if (c == C) return GR;
    SA_W64_CLASSES(X)
#undef X
    return false;
}
// fold terms in global memory: with the rings there, and from the 1024-slot window on (two CTAs of 512 threads per SM
// instead of one: 25.0 -> 26.8 GCUPS on the contract pattern).  SA_WIDE_FOLDG_FROM=<slots> moves the limit (experiments).
static bool wideFoldGlobal(int c, bool gr) {
    if (gr) return true;
    int from = 1024;
    if (const char *e = getenv("SA_WIDE_FOLDG_FROM")) from = atoi(e);
    if (kWideNSW[c] >= from) return true;
    return wideSmemBytes(false, kWideNSW[c], true) > kSmemCap;
}

// FP32 wide kernel: one instantiation per window class — <rings in global memory, window, max threads, CTAs per SM the
// registers are held to>; the window is a compile-time constant there (immediate masks and shared-memory offsets)
#define SA_W32_CLASSES(X) X(0, false, 256, 128, 6) X(1, false, 512, 256, 3) X(2, false, 1024, 256, 3) X(3, true, 2048, 256, 3) X(4, true, 4096, 512, 1)
static bool wide32GlobalRings(int c) { return c >= 3; }
static int wide32Threads(int c) {                     // a template argument of the kernel, not a launch-time choice
#define X(C, GR, NSW, NT, MB) if (c == C) return NT;
    SA_W32_CLASSES(X)
#undef X
    return 0;
}
static size_t wide32SmemBytes(int c, bool foldInSmem) {
#define X(C, GR, NSW, NT, MB) if (c == C) return WideF<GR, NSW, NT>::smemBytes(NSW, foldInSmem);
    SA_W32_CLASSES(X)
#undef X
    return 0;
}
static int wide32Occupancy(int c, int threads, size_t smem, int *bps) {
    static bool raised[kWideClasses] = {false, false, false, false, false};
#define X(C, GR, NSW, NT, MB)                                                                                                   \
    if (c == C) {                                                                                                              \
        if (smem > 48 * 1024 && !raised[C]) {                                                                                  \
            CK(cudaFuncSetAttribute(sa_wide32_kernel<GR, NSW, NT, MB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) kSmemCap)); \
            raised[C] = true;                                                                                                  \
        }                                                                                                                      \
        CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(bps, sa_wide32_kernel<GR, NSW, NT, MB>, threads, smem));             \
    }
    SA_W32_CLASSES(X)
#undef X
    return SA_OK;
}
static void wide32Launch(int c, int blocks, int threads, size_t smem, cudaStream_t st, const FParams &f) {
#define X(C, GR, NSW, NT, MB) if (c == C) sa_wide32_kernel<GR, NSW, NT, MB><<<blocks, threads, smem, st>>>(f);
    SA_W32_CLASSES(X)
#undef X
}

// widest three-diagonal span the register-resident kernel takes; wider items go to the wide kernel.  126 = the 128-slot
// class <2,2>: the 256-slot class <4,2> (four warps behind one mailbox barrier) runs at half its rate — 14.4 GCUPS on the
// long-runs / trim-14 pattern against 22.5 for the wide kernel's 256-slot window on the same reads.
// (SA_REG_MAX_SPAN=<columns> moves the limit; experiments)
static int regMaxSpan() {
    if (const char *e = getenv("SA_REG_MAX_SPAN")) return atoi(e);
    return 126;
}

// readBase: index of reads[0] in the caller's array (error messages name the caller's read)
// chunkCount: how many sub-batches of this size the caller runs back to back (scales the latency-class decision)
static int prepareBatch(const sa_model_t *m, const sa_read_t *reads, int64_t nReads, const sa_params_t *p,
                        const sa_ambig_t *ambigIn, int expectations, int64_t readBase, int64_t chunkCount,
                        sa_batch_t **out) {
    *out = nullptr;
    if (sa_gpu_device_count() == 0) {
        sa_set_error("sa_batch_prepare: no CUDA device available; libsagpu has no CPU fallback");
        return SA_ERR_CUDA;
    }
    if (p->diagonalExpansion % 2 != 0 || p->diagonalExpansion < 0 || p->traceBackDiagonals < 1 ||
        p->minDiagsBetweenTraceBack < 2 || p->traceBackDiagonals + 1 >= p->minDiagsBetweenTraceBack) {
        sa_set_error("sa_batch_prepare: parameters violate the prerequisites of getPosteriorProbsWithBanding "
                     "(pairwiseAligner.c:1460-1464)");
        return SA_ERR_INPUT;
    }
    if (m->kind == SA_EMIT_HDP && m->G == 0) {
        sa_set_error("sa_batch_prepare: HDP emission requested but the model has no .nhdp tables");
        return SA_ERR_INPUT;
    }
    int rc = uploadModel(m);
    if (rc) return rc;
    sa_ambig defAmbig;
    const sa_ambig *ambig = ambigIn;
    if (!ambig) {
        sa_ambig_default_table(defAmbig);
        ambig = &defAmbig;
    }
    std::unique_ptr<sa_batch> b(new sa_batch);
    b->model = m;
    b->params = *p;
    b->expectations = expectations != 0;
    b->nReads = nReads;
    if ((rc = poolInit())) return rc;
    int prLeast = 0, prGreatest = 0;
    CK(cudaDeviceGetStreamPriorityRange(&prLeast, &prGreatest));
    CK(cudaStreamCreateWithPriority(&b->streamOwner.s, cudaStreamNonBlocking, prLeast));
    CK(cudaStreamCreateWithPriority(&b->auxOwner.s, cudaStreamNonBlocking, prGreatest));
    CK(cudaStreamCreateWithPriority(&b->latOwner.s, cudaStreamNonBlocking, prGreatest));
    b->stream = b->streamOwner.s;
    b->aux = b->auxOwner.s;
    b->lat = b->latOwner.s;
    CK(cudaEventCreateWithFlags(&b->evFork, cudaEventDisableTiming));
    CK(cudaEventCreateWithFlags(&b->evJoin, cudaEventDisableTiming));
    CK(cudaEventCreateWithFlags(&b->evJoinSplit, cudaEventDisableTiming));
    for (auto &e : b->ev) CK(cudaEventCreate(&e));
    CK(cudaEventCreateWithFlags(&b->evPrepared, cudaEventDisableTiming));
    cudaStream_t st = b->aux;

    const bool needNoise = m->kind == SA_EMIT_GAUSS_NOISE_SCALED || m->kind == SA_EMIT_GAUSS_NOISE_DESCALE;
    const bool fastModel = modelFastEligible(m, b->expectations) && getenv("SA_FORCE_GENERAL") == nullptr;
    PinVec<int32_t> pathOff, pathKmer, pcCol, anchors32;
    PinVec<double> evMean, evNoise, evLogNoise;
    std::vector<ItemRec> &items = b->items;

    // alphabet / ambiguity classification per character for the rolling kmer index
    int charCode[256];
    bool charAmbig[256];
    for (int c = 0; c < 256; c++) {
        size_t j = m->alphabet.find((char) c);
        charCode[c] = (c != 0 && j != std::string::npos) ? (int) j : -1;
        charAmbig[c] = ambig->opts[c][0] != 0;
    }

    PhaseTimer timer("prepare");
    // ------------------------------------------------------------------ pass 1: split, events, kmer paths, anchors
    // phase A (parallel over reads): validate anchors, split into regions, count paths -> per-read item lists
    std::vector<std::vector<ItemRec>> perRead((size_t) nReads);
    std::string firstError;
    const int64_t suffixK = m->K / m->A;
#pragma omp parallel
    {
        std::vector<int64_t> splitL;
#pragma omp for schedule(dynamic, 16)
        for (int64_t r = 0; r < nReads; r++) {
            const sa_read_t &R = reads[r];
            int64_t lXt = R.targetLength > 0 ? R.targetLength - (m->k - 1) : 0;  // sequence_correctSeqLength (:523-538)
            if (lXt < 0) lXt = 0;
            int64_t lYt = R.nEvents;
            char err[512];
            err[0] = 0;
            if (lXt > 0x3fffffff || lYt > 0x3fffffff)
                snprintf(err, sizeof(err), "read %lld: sequences longer than 2^30 are not supported", (long long) (r + readBase));
            for (int64_t i = 0; i < R.nAnchors && !err[0]; i++) {
                int64_t ax = R.anchors[2 * i], ay = R.anchors[2 * i + 1];
                bool ok = ax >= 0 && ay >= 0 && ax < lXt && ay < lYt;
                if (ok && i > 0) ok = ax > R.anchors[2 * i - 2] && ay > R.anchors[2 * i - 1];
                if (!ok)
                    snprintf(err, sizeof(err), "read %lld: anchor %lld (%lld, %lld) out of range or not strictly "
                             "increasing (asserted by band_construct, pairwiseAligner.c:226-231)",
                             (long long) (r + readBase), (long long) i, (long long) ax, (long long) ay);
            }
            // every character of the target must be in the alphabet or an ambiguity code whose options are
            for (int64_t i = 0; i < R.targetLength && lXt > 0 && !err[0]; i++) {
                unsigned char ch = (unsigned char) R.target[i];
                bool ok = charCode[ch] >= 0 && !charAmbig[ch];
                if (!ok && charAmbig[ch]) {
                    ok = true;
                    for (const char *o = ambig->opts[ch]; *o; o++) ok = ok && charCode[(unsigned char) *o] >= 0;
                }
                if (!ok)
                    snprintf(err, sizeof(err), "read %lld: kmer at reference position %lld contains a character "
                             "outside the alphabet %s (the reference exits in kmer_to_word, nanopore_hdp.c:393-398)",
                             (long long) (r + readBase), (long long) std::max<int64_t>(0, i - (m->k - 1)), m->alphabet.c_str());
            }
            if (err[0]) {
#pragma omp critical
                if (firstError.empty()) firstError = err;
                continue;
            }
            sa_split_points(R.anchors, R.nAnchors, lXt, lYt, p->splitMatrixBiggerThanThis, R.raggedLeft != 0,
                            R.raggedRight != 0, splitL);
            int64_t nSplit = (int64_t) splitL.size() / 4, j = 0;
            for (int64_t sI = 0; sI < nSplit; sI++) {
                int64_t x1 = splitL[4 * sI], y1 = splitL[4 * sI + 1], x2 = splitL[4 * sI + 2], y2 = splitL[4 * sI + 3];
                ItemRec it;
                it.read = r; it.x1 = x1; it.y1 = y1; it.lX = x2 - x1; it.lY = y2 - y1; it.region = (int32_t) sI;
                bool rl = R.raggedLeft || sI > 0, rr = R.raggedRight || sI < nSplit - 1;  // :2001-2002
                it.flags = (rl ? 1 : 0) | (rr ? 2 : 0);
                it.scale = R.scale; it.shift = R.shift; it.var = R.var;
                it.anchorOff = j;   // for now: first anchor of the region inside the read's list
                while (j < R.nAnchors && R.anchors[2 * j] + R.anchors[2 * j + 1] < x2 + y2) j++;  // :1984-1997
                it.nAnchors = j - it.anchorOff;
                // number of paths of the region: 1 (NULL kmer column) + prod of option counts per column
                int64_t nPaths = 1, maxP = 1;
                int nAmbInWin = 0;
                const char *tg = R.target + x1;
                for (int q = 0; q < m->k - 1 && q < R.targetLength - x1; q++) nAmbInWin += charAmbig[(unsigned char) tg[q]];
                for (int64_t x = 1; x <= it.lX; x++) {
                    nAmbInWin += charAmbig[(unsigned char) tg[x - 1 + m->k - 1]];
                    int64_t pc = 1;
                    if (nAmbInWin)
                        for (int q = 0; q < m->k; q++) {
                            unsigned char ch = (unsigned char) tg[x - 1 + q];
                            if (charAmbig[ch]) pc *= (int64_t) strlen(ambig->opts[ch]);
                        }
                    nPaths += pc;
                    maxP = std::max(maxP, pc);
                    nAmbInWin -= charAmbig[(unsigned char) tg[x - 1]];
                }
                it.maxP = maxP;
                it.work = nPaths;  // temporarily: number of paths (re-used as the pathKmer size below)
                perRead[r].push_back(it);
            }
        }
    }
    if (!firstError.empty()) {
        sa_set_error("%s", firstError.c_str());
        return SA_ERR_INPUT;
    }
    // offsets
    int64_t evTot = 0, poTot = 0, pkTot = 0, anTot = 0;
    for (int64_t r = 0; r < nReads; r++)
        for (auto &it : perRead[r]) {
            int64_t inRead = it.anchorOff;
            it.evOff = evTot; evTot += it.lY;
            it.pathOffOff = poTot; poTot += it.lX + 2;
            it.pathKmerOff = pkTot; pkTot += it.work;
            it.anchorOff = anTot; anTot += it.nAnchors;
            it.work = inRead;  // remember where the region's anchors start in the read's list for phase B
            b->events += it.lY;
            items.push_back(it);
        }
    if (pkTot > 0x7fffffff || poTot > 0x7fffffff) {
        sa_set_error("sa_batch_prepare: batch too large (more than 2^31 kmer paths); split it");
        return SA_ERR_INPUT;
    }
    bool pinOk = evMean.resize(evTot) && pathOff.resize(poTot) && pathKmer.resize(pkTot) && pcCol.resize(pkTot) && anchors32.resize(anTot * 2);
    if (needNoise) pinOk = pinOk && evNoise.resize(evTot) && evLogNoise.resize(evTot);
    if (!pinOk) {
        sa_set_error("sa_batch_prepare: cannot allocate pinned staging memory");
        return SA_ERR_CUDA;
    }
    // phase B (parallel over items): fill the packed arrays
#pragma omp parallel
    {
        std::vector<int32_t> pathsL;
#pragma omp for schedule(dynamic, 8)
        for (int64_t i = 0; i < (int64_t) items.size(); i++) {
            ItemRec &it = items[i];
            const sa_read_t &R = reads[it.read];
            const int64_t a0 = it.work;
            it.work = 0;
            for (int64_t q = 0; q < it.nAnchors; q++) {
                anchors32[2 * (it.anchorOff + q)] = (int32_t) (R.anchors[2 * (a0 + q)] - it.x1);
                anchors32[2 * (it.anchorOff + q) + 1] = (int32_t) (R.anchors[2 * (a0 + q) + 1] - it.y1);
            }
            const double *e = R.events + it.y1 * 4;
            double *em = evMean.data() + it.evOff;
            for (int64_t y = 0; y < it.lY; y++) em[y] = e[4 * y];
            if (needNoise)
                for (int64_t y = 0; y < it.lY; y++) {
                    double noise = e[4 * y + 1];
                    evNoise[it.evOff + y] = noise;
                    if (m->kind == SA_EMIT_GAUSS_NOISE_DESCALE && noise == 0) noise = 0.000000001;
                    evLogNoise[it.evOff + y] = log(noise);
                }
            // kmer paths per matrix column (column 0 = NULL kmer)
            int32_t *po = pathOff.data() + it.pathOffOff;
            int32_t *pk = pathKmer.data() + it.pathKmerOff;
            int32_t *pcc = pcCol.data() + it.pathKmerOff;      // column of every path-cell (sa_mp_kernel)
            int64_t w = 0;
            po[0] = 0;
            pcc[w] = 0;
            pk[w++] = -1;
            po[1] = 1;
            int64_t roll = -1;
            const char *tg = R.target + it.x1;
            int nAmbInWin = 0;
            for (int q = 0; q < m->k - 1 && q < R.targetLength - it.x1; q++) nAmbInWin += charAmbig[(unsigned char) tg[q]];
            for (int64_t x = 1; x <= it.lX; x++) {
                const char *km = tg + x - 1;
                nAmbInWin += charAmbig[(unsigned char) km[m->k - 1]];
                if (nAmbInWin == 0) {
                    if (roll >= 0) roll = (roll % suffixK) * m->A + charCode[(unsigned char) km[m->k - 1]];
                    else {
                        roll = 0;
                        for (int q = 0; q < m->k; q++) roll = roll * m->A + charCode[(unsigned char) km[q]];
                    }
                    pcc[w] = (int32_t) x;
                    pk[w++] = (int32_t) roll;
                } else {
                    roll = -1;
                    sa_kmer_paths(km, *m, *ambig, pathsL);  // characters were validated in phase A
                    for (int32_t v : pathsL) { pcc[w] = (int32_t) x; pk[w++] = v; }
                }
                po[x + 1] = (int32_t) w;
                nAmbInWin -= charAmbig[(unsigned char) km[0]];
            }
        }
    }
    timer.lap("pass1 pack (split, events, kmers, anchors)");
    const int64_t n = (int64_t) items.size();
    if (n > 0x7fffffff) {
        sa_set_error("sa_batch_prepare: too many work items");
        return SA_ERR_INPUT;
    }
    if ((rc = b->dEvMean.upload(evMean, st, &b->h2dBytes))) return rc;
    if ((rc = b->dEvNoise.upload(evNoise, st, &b->h2dBytes))) return rc;
    if ((rc = b->dEvLogNoise.upload(evLogNoise, st, &b->h2dBytes))) return rc;
    if ((rc = b->dPathOff.upload(pathOff, st, &b->h2dBytes))) return rc;
    if ((rc = b->dPathKmer.upload(pathKmer, st, &b->h2dBytes))) return rc;

    if ((rc = b->dAnchors.upload(anchors32, st, &b->h2dBytes))) return rc;
    if ((rc = b->dCounters.alloc(16, st))) return rc;

    // output slots
    int64_t pairTotalCap = 0, assignTotalCap = 0;
    for (auto &it : items) {
        it.pairOff = pairTotalCap;
        it.pairCap = b->expectations ? 0 : 3 * it.lY * it.maxP + 64;
        pairTotalCap += it.pairCap;
        it.assignOff = assignTotalCap;
        it.assignCap = (b->expectations && m->kind == SA_EMIT_HDP) ? 4 * it.lY * it.maxP + 64 : 0;
        assignTotalCap += it.assignCap;
    }
    if ((rc = b->dPairCount.alloc(std::max<int64_t>(1, n), st))) return rc;
    if ((rc = b->dPairs.alloc(std::max<int64_t>(1, pairTotalCap), st))) return rc;
    if (b->expectations) {
        if ((rc = b->dExpOut.alloc(std::max<int64_t>(1, n * 10), st))) return rc;
        if ((rc = b->dAssignCount.alloc(std::max<int64_t>(1, n), st))) return rc;
        if ((rc = b->dAssignK.alloc(std::max<int64_t>(1, assignTotalCap), st))) return rc;
        if ((rc = b->dAssignE.alloc(std::max<int64_t>(1, assignTotalCap), st))) return rc;
    }

    int dev, nSM = 148;
    CK(cudaGetDevice(&dev));
    CK(cudaDeviceGetAttribute(&nSM, cudaDevAttrMultiProcessorCount, dev));
    CK(cudaStreamSynchronize(st));
    timer.lap("uploads + output allocation");

    // ------------------------------------------------------------------ pass 2: fast-kernel candidates + geometry
    FParams &f = b->fp;
    memset(&f, 0, sizeof(f));
    b->fItems.resize(n);
    std::vector<FGeo> geo(n);
    if (fastModel && n > 0) {
        int64_t tbTotal = 0;
        for (int64_t i = 0; i < n; i++) {
            ItemRec &it = items[i];
            FItem &fi = b->fItems[i];
            memset(&fi, 0, sizeof(fi));
            fi.evOff = it.evOff;
            fi.kmerOff = it.pathKmerOff + 1;   // column x reads pathKmer[.. + x] when every column has one path
            fi.anchorOff = it.anchorOff;
            fi.nAnchors = (int32_t) it.nAnchors;
            fi.tbOff = tbTotal;
            // consecutive traceback points are >= minDiags - (traceBack + 1) diagonals apart
            fi.nTbCap = (int32_t) ((it.lX + it.lY) / (p->minDiagsBetweenTraceBack - p->traceBackDiagonals - 1) + 2);
            tbTotal += fi.nTbCap;
            fi.pairOff = it.pairOff; fi.pairCap = it.pairCap;
            fi.assignOff = it.assignOff; fi.assignCap = it.assignCap;
            fi.lX = (int32_t) it.lX; fi.lY = (int32_t) it.lY;
            fi.flags = it.flags;
            fi.scale = it.scale; fi.shift = it.shift;
            fi.invVar = 1 / it.var;
            fi.logInvVar = log((1 / it.var));
            fi.var = it.var;
        }
        if ((rc = b->dFItems.upload(b->fItems, st, &b->h2dBytes))) return rc;
        if ((rc = b->dTb.alloc(std::max<int64_t>(1, tbTotal), st))) return rc;
        if ((rc = b->dGeo.alloc(n, st))) return rc;
        b->fp32 = currentPrecision() == SA_PRECISION_FP32 && !b->expectations && m->kind == SA_EMIT_GAUSS_MEANONLY_DESCALE;
        f.items = b->dFItems.p; f.nItems = (int32_t) n;
        f.evMean = b->dEvMean.p; f.kmerIdx = b->dPathKmer.p; f.anchors = b->dAnchors.p;
        f.tb = b->dTb.p; f.geo = b->dGeo.p; f.ft = m->dev->ft;
        memcpy(f.trans, m->trans, sizeof(f.trans));
        f.eX = m->gapX[0];
        f.threshold = p->threshold;
        f.logThrGate = p->threshold > 0 ? log(p->threshold) - 1e-6 : -INFINITY;
        f.minDiags = (int32_t) p->minDiagsBetweenTraceBack; f.traceBack = (int32_t) p->traceBackDiagonals;
        f.expansion = (int32_t) p->diagonalExpansion;
        f.pairs = b->dPairs.p; f.pairCount = b->dPairCount.p;
        f.expOut = b->expectations ? b->dExpOut.p : nullptr;
        if (b->expectations) { f.assignK = b->dAssignK.p; f.assignE = b->dAssignE.p; f.assignCount = b->dAssignCount.p; }
        if (m->kind == SA_EMIT_HDP) {
            f.grid = m->dev->grid; f.pp = m->dev->pp; f.slopes = m->dev->slopes; f.kmerToRow = m->dev->kmerToRow;
            f.G = (int32_t) m->G;
            f.gridDx = m->grid[1] - m->grid[0];
            f.gridInvDx = 1.0 / f.gridDx;
            f.eX = -2.3025850929940455;     // stateMachine3HDP_cellCalculate's literal X-gap emission
        }
        f.wPlanes = m->kind == SA_EMIT_HDP ? 4 : 3;
        f.wAllFull = b->expectations ? 1 : 0;
        f.wGeoAlways = (getenv("SA_WIDE_ONLY") != nullptr || regMaxSpan() < 254) ? 1 : 0;
        sa_geometry_kernel<<<(unsigned) n, 32, 0, st>>>(f);
        CK(cudaGetLastError());
        CK(cudaMemcpyAsync(geo.data(), b->dGeo.p, sizeof(FGeo) * n, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        timer.lap("geometry kernel + D2H");
    }
    static const int kW[4] = {1, 2, 4, 8};
    const int64_t planes = m->kind == SA_EMIT_HDP ? 4 : 3;   // forward row = header + M, X, Y (+ the HDP log density)
    // Which band kernel: the register-resident one while three consecutive diagonals fit its widest window (254 columns)
    // and the band never moves backwards; else the wide kernel, in the narrowest window class that holds them plus the
    // slack of its emission rings (SA_WIDE_ONLY: every item, A/B runs); the general kernel only beyond 4096 slots.
    const bool wideOnly = getenv("SA_WIDE_ONLY") != nullptr, noWide = getenv("SA_NO_WIDE") != nullptr;
    const int regSpan = regMaxSpan();
    int wideMax = kWideClasses;
    while (wideMax > 0 && wideSmemBytes(true, kWideNSW[wideMax - 1], false) > kSmemCap) wideMax--;
    for (int64_t i = 0; i < n; i++) {
        ItemRec &it = items[i];
        it.kernel = 0;
        if (fastModel && it.maxP == 1 && geo[i].nTb <= b->fItems[i].nTbCap) {
            if (geo[i].monotone && !wideOnly)
                for (int c = 0; c < 4 && !it.kernel; c++)
                    if (geo[i].maxSpan <= 32 * kW[c] - 2 && (noWide || geo[i].maxSpan <= regSpan)) it.kernel = kW[c];
            if (!it.kernel && !noWide)
                for (int c = 0; c < wideMax && !it.kernel; c++)
                    if (geo[i].maxSpan + 24 <= kWideNSW[c]) it.kernel = 100 + c;
        }
        if (it.kernel) {
            it.cells = geo[i].cells;
            it.work = geo[i].cells;
            b->bandCells += it.cells;
            b->pathCells += it.cells;
            if (it.kernel >= 100) {
                const int c = it.kernel - 100;
                b->wOrder[c].push_back((int32_t) i);
                b->wRows[c] = std::max(b->wRows[c], geo[i].maxRows);
                b->wOffCap[c] = std::max<int32_t>(b->wOffCap[c], (int32_t) ((it.lX + it.lY) / 16 + 2));
                b->wRingDoubles[c] = std::max(b->wRingDoubles[c], geo[i].maxBlockDoubles + 2 * 256 * (int64_t) planes * geo[i].maxWidth);   // + the table being written, + the tail a wrap skips
                continue;
            }
            int c = it.kernel == 1 ? 0 : it.kernel == 2 ? 1 : it.kernel == 4 ? 2 : 3;
            b->fOrder[c].push_back((int32_t) i);
            b->fRows[c] = std::max(b->fRows[c], geo[i].maxRows);
        }
    }
    b->oneCell = getenv("SA_CTA_C1") != nullptr;
    auto byWork = [&](int32_t a, int32_t c) { return items[a].work > items[c].work; };
    for (int c = 0; c < 4; c++) std::stable_sort(b->fOrder[c].begin(), b->fOrder[c].end(), byWork);
    for (int c = 0; c < kWideClasses; c++) std::stable_sort(b->wOrder[c].begin(), b->wOrder[c].end(), byWork);
    // Latency class.  A read is one CTA's sequential sweep (~2.3 us per diagonal on the two-cell mapping under load);
    // when that alone exceeds half of what the whole batch needs at full throughput (~30 GCUPS), the read decides the
    // makespan (heavy-tailed lengths, BASELINE config 5; batches smaller than the resident grid).  Those reads, at most
    // two per SM and longest first, go to the one-cell mapping, whose step is about half as long.
    if (!b->oneCell && !b->expectations && getenv("SA_NO_LATENCY_CLASS") == nullptr) {
        const double ideal = (double) b->bandCells * (double) chunkCount / 30e9;
        const char *ef = getenv("SA_LAT_FACTOR");
        const double latFactor = ef ? atof(ef) : 0.5;
        std::vector<int32_t> keep;
        for (int32_t i : b->fOrder[2]) {
            const double solo = (double) (items[i].lX + items[i].lY) * 2.3e-6;
            if ((int64_t) b->latOrder.size() < 2 * (int64_t) nSM && solo > latFactor * ideal) {
                b->latOrder.push_back(i);
                b->latRows = std::max(b->latRows, geo[i].maxRows);
            } else keep.push_back(i);
        }
        b->fOrder[2].swap(keep);
        // split mode for the longest of them while the row buffers fit (3088 B per diagonal; SA_SPLIT_GB, default 24)
        if (!b->latOrder.empty() && getenv("SA_NO_SPLIT") == nullptr) {
            const char *eg = getenv("SA_SPLIT_GB");
            const int64_t budget = (int64_t) ((eg ? atof(eg) : 24.0) * 1073741824.0 / 8.0);   // doubles
            std::vector<int32_t> tbHost((size_t) std::max<int64_t>(1, b->dTb.n));
            CK(cudaMemcpyAsync(tbHost.data(), b->dTb.p, sizeof(int32_t) * b->dTb.n, cudaMemcpyDeviceToHost, st));
            CK(cudaStreamSynchronize(st));
            int64_t fullTotal = 0, pairTotal = 0;
            std::vector<int32_t> ring;
            b->latRows = 1;
            for (int32_t i : b->latOrder) {
                const int64_t nDiag = items[i].lX + items[i].lY, need = (nDiag + 1) * (planes * 128 + 2);
                if (fullTotal + need > budget) {
                    ring.push_back(i);
                    b->latRows = std::max(b->latRows, geo[i].maxRows);
                    continue;
                }
                b->splitOrder.push_back(i);
                b->fItems[i].fullOff = fullTotal;
                fullTotal += need;
                int32_t to = 0;
                for (int32_t k = 0; k < geo[i].nTb; k++) {
                    FBlock B;
                    B.item = i;
                    B.d = tbHost[b->fItems[i].tbOff + k];
                    B.tracedBackTo = to;
                    B.tracedBackFrom = B.d - (B.d == nDiag ? 0 : (int32_t) p->traceBackDiagonals + 1);
                    B.pairOff = pairTotal;
                    B.pairCap = 4 * (int64_t) (B.tracedBackFrom - to) + 256;
                    if (const char *ec = getenv("SA_SPLIT_PAIRCAP")) B.pairCap = atoll(ec);   // tests: force the re-run path
                    pairTotal += B.pairCap;
                    to = B.tracedBackFrom;
                    b->blocks.push_back(B);
                }
            }
            b->latOrder.swap(ring);
            if (!b->splitOrder.empty()) {
                if ((rc = b->dFItems.upload(b->fItems, st, &b->h2dBytes))) return rc;   // (fullOff)
                f.items = b->dFItems.p;
                if ((rc = b->dSplitOrder.upload(b->splitOrder, st, &b->h2dBytes))) return rc;
                if ((rc = b->dBlocks.upload(b->blocks, st, &b->h2dBytes))) return rc;
                if ((rc = b->dBlockCount.alloc(b->blocks.size(), st))) return rc;
                if ((rc = b->dSplitRows.alloc((size_t) fullTotal, st))) return rc;
                if ((rc = b->dSplitPairs.alloc((size_t) std::max<int64_t>(1, pairTotal), st))) return rc;
                int bps = 1;
                if (m->kind == SA_EMIT_HDP) CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, sa_cta_block_kernel<2, 2, true>, 64, 0));
                else CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, sa_cta_block_kernel<2, 2, false>, 64, 0));
                b->blockGrid = (int) std::min<int64_t>((int64_t) nSM * std::max(1, bps), (int64_t) b->blocks.size());
            }
        }
        if (!b->latOrder.empty()) {
            if ((rc = b->dLatOrder.upload(b->latOrder, st, &b->h2dBytes))) return rc;
            b->latBlocks = (int) b->latOrder.size();
            if ((rc = b->dLatScratch.alloc((size_t) b->latBlocks * b->latRows * (planes * 128 + 2), st))) return rc;
        }
    }
    {
        int64_t need = 1;
        for (int c = 0; c < 4; c++) {
            if ((rc = b->dFOrder[c].upload(b->fOrder[c], st, &b->h2dBytes))) return rc;
            if (b->fOrder[c].empty()) continue;
            int bps = 1;
            int64_t rowD = 0, nIt = (int64_t) b->fOrder[c].size();
            if ((rc = ctaOccupancy(c, b->oneCell, b->expectations, m->kind == SA_EMIT_HDP, &bps))) return rc;
            b->fBlocks[c] = (int) std::min<int64_t>((int64_t) nSM * std::max(1, bps), nIt);
            rowD = (int64_t) b->fBlocks[c] * b->fRows[c] * (planes * 32 * kW[c] + 2);
            need = std::max(need, rowD);
        }
        if ((rc = b->dFScratch.alloc((size_t) need, st))) return rc;
        f.scratch = b->dFScratch.p;
    }
    // wide kernel: per class a persistent grid, per CTA a ring of compact forward rows + a ring of row headers
    for (int c = 0; c < kWideClasses; c++) {
        if ((rc = b->dWOrder[c].upload(b->wOrder[c], st, &b->h2dBytes))) return rc;
        if (b->wOrder[c].empty()) continue;
        const bool hdp = m->kind == SA_EMIT_HDP;
        const bool gr = b->fp32 ? wide32GlobalRings(c) : wideGlobalRings(c);
        const bool foldGlobal = b->fp32 ? (gr || kWideNSW[c] >= 1024) : wideFoldGlobal(c, gr);   // (fp32, 1024 slots: three CTAs per SM instead of two)
        const size_t smem = b->fp32 ? wide32SmemBytes(c, !foldGlobal) : wideSmemBytes(gr, kWideNSW[c], !foldGlobal);
        int bps = 1;
        if (b->fp32) {
            if ((rc = wide32Occupancy(c, wide32Threads(c), smem, &bps))) return rc;
        } else if ((rc = wideOccupancy(c, hdp, b->expectations, smem, &bps))) return rc;
        b->wBlocks[c] = (int) std::min<int64_t>((int64_t) nSM * std::max(1, bps), (int64_t) b->wOrder[c].size());
        b->wRows[c] += 4 + 256;      // + one descriptor table of headers written ahead
        b->wRingDoubles[c] = (b->wRingDoubles[c] + 64 + 15) & ~(int64_t) 15;
        if (b->wRingDoubles[c] > 0x7fffffffLL) {
            sa_set_error("sa_batch_prepare: a traceback block needs %.1f GB of forward rows (band too wide)", b->wRingDoubles[c] * 8e-9);
            return SA_ERR_INPUT;
        }
        // (the FP32 kernel's ring holds floats: the same element count in half the bytes; wRingDoubles is a multiple of 16)
        const size_t ringStride = (size_t) (b->fp32 ? b->wRingDoubles[c] / 2 : b->wRingDoubles[c]);
        {   // a class's rings take at most a fifth of the device and 40 % of what is available now (the one-shot call keeps
            // several chunks in flight, each with up to five classes): fewer resident CTAs rather than a failed call
            size_t freeB = 0, totalB = 0;
            CK(cudaMemGetInfo(&freeB, &totalB));
            cudaMemPool_t pool;
            uint64_t reserved = 0, used = 0;
            if (cudaDeviceGetDefaultMemPool(&pool, dev) == cudaSuccess &&
                cudaMemPoolGetAttribute(pool, cudaMemPoolAttrReservedMemCurrent, &reserved) == cudaSuccess &&
                cudaMemPoolGetAttribute(pool, cudaMemPoolAttrUsedMemCurrent, &used) == cudaSuccess && reserved > used)
                freeB += (size_t) (reserved - used);
            const double budget = std::min((double) totalB * 0.2, (double) freeB * 0.4);
            const int64_t fit = (int64_t) (budget / ((double) ringStride * 8.0));
            if (fit < b->wBlocks[c]) b->wBlocks[c] = (int) std::max<int64_t>(1, fit);
        }
        if ((rc = b->dWRing[c].alloc((size_t) b->wBlocks[c] * ringStride, st))) return rc;
        if ((rc = b->dWHdr[c].alloc((size_t) b->wBlocks[c] * b->wRows[c], st))) return rc;
        if (foldGlobal && (rc = b->dWFold[c].alloc((size_t) b->wBlocks[c] * 2 * kWideNSW[c], st))) return rc;
        // a stream of its own per window class: the classes are separate persistent grids, and on one stream each would
        // wait for the previous one's LAST item (a tail of one long read per SM) before its first CTA starts
        if (!b->wideOwner[c].s) {
            int prLo = 0, prHi = 0;
            CK(cudaDeviceGetStreamPriorityRange(&prLo, &prHi));
            CK(cudaStreamCreateWithPriority(&b->wideOwner[c].s, cudaStreamNonBlocking, prLo));
            CK(cudaEventCreateWithFlags(&b->evWideJoin[c], cudaEventDisableTiming));
        }
        if (gr && (rc = b->dWRings[c].alloc((size_t) b->wBlocks[c] * 6 * kWideNSW[c], st))) return rc;
        if (b->fp32 && (rc = b->dWOff[c].alloc((size_t) b->wBlocks[c] * b->wOffCap[c], st))) return rc;
    }

    timer.lap("geometry pass + fast-kernel scratch");
    // ------------------------------------------------------------------ pass 3: general kernel for the rest
    std::vector<int32_t> bandL, bandW;
    std::vector<int64_t> bandOff;
    std::vector<Diag> band;
    int64_t ringNeedMax = 1, maxDiagPC = 1, maxWidth = 1;
    b->gItems.resize(n);
    std::vector<int64_t> sub;
    for (int64_t i = 0; i < n; i++) {
        ItemRec &it = items[i];
        if (it.kernel) continue;
        GItem &g = b->gItems[i];
        memset(&g, 0, sizeof(g));
        g.evOff = it.evOff; g.pathOffOff = it.pathOffOff; g.pathKmerOff = it.pathKmerOff;
        g.pairOff = it.pairOff; g.pairCap = it.pairCap; g.assignOff = it.assignOff; g.assignCap = it.assignCap;
        g.lX = (int32_t) it.lX; g.lY = (int32_t) it.lY; g.flags = it.flags;
        g.scale = it.scale; g.shift = it.shift; g.var = it.var;
        g.invVar = 1 / it.var;
        g.logInvVar = log((1 / it.var));
        g.diagOff = (int64_t) bandL.size();
        sub.resize(it.nAnchors * 2);
        for (int64_t q = 0; q < it.nAnchors * 2; q++) sub[q] = anchors32[it.anchorOff * 2 + q];
        sa_build_band(sub.data(), it.nAnchors, it.lX, it.lY, p->diagonalExpansion, band);
        const int32_t *po = pathOff.data() + it.pathOffOff;
        int64_t nDiag = it.lX + it.lY, off = 0, tracedBackTo = 0, ringNeed = 1;
        for (int64_t d = 0; d <= nDiag; d++) {
            int64_t w = (band[d].xmyR - band[d].xmyL) / 2 + 1;
            int64_t x0 = (d + band[d].xmyL) / 2;
            int64_t pc = po[x0 + w] - po[x0];
            bandL.push_back((int32_t) band[d].xmyL);
            bandW.push_back((int32_t) w);
            bandOff.push_back(off);
            off += pc;
            it.cells += w;
            it.work += pc;
            maxDiagPC = std::max(maxDiagPC, pc);
            maxWidth = std::max(maxWidth, w);
            if (d >= 1) {
                bool atEnd = d == nDiag;
                bool tb = d >= tracedBackTo + p->minDiagsBetweenTraceBack && w <= p->diagonalExpansion * 2 + 1;
                if (atEnd || tb) {
                    ringNeed = std::max(ringNeed, off - bandOff[g.diagOff + tracedBackTo]);
                    tracedBackTo = d - (atEnd ? 0 : p->traceBackDiagonals + 1);
                }
            }
        }
        b->bandCells += it.cells;
        b->pathCells += it.work;
        ringNeedMax = std::max(ringNeedMax, ringNeed);
        b->gOrder.push_back((int32_t) i);
    }
    std::stable_sort(b->gOrder.begin(), b->gOrder.end(), byWork);
    GParams &g = b->gp;
    memset(&g, 0, sizeof(g));
    if (!b->gOrder.empty()) {
        if ((rc = b->dGItems.upload(b->gItems, st, &b->h2dBytes))) return rc;
        if ((rc = b->dGOrder.upload(b->gOrder, st, &b->h2dBytes))) return rc;
        if ((rc = b->dBandL.upload(bandL, st, &b->h2dBytes))) return rc;
        if ((rc = b->dBandW.upload(bandW, st, &b->h2dBytes))) return rc;
        if ((rc = b->dBandOff.upload(bandOff, st, &b->h2dBytes))) return rc;
        int64_t ring = nextPow2(ringNeedMax);
        const bool mpKernel = !b->expectations && getenv("SA_NO_MP") == nullptr;
        int64_t slab = ring * 3 + 9 * maxDiagPC + 2 * maxWidth + (mpKernel ? ring * 2 : 0);   // (+ the emission ring of sa_mp_kernel)
        size_t freeB = 0, totalB = 0;
        CK(cudaMemGetInfo(&freeB, &totalB));
        {   // memory the stream-ordered pool holds but does not use (the release threshold is lifted) is available too
            cudaMemPool_t pool;
            uint64_t reserved = 0, used = 0;
            if (cudaDeviceGetDefaultMemPool(&pool, dev) == cudaSuccess &&
                cudaMemPoolGetAttribute(pool, cudaMemPoolAttrReservedMemCurrent, &reserved) == cudaSuccess &&
                cudaMemPoolGetAttribute(pool, cudaMemPoolAttrUsedMemCurrent, &used) == cudaSuccess && reserved > used)
                freeB += (size_t) (reserved - used);
        }
        int nCta = (int) std::min<int64_t>((int64_t) b->gOrder.size(), (int64_t) nSM * 8);
        while (nCta > 1 && (double) nCta * slab * 8.0 > 0.6 * (double) freeB) nCta = nCta / 2;
        if ((double) nCta * slab * 8.0 > 0.9 * (double) freeB) {
            sa_set_error("sa_batch_prepare: a single work item needs %.1f GB of forward-matrix scratch (band too wide)",
                         slab * 8.0 / 1e9);
            return SA_ERR_INPUT;
        }
        b->gCta = std::max(1, nCta);
        if ((rc = b->dGScratch.alloc((size_t) b->gCta * slab, st))) return rc;
        g.items = b->dGItems.p; g.order = b->dGOrder.p; g.nItems = (int32_t) b->gOrder.size();
        g.pathOff = b->dPathOff.p; g.pathKmer = b->dPathKmer.p;
        // alignment runs: sa_mp_kernel (working set in shared memory when the widest diagonal fits; SA_NO_MP keeps the old one)
        if (mpKernel) {
            if ((rc = b->dPcCol.upload(pcCol, st, &b->h2dBytes))) return rc;
            g.pcCol = b->dPcCol.p;
            const size_t need = (size_t) 3 * maxDiagPC * 3 * sizeof(double);
            g.mpCap = need <= kSmemCap - 2048 ? (int32_t) maxDiagPC : 0;
            b->mpSmem = g.mpCap > 0 ? need : 0;
            b->mpThreads = 256;
            static size_t raised = 0;
            if (b->mpSmem > 48 * 1024 && b->mpSmem > raised) {
                CK(cudaFuncSetAttribute(sa_mp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) kSmemCap));
                raised = kSmemCap;
            }
            int bps = 1;
            CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, sa_mp_kernel, 256, b->mpSmem));
            b->gCta = std::min(b->gCta, nSM * std::max(1, bps));
        }
        g.evMean = b->dEvMean.p; g.evNoise = b->dEvNoise.p; g.evLogNoise = b->dEvLogNoise.p;
        g.bandL = b->dBandL.p; g.bandW = b->dBandW.p; g.bandOff = b->dBandOff.p;
        g.kt = m->dev->kt; g.gapX = m->dev->gapX;
        memcpy(g.trans, m->trans, sizeof(g.trans));
        g.kind = m->kind; g.A = m->A; g.suffixMod = m->K / m->A;
        {
            auto magic = [](uint64_t D, uint32_t *M, int32_t *sh) {
                int l = 0;
                while ((1ull << l) < D) l++;
                *sh = 31 + l;
                *M = (uint32_t) (((1ull << (31 + l)) + D - 1) / D);
            };
            magic((uint64_t) g.A, &g.divA_M, &g.divA_sh);
            magic((uint64_t) g.suffixMod, &g.divS_M, &g.divS_sh);
        }
        g.G = m->G; g.grid = m->dev->grid; g.pp = m->dev->pp; g.slopes = m->dev->slopes; g.kmerToRow = m->dev->kmerToRow;
        g.threshold = p->threshold;
        g.logThrGate = p->threshold > 0 ? log(p->threshold) - 1e-6 : -INFINITY;
        g.minDiags = p->minDiagsBetweenTraceBack; g.traceBack = p->traceBackDiagonals;
        g.expansion = p->diagonalExpansion;
        g.scratch = b->dGScratch.p; g.slabDoubles = slab; g.ringSize = ring; g.maxDiagPC = maxDiagPC; g.maxWidth = maxWidth;
        g.pairs = b->dPairs.p; g.pairCount = b->dPairCount.p; g.expOut = b->dExpOut.p;
        g.assignK = b->dAssignK.p; g.assignE = b->dAssignE.p; g.assignCount = b->expectations ? b->dAssignCount.p : nullptr;
    }
    CK(cudaEventRecord(b->evPrepared, st));
    CK(cudaStreamSynchronize(st));
    timer.lap("general-kernel packing");
    *out = b.release();
    return SA_OK;
}

extern "C" int sa_batch_prepare(const sa_model_t *m, const sa_read_t *reads, int64_t nReads, const sa_params_t *p,
                                const sa_ambig_t *ambig, int expectations, sa_batch_t **out) {
    return prepareBatch(m, reads, nReads, p, ambig, expectations, 0, 1, out);
}

// queue the batch's kernels on its stream; returns without waiting for them
static int launchBatch(sa_batch_t *b) {
    cudaStream_t st = b->stream;
    CK(cudaStreamWaitEvent(st, b->evPrepared, 0));
    b->lastLaunches = 0;
    CK(cudaMemsetAsync(b->dCounters.p, 0, 16 * sizeof(int32_t), st));
    CK(cudaEventRecord(b->ev[0], st));
    const bool latStream = !b->latOrder.empty() || !b->splitOrder.empty();
    if (latStream) {                           // first, so that its CTAs are resident before the persistent grid fills the SMs
        CK(cudaEventRecord(b->evFork, st));
        CK(cudaStreamWaitEvent(b->lat, b->evFork, 0));
        CK(cudaStreamWaitEvent(b->aux, b->evFork, 0));
    }
    if (!b->splitOrder.empty()) {              // on the aux stream (idle during a run): beside the ring-mode latency kernel
        FParams f = b->fp;
        f.order = b->dSplitOrder.p; f.nOrder = (int32_t) b->splitOrder.size(); f.counter = b->dCounters.p + 6;
        f.scratch = b->dSplitRows.p;
        const bool hdp = b->model->kind == SA_EMIT_HDP;
        if (hdp) sa_cta_forward_kernel<4, 1, true><<<(unsigned) b->splitOrder.size(), 128, 0, b->aux>>>(f);
        else sa_cta_forward_kernel<4, 1, false><<<(unsigned) b->splitOrder.size(), 128, 0, b->aux>>>(f);
        CK(cudaGetLastError());
        f.counter = b->dCounters.p + 7;
        f.blocks = b->dBlocks.p; f.nBlocks = (int32_t) b->blocks.size(); f.blockPairCount = b->dBlockCount.p;
        f.pairs = b->dSplitPairs.p;
        if (hdp) sa_cta_block_kernel<2, 2, true><<<b->blockGrid, 64, 0, b->aux>>>(f);
        else sa_cta_block_kernel<2, 2, false><<<b->blockGrid, 64, 0, b->aux>>>(f);
        CK(cudaGetLastError());
        b->lastLaunches += 2;
        CK(cudaEventRecord(b->evJoinSplit, b->aux));
    }
    if (!b->latOrder.empty()) {
        FParams f = b->fp;
        f.order = b->dLatOrder.p; f.nOrder = (int32_t) b->latOrder.size(); f.counter = b->dCounters.p + 5;
        f.rowsCap = b->latRows;
        f.scratch = b->dLatScratch.p;
        if (b->model->kind == SA_EMIT_HDP) sa_cta_kernel<4, 1, false, true><<<b->latBlocks, 128, 0, b->lat>>>(f);
        else sa_cta_kernel<4, 1, false, false><<<b->latBlocks, 128, 0, b->lat>>>(f);
        b->lastLaunches++;
        CK(cudaGetLastError());
    }
    if (latStream) CK(cudaEventRecord(b->evJoin, b->lat));
    bool wideFork = false;
    static const bool wideOneStream = getenv("SA_WIDE_ONE_STREAM") != nullptr;
    for (int c = kWideClasses - 1; c >= 0; c--) {   // wide kernel, widest window first (the heaviest items)
        if (b->wOrder[c].empty()) continue;
        cudaStream_t st = wideOneStream ? b->stream : b->wideOwner[c].s;
        if (!wideOneStream) {
            if (!wideFork) { CK(cudaEventRecord(b->evFork, b->stream)); wideFork = true; }
            CK(cudaStreamWaitEvent(st, b->evFork, 0));
        }
        const bool hdp = b->model->kind == SA_EMIT_HDP;
        FParams f = b->fp;
        f.order = b->dWOrder[c].p; f.nOrder = (int32_t) b->wOrder[c].size(); f.counter = b->dCounters.p + 8 + c;
        f.wNSW = kWideNSW[c]; f.wRowsCap = b->wRows[c]; f.wRingDoubles = b->wRingDoubles[c];
        f.scratch = b->dWRing[c].p; f.wHdr = b->dWHdr[c].p; f.wFold = b->dWFold[c].p; f.wRings = b->dWRings[c].p;
        if (b->fp32) {
            f.wOff = b->dWOff[c].p; f.wOffCap = b->wOffCap[c];
            wide32Launch(c, b->wBlocks[c], wide32Threads(c), wide32SmemBytes(c, b->dWFold[c].p == nullptr), st, f);
        } else {
            const bool gr = wideGlobalRings(c);
            wideLaunch(c, hdp, b->expectations, b->wBlocks[c], wideSmemBytes(gr, kWideNSW[c], b->dWFold[c].p == nullptr), st, f);
        }
        b->lastLaunches++;
        CK(cudaGetLastError());
        if (!wideOneStream) CK(cudaEventRecord(b->evWideJoin[c], st));
    }
    for (int c = 3; c >= 0; c--) {             // widest class first; [ev0, ev1) = wide + W >= 4, [ev1, ev2) = W <= 2
        if (c == 1) {
            if (!wideOneStream)
                for (int wc = 0; wc < kWideClasses; wc++)
                    if (!b->wOrder[wc].empty()) CK(cudaStreamWaitEvent(st, b->evWideJoin[wc], 0));
            if (latStream) CK(cudaStreamWaitEvent(st, b->evJoin, 0));   // counted with the 128-slot class
            if (!b->splitOrder.empty()) CK(cudaStreamWaitEvent(st, b->evJoinSplit, 0));
            CK(cudaEventRecord(b->ev[1], st));
        }
        if (b->fOrder[c].empty()) continue;
        FParams f = b->fp;
        f.order = b->dFOrder[c].p; f.nOrder = (int32_t) b->fOrder[c].size(); f.counter = b->dCounters.p + c;
        f.rowsCap = b->fRows[c];
        ctaLaunch(c, b->oneCell, b->expectations, b->model->kind == SA_EMIT_HDP, b->fBlocks[c], st, f);
        b->lastLaunches++;
        CK(cudaGetLastError());
    }
    CK(cudaEventRecord(b->ev[2], st));
    if (!b->gOrder.empty()) {
        GParams g = b->gp;
        g.counter = b->dCounters.p + 4;
        if (b->expectations) sa_general_kernel<true><<<b->gCta, 128, 0, st>>>(g);
        else if (b->mpThreads) sa_mp_kernel<<<b->gCta, b->mpThreads, b->mpSmem, st>>>(g);
        else sa_general_kernel<false><<<b->gCta, 128, 0, st>>>(g);
        b->lastLaunches++;
        CK(cudaGetLastError());
    }
    CK(cudaEventRecord(b->ev[3], st));
    return SA_OK;
}

extern "C" int sa_batch_run(sa_batch_t *b) {
    int rc = launchBatch(b);
    if (rc) return rc;
    CK(cudaStreamSynchronize(b->stream));
    // lastMs: [0] fast kernel C=4, [1] fast kernel C=2, [2] general kernel
    CK(cudaEventElapsedTime(&b->lastMs[0], b->ev[0], b->ev[1]));
    CK(cudaEventElapsedTime(&b->lastMs[1], b->ev[1], b->ev[2]));
    CK(cudaEventElapsedTime(&b->lastMs[2], b->ev[2], b->ev[3]));
    return SA_OK;
}

// meta per segment (an item, or one traceback block of a split item): source offset, END of its destination range, count,
// (x1 << 32 | y1).  The j-th pair written goes to dstEnd - 1 - j: a region's list is popped from its end.
__global__ void __launch_bounds__(128) sa_compact_pairs_kernel(const sa_pair_t *src, sa_pair_t *dst, const int64_t *meta, int64_t nSeg) {
    for (int64_t it = blockIdx.x; it < nSeg; it += gridDim.x) {
        const int64_t so = meta[4 * it], dE = meta[4 * it + 1], cnt = meta[4 * it + 2], sh = meta[4 * it + 3];
        const int32_t x1 = (int32_t) (sh >> 32), y1 = (int32_t) (sh & 0xffffffffLL);
        for (int64_t j = threadIdx.x; j < cnt; j += blockDim.x) {
            sa_pair_t pr = src[so + j];
            pr.x += x1;
            pr.y += y1;
            dst[dE - 1 - j] = pr;
        }
    }
}

static int growAndRerun(sa_batch_t *b) {
    // some item produced more pairs / assignments than its slot: re-slot with the exact counts and run again
    int64_t n = (int64_t) b->items.size();
    int64_t pt = 0, at = 0;
    for (int64_t i = 0; i < n; i++) {
        ItemRec &it = b->items[i];
        if (!b->expectations) {
            it.pairCap = std::max(it.pairCap, b->pairCountHost[i]);
            it.pairOff = pt;
            pt += it.pairCap;
        } else if (!b->assignCountHost.empty()) {
            it.assignCap = std::max(it.assignCap, b->assignCountHost[i]);
            it.assignOff = at;
            at += it.assignCap;
        }
        if (it.kernel) {
            b->fItems[i].pairOff = it.pairOff; b->fItems[i].pairCap = it.pairCap;
            b->fItems[i].assignOff = it.assignOff; b->fItems[i].assignCap = it.assignCap;
        } else {
            b->gItems[i].pairOff = it.pairOff; b->gItems[i].pairCap = it.pairCap;
            b->gItems[i].assignOff = it.assignOff; b->gItems[i].assignCap = it.assignCap;
        }
    }
    int rc;
    bool anyBand = !b->latOrder.empty() || !b->splitOrder.empty();   // (latency-class / split items live in fItems too)
    for (int c = 0; c < 4; c++) anyBand = anyBand || !b->fOrder[c].empty();
    for (int c = 0; c < kWideClasses; c++) anyBand = anyBand || !b->wOrder[c].empty();
    if (anyBand) {
        if ((rc = b->dFItems.upload(b->fItems, b->stream, &b->h2dBytes))) return rc;
        b->fp.items = b->dFItems.p;
    }
    if (!b->gOrder.empty()) {
        if ((rc = b->dGItems.upload(b->gItems, b->stream, &b->h2dBytes))) return rc;
        b->gp.items = b->dGItems.p;
    }
    if (!b->expectations) {
        if ((rc = b->dPairs.alloc(std::max<int64_t>(1, pt), b->stream))) return rc;
        b->gp.pairs = b->dPairs.p;
        b->fp.pairs = b->dPairs.p;
    } else {
        if ((rc = b->dAssignK.alloc(std::max<int64_t>(1, at), b->stream))) return rc;
        if ((rc = b->dAssignE.alloc(std::max<int64_t>(1, at), b->stream))) return rc;
        b->gp.assignK = b->dAssignK.p;
        b->gp.assignE = b->dAssignE.p;
        b->fp.assignK = b->dAssignK.p;
        b->fp.assignE = b->dAssignE.p;
    }
    return sa_batch_run(b);
}

// Growing page-locked destination of the compacted pairs: it becomes the result's `pairs` array (sa_result_free hands
// the slab back to the pool), so the dense D2H is the only copy of the result that is ever made.
struct PairSink {
    sa_pair_t *p = nullptr;
    int64_t cap = 0, used = 0;
};
static int sinkReserve(PairSink &k, int64_t need) {
    if (k.p && need <= k.cap) return SA_OK;
    int64_t cap = std::max(need, k.cap * 2);
    sa_pair_t *q = (sa_pair_t *) pinAcquire(sizeof(sa_pair_t) * (size_t) std::max<int64_t>(1, cap));
    if (!q) {
        sa_set_error("sa_batch_fetch: cannot allocate pinned memory for the result");
        return SA_ERR_CUDA;
    }
    if (k.used) memcpy(q, k.p, sizeof(sa_pair_t) * (size_t) k.used);
    pinRelease(k.p);
    k.p = q;
    k.cap = cap;
    return SA_OK;
}

// split items: per-block pair counts (blocks that out-grew their slot get the exact size and the block kernel runs again
// over the rows still in HBM); the item's count is the sum of its blocks'
static int fetchBlockCounts(sa_batch_t *b) {
    cudaStream_t st = b->stream;
    const int64_t nb = (int64_t) b->blocks.size();
    for (int attempt = 0; attempt < 3; attempt++) {
        b->blockCountHost.assign((size_t) nb, 0);
        CK(cudaMemcpyAsync(b->blockCountHost.data(), b->dBlockCount.p, sizeof(int64_t) * nb, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        b->d2hBytes += sizeof(int64_t) * nb;
        bool overflow = false;
        for (int64_t k = 0; k < nb; k++) overflow |= b->blockCountHost[k] > b->blocks[k].pairCap;
        if (!overflow) break;
        if (attempt == 2) {
            sa_set_error("sa_batch_fetch: traceback-block slots still overflow after re-run");
            return SA_ERR_CUDA;
        }
        int64_t total = 0;
        for (int64_t k = 0; k < nb; k++) {
            b->blocks[k].pairCap = std::max(b->blocks[k].pairCap, b->blockCountHost[k]);
            b->blocks[k].pairOff = total;
            total += b->blocks[k].pairCap;
        }
        int rc;
        if ((rc = b->dBlocks.upload(b->blocks, st, &b->h2dBytes))) return rc;
        if ((rc = b->dSplitPairs.alloc((size_t) std::max<int64_t>(1, total), st))) return rc;
        CK(cudaMemsetAsync(b->dCounters.p + 7, 0, sizeof(int32_t), st));
        FParams f = b->fp;
        f.scratch = b->dSplitRows.p;
        f.counter = b->dCounters.p + 7;
        f.blocks = b->dBlocks.p; f.nBlocks = (int32_t) nb; f.blockPairCount = b->dBlockCount.p;
        f.pairs = b->dSplitPairs.p;
        if (b->model->kind == SA_EMIT_HDP) sa_cta_block_kernel<2, 2, true><<<b->blockGrid, 64, 0, st>>>(f);
        else sa_cta_block_kernel<2, 2, false><<<b->blockGrid, 64, 0, st>>>(f);
        CK(cudaGetLastError());
        b->lastLaunches++;
    }
    for (int32_t i : b->splitOrder) b->pairCountHost[i] = 0;     // (the forward-only sweep writes no item count)
    for (int64_t k = 0; k < nb; k++) b->pairCountHost[b->blocks[k].item] += b->blockCountHost[k];
    for (int32_t i : b->splitOrder) b->items[i].pairCap = std::max(b->items[i].pairCap, b->pairCountHost[i]);
    return SA_OK;
}

// wait for the batch's kernels, read the per-item counts, re-run with larger slots if any overflowed
static int fetchCounts(sa_batch_t *b) {
    int64_t n = (int64_t) b->items.size();
    cudaStream_t st = b->stream;
    for (int attempt = 0; attempt < 3; attempt++) {
        bool overflow = false;
        if (!b->expectations) {
            b->pairCountHost.assign(std::max<int64_t>(1, n), 0);
            CK(cudaMemcpyAsync(b->pairCountHost.data(), b->dPairCount.p, sizeof(int64_t) * std::max<int64_t>(1, n),
                               cudaMemcpyDeviceToHost, st));
            CK(cudaStreamSynchronize(st));
            b->d2hBytes += sizeof(int64_t) * n;
            if (!b->splitOrder.empty()) {
                int rc = fetchBlockCounts(b);
                if (rc) return rc;
            }
            for (int64_t i = 0; i < n; i++) overflow |= b->pairCountHost[i] > b->items[i].pairCap;
        } else if (b->model->kind == SA_EMIT_HDP) {
            b->assignCountHost.assign(std::max<int64_t>(1, n), 0);
            CK(cudaMemcpyAsync(b->assignCountHost.data(), b->dAssignCount.p, sizeof(int64_t) * std::max<int64_t>(1, n),
                               cudaMemcpyDeviceToHost, st));
            CK(cudaStreamSynchronize(st));
            b->d2hBytes += sizeof(int64_t) * n;
            for (int64_t i = 0; i < n; i++) overflow |= b->assignCountHost[i] > b->items[i].assignCap;
        }
        if (!overflow) break;
        if (attempt == 2) {
            sa_set_error("sa_batch_fetch: output slots still overflow after re-run");
            return SA_ERR_CUDA;
        }
        int rc = growAndRerun(b);
        if (rc) return rc;
    }
    return SA_OK;
}

// Pairs of the batch, compacted on the device and copied to sink.p[sink.used ...]; pairOffsets[0 .. nReads] receive
// ABSOLUTE offsets into the sink.  The per-item slots are over-provisioned; a gather kernel packs the used prefixes
// into the final layout (items in (read, region) order, each region's list popped from its END and shifted to read
// coordinates: alignedPairCoordinateCorrectionFn, pairwiseAligner.c:2043-2050), so one dense D2H is the result.
static int fetchPairs(sa_batch_t *b, PairSink &sink, int64_t *pairOffsets) {
    int64_t n = (int64_t) b->items.size();
    cudaStream_t st = b->aux;   // (fetchCounts has waited for the band kernels)
    int64_t total = 0;
    std::vector<int64_t> meta((size_t) std::max<int64_t>(1, n) * 4), metaSplit;
    std::vector<int64_t> itemBase;           // split items: start of the item's destination range
    if (!b->splitOrder.empty()) itemBase.assign((size_t) n, -1);
    for (int32_t s : b->splitOrder) itemBase[s] = 0;
    int64_t i = 0;
    for (int64_t rd = 0; rd < b->nReads; rd++) {
        pairOffsets[rd] = sink.used + total;
        for (; i < n && b->items[i].read == rd; i++) {
            const bool split = !itemBase.empty() && itemBase[i] >= 0;
            if (split) itemBase[i] = total;
            meta[4 * i] = b->items[i].pairOff;
            meta[4 * i + 1] = total + b->pairCountHost[i];
            meta[4 * i + 2] = split ? 0 : b->pairCountHost[i];   // (split items are gathered block by block below)
            meta[4 * i + 3] = (b->items[i].x1 << 32) | (b->items[i].y1 & 0xffffffffLL);
            total += b->pairCountHost[i];
        }
    }
    if (!b->splitOrder.empty()) {
        // the item's list is its blocks' lists in block order; the whole list is popped from its end
        std::vector<int64_t> done((size_t) n, 0);
        for (size_t k = 0; k < b->blocks.size(); k++) {
            const FBlock &B = b->blocks[k];
            const int64_t it = B.item;
            metaSplit.push_back(B.pairOff);
            metaSplit.push_back(itemBase[it] + b->pairCountHost[it] - done[it]);
            metaSplit.push_back(b->blockCountHost[k]);
            metaSplit.push_back((b->items[it].x1 << 32) | (b->items[it].y1 & 0xffffffffLL));
            done[it] += b->blockCountHost[k];
        }
    }
    pairOffsets[b->nReads] = sink.used + total;
    int rc = sinkReserve(sink, sink.used + total);
    if (rc) return rc;
    if (total > 0) {
        DevBuf<int64_t> dMeta;
        DevBuf<sa_pair_t> dDense;
        if ((rc = dMeta.upload(meta, st, &b->h2dBytes))) return rc;
        if ((rc = dDense.alloc(total, st))) return rc;
        sa_compact_pairs_kernel<<<(unsigned) std::min<int64_t>(n, 148 * 16), 128, 0, st>>>(b->dPairs.p, dDense.p, dMeta.p, n);
        CK(cudaGetLastError());
        DevBuf<int64_t> dMetaSplit;
        if (!metaSplit.empty()) {
            const int64_t ns = (int64_t) metaSplit.size() / 4;
            if ((rc = dMetaSplit.upload(metaSplit, st, &b->h2dBytes))) return rc;
            sa_compact_pairs_kernel<<<(unsigned) std::min<int64_t>(ns, 148 * 16), 128, 0, st>>>(b->dSplitPairs.p, dDense.p,
                                                                                          dMetaSplit.p, ns);
            CK(cudaGetLastError());
            b->lastLaunches++;
        }
        CK(cudaMemcpyAsync(sink.p + sink.used, dDense.p, sizeof(sa_pair_t) * total, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        b->d2hBytes += sizeof(sa_pair_t) * total;
        b->lastLaunches++;
    }
    sink.used += total;
    return SA_OK;
}

extern "C" int sa_batch_fetch(sa_batch_t *b, sa_result_t **out) {
    *out = nullptr;
    PhaseTimer timer("fetch");
    int64_t n = (int64_t) b->items.size();
    cudaStream_t st = b->stream;
    b->d2hBytes = 0;
    int rc = fetchCounts(b);
    if (rc) return rc;
    timer.lap("wait for the kernels + counts");
    sa_result_t *r = (sa_result_t *) calloc(1, sizeof(sa_result_t));
    r->nReads = b->nReads;
    r->totalBandCells = b->bandCells;
    r->totalPathCells = b->pathCells;
    r->totalEvents = b->events;
    // items are created read by read, region by region: already in (read, region) order
    r->pairOffsets = (int64_t *) calloc(b->nReads + 1, sizeof(int64_t));
    if (!b->expectations) {
        PairSink sink;
        if ((rc = fetchPairs(b, sink, r->pairOffsets))) {
            pinRelease(sink.p);
            sa_result_free(r);
            return rc;
        }
        r->pairs = sink.p;
        timer.lap("compact + D2H pairs");
    } else {
        std::vector<double> e((size_t) std::max<int64_t>(1, n * 10));
        if (n > 0) {
            CK(cudaMemcpyAsync(e.data(), b->dExpOut.p, sizeof(double) * n * 10, cudaMemcpyDeviceToHost, st));
            CK(cudaStreamSynchronize(st));
            b->d2hBytes += sizeof(double) * n * 10;
        }
        r->transitions = (double *) calloc(std::max<int64_t>(1, b->nReads * 9), sizeof(double));
        r->likelihood = (double *) calloc(std::max<int64_t>(1, b->nReads), sizeof(double));
        for (int64_t i = 0; i < n; i++) {
            int64_t rd = b->items[i].read;
            for (int j = 0; j < 9; j++) r->transitions[rd * 9 + j] += e[i * 10 + j];
            r->likelihood[rd] += e[i * 10 + 9];
        }
        if (b->model->kind == SA_EMIT_HDP) {
            int64_t totalCap = 0, total = 0;
            for (int64_t i = 0; i < n; i++) {
                totalCap = std::max(totalCap, b->items[i].assignOff + b->items[i].assignCap);
                total += b->assignCountHost[i];
            }
            std::vector<int32_t> ak((size_t) std::max<int64_t>(1, totalCap)), ae((size_t) std::max<int64_t>(1, totalCap));
            if (totalCap > 0) {
                CK(cudaMemcpyAsync(ak.data(), b->dAssignK.p, sizeof(int32_t) * totalCap, cudaMemcpyDeviceToHost, st));
                CK(cudaMemcpyAsync(ae.data(), b->dAssignE.p, sizeof(int32_t) * totalCap, cudaMemcpyDeviceToHost, st));
                CK(cudaStreamSynchronize(st));
                b->d2hBytes += 2 * sizeof(int32_t) * totalCap;
            }
            r->assignOffsets = (int64_t *) calloc(b->nReads + 1, sizeof(int64_t));
            r->assignKmerPos = (int32_t *) malloc(sizeof(int32_t) * std::max<int64_t>(1, total));
            r->assignEvent = (int32_t *) malloc(sizeof(int32_t) * std::max<int64_t>(1, total));
            int64_t w = 0, i = 0;
            for (int64_t rd = 0; rd < b->nReads; rd++) {
                r->assignOffsets[rd] = w;
                for (; i < n && b->items[i].read == rd; i++) {
                    const ItemRec &it = b->items[i];
                    for (int64_t q = 0; q < b->assignCountHost[i]; q++) {
                        r->assignKmerPos[w] = ak[it.assignOff + q] + (int32_t) it.x1;
                        r->assignEvent[w] = ae[it.assignOff + q] + (int32_t) it.y1;
                        w++;
                    }
                }
            }
            r->assignOffsets[b->nReads] = w;
        }
    }
    *out = r;
    return SA_OK;
}

extern "C" int sa_batch_last_timing(const sa_batch_t *b, float *ms3, int64_t *launches) {
    if (ms3) memcpy(ms3, b->lastMs, sizeof(float) * 3);
    if (launches) *launches = b->lastLaunches;
    return SA_OK;
}
extern "C" int64_t sa_batch_band_cells(const sa_batch_t *b) { return b->bandCells; }
extern "C" void sa_batch_kernel_items(const sa_batch_t *b, int64_t *o) {
    o[0] = (int64_t) b->gOrder.size();
    o[1] = (int64_t) (b->fOrder[0].size() + b->fOrder[1].size());
    o[2] = (int64_t) (b->fOrder[2].size() + b->fOrder[3].size() + b->latOrder.size() + b->splitOrder.size());
    o[3] = 0;
    for (int c = 0; c < kWideClasses; c++) o[3] += (int64_t) b->wOrder[c].size();
}
extern "C" void sa_batch_wide_items(const sa_batch_t *b, int64_t *o) {
    for (int c = 0; c < kWideClasses; c++) o[c] = (int64_t) b->wOrder[c].size();
}
extern "C" int64_t sa_batch_h2d_bytes(const sa_batch_t *b) { return b->h2dBytes; }
extern "C" int64_t sa_batch_d2h_bytes(const sa_batch_t *b) { return b->d2hBytes; }
extern "C" void sa_batch_free(sa_batch_t *b) { delete b; }

// ---- test hook: the device logAdd on arrays (mirrors the reference's own logAdd unit test,
// tests/signalPairwiseAlignerTest.c:115-125, but compared bit-for-bit against the oracle)
__global__ void sa_logadd_kernel(const double *x, const double *y, double *o, int64_t n) {
    sa_logAdd_init();
    int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) o[i] = sa_logAdd(x[i], y[i]);
}
extern "C" int sa_debug_logadd(const double *x, const double *y, double *out, int64_t n) {
    double *dx, *dy, *dout;
    CK(cudaMalloc(&dx, n * 8)); CK(cudaMalloc(&dy, n * 8)); CK(cudaMalloc(&dout, n * 8));
    CK(cudaMemcpy(dx, x, n * 8, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dy, y, n * 8, cudaMemcpyHostToDevice));
    sa_logadd_kernel<<<(unsigned) ((n + 255) / 256), 256>>>(dx, dy, dout, n);
    CK(cudaGetLastError());
    CK(cudaMemcpy(out, dout, n * 8, cudaMemcpyDeviceToHost));
    cudaFree(dx); cudaFree(dy); cudaFree(dout);
    return SA_OK;
}

static int runBatch(const sa_model_t *m, const sa_read_t *reads, int64_t nReads, const sa_params_t *p,
                    const sa_ambig_t *ambig, int expectations, sa_result_t **out) {
    sa_batch_t *b = nullptr;
    int rc = sa_batch_prepare(m, reads, nReads, p, ambig, expectations, &b);
    if (rc) return rc;
    rc = sa_batch_run(b);
    if (!rc) rc = sa_batch_fetch(b, out);
    sa_batch_free(b);
    return rc;
}

#include <omp.h>
// host threads of the packing passes (OpenMP); launchers such as torchrun export OMP_NUM_THREADS=1, which would make
// the host side of a multi-GPU job the bottleneck: the caller hands every rank its share of the cores instead
extern "C" void sa_set_host_threads(int n) { omp_set_num_threads(n > 0 ? n : 1); }

static int64_t g_lastH2d = 0, g_lastD2h = 0, g_lastLaunches = 0;
extern "C" void sa_last_call_stats(int64_t *h2dBytes, int64_t *d2hBytes, int64_t *launches) {
    if (h2dBytes) *h2dBytes = g_lastH2d;
    if (d2hBytes) *d2hBytes = g_lastD2h;
    if (launches) *launches = g_lastLaunches;
}

// The one-shot alignment call as a pipeline over CHUNKS of the batch (contiguous in read order, each its own
// sub-batch and stream): while the band kernel of chunk k runs, the host packs chunk k+1 and its H2D / geometry pass go
// ahead; the persistent CTAs of chunk k+1 take over the SMs as those of chunk k drain; the compaction + D2H of chunk k
// overlap chunk k+1.  Host work and both copy directions are hidden behind the kernels except for the first chunk's
// pack and the last chunk's fetch.  SA_CHUNKS overrides the chunk count.
static int alignPipelined(const sa_model_t *m, const sa_read_t *reads, int64_t nReads, const sa_params_t *p,
                          const sa_ambig_t *ambig, sa_result_t **out) {
    *out = nullptr;
    int64_t nch = nReads >= 2048 ? 4 : (nReads >= 512 ? 2 : 1);
    {   // heavy-tailed lengths: a read is one CTA's sequential sweep, and every chunk would carry its own longest reads,
        // which only start once the previous chunk's persistent CTAs leave — the makespans add up.  One chunk then.
        int64_t sumEv = 0, maxEv = 0;
        for (int64_t i = 0; i < nReads; i++) { sumEv += reads[i].nEvents; maxEv = std::max(maxEv, reads[i].nEvents); }
        int nSM = 148, dev = 0;
        if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&nSM, cudaDevAttrMultiProcessorCount, dev);
        if ((double) maxEv * nSM > 0.25 * (double) sumEv) nch = 1;
    }
    if (const char *e = getenv("SA_CHUNKS")) nch = std::max<int64_t>(1, std::min<int64_t>(atoll(e), std::max<int64_t>(1, nReads)));
    PhaseTimer timer("align");
    std::vector<std::unique_ptr<sa_batch>> bs;
    std::vector<int64_t> lo((size_t) nch + 1);
    for (int64_t k = 0; k <= nch; k++) lo[k] = nReads * k / nch;
    if (const char *e = getenv("SA_CHUNK_FRACS")) {   // experiments: relative chunk sizes, e.g. "1,3,3,2"
        std::vector<double> w;
        for (const char *q = e; *q;) {
            char *end = nullptr;
            double v = strtod(q, &end);
            if (end == q) break;
            if (v > 0) w.push_back(v);
            q = *end ? end + 1 : end;
        }
        if (!w.empty() && (int64_t) w.size() <= std::max<int64_t>(1, nReads)) {
            nch = (int64_t) w.size();
            lo.assign((size_t) nch + 1, 0);
            double tot = 0, acc = 0;
            for (double v : w) tot += v;
            for (int64_t k = 0; k < nch; k++) {
                acc += w[(size_t) k];
                lo[(size_t) k + 1] = std::max<int64_t>(lo[(size_t) k] + 1, (int64_t) llround((double) nReads * acc / tot));
            }
            lo[(size_t) nch] = nReads;
            for (int64_t k = nch - 1; k > 0; k--) lo[(size_t) k] = std::min(lo[(size_t) k], lo[(size_t) k + 1] - 1);
        }
    }
    int rc = SA_OK;
    for (int64_t k = 0; k < nch; k++) {
        sa_batch_t *b = nullptr;
        if ((rc = prepareBatch(m, reads + lo[k], lo[k + 1] - lo[k], p, ambig, 0, lo[k], nch, &b))) break;
        bs.emplace_back(b);
        if ((rc = launchBatch(b))) break;
    }
    if (rc) {   // the batch is refused as a whole
        for (auto &b : bs) cudaStreamSynchronize(b->stream);
        return rc;
    }
    timer.lap("chunks packed, uploaded and queued");
    sa_result_t *r = (sa_result_t *) calloc(1, sizeof(sa_result_t));
    r->nReads = nReads;
    r->pairOffsets = (int64_t *) calloc(nReads + 1, sizeof(int64_t));
    PairSink sink;
    int64_t capAll = 1;
    for (auto &b : bs) capAll += b->events + b->events / 2 + 64 * (int64_t) b->items.size();
    g_lastH2d = g_lastD2h = g_lastLaunches = 0;
    // first guess 1.5 pairs per event (nanopore reads give 1.0 - 1.3 at threshold 0.01); the sink grows when a chunk
    // needs more (every D2H into it has completed by then)
    rc = sinkReserve(sink, capAll);
    for (int64_t k = 0; k < nch && !rc; k++) {
        sa_batch *b = bs[k].get();
        b->d2hBytes = 0;
        if ((rc = fetchCounts(b))) break;
        if ((rc = fetchPairs(b, sink, r->pairOffsets + lo[k]))) break;
        r->totalBandCells += b->bandCells;
        r->totalPathCells += b->pathCells;
        r->totalEvents += b->events;
        g_lastH2d += b->h2dBytes;
        g_lastD2h += b->d2hBytes;
        g_lastLaunches += b->lastLaunches + 1;   // + the geometry pass
    }
    if (rc) {
        for (auto &b : bs) cudaStreamSynchronize(b->stream);
        pinRelease(sink.p);
        sa_result_free(r);
        return rc;
    }
    r->pairs = sink.p;
    timer.lap("results fetched");
    if (timer.on)   // device timeline of the chunks' kernels, relative to the first chunk's start
        for (int64_t k = 0; k < nch; k++) {
            float t0 = 0, t1 = 0;
            cudaEventElapsedTime(&t0, bs[0]->ev[0], bs[k]->ev[0]);
            cudaEventElapsedTime(&t1, bs[0]->ev[0], bs[k]->ev[3]);
            fprintf(stderr, "[sa timing] align / chunk %lld kernels: start +%.1f ms, end +%.1f ms\n", (long long) k, t0, t1);
        }
    *out = r;
    return SA_OK;
}

extern "C" int sa_gpu_align_batch(const sa_model_t *m, const sa_read_t *reads, int64_t nReads, const sa_params_t *p,
                                  const sa_ambig_t *ambig, sa_result_t **out) {
    return alignPipelined(m, reads, nReads, p, ambig, out);
}
extern "C" int sa_gpu_expectations_batch(const sa_model_t *m, const sa_read_t *reads, int64_t nReads,
                                         const sa_params_t *p, const sa_ambig_t *ambig, sa_result_t **out) {
    return runBatch(m, reads, nReads, p, ambig, 1, out);
}

// ================================================================================================ per-read scaling (N2)
// signalUtils_estimateNanoporeParams (reference impl/signalMachineUtils.c:186-225) for a batch of strands: see
// sa_estimate.cuh.  Host side: pack the strands (three event columns as arrays, the event map as int32), one kernel, copy
// the 8 doubles per strand and — when the events are to be drift-adjusted — the mean column back into the caller's events.
extern "C" int sa_gpu_estimate_params_batch(const sa_model_t *m, sa_strand_t *strands, int64_t n, int adjustEvents,
                                            sa_scaling_t *out) {
    if (sa_gpu_device_count() == 0) {
        sa_set_error("sa_gpu_estimate_params_batch: no CUDA device available; libsagpu has no CPU fallback");
        return SA_ERR_CUDA;
    }
    if (n <= 0) return SA_OK;
    int rc = uploadModel(m);
    if (rc) return rc;
    if (!m->dev->raw) {
        CK(cudaMalloc(&m->dev->raw, sizeof(double) * m->match.size()));
        CK(cudaMemcpy(m->dev->raw, m->match.data(), sizeof(double) * m->match.size(), cudaMemcpyHostToDevice));
    }
    if ((rc = poolInit())) return rc;
    std::vector<EstStrand> es((size_t) n);
    int64_t rt = 0, et = 0;
    for (int64_t i = 0; i < n; i++) {
        const sa_strand_t &S = strands[i];
        if (S.readLength < 0 || S.nEvents < 0 || S.readLength > 0x3fffffff || S.nEvents > 0x3fffffff) {
            sa_set_error("sa_gpu_estimate_params_batch: strand %lld: bad lengths", (long long) i);
            return SA_ERR_INPUT;
        }
        es[(size_t) i].readOff = rt; es[(size_t) i].mapOff = rt; es[(size_t) i].evOff = et;
        es[(size_t) i].readLength = (int32_t) S.readLength; es[(size_t) i].nEvents = (int32_t) S.nEvents;
        rt += S.readLength; et += S.nEvents;
    }
    PinVec<char> hRead;
    PinVec<int32_t> hMap;
    PinVec<double> hMean, hSd, hTime, hOut;
    if (!hRead.resize((size_t) std::max<int64_t>(1, rt)) || !hMap.resize((size_t) std::max<int64_t>(1, rt)) ||
        !hMean.resize((size_t) std::max<int64_t>(1, et)) || !hSd.resize((size_t) std::max<int64_t>(1, et)) ||
        !hTime.resize((size_t) std::max<int64_t>(1, et)) || !hOut.resize((size_t) n * 8)) {
        sa_set_error("sa_gpu_estimate_params_batch: cannot allocate pinned staging memory");
        return SA_ERR_CUDA;
    }
    bool mapBad = false;
#pragma omp parallel for schedule(dynamic, 8) reduction(|| : mapBad)
    for (int64_t i = 0; i < n; i++) {
        const sa_strand_t &S = strands[i];
        const EstStrand &E = es[(size_t) i];
        memcpy(hRead.data() + E.readOff, S.read, (size_t) S.readLength);
        for (int64_t q = 0; q < S.readLength; q++) {
            const int64_t e = S.strandEventMap[q];
            mapBad = mapBad || e < 0 || e > 0x7fffffff;
            hMap[(size_t) (E.mapOff + q)] = (int32_t) e;
        }
        for (int64_t y = 0; y < S.nEvents; y++) {
            hMean[(size_t) (E.evOff + y)] = S.events[4 * y];
            hSd[(size_t) (E.evOff + y)] = S.events[4 * y + 1];
            hTime[(size_t) (E.evOff + y)] = S.events[4 * y + 3];
        }
    }
    if (mapBad) {
        sa_set_error("sa_gpu_estimate_params_batch: negative entry in a strand event map");
        return SA_ERR_INPUT;
    }
    cudaStream_t st;
    CK(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
    struct StreamGuard { cudaStream_t s; ~StreamGuard() { cudaStreamSynchronize(s); cudaStreamDestroy(s); } } guard{st};
    DevBuf<EstStrand> dS;
    DevBuf<char> dRead;
    DevBuf<int32_t> dMap, dErr;
    DevBuf<double> dMean, dSd, dTime, dOut;
    int64_t bytes = 0;
    if ((rc = dS.upload(es, st, &bytes)) || (rc = dRead.upload(hRead, st, &bytes)) || (rc = dMap.upload(hMap, st, &bytes)) ||
        (rc = dMean.upload(hMean, st, &bytes)) || (rc = dSd.upload(hSd, st, &bytes)) || (rc = dTime.upload(hTime, st, &bytes)) ||
        (rc = dOut.alloc((size_t) n * 8, st)) || (rc = dErr.alloc(1, st)))
        return rc;
    CK(cudaMemsetAsync(dErr.p, 0, sizeof(int32_t), st));
    CK(cudaMemsetAsync(dOut.p, 0, sizeof(double) * n * 8, st));
    EstParams P;
    memset(&P, 0, sizeof(P));
    P.strands = dS.p; P.reads = dRead.p; P.maps = dMap.p; P.mean = dMean.p; P.sd = dSd.p; P.time = dTime.p;
    P.table = m->dev->raw;
    for (int c = 0; c < 256; c++) {
        const size_t j = m->alphabet.find((char) c);
        P.code[c] = (c != 0 && j != std::string::npos) ? (int32_t) j : -1;
    }
    P.A = m->A; P.k = m->k; P.adjust = adjustEvents ? 1 : 0;
    P.out = dOut.p; P.err = dErr.p;
    sa_estimate_kernel<256><<<(unsigned) n, 256, 0, st>>>(P);
    CK(cudaGetLastError());
    int32_t err = 0;
    CK(cudaMemcpyAsync(hOut.data(), dOut.p, sizeof(double) * n * 8, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(&err, dErr.p, sizeof(err), cudaMemcpyDeviceToHost, st));
    if (adjustEvents && et > 0) CK(cudaMemcpyAsync(hMean.data(), dMean.p, sizeof(double) * et, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    if (err) {
        sa_set_error(err == 1 ? "sa_gpu_estimate_params_batch: a base-called strand contains a character outside the alphabet "
                                "(the reference exits in kmer_to_word, nanopore_hdp.c:393-398)"
                   : err == 3 ? "sa_gpu_estimate_params_batch: a strand event map points past the strand's events"
                              : "sa_gpu_estimate_params_batch: a strand has no assignments or a singular system (the reference "
                                "aborts: nanopore.c:762, :700)");
        return SA_ERR_INPUT;
    }
    for (int64_t i = 0; i < n; i++) {
        const double *o = hOut.data() + i * 8;
        sa_scaling_t &q = out[i];
        q.scale = o[0]; q.shift = o[1]; q.var = o[2]; q.shift_sd = o[3]; q.scale_sd = o[4]; q.var_sd = o[5]; q.drift = o[6];
        q.nAssignments = (int64_t) o[7];
    }
    if (adjustEvents) {
#pragma omp parallel for schedule(dynamic, 8)
        for (int64_t i = 0; i < n; i++)
            for (int64_t y = 0; y < strands[i].nEvents; y++) strands[i].events[4 * y] = hMean[(size_t) (es[(size_t) i].evOff + y)];
    }
    return SA_OK;
}

// ================================================================================================ SURVEY 8f N3
// sa_gpu_marginalize_variants: variant calls of aligned pairs -> per-position marginals (sa_variants.cuh)
extern "C" void sa_marginals_free(sa_marginals_t *r) {
    if (!r) return;
    free(r->positionOffsets); free(r->positions); free(r->probs); free(r->perRead);
    free(r);
}
extern "C" int sa_gpu_marginalize_variants(const sa_model_t *m, const sa_variant_read_t *reads, int64_t n, const char *variants,
                                           int roundLikeTsv, sa_marginals_t **out) {
    *out = nullptr;
    if (sa_gpu_device_count() == 0) {
        sa_set_error("sa_gpu_marginalize_variants: no CUDA device available; libsagpu has no CPU fallback");
        return SA_ERR_CUDA;
    }
    std::string vs(variants ? variants : "");
    std::sort(vs.begin(), vs.end());
    vs.erase(std::unique(vs.begin(), vs.end()), vs.end());
    if (vs.empty() || vs.size() > 31 || m->k > 15) {
        sa_set_error("sa_gpu_marginalize_variants: need 1..31 distinct variant characters and a kmer length <= 15");
        return SA_ERR_INPUT;
    }
    const int V = (int) vs.size();
    int rc = poolInit();
    if (rc) return rc;
    std::vector<VarRead> vr((size_t) std::max<int64_t>(1, n));
    int64_t pairTot = 0, rankTot = 0, massTot = 0;
    for (int64_t i = 0; i < n; i++) {
        const sa_variant_read_t &R = reads[i];
        if (R.nPairs < 0 || R.targetLength < 0 || (R.nPairs > 0 && !R.pairs) || (R.targetLength > 0 && !R.target)) {
            sa_set_error("sa_gpu_marginalize_variants: read %lld: bad sizes", (long long) i);
            return SA_ERR_INPUT;
        }
        VarRead &q = vr[(size_t) i];
        q.pairOff = pairTot; q.nPairs = R.nPairs; q.rankOff = rankTot; q.targetLen = R.targetLength; q.massOff = massTot;
        int32_t nX = 0;
        for (int64_t t = 0; t < R.targetLength; t++) nX += R.target[t] == 'X';
        q.nX = nX; q.pad = 0;
        pairTot += R.nPairs; rankTot += R.targetLength; massTot += (int64_t) nX * V;
    }
    PinVec<sa_pair_t> hPairs;
    PinVec<int32_t> hRank;
    PinVec<double> hMass;
    if (!hPairs.resize((size_t) std::max<int64_t>(1, pairTot)) || !hRank.resize((size_t) std::max<int64_t>(1, rankTot)) ||
        !hMass.resize((size_t) std::max<int64_t>(1, massTot))) {
        sa_set_error("sa_gpu_marginalize_variants: cannot allocate pinned staging memory");
        return SA_ERR_CUDA;
    }
#pragma omp parallel for schedule(dynamic, 8)
    for (int64_t i = 0; i < n; i++) {
        const sa_variant_read_t &R = reads[i];
        const VarRead &q = vr[(size_t) i];
        if (R.nPairs) memcpy(hPairs.data() + q.pairOff, R.pairs, sizeof(sa_pair_t) * (size_t) R.nPairs);
        int32_t r = 0;
        for (int64_t t = 0; t < R.targetLength; t++) hRank[(size_t) (q.rankOff + t)] = R.target[t] == 'X' ? r++ : -1;
    }
    sa_marginals_t *res = (sa_marginals_t *) calloc(1, sizeof(sa_marginals_t));
    res->nReads = n; res->nVariants = V;
    memcpy(res->variants, vs.c_str(), vs.size() + 1);
    res->positionOffsets = (int64_t *) calloc((size_t) n + 1, sizeof(int64_t));
    res->perRead = (double *) calloc((size_t) std::max<int64_t>(1, n) * V, sizeof(double));
    if (n > 0 && massTot > 0) {
        cudaStream_t st;
        CK(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
        struct StreamGuard { cudaStream_t s; ~StreamGuard() { cudaStreamSynchronize(s); cudaStreamDestroy(s); } } guard{st};
        DevBuf<VarRead> dR;
        DevBuf<sa_pair_t> dPairs;
        DevBuf<int32_t> dRank, dErr;
        DevBuf<double> dMass;
        int64_t bytes = 0;
        if ((rc = dR.upload(vr, st, &bytes)) || (rc = dPairs.upload(hPairs, st, &bytes)) || (rc = dRank.upload(hRank, st, &bytes)) ||
            (rc = dMass.alloc((size_t) massTot, st)) || (rc = dErr.alloc(1, st))) {
            sa_marginals_free(res);
            return rc;
        }
        CK(cudaMemsetAsync(dErr.p, 0, sizeof(int32_t), st));
        CK(cudaMemsetAsync(dMass.p, 0, sizeof(double) * massTot, st));
        VarParams P;
        memset(&P, 0, sizeof(P));
        P.reads = dR.p; P.pairs = dPairs.p; P.xRank = dRank.p; P.mass = dMass.p; P.err = dErr.p;
        P.A = m->A; P.k = m->k; P.V = V; P.roundLikeTsv = roundLikeTsv ? 1 : 0;
        for (int d = 0; d < 256; d++) {
            P.digitToVariant[d] = -1;
            if (d < m->A) {
                const size_t j = vs.find(m->alphabet[(size_t) d]);
                if (j != std::string::npos) P.digitToVariant[d] = (int32_t) j;
            }
        }
        sa_variant_mass_kernel<256><<<(unsigned) n, 256, 0, st>>>(P);
        CK(cudaGetLastError());
        int32_t err = 0;
        CK(cudaMemcpyAsync(hMass.data(), dMass.p, sizeof(double) * massTot, cudaMemcpyDeviceToHost, st));
        CK(cudaMemcpyAsync(&err, dErr.p, sizeof(err), cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        if (err) {
            sa_set_error(err == 1 ? "sa_gpu_marginalize_variants: a pair calls a base that is not among the variants (the reference "
                                    "raises in MarginalizeVariants.get_data, variantCaller.py:74)"
                                  : "sa_gpu_marginalize_variants: an aligned pair lies outside its target sequence");
            sa_marginals_free(res);
            return SA_ERR_INPUT;
        }
    }
    // normalise and compact (a few hundred positions per read): positions with no row are absent, order ascending
    int64_t nPosTot = 0;
    std::vector<int64_t> nPos((size_t) std::max<int64_t>(1, n), 0);
    for (int64_t i = 0; i < n; i++) {
        const VarRead &q = vr[(size_t) i];
        for (int32_t r = 0; r < q.nX; r++) {
            double tot = 0;
            for (int v = 0; v < V; v++) tot += hMass[(size_t) (q.massOff + (int64_t) r * V + v)];
            nPos[(size_t) i] += tot > 0;
        }
        res->positionOffsets[i + 1] = res->positionOffsets[i] + nPos[(size_t) i];
        nPosTot += nPos[(size_t) i];
    }
    res->positions = (int64_t *) calloc((size_t) std::max<int64_t>(1, nPosTot), sizeof(int64_t));
    res->probs = (double *) calloc((size_t) std::max<int64_t>(1, nPosTot) * V, sizeof(double));
    for (int64_t i = 0; i < n; i++) {
        const sa_variant_read_t &R = reads[i];
        const VarRead &q = vr[(size_t) i];
        const bool same = (!R.strandIsComplement && R.forward) || (R.strandIsComplement && !R.forward);
        int64_t w = res->positionOffsets[i];
        // ranks ascend with t; the reported position ascends with t (same orientation) or descends (mirrored)
        std::vector<int64_t> tOf((size_t) q.nX);
        {
            int32_t r = 0;
            for (int64_t t = 0; t < R.targetLength; t++)
                if (R.target[t] == 'X') tOf[(size_t) r++] = t;
        }
        for (int32_t rr = 0; rr < q.nX; rr++) {
            const int32_t r = same ? rr : q.nX - 1 - rr;
            const double *ms = hMass.data() + q.massOff + (int64_t) r * V;
            double tot = 0;
            for (int v = 0; v < V; v++) tot += ms[v];
            if (!(tot > 0)) continue;
            res->positions[w] = same ? tOf[(size_t) r] + R.referenceSequenceOffset : R.referenceSequenceOffset - 1 - tOf[(size_t) r];
            for (int v = 0; v < V; v++) {
                res->probs[w * V + v] = ms[v] / tot;
                res->perRead[i * V + v] += ms[v] / tot;
            }
            w++;
        }
        if (nPos[(size_t) i] > 0)
            for (int v = 0; v < V; v++) res->perRead[i * V + v] /= (double) nPos[(size_t) i];
    }
    *out = res;
    return SA_OK;
}

// ================================================================================================ SURVEY 8f N4
// sa_gpu_event_align_batch: adaptive banded event -> kmer alignment (sa_eventalign.cuh)
extern "C" void sa_event_alignment_free(sa_event_alignment_t *r) {
    if (!r) return;
    free(r->pairOffsets); free(r->kmerIdx); free(r->eventIdx); free(r->scale); free(r->shift); free(r->var);
    free(r->avgLogEmission); free(r->maxGap); free(r->spanned); free(r->failed); free(r->nAlignedBeforeQc);
    free(r);
}
extern "C" int sa_gpu_event_align_batch(const sa_model_t *m, const sa_event_read_t *reads, int64_t n, sa_event_alignment_t **out) {
    *out = nullptr;
    if (sa_gpu_device_count() == 0) {
        sa_set_error("sa_gpu_event_align_batch: no CUDA device available; libsagpu has no CPU fallback");
        return SA_ERR_CUDA;
    }
    if (m->match.size() != (size_t) m->K * 5) {
        sa_set_error("sa_gpu_event_align_batch: the model has no Gaussian match table");
        return SA_ERR_INPUT;
    }
    int rc = uploadModel(m);
    if (rc) return rc;
    if (!m->dev->raw) {
        CK(cudaMalloc(&m->dev->raw, sizeof(double) * m->match.size()));
        CK(cudaMemcpy(m->dev->raw, m->match.data(), sizeof(double) * m->match.size(), cudaMemcpyHostToDevice));
    }
    if ((rc = poolInit())) return rc;
    const int k = m->k;
    std::vector<EaRead> er((size_t) std::max<int64_t>(1, n));
    int64_t kTot = 0, eTot = 0, bTot = 0, pTot = 0;
    for (int64_t i = 0; i < n; i++) {
        const sa_event_read_t &R = reads[i];
        const int64_t nK = R.sequenceLength - k + 1;
        if (nK < 1 || R.nEvents < 1 || nK > 0x3fffffff || R.nEvents > 0x3fffffff || !R.sequence || !R.eventMean) {
            sa_set_error("sa_gpu_event_align_batch: read %lld: needs at least one kmer and one event", (long long) i);
            return SA_ERR_INPUT;
        }
        EaRead &q = er[(size_t) i];
        q.kmerOff = kTot; q.evOff = eTot; q.bandOff = bTot; q.traceOff = bTot * kEaBand; q.pairOff = 2 * pTot;
        q.nKmers = (int32_t) nK; q.nEvents = (int32_t) R.nEvents;
        kTot += nK; eTot += R.nEvents; bTot += nK + R.nEvents + 2; pTot += nK + R.nEvents;
    }
    PinVec<int32_t> hKid, hPairs;
    PinVec<float> hEv;
    PinVec<double> hQc;
    if (!hKid.resize((size_t) std::max<int64_t>(1, kTot)) || !hEv.resize((size_t) std::max<int64_t>(1, eTot)) ||
        !hPairs.resize((size_t) std::max<int64_t>(1, 2 * pTot)) || !hQc.resize((size_t) std::max<int64_t>(1, n) * 4)) {
        sa_set_error("sa_gpu_event_align_batch: cannot allocate pinned staging memory");
        return SA_ERR_CUDA;
    }
    int32_t code[256];
    for (int c = 0; c < 256; c++) {
        const size_t j = m->alphabet.find((char) c);
        code[c] = (c != 0 && j != std::string::npos) ? (int32_t) j : -1;
    }
    code[(unsigned char) 'U'] = code[(unsigned char) 'T'];        // build_kmer_list: stString_replace(sequence, "U", "T")
    bool bad = false;
#pragma omp parallel for schedule(dynamic, 4) reduction(|| : bad)
    for (int64_t i = 0; i < n; i++) {
        const sa_event_read_t &R = reads[i];
        EaRead &q = er[(size_t) i];
        for (int64_t x = 0; x < q.nKmers; x++) {
            int64_t id = 0;
            for (int j = 0; j < k; j++) {
                const int32_t c = code[(unsigned char) R.sequence[R.rna ? x + k - 1 - j : x + j]];     // RNA: the kmer reversed
                bad = bad || c < 0;
                id = id * m->A + (c < 0 ? 0 : c);
            }
            hKid[(size_t) (q.kmerOff + x)] = (int32_t) id;
        }
        memcpy(hEv.data() + q.evOff, R.eventMean, sizeof(float) * (size_t) R.nEvents);
        if (R.useMom) {            // estimate_scalings_using_mom (impl/eventAligner.c:792-842): sequential double sums
            double evSum = 0.0, kSum = 0.0, kSq = 0.0, evSq = 0.0;
            for (int64_t y = 0; y < R.nEvents; y++) evSum += (double) R.eventMean[y];
            for (int64_t x = 0; x < q.nKmers; x++) {
                const double lm = m->match[(size_t) hKid[(size_t) (q.kmerOff + x)] * 5];
                kSum += lm;
                kSq += lm * lm;
            }
            const double shift = evSum / (double) R.nEvents - kSum / (double) q.nKmers;
            for (int64_t y = 0; y < R.nEvents; y++) {
                const double dlt = (double) R.eventMean[y] - shift;
                evSq += dlt * dlt;
            }
            q.scale = (evSq / (double) R.nEvents) / (kSq / (double) q.nKmers);
            q.shift = shift;
            q.var = 1.0;
        } else {
            q.scale = R.scale; q.shift = R.shift; q.var = R.var;
        }
        q.logInvVar = log(1 / q.var);
        q.eventsPerKmer = (double) R.nEvents / (double) q.nKmers;
        const double pStay = 1 - (1 / (q.eventsPerKmer + 1));
        q.lpSkip = log(1e-10);
        q.lpStay = log(pStay);
        q.lpStep = log(1.0 - exp(q.lpSkip) - exp(q.lpStay));
        q.lpTrim = log(0.01);
    }
    if (bad) {
        sa_set_error("sa_gpu_event_align_batch: a sequence contains a character outside the model's alphabet");
        return SA_ERR_INPUT;
    }
    std::vector<double> c1((size_t) m->K);
    for (int64_t j = 0; j < m->K; j++) c1[(size_t) j] = -0.91893853320467267 - log(m->match[(size_t) j * 5 + 1]);
    sa_event_alignment_t *res = (sa_event_alignment_t *) calloc(1, sizeof(sa_event_alignment_t));
    const size_t nn = (size_t) std::max<int64_t>(1, n);
    res->nReads = n;
    res->pairOffsets = (int64_t *) calloc((size_t) n + 1, sizeof(int64_t));
    res->scale = (double *) calloc(nn, sizeof(double)); res->shift = (double *) calloc(nn, sizeof(double));
    res->var = (double *) calloc(nn, sizeof(double)); res->avgLogEmission = (double *) calloc(nn, sizeof(double));
    res->maxGap = (int32_t *) calloc(nn, sizeof(int32_t)); res->spanned = (int32_t *) calloc(nn, sizeof(int32_t));
    res->failed = (int32_t *) calloc(nn, sizeof(int32_t)); res->nAlignedBeforeQc = (int32_t *) calloc(nn, sizeof(int32_t));
    if (n > 0) {
        cudaStream_t st;
        CK(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
        struct StreamGuard { cudaStream_t s; ~StreamGuard() { cudaStreamSynchronize(s); cudaStreamDestroy(s); } } guard{st};
        DevBuf<EaRead> dR;
        DevBuf<int32_t> dKid, dLlE, dLlK, dPairs;
        DevBuf<float> dEv;
        DevBuf<double> dC1, dLast, dQc;
        DevBuf<uint8_t> dTrace;
        int64_t bytes = 0;
        if ((rc = dR.upload(er, st, &bytes)) || (rc = dKid.upload(hKid, st, &bytes)) || (rc = dEv.upload(hEv, st, &bytes)) ||
            (rc = dC1.upload(c1, st, &bytes)) || (rc = dLlE.alloc((size_t) bTot, st)) || (rc = dLlK.alloc((size_t) bTot, st)) ||
            (rc = dTrace.alloc((size_t) bTot * kEaBand, st)) || (rc = dLast.alloc((size_t) eTot, st)) ||
            (rc = dPairs.alloc((size_t) 2 * pTot, st)) || (rc = dQc.alloc((size_t) n * 4, st))) {
            sa_event_alignment_free(res);
            return rc;
        }
        EaParams P;
        memset(&P, 0, sizeof(P));
        P.reads = dR.p; P.kmerId = dKid.p; P.ev = dEv.p; P.table = m->dev->raw; P.c1 = dC1.p; P.trace = dTrace.p;
        P.llE = dLlE.p; P.llK = dLlK.p; P.lastCol = dLast.p; P.pairs = dPairs.p; P.qc = dQc.p;
        sa_event_align_kernel<<<(unsigned) n, 128, 0, st>>>(P);
        CK(cudaGetLastError());
        CK(cudaMemcpyAsync(hPairs.data(), dPairs.p, sizeof(int32_t) * 2 * pTot, cudaMemcpyDeviceToHost, st));
        CK(cudaMemcpyAsync(hQc.data(), dQc.p, sizeof(double) * n * 4, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
    }
    for (int64_t i = 0; i < n; i++) {
        const EaRead &q = er[(size_t) i];
        const double *qc = hQc.data() + i * 4;
        const int64_t na = (int64_t) qc[0];
        const double avg = qc[1] / qc[0];               // sum_emission / n_aligned_events (:1199)
        res->scale[i] = q.scale; res->shift[i] = q.shift; res->var[i] = q.var;
        res->avgLogEmission[i] = avg; res->maxGap[i] = (int32_t) qc[2]; res->spanned[i] = qc[3] != 0.0;
        res->nAlignedBeforeQc[i] = (int32_t) na;
        res->failed[i] = (avg < -5.2 || !(qc[3] != 0.0) || qc[2] > 50 || q.eventsPerKmer > 5.0) ? 1 : 0;      // :1205
        res->pairOffsets[i + 1] = res->pairOffsets[i] + (res->failed[i] ? 0 : na);
    }
    const size_t tot = (size_t) std::max<int64_t>(1, res->pairOffsets[n]);
    res->kmerIdx = (int32_t *) calloc(tot, sizeof(int32_t));
    res->eventIdx = (int32_t *) calloc(tot, sizeof(int32_t));
    for (int64_t i = 0; i < n; i++) {
        if (res->failed[i]) continue;
        const EaRead &q = er[(size_t) i];
        const int64_t na = res->nAlignedBeforeQc[i], cap = (int64_t) q.nKmers + q.nEvents;
        const int32_t *pp = hPairs.data() + q.pairOff + 2 * (cap - na);
        for (int64_t j = 0; j < na; j++) {
            res->kmerIdx[res->pairOffsets[i] + j] = pp[2 * j];
            res->eventIdx[res->pairOffsets[i] + j] = pp[2 * j + 1];
        }
    }
    *out = res;
    return SA_OK;
}

// ================================================================================================ allocation-free call
// sa_gpu_align_noalloc — ONE read through the wide band kernel (sa_wide.cuh, 512-slot window) WITHOUT any device
// allocation.  No counterpart in the reference: added for callers that cannot call
// cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
//
// Everything the regular path allocates per batch — the staged inputs, the work-item record, the traceback list, the
// row-header ring, the pair slots and the kernel's SCRATCH (the ring of forward rows the blocked traceback reads back) —
// lives in ONE file-scope device array of SA_NOALLOC_SCRATCH_BYTES (16 MiB), carved up below.  Every GPU that loads the
// library holds that array until the process exits.  Every call shares it and finds the previous call's contents: the
// work-queue counter is cleared on each call, everything else is overwritten before it is read.  Calls that overlap on the
// GPU (two host threads, two streams) RACE on it: one call at a time per device.
//
// In host code the array's name is a host-side stand-in, not a device address: its device address is taken once per
// device, with that device current, by cudaGetSymbolAddress, and only that pointer is handed to kernels.
static __device__ __align__(256) unsigned char sa_noalloc_scratch[SA_NOALLOC_SCRATCH_BYTES];

namespace noalloc {
constexpr int64_t kMaxEvents = SA_NOALLOC_MAX_EVENTS, kMaxKmers = SA_NOALLOC_MAX_KMERS, kMaxAnchors = SA_NOALLOC_MAX_KMERS;
constexpr int64_t kTbCap = 4096, kHdrCap = 8192, kPairCap = 131072;
constexpr int kNSW = 512, kNT = 256;
// byte offsets inside the scratch array (all multiples of 256)
constexpr size_t oCounter = 0, oItem = 256, oOrder = 1024, oGeo = 1280, oPairCount = 1536;
constexpr size_t oTb = 4096;                                   // int32[kTbCap]
constexpr size_t oHdr = oTb + kTbCap * 4;                      // int4[kHdrCap]
constexpr size_t oEv = oHdr + kHdrCap * 16;                    // double[kMaxEvents]
constexpr size_t oKmer = oEv + kMaxEvents * 8;                 // int32[kMaxKmers]
constexpr size_t oAnch = oKmer + kMaxKmers * 4;                // int32[2 * kMaxAnchors]
constexpr size_t oPairs = oAnch + kMaxAnchors * 8;             // sa_pair_t[kPairCap]
constexpr size_t oRing = oPairs + kPairCap * 16;               // double[...]: the forward-row ring
constexpr int64_t kRingDoubles = (int64_t) ((SA_NOALLOC_SCRATCH_BYTES - oRing) / 8);
static_assert(sizeof(FItem) <= oOrder - oItem && sizeof(FGeo) <= oPairCount - oGeo, "control block layout");
static_assert(oRing % 256 == 0 && oRing < SA_NOALLOC_SCRATCH_BYTES / 2, "scratch layout");

static unsigned char *g_base[64] = {nullptr};   // device address of the array, per device
// host staging of the packed inputs (fixed size: no heap allocation per call either)
static double g_hEv[kMaxEvents];
static int32_t g_hKmer[kMaxKmers], g_hAnch[2 * kMaxAnchors];

static int setup(int dev, unsigned char **base) {
    if (dev < 0 || dev >= 64) { sa_set_error("sa_gpu_align_noalloc: device index %d not supported", dev); return SA_ERR_CUDA; }
    if (!g_base[dev]) {
        unsigned char *ptr = nullptr;
        CK(cudaGetSymbolAddress((void **) &ptr, sa_noalloc_scratch));     // with device `dev` current
        CK(cudaFuncSetAttribute(sa_wide_kernel<false, false, false, kNSW, kNT, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) kSmemCap));
        g_base[dev] = ptr;
    }
    *base = g_base[dev];
    return SA_OK;
}
}  // namespace noalloc

extern "C" int sa_gpu_align_noalloc(const sa_model_t *m, const sa_read_t *R, const sa_params_t *p, sa_pair_t *pairsOut,
                                    int64_t pairCap, int64_t *nPairsOut) {
    using namespace noalloc;
    if (!m || !R || !p || !nPairsOut || (pairCap > 0 && !pairsOut)) { sa_set_error("sa_gpu_align_noalloc: null argument"); return SA_ERR_INPUT; }
    *nPairsOut = 0;
    if (sa_gpu_device_count() == 0) {
        sa_set_error("sa_gpu_align_noalloc: no CUDA device available; libsagpu has no CPU fallback");
        return SA_ERR_CUDA;
    }
    if (p->diagonalExpansion % 2 != 0 || p->diagonalExpansion < 0 || p->traceBackDiagonals < 1 ||
        p->minDiagsBetweenTraceBack < 2 || p->traceBackDiagonals + 1 >= p->minDiagsBetweenTraceBack) {
        sa_set_error("sa_gpu_align_noalloc: parameters violate the prerequisites of getPosteriorProbsWithBanding "
                     "(pairwiseAligner.c:1460-1464)");
        return SA_ERR_INPUT;
    }
    if (m->kind != SA_EMIT_GAUSS_MEANONLY_DESCALE || !modelFastEligible(m, false)) {
        sa_set_error("sa_gpu_align_noalloc: only the production Gaussian emission (MeanOnly, de-scaled) is supported here");
        return SA_ERR_INPUT;
    }
    int dev = 0;
    CK(cudaGetDevice(&dev));
    if (!m->dev || m->devDevice != dev) {
        sa_set_error("sa_gpu_align_noalloc: the model's tables are not resident on device %d and this entry point allocates "
                     "nothing: run any regular call with this model first (sa_batch_prepare, even with zero reads, uploads it)", dev);
        return SA_ERR_INPUT;
    }
    const int64_t lX = R->targetLength > 0 ? std::max<int64_t>(0, R->targetLength - (m->k - 1)) : 0, lY = R->nEvents;
    if (lX > kMaxKmers || lY > kMaxEvents || R->nAnchors > kMaxAnchors) {
        sa_set_error("sa_gpu_align_noalloc: read too large for the static scratch (%lld kmers, %lld events, %lld anchors; limits "
                     "%lld / %lld / %lld): use sa_gpu_align_batch", (long long) lX, (long long) lY, (long long) R->nAnchors,
                     (long long) kMaxKmers, (long long) kMaxEvents, (long long) kMaxAnchors);
        return SA_ERR_INPUT;
    }
    const int64_t nDiag = lX + lY;
    const int64_t nTbCap = nDiag / (p->minDiagsBetweenTraceBack - p->traceBackDiagonals - 1) + 2;
    if (nTbCap > kTbCap) {
        sa_set_error("sa_gpu_align_noalloc: %lld traceback blocks exceed the limit %lld", (long long) nTbCap, (long long) kTbCap);
        return SA_ERR_INPUT;
    }
    // anchors: in range, strictly increasing (band_construct asserts it), and no gap that getSplitPoints would cut at
    // (pairwiseAligner.c:1886-1937): the regions of a split read are separate work items of the regular path
    {
        int64_t x2 = 0, y2 = 0;
        for (int64_t i = 0; i <= R->nAnchors; i++) {
            const int64_t x3 = i < R->nAnchors ? R->anchors[2 * i] : lX, y3 = i < R->nAnchors ? R->anchors[2 * i + 1] : lY;
            if (i < R->nAnchors && (x3 < x2 || y3 < y2 || x3 >= lX || y3 >= lY)) {
                sa_set_error("sa_gpu_align_noalloc: anchor %lld (%lld, %lld) out of range or not strictly increasing",
                             (long long) i, (long long) x3, (long long) y3);
                return SA_ERR_INPUT;
            }
            if ((x3 - x2) * (y3 - y2) > p->splitMatrixBiggerThanThis) {
                sa_set_error("sa_gpu_align_noalloc: the read has an anchor gap that getSplitPoints cuts at; use sa_gpu_align_batch");
                return SA_ERR_INPUT;
            }
            if (i < R->nAnchors) {
                g_hAnch[2 * i] = (int32_t) x3; g_hAnch[2 * i + 1] = (int32_t) y3;
                x2 = x3 + 1; y2 = y3 + 1;
            }
        }
    }
    // rolling kmer index; every character must be in the alphabet (ambiguity codes need the general kernel)
    {
        int code[256];
        for (int c = 0; c < 256; c++) {
            const size_t j = m->alphabet.find((char) c);
            code[c] = (c != 0 && j != std::string::npos) ? (int) j : -1;
        }
        const int64_t suffixK = m->K / m->A;
        int64_t roll = 0;
        for (int64_t i = 0; i < (lX > 0 ? R->targetLength : 0); i++) {
            const int cc = code[(unsigned char) R->target[i]];
            if (cc < 0) {
                sa_set_error("sa_gpu_align_noalloc: character '%c' at reference position %lld is outside the alphabet %s "
                             "(ambiguity codes: use sa_gpu_align_batch)", R->target[i], (long long) i, m->alphabet.c_str());
                return SA_ERR_INPUT;
            }
            roll = (roll % suffixK) * m->A + cc;
            if (i >= m->k - 1) g_hKmer[i - (m->k - 1)] = (int32_t) roll;
        }
    }
    for (int64_t y = 0; y < lY; y++) g_hEv[y] = R->events[4 * y];
    unsigned char *base = nullptr;
    int rc = setup(dev, &base);
    if (rc) return rc;
    cudaStream_t st = 0;
    const int64_t slotCap = std::min<int64_t>(kPairCap, 3 * lY + 64);

    FItem fi;
    memset(&fi, 0, sizeof(fi));
    fi.nAnchors = (int32_t) R->nAnchors; fi.nTbCap = (int32_t) nTbCap;
    fi.pairCap = slotCap;
    fi.lX = (int32_t) lX; fi.lY = (int32_t) lY;
    fi.flags = (R->raggedLeft ? 1 : 0) | (R->raggedRight ? 2 : 0);
    fi.scale = R->scale; fi.shift = R->shift; fi.var = R->var;
    fi.invVar = 1 / R->var;
    fi.logInvVar = log((1 / R->var));

    FParams f;
    memset(&f, 0, sizeof(f));
    f.items = (const FItem *) (base + oItem); f.nItems = 1;
    f.counter = (int32_t *) (base + oCounter);
    f.order = (const int32_t *) (base + oOrder); f.nOrder = 1;
    f.evMean = (const double *) (base + oEv); f.kmerIdx = (const int32_t *) (base + oKmer);
    f.anchors = (const int32_t *) (base + oAnch);
    f.tb = (int32_t *) (base + oTb); f.geo = (FGeo *) (base + oGeo);
    f.ft = m->dev->ft;
    memcpy(f.trans, m->trans, sizeof(f.trans));
    f.eX = m->gapX[0];
    f.threshold = p->threshold;
    f.logThrGate = p->threshold > 0 ? log(p->threshold) - 1e-6 : -INFINITY;
    f.minDiags = (int32_t) p->minDiagsBetweenTraceBack; f.traceBack = (int32_t) p->traceBackDiagonals;
    f.expansion = (int32_t) p->diagonalExpansion;
    f.pairs = (sa_pair_t *) (base + oPairs); f.pairCount = (int64_t *) (base + oPairCount);
    f.wPlanes = 3; f.wAllFull = 0; f.wGeoAlways = 1;
    f.wNSW = kNSW; f.wRowsCap = (int32_t) kHdrCap; f.wRingDoubles = kRingDoubles;
    f.scratch = (double *) (base + oRing); f.wHdr = (int4 *) (base + oHdr);

    // what must start at zero: the work-queue counter, the item order (item 0) and the pair count (control block)
    CK(cudaMemsetAsync(base, 0, oTb, st));
    CK(cudaMemcpyAsync(base + oItem, &fi, sizeof(fi), cudaMemcpyHostToDevice, st));
    if (lY > 0) CK(cudaMemcpyAsync(base + oEv, g_hEv, sizeof(double) * lY, cudaMemcpyHostToDevice, st));
    if (lX > 0) CK(cudaMemcpyAsync(base + oKmer, g_hKmer, sizeof(int32_t) * lX, cudaMemcpyHostToDevice, st));
    if (R->nAnchors > 0) CK(cudaMemcpyAsync(base + oAnch, g_hAnch, sizeof(int32_t) * 2 * R->nAnchors, cudaMemcpyHostToDevice, st));
    sa_geometry_kernel<<<1, 32, 0, st>>>(f);
    CK(cudaGetLastError());
    FGeo geo;
    CK(cudaMemcpyAsync(&geo, base + oGeo, sizeof(geo), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    if (nDiag > 0) {
        const int64_t ringNeed = geo.maxBlockDoubles + 2 * 256 * (int64_t) 3 * geo.maxWidth + 64;
        if (geo.maxSpan + 24 > kNSW || geo.nTb > nTbCap || ringNeed > kRingDoubles || geo.maxRows + 260 > kHdrCap) {
            sa_set_error("sa_gpu_align_noalloc: the read's band does not fit the static scratch (3 diagonals span %d columns, limit "
                         "%d; forward rows of a traceback block need %lld doubles, limit %lld; %d rows, limit %lld): use "
                         "sa_gpu_align_batch", geo.maxSpan, kNSW - 24, (long long) ringNeed, (long long) kRingDoubles,
                         geo.maxRows, (long long) (kHdrCap - 260));
            return SA_ERR_INPUT;
        }
    }
    sa_wide_kernel<false, false, false, kNSW, kNT, 3><<<1, kNT, WideK<false, false, false>::smemBytes(kNSW, true), st>>>(f);
    CK(cudaGetLastError());
    int64_t n = 0;
    CK(cudaMemcpyAsync(&n, base + oPairCount, sizeof(n), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    *nPairsOut = n;
    if (n > slotCap || n > pairCap) {
        sa_set_error("sa_gpu_align_noalloc: the read produces %lld pairs; the caller's buffer holds %lld, the static slot %lld",
                     (long long) n, (long long) pairCap, (long long) slotCap);
        return SA_ERR_INPUT;
    }
    if (n > 0) {
        CK(cudaMemcpyAsync(pairsOut, base + oPairs, sizeof(sa_pair_t) * n, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        // the reference pops a region's list from its END (alignedPairCoordinateCorrectionFn, pairwiseAligner.c:2043-2050)
        std::reverse(pairsOut, pairsOut + n);
    }
    return SA_OK;
}
