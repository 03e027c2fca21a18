// sa_comm.cpp — the ONE collective of the path behind the C-ABI: the sum of the per-GPU expectation tables of an EM
// iteration (what the reference does by adding per-read expectation FILES in Python, src/signalalign/hiddenMarkovModel.py:
// 424-486), as an NCCL all-reduce over NVLink / NVSwitch, plus the gather of the HDP assignment lists
// (impl/continuousHmm.c:510-514, written out by :571-628).  No PyTorch on this path: the caller only has to carry the
// 128-byte ncclUniqueId from rank 0 to the other ranks (any launcher can: MPI, a file, torch's store).
//
// libnccl is loaded at run time (dlopen) so that libsagpu.so itself has no link-time dependency on it: single-GPU users
// never touch it; when the process already holds a libnccl.so.2 (e.g. torch's), the loader hands back that one.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <dlfcn.h>
#include <cuda_runtime.h>
#include <nccl.h>
#include "sa_internal.h"

namespace {
struct NcclApi {
    void *lib = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
};
NcclApi g_nccl;

int loadNccl() {
    if (g_nccl.lib) return SA_OK;
    const char *names[] = {getenv("SA_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
    void *h = nullptr;
    for (const char *n : names)
        if (n && n[0] && (h = dlopen(n, RTLD_NOW | RTLD_GLOBAL))) break;
    if (!h) {
        sa_set_error("sa_gpu_comm: cannot load libnccl.so.2 (%s); set SA_NCCL_LIB", dlerror());
        return SA_ERR_CUDA;
    }
#define SA_SYM(field, name)                                              \
    *(void **) (&g_nccl.field) = dlsym(h, name);                         \
    if (!g_nccl.field) {                                                 \
        sa_set_error("sa_gpu_comm: libnccl has no symbol %s", name);     \
        return SA_ERR_CUDA;                                              \
    }
    SA_SYM(GetUniqueId, "ncclGetUniqueId")
    SA_SYM(CommInitRank, "ncclCommInitRank")
    SA_SYM(CommDestroy, "ncclCommDestroy")
    SA_SYM(AllReduce, "ncclAllReduce")
    SA_SYM(AllGather, "ncclAllGather")
    SA_SYM(GetErrorString, "ncclGetErrorString")
#undef SA_SYM
    g_nccl.lib = h;
    return SA_OK;
}
}  // namespace

struct sa_comm {
    ncclComm_t comm = nullptr;
    bool owned = false;
    int nRanks = 1, rank = 0;
    cudaStream_t stream = nullptr;
};

#define NK(call)                                                                                          \
    do {                                                                                                  \
        ncclResult_t r_ = (call);                                                                         \
        if (r_ != ncclSuccess) {                                                                          \
            sa_set_error("NCCL error %s at %s:%d: %s", #call, __FILE__, __LINE__, g_nccl.GetErrorString(r_)); \
            return SA_ERR_CUDA;                                                                           \
        }                                                                                                 \
    } while (0)
#define CKC(call)                                                                                         \
    do {                                                                                                  \
        cudaError_t e_ = (call);                                                                          \
        if (e_ != cudaSuccess) {                                                                          \
            sa_set_error("CUDA error %s at %s:%d: %s", #call, __FILE__, __LINE__, cudaGetErrorString(e_)); \
            return SA_ERR_CUDA;                                                                           \
        }                                                                                                 \
    } while (0)

extern "C" int sa_gpu_comm_unique_id(void *id128) {
    static_assert(sizeof(ncclUniqueId) == SA_COMM_UNIQUE_ID_BYTES, "ncclUniqueId is 128 bytes");
    int rc = loadNccl();
    if (rc) return rc;
    ncclUniqueId id;
    NK(g_nccl.GetUniqueId(&id));
    memcpy(id128, &id, sizeof(id));
    return SA_OK;
}

extern "C" int sa_gpu_comm_init(const void *id128, int nRanks, int rank, sa_comm_t **out) {
    *out = nullptr;
    int rc = loadNccl();
    if (rc) return rc;
    ncclUniqueId id;
    memcpy(&id, id128, sizeof(id));
    sa_comm *c = new sa_comm;
    c->nRanks = nRanks; c->rank = rank; c->owned = true;
    ncclResult_t r = g_nccl.CommInitRank(&c->comm, nRanks, id, rank);      // on the device sa_gpu_init selected
    if (r != ncclSuccess) {
        sa_set_error("sa_gpu_comm_init: ncclCommInitRank: %s", g_nccl.GetErrorString(r));
        delete c;
        return SA_ERR_CUDA;
    }
    if (cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess) {
        sa_set_error("sa_gpu_comm_init: cannot create a stream");
        g_nccl.CommDestroy(c->comm);
        delete c;
        return SA_ERR_CUDA;
    }
    *out = c;
    return SA_OK;
}

extern "C" int sa_gpu_comm_from_nccl(void *ncclComm, int nRanks, int rank, sa_comm_t **out) {
    *out = nullptr;
    int rc = loadNccl();
    if (rc) return rc;
    sa_comm *c = new sa_comm;
    c->comm = (ncclComm_t) ncclComm; c->nRanks = nRanks; c->rank = rank; c->owned = false;
    if (cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess) {
        sa_set_error("sa_gpu_comm_from_nccl: cannot create a stream");
        delete c;
        return SA_ERR_CUDA;
    }
    *out = c;
    return SA_OK;
}

extern "C" void sa_gpu_comm_free(sa_comm_t *c) {
    if (!c) return;
    if (c->stream) { cudaStreamSynchronize(c->stream); cudaStreamDestroy(c->stream); }
    if (c->owned && c->comm) g_nccl.CommDestroy(c->comm);
    delete c;
}

// in place on a HOST table of n doubles: H2D, ncclAllReduce(sum, double) over NVLink, D2H.  Every rank must pass the
// same n (10 for the transitions + likelihood the C path accumulates; 10 + 3 K for the full file layout).
extern "C" int sa_gpu_allreduce_expectations(sa_comm_t *c, double *table, int64_t n) {
    if (!c || n < 0) { sa_set_error("sa_gpu_allreduce_expectations: bad arguments"); return SA_ERR_INPUT; }
    if (n == 0) return SA_OK;
    double *d = nullptr;
    CKC(cudaMalloc(&d, sizeof(double) * n));
    CKC(cudaMemcpyAsync(d, table, sizeof(double) * n, cudaMemcpyHostToDevice, c->stream));
    NK(g_nccl.AllReduce(d, d, (size_t) n, ncclDouble, ncclSum, c->comm, c->stream));
    CKC(cudaMemcpyAsync(table, d, sizeof(double) * n, cudaMemcpyDeviceToHost, c->stream));
    CKC(cudaStreamSynchronize(c->stream));
    cudaFree(d);
    return SA_OK;
}

// HDP assignment lists of all ranks, concatenated in rank order on every rank (counts all-gathered first, then the
// lists padded to the longest).  *allKmerPos / *allEvent are malloc'ed; countsPerRank has nRanks entries.
extern "C" int sa_gpu_allgather_assignments(sa_comm_t *c, const int32_t *kmerPos, const int32_t *event, int64_t n,
                                            int32_t **allKmerPos, int32_t **allEvent, int64_t *countsPerRank) {
    if (!c || n < 0) { sa_set_error("sa_gpu_allgather_assignments: bad arguments"); return SA_ERR_INPUT; }
    const int R = c->nRanks;
    int64_t *dCounts = nullptr;
    CKC(cudaMalloc(&dCounts, sizeof(int64_t) * (R + 1)));
    CKC(cudaMemcpyAsync(dCounts + R, &n, sizeof(int64_t), cudaMemcpyHostToDevice, c->stream));
    NK(g_nccl.AllGather(dCounts + R, dCounts, 1, ncclInt64, c->comm, c->stream));
    CKC(cudaMemcpyAsync(countsPerRank, dCounts, sizeof(int64_t) * R, cudaMemcpyDeviceToHost, c->stream));
    CKC(cudaStreamSynchronize(c->stream));
    cudaFree(dCounts);
    int64_t mx = 1, total = 0;
    for (int r = 0; r < R; r++) { mx = countsPerRank[r] > mx ? countsPerRank[r] : mx; total += countsPerRank[r]; }
    int32_t *dSend = nullptr, *dRecv = nullptr;
    CKC(cudaMalloc(&dSend, sizeof(int32_t) * 2 * mx));
    CKC(cudaMalloc(&dRecv, sizeof(int32_t) * 2 * mx * R));
    CKC(cudaMemsetAsync(dSend, 0, sizeof(int32_t) * 2 * mx, c->stream));
    if (n > 0) {
        CKC(cudaMemcpyAsync(dSend, kmerPos, sizeof(int32_t) * n, cudaMemcpyHostToDevice, c->stream));
        CKC(cudaMemcpyAsync(dSend + mx, event, sizeof(int32_t) * n, cudaMemcpyHostToDevice, c->stream));
    }
    NK(g_nccl.AllGather(dSend, dRecv, (size_t) (2 * mx), ncclInt32, c->comm, c->stream));
    std::vector<int32_t> host((size_t) (2 * mx * R));
    CKC(cudaMemcpyAsync(host.data(), dRecv, sizeof(int32_t) * host.size(), cudaMemcpyDeviceToHost, c->stream));
    CKC(cudaStreamSynchronize(c->stream));
    cudaFree(dSend); cudaFree(dRecv);
    *allKmerPos = (int32_t *) malloc(sizeof(int32_t) * (size_t) (total > 0 ? total : 1));
    *allEvent = (int32_t *) malloc(sizeof(int32_t) * (size_t) (total > 0 ? total : 1));
    int64_t w = 0;
    for (int r = 0; r < R; r++)
        for (int64_t i = 0; i < countsPerRank[r]; i++) {
            (*allKmerPos)[w] = host[(size_t) (2 * mx * r + i)];
            (*allEvent)[w] = host[(size_t) (2 * mx * r + mx + i)];
            w++;
        }
    return SA_OK;
}
