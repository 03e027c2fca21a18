// sa_internal.h — shared host/device declarations of libsagpu (not part of the C-ABI).
#pragma once
#include <cstdint>
#include <string>
#include <vector>
#include "signalalign_b200.h"

// order of the 9 log-transitions == struct _StateMachine3 (reference inc/stateMachine.h:154-162)
enum {
    T_MATCH_CONTINUE = 0, T_MATCH_FROM_GAP_X, T_MATCH_FROM_GAP_Y, T_GAP_OPEN_X, T_GAP_OPEN_Y,
    T_GAP_EXTEND_X, T_GAP_EXTEND_Y, T_GAP_SWITCH_TO_X, T_GAP_SWITCH_TO_Y
};

// derived per-kmer emission constants uploaded to the device: 6 doubles for the match table, 6 for gap-Y
//   [0] level mean  [1] level sd  [2] log(level sd) (host libm, so the device needs no log for Gaussian kinds)
//   [3] noise mean  [4] noise lambda  [5] log(noise lambda)
#define SA_KT_STRIDE 12

struct sa_ambig {
    char opts[256][16];
};

struct DeviceModel;  // sa_gpu.cu

struct sa_model {
    int kind = SA_EMIT_GAUSS_MEANONLY_DESCALE;
    int A = 0;  // alphabet size
    int k = 0;
    int64_t K = 0;
    std::string alphabet;  // sorted
    double trans[9];
    std::vector<double> match, gapY, gapX;  // K*5, K*5, K
    // HDP
    int64_t G = 0;
    std::vector<double> grid;
    std::vector<int32_t> kmerToRow;   // K: row of the first observed DP on the parent walk
    std::vector<int32_t> ownRow;      // K: row of the kmer's own DP if observed, else -1
    int64_t nRows = 0;
    std::vector<double> postPred, slopes;  // nRows * G
    // device copy (built lazily, dropped when tables change)
    mutable DeviceModel *dev = nullptr;
    mutable int devDevice = -1;
};

void sa_set_error(const char *fmt, ...);

// ---- host-side integer preparation (sa_host.cpp) ---------------------------------------------------------
struct Diag {
    int64_t xay, xmyL, xmyR;
};
// band_construct, reference impl/pairwiseAligner.c:195-246
void sa_build_band(const int64_t *anchors, int64_t nA, int64_t lX, int64_t lY, int64_t e, std::vector<Diag> &out);
// getSplitPoints, reference impl/pairwiseAligner.c:1886-1937; out: x1,y1,x2,y2 per region
void sa_split_points(const int64_t *anchors, int64_t nA, int64_t lX, int64_t lY, int64_t bigger, bool rl, bool rr,
                     std::vector<int64_t> &out);
// hdCell_construct2 path enumeration (reference impl/pairwiseAligner.c:723-801) -> concrete kmer indices.
// returns false if a character is outside the alphabet
bool sa_kmer_paths(const char *kmer, const sa_model &m, const sa_ambig &a, std::vector<int32_t> &out);
void sa_ambig_default_table(sa_ambig &a);

// one sub-region of one read after getSplitPoints: the unit the device works on
struct WorkItem {
    int64_t read;            // index into the caller's reads
    int64_t x1, y1;          // region offset inside the read (coordinate correction, pairwiseAligner.c:2043)
    int64_t lX, lY;
    bool raggedL, raggedR;
    std::vector<int64_t> anchors;  // region-relative (x, y)
};
