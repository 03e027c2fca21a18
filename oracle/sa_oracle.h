/*
 * sa_oracle.h — TEST INFRASTRUCTURE ONLY (CPU restatement of the reference hot path).
 *
 * A from-scratch, flat-array C restatement of signalAlign's banded forward/backward posterior-match DP
 * (/root/reference/impl/pairwiseAligner.c:1450-1590 and callees) used as the parity checker for the CUDA
 * path.  It is pinned against the compiled reference itself (oracle/_ref/libsaref.so) and against the
 * reference's golden pair counts (tests/test_oracle_vs_reference.py, tests/golden/).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may load this library.
 */
#ifndef SA_ORACLE_H
#define SA_ORACLE_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* emission kinds (which reference function is being restated) */
enum {
    SAO_EMIT_GAUSS_MEANONLY_DESCALE = 0, /* stateMachine.c:557  (production, signalMachine.c:347) */
    SAO_EMIT_HDP = 1,                    /* stateMachine.c:527  (production with -d / .nhdp) */
    SAO_EMIT_GAUSS_NOISE_SCALED = 2,     /* stateMachine.c:659  strawManGetKmerEventMatchProb (tests) */
    SAO_EMIT_GAUSS_NOISE_DESCALE = 3     /* stateMachine.c:607  ...WithDescaling (tests) */
};

/* order of the 9 log-transitions, same as struct _StateMachine3 (inc/stateMachine.h:154-162) */
enum {
    SAO_T_MATCH_CONTINUE = 0, SAO_T_MATCH_FROM_GAP_X, SAO_T_MATCH_FROM_GAP_Y, SAO_T_GAP_OPEN_X, SAO_T_GAP_OPEN_Y,
    SAO_T_GAP_EXTEND_X, SAO_T_GAP_EXTEND_Y, SAO_T_GAP_SWITCH_TO_X, SAO_T_GAP_SWITCH_TO_Y
};

typedef struct {
    int32_t kind;
    int32_t alphabetSize;
    int32_t kmerLength;
    int32_t pad;
    int64_t K;                /* alphabetSize ^ kmerLength */
    char alphabet[64];        /* sorted, NUL terminated */
    double trans[9];          /* log transitions */
    const double *matchTable; /* K*5: level_mean level_sd noise_mean noise_sd noise_lambda */
    const double *gapYTable;  /* K*5 */
    const double *gapXTable;  /* K (all log 0.1 in practice) */
    double scale, shift, var; /* per-read */
    /* HDP (kind == SAO_EMIT_HDP): uniform grid + per-kmer row of posterior predictive / slopes */
    int64_t gridLength;
    const double *grid;        /* gridLength */
    const int32_t *kmerToRow;  /* K -> row (parent walk already resolved), -1 = none */
    const double *postPred;    /* rows * gridLength */
    const double *slopes;      /* rows * gridLength */
} sao_model;

typedef struct {
    double threshold;
    int64_t minDiagsBetweenTraceBack;
    int64_t traceBackDiagonals;
    int64_t diagonalExpansion;
    int64_t splitMatrixBiggerThanThis;
} sao_params;

/* ambiguity table: for char c, ambig[c] is a NUL-terminated list of replacement chars ("" = not ambiguous) */
typedef struct { char opts[256][16]; } sao_ambig;
void sao_ambig_default(sao_ambig *a);                       /* pairwiseAligner.c:32-65 */
int sao_ambig_from_file(sao_ambig *a, const char *path);    /* pairwiseAligner.c:68-92 */

double sao_logAdd(double x, double y);                      /* pairwiseAligner.c:301-318 */
int64_t sao_kmer_id(const char *kmer, const char *alphabet, int64_t alphabetSize, int64_t k); /* nanopore_hdp.c:377-410; -1 if bad char */
double sao_emission(const sao_model *m, int64_t kmerIdx, double eventMean, double eventNoise, int match);

/* band_construct (pairwiseAligner.c:195-246): out[3*i] = xay, xmyL, xmyR for i in [0, lX+lY] */
void sao_band(const int64_t *anchors, int64_t nAnchors, int64_t lX, int64_t lY, int64_t expansion, int64_t *out);
/* getSplitPoints (pairwiseAligner.c:1910-1937): out[4*i] = x1,y1,x2,y2; returns count */
int64_t sao_split_points(const int64_t *anchors, int64_t nAnchors, int64_t lX, int64_t lY, int64_t splitBigger,
                         int raggedL, int raggedR, int64_t *out, int64_t cap);
/* number of paths + concrete kmer indices of the k-window starting at `kmer` (hdCell_construct2, :723-801).
 * returns nPaths, or -1 if a character is outside the alphabet */
int64_t sao_paths(const char *kmer, const sao_model *m, const sao_ambig *a, int32_t *idxOut, int64_t cap);

/* getAlignedPairsUsingAnchors (pairwiseAligner.c:2052-2080).
 * out: cap*4 int64 (floor(p*1e7), x, y, kmerIdx) in the reference's list order. returns #pairs (may exceed cap) or <0 on error.
 * stats (optional, 4 doubles): band cells, path-cells, #tracebacks, last total log-probability */
int64_t sao_align(const sao_model *m, const char *target, int64_t lX, const double *events, int64_t lY,
                  const int64_t *anchors, int64_t nAnchors, const sao_params *p, int raggedL, int raggedR,
                  const sao_ambig *ambig, int64_t *out, int64_t cap, double *stats);

/* getExpectationsUsingAnchors (pairwiseAligner.c:2164-2186) with diagonalCalculation_Expectations (:1423).
 * transitions9 and *likelihood are ACCUMULATED into (caller pre-loads pseudocounts).
 * HDP: assignments (kmer offset in target, event index) appended when to==match && p>=threshold. */
int64_t sao_expectations(const sao_model *m, const char *target, int64_t lX, const double *events, int64_t lY,
                         const int64_t *anchors, int64_t nAnchors, const sao_params *p, int raggedL, int raggedR,
                         const sao_ambig *ambig, double *transitions9, double *likelihood,
                         int64_t *assignKmerOff, int64_t *assignEvent, int64_t cap);

/* getAlignedPairsWithoutBanding (pairwiseAligner.c:2104-2162) */
int64_t sao_align_noband(const sao_model *m, const char *target, int64_t lX, const double *events, int64_t lY,
                         const sao_params *p, int raggedL, int raggedR, const sao_ambig *ambig,
                         int64_t *out, int64_t cap);

#ifdef __cplusplus
}
#endif
/* signalUtils_estimateNanoporeParams (impl/signalMachineUtils.c:186-225): see sa_oracle.c */
int64_t sao_estimate_params(const double *model, const char *alphabet, int64_t alphabetSize, int64_t k, const char *read,
                            int64_t readLength, const int64_t *eventMap, double *events, int64_t nEvents, int adjustEvents,
                            double *out7);

#endif
