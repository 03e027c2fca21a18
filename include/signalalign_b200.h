/*
 * signalalign_b200.h — C-ABI of the B200-native signalAlign hot path (libsagpu.so).
 *
 * One data-parallel path is exported: the anchor-banded forward/backward posterior-match DP of
 * signalAlign's 3-state pair-HMM with its Gaussian-table / HDP emissions.  Every entry point is
 * `extern "C"`, takes plain pointers and sizes, and names the reference interface it replaces
 * (paths relative to the signalAlign source tree).
 *
 * There is NO CPU fallback: every compute entry point fails (non-zero return + sa_last_error())
 * when no CUDA device is usable.
 */
#ifndef SIGNALALIGN_B200_H
#define SIGNALALIGN_B200_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SA_OK 0
#define SA_ERR_CUDA 1      /* CUDA runtime failure (message in sa_last_error) */
#define SA_ERR_INPUT 2     /* malformed input (bad kmer character, anchors out of range, ...) — the reference st_errAbort()s */
#define SA_ERR_IO 3        /* model file unreadable */

/* ---------------------------------------------------------------------------------------------- models
 * Replaces: stateMachine3_loadFromFile (impl/stateMachine.c:1440-1538), deserialize_nhdp
 * (impl/nanopore_hdp.c:1088-1115) -> deserialize_hdp (impl/hdp.c:3052-3322),
 * stateMachine3_setModelToHdpExpectedValues (impl/stateMachine.c:1275-1304).
 */
typedef struct sa_model sa_model_t;

/* which reference emission function the model evaluates (inc/stateMachine.h:303-309) */
#define SA_EMIT_GAUSS_MEANONLY_DESCALE 0 /* emissions_signal_strawManGetKmerEventMatchProbWithDescaling_MeanOnly (signalMachine production) */
#define SA_EMIT_HDP 1                    /* emissions_signal_getHdpKmerDensity (production with an .nhdp) */
#define SA_EMIT_GAUSS_NOISE_SCALED 2     /* emissions_signal_strawManGetKmerEventMatchProb (getStateMachine3) */
#define SA_EMIT_GAUSS_NOISE_DESCALE 3    /* emissions_signal_strawManGetKmerEventMatchProbWithDescaling (getStateMachine3_descaled) */

/* .model (3 lines) [+ .nhdp]; nhdpFile may be NULL.  With an .nhdp the kind is forced to SA_EMIT_HDP. */
sa_model_t *sa_model_load(const char *modelFile, const char *nhdpFile, int emissionKind);

/* From the raw tables of an existing `StateMachine` (inc/stateMachine.h:85-118):
 *   alphabet (any order; sorted internally as sequence_prepareAlphabet, impl/pairwiseAligner.c:366),
 *   logTransitions[9] in `struct _StateMachine3` order (inc/stateMachine.h:154-162),
 *   matchTable / gapYTable = EMISSION_MATCH_MATRIX / EMISSION_GAP_Y_MATRIX (K*5 doubles),
 *   gapXTable = EMISSION_GAP_X_PROBS (K doubles, may be NULL = log 0.1). */
sa_model_t *sa_model_from_tables(const char *alphabet, int32_t kmerLength, const double *logTransitions,
                                 const double *matchTable, const double *gapYTable, const double *gapXTable,
                                 int emissionKind);
/* attach HDP spline tables.  kmerToRow: K entries, the parent walk of dir_proc_density (impl/hdp.c:2588-2612) already
 * resolved; ownRow: K entries, the row of the kmer's OWN Dirichlet process when it is observed, else -1 (decides which
 * kmers stateMachine3_setModelToHdpExpectedValues replaces); NULL = every kmer's own DP is observed */
int sa_model_set_hdp(sa_model_t *m, int64_t gridLength, const double *grid, const int32_t *kmerToRow,
                     const int32_t *ownRow, int64_t nRows, const double *postPred, const double *slopes);
/* attach the HDP of an .nhdp file (what deserialize_nhdp reads, impl/nanopore_hdp.c:1088-1115) to a model built with
 * sa_model_from_tables; the emission kind becomes SA_EMIT_HDP.  The drop-in shim uses it with the file the reference's
 * own serialize_nhdp (impl/nanopore_hdp.c:1076-1086) writes from a live StateMachine3_HDP (%.17lg: exact round trip) */
int sa_model_attach_nhdp(sa_model_t *m, const char *nhdpFile);
int sa_model_set_hdp_expected_values(sa_model_t *m);
/* emissions_signal_scaleNoise (impl/stateMachine.c:721-741) / emissions_signal_scaleModel (:743-779) */
int sa_model_scale_noise(sa_model_t *m, double scale_sd, double var_sd);
int sa_model_scale_model(sa_model_t *m, double scale, double shift, double var, double scale_sd, double var_sd);
void sa_model_free(sa_model_t *m);

/* introspection (tests, drop-in shim) */
int64_t sa_model_num_kmers(const sa_model_t *m);
int32_t sa_model_kmer_length(const sa_model_t *m);
const char *sa_model_alphabet(const sa_model_t *m);               /* sorted */
void sa_model_get_tables(const sa_model_t *m, double *logTransitions9, double *matchTable, double *gapYTable);
/* host evaluation of the .nhdp spline density the device uses (get_nanopore_kmer_density, nanopore_hdp.c:420) */
double sa_model_hdp_density(const sa_model_t *m, int64_t kmerIdx, double x);
/* the HDP spline tables as loaded from the .nhdp: dims[0] = grid length G, dims[1] = number of observed DPs (rows);
 * any output pointer may be NULL.  grid: G, kmerToRow / ownRow: K, postPred / slopes: rows * G */
void sa_model_hdp_dims(const sa_model_t *m, int64_t *dims2);
void sa_model_get_hdp(const sa_model_t *m, double *grid, int32_t *kmerToRow, int32_t *ownRow, double *postPred,
                      double *slopes);

/* ---------------------------------------------------------------------------------------------- inputs
 * One sa_read_t = the arguments of one getAlignedPairsUsingAnchors call
 * (inc/pairwiseAligner.h:406-414; call site impl/signalMachine.c:436-437).
 */
typedef struct {
    const char *target;      /* nucleotides, lX + k - 1 chars (need not be NUL terminated); Sequence sX->elements */
    int64_t targetLength;    /* number of nucleotides */
    const double *events;    /* nEvents * 4 doubles: mean, sd, duration, start (NB_EVENT_PARAMS, inc/nanopore.h:5) */
    int64_t nEvents;         /* lY */
    const int64_t *anchors;  /* nAnchors (x, y) pairs, 0-based, strictly increasing in both (stList of stIntTuple) */
    int64_t nAnchors;
    double scale, shift, var; /* sM->scale / shift / var (inc/stateMachine.h:92-94) */
    int32_t raggedLeft, raggedRight;
} sa_read_t;

/* PairwiseAlignmentParameters fields the path reads (inc/pairwiseAligner.h:143-154) */
typedef struct {
    double threshold;
    int64_t minDiagsBetweenTraceBack;
    int64_t traceBackDiagonals;
    int64_t diagonalExpansion;
    int64_t splitMatrixBiggerThanThis;
} sa_params_t;
/* pairwiseAlignmentBandingParameters_construct defaults (impl/pairwiseAligner.c:2022-2037) */
void sa_params_default(sa_params_t *p);

/* ambiguity table: create_ambig_bases / create_ambig_bases2 (impl/pairwiseAligner.c:32-92); file may be NULL */
typedef struct sa_ambig sa_ambig_t;
sa_ambig_t *sa_ambig_load(const char *file);
void sa_ambig_free(sa_ambig_t *a);
/* the same table built entry by entry from a caller's own code -> bases map (Sequence.ambigBases, the stHash that
 * create_ambig_bases2 returns and hdCell_construct2 looks single characters up in, impl/pairwiseAligner.c:755):
 * sa_ambig_create() is EMPTY (no code is ambiguous), sa_ambig_set adds / replaces one code (at most 15 bases) */
sa_ambig_t *sa_ambig_create(void);
int sa_ambig_set(sa_ambig_t *a, char code, const char *bases);

/* ---------------------------------------------------------------------------------------------- outputs */
/* one aligned pair = stIntTuple4(floor(p*1e7), x, y, kmer*) (impl/pairwiseAligner.c:1405-1408);
 * the kmer pointer becomes the kmer's index (sa_kmer_string turns it back into text) */
typedef struct {
    int32_t p;        /* floor(posterior * 1e7) */
    int32_t x;        /* reference kmer index (0-based) */
    int32_t y;        /* event index (0-based) */
    int32_t kmerIdx;  /* path kmer */
} sa_pair_t;

typedef struct {
    int64_t nReads;
    int64_t *pairOffsets;    /* nReads + 1 */
    sa_pair_t *pairs;        /* pairOffsets[nReads] pairs, per read in the reference's list order */
    /* expectations (getExpectationsUsingAnchors): per read 9 transition sums (from*3+to) + likelihood */
    double *transitions;     /* nReads * 9, NULL for alignment runs */
    double *likelihood;      /* nReads */
    int64_t *assignOffsets;  /* nReads + 1 (HDP assignments), NULL unless HDP expectations */
    int32_t *assignKmerPos;  /* offset of the assigned kmer in the read's target */
    int32_t *assignEvent;    /* event index */
    /* accounting (same integer band logic as the reference): */
    int64_t totalBandCells;  /* sum over diagonals of band width */
    int64_t totalPathCells;  /* band cells weighted by the number of paths */
    int64_t totalEvents;
} sa_result_t;
void sa_result_free(sa_result_t *r);
void sa_kmer_string(const sa_model_t *m, int64_t kmerIdx, char *out /* k+1 bytes */);

/* ---------------------------------------------------------------------------------------------- device */
int sa_gpu_init(int device);                 /* select device for this process/thread; SA_ERR_CUDA if none */
int sa_gpu_shutdown(void);                   /* wait for the device; give cached pinned / pool memory back */
const char *sa_last_error(void);
/* host threads the packing passes of this process may use (default: OpenMP's, i.e. OMP_NUM_THREADS or all cores) */
void sa_set_host_threads(int n);
int sa_gpu_device_count(void);

/* Arithmetic of the wide band kernel (bands wider than 254 columns) on alignment runs with the production Gaussian
 * emission.  SA_PRECISION_FP64 (default): the reference's double-precision arithmetic, operation for operation (posteriors
 * within 2e-6 of the reference).  SA_PRECISION_FP32: DP state, stored rows and every logAdd in single precision, each value
 * relative to a per-diagonal integer offset kept in double (sa_wide32.cuh): coordinates and kmers still bit-exact,
 * posteriors within the contract's 1e-4 (measured: 3e-5), about twice the throughput.  The reference has one precision
 * (double, impl/pairwiseAligner.c:298-318); BASELINE's north_star asks for the FP32 variant.  Process-wide; the environment
 * variable SA_PRECISION=fp32 sets the default.  Expectations, the HDP emission and narrow bands always run in FP64. */
#define SA_PRECISION_FP64 0
#define SA_PRECISION_FP32 1
int sa_gpu_set_precision(int precision);
int sa_gpu_get_precision(void);

/* One-shot drop-in for a batch of getAlignedPairsUsingAnchors calls (host buffers in, host result out).
 * Large batches run as a pipeline of chunks: host packing, H2D, the band kernels and compaction + D2H of consecutive
 * chunks overlap.  result->pairs is page-locked memory of libsagpu's pool: release it with sa_result_free only. */
int sa_gpu_align_batch(const sa_model_t *m, const sa_read_t *reads, int64_t nReads, const sa_params_t *p,
                       const sa_ambig_t *ambig, sa_result_t **out);
/* accounting of the last sa_gpu_align_batch of this process: bytes copied each way and kernels launched */
void sa_last_call_stats(int64_t *h2dBytes, int64_t *d2hBytes, int64_t *launches);
/* ... and of getExpectationsUsingAnchors (inc/pairwiseAligner.h:417-429; call site signalMachine.c:479).
 * Results are the per-read expectation sums WITHOUT pseudocounts. */
int sa_gpu_expectations_batch(const sa_model_t *m, const sa_read_t *reads, int64_t nReads, const sa_params_t *p,
                              const sa_ambig_t *ambig, sa_result_t **out);

/* Per-read scaling parameters for a batch of strands.
 * Replaces: signalUtils_estimateNanoporeParams (impl/signalMachineUtils.c:186-225; call sites impl/signalMachine.c:778, 855)
 * with assignmentFunction = signalUtils_templateOneDAssignmentsFromRead / ...complement... and the drift adjustment
 * nanopore_adjust{Template,Complement}EventsForDrift: kmer <-> event assignments from the strand's event map
 * (impl/nanopore.c:601-631), weighted least squares for shift / scale / drift and var (:756-826), for shift_sd / scale_sd /
 * var_sd (:878-946), and `mean -= start_time * drift` on every event (:633-638) when adjustEvents != 0 (the mean column of
 * the caller's events array is updated in place).  Results are bit-identical to the reference's (its summation order is
 * kept).  The model's match table is used as loaded (call it before emissions_signal_scaleNoise, like the reference). */
typedef struct {
    const char *read;              /* base-called strand: NanoporeRead.templateRead / complementRead */
    int64_t readLength;
    const int64_t *strandEventMap; /* readLength entries: NanoporeRead.templateStrandEventMap / complementStrandEventMap */
    double *events;                /* nEvents * 4 doubles: mean, sd, duration, start (NanoporeRead.templateEvents) */
    int64_t nEvents;
} sa_strand_t;
typedef struct {                   /* NanoporeReadAdjustmentParameters (inc/nanopore.h:13-21) + the number of assignments */
    double scale, shift, var, shift_sd, scale_sd, var_sd, drift;
    int64_t nAssignments;
} sa_scaling_t;
int sa_gpu_estimate_params_batch(const sa_model_t *m, sa_strand_t *strands, int64_t nStrands, int adjustEvents,
                                 sa_scaling_t *out);

/* SURVEY.md 8f N3 — variant calls of aligned pairs and their per-position marginals, on the device.
 * Replaces writePosteriorProbsVC (impl/signalMachine.c:161-232: one TSV row per ambiguity character X of the reference
 * kmer of every aligned pair) followed by MarginalizeVariants.get_data (src/signalalign/variantCaller.py:37-89: per
 * reference position the posterior mass of every called base normalised over the bases, per read the mean over its
 * positions), the TSV round trip between the two included: with roundLikeTsv != 0 every posterior is first rounded to
 * the six decimals a "%f" field carries.  One entry per (read, strand); variants = the bases to track, e.g. "CE"
 * (columns come out in sorted order, as the reference sorts them).  A called base outside `variants` is an error
 * (the reference raises).  Positions come out ascending; positions no pair reports on are absent, as in the reference. */
typedef struct {
    const sa_pair_t *pairs;        /* the read's aligned pairs as sa_result_t holds them (kmerIdx = the path kmer) */
    int64_t nPairs;
    const char *target;            /* the sequence the strand was aligned to, X at the ambiguous positions */
    int64_t targetLength;
    int32_t strandIsComplement;    /* Strand: template 0, complement 1 */
    int32_t forward;               /* the `forward` flag signalMachine passes to its writers */
    int64_t referenceSequenceOffset;
} sa_variant_read_t;
typedef struct {
    int64_t nReads;
    int32_t nVariants;
    char variants[32];             /* sorted, NUL-terminated: the column order of probs / perRead */
    int64_t *positionOffsets;      /* nReads + 1: read i owns positions[positionOffsets[i] .. positionOffsets[i + 1]) */
    int64_t *positions;            /* reference positions, ascending within a read */
    double *probs;                 /* [total positions][nVariants] */
    double *perRead;               /* [nReads][nVariants]: mean of the read's rows of probs (zeros for a read without positions) */
} sa_marginals_t;
int sa_gpu_marginalize_variants(const sa_model_t *m, const sa_variant_read_t *reads, int64_t nReads, const char *variants,
                                int roundLikeTsv, sa_marginals_t **out);
void sa_marginals_free(sa_marginals_t *r);

/* SURVEY.md 8f N4 — adaptive banded event -> kmer alignment, on the device.
 * Replaces adaptive_banded_simple_event_align2 (impl/eventAligner.c:899-1235; band of 100 cells moved by Suzuki's rule,
 * Viterbi over step / stay / skip with the ..._MeanOnly emission, trim states, backtrack, QC) and, with useMom != 0, the
 * scalings of estimate_scalings_using_mom (:792-842) — what load_from_raw2 (:1240-1301) runs between event detection and
 * the embedding of the event table.  The aligned (kmer, event) pairs are bit-identical to the reference's; a read its QC
 * rejects (average log emission < -5.2, not spanning the kmers, a gap > 50, more than 5 events per kmer) comes back with
 * zero pairs and failed = 1, as the reference returns an empty list.  Models with the Gaussian match table only. */
typedef struct {
    const char *sequence;          /* nucleotides; U is read as T */
    int64_t sequenceLength;
    const float *eventMean;        /* event_t.mean of the detected events (scrappie_structures.h:8-15), in signal order */
    int64_t nEvents;
    int32_t rna;                   /* kmers reversed, as build_kmer_list does for RNA (eventAligner.c:774-790) */
    int32_t useMom;                /* scale / shift by the method of moments (var = 1); else the three below */
    double scale, shift, var;
} sa_event_read_t;
typedef struct {
    int64_t nReads;
    int64_t *pairOffsets;          /* nReads + 1 */
    int32_t *kmerIdx, *eventIdx;   /* aligned pairs in the reference's order (ascending) */
    double *scale, *shift, *var;   /* per read, as the alignment used them */
    double *avgLogEmission;        /* per read QC figures (also for failed reads) */
    int32_t *maxGap, *spanned, *failed, *nAlignedBeforeQc;
} sa_event_alignment_t;
int sa_gpu_event_align_batch(const sa_model_t *m, const sa_event_read_t *reads, int64_t nReads, sa_event_alignment_t **out);
void sa_event_alignment_free(sa_event_alignment_t *r);

/* ONE read, NO device allocation.  This entry point has no counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Same result as sa_gpu_align_batch on that read (the wide band kernel, 512-slot window):
 * pairs in the reference's list order into the caller's HOST buffer pairsOut[pairCap]; *nPairs = how many the read
 * produces (SA_ERR_INPUT when that exceeds pairCap: nothing is written).  All device memory of the call — staged inputs,
 * work-item record, traceback list, row headers, pair slots and the kernel's scratch (the ring of forward rows) — is one
 * static device array of SA_NOALLOC_SCRATCH_BYTES that every GPU loading the library holds until exit; the call makes no
 * cudaMalloc / cudaFree / pool allocation.  Limits (SA_ERR_INPUT with a message beyond them): SA_NOALLOC_MAX_KMERS reference
 * kmers and anchors, SA_NOALLOC_MAX_EVENTS events, three consecutive diagonals spanning at most 488 columns, the forward rows
 * of one traceback block within the array (about 1.7 M doubles: any -x 100 read with dense or CLI-trimmed anchors), no
 * ambiguity codes, no anchor gap that getSplitPoints would cut at, production Gaussian emission (SA_EMIT_GAUSS_MEANONLY_DESCALE)
 * only.  The model's tables must already be on the device (any earlier regular call with the model — sa_batch_prepare with
 * zero reads suffices — uploads them; that upload allocates, this call does not).  Every call shares the array: calls that
 * overlap on the GPU (two host threads, two processes' streams on one context) race; one call at a time per device. */
#define SA_NOALLOC_SCRATCH_BYTES (16u << 20)
#define SA_NOALLOC_MAX_EVENTS 32768
#define SA_NOALLOC_MAX_KMERS 32768
int sa_gpu_align_noalloc(const sa_model_t *m, const sa_read_t *read, const sa_params_t *p, sa_pair_t *pairsOut,
                         int64_t pairCap, int64_t *nPairs);

/* Staged form of the same call, so that callers (and bench.py) can keep a batch resident in HBM:
 *   prepare  = host packing + H2D of inputs (+ device geometry pass + scratch allocation)
 *   run      = kernels only, inputs and outputs resident in HBM; may be called repeatedly
 *   fetch    = D2H of the compacted results
 */
typedef struct sa_batch sa_batch_t;
int sa_batch_prepare(const sa_model_t *m, const sa_read_t *reads, int64_t nReads, const sa_params_t *p,
                     const sa_ambig_t *ambig, int expectations, sa_batch_t **out);
int sa_batch_run(sa_batch_t *b);
int sa_batch_fetch(sa_batch_t *b, sa_result_t **out);
/* device time of the last sa_batch_run measured with CUDA events on the batch's stream (ms);
 * kernelMs[0] = wide band kernel + register-resident band kernel, 128- and 256-slot classes, [1] = register-resident band
 * kernel, 32- and 64-slot classes, [2] = general kernel;
 * launches = #kernel launches of the run */
int sa_batch_last_timing(const sa_batch_t *b, float *kernelMs3, int64_t *launches);
int64_t sa_batch_band_cells(const sa_batch_t *b);
/* how many work items each kernel handles: out[0] general kernel, out[1] register-resident band kernel 32/64 slots,
 * out[2] the same, 128/256 slots (incl. latency class and split mode), out[3] wide band kernel (windows of 256 .. 4096 slots) */
void sa_batch_kernel_items(const sa_batch_t *b, int64_t *out4);
/* ... and the wide kernel's items by window class: out[0..4] = 256, 512, 1024, 2048, 4096 slots */
void sa_batch_wide_items(const sa_batch_t *b, int64_t *out5);
int64_t sa_batch_h2d_bytes(const sa_batch_t *b);
int64_t sa_batch_d2h_bytes(const sa_batch_t *b);
void sa_batch_free(sa_batch_t *b);

/* test hook: element-wise device logAdd (reference logAdd, impl/pairwiseAligner.c:314-318) */
int sa_debug_logadd(const double *x, const double *y, double *out, int64_t n);

/* test hooks for the host-side integer preparation (no device needed):
 *   band_construct (impl/pairwiseAligner.c:195-246): out[3*i] = xay, xmyL, xmyR for i in [0, lX+lY]
 *   getSplitPoints (impl/pairwiseAligner.c:1910-1937): out[4*i] = x1,y1,x2,y2; returns the number of regions
 *   hdCell_construct2 path enumeration (impl/pairwiseAligner.c:723-801): concrete kmer indices of the k-window
 *   starting at `kmer`, in the reference's path order; returns the number of paths or -1 (character outside the alphabet) */
void sa_debug_band(const int64_t *anchors, int64_t nAnchors, int64_t lX, int64_t lY, int64_t expansion, int64_t *out);
int64_t sa_debug_split_points(const int64_t *anchors, int64_t nAnchors, int64_t lX, int64_t lY, int64_t splitBigger,
                              int raggedLeft, int raggedRight, int64_t *out, int64_t cap);
int64_t sa_debug_kmer_paths(const sa_model_t *m, const sa_ambig_t *ambig, const char *kmer, int32_t *out, int64_t cap);

/* ---------------------------------------------------------------------------------------------- multi-GPU
 * The one collective of the path: the element-wise sum of the per-GPU expectation tables of an EM
 * iteration — what the reference does by adding per-read expectation files in Python
 * (src/signalalign/hiddenMarkovModel.py:424-486, HmmModel.add_expectations_file) — as ncclAllReduce(sum, double) over
 * NVLink, and the concatenation of the HDP assignment lists (impl/continuousHmm.c:510-514, 571-628).  One process per GPU;
 * rank 0 creates the 128-byte id, the launcher carries it to the other ranks (MPI, a file, any key-value store), every
 * rank then calls sa_gpu_comm_init on the device it selected with sa_gpu_init.  libnccl.so.2 is loaded on first use. */
#define SA_COMM_UNIQUE_ID_BYTES 128
typedef struct sa_comm sa_comm_t;
int sa_gpu_comm_unique_id(void *id128);                                    /* ncclGetUniqueId */
int sa_gpu_comm_init(const void *id128, int nRanks, int rank, sa_comm_t **out);   /* ncclCommInitRank */
/* adopt a communicator the caller already owns (an ncclComm_t); it is not destroyed by sa_gpu_comm_free */
int sa_gpu_comm_from_nccl(void *ncclComm, int nRanks, int rank, sa_comm_t **out);
void sa_gpu_comm_free(sa_comm_t *c);
/* in place on a host table of n doubles (same n on every rank): 10 = what sa_result_sum_expectations returns (9
 * transitions + likelihood, all the reference's C path accumulates: the emission slots of the expectations file keep
 * their pseudocounts, impl/pairwiseAligner.c:924-943), or 10 + 3 K for the file's full layout (continuousHmm.c:353-407) */
int sa_gpu_allreduce_expectations(sa_comm_t *c, double *table, int64_t n);
/* every rank's (kmer position, event) assignment list, concatenated in rank order on every rank; *allKmerPos and
 * *allEvent are malloc'ed (free them), countsPerRank has nRanks entries */
int sa_gpu_allgather_assignments(sa_comm_t *c, const int32_t *kmerPos, const int32_t *event, int64_t n,
                                 int32_t **allKmerPos, int32_t **allEvent, int64_t *countsPerRank);

/* expectation-table reduction across reads of one GPU (the per-GPU table NCCL then all-reduces):
 * out[0..8] = sum of transitions, out[9] = sum of likelihood */
void sa_result_sum_expectations(const sa_result_t *r, double *out10);

#ifdef __cplusplus
}
#endif
#endif
