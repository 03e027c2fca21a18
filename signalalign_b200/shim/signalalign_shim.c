/*
 * signalalign_shim.c — the reference-side binding of libsagpu: the two entry points signalMachine calls,
 *
 *     getAlignedPairsUsingAnchors   (reference inc/pairwiseAligner.h:406-414, call site impl/signalMachine.c:436-437)
 *     getExpectationsUsingAnchors   (reference inc/pairwiseAligner.h:417-429, call site impl/signalMachine.c:479-480)
 *
 * with the reference's own prototypes, implemented on the C-ABI of include/signalalign_b200.h.  This file is compiled
 * INSIDE the reference tree (it needs inc/stateMachine.h, inc/pairwiseAligner.h and sonLib) next to pairwiseAligner.c,
 * whose own two definitions are renamed at compile time (-DgetAlignedPairsUsingAnchors=cpu_getAlignedPairsUsingAnchors
 * -DgetExpectationsUsingAnchors=cpu_getExpectationsUsingAnchors on that one translation unit), so every caller —
 * signalMachine.c, the C test suites — links against the GPU path unchanged.  See INTEGRATION.md.
 *
 * Behaviour kept from the reference: inputs are borrowed; the returned stList owns stIntTuple4(floor(p*1e7), x, y,
 * (int64_t) char *kmer) with destructor stIntTuple_destruct and kmer strings that are never freed
 * (impl/pairwiseAligner.c:665-668, 2064); expectations are added into the caller's Hmm through its own function
 * pointers; any failure aborts the process through st_errAbort (no return codes, no CPU fallback).
 * StateMachine3_HDP: hdp.c keeps its structs private, so the machine's NanoporeHDP travels through the reference's own
 * serialize_nhdp (impl/nanopore_hdp.c:1076-1086, %.17lg: exact round trip) into a temporary file that
 * sa_model_attach_nhdp parses; HDP expectations append their (kmer, event) assignments through HdpHmm::addToAssignments
 * with pointers into sX->elements / sY->elements, as cell_signal_updateExpectationsAndAssignments does
 * (impl/pairwiseAligner.c:946-968).
 * The ambiguity table of the call is the one the caller put into the target Sequence (SsX->ambigBases: the stHash of
 * create_ambig_bases2(ambig_path), impl/signalMachine.c:431, 474), converted entry by entry; a Sequence without one gets
 * the built-in table.
 * The device model of a StateMachine is cached by CONTENT (transitions, emission tables, scale-independent): a machine
 * changed in place between two calls (hmmContinuous_loadIntoStateMachine, emissions_signal_scaleNoise, ...), or a new
 * machine at a recycled address, is uploaded again.
 */
#include <stdlib.h>
#include <string.h>
#include <unistd.h>

#include "signalalign_b200.h"

#include "pairwiseAligner.h"
#include "stateMachine.h"
#include "continuousHmm.h"
#include "signalalign_shim.h"

/* device models, keyed by the machine's CONTENT: type, emission function, kmer geometry, the nine transitions and a hash
 * of the three emission tables (+ the address of the NanoporeHDP, whose private structs cannot be hashed from here) */
#define SHIM_CACHE 8
static struct {
    uint64_t key;
    const void *hdp;
    sa_model_t *model;
} g_cache[SHIM_CACHE];
static int g_cacheNext = 0, g_inited = 0;

static uint64_t shimMix(uint64_t h, const void *data, size_t bytes) {
    const uint64_t *w = (const uint64_t *) data;
    for (size_t i = 0; i < bytes / 8; i++) {
        h ^= w[i];
        h *= 0x9E3779B97F4A7C15ULL;
        h ^= h >> 29;
    }
    return h;
}
static uint64_t shimKey(const StateMachine *sM, const double *t9, const void *fn) {
    uint64_t h = 0xcbf29ce484222325ULL;
    uint64_t head[4] = {(uint64_t) sM->type, (uint64_t) sM->kmerLength, (uint64_t) sM->parameterSetSize, (uint64_t) (uintptr_t) fn};
    h = shimMix(h, head, sizeof(head));
    h = shimMix(h, t9, 9 * sizeof(double));
    for (int64_t i = 0; i < sM->alphabetSize; i++) h = (h ^ (uint64_t) (unsigned char) sM->alphabet[i]) * 0x100000001b3ULL;
    h = shimMix(h, sM->EMISSION_MATCH_MATRIX, (size_t) sM->parameterSetSize * MODEL_PARAMS * sizeof(double));
    h = shimMix(h, sM->EMISSION_GAP_Y_MATRIX, (size_t) sM->parameterSetSize * MODEL_PARAMS * sizeof(double));
    h = shimMix(h, sM->EMISSION_GAP_X_PROBS, (size_t) sM->parameterSetSize * sizeof(double));
    return h;
}

sa_model_t *sa_shim_model(StateMachine *sM) {
    if (!g_inited) {
        const char *dev = getenv("SA_SHIM_DEVICE");
        if (sa_gpu_init(dev ? atoi(dev) : 0) != SA_OK) st_errAbort("libsagpu: %s\n", sa_last_error());
        g_inited = 1;
    }
    const void *hdp = sM->type == threeStateHdp ? (const void *) ((StateMachine3_HDP *) sM)->hdpModel : NULL;
    uint64_t key = 0;
    if (sM->type == threeStateHdp) {
        StateMachine3_HDP *sh = (StateMachine3_HDP *) sM;
        key = shimKey(sM, &sh->TRANSITION_MATCH_CONTINUE, (const void *) sh->getMatchProbFcn);
    } else if (sM->type == threeState) {
        StateMachine3 *s3 = (StateMachine3 *) sM;
        key = shimKey(sM, &s3->TRANSITION_MATCH_CONTINUE, (const void *) s3->getMatchProbFcn);
    }
    for (int i = 0; i < SHIM_CACHE; i++)
        if (g_cache[i].model && g_cache[i].key == key && g_cache[i].hdp == hdp) return g_cache[i].model;
    sa_model_t *m = NULL;
    if (sM->type == threeStateHdp) {
        /* StateMachine3_HDP keeps the transition block at the same place as StateMachine3 (inc/stateMachine.h:150-191);
         * the tables are taken as they are (stateMachine3_setModelToHdpExpectedValues has or has not been applied by
         * the caller, impl/stateMachine.c:1275-1304) */
        StateMachine3_HDP *sh = (StateMachine3_HDP *) sM;
        if (sh->getMatchProbFcn != emissions_signal_getHdpKmerDensity || sh->hdpModel == NULL)
            st_errAbort("signalalign_shim: unknown emission function in StateMachine3_HDP\n");
        double t[9] = {sh->TRANSITION_MATCH_CONTINUE, sh->TRANSITION_MATCH_FROM_GAP_X, sh->TRANSITION_MATCH_FROM_GAP_Y,
                       sh->TRANSITION_GAP_OPEN_X, sh->TRANSITION_GAP_OPEN_Y, sh->TRANSITION_GAP_EXTEND_X,
                       sh->TRANSITION_GAP_EXTEND_Y, sh->TRANSITION_GAP_SWITCH_TO_X, sh->TRANSITION_GAP_SWITCH_TO_Y};
        m = sa_model_from_tables(sM->alphabet, (int32_t) sM->kmerLength, t, sM->EMISSION_MATCH_MATRIX,
                                 sM->EMISSION_GAP_Y_MATRIX, sM->EMISSION_GAP_X_PROBS, SA_EMIT_HDP);
        if (!m) st_errAbort("libsagpu: %s\n", sa_last_error());
        const char *dir = getenv("TMPDIR");
        char path[4096];
        snprintf(path, sizeof(path), "%s/sagpu_nhdp_XXXXXX", dir && dir[0] ? dir : "/tmp");
        int fd = mkstemp(path);
        if (fd < 0) st_errAbort("signalalign_shim: cannot create %s\n", path);
        close(fd);
        serialize_nhdp(sh->hdpModel, path);
        int rc = sa_model_attach_nhdp(m, path);
        unlink(path);
        if (rc != SA_OK) st_errAbort("libsagpu: %s\n", sa_last_error());
    } else if (sM->type == threeState) {
        StateMachine3 *s3 = (StateMachine3 *) sM;
        double t[9] = {s3->TRANSITION_MATCH_CONTINUE, s3->TRANSITION_MATCH_FROM_GAP_X, s3->TRANSITION_MATCH_FROM_GAP_Y,
                       s3->TRANSITION_GAP_OPEN_X, s3->TRANSITION_GAP_OPEN_Y, s3->TRANSITION_GAP_EXTEND_X,
                       s3->TRANSITION_GAP_EXTEND_Y, s3->TRANSITION_GAP_SWITCH_TO_X, s3->TRANSITION_GAP_SWITCH_TO_Y};
        int kind;
        if (s3->getMatchProbFcn == emissions_signal_strawManGetKmerEventMatchProbWithDescaling_MeanOnly)
            kind = SA_EMIT_GAUSS_MEANONLY_DESCALE;
        else if (s3->getMatchProbFcn == emissions_signal_strawManGetKmerEventMatchProbWithDescaling)
            kind = SA_EMIT_GAUSS_NOISE_DESCALE;
        else if (s3->getMatchProbFcn == emissions_signal_strawManGetKmerEventMatchProb)
            kind = SA_EMIT_GAUSS_NOISE_SCALED;
        else {
            st_errAbort("signalalign_shim: unknown emission function in StateMachine3\n");
            return NULL;
        }
        m = sa_model_from_tables(sM->alphabet, (int32_t) sM->kmerLength, t, sM->EMISSION_MATCH_MATRIX,
                                 sM->EMISSION_GAP_Y_MATRIX, sM->EMISSION_GAP_X_PROBS, kind);
        if (!m) st_errAbort("libsagpu: %s\n", sa_last_error());
    } else {
        st_errAbort("signalalign_shim: only threeState and threeStateHdp machines are bound (signalMachine builds no other)\n");
    }
    int slot = g_cacheNext++ % SHIM_CACHE;
    if (g_cache[slot].model) sa_model_free(g_cache[slot].model);
    g_cache[slot].key = key;
    g_cache[slot].hdp = hdp;
    g_cache[slot].model = m;
    return m;
}

/* the flat view of one call's inputs; *anchorsOut is malloc'ed */
static void shimRead(StateMachine *sM, Sequence *sX, Sequence *sY, stList *anchorPairs, bool raggedL, bool raggedR,
                     sa_read_t *r, int64_t **anchorsOut) {
    int64_t nA = stList_length(anchorPairs);
    int64_t *an = st_malloc(sizeof(int64_t) * 2 * (nA + 1));
    for (int64_t i = 0; i < nA; i++) {
        stIntTuple *a = stList_get(anchorPairs, i);
        an[2 * i] = stIntTuple_get(a, 0);
        an[2 * i + 1] = stIntTuple_get(a, 1);
    }
    memset(r, 0, sizeof(*r));
    r->target = (const char *) sX->elements;
    r->targetLength = sX->length > 0 ? sX->length + sM->kmerLength - 1 : 0;   /* sX->length counts kmers */
    r->events = (const double *) sY->elements;
    r->nEvents = sY->length;
    r->anchors = an;
    r->nAnchors = nA;
    r->scale = sM->scale; r->shift = sM->shift; r->var = sM->var;
    r->raggedLeft = raggedL; r->raggedRight = raggedR;
    *anchorsOut = an;
}

/* SsX->ambigBases (stHash: code string -> bases string) as an sa_ambig_t; NULL = the built-in table.  hdCell_construct2
 * looks up one character at a time (impl/pairwiseAligner.c:755), so only one-character codes can ever match. */
sa_ambig_t *sa_shim_ambig(stHash *ambigBases) {
    if (ambigBases == NULL) return NULL;
    sa_ambig_t *a = sa_ambig_create();
    stHashIterator *it = stHash_getIterator(ambigBases);
    char *code;
    while ((code = stHash_getNext(it)) != NULL) {
        const char *bases = stHash_search(ambigBases, code);
        if (strlen(code) != 1 || bases == NULL) continue;
        if (sa_ambig_set(a, code[0], bases) != SA_OK) st_errAbort("libsagpu: %s\n", sa_last_error());
    }
    stHash_destructIterator(it);
    return a;
}

static void shimParams(const PairwiseAlignmentParameters *p, sa_params_t *sp) {
    sa_params_default(sp);
    sp->threshold = p->threshold;
    sp->minDiagsBetweenTraceBack = p->minDiagsBetweenTraceBack;
    sp->traceBackDiagonals = p->traceBackDiagonals;
    sp->diagonalExpansion = p->diagonalExpansion;
    sp->splitMatrixBiggerThanThis = p->splitMatrixBiggerThanThis;
}

stList *getAlignedPairsUsingAnchors(StateMachine *sM, Sequence *SsX, Sequence *SsY, stList *anchorPairs,
                                    PairwiseAlignmentParameters *p,
                                    void (*diagonalPosteriorProbFn)(StateMachine *sM, int64_t xay, DpMatrix *forwardDpMatrix,
                                                                    DpMatrix *backwardDpMatrix, Sequence *sX, Sequence *sY,
                                                                    double totalProbability,
                                                                    PairwiseAlignmentParameters *p, void *extraArgs),
                                    bool alignmentHasRaggedLeftEnd, bool alignmentHasRaggedRightEnd) {
    if (diagonalPosteriorProbFn != diagonalCalculationPosteriorMatchProbs)
        st_errAbort("signalalign_shim: getAlignedPairsUsingAnchors is bound for diagonalCalculationPosteriorMatchProbs only\n");
    sa_model_t *m = sa_shim_model(sM);
    sa_read_t r;
    int64_t *an;
    shimRead(sM, SsX, SsY, anchorPairs, alignmentHasRaggedLeftEnd, alignmentHasRaggedRightEnd, &r, &an);
    sa_params_t sp;
    shimParams(p, &sp);
    sa_result_t *res = NULL;
    sa_ambig_t *amb = sa_shim_ambig(SsX->ambigBases);
    if (sa_gpu_align_batch(m, &r, 1, &sp, amb, &res) != SA_OK) st_errAbort("libsagpu: %s\n", sa_last_error());
    sa_ambig_free(amb);
    stList *out = stList_construct3(0, (void (*)(void *)) stIntTuple_destruct);
    const int64_t n = res->pairOffsets[1], k = sM->kmerLength;
    /* one arena for the call's kmer strings; like the reference's, they outlive the list and are never freed */
    char *arena = n > 0 ? st_malloc((size_t) n * (k + 1)) : NULL;
    for (int64_t i = 0; i < n; i++) {
        char *km = arena + i * (k + 1);
        sa_kmer_string(m, res->pairs[i].kmerIdx, km);
        stList_append(out, stIntTuple_construct4(res->pairs[i].p, res->pairs[i].x, res->pairs[i].y, (int64_t) km));
    }
    sa_result_free(res);
    free(an);
    return out;
}

void getExpectationsUsingAnchors(StateMachine *sM, Hmm *hmmExpectations, Sequence *SsX, Sequence *SsY, stList *anchorPairs,
                                 PairwiseAlignmentParameters *p,
                                 void (*diagonalCalcExpectationFcn)(StateMachine *sM, int64_t xay, DpMatrix *forwardDpMatrix,
                                                                    DpMatrix *backwardDpMatrix, Sequence *sX, Sequence *sY,
                                                                    double totalProbability,
                                                                    PairwiseAlignmentParameters *p, void *extraArgs),
                                 bool alignmentHasRaggedLeftEnd, bool alignmentHasRaggedRightEnd) {
    if (diagonalCalcExpectationFcn != diagonalCalculation_Expectations)
        st_errAbort("signalalign_shim: getExpectationsUsingAnchors is bound for diagonalCalculation_Expectations only\n");
    sa_model_t *m = sa_shim_model(sM);
    sa_read_t r;
    int64_t *an;
    shimRead(sM, SsX, SsY, anchorPairs, alignmentHasRaggedLeftEnd, alignmentHasRaggedRightEnd, &r, &an);
    sa_params_t sp;
    shimParams(p, &sp);
    /* the assignment gate is the HdpHmm's own threshold (impl/pairwiseAligner.c:963), not the alignment parameter's */
    if (sM->type == threeStateHdp) sp.threshold = ((HdpHmm *) hmmExpectations)->threshold;
    sa_result_t *res = NULL;
    sa_ambig_t *amb = sa_shim_ambig(SsX->ambigBases);
    if (sa_gpu_expectations_batch(m, &r, 1, &sp, amb, &res) != SA_OK) st_errAbort("libsagpu: %s\n", sa_last_error());
    sa_ambig_free(amb);
    /* cell_signal_updateExpectations adds every term through the Hmm's own function (continuousHmm.c:147-149);
     * the likelihood is a plain field (pairwiseAligner.c:1432) */
    for (int64_t from = 0; from < 3; from++)
        for (int64_t to = 0; to < 3; to++)
            hmmExpectations->addToTransitionExpectationFcn(hmmExpectations, from, to, res->transitions[from * 3 + to]);
    hmmExpectations->likelihood += res->likelihood[0];
    if (sM->type == threeStateHdp && res->assignOffsets != NULL) {
        /* the list holds POINTERS into the caller's sequences (continuousHmm.c:510-514), in the reference's order */
        HdpHmm *hh = (HdpHmm *) hmmExpectations;
        char *target = (char *) SsX->elements;
        double *events = (double *) SsY->elements;
        for (int64_t i = res->assignOffsets[0]; i < res->assignOffsets[1]; i++)
            hh->addToAssignments(hh, target + res->assignKmerPos[i], events + (int64_t) res->assignEvent[i] * NB_EVENT_PARAMS);
    }
    sa_result_free(res);
    free(an);
}
