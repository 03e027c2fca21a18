/*
 * ref_harness.c — TEST INFRASTRUCTURE ONLY.
 *
 * Flat, ctypes-friendly entry points around the UNMODIFIED reference hot path
 * (compiled from /root/reference by oracle/Makefile into oracle/_ref/libsaref.so).
 * Nothing here is reference code: every function below only *calls* the reference's
 * public functions (inc/pairwiseAligner.h, inc/stateMachine.h, inc/nanopore.h,
 * inc/continuousHmm.h, inc/nanopore_hdp.h) so that tests, golden-vector generation
 * and bench.py's `--impl reference` arm can drive them with plain arrays.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / reference arm
 * may load this library.  The product (signalalign_b200/) never does.
 */
#define _GNU_SOURCE
#include <stdlib.h>
#include <stdio.h>
#include <string.h>
#include <stdint.h>
#include <time.h>
#include <unistd.h>
#include "sonLib.h"
#include "pairwiseAlignment.h"
#include "pairwiseAligner.h"
#include "stateMachine.h"
#include "continuousHmm.h"
#include "nanopore.h"
#include "nanopore_hdp.h"
#include "signalMachineUtils.h"

/* ------------------------------------------------------------------ state machines */

/* kind: 0 = production signalMachine builder (buildStateMachine2, signalMachine.c:336-362):
 *           Gaussian MeanOnly emission with de-scaling, or HDP density when nhdpFile != NULL
 *       1 = getStateMachine3 (stateMachine.c:1757): "scaled model" emission incl. inverse-Gaussian noise
 *       2 = getStateMachine3_descaled emission fn (stateMachine.c:1738): de-scaling incl. noise term
 */
void *ref_sm_load(const char *modelFile, int kind, const char *nhdpFile) {
    if (nhdpFile != NULL && nhdpFile[0] != '\0') {
        NanoporeHDP *nHdp = deserialize_nhdp(nhdpFile);
        return stateMachine3_loadFromFile(modelFile, threeStateHdp, NULL, emissions_signal_getHdpKmerDensity,
                                          stateMachine3_loadTransitionsFromFile, nHdp);
    }
    switch (kind) {
    case 0:
        return stateMachine3_loadFromFile(modelFile, threeState, emissions_kmer_getGapProb,
                                          emissions_signal_strawManGetKmerEventMatchProbWithDescaling_MeanOnly,
                                          stateMachine3_loadTransitionsFromFile, NULL);
    case 1:
        return getStateMachine3(modelFile);
    case 2:
        return stateMachine3_loadFromFile(modelFile, threeState, emissions_kmer_getGapProb,
                                          emissions_signal_strawManGetKmerEventMatchProbWithDescaling,
                                          stateMachine3_loadTransitionsFromFile, NULL);
    }
    return NULL;
}

void ref_sm_set_scaling(void *h, double scale, double shift, double var) {
    StateMachine *sM = h;
    sM->scale = scale; sM->shift = shift; sM->var = var;
}

void ref_sm_scale_model(void *h, double scale, double shift, double var, double scale_sd, double var_sd) {
    emissions_signal_scaleModel((StateMachine *) h, scale, shift, var, scale_sd, var_sd);
}

void ref_sm_scale_noise(void *h, double scale_sd, double var_sd) {
    NanoporeReadAdjustmentParameters npp;
    memset(&npp, 0, sizeof(npp));
    npp.scale_sd = scale_sd; npp.var_sd = var_sd;
    emissions_signal_scaleNoise((StateMachine *) h, npp);
}

void ref_sm_hdp_expected_values(void *h) {
    StateMachine3_HDP *sM = h;
    stateMachine3_setModelToHdpExpectedValues((StateMachine *) h, sM->hdpModel);
}

/* info[0..5] = type, K, alphabetSize, kmerLength, stateNumber, matchState */
void ref_sm_info(void *h, int64_t *info, char *alphabet, double *scaling3, double *transitions9) {
    StateMachine *sM = h;
    StateMachine3 *s3 = h; /* StateMachine3 and StateMachine3_HDP share the transition layout (stateMachine.h:150-185) */
    info[0] = sM->type; info[1] = sM->parameterSetSize; info[2] = sM->alphabetSize;
    info[3] = sM->kmerLength; info[4] = sM->stateNumber; info[5] = sM->matchState;
    memcpy(alphabet, sM->alphabet, sM->alphabetSize); alphabet[sM->alphabetSize] = 0;
    scaling3[0] = sM->scale; scaling3[1] = sM->shift; scaling3[2] = sM->var;
    transitions9[0] = s3->TRANSITION_MATCH_CONTINUE;  transitions9[1] = s3->TRANSITION_MATCH_FROM_GAP_X;
    transitions9[2] = s3->TRANSITION_MATCH_FROM_GAP_Y; transitions9[3] = s3->TRANSITION_GAP_OPEN_X;
    transitions9[4] = s3->TRANSITION_GAP_OPEN_Y;       transitions9[5] = s3->TRANSITION_GAP_EXTEND_X;
    transitions9[6] = s3->TRANSITION_GAP_EXTEND_Y;     transitions9[7] = s3->TRANSITION_GAP_SWITCH_TO_X;
    transitions9[8] = s3->TRANSITION_GAP_SWITCH_TO_Y;
}

/* copies K*5 doubles of each table */
void ref_sm_tables(void *h, double *match, double *gapY, double *gapX) {
    StateMachine *sM = h;
    memcpy(match, sM->EMISSION_MATCH_MATRIX, sizeof(double) * sM->parameterSetSize * MODEL_PARAMS);
    memcpy(gapY, sM->EMISSION_GAP_Y_MATRIX, sizeof(double) * sM->parameterSetSize * MODEL_PARAMS);
    memcpy(gapX, sM->EMISSION_GAP_X_PROBS, sizeof(double) * sM->parameterSetSize);
}

/* emission probes: direct calls of the v-table emission functions */
double ref_sm_emission(void *h, const char *kmer, double eventMean, double eventNoise, int match) {
    StateMachine *sM = h;
    double ev[4] = { eventMean, eventNoise, 0.0, 0.0 };
    if (sM->type == threeStateHdp) {
        return ((StateMachine3_HDP *) h)->getMatchProbFcn(sM, (void *) kmer, ev, match);
    }
    StateMachine3 *s3 = h;
    return match ? s3->getMatchProbFcn(sM, (void *) kmer, ev, 1) : s3->getYGapProbFcn(sM, (void *) kmer, ev, 0);
}

double ref_hdp_density(void *h, const char *kmer, double x) {
    StateMachine3_HDP *sM = h;
    return get_nanopore_kmer_density(sM->hdpModel, (void *) kmer, &x);
}

int64_t ref_kmer_id(const char *kmer, const char *alphabet, int64_t alphabetSize, int64_t k) {
    return kmer_id((char *) kmer, (char *) alphabet, alphabetSize, k);
}

double ref_logAdd(double x, double y) { return logAdd(x, y); }

/* ------------------------------------------------------------------ band / split */

/* out: 3*(lX+lY+1) int64 (xay, xmyL, xmyR) — band_construct, pairwiseAligner.c:195 */
void ref_band(const int64_t *anchors, int64_t nAnchors, int64_t lX, int64_t lY, int64_t expansion, int64_t *out) {
    stList *l = stList_construct3(0, (void (*)(void *)) stIntTuple_destruct);
    for (int64_t i = 0; i < nAnchors; i++) stList_append(l, stIntTuple_construct2(anchors[2 * i], anchors[2 * i + 1]));
    Band *band = band_construct(l, lX, lY, expansion);
    BandIterator *it = bandIterator_construct(band);
    for (int64_t i = 0; i <= lX + lY; i++) {
        Diagonal d = bandIterator_getNext(it);
        out[3 * i] = diagonal_getXay(d); out[3 * i + 1] = diagonal_getMinXmy(d); out[3 * i + 2] = diagonal_getMaxXmy(d);
    }
    bandIterator_destruct(it);
    band_destruct(band);
    stList_destruct(l);
}

/* out: 4 int64 per split region; returns count — getSplitPoints, pairwiseAligner.c:1910 */
int64_t ref_split_points(const int64_t *anchors, int64_t nAnchors, int64_t lX, int64_t lY, int64_t splitBigger,
                         int raggedL, int raggedR, int64_t *out, int64_t cap) {
    stList *l = stList_construct3(0, (void (*)(void *)) stIntTuple_destruct);
    for (int64_t i = 0; i < nAnchors; i++) stList_append(l, stIntTuple_construct2(anchors[2 * i], anchors[2 * i + 1]));
    stList *sp = getSplitPoints(l, lX, lY, splitBigger, raggedL, raggedR);
    int64_t n = stList_length(sp);
    for (int64_t i = 0; i < n && i < cap; i++) {
        stIntTuple *t = stList_get(sp, i);
        for (int j = 0; j < 4; j++) out[4 * i + j] = stIntTuple_get(t, j);
    }
    stList_destruct(sp);
    stList_destruct(l);
    return n;
}

/* ------------------------------------------------------------------ the hot path */

static PairwiseAlignmentParameters *mkParams(double threshold, int64_t minDiags, int64_t traceBack, int64_t expansion,
                                             int64_t splitBigger) {
    PairwiseAlignmentParameters *p = pairwiseAlignmentBandingParameters_construct();
    p->threshold = threshold;
    if (minDiags > 0) p->minDiagsBetweenTraceBack = minDiags;
    if (traceBack > 0) p->traceBackDiagonals = traceBack;
    if (expansion >= 0) p->diagonalExpansion = expansion;
    if (splitBigger > 0) p->splitMatrixBiggerThanThis = splitBigger;
    return p;
}

static stList *mkAnchors(const int64_t *anchors, int64_t nAnchors) {
    stList *l = stList_construct3(0, (void (*)(void *)) stIntTuple_destruct);
    for (int64_t i = 0; i < nAnchors; i++) stList_append(l, stIntTuple_construct2(anchors[2 * i], anchors[2 * i + 1]));
    return l;
}

/*
 * getAlignedPairsUsingAnchors (pairwiseAligner.c:2052) on flat inputs.
 *   target: NUL-terminated nucleotide string with lX + k - 1 characters
 *   events: lY * 4 doubles
 *   out:    cap * 3 int64  (floor(p*1e7), x, y) in the reference's emission order
 *   outKmers: cap * (k+1) chars, the path kmer of each pair
 * returns number of pairs (may exceed cap; only cap are written); *seconds = wall time of the call.
 */
int64_t ref_align(void *h, const char *target, int64_t lX, double *events, int64_t lY,
                  const int64_t *anchors, int64_t nAnchors,
                  double threshold, int64_t minDiags, int64_t traceBack, int64_t expansion, int64_t splitBigger,
                  int raggedL, int raggedR, const char *ambigFile,
                  int64_t *out, char *outKmers, int64_t cap, double *seconds) {
    StateMachine *sM = h;
    PairwiseAlignmentParameters *p = mkParams(threshold, minDiags, traceBack, expansion, splitBigger);
    stList *anchorPairs = mkAnchors(anchors, nAnchors);
    stHash *ambig = create_ambig_bases2((ambigFile && ambigFile[0]) ? (char *) ambigFile : NULL);
    Sequence *sX = sequence_constructKmerSequence(lX, (void *) target, sequence_getKmer,
                                                  sequence_sliceNucleotideSequence, kmer, ambig);
    Sequence *sY = sequence_constructEventSequence(lY, events);
    struct timespec t0, t1;
    clock_gettime(CLOCK_MONOTONIC, &t0);
    stList *pairs = getAlignedPairsUsingAnchors(sM, sX, sY, anchorPairs, p, diagonalCalculationPosteriorMatchProbs,
                                                raggedL, raggedR);
    clock_gettime(CLOCK_MONOTONIC, &t1);
    if (seconds) *seconds = (t1.tv_sec - t0.tv_sec) + 1e-9 * (t1.tv_nsec - t0.tv_nsec);
    int64_t n = stList_length(pairs);
    int64_t k = sM->kmerLength;
    for (int64_t i = 0; i < n && i < cap; i++) {
        stIntTuple *t = stList_get(pairs, i);
        out[3 * i] = stIntTuple_get(t, 0); out[3 * i + 1] = stIntTuple_get(t, 1); out[3 * i + 2] = stIntTuple_get(t, 2);
        if (outKmers) {
            const char *km = (const char *) stIntTuple_get(t, 3);
            memset(outKmers + i * (k + 1), 0, k + 1);
            if (km) strncpy(outKmers + i * (k + 1), km, k);
        }
    }
    stList_destruct(pairs);
    stList_destruct(anchorPairs);
    sequence_destruct(sX);
    sequence_destruct(sY);
    stHash_destruct(ambig);
    pairwiseAlignmentBandingParameters_destruct(p);
    return n;
}

/* getAlignedPairsWithoutBanding (pairwiseAligner.c:2104): the reference's own un-banded second oracle */
int64_t ref_align_noband(void *h, const char *target, int64_t lX, double *events, int64_t lY, double threshold,
                         int raggedL, int raggedR, int64_t *out, int64_t cap) {
    StateMachine *sM = h;
    PairwiseAlignmentParameters *p = mkParams(threshold, 0, 0, -1, 0);
    stList *pairs = getAlignedPairsWithoutBanding(sM, (void *) target, events, lX, lY, p, sequence_getKmer,
                                                  sequence_getEvent, diagonalCalculationPosteriorMatchProbs,
                                                  raggedL, raggedR);
    int64_t n = stList_length(pairs);
    for (int64_t i = 0; i < n && i < cap; i++) {
        stIntTuple *t = stList_get(pairs, i);
        out[3 * i] = stIntTuple_get(t, 0); out[3 * i + 1] = stIntTuple_get(t, 1); out[3 * i + 2] = stIntTuple_get(t, 2);
    }
    stList_destruct(pairs);
    pairwiseAlignmentBandingParameters_destruct(p);
    return n;
}

/*
 * getExpectationsUsingAnchors (pairwiseAligner.c:2164) with the production expectations HMM
 * (hmmContinuous_getExpectationsHmm, continuousHmm.c:841; pseudocounts as signalMachine.c:785).
 *   transitions9: row-major from*3+to (incl. pseudocount), *likelihood: hmm->likelihood
 *   HDP machines: assignKmerIdx / assignEventIdx receive, per assignment, the offset of the kmer in `target`
 *   and the event index (pointers turned back into indices); returns number of assignments (0 for Gaussian).
 */
int64_t ref_expectations(void *h, const char *target, int64_t lX, double *events, int64_t lY,
                         const int64_t *anchors, int64_t nAnchors,
                         double threshold, int64_t minDiags, int64_t traceBack, int64_t expansion, int64_t splitBigger,
                         int raggedL, int raggedR, const char *ambigFile, double pseudoT, double pseudoE,
                         double *transitions9, double *likelihood,
                         int64_t *assignKmerIdx, int64_t *assignEventIdx, int64_t cap, double *seconds) {
    StateMachine *sM = h;
    PairwiseAlignmentParameters *p = mkParams(threshold, minDiags, traceBack, expansion, splitBigger);
    stList *anchorPairs = mkAnchors(anchors, nAnchors);
    stHash *ambig = create_ambig_bases2((ambigFile && ambigFile[0]) ? (char *) ambigFile : NULL);
    Sequence *sX = sequence_constructKmerSequence(lX, (void *) target, sequence_getKmer,
                                                  sequence_sliceNucleotideSequence, kmer, ambig);
    Sequence *sY = sequence_constructEventSequence(lY, events);
    Hmm *hmm = hmmContinuous_getExpectationsHmm(sM, threshold, pseudoT, pseudoE);
    struct timespec t0, t1;
    clock_gettime(CLOCK_MONOTONIC, &t0);
    getExpectationsUsingAnchors(sM, hmm, sX, sY, anchorPairs, p, diagonalCalculation_Expectations, raggedL, raggedR);
    clock_gettime(CLOCK_MONOTONIC, &t1);
    if (seconds) *seconds = (t1.tv_sec - t0.tv_sec) + 1e-9 * (t1.tv_nsec - t0.tv_nsec);
    for (int i = 0; i < 9; i++) transitions9[i] = hmm->transitions[i];
    *likelihood = hmm->likelihood;
    int64_t n = 0;
    if (sM->type == threeStateHdp) {
        HdpHmm *hh = (HdpHmm *) hmm;
        n = hh->numberOfAssignments;
        for (int64_t i = 0; i < n && i < cap; i++) {
            assignKmerIdx[i] = (const char *) stList_get(hh->kmerAssignments, i) - target;
            assignEventIdx[i] = ((double *) stList_get(hh->eventAssignments, i) - events) / NB_EVENT_PARAMS;
        }
    }
    stList_destruct(anchorPairs);
    sequence_destruct(sX);
    sequence_destruct(sY);
    pairwiseAlignmentBandingParameters_destruct(p);
    return n;
}

/* ------------------------------------------------------------------ .npRead + anchors */

void *ref_npread_load(const char *path, const char *alphabet, int64_t alphabetSize) {
    return nanopore_loadNanoporeReadFromFile(path, (char *) alphabet, alphabetSize);
}

/* lens[0..4] = readLength, templateReadLength, complementReadLength, nbTemplateEvents, nbComplementEvents
 * tparams[0..6] = scale, shift, var, shift_sd, scale_sd, var_sd, drift (template) */
void ref_npread_info(void *h, int64_t *lens, double *tparams) {
    NanoporeRead *r = h;
    lens[0] = r->readLength; lens[1] = r->templateReadLength; lens[2] = r->complementReadLength;
    lens[3] = r->nbTemplateEvents; lens[4] = r->nbComplementEvents;
    NanoporeReadAdjustmentParameters *q = &r->templateParams;
    tparams[0] = q->scale; tparams[1] = q->shift; tparams[2] = q->var; tparams[3] = q->shift_sd;
    tparams[4] = q->scale_sd; tparams[5] = q->var_sd; tparams[6] = q->drift;
}

const char *ref_npread_twoD(void *h) { return ((NanoporeRead *) h)->twoDread; }
const char *ref_npread_template_read(void *h) { return ((NanoporeRead *) h)->templateRead; }
double *ref_npread_template_events(void *h) { return ((NanoporeRead *) h)->templateEvents; }
int64_t *ref_npread_template_event_map(void *h) { return ((NanoporeRead *) h)->templateEventMap; }
int64_t *ref_npread_template_strand_event_map(void *h) { return ((NanoporeRead *) h)->templateStrandEventMap; }
void ref_npread_descale(void *h) { nanopore_descaleNanoporeRead((NanoporeRead *) h); }

/* signalUtils_estimateNanoporeParams (signalMachineUtils.c:186) on the template strand */
void ref_estimate_params_template(void *smh, void *rh) {
    NanoporeRead *r = rh;
    signalUtils_estimateNanoporeParams((StateMachine *) smh, r, &r->templateParams, 0.0,
                                       signalUtils_templateOneDAssignmentsFromRead,
                                       nanopore_adjustTemplateEventsForDrift);
}

/*
 * Anchors exactly as the reference tests make them (tests/stateMachineTests.c:130-144):
 * lastz pairs between refSeq and readSeq -> remap through eventMap -> filterToRemoveOverlap.
 * Needs ./cPecanLastz in the current directory (pairwiseAligner.c:1694); `lastzDir` is chdir'ed into for the call.
 */
int64_t ref_lastz_anchors(const char *refSeq, const char *readSeq, const int64_t *eventMap, const char *lastzDir,
                          int64_t *out, int64_t cap) {
    char cwd[4096];
    if (!getcwd(cwd, sizeof(cwd))) return -1;
    if (chdir(lastzDir) != 0) return -1;
    PairwiseAlignmentParameters *p = pairwiseAlignmentBandingParameters_construct();
    stList *anchorPairs = getBlastPairsForPairwiseAlignmentParameters((void *) refSeq, (void *) readSeq, p);
    if (chdir(cwd) != 0) return -1;
    stList *remapped = nanopore_remapAnchorPairs(anchorPairs, (int64_t *) eventMap);
    stList *filtered = filterToRemoveOverlap(remapped);
    int64_t n = stList_length(filtered);
    for (int64_t i = 0; i < n && i < cap; i++) {
        stIntTuple *t = stList_get(filtered, i);
        out[2 * i] = stIntTuple_get(t, 0); out[2 * i + 1] = stIntTuple_get(t, 1);
    }
    stList_destruct(filtered);
    stList_destruct(remapped);
    stList_destruct(anchorPairs);
    pairwiseAlignmentBandingParameters_destruct(p);
    return n;
}

/*
 * Production anchors from an exonerate cigar line (signalMachine.c:665-752, 429):
 * cigarRead -> signalUtils_guideAlignmentToRebasedAnchorPairs -> signalUtils_getRemappedAnchorPairs.
 * info[0..5] = start1, end1, strand1, start2, end2, strand2 (as read, before rebasing).
 */
int64_t ref_cigar_anchors(const char *cigarFile, int64_t trim, const int64_t *eventMap, int64_t *info,
                          int64_t *out, int64_t cap) {
    FILE *fh = fopen(cigarFile, "r");
    if (!fh) return -1;
    struct PairwiseAlignment *pA = cigarRead(fh);
    fclose(fh);
    if (!pA) return -1;
    info[0] = pA->start1; info[1] = pA->end1; info[2] = pA->strand1;
    info[3] = pA->start2; info[4] = pA->end2; info[5] = pA->strand2;
    PairwiseAlignmentParameters *p = pairwiseAlignmentBandingParameters_construct();
    p->constraintDiagonalTrim = trim;
    stList *anchorPairs = signalUtils_guideAlignmentToRebasedAnchorPairs(pA, p);
    stList *filtered = signalUtils_getRemappedAnchorPairs(anchorPairs, (int64_t *) eventMap, info[3]);
    int64_t n = stList_length(filtered);
    for (int64_t i = 0; i < n && i < cap; i++) {
        stIntTuple *t = stList_get(filtered, i);
        out[2 * i] = stIntTuple_get(t, 0); out[2 * i + 1] = stIntTuple_get(t, 1);
    }
    stList_destruct(filtered);
    stList_destruct(anchorPairs);
    pairwiseAlignmentBandingParameters_destruct(p);
    destructPairwiseAlignment(pA);
    return n;
}
