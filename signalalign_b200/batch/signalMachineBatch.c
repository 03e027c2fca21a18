/*
 * signalMachineBatch — the batch driver that turns kernel throughput into reads per second.
 *
 * The reference launches ONE `signalMachine` process per read (src/signalalign/signalAlignment.py:435-507, worker pool
 * :740-848): every read pays process start, the .model / .nhdp parse and one tiny alignment call.  This driver takes a
 * MANIFEST of reads, parses the model once, prepares every read exactly as impl/signalMachine.c:main does (:650-760,
 * :848-868: guide cigar -> anchors, .npRead, reference slice, event slice, scale/shift/var re-estimation with drift
 * adjustment), aligns them in batches through ONE sa_gpu_align_batch call each, and writes every read's output file
 * with the reference's own writers (outputAlignment -> writePosteriorProbsFull / writePosteriorProbsVC /
 * writeAssignments, impl/signalMachine.c:89-364), its stdout summary line (:917-923) and its stderr SUCCESS line (:937,
 * what signalAlignment.py:480-491 looks for).  Output files are byte-identical to those of `signalMachine` linked
 * against the same library.
 *
 * Like the shim it is compiled INSIDE the reference tree: impl/signalMachine.c is compiled unmodified with
 * -Dmain=signalMachine_cli_main so that its writers and helpers link here (`make -C oracle cli`).
 *
 *   signalMachineBatch --manifest reads.tsv -T template.model [-v template.nhdp] [-s 0|1|2] [-x 50] [-D 0.01] [-m 14]
 *                      [-g 50] [-a ambig.tsv] [--batch 2048] [--rna] [--host-estimate]
 *   the per-read scale / shift / var / drift re-estimation (signalUtils_estimateNanoporeParams) runs on the device for the
 *   whole batch at once (sa_gpu_estimate_params_batch, bit-identical); --host-estimate keeps the reference's function
 *   manifest line:  npRead <TAB> guide cigar file <TAB> forward reference FASTA <TAB> sequence name <TAB> read label <TAB> output file
 *
 * 1-D (template strand) reads; 2-D complement alignment and the expectations pass stay with `signalMachine`.
 */
#include <getopt.h>
#include <string.h>
#include "fasta_handler.h"
#include "signalalign_shim.h"

/* from impl/signalMachine.c (compiled with -Dmain=signalMachine_cli_main) */
typedef enum { full = 0, variantCaller = 1, assignments = 2, both = 3 } OutputFormat;
void outputAlignment(OutputFormat fmt, char *posteriorProbsFile, char *readLabel, StateMachine *sM,
                     NanoporeReadAdjustmentParameters npp, double *events, char *target, bool forward, char *contig,
                     int64_t eventSequenceOffset, int64_t referenceSequenceOffset, stList *alignedPairs,
                     double posteriorScore, Strand strand, bool rna, char *posteriorProbsFile2);
Sequence *makeEventSequenceFromPairwiseAlignment(double *events, int64_t queryStart, int64_t queryEnd, int64_t *eventMap);
double scoreByPosteriorProbabilityIgnoringGaps(stList *alignedPairs);
StateMachine *buildStateMachine2(const char *modelFile, StateMachineType type, NanoporeHDP *nHdp);

#define ASSIGNMENT_THRESHOLD 0.1

typedef struct {
    char *npReadFile, *cigarFile, *fasta, *seqName, *label, *out;
    /* prepared */
    NanoporeRead *npRead;
    ReferenceSequence *R;
    Sequence *events;
    stList *anchors;         /* remapped (kmer, event) anchors */
    int64_t nGuideAnchors;
    int64_t tShift, rShift;
    bool forward;
    char *contig;
    double scale, shift, var;
    int64_t *flatAnchors;
} BatchRead;

static char *field(char **cursor) {
    char *s = *cursor, *e = s;
    if (!s) return NULL;
    while (*e && *e != '\t' && *e != '\n' && *e != '\r') e++;
    *cursor = (*e == '\t') ? e + 1 : NULL;
    *e = 0;
    return stString_copy(s);
}

/* everything impl/signalMachine.c:main does for one read before performSignalAlignment (:650-760, :848-861) */
static void prepareRead(BatchRead *b, StateMachine *sM, NanoporeHDP *nHdp, const double *pristineMatch,
                        const double *pristineGapY, PairwiseAlignmentParameters *p, bool rna, bool hostEstimate) {
    if (!stFile_exists(b->cigarFile)) st_errAbort("[signalMachineBatch]ERROR: no guide alignment %s\n", b->cigarFile);
    FILE *fh = fopen(b->cigarFile, "r");
    struct PairwiseAlignment *pA = cigarRead(fh);
    fclose(fh);
    const size_t tableBytes = sizeof(double) * (size_t) sM->parameterSetSize * MODEL_PARAMS;
    memcpy(sM->EMISSION_MATCH_MATRIX, pristineMatch, tableBytes);   /* a fresh process starts from the file's tables */
    memcpy(sM->EMISSION_GAP_Y_MATRIX, pristineGapY, tableBytes);
    b->npRead = nanopore_loadNanoporeReadFromFile(b->npReadFile, sM->alphabet, sM->alphabetSize);
    if (rna) {
        int64_t tmp = pA->start2;
        pA->start2 = b->npRead->templateReadLength - pA->end2;
        pA->end2 = b->npRead->templateReadLength - tmp;
    }
    b->R = fastaHandler_ReferenceSequenceConstructFull(b->fasta, NULL, pA, b->seqName, rna);
    b->events = makeEventSequenceFromPairwiseAlignment(b->npRead->templateEvents, pA->start2, pA->end2,
                                                       b->npRead->templateStrandEventMap);
    b->tShift = b->npRead->templateStrandEventMap[pA->start2];
    b->rShift = pA->start1;
    b->forward = pA->strand1;
    b->contig = stString_copy(pA->contig1);
    stList *guide = signalUtils_guideAlignmentToRebasedAnchorPairs(pA, p);
    b->nGuideAnchors = stList_length(guide);
    sM->scale = b->npRead->templateParams.scale;
    sM->shift = b->npRead->templateParams.shift;
    sM->var = b->npRead->templateParams.var;
    if (hostEstimate)
        signalUtils_estimateNanoporeParams(sM, b->npRead, &b->npRead->templateParams, ASSIGNMENT_THRESHOLD,
                                           signalUtils_templateOneDAssignmentsFromRead, nanopore_adjustTemplateEventsForDrift);
    if (sM->type == threeStateHdp) stateMachine3_setModelToHdpExpectedValues(sM, nHdp);
    b->scale = sM->scale; b->shift = sM->shift; b->var = sM->var;
    b->anchors = signalUtils_getRemappedAnchorPairs(guide, b->npRead->templateStrandEventMap, pA->start2);
    stList_destruct(guide);
    const int64_t nA = stList_length(b->anchors);
    b->flatAnchors = st_malloc(sizeof(int64_t) * 2 * (nA + 1));
    for (int64_t i = 0; i < nA; i++) {
        stIntTuple *a = stList_get(b->anchors, i);
        b->flatAnchors[2 * i] = stIntTuple_get(a, 0);
        b->flatAnchors[2 * i + 1] = stIntTuple_get(a, 1);
    }
    destructPairwiseAlignment(pA);
}

static void releaseRead(BatchRead *b) {
    nanopore_nanoporeReadDestruct(b->npRead);
    signalUtils_ReferenceSequenceDestruct(b->R);
    sequence_destruct(b->events);
    stList_destruct(b->anchors);
    free(b->flatAnchors);
    free(b->contig);
    b->npRead = NULL;
}

int main(int argc, char *argv[]) {
    int64_t diagExpansion = 50, constraintTrim = 14, traceBackDiagonals = 50, outFmt = 0, batchSize = 2048;
    double threshold = 0.01;
    bool rna = FALSE, hostEstimate = FALSE;
    char *manifest = NULL, *modelFile = NULL, *hdpFile = NULL, *ambigFile = NULL;
    static struct option longOptions[] = {
        {"manifest", required_argument, 0, 'M'}, {"templateModel", required_argument, 0, 'T'},
        {"templateHdp", required_argument, 0, 'v'}, {"sparse_output", required_argument, 0, 's'},
        {"diagonalExpansion", required_argument, 0, 'x'}, {"threshold", required_argument, 0, 'D'},
        {"constraintTrim", required_argument, 0, 'm'}, {"traceBackDiagonals", required_argument, 0, 'g'},
        {"ambig_model", required_argument, 0, 'a'}, {"batch", required_argument, 0, 'B'}, {"rna", no_argument, 0, 'r'},
        {"host-estimate", no_argument, 0, 'H'}, {0, 0, 0, 0}};
    int key;
    while ((key = getopt_long(argc, argv, "M:T:v:s:x:D:m:g:a:B:rH", longOptions, NULL)) != -1) {
        switch (key) {
            case 'M': manifest = stString_copy(optarg); break;
            case 'T': modelFile = stString_copy(optarg); break;
            case 'v': hdpFile = stString_copy(optarg); break;
            case 'a': ambigFile = stString_copy(optarg); break;
            case 's': outFmt = atoll(optarg); break;
            case 'x': diagExpansion = atoll(optarg); break;
            case 'D': threshold = atof(optarg); break;
            case 'm': constraintTrim = atoll(optarg); break;
            case 'g': traceBackDiagonals = atoll(optarg); break;
            case 'B': batchSize = atoll(optarg); break;
            case 'r': rna = TRUE; break;
            case 'H': hostEstimate = TRUE; break;
            default:
                fprintf(stderr, "usage: signalMachineBatch --manifest reads.tsv -T model [-v nhdp] [-s 0|1|2] [-x n] [-D thr] "
                                "[-m trim] [-g traceBack] [-a ambig] [--batch n] [--rna]\n");
                return 1;
        }
    }
    if (!manifest || !modelFile) st_errAbort("[signalMachineBatch]ERROR: need --manifest and -T\n");
    if (outFmt < 0 || outFmt > 2) st_errAbort("[signalMachineBatch]ERROR: output format 0 (full), 1 (variantCaller) or 2 (assignments)\n");
    if (batchSize < 1) batchSize = 1;

    /* alignment parameters and the model: once (impl/signalMachine.c:672-706) */
    PairwiseAlignmentParameters *p = pairwiseAlignmentBandingParameters_construct();
    p->threshold = threshold;
    p->constraintDiagonalTrim = constraintTrim;
    p->diagonalExpansion = diagExpansion % 2 == 0 ? diagExpansion : diagExpansion + 1;
    p->traceBackDiagonals = traceBackDiagonals;
    NanoporeHDP *nHdp = hdpFile ? deserialize_nhdp(hdpFile) : NULL;
    StateMachine *sM = buildStateMachine2(modelFile, hdpFile ? threeStateHdp : threeState, nHdp);
    const size_t tableBytes = sizeof(double) * (size_t) sM->parameterSetSize * MODEL_PARAMS;
    double *pristineMatch = st_malloc(tableBytes), *pristineGapY = st_malloc(tableBytes);
    memcpy(pristineMatch, sM->EMISSION_MATCH_MATRIX, tableBytes);
    memcpy(pristineGapY, sM->EMISSION_GAP_Y_MATRIX, tableBytes);
    stHash *ambigHash = create_ambig_bases2(ambigFile);
    sa_ambig_t *ambig = sa_shim_ambig(ambigHash);
    sa_params_t sp;
    sa_params_default(&sp);
    sp.threshold = p->threshold;
    sp.minDiagsBetweenTraceBack = p->minDiagsBetweenTraceBack;
    sp.traceBackDiagonals = p->traceBackDiagonals;
    sp.diagonalExpansion = p->diagonalExpansion;
    sp.splitMatrixBiggerThanThis = p->splitMatrixBiggerThanThis;

    /* manifest */
    stList *lines = stList_construct3(0, free);
    {
        FILE *fh = fopen(manifest, "r");
        if (!fh) st_errAbort("[signalMachineBatch]ERROR: cannot open manifest %s\n", manifest);
        char *line = NULL;
        size_t cap = 0;
        while (getline(&line, &cap, fh) >= 0)
            if (line[0] != '#' && line[0] != '\n' && line[0] != 0) stList_append(lines, stString_copy(line));
        free(line);
        fclose(fh);
    }
    const int64_t nReads = stList_length(lines);
    BatchRead *reads = st_calloc(batchSize, sizeof(BatchRead));
    sa_read_t *flat = st_calloc(batchSize, sizeof(sa_read_t));
    sa_model_t *model = NULL;
    /* the estimation reads the match table AS LOADED (before scaleNoise / the HDP's expected values touch it) */
    sa_model_t *estModel = hostEstimate ? NULL : sa_shim_model(sM);
    sa_strand_t *strands = st_calloc(batchSize, sizeof(sa_strand_t));
    sa_scaling_t *scalings = st_calloc(batchSize, sizeof(sa_scaling_t));
    int64_t done = 0;
    while (done < nReads) {
        const int64_t n = nReads - done < batchSize ? nReads - done : batchSize;
        for (int64_t i = 0; i < n; i++) {
            BatchRead *b = &reads[i];
            memset(b, 0, sizeof(*b));
            char *cursor = stList_get(lines, done + i);
            b->npReadFile = field(&cursor); b->cigarFile = field(&cursor); b->fasta = field(&cursor);
            b->seqName = field(&cursor); b->label = field(&cursor); b->out = field(&cursor);
            if (!b->out) st_errAbort("[signalMachineBatch]ERROR: manifest line %lld needs 6 tab-separated fields\n", (long long) (done + i + 1));
            prepareRead(b, sM, nHdp, pristineMatch, pristineGapY, p, rna, hostEstimate);
            strands[i].read = b->npRead->templateRead;
            strands[i].readLength = b->npRead->templateReadLength;
            strands[i].strandEventMap = b->npRead->templateStrandEventMap;
            strands[i].events = b->npRead->templateEvents;
            strands[i].nEvents = b->npRead->nbTemplateEvents;
            /* the device model: the emission the path reads is columns 0 / 1 of the tables (and the HDP), which the per-read
             * noise scaling never touches, so one upload serves every read */
            if (!model) model = sa_shim_model(sM);
        }
        if (!hostEstimate) {
            /* signalUtils_estimateNanoporeParams for the whole batch: parameters into every read's templateParams, event
             * means drift-adjusted in place (impl/signalMachineUtils.c:186-225) */
            if (sa_gpu_estimate_params_batch(estModel, strands, n, 1, scalings) != SA_OK) st_errAbort("libsagpu: %s\n", sa_last_error());
            for (int64_t i = 0; i < n; i++) {
                NanoporeReadAdjustmentParameters *q = &reads[i].npRead->templateParams;
                q->scale = scalings[i].scale; q->shift = scalings[i].shift; q->var = scalings[i].var;
                q->shift_sd = scalings[i].shift_sd; q->scale_sd = scalings[i].scale_sd; q->var_sd = scalings[i].var_sd;
                q->drift = scalings[i].drift;
                reads[i].scale = q->scale; reads[i].shift = q->shift; reads[i].var = q->var;
            }
        }
        for (int64_t i = 0; i < n; i++) {
            BatchRead *b = &reads[i];
            char *target = b->R->getTemplateTargetSequence(b->R);
            memset(&flat[i], 0, sizeof(sa_read_t));
            flat[i].target = target;
            flat[i].targetLength = (int64_t) strlen(target);
            flat[i].events = (const double *) b->events->elements;
            flat[i].nEvents = b->events->length;
            flat[i].anchors = b->flatAnchors;
            flat[i].nAnchors = stList_length(b->anchors);
            flat[i].scale = b->scale; flat[i].shift = b->shift; flat[i].var = b->var;
            flat[i].raggedLeft = 1; flat[i].raggedRight = 1;       /* performSignalAlignment, impl/signalMachine.c:436-437 */
        }
        sa_result_t *res = NULL;
        if (sa_gpu_align_batch(model, flat, n, &sp, ambig, &res) != SA_OK) st_errAbort("libsagpu: %s\n", sa_last_error());
        for (int64_t i = 0; i < n; i++) {
            BatchRead *b = &reads[i];
            /* the machine as the read's own process would hold it when it writes (noise columns: impl/stateMachine.c:721-741) */
            memcpy(sM->EMISSION_MATCH_MATRIX, pristineMatch, tableBytes);
            memcpy(sM->EMISSION_GAP_Y_MATRIX, pristineGapY, tableBytes);
            emissions_signal_scaleNoise(sM, b->npRead->templateParams);
            if (sM->type == threeStateHdp) stateMachine3_setModelToHdpExpectedValues(sM, nHdp);
            sM->scale = b->scale; sM->shift = b->shift; sM->var = b->var;
            stList *pairs = stList_construct3(0, (void (*)(void *)) stIntTuple_destruct);
            const int64_t lo = res->pairOffsets[i], hi = res->pairOffsets[i + 1], k = sM->kmerLength;
            char *arena = hi > lo ? st_malloc((size_t) (hi - lo) * (k + 1)) : NULL;
            for (int64_t q = lo; q < hi; q++) {
                char *km = arena + (q - lo) * (k + 1);
                sa_kmer_string(model, res->pairs[q].kmerIdx, km);
                stList_append(pairs, stIntTuple_construct4(res->pairs[q].p, res->pairs[q].x, res->pairs[q].y, (int64_t) km));
            }
            const double score = scoreByPosteriorProbabilityIgnoringGaps(pairs);
            stList_sort(pairs, sortByXPlusYCoordinate2);
            outputAlignment((OutputFormat) outFmt, b->out, b->label, sM, b->npRead->templateParams, b->npRead->templateEvents,
                            b->R->getTemplateTargetSequence(b->R), b->forward, b->contig, b->tShift, b->rShift, pairs, score,
                            template, rna, NULL);
            fprintf(stdout, "%s %" PRId64 "\t%" PRId64 "(%f)\t\n", b->label, b->nGuideAnchors, stList_length(pairs), score);
            fprintf(stderr, "signalAlign - SUCCESS: finished alignment of query %s, exiting\n", b->label);
            stList_destruct(pairs);
            free(arena);
            releaseRead(b);
        }
        sa_result_free(res);
        done += n;
    }
    fprintf(stderr, "signalMachineBatch - %" PRId64 " reads aligned\n", nReads);
    sa_ambig_free(ambig);
    return 0;
}
