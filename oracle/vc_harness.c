/* TEST INFRASTRUCTURE ONLY.  Flat entry point around the reference's OWN variantCaller writer
 * (writePosteriorProbsVC, impl/signalMachine.c:161-232, compiled unmodified as clib_signalMachine.o): builds the
 * stList of stIntTuple4(p, x, y, (int64_t) kmer*) the writer consumes from plain arrays and calls it.  Used by
 * tests/golden/make_variants_golden.py to produce the rows our row generator is pinned against. */
#include <stdlib.h>
#include <string.h>
#include "stateMachine.h"
#include "pairwiseAligner.h"
#include "sonLib.h"

void writePosteriorProbsVC(char *posteriorProbsFile, char *readLabel, StateMachine *sM, char *target, bool forward,
                           int64_t eventSequenceOffset, int64_t referenceSequenceOffset, stList *alignedPairs,
                           Strand strand, double posteriorScore, bool rna, char *contig);

/* kmers: nPairs strings of kmerLength characters, concatenated (no terminators) */
int ref_write_vc(const char *path, const char *readLabel, int64_t kmerLength, const char *target, int forward,
                 int64_t eventSequenceOffset, int64_t referenceSequenceOffset, int64_t nPairs, const int64_t *p,
                 const int64_t *x, const int64_t *y, const char *kmers, int strandIsComplement, double posteriorScore,
                 int rna, const char *contig) {
    StateMachine *sM = calloc(1, sizeof(StateMachine3_HDP) + 4096);   /* only sM->kmerLength is read by the writer */
    sM->kmerLength = kmerLength;
    char *arena = malloc((size_t) nPairs * (kmerLength + 1) + 1);
    stList *pairs = stList_construct3(0, (void (*)(void *)) stIntTuple_destruct);
    for (int64_t i = 0; i < nPairs; i++) {
        char *k = arena + i * (kmerLength + 1);
        memcpy(k, kmers + i * kmerLength, (size_t) kmerLength);
        k[kmerLength] = '\0';
        stList_append(pairs, stIntTuple_construct4(p[i], x[i], y[i], (int64_t) k));
    }
    writePosteriorProbsVC((char *) path, (char *) readLabel, sM, (char *) target, forward != 0, eventSequenceOffset,
                          referenceSequenceOffset, pairs, strandIsComplement ? complement : template, posteriorScore,
                          rna != 0, (char *) contig);
    stList_destruct(pairs);
    free(arena);
    free(sM);
    return 0;
}
