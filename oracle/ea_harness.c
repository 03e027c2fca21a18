/* TEST INFRASTRUCTURE ONLY.  Flat entry point around the reference's OWN adaptive banded event -> kmer alignment
 * (adaptive_banded_simple_event_align2, impl/eventAligner.c:899-1235; scalings by estimate_scalings_using_mom :792-842),
 * compiled UNMODIFIED against oracle/hdf5_stub/hdf5.h (HDF5 is not in this image; none of the fast5 functions of that file
 * is ever called).  Used by tests/golden/make_event_align_golden.py. */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <hdf5.h>
#include "eventAligner.h"
#include "signalMachineUtils.h"
#include "sonLib.h"

/* useMom != 0: scale / shift by the reference's method of moments (var = 1), else the given ones.
 * scalingsOut: scale, shift, var as the alignment used them.  Returns the number of aligned (kmer, event) pairs (0 when the
 * reference's QC rejects the alignment), -1 when cap is too small. */
int64_t ref_event_align(const char *modelFile, const char *sequence, int rna, int64_t nEvents, const float *mean, const float *stdv,
                        int useMom, double scale, double shift, double var, int64_t cap, int64_t *kmerIdx, int64_t *eventIdx,
                        double *scalingsOut) {
    StateMachine *sM = stateMachine3_loadFromFile(modelFile, threeState, emissions_kmer_getGapProb,
                                                  emissions_signal_strawManGetKmerEventMatchProbWithDescaling_MeanOnly,
                                                  stateMachine3_loadTransitionsFromFile, NULL);
    stList *kmers = build_kmer_list(sequence, sM->kmerLength, rna != 0);
    event_table et;
    et.n = (size_t) nEvents; et.start = 0; et.end = (size_t) nEvents;
    et.event = calloc((size_t) nEvents + 1, sizeof(event_t));
    for (int64_t i = 0; i < nEvents; i++) { et.event[i].mean = mean[i]; et.event[i].stdv = stdv[i]; et.event[i].start = (uint64_t) i; et.event[i].length = 1.f; }
    if (useMom) {
        NanoporeReadAdjustmentParameters npp = estimate_scalings_using_mom(kmers, *sM, et);
        update_SignalMachineWithNanoporeParameters(npp, sM);
    } else {
        sM->scale = scale; sM->shift = shift; sM->var = var;
    }
    scalingsOut[0] = sM->scale; scalingsOut[1] = sM->shift; scalingsOut[2] = sM->var;
    stList *aln = adaptive_banded_simple_event_align2(et, sM, kmers, false);
    int64_t n = stList_length(aln);
    if (n > cap) n = -1;
    for (int64_t i = 0; i < n; i++) {
        struct AlignedPair *a = stList_get(aln, i);
        kmerIdx[i] = a->ref_pos; eventIdx[i] = a->read_pos;
    }
    free(et.event);
    return n;
}
