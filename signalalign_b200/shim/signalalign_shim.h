/*
 * signalalign_shim.h — the two conversions of the drop-in shim that other reference-side code (signalMachineBatch)
 * re-uses: a live StateMachine -> device model handle (cached by content), and the reference's ambiguity stHash
 * (create_ambig_bases2, impl/pairwiseAligner.c:68-92) -> sa_ambig_t.  Compiled inside the reference tree.
 */
#ifndef SIGNALALIGN_SHIM_H
#define SIGNALALIGN_SHIM_H
#include "signalalign_b200.h"
#include "stateMachine.h"
#include "sonLib.h"

/* threeState / threeStateHdp machines only; aborts through st_errAbort on anything else.  The handle stays owned by the
 * shim's cache (do not free it). */
sa_model_t *sa_shim_model(StateMachine *sM);
/* NULL in -> NULL out (= the built-in table); the caller frees the result with sa_ambig_free */
sa_ambig_t *sa_shim_ambig(stHash *ambigBases);
#endif
