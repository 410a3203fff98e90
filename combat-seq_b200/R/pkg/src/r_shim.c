/* r_shim.c — .Call bridge from R to the C ABI of libcombatseq.so (include/combatseq.h).
 * Replaces the in-process R body of ComBat_seq() (/root/reference/ComBat_seq.R:33-203); no logic lives here.
 * Build where R exists:  R CMD INSTALL combat-seq_b200/R/pkg   (src/Makevars links libcombatseq.so), or
 *                        R CMD SHLIB r_shim.c -I../../../../include -L../../.. -lcombatseq -o cbs_r_shim.so
 * NOT COMPILED AGAINST R in the build environment (no R there); tests/test_abi_host.py compiles it against a minimal
 * stand-in of R's C API (tests/r_stub/) so that at least the C and the use of include/combatseq.h are checked.  R is single-threaded: the R API is only touched
 * on the calling thread, nothing longjmps through C++ frames (cbs_* never throw; the error text is copied before
 * Rf_error), and the one allocation is PROTECTed. */
#include <R.h>
#include <Rinternals.h>
#include <string.h>
#include "combatseq.h"

SEXP cbs_R_combat_seq(SEXP counts, SEXP batch_idx, SEXP n_batch, SEXP mod, SEXP shrink, SEXP shrink_disp,
                      SEXP gene_subset_n, SEXP mc_stable, SEXP seed, SEXP n_gpus) {
    SEXP dim = Rf_getAttrib(counts, R_DimSymbol);
    SEXP mdim = Rf_getAttrib(mod, R_DimSymbol);
    cbs_input in;
    cbs_output out;
    memset(&in, 0, sizeof in);
    memset(&out, 0, sizeof out);
    in.G0 = INTEGER(dim)[0];
    in.N = INTEGER(dim)[1];
    in.B = Rf_asInteger(n_batch);
    in.p_mod = INTEGER(mdim)[1];
    in.counts_f64 = REAL(counts);            /* column-major double: R's own layout, no copy */
    in.batch_idx = INTEGER(batch_idx);
    in.mod = REAL(mod);
    in.shrink = Rf_asLogical(shrink);
    in.shrink_disp = Rf_asLogical(shrink_disp);
    in.gene_subset_n = Rf_asInteger(gene_subset_n);
    in.seed = (uint64_t)Rf_asReal(seed);
    in.mc_mode = Rf_asInteger(mc_stable) ? CBS_MC_STABLE : CBS_MC_REFERENCE;
    in.device = -1;
    in.n_gpus = Rf_asInteger(n_gpus);        /* 0 = every visible GPU: one R process drives them all (single-process sharding) */
    SEXP ans = PROTECT(Rf_allocMatrix(REALSXP, in.G0, in.N));
    out.adjusted = REAL(ans);
    int rc = cbs_combat_seq(&in, &out);
    if (rc != CBS_OK) {
        char msg[1024];
        strncpy(msg, cbs_last_error(), sizeof msg - 1);
        msg[sizeof msg - 1] = 0;
        UNPROTECT(1);
        Rf_error("%s", msg);                  /* same wording as the reference's stop() for the shared conditions */
    }
    UNPROTECT(1);
    return ans;
}
