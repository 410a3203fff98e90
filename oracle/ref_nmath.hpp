// ORACLE — TEST INFRASTRUCTURE ONLY.  Never linked into, imported by, or executed from the product
// (combat-seq_b200/).  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
// legs may load liboracle.so.
//
// CPU restatement of the R nmath routines that /root/reference/helper_seq.R calls:
//   dnbinom (helper_seq.R:28), pnbinom (helper_seq.R:53), qnbinom (helper_seq.R:58).
// R itself is NOT under /root/reference and is not installed in this image; the algorithms below restate
// R 3.5.x src/nmath (dnbinom.c, dbinom.c, bd0.c, stirlerr.c, pnbinom.c, qnbinom.c, qnorm.c, toms708.c)
// from their published form.  PARITY UNPINNED against real R: the pins are mpmath/scipy known-answer
// vectors under tests/golden/ (see oracle/gen_golden.py).
#pragma once
#include <cmath>
#include <cfloat>

namespace ref {

double stirlerr(double n);                       // nmath/stirlerr.c
double bd0(double x, double np);                 // nmath/bd0.c
double dbinom_raw(double x, double n, double p, double q);   // nmath/dbinom.c (give_log = FALSE)
double dnbinom_mu(double x, double size, double mu);         // nmath/dnbinom.c
// regularised incomplete beta I_x(a,b) and its complement; y = 1-x is passed separately as in toms708.c
void   bratio(double a, double b, double x, double y, double* w, double* w1);
double pnbinom_mu(double q, double size, double mu);         // nmath/pnbinom.c  pnbinom_mu
double pnbinom_prob(double q, double size, double prob);     // nmath/pnbinom.c  pnbinom -> pbeta
double qnorm01(double p);                                    // nmath/qnorm.c (AS 241), lower tail
double qnbinom_prob(double p, double size, double prob);     // nmath/qnbinom.c (R 3.5: 64*eps fuzz + search)
double qnbinom_mu(double p, double size, double mu);         // nmath/qnbinom.c  qnbinom_mu
void   set_legacy_qnbinom_search(int on);                    // 1 = reproduce the R < 4.2 large-y search quirk
// independent cross-check for tests: direct long-double summation of the pmf recurrence
double pnbinom_mu_bysum(double q, double size, double mu);

}  // namespace ref
