// Link stubs for the three est_hsq.cpp entry points that option.cpp references
// (option.cpp:624, 650, 655).  est_hsq.cpp needs real Eigen arithmetic and is not on
// the --make-grm path, so the oracle build leaves it out.  TEST INFRASTRUCTURE ONLY.
#include "gcta.h"
void gcta::fit_reml(string, string, string, string, string, string, string, string, string, int, double, double, int, bool, bool, bool, int, int, vector<double>, vector<double>, vector<int>, bool, double, bool, bool)
{ throw("Error: REML is not available in the oracle build."); }
void gcta::blup_snp_geno() { throw("Error: BLUP is not available in the oracle build."); }
void gcta::blup_snp_dosage() { throw("Error: BLUP is not available in the oracle build."); }
