// Stand-in for <Rcpp.h> (test infrastructure): everything lives in RcppEigen.h here.
#include "RcppEigen.h"
