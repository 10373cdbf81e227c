// Rcpp shim: R matrices -> C ABI (include/ppls_b200.h) -> R list.
// Replaces, for the PPLS_simult loop, the generated glue of the reference:
//   RcppExport SEXP PPLS_loglC_fast(SEXP x11)   Package/PPLS/src/RcppExports.cpp:79-99
//   .Call('PPLS_loglC_fast', PACKAGE='PPLS',..) Package/PPLS/R/RcppExports.R:32-34
// Same conventions as that boundary: inputs are R-owned and never mutated, the result is a fresh
// R-owned list, C++/engine failures become R errors (Rcpp::stop inside BEGIN_RCPP/END_RCPP),
// called from R's main thread only.  No numerics here: marshalling only.
//
// Build (where R + Rcpp exist; not in this repository's containers):
//   PKG_CPPFLAGS = -I<repo>/include ; PKG_LIBS = -L<repo>/ppls_b200 -lppls_b200 -Wl,-rpath,<repo>/ppls_b200
#include <Rcpp.h>
#include <string>
#include <vector>
#include "ppls_b200.h"

namespace {
// A device-resident dataset: context (one GPU, or several GPUs of one node driven from this R process:
// ppls_ctx_create_multi) + the uploaded, re-tiled X and Y.  Held by R as an external pointer (ppls_b200_open) so that
// repeated PPLS_simult_b200 calls on the same X, Y do not pay the host-to-device upload again; released by the
// finalizer when R collects the pointer, or at once by ppls_b200_close.
struct Handles {
  ppls_ctx* ctx = nullptr;
  ppls_data* data = nullptr;
  int64_t N = 0, p = 0, q = 0;
  void release() { if (data) ppls_data_destroy(data); if (ctx) ppls_ctx_destroy(ctx); data = nullptr; ctx = nullptr; }
  ~Handles() { release(); }
};
inline void check(int rc, const char* what) {
  if (rc != PPLS_OK) Rcpp::stop("%s failed (%d): %s", what, rc, ppls_last_error());
}
}  // namespace

// [[Rcpp::export]]
// devices: integer vector of CUDA device ids, e.g. 0L or 0:7 (rows of X and Y are then split over the GPUs).
RcppExport SEXP ppls_b200_open(SEXP Xs, SEXP Ys, SEXP devicess) {
BEGIN_RCPP
  Rcpp::NumericMatrix X(Xs), Y(Ys);                        // views on R memory (REALSXP), no copies
  Rcpp::IntegerVector dev(devicess);
  if (Y.nrow() != X.nrow()) Rcpp::stop("X and Y must have the same number of rows");
  if (dev.size() < 1) Rcpp::stop("need at least one device");
  Rcpp::XPtr<Handles> h(new Handles(), true);              // finalizer: delete -> ~Handles
  h->N = X.nrow(); h->p = X.ncol(); h->q = Y.ncol();
  std::vector<int> ids(dev.begin(), dev.end());
  check(ppls_ctx_create_multi(&h->ctx, ids.data(), (int)ids.size()), "ppls_ctx_create_multi");
  check(ppls_data_from_host(h->ctx, X.begin(), h->N, Y.begin(), h->N, h->N, h->p, h->q, &h->data), "ppls_data_from_host");
  return h;
END_RCPP
}

// [[Rcpp::export]]
RcppExport SEXP ppls_b200_close(SEXP hs) {
BEGIN_RCPP
  Rcpp::XPtr<Handles> h(hs);
  h->release();
  return R_NilValue;
END_RCPP
}

// [[Rcpp::export]]
RcppExport SEXP PPLS_simult_b200_run(SEXP hs, SEXP W0s, SEXP C0s, SEXP B0s, SEXP sigE0s, SEXP sigF0s,
                                     SEXP sigH0s, SEXP sigT0s, SEXP EMstepss, SEXP atols, SEXP types) {
BEGIN_RCPP
  Rcpp::XPtr<Handles> h(hs);
  if (!h->ctx || !h->data) Rcpp::stop("the ppls_b200 handle has been closed");
  Rcpp::NumericMatrix W0(W0s), C0(C0s);
  Rcpp::NumericVector B0(B0s), T0(sigT0s);
  const int64_t N = h->N, p = h->p, q = h->q;
  const int r = W0.ncol();
  if (W0.nrow() != p || C0.nrow() != q || C0.ncol() != r || B0.size() != r || T0.size() != r)
    Rcpp::stop("start values have the wrong shape");
  const int steps = Rcpp::as<int>(EMstepss);
  const std::string type = Rcpp::as<std::string>(types);

  std::vector<double> W0c(W0.begin(), W0.end()), C0c(C0.begin(), C0.end()), B0c(B0.begin(), B0.end()),
      T0c(T0.begin(), T0.end());                            // ppls_params wants non-const pointers
  ppls_params init{W0c.data(), C0c.data(), B0c.data(), T0c.data(), Rcpp::as<double>(sigE0s),
                   Rcpp::as<double>(sigF0s), Rcpp::as<double>(sigH0s)};
  ppls_opts opts{};
  opts.r = r; opts.max_steps = steps; opts.atol = Rcpp::as<double>(atols);
  opts.orth_type = (type == "QR") ? PPLS_ORTH_QR : PPLS_ORTH_SVD;

  Rcpp::NumericMatrix W(p, r), C(q, r), muT(N, r), muU(N, r), Chh(r, r);
  Rcpp::NumericVector B(r), sigT(r), loglik(steps), mom(3 * r + 2 + r * r);
  ppls_params est{W.begin(), C.begin(), B.begin(), sigT.begin(), 0.0, 0.0, 0.0};
  int32_t done = 0;
  check(ppls_simult_run(h->ctx, h->data, &opts, &init, &est, loglik.begin(), &done, mom.begin(), muT.begin(),
                        muU.begin()), "ppls_simult_run");
  Rcpp::NumericVector ll(loglik.begin(), loglik.begin() + done);
  for (int e = 0; e < r * r; ++e) Chh[e] = mom[3 * r + 2 + e];
  return Rcpp::List::create(
      Rcpp::Named("W") = W, Rcpp::Named("C") = C, Rcpp::Named("B") = B, Rcpp::Named("sigT") = sigT,
      Rcpp::Named("sigE") = est.sigE, Rcpp::Named("sigF") = est.sigF, Rcpp::Named("sigH") = est.sigH,
      Rcpp::Named("loglik") = ll, Rcpp::Named("mu_T") = muT, Rcpp::Named("mu_U") = muU,
      Rcpp::Named("Ctt") = Rcpp::NumericVector(mom.begin(), mom.begin() + r),
      Rcpp::Named("Cuu") = Rcpp::NumericVector(mom.begin() + r, mom.begin() + 2 * r),
      Rcpp::Named("Cut") = Rcpp::NumericVector(mom.begin() + 2 * r, mom.begin() + 3 * r),
      Rcpp::Named("Cee") = mom[3 * r], Rcpp::Named("Cff") = mom[3 * r + 1], Rcpp::Named("Chh") = Chh);
END_RCPP
}
