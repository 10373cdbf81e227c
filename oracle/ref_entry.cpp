// C entry points over the reference's OWN native functions -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// oracle/Makefile (target `ref`) compiles /root/reference/Package/PPLS/src/loglC.cpp UNMODIFIED, from where it
// lies, against the stand-in headers in oracle/ref_shim/ (Eigen, Rcpp and R are not installed here), and links it
// with this file into oracle/_ref/libppls_ref.so.  The declarations below repeat the reference's signatures
// (loglC.cpp:37-39, :117-119, :269-270, :319-321, :341-342, :400-401, :452) so the linker binds to its code.
// Only tests/ (and, if ever wanted, bench.py's CPU legs) may load the library; nothing under ppls_b200/ does.
// List results are kept in a per-thread "last result" and fetched by key (column-major doubles).
#include <RcppEigen.h>
#include <cstring>
#include <string>

using Eigen::MatrixXd;
using Eigen::VectorXd;
using Rcpp::List;

List loglC(MatrixXd W, MatrixXd C, MatrixXd P_Yosc, MatrixXd P_Xosc, MatrixXd B_T, MatrixXd Dat, double sigX, double sigY,
           double sigH, MatrixXd sigT, MatrixXd sigTo, MatrixXd sigUo, double c1, double c2, double c3);
List EMstepC(MatrixXd W, MatrixXd C, MatrixXd P_Yosc, MatrixXd P_Xosc, MatrixXd B_T, MatrixXd Dat, double sigX, double sigY,
             double sigH, MatrixXd sigT, MatrixXd sigTo, MatrixXd sigUo, MatrixXd invS);
MatrixXd simulC(int N, MatrixXd W, MatrixXd C, MatrixXd P_Yosc, MatrixXd P_Xosc, MatrixXd B_T, double sigX, double sigY,
                double sigH, MatrixXd sigT, MatrixXd sigTo, MatrixXd sigUo);
double loglC_fast(MatrixXd W, MatrixXd C, MatrixXd X, MatrixXd Y, double sigX, double sigY, VectorXd sig2T, VectorXd c1,
                  VectorXd c2, VectorXd c3, VectorXd Kc);
List EMstepC_fast(VectorXd W, VectorXd C, double B, MatrixXd X, MatrixXd Y, double sigX, double sigY, double sigH, double sigT,
                  double c1, double c2, double c3);
List meta_Estep(VectorXd W, VectorXd C, double B, MatrixXd X, MatrixXd Y, double sigX, double sigY, double sigH, double sigT,
                double c1, double c2, double c3);
List meta_Mstep(List ret);

namespace {
thread_local List g_last;
thread_local std::string g_err;

MatrixXd mat(const double* a, long rows, long cols) {     // column-major, leading dimension = rows
  MatrixXd m(rows, cols);
  if (rows * cols > 0) std::memcpy(m.data(), a, sizeof(double) * (size_t)(rows * cols));
  return m;
}
VectorXd vec(const double* a, long n) { return VectorXd(mat(a, n, 1)); }
template <class F>
int guarded(F f) {
  try { f(); return 0; }
  catch (const std::exception& e) { g_err = e.what(); return -1; }
  catch (...) { g_err = "unknown C++ exception"; return -1; }
}
}  // namespace

extern "C" {

const char* ppls_ref_error(void) { return g_err.c_str(); }
const char* ppls_ref_source(void) { return "Package/PPLS/src/loglC.cpp (unmodified) + oracle/ref_shim stand-in headers"; }
void ppls_ref_seed(uint64_t seed) { Rcpp::shim_rng().seed(seed); }

// number of doubles of element `key` of the last List result (-1: no such element)
long ppls_ref_last_len(const char* key) {
  if (!g_last.has(key)) return -1;
  const Rcpp::Value& v = g_last[key];
  return v.is_matrix() ? static_cast<Eigen::Mat>(v).size() : 1;
}
int ppls_ref_last_get(const char* key, double* out) {
  if (!g_last.has(key)) { g_err = std::string("no list element ") + key; return -1; }
  const Eigen::Mat m = static_cast<Eigen::Mat>(g_last[key]);
  if (m.size() > 0) std::memcpy(out, m.data(), sizeof(double) * (size_t)m.size());
  return 0;
}

// loglC_fast, loglC.cpp:319-338
int ppls_ref_loglC_fast(const double* W, const double* C, const double* X, const double* Y, long N, long p, long q, long a,
                        double sigX, double sigY, const double* sig2T, const double* c1, const double* c2, const double* c3,
                        const double* Kc, double* out) {
  return guarded([&] {
    *out = loglC_fast(mat(W, p, a), mat(C, q, a), mat(X, N, p), mat(Y, N, q), sigX, sigY, vec(sig2T, a), vec(c1, a), vec(c2, a),
                      vec(c3, a), vec(Kc, a));
  });
}

// EMstepC_fast, loglC.cpp:341-397 (one component).  Result: ppls_ref_last_get("mu_T" | "mu_U" | "W" | "C" | "B" |
// "sighat" | "siglathat" | "Cut" | "Ctt" | "Cuu" | "Cee" | "Cff" | "Chh").
int ppls_ref_EMstepC_fast(const double* W, const double* C, double B, const double* X, const double* Y, long N, long p, long q,
                          double sigX, double sigY, double sigH, double sigT, double c1, double c2, double c3) {
  return guarded([&] { g_last = EMstepC_fast(vec(W, p), vec(C, q), B, mat(X, N, p), mat(Y, N, q), sigX, sigY, sigH, sigT, c1, c2, c3); });
}

// meta_Estep, loglC.cpp:400-449, then meta_Mstep, :452-474, on its result when `mstep` != 0
int ppls_ref_meta_step(const double* W, const double* C, double B, const double* X, const double* Y, long N, long p, long q,
                       double sigX, double sigY, double sigH, double sigT, double c1, double c2, double c3, int mstep) {
  return guarded([&] {
    List e = meta_Estep(vec(W, p), vec(C, q), B, mat(X, N, p), mat(Y, N, q), sigX, sigY, sigH, sigT, c1, c2, c3);
    g_last = mstep ? meta_Mstep(e) : e;
  });
}

// loglC, loglC.cpp:37-113 (dense covariance + Cholesky when c1 = c2 = c3 = 0).  Result: "value", "invS".
int ppls_ref_loglC(const double* W, const double* C, const double* PYo, const double* PXo, const double* B_T, const double* Dat,
                   long N, long p, long q, long a, long nx, long ny, double sigX, double sigY, double sigH, const double* sigT,
                   const double* sigTo, const double* sigUo, double c1, double c2, double c3) {
  return guarded([&] {
    g_last = loglC(mat(W, p, a), mat(C, q, a), mat(PYo, p, nx), mat(PXo, q, ny), mat(B_T, a, a), mat(Dat, N, p + q), sigX, sigY,
                   sigH, mat(sigT, a, a), mat(sigTo, nx, nx), mat(sigUo, ny, ny), c1, c2, c3);
  });
}

// EMstepC, loglC.cpp:117-250 (dense E-step).  Result: "Cxt", "Ctt", "Cyu", "Cuu", "Cut", "diag_Cee", "diag_Cff", "Chh", ...
int ppls_ref_EMstepC(const double* W, const double* C, const double* PYo, const double* PXo, const double* B_T,
                     const double* Dat, long N, long p, long q, long a, long nx, long ny, double sigX, double sigY, double sigH,
                     const double* sigT, const double* sigTo, const double* sigUo) {
  return guarded([&] {
    g_last = EMstepC(mat(W, p, a), mat(C, q, a), mat(PYo, p, nx), mat(PXo, q, ny), mat(B_T, a, a), mat(Dat, N, p + q), sigX, sigY,
                     sigH, mat(sigT, a, a), mat(sigTo, nx, nx), mat(sigUo, ny, ny), MatrixXd::Zero(1, 1));
  });
}

// simulC, loglC.cpp:269-315 (normals from the stand-in rnorm; see ref_shim/RcppEigen.h).  Dat: N x (p+q) column-major.
int ppls_ref_simulC(long N, const double* W, const double* C, const double* PYo, const double* PXo, const double* B_T, long p,
                    long q, long a, long nx, long ny, double sigX, double sigY, double sigH, const double* sigT,
                    const double* sigTo, const double* sigUo, double* Dat) {
  return guarded([&] {
    MatrixXd D = simulC((int)N, mat(W, p, a), mat(C, q, a), mat(PYo, p, nx), mat(PXo, q, ny), mat(B_T, a, a), sigX, sigY, sigH,
                        mat(sigT, a, a), mat(sigTo, nx, nx), mat(sigUo, ny, ny));
    std::memcpy(Dat, D.data(), sizeof(double) * (size_t)D.size());
  });
}

}  // extern "C"
