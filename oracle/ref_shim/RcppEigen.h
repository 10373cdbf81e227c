// Stand-in for <RcppEigen.h> / <Rcpp.h> -- TEST INFRASTRUCTURE, NOT PRODUCT CODE (and not Rcpp).
// Covers what /root/reference/Package/PPLS/src/loglC.cpp uses of Rcpp: List (string-keyed, insertion order,
// holding doubles and matrices), NumericVector / NumericMatrix with M(_, j) = v column assignment, rnorm(n),
// as<Map<MatrixXd>>(NumericMatrix) and RNGScope.  R's RNG cannot be reproduced outside R: rnorm() draws from
// std::mt19937_64 + Box-Muller, seeded through ppls_ref_seed() (oracle/ref_entry.cpp), so simulC is pinned
// in distribution and model structure only.
#ifndef PPLS_REF_SHIM_RCPPEIGEN
#define PPLS_REF_SHIM_RCPPEIGEN
#include <cmath>
#include <cstdint>
#include <random>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>
#include "Eigen/Dense"

namespace Rcpp {

class Value {
 public:
  Value() : is_mat_(false), d_(0.0) {}
  Value& operator=(double x) { is_mat_ = false; d_ = x; return *this; }
  Value& operator=(const Eigen::Mat& m) { is_mat_ = true; m_ = m; return *this; }
  operator double() const { if (is_mat_) throw std::runtime_error("Rcpp stand-in: list element is not a scalar"); return d_; }
  operator Eigen::Mat() const { if (!is_mat_) { Eigen::Mat m(1, 1); m(0, 0) = d_; return m; } return m_; }
  operator Eigen::VectorXd() const { return Eigen::VectorXd(static_cast<Eigen::Mat>(*this)); }
  bool is_matrix() const { return is_mat_; }
 private:
  bool is_mat_;
  double d_;
  Eigen::Mat m_;
};

class List {
 public:
  Value& operator[](const std::string& key) {
    for (auto& kv : items_) if (kv.first == key) return kv.second;
    items_.emplace_back(key, Value());
    return items_.back().second;
  }
  bool has(const std::string& key) const { for (auto& kv : items_) if (kv.first == key) return true; return false; }
  const std::vector<std::pair<std::string, Value>>& items() const { return items_; }
 private:
  std::vector<std::pair<std::string, Value>> items_;
};

struct Placeholder {};
static const Placeholder _ = Placeholder();

class NumericVector {
 public:
  explicit NumericVector(long n = 0) : v((size_t)n, 0.0) {}
  std::vector<double> v;
};
inline NumericVector operator*(double s, const NumericVector& a) { NumericVector r(a); for (double& x : r.v) x = s * x; return r; }

class NumericMatrix {
 public:
  NumericMatrix(long rows, long cols) : m(rows, cols) {}
  struct Column {
    Eigen::Mat* m; long j;
    Column& operator=(const NumericVector& x) {
      if ((long)x.v.size() != m->rows()) throw std::runtime_error("Rcpp stand-in: column assignment of the wrong length");
      for (long i = 0; i < m->rows(); ++i) (*m)(i, j) = x.v[(size_t)i];
      return *this;
    }
  };
  Column operator()(const Placeholder&, long j) { return Column{&m, j}; }
  Eigen::Mat m;
};

inline std::mt19937_64& shim_rng() { static std::mt19937_64 g(20261104u); return g; }
inline NumericVector rnorm(long n) {
  NumericVector r(n);
  std::uniform_real_distribution<double> U(0.0, 1.0);
  for (long i = 0; i < n; i += 2) {
    double u1 = U(shim_rng()), u2 = U(shim_rng());
    if (u1 < 1e-300) u1 = 1e-300;
    const double rad = std::sqrt(-2.0 * std::log(u1)), ang = 6.283185307179586476925 * u2;
    r.v[(size_t)i] = rad * std::cos(ang);
    if (i + 1 < n) r.v[(size_t)i + 1] = rad * std::sin(ang);
  }
  return r;
}

template <class T> struct AsImpl;
template <> struct AsImpl<Eigen::Map<Eigen::MatrixXd> > {
  static Eigen::Map<Eigen::MatrixXd> get(const NumericMatrix& x) { return Eigen::Map<Eigen::MatrixXd>(x.m); }
};
template <class T> T as(const NumericMatrix& x) { return AsImpl<T>::get(x); }

struct RNGScope {};

}  // namespace Rcpp
#endif
