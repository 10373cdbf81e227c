/* CPU oracle, C part -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * (1) The deterministic counter-based synthetic-data generator (Philox4x32-10 +
 *     AS241 inverse normal with an exactly reproducible log), restated on the host
 *     so that tests can check the device generator bit for bit.  The model is the
 *     one the reference simulates (Package/PPLS/src/loglC.cpp:288-311 simulC with the
 *     orthogonal parts zero as in Package/EM_Cpp.R:24-25,42), generalised to r
 *     components.
 * (2) A fused single-pass restatement of the data-dependent part of one EM
 *     iteration of PPLS_simult (Package/PPLS/R/EM_W_multi.R:689-694 projections and
 *     posterior means, :732-733 t(X)%*%mu_T / t(Y)%*%mu_U, and the sums of
 *     Package/PPLS/src/loglC.cpp:332-335), with optional "virtual ranks" that shard
 *     rows and add the partial sums in rank order (validates the sharded
 *     arithmetic without GPUs).  OpenMP; also the "best-effort fused CPU" baseline.
 *
 * PARITY UNPINNED (see oracle/ppls_ref.py header).  Only tests/, smoke() and the
 * cpu_baseline / --impl reference legs of bench.py may load this library.
 *
 * Build: make -C oracle   (gcc -O2 -ffp-contract=off -fopenmp -shared -fPIC)
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>
#ifdef _OPENMP
#include <omp.h>
#endif

/* ------------------------------------------------------------------ Philox4x32-10 */
static inline void philox4x32_10(uint32_t c[4], uint32_t k0, uint32_t k1) {
  for (int round = 0; round < 10; ++round) {
    uint64_t p0 = (uint64_t)0xD2511F53u * c[0];
    uint64_t p1 = (uint64_t)0xCD9E8D57u * c[2];
    uint32_t n0 = (uint32_t)(p1 >> 32) ^ c[1] ^ k0;
    uint32_t n1 = (uint32_t)p1;
    uint32_t n2 = (uint32_t)(p0 >> 32) ^ c[3] ^ k1;
    uint32_t n3 = (uint32_t)p0;
    c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
}

/* log(x) for normal positive x from + - * / only (no libm, no FMA): identical bits on
 * any IEEE-754 machine.  x = m 2^e, m in [sqrt(.5), sqrt(2)); log m = 2 atanh((m-1)/(m+1)). */
static inline double det_log(double x) {
  uint64_t bits; memcpy(&bits, &x, 8);
  int e = (int)((bits >> 52) & 0x7FF) - 1022;
  bits = (bits & 0x000FFFFFFFFFFFFFull) | 0x3FE0000000000000ull;   /* m in [0.5,1) */
  double m; memcpy(&m, &bits, 8);
  if (m < 0.70710678118654752440) { m = m * 2.0; e -= 1; }
  double s = (m - 1.0) / (m + 1.0);
  double s2 = s * s;
  double P = 1.0 / 27.0;
  P = P * s2 + 1.0 / 25.0; P = P * s2 + 1.0 / 23.0; P = P * s2 + 1.0 / 21.0;
  P = P * s2 + 1.0 / 19.0; P = P * s2 + 1.0 / 17.0; P = P * s2 + 1.0 / 15.0;
  P = P * s2 + 1.0 / 13.0; P = P * s2 + 1.0 / 11.0; P = P * s2 + 1.0 / 9.0;
  P = P * s2 + 1.0 / 7.0;  P = P * s2 + 1.0 / 5.0;  P = P * s2 + 1.0 / 3.0;
  P = P * s2 + 1.0;
  return 2.0 * s * P + (double)e * 0.69314718055994530942;
}

/* Wichura AS241 PPND16 inverse normal CDF, Horner with separate multiply and add. */
static inline double det_ndtri(double u) {
  double q = u - 0.5, r, num, den, z;
  if (fabs(q) <= 0.425) {
    r = 0.180625 - q * q;
    num = 2.5090809287301226727e+3;
    num = num * r + 3.3430575583588128105e+4; num = num * r + 6.7265770927008700853e+4;
    num = num * r + 4.5921953931549871457e+4; num = num * r + 1.3731693765509461125e+4;
    num = num * r + 1.9715909503065514427e+3; num = num * r + 1.3314166789178437745e+2;
    num = num * r + 3.3871328727963666080e0;
    den = 5.2264952788528545610e+3;
    den = den * r + 2.8729085735721942674e+4; den = den * r + 3.9307895800092710610e+4;
    den = den * r + 2.1213794301586595867e+4; den = den * r + 5.3941960214247511077e+3;
    den = den * r + 6.8718700749205790830e+2; den = den * r + 4.2313330701600911252e+1;
    den = den * r + 1.0;
    return q * num / den;
  }
  r = (q < 0.0) ? u : 1.0 - u;
  r = sqrt(-det_log(r));
  if (r <= 5.0) {
    r = r - 1.6;
    num = 7.74545014278341407640e-4;
    num = num * r + 2.27238449892691845833e-2; num = num * r + 2.41780725177450611770e-1;
    num = num * r + 1.27045825245236838258e0;  num = num * r + 3.64784832476320460504e0;
    num = num * r + 5.76949722146069140550e0;  num = num * r + 4.63033784615654529590e0;
    num = num * r + 1.42343711074968357734e0;
    den = 1.05075007164441684324e-9;
    den = den * r + 5.47593808499534494600e-4; den = den * r + 1.51986665636164571966e-2;
    den = den * r + 1.48103976427480074590e-1; den = den * r + 6.89767334985100004550e-1;
    den = den * r + 1.67638483018380384940e0;  den = den * r + 2.05319162663775882187e0;
    den = den * r + 1.0;
  } else {
    r = r - 5.0;
    num = 2.01033439929228813265e-7;
    num = num * r + 2.71155556874348757815e-5; num = num * r + 1.24266094738807843860e-3;
    num = num * r + 2.65321895265761230930e-2; num = num * r + 2.96560571828504891230e-1;
    num = num * r + 1.78482653991729133580e0;  num = num * r + 5.46378491116411436990e0;
    num = num * r + 6.65790464350110377720e0;
    den = 2.04426310338993978564e-15;
    den = den * r + 1.42151175831644588870e-7; den = den * r + 1.84631831751005468180e-5;
    den = den * r + 7.86869131145613259100e-4; den = den * r + 1.48753612908506148525e-2;
    den = den * r + 1.36929880922735805310e-1; den = den * r + 5.99832206555887937690e-1;
    den = den * r + 1.0;
  }
  z = num / den;
  return (q < 0.0) ? -z : z;
}

/* One standard normal per (seed, stream, row, col).  Streams: 0 Z_T, 1 Z_H, 2 Z_E,
 * 3 Z_F, 10/11 start-value noise for W/C (+ 16*s for multi-start s). */
double ppls_gen_normal(uint64_t seed, uint32_t stream, uint64_t row, uint32_t col) {
  uint32_t c[4] = {(uint32_t)row, (uint32_t)(row >> 32), col, stream};
  philox4x32_10(c, (uint32_t)seed, (uint32_t)(seed >> 32));
  double u = ((double)(c[0] >> 5) * 67108864.0 + (double)(c[1] >> 6) + 0.5) / 9007199254740992.0;
  return det_ndtri(u);
}

/* One uniform in (0, 1) from the same counter (the draw the normal above is made from). */
double ppls_gen_uniform(uint64_t seed, uint32_t stream, uint64_t row, uint32_t col) {
  uint32_t c[4] = {(uint32_t)row, (uint32_t)(row >> 32), col, stream};
  philox4x32_10(c, (uint32_t)seed, (uint32_t)(seed >> 32));
  return ((double)(c[0] >> 5) * 67108864.0 + (double)(c[1] >> 6) + 0.5) / 9007199254740992.0;
}

void ppls_gen_normals(uint64_t seed, uint32_t stream, uint64_t row0, int64_t nrows, int32_t ncols,
                      double* out /* nrows x ncols column-major */) {
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < nrows; ++i)
    for (int32_t j = 0; j < ncols; ++j)
      out[i + (int64_t)j * nrows] = ppls_gen_normal(seed, stream, row0 + (uint64_t)i, (uint32_t)j);
}

/* Synthetic PPLS data, rows row0 .. row0+nrows-1 of the global sample:
 *   T = Z_T diag(t),  U = T diag(b) + sigH Z_H,  X = T W' + sigE Z_E,  Y = U C' + sigF Z_F.
 * W (p x r), C (q x r), X (ldX x p), Y (ldY x q): column-major. */
void ppls_gen_data(uint64_t seed, uint64_t row0, int64_t nrows, int32_t p, int32_t q, int32_t r,
                   const double* W, const double* C, const double* b, const double* t,
                   double sigE, double sigF, double sigH,
                   double* X, int64_t ldX, double* Y, int64_t ldY) {
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < nrows; ++i) {
    double T[16], U[16];
    uint64_t row = row0 + (uint64_t)i;
    for (int k = 0; k < r; ++k) {
      T[k] = t[k] * ppls_gen_normal(seed, 0u, row, (uint32_t)k);
      U[k] = T[k] * b[k] + sigH * ppls_gen_normal(seed, 1u, row, (uint32_t)k);
    }
    for (int32_t j = 0; j < p; ++j) {
      double acc = 0.0;
      for (int k = 0; k < r; ++k) acc = acc + T[k] * W[j + (int64_t)k * p];
      X[i + (int64_t)j * ldX] = acc + sigE * ppls_gen_normal(seed, 2u, row, (uint32_t)j);
    }
    for (int32_t j = 0; j < q; ++j) {
      double acc = 0.0;
      for (int k = 0; k < r; ++k) acc = acc + U[k] * C[j + (int64_t)k * q];
      Y[i + (int64_t)j * ldY] = acc + sigF * ppls_gen_normal(seed, 3u, row, (uint32_t)j);
    }
  }
}

/* The same sample with simulC's orthogonal parts (Package/PPLS/src/loglC.cpp:288-311):
 *   X = T W' + TYo P_Yosc' + sigE Z_E,  Y = U C' + UXo P_Xosc' + sigF Z_F,
 *   TYo = Z_To diag(sigTo) (stream 4), UXo = Z_Uo diag(sigUo) (stream 5).
 * P_Yosc (p x nx), P_Xosc (q x ny) column-major.  nx = ny = 0 gives ppls_gen_data's bits. */
void ppls_gen_data_orth(uint64_t seed, uint64_t row0, int64_t nrows, int32_t p, int32_t q, int32_t r,
                        int32_t nx, int32_t ny, const double* W, const double* C, const double* PYo,
                        const double* PXo, const double* b, const double* t, const double* sigTo,
                        const double* sigUo, double sigE, double sigF, double sigH,
                        double* X, int64_t ldX, double* Y, int64_t ldY) {
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < nrows; ++i) {
    double T[16], U[16], To[16], Uo[16];
    uint64_t row = row0 + (uint64_t)i;
    for (int k = 0; k < r; ++k) {
      T[k] = t[k] * ppls_gen_normal(seed, 0u, row, (uint32_t)k);
      U[k] = T[k] * b[k] + sigH * ppls_gen_normal(seed, 1u, row, (uint32_t)k);
    }
    for (int k = 0; k < nx; ++k) To[k] = sigTo[k] * ppls_gen_normal(seed, 4u, row, (uint32_t)k);
    for (int k = 0; k < ny; ++k) Uo[k] = sigUo[k] * ppls_gen_normal(seed, 5u, row, (uint32_t)k);
    for (int32_t j = 0; j < p; ++j) {
      double acc = 0.0;
      for (int k = 0; k < r; ++k) acc = acc + T[k] * W[j + (int64_t)k * p];
      if (nx > 0) {
        double ao = 0.0;
        for (int k = 0; k < nx; ++k) ao = ao + To[k] * PYo[j + (int64_t)k * p];
        acc = acc + ao;
      }
      X[i + (int64_t)j * ldX] = acc + sigE * ppls_gen_normal(seed, 2u, row, (uint32_t)j);
    }
    for (int32_t j = 0; j < q; ++j) {
      double acc = 0.0;
      for (int k = 0; k < r; ++k) acc = acc + U[k] * C[j + (int64_t)k * q];
      if (ny > 0) {
        double ao = 0.0;
        for (int k = 0; k < ny; ++k) ao = ao + Uo[k] * PXo[j + (int64_t)k * q];
        acc = acc + ao;
      }
      Y[i + (int64_t)j * ldY] = acc + sigF * ppls_gen_normal(seed, 3u, row, (uint32_t)j);
    }
  }
}

/* One fused pass over X and Y (the data-dependent part of one EM iteration).
 *   Xw = X W, Yc = Y C                                   EM_W_multi.R:689-690
 *   mu_T[:,k] = aT[k] Xw[:,k] + bT[k] Yc[:,k]            :691-692 (diagonal algebra collapsed)
 *   mu_U[:,k] = aU[k] Xw[:,k] + bU[k] Yc[:,k]            :693-694
 *   G1 = X' mu_T (p x r), G2 = Y' mu_U (q x r)           :732-733
 *   Gam = [Xw Yc]' [Xw Yc] (2r x 2r)                     every crossprod in :696-712 and the
 *                                                        sums of loglC.cpp:335 are entries of it
 *   ssq[0] = ||X||_F^2, ssq[1] = ||Y||_F^2               loglC.cpp:334
 * Rows are split into `nranks` contiguous virtual shards whose partial results are
 * added in rank order.  All matrices column-major; mu_T/mu_U (N x r) may be NULL. */
#define RB 128
void ppls_oracle_moments(const double* X, int64_t ldX, const double* Y, int64_t ldY,
                         int64_t N, int32_t p, int32_t q, int32_t r,
                         const double* W, const double* C,
                         const double* aT, const double* bT, const double* aU, const double* bU,
                         int32_t nranks, double* G1, double* G2, double* Gam, double* ssq,
                         double* mu_T, double* mu_U) {
  const int r2 = 2 * r;
  const size_t nacc = (size_t)(p + q) * r + (size_t)r2 * r2 + 2;
  memset(G1, 0, sizeof(double) * (size_t)p * r);
  memset(G2, 0, sizeof(double) * (size_t)q * r);
  memset(Gam, 0, sizeof(double) * (size_t)r2 * r2);
  ssq[0] = ssq[1] = 0.0;
  if (nranks < 1) nranks = 1;
  for (int rk = 0; rk < nranks; ++rk) {
    int64_t lo = N * rk / nranks, hi = N * (rk + 1) / nranks;
    int64_t nblk = (hi - lo + RB - 1) / RB;
    int nthr = 1;
#ifdef _OPENMP
    nthr = omp_get_max_threads();
#endif
    double* accs = (double*)calloc(nacc * (size_t)nthr, sizeof(double));
#pragma omp parallel num_threads(nthr)
    {
      int tid = 0;
#ifdef _OPENMP
      tid = omp_get_thread_num();
#endif
      double* a1 = accs + nacc * (size_t)tid;         /* G1 */
      double* a2 = a1 + (size_t)p * r;                /* G2 */
      double* ag = a2 + (size_t)q * r;                /* Gam */
      double* as = ag + (size_t)r2 * r2;              /* ssq */
      double* pr = (double*)malloc(sizeof(double) * RB * (size_t)r2 * 2);
      double* xw = pr;                                /* RB x 2r: [Xw Yc] */
      double* mu = pr + RB * (size_t)r2;              /* RB x 2r: [mu_T mu_U] */
#pragma omp for schedule(static)
      for (int64_t blk = 0; blk < nblk; ++blk) {
        int64_t i0 = lo + blk * RB;
        int nb = (int)((hi - i0 < RB) ? hi - i0 : RB);
        for (int k = 0; k < r2 * RB; ++k) xw[k] = 0.0;
        for (int32_t j = 0; j < p; ++j) {
          const double* xj = X + i0 + (int64_t)j * ldX;
          double s = 0.0;
          for (int i = 0; i < nb; ++i) s += xj[i] * xj[i];
          as[0] += s;
          for (int k = 0; k < r; ++k) {
            double w = W[j + (int64_t)k * p];
            double* d = xw + (size_t)k * RB;
            for (int i = 0; i < nb; ++i) d[i] += xj[i] * w;
          }
        }
        for (int32_t j = 0; j < q; ++j) {
          const double* yj = Y + i0 + (int64_t)j * ldY;
          double s = 0.0;
          for (int i = 0; i < nb; ++i) s += yj[i] * yj[i];
          as[1] += s;
          for (int k = 0; k < r; ++k) {
            double c = C[j + (int64_t)k * q];
            double* d = xw + (size_t)(r + k) * RB;
            for (int i = 0; i < nb; ++i) d[i] += yj[i] * c;
          }
        }
        for (int k = 0; k < r; ++k)
          for (int i = 0; i < nb; ++i) {
            double a = xw[(size_t)k * RB + i], c = xw[(size_t)(r + k) * RB + i];
            mu[(size_t)k * RB + i] = aT[k] * a + bT[k] * c;
            mu[(size_t)(r + k) * RB + i] = aU[k] * a + bU[k] * c;
          }
        if (mu_T) for (int k = 0; k < r; ++k) for (int i = 0; i < nb; ++i) {
          mu_T[i0 + i + (int64_t)k * N] = mu[(size_t)k * RB + i];
          mu_U[i0 + i + (int64_t)k * N] = mu[(size_t)(r + k) * RB + i];
        }
        for (int a = 0; a < r2; ++a)
          for (int c = 0; c < r2; ++c) {
            double s = 0.0;
            for (int i = 0; i < nb; ++i) s += xw[(size_t)a * RB + i] * xw[(size_t)c * RB + i];
            ag[a + c * r2] += s;
          }
        for (int32_t j = 0; j < p; ++j) {
          const double* xj = X + i0 + (int64_t)j * ldX;
          for (int k = 0; k < r; ++k) {
            const double* m = mu + (size_t)k * RB;
            double s = 0.0;
            for (int i = 0; i < nb; ++i) s += xj[i] * m[i];
            a1[j + (size_t)k * p] += s;
          }
        }
        for (int32_t j = 0; j < q; ++j) {
          const double* yj = Y + i0 + (int64_t)j * ldY;
          for (int k = 0; k < r; ++k) {
            const double* m = mu + (size_t)(r + k) * RB;
            double s = 0.0;
            for (int i = 0; i < nb; ++i) s += yj[i] * m[i];
            a2[j + (size_t)k * q] += s;
          }
        }
      }
      free(pr);
    }
    /* thread partials in thread order -> this rank's partial; ranks in rank order */
    double* rsum = (double*)calloc(nacc, sizeof(double));
    for (int th = 0; th < nthr; ++th)
      for (size_t e = 0; e < nacc; ++e) rsum[e] += accs[nacc * (size_t)th + e];
    for (size_t e = 0; e < (size_t)p * r; ++e) G1[e] += rsum[e];
    for (size_t e = 0; e < (size_t)q * r; ++e) G2[e] += rsum[(size_t)p * r + e];
    for (size_t e = 0; e < (size_t)r2 * r2; ++e) Gam[e] += rsum[(size_t)(p + q) * r + e];
    ssq[0] += rsum[nacc - 2]; ssq[1] += rsum[nacc - 1];
    free(rsum); free(accs);
  }
}

int ppls_oracle_num_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

/* torchrun exports OMP_NUM_THREADS=1 to its workers; the CPU legs of bench.py restore the core count here */
void ppls_oracle_set_threads(int n) {
#ifdef _OPENMP
  if (n > 0) omp_set_num_threads(n);
#else
  (void)n;
#endif
}
