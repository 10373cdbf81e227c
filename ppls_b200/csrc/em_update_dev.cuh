// Device code of the update step shared by K2 (em_update.cu) and by the fused tail of the streaming kernel
// (em_tail.cuh): closed-form coefficients, the scalar part of the update (log-likelihood, E-step moments, break
// rule, M-step scalars), the r x r inverse square root, and the receive side of the peer exchange.
// Every translation unit that includes this header is compiled with -fmad=false: the closed-form coefficients
// cancel catastrophically for small noise variances and must be evaluated exactly as the reference writes them
// (ppls_b200/build.py).
#pragma once
#include "common.cuh"
#include <math.h>

namespace ppls {

// Integer powers with one rounding (error-free products + a correction term): what a correctly
// rounded pow(x, n) -- R's and numpy's `^` / `**` go through libm pow -- returns in all but
// near-tie cases, at a few flops instead of a few hundred.
__device__ __forceinline__ double pow2_err(double x, double* lo) { const double h = x * x; *lo = fma(x, x, -h); return h; }
__device__ __forceinline__ double pow4(double x) {
  double ae; const double a = pow2_err(x, &ae);
  const double b = a * a, be = fma(a, a, -b) + 2.0 * a * ae;
  return b + be;
}
__device__ __forceinline__ double pow6(double x) {
  double ae; const double a = pow2_err(x, &ae);
  const double b = a * a, be = fma(a, a, -b) + 2.0 * a * ae;
  const double c = b * a, ce = fma(b, a, -c) + b * ae + be * a;
  return c + ce;
}
__device__ __forceinline__ double powm2(double x) {      // x^-2
  double e; const double s2 = pow2_err(x, &e);
  const double r = 1.0 / s2;
  const double res = fma(-r, s2, 1.0) - r * e;
  return fma(r, res, r);
}

// Closed-form inverse coefficients.  sqrt_g = false: as Expect_M evaluates them (:672-680);
// sqrt_g = true: as logl_W does (:312-320; g is stored as a square root and re-squared).
__device__ inline void closed_form_coef(bool sqrt_g, double b, double t, double sE, double sF, double sH,
                                 double* c1, double* c2, double* c3, double* Kc_out) {
  const double t2 = t * t, t4 = pow4(t), t6 = pow6(t), b2 = b * b;
  const double sE2 = sE * sE, sF2 = sF * sF, sH2 = sH * sH;
  double g = t2 * b2 + sH2;
  if (sqrt_g) { g = sqrt(g); g = g * g; }
  const double Kw = t2 - t4 * b2 / sF2 + t4 * b2 * g / (sF2 * (g + sF2));
  const double Kc = g - t4 * b2 / sE2 + t6 * b2 / (sE2 * (t2 + sE2));
  const double Kwc = t2 * b / (sE2 * sF2) - Kc * t2 * b / (sE2 * sF2 * (Kc + sF2)) -
                     t4 * b / (sE2 * sF2 * (t2 + sE2)) +
                     Kc * t4 * b / (sE2 * sF2 * (Kc + sF2) * (t2 + sE2));
  *c1 = Kw / (sE2 * (Kw + sE2));
  *c3 = Kc / (sF2 * (Kc + sF2));
  *c2 = Kwc;
  *Kc_out = Kc;
}

// Coefficients of the posterior means for the streaming kernel (EM_W_multi.R:691-694), component k:
//   mu_T[:,k] = aT Xw[:,k] + bT Yc[:,k],  mu_U[:,k] = aU Xw[:,k] + bU Yc[:,k]
__device__ inline void posterior_coef_k(FitState* st, int k) {
  const double sE = st->sigE, sF = st->sigF, sH = st->sigH;
  const double isE2 = powm2(sE), isF2 = powm2(sF);
  const double b = st->b[k], t = st->t[k];
  double c1, c2, c3, Kc;
  closed_form_coef(false, b, t, sE, sF, sH, &c1, &c2, &c3, &Kc);
  const double t2 = t * t, varU = t2 * (b * b) + sH * sH;
  st->aT[k] = isE2 * t2 - c1 * t2 - c2 * t2 * b;
  st->bT[k] = isF2 * (t2 * b) - c2 * t2 - c3 * b * t2;
  st->aU[k] = isE2 * (t2 * b) - c1 * t2 * b - c2 * varU;
  st->bU[k] = isF2 * varU - c2 * t2 * b - c3 * varU;
}
__device__ inline void posterior_coef(FitState* st, int R) {
  for (int k = 0; k < R; ++k) posterior_coef_k(st, k);
}

// Cyclic Jacobi eigen-decomposition of a symmetric R x R matrix (one thread).
// On exit A is (numerically) diagonal and the columns of V are the eigenvectors.
// Cyclic Jacobi eigen-decomposition of a symmetric R x R matrix by ONE thread: A -> diagonal, V = eigenvectors.
// Slow (serial rotations) and robust: the fallback of block_inv_sqrt for ill-conditioned Gram matrices.
template <int R>
__device__ void jacobi_eig(double* A, double* V) {
  for (int i = 0; i < R; ++i)
    for (int j = 0; j < R; ++j) V[i * R + j] = (i == j) ? 1.0 : 0.0;
  for (int sweep = 0; sweep < 40; ++sweep) {
    double off = 0.0, diag = 0.0;
    for (int i = 0; i < R; ++i) {
      diag += A[i * R + i] * A[i * R + i];
      for (int j = i + 1; j < R; ++j) off += A[i * R + j] * A[i * R + j];
    }
    if (off <= 1e-40 * diag || off == 0.0) break;
    for (int p = 0; p < R - 1; ++p)
      for (int q = p + 1; q < R; ++q) {
        const double apq = A[p * R + q];
        if (apq == 0.0) continue;
        const double theta = (A[q * R + q] - A[p * R + p]) / (2.0 * apq);
        const double tt = (theta >= 0.0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
        const double c = 1.0 / sqrt(tt * tt + 1.0), s = tt * c;
        for (int k = 0; k < R; ++k) {  // A <- A J
          const double akp = A[k * R + p], akq = A[k * R + q];
          A[k * R + p] = c * akp - s * akq;
          A[k * R + q] = s * akp + c * akq;
        }
        for (int k = 0; k < R; ++k) {  // A <- J' A
          const double apk = A[p * R + k], aqk = A[q * R + k];
          A[p * R + k] = c * apk - s * aqk;
          A[q * R + k] = s * apk + c * aqk;
        }
        for (int k = 0; k < R; ++k) {
          const double vkp = V[k * R + p], vkq = V[k * R + q];
          V[k * R + p] = c * vkp - s * vkq;
          V[k * R + q] = s * vkp + c * vkq;
        }
      }
  }
}

// Mat[m] = S[m]^{-1/2} for NM symmetric positive definite R x R matrices in shared memory, by the WHOLE block
// (every thread must call it; blockDim.x >= 64 NM).  Coupled Newton-Schulz iteration on Y0 = S / trace(S), Z0 = I:
//   T = (3 I - Z Y) / 2,  Y <- Y T,  Z <- T Z;   Y -> (S/c)^{1/2},  Z -> (S/c)^{-1/2}   (eigenvalues of S/c lie in (0, 1])
// One thread per matrix element, three block barriers per step; quadratic convergence once |I - Z Y| < 1, about
// log_2.25(cond) steps before that.  Replaces the one-thread Jacobi eigen-decomposition (R = 5: ~30 us of serial
// rotations with the matrices in local memory) by ~2 us; the Newton-Schulz polish that follows in the callers is
// unchanged.  The iteration is accurate to ~ cond(S) eps and stalls above cond(S) ~ 1e9, so it is trusted only when
// it converges within NS_STEPS steps (cond(S) up to ~1e3: error <= 2e-13); otherwise one thread per matrix runs the
// Jacobi eigen-decomposition (S^{-1/2} = V diag(lambda^{-1/2}) V'), which handles anything short of a singular
// Gram matrix -- and that one yields inf/NaN, which the callers' finiteness checks turn into PPLS_E_NONFINITE.
template <int R, int NM>
__device__ void block_inv_sqrt(const double* S /*[NM][R*R]*/, double* Mat /*[NM][R*R]*/, double* work /*[NM][3][R*R]*/) {
  static_assert(R * R <= 64, "one thread per element, 64 threads per matrix");
  const int m = threadIdx.x >> 6, e = threadIdx.x & 63;
  const bool act = (m < NM) && (e < R * R);
  const int i = act ? e / R : 0, j = act ? e - i * R : 0;
  double* Y = work + (size_t)(act ? m : 0) * 3 * R * R;
  double* Z = Y + R * R;
  double* T = Z + R * R;
  double c = 0.0;
  if (act) {
    const double* Sm = S + (size_t)m * R * R;
#pragma unroll
    for (int k = 0; k < R; ++k) c += Sm[k * R + k];
    Y[e] = Sm[e] / c;
    Z[e] = (i == j) ? 1.0 : 0.0;
  }
  __syncthreads();
  bool converged = false;
  constexpr int NS_STEPS = 16;
  for (int it = 0; it < NS_STEPS; ++it) {
    double err = 0.0;
    if (act) {
      double zy = 0.0;
#pragma unroll
      for (int k = 0; k < R; ++k) zy = fma(Z[i * R + k], Y[k * R + j], zy);
      const double d = (i == j) ? 1.0 : 0.0;
      err = fabs(d - zy);
      T[e] = 0.5 * (3.0 * d - zy);
    }
    const int go = __syncthreads_or(!(err < 1e-13));      // (NaN counts as not converged)
    double yn = 0.0, zn = 0.0;
    if (act) {
#pragma unroll
      for (int k = 0; k < R; ++k) { yn = fma(Y[i * R + k], T[k * R + j], yn); zn = fma(T[i * R + k], Z[k * R + j], zn); }
    }
    __syncthreads();
    if (act) { Y[e] = yn; Z[e] = zn; }
    __syncthreads();
    if (!go) { converged = true; break; }                  // the step just taken started from |I - Z Y| < 1e-13
  }
  if (converged) {
    if (act) Mat[(size_t)m * R * R + e] = Z[e] / sqrt(c);
  } else if (m < NM && e == 0) {                           // (block-uniform branch; one thread per matrix)
    double A[R * R], V[R * R];
    for (int x = 0; x < R * R; ++x) A[x] = S[(size_t)m * R * R + x];
    jacobi_eig<R>(A, V);
    for (int x = 0; x < R; ++x)
      for (int y = 0; y < R; ++y) {
        double sacc = 0.0;
        for (int k = 0; k < R; ++k) sacc += V[x * R + k] * (1.0 / sqrt(A[k * R + k])) * V[y * R + k];
        Mat[(size_t)m * R * R + x * R + y] = sacc;
      }
  }
  __syncthreads();
}

struct UpdateArgs {
  const double* momvec;   // [ld*R | Gxx | Gxy | Gyy]
  double* L;              // [ld][R] loadings, overwritten with theta_{k+1}
  double* Lscratch;       // [ld][R] work space (QR)
  FitState* st;
  double* trace;          // [max_steps]
  double* gscratch;       // grid kernel: [3 phases][row blocks][2 matrices][R*R] Gram partials
  int32_t p, q, p2, q2, ld;
  int32_t use_x;          // 1: momvec comes out of the peer-exchange inbox (x)
  XchgDev x;
};

// ---- peer exchange, receive side -----------------------------------------------------------------
__device__ inline unsigned long long ld_acquire_sys(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ inline void st_release_sys(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ inline unsigned long long global_timer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
__device__ inline unsigned long long ld_relaxed_sys(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ inline void st_relaxed_sys(unsigned long long* p, unsigned long long v) {
  asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
// ONE WARP (all 32 lanes call): wait until every rank's vector of exchange `next` has landed in the local
// inbox.  Lane r polls rank r's flag (acquire loads at system scope, all ranks in parallel), which orders the inbox
// reads of the whole block (the caller's __syncthreads follows) after the flags.
__device__ inline void xchg_wait(const XchgDev& x, unsigned long long next) {
  const int lane = threadIdx.x & 31;
  const unsigned long long* fl = x.flags[x.rank] + (next & 1ull) * XCHG_MAXW;
  const unsigned long long t0 = global_timer_ns();
  if (lane < x.world) {
    unsigned int spins = 0;                                  // acquire polls: see xchg_wait_cta
    while (ld_acquire_sys(fl + lane) < next) {
      if ((++spins & 15u) == 0u && global_timer_ns() - t0 > x.timeout_ns) { atomicExch(x.failed, 1); break; }   // (20 s by default) a peer is gone
    }
  }
  __syncwarp();
}
// The same wait on per-CTA flags (fused tail): flag (parity, source rank r, CTA `cta`) of THIS rank's block.
__device__ inline void xchg_wait_cta(const XchgDev& x, unsigned long long next, int cta) {
  const int lane = threadIdx.x & 31;
  const unsigned long long* fl = x.cflags[x.rank] + ((next & 1ull) * XCHG_MAXW) * XCHG_MAXCTA + cta;
  const unsigned long long t0 = global_timer_ns();
  if (lane < x.world) {
    // acquire polls (system scope): the inbox reads that follow -- after the caller's bar.sync -- are ordered behind the flag
    // by the acquire itself.  A relaxed poll + a trailing system-scope fence cost ~4 us more per exchange (measured on 2 GPUs).
    unsigned int spins = 0;
    while (ld_acquire_sys(fl + (size_t)lane * XCHG_MAXCTA) < next) {
      if ((++spins & 15u) == 0u && global_timer_ns() - t0 > x.timeout_ns) { atomicExch(x.failed, 1); break; }   // (20 s by default) a peer is gone
    }
  }
  __syncwarp();
}
// element e of the summed vector, ranks added in rank order (the same bits on every rank)
__device__ inline double xchg_sum(const XchgDev& x, unsigned long long next, long long e) {
  const double* in = x.inbox[x.rank] + (size_t)(next & 1ull) * x.world * x.slot + e;
  double s = __ldcg(in);
  for (int r = 1; r < x.world; ++r) s += __ldcg(in + (size_t)r * x.slot);
  return s;
}

__device__ inline bool finite_d(double x) { return isfinite(x); }

// Scalar part, executed by TWO FULL WARPS (threads 0..63 of the block; lane k < R owns component k; every sum over
// components is added by lane 0 in component order, as the sequential reference does).  Warp 1: loglik(theta_k) (logl_W +
// loglC_fast); warp 0, at the same time: the E-step moments; then warp 0: break rule and M-step scalars.  Returns 1 (on all
// 64 threads) when the run ends with this pass.
// The state is STAGED in shared memory for the whole scalar part: it is a chain of a few dozen dependent reads and writes
// of FitState fields, and each one is an L2 round trip (~0.5-1 us) when done on the global copy.  The two warps copy the
// struct in (all loads first, then the stores: one round trip), work on the copy, and copy it back before returning (nobody
// else writes the struct meanwhile: row blocks only read `done`, which the CALLER raises after this function).
// Measured at C2 (r = 3) inside the fused tail: own sums 2.9 us + staging 5.0 + arithmetic 12-13 with one warp on the global
// struct's shared copy, word-by-word staging.
struct ScalarShared {
  double c1[MAXR], c2[MAXR], c3[MAXR], tl[MAXR], lt[MAXR], lk[MAXR], g3[3 * MAXR * MAXR];
  double prev_ll;
  int done, rc;
  unsigned long long ts[4];
  FitState stc;
};

__device__ __forceinline__ void scal_bar64() { asm volatile("bar.sync 2, 64;" ::: "memory"); }

// warp 1: log-likelihood at theta_k: logl_W :312-320 + loglC_fast loglC.cpp:331-336
template <int R>
__device__ void scal_loglik(const UpdateArgs& a, ScalarShared* sh, FitState* st, const double* Gxx) {
  const int lane = threadIdx.x & 31, k = lane;
  const double* Gxy = Gxx + R * R;
  const double* Gyy = Gxy + R * R;
  const double N = st->Ntot, p = (double)a.p, q = (double)a.q;
  const double sE = st->sigE, sF = st->sigF, sH = st->sigH;
  const double sE2 = sE * sE, sF2 = sF * sF;
  const int k_pass = st->pass;
  if (k < R) {
    double c1, c2, c3, Kc;
    closed_form_coef(true, st->b[k], st->t[k], sE, sF, sH, &c1, &c2, &c3, &Kc);
    sh->lt[k] = log(sE2 + st->t[k] * st->t[k]);
    sh->lk[k] = log(sF2 + Kc);
    sh->tl[k] = -c1 * Gxx[k * R + k] - 2.0 * c2 * Gxy[k * R + k] - c3 * Gyy[k * R + k];
  }
  // (the three logarithms of the constant terms on other lanes, meanwhile)
  double lg = 0.0;
  if (lane == 8) lg = log(sE2);
  if (lane == 9) lg = log(sF2);
  if (lane == 10) lg = log(2.0 * 3.14159265358979323846);
  const double lgE = __shfl_sync(0xffffffffu, lg, 8), lgF = __shfl_sync(0xffffffffu, lg, 9), lg2pi = __shfl_sync(0xffffffffu, lg, 10);
  __syncwarp();
  if (lane == 0) {
    double sumT = 0.0, sumK = 0.0;
    double traceL = 1.0 / sE2 * st->ssqX + 1.0 / sF2 * st->ssqY;
    for (int j = 0; j < R; ++j) { sumT += sh->lt[j]; sumK += sh->lk[j]; traceL += sh->tl[j]; }
    const double Logdiag = sumT + (p - R) * lgE + sumK + (q - R) * lgF;
    const double ll = -0.5 * N * (p + q) * lg2pi - 0.5 * N * Logdiag - 0.5 * traceL;
    sh->prev_ll = st->loglik_last;                        // loglik(theta_{k-1}) = trace[k_pass - 2]: for the break rule
    st->loglik_last = ll;
    if (k_pass >= 1) a.trace[k_pass - 1] = ll;
  }
}

// warp 0: E-step second moments at theta_k: Expect_M :672-712
template <int R>
__device__ void scal_estep(const UpdateArgs& a, ScalarShared* sh, FitState* st, const double* Gxx) {
  const int lane = threadIdx.x & 31, k = lane;
  const double* Gxy = Gxx + R * R;
  const double* Gyy = Gxy + R * R;
  const double N = st->Ntot, p = (double)a.p, q = (double)a.q;
  const double sE = st->sigE, sF = st->sigF, sH = st->sigH;
  const double sE2 = sE * sE, sF2 = sF * sF, sH2 = sH * sH;
  const double isE2 = powm2(sE), isF2 = powm2(sF);
  const double sE4 = pow4(sE), sF4 = pow4(sF), sH4 = pow4(sH);
  if (k < R) {
    double c1, c2, c3, Kc;
    closed_form_coef(false, st->b[k], st->t[k], sE, sF, sH, &c1, &c2, &c3, &Kc);
    sh->c1[k] = c1; sh->c2[k] = c2; sh->c3[k] = c3;
    const double b = st->b[k], t = st->t[k], t2 = t * t, t4 = pow4(t), b2 = b * b;
    const double varU = t2 * b2 + sH2;
    const double aT = st->aT[k], bT = st->bT[k], aU = st->aU[k], bU = st->bU[k];
    const double xx = Gxx[k * R + k], xy = Gxy[k * R + k], yy = Gyy[k * R + k];
    const double mTT = aT * aT * xx + 2.0 * aT * bT * xy + bT * bT * yy;     // crossprod(mu_T)[k,k]
    const double mUU = aU * aU * xx + 2.0 * aU * bU * xy + bU * bU * yy;     // crossprod(mu_U)[k,k]
    const double mUT = aU * aT * xx + (aU * bT + bU * aT) * xy + bU * bT * yy;
    const double Ctt = t2 - isE2 * t4 - isF2 * (t4 * b2) + t4 * c1 + 2.0 * t4 * b * c2 + t4 * b2 * c3 + mTT / N;
    const double Cuu = varU - isE2 * (t4 * b2) - isF2 * (varU * varU) + t4 * b2 * c1 +
                       2.0 * t2 * b * varU * c2 + varU * varU * c3 + mUU / N;
    const double Cut = t2 * b - isE2 * (t4 * b) - isF2 * (t2 * b * varU) + t4 * b * c1 + t2 * varU * c2 +
                       t4 * b2 * c2 + t2 * b * varU * c3 + mUT / N;
    st->Ctt[k] = fabs(Ctt);     // masks of :715-716
    st->Cuu[k] = fabs(Cuu);
    st->Cut[k] = Cut;
  }
  __syncwarp();
  const double* c1v = sh->c1; const double* c2v = sh->c2; const double* c3v = sh->c3;
  // Chh (:711-712): mu_H[:,k] = hX_k Xw[:,k] + hY_k Yc[:,k]; full r x r, abs, not masked
  for (int e = lane; e < R * R; e += 32) {
    const int kk = e / R, l = e - kk * R;
    const double hXk = -sH2 * c2v[kk], hYk = isF2 * sH2 - sH2 * c3v[kk];
    const double hXl = -sH2 * c2v[l], hYl = isF2 * sH2 - sH2 * c3v[l];
    double v = (hXk * hXl * Gxx[kk * R + l] + hXk * hYl * Gxy[kk * R + l] + hYk * hXl * Gxy[l * R + kk] +
                hYk * hYl * Gyy[kk * R + l]) / N;
    if (kk == l) v += (sH2 - sH4 / sF2) + sH4 * c3v[kk];
    st->Chh[kk + l * R] = fabs(v);
  }
  __syncwarp();
  if (lane == 0) {
    // residual sums of squares, algebraically from the Gram matrices (exact for any W, C):
    //   mu_E = X - M W',  M = Xw diag(sE2 c1) + Yc diag(sE2 c2)                      (:703)
    //   ||mu_E||^2 = ||X||^2 - 2 tr(M'Xw) + tr(M'M W'W)
    double sumc1 = 0.0, sumc3 = 0.0;
    for (int j = 0; j < R; ++j) { sumc1 += c1v[j]; sumc3 += c3v[j]; }
    double ssqE = st->ssqX, ssqF = st->ssqY;
    for (int kk = 0; kk < R; ++kk) {
      const double Ak = sE2 * c1v[kk], Bk = sE2 * c2v[kk];       // on Xw, on Yc
      const double Fk = sF2 * c3v[kk], Hk = sF2 * c2v[kk];       // mu_F: on Yc, on Xw     (:708)
      ssqE -= 2.0 * (Ak * Gxx[kk * R + kk] + Bk * Gxy[kk * R + kk]);
      ssqF -= 2.0 * (Fk * Gyy[kk * R + kk] + Hk * Gxy[kk * R + kk]);
      for (int l = 0; l < R; ++l) {
        const double Al = sE2 * c1v[l], Bl = sE2 * c2v[l];
        const double Fl = sF2 * c3v[l], Hl = sF2 * c2v[l];
        const double MtM = Ak * Al * Gxx[kk * R + l] + Ak * Bl * Gxy[kk * R + l] + Bk * Al * Gxy[l * R + kk] +
                           Bk * Bl * Gyy[kk * R + l];
        const double MtMf = Fk * Fl * Gyy[kk * R + l] + Fk * Hl * Gxy[l * R + kk] + Hk * Fl * Gxy[kk * R + l] +
                            Hk * Hl * Gxx[kk * R + l];
        ssqE += MtM * st->WtW[l * R + kk];
        ssqF += MtMf * st->CtC[l * R + kk];
      }
    }
    // The reference sums squares (ssq(mu_E), :706), which cannot be negative; the expanded form above can cancel to a
    // tiny negative number when the residual is ~1e-16 of ||X||^2 (noise-free data).  Clamp instead of handing sqrt a
    // negative argument: the run then goes on, as the reference's does.
    if (ssqE < 0.0) ssqE = 0.0;
    if (ssqF < 0.0) ssqF = 0.0;
    const double Cee = (p * sE2 - p * sE2 + sE4 * sumc1 + ssqE / N) / p;     // :706
    const double Cff = (q * sF2 - q * sF2 + sF4 * sumc3 + ssqF / N) / q;     // :709
    st->Cee = Cee; st->Cff = Cff;
  }
  __syncwarp();
}

// warp 0, after both: break rule (:792) / EMsteps exhausted, then the M-step scalars (:734-738) -> theta_{k+1}
template <int R>
__device__ int scal_finish(ScalarShared* sh, FitState* st) {
  const int lane = threadIdx.x & 31, k = lane;
  const int k_pass = st->pass;
  if (lane == 0) {
    bool done = (k_pass >= st->max_steps);
    if (k_pass >= 2 && (st->loglik_last - sh->prev_ll) < st->atol) done = true;    // (the values of trace[k-1], trace[k-2])
    if (!finite_d(st->loglik_last)) { st->status = PPLS_E_NONFINITE; done = true; }
    if (done) st->steps_done = k_pass;                   // the CALLER raises st->done (see the grid kernel)
    sh->done = done ? 1 : 0;
  }
  __syncwarp();
  if (sh->done) return 1;
  bool ok = true;
  if (k < R) {
    const double bn = st->Cut[k] * (1.0 / st->Ctt[k]);       // Cut %*% solve(Ctt) * I
    const double tn = sqrt(st->Ctt[k]);
    st->b[k] = bn; st->t[k] = tn;
    ok = finite_d(bn) && finite_d(tn);
  }
  if (lane == 0) {
    double trChh = 0.0;
    for (int j = 0; j < R; ++j) trChh += st->Chh[j + j * R];
    st->sigE = sqrt(st->Cee); st->sigF = sqrt(st->Cff); st->sigH = sqrt(trChh / R);
    ok = ok && finite_d(st->sigE) && finite_d(st->sigF) && finite_d(st->sigH);
  }
  __syncwarp();
  if (k < R) {
    posterior_coef_k(st, k);
    ok = ok && finite_d(st->aT[k]) && finite_d(st->bT[k]) && finite_d(st->aU[k]) && finite_d(st->bU[k]);
  }
  ok = __all_sync(0xffffffffu, ok);
  if (lane == 0) {
    st->pass = k_pass + 1;
    if (!ok) { st->status = PPLS_E_NONFINITE; st->steps_done = k_pass; }
  }
  return ok ? 0 : 1;
}

// Threads 0..63 of the block call this (both warps whole).
template <int R>
__device__ int update_scalars(const UpdateArgs& a, ScalarShared* sh) {
  static_assert(sizeof(FitState) % 8 == 0, "FitState is copied as 8-byte words");
  const int tid = threadIdx.x, wrp = tid >> 5;
  constexpr int NW = (int)(sizeof(FitState) / 8), NB = (NW + 63) / 64;
  const unsigned long long* src = reinterpret_cast<const unsigned long long*>(a.st);
  unsigned long long* cpy = reinterpret_cast<unsigned long long*>(&sh->stc);
  if (tid == 0) sh->ts[0] = global_timer_ns();
  {
    unsigned long long v[NB];
#pragma unroll
    for (int j = 0; j < NB; ++j) { const int e = tid + 64 * j; v[j] = (e < NW) ? src[e] : 0ull; }
    const double gv = (tid < 3 * R * R) ? a.momvec[(size_t)a.ld * R + tid] : 0.0;      // Gxx | Gxy | Gyy (3 R^2 <= 192: see below)
#pragma unroll
    for (int j = 0; j < NB; ++j) { const int e = tid + 64 * j; if (e < NW) cpy[e] = v[j]; }
    if (tid < 3 * R * R) sh->g3[tid] = gv;
    for (int e = tid + 64; e < 3 * R * R; e += 64) sh->g3[e] = a.momvec[(size_t)a.ld * R + e];
  }
  scal_bar64();
  if (tid == 0) sh->ts[1] = global_timer_ns();
  FitState* st = &sh->stc;
  if (wrp == 1) scal_loglik<R>(a, sh, st, sh->g3);
  else scal_estep<R>(a, sh, st, sh->g3);
  scal_bar64();
  if (wrp == 0) { const int rc = scal_finish<R>(sh, st); if (tid == 0) sh->rc = rc; }
  scal_bar64();
  if (tid == 0) sh->ts[2] = global_timer_ns();
  unsigned long long* dst = reinterpret_cast<unsigned long long*>(a.st);
  for (int e = tid; e < NW; e += 64) dst[e] = cpy[e];
  const int rc = sh->rc;
  scal_bar64();
  return rc;
}

}  // namespace ppls
