// K2 "em_update": the tiny replicated update between two data passes, FP64, sm_100a.
// Compiled with -fmad=false: the closed-form coefficients below cancel catastrophically
// for small noise variances and must be evaluated exactly as the reference writes them.
//
// From the reduced moment vector of pass k (parameters theta_k) it computes
//   loglik(theta_k)            logl_W, Package/PPLS/R/EM_W_multi.R:297-323 (coefficients :312-320)
//                              + loglC_fast, Package/PPLS/src/loglC.cpp:331-336
//   E-step second moments      Expect_M fast branch, EM_W_multi.R:672-680, :696-712, masks :715-716
//   the break rule             PPLS_simult, :792 (signed increment, i > 1)
//   M-step -> theta_{k+1}      Maximiz_M, :729-742; orth() = OmicsPLS::orth (polar factor /
//                              Householder QR, historical copy Package/functions.R:252-260)
//   coefficients of mu_T, mu_U for the next pass (:691-694 with the diagonal algebra collapsed)
#include "common.cuh"
#include <math.h>
#include <string.h>
#include <cooperative_groups.h>
namespace cg = cooperative_groups;

namespace ppls {

constexpr int UPD_THREADS = 512;
constexpr int UPD_WARPS = UPD_THREADS / 32;

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
__device__ void closed_form_coef(bool sqrt_g, double b, double t, double sE, double sF, double sH,
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
__device__ void posterior_coef_k(FitState* st, int k) {
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
__device__ void posterior_coef(FitState* st, int R) {
  for (int k = 0; k < R; ++k) posterior_coef_k(st, k);
}

// ---- deterministic block reductions (fixed thread partition, fixed tree) -------------------
// Sums nv (<= 36) per-thread values over the CTA; result in out[0..nv) (shared), all threads sync.
__device__ void block_reduce_vec(double* vals, int nv, double* out, double* scratch /*[UPD_WARPS*36]*/) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int v = 0; v < nv; ++v) {
    double x = vals[v];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
    if (lane == 0) scratch[warp * 36 + v] = x;
  }
  __syncthreads();
  if ((int)threadIdx.x < nv) {
    double s = 0.0;
    for (int w = 0; w < UPD_WARPS; ++w) s += scratch[w * 36 + threadIdx.x];
    out[threadIdx.x] = s;
  }
  __syncthreads();
}

// Gram matrix A'A of A [n][R] row-major -> S (shared, R*R, symmetric, both triangles).
template <int R>
__device__ void block_gram(const double* A, int n, double* S, double* scratch) {
  constexpr int NV = R * (R + 1) / 2;
  double v[NV];
#pragma unroll
  for (int e = 0; e < NV; ++e) v[e] = 0.0;
  for (int j = threadIdx.x; j < n; j += UPD_THREADS) {
    double row[R];
#pragma unroll
    for (int k = 0; k < R; ++k) row[k] = A[(size_t)j * R + k];
    int e = 0;
#pragma unroll
    for (int k = 0; k < R; ++k)
#pragma unroll
      for (int l = k; l < R; ++l) v[e++] += row[k] * row[l];
  }
  __shared__ double packed[36];
  block_reduce_vec(v, NV, packed, scratch);
  if (threadIdx.x == 0) {
    int e = 0;
    for (int k = 0; k < R; ++k)
      for (int l = k; l < R; ++l) { S[k * R + l] = packed[e]; S[l * R + k] = packed[e]; ++e; }
  }
  __syncthreads();
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

// out[j][:] = A[j][:] * Mat (R x R, shared) for all rows j.
template <int R>
__device__ void rows_times(const double* A, int n, const double* Mat, double* out) {
  for (int j = threadIdx.x; j < n; j += UPD_THREADS) {
    double row[R], res[R];
#pragma unroll
    for (int k = 0; k < R; ++k) row[k] = A[(size_t)j * R + k];
#pragma unroll
    for (int l = 0; l < R; ++l) {
      double s = 0.0;
#pragma unroll
      for (int k = 0; k < R; ++k) s += row[k] * Mat[k * R + l];
      res[l] = s;
    }
#pragma unroll
    for (int l = 0; l < R; ++l) out[(size_t)j * R + l] = res[l];
  }
  __syncthreads();
}

// orth(G, "SVD"): the polar factor U V' = G (G'G)^{-1/2}, then one Newton-Schulz step
// W <- W (3I - W'W)/2 that restores orthonormality to rounding regardless of cond(G)^2 lost
// in the Gram matrix.  G, out: [n][R] row-major (out may alias neither G nor itself).
template <int R>
__device__ void orth_polar(const double* G, int n, double* out, double* S, double* Mat,
                           double* scratch) {
  block_gram<R>(G, n, S, scratch);
  __syncthreads();
  {
    __shared__ double nswork[3 * R * R];
    block_inv_sqrt<R, 1>(S, Mat, nswork);
  }
  rows_times<R>(G, n, Mat, out);
  block_gram<R>(out, n, S, scratch);
  if (threadIdx.x == 0)
    for (int i = 0; i < R; ++i)
      for (int j = 0; j < R; ++j) Mat[i * R + j] = (i == j ? 1.5 : 0.0) - 0.5 * S[i * R + j];
  __syncthreads();
  rows_times<R>(out, n, Mat, out);
}

// orth(G, "QR"): qr.Q(qr(G)) with the Householder sign convention R_ll = -sign(x_ll)||x_l||
// (LINPACK dqrdc2 and LAPACK dgeqr2 agree on it), times sign(e[,1]'G[,1]) = sign(R_11).
// A: working copy of G [n][R] (destroyed), Q: output [n][R].
template <int R>
__device__ void orth_qr(double* A, int n, double* Q, double* scratch) {
  __shared__ double tau[R], beta0, red[36];
  for (int l = 0; l < R; ++l) {
    double v[1 + R];
    for (int e = 0; e <= R; ++e) v[e] = 0.0;
    for (int j = l + 1 + threadIdx.x; j < n; j += UPD_THREADS) v[0] += A[(size_t)j * R + l] * A[(size_t)j * R + l];
    block_reduce_vec(v, 1, red, scratch);
    const double alpha = A[(size_t)l * R + l], xnorm2 = red[0];
    __syncthreads();
    double beta, tl, scale;
    if (xnorm2 == 0.0) { beta = alpha; tl = 0.0; scale = 0.0; }
    else {
      beta = -copysign(sqrt(alpha * alpha + xnorm2), alpha);
      tl = (beta - alpha) / beta;
      scale = 1.0 / (alpha - beta);
    }
    for (int j = l + 1 + threadIdx.x; j < n; j += UPD_THREADS) A[(size_t)j * R + l] *= scale;
    __syncthreads();
    if (threadIdx.x == 0) { tau[l] = tl; if (l == 0) beta0 = beta; A[(size_t)l * R + l] = beta; }
    // apply H_l to the trailing columns
    for (int e = 0; e <= R; ++e) v[e] = 0.0;
    for (int j = l + threadIdx.x; j < n; j += UPD_THREADS) {
      const double vj = (j == l) ? 1.0 : A[(size_t)j * R + l];
      for (int c = l + 1; c < R; ++c) v[c] += vj * A[(size_t)j * R + c];
    }
    block_reduce_vec(v, R, red, scratch);
    for (int j = l + threadIdx.x; j < n; j += UPD_THREADS) {
      const double vj = (j == l) ? 1.0 : A[(size_t)j * R + l];
      for (int c = l + 1; c < R; ++c) A[(size_t)j * R + c] -= tl * vj * red[c];
    }
    __syncthreads();
  }
  // Q = H_0 ... H_{R-1} [I; 0]
  for (int j = threadIdx.x; j < n; j += UPD_THREADS)
    for (int c = 0; c < R; ++c) Q[(size_t)j * R + c] = (j == c) ? 1.0 : 0.0;
  __syncthreads();
  for (int l = R - 1; l >= 0; --l) {
    double v[1 + R];
    for (int e = 0; e <= R; ++e) v[e] = 0.0;
    for (int j = l + threadIdx.x; j < n; j += UPD_THREADS) {
      const double vj = (j == l) ? 1.0 : A[(size_t)j * R + l];
      for (int c = l; c < R; ++c) v[c] += vj * Q[(size_t)j * R + c];
    }
    block_reduce_vec(v, R, red, scratch);
    const double tl = tau[l];
    for (int j = l + threadIdx.x; j < n; j += UPD_THREADS) {
      const double vj = (j == l) ? 1.0 : A[(size_t)j * R + l];
      for (int c = l; c < R; ++c) Q[(size_t)j * R + c] -= tl * vj * red[c];
    }
    __syncthreads();
  }
  const double sgn = (beta0 > 0.0) ? 1.0 : (beta0 < 0.0 ? -1.0 : 0.0);
  for (int j = threadIdx.x; j < n; j += UPD_THREADS)
    for (int c = 0; c < R; ++c) Q[(size_t)j * R + c] *= sgn;
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
// inbox.  Lane r polls rank r's flag (relaxed loads, all ranks in parallel); one system-scope fence then
// orders the inbox reads of the whole block (the caller's __syncthreads follows) after the flags.
__device__ inline void xchg_wait(const XchgDev& x, unsigned long long next) {
  const int lane = threadIdx.x & 31;
  const unsigned long long* fl = x.flags[x.rank] + (next & 1ull) * XCHG_MAXW;
  const unsigned long long t0 = global_timer_ns();
  if (lane < x.world) {
    while (ld_relaxed_sys(fl + lane) < next) {
      if (global_timer_ns() - t0 > 20000000000ull) { atomicExch(x.failed, 1); break; }   // 20 s: a peer is gone
    }
  }
  __syncwarp();
  __threadfence_system();
}
// element e of the summed vector, ranks added in rank order (the same bits on every rank)
__device__ inline double xchg_sum(const XchgDev& x, unsigned long long next, long long e) {
  const double* in = x.inbox[x.rank] + (size_t)(next & 1ull) * x.world * x.slot + e;
  double s = __ldcg(in);
  for (int r = 1; r < x.world; ++r) s += __ldcg(in + (size_t)r * x.slot);
  return s;
}

__device__ inline bool finite_d(double x) { return isfinite(x); }

// Scalar part, executed by ONE FULL WARP (lane k < R owns component k; every sum over components is
// added by lane 0 in component order, as the sequential reference does): loglik(theta_k), E-step
// moments, break rule, M-step scalars.  Returns 1 (on all lanes) when the run ends with this pass.
struct ScalarShared { double c1[MAXR], c2[MAXR], c3[MAXR], tl[MAXR], lt[MAXR], lk[MAXR]; int done; };

template <int R>
__device__ int update_scalars(const UpdateArgs& a, ScalarShared* sh) {
  FitState* st = a.st;
  const int lane = threadIdx.x & 31;
  const double* Gxx = a.momvec + (size_t)a.ld * R;
  const double* Gxy = Gxx + R * R;
  const double* Gyy = Gxy + R * R;
  const double N = st->Ntot, p = (double)a.p, q = (double)a.q;
  const double sE = st->sigE, sF = st->sigF, sH = st->sigH;
  const double sE2 = sE * sE, sF2 = sF * sF, sH2 = sH * sH;
  const int k_pass = st->pass;
  const int k = lane;

  // ---- log-likelihood at theta_k: logl_W :312-320 + loglC_fast loglC.cpp:331-336 ----------
  if (k < R) {
    double c1, c2, c3, Kc;
    closed_form_coef(true, st->b[k], st->t[k], sE, sF, sH, &c1, &c2, &c3, &Kc);
    sh->lt[k] = log(sE2 + st->t[k] * st->t[k]);
    sh->lk[k] = log(sF2 + Kc);
    sh->tl[k] = -c1 * Gxx[k * R + k] - 2.0 * c2 * Gxy[k * R + k] - c3 * Gyy[k * R + k];
  }
  __syncwarp();
  if (lane == 0) {
    double sumT = 0.0, sumK = 0.0;
    double traceL = 1.0 / sE2 * st->ssqX + 1.0 / sF2 * st->ssqY;
    for (int j = 0; j < R; ++j) { sumT += sh->lt[j]; sumK += sh->lk[j]; traceL += sh->tl[j]; }
    const double Logdiag = sumT + (p - R) * log(sE2) + sumK + (q - R) * log(sF2);
    const double ll = -0.5 * N * (p + q) * log(2.0 * 3.14159265358979323846) - 0.5 * N * Logdiag - 0.5 * traceL;
    st->loglik_last = ll;
    if (k_pass >= 1) a.trace[k_pass - 1] = ll;
  }

  // ---- E-step second moments at theta_k: Expect_M :672-712 ----------------------------------
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
    const double Cee = (p * sE2 - p * sE2 + sE4 * sumc1 + ssqE / N) / p;     // :706
    const double Cff = (q * sF2 - q * sF2 + sF4 * sumc3 + ssqF / N) / q;     // :709
    st->Cee = Cee; st->Cff = Cff;
    // ---- break rule (:792) / EMsteps exhausted -------------------------------------------------
    bool done = (k_pass >= st->max_steps);
    if (k_pass >= 2 && (a.trace[k_pass - 1] - a.trace[k_pass - 2]) < st->atol) done = true;
    if (!finite_d(st->loglik_last)) { st->status = PPLS_E_NONFINITE; done = true; }
    if (done) st->steps_done = k_pass;                   // the CALLER raises st->done (see the grid kernel)
    sh->done = done ? 1 : 0;
  }
  __syncwarp();
  if (sh->done) return 1;

  // ---- M-step scalars (:734-738) -> theta_{k+1} -----------------------------------------------
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

template <int R>
__device__ void update_body(const UpdateArgs& a) {
  __shared__ double scratch[UPD_WARPS * 36], S[R * R], Mat[R * R];
  __shared__ int s_done;
  __shared__ ScalarShared ssh;
  FitState* st = a.st;
  if (st->done) return;
  if (threadIdx.x < 32) { const int d = update_scalars<R>(a, &ssh); if (threadIdx.x == 0) { s_done = d; if (d) raise_done(st); } }
  __syncthreads();
  if (s_done) return;
  // ---- loadings: W = orth(t(X) mu_T), C = orth(t(Y) mu_U)  (:732-733) ----------------------------
  const double* G1 = a.momvec;
  const double* G2 = a.momvec + (size_t)a.p2 * R;
  double* Wn = a.L;
  double* Cn = a.L + (size_t)a.p2 * R;
  if (st->orth_type == PPLS_ORTH_QR) {
    for (size_t e = threadIdx.x; e < (size_t)a.ld * R; e += UPD_THREADS) a.Lscratch[e] = a.momvec[e];
    __syncthreads();
    orth_qr<R>(a.Lscratch, a.p2, Wn, scratch);
    orth_qr<R>(a.Lscratch + (size_t)a.p2 * R, a.q2, Cn, scratch);
  } else {
    orth_polar<R>(G1, a.p2, Wn, S, Mat, scratch);
    orth_polar<R>(G2, a.q2, Cn, S, Mat, scratch);
  }
  block_gram<R>(Wn, a.p2, S, scratch);
  if (threadIdx.x < R * R) st->WtW[threadIdx.x] = S[threadIdx.x];
  __syncthreads();
  block_gram<R>(Cn, a.q2, S, scratch);
  if (threadIdx.x < R * R) st->CtC[threadIdx.x] = S[threadIdx.x];
  if (threadIdx.x == 0) {
    bool ok = true;
    for (int e = 0; e < R * R; ++e) ok = ok && finite_d(st->WtW[e]) && finite_d(S[e]);
    if (!ok) { st->status = PPLS_E_NONFINITE; raise_done(st); st->steps_done = st->pass - 1; }
  }
}

template <int R>
__global__ void __launch_bounds__(UPD_THREADS, 1) em_update_kernel(const UpdateArgs a) {
  if (a.use_x) {
    if (a.st->done) return;
    __shared__ int bad;
    const unsigned long long next = *a.x.epoch + 1;
    if (threadIdx.x < 32) { xchg_wait(a.x, next); if (threadIdx.x == 0) bad = *a.x.failed; }
    __syncthreads();
    if (bad) { if (threadIdx.x == 0) { a.st->status = PPLS_E_NCCL; raise_done(a.st); } return; }
    double* mv = const_cast<double*>(a.momvec);
    const long long len = (long long)a.ld * R + 3 * R * R;
    for (long long e = threadIdx.x; e < len; e += blockDim.x) mv[e] = xchg_sum(a.x, next, e);
    __syncthreads();
    if (threadIdx.x == 0) *a.x.epoch = next;
  }
  update_body<R>(a);
}

// ---- batched multi-start: one block per start ------------------------------------------------------
// The streaming kernel runs `g` starts at once as g*R "components" (the E-step is diagonal in the
// components, so a pack of starts is just a wider W); this kernel cuts every start's slice out of the
// pack's moment vector, runs the ordinary update on it and writes the new loadings / coefficients back
// into the pack.  prepare = 1: start state (coefficients, W'W, C'C) instead of an update.
struct BatchArgs {
  const double* momvec_pack;   // [groups][ld*RP + 3*RP*RP]
  double* L_pack;              // [groups][ld][RP]
  FitState* pack_st;           // [groups]  what the streaming kernel reads (coefficients, done)
  FitState* st;                // [S]
  double* momvec_s;            // [S][ld*R + 3*R*R]
  double* L_s;                 // [S][ld][R]
  double* Lscratch_s;          // [S][ld][R]
  double* trace;               // [S][tracelen]
  int32_t RP, g, S, tracelen, prepare;
  int32_t p, q, p2, q2, ld;
  double* coef;                // GEMM path (em_batch.cu): [groups][4][RP] aT bT aU bU instead of pack_st's arrays
};

template <int R>
__device__ void prepare_body(const UpdateArgs& a);

template <int R>
__global__ void __launch_bounds__(UPD_THREADS, 1) em_update_batch_kernel(const BatchArgs b) {
  const int sidx = blockIdx.x, group = sidx / b.g, k0 = (sidx - group * b.g) * R, RP = b.RP;
  const size_t veclen = (size_t)b.ld * R + 3 * R * R, veclenP = (size_t)b.ld * RP + 3 * RP * RP;
  UpdateArgs a;
  a.momvec = b.momvec_s + sidx * veclen; a.L = b.L_s + (size_t)sidx * b.ld * R;
  a.Lscratch = b.Lscratch_s + (size_t)sidx * b.ld * R; a.st = b.st + sidx; a.trace = b.trace + (size_t)sidx * b.tracelen;
  a.gscratch = nullptr; a.p = b.p; a.q = b.q; a.p2 = b.p2; a.q2 = b.q2; a.ld = b.ld; a.use_x = 0;
  FitState* st = a.st;
  FitState* ps = b.pack_st + group;
  double* Lp = b.L_pack + (size_t)group * b.ld * RP;
  if (b.prepare) {
    prepare_body<R>(a);
  } else {
    if (st->done) return;
    const double* mp = b.momvec_pack + group * veclenP;
    double* ms = b.momvec_s + sidx * veclen;
    for (int e = threadIdx.x; e < b.ld * R; e += UPD_THREADS) { const int j = e / R, k = e - j * R; ms[e] = mp[(size_t)j * RP + k0 + k]; }
    for (int e = threadIdx.x; e < 3 * R * R; e += UPD_THREADS) {
      const int which = e / (R * R), kl = e - which * R * R, k = kl / R, l = kl - k * R;
      ms[(size_t)b.ld * R + e] = mp[(size_t)b.ld * RP + (size_t)which * RP * RP + (size_t)(k0 + k) * RP + k0 + l];
    }
    __syncthreads();
    update_body<R>(a);
  }
  __syncthreads();
  for (int e = threadIdx.x; e < b.ld * R; e += UPD_THREADS) { const int j = e / R, k = e - j * R; Lp[(size_t)j * RP + k0 + k] = a.L[e]; }
  if (threadIdx.x < R) {
    const int k = threadIdx.x;
    if (b.coef) {
      double* cf = b.coef + (size_t)group * 4 * RP;
      cf[k0 + k] = st->aT[k]; cf[RP + k0 + k] = st->bT[k]; cf[2 * RP + k0 + k] = st->aU[k]; cf[3 * RP + k0 + k] = st->bU[k];
    } else {
      ps->aT[k0 + k] = st->aT[k]; ps->bT[k0 + k] = st->bT[k]; ps->aU[k0 + k] = st->aU[k]; ps->bU[k0 + k] = st->bU[k];
    }
  }
}

// a pack is done when every member is
__global__ void batch_done_kernel(FitState* pack_st, const FitState* st, int g, int S) {
  const int group = blockIdx.x * blockDim.x + threadIdx.x;
  if (group * g >= S) return;
  int all = 1;
  for (int m = group * g; m < (group + 1) * g && m < S; ++m) all = all && st[m].done;
  pack_st[group].done = all;
}


// ---- grid version of K2 (type = "SVD"): the O((p+q) r^2) work spread over row blocks ---------------
// Block 0 does the scalar part; blocks 1..nb own rows j = (b-1)*GT + tid (+ nb*GT ...) of [G1; G2].
// Three grid-wide syncs separate: Gram of G | G (G'G)^{-1/2} and its Gram | Newton-Schulz polish and
// the Gram of the new loadings.  Partial Grams are added in block order: run-to-run deterministic.
constexpr int GT = 256, GW = GT / 32;

template <int R>
__device__ void grid_partial_gram(const double* A, int row_lo, int row_hi, int rb, int nb, double* dst /*[R*R]*/,
                                  double* scratch /*[GW*36]*/) {
  constexpr int NV = R * (R + 1) / 2;
  double v[NV];
#pragma unroll
  for (int e = 0; e < NV; ++e) v[e] = 0.0;
  // same thread <-> row mapping as the loops that WRITE the rows in the same phase (no grid sync between)
  for (int j = rb * GT + threadIdx.x; j < row_hi; j += nb * GT) {
    if (j < row_lo) continue;
    double row[R];
#pragma unroll
    for (int k = 0; k < R; ++k) row[k] = A[(size_t)j * R + k];
    int e = 0;
#pragma unroll
    for (int k = 0; k < R; ++k)
#pragma unroll
      for (int l = k; l < R; ++l) v[e++] += row[k] * row[l];
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int e = 0; e < NV; ++e) {
    double x = v[e];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
    if (lane == 0) scratch[warp * 36 + e] = x;
  }
  __syncthreads();
  if ((int)threadIdx.x < NV) {
    double sacc = 0.0;
    for (int w = 0; w < GW; ++w) sacc += scratch[w * 36 + threadIdx.x];
    dst[threadIdx.x] = sacc;              // packed upper triangle
  }
  __syncthreads();
}

// S (full symmetric R x R, shared) = sum over row blocks of the packed partials
template <int R>
__device__ void grid_sum_gram(const double* parts /*[nb][2][R*R]*/, int nb, int which, double* S) {
  constexpr int NV = R * (R + 1) / 2;
  if ((int)threadIdx.x < NV) {
    double sacc = 0.0;
    for (int b = 0; b < nb; ++b) sacc += parts[((size_t)b * 2 + which) * R * R + threadIdx.x];
    int e = threadIdx.x, k = 0;
    while (e >= R - k) { e -= R - k; ++k; }
    const int l = k + e;
    S[k * R + l] = sacc; S[l * R + k] = sacc;
  }
}

template <int R>
__global__ void __launch_bounds__(GT) em_update_grid_kernel(const UpdateArgs a) {
  cg::grid_group grid = cg::this_grid();
  __shared__ double scratch[GW * 36], S[2][R * R], Mat[2][R * R];
  FitState* st = a.st;
  if (st->done) return;                                   // same on every block: no sync is skipped
  const int nb = gridDim.x - 1, rb = (int)blockIdx.x - 1;
  double* ph0 = a.gscratch;
  double* ph1 = ph0 + (size_t)nb * 2 * R * R;
  const double* G = a.momvec;
  // ---- phase 0 (multi-GPU peer exchange): wait for the world's vectors, sum them in rank order ------
  if (a.use_x) {
    const unsigned long long next = *a.x.epoch + 1;       // written below, after every block has read it
    if (threadIdx.x < 32) xchg_wait(a.x, next);
    __syncthreads();
    double* mv = const_cast<double*>(a.momvec);
    const long long len = (long long)a.ld * R + 3 * R * R;
    for (long long e = (long long)blockIdx.x * GT + threadIdx.x; e < len; e += (long long)gridDim.x * GT)
      mv[e] = xchg_sum(a.x, next, e);
    grid.sync();
    if (*a.x.failed) {                                    // same on every block (raised before the sync)
      if (blockIdx.x == 0 && threadIdx.x == 0) { st->status = PPLS_E_NCCL; raise_done(st); }
      return;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) *a.x.epoch = next;
  }
  // ---- phase 1 + 2.  Block 0: the scalar part (log-likelihood, E-step moments, break rule, M-step scalars), which
  // needs only the Gram tail.  Row blocks, meanwhile: Gram of G1, G2 | barrier among the ROW blocks only (a counter;
  // all blocks are co-resident under the cooperative launch) | T = G (G'G)^{-1/2} into Lscratch and its Gram.
  // Whether the run ends with this pass is known only after block 0 is through, so theta_k's loadings in L stay
  // untouched until the grid-wide sync.  st->done is read by every block at entry, so it may only be raised after
  // that sync; until then the decision travels in a scratch flag.
  int* done_now = reinterpret_cast<int*>(a.gscratch + (size_t)3 * nb * 2 * R * R);
  unsigned int* rowbar = reinterpret_cast<unsigned int*>(a.gscratch + (size_t)3 * nb * 2 * R * R + 2);
  double* T = a.Lscratch;
  if (rb < 0) {
    __shared__ ScalarShared ssh;
    if (threadIdx.x < 32) { const int d = update_scalars<R>(a, &ssh); if (threadIdx.x == 0) *done_now = d; }
  } else {
    grid_partial_gram<R>(G, 0, a.p2, rb, nb, ph0 + ((size_t)rb * 2 + 0) * R * R, scratch);
    grid_partial_gram<R>(G, a.p2, a.ld, rb, nb, ph0 + ((size_t)rb * 2 + 1) * R * R, scratch);
    __syncthreads();
    if (threadIdx.x == 0) {
      __threadfence();
      atomicAdd(rowbar, 1u);
      while (*reinterpret_cast<volatile unsigned int*>(rowbar) < (unsigned int)nb) { }
      __threadfence();
    }
    __syncthreads();
    grid_sum_gram<R>(ph0, nb, 0, S[0]);
    grid_sum_gram<R>(ph0, nb, 1, S[1]);
    __syncthreads();
    {
      __shared__ double nswork[2 * 3 * R * R];
      block_inv_sqrt<R, 2>(&S[0][0], &Mat[0][0], nswork);     // (G1'G1)^{-1/2} and (G2'G2)^{-1/2}, every row block its own copy
    }
    for (int j = rb * GT + threadIdx.x; j < a.ld; j += nb * GT) {
      const double* M = Mat[j >= a.p2 ? 1 : 0];
      double row[R];
#pragma unroll
      for (int k = 0; k < R; ++k) row[k] = G[(size_t)j * R + k];
#pragma unroll
      for (int l = 0; l < R; ++l) {
        double sacc = 0.0;
#pragma unroll
        for (int k = 0; k < R; ++k) sacc += row[k] * M[k * R + l];
        T[(size_t)j * R + l] = sacc;
      }
    }
    __syncthreads();
    grid_partial_gram<R>(T, 0, a.p2, rb, nb, ph1 + ((size_t)rb * 2 + 0) * R * R, scratch);
    grid_partial_gram<R>(T, a.p2, a.ld, rb, nb, ph1 + ((size_t)rb * 2 + 1) * R * R, scratch);
  }
  grid.sync();
  if (rb < 0 && threadIdx.x == 0) *rowbar = 0u;             // every row block is past the barrier: arm it for the next launch
  if (*done_now) {                                        // the run ended with this pass: keep theta_k
    if (rb < 0 && threadIdx.x == 0) raise_done(st);
    return;
  }
  // ---- phase 3: one Newton-Schulz step L <- L M, M = (3I - S)/2 with S = L'L --------------------------
  // The Gram of the result is M S M exactly (L is well conditioned here, so S carries full accuracy):
  // block 0 forms W'W, C'C from it while the row blocks apply the step -- no further pass, no further sync.
  grid_sum_gram<R>(ph1, nb, 0, S[0]);
  grid_sum_gram<R>(ph1, nb, 1, S[1]);
  __syncthreads();
  if (threadIdx.x < 2 * R * R) {
    const int w = threadIdx.x / (R * R), e = threadIdx.x - w * R * R;
    Mat[w][e] = ((e / R) == (e % R) ? 1.5 : 0.0) - 0.5 * S[w][e];
  }
  __syncthreads();
  if (rb >= 0) {
    for (int j = rb * GT + threadIdx.x; j < a.ld; j += nb * GT) {
      const double* M = Mat[j >= a.p2 ? 1 : 0];
      double row[R], res[R];
#pragma unroll
      for (int k = 0; k < R; ++k) row[k] = T[(size_t)j * R + k];
#pragma unroll
      for (int l = 0; l < R; ++l) {
        double sacc = 0.0;
#pragma unroll
        for (int k = 0; k < R; ++k) sacc += row[k] * M[k * R + l];
        res[l] = sacc;
      }
#pragma unroll
      for (int l = 0; l < R; ++l) a.L[(size_t)j * R + l] = res[l];
    }
  } else {
    __shared__ double SM[2][R * R];
    if (threadIdx.x < 2 * R * R) {                          // S M
      const int w = threadIdx.x / (R * R), e = threadIdx.x - w * R * R, i = e / R, j = e % R;
      double sacc = 0.0;
      for (int k = 0; k < R; ++k) sacc += S[w][i * R + k] * Mat[w][k * R + j];
      SM[w][e] = sacc;
    }
    __syncthreads();
    if (threadIdx.x < 2 * R * R) {                          // M (S M); M is symmetric
      const int w = threadIdx.x / (R * R), e = threadIdx.x - w * R * R, i = e / R, j = e % R;
      double sacc = 0.0;
      for (int k = 0; k < R; ++k) sacc += Mat[w][i * R + k] * SM[w][k * R + j];
      (w ? st->CtC : st->WtW)[e] = sacc;
      S[w][e] = sacc;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      bool ok = true;
      for (int e = 0; e < R * R; ++e) ok = ok && finite_d(S[0][e]) && finite_d(S[1][e]);
      if (!ok) { st->status = PPLS_E_NONFINITE; raise_done(st); st->steps_done = st->pass - 1; }
    }
  }
}

// orth(G, type) alone (OmicsPLS::orth as called at EM_W_multi.R:732-733 and by variances.PPLS_simult, :831-832):
// out = polar factor of G [n][R] (SVD) or qr.Q(qr(G)) with the sign of column 1 (QR; `work` = scratch copy of G).
template <int R>
__global__ void __launch_bounds__(UPD_THREADS, 1) orth_only_kernel(const double* G, int n, double* out, int orth_type, double* work) {
  __shared__ double scratch[UPD_WARPS * 36], S[R * R], Mat[R * R];
  if (orth_type == PPLS_ORTH_QR) {
    for (size_t e = threadIdx.x; e < (size_t)n * R; e += UPD_THREADS) work[e] = G[e];
    __syncthreads();
    orth_qr<R>(work, n, out, scratch);
  } else {
    orth_polar<R>(G, n, out, S, Mat, scratch);
  }
}
cudaError_t launch_orth(int R, const double* G, int n, double* out, int orth_type, double* work, cudaStream_t stream) {
  if (orth_type == PPLS_ORTH_QR && !work) return cudaErrorInvalidValue;
  switch (R) {
    case 1: orth_only_kernel<1><<<1, UPD_THREADS, 0, stream>>>(G, n, out, orth_type, work); break;
    case 2: orth_only_kernel<2><<<1, UPD_THREADS, 0, stream>>>(G, n, out, orth_type, work); break;
    case 3: orth_only_kernel<3><<<1, UPD_THREADS, 0, stream>>>(G, n, out, orth_type, work); break;
    case 4: orth_only_kernel<4><<<1, UPD_THREADS, 0, stream>>>(G, n, out, orth_type, work); break;
    case 5: orth_only_kernel<5><<<1, UPD_THREADS, 0, stream>>>(G, n, out, orth_type, work); break;
    case 6: orth_only_kernel<6><<<1, UPD_THREADS, 0, stream>>>(G, n, out, orth_type, work); break;
    case 7: orth_only_kernel<7><<<1, UPD_THREADS, 0, stream>>>(G, n, out, orth_type, work); break;
    case 8: orth_only_kernel<8><<<1, UPD_THREADS, 0, stream>>>(G, n, out, orth_type, work); break;
    default: return cudaErrorInvalidValue;
  }
  return cudaGetLastError();
}
cudaError_t launch_orth_polar(int R, const double* G, int n, double* out, cudaStream_t stream) {
  return launch_orth(R, G, n, out, PPLS_ORTH_SVD, nullptr, stream);
}

int update_grid_blocks(int ld) {
  int nb = (ld + GT - 1) / GT;
  if (nb > 96) nb = 96;
  if (nb < 1) nb = 1;
  return nb + 1;
}
size_t update_grid_scratch_doubles(int ld, int R) { return (size_t)3 * (update_grid_blocks(ld) - 1) * 2 * R * R + 4; }

// Start state: coefficients and Gram matrices for theta_0.
template <int R>
__device__ void prepare_body(const UpdateArgs& a) {
  __shared__ double scratch[UPD_WARPS * 36], S[R * R];
  FitState* st = a.st;
  block_gram<R>(a.L, a.p2, S, scratch);
  if (threadIdx.x < R * R) st->WtW[threadIdx.x] = S[threadIdx.x];
  __syncthreads();
  block_gram<R>(a.L + (size_t)a.p2 * R, a.q2, S, scratch);
  if (threadIdx.x < R * R) st->CtC[threadIdx.x] = S[threadIdx.x];
  if (threadIdx.x == 0) {
    posterior_coef(st, R);
    st->pass = 0; st->done = 0; st->steps_done = 0; st->status = PPLS_OK;
    bool ok = true;
    for (int k = 0; k < R; ++k) ok = ok && finite_d(st->aT[k]) && finite_d(st->bT[k]) && finite_d(st->aU[k]) && finite_d(st->bU[k]);
    if (!ok) { st->status = PPLS_E_NONFINITE; raise_done(st); }
  }
}
template <int R>
__global__ void __launch_bounds__(UPD_THREADS, 1) em_prepare_kernel(const UpdateArgs a) { prepare_body<R>(a); }

template <int R>
static void launch_batch_R(const BatchArgs& b, cudaStream_t stream) {
  em_update_batch_kernel<R><<<b.S, UPD_THREADS, 0, stream>>>(b);
}
cudaError_t launch_em_update_batch(const DataShape& s, int R, int RP, int g, int S, int tracelen, int prepare,
                                   const double* momvec_pack, double* L_pack, FitState* pack_st, FitState* st,
                                   double* momvec_s, double* L_s, double* Lscratch_s, double* trace, cudaStream_t stream,
                                   double* coef) {
  BatchArgs b;
  b.coef = coef;
  b.momvec_pack = momvec_pack; b.L_pack = L_pack; b.pack_st = pack_st; b.st = st; b.momvec_s = momvec_s; b.L_s = L_s;
  b.Lscratch_s = Lscratch_s; b.trace = trace; b.RP = RP; b.g = g; b.S = S; b.tracelen = tracelen; b.prepare = prepare;
  b.p = s.p; b.q = s.q; b.p2 = s.p2; b.q2 = s.q2; b.ld = s.ld;
  switch (R) {
    case 1: launch_batch_R<1>(b, stream); break;
    case 2: launch_batch_R<2>(b, stream); break;
    case 3: launch_batch_R<3>(b, stream); break;
    case 4: launch_batch_R<4>(b, stream); break;
    case 5: launch_batch_R<5>(b, stream); break;
    case 6: launch_batch_R<6>(b, stream); break;
    case 7: launch_batch_R<7>(b, stream); break;
    case 8: launch_batch_R<8>(b, stream); break;
    default: return cudaErrorInvalidValue;
  }
  // (the GEMM path, coef != null, never reads a pack's done flag: every start checks its own in the update block)
  if (!prepare && !coef) batch_done_kernel<<<(S / g + 64) / 64, 64, 0, stream>>>(pack_st, st, g, S);
  return cudaGetLastError();
}

// out[e] = sum over CTAs of partials[c][e], in a fixed order: slice y of 8 adds parts y, y+8, ...,
// then the 8 slices are added in order (run-to-run deterministic, 8x the parallelism of one thread
// per element).
__global__ void __launch_bounds__(256) reduce_partials_kernel(const double* __restrict__ partials, int nparts, int64_t len,
                                                              double* __restrict__ out, const FitState* st) {
  if (st->done) return;
  __shared__ double sh[8][33];
  const int64_t e = (int64_t)blockIdx.x * 32 + threadIdx.x;
  double s0 = 0.0, s1 = 0.0;
  if (e < len) {
    int c = threadIdx.y;
    for (; c + 8 < nparts; c += 16) { s0 += partials[(size_t)c * len + e]; s1 += partials[(size_t)(c + 8) * len + e]; }
    if (c < nparts) s0 += partials[(size_t)c * len + e];
  }
  sh[threadIdx.y][threadIdx.x] = s0 + s1;
  __syncthreads();
  if (threadIdx.y == 0 && e < len) {
    double t = 0.0;
#pragma unroll
    for (int y = 0; y < 8; ++y) t += sh[y][threadIdx.x];
    out[e] = t;
  }
}

// The same reduction, with the result stored straight into every rank's inbox slot for this rank (peer
// stores over NVLink; thread row y serves destination ranks y, y+8, ...).  The last block to finish raises
// this rank's flag on every rank (fence / counter / fence: all blocks' stores are visible system-wide first).
__global__ void __launch_bounds__(256) xsend_kernel(const double* __restrict__ partials, int nparts, int64_t len,
                                                    const XchgDev x, const FitState* st) {
  if (st->done) return;
  __shared__ double sh[8][33];
  const int64_t e = (int64_t)blockIdx.x * 32 + threadIdx.x;
  double s0 = 0.0, s1 = 0.0;
  if (e < len) {
    int c = threadIdx.y;
    for (; c + 8 < nparts; c += 16) { s0 += partials[(size_t)c * len + e]; s1 += partials[(size_t)(c + 8) * len + e]; }
    if (c < nparts) s0 += partials[(size_t)c * len + e];
  }
  sh[threadIdx.y][threadIdx.x] = s0 + s1;
  __syncthreads();
  const unsigned long long next = *x.epoch + 1;           // epoch moves only in the update kernel (stream order)
  if (e < len) {
    double t = 0.0;
#pragma unroll
    for (int y = 0; y < 8; ++y) t += sh[y][threadIdx.x];
    const size_t off = ((size_t)(next & 1ull) * x.world + x.rank) * x.slot + e;
    for (int r = threadIdx.y; r < x.world; r += 8) x.inbox[r][off] = t;
  }
  __syncthreads();
  __shared__ int last;
  if (threadIdx.x == 0 && threadIdx.y == 0) {
    __threadfence_system();                                 // cumulative: orders the whole block's stores (bar.sync above)
    const unsigned int prev = atomicAdd(x.counter, 1u);
    last = (prev == gridDim.x - 1);
    if (last) { *x.counter = 0; __threadfence_system(); }  // every block's stores are now ordered before the flags
  }
  __syncthreads();
  if (last && threadIdx.y == 0 && threadIdx.x < x.world) {  // one lane per destination rank, in parallel
    __threadfence_system();                                 // release pattern on the storing thread itself
    st_relaxed_sys(x.flags[threadIdx.x] + (next & 1ull) * XCHG_MAXW + x.rank, next);
  }
}

cudaError_t launch_xsend(const double* partials, int nparts, int64_t len, const XchgDev& x, const FitState* st,
                         cudaStream_t stream) {
  xsend_kernel<<<(unsigned)((len + 31) / 32), dim3(32, 8), 0, stream>>>(partials, nparts, len, x, st);
  return cudaGetLastError();
}

cudaError_t launch_reduce_partials(const double* partials, int nparts, int64_t len, double* momvec,
                                   const FitState* st, cudaStream_t stream) {
  reduce_partials_kernel<<<(unsigned)((len + 31) / 32), dim3(32, 8), 0, stream>>>(partials, nparts, len, momvec, st);
  return cudaGetLastError();
}

template <int R>
static void launch_update_R(const UpdateArgs& a, bool prepare, int orth_type, cudaStream_t stream) {
  if (prepare) em_prepare_kernel<R><<<1, UPD_THREADS, 0, stream>>>(a);
  else if (orth_type == PPLS_ORTH_SVD && a.gscratch && a.Lscratch) {
    void* args[] = {(void*)&a};
    cudaLaunchCooperativeKernel((const void*)em_update_grid_kernel<R>, dim3(update_grid_blocks(a.ld)), dim3(GT), args, 0, stream);
  } else em_update_kernel<R><<<1, UPD_THREADS, 0, stream>>>(a);
}

static cudaError_t dispatch_update(const DataShape& s, int R, const double* momvec, double* L,
                                   double* Lscratch, FitState* st, double* trace, bool prepare,
                                   int orth_type, double* gscratch, cudaStream_t stream, const XchgDev* xchg = nullptr) {
  UpdateArgs a;
  a.use_x = xchg ? 1 : 0;
  if (xchg) a.x = *xchg; else memset(&a.x, 0, sizeof(a.x));
  a.momvec = momvec; a.L = L; a.Lscratch = Lscratch; a.st = st; a.trace = trace; a.gscratch = gscratch;
  a.p = s.p; a.q = s.q; a.p2 = s.p2; a.q2 = s.q2; a.ld = s.ld;
  switch (R) {
    case 1: launch_update_R<1>(a, prepare, orth_type, stream); break;
    case 2: launch_update_R<2>(a, prepare, orth_type, stream); break;
    case 3: launch_update_R<3>(a, prepare, orth_type, stream); break;
    case 4: launch_update_R<4>(a, prepare, orth_type, stream); break;
    case 5: launch_update_R<5>(a, prepare, orth_type, stream); break;
    case 6: launch_update_R<6>(a, prepare, orth_type, stream); break;
    case 7: launch_update_R<7>(a, prepare, orth_type, stream); break;
    case 8: launch_update_R<8>(a, prepare, orth_type, stream); break;
    default: return cudaErrorInvalidValue;
  }
  return cudaGetLastError();
}

cudaError_t launch_em_update(const DataShape& s, int R, const double* momvec, double* L, double* Lscratch,
                             FitState* st, double* trace, int orth_type, double* gscratch, cudaStream_t stream,
                             const XchgDev* xchg) {
  return dispatch_update(s, R, momvec, L, Lscratch, st, trace, false, orth_type, gscratch, stream, xchg);
}
cudaError_t launch_em_prepare(const DataShape& s, int R, const double* L, FitState* st, cudaStream_t stream) {
  return dispatch_update(s, R, nullptr, const_cast<double*>(L), nullptr, st, nullptr, true, 0, nullptr, stream);
}

}  // namespace ppls
