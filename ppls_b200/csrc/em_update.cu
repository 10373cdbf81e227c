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
#include "em_update_dev.cuh"
#include <math.h>
#include <string.h>
#include <cooperative_groups.h>
namespace cg = cooperative_groups;

namespace ppls {

constexpr int UPD_THREADS = 512;
constexpr int UPD_WARPS = UPD_THREADS / 32;

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

template <int R>
__device__ void update_body(const UpdateArgs& a) {
  __shared__ double scratch[UPD_WARPS * 36], S[R * R], Mat[R * R];
  __shared__ int s_done;
  __shared__ ScalarShared ssh;
  FitState* st = a.st;
  if (st->done) return;
  if (threadIdx.x < 64) { const int d = update_scalars<R>(a, &ssh); if (threadIdx.x == 0) { s_done = d; if (d) raise_done(st); } }
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
__device__ void grid_sum_gram(const double* parts /*[nb][2][R*R]*/, int nb, double* S0, double* S1) {
  // Both sides at once, EIGHT threads per element: thread `sub` of an element adds blocks sub, sub + 8, ... (independent
  // loads, so their L2 latencies overlap), then a fixed-order butterfly.  One thread per element walking all the blocks
  // (the first version) cost ~10 us per call at C4, where 96 row blocks take part -- four calls per update.
  constexpr int NV = R * (R + 1) / 2;
  constexpr int TOTAL = 2 * NV * 8;
  for (int base = 0; base < TOTAL; base += GT) {               // (whole warps take every trip: the shuffles need all lanes)
    const int idx = base + (int)threadIdx.x;
    const bool valid = idx < TOTAL;
    const int which = valid ? idx / (NV * 8) : 0, e = valid ? (idx >> 3) % NV : 0, sub = idx & 7;
    double s0 = 0.0, s1 = 0.0;
    if (valid) {
      const double* src = parts + (size_t)which * R * R + e;
      int b = sub;
#pragma unroll 4
      for (; b + 8 < nb; b += 16) { s0 += src[(size_t)b * 2 * R * R]; s1 += src[(size_t)(b + 8) * 2 * R * R]; }
      if (b < nb) s0 += src[(size_t)b * 2 * R * R];
    }
    double sacc = s0 + s1;
    sacc += __shfl_xor_sync(0xffffffffu, sacc, 1);
    sacc += __shfl_xor_sync(0xffffffffu, sacc, 2);
    sacc += __shfl_xor_sync(0xffffffffu, sacc, 4);
    if (valid && sub == 0) {
      int ee = e, k = 0;
      while (ee >= R - k) { ee -= R - k; ++k; }
      const int l = k + ee;
      double* S = which ? S1 : S0;
      S[k * R + l] = sacc; S[l * R + k] = sacc;
    }
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
    if (threadIdx.x < 64) { const int d = update_scalars<R>(a, &ssh); if (threadIdx.x == 0) *done_now = d; }
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
    grid_sum_gram<R>(ph0, nb, S[0], S[1]);
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
  grid_sum_gram<R>(ph1, nb, S[0], S[1]);
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

// ---- orth(G, type) without any allocation (ppls_orth_noalloc) --------------------------------------------------------
// The QR form needs a working copy of G (destroyed by the Householder sweeps): `work` above, which ppls_orth gets from
// cudaMalloc.  Callers that may not allocate (code under CUDA graph capture, tight serving loops) get the same kernel
// with that scratch in a file-scope device array of ORTH_STATIC_MAX doubles (512 KB; every GPU that loads the library
// holds it until exit).  All calls share the array: calls that overlap on the GPU race.  Nothing in it has to start at
// zero (the kernel overwrites the part it uses with a copy of G before reading it).
constexpr long long ORTH_STATIC_MAX = PPLS_ORTH_NOALLOC_MAX_ELEMS;
static __device__ __align__(256) double g_orth_scratch[ORTH_STATIC_MAX];

// Device address of the scratch on the CURRENT device.  In host code `g_orth_scratch` is a host-side stand-in, not a
// device address: the address comes from cudaGetSymbolAddress, once per device, with that device current.
static cudaError_t orth_scratch_ptr(double** out) {
  static double* ptrs[64] = {};
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
  if (!ptrs[dev]) {
    double* p = nullptr;
    e = cudaGetSymbolAddress((void**)&p, g_orth_scratch);
    if (e != cudaSuccess) return e;
    ptrs[dev] = p;
  }
  *out = ptrs[dev];
  return cudaSuccess;
}

long long orth_static_max_elems() { return ORTH_STATIC_MAX; }

// G, out: DEVICE pointers, [n][R] row-major.  No cudaMalloc / cudaFree / allocating library call.
cudaError_t launch_orth_static(int R, const double* G, int n, double* out, int orth_type, cudaStream_t stream) {
  if ((long long)n * R > ORTH_STATIC_MAX) return cudaErrorInvalidValue;
  double* work = nullptr;
  cudaError_t e = orth_scratch_ptr(&work);
  if (e != cudaSuccess) return e;
  return launch_orth(R, G, n, out, orth_type, work, stream);
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
