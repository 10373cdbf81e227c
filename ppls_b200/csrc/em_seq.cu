// Update kernel of the SEQUENTIAL (one component at a time) PPLS fit -- the path PPLS_simult uses for its
// start values (Package/PPLS/R/EM_W_multi.R:762; SURVEY.md 8f "next" row 1).  FP64, -fmad=false.
//   PPLSi loop                 Package/PPLS/R/EM_W_multi.R:150-174 (logvalue, break rule, degenerate-variance exit)
//   EMstep_W coefficients      :62-70
//   EMstepC_fast               Package/PPLS/src/loglC.cpp:341-397 (expectations + maximisation, r = 1)
//   logl_W + loglC_fast        :297-323, loglC.cpp:319-338
//   deflation                  PPLS, :270-271 -- applied IMPLICITLY: the data pass runs on X with loadings that are
//                              orthogonal to the earlier components, and X'mu is projected the same way, so no
//                              deflated N x p copy is ever made.
// The data pass is the ordinary streaming kernel with R = 1 (em_stream / em_wide): pass k yields, for the
// parameters theta_k, G = [X'mu_T ; Y'mu_U] and the 2x2 Gram of (Xc w, Yc c).
#include "common.cuh"
#include <math.h>

namespace ppls {

constexpr int SEQ_THREADS = 512;

struct SeqArgs {
  const double* momvec;   // [ld | Gxx Gxy Gyy]
  double* L;              // [ld] loadings of the next pass (w rows then c rows)
  FitState* st;           // R = 1 state; ssqX / ssqY hold the DEFLATED sums of squares
  double* trace;          // logvalue[0 .. steps]
  const double* prev;     // [nprev][ld] loadings of the earlier components (unit norm, mutually orthogonal)
  int32_t nprev, p, q, p2, q2, ld, prepare;
  SeqCons cons;
};

__device__ double seq_block_sum(double v, double* scratch) {       // fixed tree: deterministic
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  __syncthreads();
  if (lane == 0) scratch[warp] = v;
  __syncthreads();
  double s = 0.0;
  for (int w = 0; w < SEQ_THREADS / 32; ++w) s += scratch[w];
  return s;
}

// v (rows [lo, hi) of a length-ld vector) <- (I - P P') v for the earlier components, one after the other
__device__ void seq_deflate(double* v, const double* prev, int nprev, int ld, int lo, int hi, double* scratch) {
  for (int j = 0; j < nprev; ++j) {
    const double* wj = prev + (size_t)j * ld;
    double d = 0.0;
    for (int e = lo + threadIdx.x; e < hi; e += SEQ_THREADS) d += wj[e] * v[e];
    d = seq_block_sum(d, scratch);
    for (int e = lo + threadIdx.x; e < hi; e += SEQ_THREADS) v[e] -= wj[e] * d;
    __syncthreads();
  }
}

// coefficients of mu_T, mu_U as EMstep_W (:62-70) + EMstepC_fast (loglC.cpp:357,362) spell them
__device__ void seq_coefficients(FitState* st) {
  const double B = st->b[0], sigT = st->t[0], sigX = st->sigE, sigY = st->sigF, sigH = st->sigH;
  const double sT2 = sigT * sigT, sT4 = sT2 * sT2, sT6 = sT4 * sT2, B2 = B * B;
  const double sX2 = sigX * sigX, sY2 = sigY * sigY, sH2 = sigH * sigH;
  const double g = sT2 * B2 + sH2;
  const double Kw = sT2 - sT4 * B2 / sY2 + sT4 * B2 * g / (sY2 * (g + sY2));
  const double Kc = g - sT4 * B2 / sX2 + sT6 * B2 / (sX2 * (sT2 + sX2));
  const double Kwc = sT2 * B / (sX2 * sY2) - Kc * sT2 * B / (sX2 * sY2 * (Kc + sY2)) -
                     sT4 * B / (sX2 * sY2 * (sT2 + sX2)) + Kc * sT4 * B / (sX2 * sY2 * (Kc + sY2) * (sT2 + sX2));
  const double c1 = Kw / (sX2 * (Kw + sX2)), c3 = Kc / (sY2 * (Kc + sY2)), c2 = Kwc;
  const double varU = sT2 * B * B + sH2;
  st->aT[0] = sT2 * (-c1 + -c2 * B + 1 / sX2);
  st->bT[0] = sT2 * (-c2 + -c3 * B + 1 / sY2 * B);
  st->aU[0] = -sT2 * B * c1 + -c2 * varU + 1 / sX2 * B * sT2;
  st->bU[0] = -c2 * B * sT2 + -c3 * varU + 1 / sY2 * varU;
  // kept for the expectations of the same pass
  st->WtW[0] = c1; st->WtW[1] = c2; st->WtW[2] = c3;
}

// logl_W(Xc, Yc, theta) (EM_W_multi.R:297-323: coefficients :312-320 with g re-squared; loglC.cpp:319-338) from the
// Gram of (Xw, Yc); st->ssqX / ssqY = ||Xc||^2, ||Yc||^2, N = rows
__device__ double seq_loglik(const FitState* st, double xx, double xy, double yy, double N, double p, double q) {
  const double B = st->b[0], sigT = st->t[0], sigX = st->sigE, sigY = st->sigF, sigH = st->sigH;
  const double sT2 = sigT * sigT, sT4 = sT2 * sT2, sT6 = sT4 * sT2, B2 = B * B;
  const double sX2 = sigX * sigX, sY2 = sigY * sigY, sH2 = sigH * sigH;
  double g = sqrt(sT2 * B2 + sH2); g = g * g;
  const double Kw = sT2 - sT4 * B2 / sY2 + sT4 * B2 * g / (sY2 * (g + sY2));
  const double Kc = g - sT4 * B2 / sX2 + sT6 * B2 / (sX2 * (sT2 + sX2));
  const double Kwc = sT2 * B / (sX2 * sY2) - Kc * sT2 * B / (sX2 * sY2 * (Kc + sY2)) -
                     sT4 * B / (sX2 * sY2 * (sT2 + sX2)) + Kc * sT4 * B / (sX2 * sY2 * (Kc + sY2) * (sT2 + sX2));
  const double c1 = Kw / (sX2 * (Kw + sX2)), c3 = Kc / (sY2 * (Kc + sY2)), c2 = Kwc;
  const double Logdiag = log(sX2 + sT2) + (p - 1) * log(sX2) + log(sY2 + Kc) + (q - 1) * log(sY2);
  double traceL = 1 / sX2 * st->ssqX + 1 / sY2 * st->ssqY;
  traceL += -c1 * xx - 2 * c2 * xy - c3 * yy;
  return -0.5 * N * (p + q) * log(2 * 3.14159265358979323846) - 0.5 * N * Logdiag - 0.5 * traceL;
}

// EMstepC_fast / meta_Estep expectations at theta_k (loglC.cpp:357-375 / :414-434) from the Gram of (Xw, Yc), then the
// maximisation of the scalars (loglC.cpp:377-386 / meta_Mstep :452-474) -> theta_{k+1} in st.  c1, c2, c3 of theta_k are
// in st->WtW[0..2], the mu coefficients in st->aT.. (seq_coefficients).
__device__ void seq_em_scalars(FitState* st, double xx, double xy, double yy, double N, double p, double q) {
  const double B = st->b[0], sigT = st->t[0], sigX = st->sigE, sigY = st->sigF, sigH = st->sigH;
  const double sT2 = sigT * sigT, sX2 = sigX * sigX, sY2 = sigY * sigY, sH2 = sigH * sigH;
  const double c1 = st->WtW[0], c2 = st->WtW[1], c3 = st->WtW[2];
  const double aT = st->aT[0], bT = st->bT[0], aU = st->aU[0], bU = st->bU[0];
  const double varU = sT2 * B * B + sH2;
  const double mTT = aT * aT * xx + 2.0 * aT * bT * xy + bT * bT * yy;
  const double mUU = aU * aU * xx + 2.0 * aU * bU * xy + bU * bU * yy;
  const double mUT = aU * aT * xx + (aU * bT + bU * aT) * xy + bU * bT * yy;
  const double Ctt = sT2 - sT2 * sT2 * (-c1 - 2 * B * c2 - B * B * (c3 - 1 / sY2) + 1 / sX2) + mTT / N;
  const double Cuu = varU - (-(c1 - 1 / sX2) * sT2 * sT2 * B * B - 2 * sT2 * B * varU * c2 - varU * varU * (c3 - 1 / sY2)) + mUU / N;
  const double Cut = sT2 * B - (-sT2 * sT2 * B * (c1 - 1 / sX2) - sT2 * sT2 * B * B * c2 - sT2 * varU * c2 -
                                varU * sT2 * B * (c3 - 1 / sY2)) + mUT / N;
  const double Ceetmp = c1 * c1 * sX2 * sX2 * xx + st->ssqX + c2 * c2 * sX2 * sX2 * yy - 2 * c1 * sX2 * xx +
                        2 * c1 * c2 * sX2 * sX2 * xy - 2 * c2 * sX2 * xy;
  const double Cee = sX2 - (-sX2 * sX2 * c1 + p * sX2) / p + Ceetmp / N / p;
  const double Cfftmp = c3 * c3 * sY2 * sY2 * yy + st->ssqY + c2 * c2 * sY2 * sY2 * xx - 2 * c3 * sY2 * yy +
                        2 * c3 * c2 * sY2 * sY2 * xy - 2 * c2 * sY2 * xy;
  const double Cff = sY2 - (-sY2 * sY2 * c3 + q * sY2) / q + Cfftmp / N / q;
  const double hX = -c2 * sH2, hY = -(c3 - 1 / sY2) * sH2;
  const double mHH = hX * hX * xx + 2.0 * hX * hY * xy + hY * hY * yy;
  const double Chh = sH2 - (-sH2 * sH2 * (c3 - 1 / sY2)) + mHH / N;
  st->Ctt[0] = Ctt; st->Cuu[0] = Cuu; st->Cut[0] = Cut; st->Cee = Cee; st->Cff = Cff; st->Chh[0] = Chh;
  st->b[0] = Cut / Ctt;
  st->sigE = sqrt(Cee); st->sigF = sqrt(Cff); st->sigH = sqrt(Chh); st->t[0] = sqrt(Ctt);
}

__global__ void __launch_bounds__(SEQ_THREADS, 1) seq_update_kernel(const SeqArgs a) {
  __shared__ double scratch[SEQ_THREADS / 32];
  __shared__ int s_done;
  FitState* st = a.st;
  if (a.prepare) {
    // start of a component: make the start loadings orthogonal to the earlier components (the reference
    // multiplies the DEFLATED data by them, :150 / :157), set up the coefficients
    seq_deflate(a.L, a.prev, a.nprev, a.ld, 0, a.p2, scratch);
    seq_deflate(a.L, a.prev, a.nprev, a.ld, a.p2, a.ld, scratch);
    if (threadIdx.x == 0) { seq_coefficients(st); st->pass = 0; st->done = 0; st->steps_done = 0; st->status = PPLS_OK; }
    return;
  }
  // one thread reads the flag for the block: thread 0 raises it further down
  if (threadIdx.x == 0) s_done = st->done;
  __syncthreads();
  if (s_done) return;
  __syncthreads();
  const double xx = a.momvec[a.ld], xy = a.momvec[a.ld + 1], yy = a.momvec[a.ld + 2];
  const double N = st->Ntot, p = (double)a.p, q = (double)a.q;
  if (threadIdx.x == 0) {
    const int k = st->pass;
    const double B = st->b[0], sigT = st->t[0], sigX = st->sigE, sigY = st->sigF, sigH = st->sigH;
    const double sT2 = sigT * sigT, sT4 = sT2 * sT2, sT6 = sT4 * sT2, B2 = B * B;
    const double sX2 = sigX * sigX, sY2 = sigY * sigY, sH2 = sigH * sigH;
    st->loglik_last = seq_loglik(st, xx, xy, yy, N, p, q);      // logvalue[k] = logl_W(Xc, Yc, theta_k), :150 / :171
    a.trace[k] = st->loglik_last;
    // ---- end of the component?  (:172 break rule, EMsteps exhausted, :153-155 vanishing noise variance) ----
    bool done = false;
    if (k >= 1) {                                        // critfunc(logvalue[i+1] - logvalue[i]) < atol, critfunc = I or abs
      const double incr = a.trace[k] - a.trace[k - 1];
      if ((a.cons.crit_abs ? fabs(incr) : incr) < st->atol) done = true;
    }
    if (k >= st->max_steps) done = true;
    if (!isfinite(st->loglik_last)) { st->status = PPLS_E_NONFINITE; done = true; }
    if (!done && (sigX < 100 * 2.220446049250313e-16 || sigY < 100 * 2.220446049250313e-16)) {
      st->status = PPLS_E_DEGENERATE; done = true;
    }
    st->gram_last[0] = xx; st->gram_last[1] = xy; st->gram_last[2] = yy;
    if (done) { raise_done(st); st->steps_done = k; }
    s_done = done ? 1 : 0;
    if (!done) {
      seq_em_scalars(st, xx, xy, yy, N, p, q);
      // constraints (:160-167): fixed values replace the estimates before the log-likelihood of the step is taken
      if (a.cons.mask & 4u) st->b[0] = a.cons.B;
      if (a.cons.mask & 8u) st->sigE = a.cons.sigE;
      if (a.cons.mask & 16u) st->sigF = a.cons.sigF;
      if (a.cons.mask & 32u) st->sigH = a.cons.sigH;
      if (a.cons.mask & 64u) st->t[0] = a.cons.sigT;
      st->pass = k + 1;
      const bool ok = isfinite(st->b[0]) && isfinite(st->sigE) && isfinite(st->sigF) && isfinite(st->sigH) && isfinite(st->t[0]);
      if (!ok) { st->status = PPLS_E_NONFINITE; raise_done(st); st->steps_done = k; s_done = 1; }
      else seq_coefficients(st);
    }
  }
  __syncthreads();
  if (s_done) return;
  // ---- W = normalized(t(Xc) mu_T / N), C = normalized(t(Yc) mu_U / N)  (loglC.cpp:358,363,381-382) ---------
  // (a constrained side keeps its fixed loadings, deflated like the data they multiply: Xc w = X (I - P P') w)
  for (int e = threadIdx.x; e < a.ld; e += SEQ_THREADS) {
    const bool fixed = (e < a.p2) ? (a.cons.mask & 1u) != 0 : (a.cons.mask & 2u) != 0;
    a.L[e] = fixed ? a.cons.L[e] : a.momvec[e];
  }
  __syncthreads();
  seq_deflate(a.L, a.prev, a.nprev, a.ld, 0, a.p2, scratch);          // t(Xc) = (I - P P') t(X)
  seq_deflate(a.L, a.prev, a.nprev, a.ld, a.p2, a.ld, scratch);
  for (int part = 0; part < 2; ++part) {
    const int lo = part ? a.p2 : 0, hi = part ? a.ld : a.p2;
    if (a.cons.mask & (part ? 2u : 1u)) continue;                    // block-uniform
    double s = 0.0;
    for (int e = lo + threadIdx.x; e < hi; e += SEQ_THREADS) { const double v = a.L[e] / N; a.L[e] = v; s += v * v; }
    s = seq_block_sum(s, scratch);
    const double nrm = sqrt(s);
    for (int e = lo + threadIdx.x; e < hi; e += SEQ_THREADS) a.L[e] = a.L[e] / nrm;
    __syncthreads();
  }
}

cudaError_t launch_seq_update(const DataShape& s, const double* momvec, double* L, FitState* st, double* trace,
                              const double* prev, int nprev, int prepare, cudaStream_t stream, const SeqCons* cons) {
  SeqArgs a;
  if (cons) a.cons = *cons;
  a.momvec = momvec; a.L = L; a.st = st; a.trace = trace; a.prev = prev; a.nprev = nprev; a.prepare = prepare;
  a.p = s.p; a.q = s.q; a.p2 = s.p2; a.q2 = s.q2; a.ld = s.ld;
  seq_update_kernel<<<1, SEQ_THREADS, 0, stream>>>(a);
  return cudaGetLastError();
}

// ---- multi-population ("meta") fit: meta_PPLSi / meta_EMstep (EM_W_multi.R:509-589 / :446-485), -----------------
// meta_Estep / meta_Mstep (loglC.cpp:399-474).  One component, loadings shared by all populations, scalars per
// population.  Pass k: the streaming kernel ran once per population (its own coefficients, the shared loadings)
// and left momvec_j = [X_j'mu_T ; Y_j'mu_U | Gram]; this kernel then does, for every population j, the scalar part
// of EMstepC_fast on (Gram_j, ssq_j, N_j), and pools the loadings:
//   W = orth(sum_j s_j Cxt_j), C = orth(sum_j s_j Cyu_j), Cxt_j = X_j'mu_T / N_j, s_j = sign(Cxt_1'Cxt_j)   (:481-482)
constexpr int META_MAXJ = 64;
struct MetaArgs {
  const double* momvecs;  // [J][ld + 3]
  double* L;              // [ld]
  FitState* sts;          // [J]; sts[0] carries the control words (pass, steps_done, status, max_steps, atol)
  double* trace;          // [max_steps + 1][J]: trace[k][j] = logl_W(X_j, Y_j, theta_k)
  int32_t J, p, q, p2, q2, ld, prepare;
};

__global__ void __launch_bounds__(SEQ_THREADS, 1) meta_update_kernel(const MetaArgs a) {
  __shared__ double scratch[SEQ_THREADS / 32];
  __shared__ double sgn[META_MAXJ];
  __shared__ int s_done, s_k;
  const int J = a.J, veclen = a.ld + 3;
  if (a.prepare) {
    if ((int)threadIdx.x < J) {
      FitState* st = a.sts + threadIdx.x;
      seq_coefficients(st);
      st->pass = 0; st->done = 0; st->steps_done = 0; st->status = PPLS_OK;
    }
    return;
  }
  if (threadIdx.x == 0) { s_done = a.sts[0].done; s_k = a.sts[0].pass; }
  __syncthreads();
  if (s_done) return;
  const int k = s_k;
  const double p = (double)a.p, q = (double)a.q;
  if ((int)threadIdx.x < J) {                                                                  // :574-576
    FitState* st = a.sts + threadIdx.x;
    const double* g = a.momvecs + (size_t)threadIdx.x * veclen + a.ld;
    st->loglik_last = seq_loglik(st, g[0], g[1], g[2], st->Ntot, p, q);
    a.trace[(size_t)k * J + threadIdx.x] = st->loglik_last;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    bool done = false, finite = true;
    double cur = 0.0, prev = 0.0;
    for (int j = 0; j < J; ++j) { cur += a.trace[(size_t)k * J + j]; finite = finite && isfinite(a.trace[(size_t)k * J + j]); }
    if (k >= 1) {
      for (int j = 0; j < J; ++j) prev += a.trace[(size_t)(k - 1) * J + j];
      if (k == 1) prev = prev * J;        // logvalue[1,] = rep(full-data log-likelihood, nlevels)  (:548)
      if (cur - prev < a.sts[0].atol) done = true;                                              // :578
    }
    if (k >= a.sts[0].max_steps) done = true;
    if (!finite) { a.sts[0].status = PPLS_E_NONFINITE; done = true; }
    if (done) { for (int j = 0; j < J; ++j) a.sts[j].done = 1; a.sts[0].steps_done = k; }
    s_done = done ? 1 : 0;
  }
  __syncthreads();
  if (s_done) return;
  if ((int)threadIdx.x < J) {                                        // meta_Estep + meta_Mstep scalars, next coefficients
    FitState* st = a.sts + threadIdx.x;
    const double* g = a.momvecs + (size_t)threadIdx.x * veclen + a.ld;
    seq_em_scalars(st, g[0], g[1], g[2], st->Ntot, p, q);
    const bool ok = isfinite(st->b[0]) && isfinite(st->sigE) && isfinite(st->sigF) && isfinite(st->sigH) && isfinite(st->t[0]);
    if (!ok) a.sts[0].status = PPLS_E_NONFINITE;
    seq_coefficients(st);
    st->pass = k + 1;
  }
  // signs from the X side
  const double N0 = a.sts[0].Ntot;
  for (int j = 0; j < J; ++j) {
    const double Nj = a.sts[j].Ntot;
    const double* g0 = a.momvecs;
    const double* gj = a.momvecs + (size_t)j * veclen;
    double d = 0.0;
    for (int e = threadIdx.x; e < a.p2; e += SEQ_THREADS) d += (g0[e] / N0) * (gj[e] / Nj);
    d = seq_block_sum(d, scratch);
    if (threadIdx.x == 0) sgn[j] = (d > 0.0) ? 1.0 : ((d < 0.0) ? -1.0 : 0.0);
  }
  __syncthreads();
  for (int part = 0; part < 2; ++part) {
    const int lo = part ? a.p2 : 0, hi = part ? a.ld : a.p2;
    double ss = 0.0;
    for (int e = lo + threadIdx.x; e < hi; e += SEQ_THREADS) {
      double v = 0.0;
      for (int j = 0; j < J; ++j) v += sgn[j] * (a.momvecs[(size_t)j * veclen + e] / a.sts[j].Ntot);
      a.L[e] = v; ss += v * v;
    }
    ss = seq_block_sum(ss, scratch);
    const double nrm = sqrt(ss);
    for (int e = lo + threadIdx.x; e < hi; e += SEQ_THREADS) a.L[e] = a.L[e] / nrm;
    __syncthreads();
  }
  if (threadIdx.x == 0 && a.sts[0].status != PPLS_OK) { for (int j = 0; j < J; ++j) a.sts[j].done = 1; a.sts[0].steps_done = k; }
}

cudaError_t launch_meta_update(const DataShape& s, int J, const double* momvecs, double* L, FitState* sts, double* trace,
                               int prepare, cudaStream_t stream) {
  if (J < 1 || J > META_MAXJ) return cudaErrorInvalidValue;
  MetaArgs a;
  a.momvecs = momvecs; a.L = L; a.sts = sts; a.trace = trace; a.J = J; a.prepare = prepare;
  a.p = s.p; a.q = s.q; a.p2 = s.p2; a.q2 = s.q2; a.ld = s.ld;
  meta_update_kernel<<<1, SEQ_THREADS, 0, stream>>>(a);
  return cudaGetLastError();
}

}  // namespace ppls
