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
    // ---- logvalue[k] = logl_W(Xc, Yc, theta_k)  (:150 / :171; coefficients :312-320 with g re-squared) ----
    {
      double g = sqrt(sT2 * B2 + sH2); g = g * g;
      const double Kw = sT2 - sT4 * B2 / sY2 + sT4 * B2 * g / (sY2 * (g + sY2));
      const double Kc = g - sT4 * B2 / sX2 + sT6 * B2 / (sX2 * (sT2 + sX2));
      const double Kwc = sT2 * B / (sX2 * sY2) - Kc * sT2 * B / (sX2 * sY2 * (Kc + sY2)) -
                         sT4 * B / (sX2 * sY2 * (sT2 + sX2)) + Kc * sT4 * B / (sX2 * sY2 * (Kc + sY2) * (sT2 + sX2));
      const double c1 = Kw / (sX2 * (Kw + sX2)), c3 = Kc / (sY2 * (Kc + sY2)), c2 = Kwc;
      const double Logdiag = log(sX2 + sT2) + (p - 1) * log(sX2) + log(sY2 + Kc) + (q - 1) * log(sY2);
      double traceL = 1 / sX2 * st->ssqX + 1 / sY2 * st->ssqY;
      traceL += -c1 * xx - 2 * c2 * xy - c3 * yy;
      const double ll = -0.5 * N * (p + q) * log(2 * 3.14159265358979323846) - 0.5 * N * Logdiag - 0.5 * traceL;
      st->loglik_last = ll;
      a.trace[k] = ll;
    }
    // ---- end of the component?  (:172 break rule, EMsteps exhausted, :153-155 vanishing noise variance) ----
    bool done = false;
    if (k >= 1 && (a.trace[k] - a.trace[k - 1]) < st->atol) done = true;
    if (k >= st->max_steps) done = true;
    if (!isfinite(st->loglik_last)) { st->status = PPLS_E_NONFINITE; done = true; }
    if (!done && (sigX < 100 * 2.220446049250313e-16 || sigY < 100 * 2.220446049250313e-16)) {
      st->status = PPLS_E_DEGENERATE; done = true;
    }
    if (done) { st->done = 1; st->steps_done = k; }
    s_done = done ? 1 : 0;
    if (!done) {
      // ---- EMstepC_fast expectations at theta_k (loglC.cpp:357-375) from the Gram of (Xw, Yc) -----------
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
      // ---- maximisation (loglC.cpp:377-386) -> theta_{k+1} -----------------------------------------------
      st->b[0] = Cut / Ctt;
      st->sigE = sqrt(Cee); st->sigF = sqrt(Cff); st->sigH = sqrt(Chh); st->t[0] = sqrt(Ctt);
      st->pass = k + 1;
      const bool ok = isfinite(st->b[0]) && isfinite(st->sigE) && isfinite(st->sigF) && isfinite(st->sigH) && isfinite(st->t[0]);
      if (!ok) { st->status = PPLS_E_NONFINITE; st->done = 1; st->steps_done = k; s_done = 1; }
      else seq_coefficients(st);
    }
  }
  __syncthreads();
  if (s_done) return;
  // ---- W = normalized(t(Xc) mu_T / N), C = normalized(t(Yc) mu_U / N)  (loglC.cpp:358,363,381-382) ---------
  for (int e = threadIdx.x; e < a.ld; e += SEQ_THREADS) a.L[e] = a.momvec[e];
  __syncthreads();
  seq_deflate(a.L, a.prev, a.nprev, a.ld, 0, a.p2, scratch);          // t(Xc) = (I - P P') t(X)
  seq_deflate(a.L, a.prev, a.nprev, a.ld, a.p2, a.ld, scratch);
  for (int part = 0; part < 2; ++part) {
    const int lo = part ? a.p2 : 0, hi = part ? a.ld : a.p2;
    double s = 0.0;
    for (int e = lo + threadIdx.x; e < hi; e += SEQ_THREADS) { const double v = a.L[e] / N; a.L[e] = v; s += v * v; }
    s = seq_block_sum(s, scratch);
    const double nrm = sqrt(s);
    for (int e = lo + threadIdx.x; e < hi; e += SEQ_THREADS) a.L[e] = a.L[e] / nrm;
    __syncthreads();
  }
}

cudaError_t launch_seq_update(const DataShape& s, const double* momvec, double* L, FitState* st, double* trace,
                              const double* prev, int nprev, int prepare, cudaStream_t stream) {
  SeqArgs a;
  a.momvec = momvec; a.L = L; a.st = st; a.trace = trace; a.prev = prev; a.nprev = nprev; a.prepare = prepare;
  a.p = s.p; a.q = s.q; a.p2 = s.p2; a.q2 = s.q2; a.ld = s.ld;
  seq_update_kernel<<<1, SEQ_THREADS, 0, stream>>>(a);
  return cudaGetLastError();
}

}  // namespace ppls
