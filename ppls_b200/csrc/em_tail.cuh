// Fused tail of the streaming kernel: ONE launch per EM iteration.
//
// After its last tile every CTA of em_stream_kernel has written its partial moment vector.  Instead of ending the
// launch there and running two more kernels (cross-CTA reduction / peer send, then the update K2), the CTAs -- all
// co-resident: the grid is at most one CTA per SM and the launch is cooperative -- go on together:
//   barrier | every row CTA sums ITS rows of G = [X'mu_T ; Y'mu_U] over the CTAs' partials in fixed order (CTA 0: the
//   Gram tail Gxx, Gxy, Gyy) | multi-GPU: the sums go straight into every rank's inbox by peer stores; CTA b raises ITS flag
//   on every rank and waits for CTA b of the peers (the same rows everywhere: no grid barrier), adds the ranks in rank order | CTA 0: the scalar part of the update (log-likelihood,
//   E-step moments, break rule, M-step scalars: update_scalars of em_update_dev.cuh) while the row CTAs form the
//   partial Gram of their rows | barrier | T = G (G'G)^{-1/2} on the rows held in shared memory, partial Gram of T |
//   barrier | one Newton-Schulz step L = T (3I - T'T)/2 written to the loadings (Maximiz_M, EM_W_multi.R:732-733, type
//   "SVD"), W'W and C'C by CTA 0.
// Same arithmetic as em_update_grid_kernel (em_update.cu), same fixed summation orders (run-to-run deterministic,
// bitwise identical on all ranks), three grid barriers (four with the exchange) instead of two kernel boundaries
// and three grid syncs.  Compiled with -fmad=false like everything that evaluates the closed-form coefficients.
#pragma once
#include "em_update_dev.cuh"

namespace ppls {

struct TailArgs {
  double* momvec;               // [ld*R + 3R^2] this pass' reduced vector (kept for diagnostics / ppls_fit users)
  double* L;                    // [ld][R] loadings: read by the streaming phase, overwritten with theta_{k+1}
  FitState* st;
  double* trace;
  double* gscratch;             // [2][G][2][R*R] Gram partials | u64 control words: arrivals of all CTAs, arrivals of the row
                                // CTAs, scalar-part flag | clocks: streaming ns, tail ns, launches, start stamp, scalar ns
  int32_t p, q;
  int32_t use_x;
  XchgDev x;
};

__host__ __device__ inline size_t tail_scratch_doubles(int grid, int R) { return (size_t)2 * grid * 2 * R * R + 16; }
__host__ __device__ inline size_t tail_clock_offset(int grid, int R) { return (size_t)2 * grid * 2 * R * R + 3; }   // doubles
__host__ __device__ inline int tail_rows_per_cta(int ld, int grid) { return (ld + grid - 2) / (grid - 1); }
// dynamic shared memory the tail needs (it reuses the streaming phase's carve-up from offset 0)
__host__ __device__ inline size_t tail_smem_bytes(int R, int ld, int grid, int threads) {
  const size_t rows = (size_t)tail_rows_per_cta(ld, grid);
  const size_t nw = (size_t)threads / 32;
  size_t d = 2 * rows * R            // Gs, Ts
             + nw * 2 * R * R        // per-warp reduction scratch
             + 2 * 2 * R * R         // S[2], Mat[2]
             + 2 * 3 * R * R         // Newton-Schulz work
             + 2 * R * R             // SM (CTA 0)
             + sizeof(ScalarShared) / 8 + 2;
  return d * 8 + 64;
}

// Barriers of the tail.  All CTAs are co-resident (cooperative launch, one CTA per SM).  The counters only grow: pass
// k of a fit (every launch that gets here runs the same number of barriers) waits for a target derived from k, so
// nothing is ever reset and a CTA that only WAITS (CTA 0 on the row CTAs' counter) needs no arrival of its own.
__device__ __forceinline__ void tail_arrive_wait(unsigned long long* cnt, unsigned long long target, bool sys) {
  __syncthreads();
  if (threadIdx.x == 0) {
    if (sys) __threadfence_system(); else __threadfence();
    atomicAdd(cnt, 1ull);
    while (*reinterpret_cast<volatile unsigned long long*>(cnt) < target) { }
    __threadfence();
  }
  __syncthreads();
}
__device__ __forceinline__ void tail_wait(const unsigned long long* cnt, unsigned long long target) {
  if (threadIdx.x == 0) {
    while (*reinterpret_cast<const volatile unsigned long long*>(cnt) < target) { }
    __threadfence();
  }
  __syncthreads();
}

// Sum over the CTA of nv (<= 2 R^2) per-thread values, fixed tree: warp shuffles, then the warps in order.
// Result in out[0..nv) (shared); every thread of the CTA must call.
__device__ __forceinline__ void tail_block_reduce(const double* vals, int nv, double* out, double* scratch /*[nwarps][nv]*/) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  for (int v = 0; v < nv; ++v) {
    double x = vals[v];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
    if (lane == 0) scratch[warp * nv + v] = x;
  }
  __syncthreads();
  for (int v = threadIdx.x; v < nv; v += blockDim.x) {
    double s = 0.0;
    for (int w = 0; w < nw; ++w) s += scratch[w * nv + v];
    out[v] = s;
  }
  __syncthreads();
}

// val(e) summed over c = 0 .. n-1 for elements e = 0 .. nelem-1, 8 lanes per element: lane y adds c = y, y+8, ... in
// order, then the 8 partial sums are added by a fixed butterfly (every lane of the group ends with the same bits).
// f(e, c) loads one term; sink(e, total, y) is called on all 8 lanes of the group.
template <class Load, class Sink>
__device__ __forceinline__ void tail_sum8(int nelem, int n, Load f, Sink sink) {
  const int y = threadIdx.x & 7, g = threadIdx.x >> 3, ng = blockDim.x >> 3;
  for (int base = 0; base < nelem; base += ng) {
    const int e = base + g;
    const bool act = e < nelem;
    double s0 = 0.0, s1 = 0.0;
    if (act) {
      for (int c0 = y; c0 < n; c0 += 64) {        // 8 loads in flight per lane, then added in the fixed order
        double v[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) { const int c = c0 + 8 * j; v[j] = (c < n) ? f(e, c) : 0.0; }
#pragma unroll
        for (int j = 0; j < 8; j += 2) { s0 += v[j]; s1 += v[j + 1]; }
      }
    }
    double s = s0 + s1;
    s += __shfl_xor_sync(0xffffffffu, s, 1);
    s += __shfl_xor_sync(0xffffffffu, s, 2);
    s += __shfl_xor_sync(0xffffffffu, s, 4);
    if (act) sink(e, s, y);
  }
}

// packed upper triangle index -> (k, l), k <= l
template <int R>
__device__ __forceinline__ void tail_unpack(int e, int& k, int& l) {
  k = 0;
  while (e >= R - k) { e -= R - k; ++k; }
  l = k + e;
}

template <int R>
__device__ __noinline__ void stream_tail(const TailArgs& t, const double* partials, int ld, int p2, unsigned char* smem) {
  constexpr int NV = R * (R + 1) / 2, RR = R * R;
  const int tid = threadIdx.x, nt = blockDim.x, nw = nt >> 5;
  const int G = gridDim.x, b = blockIdx.x, nrb = G - 1, rb = b - 1;
  const int rows_per = tail_rows_per_cta(ld, G);
  const int row_lo = rb >= 0 ? min(rb * rows_per, ld) : 0;
  const int row_hi = rb >= 0 ? min(row_lo + rows_per, ld) : 0;
  const int nrows = row_hi - row_lo;
  const long long len = (long long)ld * R + 3 * RR;
  FitState* st = t.st;
  // ---- shared-memory carve-up (the streaming phase is over: its ring is free) -----------------------------
  double* sm = reinterpret_cast<double*>(smem);
  double* Gs = sm;            sm += (size_t)rows_per * R;
  double* Ts = sm;            sm += (size_t)rows_per * R;
  double* red = sm;           sm += (size_t)nw * 2 * RR;
  double* S = sm;             sm += 2 * RR;
  double* Mat = sm;           sm += 2 * RR;
  double* nswork = sm;        sm += 2 * 3 * RR;
  double* SMm = sm;           sm += 2 * RR;
  ScalarShared* ssh = reinterpret_cast<ScalarShared*>(sm);
  // ---- global scratch: Gram partials, then the control words (u64) ------------------------------------------
  double* ph0 = t.gscratch;
  double* ph1 = ph0 + (size_t)G * 2 * RR;
  unsigned long long* ctl = reinterpret_cast<unsigned long long*>(ph1 + (size_t)G * 2 * RR);
  unsigned long long* cntA = ctl + 0;     // arrivals of ALL CTAs
  unsigned long long* cntR = ctl + 1;     // arrivals of the row CTAs
  unsigned long long* flagS = ctl + 2;    // 2 (pass + 1) + done, stored by CTA 0 when the scalar part is through
  unsigned long long* clk = ctl + 3;      // streaming ns, tail ns, launches, start stamp, scalar-part ns
  const bool xon = t.use_x != 0;
  const unsigned long long pass = (unsigned long long)st->pass;     // CTA 0 moves it only after barrier A
  const unsigned long long next = xon ? (*t.x.epoch + 1) : 0ull;    // ... and the epoch likewise
  const unsigned long long NA = 1ull;

  unsigned long long t1 = 0;
  tail_arrive_wait(cntA, (unsigned long long)G * (NA * pass + 1), false);     // A: every CTA's partial vector is in global memory
  if (b == 0 && tid == 0) { t1 = global_timer_ns(); clk[0] += t1 - clk[3]; }

  // ---- own elements of the reduced vector: rows [row_lo, row_hi) of G (row CTAs), the Gram tail (CTA 0) --------
  const long long e0 = rb >= 0 ? (long long)row_lo * R : (long long)ld * R;
  const int nelem = rb >= 0 ? nrows * R : 3 * RR;
  double* mine = rb >= 0 ? Gs : red;                                // CTA 0 parks the tail in `red` for a moment
  if (!xon) {
    tail_sum8(nelem, G,
              [&](int e, int c) { return __ldcg(partials + (size_t)c * len + e0 + e); },
              [&](int e, double s, int y) { if (y == 0) { mine[e] = s; t.momvec[e0 + e] = s; } });
  } else {
    const size_t off = ((size_t)(next & 1ull) * t.x.world + t.x.rank) * t.x.slot + (size_t)e0;
    tail_sum8(nelem, G,
              [&](int e, int c) { return __ldcg(partials + (size_t)c * len + e0 + e); },
              [&](int e, double s, int y) { for (int r = y; r < t.x.world; r += 8) t.x.inbox[r][off + e] = s; });
    // CTA b of every rank owns the SAME elements: it needs CTA b of the peers only -- per-CTA flags, no grid barrier.
    __syncthreads();                                                // the whole CTA's peer stores are issued
    unsigned long long tx = 0;                                      // (CTA 0, thread 0: stamps of the exchange, clk[8..11])
    if (b == 0 && tid == 0) { tx = global_timer_ns(); clk[8] += tx - t1; }
    if (tid < t.x.world) {
      __threadfence_system();                                       // (cumulative over the bar.sync: orders them before the flag)
      st_relaxed_sys(t.x.cflags[tid] + ((next & 1ull) * XCHG_MAXW + t.x.rank) * XCHG_MAXCTA + b, next);
    }
    if (b == 0 && tid == 0) { const unsigned long long ty = global_timer_ns(); clk[9] += ty - tx; tx = ty; }
    if (tid < 32) xchg_wait_cta(t.x, next, b);                      // the peers' CTA b has delivered
    if (b == 0 && tid == 0) { const unsigned long long ty = global_timer_ns(); clk[10] += ty - tx; tx = ty; }
    __syncthreads();
    for (int e = tid; e < nelem; e += nt) {                         // ranks added in rank order: same bits everywhere
      const double s = xchg_sum(t.x, next, e0 + e);
      mine[e] = s; t.momvec[e0 + e] = s;
    }
    if (b == 0 && tid == 0) clk[11] += global_timer_ns() - tx;
  }
  __syncthreads();

  if (rb < 0) {
    // ================= CTA 0: the scalar part, off the row CTAs' path =================
    volatile int* s_done = reinterpret_cast<volatile int*>(ssh + 1);
    if (tid < 64) {
      UpdateArgs ua;
      ua.momvec = t.momvec; ua.L = t.L; ua.Lscratch = nullptr; ua.st = st; ua.trace = t.trace; ua.gscratch = nullptr;
      ua.p = t.p; ua.q = t.q; ua.p2 = p2; ua.q2 = ld - p2; ua.ld = ld; ua.use_x = 0;
      int d = update_scalars<R>(ua, ssh);
      if (tid == 0) {
        if (xon) {
          if (*reinterpret_cast<volatile int*>(t.x.failed)) { st->status = PPLS_E_NCCL; d = 1; }   // a peer never answered
          *t.x.epoch = next;
        }
        *s_done = d;
        __threadfence();
        *reinterpret_cast<volatile unsigned long long*>(flagS) = 2ull * (pass + 1) + (d ? 1ull : 0ull);
        clk[4] += global_timer_ns() - t1;
        clk[5] += ssh->ts[0] - t1; clk[6] += ssh->ts[1] - ssh->ts[0]; clk[7] += ssh->ts[2] - ssh->ts[1];   // own sums | staging | arithmetic
      }
    }
    __syncthreads();
    if (*s_done) {                                                  // the run ended with this pass: theta_k stays
      if (tid == 0) { raise_done(st); clk[1] += global_timer_ns() - t1; clk[2] += 1ull; }
      return;
    }
    // W'W, C'C of the new loadings from the row CTAs' second Gram: M S M with M = (3I - S)/2
    tail_wait(cntR, (unsigned long long)nrb * (2 * pass + 2));
    tail_sum8(2 * NV, nrb,
              [&](int e, int c) { return __ldcg(ph1 + ((size_t)c * 2 + e / NV) * RR + e % NV); },
              [&](int e, double s, int y) {
                if (y == 0) { int k, l; tail_unpack<R>(e % NV, k, l); double* Sw = S + (e / NV) * RR; Sw[k * R + l] = s; Sw[l * R + k] = s; }
              });
    __syncthreads();
    for (int e = tid; e < 2 * RR; e += nt) {
      const int i = (e % RR) / R, j = (e % RR) % R;
      Mat[e] = (i == j ? 1.5 : 0.0) - 0.5 * S[e];
    }
    __syncthreads();
    for (int e = tid; e < 2 * RR; e += nt) {                         // S M
      const int w = e / RR, i = (e % RR) / R, j = (e % RR) % R;
      double sacc = 0.0;
      for (int k = 0; k < R; ++k) sacc += S[w * RR + i * R + k] * Mat[w * RR + k * R + j];
      SMm[e] = sacc;
    }
    __syncthreads();
    for (int e = tid; e < 2 * RR; e += nt) {                         // M (S M); M is symmetric
      const int w = e / RR, i = (e % RR) / R, j = (e % RR) % R;
      double sacc = 0.0;
      for (int k = 0; k < R; ++k) sacc += Mat[w * RR + i * R + k] * SMm[w * RR + k * R + j];
      (w ? st->CtC : st->WtW)[e % RR] = sacc;
      nswork[e] = sacc;
    }
    __syncthreads();
    if (tid == 0) {
      bool ok = true;
      for (int e = 0; e < 2 * RR; ++e) ok = ok && finite_d(nswork[e]);
      if (!ok) { st->status = PPLS_E_NONFINITE; raise_done(st); st->steps_done = st->pass - 1; }
      clk[1] += global_timer_ns() - t1; clk[2] += 1ull;
    }
    return;
  }

  // ================= row CTAs: W = orth(X'mu_T), C = orth(Y'mu_U) on their rows (Maximiz_M, :732-733) =================
  {
    double v[2 * NV];
#pragma unroll
    for (int e = 0; e < 2 * NV; ++e) v[e] = 0.0;
    for (int j = tid; j < nrows; j += nt) {
      const int w = (row_lo + j >= p2) ? NV : 0;
      double row[R];
#pragma unroll
      for (int k = 0; k < R; ++k) row[k] = Gs[j * R + k];
      int e = 0;
#pragma unroll
      for (int k = 0; k < R; ++k)
#pragma unroll
        for (int l = k; l < R; ++l) { v[w + e] += row[k] * row[l]; ++e; }
    }
    tail_block_reduce(v, 2 * NV, S, red);
    if (tid < 2 * NV) ph0[((size_t)rb * 2 + tid / NV) * RR + tid % NV] = S[tid];
  }
  tail_arrive_wait(cntR, (unsigned long long)nrb * (2 * pass + 1), false);     // B: every row CTA's Gram partial of G

  // ---- T = G (G'G)^{-1/2} on the rows in shared memory, partial Gram of T ---------------------------------------
  tail_sum8(2 * NV, nrb,
            [&](int e, int c) { return __ldcg(ph0 + ((size_t)c * 2 + e / NV) * RR + e % NV); },
            [&](int e, double s, int y) {
              if (y == 0) { int k, l; tail_unpack<R>(e % NV, k, l); double* Sw = S + (e / NV) * RR; Sw[k * R + l] = s; Sw[l * R + k] = s; }
            });
  __syncthreads();
  block_inv_sqrt<R, 2>(S, Mat, nswork);
  {
    double v[2 * NV];
#pragma unroll
    for (int e = 0; e < 2 * NV; ++e) v[e] = 0.0;
    for (int j = tid; j < nrows; j += nt) {
      const int wm = (row_lo + j >= p2) ? 1 : 0;
      const double* M = Mat + wm * RR;
      double row[R], res[R];
#pragma unroll
      for (int k = 0; k < R; ++k) row[k] = Gs[j * R + k];
#pragma unroll
      for (int l = 0; l < R; ++l) {
        double sacc = 0.0;
#pragma unroll
        for (int k = 0; k < R; ++k) sacc += row[k] * M[k * R + l];
        res[l] = sacc;
      }
      int e = 0;
#pragma unroll
      for (int k = 0; k < R; ++k) {
        Ts[j * R + k] = res[k];
#pragma unroll
        for (int l = k; l < R; ++l) { v[wm * NV + e] += res[k] * res[l]; ++e; }
      }
    }
    tail_block_reduce(v, 2 * NV, S, red);
    if (tid < 2 * NV) ph1[((size_t)rb * 2 + tid / NV) * RR + tid % NV] = S[tid];
  }
  tail_arrive_wait(cntR, (unsigned long long)nrb * (2 * pass + 2), false);     // C: every row CTA's Gram partial of T

  // ---- one Newton-Schulz step L = T M, M = (3I - S)/2, S = T'T --------------------------------------------------
  tail_sum8(2 * NV, nrb,
            [&](int e, int c) { return __ldcg(ph1 + ((size_t)c * 2 + e / NV) * RR + e % NV); },
            [&](int e, double s, int y) {
              if (y == 0) { int k, l; tail_unpack<R>(e % NV, k, l); double* Sw = S + (e / NV) * RR; Sw[k * R + l] = s; Sw[l * R + k] = s; }
            });
  __syncthreads();
  for (int e = tid; e < 2 * RR; e += nt) {
    const int i = (e % RR) / R, j = (e % RR) % R;
    Mat[e] = (i == j ? 1.5 : 0.0) - 0.5 * S[e];
  }
  // theta_k's loadings stay untouched until CTA 0 has decided whether the run ends with this pass
  tail_wait(flagS, 2ull * (pass + 1));
  if (*reinterpret_cast<volatile unsigned long long*>(flagS) & 1ull) return;
  for (int j = tid; j < nrows; j += nt) {
    const double* M = Mat + ((row_lo + j >= p2) ? RR : 0);
    double row[R], res[R];
#pragma unroll
    for (int k = 0; k < R; ++k) row[k] = Ts[j * R + k];
#pragma unroll
    for (int l = 0; l < R; ++l) {
      double sacc = 0.0;
#pragma unroll
      for (int k = 0; k < R; ++k) sacc += row[k] * M[k * R + l];
      res[l] = sacc;
    }
#pragma unroll
    for (int l = 0; l < R; ++l) t.L[(size_t)(row_lo + j) * R + l] = res[l];
  }
}

}  // namespace ppls
