// K1 "em_stream": one pass over a row shard of [X Y] per EM iteration, FP64, sm_100a.
//
// Replaces, for the parameters theta_k held in FitState/L:
//   Xw = X W, Yc = Y C                         Package/PPLS/R/EM_W_multi.R:689-690
//   mu_T, mu_U (posterior latent means)        :691-694
//   t(X) %*% mu_T, t(Y) %*% mu_U               :732-733 (Maximiz_M)
//   every crossprod(mu_.) of :696-712 and the XW/YC sums of
//   loglC_fast, Package/PPLS/src/loglC.cpp:332-335, through the 2r x 2r Gram of [Xw Yc].
// The reference touches X and Y ~24 times per iteration for this; here each byte is read
// from HBM exactly once (cp.async.bulk -> shared-memory ring), see DESIGN.md.
//
// Mapping.  A tile is T consecutive rows = one contiguous T*ld*8-byte block of HBM.
// Warps split the COLUMNS: warp w owns M chunks of 64 columns (lane owns a column pair per
// chunk), X warps and Y warps separately, and keeps for its columns the loadings (w) and
// the running G = X' mu accumulators (acc) in registers for the whole kernel.
//   phase A: partial projections part[i][k] += x[i,j] w[j,k] over the thread's columns,
//            reduced over the CTA through shared memory (transpose-reduce per warp, then
//            across warps by warp 0), which also forms mu_T/mu_U and the Gram terms;
//   phase B: acc[j][k] += x[i,j] mu[i,k] re-reading the tile from shared memory.
#include "common.cuh"
#include <cstdlib>
#include <cstdio>

namespace ppls {

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
      ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  do {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
  } while (!ok);
}
__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

constexpr int RPITCH = 34;  // pitch (doubles) of the per-warp transpose buffer: 16B rows, conflict-free

struct StreamArgs {
  const double* D;        // [n][ld]
  const double* L;        // [ld][R] loadings (W rows then C rows)
  const FitState* st;
  double* partials;       // [grid][momvec_len]
  double* muT;            // [R][n] (column-major n x R) or null
  double* muU;
  int64_t n, ntiles;
  int32_t ld, p2, q2, nwX, nwY, nstage, write_mu;
};

// Upper bound on threads per CTA incl. the reducer warp (register budget: 255 regs at 256
// threads, 168 at 384, 128 at 512).
__host__ __device__ constexpr int max_threads_for(int M, int R) {
  return M >= 4 ? 256 : (M == 2 ? (R <= 3 ? 384 : 256) : (R <= 3 ? 512 : 384));
}

__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void named_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}
__device__ __forceinline__ void named_arrive(int id, int nthreads) {
  asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}
// named barriers: 1,2 = "projection partials of tile t ready" (t even/odd), 3,4 = "mu of tile t ready"
constexpr int BAR_P = 1, BAR_M = 3;

// Warp roles.  Warps 0..nwc-1 (nwc = nwX + nwY) are COMPUTE warps: each owns M chunks of 64
// columns of X (or of Y) and keeps the loadings and the G accumulators of its columns in
// registers.  Warp nwc is the REDUCER/PRODUCER: it sums the warps' projection partials, forms
// mu_T/mu_U and the Gram terms, and keeps the shared-memory ring fed with cp.async.bulk.
// Compute warps are software-pipelined -- A(t) then B(t-1) -- so the reducer's work on tile t
// overlaps phase B of tile t-1 and never sits on the critical path.
template <int R, int M, int T>
__global__ void __launch_bounds__(max_threads_for(M, R), 1) em_stream_kernel(const StreamArgs a) {
  constexpr int V = T * R;                       // projection values per tile
  static_assert(V <= 32, "T*R must fit one warp");
  constexpr int H = (V <= 8) ? 4 : (V <= 16 ? 2 : 1);   // lane groups summing one value
  constexpr int SEG = 32 / H;                    // staged lane-partials each group adds up
  if (a.st->done && !a.write_mu) return;         // the posterior-means pass runs after the end

  extern __shared__ __align__(128) unsigned char smem_raw[];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int nwc = a.nwX + a.nwY;
  const int nbar = (nwc + 1) * 32;               // participants of the named barriers
  const int ld = a.ld;
  const size_t stage_doubles = (size_t)T * ld;

  double* stages = reinterpret_cast<double*>(smem_raw);
  double* red = stages + stage_doubles * a.nstage;            // [nwc][V][RPITCH]
  double* wsum = red + (size_t)nwc * V * RPITCH;               // [2][nwc][V]
  double* mu_s = wsum + (size_t)2 * nwc * V;                   // [2][2][V]  (buf, T/U)
  double* xy_s = mu_s + 4 * V;                                 // [2][V]     (reducer scratch: Xw, Yc)
  double* coef = xy_s + 2 * V;                                 // aT,bT,aU,bU [4][R]
  uint64_t* full = reinterpret_cast<uint64_t*>(coef + 4 * R);  // [nstage]
  uint64_t* empty = full + a.nstage;                           // [nstage]

  if (tid < R) {
    coef[0 * R + tid] = a.st->aT[tid];
    coef[1 * R + tid] = a.st->bT[tid];
    coef[2 * R + tid] = a.st->aU[tid];
    coef[3 * R + tid] = a.st->bU[tid];
  }
  // zero the ring once: rows of a short last tile that are never loaded must stay finite
  for (size_t e = tid; e < stage_doubles * a.nstage; e += blockDim.x) stages[e] = 0.0;
  if (tid == 0) {
    for (int s = 0; s < a.nstage; ++s) { mbar_init(smem_u32(&full[s]), 1); mbar_init(smem_u32(&empty[s]), nwc); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  fence_proxy_async();
  __syncthreads();

  const int64_t my_tiles = (a.ntiles > (int64_t)blockIdx.x)
                               ? (a.ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
  double* out = a.partials + (size_t)blockIdx.x * momvec_len(ld, R);

  if (warp == nwc) {
    // =============================== reducer / producer warp ===============================
    auto issue = [&](int64_t it) {  // lane 0: bulk copy of this CTA's it-th tile
      const int64_t tile = (int64_t)blockIdx.x + it * gridDim.x;
      const int s = (int)(it % a.nstage);
      const int64_t row0 = tile * T;
      const int rows = (int)((a.n - row0 < T) ? (a.n - row0) : T);
      const uint32_t bytes = (uint32_t)((size_t)rows * ld * sizeof(double));
      const uint32_t bar = smem_u32(&full[s]);
      mbar_expect_tx(bar, bytes);
      bulk_g2s(smem_u32(stages + stage_doubles * s), a.D + (size_t)row0 * ld, bytes, bar);
    };
    if (lane == 0)
      for (int64_t it = 0; it < a.nstage && it < my_tiles; ++it) issue(it);

    const int i = lane / R, k = lane - i * R;     // lane (i,k) owns Xw[i,k], Yc[i,k] of the tile
    const bool act = lane < V;
    double cT0 = 0, cT1 = 0, cU0 = 0, cU1 = 0;
    if (act) { cT0 = coef[k]; cT1 = coef[R + k]; cU0 = coef[2 * R + k]; cU1 = coef[3 * R + k]; }
    double gxx[R], gxy[R], gyy[R];
#pragma unroll
    for (int l = 0; l < R; ++l) gxx[l] = gxy[l] = gyy[l] = 0.0;

    for (int64_t it = 0; it < my_tiles; ++it) {
      const int buf = (int)(it & 1);
      const int64_t row0 = ((int64_t)blockIdx.x + it * gridDim.x) * T;
      const int rows = (int)((a.n - row0 < T) ? (a.n - row0) : T);
      named_sync(BAR_P + buf, nbar);              // all compute warps delivered tile `it`
      double xw = 0.0, yc = 0.0;
      if (act && i < rows) {
        const double* ws = wsum + (size_t)buf * nwc * V + lane;
        double x1 = 0.0, y1 = 0.0;
        int ww = 0;
        for (; ww + 1 < a.nwX; ww += 2) { xw += ws[ww * V]; x1 += ws[(ww + 1) * V]; }
        if (ww < a.nwX) xw += ws[ww * V];
        for (ww = a.nwX; ww + 1 < nwc; ww += 2) { yc += ws[ww * V]; y1 += ws[(ww + 1) * V]; }
        if (ww < nwc) yc += ws[ww * V];
        xw += x1; yc += y1;
      }
      if (act) {
        const double mt = cT0 * xw + cT1 * yc;
        const double mu = cU0 * xw + cU1 * yc;
        mu_s[(buf * 2 + 0) * V + lane] = mt;
        mu_s[(buf * 2 + 1) * V + lane] = mu;
        xy_s[lane] = xw;
        xy_s[V + lane] = yc;
        if (a.write_mu && i < rows) {
          a.muT[(size_t)k * a.n + row0 + i] = mt;
          a.muU[(size_t)k * a.n + row0 + i] = mu;
        }
      }
      __syncwarp();
      named_arrive(BAR_M + buf, nbar);            // mu of tile `it` is ready
      if (act) {
#pragma unroll
        for (int l = 0; l < R; ++l) {
          const double xl = xy_s[i * R + l], yl = xy_s[V + i * R + l];
          gxx[l] = fma(xw, xl, gxx[l]);
          gxy[l] = fma(xw, yl, gxy[l]);
          gyy[l] = fma(yc, yl, gyy[l]);
        }
      }
      __syncwarp();
      // producer: refill the stage of tile it-1 as soon as every compute warp finished B(it-1)
      if (lane == 0 && it >= 1 && it - 1 + a.nstage < my_tiles) {
        const int64_t prev = it - 1;
        mbar_wait(smem_u32(&empty[prev % a.nstage]), (uint32_t)((prev / a.nstage) & 1));
        issue(prev + a.nstage);
      }
      __syncwarp();
    }
    // Gram partials of this CTA: entry (which, k, l) = sum over tile rows i of lane (i,k)'s slot l.
    // red is free once every compute warp is past its last transpose; the final __syncthreads
    // pair below orders that.
    __syncthreads();
    if (act) {
#pragma unroll
      for (int l = 0; l < R; ++l) {
        red[(0 * R + l) * 32 + lane] = gxx[l];
        red[(1 * R + l) * 32 + lane] = gxy[l];
        red[(2 * R + l) * 32 + lane] = gyy[l];
      }
    }
    __syncthreads();
  } else {
    // ===================================== compute warps =====================================
    const bool isY = warp >= a.nwX;
    const int wg = isY ? warp - a.nwX : warp;
    const int nwg = isY ? a.nwY : a.nwX;
    const int npairs = (isY ? a.q2 : a.p2) >> 1;
    const int colbase = isY ? a.p2 : 0;
    int col[M];
    bool valid[M];
    double w[M][2][R], acc[M][2][R];
#pragma unroll
    for (int m = 0; m < M; ++m) {
      int u = (m * nwg + wg) * 32 + lane;
      valid[m] = u < npairs;
      col[m] = colbase + 2 * (valid[m] ? u : (npairs - 1));
#pragma unroll
      for (int c = 0; c < 2; ++c)
#pragma unroll
        for (int k = 0; k < R; ++k) {
          w[m][c][k] = valid[m] ? a.L[(size_t)(col[m] + c) * R + k] : 0.0;
          acc[m][c][k] = 0.0;
        }
    }
    double* red_w = red + (size_t)warp * V * RPITCH;
    const int rg = lane / V, rv = lane - rg * V;      // transpose-reduce: lane sums segment rg of value rv
    const bool ract = lane < H * V;
    const double* rsrc = red_w + rv * RPITCH + (ract ? rg : 0) * SEG;

    auto phaseB = [&](int64_t it) {                   // G += x (outer) mu over this thread's columns
      const int s = (int)(it % a.nstage);
      const int buf = (int)(it & 1);
      const double* tile = stages + stage_doubles * s;
      const double* mu_g = mu_s + (buf * 2 + (isY ? 1 : 0)) * V;
      named_sync(BAR_M + buf, nbar);
#pragma unroll
      for (int i = 0; i < T; ++i) {
        double mk[R];
#pragma unroll
        for (int k = 0; k < R; ++k) mk[k] = mu_g[i * R + k];
#pragma unroll
        for (int m = 0; m < M; ++m) {
          const double2 x = *reinterpret_cast<const double2*>(tile + (size_t)i * ld + col[m]);
#pragma unroll
          for (int k = 0; k < R; ++k) {
            acc[m][0][k] = fma(x.x, mk[k], acc[m][0][k]);
            acc[m][1][k] = fma(x.y, mk[k], acc[m][1][k]);
          }
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(smem_u32(&empty[s]));   // this warp is done with the stage
    };

    for (int64_t it = 0; it < my_tiles; ++it) {
      const int s = (int)(it % a.nstage);
      const int buf = (int)(it & 1);
      const double* tile = stages + stage_doubles * s;
      mbar_wait(smem_u32(&full[s]), (uint32_t)((it / a.nstage) & 1));

      // ---- phase A: partial projections over this thread's columns ----------------------
      double part[T][R];
#pragma unroll
      for (int i = 0; i < T; ++i)
#pragma unroll
        for (int k = 0; k < R; ++k) part[i][k] = 0.0;
#pragma unroll
      for (int i = 0; i < T; ++i) {
#pragma unroll
        for (int m = 0; m < M; ++m) {
          const double2 x = *reinterpret_cast<const double2*>(tile + (size_t)i * ld + col[m]);
#pragma unroll
          for (int k = 0; k < R; ++k) {
            part[i][k] = fma(x.x, w[m][0][k], part[i][k]);
            part[i][k] = fma(x.y, w[m][1][k], part[i][k]);
          }
        }
      }
      // transpose-reduce inside the warp through shared memory
      __syncwarp();
#pragma unroll
      for (int i = 0; i < T; ++i)
#pragma unroll
        for (int k = 0; k < R; ++k) red_w[(i * R + k) * RPITCH + lane] = part[i][k];
      __syncwarp();
      double s0 = 0.0, s1 = 0.0;
      if (ract) {
#pragma unroll
        for (int j = 0; j < SEG; j += 2) {
          const double2 v2 = *reinterpret_cast<const double2*>(rsrc + j);
          s0 += v2.x; s1 += v2.y;
        }
      }
      double tot = s0 + s1;
      if (H >= 4) tot += __shfl_down_sync(0xffffffffu, tot, 2 * V);
      if (H >= 2) tot += __shfl_down_sync(0xffffffffu, tot, V);
      if (lane < V) wsum[((size_t)buf * nwc + warp) * V + lane] = tot;
      __syncwarp();
      named_arrive(BAR_P + buf, nbar);

      if (it >= 1) phaseB(it - 1);
    }
    if (my_tiles >= 1) phaseB(my_tiles - 1);

    // ---- epilogue: this CTA's partial G ----------------------------------------------------
#pragma unroll
    for (int m = 0; m < M; ++m)
      if (valid[m]) {
#pragma unroll
        for (int c = 0; c < 2; ++c)
#pragma unroll
          for (int k = 0; k < R; ++k) out[(size_t)(col[m] + c) * R + k] = acc[m][c][k];
      }
    __syncthreads();
    __syncthreads();
  }
  for (int e = tid; e < 3 * R * R; e += blockDim.x) {
    const int which = e / (R * R), kl = e - which * R * R, k = kl / R, l = kl - k * R;
    double sum = 0.0;
    for (int i = 0; i < T; ++i) sum += red[(which * R + l) * 32 + i * R + k];
    out[(size_t)ld * R + e] = sum;
  }
}

// ---------------------------------------------------------------------------------------
// Host side: plan + launch
// ---------------------------------------------------------------------------------------
static size_t stream_smem_bytes(int R, int T, int nstage, int nwarps, int ld) {
  const size_t V = (size_t)T * R;
  // nwarps = compute warps; see the carve-up at the top of the kernel
  size_t doubles = (size_t)T * ld * nstage + (size_t)nwarps * V * RPITCH + 2 * (size_t)nwarps * V +
                   4 * V + 2 * V + 4 * (size_t)R;
  return doubles * sizeof(double) + 2 * (size_t)nstage * sizeof(uint64_t);
}

static int env_int(const char* name, int dflt) {
  const char* v = getenv(name);
  return (v && *v) ? atoi(v) : dflt;
}

typedef void (*stream_kernel_t)(const StreamArgs);
template <int R, int M, int T>
static stream_kernel_t kernel_of() { return em_stream_kernel<R, M, T>; }

// (R, M, T) instantiations.  T*R <= 32.
static stream_kernel_t lookup_kernel(int R, int M, int T) {
#define PPLS_K(r, m, t) if (R == r && M == m && T == t) return kernel_of<r, m, t>();
  PPLS_K(1, 1, 8) PPLS_K(1, 2, 8) PPLS_K(1, 4, 8)
  PPLS_K(2, 1, 8) PPLS_K(2, 2, 8) PPLS_K(2, 4, 8)
  PPLS_K(3, 1, 8) PPLS_K(3, 2, 8) PPLS_K(3, 4, 8) PPLS_K(3, 1, 4) PPLS_K(3, 2, 4) PPLS_K(3, 4, 4)
  PPLS_K(4, 1, 8) PPLS_K(4, 2, 8) PPLS_K(4, 4, 8) PPLS_K(4, 2, 4) PPLS_K(4, 4, 4)
  PPLS_K(5, 1, 4) PPLS_K(5, 2, 4) PPLS_K(5, 4, 4)
  PPLS_K(6, 1, 4) PPLS_K(6, 2, 4)
  PPLS_K(7, 1, 4) PPLS_K(7, 2, 4)
  PPLS_K(8, 1, 4) PPLS_K(8, 2, 4)
#undef PPLS_K
  return nullptr;
}

cudaError_t plan_stream(const DataShape& s, int R, int sm_count, StreamPlan* plan, const char** why) {
  *why = "";
  if (R < 1 || R > MAXR) { *why = "nr_comp outside 1..8"; return cudaErrorInvalidValue; }
  const int chX = (s.p2 / 2 + 31) / 32, chY = (s.q2 / 2 + 31) / 32;
  const size_t smem_max = 227 * 1024;  // usable dynamic shared memory per CTA (1 CTA per SM)
  const int envT = env_int("PPLS_K1_T", 0), envM = env_int("PPLS_K1_M", 0), envS = env_int("PPLS_K1_STAGES", 0);
  // M: columns per thread = 2M.  More warps (small M) hide latency, fewer (large M) amortise the
  // per-warp projection reduce; the launch bound caps the warps.
  int Ms[3], nM = 0;
  if (envM) Ms[nM++] = envM;
  else if (chX + chY <= 4) { Ms[0] = 1; Ms[1] = 2; Ms[2] = 4; nM = 3; }
  else { Ms[0] = 4; Ms[1] = 2; Ms[2] = 1; nM = 3; }
  int Ts[2], nT = 0;
  if (envT) Ts[nT++] = envT; else { Ts[nT++] = 8; Ts[nT++] = 4; }
  int bestM = 0, bestT = 0, bestS = 0, best_score = -1;
  for (int im = 0; im < nM; ++im) {
    const int M = Ms[im];
    const int nw = (chX + M - 1) / M + (chY + M - 1) / M;
    if ((nw + 1) * 32 > max_threads_for(M, R)) continue;
    for (int it = 0; it < nT; ++it) {
      const int T = Ts[it];
      if (T * R > 32 || !lookup_kernel(R, M, T)) continue;
      int ns = 0;
      for (int c = envS ? envS : 8; c >= (envS ? envS : 2); --c)
        if (stream_smem_bytes(R, T, c, nw, s.ld) <= smem_max) { ns = c; break; }
      if (!ns) continue;
      // prefer >= 4 stages (>= 2 tiles in flight), then the earlier M in the list, then larger T
      const int score = (ns >= 4 ? 1000 : ns * 100) + (nM - im) * 10 + (T == 8 ? 1 : 0);
      if (score > best_score) { best_score = score; bestM = M; bestT = T; bestS = ns; }
    }
  }
  if (best_score < 0) {
    *why = "row tile does not fit shared memory / launch bound (use the column-sharded path)";
    return cudaErrorInvalidValue;
  }
  const int M = bestM, T = bestT, nstage = bestS;
  stream_kernel_t k = lookup_kernel(R, M, T);
  const int nwX = (chX + M - 1) / M, nwY = (chY + M - 1) / M;
  const int nwarps = nwX + nwY;
  const int64_t ntiles = (s.n + T - 1) / T;
  plan->R = R; plan->M = M; plan->T = T; plan->nstage = nstage;
  plan->nwX = nwX; plan->nwY = nwY; plan->threads = (nwarps + 1) * 32;
  plan->ctas_per_sm = 1;
  int64_t grid = (int64_t)sm_count;
  if (grid > ntiles) grid = ntiles;
  if (grid < 1) grid = 1;
  plan->grid = (int)grid;
  plan->smem_bytes = stream_smem_bytes(R, T, nstage, nwarps, s.ld);
  plan->ntiles = ntiles;
  cudaError_t e = cudaFuncSetAttribute((const void*)k, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)plan->smem_bytes);
  if (e != cudaSuccess) { *why = "cudaFuncSetAttribute(MaxDynamicSharedMemorySize) failed"; return e; }
  if (env_int("PPLS_K1_VERBOSE", 0))
    fprintf(stderr, "[ppls] K1 plan: R=%d M=%d T=%d stages=%d warps=%d+%d+1 grid=%d smem=%zu\n", R, M, T, nstage, nwX,
            nwY, plan->grid, plan->smem_bytes);
  return cudaSuccess;
}

cudaError_t launch_em_stream(const StreamPlan& plan, const DataShape& s, const double* D,
                             const double* L, const FitState* st, double* partials,
                             double* muT, double* muU, int write_mu, cudaStream_t stream) {
  stream_kernel_t k = lookup_kernel(plan.R, plan.M, plan.T);
  if (!k) return cudaErrorInvalidValue;
  StreamArgs a;
  a.D = D; a.L = L; a.st = st; a.partials = partials; a.muT = muT; a.muU = muU;
  a.n = s.n; a.ntiles = plan.ntiles; a.ld = s.ld; a.p2 = s.p2; a.q2 = s.q2;
  a.nwX = plan.nwX; a.nwY = plan.nwY; a.nstage = plan.nstage; a.write_mu = write_mu;
  k<<<plan.grid, plan.threads, plan.smem_bytes, stream>>>(a);
  return cudaGetLastError();
}

}  // namespace ppls
