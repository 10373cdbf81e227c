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

constexpr int RPITCH = 33;  // pitch (doubles) of the per-warp transpose buffer: conflict-free

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

// Upper bound on threads per CTA (register budget: 255 regs at 256 threads, 128 at 512).
__host__ __device__ constexpr int max_threads_for(int M, int R) {
  return (M >= 4 || (M >= 2 && R >= 4) || R >= 7) ? 256 : 512;
}

template <int R, int M, int T>
__global__ void __launch_bounds__(max_threads_for(M, R), 1) em_stream_kernel(const StreamArgs a) {
  constexpr int V = T * R;  // partial projections per thread and tile
  static_assert(V <= 32, "T*R must fit one warp");
  if (a.st->done && !a.write_mu) return;   // the posterior-means pass runs after the end

  extern __shared__ __align__(128) unsigned char smem_raw[];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int nwarps = a.nwX + a.nwY;
  const int ld = a.ld;
  const size_t stage_doubles = (size_t)T * ld;

  double* stages = reinterpret_cast<double*>(smem_raw);
  double* red = stages + stage_doubles * a.nstage;          // [nwarps][V][RPITCH]
  double* wsum = red + (size_t)nwarps * V * RPITCH;          // [nwarps][V]
  double* mu_s = wsum + (size_t)nwarps * V;                  // [2][V]
  double* coef = mu_s + 2 * V;                               // aT,bT,aU,bU [4][R]
  uint64_t* bars = reinterpret_cast<uint64_t*>(coef + 4 * R);  // [nstage]

  // ---- column ownership -------------------------------------------------------------
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
  if (tid < R) {
    coef[0 * R + tid] = a.st->aT[tid];
    coef[1 * R + tid] = a.st->bT[tid];
    coef[2 * R + tid] = a.st->aU[tid];
    coef[3 * R + tid] = a.st->bU[tid];
  }
  // zero the ring once: rows of a short last tile that are never loaded must stay finite
  for (size_t e = tid; e < stage_doubles * a.nstage; e += blockDim.x) stages[e] = 0.0;
  if (tid == 0) {
    for (int s = 0; s < a.nstage; ++s) mbar_init(smem_u32(&bars[s]), 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  fence_proxy_async();
  __syncthreads();

  const int64_t my_tiles = (a.ntiles > (int64_t)blockIdx.x)
                               ? (a.ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
  auto issue = [&](int64_t it) {  // one elected thread: bulk copy of this CTA's it-th tile
    const int64_t tile = (int64_t)blockIdx.x + it * gridDim.x;
    const int s = (int)(it % a.nstage);
    const int64_t row0 = tile * T;
    const int rows = (int)((a.n - row0 < T) ? (a.n - row0) : T);
    const uint32_t bytes = (uint32_t)((size_t)rows * ld * sizeof(double));
    const uint32_t bar = smem_u32(&bars[s]);
    mbar_expect_tx(bar, bytes);
    bulk_g2s(smem_u32(stages + stage_doubles * s), a.D + (size_t)row0 * ld, bytes, bar);
  };
  if (tid == 0)
    for (int64_t it = 0; it < a.nstage && it < my_tiles; ++it) issue(it);

  // Gram accumulators of warp 0: lane (i,k) keeps sums over this CTA's tiles
  double gxx[R], gxy[R], gyy[R];
#pragma unroll
  for (int l = 0; l < R; ++l) gxx[l] = gxy[l] = gyy[l] = 0.0;

  double* red_w = red + (size_t)warp * V * RPITCH;

  for (int64_t it = 0; it < my_tiles; ++it) {
    const int s = (int)(it % a.nstage);
    const uint32_t parity = (uint32_t)((it / a.nstage) & 1);
    const int64_t row0 = ((int64_t)blockIdx.x + it * gridDim.x) * T;
    const int rows = (int)((a.n - row0 < T) ? (a.n - row0) : T);
    const double* tile = stages + stage_doubles * s;
    mbar_wait(smem_u32(&bars[s]), parity);

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
    // transpose-reduce inside the warp: lane v (< V) ends with the warp total of value v
#pragma unroll
    for (int i = 0; i < T; ++i)
#pragma unroll
      for (int k = 0; k < R; ++k) red_w[(i * R + k) * RPITCH + lane] = part[i][k];
    __syncwarp();
    if (lane < V) {
      const double* rv = red_w + lane * RPITCH;
      double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
#pragma unroll
      for (int j = 0; j < 32; j += 4) {
        s0 += rv[j]; s1 += rv[j + 1]; s2 += rv[j + 2]; s3 += rv[j + 3];
      }
      wsum[warp * V + lane] = (s0 + s1) + (s2 + s3);
    }
    __syncthreads();  // S1: warp totals visible; every warp has left phase B of tile it-1

    if (tid == 0 && it >= 1 && it - 1 + a.nstage < my_tiles) issue(it - 1 + a.nstage);

    if (warp == 0) {
      const int i = lane / R, k = lane - i * R;
      double xw = 0.0, yc = 0.0;
      if (lane < V && i < rows) {
        for (int ww = 0; ww < a.nwX; ++ww) xw += wsum[ww * V + lane];
        for (int ww = a.nwX; ww < nwarps; ++ww) yc += wsum[ww * V + lane];
      }
      if (lane < V) {
        const double mt = coef[0 * R + k] * xw + coef[1 * R + k] * yc;
        const double mu = coef[2 * R + k] * xw + coef[3 * R + k] * yc;
        mu_s[lane] = mt;
        mu_s[V + lane] = mu;
        if (a.write_mu && i < rows) {
          a.muT[(size_t)k * a.n + row0 + i] = mt;
          a.muU[(size_t)k * a.n + row0 + i] = mu;
        }
      }
#pragma unroll
      for (int l = 0; l < R; ++l) {
        const int src = (lane < V) ? i * R + l : lane;
        const double xl = __shfl_sync(0xffffffffu, xw, src);
        const double yl = __shfl_sync(0xffffffffu, yc, src);
        gxx[l] = fma(xw, xl, gxx[l]);
        gxy[l] = fma(xw, yl, gxy[l]);
        gyy[l] = fma(yc, yl, gyy[l]);
      }
    }
    __syncthreads();  // S2: mu visible

    // ---- phase B: G += x (outer) mu over this thread's columns --------------------------
    const double* mu_g = mu_s + (isY ? V : 0);
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
  }

  // ---- epilogue: this CTA's partial moment vector ----------------------------------------
  double* out = a.partials + (size_t)blockIdx.x * momvec_len(ld, R);
#pragma unroll
  for (int m = 0; m < M; ++m)
    if (valid[m]) {
#pragma unroll
      for (int c = 0; c < 2; ++c)
#pragma unroll
        for (int k = 0; k < R; ++k) out[(size_t)(col[m] + c) * R + k] = acc[m][c][k];
    }
  __syncthreads();  // ring and red buffers are dead now; reuse `red` for the Gram reduction
  if (warp == 0 && lane < V) {
#pragma unroll
    for (int l = 0; l < R; ++l) {
      red[(0 * R + l) * 32 + lane] = gxx[l];
      red[(1 * R + l) * 32 + lane] = gxy[l];
      red[(2 * R + l) * 32 + lane] = gyy[l];
    }
  }
  __syncthreads();
  // entry (which, k, l) = sum over tile rows i of lane (i,k)'s slot l
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
  size_t doubles = (size_t)T * ld * nstage + (size_t)nwarps * V * RPITCH + (size_t)nwarps * V +
                   2 * V + 4 * (size_t)R;
  return doubles * sizeof(double) + (size_t)nstage * sizeof(uint64_t);
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
  int T = env_int("PPLS_K1_T", (R <= 4) ? 8 : 4);
  int M = env_int("PPLS_K1_M", 0);
  if (M == 0) {
    M = (chX >= 8 && R <= 5) ? 4 : (chX >= 3 ? 2 : 1);
    // widen M until the warps fit the launch bound
    for (;;) {
      int nw = (chX + M - 1) / M + (chY + M - 1) / M;
      if (nw * 32 <= max_threads_for(M, R)) break;
      if (M == 1) M = 2; else if (M == 2 && R <= 5) M = 4; else break;
    }
  }
  stream_kernel_t k = lookup_kernel(R, M, T);
  if (!k) { *why = "no streaming-kernel instantiation for (R,M,T)"; return cudaErrorInvalidValue; }
  const int nwX = (chX + M - 1) / M, nwY = (chY + M - 1) / M;
  const int nwarps = nwX + nwY;
  if (nwarps * 32 > max_threads_for(M, R)) {
    *why = "row too wide for the row-streaming kernel (use the column-sharded path)";
    return cudaErrorInvalidValue;
  }
  const size_t smem_sm = 227 * 1024;  // usable per SM; each CTA also reserves 1 KB
  int ctas = env_int("PPLS_K1_CTAS", 0), nstage = env_int("PPLS_K1_STAGES", 0);
  const int64_t ntiles = (s.n + T - 1) / T;
  auto fits = [&](int c, int ns) {
    return (stream_smem_bytes(R, T, ns, nwarps, s.ld) + 1024) * (size_t)c <= smem_sm + 1024 &&
           stream_smem_bytes(R, T, ns, nwarps, s.ld) <= smem_sm;
  };
  if (ctas == 0 || nstage == 0) {
    int best_c = 0, best_ns = 0;
    for (int c = (ctas ? ctas : 2); c >= (ctas ? ctas : 1) && !best_c; --c) {
      int lo = nstage ? nstage : 2, hi = nstage ? nstage : (c == 2 ? 4 : 6);
      for (int ns = hi; ns >= lo; --ns)
        if (fits(c, ns)) { best_c = c; best_ns = ns; break; }
    }
    if (!best_c) { *why = "row tile does not fit shared memory (use the column-sharded path)"; return cudaErrorInvalidValue; }
    ctas = best_c; nstage = best_ns;
  } else if (!fits(ctas, nstage)) {
    *why = "requested PPLS_K1_CTAS/PPLS_K1_STAGES do not fit shared memory"; return cudaErrorInvalidValue;
  }
  plan->R = R; plan->M = M; plan->T = T; plan->nstage = nstage;
  plan->nwX = nwX; plan->nwY = nwY; plan->threads = nwarps * 32;
  plan->ctas_per_sm = ctas;
  int64_t grid = (int64_t)sm_count * ctas;
  if (grid > ntiles) grid = ntiles;
  if (grid < 1) grid = 1;
  plan->grid = (int)grid;
  plan->smem_bytes = stream_smem_bytes(R, T, nstage, nwarps, s.ld);
  plan->ntiles = ntiles;
  cudaError_t e = cudaFuncSetAttribute((const void*)k, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)plan->smem_bytes);
  if (e != cudaSuccess) { *why = "cudaFuncSetAttribute(MaxDynamicSharedMemorySize) failed"; return e; }
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
