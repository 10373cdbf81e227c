// Standard errors of the loadings -- SURVEY 8f row 2: variances.PPLS_simult
// (Package/PPLS/R/EM_W_multi.R:830-860).  The one N-sized piece of work there is
//   t(X) %*% diag(c(Ctt), nrow(X)) %*% X  =  Ctt * X'X                                  (:846)
// (the reference materialises an N x N diagonal matrix for it); X'X does not depend on the
// component, so it is formed ONCE by a register-tiled FP64 Gram kernel (DFMA-bound), next to
// the tall-skinny X'mu (HBM-bound).  Everything after that is w x w (w = p or q): assembly
// of B_exp - SSt_exp per component and its inverse by Gauss-Jordan with partial pivoting.
#include "common.cuh"
#include <math.h>

namespace ppls {

constexpr int VT = 128;        // Gram tile: VT x VT outputs per CTA
constexpr int VK = 32;         // rows per pipeline stage (fewer barriers: +2 %)
constexpr int VTHREADS = 256;  // 8 warps: 2 x 4, each a 64 x 32 block of the tile = 8 x 4 DMMA tiles of 8 x 8
constexpr int VP = VT + 4;     // smem pitch: the 4 k-rows of a DMMA fragment fall into disjoint banks

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc, int src_bytes) {
  const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(d), "l"(gsrc), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
// D (8x8) += A (8x4, row) * B (4x8, col) on the FP64 tensor path.  Lane l: A[l>>2][l&3], B[l&3][l>>2], C[l>>2][2(l&3) + {0,1}].
// Measured on B200 (tools/dmma_probe.cu): 37.0 TFLOP/s at 8 warps/SM; the same loop with DFMA tops out at 33 with ideal
// operand reuse and at 25 inside a shared-memory-fed 8x8 register tile (register bandwidth: 3 operands per FMA).
__device__ __forceinline__ void dmma884(double (&c)[2], double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c[0]), "+d"(c[1]) : "d"(a), "d"(b));
}
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

struct GramArgs {
  const double* D;
  size_t pitch;        // doubles between consecutive rows of a column block
  size_t block_stride; // doubles between consecutive 128-column blocks
  size_t base;         // offset of block 0
  int masked;          // 1: columns >= w2 of the last block belong to the other side -> read as zero
  int w2;              // padded width of the side
  int nb, ntiles;
  int64_t n, rows_per_split;
  double* partial;     // [nsplit][ntiles][VT*VT]
};

// partial[split][tile (bi<=bj)] = sum over the split's rows of side[:, bi-block]' side[:, bj-block]
__global__ void __launch_bounds__(VTHREADS, 1) var_gram_kernel(const GramArgs a) {
  extern __shared__ __align__(16) double sm[];          // [2 stages][2 operands][VK][VP]
  int bi = 0, rem = blockIdx.x;
  while (rem >= a.nb - bi) { rem -= a.nb - bi; ++bi; }
  const int bj = bi + rem;
  const bool diag = (bi == bj);
  const int64_t r0 = (int64_t)blockIdx.y * a.rows_per_split;
  const int64_t r1 = (r0 + a.rows_per_split < a.n) ? r0 + a.rows_per_split : a.n;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, gid = lane >> 2, tig = lane & 3;
  const int wm = (warp >> 2) * 64, wn = (warp & 3) * 32;         // this warp's block of the tile
  const int nst = (r1 > r0) ? (int)((r1 - r0 + VK - 1) / VK) : 0;
  const double* baseA = a.D + a.base + (size_t)bi * a.block_stride;
  const double* baseB = a.D + a.base + (size_t)bj * a.block_stride;

  auto issue = [&](int s, int buf) {
    double* dstA = sm + (size_t)(buf * 2 + 0) * VK * VP;
    double* dstB = sm + (size_t)(buf * 2 + 1) * VK * VP;
#pragma unroll
    for (int c = 0; c < (VK * VT / 2) / VTHREADS; ++c) {
      const int chunk = tid + c * VTHREADS, row = chunk >> 6, col = (chunk & 63) * 2;
      const int64_t grow = r0 + (int64_t)s * VK + row;
      const bool rok = grow < r1;
      const bool okA = rok && (!a.masked || bi * VT + col < a.w2);
      cp_async16(dstA + row * VP + col, okA ? baseA + (size_t)grow * a.pitch + col : a.D, okA ? 16 : 0);
      if (!diag) {
        const bool okB = rok && (!a.masked || bj * VT + col < a.w2);
        cp_async16(dstB + row * VP + col, okB ? baseB + (size_t)grow * a.pitch + col : a.D, okB ? 16 : 0);
      }
    }
    cp_async_commit();
  };

  double acc[8][4][2];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) { acc[i][j][0] = 0.0; acc[i][j][1] = 0.0; }

  if (nst > 0) issue(0, 0);
  for (int s = 0; s < nst; ++s) {
    if (s + 1 < nst) { issue(s + 1, (s + 1) & 1); cp_async_wait<1>(); } else cp_async_wait<0>();
    __syncthreads();
    const double* As = sm + (size_t)((s & 1) * 2 + 0) * VK * VP;
    const double* Bs = diag ? As : As + VK * VP;
#pragma unroll
    for (int k4 = 0; k4 < VK / 4; ++k4) {
      const double* ar = As + (k4 * 4 + tig) * VP + wm + gid;
      const double* br = Bs + (k4 * 4 + tig) * VP + wn + gid;
      double af[8], bf[4];
#pragma unroll
      for (int i = 0; i < 8; ++i) af[i] = ar[i * 8];
#pragma unroll
      for (int j = 0; j < 4; ++j) bf[j] = br[j * 8];
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) dmma884(acc[i][j], af[i], bf[j]);
    }
    __syncthreads();
  }
  double* out = a.partial + ((size_t)blockIdx.y * a.ntiles + blockIdx.x) * VT * VT;
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j)
      *reinterpret_cast<double2*>(out + (wm + i * 8 + gid) * VT + wn + j * 8 + 2 * tig) = make_double2(acc[i][j][0], acc[i][j][1]);
}

// S[j][l] = S[l][j] = sum over splits (fixed order) of the tile partials, for j <= l < w
__global__ void __launch_bounds__(256) var_gram_finalize_kernel(const double* partial, int nsplit, int ntiles, int nb, int w,
                                                                double* S) {
  int bi = 0, rem = blockIdx.x;
  while (rem >= nb - bi) { rem -= nb - bi; ++bi; }
  const int bj = bi + rem;
  for (int e = blockIdx.y * 256 + threadIdx.x; e < VT * VT; e += gridDim.y * 256) {
    const int jj = e / VT, ll = e % VT, j = bi * VT + jj, l = bj * VT + ll;
    if (j >= w || l >= w || j > l) continue;
    double s = 0.0;
    for (int sp = 0; sp < nsplit; ++sp) s += partial[((size_t)sp * ntiles + blockIdx.x) * VT * VT + e];
    S[(size_t)j * w + l] = s;
    S[(size_t)l * w + j] = s;
  }
}

// ---- G = side' mu, cp = colSums(mu^2) ---------------------------------------------------------------
// CTA = one 128-column block x one row split; thread t: column (t & 127), rows of parity (t >> 7).
template <int R>
__global__ void __launch_bounds__(256) var_xtmu_kernel(const GramArgs a, const double* __restrict__ mu, int w2side,
                                                       double* partialG /*[gsplit][nb*128][R]*/, double* partialC /*[gsplit][R]*/) {
  __shared__ double red[128 * R];
  const int b = blockIdx.x, col = threadIdx.x & 127, h = threadIdx.x >> 7;
  const int64_t r0 = (int64_t)blockIdx.y * a.rows_per_split;
  const int64_t r1 = (r0 + a.rows_per_split < a.n) ? r0 + a.rows_per_split : a.n;
  const bool ok = !a.masked || b * VT + col < a.w2;
  const double* src = a.D + a.base + (size_t)b * a.block_stride + col;
  double acc[R], sq[R];
#pragma unroll
  for (int k = 0; k < R; ++k) { acc[k] = 0.0; sq[k] = 0.0; }
  if (ok) {
    int64_t i = r0 + h;
    for (; i + 6 < r1; i += 8) {                           // 4 rows in flight per thread
      double x[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) x[u] = src[(size_t)(i + 2 * u) * a.pitch];
#pragma unroll
      for (int u = 0; u < 4; ++u)
#pragma unroll
        for (int k = 0; k < R; ++k) acc[k] = fma(x[u], __ldg(mu + (size_t)(i + 2 * u) * R + k), acc[k]);
    }
    for (; i < r1; i += 2) {
      const double x = src[(size_t)i * a.pitch];
#pragma unroll
      for (int k = 0; k < R; ++k) acc[k] = fma(x, __ldg(mu + (size_t)i * R + k), acc[k]);
    }
  }
  if (h == 1) {
#pragma unroll
    for (int k = 0; k < R; ++k) red[col * R + k] = acc[k];
  }
  __syncthreads();
  if (h == 0) {
    double* out = partialG + ((size_t)blockIdx.y * gridDim.x * VT + (size_t)b * VT + col) * R;
#pragma unroll
    for (int k = 0; k < R; ++k) out[k] = acc[k] + red[col * R + k];
  }
  (void)w2side;
  // column block 0 also sums mu^2 over the split's rows
  if (b == 0) {
    __syncthreads();
    for (int64_t i = r0 + threadIdx.x; i < r1; i += 256)
#pragma unroll
      for (int k = 0; k < R; ++k) { const double m = __ldg(mu + (size_t)i * R + k); sq[k] = fma(m, m, sq[k]); }
    __shared__ double wsum[8][R];
#pragma unroll
    for (int k = 0; k < R; ++k) {
      double v = sq[k];
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5][k] = v;
    }
    __syncthreads();
    if (threadIdx.x < R) {
      double s = 0.0;
      for (int wv = 0; wv < 8; ++wv) s += wsum[wv][threadIdx.x];
      partialC[(size_t)blockIdx.y * R + threadIdx.x] = s;
    }
  }
}

__global__ void __launch_bounds__(256) var_xtmu_finalize_kernel(const double* partialG, const double* partialC, int gsplit,
                                                                int nbw /*nb*128*/, int w, int R, double* G, double* cp) {
  const int e = blockIdx.x * 256 + threadIdx.x;
  if (e < w * R) {
    double s = 0.0;
    for (int sp = 0; sp < gsplit; ++sp) s += partialG[(size_t)sp * nbw * R + e];
    G[e] = s;
  }
  if (blockIdx.x == 0 && threadIdx.x < R) {
    double s = 0.0;
    for (int sp = 0; sp < gsplit; ++sp) s += partialC[(size_t)sp * R + threadIdx.x];
    cp[threadIdx.x] = s;
  }
}

// ---- per component: [A | I] with A = B_exp - SSt_exp  (EM_W_multi.R:844-849,:855) -------------------------
// Ctt here is N * Ctt[i,i] (:840), Vt = Ctt - crossprod(mu) (:842), Cxt = G[:, i] (:843).
struct AsmArgs {
  const double* S; const double* G; const double* W;
  double* aug; double* SSt_exp; double* SSt_star;
  int w, R, comp;
  double Ctt, c1, c2, sig4, N, Bstar;
};
__global__ void __launch_bounds__(256) var_assemble_kernel(const AsmArgs a) {
  const size_t e = (size_t)blockIdx.x * 256 + threadIdx.x;
  if (e >= (size_t)a.w * a.w) return;
  const int j = (int)(e / a.w), l = (int)(e % a.w);
  const double cj = a.G[(size_t)j * a.R + a.comp], cl = a.G[(size_t)l * a.R + a.comp];
  const double wj = a.W[(size_t)j * a.R + a.comp], wl = a.W[(size_t)l * a.R + a.comp];
  // t(X) diag(Ctt) X - Cxt (Ctt+2Vt) t(W) - W (Ctt+2Vt) t(Cxt) + W (Ctt^2 + 4 cp Vt + 2 Vt^2) t(W)   (:846-847)
  double v = a.Ctt * a.S[e] - (cj * a.c1) * wl;
  v = v - (wj * a.c1) * cl;
  v = v + (wj * a.c2) * wl;
  v = v / a.sig4 / a.N;                                                           // :848
  const double star = (cj - wj * a.Ctt) * (cl - wl * a.Ctt) / a.sig4;             // :850-851
  a.aug[(size_t)j * 2 * a.w + l] = (j == l ? a.Bstar : 0.0) - v;                  // :855  B_exp - SSt_exp
  a.aug[(size_t)j * 2 * a.w + a.w + l] = (j == l) ? 1.0 : 0.0;
  if (a.SSt_exp) a.SSt_exp[(size_t)j + (size_t)l * a.w] = v;
  if (a.SSt_star) a.SSt_star[(size_t)j + (size_t)l * a.w] = star;
}

// ---- Gauss-Jordan on [A | I] (w x 2w row-major), partial pivoting, one column per step -------------------
// Step k, kernel 1 (one block): pivot row r = argmax_{i>=k} |A[i][k]| (smallest index on ties); saves the
// scaled pivot row, the old row k and the old column k, so that kernel 2 can update all rows in place
// without reading anything another block is writing.
// rowbuf: [prow 2w | krow 2w | fcol w]; info: [0] = 1 + first column with a zero pivot, [1] = r.
__global__ void __launch_bounds__(1024) gj_pivot_kernel(const double* aug, int w, int k, double* rowbuf, int* info) {
  __shared__ double bv[32];
  __shared__ int bi_[32];
  __shared__ int rsel;
  double best = -1.0; int besti = w;
  for (int i = k + threadIdx.x; i < w; i += 1024) {
    const double v = fabs(aug[(size_t)i * 2 * w + k]);
    if (v > best) { best = v; besti = i; }
  }
  for (int o = 16; o > 0; o >>= 1) {
    const double ov = __shfl_xor_sync(0xffffffffu, best, o);
    const int oi = __shfl_xor_sync(0xffffffffu, besti, o);
    if (ov > best || (ov == best && oi < besti)) { best = ov; besti = oi; }
  }
  if ((threadIdx.x & 31) == 0) { bv[threadIdx.x >> 5] = best; bi_[threadIdx.x >> 5] = besti; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int q = 1; q < 32; ++q)
      if (bv[q] > best || (bv[q] == best && bi_[q] < besti)) { best = bv[q]; besti = bi_[q]; }
    if (!(best > 0.0)) { if (info[0] == 0) info[0] = k + 1; besti = k; }
    rsel = besti; info[1] = besti;
  }
  __syncthreads();
  const int r = rsel;
  const double piv = aug[(size_t)r * 2 * w + k];
  double* prow = rowbuf; double* krow = rowbuf + 2 * w; double* fcol = rowbuf + 4 * w;
  for (int j = threadIdx.x; j < 2 * w; j += 1024) {
    prow[j] = aug[(size_t)r * 2 * w + j] / piv;
    krow[j] = aug[(size_t)k * 2 * w + j];
  }
  for (int i = threadIdx.x; i < w; i += 1024) fcol[i] = aug[(size_t)i * 2 * w + k];
}
__global__ void __launch_bounds__(256) gj_elim_kernel(double* aug, int w, int k, const double* rowbuf, const int* info) {
  const int r = info[1];
  const int i = blockIdx.y;
  const double* prow = rowbuf; const double* krow = rowbuf + 2 * w; const double* fcol = rowbuf + 4 * w;
  double* row = aug + (size_t)i * 2 * w;
  // columns j < k of the left half are unit vectors already and the pivot row is zero there: nothing changes
  for (int j = k + blockIdx.x * 256 + threadIdx.x; j < 2 * w; j += gridDim.x * 256) {
    if (i == k) row[j] = prow[j];
    else if (i == r) row[j] = krow[j] - krow[k] * prow[j];        // the old row k lands in row r, eliminated
    else row[j] = row[j] - fcol[i] * prow[j];
  }
}
// var = -inv(A) (:855), se = sqrt(diag(var)) (:856)
__global__ void __launch_bounds__(256) var_extract_kernel(const double* aug, int w, double* var_out, double* se_out) {
  const size_t e = (size_t)blockIdx.x * 256 + threadIdx.x;
  if (e >= (size_t)w * w) return;
  const int j = (int)(e / w), l = (int)(e % w);
  const double v = -aug[(size_t)j * 2 * w + w + l];
  if (var_out) var_out[(size_t)j + (size_t)l * w] = v;
  if (j == l) se_out[j] = sqrt(v);
}

// ---- host side ---------------------------------------------------------------------------------------
void plan_var(const DataShape& s, int which, int sm_count, VarPlan* p) {
  p->w = which ? s.q : s.p;
  p->c0 = which ? s.p2 : 0;
  const int w2 = which ? s.q2 : s.p2;
  p->nb = (w2 + VT - 1) / VT;
  p->ntiles = p->nb * (p->nb + 1) / 2;
  // Gram: one CTA per SM (232 registers); the row split is chosen so that tiles x splits fills whole waves
  // of sm_count CTAs (>= 4 waves when the rows allow), at least 8 stages of rows per CTA
  const int64_t max_ns = (s.n + 8 * VK - 1) / (8 * VK);
  int64_t lo = (4 * (int64_t)sm_count + p->ntiles - 1) / p->ntiles, ns = 1;
  if (lo < 1) lo = 1;
  double best = -1.0;
  for (int64_t c = lo; c < lo + 64; ++c) {
    const int64_t cc = c > max_ns ? max_ns : c;
    if (cc < 1) continue;
    const int64_t ctas = cc * p->ntiles, waves = (ctas + sm_count - 1) / sm_count;
    const double eff = (double)ctas / (double)(waves * sm_count);
    if (eff > best + 1e-9) { best = eff; ns = cc; }
    if (cc == max_ns) break;
  }
  int64_t rps = (s.n + ns - 1) / ns;
  rps = (rps + VK - 1) / VK * VK;
  p->nsplit = (int)((s.n + rps - 1) / rps);
  p->rows_per_split = rps;
  // X'mu: about 8 CTAs of 256 threads per SM
  int64_t gs = (8 * (int64_t)sm_count + p->nb - 1) / p->nb;
  const int64_t max_gs = (s.n + 63) / 64;
  if (gs > max_gs) gs = max_gs;
  if (gs < 1) gs = 1;
  int64_t grs = (s.n + gs - 1) / gs;
  grs = (grs + 1) / 2 * 2;
  p->gsplit = (int)((s.n + grs - 1) / grs);
  p->grows_per_split = grs;
}

size_t var_workspace_doubles(const VarPlan& p, int R) {
  const size_t gram = (size_t)p.nsplit * p.ntiles * VT * VT;
  const size_t xtmu = (size_t)p.gsplit * p.nb * VT * R + (size_t)p.gsplit * R;
  return gram > xtmu ? gram : xtmu;
}

static GramArgs gram_args(const VarPlan& p, const DataShape& s, const double* D, int which) {
  GramArgs a{};
  a.D = D; a.n = s.n; a.nb = p.nb; a.ntiles = p.ntiles;
  a.w2 = which ? s.q2 : s.p2;
  if (s.panel) {
    a.pitch = PANEL_PITCH; a.block_stride = (size_t)s.n * PANEL_PITCH;
    a.base = which ? (size_t)s.npX * s.n * PANEL_PITCH : 0; a.masked = 0;
  } else {
    a.pitch = (size_t)s.ld; a.block_stride = VT; a.base = (size_t)p.c0; a.masked = 1;
  }
  return a;
}

cudaError_t launch_var_gram(const VarPlan& p, const DataShape& s, const double* D, double* partial, double* S, cudaStream_t stream) {
  GramArgs a = gram_args(p, s, D, p.c0 != 0);
  a.rows_per_split = p.rows_per_split; a.partial = partial;
  const size_t smem = (size_t)2 * 2 * VK * VP * sizeof(double);
  cudaError_t e = cudaFuncSetAttribute((const void*)var_gram_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  var_gram_kernel<<<dim3(p.ntiles, p.nsplit), VTHREADS, smem, stream>>>(a);
  var_gram_finalize_kernel<<<dim3(p.ntiles, 16), 256, 0, stream>>>(partial, p.nsplit, p.ntiles, p.nb, p.w, S);
  return cudaGetLastError();
}

template <int R>
static void launch_xtmu_R(const GramArgs& a, const VarPlan& p, const double* mu, double* partial, cudaStream_t stream) {
  double* pc = partial + (size_t)p.gsplit * p.nb * VT * R;
  var_xtmu_kernel<R><<<dim3(p.nb, p.gsplit), 256, 0, stream>>>(a, mu, a.w2, partial, pc);
}
cudaError_t launch_var_xtmu(const VarPlan& p, const DataShape& s, const double* D, const double* mu, int R, double* partial,
                            double* G, double* cp, cudaStream_t stream) {
  GramArgs a = gram_args(p, s, D, p.c0 != 0);
  a.rows_per_split = p.grows_per_split;
  switch (R) {
    case 1: launch_xtmu_R<1>(a, p, mu, partial, stream); break;
    case 2: launch_xtmu_R<2>(a, p, mu, partial, stream); break;
    case 3: launch_xtmu_R<3>(a, p, mu, partial, stream); break;
    case 4: launch_xtmu_R<4>(a, p, mu, partial, stream); break;
    case 5: launch_xtmu_R<5>(a, p, mu, partial, stream); break;
    case 6: launch_xtmu_R<6>(a, p, mu, partial, stream); break;
    case 7: launch_xtmu_R<7>(a, p, mu, partial, stream); break;
    case 8: launch_xtmu_R<8>(a, p, mu, partial, stream); break;
    default: return cudaErrorInvalidValue;
  }
  const double* pc = partial + (size_t)p.gsplit * p.nb * VT * R;
  var_xtmu_finalize_kernel<<<(p.w * R + 255) / 256, 256, 0, stream>>>(partial, pc, p.gsplit, p.nb * VT, p.w, R, G, cp);
  return cudaGetLastError();
}

cudaError_t launch_var_assemble(int w, int R, int comp, const double* S, const double* G, const double* Wpol, double CttN,
                                double cp, double sigE, double Ntot, double* aug, double* SSt_exp, double* SSt_star,
                                cudaStream_t stream) {
  AsmArgs a{};
  a.S = S; a.G = G; a.W = Wpol; a.aug = aug; a.SSt_exp = SSt_exp; a.SSt_star = SSt_star;
  a.w = w; a.R = R; a.comp = comp;
  const double Vt = CttN - cp;                                   // :842
  a.Ctt = CttN; a.c1 = CttN + 2 * Vt; a.c2 = CttN * CttN + 4 * cp * Vt + 2 * Vt * Vt;
  a.sig4 = (sigE * sigE) * (sigE * sigE); a.N = Ntot;
  a.Bstar = CttN / (sigE * sigE) * 1.0 / Ntot;                   // :844
  const size_t ne = (size_t)w * w;
  var_assemble_kernel<<<(unsigned)((ne + 255) / 256), 256, 0, stream>>>(a);
  return cudaGetLastError();
}

cudaError_t launch_var_invert(int w, double* aug, double* rowbuf, int* info, double* var_out, double* se_out, cudaStream_t stream) {
  cudaError_t e = cudaMemsetAsync(info, 0, 2 * sizeof(int), stream);
  if (e != cudaSuccess) return e;
  int gx = (2 * w + 255) / 256;
  if (gx > 8) gx = 8;
  for (int k = 0; k < w; ++k) {
    gj_pivot_kernel<<<1, 1024, 0, stream>>>(aug, w, k, rowbuf, info);
    gj_elim_kernel<<<dim3(gx, w), 256, 0, stream>>>(aug, w, k, rowbuf, info);
  }
  const size_t ne = (size_t)w * w;
  var_extract_kernel<<<(unsigned)((ne + 255) / 256), 256, 0, stream>>>(aug, w, var_out, se_out);
  return cudaGetLastError();
}

}  // namespace ppls
