// Wide-row path ("K4"): two data passes per EM iteration for shapes whose row (p+q columns) does not
// fit the shared-memory tile of the fused streaming kernel (wide omics, p >> N).
//   pass 1  wide_kernel<MODE_PROJECT>   partial projections  Xw = X W, Yc = Y C      EM_W_multi.R:689-690
//           wide_mu_kernel              sum over column slabs, mu_T / mu_U (:691-694), Gram terms
//   pass 2  wide_kernel<MODE_ACCUM>     G = [X' mu_T ; Y' mu_U]                      :732-733
//           wide_finalize_kernel        fixed-order sums -> the same moment vector the fused kernel makes
// A CTA owns one column SLAB (<= 512 columns of X or of Y) and a contiguous range of rows; it streams
// its slab in tiles of 16 row segments (cp.async.bulk per segment -> shared-memory ring).  Warps split
// the tile's rows, lanes stride over the slab's column pairs and keep the slab's loadings (pass 1) or G
// accumulators (pass 2) in registers.  The same layout serves a column-sharded multi-GPU run: a rank's
// local matrix is its column block, and the N x 2r projection sums are what ranks exchange.
#include "em_stream_kernel.cuh"
#include <cstdio>
#include <cstdlib>

namespace ppls {

constexpr int WIDE_TR = 16;          // rows per tile
constexpr int WIDE_WARPS = 8;        // compute warps (each handles rows w and w+8 of a tile)
constexpr int MODE_PROJECT = 1, MODE_ACCUM = 2;

struct WideArgs {
  const double* D; const double* L; const FitState* st;
  double* ppart;            // [nslab][n][R]   (pass 1 out)
  const double* mu;         // [2][n][R]       (pass 2 in: mu_T then mu_U)
  double* gpart;            // [nrg][ld][R]    (pass 2 out)
  int64_t n, npad;          // npad: n rounded up to the tile height (stride of the mu arrays)
  int32_t ld, p2, q2, scX, scY, nslabX, nslabY, nrg, rows_per_rg, nstage;
  int32_t panel;            // panel-major storage: slab = one panel, a tile is one contiguous block
};

// Transposing butterfly: on entry every lane holds NV values; on exit value index
// ((lane >> log2(32/NVP)) & (NVP-1)) of the warp-wide sums is in v[0] (NVP = NV rounded up to a power of 2).
template <int NVP>
__device__ __forceinline__ void warp_reduce_scatter(double* v, int lane) {
  int cnt = NVP;
#pragma unroll
  for (int off = 16; off >= 1; off >>= 1) {
    if (cnt > 1) {
      const int half = cnt / 2;
      const bool hi = (lane & off) != 0;
#pragma unroll
      for (int e = 0; e < NVP / 2; ++e) {
        if (e < half) {
          const double keep = hi ? v[e + half] : v[e];
          const double send = hi ? v[e] : v[e + half];
          v[e] = keep + __shfl_xor_sync(0xffffffffu, send, off);
        }
      }
      cnt = half;
    } else {
      v[0] += __shfl_xor_sync(0xffffffffu, v[0], off);
    }
  }
}
__host__ __device__ constexpr int pow2_ceil(int x) { return x <= 1 ? 1 : x <= 2 ? 2 : x <= 4 ? 4 : x <= 8 ? 8 : 16; }
__host__ __device__ constexpr int ilog2(int x) { return x <= 1 ? 0 : 1 + ilog2(x / 2); }

// Register cap: 9 warps per CTA land 5 + 4 on the SM's four schedulers when two CTAs are resident, and a scheduler's
// 16384 registers hold 5 warps only up to 102 per thread -- at 104+ the second CTA does not fit (measured: pass time
// 155 -> 219 us).  Narrow slabs stay below the cap anyway.
template <int R, int NCH, int MODE>
__global__ void __maxnreg__(NCH * R <= 10 ? 96 : 255) wide_kernel(const WideArgs a) {
  if (a.st->done) return;
  extern __shared__ __align__(128) unsigned char smem[];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  // Persistent: the grid is the number of CTAs resident at once; CTA b walks the work items b, b + grid, ...
  // (item = (column slab, row group)) through ONE continuous ring -- the copies of the next item are in flight while
  // the compute warps finish the current one, and the ring / barriers are set up once per launch.
  const int nslab = a.nslabX + a.nslabY;
  const int nitem = nslab * a.nrg;
  const uint32_t pitch = a.panel ? PANEL_PITCH * 8u : (uint32_t)NCH * 64u * 8u;   // smem row pitch (bytes): a stored panel row / the slab width
  const uint32_t mubytes = (MODE == MODE_ACCUM) ? WIDE_TR * R * 8u : 0u;
  const uint32_t stage_bytes = WIDE_TR * pitch + mubytes;
  const uint32_t sbase = smem_u32(smem);
  const uint32_t gs_off = stage_bytes * a.nstage;                             // ACCUM: warp-sum buffer [NCH*64*2... cols][R]
  const uint32_t bars = sbase + gs_off + ((MODE == MODE_ACCUM) ? (uint32_t)NCH * 64u * R * 8u : 0u);   // full[ns], empty[ns]
  // zero the ring once: lanes beyond a slab's width read (and multiply by zero loadings) whatever is there
  for (uint32_t e = tid; e < stage_bytes * a.nstage / 8u; e += blockDim.x) reinterpret_cast<double*>(smem)[e] = 0.0;
  if (tid == 0) {
    for (int s = 0; s < a.nstage; ++s) { mbar_init(bars + 8u * s, 1); mbar_init(bars + 8u * (a.nstage + s), WIDE_WARPS); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  fence_proxy_async();
  __syncthreads();

  struct Item { int slab, width, cbeg, ntile; bool isY; int64_t rbeg, rend; int rg; };
  auto item_of = [&](int it) {
    Item q;
    q.slab = it % nslab; q.rg = it / nslab;
    q.isY = q.slab >= a.nslabX;
    const int sc = q.isY ? a.scY : a.scX;
    q.cbeg = q.isY ? a.p2 + (q.slab - a.nslabX) * a.scY : q.slab * a.scX;     // first global column
    const int cend_type = q.isY ? a.p2 + a.q2 : a.p2;
    q.width = (cend_type - q.cbeg < sc) ? (cend_type - q.cbeg) : sc;          // even, >= 2
    q.rbeg = (int64_t)q.rg * a.rows_per_rg;
    q.rend = (q.rbeg + a.rows_per_rg < a.n) ? q.rbeg + a.rows_per_rg : a.n;
    q.ntile = (int)((q.rend - q.rbeg + WIDE_TR - 1) / WIDE_TR);
    return q;
  };

  if (warp == WIDE_WARPS) {
    // ------------------------------------ producer ------------------------------------------
    if (lane == 0) {
      uint32_t s = 0, par = 1;
      int64_t gt = 0;                                                          // tiles issued so far (all items)
      for (int it = blockIdx.x; it < nitem; it += gridDim.x) {
        const Item q = item_of(it);
        const uint32_t segbytes = a.panel ? PANEL_PITCH * 8u : (uint32_t)q.width * 8u;   // a panel is stored whole (zero-padded)
        for (int t = 0; t < q.ntile; ++t, ++gt) {
          if (gt >= a.nstage) mbar_wait(bars + 8u * (a.nstage + s), par);
          const int64_t row0 = q.rbeg + (int64_t)(MODE == MODE_ACCUM ? q.ntile - 1 - t : t) * WIDE_TR;   // pass 2 walks back
          const int rows = (int)((q.rend - row0 < WIDE_TR) ? (q.rend - row0) : WIDE_TR);
          const uint32_t full = bars + 8u * s, dst = sbase + s * stage_bytes;
          mbar_expect_tx(full, (uint32_t)rows * segbytes + mubytes);
          if (a.panel) {
            bulk_g2s(dst, a.D + ((size_t)q.slab * a.n + row0) * PANEL_PITCH, (uint32_t)rows * segbytes, full);
          } else {
            for (int i = 0; i < rows; ++i)
              bulk_g2s(dst + (uint32_t)i * pitch, a.D + (size_t)(row0 + i) * a.ld + q.cbeg, segbytes, full);
          }
          if (MODE == MODE_ACCUM)
            bulk_g2s(dst + WIDE_TR * pitch, a.mu + (q.isY ? (size_t)a.npad * R : 0) + (size_t)row0 * R, mubytes, full);   // whole tile: the arrays are padded
          if (++s == (uint32_t)a.nstage) { s = 0; par ^= 1u; }
        }
      }
    }
  } else {
    // ------------------------------------ compute warps --------------------------------------
    const uint32_t xoff = (uint32_t)lane * 16u;
    uint32_t s = 0, par = 0;
    for (int it = blockIdx.x; it < nitem; it += gridDim.x) {
    const Item q = item_of(it);
    const int slab = q.slab, rg = q.rg, width = q.width, cbeg = q.cbeg, ntile = q.ntile;
    const int64_t rbeg = q.rbeg, rend = q.rend;
    double wv[NCH][2][R];            // MODE_PROJECT: loadings of my columns; MODE_ACCUM: G accumulators
#pragma unroll
    for (int j = 0; j < NCH; ++j) {
      const int c = 2 * (j * 32 + lane);
#pragma unroll
      for (int h = 0; h < 2; ++h)
#pragma unroll
        for (int k = 0; k < R; ++k)
          wv[j][h][k] = (MODE == MODE_PROJECT && c < width) ? a.L[(size_t)(cbeg + c + h) * R + k] : 0.0;
    }
    for (int t = 0; t < ntile; ++t) {
      const int64_t row0 = rbeg + (int64_t)(MODE == MODE_ACCUM ? ntile - 1 - t : t) * WIDE_TR;   // pass 2 walks back
      const int rows = (int)((rend - row0 < WIDE_TR) ? (rend - row0) : WIDE_TR);
      mbar_wait(bars + 8u * s, par);
      const uint32_t tile = sbase + s * stage_bytes;
      // rows warp and warp + 8 of the tile (a short last tile keeps stale finite rows: masked below)
      const uint32_t ra = tile + (uint32_t)warp * pitch + xoff, rb = ra + 8u * pitch;
      if (MODE == MODE_PROJECT) {
        constexpr int NVP = pow2_ceil(2 * R);
        double v[NVP];
#pragma unroll
        for (int e = 0; e < NVP; ++e) v[e] = 0.0;
#pragma unroll
        for (int j = 0; j < NCH; ++j) {
          double x0, x1, y0, y1;
          lds128<0>(x0, x1, ra + j * 512);
          lds128<0>(y0, y1, rb + j * 512);
#pragma unroll
          for (int k = 0; k < R; ++k) {
            v[k] = fma(x0, wv[j][0][k], v[k]);     v[k] = fma(x1, wv[j][1][k], v[k]);
            v[R + k] = fma(y0, wv[j][0][k], v[R + k]); v[R + k] = fma(y1, wv[j][1][k], v[R + k]);
          }
        }
        warp_reduce_scatter<NVP>(v, lane);
        constexpr int SH = ilog2(32 / NVP);
        const int idx = (lane >> SH) & (NVP - 1);
        if ((lane & ((1 << SH) - 1)) == 0 && idx < 2 * R) {
          const int rr = idx / R, k = idx - rr * R;
          const int i = warp + 8 * rr;
          if (i < rows) a.ppart[((size_t)slab * a.n + row0 + i) * R + k] = v[0];
        }
      } else {
        const uint32_t mua = tile + WIDE_TR * pitch;
        double m0[R], m1[R];
#pragma unroll
        for (int k = 0; k < R; ++k) {
          double t0, t1;
          asm volatile("ld.shared.f64 %0, [%1];" : "=d"(t0) : "r"(mua + (uint32_t)(warp * R + k) * 8u));
          asm volatile("ld.shared.f64 %0, [%1];" : "=d"(t1) : "r"(mua + (uint32_t)((warp + 8) * R + k) * 8u));
          m0[k] = (warp < rows) ? t0 : 0.0;
          m1[k] = (warp + 8 < rows) ? t1 : 0.0;
        }
#pragma unroll
        for (int j = 0; j < NCH; ++j) {
          double x0, x1, y0, y1;
          lds128<0>(x0, x1, ra + j * 512);
          lds128<0>(y0, y1, rb + j * 512);
#pragma unroll
          for (int k = 0; k < R; ++k) {
            wv[j][0][k] = fma(x0, m0[k], wv[j][0][k]); wv[j][1][k] = fma(x1, m0[k], wv[j][1][k]);
            wv[j][0][k] = fma(y0, m1[k], wv[j][0][k]); wv[j][1][k] = fma(y1, m1[k], wv[j][1][k]);
          }
        }
      }
      // release the stage only once every load from it has returned (zero_once_known, em_stream_kernel.cuh)
      uint32_t dep;
      if (MODE == MODE_PROJECT) {
        dep = zero_once_known(hi_bits(wv[0][0][0]));         // (the reduction above has consumed every load already)
      } else {
        uint32_t hx = 0;
#pragma unroll
        for (int j = 0; j < NCH; ++j) hx ^= hi_bits(wv[j][0][0]) ^ hi_bits(wv[j][1][0]);
#pragma unroll
        for (int k = 1; k < R; ++k) hx ^= hi_bits(wv[NCH - 1][0][k]);
        dep = zero_once_known(hx);
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(bars + 8u * (a.nstage + s) + dep);
      if (++s == (uint32_t)a.nstage) { s = 0; par ^= 1u; }
    }
    if (MODE == MODE_ACCUM) {
      // this item's partial G: add the 8 warps' accumulators in warp order through shared memory (a buffer of
      // its own: the producer is already filling the ring for the next item)
      asm volatile("bar.sync 1, %0;" ::"n"(WIDE_WARPS * 32) : "memory");
      double* gs = reinterpret_cast<double*>(smem + gs_off);        // [NCH*64][R]
      for (int w = 0; w < WIDE_WARPS; ++w) {
        if (warp == w) {
#pragma unroll
          for (int j = 0; j < NCH; ++j)
#pragma unroll
            for (int h = 0; h < 2; ++h)
#pragma unroll
              for (int k = 0; k < R; ++k) {
                double* g = gs + (size_t)(2 * (j * 32 + lane) + h) * R + k;
                *g = (w == 0) ? wv[j][h][k] : (*g + wv[j][h][k]);
              }
        }
        asm volatile("bar.sync 1, %0;" ::"n"(WIDE_WARPS * 32) : "memory");
      }
      double* out = a.gpart + ((size_t)rg * a.ld + cbeg) * R;
      for (int e = tid; e < width * R; e += WIDE_WARPS * 32) out[e] = gs[e];
    }
    }
  }
}

// proj[i][0..R) = sum over X slabs of the partial projections, proj[i][R..2R) = sum over Y slabs.
// Fixed order: slice y of 8 adds slabs y, y+8, ...; the 8 slices are then added in order.
__global__ void __launch_bounds__(256) wide_proj_sum_kernel(const double* __restrict__ ppart, int64_t n, int R, int nslabX,
                                                            int nslabY, double* __restrict__ proj, const FitState* st) {
  if (st->done) return;
  __shared__ double sh[2][8][33];
  const int64_t e = (int64_t)blockIdx.x * 32 + threadIdx.x;
  const size_t slab_stride = (size_t)n * R;
  double sx = 0.0, sy = 0.0;
  if (e < n * R) {
    for (int s = threadIdx.y; s < nslabX; s += 8) sx += ppart[(size_t)s * slab_stride + e];
    for (int s = nslabX + threadIdx.y; s < nslabX + nslabY; s += 8) sy += ppart[(size_t)s * slab_stride + e];
  }
  sh[0][threadIdx.y][threadIdx.x] = sx; sh[1][threadIdx.y][threadIdx.x] = sy;
  __syncthreads();
  if (threadIdx.y < 2 && e < n * R) {
    double t = 0.0;
#pragma unroll
    for (int y = 0; y < 8; ++y) t += sh[threadIdx.y][y][threadIdx.x];
    const int64_t i = e / R; const int k = (int)(e - i * R);
    proj[i * 2 * R + threadIdx.y * R + k] = t;
  }
}

// mu_T, mu_U of every row (EM_W_multi.R:691-694) and per-block partial Gram sums of [Xw Yc].
constexpr int WMU_ROWS = 128;
__global__ void __launch_bounds__(WMU_ROWS) wide_mu_kernel(const double* __restrict__ proj, int64_t n, int R,
                                                           double* __restrict__ mu /*[2][npad][R]*/, int64_t npad,
                                                           double* __restrict__ gram_part /*[nblk][3R^2]*/,
                                                           const FitState* st, double* __restrict__ mu8 /*[2][npad][8] or null*/) {
  if (st->done) return;
  __shared__ double pj[WMU_ROWS][2 * MAXR];
  const int64_t i = (int64_t)blockIdx.x * WMU_ROWS + threadIdx.x;
  const int rows = (int)((n - (int64_t)blockIdx.x * WMU_ROWS < WMU_ROWS) ? (n - (int64_t)blockIdx.x * WMU_ROWS) : WMU_ROWS);
  if (i < n) {
    for (int k = 0; k < R; ++k) {
      const double xw = proj[i * 2 * R + k], yc = proj[i * 2 * R + R + k];
      pj[threadIdx.x][k] = xw; pj[threadIdx.x][R + k] = yc;
      const double mt = st->aT[k] * xw + st->bT[k] * yc, mw = st->aU[k] * xw + st->bU[k] * yc;
      mu[(size_t)i * R + k] = mt;
      mu[(size_t)(npad + i) * R + k] = mw;
      if (mu8) { mu8[(size_t)i * 8 + k] = mt; mu8[(size_t)(npad + i) * 8 + k] = mw; }     // (pad components / rows stay zero)
    }
  }
  __syncthreads();
  for (int e = threadIdx.x; e < 3 * R * R; e += WMU_ROWS) {
    const int which = e / (R * R), kl = e - which * R * R, k = kl / R, l = kl - k * R;
    const int ca = (which == 2 ? R : 0) + k, cb = (which == 0 ? 0 : R) + l;   // Gxx: x,x  Gxy: x,y  Gyy: y,y
    double s = 0.0;
    for (int r = 0; r < rows; ++r) s = fma(pj[r][ca], pj[r][cb], s);
    gram_part[(size_t)blockIdx.x * 3 * R * R + e] = s;
  }
}

// moment vector = [ G (sum over row groups) | Gxx | Gxy | Gyy (sum over the Gram partials) ]
// The first blocks sum G, one thread per element; the last ones sum the Gram partials, one warp per element (the single-pass
// form leaves one partial per row: lanes stride over them with two chains each, then a fixed-order butterfly).
__global__ void wide_finalize_kernel(const double* __restrict__ gpart, int nrg, int64_t glen,
                                     const double* __restrict__ gram_part, int nblk, int glen2,
                                     double* __restrict__ momvec, const FitState* st, int nb_g) {
  if (st->done) return;
  if ((int)blockIdx.x < nb_g) {
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e < glen) {
      double s = 0.0;
      for (int g = 0; g < nrg; ++g) s += gpart[(size_t)g * glen + e];
      momvec[e] = s;
    }
    return;
  }
  const int f = ((int)blockIdx.x - nb_g) * (int)(blockDim.x >> 5) + (int)(threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (f >= glen2) return;
  double s0 = 0.0, s1 = 0.0;
  int b = lane;
  for (; b + 32 < nblk; b += 64) { s0 += gram_part[(size_t)b * glen2 + f]; s1 += gram_part[(size_t)(b + 32) * glen2 + f]; }
  if (b < nblk) s0 += gram_part[(size_t)b * glen2 + f];
  double s = s0 + s1;
#pragma unroll
  for (int o = 16; o >= 1; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if (lane == 0) momvec[glen + f] = s;
}

// ---------------------------------------------------------------------------------------------
typedef void (*wide_kernel_t)(const WideArgs);
template <int R, int NCH, int MODE>
static wide_kernel_t wide_if() {
  if constexpr (NCH * R <= 24) return wide_kernel<R, NCH, MODE>;
  else return nullptr;
}
template <int R>
static wide_kernel_t wide_lookup_r(int NCH, int mode) {
#define PPLS_W(c) if (NCH == c) return mode == MODE_PROJECT ? wide_if<R, c, MODE_PROJECT>() : wide_if<R, c, MODE_ACCUM>();
  PPLS_W(1) PPLS_W(2) PPLS_W(3) PPLS_W(4) PPLS_W(6) PPLS_W(8)
#undef PPLS_W
  return nullptr;
}
static wide_kernel_t wide_lookup(int R, int NCH, int mode) {
  switch (R) {
    case 1: return wide_lookup_r<1>(NCH, mode);
    case 2: return wide_lookup_r<2>(NCH, mode);
    case 3: return wide_lookup_r<3>(NCH, mode);
    case 4: return wide_lookup_r<4>(NCH, mode);
    case 5: return wide_lookup_r<5>(NCH, mode);
    case 6: return wide_lookup_r<6>(NCH, mode);
    case 7: return wide_lookup_r<7>(NCH, mode);
    case 8: return wide_lookup_r<8>(NCH, mode);
  }
  return nullptr;
}

static size_t wide_smem(int R, int NCH, int nstage, int mode, int panel = 0) {
  const size_t stage = (size_t)WIDE_TR * (panel ? PANEL_PITCH : NCH * 64) * 8 + (mode == MODE_ACCUM ? (size_t)WIDE_TR * R * 8 : 0);
  return stage * nstage + (mode == MODE_ACCUM ? (size_t)NCH * 64 * R * 8 : 0) + 2 * (size_t)nstage * 8;
}

cudaError_t plan_wide(const DataShape& s, int R, int sm_count, WidePlan* w, const char** why) {
  *why = "";
  const int nch_cap = (24 / R) < 8 ? (24 / R) : 8;                 // register budget: NCH * R <= 24
  const int allowed[6] = {1, 2, 3, 4, 6, 8};
  int slots = sm_count;                                              // CTAs resident at once
  if (s.panel) {
    // panel-major storage: one slab per stored panel (PANEL_W = 128 columns, NCH = 2)
    w->scX = w->scY = PANEL_W; w->nslabX = s.npX; w->nslabY = s.npY; w->NCH = PANEL_W / 64;
    slots = 2 * sm_count;
  } else {
    int nch_max = 1;
    for (int c : allowed) if (c <= nch_cap) nch_max = c;
    const int sc_max = nch_max * 64;
    const int64_t cols = (int64_t)s.p2 + s.q2;
    int waves = (int)((cols + (int64_t)sm_count * sc_max - 1) / ((int64_t)sm_count * sc_max));
    if (waves < 1) waves = 1;
    int target = sm_count * waves;
    if (target > cols / 2) target = (int)(cols / 2);
    int nY = (int)((double)target * s.q2 / (double)cols + 0.5);
    if (nY < 1) nY = 1;
    if (nY > s.q2 / 2) nY = s.q2 / 2;
    int nX = target - nY;
    if (nX < 1) nX = 1;
    if (nX > s.p2 / 2) nX = s.p2 / 2;
    w->scX = 2 * ((s.p2 / 2 + nX - 1) / nX); w->nslabX = (s.p2 + w->scX - 1) / w->scX;
    w->scY = 2 * ((s.q2 / 2 + nY - 1) / nY); w->nslabY = (s.q2 + w->scY - 1) / w->scY;
    const int scm = w->scX > w->scY ? w->scX : w->scY;
    if (scm > sc_max) { *why = "wide path: column slab exceeds the register budget"; return cudaErrorInvalidValue; }
    w->NCH = 8;
    for (int i = 5; i >= 0; --i) if (allowed[i] * 64 >= scm) w->NCH = allowed[i];
  }
  w->R = R;
  w->nstage = 4;
  while (w->nstage > 2 && wide_smem(R, w->NCH, w->nstage, MODE_ACCUM, s.panel) > 200 * 1024) --w->nstage;
  // CTAs resident at once, as the runtime sees it for these kernels (registers and shared memory)
  int occ_min = 1 << 30;
  for (int mode = MODE_PROJECT; mode <= MODE_ACCUM; ++mode) {
    wide_kernel_t k = wide_lookup(R, w->NCH, mode);
    if (!k) { *why = "wide path: no kernel for this (nr_comp, slab width)"; return cudaErrorInvalidValue; }
    const size_t smem = wide_smem(R, w->NCH, w->nstage, mode, s.panel);
    cudaError_t e = cudaFuncSetAttribute((const void*)k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) { *why = "cudaFuncSetAttribute failed (wide path)"; return e; }
    int occ = 0;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, (const void*)k, (WIDE_WARPS + 1) * 32, smem);
    if (e != cudaSuccess || occ < 1) { *why = "wide path: the kernel does not fit an SM"; return e != cudaSuccess ? e : cudaErrorInvalidValue; }
    if (occ < occ_min) occ_min = occ;
  }
  if (occ_min > 2) occ_min = 2;                                       // two rings per SM already keep > 128 KB in flight
  if (!s.panel) occ_min = 1;                                          // the slab split above is sized for one CTA per SM
  slots = occ_min * sm_count;
  // row groups: fill the resident-CTA slots in whole waves (each group keeps >= 4 tiles of rows)
  const int nslab = w->nslabX + w->nslabY;
  const int64_t max_rg = (s.n + 4 * WIDE_TR - 1) / (4 * WIDE_TR);
  int nrg = 1; double best_eff = -1.0;
  for (int c = 1; c <= 12 && c <= max_rg; ++c) {
    const int64_t ctas = (int64_t)nslab * c;
    const double eff = (double)ctas / (double)(((ctas + slots - 1) / slots) * slots) - 0.004 * (c - 1);
    if (eff > best_eff) { best_eff = eff; nrg = c; }
  }
  w->rows_per_rg = (int)(((s.n + nrg - 1) / nrg + WIDE_TR - 1) / WIDE_TR * WIDE_TR);
  w->nrg = (int)((s.n + w->rows_per_rg - 1) / w->rows_per_rg);
  { const char* ep = getenv("PPLS_WIDE_PERSIST"); w->slots = (ep && *ep && atoi(ep) == 0) ? 0 : slots; }   // 0: one CTA per work item
  w->nblk_mu = (int)((s.n + WMU_ROWS - 1) / WMU_ROWS);
  w->v2 = 0; w->nstage2 = 0; w->sm_count = sm_count; w->v3 = 0;
  {
    const char* ed = getenv("PPLS_WIDE_DMMA");
    if (s.panel && !(ed && *ed && atoi(ed) == 0)) plan_wide2(s, sm_count, w);      // tensor-pipe form of both passes
  }
  if (getenv("PPLS_K1_VERBOSE") && atoi(getenv("PPLS_K1_VERBOSE")))
    fprintf(stderr, "[ppls] wide plan: R=%d slabs X %d x %d cols, Y %d x %d cols, NCH=%d, row groups %d x %d rows, stages=%d, resident CTAs %d\n",
            R, w->nslabX, w->scX, w->nslabY, w->scY, w->NCH, w->nrg, w->rows_per_rg, w->nstage, slots);
  return cudaSuccess;
}

// Single-pass form: the pipelined kernel, then the same fixed-order sums as the two-pass form.
cudaError_t launch_wide_single(const WidePlan& w, const DataShape& s, const double* D, const double* L, const FitState* st,
                               double* ws, const WideOffsets& o, double* momvec, cudaStream_t stream) {
  // one row lane: the kernel's G is final, it goes straight into the moment vector and only the Gram partials are summed
  const bool direct = w.nrg == 1;
  cudaError_t e = launch_wide3(w, s, D, L, st, ws + o.ppart, ws + o.mu, ws + o.mu8, direct ? momvec : ws + o.gpart, ws + o.gram, ws + o.cnt, stream);
  if (e != cudaSuccess) return e;
  const int64_t glen = (int64_t)s.ld * w.R; const int glen2 = 3 * w.R * w.R;
  const int nb_g = direct ? 0 : (int)((glen + 255) / 256);
  wide_finalize_kernel<<<(unsigned)(nb_g + (glen2 + 7) / 8), 256, 0, stream>>>(ws + o.gpart, w.nrg, glen, ws + o.gram, w.nblk_mu, glen2,
                                                                          momvec, st, nb_g);
  return cudaGetLastError();
}

int64_t wide_npad(int64_t n) { return (n + 63) / 64 * 64; }      // whole tiles of either form (16 / 64 rows)

size_t wide_workspace_doubles(const DataShape& s, const WidePlan& w, WideOffsets* o) {
  size_t off = 0;
  auto take = [&off](size_t cnt) { size_t at = off; off += (cnt + 1) & ~(size_t)1; return at; };   // 16-byte aligned blocks
  {
    size_t pp = (size_t)(w.nslabX + w.nslabY) * s.n * w.R;
    if (w.v3) pp = wide3_ring_doubles(s, w);                       // single-pass form: only the chunks in flight
    o->ppart = take(pp);
  }
  o->proj = take((size_t)s.n * 2 * w.R);
  o->mu = take((size_t)2 * wide_npad(s.n) * w.R);
  o->gpart = take((size_t)w.nrg * s.ld * w.R);
  o->gram = take((size_t)w.nblk_mu * 3 * w.R * w.R);
  o->mu8 = take((w.v2 || w.v3) ? (size_t)2 * wide_npad(s.n) * 8 : 0);      // DMMA forms: mu with the components padded to 8
  o->cnt = take(w.v3 ? wide3_counter_doubles(w) : 0);               // single-pass form: chunk arrival counters
  return off;
}

// pass 1 up to (and including) the slab sum: proj[n][2R] is what column-sharded ranks exchange
cudaError_t launch_wide_project(const WidePlan& w, const DataShape& s, const double* D, const double* L,
                                const FitState* st, double* ws, const WideOffsets& o, cudaStream_t stream) {
  WideArgs a{};
  a.D = D; a.L = L; a.st = st; a.ppart = ws + o.ppart; a.mu = ws + o.mu; a.gpart = ws + o.gpart;
  a.n = s.n; a.npad = wide_npad(s.n); a.ld = s.ld; a.p2 = s.p2; a.q2 = s.q2; a.scX = w.scX; a.scY = w.scY; a.nslabX = w.nslabX;
  a.nslabY = w.nslabY; a.nrg = w.nrg; a.rows_per_rg = w.rows_per_rg; a.nstage = w.nstage; a.panel = s.panel;
  const int nitem = (w.nslabX + w.nslabY) * w.nrg;
  const int grid = (w.slots > 0 && w.slots < nitem) ? w.slots : nitem;
  if (w.v2) {
    cudaError_t e2 = launch_wide2(1, w, s, D, L, st, ws + o.ppart, nullptr, nullptr, w.sm_count, stream);
    if (e2 != cudaSuccess) return e2;
  } else
  wide_lookup(w.R, w.NCH, MODE_PROJECT)<<<grid, (WIDE_WARPS + 1) * 32, wide_smem(w.R, w.NCH, w.nstage, MODE_PROJECT, s.panel), stream>>>(a);
  const int64_t ne = s.n * w.R;
  wide_proj_sum_kernel<<<(unsigned)((ne + 31) / 32), dim3(32, 8), 0, stream>>>(ws + o.ppart, s.n, w.R, w.nslabX, w.nslabY, ws + o.proj, st);
  return cudaGetLastError();
}

// mu, pass 2 and the moment vector
cudaError_t launch_wide_accum(const WidePlan& w, const DataShape& s, const double* D, const double* L,
                              const FitState* st, double* ws, const WideOffsets& o, double* momvec,
                              cudaStream_t stream) {
  wide_mu_kernel<<<w.nblk_mu, WMU_ROWS, 0, stream>>>(ws + o.proj, s.n, w.R, ws + o.mu, wide_npad(s.n), ws + o.gram, st,
                                                   w.v2 ? ws + o.mu8 : nullptr);
  WideArgs a{};
  a.D = D; a.L = L; a.st = st; a.ppart = ws + o.ppart; a.mu = ws + o.mu; a.gpart = ws + o.gpart;
  a.n = s.n; a.npad = wide_npad(s.n); a.ld = s.ld; a.p2 = s.p2; a.q2 = s.q2; a.scX = w.scX; a.scY = w.scY; a.nslabX = w.nslabX;
  a.nslabY = w.nslabY; a.nrg = w.nrg; a.rows_per_rg = w.rows_per_rg; a.nstage = w.nstage; a.panel = s.panel;
  const int nitem = (w.nslabX + w.nslabY) * w.nrg;
  const int grid = (w.slots > 0 && w.slots < nitem) ? w.slots : nitem;
  if (w.v2) {
    cudaError_t e2 = launch_wide2(2, w, s, D, L, st, nullptr, ws + o.mu8, ws + o.gpart, w.sm_count, stream);
    if (e2 != cudaSuccess) return e2;
  } else
  wide_lookup(w.R, w.NCH, MODE_ACCUM)<<<grid, (WIDE_WARPS + 1) * 32, wide_smem(w.R, w.NCH, w.nstage, MODE_ACCUM, s.panel), stream>>>(a);
  const int64_t glen = (int64_t)s.ld * w.R; const int glen2 = 3 * w.R * w.R;
  const int nb_g = (int)((glen + 255) / 256);
  wide_finalize_kernel<<<(unsigned)(nb_g + (glen2 + 7) / 8), 256, 0, stream>>>(ws + o.gpart, w.nrg, glen, ws + o.gram, w.nblk_mu, glen2,
                                                                          momvec, st, nb_g);
  return cudaGetLastError();
}

}  // namespace ppls
