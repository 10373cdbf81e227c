// Wide-row path on the FP64 tensor pipe ("K4 v2"): the two data passes of em_wide.cu for PANEL-MAJOR storage
// (D[panel][n][128], wide omics shapes p >> N) with the column / row reductions done INSIDE mma.sync m8n8k4 f64 (DMMA).
//   pass 1 (project)     Xw = X W, Yc = Y C per panel      EM_W_multi.R:689-690      C[8 rows][8 comps] += x[8 x 4] W[4 x 8]
//   pass 2 (accumulate)  G = [X' mu_T ; Y' mu_U]           :732-733 (Maximiz_M)      C[8 cols][8 comps] += x'[8 x 4] mu[4 x 8]
// The DFMA form (em_wide.cu) spends ~3200 warp instructions per 16 KB of data, most of them in the 16-value shuffle
// butterfly that sums a row's projection over a warp's lanes -- more than the issue slots a tile's HBM time offers.
// Here a warp owns 8 ROWS of a 64-row tile in pass 1 (the sum over the panel's 128 columns is the k loop of the
// tensor instruction) and 16 COLUMNS in pass 2 (the sum over the item's rows likewise): no shuffles, no cross-warp sums,
// ~80 / ~110 instructions per warp per 64 KB tile.  nr_comp <= 8 is the N dimension of the instruction (components
// beyond nr_comp are zero operands), so one kernel serves every nr_comp.
// Panel rows are STORED with a pitch of 1024 + 32 bytes (common.cuh PANEL_PITCH), so ONE bulk copy lands a whole tile in
// shared memory with that pitch: a fragment load (8 rows x 4 consecutive doubles in pass 1, 4 rows x 8 consecutive doubles
// in pass 2) then touches 32 distinct banks per half-warp.  (Per-row 1 KB copies into a padded pitch measured slower than
// the DFMA form: 175 us per pass against 155 / 137.)
#include "em_stream_kernel.cuh"
#include <cstdio>
#include <cstdlib>

namespace ppls {

constexpr int W2_TR = 64;                       // rows per tile
constexpr int W2_WARPS = 8;                     // compute warps
constexpr uint32_t W2_PITCH = PANEL_PITCH * 8;   // shared-memory row pitch (bytes) = stored pitch
constexpr uint32_t W2_MU_BYTES = W2_TR * 8 * 8; // mu tile of pass 2: [64 rows][8 components]

struct Wide2Args {
  const double* D; const double* L; const FitState* st;
  double* ppart;            // [npanel][n][R]      (pass 1 out)
  const double* mu8;        // [2][npad][8]        (pass 2 in: mu_T then mu_U, components padded to 8 with zeros)
  double* gpart;            // [nrg][ld][R]        (pass 2 out)
  int64_t n, npad;
  int32_t ld, p2, npX, npanel, nrg, rows_per_rg, nstage, R;
};

__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
__device__ __forceinline__ double lds64(uint32_t addr) {
  double v;
  asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(addr));
  return v;
}

template <int MODE>   // 1 = project, 2 = accumulate
__global__ void __launch_bounds__((W2_WARPS + 1) * 32, 1) wide2_kernel(const Wide2Args a) {
  if (a.st->done) return;
  extern __shared__ __align__(128) unsigned char smem[];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int g = lane >> 2, t = lane & 3;
  const int R = a.R;
  const int nitem = a.npanel * a.nrg;
  const uint32_t mubytes = (MODE == 2) ? W2_MU_BYTES : 0u;
  const uint32_t stage_bytes = W2_TR * W2_PITCH + mubytes;
  const uint32_t sbase = smem_u32(smem);
  const uint32_t bars = sbase + stage_bytes * a.nstage;                  // full[ns], empty[ns]
  // zero the ring once: rows of a short last tile keep whatever is there (finite), their mu is zero / they are not stored
  for (uint32_t e = tid; e < stage_bytes * a.nstage / 8u; e += blockDim.x) reinterpret_cast<double*>(smem)[e] = 0.0;
  if (tid == 0) {
    for (int s = 0; s < a.nstage; ++s) { mbar_init(bars + 8u * s, 1); mbar_init(bars + 8u * (a.nstage + s), W2_WARPS); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  fence_proxy_async();
  __syncthreads();

  // Persistent: CTA b walks the items b, b + grid, ... (item = (panel, row group)) through one continuous ring.  Pass 2
  // walks the CTA's list -- and the tiles of an item -- BACKWARDS: what pass 1 read last is what the L2 still holds.
  const int nmine = (nitem > (int)blockIdx.x) ? (nitem - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
  auto item_at = [&](int j) { return (int)blockIdx.x + ((MODE == 2) ? (nmine - 1 - j) : j) * (int)gridDim.x; };

  if (warp == W2_WARPS) {
    // ------------------------------------ producer: one bulk copy per tile (+ its mu rows in pass 2) ------------
    if (lane == 0) {
      uint32_t s = 0, par = 1;
      int64_t gt = 0;
      for (int j = 0; j < nmine; ++j) {
        const int it = item_at(j);
        const int panel = it % a.npanel, rg = it / a.npanel;
        const int64_t rbeg = (int64_t)rg * a.rows_per_rg;
        const int64_t rend = (rbeg + a.rows_per_rg < a.n) ? rbeg + a.rows_per_rg : a.n;
        const int ntile = (int)((rend - rbeg + W2_TR - 1) / W2_TR);
        const bool isY = panel >= a.npX;
        for (int tt = 0; tt < ntile; ++tt, ++gt) {
          if (gt >= a.nstage) mbar_wait(bars + 8u * (a.nstage + s), par);
          const int64_t row0 = rbeg + (int64_t)(MODE == 2 ? ntile - 1 - tt : tt) * W2_TR;
          const int rows = (int)((rend - row0 < W2_TR) ? (rend - row0) : W2_TR);
          const uint32_t full = bars + 8u * s, dst = sbase + s * stage_bytes;
          mbar_expect_tx(full, (uint32_t)rows * W2_PITCH + mubytes);
          bulk_g2s(dst, a.D + ((size_t)panel * a.n + row0) * PANEL_PITCH, (uint32_t)rows * W2_PITCH, full);
          if (MODE == 2)
            bulk_g2s(dst + W2_TR * W2_PITCH, a.mu8 + ((isY ? (size_t)a.npad : 0) + (size_t)row0) * 8, mubytes, full);
          if (++s == (uint32_t)a.nstage) { s = 0; par ^= 1u; }
        }
      }
    }
  } else {
    // ------------------------------------ compute warps ------------------------------------------------------
    uint32_t s = 0, par = 0;
    for (int j = 0; j < nmine; ++j) {
      const int it = item_at(j);
      const int panel = it % a.npanel, rg = it / a.npanel;
      const int64_t rbeg = (int64_t)rg * a.rows_per_rg;
      const int64_t rend = (rbeg + a.rows_per_rg < a.n) ? rbeg + a.rows_per_rg : a.n;
      const int ntile = (int)((rend - rbeg + W2_TR - 1) / W2_TR);
      const bool isY = panel >= a.npX;
      const int cbeg = isY ? a.p2 + (panel - a.npX) * PANEL_W : panel * PANEL_W;      // first global column of the panel
      const int cend = isY ? a.ld : a.p2;
      if (MODE == 1) {
        // B fragments of the whole panel: W[col 4s + t][component g], zero beyond the panel's columns / nr_comp
        double wb[32];
#pragma unroll
        for (int k = 0; k < 32; ++k) {
          const int c = cbeg + 4 * k + t;
          wb[k] = (c < cend && g < R) ? a.L[(size_t)c * R + g] : 0.0;
        }
        for (int tt = 0; tt < ntile; ++tt) {
          const int64_t row0 = rbeg + (int64_t)tt * W2_TR;
          const int rows = (int)((rend - row0 < W2_TR) ? (rend - row0) : W2_TR);
          mbar_wait(bars + 8u * s, par);
          const uint32_t xa = sbase + s * stage_bytes + (uint32_t)(8 * warp + g) * W2_PITCH + (uint32_t)t * 8u;
          double c0 = 0.0, c1 = 0.0, d0 = 0.0, d1 = 0.0;          // two independent accumulation chains
#pragma unroll
          for (int k = 0; k < 32; k += 2) {
            const double x0 = lds64(xa + k * 32), x1 = lds64(xa + (k + 1) * 32);
            dmma884(c0, c1, x0, wb[k]);
            dmma884(d0, d1, x1, wb[k + 1]);
          }
          const uint32_t dep = zero_once_known(hi_bits(c0) ^ hi_bits(d0));        // every load of the tile has returned
          __syncwarp();
          if (lane == 0) mbar_arrive(bars + 8u * (a.nstage + s) + dep);
          if (++s == (uint32_t)a.nstage) { s = 0; par ^= 1u; }
          // lane (g, t) holds components 2t, 2t + 1 of tile row 8 warp + g
          const int i = 8 * warp + g;
          if (i < rows) {
            double* out = a.ppart + ((size_t)panel * a.n + row0 + i) * R;
            if (2 * t < R) out[2 * t] = c0 + d0;
            if (2 * t + 1 < R) out[2 * t + 1] = c1 + d1;
          }
        }
      } else {
        // accumulators of my 16 columns (two 8-column blocks), components 2t, 2t + 1 of column 8 c + g
        double c00 = 0.0, c01 = 0.0, c10 = 0.0, c11 = 0.0, d00 = 0.0, d01 = 0.0, d10 = 0.0, d11 = 0.0;
        for (int tt = 0; tt < ntile; ++tt) {
          mbar_wait(bars + 8u * s, par);
          const uint32_t tile = sbase + s * stage_bytes;
          const uint32_t xa = tile + (uint32_t)t * W2_PITCH + (uint32_t)(16 * warp + g) * 8u;
          const uint32_t ma = tile + W2_TR * W2_PITCH + (uint32_t)(t * 8 + g) * 8u;
#pragma unroll
          for (int k = 0; k < 16; k += 2) {
            const double m0 = lds64(ma + k * 256), m1 = lds64(ma + (k + 1) * 256);
            const double x00 = lds64(xa + k * 4 * W2_PITCH), x01 = lds64(xa + k * 4 * W2_PITCH + 64);
            const double x10 = lds64(xa + (k + 1) * 4 * W2_PITCH), x11 = lds64(xa + (k + 1) * 4 * W2_PITCH + 64);
            dmma884(c00, c01, x00, m0);
            dmma884(c10, c11, x01, m0);
            dmma884(d00, d01, x10, m1);
            dmma884(d10, d11, x11, m1);
          }
          const uint32_t dep = zero_once_known(hi_bits(c00) ^ hi_bits(c10) ^ hi_bits(d00) ^ hi_bits(d10));   // (as above)
          __syncwarp();
          if (lane == 0) mbar_arrive(bars + 8u * (a.nstage + s) + dep);
          if (++s == (uint32_t)a.nstage) { s = 0; par ^= 1u; }
        }
        // this item's rows of G for the panel's columns (pad columns of a last panel stay out)
        double* out = a.gpart + (size_t)rg * a.ld * R;
        const int ca = cbeg + 16 * warp + g, cb = ca + 8;
        if (ca < cend) {
          if (2 * t < R) out[(size_t)ca * R + 2 * t] = c00 + d00;
          if (2 * t + 1 < R) out[(size_t)ca * R + 2 * t + 1] = c01 + d01;
        }
        if (cb < cend) {
          if (2 * t < R) out[(size_t)cb * R + 2 * t] = c10 + d10;
          if (2 * t + 1 < R) out[(size_t)cb * R + 2 * t + 1] = c11 + d11;
        }
      }
    }
  }
}

static size_t wide2_smem(int nstage, int mode) {
  return ((size_t)W2_TR * W2_PITCH + (mode == 2 ? W2_MU_BYTES : 0)) * nstage + 2 * (size_t)nstage * 8;
}

// Row groups for the DMMA form: whole 64-row tiles per group, as many groups as make the persistent CTAs' item lists
// even (item = (panel, group); every CTA walks ceil(items / grid) of them).
void plan_wide2(const DataShape& s, int sm_count, WidePlan* w) {
  const int npanel = s.npX + s.npY;
  const int64_t ntiles = (s.n + W2_TR - 1) / W2_TR;
  int best = 1; double best_cost = 1e300;
  for (int nrg = 1; nrg <= 64 && nrg <= ntiles; ++nrg) {
    const int64_t tpg = (ntiles + nrg - 1) / nrg;                     // tiles per group
    const int64_t groups = (ntiles + tpg - 1) / tpg;
    const int64_t items = (int64_t)npanel * groups;
    const double cost = (double)((items + sm_count - 1) / sm_count) * ((double)tpg + 0.25);   // + per-item set-up
    if (cost < best_cost - 1e-9) { best_cost = cost; best = (int)groups; w->rows_per_rg = (int)(tpg * W2_TR); }
  }
  w->nrg = best;
  w->v2 = 1;
  w->nstage2 = 3;
  cudaFuncSetAttribute((const void*)wide2_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)wide2_smem(3, 1));
  cudaFuncSetAttribute((const void*)wide2_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)wide2_smem(3, 2));
  if (getenv("PPLS_K1_VERBOSE") && atoi(getenv("PPLS_K1_VERBOSE")))
    fprintf(stderr, "[ppls] wide plan (DMMA form): %d panels x %d row groups of %d rows, %d CTAs, 3 stages of %d rows\n",
            npanel, w->nrg, w->rows_per_rg, sm_count, W2_TR);
}

cudaError_t launch_wide2(int mode, const WidePlan& w, const DataShape& s, const double* D, const double* L, const FitState* st,
                         double* ppart, const double* mu8, double* gpart, int sm_count, cudaStream_t stream) {
  Wide2Args a{};
  a.D = D; a.L = L; a.st = st; a.ppart = ppart; a.mu8 = mu8; a.gpart = gpart;
  a.n = s.n; a.npad = wide_npad(s.n); a.ld = s.ld; a.p2 = s.p2; a.npX = s.npX; a.npanel = s.npX + s.npY;
  a.nrg = w.nrg; a.rows_per_rg = w.rows_per_rg; a.nstage = w.nstage2; a.R = w.R;
  const int nitem = a.npanel * a.nrg;
  const int grid = nitem < sm_count ? nitem : sm_count;
  if (mode == 1) wide2_kernel<1><<<grid, (W2_WARPS + 1) * 32, wide2_smem(a.nstage, 1), stream>>>(a);
  else wide2_kernel<2><<<grid, (W2_WARPS + 1) * 32, wide2_smem(a.nstage, 2), stream>>>(a);
  return cudaGetLastError();
}

}  // namespace ppls
