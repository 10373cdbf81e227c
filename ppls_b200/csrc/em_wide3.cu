// Wide-row path in ONE data pass from HBM ("K4 v3"): a persistent cooperative kernel that pipelines the two passes of
// em_wide2.cu over row CHUNKS so that the second pass finds its data in the L2 cache.
//
// The two-pass form reads X and Y twice from HBM because mu_T, mu_U of a row need the projections of ALL its columns
// (EM_W_multi.R:689-694) before G = [X'mu_T ; Y'mu_U] (:732-733) can be accumulated, and a 400 KB row does not fit one
// SM.  But it does fit the GPU: a chunk of 32 rows of every panel is 13.5 MB at C4 (N = 2000, p = 50000, q = 1000),
// a tenth of the 126 MB L2.  So every CTA owns whole panels (panel % CTAs; accumulators of its <= 6 panels in registers
// for the whole kernel) and all CTAs walk the chunks together, LAG chunks apart:
//   compute warps, iteration w:
//                  P(w)      project the CTA's panel tiles of chunk w   (DMMA; one partial per CTA to a small ring buffer in L2)
//                  A(w-LAG)  re-read the same tiles -- from L2 -- with the chunk's mu (which the producer waited for) and
//                            accumulate G
//   signal warp    publishes "this CTA's projections of chunk w are stored" to the chunk's arrival counter (the gpu-scope
//                  fence this takes would stall the tensor pipe if a compute warp executed it)
//   reducer warps  the CTA's share of the 32 one-row duties of every chunk: wait until every CTA's projections are in, sum
//                  them in fixed order, form mu_T, mu_U, add the row's Gram terms to the CTA's Gram partial, count the duty
//   producer warp  bulk copies into the ring, as far ahead as the ring allows; the first read of a tile asks the L2 to keep
//                  it (evict_last), the second is its last (evict_first)
//   mu warp        waits for the chunk's 32 duties and stages its mu_T / mu_U tile in shared memory for the compute warps
// Nothing stops: the waits are on per-chunk arrival counters that only grow (targets derived from the fit's pass number),
// the compute warps never wait on another CTA directly, only on their own ring.  HBM traffic per EM iteration is ONE read of
// the data (algorithmic bytes 8 n (p+q), half of the two-pass form; measured 884 MB at C4 against 816 MB); the second read
// is L2 bandwidth.
// Shapes with fewer panels than CTAs split the rows into `lanes` (row groups) with a pipeline each.
// Column-sharded multi-GPU runs keep the two-kernel form (the projections are exchanged between the passes).
#include "em_stream_kernel.cuh"
#include <cstdio>
#include <cstdlib>

namespace ppls {

constexpr int W3_TR = 32;                        // rows per tile = rows per chunk
constexpr int W3_WARPS = 8;                      // compute warps
constexpr uint32_t W3_PITCH = PANEL_PITCH * 8;   // shared-memory row pitch (bytes) = stored pitch (conflict-free fragments)
constexpr uint32_t W3_MU_BYTES = W3_TR * 8 * 8;  // mu tile: [32 rows][8 components]
constexpr uint32_t W3_STAGE = W3_TR * W3_PITCH;  // one tile: 32 rows of one panel
constexpr int W3_MUSLOTS = 2;                    // chunks of mu staged ahead in shared memory ([mu_T tile | mu_U tile] each)
constexpr int W3_WPITCH = PANEL_W + 4;           // loadings in shared memory: [8 components][132] per slot
constexpr int W3_NRED = 32;                      // reducer duties per chunk (1 row each: 8 components x 16 sub-lanes)
constexpr int W3_RWARPS = 4;                     // reducer warps per CTA (their own instruction stream: duties never stall the tensor pipe)
constexpr int W3_THREADS = (W3_WARPS + 3 + W3_RWARPS) * 32;   // compute | producer, signal, mu | reducers
constexpr int W3_MAXSLOT = 6;

struct Wide3Args {
  const double* D; const double* L; const FitState* st;
  double* pring;                 // [lanes][LAG+1][CTAs per lane][2 halves][X|Y][32][R]  projections of the chunks in flight
  double* mu;                    // [2][npad][R]              mu_T, mu_U (kept for ppls_fit_result)
  double* mu8;                   // [2][npad][8]              the same, components padded to 8 (operand of the A tiles)
  double* gpart;                 // [lanes][ld][R]            G per lane
  double* gram_part;             // [CTAs][3 R^2]             Gram partials of the reducers (one per CTA: its duties' rows)
  unsigned long long* cnt;       // [2][nchunk]               arrivals: CTAs whose projections are in, duties done
  unsigned long long* trace;     // PPLS_WIDE_TRACE=1: [nchunk][12] globaltimer stamps of the chain of one chunk (else null)
  int64_t n, npad;
  int32_t ld, p2, npX, npanel, R, nstage, lag;
  int32_t lanes, ctas_per_lane, chunks_per_lane, nchunk, hint;
};

__device__ __forceinline__ void w3_dmma(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
__device__ __forceinline__ double w3_lds(uint32_t addr) {
  double v;
  asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ void w3_bar_red() { asm volatile("bar.sync 2, %0;" ::"n"(W3_RWARPS * 32) : "memory"); }
// Arrival counters: release-add (fire and forget: orders the arriving thread's earlier writes -- and, cumulatively, those it
// has observed through __syncwarp / bar.sync -- before the increment, without stalling it on a fence) and acquire-poll.
__device__ __forceinline__ void w3_arrive(unsigned long long* c) {
  asm volatile("red.release.gpu.global.add.u64 [%0], 1;" ::"l"(c) : "memory");
}
__device__ __forceinline__ void w3_wait(const unsigned long long* c, unsigned long long target) {
  unsigned long long v;
  do {
    asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(c) : "memory");
  } while (v < target);
}

__device__ __forceinline__ void w3_stamp(unsigned long long* trace, int chunk, int what) {
  if (trace) trace[(size_t)chunk * 12 + what] = global_timer_ns();
}

struct Wide3Smem { uint32_t ring, ws, mu, comb, xy, bars, total; };
__host__ __device__ inline Wide3Smem wide3_smem(int NS, int nstage) {
  Wide3Smem s;
  uint32_t o = 0;
  s.ring = o; o += (uint32_t)nstage * W3_STAGE;
  s.ws = o;   o += (uint32_t)NS * 8u * W3_WPITCH * 8u;
  s.mu = o;   o += (uint32_t)W3_MUSLOTS * 2u * W3_MU_BYTES;
  s.comb = o; o += 64u;                                  // 16 counters (chunk % 16) of warps that have stored a chunk's projections
  s.xy = o;   o += 64u * 8u;                             // reducer scratch: Xw, Yc of 2 rows (+ coefficients)
  s.bars = o; o += (2u * (uint32_t)nstage + 2u * W3_MUSLOTS) * 8u;     // ring full/empty, mu full/empty
  s.total = o;
  return s;
}

template <int NS>
__global__ void __launch_bounds__(W3_THREADS, 1) wide3_kernel(const Wide3Args a) {
  if (a.st->done) return;
  extern __shared__ __align__(128) unsigned char smem[];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int g = lane >> 2, t = lane & 3;
  const int R = a.R;
  const int ln = (int)blockIdx.x / a.ctas_per_lane, pos = (int)blockIdx.x - ln * a.ctas_per_lane;
  if (ln >= a.lanes) return;                                                   // leftover CTAs
  const int ns = (a.npanel > pos) ? (a.npanel - pos + a.ctas_per_lane - 1) / a.ctas_per_lane : 0;   // my panels: pos, pos + ctas_per_lane, ...
  const int c_lo = ln * a.chunks_per_lane;
  const int nch = (a.nchunk - c_lo < a.chunks_per_lane) ? (a.nchunk - c_lo) : a.chunks_per_lane;
  const unsigned long long pass = (unsigned long long)a.st->pass;               // moves only in the update kernel that follows
  const unsigned long long targetP = (unsigned long long)((a.npanel < a.ctas_per_lane) ? a.npanel : a.ctas_per_lane) * (pass + 1);
  const unsigned long long targetR = (unsigned long long)W3_NRED * (pass + 1);
  unsigned long long* cntP = a.cnt;
  unsigned long long* cntR = a.cnt + a.nchunk;
  const Wide3Smem lay = wide3_smem(NS, a.nstage);
  const uint32_t sbase = smem_u32(smem);
  const uint32_t bars = sbase + lay.bars;
  const uint32_t mubars = bars + 16u * (uint32_t)a.nstage;          // mu full[W3_MUSLOTS], mu empty[W3_MUSLOTS]
  double* Ws = reinterpret_cast<double*>(smem + lay.ws);
  unsigned int* sig = reinterpret_cast<unsigned int*>(smem + lay.comb);
  double* xy = reinterpret_cast<double*>(smem + lay.xy);

  // zero the ring once (rows of a short last tile keep finite stale data; their mu is zero / they are not stored)
  for (uint32_t e = tid; e < (uint32_t)a.nstage * W3_STAGE / 8u; e += blockDim.x) reinterpret_cast<double*>(smem + lay.ring)[e] = 0.0;
  // loadings of my panels, component-major with a conflict-free pitch: Ws[slot][k][j] = L[(cbeg + j)][k]
  for (int e = tid; e < NS * 8 * W3_WPITCH; e += blockDim.x) {
    const int sl = e / (8 * W3_WPITCH), rem = e - sl * 8 * W3_WPITCH, k = rem / W3_WPITCH, j = rem - k * W3_WPITCH;
    double v = 0.0;
    if (sl < ns && k < R && j < PANEL_W) {
      const int panel = pos + sl * a.ctas_per_lane;
      const bool isY = panel >= a.npX;
      const int c = (isY ? a.p2 + (panel - a.npX) * PANEL_W : panel * PANEL_W) + j;
      if (c < (isY ? a.ld : a.p2)) v = a.L[(size_t)c * R + k];
    }
    Ws[e] = v;
  }
  if (tid < 16) sig[tid] = 0u;
  if (tid == 0) {
    for (int s = 0; s < a.nstage; ++s) { mbar_init(bars + 8u * s, 1); mbar_init(bars + 8u * (a.nstage + s), W3_WARPS); }
    for (int s = 0; s < W3_MUSLOTS; ++s) { mbar_init(mubars + 8u * s, 1); mbar_init(mubars + 8u * (W3_MUSLOTS + s), W3_WARPS); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  fence_proxy_async();
  __syncthreads();

  const int nit = nch + a.lag;
  if (warp == W3_WARPS) {
    // ------------------------------------ producer (one lane) -------------------------------------------------------
    // Only ever waits for a free stage: the tiles of both phases are issued as far ahead as the ring allows (the chunk's
    // mu reaches the compute warps through the mu warp, not through the ring).
    if (lane == 0 && ns > 0) {
      uint32_t s = 0, par = 1;
      int64_t gt = 0;
      // L2 policy: the first read of a tile should stay in the L2 until its second read LAG chunks later, which is its last
      unsigned long long pol_keep, pol_last;
      asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol_keep));
      asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol_last));
      auto issue = [&](int panel, int chunk, bool second) {
        const int64_t row0 = (int64_t)chunk * W3_TR;
        const int rows = (int)((a.n - row0 < W3_TR) ? (a.n - row0) : W3_TR);
        const uint32_t full = bars + 8u * s, dst = sbase + lay.ring + s * W3_STAGE;
        if (gt >= a.nstage) mbar_wait(bars + 8u * (a.nstage + s), par);
        mbar_expect_tx(full, (uint32_t)rows * W3_PITCH);
        if (a.hint & 1)
          asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
                       ::"r"(dst), "l"(a.D + ((size_t)panel * a.n + row0) * PANEL_PITCH), "r"((uint32_t)rows * W3_PITCH), "r"(full),
                         "l"(second ? pol_last : pol_keep) : "memory");
        else
          bulk_g2s(dst, a.D + ((size_t)panel * a.n + row0) * PANEL_PITCH, (uint32_t)rows * W3_PITCH, full);
        ++gt;
        if (++s == (uint32_t)a.nstage) { s = 0; par ^= 1u; }
      };
      for (int w = 0; w < nit; ++w) {
        if (w < nch)
          for (int sl = 0; sl < ns; ++sl) issue(pos + sl * a.ctas_per_lane, c_lo + w, false);
        if (w >= a.lag)
          for (int sl = 0; sl < ns; ++sl) issue(pos + sl * a.ctas_per_lane, c_lo + w - a.lag, true);
      }
    }
    return;
  }
  if (warp == W3_WARPS + 2) {
    // ------------------------------------ mu warp --------------------------------------------------------------------
    // Stages the mu_T (and, for a CTA with a Y panel, mu_U) tile of chunk c in shared memory as soon as the chunk's 32
    // duties are counted: generic loads (written by other SMs' generic stores: no proxy crossing), W3_MUSLOTS chunks ahead.
    if (ns > 0) {
      const bool own_y = pos + (ns - 1) * a.ctas_per_lane >= a.npX;
      for (int c = 0; c < nch; ++c) {
        const int slot = c % W3_MUSLOTS;
        if (lane == 0) {
          if (c >= W3_MUSLOTS) mbar_wait(mubars + 8u * (W3_MUSLOTS + slot), (uint32_t)((c / W3_MUSLOTS - 1) & 1));
          w3_wait(cntR + c_lo + c, targetR);                                // the chunk's mu_T, mu_U are complete
          if (blockIdx.x == 0) w3_stamp(a.trace, c_lo + c, 5);
        }
        __syncwarp();
        const double* m0 = a.mu8 + (size_t)(c_lo + c) * W3_TR * 8;
        double* md = reinterpret_cast<double*>(smem + lay.mu + (uint32_t)slot * 2u * W3_MU_BYTES);
        double vt[8], vu[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) vt[j] = __ldcg(m0 + lane + 32 * j);
        if (own_y) {
#pragma unroll
          for (int j = 0; j < 8; ++j) vu[j] = __ldcg(m0 + (size_t)a.npad * 8 + lane + 32 * j);
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) md[lane + 32 * j] = vt[j];
        if (own_y) {
#pragma unroll
          for (int j = 0; j < 8; ++j) md[256 + lane + 32 * j] = vu[j];
        }
        __syncwarp();
        if (lane == 0) { mbar_arrive(mubars + 8u * slot); if (blockIdx.x == 0) w3_stamp(a.trace, c_lo + c, 6); }   // (release: the stores above are in)
      }
    }
    return;
  }
  if (warp == W3_WARPS + 1) {
    // ------------------------------------ signal warp --------------------------------------------------------------
    // Publishes "this CTA's projections of chunk w are in global memory" to the grid.  The gpu-scope fence that this takes
    // (~1-2 us: it waits for the stores to reach the L2) would stall the tensor pipe if the compute warps executed it.
    if (lane == 0 && ns > 0) {
      for (int w = 0; w < nch; ++w) {
        // the warps drift apart by up to a ring of tiles, so arrivals are counted per chunk (16 counters, revisited every
        // 16 chunks: a warp cannot be that far ahead), never as one running total
        unsigned int v;
        const uint32_t sa = sbase + lay.comb + 4u * (uint32_t)(w & 15);
        do { asm volatile("ld.acquire.cta.shared.u32 %0, [%1];" : "=r"(v) : "r"(sa) : "memory"); } while (v < (unsigned int)(W3_WARPS * ((w >> 4) + 1)));
        if (blockIdx.x == 0) w3_stamp(a.trace, c_lo + w, 0);       // this CTA's projections of the chunk are stored
        __threadfence();
        atomicAdd(cntP + c_lo + w, 1ull);
        if (blockIdx.x == 0) w3_stamp(a.trace, c_lo + w, 1);
      }
    }
    return;
  }

  if (warp >= W3_WARPS + 3) {
    // ------------------------------------ reducer warps ------------------------------------------------------------
    // One row per duty: component k = 16-lane half, 16 sub-lanes per component sum the CTAs' partials in a fixed order.
    const int tr = tid - (W3_WARPS + 3) * 32;            // 0 .. 127
    const int k = tr >> 4, y = tr & 15;
    const bool act = k < R;
    const int ring_slots = a.lag + 1;
    const int nact = (a.npanel < a.ctas_per_lane) ? a.npanel : a.ctas_per_lane;          // CTAs of the lane that own panels
    const int nyown = (a.npanel - a.npX < a.ctas_per_lane) ? a.npanel - a.npX : a.ctas_per_lane;   // ... that own Y panels
    double gs[2] = {0.0, 0.0};                         // this CTA's Gram sums over its duties: elements tr and tr + 128 of [3][R][R]
    for (int d = pos; d < nch * W3_NRED; d += a.ctas_per_lane) {                // my duties, in the order the chunks complete
      const int wl = d / W3_NRED, i = d - wl * W3_NRED;
      const int chunk = c_lo + wl;
      const int64_t row = (int64_t)chunk * W3_TR + i;
      const double* base = a.pring + ((size_t)(ln * ring_slots + wl % ring_slots) * a.ctas_per_lane) * 4 * W3_TR * R + (size_t)i * R + k;
      const size_t cstride = (size_t)4 * W3_TR * R, hstride = (size_t)2 * W3_TR * R, sstride = (size_t)W3_TR * R;
      const bool live = act && row < a.n;
      if (tr == 0) { w3_wait(cntP + chunk, targetP); if (i == 0) w3_stamp(a.trace, chunk, 2); }   // every CTA's projections of the chunk are in
      w3_bar_red();
      double sx0 = 0.0, sx1 = 0.0, sy = 0.0;
      for (int c0 = y; c0 < nact; c0 += 160) {                                  // all loads of a sweep in flight, added in a fixed order
        double vx[20], vy[4];
#pragma unroll
        for (int j = 0; j < 10; ++j) {
          const int c = c0 + 16 * j;
          const bool hx = live && c < nact && c < a.npX;                        // CTA c owns an X panel (its first panel is c)
          vx[2 * j] = hx ? __ldcg(base + (size_t)c * cstride) : 0.0;
          vx[2 * j + 1] = hx ? __ldcg(base + (size_t)c * cstride + hstride) : 0.0;
        }
        // CTAs that own Y panels: (npX + m) mod CTAs-per-lane, m < nyown; the first 32 of them ride along with the first sweep
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          const int m = y + 16 * j;
          const bool hy = live && c0 == y && m < nyown;
          const int c = hy ? (a.npX + m) % a.ctas_per_lane : 0;
          vy[2 * j] = hy ? __ldcg(base + (size_t)c * cstride + sstride) : 0.0;
          vy[2 * j + 1] = hy ? __ldcg(base + (size_t)c * cstride + hstride + sstride) : 0.0;
        }
#pragma unroll
        for (int j = 0; j < 20; j += 2) { sx0 += vx[j]; sx1 += vx[j + 1]; }
        sy += (vy[0] + vy[1]) + (vy[2] + vy[3]);
      }
      if (live)
        for (int m = y + 32; m < nyown; m += 16) {                              // (more than 32 Y owners: q > 4096)
          const int c = (a.npX + m) % a.ctas_per_lane;
          sy += __ldcg(base + (size_t)c * cstride + sstride) + __ldcg(base + (size_t)c * cstride + hstride + sstride);
        }
      double xw = sx0 + sx1, yc = sy;
#pragma unroll
      for (int o = 1; o < 16; o <<= 1) { xw += __shfl_xor_sync(0xffffffffu, xw, o); yc += __shfl_xor_sync(0xffffffffu, yc, o); }
      if (tr == 0 && i == 0) w3_stamp(a.trace, chunk, 3);
      if (act && y == 0) {
        const double mt = a.st->aT[k] * xw + a.st->bT[k] * yc, mw = a.st->aU[k] * xw + a.st->bU[k] * yc;
        a.mu[(size_t)row * R + k] = mt;
        a.mu[(size_t)(a.npad + row) * R + k] = mw;
        a.mu8[(size_t)row * 8 + k] = mt;
        a.mu8[(size_t)(a.npad + row) * 8 + k] = mw;
        xy[k] = xw; xy[16 + k] = yc;
        __threadfence();                                   // the writer itself publishes its stores before the count goes up
      }
      w3_bar_red();
#pragma unroll
      for (int m = 0; m < 2; ++m) {
        const int e = tr + m * W3_RWARPS * 32;
        if (e < 3 * R * R) {
          const int which = e / (R * R), kl = e - which * R * R, kk = kl / R, l = kl - kk * R;
          const int oa = (which == 2 ? 16 : 0), ob = (which == 0 ? 0 : 16);    // Gxx: x,x  Gxy: x,y  Gyy: y,y
          gs[m] += xy[oa + kk] * xy[ob + l];                                   // (rows past n have xw = yc = 0)
        }
      }
      if (tr == 0) { atomicAdd(cntR + chunk, 1ull); if (i == 0) w3_stamp(a.trace, chunk, 4); }   // (the Gram partials are read by the next kernel)
      w3_bar_red();                                                             // xy is free again
    }
    // one Gram partial per CTA (its duties in their fixed order), summed over the CTAs by wide_finalize_kernel
#pragma unroll
    for (int m = 0; m < 2; ++m) {
      const int e = tr + m * W3_RWARPS * 32;
      if (e < 3 * R * R) a.gram_part[(size_t)blockIdx.x * 3 * R * R + e] = gs[m];
    }
    return;
  }

  // -------------------------------------- compute warps -------------------------------------------------------
  const int rq = warp & 3, h = warp >> 2;               // P tiles: rows 8 rq .., column half h
  double acc[NS][8];
#pragma unroll
  for (int sl = 0; sl < NS; ++sl)
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[sl][e] = 0.0;
  uint32_t s = 0, par = 0;
  const int ring_slots = a.lag + 1;

  for (int w = 0; w < nit; ++w) {
    // ---- P(w): projections of my panel tiles of chunk w ------------------------------------------------------
    if (w < nch) {
      const int chunk = c_lo + w;
      if (blockIdx.x == 0 && tid == 0) w3_stamp(a.trace, chunk, 8);             // start of the chunk's projection
      const int64_t row0 = (int64_t)chunk * W3_TR;
      const int rows = (int)((a.n - row0 < W3_TR) ? (a.n - row0) : W3_TR);
      // this CTA's slot of the ring: [half h][side X|Y][32 rows][R] -- the projections are summed over the CTA's own panels
      // before they leave it (one partial per CTA, column half and side instead of one per panel)
      double* pr = a.pring + ((size_t)(ln * ring_slots + w % ring_slots) * a.ctas_per_lane + pos) * 4 * W3_TR * R;
      double ax0 = 0.0, ax1 = 0.0, ay0 = 0.0, ay1 = 0.0;
#pragma unroll
      for (int sl = 0; sl < NS; ++sl) {
        if (sl < ns) {
          mbar_wait(bars + 8u * s, par);
          const uint32_t xa = sbase + lay.ring + s * W3_STAGE + (uint32_t)(8 * rq + g) * W3_PITCH + (uint32_t)(64 * h + t) * 8u;
          const uint32_t wa = sbase + lay.ws + (uint32_t)((sl * 8 + g) * W3_WPITCH + 64 * h + t) * 8u;
          double c0 = 0.0, c1 = 0.0, d0 = 0.0, d1 = 0.0, e0 = 0.0, e1 = 0.0, f0 = 0.0, f1 = 0.0;     // four independent chains
#pragma unroll
          for (int k = 0; k < 16; k += 4) {
            const double x0 = w3_lds(xa + k * 32), x1 = w3_lds(xa + (k + 1) * 32), x2 = w3_lds(xa + (k + 2) * 32), x3 = w3_lds(xa + (k + 3) * 32);
            const double b0 = w3_lds(wa + k * 32), b1 = w3_lds(wa + (k + 1) * 32), b2 = w3_lds(wa + (k + 2) * 32), b3 = w3_lds(wa + (k + 3) * 32);
            w3_dmma(c0, c1, x0, b0);
            w3_dmma(d0, d1, x1, b1);
            w3_dmma(e0, e1, x2, b2);
            w3_dmma(f0, f1, x3, b3);
          }
          c0 = (c0 + d0) + (e0 + f0); c1 = (c1 + d1) + (e1 + f1);
          const uint32_t dep = zero_once_known(hi_bits(c0) ^ hi_bits(c1));      // every load of the tile has returned
          __syncwarp();
          if (lane == 0) mbar_arrive(bars + 8u * (a.nstage + s) + dep);
          if (++s == (uint32_t)a.nstage) { s = 0; par ^= 1u; }
          if (pos + sl * a.ctas_per_lane >= a.npX) { ay0 += c0; ay1 += c1; } else { ax0 += c0; ax1 += c1; }
        }
      }
      if (ns > 0) {
        const int i = 8 * rq + g;                            // lane (g, t): components 2t, 2t + 1 of tile row i
        if (i < rows) {
          double* ox = pr + ((size_t)(2 * h + 0) * W3_TR + i) * R;
          double* oy = pr + ((size_t)(2 * h + 1) * W3_TR + i) * R;
          if (2 * t < R) { ox[2 * t] = ax0; oy[2 * t] = ay0; }
          if (2 * t + 1 < R) { ox[2 * t + 1] = ax1; oy[2 * t + 1] = ay1; }
        }
        __syncwarp();                                        // cue for the signal warp: my stores of this chunk are issued
        if (lane == 0) asm volatile("red.release.cta.shared.add.u32 [%0], 1;" ::"r"(sbase + lay.comb + 4u * (uint32_t)(w & 15)) : "memory");
      }
    }
    // ---- A(w - lag): accumulate G from the same tiles (L2) with the chunk's mu ---------------------------------
    if (w >= a.lag && ns > 0) {
      const int c = w - a.lag, mslot = c % W3_MUSLOTS;
      if (blockIdx.x == 0 && tid == 0) w3_stamp(a.trace, c_lo + c, 9);          // (about to wait for the chunk's mu)
      mbar_wait(mubars + 8u * mslot, (uint32_t)((c / W3_MUSLOTS) & 1));          // the mu warp has staged the chunk's mu
      if (blockIdx.x == 0 && tid == 0) w3_stamp(a.trace, c_lo + c, 7);
      const uint32_t mbase = sbase + lay.mu + (uint32_t)mslot * 2u * W3_MU_BYTES + (uint32_t)(t * 8 + g) * 8u;
      uint32_t dep = 0;
#pragma unroll
      for (int sl = 0; sl < NS; ++sl) {
        if (sl < ns) {
          mbar_wait(bars + 8u * s, par);
          const uint32_t tile = sbase + lay.ring + s * W3_STAGE;
          const uint32_t xa = tile + (uint32_t)t * W3_PITCH + (uint32_t)(16 * warp + g) * 8u;
          const uint32_t ma = mbase + ((pos + sl * a.ctas_per_lane >= a.npX) ? W3_MU_BYTES : 0u);   // a Y panel takes mu_U
#pragma unroll
          for (int k = 0; k < 8; k += 2) {
            const double m0 = w3_lds(ma + k * 256), m1 = w3_lds(ma + (k + 1) * 256);
            const double x00 = w3_lds(xa + k * 4 * W3_PITCH), x01 = w3_lds(xa + k * 4 * W3_PITCH + 64);
            const double x10 = w3_lds(xa + (k + 1) * 4 * W3_PITCH), x11 = w3_lds(xa + (k + 1) * 4 * W3_PITCH + 64);
            w3_dmma(acc[sl][0], acc[sl][1], x00, m0);
            w3_dmma(acc[sl][2], acc[sl][3], x01, m0);
            w3_dmma(acc[sl][4], acc[sl][5], x10, m1);
            w3_dmma(acc[sl][6], acc[sl][7], x11, m1);
          }
          // release the stage only once every load of the tile has returned (zero_once_known, em_stream_kernel.cuh): the
          // results of the last four DMMAs depend on the last loads
          dep = zero_once_known(hi_bits(acc[sl][0]) ^ hi_bits(acc[sl][2]) ^ hi_bits(acc[sl][4]) ^ hi_bits(acc[sl][6]));
          __syncwarp();
          if (lane == 0) mbar_arrive(bars + 8u * (a.nstage + s) + dep);
          if (++s == (uint32_t)a.nstage) { s = 0; par ^= 1u; }
        }
      }
      if (lane == 0) mbar_arrive(mubars + 8u * (W3_MUSLOTS + mslot) + dep);      // ... and the chunk's mu slot (same dependency)
    }
  }
  // ---- this lane's rows of G for my panels' columns --------------------------------------------------------------
  double* out = a.gpart + (size_t)ln * a.ld * R;
#pragma unroll
  for (int sl = 0; sl < NS; ++sl) {
    if (sl < ns) {
      const int panel = pos + sl * a.ctas_per_lane;
      const bool isY = panel >= a.npX;
      const int cbeg = isY ? a.p2 + (panel - a.npX) * PANEL_W : panel * PANEL_W;
      const int cend = isY ? a.ld : a.p2;
      const int ca = cbeg + 16 * warp + g, cbk = ca + 8;
      if (ca < cend) {
        if (2 * t < R) out[(size_t)ca * R + 2 * t] = acc[sl][0] + acc[sl][4];
        if (2 * t + 1 < R) out[(size_t)ca * R + 2 * t + 1] = acc[sl][1] + acc[sl][5];
      }
      if (cbk < cend) {
        if (2 * t < R) out[(size_t)cbk * R + 2 * t] = acc[sl][2] + acc[sl][6];
        if (2 * t + 1 < R) out[(size_t)cbk * R + 2 * t + 1] = acc[sl][3] + acc[sl][7];
      }
    }
  }
}

typedef void (*wide3_kernel_t)(const Wide3Args);
static wide3_kernel_t wide3_lookup(int ns, int* NS) {
  if (ns <= 1) { *NS = 1; return wide3_kernel<1>; }
  if (ns <= 2) { *NS = 2; return wide3_kernel<2>; }
  if (ns <= 3) { *NS = 3; return wide3_kernel<3>; }
  if (ns <= 4) { *NS = 4; return wide3_kernel<4>; }
  if (ns <= W3_MAXSLOT) { *NS = W3_MAXSLOT; return wide3_kernel<W3_MAXSLOT>; }
  *NS = 0;
  return nullptr;
}

static int w3_env(const char* name, int dflt) { const char* v = getenv(name); return (v && *v) ? atoi(v) : dflt; }

// Plan of the single-pass form; false: keep the two-pass form (too many panels per CTA, kernel does not fit, switched off).
bool plan_wide3(const DataShape& s, int R, int sm_count, WidePlan* w) {
  w->v3 = 0;
  if (!s.panel || !w3_env("PPLS_WIDE_SINGLE", 1)) return false;
  const int npanel = s.npX + s.npY;
  const int64_t nchunk = (s.n + W3_TR - 1) / W3_TR;
  int lanes = sm_count / npanel;
  if (lanes < 1) lanes = 1;
  if (lanes > nchunk) lanes = (int)nchunk;
  const int cpl = sm_count / lanes;                               // CTAs per lane
  const int ns = (npanel + cpl - 1) / cpl;
  int NS = 0;
  wide3_kernel_t k = wide3_lookup(ns, &NS);
  if (!k) return false;
  const int lag = w3_env("PPLS_WIDE_LAG", 4);
  int nstage = w3_env("PPLS_WIDE_STAGES", 4);
  while (nstage > 2 && wide3_smem(NS, nstage).total > 220 * 1024) --nstage;
  const size_t smem = wide3_smem(NS, nstage).total;
  if (smem > 227 * 1024 || lag < 2 || lag > 8) return false;
  if (cudaFuncSetAttribute((const void*)k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) { cudaGetLastError(); return false; }
  int occ = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, (const void*)k, W3_THREADS, smem) != cudaSuccess || occ < 1) { cudaGetLastError(); return false; }
  w->v3 = 1; w->v3_lanes = lanes; w->v3_cpl = cpl; w->v3_lag = lag; w->v3_nstage = nstage; w->v3_ns = NS;
  w->v3_chunks_per_lane = (int)((nchunk + lanes - 1) / lanes); w->v3_nchunk = (int)nchunk;
  // what the shared finalize kernel sees: one G per lane, one Gram partial per CTA
  w->nrg = lanes; w->nblk_mu = lanes * cpl;
  if (w3_env("PPLS_K1_VERBOSE", 0))
    fprintf(stderr, "[ppls] wide plan (single pass, L2-pipelined): %d panels, %d lane(s) x %d CTAs, <= %d panels per CTA, %lld chunks of %d rows, "
                    "lag %d, %d stages, %zu B shared memory\n", npanel, lanes, cpl, ns, (long long)nchunk, W3_TR, lag, nstage, smem);
  return true;
}

size_t wide3_ring_doubles(const DataShape& s, const WidePlan& w) { (void)s; return (size_t)w.v3_lanes * (w.v3_lag + 1) * w.v3_cpl * 4 * W3_TR * w.R; }
size_t wide3_counter_doubles(const WidePlan& w) { return (size_t)2 * w.v3_nchunk + (w3_env("PPLS_WIDE_TRACE", 0) ? (size_t)12 * w.v3_nchunk : 0); }

// PPLS_WIDE_TRACE=1: the chain of one chunk, relative to "CTA 0's projections stored" (us)
__global__ void wide3_trace_print(const unsigned long long* tr, int nchunk, const FitState* st) {
  if (st->pass != 3) return;
  printf("[ppls] wide3 chain per chunk (us after CTA 0 stored its projections): published | reducer sees all | sums done | duty counted | "
         "mu warp sees 32 duties | mu staged || CTA 0: projection start, wants mu, has mu (relative to the previous chunk's)\n");
  for (int c = 2; c < nchunk; c += (nchunk > 24 ? nchunk / 12 : 1)) {
    const unsigned long long* t = tr + (size_t)c * 12;
    const double t0 = (double)t[0];
    printf("[ppls]   chunk %3d: %5.1f | %5.1f | %5.1f | %5.1f | %5.1f | %5.1f || P start %6.1f  wants mu %6.1f  has mu %6.1f  (period %5.1f)\n", c,
           ((double)t[1] - t0) * 1e-3, ((double)t[2] - t0) * 1e-3, ((double)t[3] - t0) * 1e-3, ((double)t[4] - t0) * 1e-3,
           ((double)t[5] - t0) * 1e-3, ((double)t[6] - t0) * 1e-3, ((double)t[8] - t0) * 1e-3, ((double)t[9] - t0) * 1e-3, ((double)t[7] - t0) * 1e-3,
           ((double)t[0] - (double)(t - 12)[0]) * 1e-3);
  }
}

cudaError_t launch_wide3(const WidePlan& w, const DataShape& s, const double* D, const double* L, const FitState* st, double* pring,
                         double* mu, double* mu8, double* gpart, double* gram_part, double* counters, cudaStream_t stream) {
  Wide3Args a{};
  a.D = D; a.L = L; a.st = st; a.pring = pring; a.mu = mu; a.mu8 = mu8; a.gpart = gpart; a.gram_part = gram_part;
  a.cnt = reinterpret_cast<unsigned long long*>(counters);
  a.trace = w3_env("PPLS_WIDE_TRACE", 0) ? a.cnt + 2 * w.v3_nchunk : nullptr;
  a.n = s.n; a.npad = wide_npad(s.n); a.ld = s.ld; a.p2 = s.p2; a.npX = s.npX; a.npanel = s.npX + s.npY; a.R = w.R;
  a.nstage = w.v3_nstage; a.lag = w.v3_lag; a.lanes = w.v3_lanes; a.ctas_per_lane = w.v3_cpl;
  a.chunks_per_lane = w.v3_chunks_per_lane; a.nchunk = w.v3_nchunk; a.hint = w3_env("PPLS_WIDE_HINT", 1);   // bit 0: L2 eviction policies
  int NS = 0;
  wide3_kernel_t k = wide3_lookup(w.v3_ns, &NS);
  if (!k) return cudaErrorInvalidValue;
  void* args[] = {(void*)&a};
  // cooperative: the CTAs wait on one another's chunk counters, so all of them must be resident (<= 1 per SM by construction)
  cudaError_t e = cudaLaunchCooperativeKernel((const void*)k, dim3(w.v3_lanes * w.v3_cpl), dim3(W3_THREADS), args,
                                              wide3_smem(NS, w.v3_nstage).total, stream);
  if (e == cudaSuccess && a.trace) wide3_trace_print<<<1, 1, 0, stream>>>(a.trace, w.v3_nchunk, st);
  return e;
}

}  // namespace ppls
