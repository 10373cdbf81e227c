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
// COMPUTE warps split the COLUMNS: a warp owns up to M consecutive chunks of 64 columns of X
// (or of Y); a lane owns one column pair per chunk and keeps, for its columns, the loadings
// (w) and the running G = X' mu accumulators (acc) in registers for the whole kernel.
//   phase A(t): partial projections part[i][k] += x[i,j] w[j,k] over the thread's columns,
//               summed over the warp's lanes through a transposing shared-memory buffer;
//   REDUCER warp: adds the warps' partials, forms mu_T/mu_U of the tile and the Gram terms;
//   phase B(t): acc[j][k] += x[i,j] mu[i,k], re-reading the tile from shared memory.
// Compute warps run A(t) then B(t-1), so the reducer's latency hides behind phase B.
// A PRODUCER warp keeps the ring fed with cp.async.bulk.  All hand-offs are mbarriers: no
// CTA-wide barrier inside the loop, warps drift freely within the 2-tile window.
#pragma once
#include "common.cuh"
#include "em_tail.cuh"

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
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
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
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok) : "r"(bar), "r"(parity), "r"(0x989680u) /* suspend-time hint (ns): sleep, don't spin */ : "memory");
  } while (!ok);
}
// Stage release vs. outstanding loads.  An mbarrier arrive that frees a TMA-filled stage must not be ISSUED while loads from
// that stage are still outstanding: ptxas orders the arrive after the issue of the loads, not after their return (it was
// seen ahead of the last DMMAs that consume them), and on B200 a refill of the stage was measured to overtake such loads
// (em_wide3.cu: wrong 32-byte sectors in 1 of ~10^6 tiles).  A value that is 0 but only known once `x` is -- a population
// count is at most 32 -- added to the barrier address keeps the arrive behind everything `x` depends on.
// tools/sass_arrive_audit.py checks the compiled kernels for this.
__device__ __forceinline__ uint32_t zero_once_known(uint32_t x) {
  uint32_t z;
  asm volatile("{\n\t.reg .b32 t;\n\tpopc.b32 t, %1;\n\tshr.u32 %0, t, 6;\n\t}" : "=r"(z) : "r"(x));
  return z;
}
__device__ __forceinline__ uint32_t hi_bits(double v) { return (uint32_t)__double2hiint(v); }

__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

constexpr int RING_PAD = 512;    // zeroed bytes after the ring: lanes of a partial last chunk over-read it

struct StreamArgs {
  const double* D;        // [n][ld]
  const double* L;        // [ld][R] loadings (W rows then C rows)
  const FitState* st;
  double* partials;       // [grid][momvec_len]
  double* muT;            // [R][n] (column-major n x R) or null
  double* muU;
  int64_t n, ntiles;
  int32_t ld, p2, q2, nwX, nwY, nstage, write_mu;
  // chunk split: X warps 0..nwX-1; the first bigX of them own cX+1 chunks, the rest cX (same for Y)
  int32_t cX, bigX, cY, bigY;
  // fuse != 0: the launch is cooperative and goes on, after the last tile, with the cross-CTA reduction, the
  // (multi-GPU) exchange and the whole update of the iteration (em_tail.cuh)
  int32_t fuse;
  TailArgs tail;
};

// Launch bound = the most threads the planner may ask for with M chunks per warp: compute warps
// capped per M (wide M -> few fat warps) and by the register file.  Per thread: loadings 4MR +
// accumulators 4MR + partial projections 2TR (phase A) / mu 2TR (phase B) + ~56 for addresses,
// staging and the loads in flight.
// KEEP (the tile's x values stay in registers from phase A to phase B): 4MT more.
__host__ __device__ constexpr int regs_estimate(int R, int M, int T, bool keep = false) {
  return 8 * M * R + (keep ? 4 * M * T : 0) + 2 * T * R + 56;
}
__host__ __device__ constexpr int max_threads_for(int R, int M, int T, bool keep = false) {
  int r = regs_estimate(R, M, T, keep);
  if (r > 255) r = 255;
  int by_regs = (65536 / r) / 128 * 128;         // ptxas budgets registers for the CTA size rounded up to 128
  int by_m = ((M >= 4 ? 6 : (M == 3 ? 8 : (M == 2 ? 10 : 20))) + 2) * 32;
  int t = by_regs < by_m ? by_regs : by_m;
  return t < 128 ? 128 : t;
}

template <int V>
struct RedShape {
  static constexpr int NP = V / 2;                                   // value pairs
  static constexpr int H = (NP * 16 <= 32) ? 16 : (NP * 8 <= 32) ? 8 : (NP * 4 <= 32) ? 4 : (NP * 2 <= 32) ? 2 : 1;
  static constexpr int SEG = 32 / H;                                 // lanes each reader adds up
  // pitch (double2 units) of one value-pair row of the transpose buffer.  LDS.128 is served in
  // groups of 8 lanes; a group holds 8/H value pairs x H lane groups, so pitch = H (mod 8) makes
  // its 8 addresses hit 8 different 16-byte bank groups.
  static constexpr int RP2 = 32 + (H < 8 ? H : 1);
};

// 16-byte shared-memory load at [addr + IMM] (32-bit shared address, immediate offset).
// volatile: stays behind the mbarrier wait that guards the tile; ptxas schedules it freely.
template <int IMM>
__device__ __forceinline__ void lds128(double& a, double& b, uint32_t addr) {
  asm volatile("ld.shared.v2.f64 {%0, %1}, [%2+%3];" : "=d"(a), "=d"(b) : "r"(addr), "n"(IMM));
}

__host__ __device__ constexpr int red_pitch(int V) {
  const int NP = V / 2;
  const int H = (NP * 16 <= 32) ? 16 : (NP * 8 <= 32) ? 8 : (NP * 4 <= 32) ? 4 : (NP * 2 <= 32) ? 2 : 1;
  return 32 + (H < 8 ? H : 1);
}

struct StreamSmem {            // byte offsets of the dynamic shared-memory carve-up
  uint32_t ring, red, wsum, mu, xy, coef, bars, total;
};
__host__ __device__ inline StreamSmem stream_smem_layout(int R, int T, int nstage, int nwc, int ld) {
  const uint32_t V = (uint32_t)T * R, NP = V / 2;
  StreamSmem s;
  uint32_t o = 0;
  s.ring = o; o += (uint32_t)T * ld * 8u * nstage + RING_PAD;
  s.red = o;  o += (uint32_t)nwc * NP * red_pitch(V) * 16u;
  s.wsum = o; o += 2u * nwc * V * 8u;
  s.mu = o;   o += 4u * V * 8u;           // [buf][T/U][V]
  s.xy = o;   o += 2u * V * 8u;           // reducer scratch: Xw, Yc of the tile
  s.coef = o; o += 4u * R * 8u;
  o = (o + 15u) & ~15u;
  s.bars = o; o += (2u * nstage + 4u) * 8u;   // full[ns], empty[ns], part_ready[2], mu_ready[2]
  s.total = o;
  return s;
}

// ------------------------------------------------------------------------------------------
// Compute-warp loop, specialised on the number of 64-column chunks CNT this warp owns.
// ------------------------------------------------------------------------------------------
template <int R, int T, int CNT, bool KEEP>
__device__ __forceinline__ void compute_warp(const StreamArgs& a, unsigned char* smem, const StreamSmem& lay,
                                             int warp, int lane, int nwc, int my_tiles, bool isY, int c0,
                                             double* out) {
  constexpr int V = T * R;
  using RS = RedShape<V>;
  constexpr int NP = RS::NP, H = RS::H, SEG = RS::SEG, RP2 = RS::RP2;
  static_assert(RP2 == red_pitch(V), "pitch");
  const uint32_t rowbytes = (uint32_t)a.ld * 8u;
  const uint32_t stage_bytes = rowbytes * T;
  const uint32_t ring_bytes = stage_bytes * (uint32_t)a.nstage;
  const int npairs = (isY ? a.q2 : a.p2) >> 1;
  const int pair0 = c0 * 32 + lane;
  const int col0 = (isY ? a.p2 : 0) + 2 * pair0;

  uint32_t validmask = 0;
  double w[CNT][2][R], acc[CNT][2][R];
#pragma unroll
  for (int m = 0; m < CNT; ++m) {
    const bool v = pair0 + 32 * m < npairs;
    validmask |= (v ? 1u : 0u) << m;
#pragma unroll
    for (int c = 0; c < 2; ++c)
#pragma unroll
      for (int k = 0; k < R; ++k) {
        w[m][c][k] = v ? a.L[(size_t)(col0 + 64 * m + c) * R + k] : 0.0;
        acc[m][c][k] = 0.0;
      }
  }
  const uint32_t sbase = smem_u32(smem);
  const uint32_t xaddr = sbase + lay.ring + (uint32_t)col0 * 8u;          // this thread's first column, stage 0
  const uint32_t full0 = sbase + lay.bars, empty0 = full0 + 8u * a.nstage;
  const uint32_t part0 = empty0 + 8u * a.nstage, muready0 = part0 + 16u;
  double2* red_w = reinterpret_cast<double2*>(smem + lay.red) + (size_t)warp * NP * RP2;
  const int rvp = lane / H, rh = lane - rvp * H;                           // reader: value pair, lane group
  const bool ract = lane < NP * H;
  const double2* rsrc = red_w + (ract ? rvp : 0) * RP2 + rh;
  double2* wsum2 = reinterpret_cast<double2*>(smem + lay.wsum);
  const uint32_t mu_addr0 = sbase + lay.mu + (isY ? V * 8u : 0u);

  uint32_t offA = 0, sA = 0, phA = 0;      // stage of tile `it` (phase A)
  uint32_t offB = 0, sB = 0;               // stage of tile `it-1` (phase B)

  auto phaseB = [&](int t) {               // G += x (outer) mu over this thread's columns
    const uint32_t buf = (uint32_t)t & 1u;
    mbar_wait(muready0 + 8u * buf, ((uint32_t)t >> 1) & 1u);
    double mus[V];
    const uint32_t ma = mu_addr0 + buf * (2u * V * 8u);
#pragma unroll
    for (int v = 0; v < V; v += 2) {
      if (v == 0) lds128<0>(mus[v], mus[v + 1], ma);
      else lds128<0>(mus[v], mus[v + 1], ma + v * 8);
    }
    const uint32_t rb = xaddr + offB;
#pragma unroll
    for (int i = 0; i < T; ++i) {
      const uint32_t ra = rb + (uint32_t)i * rowbytes;
#pragma unroll
      for (int m = 0; m < CNT; ++m) {
        double x0, x1;
        switch (m) {
          case 0: lds128<0>(x0, x1, ra); break;
          case 1: lds128<512>(x0, x1, ra); break;
          case 2: lds128<1024>(x0, x1, ra); break;
          case 3: lds128<1536>(x0, x1, ra); break;
          case 4: lds128<2048>(x0, x1, ra); break;
          default: lds128<2560>(x0, x1, ra); break;
        }
#pragma unroll
        for (int k = 0; k < R; ++k) {
          acc[m][0][k] = fma(x0, mus[i * R + k], acc[m][0][k]);
          acc[m][1][k] = fma(x1, mus[i * R + k], acc[m][1][k]);
        }
      }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(empty0 + 8u * sB);     // this warp is done with the stage
    offB += stage_bytes; ++sB;
    if (sB == (uint32_t)a.nstage) { sB = 0; offB = 0; }
  };

  if constexpr (KEEP) {
    // ---- KEEP form: the tile is read from shared memory ONCE.  Its values stay in registers from the projection
    // (phase A) to the accumulation (phase B) of the SAME tile; the ring stage is released as soon as phase A has
    // consumed the loads, so the copies of later tiles run under the reducer's turnaround.
    for (int it = 0; it < my_tiles; ++it) {
      const uint32_t buf = (uint32_t)it & 1u;
      mbar_wait(full0 + 8u * sA, phA);
      double x[T][CNT][2];
      const uint32_t rb = xaddr + offA;
#pragma unroll
      for (int i = 0; i < T; ++i) {
        const uint32_t ra = rb + (uint32_t)i * rowbytes;
#pragma unroll
        for (int m = 0; m < CNT; ++m) {
          switch (m) {
            case 0: lds128<0>(x[i][m][0], x[i][m][1], ra); break;
            case 1: lds128<512>(x[i][m][0], x[i][m][1], ra); break;
            case 2: lds128<1024>(x[i][m][0], x[i][m][1], ra); break;
            case 3: lds128<1536>(x[i][m][0], x[i][m][1], ra); break;
            case 4: lds128<2048>(x[i][m][0], x[i][m][1], ra); break;
            default: lds128<2560>(x[i][m][0], x[i][m][1], ra); break;
          }
        }
      }
      double pf[V];
#pragma unroll
      for (int v = 0; v < V; ++v) pf[v] = 0.0;
#pragma unroll
      for (int i = 0; i < T; ++i)
#pragma unroll
        for (int m = 0; m < CNT; ++m)
#pragma unroll
          for (int k = 0; k < R; ++k) {
            pf[i * R + k] = fma(x[i][m][0], w[m][0][k], pf[i * R + k]);
            pf[i * R + k] = fma(x[i][m][1], w[m][1][k], pf[i * R + k]);
          }
      __syncwarp();
      if (lane == 0) mbar_arrive(empty0 + 8u * sA);      // every lane's loads have been consumed: the stage is free
      offA += stage_bytes; ++sA;
      if (sA == (uint32_t)a.nstage) { sA = 0; offA = 0; phA ^= 1u; }
#pragma unroll
      for (int vp = 0; vp < NP; ++vp) red_w[vp * RP2 + lane] = make_double2(pf[2 * vp], pf[2 * vp + 1]);
      __syncwarp();
      double s0 = 0.0, s1 = 0.0, u0 = 0.0, u1 = 0.0;
      if (ract) {
#pragma unroll
        for (int j = 0; j < SEG; j += 2) {
          const double2 e = rsrc[j * H];
          s0 += e.x; s1 += e.y;
          if (j + 1 < SEG) { const double2 f = rsrc[(j + 1) * H]; u0 += f.x; u1 += f.y; }
        }
      }
      s0 += u0; s1 += u1;
#pragma unroll
      for (int o = H / 2; o > 0; o >>= 1) {
        s0 += __shfl_xor_sync(0xffffffffu, s0, o);
        s1 += __shfl_xor_sync(0xffffffffu, s1, o);
      }
      if (ract && rh == 0) wsum2[((size_t)buf * nwc + warp) * NP + rvp] = make_double2(s0, s1);
      __syncwarp();
      if (lane == 0) mbar_arrive(part0 + 8u * buf);
      // ---- phase B of the same tile, from registers ----------------------------------------------------------
      mbar_wait(muready0 + 8u * buf, ((uint32_t)it >> 1) & 1u);
      double mus[V];
      const uint32_t ma = mu_addr0 + buf * (2u * V * 8u);
#pragma unroll
      for (int v = 0; v < V; v += 2) lds128<0>(mus[v], mus[v + 1], ma + v * 8);
#pragma unroll
      for (int i = 0; i < T; ++i)
#pragma unroll
        for (int m = 0; m < CNT; ++m)
#pragma unroll
          for (int k = 0; k < R; ++k) {
            acc[m][0][k] = fma(x[i][m][0], mus[i * R + k], acc[m][0][k]);
            acc[m][1][k] = fma(x[i][m][1], mus[i * R + k], acc[m][1][k]);
          }
    }
  } else {
  for (int it = 0; it < my_tiles; ++it) {
    const uint32_t buf = (uint32_t)it & 1u;
    mbar_wait(full0 + 8u * sA, phA);
    // ---- phase A: partial projections over this thread's columns ------------------------
    double pf[V];
#pragma unroll
    for (int v = 0; v < V; ++v) pf[v] = 0.0;
    const uint32_t rb = xaddr + offA;
#pragma unroll
    for (int i = 0; i < T; ++i) {
      const uint32_t ra = rb + (uint32_t)i * rowbytes;
#pragma unroll
      for (int m = 0; m < CNT; ++m) {
        double x0, x1;
        switch (m) {
          case 0: lds128<0>(x0, x1, ra); break;
          case 1: lds128<512>(x0, x1, ra); break;
          case 2: lds128<1024>(x0, x1, ra); break;
          case 3: lds128<1536>(x0, x1, ra); break;
          case 4: lds128<2048>(x0, x1, ra); break;
          default: lds128<2560>(x0, x1, ra); break;
        }
#pragma unroll
        for (int k = 0; k < R; ++k) {
          pf[i * R + k] = fma(x0, w[m][0][k], pf[i * R + k]);
          pf[i * R + k] = fma(x1, w[m][1][k], pf[i * R + k]);
        }
      }
    }
    offA += stage_bytes; ++sA;
    if (sA == (uint32_t)a.nstage) { sA = 0; offA = 0; phA ^= 1u; }
    // ---- sum over the warp's lanes through the transposing buffer -------------------------
#pragma unroll
    for (int vp = 0; vp < NP; ++vp) red_w[vp * RP2 + lane] = make_double2(pf[2 * vp], pf[2 * vp + 1]);
    __syncwarp();
    double s0 = 0.0, s1 = 0.0, u0 = 0.0, u1 = 0.0;
    if (ract) {
#pragma unroll
      for (int j = 0; j < SEG; j += 2) {
        const double2 e = rsrc[j * H];
        s0 += e.x; s1 += e.y;
        if (j + 1 < SEG) { const double2 f = rsrc[(j + 1) * H]; u0 += f.x; u1 += f.y; }
      }
    }
    s0 += u0; s1 += u1;
#pragma unroll
    for (int o = H / 2; o > 0; o >>= 1) {
      s0 += __shfl_xor_sync(0xffffffffu, s0, o);
      s1 += __shfl_xor_sync(0xffffffffu, s1, o);
    }
    if (ract && rh == 0) wsum2[((size_t)buf * nwc + warp) * NP + rvp] = make_double2(s0, s1);
    __syncwarp();
    if (lane == 0) mbar_arrive(part0 + 8u * buf);
    if (it >= 1) phaseB(it - 1);
  }
  if (my_tiles >= 1) phaseB(my_tiles - 1);
  }

  // ---- epilogue: this CTA's partial G ------------------------------------------------------
#pragma unroll
  for (int m = 0; m < CNT; ++m)
    if ((validmask >> m) & 1u) {
#pragma unroll
      for (int c = 0; c < 2; ++c)
#pragma unroll
        for (int k = 0; k < R; ++k) out[(size_t)(col0 + 64 * m + c) * R + k] = acc[m][c][k];
    }
}

// Warp roles.  Warps 0..nwc-1 (nwc = nwX + nwY): COMPUTE (above).  Warp nwc: REDUCER -- sums
// the warps' projection partials, forms mu_T/mu_U and the Gram terms.  Warp nwc+1: PRODUCER --
// keeps the shared-memory ring fed with cp.async.bulk.
template <int R, int M, int T, bool KEEP>
__global__ void __launch_bounds__(max_threads_for(R, M, T, KEEP), 1) em_stream_kernel(const StreamArgs a) {
  constexpr int V = T * R;
  static_assert(V <= 32 && V % 2 == 0, "T*R must be even and fit one warp");
  if (a.st->done && !a.write_mu) return;         // the posterior-means pass runs after the end
  if (a.fuse && blockIdx.x == 0 && threadIdx.x == 0)
    reinterpret_cast<unsigned long long*>(a.tail.gscratch + tail_clock_offset(gridDim.x, R))[3] = global_timer_ns();

  extern __shared__ __align__(128) unsigned char smem[];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int nwc = a.nwX + a.nwY;
  const StreamSmem lay = stream_smem_layout(R, T, a.nstage, nwc, a.ld);
  const uint32_t rowbytes = (uint32_t)a.ld * 8u, stage_bytes = rowbytes * T;
  double* red = reinterpret_cast<double*>(smem + lay.red);
  double* wsum = reinterpret_cast<double*>(smem + lay.wsum);
  double* mu_s = reinterpret_cast<double*>(smem + lay.mu);
  double* xy_s = reinterpret_cast<double*>(smem + lay.xy);
  double* coef = reinterpret_cast<double*>(smem + lay.coef);
  const uint32_t sbase = smem_u32(smem);
  const uint32_t full0 = sbase + lay.bars, empty0 = full0 + 8u * a.nstage;
  const uint32_t part0 = empty0 + 8u * a.nstage, muready0 = part0 + 16u;

  if (tid < R) {
    coef[0 * R + tid] = a.st->aT[tid];
    coef[1 * R + tid] = a.st->bT[tid];
    coef[2 * R + tid] = a.st->aU[tid];
    coef[3 * R + tid] = a.st->bU[tid];
  }
  // zero the ring (+pad) once: rows of a short last tile that are never loaded, and the bytes a
  // partial last chunk over-reads, must stay finite
  {
    double* z = reinterpret_cast<double*>(smem + lay.ring);
    const uint32_t nz = (stage_bytes * (uint32_t)a.nstage + RING_PAD) / 8u;
    for (uint32_t e = tid; e < nz; e += blockDim.x) z[e] = 0.0;
  }
  if (tid == 0) {
    for (int s = 0; s < a.nstage; ++s) { mbar_init(full0 + 8u * s, 1); mbar_init(empty0 + 8u * s, nwc); }
    for (int b = 0; b < 2; ++b) { mbar_init(part0 + 8u * b, nwc); mbar_init(muready0 + 8u * b, 1); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  fence_proxy_async();
  __syncthreads();

  const int my_tiles = (a.ntiles > (int64_t)blockIdx.x)
                           ? (int)((a.ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x) : 0;
  double* out = a.partials + (size_t)blockIdx.x * momvec_len(a.ld, R);
  double gxx[R], gxy[R], gyy[R];                 // reducer lanes only
#pragma unroll
  for (int l = 0; l < R; ++l) gxx[l] = gxy[l] = gyy[l] = 0.0;

  if (warp < nwc) {
    // ===================================== compute warps =====================================
    const bool isY = warp >= a.nwX;
    const int wg = isY ? warp - a.nwX : warp;
    const int cb = isY ? a.cY : a.cX, big = isY ? a.bigY : a.bigX;
    const int cnt = cb + (wg < big ? 1 : 0);
    const int c0 = wg * cb + (wg < big ? wg : big);
#define PPLS_CW(c) compute_warp<R, T, (M >= c ? c : 1), KEEP>(a, smem, lay, warp, lane, nwc, my_tiles, isY, c0, out)
    if (M >= 6 && cnt == 6) PPLS_CW(6);
    else if (M >= 5 && cnt == 5) PPLS_CW(5);
    else if (M >= 4 && cnt == 4) PPLS_CW(4);
    else if (M >= 3 && cnt == 3) PPLS_CW(3);
    else if (M >= 2 && cnt == 2) PPLS_CW(2);
    else PPLS_CW(1);
#undef PPLS_CW
  } else if (warp == nwc) {
    // ==================================== reducer warp =======================================
    const int i = lane / R, k = lane - i * R;     // lane (i,k) owns Xw[i,k], Yc[i,k] of the tile
    const bool act = lane < V;
    double cT0 = 0, cT1 = 0, cU0 = 0, cU1 = 0;
    if (act) { cT0 = coef[k]; cT1 = coef[R + k]; cU0 = coef[2 * R + k]; cU1 = coef[3 * R + k]; }
    for (int it = 0; it < my_tiles; ++it) {
      const int buf = it & 1;
      const int64_t row0 = ((int64_t)blockIdx.x + (int64_t)it * gridDim.x) * T;
      const int rows = (int)((a.n - row0 < T) ? (a.n - row0) : T);
      mbar_wait(part0 + 8u * buf, ((uint32_t)it >> 1) & 1u);     // all compute warps delivered tile `it`
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
      const double mt = cT0 * xw + cT1 * yc;
      const double mu = cU0 * xw + cU1 * yc;
      if (act) {
        mu_s[(buf * 2 + 0) * V + lane] = mt;
        mu_s[(buf * 2 + 1) * V + lane] = mu;
        xy_s[lane] = xw;
        xy_s[V + lane] = yc;
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(muready0 + 8u * buf);           // mu of tile `it` is ready
      if (act) {
        if (a.write_mu && i < rows) {
          a.muT[(size_t)k * a.n + row0 + i] = mt;
          a.muU[(size_t)k * a.n + row0 + i] = mu;
        }
#pragma unroll
        for (int l = 0; l < R; ++l) {
          const double xl = xy_s[i * R + l], yl = xy_s[V + i * R + l];
          gxx[l] = fma(xw, xl, gxx[l]);
          gxy[l] = fma(xw, yl, gxy[l]);
          gyy[l] = fma(yc, yl, gyy[l]);
        }
      }
      __syncwarp();
    }
  } else if (warp == nwc + 1) {
    // ==================================== producer warp ======================================
    if (lane == 0) {
      uint32_t s = 0, off = 0, par = 1;           // par: parity of the PREVIOUS use of the stage
      for (int it = 0; it < my_tiles; ++it) {
        if (it >= a.nstage) mbar_wait(empty0 + 8u * s, par);     // every compute warp finished B on it
        const int64_t row0 = ((int64_t)blockIdx.x + (int64_t)it * gridDim.x) * T;
        const int rows = (int)((a.n - row0 < T) ? (a.n - row0) : T);
        const uint32_t bytes = (uint32_t)rows * rowbytes;
        mbar_expect_tx(full0 + 8u * s, bytes);
        bulk_g2s(sbase + lay.ring + off, a.D + (size_t)row0 * a.ld, bytes, full0 + 8u * s);
        off += stage_bytes; ++s;
        if (s == (uint32_t)a.nstage) { s = 0; off = 0; par ^= 1u; }
      }
    }
  }
  // Gram partials of this CTA: entry (which, k, l) = sum over tile rows i of lane (i,k)'s slot l.
  __syncthreads();                                // every compute warp is past its last transpose
  if (warp == nwc && lane < V) {
#pragma unroll
    for (int l = 0; l < R; ++l) {
      red[(0 * R + l) * 32 + lane] = gxx[l];
      red[(1 * R + l) * 32 + lane] = gxy[l];
      red[(2 * R + l) * 32 + lane] = gyy[l];
    }
  }
  __syncthreads();
  for (int e = tid; e < 3 * R * R; e += blockDim.x) {
    const int which = e / (R * R), kl = e - which * R * R, k = kl / R, l = kl - k * R;
    double sum = 0.0;
    for (int i = 0; i < T; ++i) sum += red[(which * R + l) * 32 + i * R + k];
    out[(size_t)a.ld * R + e] = sum;
  }
  if (a.fuse) stream_tail<R>(a.tail, a.partials, a.ld, a.p2, smem);
}

typedef void (*stream_kernel_t)(const StreamArgs);
// one translation unit per R (em_stream_inst.cu with -DPPLS_INST_R=r) defines these
stream_kernel_t lookup_kernel_r1(int M, int T, int keep);
stream_kernel_t lookup_kernel_r2(int M, int T, int keep);
stream_kernel_t lookup_kernel_r3(int M, int T, int keep);
stream_kernel_t lookup_kernel_r4(int M, int T, int keep);
stream_kernel_t lookup_kernel_r5(int M, int T, int keep);
stream_kernel_t lookup_kernel_r6(int M, int T, int keep);
stream_kernel_t lookup_kernel_r7(int M, int T, int keep);
stream_kernel_t lookup_kernel_r8(int M, int T, int keep);

}  // namespace ppls
