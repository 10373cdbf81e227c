// Batched multi-start EM (C5) as two FP64 GEMMs per pass -- SURVEY 8e "batched multi-start", kernel K5.
// With S starts of r components the data-dependent part of one EM pass is, for ALL starts at once (RP = S*r
// "components", padded to the 128-wide tile):
//   P  = [X W | Y C]                      (n x 2RP)   bproj_kernel    GEMM "NN": rows x K(columns) x comps
//   mu = [aT.Xw + bT.Yc | aU.Xw + bU.Yc]  (n x 2RP)   bmu_kernel      elementwise + the r x r Gram blocks per start
//   G  = [X' mu_T ; Y' mu_U]              (ld x RP)   baccum_kernel   GEMM "TN": columns x K(rows) x comps
// 4 (p+q) r S flop per row against 8 (p+q) bytes: S r / 2 flop per byte -- FP64-pipe bound for S r >= 32, so the
// streaming kernel K1 (built for the HBM-bound r <= 8 case, 4 rows per tile at r = 8) is the wrong tool here.
// FP64 tensor-core kernels (mma.sync m8n8k4 f64 -- there is no FP64 kind on tcgen05): 128 x 128 tile per CTA, 8 warps
// of 64 x 32 (8 x 4 DMMA tiles, accumulators in registers), cp.async double-buffered operand stages in shared memory
// with pitches chosen so that the 4 k-rows of a fragment fall into disjoint banks.  tools/dmma_probe.cu: DMMA
// sustains 37 TFLOP/s where a shared-memory-fed DFMA register tile stops at 25 (3 operand reads per FMA).
#include "common.cuh"
#include <math.h>
#include <stdlib.h>

namespace ppls {

constexpr int BT = 128;         // tile edge
constexpr int BK = 16;          // K per stage
constexpr int BTHREADS = 256;
constexpr int BACC_WAVES_DEFAULT = 1;     // least number of waves (one CTA per SM) the accumulate GEMM's row splits fill (PPLS_BACC_WAVES)
constexpr int BACC_STAGES_DEFAULT = 3;   // cp.async stages of the accumulate GEMM (PPLS_BACC_STAGES=2..4 overrides)
constexpr int BAP = BK + 4;     // pitch of the row-major A tile of bproj: 4 consecutive rows -> disjoint banks
constexpr int BP = BT + 4;      // pitch of the k-major tiles: 4 consecutive k-rows -> disjoint banks

__device__ __forceinline__ void bcp16(void* smem_dst, const void* gsrc, int src_bytes) {
  const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(d), "l"(gsrc), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void bcommit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bdmma(double (&c)[2], double a, double b) {   // lane l: A[l>>2][l&3], B[l&3][l>>2], C[l>>2][2(l&3)+{0,1}]
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c[0]), "+d"(c[1]) : "d"(a), "d"(b));
}
// m16n8k16: rows {gid, gid+8} x k {tig, tig+4, tig+8, tig+12} -- the same lane mapping as eight m8n8k4 (two row tiles x four
// k-steps); one statement per K = 16 stage (ptxas for sm_100a expands it to eight DMMA.8x8x4, the only FP64 tensor
// instruction of the part, so the rate is the same).  a[v]: row gid + 8 (v & 1), k = tig + 4 (v >> 1);
// b[v]: k = tig + 4 v, n = gid; c: {row gid: cols 2 tig, 2 tig + 1; row gid + 8: the same}.
__device__ __forceinline__ void bdmma16(double (&c0)[2], double (&c1)[2], const double (&a)[8], const double (&b)[4]) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7,%8,%9,%10,%11}, {%12,%13,%14,%15}, {%0,%1,%2,%3};"
               : "+d"(c0[0]), "+d"(c0[1]), "+d"(c1[0]), "+d"(c1[1])
               : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(a[4]), "d"(a[5]), "d"(a[6]), "d"(a[7]),
                 "d"(b[0]), "d"(b[1]), "d"(b[2]), "d"(b[3]));
}
template <int N>
__device__ __forceinline__ void bwait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// Warp-specialised form of the two GEMM kernels (WS = true): 8 consumer warps (fragments + DMMA only) and a
// PRODUCER warpgroup (4 warps) that issues every cp.async of a stage; hand-offs through mbarriers -- full[slot] completes when the
// producer lanes' copies have landed (cp.async.mbarrier.arrive.noinc, one arrival per lane), empty[slot] when the 8
// consumer warps have multiplied the stage.  No CTA-wide barrier and no address arithmetic in the consumers' loop:
// ncu of the __syncthreads form showed the DMMA pipe 60-74 % busy, every warp doing the copies' integer work at the
// same time right after each barrier.
constexpr int WS_THREADS = BTHREADS + 128;   // consumers = warpgroups 0, 1; producers = warpgroup 2
// register budget per warpgroup (setmaxnreg works on whole warpgroups): 256 x 224 + 128 x 56 = 64512 <= 65536
__device__ __forceinline__ void breg_consumer() { asm volatile("setmaxnreg.inc.sync.aligned.u32 224;" ::: "memory"); }
__device__ __forceinline__ void breg_producer() { asm volatile("setmaxnreg.dec.sync.aligned.u32 56;" ::: "memory"); }
__device__ __forceinline__ uint32_t bsm32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void bbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void bbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void bbar_arrive_cpasync(uint32_t bar) {      // arrives when this thread's earlier cp.async have landed
  asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void bbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  do {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
  } while (!ok);
}

// ---- P[row][side*RP + c] = sum_j D[row][c0 + j] * L[c0 + j][c]   (side 0: X columns, side 1: Y columns) -------
struct ProjArgs {
  const double* D; const double* L; double* P;
  int64_t n;
  int ld, p2, q2, RP;            // RP: multiple of BT
};
__global__ void __launch_bounds__(BTHREADS, 1) bproj_kernel(const ProjArgs a) {
  extern __shared__ __align__(16) double sm[];    // [2 stages][ A: BT x BAP | B: BK x BP ]
  const int side = blockIdx.z, cb = blockIdx.y;
  const int64_t row0 = (int64_t)blockIdx.x * BT;
  const int c0 = side ? a.p2 : 0, K = side ? a.q2 : a.p2;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, gid = lane >> 2, tig = lane & 3;
  const int wm = (warp >> 2) * 64, wn = (warp & 3) * 32;         // this warp's rows / components of the tile
  const int nst = (K + BK - 1) / BK;
  constexpr int STAGE = BT * BAP + BK * BP;

  auto issue = [&](int s, int buf) {
    double* As = sm + (size_t)buf * STAGE;
    double* Bs = As + BT * BAP;
    const int k0 = s * BK;
#pragma unroll
    for (int c = 0; c < (BT * BK / 2) / BTHREADS; ++c) {          // A: 128 rows x 8 chunks (row-major, as the data lie)
      const int chunk = tid + c * BTHREADS, r = chunk >> 3, kc = (chunk & 7) * 2;
      const bool ok = (row0 + r < a.n) && (k0 + kc < K);
      bcp16(As + r * BAP + kc, ok ? a.D + (size_t)(row0 + r) * a.ld + c0 + k0 + kc : a.D, ok ? 16 : 0);
    }
#pragma unroll
    for (int c = 0; c < (BK * BT / 2) / BTHREADS; ++c) {          // B: 16 k x 64 pairs of components
      const int chunk = tid + c * BTHREADS, kk = chunk >> 6, pr = chunk & 63;
      const bool ok = k0 + kk < K;
      bcp16(Bs + kk * BP + pr * 2, ok ? a.L + (size_t)(c0 + k0 + kk) * a.RP + cb * BT + pr * 2 : a.L, ok ? 16 : 0);
    }
    bcommit();
  };

  double acc[8][4][2];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) { acc[i][j][0] = 0.0; acc[i][j][1] = 0.0; }
  issue(0, 0);
  for (int s = 0; s < nst; ++s) {
    if (s + 1 < nst) { issue(s + 1, (s + 1) & 1); bwait<1>(); } else bwait<0>();
    __syncthreads();
    const double* As = sm + (size_t)(s & 1) * STAGE;
    const double* Bs = As + BT * BAP;
#pragma unroll
    for (int k4 = 0; k4 < BK / 4; ++k4) {
      const double* ar = As + (wm + gid) * BAP + k4 * 4 + tig;      // A[m = row][k = data column]
      const double* br = Bs + (k4 * 4 + tig) * BP + wn + gid;       // B[k][n = component]
      double af[8], bf[4];
#pragma unroll
      for (int i = 0; i < 8; ++i) af[i] = ar[i * 8 * BAP];
#pragma unroll
      for (int j = 0; j < 4; ++j) bf[j] = br[j * 8];
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) bdmma(acc[i][j], af[i], bf[j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int64_t row = row0 + wm + i * 8 + gid;
    if (row >= a.n) continue;
    double* out = a.P + (size_t)row * 2 * a.RP + (size_t)side * a.RP + cb * BT + wn + 2 * tig;
#pragma unroll
    for (int j = 0; j < 4; ++j) *reinterpret_cast<double2*>(out + j * 8) = make_double2(acc[i][j][0], acc[i][j][1]);
  }
}

// ---- mu and the per-start Gram blocks ------------------------------------------------------------------------
// P -> mu in place: P[row][c] = aT[c] Xw + bT[c] Yc, P[row][RP + c] = aU[c] Xw + bU[c] Yc.
// Gram: gpart[chunk][start][3][R*R] = sum over the chunk's rows of Xw_k Xw_l, Xw_k Yc_l, Yc_k Yc_l.
// Lane = one component (coalesced 8-byte accesses along a row of P); a warp covers 32/R whole starts, lane (s, k)
// keeps row k of its start's three blocks and gets the other components of the start by shuffles; the block's
// 8 warps take every 8th row of the chunk.
struct MuArgs {
  double* P; const double* coef /*[4][RP]: aT bT aU bU*/; double* gpart;
  int64_t n, rows_per_chunk;
  int RP, R, nstart;             // nstart: starts in this group
};
template <int R>
__global__ void __launch_bounds__(256) bmu_kernel(const MuArgs a) {
  constexpr int SW = 32 / R;                      // starts per warp
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int sl = lane / R, k = lane - sl * R;
  const int s = blockIdx.x * SW + sl;
  const bool active = (sl < SW) && (s < a.nstart);
  const int c = active ? s * R + k : 0;
  const int64_t r0 = (int64_t)blockIdx.y * a.rows_per_chunk;
  const int64_t r1 = (r0 + a.rows_per_chunk < a.n) ? r0 + a.rows_per_chunk : a.n;
  const double aT = a.coef[c], bT = a.coef[a.RP + c], aU = a.coef[2 * a.RP + c], bU = a.coef[3 * a.RP + c];
  double gxx[R], gxy[R], gyy[R];
#pragma unroll
  for (int l = 0; l < R; ++l) { gxx[l] = 0.0; gxy[l] = 0.0; gyy[l] = 0.0; }
  const int base = sl * R;                        // first lane of this start
  for (int64_t i = r0 + warp; i < r1; i += 8) {
    double* px = a.P + (size_t)i * 2 * a.RP + c;
    double* py = px + a.RP;
    const double xw = active ? *px : 0.0, yc = active ? *py : 0.0;
#pragma unroll
    for (int l = 0; l < R; ++l) {
      const double xl = __shfl_sync(0xffffffffu, xw, (base + l) & 31), yl = __shfl_sync(0xffffffffu, yc, (base + l) & 31);
      gxx[l] = fma(xw, xl, gxx[l]); gxy[l] = fma(xw, yl, gxy[l]); gyy[l] = fma(yc, yl, gyy[l]);
    }
    if (active) { *px = aT * xw + bT * yc; *py = aU * xw + bU * yc; }
  }
  __shared__ double fold[8][32][3 * R];
#pragma unroll
  for (int l = 0; l < R; ++l) { fold[warp][lane][l] = gxx[l]; fold[warp][lane][R + l] = gxy[l]; fold[warp][lane][2 * R + l] = gyy[l]; }
  __syncthreads();
  if (warp == 0 && active) {                       // warps added in order: deterministic
    double* out = a.gpart + ((size_t)blockIdx.y * a.nstart + s) * 3 * R * R;
#pragma unroll
    for (int v = 0; v < 3 * R; ++v) {
      double t = 0.0;
#pragma unroll
      for (int w = 0; w < 8; ++w) t += fold[w][lane][v];
      out[(v / R) * R * R + k * R + (v % R)] = t;
    }
  }
}

// ---- fused projection + posterior means + Gram blocks (R divides 8) ---------------------------------------------
// One CTA = 128 rows x 64 components, BOTH sides: the K loop runs over the X columns into accx, then over the Y
// columns into accy (one continuous cp.async pipeline, FSTAGES deep), so a thread ends up holding Xw and Yc of the
// same (row, component) and the epilogue writes mu_T / mu_U straight from registers -- P never makes the round trip
// through HBM that bproj_kernel + bmu_kernel cost (n x 2RP doubles written, read and written again).  The r x r Gram
// blocks of a start come from the same registers: R <= 2 keeps a start inside one thread (components 2 tig, 2 tig + 1
// of a DMMA tile), R = 4, 8 fetch the start's other components from the neighbouring lanes of the quad.
constexpr int FT_N = 64;          // components per CTA
constexpr int FSTAGES = 4;
constexpr int FBP = FT_N + 4;     // pitch of the k-major loadings tile
constexpr int FSTAGE = BT * BAP + BK * FBP;
struct ProjMuArgs {
  const double* D; const double* L; const double* coef; double* P; double* gpart;
  int64_t n;
  int ld, p2, q2, RP, nstart, rowtiles;
  int dbg;                       // timing experiments only (PPLS_BGEMM_DBG): 1 = no copies, 2 = no DMMA, 3 = no epilogue, 4 = no Gram, 5 = no mu stores
};
template <int R>
__global__ void __launch_bounds__(WS_THREADS, 1) bprojmu_kernel(const ProjMuArgs a) {
  static_assert(8 % R == 0, "a start must not straddle a DMMA tile");
  static_assert(4 * FT_N == BTHREADS, "one coefficient per consumer thread");
  extern __shared__ __align__(16) double sm[];    // [FSTAGES][ A: BT x BAP | B: BK x FBP ] | gsh [4][FT_N][3 R] | coef [4][FT_N] | mbarriers
  const int ncb2 = a.RP / FT_N;
  const int64_t ntile = (int64_t)ncb2 * a.rowtiles;               // tile t = (row tile t / ncb2, component block t % ncb2)
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, gid = lane >> 2, tig = lane & 3;
  const int wr = warp >> 1, wm = wr * 32, wn = (warp & 1) * 32;  // this warp's rows / components of the tile
  const int nsX = (a.p2 + BK - 1) / BK, nsY = (a.q2 + BK - 1) / BK, nst = nsX + nsY;
  double* gsh = sm + (size_t)FSTAGES * FSTAGE;                    // Gram fold buffer (not the ring: the producer runs ahead)
  double* csh = gsh + 4 * FT_N * 3 * R;                           // coefficients aT, bT, aU, bU of the tile's components
  double* gacc = csh + 4 * FT_N;                                  // [ncb2][FT_N][3 R]: this CTA's Gram sums over all its tiles
  const uint32_t bars = bsm32(gacc + (size_t)ncb2 * FT_N * 3 * R);   // full[FSTAGES], empty[FSTAGES]
  for (int e = tid; e < ncb2 * FT_N * 3 * R; e += WS_THREADS) gacc[e] = 0.0;
  if (tid == 0) {
    for (int b = 0; b < FSTAGES; ++b) { bbar_init(bars + 8u * b, 128); bbar_init(bars + 8u * (FSTAGES + b), BTHREADS / 32); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  if (warp >= BTHREADS / 32) {
    // ================================ producer warpgroup (warps 8..11) ================================
    // One continuous stream of stages over all of this CTA's tiles: the copies of the next tile are in flight while
    // the consumers run the epilogue of the current one.
    breg_producer();
    const int ctid = tid - BTHREADS;
    uint32_t g = 0;                                               // stages issued so far
    for (int64_t t = blockIdx.x; t < ntile; t += gridDim.x) {
      const int cb = (int)(t % ncb2);
      const int64_t row0 = (t / ncb2) * BT;
      for (int s = 0; s < nst; ++s, ++g) {
        const uint32_t slot = g % FSTAGES;
        if (g >= FSTAGES) bbar_wait(bars + 8u * (FSTAGES + slot), (g / FSTAGES - 1u) & 1u);
        double* As = sm + (size_t)slot * FSTAGE;
        double* Bs = As + BT * BAP;
        const bool sideY = s >= nsX;
        const int c0 = sideY ? a.p2 : 0, K = sideY ? a.q2 : a.p2, k0 = (sideY ? s - nsX : s) * BK;
        if (a.dbg != 1) {
#pragma unroll 4
        for (int c = 0; c < (BT * BK / 2) / 128; ++c) {           // A: 128 rows x 8 chunks of 2 columns
          const int chunk = ctid + c * 128, r = chunk >> 3, kc = (chunk & 7) * 2;
          const bool ok = (row0 + r < a.n) && (k0 + kc < K);
          bcp16(As + r * BAP + kc, ok ? a.D + (size_t)(row0 + r) * a.ld + c0 + k0 + kc : a.D, ok ? 16 : 0);
        }
#pragma unroll 4
        for (int c = 0; c < (BK * FT_N / 2) / 128; ++c) {         // B: 16 k x 32 pairs of components
          const int chunk = ctid + c * 128, kk = chunk >> 5, pr = chunk & 31;
          const bool ok = k0 + kk < K;
          bcp16(Bs + kk * FBP + pr * 2, ok ? a.L + (size_t)(c0 + k0 + kk) * a.RP + cb * FT_N + pr * 2 : a.L, ok ? 16 : 0);
        }
        }
        bbar_arrive_cpasync(bars + 8u * slot);
      }
    }
    bwait<0>();
    return;
  }

  // ==================================== consumer warps 0..7 ========================================
  breg_consumer();
  auto csync = [&]() { asm volatile("bar.sync 1, 256;" ::: "memory"); };   // the 256 consumer threads
  uint32_t g = 0;                                                 // stages consumed so far
  for (int64_t t = blockIdx.x; t < ntile; t += gridDim.x) {
    const int cb = (int)(t % ncb2);
    const int64_t row0 = (t / ncb2) * BT;
    // (read after the barrier that follows the K loop; the previous tile's readers are past its second barrier)
    csh[tid] = a.coef[(size_t)(tid / FT_N) * a.RP + cb * FT_N + (tid % FT_N)];
    double accx[4][4][2], accy[4][4][2];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) { accx[i][j][0] = 0.0; accx[i][j][1] = 0.0; accy[i][j][0] = 0.0; accy[i][j][1] = 0.0; }
    for (int s = 0; s < nst; ++s, ++g) {
      const uint32_t slot = g % FSTAGES;
      bbar_wait(bars + 8u * slot, (g / FSTAGES) & 1u);
      const double* As = sm + (size_t)slot * FSTAGE;
      const double* ar = As + (wm + gid) * BAP + tig;               // A[m = row][k = data column]
      const double* br = As + BT * BAP + tig * FBP + wn + gid;      // B[k][n = component]
      // k-steps of 4 columns that hold data: the last stage of a side is usually short (C5: 200 = 12 x 16 + 8 columns
      // of X, 50 = 3 x 16 + 2 of Y), and skipping its empty k-steps saves 5 of 68 per tile
      const bool sideY = s >= nsX;
      const int kleft = (sideY ? a.q2 - (s - nsX) * BK : a.p2 - s * BK);
      const int nk4 = kleft >= BK ? BK / 4 : (kleft + 3) >> 2;
      if (a.dbg != 2)
#pragma unroll
      for (int k4 = 0; k4 < BK / 4; ++k4) {
        if (k4 < nk4) {
          double af[4], bf[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) af[i] = ar[i * 8 * BAP + k4 * 4];
#pragma unroll
          for (int j = 0; j < 4; ++j) bf[j] = br[k4 * 4 * FBP + j * 8];
          if (!sideY) {
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
              for (int j = 0; j < 4; ++j) bdmma(accx[i][j], af[i], bf[j]);
          } else {
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
              for (int j = 0; j < 4; ++j) bdmma(accy[i][j], af[i], bf[j]);
          }
        }
      }
      __syncwarp();
      if (lane == 0) bbar_arrive(bars + 8u * (FSTAGES + slot));
    }
    csync();                                   // the fold of the previous tile has read gsh

    // ---- Gram blocks: entry (component c, which, l) = sum over the tile's rows of v_c * w_{start(c), l} ------------
    // rows past n were zero-filled, so they add nothing
    if (a.dbg != 3 && a.dbg != 4)                        // (3: no epilogue at all, 4: no Gram blocks -- timing experiments)
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        double gg[3][R];
#pragma unroll
        for (int l = 0; l < R; ++l) { gg[0][l] = 0.0; gg[1][l] = 0.0; gg[2][l] = 0.0; }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const double xw = accx[i][j][e], yc = accy[i][j][e];
#pragma unroll
          for (int l = 0; l < R; ++l) {
            double xl, yl;
            if constexpr (R == 1) { xl = xw; yl = yc; }
            else if constexpr (R == 2) { xl = accx[i][j][l]; yl = accy[i][j][l]; }
            else {
              const int c8 = 2 * tig + e, cl = c8 - (c8 % R) + l, src = (lane & ~3) | (cl >> 1);
              const double x0 = __shfl_sync(0xffffffffu, accx[i][j][0], src), x1 = __shfl_sync(0xffffffffu, accx[i][j][1], src);
              const double y0 = __shfl_sync(0xffffffffu, accy[i][j][0], src), y1 = __shfl_sync(0xffffffffu, accy[i][j][1], src);
              xl = (cl & 1) ? x1 : x0; yl = (cl & 1) ? y1 : y0;
            }
            gg[0][l] = fma(xw, xl, gg[0][l]); gg[1][l] = fma(xw, yl, gg[1][l]); gg[2][l] = fma(yc, yl, gg[2][l]);
          }
        }
#pragma unroll
        for (int w = 0; w < 3; ++w)
#pragma unroll
          for (int l = 0; l < R; ++l) {
            double v = gg[w][l];
            v += __shfl_xor_sync(0xffffffffu, v, 4);
            v += __shfl_xor_sync(0xffffffffu, v, 8);
            v += __shfl_xor_sync(0xffffffffu, v, 16);
            if (gid == 0) gsh[((size_t)wr * FT_N + wn + j * 8 + 2 * tig + e) * 3 * R + w * R + l] = v;
          }
      }

    // ---- mu_T, mu_U of the tile, straight from the accumulators ------------------------------------------------
    if (a.dbg != 3 && a.dbg != 5)                        // (5: no mu stores -- timing experiment)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int cl = wn + j * 8 + 2 * tig, c = cb * FT_N + cl;
      const double2 aT = *reinterpret_cast<const double2*>(csh + cl), bT = *reinterpret_cast<const double2*>(csh + FT_N + cl);
      const double2 aU = *reinterpret_cast<const double2*>(csh + 2 * FT_N + cl), bU = *reinterpret_cast<const double2*>(csh + 3 * FT_N + cl);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int64_t row = row0 + wm + i * 8 + gid;
        if (row >= a.n) continue;
        double* out = a.P + (size_t)row * 2 * a.RP + c;
        const double x0 = accx[i][j][0], x1 = accx[i][j][1], y0 = accy[i][j][0], y1 = accy[i][j][1];
        *reinterpret_cast<double2*>(out) = make_double2(aT.x * x0 + bT.x * y0, aT.y * x1 + bT.y * y1);
        *reinterpret_cast<double2*>(out + a.RP) = make_double2(aU.x * x0 + bU.x * y0, aU.y * x1 + bU.y * y1);
      }
    }
    csync();
    // fold the 4 row warps in fixed order and add to this CTA's running sums (element e always belongs to the same thread)
    for (int e = tid; e < FT_N * 3 * R; e += BTHREADS) {
      double tsum = 0.0;
#pragma unroll
      for (int q = 0; q < 4; ++q) tsum += gsh[(size_t)q * FT_N * 3 * R + e];
      gacc[(size_t)cb * FT_N * 3 * R + e] += tsum;
    }
  }
  // this CTA's Gram partial -> gpart[CTA][start][3][R*R] (tiles were added in the order the CTA walked them)
  csync();
  for (int e = tid; e < ncb2 * FT_N * 3 * R; e += BTHREADS) {
    const int c = e / (3 * R), wl = e - c * 3 * R, w = wl / R, l = wl - w * R;
    const int st = c / R, k = c - st * R;
    if (st < a.nstart) a.gpart[((size_t)blockIdx.x * a.nstart + st) * 3 * R * R + w * R * R + k * R + l] = gacc[e];
  }
}

// ---- the same kernel as two HALF-CTAs that work out of phase ("ping-pong") ---------------------------------------------
// bprojmu_kernel's epilogue (Gram blocks + mu stores) is 19 % of its time (282 -> 228 us with the epilogue switched off,
// PPLS_BGEMM_DBG=3), and the tensor pipe idles through it because all 8 consumer warps are in it together.  Here the
// CTA is two independent halves -- consumer warpgroup h (4 warps, a 64-row x 64-component tile, the same 32 x 32 warp
// tiles), two producer warps and a ring of its own each -- walking alternate tiles; half 1 starts half a tile late, so
// one half's epilogue runs under the other half's K loop.  Same arithmetic per element, same fixed-order sums.
// RESULT: slower than bprojmu_kernel (see bproj_pingpong below); kept as an opt-in experiment, parity-tested.
constexpr int HT = 64;                       // rows of a half-CTA's tile
#ifndef PPLS_HSTAGES
#define PPLS_HSTAGES 3
#endif
constexpr int HSTAGES = PPLS_HSTAGES;
constexpr int HSTAGE = HT * BAP + BK * FBP;  // doubles per stage of one half
template <int R>
__global__ void __launch_bounds__(WS_THREADS, 1) bprojmu2_kernel(const ProjMuArgs a) {
  static_assert(8 % R == 0, "a start must not straddle a DMMA tile");
  extern __shared__ __align__(16) double sm[];   // per half: [HSTAGES][A: HT x BAP | B: BK x FBP] | gsh [2][FT_N][3R] | coef [4][FT_N] | gacc | then the mbarriers
  const int ncb2 = a.RP / FT_N;
  const int rowtiles = (int)((a.n + HT - 1) / HT);
  const int64_t ntile = (int64_t)ncb2 * rowtiles;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, gid = lane >> 2, tig = lane & 3;
  const int nsX = (a.p2 + BK - 1) / BK, nsY = (a.q2 + BK - 1) / BK, nst = nsX + nsY;
  const size_t half_doubles = (size_t)HSTAGES * HSTAGE + 2 * FT_N * 3 * R + 4 * FT_N + (size_t)ncb2 * FT_N * 3 * R;
  const int half = (warp < 8) ? (warp >> 2) : ((warp - 8) >> 1);             // consumer warpgroup / producer warp pair
  double* hs = sm + (size_t)half * half_doubles;
  double* gsh = hs + (size_t)HSTAGES * HSTAGE;
  double* csh = gsh + 2 * FT_N * 3 * R;
  double* gacc = csh + 4 * FT_N;
  const uint32_t bars0 = bsm32(sm + 2 * half_doubles);                       // [2 halves][full HSTAGES | empty HSTAGES] | start
  const uint32_t bars = bars0 + (uint32_t)half * 2u * HSTAGES * 8u;
  const uint32_t startbar = bars0 + 4u * HSTAGES * 8u;
  for (int e = tid; e < 2 * ncb2 * FT_N * 3 * R; e += WS_THREADS) {
    const int h = e / (ncb2 * FT_N * 3 * R), r2 = e - h * (ncb2 * FT_N * 3 * R);
    (sm + (size_t)h * half_doubles + (size_t)HSTAGES * HSTAGE + 2 * FT_N * 3 * R + 4 * FT_N)[r2] = 0.0;
  }
  if (tid == 0) {
    for (int h = 0; h < 2; ++h)
      for (int b = 0; b < HSTAGES; ++b) {
        bbar_init(bars0 + (uint32_t)(h * 2 * HSTAGES + b) * 8u, 64);                  // full: the half's 64 producer lanes
        bbar_init(bars0 + (uint32_t)(h * 2 * HSTAGES + HSTAGES + b) * 8u, 4);         // empty: its 4 consumer warps
      }
    bbar_init(startbar, 4);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const int64_t tfirst = 2 * (int64_t)blockIdx.x + half, tstep = 2 * (int64_t)gridDim.x;

  if (warp >= 8) {
    // ================================ producers: warps 8, 9 -> half 0; 10, 11 -> half 1 ================================
    breg_producer();
    const int ctid = (tid - BTHREADS) & 63;
    uint32_t g = 0;
    for (int64_t t = tfirst; t < ntile; t += tstep) {
      const int cb = (int)(t % ncb2);
      const int64_t row0 = (t / ncb2) * HT;
      for (int s = 0; s < nst; ++s, ++g) {
        const uint32_t slot = g % HSTAGES;
        if (g >= HSTAGES) bbar_wait(bars + 8u * (HSTAGES + slot), (g / HSTAGES - 1u) & 1u);
        double* As = hs + (size_t)slot * HSTAGE;
        double* Bs = As + HT * BAP;
        const bool sideY = s >= nsX;
        const int c0 = sideY ? a.p2 : 0, K = sideY ? a.q2 : a.p2, k0 = (sideY ? s - nsX : s) * BK;
        if (a.dbg != 1) {
#pragma unroll 4
          for (int c = 0; c < (HT * BK / 2) / 64; ++c) {            // A: 64 rows x 8 chunks of 2 columns
            const int chunk = ctid + c * 64, r = chunk >> 3, kc = (chunk & 7) * 2;
            const bool ok = (row0 + r < a.n) && (k0 + kc < K);
            bcp16(As + r * BAP + kc, ok ? a.D + (size_t)(row0 + r) * a.ld + c0 + k0 + kc : a.D, ok ? 16 : 0);
          }
#pragma unroll 4
          for (int c = 0; c < (BK * FT_N / 2) / 64; ++c) {          // B: 16 k x 32 pairs of components
            const int chunk = ctid + c * 64, kk = chunk >> 5, pr = chunk & 31;
            const bool ok = k0 + kk < K;
            bcp16(Bs + kk * FBP + pr * 2, ok ? a.L + (size_t)(c0 + k0 + kk) * a.RP + cb * FT_N + pr * 2 : a.L, ok ? 16 : 0);
          }
        }
        bbar_arrive_cpasync(bars + 8u * slot);
      }
    }
    bwait<0>();
    return;
  }

  // ==================================== consumers: warpgroup `half` ========================================
  breg_consumer();
  const int hw = warp & 3, htid = tid & 127;
  const int wr = hw >> 1, wm = wr * 32, wn = (hw & 1) * 32;
  auto hsync = [&]() { if (half == 0) asm volatile("bar.sync 1, 128;" ::: "memory"); else asm volatile("bar.sync 2, 128;" ::: "memory"); };
  uint32_t g = 0;
  bool first = true;
  for (int64_t t = tfirst; t < ntile; t += tstep) {
    const int cb = (int)(t % ncb2);
    const int64_t row0 = (t / ncb2) * HT;
    for (int e = htid; e < 4 * FT_N; e += 128) csh[e] = a.coef[(size_t)(e / FT_N) * a.RP + cb * FT_N + (e % FT_N)];
    double accx[4][4][2], accy[4][4][2];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) { accx[i][j][0] = 0.0; accx[i][j][1] = 0.0; accy[i][j][0] = 0.0; accy[i][j][1] = 0.0; }
    if (first && half == 1 && a.dbg != 6) bbar_wait(startbar, 0); // half 0 is half way through its first tile (6: no stagger, experiment)
    for (int s = 0; s < nst; ++s, ++g) {
      const uint32_t slot = g % HSTAGES;
      bbar_wait(bars + 8u * slot, (g / HSTAGES) & 1u);
      const double* As = hs + (size_t)slot * HSTAGE;
      const double* ar = As + (wm + gid) * BAP + tig;
      const double* br = As + HT * BAP + tig * FBP + wn + gid;
      const bool sideY = s >= nsX;
      const int kleft = (sideY ? a.q2 - (s - nsX) * BK : a.p2 - s * BK);
      const int nk4 = kleft >= BK ? BK / 4 : (kleft + 3) >> 2;
      if (a.dbg != 2)
#pragma unroll
      for (int k4 = 0; k4 < BK / 4; ++k4) {
        if (k4 < nk4) {
          double af[4], bf[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) af[i] = ar[i * 8 * BAP + k4 * 4];
#pragma unroll
          for (int j = 0; j < 4; ++j) bf[j] = br[k4 * 4 * FBP + j * 8];
          if (!sideY) {
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
              for (int j = 0; j < 4; ++j) bdmma(accx[i][j], af[i], bf[j]);
          } else {
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
              for (int j = 0; j < 4; ++j) bdmma(accy[i][j], af[i], bf[j]);
          }
        }
      }
      __syncwarp();
      if (lane == 0) {
        bbar_arrive(bars + 8u * (HSTAGES + slot));
        if (first && half == 0 && s == nst / 2) bbar_arrive(startbar);
      }
    }
    first = false;
    hsync();                                   // the fold of the previous tile has read gsh; csh is complete

    // NOTE: a warp's 32 x 32 tile is rows wm + {0, 8, 16, 24} + gid -- but a half tile has only 64 rows and 2 row warps,
    // so i runs over the warp's FOUR 8-row blocks as in the full-size kernel.
    if (a.dbg != 3 && a.dbg != 4)
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        double gg[3][R];
#pragma unroll
        for (int l = 0; l < R; ++l) { gg[0][l] = 0.0; gg[1][l] = 0.0; gg[2][l] = 0.0; }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const double xw = accx[i][j][e], yc = accy[i][j][e];
#pragma unroll
          for (int l = 0; l < R; ++l) {
            double xl, yl;
            if constexpr (R == 1) { xl = xw; yl = yc; }
            else if constexpr (R == 2) { xl = accx[i][j][l]; yl = accy[i][j][l]; }
            else {
              const int c8 = 2 * tig + e, cl = c8 - (c8 % R) + l, src = (lane & ~3) | (cl >> 1);
              const double x0 = __shfl_sync(0xffffffffu, accx[i][j][0], src), x1 = __shfl_sync(0xffffffffu, accx[i][j][1], src);
              const double y0 = __shfl_sync(0xffffffffu, accy[i][j][0], src), y1 = __shfl_sync(0xffffffffu, accy[i][j][1], src);
              xl = (cl & 1) ? x1 : x0; yl = (cl & 1) ? y1 : y0;
            }
            gg[0][l] = fma(xw, xl, gg[0][l]); gg[1][l] = fma(xw, yl, gg[1][l]); gg[2][l] = fma(yc, yl, gg[2][l]);
          }
        }
#pragma unroll
        for (int w = 0; w < 3; ++w)
#pragma unroll
          for (int l = 0; l < R; ++l) {
            double v = gg[w][l];
            v += __shfl_xor_sync(0xffffffffu, v, 4);
            v += __shfl_xor_sync(0xffffffffu, v, 8);
            v += __shfl_xor_sync(0xffffffffu, v, 16);
            if (gid == 0) gsh[((size_t)wr * FT_N + wn + j * 8 + 2 * tig + e) * 3 * R + w * R + l] = v;
          }
      }
    if (a.dbg != 3 && a.dbg != 5)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int cl = wn + j * 8 + 2 * tig, c = cb * FT_N + cl;
      const double2 aT = *reinterpret_cast<const double2*>(csh + cl), bT = *reinterpret_cast<const double2*>(csh + FT_N + cl);
      const double2 aU = *reinterpret_cast<const double2*>(csh + 2 * FT_N + cl), bU = *reinterpret_cast<const double2*>(csh + 3 * FT_N + cl);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int64_t row = row0 + wm + i * 8 + gid;
        if (row >= a.n) continue;
        double* out = a.P + (size_t)row * 2 * a.RP + c;
        const double x0 = accx[i][j][0], x1 = accx[i][j][1], y0 = accy[i][j][0], y1 = accy[i][j][1];
        *reinterpret_cast<double2*>(out) = make_double2(aT.x * x0 + bT.x * y0, aT.y * x1 + bT.y * y1);
        *reinterpret_cast<double2*>(out + a.RP) = make_double2(aU.x * x0 + bU.x * y0, aU.y * x1 + bU.y * y1);
      }
    }
    hsync();
    for (int e = htid; e < FT_N * 3 * R; e += 128) {
      double tsum = 0.0;
#pragma unroll
      for (int q = 0; q < 2; ++q) tsum += gsh[(size_t)q * FT_N * 3 * R + e];
      gacc[(size_t)cb * FT_N * 3 * R + e] += tsum;
    }
  }
  // a half that never got a tile must still release the other: (half 0 always has one when half 1 has)
  asm volatile("bar.sync 3, 256;" ::: "memory");              // both halves' sums are complete
  const double* g0 = sm + (size_t)HSTAGES * HSTAGE + 2 * FT_N * 3 * R + 4 * FT_N;
  const double* g1 = g0 + half_doubles;
  for (int e = tid; e < ncb2 * FT_N * 3 * R; e += BTHREADS) {
    const int c = e / (3 * R), wl = e - c * 3 * R, w = wl / R, l = wl - w * R;
    const int st = c / R, k = c - st * R;
    if (st < a.nstart) a.gpart[((size_t)blockIdx.x * a.nstart + st) * 3 * R * R + w * R * R + k * R + l] = g0[e] + g1[e];
  }
}

// ---- partial[split][tile] = D[:, col block]' mu_side[:, comp block] over the split's rows ----------------------
struct AccArgs {
  const double* D; const double* MU; double* partial;
  int64_t n, rows_per_split, rows_per_split_mixed;
  int ntiles, nsplit, nsplit_mixed, mixed_db;   // 1-D grid: the plain tiles' splits (tile fastest), then the mixed block's
  int ld, p2, q2, RP, nbx, nby, ncb;     // column blocks of X / Y (BT wide), comp blocks
  int unified;                           // 1: blocks run over the whole row [0, ld); a block may straddle the X|Y boundary
  int dbg;                               // timing experiments only (PPLS_BGEMM_DBG): 1 = no copies, 2 = no DMMA
};
template <int NST, bool WS>
__global__ void __launch_bounds__(WS ? WS_THREADS : BTHREADS, 1) baccum_kernel(const AccArgs a) {
  extern __shared__ __align__(16) double sm[];    // [NST stages][3 operands: D, mu, mu_U (mixed block only)][BK][BP]
  // CTA -> (tile, row split): plain tiles first (tile index fastest), then the tiles of the straddling column block
  int tile, split;
  int64_t rps;
  {
    const int n_mixed = a.mixed_db >= 0 ? a.ncb : 0, n_plain = a.ntiles - n_mixed, bid = (int)blockIdx.x;
    if (bid < n_plain * a.nsplit) {
      const int pt = bid % n_plain;
      split = bid / n_plain;
      tile = (a.mixed_db >= 0 && pt >= a.mixed_db * a.ncb) ? pt + a.ncb : pt;
      rps = a.rows_per_split;
    } else {
      const int r = bid - n_plain * a.nsplit;
      split = r / n_mixed;
      tile = a.mixed_db * a.ncb + r % n_mixed;
      rps = a.rows_per_split_mixed;
    }
  }
  const int cb = tile % a.ncb, db = tile / a.ncb;                           // db: data column block
  // separate mode: X blocks, then Y blocks, each side padded to the tile.  unified mode (p2 % 8 == 0): blocks over
  // the whole row; in the block that straddles the X|Y boundary every 8-column DMMA tile still lies on one side,
  // so it takes its B fragment from mu_T or from mu_U (both staged) -- C5: 250 -> 256 columns instead of 384.
  int side, col0, cend;
  bool mixed = false;
  if (a.unified) {
    col0 = db * BT; cend = a.ld;
    side = col0 >= a.p2;
    mixed = (col0 < a.p2) && (col0 + BT > a.p2);
  } else {
    side = db >= a.nbx;
    col0 = side ? a.p2 + (db - a.nbx) * BT : db * BT;
    cend = side ? a.ld : a.p2;
  }
  const double* Bbase = a.MU + (size_t)side * a.RP + (size_t)cb * BT;       // mixed: side = 0 here, mu_U = Bbase + RP
  const int64_t r0 = (int64_t)split * rps;
  const int64_t r1 = (r0 + rps < a.n) ? r0 + rps : a.n;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, gid = lane >> 2, tig = lane & 3;
  const int wm = (warp >> 2) * 64, wn = (warp & 3) * 32;         // this warp's data columns / components of the tile
  const int nst = (r1 > r0) ? (int)((r1 - r0 + BK - 1) / BK) : 0;

  constexpr int NCOPY = WS ? 128 : BTHREADS;        // threads that issue the copies of a stage
  const int ctid = WS ? tid - BTHREADS : tid;
  auto issue = [&](int s, int buf) {
    double* As = sm + (size_t)(buf * 3 + 0) * BK * BP;
    double* Bs = As + BK * BP;
#pragma unroll 4
    for (int c = 0; c < (BK * BT / 2) / NCOPY; ++c) {
      const int chunk = ctid + c * NCOPY, kk = chunk >> 6, pr = chunk & 63;
      const int64_t grow = r0 + (int64_t)s * BK + kk;
      const bool rok = grow < r1;
      const bool okA = rok && (col0 + pr * 2 < cend);
      bcp16(As + kk * BP + pr * 2, okA ? a.D + (size_t)grow * a.ld + col0 + pr * 2 : a.D, okA ? 16 : 0);
      bcp16(Bs + kk * BP + pr * 2, rok ? Bbase + (size_t)grow * 2 * a.RP + pr * 2 : a.D, rok ? 16 : 0);
      if (mixed) bcp16(Bs + BK * BP + kk * BP + pr * 2, rok ? Bbase + a.RP + (size_t)grow * 2 * a.RP + pr * 2 : a.D, rok ? 16 : 0);
    }
  };
  // per DMMA tile of this warp: does it take mu_U?  (bit i: data columns col0 + wm + 8 i .. + 7)
  unsigned ymask = 0;
  if (mixed)
    for (int i = 0; i < 8; ++i) if (col0 + wm + i * 8 >= a.p2) ymask |= 1u << i;
  [[maybe_unused]] uint32_t bars = 0;                                // WS: full[NST], empty[NST]
  if constexpr (WS) {
    bars = bsm32(sm + (size_t)NST * 3 * BK * BP);
    if (tid == 0) {
      for (int b = 0; b < NST; ++b) { bbar_init(bars + 8u * b, 128); bbar_init(bars + 8u * (NST + b), BTHREADS / 32); }
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (warp >= BTHREADS / 32) {                                    // ---- producer warpgroup (warps 8..11)
      breg_producer();
      for (int s = 0; s < nst; ++s) {
        const int slot = s % NST;
        if (s >= NST) bbar_wait(bars + 8u * (NST + slot), (uint32_t)(s / NST - 1) & 1u);
        if (a.dbg != 1) issue(s, slot);
        bbar_arrive_cpasync(bars + 8u * slot);
      }
      bwait<0>();
      return;
    }
    breg_consumer();
  }
  double acc[8][4][2];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) { acc[i][j][0] = 0.0; acc[i][j][1] = 0.0; }
  auto multiply = [&](int s) {
    static_assert(BK == 16, "one m16n8k16 step per stage");
    const double* As = sm + (size_t)((s % NST) * 3 + 0) * BK * BP;
    const double* ar = As + tig * BP + wm + gid;                    // A[m = data column][k = row]
    const double* br = As + BK * BP + tig * BP + wn + gid;          // B[k = row][n = component]
    const bool uniform = (ymask == 0u || ymask == 0xffu);           // all of this warp's columns on one side
    if (ymask == 0xffu) br += BK * BP;
    double bf[4][4];
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
      for (int v = 0; v < 4; ++v) bf[j][v] = br[v * 4 * BP + j * 8];
    if (uniform) {
#pragma unroll
      for (int ip = 0; ip < 4; ++ip) {                              // column tiles 2 ip, 2 ip + 1
        double af[8];
#pragma unroll
        for (int v = 0; v < 8; ++v) af[v] = ar[(v >> 1) * 4 * BP + (2 * ip + (v & 1)) * 8];
#pragma unroll
        for (int j = 0; j < 4; ++j) bdmma16(acc[2 * ip][j], acc[2 * ip + 1][j], af, bf[j]);
      }
    } else {
      // the X|Y boundary runs through this warp's columns: every 8-column tile still lies on one side; a tile on the
      // other side than its partner gets zeros in the partner's A rows of the other product
#pragma unroll
      for (int ip = 0; ip < 4; ++ip) {
        const bool y0 = (ymask >> (2 * ip)) & 1u, y1 = (ymask >> (2 * ip + 1)) & 1u;
        double af[8];
#pragma unroll
        for (int v = 0; v < 8; ++v) af[v] = ar[(v >> 1) * 4 * BP + (2 * ip + (v & 1)) * 8];
        if (y0 == y1) {
          if (!y0) {
#pragma unroll
            for (int j = 0; j < 4; ++j) bdmma16(acc[2 * ip][j], acc[2 * ip + 1][j], af, bf[j]);
          } else {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              double bg[4];
#pragma unroll
              for (int v = 0; v < 4; ++v) bg[v] = br[BK * BP + v * 4 * BP + j * 8];
              bdmma16(acc[2 * ip][j], acc[2 * ip + 1][j], af, bg);
            }
          }
        } else {                                                    // (ymask is monotone: tile 2 ip on X, 2 ip + 1 on Y)
          double ax[8], ay[8];
#pragma unroll
          for (int v = 0; v < 8; ++v) { ax[v] = (v & 1) ? 0.0 : af[v]; ay[v] = (v & 1) ? af[v] : 0.0; }
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            double bg[4];
#pragma unroll
            for (int v = 0; v < 4; ++v) bg[v] = br[BK * BP + v * 4 * BP + j * 8];
            bdmma16(acc[2 * ip][j], acc[2 * ip + 1][j], ax, bf[j]);
            bdmma16(acc[2 * ip][j], acc[2 * ip + 1][j], ay, bg);
          }
        }
      }
    }
  };
  if constexpr (WS) {
    for (int s = 0; s < nst; ++s) {
      const int slot = s % NST;
      bbar_wait(bars + 8u * slot, (uint32_t)(s / NST) & 1u);
      if (a.dbg != 2) multiply(s);
      __syncwarp();
      if (lane == 0) bbar_arrive(bars + 8u * (NST + slot));
    }
  } else {
#pragma unroll
    for (int s = 0; s < NST - 1; ++s) { if (s < nst) issue(s, s); bcommit(); }
    for (int s = 0; s < nst; ++s) {
      bwait<NST - 2>();                        // stage s has landed (one group is committed per stage slot)
      __syncthreads();                         // ... for every thread; and everyone is done with stage s-1's buffer
      if (s + NST - 1 < nst) issue(s + NST - 1, (s + NST - 1) % NST);
      bcommit();
      multiply(s);
    }
  }
  double* out = a.partial + ((size_t)split * a.ntiles + tile) * BT * BT;
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j)
      *reinterpret_cast<double2*>(out + (wm + i * 8 + gid) * BT + wn + j * 8 + 2 * tig) = make_double2(acc[i][j][0], acc[i][j][1]);
}

// momvec_pack = [G (ld x RP) | Gxx (RP x RP) | Gxy | Gyy]: G summed over the row splits in fixed order, the Gram
// blocks of every start summed over the row chunks and written on the block diagonal (nothing else is read).
struct FinArgs {
  const double* partial; const double* gpart; double* momvec;
  int nsplit, ntiles, ncb, nbx, ld, p2, RP, R, nstart, nchunk, unified, mixed_db, nsplit_mixed;
};
__global__ void __launch_bounds__(256) bfinal_kernel(const FinArgs a) {
  __shared__ double sh[8][33];
  const size_t e = (size_t)blockIdx.x * 32 + threadIdx.x;
  const size_t ng = (size_t)a.ld * a.RP;
  const int R2 = a.R * a.R;
  const size_t nall = ng + (size_t)a.nstart * 3 * R2;
  double s0 = 0.0;
  size_t dst = 0;
  if (e < ng) {
    const int j = (int)(e / a.RP), c = (int)(e % a.RP);
    const int side = j >= a.p2, jj = a.unified ? j : (side ? j - a.p2 : j);
    const int db = (a.unified ? 0 : (side ? a.nbx : 0)) + jj / BT, cb = c / BT, tile = db * a.ncb + cb;
    const size_t off = (size_t)tile * BT * BT + (size_t)(jj % BT) * BT + (c % BT);
    const int ns_t = (db == a.mixed_db) ? a.nsplit_mixed : a.nsplit;
    for (int sp = threadIdx.y; sp < ns_t; sp += 8) s0 += a.partial[(size_t)sp * a.ntiles * BT * BT + off];
    dst = e;
  } else if (e < nall) {
    const size_t f = e - ng;
    const int s = (int)(f / (3 * R2)), w = (int)(f % (3 * R2)), which = w / R2, kl = w % R2, k = kl / a.R, l = kl % a.R;
    for (int ch = threadIdx.y; ch < a.nchunk; ch += 8) s0 += a.gpart[((size_t)ch * a.nstart + s) * 3 * R2 + w];
    dst = ng + (size_t)which * a.RP * a.RP + (size_t)(s * a.R + k) * a.RP + s * a.R + l;
  }
  sh[threadIdx.y][threadIdx.x] = s0;
  __syncthreads();
  if (threadIdx.y == 0 && e < nall) {
    double t = 0.0;
#pragma unroll
    for (int y = 0; y < 8; ++y) t += sh[y][threadIdx.x];
    a.momvec[dst] = t;
  }
}

// ---- host side ---------------------------------------------------------------------------------------------------
void plan_bgemm(const DataShape& s, int RP, int R, int nstart, int sm_count, BGemmPlan* p) {
  p->RP = RP; p->R = R; p->nstart = nstart; p->sm_count = sm_count;
  p->ncb = RP / BT;
  p->nbx = (s.p2 + BT - 1) / BT; p->nby = (s.q2 + BT - 1) / BT;
  p->unified = (s.p2 % 8 == 0) && ((s.ld + BT - 1) / BT < p->nbx + p->nby);
  p->ntiles = (p->unified ? (s.ld + BT - 1) / BT : p->nbx + p->nby) * p->ncb;
  p->rowtiles = (int)((s.n + BT - 1) / BT);
  // accumulate GEMM: row splits so that tiles x splits fills whole waves of one CTA per SM
  const int64_t max_ns = (s.n + 8 * BK - 1) / (8 * BK);
  const char* ew = getenv("PPLS_BACC_WAVES");
  const int waves_min = (ew && *ew && atoi(ew) > 0) ? atoi(ew) : BACC_WAVES_DEFAULT;
  int64_t lo = (waves_min * (int64_t)sm_count + p->ntiles - 1) / p->ntiles, ns = 1;
  double best = -1.0;
  for (int64_t c = lo < 1 ? 1 : lo; c < lo + 64; ++c) {
    const int64_t cc = c > max_ns ? max_ns : c;
    if (cc < 1) continue;
    const int64_t ctas = cc * p->ntiles, waves = (ctas + sm_count - 1) / sm_count;
    const double eff = (double)ctas / (double)(waves * sm_count);
    if (eff > best + 1e-9) { best = eff; ns = cc; }
    if (cc == max_ns) break;
  }
  int64_t rps = (s.n + ns - 1) / ns;
  rps = (rps + BK - 1) / BK * BK;
  p->nsplit = (int)((s.n + rps - 1) / rps);
  p->rows_per_split = rps;
  p->mixed_db = -1; p->nsplit_mixed = p->nsplit; p->rows_per_split_mixed = rps;
  {
    // uneven splits for the straddling block (PPLS_BACC_MIXW = its cost per row relative to a plain block; 1 = even splits)
    const char* em = getenv("PPLS_BACC_MIXW");
    const double c = (em && *em) ? atof(em) : 1.2;
    const bool has_mixed = p->unified && (s.p2 % BT != 0) && (s.p2 < s.ld);
    const int n_mixed = p->ncb, n_plain = p->ntiles - p->ncb;
    const int64_t total = (int64_t)p->nsplit * p->ntiles;
    if (has_mixed && c > 1.0 && total <= sm_count && n_plain >= 1) {
      int64_t nsp = (int64_t)((double)total / ((double)n_plain + c * (double)n_mixed));
      if (nsp < 1) nsp = 1;
      int64_t nsm = (total - nsp * n_plain) / n_mixed;
      if (nsm > max_ns) nsm = max_ns;
      if (nsm >= nsp && nsm >= 1) {
        int64_t rp = ((s.n + nsp - 1) / nsp + BK - 1) / BK * BK, rm = ((s.n + nsm - 1) / nsm + BK - 1) / BK * BK;
        p->mixed_db = s.p2 / BT;
        p->nsplit = (int)((s.n + rp - 1) / rp); p->rows_per_split = rp;
        p->nsplit_mixed = (int)((s.n + rm - 1) / rm); p->rows_per_split_mixed = rm;
      }
    }
  }
  // mu kernel: (32/R starts per warp) x row chunks, about 8 blocks per SM
  const int sb = (nstart + (32 / R) - 1) / (32 / R);
  int64_t nc = (4 * (int64_t)sm_count + sb - 1) / sb;
  const int64_t max_nc = (s.n + 255) / 256;
  if (nc > max_nc) nc = max_nc;
  if (nc < 1) nc = 1;
  int64_t rpc = (s.n + nc - 1) / nc;
  p->nchunk = (int)((s.n + rpc - 1) / rpc);
  p->rows_per_chunk = rpc;
  // fused projection + mu kernel (R divides 8): one Gram partial per row tile
  const char* ef = getenv("PPLS_BATCH_FUSED");
  p->fused = (8 % R == 0) && !(ef && *ef && atoi(ef) == 0);
  if (p->fused) {
    // persistent: one CTA per SM walks the (row tile, component block) tiles (PPLS_BPROJ_PERSIST=0: one CTA per tile);
    // every CTA leaves ONE Gram partial, so nchunk = the grid of bprojmu_kernel
    const char* ep = getenv("PPLS_BPROJ_PERSIST");
    int64_t grid = (int64_t)(RP / FT_N) * p->rowtiles;
    if (!(ep && *ep && atoi(ep) == 0) && sm_count > 0 && grid > sm_count) grid = sm_count;
    p->nchunk = (int)grid; p->rows_per_chunk = BT;
  }
}
size_t bgemm_workspace_doubles(const DataShape& s, const BGemmPlan& p) {
  return (size_t)s.n * 2 * p.RP                                   // P / mu
         + (size_t)(p.nsplit > p.nsplit_mixed ? p.nsplit : p.nsplit_mixed) * p.ntiles * BT * BT     // accumulate partials
         + (size_t)p.nchunk * p.nstart * 3 * p.R * p.R;            // Gram partials
}

template <int R>
static void launch_bmu(const MuArgs& m, const BGemmPlan& p, cudaStream_t stream) {
  constexpr int SW = 32 / R;
  bmu_kernel<R><<<dim3((p.nstart + SW - 1) / SW, p.nchunk), 256, 0, stream>>>(m);
}

// PPLS_BGEMM_WS=0: the __syncthreads form of the two GEMM kernels (kept for comparison); default: warp-specialised
static int bgemm_ws() {
  const char* e = getenv("PPLS_BGEMM_WS");
  return (e && *e) ? (atoi(e) != 0) : 1;
}
template <int R>
static cudaError_t launch_bprojmu_r(const ProjMuArgs& fa, const BGemmPlan& p, cudaStream_t stream) {
  const size_t smem = ((size_t)FSTAGES * FSTAGE + 4 * FT_N * 3 * R + 4 * FT_N + (size_t)(p.RP / FT_N) * FT_N * 3 * R + 2 * FSTAGES) * sizeof(double);
  cudaError_t e = cudaFuncSetAttribute((const void*)bprojmu_kernel<R>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  const size_t grid = (size_t)p.nchunk;            // = CTAs of this kernel (plan_bgemm): one Gram partial per CTA
  bprojmu_kernel<R><<<(unsigned)grid, WS_THREADS, smem, stream>>>(fa);
  return cudaGetLastError();
}
// PPLS_BPROJ_PP=1: two half-CTAs out of phase (bprojmu2_kernel); default: one 128-row tile per CTA at a time (bprojmu_kernel).
// Measured at C5: 307-318 us against 281 us -- the half-size tiles cost more (twice the loadings traffic and per-stage
// hand-offs per flop) than the overlapped epilogue returns, with or without the stagger, with 3 or 4 stages.
static int bproj_pingpong() {
  const char* e = getenv("PPLS_BPROJ_PP");
  return (e && *e) ? (atoi(e) != 0) : 0;
}
template <int R>
static cudaError_t launch_bprojmu2_r(const ProjMuArgs& fa, const BGemmPlan& p, cudaStream_t stream) {
  const size_t half = (size_t)HSTAGES * HSTAGE + 2 * FT_N * 3 * R + 4 * FT_N + (size_t)(p.RP / FT_N) * FT_N * 3 * R;
  const size_t smem = (2 * half + 4 * HSTAGES + 1) * sizeof(double);
  if (smem > 227 * 1024) return launch_bprojmu_r<R>(fa, p, stream);
  cudaError_t e = cudaFuncSetAttribute((const void*)bprojmu2_kernel<R>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  bprojmu2_kernel<R><<<(unsigned)p.nchunk, WS_THREADS, smem, stream>>>(fa);
  return cudaGetLastError();
}
static cudaError_t launch_bprojmu(const ProjMuArgs& fa, const BGemmPlan& p, cudaStream_t stream) {
  if (bproj_pingpong() && bgemm_ws())
    switch (p.R) {
      case 1: return launch_bprojmu2_r<1>(fa, p, stream);
      case 2: return launch_bprojmu2_r<2>(fa, p, stream);
      case 4: return launch_bprojmu2_r<4>(fa, p, stream);
      case 8: return launch_bprojmu2_r<8>(fa, p, stream);
      default: return cudaErrorInvalidValue;
    }
  switch (p.R) {
    case 1: return launch_bprojmu_r<1>(fa, p, stream);
    case 2: return launch_bprojmu_r<2>(fa, p, stream);
    case 4: return launch_bprojmu_r<4>(fa, p, stream);
    case 8: return launch_bprojmu_r<8>(fa, p, stream);
    default: return cudaErrorInvalidValue;
  }
}

cudaError_t launch_bgemm_pass(const BGemmPlan& p, const DataShape& s, const double* D, const double* Lpack, const double* coef,
                              double* ws, double* momvec_pack, cudaStream_t stream) {
  double* P = ws;
  double* partial = P + (size_t)s.n * 2 * p.RP;
  double* gpart = partial + (size_t)(p.nsplit > p.nsplit_mixed ? p.nsplit : p.nsplit_mixed) * p.ntiles * BT * BT;
  const char* es = getenv("PPLS_BACC_STAGES");
  const int nstg = (es && *es) ? atoi(es) : BACC_STAGES_DEFAULT;
  if (nstg < 2 || nstg > 4) return cudaErrorInvalidValue;
  const int wsp = bgemm_ws();
  const void* bacc_tab[3][2] = {{(const void*)baccum_kernel<2, false>, (const void*)baccum_kernel<2, true>},
                                {(const void*)baccum_kernel<3, false>, (const void*)baccum_kernel<3, true>},
                                {(const void*)baccum_kernel<4, false>, (const void*)baccum_kernel<4, true>}};
  const void* bacc = bacc_tab[nstg - 2][wsp];
  const size_t smemA = (size_t)2 * (BT * BAP + BK * BP) * sizeof(double);
  const size_t smemB = (size_t)nstg * 3 * BK * BP * sizeof(double) + (wsp ? 2 * nstg * 8 : 0);
  cudaError_t e = cudaFuncSetAttribute((const void*)bproj_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smemA);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(bacc, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smemB);
  if (e != cudaSuccess) return e;
  if (p.fused) {
    ProjMuArgs fa{};
    fa.D = D; fa.L = Lpack; fa.coef = coef; fa.P = P; fa.gpart = gpart; fa.n = s.n;
    fa.ld = s.ld; fa.p2 = s.p2; fa.q2 = s.q2; fa.RP = p.RP; fa.nstart = p.nstart; fa.rowtiles = p.rowtiles;
    { const char* ed = getenv("PPLS_BGEMM_DBG"); fa.dbg = (ed && *ed) ? atoi(ed) : 0; }
    e = launch_bprojmu(fa, p, stream);
    if (e != cudaSuccess) return e;
  } else {
  ProjArgs pa{};
  pa.D = D; pa.L = Lpack; pa.P = P; pa.n = s.n; pa.ld = s.ld; pa.p2 = s.p2; pa.q2 = s.q2; pa.RP = p.RP;
  bproj_kernel<<<dim3(p.rowtiles, p.ncb, 2), BTHREADS, smemA, stream>>>(pa);
  MuArgs m{};
  m.P = P; m.coef = coef; m.gpart = gpart; m.n = s.n; m.rows_per_chunk = p.rows_per_chunk; m.RP = p.RP; m.R = p.R; m.nstart = p.nstart;
  switch (p.R) {
    case 1: launch_bmu<1>(m, p, stream); break;
    case 2: launch_bmu<2>(m, p, stream); break;
    case 3: launch_bmu<3>(m, p, stream); break;
    case 4: launch_bmu<4>(m, p, stream); break;
    case 5: launch_bmu<5>(m, p, stream); break;
    case 6: launch_bmu<6>(m, p, stream); break;
    case 7: launch_bmu<7>(m, p, stream); break;
    case 8: launch_bmu<8>(m, p, stream); break;
    default: return cudaErrorInvalidValue;
  }
  }
  AccArgs aa{};
  aa.D = D; aa.MU = P; aa.partial = partial; aa.n = s.n; aa.rows_per_split = p.rows_per_split;
  aa.rows_per_split_mixed = p.rows_per_split_mixed; aa.ntiles = p.ntiles; aa.nsplit = p.nsplit; aa.nsplit_mixed = p.nsplit_mixed;
  aa.mixed_db = p.mixed_db;
  aa.ld = s.ld; aa.p2 = s.p2; aa.q2 = s.q2; aa.RP = p.RP; aa.nbx = p.nbx; aa.nby = p.nby; aa.ncb = p.ncb; aa.unified = p.unified;
  { const char* ed = getenv("PPLS_BGEMM_DBG"); aa.dbg = (ed && *ed) ? atoi(ed) : 0; }
  {
    void* kargs[] = {(void*)&aa};
    const int n_mixed = p.mixed_db >= 0 ? p.ncb : 0;
    const unsigned grid = (unsigned)((p.ntiles - n_mixed) * p.nsplit + n_mixed * p.nsplit_mixed);
    e = cudaLaunchKernel(bacc, dim3(grid), dim3(wsp ? WS_THREADS : BTHREADS), kargs, smemB, stream);
    if (e != cudaSuccess) return e;
  }
  FinArgs f{};
  f.partial = partial; f.gpart = gpart; f.momvec = momvec_pack; f.nsplit = p.nsplit; f.ntiles = p.ntiles; f.ncb = p.ncb;
  f.mixed_db = p.mixed_db; f.nsplit_mixed = p.nsplit_mixed;
  f.nbx = p.nbx; f.ld = s.ld; f.p2 = s.p2; f.RP = p.RP; f.R = p.R; f.nstart = p.nstart; f.nchunk = p.nchunk; f.unified = p.unified;
  const size_t ne = (size_t)s.ld * p.RP + (size_t)p.nstart * 3 * p.R * p.R;
  bfinal_kernel<<<(unsigned)((ne + 31) / 32), dim3(32, 8), 0, stream>>>(f);
  return cudaGetLastError();
}

}  // namespace ppls
