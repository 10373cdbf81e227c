// Internal declarations shared by the translation units of libppls_b200.so.
// Not part of the C ABI (that is include/ppls_b200.h).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/ppls_b200.h"

namespace ppls {

constexpr int MAXR = PPLS_MAX_R;

// ---------------------------------------------------------------------------------------
// HBM layout of a data shard (DESIGN.md "Data layout"):
//   D[n][ld] row-major FP64, ld = p2 + q2, p2 = p rounded up to even, q2 likewise.
//   Row i = [ x_i (p values) | 0-pad | y_i (q values) | 0-pad ].  A tile of T rows is one
//   contiguous T*ld*8-byte block -> one cp.async.bulk per tile, perfectly sequential DRAM.
// Loadings live in the same column space: L[ld][R] row-major (W rows then C rows).
// ---------------------------------------------------------------------------------------
struct DataShape {
  int64_t n;        // local rows
  int32_t p, q;     // logical columns
  int32_t p2, q2;   // padded (even)
  int32_t ld;       // p2 + q2
  // Wide shapes (ld > PANEL_LD_MIN) are stored PANEL-MAJOR instead: the padded column space is cut into
  // panels of PANEL_W columns (X panels, then Y panels; the last panel of each is zero-padded) and
  // D[panel][n][PANEL_PITCH] -- a CTA of the wide-row path then streams ONE contiguous block of n panel rows.  A stored
  // panel row is PANEL_PITCH = 128 + 4 doubles (4 zero pad values, +3.1 % bytes): one bulk copy lands a whole tile in
  // shared memory with a row pitch of 1056 bytes, which is what makes the FP64 tensor-core fragment loads of
  // em_wide2.cu conflict-free (8 rows x 32 bytes and 4 rows x 64 bytes both touch 32 distinct banks per half-warp).
  int32_t panel, npX, npY;
};
constexpr int PANEL_W = 128;
constexpr int PANEL_PITCH = PANEL_W + 4;
constexpr int PANEL_LD_MIN = 2304;
__host__ __device__ inline size_t data_index(const DataShape& s, int64_t i, int c) {
  if (!s.panel) return (size_t)i * s.ld + c;
  const int cy = c - s.p2;
  const int P = (c < s.p2) ? c / PANEL_W : s.npX + cy / PANEL_W;
  const int j = (c < s.p2) ? c % PANEL_W : cy % PANEL_W;
  return ((size_t)P * s.n + i) * PANEL_PITCH + j;
}
__host__ __device__ inline size_t data_elems(const DataShape& s) {
  return s.panel ? (size_t)(s.npX + s.npY) * s.n * PANEL_PITCH : (size_t)s.n * s.ld;
}

// Layout of the per-CTA partial vector and of the reduced vector ("moment vector"):
//   [ G (ld x R, row-major: G1 rows then G2 rows) | Gxx (R x R) | Gxy (R x R) | Gyy (R x R) ]
// Gxx[k][l] = sum_i Xw[i,k] Xw[i,l];  Gxy[k][l] = sum_i Xw[i,k] Yc[i,l];  Gyy likewise.
__host__ __device__ inline int64_t momvec_len(int ld, int R) { return (int64_t)ld * R + 3 * R * R; }

// Device-resident state of one EM run.  Written only by the update kernel (K2).
struct FitState {
  // parameters theta_k (loop order)
  double b[MAXR], t[MAXR];
  double sigE, sigF, sigH;
  // coefficients the streaming kernel needs for theta_k:
  //   mu_T[:,k] = aT[k] Xw[:,k] + bT[k] Yc[:,k];  mu_U[:,k] = aU[k] Xw[:,k] + bU[k] Yc[:,k]
  double aT[MAXR], bT[MAXR], aU[MAXR], bU[MAXR];
  // Gram matrices of the current loadings (exact algebraic residuals for any W, C)
  double WtW[MAXR * MAXR], CtC[MAXR * MAXR];
  // E-step second moments at the parameters of the last pass
  double Ctt[MAXR], Cuu[MAXR], Cut[MAXR], Cee, Cff, Chh[MAXR * MAXR];
  double loglik_last;
  // data constants
  double ssqX, ssqY;     // global ||X||_F^2, ||Y||_F^2
  double Ntot;           // global row count
  // control
  int32_t pass;          // index k of the NEXT pass to run
  int32_t done;          // 1 once the break rule or max_steps ended the run
  int32_t steps_done;    // EM steps completed (= length of the loglik trace)
  int32_t status;        // PPLS_OK or PPLS_E_NONFINITE
  int32_t max_steps, orth_type, R, pad_;
  double atol;
  // Gram (xx, xy, yy) of the pass that ENDED a sequential fit: the deflation of the next component reads it from
  // here (momvec is rewritten / re-reduced by the no-op passes the host may still have queued)
  double gram_last[3];
  // host-mapped (pinned, zero-copy) word the device sets when it raises `done`, so that the host stops
  // queueing passes without a stream synchronisation; 0 = none
  int32_t* done_host;
};
// every kernel raises `done` through this
__device__ __forceinline__ void raise_done(FitState* st) {
  st->done = 1;
  if (st->done_host) { *reinterpret_cast<volatile int32_t*>(st->done_host) = 1; __threadfence_system(); }
}

// ---- launchers (host functions defined in the .cu files) -------------------------------
struct StreamPlan {          // how K1 is launched for one shape
  int R, M, T, nstage;
  int keep;                  // KEEP form: the tile is read from shared memory once, x stays in registers (em_stream_kernel.cuh)
  int nwX, nwY;              // warps owning X column chunks / Y column chunks
  int cX, bigX, cY, bigY;    // chunks per warp: the first bigX X-warps own cX+1, the rest cX (Y alike)
  int threads, grid, ctas_per_sm;
  size_t smem_bytes;
  int64_t ntiles;
};

struct WidePlan {            // two-pass wide-row path (em_wide.cu)
  int R, NCH, nstage;
  int scX, scY, nslabX, nslabY;      // slab widths (columns) and counts
  int nrg, rows_per_rg, nblk_mu;
  int slots;                         // CTAs resident at once = grid of the persistent wide kernel (0: one CTA per work item)
  int v2, nstage2;                   // DMMA form (em_wide2.cu; panel-major storage): nrg / rows_per_rg are whole 64-row tiles
  int sm_count;
  // single-pass, L2-pipelined form (em_wide3.cu): lanes (row groups with a pipeline each) x CTAs per lane
  int v3, v3_lanes, v3_cpl, v3_lag, v3_nstage, v3_ns, v3_chunks_per_lane, v3_nchunk;
};
struct WideOffsets { size_t ppart, proj, mu, gpart, gram, mu8, cnt; };   // offsets (doubles) into the wide workspace
cudaError_t plan_wide(const DataShape& s, int R, int sm_count, WidePlan* w, const char** why);
int64_t wide_npad(int64_t n);
// em_wide3.cu: ONE pass from HBM (persistent cooperative kernel, second pass out of the L2)
bool plan_wide3(const DataShape& s, int R, int sm_count, WidePlan* w);
size_t wide3_ring_doubles(const DataShape& s, const WidePlan& w);
size_t wide3_counter_doubles(const WidePlan& w);
cudaError_t launch_wide3(const WidePlan& w, const DataShape& s, const double* D, const double* L, const FitState* st, double* pring,
                         double* mu, double* mu8, double* gpart, double* gram_part, double* counters, cudaStream_t stream);
// the whole iteration of the wide path when the single-pass form is planned (kernel + fixed-order sums -> momvec)
cudaError_t launch_wide_single(const WidePlan& w, const DataShape& s, const double* D, const double* L, const FitState* st,
                               double* ws, const WideOffsets& o, double* momvec, cudaStream_t stream);
// em_wide2.cu: the two passes on the FP64 tensor pipe (panel-major storage)
void plan_wide2(const DataShape& s, int sm_count, WidePlan* w);
cudaError_t launch_wide2(int mode, const WidePlan& w, const DataShape& s, const double* D, const double* L, const FitState* st,
                         double* ppart, const double* mu8, double* gpart, int sm_count, cudaStream_t stream);
size_t wide_workspace_doubles(const DataShape& s, const WidePlan& w, WideOffsets* o);
cudaError_t launch_wide_project(const WidePlan& w, const DataShape& s, const double* D, const double* L,
                                const FitState* st, double* ws, const WideOffsets& o, cudaStream_t stream);
cudaError_t launch_wide_accum(const WidePlan& w, const DataShape& s, const double* D, const double* L,
                              const FitState* st, double* ws, const WideOffsets& o, double* momvec,
                              cudaStream_t stream);

cudaError_t plan_stream(const DataShape& s, int R, int sm_count, StreamPlan* plan, const char** why);
// One data pass: per-CTA partials -> `partials` [grid][momvec_len].
struct XchgDev;
// fused tail (em_tail.cuh): with `ft` the launch also reduces the partials, exchanges them (xchg != null) and runs
// the whole update of the iteration -- one launch per EM iteration
struct FusedTail {
  double* momvec; double* L; FitState* st; double* trace; double* gscratch;
  const XchgDev* xchg;
};
bool stream_tail_fits(const StreamPlan& plan, const DataShape& s);
size_t stream_tail_scratch_doubles(const StreamPlan& plan);
size_t stream_tail_clock_offset(const StreamPlan& plan);      // doubles; then u64 x 3: streaming ns, tail ns, launches
cudaError_t launch_em_stream(const StreamPlan& plan, const DataShape& s, const double* D,
                             const double* L, const FitState* st, double* partials,
                             double* muT, double* muU, int write_mu, cudaStream_t stream, const FusedTail* ft = nullptr);
// Sum the per-CTA partials in CTA order -> momvec.
// One-shot peer exchange of the moment vector (multi-GPU, one node): every rank stores its reduced vector
// straight into every peer's inbox over NVLink (peer-mapped cudaIpc memory), raises a flag there, and the
// update kernel of each rank sums the world's inbox slots in rank order (bitwise identical on all ranks).
constexpr int XCHG_MAXW = 16;
constexpr int XCHG_MAXCTA = 160;          // per-CTA flags of the fused tail: one word per (parity, source rank, CTA)
// exchange block of a rank: [control words 1 KB | per-CTA flags | inbox]
constexpr size_t XCHG_CTRL_BYTES = 1024;
constexpr size_t XCHG_CFLAG_BYTES = (size_t)2 * XCHG_MAXW * XCHG_MAXCTA * 8;
constexpr size_t XCHG_INBOX_OFFSET = XCHG_CTRL_BYTES + XCHG_CFLAG_BYTES;
struct XchgDev {
  double* inbox[XCHG_MAXW];               // rank r's inbox [2 parities][world][slot] (own rank: local pointer)
  unsigned long long* flags[XCHG_MAXW];   // rank r's flag words [2 parities][XCHG_MAXW]: epoch of the last arrival
  unsigned long long* cflags[XCHG_MAXW];  // rank r's per-CTA flag words [2 parities][XCHG_MAXW][XCHG_MAXCTA] (fused tail)
  unsigned long long* epoch;              // local: exchanges completed
  unsigned int* counter;                  // local: send blocks finished (threadfence-reduction pattern)
  int* failed;                            // local: a wait timed out (a peer died)
  int rank, world;
  long long slot;                         // doubles per (parity, source rank) slot
  unsigned long long timeout_ns;          // a wait longer than this marks the exchange failed (20 s; PPLS_XCHG_TIMEOUT_MS)
};
// reduce the per-CTA partial vectors (as launch_reduce_partials) and send the result to every rank's inbox
cudaError_t launch_xsend(const double* partials, int nparts, int64_t len, const XchgDev& x, const FitState* st,
                         cudaStream_t stream);
cudaError_t launch_reduce_partials(const double* partials, int nparts, int64_t len, double* momvec,
                                   const FitState* st, cudaStream_t stream);
// K2: loglik, E-step moments, break rule, M-step, next coefficients.
// orth_type SVD + gscratch != null: the grid (cooperative) kernel; else the single-CTA kernel.
// xchg != null: momvec is first assembled from the peer-exchange inbox (wait for the world's flags, sum).
cudaError_t launch_em_update(const DataShape& s, int R, const double* momvec, double* L, double* Lscratch,
                             FitState* st, double* trace, int orth_type, double* gscratch, cudaStream_t stream,
                             const XchgDev* xchg = nullptr);
// batched multi-start: S starts, g per pack of RP = g*R components (em_update.cu)
cudaError_t launch_em_update_batch(const DataShape& s, int R, int RP, int g, int S, int tracelen, int prepare,
                                   const double* momvec_pack, double* L_pack, FitState* pack_st, FitState* st,
                                   double* momvec_s, double* L_s, double* Lscratch_s, double* trace, cudaStream_t stream,
                                   double* coef = nullptr);
int update_grid_blocks(int ld);
// batched multi-start as two FP64 GEMMs per pass (em_batch.cu)
struct BGemmPlan {
  int RP, R, nstart;            // components of the group (multiple of 128), per start, starts in the group
  int ncb, nbx, nby, ntiles, rowtiles, unified;
  int nsplit; int64_t rows_per_split;
  // the column block that straddles the X|Y boundary costs more per row (it stages mu_T and mu_U and multiplies the tile pair
  // on the boundary twice): it gets MORE, shorter row splits so that all CTAs of the single wave end together
  int mixed_db; int nsplit_mixed; int64_t rows_per_split_mixed;       // mixed_db < 0: no such block / even splits
  int nchunk; int64_t rows_per_chunk;
  int fused;                    // projection + mu + Gram blocks in one kernel (R divides 8): nchunk = rowtiles
  int sm_count;
};
void plan_bgemm(const DataShape& s, int RP, int R, int nstart, int sm_count, BGemmPlan* p);
size_t bgemm_workspace_doubles(const DataShape& s, const BGemmPlan& p);
// one data pass for a group: Lpack [ld][RP], coef [4][RP] -> momvec_pack [ld*RP + 3*RP*RP] (block diagonal of the Grams)
cudaError_t launch_bgemm_pass(const BGemmPlan& p, const DataShape& s, const double* D, const double* Lpack, const double* coef,
                              double* ws, double* momvec_pack, cudaStream_t stream);
// multi-population fit (em_seq.cu): J populations, momvecs [J][ld+3], sts [J]
cudaError_t launch_meta_update(const DataShape& s, int J, const double* momvecs, double* L, FitState* sts, double* trace,
                               int prepare, cudaStream_t stream);
cudaError_t launch_orth_polar(int R, const double* G, int n, double* out, cudaStream_t stream);
cudaError_t launch_orth(int R, const double* G, int n, double* out, int orth_type, double* work, cudaStream_t stream);
// the same kernel with its scratch in a static device array (no allocation): device pointers, n * R <= orth_static_max_elems()
cudaError_t launch_orth_static(int R, const double* G, int n, double* out, int orth_type, cudaStream_t stream);
long long orth_static_max_elems();

// ---- standard errors of the loadings (em_var.cu; variances.PPLS_simult, EM_W_multi.R:830-860) -----------
struct VarPlan {
  int w;              // columns of the side (p or q)
  int c0;             // first column of the side in a data row (0 or p2)
  int nb, ntiles;     // 128-column blocks, block pairs bi <= bj
  int nsplit;         // row splits of the Gram kernel
  int64_t rows_per_split;
  int gsplit;         // row splits of the X'mu kernel
  int64_t grows_per_split;
};
void plan_var(const DataShape& s, int which, int sm_count, VarPlan* plan);
size_t var_workspace_doubles(const VarPlan& p, int R);
// S = side' side (w x w, row-major, both triangles), summed over this rank's rows
cudaError_t launch_var_gram(const VarPlan& p, const DataShape& s, const double* D, double* partial, double* S, cudaStream_t stream);
// G = side' mu (w x R row-major), cp[k] = sum_i mu[i][k]^2; mu: [n][R] row-major on the device
cudaError_t launch_var_xtmu(const VarPlan& p, const DataShape& s, const double* D, const double* mu, int R, double* partial,
                            double* G, double* cp, cudaStream_t stream);
// per component: A = B_exp I - SSt_exp into the augmented matrix [A | I] (w x 2w), optional copies of SSt_exp, SSt_star
cudaError_t launch_var_assemble(int w, int R, int comp, const double* S, const double* G, const double* Wpol, double CttN,
                                double cp, double sigE, double Ntot, double* aug, double* SSt_exp, double* SSt_star,
                                cudaStream_t stream);
// in-place Gauss-Jordan with partial pivoting on [A | I]; then var = -inv(A) (column-major w x w, optional), se = sqrt(diag)
cudaError_t launch_var_invert(int w, double* aug, double* rowbuf, int* info, double* var_out, double* se_out, cudaStream_t stream);
size_t update_grid_scratch_doubles(int ld, int R);
// Prepare coefficients + Gram matrices for the start state (pass 0).
cudaError_t launch_em_prepare(const DataShape& s, int R, const double* L, FitState* st, cudaStream_t stream);

// em_seq.cu: update of the sequential one-component fit (R = 1; `prev` = earlier components' loadings)
// Constraints of PPLSi (fconstraint, EM_W_multi.R:85-92, applied at :142-146 and :160-167): fixed values that replace the
// estimates after every EM step; critfunc (:172) = identity or abs.
struct SeqCons {
  uint32_t mask = 0;              // bit 0 W, 1 C, 2 B, 3 sigE, 4 sigF, 5 sigH, 6 sigT
  int32_t crit_abs = 0;           // critfunc = abs
  const double* L = nullptr;      // device [ld]: fixed loadings (W rows then C rows) where the mask says so
  double B = 0, sigE = 0, sigF = 0, sigH = 0, sigT = 0;
};
cudaError_t launch_seq_update(const DataShape& s, const double* momvec, double* L, FitState* st, double* trace,
                              const double* prev, int nprev, int prepare, cudaStream_t stream, const SeqCons* cons = nullptr);

// data.cu
cudaError_t launch_retile(const double* colmajor, int64_t ldc, int64_t nrows, int ncols,
                          double* D, const DataShape& s, int col0, int64_t row0, cudaStream_t stream);
cudaError_t launch_untile(const double* D, const DataShape& s, int col0, int64_t row0, int64_t nrows, int ncols,
                          double* colmajor, int64_t ldc, cudaStream_t stream);
cudaError_t launch_zero_pad(double* D, const DataShape& s, cudaStream_t stream);
cudaError_t launch_ssq(const double* D, const DataShape& s, double* out2, double* scratch, int nblocks,
                       cudaStream_t stream);
// orthogonal parts of simulC (loglC.cpp:269-315): X += TYo P_Yosc', Y += UXo P_Xosc'; device pointers
struct SynthOrth {
  int nx = 0, ny = 0;
  const double* PYo = nullptr;    // [p][nx] row-major
  const double* PXo = nullptr;    // [q][ny] row-major
  const double* sigTo = nullptr;  // [nx]
  const double* sigUo = nullptr;  // [ny]
};
cudaError_t launch_synth(double* D, const DataShape& s, int64_t row0, uint32_t colX0, uint32_t colY0, int R, uint64_t seed,
                         const double* Ltruth /*[ld][R]*/, const double* b, const double* t,
                         double sigE, double sigF, double sigH, cudaStream_t stream, const SynthOrth* orth = nullptr);

// initialGuess = "random" (EM_W_multi.R:133) on the device generator: W [p], C [q], sc = {B, sigE, sigF, sigH, sigT}
cudaError_t launch_guess(double* W, int p, double* C, int q, double* sc, uint64_t seed, uint32_t comp, cudaStream_t stream);
// probe.cu: measured ceilings for bench.py's roofline lines
cudaError_t probe_read(const void* x, size_t bytes, int sm_count, int launches, double* scratch, double* gbs, cudaStream_t stream);
cudaError_t probe_dmma(int sm_count, double* scratch, double* tflops, cudaStream_t stream);

}  // namespace ppls
