/* ppls_b200.h -- C ABI of the B200-native engine for the PPLS_simult EM loop.
 *
 * Drop-in boundary for ONE path of selbouhaddani/PPLS: the loop body of
 *   PPLS_simult(X, Y, a, EMsteps, atol, type, ...)   Package/PPLS/R/EM_W_multi.R:758-807
 * i.e. lines :780-802 (Expect_M :637-717 fast branch, Maximiz_M :729-742, logl_W :297-323
 * and the only native call on that path, loglC_fast, Package/PPLS/src/loglC.cpp:319-338,
 * bound today through .Call('PPLS_loglC_fast', ...) Package/PPLS/R/RcppExports.R:32-34 /
 * RcppExport SEXP PPLS_loglC_fast(...) Package/PPLS/src/RcppExports.cpp:79-99).
 *
 * Because E- and M-step are R code in the reference, the seam is the loop, not one symbol:
 * an Rcpp shim (r/ppls_b200_shim.cpp) marshals R matrices into the
 * calls below.  Plain C types only: no C++/R/torch types, no exceptions cross this line.
 *
 * Conventions
 *  - All matrices passed by the caller are host memory, column-major FP64 (R / Eigen
 *    layout), caller-owned, read-only unless documented as output.  Outputs are
 *    caller-allocated.
 *  - Every function returns PPLS_OK (0) or a negative PPLS_E_* code; the message of the
 *    last failure on this thread is ppls_last_error().
 *  - A context is used by one host thread at a time (R's C API is single-threaded).
 *  - There is no CPU fallback: without a CUDA device every compute entry point fails
 *    with PPLS_E_CUDA.
 */
#ifndef PPLS_B200_H
#define PPLS_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PPLS_OK              0
#define PPLS_E_INVALID      -1   /* bad argument / shape (R side: stopifnot) */
#define PPLS_E_CUDA         -2   /* CUDA runtime failure, or no device        */
#define PPLS_E_NCCL         -3   /* NCCL failure / libnccl.so.2 not loadable  */
#define PPLS_E_OOM          -4   /* device or host allocation failed          */
#define PPLS_E_NONFINITE    -5   /* a parameter became NaN/Inf during EM      */
#define PPLS_E_UNSUPPORTED  -6   /* shape outside what the kernels cover      */
#define PPLS_E_DEGENERATE   -7   /* a noise s.d. fell below 100*eps (EM_W_multi.R:153) */

#define PPLS_MAX_R           8   /* nr_comp supported by the streaming kernel */
#define PPLS_NCCL_ID_BYTES 128

#define PPLS_ORTH_SVD 0          /* type = "SVD" (default, EM_W_multi.R:758,761) */
#define PPLS_ORTH_QR  1          /* type = "QR"                                  */

#define PPLS_F_TIMING      1u    /* record CUDA events around every streaming launch */
#define PPLS_F_DIRECT_RES  2u    /* (reserved) direct residual sums, debug mode      */

typedef struct ppls_ctx  ppls_ctx;   /* device, stream, NCCL communicator, workspaces */
typedef struct ppls_data ppls_data;  /* device-resident row shard of [X Y]            */
typedef struct ppls_fit  ppls_fit;   /* state of one EM run (parameters, trace)       */

/* Arguments of PPLS_simult that reach the loop (EM_W_multi.R:758). */
typedef struct {
  int32_t  r;            /* a / nr_comp                                            */
  int32_t  max_steps;    /* EMsteps                                                */
  int32_t  orth_type;    /* PPLS_ORTH_SVD | PPLS_ORTH_QR                           */
  int32_t  check_every;  /* passes enqueued between host convergence checks; 0 = default */
  double   atol;         /* break when loglik[i]-loglik[i-1] < atol (signed, :792) */
  uint32_t flags;        /* PPLS_F_*                                               */
  uint32_t reserved;
} ppls_opts;

/* The state (W., C., B., sigE., sigF., sigH., sigT.) of EM_W_multi.R:765-771 / :783-789.
 * B and sigT are diagonal in the model; only the diagonals cross the ABI. */
typedef struct {
  double* W;      /* p x r, column-major */
  double* C;      /* q x r, column-major */
  double* B;      /* r  (diag)           */
  double* sigT;   /* r  (diag)           */
  double  sigE, sigF, sigH;
} ppls_params;

/* Device-time accounting of one fit (PPLS_F_TIMING). */
typedef struct {
  double  stream_ms;      /* sum of streaming-kernel (K1) launch durations       */
  int64_t stream_launches;
  double  loop_ms;        /* first-to-last event of all iterate() calls          */
  int64_t kernel_launches;/* every kernel of ours launched by this fit           */
  int64_t passes;         /* data passes done                                    */
  int64_t bytes_per_pass; /* algorithmic bytes of one pass on this rank: 8*N*(p+q) */
  double  tail_ms;        /* sum over passes of what follows K1: reduce, exchange, update (K2) */
  int64_t peer_exchange;  /* 1: the moment vector travels by the one-shot peer exchange, 0: NCCL / single GPU */
  double  stream_phase_ms;/* fused launches (one kernel per EM iteration): device-clock time of the streaming phase
                             alone, summed; tail_ms is then the device-clock time of reduction + exchange + update */
  double  scalar_ms;      /* fused launches: of tail_ms, the time until CTA 0 has finished the scalar part (reduction of its
                             elements + exchange + log-likelihood, E-step moments, break rule, M-step scalars) */
} ppls_timing;

const char* ppls_version(void);
const char* ppls_last_error(void);

/* ---- context ------------------------------------------------------------------------- */
int ppls_ctx_create(ppls_ctx** out, int device);
/* One rank of a row-sharded multi-GPU run (one process per GPU).  `nccl_id` is the
 * PPLS_NCCL_ID_BYTES blob made by ppls_nccl_unique_id() on rank 0 and handed to the other
 * ranks by the caller (torch.distributed, MPI, a file ...). */
int ppls_nccl_unique_id(void* id_out, size_t nbytes);
int ppls_ctx_create_dist(ppls_ctx** out, int device, int rank, int world,
                         const void* nccl_id, size_t nbytes);
/* ONE process, SEVERAL GPUs of one node (what a single R session needs): the context drives the listed
 * devices from the calling thread -- one stream per device, direct peer access, the same one-shot peer exchange of the
 * moment vector as the one-process-per-GPU form, no NCCL.  ppls_data_from_host / ppls_data_synthetic then split the ROWS
 * over the devices (shard i = rows [N i/G, N (i+1)/G), uploaded concurrently), ppls_fit_* / ppls_simult_run /
 * ppls_loglik / ppls_data_export / ppls_data_info work on the whole; results are bitwise those of the
 * one-process-per-GPU form.  Every device must reach every other by peer access (NVLink / NVSwitch).  The other entry
 * points (sequential, batched, variances, meta, column-sharded) return PPLS_E_UNSUPPORTED on such a context; so do
 * shapes that take the wide-row path.  n_devices == 1 is ppls_ctx_create. */
int ppls_ctx_create_multi(ppls_ctx** out, const int* device_ids, int n_devices);
int ppls_ctx_destroy(ppls_ctx* ctx);
int ppls_ctx_info(const ppls_ctx* ctx, int* device, int* rank, int* world, int* sm_count);
/* Synchronise the context's stream (for callers that time with their own clock). */
int ppls_ctx_sync(ppls_ctx* ctx);

/* Measured ceilings of the context's GPU, for roofline reporting (no counterpart in the reference; bench.py's
 * `roofline` object): which = PPLS_PROBE_HBM_READ -> GB/s of a read-only stream over the resident dataset `d`
 * (or over a scratch buffer of up to 8 GiB when d is NULL); PPLS_PROBE_FP64_DMMA -> TFLOP/s of back-to-back FP64
 * tensor instructions (mma.sync m8n8k4 f64) from registers, 8 warps per SM. */
#define PPLS_PROBE_HBM_READ  0
#define PPLS_PROBE_FP64_DMMA 1
int ppls_probe(ppls_ctx* ctx, const ppls_data* d, int32_t which, double* value);

/* ---- data ---------------------------------------------------------------------------- */
/* Copy this rank's rows of X (n x p, leading dimension ldX) and Y (n x q, ldY) to the
 * device and re-tile them into the engine's row-major layout.  Replaces the by-value
 * Eigen::MatrixXd copies of RcppExports.cpp:87-88,96 -- done once, not once per iteration. */
int ppls_data_from_host(ppls_ctx* ctx, const double* X, int64_t ldX, const double* Y, int64_t ldY,
                        int64_t n, int64_t p, int64_t q, ppls_data** out);
/* Generate rows row0 .. row0+n-1 of the synthetic PPLS sample on the device (model of
 * simulC, loglC.cpp:288-311, orthogonal parts zero; counter-based, so identical bits for
 * any sharding).  `truth` holds the generating parameters with r components. */
int ppls_data_synthetic(ppls_ctx* ctx, int64_t n, int64_t row0, int64_t p, int64_t q, int32_t r,
                        uint64_t seed, const ppls_params* truth, ppls_data** out);
/* simulC(N, W, C, P_Yosc, P_Xosc, B_T, sigX, sigY, sigH, sigT, sigTo, sigUo) of Package/PPLS/src/loglC.cpp:269-315
 * (R/RcppExports.R:12-30) as a resident dataset: X = T W' + TYo P_Yosc' + E, Y = U C' + UXo P_Xosc' + F with
 * T = Z diag(sigT), U = T diag(B) + H, TYo = Z diag(sigTo), UXo = Z diag(sigUo); r >= 1 components (the reference's
 * code type-checks for one).  P_Yosc p x nx, P_Xosc q x ny column-major (nx, ny in 0..8); normals from the
 * counter-based generator (seed, global row, column), so any row range [row0, row0+n) gives the same sample. */
int ppls_data_simulc(ppls_ctx* ctx, int64_t n, int64_t row0, int64_t p, int64_t q, int32_t r, int32_t nx, int32_t ny,
                     uint64_t seed, const ppls_params* truth, const double* P_Yosc, const double* P_Xosc,
                     const double* sigTo, const double* sigUo, ppls_data** out);

/* Column-sharded variants (wide shapes, p >> N): every rank holds ALL n rows of a block of
 * p_local columns of X and q_local columns of Y (blocks in rank order).  Start values / estimates passed
 * to ppls_fit_* are then the rank's ROWS of W (p_local x r) and C (q_local x r); per EM iteration the
 * ranks exchange the n x 2r projection sums and the (p+q) x r moment rows.  colX0 / colY0: global index
 * of the block's first column (counter of the generator); truth_local: the block's rows of W, C. */
int ppls_data_from_host_cols(ppls_ctx* ctx, const double* X, int64_t ldX, const double* Y, int64_t ldY,
                             int64_t n, int64_t p_local, int64_t q_local, ppls_data** out);
int ppls_data_synthetic_cols(ppls_ctx* ctx, int64_t n, int64_t p_local, int64_t q_local, int64_t colX0,
                             int64_t colY0, int32_t r, uint64_t seed, const ppls_params* truth_local,
                             ppls_data** out);
/* Copy rows row0 .. row0+n-1 (local numbering) back to column-major host arrays. */
int ppls_data_export(ppls_ctx* ctx, const ppls_data* d, int64_t row0, int64_t n,
                     double* X, int64_t ldX, double* Y, int64_t ldY);
/* Local shape and the (global, all ranks) sums of squares ||X||_F^2, ||Y||_F^2. */
int ppls_data_info(const ppls_data* d, int64_t* n, int64_t* p, int64_t* q, int64_t* n_total,
                   double* ssqX, double* ssqY);
int ppls_data_destroy(ppls_data* d);

/* ---- one EM run, stepwise -------------------------------------------------------------- */
/* Enter the loop at EM_W_multi.R:780 with an explicit start state (the reference's random
 * start, :762-778, stays on the caller's side). */
int ppls_fit_create(ppls_ctx* ctx, const ppls_data* d, const ppls_opts* opts,
                    const ppls_params* init, ppls_fit** out);
/* Enqueue up to `n_passes` data passes.  Pass k (k = 0, 1, ...) streams X and Y once with
 * the parameters theta_k and yields loglik(theta_k) (= loglik[k], k >= 1; logl_W at :791)
 * and the E-step at theta_k (:782); unless the break rule (:792) or k == max_steps ends
 * the run, the M-step then makes theta_{k+1}.  Returns the number of passes done so far
 * and whether the run has ended. */
int ppls_fit_iterate(ppls_fit* fit, int32_t n_passes, int32_t* passes_done, int32_t* finished);
/* Final state in LOOP order (the sign/order normalisation of :794-799 is the caller's,
 * on these tiny outputs), loglik[0 .. steps_done-1], the second moments of the final
 * E-step (:802) as Ctt[r], Cuu[r], Cut[r], Cee, Cff, Chh[r*r column-major]
 * (3r + 2 + r*r doubles), and mu_T / mu_U (local n x r, column-major; NULL to skip). */
int ppls_fit_result(ppls_fit* fit, ppls_params* est, double* loglik, int32_t* steps_done,
                    double* moments, double* mu_T, double* mu_U);
int ppls_fit_timing(const ppls_fit* fit, ppls_timing* out);
int ppls_fit_destroy(ppls_fit* fit);

/* ---- one EM run, one call (what the Rcpp shim binds) ----------------------------------- */
/* Equivalent to fit_create + fit_iterate until finished + fit_result + fit_destroy:
 * EM_W_multi.R:780-802 from the start state `init`. */
int ppls_simult_run(ppls_ctx* ctx, const ppls_data* d, const ppls_opts* opts,
                    const ppls_params* init, ppls_params* est, double* loglik,
                    int32_t* steps_done, double* moments, double* mu_T, double* mu_U);

/* ---- sequential fit: the start-value path of PPLS_simult (EM_W_multi.R:762) ------------------------ */
/* PPLS(X, Y, nr_comp, EMsteps, atol, initialGuess, customGuess) of EM_W_multi.R:229-279: components one at
 * a time (PPLSi :116-180, EMstep_W :51-74, EMstepC_fast loglC.cpp:341-397), each on the data deflated by the
 * earlier ones (:270-271, applied implicitly on the device).  critfunc = identity, no constraints.
 * guesses: nr_comp start states with ONE component each (W p x 1, C q x 1, B[1], sigT[1], sigE, sigF, sigH);
 * est: W p x nr_comp, C q x nr_comp, B[nr_comp], sigT[nr_comp] (+ sigE, sigF, sigH of the last component);
 * sig: nr_comp x 4 column-major (sigX, sigY, sigH, sigT per component, the reference's `sig` matrix);
 * logvalues: nr_comp x (EMsteps+1), row i = logvalue[0 .. steps[i]] (may be NULL); last_incr may be NULL;
 * comps_done < nr_comp when a noise s.d. vanished (:153-155, :257-262: the earlier components are kept). */
int ppls_sequential_run(ppls_ctx* ctx, const ppls_data* d, int32_t nr_comp, int32_t EMsteps, double atol,
                        const ppls_params* guesses, ppls_params* est, double* sig, double* logvalues,
                        int32_t* steps, double* last_incr, int32_t* comps_done);

/* ... with the `constraints` and `critfunc` arguments of PPLS / PPLSi (EM_W_multi.R:116-117, :229-230).  constraints: NULL or
 * nr_comp entries (fconstraint, :85-92): the values named in `mask` replace the start value (:142-146) and the estimate after
 * every EM step (:160-167).  Fixed loadings must have unit norm and be orthogonal to the earlier components (the deflation of
 * :270-271 is applied implicitly here; PPLS_E_UNSUPPORTED otherwise).  crit_abs != 0: critfunc = abs, else the identity
 * (:172: break when critfunc(logvalue[i+1] - logvalue[i]) < atol). */
#define PPLS_FIX_W    1u
#define PPLS_FIX_C    2u
#define PPLS_FIX_B    4u
#define PPLS_FIX_SIGE 8u
#define PPLS_FIX_SIGF 16u
#define PPLS_FIX_SIGH 32u
#define PPLS_FIX_SIGT 64u
typedef struct {
  uint32_t mask;        /* PPLS_FIX_* */
  uint32_t reserved;
  const double* W;      /* p values (PPLS_FIX_W) */
  const double* C;      /* q values (PPLS_FIX_C) */
  double B, sigE, sigF, sigH, sigT;
} ppls_constraints;
int ppls_sequential_run_ex(ppls_ctx* ctx, const ppls_data* d, int32_t nr_comp, int32_t EMsteps, double atol,
                           const ppls_params* guesses, const ppls_constraints* constraints, int32_t crit_abs,
                           ppls_params* est, double* sig, double* logvalues, int32_t* steps, double* last_incr,
                           int32_t* comps_done);

/* Start state of ONE component as PPLSi draws it (EM_W_multi.R:126-141).  kind PPLS_GUESS_EQUAL: orth(rep(1,p)), orth(rep(1,q)),
 * B = 1, sigH = sigT = 1, sigE = 1/p, sigF = 1/q (:135-136).  kind PPLS_GUESS_RANDOM (:133): orth(runif(p)), orth(runif(q)),
 * B = rchisq(1,1), (sigH, sigT) = rchisq(2,100)/100, (sigE, sigF) = rchisq(2,10)/100 -- drawn ON THE DEVICE from the counter-based
 * generator keyed by (seed, component): the distribution of the reference's start, repeatable bit for bit; R's own RNG stream
 * cannot be reproduced outside R.  out: W p x 1, C q x 1, B[1], sigT[1] caller-allocated.  ("o2m" needs OmicsPLS::o2m: pass its
 * result as a custom guess.) */
#define PPLS_GUESS_EQUAL  0
#define PPLS_GUESS_RANDOM 1
int ppls_initial_guess(ppls_ctx* ctx, int32_t kind, int64_t p, int64_t q, uint64_t seed, int32_t component, ppls_params* out);

/* logl_W(X, Y, W, C, B, sigE, sigF, sigH, sigT) of EM_W_multi.R:296-326 (-> loglC_fast, loglC.cpp:318-338)
 * for r components: one streaming pass over the resident data. */
int ppls_loglik(ppls_ctx* ctx, const ppls_data* d, int32_t r, const ppls_params* theta, double* out);

/* ---- orth(G, type): OmicsPLS::orth as the M-step calls it (EM_W_multi.R:732-733; historical in-tree copy
 * Package/functions.R:252-260).  G: n x r column-major host matrix (r <= n); type PPLS_ORTH_SVD: the polar factor
 * U V' of the thin SVD; PPLS_ORTH_QR: qr.Q(qr(G)) times sign(e[,1]' G[,1]).  out: n x r column-major host. ---- */
int ppls_orth(ppls_ctx* ctx, const double* G, int64_t ldG, int64_t n, int32_t r, int32_t orth_type,
              double* out, int64_t ldOut);

/* orth(G, type) WITHOUT ANY ALLOCATION.  No counterpart in the reference: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * The same kernel as ppls_orth; the working copy of G that the QR form needs lives in a static device array of
 * PPLS_ORTH_NOALLOC_MAX_ELEMS doubles (512 KB, held by every GPU that loads the library until exit) instead of a
 * cudaMalloc'ed buffer.  G_dev, out_dev: DEVICE pointers on the context's device, n x r in the engine's ROW-major
 * layout (element (j, k) at [j * r + k]); they must not overlap.  n * r > PPLS_ORTH_NOALLOC_MAX_ELEMS -> PPLS_E_UNSUPPORTED
 * with a message that names the limit.  The call is enqueued on the context's stream and has completed on return; it
 * calls no cudaMalloc, cudaFree or allocating library routine.  All calls share the one scratch array (it keeps the last
 * call's contents; nothing in it needs to start at zero): calls that overlap on the GPU -- other contexts, other host
 * threads on the same device -- race. */
#define PPLS_ORTH_NOALLOC_MAX_ELEMS 65536
int ppls_orth_noalloc(ppls_ctx* ctx, const double* G_dev, int64_t n, int32_t r, int32_t orth_type, double* out_dev);

/* ---- standard errors of the loadings: variances.PPLS_simult(fit, data, XorY) (EM_W_multi.R:830-860) ---- */
/* which = 0: data = X, mu = fit$Expectations$mu_T (n x r column-major, host, this rank's rows), Cdiag =
 * diag(fit$Expectations$Ctt); which = 1: data = Y, mu = mu_U, Cdiag = diag(Cuu).  sigE = fit$estimates$sigE
 * for both sides, as the reference has it.  Outputs (host, column-major; w = p or q): seLoad w x r;
 * varMatrix, SSt_exp, SSt_star r blocks of w x w (each may be NULL); B_exp r scalars (B_exp = scalar * I, may
 * be NULL); W = orth(t(data) %*% mu) w x r (may be NULL). */
int ppls_variances(ppls_ctx* ctx, const ppls_data* d, int32_t which, int32_t r, const double* mu, int64_t ldmu,
                   const double* Cdiag, double sigE, double* seLoad, double* varMatrix, double* SSt_exp,
                   double* SSt_star, double* B_exp, double* W);

/* ---- multi-population ("meta") fit: meta_PPLSi(X, Y, Ipopu, EMsteps, atol, initialGuess, customGuess) ---------- */
/* EM_W_multi.R:509-589 (meta_EMstep :446-485; meta_Estep / meta_Mstep, loglC.cpp:399-474): one component, loadings
 * shared by all populations, B_T / sigX / sigY / sigH / sigT per population.  Populations are contiguous row blocks
 * of pop_rows[j] rows in row order.  guess: one-component start state used for every population.  Outputs: W[p],
 * C[q], params n_pop x 5 column-major (B_T, sigX, sigY, sigH, sigT), logvalues steps x n_pop ROW-major (the
 * reference's `log`, rows 2..steps+1 of logvalue; room for EMsteps rows; may be NULL), steps.
 * Row-sharded contexts (ppls_ctx_create_dist): pop_rows[j] is THIS rank's number of rows of population j (same
 * populations in the same order on every rank, 0 where a rank holds none); the pooled sums of :481-482 are one
 * allreduce per pass and every rank returns the same results. */
int ppls_meta_run(ppls_ctx* ctx, const ppls_data* d, int32_t n_pop, const int64_t* pop_rows, int32_t EMsteps,
                  double atol, const ppls_params* guess, double* W, double* C, double* params, double* logvalues,
                  int32_t* steps);

/* ---- batched multi-start: n_starts start states over one resident dataset ----------------- */
/* Every start is an independent run of EM_W_multi.R:780-793 (its own break rule and trace); the
 * starts share each pass over the data (floor(8/r) starts per streaming-kernel launch).
 * inits / ests: arrays of n_starts structs (ests in LOOP order, caller-allocated W, C, B, sigT);
 * loglik: n_starts x max_steps, row s holds loglik[0 .. steps_done[s]-1] of start s (may be NULL);
 * timing: loop time / launches / passes of the whole batch (may be NULL). */
int ppls_simult_run_batched(ppls_ctx* ctx, const ppls_data* d, const ppls_opts* opts, int32_t n_starts,
                            const ppls_params* inits, ppls_params* ests, double* loglik,
                            int32_t* steps_done, ppls_timing* timing);

#ifdef __cplusplus
}
#endif
#endif /* PPLS_B200_H */
