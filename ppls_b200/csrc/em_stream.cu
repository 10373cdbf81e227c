// Host side of K1 "em_stream" (kernel: em_stream_kernel.cuh): launch plan + launcher.
#include "em_stream_kernel.cuh"
#include <cstdlib>
#include <cstdio>
#include <cstring>

namespace ppls {

static int env_int(const char* name, int dflt) {
  const char* v = getenv(name);
  return (v && *v) ? atoi(v) : dflt;
}

static stream_kernel_t lookup_kernel(int R, int M, int T, int keep) {
  switch (R) {
    case 1: return lookup_kernel_r1(M, T, keep);
    case 2: return lookup_kernel_r2(M, T, keep);
    case 3: return lookup_kernel_r3(M, T, keep);
    case 4: return lookup_kernel_r4(M, T, keep);
    case 5: return lookup_kernel_r5(M, T, keep);
    case 6: return lookup_kernel_r6(M, T, keep);
    case 7: return lookup_kernel_r7(M, T, keep);
    case 8: return lookup_kernel_r8(M, T, keep);
  }
  return nullptr;
}

static int max_threads_rt(int R, int M, int T, int keep) { return max_threads_for(R, M, T, keep != 0); }

// Split chX X-chunks and chY Y-chunks (64 columns each) over nw compute warps.
static bool split_chunks(int chX, int chY, int nw, StreamPlan* p) {
  if (nw < 2) nw = 2;
  if (nw > chX + chY) nw = chX + chY;
  int nwY = (int)((double)nw * chY / (chX + chY) + 0.5);
  if (nwY < 1) nwY = 1;
  if (nwY > chY) nwY = chY;
  int nwX = nw - nwY;
  if (nwX < 1) nwX = 1;
  if (nwX > chX) nwX = chX;
  p->nwX = nwX; p->nwY = nwY;
  p->cX = chX / nwX; p->bigX = chX % nwX;
  p->cY = chY / nwY; p->bigY = chY % nwY;
  const int mX = p->cX + (p->bigX ? 1 : 0), mY = p->cY + (p->bigY ? 1 : 0);
  p->M = mX > mY ? mX : mY;
  return true;
}

cudaError_t plan_stream(const DataShape& s, int R, int sm_count, StreamPlan* plan, const char** why) {
  *why = "";
  if (R < 1 || R > MAXR) { *why = "nr_comp outside 1..8"; return cudaErrorInvalidValue; }
  if (s.panel) { *why = "panel-major (wide) data layout"; return cudaErrorInvalidValue; }
  const int chX = (s.p2 / 2 + 31) / 32, chY = (s.q2 / 2 + 31) / 32;
  const size_t smem_max = 227 * 1024;  // usable dynamic shared memory per CTA (1 CTA per SM)
  const int envT = env_int("PPLS_K1_T", 0), envW = env_int("PPLS_K1_WARPS", 0), envS = env_int("PPLS_K1_STAGES", 0);
  const int keep = env_int("PPLS_K1_KEEP", 0) ? 1 : 0;
  const size_t rowbytes = (size_t)s.ld * 8;
  // candidates: T in {8,4,2} (larger T amortises the per-tile hand-offs), compute warps from the
  // preferred count (2 per SM sub-partition) upwards until the chunks per warp fit the register budget
  const int Ts[3] = {8, 4, 2};
  double best_score = -1.0;
  StreamPlan best{};
  for (int it = 0; it < 3; ++it) {
    const int T = Ts[it];
    if (envT && T != envT) continue;
    if (T * R > 32 || (T * R) % 2) continue;
    // up to 6 chunks per warp (fewer, fatter warps = less projection-reduce traffic through shared
    // memory), warps in multiples of the 4 SM sub-partitions (measured: tools/tune_k1.py)
    const int nw0 = envW ? envW : 4 * ((chX + chY + 23) / 24);
    for (int nw = nw0; nw <= 30; ++nw) {
      StreamPlan p{};
      split_chunks(chX, chY, nw, &p);
      const int nwc = p.nwX + p.nwY;
      if (p.M > 6 || !lookup_kernel(R, p.M, T, keep) || (nwc + 2) * 32 > max_threads_rt(R, p.M, T, keep)) {
        if (envW || nwc < nw) break;           // fixed by the caller / no more chunks to spread
        continue;
      }
      // stages: the fewest that keep >= 72 KB in flight while one stage is in phase A and one waits
      // for phase B (more in flight measured slower), else as many as fit
      int ns = 0;
      for (int c = envS ? envS : 3; c <= (envS ? envS : 8); ++c) {
        if (stream_smem_layout(R, T, c, nwc, s.ld).total > smem_max) break;
        ns = c;
        if ((double)(c - 2) * T * rowbytes >= 72.0 * 1024) break;
      }
      if (!ns) break;
      const double inflight = (double)(ns - 2) * T * rowbytes;
      const double score = (inflight >= 72.0 * 1024 ? 1e9 : inflight) + T * 1e3 - nwc;
      if (score > best_score) {
        best_score = score; best = p;
        best.R = R; best.T = T; best.nstage = ns; best.threads = (nwc + 2) * 32; best.keep = keep;
        best.smem_bytes = stream_smem_layout(R, T, ns, nwc, s.ld).total;
      }
      break;
    }
  }
  if (best_score < 0) {
    *why = "row tile does not fit shared memory / register budget (use the column-sharded path)";
    return cudaErrorInvalidValue;
  }
  *plan = best;
  stream_kernel_t k = lookup_kernel(R, plan->M, plan->T, plan->keep);
  plan->ntiles = (s.n + plan->T - 1) / plan->T;
  plan->ctas_per_sm = 1;
  int64_t grid = (int64_t)sm_count;
  if (grid > plan->ntiles) grid = plan->ntiles;
  if (grid < 1) grid = 1;
  plan->grid = (int)grid;
  // the fused tail (em_tail.cuh) reuses the dynamic shared memory from offset 0: make sure it is large enough
  if (plan->grid >= 2) {
    const size_t tb = tail_smem_bytes(R, s.ld, plan->grid, plan->threads);
    if (tb > plan->smem_bytes && tb <= smem_max) plan->smem_bytes = tb;
  }
  cudaError_t e = cudaFuncSetAttribute((const void*)k, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)plan->smem_bytes);
  if (e != cudaSuccess) { *why = "cudaFuncSetAttribute(MaxDynamicSharedMemorySize) failed"; return e; }
  if (env_int("PPLS_K1_VERBOSE", 0))
    fprintf(stderr, "[ppls] K1 plan: keep=%d R=%d M=%d T=%d stages=%d warps=%d(X %dx%d+%d)+%d(Y %dx%d+%d)+2 grid=%d smem=%zu\n",
            plan->keep, R, plan->M, plan->T, plan->nstage, plan->nwX, plan->nwX, plan->cX, plan->bigX, plan->nwY, plan->nwY,
            plan->cY, plan->bigY, plan->grid, plan->smem_bytes);
  return cudaSuccess;
}

bool stream_tail_fits(const StreamPlan& plan, const DataShape& s) {
  return plan.grid >= 2 && tail_smem_bytes(plan.R, s.ld, plan.grid, plan.threads) <= plan.smem_bytes;
}
size_t stream_tail_scratch_doubles(const StreamPlan& plan) { return tail_scratch_doubles(plan.grid, plan.R); }
size_t stream_tail_clock_offset(const StreamPlan& plan) { return tail_clock_offset(plan.grid, plan.R); }

cudaError_t launch_em_stream(const StreamPlan& plan, const DataShape& s, const double* D,
                             const double* L, const FitState* st, double* partials,
                             double* muT, double* muU, int write_mu, cudaStream_t stream, const FusedTail* ft) {
  stream_kernel_t k = lookup_kernel(plan.R, plan.M, plan.T, plan.keep);
  if (!k) return cudaErrorInvalidValue;
  StreamArgs a;
  memset(&a.tail, 0, sizeof(a.tail));
  a.fuse = 0;
  a.D = D; a.L = L; a.st = st; a.partials = partials; a.muT = muT; a.muU = muU;
  a.n = s.n; a.ntiles = plan.ntiles; a.ld = s.ld; a.p2 = s.p2; a.q2 = s.q2;
  a.nwX = plan.nwX; a.nwY = plan.nwY; a.nstage = plan.nstage; a.write_mu = write_mu;
  a.cX = plan.cX; a.bigX = plan.bigX; a.cY = plan.cY; a.bigY = plan.bigY;
  if (ft) {
    // ONE launch per EM iteration: streaming pass + cross-CTA reduction + exchange + update (em_tail.cuh).  Cooperative:
    // the tail's grid barriers need every CTA resident (1 CTA per SM, grid <= SM count).
    a.fuse = 1;
    a.tail.momvec = ft->momvec; a.tail.L = ft->L; a.tail.st = ft->st; a.tail.trace = ft->trace; a.tail.gscratch = ft->gscratch;
    a.tail.p = s.p; a.tail.q = s.q; a.tail.use_x = ft->xchg ? 1 : 0;
    if (ft->xchg) a.tail.x = *ft->xchg;
    void* args[] = {(void*)&a};
    return cudaLaunchCooperativeKernel((const void*)k, dim3(plan.grid), dim3(plan.threads), args, plan.smem_bytes, stream);
  }
  k<<<plan.grid, plan.threads, plan.smem_bytes, stream>>>(a);
  return cudaGetLastError();
}

}  // namespace ppls
