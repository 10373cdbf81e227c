// libppls_b200.so -- C ABI (include/ppls_b200.h) and host orchestration of the EM loop.
// Host side of PPLS_simult's loop, Package/PPLS/R/EM_W_multi.R:780-802, in C++: dataset
// residency, the pass loop with a device-side break flag, the NCCL exchange between the
// streaming kernel and the update kernel.  No CPU fallback: every compute entry point needs
// a CUDA device and fails with PPLS_E_CUDA otherwise.
#include "common.cuh"
#include <dlfcn.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <algorithm>
#include <string>
#include <vector>
#include <thread>
#include <new>
#include <mutex>
#include <unordered_map>

// ---- guarded device allocations (PPLS_GUARD=1) ---------------------------------------------------------------------
// compute-sanitizer is not available on every pool.  With PPLS_GUARD=1 every device allocation of this file carries 4 KB
// of 0xA5 bytes on either side; when it is freed the guards are read back and any changed byte aborts the process with
// the size of the allocation and the offset of the damage.  Catches writes past (or before) every workspace, ring and
// exchange buffer the engine owns; tests/test_gpu_guard.py runs every kernel family under it.  Off by default (no cost).
namespace {
constexpr size_t GUARD_BYTES = 4096;
struct GuardRec { unsigned char* base; size_t bytes, padded; };
std::mutex g_guard_mu;
std::unordered_map<void*, GuardRec> g_guard;
bool guard_on() {
  static const int on = [] { const char* e = getenv("PPLS_GUARD"); return (e && atoi(e)) ? 1 : 0; }();
  return on != 0;
}
cudaError_t guard_malloc(void** p, size_t bytes) {
  if (!guard_on()) return cudaMalloc(p, bytes);
  const size_t padded = (bytes + 255) / 256 * 256;
  unsigned char* base = nullptr;
  cudaError_t e = cudaMalloc(&base, padded + 2 * GUARD_BYTES);
  if (e != cudaSuccess) return e;
  e = cudaMemset(base, 0xA5, GUARD_BYTES);
  if (e == cudaSuccess) e = cudaMemset(base + GUARD_BYTES + bytes, 0xA5, padded - bytes + GUARD_BYTES);
  if (e != cudaSuccess) { cudaFree(base); return e; }
  *p = base + GUARD_BYTES;
  {  // PPLS_GUARD_SELFTEST=1: damage the guard on purpose (the test of the detector itself)
    static const int selftest = [] { const char* e = getenv("PPLS_GUARD_SELFTEST"); return (e && atoi(e)) ? 1 : 0; }();
    if (selftest) cudaMemset(base + GUARD_BYTES + bytes + 3, 0, 1);
  }
  std::lock_guard<std::mutex> lk(g_guard_mu);
  g_guard[*p] = GuardRec{base, bytes, padded};
  return cudaSuccess;
}
cudaError_t guard_free(void* p) {
  if (!guard_on() || !p) return cudaFree(p);
  GuardRec r{};
  {
    std::lock_guard<std::mutex> lk(g_guard_mu);
    auto it = g_guard.find(p);
    if (it == g_guard.end()) return cudaFree(p);
    r = it->second;
    g_guard.erase(it);
  }
  cudaDeviceSynchronize();
  const size_t tail = r.padded - r.bytes + GUARD_BYTES;
  std::vector<unsigned char> lo(GUARD_BYTES), hi(tail);
  cudaMemcpy(lo.data(), r.base, GUARD_BYTES, cudaMemcpyDeviceToHost);
  cudaMemcpy(hi.data(), r.base + GUARD_BYTES + r.bytes, tail, cudaMemcpyDeviceToHost);
  for (size_t i = 0; i < GUARD_BYTES; ++i)
    if (lo[i] != 0xA5) { fprintf(stderr, "[ppls] PPLS_GUARD: write %zu bytes BEFORE an allocation of %zu bytes\n", GUARD_BYTES - i, r.bytes); abort(); }
  for (size_t i = 0; i < tail; ++i)
    if (hi[i] != 0xA5) { fprintf(stderr, "[ppls] PPLS_GUARD: write %zu bytes PAST the end of an allocation of %zu bytes\n", i, r.bytes); abort(); }
  return cudaFree(r.base);
}
template <typename T> cudaError_t guard_malloc_t(T** p, size_t bytes) { return guard_malloc(reinterpret_cast<void**>(p), bytes); }
}  // namespace
#define cudaMalloc(p, n) guard_malloc_t((p), (n))
#define cudaFree(p) guard_free((void*)(p))

using namespace ppls;

// ---- errors ---------------------------------------------------------------------------------
static thread_local std::string g_err;
static int fail(int code, const std::string& msg) { g_err = msg; return code; }
#define CU(call)                                                                          \
  do {                                                                                    \
    cudaError_t e__ = (call);                                                             \
    if (e__ != cudaSuccess)                                                               \
      return fail(e__ == cudaErrorMemoryAllocation ? PPLS_E_OOM : PPLS_E_CUDA,            \
                  std::string(#call) + ": " + cudaGetErrorString(e__));                   \
  } while (0)

// ---- NCCL, loaded lazily (single-GPU use needs no libnccl) -----------------------------------
namespace {
typedef struct { char internal[128]; } nccl_uid;
typedef void* nccl_comm;
struct NcclApi {
  void* handle = nullptr;
  int (*GetUniqueId)(nccl_uid*) = nullptr;
  int (*CommInitRank)(nccl_comm*, int, nccl_uid, int) = nullptr;
  int (*AllReduce)(const void*, void*, size_t, int, int, nccl_comm, cudaStream_t) = nullptr;
  int (*CommDestroy)(nccl_comm) = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
};
NcclApi g_nccl;
const int kNcclFloat64 = 8, kNcclSum = 0;

int load_nccl() {
  if (g_nccl.handle) return PPLS_OK;
  const char* names[] = {getenv("PPLS_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
  void* h = nullptr;
  for (const char* nm : names) {
    if (!nm || !*nm) continue;
    h = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
    if (h) break;
  }
  if (!h) return fail(PPLS_E_NCCL, std::string("cannot load libnccl.so.2: ") + dlerror());
  g_nccl.GetUniqueId = (int (*)(nccl_uid*))dlsym(h, "ncclGetUniqueId");
  g_nccl.CommInitRank = (int (*)(nccl_comm*, int, nccl_uid, int))dlsym(h, "ncclCommInitRank");
  g_nccl.AllReduce = (int (*)(const void*, void*, size_t, int, int, nccl_comm, cudaStream_t))dlsym(h, "ncclAllReduce");
  g_nccl.CommDestroy = (int (*)(nccl_comm))dlsym(h, "ncclCommDestroy");
  g_nccl.GetErrorString = (const char* (*)(int))dlsym(h, "ncclGetErrorString");
  if (!g_nccl.GetUniqueId || !g_nccl.CommInitRank || !g_nccl.AllReduce || !g_nccl.CommDestroy)
    return fail(PPLS_E_NCCL, "libnccl.so.2 lacks a required symbol");
  g_nccl.handle = h;
  return PPLS_OK;
}
int nccl_fail(int rc, const char* what) {
  return fail(PPLS_E_NCCL, std::string(what) + ": " + (g_nccl.GetErrorString ? g_nccl.GetErrorString(rc) : "nccl error"));
}
}  // namespace

// ---- objects ----------------------------------------------------------------------------------
struct ppls_ctx {
  int device = 0, rank = 0, world = 1, sm_count = 0;
  cudaStream_t stream = nullptr;
  nccl_comm comm = nullptr;
  double* d_small = nullptr;   // 8 doubles of device scratch for tiny collectives
  // one-shot peer exchange (common.cuh::XchgDev): own block + the peers' blocks mapped through CUDA IPC
  bool xchg_on = false;
  char* xbuf = nullptr;
  void* xpeer[XCHG_MAXW] = {};
  XchgDev xdev{};
  // single-process multi-device form (ppls_ctx_create_multi): the handle the caller holds is a GROUP whose `subs`
  // are ordinary per-device contexts (rank = position, world = count) that reach each other's exchange blocks by
  // direct peer access; `group` points back from a member.  No NCCL communicator exists in this form.
  std::vector<ppls_ctx*> subs;
  ppls_ctx* group = nullptr;
};

struct ppls_data {
  ppls_ctx* ctx = nullptr;
  DataShape s{};
  double* D = nullptr;
  double ssq[2] = {0, 0};      // global
  int64_t n_total = 0;         // global
  // column-sharded mode (wide shapes): this rank holds ALL rows of a column block of X and of Y
  int colshard = 0;
  int64_t p_glob = 0, q_glob = 0;          // global logical column counts
  std::vector<int32_t> xoff, yoff;         // [world+1] cumulative padded (p2 / q2) widths of the ranks' blocks
  // multi-device context: the handle is a group of row shards (one per device, in row order)
  std::vector<ppls_data*> shards;
  std::vector<int64_t> shard_row0;
};

static const size_t XCHG_CTRL = XCHG_INBOX_OFFSET, XCHG_SLOT = 65536;   // exchange block: control words, per-CTA flags | inbox of 2 x world x slot doubles

static int ctx_activate(const ppls_ctx* ctx) {
  if (!ctx) return fail(PPLS_E_INVALID, "null context");
  if (!ctx->subs.empty())        // (the entry points that serve a multi-device context never get here with the group handle)
    return fail(PPLS_E_UNSUPPORTED, "not available on a multi-device context: the row-sharded EM loop (data_from_host / synthetic / "
                                    "export, fit_*, simult_run, loglik, orth) is; use a single-device context or one process per GPU");
  CU(cudaSetDevice(ctx->device));
  return PPLS_OK;
}

extern "C" const char* ppls_version(void) { return "ppls_b200 0.1 (sm_100a)"; }
extern "C" const char* ppls_last_error(void) { return g_err.c_str(); }

static int ctx_create_common(ppls_ctx** out, int device) {
  if (!out) return fail(PPLS_E_INVALID, "null output pointer");
  *out = nullptr;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0)
    return fail(PPLS_E_CUDA, std::string("no CUDA device (there is no CPU fallback): ") + cudaGetErrorString(e));
  if (device < 0 || device >= ndev) return fail(PPLS_E_INVALID, "device index out of range");
  CU(cudaSetDevice(device));
  cudaDeviceProp prop;
  CU(cudaGetDeviceProperties(&prop, device));
  if (prop.major < 10)
    return fail(PPLS_E_UNSUPPORTED, "this library is built for sm_100a (B200) only");
  ppls_ctx* c = new (std::nothrow) ppls_ctx();
  if (!c) return fail(PPLS_E_OOM, "host allocation failed");
  c->device = device;
  c->sm_count = prop.multiProcessorCount;
  cudaError_t e1 = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking);
  cudaError_t e2 = (e1 == cudaSuccess) ? cudaMalloc(&c->d_small, 8 * sizeof(double)) : e1;
  if (e2 != cudaSuccess) {
    if (c->stream) cudaStreamDestroy(c->stream);
    delete c;
    return fail(PPLS_E_CUDA, std::string("context setup: ") + cudaGetErrorString(e2));
  }
  *out = c;
  return PPLS_OK;
}

extern "C" int ppls_ctx_create(ppls_ctx** out, int device) { return ctx_create_common(out, device); }

extern "C" int ppls_nccl_unique_id(void* id_out, size_t nbytes) {
  if (!id_out || nbytes < PPLS_NCCL_ID_BYTES) return fail(PPLS_E_INVALID, "id buffer must hold 128 bytes");
  int rc = load_nccl();
  if (rc) return rc;
  nccl_uid id;
  int nr = g_nccl.GetUniqueId(&id);
  if (nr) return nccl_fail(nr, "ncclGetUniqueId");
  memcpy(id_out, &id, sizeof(id));
  return PPLS_OK;
}

static unsigned long long xchg_timeout_ns() {
  const char* e = getenv("PPLS_XCHG_TIMEOUT_MS");
  const long long ms = (e && atoll(e) > 0) ? atoll(e) : 20000;
  return (unsigned long long)ms * 1000000ull;
}
static int allreduce_sum(ppls_ctx* ctx, double* dbuf, size_t count);

// Peer exchange set-up: every rank allocates [control 1 KB | inbox 2 x world x slot doubles], publishes its
// CUDA IPC handle (bytes carried as doubles through one NCCL sum) and maps every peer's block.  All ranks
// then agree (a second sum) whether everybody succeeded; otherwise the NCCL allreduce stays in use.
static void xchg_setup(ppls_ctx* c) {
  const char* env = getenv("PPLS_XCHG");
  if (c->world < 2 || c->world > XCHG_MAXW || (env && *env && atoi(env) == 0)) return;
  const int W = c->world;
  const size_t bytes = XCHG_CTRL + (size_t)2 * W * XCHG_SLOT * sizeof(double);
  int ok = 1;
  double* hb = nullptr;      // [W][64] handle bytes + [1] failure count
  const size_t nd = (size_t)W * 64 + 1;
  if (cudaMalloc(&c->xbuf, bytes) != cudaSuccess) { c->xbuf = nullptr; ok = 0; }
  if (ok && cudaMemset(c->xbuf, 0, bytes) != cudaSuccess) ok = 0;
  cudaIpcMemHandle_t mine;
  memset(&mine, 0, sizeof(mine));
  if (ok && cudaIpcGetMemHandle(&mine, c->xbuf) != cudaSuccess) ok = 0;
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "CUDA IPC handle size");
  std::vector<double> host(nd, 0.0);
  for (int i = 0; i < 64; ++i) host[(size_t)c->rank * 64 + i] = (double)((const unsigned char*)&mine)[i];
  host[nd - 1] = ok ? 0.0 : 1.0;
  if (cudaMalloc(&hb, nd * sizeof(double)) != cudaSuccess) { cudaGetLastError(); hb = nullptr; }
  // from here on every rank must make the same collective calls
  bool coll = hb != nullptr;
  if (coll) coll = cudaMemcpyAsync(hb, host.data(), nd * sizeof(double), cudaMemcpyHostToDevice, c->stream) == cudaSuccess;
  if (coll) coll = allreduce_sum(c, hb, nd) == PPLS_OK;
  if (coll) coll = cudaMemcpyAsync(host.data(), hb, nd * sizeof(double), cudaMemcpyDeviceToHost, c->stream) == cudaSuccess;
  if (coll) coll = cudaStreamSynchronize(c->stream) == cudaSuccess;
  if (!coll || host[nd - 1] != 0.0) ok = 0;
  if (ok) {
    for (int r = 0; r < W && ok; ++r) {
      if (r == c->rank) { c->xpeer[r] = c->xbuf; continue; }
      cudaIpcMemHandle_t h;
      for (int i = 0; i < 64; ++i) ((unsigned char*)&h)[i] = (unsigned char)host[(size_t)r * 64 + i];
      if (cudaIpcOpenMemHandle(&c->xpeer[r], h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { c->xpeer[r] = nullptr; ok = 0; }
    }
  }
  cudaGetLastError();
  if (coll) {                                   // agreement: everybody mapped everybody
    double flag = ok ? 0.0 : 1.0;
    bool c2 = cudaMemcpyAsync(hb, &flag, sizeof(double), cudaMemcpyHostToDevice, c->stream) == cudaSuccess;
    if (c2) c2 = allreduce_sum(c, hb, 1) == PPLS_OK;
    if (c2) c2 = cudaMemcpyAsync(&flag, hb, sizeof(double), cudaMemcpyDeviceToHost, c->stream) == cudaSuccess;
    if (c2) c2 = cudaStreamSynchronize(c->stream) == cudaSuccess;
    if (!c2 || flag != 0.0) ok = 0;
  }
  if (hb) cudaFree(hb);
  if (!ok) {
    for (int r = 0; r < W; ++r) if (r != c->rank && c->xpeer[r]) { cudaIpcCloseMemHandle(c->xpeer[r]); c->xpeer[r] = nullptr; }
    if (c->xbuf) { cudaFree(c->xbuf); c->xbuf = nullptr; }
    cudaGetLastError();
    return;
  }
  XchgDev& x = c->xdev;
  memset(&x, 0, sizeof(x));
  for (int r = 0; r < W; ++r) {
    char* base = (char*)c->xpeer[r];
    x.flags[r] = (unsigned long long*)base;                     // [2][XCHG_MAXW] = 256 B
    x.cflags[r] = (unsigned long long*)(base + XCHG_CTRL_BYTES);
    x.inbox[r] = (double*)(base + XCHG_CTRL);
  }
  x.epoch = (unsigned long long*)(c->xbuf + 512);
  x.counter = (unsigned int*)(c->xbuf + 520);
  x.failed = (int*)(c->xbuf + 524);
  x.rank = c->rank; x.world = W; x.slot = (long long)XCHG_SLOT; x.timeout_ns = xchg_timeout_ns();
  c->xchg_on = true;
}

extern "C" int ppls_ctx_create_dist(ppls_ctx** out, int device, int rank, int world, const void* nccl_id,
                                    size_t nbytes) {
  if (world < 1 || rank < 0 || rank >= world) return fail(PPLS_E_INVALID, "bad rank/world");
  int rc = ctx_create_common(out, device);
  if (rc) return rc;
  (*out)->rank = rank; (*out)->world = world;
  if (world == 1) return PPLS_OK;
  if (!nccl_id || nbytes < PPLS_NCCL_ID_BYTES) { ppls_ctx_destroy(*out); *out = nullptr; return fail(PPLS_E_INVALID, "missing NCCL id"); }
  rc = load_nccl();
  if (rc) { ppls_ctx_destroy(*out); *out = nullptr; return rc; }
  nccl_uid id;
  memcpy(&id, nccl_id, sizeof(id));
  int nr = g_nccl.CommInitRank(&(*out)->comm, world, id, rank);
  if (nr) { ppls_ctx_destroy(*out); *out = nullptr; return nccl_fail(nr, "ncclCommInitRank"); }
  xchg_setup(*out);
  return PPLS_OK;
}

// Single-process multi-device context (ppls_ctx_create(ppls_ctx**, const int* device_ids, int n)): one host
// thread, one stream per device, direct peer access instead of CUDA IPC, the same inbox / flag exchange.
extern "C" int ppls_ctx_create_multi(ppls_ctx** out, const int* device_ids, int n_devices) {
  if (!out) return fail(PPLS_E_INVALID, "null output pointer");
  *out = nullptr;
  if (!device_ids || n_devices < 1 || n_devices > XCHG_MAXW) return fail(PPLS_E_INVALID, "need 1..16 device ids");
  for (int i = 0; i < n_devices; ++i)
    for (int j = 0; j < i; ++j)
      if (device_ids[i] == device_ids[j]) return fail(PPLS_E_INVALID, "a device is listed twice");
  if (n_devices == 1) return ctx_create_common(out, device_ids[0]);
  ppls_ctx* g = new (std::nothrow) ppls_ctx();
  if (!g) return fail(PPLS_E_OOM, "host allocation failed");
  g->device = device_ids[0]; g->rank = 0; g->world = n_devices;
  int rc = PPLS_OK;
  for (int i = 0; i < n_devices && rc == PPLS_OK; ++i) {
    ppls_ctx* c = nullptr;
    rc = ctx_create_common(&c, device_ids[i]);
    if (rc == PPLS_OK) { c->rank = i; c->world = n_devices; c->group = g; g->subs.push_back(c); }
  }
  const size_t bytes = XCHG_CTRL + (size_t)2 * n_devices * XCHG_SLOT * sizeof(double);
  for (int i = 0; i < n_devices && rc == PPLS_OK; ++i) {
    ppls_ctx* c = g->subs[i];
    cudaSetDevice(c->device);
    for (int j = 0; j < n_devices && rc == PPLS_OK; ++j) {
      if (j == i) continue;
      int can = 0;
      cudaDeviceCanAccessPeer(&can, c->device, g->subs[j]->device);
      if (!can) { rc = fail(PPLS_E_UNSUPPORTED, "devices of a multi-device context must have peer access to each other"); break; }
      cudaError_t e = cudaDeviceEnablePeerAccess(g->subs[j]->device, 0);
      if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); e = cudaSuccess; }
      if (e != cudaSuccess) rc = fail(PPLS_E_CUDA, std::string("cudaDeviceEnablePeerAccess: ") + cudaGetErrorString(e));
    }
    if (rc == PPLS_OK) {
      cudaError_t e = cudaMalloc(&c->xbuf, bytes);
      if (e == cudaSuccess) e = cudaMemset(c->xbuf, 0, bytes);
      if (e != cudaSuccess) rc = fail(PPLS_E_OOM, std::string("exchange block: ") + cudaGetErrorString(e));
    }
  }
  if (rc != PPLS_OK) { const std::string msg = g_err; ppls_ctx_destroy(g); g_err = msg; return rc; }
  g->sm_count = g->subs[0]->sm_count;
  for (int i = 0; i < n_devices; ++i) {
    ppls_ctx* c = g->subs[i];
    XchgDev& x = c->xdev;
    memset(&x, 0, sizeof(x));
    for (int r = 0; r < n_devices; ++r) {
      char* base = g->subs[r]->xbuf;
      x.flags[r] = (unsigned long long*)base;
      x.cflags[r] = (unsigned long long*)(base + XCHG_CTRL_BYTES);
      x.inbox[r] = (double*)(base + XCHG_CTRL);
    }
    x.epoch = (unsigned long long*)(c->xbuf + 512);
    x.counter = (unsigned int*)(c->xbuf + 520);
    x.failed = (int*)(c->xbuf + 524);
    x.rank = i; x.world = n_devices; x.slot = (long long)XCHG_SLOT; x.timeout_ns = xchg_timeout_ns();
    c->xchg_on = true;
  }
  *out = g;
  return PPLS_OK;
}

extern "C" int ppls_ctx_destroy(ppls_ctx* ctx) {
  if (!ctx) return PPLS_OK;
  if (!ctx->subs.empty()) {
    for (ppls_ctx* c : ctx->subs) if (c) { cudaSetDevice(c->device); if (c->stream) cudaStreamSynchronize(c->stream); }
    for (ppls_ctx* c : ctx->subs) ppls_ctx_destroy(c);
    delete ctx;
    return PPLS_OK;
  }
  cudaSetDevice(ctx->device);
  if (ctx->stream) cudaStreamSynchronize(ctx->stream);
  for (int r = 0; r < XCHG_MAXW; ++r)
    if (r != ctx->rank && ctx->xpeer[r]) cudaIpcCloseMemHandle(ctx->xpeer[r]);
  if (ctx->comm && g_nccl.CommDestroy) g_nccl.CommDestroy(ctx->comm);
  if (ctx->xbuf) cudaFree(ctx->xbuf);
  if (ctx->d_small) cudaFree(ctx->d_small);
  if (ctx->stream) cudaStreamDestroy(ctx->stream);
  delete ctx;
  return PPLS_OK;
}

extern "C" int ppls_ctx_info(const ppls_ctx* ctx, int* device, int* rank, int* world, int* sm_count) {
  if (!ctx) return fail(PPLS_E_INVALID, "null context");
  if (device) *device = ctx->device;
  if (rank) *rank = ctx->rank;
  if (world) *world = ctx->world;
  if (sm_count) *sm_count = ctx->sm_count;
  return PPLS_OK;
}

extern "C" int ppls_ctx_sync(ppls_ctx* ctx) {
  if (ctx && !ctx->subs.empty()) {
    for (ppls_ctx* c : ctx->subs) { int r = ppls_ctx_sync(c); if (r) return r; }
    return PPLS_OK;
  }
  int rc = ctx_activate(ctx);
  if (rc) return rc;
  CU(cudaStreamSynchronize(ctx->stream));
  return PPLS_OK;
}

extern "C" int ppls_probe(ppls_ctx* ctx, const ppls_data* d, int32_t which, double* value) {
  if (ctx && !ctx->subs.empty()) { ctx = ctx->subs[0]; if (d && !d->shards.empty()) d = d->shards[0]; }
  int rc = ctx_activate(ctx);
  if (rc) return rc;
  if (!value) return fail(PPLS_E_INVALID, "null output pointer");
  if (d && d->ctx != ctx) return fail(PPLS_E_INVALID, "data belongs to another context");
  double* scratch = nullptr;
  CU(cudaMalloc(&scratch, (size_t)ctx->sm_count * 256 * sizeof(double)));
  cudaError_t e = cudaSuccess;
  if (which == PPLS_PROBE_FP64_DMMA) {
    e = probe_dmma(ctx->sm_count, scratch, value, ctx->stream);
  } else if (which == PPLS_PROBE_HBM_READ) {
    void* buf = nullptr;
    size_t bytes = 0;
    if (d) { bytes = data_elems(d->s) * sizeof(double); }
    else {
      size_t fr = 0, tot = 0;
      e = cudaMemGetInfo(&fr, &tot);
      bytes = fr / 2 < ((size_t)8 << 30) ? fr / 2 : ((size_t)8 << 30);
      if (e == cudaSuccess) e = cudaMalloc(&buf, bytes);
      if (e == cudaSuccess) e = cudaMemsetAsync(buf, 0, bytes, ctx->stream);
    }
    if (e == cudaSuccess) {
      // ~0.2 s of streaming: long enough to see the clocks the power cap allows
      int launches = (int)(1.4e12 / (double)(bytes ? bytes : 1));
      if (launches < 3) launches = 3;
      if (launches > 400) launches = 400;
      e = probe_read(d ? (const void*)d->D : buf, bytes, ctx->sm_count, launches, scratch, value, ctx->stream);
    }
    if (buf) cudaFree(buf);
  } else { cudaFree(scratch); return fail(PPLS_E_INVALID, "unknown probe"); }
  cudaFree(scratch);
  if (e != cudaSuccess) return fail(PPLS_E_CUDA, std::string("probe: ") + cudaGetErrorString(e));
  return PPLS_OK;
}

// sum `count` doubles (device, in place) over all ranks
static int allreduce_sum(ppls_ctx* ctx, double* dbuf, size_t count) {
  if (ctx->world == 1) return PPLS_OK;
  if (!ctx->comm)
    return fail(PPLS_E_UNSUPPORTED, "this path needs a collective: on a multi-device context only the row-sharded EM loop "
                                    "is available (use one process per GPU, ppls_ctx_create_dist, for the others)");
  int nr = g_nccl.AllReduce(dbuf, dbuf, count, kNcclFloat64, kNcclSum, ctx->comm, ctx->stream);
  if (nr) return nccl_fail(nr, "ncclAllReduce");
  return PPLS_OK;
}

// ---- data ---------------------------------------------------------------------------------------
static int shape_init(DataShape* s, int64_t n, int64_t p, int64_t q) {
  if (n < 1 || p < 1 || q < 1) return fail(PPLS_E_INVALID, "X and Y need at least one row and one column");
  if (p > (1 << 24) || q > (1 << 24)) return fail(PPLS_E_UNSUPPORTED, "too many columns");
  s->n = n; s->p = (int32_t)p; s->q = (int32_t)q;
  s->p2 = (int32_t)((p + 1) & ~(int64_t)1); s->q2 = (int32_t)((q + 1) & ~(int64_t)1);
  s->ld = s->p2 + s->q2;
  const char* fp = getenv("PPLS_PANEL");
  s->panel = (fp && *fp) ? (atoi(fp) != 0) : (s->ld > PANEL_LD_MIN);
  s->npX = (s->p2 + PANEL_W - 1) / PANEL_W; s->npY = (s->q2 + PANEL_W - 1) / PANEL_W;
  return PPLS_OK;
}

// ||X||^2, ||Y||^2 and the row count, summed over ranks
static int data_finalize(ppls_ctx* ctx, ppls_data* d) {
  const int nblocks = ctx->sm_count * 4;
  double* scratch = nullptr;
  CU(cudaMalloc(&scratch, sizeof(double) * 2 * nblocks));
  cudaError_t e = launch_ssq(d->D, d->s, ctx->d_small, scratch, nblocks, ctx->stream);
  if (e != cudaSuccess) { cudaFree(scratch); return fail(PPLS_E_CUDA, std::string("ssq kernel: ") + cudaGetErrorString(e)); }
  double nloc = (double)d->s.n;
  e = cudaMemcpyAsync(ctx->d_small + 2, &nloc, sizeof(double), cudaMemcpyHostToDevice, ctx->stream);
  // (member of a multi-device context: local sums only -- the group adds its shards' sums on the host, in rank order)
  int rc = (e != cudaSuccess) ? fail(PPLS_E_CUDA, cudaGetErrorString(e)) : (ctx->group ? PPLS_OK : allreduce_sum(ctx, ctx->d_small, 3));
  double h[3] = {0, 0, 0};
  if (!rc) {
    e = cudaMemcpyAsync(h, ctx->d_small, 3 * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    if (e != cudaSuccess) rc = fail(PPLS_E_CUDA, std::string("ssq readback: ") + cudaGetErrorString(e));
  }
  cudaFree(scratch);
  if (rc) return rc;
  d->ssq[0] = h[0]; d->ssq[1] = h[1]; d->n_total = (int64_t)llround(h[2]);
  d->p_glob = d->s.p; d->q_glob = d->s.q;
  if (d->colshard) {
    // every rank holds all rows: N is not summed; the ranks' block widths are exchanged once
    d->n_total = d->s.n;
    const int W = ctx->world;
    std::vector<double> hw((size_t)4 * W, 0.0);
    hw[4 * ctx->rank + 0] = d->s.p; hw[4 * ctx->rank + 1] = d->s.q; hw[4 * ctx->rank + 2] = d->s.p2; hw[4 * ctx->rank + 3] = d->s.q2;
    double* dw = nullptr;
    CU(cudaMalloc(&dw, hw.size() * sizeof(double)));
    e = cudaMemcpyAsync(dw, hw.data(), hw.size() * sizeof(double), cudaMemcpyHostToDevice, ctx->stream);
    rc = (e == cudaSuccess) ? allreduce_sum(ctx, dw, hw.size()) : fail(PPLS_E_CUDA, cudaGetErrorString(e));
    if (!rc) {
      e = cudaMemcpyAsync(hw.data(), dw, hw.size() * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream);
      if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
      if (e != cudaSuccess) rc = fail(PPLS_E_CUDA, std::string("block-width exchange: ") + cudaGetErrorString(e));
    }
    cudaFree(dw);
    if (rc) return rc;
    d->xoff.assign(W + 1, 0); d->yoff.assign(W + 1, 0);
    d->p_glob = 0; d->q_glob = 0;
    for (int g = 0; g < W; ++g) {
      d->p_glob += (int64_t)llround(hw[4 * g + 0]); d->q_glob += (int64_t)llround(hw[4 * g + 1]);
      d->xoff[g + 1] = d->xoff[g] + (int32_t)llround(hw[4 * g + 2]);
      d->yoff[g + 1] = d->yoff[g] + (int32_t)llround(hw[4 * g + 3]);
    }
  }
  return PPLS_OK;
}

static int data_from_host_impl(ppls_ctx* ctx, const double* X, int64_t ldX, const double* Y, int64_t ldY,
                               int64_t n, int64_t p, int64_t q, int colshard, ppls_data** out);

// ---- exchange recovery -------------------------------------------------------------------------------------------
// The peer exchange keeps context-wide state (epoch, flags).  A wait that timed out (a peer died or abandoned its fit
// mid-run) leaves `failed` set and the ranks' epochs apart, and every later fit on the context would fail or pair up the
// wrong exchanges.  Creating a fit is collective, so that is where the ranks compare notes: everybody contributes its
// epoch and its failed flag; if anything differs, everybody clears its own control words and flags and the ranks meet
// again before anyone sends.  Normal cost: one 17-double allreduce per fit creation.
static int xchg_recover_dist(ppls_ctx* ctx) {
  if (!ctx->xchg_on || ctx->world == 1 || ctx->group) return PPLS_OK;
  const int W = ctx->world;
  struct { unsigned long long epoch; unsigned int counter; int failed; } hw;
  CU(cudaMemcpyAsync(&hw, ctx->xbuf + 512, sizeof(hw), cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  std::vector<double> v((size_t)XCHG_MAXW + 1, 0.0);
  v[ctx->rank] = (double)hw.epoch; v[XCHG_MAXW] = hw.failed ? 1.0 : 0.0;
  double* db = nullptr;
  CU(cudaMalloc(&db, v.size() * sizeof(double)));
  cudaError_t e = cudaMemcpyAsync(db, v.data(), v.size() * sizeof(double), cudaMemcpyHostToDevice, ctx->stream);
  int rc = (e == cudaSuccess) ? allreduce_sum(ctx, db, v.size()) : fail(PPLS_E_CUDA, cudaGetErrorString(e));
  if (rc == PPLS_OK) {
    e = cudaMemcpyAsync(v.data(), db, v.size() * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    if (e != cudaSuccess) rc = fail(PPLS_E_CUDA, cudaGetErrorString(e));
  }
  bool need = v[XCHG_MAXW] != 0.0;
  for (int r = 1; r < W; ++r) need = need || v[r] != v[0];
  if (rc == PPLS_OK && need) {
    e = cudaMemsetAsync(ctx->xbuf, 0, XCHG_INBOX_OFFSET, ctx->stream);           // control words, flags, per-CTA flags
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    if (e != cudaSuccess) rc = fail(PPLS_E_CUDA, cudaGetErrorString(e));
    if (rc == PPLS_OK) rc = allreduce_sum(ctx, db, 1);                            // nobody sends before everybody has cleared
    if (rc == PPLS_OK && cudaStreamSynchronize(ctx->stream) != cudaSuccess) rc = fail(PPLS_E_CUDA, "stream sync");
    if (getenv("PPLS_K1_VERBOSE") && atoi(getenv("PPLS_K1_VERBOSE")))
      fprintf(stderr, "[ppls] rank %d: peer exchange state reset (a previous fit failed or was abandoned)\n", ctx->rank);
  }
  cudaFree(db);
  return rc;
}
// The same for a multi-device context: one host thread sees every device's words.
static int xchg_recover_group(ppls_ctx* g) {
  bool need = false;
  unsigned long long e0 = 0;
  for (size_t i = 0; i < g->subs.size(); ++i) {
    ppls_ctx* c = g->subs[i];
    struct { unsigned long long epoch; unsigned int counter; int failed; } hw;
    CU(cudaSetDevice(c->device));
    CU(cudaStreamSynchronize(c->stream));
    CU(cudaMemcpy(&hw, c->xbuf + 512, sizeof(hw), cudaMemcpyDeviceToHost));
    if (i == 0) e0 = hw.epoch;
    need = need || hw.failed || hw.epoch != e0;
  }
  if (!need) return PPLS_OK;
  for (ppls_ctx* c : g->subs) {
    CU(cudaSetDevice(c->device));
    CU(cudaMemset(c->xbuf, 0, XCHG_INBOX_OFFSET));
    CU(cudaDeviceSynchronize());
  }
  return PPLS_OK;
}

static bool is_group(const ppls_ctx* c) { return c && !c->subs.empty(); }
static int no_group(const ppls_ctx* c, const char* what) {
  if (is_group(c))
    return fail(PPLS_E_UNSUPPORTED, std::string(what) + ": not available on a multi-device context (the row-sharded EM loop is; "
                                    "use a single-device context or one process per GPU for this entry point)");
  return PPLS_OK;
}

// Multi-device context: shard i holds rows [N i / G, N (i+1) / G).  `make(i, row0, rows, &shard)` builds one shard on
// member i; the shards are built by one host thread EACH (uploads from pageable memory block their thread, so this is
// what lets the G copies run concurrently), then ||X||^2, ||Y||^2 and N are added on the host in rank order.
template <class Make>
static int group_data_build(ppls_ctx* g, int64_t n, Make make, ppls_data** out) {
  const int G = (int)g->subs.size();
  if (n < G) return fail(PPLS_E_INVALID, "fewer rows than devices");
  ppls_data* d = new (std::nothrow) ppls_data();
  if (!d) return fail(PPLS_E_OOM, "host allocation failed");
  d->ctx = g;
  d->shards.assign(G, nullptr); d->shard_row0.assign(G + 1, 0);
  for (int i = 0; i <= G; ++i) d->shard_row0[i] = n * i / G;
  std::vector<int> rcs(G, PPLS_OK);
  std::vector<std::string> errs(G);
  std::vector<std::thread> th;
  for (int i = 0; i < G; ++i)
    th.emplace_back([&, i] {
      rcs[i] = make(i, d->shard_row0[i], d->shard_row0[i + 1] - d->shard_row0[i], &d->shards[i]);
      if (rcs[i]) errs[i] = g_err;           // (thread-local message of the worker)
    });
  for (auto& t : th) t.join();
  for (int i = 0; i < G; ++i)
    if (rcs[i]) {
      const int rc = rcs[i];
      const std::string msg = "device " + std::to_string(g->subs[i]->device) + ": " + errs[i];
      ppls_data_destroy(d);
      return fail(rc, msg);
    }
  double sx = 0.0, sy = 0.0;
  for (int i = 0; i < G; ++i) { sx += d->shards[i]->ssq[0]; sy += d->shards[i]->ssq[1]; }
  for (int i = 0; i < G; ++i) { d->shards[i]->ssq[0] = sx; d->shards[i]->ssq[1] = sy; d->shards[i]->n_total = n; }
  d->s = d->shards[0]->s; d->s.n = n;
  d->ssq[0] = sx; d->ssq[1] = sy; d->n_total = n;
  d->p_glob = d->s.p; d->q_glob = d->s.q;
  *out = d;
  return PPLS_OK;
}

extern "C" int ppls_data_from_host(ppls_ctx* ctx, const double* X, int64_t ldX, const double* Y, int64_t ldY,
                                   int64_t n, int64_t p, int64_t q, ppls_data** out) {
  if (is_group(ctx)) {
    if (!out) return fail(PPLS_E_INVALID, "null output pointer");
    *out = nullptr;
    if (!X || !Y) return fail(PPLS_E_INVALID, "null X or Y");
    if (ldX < n || ldY < n) return fail(PPLS_E_INVALID, "leading dimension smaller than the row count");
    return group_data_build(ctx, n, [&](int i, int64_t row0, int64_t rows, ppls_data** sh) {
      return data_from_host_impl(ctx->subs[i], X + row0, ldX, Y + row0, ldY, rows, p, q, 0, sh);
    }, out);
  }
  return data_from_host_impl(ctx, X, ldX, Y, ldY, n, p, q, 0, out);
}
extern "C" int ppls_data_from_host_cols(ppls_ctx* ctx, const double* X, int64_t ldX, const double* Y, int64_t ldY,
                                        int64_t n, int64_t p_local, int64_t q_local, ppls_data** out) {
  { int rg = no_group(ctx, "ppls_data_from_host_cols"); if (rg) return rg; }
  return data_from_host_impl(ctx, X, ldX, Y, ldY, n, p_local, q_local, 1, out);
}
static int data_from_host_impl(ppls_ctx* ctx, const double* X, int64_t ldX, const double* Y, int64_t ldY,
                               int64_t n, int64_t p, int64_t q, int colshard, ppls_data** out) {
  if (!out) return fail(PPLS_E_INVALID, "null output pointer");
  *out = nullptr;
  int rc = ctx_activate(ctx);
  if (rc) return rc;
  if (!X || !Y) return fail(PPLS_E_INVALID, "null X or Y");
  if (ldX < n || ldY < n) return fail(PPLS_E_INVALID, "leading dimension smaller than the row count");
  ppls_data* d = new (std::nothrow) ppls_data();
  if (!d) return fail(PPLS_E_OOM, "host allocation failed");
  d->ctx = ctx; d->colshard = colshard;
  rc = shape_init(&d->s, n, p, q);
  if (rc) { delete d; return rc; }
  const size_t bytes = data_elems(d->s) * sizeof(double);
  cudaError_t e = cudaMalloc(&d->D, bytes);
  if (e == cudaSuccess && d->s.panel) e = cudaMemsetAsync(d->D, 0, bytes, ctx->stream);
  if (e != cudaSuccess) { delete d; return fail(PPLS_E_OOM, std::string("device allocation of the data shard: ") + cudaGetErrorString(e)); }
  // staging: chunks of rows, column-major as they lie in the caller's arrays.  Two staging buffers and a copy
  // stream of its own: the host->device copy of chunk k+1 runs while chunk k is re-tiled into the row-major shard.
  const size_t stage_budget = (size_t)128 << 20;
  int64_t cr = (int64_t)(stage_budget / (sizeof(double) * (size_t)(p + q)));
  if (cr > n) cr = n;
  if (cr < 32) cr = (n < 32) ? n : 32;
  const size_t stg_elems = (size_t)cr * (size_t)(p + q);
  double* stg = nullptr;
  e = cudaMalloc(&stg, 2 * sizeof(double) * stg_elems);
  if (e != cudaSuccess) { cudaFree(d->D); delete d; return fail(PPLS_E_OOM, "device allocation of the staging buffers"); }
  cudaStream_t cs = nullptr;
  cudaEvent_t copied[2] = {nullptr, nullptr}, tiled[2] = {nullptr, nullptr};
  e = cudaStreamCreateWithFlags(&cs, cudaStreamNonBlocking);
  for (int b = 0; b < 2 && e == cudaSuccess; ++b) {
    e = cudaEventCreateWithFlags(&copied[b], cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&tiled[b], cudaEventDisableTiming);
  }
  int64_t k = 0;
  for (int64_t i0 = 0; i0 < n && e == cudaSuccess; i0 += cr, ++k) {
    const int b = (int)(k & 1);
    const int64_t rows = (n - i0 < cr) ? (n - i0) : cr;
    double* stgX = stg + (size_t)b * stg_elems;
    double* stgY = stgX + (size_t)cr * p;
    if (k >= 2) e = cudaStreamWaitEvent(cs, tiled[b], 0);          // the re-tile of chunk k-2 has read this buffer
    if (e == cudaSuccess)
      e = cudaMemcpy2DAsync(stgX, rows * sizeof(double), X + i0, ldX * sizeof(double), rows * sizeof(double), (size_t)p,
                            cudaMemcpyHostToDevice, cs);
    if (e == cudaSuccess)
      e = cudaMemcpy2DAsync(stgY, rows * sizeof(double), Y + i0, ldY * sizeof(double), rows * sizeof(double), (size_t)q,
                            cudaMemcpyHostToDevice, cs);
    if (e == cudaSuccess) e = cudaEventRecord(copied[b], cs);
    if (e == cudaSuccess) e = cudaStreamWaitEvent(ctx->stream, copied[b], 0);
    if (e == cudaSuccess) e = launch_retile(stgX, rows, rows, (int)p, d->D, d->s, 0, i0, ctx->stream);
    if (e == cudaSuccess) e = launch_retile(stgY, rows, rows, (int)q, d->D, d->s, d->s.p2, i0, ctx->stream);
    if (e == cudaSuccess) e = cudaEventRecord(tiled[b], ctx->stream);
  }
  if (e == cudaSuccess) e = launch_zero_pad(d->D, d->s, ctx->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
  cudaError_t e2 = cs ? cudaStreamSynchronize(cs) : cudaSuccess;
  if (e == cudaSuccess) e = e2;
  for (int b = 0; b < 2; ++b) { if (copied[b]) cudaEventDestroy(copied[b]); if (tiled[b]) cudaEventDestroy(tiled[b]); }
  if (cs) cudaStreamDestroy(cs);
  cudaFree(stg);
  if (e != cudaSuccess) { cudaFree(d->D); delete d; return fail(PPLS_E_CUDA, std::string("upload: ") + cudaGetErrorString(e)); }
  rc = data_finalize(ctx, d);
  if (rc) { cudaFree(d->D); delete d; return rc; }
  *out = d;
  return PPLS_OK;
}

// loadings in the engine layout: L[ld][R] row-major, W rows then C rows, pad rows zero
static void pack_loadings(const DataShape& s, int R, const double* W, const double* C, std::vector<double>& L) {
  L.assign((size_t)s.ld * R, 0.0);
  for (int k = 0; k < R; ++k) {
    for (int j = 0; j < s.p; ++j) L[(size_t)j * R + k] = W[j + (size_t)k * s.p];
    for (int j = 0; j < s.q; ++j) L[(size_t)(s.p2 + j) * R + k] = C[j + (size_t)k * s.q];
  }
}

static int check_params(const ppls_params* P, const char* what) {
  if (!P || !P->W || !P->C || !P->B || !P->sigT) return fail(PPLS_E_INVALID, std::string(what) + ": null parameter pointer");
  return PPLS_OK;
}

struct HostOrth { int nx = 0, ny = 0; const double *PYo = nullptr, *PXo = nullptr, *sigTo = nullptr, *sigUo = nullptr; };
static int data_synthetic_impl(ppls_ctx* ctx, int64_t n, int64_t row0, int64_t p, int64_t q, int64_t colX0, int64_t colY0,
                               int colshard, int32_t r, uint64_t seed, const ppls_params* truth, ppls_data** out,
                               const HostOrth* ho = nullptr);
extern "C" int ppls_data_synthetic(ppls_ctx* ctx, int64_t n, int64_t row0, int64_t p, int64_t q, int32_t r,
                                   uint64_t seed, const ppls_params* truth, ppls_data** out) {
  if (is_group(ctx)) {
    if (!out) return fail(PPLS_E_INVALID, "null output pointer");
    *out = nullptr;
    return group_data_build(ctx, n, [&](int i, int64_t lo, int64_t rows, ppls_data** sh) {
      return data_synthetic_impl(ctx->subs[i], rows, row0 + lo, p, q, 0, 0, 0, r, seed, truth, sh);
    }, out);
  }
  return data_synthetic_impl(ctx, n, row0, p, q, 0, 0, 0, r, seed, truth, out);
}
extern "C" int ppls_data_synthetic_cols(ppls_ctx* ctx, int64_t n, int64_t p_local, int64_t q_local, int64_t colX0,
                                        int64_t colY0, int32_t r, uint64_t seed, const ppls_params* truth_local,
                                        ppls_data** out) {
  { int rg = no_group(ctx, "ppls_data_synthetic_cols"); if (rg) return rg; }
  return data_synthetic_impl(ctx, n, 0, p_local, q_local, colX0, colY0, 1, r, seed, truth_local, out);
}
static int data_synthetic_impl(ppls_ctx* ctx, int64_t n, int64_t row0, int64_t p, int64_t q, int64_t colX0, int64_t colY0,
                               int colshard, int32_t r, uint64_t seed, const ppls_params* truth, ppls_data** out,
                               const HostOrth* ho) {
  if (!out) return fail(PPLS_E_INVALID, "null output pointer");
  *out = nullptr;
  int rc = ctx_activate(ctx);
  if (rc) return rc;
  rc = check_params(truth, "truth");
  if (rc) return rc;
  if (r < 1 || r > MAXR) return fail(PPLS_E_INVALID, "nr_comp must be in 1..8");
  if (row0 < 0 || colX0 < 0 || colY0 < 0) return fail(PPLS_E_INVALID, "negative row / column offset");
  ppls_data* d = new (std::nothrow) ppls_data();
  if (!d) return fail(PPLS_E_OOM, "host allocation failed");
  d->ctx = ctx; d->colshard = colshard;
  rc = shape_init(&d->s, n, p, q);
  if (rc) { delete d; return rc; }
  cudaError_t e = cudaMalloc(&d->D, data_elems(d->s) * sizeof(double));
  if (e != cudaSuccess) { delete d; return fail(PPLS_E_OOM, std::string("device allocation of the data shard: ") + cudaGetErrorString(e)); }
  if (d->s.panel) cudaMemsetAsync(d->D, 0, data_elems(d->s) * sizeof(double), ctx->stream);
  std::vector<double> L;
  pack_loadings(d->s, r, truth->W, truth->C, L);
  double *dL = nullptr, *dbt = nullptr;
  e = cudaMalloc(&dL, L.size() * sizeof(double));
  if (e == cudaSuccess) e = cudaMalloc(&dbt, 2 * MAXR * sizeof(double));
  double bt[2 * MAXR] = {0};
  for (int k = 0; k < r; ++k) { bt[k] = truth->B[k]; bt[MAXR + k] = truth->sigT[k]; }
  if (e == cudaSuccess) e = cudaMemcpyAsync(dL, L.data(), L.size() * sizeof(double), cudaMemcpyHostToDevice, ctx->stream);
  if (e == cudaSuccess) e = cudaMemcpyAsync(dbt, bt, sizeof(bt), cudaMemcpyHostToDevice, ctx->stream);
  // orthogonal parts: P_Yosc (p x nx), P_Xosc (q x ny) column-major on the host -> row-major on the device
  SynthOrth so{};
  double* dO = nullptr;
  std::vector<double> ob;
  if (ho && (ho->nx > 0 || ho->ny > 0)) {
    so.nx = ho->nx; so.ny = ho->ny;
    const size_t o1 = 0, o2 = o1 + (size_t)p * so.nx, o3 = o2 + (size_t)q * so.ny, o4 = o3 + MAXR;
    ob.assign(o4 + MAXR, 0.0);
    for (int k = 0; k < so.nx; ++k) { for (int64_t j = 0; j < p; ++j) ob[o1 + (size_t)j * so.nx + k] = ho->PYo[j + (size_t)k * p]; ob[o3 + k] = ho->sigTo[k]; }
    for (int k = 0; k < so.ny; ++k) { for (int64_t j = 0; j < q; ++j) ob[o2 + (size_t)j * so.ny + k] = ho->PXo[j + (size_t)k * q]; ob[o4 + k] = ho->sigUo[k]; }
    if (e == cudaSuccess) e = cudaMalloc(&dO, ob.size() * sizeof(double));
    if (e == cudaSuccess) e = cudaMemcpyAsync(dO, ob.data(), ob.size() * sizeof(double), cudaMemcpyHostToDevice, ctx->stream);
    so.PYo = dO + o1; so.PXo = dO + o2; so.sigTo = dO + o3; so.sigUo = dO + o4;
  }
  if (e == cudaSuccess)
    e = launch_synth(d->D, d->s, row0, (uint32_t)colX0, (uint32_t)colY0, r, seed, dL, dbt, dbt + MAXR, truth->sigE, truth->sigF, truth->sigH, ctx->stream, &so);
  if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
  if (dO) cudaFree(dO);
  if (dL) cudaFree(dL);
  if (dbt) cudaFree(dbt);
  if (e != cudaSuccess) { cudaFree(d->D); delete d; return fail(PPLS_E_CUDA, std::string("synthetic generator: ") + cudaGetErrorString(e)); }
  rc = data_finalize(ctx, d);
  if (rc) { cudaFree(d->D); delete d; return rc; }
  *out = d;
  return PPLS_OK;
}

// simulC(N, W, C, P_Yosc, P_Xosc, B_T, sigX, sigY, sigH, sigT, sigTo, sigUo) (Package/PPLS/src/loglC.cpp:269-315,
// R doc R/RcppExports.R:12-30) as a resident dataset: the same model, generalised to r components, on the
// counter-based generator (R's RNG stream cannot be reproduced outside R).
extern "C" int ppls_data_simulc(ppls_ctx* ctx, int64_t n, int64_t row0, int64_t p, int64_t q, int32_t r, int32_t nx, int32_t ny,
                                uint64_t seed, const ppls_params* truth, const double* P_Yosc, const double* P_Xosc,
                                const double* sigTo, const double* sigUo, ppls_data** out) {
  { int rg = no_group(ctx, "ppls_data_simulc"); if (rg) return rg; }
  if (nx < 0 || ny < 0 || nx > MAXR || ny > MAXR) return fail(PPLS_E_INVALID, "orthogonal components outside 0..8");
  if ((nx > 0 && (!P_Yosc || !sigTo)) || (ny > 0 && (!P_Xosc || !sigUo))) return fail(PPLS_E_INVALID, "null orthogonal loadings");
  HostOrth ho;
  ho.nx = nx; ho.ny = ny; ho.PYo = P_Yosc; ho.PXo = P_Xosc; ho.sigTo = sigTo; ho.sigUo = sigUo;
  return data_synthetic_impl(ctx, n, row0, p, q, 0, 0, 0, r, seed, truth, out, &ho);
}

extern "C" int ppls_data_export(ppls_ctx* ctx, const ppls_data* d, int64_t row0, int64_t n, double* X, int64_t ldX,
                                double* Y, int64_t ldY) {
  if (d && !d->shards.empty()) {
    if (!X || !Y) return fail(PPLS_E_INVALID, "null argument");
    if (row0 < 0 || n < 0 || row0 + n > d->s.n || ldX < n || ldY < n) return fail(PPLS_E_INVALID, "row range / leading dimension");
    for (size_t i = 0; i < d->shards.size(); ++i) {       // the part of [row0, row0 + n) that lives on shard i
      const int64_t lo = std::max(row0, d->shard_row0[i]), hi = std::min(row0 + n, d->shard_row0[i + 1]);
      if (hi <= lo) continue;
      int rc = ppls_data_export(d->shards[i]->ctx, d->shards[i], lo - d->shard_row0[i], hi - lo, X + (lo - row0), ldX,
                                Y + (lo - row0), ldY);
      if (rc) return rc;
    }
    return PPLS_OK;
  }
  int rc = ctx_activate(ctx);
  if (rc) return rc;
  if (!d || !X || !Y) return fail(PPLS_E_INVALID, "null argument");
  if (row0 < 0 || n < 0 || row0 + n > d->s.n || ldX < n || ldY < n) return fail(PPLS_E_INVALID, "row range / leading dimension");
  if (n == 0) return PPLS_OK;
  const int p = d->s.p, q = d->s.q;
  int64_t cr = (int64_t)(((size_t)64 << 20) / (sizeof(double) * (size_t)(p + q)));
  if (cr > n) cr = n;
  if (cr < 32) cr = (n < 32) ? n : 32;
  double* stg = nullptr;
  CU(cudaMalloc(&stg, sizeof(double) * (size_t)cr * (size_t)(p + q)));
  cudaError_t e = cudaSuccess;
  for (int64_t i0 = 0; i0 < n && e == cudaSuccess; i0 += cr) {
    const int64_t rows = (n - i0 < cr) ? (n - i0) : cr;
    double* sX = stg; double* sY = stg + (size_t)cr * p;
    e = launch_untile(d->D, d->s, 0, row0 + i0, rows, p, sX, rows, ctx->stream);
    if (e == cudaSuccess) e = launch_untile(d->D, d->s, d->s.p2, row0 + i0, rows, q, sY, rows, ctx->stream);
    if (e == cudaSuccess)
      e = cudaMemcpy2DAsync(X + i0, ldX * sizeof(double), sX, rows * sizeof(double), rows * sizeof(double), (size_t)p,
                            cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess)
      e = cudaMemcpy2DAsync(Y + i0, ldY * sizeof(double), sY, rows * sizeof(double), rows * sizeof(double), (size_t)q,
                            cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
  }
  cudaFree(stg);
  if (e != cudaSuccess) return fail(PPLS_E_CUDA, std::string("export: ") + cudaGetErrorString(e));
  return PPLS_OK;
}

extern "C" int ppls_data_info(const ppls_data* d, int64_t* n, int64_t* p, int64_t* q, int64_t* n_total, double* ssqX,
                              double* ssqY) {
  if (!d) return fail(PPLS_E_INVALID, "null data handle");
  if (n) *n = d->s.n;
  if (p) *p = d->s.p;
  if (q) *q = d->s.q;
  if (n_total) *n_total = d->n_total;
  if (ssqX) *ssqX = d->ssq[0];
  if (ssqY) *ssqY = d->ssq[1];
  return PPLS_OK;
}

extern "C" int ppls_data_destroy(ppls_data* d) {
  if (!d) return PPLS_OK;
  for (ppls_data* sh : d->shards) ppls_data_destroy(sh);
  d->shards.clear();
  if (d->ctx) cudaSetDevice(d->ctx->device);
  if (d->D) cudaFree(d->D);
  delete d;
  return PPLS_OK;
}

// ---- column-sharded mode: local <-> global rows of the (p+q) x r objects -------------------------------
// Global layout: [X rows of rank 0 .. rank W-1 | Y rows of rank 0 .. rank W-1] (padded widths).
__global__ void rows_local_to_global(const double* __restrict__ loc, double* __restrict__ glob, int p2l, int q2l,
                                     int xoff, int yoff_abs, int R) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= (p2l + q2l) * R) return;
  const int j = e / R, k = e - j * R;
  const int gj = (j < p2l) ? xoff + j : yoff_abs + (j - p2l);
  glob[(size_t)gj * R + k] = loc[e];
}
__global__ void rows_global_to_local(const double* __restrict__ glob, double* __restrict__ loc, int p2l, int q2l,
                                     int xoff, int yoff_abs, int R) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= (p2l + q2l) * R) return;
  const int j = e / R, k = e - j * R;
  const int gj = (j < p2l) ? xoff + j : yoff_abs + (j - p2l);
  loc[e] = glob[(size_t)gj * R + k];
}

// ---- one EM run -----------------------------------------------------------------------------------
struct ppls_fit {
  ppls_ctx* ctx = nullptr;
  const ppls_data* data = nullptr;
  ppls_opts opts{};
  StreamPlan plan{};
  bool wide = false;             // two-pass wide-row path (em_wide.cu) instead of the fused kernel
  WidePlan wplan{};
  WideOffsets woff{};
  double* wide_ws = nullptr;
  double* upd_scratch = nullptr; // Gram partials of the grid update kernel
  bool fused = false;            // one launch per EM iteration: the streaming kernel carries the reduction, the exchange and the update (em_tail.cuh)
  double* tail_scratch = nullptr;
  unsigned long long* clk_host = nullptr;   // pinned: device clocks of the fused launches (streaming ns, tail ns, launches)
  bool seq = false;              // sequential one-component fit (em_seq.cu) instead of the simultaneous update
  const double* prev = nullptr;  // [nprev][ld] loadings of the earlier components (device, owned by the caller)
  int nprev = 0;
  SeqCons cons{};                // sequential fit: constraints / critfunc of this component
  bool cols = false;             // column-sharded: the update runs replicated on the GLOBAL (p+q) x r objects
  DataShape sg{};                // global shape (all ranks' column blocks)
  double *Gglob = nullptr, *Lglob = nullptr, *Lscratch_g = nullptr;
  int xoff = 0, yoff_abs = 0;    // first global row of this rank's X / Y block
  int R = 0;
  int64_t veclen = 0;
  double *L = nullptr, *Lscratch = nullptr, *partials = nullptr, *momvec = nullptr, *trace = nullptr;
  FitState* st = nullptr;        // device
  FitState* st_host = nullptr;   // pinned mirror
  int32_t* done_host = nullptr;  // pinned, mapped: set by the device when it raises `done` (FitState::done_host)
  bool finished = false;
  int32_t passes = 0;
  std::vector<ppls_fit*> parts;  // multi-device context: one fit per device, driven in lockstep
  // timing
  std::vector<cudaEvent_t> ev_free;
  std::vector<std::pair<cudaEvent_t, cudaEvent_t>> ev_k1, ev_tail;
  ppls_timing tm{};
};

static void fit_free(ppls_fit* f) {
  if (!f) return;
  for (ppls_fit* part : f->parts) fit_free(part);
  f->parts.clear();
  if (f->ctx) cudaSetDevice(f->ctx->device);
  for (auto& pr : f->ev_k1) { cudaEventDestroy(pr.first); cudaEventDestroy(pr.second); }
  for (auto& pr : f->ev_tail) cudaEventDestroy(pr.second);
  for (auto e : f->ev_free) cudaEventDestroy(e);
  if (f->clk_host) cudaFreeHost(f->clk_host);
  cudaFree(f->tail_scratch);
  cudaFree(f->wide_ws); cudaFree(f->upd_scratch); cudaFree(f->Gglob); cudaFree(f->Lglob); cudaFree(f->Lscratch_g); cudaFree(f->L); cudaFree(f->Lscratch); cudaFree(f->partials); cudaFree(f->momvec); cudaFree(f->trace);
  cudaFree(f->st);
  if (f->st_host) cudaFreeHost(f->st_host);
  if (f->done_host) cudaFreeHost(f->done_host);
  delete f;
}

static int fit_create_impl(ppls_ctx* ctx, const ppls_data* d, const ppls_opts* opts, const ppls_params* init,
                           bool seq, const double* prev, int nprev, const double* ssq_deflated, ppls_fit** out,
                           const SeqCons* cons = nullptr);
extern "C" int ppls_fit_create(ppls_ctx* ctx, const ppls_data* d, const ppls_opts* opts, const ppls_params* init,
                               ppls_fit** out) {
  if (is_group(ctx)) {
    if (!out) return fail(PPLS_E_INVALID, "null output pointer");
    *out = nullptr;
    if (!d || !opts) return fail(PPLS_E_INVALID, "null data or options");
    if (d->ctx != ctx || d->shards.size() != ctx->subs.size()) return fail(PPLS_E_INVALID, "data belongs to another context");
    ppls_fit* f = new (std::nothrow) ppls_fit();
    if (!f) return fail(PPLS_E_OOM, "host allocation failed");
    f->ctx = ctx; f->data = d; f->opts = *opts; f->R = opts->r;
    { int rr = xchg_recover_group(ctx); if (rr) { fit_free(f); return rr; } }
    for (size_t i = 0; i < ctx->subs.size(); ++i) {
      ppls_fit* part = nullptr;
      int rc = fit_create_impl(ctx->subs[i], d->shards[i], opts, init, false, nullptr, 0, nullptr, &part);
      if (rc) { const std::string msg = g_err; fit_free(f); g_err = msg; return rc; }
      f->parts.push_back(part);
    }
    *out = f;
    return PPLS_OK;
  }
  return fit_create_impl(ctx, d, opts, init, false, nullptr, 0, nullptr, out);
}
static int fit_create_impl(ppls_ctx* ctx, const ppls_data* d, const ppls_opts* opts, const ppls_params* init,
                           bool seq, const double* prev, int nprev, const double* ssq_deflated, ppls_fit** out,
                           const SeqCons* cons) {
  if (!out) return fail(PPLS_E_INVALID, "null output pointer");
  *out = nullptr;
  int rc = ctx_activate(ctx);
  if (rc) return rc;
  if (!d || !opts) return fail(PPLS_E_INVALID, "null data or options");
  if (d->ctx != ctx) return fail(PPLS_E_INVALID, "data belongs to another context");
  rc = check_params(init, "init");
  if (rc) return rc;
  const int R = opts->r;
  if (R < 1 || R > MAXR) return fail(PPLS_E_INVALID, "nr_comp must be in 1..8");
  if (!d->colshard && (R > d->s.p || R > d->s.q)) return fail(PPLS_E_INVALID, "nr_comp exceeds ncol(X) or ncol(Y)");
  if (opts->max_steps < 0) return fail(PPLS_E_INVALID, "EMsteps must be >= 0");
  if (opts->orth_type != PPLS_ORTH_SVD && opts->orth_type != PPLS_ORTH_QR) return fail(PPLS_E_INVALID, "type must be SVD or QR");
  ppls_fit* f = new (std::nothrow) ppls_fit();
  if (!f) return fail(PPLS_E_OOM, "host allocation failed");
  f->ctx = ctx; f->data = d; f->opts = *opts; f->R = R;
  f->seq = seq; f->prev = prev; f->nprev = nprev;
  if (cons) f->cons = *cons;
  if (seq && (R != 1 || d->colshard)) { fit_free(f); return fail(PPLS_E_UNSUPPORTED, "sequential fit: one component at a time, row-sharded data"); }
  const char* why = "";
  f->cols = d->colshard != 0;
  if (f->cols) {
    f->sg = d->s;
    f->sg.p = (int32_t)d->p_glob; f->sg.q = (int32_t)d->q_glob;
    f->sg.p2 = d->xoff[ctx->world]; f->sg.q2 = d->yoff[ctx->world]; f->sg.ld = f->sg.p2 + f->sg.q2;
    f->xoff = d->xoff[ctx->rank]; f->yoff_abs = f->sg.p2 + d->yoff[ctx->rank];
    if (R > d->p_glob || R > d->q_glob) { fit_free(f); return fail(PPLS_E_INVALID, "nr_comp exceeds ncol(X) or ncol(Y)"); }
  }
  const char* force_wide = getenv("PPLS_FORCE_WIDE");
  cudaError_t e = ((force_wide && atoi(force_wide)) || f->cols) ? cudaErrorInvalidValue : plan_stream(d->s, R, ctx->sm_count, &f->plan, &why);
  if (e != cudaSuccess) {
    // the row does not fit the fused kernel's shared-memory tile: two-pass wide-row path
    const std::string why1 = why;
    e = plan_wide(d->s, R, ctx->sm_count, &f->wplan, &why);
    if (e != cudaSuccess) { fit_free(f); return fail(PPLS_E_UNSUPPORTED, "streaming-kernel plan: " + why1 + "; " + why); }
    f->wide = true;
    // one pass from HBM when the data is not column-sharded (the sharded form exchanges the projections between the passes)
    if (!f->cols) { f->wplan.R = R; plan_wide3(d->s, R, ctx->sm_count, &f->wplan); }
  }
  f->veclen = momvec_len(d->s.ld, R);
  if (ctx->group && (f->wide || f->cols || seq || (size_t)f->veclen > XCHG_SLOT)) {
    fit_free(f);
    return fail(PPLS_E_UNSUPPORTED, "multi-device context: this shape takes the wide-row path, which exchanges through NCCL -- "
                                    "use one process per GPU (ppls_ctx_create_dist)");
  }
  if (!f->wide && !f->cols && !seq && (size_t)f->veclen <= XCHG_SLOT) {       // this fit exchanges through the peer inboxes
    rc = xchg_recover_dist(ctx);                                                // (collective, like this call)
    if (rc) { fit_free(f); return rc; }
  }
  {
    const char* ef = getenv("PPLS_FUSE_TAIL");
    // multi-GPU: every rank must take the same decision AND run the same grid (row CTA b of every rank owns the same rows
    // of the moment vector and exchanges them with CTA b of the peers), so the test uses global quantities only: the
    // smallest shard, floor(N / world) rows, must fill one tile per SM.
    const bool grid_ok = ctx->world == 1 ? f->plan.grid >= 32
                                         : (!f->wide && d->n_total / ctx->world >= (int64_t)ctx->sm_count * f->plan.T &&
                                            ctx->sm_count <= XCHG_MAXCTA);
    f->fused = !f->wide && !f->cols && !seq && opts->orth_type == PPLS_ORTH_SVD && grid_ok &&
               stream_tail_fits(f->plan, d->s) && !(ef && *ef && atoi(ef) == 0) &&
               (ctx->world == 1 || (ctx->xchg_on && (size_t)f->veclen <= XCHG_SLOT));
  }
  const size_t Lbytes = (size_t)d->s.ld * R * sizeof(double);
  e = cudaMalloc(&f->L, Lbytes);
  if (e == cudaSuccess && f->fused) {
    const size_t tb = stream_tail_scratch_doubles(f->plan) * sizeof(double);
    e = cudaMalloc(&f->tail_scratch, tb);
    if (e == cudaSuccess) e = cudaMemsetAsync(f->tail_scratch, 0, tb, ctx->stream);     // barrier words, clocks
    if (e == cudaSuccess) e = cudaMallocHost(&f->clk_host, 12 * sizeof(unsigned long long));
    if (e == cudaSuccess) memset(f->clk_host, 0, 12 * sizeof(unsigned long long));
  }
  if (e == cudaSuccess) e = cudaMalloc(&f->Lscratch, Lbytes);
  if (e == cudaSuccess && !f->wide) e = cudaMalloc(&f->partials, (size_t)f->plan.grid * f->veclen * sizeof(double));
  if (e == cudaSuccess && f->wide) {
    const size_t wb = wide_workspace_doubles(d->s, f->wplan, &f->woff) * sizeof(double);
    e = cudaMalloc(&f->wide_ws, wb);
    if (e == cudaSuccess) e = cudaMemsetAsync(f->wide_ws, 0, wb, ctx->stream);      // pad rows / components of the mu arrays stay zero
  }
  if (e == cudaSuccess) e = cudaMalloc(&f->momvec, (size_t)f->veclen * sizeof(double));
  if (e == cudaSuccess && !(getenv("PPLS_K2_SINGLE") && atoi(getenv("PPLS_K2_SINGLE"))))
  {
    const size_t ub = update_grid_scratch_doubles(f->cols ? f->sg.ld : d->s.ld, R) * sizeof(double);
    e = cudaMalloc(&f->upd_scratch, ub);
    if (e == cudaSuccess) e = cudaMemsetAsync(f->upd_scratch, 0, ub, ctx->stream);     // holds the row-block barrier counter
  }
  if (e == cudaSuccess && f->cols) e = cudaMalloc(&f->Gglob, (size_t)momvec_len(f->sg.ld, R) * sizeof(double));
  if (e == cudaSuccess && f->cols) e = cudaMalloc(&f->Lglob, (size_t)f->sg.ld * R * sizeof(double));
  if (e == cudaSuccess && f->cols) e = cudaMalloc(&f->Lscratch_g, (size_t)f->sg.ld * R * sizeof(double));
  if (e == cudaSuccess) e = cudaMalloc(&f->trace, (size_t)(opts->max_steps + 1) * sizeof(double));
  if (e == cudaSuccess) e = cudaMalloc(&f->st, sizeof(FitState));
  if (e == cudaSuccess) e = cudaMallocHost(&f->st_host, sizeof(FitState));
  if (e == cudaSuccess) e = cudaHostAlloc(&f->done_host, sizeof(int32_t), cudaHostAllocMapped);
  if (e != cudaSuccess) { fit_free(f); return fail(PPLS_E_OOM, std::string("fit workspace: ") + cudaGetErrorString(e)); }
  *f->done_host = 0;
  std::vector<double> L;
  pack_loadings(d->s, R, init->W, init->C, L);
  FitState* h = f->st_host;
  memset(h, 0, sizeof(FitState));
  for (int k = 0; k < R; ++k) { h->b[k] = init->B[k]; h->t[k] = init->sigT[k]; }
  h->sigE = init->sigE; h->sigF = init->sigF; h->sigH = init->sigH;
  h->ssqX = ssq_deflated ? ssq_deflated[0] : d->ssq[0]; h->ssqY = ssq_deflated ? ssq_deflated[1] : d->ssq[1];
  h->Ntot = (double)d->n_total;
  h->max_steps = opts->max_steps; h->orth_type = opts->orth_type; h->R = R; h->atol = opts->atol;
  {
    void* dp = nullptr;     // device alias of the mapped flag (same address under unified addressing)
    h->done_host = (cudaHostGetDevicePointer(&dp, f->done_host, 0) == cudaSuccess) ? (int32_t*)dp : nullptr;
  }
  e = cudaMemcpyAsync(f->L, L.data(), Lbytes, cudaMemcpyHostToDevice, ctx->stream);
  if (e == cudaSuccess) e = cudaMemcpyAsync(f->st, h, sizeof(FitState), cudaMemcpyHostToDevice, ctx->stream);
  if (e == cudaSuccess) e = cudaMemsetAsync(f->trace, 0, (size_t)(opts->max_steps + 1) * sizeof(double), ctx->stream);
  if (e == cudaSuccess && f->cols) {
    // assemble the global start loadings (each rank contributes its rows; the rest are zeros)
    e = cudaMemsetAsync(f->Lglob, 0, (size_t)f->sg.ld * R * sizeof(double), ctx->stream);
    const int ne = d->s.ld * R;
    if (e == cudaSuccess) rows_local_to_global<<<(ne + 255) / 256, 256, 0, ctx->stream>>>(f->L, f->Lglob, d->s.p2, d->s.q2, f->xoff, f->yoff_abs, R);
    if (e == cudaSuccess && allreduce_sum(ctx, f->Lglob, (size_t)f->sg.ld * R)) e = cudaErrorUnknown;
    if (e == cudaSuccess) e = launch_em_prepare(f->sg, R, f->Lglob, f->st, ctx->stream);
  } else if (e == cudaSuccess && seq) e = launch_seq_update(d->s, nullptr, f->L, f->st, f->trace, prev, nprev, 1, ctx->stream, &f->cons);
  else if (e == cudaSuccess) e = launch_em_prepare(d->s, R, f->L, f->st, ctx->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);   // L (host vector) must outlive the copy
  if (e != cudaSuccess) { fit_free(f); return fail(PPLS_E_CUDA, std::string("fit setup: ") + cudaGetErrorString(e)); }
  f->tm.kernel_launches = 1;
  f->tm.bytes_per_pass = (int64_t)((f->wide && !f->wplan.v3) ? 16 : 8) * d->s.n * ((int64_t)d->s.p + d->s.q);
  *out = f;
  return PPLS_OK;
}

static cudaEvent_t fit_event(ppls_fit* f) {
  if (!f->ev_free.empty()) { cudaEvent_t e = f->ev_free.back(); f->ev_free.pop_back(); return e; }
  cudaEvent_t e = nullptr;
  cudaEventCreate(&e);
  return e;
}

// read the control block back (after a stream sync) and fold finished timings
static int fit_poll(ppls_fit* f) {
  ppls_ctx* ctx = f->ctx;
  { int rc0 = ctx_activate(ctx); if (rc0) return rc0; }
  CU(cudaMemcpyAsync(f->st_host, f->st, sizeof(FitState), cudaMemcpyDeviceToHost, ctx->stream));
  if (f->fused)
    CU(cudaMemcpyAsync(f->clk_host, f->tail_scratch + stream_tail_clock_offset(f->plan), 12 * sizeof(unsigned long long),
                       cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  if (f->fused) {       // device clocks (globaltimer) of the fused launches: streaming phase | reduction + exchange + update
    f->tm.stream_phase_ms = (double)f->clk_host[0] * 1e-6;
    f->tm.tail_ms = (double)f->clk_host[1] * 1e-6;
    f->tm.scalar_ms = (double)f->clk_host[4] * 1e-6;
    if (getenv("PPLS_TAIL_VERBOSE") && f->clk_host[2])
      fprintf(stderr, "[ppls] fused tail per launch (us): streaming %.1f | tail %.1f = CTA 0: own sums %.1f + staging %.1f + scalar arithmetic %.1f (+ flag) ...\n",
              f->clk_host[0] * 1e-3 / f->clk_host[2], f->clk_host[1] * 1e-3 / f->clk_host[2], f->clk_host[5] * 1e-3 / f->clk_host[2],
              f->clk_host[6] * 1e-3 / f->clk_host[2], f->clk_host[7] * 1e-3 / f->clk_host[2]);
    if (getenv("PPLS_TAIL_VERBOSE") && f->clk_host[2] && f->clk_host[8])
      fprintf(stderr, "[ppls]   exchange inside 'own sums' (us): reduce + peer stores %.1f | fence + flags %.1f | wait for the peers %.1f | sum of the ranks %.1f\n",
              f->clk_host[8] * 1e-3 / f->clk_host[2], f->clk_host[9] * 1e-3 / f->clk_host[2], f->clk_host[10] * 1e-3 / f->clk_host[2],
              f->clk_host[11] * 1e-3 / f->clk_host[2]);
  }
  for (auto& pr : f->ev_tail) {       // .first is the K1 end event (owned by ev_k1, recycled below)
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, pr.first, pr.second) == cudaSuccess) f->tm.tail_ms += ms;
    f->ev_free.push_back(pr.second);
  }
  f->ev_tail.clear();
  for (auto& pr : f->ev_k1) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, pr.first, pr.second) == cudaSuccess) { f->tm.stream_ms += ms; f->tm.stream_launches += 1; }
    f->ev_free.push_back(pr.first); f->ev_free.push_back(pr.second);
  }
  f->ev_k1.clear();
  f->finished = f->st_host->done != 0;
  if (f->st_host->status != PPLS_OK && !(f->seq && f->st_host->status == PPLS_E_DEGENERATE))
    return fail(f->st_host->status, "a parameter became non-finite during EM (degenerate data or start values)");
  return PPLS_OK;
}

// Enqueue ONE data pass of fit `f` on its device (no synchronisation).
static int fit_enqueue_pass(ppls_fit* f) {
  int rc = ctx_activate(f->ctx);
  if (rc) return rc;
  ppls_ctx* ctx = f->ctx;
  const DataShape& s = f->data->s;
  const bool timing = (f->opts.flags & PPLS_F_TIMING) != 0;
  // row-sharded multi-GPU, fused-kernel path: the moment vector travels by peer stores (em_tail.cuh / em_update.cu)
  const bool use_x = ctx->xchg_on && !f->wide && !f->cols && !f->seq && (size_t)f->veclen <= XCHG_SLOT;
  cudaEvent_t a = nullptr, b = nullptr;
  if (timing) { a = fit_event(f); b = fit_event(f); cudaEventRecord(a, ctx->stream); }
  cudaError_t e;
  if (f->fused) {
    FusedTail ft{f->momvec, f->L, f->st, f->trace, f->tail_scratch, ctx->world > 1 ? &ctx->xdev : nullptr};
    e = launch_em_stream(f->plan, s, f->data->D, f->L, f->st, f->partials, nullptr, nullptr, 0, ctx->stream, &ft);
    if (timing) { cudaEventRecord(b, ctx->stream); f->ev_k1.emplace_back(a, b); }
    if (e != cudaSuccess) return fail(PPLS_E_CUDA, std::string("fused EM iteration: ") + cudaGetErrorString(e));
    f->tm.peer_exchange = ctx->world > 1 ? 1 : 0;
    f->tm.kernel_launches += 1;
    f->passes += 1;
    return PPLS_OK;
  }
  if (f->wide && f->wplan.v3) {
    e = launch_wide_single(f->wplan, s, f->data->D, f->L, f->st, f->wide_ws, f->woff, f->momvec, ctx->stream);
    if (timing) { cudaEventRecord(b, ctx->stream); f->ev_k1.emplace_back(a, b); }
  } else if (f->wide) {
    e = launch_wide_project(f->wplan, s, f->data->D, f->L, f->st, f->wide_ws, f->woff, ctx->stream);
    if (e == cudaSuccess && f->cols) {      // column-sharded: the N x 2r projection sums are what ranks exchange
      rc = allreduce_sum(ctx, f->wide_ws + f->woff.proj, (size_t)s.n * 2 * f->R);
      if (rc) return rc;
    }
    if (e == cudaSuccess) e = launch_wide_accum(f->wplan, s, f->data->D, f->L, f->st, f->wide_ws, f->woff, f->momvec, ctx->stream);
    if (timing) { cudaEventRecord(b, ctx->stream); f->ev_k1.emplace_back(a, b); }
    f->tm.kernel_launches += 3;      // 5 launches in this branch, 2 in the other; +3 more below
  } else {
    e = launch_em_stream(f->plan, s, f->data->D, f->L, f->st, f->partials, nullptr, nullptr, 0, ctx->stream);
    if (timing) { cudaEventRecord(b, ctx->stream); f->ev_k1.emplace_back(a, b); }
    if (e == cudaSuccess && use_x) e = launch_xsend(f->partials, f->plan.grid, f->veclen, ctx->xdev, f->st, ctx->stream);
    else if (e == cudaSuccess) e = launch_reduce_partials(f->partials, f->plan.grid, f->veclen, f->momvec, f->st, ctx->stream);
  }
  if (e != cudaSuccess) return fail(PPLS_E_CUDA, std::string("streaming pass: ") + cudaGetErrorString(e));
  if (f->cols) {
    // global G = every rank's rows (zeros elsewhere) summed; the Gram tail is identical on all ranks
    // (it comes from the exchanged projections), so rank 0 alone contributes it
    const int R_ = f->R, ne = s.ld * R_;
    const size_t glen = (size_t)f->sg.ld * R_, tail = (size_t)3 * R_ * R_;
    e = cudaMemsetAsync(f->Gglob, 0, (glen + tail) * sizeof(double), ctx->stream);
    if (e == cudaSuccess) rows_local_to_global<<<(ne + 255) / 256, 256, 0, ctx->stream>>>(f->momvec, f->Gglob, s.p2, s.q2, f->xoff, f->yoff_abs, R_);
    if (e == cudaSuccess && ctx->rank == 0)
      e = cudaMemcpyAsync(f->Gglob + glen, f->momvec + (size_t)s.ld * R_, tail * sizeof(double), cudaMemcpyDeviceToDevice, ctx->stream);
    if (e != cudaSuccess) return fail(PPLS_E_CUDA, std::string("column-sharded gather: ") + cudaGetErrorString(e));
    rc = allreduce_sum(ctx, f->Gglob, glen + tail);
    if (rc) return rc;
    e = launch_em_update(f->sg, R_, f->Gglob, f->Lglob, f->Lscratch_g, f->st, f->trace, f->opts.orth_type, f->upd_scratch, ctx->stream);
    if (e == cudaSuccess) rows_global_to_local<<<(ne + 255) / 256, 256, 0, ctx->stream>>>(f->Lglob, f->L, s.p2, s.q2, f->xoff, f->yoff_abs, R_);
    if (e == cudaSuccess) e = cudaGetLastError();
    f->tm.kernel_launches += 2;
  } else if (use_x) {
    e = launch_em_update(s, f->R, f->momvec, f->L, f->Lscratch, f->st, f->trace, f->opts.orth_type, f->upd_scratch, ctx->stream, &ctx->xdev);
  } else {
    rc = allreduce_sum(ctx, f->momvec, (size_t)f->veclen);
    if (rc) return rc;
    if (f->seq) e = launch_seq_update(s, f->momvec, f->L, f->st, f->trace, f->prev, f->nprev, 0, ctx->stream, &f->cons);
    else e = launch_em_update(s, f->R, f->momvec, f->L, f->Lscratch, f->st, f->trace, f->opts.orth_type, f->upd_scratch, ctx->stream);
  }
  if (e != cudaSuccess) return fail(PPLS_E_CUDA, std::string("update kernel: ") + cudaGetErrorString(e));
  if (timing) { cudaEvent_t c = fit_event(f); cudaEventRecord(c, ctx->stream); f->ev_tail.emplace_back(b, c); }
  f->tm.peer_exchange = use_x ? 1 : 0;
  f->tm.kernel_launches += 3;
  f->passes += 1;
  return PPLS_OK;
}

// A fit on a multi-device context is one part per device, driven in lockstep by this host thread: pass k is
// enqueued on every device before pass k+1 (the launches are asynchronous; the kernels of a pass meet in the peer
// exchange).  A plain fit is its own single part.
static ppls_fit** fit_parts(ppls_fit* f, ppls_fit** self, size_t* n) {
  if (f->parts.empty()) { *self = f; *n = 1; return self; }
  *n = f->parts.size();
  return f->parts.data();
}

extern "C" int ppls_fit_iterate(ppls_fit* f, int32_t n_passes, int32_t* passes_done, int32_t* finished) {
  if (!f) return fail(PPLS_E_INVALID, "null fit handle");
  ppls_fit* self = nullptr;
  size_t np = 0;
  ppls_fit** parts = fit_parts(f, &self, &np);
  ppls_fit* lead = parts[0];
  int rc;
  const bool timing = (lead->opts.flags & PPLS_F_TIMING) != 0;
  const int check_every = lead->opts.check_every > 0 ? lead->opts.check_every : 16;
  std::vector<cudaEvent_t> ev0(np, nullptr), ev1(np, nullptr);
  if (timing)
    for (size_t i = 0; i < np; ++i) {
      rc = ctx_activate(parts[i]->ctx);
      if (rc) return rc;
      ev0[i] = fit_event(parts[i]); ev1[i] = fit_event(parts[i]);
      cudaEventRecord(ev0[i], parts[i]->ctx->stream);
    }
  int32_t since_check = 0;
  // passes the device may still run: it stops by itself (done flag), the host only bounds the queue
  for (int32_t i = 0; i < n_passes && !lead->finished; ++i) {
    if (lead->passes > lead->opts.max_steps) break;
    if (*reinterpret_cast<volatile int32_t*>(lead->done_host)) break;      // the device has ended the run: stop queueing
    for (size_t k = 0; k < np; ++k) {
      rc = fit_enqueue_pass(parts[k]);
      if (rc) return rc;
    }
    if (++since_check >= check_every) {
      for (size_t k = 0; k < np; ++k) { rc = fit_poll(parts[k]); if (rc) return rc; }
      since_check = 0;
    }
  }
  for (size_t k = 0; k < np; ++k) {
    rc = ctx_activate(parts[k]->ctx);
    if (rc) return rc;
    if (timing) cudaEventRecord(ev1[k], parts[k]->ctx->stream);
  }
  for (size_t k = 0; k < np; ++k) {
    ppls_fit* g = parts[k];
    rc = ctx_activate(g->ctx);
    if (rc) return rc;
    rc = fit_poll(g);
    if (rc) return rc;
    if (timing) {
      float ms = 0.f;
      if (cudaEventElapsedTime(&ms, ev0[k], ev1[k]) == cudaSuccess) g->tm.loop_ms += ms;
      g->ev_free.push_back(ev0[k]); g->ev_free.push_back(ev1[k]);
    }
    // passes actually executed on the device (later launches were no-ops once done was set)
    g->tm.passes = g->st_host->done ? g->st_host->steps_done + 1 : g->st_host->pass;
  }
  if (f != lead) { f->finished = lead->finished; f->passes = lead->passes; }
  if (passes_done) *passes_done = (int32_t)lead->tm.passes;
  if (finished) *finished = lead->finished ? 1 : 0;
  return PPLS_OK;
}

static int fit_result_impl(ppls_fit* f, ppls_params* est, double* loglik, int32_t* steps_done, double* moments,
                           double* mu_T, double* mu_U, int64_t ldmu);
extern "C" int ppls_fit_result(ppls_fit* f, ppls_params* est, double* loglik, int32_t* steps_done, double* moments,
                               double* mu_T, double* mu_U) {
  if (!f) return fail(PPLS_E_INVALID, "null fit handle");
  if (!f->parts.empty()) {
    // parameters, trace and moments are identical on every device (rank-ordered sums): take them from the first;
    // the posterior means are per row: every device writes its rows of the caller's n x r arrays
    if ((mu_T == nullptr) != (mu_U == nullptr)) return fail(PPLS_E_INVALID, "pass both mu_T and mu_U or neither");
    int rc = fit_result_impl(f->parts[0], est, loglik, steps_done, moments, nullptr, nullptr, 0);
    if (rc || !mu_T) return rc;
    const int64_t N = f->data->s.n;
    for (size_t i = 0; i < f->parts.size(); ++i) {
      const int64_t r0 = f->data->shard_row0[i];
      rc = fit_result_impl(f->parts[i], nullptr, nullptr, nullptr, nullptr, mu_T + r0, mu_U + r0, N);
      if (rc) return rc;
    }
    return PPLS_OK;
  }
  return fit_result_impl(f, est, loglik, steps_done, moments, mu_T, mu_U, f->data->s.n);
}
// mu_T / mu_U: this fit's rows, column-major with leading dimension ldmu
static int fit_result_impl(ppls_fit* f, ppls_params* est, double* loglik, int32_t* steps_done, double* moments,
                           double* mu_T, double* mu_U, int64_t ldmu) {
  int rc = ctx_activate(f->ctx);
  if (rc) return rc;
  ppls_ctx* ctx = f->ctx;
  const DataShape& s = f->data->s;
  const int R = f->R;
  rc = fit_poll(f);
  if (rc) return rc;
  const FitState* h = f->st_host;
  if (!f->finished) return fail(PPLS_E_INVALID, "the run has not ended yet: call ppls_fit_iterate until finished");
  if ((mu_T == nullptr) != (mu_U == nullptr)) return fail(PPLS_E_INVALID, "pass both mu_T and mu_U or neither");
  if (est) {
    rc = check_params(est, "est");
    if (rc) return rc;
    std::vector<double> L((size_t)s.ld * R);
    CU(cudaMemcpyAsync(L.data(), f->L, L.size() * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    for (int k = 0; k < R; ++k) {
      for (int j = 0; j < s.p; ++j) est->W[j + (size_t)k * s.p] = L[(size_t)j * R + k];
      for (int j = 0; j < s.q; ++j) est->C[j + (size_t)k * s.q] = L[(size_t)(s.p2 + j) * R + k];
      est->B[k] = h->b[k]; est->sigT[k] = h->t[k];
    }
    est->sigE = h->sigE; est->sigF = h->sigF; est->sigH = h->sigH;
  }
  if (steps_done) *steps_done = h->steps_done;
  if (loglik && h->steps_done > 0) {
    CU(cudaMemcpyAsync(loglik, f->trace, (size_t)h->steps_done * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
  }
  if (moments) {
    double* m = moments;
    for (int k = 0; k < R; ++k) m[k] = h->Ctt[k];
    for (int k = 0; k < R; ++k) m[R + k] = h->Cuu[k];
    for (int k = 0; k < R; ++k) m[2 * R + k] = h->Cut[k];
    m[3 * R] = h->Cee; m[3 * R + 1] = h->Cff;
    for (int e = 0; e < R * R; ++e) m[3 * R + 2 + e] = h->Chh[e];
  }
  if (mu_T && f->wide) {
    // the last pass left mu_T, mu_U of the final parameters in the workspace ([2][n][R], row-major)
    const int64_t npad = wide_npad(s.n);
    std::vector<double> mu((size_t)2 * npad * R);
    CU(cudaMemcpyAsync(mu.data(), f->wide_ws + f->woff.mu, mu.size() * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    for (int64_t i = 0; i < s.n; ++i)
      for (int k = 0; k < R; ++k) {
        mu_T[i + (size_t)k * ldmu] = mu[(size_t)i * R + k];
        mu_U[i + (size_t)k * ldmu] = mu[(size_t)(npad + i) * R + k];
      }
  } else if (mu_T) {
    // One more pass with the final parameters writes the posterior means (EM_W_multi.R:802).
    double* dmu = nullptr;
    CU(cudaMalloc(&dmu, (size_t)2 * s.n * R * sizeof(double)));
    cudaError_t e = launch_em_stream(f->plan, s, f->data->D, f->L, f->st, f->partials, dmu, dmu + (size_t)s.n * R, 1, ctx->stream);
    f->tm.kernel_launches += 1;
    if (e == cudaSuccess)
      e = cudaMemcpy2DAsync(mu_T, (size_t)ldmu * sizeof(double), dmu, (size_t)s.n * sizeof(double), (size_t)s.n * sizeof(double), (size_t)R,
                            cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess)
      e = cudaMemcpy2DAsync(mu_U, (size_t)ldmu * sizeof(double), dmu + (size_t)s.n * R, (size_t)s.n * sizeof(double),
                            (size_t)s.n * sizeof(double), (size_t)R, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    cudaFree(dmu);
    if (e != cudaSuccess) return fail(PPLS_E_CUDA, std::string("posterior-means pass: ") + cudaGetErrorString(e));
  }
  return PPLS_OK;
}

extern "C" int ppls_fit_timing(const ppls_fit* f, ppls_timing* out) {
  if (!f || !out) return fail(PPLS_E_INVALID, "null argument");
  if (!f->parts.empty()) {            // multi-device: the slowest device's times, the launches of all
    *out = f->parts[0]->tm;
    for (size_t i = 1; i < f->parts.size(); ++i) {
      const ppls_timing& t = f->parts[i]->tm;
      out->stream_ms = std::max(out->stream_ms, t.stream_ms); out->loop_ms = std::max(out->loop_ms, t.loop_ms);
      out->tail_ms = std::max(out->tail_ms, t.tail_ms); out->stream_phase_ms = std::max(out->stream_phase_ms, t.stream_phase_ms);
      out->scalar_ms = std::max(out->scalar_ms, t.scalar_ms);
      out->kernel_launches += t.kernel_launches; out->bytes_per_pass += t.bytes_per_pass;
    }
    return PPLS_OK;
  }
  *out = f->tm;
  return PPLS_OK;
}

extern "C" int ppls_fit_destroy(ppls_fit* f) {
  fit_free(f);
  return PPLS_OK;
}

extern "C" int ppls_simult_run(ppls_ctx* ctx, const ppls_data* d, const ppls_opts* opts, const ppls_params* init,
                               ppls_params* est, double* loglik, int32_t* steps_done, double* moments, double* mu_T,
                               double* mu_U) {
  ppls_fit* f = nullptr;
  int rc = ppls_fit_create(ctx, d, opts, init, &f);
  if (rc) return rc;
  int32_t done = 0, fin = 0;
  while (!fin) {
    rc = ppls_fit_iterate(f, opts->max_steps + 1, &done, &fin);
    if (rc) { fit_free(f); return rc; }
    if (!fin && done > opts->max_steps) { fit_free(f); return fail(PPLS_E_CUDA, "EM loop did not terminate"); }
  }
  rc = ppls_fit_result(f, est, loglik, steps_done, moments, mu_T, mu_U);
  fit_free(f);
  return rc;
}

// ---- batched multi-start (BASELINE config 5) ------------------------------------------------------------
// S start states over ONE resident dataset.  Starts are packed g = floor(8 / r) at a time into the
// streaming kernel's component dimension (the E-step is diagonal in the components, so a pack of starts
// is a wider W): each data pass serves g starts, the update runs one block per start.
extern "C" int ppls_simult_run_batched(ppls_ctx* ctx, const ppls_data* d, const ppls_opts* opts, int32_t n_starts,
                                       const ppls_params* inits, ppls_params* ests, double* loglik,
                                       int32_t* steps_done, ppls_timing* timing) {
  int rc = ctx_activate(ctx);
  if (rc) return rc;
  if (!d || !opts || !inits || !ests || !steps_done) return fail(PPLS_E_INVALID, "null argument");
  if (d->ctx != ctx) return fail(PPLS_E_INVALID, "data belongs to another context");
  const int R = opts->r, S = n_starts;
  if (R < 1 || R > MAXR) return fail(PPLS_E_INVALID, "nr_comp must be in 1..8");
  if (S < 1) return fail(PPLS_E_INVALID, "need at least one start");
  if (R > d->s.p || R > d->s.q) return fail(PPLS_E_INVALID, "nr_comp exceeds ncol(X) or ncol(Y)");
  if (opts->max_steps < 0) return fail(PPLS_E_INVALID, "EMsteps must be >= 0");
  for (int s = 0; s < S; ++s) { rc = check_params(&inits[s], "init"); if (rc) return rc; rc = check_params(&ests[s], "est"); if (rc) return rc; }
  const DataShape& sh = d->s;
  // Two data paths: many components in flight (S*r >= 64) -> two FP64 GEMMs per pass over groups of up to 128
  // components (em_batch.cu); otherwise packs of <= 8 components through the streaming kernel K1.
  const char* eg = getenv("PPLS_BATCH_GEMM");
  const bool gemm = !sh.panel && ((eg && *eg) ? atoi(eg) != 0 : (int64_t)S * R >= 64);
  int g, RP, groups;
  if (gemm) {
    g = (128 / R) < S ? (128 / R) : S;                   // starts per group
    RP = ((g * R + 127) / 128) * 128;                    // the GEMM tile is 128 components wide (pad columns are zero)
  } else {
    g = (MAXR / R) < S ? (MAXR / R) : S; RP = g * R;
  }
  groups = (S + g - 1) / g;
  StreamPlan plan{};
  BGemmPlan bplan{};
  const char* why = "";
  if (gemm) plan_bgemm(sh, RP, R, g, ctx->sm_count, &bplan);
  else if (plan_stream(sh, RP, ctx->sm_count, &plan, &why) != cudaSuccess)
    return fail(PPLS_E_UNSUPPORTED, std::string("batched run: streaming-kernel plan: ") + why);
  const size_t veclen = (size_t)momvec_len(sh.ld, R), veclenP = (size_t)momvec_len(sh.ld, RP);
  const int tracelen = opts->max_steps + 1;
  struct Bufs {
    double *Lp = nullptr, *Ls = nullptr, *Lsc = nullptr, *mp = nullptr, *ms = nullptr, *tr = nullptr, *part = nullptr, *coef = nullptr;
    FitState *pst = nullptr, *st = nullptr;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    ~Bufs() { cudaFree(Lp); cudaFree(Ls); cudaFree(Lsc); cudaFree(mp); cudaFree(ms); cudaFree(tr); cudaFree(part); cudaFree(coef); cudaFree(pst); cudaFree(st);
              if (e0) cudaEventDestroy(e0); if (e1) cudaEventDestroy(e1); }
  } B;
  CU(cudaMalloc(&B.Lp, (size_t)groups * sh.ld * RP * sizeof(double)));
  CU(cudaMalloc(&B.Ls, (size_t)S * sh.ld * R * sizeof(double)));
  CU(cudaMalloc(&B.Lsc, (size_t)S * sh.ld * R * sizeof(double)));
  CU(cudaMalloc(&B.mp, (size_t)groups * veclenP * sizeof(double)));
  CU(cudaMalloc(&B.ms, (size_t)S * veclen * sizeof(double)));
  CU(cudaMalloc(&B.tr, (size_t)S * tracelen * sizeof(double)));
  if (gemm) {
    CU(cudaMalloc(&B.part, bgemm_workspace_doubles(sh, bplan) * sizeof(double)));
    CU(cudaMalloc(&B.coef, (size_t)groups * 4 * RP * sizeof(double)));
    CU(cudaMemsetAsync(B.coef, 0, (size_t)groups * 4 * RP * sizeof(double), ctx->stream));
    CU(cudaMemsetAsync(B.mp, 0, (size_t)groups * veclenP * sizeof(double), ctx->stream));
  } else CU(cudaMalloc(&B.part, (size_t)plan.grid * veclenP * sizeof(double)));
  CU(cudaMalloc(&B.pst, (size_t)groups * sizeof(FitState)));
  CU(cudaMalloc(&B.st, (size_t)S * sizeof(FitState)));
  CU(cudaEventCreate(&B.e0)); CU(cudaEventCreate(&B.e1));
  std::vector<FitState> hst(S);
  std::vector<double> Lh((size_t)S * sh.ld * R), L1;
  for (int s = 0; s < S; ++s) {
    FitState& h = hst[s];
    memset(&h, 0, sizeof(FitState));
    for (int k = 0; k < R; ++k) { h.b[k] = inits[s].B[k]; h.t[k] = inits[s].sigT[k]; }
    h.sigE = inits[s].sigE; h.sigF = inits[s].sigF; h.sigH = inits[s].sigH;
    h.ssqX = d->ssq[0]; h.ssqY = d->ssq[1]; h.Ntot = (double)d->n_total;
    h.max_steps = opts->max_steps; h.orth_type = opts->orth_type; h.R = R; h.atol = opts->atol;
    pack_loadings(sh, R, inits[s].W, inits[s].C, L1);
    memcpy(&Lh[(size_t)s * sh.ld * R], L1.data(), L1.size() * sizeof(double));
  }
  cudaStream_t st = ctx->stream;
  CU(cudaMemcpyAsync(B.st, hst.data(), (size_t)S * sizeof(FitState), cudaMemcpyHostToDevice, st));
  CU(cudaMemcpyAsync(B.Ls, Lh.data(), Lh.size() * sizeof(double), cudaMemcpyHostToDevice, st));
  CU(cudaMemsetAsync(B.Lp, 0, (size_t)groups * sh.ld * RP * sizeof(double), st));
  CU(cudaMemsetAsync(B.pst, 0, (size_t)groups * sizeof(FitState), st));
  CU(cudaMemsetAsync(B.tr, 0, (size_t)S * tracelen * sizeof(double), st));
  CU(launch_em_update_batch(sh, R, RP, g, S, tracelen, 1, B.mp, B.Lp, B.pst, B.st, B.ms, B.Ls, B.Lsc, B.tr, st, B.coef));
  CU(cudaStreamSynchronize(st));            // host vectors must outlive the copies
  const int check_every = opts->check_every > 0 ? opts->check_every : 16;
  int64_t launches = 1, passes = 0;
  CU(cudaEventRecord(B.e0, st));
  for (int pass = 0; pass <= opts->max_steps; ++pass) {
    for (int gr = 0; gr < groups; ++gr) {
      if (gemm) {
        BGemmPlan bp = bplan;
        bp.nstart = (S - gr * g < g) ? S - gr * g : g;
        CU(launch_bgemm_pass(bp, sh, d->D, B.Lp + (size_t)gr * sh.ld * RP, B.coef + (size_t)gr * 4 * RP, B.part,
                             B.mp + (size_t)gr * veclenP, st));
        continue;
      }
      CU(launch_em_stream(plan, sh, d->D, B.Lp + (size_t)gr * sh.ld * RP, B.pst + gr, B.part, nullptr, nullptr, 0, st));
      CU(launch_reduce_partials(B.part, plan.grid, (int64_t)veclenP, B.mp + (size_t)gr * veclenP, B.pst + gr, st));
    }
    rc = allreduce_sum(ctx, B.mp, (size_t)groups * veclenP);
    if (rc) return rc;
    CU(launch_em_update_batch(sh, R, RP, g, S, tracelen, 0, B.mp, B.Lp, B.pst, B.st, B.ms, B.Ls, B.Lsc, B.tr, st, B.coef));
    launches += (gemm ? (bplan.fused ? 3 : 4) : 2) * groups + (gemm ? 1 : 2); passes += 1;
    if ((pass + 1) % check_every == 0 && pass < opts->max_steps) {
      CU(cudaMemcpyAsync(hst.data(), B.st, (size_t)S * sizeof(FitState), cudaMemcpyDeviceToHost, st));
      CU(cudaStreamSynchronize(st));
      bool all = true;
      for (int s = 0; s < S; ++s) all = all && hst[s].done;
      if (all) break;
    }
  }
  CU(cudaEventRecord(B.e1, st));
  CU(cudaMemcpyAsync(hst.data(), B.st, (size_t)S * sizeof(FitState), cudaMemcpyDeviceToHost, st));
  CU(cudaMemcpyAsync(Lh.data(), B.Ls, Lh.size() * sizeof(double), cudaMemcpyDeviceToHost, st));
  std::vector<double> trh((size_t)S * tracelen);
  CU(cudaMemcpyAsync(trh.data(), B.tr, trh.size() * sizeof(double), cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  for (int s = 0; s < S; ++s) {
    const FitState& h = hst[s];
    if (h.status != PPLS_OK) return fail(h.status, "a parameter became non-finite during EM (batched run)");
    const double* L = &Lh[(size_t)s * sh.ld * R];
    for (int k = 0; k < R; ++k) {
      for (int j = 0; j < sh.p; ++j) ests[s].W[j + (size_t)k * sh.p] = L[(size_t)j * R + k];
      for (int j = 0; j < sh.q; ++j) ests[s].C[j + (size_t)k * sh.q] = L[(size_t)(sh.p2 + j) * R + k];
      ests[s].B[k] = h.b[k]; ests[s].sigT[k] = h.t[k];
    }
    ests[s].sigE = h.sigE; ests[s].sigF = h.sigF; ests[s].sigH = h.sigH;
    steps_done[s] = h.steps_done;
    if (loglik) for (int i = 0; i < h.steps_done; ++i) loglik[(size_t)s * opts->max_steps + i] = trh[(size_t)s * tracelen + i];
  }
  if (timing) {
    float ms = 0.f;
    cudaEventElapsedTime(&ms, B.e0, B.e1);
    memset(timing, 0, sizeof(*timing));
    timing->loop_ms = ms; timing->kernel_launches = launches; timing->passes = passes;
    timing->stream_launches = passes * groups;
    timing->bytes_per_pass = (int64_t)8 * sh.n * ((int64_t)sh.p + sh.q);
  }
  return PPLS_OK;
}


// ---- sequential fit: PPLS(X, Y, nr_comp, EMsteps, atol, ...), the start-value path of PPLS_simult ------------
// Package/PPLS/R/EM_W_multi.R:229-279 (PPLS) / :116-180 (PPLSi) / :51-74 (EMstep_W) /
// Package/PPLS/src/loglC.cpp:341-397 (EMstepC_fast).  Components one at a time; the deflation of :270-271 is
// implicit (no deflated copy of X or Y): see csrc/em_seq.cu.  critfunc = identity, no constraints.
extern "C" int ppls_sequential_run(ppls_ctx* ctx, const ppls_data* d, int32_t nr_comp, int32_t EMsteps, double atol,
                                   const ppls_params* guesses, ppls_params* est, double* sig, double* logvalues,
                                   int32_t* steps, double* last_incr, int32_t* comps_done) {
  return ppls_sequential_run_ex(ctx, d, nr_comp, EMsteps, atol, guesses, nullptr, 0, est, sig, logvalues, steps, last_incr, comps_done);
}

// ... with PPLSi's `constraints` (fconstraint, EM_W_multi.R:85-92; applied to the start, :142-146, and after every EM step,
// :160-167) and `critfunc` (:172; identity or abs).
extern "C" int ppls_sequential_run_ex(ppls_ctx* ctx, const ppls_data* d, int32_t nr_comp, int32_t EMsteps, double atol,
                                      const ppls_params* guesses, const ppls_constraints* constraints, int32_t crit_abs,
                                      ppls_params* est, double* sig, double* logvalues, int32_t* steps, double* last_incr,
                                      int32_t* comps_done) {
  int rc = ctx_activate(ctx);
  if (rc) return rc;
  if (!d || !guesses || !est || !sig || !steps || !comps_done) return fail(PPLS_E_INVALID, "null argument");
  if (d->ctx != ctx) return fail(PPLS_E_INVALID, "data belongs to another context");
  if (nr_comp <= 0) return fail(PPLS_E_INVALID, "#components must be >0");
  if (nr_comp > d->s.p || nr_comp > d->s.q) return fail(PPLS_E_INVALID, "nr_comp exceeds ncol(X) or ncol(Y)");
  if (EMsteps < 1) return fail(PPLS_E_INVALID, "EMsteps must be a strictly positive integer");
  rc = check_params(est, "est");
  if (rc) return rc;
  const DataShape& sh = d->s;
  const int a = nr_comp;
  double* dprev = nullptr;
  CU(cudaMalloc(&dprev, (size_t)a * sh.ld * sizeof(double)));
  struct Guard { double* p; ~Guard() { cudaFree(p); } } guard{dprev};
  double ssq[2] = {d->ssq[0], d->ssq[1]};
  ppls_opts opts{};
  opts.r = 1; opts.max_steps = EMsteps; opts.atol = atol; opts.orth_type = PPLS_ORTH_SVD; opts.check_every = 8;
  *comps_done = 0;
  double* dcons = nullptr;                      // fixed loadings of the current component, engine layout [ld]
  CU(cudaMalloc(&dcons, (size_t)sh.ld * sizeof(double)));
  struct Guard2 { double* p; ~Guard2() { cudaFree(p); } } guard2{dcons};
  for (int i = 0; i < a; ++i) {
    rc = check_params(&guesses[i], "guess");
    if (rc) return rc;
    // constraints of this component: fixed values replace the start (:142-146) and every later estimate (:160-167)
    SeqCons sc{};
    sc.crit_abs = crit_abs ? 1 : 0;
    ppls_params g = guesses[i];
    std::vector<double> gW, gC;
    double gB = g.B[0], gT = g.sigT[0];
    if (constraints && constraints[i].mask) {
      const ppls_constraints& c = constraints[i];
      if (c.mask & ~127u) return fail(PPLS_E_INVALID, "unknown constraint bits");
      if (((c.mask & PPLS_FIX_W) && !c.W) || ((c.mask & PPLS_FIX_C) && !c.C)) return fail(PPLS_E_INVALID, "constraint on W / C without values");
      sc.mask = c.mask; sc.B = c.B; sc.sigE = c.sigE; sc.sigF = c.sigF; sc.sigH = c.sigH; sc.sigT = c.sigT;
      if (c.mask & (PPLS_FIX_W | PPLS_FIX_C)) {
        // The deflation is implicit here (the data is never rewritten): that is the reference's Xc - (Xc w) w' only for
        // unit-norm loadings orthogonal to the earlier components -- which EM iterates are by construction, and which a
        // fixed loading vector has to be.
        std::vector<double> Lc((size_t)sh.ld, 0.0);
        for (int side = 0; side < 2; ++side) {
          if (!(c.mask & (side ? PPLS_FIX_C : PPLS_FIX_W))) continue;
          const double* v = side ? c.C : c.W;
          const int len = side ? sh.q : sh.p;
          double nn = 0.0;
          for (int j = 0; j < len; ++j) nn += v[j] * v[j];
          bool ok = fabs(nn - 1.0) < 1e-10;
          for (int k = 0; k < i && ok; ++k) {
            const double* pk = (side ? est->C : est->W) + (size_t)k * len;
            double dp = 0.0;
            for (int j = 0; j < len; ++j) dp += pk[j] * v[j];
            ok = fabs(dp) < 1e-10;
          }
          if (!ok) return fail(PPLS_E_UNSUPPORTED, "fixed loadings must have unit norm and be orthogonal to the earlier components "
                                                   "(the deflation of the sequential fit is applied implicitly)");
          for (int j = 0; j < len; ++j) Lc[(size_t)(side ? sh.p2 : 0) + j] = v[j];
        }
        CU(cudaMemcpyAsync(dcons, Lc.data(), Lc.size() * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
        CU(cudaStreamSynchronize(ctx->stream));
        sc.L = dcons;
      }
      if (c.mask & PPLS_FIX_W) g.W = const_cast<double*>(c.W);
      if (c.mask & PPLS_FIX_C) g.C = const_cast<double*>(c.C);
      if (c.mask & PPLS_FIX_B) gB = c.B;
      if (c.mask & PPLS_FIX_SIGE) g.sigE = c.sigE;
      if (c.mask & PPLS_FIX_SIGF) g.sigF = c.sigF;
      if (c.mask & PPLS_FIX_SIGH) g.sigH = c.sigH;
      if (c.mask & PPLS_FIX_SIGT) gT = c.sigT;
      g.B = &gB; g.sigT = &gT;
    }
    ppls_fit* f = nullptr;
    rc = fit_create_impl(ctx, d, &opts, &g, true, dprev, i, ssq, &f, &sc);
    if (rc) return rc;
    int32_t done = 0, fin = 0;
    while (!fin) {
      rc = ppls_fit_iterate(f, 8, &done, &fin);
      if (rc) { fit_free(f); return rc; }
      if (!fin && done > EMsteps + 1) { fit_free(f); return fail(PPLS_E_CUDA, "sequential EM loop did not terminate"); }
    }
    const FitState hs = *f->st_host;      // by value: the pinned copy goes away with the fit
    const FitState* h = &hs;
    if (h->status == PPLS_E_DEGENERATE) { fit_free(f); break; }     // :257-262: stop at this component, keep the earlier ones
    const int k = h->steps_done;
    std::vector<double> L(sh.ld), tr((size_t)k + 1);
    // Gram of the pass that ended the fit, as the update kernel recorded it (NOT momvec: the host may have queued
    // further passes, which are no-ops on the device but re-run the in-place allreduce of momvec on world > 1)
    const double gram[3] = {h->gram_last[0], h->gram_last[1], h->gram_last[2]};
    cudaError_t e = cudaMemcpyAsync(L.data(), f->L, sh.ld * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(tr.data(), f->trace, tr.size() * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(dprev + (size_t)i * sh.ld, f->L, sh.ld * sizeof(double), cudaMemcpyDeviceToDevice, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    fit_free(f);
    if (e != cudaSuccess) return fail(PPLS_E_CUDA, std::string("sequential fit readback: ") + cudaGetErrorString(e));
    for (int j = 0; j < sh.p; ++j) est->W[j + (size_t)i * sh.p] = (sc.mask & PPLS_FIX_W) ? g.W[j] : L[j];
    for (int j = 0; j < sh.q; ++j) est->C[j + (size_t)i * sh.q] = (sc.mask & PPLS_FIX_C) ? g.C[j] : L[sh.p2 + j];
    est->B[i] = h->b[0]; est->sigT[i] = h->t[0];
    sig[i + 0 * a] = h->sigE; sig[i + 1 * a] = h->sigF; sig[i + 2 * a] = h->sigH; sig[i + 3 * a] = h->t[0];
    steps[i] = k;
    if (last_incr) last_incr[i] = (k >= 1) ? tr[k] - tr[k - 1] : 0.0;
    if (logvalues) for (int j = 0; j <= k; ++j) logvalues[(size_t)i * (EMsteps + 1) + j] = tr[j];
    // Xc <- Xc - (Xc w) w':  ||Xc||^2 drops by ||Xc w||^2 (w has unit norm), the Gram of the final pass
    ssq[0] -= gram[0]; ssq[1] -= gram[2];
    *comps_done = i + 1;
  }
  if (*comps_done > 0) { est->sigE = sig[*comps_done - 1]; est->sigF = sig[*comps_done - 1 + a]; est->sigH = sig[*comps_done - 1 + 2 * a]; }
  return PPLS_OK;
}

// ---- start state of one component: PPLSi's initialGuess "equal" / "random" (EM_W_multi.R:126-141) -----------------------
extern "C" int ppls_initial_guess(ppls_ctx* ctx, int32_t kind, int64_t p, int64_t q, uint64_t seed, int32_t component,
                                  ppls_params* out) {
  if (ctx && !ctx->subs.empty()) ctx = ctx->subs[0];
  int rc = ctx_activate(ctx);
  if (rc) return rc;
  rc = check_params(out, "out");
  if (rc) return rc;
  if (p < 1 || q < 1 || p > (1 << 24) || q > (1 << 24)) return fail(PPLS_E_INVALID, "p, q out of range");
  if (component < 0) return fail(PPLS_E_INVALID, "negative component index");
  if (kind == PPLS_GUESS_EQUAL) {
    for (int64_t j = 0; j < p; ++j) out->W[j] = 1.0 / sqrt((double)p);
    for (int64_t j = 0; j < q; ++j) out->C[j] = 1.0 / sqrt((double)q);
    out->B[0] = 1.0; out->sigT[0] = 1.0; out->sigH = 1.0; out->sigE = 1.0 / (double)p; out->sigF = 1.0 / (double)q;
    return PPLS_OK;
  }
  if (kind != PPLS_GUESS_RANDOM) return fail(PPLS_E_INVALID, "initialGuess must be 'equal' or 'random' here ('o2m': pass a custom guess)");
  double* d = nullptr;
  CU(cudaMalloc(&d, (size_t)(p + q + 5) * sizeof(double)));
  cudaError_t e = launch_guess(d, (int)p, d + p, (int)q, d + p + q, seed, (uint32_t)component, ctx->stream);
  double sc[5] = {0, 0, 0, 0, 0};
  if (e == cudaSuccess) e = cudaMemcpyAsync(out->W, d, (size_t)p * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream);
  if (e == cudaSuccess) e = cudaMemcpyAsync(out->C, d + p, (size_t)q * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream);
  if (e == cudaSuccess) e = cudaMemcpyAsync(sc, d + p + q, sizeof(sc), cudaMemcpyDeviceToHost, ctx->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
  cudaFree(d);
  if (e != cudaSuccess) return fail(PPLS_E_CUDA, std::string("initial guess: ") + cudaGetErrorString(e));
  out->B[0] = sc[0]; out->sigE = sc[1]; out->sigF = sc[2]; out->sigH = sc[3]; out->sigT[0] = sc[4];
  return PPLS_OK;
}

// ---- log-likelihood of a parameter set: logl_W (EM_W_multi.R:296-326) -> loglC_fast (loglC.cpp:318-338) ------
// One streaming pass (the same kernels as an EM iteration; the moments it needs are the Gram of (XW, YC)).
// orth(G, type): OmicsPLS::orth as PPLS_simult calls it (EM_W_multi.R:732-733).  Host in, host out.
extern "C" int ppls_orth(ppls_ctx* ctx, const double* G, int64_t ldG, int64_t n, int32_t r, int32_t orth_type,
                         double* out, int64_t ldOut) {
  if (ctx && !ctx->subs.empty()) ctx = ctx->subs[0];          // context-only work: the first device does it
  int rc = ctx_activate(ctx);
  if (rc) return rc;
  if (!G || !out) return fail(PPLS_E_INVALID, "null G or out");
  if (r < 1 || r > MAXR) return fail(PPLS_E_INVALID, "nr_comp must be in 1..8");
  if (n < r || n > 0x7fffffff / MAXR) return fail(PPLS_E_INVALID, "orth: need r <= nrow(G) < 2^31 / 8");
  if (ldG < n || ldOut < n) return fail(PPLS_E_INVALID, "leading dimension smaller than the row count");
  if (orth_type != PPLS_ORTH_SVD && orth_type != PPLS_ORTH_QR) return fail(PPLS_E_INVALID, "'arg' should be one of 'SVD', 'QR'");
  std::vector<double> h((size_t)n * r);
  for (int k = 0; k < r; ++k)
    for (int64_t j = 0; j < n; ++j) h[(size_t)j * r + k] = G[j + (size_t)k * ldG];
  struct Bufs { double *g = nullptr, *o = nullptr, *w = nullptr; ~Bufs() { cudaFree(g); cudaFree(o); cudaFree(w); } } B;
  const size_t bytes = h.size() * sizeof(double);
  CU(cudaMalloc(&B.g, bytes)); CU(cudaMalloc(&B.o, bytes)); CU(cudaMalloc(&B.w, bytes));
  CU(cudaMemcpyAsync(B.g, h.data(), bytes, cudaMemcpyHostToDevice, ctx->stream));
  CU(launch_orth(r, B.g, (int)n, B.o, orth_type, B.w, ctx->stream));
  CU(cudaMemcpyAsync(h.data(), B.o, bytes, cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  for (int k = 0; k < r; ++k)
    for (int64_t j = 0; j < n; ++j) out[j + (size_t)k * ldOut] = h[(size_t)j * r + k];
  return PPLS_OK;
}

// orth(G, type) on device pointers with the kernel's scratch in a static device array: allocates nothing (em_update.cu).
extern "C" int ppls_orth_noalloc(ppls_ctx* ctx, const double* G_dev, int64_t n, int32_t r, int32_t orth_type, double* out_dev) {
  if (ctx && !ctx->subs.empty()) ctx = ctx->subs[0];
  int rc = ctx_activate(ctx);
  if (rc) return rc;
  if (!G_dev || !out_dev) return fail(PPLS_E_INVALID, "null G or out");
  if (r < 1 || r > MAXR) return fail(PPLS_E_INVALID, "nr_comp must be in 1..8");
  if (n < r) return fail(PPLS_E_INVALID, "orth: need r <= nrow(G)");
  if (orth_type != PPLS_ORTH_SVD && orth_type != PPLS_ORTH_QR) return fail(PPLS_E_INVALID, "'arg' should be one of 'SVD', 'QR'");
  if (n * (int64_t)r > orth_static_max_elems())
    return fail(PPLS_E_UNSUPPORTED, "ppls_orth_noalloc: n * r = " + std::to_string(n * (int64_t)r) + " exceeds the static scratch of " +
                                    std::to_string(orth_static_max_elems()) + " doubles (PPLS_ORTH_NOALLOC_MAX_ELEMS); use ppls_orth");
  CU(launch_orth_static(r, G_dev, (int)n, out_dev, orth_type, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  return PPLS_OK;
}

extern "C" int ppls_loglik(ppls_ctx* ctx, const ppls_data* d, int32_t r, const ppls_params* theta, double* out) {
  if (!out) return fail(PPLS_E_INVALID, "null output pointer");
  ppls_opts opts{};
  opts.r = r; opts.max_steps = 1; opts.atol = 0.0; opts.orth_type = PPLS_ORTH_SVD; opts.check_every = 1;
  ppls_fit* f = nullptr;
  int rc = ppls_fit_create(ctx, d, &opts, theta, &f);
  if (rc) return rc;
  int32_t done = 0, fin = 0;
  rc = ppls_fit_iterate(f, 1, &done, &fin);
  if (rc == PPLS_OK) *out = (f->parts.empty() ? f : f->parts[0])->st_host->loglik_last;
  fit_free(f);
  return rc;
}

// ---- standard errors of the loadings: variances.PPLS_simult (EM_W_multi.R:830-860) -----------------------
// which = 0: data = X, mu = fit$Expectations$mu_T, Cdiag = diag(fit$Expectations$Ctt);
// which = 1: data = Y, mu = mu_U, Cdiag = diag(Cuu).  sigE = fit$estimates$sigE in BOTH cases (the reference
// divides by sigE also for Y, :844,:848,:851).  Kernels: csrc/em_var.cu.
extern "C" int ppls_variances(ppls_ctx* ctx, const ppls_data* d, int32_t which, int32_t r, const double* mu, int64_t ldmu,
                              const double* Cdiag, double sigE, double* seLoad, double* varMatrix, double* SSt_exp,
                              double* SSt_star, double* B_exp, double* Wout) {
  int rc = ctx_activate(ctx);
  if (rc) return rc;
  if (!d || !mu || !Cdiag || !seLoad) return fail(PPLS_E_INVALID, "null argument");
  if (d->ctx != ctx) return fail(PPLS_E_INVALID, "data belongs to another context");
  if (d->colshard) return fail(PPLS_E_UNSUPPORTED, "variances: row-sharded data only");
  if (r < 1 || r > MAXR) return fail(PPLS_E_INVALID, "nr_comp outside 1..8");
  if (which != 0 && which != 1) return fail(PPLS_E_INVALID, "which must be 0 (X) or 1 (Y)");
  const DataShape& sh = d->s;
  if (ldmu < sh.n) return fail(PPLS_E_INVALID, "leading dimension of mu smaller than the row count");
  if (!(sigE > 0.0)) return fail(PPLS_E_INVALID, "sigE must be positive");
  VarPlan vp;
  plan_var(sh, which, ctx->sm_count, &vp);
  const int w = vp.w, R = r;
  const size_t ww = (size_t)w * w;
  struct Bufs {
    double *ws = nullptr, *S = nullptr, *G = nullptr, *cp = nullptr, *mu = nullptr, *Wp = nullptr, *aug = nullptr,
           *rowbuf = nullptr, *se = nullptr, *var = nullptr, *e1 = nullptr, *e2 = nullptr;
    int* info = nullptr;
    ~Bufs() {
      cudaFree(ws); cudaFree(S); cudaFree(G); cudaFree(cp); cudaFree(mu); cudaFree(Wp); cudaFree(aug); cudaFree(rowbuf);
      cudaFree(se); cudaFree(var); cudaFree(e1); cudaFree(e2); cudaFree(info);
    }
  } B;
  CU(cudaMalloc(&B.ws, var_workspace_doubles(vp, R) * sizeof(double)));
  CU(cudaMalloc(&B.S, ww * sizeof(double)));
  CU(cudaMalloc(&B.G, ((size_t)w * R + R) * sizeof(double)));
  B.cp = nullptr;
  CU(cudaMalloc(&B.mu, (size_t)sh.n * R * sizeof(double)));
  CU(cudaMalloc(&B.Wp, (size_t)w * R * sizeof(double)));
  CU(cudaMalloc(&B.aug, 2 * ww * sizeof(double)));
  CU(cudaMalloc(&B.rowbuf, (size_t)5 * w * sizeof(double)));
  CU(cudaMalloc(&B.se, (size_t)w * sizeof(double)));
  CU(cudaMalloc(&B.info, 2 * sizeof(int)));
  if (varMatrix) CU(cudaMalloc(&B.var, ww * sizeof(double)));
  if (SSt_exp) CU(cudaMalloc(&B.e1, ww * sizeof(double)));
  if (SSt_star) CU(cudaMalloc(&B.e2, ww * sizeof(double)));
  double* cpd = B.G + (size_t)w * R;               // cp rides behind G: one allreduce for both
  {  // mu: column-major n x r on the host -> [n][R] row-major on the device
    std::vector<double> t((size_t)sh.n * R);
    for (int k = 0; k < R; ++k)
      for (int64_t i = 0; i < sh.n; ++i) t[(size_t)i * R + k] = mu[i + (size_t)k * ldmu];
    CU(cudaMemcpyAsync(B.mu, t.data(), t.size() * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
  }
  const char* vb = getenv("PPLS_VAR_VERBOSE");
  const bool verbose = vb && *vb && atoi(vb) != 0;
  cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
  if (verbose) { for (auto& x : ev) cudaEventCreate(&x); cudaEventRecord(ev[0], ctx->stream); }
  CU(launch_var_gram(vp, sh, d->D, B.ws, B.S, ctx->stream));
  if (verbose) cudaEventRecord(ev[1], ctx->stream);
  CU(launch_var_xtmu(vp, sh, d->D, B.mu, R, B.ws, B.G, cpd, ctx->stream));
  if (verbose) cudaEventRecord(ev[2], ctx->stream);
  rc = allreduce_sum(ctx, B.S, ww);
  if (rc) return rc;
  rc = allreduce_sum(ctx, B.G, (size_t)w * R + R);
  if (rc) return rc;
  CU(launch_orth_polar(R, B.G, w, B.Wp, ctx->stream));                                    // :831-832
  double cph[MAXR];
  CU(cudaMemcpyAsync(cph, cpd, R * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  const double N = (double)d->n_total;
  for (int i = 0; i < R; ++i) {
    const double CttN = N * Cdiag[i];                                                     // :840
    CU(launch_var_assemble(w, R, i, B.S, B.G, B.Wp, CttN, cph[i], sigE, N, B.aug, B.e1, B.e2, ctx->stream));
    CU(launch_var_invert(w, B.aug, B.rowbuf, B.info, B.var, B.se, ctx->stream));
    int info[2] = {0, 0};
    CU(cudaMemcpyAsync(info, B.info, sizeof(info), cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaMemcpyAsync(seLoad + (size_t)i * w, B.se, (size_t)w * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    if (varMatrix) CU(cudaMemcpyAsync(varMatrix + (size_t)i * ww, B.var, ww * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    if (SSt_exp) CU(cudaMemcpyAsync(SSt_exp + (size_t)i * ww, B.e1, ww * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    if (SSt_star) CU(cudaMemcpyAsync(SSt_star + (size_t)i * ww, B.e2, ww * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    if (info[0]) return fail(PPLS_E_NONFINITE, "variances: B_exp - SSt_exp is singular (zero pivot)");
    if (B_exp) B_exp[i] = CttN / (sigE * sigE) * 1.0 / N;                                   // :844 (times I_p)
  }
  if (verbose) {
    cudaEventRecord(ev[3], ctx->stream);
    cudaEventSynchronize(ev[3]);
    float g = 0, x = 0, rest = 0;
    cudaEventElapsedTime(&g, ev[0], ev[1]); cudaEventElapsedTime(&x, ev[1], ev[2]); cudaEventElapsedTime(&rest, ev[2], ev[3]);
    const double flop = 2.0 * vp.ntiles * 128.0 * 128.0 * (double)sh.n, bytes = 8.0 * (double)sh.n * (which ? sh.q2 : sh.p2);
    fprintf(stderr, "[ppls] variances: n=%lld w=%d r=%d | gram %.3f ms (%d tiles x %d splits, %.2f TFLOP/s FP64) | "
                    "X'mu %.3f ms (%.0f GB/s) | %d x (assemble + invert) %.3f ms\n",
            (long long)sh.n, w, R, g, vp.ntiles, vp.nsplit, flop / (g * 1e-3) / 1e12, x, bytes / (x * 1e-3) / 1e9, R, rest);
    for (auto& e : ev) cudaEventDestroy(e);
  }
  if (Wout) {
    std::vector<double> t((size_t)w * R);
    CU(cudaMemcpyAsync(t.data(), B.Wp, t.size() * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    for (int k = 0; k < R; ++k)
      for (int j = 0; j < w; ++j) Wout[j + (size_t)k * w] = t[(size_t)j * R + k];
  }
  return PPLS_OK;
}

// ---- multi-population ("meta") fit: meta_PPLSi (EM_W_multi.R:509-589) --------------------------------------------
// Populations are contiguous row blocks (pop_rows[j] rows each, in row order -- the reference slices
// (1+Ni[i]):Ni[i+1], :455-456).  One component; critfunc = identity, no constraints.  Per pass the streaming
// kernel K1 (R = 1) runs once per population on its row block with that population's coefficients and the
// shared loadings; em_seq.cu::meta_update_kernel does the per-population scalars and pools the loadings.
extern "C" int ppls_meta_run(ppls_ctx* ctx, const ppls_data* d, int32_t n_pop, const int64_t* pop_rows, int32_t EMsteps,
                             double atol, const ppls_params* guess, double* W, double* Cout, double* params,
                             double* logvalues, int32_t* steps) {
  int rc = ctx_activate(ctx);
  if (rc) return rc;
  if (!d || !pop_rows || !guess || !W || !Cout || !params || !steps) return fail(PPLS_E_INVALID, "null argument");
  if (d->ctx != ctx) return fail(PPLS_E_INVALID, "data belongs to another context");
  if (d->colshard) return fail(PPLS_E_UNSUPPORTED, "meta fit: row-sharded data only");
  const bool multi = ctx->world > 1;          // row-sharded: pop_rows are this rank's LOCAL rows of each population (0 allowed)
  if (d->s.panel) return fail(PPLS_E_UNSUPPORTED, "meta fit: wide-row (panel-major) data is not supported");
  if (n_pop < 1 || n_pop > 64) return fail(PPLS_E_INVALID, "number of populations outside 1..64");
  if (EMsteps < 1) return fail(PPLS_E_INVALID, "EMsteps must be a strictly positive integer");
  rc = check_params(guess, "guess");
  if (rc) return rc;
  const DataShape& sh = d->s;
  const int J = n_pop;
  int64_t tot = 0;
  for (int j = 0; j < J; ++j) {
    if (pop_rows[j] < (multi ? 0 : 1)) return fail(PPLS_E_INVALID, "empty population");
    tot += pop_rows[j];
  }
  if (tot != sh.n) return fail(PPLS_E_INVALID, "nrow(X) == length(Ipopu) is required");
  const int64_t veclen = momvec_len(sh.ld, 1);
  struct PopPlan { DataShape s; StreamPlan plan; const double* D; };
  std::vector<PopPlan> pops(J);
  int maxgrid = 1;
  int64_t row = 0;
  for (int j = 0; j < J; ++j) {
    pops[j].s = sh; pops[j].s.n = pop_rows[j];
    pops[j].D = d->D + (size_t)row * sh.ld;
    if (pop_rows[j] == 0) continue;                       // (multi-GPU) none of this population's rows live here
    const char* why = "";
    if (plan_stream(pops[j].s, 1, ctx->sm_count, &pops[j].plan, &why) != cudaSuccess)
      return fail(PPLS_E_UNSUPPORTED, std::string("meta fit: no streaming plan for this shape: ") + why);
    if (pops[j].plan.grid > maxgrid) maxgrid = pops[j].plan.grid;
    row += pop_rows[j];
  }
  struct Bufs {
    double *partials = nullptr, *momvecs = nullptr, *L = nullptr, *trace = nullptr, *ssq = nullptr, *scr = nullptr;
    FitState* sts = nullptr;
    ~Bufs() { cudaFree(partials); cudaFree(momvecs); cudaFree(L); cudaFree(trace); cudaFree(ssq); cudaFree(scr); cudaFree(sts); }
  } B;
  const int nblocks = ctx->sm_count * 4;
  CU(cudaMalloc(&B.partials, (size_t)maxgrid * veclen * sizeof(double)));
  CU(cudaMalloc(&B.momvecs, (size_t)J * veclen * sizeof(double)));
  CU(cudaMalloc(&B.L, (size_t)sh.ld * sizeof(double)));
  CU(cudaMalloc(&B.trace, (size_t)(EMsteps + 1) * J * sizeof(double)));
  CU(cudaMalloc(&B.ssq, (size_t)3 * J * sizeof(double)));           // per population: ||X||^2, ||Y||^2 | row counts
  CU(cudaMalloc(&B.scr, (size_t)2 * nblocks * sizeof(double)));
  CU(cudaMalloc(&B.sts, (size_t)J * sizeof(FitState)));
  CU(cudaMemsetAsync(B.trace, 0, (size_t)(EMsteps + 1) * J * sizeof(double), ctx->stream));
  // per-population sums of squares (loglC.cpp:421,425: X.array().square().sum() of the population's rows)
  CU(cudaMemsetAsync(B.ssq, 0, (size_t)3 * J * sizeof(double), ctx->stream));
  for (int j = 0; j < J; ++j)
    if (pop_rows[j] > 0) CU(launch_ssq(pops[j].D, pops[j].s, B.ssq + 2 * j, B.scr, nblocks, ctx->stream));
  std::vector<double> ssq((size_t)3 * J);
  for (int j = 0; j < J; ++j) ssq[(size_t)2 * J + j] = (double)pop_rows[j];
  CU(cudaMemcpyAsync(B.ssq + 2 * J, ssq.data() + 2 * J, (size_t)J * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
  if (multi) { rc = allreduce_sum(ctx, B.ssq, (size_t)3 * J); if (rc) return rc; }     // population totals over the ranks
  CU(cudaMemcpyAsync(ssq.data(), B.ssq, ssq.size() * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  for (int j = 0; j < J; ++j)
    if (ssq[(size_t)2 * J + j] < 0.5) return fail(PPLS_E_INVALID, "empty population");
  std::vector<FitState> hst(J);
  std::vector<double> L0;
  pack_loadings(sh, 1, guess->W, guess->C, L0);
  for (int j = 0; j < J; ++j) {
    FitState& h = hst[j];
    memset(&h, 0, sizeof(h));
    h.b[0] = guess->B[0]; h.t[0] = guess->sigT[0]; h.sigE = guess->sigE; h.sigF = guess->sigF; h.sigH = guess->sigH;
    h.ssqX = ssq[2 * j]; h.ssqY = ssq[2 * j + 1]; h.Ntot = ssq[(size_t)2 * J + j];
    h.max_steps = EMsteps; h.atol = atol; h.R = 1; h.orth_type = PPLS_ORTH_SVD;
  }
  CU(cudaMemcpyAsync(B.sts, hst.data(), (size_t)J * sizeof(FitState), cudaMemcpyHostToDevice, ctx->stream));
  CU(cudaMemcpyAsync(B.L, L0.data(), (size_t)sh.ld * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
  CU(launch_meta_update(sh, J, B.momvecs, B.L, B.sts, B.trace, 1, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  int passes = 0;
  bool finished = false;
  while (!finished) {
    for (int rep = 0; rep < 8 && passes <= EMsteps; ++rep, ++passes) {
      for (int j = 0; j < J; ++j) {
        if (pop_rows[j] == 0) { CU(cudaMemsetAsync(B.momvecs + (size_t)j * veclen, 0, (size_t)veclen * sizeof(double), ctx->stream)); continue; }
        CU(launch_em_stream(pops[j].plan, pops[j].s, pops[j].D, B.L, B.sts + j, B.partials, nullptr, nullptr, 0, ctx->stream));
        CU(launch_reduce_partials(B.partials, pops[j].plan.grid, veclen, B.momvecs + (size_t)j * veclen, B.sts + j, ctx->stream));
      }
      // row-sharded: the pooled sums of EM_W_multi.R:481-482 run over every rank's rows (after the fit has ended the
      // kernels above no longer rewrite the vectors and this sum scales stale values that nothing reads)
      if (multi) { rc = allreduce_sum(ctx, B.momvecs, (size_t)J * veclen); if (rc) return rc; }
      CU(launch_meta_update(sh, J, B.momvecs, B.L, B.sts, B.trace, 0, ctx->stream));
    }
    CU(cudaMemcpyAsync(hst.data(), B.sts, (size_t)J * sizeof(FitState), cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    finished = hst[0].done != 0;
    if (!finished && passes > EMsteps) return fail(PPLS_E_CUDA, "meta EM loop did not terminate");
  }
  if (hst[0].status != PPLS_OK) return fail(hst[0].status, "meta fit: a parameter became non-finite");
  const int k = hst[0].steps_done;
  std::vector<double> L(sh.ld), tr((size_t)(k + 1) * J);
  CU(cudaMemcpyAsync(L.data(), B.L, (size_t)sh.ld * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaMemcpyAsync(tr.data(), B.trace, tr.size() * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
  CU(cudaStreamSynchronize(ctx->stream));
  for (int j = 0; j < sh.p; ++j) W[j] = L[j];
  for (int j = 0; j < sh.q; ++j) Cout[j] = L[sh.p2 + j];
  for (int j = 0; j < J; ++j) {       // column-major J x 5: B_T, sigX, sigY, sigH, sigT
    params[j + 0 * J] = hst[j].b[0]; params[j + 1 * J] = hst[j].sigE; params[j + 2 * J] = hst[j].sigF;
    params[j + 3 * J] = hst[j].sigH; params[j + 4 * J] = hst[j].t[0];
  }
  *steps = k;
  if (logvalues)                        // rows 1..k of the trace = logvalue[2 .. k+1, ] (:586), row-major k x J
    for (int i = 1; i <= k; ++i)
      for (int j = 0; j < J; ++j) logvalues[(size_t)(i - 1) * J + j] = tr[(size_t)i * J + j];
  return PPLS_OK;
}
