// Context: device selection, streams, caching pool, host mailbox, workspace, NCCL (dlopen),
// per-kernel-class event timing.  Replaces Epetra_Comm / Epetra_MpiComm for this path
// (epetra/src/Epetra_MpiComm.cpp:62,126,154; Epetra_SerialComm.cpp:117-152).
#include "runtime.cuh"

#include <dlfcn.h>
#include <immintrin.h>

#include <atomic>
#include <chrono>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace b200 {

static thread_local char g_err[1024] = "";
static thread_local int g_status = 0;

int fail(int code, const char* msg) {
  if (msg != g_err) std::snprintf(g_err, sizeof(g_err), "%s", msg);
  g_status = code;
  return code;
}
int failf(int code, const char* fmt, ...) {
  g_status = code;
  va_list ap;
  va_start(ap, fmt);
  std::vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

// ------------------------------------------------------------------ NCCL via dlopen
// The single-GPU path must not depend on libnccl being loadable; the multi-rank path resolves
// the handful of entry points it needs at b200_ctx_comm_init time.
struct Id128 { char b[128]; };
struct NcclApi {
  void* h = nullptr;
  int (*GetUniqueId)(void*) = nullptr;
  int (*CommInitRank)(void**, int, /*ncclUniqueId by value*/ Id128, int) = nullptr;
  int (*CommDestroy)(void*) = nullptr;
  int (*AllReduce)(const void*, void*, size_t, int, int, void*, cudaStream_t) = nullptr;
  int (*Send)(const void*, size_t, int, int, void*, cudaStream_t) = nullptr;
  int (*Recv)(void*, size_t, int, int, void*, cudaStream_t) = nullptr;
  int (*GroupStart)() = nullptr;
  int (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
};
constexpr int kNcclFloat64 = 8, kNcclInt64 = 4, kNcclSum = 0, kNcclMax = 2;

static NcclApi* load_nccl() {
  static NcclApi api;
  if (api.h) return &api;
  const char* names[] = {"libnccl.so.2", "libnccl.so"};
  for (const char* nm : names) {
    api.h = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
    if (api.h) break;
  }
  if (!api.h) return nullptr;
#define B200_SYM(field, name)                                                 \
  *(void**)(&api.field) = dlsym(api.h, name);                                 \
  if (!api.field) {                                                           \
    api.h = nullptr;                                                          \
    return nullptr;                                                           \
  }
  B200_SYM(GetUniqueId, "ncclGetUniqueId");
  B200_SYM(CommInitRank, "ncclCommInitRank");
  B200_SYM(CommDestroy, "ncclCommDestroy");
  B200_SYM(AllReduce, "ncclAllReduce");
  B200_SYM(Send, "ncclSend");
  B200_SYM(Recv, "ncclRecv");
  B200_SYM(GroupStart, "ncclGroupStart");
  B200_SYM(GroupEnd, "ncclGroupEnd");
  B200_SYM(GetErrorString, "ncclGetErrorString");
#undef B200_SYM
  return &api;
}

#define B200_NCCL(api, call)                                                              \
  do {                                                                                    \
    int r_ = (call);                                                                      \
    if (r_ != 0)                                                                          \
      return ::b200::failf(B200_ERR_NCCL, "%s:%d NCCL: %s", __FILE__, __LINE__,           \
                           (api)->GetErrorString(r_));                                    \
  } while (0)

// helpers used by capi.cu for the halo exchange / setup exchange
int nccl_allreduce_f64(Ctx* ctx, double* buf, size_t count, bool is_max, cudaStream_t s) {
  B200_NCCL(ctx->nccl, ctx->nccl->AllReduce(buf, buf, count, kNcclFloat64,
                                            is_max ? kNcclMax : kNcclSum, ctx->comm, s));
  return 0;
}
int nccl_allreduce_i64(Ctx* ctx, int64_t* buf, size_t count, cudaStream_t s) {
  B200_NCCL(ctx->nccl,
            ctx->nccl->AllReduce(buf, buf, count, kNcclInt64, kNcclSum, ctx->comm, s));
  return 0;
}
int nccl_group_start(Ctx* ctx) {
  B200_NCCL(ctx->nccl, ctx->nccl->GroupStart());
  return 0;
}
int nccl_group_end(Ctx* ctx) {
  B200_NCCL(ctx->nccl, ctx->nccl->GroupEnd());
  return 0;
}
int nccl_send(Ctx* ctx, const void* buf, size_t count, bool f64, int peer, cudaStream_t s) {
  B200_NCCL(ctx->nccl,
            ctx->nccl->Send(buf, count, f64 ? kNcclFloat64 : kNcclInt64, peer, ctx->comm, s));
  return 0;
}
int nccl_recv(Ctx* ctx, void* buf, size_t count, bool f64, int peer, cudaStream_t s) {
  B200_NCCL(ctx->nccl,
            ctx->nccl->Recv(buf, count, f64 ? kNcclFloat64 : kNcclInt64, peer, ctx->comm, s));
  return 0;
}

// ------------------------------------------------------------------ storage
Storage::~Storage() {
  // copies that were never needed: forget them at their sources (no launch ever happened)
  for (size_t c = 0; c < lazy.size(); ++c)
    if (lazy[c].src) {
      auto& d = lazy[c].src->deps[(size_t)lazy[c].col];
      for (size_t i = 0; i < d.size(); ++i)
        if (d[i].first == this && d[i].second == (int)c) {
          d.erase(d.begin() + (long)i);
          break;
        }
      --ctx->pending;
      ++ctx->lazy_saved;
    }
  for (double f : pscale)
    if (f != 1.0) --ctx->pending;
  ctx->release(ptr, bytes);
}

// ------------------------------------------------------------------ pool
int Ctx::alloc(size_t bytes, void** out) {
  bytes = (bytes + 511) & ~(size_t)511;
  if (bytes == 0) bytes = 512;
  auto it = pool.find(bytes);
  if (it != pool.end()) {
    *out = it->second;
    pool.erase(it);
    pool_bytes -= bytes;
    return 0;
  }
  cudaError_t e = cudaMalloc(out, bytes);
  if (e != cudaSuccess) {
    // drop the cache and retry once
    for (auto& kv : pool) cudaFree(kv.second);
    pool.clear();
    pool_bytes = 0;
    cudaGetLastError();
    e = cudaMalloc(out, bytes);
  }
  if (e != cudaSuccess)
    return failf(B200_ERR_CUDA, "cudaMalloc(%zu bytes): %s", bytes, cudaGetErrorString(e));
  return 0;
}

void Ctx::release(void* p, size_t bytes) {
  if (!p) return;
  bytes = (bytes + 511) & ~(size_t)511;
  if (bytes == 0) bytes = 512;
  // Stream-ordered reuse: every kernel of this context runs on `stream`, so a block handed
  // out again is only touched after all previously enqueued work on it.
  pool.emplace(bytes, p);
  pool_bytes += bytes;
}

// ------------------------------------------------------------------ timing
void Ctx::t_begin(int cls, double bytes) {
  if (!timing) return;
  Stamp s;
  auto get = [&]() {
    cudaEvent_t e;
    if (!ev_free.empty()) {
      e = ev_free.back();
      ev_free.pop_back();
    } else {
      cudaEventCreate(&e);
    }
    return e;
  };
  s.a = get();
  s.b = get();
  s.cls = cls;
  cudaEventRecord(s.a, stream);
  stamps.push_back(s);
  t_calls[cls] += 1;
  t_bytes[cls] += bytes;
}
void Ctx::t_end() {
  if (!timing || stamps.empty()) return;
  cudaEventRecord(stamps.back().b, stream);
  if (stamps.size() >= 4096) t_drain();
}
int Ctx::t_drain() {
  if (stamps.empty()) return 0;
  B200_CUDA(cudaStreamSynchronize(stream));
  for (auto& s : stamps) {
    float ms = 0.f;
    cudaEventElapsedTime(&ms, s.a, s.b);
    t_ms[s.cls] += ms;
    ev_free.push_back(s.a);
    ev_free.push_back(s.b);
  }
  stamps.clear();
  return 0;
}

// ------------------------------------------------------------------ mailbox
int Ctx::wait_flag(uint64_t s) {
  ++syncs;
  using clk = std::chrono::steady_clock;
  const auto t0 = clk::now();
  uint64_t spins = 0;
  while (true) {
    if (*flag_h >= s) break;
    _mm_pause();
    if ((++spins & 0xfffff) == 0) {
      if (flag_h[1] != 0)  // a peer-memory wait on the device gave up (spin_acquire_sys)
        return fail(B200_ERR_TIMEOUT, "a peer GPU did not arrive at an exchange within 30 s");
      cudaError_t e = cudaStreamQuery(stream);
      if (e != cudaSuccess && e != cudaErrorNotReady)
        return failf(B200_ERR_CUDA, "device error while waiting for a reduction: %s",
                     cudaGetErrorString(e));
      if (e == cudaSuccess && *flag_h < s) {
        // stream drained but the flag never arrived: should be impossible
        _mm_pause();
        if (*flag_h < s) return fail(B200_ERR_CUDA, "reduction result was never published");
      }
      if (std::chrono::duration<double>(clk::now() - t0).count() > 120.0)
        return fail(B200_ERR_TIMEOUT, "timeout waiting for a reduction result");
    }
  }
  std::atomic_thread_fence(std::memory_order_acquire);
  // a peer-memory wait that gave up earlier in the stream (halo or exchange) leaves the error word set:
  // whatever was published after it was computed from stale data.  Sticky until the context is destroyed.
  if (flag_h[1] != 0)
    return fail(B200_ERR_TIMEOUT, "a peer GPU did not arrive at an exchange within 30 s");
  return 0;
}

// ------------------------------------------------------------------ remembered dot products
void Ctx::remember_dot(uint64_t a, uint64_t b, double val, uint64_t seq_, size_t src, double mul) {
  if (!fuse) return;
  if (a > b) std::swap(a, b);
  for (int i = 0; i < kDotCache; ++i)
    if (dots[i].a == a && dots[i].b == b) {
      dots[i].val = val; dots[i].seq = seq_; dots[i].src = src; dots[i].mul = mul;
      return;
    }
  DotEntry& e = dots[dot_next];
  dot_next = (dot_next + 1) % kDotCache;
  e.a = a; e.b = b; e.val = val; e.seq = seq_; e.src = src; e.mul = mul;
}

bool Ctx::recall_dot(uint64_t a, uint64_t b, double* val, int* rc) {
  if (!fuse) return false;
  if (a > b) std::swap(a, b);
  for (int i = 0; i < kDotCache; ++i) {
    DotEntry& e = dots[i];
    if (e.a != a || e.b != b || a == 0) continue;
    if (e.seq != 0) {  // produced by a fused epilogue that may still be running
      *rc = wait_flag(e.seq);
      if (*rc) return false;
      e.val = e.mul * mbox_h[e.src];
      e.seq = 0;
      e.mul = 1.0;
    }
    *val = e.val;
    return true;
  }
  return false;
}

void Ctx::rescale_dots(uint64_t from, uint64_t to, double factor) {
  if (!fuse || from == 0) return;
  DotEntry moved[16];
  int nm = 0;
  for (int i = 0; i < kDotCache && nm < 16; ++i) {
    const DotEntry& e = dots[i];
    if (e.a != from && e.b != from) continue;
    DotEntry m = e;
    const double f = (e.a == from && e.b == from) ? factor * factor : factor;
    if (m.seq != 0) m.mul *= f; else m.val *= f;
    if (m.a == from) m.a = to;
    if (m.b == from) m.b = to;
    moved[nm++] = m;
  }
  for (int i = 0; i < nm; ++i) remember_dot(moved[i].a, moved[i].b, moved[i].val, moved[i].seq, moved[i].src, moved[i].mul);
}

void Ctx::plan_reduction(int count, bool is_max) {
  static const bool on = [] {
    const char* e = std::getenv("B200_FUSE_XCHG");
    return !(e && e[0] == '0');
  }();
  xplan_count = count;
  xplan_max = is_max;
  xplan_ok = on && nranks > 1 && p2p;
  xplan_done = false;
}

XchgPub Ctx::make_pub(bool last) {
  XchgPub pb;
  pb.seq = seq;
  if (!last) return pb;
  if (nranks <= 1) {
    pb.flag = flag_d;
    return pb;
  }
  if (!xplan_ok) return pb;  // finish_reduction() launches the exchange (or the NCCL allreduce)
  pb.flag = flag_d;
  pb.nranks = nranks;
  pb.rank = rank;
  pb.count = xplan_count;
  pb.is_max = xplan_max ? 1 : 0;
  pb.xseq = xseq + 1;   // one exchange number per chunk of kXchgMax entries
  xseq += (uint64_t)std::max(1, (xplan_count + kXchgMax - 1) / kXchgMax);
  pb.src = red_d;
  pb.dst = mbox_d;
  pb.err = flag_d + 1;
  for (int i = 0; i < nranks; ++i) pb.win[i] = xw.win[i];
  xplan_done = true;
  return pb;
}

int Ctx::finish_reduction(int count, bool is_max) {
  const bool done = xplan_done;
  xplan_ok = xplan_done = false;
  if (nranks > 1 && !done) {
    if (p2p) {
      B200_TRY(launch_xchg_publish(this, red_d, mbox_d, count, is_max, seq));
    } else {
      B200_TRY(nccl_allreduce_f64(this, red_d, (size_t)count, is_max, stream));
      ++launches;
      B200_TRY(launch_publish(this, count));
    }
  }
  return wait_flag(seq);
}

}  // namespace b200

using namespace b200;

extern "C" const char* b200_last_error(void) { return g_err; }
extern "C" int b200_last_status(void) { return g_status; }
extern "C" const char* b200_version(void) { return "trilinos_b200 0.1 (sm_100a)"; }

extern "C" int b200_ctx_create(int device, b200_ctx** out) {
  if (!out) return fail(B200_ERR_ARG, "b200_ctx_create: null out");
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0) {
    cudaGetLastError();
    return fail(B200_ERR_NO_DEVICE,
                "no CUDA device: the B200 path has no CPU fallback (cudaGetDeviceCount failed)");
  }
  if (device < 0 || device >= ndev) return fail(B200_ERR_ARG, "b200_ctx_create: bad device id");
  B200_CUDA(cudaSetDevice(device));
  auto* c = new b200_ctx();
  c->device = device;
  if (const char* nf = std::getenv("B200_NO_FUSE")) c->fuse = !(nf[0] == '1');
  B200_CUDA(cudaDeviceGetAttribute(&c->num_sms, cudaDevAttrMultiProcessorCount, device));
  B200_CUDA(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
  B200_CUDA(cudaStreamCreateWithFlags(&c->comm_stream, cudaStreamNonBlocking));
  B200_CUDA(cudaEventCreateWithFlags(&c->ev_comm, cudaEventDisableTiming));
  B200_CUDA(cudaEventCreateWithFlags(&c->ev_compute, cudaEventDisableTiming));
  void* hp = nullptr;
  B200_CUDA(cudaHostAlloc(&hp, kMailboxDoubles * sizeof(double) + 64, cudaHostAllocMapped));
  std::memset(hp, 0, kMailboxDoubles * sizeof(double) + 64);
  c->flag_h = (volatile uint64_t*)hp;
  c->mbox_h = (double*)((char*)hp + 64);
  void* dp = nullptr;
  B200_CUDA(cudaHostGetDevicePointer(&dp, hp, 0));
  c->flag_d = (uint64_t*)dp;
  c->mbox_d = (double*)((char*)dp + 64);
  B200_CUDA(cudaMalloc((void**)&c->ws, kWorkspaceBytes));
  B200_CUDA(cudaMalloc((void**)&c->ticket, kTicketBytes));
  B200_CUDA(cudaMemset(c->ticket, 0, kTicketBytes));
  B200_CUDA(cudaMalloc((void**)&c->red_d, kMailboxDoubles * sizeof(double)));
  B200_CUDA(cudaMalloc((void**)&c->aux_d, ((size_t)kAuxSlots * kAuxSlotDoubles + (size_t)kSpecSlots * kSpecSlotDoubles) * sizeof(double)));
  B200_CUDA(cudaMalloc((void**)&c->dB, kStageSlotBytes));
  B200_CUDA(cudaMalloc((void**)&c->dBpack, kPackBytes));
  B200_CUDA(cudaHostAlloc((void**)&c->stage_h, kStageSlots * kStageSlotBytes, cudaHostAllocDefault));
  for (int i = 0; i < kStageSlots; ++i)
    B200_CUDA(cudaEventCreateWithFlags(&c->stage_ev[i], cudaEventDisableTiming));
  B200_CUDA(cudaDeviceSynchronize());
  *out = c;
  return 0;
}

extern "C" int b200_ctx_destroy(b200_ctx* c) {
  if (!c) return 0;
  cudaSetDevice(c->device);
  cudaDeviceSynchronize();
  for (int r = 0; r < c->xw.nranks; ++r)
    if (r != c->xw.rank && c->xw.win[r]) cudaIpcCloseMemHandle(c->xw.win[r]);
  if (c->xwin) cudaFree(c->xwin);
  if (c->comm && c->nccl) c->nccl->CommDestroy(c->comm);
  for (auto& kv : c->pool) cudaFree(kv.second);
  c->pool.clear();
  for (auto& s : c->stamps) { cudaEventDestroy(s.a); cudaEventDestroy(s.b); }
  for (auto e : c->ev_free) cudaEventDestroy(e);
  for (int i = 0; i < kStageSlots; ++i) cudaEventDestroy(c->stage_ev[i]);
  cudaFreeHost((void*)c->flag_h);
  cudaFreeHost(c->stage_h);
  cudaFree(c->ws);
  cudaFree(c->ticket);
  cudaFree(c->red_d);
  cudaFree(c->aux_d);
  cudaFree(c->dB);
  cudaFree(c->dBpack);
  if (c->flush_buf) cudaFree(c->flush_buf);
  cudaEventDestroy(c->ev_comm);
  cudaEventDestroy(c->ev_compute);
  cudaStreamDestroy(c->stream);
  cudaStreamDestroy(c->comm_stream);
  delete c;
  return 0;
}

extern "C" int b200_ctx_sync(b200_ctx* c) {
  if (!c) return fail(B200_ERR_ARG, "null context");
  B200_CUDA(cudaStreamSynchronize(c->stream));
  B200_CUDA(cudaStreamSynchronize(c->comm_stream));
  return 0;
}

// Map every rank's exchange window into this process (CUDA IPC; the 64-byte handles travel through
// one int64 allreduce on the fresh communicator).  Any failure on any rank leaves all ranks on
// the NCCL allreduce path (the decision is itself allreduced, so it is unanimous).
static int setup_peer_windows(b200_ctx* c) {
  const int P = c->nranks, me = c->rank;
  const char* off = getenv("B200_NO_P2P");
  int ok = (P <= kMaxPeers && !(off && off[0] == '1')) ? 1 : 0;
  cudaIpcMemHandle_t mine;
  std::memset(&mine, 0, sizeof(mine));
  if (ok && cudaMalloc(&c->xwin, kXchgBytes) != cudaSuccess) ok = 0;
  if (ok && cudaMemset(c->xwin, 0, kXchgBytes) != cudaSuccess) ok = 0;
  if (ok && cudaDeviceSynchronize() != cudaSuccess) ok = 0;  // zeroed before any peer can learn the handle
  if (ok && cudaIpcGetMemHandle(&mine, c->xwin) != cudaSuccess) ok = 0;
  cudaGetLastError();
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  const size_t per = 9;  // 8 words of handle + 1 "ok" word per rank
  std::vector<int64_t> h((size_t)P * per, 0);
  std::memcpy(&h[(size_t)me * per], &mine, 64);
  h[(size_t)me * per + 8] = ok;
  int64_t* d = nullptr;
  B200_CUDA(cudaMalloc((void**)&d, h.size() * sizeof(int64_t)));
  B200_CUDA(cudaMemcpy(d, h.data(), h.size() * sizeof(int64_t), cudaMemcpyHostToDevice));
  B200_TRY(nccl_allreduce_i64(c, d, h.size(), c->stream));
  B200_CUDA(cudaMemcpyAsync(h.data(), d, h.size() * sizeof(int64_t), cudaMemcpyDeviceToHost, c->stream));
  B200_CUDA(cudaStreamSynchronize(c->stream));
  for (int r = 0; r < P; ++r) ok = ok && h[(size_t)r * per + 8] == 1;
  c->xw.nranks = P;
  c->xw.rank = me;
  for (int r = 0; r < P && ok; ++r) {
    if (r == me) {
      c->xw.win[r] = c->xwin;
      continue;
    }
    cudaIpcMemHandle_t hr;
    std::memcpy(&hr, &h[(size_t)r * per], 64);
    if (cudaIpcOpenMemHandle(&c->xw.win[r], hr, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
      cudaGetLastError();
      ok = 0;
    }
  }
  // unanimous decision (an open can fail on one rank only)
  int64_t v = ok;
  B200_CUDA(cudaMemcpy(d, &v, sizeof(v), cudaMemcpyHostToDevice));
  B200_TRY(nccl_allreduce_i64(c, d, 1, c->stream));
  B200_CUDA(cudaMemcpyAsync(&v, d, sizeof(v), cudaMemcpyDeviceToHost, c->stream));
  B200_CUDA(cudaStreamSynchronize(c->stream));
  cudaFree(d);
  c->p2p = (v == P);
  if (getenv("B200_VERBOSE"))
    fprintf(stderr, "[b200 rank %d/%d] peer windows: %s\n", me, P, c->p2p ? "on" : "off (NCCL allreduce)");
  return 0;
}

extern "C" int b200_comm_unique_id(void* id128) {
  NcclApi* api = load_nccl();
  if (!api) return fail(B200_ERR_NCCL, "libnccl.so.2 could not be loaded");
  B200_NCCL(api, api->GetUniqueId(id128));
  return 0;
}

extern "C" int b200_ctx_comm_init(b200_ctx* c, const void* id128, int nranks, int rank) {
  if (!c || !id128 || nranks < 1 || rank < 0 || rank >= nranks)
    return fail(B200_ERR_ARG, "b200_ctx_comm_init: bad arguments");
  if (nranks == 1) {
    c->rank = 0;
    c->nranks = 1;
    return 0;
  }
  NcclApi* api = load_nccl();
  if (!api) return fail(B200_ERR_NCCL, "libnccl.so.2 could not be loaded");
  B200_CUDA(cudaSetDevice(c->device));
  Id128 id;
  std::memcpy(id.b, id128, 128);
  void* comm = nullptr;
  B200_NCCL(api, api->CommInitRank(&comm, nranks, id, rank));
  c->nccl = api;
  c->comm = comm;
  c->rank = rank;
  c->nranks = nranks;
  return setup_peer_windows(c);
}

extern "C" int b200_ctx_rank(const b200_ctx* c, int* rank, int* nranks) {
  if (!c) return fail(B200_ERR_ARG, "null context");
  if (rank) *rank = c->rank;
  if (nranks) *nranks = c->nranks;
  return 0;
}

extern "C" int b200_ctx_set_seed(b200_ctx* c, unsigned int seed) {
  if (!c) return fail(B200_ERR_ARG, "null context");
  seed %= 2147483647u;
  if (seed == 0) seed = 1;
  c->seed = seed;
  return 0;
}

extern "C" int b200_ctx_counters(const b200_ctx* c, int64_t* launches, int64_t* syncs) {
  if (!c) return fail(B200_ERR_ARG, "null context");
  if (launches) *launches = c->launches;
  if (syncs) *syncs = c->syncs;
  return 0;
}

extern "C" int b200_ctx_set_fusion(b200_ctx* c, int on) {
  if (!c) return fail(B200_ERR_ARG, "null context");
  c->fuse = on != 0;
  for (auto& e : c->dots) e = Ctx::DotEntry();
  c->gmemo.valid = false;
  c->spec = Ctx::SpecState();
  c->last_block = Ctx::LastBlock();
  return 0;
}

extern "C" int b200_ctx_fusion_stats(const b200_ctx* c, int64_t* dot_hits) {
  if (!c) return fail(B200_ERR_ARG, "null context");
  if (dot_hits) *dot_hits = c->dot_hits;
  return 0;
}

extern "C" int b200_ctx_spec_gram_stats(const b200_ctx* c, int64_t* launches, int64_t* hits) {
  if (!c) return fail(B200_ERR_ARG, "null context");
  if (launches) *launches = c->gram_spec;
  if (hits) *hits = c->gram_hits;
  return 0;
}

extern "C" int b200_ctx_answered_bytes(b200_ctx* c, double* bytes, int reset) {
  if (!c) return fail(B200_ERR_ARG, "null context");
  if (bytes) *bytes = c->answered_bytes;
  if (reset) c->answered_bytes = 0.0;
  return 0;
}

extern "C" int b200_ctx_findbasis_stats(const b200_ctx* c, int64_t* gram_calls_answered) {
  if (!c) return fail(B200_ERR_ARG, "null context");
  if (gram_calls_answered) *gram_calls_answered = c->skinny_hits;
  return 0;
}

extern "C" int b200_ctx_deferral_stats(const b200_ctx* c, int64_t* lazy_copies_saved, int64_t* scales_folded) {
  if (!c) return fail(B200_ERR_ARG, "null context");
  if (lazy_copies_saved) *lazy_copies_saved = c->lazy_saved;
  if (scales_folded) *scales_folded = c->scale_folded;
  return 0;
}

extern "C" int b200_ctx_timing_enable(b200_ctx* c, int on) {
  if (!c) return fail(B200_ERR_ARG, "null context");
  if (!on) B200_TRY(c->t_drain());
  c->timing = on != 0;
  return 0;
}

extern "C" int b200_ctx_timing_read(b200_ctx* c, double* ms, int64_t* calls, double* bytes,
                                    int reset) {
  if (!c) return fail(B200_ERR_ARG, "null context");
  B200_TRY(c->t_drain());
  for (int i = 0; i < B200_NUM_KCLASS; ++i) {
    if (ms) ms[i] = c->t_ms[i];
    if (calls) calls[i] = c->t_calls[i];
    if (bytes) bytes[i] = c->t_bytes[i];
    if (reset) {
      c->t_ms[i] = 0;
      c->t_calls[i] = 0;
      c->t_bytes[i] = 0;
    }
  }
  return 0;
}

extern "C" void* b200_ctx_stream(b200_ctx* c) { return c ? (void*)c->stream : nullptr; }

extern "C" int b200_ctx_flush_l2(b200_ctx* c) {
  if (!c) return fail(B200_ERR_ARG, "null context");
  const size_t bytes = 256u << 20;  // 2 x the 126 MB L2
  if (!c->flush_buf) {
    B200_CUDA(cudaMalloc(&c->flush_buf, bytes));
    B200_CUDA(cudaMemsetAsync(c->flush_buf, 0, bytes, c->stream));
  }
  B200_TRY(launch_l2_sweep(c, (const double*)c->flush_buf, bytes, c->ws));
  return 0;
}
