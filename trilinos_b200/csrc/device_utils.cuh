// Small device helpers shared by the kernels (sm_100a).
#ifndef B200_DEVICE_UTILS_CUH
#define B200_DEVICE_UTILS_CUH

#include <cuda_runtime.h>

#include <cstdint>

namespace b200 {

// Streaming 256-bit load (sm_100 LDG.256): read-only path, no L1 allocation, evict-first in
// L2 — each Krylov-basis element is touched exactly once per pass, so it must not push the
// small re-used operand (the n x b block, which fits the 126 MB L2) out of cache.
struct double4v { double x, y, z, w; };
__device__ __forceinline__ double4v ld_stream4(const double* p) {
  double4v v;
  asm volatile("ld.global.nc.L1::no_allocate.L2::evict_first.v4.f64 {%0,%1,%2,%3}, [%4];"
               : "=d"(v.x), "=d"(v.y), "=d"(v.z), "=d"(v.w)
               : "l"(p));
  return v;
}
__device__ __forceinline__ double ld_stream1(const double* p) {
  double v;
  asm volatile("ld.global.nc.L1::no_allocate.f64 %0, [%1];" : "=d"(v) : "l"(p));
  return v;
}
// re-used operand: 256-bit load that asks L2 to keep the line
__device__ __forceinline__ double4v ld_keep4(const double* p) {
  double4v v;
  asm volatile("ld.global.nc.L2::evict_last.v4.f64 {%0,%1,%2,%3}, [%4];"
               : "=d"(v.x), "=d"(v.y), "=d"(v.z), "=d"(v.w)
               : "l"(p));
  return v;
}
// plain (coherent) 256-bit load / store for operands that are also written by the kernel
__device__ __forceinline__ double4v ld_plain4(const double* p) {
  double4v v;
  asm volatile("ld.global.v4.f64 {%0,%1,%2,%3}, [%4];"
               : "=d"(v.x), "=d"(v.y), "=d"(v.z), "=d"(v.w)
               : "l"(p));
  return v;
}
__device__ __forceinline__ void st_plain4(double* p, const double4v& v) {
  asm volatile("st.global.v4.f64 [%0], {%1,%2,%3,%4};" ::"l"(p), "d"(v.x), "d"(v.y), "d"(v.z),
               "d"(v.w)
               : "memory");
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// Publish protocol for host-visible results: the threads of the finishing CTA store the results
// into the mapped mailbox, the CTA synchronises, then ONE thread releases the sequence number the
// host is spinning on.  st.release.sys is cumulative over the stores the barrier ordered before it,
// so no other thread needs a system-scope fence (each one costs ~3 us of kernel tail).
__device__ __forceinline__ void publish_flag(uint64_t* flag, uint64_t seq) {
  __syncthreads();
  if (flag != nullptr && threadIdx.x == 0) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(flag), "l"(seq) : "memory");
  }
}

// Last-CTA election for two-stage deterministic reductions.  Returns true in every thread of
// the CTA that arrived last; the counter resets itself for the next launch.
__device__ __forceinline__ bool last_cta_arrives(unsigned int* ticket, unsigned int total) {
  __shared__ bool is_last;
  // Release: the CTA barrier orders every thread's partial-sum store before thread 0's fence,
  // and the gpu-scope fence is cumulative, so ONE fence per CTA suffices (the pattern of
  // cooperative-groups grid.sync).  A fence in every thread cost ~8 ns per CTA, serialised
  // GPU-wide: +15 us on a 2000-CTA kernel.
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    unsigned int t = atomicAdd(ticket, 1u);
    is_last = (t == total - 1);
    if (is_last) *ticket = 0u;
  }
  __syncthreads();
  if (is_last) __threadfence();
  return is_last;
}

// ---------------------------------------------------------------------------------------------
// Deterministic two-level folds of per-CTA partial sums.
//
// A single "last CTA" that walks all partials serially is latency-bound: every batch of loads
// costs one L2 round trip (~0.4 us), and a fold over 148 ... 4000 partials took 10-20 us — as
// long as the streaming part of a 1M-row kernel.  Here CTAs are grouped; the last CTA to finish
// in a group folds the group's partials (one batch of independent loads), the last group to
// finish folds the group sums.  The summation tree depends only on the launch geometry, never on
// the arrival order, so results are run-to-run reproducible.
//
// tickets[0] is the global counter, tickets[1 + g] the counter of group g (all self-resetting).

// E entries (few), nparts partials per entry (many): ws[e * nparts + i].  The E results of the
// calling grid land in out[e].  `mult` = CTAs that share one partial index (gridDim.y for the
// column reductions).  Returns true in the CTA that wrote the results.
template <int NW, bool IS_MAX = false>
__device__ __forceinline__ bool fold_partials_wide(double* ws, unsigned int* tickets, int E,
                                                   int nparts, int part, int mult, double* out) {
  constexpr int G = 64;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int ngroups = (nparts + G - 1) / G;
  const int g = part / G;
  const int gs = min(G, nparts - g * G);
  if (!last_cta_arrives(&tickets[1 + g], (unsigned int)(gs * mult))) return false;
  double* ws2 = ws + (size_t)E * nparts;
  for (int e = warp; e < E; e += NW) {
    const double* p = ws + (size_t)e * nparts + g * G;
    const double a = lane < gs ? __ldcg(p + lane) : 0.0;
    const double b = lane + 32 < gs ? __ldcg(p + lane + 32) : 0.0;
    const double t = IS_MAX ? warp_max(fmax(a, b)) : warp_sum(a + b);
    if (lane == 0) ws2[(size_t)e * ngroups + g] = t;
  }
  if (!last_cta_arrives(&tickets[0], (unsigned int)ngroups)) return false;
  for (int e = warp; e < E; e += NW) {
    const double* p = ws2 + (size_t)e * ngroups;
    double t0 = 0.0, t1 = 0.0, t2 = 0.0, t3 = 0.0;
    for (int i = lane; i < ngroups; i += 128) {
      const double v0 = __ldcg(p + i);
      const double v1 = i + 32 < ngroups ? __ldcg(p + i + 32) : 0.0;
      const double v2 = i + 64 < ngroups ? __ldcg(p + i + 64) : 0.0;
      const double v3 = i + 96 < ngroups ? __ldcg(p + i + 96) : 0.0;
      if (IS_MAX) {
        t0 = fmax(t0, v0); t1 = fmax(t1, v1); t2 = fmax(t2, v2); t3 = fmax(t3, v3);
      } else {
        t0 += v0; t1 += v1; t2 += v2; t3 += v3;
      }
    }
    const double t = IS_MAX ? warp_max(fmax(fmax(t0, t1), fmax(t2, t3)))
                            : warp_sum((t0 + t1) + (t2 + t3));
    if (lane == 0) out[e] = t;
  }
  return true;
}

// E entries (many), nparts partials per entry (<= a few hundred): ws[i * E + e] (coalesced for
// thread-per-entry access).  store(e, sum) is called once per entry by the finishing CTA.
template <int NT, class Store>
__device__ __forceinline__ bool fold_partials_tall(double* ws, unsigned int* tickets, int E,
                                                   int nparts, int part, Store store) {
  constexpr int G = 16;
  const int ngroups = (nparts + G - 1) / G;
  const int g = part / G;
  const int gs = min(G, nparts - g * G);
  if (!last_cta_arrives(&tickets[1 + g], (unsigned int)gs)) return false;
  double* ws2 = ws + (size_t)E * nparts;
  for (int e = threadIdx.x; e < E; e += NT) {
    const double* p = ws + (size_t)(g * G) * E + e;
    double v[G];
#pragma unroll
    for (int i = 0; i < G; ++i) v[i] = i < gs ? __ldcg(p + (size_t)i * E) : 0.0;
#pragma unroll
    for (int w = 1; w < G; w <<= 1)
#pragma unroll
      for (int i = 0; i < G; i += 2 * w) v[i] += v[i + w];
    ws2[(size_t)g * E + e] = v[0];
  }
  if (!last_cta_arrives(&tickets[0], (unsigned int)ngroups)) return false;
  for (int e = threadIdx.x; e < E; e += NT) {
    const double* p = ws2 + e;
    double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
    int i = 0;
    for (; i + 7 < ngroups; i += 8) {
      const double a0 = __ldcg(p + (size_t)(i + 0) * E), a1 = __ldcg(p + (size_t)(i + 1) * E);
      const double a2 = __ldcg(p + (size_t)(i + 2) * E), a3 = __ldcg(p + (size_t)(i + 3) * E);
      const double a4 = __ldcg(p + (size_t)(i + 4) * E), a5 = __ldcg(p + (size_t)(i + 5) * E);
      const double a6 = __ldcg(p + (size_t)(i + 6) * E), a7 = __ldcg(p + (size_t)(i + 7) * E);
      s0 += a0; s1 += a1; s2 += a2; s3 += a3;
      s0 += a4; s1 += a5; s2 += a6; s3 += a7;
    }
    double r[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) r[u] = i + u < ngroups ? __ldcg(p + (size_t)(i + u) * E) : 0.0;
    s0 += r[0]; s1 += r[1]; s2 += r[2]; s3 += r[3];
    s0 += r[4]; s1 += r[5]; s2 += r[6]; s3 += r[7];
    store(e, (s0 + s1) + (s2 + s3));
  }
  return true;
}


// ---------------------------------------------------------------------------------------------
// mbarrier / TMA (async proxy) helpers shared by the TMA-staged kernels (gram.cu, spmm.cu)
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  uint32_t done;
  do {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(done)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
  } while (!done);
}
__device__ __forceinline__ uint64_t l2_policy_evict_first() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ uint64_t l2_policy_evict_last() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
  return p;
}
// explicit shared-space loads from a 32-bit shared address (a pointer derived from the rounded-up
// dynamic shared base is "generic" to the compiler and would compile to slow generic LD)
__device__ __forceinline__ double lds_f64(uint32_t addr) {
  double v;
  asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ int32_t lds_s32(uint32_t addr) {
  int32_t v;
  asm volatile("ld.shared.s32 %0, [%1];" : "=r"(v) : "r"(addr));
  return v;
}
// Read-only gather load that the compiler may not reorder against other volatile asm statements,
// and a "pin" that makes a value opaque: a batch of ld_gather's followed by pin's keeps every load
// of the batch in flight before the first dependent FMA can stall the warp.
__device__ __forceinline__ double ld_gather(const double* p) {
  double v;
  asm volatile("ld.global.nc.f64 %0, [%1];" : "=d"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ void pin(double& v) { asm volatile("" : "+d"(v)); }
// Asynchronous L2 prefetch of a contiguous range (16-B aligned, size a multiple of 16 B)
__device__ __forceinline__ void bulk_prefetch_l2(const void* src, uint32_t bytes) {
  asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(src), "r"(bytes) : "memory");
}
// 1-D bulk copy global -> shared (SASS: UBLKCP), completion counted on an mbarrier.  Source and
// destination 16-B aligned, size a multiple of 16 B.
__device__ __forceinline__ void bulk_load_1d(void* dst, const void* src, uint32_t bytes,
                                             uint64_t* bar, uint64_t policy) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::
          "r"(smem_u32(dst)),
      "l"(src), "r"(bytes), "r"(smem_u32(bar)), "l"(policy)
      : "memory");
}

// 2-D TMA tensor copy global -> shared (coordinates: c0 = row, c1 = column), completion on an
// mbarrier (SASS: UTMALDG)
// `policy`: L2 eviction policy (createpolicy).  The basis A is touched once per call -> evict_first;
// the block V is re-read once per 32-column block of A -> evict_last.  Without the hints the A
// stream pushed V out of L2 and V came back from DRAM every column block (ncu: 924 MB read for
// 832 algorithmic MB at p = 100, q = 4).
__device__ __forceinline__ void tma_load_2d(void* dst, const void* tmap, int c0, int c1,
                                            uint64_t* bar, uint64_t policy) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1, {%2, %3}], [%4], %5;" ::
          "r"(smem_u32(dst)),
      "l"(reinterpret_cast<uint64_t>(tmap)), "r"(c0), "r"(c1), "r"(smem_u32(bar)), "l"(policy)
      : "memory");
}
template <int NTHREADS>
__device__ __forceinline__ void consumer_bar() {
  asm volatile("bar.sync 1, %0;" ::"n"(NTHREADS) : "memory");
}



// Acquire-spin on a flag another GPU raises over NVLink.  A peer that never arrives (a rank that
// died or left the collective call sequence) must not hang this GPU: after ~30 s the wait gives
// up, records it in the host-visible error word and returns false.
__device__ __forceinline__ bool spin_acquire_sys(const uint64_t* flag, uint64_t want, uint64_t* err) {
  uint64_t v, t0 = 0;
  uint32_t it = 0;
  while (true) {
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(flag) : "memory");
    if (v >= want) return true;
    if ((++it & 0xfffu) == 0) {
      uint64_t now;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
      if (t0 == 0) t0 = now;
      else if (now - t0 > 30000000000ull) {
        if (err) asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(err), "l"((uint64_t)1) : "memory");
        return false;
      }
    }
  }
}


// ---------------------------------------------------------------------------------------------
// Low-latency peer exchange ("LL": value and sequence tag travel in the same store).
// A double is sent as one 16-byte store {lo32, tag, hi32, tag}; each 8-byte half is written
// atomically, so a reader that sees the expected tag in both halves has the whole value -- no
// separate flag, no release fence waiting for the data to be acknowledged across NVLink, and the
// receiver polls the data itself.  (The previous protocol -- data, CTA barrier, st.release.sys per
// peer, acquire-spin on the flags -- paid one extra NVLink round trip per exchange.)
struct __align__(16) LLWord { uint32_t lo, t0, hi, t1; };
__device__ __forceinline__ void ll_store(LLWord* p, double v, uint32_t tag) {
  const uint32_t lo = (uint32_t)__double2loint(v), hi = (uint32_t)__double2hiint(v);
  asm volatile("st.volatile.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(lo), "r"(tag), "r"(hi), "r"(tag)
               : "memory");
}
// false after ~30 s without the tag (a rank died or left the collective call sequence): the wait
// gives up and records it in the host-visible error word instead of hanging the GPU
__device__ __forceinline__ bool ll_load(const LLWord* p, uint32_t tag, double* out, uint64_t* err) {
  uint32_t a, b, c, d, it = 0;
  uint64_t t0 = 0;
  while (true) {
    asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(a), "=r"(b), "=r"(c), "=r"(d) : "l"(p) : "memory");
    if (b == tag && d == tag) break;
    if ((++it & 0xfffu) == 0) {
      uint64_t now;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
      if (t0 == 0) t0 = now;
      else if (now - t0 > 30000000000ull) {
        if (err) asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(err), "l"((uint64_t)1) : "memory");
        return false;
      }
    }
  }
  *out = __hiloint2double((int)c, (int)a);
  return true;
}

// Hand a finished reduction to the host (see XchgPub in runtime.cuh).  Called by every thread of
// the finishing CTA; results of this launch were stored by this CTA, those of earlier launches of
// the same reduction by earlier kernels on the stream.
//
// nranks > 1: all-to-all LL exchange of this rank's partial result, fold in RANK ORDER (bitwise
// identical on every rank, so the replicated host-side Belos state cannot diverge), result into the
// host mailbox, then the host flag.  Window = [parity][source rank][kLL] LLWords; results longer than
// kLL entries go in chunks, one exchange number (tag) per chunk.  A window half is re-used two
// exchanges later; a peer can only be that far ahead after it has received this rank's contribution
// to the exchange in between, which this rank sends after it finished reading (CTA barrier below).
template <class Pub>
__device__ __forceinline__ void publish_result(const Pub& pb) {
  if (pb.nranks <= 1) {
    publish_flag(pb.flag, pb.seq);
    return;
  }
  constexpr int kMaxR = 8, kLL = 8192;  // kMaxPeers, kXchgMax
  __shared__ int timed_out;
  if (threadIdx.x == 0) timed_out = 0;
  __syncthreads();  // (also orders this CTA's result stores before the reads of pb.src below)
  const int P = pb.nranks, me = pb.rank;
  int ch = 0;
  for (int c0 = 0; c0 < pb.count; c0 += kLL, ++ch) {
    const uint64_t xs = pb.xseq + (uint64_t)ch;
    const uint32_t tag = (uint32_t)xs;
    const size_t half = (size_t)(xs & 1u) * kMaxR;
    const int cnt = min(kLL, pb.count - c0);
    for (int i = threadIdx.x; i < cnt; i += blockDim.x) {
      const double v = pb.src[c0 + i];
      for (int r = 0; r < P; ++r)
        if (r != me) ll_store(static_cast<LLWord*>(pb.win[r]) + (half + me) * kLL + i, v, tag);
    }
    const LLWord* mine = static_cast<const LLWord*>(pb.win[me]) + half * kLL;
    for (int i = threadIdx.x; i < cnt; i += blockDim.x) {
      double t = 0.0;
      for (int r = 0; r < P; ++r) {
        double v = 0.0;
        if (r == me) v = pb.src[c0 + i];
        else if (!ll_load(mine + (size_t)r * kLL + i, tag, &v, pb.err)) timed_out = 1;
        t = r == 0 ? v : (pb.is_max ? fmax(t, v) : t + v);
      }
      pb.dst[c0 + i] = t;
    }
    __syncthreads();
    if (timed_out) return;  // nothing is published: the host sees the error word
  }
  publish_flag(pb.flag, pb.seq);
}

}  // namespace b200
#endif
