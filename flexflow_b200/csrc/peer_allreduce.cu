// peer_allreduce.cu — AllReduce(sum) of tree-verify / prompt sized messages (64 KiB ... 16 MiB of 16-bit elements) over
// NVLink peer memory, in ONE kernel: replaces the ncclAllReduce of Kernels::AllReduce::inference_kernel_wrapper
// (src/parallel_ops/kernels/allreduce_kernels.cu:56-78) where the communicator has mapped every rank's receive area
// (comm.cu peer_setup).  Decode-sized messages stay with the fused o_proj kernel of linear_gemv.cu.
//
// Two shots, both by PUSH (remote stores; NVLink reads would pay a round trip per request):
//   1. reduce-scatter: the message is cut into one slice per rank; every rank writes slice d of its partial sums into
//      slot [my rank] of rank d's area, then raises a flag there (fence.sys + release store of the epoch);
//   2. rank d adds the nranks versions of its slice in RANK ORDER in fp32, rounds once, keeps the result and pushes it
//      into slot [d] of every rank's all-gather area, flag again;
//   3. every rank copies the nranks - 1 foreign result slices out of its own area.
// Every slice is further cut into one chunk per CTA, and CTA c of every rank handles chunk c of every slice, so a CTA
// only ever waits for the SAME CTA of its peers: no grid-wide barrier, no co-residency requirement.  One rank computes
// each element and everybody copies it: the result is bit-identical on all ranks.  Areas are double buffered by epoch
// parity (a rank can be at most one collective ahead of the slowest one); flags hold epochs and are never reset.
// Algorithmic NVLink volume per rank: 2 (nranks - 1) / nranks x message bytes out, the same in.
#include <algorithm>
#include <cstdio>
#include <cstring>

#include "common.cuh"
#include "peer.h"

namespace ffb {
namespace {

constexpr int PA_THREADS = 512;

#ifdef FFB_PA_TRACE_BUILD
// nvcc -DFFB_PA_TRACE_BUILD: CTA 0 stamps %globaltimer at its phase boundaries; the host prints them every 16th launch
__device__ long long g_pa_trace[16];
#define PA_TR(ev)                                                    \
  do {                                                               \
    if (blockIdx.x == 0 && threadIdx.x == 0) {                       \
      long long t_;                                                  \
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_));         \
      g_pa_trace[ev] = t_;                                           \
    }                                                                \
  } while (0)
#else
#define PA_TR(ev) ((void)0)
#endif

__device__ __forceinline__ void st_release_sys(int *p, int v) {
  asm volatile("st.release.sys.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
// spin on a flag a peer writes into local memory: relaxed polls, one acquire fence once it has the epoch
__device__ __forceinline__ void wait_flag(int const *p, int epoch) {
  int v;
  do {
    asm volatile("ld.relaxed.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  } while (v != epoch);
  asm volatile("fence.acq_rel.sys;" ::: "memory");
}
// data written by a peer over NVLink into LOCAL memory: L2 is its point of coherence, so an L1-bypassing load is enough
__device__ __forceinline__ uint4 ld_peer_written(void const *p) {
  uint4 r;
  asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p) : "memory");
  return r;
}

// in / out: count elements of T (count % 8 == 0); per = elements per slice, cper = elements per chunk (both % 8 == 0)
template <typename T>
__global__ void __launch_bounds__(PA_THREADS)
peer_allreduce_kernel(PeerTable const pt, T const *in, T *out, size_t count, size_t per, size_t cper) {
  int const N = pt.nranks, me = pt.rank, c = blockIdx.x, tid = threadIdx.x;
  int const epoch = *reinterpret_cast<int volatile *>(pt.big_epoch) + 1;
  size_t const slice_b = pt.big_slice_bytes;
  auto area = [&](int rank, int phase, int src) { // slot [src] of rank's area for this epoch's parity
    return pt.big[rank] + ((size_t)((epoch & 1) * 2 + phase) * N + src) * slice_b;
  };
  auto flag = [&](int rank, int phase, int src) { return pt.big_flags[rank] + ((size_t)phase * N + src) * kPeerBigChunks + c; };
  // this CTA's chunk of slice s: elements [lo(s), hi(s))
  auto lo = [&](int s) { return min(count, (size_t)s * per + min(per, (size_t)c * cper)); };
  auto hi = [&](int s) { return min(count, (size_t)s * per + min(per, (size_t)(c + 1) * cper)); };
  size_t const coff = (size_t)c * cper * sizeof(T); // byte offset of the chunk inside a slot

  PA_TR(0);
  // ---- 1. push my partial sums of every foreign slice ----
  for (int i = 1; i < N; i++) {
    int const d = (me + i) % N;
    size_t const a = lo(d), b = hi(d);
    uint4 *dst = reinterpret_cast<uint4 *>(area(d, 0, me) + coff);
    uint4 const *src = reinterpret_cast<uint4 const *>(in + a);
    for (size_t v = tid; v < (b - a) / 8; v += PA_THREADS) {
      dst[v] = src[v];
    }
  }
  PA_TR(1);
  // bar.sync orders every thread's stores before the releasing threads; st.release.sys is cumulative, so one fence per
  // destination replaces a fence.sys in each of the 512 threads (measured 7 us per phase)
  __syncthreads();
  PA_TR(2);
  if (tid < N && tid != me) {
    st_release_sys(flag(tid, 0, me), epoch);
  }
  PA_TR(3);
  // ---- 2. reduce my slice in rank order, keep it, push it to everybody ----
  if (tid < N && tid != me) {
    wait_flag(flag(me, 0, tid), epoch);
  }
  __syncthreads();
  PA_TR(4);
  {
    size_t const a = lo(me), b = hi(me);
    for (size_t v = tid; v < (b - a) / 8; v += PA_THREADS) {
      float acc[8];
#pragma unroll
      for (int e = 0; e < 8; e++) {
        acc[e] = 0.f;
      }
      for (int s = 0; s < N; s++) {
        uint4 const u = s == me ? reinterpret_cast<uint4 const *>(in + a)[v]
                                : ld_peer_written(reinterpret_cast<uint4 const *>(area(me, 0, s) + coff) + v);
        float f[8];
        unpack8<T>(u, f);
#pragma unroll
        for (int e = 0; e < 8; e++) {
          acc[e] += f[e];
        }
      }
      uint4 const r = pack8<T>(acc);
      reinterpret_cast<uint4 *>(out + a)[v] = r;
      for (int i = 1; i < N; i++) {
        int const d = (me + i) % N;
        reinterpret_cast<uint4 *>(area(d, 1, me) + coff)[v] = r;
      }
    }
  }
  PA_TR(5);
  __syncthreads();
  PA_TR(6);
  if (tid < N && tid != me) {
    st_release_sys(flag(tid, 1, me), epoch);
  }
  // ---- 3. collect the other ranks' result slices ----
  if (tid < N && tid != me) {
    wait_flag(flag(me, 1, tid), epoch);
  }
  __syncthreads();
  PA_TR(7);
  for (int i = 1; i < N; i++) {
    int const s = (me + i) % N;
    size_t const a = lo(s), b = hi(s);
    uint4 const *src = reinterpret_cast<uint4 const *>(area(me, 1, s) + coff);
    uint4 *dst = reinterpret_cast<uint4 *>(out + a);
    for (size_t v = tid; v < (b - a) / 8; v += PA_THREADS) {
      dst[v] = ld_peer_written(src + v);
    }
  }
  // the last CTA to finish closes the epoch (every CTA read it at kernel start)
  __syncthreads();
  PA_TR(8);
  if (tid == 0) {
    int const prev = atomicAdd(pt.big_done, 1);
    if (prev == (int)gridDim.x - 1) {
      *pt.big_done = 0;
      __threadfence();
      *reinterpret_cast<int volatile *>(pt.big_epoch) = epoch;
    }
  }
}

// ---- flag-in-data variant (messages up to kPeerLLMaxBytes) -------------------------------------------------------
// The flag kernel above pays, per phase, a release fence that waits for every remote store to be acknowledged plus the
// flight of a separate flag (measured: 8 us per phase on NVLink 5 / NVSwitch whatever the message size: a 22 us
// floor).  Here every 8 bytes on the wire are {4 bytes of data, epoch}: an 8-byte store is single-copy atomic, so the
// receiver simply re-reads a 16-byte line until both of its halves carry this collective's epoch.  No fence, no flag
// store, no wait between "data written" and "flag visible"; twice the bytes on the wire, which at these sizes costs
// less than the round trips it removes.  The area is used by this kernel only (a payload word of another format could
// look like an epoch), starts zeroed, and epochs start at 1.  Same slicing, same rank-order fp32 sum, same
// one-rank-computes-everybody-copies rule as above: results are bit-identical to the flag kernel's and on all ranks.
__device__ __forceinline__ void st_ll(void *p, uint32_t d0, uint32_t d1, uint32_t flag) {
  asm volatile("st.volatile.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(d0), "r"(flag), "r"(d1), "r"(flag) : "memory");
}
__device__ __forceinline__ uint2 ld_ll(void const *p, uint32_t flag) {
  uint32_t d0, f0, d1, f1;
  do {
    asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(d0), "=r"(f0), "=r"(d1), "=r"(f1) : "l"(p) : "memory");
  } while (f0 != flag || f1 != flag);
  return make_uint2(d0, d1);
}

// count % 8 == 0; per / cper as in the flag kernel.  A thread moves 8 elements (16 payload bytes = two lines) at a time.
template <typename T>
__global__ void __launch_bounds__(PA_THREADS)
peer_allreduce_ll_kernel(PeerTable const pt, T const *in, T *out, size_t count, size_t per, size_t cper) {
  int const N = pt.nranks, me = pt.rank, c = blockIdx.x, tid = threadIdx.x;
  uint32_t const epoch = (uint32_t)(*reinterpret_cast<int volatile *>(pt.ll_epoch) + 1);
  size_t const slice_b = pt.ll_slice_bytes;
  auto area = [&](int rank, int phase, int src) { return pt.ll[rank] + ((size_t)phase * N + src) * slice_b; };
  auto lo = [&](int s) { return min(count, (size_t)s * per + min(per, (size_t)c * cper)); };
  auto hi = [&](int s) { return min(count, (size_t)s * per + min(per, (size_t)(c + 1) * cper)); };
  size_t const coff = (size_t)c * cper * sizeof(T) * 2; // byte offset of the chunk's lines inside a slot
  PA_TR(0);
  // ---- 1. push my partial sums of every foreign slice ----
  for (int i = 1; i < N; i++) {
    int const d = (me + i) % N;
    size_t const a = lo(d), b = hi(d);
    char *dst = area(d, 0, me) + coff;
    uint4 const *src = reinterpret_cast<uint4 const *>(in + a);
    for (size_t v = tid; v < (b - a) / 8; v += PA_THREADS) {
      uint4 const u = src[v];
      st_ll(dst + v * 32, u.x, u.y, epoch);
      st_ll(dst + v * 32 + 16, u.z, u.w, epoch);
    }
  }
  PA_TR(1);
  // ---- 2. reduce my slice in rank order, keep it, push it to everybody ----
  {
    size_t const a = lo(me), b = hi(me);
    for (size_t v = tid; v < (b - a) / 8; v += PA_THREADS) {
      float acc[8];
#pragma unroll
      for (int e = 0; e < 8; e++) {
        acc[e] = 0.f;
      }
      for (int s = 0; s < N; s++) {
        uint4 u;
        if (s == me) {
          u = reinterpret_cast<uint4 const *>(in + a)[v];
        } else {
          char const *src = area(me, 0, s) + coff + v * 32;
          uint2 const l0 = ld_ll(src, epoch), l1 = ld_ll(src + 16, epoch);
          u = make_uint4(l0.x, l0.y, l1.x, l1.y);
        }
        float f[8];
        unpack8<T>(u, f);
#pragma unroll
        for (int e = 0; e < 8; e++) {
          acc[e] += f[e];
        }
      }
      uint4 const r = pack8<T>(acc);
      reinterpret_cast<uint4 *>(out + a)[v] = r;
      for (int i = 1; i < N; i++) {
        int const d = (me + i) % N;
        char *dst = area(d, 1, me) + coff + v * 32;
        st_ll(dst, r.x, r.y, epoch);
        st_ll(dst + 16, r.z, r.w, epoch);
      }
    }
  }
  PA_TR(2);
  // ---- 3. collect the other ranks' result slices ----
  for (int i = 1; i < N; i++) {
    int const s = (me + i) % N;
    size_t const a = lo(s), b = hi(s);
    char const *src = area(me, 1, s) + coff;
    uint4 *dst = reinterpret_cast<uint4 *>(out + a);
    for (size_t v = tid; v < (b - a) / 8; v += PA_THREADS) {
      uint2 const l0 = ld_ll(src + v * 32, epoch), l1 = ld_ll(src + v * 32 + 16, epoch);
      dst[v] = make_uint4(l0.x, l0.y, l1.x, l1.y);
    }
  }
  PA_TR(3);
  // the last CTA to finish closes the epoch (every CTA read it at kernel start)
  __syncthreads();
  if (tid == 0) {
    int const prev = atomicAdd(pt.ll_done, 1);
    if (prev == (int)gridDim.x - 1) {
      *pt.ll_done = 0;
      __threadfence();
      *reinterpret_cast<int volatile *>(pt.ll_epoch) = (int)epoch;
    }
  }
}

// ---- 128-byte lines, one flag per line (messages that fit the flag-in-data area) -------------------------------------
// The 8-byte lines above put twice the message on the wire, and the push rate of this store pattern tops out at
// ~400-500 GB/s per GPU (measured), so from ~1 MiB the wire time dominates.  Here a line is 128 bytes = fifteen
// 8-byte data words + one 8-byte flag, written by EIGHT NEIGHBOURING LANES OF ONE WARP STORE INSTRUCTION (16 bytes
// each) and read the same way: NVLink delivers the 128-byte sector of one store instruction as a unit (the property
// NCCL's LL128 protocol is built on), so a reader that sees the flag sees the fifteen words.  1.07x the message on the
// wire.  Lane piece p (0..7) of line j holds payload words 15 j + 2 p and 15 j + 2 p + 1; piece 7's second word is
// the flag (epoch | magic << 32: no stale word of the 8-byte-line format can equal it).
constexpr unsigned long long kLL128Magic = 0xA5C3F00Dull << 32;

__device__ __forceinline__ void st_l128(void *p, unsigned long long a, unsigned long long b) {
  asm volatile("st.volatile.global.v2.u64 [%0], {%1,%2};" ::"l"(p), "l"(a), "l"(b) : "memory");
}
__device__ __forceinline__ void ld_l128(void const *p, unsigned long long &a, unsigned long long &b) {
  asm volatile("ld.volatile.global.v2.u64 {%0,%1}, [%2];" : "=l"(a), "=l"(b) : "l"(p) : "memory");
}
// whole-warp spin: every 8-lane group re-reads its line until the group's flag (piece 7, second word) is `flag`
__device__ __forceinline__ void ld_l128_wait(void const *p, bool valid, unsigned long long flag, unsigned long long &a,
                                             unsigned long long &b) {
  int const lane = threadIdx.x & 31;
  bool ok;
  do {
    a = b = 0;
    if (valid) {
      ld_l128(p, a, b);
    }
    unsigned long long const f = __shfl_sync(0xffffffffu, b, lane | 7);
    ok = !valid || f == flag;
  } while (!__all_sync(0xffffffffu, ok));
}
template <typename T> __device__ __forceinline__ void unpack4(unsigned long long w, float (&f)[4]) {
  float2 const a = DT<T>::to_f2((uint32_t)w), b = DT<T>::to_f2((uint32_t)(w >> 32));
  f[0] = a.x;
  f[1] = a.y;
  f[2] = b.x;
  f[3] = b.y;
}
template <typename T> __device__ __forceinline__ unsigned long long pack4(float const (&f)[4]) {
  uint32_t const lo = DT<T>::pack(f[0], f[1]), hi = DT<T>::pack(f[2], f[3]);
  return (unsigned long long)lo | ((unsigned long long)hi << 32);
}

// lpc = lines per chunk (a slot holds gridDim.x * lpc lines)
template <typename T>
__global__ void __launch_bounds__(PA_THREADS)
peer_allreduce_l128_kernel(PeerTable const pt, T const *in, T *out, size_t count, size_t per, size_t cper, int lpc,
                           int trigger) {
  // Launched as a programmatic dependent of the kernel that produced `in` (and of the previous collective, whose epoch
  // is read below): wait for it before anything else.  Then let the next kernel of the stream - the next layer's
  // GEMM - start its prologue and weight prefetch beside this one.
  asm volatile("griddepcontrol.wait;" ::: "memory");
  if (trigger) {
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  }
  int const N = pt.nranks, me = pt.rank, c = blockIdx.x, tid = threadIdx.x;
  int const warp = tid >> 5, lane = tid & 31, piece = lane & 7;
  unsigned long long const flag = (unsigned long long)(uint32_t)(*reinterpret_cast<int volatile *>(pt.ll_epoch) + 1) | kLL128Magic;
  size_t const slice_b = pt.ll_slice_bytes;
  auto area = [&](int rank, int phase, int src) { return pt.ll[rank] + ((size_t)phase * N + src) * slice_b + (size_t)c * lpc * 128; };
  auto lo = [&](int s) { return min(count, (size_t)s * per + min(per, (size_t)c * cper)); };
  auto hi = [&](int s) { return min(count, (size_t)s * per + min(per, (size_t)(c + 1) * cper)); };
  constexpr int LPI = PA_THREADS / 8; // lines per CTA iteration
  PA_TR(0);
  // ---- 1. push my partial sums of every foreign slice ----
  for (int i = 1; i < N; i++) {
    int const d = (me + i) % N;
    size_t const a = lo(d), b = hi(d);
    int const nwords = (int)((b - a) / 4), nlines = (nwords + 14) / 15;
    unsigned long long const *src = reinterpret_cast<unsigned long long const *>(in + a);
    char *dst = area(d, 0, me);
    for (int base = warp * 4; base < nlines; base += LPI) {
      int const line = base + (lane >> 3);
      if (line < nlines) {
        int const w = line * 15 + 2 * piece;
        unsigned long long const w0 = w < nwords ? src[w] : 0ull;
        unsigned long long const w1 = piece == 7 ? flag : (w + 1 < nwords ? src[w + 1] : 0ull);
        st_l128(dst + (size_t)line * 128 + piece * 16, w0, w1);
      }
    }
  }
  PA_TR(1);
  // ---- 2. reduce my slice in rank order, keep it, push it to everybody ----
  {
    size_t const a = lo(me), b = hi(me);
    int const nwords = (int)((b - a) / 4), nlines = (nwords + 14) / 15;
    unsigned long long const *mine = reinterpret_cast<unsigned long long const *>(in + a);
    unsigned long long *res = reinterpret_cast<unsigned long long *>(out + a);
    for (int base = warp * 4; base < nlines; base += LPI) {
      int const line = base + (lane >> 3);
      bool const valid = line < nlines;
      int const w = line * 15 + 2 * piece;
      bool const has0 = valid && w < nwords, has1 = valid && piece != 7 && w + 1 < nwords;
      float acc0[4] = {0.f, 0.f, 0.f, 0.f}, acc1[4] = {0.f, 0.f, 0.f, 0.f};
      for (int s = 0; s < N; s++) {
        unsigned long long w0 = 0, w1 = 0;
        if (s == me) {
          w0 = has0 ? mine[w] : 0ull;
          w1 = has1 ? mine[w + 1] : 0ull;
        } else {
          ld_l128_wait(area(me, 0, s) + (size_t)line * 128 + piece * 16, valid, flag, w0, w1);
        }
        float f[4];
        unpack4<T>(w0, f);
#pragma unroll
        for (int e = 0; e < 4; e++) {
          acc0[e] += f[e];
        }
        unpack4<T>(w1, f); // (piece 7: the flag word; its sum is never used)
#pragma unroll
        for (int e = 0; e < 4; e++) {
          acc1[e] += f[e];
        }
      }
      unsigned long long const r0 = pack4<T>(acc0), r1 = piece == 7 ? flag : pack4<T>(acc1);
      if (has0) {
        res[w] = r0;
      }
      if (has1) {
        res[w + 1] = r1;
      }
      if (valid) {
        for (int i = 1; i < N; i++) {
          int const d = (me + i) % N;
          st_l128(area(d, 1, me) + (size_t)line * 128 + piece * 16, has0 ? r0 : 0ull, piece == 7 ? flag : (has1 ? r1 : 0ull));
        }
      }
    }
  }
  PA_TR(2);
  // ---- 3. collect the other ranks' result slices ----
  for (int i = 1; i < N; i++) {
    int const s = (me + i) % N;
    size_t const a = lo(s), b = hi(s);
    int const nwords = (int)((b - a) / 4), nlines = (nwords + 14) / 15;
    unsigned long long *res = reinterpret_cast<unsigned long long *>(out + a);
    char const *src = area(me, 1, s);
    for (int base = warp * 4; base < nlines; base += LPI) {
      int const line = base + (lane >> 3);
      bool const valid = line < nlines;
      int const w = line * 15 + 2 * piece;
      unsigned long long w0, w1;
      ld_l128_wait(src + (size_t)line * 128 + piece * 16, valid, flag, w0, w1);
      if (valid && w < nwords) {
        res[w] = w0;
      }
      if (valid && piece != 7 && w + 1 < nwords) {
        res[w + 1] = w1;
      }
    }
  }
  PA_TR(3);
  __syncthreads();
  if (tid == 0) {
    int const prev = atomicAdd(pt.ll_done, 1);
    if (prev == (int)gridDim.x - 1) {
      *pt.ll_done = 0;
      __threadfence();
      *reinterpret_cast<int volatile *>(pt.ll_epoch) = (int)(uint32_t)flag;
    }
  }
}

} // namespace

int peer_allreduce_supported(PeerTable const *pt, size_t count, int dtype) {
  return pt && pt->nranks >= 2 && pt->big_slice_bytes > 0 && (dtype == FFB_DT_HALF || dtype == FFB_DT_BF16) && count % 8 == 0 &&
         count * 2 <= kPeerBigMaxBytes && count >= 8;
}

int launch_peer_allreduce(PeerTable const *pt, void const *in, void *out, size_t count, int dtype, cudaStream_t s) {
  int const N = pt->nranks;
  // slice per rank and chunk per CTA, both multiples of 8 elements; chunks of >= 2 KiB keep the flag traffic small
  size_t const per = ((count + N - 1) / N + 7) & ~(size_t)7;
  // FFB_PEER_GRID / FFB_PEER_CHUNK: experiment knobs (CTA cap, minimum chunk bytes)
  static int const grid_cap = env("FFB_PEER_GRID") ? std::max(1, std::min(kPeerBigChunks, atoi(env("FFB_PEER_GRID")))) : 128;
  static size_t const min_chunk = env("FFB_PEER_CHUNK") ? (size_t)std::max(256, atoi(env("FFB_PEER_CHUNK"))) : 2048;
  int grid = (int)std::min<size_t>(grid_cap, std::max<size_t>(1, per * 2 / min_chunk));
  size_t const cper = ((per + grid - 1) / grid + 7) & ~(size_t)7;
  grid = (int)((per + cper - 1) / cper);
  if (per * 2 > pt->big_slice_bytes) {
    set_error("allreduce message of %zu elements exceeds the peer area", count);
    return FFB_ERR_INVALID;
  }
  // Protocol: 128-byte lines with one flag per line when the message fits the flag-in-data area (default), 8-byte
  // lines (FFB_PEER_PROTO=ll, up to kPeerLLMaxBytes), or separate flags (FFB_PEER_PROTO=flag, and anything larger).
  // Chosen once per process: the two line formats share the area and must not alternate.
  static int const proto = !env("FFB_PEER_PROTO") ? 2 : (strcmp(env("FFB_PEER_PROTO"), "ll") == 0 ? 1 : (strcmp(env("FFB_PEER_PROTO"), "flag") == 0 ? 0 : 2));
  if (proto == 2 && pt->ll_slice_bytes > 0) {
    static int const cap_env = env("FFB_PEER_LL_GRID") ? std::max(1, std::min(kPeerBigChunks, atoi(env("FFB_PEER_LL_GRID")))) : 0;
    int const cap = cap_env ? cap_env : (count * 2 > ((size_t)4 << 20) ? 2 * kNumSMs : kNumSMs); // measured, 2 ranks
    int g = (int)std::min<size_t>(cap, std::max<size_t>(1, per * 2 / min_chunk));
    size_t const lcper = ((per + g - 1) / g + 7) & ~(size_t)7;
    g = (int)((per + lcper - 1) / lcper);
    int const lpc = (int)((lcper / 4 + 14) / 15);
    if ((size_t)g * lpc * 128 <= pt->ll_slice_bytes) {
      // FFB_AR_PDL: bit 0 = launch as a programmatic dependent, bit 1 = trigger the next kernel early
      static int const ar_pdl = env("FFB_AR_PDL") ? atoi(env("FFB_AR_PDL")) : 3;
      cudaLaunchConfig_t cfg = {};
      cfg.gridDim = dim3(g);
      cfg.blockDim = dim3(PA_THREADS);
      cfg.stream = s;
      cudaLaunchAttribute attr[1];
      attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
      attr[0].val.programmaticStreamSerializationAllowed = 1;
      cfg.attrs = attr;
      cfg.numAttrs = (ar_pdl & 1) ? 1 : 0;
      if (dtype == FFB_DT_HALF) {
        FFB_CUDA_TRY(cudaLaunchKernelEx(&cfg, peer_allreduce_l128_kernel<__half>, *pt, (__half const *)in, (__half *)out, count, per,
                                        lcper, lpc, (ar_pdl >> 1) & 1));
      } else {
        FFB_CUDA_TRY(cudaLaunchKernelEx(&cfg, peer_allreduce_l128_kernel<__nv_bfloat16>, *pt, (__nv_bfloat16 const *)in,
                                        (__nv_bfloat16 *)out, count, per, lcper, lpc, (ar_pdl >> 1) & 1));
      }
      FFB_LAUNCH_CHECK();
      return FFB_OK;
    }
  }
  bool const ll = proto == 1 && pt->ll_slice_bytes > 0 && count * 2 <= kPeerLLMaxBytes && per * 4 <= pt->ll_slice_bytes;
  if (ll) {
    // measured at 2 ranks (tools/bench_allreduce.py): two CTAs per SM up to 4 MiB (26.6 us against 28.8 with one), one above
    static int const ll_cap_env = env("FFB_PEER_LL_GRID") ? std::max(1, std::min(kPeerBigChunks, atoi(env("FFB_PEER_LL_GRID")))) : 0;
    int const ll_cap = ll_cap_env ? ll_cap_env : (count * 2 <= ((size_t)4 << 20) ? 2 * kNumSMs : kNumSMs);
    grid = (int)std::min<size_t>(ll_cap, std::max<size_t>(1, per * 2 / min_chunk));
    size_t const lcper = ((per + grid - 1) / grid + 7) & ~(size_t)7;
    grid = (int)((per + lcper - 1) / lcper);
    if (dtype == FFB_DT_HALF) {
      peer_allreduce_ll_kernel<__half><<<grid, PA_THREADS, 0, s>>>(*pt, (__half const *)in, (__half *)out, count, per, lcper);
    } else {
      peer_allreduce_ll_kernel<__nv_bfloat16><<<grid, PA_THREADS, 0, s>>>(*pt, (__nv_bfloat16 const *)in, (__nv_bfloat16 *)out, count, per, lcper);
    }
  } else if (dtype == FFB_DT_HALF) {
    peer_allreduce_kernel<__half><<<grid, PA_THREADS, 0, s>>>(*pt, (__half const *)in, (__half *)out, count, per, cper);
  } else {
    peer_allreduce_kernel<__nv_bfloat16><<<grid, PA_THREADS, 0, s>>>(*pt, (__nv_bfloat16 const *)in, (__nv_bfloat16 *)out, count, per, cper);
  }
  FFB_LAUNCH_CHECK();
#ifdef FFB_PA_TRACE_BUILD
  static int calls = 0;
  if (++calls % 16 == 0) {
    long long h[16];
    cudaStreamSynchronize(s);
    cudaMemcpyFromSymbol(h, g_pa_trace, sizeof(h));
    fprintf(stderr, "PATRACE rank %d count %zu grid %d:", pt->rank, count, grid);
    for (int e = 1; e <= 8; e++) {
      fprintf(stderr, " %lld", h[e] - h[0]);
    }
    fprintf(stderr, "\n");
  }
#endif
  return FFB_OK;
}

} // namespace ffb
