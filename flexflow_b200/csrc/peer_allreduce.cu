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

#include "common.cuh"
#include "peer.h"

namespace ffb {
namespace {

constexpr int PA_THREADS = 512;

__device__ __forceinline__ void st_release_sys(int *p, int v) {
  asm volatile("st.release.sys.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ int ld_acquire_sys(int const *p) {
  int v;
  asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
// data written by a peer over NVLink: read it from L2 / HBM, never from a stale L1 line
__device__ __forceinline__ uint4 ld_peer_written(void const *p) {
  uint4 r;
  asm volatile("ld.relaxed.sys.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p) : "memory");
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
  __threadfence_system();
  __syncthreads();
  if (tid < N && tid != me) {
    st_release_sys(flag(tid, 0, me), epoch);
  }
  // ---- 2. reduce my slice in rank order, keep it, push it to everybody ----
  if (tid < N && tid != me) {
    while (ld_acquire_sys(flag(me, 0, tid)) != epoch) {
    }
  }
  __syncthreads();
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
  __threadfence_system();
  __syncthreads();
  if (tid < N && tid != me) {
    st_release_sys(flag(tid, 1, me), epoch);
  }
  // ---- 3. collect the other ranks' result slices ----
  if (tid < N && tid != me) {
    while (ld_acquire_sys(flag(me, 1, tid)) != epoch) {
    }
  }
  __syncthreads();
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
  if (tid == 0) {
    int const prev = atomicAdd(pt.big_done, 1);
    if (prev == (int)gridDim.x - 1) {
      *pt.big_done = 0;
      __threadfence();
      *reinterpret_cast<int volatile *>(pt.big_epoch) = epoch;
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
  int grid = (int)std::min<size_t>(kPeerBigChunks, std::max<size_t>(1, per * 2 / 2048));
  size_t const cper = ((per + grid - 1) / grid + 7) & ~(size_t)7;
  grid = (int)((per + cper - 1) / cper);
  if (per * 2 > pt->big_slice_bytes) {
    set_error("allreduce message of %zu elements exceeds the peer area", count);
    return FFB_ERR_INVALID;
  }
  if (dtype == FFB_DT_HALF) {
    peer_allreduce_kernel<__half><<<grid, PA_THREADS, 0, s>>>(*pt, (__half const *)in, (__half *)out, count, per, cper);
  } else {
    peer_allreduce_kernel<__nv_bfloat16><<<grid, PA_THREADS, 0, s>>>(*pt, (__nv_bfloat16 const *)in, (__nv_bfloat16 *)out, count, per, cper);
  }
  FFB_LAUNCH_CHECK();
  return FFB_OK;
}

} // namespace ffb
