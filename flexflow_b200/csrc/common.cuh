// common.cuh — shared device/host helpers for the sm_100a kernels of the FlexFlow Serve hot path.
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>

#include "../../include/ffb200.h"

#define FFB_STR2(x) #x
#define FFB_STR(x) FFB_STR2(x)

namespace ffb {

// ---- error plumbing (the reference asserts/exits; the C ABI returns codes) ----
void set_error(char const *fmt, ...);
// Cached getenv for the FFB_* override variables (tests, A/B experiments, profiling): the environment is read once per
// variable and call site, not on every hot call.  `name` must be a string literal (the cache is keyed by its address).
// A process that changes such a variable after the library has read it calls ffb_reload_env().
char const *env(char const *name);
#define FFB_CUDA_TRY(expr)                                                                  \
  do {                                                                                      \
    cudaError_t _e = (expr);                                                                \
    if (_e != cudaSuccess) {                                                                \
      ffb::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
      return FFB_ERR_CUDA;                                                                  \
    }                                                                                       \
  } while (0)
#define FFB_LAUNCH_CHECK()                                                                  \
  do {                                                                                      \
    cudaError_t _e = cudaGetLastError();                                                    \
    if (_e != cudaSuccess) {                                                                \
      ffb::set_error("kernel launch failed: %s (%s:%d)", cudaGetErrorString(_e), __FILE__, __LINE__); \
      return FFB_ERR_CUDA;                                                                  \
    }                                                                                       \
  } while (0)
#define FFB_REQUIRE(cond, ...)                                                              \
  do {                                                                                      \
    if (!(cond)) {                                                                          \
      ffb::set_error(__VA_ARGS__);                                                          \
      return FFB_ERR_INVALID;                                                               \
    }                                                                                       \
  } while (0)

// cudaFuncSetAttribute acts on the CURRENT device only, and one host process may drive several GPUs from several
// threads (the reference's Legion process model): a launch wrapper remembers, per call site, WHICH devices it has
// configured.  Two threads racing here both set the same attribute - harmless.
#define FFB_SET_MAX_SMEM_ONCE(kernel, bytes)                                                                  \
  do {                                                                                                        \
    static std::atomic<unsigned long long> _done{0};                                                          \
    int _dev = 0;                                                                                             \
    cudaGetDevice(&_dev);                                                                                     \
    unsigned long long const _bit = (_dev >= 0 && _dev < 64) ? 1ULL << _dev : 0ULL;                           \
    if (!(_done.load(std::memory_order_acquire) & _bit)) {                                                    \
      FFB_CUDA_TRY(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(bytes)));  \
      _done.fetch_or(_bit, std::memory_order_release);                                                        \
    }                                                                                                         \
  } while (0)

constexpr int kNumSMs = 148; // B200: 2 dies x 74 SMs

// ---- 16-bit storage types ----
template <typename T> struct DT;
template <> struct DT<__half> {
  using T2 = __half2;
  static __device__ __forceinline__ float to_f(__half v) { return __half2float(v); }
  static __device__ __forceinline__ __half from_f(float v) { return __float2half_rn(v); }
  static __device__ __forceinline__ float2 to_f2(uint32_t u) {
    return __half22float2(*reinterpret_cast<__half2 *>(&u));
  }
  static __device__ __forceinline__ uint32_t pack(float a, float b) {
    __half2 h = __floats2half2_rn(a, b);
    return *reinterpret_cast<uint32_t *>(&h);
  }
};
template <> struct DT<__nv_bfloat16> {
  using T2 = __nv_bfloat162;
  static __device__ __forceinline__ float to_f(__nv_bfloat16 v) { return __bfloat162float(v); }
  static __device__ __forceinline__ __nv_bfloat16 from_f(float v) { return __float2bfloat16_rn(v); }
  static __device__ __forceinline__ float2 to_f2(uint32_t u) {
    // bf16 -> f32 is a 16-bit shift
    return make_float2(__uint_as_float(u << 16), __uint_as_float(u & 0xFFFF0000u));
  }
  static __device__ __forceinline__ uint32_t pack(float a, float b) {
    __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t *>(&h);
  }
};

// unpack 8 x 16-bit values held in a uint4 into 8 floats
template <typename T> __device__ __forceinline__ void unpack8(uint4 const &u, float (&f)[8]) {
  float2 a = DT<T>::to_f2(u.x), b = DT<T>::to_f2(u.y), c = DT<T>::to_f2(u.z), d = DT<T>::to_f2(u.w);
  f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
}
template <typename T> __device__ __forceinline__ uint4 pack8(float const (&f)[8]) {
  uint4 u;
  u.x = DT<T>::pack(f[0], f[1]); u.y = DT<T>::pack(f[2], f[3]);
  u.z = DT<T>::pack(f[4], f[5]); u.w = DT<T>::pack(f[6], f[7]);
  return u;
}

// streaming 128-bit global load that does not pollute L1 (KV cache / weights are read once)
__device__ __forceinline__ uint4 ldg_stream(void const *p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}

__device__ __forceinline__ uint2 ldg_stream64(void const *p) {
  uint2 r;
  asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0,%1}, [%2];" : "=r"(r.x), "=r"(r.y) : "l"(p));
  return r;
}

// ---- per-query descriptor written by the prep kernel, consumed by the attention kernels ----
// A query attends keys [0, kv_len) of request slot `slot`; keys >= ntcs are additionally gated by
// bit `bit` of causalMask[slot].mask[key - ntcs] when bit >= 0 (tree / beam modes).
struct QueryDesc {
  int slot;
  int kv_len;
  int ntcs;
  int bit;
  // the query's row of the tree mask, TRANSPOSED by the prep kernel: bit i set <=> tree position i (key ntcs + i) is
  // visible to this query, i.e. bit `bit` of causalMask[slot].mask[i] (0 when bit is outside [0, 64))
  unsigned long long vis;
};

__device__ __forceinline__ float ffb_ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// Exact log-sum-exp merge of the split-KV partial records of one (token, head) followed by the
// reference's 1/(sum + 1e-6) normalisation (inc_…cu:538).  Called by `nthreads` threads with ids
// [0, nthreads); each thread produces elements tid, tid+nthreads, ...  Records are read with ld.cg
// (written by other blocks of the same launch).
template <typename T, int DH>
__device__ __forceinline__ void combine_record(float const *rec, int nvalid, T *out_row, int tid, int nthreads) {
  float M = -INFINITY;
  for (int i = 0; i < nvalid; i++) {
    M = fmaxf(M, __ldcg(rec + i * (DH + 2) + DH));
  }
  float L = 0.f;
  float o[(DH + 31) / 32];
#pragma unroll
  for (int j = 0; j < (DH + 31) / 32; j++) {
    o[j] = 0.f;
  }
  for (int i = 0; i < nvalid; i++) {
    float const mi = __ldcg(rec + i * (DH + 2) + DH);
    float const w = (mi == -INFINITY) ? 0.f : ffb_ex2(mi - M);
    L += __ldcg(rec + i * (DH + 2) + DH + 1) * w;
#pragma unroll
    for (int j = 0; j < (DH + 31) / 32; j++) {
      int const d = tid + j * nthreads;
      if (d < DH) {
        o[j] += __ldcg(rec + i * (DH + 2) + d) * w;
      }
    }
  }
  float const inv = __fdividef(1.f, L + 1.e-6f);
#pragma unroll
  for (int j = 0; j < (DH + 31) / 32; j++) {
    int const d = tid + j * nthreads;
    if (d < DH) {
      out_row[d] = DT<T>::from_f(o[j] * inv);
    }
  }
}

struct RopeParams {
  int apply;
  float theta;
  int llama3;
  float factor, low_ff, high_ff;
  int orig_max_pos;
};

} // namespace ffb
