// norm_argmax.cu — the two small operators either side of the attention chain that the reference's FusedOp already
// spans:
//   * (Residual)RMSNorm forward — ResidualRMSNormFusedForwardKernel (src/ops/kernels/residual_rms_norm_kernels.cu:
//     97-127) and RMSNormFusedForwardKernel (src/ops/kernels/rms_norm_kernels.cu:97-124): the input of qkv_proj.
//   * ArgMax — ArgMax::forward_kernel (src/ops/argmax.cu:64-102: cub::DeviceSegmentedReduce::ArgMax + copy_result):
//     the token ids that the accept / reject walk consumes.
// Both are latency / HBM-bound row operators.  The norm keeps the whole row in registers (one pass over HBM: the
// reference re-reads X_out and Y from global memory) and reproduces the reference's roundings AND its summation
// order, so the result is the reference's bit for bit up to the 2-ulp freedom of rsqrtf:
//   - the reference block has B = min(1024, N) threads; thread t sums x[t], x[t+B], ... sequentially in fp32 (the
//     squares of 16-bit values are exact in fp32, so FMA contraction does not matter), warps are reduced by a
//     shfl_down tree (16, 8, 4, 2, 1), the warp sums by a second tree in warp 0 (residual_rms_norm_kernels.cu:71-95);
//   - here one thread owns 8 CONSECUTIVE reference threads (one 128-bit load per B-sized segment), so four lanes
//     make one reference warp: tree steps 16 and 8 are shuffles by 2 and 1 lanes, steps 4, 2, 1 stay in registers.
#include <climits>

#include "common.cuh"

namespace ffb {
namespace {

constexpr int RN_MAXSEG = 16; // row length <= 16 * 1024 elements

template <typename T> __device__ __forceinline__ uint32_t add2_dt(uint32_t a, uint32_t b);
template <> __device__ __forceinline__ uint32_t add2_dt<__half>(uint32_t a, uint32_t b) {
  __half2 r = __hadd2(*reinterpret_cast<__half2 *>(&a), *reinterpret_cast<__half2 *>(&b)); // two independent RN adds
  return *reinterpret_cast<uint32_t *>(&r);
}
template <> __device__ __forceinline__ uint32_t add2_dt<__nv_bfloat16>(uint32_t a, uint32_t b) {
  float2 const x = DT<__nv_bfloat16>::to_f2(a), y = DT<__nv_bfloat16>::to_f2(b);
  return DT<__nv_bfloat16>::pack(x.x + y.x, x.y + y.y); // exact fp32 sum rounded once = the bf16 add
}

// One CTA per token row.  blockDim.x = B / 8 rounded up to whole warps, B = min(1024, N) (B % 32 == 0).
template <typename T, int NSEG>
__global__ void __launch_bounds__(128)
rms_norm_kernel(T const *__restrict__ x1, T const *__restrict__ x2, T const *__restrict__ weight,
                T *__restrict__ residual_out, T *__restrict__ out, int N, int B, float eps) {
  int const row = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  size_t const base = (size_t)row * N;
  bool const active = tid * 8 < B; // threads past B/8 only take part in the shuffles
  uint4 x[NSEG];
  float part[8]; // sums of reference threads 8*tid .. 8*tid+7
#pragma unroll
  for (int e = 0; e < 8; e++) {
    part[e] = 0.f;
  }
#pragma unroll
  for (int s = 0; s < NSEG; s++) {
    int const j = s * B + tid * 8;
    x[s] = make_uint4(0, 0, 0, 0);
    if (active && j < N) {
      x[s] = *reinterpret_cast<uint4 const *>(x1 + base + j);
      if (x2 != nullptr) {
        uint4 const y = *reinterpret_cast<uint4 const *>(x2 + base + j);
        x[s].x = add2_dt<T>(x[s].x, y.x);
        x[s].y = add2_dt<T>(x[s].y, y.y);
        x[s].z = add2_dt<T>(x[s].z, y.z);
        x[s].w = add2_dt<T>(x[s].w, y.w);
        *reinterpret_cast<uint4 *>(residual_out + base + j) = x[s];
      }
      float f[8];
      unpack8<T>(x[s], f);
#pragma unroll
      for (int e = 0; e < 8; e++) {
        part[e] = fmaf(f[e], f[e], part[e]);
      }
    }
  }
  // reference warp tree: lanes l and l+16, then l and l+8 -> threads +2 and +1; then elements +4, +2, +1
#pragma unroll
  for (int e = 0; e < 8; e++) {
    part[e] += __shfl_down_sync(0xffffffffu, part[e], 2);
  }
#pragma unroll
  for (int e = 0; e < 8; e++) {
    part[e] += __shfl_down_sync(0xffffffffu, part[e], 1);
  }
  float const s4a = part[0] + part[4], s4b = part[1] + part[5], s4c = part[2] + part[6], s4d = part[3] + part[7];
  float const wsum = (s4a + s4c) + (s4b + s4d); // steps 2 and 1: (l0 + l2) + (l1 + l3)
  __shared__ float v_shared[32];
  __shared__ float s_rms;
  if (tid < 32) {
    v_shared[tid] = 0.f; // reference: lanes >= blockDim/32 contribute T(0)
  }
  __syncthreads();
  if ((lane & 3) == 0 && active) {
    v_shared[warp * 8 + (lane >> 2)] = wsum; // reference warp index = tid / 4
  }
  __syncthreads();
  if (warp == 0) {
    float v = v_shared[lane];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      v += __shfl_down_sync(0xffffffffu, v, o);
    }
    if (lane == 0) {
      // rms[i] = static_cast<T>(rsqrt(sum / N + eps)) — stored in DT, read back as float
      s_rms = DT<T>::to_f(DT<T>::from_f(rsqrtf(v / (float)N + eps)));
    }
  }
  __syncthreads();
  float const rms = s_rms;
#pragma unroll
  for (int s = 0; s < NSEG; s++) {
    int const j = s * B + tid * 8;
    if (active && j < N) {
      float f[8], w[8], y[8];
      unpack8<T>(x[s], f);
      unpack8<T>(__ldg(reinterpret_cast<uint4 const *>(weight + j)), w);
#pragma unroll
      for (int e = 0; e < 8; e++) {
        float const yn = DT<T>::to_f(DT<T>::from_f(f[e] * rms)); // Y[index] in DT
        y[e] = yn * w[e];
      }
      *reinterpret_cast<uint4 *>(out + base + j) = pack8<T>(y);
    }
  }
}

// One CTA per row; ties go to the smallest index (cub::ArgMax), NaNs never win.
template <typename T>
__global__ void __launch_bounds__(1024)
argmax_kernel(T const *__restrict__ in, int *__restrict__ indices, float *__restrict__ prob, int *__restrict__ parent,
              int V) {
  int const row = blockIdx.x, tid = threadIdx.x;
  T const *p = in + (size_t)row * V;
  float best = -INFINITY;
  int bi = INT_MAX;
  int const nvec = ((size_t)row * V) % 8 == 0 ? V / 8 : 0; // 128-bit loads only when the row is 16-byte aligned
  int const nthr = blockDim.x;
  for (int c0 = tid; c0 < nvec; c0 += 4 * nthr) { // four independent 128-bit loads in flight per thread
    uint4 v[4];
#pragma unroll
    for (int u = 0; u < 4; u++) {
      int const c = c0 + u * nthr;
      v[u] = c < nvec ? ldg_stream(p + (size_t)c * 8) : make_uint4(0, 0, 0, 0);
    }
#pragma unroll
    for (int u = 0; u < 4; u++) {
      int const c = c0 + u * nthr;
      if (c < nvec) {
        float f[8];
        unpack8<T>(v[u], f);
#pragma unroll
        for (int e = 0; e < 8; e++) { // increasing index inside a thread: strict > keeps the first maximum
          if (f[e] > best || (f[e] == best && c * 8 + e < bi)) {
            best = f[e];
            bi = c * 8 + e;
          }
        }
      }
    }
  }
  for (int i = nvec * 8 + tid; i < V; i += blockDim.x) {
    float const v = DT<T>::to_f(p[i]);
    if (v > best || (v == best && i < bi)) {
      best = v;
      bi = i;
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    float const ov = __shfl_xor_sync(0xffffffffu, best, o);
    int const oi = __shfl_xor_sync(0xffffffffu, bi, o);
    if (ov > best || (ov == best && oi < bi)) {
      best = ov;
      bi = oi;
    }
  }
  __shared__ float sv[32];
  __shared__ int si[32];
  if ((tid & 31) == 0) {
    sv[tid >> 5] = best;
    si[tid >> 5] = bi;
  }
  __syncthreads();
  if (tid < 32) {
    int const nw = (blockDim.x + 31) >> 5;
    best = tid < nw ? sv[tid] : -INFINITY;
    bi = tid < nw ? si[tid] : INT_MAX;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      float const ov = __shfl_xor_sync(0xffffffffu, best, o);
      int const oi = __shfl_xor_sync(0xffffffffu, bi, o);
      if (ov > best || (ov == best && oi < bi)) {
        best = ov;
        bi = oi;
      }
    }
    if (tid == 0) {
      indices[row] = bi == INT_MAX ? 0 : bi;
      if (prob != nullptr) {
        prob[row] = best; // copy_result: static_cast<float>(d_out[i].value)
      }
      if (parent != nullptr) {
        parent[row] = 0; // cudaMemsetAsync(parent, 0, ...) in the arg-top-1 beam case (argmax.cu:77-80)
      }
    }
  }
}

// ArgTopK (src/ops/arg_topk.cu:340-449, the TensorFlow heap kernel launched with k threads per row): the k largest
// entries of each row, best first, equal values in increasing index order.  One CTA per row; round r finds the maximum
// among the entries that come strictly after round r-1's winner in that (value desc, index asc) order, so the result
// is the reference's for any input, ties included.  k <= 16 (the reference's beam widths are <= 3).
template <typename T>
__global__ void __launch_bounds__(1024)
arg_topk_kernel(T const *__restrict__ in, float *__restrict__ values, int *__restrict__ indices, int V, int k) {
  int const row = blockIdx.x, tid = threadIdx.x, nthr = blockDim.x;
  T const *p = in + (size_t)row * V;
  int const nvec = ((size_t)row * V) % 8 == 0 ? V / 8 : 0;
  __shared__ float sv[32];
  __shared__ int si[32];
  __shared__ float s_pv;
  __shared__ int s_pi;
  float pv = INFINITY; // previous winner: nothing precedes the first round
  int pi = -1;
  for (int r = 0; r < k; r++) {
    float best = -INFINITY;
    int bi = INT_MAX;
    auto consider = [&](float v, int i) {
      bool const eligible = v < pv || (v == pv && i > pi);
      if (eligible && (v > best || (v == best && i < bi))) {
        best = v;
        bi = i;
      }
    };
    for (int c = tid; c < nvec; c += nthr) {
      float f[8];
      unpack8<T>(__ldg(reinterpret_cast<uint4 const *>(p + (size_t)c * 8)), f); // rows are re-read k times: keep them in L1
#pragma unroll
      for (int e = 0; e < 8; e++) {
        consider(f[e], c * 8 + e);
      }
    }
    for (int i = nvec * 8 + tid; i < V; i += nthr) {
      consider(DT<T>::to_f(p[i]), i);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      float const ov = __shfl_xor_sync(0xffffffffu, best, o);
      int const oi = __shfl_xor_sync(0xffffffffu, bi, o);
      if (ov > best || (ov == best && oi < bi)) {
        best = ov;
        bi = oi;
      }
    }
    if ((tid & 31) == 0) {
      sv[tid >> 5] = best;
      si[tid >> 5] = bi;
    }
    __syncthreads();
    if (tid < 32) {
      int const nw = (nthr + 31) >> 5;
      best = tid < nw ? sv[tid] : -INFINITY;
      bi = tid < nw ? si[tid] : INT_MAX;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        float const ov = __shfl_xor_sync(0xffffffffu, best, o);
        int const oi = __shfl_xor_sync(0xffffffffu, bi, o);
        if (ov > best || (ov == best && oi < bi)) {
          best = ov;
          bi = oi;
        }
      }
      if (tid == 0) {
        s_pv = best;
        s_pi = bi;
        indices[(size_t)row * k + r] = bi == INT_MAX ? 0 : bi;
        if (values != nullptr) {
          values[(size_t)row * k + r] = best;
        }
      }
    }
    __syncthreads();
    pv = s_pv;
    pi = s_pi;
  }
}

// k <= KM (the reference's beam widths: <= 3): ONE pass over the row.  Every thread keeps its own KM best entries sorted
// in registers (a thread meets its entries in increasing index order, so a plain `>` keeps equal values in index order);
// then k rounds of a block-wide (value desc, index asc) maximum over the threads' current heads, the winner popping its
// head.  Same result as arg_topk_kernel, a third of its time at k = 3.
template <typename T, int KM>
__global__ void __launch_bounds__(1024)
arg_topk_small_kernel(T const *__restrict__ in, float *__restrict__ values, int *__restrict__ indices, int V, int k) {
  int const row = blockIdx.x, tid = threadIdx.x, nthr = blockDim.x;
  T const *p = in + (size_t)row * V;
  int const nvec = ((size_t)row * V) % 8 == 0 ? V / 8 : 0;
  float tv[KM];
  int ti[KM];
#pragma unroll
  for (int j = 0; j < KM; j++) {
    tv[j] = -INFINITY;
    ti[j] = INT_MAX;
  }
  auto insert = [&](float v, int i) {
    if (!(v > tv[KM - 1])) { // also rejects NaN
      return;
    }
#pragma unroll
    for (int j = KM - 1; j >= 0; j--) {
      if (v > tv[j]) { // shift entry j down, take its place (undone by the next iteration if v belongs higher)
        if (j + 1 < KM) {
          tv[j + 1] = tv[j];
          ti[j + 1] = ti[j];
        }
        tv[j] = v;
        ti[j] = i;
      }
    }
  };
  for (int c0 = tid; c0 < nvec; c0 += 4 * nthr) {
    uint4 v[4];
#pragma unroll
    for (int u = 0; u < 4; u++) {
      int const c = c0 + u * nthr;
      v[u] = c < nvec ? ldg_stream(p + (size_t)c * 8) : make_uint4(0, 0, 0, 0);
    }
#pragma unroll
    for (int u = 0; u < 4; u++) {
      int const c = c0 + u * nthr;
      if (c < nvec) {
        float f[8];
        unpack8<T>(v[u], f);
#pragma unroll
        for (int e = 0; e < 8; e++) {
          insert(f[e], c * 8 + e);
        }
      }
    }
  }
  for (int i = nvec * 8 + tid; i < V; i += nthr) {
    insert(DT<T>::to_f(p[i]), i);
  }
  __shared__ float sv[32];
  __shared__ int si[32];
  __shared__ int s_win;
  for (int r = 0; r < k; r++) {
    float best = tv[0];
    int bi = ti[0];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      float const ov = __shfl_xor_sync(0xffffffffu, best, o);
      int const oi = __shfl_xor_sync(0xffffffffu, bi, o);
      if (ov > best || (ov == best && oi < bi)) {
        best = ov;
        bi = oi;
      }
    }
    if ((tid & 31) == 0) {
      sv[tid >> 5] = best;
      si[tid >> 5] = bi;
    }
    __syncthreads();
    if (tid < 32) {
      int const nw = (nthr + 31) >> 5;
      best = tid < nw ? sv[tid] : -INFINITY;
      bi = tid < nw ? si[tid] : INT_MAX;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        float const ov = __shfl_xor_sync(0xffffffffu, best, o);
        int const oi = __shfl_xor_sync(0xffffffffu, bi, o);
        if (ov > best || (ov == best && oi < bi)) {
          best = ov;
          bi = oi;
        }
      }
      if (tid == 0) {
        s_win = bi;
        indices[(size_t)row * k + r] = bi == INT_MAX ? 0 : bi;
        if (values != nullptr) {
          values[(size_t)row * k + r] = best;
        }
      }
    }
    __syncthreads();
    if (ti[0] == s_win && s_win != INT_MAX) { // the winner pops its head (indices are unique)
#pragma unroll
      for (int j = 0; j + 1 < KM; j++) {
        tv[j] = tv[j + 1];
        ti[j] = ti[j + 1];
      }
      tv[KM - 1] = -INFINITY;
      ti[KM - 1] = INT_MAX;
    }
    __syncthreads(); // s_win is rewritten by the next round
  }
}

template <typename T>
int launch_rms_norm(void const *x1, void const *x2, void const *w, void *res, void *out, int T_rows, int N, float eps,
                    cudaStream_t s) {
  int const B = N < 1024 ? N : 1024;
  int const nseg = (N + B - 1) / B;
  int const threads = ((B / 8 + 31) / 32) * 32;
#define FFB_RN(NS)                                                                                             \
  rms_norm_kernel<T, NS><<<T_rows, threads, 0, s>>>((T const *)x1, (T const *)x2, (T const *)w, (T *)res, (T *)out, N, B, eps)
  if (nseg <= 1) FFB_RN(1);
  else if (nseg <= 2) FFB_RN(2);
  else if (nseg <= 4) FFB_RN(4);
  else if (nseg <= 8) FFB_RN(8);
  else FFB_RN(RN_MAXSEG);
#undef FFB_RN
  FFB_LAUNCH_CHECK();
  return FFB_OK;
}


// Softmax over the last dimension (the op between lm_head and ArgMax / ArgTopK, inference/models/llama.cc:277-291; the
// reference calls cudnnSoftmaxForward(ACCURATE, MODE_CHANNEL), src/ops/kernels/softmax.cu:262-288): one CTA per row,
// the row is read ONCE from HBM (up to SM_REGS_ELEMS classes it stays in registers, longer rows are re-read from L2),
// maximum and sum in fp32, exp through ex2, one rounding to DT at the end.
constexpr int SM_THREADS = 1024, SM_VEC = 4; // 4 x 8 elements per thread in registers = 32768 classes
template <typename T>
__global__ void __launch_bounds__(SM_THREADS)
softmax_kernel(T const *__restrict__ in, T *__restrict__ out, int V) {
  int const row = blockIdx.x, tid = threadIdx.x;
  T const *p = in + (size_t)row * V;
  T *q = out + (size_t)row * V;
  __shared__ float red[32];
  __shared__ float bcast;
  bool const vec_ok = ((size_t)row * V) % 8 == 0 && V % 8 == 0 && V <= SM_THREADS * SM_VEC * 8;
  auto block_reduce = [&](float v, bool is_max) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      float const w = __shfl_xor_sync(0xffffffffu, v, o);
      v = is_max ? fmaxf(v, w) : v + w;
    }
    if ((tid & 31) == 0) {
      red[tid >> 5] = v;
    }
    __syncthreads();
    if (tid < 32) {
      float x = tid < (SM_THREADS >> 5) ? red[tid] : (is_max ? -INFINITY : 0.f);
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        float const w = __shfl_xor_sync(0xffffffffu, x, o);
        x = is_max ? fmaxf(x, w) : x + w;
      }
      if (tid == 0) {
        bcast = x;
      }
    }
    __syncthreads();
    float const r = bcast;
    __syncthreads();
    return r;
  };
  constexpr float kLog2e = 1.4426950408889634f;
  if (vec_ok) {
    float f[SM_VEC][8];
    int const nvec = V / 8;
    float mx = -INFINITY;
#pragma unroll
    for (int u = 0; u < SM_VEC; u++) {
      int const c = tid + u * SM_THREADS;
      if (c < nvec) {
        unpack8<T>(ldg_stream(p + (size_t)c * 8), f[u]);
#pragma unroll
        for (int e = 0; e < 8; e++) {
          mx = fmaxf(mx, f[u][e]);
        }
      }
    }
    mx = block_reduce(mx, true);
    // (x - max) first, then the scale: exact for x = max whatever the magnitude (an fma against a rounded max * log2e
    // leaves a residue of |max| * 2^-24 in the exponent, harmless for logits, fatal for rows like -1e30)
    float const base = mx == -INFINITY ? 0.f : mx;
    float sum = 0.f;
#pragma unroll
    for (int u = 0; u < SM_VEC; u++) {
      int const c = tid + u * SM_THREADS;
      if (c < nvec) {
#pragma unroll
        for (int e = 0; e < 8; e++) {
          f[u][e] = ffb_ex2((f[u][e] - base) * kLog2e);
          sum += f[u][e];
        }
      }
    }
    sum = block_reduce(sum, false);
    float const inv = 1.f / sum;
#pragma unroll
    for (int u = 0; u < SM_VEC; u++) {
      int const c = tid + u * SM_THREADS;
      if (c < nvec) {
#pragma unroll
        for (int e = 0; e < 8; e++) {
          f[u][e] *= inv;
        }
        *reinterpret_cast<uint4 *>(q + (size_t)c * 8) = pack8<T>(f[u]);
      }
    }
    return;
  }
  // long or unaligned rows: three passes (the second and third hit L2)
  float mx = -INFINITY;
  for (int i = tid; i < V; i += SM_THREADS) {
    mx = fmaxf(mx, DT<T>::to_f(p[i]));
  }
  mx = block_reduce(mx, true);
  float const base = mx == -INFINITY ? 0.f : mx;
  float sum = 0.f;
  for (int i = tid; i < V; i += SM_THREADS) {
    sum += ffb_ex2((DT<T>::to_f(p[i]) - base) * kLog2e);
  }
  sum = block_reduce(sum, false);
  float const inv = 1.f / sum;
  for (int i = tid; i < V; i += SM_THREADS) {
    q[i] = DT<T>::from_f(ffb_ex2((DT<T>::to_f(p[i]) - base) * kLog2e) * inv);
  }
}

} // namespace
} // namespace ffb

extern "C" {

int ffb_residual_rms_norm(void const *input1_dev, void const *input2_dev, void const *weight_dev,
                          void *residual_output_dev, void *output_dev, int num_tokens, int dim, float eps, int dtype,
                          ffb_stream_t stream) {
  using namespace ffb;
  if (int rc = ffb_check_device()) {
    return rc;
  }
  FFB_REQUIRE(input1_dev && weight_dev && output_dev, "null argument");
  FFB_REQUIRE(input2_dev == nullptr || residual_output_dev != nullptr, "residual_output_dev is required with input2_dev");
  FFB_REQUIRE(dtype == FFB_DT_HALF || dtype == FFB_DT_BF16, "dtype %d not supported", dtype);
  FFB_REQUIRE(num_tokens >= 0 && dim > 0 && dim % 32 == 0 && dim <= RN_MAXSEG * 1024,
              "rms_norm needs dim %% 32 == 0 and dim <= %d (got %d)", RN_MAXSEG * 1024, dim);
  if (num_tokens == 0) {
    return FFB_OK;
  }
  if (dtype == FFB_DT_HALF) {
    return launch_rms_norm<__half>(input1_dev, input2_dev, weight_dev, residual_output_dev, output_dev, num_tokens,
                                   dim, eps, (cudaStream_t)stream);
  }
  return launch_rms_norm<__nv_bfloat16>(input1_dev, input2_dev, weight_dev, residual_output_dev, output_dev,
                                        num_tokens, dim, eps, (cudaStream_t)stream);
}

int ffb_argmax(void const *input_dev, int *indices_dev, float *prob_dev, int *parent_dev, int num_tokens, int length,
               int dtype, ffb_stream_t stream) {
  using namespace ffb;
  if (int rc = ffb_check_device()) {
    return rc;
  }
  FFB_REQUIRE(input_dev && indices_dev, "null argument");
  FFB_REQUIRE(dtype == FFB_DT_HALF || dtype == FFB_DT_BF16, "dtype %d not supported", dtype);
  FFB_REQUIRE(num_tokens >= 0 && length > 0, "bad shape %d x %d", num_tokens, length);
  if (num_tokens == 0) {
    return FFB_OK;
  }
  // few rows: 1024 threads per row (latency); many rows: 256, so that every row's CTA is resident in one wave
  int const threads = (length >= 8192 && num_tokens <= 2 * kNumSMs) ? 1024 : 256;
  if (dtype == FFB_DT_HALF) {
    argmax_kernel<__half><<<num_tokens, threads, 0, (cudaStream_t)stream>>>((__half const *)input_dev, indices_dev,
                                                                            prob_dev, parent_dev, length);
  } else {
    argmax_kernel<__nv_bfloat16><<<num_tokens, threads, 0, (cudaStream_t)stream>>>(
        (__nv_bfloat16 const *)input_dev, indices_dev, prob_dev, parent_dev, length);
  }
  FFB_LAUNCH_CHECK();
  return FFB_OK;
}

int ffb_softmax(void const *input_dev, void *output_dev, int num_tokens, int num_classes, int dtype, ffb_stream_t stream) {
  using namespace ffb;
  if (int rc = ffb_check_device()) {
    return rc;
  }
  FFB_REQUIRE(input_dev && output_dev, "null argument");
  FFB_REQUIRE(dtype == FFB_DT_HALF || dtype == FFB_DT_BF16, "dtype %d not supported", dtype);
  FFB_REQUIRE(num_tokens >= 0 && num_classes > 0, "bad shape %d x %d", num_tokens, num_classes);
  if (num_tokens == 0) {
    return FFB_OK;
  }
  if (dtype == FFB_DT_HALF) {
    softmax_kernel<__half><<<num_tokens, SM_THREADS, 0, (cudaStream_t)stream>>>((__half const *)input_dev, (__half *)output_dev,
                                                                               num_classes);
  } else {
    softmax_kernel<__nv_bfloat16><<<num_tokens, SM_THREADS, 0, (cudaStream_t)stream>>>(
        (__nv_bfloat16 const *)input_dev, (__nv_bfloat16 *)output_dev, num_classes);
  }
  FFB_LAUNCH_CHECK();
  return FFB_OK;
}

int ffb_arg_topk(void const *input_dev, float *values_dev, int *indices_dev, int num_tokens, int length, int k,
                 int dtype, ffb_stream_t stream) {
  using namespace ffb;
  if (int rc = ffb_check_device()) {
    return rc;
  }
  FFB_REQUIRE(input_dev && indices_dev, "null argument");
  FFB_REQUIRE(dtype == FFB_DT_HALF || dtype == FFB_DT_BF16, "dtype %d not supported", dtype);
  FFB_REQUIRE(num_tokens >= 0 && length > 0 && k >= 1 && k <= 16 && k <= length, "bad shape %d x %d, k = %d", num_tokens,
              length, k);
  if (num_tokens == 0) {
    return FFB_OK;
  }
  int const threads = length >= 8192 ? 1024 : 256;
  cudaStream_t const s = (cudaStream_t)stream;
  if (k <= 4) {
    if (dtype == FFB_DT_HALF) {
      arg_topk_small_kernel<__half, 4><<<num_tokens, threads, 0, s>>>((__half const *)input_dev, values_dev, indices_dev, length, k);
    } else {
      arg_topk_small_kernel<__nv_bfloat16, 4><<<num_tokens, threads, 0, s>>>((__nv_bfloat16 const *)input_dev, values_dev,
                                                                            indices_dev, length, k);
    }
  } else if (dtype == FFB_DT_HALF) {
    arg_topk_kernel<__half><<<num_tokens, threads, 0, s>>>((__half const *)input_dev, values_dev, indices_dev, length, k);
  } else {
    arg_topk_kernel<__nv_bfloat16><<<num_tokens, threads, 0, s>>>((__nv_bfloat16 const *)input_dev, values_dev, indices_dev,
                                                                  length, k);
  }
  FFB_LAUNCH_CHECK();
  return FFB_OK;
}

} // extern "C"
