// see orc_common.h for the header comment (test infrastructure only)
//
// (Residual)RMSNorm forward and ArgMax — SURVEY §8f N1 / N2.  Parity status: the reference has no golden vectors for
// these kernels either; the norm is pinned against HuggingFace's LlamaRMSNorm in tests/test_oracle_norm.py (the
// alignment path of tests/inference/inference_alignment_test.py), the argmax against numpy.
#include "orc_common.h"

namespace {

inline float bits_to_f(uint16_t b, int dt) { return dt == 1 ? half_bits_to_float(b) : bf16_bits_to_float(b); }
inline uint16_t f_to_bits(float f, int dt) { return dt == 1 ? float_to_half_bits(f) : float_to_bf16_bits(f); }

// WarpReduceSum, residual_rms_norm_kernels.cu:71-78: val += shfl_down(val, offset) for offset = 16 .. 1.  Lanes whose
// partner is out of range get their own value back; only lane 0's result is used.
inline float warp_reduce_lane0(float v[32]) {
  for (int offset = 16; offset > 0; offset >>= 1) {
    float nv[32];
    for (int l = 0; l < 32; l++) {
      nv[l] = v[l] + (l + offset < 32 ? v[l + offset] : v[l]);
    }
    std::memcpy(v, nv, sizeof(nv));
  }
  return v[0];
}

} // namespace

extern "C" {

// ResidualRMSNormFusedForwardKernel, src/ops/kernels/residual_rms_norm_kernels.cu:97-127 (x2 != NULL) and
// RMSNormFusedForwardKernel, src/ops/kernels/rms_norm_kernels.cu:97-124 (x2 == NULL).  Launch: one block per row,
// min(CUDA_NUM_THREADS = 1024, in_dim) threads (:139, cuda_helper.h:69).  16-bit patterns on the interface.
void orc_residual_rms_norm(uint16_t const *x1, uint16_t const *x2, uint16_t const *weight, uint16_t *residual_out,
                           uint16_t *out, int rows, int N, float eps, int dt) {
  int const B = std::min(1024, N);
  std::vector<float> xo(N), part(1024);
  for (int i = 0; i < rows; i++) {
    size_t const base = (size_t)i * N;
    std::fill(part.begin(), part.end(), 0.f);
    for (int t = 0; t < B; t++) { // :106-112, thread t
      float sum = 0.0f;
      for (int j = t; j < N; j += B) {
        float v = bits_to_f(x1[base + j], dt);
        if (x2) { // X_out[index] = X1[index] + X2[index] in DT: exact sum rounded once
          uint16_t const b = dt == 1 ? double_to_half_bits((double)v + (double)bits_to_f(x2[base + j], dt))
                                     : float_to_bf16_bits(v + bits_to_f(x2[base + j], dt));
          residual_out[base + j] = b;
          v = bits_to_f(b, dt);
        }
        xo[j] = v;
        sum += v * v; // exact product of two 16-bit values, one fp32 rounding in the add
      }
      part[t] = sum;
    }
    // BlockReduceSum :80-95
    float shared[32];
    int const nwarps = (B + 31) / 32;
    for (int w = 0; w < 32; w++) {
      shared[w] = 0.f;
    }
    for (int w = 0; w < nwarps; w++) {
      float v[32];
      for (int l = 0; l < 32; l++) {
        v[l] = w * 32 + l < B ? part[w * 32 + l] : 0.f;
      }
      shared[w] = warp_reduce_lane0(v);
    }
    float v[32];
    for (int l = 0; l < 32; l++) {
      v[l] = l < B / 32 ? shared[l] : 0.f; // threadIdx.x < blockDim.x / 32
    }
    float const sum = warp_reduce_lane0(v);
    // rms[i] = static_cast<T>(rsqrt(sum / N + eps)) :114 (rsqrt restated in double, rounded to float)
    float const arg = sum / (float)N + eps;
    float const r32 = (float)(1.0 / std::sqrt((double)arg));
    float const rms = bits_to_f(f_to_bits(r32, dt), dt);
    for (int j = 0; j < N; j++) { // :119-126
      float const y = bits_to_f(f_to_bits(xo[j] * rms, dt), dt);
      out[base + j] = f_to_bits(y * bits_to_f(weight[j], dt), dt);
    }
  }
}

// ArgMax::forward_kernel, src/ops/argmax.cu:64-102: cub::DeviceSegmentedReduce::ArgMax (ties -> smaller key) followed
// by copy_result (:35-47); parent is zeroed in the beam-search case (:77-80).
void orc_argmax(float const *in, int *indices, float *prob, int *parent, int rows, int length) {
  for (int i = 0; i < rows; i++) {
    float const *p = in + (size_t)i * length;
    int best = 0;
    for (int j = 1; j < length; j++) {
      if (p[j] > p[best]) {
        best = j;
      }
    }
    indices[i] = best;
    if (prob) {
      prob[i] = p[best];
    }
    if (parent) {
      parent[i] = 0;
    }
  }
}

} // extern "C"

extern "C" {

// ArgTopK::forward_kernel, src/ops/arg_topk.cu:260-371 (heapArgTopK per shard + mergeShards): the min-heap that keeps
// the k best evicts the HIGHER index among equal values (PreferIndices::kHigher, :272-276, :283-291) and the max-heap
// that emits them prefers the LOWER index (:299-303), i.e. the output is the first k entries of the row sorted by
// (value descending, index ascending).  Values are emitted as float for rank < k - 1 only (:318-321, :330-331 is
// commented out); values[rank k-1] is left as it was.
void orc_arg_topk(float const *in, float *values, int *indices, int rows, int length, int k) {
  std::vector<int> order(length);
  for (int i = 0; i < rows; i++) {
    float const *p = in + (size_t)i * length;
    for (int j = 0; j < length; j++) {
      order[j] = j;
    }
    std::partial_sort(order.begin(), order.begin() + k, order.end(),
                      [&](int a, int b) { return p[a] > p[b] || (p[a] == p[b] && a < b); });
    for (int r = 0; r < k; r++) {
      indices[(size_t)i * k + r] = order[r];
      if (values && r < k - 1) {
        values[(size_t)i * k + r] = p[order[r]];
      }
    }
  }
}

} // extern "C"
