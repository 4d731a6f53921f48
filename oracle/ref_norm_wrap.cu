// ref_norm_wrap.cu — compiles the REFERENCE's own (Residual)RMSNorm forward kernels
// (/root/reference/src/ops/kernels/residual_rms_norm_kernels.cu:58-127: WARP_SHFL_DOWN, WarpReduceSum, BlockReduceSum,
// ResidualRMSNormFusedForwardKernel; /root/reference/src/ops/kernels/rms_norm_kernels.cu:58-124: the same helpers and
// RMSNormFusedForwardKernel) into oracle/_ref/libref_norm.so, plus the cub call that ArgMax::forward_kernel makes
// (src/ops/argmax.cu:81-102) restated with the same library.  TEST INFRASTRUCTURE ONLY.  The Makefile streams each
// file's text from the first helper to the forward kernel's closing brace into a temporary include (deleted after the
// compile); nothing is copied into this repository.  Supplied here: the `template <typename T>` line that precedes
// the first helper, C10_WARP_SIZE (:26 of both files), and the launches, which restate forward_kernel of either file
// (one block per row, min(CUDA_NUM_THREADS = 1024, in_dim) threads).
#include <algorithm>
#include <cassert>
#include <cstdint>
#include <cub/cub.cuh>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#define C10_WARP_SIZE 32

namespace RefResidualRMSNorm {
template <typename T>
#include REF_RESIDUAL_RMS_INC
} // namespace RefResidualRMSNorm

namespace RefRMSNorm {
template <typename T>
#include REF_RMS_INC
} // namespace RefRMSNorm

// ArgTopK: the whole kernel section of src/ops/arg_topk.cu (from `enum class HeapType` :23 to the end of
// arg_topk_forward_kernel :371: heap, heapArgTopK, mergeShards, the kernel), streamed by the Makefile.
namespace RefArgTopK {
#include REF_ARGTOPK_INC
} // namespace RefArgTopK

namespace {
// copy_result, src/ops/argmax.cu:35-47
__global__ void ref_copy_result(cub::KeyValuePair<int, half> *d_out, int *indices, float *prob, int batch_size) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < batch_size; i += blockDim.x * gridDim.x) {
    indices[i] = d_out[i].key;
    if (prob) {
      prob[i] = static_cast<float>(d_out[i].value);
    }
  }
}
} // namespace

extern "C" {

// rms_scratch: [rows] half, norm_scratch: [rows][dim] half (m->rms_ptr, m->norm_ptr of the reference Meta)
int ref_residual_rms_norm_half(void const *x1, void const *x2, void const *w, void *residual_out, void *out,
                               void *rms_scratch, void *norm_scratch, int rows, int dim, float eps, void *stream) {
  RefResidualRMSNorm::ResidualRMSNormFusedForwardKernel<half><<<rows, std::min(1024, dim), 0, (cudaStream_t)stream>>>(
      dim, eps, (half const *)x1, (half const *)x2, (half *)residual_out, (half *)rms_scratch, (half *)norm_scratch,
      (half const *)w, (half *)out);
  return (int)cudaGetLastError();
}
int ref_rms_norm_half(void const *x, void const *w, void *out, void *rms_scratch, void *norm_scratch, int rows, int dim,
                      float eps, void *stream) {
  RefRMSNorm::RMSNormFusedForwardKernel<half><<<rows, std::min(1024, dim), 0, (cudaStream_t)stream>>>(
      dim, eps, (half const *)x, (half *)rms_scratch, (half *)norm_scratch, (half const *)w, (half *)out);
  return (int)cudaGetLastError();
}
// ArgTopK::forward_kernel's speculative-decoding launch (src/ops/arg_topk.cu:425-436): one block per row, k threads
// (num_shards = k), values written for rank < k - 1.
int ref_arg_topk_half(void const *in, float *values, int *indices, int rows, int length, int k, void *stream) {
  size_t const shared_memory_size = (size_t)(k + 1) * k * sizeof(RefArgTopK::Entry<half>);
  RefArgTopK::arg_topk_forward_kernel<half><<<rows, k, 0, (cudaStream_t)stream>>>((half const *)in, shared_memory_size, length, k,
                                                                               true, values, indices, values != nullptr);
  return (int)cudaGetLastError();
}
// offsets_dev: rows + 1 ints {0, length, 2 length, ...}; temp_dev: temp_bytes of scratch (0 / NULL: returns the size
// cub asks for through *temp_bytes_out); kv_dev: rows KeyValuePair<int, half>.
int ref_argmax_half(void const *in, int *indices, float *prob, int rows, int const *offsets_dev, void *temp_dev,
                    size_t temp_bytes, size_t *temp_bytes_out, void *kv_dev, void *stream) {
  size_t bytes = temp_bytes;
  cudaError_t e = cub::DeviceSegmentedReduce::ArgMax(temp_dev, bytes, (half const *)in,
                                                     (cub::KeyValuePair<int, half> *)kv_dev, rows, offsets_dev,
                                                     offsets_dev + 1, (cudaStream_t)stream);
  if (temp_bytes_out) {
    *temp_bytes_out = bytes;
  }
  if (e != cudaSuccess || temp_dev == nullptr) {
    return (int)e;
  }
  ref_copy_result<<<(rows + 1023) / 1024, std::min(1024, rows), 0, (cudaStream_t)stream>>>(
      (cub::KeyValuePair<int, half> *)kv_dev, indices, prob, rows);
  return (int)cudaGetLastError();
}

} // extern "C"
