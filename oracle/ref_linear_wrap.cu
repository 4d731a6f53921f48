// ref_linear_wrap.cu — the cuBLAS call of the reference's Kernels::Linear::Internal::forward_kernel
// (/root/reference/src/ops/kernels/linear_kernels.cu:502-528) RESTATED one for one (this is a library call, there is no
// reference kernel text to compile): cublasGemmEx(OP_T, OP_N, out_dim, batch, in_dim, alpha, W [out][in] ld in_dim,
// input ld in_dim, beta, output ld out_dim), all three types and the COMPUTE type equal to DT - for half that is fp16
// accumulation (compute_type = output_type, :509) - with CUBLAS_GEMM_DEFAULT_TENSOR_OP.  Built into
// oracle/_ref/libref_linear.so (needs no reference source, so it is built everywhere).  TEST INFRASTRUCTURE ONLY: the
// GPU tests compare ffb_linear_forward with it, tools/bench_linear.py times it.
#include <cublas_v2.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

static cublasHandle_t g_handle = nullptr;

extern "C" int ref_linear_forward_half(void const *input, void const *weight, void *output, int in_dim, int out_dim,
                                       int batch_size, void *stream) {
  if (g_handle == nullptr && cublasCreate(&g_handle) != CUBLAS_STATUS_SUCCESS) {
    return -1;
  }
  if (cublasSetStream(g_handle, (cudaStream_t)stream) != CUBLAS_STATUS_SUCCESS) {
    return -2;
  }
  half const alpha = __float2half(1.0f), beta = __float2half(0.0f); // DT alpha = 1.0f, beta = 0.0f (:502)
  cublasStatus_t const st =
      cublasGemmEx(g_handle, CUBLAS_OP_T, CUBLAS_OP_N, out_dim, batch_size, in_dim, &alpha, weight, CUDA_R_16F, in_dim, input,
                   CUDA_R_16F, in_dim, &beta, output, CUDA_R_16F, out_dim, CUBLAS_COMPUTE_16F, CUBLAS_GEMM_DEFAULT_TENSOR_OP);
  return st == CUBLAS_STATUS_SUCCESS ? 0 : (int)st;
}
