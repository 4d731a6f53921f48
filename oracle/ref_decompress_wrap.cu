// ref_decompress_wrap.cu — compiles the REFERENCE's own decompression kernels
// (/root/reference/src/ops/kernels/decompress_kernels.cu, read where it lies; nothing is copied into this repo) into
// oracle/_ref/libref_decompress.so, so that the GPU tests can compare ffb_dequant_int4 / _int8 bit for bit with the
// reference kernel running on the same B200, and bench tools can time it beside ours.  TEST INFRASTRUCTURE ONLY.
//
// The reference file includes three headers that drag in Legion / cuDNN / NCCL for unrelated declarations; empty
// stand-ins of the same names (oracle/ref_stubs/) come first on the include path and the three names the kernels
// actually use are supplied here instead:
//   CUDA_KERNEL_LOOP                  include/flexflow/utils/cuda_helper.h:64-66 (grid-stride loop over Legion::coord_t)
//   INT4_NUM_OF_ELEMENTS_PER_GROUP    include/flexflow/ffconst_utils.h:13
//   Legion::coord_t / Legion::Memory  named by the file's using-declarations
// The launch configuration restates src/ops/kernels/linear_kernels.cu:469-490 (GET_BLOCKS, CUDA_NUM_THREADS = 1024,
// BLOCK_SIZE_LIMIT = 32768: cuda_helper.h:69-76).
#include <cuda_fp16.h>
#include <cuda_runtime.h>

namespace Legion {
typedef long long coord_t;
struct Memory {};
} // namespace Legion
#define INT4_NUM_OF_ELEMENTS_PER_GROUP 32
#define CUDA_KERNEL_LOOP(i, n) \
  for (Legion::coord_t i = blockIdx.x * blockDim.x + threadIdx.x; i < (n); i += blockDim.x * gridDim.x)

#include REF_DECOMPRESS_CU

static int ref_blocks(long long n) {
  long long b = (n + 1023) / 1024;
  return (int)(b > 32768 ? 32768 : b);
}

extern "C" {

int ref_decompress_int4_half(void const *blob, void *out, int in_dim, int out_dim, void *stream) {
  long long const par = (long long)in_dim * out_dim / 2;
  FlexFlow::Kernels::decompress_int4_general_weights<half>
      <<<ref_blocks(par), par < 1024 ? (int)par : 1024, 0, (cudaStream_t)stream>>>((char const *)blob, (half *)out, in_dim, in_dim * out_dim);
  return (int)cudaGetLastError();
}
int ref_decompress_int8_half(void const *blob, void *out, int in_dim, int out_dim, void *stream) {
  long long const par = (long long)in_dim * out_dim;
  FlexFlow::Kernels::decompress_int8_general_weights<half>
      <<<ref_blocks(par), par < 1024 ? (int)par : 1024, 0, (cudaStream_t)stream>>>((char const *)blob, (half *)out, in_dim, in_dim * out_dim);
  return (int)cudaGetLastError();
}

} // extern "C"
