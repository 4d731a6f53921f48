// linear_tc.cu — tcgen05 / TMA projection GEMM for batched prompt / tree tokens (M >= 65).
// Replaces the cublasGemmEx call of Kernels::Linear::forward_kernel
// (src/ops/kernels/linear_kernels.cu:507-528): out[M][N] = A[M][K] * W[N][K]^T, both operands K-major.
// One 128 x BN output tile per CTA: a TMA producer thread streams 128B-swizzled A/W tiles through a
// STAGES-deep shared-memory ring (mbarrier full/empty pairs), one elected thread issues
// tcgen05.mma.cta_group::1.kind::f16 (UMMA 128 x BN x 16, fp32 accumulator in TMEM), tcgen05.commit
// releases ring slots and finally signals the four epilogue warps, which read the accumulator with
// tcgen05.ld (32 lanes x 32 columns per instruction), add the bias, convert and store.
// The reference accumulates in DT (cuBLAS compute type = output type); fp32 here.
// int8/int4 weights at large M go to linear_tcq.cu (dequantised in shared memory ahead of the MMA); only the cases
// it does not take (bf16 group meta, N not a multiple of 128) are dequantised once per call into a private DT scratch
// and fed to this kernel (what the reference does, minus its host-to-device copy); the M <= 64 path
// (linear_stream.cu) dequantises in registers.
#include <cuda.h>

#include <new>
#include <type_traits>

#include "tc_common.cuh"

extern "C" int ffb_dequant_int4(void const *, void *, int, int, int, ffb_stream_t);
extern "C" int ffb_dequant_int8(void const *, void *, int, int, int, ffb_stream_t);

namespace ffb {
using namespace tc;
namespace {

constexpr int TC_BM = 128, TC_BK = 64, TC_THREADS = 256;

template <typename T, int BN, int STAGES>
__global__ void __launch_bounds__(TC_THREADS, 1)
linear_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                 T *__restrict__ out, T const *__restrict__ bias, int M, int N, int K) {
  constexpr uint32_t A_BYTES = TC_BM * TC_BK * 2, B_BYTES = BN * TC_BK * 2;
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[STAGES], empty_bar[STAGES], tmem_full_bar;
  __shared__ uint32_t tmem_base_smem;
  // 128B-swizzled tiles need 1024-byte alignment
  uint8_t *smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t *sA = smem, *sB = smem + STAGES * A_BYTES;

  int const warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int const n_blk = blockIdx.x, m_blk = blockIdx.y;
  int const num_k = K / TC_BK;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmA) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmB) : "memory");
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < STAGES; s++) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(&tmem_full_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 2) { // TMEM: BN fp32 accumulator columns (power of two >= 32)
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_smem)), "r"((uint32_t)BN));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t const tmem_base = tmem_base_smem;

  if (warp == 0 && lane == 0) {
    // ===== TMA producer =====
    for (int kb = 0; kb < num_k; kb++) {
      int const s = kb % STAGES;
      uint32_t const ph = (kb / STAGES) & 1;
      mbar_wait(&empty_bar[s], ph ^ 1);
      mbar_expect_tx(&full_bar[s], A_BYTES + B_BYTES);
      tma_load_2d(sA + s * A_BYTES, &tmA, kb * TC_BK, m_blk * TC_BM, &full_bar[s]);
      tma_load_2d(sB + s * B_BYTES, &tmB, kb * TC_BK, n_blk * BN, &full_bar[s]);
    }
  } else if (warp == 1 && lane == 0) {
    // ===== MMA issuer =====
    // cute::UMMA::InstrDescriptor: c_format F32 (1) @4, a/b format @7/@10 (F16 0, BF16 1), K-major, N>>3 @17, M>>4 @24
    constexpr uint32_t fmt = sizeof(T) == 2 && std::is_same<T, __nv_bfloat16>::value ? 1u : 0u;
    constexpr uint32_t idesc = (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(TC_BM >> 4) << 24);
    for (int kb = 0; kb < num_k; kb++) {
      int const s = kb % STAGES;
      uint32_t const ph = (kb / STAGES) & 1;
      mbar_wait(&full_bar[s], ph);
      tc_fence_after();
      uint64_t const da = umma_desc_sw128(smem_u32(sA + s * A_BYTES));
      uint64_t const db = umma_desc_sw128(smem_u32(sB + s * B_BYTES));
#pragma unroll
      for (int k = 0; k < TC_BK / 16; k++) {
        // advance 16 elements (32 B) along K inside the swizzle atom: start-address field += 2
        tc_mma_f16(tmem_base, da + (uint64_t)(k * 2), db + (uint64_t)(k * 2), idesc, (kb | k) != 0);
      }
      tc_commit(&empty_bar[s]); // slot reusable once these MMAs have read it
    }
    tc_commit(&tmem_full_bar);
  } else if (warp >= 4) {
    // ===== epilogue: TMEM -> registers -> global =====
    mbar_wait(&tmem_full_bar, 0);
    tc_fence_after();
    int const q = warp & 3; // this warp may touch TMEM lanes [32q, 32q+32)
    int const row = m_blk * TC_BM + q * 32 + lane;
#pragma unroll 1
    for (int c0 = 0; c0 < BN; c0 += 32) {
      uint32_t r[32];
      tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)c0, r);
      if (row < M) {
        int const col = n_blk * BN + c0;
        T *dst = out + (size_t)row * N + col;
#pragma unroll
        for (int v = 0; v < 4; v++) {
          float f[8];
#pragma unroll
          for (int e = 0; e < 8; e++) {
            f[e] = __uint_as_float(r[v * 8 + e]);
            if (bias) {
              f[e] += DT<T>::to_f(bias[col + v * 8 + e]);
            }
          }
          *reinterpret_cast<uint4 *>(dst + v * 8) = pack8<T>(f);
        }
      }
    }
    tc_fence_before();
  }
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)BN));
  }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, cuuint64_t const *,
                                  cuuint64_t const *, cuuint32_t const *, cuuint32_t const *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn() {
  // function-local static: initialised once, thread-safe (C++11)
  static EncodeTiledFn const fn = []() -> EncodeTiledFn {
    void *p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess) {
      return (EncodeTiledFn)p;
    }
    return nullptr;
  }();
  return fn;
}

} // namespace

// [rows][K] row-major 16-bit matrix, box = 64 (K) x box_rows, 128B swizzle, OOB rows read as zero
int make_tensor_map_2d(CUtensorMap *tm, int dtype, void const *ptr, long rows, int K, int box_rows) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) {
    set_error("cuTensorMapEncodeTiled is not available from the driver");
    return FFB_ERR_CUDA;
  }
  cuuint64_t gdim[2] = {(cuuint64_t)K, (cuuint64_t)rows};
  cuuint64_t gstride[1] = {(cuuint64_t)K * 2};
  cuuint32_t box[2] = {(cuuint32_t)TC_BK, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(tm, dtype == FFB_DT_HALF ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2,
                  (void *)ptr, gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed with CUresult %d (ptr %p rows %ld K %d)", (int)r, ptr, rows, K);
    return FFB_ERR_CUDA;
  }
  return FFB_OK;
}

// [rows][row_bytes] row-major byte matrix (quantised codes), box = box_bytes x box_rows, no swizzle
int make_tensor_map_u8(CUtensorMap *tm, void const *ptr, long rows, long row_bytes, int box_rows, int box_bytes) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) {
    set_error("cuTensorMapEncodeTiled is not available from the driver");
    return FFB_ERR_CUDA;
  }
  cuuint64_t gdim[2] = {(cuuint64_t)row_bytes, (cuuint64_t)rows};
  cuuint64_t gstride[1] = {(cuuint64_t)row_bytes};
  cuuint32_t box[2] = {(cuuint32_t)box_bytes, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, (void *)ptr, gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled (uint8) failed with CUresult %d (ptr %p rows %ld row bytes %ld)", (int)r, ptr, rows, row_bytes);
    return FFB_ERR_CUDA;
  }
  return FFB_OK;
}

// general form: `rank` dimensions, fastest first; strides in bytes for dimensions 1 .. rank - 1; elem: 1 = uint8,
// 2 = 16-bit (uint16 when !is_half, fp16 otherwise); swizzle128 selects the 128-byte swizzle (else none)
int make_tensor_map_nd(CUtensorMap *tm, int elem, bool is_half, int rank, void const *ptr, unsigned long long const *dims,
                       unsigned long long const *strides, unsigned const *box, bool swizzle128) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) {
    set_error("cuTensorMapEncodeTiled is not available from the driver");
    return FFB_ERR_CUDA;
  }
  cuuint64_t gdim[5], gstride[4];
  cuuint32_t bx[5], estr[5];
  for (int i = 0; i < rank; i++) {
    gdim[i] = dims[i];
    bx[i] = box[i];
    estr[i] = 1;
    if (i > 0) {
      gstride[i - 1] = strides[i - 1];
    }
  }
  CUtensorMapDataType const dt = elem == 1 ? CU_TENSOR_MAP_DATA_TYPE_UINT8
                                           : (is_half ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_UINT16);
  CUresult r = fn(tm, dt, rank, (void *)ptr, gdim, gstride, bx, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  swizzle128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled (rank %d) failed with CUresult %d (ptr %p)", rank, (int)r, ptr);
    return FFB_ERR_CUDA;
  }
  return FFB_OK;
}

int linear_tc2_supported(int dtype, int M, int K, int N);
// linear_tcq.cu: quantised weights dequantised in shared memory ahead of the MMA (half activations)
int linear_tcq_supported(int dtype, int quant, int K, int N);
int launch_linear_tcq(int quant, void const *in, void *out, void const *blob, void const *bias, int M, int K, int N, cudaStream_t s);
// linear_tc4.cu: the same SM-pair tiles in clusters of two pairs sharing A by TMA multicast (FFB_TC4=1)
int linear_tc4_supported(int dtype, int M, int K, int N);
int launch_linear_tc4(int dtype, void const *in, void *out, void const *w, void const *bias, int M, int K, int N, cudaStream_t s);
int launch_linear_tc2(int dtype, void const *in, void *out, void const *w, void const *bias, int M, int K, int N,
                      void *ws, int *cnt, cudaStream_t s);
// linear_sk.cu: the same SM-pair tiles on a stream-K schedule (equal k-block ranges per pair, partial tiles merged
// through the split-K workspace)
int linear_sk_supported(int dtype, int M, int K, int N, void const *ws, int const *cnt);
int launch_linear_sk(int dtype, void const *in, void *out, void const *w, void const *bias, int M, int K, int N,
                     void *ws, int *cnt, cudaStream_t s);
// linear_w512.cu: 256 x 512 tiles per SM pair (two UMMAs per k-step against one A tile), one wave
int linear_tcc_supported(int dtype, int M, int K, int N);
int launch_linear_tcc(int dtype, void const *in, void *out, void const *w, void const *bias, int M, int K, int N, cudaStream_t s);
int linear_w512_supported(int dtype, int M, int K, int N);
int launch_linear_w512(int dtype, void const *in, void *out, void const *w, void const *bias, int M, int K, int N, cudaStream_t s);
size_t linear_tc2_ws_bytes();
size_t linear_tc2_cnt_bytes();

struct LinearTcState {
  int dtype, quant, K, N, max_batch;
  void *scratch = nullptr; // dequantised weights [N][K] DT (quantised layers only)
  void *splitk_ws = nullptr; // fp32 partial tiles of the split-K SM-pair GEMM (linear_tc2.cu)
  int *splitk_cnt = nullptr; // its arrival counters (zero between launches)
  bool attr_set[2] = {false, false};
};

int linear_tc_supported(int dtype, int quant, int K, int N) {
  (void)quant;
  if (ffb::env("FFB_DISABLE_TC")) {
    return 0;
  }
  return (dtype == FFB_DT_HALF || dtype == FFB_DT_BF16) && K % TC_BK == 0 && N % 16 == 0;
}

void linear_tc_destroy(LinearTcState *st);

LinearTcState *linear_tc_create(int dtype, int quant, int K, int N, int max_batch) {
  LinearTcState *st = new (std::nothrow) LinearTcState();
  if (!st) {
    set_error("out of host memory");
    return nullptr;
  }
  st->dtype = dtype; st->quant = quant; st->K = K; st->N = N; st->max_batch = max_batch;
  // quantised layers: linear_tcq.cu dequantises in shared memory; only shapes / types it does not take (bf16, ragged
  // N) fall back to dequantise-into-a-scratch + dense GEMM
  if (quant != FFB_DT_NONE && max_batch > 64 && !linear_tcq_supported(dtype, quant, K, N)) {
    cudaError_t e = cudaMalloc(&st->scratch, (size_t)K * N * 2);
    if (e != cudaSuccess) {
      set_error("cudaMalloc of the dequant scratch failed: %s", cudaGetErrorString(e));
      delete st;
      return nullptr;
    }
  }
  // split-K only pays when the output has fewer 256 x 256 tiles than SM pairs at some batch size <= max_batch
  if (max_batch > 128 && (N + 255) / 256 < kNumSMs / 2 && !ffb::env("FFB_DISABLE_TC2_SPLITK")) {
    if (cudaMalloc(&st->splitk_ws, linear_tc2_ws_bytes()) != cudaSuccess ||
        cudaMalloc((void **)&st->splitk_cnt, linear_tc2_cnt_bytes()) != cudaSuccess ||
        cudaMemset(st->splitk_cnt, 0, linear_tc2_cnt_bytes()) != cudaSuccess) {
      set_error("cudaMalloc of the split-K workspace failed: %s", cudaGetErrorString(cudaGetLastError()));
      linear_tc_destroy(st);
      return nullptr;
    }
  }
  return st;
}

void linear_tc_destroy(LinearTcState *st) {
  if (st) {
    if (st->scratch) {
      cudaFree(st->scratch);
    }
    if (st->splitk_ws) {
      cudaFree(st->splitk_ws);
    }
    if (st->splitk_cnt) {
      cudaFree(st->splitk_cnt);
    }
    delete st;
  }
}

int launch_linear_tc(LinearTcState *st, int dtype, int quant, void const *in, void *out, void const *w,
                     void const *bias, int M, int K, int N, cudaStream_t s) {
  void const *wd = w;
  if (quant != FFB_DT_NONE && linear_tcq_supported(dtype, quant, K, N)) {
    return launch_linear_tcq(quant, in, out, w, bias, M, K, N, s);
  }
  if (quant != FFB_DT_NONE) {
    FFB_REQUIRE(st->scratch, "quantised layer created with max_batch <= 64 has no dequant scratch");
    int rc = quant == FFB_DT_INT4 ? ffb_dequant_int4(w, st->scratch, K, N, dtype, s) : ffb_dequant_int8(w, st->scratch, K, N, dtype, s);
    if (rc != FFB_OK) {
      return rc;
    }
    wd = st->scratch;
  }
  // up to 128 tokens: one 128-feature tile per cluster, the reduction dimension split over its CTAs (linear_tcc.cu)
  if (M <= 128 && linear_tcc_supported(dtype, M, K, N)) {
    return launch_linear_tcc(dtype, in, out, wd, bias, M, K, N, s);
  }
  // M > 128 (or an N the 128-wide 1-SM tiles do not divide): SM-pair tiles (linear_tc2.cu)
  if (M > 128 && linear_tc4_supported(dtype, M, K, N)) {
    return launch_linear_tc4(dtype, in, out, wd, bias, M, K, N, s);
  }
  if (linear_w512_supported(dtype, M, K, N)) {
    return launch_linear_w512(dtype, in, out, wd, bias, M, K, N, s);
  }
  if (linear_sk_supported(dtype, M, K, N, st->splitk_ws, st->splitk_cnt)) {
    return launch_linear_sk(dtype, in, out, wd, bias, M, K, N, st->splitk_ws, st->splitk_cnt, s);
  }
  // more tokens on a narrow layer: the same cluster split per (128-token block, tile) where the pair kernel has few tiles
  if (M > 128 && !ffb::env("FFB_TC2_BN") && !ffb::env("FFB_TC2_SPLIT") && linear_tcc_supported(dtype, M, K, N)) {
    return launch_linear_tcc(dtype, in, out, wd, bias, M, K, N, s);
  }
  if ((M > 128 || N % 128 != 0) && linear_tc2_supported(dtype, M, K, N)) {
    return launch_linear_tc2(dtype, in, out, wd, bias, M, K, N, st->splitk_ws, st->splitk_cnt, s);
  }
  constexpr int BN = 128, STAGES = 6;
  CUtensorMap tmA, tmB;
  int rc = make_tensor_map_2d(&tmA, dtype, in, M, K, TC_BM);
  if (rc != FFB_OK) {
    return rc;
  }
  rc = make_tensor_map_2d(&tmB, dtype, wd, N, K, BN);
  if (rc != FFB_OK) {
    return rc;
  }
  size_t const smem = (size_t)STAGES * (TC_BM + BN) * TC_BK * 2 + 1024;
  dim3 grid(N / BN, (M + TC_BM - 1) / TC_BM);
  if (dtype == FFB_DT_HALF) {
    if (!st->attr_set[0]) {
      FFB_CUDA_TRY(cudaFuncSetAttribute(linear_tc_kernel<__half, BN, STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      st->attr_set[0] = true;
    }
    linear_tc_kernel<__half, BN, STAGES><<<grid, TC_THREADS, smem, s>>>(tmA, tmB, (__half *)out, (__half const *)bias, M, N, K);
  } else {
    if (!st->attr_set[1]) {
      FFB_CUDA_TRY(cudaFuncSetAttribute(linear_tc_kernel<__nv_bfloat16, BN, STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      st->attr_set[1] = true;
    }
    linear_tc_kernel<__nv_bfloat16, BN, STAGES><<<grid, TC_THREADS, smem, s>>>(tmA, tmB, (__nv_bfloat16 *)out, (__nv_bfloat16 const *)bias, M, N, K);
  }
  FFB_LAUNCH_CHECK();
  return FFB_OK;
}

} // namespace ffb
