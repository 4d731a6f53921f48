// capi_linear.cu — C ABI for the QKV / o_proj dense layers (include/ffb200.h).
// Mirrors Kernels::Linear::forward_kernel_wrapper / inference_kernel_wrapper
// (src/ops/kernels/linear_kernels.cu:133-140,187-195): out = in * W^T (+ bias).
#include <cstdlib>
#include <cstring>
#include <new>

#include "common.cuh"
#include "peer.h"

namespace ffb {
int launch_linear_stream(int dtype, int quant, void const *in, void *out, void const *w, void const *bias, int M,
                         int K, int N, cudaStream_t s);
// persistent bulk-copy weight stream for decode batches (linear_gemv.cu)
int linear_gemv_supported(int quant, int M, int K, int N);
int launch_linear_gemv(int dtype, void const *in, void *out, void const *w, void const *bias, int M, int K, int N,
                       cudaStream_t s);
// fused o_proj + NVLink AllReduce for decode batches (linear_gemv.cu)
int linear_qgemv_supported(int dtype, int quant, int M, int K, int N);
int launch_linear_qgemv(int quant, void const *in, void *out, void const *w, void const *bias, int M, int K, int N,
                        cudaStream_t s);
int linear_gemv_allreduce_supported(PeerTable const *pt, int quant, int M, int K, int N);
int launch_linear_gemv_allreduce(PeerTable const *pt, int dtype, void const *in, void *out, void const *w, void const *bias,
                                 int M, int K, int N, cudaStream_t s);
// tcgen05/TMA GEMM for batched prompt/tree tokens (linear_tc.cu)
int linear_tc_supported(int dtype, int quant, int K, int N);
int launch_linear_tc(struct LinearTcState *st, int dtype, int quant, void const *in, void *out, void const *w,
                     void const *bias, int M, int K, int N, cudaStream_t s);
// quantised weights, <= 64 tokens: TMA stream + in-shared-memory dequantisation + tcgen05 (linear_qtc.cu)
int linear_qtc_supported(int dtype, int quant, int M, int K, int N);
size_t linear_qtc_ws_bytes();
int launch_linear_qtc(int quant, void const *in, void *out, void const *blob, void const *bias, int M, int K, int N,
                      void *ws, cudaStream_t s);
struct LinearTcState *linear_tc_create(int dtype, int quant, int K, int N, int max_batch);
void linear_tc_destroy(struct LinearTcState *st);
} // namespace ffb

struct ffb_linear_s {
  ffb_linear_desc d;
  ffb::LinearTcState *tc = nullptr;
  void *qtc_ws = nullptr; // quantised layers: partial slots + flags of linear_qtc.cu
};

using namespace ffb;

extern "C" {

int ffb_linear_create(ffb_linear_desc const *desc, ffb_linear_t *out) {
  FFB_REQUIRE(desc && out, "null argument");
  ffb_linear_desc const &d = *desc;
  FFB_REQUIRE(d.dtype == FFB_DT_HALF || d.dtype == FFB_DT_BF16, "dtype %d not supported (half/bf16 only)", d.dtype);
  FFB_REQUIRE(d.quant_type == FFB_DT_NONE || d.quant_type == FFB_DT_INT8 || d.quant_type == FFB_DT_INT4,
              "quant_type %d not supported", d.quant_type);
  FFB_REQUIRE(d.in_dim > 0 && d.out_dim > 0 && d.in_dim % 128 == 0 && d.out_dim % 32 == 0,
              "linear needs in_dim %% 128 == 0 and out_dim %% 32 == 0 (got %d -> %d)", d.in_dim, d.out_dim);
  FFB_REQUIRE(d.max_batch > 0 && d.max_batch <= FFB_MAX_NUM_TOKENS, "max_batch %d out of range", d.max_batch);
  if (d.activation != 0) {
    set_error("activation %d: only AC_MODE_NONE is on the attention hot path", d.activation);
    return FFB_ERR_UNSUPPORTED;
  }
  int rc = ffb_check_device();
  if (rc != FFB_OK) {
    return rc;
  }
  ffb_linear_s *h = new (std::nothrow) ffb_linear_s();
  FFB_REQUIRE(h, "out of host memory");
  h->d = d;
  if (linear_tc_supported(d.dtype, d.quant_type, d.in_dim, d.out_dim)) {
    h->tc = linear_tc_create(d.dtype, d.quant_type, d.in_dim, d.out_dim, d.max_batch);
    if (!h->tc) {
      delete h;
      return FFB_ERR_CUDA;
    }
  }
  if (d.quant_type != FFB_DT_NONE && linear_qtc_supported(d.dtype, d.quant_type, 1, d.in_dim, d.out_dim)) {
    if (cudaMalloc(&h->qtc_ws, linear_qtc_ws_bytes()) != cudaSuccess || cudaMemset(h->qtc_ws, 0, linear_qtc_ws_bytes()) != cudaSuccess) {
      set_error("cudaMalloc of the quantised-projection workspace failed: %s", cudaGetErrorString(cudaGetLastError()));
      ffb_linear_destroy(h);
      return FFB_ERR_CUDA;
    }
  }
  *out = h;
  return FFB_OK;
}

int ffb_linear_destroy(ffb_linear_t h) {
  if (h) {
    if (h->qtc_ws) {
      cudaFree(h->qtc_ws);
    }
    if (h->tc) {
      linear_tc_destroy(h->tc);
    }
    delete h;
  }
  return FFB_OK;
}

int ffb_linear_forward(ffb_linear_t h, void const *in, void *out, void const *w, void const *bias, int batch,
                       ffb_stream_t stream) {
  FFB_REQUIRE(h && in && out && w, "null argument");
  FFB_REQUIRE(batch >= 0 && batch <= h->d.max_batch, "batch %d exceeds max_batch %d", batch, h->d.max_batch);
  if (batch == 0) {
    return FFB_OK;
  }
  ffb_linear_desc const &d = h->d;
  cudaStream_t s = (cudaStream_t)stream;
  // FFB_LINEAR=stream|tc forces a path (tests / profiling)
  char const *force = ffb::env("FFB_LINEAR");
  // Dense weights: up to 16 tokens the persistent bulk-copy GEMV, above that the tcgen05 kernels: at 17..128 tokens one
  // 128..256-feature tile per thread-block cluster streams the weights once, the reduction dimension split over the
  // cluster's CTAs when the layer has few tiles (linear_tcc.cu; 7B o_proj at 32 tokens: 10.7 us against 13.4 us in the
  // mma.sync stream kernel).  Quantised weights: in-register dequantisation in the stream kernel up to 64 tokens,
  // dequantise-then-GEMM above.  Env FFB_TC_MIN overrides.
  static int const tc_min_env = ffb::env("FFB_TC_MIN") ? atoi(ffb::env("FFB_TC_MIN")) : 0;
  int const tc_min = tc_min_env > 0 ? tc_min_env : (d.quant_type != FFB_DT_NONE ? 65 : 17);
  bool use_tc = h->tc && batch >= tc_min;
  if (force && strcmp(force, "stream") == 0) {
    use_tc = false;
  } else if (force && strcmp(force, "tc") == 0) {
    FFB_REQUIRE(h->tc, "FFB_LINEAR=tc but the tcgen05 path does not support this shape");
    use_tc = true;
  }
  if (use_tc) {
    return launch_linear_tc(h->tc, d.dtype, d.quant_type, in, out, w, bias, batch, d.in_dim, d.out_dim, s);
  }
  bool const force_stream = force && strcmp(force, "stream") == 0;
  if (force && strcmp(force, "qtc") == 0) { // (tests: fail instead of silently taking another kernel)
    FFB_REQUIRE(h->qtc_ws && linear_qtc_supported(d.dtype, d.quant_type, batch, d.in_dim, d.out_dim),
                "FFB_LINEAR=qtc but linear_qtc does not support this shape / batch");
  }
  if (!force_stream && !(force && strcmp(force, "qgemv") == 0) && h->qtc_ws &&
      linear_qtc_supported(d.dtype, d.quant_type, batch, d.in_dim, d.out_dim)) {
    return launch_linear_qtc(d.quant_type, in, out, w, bias, batch, d.in_dim, d.out_dim, h->qtc_ws, s);
  }
  if (!force_stream && linear_gemv_supported(d.quant_type, batch, d.in_dim, d.out_dim)) {
    return launch_linear_gemv(d.dtype, in, out, w, bias, batch, d.in_dim, d.out_dim, s);
  }
  // quantised blobs, decode batches: the persistent bulk-copy kernel with per-(group, column) reciprocals in shared
  // memory (linear_qgemv.cu) when the layer has at most one 32-feature group per SM (o_proj: 11.4 / 11.8 us for int8 /
  // int4 against 11.5 / 12.4 us in the stream kernel); wider layers (qkv_proj, 384 groups = 3 tiles on the busiest
  // CTA) are level with the stream kernel (24.3 / 26.9 against 23.2 / 26.4 us) and stay there.  FFB_LINEAR=qgemv /
  // stream force either for A/B.
  bool const force_qgemv = force && strcmp(force, "qgemv") == 0;
  if (!force_stream && (force_qgemv || d.out_dim / 32 <= kNumSMs) &&
      linear_qgemv_supported(d.dtype, d.quant_type, batch, d.in_dim, d.out_dim)) {
    return launch_linear_qgemv(d.quant_type, in, out, w, bias, batch, d.in_dim, d.out_dim, s);
  }
  size_t const elt = 2;
  for (int m0 = 0; m0 < batch; m0 += 64) {
    int const m = batch - m0 < 64 ? batch - m0 : 64;
    int rc = launch_linear_stream(d.dtype, d.quant_type, (char const *)in + (size_t)m0 * d.in_dim * elt,
                                  (char *)out + (size_t)m0 * d.out_dim * elt, w, bias, m, d.in_dim, d.out_dim, s);
    if (rc != FFB_OK) {
      return rc;
    }
  }
  return FFB_OK;
}

/* Row-parallel dense followed by the AllReduce over the TP shards (Linear -> AllReduce of
 * src/runtime/model.cc:3421-3445, run back to back inside FusedOp::inference_task, src/ops/fused.cu:448-495).
 * Decode batches (<= 16 tokens, dense weights, a communicator with an NVLink peer mapping) run as ONE kernel that
 * exchanges the partial sums over peer memory; everything else is ffb_linear_forward + ffb_allreduce (NCCL). */
int ffb_linear_forward_allreduce(ffb_linear_t h, ffb_comm_t comm, void const *in, void *out, void const *w,
                                 void const *bias, int batch, ffb_stream_t stream) {
  FFB_REQUIRE(h && comm && in && out && w, "null argument");
  FFB_REQUIRE(batch >= 0 && batch <= h->d.max_batch, "batch %d exceeds max_batch %d", batch, h->d.max_batch);
  if (batch == 0) {
    return FFB_OK;
  }
  ffb_linear_desc const &d = h->d;
  PeerTable const *pt = comm_peer_table(comm);
  char const *force = ffb::env("FFB_LINEAR");
  if (!force && !ffb::env("FFB_DISABLE_FUSED_ALLREDUCE") &&
      linear_gemv_allreduce_supported(pt, d.quant_type, batch, d.in_dim, d.out_dim)) {
    return launch_linear_gemv_allreduce(pt, d.dtype, in, out, w, bias, batch, d.in_dim, d.out_dim, (cudaStream_t)stream);
  }
  int rc = ffb_linear_forward(h, in, out, w, bias, batch, stream);
  if (rc != FFB_OK) {
    return rc;
  }
  return ffb_allreduce(comm, out, out, (size_t)batch * d.out_dim, d.dtype, stream);
}

} // extern "C"
