// linear_static.cu — ffb_linear_forward_static: the SM-pair tcgen05 projection GEMM (linear_tc2.cu) with its split-K
// workspace in a file-scope __device__ array instead of the handle's cudaMalloc'd one.  No counterpart in the
// reference; added for callers that cannot call cudaMalloc (code under CUDA
// graph capture, tight serving loops, harnesses that forbid allocation): the call allocates nothing, on the host
// side or the device side.
//
// The scratch: kStaticWsBytes (one 256 x 256 fp32 partial per SM pair: 74 x 256 KB = 19.4 MB) + one arrival flag per
// (SM pair, epilogue warp).  Every GPU that loads the library holds it until exit.  All calls share it, so two calls
// whose kernels overlap on the GPU (different streams) race; calls on one stream are ordered and safe.  The flags
// must be zero when a launch starts: the kernel leaves them zero, and every call clears them on its stream first
// anyway (an aborted launch may not have).
#include "common.cuh"

namespace ffb {
int linear_tc2_supported(int dtype, int M, int K, int N);
int launch_linear_tc2(int dtype, void const *in, void *out, void const *w, void const *bias, int M, int K, int N,
                      void *ws, int *cnt, cudaStream_t s);
size_t linear_tc2_ws_bytes();
size_t linear_tc2_cnt_bytes();
} // namespace ffb

namespace {

constexpr size_t kStaticWsBytes = (size_t)(ffb::kNumSMs / 2) * 256 * 256 * 4; // == linear_tc2_ws_bytes()
constexpr size_t kStaticCntInts = (size_t)(ffb::kNumSMs / 2) * 8;              // == linear_tc2_cnt_bytes() / 4
constexpr int kStaticMaxM = 1024, kStaticMaxK = 16384, kStaticMaxN = 16384;

static __device__ __align__(256) unsigned char g_static_ws[kStaticWsBytes];
static __device__ __align__(256) int g_static_cnt[kStaticCntInts];

// In host code g_static_ws / g_static_cnt are host-side stand-ins, not device addresses: the device addresses are
// looked up once per device, with that device current.
struct StaticScratch {
  void *ws;
  int *cnt;
};
constexpr int kMaxDevices = 64;
StaticScratch g_addr[kMaxDevices];
std::atomic<unsigned long long> g_addr_ok{0};

int static_scratch(StaticScratch *out) {
  int dev = 0;
  FFB_CUDA_TRY(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) {
    ffb::set_error("device ordinal %d out of range", dev);
    return FFB_ERR_INVALID;
  }
  if (!((g_addr_ok.load(std::memory_order_acquire) >> dev) & 1ULL)) {
    StaticScratch a{nullptr, nullptr};
    FFB_CUDA_TRY(cudaGetSymbolAddress(&a.ws, g_static_ws));
    FFB_CUDA_TRY(cudaGetSymbolAddress((void **)&a.cnt, g_static_cnt));
    g_addr[dev] = a; // (two threads racing here write the same values)
    g_addr_ok.fetch_or(1ULL << dev, std::memory_order_release);
  }
  *out = g_addr[dev];
  return FFB_OK;
}

} // namespace

extern "C" {

size_t ffb_linear_static_scratch_bytes(void) { return kStaticWsBytes + kStaticCntInts * sizeof(int); }

int ffb_linear_forward_static(void const *in, void *out, void const *w, void const *bias, int batch, int in_dim,
                              int out_dim, int dtype, ffb_stream_t stream) {
  FFB_REQUIRE(in && out && w, "null argument");
  FFB_REQUIRE(dtype == FFB_DT_HALF || dtype == FFB_DT_BF16, "dtype %d not supported (half/bf16 only)", dtype);
  FFB_REQUIRE(batch >= 0 && batch <= kStaticMaxM, "batch %d exceeds the static-scratch limit of %d tokens", batch, kStaticMaxM);
  FFB_REQUIRE(in_dim > 0 && in_dim <= kStaticMaxK && out_dim > 0 && out_dim <= kStaticMaxN,
              "in_dim %d / out_dim %d exceed the static-scratch limits (%d / %d)", in_dim, out_dim, kStaticMaxK, kStaticMaxN);
  if (batch == 0) {
    return FFB_OK;
  }
  int rc = ffb_check_device();
  if (rc != FFB_OK) {
    return rc;
  }
  if (!ffb::linear_tc2_supported(dtype, batch, in_dim, out_dim)) {
    ffb::set_error("ffb_linear_forward_static needs in_dim %% 64 == 0 and out_dim %% 16 == 0 (got %d, %d)", in_dim, out_dim);
    return FFB_ERR_UNSUPPORTED;
  }
  if (ffb::linear_tc2_ws_bytes() > kStaticWsBytes || ffb::linear_tc2_cnt_bytes() > kStaticCntInts * sizeof(int)) {
    ffb::set_error("static scratch is smaller than the kernel's workspace");
    return FFB_ERR_STATE;
  }
  StaticScratch sc;
  rc = static_scratch(&sc);
  if (rc != FFB_OK) {
    return rc;
  }
  cudaStream_t s = (cudaStream_t)stream;
  FFB_CUDA_TRY(cudaMemsetAsync(sc.cnt, 0, kStaticCntInts * sizeof(int), s));
  return ffb::launch_linear_tc2(dtype, in, out, w, bias, batch, in_dim, out_dim, sc.ws, sc.cnt, s);
}

} // extern "C"
