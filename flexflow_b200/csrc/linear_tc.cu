// linear_tc.cu — placeholder until the tcgen05/TMA GEMM lands (next commit): reports "unsupported" so
// ffb_linear_forward streams the weights in 64-row passes.
#include "common.cuh"
namespace ffb {
struct LinearTcState {};
int linear_tc_supported(int, int, int, int) { return 0; }
LinearTcState *linear_tc_create(int, int, int, int, int) { return nullptr; }
void linear_tc_destroy(LinearTcState *) {}
int launch_linear_tc(LinearTcState *, int, int, void const *, void *, void const *, void const *, int, int, int, cudaStream_t) {
  set_error("tcgen05 GEMM not built");
  return FFB_ERR_UNSUPPORTED;
}
} // namespace ffb
