// peer.h — NVLink peer-memory state shared by the communicator (comm.cu) and the fused o_proj + AllReduce kernel
// (linear_gemv.cu).  One process per GPU; every rank maps every other rank's receive area and flag array with CUDA
// IPC at communicator creation.  Replaces, for decode-sized messages, the ncclAllReduce of
// Kernels::AllReduce::inference_kernel_wrapper (src/parallel_ops/kernels/allreduce_kernels.cu:64-73).
#pragma once
#include <cuda_runtime.h>

namespace ffb {

constexpr int kPeerMaxRanks = 8;
constexpr int kPeerMaxTokens = 16;  // rows of a fused message (decode batches)
constexpr int kPeerMaxCtas = 160;   // flag slots per source rank (>= SM count)

// Passed BY VALUE to kernels.  recv[p] points at rank p's receive area, laid out
// [2 (epoch parity)][nranks (source rank)][kPeerMaxTokens][max_n] 8-byte words {fp32 value, int epoch}: the epoch
// stored with each value is its arrival flag (epochs only grow, nothing is ever reset).  flags[p] (rank p's
// [nranks][kPeerMaxCtas] ints) is spare signalling space for kernels that exchange whole tiles.
struct PeerTable {
  int nranks = 0, rank = 0, max_n = 0;
  float *recv[kPeerMaxRanks];
  int *flags[kPeerMaxRanks];
  int *epoch; // local: number of fused collectives completed on this communicator
  int *done;  // local: CTA completion counter of the collective in flight
};

} // namespace ffb

// C-ABI handle internals needed across translation units
struct ffb_comm_s;
namespace ffb {
// nullptr when the communicator has no peer mapping (single rank, IPC unavailable, adopted ncclComm_t)
PeerTable const *comm_peer_table(ffb_comm_s *c);
} // namespace ffb
