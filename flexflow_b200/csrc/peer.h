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
  // ---- large messages (peer_allreduce.cu): two-shot allreduce, reduce-scatter then all-gather, both by PUSH ----
  // big[p]: rank p's area [2 (epoch parity)][2 (0 = reduce-scatter, 1 = all-gather)][nranks (source)][big_slice_bytes];
  // big_flags[p]: rank p's [2 (phase)][nranks (source)][kPeerBigChunks] ints holding the epoch of the last chunk that
  // has landed (epochs only grow)
  char *big[kPeerMaxRanks];
  int *big_flags[kPeerMaxRanks];
  size_t big_slice_bytes = 0;
  int *big_epoch; // local
  int *big_done;  // local
  // ---- the same two shots with flag-in-data lines (peer_allreduce.cu, messages up to kPeerLLMaxBytes) ----
  // ll[p]: rank p's area [2 (phase)][nranks (source)][ll_slice_bytes] of 16-byte lines {data, epoch, data, epoch}: every
  // 8-byte half carries its own arrival flag, so no fence and no separate flag store is needed
  char *ll[kPeerMaxRanks];
  size_t ll_slice_bytes = 0;
  int *ll_epoch; // local
  int *ll_done;  // local
};
constexpr int kPeerBigChunks = 592;                        // CTAs (= flag slots per source) of the large-message kernel
constexpr size_t kPeerBigMaxBytes = (size_t)16 << 20;      // largest message: 1024 tokens x 8192 x 2 bytes
constexpr size_t kPeerLLMaxBytes = (size_t)16 << 20;       // largest flag-in-data message (its lines take twice the bytes)

} // namespace ffb

// C-ABI handle internals needed across translation units
struct ffb_comm_s;
namespace ffb {
// nullptr when the communicator has no peer mapping (single rank, IPC unavailable, adopted ncclComm_t)
PeerTable const *comm_peer_table(ffb_comm_s *c);
// two-shot allreduce of `count` 16-bit elements over peer memory (peer_allreduce.cu); in-place allowed
int peer_allreduce_supported(PeerTable const *pt, size_t count, int dtype);
int launch_peer_allreduce(PeerTable const *pt, void const *in, void *out, size_t count, int dtype, cudaStream_t s);
} // namespace ffb
