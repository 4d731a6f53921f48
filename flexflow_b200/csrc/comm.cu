// comm.cu — head-sharded AllReduce parallel op over NCCL (NVLink 5 / NVSwitch).
// Replaces Kernels::AllReduce::inference_kernel_wrapper (src/parallel_ops/kernels/allreduce_kernels.cu:56-78)
// and the communicator bring-up of src/runtime/model.cc:584-628.  NCCL is resolved at run time with
// dlopen so the library loads (and every other entry point works) in a process without NCCL, and so
// that inside a PyTorch process the already-loaded libnccl.so.2 is shared instead of duplicated.
#include <dlfcn.h>
#include <cstring>

#include <new>

#include "common.cuh"
#include "peer.h"

namespace {

typedef struct ncclComm *ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
enum { ncclSuccess = 0 };
enum { ncclFloat16 = 6, ncclFloat32 = 7, ncclBfloat16 = 9 }; // nccl.h ncclDataType_t
enum { ncclSum = 0 };

struct NcclApi {
  void *lib = nullptr;
  int (*GetUniqueId)(ncclUniqueId *) = nullptr;
  int (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
  int (*CommDestroy)(ncclComm_t) = nullptr;
  int (*AllReduce)(void const *, void *, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
  int (*AllGather)(void const *, void *, size_t, int, ncclComm_t, cudaStream_t) = nullptr;
  char const *(*GetErrorString)(int) = nullptr;
};

NcclApi load_nccl() {
  NcclApi api;
  char const *names[] = {"libnccl.so.2", "libnccl.so"};
  for (char const *n : names) {
    api.lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
    if (api.lib) {
      break;
    }
  }
  if (api.lib) {
    api.GetUniqueId = (decltype(api.GetUniqueId))dlsym(api.lib, "ncclGetUniqueId");
    api.CommInitRank = (decltype(api.CommInitRank))dlsym(api.lib, "ncclCommInitRank");
    api.CommDestroy = (decltype(api.CommDestroy))dlsym(api.lib, "ncclCommDestroy");
    api.AllReduce = (decltype(api.AllReduce))dlsym(api.lib, "ncclAllReduce");
    api.AllGather = (decltype(api.AllGather))dlsym(api.lib, "ncclAllGather");
    api.GetErrorString = (decltype(api.GetErrorString))dlsym(api.lib, "ncclGetErrorString");
  }
  return api;
}

NcclApi *nccl() {
  static NcclApi api = load_nccl(); // initialised once, thread-safe (C++11)
  if (!api.lib || !api.GetUniqueId || !api.CommInitRank || !api.CommDestroy || !api.AllReduce) {
    ffb::set_error("NCCL (libnccl.so.2) could not be loaded: %s", dlerror());
    return nullptr;
  }
  return &api;
}

#define FFB_NCCL_TRY(api, expr)                                                               \
  do {                                                                                        \
    int _r = (expr);                                                                          \
    if (_r != ncclSuccess) {                                                                  \
      ffb::set_error("%s failed: %s", #expr, (api)->GetErrorString ? (api)->GetErrorString(_r) : "?"); \
      return FFB_ERR_NCCL;                                                                    \
    }                                                                                         \
  } while (0)

} // namespace

struct ffb_comm_s {
  ncclComm_t comm = nullptr;
  int nranks = 1, rank = 0;
  bool owned = false;
  // NVLink peer mapping for the fused o_proj + AllReduce kernel (peer.h)
  bool peer_ok = false;
  ffb::PeerTable peer;
  void *peer_local = nullptr;                      // this rank's allocation (receive area + flags + counters)
  void *peer_mapped[ffb::kPeerMaxRanks] = {nullptr}; // cudaIpcOpenMemHandle results (other ranks)
};

namespace ffb {
PeerTable const *comm_peer_table(ffb_comm_s *c) { return c && c->peer_ok ? &c->peer : nullptr; }
} // namespace ffb

namespace {

constexpr int kPeerMaxN = 8192; // widest hidden size served by the fused path (LLaMA-2-70B)

// Best effort, but AGREED: every rank of a communicator ends up with the same peer_ok, because a rank that launched the
// fused peer-push kernel while another fell back to NCCL would spin forever.  All ranks therefore take part in every
// collective below whatever happened locally (the only early return, on the rank-independent conditions nranks /
// AllGather availability, is taken by all ranks alike); local failures - allocation, IPC export or open, or
// FFB_DISABLE_PEER set in this rank's environment - travel in an ok flag that is min-reduced across ranks at the end.
void peer_setup(ffb_comm_s *c, NcclApi *api) {
  using namespace ffb;
  if (c->nranks < 2 || c->nranks > kPeerMaxRanks || !api->AllGather) {
    return;
  }
  size_t const recv_bytes = (size_t)2 * c->nranks * kPeerMaxTokens * kPeerMaxN * 2 * sizeof(float); // {value, epoch}
  size_t const flag_bytes = (size_t)c->nranks * kPeerMaxCtas * sizeof(int);
  // large-message area: per rank 2 parities x 2 phases x one whole message, + its flags
  size_t const small_bytes = (recv_bytes + flag_bytes + 256 + 1023) & ~(size_t)1023;
  size_t const big_slice = ((kPeerBigMaxBytes / c->nranks + 1023) & ~(size_t)1023);
  size_t const big_bytes = (size_t)4 * c->nranks * big_slice;
  size_t const big_flag_bytes = ((size_t)2 * c->nranks * kPeerBigChunks * sizeof(int) + 256 + 1023) & ~(size_t)1023;
  // flag-in-data area: 2 phases x nranks slots x (one slice of the largest such message, doubled by its flags)
  size_t const ll_slice = ((2 * kPeerLLMaxBytes / c->nranks + 1023) & ~(size_t)1023);
  size_t const ll_bytes = (size_t)2 * c->nranks * ll_slice + 1024; // + its epoch / completion counter
  size_t const total = small_bytes + big_flag_bytes + big_bytes + ll_bytes;
  struct Rec {
    cudaIpcMemHandle_t handle;
    int ok;
  };
  Rec mine, all[kPeerMaxRanks];
  memset(&mine, 0, sizeof(mine));
  bool ok = ffb::env("FFB_DISABLE_PEER") == nullptr;
  if (ok && cudaMalloc(&c->peer_local, total) != cudaSuccess) {
    c->peer_local = nullptr;
    cudaGetLastError();
    ok = false;
  }
  ok = ok && cudaMemset(c->peer_local, 0, total) == cudaSuccess && cudaIpcGetMemHandle(&mine.handle, c->peer_local) == cudaSuccess;
  mine.ok = ok ? 1 : 0;
  // exchange {handle, ok}: a few hundred bytes of device scratch; without it this rank cannot take part at all
  Rec *scratch = nullptr;
  if (cudaMalloc(&scratch, sizeof(Rec) * (c->nranks + 1)) != cudaSuccess) {
    cudaGetLastError();
    return; // (the other ranks would wait in the AllGather: there is nothing left to do for a process this starved)
  }
  bool exchanged = cudaMemcpy(scratch + c->nranks, &mine, sizeof(mine), cudaMemcpyHostToDevice) == cudaSuccess;
  exchanged = (api->AllGather(scratch + c->nranks, scratch, sizeof(Rec), /*ncclInt8*/ 0, c->comm, nullptr) == ncclSuccess) && exchanged;
  exchanged = exchanged && cudaStreamSynchronize(nullptr) == cudaSuccess &&
              cudaMemcpy(all, scratch, sizeof(Rec) * c->nranks, cudaMemcpyDeviceToHost) == cudaSuccess;
  ok = ok && exchanged;
  for (int r = 0; exchanged && r < c->nranks; r++) {
    ok = ok && all[r].ok != 0;
  }
  for (int r = 0; ok && r < c->nranks; r++) {
    void *base = c->peer_local;
    if (r != c->rank) {
      ok = cudaIpcOpenMemHandle(&c->peer_mapped[r], all[r].handle, cudaIpcMemLazyEnablePeerAccess) == cudaSuccess;
      base = c->peer_mapped[r];
    }
    c->peer.recv[r] = (float *)base;
    c->peer.flags[r] = (int *)((char *)base + recv_bytes);
    c->peer.big_flags[r] = (int *)((char *)base + small_bytes);
    c->peer.big[r] = (char *)base + small_bytes + big_flag_bytes;
    c->peer.ll[r] = (char *)base + small_bytes + big_flag_bytes + big_bytes + 1024;
  }
  // second agreement: did EVERY rank map every area?
  int flag = ok ? 1 : 0;
  bool agreed = cudaMemcpy(scratch, &flag, sizeof(int), cudaMemcpyHostToDevice) == cudaSuccess;
  agreed = (api->AllReduce(scratch, scratch, 1, /*ncclInt32*/ 2, /*ncclMin*/ 3, c->comm, nullptr) == ncclSuccess) && agreed;
  agreed = agreed && cudaStreamSynchronize(nullptr) == cudaSuccess &&
           cudaMemcpy(&flag, scratch, sizeof(int), cudaMemcpyDeviceToHost) == cudaSuccess;
  cudaFree(scratch);
  if (!agreed || flag == 0) {
    cudaGetLastError();
    return; // mappings (if any) are released in ffb_comm_destroy
  }
  c->peer.nranks = c->nranks;
  c->peer.rank = c->rank;
  c->peer.max_n = kPeerMaxN;
  c->peer.epoch = (int *)((char *)c->peer_local + recv_bytes + flag_bytes);
  c->peer.done = c->peer.epoch + 1;
  c->peer.big_slice_bytes = big_slice;
  c->peer.big_epoch = (int *)((char *)c->peer_local + small_bytes + (size_t)2 * c->nranks * kPeerBigChunks * sizeof(int));
  c->peer.big_done = c->peer.big_epoch + 1;
  c->peer.ll_slice_bytes = ll_slice;
  c->peer.ll_epoch = (int *)((char *)c->peer_local + small_bytes + big_flag_bytes + big_bytes);
  c->peer.ll_done = c->peer.ll_epoch + 1;
  c->peer_ok = true;
}

} // namespace

extern "C" {

int ffb_comm_get_unique_id(void *id_out) {
  FFB_REQUIRE(id_out, "null argument");
  NcclApi *api = nccl();
  if (!api) {
    return FFB_ERR_NCCL;
  }
  FFB_NCCL_TRY(api, api->GetUniqueId((ncclUniqueId *)id_out));
  return FFB_OK;
}

int ffb_comm_init_rank(void const *id, int nranks, int rank, ffb_comm_t *out) {
  FFB_REQUIRE(id && out && nranks >= 1 && rank >= 0 && rank < nranks, "bad communicator arguments");
  NcclApi *api = nccl();
  if (!api) {
    return FFB_ERR_NCCL;
  }
  ffb_comm_s *c = new (std::nothrow) ffb_comm_s();
  FFB_REQUIRE(c, "out of host memory");
  c->nranks = nranks;
  c->rank = rank;
  c->owned = true;
  ncclUniqueId uid;
  memcpy(&uid, id, sizeof(uid));
  int r = api->CommInitRank(&c->comm, nranks, uid, rank);
  if (r != ncclSuccess) {
    ffb::set_error("ncclCommInitRank failed: %s", api->GetErrorString ? api->GetErrorString(r) : "?");
    delete c;
    return FFB_ERR_NCCL;
  }
  peer_setup(c, api);
  *out = c;
  return FFB_OK;
}

int ffb_comm_has_peer(ffb_comm_t c) { return c && c->peer_ok ? 1 : 0; }

int ffb_comm_from_nccl(void *nccl_comm, int nranks, int rank, ffb_comm_t *out) {
  FFB_REQUIRE(nccl_comm && out, "null argument");
  ffb_comm_s *c = new (std::nothrow) ffb_comm_s();
  FFB_REQUIRE(c, "out of host memory");
  c->comm = (ncclComm_t)nccl_comm;
  c->nranks = nranks;
  c->rank = rank;
  c->owned = false;
  *out = c;
  return FFB_OK;
}

int ffb_comm_destroy(ffb_comm_t c) {
  if (!c) {
    return FFB_OK;
  }
  for (int r = 0; r < ffb::kPeerMaxRanks; r++) {
    if (c->peer_mapped[r]) {
      cudaIpcCloseMemHandle(c->peer_mapped[r]);
    }
  }
  if (c->peer_local) {
    cudaFree(c->peer_local);
  }
  if (c->owned && c->comm) {
    NcclApi *api = nccl();
    if (api) {
      api->CommDestroy(c->comm);
    }
  }
  delete c;
  return FFB_OK;
}

int ffb_allreduce(ffb_comm_t c, void const *in, void *out, size_t count, int dtype, ffb_stream_t stream) {
  FFB_REQUIRE(c && in && out, "null argument");
  if (count == 0) {
    return FFB_OK;
  }
  int const nt = dtype == FFB_DT_HALF ? ncclFloat16 : (dtype == FFB_DT_BF16 ? ncclBfloat16 : (dtype == FFB_DT_FLOAT ? ncclFloat32 : -1));
  FFB_REQUIRE(nt >= 0, "dtype %d not supported by allreduce", dtype);
  // messages up to 16 MiB on a peer-mapped communicator: two-shot allreduce over NVLink peer memory (peer_allreduce.cu);
  // FFB_DISABLE_PEER_ALLREDUCE=1 keeps the NCCL call (A/B, parity tests).  Every rank must take the same branch: the
  // condition depends only on values that are equal on all ranks (count, dtype, the agreed peer_ok) and this variable.
  static bool const peer_off = ffb::env("FFB_DISABLE_PEER_ALLREDUCE") != nullptr;
  if (!peer_off && c->peer_ok && ffb::peer_allreduce_supported(&c->peer, count, dtype)) {
    return ffb::launch_peer_allreduce(&c->peer, in, out, count, dtype, (cudaStream_t)stream);
  }
  NcclApi *api = nccl();
  if (!api) {
    return FFB_ERR_NCCL;
  }
  // same call as allreduce_kernels.cu:67-73
  FFB_NCCL_TRY(api, api->AllReduce(in, out, count, nt, ncclSum, c->comm, (cudaStream_t)stream));
  return FFB_OK;
}

} // extern "C"
