// incr_decoding.cc — a C++ host program over the C ABI: incremental decoding of an attention-only LLaMA-style model,
// the shape of the reference's inference/incr_decoding/incr_decoding.cc + inference/models/llama.cc restricted to the
// operators of the hot path.  Everything the reference delegates to Legion is a plain loop here; every operator call
// goes through flexflow_b200/host/ffb_ops.h (the reference's wrapper names) into libffb200.so:
//   FileDataLoader            HF-export weight files -> qkv_proj / o_proj / norm / lm_head / embedding tensors
//   RequestManager            register_new_request, prepare_next_batch (serve_incr_decoding's schedule,
//                             src/runtime/request_manager.cc:3012-3081, synchronous)
//   per layer                 ResidualRMSNorm -> Linear(qkv_proj) -> IncMultiHeadSelfAttention -> Linear(o_proj)
//                             (the residual add of layer i is fused into layer i + 1's norm, llama.cc:60-93)
//   head                      ResidualRMSNorm -> Linear(lm_head) -> Softmax -> ArgMax -> InferenceResult D2H
// The embedding lookup runs on the host (one H2D of the step's hidden rows).  No MLP: it is outside the path
// (SURVEY §8); tests/test_gpu_cpp_host.py runs the same model through the Python mirror and compares the tokens.
//
//   incr_decoding <weights_folder> <layers> <hidden> <heads> <kv_heads> <vocab> <max_new_tokens> <prompts_file>
// prompts_file: one request per line, token ids separated by spaces.  Prints one line per request:
//   TOKENS <guid> <id> <id> ...
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <memory>
#include <sstream>

#include "ffb_ops.h"

namespace {

#define CUDA_OK(expr)                                                                   \
  do {                                                                                  \
    cudaError_t e_ = (expr);                                                            \
    if (e_ != cudaSuccess) {                                                            \
      fprintf(stderr, "%s failed: %s\n", #expr, cudaGetErrorString(e_));                \
      exit(2);                                                                          \
    }                                                                                   \
  } while (0)

struct DevBuf {
  void *p = nullptr;
  explicit DevBuf(size_t bytes) { CUDA_OK(cudaMalloc(&p, bytes)); }
  ~DevBuf() { cudaFree(p); }
  DevBuf(DevBuf const &) = delete;
  DevBuf &operator=(DevBuf const &) = delete;
};

std::unique_ptr<DevBuf> upload(std::vector<uint16_t> const &host) {
  auto d = std::make_unique<DevBuf>(host.size() * 2);
  CUDA_OK(cudaMemcpy(d->p, host.data(), host.size() * 2, cudaMemcpyHostToDevice));
  return d;
}

struct Layer {
  std::unique_ptr<DevBuf> norm_w, wqkv, wo;
  std::unique_ptr<ffb::LinearMeta> lin_qkv, lin_o;
  std::unique_ptr<ffb::IncMultiHeadSelfAttentionMeta> attn;
};

} // namespace

int main(int argc, char **argv) {
  if (argc != 9) {
    fprintf(stderr, "usage: %s weights_folder layers hidden heads kv_heads vocab max_new_tokens prompts_file\n", argv[0]);
    return 1;
  }
  std::string const folder = argv[1];
  int const L = atoi(argv[2]), hidden = atoi(argv[3]), nh = atoi(argv[4]), nkv = atoi(argv[5]), vocab = atoi(argv[6]);
  int const max_new = atoi(argv[7]);
  int const dh = hidden / nh, H = nh * dh;
  int const MAXR = 4, MAXT = 64, MAXSEQ = 256;
  float const eps = 1e-6f;
  try {
    ffb::check(ffb_check_device(), "ffb_check_device");
    cudaStream_t stream;
    CUDA_OK(cudaStreamCreate(&stream));

    // ---- FFHandler: device batch descriptor + pinned staging block ----
    ffb::FFHandler handler;
    CUDA_OK(cudaMalloc((void **)&handler.batch_config_metadata, sizeof(ffb_combined_batch_config_meta)));
    CUDA_OK(cudaMallocHost(&handler.staging_pinned, sizeof(ffb_combined_batch_config_meta)));

    // ---- weights from HF-export files ----
    ffb::FileDataLoader loader(folder, nh, nkv, hidden, dh, 1, FFB_DT_HALF);
    std::vector<uint16_t> emb((size_t)vocab * hidden), tmp;
    loader.load_from_file("embed_tokens.weight", emb.data(), emb.size());
    ffb::RotaryEmbeddingMeta rope;
    rope.apply_rotary_embedding = true;
    std::vector<Layer> layers((size_t)L);
    for (int i = 0; i < L; i++) {
      std::string const base = "layers." + std::to_string(i);
      tmp.assign((size_t)hidden, 0);
      loader.load_from_file(base + ".input_layernorm.weight", tmp.data(), tmp.size());
      layers[i].norm_w = upload(tmp);
      tmp.assign((size_t)3 * H * hidden, 0);
      loader.load_attention_weights_to_dense(base + ".self_attn", false, 0, tmp.data());
      layers[i].wqkv = upload(tmp);
      tmp.assign((size_t)hidden * H, 0);
      loader.load_attention_weights_to_dense(base + ".self_attn", true, 0, tmp.data());
      layers[i].wo = upload(tmp);
      layers[i].lin_qkv = std::make_unique<ffb::LinearMeta>(handler, FFB_DT_HALF, FFB_DT_NONE, hidden, 3 * H, MAXT);
      layers[i].lin_o = std::make_unique<ffb::LinearMeta>(handler, FFB_DT_HALF, FFB_DT_NONE, H, hidden, MAXT);
      layers[i].attn = std::make_unique<ffb::IncMultiHeadSelfAttentionMeta>(
          handler, FFB_INC_DECODING_MODE, FFB_DT_HALF, dh, rope, false, true, false, 1.0f, nh, nh, nkv, MAXR, MAXT, MAXSEQ, 0);
    }
    tmp.assign((size_t)hidden, 0);
    loader.load_from_file("norm.weight", tmp.data(), tmp.size());
    auto final_norm_w = upload(tmp);
    tmp.assign((size_t)vocab * hidden, 0);
    loader.load_from_file("lm_head.weight", tmp.data(), tmp.size());
    auto lm_head_w = upload(tmp);
    ffb::LinearMeta lin_head(handler, FFB_DT_HALF, FFB_DT_NONE, hidden, vocab, MAXT);

    // ---- activations ----
    DevBuf x((size_t)MAXT * hidden * 2), res((size_t)MAXT * hidden * 2), xn((size_t)MAXT * hidden * 2);
    DevBuf qkv((size_t)MAXT * 3 * H * 2), att((size_t)MAXT * H * 2), o((size_t)MAXT * hidden * 2);
    DevBuf logits((size_t)MAXT * vocab * 2), probs((size_t)MAXT * vocab * 2), idx((size_t)MAXT * sizeof(int));
    uint16_t *x_host = nullptr;
    CUDA_OK(cudaMallocHost((void **)&x_host, (size_t)MAXT * hidden * 2));

    // ---- requests ----
    ffb::RequestManager rm;
    rm.set_max_requests_per_batch(MAXR);
    rm.set_max_tokens_per_batch(MAXT);
    rm.set_max_spec_tree_token_num(0);
    rm.set_max_sequence_length(MAXSEQ);
    std::vector<size_t> guids;
    std::ifstream pf(argv[8]);
    for (std::string line; std::getline(pf, line);) {
      std::istringstream is(line);
      std::vector<int> toks;
      for (int t; is >> t;) {
        toks.push_back(t);
      }
      if (!toks.empty()) {
        guids.push_back(rm.register_new_request(toks, -1, max_new));
      }
    }

    // ---- serve_incr_decoding, synchronous ----
    auto bc = std::make_unique<ffb_batch_config>();
    auto next = std::make_unique<ffb_batch_config>();
    auto ir = std::make_unique<ffb_inference_result>();
    memset(bc.get(), 0, sizeof(ffb_batch_config));
    for (int i = 0; i < FFB_MAX_NUM_REQUESTS; i++) {
      bc->request_completed[i] = true;
    }
    ffb::Kernels::ResidualRMSNorm::ResidualRMSNormMeta norm_meta{eps, hidden};
    int steps = 0;
    for (;; steps++) {
      rm.prepare_next_batch(*bc, *ir, next.get());
      std::swap(bc, next);
      int active = 0;
      for (int i = 0; i < MAXR; i++) {
        active += bc->request_completed[i] ? 0 : 1;
      }
      if (active == 0 && rm.num_pending() == 0) {
        break;
      }
      int const T = bc->num_tokens;
      handler.load_batch_config(bc.get(), FFB_INC_DECODING_MODE, stream);
      for (int t = 0; t < T; t++) { // embedding lookup on the host, one H2D
        memcpy(x_host + (size_t)t * hidden, emb.data() + (size_t)bc->tokensInfo[t].token_id * hidden, (size_t)hidden * 2);
      }
      CUDA_OK(cudaMemcpyAsync(x.p, x_host, (size_t)T * hidden * 2, cudaMemcpyHostToDevice, stream));
      using RA = ffb::GenericTensorAccessorR;
      using WA = ffb::GenericTensorAccessorW;
      void *stream_in = x.p; // the residual stream entering the layer
      for (int i = 0; i < L; i++) {
        Layer &ly = layers[i];
        if (i == 0) {
          ffb::Kernels::RMSNorm::inference_kernel_wrapper(&norm_meta, bc.get(), RA{FFB_DT_HALF, stream_in, (size_t)hidden},
                                                          RA{FFB_DT_HALF, ly.norm_w->p, (size_t)hidden},
                                                          WA{FFB_DT_HALF, xn.p, (size_t)hidden}, stream);
        } else { // h_i = h_{i-1} + o_{i-1}, fused with the norm
          ffb::Kernels::ResidualRMSNorm::inference_kernel_wrapper(
              &norm_meta, bc.get(), RA{FFB_DT_HALF, stream_in, (size_t)hidden}, RA{FFB_DT_HALF, o.p, (size_t)hidden},
              RA{FFB_DT_HALF, ly.norm_w->p, (size_t)hidden}, WA{FFB_DT_HALF, res.p, (size_t)hidden},
              WA{FFB_DT_HALF, xn.p, (size_t)hidden}, stream);
          CUDA_OK(cudaMemcpyAsync(x.p, res.p, (size_t)T * hidden * 2, cudaMemcpyDeviceToDevice, stream));
          stream_in = x.p;
        }
        ffb::Kernels::Linear::inference_kernel_wrapper(ly.lin_qkv.get(), bc.get(), xn.p, qkv.p, ly.wqkv->p, nullptr, hidden,
                                                       3 * H, T, stream);
        ffb::IncMultiHeadSelfAttention::inference_kernel_wrapper(ly.attn.get(), bc.get(), 0, RA{FFB_DT_HALF, qkv.p, (size_t)3 * H},
                                                                 WA{FFB_DT_HALF, att.p, (size_t)H}, stream);
        ffb::Kernels::Linear::inference_kernel_wrapper(ly.lin_o.get(), bc.get(), att.p, o.p, ly.wo->p, nullptr, H, hidden, T,
                                                       stream);
      }
      ffb::Kernels::ResidualRMSNorm::inference_kernel_wrapper(
          &norm_meta, bc.get(), RA{FFB_DT_HALF, stream_in, (size_t)hidden}, RA{FFB_DT_HALF, o.p, (size_t)hidden},
          RA{FFB_DT_HALF, final_norm_w->p, (size_t)hidden}, WA{FFB_DT_HALF, res.p, (size_t)hidden},
          WA{FFB_DT_HALF, xn.p, (size_t)hidden}, stream);
      ffb::Kernels::Linear::inference_kernel_wrapper(&lin_head, bc.get(), xn.p, logits.p, lm_head_w->p, nullptr, hidden, vocab, T,
                                                     stream);
      ffb::Softmax::inference_kernel_wrapper(bc.get(), RA{FFB_DT_HALF, logits.p, (size_t)vocab}, WA{FFB_DT_HALF, probs.p, (size_t)vocab},
                                             vocab, stream);
      ffb::ArgMax::forward_kernel_wrapper(false, RA{FFB_DT_HALF, probs.p, (size_t)vocab}, (int *)idx.p, nullptr, nullptr, T, vocab,
                                          stream);
      ffb::fetch_inference_result((int const *)idx.p, T, ir.get(), stream);
    }
    for (size_t g : guids) {
      printf("TOKENS %zu", g);
      for (int t : rm.request_tokens(g)) {
        printf(" %d", t);
      }
      printf("\n");
    }
    fprintf(stderr, "incr_decoding: %zu requests, %d steps\n", guids.size(), steps);
    layers.clear();
    cudaFreeHost(x_host);
    cudaFreeHost(handler.staging_pinned);
    cudaFree(handler.batch_config_metadata);
    cudaStreamDestroy(stream);
  } catch (ffb::Error const &e) {
    fprintf(stderr, "ffb error %d: %s\n", e.code, e.what());
    return 3;
  }
  return 0;
}
