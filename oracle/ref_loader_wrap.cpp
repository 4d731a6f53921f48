// ref_loader_wrap.cpp — compiles the REFERENCE's own weight-file readers into oracle/_ref/libref_loader.so (TEST
// INFRASTRUCTURE ONLY): load_attention_weights_to_dense_v2<DT>, load_from_file<DT> and load_from_quantized_file
// (/root/reference/src/runtime/file_loader.cc:215-343, :345-377, :552-669) and join_path
// (src/runtime/inference_manager.cc:824-841), streamed by oracle/Makefile (lift_fn.awk) into a temporary include that is
// deleted after the compile.  tests/test_weights_io.py writes weight files, reads them back through the product's
// ffb_load_* entry points and through these functions and compares the buffers byte for byte.
// Supplied here: `half` as a 16-bit storage type (the readers only copy elements), the DataType enum from the
// reference's own ffconst.h, C entry points.
#include <cassert>
#include <cstdint>
#include <cstring>
#include <fstream>
#include <iostream>
#include <sstream>
#include <string>
#include <vector>

#include "flexflow/ffconst.h"

typedef uint16_t half;
using namespace std;

namespace FlexFlow {
#include REF_LOADER_FUNCS_INC
} // namespace FlexFlow

namespace {
struct Quiet {
  std::ostringstream sink;
  std::streambuf *old;
  Quiet() : old(std::cout.rdbuf(sink.rdbuf())) {}
  ~Quiet() { std::cout.rdbuf(old); }
};
} // namespace

extern "C" {

// elt_bytes 2 (half) or 4 (float); dst holds the FULL weight (all TP shards): qkv [tp][Q_r | K_r | V_r], o_proj as stored
void ref_load_attention_weights_to_dense_v2(void *dst, int elt_bytes, int num_heads, int num_kv_heads, size_t hidden_dim,
                                            size_t qkv_inner_dim, char const *layer_name, char const *weights_folder,
                                            size_t volume, int tp_degree, int load_o_proj) {
  Quiet q;
  if (elt_bytes == 2) {
    FlexFlow::load_attention_weights_to_dense_v2<uint16_t>((uint16_t *)dst, num_heads, num_kv_heads, hidden_dim, qkv_inner_dim,
                                                          layer_name, weights_folder, volume, tp_degree, load_o_proj != 0);
  } else {
    FlexFlow::load_attention_weights_to_dense_v2<float>((float *)dst, num_heads, num_kv_heads, hidden_dim, qkv_inner_dim,
                                                       layer_name, weights_folder, volume, tp_degree, load_o_proj != 0);
  }
}

void ref_load_from_file(void *dst, int elt_bytes, size_t count, char const *path) {
  Quiet q;
  if (elt_bytes == 2) {
    FlexFlow::load_from_file<uint16_t>((uint16_t *)dst, count, path);
  } else {
    FlexFlow::load_from_file<float>((float *)dst, count, path);
  }
}

// data_type: DT_INT4 / DT_INT8 of the reference's ffconst.h (46 / 47)
void ref_load_from_quantized_file(char *dst, size_t blob_bytes, char const *path, int data_type, int use_full_precision) {
  Quiet q;
  FlexFlow::load_from_quantized_file(dst, blob_bytes, path, (DataType)data_type, use_full_precision != 0);
}

} // extern "C"
