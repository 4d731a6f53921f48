// see orc_common.h for the header comment (test infrastructure only)
#include "orc_common.h"


extern "C" {

// Conversions exported for the tests (vectorised helpers so numpy does not need its own half model).
void orc_round_to_dt(float *x, size_t n, int dt) {
  for (size_t i = 0; i < n; i++) {
    x[i] = round_dt(x[i], dt);
  }
}
void orc_float_to_half_bits(float const *x, uint16_t *out, size_t n) {
  for (size_t i = 0; i < n; i++) {
    out[i] = float_to_half_bits(x[i]);
  }
}
void orc_half_bits_to_float(uint16_t const *x, float *out, size_t n) {
  for (size_t i = 0; i < n; i++) {
    out[i] = half_bits_to_float(x[i]);
  }
}

} // extern "C"


static int g_orc_threads = 1;
int orc_threads() { return g_orc_threads; }
extern "C" void orc_set_threads(int n) { g_orc_threads = n < 1 ? 1 : n; }
