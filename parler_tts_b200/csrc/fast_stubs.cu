// Placeholders for fast-family kernels that are not implemented yet: the dispatcher falls back to the simple family.
#include "ops.cuh"
namespace ptts {
bool fast_conv_supported(int, int, int) { return false; }
int dac_conv1d_fast(const float*, int, int, int, const float*, const float*, int, int, int, const float*, const float*,
                    int, float*, cudaStream_t) {
  set_error("dac_conv1d_fast: not built");
  return PTTS_E_UNSUPPORTED;
}
}  // namespace ptts
