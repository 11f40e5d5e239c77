// Placeholders for fast-family kernels that are not implemented yet: the dispatcher falls back to the simple family.
#include "ops.cuh"
namespace ptts {
bool fast_conv_supported(int, int, int) { return false; }
int dac_conv1d_fast(const float*, int, int, int, const float*, const float*, int, int, int, const float*, const float*,
                    int, float*, cudaStream_t) {
  set_error("dac_conv1d_fast: not built");
  return PTTS_E_UNSUPPORTED;
}
bool fast_linear_supported(int, int, int, int, int) { return false; }
int linear_fast(const void*, int, const void*, int, const float*, int, int, int, float, int, const float*, const int*,
                void*, int, cudaStream_t) {
  set_error("linear_fast: not built");
  return PTTS_E_UNSUPPORTED;
}
bool fast_attention_supported(int, int, int) { return false; }
int attention_fast(const void*, const void*, const void*, int, int, int, int, int, int, int, const int*, int,
                   const int*, int, void*, float*, cudaStream_t) {
  set_error("attention_fast: not built");
  return PTTS_E_UNSUPPORTED;
}
}  // namespace ptts
