// Entry points not implemented yet in this round: they fail loudly.
#include "core.h"
using namespace teb;
namespace teb {
namespace nvfp4 {
void quantize(const Tensor &, Tensor &, const QuantConfig &, const Tensor *, bool, int, cudaStream_t) {
  TEB_ERROR("NVFP4 quantization is not implemented yet");
}
void rht_amax(const Tensor &, Tensor &, int, int, cudaStream_t) { TEB_ERROR("not implemented yet"); }
void rht(const Tensor &, Tensor &, int, int, cudaStream_t) { TEB_ERROR("not implemented yet"); }
}  // namespace nvfp4
}  // namespace teb
