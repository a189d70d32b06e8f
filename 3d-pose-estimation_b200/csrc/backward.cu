// Backward pass — placeholder until the kernels land.
#include "common.cuh"
namespace smplb {
size_t backward_scratch_bytes(int B) { (void)B; return 0; }
}
extern "C" int32_t smplb_backward(const smplb_model*, const smplb_backward_io*, void*, size_t, int32_t, void*) {
  return smplb::fail(SMPLB_E_UNSUPPORTED, "smplb_backward: not built yet");
}
extern "C" int32_t smplb_project_backward(int32_t, int32_t, const float*, const float*, const float*, const float*,
                                          int32_t, const float*, float*, float*, float*, void*) {
  return smplb::fail(SMPLB_E_UNSUPPORTED, "smplb_project_backward: not built yet");
}
