// K2 blend-shape GEMM on tcgen05 (3xTF32) — placeholder until the kernel lands.
#include "common.cuh"
namespace smplb {
void free_tmap_cache(smplb_model* m) { (void)m; }
int32_t launch_blend_tc(const smplb_model*, int, const float*, const float*, float*, cudaStream_t) {
  return fail(SMPLB_E_UNSUPPORTED, "blend_tc: not built yet");
}
}  // namespace smplb
