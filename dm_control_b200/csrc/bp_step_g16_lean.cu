// Instantiation: 16 lane(s) per environment, image in global memory, LEAN feature set
// (Ctx<G, S, true>, csrc/bp_math.cuh): the plain locomotion models.
#include "bp_pipeline.cuh"

extern "C" int bp_launch_step_g16_lean(const BpParams* P, cudaStream_t stream) {
  return bp_launch_gmem<16, 1, 1>(P, stream);
}

#ifdef BP_TIMELINE
extern "C" int bp_debug_tl2_g16_lean(unsigned long long* out, int reset) {
  cudaDeviceSynchronize();
  cudaMemcpyFromSymbol(out, bp_tl2, sizeof(unsigned long long) * 64);
  if (reset) { unsigned long long z[64] = {0}; cudaMemcpyToSymbol(bp_tl2, z, sizeof(z)); }
  return 0;
}
#endif
