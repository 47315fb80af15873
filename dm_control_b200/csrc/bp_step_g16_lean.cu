// Instantiation: 16 lane(s) per environment, image in global memory, LEAN feature set
// (Ctx<G, S, true>, csrc/bp_math.cuh): the plain locomotion models.
// the only kernel of this translation unit: stages inlined at the compiler's discretion and
// lock-step on the Newton-loop barrier alone, as in the static instances (cheetah through this
// kernel 30.2 -> 37.0 M env-steps/s, walker 2.66 -> 3.09 M; profiles/r2_tuning.md)
#define BP_STAGE_AUTO_INLINE
#define BP_LOCKMASK 0x100
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
