// Instantiation: 1 lane(s) per environment, image in global memory (interleave 32).
#include "bp_pipeline.cuh"

extern "C" int bp_launch_step_g1(const BpParams* P, cudaStream_t stream) {
  return bp_launch_gmem<1, 32>(P, stream);
}

#ifdef BP_TIMELINE
extern "C" int bp_debug_timeline_g1(unsigned long long* out, int reset) {
  cudaDeviceSynchronize();
  cudaMemcpyFromSymbol(out, bp_timeline, sizeof(unsigned long long) * 32);
  if (reset) { unsigned long long z[32] = {0}; cudaMemcpyToSymbol(bp_timeline, z, sizeof(z)); }
  return 0;
}
#endif
