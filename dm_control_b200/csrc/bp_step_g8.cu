// Instantiation: 8 lane(s) per environment, image in global memory (interleave 1).
#include "bp_pipeline.cuh"

extern "C" int bp_launch_step_g8(const BpParams* P, cudaStream_t stream) {
  return bp_launch_gmem<8, 1>(P, stream);
}

extern "C" int bp_launch_seg_g8(const BpParams* P, int seg, int threads, int first, int last,
                                cudaStream_t stream) {
  return bp_launch_seg_any<8>(P, seg, threads, first, last, stream);
}

#ifdef BP_TIMELINE
extern "C" int bp_debug_timeline_g8(unsigned long long* out, int reset) {
  cudaDeviceSynchronize();
  cudaMemcpyFromSymbol(out, bp_timeline, sizeof(unsigned long long) * 32);
  if (reset) { unsigned long long z[32] = {0}; cudaMemcpyToSymbol(bp_timeline, z, sizeof(z)); }
  return 0;
}
#endif
