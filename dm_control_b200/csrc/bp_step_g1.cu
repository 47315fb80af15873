// Instantiation: 1 lane(s) per environment, image in global memory (interleave 32).
#include "bp_pipeline.cuh"

extern "C" int bp_launch_step_g1(const BpParams* P, cudaStream_t stream) {
  return bp_launch_gmem<1, 32>(P, stream);
}
