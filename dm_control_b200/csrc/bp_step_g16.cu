// Instantiation: 16 lane(s) per environment, image in global memory (interleave 1).
#include "bp_pipeline.cuh"

extern "C" int bp_launch_step_g16(const BpParams* P, cudaStream_t stream) {
  return bp_launch_gmem<16, 1>(P, stream);
}
