// Instantiation: 8 lane(s) per environment, image in global memory, MID feature set
// (Ctx<G, S, 2>, csrc/bp_math.cuh): the quadruped models.
#include "bp_pipeline.cuh"

extern "C" int bp_launch_step_g8_mid(const BpParams* P, cudaStream_t stream) {
  return bp_launch_gmem<8, 1, 2>(P, stream);
}
