// Instantiation: 4 lane(s) per environment, image in global memory, LEAN feature set
// (Ctx<G, S, true>, csrc/bp_math.cuh): the plain locomotion models.
#include "bp_pipeline.cuh"

extern "C" int bp_launch_step_g4_lean(const BpParams* P, cudaStream_t stream) {
  return bp_launch_gmem<4, 1, 1>(P, stream);
}
