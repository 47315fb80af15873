// Instantiation: 16 lane(s) per environment, image in global memory, MID feature set
// (Ctx<G, S, 2>, csrc/bp_math.cuh): the quadruped models.
// the only kernel of this translation unit: stages inlined at the compiler's discretion; the
// larger MID code keeps the stage barriers 0-3 next to the Newton-loop barrier, like the static
// quadruped-escape instance (profiles/r2_tuning.md)
#define BP_STAGE_AUTO_INLINE
#define BP_LOCKMASK 0x10f
#include "bp_pipeline.cuh"

extern "C" int bp_launch_step_g16_mid(const BpParams* P, cudaStream_t stream) {
  return bp_launch_gmem<16, 1, 2>(P, stream);
}
