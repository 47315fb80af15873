// Dispatch over the static (ahead-of-time specialised) instances of the step kernel
// (csrc/bp_step_static.cuh; one generated translation unit per instance, static/instances.inc).
#include <cuda_runtime.h>

#include "bp_types.h"

#define BP_STATIC_DECLARE
#include "static/instances.inc"
#undef BP_STATIC_DECLARE

// Returns -1 when no instance serves this structure hash, lane width and layout, else the
// launch status.
extern "C" int bp_launch_step_static(const BpParams* P, unsigned long long hash, int plain,
                                     cudaStream_t stream) {
  int e = -1;
#include "static/instances.inc"
  return e;
}
