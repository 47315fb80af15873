// Instantiation: 2 lane(s) per environment, image in global memory (interleave 1).
#include "bp_pipeline.cuh"

extern "C" int bp_launch_step_g2(const BpParams* P, cudaStream_t stream) {
  long long total = P->nenv * 2;
  const int threads = bp_block_threads(total);
  unsigned blocks = (unsigned)((total + threads - 1) / threads);
  bp_step_kernel_gmem<2, 1><<<blocks, threads, 0, stream>>>(*P);
  return (int)cudaGetLastError();
}
