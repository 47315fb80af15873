// Instantiation: 1 lane(s) per environment, image in global memory (interleave 32).
#include "bp_pipeline.cuh"

extern "C" int bp_launch_step_g1(const BpParams* P, cudaStream_t stream) {
  long long total = P->nenv * 1;
  const int threads = bp_block_threads(total);
  unsigned blocks = (unsigned)((total + threads - 1) / threads);
  bp_step_kernel_gmem<1, 32><<<blocks, threads, 0, stream>>>(*P);
  return (int)cudaGetLastError();
}
