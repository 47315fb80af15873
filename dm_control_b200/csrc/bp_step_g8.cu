// Instantiation: 8 lane(s) per environment, image in global memory (interleave 1).
#include "bp_pipeline.cuh"

extern "C" int bp_launch_step_g8(const BpParams* P, cudaStream_t stream) {
  long long total = P->nenv * 8;
  const int threads = bp_block_threads(total);
  unsigned blocks = (unsigned)((total + threads - 1) / threads);
  bp_step_kernel_gmem<8, 1><<<blocks, threads, 0, stream>>>(*P);
  return (int)cudaGetLastError();
}
