// Instantiation: one warp per environment, image in shared memory (G = 32).
#include "bp_pipeline.cuh"

// ---- G = 32: one warp per environment, image staged in shared memory ---------
__global__ void __launch_bounds__(32)
bp_step_kernel_smem(const __grid_constant__ BpParams P) {
  extern __shared__ double smem[];
  const int lane = threadIdx.x;
  const long long env = blockIdx.x;
  if (env >= P.nenv || (P.mask && !P.mask[env])) return;
  double* g = P.image + env * P.stride;
  for (int w = lane; w < P.L.persist; w += 32) smem[w] = g[w];
  __syncwarp();
  const float* hf = P.hfield ? P.hfield + env * (long long)P.m.nhfielddata : nullptr;
  const double* gp = P.env_geom_pos ? P.env_geom_pos + env * 3LL * P.m.ngeom : (const double*)P.m.geom_pos;
  const double* sp = P.env_site_pos ? P.env_site_pos + env * 3LL * P.m.nsite : (const double*)P.m.site_pos;
  const double* bp = P.env_body_pos ? P.env_body_pos + env * 3LL * P.m.nbody : (const double*)P.m.body_pos;
  const double* bq = P.env_body_quat ? P.env_body_quat + env * 4LL * P.m.nbody : (const double*)P.m.body_quat;
  Ctx<32, 1> c{P, smem, (int*)(smem + P.L.ints), lane, 0xffffffffu, true, false, nullptr, hf,
               gp, sp, bp, bq,
               P.env_wrap_prm ? P.env_wrap_prm + env * (long long)P.env_wrap_n : (const double*)P.m.wrap_prm};
  bp_load_ctrl(c, env);
  bp_run_mode(c);
  __syncwarp();
  for (int w = lane; w < P.L.persist; w += 32) g[w] = smem[w];
  if (P.rew_out) {                  // epilogue reads the image just written back
    __threadfence_block();
    __syncwarp();
    if (lane == 0) bp_task_eval(P, P.task, env, P.task_nobs, P.obs_out, P.rew_out);
  }
}

extern "C" int bp_launch_step_g32(const BpParams* P, cudaStream_t stream) {
  size_t bytes = (size_t)P->L.total * sizeof(double);
  static size_t configured = 0;
  if (bytes > configured) {
    cudaError_t e = cudaFuncSetAttribute(bp_step_kernel_smem,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)bytes);
    if (e != cudaSuccess) return (int)e;
    configured = bytes;
  }
  bp_step_kernel_smem<<<(unsigned)P->nenv, 32, bytes, stream>>>(*P);
  return (int)cudaGetLastError();
}
