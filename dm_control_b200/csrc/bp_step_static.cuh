// Static (ahead-of-time specialised) instances of the step kernel: one thread per environment
// (G = 1, warp-interleaved images, every access of a warp one coalesced 256-byte line), the
// model's structure and the data layout compile-time constants (csrc/static/*.cuh), free-running
// warps.  VERDICT r1 item 4 / DESIGN.md section 8 item 0: per-model kernels.
//
// Why one thread per env here: with the structure known at compile time the tree loops unroll
// into straight-line code whose intermediate values stay in registers (constant image offsets
// let the compiler forward stores to loads), so an env no longer needs several lanes to walk
// its index arrays, and all 32 lanes of a warp do useful work.
#ifndef BP_STEP_STATIC_CUH_
#define BP_STEP_STATIC_CUH_
#include "bp_pipeline.cuh"

#ifndef BP_STATIC_TPB
#define BP_STATIC_TPB 64
#endif

template <class MT>
__global__ void __launch_bounds__(BP_STATIC_TPB)
bp_step_kernel_static(const __grid_constant__ BpParams P) {
  const long long env = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  // lanes past the end of a ragged last warp stay: they help copy the results out below
  if (env < P.nenv && !(P.mask && !P.mask[env])) {
    using L = typename MT::Layout;
    double* wb = P.image + (env / 32) * P.stride * 32;
    double* base = wb + (env & 31);
    int* ib = (int*)(wb + (long long)L::ints * 32) + (env & 31);
    Ctx<1, 32, 1, MT> c{P, base, ib, 0, 1u << (threadIdx.x & 31), true, false, nullptr, nullptr,
                        P.m.geom_pos, P.m.site_pos, P.m.body_pos, P.m.body_quat, P.m.wrap_prm};
    bp_load_ctrl(c, env);
    bp_run_mode(c);
    if (P.rew_out) bp_task_eval(P, P.task, env, P.task_nobs, P.obs_out, P.rew_out);
  }
  // results straight to pinned host memory (BpParams.obs_host): every warp copies the rows of
  // its 32 envs with lane-contiguous stores
  if (P.obs_host || P.rew_host) {
    __syncwarp();
    const long long e0 = env & ~31LL;
    long long cnt = P.nenv - e0 < 32 ? P.nenv - e0 : 32;
    const int ln = threadIdx.x & 31;
    if (P.obs_host && P.obs_out)
      for (long long i = ln; i < cnt * P.task_nobs; i += 32)
        P.obs_host[e0 * P.task_nobs + i] = __ldcg(P.obs_out + e0 * P.task_nobs + i);
    if (P.rew_host && ln < cnt) P.rew_host[e0 + ln] = __ldcg(P.rew_out + e0 + ln);
  }
}

template <class MT>
static int launch_static(const BpParams* P, cudaStream_t stream) {
  int threads = 32;                       // one warp per block: 512 warps over 592 schedulers
  if (const char* e = getenv("BPHYS_STATIC_THREADS")) {
    int v = atoi(e);
    if (v >= 32 && v <= BP_STATIC_TPB && v % 32 == 0) threads = v;
  }
  const unsigned blocks = (unsigned)((P->nenv + threads - 1) / threads);
  bp_step_kernel_static<MT><<<blocks, threads, 0, stream>>>(*P);
  return (int)cudaGetLastError();
}

// ---- instances at the production lane widths --------------------------------------------
// The generic global-memory kernel (bp_pipeline.cuh bp_step_kernel_gmem: G lanes per env,
// contiguous images, one lock-step block per SM) instantiated with static traits: sizes, tree,
// address tables and image offsets fold into the code (UF = 1: no extra unrolling -- measured,
// profiles/r2_tuning.md: the constants are worth 11 %, unrolling costs code size).
template <int G, int SHAPE, class MT>
static int launch_static_lanes(const BpParams* P, cudaStream_t stream) {
  const long long total = P->nenv * G;
  const int threads = bp_block_threads(total, BP_TPB(G));
  const unsigned blocks = (unsigned)((total + threads - 1) / threads);
  BpParams Q = *P;
  Q.hot_doubles = 0;
  bp_step_kernel_gmem<G, 1, SHAPE, MT><<<blocks, threads, 0, stream>>>(Q);
  return (int)cudaGetLastError();
}

// One translation unit per instance (they compile in parallel).  LANES = 1 selects the
// one-thread-per-env kernel above, anything else the lock-step kernel at that lane width.
#define BP_STATIC_INSTANCE(NAME, LANES, SHAPE)                                                 \
  extern "C" int bp_launch_static_##NAME##_g##LANES(const BpParams* P, unsigned long long hash, \
                                                    int plain, cudaStream_t stream) {           \
    using T = bp_static_##NAME::Traits;                                                        \
    /* a LEAN / MID instance only serves launches that qualify for that shape (bp_api.cu) */   \
    if ((SHAPE != 0 && !plain) || hash != T::HASH || P->lanes != LANES || P->L.njmax != T::njmax || \
        P->L.nconmax != T::nconmax || P->L.total != T::Layout::total)                          \
      return -1;                                                                               \
    if constexpr (LANES == 1) return launch_static<T>(P, stream);                              \
    else return launch_static_lanes<LANES, SHAPE, T>(P, stream);                               \
  }
#endif  // BP_STEP_STATIC_CUH_
