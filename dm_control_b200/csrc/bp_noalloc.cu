// bp_rewards_tolerance_noalloc: rewards.tolerance on the device for callers that may not
// allocate (CUDA graph capture around their own code, tight serving loops, harnesses that
// forbid cudaMalloc).  Same arithmetic as bp_rewards_tolerance (bp_api.cu / bp_task.cu),
// but the device staging of x and out -- which that entry point cudaMallocs and frees per
// call -- lives in ONE file-scope device array of a fixed, modest size.  No counterpart in
// the reference (include/bphys.h).
//
// Rules of the array:
//   * every GPU that loads libbphys.so holds BP_NOALLOC_MAX * 16 bytes (1 MiB) until exit;
//   * in host code `bp_noalloc_scratch` is only a stand-in for the symbol, NOT a device
//     address: the address is taken once per device with cudaGetSymbolAddress (that device
//     current), checked, cached, and passed to the kernel as an ordinary pointer;
//   * every call shares it and it keeps the previous call's contents.  Nothing has to
//     start at zero here: the first n doubles (x) are overwritten by the H2D copy and the
//     next n (out, at BP_NOALLOC_MAX) by the kernel before anything reads them;
//   * calls that overlap on the GPU (two host threads, two processes are fine -- each
//     process has its own copy -- but two threads of one process on one device are not)
//     race on it.  The entry point serialises its own callers with a mutex and runs on the
//     legacy default stream with synchronous copies, so a call has left the array by the
//     time it returns.
#include <cuda_runtime.h>
#include <stdint.h>

#include <mutex>
#include <string>

#include "../../include/bphys.h"
#include "bp_task.cuh"

#define BP_NOALLOC_MAX 65536          /* elements per call; declared in include/bphys.h */
#define BP_NOALLOC_MAXDEV 64

static __device__ __align__(256) double bp_noalloc_scratch[2 * BP_NOALLOC_MAX];

extern "C" void bp_set_last_error_(const char* msg);    // bp_api.cu (thread-local message)

static int na_fail(int code, const std::string& msg) {
  bp_set_last_error_(msg.c_str());
  return code;
}

static std::mutex g_na_mutex;
static double* g_na_ptr[BP_NOALLOC_MAXDEV];             // device address of the array, per device

__global__ void bp_tolerance_noalloc_kernel(const double* __restrict__ x, double* __restrict__ out,
                                            int n, double lower, double upper, double margin,
                                            int sig, double v1) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = tolerance(x[i], lower, upper, margin, sig, v1);
}

extern "C" int bp_rewards_tolerance_noalloc_max(void) { return BP_NOALLOC_MAX; }

extern "C" int bp_rewards_tolerance_noalloc(int device, const double* x_host, double* out_host,
                                            int64_t n, double lower, double upper, double margin,
                                            int sigmoid, double value_at_margin) {
  if (!x_host || !out_host || n < 0) return na_fail(BP_EINVAL, "bad arguments");
  if (lower > upper) return na_fail(BP_EINVAL, "Lower bound must be <= upper bound.");
  if (margin < 0) return na_fail(BP_EINVAL, "`margin` must be non-negative.");
  if (sigmoid < 0 || sigmoid > 3) return na_fail(BP_EINVAL, "Unknown sigmoid type");
  if (n > BP_NOALLOC_MAX)
    return na_fail(BP_EINVAL, "bp_rewards_tolerance_noalloc: n = " + std::to_string(n) +
                   " exceeds the fixed scratch of " + std::to_string(BP_NOALLOC_MAX) +
                   " elements; split the call or use bp_rewards_tolerance");
  if (device < 0 || device >= BP_NOALLOC_MAXDEV)
    return na_fail(BP_EINVAL, "bp_rewards_tolerance_noalloc: device index out of range");
  if (n == 0) return BP_OK;
  std::lock_guard<std::mutex> guard(g_na_mutex);
  cudaError_t e = cudaSetDevice(device);
  if (e != cudaSuccess)
    return na_fail(BP_ECUDA, std::string("cudaSetDevice: ") + cudaGetErrorString(e));
  if (!g_na_ptr[device]) {            // set-up, once per device, with that device current
    double* p = nullptr;
    e = cudaGetSymbolAddress((void**)&p, bp_noalloc_scratch);
    if (e != cudaSuccess || !p)
      return na_fail(BP_ECUDA, std::string("cudaGetSymbolAddress(bp_noalloc_scratch): ") +
                     cudaGetErrorString(e));
    g_na_ptr[device] = p;
  }
  double* dx = g_na_ptr[device];
  double* dout = dx + BP_NOALLOC_MAX;
  e = cudaMemcpy(dx, x_host, (size_t)n * 8, cudaMemcpyHostToDevice);
  if (e != cudaSuccess)
    return na_fail(BP_ECUDA, std::string("H2D copy: ") + cudaGetErrorString(e));
  const int threads = 256, blocks = (int)((n + threads - 1) / threads);
  bp_tolerance_noalloc_kernel<<<blocks, threads, 0, 0>>>(dx, dout, (int)n, lower, upper, margin,
                                                        sigmoid, value_at_margin);
  e = cudaGetLastError();
  if (e != cudaSuccess)
    return na_fail(BP_ECUDA, std::string("tolerance launch: ") + cudaGetErrorString(e));
  e = cudaMemcpy(out_host, dout, (size_t)n * 8, cudaMemcpyDeviceToHost);
  if (e != cudaSuccess)
    return na_fail(BP_ECUDA, std::string("D2H copy: ") + cudaGetErrorString(e));
  return BP_OK;
}
