// vk_launch.h — host-side helpers shared by the C-ABI translation units.
#pragma once

#include <atomic>
#include <cstring>
#include <utility>
#include <cuda_runtime.h>
#include "vk_math.cuh"

namespace vk {
// number of kernels this library has launched since load (vk_launch_count)
extern std::atomic<unsigned long long> g_launches;
// process-wide default of vkVolume::math_mode for entry points without a volume argument
extern std::atomic<int> g_default_math_mode;

// Tuning knobs (vk_tune): launch shapes and batching thresholds, never results.
enum { VK_TUNE_RAY_CTA_THREADS = 0, VK_TUNE_RAY_BATCH_LANES = 1, VK_TUNE_RAY_BATCH_WAIT = 2,
       VK_TUNE_INTEGRATE_CTAS_PER_SM = 3, VK_TUNE_ALLOC_MODE = 4, VK_TUNE_RAY_COOP_LANES = 5, VK_TUNE_DEBUG_X = 6, VK_TUNE_RAY_FREE_RUN = 7,
       VK_TUNE_RAY_COOP_ROUNDS = 8, VK_TUNE_PDL = 9, VK_TUNE_RAY_SOA = 10, VK_TUNE_RAY_PAIR_AFTER = 11, VK_TUNE_RAY_WARM_SDF = 12, VK_TUNE_INTEGRATE_BULK = 14, VK_TUNE_TRACE_FORK = 15, VK_TUNE_COUNT = 16 };
extern std::atomic<int> g_tune[VK_TUNE_COUNT];

// rows 0..2 of a column-major 4x4 (vulcan::Matrix4f) as row-major 3x4
inline void rows34(const float* colmajor, float* out12) {
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 4; ++c) out12[r * 4 + c] = colmajor[c * 4 + r];
}

// Divisor descriptor for vk::divc.  The first use of a divisor value in the process runs the
// device check (every mantissa of the dividend against __fdiv_rn, ~20 us; see vk_runtime.cu for why
// one binade is a complete proof) on `stream` and waits for it; later uses hit a host-side cache.  Returns fast = 0 (plain IEEE division) if the
// check fails or cannot run.
Div make_div(float d, cudaStream_t stream);

// number of SMs of the current device (cached per device)
int sm_count();

// Stream capture and the side stream of vk_trace do not mix (the fork would have to be part of the captured graph,
// and a graph replay is invisible to the events that guard the bounds scratch): once a capturing stream has been
// seen, the process stays on the plain in-stream path.
extern std::atomic<int> g_capture_seen;
inline bool stream_capturing(cudaStream_t s) {
  cudaStreamCaptureStatus st = cudaStreamCaptureStatusNone;
  if (cudaStreamIsCapturing(s, &st) != cudaSuccess) { cudaGetLastError(); g_capture_seen.store(1); return true; }
  if (st != cudaStreamCaptureStatusNone) g_capture_seen.store(1);
  return g_capture_seen.load() != 0;
}

// Programmatic dependent launch (sm_90+): the kernels of a frame form a chain on one stream; a
// kernel launched through launch_chained may be scheduled while the tail of its predecessor is
// still running, and blocks in pdl_wait() until the predecessor has completed and its writes are
// visible.  Hides the launch latency between the seven kernels of a step and lets a kernel do
// work that does not depend on its predecessor first.  Every kernel launched this way calls
// pdl_wait() in all of its threads before touching anything an earlier kernel wrote.
template <typename... KArgs, typename... Args>
inline cudaError_t launch_chained(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s,
                                  Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = g_tune[VK_TUNE_PDL].load() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, std::forward<Args>(args)...);
}
// Launch of a persistent kernel whose CTAs wait for one another (a grid-wide arrival counter): a
// COOPERATIVE launch, which the runtime only starts when the whole grid can be resident at once, also
// next to other work on the device.  Chained like launch_chained when the runtime accepts both
// attributes together, plain cooperative otherwise.
template <typename... KArgs, typename... Args>
inline cudaError_t launch_cooperative(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s,
                                      Args&&... args) {
  static std::atomic<int> chain_ok{1};
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeCooperative;
  attr[0].val.cooperative = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  if (chain_ok.load() && g_tune[VK_TUNE_PDL].load()) {
    cfg.numAttrs = 2;
    const cudaError_t err = cudaLaunchKernelEx(&cfg, kernel, std::forward<Args>(args)...);
    if (err == cudaSuccess) return err;
    cudaGetLastError();
    chain_ok.store(0);   // this runtime does not combine the two: cooperative only from now on
  }
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kernel, args...);
}
}  // namespace vk

#ifdef __CUDACC__
namespace vk {
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;"); }
}  // namespace vk
#endif

#define VK_COUNT_LAUNCH() (::vk::g_launches.fetch_add(1, std::memory_order_relaxed))
