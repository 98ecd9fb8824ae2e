// vk_exchange.cu - the image bands of a block-sharded volume over peer memory (NVLink / NVSwitch), without
// a library collective.  No reference counterpart (the reference is single-GPU); interface in
// include/vulcan_b200.h (vk_trace_sharded), buffer layout and handshake in vk_shard.cuh.
#include "vk_math.cuh"
#include "vk_shard.cuh"
#include "vk_launch.h"
#include <cstring>

namespace vk {

// Image bands of a sharded volume with replicated lookup structures (vk_trace_sharded): every rank has
// rendered rows [bands[rank], bands[rank + 1]) into its own output images; it now STORES that band into
// the same place of every peer's output images (posted writes over NVLink, nothing waits on a load),
// and the last CTA to finish announces it and waits for the peers' announcements.  When the kernel
// returns, every rank's output images are complete.
__global__ void __launch_bounds__(256) vk_push_bands_kernel(const Xchg x, int w, int h) {
  pdl_wait();
  pdl_trigger();
  void* mine = x.buf[x.rank];
  const int* bands = xchg_bands(mine, x.parity);
  const int b0 = bands[x.rank], b1 = bands[x.rank + 1];
  const size_t plane = (size_t)w * h;
  const size_t row0 = (size_t)b0 * w, rows = (b1 > b0) ? (size_t)(b1 - b0) * w : 0;
  const float* src = xchg_images(mine, x.parity, w, h);
  // depth | colour | normal: three contiguous segments per band
  const size_t seg_off[3] = {row0, plane + 3 * row0, 4 * plane + 3 * row0};
  const size_t seg_n[3] = {rows, 3 * rows, 3 * rows};
  const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x, nthr = (size_t)gridDim.x * blockDim.x;
  for (int r = 0; r < x.count && xchg_stores(x.stages); ++r) {
    if (r == x.rank) continue;
    float* dst = xchg_images(x.buf[r], x.parity, w, h);
    for (int sgm = 0; sgm < 3; ++sgm) {
      const float* a = src + seg_off[sgm];
      float* b = dst + seg_off[sgm];
      const size_t n = seg_n[sgm];
      if ((((uintptr_t)a | (uintptr_t)b) & 15u) == 0) {
        const size_t n4 = n / 4;
        for (size_t i = tid; i < n4; i += nthr) reinterpret_cast<float4*>(b)[i] = __ldcg(reinterpret_cast<const float4*>(a) + i);
        for (size_t i = n4 * 4 + tid; i < n; i += nthr) b[i] = __ldcg(a + i);
      } else {
        for (size_t i = tid; i < n; i += nthr) b[i] = __ldcg(a + i);
      }
    }
  }
  if (x.stages == 4) return;   // store-only launch: the handshake follows in the next one (kernel boundary = fence)
  __shared__ bool last;
  if (!(x.stages & 8)) __threadfence_system();
  __syncthreads();
  int* ticket = xchg_ticket(mine) + 1;
  if (threadIdx.x == 0) last = (atomicAdd(ticket, 1) == (int)gridDim.x - 1);
  __syncthreads();
  if (!last) return;
  if (threadIdx.x == 0) *ticket = 0;
  xchg_handshake(x, true);
}

int exchange_push_bands(const Xchg& x, int w, int h, cudaStream_t s) {
  // (wait-only stage of ranks that share a GPU: one CTA, the peers have all announced by then)
  if (x.stages & 8) vk_push_bands_kernel<<<1, 256, 0, s>>>(x, w, h);   // plain launch: see trace_sharded_bounds
  else launch_chained(vk_push_bands_kernel, dim3(xchg_stores(x.stages) ? sm_count() * 2 : 1), dim3(256), 0, s, x, w, h);
  VK_COUNT_LAUNCH();
  return (int)cudaGetLastError();
}

}  // namespace vk
