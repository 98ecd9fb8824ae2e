// vk_exchange.cu - exchange of the rendered image bands of a block-sharded volume over peer
// memory (NVLink / NVSwitch), without a library collective.  No reference counterpart (the
// reference is single-GPU); interface in include/vulcan_b200.h.
//
// Every rank renders its band into its own exchange buffer.  One kernel per rank then
//   1. announces "my band of view `epoch` is complete" by storing epoch into its flag in every
//      peer's buffer (the band was written by earlier kernels of the same stream, so it is
//      complete; __threadfence_system orders it before the flag for observers on other GPUs),
//   2. waits until the flags of all peers in the LOCAL buffer have reached epoch,
//   3. copies each peer's band straight out of that peer's buffer into the caller's images.
// Step 1 of a rank never waits for anything and precedes its step 2, so the ranks cannot wait for
// each other in a cycle.  Latency: two NVLink round trips plus the copy at link bandwidth, against
// ~40 us per call for a small NCCL collective.
#include "vk_math.cuh"
#include "vk_shard.cuh"
#include "vk_launch.h"
#include <cstring>

namespace vk {

// buffer layout: [depth H*W][colour H*W*3][normal H*W*3] floats, 64 ints (band flags, one per
// rank), then VK_MAX_SHARDS ray-bounds grids (one slot per rank, 80 x 60 x 2 floats each) and 64
// more ints (bounds flags)
constexpr int kBoundsFloats = VK_BOUNDS_W * VK_BOUNDS_H * 2;
__host__ __device__ inline size_t exchange_floats(int w, int h) { return (size_t)w * h * 7; }
__host__ __device__ inline size_t exchange_bytes(int w, int h) {
  return exchange_floats(w, h) * sizeof(float) + 64 * sizeof(int) +
         (size_t)VK_MAX_SHARDS * kBoundsFloats * sizeof(float) + 64 * sizeof(int);
}

struct ExchangeParams {
  int count, rank, w, h, epoch;
  float* buffers[VK_MAX_SHARDS];
  int row_begin[VK_MAX_SHARDS], row_end[VK_MAX_SHARDS];
  float *depth, *color, *normal;
};

__device__ __forceinline__ int* exchange_flags(float* buffer, int w, int h) {
  return reinterpret_cast<int*>(buffer + exchange_floats(w, h));
}

__device__ __forceinline__ float* exchange_bounds_slot(float* buffer, int w, int h, int rank) {
  return reinterpret_cast<float*>(exchange_flags(buffer, w, h) + 64) + (size_t)rank * kBoundsFloats;
}
__device__ __forceinline__ int* exchange_bounds_flags(float* buffer, int w, int h) {
  return reinterpret_cast<int*>(exchange_bounds_slot(buffer, w, h, VK_MAX_SHARDS));
}

__global__ void __launch_bounds__(256) vk_exchange_kernel(const ExchangeParams p) {
  // 1. announce (one CTA does it; every CTA waits below)
  if (blockIdx.x == 0 && threadIdx.x < p.count && threadIdx.x != p.rank) {
    __threadfence_system();
    st_release_sys(exchange_flags(p.buffers[threadIdx.x], p.w, p.h) + p.rank, p.epoch);
  }
  // 2. wait for the peers
  if (threadIdx.x < p.count && threadIdx.x != p.rank) {
    const int* flag = exchange_flags(p.buffers[p.rank], p.w, p.h) + threadIdx.x;
    while (ld_acquire_sys(flag) < p.epoch) __nanosleep(100);
  }
  __syncthreads();
  // 3. copy the bands: rows are contiguous in all three planes
  const size_t plane = (size_t)p.w * p.h;
  for (int r = 0; r < p.count; ++r) {
    const size_t row0 = (size_t)p.row_begin[r] * p.w, rows = (size_t)(p.row_end[r] - p.row_begin[r]) * p.w;
    const float* src = p.buffers[r];
    // three segments of the band, in units of float4 where alignment allows (w * rows * 4 bytes;
    // image widths are multiples of 4 in every configuration, a scalar loop handles the rest)
    const float* seg_src[3] = {src + row0, src + plane + 3 * row0, src + 4 * plane + 3 * row0};
    float* seg_dst[3] = {p.depth + row0, p.color + 3 * row0, p.normal ? p.normal + 3 * row0 : nullptr};
    const size_t seg_n[3] = {rows, 3 * rows, 3 * rows};
    for (int s = 0; s < 3; ++s) {
      if (!seg_dst[s]) continue;
      const float* a = seg_src[s];
      float* b = seg_dst[s];
      const size_t n = seg_n[s];
      if (r == p.rank && a == b) continue;
      const bool vec = ((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b)) & 15u) == 0;
      if (vec) {
        const size_t n4 = n / 4;
        for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x)
          reinterpret_cast<float4*>(b)[i] = __ldcg(reinterpret_cast<const float4*>(a) + i);
        for (size_t i = n4 * 4 + (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
          b[i] = __ldcg(a + i);
      } else {
        for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
          b[i] = __ldcg(a + i);
      }
    }
  }
}

// Ray-bounds grids: every rank stores its grid into its slot of every rank's buffer, announces
// it, waits for the others and leaves the element-wise minimum of all grids in `bounds` (cells
// are (min, -max), so the minimum combines both).  One CTA: the grid is 38 KB.  Like the NCCL
// all-reduce it replaces, this is also the barrier that orders every rank's integration (earlier
// in its stream) before any rank's peer reads (later in theirs).
__global__ void __launch_bounds__(1024) vk_exchange_bounds_kernel(const ExchangeParams p, float* bounds) {
  for (int r = 0; r < p.count; ++r) {
    float* slot = exchange_bounds_slot(p.buffers[r], p.w, p.h, p.rank);
    for (int i = threadIdx.x; i < kBoundsFloats; i += blockDim.x) slot[i] = bounds[i];
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x < p.count && threadIdx.x != p.rank)
    st_release_sys(exchange_bounds_flags(p.buffers[threadIdx.x], p.w, p.h) + p.rank, p.epoch);
  if (threadIdx.x < p.count && threadIdx.x != p.rank) {
    const int* flag = exchange_bounds_flags(p.buffers[p.rank], p.w, p.h) + threadIdx.x;
    while (ld_acquire_sys(flag) < p.epoch) __nanosleep(100);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < kBoundsFloats; i += blockDim.x) {
    float m = bounds[i];
    for (int r = 0; r < p.count; ++r)
      if (r != p.rank) m = fminf(m, __ldcg(exchange_bounds_slot(p.buffers[p.rank], p.w, p.h, r) + i));
    bounds[i] = m;
  }
}

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
  for (int r = 0; r < x.count && (x.stages & 1); ++r) {
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
  __shared__ bool last;
  __threadfence_system();
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
  launch_chained(vk_push_bands_kernel, dim3((x.stages & 1) ? sm_count() * 2 : 1), dim3(256), 0, s, x, w, h);
  VK_COUNT_LAUNCH();
  return (int)cudaGetLastError();
}

}  // namespace vk

using namespace vk;

static int fill_exchange_params(ExchangeParams& p, const vkExchange* x, int epoch) {
  if (!x || x->count < 1 || x->count > VK_MAX_SHARDS || x->rank < 0 || x->rank >= x->count || x->width <= 0 ||
      x->height <= 0 || epoch <= 0)
    return VK_E_INVALID;
  p.count = x->count; p.rank = x->rank; p.w = x->width; p.h = x->height; p.epoch = epoch;
  for (int r = 0; r < VK_MAX_SHARDS; ++r) {
    p.buffers[r] = (r < x->count) ? static_cast<float*>(x->buffers[r]) : nullptr;
    p.row_begin[r] = (r < x->count) ? x->row_begin[r] : 0;
    p.row_end[r] = (r < x->count) ? x->row_end[r] : 0;
    if (r < x->count && (!p.buffers[r] || p.row_begin[r] < 0 || p.row_end[r] > x->height || p.row_begin[r] > p.row_end[r]))
      return VK_E_INVALID;
  }
  p.depth = p.color = p.normal = nullptr;
  return 0;
}

extern "C" int vk_exchange_bounds(const vkExchange* x, int epoch, float* bounds, void* stream) {
  ExchangeParams p;
  const int rc = fill_exchange_params(p, x, epoch);
  if (rc) return rc;
  if (!bounds) return VK_E_INVALID;
  vk_exchange_bounds_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(p, bounds);
  VK_COUNT_LAUNCH();
  return (int)cudaGetLastError();
}

extern "C" int vk_exchange_alloc(int width, int height, void** buffer) {
  if (width <= 0 || height <= 0 || !buffer) return VK_E_INVALID;
  void* p = nullptr;
  cudaError_t err = cudaMalloc(&p, exchange_bytes(width, height));
  if (err != cudaSuccess) return (int)err;
  err = cudaMemset(p, 0, exchange_bytes(width, height));
  if (err != cudaSuccess) { cudaFree(p); return (int)err; }
  *buffer = p;
  return 0;
}

extern "C" int vk_exchange_free(void* buffer) { return buffer ? (int)cudaFree(buffer) : 0; }

extern "C" int vk_exchange_export(const void* buffer, unsigned char handle[64]) {
  if (!buffer || !handle) return VK_E_INVALID;
  cudaIpcMemHandle_t h;
  cudaError_t err = cudaIpcGetMemHandle(&h, const_cast<void*>(buffer));
  if (err != cudaSuccess) return (int)err;
  memcpy(handle, &h, sizeof(h));
  return 0;
}

extern "C" int vk_exchange_open(const unsigned char handle[64], void** buffer) {
  if (!handle || !buffer) return VK_E_INVALID;
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, sizeof(h));
  return (int)cudaIpcOpenMemHandle(buffer, h, cudaIpcMemLazyEnablePeerAccess);
}

extern "C" int vk_exchange_close(void* buffer) { return buffer ? (int)cudaIpcCloseMemHandle(buffer) : 0; }

extern "C" float* vk_exchange_depth(void* buffer, int, int) { return static_cast<float*>(buffer); }
extern "C" float* vk_exchange_color(void* buffer, int width, int height) {
  return static_cast<float*>(buffer) + (size_t)width * height;
}
extern "C" float* vk_exchange_normal(void* buffer, int width, int height) {
  return static_cast<float*>(buffer) + (size_t)width * height * 4;
}

extern "C" int vk_exchange_bands(const vkExchange* x, int epoch, float* out_depth, float* out_color,
                                 float* out_normal, void* stream) {
  ExchangeParams p;
  const int rc = fill_exchange_params(p, x, epoch);
  if (rc) return rc;
  if (!out_depth || !out_color) return VK_E_INVALID;
  p.depth = out_depth; p.color = out_color; p.normal = out_normal;
  // every CTA spins on the flags, so all of them must be resident at once: one per SM
  vk_exchange_kernel<<<sm_count(), 256, 0, (cudaStream_t)stream>>>(p);
  VK_COUNT_LAUNCH();
  return (int)cudaGetLastError();
}
