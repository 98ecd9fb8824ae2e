// vk_shard.cuh - device-side view of a block-sharded volume's exchange buffers (no reference
// counterpart: the reference is single-GPU).  Interface: include/vulcan_b200.h, "sharding".
//
// Every rank owns one exchange buffer and maps the buffers of its peers (CUDA IPC over NVLink /
// NVSwitch, or plain pointers when the ranks share a process).  Layout of a buffer:
//
//   ints   [0, 16)    bounds_flag[2][8]   epoch of the last ray-bounds grid rank r stored here, per epoch parity
//   ints   [16, 32)   bands_flag[2][8]    epoch of the last image band rank r stored here, per epoch parity
//   ints   [32, 64)   bands[2][16]        row bands of the current view: rank r renders rows [bands[r], bands[r+1])
//   ints   [64, 256)  reserved (ticket at 64)
//   floats            bounds_slot[2][8][80*60*2]   ray-bounds grid of rank r, per parity
//   floats            grid[80*60*2]                 this rank's working grid
//   floats            images[2][7*W*H]              depth | colour | normal of a view, per parity
//
// Everything that is written by a peer is double-buffered by the parity of the view's epoch, so a rank
// that is one collective ahead of another never overwrites what the slower one still reads (a rank
// cannot get two collectives ahead: each collective waits for every rank's announcement).
#pragma once

#include "vk_math.cuh"

namespace vk {

constexpr int kGridFloats = VK_BOUNDS_W * VK_BOUNDS_H * 2;
constexpr int kXchgHeaderInts = 256;

__host__ __device__ inline size_t xchg_image_floats(int w, int h) { return (size_t)w * h * 7; }
__host__ __device__ inline size_t xchg_bytes(int w, int h) {
  return kXchgHeaderInts * sizeof(int) + ((size_t)2 * VK_MAX_SHARDS + 1) * kGridFloats * sizeof(float) +
         2 * xchg_image_floats(w, h) * sizeof(float);
}
__host__ __device__ inline int* xchg_bounds_flags(void* buf, int parity) { return static_cast<int*>(buf) + parity * 8; }
__host__ __device__ inline int* xchg_bands_flags(void* buf, int parity) { return static_cast<int*>(buf) + 16 + parity * 8; }
__host__ __device__ inline int* xchg_bands(void* buf, int parity) { return static_cast<int*>(buf) + 32 + parity * 16; }
__host__ __device__ inline int* xchg_ticket(void* buf) { return static_cast<int*>(buf) + 64; }
__host__ __device__ inline int* xchg_timeouts(void* buf) { return static_cast<int*>(buf) + 68; }  // handshakes given up
__host__ __device__ inline float* xchg_floats(void* buf) { return reinterpret_cast<float*>(static_cast<int*>(buf) + kXchgHeaderInts); }
__host__ __device__ inline float* xchg_bounds_slot(void* buf, int parity, int rank) {
  return xchg_floats(buf) + ((size_t)parity * VK_MAX_SHARDS + rank) * kGridFloats;
}
__host__ __device__ inline float* xchg_grid(void* buf) { return xchg_floats(buf) + (size_t)2 * VK_MAX_SHARDS * kGridFloats; }
__host__ __device__ inline float* xchg_images(void* buf, int parity, int w, int h) {
  return xchg_grid(buf) + kGridFloats + (size_t)parity * xchg_image_floats(w, h);
}
__host__ __device__ inline float* xchg_depth(void* buf, int parity, int w, int h) { return xchg_images(buf, parity, w, h); }
__host__ __device__ inline float* xchg_color(void* buf, int parity, int w, int h) { return xchg_images(buf, parity, w, h) + (size_t)w * h; }
__host__ __device__ inline float* xchg_normal(void* buf, int parity, int w, int h) { return xchg_images(buf, parity, w, h) + (size_t)w * h * 4; }

// kernel-parameter view of the ranks' buffers for one collective
struct Xchg {
  int count, rank, epoch, parity;
  int stages;   // bit 0: store + announce, bit 1: wait + consume.  Ranks that SHARE a GPU run the two halves as
                // separate launches, all announcements first (vk_trace_sharded_staged): kernels of one device must
                // never wait for one another.  One rank per GPU (vk_trace_sharded) uses two launches as well, for
                // another reason: bit 2 (alone) = store only, then a second launch with bits 0 | 1 | 3 =
                // announce + wait + consume, where bit 3 says "what is announced was stored by an EARLIER launch":
                // the kernel boundary orders those stores before the flags, so no system-scope fence is needed.
                // (A system-scope fence waits for the GPU's outstanding PCIe traffic too: with a download of the
                // previous frame in flight it cost the streamed pipeline 30-70 us per view.)
  void* buf[VK_MAX_SHARDS];
};

#ifdef __CUDACC__
__device__ __forceinline__ int ld_acquire_sys(const int* p) {
  int v;
  asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(int* p, int v) {
  asm volatile("st.release.sys.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void st_relaxed_sys(int* p, int v) {
  asm volatile("st.relaxed.sys.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
// what a launch with these stage bits stores into the peers itself
__host__ __device__ inline bool xchg_stores(int stages) { return (stages & 4) || ((stages & 1) && !(stages & 8)); }

// One CTA: "what this rank stored into its peers' buffers for `epoch` is complete" -> every peer, then
// wait until every peer has said the same to this rank.  The stores being announced were issued by
// this CTA before the call or by earlier work of the stream; the system-scope fence orders them before
// the flags for observers on other GPUs.  Announcing never waits and precedes the wait, so ranks
// cannot wait for each other in a cycle.
__device__ __forceinline__ void xchg_handshake(const Xchg& x, bool bands) {
  const bool ordered = (x.stages & 8) != 0;   // the announced stores belong to an earlier launch
  if (!ordered) {
    if (x.count > 1) __threadfence_system();   // (a lone rank has nobody to publish to)
    else __threadfence();
  }
  __syncthreads();
  const int t = threadIdx.x;
  if ((x.stages & 1) && t < x.count && t != x.rank) {
    int* theirs = (bands ? xchg_bands_flags(x.buf[t], x.parity) : xchg_bounds_flags(x.buf[t], x.parity)) + x.rank;
    if (ordered) st_relaxed_sys(theirs, x.epoch);
    else st_release_sys(theirs, x.epoch);
  }
  if ((x.stages & 2) && t < x.count && t != x.rank) {
    const int* mine = (bands ? xchg_bands_flags(x.buf[x.rank], x.parity) : xchg_bounds_flags(x.buf[x.rank], x.parity)) + t;
    // a peer that never arrives (crashed process, mismatched call sequence) must not hang the GPU: after
    // ~2 s the wait is given up and counted; vk_shard_band / vk_shard_status report VK_E_TIMEOUT
    long long spins = 0;
    while (ld_acquire_sys(mine) < x.epoch) {
      __nanosleep(100);
      if (++spins > (1ll << 24)) {
        atomicAdd(xchg_timeouts(x.buf[x.rank]), 1);
        break;
      }
    }
  }
  __syncthreads();
}
#endif

}  // namespace vk
