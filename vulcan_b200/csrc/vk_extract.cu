// vk_extract.cu — surface points of the fused volume (first slice).
//
// The reference's extractor is unfinished: its cube-state / vertex-edge kernels stop at edge midpoints of
// one block (`points[index] = 0.5f * ap + 0.5f * bp`, src/extractor.cu:386), the index and face kernels
// are empty (src/extractor.cu:391-413) and the app's --output flag is never used (apps/vulcan/vulcan.cu:20).
// There is therefore no reference output to be identical to; this file implements the part its design
// fixes - the marching-cubes VERTICES, i.e. the zero crossings of the TSDF on the three lattice edges
// every voxel owns (edges 0, 3 and 8 of its cube, the `edge_indices` of src/extractor.cu:52-117) - over ALL
// allocated blocks (the reference's own TODO, src/extractor.cu:490,673) and across block faces, with
// linear interpolation of position and colour.  Parity is pinned to the CPU oracle only
// (oracle/vulcan_oracle.cpp: vo_extract_points), bit for bit as sorted sets ("parity unpinned" with
// respect to the reference).
//
// One CTA takes 16 hash entries at a time (grid-stride: small chunks balance the CTAs), lists the allocated ones in shared memory, and for each
// stages the block's 512 voxels (10 KB, coalesced 4-byte words) in shared memory; thread t owns voxel t.
// Neighbours inside the block come from the tile, neighbours across a face from a hash lookup of the
// adjacent block.  Points are appended with one atomicAdd per block (CTA-wide scan of the per-voxel
// counts); the order of blocks in the output is arbitrary.
#include <cuda_runtime.h>
#include "vk_launch.h"
#include "vk_math.cuh"

namespace vk {

struct ExtractParams {
  const vkHashEntry* entries;
  const vkVoxel* voxels;
  int total_entries;
  Mod K;
  float voxel_length;
  float block_length;
  float* points;
  float* colors;
  int capacity;
  int* count;
  int chunk;  // hash entries a CTA lists at a time (<= kExtractChunkMax)
};

__device__ __forceinline__ int extract_lookup(const ExtractParams& p, int bx, int by, int bz) {
  if (bx > 32767 || by > 32767 || bz > 32767) return -1;
  const unsigned long long key = pack_block(bx, by, bz);
  int4 e = ld_entry(p.entries + block_hash(bx, by, bz, p.K));
  while (entry_block(e) != key && e.w != -1) e = ld_entry(p.entries + e.w);
  return (entry_block(e) == key) ? e.z : -1;
}

struct Sample { float d, r, g, b; int w; };

constexpr int kExtractChunkMax = 128;

__global__ void __launch_bounds__(512, 3) vk_extract_points_kernel(ExtractParams p) {
  __shared__ uint32_t tile[VK_BLOCK_VOXELS * 5];
  __shared__ int list[kExtractChunkMax];
  __shared__ int list_count;
  __shared__ int warp_totals[16];
  __shared__ int block_base;
  const int tid = threadIdx.x;
  const int vx = tid & 7, vy = (tid >> 3) & 7, vz = tid >> 6;
  const int chunks = (p.total_entries + p.chunk - 1) / p.chunk;
  for (int chunk = blockIdx.x; chunk < chunks; chunk += gridDim.x) {
    if (tid == 0) list_count = 0;
    __syncthreads();
    const int ei = chunk * p.chunk + tid;
    if (tid < p.chunk && ei < p.total_entries && __ldg(&p.entries[ei].data) != -1) list[atomicAdd(&list_count, 1)] = ei;
    __syncthreads();
    const int n = list_count;
    // the tile of block li + 1 is fetched into registers (five independent 4-byte words per thread) while
    // block li is processed, so the DRAM latency of a block hides behind its predecessor's work
    uint32_t r[5];
    int4 e_next = make_int4(0, 0, -1, -1);
    if (n > 0) {
      e_next = ld_entry(p.entries + list[0]);
      const uint32_t* src = reinterpret_cast<const uint32_t*>(p.voxels + (size_t)e_next.z * VK_BLOCK_VOXELS);
#pragma unroll
      for (int k = 0; k < 5; ++k) r[k] = __ldg(src + tid + 512 * k);
    }
    for (int li = 0; li < n; ++li) {
      const int4 e = e_next;
      int bx, by, bz;
      entry_origin(e, bx, by, bz);
      __syncthreads();  // the previous block's tile is no longer read
#pragma unroll
      for (int k = 0; k < 5; ++k) tile[tid + 512 * k] = r[k];
      __syncthreads();
      if (li + 1 < n) {
        e_next = ld_entry(p.entries + list[li + 1]);
        const uint32_t* src = reinterpret_cast<const uint32_t*>(p.voxels + (size_t)e_next.z * VK_BLOCK_VOXELS);
#pragma unroll
        for (int k = 0; k < 5; ++k) r[k] = __ldg(src + tid + 512 * k);
      }
      const float d0 = __uint_as_float(tile[5 * tid]);
      const int w0 = (int)(int16_t)(tile[5 * tid + 4] & 0xffffu);
      const float c0[3] = {__uint_as_float(tile[5 * tid + 1]), __uint_as_float(tile[5 * tid + 2]),
                           __uint_as_float(tile[5 * tid + 3])};
      // voxel centre: L * b + l * (v + 0.5)   (src/depth_integrator.cu:44-52)
      const float base[3] = {add(mul((float)bx, p.block_length), mul(add((float)vx, 0.5f), p.voxel_length)),
                             add(mul((float)by, p.block_length), mul(add((float)vy, 0.5f), p.voxel_length)),
                             add(mul((float)bz, p.block_length), mul(add((float)vz, 0.5f), p.voxel_length))};
      float tpar[3];      // interpolation parameter per axis
      float ncol[3][3];   // neighbour colour per axis
      unsigned mask = 0;  // axes with a crossing
      if (w0 > 0) {
#pragma unroll
        for (int a = 0; a < 3; ++a) {
          const int nx = vx + (a == 0), ny = vy + (a == 1), nz = vz + (a == 2);
          Sample s;
          if (nx < 8 && ny < 8 && nz < 8) {
            const int j = (nz << 6) | (ny << 3) | nx;
            s.d = __uint_as_float(tile[5 * j]);
            s.r = __uint_as_float(tile[5 * j + 1]);
            s.g = __uint_as_float(tile[5 * j + 2]);
            s.b = __uint_as_float(tile[5 * j + 3]);
            s.w = (int)(int16_t)(tile[5 * j + 4] & 0xffffu);
          } else {
            const int data = extract_lookup(p, bx + (nx >> 3), by + (ny >> 3), bz + (nz >> 3));
            if (data == -1) continue;
            const int j = ((nz & 7) << 6) | ((ny & 7) << 3) | (nx & 7);
            const vkVoxel* q = p.voxels + (size_t)data * VK_BLOCK_VOXELS + j;
            s.d = __ldg(&q->distance);
            s.r = __ldg(&q->color[0]);
            s.g = __ldg(&q->color[1]);
            s.b = __ldg(&q->color[2]);
            s.w = (int)__ldg(&q->distance_weight);
          }
          if (s.w <= 0 || ((d0 > 0) == (s.d > 0))) continue;
          tpar[a] = fdiv(d0, sub(d0, s.d));
          ncol[a][0] = s.r; ncol[a][1] = s.g; ncol[a][2] = s.b;
          mask |= 1u << a;
        }
      }
      // one atomicAdd per block: exclusive scan of the per-thread counts over the CTA
      const int mine = __popc(mask);
      int incl = mine;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int up = __shfl_up_sync(0xffffffffu, incl, o);
        if ((tid & 31) >= o) incl += up;
      }
      if ((tid & 31) == 31) warp_totals[tid >> 5] = incl;
      __syncthreads();
      if (tid < 32) {
        int wt = tid < 16 ? warp_totals[tid] : 0;
        int wi = wt;
#pragma unroll
        for (int o = 1; o < 16; o <<= 1) {
          const int up = __shfl_up_sync(0xffffffffu, wi, o);
          if (tid >= o) wi += up;
        }
        if (tid < 16) warp_totals[tid] = wi - wt;  // exclusive offset of the warp
        if (tid == 15) block_base = wi ? atomicAdd(p.count, wi) : 0;
      }
      __syncthreads();
      int at = block_base + warp_totals[tid >> 5] + incl - mine;
#pragma unroll
      for (int a = 0; a < 3; ++a) {
        if (!((mask >> a) & 1u)) continue;
        if (at < p.capacity) {
          const size_t o = 3 * (size_t)at;
          p.points[o] = a == 0 ? add(base[0], mul(tpar[0], p.voxel_length)) : base[0];
          p.points[o + 1] = a == 1 ? add(base[1], mul(tpar[1], p.voxel_length)) : base[1];
          p.points[o + 2] = a == 2 ? add(base[2], mul(tpar[2], p.voxel_length)) : base[2];
          if (p.colors) {
            p.colors[o] = add(c0[0], mul(tpar[a], sub(ncol[a][0], c0[0])));
            p.colors[o + 1] = add(c0[1], mul(tpar[a], sub(ncol[a][1], c0[1])));
            p.colors[o + 2] = add(c0[2], mul(tpar[a], sub(ncol[a][2], c0[2])));
          }
        }
        ++at;
      }
    }
    __syncthreads();
  }
}

}  // namespace vk

using namespace vk;

extern "C" int vk_extract_points(const vkVolume* v, float* points, float* colors, int capacity, int* count,
                                 void* stream) {
  if (!v || !v->voxels || !points || !count || capacity < 0) return VK_E_INVALID;
  // A shard resolves face neighbours in its own table only: crossings into another rank's blocks would be
  // dropped silently.  Extract from an unsharded volume (or gather the shards first).
  if (v->shard_count > 1) return VK_E_INVALID;
  cudaStream_t s = (cudaStream_t)stream;
  int rc = (int)cudaMemsetAsync(count, 0, sizeof(int), s);
  if (rc) return rc;
  ExtractParams p;
  p.entries = (const vkHashEntry*)v->entries;
  p.voxels = (const vkVoxel*)v->voxels;
  p.total_entries = v->main_blocks + v->excess_blocks;
  p.K = make_mod((uint32_t)v->main_blocks);
  p.voxel_length = v->voxel_length;
  p.block_length = VK_BLOCK_RES * v->voxel_length;
  p.points = points;
  p.colors = colors;
  p.capacity = capacity;
  p.count = count;
  p.chunk = 16;  // measured on config 1 (33.6 k blocks): 128 -> 0.77 ms, 64 -> 0.58, 32 -> 0.49, 16 -> 0.47, 8 -> 0.50
  const int chunks = (p.total_entries + p.chunk - 1) / p.chunk;
  const int cap = 3 * sm_count();  // three 512-thread CTAs per SM are resident
  vk_extract_points_kernel<<<chunks < cap ? chunks : cap, 512, 0, s>>>(p);
  VK_COUNT_LAUNCH();
  return (int)cudaGetLastError();
}
