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
// One CTA takes 512 hash entries at a time, lists the allocated ones in shared memory, and for each
// stages the block's 512 voxels (10 KB, coalesced 4-byte words) in shared memory; thread t owns voxel t.
// Neighbours inside the block come from the tile, neighbours across a face from a hash lookup of the
// adjacent block.  Points are appended with one atomicAdd per emitting thread; order is arbitrary.
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
};

__device__ __forceinline__ int extract_lookup(const ExtractParams& p, int bx, int by, int bz) {
  if (bx > 32767 || by > 32767 || bz > 32767) return -1;
  const unsigned long long key = pack_block(bx, by, bz);
  int4 e = ld_entry(p.entries + block_hash(bx, by, bz, p.K));
  while (entry_block(e) != key && e.w != -1) e = ld_entry(p.entries + e.w);
  return (entry_block(e) == key) ? e.z : -1;
}

struct Sample { float d, r, g, b; int w; };

__global__ void __launch_bounds__(512) vk_extract_points_kernel(ExtractParams p) {
  __shared__ uint32_t tile[VK_BLOCK_VOXELS * 5];
  __shared__ int list[512];
  __shared__ int list_count;
  const int tid = threadIdx.x;
  const int vx = tid & 7, vy = (tid >> 3) & 7, vz = tid >> 6;
  const int chunks = (p.total_entries + 511) / 512;
  for (int chunk = blockIdx.x; chunk < chunks; chunk += gridDim.x) {
    if (tid == 0) list_count = 0;
    __syncthreads();
    const int ei = chunk * 512 + tid;
    if (ei < p.total_entries && __ldg(&p.entries[ei].data) != -1) list[atomicAdd(&list_count, 1)] = ei;
    __syncthreads();
    const int n = list_count;
    for (int li = 0; li < n; ++li) {
      const int4 e = ld_entry(p.entries + list[li]);
      int bx, by, bz;
      entry_origin(e, bx, by, bz);
      const uint32_t* src = reinterpret_cast<const uint32_t*>(p.voxels + (size_t)e.z * VK_BLOCK_VOXELS);
      __syncthreads();  // the previous block's tile is no longer read
      for (int i = tid; i < VK_BLOCK_VOXELS * 5; i += 512) tile[i] = __ldg(src + i);
      __syncthreads();
      const float d0 = __uint_as_float(tile[5 * tid]);
      const int w0 = (int)(int16_t)(tile[5 * tid + 4] & 0xffffu);
      if (w0 <= 0) continue;
      const float c0[3] = {__uint_as_float(tile[5 * tid + 1]), __uint_as_float(tile[5 * tid + 2]),
                           __uint_as_float(tile[5 * tid + 3])};
      // voxel centre: L * b + l * (v + 0.5)   (src/depth_integrator.cu:44-52)
      const float base[3] = {add(mul((float)bx, p.block_length), mul(add((float)vx, 0.5f), p.voxel_length)),
                             add(mul((float)by, p.block_length), mul(add((float)vy, 0.5f), p.voxel_length)),
                             add(mul((float)bz, p.block_length), mul(add((float)vz, 0.5f), p.voxel_length))};
      float out[3][6];
      int emitted = 0;
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
        const float t = fdiv(d0, sub(d0, s.d));
        float* o = out[emitted++];
        o[0] = base[0]; o[1] = base[1]; o[2] = base[2];
        o[a] = add(base[a], mul(t, p.voxel_length));
        o[3] = add(c0[0], mul(t, sub(s.r, c0[0])));
        o[4] = add(c0[1], mul(t, sub(s.g, c0[1])));
        o[5] = add(c0[2], mul(t, sub(s.b, c0[2])));
      }
      if (emitted) {
        const int at = atomicAdd(p.count, emitted);
        for (int k = 0; k < emitted; ++k) {
          if (at + k >= p.capacity) break;
          const size_t o = 3 * (size_t)(at + k);
          p.points[o] = out[k][0]; p.points[o + 1] = out[k][1]; p.points[o + 2] = out[k][2];
          if (p.colors) { p.colors[o] = out[k][3]; p.colors[o + 1] = out[k][4]; p.colors[o + 2] = out[k][5]; }
        }
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
  const int chunks = (p.total_entries + 511) / 512;
  const int cap = 4 * sm_count();
  vk_extract_points_kernel<<<chunks < cap ? chunks : cap, 512, 0, s>>>(p);
  VK_COUNT_LAUNCH();
  return (int)cudaGetLastError();
}
