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
#include "vk_mc_tables.h"

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
  // mesh extraction (vk_extract_mesh): every block gets a slot, every lattice edge of the block the index of
  // its vertex (or -1) at edge_index[slot * 1536 + voxel * 3 + axis]; all NULL for points only
  int* block_slot;     // [M + E] slot of voxel block `data`
  int* slot_counter;
  int* edge_index;
  int* faces;          // faces kernel: vertex index triples
  int face_capacity;
  int* face_count;
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
  __shared__ int block_slot_s;
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
        if (tid == 0 && p.edge_index) {
          block_slot_s = atomicAdd(p.slot_counter, 1);
          p.block_slot[e.z] = block_slot_s;
        }
      }
      __syncthreads();
      int at = block_base + warp_totals[tid >> 5] + incl - mine;
      if (p.edge_index) {   // vertex index of each of the voxel's three lattice edges, for the faces
        int* row = p.edge_index + ((size_t)block_slot_s * VK_BLOCK_VOXELS + tid) * 3;
        int k = at;
#pragma unroll
        for (int a = 0; a < 3; ++a) row[a] = ((mask >> a) & 1u) ? k++ : -1;
      }
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

// ---- triangles (what the reference's empty ExtractVertexIndicesKernel / ExtractFacesKernel,
// src/extractor.cu:392-413, leave open) ----------------------------------------------------------------------
// One marching-cubes cube per voxel: the cube spanned by the voxel and its +x / +y / +z neighbours (corner c at
// offset (c & 1, (c >> 1) & 1, c >> 2)), across block faces.  A cube is meshed when all eight corners are
// observed (distance_weight > 0); its case index has bit c set where !(distance > 0) - the sign test of the
// vertices - and kMcTriTable (generated, tools/gen_mc_tables.py) lists its triangles as cube-edge triples.
// Every cube edge is the lattice edge some voxel owns, so a triangle corner is just the vertex index
// vk_extract_points_kernel left in edge_index for (owner voxel, axis): vertices are shared between cubes and
// blocks by construction, and the mesh is watertight wherever the surface is surrounded by observed voxels.
__constant__ signed char c_tri_table[256][16];
__constant__ signed char c_edge_corner[12];
__constant__ signed char c_edge_axis[12];

__global__ void __launch_bounds__(512, 3) vk_extract_faces_kernel(ExtractParams p) {
  __shared__ float s_dist[VK_BLOCK_VOXELS];
  __shared__ short s_weight[VK_BLOCK_VOXELS];
  __shared__ int list[kExtractChunkMax];
  __shared__ int list_count;
  __shared__ int nb_data[8], nb_slot[8];   // the block (index 0) and its seven +x / +y / +z neighbours
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
    for (int li = 0; li < n; ++li) {
      const int4 e = ld_entry(p.entries + list[li]);
      int bx, by, bz;
      entry_origin(e, bx, by, bz);
      __syncthreads();  // the previous block's tile is no longer read
      {
        const uint32_t* src = reinterpret_cast<const uint32_t*>(p.voxels + (size_t)e.z * VK_BLOCK_VOXELS) + 5 * tid;
        s_dist[tid] = __uint_as_float(__ldg(src));
        s_weight[tid] = (short)(__ldg(src + 4) & 0xffffu);
      }
      if (tid < 8) {
        const int d = (tid == 0) ? e.z : extract_lookup(p, bx + (tid & 1), by + ((tid >> 1) & 1), bz + (tid >> 2));
        nb_data[tid] = d;
        nb_slot[tid] = (d >= 0) ? p.block_slot[d] : -1;
      }
      __syncthreads();
      // the eight corners
      unsigned cube = 0;
      bool ok = true;
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        const int x = vx + (c & 1), y = vy + ((c >> 1) & 1), z = vz + (c >> 2);
        float d;
        int w;
        if (x < 8 && y < 8 && z < 8) {
          const int j = (z << 6) | (y << 3) | x;
          d = s_dist[j];
          w = s_weight[j];
        } else {
          const int nb = nb_data[(x >> 3) | ((y >> 3) << 1) | ((z >> 3) << 2)];
          if (nb < 0) { ok = false; d = 1.f; w = 0; }
          else {
            const vkVoxel* q = p.voxels + (size_t)nb * VK_BLOCK_VOXELS + (((z & 7) << 6) | ((y & 7) << 3) | (x & 7));
            d = __ldg(&q->distance);
            w = (int)__ldg(&q->distance_weight);
          }
        }
        if (w <= 0) ok = false;
        if (!(d > 0)) cube |= 1u << c;
      }
      if (!ok) cube = 0;
      int mine = 0;
      while (mine < 5 && c_tri_table[cube][3 * mine] >= 0) ++mine;
      // one atomicAdd per block: exclusive scan of the per-thread triangle counts over the CTA
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
        if (tid < 16) warp_totals[tid] = wi - wt;
        if (tid == 15) block_base = wi ? atomicAdd(p.face_count, wi) : 0;
      }
      __syncthreads();
      int at = block_base + warp_totals[tid >> 5] + incl - mine;
      for (int k = 0; k < mine; ++k, ++at) {
        int ids[3];
#pragma unroll
        for (int j = 0; j < 3; ++j) {
          const int edge = c_tri_table[cube][3 * k + j];
          const int corner = c_edge_corner[edge], axis = c_edge_axis[edge];
          const int x = vx + (corner & 1), y = vy + ((corner >> 1) & 1), z = vz + (corner >> 2);
          const int slot = nb_slot[(x >> 3) | ((y >> 3) << 1) | ((z >> 3) << 2)];
          const int voxel = ((z & 7) << 6) | ((y & 7) << 3) | (x & 7);
          ids[j] = __ldg(p.edge_index + ((size_t)slot * VK_BLOCK_VOXELS + voxel) * 3 + axis);
        }
        if (at < p.face_capacity) {
          p.faces[3 * (size_t)at] = ids[0];
          p.faces[3 * (size_t)at + 1] = ids[1];
          p.faces[3 * (size_t)at + 2] = ids[2];
        }
      }
    }
    __syncthreads();
  }
}

// the generated tables (tools/gen_mc_tables.py) go to constant memory once per DEVICE, with that device current
constexpr int kMaxDevices = 64;
static bool g_mc_tables_loaded[kMaxDevices];

static int load_mc_tables() {
  int dev = -1;
  cudaError_t err = cudaGetDevice(&dev);
  if (err != cudaSuccess) return (int)err;
  if (dev < 0 || dev >= kMaxDevices) return VK_E_NODEVICE;
  if (g_mc_tables_loaded[dev]) return 0;
  err = cudaMemcpyToSymbol(c_tri_table, kMcTriTable, sizeof(kMcTriTable));
  if (err == cudaSuccess) err = cudaMemcpyToSymbol(c_edge_corner, kMcEdgeCorner, sizeof(kMcEdgeCorner));
  if (err == cudaSuccess) err = cudaMemcpyToSymbol(c_edge_axis, kMcEdgeAxis, sizeof(kMcEdgeAxis));
  if (err != cudaSuccess) return (int)err;
  g_mc_tables_loaded[dev] = true;
  return 0;
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
  p.block_slot = nullptr; p.slot_counter = nullptr; p.edge_index = nullptr;
  p.faces = nullptr; p.face_capacity = 0; p.face_count = nullptr;
  p.chunk = 16;  // measured on config 1 (33.6 k blocks): 128 -> 0.77 ms, 64 -> 0.58, 32 -> 0.49, 16 -> 0.47, 8 -> 0.50
  const int chunks = (p.total_entries + p.chunk - 1) / p.chunk;
  const int cap = 3 * sm_count();  // three 512-thread CTAs per SM are resident
  vk_extract_points_kernel<<<chunks < cap ? chunks : cap, 512, 0, s>>>(p);
  VK_COUNT_LAUNCH();
  return (int)cudaGetLastError();
}

extern "C" int vk_extract_mesh(const vkVolume* v, float* points, float* colors, int point_capacity, int* point_count,
                               int* faces, int face_capacity, int* face_count, void* stream) {
  if (!v || !v->voxels || !points || !point_count || !faces || !face_count || point_capacity < 0 || face_capacity < 0)
    return VK_E_INVALID;
  if (v->shard_count > 1) return VK_E_INVALID;   // see vk_extract_points
  cudaStream_t s = (cudaStream_t)stream;
  {
    const int trc = load_mc_tables();   // constant memory is per device
    if (trc) return trc;
  }
  // blocks in use -> size of the per-edge vertex index table (6 KB per block)
  int allocated = 0;
  cudaError_t err = cudaMemcpyAsync(&allocated, v->counters + VK_CNT_ALLOCATED, sizeof(int), cudaMemcpyDeviceToHost, s);
  if (err == cudaSuccess) err = cudaStreamSynchronize(s);
  if (err != cudaSuccess) return (int)err;
  const size_t B = (size_t)v->main_blocks + (size_t)v->excess_blocks;
  int *block_slot = nullptr, *slot_counter = nullptr, *edge_index = nullptr;
  err = cudaMalloc((void**)&block_slot, (B + 1) * sizeof(int));
  if (err == cudaSuccess) err = cudaMalloc((void**)&edge_index, ((size_t)allocated + 1) * VK_BLOCK_VOXELS * 3 * sizeof(int));
  int rc = 0;
  if (err == cudaSuccess) {
    slot_counter = block_slot + B;
    err = cudaMemsetAsync(slot_counter, 0, sizeof(int), s);
    if (err == cudaSuccess) err = cudaMemsetAsync(point_count, 0, sizeof(int), s);
    if (err == cudaSuccess) err = cudaMemsetAsync(face_count, 0, sizeof(int), s);
  }
  if (err == cudaSuccess) {
    ExtractParams p;
    p.entries = (const vkHashEntry*)v->entries;
    p.voxels = (const vkVoxel*)v->voxels;
    p.total_entries = v->main_blocks + v->excess_blocks;
    p.K = make_mod((uint32_t)v->main_blocks);
    p.voxel_length = v->voxel_length;
    p.block_length = VK_BLOCK_RES * v->voxel_length;
    p.points = points;
    p.colors = colors;
    p.capacity = point_capacity;
    p.count = point_count;
    p.block_slot = block_slot;
    p.slot_counter = slot_counter;
    p.edge_index = edge_index;
    p.faces = faces;
    p.face_capacity = face_capacity;
    p.face_count = face_count;
    p.chunk = 16;
    const int chunks = (p.total_entries + p.chunk - 1) / p.chunk;
    const int cap = 3 * sm_count();
    vk_extract_points_kernel<<<chunks < cap ? chunks : cap, 512, 0, s>>>(p);
    VK_COUNT_LAUNCH();
    vk_extract_faces_kernel<<<chunks < cap ? chunks : cap, 512, 0, s>>>(p);
    VK_COUNT_LAUNCH();
    err = cudaGetLastError();
    if (err == cudaSuccess) err = cudaStreamSynchronize(s);   // the workspace is freed below
  }
  cudaFree(block_slot);
  cudaFree(edge_index);
  return rc ? rc : (int)err;
}

// ---- vk_extract_mesh_static: the same mesh without any allocation (include/vulcan_b200.h) ------------------
// For callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
// that forbid it): the workspace vk_extract_mesh allocates per call - block -> slot table, slot counter and
// the per-edge vertex index table - lives in ONE file-scope device array of fixed size instead, held by
// every GPU that loads the library until the process exits.  Layout (ints):
//   [0, 64)                                   header; [0] = slot counter (cleared by every call)
//   [64, 64 + VK_MESH_STATIC_MAX_POOL)        block_slot: slot of voxel block `data`
//   then VK_MESH_STATIC_MAX_BLOCKS * 1536     edge_index rows of the allocated blocks, by slot
// block_slot and edge_index keep the previous call's contents; both kernels only read the rows the points
// kernel of the SAME call wrote (every allocated block gets a slot and a full row), so only the counter must
// start at zero.  All calls share the array: calls that overlap on the GPU (two streams, two host threads)
// race - serialise them.
namespace vk {

constexpr size_t kMeshStaticInts = 64 + (size_t)VK_MESH_STATIC_MAX_POOL +
                                   (size_t)VK_MESH_STATIC_MAX_BLOCKS * VK_BLOCK_VOXELS * 3;
static __device__ __align__(256) int g_mesh_static[kMeshStaticInts];

static int* g_mesh_static_addr[kMaxDevices];      // device address of g_mesh_static, per device (set-up below)

// Once per device, with that device current: in host code `g_mesh_static` is only the host-side stand-in
// of the symbol, NOT a device address - the address the kernels get comes from cudaGetSymbolAddress.
static int mesh_static_setup(int** scratch) {
  int dev = -1;
  cudaError_t err = cudaGetDevice(&dev);
  if (err != cudaSuccess) return (int)err;
  if (dev < 0 || dev >= kMaxDevices) return VK_E_NODEVICE;
  if (!g_mesh_static_addr[dev]) {
    int* ptr = nullptr;
    err = cudaGetSymbolAddress((void**)&ptr, g_mesh_static);
    if (err != cudaSuccess) return (int)err;
    if (!ptr) return VK_E_NODEVICE;
    g_mesh_static_addr[dev] = ptr;
  }
  const int trc = load_mc_tables();
  if (trc) return trc;
  *scratch = g_mesh_static_addr[dev];
  return 0;
}

}  // namespace vk

extern "C" int vk_extract_mesh_static(const vkVolume* v, float* points, float* colors, int point_capacity,
                                      int* point_count, int* faces, int face_capacity, int* face_count, void* stream) {
  if (!v || !v->voxels || !points || !point_count || !faces || !face_count || point_capacity < 0 || face_capacity < 0)
    return VK_E_INVALID;
  if (v->shard_count > 1) return VK_E_INVALID;   // see vk_extract_points
  const size_t B = (size_t)v->main_blocks + (size_t)v->excess_blocks;
  if (B > (size_t)VK_MESH_STATIC_MAX_POOL) return VK_E_TOO_LARGE;
  cudaStream_t s = (cudaStream_t)stream;
  int* scratch = nullptr;
  int rc = mesh_static_setup(&scratch);
  if (rc) return rc;
  // blocks in use: each takes one 6 KB row of the static vertex index table
  int allocated = 0;
  cudaError_t err = cudaMemcpyAsync(&allocated, v->counters + VK_CNT_ALLOCATED, sizeof(int), cudaMemcpyDeviceToHost, s);
  if (err == cudaSuccess) err = cudaStreamSynchronize(s);
  if (err != cudaSuccess) return (int)err;
  if (allocated < 0 || allocated > VK_MESH_STATIC_MAX_BLOCKS) return VK_E_TOO_LARGE;
  err = cudaMemsetAsync(scratch, 0, 64 * sizeof(int), s);   // the slot counter starts at zero
  if (err == cudaSuccess) err = cudaMemsetAsync(point_count, 0, sizeof(int), s);
  if (err == cudaSuccess) err = cudaMemsetAsync(face_count, 0, sizeof(int), s);
  if (err != cudaSuccess) return (int)err;
  ExtractParams p;
  p.entries = (const vkHashEntry*)v->entries;
  p.voxels = (const vkVoxel*)v->voxels;
  p.total_entries = v->main_blocks + v->excess_blocks;
  p.K = make_mod((uint32_t)v->main_blocks);
  p.voxel_length = v->voxel_length;
  p.block_length = VK_BLOCK_RES * v->voxel_length;
  p.points = points;
  p.colors = colors;
  p.capacity = point_capacity;
  p.count = point_count;
  p.slot_counter = scratch;
  p.block_slot = scratch + 64;
  p.edge_index = scratch + 64 + VK_MESH_STATIC_MAX_POOL;
  p.faces = faces;
  p.face_capacity = face_capacity;
  p.face_count = face_count;
  p.chunk = 16;
  const int chunks = (p.total_entries + p.chunk - 1) / p.chunk;
  const int cap = 3 * sm_count();
  vk_extract_points_kernel<<<chunks < cap ? chunks : cap, 512, 0, s>>>(p);
  VK_COUNT_LAUNCH();
  vk_extract_faces_kernel<<<chunks < cap ? chunks : cap, 512, 0, s>>>(p);
  VK_COUNT_LAUNCH();
  return (int)cudaGetLastError();
}
