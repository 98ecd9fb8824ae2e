// vk_trace.cu — Tracer::Trace rebuilt for sm_100a.
//
// Reference (src/tracer.cpp:41-47): ComputePatches (20-barrier scan + blocking D2H read
// of the patch count) -> FillKernel -> ComputeBounds (float CAS loops) -> ComputePoints
// (16x16 tiles, scalar loads, two inlined copies of the 8-tap interpolation with colour, a full
// hash walk per trilinear tap on block faces) -> ComputeNormals.
//
// Here:
//   vk_bounds_kernel    the visible blocks are rasterised straight into the 80x60 bounds grid:
//                       per-CTA privatised grid in shared memory (native integer atomicMin/Max on
//                       float bits), merged into the global grid by the cells a CTA touched.  No
//                       patch list, no host read-back of a count.
//   vk_raycast_kernel   a warp marches an 8x4 pixel tile.  The march is a per-lane state machine
//                       (MARCH -> REFINE -> DONE) so that the trilinear interpolation exists once in
//                       the instruction stream and lanes at different phases share it; the
//                       interpolation fetches only the 8 distances (the reference's colour blend is
//                       evaluated once per ray, for the last interpolation, after the march);
//                       taps that cross a block face resolve each distinct neighbour block once;
//                       the hash modulo and the divisions by block / voxel length are strength
//                       reduced (vk_math.cuh) - all bit-exact with the reference's arithmetic.
//   vk_normals_kernel   32x8 tile + 2 px halo in shared memory.
// The step-wise entry points of include/vulcan/tracer.cuh are kept as well.
#include "vk_math.cuh"
#include "vk_shard.cuh"
#include "vk_launch.h"
#include <algorithm>
#include <cfloat>

namespace vk {

// ---- ray bounds -------------------------------------------------------------------

struct BlockBox {
  int bminx, bminy, bmaxx, bmaxy;
  float zlo, zhi;
  int gx, gy;
};

// float atomic min/max without a CAS loop (include/vulcan/util.cuh:20-50 uses one):
// non-negative floats order like signed ints, negative floats like reversed unsigned ints
__device__ __forceinline__ void atomic_min_float(float* addr, float value) {
  if (value >= 0) atomicMin(reinterpret_cast<int*>(addr), __float_as_int(value));
  else atomicMax(reinterpret_cast<unsigned int*>(addr), __float_as_uint(value));
}
__device__ __forceinline__ void atomic_max_float(float* addr, float value) {
  if (value >= 0) atomicMax(reinterpret_cast<int*>(addr), __float_as_int(value));
  else atomicMin(reinterpret_cast<unsigned int*>(addr), __float_as_uint(value));
}

template <typename T>
__device__ __forceinline__ T vclamp(T v, T lo, T hi) {
  // vulcan::clamp (include/vulcan/math.h:23-28): min(vmax, max(vmin, v))
  const T a = (lo > v) ? lo : v;
  return (a < hi) ? a : hi;
}

// `(short)floorf(u)` as the reference's machine code does it: F2I to a saturating s32
// (NaN -> 0), then the low 16 bits (VIMNMX.S16x2 on the s32 register).
__device__ __forceinline__ int to_short(float f) { return (int)(short)(int)f; }

struct BoundsParams {
  const int32_t* visible;
  const int32_t* count;
  const vkHashEntry* entries;
  float Tcw[12];
  vkProjection proj;
  float block_length, min_depth, max_depth;
  Div image_w, image_h;
  int bw, bh;
  float* bounds;
  const uint32_t* warm;  // optional: table the ray marcher will gather from (vkVolume::space_bits)
  int warm_lines;        // its size in 128-byte lines
  const float* warm_sdf; // optional: vkVolume::sdf - the distance tiles of the visible blocks are pulled into L2
  bool negate_hi;        // cells hold (min, -max): an element-wise MIN over ranks combines both
  // sharded volume (vk_trace_sharded): the last CTA combines the grid with the other ranks' grids over
  // peer memory and splits the image rows into bands; has_x = 0 otherwise
  int has_x, image_rows;
  Xchg x;
};

// body of ComputePatchesKernel (src/tracer.cu:13-66) for one block.
// FMA map: corner = L * (i + origin) FMUL; Tcw * corner rows as xform1_xfirst; Project;
//          u' = bounds_width * u / image_width   FMUL then IEEE division.
template <bool F>
__device__ __forceinline__ BlockBox block_box(const int4& ent, const Mat34& Tcw, const BoundsParams& p) {
  int ox, oy, oz;
  entry_origin(ent, ox, oy, oz);
  int bmaxx = -1, bmaxy = -1;
  int bminx = (short)p.bw, bminy = (short)p.bh;
  float zlo = +FLT_MAX, zhi = -FLT_MAX;
  const int cw = (short)(p.bw - 1), ch = (short)(p.bh - 1);
  for (int z = 0; z <= 1; ++z) {
    const float Z = mul(p.block_length, (float)(z + oz));
    for (int y = 0; y <= 1; ++y) {
      const float Y = mul(p.block_length, (float)(y + oy));
      for (int x = 0; x <= 1; ++x) {
        const float X = mul(p.block_length, (float)(x + ox));
        const Vec3 Xc = xform1_xfirst<F>(Tcw, X, Y, Z);
        float u, v;
        project<F>(p.proj, Xc, u, v);
        u = divc(mul((float)p.bw, u), p.image_w);
        v = divc(mul((float)p.bh, v), p.image_h);
        const int fu = to_short(floorf(u)), fv = to_short(floorf(v));
        const int cu = to_short(ceilf(u)), cv = to_short(ceilf(v));
        bminx = vclamp<int>((fu < bminx) ? fu : bminx, 0, cw);
        bminy = vclamp<int>((fv < bminy) ? fv : bminy, 0, ch);
        bmaxx = vclamp<int>((cu > bmaxx) ? cu : bmaxx, 0, cw);
        bmaxy = vclamp<int>((cv > bmaxy) ? cv : bmaxy, 0, ch);
        zlo = vclamp<float>(vmin(Xc.z, zlo), p.min_depth, p.max_depth);
        zhi = vclamp<float>(vmax(Xc.z, zhi), p.min_depth, p.max_depth);
      }
    }
  }
  BlockBox b;
  b.bminx = bminx; b.bminy = bminy; b.bmaxx = bmaxx; b.bmaxy = bmaxy;
  b.zlo = zlo; b.zhi = zhi;
  const int rx = bmaxx - bminx, ry = bmaxy - bminy;
  b.gx = (rx + 15) / 16;   // Q12: rx == 0 -> no patch at all
  b.gy = (ry + 15) / 16;
  if (!(zhi > zlo)) { b.gx = 0; b.gy = 0; }
  return b;
}

// Ray bounds of a sharded volume (one CTA; vk_trace_sharded).  Stage bit 0: this rank's grid goes into its
// slot of every peer's buffer (posted stores over NVLink) and is announced.  Stage bit 1: wait for the peers'
// announcements, leave the element-wise minimum in place (cells are (min, -max)) and split the image rows into
// bands of equal active-cell counts.  The handshake is also the barrier that orders everything a rank stored
// into its peers' replicas before this kernel - its allocation and integration of the frame - before any
// rank's ray cast.
__device__ __forceinline__ void bounds_exchange(const BoundsParams& p, float* s_cells) {
  const int ncell = p.bw * p.bh, nf = 2 * ncell;
  if (xchg_stores(p.x.stages))
    for (int r = 0; r < p.x.count; ++r) {
      if (r == p.x.rank) continue;
      float* slot = xchg_bounds_slot(p.x.buf[r], p.x.parity, p.x.rank);
      for (int i = threadIdx.x; i < nf; i += blockDim.x) slot[i] = __ldcg(p.bounds + i);
    }
  if (p.x.stages == 4) return;   // store-only launch: the announcement follows in the next one
  xchg_handshake(p.x, false);
  if (!(p.x.stages & 2)) return;
  // the same grid on every rank
  for (int i = threadIdx.x; i < nf; i += blockDim.x) {
    float m = __ldcg(p.bounds + i);
    for (int r = 0; r < p.x.count; ++r)
      if (r != p.x.rank) m = fminf(m, __ldcg(xchg_bounds_slot(p.x.buf[p.x.rank], p.x.parity, r) + i));
    p.bounds[i] = m;
    s_cells[i] = m;
  }
  __syncthreads();
  // row bands with equal numbers of active cells (cells some ray has to march): the same split on every
  // rank, computed from the same grid; rank r renders pixel rows [bands[r], bands[r + 1])
  int* row_active = reinterpret_cast<int*>(s_cells + nf);
  if (threadIdx.x < p.bh) {
    int cnt = 0;
    for (int cx = 0; cx < p.bw; ++cx) {
      const float lo = s_cells[2 * (threadIdx.x * p.bw + cx)], nhi = s_cells[2 * (threadIdx.x * p.bw + cx) + 1];
      cnt += (lo < -nhi) ? 1 : 0;
    }
    row_active[threadIdx.x] = cnt;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    int* bands = xchg_bands(p.x.buf[p.x.rank], p.x.parity);
    long long total = 0;
    for (int r = 0; r < p.bh; ++r) total += row_active[r];
    long long acc = 0;
    int k = 1;
    bands[0] = 0;
    for (int r = 0; r < p.bh && k < p.x.count; ++r) {
      acc += row_active[r];
      // first pixel row of bounds row r + 1 (a pixel row y lies in bounds row bh * y / image_rows)
      while (k < p.x.count && acc * p.x.count >= total * k) bands[k++] = ((r + 1) * p.image_rows + p.bh - 1) / p.bh;
    }
    while (k <= p.x.count) bands[k++] = p.image_rows;
  }
}

// K13 + K10 + K11 fused: the union of a block's patches is the cell rectangle
// [bmin, min(bmax, bmin + 16*g - 1)] (src/tracer.cu:67-86).  The global grid must hold
// (+FLT_MAX, -FLT_MAX) on entry (vk_reset_bounds).  Each CTA accumulates into a private copy
// of the grid in shared memory and then merges the cells it touched.
template <bool F>
__global__ void __launch_bounds__(512) vk_bounds_kernel(const BoundsParams p) {
  pdl_wait();     // the previous kernel of the chain has completed (no-op for a plain launch)
  pdl_trigger();  // the next one may be scheduled as this grid drains
  extern __shared__ float s_cells[];  // bw*bh*2
  // The ray marcher that follows walks free space through the occupancy filter with dependent
  // gathers; its longest rays set the duration of the whole trace.  Pull the filter into L2 now
  // (this kernel leaves most of the machine idle) so those gathers are L2 hits, not DRAM misses.
  if (p.warm)
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < p.warm_lines; i += gridDim.x * blockDim.x)
      asm volatile("prefetch.global.L2 [%0];" ::"l"(p.warm + (size_t)i * 32));
  const int ncell = p.bw * p.bh;
  for (int i = threadIdx.x; i < ncell; i += blockDim.x) {
    s_cells[2 * i] = +FLT_MAX;
    s_cells[2 * i + 1] = -FLT_MAX;
  }
  __syncthreads();
  const int n = *p.count;
  Mat34 Tcw;
  Tcw.load(p.Tcw);
  // contiguous chunk of the visible list per CTA: neighbouring list entries are neighbouring
  // blocks (the list is built tile by tile), so a CTA touches few cells
  const int per_cta = (n + gridDim.x - 1) / gridDim.x;
  const int begin = blockIdx.x * per_cta;
  const int end = min(n, begin + per_cta);
  for (int i = begin + threadIdx.x; i < end; i += blockDim.x) {
    const int e = __ldg(p.visible + i);
    const int4 ent = ld_entry(p.entries + e);
    // The ray marcher samples the distances of these very blocks through chains of dependent loads; its
    // longest rays set the duration of the whole trace.  Their 2 KB distance tiles go to L2 now (16 lines
    // per block, ~30 MB per view, nobody waits for them): a sample then costs an L2 hit, not a DRAM miss.
    if (p.warm_sdf && ent.z >= 0) {
      const float* tile = p.warm_sdf + (size_t)ent.z * VK_BLOCK_VOXELS;
#pragma unroll
      for (int l = 0; l < VK_BLOCK_VOXELS / 32; ++l) asm volatile("prefetch.global.L2 [%0];" ::"l"(tile + 32 * l));
    }
    const BlockBox b = block_box<F>(ent, Tcw, p);
    if (b.gx * b.gy == 0) continue;
    const int x1 = min(b.bmaxx, b.bminx + 16 * b.gx - 1);
    const int y1 = min(b.bmaxy, b.bminy + 16 * b.gy - 1);
    for (int y = b.bminy; y <= y1; ++y)
      for (int x = b.bminx; x <= x1; ++x) {
        float* cell = s_cells + 2 * (y * p.bw + x);
        atomic_min_float(cell, b.zlo);
        atomic_max_float(cell + 1, b.zhi);
      }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < ncell; i += blockDim.x) {
    const float lo = s_cells[2 * i], hi = s_cells[2 * i + 1];
    if (lo != +FLT_MAX) atomic_min_float(p.bounds + 2 * i, lo);
    if (hi != -FLT_MAX) {
      if (p.negate_hi) atomic_min_float(p.bounds + 2 * i + 1, -hi);
      else atomic_max_float(p.bounds + 2 * i + 1, hi);
    }
  }
  if (!p.has_x) return;
  // ---- sharded volume: the last CTA to arrive holds the rank's complete grid -------------------------
  __shared__ bool last;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) last = (atomicAdd(xchg_ticket(p.x.buf[p.x.rank]), 1) == (int)gridDim.x - 1);
  __syncthreads();
  if (!last) return;
  if (threadIdx.x == 0) *xchg_ticket(p.x.buf[p.x.rank]) = 0;
  __threadfence();
  bounds_exchange(p, s_cells);
}

// The second half of bounds_exchange as a launch of its own (one CTA), for ranks that share a GPU
// (vk_trace_sharded_staged): by then every rank's first half has run, so nothing waits.
__global__ void __launch_bounds__(512) vk_bounds_merge_kernel(const BoundsParams p) {
  pdl_wait();
  pdl_trigger();
  extern __shared__ float s_cells[];
  bounds_exchange(p, s_cells);
}

__global__ void vk_reset_bounds_kernel(float2* bounds, int count, float hi_init) {
  pdl_wait();     // the previous kernel of the chain has completed (no-op for a plain launch)
  pdl_trigger();  // the next one may be scheduled as this grid drains
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < count) bounds[i] = make_float2(+FLT_MAX, hi_init);
}

// ComputePatchesKernel (src/tracer.cu:13-87) with a warp-aggregated append in place of
// the shared-memory scan; patch order is unspecified in the reference too (Q7-like).
template <bool F>
__global__ void __launch_bounds__(256) vk_patches_kernel(const int* indices, const vkHashEntry* entries,
                                                         const BoundsParams p, int block_count,
                                                         vkPatch* patches, int* patch_count) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int lane = threadIdx.x & 31;
  Mat34 Tcw;
  Tcw.load(p.Tcw);
  BlockBox b;
  b.gx = b.gy = 0;
  if (i < block_count) {
    const int4 ent = ld_entry(entries + indices[i]);
    b = block_box<F>(ent, Tcw, p);
  }
  const int cnt = b.gx * b.gy;
  // inclusive warp scan of cnt
  int incl = cnt;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const int o = __shfl_up_sync(0xffffffffu, incl, d);
    if (lane >= d) incl += o;
  }
  const int total = __shfl_sync(0xffffffffu, incl, 31);
  int base = 0;
  if (lane == 0 && total > 0) base = atomicAdd(patch_count, total);
  base = __shfl_sync(0xffffffffu, base, 0);
  const int offset = base + incl - cnt;
  for (int ii = 0; ii < b.gy; ++ii)
    for (int j = 0; j < b.gx; ++j) {
      vkPatch pt;
      pt.origin[0] = (short)(b.bminx + 16 * j);
      pt.origin[1] = (short)(b.bminy + 16 * ii);
      pt.size[0] = (short)min(16, b.bmaxx - pt.origin[0] + 1);
      pt.size[1] = (short)min(16, b.bmaxy - pt.origin[1] + 1);
      pt.bounds[0] = b.zlo;
      pt.bounds[1] = b.zhi;
      patches[offset + ii * b.gx + j] = pt;
    }
}

// ComputeBoundsKernel (src/tracer.cu:89-112)
__global__ void __launch_bounds__(256) vk_patch_bounds_kernel(const vkPatch* patches, float* bounds, int bw,
                                                              int patch_count) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= patch_count) return;
  const vkPatch pt = patches[i];
  for (int a = 0; a < pt.size[1]; ++a)
    for (int b = 0; b < pt.size[0]; ++b) {
      float* cell = bounds + 2 * ((pt.origin[1] + a) * bw + pt.origin[0] + b);
      atomic_min_float(cell, pt.bounds[0]);
      atomic_max_float(cell + 1, pt.bounds[1]);
    }
}

}  // namespace vk

// ---- ray marcher ------------------------------------------------------------------

namespace vk {

struct TraceParams {
  const vkHashEntry* entries;
  const vkVoxel* voxels;
  const float* sdf;             // vkVolume::sdf: distances alone, block-major (SOA kernels); NULL for raw-pointer calls
  const float* bounds;
  const uint32_t* bucket_bits;  // optional: bit h set <=> bucket head h allocated (vkVolume::bucket_bits)
  const uint32_t* space_bits;   // optional: free-space filter (vkVolume::space_bits)
  int free_run;                 // max free-space steps per lane per trip (0 = filter off)
  bool negated_hi;              // bounds cells hold (min, -max) (vk_trace_bounds / vk_trace_rows)
  int coop_lanes, coop_rounds;  // cooperative free-space runs when at most coop_lanes lanes are in free space
  int pair_after;               // trips after which a warp takes its free-space steps two at a time (0 = never)
  Mod K;
  Div block_length, voxel_length;
  float trunc;
  float Twc[12], Tcw[12];
  vkProjection proj;
  float* depths;
  float* colors;
  unsigned long long* keys;
  int w, h, bw, bh;
  int shard_rank, shard_count, shard_shift;
  int batch_lanes, batch_wait;  // phase-B batching of the ray marcher (see vk_raycast_kernel)
  // Sharded volumes: the tables of every rank (own pointers in slot peer_rank, the others mapped
  // over NVLink).  A block id in the marcher is (rank << kPeerShift) | voxel-block id.
  const vkHashEntry* peer_entries[VK_MAX_SHARDS];
  const vkVoxel* peer_voxels[VK_MAX_SHARDS];
  const float* peer_sdf[VK_MAX_SHARDS];
  int peer_count, peer_rank;
  int y_begin, y_end;           // image rows marched by this launch
  const int* bands;             // sharded volume: device table of row bands (this rank marches bands[band_rank] - 2 ..
  int band_rank;                // bands[band_rank + 1] + 2); NULL: y_begin / y_end
  int with_color;               // 0: the volume holds no colour (every color_weight is 0): skip the colour gather
  int debug_flags;              // diagnostics: 1 = skip the filter load, 2 = skip the block division
  int debug_only_y;
  int debug_only_x;             // diagnostics: march only the tile column starting at this x
  unsigned long long* timeline; // diagnostics: per-warp (start, end) globaltimer, or NULL
};

// Per-lane cache of block lookups in shared memory: kCacheSlots (coordinates -> voxel-block id,
// or -1 = no such block) per lane, direct mapped.  The hash table is read-only during a trace, so
// replaying a lookup from the cache is exact.  A ray looks up the same few blocks over and over
// (the block under it for ~8 steps, and its face neighbours for the interpolation taps); the
// cache turns those hash walks into one shared-memory probe, and lets the look-ahead prefetch
// resolve blocks without ever waiting on the table.
constexpr int kCacheSlots = 8;
constexpr int kMaxRayThreads = 64;
struct BlockCache {
  unsigned long long* keys;  // [kCacheSlots][blockDim.x], this lane's column
  int* data;
  int stride;
  __device__ __forceinline__ static int slot(int bx, int by, int bz) { return (bx + 3 * by + 5 * bz) & (kCacheSlots - 1); }
  __device__ __forceinline__ bool peek(int bx, int by, int bz, int& out) const {
    const int s = slot(bx, by, bz) * stride;
    if (keys[s] != pack_block(bx, by, bz)) return false;
    out = data[s];
    return true;
  }
  __device__ __forceinline__ void put(int bx, int by, int bz, int value) {
    const int s = slot(bx, by, bz) * stride;
    keys[s] = pack_block(bx, by, bz);
    data[s] = value;
  }
};

// `while (entry.block != b && entry.HasNext()) entry = entries[entry.next];`
// `entry.block == b && entry.IsAllocated()`  (src/tracer.cu:363-371, :151-175)
constexpr int kPeerShift = 26;

__device__ __forceinline__ int walk_chain(const vkHashEntry* entries, uint32_t h, unsigned long long key) {
  int4 e = ld_entry(entries + h);
  while (entry_block(e) != key && e.w != -1) e = ld_entry(entries + e.w);
  return (entry_block(e) == key) ? e.z : -1;
}

template <bool PEERS>
__device__ __forceinline__ int lookup_block(const TraceParams& p, int bx, int by, int bz) {
  const unsigned long long key = pack_block(bx, by, bz);
  const uint32_t h = block_hash(bx, by, bz, p.K);
  if (PEERS) {
    // the owner's table: the local one, or a peer's over NVLink (peer_entries[rank] is local)
    const int owner = shard_owner(bx, by, bz, p.shard_shift, p.peer_count);
    const int data = walk_chain(p.peer_entries[owner], h, key);
    return (data != -1) ? ((owner << kPeerShift) | data) : -1;
  }
  // an unallocated bucket head has no chain: most steps through free space end here, on a
  // 32 KB bitmap that stays in L1, instead of on a 16-byte entry fetched from L2
  if (p.bucket_bits && !p.space_bits && !((__ldg(p.bucket_bits + (h >> 5)) >> (h & 31)) & 1u)) return -1;
  int4 e = ld_entry(p.entries + h);
  while (entry_block(e) != key && e.w != -1) e = ld_entry(p.entries + e.w);
  return (entry_block(e) == key) ? e.z : -1;
}
template <bool PEERS>
__device__ __forceinline__ int find_data(const TraceParams& p, BlockCache& cache, int bx, int by, int bz) {
  int data;
  if (cache.peek(bx, by, bz, data)) return data;
  data = lookup_block<PEERS>(p, bx, by, bz);
  cache.put(bx, by, bz, data);
  return data;
}

template <bool PEERS>
__device__ __forceinline__ const uint32_t* voxel_words(const TraceParams& p, int data, int voxel_offset) {
  // Voxel is 20 bytes, 4-byte aligned: distance, rgb, (dw | cw << 16)
  if (!PEERS) return reinterpret_cast<const uint32_t*>(p.voxels) + (size_t)data * (VK_BLOCK_VOXELS * 5) + voxel_offset * 5;
  const vkVoxel* pool = p.peer_voxels[data >> kPeerShift];
  return reinterpret_cast<const uint32_t*>(pool) + (size_t)(data & ((1 << kPeerShift) - 1)) * (VK_BLOCK_VOXELS * 5) +
         voxel_offset * 5;
}

// Voxel::distance of voxel `voxel_offset` of block `data`.  SOA: from the block-major distance mirror
// (2 KB per block: a 128-byte line holds 32 voxels, against 1.6 voxels per 32-byte sector of the 20-byte
// AoS voxels), which the integrators keep in step with the voxel pool; otherwise from the voxel itself.
// Q22: voxel_offset may exceed 511 on an exact block face; both layouts then read the same voxel of
// the next block of the pool.
template <bool PEERS, bool SOA>
__device__ __forceinline__ float load_distance(const TraceParams& p, int data, int voxel_offset) {
  if (!SOA) return __uint_as_float(__ldg(voxel_words<PEERS>(p, data, voxel_offset)));
  if (!PEERS) return __ldg(p.sdf + (size_t)data * VK_BLOCK_VOXELS + voxel_offset);
  return __ldg(p.peer_sdf[data >> kPeerShift] + (size_t)(data & ((1 << kPeerShift) - 1)) * VK_BLOCK_VOXELS + voxel_offset);
}

// The 8 taps of GetInterpolatedDistance (src/tracer.cu:190-256): per axis the tap pair
// (i0, i0 + 1), i0 = floor(w - 0.5) in [-1, 7]; a tap outside [0, 8) lives in the neighbouring
// block (GetVoxel, src/tracer.cu:114-188: one wrap per axis), which is looked up ONCE per
// distinct neighbour instead of once per tap; a missing neighbour reads as Voxel::Empty() (Q16).
struct Taps {
  int data[8];   // voxel-block id per tap (index = z*4 + y*2 + x), -1 = empty
  int off[8];    // voxel offset inside that block
  float w1x, w1y, w1z;
};

template <bool PEERS>
__device__ __forceinline__ void resolve_taps(const TraceParams& p, BlockCache& cache, int bx, int by, int bz,
                                             int home, float wx, float wy, float wz, Taps& t) {
  const int i0x = (int)floorf(add(wx, -0.5f));
  const int i0y = (int)floorf(add(wy, -0.5f));
  const int i0z = (int)floorf(add(wz, -0.5f));
  // w1 = w - (i0 + 0.5)  (src/tracer.cu:258-261)
  t.w1x = sub(wx, add((float)i0x, 0.5f));
  t.w1y = sub(wy, add((float)i0y, 0.5f));
  t.w1z = sub(wz, add((float)i0z, 0.5f));
  // block offset and in-block coordinate of tap 0 / tap 1 per axis
  int ox[2], oy[2], oz[2], cx[2], cy[2], cz[2];
#define VK_AXIS(i0, o, c)                                   \
  o[0] = (i0 < 0) ? -1 : 0;  c[0] = (i0 < 0) ? i0 + 8 : i0; \
  o[1] = (i0 + 1 >= 8) ? 1 : 0;  c[1] = (i0 + 1 >= 8) ? i0 + 1 - 8 : i0 + 1;
  VK_AXIS(i0x, ox, cx)
  VK_AXIS(i0y, oy, cy)
  VK_AXIS(i0z, oz, cz)
#undef VK_AXIS
  // axes on which the two taps lie in different blocks
  const int cross = ((ox[0] != ox[1]) ? 1 : 0) | ((oy[0] != oy[1]) ? 2 : 0) | ((oz[0] != oz[1]) ? 4 : 0);
#pragma unroll
  for (int c = 0; c < 8; ++c) {
    const int tx = c & 1, ty = (c >> 1) & 1, tz = c >> 2;
    t.off[c] = cz[tz] * 64 + cy[ty] * 8 + cx[tx];
    const int r = c & cross;  // representative tap: same block as tap c
    int d;
    if (r == c) {
      const int dx = ox[tx], dy = oy[ty], dz = oz[tz];
      if ((dx | dy | dz) == 0) d = home;
      else d = find_data<PEERS>(p, cache, bx + dx, by + dy, bz + dz);
    } else {
      d = t.data[0];
#pragma unroll
      for (int s = 1; s < 8; ++s)
        if (s < c && (s & c) == s && r == s) d = t.data[s];
    }
    t.data[c] = d;
  }
}

// trilinear SDF (src/tracer.cu:265-280,312).
// FMA map: n00 = v000*w0x + v001*w1x -> FMUL v001*w1x ; FFMA v000*w0x + .   (same for y and z)
template <bool F, bool PEERS, bool SOA>
__device__ __forceinline__ float interp_distance(const TraceParams& p, const Taps& t) {
  float n[8];
#pragma unroll
  for (int c = 0; c < 8; ++c) {
    n[c] = 1.0f;  // Voxel::Empty()
    if (t.data[c] != -1) n[c] = load_distance<PEERS, SOA>(p, t.data[c], t.off[c]);
  }
  const float w0x = sub(1.0f, t.w1x), w0y = sub(1.0f, t.w1y), w0z = sub(1.0f, t.w1z);
  const float n00 = FP<F>::mad(n[0], w0x, mul(n[1], t.w1x));
  const float n01 = FP<F>::mad(n[2], w0x, mul(n[3], t.w1x));
  const float n10 = FP<F>::mad(n[4], w0x, mul(n[5], t.w1x));
  const float n11 = FP<F>::mad(n[6], w0x, mul(n[7], t.w1x));
  const float n0 = FP<F>::mad(n00, w0y, mul(n01, t.w1y));
  const float n1 = FP<F>::mad(n10, w0y, mul(n11, t.w1y));
  return FP<F>::mad(n0, w0z, mul(n1, t.w1z));
}

// colour blend of the same 8 taps (src/tracer.cu:282-310), Q15: each weight is 1 if the trilinear
// weight is non-zero and the voxel has colour, else 0.
// FMA map: color += w * c -> FFMA chain from 0; color /= total -> reciprocal, FMUL.
template <bool F, bool PEERS>
__device__ __forceinline__ Vec3 interp_color(const TraceParams& p, const Taps& t) {
  const float w0x = sub(1.0f, t.w1x), w0y = sub(1.0f, t.w1y), w0z = sub(1.0f, t.w1z);
  float total = 0.0f;
  float cr = 0.0f, cg = 0.0f, cb = 0.0f;
#pragma unroll
  for (int c = 0; c < 8; ++c) {
    float r = 0.f, g = 0.f, b = 0.f;
    int cw = 0;
    if (t.data[c] != -1) {
      const uint32_t* w = voxel_words<PEERS>(p, t.data[c], t.off[c]);
      r = __uint_as_float(__ldg(w + 1));
      g = __uint_as_float(__ldg(w + 2));
      b = __uint_as_float(__ldg(w + 3));
      cw = (int)(int16_t)(__ldg(w + 4) >> 16);
    }
    const float wz_ = (c >> 2) ? t.w1z : w0z;
    const float wy_ = ((c >> 1) & 1) ? t.w1y : w0y;
    const float wx_ = (c & 1) ? t.w1x : w0x;
    const float prod = mul(mul(mul(wz_, wy_), wx_), (cw > 0) ? 1.0f : 0.0f);
    const float wgt = (prod != 0.0f) ? 1.0f : 0.0f;  // `a * b * c * (cw > 0) ? 1 : 0`
    total = add(total, wgt);
    cr = FP<F>::mad(wgt, r, cr);
    cg = FP<F>::mad(wgt, g, cg);
    cb = FP<F>::mad(wgt, b, cb);
  }
  if (total > 0) {
    const float inv = rcp(total);
    cr = mul(cr, inv); cg = mul(cg, inv); cb = mul(cb, inv);
  }
  Vec3 o = {cr, cg, cb};
  return o;
}

}  // namespace vk

namespace vk {



__device__ __forceinline__ bool space_bit(const TraceParams& p, int bx, int by, int bz) {
  uint32_t sw, sb;
  space_bit_index(bx, by, bz, sw, sb);
  return (__ldg(p.space_bits + sw) >> sb) & 1u;
}

enum { ST_MARCH = 0, ST_REFINE = 1, ST_DONE = 2 };

// ComputePointsKernel (src/tracer.cu:317-451).
// FMA map (reference release SASS):
//   Xcp = lo * Unproject(uv)             FMUL
//   p   = Twc * (Xcp, 1), dir = Twc * (Xcp, 0)   shared row sum; FADD m3 / FFMA 0*m3 + .
//   b   = floor(p / L)                   IEEE division
//   w   = (p - b*L) / l                  FFMA -b*L + p ; IEEE division
//   p  += dir * s                        FFMA dir*s + p   (s = L | max(l, trunc*sdf) | trunc*sdf)
//   depth = (Tcw * (p,1)).z              xform1 row 2
//
// One warp marches an 8x4 pixel tile.  Every lane is a small state machine; a loop trip is
//   phase A (cheap, every running lane): block under p, hash lookup, nearest-voxel sample;
//   phase B (expensive, ~4x phase A): the 8-tap interpolation for lanes that asked for one.
// Lanes that need phase B PARK (keep their block / voxel coordinates in registers) while the
// other lanes keep taking cheap steps; phase B runs for all parked lanes at once when at least
// `batch_lanes` are parked, when nobody else can advance, or after `batch_wait` trips - so the
// expensive instructions run with most lanes active instead of 7-9 of 32.  The order of
// operations of each ray is unchanged, so the result is bit-identical to the per-ray loop.
// DIAG: per-warp timeline / trip counters for tools/tune_bench.py (compiled out of the product kernels)
template <bool F, bool KEYS, bool PEERS, bool DIAG, bool SOA>
__global__ void __launch_bounds__(kMaxRayThreads, 16) vk_raycast_kernel(const TraceParams p) {
  pdl_wait();     // the previous kernel of the chain has completed (no-op for a plain launch)
  pdl_trigger();  // the next one may be scheduled as this grid drains
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int wpc = blockDim.x >> 5;
  // warp tile 8x4 pixels, CTA = wpc tiles stacked in y
  const int x = blockIdx.x * 8 + (lane & 7);
  const int y = p.y_begin + (blockIdx.y * wpc + warp) * 4 + (lane >> 3);
  int y_lo = p.y_begin, y_hi = p.y_end;
  if (p.bands) {  // rows of this rank's band plus the two rows either side its normals need
    const int b0 = p.bands[p.band_rank], b1 = p.bands[p.band_rank + 1];
    y_lo = (b1 > b0) ? max(0, b0 - 2) : 0;
    y_hi = (b1 > b0) ? min(p.h, b1 + 2) : 0;
  }
  const bool inside = (x < p.w && y >= y_lo && y < y_hi);
  unsigned long long* const timeline = DIAG ? p.timeline : nullptr;
  unsigned long long t_start = 0;
  if (timeline) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_start));
  const long long c_start = DIAG ? clock64() : 0;
  int dbg_trips = 0;
  int trips = 0;   // loop trips of this warp (warp-uniform)
  long long c_free = 0, c_a = 0, c_b = 0, c_t = 0, c_la = 0;
  int dbg_alive_sum = 0;
  int dbg_steps = 0, dbg_rounds = 0, dbg_pair = 0, dbg_single = 0;
  long long dbg_c1 = 0, dbg_c2 = 0;

  float lo = 0.f, hi = 0.f;
  if (inside) {
    const int px = p.bw * x / p.w;
    const int py = p.bh * y / p.h;
    const float2 bound = __ldg(reinterpret_cast<const float2*>(p.bounds) + py * p.bw + px);
    lo = bound.x;
    hi = p.negated_hi ? -bound.y : bound.y;
  }

  int state = (inside && lo < hi) ? ST_MARCH : ST_DONE;
  if (p.debug_only_x && (int)(blockIdx.x * 8) != p.debug_only_x) state = ST_DONE;  // diagnostics
  if (p.debug_only_y && (y & ~3) != p.debug_only_y) state = ST_DONE;
  float final_depth = 0.f;
  bool red = false;         // iteration cap hit: colour (1, 0, 0)
  bool has_interp = false;  // an interpolation happened: colour comes from its taps
  bool owned_hit = false;
  Vec3 pt = {0.f, 0.f, 0.f}, dir = {0.f, 0.f, 0.f};
  Vec3 ipt = {0.f, 0.f, 0.f};  // point of the last interpolation
  Mat34 Tcw;
  Tcw.load(p.Tcw);
  int iters = 0;
  __shared__ unsigned long long s_keys[kCacheSlots * kMaxRayThreads];
  __shared__ int s_data[kCacheSlots * kMaxRayThreads];
  BlockCache cache;
  cache.keys = s_keys + threadIdx.x;
  cache.data = s_data + threadIdx.x;
  cache.stride = blockDim.x;
#pragma unroll
  for (int i = 0; i < kCacheSlots; ++i) cache.keys[i * cache.stride] = ~0ull;

  if (state == ST_MARCH) {
    Mat34 Twc;
    Twc.load(p.Twc);
    Unprojector un;
    un.init(p.proj);
    const float ux = un.x<F>(add((float)x, 0.5f)), uy = un.y<F>(add((float)y, 0.5f));
    const float cx = mul(ux, lo), cy = mul(uy, lo), cz = lo;
    pt = xform1<F>(Twc, cx, cy, cz);
    dir = normalized<F>(xform0<F>(Twc, cx, cy, cz));
  }

  // parked lane: waits for phase B with the block / voxel coordinates of its point
  bool parked = false;
  int kbx = 0, kby = 0, kbz = 0, kdata = -1;
  float kwx = 0.f, kwy = 0.f, kwz = 0.f;
  int waited = 0;  // warp-uniform: trips since a lane parked without phase B running

  // transitions of a lane whose block was found, given the distance the reference would hold in
  // `sdf` at that point (src/tracer.cu:398-437)
  auto advance = [&](float sdf, int bx, int by, int bz) -> bool {
    if (state == ST_MARCH) {
      if (sdf <= 0.0f) {
        if (KEYS) owned_hit = (shard_owner(bx, by, bz, p.shard_shift, p.shard_count) == p.shard_rank);
        const float s = mul(p.trunc, sdf);
        pt.x = FP<F>::mad(dir.x, s, pt.x);
        pt.y = FP<F>::mad(dir.y, s, pt.y);
        pt.z = FP<F>::mad(dir.z, s, pt.z);
        state = ST_REFINE;
        return false;
      }
      const float s = vmax(p.voxel_length.d, mul(p.trunc, sdf));
      pt.x = FP<F>::mad(dir.x, s, pt.x);
      pt.y = FP<F>::mad(dir.y, s, pt.y);
      pt.z = FP<F>::mad(dir.z, s, pt.z);
      return true;  // stepped: depth / iteration test follows
    }
    // ST_REFINE, block found: second refinement step, then the final depth
    const float s = mul(p.trunc, sdf);
    pt.x = FP<F>::mad(dir.x, s, pt.x);
    pt.y = FP<F>::mad(dir.y, s, pt.y);
    pt.z = FP<F>::mad(dir.z, s, pt.z);
    final_depth = xform1_z<F>(Tcw, pt.x, pt.y, pt.z);
    state = ST_DONE;
    return false;
  };
  auto after_step = [&]() {
    const float depth = xform1_z<F>(Tcw, pt.x, pt.y, pt.z);
    if (++iters >= 500) { red = true; state = ST_DONE; }
    else if (!(depth < hi)) state = ST_DONE;
  };

  auto block_of = [&](const Vec3& q, int& ix, int& iy, int& iz) {
    float fx, fy, fz;
    divc3(q.x, q.y, q.z, p.block_length, fx, fy, fz);
    ix = (int)floorf(fx);
    iy = (int)floorf(fy);
    iz = (int)floorf(fz);
  };

  // up to max_steps free-space steps of every lane that is in free space, two per pass (see the long-warp
  // comment in the loop below): the second step's position, depth, block and filter bit are evaluated
  // alongside the first one's and used only if the first step leaves the ray in free space
  auto pair_steps = [&](bool& freeing, int& bx, int& by, int& bz, int max_steps) {
    for (int n = 0; n < max_steps && __any_sync(0xffffffffu, freeing); n += 2) {
      if (DIAG) ++dbg_pair;
      if (freeing) {
        Vec3 p1, p2;
        p1.x = FP<F>::mad(dir.x, p.block_length.d, pt.x);
        p1.y = FP<F>::mad(dir.y, p.block_length.d, pt.y);
        p1.z = FP<F>::mad(dir.z, p.block_length.d, pt.z);
        p2.x = FP<F>::mad(dir.x, p.block_length.d, p1.x);
        p2.y = FP<F>::mad(dir.y, p.block_length.d, p1.y);
        p2.z = FP<F>::mad(dir.z, p.block_length.d, p1.z);
        const float d1 = xform1_z<F>(Tcw, p1.x, p1.y, p1.z), d2 = xform1_z<F>(Tcw, p2.x, p2.y, p2.z);
        int b1x, b1y, b1z, b2x, b2y, b2z;
        block_of(p1, b1x, b1y, b1z);
        block_of(p2, b2x, b2y, b2z);
        const bool occ1 = space_bit(p, b1x, b1y, b1z), occ2 = space_bit(p, b2x, b2y, b2z);
        // first step (after_step of the reference's trip)
        pt = p1;
        bx = b1x; by = b1y; bz = b1z;
        freeing = false;
        if (++iters >= 500) { red = true; state = ST_DONE; }
        else if (!(d1 < hi)) state = ST_DONE;
        else freeing = !occ1;
        if (freeing) {   // second step
          pt = p2;
          bx = b2x; by = b2y; bz = b2z;
          freeing = false;
          if (++iters >= 500) { red = true; state = ST_DONE; }
          else if (!(d2 < hi)) state = ST_DONE;
          else freeing = !occ2;
        }
      }
    }
  };

  while (__any_sync(0xffffffffu, state != ST_DONE)) {
    ++dbg_trips;
    ++trips;
    if (timeline) c_t = clock64();
    // ---- phase A: block under p, nearest voxel ---------------------------------------------
    int bx = 0, by = 0, bz = 0;
    if (state != ST_DONE && !parked) {
      block_of(pt, bx, by, bz);
    }
    // Free-space run.  While the occupancy filter proves that no block exists under p, the
    // reference's trip is "block not found": p += L * dir, then the depth / iteration test.  All
    // lanes in free space take those steps TOGETHER in a lean warp-uniform loop (the duration of
    // the kernel is its longest ray, a lone warp issues one instruction per ~5 cycles, and lanes
    // that loop on their own would serialise), up to free_run steps per trip so that lanes
    // waiting in other states are not starved.
    if (p.free_run > 0) {
      bool freeing = (state == ST_MARCH && !parked && !space_bit(p, bx, by, bz));
      unsigned fmask = __ballot_sync(0xffffffffu, freeing);
      // Few lanes in free space (the tail of the kernel: a handful of rays that graze the
      // surface and then cross metres of empty space): all 32 lanes evaluate the next 32
      // positions of ONE such ray at a time - lane j takes the position after j+1 steps, computed
      // by the same chain of FMAs the ray itself would run - and the ray jumps to the first
      // position that is occupied, out of range or at the iteration cap.
      const long long c_coop0 = timeline ? clock64() : 0;
      if (timeline) dbg_alive_sum += __popc(__ballot_sync(0xffffffffu, state != ST_DONE));
      for (int round = 0; round < p.coop_rounds && fmask != 0 && __popc(fmask) <= p.coop_lanes; ++round) {
        unsigned todo = fmask;
        while (todo) {
          ++dbg_rounds;
          const int f = __ffs(todo) - 1;
          todo &= todo - 1;
          Vec3 c = {__shfl_sync(0xffffffffu, pt.x, f), __shfl_sync(0xffffffffu, pt.y, f),
                    __shfl_sync(0xffffffffu, pt.z, f)};
          const float dx = __shfl_sync(0xffffffffu, dir.x, f), dy = __shfl_sync(0xffffffffu, dir.y, f),
                      dz = __shfl_sync(0xffffffffu, dir.z, f);
          const float fhi = __shfl_sync(0xffffffffu, hi, f);
          const int fit = __shfl_sync(0xffffffffu, iters, f);
#pragma unroll 4
          for (int k = 0; k < 32; ++k) {
            if (k <= lane) {
              c.x = FP<F>::mad(dx, p.block_length.d, c.x);
              c.y = FP<F>::mad(dy, p.block_length.d, c.y);
              c.z = FP<F>::mad(dz, p.block_length.d, c.z);
            }
          }
          const float depth = xform1_z<F>(Tcw, c.x, c.y, c.z);
          int cbx, cby, cbz;
          block_of(c, cbx, cby, cbz);
          const bool occ = space_bit(p, cbx, cby, cbz);
          const bool cap = (fit + lane + 1 >= 500);
          const bool out = !(depth < fhi);
          const unsigned sm = __ballot_sync(0xffffffffu, cap || out || occ);
          const int w = sm ? __ffs(sm) - 1 : 31;  // first position where the run ends, else the last one
          const float rx = __shfl_sync(0xffffffffu, c.x, w), ry = __shfl_sync(0xffffffffu, c.y, w),
                      rz = __shfl_sync(0xffffffffu, c.z, w);
          const int rbx = __shfl_sync(0xffffffffu, cbx, w), rby = __shfl_sync(0xffffffffu, cby, w),
                    rbz = __shfl_sync(0xffffffffu, cbz, w);
          const int rflags = __shfl_sync(0xffffffffu, (cap ? 1 : 0) | (out ? 2 : 0) | (occ ? 4 : 0), w);
          if (lane == f) {
            dbg_steps += w + 1;
            pt.x = rx; pt.y = ry; pt.z = rz;
            bx = rbx; by = rby; bz = rbz;
            iters = fit + w + 1;
            if (rflags & 1) { red = true; state = ST_DONE; freeing = false; }
            else if (rflags & 2) { state = ST_DONE; freeing = false; }
            else if (rflags & 4) freeing = false;
          }
        }
        fmask = __ballot_sync(0xffffffffu, freeing);
      }
      if (timeline) c_la += clock64() - c_coop0;
      // Many lanes in free space: they take the steps together, one step per loop trip.
      // (Measured and rejected on B200, see profiles/r01_raycast_tail.md and profiles/r02_raycast_tail.md:
      // 4 positions per lane and trip, a brick-level map for 4-step jumps, one lane group per ray in the
      // cooperative runs, several sampling steps per trip, a second pass for the rays that outlive their
      // tile, paired steps before the cooperative runs, deferring long free flights until no lane is near
      // the surface.)
      // A warp that is still marching after pair_after trips is one of the few long ones (the average warp is
      // done after four): the machine is emptying, its own instruction stream is what the kernel waits for, and
      // that stream is a chain of dependent operations per step.  Such a warp takes its free-space steps in
      // PAIRS: the position, depth, block and filter bit of the step after next are evaluated alongside the
      // next one's (independent work for the same pipeline latency) and used only if the first step leaves the
      // ray in free space - the same operations in the same order per ray, hence the same result.
      if (p.pair_after > 0 && trips > p.pair_after) pair_steps(freeing, bx, by, bz, p.free_run);
      for (int n = 0; n < p.free_run && __any_sync(0xffffffffu, freeing); ++n) {
        if (DIAG) ++dbg_single;
        const long long d0 = (DIAG && timeline) ? clock64() : 0;
        long long d1 = d0;
        if (freeing) {
          ++dbg_steps;
          pt.x = FP<F>::mad(dir.x, p.block_length.d, pt.x);
          pt.y = FP<F>::mad(dir.y, p.block_length.d, pt.y);
          pt.z = FP<F>::mad(dir.z, p.block_length.d, pt.z);
          after_step();
          freeing = false;
          if (state == ST_MARCH) {
            block_of(pt, bx, by, bz);
            if (DIAG && timeline) d1 = clock64() + (bx & 0);
            freeing = !space_bit(p, bx, by, bz);
          }
        }
        if (DIAG && timeline) {
          const long long d2 = clock64() + (freeing ? 0 : 0);
          dbg_c1 += __reduce_max_sync(0xffffffffu, (int)(d1 - d0));
          dbg_c2 += (int)(d2 - d0);
        }
      }
    }
    if (timeline) { const long long c = clock64(); c_free += c - c_t; c_t = c; }
    if (state != ST_DONE && !parked) {
      const int data = find_data<PEERS>(p, cache, bx, by, bz);
      if (data != -1) {
        float wx, wy, wz;
        divc3(FP<F>::nmad((float)bx, p.block_length.d, pt.x), FP<F>::nmad((float)by, p.block_length.d, pt.y),
              FP<F>::nmad((float)bz, p.block_length.d, pt.z), p.voxel_length, wx, wy, wz);
        bool want = true;
        float sdf = 0.f;
        if (state == ST_MARCH) {
          const int vx = (int)wx, vy = (int)wy, vz = (int)wz;  // Q22: may be 8 on a face
          sdf = load_distance<PEERS, SOA>(p, data, vz * 64 + vy * 8 + vx);
          want = (sdf <= 0.1f && sdf >= -0.5f);
        }
        if (want) {
          parked = true;
          kbx = bx; kby = by; kbz = bz; kdata = data;
          kwx = wx; kwy = wy; kwz = wz;
        } else if (advance(sdf, bx, by, bz)) {
          after_step();
        }
      } else if (state == ST_MARCH) {
        pt.x = FP<F>::mad(dir.x, p.block_length.d, pt.x);
        pt.y = FP<F>::mad(dir.y, p.block_length.d, pt.y);
        pt.z = FP<F>::mad(dir.z, p.block_length.d, pt.z);
        after_step();
      } else {
        // REFINE without a block under the backed-up point: no second interpolation
        final_depth = xform1_z<F>(Tcw, pt.x, pt.y, pt.z);
        state = ST_DONE;
      }
    }
    // ---- phase B: interpolation for the parked lanes, batched ------------------------------
    if (timeline) { const long long c = clock64(); c_a += c - c_t; c_t = c; }
    const unsigned pm = __ballot_sync(0xffffffffu, parked);
    if (pm) {
      const unsigned fm = __ballot_sync(0xffffffffu, state != ST_DONE && !parked);
      if (fm == 0 || __popc(pm) >= p.batch_lanes || waited >= p.batch_wait) {
        if (parked) {
          Taps taps;
          resolve_taps<PEERS>(p, cache, kbx, kby, kbz, kdata, kwx, kwy, kwz, taps);
          const float sdf = interp_distance<F, PEERS, SOA>(p, taps);
          ipt = pt;
          has_interp = true;
          parked = false;
          if (advance(sdf, kbx, kby, kbz)) after_step();
        }
        waited = 0;
      } else {
        ++waited;
      }
    }
    if (timeline) { const long long c = clock64(); c_b += c - c_t; c_t = c; }
  }

  if (timeline) {
    dbg_steps = __reduce_max_sync(0xffffffffu, dbg_steps);
    dbg_rounds = __reduce_max_sync(0xffffffffu, dbg_rounds);
  }
  if (timeline && lane == 0) {
    unsigned long long t_end;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_end));
    const size_t wid = ((size_t)blockIdx.y * gridDim.x + blockIdx.x) * wpc + warp;
    unsigned smid;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    timeline[8 * wid] = t_start;
    timeline[8 * wid + 1] = t_end;
    timeline[8 * wid + 2] = ((unsigned long long)smid << 32) | (unsigned)dbg_trips;
    timeline[8 * wid + 3] = (unsigned long long)(clock64() - c_start) | ((unsigned long long)dbg_alive_sum << 40);
    timeline[8 * wid + 4] = (unsigned long long)c_free | ((unsigned long long)c_la << 40) ;
    timeline[8 * wid + 5] = (unsigned long long)c_a | ((unsigned long long)(dbg_c1 >> 4) << 40);
    timeline[8 * wid + 6] = (unsigned long long)c_b | ((unsigned long long)(dbg_c2 >> 4) << 40);
    timeline[8 * wid + 7] = ((unsigned long long)(dbg_pair & 0xffff) << 48) | ((unsigned long long)(dbg_single & 0xffff) << 32) |
                            ((unsigned long long)(dbg_rounds & 0xffff) << 16) | (unsigned)(dbg_steps & 0xffff);
  }
  if (!inside) return;
  // colour of the last interpolation (Q13: kept even on a miss), evaluated once per ray
  Vec3 color = {0.f, 0.f, 0.f};
  if (red) {
    color.x = 1.f;
  } else if (has_interp && p.with_color) {
    const int bx = (int)floorf(divc(ipt.x, p.block_length));
    const int by = (int)floorf(divc(ipt.y, p.block_length));
    const int bz = (int)floorf(divc(ipt.z, p.block_length));
    const int data = find_data<PEERS>(p, cache, bx, by, bz);
    const float wx = divc(FP<F>::nmad((float)bx, p.block_length.d, ipt.x), p.voxel_length);
    const float wy = divc(FP<F>::nmad((float)by, p.block_length.d, ipt.y), p.voxel_length);
    const float wz = divc(FP<F>::nmad((float)bz, p.block_length.d, ipt.z), p.voxel_length);
    Taps taps;
    resolve_taps<PEERS>(p, cache, bx, by, bz, data, wx, wy, wz, taps);
    color = interp_color<F, PEERS>(p, taps);
  }

  const int pixel = y * p.w + x;
  if (KEYS) {
    unsigned long long key = ((unsigned long long)0x7f800000u << 32);  // +inf, no payload
    if (owned_hit && final_depth > 0.0f) {
      const uint32_t r8 = (uint32_t)fminf(fmaxf(color.x * 255.0f + 0.5f, 0.f), 255.f);
      const uint32_t g8 = (uint32_t)fminf(fmaxf(color.y * 255.0f + 0.5f, 0.f), 255.f);
      const uint32_t b8 = (uint32_t)fminf(fmaxf(color.z * 255.0f + 0.5f, 0.f), 255.f);
      key = ((unsigned long long)__float_as_uint(final_depth) << 32) | ((unsigned long long)r8 << 24) |
            ((unsigned long long)g8 << 16) | ((unsigned long long)b8 << 8) | (unsigned long long)(p.shard_rank & 0xff);
    }
    p.keys[pixel] = key;
  } else {
    p.depths[pixel] = final_depth;
    p.colors[3 * (size_t)pixel + 0] = color.x;
    p.colors[3 * (size_t)pixel + 1] = color.y;
    p.colors[3 * (size_t)pixel + 2] = color.z;
  }
}

__global__ void __launch_bounds__(256) vk_resolve_keys_kernel(const unsigned long long* keys, int n, float* depths,
                                                              float* colors) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const unsigned long long k = keys[i];
  const uint32_t hi = (uint32_t)(k >> 32);
  const bool hit = hi < 0x7f800000u;
  depths[i] = hit ? __uint_as_float(hi) : 0.0f;
  const float s = 1.0f / 255.0f;
  colors[3 * (size_t)i + 0] = hit ? (float)((k >> 24) & 0xff) * s : 0.f;
  colors[3 * (size_t)i + 1] = hit ? (float)((k >> 16) & 0xff) * s : 0.f;
  colors[3 * (size_t)i + 2] = hit ? (float)((k >> 8) & 0xff) * s : 0.f;
}

}  // namespace vk

// ---- image-space kernels ------------------------------------------------------------

namespace vk {

// normal of pixel (x, y) from a shared-memory depth tile with a 2-pixel margin
// (ComputeNormalsKernel<16>, src/frame.cu:7-122)
template <bool F>
__device__ __forceinline__ Vec3 normal_from_tile(const float* tile, int stride, int tx, int ty, int x, int y,
                                                 const vkProjection& proj) {
  constexpr int PAD = 2;
  const float depth = tile[ty * stride + tx];
  Vec3 out = {0.f, 0.f, 0.f};
  if (depth > 0) {
    Unprojector un;
    un.init(proj);
    const Vec3 z0 = {mul(un.x<F>(add((float)x, 0.5f)), depth), mul(un.y<F>(add((float)y, 0.5f)), depth), depth};
    Vec3 q[4];
    const int ox[4] = {-PAD, +PAD, 0, 0}, oy[4] = {0, 0, -PAD, +PAD};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float d = tile[(ty + oy[k]) * stride + tx + ox[k]];
      if (d == 0) {
        q[k] = z0;
      } else {
        q[k].x = mul(un.x<F>(add((float)(x + ox[k]), 0.5f)), d);
        q[k].y = mul(un.y<F>(add((float)(y + oy[k]), 0.5f)), d);
        q[k].z = d;
      }
    }
    const Vec3 dx = {sub(q[0].x, q[1].x), sub(q[0].y, q[1].y), sub(q[0].z, q[1].z)};
    const Vec3 dy = {sub(q[2].x, q[3].x), sub(q[2].y, q[3].y), sub(q[2].z, q[3].z)};
    if (dot3<F>(dx, dx) > 0 && dot3<F>(dy, dy) > 0) {
      Vec3 n;  // dy.Cross(dx)
      n.x = FP<F>::msub(dy.y, dx.z, mul(dy.z, dx.y));
      n.y = FP<F>::msub(dy.z, dx.x, mul(dy.x, dx.z));
      n.z = FP<F>::msub(dy.x, dx.y, mul(dy.y, dx.x));
      out = normalized<F>(n);
    }
  }
  return out;
}

// ComputeNormalsKernel<16> (src/frame.cu:7-122): 32x8 tile + 2 px halo, zero outside.
// FMA map: Unproject as in vk_math.cuh; dot FFMA chains; cross a.y*b.z - a.z*b.y ->
// FMUL a.z*b.y ; FFMA a.y*b.z - .   (include/vulcan/matrix.h:174-183)
template <bool F>
__global__ void __launch_bounds__(256) vk_normals_kernel(const float* depths, const vkProjection proj,
                                                         float* normals, int w, int h, int y_begin, int y_end,
                                                         const int* bands, int band_rank) {
  pdl_wait();     // the previous kernel of the chain has completed (no-op for a plain launch)
  pdl_trigger();  // the next one may be scheduled as this grid drains
  constexpr int TW = 32, TH = 8, PAD = 2, SW = TW + 2 * PAD, SH = TH + 2 * PAD;
  __shared__ float tile[SH * SW];
  const int x0 = blockIdx.x * TW, y0 = y_begin + blockIdx.y * TH;
  int y_lo = y_begin;
  if (bands) {  // sharded volume: the grid covers the whole image, this rank writes the rows of its band
    y_lo = bands[band_rank];
    y_end = bands[band_rank + 1];
    if (y0 >= y_end || y0 + TH <= y_lo) return;
  }
  for (int s = threadIdx.x; s < SH * SW; s += TW * TH) {
    const int gx = x0 - PAD + (s % SW), gy = y0 - PAD + (s / SW);
    float d = 0;
    if (gx >= 0 && gx < w && gy >= 0 && gy < h) d = __ldg(depths + gy * w + gx);
    tile[s] = d;
  }
  __syncthreads();
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int x = x0 + tx, y = y0 + ty;
  if (x >= w || y >= y_end || y < y_lo) return;
  const Vec3 n = normal_from_tile<F>(tile, SW, tx + PAD, ty + PAD, x, y, proj);
  const float nx = n.x, ny = n.y, nz = n.z;
  const size_t o = 3 * ((size_t)y * w + x);
  normals[o] = nx; normals[o + 1] = ny; normals[o + 2] = nz;
}

// FilterDepthsKernel<16> (src/frame.cu:124-181): 7x7 bilateral, sigma_r^2 = 0.0004.
// FMA map (reference release SASS): only dn += ww * dk contracts; the spatial weight
// expf(-(i*i + j*j) / 9) is a compile-time constant in both builds.
template <bool F>
__device__ __forceinline__ float bilateral7(const float* tile, int stride, int tx, int ty) {
  constexpr int PAD = 3;
  const float d0 = tile[ty * stride + tx];
  float dn = 0, wsum = 0;
#pragma unroll
  for (int i = -PAD; i <= PAD; ++i)
#pragma unroll
    for (int j = -PAD; j <= PAD; ++j) {
      const float dk = tile[(ty + i) * stride + tx + j];
      const float delta = sub(d0, dk);
      // Vector2f(i, j).SquaredNorm() / (pad * pad): float / int -> true division
      const float wr = expf(-(float)(i * i + j * j) / (float)(PAD * PAD));
      const float ws = expf(fdiv(-mul(delta, delta), 0.0004f));
      const float ww = mul(wr, ws);
      dn = FP<F>::mad(ww, dk, dn);
      wsum = add(wsum, ww);
    }
  return fdiv(dn, wsum);
}

template <bool F>
__global__ void __launch_bounds__(256) vk_filter_depths_kernel(int w, int h, const float* src, float* dst) {
  constexpr int TW = 32, TH = 8, PAD = 3, SW = TW + 2 * PAD, SH = TH + 2 * PAD;
  __shared__ float tile[SH * SW];
  const int x0 = blockIdx.x * TW, y0 = blockIdx.y * TH;
  for (int s = threadIdx.x; s < SH * SW; s += TW * TH) {
    const int gx = x0 - PAD + (s % SW), gy = y0 - PAD + (s / SW);
    float d = 0;
    if (gx >= 0 && gx < w && gy >= 0 && gy < h) d = __ldg(src + gy * w + gx);
    tile[s] = d;
  }
  __syncthreads();
  const int tx = (threadIdx.x & 31) + PAD, ty = (threadIdx.x >> 5) + PAD;
  const int x = x0 + tx - PAD, y = y0 + ty - PAD;
  if (x >= w || y >= h) return;
  dst[y * w + x] = bilateral7<F>(tile, SW, tx, ty);
}

// Input-side pre-pass of a sensor frame in ONE kernel: the bilateral filter of
// Frame::FilterDepths followed by Frame::ComputeNormals on the filtered depth.  The raw depth
// tile is read once with a 5-pixel margin, the filtered depth is kept in shared memory for the
// tile plus the 2-pixel margin the normals need (0 outside the image, like the separate kernels
// see it), and both images are written.  Results are bit-identical to the two separate calls.
template <bool F>
__global__ void __launch_bounds__(256) vk_prepare_frame_kernel(int w, int h, const float* src, const vkProjection proj,
                                                               float* dst_depth, float* dst_normals) {
  constexpr int TW = 32, TH = 8, P1 = 3, P2 = 2, PAD = P1 + P2;
  constexpr int RW = TW + 2 * PAD, RH = TH + 2 * PAD;  // raw tile
  constexpr int FW = TW + 2 * P2, FH = TH + 2 * P2;    // filtered tile
  __shared__ float raw[RH * RW];
  __shared__ float fil[FH * FW];
  const int x0 = blockIdx.x * TW, y0 = blockIdx.y * TH;
  for (int s = threadIdx.x; s < RH * RW; s += TW * TH) {
    const int gx = x0 - PAD + (s % RW), gy = y0 - PAD + (s / RW);
    float d = 0;
    if (gx >= 0 && gx < w && gy >= 0 && gy < h) d = __ldg(src + gy * w + gx);
    raw[s] = d;
  }
  __syncthreads();
  for (int s = threadIdx.x; s < FH * FW; s += TW * TH) {
    const int fx = s % FW, fy = s / FW;
    const int gx = x0 - P2 + fx, gy = y0 - P2 + fy;
    float d = 0;
    if (gx >= 0 && gx < w && gy >= 0 && gy < h) d = bilateral7<F>(raw, RW, fx + P1, fy + P1);
    fil[s] = d;
  }
  __syncthreads();
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int x = x0 + tx, y = y0 + ty;
  if (x >= w || y >= h) return;
  dst_depth[y * w + x] = fil[(ty + P2) * FW + tx + P2];
  const Vec3 n = normal_from_tile<F>(fil, FW, tx + P2, ty + P2, x, y, proj);
  const size_t o = 3 * ((size_t)y * w + x);
  dst_normals[o] = n.x; dst_normals[o + 1] = n.y; dst_normals[o + 2] = n.z;
}

}  // namespace vk

// ---- C ABI --------------------------------------------------------------------------

using namespace vk;

// diagnostics hook (tools/tune_bench.py --timeline): device buffer of 2 x u64 per ray-march warp
static std::atomic<unsigned long long*> g_trace_timeline{nullptr};
extern "C" void vk_debug_trace_timeline(unsigned long long* device_buffer) { g_trace_timeline.store(device_buffer); }

static int reset_bounds(float* bounds, int count, float hi_init, cudaStream_t s) {
  if (!bounds || count < 0) return VK_E_INVALID;
  if (count == 0) return 0;
  launch_chained(vk_reset_bounds_kernel, dim3((count + 255) / 256), dim3(256), 0, s, reinterpret_cast<float2*>(bounds), count,
                 hi_init);
  VK_COUNT_LAUNCH();
  return (int)cudaGetLastError();
}

extern "C" int vk_reset_bounds(float* bounds, int count, void* stream) {
  return reset_bounds(bounds, count, -FLT_MAX, (cudaStream_t)stream);
}

static void fill_bounds_params(BoundsParams& b, const float* Tcw_colmajor, const vkProjection* proj,
                               float block_length, float min_depth, float max_depth, int image_w, int image_h,
                               int bw, int bh, float* bounds, cudaStream_t s) {
  rows34(Tcw_colmajor, b.Tcw);
  b.proj = *proj;
  b.block_length = block_length;
  b.min_depth = min_depth;
  b.max_depth = max_depth;
  b.image_w = make_div((float)image_w, s);
  b.image_h = make_div((float)image_h, s);
  b.bw = bw; b.bh = bh;
  b.bounds = bounds;
  b.visible = nullptr; b.count = nullptr; b.entries = nullptr;
  b.warm = nullptr; b.warm_lines = 0;
  b.warm_sdf = nullptr;
  b.negate_hi = false;
  b.has_x = 0; b.image_rows = image_h;
  memset(&b.x, 0, sizeof(b.x));
}

extern "C" int vk_compute_patches(const int* indices, const vkHashEntry* entries, const vkTransform* Tcw,
                                  const vkProjection* proj, float block_length, float min_depth,
                                  float max_depth, int block_count, int image_w, int image_h, int bounds_w,
                                  int bounds_h, vkPatch* patches, int* patch_count, void* stream) {
  if (!indices || !entries || !Tcw || !proj || !patches || !patch_count || block_count < 0) return VK_E_INVALID;
  if (block_count == 0) return 0;
  cudaStream_t s = (cudaStream_t)stream;
  BoundsParams b;
  fill_bounds_params(b, Tcw->m, proj, block_length, min_depth, max_depth, image_w, image_h, bounds_w, bounds_h,
                     nullptr, s);
  const int grid = (block_count + 255) / 256;
  if (g_default_math_mode.load() == VK_MATH_STRICT)
    vk_patches_kernel<false><<<grid, 256, 0, s>>>(indices, entries, b, block_count, patches, patch_count);
  else
    vk_patches_kernel<true><<<grid, 256, 0, s>>>(indices, entries, b, block_count, patches, patch_count);
  VK_COUNT_LAUNCH();
  return (int)cudaGetLastError();
}

extern "C" int vk_compute_bounds(const vkPatch* patches, float* bounds, int bounds_w, int patch_count,
                                 void* stream) {
  if (!patches || !bounds || patch_count < 0) return VK_E_INVALID;
  if (patch_count == 0) return 0;
  vk_patch_bounds_kernel<<<(patch_count + 255) / 256, 256, 0, (cudaStream_t)stream>>>(patches, bounds, bounds_w,
                                                                                      patch_count);
  VK_COUNT_LAUNCH();
  return (int)cudaGetLastError();
}

static void fill_trace_params(TraceParams& t, const vkHashEntry* entries, const vkVoxel* voxels,
                              const float* bounds, const uint32_t* bucket_bits, int K, float block_length, float voxel_length, float trunc,
                              const vkTransform* Twc, const vkProjection* proj, int w, int h, int bw, int bh,
                              cudaStream_t s) {
  t.entries = entries;
  t.voxels = voxels;
  t.sdf = nullptr;
  t.bounds = bounds;
  t.bucket_bits = bucket_bits;
  t.space_bits = nullptr;
  t.free_run = 0;
  t.coop_lanes = g_tune[VK_TUNE_RAY_COOP_LANES].load();
  t.coop_rounds = g_tune[VK_TUNE_RAY_COOP_ROUNDS].load();
  t.pair_after = g_tune[VK_TUNE_RAY_PAIR_AFTER].load();
  t.negated_hi = false;
  t.K = make_mod((uint32_t)K);
  t.block_length = make_div(block_length, s);
  t.voxel_length = make_div(voxel_length, s);
  t.trunc = trunc;
  rows34(Twc->m, t.Twc);
  rows34(Twc->inv, t.Tcw);
  t.proj = *proj;
  t.depths = nullptr; t.colors = nullptr; t.keys = nullptr;
  t.w = w; t.h = h; t.bw = bw; t.bh = bh;
  t.shard_rank = 0; t.shard_count = 1; t.shard_shift = 0;
  t.peer_count = 1; t.peer_rank = 0;
  for (int i = 0; i < VK_MAX_SHARDS; ++i) { t.peer_entries[i] = entries; t.peer_voxels[i] = voxels; t.peer_sdf[i] = nullptr; }
  t.y_begin = 0; t.y_end = h;
  t.bands = nullptr; t.band_rank = 0;
  t.with_color = 1;
  t.batch_lanes = g_tune[VK_TUNE_RAY_BATCH_LANES].load();
  t.batch_wait = g_tune[VK_TUNE_RAY_BATCH_WAIT].load();
  t.debug_only_x = g_tune[6].load() & 0x3ff;
  t.debug_only_y = (g_tune[6].load() >> 10) & 0x3ff;
  t.debug_flags = g_tune[6].load() >> 20;
  t.timeline = g_trace_timeline.load();
}

template <bool KEYS>
static int launch_raycast(const TraceParams& t, int math_mode, cudaStream_t s, bool sharded = false) {
  const int threads = (g_tune[VK_TUNE_RAY_CTA_THREADS].load() == 64) ? 64 : 32;
  const int rows = 4 * (threads / 32);
  if (t.y_end <= t.y_begin) return 0;
  dim3 grid((t.w + 7) / 8, (t.y_end - t.y_begin + rows - 1) / rows);
  const bool strict = (math_mode == VK_MATH_STRICT);
  // kernels of a volume (t.sdf set) sample the distance mirror; raw-pointer calls read the voxels
#define VK_RAY(K, P, D, S)                                                                                    \
  (strict ? launch_chained(vk_raycast_kernel<false, K, P, D, S>, grid, dim3(threads), 0, s, t)                \
          : launch_chained(vk_raycast_kernel<true, K, P, D, S>, grid, dim3(threads), 0, s, t))
  if (!t.sdf) {
    if (KEYS || t.peer_count > 1) return VK_E_INVALID;
    VK_RAY(false, false, false, false);
  } else if (!KEYS && (t.peer_count > 1 || sharded)) {  // sharded volume: lookups go to the owner's table (or its replica)
    VK_RAY(false, true, false, true);
  } else if (!KEYS && t.timeline) {        // diagnostics build of the plain kernel
    VK_RAY(false, false, true, true);
  } else {
    VK_RAY(KEYS, false, false, true);
  }
#undef VK_RAY
  VK_COUNT_LAUNCH();
  return (int)cudaGetLastError();
}

static int compute_points_mode(const vkHashEntry* entries, const vkVoxel* voxels, const float* sdf, const float* bounds,
                               const uint32_t* bucket_bits, const uint32_t* space_bits, int block_count, float block_length,
                               float voxel_length, float trunc_length,
                               const vkTransform* Twc, const vkProjection* proj, float* depths, float* colors,
                               int image_w, int image_h, int bounds_w, int bounds_h, int math_mode, cudaStream_t s,
                               int with_color = 1) {
  if (!entries || !voxels || !bounds || !Twc || !proj || !depths || !colors || image_w <= 0 || image_h <= 0)
    return VK_E_INVALID;
  TraceParams t;
  fill_trace_params(t, entries, voxels, bounds, bucket_bits, block_count, block_length, voxel_length, trunc_length,
                    Twc, proj, image_w, image_h, bounds_w, bounds_h, s);
  t.depths = depths;
  t.colors = colors;
  t.with_color = with_color;
  t.sdf = g_tune[VK_TUNE_RAY_SOA].load() ? sdf : nullptr;   // A/B knob: 0 = sample the 20-byte voxels
  if (space_bits) {
    t.space_bits = space_bits;
    t.free_run = g_tune[VK_TUNE_RAY_FREE_RUN].load();
  }
  return launch_raycast<false>(t, math_mode, s);
}

extern "C" int vk_compute_points(const vkHashEntry* entries, const vkVoxel* voxels, const float* bounds,
                                 int block_count, float block_length, float voxel_length, float trunc_length,
                                 const vkTransform* Twc, const vkProjection* proj, float* depths, float* colors,
                                 int image_w, int image_h, int bounds_w, int bounds_h, void* stream) {
  return compute_points_mode(entries, voxels, nullptr, bounds, nullptr, nullptr, block_count, block_length, voxel_length, trunc_length, Twc,
                             proj, depths, colors, image_w, image_h, bounds_w, bounds_h,
                             g_default_math_mode.load(), (cudaStream_t)stream);
}

static int normals_mode(const float* depths, const vkProjection* proj, float* normals, int image_w, int image_h,
                        int math_mode, cudaStream_t s, int y_begin = 0, int y_end = -1, const int* bands = nullptr,
                        int band_rank = 0) {
  if (!depths || !proj || !normals || image_w <= 0 || image_h <= 0) return VK_E_INVALID;
  if (y_end < 0) y_end = image_h;
  if (y_end <= y_begin) return 0;
  dim3 grid((image_w + 31) / 32, (y_end - y_begin + 7) / 8);
  if (math_mode == VK_MATH_STRICT)
    launch_chained(vk_normals_kernel<false>, grid, dim3(256), 0, s, depths, *proj, normals, image_w, image_h, y_begin, y_end,
                   bands, band_rank);
  else
    launch_chained(vk_normals_kernel<true>, grid, dim3(256), 0, s, depths, *proj, normals, image_w, image_h, y_begin, y_end,
                   bands, band_rank);
  VK_COUNT_LAUNCH();
  return (int)cudaGetLastError();
}

extern "C" int vk_compute_normals(const float* depths, const vkProjection* proj, float* normals, int image_w,
                                  int image_h, void* stream) {
  return normals_mode(depths, proj, normals, image_w, image_h, g_default_math_mode.load(), (cudaStream_t)stream);
}

extern "C" int vk_filter_depths(int image_w, int image_h, const float* src, float* dst, void* stream) {
  if (!src || !dst || image_w <= 0 || image_h <= 0) return VK_E_INVALID;
  dim3 grid((image_w + 31) / 32, (image_h + 7) / 8);
  if (g_default_math_mode.load() == VK_MATH_STRICT)
    vk_filter_depths_kernel<false><<<grid, 256, 0, (cudaStream_t)stream>>>(image_w, image_h, src, dst);
  else
    vk_filter_depths_kernel<true><<<grid, 256, 0, (cudaStream_t)stream>>>(image_w, image_h, src, dst);
  VK_COUNT_LAUNCH();
  return (int)cudaGetLastError();
}

extern "C" int vk_prepare_frame(int image_w, int image_h, const float* src, const vkProjection* proj,
                                float* dst_depth, float* dst_normals, void* stream) {
  if (!src || !proj || !dst_depth || !dst_normals || src == dst_depth || image_w <= 0 || image_h <= 0)
    return VK_E_INVALID;
  dim3 grid((image_w + 31) / 32, (image_h + 7) / 8);
  if (g_default_math_mode.load() == VK_MATH_STRICT)
    vk_prepare_frame_kernel<false><<<grid, 256, 0, (cudaStream_t)stream>>>(image_w, image_h, src, *proj, dst_depth,
                                                                           dst_normals);
  else
    vk_prepare_frame_kernel<true><<<grid, 256, 0, (cudaStream_t)stream>>>(image_w, image_h, src, *proj, dst_depth,
                                                                          dst_normals);
  VK_COUNT_LAUNCH();
  return (int)cudaGetLastError();
}

static const int kBoundsGrid = 32;

static int trace_bounds(const vkVolume* v, const vkFrame* view, float min_depth, float max_depth, float* bounds,
                        cudaStream_t s, bool negate_hi = false, const Xchg* x = nullptr) {
  // Tracer::ComputePatches + ComputeBounds (src/tracer.cpp:49-76), 80x60 grid hard coded
  int rc = reset_bounds(bounds, VK_BOUNDS_W * VK_BOUNDS_H, negate_hi ? +FLT_MAX : -FLT_MAX, s);
  if (rc) return rc;
  BoundsParams b;
  fill_bounds_params(b, view->depth_to_world.inv, &view->depth_proj, VK_BLOCK_RES * v->voxel_length, min_depth,
                     max_depth, view->depth_w, view->depth_h, VK_BOUNDS_W, VK_BOUNDS_H, bounds, s);
  b.visible = v->visible[v->visible_slot];
  b.count = v->counters + VK_CNT_VISIBLE0 + v->visible_slot;
  b.entries = v->entries;
  b.warm = v->space_bits;
  b.warm_lines = VK_SPACE_WORDS / 32;
  b.warm_sdf = g_tune[VK_TUNE_RAY_WARM_SDF].load() ? v->sdf : nullptr;
  b.negate_hi = negate_hi;
  if (x) { b.has_x = 1; b.x = *x; }
  const size_t smem = (size_t)VK_BOUNDS_W * VK_BOUNDS_H * 2 * sizeof(float) + 64 * sizeof(int);
  if (v->math_mode == VK_MATH_STRICT) launch_chained(vk_bounds_kernel<false>, dim3(kBoundsGrid), dim3(512), smem, s, b);
  else launch_chained(vk_bounds_kernel<true>, dim3(kBoundsGrid), dim3(512), smem, s, b);
  VK_COUNT_LAUNCH();
  return (int)cudaGetLastError();
}

// end of a ray cast on stream s: from here on the bounds scratch may be rewritten (see vk_trace)
static int trace_done(const vkVolume* v, cudaStream_t s) {
  if (!v->ev_trace || stream_capturing(s)) return 0;
  return (int)cudaEventRecord((cudaEvent_t)v->ev_trace, s);
}

extern "C" int vk_trace(const vkVolume* v, const vkFrame* view, float min_depth, float max_depth, float* bounds,
                        float* out_depth, float* out_color, float* out_normal, void* stream) {
  if (!v || !view || !v->voxels || !bounds || !out_depth || !out_color) return VK_E_INVALID;
  cudaStream_t s = (cudaStream_t)stream;
  // The ray bounds depend on the view's visible list (vk_set_view) and the pose only, not on the voxels: they are
  // computed on the volume's side stream, where they overlap whatever the caller's stream is still doing
  // (normally this frame's integration), and the ray march joins them.  ev_trace keeps the bounds scratch of the
  // previous ray cast safe.
  const bool capturing = stream_capturing(s);
  const bool fork = !capturing && v->aux_valid && v->aux_stream && g_tune[VK_TUNE_TRACE_FORK].load() != 0;
  cudaStream_t bs = s;
  if (fork) {
    bs = (cudaStream_t)v->aux_stream;
    cudaError_t e = cudaStreamWaitEvent(bs, (cudaEvent_t)v->ev_view, 0);
    if (e == cudaSuccess) e = cudaStreamWaitEvent(bs, (cudaEvent_t)v->ev_trace, 0);
    if (e != cudaSuccess) return (int)e;
  }
  int rc = trace_bounds(v, view, min_depth, max_depth, bounds, bs);
  if (rc) return rc;
  if (fork) {
    cudaError_t e = cudaEventRecord((cudaEvent_t)v->ev_bounds, bs);
    if (e == cudaSuccess) e = cudaStreamWaitEvent(s, (cudaEvent_t)v->ev_bounds, 0);
    if (e != cudaSuccess) return (int)e;
  }
  rc = compute_points_mode(v->entries, v->voxels, v->sdf, bounds, v->bucket_bits, v->space_bits, v->main_blocks, VK_BLOCK_RES * v->voxel_length,
                           v->voxel_length, v->trunc_length, &view->depth_to_world, &view->depth_proj, out_depth,
                           out_color, view->depth_w, view->depth_h, VK_BOUNDS_W, VK_BOUNDS_H, v->math_mode, s,
                           v->has_color);
  if (rc) return rc;
  if (out_normal)
    rc = normals_mode(out_depth, &view->depth_proj, out_normal, view->depth_w, view->depth_h, v->math_mode, s);
  if (rc) return rc;
  return trace_done(v, s);
}

extern "C" int vk_trace_keys(const vkVolume* v, const vkFrame* view, float min_depth, float max_depth,
                             float* bounds, unsigned long long* out_keys, void* stream) {
  if (!v || !view || !v->voxels || !bounds || !out_keys) return VK_E_INVALID;
  cudaStream_t s = (cudaStream_t)stream;
  int rc = trace_bounds(v, view, min_depth, max_depth, bounds, s);
  if (rc) return rc;
  TraceParams t;
  fill_trace_params(t, v->entries, v->voxels, bounds, v->bucket_bits, v->main_blocks, VK_BLOCK_RES * v->voxel_length,
                    v->voxel_length, v->trunc_length, &view->depth_to_world, &view->depth_proj, view->depth_w,
                    view->depth_h, VK_BOUNDS_W, VK_BOUNDS_H, s);
  t.keys = out_keys;
  t.sdf = v->sdf;
  t.space_bits = v->space_bits;
  t.free_run = g_tune[VK_TUNE_RAY_FREE_RUN].load();
  t.shard_rank = v->shard_rank;
  t.shard_count = v->shard_count;
  t.shard_shift = v->shard_shift;
  rc = launch_raycast<true>(t, v->math_mode, s);
  if (rc) return rc;
  return trace_done(v, s);
}

extern "C" int vk_resolve_keys(const unsigned long long* keys, const vkFrame* view, float* out_depth,
                               float* out_color, float* out_normal, void* stream) {
  if (!keys || !view || !out_depth || !out_color) return VK_E_INVALID;
  const int n = view->depth_w * view->depth_h;
  vk_resolve_keys_kernel<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(keys, n, out_depth, out_color);
  VK_COUNT_LAUNCH();
  int rc = (int)cudaGetLastError();
  if (rc) return rc;
  if (out_normal) rc = vk_compute_normals(out_depth, &view->depth_proj, out_normal, view->depth_w, view->depth_h, stream);
  return rc;
}

// ---- sharded volumes: exact ray casting over peer memory ---------------------------------------

extern "C" int vk_trace_bounds(const vkVolume* v, const vkFrame* view, float min_depth, float max_depth,
                               float* bounds, void* stream) {
  if (!v || !view || !v->voxels || !bounds) return VK_E_INVALID;
  // cells are (min depth, -max depth): see include/vulcan_b200.h
  return trace_bounds(v, view, min_depth, max_depth, bounds, (cudaStream_t)stream, true);
}

extern "C" int vk_trace_rows(const vkVolume* v, const vkPeerTable* peers, const vkFrame* view, const float* bounds,
                             int row_begin, int row_end, float* out_depth, float* out_color, float* out_normal,
                             void* stream) {
  if (!v || !view || !v->voxels || !bounds || !out_depth || !out_color) return VK_E_INVALID;
  if (row_begin < 0 || row_end > view->depth_h || row_begin > row_end) return VK_E_INVALID;
  if (peers && (peers->count < 1 || peers->count > VK_MAX_SHARDS || peers->rank < 0 || peers->rank >= peers->count))
    return VK_E_INVALID;
  // owners are computed from the peer table's rank count and the volume's brick shift: they must
  // describe the same partition
  if (peers && peers->count > 1 && (peers->count != v->shard_count || peers->rank != v->shard_rank)) return VK_E_INVALID;
  cudaStream_t s = (cudaStream_t)stream;
  TraceParams t;
  fill_trace_params(t, v->entries, v->voxels, bounds, v->bucket_bits, v->main_blocks, VK_BLOCK_RES * v->voxel_length,
                    v->voxel_length, v->trunc_length, &view->depth_to_world, &view->depth_proj, view->depth_w,
                    view->depth_h, VK_BOUNDS_W, VK_BOUNDS_H, s);
  t.depths = out_depth;
  t.colors = out_color;
  t.sdf = v->sdf;
  t.negated_hi = true;
  t.space_bits = v->space_bits;
  t.free_run = g_tune[VK_TUNE_RAY_FREE_RUN].load();
  t.shard_shift = v->shard_shift;
  if (peers && peers->count > 1) {
    t.peer_count = peers->count;
    t.peer_rank = peers->rank;
    for (int i = 0; i < peers->count; ++i) {
      if (!peers->entries[i] || !peers->voxels[i] || !peers->sdf[i]) return VK_E_INVALID;
      t.peer_entries[i] = peers->entries[i];
      t.peer_voxels[i] = peers->voxels[i];
      t.peer_sdf[i] = peers->sdf[i];
    }
  }
  // the normals of the band need the depth of two more rows on each side (src/frame.cu:23-32)
  t.y_begin = (row_begin - 2 > 0) ? row_begin - 2 : 0;
  t.y_end = (row_end + 2 < view->depth_h) ? row_end + 2 : view->depth_h;
  if (row_end == row_begin) return 0;
  int rc = launch_raycast<false>(t, v->math_mode, s);
  if (rc) return rc;
  if (out_normal)
    rc = normals_mode(out_depth, &view->depth_proj, out_normal, view->depth_w, view->depth_h, v->math_mode, s,
                      row_begin, row_end);
  if (rc) return rc;
  return trace_done(v, s);
}

// ---- block-sharded volume with replicated lookup structures (vk_shard.cu) --------------------------

namespace vk {

int trace_sharded_bounds(const vkVolume* v, const vkFrame* view, float min_depth, float max_depth, const Xchg& x,
                         cudaStream_t s) {
  // the working grid lives in this rank's exchange buffer, cells as (min depth, -max depth)
  if (x.stages == 4) {
    // One rank per GPU, store-only launch: like vk_trace, the rank's own bounds (and their stores into the peers'
    // slots) depend on vk_set_view only and run on the volume's side stream, beside the integration; the announce +
    // wait + merge launch that follows on `s` is ordered after both.
    const bool fork = !stream_capturing(s) && v->aux_valid && v->aux_stream && g_tune[VK_TUNE_TRACE_FORK].load() != 0;
    cudaStream_t bs = s;
    if (fork) {
      bs = (cudaStream_t)v->aux_stream;
      cudaError_t e = cudaStreamWaitEvent(bs, (cudaEvent_t)v->ev_view, 0);
      if (e == cudaSuccess) e = cudaStreamWaitEvent(bs, (cudaEvent_t)v->ev_trace, 0);
      if (e != cudaSuccess) return (int)e;
    }
    const int rc = trace_bounds(v, view, min_depth, max_depth, xchg_grid(x.buf[x.rank]), bs, true, &x);
    if (rc || !fork) return rc;
    cudaError_t e = cudaEventRecord((cudaEvent_t)v->ev_bounds, bs);
    if (e == cudaSuccess) e = cudaStreamWaitEvent(s, (cudaEvent_t)v->ev_bounds, 0);
    return (int)e;
  }
  if (xchg_stores(x.stages)) return trace_bounds(v, view, min_depth, max_depth, xchg_grid(x.buf[x.rank]), s, true, &x);
  // second half only (the first half ran in an earlier launch)
  BoundsParams b;
  fill_bounds_params(b, view->depth_to_world.inv, &view->depth_proj, VK_BLOCK_RES * v->voxel_length, min_depth, max_depth,
                     view->depth_w, view->depth_h, VK_BOUNDS_W, VK_BOUNDS_H, xchg_grid(x.buf[x.rank]), s);
  b.negate_hi = true;
  b.has_x = 1;
  b.x = x;
  const size_t smem = (size_t)VK_BOUNDS_W * VK_BOUNDS_H * 2 * sizeof(float) + 64 * sizeof(int);
  // "ordered" second launch (bit 3): a plain launch - the full kernel boundary is what orders the first launch's
  // peer stores before the flags; a programmatic dependent launch only promises visibility to this grid
  if (x.stages & 8) vk_bounds_merge_kernel<<<1, 512, smem, s>>>(b);
  else launch_chained(vk_bounds_merge_kernel, dim3(1), dim3(512), smem, s, b);
  VK_COUNT_LAUNCH();
  return (int)cudaGetLastError();
}

int trace_sharded_band(const vkVolume* v, const vkFrame* view, const Xchg& x, float* depth, float* color, float* normal,
                       cudaStream_t s) {
  const vkShard* sh = v->shard;
  TraceParams t;
  fill_trace_params(t, v->entries, v->voxels, xchg_grid(x.buf[x.rank]), v->bucket_bits, v->main_blocks,
                    VK_BLOCK_RES * v->voxel_length, v->voxel_length, v->trunc_length, &view->depth_to_world,
                    &view->depth_proj, view->depth_w, view->depth_h, VK_BOUNDS_W, VK_BOUNDS_H, s);
  t.depths = depth;
  t.colors = color;
  t.sdf = v->sdf;
  t.negated_hi = true;
  t.with_color = v->has_color;
  t.space_bits = v->space_bits;
  t.free_run = g_tune[VK_TUNE_RAY_FREE_RUN].load();
  t.shard_shift = v->shard_shift;
  // every table the march reads is LOCAL: the volume's own for blocks it owns, the replicas otherwise;
  // only the colour taps of a hit go to the owner's voxel pool
  t.peer_count = sh->count;
  t.peer_rank = sh->rank;
  for (int r = 0; r < sh->count; ++r) {
    t.peer_entries[r] = sh->entries[r];
    t.peer_sdf[r] = sh->sdf[r];
    t.peer_voxels[r] = sh->voxels[r];
  }
  t.bands = xchg_bands(x.buf[x.rank], x.parity);
  t.band_rank = x.rank;
  int rc = launch_raycast<false>(t, v->math_mode, s, true);
  if (rc) return rc;
  rc = normals_mode(depth, &view->depth_proj, normal, view->depth_w, view->depth_h, v->math_mode, s, 0, -1, t.bands,
                    t.band_rank);
  if (rc) return rc;
  return trace_done(v, s);   // the working grid may be rewritten from here on
}

}  // namespace vk

extern "C" int vk_ipc_export(const vkVolume* v, vkIpcHandles* out) {
  if (!v || !v->entries || !v->voxels || !v->sdf || !out) return VK_E_INVALID;
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "vkIpcHandles holds CUDA IPC handles");
  cudaError_t err = cudaIpcGetMemHandle(reinterpret_cast<cudaIpcMemHandle_t*>(out->entries), v->entries);
  if (err != cudaSuccess) return (int)err;
  err = cudaIpcGetMemHandle(reinterpret_cast<cudaIpcMemHandle_t*>(out->voxels), v->voxels);
  if (err != cudaSuccess) return (int)err;
  err = cudaIpcGetMemHandle(reinterpret_cast<cudaIpcMemHandle_t*>(out->sdf), v->sdf);
  return (int)err;
}

extern "C" int vk_ipc_open(const vkIpcHandles* handles, const vkHashEntry** entries, const vkVoxel** voxels,
                           const float** sdf) {
  if (!handles || !entries || !voxels || !sdf) return VK_E_INVALID;
  const unsigned char* blobs[3] = {handles->entries, handles->voxels, handles->sdf};
  void* ptrs[3] = {nullptr, nullptr, nullptr};
  for (int i = 0; i < 3; ++i) {
    cudaIpcMemHandle_t h;
    memcpy(&h, blobs[i], sizeof(h));
    const cudaError_t err = cudaIpcOpenMemHandle(&ptrs[i], h, cudaIpcMemLazyEnablePeerAccess);
    if (err != cudaSuccess) {
      for (int j = 0; j < i; ++j) cudaIpcCloseMemHandle(ptrs[j]);
      return (int)err;
    }
  }
  *entries = static_cast<const vkHashEntry*>(ptrs[0]);
  *voxels = static_cast<const vkVoxel*>(ptrs[1]);
  *sdf = static_cast<const float*>(ptrs[2]);
  return 0;
}

extern "C" int vk_ipc_close(const vkHashEntry* entries, const vkVoxel* voxels, const float* sdf) {
  cudaError_t e1 = entries ? cudaIpcCloseMemHandle(const_cast<vkHashEntry*>(entries)) : cudaSuccess;
  cudaError_t e2 = voxels ? cudaIpcCloseMemHandle(const_cast<vkVoxel*>(voxels)) : cudaSuccess;
  cudaError_t e3 = sdf ? cudaIpcCloseMemHandle(const_cast<float*>(sdf)) : cudaSuccess;
  return (int)(e1 != cudaSuccess ? e1 : (e2 != cudaSuccess ? e2 : e3));
}
