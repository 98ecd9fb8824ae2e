// vk_alloc.cu — Volume::SetView rebuilt for sm_100a.
//
// Reference (src/volume.cu:430-437) = thrust::replace over all B visibility bytes
// + per-pixel request kernel (one surviving request per bucket per pass, racy)
// + per-bucket handler + a 20-barrier scan over all B entries + two blocking
// symbol memcpys, and the caller has to repeat the whole thing until allocation
// converges.  Here one call reaches that fixed point with two launches (the second overlaps the
// tail of the first through programmatic dependent launch) and no host synchronisation:
//
//   vk_alloc_kernel    one warp per 8x4 pixel tile walks the truncation band of its
//                      32 rays in lock step; identical block coordinates inside the
//                      warp are merged with __match_any_sync so one lane per distinct
//                      block does a find-or-insert in the hash (per-bucket lock only
//                      on the insert path).  The first lane to touch an entry in
//                      this view (atomicExch on a per-entry view stamp) marks it
//                      VISIBILITY_TRUE and appends it to the new visible list.
//   vk_carry_kernel    walks the *previous* visible list only (the UNKNOWN entries of
//                      the reference are exactly that list): entries not touched in
//                      this view get the 8-corner frustum test (Q6) and are carried
//                      over (stay UNKNOWN, Q5) or dropped (FALSE).  O(V) instead of
//                      O(M+E).
#include "vk_math.cuh"
#include "vk_launch.h"

namespace vk {

struct AllocParams {
  vkHashEntry* entries;
  uint8_t* visibility;
  int32_t* stamps;
  int32_t* locks;
  unsigned long long* lock_tags;   // per bucket: block (and view) its last / current inserter works on
  uint32_t* bucket_bits;
  uint32_t* space_bits;
  const int32_t* free_list;
  int32_t* counters;
  int32_t* visible_out;
  int cnt_slot;
  const float* depth;
  int w, h;
  vkProjection proj;
  float Twd[12];  // rows 0..2 of depth_to_world, row-major 3x4
  Mod K;
  int B;
  float block_length, inv_block_length, trunc, min_depth, max_depth;
  int view_id;
  int shard_rank, shard_count, shard_shift, shard_halo;
  // sharded volume with replicated lookup structures: this rank's hash table as its peers hold it
  // (peer memory); every entry written here is stored there too.  push_count = 0: no replicas.
  int push_count;
  vkHashEntry* push_entries[VK_MAX_SHARDS];
};

// chain walk; returns entry index of block `key` or -1.  `tail` receives the last
// entry visited, `head_empty` whether the bucket head is unallocated.
__device__ __forceinline__ int find_block(const vkHashEntry* entries, uint32_t h,
                                          unsigned long long key, int& tail, bool& head_empty) {
  int idx = (int)h;
  int4 e = ld_entry_cg(entries + idx);
  head_empty = (e.z == -1);
  if (head_empty) { tail = idx; return -1; }
  while (true) {
    if (entry_block(e) == key && e.z != -1) return idx;
    if (e.w == -1) { tail = idx; return -1; }
    idx = e.w;
    e = ld_entry_cg(entries + idx);
  }
}

// New entries of the visible list are collected per CTA in shared memory and appended with ONE atomicAdd on the
// list's counter when the CTA retires (~16 k same-address atomics with a dependent store each
// were serialising in L2).  A CTA that finds more than kNewCap falls back to direct appends for the rest.
constexpr int kNewCap = 384;

__device__ __forceinline__ void touch_entry(const AllocParams& p, int idx, int* s_new, int* s_count) {
  if (__ldcg(p.stamps + idx) != p.view_id) {
    const int old = atomicExch(p.stamps + idx, p.view_id);
    if (old != p.view_id) {
      p.visibility[idx] = VK_VIS_TRUE;
      const int at = atomicAdd(s_count, 1);
      if (at < kNewCap) {
        s_new[at] = idx;
      } else {
        const int pos = atomicAdd(p.counters + p.cnt_slot, 1);
        p.visible_out[pos] = idx;
      }
    }
  }
}

// Sharded volumes: every rank walks every ray, so every rank sees every block of the scene; blocks
// it does not store still go into its occupancy filter, which therefore equals the filter of the
// unsharded volume on all ranks (the ray marcher relies on "bit clear => no such block anywhere").
__device__ __forceinline__ void mark_space(const AllocParams& p, int bx, int by, int bz) {
  uint32_t sw, sb;
  space_bit_index(bx, by, bz, sw, sb);
  if (!((__ldcg(p.space_bits + sw) >> sb) & 1u)) atomicOr(p.space_bits + sw, 1u << sb);
}

__device__ __noinline__ int insert_block(const AllocParams& p, uint32_t h, int bx, int by, int bz,
                                         unsigned long long key) {
  int result = -1;
  bool done = false;
  // what this lane is about to insert, readable by the lanes it keeps out: a ray that finds the bucket
  // locked by an inserter of ITS block (in this view) has nothing left to do - the inserter also marks the
  // block visible - instead of spinning until the lock is free
  const unsigned long long tag = key | ((unsigned long long)(p.view_id & 0xffff) << 48);
  while (!done) {
    if (atomicCAS(p.locks + h, 0, 1) == 0) {
      __stcg(p.lock_tags + h, tag);
      __threadfence();
      int tail;
      bool head_empty;
      int idx = find_block(p.entries, h, key, tail, head_empty);
      if (idx < 0) {
        // HandleAllocationRequestsKernel (src/volume.cu:303-368): MAIN request writes the
        // bucket head, EXCESS request bump-allocates an entry and links it to the tail.
        int slot = tail;
        if (!head_empty) slot = atomicAdd(p.counters + VK_CNT_EXCESS_PTR, 1);
        const int vtop = atomicSub(p.counters + VK_CNT_VOXEL_PTR, 1);
        if (slot < p.B && vtop >= 0) {
          const int data = p.free_list[vtop];
          {  // set before the entry is published: a reader that can find the block sees the bit
            uint32_t sw, sb;
            space_bit_index(bx, by, bz, sw, sb);
            atomicOr(p.space_bits + sw, 1u << sb);
          }
          int4 ne;
          ne.x = (int)((uint32_t)(uint16_t)(int16_t)bx | ((uint32_t)(uint16_t)(int16_t)by << 16));
          ne.y = (int)(uint32_t)(uint16_t)(int16_t)bz;
          ne.w = -1;
          ne.z = data;
          if (head_empty) {
            // publish coordinates first, then the data id that makes the head "allocated"
            int2* dst = reinterpret_cast<int2*>(p.entries + slot);
            __stcg(dst, make_int2(ne.x, ne.y));
            __threadfence();
            __stcg(reinterpret_cast<int*>(p.entries + slot) + 2, data);
            atomicOr(p.bucket_bits + (h >> 5), 1u << (h & 31));
          } else {
            __stcg(reinterpret_cast<int4*>(p.entries + slot), ne);
            __threadfence();
            __stcg(reinterpret_cast<int*>(p.entries + tail) + 3, slot);
          }
          // the peers' copies of this table: same entry, same link (posted stores; the peers read them
          // only after the next ray-bounds handshake, so their order does not matter)
          for (int r = 0; r < p.push_count; ++r) {
            if (r == p.shard_rank) continue;
            vkHashEntry* rep = p.push_entries[r];
            __stcg(reinterpret_cast<int4*>(rep + slot), ne);
            if (!head_empty) __stcg(reinterpret_cast<int*>(rep + tail) + 3, slot);
          }
          atomicAdd(p.counters + VK_CNT_ALLOCATED, 1);
          idx = slot;
        } else {
          atomicAdd(p.counters + VK_CNT_OVERFLOW, 1);
          idx = -1;
        }
      }
      result = idx;
      __threadfence();
      atomicExch(p.locks + h, 0);
      done = true;
    } else {
      // another lane is inserting into this bucket: it may be inserting our block
      if (__ldcg(p.lock_tags + h) == tag) { result = -2; done = true; continue; }
      int tail;
      bool head_empty;
      const int idx = find_block(p.entries, h, key, tail, head_empty);
      if (idx >= 0) { result = idx; done = true; }
    }
  }
  return result;
}

// CreateAllocationRequestsKernel (src/volume.cu:86-301) + HandleAllocationRequestsKernel
// (src/volume.cu:303-368) to their common fixed point.
// FMA map (reference release SASS of CreateAllocationRequestsKernel):
//   dir   = Twd * (unproj, 0)      FMUL y*m1; FFMA x*m0 + .; FADD m2 (z == 1); FFMA 0*m3 + .
//   X     = origin + depth * dir   FFMA dir*depth + origin
//   begin = X - trunc * dir        FFMA -dir*trunc + X     end = FFMA dir*trunc + X
//   tmax  = (L*(b + max(0,step)) - begin) / dir     FFMA bf*L - begin ; IEEE div
template <bool F>
__global__ void __launch_bounds__(256) vk_alloc_kernel(const AllocParams p) {
  __shared__ int s_new[kNewCap];
  __shared__ int s_count, s_base;
  if (threadIdx.x == 0) s_count = 0;
  __syncthreads();
  pdl_wait();     // the previous kernel of the chain has completed (no-op for a plain launch)
  pdl_trigger();  // the next one may be scheduled as this grid drains
  const int lane = threadIdx.x & 31;
  // warp tile 8x4 pixels; the warps of the grid take the tiles row by row
  const int tile = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int tiles_x = (p.w + 7) >> 3;
  const int x = (tile % tiles_x) * 8 + (lane & 7);
  const int y = (tile / tiles_x) * 4 + (lane >> 3);

  bool active = (x < p.w && y < p.h);
  int bx = 0, by = 0, bz = 0, ex = 0, ey = 0, ez = 0, sx = 1, sy = 1, sz = 1;
  float tmx = 0, tmy = 0, tmz = 0, tdx = 0, tdy = 0, tdz = 0;

  if (active) {
    const float depth = __ldg(p.depth + y * p.w + x);
    if (depth < p.min_depth || depth > p.max_depth) active = false;
    if (active) {
      const float u = add((float)x, 0.5f), v = add((float)y, 0.5f);
      Unprojector un;
      un.init(p.proj);
      const float d0x = un.x<F>(u), d0y = un.y<F>(v);
      Mat34 Twd;
      Twd.load(p.Twd);
      Vec3 dir = xform0<F>(Twd, d0x, d0y, 1.0f);
      const float Xx = FP<F>::mad(dir.x, depth, p.Twd[3]);
      const float Xy = FP<F>::mad(dir.y, depth, p.Twd[7]);
      const float Xz = FP<F>::mad(dir.z, depth, p.Twd[11]);
      dir = normalized<F>(dir);
      const float bgx = FP<F>::nmad(dir.x, p.trunc, Xx), bgy = FP<F>::nmad(dir.y, p.trunc, Xy),
                  bgz = FP<F>::nmad(dir.z, p.trunc, Xz);
      const float enx = FP<F>::mad(dir.x, p.trunc, Xx), eny = FP<F>::mad(dir.y, p.trunc, Xy),
                  enz = FP<F>::mad(dir.z, p.trunc, Xz);
      sx = (dir.x < 0) ? -1 : 1;
      sy = (dir.y < 0) ? -1 : 1;
      sz = (dir.z < 0) ? -1 : 1;
      const float inv = p.inv_block_length;
      bx = (int)floorf(mul(bgx, inv));
      by = (int)floorf(mul(bgy, inv));
      bz = (int)floorf(mul(bgz, inv));
      ex = (int)floorf(mul(enx, inv));
      ey = (int)floorf(mul(eny, inv));
      ez = (int)floorf(mul(enz, inv));
      const float ox = FP<F>::msub((float)(bx + max(0, sx)), p.block_length, bgx);
      const float oy = FP<F>::msub((float)(by + max(0, sy)), p.block_length, bgy);
      const float oz = FP<F>::msub((float)(bz + max(0, sz)), p.block_length, bgz);
      tmx = fdiv(ox, dir.x);
      tmy = fdiv(oy, dir.y);
      tmz = fdiv(oz, dir.z);
      if (dir.x == 0) tmx = 1E20f;
      if (dir.y == 0) tmy = 1E20f;
      if (dir.z == 0) tmz = 1E20f;
      tdx = fdiv(mul((float)sx, p.block_length), dir.x);
      tdy = fdiv(mul((float)sy, p.block_length), dir.y);
      tdz = fdiv(mul((float)sz, p.block_length), dir.z);
    }
  }

  unsigned long long last_key = ~0ull;  // block this lane handled last as leader
  while (__any_sync(0xffffffffu, active)) {
    const unsigned long long key = active ? pack_block(bx, by, bz) : ~0ull;
    const unsigned peers = __match_any_sync(0xffffffffu, key);
    const bool leader = active && (lane == (__ffs(peers) - 1)) && key != last_key;
    if (leader) {
      last_key = key;
      const uint32_t h = block_hash(bx, by, bz, p.K);
      // most blocks sit at the head of their bucket, whose index is known before the entry
      // arrives: its stamp is fetched together with the entry instead of after it
      const int head_stamp = __ldcg(p.stamps + h);
      int tail;
      bool head_empty;
      int idx = find_block(p.entries, h, key, tail, head_empty);
      if (idx < 0) {
        if (shard_stores(bx, by, bz, p.shard_shift, p.shard_count, p.shard_rank, p.shard_halo)) idx = insert_block(p, h, bx, by, bz, key);
        else mark_space(p, bx, by, bz);  // exists on another rank: the filter is global
      }
      if (idx >= 0 && !(idx == (int)h && head_stamp == p.view_id)) touch_entry(p, idx, s_new, &s_count);
    }
    if (active) {
      // Amanatides-Woo step with the reference's comparison tree (src/volume.cu:241-295)
      if (tmx < tmy) {
        if (tmx < tmz) { bx += sx; if (bx == ex + sx) active = false; tmx = add(tmx, tdx); }
        else           { bz += sz; if (bz == ez + sz) active = false; tmz = add(tmz, tdz); }
      } else {
        if (tmy < tmz) { by += sy; if (by == ey + sy) active = false; tmy = add(tmy, tdy); }
        else           { bz += sz; if (bz == ez + sz) active = false; tmz = add(tmz, tdz); }
      }
    }
  }
  // the CTA's new entries: one append
  __syncthreads();
  const int n_new = min(s_count, kNewCap);
  if (threadIdx.x == 0 && n_new > 0) s_base = atomicAdd(p.counters + p.cnt_slot, n_new);
  __syncthreads();
  for (int i = threadIdx.x; i < n_new; i += blockDim.x) p.visible_out[s_base + i] = s_new[i];
}

struct CarryParams {
  const vkHashEntry* entries;
  uint8_t* visibility;
  const int32_t* stamps;
  int32_t* counters;
  const int32_t* prev_list;
  int32_t* visible_out;
  int prev_slot, cnt_slot;
  int w, h;
  vkProjection proj;
  float Tdw[12];
  float block_length;
  int view_id;
  int* ticket;
};

// UpdateBlockVisibilityKernel (src/volume.cu:23-84) restricted to the entries that can
// be VISIBILITY_UNKNOWN, i.e. the previous visible list.
template <bool F>
__global__ void __launch_bounds__(256) vk_carry_kernel(const CarryParams p) {
  // The previous list, its entries and the frustum test do not depend on the allocation kernel
  // that precedes this one in the chain: they are evaluated while that kernel drains; only the
  // stamps (is the entry already on the new list?) need it to have completed.
  pdl_trigger();
  bool waited = false;
  const int n_prev = p.counters[p.prev_slot];
  const int lane = threadIdx.x & 31;
  for (int base = blockIdx.x * blockDim.x; base < n_prev; base += gridDim.x * blockDim.x) {
    const int i = base + threadIdx.x;
    bool keep = false;
    int e = -1;
    int4 ent = make_int4(0, 0, 0, 0);
    if (i < n_prev) {
      e = p.prev_list[i];
      ent = ld_entry(p.entries + e);
    }
    if (!waited) { pdl_wait(); waited = true; }
    if (i < n_prev) {
      if (p.stamps[e] != p.view_id) {
        int ox, oy, oz;
        entry_origin(ent, ox, oy, oz);
        Mat34 M;
        M.load(p.Tdw);
        for (int c = 0; c < 8; ++c) {
          const float X = mul(p.block_length, (float)(ox + ((c & 1) >> 0)));
          const float Y = mul(p.block_length, (float)(oy + ((c & 2) >> 1)));
          const float Z = mul(p.block_length, (float)(oz + ((c & 4) >> 2)));
          const Vec3 Xd = xform1<F>(M, X, Y, Z);
          if (Xd.z < 0) continue;
          float u, v;
          project<F>(p.proj, Xd, u, v);
          if (u >= 0 && u <= (float)p.w && v >= 0 && v <= (float)p.h) { keep = true; break; }
        }
        p.visibility[e] = keep ? VK_VIS_UNKNOWN : VK_VIS_FALSE;
      }
    }
    // warp-aggregated append
    const unsigned m = __ballot_sync(0xffffffffu, keep);
    if (m) {
      int pos = 0;
      if (lane == 0) pos = atomicAdd(p.counters + p.cnt_slot, __popc(m));
      pos = __shfl_sync(0xffffffffu, pos, 0);
      if (keep) p.visible_out[pos + __popc(m & ((1u << lane) - 1))] = e;
    }
  }
  if (!waited) pdl_wait();
  // last CTA out: retire the previous list and mirror the counters to the host
  __shared__ bool last;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) last = (atomicAdd(p.ticket, 1) == (int)gridDim.x - 1);
  __syncthreads();
  if (last) {
    __threadfence();
    if (threadIdx.x == 0) {
      p.counters[p.prev_slot] = 0;
      *p.ticket = 0;
    }
  }
}

// pool fills of Volume::Initialize (src/volume.cu:552-627)
__global__ void vk_fill_pools_kernel(uint4* voxels16, size_t n16, float4* sdf4, vkHashEntry* entries, int32_t* free_list,
                                     uint8_t* visibility, int32_t* stamps, int32_t* locks, int32_t* counters, int B, int M) {
  const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  // counters (and the ticket behind them) are initialised HERE, in stream order, not copied from the
  // pinned host mirror: a SetView still in flight on the stream writes that mirror when it retires
  if (tid < VK_CNT_COUNT + 4) {
    int32_t val = 0;
    if (tid == VK_CNT_VOXEL_PTR) val = B - 1;     // src/volume.cu:595-601
    if (tid == VK_CNT_EXCESS_PTR) val = M;        // src/volume.cu:579-584
    counters[tid] = val;
  }
  // Voxel::Empty() = {1.0f, 0,0,0, 0,0}: 5 words; a 16-byte chunk q starts at word 4q.
  for (size_t q = tid; q < n16; q += stride) {
    const uint32_t r = (uint32_t)((q * 4) % 5);
    uint4 val;
    val.x = (r == 0) ? 0x3f800000u : 0u;
    val.y = (r == 4) ? 0x3f800000u : 0u;
    val.z = (r == 3) ? 0x3f800000u : 0u;
    val.w = (r == 2) ? 0x3f800000u : 0u;
    voxels16[q] = val;
  }
  // distance mirror: Voxel::Empty().distance everywhere
  const size_t n4 = (size_t)B * (VK_BLOCK_VOXELS / 4);
  for (size_t q = tid; q < n4; q += stride) sdf4[q] = make_float4(1.0f, 1.0f, 1.0f, 1.0f);
  for (size_t i = tid; i < (size_t)B; i += stride) {
    reinterpret_cast<int4*>(entries)[i] = make_int4(0, 0, -1, -1);
    free_list[i] = (int)i;
    visibility[i] = VK_VIS_FALSE;
    stamps[i] = 0;
    if (i < (size_t)M) locks[i] = 0;
  }
}

// vkVolume::sdf rebuilt from the voxel pool: one warp per allocated entry (strided 4-byte reads of
// Voxel::distance, coalesced writes).  Maintenance path only (vk_volume_sync_sdf).
__global__ void __launch_bounds__(256) vk_sdf_rebuild_kernel(const vkHashEntry* entries, const vkVoxel* voxels, float* sdf,
                                                             int B) {
  const int lane = threadIdx.x & 31;
  const int warps = (gridDim.x * blockDim.x) >> 5;
  for (int e = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; e < B; e += warps) {
    const int data = ld_entry_cg(entries + e).z;
    if (data < 0) continue;
    const uint32_t* src = reinterpret_cast<const uint32_t*>(voxels) + (size_t)data * (VK_BLOCK_VOXELS * 5);
    float* dst = sdf + (size_t)data * VK_BLOCK_VOXELS;
#pragma unroll 4
    for (int i = lane; i < VK_BLOCK_VOXELS; i += 32) dst[i] = __uint_as_float(src[5 * i]);
  }
}

}  // namespace vk

using namespace vk;

extern "C" int vk_volume_sync_sdf(vkVolume* v, void* stream) {
  if (!v || !v->voxels || !v->sdf) return VK_E_INVALID;
  const int B = v->main_blocks + v->excess_blocks;
  v->has_color = 1;   // voxels were written from outside: they may carry colour
  vk_sdf_rebuild_kernel<<<sm_count() * 8, 256, 0, (cudaStream_t)stream>>>(v->entries, v->voxels, v->sdf, B);
  VK_COUNT_LAUNCH();
  return (int)cudaGetLastError();
}

namespace vk { int shard_reset_replicas(vkShard* s, cudaStream_t st); }

extern "C" int vk_volume_reset(vkVolume* v, void* stream) {
  if (!v || !v->voxels) return VK_E_INVALID;
  cudaStream_t s = (cudaStream_t)stream;
  if (v->shard) {
    const int rc = shard_reset_replicas(v->shard, s);
    if (rc) return rc;
  }
  const int B = v->main_blocks + v->excess_blocks;
  const size_t n16 = (size_t)B * VK_BLOCK_VOXELS * sizeof(vkVoxel) / 16;
  vk_fill_pools_kernel<<<148 * 8, 256, 0, s>>>(reinterpret_cast<uint4*>(v->voxels), n16, reinterpret_cast<float4*>(v->sdf), v->entries,
                                               v->free_list, v->visibility, v->stamps, v->locks, v->counters, B,
                                               v->main_blocks);
  VK_COUNT_LAUNCH();
  cudaError_t err = cudaMemsetAsync(v->lock_tags, 0, (size_t)v->main_blocks * sizeof(unsigned long long), s);
  if (err != cudaSuccess) return (int)err;
  err = cudaMemsetAsync(v->bucket_bits, 0, ((size_t)v->main_blocks + 31) / 32 * sizeof(uint32_t), s);
  if (err != cudaSuccess) return (int)err;
  err = cudaMemsetAsync(v->space_bits, 0, VK_SPACE_WORDS * sizeof(uint32_t), s);
  if (err != cudaSuccess) return (int)err;
  err = cudaStreamSynchronize(s);  // everything that could still write the host mirror has retired
  if (err != cudaSuccess) return (int)err;
  for (int i = 0; i < VK_CNT_COUNT; ++i) v->host_counters[i] = 0;
  v->host_counters[VK_CNT_VOXEL_PTR] = B - 1;
  v->host_counters[VK_CNT_EXCESS_PTR] = v->main_blocks;
  v->visible_slot = 0;
  v->view_id = 0;
  v->has_color = 0;
  v->aux_valid = 0;
  return 0;
}

extern "C" int vk_volume_create(vkVolume* v, int main_blocks, int excess_blocks, void* stream) {
  if (!v || main_blocks <= 0 || excess_blocks < 0) return VK_E_INVALID;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { cudaGetLastError(); return VK_E_NODEVICE; }
  // Q20: voxel indices are 32-bit ints in the reference (B <= 4194303); ours are 64-bit
  // byte offsets, the limit here is only that ids fit an int.
  const size_t B = (size_t)main_blocks + (size_t)excess_blocks;
  memset(v, 0, sizeof(*v));
  v->main_blocks = main_blocks;
  v->excess_blocks = excess_blocks;
  v->voxel_length = 0.008f;   // src/volume.cu:374-376
  v->trunc_length = 0.04f;
  v->min_depth = 0.1f;
  v->max_depth = 5.0f;
  v->shard_rank = 0;
  v->shard_count = 1;
  v->shard_shift = 3;
  v->shard_halo = 0;
  v->math_mode = g_default_math_mode.load();
  cudaError_t err;
#define VK_TRY(x) if ((err = (x)) != cudaSuccess) { vk_volume_destroy(v); return (int)err; }
  VK_TRY(cudaMalloc((void**)&v->voxels, B * VK_BLOCK_VOXELS * sizeof(vkVoxel)));
  VK_TRY(cudaMalloc((void**)&v->entries, B * sizeof(vkHashEntry)));
  VK_TRY(cudaMalloc((void**)&v->free_list, B * sizeof(int32_t)));
  VK_TRY(cudaMalloc((void**)&v->visibility, B));
  VK_TRY(cudaMalloc((void**)&v->visible[0], B * sizeof(int32_t)));
  VK_TRY(cudaMalloc((void**)&v->visible[1], B * sizeof(int32_t)));
  VK_TRY(cudaMalloc((void**)&v->stamps, B * sizeof(int32_t)));
  VK_TRY(cudaMalloc((void**)&v->locks, (size_t)main_blocks * sizeof(int32_t)));
  VK_TRY(cudaMalloc((void**)&v->lock_tags, (size_t)main_blocks * sizeof(unsigned long long)));
  VK_TRY(cudaMalloc((void**)&v->bucket_bits, ((size_t)main_blocks + 31) / 32 * sizeof(uint32_t)));
  VK_TRY(cudaMalloc((void**)&v->space_bits, VK_SPACE_WORDS * sizeof(uint32_t)));
  VK_TRY(cudaMalloc((void**)&v->sdf, B * VK_BLOCK_VOXELS * sizeof(float)));
  VK_TRY(cudaMalloc((void**)&v->counters, (VK_CNT_COUNT + 4) * sizeof(int32_t)));
  VK_TRY(cudaHostAlloc((void**)&v->host_counters, VK_CNT_COUNT * sizeof(int32_t), cudaHostAllocMapped));
  {
    cudaStream_t aux = nullptr;
    cudaEvent_t ev[3] = {nullptr, nullptr, nullptr};
    VK_TRY(cudaStreamCreateWithFlags(&aux, cudaStreamNonBlocking));
    v->aux_stream = aux;
    for (int i = 0; i < 3; ++i) VK_TRY(cudaEventCreateWithFlags(&ev[i], cudaEventDisableTiming));
    v->ev_view = ev[0]; v->ev_bounds = ev[1]; v->ev_trace = ev[2];
  }
#undef VK_TRY
  return vk_volume_reset(v, stream);
}

extern "C" int vk_volume_destroy(vkVolume* v) {
  if (!v) return VK_E_INVALID;
  cudaFree(v->voxels);
  cudaFree(v->entries);
  cudaFree(v->free_list);
  cudaFree(v->visibility);
  cudaFree(v->visible[0]);
  cudaFree(v->visible[1]);
  cudaFree(v->stamps);
  cudaFree(v->locks);
  cudaFree(v->lock_tags);
  cudaFree(v->bucket_bits);
  cudaFree(v->space_bits);
  cudaFree(v->sdf);
  cudaFree(v->counters);
  if (v->host_counters) cudaFreeHost(v->host_counters);
  if (v->aux_stream) {
    cudaStreamSynchronize((cudaStream_t)v->aux_stream);
    cudaStreamDestroy((cudaStream_t)v->aux_stream);
  }
  if (v->ev_view) cudaEventDestroy((cudaEvent_t)v->ev_view);
  if (v->ev_bounds) cudaEventDestroy((cudaEvent_t)v->ev_bounds);
  if (v->ev_trace) cudaEventDestroy((cudaEvent_t)v->ev_trace);
  const int m = v->main_blocks, e = v->excess_blocks;
  memset(v, 0, sizeof(*v));
  v->main_blocks = m;
  v->excess_blocks = e;
  return 0;
}

extern "C" int vk_set_view(vkVolume* v, const vkFrame* f, void* stream) {
  if (!v || !f || !v->voxels || !f->depth || f->depth_w <= 0 || f->depth_h <= 0) return VK_E_INVALID;
  cudaStream_t s = (cudaStream_t)stream;
  const int prev = v->visible_slot, cur = prev ^ 1;
  v->view_id += 1;

  AllocParams a;
  a.entries = v->entries;
  a.visibility = v->visibility;
  a.stamps = v->stamps;
  a.locks = v->locks;
  a.lock_tags = reinterpret_cast<unsigned long long*>(v->lock_tags);
  a.bucket_bits = v->bucket_bits;
  a.space_bits = v->space_bits;
  a.free_list = v->free_list;
  a.counters = v->counters;
  a.visible_out = v->visible[cur];
  a.cnt_slot = VK_CNT_VISIBLE0 + cur;
  a.depth = f->depth;
  a.w = f->depth_w;
  a.h = f->depth_h;
  a.proj = f->depth_proj;
  rows34(f->depth_to_world.m, a.Twd);
  a.K = make_mod((uint32_t)v->main_blocks);
  a.B = v->main_blocks + v->excess_blocks;
  a.block_length = VK_BLOCK_RES * v->voxel_length;
  a.inv_block_length = 1.0f / a.block_length;   // src/volume.cu:131 (1.0f / block_length, IEEE)
  a.trunc = v->trunc_length;
  a.min_depth = v->min_depth;
  a.max_depth = v->max_depth;
  a.view_id = v->view_id;
  a.shard_rank = v->shard_rank;
  a.shard_count = v->shard_count;
  a.shard_shift = v->shard_shift;
  a.shard_halo = v->shard_halo;
  a.push_count = 0;
  for (int r = 0; r < VK_MAX_SHARDS; ++r) a.push_entries[r] = nullptr;
  if (v->shard && v->shard->count > 1) {
    if (!v->shard->connected) return VK_E_INVALID;
    a.push_count = v->shard->count;
    for (int r = 0; r < v->shard->count; ++r) a.push_entries[r] = v->shard->push_entries[r];
  }
  const bool strict = (v->math_mode == VK_MATH_STRICT);
  {
    const int threads = 256;  // (64 / 128 / 256 measured the same on B200)
    const int tiles = ((a.w + 7) / 8) * ((a.h + 3) / 4), wpc = threads / 32;
    const dim3 wgrid((tiles + wpc - 1) / wpc);
    if (strict) launch_chained(vk_alloc_kernel<false>, wgrid, dim3(threads), 0, s, a);
    else launch_chained(vk_alloc_kernel<true>, wgrid, dim3(threads), 0, s, a);
  }
  VK_COUNT_LAUNCH();

  CarryParams c;
  c.entries = v->entries;
  c.visibility = v->visibility;
  c.stamps = v->stamps;
  c.counters = v->counters;
  c.prev_list = v->visible[prev];
  c.visible_out = v->visible[cur];
  c.prev_slot = VK_CNT_VISIBLE0 + prev;
  c.cnt_slot = VK_CNT_VISIBLE0 + cur;
  c.w = f->depth_w;
  c.h = f->depth_h;
  c.proj = f->depth_proj;
  rows34(f->depth_to_world.inv, c.Tdw);
  c.block_length = a.block_length;
  c.view_id = v->view_id;
  c.ticket = v->counters + VK_CNT_COUNT;
  if (strict) launch_chained(vk_carry_kernel<false>, dim3(sm_count()), dim3(256), 0, s, c);
  else launch_chained(vk_carry_kernel<true>, dim3(sm_count()), dim3(256), 0, s, c);
  VK_COUNT_LAUNCH();
  v->visible_slot = cur;
  // vk_trace may compute this view's ray bounds on the side stream as soon as the list above exists
  v->aux_valid = 0;
  if (v->aux_stream && v->ev_view && !stream_capturing(s) && cudaEventRecord((cudaEvent_t)v->ev_view, s) == cudaSuccess)
    v->aux_valid = 1;
  return (int)cudaGetLastError();
}

extern "C" int vk_volume_sync_counters(vkVolume* v, int* visible, void* stream) {
  if (!v || !v->counters) return VK_E_INVALID;
  cudaStream_t s = (cudaStream_t)stream;
  cudaError_t err = cudaMemcpyAsync(v->host_counters, v->counters, VK_CNT_COUNT * sizeof(int32_t),
                                    cudaMemcpyDeviceToHost, s);
  if (err != cudaSuccess) return (int)err;
  err = cudaStreamSynchronize(s);
  if (err != cudaSuccess) return (int)err;
  if (visible) *visible = v->host_counters[VK_CNT_VISIBLE0 + v->visible_slot];
  return v->host_counters[VK_CNT_OVERFLOW] ? VK_E_OVERFLOW : 0;
}
