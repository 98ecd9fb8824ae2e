// vk_alloc.cu — Volume::SetView rebuilt for sm_100a.
//
// Reference (src/volume.cu:430-437) = thrust::replace over all B visibility bytes
// + per-pixel request kernel (one surviving request per bucket per pass, racy)
// + per-bucket handler + a 20-barrier scan over all B entries + two blocking
// symbol memcpys, and the caller has to repeat the whole thing until allocation
// converges.  Here one call reaches that fixed point with ONE launch and no host
// synchronisation (vk_setview_kernel below).
#include "vk_math.cuh"
#include "vk_launch.h"

namespace vk {

struct AllocParams {
  vkHashEntry* entries;
  uint8_t* visibility;
  int32_t* stamps;
  int32_t* locks;
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
  while (!done) {
    if (atomicCAS(p.locks + h, 0, 1) == 0) {
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
      int tail;
      bool head_empty;
      const int idx = find_block(p.entries, h, key, tail, head_empty);
      if (idx >= 0) { result = idx; done = true; }
    }
  }
  return result;
}

// One launch per SetView.  CreateAllocationRequestsKernel (src/volume.cu:86-301) +
// HandleAllocationRequestsKernel (src/volume.cu:303-368) to their common fixed point, then
// UpdateBlockVisibilityKernel (src/volume.cu:23-84) restricted to the entries that can be
// VISIBILITY_UNKNOWN, i.e. the previous visible list.
//
// A persistent grid (every CTA resident) in three phases:
//   1. pixels.  A thread walks the truncation band of one depth pixel with the reference's
//      Amanatides-Woo comparison tree.  A block costs two independent loads - its bucket head and that
//      head's view stamp, issued one DDA step AHEAD of their use - and in the steady state (block at the
//      head of its bucket, already stamped by a neighbouring ray) nothing else: no warp-wide matching, no
//      atomics.  Lanes of a warp that ask for the same block load the same addresses (one transaction).
//      Only the first ray to touch an entry in this view appends it (atomicExch on the stamp;
//      warp-aggregated append), and only chain walks and inserts take the slow path.
//   2. a grid-wide arrival counter (all CTAs are resident by construction: the grid is sized from the
//      occupancy of this kernel): every entry touched by a ray is now stamped.
//   3. the previous visible list: entries not stamped in this view get the 8-corner frustum test (Q6)
//      and are carried over (stay UNKNOWN, Q5) or dropped (FALSE).  The corner tests themselves do not
//      depend on phase 1 and run BEFORE the counter wait.  The last CTA mirrors the counters to the host.
// FMA map (reference release SASS of CreateAllocationRequestsKernel):
//   dir   = Twd * (unproj, 0)      FMUL y*m1; FFMA x*m0 + .; FADD m2 (z == 1); FFMA 0*m3 + .
//   X     = origin + depth * dir   FFMA dir*depth + origin
//   begin = X - trunc * dir        FFMA -dir*trunc + X     end = FFMA dir*trunc + X
//   tmax  = (L*(b + max(0,step)) - begin) / dir     FFMA bf*L - begin ; IEEE div
struct CarryParams {
  uint8_t* visibility;
  int32_t* host_counters;
  const int32_t* prev_list;
  int prev_slot;
  float Tdw[12];
  int* ticket;   // [0] phase counter, [1] exit counter
};

// first toucher of an entry in this view: marks it visible and appends it.  The append is aggregated over
// the lanes that arrive here together (one atomic per group).
__device__ __forceinline__ void touch_entry(const AllocParams& p, int idx) {
  const int old = atomicExch(p.stamps + idx, p.view_id);
  const bool first = (old != p.view_id);
  if (first) p.visibility[idx] = VK_VIS_TRUE;
  const unsigned group = __activemask();
  const unsigned firsts = __ballot_sync(group, first);
  if (firsts == 0) return;
  const int lane = threadIdx.x & 31;
  const int leader = __ffs(firsts) - 1;
  int base = 0;
  if (lane == leader) base = atomicAdd(p.counters + p.cnt_slot, __popc(firsts));
  base = __shfl_sync(group, base, leader);
  if (first) p.visible_out[base + __popc(firsts & ((1u << lane) - 1))] = idx;
}

// everything that is not "block found at the head of its bucket": chain walk, insert, or a block that
// another rank owns
__device__ __noinline__ void resolve_slow(const AllocParams& p, uint32_t h, int bx, int by, int bz, unsigned long long key) {
  int tail;
  bool head_empty;
  int idx = find_block(p.entries, h, key, tail, head_empty);
  if (idx < 0) idx = insert_block(p, h, bx, by, bz, key);
  if (idx >= 0 && __ldcg(p.stamps + idx) != p.view_id) touch_entry(p, idx);
}

template <bool F>
__global__ void __launch_bounds__(256, 4) vk_setview_kernel(const AllocParams p, const CarryParams c) {
  pdl_wait();     // the previous kernel of the chain has completed (no-op for a plain launch)
  pdl_trigger();  // the next one may be scheduled as this grid drains
  const int lane = threadIdx.x & 31;
  const int tiles_x = (p.w + 7) >> 3, tiles = tiles_x * ((p.h + 3) >> 2);
  const int warps = gridDim.x * (blockDim.x >> 5);
  Mat34 Twd;
  Twd.load(p.Twd);
  Unprojector un;
  un.init(p.proj);
  const bool sharded = p.shard_count > 1;

  // ---- phase 1: warp tiles of 8x4 pixels, taken grid-stride ---------------------------------------
  for (int tile = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); tile < tiles; tile += warps) {
    const int x = (tile % tiles_x) * 8 + (lane & 7);
    const int y = (tile / tiles_x) * 4 + (lane >> 3);
    if (x >= p.w || y >= p.h) continue;
    const float depth = __ldg(p.depth + y * p.w + x);
    if (depth < p.min_depth || depth > p.max_depth) continue;
    const float u = add((float)x, 0.5f), v = add((float)y, 0.5f);
    const float d0x = un.x<F>(u), d0y = un.y<F>(v);
    Vec3 dir = xform0<F>(Twd, d0x, d0y, 1.0f);
    const float Xx = FP<F>::mad(dir.x, depth, p.Twd[3]);
    const float Xy = FP<F>::mad(dir.y, depth, p.Twd[7]);
    const float Xz = FP<F>::mad(dir.z, depth, p.Twd[11]);
    dir = normalized<F>(dir);
    const float bgx = FP<F>::nmad(dir.x, p.trunc, Xx), bgy = FP<F>::nmad(dir.y, p.trunc, Xy),
                bgz = FP<F>::nmad(dir.z, p.trunc, Xz);
    const float enx = FP<F>::mad(dir.x, p.trunc, Xx), eny = FP<F>::mad(dir.y, p.trunc, Xy),
                enz = FP<F>::mad(dir.z, p.trunc, Xz);
    const int sx = (dir.x < 0) ? -1 : 1, sy = (dir.y < 0) ? -1 : 1, sz = (dir.z < 0) ? -1 : 1;
    const float inv = p.inv_block_length;
    int bx = (int)floorf(mul(bgx, inv)), by = (int)floorf(mul(bgy, inv)), bz = (int)floorf(mul(bgz, inv));
    const int ex = (int)floorf(mul(enx, inv)), ey = (int)floorf(mul(eny, inv)), ez = (int)floorf(mul(enz, inv));
    float tmx = fdiv(FP<F>::msub((float)(bx + max(0, sx)), p.block_length, bgx), dir.x);
    float tmy = fdiv(FP<F>::msub((float)(by + max(0, sy)), p.block_length, bgy), dir.y);
    float tmz = fdiv(FP<F>::msub((float)(bz + max(0, sz)), p.block_length, bgz), dir.z);
    if (dir.x == 0) tmx = 1E20f;
    if (dir.y == 0) tmy = 1E20f;
    if (dir.z == 0) tmz = 1E20f;
    const float tdx = fdiv(mul((float)sx, p.block_length), dir.x);
    const float tdy = fdiv(mul((float)sy, p.block_length), dir.y);
    const float tdz = fdiv(mul((float)sz, p.block_length), dir.z);

    // lookup of the current block: issued here, used one DDA step later
    uint32_t h = block_hash(bx, by, bz, p.K);
    int4 ent = ld_entry_cg(p.entries + h);
    int stamp = __ldcg(p.stamps + h);
    bool more = true;
    while (more) {
      const int cx = bx, cy = by, cz = bz;
      const uint32_t ch = h;
      const int4 cent = ent;
      const int cstamp = stamp;
      // Amanatides-Woo step with the reference's comparison tree (src/volume.cu:241-295)
      if (tmx < tmy) {
        if (tmx < tmz) { bx += sx; if (bx == ex + sx) more = false; tmx = add(tmx, tdx); }
        else           { bz += sz; if (bz == ez + sz) more = false; tmz = add(tmz, tdz); }
      } else {
        if (tmy < tmz) { by += sy; if (by == ey + sy) more = false; tmy = add(tmy, tdy); }
        else           { bz += sz; if (bz == ez + sz) more = false; tmz = add(tmz, tdz); }
      }
      if (more) {  // the next block's bucket head and stamp are on their way while this one is handled
        h = block_hash(bx, by, bz, p.K);
        ent = ld_entry_cg(p.entries + h);
        stamp = __ldcg(p.stamps + h);
      }
      const unsigned long long key = pack_block(cx, cy, cz);
      if (sharded && !shard_stores(cx, cy, cz, p.shard_shift, p.shard_count, p.shard_rank, p.shard_halo)) {
        mark_space(p, cx, cy, cz);  // exists on another rank: the free-space filter is global
      } else if (entry_block(cent) == key && cent.z != -1) {
        if (cstamp != p.view_id) touch_entry(p, (int)ch);
      } else {
        resolve_slow(p, ch, cx, cy, cz, key);
      }
    }
  }

  // ---- phase 3a: corner tests of this thread's share of the previous list (independent of phase 1) ----
  const int n_prev = p.counters[c.prev_slot];
  const int gtid = blockIdx.x * blockDim.x + threadIdx.x, gsize = gridDim.x * blockDim.x;
  constexpr int kMaxShare = 4;   // entries per thread held in registers across the wait; more go through the loop below
  int share_e[kMaxShare];
  bool share_keep[kMaxShare];
  Mat34 M;
  M.load(c.Tdw);
  auto in_frustum = [&](const int4& e4) {
    int ox, oy, oz;
    entry_origin(e4, ox, oy, oz);
    for (int k = 0; k < 8; ++k) {
      const float X = mul(p.block_length, (float)(ox + ((k & 1) >> 0)));
      const float Y = mul(p.block_length, (float)(oy + ((k & 2) >> 1)));
      const float Z = mul(p.block_length, (float)(oz + ((k & 4) >> 2)));
      const Vec3 Xd = xform1<F>(M, X, Y, Z);
      if (Xd.z < 0) continue;
      float uu, vv;
      project<F>(p.proj, Xd, uu, vv);
      if (uu >= 0 && uu <= (float)p.w && vv >= 0 && vv <= (float)p.h) return true;
    }
    return false;
  };
#pragma unroll
  for (int k = 0; k < kMaxShare; ++k) {
    const int i = gtid + k * gsize;
    share_e[k] = -1;
    share_keep[k] = false;
    if (i < n_prev) {
      share_e[k] = c.prev_list[i];
      share_keep[k] = in_frustum(ld_entry(p.entries + share_e[k]));
    }
  }

  // ---- phase 2: every CTA has finished its pixels ----------------------------------------------------
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    atomicAdd(c.ticket, 1);
    // (the launch is cooperative: all CTAs are resident; the bound only keeps a broken device from hanging,
    // and is reported through the overflow status)
    long long spins = 0;
    while (*reinterpret_cast<volatile int*>(c.ticket) < (int)gridDim.x) {
      __nanosleep(32);
      if (++spins > (1ll << 24)) { atomicAdd(p.counters + VK_CNT_OVERFLOW, 1); break; }
    }
    __threadfence();
  }
  __syncthreads();

  // ---- phase 3b: carry over / drop --------------------------------------------------------------------
  auto carry = [&](int e, bool keep_if_untouched) {
    bool keep = false;
    if (e >= 0 && __ldcg(p.stamps + e) != p.view_id) {
      keep = keep_if_untouched;
      c.visibility[e] = keep ? VK_VIS_UNKNOWN : VK_VIS_FALSE;
    }
    const unsigned m = __ballot_sync(0xffffffffu, keep);
    if (m) {
      int pos = 0;
      if (lane == 0) pos = atomicAdd(p.counters + p.cnt_slot, __popc(m));
      pos = __shfl_sync(0xffffffffu, pos, 0);
      if (keep) p.visible_out[pos + __popc(m & ((1u << lane) - 1))] = e;
    }
  };
#pragma unroll
  for (int k = 0; k < kMaxShare; ++k) carry(share_e[k], share_keep[k]);
  for (int base = kMaxShare * gsize; base < n_prev; base += gsize) {   // very long lists only
    const int i = base + gtid;
    int e = -1;
    bool keep = false;
    if (i < n_prev) {
      e = c.prev_list[i];
      keep = in_frustum(ld_entry(p.entries + e));
    }
    carry(e, keep);
  }

  // ---- last CTA out: retire the previous list, mirror the counters to the host ------------------------
  __shared__ bool last;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) last = (atomicAdd(c.ticket + 1, 1) == (int)gridDim.x - 1);
  __syncthreads();
  if (last) {
    __threadfence();
    if (threadIdx.x == 0) {
      p.counters[c.prev_slot] = 0;
      c.ticket[0] = 0;
      c.ticket[1] = 0;
    }
    if (threadIdx.x < VK_CNT_COUNT && c.host_counters) {
      int val = __ldcg(p.counters + threadIdx.x);
      if (threadIdx.x == c.prev_slot) val = 0;
      c.host_counters[threadIdx.x] = val;
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
  cudaError_t err = cudaMemsetAsync(v->bucket_bits, 0, ((size_t)v->main_blocks + 31) / 32 * sizeof(uint32_t), s);
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
  VK_TRY(cudaMalloc((void**)&v->bucket_bits, ((size_t)main_blocks + 31) / 32 * sizeof(uint32_t)));
  VK_TRY(cudaMalloc((void**)&v->space_bits, VK_SPACE_WORDS * sizeof(uint32_t)));
  VK_TRY(cudaMalloc((void**)&v->sdf, B * VK_BLOCK_VOXELS * sizeof(float)));
  VK_TRY(cudaMalloc((void**)&v->counters, (VK_CNT_COUNT + 4) * sizeof(int32_t)));
  VK_TRY(cudaHostAlloc((void**)&v->host_counters, VK_CNT_COUNT * sizeof(int32_t), cudaHostAllocMapped));
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
  cudaFree(v->bucket_bits);
  cudaFree(v->space_bits);
  cudaFree(v->sdf);
  cudaFree(v->counters);
  if (v->host_counters) cudaFreeHost(v->host_counters);
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
  CarryParams c;
  c.visibility = v->visibility;
  c.host_counters = v->host_counters;
  c.prev_list = v->visible[prev];
  c.prev_slot = VK_CNT_VISIBLE0 + prev;
  rows34(f->depth_to_world.inv, c.Tdw);
  c.ticket = v->counters + VK_CNT_COUNT;
  // persistent grid: exactly the CTAs that are resident together (phase 2 waits for all of them)
  static int per_sm[2] = {0, 0};
  if (per_sm[strict] == 0) {
    int n = 0;
    cudaError_t oe = strict ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, vk_setview_kernel<false>, 256, 0)
                            : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, vk_setview_kernel<true>, 256, 0);
    per_sm[strict] = (oe == cudaSuccess && n > 0) ? (n < 4 ? n : 4) : 1;
  }
  const int forced = g_tune[VK_TUNE_ALLOC_MODE].load();   // experiments: CTAs per SM (1..4)
  const dim3 grid(sm_count() * ((forced >= 1 && forced <= per_sm[strict]) ? forced : per_sm[strict]));
  const cudaError_t le = strict ? launch_cooperative(vk_setview_kernel<false>, grid, dim3(256), 0, s, a, c)
                                : launch_cooperative(vk_setview_kernel<true>, grid, dim3(256), 0, s, a, c);
  if (le != cudaSuccess) return (int)le;
  VK_COUNT_LAUNCH();
  v->visible_slot = cur;
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
