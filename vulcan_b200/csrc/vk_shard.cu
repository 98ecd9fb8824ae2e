// vk_shard.cu - host side of the block-sharded volume with replicated lookup structures
// (include/vulcan_b200.h, "block-sharded volume ..."; no reference counterpart - the reference is
// single-GPU).  Replica allocation, CUDA IPC plumbing and the per-view call vk_trace_sharded; the
// kernels that fill the replicas are the allocation and integration kernels themselves
// (vk_alloc.cu, vk_integrate.cu), the collectives live in vk_trace.cu (ray bounds) and
// vk_exchange.cu (image bands).
#include "vk_shard.cuh"
#include "vk_launch.h"

#include <cstring>

namespace vk {

// a replica of a peer's hash table + distance mirror in its initial state (entries invalid, distances 1)
__global__ void __launch_bounds__(256) vk_fill_replica_kernel(vkHashEntry* entries, float4* sdf4, int B) {
  const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = tid; i < (size_t)B; i += stride) reinterpret_cast<int4*>(entries)[i] = make_int4(0, 0, -1, -1);
  const size_t n4 = (size_t)B * (VK_BLOCK_VOXELS / 4);
  for (size_t q = tid; q < n4; q += stride) sdf4[q] = make_float4(1.0f, 1.0f, 1.0f, 1.0f);
}

// implemented in vk_trace.cu / vk_exchange.cu
int trace_sharded_bounds(const vkVolume* v, const vkFrame* view, float min_depth, float max_depth, const Xchg& x,
                         cudaStream_t s);
int trace_sharded_band(const vkVolume* v, const vkFrame* view, const Xchg& x, float* depth, float* color, float* normal,
                       cudaStream_t s);
int exchange_push_bands(const Xchg& x, int w, int h, cudaStream_t s);

// Volume::Reset of a sharded volume (collective: every rank resets): the local copies of the peers'
// tables go back to their initial state together with the volume itself
int shard_reset_replicas(vkShard* s, cudaStream_t st) {
  const size_t B = (size_t)s->main_blocks + (size_t)s->excess_blocks;
  for (int r = 0; r < s->count; ++r) {
    if (r == s->rank) continue;
    vk_fill_replica_kernel<<<sm_count() * 8, 256, 0, st>>>(s->entries[r], reinterpret_cast<float4*>(s->sdf[r]), (int)B);
    VK_COUNT_LAUNCH();
  }
  return (int)cudaGetLastError();
}

static void free_shard(vkShard* s) {
  for (int r = 0; r < VK_MAX_SHARDS; ++r) {
    if (r == s->rank) continue;
    if (s->connected == 1) {
      if (s->voxels[r]) cudaIpcCloseMemHandle(const_cast<vkVoxel*>(s->voxels[r]));
      if (s->push_entries[r]) cudaIpcCloseMemHandle(s->push_entries[r]);
      if (s->push_sdf[r]) cudaIpcCloseMemHandle(s->push_sdf[r]);
      if (s->exchange[r]) cudaIpcCloseMemHandle(s->exchange[r]);
    }
    if (s->entries[r]) cudaFree(s->entries[r]);
    if (s->sdf[r]) cudaFree(s->sdf[r]);
  }
  if (s->rank >= 0 && s->rank < VK_MAX_SHARDS && s->exchange[s->rank]) cudaFree(s->exchange[s->rank]);
  cudaGetLastError();
}

}  // namespace vk

using namespace vk;

extern "C" int vk_shard_create(vkShard* s, vkVolume* v, int rank, int count, int brick_shift, int width, int height,
                               void* stream) {
  if (!s || !v || !v->voxels || !v->sdf || count < 1 || count > VK_MAX_SHARDS || rank < 0 || rank >= count ||
      brick_shift < 0 || brick_shift > 15 || width <= 0 || height <= 0)
    return VK_E_INVALID;
  if (v->view_id != 0 || v->shard) return VK_E_INVALID;   // the replicas start empty: shard an empty volume
  cudaStream_t st = (cudaStream_t)stream;
  memset(s, 0, sizeof(*s));
  s->count = count; s->rank = rank; s->width = width; s->height = height;
  s->main_blocks = v->main_blocks; s->excess_blocks = v->excess_blocks;
  const size_t B = (size_t)v->main_blocks + (size_t)v->excess_blocks;
  cudaError_t err = cudaSuccess;
  for (int r = 0; r < count && err == cudaSuccess; ++r) {
    if (r == rank) {
      s->entries[r] = v->entries;
      s->sdf[r] = v->sdf;
      s->voxels[r] = v->voxels;
      continue;
    }
    err = cudaMalloc((void**)&s->entries[r], B * sizeof(vkHashEntry));
    if (err == cudaSuccess) err = cudaMalloc((void**)&s->sdf[r], B * VK_BLOCK_VOXELS * sizeof(float));
    if (err == cudaSuccess) {
      vk_fill_replica_kernel<<<sm_count() * 8, 256, 0, st>>>(s->entries[r], reinterpret_cast<float4*>(s->sdf[r]), (int)B);
      VK_COUNT_LAUNCH();
      err = cudaGetLastError();
    }
  }
  if (err == cudaSuccess) err = cudaMalloc(&s->exchange[rank], xchg_bytes(width, height));
  if (err == cudaSuccess) err = cudaMemsetAsync(s->exchange[rank], 0, xchg_bytes(width, height), st);
  if (err == cudaSuccess) err = cudaStreamSynchronize(st);
  if (err != cudaSuccess) {
    free_shard(s);
    memset(s, 0, sizeof(*s));
    return (int)err;
  }
  v->shard_rank = rank; v->shard_count = count; v->shard_shift = brick_shift; v->shard_halo = 0;
  v->shard = s;
  if (count == 1) s->connected = 2;
  return 0;
}

extern "C" int vk_shard_export(const vkShard* s, const vkVolume* v, vkShardHandles* out) {
  if (!s || !v || !out || v->shard != s) return VK_E_INVALID;
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "vkShardHandles holds CUDA IPC handles");
  memset(out, 0, sizeof(*out));
  cudaError_t err = cudaIpcGetMemHandle(reinterpret_cast<cudaIpcMemHandle_t*>(out->voxels), v->voxels);
  if (err == cudaSuccess) err = cudaIpcGetMemHandle(reinterpret_cast<cudaIpcMemHandle_t*>(out->exchange), s->exchange[s->rank]);
  for (int r = 0; r < s->count && err == cudaSuccess; ++r) {
    if (r == s->rank) continue;
    err = cudaIpcGetMemHandle(reinterpret_cast<cudaIpcMemHandle_t*>(out->rep_entries[r]), s->entries[r]);
    if (err == cudaSuccess) err = cudaIpcGetMemHandle(reinterpret_cast<cudaIpcMemHandle_t*>(out->rep_sdf[r]), s->sdf[r]);
  }
  return (int)err;
}

static cudaError_t ipc_open(const unsigned char blob[64], void** ptr) {
  cudaIpcMemHandle_t h;
  memcpy(&h, blob, sizeof(h));
  return cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess);
}

extern "C" int vk_shard_connect(vkShard* s, const vkShardHandles* all) {
  if (!s || !all || s->connected) return VK_E_INVALID;
  cudaError_t err = cudaSuccess;
  s->connected = 1;   // from here on vk_shard_destroy closes whatever was opened
  for (int r = 0; r < s->count && err == cudaSuccess; ++r) {
    if (r == s->rank) continue;
    void* p = nullptr;
    err = ipc_open(all[r].voxels, &p);
    if (err != cudaSuccess) break;
    s->voxels[r] = static_cast<const vkVoxel*>(p);
    err = ipc_open(all[r].exchange, &s->exchange[r]);
    if (err != cudaSuccess) break;
    // rank r's replica of MY tables: what my allocation / integration kernels store into
    err = ipc_open(all[r].rep_entries[s->rank], &p);
    if (err != cudaSuccess) break;
    s->push_entries[r] = static_cast<vkHashEntry*>(p);
    err = ipc_open(all[r].rep_sdf[s->rank], &p);
    if (err != cudaSuccess) break;
    s->push_sdf[r] = static_cast<float*>(p);
  }
  return (int)err;
}

extern "C" int vk_shard_connect_local(vkShard* const* shards, vkVolume* const* volumes, int count) {
  if (!shards || !volumes || count < 1 || count > VK_MAX_SHARDS) return VK_E_INVALID;
  for (int a = 0; a < count; ++a)
    if (!shards[a] || !volumes[a] || shards[a]->count != count || shards[a]->rank != a || volumes[a]->shard != shards[a] ||
        shards[a]->main_blocks != shards[0]->main_blocks || shards[a]->excess_blocks != shards[0]->excess_blocks)
      return VK_E_INVALID;
  for (int a = 0; a < count; ++a) {
    vkShard* s = shards[a];
    for (int r = 0; r < count; ++r) {
      if (r == a) continue;
      s->voxels[r] = volumes[r]->voxels;
      s->exchange[r] = shards[r]->exchange[r];
      s->push_entries[r] = shards[r]->entries[a];
      s->push_sdf[r] = shards[r]->sdf[a];
    }
    s->connected = 2;
  }
  return 0;
}

extern "C" int vk_shard_destroy(vkShard* s, vkVolume* v) {
  if (!s) return VK_E_INVALID;
  free_shard(s);
  if (v && v->shard == s) {
    v->shard = nullptr;
    v->shard_rank = 0;
    v->shard_count = 1;
  }
  memset(s, 0, sizeof(*s));
  return 0;
}

// stage: 0 = the whole view; 1, 2, 3 = its three parts for ranks that share a GPU (see the header)
static int trace_sharded(vkVolume* v, const vkFrame* view, float min_depth, float max_depth, int gather, int stage,
                         float** out_depth, float** out_color, float** out_normal, void* stream) {
  if (!v || !view || !v->voxels || !v->shard || !out_depth || !out_color || !out_normal || stage < 0 || stage > 3)
    return VK_E_INVALID;
  vkShard* s = v->shard;
  if (!s->connected || view->depth_w != s->width || view->depth_h != s->height) return VK_E_INVALID;
  cudaStream_t st = (cudaStream_t)stream;
  if (stage <= 1) s->epoch += 1;
  Xchg x;
  x.count = s->count; x.rank = s->rank; x.epoch = s->epoch; x.parity = s->epoch & 1;
  x.stages = 3;
  for (int r = 0; r < VK_MAX_SHARDS; ++r) x.buf[r] = s->exchange[r];
  void* mine = s->exchange[s->rank];
  float* depth = xchg_depth(mine, x.parity, s->width, s->height);
  float* color = xchg_color(mine, x.parity, s->width, s->height);
  float* normal = xchg_normal(mine, x.parity, s->width, s->height);
  int rc = 0;
  // 1. ray bounds of the local visible blocks, combined over the ranks; row bands (also the barrier that
  //    orders every rank's replica stores before any rank's ray cast)
  //    One rank per GPU: the grid is stored into the peers by one launch and announced by the next - the kernel
  //    boundary orders the stores before the flags without a system-scope fence (vk_shard.cuh).
  if (stage == 0 || stage == 1) {
    x.stages = (stage == 0) ? 4 : 1;
    rc = trace_sharded_bounds(v, view, min_depth, max_depth, x, st);
    if (rc) return rc;
  }
  if (stage == 0 || stage == 2) {
    x.stages = (stage == 0) ? (1 | 2 | 8) : 2;
    rc = trace_sharded_bounds(v, view, min_depth, max_depth, x, st);
    if (rc) return rc;
    // 2. this rank's band: ray march (+2 rows either side) and normals, all reads local
    rc = trace_sharded_band(v, view, x, depth, color, normal, st);
    if (rc) return rc;
    // 3. the band into the peers' images, handshake
    if (gather) {   // (a single rank: nothing to store, the handshake is trivial)
      x.stages = (stage == 0) ? 4 : 1;
      rc = exchange_push_bands(x, s->width, s->height, st);
      if (rc) return rc;
      if (stage == 0) {
        x.stages = 1 | 2 | 8;
        rc = exchange_push_bands(x, s->width, s->height, st);
        if (rc) return rc;
      }
    }
  }
  if (stage == 3 && gather) {
    x.stages = 2;
    rc = exchange_push_bands(x, s->width, s->height, st);
    if (rc) return rc;
  }
  *out_depth = depth; *out_color = color; *out_normal = normal;
  return 0;
}

extern "C" int vk_trace_sharded(vkVolume* v, const vkFrame* view, float min_depth, float max_depth, int gather,
                                float** out_depth, float** out_color, float** out_normal, void* stream) {
  return trace_sharded(v, view, min_depth, max_depth, gather, 0, out_depth, out_color, out_normal, stream);
}

extern "C" int vk_trace_sharded_staged(vkVolume* v, const vkFrame* view, float min_depth, float max_depth, int gather,
                                       int stage, float** out_depth, float** out_color, float** out_normal, void* stream) {
  if (stage < 1 || stage > 3) return VK_E_INVALID;
  return trace_sharded(v, view, min_depth, max_depth, gather, stage, out_depth, out_color, out_normal, stream);
}

extern "C" int vk_shard_band(const vkVolume* v, int* band_begin, int* band_end, void* stream) {
  if (!v || !v->shard || !band_begin || !band_end) return VK_E_INVALID;
  const vkShard* s = v->shard;
  if (s->epoch <= 0) return VK_E_INVALID;
  int bands[16], timeouts = 0;
  cudaError_t err = cudaMemcpyAsync(bands, xchg_bands(s->exchange[s->rank], s->epoch & 1), sizeof(bands),
                                    cudaMemcpyDeviceToHost, (cudaStream_t)stream);
  if (err == cudaSuccess)
    err = cudaMemcpyAsync(&timeouts, xchg_timeouts(s->exchange[s->rank]), sizeof(int), cudaMemcpyDeviceToHost,
                          (cudaStream_t)stream);
  if (err == cudaSuccess) err = cudaStreamSynchronize((cudaStream_t)stream);
  if (err != cudaSuccess) return (int)err;
  *band_begin = bands[s->rank];
  *band_end = bands[s->rank + 1];
  return timeouts ? VK_E_TIMEOUT : 0;
}

extern "C" int vk_shard_band_async(const vkVolume* v, int* host_bands, void* stream) {
  if (!v || !v->shard || !host_bands) return VK_E_INVALID;
  const vkShard* s = v->shard;
  if (s->epoch <= 0) return VK_E_INVALID;
  cudaError_t err = cudaMemcpyAsync(host_bands, xchg_bands(s->exchange[s->rank], s->epoch & 1), 16 * sizeof(int),
                                    cudaMemcpyDeviceToHost, (cudaStream_t)stream);
  if (err == cudaSuccess)
    err = cudaMemcpyAsync(host_bands + 16, xchg_timeouts(s->exchange[s->rank]), sizeof(int), cudaMemcpyDeviceToHost,
                          (cudaStream_t)stream);
  return (int)err;
}

