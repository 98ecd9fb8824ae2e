// vk_snapshot.cu - volume snapshots (the reference has no volume I/O).
//
// vk_volume_save writes everything a volume needs to continue exactly where it was: the hash
// entries, free list, visibility, both visible lists, stamps, occupancy bitmaps and counters in
// full (a few MB), and the voxel blocks that some entry refers to (10 KB each; the other blocks
// of the pool still hold Voxel::Empty()).  vk_volume_load restores them into a volume created
// with the same block counts.  The file is little-endian and versioned; voxel blocks travel
// through a pinned staging buffer in chunks, gathered / scattered by a kernel, so a 43 GB pool
// with a few thousand allocated blocks costs a few MB, not 43 GB.  A load reads and validates the
// whole file on the host first (sizes, every index range, the block list against the table) and
// touches the volume only then; it re-empties only the blocks the volume uses at that moment, and
// a failure after that point leaves the volume reset, never half loaded.
#include "vk_math.cuh"
#include "vk_launch.h"
#include <cstdio>
#include <vector>

namespace vk {

struct SnapshotHeader {
  char magic[8];  // "VKVOL\0\0\1"
  int32_t version;
  int32_t main_blocks, excess_blocks;
  float voxel_length, trunc_length, min_depth, max_depth;
  int32_t shard_rank, shard_count, shard_shift, shard_halo;
  int32_t visible_slot, view_id, math_mode;
  int32_t counters[VK_CNT_COUNT + 4];
  int64_t stored_blocks;  // voxel blocks that follow
};

constexpr int kSnapshotVersion = 1;
constexpr size_t kChunkBlocks = 4096;  // 40 MB of voxels per staging round

// dst[i] = pool[ids[i]] (gather) or pool[ids[i]] = src[i] (scatter), 10 KB per block, 16 B per thread
__global__ void __launch_bounds__(256) vk_move_blocks_kernel(uint4* pool, uint4* staging, const int32_t* ids, int n,
                                                             int scatter) {
  constexpr int kQuads = VK_BLOCK_VOXELS * sizeof(vkVoxel) / 16;  // 640
  for (int b = blockIdx.x; b < n; b += gridDim.x) {
    uint4* p = pool + (size_t)ids[b] * kQuads;
    uint4* s = staging + (size_t)b * kQuads;
    for (int q = threadIdx.x; q < kQuads; q += blockDim.x) {
      if (scatter) p[q] = s[q];
      else s[q] = p[q];
    }
  }
}

struct File {
  FILE* f = nullptr;
  ~File() { if (f) fclose(f); }
};

static bool put(FILE* f, const void* p, size_t n) { return n == 0 || fwrite(p, 1, n, f) == n; }
static bool get(FILE* f, void* p, size_t n) { return n == 0 || fread(p, 1, n, f) == n; }

// device array <-> file through a host vector
static int save_array(FILE* f, const void* dev, size_t bytes, cudaStream_t s) {
  std::vector<char> host(bytes);
  cudaError_t err = cudaMemcpyAsync(host.data(), dev, bytes, cudaMemcpyDeviceToHost, s);
  if (err == cudaSuccess) err = cudaStreamSynchronize(s);
  if (err != cudaSuccess) return (int)err;
  return put(f, host.data(), bytes) ? 0 : VK_E_IO;
}
}  // namespace vk

using namespace vk;

extern "C" int vk_volume_save(const vkVolume* v, const char* path, void* stream) {
  if (!v || !v->voxels || !path) return VK_E_INVALID;
  cudaStream_t s = (cudaStream_t)stream;
  const size_t M = (size_t)v->main_blocks, B = M + (size_t)v->excess_blocks;
  File file;
  file.f = fopen(path, "wb");
  if (!file.f) return VK_E_IO;

  // entries first (on the host): they say which voxel blocks are in use
  std::vector<vkHashEntry> entries(B);
  cudaError_t err = cudaMemcpyAsync(entries.data(), v->entries, B * sizeof(vkHashEntry), cudaMemcpyDeviceToHost, s);
  if (err == cudaSuccess) err = cudaStreamSynchronize(s);
  if (err != cudaSuccess) return (int)err;
  std::vector<int32_t> ids;
  for (size_t i = 0; i < B; ++i) {
    int32_t data;
    memcpy(&data, reinterpret_cast<const char*>(&entries[i]) + 8, sizeof(data));  // HashEntry::data (hash.h:49-55)
    if (data >= 0) ids.push_back(data);
  }

  SnapshotHeader h;
  memset(&h, 0, sizeof(h));
  memcpy(h.magic, "VKVOL\0\0\1", 8);
  h.version = kSnapshotVersion;
  h.main_blocks = v->main_blocks; h.excess_blocks = v->excess_blocks;
  h.voxel_length = v->voxel_length; h.trunc_length = v->trunc_length;
  h.min_depth = v->min_depth; h.max_depth = v->max_depth;
  h.shard_rank = v->shard_rank; h.shard_count = v->shard_count; h.shard_shift = v->shard_shift;
  h.shard_halo = v->shard_halo;
  h.visible_slot = v->visible_slot; h.view_id = v->view_id; h.math_mode = v->math_mode;
  err = cudaMemcpyAsync(h.counters, v->counters, sizeof(h.counters), cudaMemcpyDeviceToHost, s);
  if (err == cudaSuccess) err = cudaStreamSynchronize(s);
  if (err != cudaSuccess) return (int)err;
  h.stored_blocks = (int64_t)ids.size();
  if (!put(file.f, &h, sizeof(h))) return VK_E_IO;
  if (!put(file.f, entries.data(), B * sizeof(vkHashEntry))) return VK_E_IO;

  int rc;
  if ((rc = save_array(file.f, v->free_list, B * sizeof(int32_t), s))) return rc;
  if ((rc = save_array(file.f, v->visibility, B, s))) return rc;
  if ((rc = save_array(file.f, v->visible[0], B * sizeof(int32_t), s))) return rc;
  if ((rc = save_array(file.f, v->visible[1], B * sizeof(int32_t), s))) return rc;
  if ((rc = save_array(file.f, v->stamps, B * sizeof(int32_t), s))) return rc;
  if ((rc = save_array(file.f, v->bucket_bits, (M + 31) / 32 * sizeof(uint32_t), s))) return rc;
  if ((rc = save_array(file.f, v->space_bits, VK_SPACE_WORDS * sizeof(uint32_t), s))) return rc;
  if (!put(file.f, ids.data(), ids.size() * sizeof(int32_t))) return VK_E_IO;

  // voxel blocks, gathered chunk by chunk
  const size_t block_bytes = VK_BLOCK_VOXELS * sizeof(vkVoxel);
  const size_t chunk = ids.size() < kChunkBlocks ? ids.size() : kChunkBlocks;
  if (chunk) {
    uint4* d_stage = nullptr;
    int32_t* d_ids = nullptr;
    char* h_stage = nullptr;
    err = cudaMalloc((void**)&d_stage, chunk * block_bytes);
    if (err == cudaSuccess) err = cudaMalloc((void**)&d_ids, chunk * sizeof(int32_t));
    if (err == cudaSuccess) err = cudaMallocHost((void**)&h_stage, chunk * block_bytes);
    for (size_t at = 0; err == cudaSuccess && rc == 0 && at < ids.size(); at += chunk) {
      const size_t n = (ids.size() - at < chunk) ? ids.size() - at : chunk;
      err = cudaMemcpyAsync(d_ids, ids.data() + at, n * sizeof(int32_t), cudaMemcpyHostToDevice, s);
      if (err != cudaSuccess) break;
      vk_move_blocks_kernel<<<sm_count() * 4, 256, 0, s>>>(reinterpret_cast<uint4*>(v->voxels), d_stage, d_ids, (int)n, 0);
      VK_COUNT_LAUNCH();
      err = cudaMemcpyAsync(h_stage, d_stage, n * block_bytes, cudaMemcpyDeviceToHost, s);
      if (err == cudaSuccess) err = cudaStreamSynchronize(s);
      if (err == cudaSuccess && !put(file.f, h_stage, n * block_bytes)) rc = VK_E_IO;
    }
    cudaFree(d_stage);
    cudaFree(d_ids);
    cudaFreeHost(h_stage);
    if (err != cudaSuccess) return (int)err;
    if (rc) return rc;
  }
  return fflush(file.f) == 0 ? 0 : VK_E_IO;
}

// blocks back to Voxel::Empty() (the pattern of vk_fill_pools_kernel: distance 1, everything else 0)
__global__ void __launch_bounds__(256) vk_empty_blocks_kernel(uint4* pool, float4* sdf4, const int32_t* ids, int n) {
  constexpr int kQuads = VK_BLOCK_VOXELS * sizeof(vkVoxel) / 16;  // 640
  for (int b = blockIdx.x; b < n; b += gridDim.x) {
    uint4* p = pool + (size_t)ids[b] * kQuads;
    float4* m = sdf4 + (size_t)ids[b] * (VK_BLOCK_VOXELS / 4);  // the distance mirror of the block
    for (int q = threadIdx.x; q < VK_BLOCK_VOXELS / 4; q += blockDim.x) m[q] = make_float4(1.0f, 1.0f, 1.0f, 1.0f);
    for (int q = threadIdx.x; q < kQuads; q += blockDim.x) {
      const uint32_t r = (uint32_t)((q * 4) % 5);
      p[q] = make_uint4((r == 0) ? 0x3f800000u : 0u, (r == 4) ? 0x3f800000u : 0u, (r == 3) ? 0x3f800000u : 0u,
                        (r == 2) ? 0x3f800000u : 0u);
    }
  }
}

static int upload(void* dev, const void* host, size_t bytes, cudaStream_t s) {
  if (bytes == 0) return 0;
  cudaError_t err = cudaMemcpyAsync(dev, host, bytes, cudaMemcpyHostToDevice, s);
  if (err == cudaSuccess) err = cudaStreamSynchronize(s);
  return (int)err;
}

// voxel-block ids some entry of a host copy of the table refers to
static std::vector<int32_t> ids_in_use(const std::vector<vkHashEntry>& entries) {
  std::vector<int32_t> ids;
  for (const vkHashEntry& e : entries)
    if (e.data >= 0) ids.push_back(e.data);
  return ids;
}

extern "C" int vk_volume_load(vkVolume* v, const char* path, void* stream) {
  if (!v || !v->voxels || !path) return VK_E_INVALID;
  // the peers' copies of a sharded volume's tables could not follow a load
  if (v->shard && v->shard->count > 1) return VK_E_INVALID;
  cudaStream_t s = (cudaStream_t)stream;
  File file;
  file.f = fopen(path, "rb");
  if (!file.f) return VK_E_IO;
  SnapshotHeader h;
  if (!get(file.f, &h, sizeof(h))) return VK_E_IO;
  if (memcmp(h.magic, "VKVOL\0\0\1", 8) != 0 || h.version != kSnapshotVersion) return VK_E_IO;
  if (h.main_blocks != v->main_blocks || h.excess_blocks != v->excess_blocks) return VK_E_INVALID;
  // a snapshot of one shard only fits a volume set up as that shard: the caller's sharding is not overridden
  if (h.shard_rank != v->shard_rank || h.shard_count != v->shard_count || h.shard_shift != v->shard_shift ||
      h.shard_halo != v->shard_halo)
    return VK_E_INVALID;
  const size_t M = (size_t)v->main_blocks, B = M + (size_t)v->excess_blocks;
  if (h.stored_blocks < 0 || (size_t)h.stored_blocks > B) return VK_E_IO;

  // ---- 1. the whole file is read and validated on the host before the volume is touched ----------
  std::vector<vkHashEntry> entries(B);
  std::vector<int32_t> free_list(B), visible0(B), visible1(B), stamps(B);
  std::vector<uint8_t> visibility(B);
  std::vector<uint32_t> bucket_bits((M + 31) / 32), space_bits(VK_SPACE_WORDS);
  std::vector<int32_t> ids((size_t)h.stored_blocks);
  if (!get(file.f, entries.data(), B * sizeof(vkHashEntry)) || !get(file.f, free_list.data(), B * sizeof(int32_t)) ||
      !get(file.f, visibility.data(), B) || !get(file.f, visible0.data(), B * sizeof(int32_t)) ||
      !get(file.f, visible1.data(), B * sizeof(int32_t)) || !get(file.f, stamps.data(), B * sizeof(int32_t)) ||
      !get(file.f, bucket_bits.data(), bucket_bits.size() * sizeof(uint32_t)) ||
      !get(file.f, space_bits.data(), space_bits.size() * sizeof(uint32_t)) ||
      !get(file.f, ids.data(), ids.size() * sizeof(int32_t)))
    return VK_E_IO;
  const size_t block_bytes = VK_BLOCK_VOXELS * sizeof(vkVoxel);
  {  // the voxel blocks must all be there: header + arrays + stored_blocks * 10 KB is the file size
    const long at = ftell(file.f);
    if (at < 0 || fseek(file.f, 0, SEEK_END) != 0) return VK_E_IO;
    const long end = ftell(file.f);
    if (end < 0 || (unsigned long long)(end - at) != (unsigned long long)ids.size() * block_bytes) return VK_E_IO;
    if (fseek(file.f, at, SEEK_SET) != 0) return VK_E_IO;
  }
  const int iB = (int)B, iM = (int)M;
  size_t in_use = 0;
  for (size_t i = 0; i < B; ++i) {
    const vkHashEntry& e = entries[i];
    if (e.data < -1 || e.data >= iB) return VK_E_IO;
    if (e.next != -1 && (e.next < iM || e.next >= iB)) return VK_E_IO;   // chains live in the excess range
    if (e.data >= 0) ++in_use;
    if (free_list[i] < 0 || free_list[i] >= iB) return VK_E_IO;
    if (visibility[i] > VK_VIS_TRUE) return VK_E_IO;
  }
  if (in_use != ids.size() || ids != ids_in_use(entries)) return VK_E_IO;   // written in entry order by vk_volume_save
  {
    std::vector<uint8_t> seen(B, 0);
    for (int32_t id : ids) {
      if (seen[(size_t)id]) return VK_E_IO;   // two entries share a voxel block
      seen[(size_t)id] = 1;
    }
  }
  const int32_t* c = h.counters;
  if (h.visible_slot < 0 || h.visible_slot > 1 || h.view_id < 0) return VK_E_IO;
  if (c[VK_CNT_VISIBLE0] < 0 || c[VK_CNT_VISIBLE0] > iB || c[VK_CNT_VISIBLE1] < 0 || c[VK_CNT_VISIBLE1] > iB) return VK_E_IO;
  if (c[VK_CNT_VOXEL_PTR] >= iB || c[VK_CNT_EXCESS_PTR] < iM || c[VK_CNT_ALLOCATED] != (int32_t)in_use) return VK_E_IO;
  for (int k = 0; k < 2; ++k) {
    const std::vector<int32_t>& list = k ? visible1 : visible0;
    for (int i = 0; i < c[VK_CNT_VISIBLE0 + k]; ++i)
      if (list[(size_t)i] < 0 || list[(size_t)i] >= iB || entries[(size_t)list[(size_t)i]].data < 0) return VK_E_IO;
  }

  // ---- 2. the volume: re-empty the blocks it uses NOW (not the whole pool), then the saved state ------
  std::vector<vkHashEntry> current(B);
  cudaError_t err = cudaMemcpyAsync(current.data(), v->entries, B * sizeof(vkHashEntry), cudaMemcpyDeviceToHost, s);
  if (err == cudaSuccess) err = cudaStreamSynchronize(s);
  if (err != cudaSuccess) return (int)err;
  const std::vector<int32_t> stale = ids_in_use(current);
  const size_t most = stale.size() > kChunkBlocks ? stale.size() : kChunkBlocks;
  uint4* d_stage = nullptr;
  int32_t* d_ids = nullptr;
  char* h_stage = nullptr;
  int rc = 0;
  err = cudaMalloc((void**)&d_ids, most * sizeof(int32_t));
  if (err == cudaSuccess) err = cudaMalloc((void**)&d_stage, kChunkBlocks * block_bytes);
  if (err == cudaSuccess) err = cudaMallocHost((void**)&h_stage, kChunkBlocks * block_bytes);
  if (err == cudaSuccess && !stale.empty()) {
    err = cudaMemcpyAsync(d_ids, stale.data(), stale.size() * sizeof(int32_t), cudaMemcpyHostToDevice, s);
    if (err == cudaSuccess) {
      vk_empty_blocks_kernel<<<sm_count() * 4, 256, 0, s>>>(reinterpret_cast<uint4*>(v->voxels),
                                                            reinterpret_cast<float4*>(v->sdf), d_ids, (int)stale.size());
      VK_COUNT_LAUNCH();
      err = cudaStreamSynchronize(s);
    }
  }
  if (err == cudaSuccess) {
    // from here on a failure leaves a half-written volume: it is reset (empty, consistent) on the way out
    if (!rc) rc = upload(v->entries, entries.data(), B * sizeof(vkHashEntry), s);
    if (!rc) rc = upload(v->free_list, free_list.data(), B * sizeof(int32_t), s);
    if (!rc) rc = upload(v->visibility, visibility.data(), B, s);
    if (!rc) rc = upload(v->visible[0], visible0.data(), B * sizeof(int32_t), s);
    if (!rc) rc = upload(v->visible[1], visible1.data(), B * sizeof(int32_t), s);
    if (!rc) rc = upload(v->stamps, stamps.data(), B * sizeof(int32_t), s);
    if (!rc) rc = upload(v->bucket_bits, bucket_bits.data(), bucket_bits.size() * sizeof(uint32_t), s);
    if (!rc) rc = upload(v->space_bits, space_bits.data(), space_bits.size() * sizeof(uint32_t), s);
    if (!rc) rc = upload(v->counters, h.counters, sizeof(h.counters), s);
    for (size_t at = 0; rc == 0 && at < ids.size(); at += kChunkBlocks) {
      const size_t n = (ids.size() - at < kChunkBlocks) ? ids.size() - at : kChunkBlocks;
      if (!get(file.f, h_stage, n * block_bytes)) { rc = VK_E_IO; break; }
      err = cudaMemcpyAsync(d_ids, ids.data() + at, n * sizeof(int32_t), cudaMemcpyHostToDevice, s);
      if (err == cudaSuccess) err = cudaMemcpyAsync(d_stage, h_stage, n * block_bytes, cudaMemcpyHostToDevice, s);
      if (err != cudaSuccess) break;
      vk_move_blocks_kernel<<<sm_count() * 4, 256, 0, s>>>(reinterpret_cast<uint4*>(v->voxels), d_stage, d_ids, (int)n, 1);
      VK_COUNT_LAUNCH();
      err = cudaStreamSynchronize(s);
      if (err != cudaSuccess) break;
    }
  }
  cudaFree(d_stage);
  cudaFree(d_ids);
  cudaFreeHost(h_stage);
  if (err != cudaSuccess && rc == 0) rc = (int)err;
  if (!rc) rc = vk_volume_sync_sdf(v, stream);   // the distance mirror of the restored blocks
  if (!rc) rc = (int)cudaStreamSynchronize(s);
  if (rc) {
    vk_volume_reset(v, stream);
    return rc;
  }
  // properties of the stored data; the caller's math mode and sharding stay as they are
  v->voxel_length = h.voxel_length; v->trunc_length = h.trunc_length;
  v->min_depth = h.min_depth; v->max_depth = h.max_depth;
  v->visible_slot = h.visible_slot; v->view_id = h.view_id;
  v->aux_valid = 0;   // the visible list was replaced: no side-stream bounds until the next vk_set_view
  memcpy(v->host_counters, h.counters, VK_CNT_COUNT * sizeof(int32_t));
  return 0;
}
