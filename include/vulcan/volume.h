// vulcan/volume.h - the voxel-hash volume (reference: include/vulcan/volume.h:16-117,
// src/volume.cu:370-463).  Same constructor, setters, SetView and buffer getters; the pools live in
// a vkVolume owned through the C ABI and are exposed as Buffer<T> views.
//
// Differences that are part of the design:
//   - one SetView call reaches the fixed point the reference needs repeated calls for; calling it
//     several times like the reference's callers do (apps/vulcan/vulcan.cu:316-318) is harmless;
//   - allocator counters are per volume (the reference keeps them in process-wide __device__
//     globals, src/volume.cu:17-21, so that two Volume objects corrupt each other);
//   - the visible count stays on the device: GetVisibleBlocks() synchronises to learn its size.
#pragma once

#include <string>

#include <cstdint>
#include <vulcan/buffer.h>
#include <vulcan/matrix.h>
#include <vulcan/types.h>
#include <vulcan_b200.h>

namespace vulcan
{

class Block;
struct Frame;
class HashEntry;
class Voxel;

class Volume
{
  public:

    Volume(int main_block_count, int excess_block_count);
    ~Volume();

    int GetMainBlockCount() const;
    int GetExcessBlockCount() const;

    const Vector2f& GetDepthRange() const;
    void SetDepthRange(const Vector2f& range);
    void SetDepthRange(float min, float max);

    float GetVoxelLength() const;
    void SetVoxelLength(float length);

    float GetTruncationLength() const;
    void SetTruncationLength(float length);

    void SetView(const Frame& frame);

    const Buffer<HashEntry>& GetHashEntries() const;
    const Buffer<int>& GetAllocatedBlocks() const;  // throws "not implemented" like src/volume.cu:444-448
    const Buffer<int>& GetVisibleBlocks() const;
    const Buffer<Voxel>& GetVoxels() const;
    Buffer<Voxel>& GetVoxels();

    // ---- not in the reference ------------------------------------------------------------
    const Buffer<Visibility>& GetBlockVisibility() const;  // protected member block_visibility_ there
    int GetVisibleBlockCount() const;                      // synchronises the stream
    int GetAllocatedBlockCount() const;                    // synchronises the stream
    void Reset();                                          // back to the freshly constructed state
    void SetSharding(int rank, int count, int brick_shift = 3, int halo = 0);  // spatial block sharding (multi-GPU)
    void Save(const std::string& path) const;   // snapshot of the complete state (addition)
    void Load(const std::string& path);         // into a volume of the same block counts
    void SetStream(cudaStream_t stream) { stream_ = stream; }
    cudaStream_t GetStream() const { return stream_; }
    // The ray marcher samples a block-major mirror of Voxel::distance (vkVolume::sdf).  The non-const
    // GetVoxels() hands out write access to the pool like the reference does, so it marks the mirror
    // stale; Tracer::Trace rebuilds it (vk_volume_sync_sdf) before marching.  Callers that write the
    // pool through a pointer obtained earlier call this themselves.
    void MarkVoxelsWritten() const { sdf_stale_ = true; }
    void SyncDistances() const;
    vkVolume* GetHandle() { return &volume_; }
    const vkVolume* GetHandle() const { return &volume_; }

  private:

    Volume(const Volume&);
    Volume& operator=(const Volume&);
    void RefreshViews() const;

  protected:

    mutable vkVolume volume_;
    mutable bool sdf_stale_ = false;
    cudaStream_t stream_;
    Vector2f depth_range_;
    mutable Buffer<Voxel> voxels_;
    mutable Buffer<HashEntry> hash_entries_;
    mutable Buffer<Visibility> block_visibility_;
    mutable Buffer<int> visible_blocks_;
    Buffer<int> allocated_blocks_;
};

} // namespace vulcan
