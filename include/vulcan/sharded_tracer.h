// vulcan/sharded_tracer.h - one rank of a block-sharded volume (addition; the reference is single-GPU).
// C++ host over vk_shard_* (include/vulcan_b200.h): the rank stores, allocates and integrates only the blocks
// it owns; hash entries and block distances are replicated into every rank as they are written, so every rank
// ray-casts its band of rows from local memory, bit-identically to an unsharded volume.
//
//   one process per GPU                        ranks that share a process
//   -------------------                        --------------------------
//   ShardedTracer st(volume, rank, n, w, h);   ShardedTracer a(va, 0, 2, w, h), b(vb, 1, 2, w, h);
//   all = exchange(st.Export());  // any byte  ShardedTracer::ConnectLocal({&a, &b});
//   st.Connect(all);              // transport
//   volume->SetView(f); integrator.Integrate(f);   // as usual, on every rank
//   st.Trace(view, images);                        // collective
#pragma once

#include <memory>
#include <vector>
#include <vulcan/frame.h>
#include <vulcan/volume.h>

namespace vulcan
{

class ShardedTracer
{
  public:

    // device images of a traced view inside the rank's exchange buffer (valid until the second-next Trace)
    struct Images
    {
      float* depth;
      float* color;
      float* normal;
      int width, height;
    };

    // must be constructed before the volume allocates anything (its peers' copies start empty)
    ShardedTracer(std::shared_ptr<Volume> volume, int rank, int count, int width, int height, int brick_shift = 3);
    ~ShardedTracer();

    int GetRank() const { return shard_.rank; }
    int GetCount() const { return shard_.count; }
    const Vector2f& GetDepthRange() const { return depth_range_; }
    void SetDepthRange(float min, float max) { depth_range_ = Vector2f(min, max); }

    // CUDA IPC handles of this rank's buffers, to be sent to every peer
    vkShardHandles Export() const;
    // handles of all ranks in rank order (own slot ignored)
    void Connect(const std::vector<vkShardHandles>& all_ranks);
    // all ranks live in this process
    static void ConnectLocal(const std::vector<ShardedTracer*>& ranks);

    // Tracer::Trace of the sharded volume (collective).  gather = false: only this rank's band of rows
    // (GetBand) is rendered and nothing but the 38 KB ray-bounds grids is exchanged (the consumer is the host).
    void Trace(const Frame& view, Images& images, bool gather = true);
    // the same in three parts for ranks that share a GPU: every rank part 1, then every rank part 2, then 3
    void TraceStage(const Frame& view, Images& images, int stage, bool gather = true);
    // rows [begin, end) rendered by this rank in the last Trace (synchronises the volume's stream)
    void GetBand(int& begin, int& end) const;

  private:

    ShardedTracer(const ShardedTracer&);
    ShardedTracer& operator=(const ShardedTracer&);

    std::shared_ptr<Volume> volume_;
    vkShard shard_;
    Vector2f depth_range_;
};

} // namespace vulcan
