// vulcan::ShardedTracer (include/vulcan/sharded_tracer.h): C++ host of the block-sharded mode over vk_shard_*.
#include <vulcan/vulcan.h>
#include <vulcan/sharded_tracer.h>

#include <cstring>

namespace vulcan
{

namespace
{

vkFrame PoseOnly(const Frame& view, int w, int h)
{
  vkFrame f;
  std::memset(&f, 0, sizeof(f));
  f.depth_w = w;
  f.depth_h = h;
  std::memcpy(&f.depth_proj, &view.depth_projection, sizeof(vkProjection));
  std::memcpy(&f.depth_to_world, &view.depth_to_world_transform, sizeof(vkTransform));
  return f;
}

} // namespace

ShardedTracer::ShardedTracer(std::shared_ptr<Volume> volume, int rank, int count, int width, int height, int brick_shift) :
  volume_(volume), depth_range_(0.1f, 5.0f)
{
  VULCAN_ASSERT(volume_ != nullptr);
  VK_ASSERT(vk_shard_create(&shard_, volume_->GetHandle(), rank, count, brick_shift, width, height, volume_->GetStream()));
}

ShardedTracer::~ShardedTracer()
{
  cudaStreamSynchronize(volume_->GetStream());
  vk_shard_destroy(&shard_, volume_->GetHandle());
}

vkShardHandles ShardedTracer::Export() const
{
  vkShardHandles out;
  VK_ASSERT(vk_shard_export(&shard_, volume_->GetHandle(), &out));
  return out;
}

void ShardedTracer::Connect(const std::vector<vkShardHandles>& all_ranks)
{
  VULCAN_ASSERT_MSG((int)all_ranks.size() == shard_.count, "one set of handles per rank");
  VK_ASSERT(vk_shard_connect(&shard_, all_ranks.data()));
}

void ShardedTracer::ConnectLocal(const std::vector<ShardedTracer*>& ranks)
{
  std::vector<vkShard*> shards;
  std::vector<vkVolume*> volumes;
  for (ShardedTracer* r : ranks)
  {
    shards.push_back(&r->shard_);
    volumes.push_back(r->volume_->GetHandle());
  }
  VK_ASSERT(vk_shard_connect_local(shards.data(), volumes.data(), (int)ranks.size()));
}

void ShardedTracer::Trace(const Frame& view, Images& images, bool gather)
{
  const vkFrame f = PoseOnly(view, shard_.width, shard_.height);
  volume_->SyncDistances();
  VK_ASSERT(vk_trace_sharded(volume_->GetHandle(), &f, depth_range_[0], depth_range_[1], gather ? 1 : 0, &images.depth,
                             &images.color, &images.normal, volume_->GetStream()));
  images.width = shard_.width;
  images.height = shard_.height;
}

void ShardedTracer::TraceStage(const Frame& view, Images& images, int stage, bool gather)
{
  const vkFrame f = PoseOnly(view, shard_.width, shard_.height);
  if (stage == 1) volume_->SyncDistances();
  VK_ASSERT(vk_trace_sharded_staged(volume_->GetHandle(), &f, depth_range_[0], depth_range_[1], gather ? 1 : 0, stage,
                                    &images.depth, &images.color, &images.normal, volume_->GetStream()));
  images.width = shard_.width;
  images.height = shard_.height;
}

void ShardedTracer::GetBand(int& begin, int& end) const
{
  VK_ASSERT(vk_shard_band(volume_->GetHandle(), &begin, &end, volume_->GetStream()));
}

} // namespace vulcan
