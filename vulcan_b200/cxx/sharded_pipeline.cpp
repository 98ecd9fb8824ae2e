// vulcan::ShardedPipeline (include/vulcan/sharded_pipeline.h) and its C entry points: host frames in, host image
// bands out, for one rank of a block-sharded volume.  Plain C++ over the drop-in classes and ShardedTracer.
#include <vulcan/vulcan.h>
#include <vulcan/frame_pipeline.h>
#include <vulcan/sharded_pipeline.h>

#include <chrono>
#include <cstdlib>
#include <cstring>
#include <string>

namespace vulcan
{

struct ShardedPipeline::Slot
{
  Frame in;                      // device depth image of the slot's frame
  ShardedTracer::Images out;     // the tracer's images of that view (inside its exchange buffer)
  float* h_depth;
  float* h_color;
  float* h_normal;
  int* h_bands;                  // [17] band table + handshakes given up
  cudaEvent_t uploaded, integrated, computed, bandinfo, downloaded;
  int frame;
  bool issued;                   // the band copy of `frame` has been enqueued
  int begin, end;
};

ShardedPipeline::ShardedPipeline(std::shared_ptr<Volume> volume, std::shared_ptr<DepthIntegrator> integrator,
    std::shared_ptr<ShardedTracer> tracer, int width, int height, int depth) :
  volume_(volume), integrator_(integrator), tracer_(tracer), width_(width), height_(height), next_(0)
{
  VULCAN_ASSERT_MSG(depth >= 3, "a sharded pipeline needs at least 3 slots");
  VULCAN_ASSERT_MSG(width > 0 && height > 0, "invalid image size");
  CUDA_ASSERT(cudaStreamCreateWithFlags(&upload_, cudaStreamNonBlocking));
  CUDA_ASSERT(cudaStreamCreateWithFlags(&compute_, cudaStreamNonBlocking));
  CUDA_ASSERT(cudaStreamCreateWithFlags(&download_, cudaStreamNonBlocking));
  CUDA_ASSERT(cudaStreamCreateWithFlags(&info_, cudaStreamNonBlocking));
  previous_ = volume_->GetStream();
  CUDA_ASSERT(cudaStreamSynchronize(previous_));
  volume_->SetStream(compute_);
  const size_t n = (size_t)width * height;
  for (int i = 0; i < depth; ++i)
  {
    Slot* s = new Slot();
    s->in.depth_image = std::make_shared<Image>(width, height);
    s->h_depth = FramePipeline::AllocHost(n);
    s->h_color = FramePipeline::AllocHost(3 * n);
    s->h_normal = FramePipeline::AllocHost(3 * n);
    s->h_bands = reinterpret_cast<int*>(FramePipeline::AllocHost(32));
    std::memset(s->h_bands, 0, 32 * sizeof(float));
    CUDA_ASSERT(cudaEventCreateWithFlags(&s->uploaded, cudaEventDisableTiming));
    CUDA_ASSERT(cudaEventCreateWithFlags(&s->integrated, cudaEventDisableTiming));
    CUDA_ASSERT(cudaEventCreateWithFlags(&s->computed, cudaEventDisableTiming));
    CUDA_ASSERT(cudaEventCreateWithFlags(&s->bandinfo, cudaEventDisableTiming));
    CUDA_ASSERT(cudaEventCreateWithFlags(&s->downloaded, cudaEventDisableTiming));
    s->frame = -1;
    s->issued = true;
    s->begin = s->end = 0;
    slots_.push_back(s);
  }
}

ShardedPipeline::~ShardedPipeline()
{
  cudaStreamSynchronize(upload_);
  cudaStreamSynchronize(compute_);
  cudaStreamSynchronize(download_);
  cudaStreamSynchronize(info_);
  volume_->SetStream(previous_);
  for (Slot* s : slots_)
  {
    FramePipeline::FreeHost(s->h_depth);
    FramePipeline::FreeHost(s->h_color);
    FramePipeline::FreeHost(s->h_normal);
    FramePipeline::FreeHost(reinterpret_cast<float*>(s->h_bands));
    cudaEventDestroy(s->uploaded);
    cudaEventDestroy(s->integrated);
    cudaEventDestroy(s->computed);
    cudaEventDestroy(s->bandinfo);
    cudaEventDestroy(s->downloaded);
    delete s;
  }
  cudaStreamDestroy(upload_);
  cudaStreamDestroy(compute_);
  cudaStreamDestroy(download_);
  cudaStreamDestroy(info_);
}

// Enqueues the copy of frame k's band; needs its band table on the host (waits for that 68-byte copy).
void ShardedPipeline::IssueDownload(int k)
{
  Slot& s = *slots_[k % slots_.size()];
  if (s.frame != k || s.issued) return;
  CUDA_ASSERT(cudaEventSynchronize(s.bandinfo));
  VULCAN_ASSERT_MSG(s.h_bands[16] == 0, "sharded volume: a peer did not arrive at a handshake (VK_E_TIMEOUT)");
  s.begin = s.h_bands[tracer_->GetRank()];
  s.end = s.h_bands[tracer_->GetRank() + 1];
  // diagnostics only (VULCAN_PIPELINE_DEBUG): bit 0 = skip the downloads, bit 1 = skip the uploads
  static const int debug = std::getenv("VULCAN_PIPELINE_DEBUG") ? std::atoi(std::getenv("VULCAN_PIPELINE_DEBUG")) : 0;
  // The copy starts when the NEXT frame (if one has been submitted) has finished integrating: allocation and
  // integration store into the peers' replicas, and those stores queue behind download traffic on their way out
  // of the GPU (measured: +50 us per frame at 2 GPUs); the ray cast that follows reads local memory only.
  if (k + 1 < next_) CUDA_ASSERT(cudaStreamWaitEvent(download_, slots_[(k + 1) % slots_.size()]->integrated, 0));
  if (s.end > s.begin && !(debug & 1))
  {
    const size_t first = (size_t)s.begin * width_, count = (size_t)(s.end - s.begin) * width_;
    CUDA_ASSERT(cudaMemcpyAsync(s.h_depth + first, s.out.depth + first, count * sizeof(float), cudaMemcpyDeviceToHost, download_));
    CUDA_ASSERT(cudaMemcpyAsync(s.h_color + 3 * first, s.out.color + 3 * first, 3 * count * sizeof(float), cudaMemcpyDeviceToHost, download_));
    CUDA_ASSERT(cudaMemcpyAsync(s.h_normal + 3 * first, s.out.normal + 3 * first, 3 * count * sizeof(float), cudaMemcpyDeviceToHost, download_));
  }
  CUDA_ASSERT(cudaEventRecord(s.downloaded, download_));
  s.issued = true;
}

int ShardedPipeline::Submit(const float* host_depth, const Frame& frame)
{
  VULCAN_ASSERT_MSG(host_depth != nullptr, "missing depth image");
  const int k = next_;
  Slot& s = *slots_[k % slots_.size()];
  if (s.frame >= 0)
  {
    IssueDownload(s.frame);
    CUDA_ASSERT(cudaEventSynchronize(s.downloaded));   // the slot's previous frame has left the device
  }
  // the band of frame k-2 must be on its way before Trace(k) reuses that view's output images
  if (k >= 2) IssueDownload(k - 2);
  s.frame = k;
  s.issued = false;
  const size_t n = (size_t)width_ * height_;
  static const int debug = std::getenv("VULCAN_PIPELINE_DEBUG") ? std::atoi(std::getenv("VULCAN_PIPELINE_DEBUG")) : 0;
  if (!(debug & 2) || k < (int)slots_.size())
    CUDA_ASSERT(cudaMemcpyAsync(s.in.depth_image->GetData(), host_depth, n * sizeof(float), cudaMemcpyHostToDevice, upload_));
  CUDA_ASSERT(cudaEventRecord(s.uploaded, upload_));
  s.in.depth_projection = frame.depth_projection;
  s.in.depth_to_world_transform = frame.depth_to_world_transform;
  CUDA_ASSERT(cudaStreamWaitEvent(compute_, s.uploaded, 0));
  if (k >= 2) CUDA_ASSERT(cudaStreamWaitEvent(compute_, slots_[(k - 2) % slots_.size()]->downloaded, 0));
  volume_->SetView(s.in);
  integrator_->Integrate(s.in);
  CUDA_ASSERT(cudaEventRecord(s.integrated, compute_));
  tracer_->Trace(s.in, s.out, false);
  CUDA_ASSERT(cudaEventRecord(s.computed, compute_));
  // the band table travels on its own stream: on the download stream it would queue the band of frame k-1 (issued
  // by the next Submit) behind this frame's kernels
  CUDA_ASSERT(cudaStreamWaitEvent(info_, s.computed, 0));
  VK_ASSERT(vk_shard_band_async(volume_->GetHandle(), s.h_bands, info_));
  CUDA_ASSERT(cudaEventRecord(s.bandinfo, info_));
  next_ = k + 1;
  return k;
}

ShardedPipeline::Band ShardedPipeline::Result(int k)
{
  VULCAN_ASSERT_MSG(k >= 0 && k < next_, "frame has not been submitted");
  Slot& s = *slots_[k % slots_.size()];
  VULCAN_ASSERT_MSG(s.frame == k, "frame is no longer in the pipeline");
  IssueDownload(k);
  CUDA_ASSERT(cudaEventSynchronize(s.downloaded));
  Band out = {s.begin, s.end, s.h_depth, s.h_color, s.h_normal};
  return out;
}

void ShardedPipeline::Drain()
{
  const int first = next_ > (int)slots_.size() ? next_ - (int)slots_.size() : 0;
  for (int k = first; k < next_; ++k) IssueDownload(k);
  CUDA_ASSERT(cudaStreamSynchronize(upload_));
  CUDA_ASSERT(cudaStreamSynchronize(compute_));
  CUDA_ASSERT(cudaStreamSynchronize(download_));
  CUDA_ASSERT(cudaStreamSynchronize(info_));
}

} // namespace vulcan

// ---- C entry points (what bench.py --gpus N and other FFI hosts bind) --------------------------------------

namespace
{

struct ShardedHandle
{
  std::shared_ptr<vulcan::Volume> volume;
  std::shared_ptr<vulcan::DepthIntegrator> integrator;
  std::shared_ptr<vulcan::ShardedTracer> tracer;
  std::unique_ptr<vulcan::ShardedPipeline> pipeline;
  int width, height, depth;
  int taken;
  std::string error;
};

vulcan::Frame PoseFrame(const float proj4[4], const float pose32[32])
{
  vulcan::Frame f;
  f.depth_projection.SetFocalLength(proj4[0], proj4[1]);
  f.depth_projection.SetCenterPoint(proj4[2], proj4[3]);
  vulcan::Matrix4f m, inv;
  std::memcpy(static_cast<void*>(&m), pose32, 64);
  std::memcpy(static_cast<void*>(&inv), pose32 + 16, 64);
  f.depth_to_world_transform = vulcan::Transform::FromMatrices(m, inv);
  return f;
}

} // namespace

extern "C" {

int vulcan_shard_handles_bytes(void) { return (int)sizeof(vkShardHandles); }

// One rank of a sharded volume + DepthIntegrator + ShardedTracer.  The pipeline itself is created by
// vulcan_sharded_pipeline_connect, once the ranks have exchanged their handles.  NULL on failure.
void* vulcan_sharded_pipeline_create(int main_blocks, int excess_blocks, float voxel_length, float trunc_length, int width,
                                     int height, int depth, int rank, int count)
{
  try
  {
    ShardedHandle* h = new ShardedHandle();
    h->volume = std::make_shared<vulcan::Volume>(main_blocks, excess_blocks);
    h->volume->SetVoxelLength(voxel_length);
    h->volume->SetTruncationLength(trunc_length);
    h->integrator = std::make_shared<vulcan::DepthIntegrator>(h->volume);
    h->tracer = std::make_shared<vulcan::ShardedTracer>(h->volume, rank, count, width, height);
    h->width = width; h->height = height; h->depth = depth;
    h->taken = 0;
    return h;
  }
  catch (...)
  {
    return nullptr;
  }
}

// this rank's CUDA IPC handles (vulcan_shard_handles_bytes() bytes), to be sent to every rank
int vulcan_sharded_pipeline_export(void* handle, void* out)
{
  ShardedHandle* h = (ShardedHandle*)handle;
  try
  {
    const vkShardHandles hs = h->tracer->Export();
    std::memcpy(out, &hs, sizeof(hs));
    return 0;
  }
  catch (const std::exception& e) { h->error = e.what(); return -1; }
}

// the handles of all ranks in rank order (count x vulcan_shard_handles_bytes() bytes); creates the pipeline
int vulcan_sharded_pipeline_connect(void* handle, const void* all_ranks)
{
  ShardedHandle* h = (ShardedHandle*)handle;
  try
  {
    std::vector<vkShardHandles> all(h->tracer->GetCount());
    std::memcpy(all.data(), all_ranks, all.size() * sizeof(vkShardHandles));
    if (h->tracer->GetCount() > 1) h->tracer->Connect(all);
    h->pipeline.reset(new vulcan::ShardedPipeline(h->volume, h->integrator, h->tracer, h->width, h->height, h->depth));
    return 0;
  }
  catch (const std::exception& e) { h->error = e.what(); return -1; }
}

// The streaming loop of a host application in one call (cf. vulcan_pipeline_run): `frames` frames are submitted in
// order; after submitting frame k the host takes the band of frame k - lag (lag >= 2; each frame exactly once, in
// order) and reads the centre pixel of the band's middle row in each image; band_rows[i] (may be NULL) receives
// the number of rows this rank rendered of frame i.  Every rank makes the same call.  Returns 0, or -1.
int vulcan_sharded_pipeline_run(void* handle, int frames, const float* const* host_depths, const float* proj4,
                                const float* poses32, int lag, double* checksum, double* submit_us, int* band_rows)
{
  ShardedHandle* h = (ShardedHandle*)handle;
  try
  {
    VULCAN_ASSERT_MSG(h->pipeline != nullptr, "vulcan_sharded_pipeline_connect has not been called");
    VULCAN_ASSERT_MSG(lag >= 2 && lag < h->depth, "lag must be in [2, depth)");
    vulcan::ShardedPipeline& p = *h->pipeline;
    const int w = p.GetWidth();
    double acc = 0;
    int base = -1;
    auto take = [&](int k)
    {
      const vulcan::ShardedPipeline::Band b = p.Result(k);
      if (band_rows && base >= 0 && k >= base) band_rows[k - base] = b.end - b.begin;
      if (b.end > b.begin)
      {
        const size_t centre = (size_t)((b.begin + b.end) / 2) * w + w / 2;
        acc += b.depth[centre] + b.color[3 * centre] + b.normal[3 * centre + 2];
      }
    };
    int last = -1;
    for (int i = 0; i < frames; ++i)
    {
      const auto t0 = std::chrono::steady_clock::now();
      const int k = p.Submit(host_depths[i], PoseFrame(proj4, poses32 + 32 * (size_t)i));
      const auto t1 = std::chrono::steady_clock::now();
      if (submit_us) submit_us[i] = std::chrono::duration<double, std::micro>(t1 - t0).count();
      if (base < 0) base = k;
      last = k;
      while (h->taken <= k - lag) take(h->taken++);
    }
    while (h->taken <= last) take(h->taken++);
    p.Drain();
    if (checksum) *checksum = acc;
    return 0;
  }
  catch (const std::exception& e) { h->error = e.what(); return -1; }
}

// (collective in effect: a rank's buffers are mapped by its peers - let every rank finish its work first)
void vulcan_sharded_pipeline_destroy(void* handle)
{
  ShardedHandle* h = (ShardedHandle*)handle;
  if (!h) return;
  h->pipeline.reset();
  h->tracer.reset();
  h->integrator.reset();
  h->volume.reset();
  delete h;
}

const char* vulcan_sharded_pipeline_error(void* handle) { return ((ShardedHandle*)handle)->error.c_str(); }

}  // extern "C"
