// vulcan::FramePipeline (include/vulcan/frame_pipeline.h) and its C entry points: host frames in,
// host images out, the copies of neighbouring frames overlapped with the kernels.  Plain C++ over the
// drop-in classes - the same SetView / Integrate / Trace calls a caller of the reference makes.
#include <vulcan/vulcan.h>
#include <vulcan/frame_pipeline.h>

#include <chrono>
#include <cstdlib>
#include <cstring>

namespace vulcan
{

struct FramePipeline::Slot
{
  Frame in;    // device input images of the slot's frame
  Frame out;   // traced images
  float* h_depth;
  float* h_color;
  float* h_normal;
  cudaEvent_t uploaded, computed, downloaded;
  int frame;
};

FramePipeline::FramePipeline(std::shared_ptr<Volume> volume, std::shared_ptr<Integrator> integrator,
    std::shared_ptr<Tracer> tracer, int width, int height, int depth) :
  volume_(volume), integrator_(integrator), tracer_(tracer), width_(width), height_(height),
  with_color_(dynamic_cast<ColorIntegrator*>(integrator.get()) != nullptr), next_(0)
{
  VULCAN_ASSERT_MSG(depth >= 1, "pipeline depth must be at least 1");
  VULCAN_ASSERT_MSG(width > 0 && height > 0, "invalid image size");
  VULCAN_ASSERT_MSG(with_color_ || dynamic_cast<DepthIntegrator*>(integrator.get()) != nullptr,
      "FramePipeline drives a DepthIntegrator or a ColorIntegrator");
  CUDA_ASSERT(cudaStreamCreateWithFlags(&upload_, cudaStreamNonBlocking));
  CUDA_ASSERT(cudaStreamCreateWithFlags(&compute_, cudaStreamNonBlocking));
  CUDA_ASSERT(cudaStreamCreateWithFlags(&download_, cudaStreamNonBlocking));
  // everything enqueued so far on the volume's stream (creation, settings) comes first
  previous_ = volume_->GetStream();
  CUDA_ASSERT(cudaStreamSynchronize(previous_));
  volume_->SetStream(compute_);
  const size_t n = (size_t)width * height;
  for (int i = 0; i < depth; ++i)
  {
    Slot* s = new Slot();
    s->in.depth_image = std::make_shared<Image>(width, height);
    if (with_color_) s->in.color_image = std::make_shared<ColorImage>(width, height);
    s->out.depth_image = std::make_shared<Image>(width, height);
    s->out.color_image = std::make_shared<ColorImage>(width, height);
    s->out.normal_image = std::make_shared<ColorImage>(width, height);
    s->h_depth = AllocHost(n);
    s->h_color = AllocHost(3 * n);
    s->h_normal = AllocHost(3 * n);
    CUDA_ASSERT(cudaEventCreateWithFlags(&s->uploaded, cudaEventDisableTiming));
    CUDA_ASSERT(cudaEventCreateWithFlags(&s->computed, cudaEventDisableTiming));
    CUDA_ASSERT(cudaEventCreateWithFlags(&s->downloaded, cudaEventDisableTiming));
    s->frame = -1;
    slots_.push_back(s);
  }
}

FramePipeline::~FramePipeline()
{
  cudaStreamSynchronize(upload_);
  cudaStreamSynchronize(compute_);
  cudaStreamSynchronize(download_);
  volume_->SetStream(previous_);
  for (Slot* s : slots_)
  {
    FreeHost(s->h_depth);
    FreeHost(s->h_color);
    FreeHost(s->h_normal);
    cudaEventDestroy(s->uploaded);
    cudaEventDestroy(s->computed);
    cudaEventDestroy(s->downloaded);
    delete s;
  }
  cudaStreamDestroy(upload_);
  cudaStreamDestroy(compute_);
  cudaStreamDestroy(download_);
}

float* FramePipeline::AllocHost(size_t floats)
{
  float* p = nullptr;
  CUDA_ASSERT(cudaMallocHost((void**)&p, floats * sizeof(float)));
  return p;
}

void FramePipeline::FreeHost(float* ptr)
{
  if (ptr) cudaFreeHost(ptr);
}

int FramePipeline::Submit(const float* host_depth, const Frame& frame, const float* host_color)
{
  VULCAN_ASSERT_MSG(host_depth != nullptr, "missing depth image");
  VULCAN_ASSERT_MSG(!with_color_ || host_color != nullptr, "missing color image");
  const int k = next_;
  // diagnostics only (VULCAN_PIPELINE_DEBUG): bit 0 = skip the downloads, bit 1 = skip the uploads
  static const int debug = std::getenv("VULCAN_PIPELINE_DEBUG") ? std::atoi(std::getenv("VULCAN_PIPELINE_DEBUG")) : 0;
  Slot& s = *slots_[k % slots_.size()];
  if (s.frame >= 0) CUDA_ASSERT(cudaEventSynchronize(s.downloaded));  // the slot's previous frame has left the device
  s.frame = k;
  const size_t n = (size_t)width_ * height_;
  // upload
  if (!(debug & 2) || k < (int)slots_.size())
    CUDA_ASSERT(cudaMemcpyAsync(s.in.depth_image->GetData(), host_depth, n * sizeof(float), cudaMemcpyHostToDevice, upload_));
  if (with_color_)
    CUDA_ASSERT(cudaMemcpyAsync(s.in.color_image->GetData(), host_color, 3 * n * sizeof(float), cudaMemcpyHostToDevice, upload_));
  CUDA_ASSERT(cudaEventRecord(s.uploaded, upload_));
  // compute: the calls a caller of the reference makes (apps/vulcan/vulcan.cu:316-325)
  s.in.depth_projection = frame.depth_projection;
  s.in.color_projection = frame.color_projection;
  s.in.depth_to_world_transform = frame.depth_to_world_transform;
  s.in.depth_to_color_transform = frame.depth_to_color_transform;
  s.out.depth_projection = frame.depth_projection;
  s.out.depth_to_world_transform = frame.depth_to_world_transform;
  CUDA_ASSERT(cudaStreamWaitEvent(compute_, s.uploaded, 0));
  volume_->SetView(s.in);
  integrator_->Integrate(s.in);
  tracer_->Trace(s.out);
  CUDA_ASSERT(cudaEventRecord(s.computed, compute_));
  // download
  CUDA_ASSERT(cudaStreamWaitEvent(download_, s.computed, 0));
  if (!(debug & 1))
  {
    CUDA_ASSERT(cudaMemcpyAsync(s.h_depth, s.out.depth_image->GetData(), n * sizeof(float), cudaMemcpyDeviceToHost, download_));
    CUDA_ASSERT(cudaMemcpyAsync(s.h_color, s.out.color_image->GetData(), 3 * n * sizeof(float), cudaMemcpyDeviceToHost, download_));
    CUDA_ASSERT(cudaMemcpyAsync(s.h_normal, s.out.normal_image->GetData(), 3 * n * sizeof(float), cudaMemcpyDeviceToHost, download_));
  }
  CUDA_ASSERT(cudaEventRecord(s.downloaded, download_));
  next_ = k + 1;
  return k;
}

FramePipeline::Images FramePipeline::Result(int k)
{
  VULCAN_ASSERT_MSG(k >= 0 && k < next_, "frame has not been submitted");
  Slot& s = *slots_[k % slots_.size()];
  VULCAN_ASSERT_MSG(s.frame == k, "frame is no longer in the pipeline");
  CUDA_ASSERT(cudaEventSynchronize(s.downloaded));
  Images out = {s.h_depth, s.h_color, s.h_normal};
  return out;
}

void FramePipeline::Drain()
{
  CUDA_ASSERT(cudaStreamSynchronize(upload_));
  CUDA_ASSERT(cudaStreamSynchronize(compute_));
  CUDA_ASSERT(cudaStreamSynchronize(download_));
}

} // namespace vulcan

// ---- C entry points (what bench.py's end-to-end leg and other FFI hosts bind) ------------------------

namespace
{

struct PipelineHandle
{
  std::shared_ptr<vulcan::Volume> volume;
  std::shared_ptr<vulcan::DepthIntegrator> integrator;
  std::shared_ptr<vulcan::Tracer> tracer;
  std::unique_ptr<vulcan::FramePipeline> pipeline;
  int taken;   // frames whose images the host has taken so far (vulcan_pipeline_run)
  std::string error;
};

vulcan::Frame MakeFrame(const float proj4[4], const float pose32[32])
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

// A volume + DepthIntegrator + Tracer + FramePipeline (frames in flight = depth).  NULL on failure.
void* vulcan_pipeline_create(int main_blocks, int excess_blocks, float voxel_length, float trunc_length, int width,
                             int height, int depth)
{
  try
  {
    PipelineHandle* h = new PipelineHandle();
    h->volume = std::make_shared<vulcan::Volume>(main_blocks, excess_blocks);
    h->volume->SetVoxelLength(voxel_length);
    h->volume->SetTruncationLength(trunc_length);
    h->integrator = std::make_shared<vulcan::DepthIntegrator>(h->volume);
    h->tracer = std::make_shared<vulcan::Tracer>(h->volume);
    h->pipeline.reset(new vulcan::FramePipeline(h->volume, h->integrator, h->tracer, width, height, depth));
    h->taken = 0;
    return h;
  }
  catch (...)
  {
    return nullptr;
  }
}

void vulcan_pipeline_destroy(void* handle)
{
  delete (PipelineHandle*)handle;
}

float* vulcan_pipeline_alloc_host(size_t floats)
{
  try { return vulcan::FramePipeline::AllocHost(floats); } catch (...) { return nullptr; }
}

void vulcan_pipeline_free_host(float* ptr) { vulcan::FramePipeline::FreeHost(ptr); }

// proj4 = fx fy cx cy; pose32 = depth-to-world matrix and its inverse, column major (byte image of
// vulcan::Transform).  Returns the frame's sequence number, or -1.
int vulcan_pipeline_submit(void* handle, const float* host_depth, const float* proj4, const float* pose32)
{
  PipelineHandle* h = (PipelineHandle*)handle;
  try { return h->pipeline->Submit(host_depth, MakeFrame(proj4, pose32)); }
  catch (const std::exception& e) { h->error = e.what(); return -1; }
}

int vulcan_pipeline_result(void* handle, int k, const float** depth, const float** color, const float** normal)
{
  PipelineHandle* h = (PipelineHandle*)handle;
  try
  {
    const vulcan::FramePipeline::Images im = h->pipeline->Result(k);
    *depth = im.depth; *color = im.color; *normal = im.normal;
    return 0;
  }
  catch (const std::exception& e) { h->error = e.what(); return -1; }
}

int vulcan_pipeline_drain(void* handle)
{
  PipelineHandle* h = (PipelineHandle*)handle;
  try { h->pipeline->Drain(); return 0; } catch (const std::exception& e) { h->error = e.what(); return -1; }
}

// The streaming loop of a host application in one call: `frames` frames are submitted in order
// (host_depths[i]: page-locked width * height floats, poses32 + 32 * i); after submitting frame k the host
// takes the images of frame k - lag (each frame exactly once, in order) and reads the centre pixel of each
// image; the remaining frames are taken at the end and the pipeline is drained.  checksum receives the sum
// of the pixels read, submit_us[i] the host time of Submit for frame i (may be NULL).  Frame numbers
// continue over successive calls.  Returns 0, or -1 (vulcan_pipeline_error).
int vulcan_pipeline_run(void* handle, int frames, const float* const* host_depths, const float* proj4,
                        const float* poses32, int lag, double* checksum, double* submit_us)
{
  PipelineHandle* h = (PipelineHandle*)handle;
  try
  {
    vulcan::FramePipeline& p = *h->pipeline;
    const int w = p.GetWidth(), ht = p.GetHeight();
    const size_t centre = (size_t)(ht / 2) * w + w / 2;
    double acc = 0;
    auto take = [&](int k)
    {
      const vulcan::FramePipeline::Images im = p.Result(k);
      acc += im.depth[centre] + im.color[3 * centre] + im.normal[3 * centre + 2];
    };
    int first = -1, last = -1;
    for (int i = 0; i < frames; ++i)
    {
      const auto t0 = std::chrono::steady_clock::now();
      const int k = p.Submit(host_depths[i], MakeFrame(proj4, poses32 + 32 * (size_t)i));
      const auto t1 = std::chrono::steady_clock::now();
      if (submit_us) submit_us[i] = std::chrono::duration<double, std::micro>(t1 - t0).count();
      if (first < 0) first = k;
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

int vulcan_pipeline_visible_blocks(void* handle)
{
  PipelineHandle* h = (PipelineHandle*)handle;
  try { return h->volume->GetVisibleBlockCount(); } catch (...) { return -1; }
}

const char* vulcan_pipeline_error(void* handle) { return ((PipelineHandle*)handle)->error.c_str(); }

}  // extern "C"
