// vulcan/sharded_pipeline.h - host frames in, host image BANDS out, for one rank of a block-sharded volume
// (addition; the reference is single-GPU).  The streamed counterpart of FramePipeline for ShardedTracer.
//
// The consumer is the host, so no image is exchanged between the GPUs: every rank uploads the depth image,
// ray-casts its band of rows (ShardedTracer::Trace with gather = false) and downloads that band only - the ranks'
// bands together are one image set.  Three streams per rank:
//
//     upload   : pinned host depth of frame k+1              -> device slot (k+1) % depth
//     compute  : SetView, Integrate, sharded Trace of frame k
//     download : the band of frame k-1 -> pinned host slot   (+ a fourth stream for the 68-byte band tables)
//
// Which rows a rank renders is decided on the device (the bands are balanced from the merged ray bounds), so a
// frame's band table is fetched asynchronously (vk_shard_band_async) and its band is copied one Submit later,
// while the next frame computes.  The tracer's output images of a view stay valid until the second-next Trace,
// which therefore waits for that copy.  Every rank submits the same frames in the same order (Trace is
// collective).
#pragma once

#include <memory>
#include <vector>
#include <vulcan/depth_integrator.h>
#include <vulcan/frame.h>
#include <vulcan/sharded_tracer.h>
#include <vulcan/volume.h>

namespace vulcan
{

class ShardedPipeline
{
  public:

    // full-size page-locked host images of which rows [begin, end) are this rank's band of the frame
    struct Band
    {
      int begin, end;
      const float* depth;   // width * height
      const float* color;   // width * height * 3
      const float* normal;  // width * height * 3
    };

    // depth >= 3 slots.  While the pipeline lives the volume works on the pipeline's compute stream.
    ShardedPipeline(std::shared_ptr<Volume> volume, std::shared_ptr<DepthIntegrator> integrator,
        std::shared_ptr<ShardedTracer> tracer, int width, int height, int depth = 3);

    ~ShardedPipeline();

    int GetWidth() const { return width_; }
    int GetHeight() const { return height_; }

    // Enqueues one frame (host_depth: width * height floats, page-locked: FramePipeline::AllocHost); `frame`
    // supplies the projection and the pose.  Returns the frame's sequence number.
    int Submit(const float* host_depth, const Frame& frame);

    // Blocks until frame k's band is in host memory.  Throws if a peer never arrived at a handshake.
    Band Result(int k);

    // Waits for everything submitted so far.
    void Drain();

  private:

    ShardedPipeline(const ShardedPipeline&);
    ShardedPipeline& operator=(const ShardedPipeline&);

    struct Slot;
    void IssueDownload(int k);

    std::shared_ptr<Volume> volume_;
    std::shared_ptr<DepthIntegrator> integrator_;
    std::shared_ptr<ShardedTracer> tracer_;
    int width_, height_;
    std::vector<Slot*> slots_;
    cudaStream_t upload_, compute_, download_, info_, previous_;
    int next_;
};

} // namespace vulcan
