// vulcan/frame_pipeline.h - streaming front end for host frames (addition; the reference's app does
// upload -> SetView -> Integrate -> Trace -> download on the legacy stream, one frame at a time,
// apps/vulcan/vulcan.cu:297-366).
//
// At 640x480 the traced depth + colour + normal images are 8.6 MB per frame: copying them to the host
// takes about as long as all kernels of the frame.  FramePipeline keeps three streams busy instead:
//
//     upload   : pinned host depth (+ colour) of frame k+1  -> device slot (k+1) % depth
//     compute  : SetView, Integrate, Trace of frame k        -> output slot k % depth
//     download : traced images of frame k-1                  -> pinned host slot (k-1) % depth
//
// Every frame's copies still happen (nothing is skipped or cached); only their latency is hidden behind
// the kernels of the neighbouring frames.  Submit() enqueues a frame and never blocks on the GPU unless
// all slots are in flight; Result(k) blocks until frame k's images are in host memory and returns them
// (valid until frame k + depth is submitted).
#pragma once

#include <memory>
#include <vector>
#include <vulcan/frame.h>
#include <vulcan/integrator.h>
#include <vulcan/tracer.h>
#include <vulcan/volume.h>

namespace vulcan
{

class FramePipeline
{
  public:

    struct Images
    {
      const float* depth;   // width * height
      const float* color;   // width * height * 3
      const float* normal;  // width * height * 3
    };

    // `integrator` is a DepthIntegrator or a ColorIntegrator on `volume`.  While the pipeline lives the
    // volume works on the pipeline's compute stream (Volume::SetStream); the previous stream is restored
    // by the destructor.
    FramePipeline(std::shared_ptr<Volume> volume, std::shared_ptr<Integrator> integrator,
        std::shared_ptr<Tracer> tracer, int width, int height, int depth = 3);

    ~FramePipeline();

    int GetWidth() const { return width_; }
    int GetHeight() const { return height_; }
    int GetDepth() const { return (int)slots_.size(); }

    // Enqueues one frame: `host_depth` (width * height floats; page-locked memory for the copy to be
    // asynchronous, see AllocHost) is this frame's sensor image, `host_color` (width * height * 3) is
    // required with a ColorIntegrator; `frame` supplies the projections and transforms (its image
    // members are ignored).  Returns the frame's sequence number.
    int Submit(const float* host_depth, const Frame& frame, const float* host_color = nullptr);

    // Blocks until frame k's traced images are in host memory.
    Images Result(int k);

    // Waits for everything submitted so far.
    void Drain();

    // page-locked host memory for Submit
    static float* AllocHost(size_t floats);
    static void FreeHost(float* ptr);

  private:

    FramePipeline(const FramePipeline&);
    FramePipeline& operator=(const FramePipeline&);

    struct Slot;

    std::shared_ptr<Volume> volume_;
    std::shared_ptr<Integrator> integrator_;
    std::shared_ptr<Tracer> tracer_;
    int width_, height_;
    bool with_color_;
    std::vector<Slot*> slots_;
    cudaStream_t upload_, compute_, download_, previous_;
    int next_;
};

} // namespace vulcan
