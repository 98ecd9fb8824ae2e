// vulcan/tracker.h - frame-to-keyframe pose tracker base (reference: include/vulcan/tracker.h:16-103,
// src/tracker.cpp).  Public surface as in the reference.  The reference's protected hooks that pass
// Eigen types (ComputeUpdate / ApplyUpdate with Eigen::VectorXf) are not reproduced: the solve and
// the pose update run inside the iteration kernel (vk_track_depth), one kernel per iteration.
#pragma once

#include <memory>
#include <vulcan/buffer.h>
#include <vulcan/device.h>
#include <vulcan/matrix.h>

namespace vulcan
{

struct Frame;

class Tracker
{
  public:

    Tracker();
    virtual ~Tracker();

    std::shared_ptr<const Frame> GetKeyframe() const { return keyframe_; }
    void SetKeyframe(std::shared_ptr<const Frame> keyframe) { keyframe_ = keyframe; }

    bool GetTranslationEnabled() const { return translation_enabled_; }
    void SetTranslationEnabled(bool enabled) { translation_enabled_ = enabled; }

    int GetMaxIterations() const { return max_iterations_; }
    void SetMaxIterations(int iterations);

    // updates frame.depth_to_world_transform in place (src/tracker.cpp:57-67)
    void Track(Frame& frame);

    // additions: systems solved by the last Track call and the norm of its last update
    int GetIterations() const { return iterations_; }
    float GetLastUpdateNorm() const { return last_update_norm_; }

  protected:

    void ValidateKeyframe() const;
    void ValidateFrame(const Frame& frame) const;
    virtual void Solve(Frame& frame) = 0;
    int GetParameterCount() const { return translation_enabled_ ? 6 : 3; }

    bool translation_enabled_;
    int max_iterations_;
    int iterations_;
    float last_update_norm_;
    std::shared_ptr<const Frame> keyframe_;
    Buffer<float> workspace_;
};

} // namespace vulcan
