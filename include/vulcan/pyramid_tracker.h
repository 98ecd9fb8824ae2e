// vulcan/pyramid_tracker.h - coarse-to-fine schedule over a tracker (reference:
// include/vulcan/pyramid_tracker.h:11-50, src/pyramid_tracker.cpp:52-90): both pyramids are built,
// the half-size level runs 15 iterations, the full size 20.  Header-only here; the reference
// instantiates it for its three trackers, of which DepthTracker exists in this library.
#pragma once

#include <memory>
#include <vulcan/exception.h>
#include <vulcan/frame.h>

namespace vulcan
{

template <typename Tracker>
class PyramidTracker
{
  public:

    PyramidTracker() :
      tracker_(std::make_shared<Tracker>()),
      half_keyframe_(std::make_shared<Frame>()),
      quarter_keyframe_(std::make_shared<Frame>()),
      iter_(0)
    {
    }

    PyramidTracker(std::shared_ptr<Tracker> tracker) :
      tracker_(tracker),
      half_keyframe_(std::make_shared<Frame>()),
      quarter_keyframe_(std::make_shared<Frame>()),
      iter_(0)
    {
    }

    virtual ~PyramidTracker() {}

    std::shared_ptr<const Tracker> GetTracker() const { return tracker_; }
    std::shared_ptr<const Frame> GetKeyframe() const { return keyframe_; }
    void SetKeyframe(std::shared_ptr<const Frame> keyframe) { keyframe_ = keyframe; }

    void Track(Frame& frame)
    {
      VULCAN_ASSERT_MSG(keyframe_ != nullptr, "keyframe has not been assigned");
      Frame half_frame;
      Frame quarter_frame;
      frame.Downsample(half_frame);
      half_frame.Downsample(quarter_frame);
      keyframe_->Downsample(*half_keyframe_);
      half_keyframe_->Downsample(*quarter_keyframe_);

      tracker_->SetMaxIterations(15);
      tracker_->SetTranslationEnabled(true);
      half_frame.depth_to_world_transform = quarter_frame.depth_to_world_transform;
      tracker_->SetKeyframe(half_keyframe_);
      tracker_->Track(half_frame);

      tracker_->SetMaxIterations(20);
      tracker_->SetTranslationEnabled(true);
      frame.depth_to_world_transform = half_frame.depth_to_world_transform;
      tracker_->SetKeyframe(keyframe_);
      tracker_->Track(frame);
    }

  protected:

    std::shared_ptr<Tracker> tracker_;
    std::shared_ptr<const Frame> keyframe_;
    std::shared_ptr<Frame> half_keyframe_;
    std::shared_ptr<Frame> quarter_keyframe_;
    int iter_;
};

} // namespace vulcan
