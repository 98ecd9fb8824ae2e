// vulcan/depth_tracker.h - projective point-to-plane depth tracker (reference:
// include/vulcan/depth_tracker.h:9-27, src/depth_tracker.cpp, src/depth_tracker.cu)
#pragma once

#include <vulcan/matrix.h>
#include <vulcan/tracker.h>

namespace vulcan
{

class DepthTracker : public Tracker
{
  public:

    DepthTracker();
    virtual ~DepthTracker();

    // one float / one Vector6f per frame pixel (src/depth_tracker.cu:279-344)
    void ComputeResiduals(const Frame& frame, Buffer<float>& residuals) const;
    void ComputeJacobian(const Frame& frame, Buffer<Vector6f>& jacobian) const;

    // addition: packed lower triangle of J^T J (21 or 6 floats) and J^T r (6 or 3), device buffers
    void ComputeSystem(const Frame& frame, Buffer<float>& hessian, Buffer<float>& gradient);

  protected:

    void Solve(Frame& frame) override;
};

} // namespace vulcan
