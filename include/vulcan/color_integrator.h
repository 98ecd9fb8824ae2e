// vulcan/color_integrator.h (reference: include/vulcan/color_integrator.h:8-21, src/color_integrator.cu).
// Integrate = depth update + colour update of the same voxels in ONE pass over the voxel tiles
// (the reference streams them twice through its protected IntegrateDepth / IntegrateColor pair).
// IntegrateDepth is kept for subclasses; the colour half is not separable from the fused pass.
#pragma once

#include <vulcan/integrator.h>

namespace vulcan
{

class ColorIntegrator : public Integrator
{
  public:

    ColorIntegrator(std::shared_ptr<Volume> volume);
    void Integrate(const Frame& frame) override;

  protected:

    void IntegrateDepth(const Frame& frame);
};

} // namespace vulcan
