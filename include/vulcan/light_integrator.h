// vulcan/light_integrator.h (reference: include/vulcan/light_integrator.h:10-37, src/light_integrator.cu):
// frame mask, depth update and albedo update (colour divided by the point-light shading).
#pragma once

#include <vulcan/image.h>
#include <vulcan/integrator.h>
#include <vulcan/light.h>

namespace vulcan
{

class LightIntegrator : public Integrator
{
  public:

    LightIntegrator(std::shared_ptr<Volume> volume);

    const Light& GetLight() const;
    void SetLight(const Light& light);

    void Integrate(const Frame& frame) override;

  protected:

    void ComputeFrameMask(const Frame& frame);
    void IntegrateDepth(const Frame& frame);

  protected:

    Light light_;
    Image frame_mask_;
    float depth_threshold_;
};

} // namespace vulcan
