// vulcan/depth_integrator.h (reference: include/vulcan/depth_integrator.h:8-15, src/depth_integrator.cu)
#pragma once

#include <vulcan/integrator.h>

namespace vulcan
{

class DepthIntegrator : public Integrator
{
  public:

    DepthIntegrator(std::shared_ptr<Volume> volume);
    void Integrate(const Frame& frame) override;
};

} // namespace vulcan
