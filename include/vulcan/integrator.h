// vulcan/integrator.h - base of the per-frame integrators (reference: include/vulcan/integrator.h:16-45,
// src/integrator.cu:7-56): shared volume, depth range (0.1, 5), weight caps 16 / 16.
#pragma once

#include <memory>
#include <vulcan/matrix.h>

namespace vulcan
{

struct Frame;
class Volume;

class Integrator
{
  public:

    Integrator(std::shared_ptr<Volume> volume);
    virtual ~Integrator() {}

    std::shared_ptr<Volume> GetVolume() const;

    const Vector2f& GetDepthRange() const;
    void SetDepthRange(const Vector2f& range);
    void SetDepthRange(float min, float max);

    float GetMaxDistanceWeight() const;
    void SetMaxDistanceWeight(float weight);

    float GetMaxColorWeight() const;
    void SetMaxColorWeight(float weight);

    virtual void Integrate(const Frame& frame) = 0;

  protected:

    std::shared_ptr<Volume> volume_;
    Vector2f depth_range_;
    float max_distance_weight_;
    float max_color_weight_;
};

} // namespace vulcan
