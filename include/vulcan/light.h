// vulcan/light.h - point light of the LightIntegrator, 16 bytes (reference: include/vulcan/light.h:10-67;
// byte image = vkLight).
#pragma once

#include <vulcan/device.h>
#include <vulcan/matrix.h>

namespace vulcan
{

class Light
{
  public:

    VULCAN_HOST_DEVICE Light() : intensity_(1.0f), position_(0, 0, 0) {}

    VULCAN_HOST_DEVICE float GetIntensity() const { return intensity_; }
    VULCAN_HOST_DEVICE void SetIntensity(float intensity) { VULCAN_ASSERT(intensity >= 0); intensity_ = intensity; }
    VULCAN_HOST_DEVICE const Vector3f& GetPosition() const { return position_; }
    VULCAN_HOST_DEVICE void SetPosition(const Vector3f& position) { position_ = position; }
    VULCAN_HOST_DEVICE void SetPosition(float x, float y, float z) { SetPosition(Vector3f(x, y, z)); }

    // intensity * cos(theta) / distance^2 for a surface point and normal in the light's frame
    VULCAN_HOST_DEVICE float GetShading(const Vector3f& point, const Vector3f& normal) const
    {
      const Vector3f delta = position_ - point;
      const Vector3f direction = delta.Normalized();
      return intensity_ * normal.Dot(direction) / delta.SquaredNorm();
    }

  protected:

    float intensity_;
    Vector3f position_;
};

static_assert(sizeof(Light) == 16, "Light is passed to the C ABI as a byte image of vkLight");

} // namespace vulcan
