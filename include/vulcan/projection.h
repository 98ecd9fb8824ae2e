// vulcan/projection.h - pinhole intrinsics, 16 bytes (reference: include/vulcan/projection.h:11-114;
// byte image = vkProjection).  Pixel centres are at +0.5.
#pragma once

#include <vulcan/device.h>
#include <vulcan/math.h>
#include <vulcan/matrix.h>

namespace vulcan
{

class Projection
{
  public:

    VULCAN_HOST_DEVICE Projection() : focal_length_(500, 500), center_point_(320, 240) {}

    VULCAN_HOST_DEVICE const Vector2f& GetFocalLength() const { return focal_length_; }

    VULCAN_HOST_DEVICE void SetFocalLength(const Vector2f& length)
    {
      VULCAN_DEBUG(length[0] > 0 && length[1] > 0);
      focal_length_ = length;
    }

    VULCAN_HOST_DEVICE void SetFocalLength(float w, float h) { SetFocalLength(Vector2f(w, h)); }

    VULCAN_HOST_DEVICE const Vector2f& GetCenterPoint() const { return center_point_; }

    VULCAN_HOST_DEVICE void SetCenterPoint(const Vector2f& point)
    {
      VULCAN_DEBUG(!isnan(point[0]) && !isnan(point[1]));
      center_point_ = point;
    }

    VULCAN_HOST_DEVICE void SetCenterPoint(float w, float h) { SetCenterPoint(Vector2f(w, h)); }

    // (fx X/Z + cx, fy Y/Z + cy), evaluated as (1/Z) * f * X + c
    VULCAN_HOST_DEVICE Vector2f Project(const Vector3f& Xcp) const
    {
      const float inv_w = 1.0f / Xcp[2];
      return Vector2f(inv_w * focal_length_[0] * Xcp[0] + center_point_[0],
                      inv_w * focal_length_[1] * Xcp[1] + center_point_[1]);
    }

    VULCAN_HOST_DEVICE Vector2f Project(float x, float y, float z) const { return Project(Vector3f(x, y, z)); }

    // point at depth 1: ((1/fx) u - cx (1/fx), (1/fy) v - cy (1/fy), 1)
    VULCAN_HOST_DEVICE Vector3f Unproject(const Vector2f& uv) const
    {
      const float ifx = 1.0f / focal_length_[0], ify = 1.0f / focal_length_[1];
      return Vector3f(ifx * uv[0] - center_point_[0] * ifx, ify * uv[1] - center_point_[1] * ify, 1.0f);
    }

    VULCAN_HOST_DEVICE Vector3f Unproject(float u, float v) const { return Unproject(Vector2f(u, v)); }
    VULCAN_HOST_DEVICE Vector3f Unproject(const Vector2f& uv, float d) const { return d * Unproject(uv); }
    VULCAN_HOST_DEVICE Vector3f Unproject(float u, float v, float d) const { return Unproject(Vector2f(u, v), d); }

  protected:

    Vector2f focal_length_;
    Vector2f center_point_;
};

static_assert(sizeof(Projection) == 16, "Projection is passed to the C ABI as a byte image of vkProjection");

} // namespace vulcan
