// vulcan/voxel.h - one TSDF voxel, 20 bytes AoS (reference: include/vulcan/voxel.h:9-50;
// byte image = vkVoxel of include/vulcan_b200.h).  A block is 512 contiguous voxels, z*64 + y*8 + x.
#pragma once

#include <vulcan/device.h>
#include <vulcan/matrix.h>

namespace vulcan
{

class Voxel
{
  public:

    VULCAN_HOST_DEVICE Voxel() {}

    VULCAN_HOST_DEVICE const Vector3f& GetColor() const { return color; }
    VULCAN_HOST_DEVICE void SetColor(const Vector3f& c) { color = c; }

    // distance 1 (in truncation lengths), no colour, both weights 0 (tests/voxel_test.cpp:9-19)
    VULCAN_HOST_DEVICE static Voxel Empty()
    {
      Voxel v;
      v.distance = 1;
      v.color = Vector3f::Zeros();
      v.distance_weight = 0;
      v.color_weight = 0;
      return v;
    }

  public:

    float distance;
    Vector3f color;
    short distance_weight;
    short color_weight;
};

static_assert(sizeof(Voxel) == 20, "Voxel is part of the ABI");

} // namespace vulcan
