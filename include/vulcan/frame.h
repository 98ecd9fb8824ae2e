// vulcan/frame.h - one RGB-D frame: two projections, two transforms, three shared device images
// (reference: include/vulcan/frame.h:11-32, src/frame.cpp:8-36).
#pragma once

#include <memory>
#include <vulcan/image.h>
#include <vulcan/projection.h>
#include <vulcan/transform.h>

namespace vulcan
{

struct Frame
{
  Projection depth_projection;
  Projection color_projection;
  Transform depth_to_world_transform;
  Transform depth_to_color_transform;
  std::shared_ptr<Image> depth_image;
  std::shared_ptr<ColorImage> color_image;
  std::shared_ptr<ColorImage> normal_image;

  // 7x7 bilateral filter of the depth image, in place (src/frame.cu:124-181)
  void FilterDepths();

  // normals from depth; creates / resizes normal_image (src/frame.cu:7-122, src/frame.cpp:21-36)
  void ComputeNormals();

  // addition: FilterDepths followed by ComputeNormals in one kernel; replaces depth_image by the
  // filtered image (a new Image object: other holders of the old pointer keep the raw depth)
  void Prepare();

  // half-size pyramid level into `frame`: depth and normals by nearest sample, colour by 2x2 average,
  // intrinsics halved, transforms copied; one kernel for the three images (src/frame.cpp:38-57)
  void Downsample(Frame& frame) const;
};

} // namespace vulcan
