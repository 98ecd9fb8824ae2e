// vulcan/mesh.h (reference: include/vulcan/mesh.h:8-21).  `colors` is an addition: the fused voxel
// colour interpolated at each vertex (empty when the extractor did not fill it).
#pragma once

#include <vector>
#include <vulcan/buffer.h>
#include <vulcan/matrix.h>

namespace vulcan
{

struct Mesh
{
  std::vector<Vector3f> points;
  std::vector<Vector3i> faces;
  std::vector<Vector3f> colors;
};

struct DeviceMesh
{
  Buffer<Vector3f> points;
  Buffer<Vector3i> faces;
  Buffer<Vector3f> colors;
};

} // namespace vulcan
