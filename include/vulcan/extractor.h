// vulcan/extractor.h - surface extraction (reference: include/vulcan/extractor.h:116-137,
// src/extractor.cu:660-740).  The reference's extractor is unfinished (edge midpoints of one block, empty
// index / face kernels); this one returns the interpolated marching-cubes vertices of ALL allocated
// blocks with their colours (vk_extract_points) and, like the reference today, no faces.
#pragma once

#include <memory>
#include <vulcan/mesh.h>

namespace vulcan
{

class Volume;

class Extractor
{
  public:

    Extractor(std::shared_ptr<const Volume> volume);
    std::shared_ptr<const Volume> GetVolume() const { return volume_; }
    void Extract(DeviceMesh& mesh) const;
    void Extract(Mesh& mesh) const;

  protected:

    std::shared_ptr<const Volume> volume_;
};

} // namespace vulcan
