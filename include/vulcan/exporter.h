// vulcan/exporter.h - ASCII PLY writer (reference: include/vulcan/exporter.h, src/exporter.cpp:19-71).
// Same header as the reference's file; vertex colours are mesh.colors (grey when absent) instead of the
// reference's debug z ramp (marked REMOVE in src/exporter.cpp:30-32).
#pragma once

#include <string>

namespace vulcan
{

struct Mesh;

class Exporter
{
  public:

    Exporter(const std::string& file);
    const std::string& GetFile() const { return file_; }
    void Export(const Mesh& mesh) const;

  protected:

    std::string file_;
};

} // namespace vulcan
