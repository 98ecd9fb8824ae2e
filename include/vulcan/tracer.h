// vulcan/tracer.h - ray caster of the volume (reference: include/vulcan/tracer.h:13-76,
// src/tracer.cpp:13-147).  Trace = ray bounds from the visible blocks + per-pixel ray march +
// normals; the frame's depth image must be pre-sized by the caller, colour / normal images are
// created or resized (src/tracer.cpp:83-88, src/frame.cpp:25-28).
#pragma once

#include <memory>
#include <vulcan/buffer.h>
#include <vulcan/matrix.h>

namespace vulcan
{

struct Frame;
class Volume;

struct Patch
{
  static const int max_size = 16;
  Vector2s origin;
  Vector2s size;
  Vector2f bounds;
};

static_assert(sizeof(Patch) == 16, "Patch is part of the ABI");

class Tracer
{
  public:

    Tracer(std::shared_ptr<const Volume> volume);

    std::shared_ptr<const Volume> GetVolume() const;

    const Vector2f& GetDepthRange() const;
    void SetDepthRange(const Vector2f& range);
    void SetDepthRange(float min, float max);

    void Trace(Frame& frame);

    // ---- not in the reference: sharded volumes (one rank per GPU) -----------------------------
    // per-pixel merge keys (depth bits << 32 | rgb8 << 8 | rank) of the hits this rank owns;
    // ranks combine them with a 64-bit MIN all-reduce, then ResolveKeys fills the frame.
    void TraceKeys(const Frame& frame, unsigned long long* device_keys);
    void ResolveKeys(Frame& frame, const unsigned long long* device_keys);

  protected:

    Buffer<Vector2f> bounds_;
    std::shared_ptr<const Volume> volume_;
    Vector2f depth_range_;
};

} // namespace vulcan
