// ref_shim.cu — extern "C" handle API over the UNMODIFIED reference classes
// (vulcan::Volume, Depth/Color/LightIntegrator, Tracer, Frame), so that the tests and
// bench.py can drive the reference's own CUDA implementation through ctypes.
//
// TEST INFRASTRUCTURE ONLY (lives under oracle/).  This file contains no reference code;
// oracle/build_ref.sh compiles it together with the reference sources where they lie
// under /root/reference into oracle/_ref/libvulcan_ref*.so.
#include <vulcan/volume.h>
#include <vulcan/block.h>
#include <vulcan/hash.h>
#include <vulcan/voxel.h>
#include <vulcan/frame.h>
#include <vulcan/frame.cuh>
#include <vulcan/image.h>
#include <vulcan/depth_integrator.h>
#include <vulcan/color_integrator.h>
#include <vulcan/light_integrator.h>
#include <vulcan/light.h>
#include <vulcan/tracer.h>
#include <cstring>
#include <memory>

namespace {

using namespace vulcan;

// Transform's (matrix, inverse) constructor is protected (include/vulcan/transform.h:158-164)
struct RawTransform : public Transform {
  explicit RawTransform(const float* m32) {
    std::memcpy(&matrix_, m32, 64);
    std::memcpy(&inv_matrix_, m32 + 16, 64);
  }
};

struct OpenVolume : public Volume {
  OpenVolume(int m, int e) : Volume(m, e) {}
  const void* visibility() const { return block_visibility_.GetData(); }
};

struct Ref {
  std::shared_ptr<OpenVolume> volume;
  std::unique_ptr<DepthIntegrator> depth_integrator;
  std::unique_ptr<ColorIntegrator> color_integrator;
  std::unique_ptr<LightIntegrator> light_integrator;
  std::unique_ptr<Tracer> tracer;
  Frame frame;   // input frame
  Frame trace;   // traced frame
};

void set_pose(Frame& f, const float* dproj, const float* cproj, const float* Twd, const float* Tcd) {
  f.depth_projection.SetFocalLength(dproj[0], dproj[1]);
  f.depth_projection.SetCenterPoint(dproj[2], dproj[3]);
  if (cproj) {
    f.color_projection.SetFocalLength(cproj[0], cproj[1]);
    f.color_projection.SetCenterPoint(cproj[2], cproj[3]);
  }
  f.depth_to_world_transform = RawTransform(Twd);
  if (Tcd) f.depth_to_color_transform = RawTransform(Tcd);
}

}  // namespace

extern "C" {

void* ref_create(int main_blocks, int excess_blocks, float voxel_length, float trunc_length, float min_depth,
                 float max_depth) {
  try {
    Ref* r = new Ref();
    r->volume = std::make_shared<OpenVolume>(main_blocks, excess_blocks);
    r->volume->SetVoxelLength(voxel_length);
    r->volume->SetTruncationLength(trunc_length);
    r->volume->SetDepthRange(min_depth, max_depth);
    r->depth_integrator.reset(new DepthIntegrator(r->volume));
    r->color_integrator.reset(new ColorIntegrator(r->volume));
    r->light_integrator.reset(new LightIntegrator(r->volume));
    r->tracer.reset(new Tracer(r->volume));
    r->frame.depth_image = std::make_shared<Image>();
    r->frame.color_image = std::make_shared<ColorImage>();
    r->frame.normal_image = std::make_shared<ColorImage>();
    r->trace.depth_image = std::make_shared<Image>();
    r->trace.color_image = std::make_shared<ColorImage>();
    r->trace.normal_image = std::make_shared<ColorImage>();
    return r;
  } catch (...) {
    return nullptr;
  }
}

void ref_destroy(void* h) { delete (Ref*)h; }

void ref_set_ranges(void* h, float imin, float imax, float max_dw, float max_cw, float tmin, float tmax) {
  Ref* r = (Ref*)h;
  Integrator* ints[3] = {r->depth_integrator.get(), r->color_integrator.get(), r->light_integrator.get()};
  for (Integrator* i : ints) {
    i->SetDepthRange(imin, imax);
    i->SetMaxDistanceWeight(max_dw);
    i->SetMaxColorWeight(max_cw);
  }
  r->tracer->SetDepthRange(tmin, tmax);
}

void ref_set_light(void* h, float intensity, float x, float y, float z) {
  Light l;
  l.SetIntensity(intensity);
  l.SetPosition(x, y, z);
  ((Ref*)h)->light_integrator->SetLight(l);
}

// Host -> device upload of one frame (what Image::Load does after decoding,
// include/vulcan/image.h:107-109) plus pose / intrinsics.  color / normal may be NULL.
int ref_set_frame(void* h, const float* depth, int w, int hgt, const float* color, int cw, int ch,
                  const float* normal, const float* dproj, const float* cproj, const float* Twd,
                  const float* Tcd) {
  Ref* r = (Ref*)h;
  try {
    r->frame.depth_image->Resize(w, hgt);
    cudaMemcpy(r->frame.depth_image->GetData(), depth, sizeof(float) * w * hgt, cudaMemcpyHostToDevice);
    if (color) {
      r->frame.color_image->Resize(cw, ch);
      cudaMemcpy(r->frame.color_image->GetData(), color, sizeof(float) * 3 * cw * ch, cudaMemcpyHostToDevice);
    }
    if (normal) {
      r->frame.normal_image->Resize(w, hgt);
      cudaMemcpy(r->frame.normal_image->GetData(), normal, sizeof(float) * 3 * w * hgt, cudaMemcpyHostToDevice);
    }
    set_pose(r->frame, dproj, cproj, Twd, Tcd);
    return (int)cudaGetLastError();
  } catch (...) {
    return -1;
  }
}

// pose / intrinsics only (device-resident inputs: the depth image is copied device->device)
int ref_set_frame_device(void* h, const float* depth_dev, int w, int hgt, const float* dproj, const float* Twd) {
  Ref* r = (Ref*)h;
  r->frame.depth_image->Resize(w, hgt);
  if (depth_dev)
    cudaMemcpyAsync(r->frame.depth_image->GetData(), depth_dev, sizeof(float) * w * hgt, cudaMemcpyDeviceToDevice, 0);
  set_pose(r->frame, dproj, nullptr, Twd, nullptr);
  return (int)cudaGetLastError();
}

// vulcan::FilterDepths (src/frame.cu:194-203) on caller-provided device buffers.  (Frame::FilterDepths
// discards its result in a release build: the copy back is inside CUDA_DEBUG, src/frame.cpp:18.)
int ref_filter_depths(int w, int hgt, const float* src_dev, float* dst_dev) {
  vulcan::FilterDepths(w, hgt, src_dev, dst_dev);
  return (int)cudaDeviceSynchronize();
}

// vulcan::ComputeNormals (src/frame.cu:183-192) on caller-provided device buffers
int ref_compute_normals(const float* depth_dev, const float* dproj, float* normals_dev, int w, int hgt) {
  vulcan::Projection P;
  P.SetFocalLength(dproj[0], dproj[1]);
  P.SetCenterPoint(dproj[2], dproj[3]);
  vulcan::ComputeNormals(depth_dev, P, reinterpret_cast<vulcan::Vector3f*>(normals_dev), w, hgt);
  return (int)cudaDeviceSynchronize();
}

int ref_compute_input_normals(void* h) {
  try { ((Ref*)h)->frame.ComputeNormals(); return 0; } catch (...) { return -1; }
}

// one Volume::SetView; returns the visible count (host side, like Buffer::GetSize)
int ref_set_view(void* h) {
  Ref* r = (Ref*)h;
  try {
    r->volume->SetView(r->frame);
    return (int)r->volume->GetVisibleBlocks().GetSize();
  } catch (...) {
    return -1;
  }
}

int ref_integrate(void* h, int kind) {
  Ref* r = (Ref*)h;
  try {
    if (kind == 0) r->depth_integrator->Integrate(r->frame);
    else if (kind == 1) r->color_integrator->Integrate(r->frame);
    else r->light_integrator->Integrate(r->frame);
    return 0;
  } catch (...) {
    return -1;
  }
}

// Tracer::Trace into the internal trace frame from the given pose
int ref_trace(void* h, int w, int hgt, const float* dproj, const float* Twd) {
  Ref* r = (Ref*)h;
  try {
    r->trace.depth_image->Resize(w, hgt);
    set_pose(r->trace, dproj, nullptr, Twd, nullptr);
    r->tracer->Trace(r->trace);
    return 0;
  } catch (...) {
    return -1;
  }
}

int ref_sync() { return (int)cudaDeviceSynchronize(); }

// raw device pointers: 0 entries, 1 voxels, 2 visibility, 3 visible list, 4 traced depth,
// 5 traced colour, 6 traced normals, 7 input depth, 8 input normals
const void* ref_device_ptr(void* h, int which) {
  Ref* r = (Ref*)h;
  switch (which) {
    case 0: return r->volume->GetHashEntries().GetData();
    case 1: return r->volume->GetVoxels().GetData();
    case 2: return r->volume->visibility();
    case 3: return r->volume->GetVisibleBlocks().GetData();
    case 4: return r->trace.depth_image->GetData();
    case 5: return r->trace.color_image->GetData();
    case 6: return r->trace.normal_image ? r->trace.normal_image->GetData() : nullptr;
    case 7: return r->frame.depth_image->GetData();
    case 8: return r->frame.normal_image->GetData();
  }
  return nullptr;
}

int ref_visible_count(void* h) { return (int)((Ref*)h)->volume->GetVisibleBlocks().GetSize(); }

int ref_download(void* dst, const void* src_dev, size_t bytes) {
  return (int)cudaMemcpy(dst, src_dev, bytes, cudaMemcpyDeviceToHost);
}

}  // extern "C"
