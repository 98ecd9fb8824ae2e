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
#include <vulcan/tracer.cuh>
#include <vulcan/depth_tracker.h>
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

struct OpenTracer : public Tracer {
  explicit OpenTracer(std::shared_ptr<const Volume> v) : Tracer(v) {}
  const void* bounds() const { return bounds_.GetData(); }
  const void* patches() const { return patches_.GetData(); }
  int patch_count() const { return (int)patches_.GetSize(); }
};

struct Ref {
  std::shared_ptr<OpenVolume> volume;
  std::unique_ptr<DepthIntegrator> depth_integrator;
  std::unique_ptr<ColorIntegrator> color_integrator;
  std::unique_ptr<LightIntegrator> light_integrator;
  std::unique_ptr<OpenTracer> tracer;
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
    r->tracer.reset(new OpenTracer(r->volume));
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

// Tracer::bounds_ (80 x 60 x Vector2f) and Tracer::patches_ as the last Trace left them
const void* ref_tracer_bounds(void* h) { return ((Ref*)h)->tracer->bounds(); }
const void* ref_tracer_patches(void* h) { return ((Ref*)h)->tracer->patches(); }
int ref_tracer_patch_count(void* h) { return ((Ref*)h)->tracer->patch_count(); }

// The step-wise free functions of include/vulcan/tracer.cuh:14-32 on caller-provided device buffers
// (what tests/tracer_test.cu:135,234-235,326-337 call).  Tcw32 / Twc32: matrix + inverse, 32 floats.
int ref_compute_patches(const int* indices, const void* entries, const float* Tcw32, const float* proj4,
                        float block_length, float min_depth, float max_depth, int block_count, int image_w,
                        int image_h, int bounds_w, int bounds_h, void* patches, int* patch_count) {
  try {
    Projection P;
    P.SetFocalLength(proj4[0], proj4[1]);
    P.SetCenterPoint(proj4[2], proj4[3]);
    vulcan::ComputePatches(indices, (const HashEntry*)entries, RawTransform(Tcw32), P, block_length, min_depth,
                           max_depth, block_count, image_w, image_h, bounds_w, bounds_h, (Patch*)patches, patch_count);
    return (int)cudaDeviceSynchronize();
  } catch (...) { return -1; }
}
int ref_compute_bounds(const void* patches, float* bounds, int bounds_w, int patch_count) {
  try {
    vulcan::ComputeBounds((const Patch*)patches, (Vector2f*)bounds, bounds_w, patch_count);
    return (int)cudaDeviceSynchronize();
  } catch (...) { return -1; }
}
int ref_reset_bounds(float* bounds, int count) {
  try {
    vulcan::ResetBoundsBuffer((Vector2f*)bounds, count);
    return (int)cudaDeviceSynchronize();
  } catch (...) { return -1; }
}
int ref_compute_points(const void* entries, const void* voxels, const float* bounds, int block_count,
                       float block_length, float voxel_length, float trunc_length, const float* Twc32,
                       const float* proj4, float* depths, float* colors, int image_w, int image_h, int bounds_w,
                       int bounds_h) {
  try {
    Projection P;
    P.SetFocalLength(proj4[0], proj4[1]);
    P.SetCenterPoint(proj4[2], proj4[3]);
    vulcan::ComputePoints((const HashEntry*)entries, (const Voxel*)voxels, (const Vector2f*)bounds, block_count,
                          block_length, voxel_length, trunc_length, RawTransform(Twc32), P, depths, (Vector3f*)colors,
                          image_w, image_h, bounds_w, bounds_h);
    return (int)cudaDeviceSynchronize();
  } catch (...) { return -1; }
}

int ref_visible_count(void* h) { return (int)((Ref*)h)->volume->GetVisibleBlocks().GetSize(); }

int ref_download(void* dst, const void* src_dev, size_t bytes) {
  return (int)cudaMemcpy(dst, src_dev, bytes, cudaMemcpyDeviceToHost);
}


// ---- image pyramid + DepthTracker (src/frame.cpp:38-57, src/tracker.cpp, src/depth_tracker.{cpp,cu}) ----

namespace {
struct OpenDepthTracker : public DepthTracker {
  void system(const Frame& f) {
    BeginSolve(f);  // ResizeBuffers (src/tracker.cpp:75-81)
    ComputeSystem(f);
  }
  const float* hessian() const { return hessian_.GetData(); }
  const float* gradient() const { return gradient_.GetData(); }
  int iteration() const { return iteration_; }
};
struct RefTrack {
  OpenDepthTracker tracker;
  std::shared_ptr<Frame> frames[2];  // 0 keyframe, 1 frame
  Frame half, quarter;               // last ref_track_downsample results
};
void upload_frame(Frame& f, const float* depth, const float* normal, const float* color, int w, int h,
                  const float* dproj, const float* Twd) {
  if (!f.depth_image) f.depth_image = std::make_shared<Image>();
  if (!f.normal_image) f.normal_image = std::make_shared<ColorImage>();
  if (!f.color_image) f.color_image = std::make_shared<ColorImage>();
  f.depth_image->Resize(w, h);
  f.normal_image->Resize(w, h);
  f.color_image->Resize(w, h);
  cudaMemcpy(f.depth_image->GetData(), depth, sizeof(float) * w * h, cudaMemcpyHostToDevice);
  cudaMemcpy(f.normal_image->GetData(), normal, sizeof(float) * 3 * w * h, cudaMemcpyHostToDevice);
  if (color) cudaMemcpy(f.color_image->GetData(), color, sizeof(float) * 3 * w * h, cudaMemcpyHostToDevice);
  else cudaMemset(f.color_image->GetData(), 0, sizeof(float) * 3 * w * h);
  set_pose(f, dproj, dproj, Twd, nullptr);
}
void get_pose(const Frame& f, float* out32) {
  std::memcpy(out32, &f.depth_to_world_transform.GetMatrix(), 64);
  std::memcpy(out32 + 16, &f.depth_to_world_transform.GetInverseMatrix(), 64);
}
}  // namespace

void* ref_track_create() {
  try {
    RefTrack* t = new RefTrack();
    t->frames[0] = std::make_shared<Frame>();
    t->frames[1] = std::make_shared<Frame>();
    return t;
  } catch (...) { return nullptr; }
}
void ref_track_destroy(void* h) { delete (RefTrack*)h; }

// which: 0 keyframe, 1 frame; host images of one size; colour may be NULL (zeros)
int ref_track_set_frame(void* h, int which, const float* depth, const float* normal, const float* color, int w,
                        int hgt, const float* dproj, const float* Twd) {
  RefTrack* t = (RefTrack*)h;
  try {
    upload_frame(*t->frames[which], depth, normal, color, w, hgt, dproj, Twd);
    t->tracker.SetKeyframe(t->frames[0]);
    return (int)cudaGetLastError();
  } catch (...) { return -1; }
}

int ref_track_residuals(void* h, float* out_host) {
  RefTrack* t = (RefTrack*)h;
  try {
    Buffer<float> r;
    t->tracker.ComputeResiduals(*t->frames[1], r);
    return (int)cudaMemcpy(out_host, r.GetData(), sizeof(float) * r.GetSize(), cudaMemcpyDeviceToHost);
  } catch (...) { return -1; }
}

int ref_track_jacobian(void* h, int translation, float* out_host) {
  RefTrack* t = (RefTrack*)h;
  try {
    Buffer<Vector6f> j;
    t->tracker.SetTranslationEnabled(translation != 0);
    t->tracker.ComputeJacobian(*t->frames[1], j);
    return (int)cudaMemcpy(out_host, j.GetData(), sizeof(float) * 6 * j.GetSize(), cudaMemcpyDeviceToHost);
  } catch (...) { return -1; }
}

// packed lower triangle (21 or 6) and gradient (6 or 3) as the reference's kernel leaves them
int ref_track_system(void* h, int translation, float* hessian_host, float* gradient_host) {
  RefTrack* t = (RefTrack*)h;
  try {
    t->tracker.SetTranslationEnabled(translation != 0);
    t->tracker.system(*t->frames[1]);
    cudaMemcpy(hessian_host, t->tracker.hessian(), sizeof(float) * (translation ? 21 : 6), cudaMemcpyDeviceToHost);
    return (int)cudaMemcpy(gradient_host, t->tracker.gradient(), sizeof(float) * (translation ? 6 : 3),
                           cudaMemcpyDeviceToHost);
  } catch (...) { return -1; }
}

// Tracker::Track on the frame; pose_out = matrix + inverse (32 floats)
int ref_track(void* h, int translation, int max_iterations, float* pose_out) {
  RefTrack* t = (RefTrack*)h;
  try {
    t->tracker.SetTranslationEnabled(translation != 0);
    t->tracker.SetMaxIterations(max_iterations);
    t->tracker.SetKeyframe(t->frames[0]);
    t->tracker.Track(*t->frames[1]);
    cudaDeviceSynchronize();
    get_pose(*t->frames[1], pose_out);
    return (int)cudaGetLastError();
  } catch (...) { return -1; }
}

// Frame::Downsample of frame `which` (twice: half and quarter); level 1 = half, 2 = quarter copied to host.
// proj_out = depth fx fy cx cy
int ref_track_downsample(void* h, int which, int level, float* depth_out, float* color_out, float* normal_out,
                         float* proj_out) {
  RefTrack* t = (RefTrack*)h;
  try {
    t->frames[which]->Downsample(t->half);
    t->half.Downsample(t->quarter);
    const Frame& f = level == 1 ? t->half : t->quarter;
    const size_t n = (size_t)f.depth_image->GetWidth() * f.depth_image->GetHeight();
    cudaMemcpy(depth_out, f.depth_image->GetData(), 4 * n, cudaMemcpyDeviceToHost);
    cudaMemcpy(color_out, f.color_image->GetData(), 12 * n, cudaMemcpyDeviceToHost);
    cudaMemcpy(normal_out, f.normal_image->GetData(), 12 * n, cudaMemcpyDeviceToHost);
    proj_out[0] = f.depth_projection.GetFocalLength()[0];
    proj_out[1] = f.depth_projection.GetFocalLength()[1];
    proj_out[2] = f.depth_projection.GetCenterPoint()[0];
    proj_out[3] = f.depth_projection.GetCenterPoint()[1];
    return (int)cudaGetLastError();
  } catch (...) { return -1; }
}

// The live part of PyramidTracker<DepthTracker>::Track (src/pyramid_tracker.cpp:52-90), whose own source
// file cannot be compiled here (it instantiates the colour and light trackers): both pyramids, 15
// iterations at half size, 20 at full size.
int ref_track_pyramid(void* h, float* pose_out) {
  RefTrack* t = (RefTrack*)h;
  try {
    Frame half_frame, quarter_frame;
    auto half_key = std::make_shared<Frame>();
    auto quarter_key = std::make_shared<Frame>();
    t->frames[1]->Downsample(half_frame);
    half_frame.Downsample(quarter_frame);
    t->frames[0]->Downsample(*half_key);
    half_key->Downsample(*quarter_key);
    t->tracker.SetMaxIterations(15);
    t->tracker.SetTranslationEnabled(true);
    half_frame.depth_to_world_transform = quarter_frame.depth_to_world_transform;
    t->tracker.SetKeyframe(half_key);
    t->tracker.Track(half_frame);
    t->tracker.SetMaxIterations(20);
    t->frames[1]->depth_to_world_transform = half_frame.depth_to_world_transform;
    t->tracker.SetKeyframe(t->frames[0]);
    t->tracker.Track(*t->frames[1]);
    cudaDeviceSynchronize();
    get_pose(*t->frames[1], pose_out);
    return (int)cudaGetLastError();
  } catch (...) { return -1; }
}

}  // extern "C"
