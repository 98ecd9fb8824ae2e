// libvulcan.so - the reference's public C++ classes (namespace vulcan) as thin host wrappers over
// the C ABI of the B200 kernels (include/vulcan_b200.h).  No kernel code here: plain C++ + the
// CUDA runtime for memory.  Status codes of the C ABI become vulcan::Exception like CUDA_ASSERT
// does in the reference (include/vulcan/device.h:14-17).
#include <vulcan/vulcan.h>

#include <cstdio>
#include <cstring>

namespace vulcan
{

namespace
{

// byte images: the value types are laid out like their C-ABI mirrors (static_asserts in the headers)
vkTransform ToC(const Transform& t)
{
  vkTransform out;
  std::memcpy(&out, &t, sizeof(out));
  return out;
}

vkProjection ToC(const Projection& p)
{
  vkProjection out;
  std::memcpy(&out, &p, sizeof(out));
  return out;
}

vkFrame ToC(const Frame& frame)
{
  VULCAN_ASSERT_MSG(frame.depth_image != nullptr, "missing depth image");
  vkFrame f;
  std::memset(&f, 0, sizeof(f));
  f.depth = frame.depth_image->GetData();
  f.depth_w = frame.depth_image->GetWidth();
  f.depth_h = frame.depth_image->GetHeight();
  if (frame.color_image)
  {
    f.color = reinterpret_cast<const float*>(frame.color_image->GetData());
    f.color_w = frame.color_image->GetWidth();
    f.color_h = frame.color_image->GetHeight();
  }
  if (frame.normal_image) f.normal = reinterpret_cast<const float*>(frame.normal_image->GetData());
  f.depth_proj = ToC(frame.depth_projection);
  f.color_proj = ToC(frame.color_projection);
  f.depth_to_world = ToC(frame.depth_to_world_transform);
  f.depth_to_color = ToC(frame.depth_to_color_transform);
  return f;
}

void EnsureColorImage(std::shared_ptr<ColorImage>& image, int w, int h)
{
  if (!image) image = std::make_shared<ColorImage>();
  image->Resize(w, h);
}

} // namespace

// ---- Frame (src/frame.cpp:8-58) ----------------------------------------------------------------

void Frame::FilterDepths()
{
  VULCAN_ASSERT_MSG(depth_image != nullptr, "missing depth image");
  const int w = depth_image->GetWidth(), h = depth_image->GetHeight();
  Image temp(w, h);
  VK_ASSERT(vk_filter_depths(w, h, depth_image->GetData(), temp.GetData(), nullptr));
  CUDA_ASSERT(cudaMemcpyAsync(depth_image->GetData(), temp.GetData(), temp.GetBytes(), cudaMemcpyDeviceToDevice, 0));
  CUDA_ASSERT(cudaStreamSynchronize(0));  // temp dies here
}

// FilterDepths + ComputeNormals in one kernel (vk_prepare_frame); not in the reference.
void Frame::Prepare()
{
  VULCAN_ASSERT_MSG(depth_image != nullptr, "missing depth image");
  const int w = depth_image->GetWidth(), h = depth_image->GetHeight();
  auto filtered = std::make_shared<Image>(w, h);
  EnsureColorImage(normal_image, w, h);
  const vkProjection proj = ToC(depth_projection);
  VK_ASSERT(vk_prepare_frame(w, h, depth_image->GetData(), &proj, filtered->GetData(),
                             reinterpret_cast<float*>(normal_image->GetData()), nullptr));
  CUDA_ASSERT(cudaStreamSynchronize(0));  // the unfiltered image may die here
  depth_image = filtered;
}

void Frame::ComputeNormals()
{
  VULCAN_ASSERT_MSG(depth_image != nullptr, "missing depth image");
  const int w = depth_image->GetWidth(), h = depth_image->GetHeight();
  EnsureColorImage(normal_image, w, h);
  const vkProjection proj = ToC(depth_projection);
  VK_ASSERT(vk_compute_normals(depth_image->GetData(), &proj, reinterpret_cast<float*>(normal_image->GetData()), w, h,
                               nullptr));
}

void Frame::Downsample(Frame& frame) const
{
  VULCAN_ASSERT_MSG(depth_image && color_image && normal_image, "depth, color and normal images required");
  if (!frame.depth_image) frame.depth_image = std::make_shared<Image>();
  if (!frame.color_image) frame.color_image = std::make_shared<ColorImage>();
  if (!frame.normal_image) frame.normal_image = std::make_shared<ColorImage>();
  frame.depth_image->Resize(depth_image->GetWidth() / 2, depth_image->GetHeight() / 2);
  frame.normal_image->Resize(depth_image->GetWidth() / 2, depth_image->GetHeight() / 2);
  frame.color_image->Resize(color_image->GetWidth() / 2, color_image->GetHeight() / 2);
  const vkFrame src = ToC(*this);
  vkFrame dst = ToC(frame);
  const int status = vk_downsample_frame(&src, &dst, nullptr);
  VULCAN_ASSERT_MSG(status != VK_E_INVALID, "even image dimensions required");  // src/image.cu:185-186
  VK_ASSERT(status);
  std::memcpy(static_cast<void*>(&frame.depth_projection), &dst.depth_proj, sizeof(dst.depth_proj));
  std::memcpy(static_cast<void*>(&frame.color_projection), &dst.color_proj, sizeof(dst.color_proj));
  frame.depth_to_world_transform = depth_to_world_transform;
  frame.depth_to_color_transform = depth_to_color_transform;
}

void Image::Downsample(Image& image, bool nearest) const
{
  VULCAN_ASSERT_MSG(GetWidth() % 2 == 0 && GetHeight() % 2 == 0, "even image dimensions required");
  image.Resize(GetWidth() / 2, GetHeight() / 2);
  VK_ASSERT(vk_downsample_image(GetData(), GetWidth(), GetHeight(), 1, nearest, image.GetData(), nullptr));
}

void ColorImage::Downsample(ColorImage& image, bool nearest) const
{
  VULCAN_ASSERT_MSG(GetWidth() % 2 == 0 && GetHeight() % 2 == 0, "even image dimensions required");
  image.Resize(GetWidth() / 2, GetHeight() / 2);
  VK_ASSERT(vk_downsample_image(reinterpret_cast<const float*>(GetData()), GetWidth(), GetHeight(), 3, nearest,
                                reinterpret_cast<float*>(image.GetData()), nullptr));
}

// ---- Tracker / DepthTracker (src/tracker.cpp, src/depth_tracker.cpp, src/depth_tracker.cu) --------

Tracker::Tracker() : translation_enabled_(true), max_iterations_(20), iterations_(0), last_update_norm_(0)
{
  workspace_.Resize(vk_track_workspace_floats());
}

Tracker::~Tracker() {}

void Tracker::SetMaxIterations(int iterations)
{
  VULCAN_ASSERT(iterations > 0);
  max_iterations_ = iterations;
}

void Tracker::ValidateKeyframe() const
{
  VULCAN_ASSERT_MSG(keyframe_ != nullptr, "keyframe has not been assigned");
  VULCAN_ASSERT_MSG(keyframe_->depth_image != nullptr, "keyframe missing depth image");
  VULCAN_ASSERT_MSG(keyframe_->normal_image != nullptr, "keyframe missing normal image");
  const int dw = keyframe_->depth_image->GetWidth(), dh = keyframe_->depth_image->GetHeight();
  VULCAN_ASSERT_MSG(dw > 0 && dh > 0, "invalid keyframe depth image size");
  VULCAN_ASSERT_MSG(dw == keyframe_->normal_image->GetWidth() && dh == keyframe_->normal_image->GetHeight(),
                    "keyframe image size mismatch");
}

void Tracker::ValidateFrame(const Frame& frame) const
{
  VULCAN_ASSERT_MSG(frame.depth_image != nullptr, "frame missing depth image");
  VULCAN_ASSERT_MSG(frame.normal_image != nullptr, "frame missing normal image");
  const int dw = frame.depth_image->GetWidth(), dh = frame.depth_image->GetHeight();
  VULCAN_ASSERT_MSG(dw > 0 && dh > 0, "invalid frame depth image size");
  VULCAN_ASSERT_MSG(dw == frame.normal_image->GetWidth() && dh == frame.normal_image->GetHeight(),
                    "frame image size mismatch");
}

void Tracker::Track(Frame& frame)
{
  ValidateKeyframe();
  ValidateFrame(frame);
  Solve(frame);
}

DepthTracker::DepthTracker() {}
DepthTracker::~DepthTracker() {}

void DepthTracker::Solve(Frame& frame)
{
  const vkFrame key = ToC(*keyframe_);
  vkFrame f = ToC(frame);
  VK_ASSERT(vk_track_depth(&key, &f, translation_enabled_, max_iterations_, workspace_.GetData(), &iterations_,
                           &last_update_norm_, nullptr));
  std::memcpy(static_cast<void*>(&frame.depth_to_world_transform), &f.depth_to_world, sizeof(f.depth_to_world));
}

void DepthTracker::ComputeResiduals(const Frame& frame, Buffer<float>& residuals) const
{
  ValidateKeyframe();
  ValidateFrame(frame);
  residuals.Resize((size_t)frame.depth_image->GetWidth() * frame.depth_image->GetHeight());
  const vkFrame key = ToC(*keyframe_), f = ToC(frame);
  VK_ASSERT(vk_track_residuals(&key, &f, residuals.GetData(), nullptr));
}

void DepthTracker::ComputeJacobian(const Frame& frame, Buffer<Vector6f>& jacobian) const
{
  ValidateKeyframe();
  ValidateFrame(frame);
  jacobian.Resize((size_t)frame.depth_image->GetWidth() * frame.depth_image->GetHeight());
  const vkFrame key = ToC(*keyframe_), f = ToC(frame);
  VK_ASSERT(vk_track_jacobian(&key, &f, translation_enabled_, reinterpret_cast<float*>(jacobian.GetData()), nullptr));
}

void DepthTracker::ComputeSystem(const Frame& frame, Buffer<float>& hessian, Buffer<float>& gradient)
{
  ValidateKeyframe();
  ValidateFrame(frame);
  const int n = GetParameterCount();
  hessian.Resize(n * (n + 1) / 2);
  gradient.Resize(n);
  const vkFrame key = ToC(*keyframe_), f = ToC(frame);
  VK_ASSERT(vk_track_system(&key, &f, translation_enabled_, workspace_.GetData(), hessian.GetData(),
                            gradient.GetData(), nullptr));
}

// ---- Extractor / Exporter (src/extractor.cu:660-740, src/exporter.cpp:19-71) ------------------

Extractor::Extractor(std::shared_ptr<const Volume> volume) : volume_(volume)
{
  VULCAN_ASSERT(volume_ != nullptr);
}

void Extractor::Extract(DeviceMesh& mesh) const
{
  // vertices (zero crossings on the lattice edges) and marching-cubes faces over them; the kernels count on
  // past the capacities, so a second round with room for all follows when the first guess was too small
  Buffer<int> count(2);
  size_t capacity = mesh.points.GetSize() > 0 ? mesh.points.GetSize() : (size_t)1 << 20;
  size_t face_capacity = mesh.faces.GetSize() > 0 ? mesh.faces.GetSize() : 2 * capacity;
  for (;;)
  {
    mesh.points.Resize(capacity);
    mesh.colors.Resize(capacity);
    mesh.faces.Resize(face_capacity);
    VK_ASSERT(vk_extract_mesh(volume_->GetHandle(), reinterpret_cast<float*>(mesh.points.GetData()),
                              reinterpret_cast<float*>(mesh.colors.GetData()), (int)capacity, count.GetData(),
                              reinterpret_cast<int*>(mesh.faces.GetData()), (int)face_capacity, count.GetData() + 1,
                              volume_->GetStream()));
    int found[2] = {0, 0};
    count.CopyToHost(found);
    if ((size_t)found[0] <= capacity && (size_t)found[1] <= face_capacity)
    {
      mesh.points.Resize(found[0]);
      mesh.colors.Resize(found[0]);
      mesh.faces.Resize(found[1]);
      break;
    }
    if ((size_t)found[0] > capacity) capacity = found[0];
    if ((size_t)found[1] > face_capacity) face_capacity = found[1];
  }
}

void Extractor::Extract(Mesh& mesh) const
{
  DeviceMesh device;
  Extract(device);
  mesh.points.resize(device.points.GetSize());
  mesh.colors.resize(device.colors.GetSize());
  mesh.faces.resize(device.faces.GetSize());
  if (!mesh.points.empty())
  {
    device.points.CopyToHost(mesh.points.data());
    device.colors.CopyToHost(mesh.colors.data());
  }
  if (!mesh.faces.empty()) device.faces.CopyToHost(mesh.faces.data());
}

Exporter::Exporter(const std::string& file) : file_(file) {}

void Exporter::Export(const Mesh& mesh) const
{
  FILE* f = fopen(file_.c_str(), "w");
  VULCAN_ASSERT_MSG(f != nullptr, "cannot open output file");
  fprintf(f, "ply\nformat ascii 1.0\nelement vertex %zu\n", mesh.points.size());
  fprintf(f, "property float x\nproperty float y\nproperty float z\n");
  fprintf(f, "property uchar red\nproperty uchar green\nproperty uchar blue\n");
  fprintf(f, "element face %zu\nproperty list uchar int vertex_indices\nend_header\n", mesh.faces.size());
  const bool colored = mesh.colors.size() == mesh.points.size();
  for (size_t i = 0; i < mesh.points.size(); ++i)
  {
    int rgb[3] = {128, 128, 128};
    if (colored)
      for (int c = 0; c < 3; ++c)
      {
        const float v = 255.0f * mesh.colors[i][c] + 0.5f;
        rgb[c] = v < 0 ? 0 : v > 255 ? 255 : (int)v;
      }
    fprintf(f, "%.9g %.9g %.9g %d %d %d\n", mesh.points[i][0], mesh.points[i][1], mesh.points[i][2], rgb[0], rgb[1],
            rgb[2]);
  }
  for (const Vector3i& t : mesh.faces) fprintf(f, "3 %d %d %d\n", t[0], t[1], t[2]);
  fclose(f);
}

// ---- Volume (src/volume.cu:370-463) ----------------------------------------------------------

Volume::Volume(int main_block_count, int excess_block_count) : stream_(nullptr), depth_range_(0.1f, 5.0f)
{
  VULCAN_ASSERT(main_block_count > 0 && excess_block_count >= 0);
  VK_ASSERT(vk_volume_create(&volume_, main_block_count, excess_block_count, stream_));
  depth_range_ = Vector2f(volume_.min_depth, volume_.max_depth);
  RefreshViews();
}

Volume::~Volume()
{
  vk_volume_destroy(&volume_);
}

void Volume::RefreshViews() const
{
  const size_t B = (size_t)volume_.main_blocks + (size_t)volume_.excess_blocks;
  voxels_.Adopt(reinterpret_cast<Voxel*>(volume_.voxels), B * Block::voxel_count, B * Block::voxel_count);
  hash_entries_.Adopt(reinterpret_cast<HashEntry*>(volume_.entries), B, B);
  block_visibility_.Adopt(reinterpret_cast<Visibility*>(volume_.visibility), B, B);
  visible_blocks_.Adopt(volume_.visible[volume_.visible_slot], B, visible_blocks_.GetSize() <= B ? visible_blocks_.GetSize() : 0);
}

int Volume::GetMainBlockCount() const { return volume_.main_blocks; }
int Volume::GetExcessBlockCount() const { return volume_.excess_blocks; }

const Vector2f& Volume::GetDepthRange() const { return depth_range_; }

void Volume::SetDepthRange(const Vector2f& range)
{
  VULCAN_DEBUG(range[0] > 0 && range[0] < range[1]);
  depth_range_ = range;
  volume_.min_depth = range[0];
  volume_.max_depth = range[1];
}

void Volume::SetDepthRange(float min, float max) { SetDepthRange(Vector2f(min, max)); }

float Volume::GetVoxelLength() const { return volume_.voxel_length; }

void Volume::SetVoxelLength(float length)
{
  VULCAN_DEBUG(length > 0);
  volume_.voxel_length = length;
  // the exact fast division by the voxel and block length is proven here (first use of a value
  // synchronises the stream), not inside the first frame
  vk_div_prepare(length, stream_);
  vk_div_prepare(Block::resolution * length, stream_);
}

float Volume::GetTruncationLength() const { return volume_.trunc_length; }

void Volume::SetTruncationLength(float length)
{
  VULCAN_DEBUG(length > 0);
  volume_.trunc_length = length;
  vk_div_prepare(length, stream_);
}

void Volume::SetView(const Frame& frame)
{
  const vkFrame f = ToC(frame);
  VK_ASSERT(vk_set_view(&volume_, &f, stream_));
}

const Buffer<HashEntry>& Volume::GetHashEntries() const { return hash_entries_; }

const Buffer<int>& Volume::GetAllocatedBlocks() const
{
  VULCAN_THROW("not implemented");
}

int Volume::GetVisibleBlockCount() const
{
  int count = 0;
  const int status = vk_volume_sync_counters(&volume_, &count, stream_);
  VULCAN_ASSERT_MSG(status == 0 || status == VK_E_OVERFLOW, GetStatusString(status));
  return count;
}

int Volume::GetAllocatedBlockCount() const
{
  GetVisibleBlockCount();
  return volume_.host_counters[VK_CNT_ALLOCATED];
}

const Buffer<int>& Volume::GetVisibleBlocks() const
{
  const int count = GetVisibleBlockCount();
  const size_t B = (size_t)volume_.main_blocks + (size_t)volume_.excess_blocks;
  visible_blocks_.Adopt(volume_.visible[volume_.visible_slot], B, (size_t)count);
  return visible_blocks_;
}

const Buffer<Voxel>& Volume::GetVoxels() const { return voxels_; }
Buffer<Voxel>& Volume::GetVoxels()
{
  sdf_stale_ = true;  // the caller may write the pool: the distance mirror is rebuilt before the next trace
  return voxels_;
}

void Volume::SyncDistances() const
{
  if (!sdf_stale_) return;
  VK_ASSERT(vk_volume_sync_sdf(&volume_, stream_));
  sdf_stale_ = false;
}
const Buffer<Visibility>& Volume::GetBlockVisibility() const { return block_visibility_; }

void Volume::Reset()
{
  VK_ASSERT(vk_volume_reset(&volume_, stream_));
  visible_blocks_.SetSize(0);
  RefreshViews();
}

void Volume::Save(const std::string& path) const
{
  VK_ASSERT(vk_volume_save(&volume_, path.c_str(), stream_));
}

void Volume::Load(const std::string& path)
{
  VK_ASSERT(vk_volume_load(&volume_, path.c_str(), stream_));
  depth_range_ = Vector2f(volume_.min_depth, volume_.max_depth);
  visible_blocks_.SetSize(0);  // refreshed from the device counters on the next GetVisibleBlocks()
  RefreshViews();
}

void Volume::SetSharding(int rank, int count, int brick_shift, int halo)
{
  VULCAN_ASSERT(count >= 1 && rank >= 0 && rank < count && brick_shift >= 0 && brick_shift < 15);
  VULCAN_ASSERT(halo == 0 || halo == 1);
  volume_.shard_rank = rank;
  volume_.shard_count = count;
  volume_.shard_shift = brick_shift;
  volume_.shard_halo = halo;
}

// ---- Integrators (src/integrator.cu, depth_/color_/light_integrator.cu) -----------------------------

Integrator::Integrator(std::shared_ptr<Volume> volume) :
  volume_(volume), depth_range_(0.1f, 5.0f), max_distance_weight_(16), max_color_weight_(16)
{
}

std::shared_ptr<Volume> Integrator::GetVolume() const { return volume_; }
const Vector2f& Integrator::GetDepthRange() const { return depth_range_; }

void Integrator::SetDepthRange(const Vector2f& range)
{
  VULCAN_DEBUG(range[0] > 0 && range[0] < range[1]);
  depth_range_ = range;
}

void Integrator::SetDepthRange(float min, float max) { SetDepthRange(Vector2f(min, max)); }
float Integrator::GetMaxDistanceWeight() const { return max_distance_weight_; }

void Integrator::SetMaxDistanceWeight(float weight)
{
  VULCAN_DEBUG(weight > 0);
  max_distance_weight_ = weight;
}

float Integrator::GetMaxColorWeight() const { return max_color_weight_; }

void Integrator::SetMaxColorWeight(float weight)
{
  VULCAN_DEBUG(weight > 0);
  max_color_weight_ = weight;
}

DepthIntegrator::DepthIntegrator(std::shared_ptr<Volume> volume) : Integrator(volume) {}

void DepthIntegrator::Integrate(const Frame& frame)
{
  const vkFrame f = ToC(frame);
  VK_ASSERT(vk_integrate_depth(volume_->GetHandle(), &f, depth_range_[0], depth_range_[1], max_distance_weight_,
                               volume_->GetStream()));
}

ColorIntegrator::ColorIntegrator(std::shared_ptr<Volume> volume) : Integrator(volume) {}

void ColorIntegrator::Integrate(const Frame& frame)
{
  VULCAN_ASSERT_MSG(frame.color_image != nullptr, "missing color image");
  const vkFrame f = ToC(frame);
  VK_ASSERT(vk_integrate_color(volume_->GetHandle(), &f, depth_range_[0], depth_range_[1], max_distance_weight_,
                               max_color_weight_, volume_->GetStream()));
}

void ColorIntegrator::IntegrateDepth(const Frame& frame)
{
  const vkFrame f = ToC(frame);
  VK_ASSERT(vk_integrate_depth(volume_->GetHandle(), &f, depth_range_[0], depth_range_[1], max_distance_weight_,
                               volume_->GetStream()));
}

LightIntegrator::LightIntegrator(std::shared_ptr<Volume> volume) : Integrator(volume), depth_threshold_(0.2f) {}

const Light& LightIntegrator::GetLight() const { return light_; }
void LightIntegrator::SetLight(const Light& light) { light_ = light; }

void LightIntegrator::ComputeFrameMask(const Frame& frame)
{
  VULCAN_ASSERT_MSG(frame.color_image != nullptr, "missing color image");
  const vkFrame f = ToC(frame);
  frame_mask_.Resize(f.depth_w, f.depth_h);
  VK_ASSERT(vk_compute_frame_mask(&f, depth_threshold_, frame_mask_.GetData(), volume_->GetStream()));
}

void LightIntegrator::IntegrateDepth(const Frame& frame)
{
  const vkFrame f = ToC(frame);
  VK_ASSERT(vk_integrate_depth(volume_->GetHandle(), &f, depth_range_[0], depth_range_[1], max_distance_weight_,
                               volume_->GetStream()));
}

void LightIntegrator::Integrate(const Frame& frame)
{
  VULCAN_ASSERT_MSG(frame.color_image != nullptr, "missing color image");
  VULCAN_ASSERT_MSG(frame.normal_image != nullptr, "missing normal image");
  const vkFrame f = ToC(frame);
  frame_mask_.Resize(f.depth_w, f.depth_h);
  vkLight light;
  std::memcpy(&light, &light_, sizeof(light));
  VK_ASSERT(vk_integrate_light(volume_->GetHandle(), &f, depth_range_[0], depth_range_[1], max_distance_weight_,
                               max_color_weight_, &light, depth_threshold_, frame_mask_.GetData(),
                               volume_->GetStream()));
}

// ---- Tracer (src/tracer.cpp:13-147) --------------------------------------------------------------

Tracer::Tracer(std::shared_ptr<const Volume> volume) :
  bounds_(VK_BOUNDS_W * VK_BOUNDS_H), volume_(volume), depth_range_(0.1f, 5.0f)
{
}

std::shared_ptr<const Volume> Tracer::GetVolume() const { return volume_; }
const Vector2f& Tracer::GetDepthRange() const { return depth_range_; }

void Tracer::SetDepthRange(const Vector2f& range)
{
  VULCAN_DEBUG(range[0] > 0 && range[0] < range[1]);
  depth_range_ = range;
}

void Tracer::SetDepthRange(float min, float max) { SetDepthRange(Vector2f(min, max)); }

void Tracer::Trace(Frame& frame)
{
  VULCAN_ASSERT_MSG(frame.depth_image != nullptr, "missing depth image");
  const int w = frame.depth_image->GetWidth(), h = frame.depth_image->GetHeight();
  EnsureColorImage(frame.color_image, w, h);
  EnsureColorImage(frame.normal_image, w, h);
  const vkFrame f = ToC(frame);
  volume_->SyncDistances();
  VK_ASSERT(vk_trace(volume_->GetHandle(), &f, depth_range_[0], depth_range_[1],
                     reinterpret_cast<float*>(bounds_.GetData()), frame.depth_image->GetData(),
                     reinterpret_cast<float*>(frame.color_image->GetData()),
                     reinterpret_cast<float*>(frame.normal_image->GetData()), volume_->GetStream()));
}

void Tracer::TraceKeys(const Frame& frame, unsigned long long* device_keys)
{
  const vkFrame f = ToC(frame);
  volume_->SyncDistances();
  VK_ASSERT(vk_trace_keys(volume_->GetHandle(), &f, depth_range_[0], depth_range_[1],
                          reinterpret_cast<float*>(bounds_.GetData()), device_keys, volume_->GetStream()));
}

void Tracer::ResolveKeys(Frame& frame, const unsigned long long* device_keys)
{
  VULCAN_ASSERT_MSG(frame.depth_image != nullptr, "missing depth image");
  const int w = frame.depth_image->GetWidth(), h = frame.depth_image->GetHeight();
  EnsureColorImage(frame.color_image, w, h);
  EnsureColorImage(frame.normal_image, w, h);
  const vkFrame f = ToC(frame);
  VK_ASSERT(vk_resolve_keys(device_keys, &f, frame.depth_image->GetData(),
                            reinterpret_cast<float*>(frame.color_image->GetData()),
                            reinterpret_cast<float*>(frame.normal_image->GetData()), volume_->GetStream()));
}

// ---- step-wise functions (include/vulcan/tracer.cuh:14-32), legacy default stream ------------------

void ComputePatches(const int* indices, const HashEntry* entries, const Transform& Tcw, const Projection& projection,
    float block_length, float min_depth, float max_depth, int block_count, int image_width, int image_height,
    int bounds_width, int bounds_height, Patch* patches, int* patch_count)
{
  const vkTransform T = ToC(Tcw);
  const vkProjection P = ToC(projection);
  VK_ASSERT(vk_compute_patches(indices, reinterpret_cast<const vkHashEntry*>(entries), &T, &P, block_length, min_depth,
                               max_depth, block_count, image_width, image_height, bounds_width, bounds_height,
                               reinterpret_cast<vkPatch*>(patches), patch_count, nullptr));
}

void ComputeBounds(const Patch* patches, Vector2f* bounds, int bounds_width, int patch_count)
{
  VK_ASSERT(vk_compute_bounds(reinterpret_cast<const vkPatch*>(patches), reinterpret_cast<float*>(bounds), bounds_width,
                              patch_count, nullptr));
}

void ComputePoints(const HashEntry* hash_entries, const Voxel* voxels, const Vector2f* bounds, int block_count,
    float block_length, float voxel_length, float trunc_length, const Transform& Twc, const Projection& projection,
    float* depths, Vector3f* colors, int image_width, int image_height, int bounds_width, int bounds_height)
{
  const vkTransform T = ToC(Twc);
  const vkProjection P = ToC(projection);
  VK_ASSERT(vk_compute_points(reinterpret_cast<const vkHashEntry*>(hash_entries), reinterpret_cast<const vkVoxel*>(voxels),
                              reinterpret_cast<const float*>(bounds), block_count, block_length, voxel_length,
                              trunc_length, &T, &P, depths, reinterpret_cast<float*>(colors), image_width, image_height,
                              bounds_width, bounds_height, nullptr));
}

void ComputeNormals(const float* depths, const Projection& projection, Vector3f* normals, int image_width,
    int image_height)
{
  const vkProjection P = ToC(projection);
  VK_ASSERT(vk_compute_normals(depths, &P, reinterpret_cast<float*>(normals), image_width, image_height, nullptr));
}

void ResetBoundsBuffer(Vector2f* bounds, int count)
{
  VK_ASSERT(vk_reset_bounds(reinterpret_cast<float*>(bounds), count, nullptr));
}

void FilterDepths(int image_width, int image_height, const float* src, float* dst)
{
  VK_ASSERT(vk_filter_depths(image_width, image_height, src, dst, nullptr));
}

} // namespace vulcan
