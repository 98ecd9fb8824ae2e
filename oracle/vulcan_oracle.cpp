// vulcan_oracle.cpp — CPU restatement of Vulcan's TSDF-fusion hot path.
//
// TEST INFRASTRUCTURE ONLY.  Nothing in the product (vulcan_b200/, include/) links,
// imports or calls this file; it is used by tests/, __graft_entry__.smoke() and the
// cpu_baseline / --impl reference legs of bench.py as the checker and as the timed CPU
// baseline.
//
// Each function restates one reference kernel with plain IEEE float arithmetic (build
// with -ffp-contract=off so that the compiler never fuses anything by itself) and cites the
// reference lines it follows (paths relative to /root/reference).  Two arithmetic modes
// (vo_set_math_mode), same association order in both:
//   0  "reference": a*b+c pairs are fused (std::fmaf) exactly where nvcc 12.9 -O3 with its default
//      -fmad=true fuses them when it builds the reference's sources for sm_100a - the pattern
//      was read off the SASS of that build (DESIGN.md "FMA map") and is pinned by the golden
//      vectors in tests/golden/ (outputs of that build on a B200);
//   1  "strict": one rounding per source operation (the reference built with -fmad=false).
// The reference has no CPU implementation of this path and stores no golden vectors
// (SURVEY.md 8c); parity of this oracle is pinned by running the reference's own CUDA
// sources (oracle/_ref, built by oracle/build_ref.sh) on a B200 on identical inputs —
// see tests/test_parity_gpu.py and tests/golden/ (vectors produced by that build).
//
// Build:  g++ -O2 -ffp-contract=off -fopenmp -shared -fPIC -o _build/libvulcan_oracle.so vulcan_oracle.cpp
#include <cfloat>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <vector>

namespace {

struct Voxel {  // include/vulcan/voxel.h:41-49
  float distance;
  float color[3];
  int16_t distance_weight;
  int16_t color_weight;
};
static_assert(sizeof(Voxel) == 20, "Voxel layout");

struct HashEntry {  // include/vulcan/hash.h:49-55, include/vulcan/block.h:79-85
  int16_t block[3];
  int16_t pad;
  int32_t data;
  int32_t next;
};
static_assert(sizeof(HashEntry) == 16, "HashEntry layout");

struct Patch {  // include/vulcan/tracer.h:13-22
  int16_t origin[2];
  int16_t size[2];
  float bounds[2];
};

enum { VIS_UNKNOWN = 0, VIS_FALSE = 1, VIS_TRUE = 2 };      // include/vulcan/types.h:6-11
enum { ALLOC_NONE = 0, ALLOC_MAIN = 1, ALLOC_EXCESS = 2 };  // include/vulcan/types.h:13-18

struct V3 { float x, y, z; };

int g_fma = 1;  // math mode 0 (reference contraction pattern) by default
inline float mad(float a, float b, float c) { return g_fma ? std::fmaf(a, b, c) : a * b + c; }    // a*b + c
inline float nmad(float a, float b, float c) { return g_fma ? std::fmaf(-a, b, c) : c - a * b; }  // c - a*b
inline float msub(float a, float b, float c) { return g_fma ? std::fmaf(a, b, -c) : a * b - c; }  // a*b - c

// column-major 4x4, include/vulcan/matrix.h:320-333
inline float M(const float* m, int r, int c) { return m[c * 4 + r]; }

// Transform::operator*(Vector4f), include/vulcan/transform.h:43-58
// M(r,0)*x + M(r,1)*y + M(r,2)*z + M(r,3)*w, left to right.  Reference SASS: FMUL y*m1;
// FFMA x*m0 + .; FFMA z*m2 + .; then FADD m3 (w == 1) or FFMA 0*m3 + . (w == 0).
inline float row_acc(const float* m, int r, float x, float y, float z) {
  return mad(z, M(m, r, 2), mad(x, M(m, r, 0), y * M(m, r, 1)));
}
inline V3 xform(const float* m, float x, float y, float z, float w) {
  V3 o;
  if (w == 1.0f) {
    o.x = row_acc(m, 0, x, y, z) + M(m, 0, 3);
    o.y = row_acc(m, 1, x, y, z) + M(m, 1, 3);
    o.z = row_acc(m, 2, x, y, z) + M(m, 2, 3);
  } else {  // w == 0
    o.x = mad(0.0f, M(m, 0, 3), row_acc(m, 0, x, y, z));
    o.y = mad(0.0f, M(m, 1, 3), row_acc(m, 1, x, y, z));
    o.z = mad(0.0f, M(m, 2, 3), row_acc(m, 2, x, y, z));
  }
  return o;
}

struct Proj { float fx, fy, cx, cy; };

// Projection::Project, include/vulcan/projection.h:63-71
inline void project(const Proj& P, const V3& X, float& u, float& v) {
  const float inv_w = 1.0f / X.z;
  u = mad(inv_w * P.fx, X.x, P.cx);   // FMUL inv_w*fx ; FFMA . * X.x + cx
  v = mad(inv_w * P.fy, X.y, P.cy);
}

// Projection::Unproject, include/vulcan/projection.h:80-89
inline V3 unproject(const Proj& P, float u, float v) {
  const float ifx = 1.0f / P.fx;
  const float ify = 1.0f / P.fy;
  V3 o;
  o.x = msub(ifx, u, P.cx * ifx);      // FMUL cx*ifx ; FFMA ifx*u - .
  o.y = msub(ify, v, P.cy * ify);
  o.z = 1;
  return o;
}

// Matrix::Dot, include/vulcan/matrix.h:158-171
inline float dot(const V3& a, const V3& b) {
  float r = 0;                  // FFMA chain from 0
  r = mad(a.x, b.x, r);
  r = mad(a.y, b.y, r);
  r = mad(a.z, b.z, r);
  return r;
}

// Matrix::Normalize, include/vulcan/matrix.h:118-141
inline V3 normalized(const V3& a) {
  const float inv_norm = 1.0f / std::sqrt(dot(a, a));
  V3 o = {a.x * inv_norm, a.y * inv_norm, a.z * inv_norm};
  return o;
}

inline V3 scale(const V3& a, float s) { V3 o = {a.x * s, a.y * s, a.z * s}; return o; }
inline V3 plus(const V3& a, const V3& b) { V3 o = {a.x + b.x, a.y + b.y, a.z + b.z}; return o; }
inline V3 minus(const V3& a, const V3& b) { V3 o = {a.x - b.x, a.y - b.y, a.z - b.z}; return o; }

template <typename T> inline T vmin(T a, T b) { return (b < a) ? b : a; }   // include/vulcan/math.h:9-14
template <typename T> inline T vmax(T a, T b) { return (b > a) ? b : a; }   // include/vulcan/math.h:16-21
template <typename T> inline T vclamp(T v, T lo, T hi) { return vmin(hi, vmax(lo, v)); }  // :23-28

// Q1 hash, src/volume.cu:168-180
inline uint32_t block_hash(int bx, int by, int bz, uint32_t K) {
  return (((uint32_t)bx * 73856093u) ^ ((uint32_t)by * 19349669u) ^ ((uint32_t)bz * 83492791u)) % K;
}

inline bool same_block(const HashEntry& e, int bx, int by, int bz) {
  return e.block[0] == (int16_t)bx && e.block[1] == (int16_t)by && e.block[2] == (int16_t)bz;
}

struct Volume {
  int M_, E_, B_;
  float voxel_length, trunc_length, min_depth, max_depth;
  Voxel* voxels;
  std::vector<HashEntry> entries;
  std::vector<int32_t> free_list;
  std::vector<uint8_t> alloc_types;
  std::vector<HashEntry> alloc_blocks;  // only .block used
  std::vector<uint8_t> visibility;
  std::vector<int32_t> visible;
  int voxel_pointer, excess_pointer;
  long long dropped;
  // work counters of the last trace (for algorithmic bytes)
  long long voxel_fetches, entry_fetches;
};

}  // namespace

extern "C" {

// Volume::Volume + Initialize, src/volume.cu:370-380,552-627
void* vo_create(int main_blocks, int excess_blocks) {
  Volume* v = new Volume();
  v->M_ = main_blocks;
  v->E_ = excess_blocks;
  v->B_ = main_blocks + excess_blocks;
  v->voxel_length = 0.008f;
  v->trunc_length = 0.04f;
  v->min_depth = 0.1f;
  v->max_depth = 5.0f;
  const size_t nvox = (size_t)v->B_ * 512;
  v->voxels = (Voxel*)std::malloc(nvox * sizeof(Voxel));
  Voxel empty;  // Voxel::Empty(), include/vulcan/voxel.h:29-37
  empty.distance = 1;
  empty.color[0] = empty.color[1] = empty.color[2] = 0;
  empty.distance_weight = 0;
  empty.color_weight = 0;
#pragma omp parallel for schedule(static)
  for (long long i = 0; i < (long long)nvox; ++i) v->voxels[i] = empty;
  HashEntry he;
  he.block[0] = he.block[1] = he.block[2] = 0;
  he.pad = 0;
  he.data = -1;
  he.next = -1;
  v->entries.assign(v->B_, he);
  v->free_list.resize(v->B_);
  for (int i = 0; i < v->B_; ++i) v->free_list[i] = i;
  v->alloc_types.assign(v->M_, ALLOC_NONE);
  v->alloc_blocks.assign(v->M_, he);
  v->visibility.assign(v->B_, VIS_FALSE);
  v->visible.reserve(v->B_);
  v->voxel_pointer = v->B_ - 1;
  v->excess_pointer = v->M_;
  v->dropped = 0;
  v->voxel_fetches = v->entry_fetches = 0;
  return v;
}

void vo_destroy(void* h) {
  Volume* v = (Volume*)h;
  if (!v) return;
  std::free(v->voxels);
  delete v;
}

// 0 = reference contraction pattern (default), 1 = strict (one rounding per operation)
void vo_set_math_mode(int mode) { g_fma = (mode == 0) ? 1 : 0; }
int vo_get_math_mode(void) { return g_fma ? 0 : 1; }

void vo_set_params(void* h, float voxel_length, float trunc_length, float min_depth, float max_depth) {
  Volume* v = (Volume*)h;
  v->voxel_length = voxel_length;
  v->trunc_length = trunc_length;
  v->min_depth = min_depth;
  v->max_depth = max_depth;
}

void* vo_voxels(void* h) { return ((Volume*)h)->voxels; }
void* vo_entries(void* h) { return ((Volume*)h)->entries.data(); }
void* vo_visibility(void* h) { return ((Volume*)h)->visibility.data(); }
void* vo_visible(void* h) { return ((Volume*)h)->visible.data(); }
int vo_visible_count(void* h) { return (int)((Volume*)h)->visible.size(); }
int vo_voxel_pointer(void* h) { return ((Volume*)h)->voxel_pointer; }
int vo_excess_pointer(void* h) { return ((Volume*)h)->excess_pointer; }
long long vo_dropped(void* h) { return ((Volume*)h)->dropped; }
long long vo_voxel_fetches(void* h) { return ((Volume*)h)->voxel_fetches; }
long long vo_entry_fetches(void* h) { return ((Volume*)h)->entry_fetches; }

}  // extern "C"

// ---- Volume::SetView ---------------------------------------------------------------

namespace {

// ResetBlockVisibility, src/volume.cu:465-471
void reset_visibility(Volume* v) {
  for (int i = 0; i < v->B_; ++i)
    if (v->visibility[i] == VIS_TRUE) v->visibility[i] = VIS_UNKNOWN;
}

// CreateAllocationRequestsKernel, src/volume.cu:86-301.  Pixels are visited in raster
// order, so the "last writer" of a contended bucket is the last pixel in that order; the
// reference leaves the winner unspecified (Q3).  Returns the number of requests written.
long long create_requests(Volume* v, const float* depths, int w, int h, const Proj& P, const float* Twd) {
  long long requests = 0;
  const float block_length = 8 * v->voxel_length;
  const float trunc = v->trunc_length;
  const uint32_t K = (uint32_t)v->M_;
  for (int y = 0; y < h; ++y)
    for (int x = 0; x < w; ++x) {
      V3 direction = unproject(P, x + 0.5f, y + 0.5f);
      direction = xform(Twd, direction.x, direction.y, direction.z, 0);
      const V3 origin = {M(Twd, 0, 3), M(Twd, 1, 3), M(Twd, 2, 3)};
      const float depth = depths[y * w + x];
      if (depth < v->min_depth || depth > v->max_depth) continue;
      // origin + depth * direction: FFMA dir*depth + origin
      const V3 Xwp = {mad(direction.x, depth, origin.x), mad(direction.y, depth, origin.y),
                      mad(direction.z, depth, origin.z)};
      direction = normalized(direction);
      // X -/+ trunc * direction: FFMA -/+dir*trunc + X
      const V3 begin = {nmad(direction.x, trunc, Xwp.x), nmad(direction.y, trunc, Xwp.y),
                        nmad(direction.z, trunc, Xwp.z)};
      const V3 end = {mad(direction.x, trunc, Xwp.x), mad(direction.y, trunc, Xwp.y), mad(direction.z, trunc, Xwp.z)};
      const int step_x = (direction.x < 0) ? -1 : 1;
      const int step_y = (direction.y < 0) ? -1 : 1;
      const int step_z = (direction.z < 0) ? -1 : 1;
      const float inv_block_length = 1.0f / block_length;
      int bx = (int)floorf(begin.x * inv_block_length);
      int by = (int)floorf(begin.y * inv_block_length);
      int bz = (int)floorf(begin.z * inv_block_length);
      const int ex = (int)floorf(end.x * inv_block_length);
      const int ey = (int)floorf(end.y * inv_block_length);
      const int ez = (int)floorf(end.z * inv_block_length);
      // block_length * (b + max(0, step)) - begin: FFMA bf*L - begin
      const float ox = msub((float)(bx + vmax(0, step_x)), block_length, begin.x);
      const float oy = msub((float)(by + vmax(0, step_y)), block_length, begin.y);
      const float oz = msub((float)(bz + vmax(0, step_z)), block_length, begin.z);
      float tmax_x = ox / direction.x;
      float tmax_y = oy / direction.y;
      float tmax_z = oz / direction.z;
      if (direction.x == 0) tmax_x = 1E20f;
      if (direction.y == 0) tmax_y = 1E20f;
      if (direction.z == 0) tmax_z = 1E20f;
      const float tdelta_x = (step_x * block_length) / direction.x;
      const float tdelta_y = (step_y * block_length) / direction.y;
      const float tdelta_z = (step_z * block_length) / direction.z;
      while (true) {
        const uint32_t hash_code = block_hash(bx, by, bz, K);
        HashEntry entry = v->entries[hash_code];
        if (same_block(entry, bx, by, bz)) {
          v->visibility[hash_code] = VIS_TRUE;  // also Q23 for block (0,0,0) in an empty bucket 0
        } else if (entry.data == -1) {
          v->visibility[hash_code] = VIS_TRUE;
          v->alloc_types[hash_code] = ALLOC_MAIN;
          v->alloc_blocks[hash_code].block[0] = (int16_t)bx;
          v->alloc_blocks[hash_code].block[1] = (int16_t)by;
          v->alloc_blocks[hash_code].block[2] = (int16_t)bz;
          ++requests;
        } else {
          bool found = false;
          while (entry.next != -1) {
            const int index = entry.next;
            entry = v->entries[index];
            if (same_block(entry, bx, by, bz)) {
              v->visibility[index] = VIS_TRUE;
              found = true;
              break;
            }
          }
          if (!found) {
            v->alloc_types[hash_code] = ALLOC_EXCESS;
            v->alloc_blocks[hash_code].block[0] = (int16_t)bx;
            v->alloc_blocks[hash_code].block[1] = (int16_t)by;
            v->alloc_blocks[hash_code].block[2] = (int16_t)bz;
            ++requests;
          }
        }
        if (tmax_x < tmax_y) {
          if (tmax_x < tmax_z) { bx += step_x; if (bx == ex + step_x) break; tmax_x += tdelta_x; }
          else                 { bz += step_z; if (bz == ez + step_z) break; tmax_z += tdelta_z; }
        } else {
          if (tmax_y < tmax_z) { by += step_y; if (by == ey + step_y) break; tmax_y += tdelta_y; }
          else                 { bz += step_z; if (bz == ez + step_z) break; tmax_z += tdelta_z; }
        }
      }
    }
  return requests;
}

// HandleAllocationRequestsKernel, src/volume.cu:303-368 (buckets visited in index order)
void handle_requests(Volume* v) {
  for (int index = 0; index < v->M_; ++index) {
    const uint8_t type = v->alloc_types[index];
    if (type == ALLOC_NONE) continue;
    HashEntry entry;
    entry.block[0] = v->alloc_blocks[index].block[0];
    entry.block[1] = v->alloc_blocks[index].block[1];
    entry.block[2] = v->alloc_blocks[index].block[2];
    entry.pad = 0;
    entry.data = -1;
    entry.next = -1;
    int entry_index = index;
    if (type == ALLOC_EXCESS) {
      int other_index = index;
      HashEntry other = v->entries[other_index];
      while (other.next != -1) {
        other_index = other.next;
        other = v->entries[other_index];
      }
      entry_index = v->excess_pointer++;
      if (entry_index < v->B_) {
        v->entries[other_index].next = entry_index;
        v->visibility[entry_index] = VIS_TRUE;
      }
    }
    const int voxel_index = v->voxel_pointer--;
    if (entry_index < v->B_ && voxel_index >= 0) {
      entry.data = v->free_list[voxel_index];
      v->entries[entry_index] = entry;
    } else {
      v->dropped++;
    }
    v->alloc_types[index] = ALLOC_NONE;
  }
}

// UpdateBlockVisibilityKernel, src/volume.cu:23-84 (entries visited in index order)
void update_visibility(Volume* v, int w, int h, const Proj& P, const float* Tdw) {
  const float block_length = 8 * v->voxel_length;
  v->visible.clear();
  for (int index = 0; index < v->B_; ++index) {
    const uint8_t vis = v->visibility[index];
    bool visible = (vis == VIS_TRUE);
    if (vis == VIS_UNKNOWN) {
      const HashEntry& e = v->entries[index];
      for (int i = 0; i < 8; ++i) {
        const float X = block_length * (e.block[0] + ((i & 1) >> 0));
        const float Y = block_length * (e.block[1] + ((i & 2) >> 1));
        const float Z = block_length * (e.block[2] + ((i & 4) >> 2));
        const V3 Xdp = xform(Tdw, X, Y, Z, 1);
        if (Xdp.z < 0) continue;
        float u, vv;
        project(P, Xdp, u, vv);
        if (u >= 0 && u <= w && vv >= 0 && vv <= h) { visible = true; break; }
      }
      if (!visible) v->visibility[index] = VIS_FALSE;
    }
    if (visible) v->visible.push_back(index);
  }
}

}  // namespace

extern "C" {

// One Volume::SetView pass (src/volume.cu:430-437).  T = 32 floats: column-major
// depth_to_world matrix followed by its inverse.  Returns the number of allocation
// requests the pass generated (0 => the pass was at the fixed point).
long long vo_set_view(void* h, const float* depth, int w, int hgt, const float* proj4, const float* T) {
  Volume* v = (Volume*)h;
  const Proj P = {proj4[0], proj4[1], proj4[2], proj4[3]};
  reset_visibility(v);
  const long long req = create_requests(v, depth, w, hgt, P, T);
  handle_requests(v);
  update_visibility(v, w, hgt, P, T + 16);
  return req;
}

// SetView repeated until a pass generates no request (apps/vulcan/vulcan.cu:316-318 calls it
// three times; tests/tracer_test.cu:298-303 loops to the fixed point).  Returns the passes run.
int vo_set_view_converged(void* h, const float* depth, int w, int hgt, const float* proj4, const float* T,
                          int max_passes) {
  int passes = 0;
  while (passes < max_passes) {
    ++passes;
    if (vo_set_view(h, depth, w, hgt, proj4, T) == 0) break;
  }
  return passes;
}

}  // extern "C"

// ---- integrators -------------------------------------------------------------------

namespace {

// Matrix::operator*(Matrix), include/vulcan/matrix.h:296-316 (column-major 4x4)
void matmul4(const float* A, const float* Bm, float* C) {
  for (int p = 0; p < 4; ++p)
    for (int m = 0; m < 4; ++m) {
      float acc = 0;
      for (int n = 0; n < 4; ++n) acc += A[n * 4 + m] * Bm[p * 4 + n];
      C[p * 4 + m] = acc;
    }
}

inline V3 voxel_world(const HashEntry& e, float block_length, float voxel_length, int x, int y, int z) {
  // src/depth_integrator.cu:36-39
  V3 Xwp;
  // FMUL origin*L ; FFMA (v + 0.5)*l + .
  Xwp.x = mad((float)x + 0.5f, voxel_length, block_length * (float)e.block[0]);
  Xwp.y = mad((float)y + 0.5f, voxel_length, block_length * (float)e.block[1]);
  Xwp.z = mad((float)z + 0.5f, voxel_length, block_length * (float)e.block[2]);
  return Xwp;
}

// IntegrateKernel, src/depth_integrator.cu:17-80 (= IntegrateDepthKernel of the colour and
// light integrators, src/color_integrator.cu:17-80, src/light_integrator.cu:105-168)
void integrate_depth(Volume* v, const float* depths, int w, int h, const Proj& P, const float* Tdw,
                     float min_depth, float max_depth, float max_dist_weight) {
  const float voxel_length = v->voxel_length;
  const float block_length = 8 * voxel_length;
  const float trunc = v->trunc_length;
  const int n = (int)v->visible.size();
#pragma omp parallel for schedule(dynamic, 16)
  for (int b = 0; b < n; ++b) {
    const HashEntry entry = v->entries[v->visible[b]];
    for (int z = 0; z < 8; ++z)
      for (int y = 0; y < 8; ++y)
        for (int x = 0; x < 8; ++x) {
          const V3 Xwp = voxel_world(entry, block_length, voxel_length, x, y, z);
          const V3 Xdp = xform(Tdw, Xwp.x, Xwp.y, Xwp.z, 1);
          float u, vv;
          project(P, Xdp, u, vv);
          if (u >= 0 && u < w && vv >= 0 && vv < h) {  // Q9a: no z > 0 test
            const float depth = depths[(int)vv * w + (int)u];
            if (depth < min_depth || depth > max_depth) continue;
            const float distance = depth - Xdp.z;
            if (distance > -trunc) {
              Voxel& voxel = v->voxels[(long long)entry.data * 512 + z * 64 + y * 8 + x];
              const float curr_dist = vmin(1.0f, distance / trunc);
              const float dist_weight = voxel.distance_weight + 1;
              // (w*D + cur) / (w + 1): FFMA D*w + cur ; IEEE division (Q9b)
              voxel.distance = mad(voxel.distance, (float)voxel.distance_weight, curr_dist) / dist_weight;
              voxel.distance_weight = (int16_t)vmin(max_dist_weight, dist_weight);
            }
          }
        }
  }
}

// IntegrateColorKernel, src/color_integrator.cu:82-135
void integrate_color(Volume* v, const float* colors, int w, int h, const Proj& P, const float* Tcw,
                     float max_color_weight) {
  const float voxel_length = v->voxel_length;
  const float block_length = 8 * voxel_length;
  const int n = (int)v->visible.size();
#pragma omp parallel for schedule(dynamic, 16)
  for (int b = 0; b < n; ++b) {
    const HashEntry entry = v->entries[v->visible[b]];
    for (int z = 0; z < 8; ++z)
      for (int y = 0; y < 8; ++y)
        for (int x = 0; x < 8; ++x) {
          const V3 Xwp = voxel_world(entry, block_length, voxel_length, x, y, z);
          const V3 Xcp = xform(Tcw, Xwp.x, Xwp.y, Xwp.z, 1);
          float u, vv;
          project(P, Xcp, u, vv);
          if (u >= 0 && u < w && vv >= 0 && vv < h) {
            Voxel& voxel = v->voxels[(long long)entry.data * 512 + z * 64 + y * 8 + x];
            if (fabsf(voxel.distance) < 1.0) {  // Q10
              const float* c = colors + 3 * ((long long)(int)vv * w + (int)u);
              const float cw = voxel.color_weight;
              const float color_weight = voxel.color_weight + 1;
              const float inv = 1.0f / color_weight;  // Matrix::operator/=, include/vulcan/matrix.h:288-295
              for (int k = 0; k < 3; ++k) voxel.color[k] = mad(voxel.color[k], cw, c[k]) * inv;  // FFMA C*cw + c
              voxel.color_weight = (int16_t)vmin(max_color_weight, color_weight);
            }
          }
        }
  }
}

// ComputeFrameMaskKernel<16,3>, src/light_integrator.cu:17-103 (Q14: tile filled from image
// offset -1, window read at +3 +/- 3 => window = image offsets [-1, +5] around the pixel)
void frame_mask(int w, int h, const float* depths, const float* colors, float threshold, float* mask) {
#pragma omp parallel for schedule(static)
  for (int y = 0; y < h; ++y)
    for (int x = 0; x < w; ++x) {
      const int index = y * w + x;
      const float* c = colors + 3 * (long long)index;
      if (c[0] < 0.02f || c[0] > 0.98f || c[1] < 0.02f || c[1] > 0.98f || c[2] < 0.02f || c[2] > 0.98f) {
        mask[index] = 0;
        continue;
      }
      float dmin = +FLT_MAX, dmax = -FLT_MAX;
      for (int i = -3; i <= 3; ++i)
        for (int j = -3; j <= 3; ++j) {
          const int vx = x - 1 + 3 + j, vy = y - 1 + 3 + i;
          float d = 0;
          if (vx >= 0 && vx < w && vy >= 0 && vy < h) d = depths[vy * w + vx];
          dmin = fminf(d, dmin);
          dmax = fmaxf(d, dmax);
        }
      mask[index] = (dmax - dmin <= threshold) ? 1.0f : 0.0f;
    }
}

// IntegrateColorKernel of the light integrator, src/light_integrator.cu:170-250 with
// Light::GetShading, include/vulcan/light.h:52-60
void integrate_light_color(Volume* v, const float* mask, const float* colors, const float* normals, int w, int h,
                           const Proj& Pd, const Proj& Pc, const float* Tdw, const float* Tcw, const float* Tcd,
                           float intensity, const float* lpos, float max_color_weight) {
  const float voxel_length = v->voxel_length;
  const float block_length = 8 * voxel_length;
  const int n = (int)v->visible.size();
#pragma omp parallel for schedule(dynamic, 16)
  for (int b = 0; b < n; ++b) {
    const HashEntry entry = v->entries[v->visible[b]];
    for (int z = 0; z < 8; ++z)
      for (int y = 0; y < 8; ++y)
        for (int x = 0; x < 8; ++x) {
          const V3 Xwp = voxel_world(entry, block_length, voxel_length, x, y, z);
          const V3 Xdp = xform(Tdw, Xwp.x, Xwp.y, Xwp.z, 1);
          const V3 Xcp = xform(Tcw, Xwp.x, Xwp.y, Xwp.z, 1);
          float du, dv, cu, cv;
          project(Pd, Xdp, du, dv);
          project(Pc, Xcp, cu, cv);
          if (!(du >= 0 && du < w && dv >= 0 && dv < h && cu >= 0 && cu < w && cv >= 0 && cv < h)) continue;
          const int depth_index = (int)dv * w + (int)du;
          const int color_index = (int)cv * w + (int)cu;
          if (!(mask[depth_index] > 0.5f)) continue;
          Voxel& voxel = v->voxels[(long long)entry.data * 512 + z * 64 + y * 8 + x];
          if (!(fabsf(voxel.distance) < 1.0)) continue;
          float cur[3] = {colors[3 * (long long)color_index], colors[3 * (long long)color_index + 1],
                          colors[3 * (long long)color_index + 2]};
          const float* nrm = normals + 3 * (long long)depth_index;
          const V3 Xcn = xform(Tcd, nrm[0], nrm[1], nrm[2], 0);
          const V3 delta = {lpos[0] - Xcp.x, lpos[1] - Xcp.y, lpos[2] - Xcp.z};
          const V3 direction = normalized(delta);
          const float distance_squared = dot(delta, delta);
          const float cos_theta = dot(Xcn, direction);
          const float shading = intensity * cos_theta / distance_squared;
          if (shading > 0.05f) {
            const float inv_s = 1.0f / shading;
            const float cw = voxel.color_weight;
            const float color_weight = voxel.color_weight + 1;
            const float inv = 1.0f / color_weight;
            // prev + curr/shading: FMUL C*cw ; FFMA cur*(1/s) + .
            for (int k = 0; k < 3; ++k) voxel.color[k] = mad(cur[k], inv_s, voxel.color[k] * cw) * inv;
            voxel.color_weight = (int16_t)vmin(max_color_weight, color_weight);
          }
        }
  }
}

}  // namespace

extern "C" {

// DepthIntegrator::Integrate, src/depth_integrator.cu:89-115
void vo_integrate_depth(void* h, const float* depth, int w, int hgt, const float* proj4, const float* T,
                        float min_depth, float max_depth, float max_dist_weight) {
  const Proj P = {proj4[0], proj4[1], proj4[2], proj4[3]};
  integrate_depth((Volume*)h, depth, w, hgt, P, T + 16, min_depth, max_depth, max_dist_weight);
}

// ColorIntegrator::Integrate, src/color_integrator.cu:144-204.  Tcd = 32 floats like T.
void vo_integrate_color(void* h, const float* depth, int w, int hgt, const float* dproj4, const float* T,
                        const float* color, int cw, int ch, const float* cproj4, const float* Tcd,
                        float min_depth, float max_depth, float max_dist_weight, float max_color_weight) {
  const Proj Pd = {dproj4[0], dproj4[1], dproj4[2], dproj4[3]};
  const Proj Pc = {cproj4[0], cproj4[1], cproj4[2], cproj4[3]};
  integrate_depth((Volume*)h, depth, w, hgt, Pd, T + 16, min_depth, max_depth, max_dist_weight);
  float Tcw[16];
  matmul4(Tcd, T + 16, Tcw);  // Tcd * Twd.Inverse(), src/color_integrator.cu:191-193
  integrate_color((Volume*)h, color, cw, ch, Pc, Tcw, max_color_weight);
}

void vo_frame_mask(int w, int hgt, const float* depth, const float* color, float threshold, float* mask) {
  frame_mask(w, hgt, depth, color, threshold, mask);
}

// LightIntegrator::Integrate, src/light_integrator.cu:270-354
void vo_integrate_light(void* h, const float* depth, int w, int hgt, const float* dproj4, const float* T,
                        const float* color, const float* normal, const float* cproj4, const float* Tcd,
                        float min_depth, float max_depth, float max_dist_weight, float max_color_weight,
                        float intensity, const float* light_pos, float depth_threshold, float* mask) {
  const Proj Pd = {dproj4[0], dproj4[1], dproj4[2], dproj4[3]};
  const Proj Pc = {cproj4[0], cproj4[1], cproj4[2], cproj4[3]};
  frame_mask(w, hgt, depth, color, depth_threshold, mask);
  integrate_depth((Volume*)h, depth, w, hgt, Pd, T + 16, min_depth, max_depth, max_dist_weight);
  float Tcw[16];
  matmul4(Tcd, T + 16, Tcw);
  integrate_light_color((Volume*)h, mask, color, normal, w, hgt, Pd, Pc, T + 16, Tcw, Tcd, intensity, light_pos,
                        max_color_weight);
}

}  // extern "C"

// ---- Tracer --------------------------------------------------------------------------

namespace {

// `(short)floorf(u)` as the reference's machine code does it (F2I to a saturating s32,
// NaN -> 0, then 16-bit VIMNMX on the low half of that register): in-range values truncate
// like a C cast, wild values (blocks behind the camera) wrap instead of saturating
inline int16_t to_short(float f) {
  int32_t i;
  if (f != f) i = 0;
  else if (f >= 2147483648.0f) i = INT32_MAX;
  else if (f <= -2147483648.0f) i = INT32_MIN;
  else i = (int32_t)f;
  return (int16_t)(uint16_t)((uint32_t)i & 0xffffu);
}

// ComputePatchesKernel, src/tracer.cu:13-87, for one block; returns patch count
int block_patches(const HashEntry& entry, const float* Tcw, const Proj& P, float block_length, float min_depth,
                  float max_depth, int image_w, int image_h, int bw, int bh, Patch* out) {
  int16_t bmax[2] = {-1, -1};
  int16_t bmin[2] = {(int16_t)bw, (int16_t)bh};
  float db[2] = {+FLT_MAX, -FLT_MAX};
  for (int z = 0; z <= 1; ++z) {
    const float Z = block_length * (z + entry.block[2]);
    for (int y = 0; y <= 1; ++y) {
      const float Y = block_length * (y + entry.block[1]);
      for (int x = 0; x <= 1; ++x) {
        const float X = block_length * (x + entry.block[0]);
        // rows fused the other way round here (x and y products hoisted out of the corner loops):
        // FMUL x*m0 ; FFMA y*m1 + . ; FFMA z*m2 + . ; FADD m3
        V3 Xcp;
        Xcp.x = mad(Z, M(Tcw, 0, 2), mad(Y, M(Tcw, 0, 1), X * M(Tcw, 0, 0))) + M(Tcw, 0, 3);
        Xcp.y = mad(Z, M(Tcw, 1, 2), mad(Y, M(Tcw, 1, 1), X * M(Tcw, 1, 0))) + M(Tcw, 1, 3);
        Xcp.z = mad(Z, M(Tcw, 2, 2), mad(Y, M(Tcw, 2, 1), X * M(Tcw, 2, 0))) + M(Tcw, 2, 3);
        float u, v;
        project(P, Xcp, u, v);
        u = bw * u / image_w;
        v = bh * v / image_h;
        bmin[0] = vclamp<int16_t>(vmin(to_short(floorf(u)), bmin[0]), 0, bw - 1);
        bmin[1] = vclamp<int16_t>(vmin(to_short(floorf(v)), bmin[1]), 0, bh - 1);
        bmax[0] = vclamp<int16_t>(vmax(to_short(ceilf(u)), bmax[0]), 0, bw - 1);
        bmax[1] = vclamp<int16_t>(vmax(to_short(ceilf(v)), bmax[1]), 0, bh - 1);
        db[0] = vclamp(vmin(Xcp.z, db[0]), min_depth, max_depth);
        db[1] = vclamp(vmax(Xcp.z, db[1]), min_depth, max_depth);
      }
    }
  }
  const int rx = bmax[0] - bmin[0];
  const int ry = bmax[1] - bmin[1];
  const int gx = (rx + 16 - 1) / 16;  // Q12
  const int gy = (ry + 16 - 1) / 16;
  const int count = (db[1] > db[0]) ? gx * gy : 0;
  if (count == 0) return 0;
  for (int i = 0; i < gy; ++i)
    for (int j = 0; j < gx; ++j) {
      Patch& p = out[i * gx + j];
      p.origin[0] = (int16_t)(bmin[0] + 16 * j);
      p.origin[1] = (int16_t)(bmin[1] + 16 * i);
      p.size[0] = (int16_t)vmin(16, bmax[0] - p.origin[0] + 1);
      p.size[1] = (int16_t)vmin(16, bmax[1] - p.origin[1] + 1);
      p.bounds[0] = db[0];
      p.bounds[1] = db[1];
    }
  return count;
}

struct Tap { float distance; float color[3]; int color_weight; };

struct Fetch { long long voxels, entries; };

// lookup used by ComputePointsKernel (src/tracer.cu:363-371) and GetVoxel (:151-175)
inline int find_data(const Volume* v, int bx, int by, int bz, Fetch& f) {
  HashEntry entry = v->entries[block_hash(bx, by, bz, (uint32_t)v->M_)];
  ++f.entries;
  while (!same_block(entry, bx, by, bz) && entry.next != -1) {
    entry = v->entries[entry.next];
    ++f.entries;
  }
  return same_block(entry, bx, by, bz) ? entry.data : -1;
}

inline Tap tap_at(const Volume* v, int data, int offset, Fetch& f) {
  const Voxel& vx = v->voxels[(long long)data * 512 + offset];
  ++f.voxels;
  Tap t = {vx.distance, {vx.color[0], vx.color[1], vx.color[2]}, vx.color_weight};
  return t;
}

// GetVoxel, src/tracer.cu:114-188
Tap get_voxel(const Volume* v, int bx, int by, int bz, int vx, int vy, int vz, Fetch& f) {
  if (vx < 0) { --bx; vx = 8 + vx; } else if (vx >= 8) { ++bx; vx = vx - 8; }
  if (vy < 0) { --by; vy = 8 + vy; } else if (vy >= 8) { ++by; vy = vy - 8; }
  if (vz < 0) { --bz; vz = 8 + vz; } else if (vz >= 8) { ++bz; vz = vz - 8; }
  const int data = find_data(v, bx, by, bz, f);
  if (data != -1) return tap_at(v, data, vz * 64 + vy * 8 + vx, f);
  Tap t = {1.0f, {0, 0, 0}, 0};  // Voxel::Empty(), Q16
  return t;
}

// GetInterpolatedDistance, src/tracer.cu:190-315
void interpolate(const Volume* v, float block_length, float voxel_length, int bx, int by, int bz, int data,
                 const V3& p, float& sdf, float* color, Fetch& f) {
  // (p - b*L) / l: FFMA -b*L + p ; IEEE division
  const float wx = nmad((float)bx, block_length, p.x) / voxel_length;
  const float wy = nmad((float)by, block_length, p.y) / voxel_length;
  const float wz = nmad((float)bz, block_length, p.z) / voxel_length;
  const int i0x = (int)floorf(wx - 0.5f);
  const int i0y = (int)floorf(wy - 0.5f);
  const int i0z = (int)floorf(wz - 0.5f);
  Tap t[8];  // index z*4 + y*2 + x  (v000, v001, v010, v011, v100, ...: last digit = x)
  if (i0x >= 0 && i0y >= 0 && i0z >= 0 && i0x < 7 && i0y < 7 && i0z < 7) {
    for (int c = 0; c < 8; ++c)
      t[c] = tap_at(v, data, (i0z + (c >> 2)) * 64 + (i0y + ((c >> 1) & 1)) * 8 + (i0x + (c & 1)), f);
  } else {
    for (int c = 0; c < 8; ++c)
      t[c] = get_voxel(v, bx, by, bz, i0x + (c & 1), i0y + ((c >> 1) & 1), i0z + (c >> 2), f);
  }
  const float w1[3] = {wx - (i0x + 0.5f), wy - (i0y + 0.5f), wz - (i0z + 0.5f)};
  const float w0[3] = {1.0f - w1[0], 1.0f - w1[1], 1.0f - w1[2]};
  // a*w0 + b*w1: FMUL b*w1 ; FFMA a*w0 + .
  const float n00 = mad(t[0].distance, w0[0], t[1].distance * w1[0]);
  const float n01 = mad(t[2].distance, w0[0], t[3].distance * w1[0]);
  const float n10 = mad(t[4].distance, w0[0], t[5].distance * w1[0]);
  const float n11 = mad(t[6].distance, w0[0], t[7].distance * w1[0]);
  const float n0 = mad(n00, w0[1], n01 * w1[1]);
  const float n1 = mad(n10, w0[1], n11 * w1[1]);
  float total = 0;
  float col[3] = {0, 0, 0};
  for (int c = 0; c < 8; ++c) {
    const float wz_ = (c >> 2) ? w1[2] : w0[2];
    const float wy_ = ((c >> 1) & 1) ? w1[1] : w0[1];
    const float wx_ = (c & 1) ? w1[0] : w0[0];
    // Q15: `wz * wy * wx * (cw > 0) ? 1 : 0`
    const float weight = (wz_ * wy_ * wx_ * (t[c].color_weight > 0)) ? 1 : 0;
    total += weight;
    for (int k = 0; k < 3; ++k) col[k] = mad(weight, t[c].color[k], col[k]);  // FFMA chain from 0
  }
  if (total > 0) {
    const float inv = 1.0f / total;
    for (int k = 0; k < 3; ++k) col[k] *= inv;
  }
  color[0] = col[0]; color[1] = col[1]; color[2] = col[2];
  sdf = mad(n0, w0[2], n1 * w1[2]);
}

// p += dir * s (include/vulcan/matrix.h:212): FFMA dir*s + p
inline V3 step(const V3& p, const V3& dir, float s) {
  V3 o = {mad(dir.x, s, p.x), mad(dir.y, s, p.y), mad(dir.z, s, p.z)};
  return o;
}

// ComputeNormalsKernel<16>, src/frame.cu:7-122
void normals(const float* depths, const Proj& P, float* out, int w, int h) {
  auto at = [&](int x, int y) -> float { return (x >= 0 && x < w && y >= 0 && y < h) ? depths[y * w + x] : 0.0f; };
#pragma omp parallel for schedule(static)
  for (int y = 0; y < h; ++y)
    for (int x = 0; x < w; ++x) {
      float n[3] = {0, 0, 0};
      const float depth = at(x, y);
      if (depth > 0) {
        const V3 z0 = scale(unproject(P, (x + 0) + 0.5f, (y + 0) + 0.5f), depth);
        const int ox[4] = {-2, +2, 0, 0}, oy[4] = {0, 0, -2, +2};
        V3 q[4];
        for (int k = 0; k < 4; ++k) {
          const float d = at(x + ox[k], y + oy[k]);
          q[k] = (d == 0) ? z0 : scale(unproject(P, (x + ox[k]) + 0.5f, (y + oy[k]) + 0.5f), d);
        }
        const V3 dx = minus(q[0], q[1]);
        const V3 dy = minus(q[2], q[3]);
        if (dot(dx, dx) > 0 && dot(dy, dy) > 0) {
          V3 c;  // dy.Cross(dx), include/vulcan/matrix.h:174-183: FMUL a.z*b.y ; FFMA a.y*b.z - .
          c.x = msub(dy.y, dx.z, dy.z * dx.y);
          c.y = msub(dy.z, dx.x, dy.x * dx.z);
          c.z = msub(dy.x, dx.y, dy.y * dx.x);
          c = normalized(c);
          n[0] = c.x; n[1] = c.y; n[2] = c.z;
        }
      }
      float* o = out + 3 * ((long long)y * w + x);
      o[0] = n[0]; o[1] = n[1]; o[2] = n[2];
    }
}

}  // namespace

extern "C" {

// Tracer::ComputePatches for an explicit index list (tests/tracer_test.cu:22-166 scenario).
// Tcw = 16 floats column-major.  Returns the patch count.
int vo_compute_patches(void* h, const int* indices, int count, const float* Tcw, const float* proj4,
                       float block_length, float min_depth, float max_depth, int image_w, int image_h, int bw,
                       int bh, void* patches_out) {
  Volume* v = (Volume*)h;
  const Proj P = {proj4[0], proj4[1], proj4[2], proj4[3]};
  Patch* out = (Patch*)patches_out;
  int total = 0;
  for (int i = 0; i < count; ++i)
    total += block_patches(v->entries[indices[i]], Tcw, P, block_length, min_depth, max_depth, image_w, image_h, bw,
                           bh, out + total);
  return total;
}

// ComputeBoundsKernel + FillKernel, src/tracer.cu:89-112,494-500
void vo_compute_bounds(const void* patches_in, int count, float* bounds, int bw, int bh, int reset) {
  const Patch* patches = (const Patch*)patches_in;
  if (reset)
    for (int i = 0; i < bw * bh; ++i) { bounds[2 * i] = +FLT_MAX; bounds[2 * i + 1] = -FLT_MAX; }
  for (int k = 0; k < count; ++k) {
    const Patch& p = patches[k];
    for (int i = 0; i < p.size[1]; ++i)
      for (int j = 0; j < p.size[0]; ++j) {
        const int pixel = (p.origin[1] + i) * bw + p.origin[0] + j;
        bounds[2 * pixel] = vmin(p.bounds[0], bounds[2 * pixel]);
        bounds[2 * pixel + 1] = vmax(p.bounds[1], bounds[2 * pixel + 1]);
      }
  }
}

void vo_compute_normals(const float* depth, const float* proj4, float* out, int w, int hgt) {
  const Proj P = {proj4[0], proj4[1], proj4[2], proj4[3]};
  normals(depth, P, out, w, hgt);
}

// diagnostics: per-pixel march iteration counts of the next vo_compute_points (tools/ray_stats.py)
static int* g_iter_out = nullptr;
void vo_set_iter_output(int* buf) { g_iter_out = buf; }

// ComputePointsKernel, src/tracer.cu:317-451
void vo_compute_points(void* h, const float* bounds, const float* T, const float* proj4, float* depths,
                       float* colors, int w, int hgt, int bw, int bh) {
  Volume* v = (Volume*)h;
  const Proj P = {proj4[0], proj4[1], proj4[2], proj4[3]};
  const float* Twc = T;
  const float* Tcw = T + 16;
  const float voxel_length = v->voxel_length;
  const float block_length = 8 * voxel_length;
  const float trunc = v->trunc_length;
  long long tot_vox = 0, tot_ent = 0;
#pragma omp parallel for schedule(dynamic, 4) reduction(+ : tot_vox, tot_ent)
  for (int y = 0; y < hgt; ++y)
    for (int x = 0; x < w; ++x) {
      Fetch f = {0, 0};
      const int px = bw * x / w;
      const int py = bh * y / hgt;
      const float lo = bounds[2 * (py * bw + px)], hi = bounds[2 * (py * bw + px) + 1];
      float depth = 0, final_depth = 0;
      float color[3] = {0, 0, 0};
      int iters_done = 0;
      if (lo < hi) {
        const V3 Xcp = scale(unproject(P, x + 0.5f, y + 0.5f), lo);
        const V3 Xwp = xform(Twc, Xcp.x, Xcp.y, Xcp.z, 1.0f);
        const V3 dir = normalized(xform(Twc, Xcp.x, Xcp.y, Xcp.z, 0.0f));
        V3 p = Xwp;
        depth = lo;
        int iters = 0;
        do {
          const int bx = (int)floorf(p.x / block_length);
          const int by = (int)floorf(p.y / block_length);
          const int bz = (int)floorf(p.z / block_length);
          const int data = find_data(v, bx, by, bz, f);
          if (data != -1) {
            const float wx = nmad((float)bx, block_length, p.x) / voxel_length;
            const float wy = nmad((float)by, block_length, p.y) / voxel_length;
            const float wz = nmad((float)bz, block_length, p.z) / voxel_length;
            const int vx = (int)wx, vy = (int)wy, vz = (int)wz;  // Q22
            float sdf = v->voxels[(long long)data * 512 + vz * 64 + vy * 8 + vx].distance;
            ++f.voxels;
            if (sdf <= 0.1f && sdf >= -0.5f) interpolate(v, block_length, voxel_length, bx, by, bz, data, p, sdf, color, f);
            if (sdf <= 0.0f) {
              p = step(p, dir, trunc * sdf);
              const int bx2 = (int)floorf(p.x / block_length);
              const int by2 = (int)floorf(p.y / block_length);
              const int bz2 = (int)floorf(p.z / block_length);
              const int data2 = find_data(v, bx2, by2, bz2, f);
              if (data2 != -1) {
                interpolate(v, block_length, voxel_length, bx2, by2, bz2, data2, p, sdf, color, f);
                p = step(p, dir, trunc * sdf);
              }
              final_depth = xform(Tcw, p.x, p.y, p.z, 1.0f).z;
              break;
            } else {
              p = step(p, dir, vmax(voxel_length, trunc * sdf));
            }
          } else {
            p = step(p, dir, block_length);
          }
          depth = xform(Tcw, p.x, p.y, p.z, 1.0f).z;
          if (++iters >= 500) { color[0] = 1; color[1] = 0; color[2] = 0; break; }
        } while (depth < hi);
        iters_done = iters + 1;
      }
      if (g_iter_out) g_iter_out[(long long)y * w + x] = iters_done;
      const long long pixel = (long long)y * w + x;
      depths[pixel] = final_depth;
      colors[3 * pixel] = color[0]; colors[3 * pixel + 1] = color[1]; colors[3 * pixel + 2] = color[2];
      tot_vox += f.voxels;
      tot_ent += f.entries;
    }
  v->voxel_fetches = tot_vox;
  v->entry_fetches = tot_ent;
}

// Tracer::Trace, src/tracer.cpp:41-47, with the hard-coded 80x60 bounds grid
void vo_trace(void* h, const float* T, const float* proj4, int w, int hgt, float min_depth, float max_depth,
              float* bounds /* 80*60*2 */, float* out_depth, float* out_color, float* out_normal) {
  Volume* v = (Volume*)h;
  const int bw = 80, bh = 60;
  std::vector<Patch> patches;
  patches.resize(1);
  std::vector<Patch> all;
  const Proj P = {proj4[0], proj4[1], proj4[2], proj4[3]};
  Patch tmp[256 + 64];
  for (size_t i = 0; i < v->visible.size(); ++i) {
    const int n = block_patches(v->entries[v->visible[i]], T + 16, P, 8 * v->voxel_length, min_depth, max_depth, w, hgt,
                                bw, bh, tmp);
    all.insert(all.end(), tmp, tmp + n);
  }
  vo_compute_bounds(all.data(), (int)all.size(), bounds, bw, bh, 1);
  vo_compute_points(h, bounds, T, proj4, out_depth, out_color, w, hgt, bw, bh);
  if (out_normal) normals(out_depth, P, out_normal, w, hgt);
}

}  // extern "C"

// ---- image pyramid + DepthTracker ----------------------------------------

namespace {

// Evaluate<>, src/depth_tracker.cu:16-99.  Returns false when the pixel contributes nothing.
// Twc, Twm, Tmw: column-major 4x4 (frame pose, keyframe pose, keyframe pose inverse).
bool track_evaluate(const float* Twc, const float* Twm, const float* Tmw, const float* key_depth,
                    const float* key_normal, const Proj& kP, int kw, int kh, const float* depth, const float* normal,
                    const Proj& P, int w, int x, int y, float& residual, float* J) {
  const int index = y * w + x;
  const float d = depth[index];
  if (!(d > 0)) return false;
  const V3 Xc = scale(unproject(P, x + 0.5f, y + 0.5f), d);           // :38-40
  const V3 Xw = xform(Twc, Xc.x, Xc.y, Xc.z, 1.0f);                    // :41
  const V3 Xm = xform(Tmw, Xw.x, Xw.y, Xw.z, 1.0f);                    // :42
  float ku, kv;
  project(kP, Xm, ku, kv);                                            // :43
  if (!(ku >= 0 && ku < kw && kv >= 0 && kv < kh)) return false;      // :45-46
  const int kindex = (int)kv * kw + (int)ku;                          // :48-50
  const float kd = key_depth[kindex];
  if (!(kd > 0)) return false;                                        // :53
  const V3 fn = xform(Twc, normal[3 * index], normal[3 * index + 1], normal[3 * index + 2], 0.0f);             // :55-56
  const V3 kn = xform(Twm, key_normal[3 * kindex], key_normal[3 * kindex + 1], key_normal[3 * kindex + 2], 0.0f);  // :58-59
  if (!(dot(kn, kn) > 0.0f && dot(fn, kn) > 0.5f)) return false;      // :61-62
  const V3 Ym = scale(unproject(kP, std::floor(ku) + 0.5f, std::floor(kv) + 0.5f), kd);  // :64-67
  const V3 Yw = xform(Twm, Ym.x, Ym.y, Ym.z, 1.0f);                    // :68
  const V3 delta = minus(Xw, Yw);
  if (!(dot(delta, delta) < 0.05f)) return false;                     // :71
  residual = dot(delta, kn);                                          // :73
  J[0] = msub(kn.z, Xw.y, kn.y * Xw.z);                                // :77-79
  J[1] = msub(kn.x, Xw.z, kn.z * Xw.x);
  J[2] = msub(kn.y, Xw.x, kn.x * Xw.y);
  J[3] = kn.x; J[4] = kn.y; J[5] = kn.z;                               // :83-85
  return true;
}

// 4x4 column-major product, include/vulcan/matrix.h:299-317
void mat4_mul(const float* A, const float* B, float* out) {
  float r[16];
  for (int p = 0; p < 4; ++p)
    for (int m = 0; m < 4; ++m) {
      float acc = 0;
      for (int n = 0; n < 4; ++n) acc += A[n * 4 + m] * B[p * 4 + n];
      r[p * 4 + m] = acc;
    }
  std::memcpy(out, r, sizeof(r));
}

// H x = -g by LDL^T in double; H packed lower triangle (src/depth_tracker.cu:214-223,
// src/tracker.cpp:133-147 solve it with Eigen::LDLT<float>, absent here)
void track_solve(const double* packed, const double* g, int n, float* x) {
  double L[6][6], D[6], y[6];
  for (int r = 0; r < n; ++r)
    for (int c = 0; c <= r; ++c) L[r][c] = packed[r * (r + 1) / 2 + c];
  for (int j = 0; j < n; ++j) {
    double d = L[j][j];
    for (int k = 0; k < j; ++k) d -= L[j][k] * L[j][k] * D[k];
    D[j] = d;
    const double inv = std::fabs(d) > 1e-30 ? 1.0 / d : 0.0;
    for (int i = j + 1; i < n; ++i) {
      double s = L[i][j];
      for (int k = 0; k < j; ++k) s -= L[i][k] * L[j][k] * D[k];
      L[i][j] = s * inv;
    }
  }
  for (int i = 0; i < n; ++i) {
    double s = -g[i];
    for (int k = 0; k < i; ++k) s -= L[i][k] * y[k];
    y[i] = s;
  }
  for (int i = 0; i < n; ++i) y[i] = std::fabs(D[i]) > 1e-30 ? y[i] / D[i] : 0.0;
  for (int i = n - 1; i >= 0; --i) {
    double s = y[i];
    for (int k = i + 1; k < n; ++k) s -= L[k][i] * y[k];
    y[i] = s;
  }
  for (int i = 0; i < 6; ++i) x[i] = i < n ? (float)y[i] : 0.0f;
}

// DepthTracker::ApplyUpdate, src/depth_tracker.cpp:23-88.  T = matrix + inverse (32 floats).
void track_apply_update(const float* u, float* T) {
  float Tinc[16];
  for (int i = 0; i < 16; ++i) Tinc[i] = (i % 5 == 0) ? 1.0f : 0.0f;
#define AT(m, r, c) (m)[(c) * 4 + (r)]
  AT(Tinc, 0, 1) = -u[2]; AT(Tinc, 0, 2) = +u[1]; AT(Tinc, 0, 3) = +u[3];
  AT(Tinc, 1, 0) = +u[2]; AT(Tinc, 1, 2) = +u[0]; AT(Tinc, 1, 3) = +u[4];   // :38-41 (the sign of (1,2) is the reference's)
  AT(Tinc, 2, 0) = -u[1]; AT(Tinc, 2, 1) = +u[0]; AT(Tinc, 2, 3) = +u[5];
  float Mx[16];
  mat4_mul(Tinc, T, Mx);                                                    // :53
  V3 X = {AT(Mx, 0, 0), AT(Mx, 1, 0), AT(Mx, 2, 0)};
  V3 Y = {AT(Mx, 0, 1), AT(Mx, 1, 1), AT(Mx, 2, 1)};
  auto plain_dot = [](const V3& a, const V3& b) { float r = 0; r += a.x * b.x; r += a.y * b.y; r += a.z * b.z; return r; };
  auto norm = [&](V3& a) { const float inv = 1.0f / std::sqrt(plain_dot(a, a)); a.x *= inv; a.y *= inv; a.z *= inv; };
  auto cross = [](const V3& a, const V3& b) {
    V3 o = {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
    return o;
  };
  norm(X); norm(Y);                                                         // :59-60
  const V3 Z = cross(X, Y);                                                 // :61
  Y = cross(Z, X);                                                          // :62
  float R4[16], Rt4[16], Tt[16], Tti[16];
  for (int i = 0; i < 16; ++i) R4[i] = Rt4[i] = Tt[i] = Tti[i] = (i % 5 == 0) ? 1.0f : 0.0f;
  const float xs[3] = {X.x, X.y, X.z}, ys[3] = {Y.x, Y.y, Y.z}, zs[3] = {Z.x, Z.y, Z.z};
  for (int r = 0; r < 3; ++r) {
    AT(R4, r, 0) = xs[r]; AT(R4, r, 1) = ys[r]; AT(R4, r, 2) = zs[r];
    AT(Rt4, 0, r) = xs[r]; AT(Rt4, 1, r) = ys[r]; AT(Rt4, 2, r) = zs[r];
    AT(Tt, r, 3) = AT(Mx, r, 3);
    AT(Tti, r, 3) = -AT(Mx, r, 3);
  }
#undef AT
  mat4_mul(Tt, R4, T);            // Translate(t) * Rotate(R), include/vulcan/transform.h:60-64
  mat4_mul(Rt4, Tti, T + 16);
}

void track_system(const float* T, const float* Tk, const float* key_depth, const float* key_normal, const Proj& kP,
                  int kw, int kh, const float* depth, const float* normal, const Proj& P, int w, int h,
                  int translation, double* H, double* g) {
  const int n = translation ? 6 : 3;
  for (int i = 0; i < 21; ++i) H[i] = 0;
  for (int i = 0; i < 6; ++i) g[i] = 0;
  for (int y = 0; y < h; ++y)
    for (int x = 0; x < w; ++x) {
      float r = 0, J[6];
      if (!track_evaluate(T, Tk, Tk + 16, key_depth, key_normal, kP, kw, kh, depth, normal, P, w, x, y, r, J)) continue;
      for (int a = 0, k = 0; a < n; ++a)
        for (int b = 0; b <= a; ++b, ++k) H[k] += (double)(J[a] * J[b]);   // float products, as the kernel forms them
      for (int a = 0; a < n; ++a) g[a] += (double)(J[a] * r);
    }
}

}  // namespace

extern "C" {

// DownsampleKernel, src/image.cu:101-165 (channels 1 or 3; dst is (w/2) x (h/2))
void vo_downsample(const float* src, int w, int h, int channels, int nearest, float* dst) {
  const int dw = w / 2, dh = h / 2;
  for (int y = 0; y < dh; ++y)
    for (int x = 0; x < dw; ++x)
      for (int c = 0; c < channels; ++c) {
        const float* s = src + ((size_t)(2 * y) * w + 2 * x) * channels + c;
        float v;
        if (nearest) v = s[0];
        else {
          v = 0;
          v += s[channels]; v += s[0]; v += s[(size_t)w * channels + channels]; v += s[(size_t)w * channels];
          v *= 0.25f;
        }
        dst[((size_t)y * dw + x) * channels + c] = v;
      }
}

// DepthTracker::ComputeResiduals / ComputeJacobian, src/depth_tracker.cu:101-144.  T, Tk: frame / keyframe
// depth_to_world matrix + inverse (32 floats each); proj4 fx fy cx cy.  Either output may be NULL.
void vo_track_pixels(const float* T, const float* Tk, const float* key_depth, const float* key_normal,
                     const float* kproj4, int kw, int kh, const float* depth, const float* normal, const float* proj4,
                     int w, int h, int translation, float* residuals, float* jacobian) {
  const Proj kP = {kproj4[0], kproj4[1], kproj4[2], kproj4[3]}, P = {proj4[0], proj4[1], proj4[2], proj4[3]};
  for (int y = 0; y < h; ++y)
    for (int x = 0; x < w; ++x) {
      float r = 0, J[6] = {0, 0, 0, 0, 0, 0};
      const bool ok = track_evaluate(T, Tk, Tk + 16, key_depth, key_normal, kP, kw, kh, depth, normal, P, w, x, y, r, J);
      const size_t i = (size_t)y * w + x;
      if (residuals) residuals[i] = ok ? r : 0.0f;
      if (jacobian)
        for (int k = 0; k < 6; ++k) jacobian[6 * i + k] = (ok && (translation || k < 3)) ? J[k] : 0.0f;
    }
}

// DepthTracker::ComputeSystem, src/depth_tracker.cu:146-274, accumulated in double
void vo_track_system(const float* T, const float* Tk, const float* key_depth, const float* key_normal,
                     const float* kproj4, int kw, int kh, const float* depth, const float* normal, const float* proj4,
                     int w, int h, int translation, double* hessian21, double* gradient6) {
  const Proj kP = {kproj4[0], kproj4[1], kproj4[2], kproj4[3]}, P = {proj4[0], proj4[1], proj4[2], proj4[3]};
  track_system(T, Tk, key_depth, key_normal, kP, kw, kh, depth, normal, P, w, h, translation, hessian21, gradient6);
}

// Tracker::Track, src/tracker.cpp:57-67,112-153 with DepthTracker::ApplyUpdate.  T (32 floats) is updated in
// place; returns the number of systems solved.
int vo_track_depth(float* T, const float* Tk, const float* key_depth, const float* key_normal, const float* kproj4,
                   int kw, int kh, const float* depth, const float* normal, const float* proj4, int w, int h,
                   int translation, int max_iterations) {
  const Proj kP = {kproj4[0], kproj4[1], kproj4[2], kproj4[3]}, P = {proj4[0], proj4[1], proj4[2], proj4[3]};
  int iteration = 0, solved = 0;
  while (iteration < max_iterations) {
    double H[21], g[6];
    track_system(T, Tk, key_depth, key_normal, kP, kw, kh, depth, normal, P, w, h, translation, H, g);
    float x[6];
    track_solve(H, g, translation ? 6 : 3, x);
    track_apply_update(x, T);
    ++solved;
    float nn = 0;
    for (int i = 0; i < 6; ++i) nn += x[i] * x[i];
    if (std::sqrt(nn) < 1e-6f) iteration = max_iterations;
    ++iteration;
  }
  return solved;
}

// DepthTracker::ApplyUpdate alone (for the host-logic tests)
void vo_track_apply_update(const float* update6, float* T) { track_apply_update(update6, T); }

}  // extern "C"

// ---- surface points (first slice) -------------------------------------------
// The reference's extractor is unfinished (src/extractor.cu:386 emits edge midpoints of one block, the
// index / face kernels at :391-413 are empty), so this restates the DEFINITION used by
// vulcan_b200/csrc/vk_extract.cu, not reference code: zero crossings of the TSDF on the three lattice
// edges each voxel owns (+x, +y, +z neighbour; edges 0, 3, 8 of its cube, src/extractor.cu:52-117), both
// end points observed (distance_weight > 0) and of different sign (`distance > 0`, src/extractor.cu:199),
// position and colour interpolated linearly.  "Parity unpinned" with respect to the reference.
extern "C" long long vo_extract_points(void* h, float* points, float* colors, long long capacity) {
  Volume* v = (Volume*)h;
  const float l = v->voxel_length, L = 8 * v->voxel_length;
  long long n = 0;
  Fetch f = {0, 0};
  for (int ei = 0; ei < v->B_; ++ei) {
    const HashEntry& e = v->entries[ei];
    if (e.data == -1) continue;
    const int bx = e.block[0], by = e.block[1], bz = e.block[2];
    for (int t = 0; t < 512; ++t) {
      const int vx = t & 7, vy = (t >> 3) & 7, vz = t >> 6;
      const Voxel& a = v->voxels[(long long)e.data * 512 + t];
      if (a.distance_weight <= 0) continue;
      const float base[3] = {(float)bx * L + ((float)vx + 0.5f) * l, (float)by * L + ((float)vy + 0.5f) * l,
                             (float)bz * L + ((float)vz + 0.5f) * l};
      for (int ax = 0; ax < 3; ++ax) {
        const int nx = vx + (ax == 0), ny = vy + (ax == 1), nz = vz + (ax == 2);
        int data = e.data;
        if (nx > 7 || ny > 7 || nz > 7) {
          const int cx = bx + (nx >> 3), cy = by + (ny >> 3), cz = bz + (nz >> 3);
          if (cx > 32767 || cy > 32767 || cz > 32767) continue;
          data = find_data(v, cx, cy, cz, f);
          if (data == -1) continue;
        }
        const Voxel& b = v->voxels[(long long)data * 512 + (((nz & 7) << 6) | ((ny & 7) << 3) | (nx & 7))];
        if (b.distance_weight <= 0 || ((a.distance > 0) == (b.distance > 0))) continue;
        const float tt = a.distance / (a.distance - b.distance);
        if (n < capacity) {
          float* o = points + 3 * n;
          o[0] = base[0]; o[1] = base[1]; o[2] = base[2];
          o[ax] = base[ax] + tt * l;
          if (colors)
            for (int c = 0; c < 3; ++c) colors[3 * n + c] = a.color[c] + tt * (b.color[c] - a.color[c]);
        }
        ++n;
      }
    }
  }
  return n;
}

// ---- triangles over the vertices of vo_extract_points (no finished reference counterpart: the reference's
// ExtractVertexIndicesKernel / ExtractFacesKernel are empty, src/extractor.cu:392-413).  Restates the definition
// of include/vulcan_b200.h (vk_extract_mesh): one marching-cubes cube per voxel (corner c at offset
// (c & 1, (c >> 1) & 1, c >> 2), across block faces), meshed when all eight corners are observed; case bit c set
// where !(distance > 0); triangles from the generated table (tools/gen_mc_tables.py), a triangle corner being
// the vertex on the lattice edge (owner voxel, axis).  "Parity unpinned" with respect to the reference. --------
#include "../vulcan_b200/csrc/vk_mc_tables.h"
#include <unordered_map>

extern "C" long long vo_extract_mesh(void* h, float* points, float* colors, long long point_capacity, int* faces,
                                     long long face_capacity, long long* face_count) {
  Volume* v = (Volume*)h;
  const float l = v->voxel_length, L = 8 * v->voxel_length;
  Fetch f = {0, 0};
  std::unordered_map<long long, long long> vertex;   // (voxel-block id * 512 + voxel) * 3 + axis -> vertex index
  long long n = 0;
  // vertices: exactly the loop of vo_extract_points, remembering where each one went
  for (int ei = 0; ei < v->B_; ++ei) {
    const HashEntry& e = v->entries[ei];
    if (e.data == -1) continue;
    const int bx = e.block[0], by = e.block[1], bz = e.block[2];
    for (int t = 0; t < 512; ++t) {
      const int vx = t & 7, vy = (t >> 3) & 7, vz = t >> 6;
      const Voxel& a = v->voxels[(long long)e.data * 512 + t];
      if (a.distance_weight <= 0) continue;
      const float base[3] = {(float)bx * L + ((float)vx + 0.5f) * l, (float)by * L + ((float)vy + 0.5f) * l,
                             (float)bz * L + ((float)vz + 0.5f) * l};
      for (int ax = 0; ax < 3; ++ax) {
        const int nx = vx + (ax == 0), ny = vy + (ax == 1), nz = vz + (ax == 2);
        int data = e.data;
        if (nx > 7 || ny > 7 || nz > 7) {
          const int cx = bx + (nx >> 3), cy = by + (ny >> 3), cz = bz + (nz >> 3);
          if (cx > 32767 || cy > 32767 || cz > 32767) continue;
          data = find_data(v, cx, cy, cz, f);
          if (data == -1) continue;
        }
        const Voxel& b = v->voxels[(long long)data * 512 + (((nz & 7) << 6) | ((ny & 7) << 3) | (nx & 7))];
        if (b.distance_weight <= 0 || ((a.distance > 0) == (b.distance > 0))) continue;
        const float tt = a.distance / (a.distance - b.distance);
        if (n < point_capacity) {
          float* o = points + 3 * n;
          o[0] = base[0]; o[1] = base[1]; o[2] = base[2];
          o[ax] = base[ax] + tt * l;
          if (colors)
            for (int c = 0; c < 3; ++c) colors[3 * n + c] = a.color[c] + tt * (b.color[c] - a.color[c]);
        }
        vertex[((long long)e.data * 512 + t) * 3 + ax] = n;
        ++n;
      }
    }
  }
  // faces
  long long nf = 0;
  for (int ei = 0; ei < v->B_; ++ei) {
    const HashEntry& e = v->entries[ei];
    if (e.data == -1) continue;
    const int bx = e.block[0], by = e.block[1], bz = e.block[2];
    int nb[8];
    for (int k = 0; k < 8; ++k) {
      const int cx = bx + (k & 1), cy = by + ((k >> 1) & 1), cz = bz + (k >> 2);
      nb[k] = (k == 0) ? e.data : ((cx > 32767 || cy > 32767 || cz > 32767) ? -1 : find_data(v, cx, cy, cz, f));
    }
    for (int t = 0; t < 512; ++t) {
      const int vx = t & 7, vy = (t >> 3) & 7, vz = t >> 6;
      unsigned cube = 0;
      bool ok = true;
      for (int c = 0; c < 8 && ok; ++c) {
        const int x = vx + (c & 1), y = vy + ((c >> 1) & 1), z = vz + (c >> 2);
        const int data = nb[(x >> 3) | ((y >> 3) << 1) | ((z >> 3) << 2)];
        if (data < 0) { ok = false; break; }
        const Voxel& q = v->voxels[(long long)data * 512 + (((z & 7) << 6) | ((y & 7) << 3) | (x & 7))];
        if (q.distance_weight <= 0) ok = false;
        if (!(q.distance > 0)) cube |= 1u << c;
      }
      if (!ok) continue;
      for (int k = 0; k < 5 && kMcTriTable[cube][3 * k] >= 0; ++k) {
        int ids[3];
        for (int j = 0; j < 3; ++j) {
          const int edge = kMcTriTable[cube][3 * k + j];
          const int corner = kMcEdgeCorner[edge], axis = kMcEdgeAxis[edge];
          const int x = vx + (corner & 1), y = vy + ((corner >> 1) & 1), z = vz + (corner >> 2);
          const int data = nb[(x >> 3) | ((y >> 3) << 1) | ((z >> 3) << 2)];
          const long long key = ((long long)data * 512 + (((z & 7) << 6) | ((y & 7) << 3) | (x & 7))) * 3 + axis;
          const auto it = vertex.find(key);
          ids[j] = (it == vertex.end()) ? -1 : (int)it->second;
        }
        if (nf < face_capacity) { faces[3 * nf] = ids[0]; faces[3 * nf + 1] = ids[1]; faces[3 * nf + 2] = ids[2]; }
        ++nf;
      }
    }
  }
  if (face_count) *face_count = nf;
  return n;
}
