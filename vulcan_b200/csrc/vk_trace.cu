// vk_trace.cu — Tracer::Trace rebuilt for sm_100a.
//
// Reference (src/tracer.cpp:41-47): ComputePatches (20-barrier scan + blocking D2H read
// of the patch count) -> FillKernel -> ComputeBounds (float CAS loops) -> ComputePoints
// (16x16 tiles, scalar loads, a full hash walk per trilinear tap on block faces) ->
// ComputeNormals.
//
// Here: the visible blocks are rasterised straight into the 80x60 bounds grid with
// native integer atomicMin/atomicMax (positive floats order like ints), no patch list
// and no host read-back; the ray marcher maps a warp to an 8x4 pixel tile, loads hash
// entries and voxels with 128-bit / read-only loads and memoises the last block lookup
// per lane; the step-wise entry points of include/vulcan/tracer.cuh are kept as well.
#include "vk_math.cuh"
#include "vk_launch.h"
#include <cfloat>

namespace vk {

// ---- ray bounds -------------------------------------------------------------------

struct BlockBox {
  short bminx, bminy, bmaxx, bmaxy;
  float zlo, zhi;
  int gx, gy;
};

// float atomic min/max without a CAS loop (include/vulcan/util.cuh:20-50 uses one):
// non-negative floats order like signed ints, negative floats like reversed unsigned ints
__device__ __forceinline__ void atomic_min_float(float* addr, float value) {
  if (value >= 0) atomicMin(reinterpret_cast<int*>(addr), __float_as_int(value));
  else atomicMax(reinterpret_cast<unsigned int*>(addr), __float_as_uint(value));
}
__device__ __forceinline__ void atomic_max_float(float* addr, float value) {
  if (value >= 0) atomicMax(reinterpret_cast<int*>(addr), __float_as_int(value));
  else atomicMin(reinterpret_cast<unsigned int*>(addr), __float_as_uint(value));
}

template <typename T>
__device__ __forceinline__ T vclamp(T v, T lo, T hi) {
  // vulcan::clamp (include/vulcan/math.h:23-28): min(vmax, max(vmin, v))
  const T a = (lo > v) ? lo : v;
  return (a < hi) ? a : hi;
}

// body of ComputePatchesKernel (src/tracer.cu:13-66) for one block
__device__ __forceinline__ BlockBox block_box(const int4& ent, const Mat34& Tcw, const vkProjection& proj,
                                              float block_length, float min_depth, float max_depth,
                                              int image_w, int image_h, int bw, int bh) {
  const int ox = (int)(int16_t)(ent.x & 0xffff);
  const int oy = (int)(int16_t)((uint32_t)ent.x >> 16);
  const int oz = (int)(int16_t)(ent.y & 0xffff);
  short bmaxx = -1, bmaxy = -1;
  short bminx = (short)bw, bminy = (short)bh;
  float zlo = +FLT_MAX, zhi = -FLT_MAX;
  for (int z = 0; z <= 1; ++z) {
    const float Z = mul(block_length, (float)(z + oz));
    for (int y = 0; y <= 1; ++y) {
      const float Y = mul(block_length, (float)(y + oy));
      for (int x = 0; x <= 1; ++x) {
        const float X = mul(block_length, (float)(x + ox));
        const Vec3 Xc = xform(Tcw, X, Y, Z, 1.0f);
        float u, v;
        project(proj, Xc, u, v);
        u = fdiv(mul((float)bw, u), (float)image_w);
        v = fdiv(mul((float)bh, v), (float)image_h);
        const short fu = (short)floorf(u), fv = (short)floorf(v);
        const short cu = (short)ceilf(u), cv = (short)ceilf(v);
        bminx = vclamp<short>((fu < bminx) ? fu : bminx, 0, (short)(bw - 1));
        bminy = vclamp<short>((fv < bminy) ? fv : bminy, 0, (short)(bh - 1));
        bmaxx = vclamp<short>((cu > bmaxx) ? cu : bmaxx, 0, (short)(bw - 1));
        bmaxy = vclamp<short>((cv > bmaxy) ? cv : bmaxy, 0, (short)(bh - 1));
        zlo = vclamp<float>(vmin(Xc.z, zlo), min_depth, max_depth);
        zhi = vclamp<float>(vmax(Xc.z, zhi), min_depth, max_depth);
      }
    }
  }
  BlockBox b;
  b.bminx = bminx; b.bminy = bminy; b.bmaxx = bmaxx; b.bmaxy = bmaxy;
  b.zlo = zlo; b.zhi = zhi;
  const int rx = bmaxx - bminx, ry = bmaxy - bminy;
  b.gx = (rx + 15) / 16;   // Q12: rx == 0 -> no patch at all
  b.gy = (ry + 15) / 16;
  if (!(zhi > zlo)) { b.gx = 0; b.gy = 0; }
  return b;
}

struct BoundsParams {
  const int32_t* visible;
  const int32_t* count;
  const vkHashEntry* entries;
  float Tcw[12];
  vkProjection proj;
  float block_length, min_depth, max_depth;
  int image_w, image_h, bw, bh;
  float* bounds;
};

// K10 + K11 fused: the union of a block's patches is the cell rectangle
// [bmin, min(bmax, bmin + 16*g - 1)] (src/tracer.cu:67-86), merged with integer atomics.
__global__ void __launch_bounds__(256) vk_bounds_kernel(const BoundsParams p) {
  const int n = *p.count;
  Mat34 Tcw;
#pragma unroll
  for (int r = 0; r < 3; ++r)
#pragma unroll
    for (int c = 0; c < 4; ++c) Tcw.r[r][c] = p.Tcw[r * 4 + c];
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const int e = __ldg(p.visible + i);
    const int4 ent = ld_entry(p.entries + e);
    const BlockBox b = block_box(ent, Tcw, p.proj, p.block_length, p.min_depth, p.max_depth, p.image_w,
                                 p.image_h, p.bw, p.bh);
    if (b.gx * b.gy == 0) continue;
    const int x1 = min((int)b.bmaxx, (int)b.bminx + 16 * b.gx - 1);
    const int y1 = min((int)b.bmaxy, (int)b.bminy + 16 * b.gy - 1);
    for (int y = b.bminy; y <= y1; ++y)
      for (int x = b.bminx; x <= x1; ++x) {
        float* cell = p.bounds + 2 * (y * p.bw + x);
        atomic_min_float(cell, b.zlo);
        atomic_max_float(cell + 1, b.zhi);
      }
  }
}

__global__ void vk_reset_bounds_kernel(float2* bounds, int count) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < count) bounds[i] = make_float2(+FLT_MAX, -FLT_MAX);
}

// ComputePatchesKernel (src/tracer.cu:13-87) with a warp-aggregated append in place of
// the shared-memory scan; patch order is unspecified in the reference too (Q7-like).
__global__ void __launch_bounds__(256) vk_patches_kernel(const int* indices, const vkHashEntry* entries,
                                                         const BoundsParams p, int block_count,
                                                         vkPatch* patches, int* patch_count) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int lane = threadIdx.x & 31;
  Mat34 Tcw;
#pragma unroll
  for (int r = 0; r < 3; ++r)
#pragma unroll
    for (int c = 0; c < 4; ++c) Tcw.r[r][c] = p.Tcw[r * 4 + c];
  BlockBox b;
  b.gx = b.gy = 0;
  if (i < block_count) {
    const int4 ent = ld_entry(entries + indices[i]);
    b = block_box(ent, Tcw, p.proj, p.block_length, p.min_depth, p.max_depth, p.image_w, p.image_h, p.bw, p.bh);
  }
  const int cnt = b.gx * b.gy;
  // inclusive warp scan of cnt
  int incl = cnt;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const int o = __shfl_up_sync(0xffffffffu, incl, d);
    if (lane >= d) incl += o;
  }
  const int total = __shfl_sync(0xffffffffu, incl, 31);
  int base = 0;
  if (lane == 0 && total > 0) base = atomicAdd(patch_count, total);
  base = __shfl_sync(0xffffffffu, base, 0);
  const int offset = base + incl - cnt;
  for (int ii = 0; ii < b.gy; ++ii)
    for (int j = 0; j < b.gx; ++j) {
      vkPatch pt;
      pt.origin[0] = (short)(b.bminx + 16 * j);
      pt.origin[1] = (short)(b.bminy + 16 * ii);
      pt.size[0] = (short)min(16, b.bmaxx - pt.origin[0] + 1);
      pt.size[1] = (short)min(16, b.bmaxy - pt.origin[1] + 1);
      pt.bounds[0] = b.zlo;
      pt.bounds[1] = b.zhi;
      patches[offset + ii * b.gx + j] = pt;
    }
}

// ComputeBoundsKernel (src/tracer.cu:89-112)
__global__ void __launch_bounds__(256) vk_patch_bounds_kernel(const vkPatch* patches, float* bounds, int bw,
                                                              int patch_count) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= patch_count) return;
  const vkPatch pt = patches[i];
  for (int a = 0; a < pt.size[1]; ++a)
    for (int b = 0; b < pt.size[0]; ++b) {
      float* cell = bounds + 2 * ((pt.origin[1] + a) * bw + pt.origin[0] + b);
      atomic_min_float(cell, pt.bounds[0]);
      atomic_max_float(cell + 1, pt.bounds[1]);
    }
}

}  // namespace vk

// ---- ray marcher ------------------------------------------------------------------

namespace vk {

struct TraceParams {
  const vkHashEntry* entries;
  const vkVoxel* voxels;
  const float* bounds;
  uint32_t K;
  float block_length, voxel_length, trunc;
  float Twc[12], Tcw[12];
  vkProjection proj;
  float* depths;
  float* colors;
  unsigned long long* keys;
  int w, h, bw, bh;
  int shard_rank, shard_count, shard_shift;
};

// memo of the last hash lookup of this lane; the table is read-only during a trace so
// replaying a lookup from the memo is exact
struct Lookup {
  unsigned long long key;
  int data;
};

// `while (entry.block != b && entry.HasNext()) entry = entries[entry.next];`
// `entry.block == b && entry.IsAllocated()`  (src/tracer.cu:363-371, :151-175)
__device__ __forceinline__ int find_data(const TraceParams& p, Lookup& memo, int bx, int by, int bz) {
  const unsigned long long key = pack_block(bx, by, bz);
  if (key == memo.key) return memo.data;
  int4 e = ld_entry(p.entries + block_hash(bx, by, bz, p.K));
  while (entry_block(e) != key && e.w != -1) e = ld_entry(p.entries + e.w);
  const int data = (entry_block(e) == key) ? e.z : -1;
  memo.key = key;
  memo.data = data;
  return data;
}

struct Tap {
  float distance;
  float r, g, b;
  int cw;
};

__device__ __forceinline__ Tap load_tap(const TraceParams& p, int data, int voxel_offset) {
  // Voxel is 20 bytes, 4-byte aligned: distance, rgb, (dw | cw << 16)
  const uint32_t* w = reinterpret_cast<const uint32_t*>(p.voxels) +
                      ((size_t)data * VK_BLOCK_VOXELS + (size_t)voxel_offset) * 5;
  Tap t;
  t.distance = __uint_as_float(__ldg(w));
  t.r = __uint_as_float(__ldg(w + 1));
  t.g = __uint_as_float(__ldg(w + 2));
  t.b = __uint_as_float(__ldg(w + 3));
  t.cw = (int)(int16_t)(__ldg(w + 4) >> 16);
  return t;
}

// GetVoxel (src/tracer.cu:114-188)
__device__ __forceinline__ Tap get_voxel(const TraceParams& p, Lookup& memo, int bx, int by, int bz, int vx,
                                         int vy, int vz) {
  if (vx < 0) { --bx; vx += 8; } else if (vx >= 8) { ++bx; vx -= 8; }
  if (vy < 0) { --by; vy += 8; } else if (vy >= 8) { ++by; vy -= 8; }
  if (vz < 0) { --bz; vz += 8; } else if (vz >= 8) { ++bz; vz -= 8; }
  const int data = find_data(p, memo, bx, by, bz);
  if (data != -1) return load_tap(p, data, vz * 64 + vy * 8 + vx);
  Tap t;  // Voxel::Empty()
  t.distance = 1.0f; t.r = t.g = t.b = 0.0f; t.cw = 0;
  return t;
}

// GetInterpolatedDistance (src/tracer.cu:190-315); Q15 colour weights, Q16 empty taps
__device__ __forceinline__ void interpolate(const TraceParams& p, Lookup& memo, int bx, int by, int bz, int data,
                                            const Vec3& pt, float& sdf, Vec3& color) {
  const float wx = fdiv(sub(pt.x, mul((float)bx, p.block_length)), p.voxel_length);
  const float wy = fdiv(sub(pt.y, mul((float)by, p.block_length)), p.voxel_length);
  const float wz = fdiv(sub(pt.z, mul((float)bz, p.block_length)), p.voxel_length);
  const int i0x = (int)floorf(sub(wx, 0.5f));
  const int i0y = (int)floorf(sub(wy, 0.5f));
  const int i0z = (int)floorf(sub(wz, 0.5f));
  Tap v[8];  // index = z*4 + y*2 + x
  if (i0x >= 0 && i0y >= 0 && i0z >= 0 && i0x < 7 && i0y < 7 && i0z < 7) {
#pragma unroll
    for (int c = 0; c < 8; ++c)
      v[c] = load_tap(p, data, (i0z + (c >> 2)) * 64 + (i0y + ((c >> 1) & 1)) * 8 + (i0x + (c & 1)));
  } else {
#pragma unroll
    for (int c = 0; c < 8; ++c)
      v[c] = get_voxel(p, memo, bx, by, bz, i0x + (c & 1), i0y + ((c >> 1) & 1), i0z + (c >> 2));
  }
  const float w1x = sub(wx, add((float)i0x, 0.5f));
  const float w1y = sub(wy, add((float)i0y, 0.5f));
  const float w1z = sub(wz, add((float)i0z, 0.5f));
  const float w0x = sub(1.0f, w1x), w0y = sub(1.0f, w1y), w0z = sub(1.0f, w1z);

  const float n00 = add(mul(v[0].distance, w0x), mul(v[1].distance, w1x));
  const float n01 = add(mul(v[2].distance, w0x), mul(v[3].distance, w1x));
  const float n10 = add(mul(v[4].distance, w0x), mul(v[5].distance, w1x));
  const float n11 = add(mul(v[6].distance, w0x), mul(v[7].distance, w1x));
  const float n0 = add(mul(n00, w0y), mul(n01, w1y));
  const float n1 = add(mul(n10, w0y), mul(n11, w1y));

  float total = 0.0f;
  float cr = 0.0f, cg = 0.0f, cb = 0.0f;
#pragma unroll
  for (int c = 0; c < 8; ++c) {
    const float wz_ = (c >> 2) ? w1z : w0z;
    const float wy_ = ((c >> 1) & 1) ? w1y : w0y;
    const float wx_ = (c & 1) ? w1x : w0x;
    const float prod = mul(mul(mul(wz_, wy_), wx_), (v[c].cw > 0) ? 1.0f : 0.0f);
    const float wgt = (prod != 0.0f) ? 1.0f : 0.0f;  // `a * b * c * (cw > 0) ? 1 : 0`
    total = add(total, wgt);
    cr = add(cr, mul(v[c].r, wgt));
    cg = add(cg, mul(v[c].g, wgt));
    cb = add(cb, mul(v[c].b, wgt));
  }
  if (total > 0) {
    const float inv = fdiv(1.0f, total);
    cr = mul(cr, inv); cg = mul(cg, inv); cb = mul(cb, inv);
  }
  color.x = cr; color.y = cg; color.z = cb;
  sdf = add(mul(n0, w0z), mul(n1, w1z));
}

// ComputePointsKernel (src/tracer.cu:317-451)
template <bool KEYS>
__global__ void __launch_bounds__(256) vk_raycast_kernel(const TraceParams p) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int x = blockIdx.x * 32 + (warp & 3) * 8 + (lane & 7);
  const int y = blockIdx.y * 8 + (warp >> 2) * 4 + (lane >> 3);
  if (x >= p.w || y >= p.h) return;

  const int px = p.bw * x / p.w;
  const int py = p.bh * y / p.h;
  const float2 bound = __ldg(reinterpret_cast<const float2*>(p.bounds) + py * p.bw + px);

  float depth = 0, final_depth = 0;
  Vec3 color = {0.f, 0.f, 0.f};
  bool owned_hit = false;

  if (bound.x < bound.y) {
    Mat34 Twc, Tcw;
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
      for (int c = 0; c < 4; ++c) { Twc.r[r][c] = p.Twc[r * 4 + c]; Tcw.r[r][c] = p.Tcw[r * 4 + c]; }
    const Vec3 un = unproject(p.proj, (float)x + 0.5f, (float)y + 0.5f);
    const Vec3 Xc = {mul(un.x, bound.x), mul(un.y, bound.x), mul(un.z, bound.x)};
    Vec3 pt = xform(Twc, Xc.x, Xc.y, Xc.z, 1.0f);
    const Vec3 dir = normalized(xform(Twc, Xc.x, Xc.y, Xc.z, 0.0f));
    Lookup memo;
    memo.key = ~0ull;
    memo.data = -1;
    depth = bound.x;
    int iters = 0;
    do {
      const int bx = (int)floorf(fdiv(pt.x, p.block_length));
      const int by = (int)floorf(fdiv(pt.y, p.block_length));
      const int bz = (int)floorf(fdiv(pt.z, p.block_length));
      const int data = find_data(p, memo, bx, by, bz);
      if (data != -1) {
        const float wx = fdiv(sub(pt.x, mul((float)bx, p.block_length)), p.voxel_length);
        const float wy = fdiv(sub(pt.y, mul((float)by, p.block_length)), p.voxel_length);
        const float wz = fdiv(sub(pt.z, mul((float)bz, p.block_length)), p.voxel_length);
        const int vx = (int)wx, vy = (int)wy, vz = (int)wz;  // Q22: may be 8 on a face
        const uint32_t* vw = reinterpret_cast<const uint32_t*>(p.voxels) +
                             ((size_t)data * VK_BLOCK_VOXELS + (size_t)(vz * 64 + vy * 8 + vx)) * 5;
        float sdf = __uint_as_float(__ldg(vw));
        if (sdf <= 0.1f && sdf >= -0.5f) interpolate(p, memo, bx, by, bz, data, pt, sdf, color);
        if (sdf <= 0.0f) {
          if (KEYS) owned_hit = (shard_owner(bx, by, bz, p.shard_shift, p.shard_count) == p.shard_rank);
          float s = mul(p.trunc, sdf);
          pt.x = add(pt.x, mul(dir.x, s)); pt.y = add(pt.y, mul(dir.y, s)); pt.z = add(pt.z, mul(dir.z, s));
          const int bx2 = (int)floorf(fdiv(pt.x, p.block_length));
          const int by2 = (int)floorf(fdiv(pt.y, p.block_length));
          const int bz2 = (int)floorf(fdiv(pt.z, p.block_length));
          const int data2 = find_data(p, memo, bx2, by2, bz2);
          if (data2 != -1) {
            interpolate(p, memo, bx2, by2, bz2, data2, pt, sdf, color);
            s = mul(p.trunc, sdf);
            pt.x = add(pt.x, mul(dir.x, s)); pt.y = add(pt.y, mul(dir.y, s)); pt.z = add(pt.z, mul(dir.z, s));
          }
          final_depth = xform(Tcw, pt.x, pt.y, pt.z, 1.0f).z;
          break;
        } else {
          const float s = vmax(p.voxel_length, mul(p.trunc, sdf));
          pt.x = add(pt.x, mul(dir.x, s)); pt.y = add(pt.y, mul(dir.y, s)); pt.z = add(pt.z, mul(dir.z, s));
        }
      } else {
        pt.x = add(pt.x, mul(dir.x, p.block_length));
        pt.y = add(pt.y, mul(dir.y, p.block_length));
        pt.z = add(pt.z, mul(dir.z, p.block_length));
      }
      depth = xform(Tcw, pt.x, pt.y, pt.z, 1.0f).z;
      if (++iters >= 500) { color.x = 1.f; color.y = 0.f; color.z = 0.f; break; }
    } while (depth < bound.y);
  }

  const int pixel = y * p.w + x;
  if (KEYS) {
    unsigned long long key = ((unsigned long long)0x7f800000u << 32);  // +inf, no payload
    if (owned_hit && final_depth > 0.0f) {
      const uint32_t r8 = (uint32_t)fminf(fmaxf(color.x * 255.0f + 0.5f, 0.f), 255.f);
      const uint32_t g8 = (uint32_t)fminf(fmaxf(color.y * 255.0f + 0.5f, 0.f), 255.f);
      const uint32_t b8 = (uint32_t)fminf(fmaxf(color.z * 255.0f + 0.5f, 0.f), 255.f);
      key = ((unsigned long long)__float_as_uint(final_depth) << 32) | ((unsigned long long)r8 << 24) |
            ((unsigned long long)g8 << 16) | ((unsigned long long)b8 << 8) | (unsigned long long)(p.shard_rank & 0xff);
    }
    p.keys[pixel] = key;
  } else {
    p.depths[pixel] = final_depth;
    p.colors[3 * (size_t)pixel + 0] = color.x;
    p.colors[3 * (size_t)pixel + 1] = color.y;
    p.colors[3 * (size_t)pixel + 2] = color.z;
  }
}

__global__ void __launch_bounds__(256) vk_resolve_keys_kernel(const unsigned long long* keys, int n, float* depths,
                                                              float* colors) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const unsigned long long k = keys[i];
  const uint32_t hi = (uint32_t)(k >> 32);
  const bool hit = hi < 0x7f800000u;
  depths[i] = hit ? __uint_as_float(hi) : 0.0f;
  const float s = 1.0f / 255.0f;
  colors[3 * (size_t)i + 0] = hit ? (float)((k >> 24) & 0xff) * s : 0.f;
  colors[3 * (size_t)i + 1] = hit ? (float)((k >> 16) & 0xff) * s : 0.f;
  colors[3 * (size_t)i + 2] = hit ? (float)((k >> 8) & 0xff) * s : 0.f;
}

}  // namespace vk

// ---- image-space kernels ------------------------------------------------------------

namespace vk {

// ComputeNormalsKernel<16> (src/frame.cu:7-122): 32x8 tile + 2 px halo, zero outside
__global__ void __launch_bounds__(256) vk_normals_kernel(const float* depths, const vkProjection proj,
                                                         float* normals, int w, int h) {
  constexpr int TW = 32, TH = 8, PAD = 2, SW = TW + 2 * PAD, SH = TH + 2 * PAD;
  __shared__ float tile[SH * SW];
  const int x0 = blockIdx.x * TW, y0 = blockIdx.y * TH;
  for (int s = threadIdx.x; s < SH * SW; s += TW * TH) {
    const int gx = x0 - PAD + (s % SW), gy = y0 - PAD + (s / SW);
    float d = 0;
    if (gx >= 0 && gx < w && gy >= 0 && gy < h) d = __ldg(depths + gy * w + gx);
    tile[s] = d;
  }
  __syncthreads();
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int x = x0 + tx, y = y0 + ty;
  if (x >= w || y >= h) return;
  const float depth = tile[(ty + PAD) * SW + tx + PAD];
  float nx = 0, ny = 0, nz = 0;
  if (depth > 0) {
    const float fx = (float)x + 0.5f, fy = (float)y + 0.5f;
    const Vec3 u0 = unproject(proj, fx, fy);
    const Vec3 z0 = {mul(u0.x, depth), mul(u0.y, depth), mul(u0.z, depth)};
    Vec3 q[4];
    const int ox[4] = {-PAD, +PAD, 0, 0}, oy[4] = {0, 0, -PAD, +PAD};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float d = tile[(ty + PAD + oy[k]) * SW + tx + PAD + ox[k]];
      if (d == 0) {
        q[k] = z0;
      } else {
        const Vec3 un = unproject(proj, (float)(x + ox[k]) + 0.5f, (float)(y + oy[k]) + 0.5f);
        q[k].x = mul(un.x, d); q[k].y = mul(un.y, d); q[k].z = mul(un.z, d);
      }
    }
    const Vec3 dx = {sub(q[0].x, q[1].x), sub(q[0].y, q[1].y), sub(q[0].z, q[1].z)};
    const Vec3 dy = {sub(q[2].x, q[3].x), sub(q[2].y, q[3].y), sub(q[2].z, q[3].z)};
    if (dot3(dx, dx) > 0 && dot3(dy, dy) > 0) {
      // dy.Cross(dx) (include/vulcan/matrix.h:174-183)
      Vec3 n;
      n.x = sub(mul(dy.y, dx.z), mul(dy.z, dx.y));
      n.y = sub(mul(dy.z, dx.x), mul(dy.x, dx.z));
      n.z = sub(mul(dy.x, dx.y), mul(dy.y, dx.x));
      n = normalized(n);
      nx = n.x; ny = n.y; nz = n.z;
    }
  }
  const size_t o = 3 * ((size_t)y * w + x);
  normals[o] = nx; normals[o + 1] = ny; normals[o + 2] = nz;
}

// FilterDepthsKernel<16> (src/frame.cu:124-181): 7x7 bilateral, sigma_r^2 = 0.0004
__global__ void __launch_bounds__(256) vk_filter_depths_kernel(int w, int h, const float* src, float* dst) {
  constexpr int TW = 32, TH = 8, PAD = 3, SW = TW + 2 * PAD, SH = TH + 2 * PAD;
  __shared__ float tile[SH * SW];
  const int x0 = blockIdx.x * TW, y0 = blockIdx.y * TH;
  for (int s = threadIdx.x; s < SH * SW; s += TW * TH) {
    const int gx = x0 - PAD + (s % SW), gy = y0 - PAD + (s / SW);
    float d = 0;
    if (gx >= 0 && gx < w && gy >= 0 && gy < h) d = __ldg(src + gy * w + gx);
    tile[s] = d;
  }
  __syncthreads();
  const int tx = (threadIdx.x & 31) + PAD, ty = (threadIdx.x >> 5) + PAD;
  const int x = x0 + tx - PAD, y = y0 + ty - PAD;
  if (x >= w || y >= h) return;
  const float d0 = tile[ty * SW + tx];
  float dn = 0, wsum = 0;
  for (int i = -PAD; i <= PAD; ++i)
    for (int j = -PAD; j <= PAD; ++j) {
      const float dk = tile[(ty + i) * SW + tx + j];
      const float delta = sub(d0, dk);
      // Vector2f(i, j).SquaredNorm() / (pad * pad): float / int -> true division
      const float sq = add(add(0.0f, mul((float)i, (float)i)), mul((float)j, (float)j));
      const float wr = expf(fdiv(-sq, (float)(PAD * PAD)));
      const float ws = expf(fdiv(-mul(delta, delta), 0.0004f));
      const float ww = mul(wr, ws);
      dn = add(dn, mul(ww, dk));
      wsum = add(wsum, ww);
    }
  dst[y * w + x] = fdiv(dn, wsum);
}

}  // namespace vk

// ---- C ABI --------------------------------------------------------------------------

using namespace vk;

static void rows34(const float* colmajor, float* out12) {
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 4; ++c) out12[r * 4 + c] = colmajor[c * 4 + r];
}

extern "C" int vk_reset_bounds(float* bounds, int count, void* stream) {
  if (!bounds || count < 0) return VK_E_INVALID;
  if (count == 0) return 0;
  vk_reset_bounds_kernel<<<(count + 255) / 256, 256, 0, (cudaStream_t)stream>>>(
      reinterpret_cast<float2*>(bounds), count);
  VK_COUNT_LAUNCH();
  return (int)cudaGetLastError();
}

static void fill_bounds_params(BoundsParams& b, const vkTransform* Tcw_is_inverse_of, bool use_inverse,
                               const vkProjection* proj, float block_length, float min_depth, float max_depth,
                               int image_w, int image_h, int bw, int bh, float* bounds) {
  rows34(use_inverse ? Tcw_is_inverse_of->inv : Tcw_is_inverse_of->m, b.Tcw);
  b.proj = *proj;
  b.block_length = block_length;
  b.min_depth = min_depth;
  b.max_depth = max_depth;
  b.image_w = image_w; b.image_h = image_h; b.bw = bw; b.bh = bh;
  b.bounds = bounds;
  b.visible = nullptr; b.count = nullptr; b.entries = nullptr;
}

extern "C" int vk_compute_patches(const int* indices, const vkHashEntry* entries, const vkTransform* Tcw,
                                  const vkProjection* proj, float block_length, float min_depth,
                                  float max_depth, int block_count, int image_w, int image_h, int bounds_w,
                                  int bounds_h, vkPatch* patches, int* patch_count, void* stream) {
  if (!indices || !entries || !Tcw || !proj || !patches || !patch_count || block_count < 0) return VK_E_INVALID;
  if (block_count == 0) return 0;
  BoundsParams b;
  fill_bounds_params(b, Tcw, false, proj, block_length, min_depth, max_depth, image_w, image_h, bounds_w,
                     bounds_h, nullptr);
  vk_patches_kernel<<<(block_count + 255) / 256, 256, 0, (cudaStream_t)stream>>>(indices, entries, b, block_count,
                                                                                patches, patch_count);
  VK_COUNT_LAUNCH();
  return (int)cudaGetLastError();
}

extern "C" int vk_compute_bounds(const vkPatch* patches, float* bounds, int bounds_w, int patch_count,
                                 void* stream) {
  if (!patches || !bounds || patch_count < 0) return VK_E_INVALID;
  if (patch_count == 0) return 0;
  vk_patch_bounds_kernel<<<(patch_count + 255) / 256, 256, 0, (cudaStream_t)stream>>>(patches, bounds, bounds_w,
                                                                                      patch_count);
  VK_COUNT_LAUNCH();
  return (int)cudaGetLastError();
}

static void fill_trace_params(TraceParams& t, const vkHashEntry* entries, const vkVoxel* voxels,
                              const float* bounds, int K, float block_length, float voxel_length, float trunc,
                              const vkTransform* Twc, const vkProjection* proj, int w, int h, int bw, int bh) {
  t.entries = entries;
  t.voxels = voxels;
  t.bounds = bounds;
  t.K = (uint32_t)K;
  t.block_length = block_length;
  t.voxel_length = voxel_length;
  t.trunc = trunc;
  rows34(Twc->m, t.Twc);
  rows34(Twc->inv, t.Tcw);
  t.proj = *proj;
  t.depths = nullptr; t.colors = nullptr; t.keys = nullptr;
  t.w = w; t.h = h; t.bw = bw; t.bh = bh;
  t.shard_rank = 0; t.shard_count = 1; t.shard_shift = 0;
}

extern "C" int vk_compute_points(const vkHashEntry* entries, const vkVoxel* voxels, const float* bounds,
                                 int block_count, float block_length, float voxel_length, float trunc_length,
                                 const vkTransform* Twc, const vkProjection* proj, float* depths, float* colors,
                                 int image_w, int image_h, int bounds_w, int bounds_h, void* stream) {
  if (!entries || !voxels || !bounds || !Twc || !proj || !depths || !colors || image_w <= 0 || image_h <= 0)
    return VK_E_INVALID;
  TraceParams t;
  fill_trace_params(t, entries, voxels, bounds, block_count, block_length, voxel_length, trunc_length, Twc, proj,
                    image_w, image_h, bounds_w, bounds_h);
  t.depths = depths;
  t.colors = colors;
  dim3 grid((image_w + 31) / 32, (image_h + 7) / 8);
  vk_raycast_kernel<false><<<grid, 256, 0, (cudaStream_t)stream>>>(t);
  VK_COUNT_LAUNCH();
  return (int)cudaGetLastError();
}

extern "C" int vk_compute_normals(const float* depths, const vkProjection* proj, float* normals, int image_w,
                                  int image_h, void* stream) {
  if (!depths || !proj || !normals || image_w <= 0 || image_h <= 0) return VK_E_INVALID;
  dim3 grid((image_w + 31) / 32, (image_h + 7) / 8);
  vk_normals_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(depths, *proj, normals, image_w, image_h);
  VK_COUNT_LAUNCH();
  return (int)cudaGetLastError();
}

extern "C" int vk_filter_depths(int image_w, int image_h, const float* src, float* dst, void* stream) {
  if (!src || !dst || image_w <= 0 || image_h <= 0) return VK_E_INVALID;
  dim3 grid((image_w + 31) / 32, (image_h + 7) / 8);
  vk_filter_depths_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(image_w, image_h, src, dst);
  VK_COUNT_LAUNCH();
  return (int)cudaGetLastError();
}

static int trace_bounds(const vkVolume* v, const vkFrame* view, float min_depth, float max_depth, float* bounds,
                        cudaStream_t s) {
  // Tracer::ComputePatches + ComputeBounds (src/tracer.cpp:49-76), 80x60 grid hard coded
  int rc = vk_reset_bounds(bounds, VK_BOUNDS_W * VK_BOUNDS_H, s);
  if (rc) return rc;
  BoundsParams b;
  fill_bounds_params(b, &view->depth_to_world, true, &view->depth_proj, VK_BLOCK_RES * v->voxel_length, min_depth,
                     max_depth, view->depth_w, view->depth_h, VK_BOUNDS_W, VK_BOUNDS_H, bounds);
  b.visible = v->visible[v->visible_slot];
  b.count = v->counters + VK_CNT_VISIBLE0 + v->visible_slot;
  b.entries = v->entries;
  vk_bounds_kernel<<<148, 256, 0, s>>>(b);
  VK_COUNT_LAUNCH();
  return (int)cudaGetLastError();
}

extern "C" int vk_trace(const vkVolume* v, const vkFrame* view, float min_depth, float max_depth, float* bounds,
                        float* out_depth, float* out_color, float* out_normal, void* stream) {
  if (!v || !view || !v->voxels || !bounds || !out_depth || !out_color) return VK_E_INVALID;
  cudaStream_t s = (cudaStream_t)stream;
  int rc = trace_bounds(v, view, min_depth, max_depth, bounds, s);
  if (rc) return rc;
  rc = vk_compute_points(v->entries, v->voxels, bounds, v->main_blocks, VK_BLOCK_RES * v->voxel_length,
                         v->voxel_length, v->trunc_length, &view->depth_to_world, &view->depth_proj, out_depth,
                         out_color, view->depth_w, view->depth_h, VK_BOUNDS_W, VK_BOUNDS_H, stream);
  if (rc) return rc;
  if (out_normal) rc = vk_compute_normals(out_depth, &view->depth_proj, out_normal, view->depth_w, view->depth_h, stream);
  return rc;
}

extern "C" int vk_trace_keys(const vkVolume* v, const vkFrame* view, float min_depth, float max_depth,
                             float* bounds, unsigned long long* out_keys, void* stream) {
  if (!v || !view || !v->voxels || !bounds || !out_keys) return VK_E_INVALID;
  cudaStream_t s = (cudaStream_t)stream;
  int rc = trace_bounds(v, view, min_depth, max_depth, bounds, s);
  if (rc) return rc;
  TraceParams t;
  fill_trace_params(t, v->entries, v->voxels, bounds, v->main_blocks, VK_BLOCK_RES * v->voxel_length,
                    v->voxel_length, v->trunc_length, &view->depth_to_world, &view->depth_proj, view->depth_w,
                    view->depth_h, VK_BOUNDS_W, VK_BOUNDS_H);
  t.keys = out_keys;
  t.shard_rank = v->shard_rank;
  t.shard_count = v->shard_count;
  t.shard_shift = v->shard_shift;
  dim3 grid((view->depth_w + 31) / 32, (view->depth_h + 7) / 8);
  vk_raycast_kernel<true><<<grid, 256, 0, s>>>(t);
  VK_COUNT_LAUNCH();
  return (int)cudaGetLastError();
}

extern "C" int vk_resolve_keys(const unsigned long long* keys, const vkFrame* view, float* out_depth,
                               float* out_color, float* out_normal, void* stream) {
  if (!keys || !view || !out_depth || !out_color) return VK_E_INVALID;
  const int n = view->depth_w * view->depth_h;
  vk_resolve_keys_kernel<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(keys, n, out_depth, out_color);
  VK_COUNT_LAUNCH();
  int rc = (int)cudaGetLastError();
  if (rc) return rc;
  if (out_normal) rc = vk_compute_normals(out_depth, &view->depth_proj, out_normal, view->depth_w, view->depth_h, stream);
  return rc;
}

namespace vk { std::atomic<unsigned long long> g_launches{0}; }
extern "C" int vk_version(void) { return 100; }
extern "C" unsigned long long vk_launch_count(void) { return vk::g_launches.load(); }
