// vk_math.cuh — device math with a fixed, contraction-free evaluation order.
//
// Every formula on the hot path is a knife-edge decision somewhere (which block a
// DDA step lands in, which pixel int(uv) picks, `distance > -tau`), so results are
// only reproducible if each multiply and add rounds exactly where the reference's
// source says it does.  The __f*_rn intrinsics are never fused by nvcc, so the
// functions below evaluate the reference's expressions operation by operation in
// source order; the CPU oracle (oracle/, built with -ffp-contract=off) and the
// reference built with -fmad=false produce the same bits.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/vulcan_b200.h"

namespace vk {

__device__ __forceinline__ float mul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float add(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float sub(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float fdiv(float a, float b) { return __fdiv_rn(a, b); }

struct Vec3 { float x, y, z; };

// Rows 0..2 of a column-major 4x4 matrix, loaded once per kernel.
struct Mat34 {
  float r[3][4];
  __device__ __forceinline__ void load(const float* m) {
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) r[i][j] = m[j * 4 + i];
  }
};

// Transform::operator*(Vector4f) (include/vulcan/transform.h:43-58):
// M(r,0)*x + M(r,1)*y + M(r,2)*z + M(r,3)*w, left to right.
__device__ __forceinline__ Vec3 xform(const Mat34& M, float x, float y, float z, float w) {
  Vec3 o;
  o.x = add(add(add(mul(M.r[0][0], x), mul(M.r[0][1], y)), mul(M.r[0][2], z)), mul(M.r[0][3], w));
  o.y = add(add(add(mul(M.r[1][0], x), mul(M.r[1][1], y)), mul(M.r[1][2], z)), mul(M.r[1][3], w));
  o.z = add(add(add(mul(M.r[2][0], x), mul(M.r[2][1], y)), mul(M.r[2][2], z)), mul(M.r[2][3], w));
  return o;
}

// Projection::Project (include/vulcan/projection.h:63-71)
__device__ __forceinline__ void project(const vkProjection& P, const Vec3& X, float& u, float& v) {
  const float inv_w = fdiv(1.0f, X.z);
  u = add(mul(mul(inv_w, P.fx), X.x), P.cx);
  v = add(mul(mul(inv_w, P.fy), X.y), P.cy);
}

// Projection::Unproject(uv) (include/vulcan/projection.h:80-89)
__device__ __forceinline__ Vec3 unproject(const vkProjection& P, float u, float v) {
  const float ifx = fdiv(1.0f, P.fx);
  const float ify = fdiv(1.0f, P.fy);
  Vec3 o;
  o.x = sub(mul(ifx, u), mul(P.cx, ifx));
  o.y = sub(mul(ify, v), mul(P.cy, ify));
  o.z = 1.0f;
  return o;
}

// Vector3f::Dot (include/vulcan/matrix.h:158-171): result starts at 0.
__device__ __forceinline__ float dot3(const Vec3& a, const Vec3& b) {
  float r = 0.0f;
  r = add(r, mul(a.x, b.x));
  r = add(r, mul(a.y, b.y));
  r = add(r, mul(a.z, b.z));
  return r;
}

// Vector3f::Normalize (include/vulcan/matrix.h:132-141): scale by 1/sqrt(dot).
__device__ __forceinline__ Vec3 normalized(const Vec3& a) {
  const float inv = fdiv(1.0f, __fsqrt_rn(dot3(a, a)));
  Vec3 o = {mul(a.x, inv), mul(a.y, inv), mul(a.z, inv)};
  return o;
}

// Block hash, Q1 (src/volume.cu:168-180, src/tracer.cu:151-155)
__device__ __forceinline__ uint32_t block_hash(int bx, int by, int bz, uint32_t K) {
  return (((uint32_t)bx * 73856093u) ^ ((uint32_t)by * 19349669u) ^ ((uint32_t)bz * 83492791u)) % K;
}

// vulcan::min / vulcan::max (include/vulcan/math.h:9-21)
__device__ __forceinline__ float vmin(float a, float b) { return (b < a) ? b : a; }
__device__ __forceinline__ float vmax(float a, float b) { return (b > a) ? b : a; }

// 64-bit image of a block coordinate triple (pad forced to 0) for register compares.
__device__ __forceinline__ unsigned long long pack_block(int bx, int by, int bz) {
  return (unsigned long long)(uint16_t)(int16_t)bx |
         ((unsigned long long)(uint16_t)(int16_t)by << 16) |
         ((unsigned long long)(uint16_t)(int16_t)bz << 32);
}

// L2-coherent 16-byte load of a hash entry (entries are written by other CTAs of
// the same launch during allocation, so L1 must be bypassed there).
__device__ __forceinline__ int4 ld_entry_cg(const vkHashEntry* e) {
  return __ldcg(reinterpret_cast<const int4*>(e));
}
__device__ __forceinline__ int4 ld_entry(const vkHashEntry* e) {
  return __ldg(reinterpret_cast<const int4*>(e));
}
__device__ __forceinline__ unsigned long long entry_block(const int4& e) {
  // x = block.x | block.y<<16, y = block.z | pad<<16 ; pad is not initialised by the
  // reference (include/vulcan/block.h:83-85), always 0 here.
  return (unsigned long long)(uint32_t)e.x | ((unsigned long long)((uint32_t)e.y & 0xffffu) << 32);
}

// ---- spatial sharding (no reference counterpart) ------------------
// Ownership is a pure function of the brick a block lies in.
__device__ __host__ __forceinline__ uint32_t shard_mix(int x, int y, int z) {
  uint32_t h = (uint32_t)x * 0x9E3779B1u ^ (uint32_t)y * 0x85EBCA77u ^ (uint32_t)z * 0xC2B2AE3Du;
  h ^= h >> 16; h *= 0x7FEB352Du; h ^= h >> 15; h *= 0x846CA68Bu; h ^= h >> 16;
  return h;
}
__device__ __host__ __forceinline__ int shard_owner(int bx, int by, int bz, int shift, int count) {
  return (int)(shard_mix(bx >> shift, by >> shift, bz >> shift) % (uint32_t)count);
}
// A rank stores a block if it owns it or any of its 26 neighbours (1-block halo).
__device__ __forceinline__ bool shard_stores(int bx, int by, int bz, int shift, int count, int rank) {
  if (count <= 1) return true;
  for (int dz = -1; dz <= 1; ++dz)
    for (int dy = -1; dy <= 1; ++dy)
      for (int dx = -1; dx <= 1; ++dx)
        if (shard_owner(bx + dx, by + dy, bz + dz, shift, count) == rank) return true;
  return false;
}

}  // namespace vk
