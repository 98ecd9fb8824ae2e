// vk_math.cuh — device math with a fixed evaluation order and a fixed fusion pattern.
//
// Every formula on the hot path is a knife-edge decision somewhere (which block a DDA
// step lands in, which pixel int(uv) picks, `distance > -tau`), so results are only
// reproducible if each multiply and add rounds exactly where the reference's machine code
// rounds.  Nothing here is left to the compiler: every operation is an explicit
// __fmul_rn / __fadd_rn / __fmaf_rn (never re-fused or split by nvcc), in one of two modes
//
//   FP<true>   VK_MATH_REFERENCE: the mul+add pairs that nvcc 12.9 contracts into FFMA when it
//              builds the reference's own sources with its default flags (-O3, -fmad=true) are
//              fused here in the same places.  The pattern was read off the SASS of that build
//              (oracle/_ref/libvulcan_ref.so, see DESIGN.md "FMA map"); each helper below cites
//              the reference expression it mirrors.
//   FP<false>  VK_MATH_STRICT: one rounding per source operation (the reference built with
//              -fmad=false, and the CPU oracle built with -ffp-contract=off).
//
// Both modes use the same association order, so a helper is written once in terms of
// mad / nmad / msub.  Division, reciprocal and square root are IEEE correctly rounded in the
// reference (nvcc defaults -prec-div=true -prec-sqrt=true) and here.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/vulcan_b200.h"

namespace vk {

__device__ __forceinline__ float mul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float add(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float sub(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float fdiv(float a, float b) { return __fdiv_rn(a, b); }
__device__ __forceinline__ float rcp(float a) { return __frcp_rn(a); }   // == 1.0f / a, correctly rounded

template <bool FMA>
struct FP {
  // a * b + c
  static __device__ __forceinline__ float mad(float a, float b, float c) {
    return FMA ? __fmaf_rn(a, b, c) : __fadd_rn(__fmul_rn(a, b), c);
  }
  // c - a * b
  static __device__ __forceinline__ float nmad(float a, float b, float c) {
    return FMA ? __fmaf_rn(-a, b, c) : __fsub_rn(c, __fmul_rn(a, b));
  }
  // a * b - c
  static __device__ __forceinline__ float msub(float a, float b, float c) {
    return FMA ? __fmaf_rn(a, b, -c) : __fsub_rn(__fmul_rn(a, b), c);
  }
};

struct Vec3 { float x, y, z; };

// Rows 0..2 of a 4x4 transform, row-major 3x4 (the host passes rows, see rows34()).
struct Mat34 {
  float r[3][4];
  __device__ __forceinline__ void load(const float* rows) {
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) r[i][j] = rows[i * 4 + j];
  }
};

// Transform::operator*(Vector4f) (include/vulcan/transform.h:43-58), one row:
//   M(r,0)*x + M(r,1)*y + M(r,2)*z + M(r,3)*w   evaluated left to right.
// Reference SASS: FMUL y*m1; FFMA x*m0 + .; FFMA z*m2 + .; then w == 1: FADD m3,
// w == 0: FFMA 0*m3 + .
template <bool F>
__device__ __forceinline__ float row_acc(const float* m, float x, float y, float z) {
  return FP<F>::mad(z, m[2], FP<F>::mad(x, m[0], mul(y, m[1])));
}
template <bool F>
__device__ __forceinline__ Vec3 xform1(const Mat34& M, float x, float y, float z) {  // w = 1
  Vec3 o;
  o.x = add(row_acc<F>(M.r[0], x, y, z), M.r[0][3]);
  o.y = add(row_acc<F>(M.r[1], x, y, z), M.r[1][3]);
  o.z = add(row_acc<F>(M.r[2], x, y, z), M.r[2][3]);
  return o;
}
// Same sum in ComputePatchesKernel (src/tracer.cu:31-49), where the compiler hoists the x and y
// products out of the corner loops and fuses the other way round:
//   FMUL x*m0; FFMA y*m1 + .; FFMA z*m2 + .; FADD m3
template <bool F>
__device__ __forceinline__ Vec3 xform1_xfirst(const Mat34& M, float x, float y, float z) {
  Vec3 o;
#pragma unroll
  for (int r = 0; r < 3; ++r) {
    const float acc = FP<F>::mad(z, M.r[r][2], FP<F>::mad(y, M.r[r][1], mul(x, M.r[r][0])));
    (r == 0 ? o.x : r == 1 ? o.y : o.z) = add(acc, M.r[r][3]);
  }
  return o;
}
template <bool F>
__device__ __forceinline__ float xform1_z(const Mat34& M, float x, float y, float z) {
  return add(row_acc<F>(M.r[2], x, y, z), M.r[2][3]);
}
template <bool F>
__device__ __forceinline__ Vec3 xform0(const Mat34& M, float x, float y, float z) {  // w = 0
  Vec3 o;
  o.x = FP<F>::mad(0.0f, M.r[0][3], row_acc<F>(M.r[0], x, y, z));
  o.y = FP<F>::mad(0.0f, M.r[1][3], row_acc<F>(M.r[1], x, y, z));
  o.z = FP<F>::mad(0.0f, M.r[2][3], row_acc<F>(M.r[2], x, y, z));
  return o;
}

// Projection::Project (include/vulcan/projection.h:63-71):
//   inv_w = 1.0f / X.z;  u = inv_w * fx * X.x + cx     (FMUL inv_w*fx; FFMA . * X.x + cx)
template <bool F>
__device__ __forceinline__ void project(const vkProjection& P, const Vec3& X, float& u, float& v) {
  const float inv_w = rcp(X.z);
  u = FP<F>::mad(mul(inv_w, P.fx), X.x, P.cx);
  v = FP<F>::mad(mul(inv_w, P.fy), X.y, P.cy);
}

// Projection::Unproject(uv) (include/vulcan/projection.h:80-89):
//   ifx * u - cx * ifx                                  (FMUL cx*ifx; FFMA ifx*u - .)
struct Unprojector {
  float ifx, ify, cxi, cyi;
  __device__ __forceinline__ void init(const vkProjection& P) {
    ifx = rcp(P.fx);
    ify = rcp(P.fy);
    cxi = mul(P.cx, ifx);
    cyi = mul(P.cy, ify);
  }
  template <bool F>
  __device__ __forceinline__ float x(float u) const { return FP<F>::msub(ifx, u, cxi); }
  template <bool F>
  __device__ __forceinline__ float y(float v) const { return FP<F>::msub(ify, v, cyi); }
};

// Vector3f::Dot (include/vulcan/matrix.h:158-171): r = 0; r += a[i] * b[i]  (FFMA chain from 0)
template <bool F>
__device__ __forceinline__ float dot3(const Vec3& a, const Vec3& b) {
  float r = FP<F>::mad(a.x, b.x, 0.0f);
  r = FP<F>::mad(a.y, b.y, r);
  r = FP<F>::mad(a.z, b.z, r);
  return r;
}

// Vector3f::Normalize (include/vulcan/matrix.h:132-141): scale by 1 / sqrt(dot)
template <bool F>
__device__ __forceinline__ Vec3 normalized(const Vec3& a) {
  const float inv = rcp(__fsqrt_rn(dot3<F>(a, a)));
  Vec3 o = {mul(a.x, inv), mul(a.y, inv), mul(a.z, inv)};
  return o;
}

// vulcan::min / vulcan::max (include/vulcan/math.h:9-21)
__device__ __forceinline__ float vmin(float a, float b) { return (b < a) ? b : a; }
__device__ __forceinline__ float vmax(float a, float b) { return (b > a) ? b : a; }

// ---- correctly rounded division by a loop-invariant divisor ------------------------------
// q = RN(x / d) from r = RN(1 / d) with one FMA residual step:
//   q0 = RN(x * r);  e = x - q0 * d (exact, FMA);  q = RN(q0 + e * r)
// For a given d this is either right for every x in the guarded range or it is not; the host
// proves it per divisor by running it against __fdiv_rn over every mantissa of x once (one binade
// is a complete proof inside the guarded range, where every step scales exactly with the exponent
// of x: vk_divcheck_kernel in vk_runtime.cu) and clears `fast` otherwise.  Outside the guarded
// range (tiny / huge / non-finite quotients) the IEEE division instruction sequence is used.
struct Div {
  float d, r;
  int fast;
};
__device__ __forceinline__ float divc(float x, const Div& dv) {
  const float q0 = __fmul_rn(x, dv.r);
  // dividend exponent in [2^-40, 2^40) and (make_div) divisor in [2^-63, 2^64): quotient, residual
  // and correction are all normal numbers far from overflow, so e is exact
  const uint32_t ex = (__float_as_uint(x) >> 23) & 0xffu;
  if (dv.fast && (ex - 87u) < 80u) {
    const float e = __fmaf_rn(-q0, dv.d, x);
    return __fmaf_rn(e, dv.r, q0);
  }
  return __fdiv_rn(x, dv.d);
}

// Three quotients by the same divisor with ONE range guard instead of three (the guard is the
// same dividend range divc was proven on; zero, tiny, huge and NaN components take the IEEE path).
__device__ __forceinline__ void divc3(float x, float y, float z, const Div& dv, float& qx, float& qy, float& qz) {
  const float ax = fabsf(x), ay = fabsf(y), az = fabsf(z);
  const float lo = fminf(fminf(ax, ay), az), hi = fmaxf(fmaxf(ax, ay), az);
  const float x0 = __fmul_rn(x, dv.r), y0 = __fmul_rn(y, dv.r), z0 = __fmul_rn(z, dv.r);
  qx = __fmaf_rn(__fmaf_rn(-x0, dv.d, x), dv.r, x0);
  qy = __fmaf_rn(__fmaf_rn(-y0, dv.d, y), dv.r, y0);
  qz = __fmaf_rn(__fmaf_rn(-z0, dv.d, z), dv.r, z0);
  // 2^-40 <= |.| < 2^40 for all three (exponent field 87..166, as in divc)
  if (!(dv.fast && lo >= 9.094947017729282e-13f && hi < 1099511627776.0f)) {
    qx = __fdiv_rn(x, dv.d);
    qy = __fdiv_rn(y, dv.d);
    qz = __fdiv_rn(z, dv.d);
  }
}

// ---- block hash, Q1 (src/volume.cu:168-180, src/tracer.cu:151-155) -------------------------
// h = (bx*P1 ^ by*P2 ^ bz*P3) % K with a precomputed multiplier instead of a hardware-less
// 32-bit modulo (~25 instructions): m = floor(2^32 / K); q = umulhi(h, m) <= h / K < q + 2.
struct Mod {
  uint32_t K, m;
};
__host__ __device__ __forceinline__ Mod make_mod(uint32_t K) {
  Mod o;
  o.K = K;
  o.m = (K <= 1) ? 0xffffffffu : (uint32_t)(0x100000000ull / K);
  return o;
}
__device__ __forceinline__ uint32_t fastmod(uint32_t h, const Mod& md) {
  // 2^32 = m*K + g, 0 <= g < K  =>  h*m/2^32 = h/K - h*g/(K*2^32) > h/K - 1  =>  r in [0, 2K)
  uint32_t r = h - __umulhi(h, md.m) * md.K;
  if (r >= md.K) r -= md.K;
  return r;
}
__device__ __forceinline__ uint32_t block_hash(int bx, int by, int bz, const Mod& md) {
  return fastmod(((uint32_t)bx * 73856093u) ^ ((uint32_t)by * 19349669u) ^ ((uint32_t)bz * 83492791u), md);
}

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
__device__ __forceinline__ void entry_origin(const int4& e, int& ox, int& oy, int& oz) {
  ox = (int)(int16_t)(e.x & 0xffff);
  oy = (int)(int16_t)((uint32_t)e.x >> 16);
  oz = (int)(int16_t)(e.y & 0xffff);
}

// ---- free-space filter (vkVolume::space_bits) ------------------------------------------------
// One bit per block coordinate triple modulo 128, set when a block with those low bits is
// allocated (never cleared).  Unlike the hash-bucket bitmap it has spatial locality: a 32-bit word
// covers 4x4x2 blocks and a 32-byte sector 8x8x4 blocks, so a ray stepping block by block through
// free space, and its neighbours in the warp, keep hitting the same L1 sectors.  A clear bit
// proves the block is not in the table; a set bit means "ask the table".
__host__ __device__ __forceinline__ void space_bit_index(int bx, int by, int bz, uint32_t& word, uint32_t& bit) {
  const uint32_t x = (uint32_t)bx & 127u, y = (uint32_t)by & 127u, z = (uint32_t)bz & 127u;
  bit = (x & 3u) | ((y & 3u) << 2) | ((z & 1u) << 4);
  const uint32_t wx = x >> 2, wy = y >> 2, wz = z >> 1;  // 32 x 32 x 64 words
  word = ((wx & 1u) | ((wy & 1u) << 1) | ((wz & 1u) << 2)) | (((wx >> 1) + 16u * ((wy >> 1) + 16u * (wz >> 1))) << 3);
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
// A rank stores the blocks it owns and, with halo = 1, their 26 neighbours.
__device__ __forceinline__ bool shard_stores(int bx, int by, int bz, int shift, int count, int rank, int halo) {
  if (count <= 1) return true;
  if (halo <= 0) return shard_owner(bx, by, bz, shift, count) == rank;
  for (int dz = -1; dz <= 1; ++dz)
    for (int dy = -1; dy <= 1; ++dy)
      for (int dx = -1; dx <= 1; ++dx)
        if (shard_owner(bx + dx, by + dy, bz + dz, shift, count) == rank) return true;
  return false;
}

}  // namespace vk
