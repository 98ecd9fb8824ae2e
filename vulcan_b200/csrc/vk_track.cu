// vk_track.cu — the caller side of Tracer::Trace: image pyramids and the projective
// point-to-plane depth tracker.
//
//   vk_downsample_frame   Frame::Downsample (src/frame.cpp:38-57; kernels src/image.cu:101-165): depth and
//                         normals by nearest sample, colour by 2x2 average, all three images in ONE launch
//                         (the reference launches three kernels).
//   vk_track_residuals    DepthTracker::ComputeResiduals (src/depth_tracker.cu:101-121,279-306)
//   vk_track_jacobian     DepthTracker::ComputeJacobian  (src/depth_tracker.cu:123-144,308-344)
//   vk_track_system       DepthTracker::ComputeSystem    (src/depth_tracker.cu:146-274,346-378)
//   vk_track_depth        Tracker::Track (src/tracker.cpp:57-67,112-165) with DepthTracker::ApplyUpdate
//                         (src/depth_tracker.cpp:23-88): every Gauss-Newton iteration is ONE kernel.  The
//                         last CTA to finish folds the per-CTA partial sums in a fixed order, solves the
//                         6x6 (or 3x3) normal equations, applies the update to the pose held in device
//                         memory and raises the `done` flag on convergence; the next iteration's kernel
//                         reads the pose from there.  A Track call is max_iterations launches back to
//                         back on one stream with no host round trip (the reference does two thrust fills,
//                         a kernel, two device->host copies and an Eigen solve per iteration), and the
//                         sums are deterministic (the reference adds CTA sums with atomicAdd in arrival
//                         order).
//
// Arithmetic: per-pixel geometry goes through vk_math.cuh.  VK_MATH_STRICT is bit-identical to the
// reference built with -fmad=false; VK_MATH_REFERENCE fuses with the helpers' usual pattern, which is not
// verified against the SASS of this particular reference kernel (tests hold it to 1e-6 instead).
#include <cuda_runtime.h>
#include <math.h>
#include "vk_launch.h"
#include "vk_math.cuh"

namespace vk {

// ---- Frame::Downsample ------------------------------------------------------------------------------

// Image / ColorImage DownsampleKernel (src/image.cu:101-165).  Average: the four samples are added to 0
// in the order (0,+1) (0,0) (+1,+1) (+1,0) and scaled by 0.25f.
__device__ __forceinline__ float avg4(float s01, float s00, float s11, float s10) {
  return mul(add(add(add(add(0.0f, s01), s00), s11), s10), 0.25f);
}

__global__ void __launch_bounds__(256) vk_downsample_frame_kernel(
    const float* __restrict__ depth, const float* __restrict__ normal, int dw, int dh, float* __restrict__ out_depth,
    float* __restrict__ out_normal, const float* __restrict__ color, int cw, int ch, float* __restrict__ out_color,
    int depth_nearest, int normal_nearest, int color_nearest) {
  const int odw = dw >> 1, odh = dh >> 1, ocw = cw >> 1, och = ch >> 1;
  const int nd = odw * odh, nc = ocw * och;
  const int total = nd > nc ? nd : nc;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    if (i < nd) {
      const int x = i % odw, y = i / odw;
      const size_t s00 = (size_t)(2 * y) * dw + 2 * x, s10 = s00 + dw;
      if (depth) {
        out_depth[i] = depth_nearest ? __ldg(depth + s00)
                                     : avg4(__ldg(depth + s00 + 1), __ldg(depth + s00), __ldg(depth + s10 + 1),
                                            __ldg(depth + s10));
      }
      if (normal) {
#pragma unroll
        for (int c = 0; c < 3; ++c)
          out_normal[3 * (size_t)i + c] =
              normal_nearest ? __ldg(normal + 3 * s00 + c)
                             : avg4(__ldg(normal + 3 * (s00 + 1) + c), __ldg(normal + 3 * s00 + c),
                                    __ldg(normal + 3 * (s10 + 1) + c), __ldg(normal + 3 * s10 + c));
      }
    }
    if (color && i < nc) {
      const int x = i % ocw, y = i / ocw;
      const size_t s00 = (size_t)(2 * y) * cw + 2 * x, s10 = s00 + cw;
      // the two source pixels of a row are 24 contiguous bytes
      float a[6], b[6];
#pragma unroll
      for (int c = 0; c < 6; ++c) {
        a[c] = __ldg(color + 3 * s00 + c);
        b[c] = color_nearest ? 0.0f : __ldg(color + 3 * s10 + c);
      }
#pragma unroll
      for (int c = 0; c < 3; ++c)
        out_color[3 * (size_t)i + c] = color_nearest ? a[c] : avg4(a[3 + c], a[c], b[3 + c], b[c]);
    }
  }
}

// ---- DepthTracker -------------------------------------------------------------------------------------

struct TrackImages {
  const float* key_depth;
  const float* key_normal;
  const float* depth;
  const float* normal;
  int kw, kh, w, h;
  vkProjection key_proj, proj;
  float Twm[12];      // keyframe depth_to_world, rows
  float Twm_inv[12];  // its inverse, rows
};

// Device-resident solver state of one Track call (vk_track_depth).  Layout is part of the C ABI only as
// an opaque VK_TRACK_STATE_FLOATS-float buffer; the host reads it back through vk_track_depth itself.
struct TrackState {
  float Twc[16];      // current frame pose, column major (vulcan::Matrix4f)
  float Twc_inv[16];
  float system[27];   // 21 packed lower-triangle Hessian entries + 6 gradient entries of the last iteration
  float update_norm;
  int iteration;
  int done;
  unsigned ticket;
  int max_iterations;
  int evaluations;    // systems solved
  int pad_;
  unsigned long long stamps[6];  // %globaltimer of the last iteration's last CTA: entry, pixels done, ticket,
                                 // sums folded, solved; [5] = entry of CTA 0 (tools/track_once.py)
};
static_assert(sizeof(TrackState) <= VK_TRACK_STATE_FLOATS * sizeof(float), "VK_TRACK_STATE_FLOATS too small");

// depth and normal of one frame pixel
struct FramePixel { float d, nx, ny, nz; };
__device__ __forceinline__ FramePixel load_frame_pixel(const TrackImages& im, int x, int y) {
  FramePixel p = {0.0f, 0.0f, 0.0f, 0.0f};
  if (x < im.w && y < im.h) {
    const size_t index = (size_t)y * im.w + x;
    p.d = __ldg(im.depth + index);
    p.nx = __ldg(im.normal + 3 * index);
    p.ny = __ldg(im.normal + 3 * index + 1);
    p.nz = __ldg(im.normal + 3 * index + 2);
  }
  return p;
}

// Evaluate<> (src/depth_tracker.cu:16-99).  Returns false when the pixel contributes nothing
// (residual 0, Jacobian 0).
template <bool F>
__device__ __forceinline__ bool track_evaluate(const TrackImages& im, const Mat34& Twc, const Mat34& Twm,
                                               const Mat34& Tmw, const Unprojector& up, const Unprojector& kup, int x,
                                               int y, const FramePixel& px, float& residual, float* J) {
  // the frame's depth and normal are fetched together by the caller (load_frame_pixel, one tile ahead in
  // the iteration kernel), the keyframe's depth and normal together here (the normal speculatively: kindex
  // is inside the image once the bounds test passed): two dependent round trips per pixel instead of three
  const float d = px.d, fnx = px.nx, fny = px.ny, fnz = px.nz;
  if (!(d > 0)) return false;
  const float u = (float)x + 0.5f, v = (float)y + 0.5f;
  // Unproject(u, v, d) = d * Unproject(uv)  (include/vulcan/projection.h:80-101)
  const float cx = mul(up.x<F>(u), d), cy = mul(up.y<F>(v), d), cz = mul(1.0f, d);
  const Vec3 Xw = xform1<F>(Twc, cx, cy, cz);
  const Vec3 Xm = xform1<F>(Tmw, Xw.x, Xw.y, Xw.z);
  float ku, kv;
  project<F>(im.key_proj, Xm, ku, kv);
  if (!(ku >= 0 && ku < (float)im.kw && kv >= 0 && kv < (float)im.kh)) return false;
  const int kx = (int)ku, ky = (int)kv;
  const int kindex = ky * im.kw + kx;
  const float kd = __ldg(im.key_depth + kindex);
  const float knx = __ldg(im.key_normal + 3 * (size_t)kindex), kny = __ldg(im.key_normal + 3 * (size_t)kindex + 1),
              knz = __ldg(im.key_normal + 3 * (size_t)kindex + 2);
  if (!(kd > 0)) return false;
  const Vec3 fn = xform0<F>(Twc, fnx, fny, fnz);
  const Vec3 kn = xform0<F>(Twm, knx, kny, knz);
  if (!(dot3<F>(kn, kn) > 0.0f && dot3<F>(fn, kn) > 0.5f)) return false;
  const float fu = floorf(ku) + 0.5f, fv = floorf(kv) + 0.5f;
  const float yx = mul(kup.x<F>(fu), kd), yy = mul(kup.y<F>(fv), kd), yz = mul(1.0f, kd);
  const Vec3 Yw = xform1<F>(Twm, yx, yy, yz);
  const Vec3 delta = {sub(Xw.x, Yw.x), sub(Xw.y, Yw.y), sub(Xw.z, Yw.z)};
  if (!(dot3<F>(delta, delta) < 0.05f)) return false;
  residual = dot3<F>(delta, kn);
  J[0] = FP<F>::msub(kn.z, Xw.y, mul(kn.y, Xw.z));
  J[1] = FP<F>::msub(kn.x, Xw.z, mul(kn.z, Xw.x));
  J[2] = FP<F>::msub(kn.y, Xw.x, mul(kn.x, Xw.y));
  J[3] = kn.x;
  J[4] = kn.y;
  J[5] = kn.z;
  return true;
}

__device__ __forceinline__ unsigned long long gtime() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

__device__ __forceinline__ void load_rows_colmajor(Mat34& M, const float* colmajor) {
#pragma unroll
  for (int r = 0; r < 3; ++r)
#pragma unroll
    for (int c = 0; c < 4; ++c) M.r[r][c] = colmajor[c * 4 + r];
}

// mode 0: residuals (float per pixel), mode 1: Jacobian rows (6 floats per pixel)
template <bool F, int MODE>
__global__ void __launch_bounds__(256) vk_track_pixels_kernel(TrackImages im, const float* __restrict__ Twc_colmajor,
                                                              int translation, float* __restrict__ out) {
  Mat34 Twc, Twm, Tmw;
  load_rows_colmajor(Twc, Twc_colmajor);
  Twm.load(im.Twm);
  Tmw.load(im.Twm_inv);
  Unprojector up, kup;
  up.init(im.proj);
  kup.init(im.key_proj);
  const int n = im.w * im.h;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    float r = 0, J[6] = {0, 0, 0, 0, 0, 0};
    const FramePixel px = load_frame_pixel(im, i % im.w, i / im.w);
    const bool ok = track_evaluate<F>(im, Twc, Twm, Tmw, up, kup, i % im.w, i / im.w, px, r, J);
    if (MODE == 0) {
      out[i] = ok ? r : 0.0f;
    } else {
#pragma unroll
      for (int k = 0; k < 6; ++k) out[6 * (size_t)i + k] = (ok && (translation || k < 3)) ? J[k] : 0.0f;
    }
  }
}

// 6x6 (n = 6) or 3x3 (n = 3) LDL^T solve of H x = -g in double; H is the packed lower triangle
// (row r, column c <= r at r (r + 1) / 2 + c), as the reference's kernel stores it
// (src/depth_tracker.cu:214-223, unpacked in src/tracker.cpp:133-143).  A vanishing pivot zeroes its
// component (no valid pixel -> no motion).
template <int N>
__device__ __forceinline__ void track_solve(const float* packed, const float* g, float* x) {
  // fully unrolled: L, D and y live in registers
  double L[N][N], D[N], invD[N], y[N];
#pragma unroll
  for (int r = 0; r < N; ++r)
#pragma unroll
    for (int c = 0; c <= r; ++c) L[r][c] = (double)packed[r * (r + 1) / 2 + c];
#pragma unroll
  for (int j = 0; j < N; ++j) {
    double d = L[j][j];
#pragma unroll
    for (int k = 0; k < j; ++k) d -= L[j][k] * L[j][k] * D[k];
    D[j] = d;
    invD[j] = fabs(d) > 1e-30 ? 1.0 / d : 0.0;
#pragma unroll
    for (int i = j + 1; i < N; ++i) {
      double s = L[i][j];
#pragma unroll
      for (int k = 0; k < j; ++k) s -= L[i][k] * L[j][k] * D[k];
      L[i][j] = s * invD[j];
    }
  }
#pragma unroll
  for (int i = 0; i < N; ++i) {
    double s = -(double)g[i];
#pragma unroll
    for (int k = 0; k < i; ++k) s -= L[i][k] * y[k];
    y[i] = s;
  }
#pragma unroll
  for (int i = 0; i < N; ++i) y[i] = y[i] * invD[i];
#pragma unroll
  for (int i = N - 1; i >= 0; --i) {
    double s = y[i];
#pragma unroll
    for (int k = i + 1; k < N; ++k) s -= L[k][i] * y[k];
    y[i] = s;
  }
#pragma unroll
  for (int i = 0; i < 6; ++i) x[i] = i < N ? (float)y[i] : 0.0f;
}

// DepthTracker::ApplyUpdate (src/depth_tracker.cpp:23-88) in single precision, one rounding per
// operation like the reference's host code (x86-64 without FMA contraction).
__device__ void track_apply_update(const float* u, float* Twc, float* Twc_inv) {
#define MAT(m, r, c) (m)[(c) * 4 + (r)]
  float T[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) T[i] = (i % 5 == 0) ? 1.0f : 0.0f;
  MAT(T, 0, 1) = -u[2]; MAT(T, 0, 2) = +u[1]; MAT(T, 0, 3) = +u[3];
  MAT(T, 1, 0) = +u[2]; MAT(T, 1, 2) = +u[0]; MAT(T, 1, 3) = +u[4];  // sign of (1,2) as in the reference
  MAT(T, 2, 0) = -u[1]; MAT(T, 2, 1) = +u[0]; MAT(T, 2, 3) = +u[5];
  float M[16];
#pragma unroll
  for (int p = 0; p < 4; ++p)
#pragma unroll
    for (int m = 0; m < 4; ++m) {
      float acc = 0;
#pragma unroll
      for (int n = 0; n < 4; ++n) acc = add(acc, mul(MAT(T, m, n), MAT(Twc, n, p)));
      MAT(M, m, p) = acc;
    }
  float X[3] = {MAT(M, 0, 0), MAT(M, 1, 0), MAT(M, 2, 0)};
  float Y[3] = {MAT(M, 0, 1), MAT(M, 1, 1), MAT(M, 2, 1)};
  float Z[3];
  {
    float s = 0;
#pragma unroll
    for (int i = 0; i < 3; ++i) s = add(s, mul(X[i], X[i]));
    const float inv = fdiv(1.0f, __fsqrt_rn(s));
#pragma unroll
    for (int i = 0; i < 3; ++i) X[i] = mul(X[i], inv);
    s = 0;
#pragma unroll
    for (int i = 0; i < 3; ++i) s = add(s, mul(Y[i], Y[i]));
    const float invy = fdiv(1.0f, __fsqrt_rn(s));
#pragma unroll
    for (int i = 0; i < 3; ++i) Y[i] = mul(Y[i], invy);
  }
  Z[0] = sub(mul(X[1], Y[2]), mul(X[2], Y[1]));
  Z[1] = sub(mul(X[2], Y[0]), mul(X[0], Y[2]));
  Z[2] = sub(mul(X[0], Y[1]), mul(X[1], Y[0]));
  Y[0] = sub(mul(Z[1], X[2]), mul(Z[2], X[1]));
  Y[1] = sub(mul(Z[2], X[0]), mul(Z[0], X[2]));
  Y[2] = sub(mul(Z[0], X[1]), mul(Z[1], X[0]));
  const float t[3] = {MAT(M, 0, 3), MAT(M, 1, 3), MAT(M, 2, 3)};
  // Transform::Translate(t) * Transform::Rotate(R): matrix = [R | t]; inverse = R^T * Translate(-t)
  // evaluated as the 4x4 product the reference forms (include/vulcan/transform.h:60-64, matrix.h:299-317)
  float R4[16], Rt4[16], Tt[16], Tti[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) R4[i] = Rt4[i] = Tt[i] = Tti[i] = (i % 5 == 0) ? 1.0f : 0.0f;
#pragma unroll
  for (int r = 0; r < 3; ++r) {
    MAT(R4, r, 0) = X[r]; MAT(R4, r, 1) = Y[r]; MAT(R4, r, 2) = Z[r];
    MAT(Rt4, 0, r) = X[r]; MAT(Rt4, 1, r) = Y[r]; MAT(Rt4, 2, r) = Z[r];
    MAT(Tt, r, 3) = t[r];
    MAT(Tti, r, 3) = -t[r];
  }
#pragma unroll
  for (int p = 0; p < 4; ++p)
#pragma unroll
    for (int m = 0; m < 4; ++m) {
      float a = 0, b = 0;
#pragma unroll
      for (int n = 0; n < 4; ++n) {
        a = add(a, mul(MAT(Tt, m, n), MAT(R4, n, p)));
        b = add(b, mul(MAT(Rt4, m, n), MAT(Tti, n, p)));
      }
      MAT(Twc, m, p) = a;
      MAT(Twc_inv, m, p) = b;
    }
#undef MAT
}

// Warp sums of K <= 32 values per lane with 31 shuffles instead of 5 K: at every butterfly step a lane
// hands the half of its values that its partner keeps and adds the partner's half of the ones it keeps, so
// the number of live values halves with the distance.  Afterwards lane L holds the sum over the warp of
// value L.  The tree is fixed: the same inputs give the same bits.
template <int K>
__device__ __forceinline__ float warp_sums_transposed(const float* acc, int lane) {
  float v[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) v[i] = i < K ? acc[i] : 0.0f;
#pragma unroll
  for (int o = 16; o >= 1; o >>= 1) {
    const bool up = (lane & o) != 0;
#pragma unroll
    for (int i = 0; i < o; ++i) {
      const float send = up ? v[i] : v[i + o];
      const float keep = up ? v[i + o] : v[i];
      v[i] = keep + __shfl_xor_sync(0xffffffffu, send, o);
    }
  }
  return v[0];
}

// One Gauss-Newton iteration (or, with solve == 0, only ComputeSystem).  Each thread accumulates its
// pixels' J^T r and the packed lower triangle of J^T J in registers; CTA sums go to partials[cta][27];
// the last CTA to arrive adds them in CTA order and, when solving, advances the state.
template <bool F, bool TRANS>
__global__ void __launch_bounds__(256, 3) vk_track_iter_kernel(TrackImages im, TrackState* __restrict__ st,
                                                            float* __restrict__ partials, int solve) {
  constexpr int P = TRANS ? 6 : 3, H = TRANS ? 21 : 6, K = H + P;
  pdl_wait();  // the previous iteration's kernel wrote the state
  if (solve && *(volatile int*)&st->done) return;
  unsigned long long t0 = 0, t1 = 0, t2 = 0, t3 = 0;
  if (threadIdx.x == 0) {
    t0 = gtime();
    if (blockIdx.x == 0) st->stamps[5] = t0;
  }
  Mat34 Twc, Twm, Tmw;
  load_rows_colmajor(Twc, st->Twc);
  Twm.load(im.Twm);
  Tmw.load(im.Twm_inv);
  Unprojector up, kup;
  up.init(im.proj);
  kup.init(im.key_proj);
  float acc[K];
#pragma unroll
  for (int k = 0; k < K; ++k) acc[k] = 0;
  // 32 x 8 pixel tiles per CTA pass: neighbouring lanes gather neighbouring keyframe pixels
  const int tiles_x = (im.w + 31) >> 5, tiles_y = (im.h + 7) >> 3;
  const int lx = threadIdx.x & 31, ly = threadIdx.x >> 5;
  const int tiles = tiles_x * tiles_y;
  // the frame pixel of the next pass is fetched before the current one is evaluated
  FramePixel px = {0.0f, 0.0f, 0.0f, 0.0f};
  if ((int)blockIdx.x < tiles) px = load_frame_pixel(im, (blockIdx.x % tiles_x) * 32 + lx, (blockIdx.x / tiles_x) * 8 + ly);
  for (int t = blockIdx.x; t < tiles; t += gridDim.x) {
    const int x = (t % tiles_x) * 32 + lx, y = (t / tiles_x) * 8 + ly;
    const FramePixel cur = px;
    const int tn = t + gridDim.x;
    if (tn < tiles) px = load_frame_pixel(im, (tn % tiles_x) * 32 + lx, (tn / tiles_x) * 8 + ly);
    if (x >= im.w || y >= im.h) continue;
    float r = 0, J[6];
    if (!track_evaluate<F>(im, Twc, Twm, Tmw, up, kup, x, y, cur, r, J)) continue;
#pragma unroll
    for (int a = 0, k = 0; a < P; ++a)
#pragma unroll
      for (int b = 0; b <= a; ++b, ++k) acc[k] = __fmaf_rn(J[a], J[b], acc[k]);
#pragma unroll
    for (int a = 0; a < P; ++a) acc[H + a] = __fmaf_rn(J[a], r, acc[H + a]);
  }
  if (threadIdx.x == 0) t1 = gtime();
  __shared__ float warp_sums[8][K];
  __shared__ float s_system[27];
  __shared__ bool last;
  {
    const float mine = warp_sums_transposed<K>(acc, lx);
    if (lx < K) warp_sums[ly][lx] = mine;
  }
  __syncthreads();
  // partials are stored value-major ([value][CTA]) so that the fold below reads them coalesced
  if (threadIdx.x < K) {
    float s = 0;
#pragma unroll
    for (int w = 0; w < 8; ++w) s += warp_sums[w][threadIdx.x];
    partials[(size_t)threadIdx.x * gridDim.x + blockIdx.x] = s;
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) last = (atomicAdd(&st->ticket, 1u) == gridDim.x - 1);
  __syncthreads();
  if (!last) return;
  __threadfence();
  if (threadIdx.x == 0) t2 = gtime();
  // the last CTA folds the CTA sums: thread t takes CTAs t and t + 256 of every value (all loads
  // independent and coalesced: one round trip to L2), then the same fixed reduction tree as above - every
  // run adds the same numbers in the same order.  (track_grid keeps the grid below 512 CTAs.)
#pragma unroll
  for (int k = 0; k < K; ++k) {
    const float* col = partials + (size_t)k * gridDim.x;
    const float p0 = threadIdx.x < gridDim.x ? __ldcg(col + threadIdx.x) : 0.0f;
    const float p1 = threadIdx.x + 256 < gridDim.x ? __ldcg(col + threadIdx.x + 256) : 0.0f;
    acc[k] = p0 + p1;
  }
  __syncthreads();  // warp_sums is reused
  {
    const float mine = warp_sums_transposed<K>(acc, lx);
    if (lx < K) warp_sums[ly][lx] = mine;
  }
  __syncthreads();
  if (threadIdx.x < K) {
    float s = 0;
#pragma unroll
    for (int w = 0; w < 8; ++w) s += warp_sums[w][threadIdx.x];
    // Hessian entries first, gradient at 21.. (both layouts of the reference's buffers)
    const int slot = threadIdx.x < H ? threadIdx.x : 21 + (threadIdx.x - H);
    s_system[slot] = s;
    st->system[slot] = s;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    t3 = gtime();
    st->ticket = 0;
    if (solve) {
      float x[6];
      track_solve<P>(s_system, s_system + 21, x);
      // pose in registers: one round trip in, one out
      float T[16], Ti[16];
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        T[i] = st->Twc[i];
        Ti[i] = st->Twc_inv[i];
      }
      track_apply_update(x, T, Ti);
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        st->Twc[i] = T[i];
        st->Twc_inv[i] = Ti[i];
      }
      float nn = 0;
      for (int i = 0; i < P; ++i) nn += x[i] * x[i];
      const float norm = sqrtf(nn);
      st->update_norm = norm;
      // Tracker::ComputeUpdate / UpdateState (src/tracker.cpp:147-153): converged -> iteration = max, then ++
      int it = st->iteration;
      if (norm < 1e-6f) it = st->max_iterations;
      ++it;
      st->iteration = it;
      st->evaluations += 1;
      if (it >= st->max_iterations) st->done = 1;
    }
    st->stamps[0] = t0; st->stamps[1] = t1; st->stamps[2] = t2; st->stamps[3] = t3; st->stamps[4] = gtime();
  }
}

}  // namespace vk

using namespace vk;

static int fill_images(TrackImages& im, const vkFrame* key, const vkFrame* frame) {
  if (!key || !frame || !key->depth || !key->normal || !frame->depth || !frame->normal) return VK_E_INVALID;
  if (key->depth_w <= 0 || key->depth_h <= 0 || frame->depth_w <= 0 || frame->depth_h <= 0) return VK_E_INVALID;
  im.key_depth = (const float*)key->depth;
  im.key_normal = (const float*)key->normal;
  im.depth = (const float*)frame->depth;
  im.normal = (const float*)frame->normal;
  im.kw = key->depth_w; im.kh = key->depth_h;
  im.w = frame->depth_w; im.h = frame->depth_h;
  im.key_proj = key->depth_proj;
  im.proj = frame->depth_proj;
  rows34(key->depth_to_world.m, im.Twm);
  rows34(key->depth_to_world.inv, im.Twm_inv);
  return 0;
}

static int track_grid(const TrackImages& im) {
  // 84 registers x 256 threads: three CTAs per SM are resident.  The tiles are spread evenly over the
  // fewest passes that fit (a grid of exactly `cap` CTAs would leave a nearly empty last pass).
  const int tiles = ((im.w + 31) / 32) * ((im.h + 7) / 8);
  int cap = 3 * sm_count();
  if (cap > 512) cap = 512;  // the fold of the CTA sums takes two CTAs per thread
  const int passes = (tiles + cap - 1) / cap;
  return (tiles + passes - 1) / passes;
}

extern "C" int vk_downsample_frame(const vkFrame* src, vkFrame* dst, void* stream) {
  if (!src || !dst) return VK_E_INVALID;
  const bool has_d = src->depth != nullptr, has_n = src->normal != nullptr, has_c = src->color != nullptr;
  if ((has_d && !dst->depth) || (has_n && !dst->normal) || (has_c && !dst->color)) return VK_E_INVALID;
  if ((has_d || has_n) && (src->depth_w <= 0 || src->depth_h <= 0 || (src->depth_w & 1) || (src->depth_h & 1)))
    return VK_E_INVALID;  // "even image dimensions required" (src/image.cu:185-186)
  if (has_c && (src->color_w <= 0 || src->color_h <= 0 || (src->color_w & 1) || (src->color_h & 1))) return VK_E_INVALID;
  dst->depth_w = src->depth_w / 2; dst->depth_h = src->depth_h / 2;
  dst->color_w = src->color_w / 2; dst->color_h = src->color_h / 2;
  // src/frame.cpp:51-56
  dst->depth_proj.fx = src->depth_proj.fx / 2; dst->depth_proj.fy = src->depth_proj.fy / 2;
  dst->depth_proj.cx = src->depth_proj.cx / 2; dst->depth_proj.cy = src->depth_proj.cy / 2;
  dst->color_proj.fx = src->color_proj.fx / 2; dst->color_proj.fy = src->color_proj.fy / 2;
  dst->color_proj.cx = src->color_proj.cx / 2; dst->color_proj.cy = src->color_proj.cy / 2;
  dst->depth_to_world = src->depth_to_world;
  dst->depth_to_color = src->depth_to_color;
  const int nd = (has_d || has_n) ? dst->depth_w * dst->depth_h : 0, nc = has_c ? dst->color_w * dst->color_h : 0;
  const int total = nd > nc ? nd : nc;
  if (total == 0) return 0;
  int grid = (total + 255) / 256;
  const int cap = 8 * sm_count();
  if (grid > cap) grid = cap;
  vk_downsample_frame_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(
      (const float*)src->depth, (const float*)src->normal, (has_d || has_n) ? src->depth_w : 0,
      (has_d || has_n) ? src->depth_h : 0, (float*)dst->depth, (float*)dst->normal, (const float*)src->color,
      has_c ? src->color_w : 0, has_c ? src->color_h : 0, (float*)dst->color, 1, 1, 0);
  VK_COUNT_LAUNCH();
  return (int)cudaGetLastError();
}

extern "C" int vk_downsample_image(const float* src, int src_w, int src_h, int channels, int nearest, float* dst,
                                   void* stream) {
  if (!src || !dst || src_w <= 0 || src_h <= 0 || (src_w & 1) || (src_h & 1) || (channels != 1 && channels != 3))
    return VK_E_INVALID;
  const int total = (src_w / 2) * (src_h / 2);
  int grid = (total + 255) / 256;
  const int cap = 8 * sm_count();
  if (grid > cap) grid = cap;
  if (channels == 1)
    vk_downsample_frame_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(src, nullptr, src_w, src_h, dst, nullptr, nullptr,
                                                                       0, 0, nullptr, nearest, 1, 0);
  else
    vk_downsample_frame_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(nullptr, src, src_w, src_h, nullptr, dst, nullptr,
                                                                       0, 0, nullptr, 1, nearest, 0);
  VK_COUNT_LAUNCH();
  return (int)cudaGetLastError();
}

static int track_pixels(const vkFrame* key, const vkFrame* frame, int translation, float* out, int mode, void* stream) {
  TrackImages im;
  int rc = fill_images(im, key, frame);
  if (rc || !out) return VK_E_INVALID;
  cudaStream_t s = (cudaStream_t)stream;
  float* pose = nullptr;
  if ((rc = (int)cudaMallocAsync(&pose, 64, s))) return rc;
  if ((rc = (int)cudaMemcpyAsync(pose, frame->depth_to_world.m, 64, cudaMemcpyHostToDevice, s))) {
    cudaFreeAsync(pose, s);
    return rc;
  }
  const int n = im.w * im.h;
  int grid = (n + 255) / 256;
  const int cap = 8 * sm_count();
  if (grid > cap) grid = cap;
  const bool strict = g_default_math_mode.load() == VK_MATH_STRICT;
  if (mode == 0) {
    if (strict) vk_track_pixels_kernel<false, 0><<<grid, 256, 0, s>>>(im, pose, translation, out);
    else vk_track_pixels_kernel<true, 0><<<grid, 256, 0, s>>>(im, pose, translation, out);
  } else {
    if (strict) vk_track_pixels_kernel<false, 1><<<grid, 256, 0, s>>>(im, pose, translation, out);
    else vk_track_pixels_kernel<true, 1><<<grid, 256, 0, s>>>(im, pose, translation, out);
  }
  VK_COUNT_LAUNCH();
  rc = (int)cudaGetLastError();
  cudaFreeAsync(pose, s);
  return rc;
}

extern "C" int vk_track_residuals(const vkFrame* keyframe, const vkFrame* frame, float* residuals, void* stream) {
  return track_pixels(keyframe, frame, 0, residuals, 0, stream);
}

extern "C" int vk_track_jacobian(const vkFrame* keyframe, const vkFrame* frame, int translation_enabled, float* jacobian,
                                 void* stream) {
  return track_pixels(keyframe, frame, translation_enabled, jacobian, 1, stream);
}

static void launch_iter(const TrackImages& im, TrackState* st, float* partials, int translation, int solve, int grid,
                        cudaStream_t s) {
  const bool strict = g_default_math_mode.load() == VK_MATH_STRICT;
  if (translation) {
    if (strict) launch_chained(vk_track_iter_kernel<false, true>, dim3(grid), dim3(256), 0, s, im, st, partials, solve);
    else launch_chained(vk_track_iter_kernel<true, true>, dim3(grid), dim3(256), 0, s, im, st, partials, solve);
  } else {
    if (strict) launch_chained(vk_track_iter_kernel<false, false>, dim3(grid), dim3(256), 0, s, im, st, partials, solve);
    else launch_chained(vk_track_iter_kernel<true, false>, dim3(grid), dim3(256), 0, s, im, st, partials, solve);
  }
  VK_COUNT_LAUNCH();
}

static int init_state(TrackState* dev, const vkFrame* frame, int max_iterations, cudaStream_t s) {
  TrackState h;
  memset(&h, 0, sizeof(h));
  memcpy(h.Twc, frame->depth_to_world.m, 64);
  memcpy(h.Twc_inv, frame->depth_to_world.inv, 64);
  h.max_iterations = max_iterations;
  // pageable source: the copy is staged before the call returns, so the stack object may go away
  return (int)cudaMemcpyAsync(dev, &h, sizeof(h), cudaMemcpyHostToDevice, s);
}

extern "C" int vk_track_workspace_floats(void) { return VK_TRACK_STATE_FLOATS + 27 * 3 * sm_count(); }

extern "C" int vk_track_system(const vkFrame* keyframe, const vkFrame* frame, int translation_enabled, float* workspace,
                               float* hessian, float* gradient, void* stream) {
  TrackImages im;
  int rc = fill_images(im, keyframe, frame);
  if (rc || !workspace || !hessian || !gradient) return VK_E_INVALID;
  cudaStream_t s = (cudaStream_t)stream;
  TrackState* st = (TrackState*)workspace;
  if ((rc = init_state(st, frame, 1, s))) return rc;
  launch_iter(im, st, workspace + VK_TRACK_STATE_FLOATS, translation_enabled, 0, track_grid(im), s);
  if ((rc = (int)cudaGetLastError())) return rc;
  const int P = translation_enabled ? 6 : 3, H = translation_enabled ? 21 : 6;
  if ((rc = (int)cudaMemcpyAsync(hessian, st->system, H * sizeof(float), cudaMemcpyDeviceToDevice, s))) return rc;
  return (int)cudaMemcpyAsync(gradient, st->system + 21, P * sizeof(float), cudaMemcpyDeviceToDevice, s);
}

extern "C" int vk_track_depth(const vkFrame* keyframe, vkFrame* frame, int translation_enabled, int max_iterations,
                              float* workspace, int* iterations, float* last_update_norm, void* stream) {
  TrackImages im;
  int rc = fill_images(im, keyframe, frame);
  if (rc || !workspace || max_iterations < 0) return VK_E_INVALID;
  cudaStream_t s = (cudaStream_t)stream;
  TrackState* st = (TrackState*)workspace;
  if ((rc = init_state(st, frame, max_iterations, s))) return rc;
  const int grid = track_grid(im);
  for (int i = 0; i < max_iterations; ++i) launch_iter(im, st, workspace + VK_TRACK_STATE_FLOATS, translation_enabled, 1, grid, s);
  if ((rc = (int)cudaGetLastError())) return rc;
  TrackState h;
  if ((rc = (int)cudaMemcpyAsync(&h, st, sizeof(h), cudaMemcpyDeviceToHost, s))) return rc;
  if ((rc = (int)cudaStreamSynchronize(s))) return rc;
  memcpy(frame->depth_to_world.m, h.Twc, 64);
  memcpy(frame->depth_to_world.inv, h.Twc_inv, 64);
  if (iterations) *iterations = h.evaluations;
  if (last_update_norm) *last_update_norm = h.update_norm;
  return 0;
}
