// vk_integrate.cu — DepthIntegrator / ColorIntegrator / LightIntegrator for sm_100a.
//
// Reference: one CUDA block (8,8,8) per visible voxel block, one thread per voxel,
// scalar 4-byte accesses at a 20-byte stride, and a second full pass over the same
// voxel tiles for colour (src/depth_integrator.cu:17-80, src/color_integrator.cu:82-135,
// src/light_integrator.cu:170-250).
//
// Here: persistent CTAs of 128 threads grid-stride over the device-resident visible
// list (no host read-back of the count).  A thread owns 4 consecutive voxels of a row
// (80 bytes = 5 aligned 16-byte chunks of the AoS tile), decides from the projection
// and the depth image alone whether any of them can change, and only then moves the
// 80 bytes with 128-bit loads; depth and colour updates happen on the registers in
// one pass, and only the 16-byte chunks that changed are written back.
// The loop is software-pipelined: the image tests of the next block run under the tile loads of
// the current one (see the kernel).  A cp.async double-buffered variant that staged the 80-byte
// pieces through shared memory one block ahead was measured on B200 and was slower (68 us against
// 55 us per frame on the bench workload).
#include "vk_math.cuh"
#include "vk_launch.h"
#include <cfloat>

namespace vk {

enum { MODE_DEPTH = 0, MODE_COLOR = 1, MODE_LIGHT = 2 };

struct IntegrateParams {
  const int32_t* visible;
  const int32_t* count;
  const vkHashEntry* entries;
  vkVoxel* voxels;
  float* sdf;   // vkVolume::sdf, the block-major mirror of Voxel::distance the ray marcher reads
  int push_count, push_rank;          // sharded volume: the peers' copies of this rank's mirror (peer memory),
  float* push_sdf[VK_MAX_SHARDS];     // written together with the mirror; push_count = 0: none
  const float* depth;
  const float* color;
  const float* normal;
  const float* mask;
  unsigned long long* updated;  // counters[VK_CNT_UPDATED..+1]: voxels written since reset
  int dw, dh, cw, ch;
  vkProjection dproj, cproj;
  float Tdw[12], Tcw[12], Tcd[12];
  float voxel_length, block_length;
  Div trunc;
  float min_depth, max_depth, max_dw, max_cw;
  vkLight light;
};

// FMA map (reference release SASS, see vk_math.cuh):
//   Xw      = L*o + l*(v+0.5)          FMUL o*L ; FFMA (v+0.5)*l + .      depth_integrator.cu:36-39
//   D'      = (w*D + cur) / (w+1)      FFMA D*w + cur ; IEEE div         depth_integrator.cu:68-76
//   C'      = (cw*C + c) * (1/(cw+1))  FFMA C*cw + c ; FMUL              color_integrator.cu:121-126
//   C'light = (cw*C + c*(1/s)) * ...   FMUL C*cw ; FFMA c*(1/s) + . ; FMUL  light_integrator.cu:238-241
// Outcome of the geometric tests of a thread's 4 voxels against the images: everything that can be
// decided without touching the voxel tile.
template <int MODE>
struct VoxelTests {
  uint32_t dmask;  // bit k: depth update applies to voxel k
  uint32_t cmask;  // bit k: colour-side geometric test passes
  float dcur[4];   // depth - Xd.z
  int cidx[MODE == MODE_DEPTH ? 1 : 4];   // colour pixel
  int didx[MODE == MODE_LIGHT ? 4 : 1];   // depth pixel (light mode)
  Vec3 Xc[MODE == MODE_LIGHT ? 4 : 1];    // voxel centre in the colour camera (light mode)
};

template <int MODE, bool F>
__device__ __forceinline__ void test_voxels(const IntegrateParams& p, const Mat34& Tdw, const Mat34& Tcw,
                                            const int4 ent, int vx0, float fy, float fz, VoxelTests<MODE>& r) {
  int iox, ioy, ioz;
  entry_origin(ent, iox, ioy, ioz);
  const float Xy = FP<F>::mad(fy, p.voxel_length, mul((float)ioy, p.block_length));
  const float Xz = FP<F>::mad(fz, p.voxel_length, mul((float)ioz, p.block_length));
  const float bxl = mul((float)iox, p.block_length);
  r.dmask = 0;
  r.cmask = 0;
  // all four depth gathers are issued before any of them is looked at (pixel 0 stands in for
  // voxels that project outside the image; its value is not used)
  float Xx[4], zd[4], d[4];
  int di[4];
  uint32_t inside = 0;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    Xx[k] = FP<F>::mad(add((float)(vx0 + k), 0.5f), p.voxel_length, bxl);
    const Vec3 Xd = xform1<F>(Tdw, Xx[k], Xy, Xz);
    float u, v;
    project<F>(p.dproj, Xd, u, v);
    const bool din = (u >= 0 && u < (float)p.dw && v >= 0 && v < (float)p.dh);
    zd[k] = Xd.z;
    di[k] = din ? (int)v * p.dw + (int)u : 0;
    if (din) inside |= 1u << k;
  }
#pragma unroll
  for (int k = 0; k < 4; ++k) d[k] = __ldg(p.depth + di[k]);
  float m[4];
  if (MODE == MODE_LIGHT) {
#pragma unroll
    for (int k = 0; k < 4; ++k) m[k] = __ldg(p.mask + di[k]);
  }
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const bool din = (inside >> k) & 1u;
    const float s = sub(d[k], zd[k]);
    const bool pass = din && !(d[k] < p.min_depth || d[k] > p.max_depth) && s > -p.trunc.d;
    r.dcur[k] = pass ? s : 0.f;
    if (pass) r.dmask |= 1u << k;
    if (MODE == MODE_COLOR) {
      const Vec3 Xc = xform1<F>(Tcw, Xx[k], Xy, Xz);
      float cu, cv;
      project<F>(p.cproj, Xc, cu, cv);
      const bool ok = (cu >= 0 && cu < (float)p.cw && cv >= 0 && cv < (float)p.ch);
      r.cidx[k] = ok ? (int)cv * p.cw + (int)cu : 0;
      if (ok) r.cmask |= 1u << k;
    } else if (MODE == MODE_LIGHT) {
      // both projections are tested against the depth image size (src/light_integrator.cu:207-215,335-336)
      const Vec3 Xc = xform1<F>(Tcw, Xx[k], Xy, Xz);
      r.Xc[k] = Xc;
      float cu, cv;
      project<F>(p.cproj, Xc, cu, cv);
      const bool ok = din && (cu >= 0 && cu < (float)p.dw && cv >= 0 && cv < (float)p.dh) && m[k] > 0.5f;
      r.didx[k] = di[k];
      r.cidx[k] = ok ? (int)cv * p.dw + (int)cu : 0;
      if (ok) r.cmask |= 1u << k;
    }
  }
}

// Tile movement.  A thread's 4 voxels are 80 contiguous bytes, so direct per-thread 128-bit
// accesses run at an 80-byte stride: every warp instruction touches 20 cache lines, and a plain
// read of the same scattered tiles then tops out near 4.3 TB/s on B200 against 8 TB/s for
// lane-contiguous accesses (tools/microbench/tile_copy.cu).  So a warp moves its 2560 bytes of
// the tile lane-contiguously (5 x 512 B) through a private shared-memory buffer: cp.async in,
// per-thread 80-byte rows out of shared memory (an 80-byte stride is conflict-free for 128-bit
// shared accesses), and the changed 16-byte chunks back the same way.  Only chunks whose owner
// thread has something to update are fetched, only changed chunks are stored.
constexpr int kWarpChunks = 160;  // 16-byte chunks of a warp's part of a tile (32 threads x 5)

__device__ __forceinline__ void cp_async16(uint4* smem_dst, const uint4* gsrc) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)),
               "l"(gsrc));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

// The update of a thread's 4 voxels (80 bytes = 20 words in w) given the outcome of the image tests: TSDF
// running average and, in the same pass, colour / albedo; returns the mask of changed words.  Also keeps the
// block's distance mirror (and its copies on the peers of a sharded volume) in step.
template <int MODE, bool F>
__device__ __forceinline__ uint32_t update_piece(const IntegrateParams& p, const VoxelTests<MODE>& q, uint32_t (&w)[20],
                                                 int data, int t) {
  uint32_t dirty = 0;  // bit per word of the thread's 80 bytes
  const uint4 d_old = make_uint4(w[0], w[5], w[10], w[15]);
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    float D = __uint_as_float(w[5 * k]);
    int dwt = (int)(int16_t)(w[5 * k + 4] & 0xffff);
    int cwt = (int)(int16_t)(w[5 * k + 4] >> 16);
    if ((q.dmask >> k) & 1u) {
      // src/depth_integrator.cu:68-76
      const float c = vmin(1.0f, divc(q.dcur[k], p.trunc));
      const float wn = (float)(dwt + 1);
      D = fdiv(FP<F>::mad(D, (float)dwt, c), wn);
      dwt = (int)(short)vmin(p.max_dw, wn);
      w[5 * k] = __float_as_uint(D);
      dirty |= (1u << (5 * k)) | (1u << (5 * k + 4));
    }
    if (MODE != MODE_DEPTH) {
      if (((q.cmask >> k) & 1u) && fabsf(D) < 1.0f) {
        const size_t ci = 3 * (size_t)q.cidx[MODE == MODE_DEPTH ? 0 : k];
        const float cr = __ldg(p.color + ci + 0);
        const float cg = __ldg(p.color + ci + 1);
        const float cb = __ldg(p.color + ci + 2);
        const float fw = (float)cwt;
        const float wn = (float)(cwt + 1);
        const float r0 = __uint_as_float(w[5 * k + 1]);
        const float g0 = __uint_as_float(w[5 * k + 2]);
        const float b0 = __uint_as_float(w[5 * k + 3]);
        bool ok = true;
        float sr, sg, sb;
        if (MODE == MODE_LIGHT) {
          // src/light_integrator.cu:229-241, include/vulcan/light.h:52-60
          const size_t ni = 3 * (size_t)q.didx[MODE == MODE_LIGHT ? k : 0];
          const float nx = __ldg(p.normal + ni + 0);
          const float ny = __ldg(p.normal + ni + 1);
          const float nz = __ldg(p.normal + ni + 2);
          Mat34 Tcd;
          Tcd.load(p.Tcd);
          const Vec3 Xcn = xform0<F>(Tcd, nx, ny, nz);
          const Vec3 Xc = q.Xc[MODE == MODE_LIGHT ? k : 0];
          const Vec3 delta = {sub(p.light.position[0], Xc.x), sub(p.light.position[1], Xc.y),
                              sub(p.light.position[2], Xc.z)};
          const float dist2 = dot3<F>(delta, delta);
          const float invn = rcp(__fsqrt_rn(dist2));
          const Vec3 dirn = {mul(delta.x, invn), mul(delta.y, invn), mul(delta.z, invn)};
          const float cos_theta = dot3<F>(Xcn, dirn);
          const float shading = fdiv(mul(p.light.intensity, cos_theta), dist2);
          ok = shading > 0.05f;
          const float inv_s = rcp(shading);
          sr = FP<F>::mad(cr, inv_s, mul(r0, fw));
          sg = FP<F>::mad(cg, inv_s, mul(g0, fw));
          sb = FP<F>::mad(cb, inv_s, mul(b0, fw));
        } else {
          sr = FP<F>::mad(r0, fw, cr);
          sg = FP<F>::mad(g0, fw, cg);
          sb = FP<F>::mad(b0, fw, cb);
        }
        if (ok) {
          const float inv = rcp(wn);
          w[5 * k + 1] = __float_as_uint(mul(sr, inv));
          w[5 * k + 2] = __float_as_uint(mul(sg, inv));
          w[5 * k + 3] = __float_as_uint(mul(sb, inv));
          cwt = (int)(short)vmin(p.max_cw, wn);
          dirty |= (0xeu << (5 * k)) | (1u << (5 * k + 4));
        }
      }
    }
    w[5 * k + 4] = ((uint32_t)(uint16_t)(int16_t)dwt) | ((uint32_t)(uint16_t)(int16_t)cwt << 16);
  }
  // the thread's 4 distances are one aligned 16-byte piece of the block's distance mirror
  // (voxel index 4t): a warp writes 512 contiguous bytes
  // - written only when a distance really changed: free-space voxels in front of the surface keep
  // D = 1 under the running average, and those are half of the updates
  const uint4 d4 = make_uint4(w[0], w[5], w[10], w[15]);
  if (q.dmask && ((d4.x ^ d_old.x) | (d4.y ^ d_old.y) | (d4.z ^ d_old.z) | (d4.w ^ d_old.w))) {
    reinterpret_cast<uint4*>(p.sdf + (size_t)data * VK_BLOCK_VOXELS)[t] = d4;
    for (int r = 0; r < p.push_count; ++r)   // posted stores over NVLink; read after the next handshake
      if (r != p.push_rank) reinterpret_cast<uint4*>(p.push_sdf[r] + (size_t)data * VK_BLOCK_VOXELS)[t] = d4;
  }
  return dirty;
}

// Software pipeline over the CTA's blocks: the tile fetch of block i is issued, then the image
// tests of block i+1 run (projection + depth gathers, independent of the tile), then block i is
// updated in registers and its changed chunks are stored.  The entry of block i+2 is in flight
// meanwhile.  One iteration costs max(tile latency, gather latency) instead of their sum.
// resident CTAs per SM the register allocation is held to (measured on B200: the depth kernel
// lands at 55 registers = 9 CTAs/SM; forcing 10 CTAs spills and is slower; colour 70 registers / 7 CTAs;
// light 80 registers / 6 CTAs with 72 bytes of spills, 7 % faster than 110 registers / 4 CTAs)
template <int MODE>
constexpr int integrate_min_ctas() { return MODE == MODE_DEPTH ? 8 : (MODE == MODE_COLOR ? 7 : 6); }

template <int MODE, bool F>
__global__ void __launch_bounds__(128, integrate_min_ctas<MODE>()) vk_integrate_kernel(const IntegrateParams p) {
  pdl_wait();     // the previous kernel of the chain has completed (no-op for a plain launch)
  pdl_trigger();  // the next one may be scheduled as this grid drains
  const int n = *p.count;
  const int t = threadIdx.x;
  const int vz = t >> 4, vy = (t >> 1) & 7, vx0 = (t & 1) * 4;
  Mat34 Tdw, Tcw;
  Tdw.load(p.Tdw);
  if (MODE != MODE_DEPTH) Tcw.load(p.Tcw);
  unsigned n_updated = 0;
  const float fy = add((float)vy, 0.5f), fz = add((float)vz, 0.5f);
  const int G = gridDim.x;
  __shared__ uint4 s_tile[4][kWarpChunks];
  const int lane = t & 31, warp = t >> 5;
  uint4* const stage = s_tile[warp];

  int bi = blockIdx.x;
  if (bi >= n) return;
  // list slots are read three blocks ahead and entries two blocks ahead, so that neither load
  // is waited for where it is issued
  int4 ent = ld_entry(p.entries + __ldg(p.visible + bi));
  int4 ent_n = make_int4(0, 0, 0, 0), ent_nn = make_int4(0, 0, 0, 0);
  if (bi + G < n) ent_n = ld_entry(p.entries + __ldg(p.visible + bi + G));
  int slot_nn = (bi + 2 * G < n) ? __ldg(p.visible + bi + 2 * G) : 0;
  VoxelTests<MODE> cur, nxt;
  test_voxels<MODE, F>(p, Tdw, Tcw, ent, vx0, fy, fz, cur);

  for (; bi < n; bi += G) {
    const bool need = (cur.dmask | cur.cmask) != 0;
    const uint32_t need_mask = __ballot_sync(0xffffffffu, need);
    // the warp's 2560 bytes of the tile
    uint4* const tile = reinterpret_cast<uint4*>(reinterpret_cast<char*>(p.voxels) +
                                                 (size_t)ent.z * (VK_BLOCK_VOXELS * sizeof(vkVoxel)) +
                                                 (size_t)warp * (kWarpChunks * 16));
    if (need_mask) {
#pragma unroll
      for (int c = 0; c < 5; ++c) {
        const int j = c * 32 + lane;         // chunk of the warp's span
        const int owner = (j * 205) >> 10;   // j / 5: the lane whose voxels live in this chunk
        if ((need_mask >> owner) & 1u) cp_async16(stage + j, tile + j);
      }
    }
    cp_async_commit();
    if (bi + 2 * G < n) ent_nn = ld_entry(p.entries + slot_nn);
    if (bi + 3 * G < n) slot_nn = __ldg(p.visible + bi + 3 * G);
    if (bi + G < n) test_voxels<MODE, F>(p, Tdw, Tcw, ent_n, vx0, fy, fz, nxt);
    cp_async_wait_all();
    __syncwarp();

    uint32_t dirty = 0;  // bit per word of the thread's 80 bytes
    if (need) {
      uint32_t w[20];
#pragma unroll
      for (int c = 0; c < 5; ++c) {
        const uint4 q = stage[lane * 5 + c];
        w[4 * c + 0] = q.x; w[4 * c + 1] = q.y; w[4 * c + 2] = q.z; w[4 * c + 3] = q.w;
      }
      dirty = update_piece<MODE, F>(p, cur, w, ent.z, t);
#pragma unroll
      for (int c = 0; c < 5; ++c) {
        if ((dirty >> (4 * c)) & 0xfu)
          stage[lane * 5 + c] = make_uint4(w[4 * c], w[4 * c + 1], w[4 * c + 2], w[4 * c + 3]);
      }
      // voxels written = voxels whose weight word is dirty (bit 5k+4)
      n_updated += __popc(dirty & 0x84210u);
    }
    // changed chunks go back lane-contiguously
    uint32_t chunk_dirty = 0;  // bit c: the thread's chunk c changed
#pragma unroll
    for (int c = 0; c < 5; ++c) chunk_dirty |= ((dirty >> (4 * c)) & 0xfu) ? (1u << c) : 0u;
    if (__any_sync(0xffffffffu, chunk_dirty != 0)) {
      __syncwarp();
#pragma unroll
      for (int c = 0; c < 5; ++c) {
        const int j = c * 32 + lane;
        const int owner = (j * 205) >> 10;
        const uint32_t m = __shfl_sync(0xffffffffu, chunk_dirty, owner);
        if ((m >> (j - owner * 5)) & 1u) tile[j] = stage[j];
      }
    }
    __syncwarp();  // the buffer is refilled by the next iteration
    cur = nxt;
    ent = ent_n;
    ent_n = ent_nn;
  }
  n_updated = __reduce_add_sync(0xffffffffu, n_updated);
  if ((threadIdx.x & 31) == 0 && n_updated) atomicAdd(p.updated, (unsigned long long)n_updated);
}

// ---- bulk-copy variant (TMA 1-D) --------------------------------------------------------------------
// The same update, with the tile movement handed to the copy engine: a block's 10 240-byte tile is
// fetched whole by ONE cp.async.bulk (global -> shared, completion on an mbarrier) issued by one thread,
// kBulkStages tiles deep - the tiles of blocks i+1 and i+2 are in flight while block i is updated - so no
// thread spends issue slots or registers on loads.  The price: 100 % of every visited tile crosses the
// memory system, where the selective path above fetches the ~58 % of 16-byte pieces some voxel needs.
// Changed pieces go back like above (lane-contiguous 16-byte stores of the dirty chunks only).
// Selected with tuning knob VK_TUNE_INTEGRATE_BULK; A/B numbers in DESIGN.md 4.2.
constexpr int kBulkStages = 3;
constexpr int kTileBytes = VK_BLOCK_VOXELS * (int)sizeof(vkVoxel);   // 10 240
constexpr int kTileChunks = kTileBytes / 16;                         // 640

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_load(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(smem_dst)),
               "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

template <int MODE, bool F>
__global__ void __launch_bounds__(128, 6) vk_integrate_bulk_kernel(const IntegrateParams p) {
  pdl_wait();     // the previous kernel of the chain has completed (no-op for a plain launch)
  pdl_trigger();  // the next one may be scheduled as this grid drains
  extern __shared__ __align__(128) unsigned char s_raw[];
  uint4* const tiles = reinterpret_cast<uint4*>(s_raw);                                           // [kBulkStages][640]
  uint64_t* const full = reinterpret_cast<uint64_t*>(s_raw + (size_t)kBulkStages * kTileBytes);   // [kBulkStages]
  const int n = *p.count;
  const int t = threadIdx.x;
  const int lane = t & 31, warp = t >> 5;
  const int vz = t >> 4, vy = (t >> 1) & 7, vx0 = (t & 1) * 4;
  const float fy = add((float)vy, 0.5f), fz = add((float)vz, 0.5f);
  const int G = gridDim.x;
  Mat34 Tdw, Tcw;
  Tdw.load(p.Tdw);
  if (MODE != MODE_DEPTH) Tcw.load(p.Tcw);
  if (t == 0) {
#pragma unroll
    for (int s = 0; s < kBulkStages; ++s) mbar_init(full + s, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  int bi = blockIdx.x;
  if (bi >= n) return;
  unsigned n_updated = 0;
  auto tile_of = [&](int data) {
    return reinterpret_cast<const char*>(p.voxels) + (size_t)data * kTileBytes;
  };
  // entries one, two and three blocks ahead (every thread holds them: the loads are broadcasts)
  int4 ent = ld_entry(p.entries + __ldg(p.visible + bi));
  int4 ent_n = make_int4(0, 0, 0, 0), ent_nn = make_int4(0, 0, 0, 0);
  if (bi + G < n) ent_n = ld_entry(p.entries + __ldg(p.visible + bi + G));
  if (bi + 2 * G < n) ent_nn = ld_entry(p.entries + __ldg(p.visible + bi + 2 * G));
  int slot_nnn = (bi + 3 * G < n) ? __ldg(p.visible + bi + 3 * G) : 0;
  if (t == 0) {   // prologue: the first two tiles
    mbar_expect_tx(full + 0, kTileBytes);
    bulk_load(tiles, tile_of(ent.z), kTileBytes, full + 0);
    if (bi + G < n) {
      mbar_expect_tx(full + 1, kTileBytes);
      bulk_load(tiles + kTileChunks, tile_of(ent_n.z), kTileBytes, full + 1);
    }
  }
  VoxelTests<MODE> cur, nxt;
  test_voxels<MODE, F>(p, Tdw, Tcw, ent, vx0, fy, fz, cur);
  int k = 0;
  for (; bi < n; bi += G, ++k) {
    const int stage = k % kBulkStages;
    // tile of block k + 2 into the stage block k - 1 used (all of its readers passed the barrier at the end
    // of the previous iteration; the proxy fence orders their generic-proxy accesses before the copy engine's)
    if (t == 0 && bi + 2 * G < n) {
      const int s2 = (k + 2) % kBulkStages;
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      mbar_expect_tx(full + s2, kTileBytes);
      bulk_load(tiles + (size_t)s2 * kTileChunks, tile_of(ent_nn.z), kTileBytes, full + s2);
    }
    int4 ent_nnn = make_int4(0, 0, 0, 0);
    if (bi + 3 * G < n) ent_nnn = ld_entry(p.entries + slot_nnn);
    if (bi + 4 * G < n) slot_nnn = __ldg(p.visible + bi + 4 * G);
    if (bi + G < n) test_voxels<MODE, F>(p, Tdw, Tcw, ent_n, vx0, fy, fz, nxt);
    mbar_wait(full + stage, (uint32_t)((k / kBulkStages) & 1));
    uint4* const stage_tile = tiles + (size_t)stage * kTileChunks;
    uint4* const mine = stage_tile + warp * kWarpChunks;   // the warp's 2560 bytes
    uint4* const gtile = reinterpret_cast<uint4*>(reinterpret_cast<char*>(p.voxels) + (size_t)ent.z * kTileBytes) +
                         warp * kWarpChunks;
    const bool need = (cur.dmask | cur.cmask) != 0;
    uint32_t dirty = 0;
    if (need) {
      uint32_t w[20];
#pragma unroll
      for (int c = 0; c < 5; ++c) {
        const uint4 q = mine[lane * 5 + c];
        w[4 * c + 0] = q.x; w[4 * c + 1] = q.y; w[4 * c + 2] = q.z; w[4 * c + 3] = q.w;
      }
      dirty = update_piece<MODE, F>(p, cur, w, ent.z, t);
#pragma unroll
      for (int c = 0; c < 5; ++c) {
        if ((dirty >> (4 * c)) & 0xfu) mine[lane * 5 + c] = make_uint4(w[4 * c], w[4 * c + 1], w[4 * c + 2], w[4 * c + 3]);
      }
      n_updated += __popc(dirty & 0x84210u);
    }
    uint32_t chunk_dirty = 0;
#pragma unroll
    for (int c = 0; c < 5; ++c) chunk_dirty |= ((dirty >> (4 * c)) & 0xfu) ? (1u << c) : 0u;
    if (__any_sync(0xffffffffu, chunk_dirty != 0)) {
      __syncwarp();
#pragma unroll
      for (int c = 0; c < 5; ++c) {
        const int j = c * 32 + lane;
        const int owner = (j * 205) >> 10;
        const uint32_t m = __shfl_sync(0xffffffffu, chunk_dirty, owner);
        if ((m >> (j - owner * 5)) & 1u) gtile[j] = mine[j];
      }
    }
    __syncthreads();   // every warp is done with this stage before it is refilled (next iteration)
    cur = nxt;
    ent = ent_n;
    ent_n = ent_nn;
    ent_nn = ent_nnn;
  }
  n_updated = __reduce_add_sync(0xffffffffu, n_updated);
  if ((threadIdx.x & 31) == 0 && n_updated) atomicAdd(p.updated, (unsigned long long)n_updated);
}

// ComputeFrameMaskKernel<16,3> (src/light_integrator.cu:17-103) including Q14: the
// shared tile is filled from image offset -1 (not -KERNEL_SIZE) while the window is read
// at [-3,+3] around (t + 3), so the window covers image offsets [-1,+5] of the pixel.
__global__ void __launch_bounds__(256) vk_frame_mask_kernel(int width, int height, const float* depths,
                                                            const float* colors, float threshold, float* mask) {
  pdl_wait();     // the previous kernel of the chain has completed (no-op for a plain launch)
  pdl_trigger();  // the next one may be scheduled as this grid drains
  constexpr int BD = 16, KS = 3, DIM = BD + 2 * KS;
  __shared__ float buf[DIM * DIM];
  const int x = blockIdx.x * BD + threadIdx.x;
  const int y = blockIdx.y * BD + threadIdx.y;
  for (int s = threadIdx.y * BD + threadIdx.x; s < DIM * DIM; s += BD * BD) {
    const int vx = blockIdx.x * BD - 1 + (s % DIM);
    const int vy = blockIdx.y * BD - 1 + (s / DIM);
    float d = 0;
    if (vx >= 0 && vx < width && vy >= 0 && vy < height) d = __ldg(depths + vy * width + vx);
    buf[s] = d;
  }
  __syncthreads();
  if (x < width && y < height) {
    const int index = y * width + x;
    const float r = colors[3 * (size_t)index], g = colors[3 * (size_t)index + 1], b = colors[3 * (size_t)index + 2];
    if (r < 0.02f || r > 0.98f || g < 0.02f || g > 0.98f || b < 0.02f || b > 0.98f) {
      mask[index] = 0.0f;
      return;
    }
    float dmin = +FLT_MAX, dmax = -FLT_MAX;
    const int cx = threadIdx.x + KS, cy = threadIdx.y + KS;
    for (int i = -KS; i <= KS; ++i)
      for (int j = -KS; j <= KS; ++j) {
        const float d = buf[(cy + i) * DIM + (cx + j)];
        dmin = fminf(d, dmin);
        dmax = fmaxf(d, dmax);
      }
    mask[index] = (sub(dmax, dmin) <= threshold) ? 1.0f : 0.0f;
  }
}

}  // namespace vk

using namespace vk;

// Matrix::operator* (include/vulcan/matrix.h:296-316), evaluated on the host like the
// reference does for Tcw = Tcd * Tdw (src/color_integrator.cu:191-193; host code, no FMA).
static void matmul4(const float* A, const float* Bm, float* C) {
  for (int p = 0; p < 4; ++p)
    for (int m = 0; m < 4; ++m) {
      volatile float acc = 0;
      for (int n = 0; n < 4; ++n) {
        volatile float prod = A[n * 4 + m] * Bm[p * 4 + n];
        acc = acc + prod;
      }
      C[p * 4 + m] = acc;
    }
}

static int fill_params(IntegrateParams& p, vkVolume* v, const vkFrame* f, float min_depth, float max_depth,
                       float max_dw, float max_cw, cudaStream_t s) {
  if (!v || !f || !v->voxels || !v->sdf || !f->depth) return VK_E_INVALID;
  p.visible = v->visible[v->visible_slot];
  p.count = v->counters + VK_CNT_VISIBLE0 + v->visible_slot;
  p.entries = v->entries;
  p.voxels = v->voxels;
  p.sdf = v->sdf;
  p.push_count = 0; p.push_rank = 0;
  for (int r = 0; r < VK_MAX_SHARDS; ++r) p.push_sdf[r] = nullptr;
  if (v->shard && v->shard->count > 1) {
    if (!v->shard->connected) return VK_E_INVALID;
    p.push_count = v->shard->count;
    p.push_rank = v->shard->rank;
    for (int r = 0; r < v->shard->count; ++r) p.push_sdf[r] = v->shard->push_sdf[r];
  }
  p.depth = f->depth;
  p.color = f->color;
  p.normal = f->normal;
  p.mask = nullptr;
  p.updated = reinterpret_cast<unsigned long long*>(v->counters + VK_CNT_UPDATED);
  p.dw = f->depth_w; p.dh = f->depth_h; p.cw = f->color_w; p.ch = f->color_h;
  p.dproj = f->depth_proj;
  p.cproj = f->color_proj;
  rows34(f->depth_to_world.inv, p.Tdw);
  float Tcw[16];
  matmul4(f->depth_to_color.m, f->depth_to_world.inv, Tcw);
  rows34(Tcw, p.Tcw);
  rows34(f->depth_to_color.m, p.Tcd);
  p.voxel_length = v->voxel_length;
  p.block_length = VK_BLOCK_RES * v->voxel_length;
  p.trunc = make_div(v->trunc_length, s);
  p.min_depth = min_depth; p.max_depth = max_depth;
  p.max_dw = max_dw; p.max_cw = max_cw;
  p.light.intensity = 1.0f;
  p.light.position[0] = p.light.position[1] = p.light.position[2] = 0.f;
  return 0;
}

// persistent grid: one wave of resident CTAs
template <int MODE, bool F>
static int integrate_grid() {
  static int grid = 0;
  if (grid == 0) {
    int per_sm = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, vk_integrate_kernel<MODE, F>, 128, 0) != cudaSuccess ||
        per_sm <= 0)
      per_sm = 8;
    grid = per_sm * sm_count();
  }
  const int forced = g_tune[VK_TUNE_INTEGRATE_CTAS_PER_SM].load();
  if (forced > 0) return forced * sm_count();
  return grid;
}

template <int MODE, bool F>
static int launch_integrate_bulk(const IntegrateParams& p, cudaStream_t s) {
  const size_t smem = (size_t)kBulkStages * kTileBytes + kBulkStages * sizeof(uint64_t);
  static int grid = 0;
  if (grid == 0) {
    cudaFuncSetAttribute(vk_integrate_bulk_kernel<MODE, F>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    int per_sm = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, vk_integrate_bulk_kernel<MODE, F>, 128, smem) != cudaSuccess ||
        per_sm <= 0)
      per_sm = 4;
    grid = per_sm * sm_count();
  }
  const int forced = g_tune[VK_TUNE_INTEGRATE_CTAS_PER_SM].load();
  launch_chained(vk_integrate_bulk_kernel<MODE, F>, dim3(forced > 0 ? forced * sm_count() : grid), dim3(128), smem, s, p);
  VK_COUNT_LAUNCH();
  return (int)cudaGetLastError();
}

template <int MODE>
static int launch_integrate(const IntegrateParams& p, int math_mode, cudaStream_t s) {
  if (g_tune[VK_TUNE_INTEGRATE_BULK].load())   // A/B: tiles fetched whole by the copy engine
    return (math_mode == VK_MATH_STRICT) ? launch_integrate_bulk<MODE, false>(p, s) : launch_integrate_bulk<MODE, true>(p, s);
  if (math_mode == VK_MATH_STRICT)
    launch_chained(vk_integrate_kernel<MODE, false>, dim3(integrate_grid<MODE, false>()), dim3(128), 0, s, p);
  else
    launch_chained(vk_integrate_kernel<MODE, true>, dim3(integrate_grid<MODE, true>()), dim3(128), 0, s, p);
  VK_COUNT_LAUNCH();
  return (int)cudaGetLastError();
}

extern "C" int vk_integrate_depth(vkVolume* v, const vkFrame* f, float min_depth, float max_depth,
                                  float max_distance_weight, void* stream) {
  IntegrateParams p;
  int rc = fill_params(p, v, f, min_depth, max_depth, max_distance_weight, 16.f, (cudaStream_t)stream);
  if (rc) return rc;
  return launch_integrate<MODE_DEPTH>(p, v->math_mode, (cudaStream_t)stream);
}

extern "C" int vk_integrate_color(vkVolume* v, const vkFrame* f, float min_depth, float max_depth,
                                  float max_distance_weight, float max_color_weight, void* stream) {
  IntegrateParams p;
  int rc = fill_params(p, v, f, min_depth, max_depth, max_distance_weight, max_color_weight, (cudaStream_t)stream);
  if (rc) return rc;
  if (!f->color || f->color_w <= 0 || f->color_h <= 0) return VK_E_INVALID;
  v->has_color = 1;
  return launch_integrate<MODE_COLOR>(p, v->math_mode, (cudaStream_t)stream);
}

extern "C" int vk_compute_frame_mask(const vkFrame* f, float depth_threshold, float* mask, void* stream) {
  if (!f || !f->depth || !f->color || !mask) return VK_E_INVALID;
  dim3 grid((f->depth_w + 15) / 16, (f->depth_h + 15) / 16), block(16, 16);
  launch_chained(vk_frame_mask_kernel, grid, block, 0, (cudaStream_t)stream, f->depth_w, f->depth_h, f->depth, f->color,
                 depth_threshold, mask);
  VK_COUNT_LAUNCH();
  return (int)cudaGetLastError();
}

extern "C" int vk_integrate_light(vkVolume* v, const vkFrame* f, float min_depth, float max_depth,
                                  float max_distance_weight, float max_color_weight, const vkLight* light,
                                  float depth_threshold, float* mask, void* stream) {
  IntegrateParams p;
  int rc = fill_params(p, v, f, min_depth, max_depth, max_distance_weight, max_color_weight, (cudaStream_t)stream);
  if (rc) return rc;
  if (!f->color || !f->normal || !mask || !light) return VK_E_INVALID;
  rc = vk_compute_frame_mask(f, depth_threshold, mask, stream);
  if (rc) return rc;
  p.mask = mask;
  p.light = *light;
  v->has_color = 1;
  return launch_integrate<MODE_LIGHT>(p, v->math_mode, (cudaStream_t)stream);
}
