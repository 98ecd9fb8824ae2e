// Micro-benchmark: cost of one ray-march step for a lone warp (diagnostics for the ray-cast tail).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o /tmp/free_run tools/microbench/free_run.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#include "../../vulcan_b200/csrc/vk_math.cuh"

using namespace vk;

struct P {
  const uint32_t* bits;
  Div L;
  float Tz[4];
  float hi;
  long long* out;
  int steps;
};

template <int VARIANT, int UN>
__global__ void k(const P p) {
  const int lane = threadIdx.x & 31;
  if (lane >= 4 * 8 && VARIANT < 100) return;
  const bool act = (lane % 8) == 0;  // 4 active lanes like the tail tiles
  Vec3 pt = {0.3f + lane * 0.01f, -0.2f, 0.5f};
  Vec3 dir = {0.05f, 0.02f, 0.998f};
  int iters = 0;
  int acc = 0;
  const long long t0 = clock64();
  if (act) {
    for (int n = 0; n < p.steps / UN; ++n) {
      Vec3 q[UN];
      float qd[UN];
      int qb[UN][3];
      bool occ[UN];
      Vec3 c = pt;
#pragma unroll
      for (int kk = 0; kk < UN; ++kk) {
        c.x = __fmaf_rn(dir.x, p.L.d, c.x);
        c.y = __fmaf_rn(dir.y, p.L.d, c.y);
        c.z = __fmaf_rn(dir.z, p.L.d, c.z);
        q[kk] = c;
      }
#pragma unroll
      for (int kk = 0; kk < UN; ++kk) {
        qd[kk] = add(FP<true>::mad(q[kk].z, p.Tz[2], FP<true>::mad(q[kk].x, p.Tz[0], mul(q[kk].y, p.Tz[1]))), p.Tz[3]);
        float fx, fy, fz;
        if (VARIANT == 0) divc3(q[kk].x, q[kk].y, q[kk].z, p.L, fx, fy, fz);
        if (VARIANT == 1) { fx = divc(q[kk].x, p.L); fy = divc(q[kk].y, p.L); fz = divc(q[kk].z, p.L); }
        if (VARIANT == 2) { fx = q[kk].x * p.L.r; fy = q[kk].y * p.L.r; fz = q[kk].z * p.L.r; }
        if (VARIANT == 3) { fx = __fdiv_rn(q[kk].x, p.L.d); fy = __fdiv_rn(q[kk].y, p.L.d); fz = __fdiv_rn(q[kk].z, p.L.d); }
        qb[kk][0] = (int)floorf(fx); qb[kk][1] = (int)floorf(fy); qb[kk][2] = (int)floorf(fz);
        uint32_t sw, sb;
        space_bit_index(qb[kk][0], qb[kk][1], qb[kk][2], sw, sb);
        occ[kk] = (__ldg(p.bits + sw) >> sb) & 1u;
      }
      bool stop = false;
#pragma unroll
      for (int kk = 0; kk < UN; ++kk) {
        if (!stop) {
          pt = q[kk];
          acc += qb[kk][0] ^ qb[kk][1] ^ qb[kk][2];
          if (++iters >= 100000) stop = true;
          else if (!(qd[kk] < p.hi)) stop = true;
          else if (occ[kk]) stop = true;
        }
      }
      if (stop) break;
    }
  }
  const long long t1 = clock64();
  if (lane == 0) { p.out[0] = t1 - t0; p.out[1] = iters; p.out[2] = acc; }
}

template <int V, int UN>
void run(const P& p, const char* name) {
  long long h[3];
  for (int rep = 0; rep < 2; ++rep) {
    k<V, UN><<<1, 32>>>(p);
    cudaDeviceSynchronize();
  }
  cudaMemcpy(h, p.out, sizeof(h), cudaMemcpyDeviceToHost);
  printf("%-28s unroll %d: %lld cycles for %lld steps = %.1f cycles/step (err %s)\n", name, UN, h[0], h[1],
         (double)h[0] / (double)h[1], cudaGetErrorString(cudaGetLastError()));
}

int main() {
  P p;
  uint32_t* bits;
  cudaMalloc(&bits, 65536 * 4);
  cudaMemset(bits, 0, 65536 * 4);
  cudaMalloc(&p.out, 64);
  p.bits = bits;
  p.L.d = 0.032f; p.L.r = 1.0f / 0.032f; p.L.fast = 1;
  p.Tz[0] = 0.1f; p.Tz[1] = 0.05f; p.Tz[2] = 0.99f; p.Tz[3] = 0.2f;
  p.hi = 1e9f;
  p.steps = 2000;
  run<0, 4>(p, "divc3 (combined guard)");
  run<0, 1>(p, "divc3 (combined guard)");
  run<1, 4>(p, "divc x3 (3 guards)");
  run<1, 1>(p, "divc x3 (3 guards)");
  run<2, 4>(p, "plain multiply");
  run<2, 1>(p, "plain multiply");
  run<3, 4>(p, "__fdiv_rn");
  run<3, 1>(p, "__fdiv_rn");
  p.L.fast = 0;
  run<0, 4>(p, "divc3 with fast=0");
  return 0;
}
