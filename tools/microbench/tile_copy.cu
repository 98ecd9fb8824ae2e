// Micro-benchmark: how fast can 10 KiB voxel tiles scattered over a 3.3 GB pool be read, modified
// and written back by a persistent grid?  This is the data movement of vk_integrate_kernel with
// the arithmetic and the image tests removed - the practical ceiling for that kernel.
//   mode 0: thread t owns bytes [80 t, 80 t + 80) of the tile (the kernel's layout), 5 x 16 B
//   mode 1: warp-contiguous: lane l moves 16 B at 512 i + 16 l (perfectly coalesced)
//   mode 2: mode 0 with 58 % of the threads active (as measured on the bench workload)
//   mode 3: mode 0, two tiles per iteration (more bytes in flight per thread)
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o _bin/tile_copy tile_copy.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include <vector>
#include <algorithm>
#include <random>

constexpr int kTileBytes = 10240;

__device__ __forceinline__ bool active58(int tile, int t) {
  unsigned h = (unsigned)tile * 2654435761u ^ ((unsigned)(t >> 3) * 40503u);  // groups of 8 threads
  h ^= h >> 15; h *= 2246822519u; h ^= h >> 13;
  return (h % 100u) < 58u;
}

template <int MODE>
__global__ void __launch_bounds__(128) tile_kernel(char* pool, const int* list, int n, int write) {
  const int t = threadIdx.x;
  const int step = (MODE == 3) ? 2 * gridDim.x : gridDim.x;
  for (int bi = blockIdx.x; bi < n; bi += step) {
    if (MODE == 3) {
      const int b2 = bi + gridDim.x;
      uint4* a = reinterpret_cast<uint4*>(pool + (size_t)list[bi] * kTileBytes + t * 80);
      uint4* b = reinterpret_cast<uint4*>(pool + (size_t)list[b2 < n ? b2 : bi] * kTileBytes + t * 80);
      uint4 qa[5], qb[5];
#pragma unroll
      for (int c = 0; c < 5; ++c) qa[c] = a[c];
#pragma unroll
      for (int c = 0; c < 5; ++c) qb[c] = b[c];
      if (write) {
#pragma unroll
        for (int c = 0; c < 5; ++c) { qa[c].x += 1; a[c] = qa[c]; }
        if (b2 < n) {
#pragma unroll
          for (int c = 0; c < 5; ++c) { qb[c].x += 1; b[c] = qb[c]; }
        }
      } else if (qa[0].x == 0xdeadbeefu && qb[1].y == 0x1234567u) a[0] = qb[0];
      continue;
    }
    const int tile = list[bi];
    if (MODE == 2 && !active58(tile, t)) continue;
    uint4* base = reinterpret_cast<uint4*>(pool + (size_t)tile * kTileBytes);
    uint4 q[5];
    if (MODE == 1) {
      const int warp = t >> 5, lane = t & 31;
#pragma unroll
      for (int c = 0; c < 5; ++c) q[c] = base[(warp * 5 + c) * 32 + lane];
      if (write) {
#pragma unroll
        for (int c = 0; c < 5; ++c) { q[c].x += 1; base[(warp * 5 + c) * 32 + lane] = q[c]; }
      } else if (q[0].x == 0xdeadbeefu && q[4].y == 0x1234567u) base[0] = q[1];
    } else {
      uint4* p = base + t * 5;
#pragma unroll
      for (int c = 0; c < 5; ++c) q[c] = p[c];
      if (write) {
#pragma unroll
        for (int c = 0; c < 5; ++c) { q[c].x += 1; p[c] = q[c]; }
      } else if (q[0].x == 0xdeadbeefu && q[4].y == 0x1234567u) p[0] = q[1];
    }
  }
}

int main() {
  const int pool_tiles = 327680, n = 17916;
  char *pool, *flush;
  int* list;
  cudaMalloc(&pool, (size_t)pool_tiles * kTileBytes);
  cudaMalloc(&flush, 256u << 20);
  cudaMalloc(&list, n * 4);
  cudaMemset(pool, 0, (size_t)pool_tiles * kTileBytes);
  std::vector<int> ids(pool_tiles);
  for (int i = 0; i < pool_tiles; ++i) ids[i] = i;
  std::mt19937 rng(7);
  std::shuffle(ids.begin(), ids.end(), rng);
  cudaMemcpy(list, ids.data(), n * 4, cudaMemcpyHostToDevice);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  for (int mode = 0; mode < 4; ++mode)
    for (int write = 0; write < 2; ++write)
      for (int per_sm : {4, 8, 12, 16}) {
        std::vector<float> us;
        for (int rep = 0; rep < 12; ++rep) {
          cudaMemsetAsync(flush, rep, 256u << 20);
          cudaEventRecord(e0);
          const int grid = 148 * per_sm;
          if (mode == 0) tile_kernel<0><<<grid, 128>>>(pool, list, n, write);
          if (mode == 1) tile_kernel<1><<<grid, 128>>>(pool, list, n, write);
          if (mode == 2) tile_kernel<2><<<grid, 128>>>(pool, list, n, write);
          if (mode == 3) tile_kernel<3><<<grid, 128>>>(pool, list, n, write);
          cudaEventRecord(e1);
          cudaEventSynchronize(e1);
          float ms;
          cudaEventElapsedTime(&ms, e0, e1);
          if (rep >= 2) us.push_back(ms * 1e3f);
        }
        std::sort(us.begin(), us.end());
        const double frac = (mode == 2) ? 0.58 : 1.0;
        const double bytes = (double)n * kTileBytes * frac * (write ? 2 : 1);
        printf("mode %d write %d ctas/sm %2d: median %.1f us  %.0f GB/s\n", mode, write, per_sm, us[us.size() / 2],
               bytes / us[us.size() / 2] * 1e-3);
      }
  cudaError_t err = cudaDeviceSynchronize();
  printf("status %s\n", cudaGetErrorString(err));
  return 0;
}
