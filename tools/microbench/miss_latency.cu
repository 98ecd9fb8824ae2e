// Micro-benchmark: latency of dependent cold loads for a lone warp, with and without a preceding
// 256 MiB memset (the "L2 flush" of the bench) - diagnostics for the ray-cast tail.
#include <cstdio>
#include <cuda_runtime.h>
#include <vector>
#include <algorithm>
#include <random>

__global__ void chase(const unsigned* buf, int n, long long* out, int use_ldg) {
  unsigned idx = 0;
  const long long t0 = clock64();
  for (int i = 0; i < n; ++i) idx = use_ldg ? __ldg(buf + idx) : __ldcg(buf + idx);
  const long long t1 = clock64();
  out[0] = t1 - t0;
  out[1] = idx;
}

int main() {
  const int words = 65536;          // 256 KiB like vkVolume::space_bits
  const int stride = 8;             // one hop per 32-byte sector
  const int hops = words / stride;  // 8192 sectors
  std::vector<unsigned> h(words, 0);
  std::vector<int> order(hops);
  for (int i = 0; i < hops; ++i) order[i] = i;
  std::mt19937 rng(1);
  std::shuffle(order.begin() + 1, order.end(), rng);
  for (int i = 0; i < hops; ++i) h[order[i] * stride] = order[(i + 1) % hops] * stride;
  unsigned* d;
  long long* out;
  char* flush;
  cudaMalloc(&d, words * 4);
  cudaMalloc(&out, 16);
  cudaMalloc(&flush, 256u << 20);
  cudaMemcpy(d, h.data(), words * 4, cudaMemcpyHostToDevice);
  long long r[2];
  const int n = 2000;
  for (int mode = 0; mode < 6; ++mode) {
    const bool do_flush = mode >= 2;
    const bool read_flush = mode >= 4;
    const int ldg = mode & 1;
    if (do_flush) cudaMemsetAsync(flush, mode, 256u << 20);
    if (read_flush) cudaMemcpyAsync(flush, flush + (128u << 20), 128u << 20, cudaMemcpyDeviceToDevice);
    chase<<<1, 32>>>(d, n, out, ldg);
    cudaDeviceSynchronize();
    cudaMemcpy(r, out, 16, cudaMemcpyDeviceToHost);
    printf("%s %s: %.0f cycles per dependent load\n",
           read_flush ? "after memset + 128 MiB copy" : do_flush ? "after 256 MiB memset       " : "no flush (L2 warm)         ",
           ldg ? "ld.global.nc" : "ld.global.cg", (double)r[0] / n);
  }
  return 0;
}
