// vk_runtime.cu — library-wide host state: launch counter, default math mode, SM count and
// the per-divisor proof that vk::divc is correctly rounded.
#include "vk_launch.h"

#include <cstdio>
#include <cstdlib>
#include <map>
#include <mutex>

namespace vk {

std::atomic<unsigned long long> g_launches{0};
std::atomic<int> g_default_math_mode{VK_MATH_REFERENCE};
std::atomic<int> g_tune[VK_TUNE_COUNT] = {{64}, {12}, {3}, {0}, {0}, {4}, {0}, {8}, {4}, {1}, {1}, {8}, {1}, {0}, {0}, {1}};
std::atomic<int> g_capture_seen{0};

// Compares the one-step FMA division against the IEEE division.
//
// `binade` != 0 (what make_div runs): the 2^24 dividends +-[1, 2).  That is a complete proof for the
// range divc uses the fast path on.  Inside that range (dividend exponent field 87..166, divisor
// exponent field 64..190, checked by divc / make_div) the quotient, the exact residual
// e = x - q0 * d and the rounded sum q0 + e * r are all normal numbers far from overflow, so for
// x' = x * 2^k every operation of both sequences scales exactly by 2^k: the fast path agrees with
// __fdiv_rn for x' if and only if it agrees for x.  One binade covers every mantissa.
// `binade` == 0: all 2^32 bit patterns (6 ms; kept as a test hook that validates the argument above).
__global__ void __launch_bounds__(256) vk_divcheck_kernel(float d, float r, unsigned int* mismatches, int binade) {
  Div dv;
  dv.d = d;
  dv.r = r;
  dv.fast = 1;
  unsigned int bad = 0;
  const unsigned int stride = gridDim.x * blockDim.x;
  unsigned int i = blockIdx.x * blockDim.x + threadIdx.x;
  // grid-strided; stride divides the pattern count so every thread runs the same trip count
  const unsigned long long total = binade ? (1ull << 24) : (1ull << 32);
  const unsigned int trips = (unsigned int)(total / stride);
  for (unsigned int k = 0; k < trips; ++k, i += stride) {
    const unsigned int bits = binade ? (((i >> 23) << 31) | (127u << 23) | (i & 0x7fffffu)) : i;
    const float x = __uint_as_float(bits);
    const float a = divc(x, dv);
    const float b = __fdiv_rn(x, d);
    // NaNs compare by class, everything else by bits
    const bool same = (__float_as_uint(a) == __float_as_uint(b)) || (a != a && b != b);
    bad += same ? 0u : 1u;
    if (!same) atomicMin(mismatches + 1, bits);  // smallest failing pattern, for diagnostics
  }
  bad = __reduce_add_sync(0xffffffffu, bad);
  if ((threadIdx.x & 31) == 0 && bad) atomicAdd(mismatches, bad);
}

// VK_TUNE="knob=value,knob=value" presets the tuning knobs at load (experiments only)
static const bool g_tune_env = [] {
  const char* e = getenv("VK_TUNE");
  while (e && *e) {
    int k = -1, v = 0, used = 0;
    if (sscanf(e, "%d=%d%n", &k, &v, &used) == 2 && k >= 0 && k < VK_TUNE_COUNT) g_tune[k].store(v);
    e = strchr(e, ',');
    if (e) ++e;
  }
  return true;
}();

static std::mutex g_div_mutex;
static std::map<uint32_t, int> g_div_cache;
static unsigned int g_div_last[2] = {0, 0};  // mismatch count / smallest failing dividend of the last check

// runs the check for d on `stream` and waits for it; 1 = the fast path is proven
static int run_divcheck(float d, float r, cudaStream_t stream, int binade) {
  unsigned int* dev = nullptr;
  unsigned int host[2] = {1u, 0xffffffffu};
  int ok = 0;
  if (cudaMalloc((void**)&dev, 2 * sizeof(unsigned int)) == cudaSuccess) {
    if (cudaMemcpyAsync(dev, host, sizeof(host), cudaMemcpyHostToDevice, stream) == cudaSuccess &&
        cudaMemsetAsync(dev, 0, sizeof(unsigned int), stream) == cudaSuccess) {
      if (binade) vk_divcheck_kernel<<<256, 256, 0, stream>>>(d, r, dev, 1);   // 2^16 threads x 256 trips
      else vk_divcheck_kernel<<<4096, 256, 0, stream>>>(d, r, dev, 0);         // 2^20 threads x 4096 trips
      VK_COUNT_LAUNCH();
      if (cudaMemcpyAsync(host, dev, sizeof(host), cudaMemcpyDeviceToHost, stream) == cudaSuccess &&
          cudaStreamSynchronize(stream) == cudaSuccess)
        ok = (host[0] == 0) ? 1 : 0;
    }
    cudaFree(dev);
  }
  g_div_last[0] = host[0];
  g_div_last[1] = host[1];
  cudaGetLastError();
  return ok;
}

Div make_div(float d, cudaStream_t stream) {
  Div dv;
  dv.d = d;
  dv.r = 1.0f / d;  // host division is IEEE correctly rounded
  dv.fast = 0;
  uint32_t bits;
  memcpy(&bits, &d, 4);
  const uint32_t mant = bits & 0x7fffffu;
  const uint32_t expo = (bits >> 23) & 0xffu;
  if (!(d == d) || expo < 64 || expo > 190 || mant == 0x7fffffu) return dv;  // keep it simple: moderate, normal
  std::lock_guard<std::mutex> lock(g_div_mutex);
  auto it = g_div_cache.find(bits);
  if (it != g_div_cache.end()) {
    dv.fast = it->second;
    return dv;
  }
  // First use of this divisor value in the process: ~20 us of GPU time and one stream
  // synchronisation (not capturable into a graph - the host layers therefore prime the divisors of a
  // volume when its lengths are set, see vk_div_prepare).
  const int ok = run_divcheck(d, dv.r, stream, 1);
  g_div_cache[bits] = ok;
  dv.fast = ok;
  return dv;
}

int sm_count() {
  static std::atomic<int> cached{0};
  int n = cached.load();
  if (n > 0) return n;
  int dev = 0;
  cudaGetDevice(&dev);
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, dev) != cudaSuccess) return 148;
  cached.store(prop.multiProcessorCount);
  return prop.multiProcessorCount;
}

}  // namespace vk

extern "C" int vk_shard_owner(int bx, int by, int bz, int shift, int count) {
  if (count <= 0 || shift < 0 || shift > 15) return VK_E_INVALID;
  return vk::shard_owner(bx, by, bz, shift, count);
}
extern "C" int vk_version(void) { return 110; }
extern "C" unsigned long long vk_launch_count(void) { return vk::g_launches.load(); }
extern "C" int vk_set_default_math_mode(int mode) {
  if (mode != VK_MATH_REFERENCE && mode != VK_MATH_STRICT) return VK_E_INVALID;
  vk::g_default_math_mode.store(mode);
  return 0;
}
// Experiment hook (not part of the ABI header): sets a tuning knob, returns its previous value.
extern "C" int vk_tune(int knob, int value) {
  if (knob < 0 || knob >= vk::VK_TUNE_COUNT) return VK_E_INVALID;
  return vk::g_tune[knob].exchange(value);
}
extern "C" int vk_get_default_math_mode(void) { return vk::g_default_math_mode.load(); }

// Test hook: runs the exhaustive division check for `d` and reports whether the fast path is on.
extern "C" int vk_div_prepare(float d, void* stream) { return vk::make_div(d, (cudaStream_t)stream).fast; }
// Test hook (not in the ABI header): the same verdict from all 2^32 dividends, uncached.
extern "C" int vk_div_check_exhaustive(float d, void* stream) {
  std::lock_guard<std::mutex> lock(vk::g_div_mutex);
  return vk::run_divcheck(d, 1.0f / d, (cudaStream_t)stream, 0);
}
extern "C" void vk_div_last_check(unsigned int out[2]) { out[0] = vk::g_div_last[0]; out[1] = vk::g_div_last[1]; }
