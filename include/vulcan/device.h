// vulcan/device.h - CUDA runtime glue of the drop-in C++ API (reference: include/vulcan/device.h).
// The host classes only need the CUDA runtime for memory; every kernel lives behind the C ABI
// (include/vulcan_b200.h), so these headers compile with a plain C++ compiler.
#pragma once

#include <cuda_runtime_api.h>
#include <string>
#include <vulcan/exception.h>

#if defined(__CUDACC__)
#define VULCAN_HOST_DEVICE __host__ __device__
#define VULCAN_HOST __host__
#define VULCAN_DEVICE __device__
#define VULCAN_GLOBAL __global__
#else
#define VULCAN_HOST_DEVICE
#define VULCAN_HOST
#define VULCAN_DEVICE
#define VULCAN_GLOBAL
#endif

namespace vulcan
{

// "<cudaGetErrorString> [cuda error N]" like the reference (include/vulcan/device.h:97-101)
inline std::string GetCudaErrorString(cudaError_t code)
{
  return std::string(cudaGetErrorString(code)) + " [cuda error " + std::to_string((int)code) + "]";
}

// status of a C-ABI call (include/vulcan_b200.h): 0 ok, >0 cudaError_t, <0 VK_E_*
inline std::string GetStatusString(int status)
{
  if (status > 0) return GetCudaErrorString((cudaError_t)status);
  switch (status)
  {
    case -1: return "invalid argument [vulcan error -1]";
    case -2: return "no CUDA device [vulcan error -2]";
    case -3: return "voxel or excess pool exhausted [vulcan error -3]";
  }
  return "vulcan error " + std::to_string(status);
}

// grid size helper with the reference's name (include/vulcan/device.h:111-123)
inline int GetKernelBlocks(int total, int threads)
{
  return (total + threads - 1) / threads;
}

} // namespace vulcan

#define CUDA_ASSERT(cmd) do { const cudaError_t vulcan_code_ = (cmd); \
  VULCAN_ASSERT_MSG(vulcan_code_ == cudaSuccess, ::vulcan::GetCudaErrorString(vulcan_code_)); } while (0)
#define VK_ASSERT(cmd) do { const int vulcan_status_ = (cmd); \
  VULCAN_ASSERT_MSG(vulcan_status_ == 0, ::vulcan::GetStatusString(vulcan_status_)); } while (0)

#ifdef NDEBUG
#define CUDA_DEBUG(cmd) (void)(cmd)
#else
#define CUDA_DEBUG CUDA_ASSERT
#endif
