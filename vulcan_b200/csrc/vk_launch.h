// vk_launch.h — host-side helpers shared by the C-ABI translation units.
#pragma once

#include <atomic>
#include <cstring>
#include <cuda_runtime.h>

namespace vk {
// number of kernels this library has launched since load (vk_launch_count)
extern std::atomic<unsigned long long> g_launches;
}  // namespace vk

#define VK_COUNT_LAUNCH() (::vk::g_launches.fetch_add(1, std::memory_order_relaxed))
