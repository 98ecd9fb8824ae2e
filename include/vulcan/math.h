// vulcan/math.h - scalar helpers with the reference's names and tie behaviour
// (include/vulcan/math.h:9-44): min(a,b) = (b<a)?b:a, max(a,b) = (b>a)?b:a, clamp = min(hi, max(lo, v)).
#pragma once

#include <cmath>
#include <vulcan/device.h>

namespace vulcan
{

template <typename T> VULCAN_HOST_DEVICE inline T min(T a, T b) { return (b < a) ? b : a; }
template <typename T> VULCAN_HOST_DEVICE inline T max(T a, T b) { return (b > a) ? b : a; }
template <typename T> VULCAN_HOST_DEVICE inline T clamp(T v, T vmin, T vmax) { return min(vmax, max(vmin, v)); }
template <typename T> VULCAN_HOST_DEVICE inline T sqrt(T value) { return ::std::sqrt(value); }
template <typename T> VULCAN_HOST_DEVICE inline bool isnan(T value) { return value != value; }

} // namespace vulcan
