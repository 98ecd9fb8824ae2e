// vulcan/hash.h - one entry of the voxel hash, 16 bytes (reference: include/vulcan/hash.h:10-56;
// byte image = vkHashEntry of include/vulcan_b200.h).
#pragma once

#include <vulcan/block.h>
#include <vulcan/device.h>

namespace vulcan
{

class HashEntry
{
  public:

    static const int invalid = -1;

    VULCAN_HOST_DEVICE HashEntry() : data(invalid), next(invalid) {}

    VULCAN_HOST_DEVICE bool IsAllocated() const { return data != invalid; }
    VULCAN_HOST_DEVICE void InvalidateData() { data = invalid; }
    VULCAN_HOST_DEVICE bool HasNext() const { return next != invalid; }
    VULCAN_HOST_DEVICE void InvalidateNext() { next = invalid; }

  public:

    Block block;
    int data;  // voxel-block id (voxels at data * 512), -1 = not allocated
    int next;  // entry index of the next chain element (>= main block count), -1 = end
};

static_assert(sizeof(HashEntry) == 16, "HashEntry is part of the ABI (tests/hash_test.cpp:9-12)");

} // namespace vulcan
