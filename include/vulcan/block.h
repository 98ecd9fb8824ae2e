// vulcan/block.h - coordinates of one 8x8x8 voxel block: 3 x int16 + pad, 8 bytes
// (reference: include/vulcan/block.h:9-86; byte image = vkHashEntry::block/pad).
#pragma once

#include <vulcan/device.h>
#include <vulcan/matrix.h>

namespace vulcan
{

class Block
{
  public:

    static const int resolution = 8;
    static const int voxel_count = 512;

    VULCAN_HOST_DEVICE Block() : origin_(0, 0, 0), pad_(0) {}
    VULCAN_HOST_DEVICE Block(short x, short y, short z) : origin_(x, y, z), pad_(0) {}
    VULCAN_HOST_DEVICE Block(const Vector3s& origin) : origin_(origin), pad_(0) {}

    VULCAN_HOST_DEVICE const Vector3s& GetOrigin() const { return origin_; }
    VULCAN_HOST_DEVICE bool operator==(const Block& block) const { return origin_ == block.origin_; }
    VULCAN_HOST_DEVICE bool operator!=(const Block& block) const { return !(origin_ == block.origin_); }
    VULCAN_HOST_DEVICE const short& operator[](int index) const { return origin_[index]; }
    VULCAN_HOST_DEVICE short& operator[](int index) { return origin_[index]; }

  protected:

    Vector3s origin_;

  private:

    short pad_;  // the reference leaves it uninitialised; always 0 here
};

static_assert(sizeof(Block) == 8, "Block is part of the ABI (tests/block_test.cpp:9-12)");

} // namespace vulcan
