// vulcan/tracer.cuh - step-wise tracer functions on raw device pointers
// (reference: include/vulcan/tracer.cuh:14-32, used directly by tests/tracer_test.cu:135,234,326).
#pragma once

#include <vulcan/hash.h>
#include <vulcan/projection.h>
#include <vulcan/tracer.h>
#include <vulcan/transform.h>
#include <vulcan/voxel.h>

namespace vulcan
{

void ComputePatches(const int* indices, const HashEntry* entries, const Transform& Tcw, const Projection& projection,
    float block_length, float min_depth, float max_depth, int block_count, int image_width, int image_height,
    int bounds_width, int bounds_height, Patch* patches, int* patch_count);

void ComputeBounds(const Patch* patches, Vector2f* bounds, int bounds_width, int patch_count);

void ComputePoints(const HashEntry* hash_entries, const Voxel* voxels, const Vector2f* bounds, int block_count,
    float block_length, float voxel_length, float trunc_length, const Transform& Twc, const Projection& projection,
    float* depths, Vector3f* colors, int image_width, int image_height, int bounds_width, int bounds_height);

void ComputeNormals(const float* depths, const Projection& projection, Vector3f* normals, int image_width,
    int image_height);

void ResetBoundsBuffer(Vector2f* bounds, int count);

void FilterDepths(int image_width, int image_height, const float* src, float* dst);

} // namespace vulcan
