// vulcan/transform.h - rigid transform kept as a 4x4 matrix plus its inverse, 128 bytes
// (reference: include/vulcan/transform.h:9-171; byte image = vkTransform).  Inverse() swaps the
// pair, operator* multiplies both pairs, operator*(Vector4f) evaluates rows 0..2 left to right and
// passes w through - the evaluation order the kernels reproduce.
#pragma once

#include <vulcan/device.h>
#include <vulcan/matrix.h>

namespace vulcan
{

class Transform
{
  public:

    VULCAN_HOST_DEVICE Transform() : matrix_(Matrix4f::Identity()), inv_matrix_(Matrix4f::Identity()) {}

    VULCAN_HOST_DEVICE const Matrix4f& GetMatrix() const { return matrix_; }
    VULCAN_HOST_DEVICE const Matrix4f& GetInverseMatrix() const { return inv_matrix_; }

    VULCAN_HOST_DEVICE Vector3f GetTranslation() const { return Vector3f(matrix_(0, 3), matrix_(1, 3), matrix_(2, 3)); }

    VULCAN_HOST_DEVICE Vector4f operator*(const Vector4f& p) const
    {
      const Matrix4f& M = matrix_;
      Vector4f r;
      for (int i = 0; i < 3; ++i) r[i] = M(i, 0) * p[0] + M(i, 1) * p[1] + M(i, 2) * p[2] + M(i, 3) * p[3];
      r[3] = p[3];
      return r;
    }

    VULCAN_HOST_DEVICE Transform operator*(const Transform& t) const
    {
      return Transform(matrix_ * t.matrix_, t.inv_matrix_ * inv_matrix_);
    }

    VULCAN_HOST_DEVICE Transform Inverse() const { return Transform(inv_matrix_, matrix_); }

    // rotation from a 3x3 matrix (taken as is: no re-orthogonalisation, like the reference)
    VULCAN_HOST_DEVICE static Transform Rotate(const Matrix3f& R)
    {
      Matrix4f m = Matrix4f::Identity();
      for (int c = 0; c < 3; ++c) for (int r = 0; r < 3; ++r) m(r, c) = R(r, c);
      return Transform(m, m.Transpose());
    }

    // rotation from a unit quaternion (w, x, y, z)
    VULCAN_HOST_DEVICE static Transform Rotate(float w, float x, float y, float z)
    {
      Matrix4f m = Matrix4f::Identity();
      m(0, 0) = 1 - 2 * (y * y + z * z); m(0, 1) = 2 * (x * y - w * z);     m(0, 2) = 2 * (x * z + w * y);
      m(1, 0) = 2 * (x * y + w * z);     m(1, 1) = 1 - 2 * (x * x + z * z); m(1, 2) = 2 * (y * z - w * x);
      m(2, 0) = 2 * (x * z - w * y);     m(2, 1) = 2 * (y * z + w * x);     m(2, 2) = 1 - 2 * (x * x + y * y);
      return Transform(m, m.Transpose());
    }

    VULCAN_HOST_DEVICE static Transform Rotate(const Vector4f& q) { return Rotate(q[0], q[1], q[2], q[3]); }

    VULCAN_HOST_DEVICE static Transform Translate(float x, float y, float z)
    {
      Matrix4f m = Matrix4f::Identity(), i = Matrix4f::Identity();
      m(0, 3) = x;  m(1, 3) = y;  m(2, 3) = z;
      i(0, 3) = -x; i(1, 3) = -y; i(2, 3) = -z;
      return Transform(m, i);
    }

    VULCAN_HOST_DEVICE static Transform Translate(const Vector3f& t) { return Translate(t[0], t[1], t[2]); }

    // Not in the reference: a transform from an explicit matrix / inverse pair (poses that were
    // computed in double precision and rounded once; used by the synthetic-orbit harnesses).
    VULCAN_HOST_DEVICE static Transform FromMatrices(const Matrix4f& matrix, const Matrix4f& inv_matrix)
    {
      return Transform(matrix, inv_matrix);
    }

  protected:

    VULCAN_HOST_DEVICE Transform(const Matrix4f& matrix, const Matrix4f& inv_matrix) :
      matrix_(matrix), inv_matrix_(inv_matrix) {}

    Matrix4f matrix_;
    Matrix4f inv_matrix_;
};

static_assert(sizeof(Transform) == 128, "Transform is passed to the C ABI as a byte image of vkTransform");

} // namespace vulcan
