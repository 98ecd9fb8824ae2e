// vulcan/matrix.h - small fixed-size column-major matrices and vectors.
// Public surface of the reference's Matrix<T,M,N> (include/vulcan/matrix.h:12-468): same type
// names, storage order (column-major, operator()(row, col)), methods and evaluation order of the
// arithmetic that the hot path's parity depends on (Dot accumulates from 0, Normalize scales by
// 1/sqrt(dot), operator/ multiplies by the reciprocal, operator* accumulates over the inner index).
#pragma once

#include <cmath>
#include <ostream>
#include <vulcan/device.h>

namespace vulcan
{

template <typename T, int M, int N>
class Matrix
{
  static_assert(M > 0 && N > 0, "invalid matrix size");
  static const bool is_vector = (M == 1 || N == 1);

  public:

    VULCAN_HOST_DEVICE Matrix() { for (int i = 0; i < M * N; ++i) data_[i] = T(0); }

    VULCAN_HOST_DEVICE Matrix(T x, T y)
    {
      static_assert(M * N == 2, "2D vector required");
      data_[0] = x; data_[1] = y;
    }

    VULCAN_HOST_DEVICE Matrix(T x, T y, T z)
    {
      static_assert(M * N == 3, "3D vector required");
      data_[0] = x; data_[1] = y; data_[2] = z;
    }

    VULCAN_HOST_DEVICE Matrix(T x, T y, T z, T w)
    {
      static_assert(M * N == 4 && is_vector, "4D vector required");
      data_[0] = x; data_[1] = y; data_[2] = z; data_[3] = w;
    }

    // element-wise conversion
    template <typename U>
    VULCAN_HOST_DEVICE explicit Matrix(const Matrix<U, M, N>& other)
    {
      for (int c = 0; c < N; ++c) for (int r = 0; r < M; ++r) (*this)(r, c) = T(other(r, c));
    }

    // vector from a vector of the same length but other orientation
    template <int P, int Q>
    VULCAN_HOST_DEVICE explicit Matrix(const Matrix<T, P, Q>& v)
    {
      static_assert(is_vector && (P == 1 || Q == 1) && P * Q == M * N, "vectors of equal length required");
      for (int i = 0; i < M * N; ++i) data_[i] = v[i];
    }

    // homogeneous extension: (v, s)
    template <int P, int Q>
    VULCAN_HOST_DEVICE Matrix(const Matrix<T, P, Q>& v, T s)
    {
      static_assert(is_vector && (P == 1 || Q == 1) && M * N == P * Q + 1, "invalid vector length");
      for (int i = 0; i < P * Q; ++i) data_[i] = v[i];
      data_[P * Q] = s;
    }

    VULCAN_HOST_DEVICE int GetRows() const { return M; }
    VULCAN_HOST_DEVICE int GetColumns() const { return N; }
    VULCAN_HOST_DEVICE int GetTotal() const { return M * N; }

    VULCAN_HOST_DEVICE T SquaredNorm() const
    {
      T sum = T(0);
      for (int i = 0; i < M * N; ++i) sum += data_[i] * data_[i];
      return sum;
    }

    VULCAN_HOST_DEVICE T Norm() const { return T(sqrtf(float(SquaredNorm()))); }

    VULCAN_HOST_DEVICE void Normalize()
    {
      const T scale = T(1) / Norm();
      for (int i = 0; i < M * N; ++i) data_[i] *= scale;
    }

    VULCAN_HOST_DEVICE Matrix Normalized() const { Matrix r(*this); r.Normalize(); return r; }

    VULCAN_HOST_DEVICE float Dot(const Matrix& rhs) const
    {
      static_assert(is_vector, "vector required");
      float sum = 0;
      for (int i = 0; i < M * N; ++i) sum += data_[i] * rhs.data_[i];
      return sum;
    }

    VULCAN_HOST_DEVICE Matrix Cross(const Matrix& b) const
    {
      static_assert(M * N == 3, "3D vector required");
      const Matrix& a = *this;
      return Matrix(a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0]);
    }

    VULCAN_HOST_DEVICE Matrix<T, N, M> Transpose() const
    {
      Matrix<T, N, M> r;
      for (int c = 0; c < N; ++c) for (int row = 0; row < M; ++row) r(c, row) = (*this)(row, c);
      return r;
    }

    VULCAN_HOST_DEVICE Matrix& operator+=(const Matrix& m) { for (int i = 0; i < M * N; ++i) data_[i] += m.data_[i]; return *this; }
    VULCAN_HOST_DEVICE Matrix& operator-=(const Matrix& m) { for (int i = 0; i < M * N; ++i) data_[i] -= m.data_[i]; return *this; }
    VULCAN_HOST_DEVICE const Matrix operator+(const Matrix& m) const { Matrix r(*this); r += m; return r; }
    VULCAN_HOST_DEVICE const Matrix operator-(const Matrix& m) const { Matrix r(*this); r -= m; return r; }

    template <typename S> VULCAN_HOST_DEVICE Matrix& operator+=(S s) { for (int i = 0; i < M * N; ++i) data_[i] += s; return *this; }
    template <typename S> VULCAN_HOST_DEVICE const Matrix operator+(S s) const { Matrix r(*this); r += s; return r; }
    template <typename S> VULCAN_HOST_DEVICE Matrix& operator*=(S s) { for (int i = 0; i < M * N; ++i) data_[i] *= s; return *this; }
    template <typename S> VULCAN_HOST_DEVICE const Matrix operator*(S s) const { Matrix r(*this); r *= s; return r; }
    // division multiplies by the reciprocal (the hot path relies on this rounding)
    template <typename S> VULCAN_HOST_DEVICE Matrix& operator/=(S s) { return (*this) *= (1.0f / s); }
    template <typename S> VULCAN_HOST_DEVICE const Matrix operator/(S s) const { Matrix r(*this); r /= s; return r; }

    template <int P>
    VULCAN_HOST_DEVICE const Matrix<T, M, P> operator*(const Matrix<T, N, P>& rhs) const
    {
      Matrix<T, M, P> r;
      for (int p = 0; p < P; ++p)
        for (int m = 0; m < M; ++m)
        {
          T sum = T(0);
          for (int n = 0; n < N; ++n) sum += (*this)(m, n) * rhs(n, p);
          r(m, p) = sum;
        }
      return r;
    }

    VULCAN_HOST_DEVICE const T& operator()(int row, int col) const { return data_[col * M + row]; }
    VULCAN_HOST_DEVICE T& operator()(int row, int col) { return data_[col * M + row]; }

    VULCAN_HOST_DEVICE const T& operator[](int index) const { static_assert(is_vector, "vector required"); return data_[index]; }
    VULCAN_HOST_DEVICE T& operator[](int index) { static_assert(is_vector, "vector required"); return data_[index]; }

    VULCAN_HOST_DEVICE bool operator==(const Matrix& m) const
    {
      for (int i = 0; i < M * N; ++i) if (data_[i] != m.data_[i]) return false;
      return true;
    }
    VULCAN_HOST_DEVICE bool operator!=(const Matrix& m) const { return !(*this == m); }

    VULCAN_HOST_DEVICE const T* GetData() const { return data_; }
    VULCAN_HOST_DEVICE T* GetData() { return data_; }

    VULCAN_HOST_DEVICE static Matrix Zeros() { return Matrix(); }
    VULCAN_HOST_DEVICE static Matrix Ones() { Matrix r; for (int i = 0; i < M * N; ++i) r.data_[i] = T(1); return r; }
    VULCAN_HOST_DEVICE static Matrix Identity()
    {
      static_assert(M == N, "square matrix required");
      Matrix r;
      for (int i = 0; i < M; ++i) r(i, i) = T(1);
      return r;
    }

  protected:

    T data_[M * N];
};

template <typename S, typename T, int M, int N>
VULCAN_HOST_DEVICE inline Matrix<T, M, N> operator+(S s, const Matrix<T, M, N>& m) { return m + s; }

template <typename T, int M, int N>
VULCAN_HOST_DEVICE inline Matrix<T, M, N> operator*(float s, const Matrix<T, M, N>& m) { return m * s; }

template <typename T, int M, int N>
std::ostream& operator<<(std::ostream& out, const Matrix<T, M, N>& m)
{
  for (int r = 0; r < M; ++r)
  {
    for (int c = 0; c < N; ++c) out << (c ? " " : "") << m(r, c);
    if (r + 1 < M) out << "\n";
  }
  return out;
}

template <typename T, int M> using Vector = Matrix<T, M, 1>;

typedef Vector<char, 2> Vector2c;
typedef Vector<char, 3> Vector3c;
typedef Vector<char, 4> Vector4c;
typedef Vector<int, 2> Vector2i;
typedef Vector<int, 3> Vector3i;
typedef Vector<int, 4> Vector4i;
typedef Vector<short, 2> Vector2s;
typedef Vector<short, 3> Vector3s;
typedef Vector<short, 4> Vector4s;
typedef Vector<float, 2> Vector2f;
typedef Vector<float, 3> Vector3f;
typedef Vector<float, 4> Vector4f;
typedef Vector<float, 6> Vector6f;
typedef Vector<double, 2> Vector2d;
typedef Vector<double, 3> Vector3d;
typedef Vector<double, 4> Vector4d;
typedef Vector<double, 6> Vector6d;
typedef Matrix<float, 2, 2> Matrix2f;
typedef Matrix<float, 3, 3> Matrix3f;
typedef Matrix<float, 4, 4> Matrix4f;
typedef Matrix<double, 2, 2> Matrix2d;
typedef Matrix<double, 3, 3> Matrix3d;
typedef Matrix<double, 4, 4> Matrix4d;

} // namespace vulcan
