// gpu_voxels/helpers/cuda_datatypes.h -- host-side value types of the reference's public API with the same names,
// member names and memory layout (GV/helpers/cuda_vectors.h:42-57,105-115,192-202; GV/helpers/cuda_matrices.h:429-434).
// Layout is contract: Vector3f is 12 packed bytes (point clouds are arrays of it), Matrix4f is 16 row-major floats.
#ifndef GPU_VOXELS_B200_HELPERS_CUDA_DATATYPES_H
#define GPU_VOXELS_B200_HELPERS_CUDA_DATATYPES_H

#include <cmath>
#include <cstdio>
#include <cstring>
#include <iomanip>
#include <ostream>
#include <utility>
#include <stdint.h>

namespace gpu_voxels {

template <class T>
struct Vec3T
{
  T x, y, z;
  Vec3T() : x(0), y(0), z(0) {}
  explicit Vec3T(T v) : x(v), y(v), z(v) {}
  Vec3T(T x_, T y_, T z_) : x(x_), y(y_), z(z_) {}
  template <class U> explicit Vec3T(const Vec3T<U>& o) : x(T(o.x)), y(T(o.y)), z(T(o.z)) {}
  Vec3T operator+(const Vec3T& o) const { return Vec3T(x + o.x, y + o.y, z + o.z); }
  Vec3T operator-(const Vec3T& o) const { return Vec3T(x - o.x, y - o.y, z - o.z); }
  Vec3T operator*(T s) const { return Vec3T(x * s, y * s, z * s); }
  Vec3T operator/(T s) const { return Vec3T(x / s, y / s, z / s); }
  Vec3T& operator+=(const Vec3T& o) { x += o.x; y += o.y; z += o.z; return *this; }
  Vec3T& operator-=(const Vec3T& o) { x -= o.x; y -= o.y; z -= o.z; return *this; }
  // GV/helpers/cuda_vectors.h:268-272
  float length() const { return std::sqrt(float((x * x) + (y * y) + (z * z))); }
  // the rest of the reference's Vector3f value API (GV/helpers/cuda_vectors.h:247-331), same expressions and types
  static bool compVec(const Vec3T& i, const Vec3T& j)
  {
    return i.x < j.x || (i.x == j.x && i.y < j.y) || (i.x == j.x && i.y == j.y && i.z < j.z);
  }
  static bool eqlVec(const Vec3T& i, const Vec3T& j) { return i == j; }
  bool apprx_equal(const Vec3T& b, double epsilon) const
  {
    return std::fabs(x - b.x) < epsilon && std::fabs(y - b.y) < epsilon && std::fabs(z - b.z) < epsilon;
  }
  void normalize() { x = x / length(); y = y / length(); z = z / length(); }  // yes: the length changes in between
  Vec3T normalized() const { return Vec3T(x / length(), y / length(), z / length()); }
  Vec3T abs() const { return Vec3T(T(std::fabs(x)), T(std::fabs(y)), T(std::fabs(z))); }
  T dot(const Vec3T& o) const { return x * o.x + y * o.y + z * o.z; }
  Vec3T cross(const Vec3T& o) const { return Vec3T(y * o.z - z * o.y, z * o.x - x * o.z, x * o.y - y * o.x); }
  float angleBetween(const Vec3T& other) const
  {
    float rad = this->normalized().dot(other.normalized());
    if (rad < -1.0) rad = -1.0;
    else if (rad > 1.0) rad = 1.0;
    return std::acos(rad);
  }
  bool operator==(const Vec3T& o) const { return x == o.x && y == o.y && z == o.z; }
  bool operator!=(const Vec3T& o) const { return !(*this == o); }
};
template <class T> inline Vec3T<T> operator*(T s, const Vec3T<T>& v) { return v * s; }
template <class T> inline std::ostream& operator<<(std::ostream& os, const Vec3T<T>& v)
{
  return os << "(x, y, z) = (" << v.x << ", " << v.y << ", " << v.z << ")" << std::endl;  // ends the line, as there
}

typedef Vec3T<float> Vector3f;
typedef Vec3T<uint32_t> Vector3ui;
typedef Vec3T<int32_t> Vector3i;

// Gauss-Jordan inversion in double as the reference's host code runs it (GV/helpers/cuda_matrices.h:83-149): rows are
// swapped only for an exactly zero pivot, forward elimination, determinant test (== 0 -> not invertible), backward
// elimination, then every row is divided by its pivot. Same sequence of operations, so the same doubles come out.
namespace detail {
template <std::size_t N> inline bool gaussJordanInverse(const double (&mat)[N][N], double (&inv)[N][N])
{
  double w[N][2 * N];
  for (std::size_t r = 0; r < N; ++r)
    for (std::size_t c = 0; c < 2 * N; ++c) w[r][c] = c < N ? mat[r][c] : (c - N == r ? 1.0 : 0.0);
  for (std::size_t k = 0; k + 1 < N; ++k)
  {
    if (w[k][k] == 0.0)
    {
      std::size_t other = k + 1;
      while (other < N && w[other][k] == 0.0) ++other;
      if (other == N) return false;
      for (std::size_t c = 0; c < 2 * N; ++c) std::swap(w[k][c], w[other][c]);
    }
    for (std::size_t r = k + 1; r < N; ++r)
    {
      const double p = w[r][k] / w[k][k];
      for (std::size_t c = k; c < 2 * N; ++c) w[r][c] -= w[k][c] * p;
    }
  }
  double det = 1.0;
  for (std::size_t k = 0; k < N; ++k) det *= w[k][k];
  if (det == 0.0) return false;
  for (std::size_t k = N - 1; k > 0; --k)
    for (std::size_t r = k; r-- > 0;)
    {
      const double p = w[r][k] / w[k][k];
      for (std::size_t c = k; c < 2 * N; ++c) w[r][c] -= w[k][c] * p;
    }
  for (std::size_t r = 0; r < N; ++r)
  {
    const double f = w[r][r];
    for (std::size_t c = 0; c < N; ++c) inv[r][c] = w[r][N + c] / f;
  }
  return true;
}
}  // namespace detail

struct Matrix3f
{
  float a11, a12, a13, a21, a22, a23, a31, a32, a33;
  Matrix3f() : a11(0), a12(0), a13(0), a21(0), a22(0), a23(0), a31(0), a32(0), a33(0) {}
  Matrix3f(float b11, float b12, float b13, float b21, float b22, float b23, float b31, float b32, float b33)
    : a11(b11), a12(b12), a13(b13), a21(b21), a22(b22), a23(b23), a31(b31), a32(b32), a33(b33) {}
  static Matrix3f createIdentity() { return Matrix3f(1, 0, 0, 0, 1, 0, 0, 0, 1); }
  // single-axis rotations and their products exactly as the reference's host code builds them
  // (GV/helpers/cuda_matrices.h:185-277): float sin / cos of the float angle, every entry of a product is
  // (x_i1*y_1j + x_i2*y_2j) + x_i3*y_3j in fp32 without contraction -- pinned by tests/golden/geometry.npz
  static Matrix3f createFromRoll(float angle)
  {
    return Matrix3f(1, 0, 0, 0, cosf(angle), -sinf(angle), 0, sinf(angle), cosf(angle));
  }
  static Matrix3f createFromPitch(float angle)
  {
    return Matrix3f(cosf(angle), 0, sinf(angle), 0, 1, 0, -sinf(angle), 0, cosf(angle));
  }
  static Matrix3f createFromYaw(float angle)
  {
    return Matrix3f(cosf(angle), -sinf(angle), 0, sinf(angle), cosf(angle), 0, 0, 0, 1);
  }
  Matrix3f operator*(const Matrix3f& rhs) const
  {
    const float* x = &a11;
    const float* y = &rhs.a11;
    Matrix3f r;
    float* o = &r.a11;
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j)
      {
        volatile float p0 = x[3 * i] * y[j], p1 = x[3 * i + 1] * y[3 + j], p2 = x[3 * i + 2] * y[6 + j];
        volatile float sum = p0 + p1;
        sum = sum + p2;
        o[3 * i + j] = sum;
      }
    return r;
  }
  // rotation about x by roll, then y by pitch, then z by yaw: R = Rz(yaw) * Ry(pitch) * Rx(roll)
  static Matrix3f createFromRPY(float roll, float pitch, float yaw)
  {
    return createFromYaw(yaw) * createFromPitch(pitch) * createFromRoll(roll);
  }
  static Matrix3f createFromYPR(float roll, float pitch, float yaw)
  {
    return createFromRoll(roll) * createFromPitch(pitch) * createFromYaw(yaw);
  }
  static Matrix3f createFromRPY(const Vector3f& rpy) { return createFromRPY(rpy.x, rpy.y, rpy.z); }
  static Matrix3f createFromYPR(const Vector3f& rpy) { return createFromYPR(rpy.x, rpy.y, rpy.z); }

  // the two roll / pitch / yaw solutions of a rotation matrix (GV/helpers/cuda_matrices.h:285-327): solution 1 for
  // solution_number == 1, the other one for every other value; near the pitch singularity yaw is set to 0
  Vector3f toRPY(unsigned int solution_number = 1) const
  {
    Vector3f out1, out2;
    if (1.0 - std::fabs(a31) < 0.00001)
    {
      const bool down = a31 < 0;
      const float delta = down ? std::atan2(a12, a13) : std::atan2(-a12, -a13);
      out1.y = out2.y = float(down ? M_PI_2 : -M_PI_2);
      out1.x = out2.x = delta;
    }
    else
    {
      out1.y = -std::asin(a31);
      out2.y = float(M_PI - out1.y);
      out1.x = std::atan2(a32 / std::cos(out1.y), a33 / std::cos(out1.y));
      out2.x = std::atan2(a32 / std::cos(out2.y), a33 / std::cos(out2.y));
      out1.z = std::atan2(a21 / std::cos(out1.y), a11 / std::cos(out1.y));
      out2.z = std::atan2(a21 / std::cos(out2.y), a11 / std::cos(out2.y));
    }
    return solution_number == 1 ? out1 : out2;
  }
  // small-angle difference of two orientations: asin of half the summed column cross products (:335-353)
  Vector3f orientationMatrixDiff(const Matrix3f& other) const
  {
    Vector3f d = Vector3f(a11, a21, a31).cross(Vector3f(other.a11, other.a21, other.a31));
    d = d + Vector3f(a12, a22, a32).cross(Vector3f(other.a12, other.a22, other.a32));
    d = d + Vector3f(a13, a23, a33).cross(Vector3f(other.a13, other.a23, other.a33));
    return Vector3f(float(std::asin(d.x / 2.0)), float(std::asin(d.y / 2.0)), float(std::asin(d.z / 2.0)));
  }
  bool invertMatrix(Matrix3f& inverse_matrix) const
  {
    double m[3][3], inv[3][3];
    const float* in = &a11;
    for (int i = 0; i < 9; ++i) m[i / 3][i % 3] = in[i];
    if (!detail::gaussJordanInverse<3>(m, inv)) return false;
    float* out = &inverse_matrix.a11;
    for (int i = 0; i < 9; ++i) out[i] = float(inv[i / 3][i % 3]);
    return true;
  }
  friend std::ostream& operator<<(std::ostream& out, const Matrix3f& m)
  {
    const float* v = &m.a11;
    out.precision(3);
    out << "\n" << std::fixed;
    for (int r = 0; r < 3; ++r)
    {
      out << (r ? " " : "[");
      for (int c = 0; c < 3; ++c) out << std::setw(10) << v[3 * r + c] << (c < 2 ? ", " : (r < 2 ? ",\n" : "]"));
    }
    return out << std::endl;
  }
};
inline Vector3f operator*(const Matrix3f& m, const Vector3f& v)
{
  return Vector3f(m.a11 * v.x + m.a12 * v.y + m.a13 * v.z, m.a21 * v.x + m.a22 * v.y + m.a23 * v.z,
                  m.a31 * v.x + m.a32 * v.y + m.a33 * v.z);
}

struct Matrix4f
{
  float a11, a12, a13, a14;
  float a21, a22, a23, a24;
  float a31, a32, a33, a34;
  float a41, a42, a43, a44;

  Matrix4f() { std::memset(this, 0, sizeof(*this)); }
  Matrix4f(float b11, float b12, float b13, float b14, float b21, float b22, float b23, float b24,
           float b31, float b32, float b33, float b34, float b41, float b42, float b43, float b44)
    : a11(b11), a12(b12), a13(b13), a14(b14), a21(b21), a22(b22), a23(b23), a24(b24),
      a31(b31), a32(b32), a33(b33), a34(b34), a41(b41), a42(b42), a43(b43), a44(b44) {}
  static Matrix4f createIdentity() { return Matrix4f(1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1); }
  static Matrix4f createFromRotationAndTranslation(const Matrix3f& r, const Vector3f& t)
  {
    return Matrix4f(r.a11, r.a12, r.a13, t.x, r.a21, r.a22, r.a23, t.y, r.a31, r.a32, r.a33, t.z, 0, 0, 0, 1);
  }
  void setIdentity() { *this = createIdentity(); }
  // GV/helpers/cuda_matrices.h:480-660
  Matrix3f getRotation() const { return Matrix3f(a11, a12, a13, a21, a22, a23, a31, a32, a33); }
  Vector3f getTranslation() const { return Vector3f(a14, a24, a34); }
  void setRotation(const Matrix3f& r)
  {
    a11 = r.a11; a12 = r.a12; a13 = r.a13; a21 = r.a21; a22 = r.a22; a23 = r.a23; a31 = r.a31; a32 = r.a32; a33 = r.a33;
  }
  void setTranslation(const Vector3f& t) { a14 = t.x; a24 = t.y; a34 = t.z; }
  Matrix4f transpose() const
  {
    return Matrix4f(a11, a21, a31, a41, a12, a22, a32, a42, a13, a23, a33, a43, a14, a24, a34, a44);
  }
  inline Matrix4f& leftMultiply(const Matrix4f& other);  // *this = *this * other (the name is the reference's)
  bool apprx_equal(const Matrix4f& b, double epsilon) const
  {
    const float* p = &a11;
    const float* q = &b.a11;
    for (int i = 0; i < 16; ++i)
      if (!(std::fabs(p[i] - q[i]) < epsilon)) return false;
    return true;
  }
  bool invertMatrix(Matrix4f& inverse_matrix) const
  {
    double m[4][4], inv[4][4];
    const float* in = &a11;
    for (int i = 0; i < 16; ++i) m[i / 4][i % 4] = in[i];
    if (!detail::gaussJordanInverse<4>(m, inv)) return false;
    float* out = &inverse_matrix.a11;
    for (int i = 0; i < 16; ++i) out[i] = float(inv[i / 4][i % 4]);
    return true;
  }
  void print() const
  {
    const float* v = &a11;
    for (int r = 0; r < 4; ++r)
      std::printf("  %0.7lf  %0.7lf  %0.7lf  %0.7lf\n%s", v[4 * r], v[4 * r + 1], v[4 * r + 2], v[4 * r + 3], r == 3 ? "\n" : "");
  }
  friend std::ostream& operator<<(std::ostream& out, const Matrix4f& m)
  {
    const float* v = &m.a11;
    out.precision(3);
    out << "\n" << std::fixed;
    for (int r = 0; r < 4; ++r)
    {
      out << (r ? " " : "[");
      for (int c = 0; c < 4; ++c) out << std::setw(10) << v[4 * r + c] << (c < 3 ? ", " : (r < 3 ? ",\n" : "]"));
    }
    return out << std::endl;
  }
  const float* data() const { return &a11; }
  float* data() { return &a11; }
  // element-wise float comparison (GV/helpers/cuda_matrices.h:560-567): -0 equals +0, NaN equals nothing
  bool operator==(const Matrix4f& o) const
  {
    for (int i = 0; i < 16; ++i)
      if (!(data()[i] == o.data()[i])) return false;
    return true;
  }
};

// Matrix4f * Matrix4f as the reference's HOST code evaluates it (GV/helpers/cuda_matrices.h:772-803): every entry is
// ((x_i1*y_1j + x_i2*y_2j) + x_i3*y_3j) + x_i4*y_4j in fp32, left to right, no FMA contraction (g++ -ffp-contract=off
// semantics: each product is rounded before it is added). Used to build kinematic chains on the host.
inline Matrix4f operator*(const Matrix4f& x, const Matrix4f& y)
{
  Matrix4f r;
  const float* a = x.data();
  const float* b = y.data();
  float* o = r.data();
  for (int i = 0; i < 4; ++i)
    for (int j = 0; j < 4; ++j)
    {
      volatile float p0 = a[4 * i] * b[j], p1 = a[4 * i + 1] * b[4 + j], p2 = a[4 * i + 2] * b[8 + j],
                     p3 = a[4 * i + 3] * b[12 + j];
      volatile float s = p0 + p1;
      s = s + p2;
      s = s + p3;
      o[4 * i + j] = s;
    }
  return r;
}

inline Matrix4f& Matrix4f::leftMultiply(const Matrix4f& other) { *this = *this * other; return *this; }
inline Matrix4f operator-(const Matrix4f& a, const Matrix4f& b)
{
  Matrix4f r;
  for (int i = 0; i < 16; ++i) r.data()[i] = a.data()[i] - b.data()[i];
  return r;
}
// affine apply on the host (GV/helpers/cuda_matrices.h:740-746); the device kernels use the contraction nvcc emits
inline Vector3f operator*(const Matrix4f& m, const Vector3f& v)
{
  return Vector3f(m.a11 * v.x + m.a12 * v.y + m.a13 * v.z + m.a14, m.a21 * v.x + m.a22 * v.y + m.a23 * v.z + m.a24,
                  m.a31 * v.x + m.a32 * v.y + m.a33 * v.z + m.a34);
}

// box given by centre, HALF side lengths and roll / pitch / yaw (GV/helpers/cuda_datatypes.h:92-97)
struct OrientedBoxParams
{
  Vector3f dim;
  Vector3f center;
  Vector3f rot;
};

// depth sensor description (GV/helpers/cuda_datatypes.h:45-69)
struct Sensor
{
  Sensor() : data_width(0), data_height(0), data_size(0) {}
  Sensor(Vector3f _position, Matrix3f _orientation, uint32_t _data_width, uint32_t _data_height)
    : position(_position), orientation(_orientation), data_width(_data_width), data_height(_data_height),
      data_size(_data_width * _data_height) {}
  Vector3f position;
  Matrix3f orientation;
  uint32_t data_width;
  uint32_t data_height;
  uint32_t data_size;
};

static_assert(sizeof(Vector3f) == 12, "Vector3f must be 12 packed bytes");
static_assert(sizeof(Matrix4f) == 64, "Matrix4f must be 16 packed floats");

}  // namespace gpu_voxels
#endif
