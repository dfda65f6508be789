// gpu_voxels/helpers/cuda_datatypes.h -- host-side value types of the reference's public API with the same names,
// member names and memory layout (GV/helpers/cuda_vectors.h:42-57,105-115,192-202; GV/helpers/cuda_matrices.h:429-434).
// Layout is contract: Vector3f is 12 packed bytes (point clouds are arrays of it), Matrix4f is 16 row-major floats.
#ifndef GPU_VOXELS_B200_HELPERS_CUDA_DATATYPES_H
#define GPU_VOXELS_B200_HELPERS_CUDA_DATATYPES_H

#include <cmath>
#include <cstring>
#include <ostream>
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
  bool operator==(const Vec3T& o) const { return x == o.x && y == o.y && z == o.z; }
  bool operator!=(const Vec3T& o) const { return !(*this == o); }
};
template <class T> inline Vec3T<T> operator*(T s, const Vec3T<T>& v) { return v * s; }
template <class T> inline std::ostream& operator<<(std::ostream& os, const Vec3T<T>& v)
{
  return os << "(x, y, z) = (" << v.x << ", " << v.y << ", " << v.z << ")";
}

typedef Vec3T<float> Vector3f;
typedef Vec3T<uint32_t> Vector3ui;
typedef Vec3T<int32_t> Vector3i;

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
};

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
  const float* data() const { return &a11; }
  float* data() { return &a11; }
  bool operator==(const Matrix4f& o) const { return std::memcmp(this, &o, sizeof(*this)) == 0; }
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
