/*
 * t4_math.h -- small fixed-size vector / rigid-transform types shared by every kernel body.
 *
 * All device code of the engine is written as `T4_HD` functions templated on the scalar type so
 * that (a) the bulk FP32 path and the FP64 refinement path share one source and (b) the very same
 * kernel bodies can be compiled by g++ into the TEST-ONLY host simulator (tests/hostsim, built with
 * -DT4C_HOSTSIM) to debug numerics on a machine without a GPU.  The product library (libt4c.so) is
 * always built by nvcc for sm_100a and contains no host execution path for these bodies.
 */
#ifndef T4_MATH_H_
#define T4_MATH_H_

#include <math.h>
#include <stdint.h>
#include <float.h>

#if defined(__CUDACC__) && !defined(T4C_HOSTSIM)
#define T4_HD __host__ __device__ __forceinline__
#define T4_HDN __host__ __device__ __noinline__
#define T4_KD __device__ __forceinline__ /* kernel bodies: may use device-only primitives */
#define T4_KDN static __device__ __noinline__
#define T4_KM __device__ __forceinline__ /* member functions of device-only classes */
#define T4_KMN __device__ __noinline__
#define T4_DEVICE_BUILD 1
#else
#define T4_HD inline
#define T4_HDN
#define T4_KD inline
#define T4_KDN inline
#define T4_KM inline
#define T4_KMN
#endif

namespace t4
{
struct alignas(16) f4
{
  float x, y, z, w;
};

template <class T>
struct Eps;
template <>
struct Eps<float>
{
  static T4_HD float eps() { return FLT_EPSILON; }
  static T4_HD float large() { return 1e30f; }
  // GJK relative duality-gap tolerance (Bullet REL_ERROR2, single / double: tesseract_gjk_pair_detector.cpp:36-42)
  static T4_HD float rel2() { return 1e-6f * 1e-6f * 25.0f; }  // gap/|v|^2 <= 2.5e-11 is below FP32 noise: acts as "converged"
  static T4_HD float touch2() { return 1e-12f; }               // |v|^2 below this: cores touch -> penetration path
};
template <>
struct Eps<double>
{
  static T4_HD double eps() { return DBL_EPSILON; }
  static T4_HD double large() { return 1e30; }
  static T4_HD double rel2() { return 1e-12; }
  static T4_HD double touch2() { return 1e-20; }
};

T4_HD float t4sqrt(float x) { return sqrtf(x); }
T4_HD double t4sqrt(double x) { return sqrt(x); }
T4_HD float t4abs(float x) { return fabsf(x); }
T4_HD double t4abs(double x) { return fabs(x); }
template <class T>
T4_HD T t4min(T a, T b)
{
  return a < b ? a : b;
}
template <class T>
T4_HD T t4max(T a, T b)
{
  return a > b ? a : b;
}

template <class T>
struct V3
{
  T x, y, z;
  T4_HD V3() : x(0), y(0), z(0) {}
  T4_HD V3(T a, T b, T c) : x(a), y(b), z(c) {}
  template <class U>
  T4_HD explicit V3(const V3<U>& o) : x((T)o.x), y((T)o.y), z((T)o.z)
  {
  }
  T4_HD V3 operator+(const V3& o) const { return V3(x + o.x, y + o.y, z + o.z); }
  T4_HD V3 operator-(const V3& o) const { return V3(x - o.x, y - o.y, z - o.z); }
  T4_HD V3 operator-() const { return V3(-x, -y, -z); }
  T4_HD V3 operator*(T s) const { return V3(x * s, y * s, z * s); }
  T4_HD V3& operator+=(const V3& o)
  {
    x += o.x;
    y += o.y;
    z += o.z;
    return *this;
  }
  T4_HD V3& operator-=(const V3& o)
  {
    x -= o.x;
    y -= o.y;
    z -= o.z;
    return *this;
  }
  T4_HD bool operator==(const V3& o) const { return x == o.x && y == o.y && z == o.z; }
  T4_HD T dot(const V3& o) const { return x * o.x + y * o.y + z * o.z; }
  T4_HD V3 cross(const V3& o) const { return V3(y * o.z - z * o.y, z * o.x - x * o.z, x * o.y - y * o.x); }
  T4_HD T len2() const { return x * x + y * y + z * z; }
  T4_HD T len() const { return t4sqrt(len2()); }
  T4_HD T get(int i) const { return i == 0 ? x : (i == 1 ? y : z); }
  T4_HD void set(int i, T v)
  {
    if (i == 0)
      x = v;
    else if (i == 1)
      y = v;
    else
      z = v;
  }
};

// rigid transform, rotation row-major
template <class T>
struct Xf
{
  T r[9];
  V3<T> p;
  T4_HD Xf()
  {
    r[0] = r[4] = r[8] = 1;
    r[1] = r[2] = r[3] = r[5] = r[6] = r[7] = 0;
  }
  T4_HD V3<T> rot(const V3<T>& v) const
  {
    return V3<T>(r[0] * v.x + r[1] * v.y + r[2] * v.z, r[3] * v.x + r[4] * v.y + r[5] * v.z,
                 r[6] * v.x + r[7] * v.y + r[8] * v.z);
  }
  // R^T v  (Bullet's "v * basis")
  T4_HD V3<T> rotT(const V3<T>& v) const
  {
    return V3<T>(r[0] * v.x + r[3] * v.y + r[6] * v.z, r[1] * v.x + r[4] * v.y + r[7] * v.z,
                 r[2] * v.x + r[5] * v.y + r[8] * v.z);
  }
  T4_HD V3<T> operator()(const V3<T>& v) const { return rot(v) + p; }
  T4_HD V3<T> invApply(const V3<T>& v) const { return rotT(v - p); }
  T4_HD V3<T> col(int j) const { return V3<T>(r[j], r[3 + j], r[6 + j]); }
  T4_HD Xf operator*(const Xf& o) const
  {
    Xf out;
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j)
        out.r[3 * i + j] = r[3 * i] * o.r[j] + r[3 * i + 1] * o.r[3 + j] + r[3 * i + 2] * o.r[6 + j];
    out.p = rot(o.p) + p;
    return out;
  }
  // this^-1 * o
  T4_HD Xf invTimes(const Xf& o) const
  {
    Xf out;
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j)
        out.r[3 * i + j] = r[i] * o.r[j] + r[3 + i] * o.r[3 + j] + r[6 + i] * o.r[6 + j];
    out.p = rotT(o.p - p);
    return out;
  }
  // from 3x4 row-major [R|p]
  template <class U>
  T4_HD static Xf from12(const U* a)
  {
    Xf t;
    for (int i = 0; i < 3; ++i)
    {
      for (int j = 0; j < 3; ++j)
        t.r[3 * i + j] = (T)a[4 * i + j];
    }
    t.p = V3<T>((T)a[3], (T)a[7], (T)a[11]);
    return t;
  }
  template <class U>
  T4_HD void to12(U* a) const
  {
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j)
        a[4 * i + j] = (U)r[3 * i + j];
    a[3] = (U)p.x;
    a[7] = (U)p.y;
    a[11] = (U)p.z;
  }
  template <class U>
  T4_HD Xf<U> cast() const
  {
    Xf<U> o;
    for (int i = 0; i < 9; ++i)
      o.r[i] = (U)r[i];
    o.p = V3<U>(p);
    return o;
  }
};

template <class T>
T4_HD V3<T> normalizedOrDefault(const V3<T>& d)
{
  // btConvexInternalShape::localGetSupportingVertex: zero direction -> (-1,-1,-1)
  V3<T> n = d;
  if (n.len2() < Eps<T>::eps() * Eps<T>::eps())
    n = V3<T>(-1, -1, -1);
  return n * (T(1) / n.len());
}

}  // namespace t4
#endif
