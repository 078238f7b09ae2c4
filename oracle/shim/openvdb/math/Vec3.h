// TEST INFRASTRUCTURE ONLY.  Minimal stand-in for <openvdb/math/Vec3.h> (OpenVDB, the un-vendored dependency of
// /root/reference/GPURenderer/ShadowFX.cpp; the reference's Visual Studio project links OpenVDB 5.x/6.x, no version
// pinned in the tree) so that the reference's ShadowFX.cpp compiles IN PLACE into oracle/_ref (oracle/Makefile, target
// _ref).  Only the members ShadowFX.cpp uses, restating OpenVDB's published arithmetic:
//   length() = sqrt(x*x + y*y + z*z);  normalize(eps = 1e-7): if |length| <= eps leave unchanged, else *= 1/length;
//   dot: x*x' + y*y' + z*z';  cross: (y*z' - z*y', z*x' - x*z', x*y' - y*x');  scalar ops component-wise in T.
#pragma once
#include <cmath>

namespace openvdb {
namespace math {

template <typename T>
class Vec3 {
public:
  T mm[3];
  Vec3() : mm{T(0), T(0), T(0)} {}
  template <typename A, typename B, typename C>
  Vec3(A x, B y, C z) : mm{static_cast<T>(x), static_cast<T>(y), static_cast<T>(z)} {}
  T x() const { return mm[0]; }
  T y() const { return mm[1]; }
  T z() const { return mm[2]; }
  T operator[](int i) const { return mm[i]; }
  T length() const { return static_cast<T>(std::sqrt(mm[0] * mm[0] + mm[1] * mm[1] + mm[2] * mm[2])); }
  bool normalize(T eps = static_cast<T>(1.0e-7)) {
    const T d = length();
    if (!(std::fabs(d - T(0)) > eps)) return false;   // isApproxEqual(d, 0, eps)
    *this *= (T(1) / d);
    return true;
  }
  template <typename S>
  Vec3 &operator*=(S s) {
    mm[0] = static_cast<T>(mm[0] * s);
    mm[1] = static_cast<T>(mm[1] * s);
    mm[2] = static_cast<T>(mm[2] * s);
    return *this;
  }
  T dot(const Vec3 &v) const { return mm[0] * v.mm[0] + mm[1] * v.mm[1] + mm[2] * v.mm[2]; }
  Vec3 cross(const Vec3 &v) const {
    return Vec3(mm[1] * v.mm[2] - mm[2] * v.mm[1], mm[2] * v.mm[0] - mm[0] * v.mm[2], mm[0] * v.mm[1] - mm[1] * v.mm[0]);
  }
};
template <typename T, typename S>
inline Vec3<T> operator*(const Vec3<T> &v, S s) {
  Vec3<T> r(v);
  r *= s;
  return r;
}
template <typename T>
inline Vec3<T> operator+(const Vec3<T> &a, const Vec3<T> &b) {
  return Vec3<T>(a.mm[0] + b.mm[0], a.mm[1] + b.mm[1], a.mm[2] + b.mm[2]);
}
template <typename T>
inline Vec3<T> operator-(const Vec3<T> &a, const Vec3<T> &b) {
  return Vec3<T>(a.mm[0] - b.mm[0], a.mm[1] - b.mm[1], a.mm[2] - b.mm[2]);
}

}  // namespace math
}  // namespace openvdb
