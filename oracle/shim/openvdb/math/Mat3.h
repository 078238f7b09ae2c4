// TEST INFRASTRUCTURE ONLY.  Minimal stand-in for <openvdb/math/Mat3.h> (see Vec3.h in this directory).
//   Mat3(v1, v2, v3, rows = true): the vectors are the rows, or the COLUMNS when rows == false;
//   Mat3 * Vec3: result[i] = v[0]*m[i][0] + v[1]*m[i][1] + v[2]*m[i][2].
#pragma once
#include "Vec3.h"

namespace openvdb {
namespace math {

template <typename T>
class Mat3 {
public:
  T m[3][3];
  Mat3(const Vec3<T> &v1, const Vec3<T> &v2, const Vec3<T> &v3, bool rows = true) {
    const Vec3<T> *v[3] = {&v1, &v2, &v3};
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j) {
        if (rows) m[i][j] = v[i]->mm[j];
        else m[j][i] = v[i]->mm[j];
      }
  }
};
template <typename T>
inline Vec3<T> operator*(const Mat3<T> &a, const Vec3<T> &v) {
  return Vec3<T>(v.mm[0] * a.m[0][0] + v.mm[1] * a.m[0][1] + v.mm[2] * a.m[0][2],
                 v.mm[0] * a.m[1][0] + v.mm[1] * a.m[1][1] + v.mm[2] * a.m[1][2],
                 v.mm[0] * a.m[2][0] + v.mm[1] * a.m[2][1] + v.mm[2] * a.m[2][2]);
}

}  // namespace math
}  // namespace openvdb
