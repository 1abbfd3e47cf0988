// Small fixed-size linear algebra shared by device kernels and the CPU oracle.
// Restates, with an explicit operation order, the pieces of Eigen 3.3 the
// reference's hot path leans on (the reference does not vendor Eigen; the
// orders below follow Eigen 3.3.7's un-vectorised fixed-size evaluators as
// best known: 3-term reductions are a0 + (a1 + a2), 4-term are (a0+a1)+(a2+a3)).
// Everything is plain IEEE arithmetic: build with -ffp-contract=off / -fmad=false.
#pragma once
#include "detmath.h"

namespace lsb {

template <typename T>
struct Vec3 {
  T x, y, z;
  LSB_HD Vec3() : x(0), y(0), z(0) {}
  LSB_HD Vec3(T a, T b, T c) : x(a), y(b), z(c) {}
  LSB_HD T operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
  LSB_HD void set(int i, T v) { if (i == 0) x = v; else if (i == 1) y = v; else z = v; }
};
using Vec3f = Vec3<float>;
using Vec3d = Vec3<double>;

template <typename T> LSB_HD Vec3<T> operator+(const Vec3<T>& a, const Vec3<T>& b) { return Vec3<T>(a.x + b.x, a.y + b.y, a.z + b.z); }
template <typename T> LSB_HD Vec3<T> operator-(const Vec3<T>& a, const Vec3<T>& b) { return Vec3<T>(a.x - b.x, a.y - b.y, a.z - b.z); }
template <typename T> LSB_HD Vec3<T> operator-(const Vec3<T>& a) { return Vec3<T>(-a.x, -a.y, -a.z); }
template <typename T> LSB_HD Vec3<T> operator*(T s, const Vec3<T>& a) { return Vec3<T>(s * a.x, s * a.y, s * a.z); }
template <typename T> LSB_HD Vec3<T> operator*(const Vec3<T>& a, T s) { return Vec3<T>(a.x * s, a.y * s, a.z * s); }
template <typename T> LSB_HD Vec3<T> operator/(const Vec3<T>& a, T s) { return Vec3<T>(a.x / s, a.y / s, a.z / s); }

// Eigen redux order for 3 terms: c0 + (c1 + c2)
template <typename T> LSB_HD T sum3(T c0, T c1, T c2) { return c0 + (c1 + c2); }
template <typename T> LSB_HD T dot(const Vec3<T>& a, const Vec3<T>& b) { return sum3(a.x * b.x, a.y * b.y, a.z * b.z); }
template <typename T> LSB_HD T sqnorm(const Vec3<T>& a) { return dot(a, a); }
template <typename T> LSB_HD T norm(const Vec3<T>& a) { return dm::t_sqrt(sqnorm(a)); }
template <typename T> LSB_HD Vec3<T> cross(const Vec3<T>& a, const Vec3<T>& b)
{
  return Vec3<T>(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
// Eigen::MatrixBase::normalized(): divide by sqrt(squaredNorm) if > 0
template <typename T> LSB_HD Vec3<T> normalized(const Vec3<T>& a)
{
  const T z = sqnorm(a);
  if (z > T(0)) return a / dm::t_sqrt(z);
  return a;
}

// Row-major 3x3
template <typename T>
struct Mat3 {
  T m[9];
  LSB_HD T& operator()(int r, int c) { return m[r * 3 + c]; }
  LSB_HD T operator()(int r, int c) const { return m[r * 3 + c]; }
  LSB_HD Vec3<T> row(int r) const { return Vec3<T>(m[r * 3], m[r * 3 + 1], m[r * 3 + 2]); }
  LSB_HD Vec3<T> col(int c) const { return Vec3<T>(m[c], m[3 + c], m[6 + c]); }
  LSB_HD void setcol(int c, const Vec3<T>& v) { m[c] = v.x; m[3 + c] = v.y; m[6 + c] = v.z; }
  LSB_HD static Mat3 identity()
  {
    Mat3 r;
    for (int i = 0; i < 9; ++i) r.m[i] = T(0);
    r.m[0] = r.m[4] = r.m[8] = T(1);
    return r;
  }
};
using Mat3f = Mat3<float>;
using Mat3d = Mat3<double>;

template <typename T> LSB_HD Vec3<T> operator*(const Mat3<T>& A, const Vec3<T>& v)
{
  return Vec3<T>(sum3(A.m[0] * v.x, A.m[1] * v.y, A.m[2] * v.z),
                 sum3(A.m[3] * v.x, A.m[4] * v.y, A.m[5] * v.z),
                 sum3(A.m[6] * v.x, A.m[7] * v.y, A.m[8] * v.z));
}
template <typename T> LSB_HD Mat3<T> operator*(const Mat3<T>& A, const Mat3<T>& B)
{
  Mat3<T> C;
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j)
      C.m[i * 3 + j] = sum3(A.m[i * 3] * B.m[j], A.m[i * 3 + 1] * B.m[3 + j], A.m[i * 3 + 2] * B.m[6 + j]);
  return C;
}
template <typename T> LSB_HD Mat3<T> transpose(const Mat3<T>& A)
{
  Mat3<T> C;
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) C.m[i * 3 + j] = A.m[j * 3 + i];
  return C;
}

// ---------------------------------------------------------------------------
// Quaternion (x,y,z,w) and rigid isometry, double only (pose bookkeeping)
// ---------------------------------------------------------------------------
struct Quatd {
  double x, y, z, w;
  LSB_HD Quatd() : x(0), y(0), z(0), w(1) {}
  LSB_HD Quatd(double w_, double x_, double y_, double z_) : x(x_), y(y_), z(z_), w(w_) {}
};

// Eigen::QuaternionBase::toRotationMatrix()
LSB_HD Mat3d quat_to_mat(const Quatd& q)
{
  const double tx = 2.0 * q.x, ty = 2.0 * q.y, tz = 2.0 * q.z;
  const double twx = tx * q.w, twy = ty * q.w, twz = tz * q.w;
  const double txx = tx * q.x, txy = ty * q.x, txz = tz * q.x;
  const double tyy = ty * q.y, tyz = tz * q.y, tzz = tz * q.z;
  Mat3d R;
  R(0, 0) = 1.0 - (tyy + tzz); R(0, 1) = txy - twz;         R(0, 2) = txz + twy;
  R(1, 0) = txy + twz;         R(1, 1) = 1.0 - (txx + tzz); R(1, 2) = tyz - twx;
  R(2, 0) = txz - twy;         R(2, 1) = tyz + twx;         R(2, 2) = 1.0 - (txx + tyy);
  return R;
}

// Eigen quaternionbase_assign_impl<Matrix3> (Shepperd)
LSB_HD Quatd mat_to_quat(const Mat3d& M)
{
  Quatd q;
  double t = sum3(M(0, 0), M(1, 1), M(2, 2));
  if (t > 0.0) {
    t = ::sqrt(t + 1.0);
    q.w = 0.5 * t;
    t = 0.5 / t;
    q.x = (M(2, 1) - M(1, 2)) * t;
    q.y = (M(0, 2) - M(2, 0)) * t;
    q.z = (M(1, 0) - M(0, 1)) * t;
  } else {
    int i = 0;
    if (M(1, 1) > M(0, 0)) i = 1;
    if (M(2, 2) > M(i, i)) i = 2;
    const int j = (i + 1) % 3, k = (j + 1) % 3;
    t = ::sqrt(M(i, i) - M(j, j) - M(k, k) + 1.0);
    double c[3];
    c[i] = 0.5 * t;
    t = 0.5 / t;
    q.w = (M(k, j) - M(j, k)) * t;
    c[j] = (M(j, i) + M(i, j)) * t;
    c[k] = (M(k, i) + M(i, k)) * t;
    q.x = c[0]; q.y = c[1]; q.z = c[2];
  }
  return q;
}

LSB_HD Quatd quat_mul(const Quatd& a, const Quatd& b)
{
  return Quatd(a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z,
               a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y,
               a.w * b.y + a.y * b.w + a.z * b.x - a.x * b.z,
               a.w * b.z + a.z * b.w + a.x * b.y - a.y * b.x);
}

// Eigen::QuaternionBase::slerp (not re-normalised; t may be outside [0,1])
LSB_HD Quatd slerp(const Quatd& a, double t, const Quatd& b)
{
  const double one = 1.0 - 2.220446049250313e-16;
  const double d = (a.x * b.x + a.z * b.z) + (a.y * b.y + a.w * b.w);
  const double absD = (d < 0.0) ? -d : d;
  double s0, s1;
  if (absD >= one) {
    s0 = 1.0 - t;
    s1 = t;
  } else {
    const double theta = dm::acos(absD);
    const double sinTheta = dm::sin(theta);
    s0 = dm::sin((1.0 - t) * theta) / sinTheta;
    s1 = dm::sin(t * theta) / sinTheta;
  }
  if (d < 0.0) s1 = -s1;
  Quatd r;
  r.x = s0 * a.x + s1 * b.x;
  r.y = s0 * a.y + s1 * b.y;
  r.z = s0 * a.z + s1 * b.z;
  r.w = s0 * a.w + s1 * b.w;
  return r;
}

struct Iso3d {
  Mat3d R;
  Vec3d t;
  LSB_HD static Iso3d identity()
  {
    Iso3d r;
    r.R = Mat3d::identity();
    r.t = Vec3d(0, 0, 0);
    return r;
  }
};

// Transform*Transform (affine/isometry mode): linear product + rotated translation
LSB_HD Iso3d operator*(const Iso3d& a, const Iso3d& b)
{
  Iso3d r;
  r.R = a.R * b.R;
  r.t = a.R * b.t + a.t;
  return r;
}
// Isometry inverse: R^T, -(R^T t)
LSB_HD Iso3d inverse(const Iso3d& a)
{
  Iso3d r;
  r.R = transpose(a.R);
  r.t = -(r.R * a.t);
  return r;
}
// Isometry3d * Vector3d  (KeypointsMatcher.cxx:118): linear*v + translation
LSB_HD Vec3d apply(const Iso3d& T, const Vec3d& v) { return T.R * v + T.t; }

// Utils::TransformPoint (Utilities.h:274-278): homogeneous 4-vector product in
// double, (r0*x + r1*y) + (r2*z + t*1), then cast to float.
LSB_HD void transform_point_f(const Iso3d& T, float& x, float& y, float& z)
{
  const double dx = (double)x, dy = (double)y, dz = (double)z;
  const double ox = (T.R.m[0] * dx + T.R.m[1] * dy) + (T.R.m[2] * dz + T.t.x * 1.0);
  const double oy = (T.R.m[3] * dx + T.R.m[4] * dy) + (T.R.m[5] * dz + T.t.y * 1.0);
  const double oz = (T.R.m[6] * dx + T.R.m[7] * dy) + (T.R.m[8] * dz + T.t.z * 1.0);
  x = (float)ox; y = (float)oy; z = (float)oz;
}

// Translation * Quaternion -> Isometry
LSB_HD Iso3d make_iso(const Vec3d& t, const Quatd& q)
{
  Iso3d r;
  r.R = quat_to_mat(q);
  r.t = t;
  return r;
}

// Transform::isApprox on the 4x4 matrices, prec = 1e-12
LSB_HD bool iso_is_approx(const Iso3d& a, const Iso3d& b)
{
  double d2 = 0.0, na = 1.0, nb = 1.0;  // the (3,3) entries are both 1
  for (int i = 0; i < 9; ++i) {
    const double d = a.R.m[i] - b.R.m[i];
    d2 += d * d; na += a.R.m[i] * a.R.m[i]; nb += b.R.m[i] * b.R.m[i];
  }
  for (int i = 0; i < 3; ++i) {
    const double d = a.t[i] - b.t[i];
    d2 += d * d; na += a.t[i] * a.t[i]; nb += b.t[i] * b.t[i];
  }
  const double mn = (na < nb) ? na : nb;
  return d2 <= 1e-12 * 1e-12 * mn;
}

// Utils::RPYtoRotationMatrix (Utilities.cxx:33-38): AngleAxis(z)*AngleAxis(y)*AngleAxis(x)
LSB_HD Mat3d rpy_to_mat(double roll, double pitch, double yaw)
{
  const Quatd qz(dm::cos(0.5 * yaw), 0.0, 0.0, dm::sin(0.5 * yaw));
  const Quatd qy(dm::cos(0.5 * pitch), 0.0, dm::sin(0.5 * pitch), 0.0);
  const Quatd qx(dm::cos(0.5 * roll), dm::sin(0.5 * roll), 0.0, 0.0);
  return quat_to_mat(quat_mul(quat_mul(qz, qy), qx));
}
// Utils::RotationMatrixToRPY (Utilities.cxx:41-59)
LSB_HD Vec3d mat_to_rpy(const Mat3d& R)
{
  return Vec3d(dm::atan2(R(2, 1), R(2, 2)), -dm::asin(R(2, 0)), dm::atan2(R(1, 0), R(0, 0)));
}
LSB_HD void iso_to_xyzrpy(const Iso3d& T, double* x6)
{
  const Vec3d rpy = mat_to_rpy(T.R);
  x6[0] = T.t.x; x6[1] = T.t.y; x6[2] = T.t.z;
  x6[3] = rpy.x; x6[4] = rpy.y; x6[5] = rpy.z;
}
LSB_HD Iso3d xyzrpy_to_iso(const double* x6)
{
  Iso3d r;
  r.R = rpy_to_mat(x6[3], x6[4], x6[5]);
  r.t = Vec3d(x6[0], x6[1], x6[2]);
  return r;
}

// CeresCostFunctions.h:67-79 RotationMatrixFromRPY, explicit 9 entries
LSB_HD Mat3d rot_from_rpy_explicit(double rx, double ry, double rz)
{
  const double cx = dm::cos(rx), sx = dm::sin(rx);
  const double cy = dm::cos(ry), sy = dm::sin(ry);
  const double cz = dm::cos(rz), sz = dm::sin(rz);
  Mat3d R;
  R(0, 0) = cy * cz; R(0, 1) = sx * sy * cz - cx * sz; R(0, 2) = cx * sy * cz + sx * sz;
  R(1, 0) = cy * sz; R(1, 1) = sx * sy * sz + cx * cz; R(1, 2) = cx * sy * sz - sx * cz;
  R(2, 0) = -sy;     R(2, 1) = sx * cy;                R(2, 2) = cx * cy;
  return R;
}

// AngleAxisd(Matrix3d).angle()  (Slam.cxx:1183)
LSB_HD double rotation_angle(const Mat3d& R)
{
  const Quatd q = mat_to_quat(R);
  const double n = ::sqrt(sum3(q.x * q.x, q.y * q.y, q.z * q.z));
  if (n != 0.0) return 2.0 * dm::atan2(n, (q.w < 0.0) ? -q.w : q.w);
  return 0.0;
}

// LinearInterpolation (MotionModel.cxx:26-34): returns H1 when degenerate
LSB_HD Iso3d linear_interpolation(const Iso3d& H0, const Iso3d& H1, double t, double t0, double t1)
{
  if (t0 == t1 || iso_is_approx(H0, H1)) return H1;
  const double time = (t - t0) / (t1 - t0);
  const Quatd rot = slerp(mat_to_quat(H0.R), time, mat_to_quat(H1.R));
  const Vec3d trans = H0.t + time * (H1.t - H0.t);
  return make_iso(trans, rot);
}

// LinearTransformInterpolator<double> (MotionModel.h:36-136). POD so that it
// can live in device memory.
struct Interpolator {
  int invalid;
  double t0, t1;
  Quatd q0, q1;
  Vec3d p0, p1;

  LSB_HD Iso3d H0() const { return make_iso(p0, q0); }
  LSB_HD Iso3d H1() const { return make_iso(p1, q1); }
  LSB_HD void revalidate() { invalid = (t0 == t1 || iso_is_approx(H0(), H1())) ? 1 : 0; }
  LSB_HD void set_transforms(const Iso3d& a, const Iso3d& b)
  {
    q0 = mat_to_quat(a.R); p0 = a.t;
    q1 = mat_to_quat(b.R); p1 = b.t;
    revalidate();
  }
  LSB_HD void set_times(double a, double b) { t0 = a; t1 = b; revalidate(); }
  LSB_HD void reset()
  {
    t0 = 0.0; t1 = 1.0;
    set_transforms(Iso3d::identity(), Iso3d::identity());
  }
  LSB_HD Iso3d at(double t) const
  {
    if (invalid) return H0();
    const double time = (t - t0) / (t1 - t0);
    return make_iso(p0 + time * (p1 - p0), slerp(q0, time, q1));
  }
};

}  // namespace lsb
