// Restatement of the two PCL routines behind Utils::ComputeMeanAndPCA
// (reference slam_lib/include/LidarSlam/Utilities.h:247-262):
//   pcl::computeMeanAndCovarianceMatrix (PCL <= 1.10 form, single pass, unshifted)
//   pcl::eigen33 (closed-form symmetric 3x3, ascending eigenvalues)
// PCL is NOT under /root/reference; this follows the published PCL 1.10
// common/impl/centroid.hpp and common/impl/eigen.hpp algorithms.
// Scalar = float in the keypoint extractor, double in the matcher.
#pragma once
#include "linalg.h"

namespace lsb {

template <typename T> struct NumEps;
template <> struct NumEps<float> { LSB_HD static float eps() { return 1.1920928955078125e-07f; } LSB_HD static float tiny() { return 1.17549435e-38f; } LSB_HD static float dummy() { return 1e-5f; } };
template <> struct NumEps<double> { LSB_HD static double eps() { return 2.220446049250313e-16; } LSB_HD static double tiny() { return 2.2250738585072014e-308; } LSB_HD static double dummy() { return 1e-12; } };

// Accumulator for computeMeanAndCovarianceMatrix; feed points in index order.
template <typename T>
struct CovAccum {
  T a[9];
  int n;
  LSB_HD CovAccum() : n(0) { for (int i = 0; i < 9; ++i) a[i] = T(0); }
  LSB_HD void add(float x, float y, float z)
  {
    // products are float*float rounded to float, THEN converted to T
    a[0] += (T)(x * x); a[1] += (T)(x * y); a[2] += (T)(x * z);
    a[3] += (T)(y * y); a[4] += (T)(y * z); a[5] += (T)(z * z);
    a[6] += (T)x;       a[7] += (T)y;       a[8] += (T)z;
    ++n;
  }
  // outputs centroid and symmetric covariance (population, /n)
  LSB_HD void finish(Vec3<T>& centroid, Mat3<T>& cov)
  {
    const T cnt = (T)n;
    for (int i = 0; i < 9; ++i) a[i] = a[i] / cnt;
    centroid = Vec3<T>(a[6], a[7], a[8]);
    cov(0, 0) = a[0] - a[6] * a[6];
    cov(0, 1) = a[1] - a[6] * a[7];
    cov(0, 2) = a[2] - a[6] * a[8];
    cov(1, 1) = a[3] - a[7] * a[7];
    cov(1, 2) = a[4] - a[7] * a[8];
    cov(2, 2) = a[5] - a[8] * a[8];
    cov(1, 0) = cov(0, 1); cov(2, 0) = cov(0, 2); cov(2, 1) = cov(1, 2);
  }
};

template <typename T>
LSB_HD void compute_roots2(T b, T c, T* roots)
{
  roots[0] = T(0);
  // Scalar(b*b - 4.0*c): b*b in T, the subtraction in double
  T d = (T)((double)(b * b) - 4.0 * (double)c);
  if (d < T(0)) d = T(0);
  const T sd = dm::t_sqrt(d);
  roots[2] = T(0.5) * (b + sd);
  roots[1] = T(0.5) * (b - sd);
}

template <typename T>
LSB_HD void compute_roots(const Mat3<T>& m, T* roots)
{
  const T c0 = m(0, 0) * m(1, 1) * m(2, 2) + T(2) * m(0, 1) * m(0, 2) * m(1, 2) - m(0, 0) * m(1, 2) * m(1, 2) -
               m(1, 1) * m(0, 2) * m(0, 2) - m(2, 2) * m(0, 1) * m(0, 1);
  const T c1 = m(0, 0) * m(1, 1) - m(0, 1) * m(0, 1) + m(0, 0) * m(2, 2) - m(0, 2) * m(0, 2) + m(1, 1) * m(2, 2) -
               m(1, 2) * m(1, 2);
  const T c2 = m(0, 0) + m(1, 1) + m(2, 2);
  if (dm::t_abs(c0) < NumEps<T>::eps()) {
    compute_roots2(c2, c1, roots);
    return;
  }
  const T s_inv3 = (T)(1.0 / 3.0);
  const T s_sqrt3 = dm::t_sqrt(T(3.0));
  const T c2_over_3 = c2 * s_inv3;
  T a_over_3 = (c1 - c2 * c2_over_3) * s_inv3;
  if (a_over_3 > T(0)) a_over_3 = T(0);
  const T half_b = T(0.5) * (c0 + c2_over_3 * (T(2) * c2_over_3 * c2_over_3 - c1));
  T q = half_b * half_b + a_over_3 * a_over_3 * a_over_3;
  if (q > T(0)) q = T(0);
  const T rho = dm::t_sqrt(-a_over_3);
  const T theta = dm::t_atan2(dm::t_sqrt(-q), half_b) * s_inv3;
  const T cos_theta = dm::t_cos(theta);
  const T sin_theta = dm::t_sin(theta);
  roots[0] = c2_over_3 + T(2) * rho * cos_theta;
  roots[1] = c2_over_3 - rho * (cos_theta + s_sqrt3 * sin_theta);
  roots[2] = c2_over_3 - rho * (cos_theta - s_sqrt3 * sin_theta);
  T tmp;
  if (roots[0] >= roots[1]) { tmp = roots[0]; roots[0] = roots[1]; roots[1] = tmp; }
  if (roots[1] >= roots[2]) {
    tmp = roots[1]; roots[1] = roots[2]; roots[2] = tmp;
    if (roots[0] >= roots[1]) { tmp = roots[0]; roots[0] = roots[1]; roots[1] = tmp; }
  }
  if (roots[0] <= T(0)) compute_roots2(c2, c1, roots);
}

// Eigen::MatrixBase::unitOrthogonal() for 3-vectors
template <typename T>
LSB_HD Vec3<T> unit_orthogonal(const Vec3<T>& s)
{
  const T prec = NumEps<T>::dummy();
  const bool xsmall = dm::t_abs(s.x) <= dm::t_abs(s.z) * prec;
  const bool ysmall = dm::t_abs(s.y) <= dm::t_abs(s.z) * prec;
  if (!xsmall || !ysmall) {
    const T invnm = T(1) / dm::t_sqrt(s.x * s.x + s.y * s.y);
    return Vec3<T>(-s.y * invnm, s.x * invnm, T(0));
  }
  const T invnm = T(1) / dm::t_sqrt(s.y * s.y + s.z * s.z);
  return Vec3<T>(T(0), -s.z * invnm, s.y * invnm);
}

// "largest of the three row cross products" of (m - ev*I); returns the
// normalised vector and its squared length.
template <typename T>
LSB_HD Vec3<T> eig_vec_from_rows(const Mat3<T>& sm, T ev, T& len)
{
  Mat3<T> tmp = sm;
  tmp(0, 0) -= ev; tmp(1, 1) -= ev; tmp(2, 2) -= ev;
  const Vec3<T> v1 = cross(tmp.row(0), tmp.row(1));
  const Vec3<T> v2 = cross(tmp.row(0), tmp.row(2));
  const Vec3<T> v3 = cross(tmp.row(1), tmp.row(2));
  const T l1 = sqnorm(v1), l2 = sqnorm(v2), l3 = sqnorm(v3);
  if (l1 >= l2 && l1 >= l3) { len = l1; return v1 / dm::t_sqrt(l1); }
  if (l2 >= l1 && l2 >= l3) { len = l2; return v2 / dm::t_sqrt(l2); }
  len = l3;
  return v3 / dm::t_sqrt(l3);
}

// pcl::eigen33(mat, evecs, evals): evals ascending, evecs columns matching.
template <typename T>
LSB_HD void eigen33(const Mat3<T>& mat, Mat3<T>& evecs, T* evals)
{
  T scale = T(0);
  for (int i = 0; i < 9; ++i) {
    const T a = dm::t_abs(mat.m[i]);
    if (a > scale) scale = a;   // maxCoeff: NaN never wins a '>' compare
  }
  if (scale <= NumEps<T>::tiny()) scale = T(1);
  Mat3<T> sm;
  for (int i = 0; i < 9; ++i) sm.m[i] = mat.m[i] / scale;

  compute_roots(sm, evals);

  const T eps = NumEps<T>::eps();
  if ((evals[2] - evals[0]) <= eps) {
    evecs = Mat3<T>::identity();
  } else if ((evals[1] - evals[0]) <= eps) {
    T l;
    const Vec3<T> e2 = eig_vec_from_rows(sm, evals[2], l);
    const Vec3<T> e1 = unit_orthogonal(e2);
    const Vec3<T> e0 = cross(e1, e2);
    evecs.setcol(0, e0); evecs.setcol(1, e1); evecs.setcol(2, e2);
  } else if ((evals[2] - evals[1]) <= eps) {
    T l;
    const Vec3<T> e0 = eig_vec_from_rows(sm, evals[0], l);
    const Vec3<T> e1 = unit_orthogonal(e0);
    const Vec3<T> e2 = cross(e0, e1);
    evecs.setcol(0, e0); evecs.setcol(1, e1); evecs.setcol(2, e2);
  } else {
    T mmax[3];
    Vec3<T> e[3];
    unsigned int min_el = 2, max_el = 2;
    e[2] = eig_vec_from_rows(sm, evals[2], mmax[2]);
    e[1] = eig_vec_from_rows(sm, evals[1], mmax[1]);
    min_el = (mmax[1] <= mmax[min_el]) ? 1u : min_el;
    max_el = (mmax[1] > mmax[max_el]) ? 1u : max_el;
    e[0] = eig_vec_from_rows(sm, evals[0], mmax[0]);
    min_el = (mmax[0] <= mmax[min_el]) ? 0u : min_el;
    max_el = (mmax[0] > mmax[max_el]) ? 0u : max_el;
    if (min_el != max_el) {  // equal only when NaNs poisoned the compares (UB upstream)
      const unsigned int mid_el = 3u - min_el - max_el;
      e[min_el] = normalized(cross(e[(min_el + 1) % 3], e[(min_el + 2) % 3]));
      e[mid_el] = normalized(cross(e[(mid_el + 1) % 3], e[(mid_el + 2) % 3]));
    }
    evecs.setcol(0, e[0]); evecs.setcol(1, e[1]); evecs.setcol(2, e[2]);
  }
  evals[0] *= scale; evals[1] *= scale; evals[2] *= scale;
}

}  // namespace lsb
