// LocalOptimizer::EstimateRegistrationError (reference LocalOptimizer.cxx:112-140):
// ceres::Covariance(DENSE_SVD, apply_loss_function, null_space_rank = -1) of the
// 6x6 block at the solution, then the max eigenpairs of the position and
// orientation 3x3 blocks.  Host-side code (runs once per frame on 28 doubles read
// back from the device); used by the product driver and by the oracle.
//
// Ceres' DENSE_SVD covariance is V diag(1/sigma_i^2) V^T over the singular values
// of J~ with sigma_i/sigma_max >= sqrt(1e-14); equivalently over the eigenpairs
// (lambda_i = sigma_i^2) of J~^T J~, which is what is available here.
// Eigenvector signs are arbitrary upstream (Eigen::SelfAdjointEigenSolver); the
// canonical sign here makes the largest-magnitude component positive.
#pragma once
#include <cmath>
#include <cstring>
#include "../../../include/lidarslam_b200.h"

namespace lsb {

// cyclic Jacobi eigen-decomposition of a symmetric n x n matrix (n <= 6), row-major.
// On return a holds eigenvalues on its diagonal, v the eigenvectors in columns.
inline void jacobi_eigen(int n, double* a, double* v)
{
  for (int i = 0; i < n; ++i)
    for (int j = 0; j < n; ++j) v[i * n + j] = (i == j) ? 1.0 : 0.0;
  for (int sweep = 0; sweep < 64; ++sweep) {
    double off = 0.0;
    for (int i = 0; i < n; ++i)
      for (int j = i + 1; j < n; ++j) off += a[i * n + j] * a[i * n + j];
    if (off == 0.0) break;
    for (int p = 0; p < n; ++p)
      for (int q = p + 1; q < n; ++q) {
        const double apq = a[p * n + q];
        if (apq == 0.0) continue;
        const double theta = (a[q * n + q] - a[p * n + p]) / (2.0 * apq);
        const double t = ((theta >= 0.0) ? 1.0 : -1.0) / (std::fabs(theta) + std::sqrt(theta * theta + 1.0));
        const double c = 1.0 / std::sqrt(t * t + 1.0), s = t * c;
        for (int k = 0; k < n; ++k) {
          const double akp = a[k * n + p], akq = a[k * n + q];
          a[k * n + p] = c * akp - s * akq;
          a[k * n + q] = s * akp + c * akq;
        }
        for (int k = 0; k < n; ++k) {
          const double apk = a[p * n + k], aqk = a[q * n + k];
          a[p * n + k] = c * apk - s * aqk;
          a[q * n + k] = s * apk + c * aqk;
        }
        for (int k = 0; k < n; ++k) {
          const double vkp = v[k * n + p], vkq = v[k * n + q];
          v[k * n + p] = c * vkp - s * vkq;
          v[k * n + q] = s * vkp + c * vkq;
        }
      }
  }
}

inline void max_eigenpair3(const double* cov6_colmajor, int off, double& val, double* dir)
{
  double a[9], v[9];
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) a[i * 3 + j] = cov6_colmajor[(off + j) * 6 + (off + i)];
  jacobi_eigen(3, a, v);
  int m = 0;
  for (int i = 1; i < 3; ++i) if (a[i * 3 + i] > a[m * 3 + m]) m = i;
  val = a[m * 3 + m];
  int big = 0;
  for (int i = 1; i < 3; ++i) if (std::fabs(v[i * 3 + m]) > std::fabs(v[big * 3 + m])) big = i;
  const double sgn = (v[big * 3 + m] < 0.0) ? -1.0 : 1.0;
  for (int i = 0; i < 3; ++i) dir[i] = sgn * v[i * 3 + m];
}

// H21 = upper triangle (row-major) of J~^T J~ at the solution.
inline void registration_error_from_normal_matrix(const double* H21, int two_d, lsb_registration_error* e)
{
  int map[6], n = 0;
  for (int i = 0; i < 6; ++i) if (!(two_d && (i == 2 || i == 3 || i == 4))) map[n++] = i;
  double a[36], v[36];
  auto hidx = [](int i, int j) { return i * 6 - (i * (i - 1)) / 2 + (j - i); };
  for (int i = 0; i < n; ++i)
    for (int j = 0; j < n; ++j) {
      const int I = map[i], J = map[j];
      a[i * n + j] = (I <= J) ? H21[hidx(I, J)] : H21[hidx(J, I)];
    }
  jacobi_eigen(n, a, v);
  double lmax = 0.0;
  for (int i = 0; i < n; ++i) lmax = (a[i * n + i] > lmax) ? a[i * n + i] : lmax;
  std::memset(e->covariance, 0, sizeof(e->covariance));
  for (int k = 0; k < n; ++k) {
    const double l = a[k * n + k];
    if (!(l > 0.0) || std::sqrt(l / lmax) < 1e-7) continue;   // min_reciprocal_condition_number = 1e-14
    const double inv = 1.0 / l;
    for (int i = 0; i < n; ++i)
      for (int j = 0; j < n; ++j) e->covariance[map[j] * 6 + map[i]] += v[i * n + k] * inv * v[j * n + k];
  }
  double pv, ov;
  max_eigenpair3(e->covariance, 0, pv, e->position_error_direction);
  max_eigenpair3(e->covariance, 3, ov, e->orientation_error_direction);
  e->position_error = std::sqrt(pv);
  e->orientation_error = std::sqrt(ov) / M_PI * 180.;
}

}  // namespace lsb
