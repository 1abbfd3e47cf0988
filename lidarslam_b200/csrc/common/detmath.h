// Deterministic elementary functions shared by the CUDA kernels (nvcc, sm_100a)
// and the CPU oracle (g++).
//
// Why: the reference calls libm's atan2/cos/sin inside pcl::eigen33 (called from
// Utils::ComputeMeanAndPCA, reference slam_lib/include/LidarSlam/Utilities.h:247-262),
// acos/sin inside Eigen's slerp (MotionModel.h:115-124) and sin/cos in
// RotationMatrixFromRPY (CeresCostFunctions.h:67-79).  glibc and CUDA libm differ
// in last-ulp results, and keypoint labels / match statuses are decided by
// threshold compares on values derived from them.  To make GPU and oracle
// bit-identical both sides evaluate THESE functions, which use only IEEE-754
// +,-,*,/,sqrt,floor in a fixed order (compile with -ffp-contract=off / -fmad=false).
// Accuracy is ~1 ulp in double (checked against libm in tests/test_detmath.py);
// the float variants evaluate in double and round once, i.e. they are
// (almost always) correctly rounded like glibc's sinf/cosf/atan2f.
#pragma once
#include <math.h>
#include <string.h>

#if defined(__CUDACC__)
#define LSB_HD __host__ __device__ __forceinline__
#else
#define LSB_HD inline
#endif

namespace lsb {
namespace dm {

// --- kernels on |y| <= pi/4 (fdlibm-style minimax polynomials) ---------------
LSB_HD double ksin(double y)
{
  const double S1 = -1.66666666666666324348e-01, S2 = 8.33333333332248946124e-03,
               S3 = -1.98412698298579493134e-04, S4 = 2.75573137070700676789e-06,
               S5 = -2.50507602534068634195e-08, S6 = 1.58969099521155010221e-10;
  const double z = y * y;
  const double v = z * y;
  const double r = S2 + z * (S3 + z * (S4 + z * (S5 + z * S6)));
  return y + v * (S1 + z * r);
}

LSB_HD double kcos(double y)
{
  const double C1 = 4.16666666666666019037e-02, C2 = -1.38888888888741095749e-03,
               C3 = 2.48015872894767294178e-05, C4 = -2.75573143513906633035e-07,
               C5 = 2.08757232129817482790e-09, C6 = -1.13596475577881948265e-11;
  const double z = y * y;
  const double r = z * (C1 + z * (C2 + z * (C3 + z * (C4 + z * (C5 + z * C6)))));
  const double hz = 0.5 * z;
  const double w = 1.0 - hz;
  // 1 - (hz - z*r) with the rounding error of (1 - hz) folded back in
  return w + (((1.0 - w) - hz) + z * r);
}

// Cody-Waite reduction x = n*(pi/2) + y, |y| <= pi/4 (+eps). Valid for |x| < ~1e6.
LSB_HD double reduce_pio2(double x, int& q)
{
  const double invpio2 = 6.36619772367581382433e-01;
  const double pio2_1 = 1.57079632673412561417e+00;   // first 33 bits of pi/2
  const double pio2_2 = 6.07710050630396597660e-11;   // next 33 bits
  const double pio2_3 = 2.02226624871116645580e-21;   // next 33 bits
  const double pio2_3t = 8.47842766036889956997e-32;  // tail
  const double fn = floor(x * invpio2 + 0.5);
  double y = x - fn * pio2_1;
  y = y - fn * pio2_2;
  y = y - fn * pio2_3;
  y = y - fn * pio2_3t;
  // n mod 4, robust to negative fn
  const double m = fn - 4.0 * floor(fn * 0.25);
  q = (int)m;
  return y;
}

LSB_HD double sin(double x)
{
  int q;
  const double y = reduce_pio2(x, q);
  if (q == 0) return ksin(y);
  if (q == 1) return kcos(y);
  if (q == 2) return -ksin(y);
  return -kcos(y);
}

LSB_HD double cos(double x)
{
  int q;
  const double y = reduce_pio2(x, q);
  if (q == 0) return kcos(y);
  if (q == 1) return -ksin(y);
  if (q == 2) return -kcos(y);
  return ksin(y);
}

// atan for x >= 0 (fdlibm breakpoints 7/16, 11/16, 19/16, 39/16)
LSB_HD double atan_pos(double x)
{
  const double hi0 = 4.63647609000806093515e-01, lo0 = 2.26987774529616870924e-17;
  const double hi1 = 7.85398163397448278999e-01, lo1 = 3.06161699786838301793e-17;
  const double hi2 = 9.82793723247329054082e-01, lo2 = 1.39033110312309984516e-17;
  const double hi3 = 1.57079632679489655800e+00, lo3 = 6.12323399573676603587e-17;
  const double aT0 = 3.33333333333329318027e-01, aT1 = -1.99999999998764832476e-01,
               aT2 = 1.42857142725034663711e-01, aT3 = -1.11111104054623557880e-01,
               aT4 = 9.09088713343650656196e-02, aT5 = -7.69187620504482999495e-02,
               aT6 = 6.66107313738753120669e-02, aT7 = -5.83357013379057348645e-02,
               aT8 = 4.97687799461593236017e-02, aT9 = -3.65315727442169155270e-02,
               aT10 = 1.62858201153657823623e-02;
  int id;
  double t;
  if (x < 0.4375) { id = -1; t = x; }
  else if (x < 0.6875) { id = 0; t = (2.0 * x - 1.0) / (2.0 + x); }
  else if (x < 1.1875) { id = 1; t = (x - 1.0) / (x + 1.0); }
  else if (x < 2.4375) { id = 2; t = (x - 1.5) / (1.0 + 1.5 * x); }
  else if (x < 1.0e300) { id = 3; t = -1.0 / x; }
  else { return hi3 + lo3; }
  const double z = t * t;
  const double w = z * z;
  const double s1 = z * (aT0 + w * (aT2 + w * (aT4 + w * (aT6 + w * (aT8 + w * aT10)))));
  const double s2 = w * (aT1 + w * (aT3 + w * (aT5 + w * (aT7 + w * aT9))));
  if (id < 0) return t - t * (s1 + s2);
  double hi, lo;
  if (id == 0) { hi = hi0; lo = lo0; }
  else if (id == 1) { hi = hi1; lo = lo1; }
  else if (id == 2) { hi = hi2; lo = lo2; }
  else { hi = hi3; lo = lo3; }
  return hi - ((t * (s1 + s2) - lo) - t);
}

LSB_HD double atan(double x) { return (x < 0.0) ? -atan_pos(-x) : atan_pos(x); }

// atan2 with the usual quadrant conventions; (0,0) -> 0 ; NaN propagates.
LSB_HD double atan2(double y, double x)
{
  const double pi = 3.14159265358979311600e+00, pi_lo = 1.22464679914735317720e-16;
  const double pio2 = 1.57079632679489655800e+00;
  if (x != x || y != y) return x + y;
  if (y == 0.0) return (x < 0.0) ? ((1.0 / y < 0.0) ? -pi : pi) : y;  // keeps sign of zero for x>=0
  if (x == 0.0) return (y < 0.0) ? -pio2 : pio2;
  const double ay = (y < 0.0) ? -y : y;
  const double ax = (x < 0.0) ? -x : x;
  double z = atan_pos(ay / ax);
  if (x < 0.0) z = pi - (z - pi_lo);
  return (y < 0.0) ? -z : z;
}

LSB_HD double asin(double x)
{
  // asin(x) = atan2(x, sqrt((1-x)(1+x))) ; |x|>1 -> NaN via sqrt of negative
  return atan2(x, ::sqrt((1.0 - x) * (1.0 + x)));
}

LSB_HD double acos(double x)
{
  return atan2(::sqrt((1.0 - x) * (1.0 + x)), x);
}

// float variants: evaluate in double, round once
// exp(x): Cody-Waite reduction x = n ln2 + r, |r| <= ln2/2, Taylor polynomial of degree 13 in
// Horner form (remainder < 1e-17 relative), scaling by 2^n through the exponent bits.  Used by the
// overlap estimator (std::exp in ConfidenceEstimators.cxx:55); inputs there are <= 0.
LSB_HD double exp(double x)
{
  if (!(x == x)) return x;
  if (x > 709.0) return INFINITY;
  if (x < -745.0) return 0.0;
  const double ln2_hi = 6.93147180369123816490e-01, ln2_lo = 1.90821492927058770002e-10, inv_ln2 = 1.44269504088896338700e+00;
  const double fn = floor(x * inv_ln2 + 0.5);
  const double r = (x - fn * ln2_hi) - fn * ln2_lo;
  double p = 1.0 / 6227020800.0;
  p = p * r + 1.0 / 479001600.0;
  p = p * r + 1.0 / 39916800.0;
  p = p * r + 1.0 / 3628800.0;
  p = p * r + 1.0 / 362880.0;
  p = p * r + 1.0 / 40320.0;
  p = p * r + 1.0 / 5040.0;
  p = p * r + 1.0 / 720.0;
  p = p * r + 1.0 / 120.0;
  p = p * r + 1.0 / 24.0;
  p = p * r + 1.0 / 6.0;
  p = p * r + 0.5;
  p = p * r + 1.0;
  p = p * r + 1.0;
  // 2^n in two exact factors so that results in the subnormal range round once
  int n = (int)fn;
  double scale = 1.0;
  if (n < -1000) { scale = 1.4916681462400413e-154; n += 512; }   // 2^-512
#ifdef __CUDA_ARCH__
  const double two_n = __longlong_as_double((long long)(n + 1023) << 52);
#else
  const long long bits = (long long)(n + 1023) << 52;
  double two_n;
  memcpy(&two_n, &bits, 8);
#endif
  return (p * two_n) * scale;
}
LSB_HD float expf(float x) { return (float)exp((double)x); }

LSB_HD float sinf(float x) { return (float)sin((double)x); }
LSB_HD float cosf(float x) { return (float)cos((double)x); }
LSB_HD float atan2f(float y, float x) { return (float)atan2((double)y, (double)x); }
LSB_HD float acosf(float x) { return (float)acos((double)x); }

// templated front-ends used by the PCA code
template <typename T> LSB_HD T t_sin(T x);
template <> LSB_HD double t_sin<double>(double x) { return sin(x); }
template <> LSB_HD float t_sin<float>(float x) { return sinf(x); }
template <typename T> LSB_HD T t_cos(T x);
template <> LSB_HD double t_cos<double>(double x) { return cos(x); }
template <> LSB_HD float t_cos<float>(float x) { return cosf(x); }
template <typename T> LSB_HD T t_atan2(T y, T x);
template <> LSB_HD double t_atan2<double>(double y, double x) { return atan2(y, x); }
template <> LSB_HD float t_atan2<float>(float y, float x) { return atan2f(y, x); }
template <typename T> LSB_HD T t_sqrt(T x);
template <> LSB_HD double t_sqrt<double>(double x) { return ::sqrt(x); }
template <> LSB_HD float t_sqrt<float>(float x) { return ::sqrtf(x); }
template <typename T> LSB_HD T t_abs(T x) { return (x < T(0)) ? -x : x; }

}  // namespace dm
}  // namespace lsb
