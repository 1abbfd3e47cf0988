// Gauss-Newton / Levenberg-Marquardt core replacing Ceres on the ICP path
// (reference slam_lib/src/LocalOptimizer.cxx:74-102 configures ceres::Solve with
// TRUST_REGION + LEVENBERG_MARQUARDT + DENSE_QR, max_num_iterations = LMMaxIter,
// every other option default; the residual is CeresCostFunctions.h:105-152 with
// ScaledLoss(TukeyLoss) from KeypointsMatcher.cxx:78-103).
//
// Ceres is NOT under /root/reference.  This restates the published Ceres 2.x
// algorithm (trust_region_minimizer.cc, levenberg_marquardt_strategy.cc,
// corrector.cc, loss_function.cc): Jacobi column scaling frozen at iteration 0,
// LM diagonal clamp [1e-6, 1e32], step-quality radius update, function /
// parameter / gradient tolerances 1e-6 / 1e-8 / 1e-10.  Deviation: the damped
// system is solved by Cholesky on the 6x6 normal equations instead of QR on the
// stacked Jacobian (same minimiser, different rounding), and the Jacobian is
// analytic instead of autodiff Jets (same values up to rounding).
//
// Shared by the CUDA LM kernel and the CPU oracle so that both take identical
// accept/reject decisions; compiled with -fmad=false / -ffp-contract=off.
#pragma once
#include "linalg.h"

namespace lsb {

// 28 reduced quantities of one cost/Jacobian evaluation:
// [0..20] upper triangle of J~^T J~ (row-major: 00 01 02 03 04 05 11 12 ...),
// [21..26] J~^T r~, [27] cost = sum 0.5*rho.
#define LSB_LM_NSUM 28
// Canonical reduction order (GPU and oracle): a radix-32 tree.  fold32(v[0..31]) is the pairwise
// tree  for off in 16,8,4,2,1: v[i] += v[i+off] (i < off);  missing leaves are +0.0.
//   group sum  G_g = fold32 of the records 32g .. 32g+31
//   chunk sum  C_c = fold32 of G_32c .. G_32c+31          (1024 records, one CTA)
//   block sum  B_b = fold32 of C_32b .. C_32b+31
//   E          = B_0 + B_1 + ...                          (in order)
#define LSB_LM_CHUNK 1024

// One residual record (13 doubles): the residual of the reference is r = A (R X + t - P)
// (CeresCostFunctions.h:105-152); only M = A^T A enters the normal equations and the cost
// (r.r = d^T M d, J^T J = Jd^T M Jd, J^T r = Jd^T M d with d = R X + t - P), so the record keeps
// the 6 unique entries of M.  w < 0 marks "no residual" (match failed): contributes exactly zero.
struct LmRecord {
  double M[6];   // 00 01 02 11 12 22
  double P[3];
  double X[3];
  double w;
};

// M = A^T A, each entry summed over the rows of A in order (shared by k_fit and the oracle)
LSB_HD void lm_record_set(LmRecord& r, const double* A, const double* P, const double* X, double w)
{
  int o = 0;
  for (int i = 0; i < 3; ++i)
    for (int j = i; j < 3; ++j) r.M[o++] = A[0 + i] * A[0 + j] + A[3 + i] * A[3 + j] + A[6 + i] * A[6 + j];
  for (int k = 0; k < 3; ++k) { r.P[k] = P[k]; r.X[k] = X[k]; }
  r.w = w;
}

// Quantities that depend only on the current parameters.  R = Rz(rz) Ry(ry) Rx(rx); its
// derivatives are cross products with the rotated point Y = R X:
//   dR/drx X = (first column of R) x Y,   dR/dry X = (-sz, cz, 0) x Y,   dR/drz X = (0, 0, 1) x Y.
struct LmPoint {
  Mat3d R;
  Vec3d t;
  double sz, cz;
};

// trig[6] = cos rx, sin rx, cos ry, sin ry, cos rz, sin rz  (entry j: angle x[3 + j/2], cos if j even)
LSB_HD double lm_point_trig(const double* x, int j)
{
  const double a = x[3 + (j >> 1)];
  return (j & 1) ? dm::sin(a) : dm::cos(a);
}

LSB_HD void lm_make_point_from_trig(const double* x, const double* trig, LmPoint& p)
{
  const double cx = trig[0], sx = trig[1];
  const double cy = trig[2], sy = trig[3];
  const double cz = trig[4], sz = trig[5];
  p.t = Vec3d(x[0], x[1], x[2]);
  Mat3d& R = p.R;
  R(0, 0) = cy * cz; R(0, 1) = sx * sy * cz - cx * sz; R(0, 2) = cx * sy * cz + sx * sz;
  R(1, 0) = cy * sz; R(1, 1) = sx * sy * sz + cx * cz; R(1, 2) = cx * sy * sz - sx * cz;
  R(2, 0) = -sy;     R(2, 1) = sx * cy;                R(2, 2) = cx * cy;
  p.sz = sz;
  p.cz = cz;
}

LSB_HD void lm_make_point(const double* x, LmPoint& p)
{
  double trig[6];
  for (int j = 0; j < 6; ++j) trig[j] = lm_point_trig(x, j);
  lm_make_point_from_trig(x, trig, p);
}

// Adds nothing: WRITES the 28 terms of one record into out[] (zeros if w < 0).
LSB_HD void lm_eval_record(const LmRecord& rec, const LmPoint& pt, double sat, double* out)
{
  if (!(rec.w >= 0.0)) {
    for (int i = 0; i < LSB_LM_NSUM; ++i) out[i] = 0.0;
    return;
  }
  const double m00 = rec.M[0], m01 = rec.M[1], m02 = rec.M[2], m11 = rec.M[3], m12 = rec.M[4], m22 = rec.M[5];
  const Vec3d X(rec.X[0], rec.X[1], rec.X[2]);
  const Vec3d Y = pt.R * X;                                   // rotated point (no translation)
  const double d0 = (Y.x + pt.t.x) - rec.P[0], d1 = (Y.y + pt.t.y) - rec.P[1], d2 = (Y.z + pt.t.z) - rec.P[2];
  // u = M d,  s = d.u = |A d|^2
  const double u0 = m00 * d0 + m01 * d1 + m02 * d2;
  const double u1 = m01 * d0 + m11 * d1 + m12 * d2;
  const double u2 = m02 * d0 + m12 * d1 + m22 * d2;
  const double s = d0 * u0 + d1 * u1 + d2 * u2;
  // TukeyLoss(sat) scaled by w (Ceres >= 2.0 form, KeypointsMatcher.cxx:96-101)
  const double a2 = sat * sat;
  double rho0, rho1;
  if (s <= a2) {
    const double v = 1.0 - s / a2;
    const double v2 = v * v;
    rho0 = a2 / 3.0 * (1.0 - v2 * v);
    rho1 = v2;
  } else {
    rho0 = a2 / 3.0;
    rho1 = 0.0;
  }
  rho0 = rec.w * rho0;
  rho1 = rec.w * rho1;
  // Ceres corrector with rho'' <= 0: residual and Jacobian scaled by sqrt(rho'), i.e. the normal
  // equations scaled by rho'.  Rotation columns of Jd = d(d)/d(rx, ry, rz):
  const double ax = pt.R.m[0], ay = pt.R.m[3], az = pt.R.m[6];            // first column of R
  const double cx0 = ay * Y.z - az * Y.y, cx1 = az * Y.x - ax * Y.z, cx2 = ax * Y.y - ay * Y.x;
  const double cy0 = pt.cz * Y.z, cy1 = pt.sz * Y.z, cy2 = -(pt.sz * Y.y) - pt.cz * Y.x;
  const double cz0 = -Y.y, cz1 = Y.x;                                      // cz2 = 0
  // v_k = M c_k
  const double vx0 = m00 * cx0 + m01 * cx1 + m02 * cx2, vx1 = m01 * cx0 + m11 * cx1 + m12 * cx2, vx2 = m02 * cx0 + m12 * cx1 + m22 * cx2;
  const double vy0 = m00 * cy0 + m01 * cy1 + m02 * cy2, vy1 = m01 * cy0 + m11 * cy1 + m12 * cy2, vy2 = m02 * cy0 + m12 * cy1 + m22 * cy2;
  const double vz0 = m00 * cz0 + m01 * cz1, vz1 = m01 * cz0 + m11 * cz1, vz2 = m02 * cz0 + m12 * cz1;
  // J~^T J~ (upper triangle, parameter order tx ty tz rx ry rz)
  out[0] = rho1 * m00; out[1] = rho1 * m01; out[2] = rho1 * m02; out[3] = rho1 * vx0; out[4] = rho1 * vy0; out[5] = rho1 * vz0;
  out[6] = rho1 * m11; out[7] = rho1 * m12; out[8] = rho1 * vx1; out[9] = rho1 * vy1; out[10] = rho1 * vz1;
  out[11] = rho1 * m22; out[12] = rho1 * vx2; out[13] = rho1 * vy2; out[14] = rho1 * vz2;
  out[15] = rho1 * (cx0 * vx0 + cx1 * vx1 + cx2 * vx2);
  out[16] = rho1 * (cx0 * vy0 + cx1 * vy1 + cx2 * vy2);
  out[17] = rho1 * (cx0 * vz0 + cx1 * vz1 + cx2 * vz2);
  out[18] = rho1 * (cy0 * vy0 + cy1 * vy1 + cy2 * vy2);
  out[19] = rho1 * (cy0 * vz0 + cy1 * vz1 + cy2 * vz2);
  out[20] = rho1 * (cz0 * vz0 + cz1 * vz1);
  // J~^T r~
  out[21] = rho1 * u0; out[22] = rho1 * u1; out[23] = rho1 * u2;
  out[24] = rho1 * (cx0 * u0 + cx1 * u1 + cx2 * u2);
  out[25] = rho1 * (cy0 * u0 + cy1 * u1 + cy2 * u2);
  out[26] = rho1 * (cz0 * u0 + cz1 * u1);
  out[27] = 0.5 * rho0;
}

// Cholesky solve of the symmetric 6x6 system M y = b (M full, row-major; only the lower
// triangle is read).  Returns false if M is not positive definite or the result is not finite.
// Canonical operation order (the CUDA control step runs the same operations spread over the
// lanes of a warp, see lm_propose_warp in cuda/match.cu, and must round identically):
//   L[i][j] = (M[i][j] - sum_{k<j, ascending} L[i][k] L[j][k]) * (1 / L[j][j])
//   z[i]    = (b[i]    - sum_{k<i, ascending}  L[i][k] z[k])   * (1 / L[i][i])
//   y[i]    = (z[i]    - sum_{k>i, DESCENDING} L[k][i] y[k])   * (1 / L[i][i])
LSB_HD bool lm_chol_solve6(const double* M, const double* b, double* y)
{
  double L[36], inv[6];
  bool ok = true;
#pragma unroll
  for (int i = 0; i < 36; ++i) L[i] = 0.0;
#pragma unroll
  for (int j = 0; j < 6; ++j) {
    double d = M[j * 6 + j];
#pragma unroll
    for (int k = 0; k < j; ++k) d = d - L[j * 6 + k] * L[j * 6 + k];
    if (!(d > 0.0)) ok = false;
    const double ljj = ::sqrt(d);
    inv[j] = 1.0 / ljj;
    L[j * 6 + j] = ljj;
#pragma unroll
    for (int i = j + 1; i < 6; ++i) {
      double s = M[i * 6 + j];
#pragma unroll
      for (int k = 0; k < j; ++k) s = s - L[i * 6 + k] * L[j * 6 + k];
      L[i * 6 + j] = s * inv[j];
    }
  }
  if (!ok) return false;
  double z[6];
#pragma unroll
  for (int i = 0; i < 6; ++i) {
    double s = b[i];
#pragma unroll
    for (int k = 0; k < i; ++k) s = s - L[i * 6 + k] * z[k];
    z[i] = s * inv[i];
  }
#pragma unroll
  for (int i = 5; i >= 0; --i) {
    double s = z[i];
#pragma unroll
    for (int k = 5; k > i; --k) s = s - L[k * 6 + i] * y[k];
    y[i] = s * inv[i];
  }
#pragma unroll
  for (int i = 0; i < 6; ++i)
    if (!(y[i] - y[i] == 0.0)) ok = false;   // NaN / inf
  return ok;
}

enum { LSB_LM_RUNNING = 0, LSB_LM_CONV_GRADIENT, LSB_LM_CONV_PARAMETER, LSB_LM_CONV_FUNCTION, LSB_LM_CONV_RADIUS,
       LSB_LM_MAX_ITER, LSB_LM_FAIL_INVALID, LSB_LM_SKIPPED /* not enough matches: never started */ };

// Trust-region LM state machine.  Driver:
//   lm_begin(s, x0, max_iter, two_d); E = evaluate(s.x); lm_first_eval(s, E);
//   while (!s.done) { E = evaluate(s.xc); lm_next_eval(s, E); }
// The solution is s.x; s.H/s.g/s.cost always describe s.x.
struct LmState {
  double x[6];
  double xc[6];           // candidate point awaiting evaluation
  double cost;
  double H[21];           // J~^T J~ at x (upper, row-major)
  double g[6];            // J~^T r~ at x
  double S[6];            // Jacobi scaling, frozen at iteration 0
  double radius, decrease_factor;
  double model_cost_change;
  double step_norm;
  int iter;               // LM iterations started (Ceres iteration index)
  int max_iter;
  int num_successful;     // accepted steps (iteration 0 NOT counted)
  int num_invalid;        // consecutive invalid steps
  int two_d;              // hold z, rx, ry (LocalOptimizer.cxx:89-90)
  int done;
  int term;
};

LSB_HD bool lm_free(const LmState& s, int i) { return !(s.two_d && (i == 2 || i == 3 || i == 4)); }

LSB_HD int lm_hidx(int i, int j)  // index in the 21-entry upper triangle, i <= j
{
  return i * 6 - (i * (i - 1)) / 2 + (j - i);
}

// Computes the next candidate (or terminates).  Loops over invalid steps, which
// consume iterations without needing an evaluation.
LSB_HD void lm_propose(LmState& s)
{
  while (true) {
    if (s.iter >= s.max_iter) { s.done = 1; s.term = LSB_LM_MAX_ITER; return; }
    if (s.radius < 1e-32) { s.done = 1; s.term = LSB_LM_CONV_RADIUS; return; }
    s.iter += 1;
    double M[36], b[6], Hs[36], gs[6];
#pragma unroll
    for (int i = 0; i < 6; ++i) {
#pragma unroll
      for (int j = 0; j < 6; ++j) {
        const double h = (i <= j) ? s.H[lm_hidx(i, j)] : s.H[lm_hidx(j, i)];
        Hs[i * 6 + j] = (lm_free(s, i) && lm_free(s, j)) ? h * s.S[i] * s.S[j] : 0.0;
        M[i * 6 + j] = Hs[i * 6 + j];
      }
      gs[i] = lm_free(s, i) ? s.g[i] * s.S[i] : 0.0;
      b[i] = -gs[i];
    }
#pragma unroll
    for (int i = 0; i < 6; ++i) {
      if (!lm_free(s, i)) { M[i * 6 + i] = 1.0; continue; }   // decoupled dummy row -> delta_i = 0
      double d = Hs[i * 6 + i];
      d = (d < 1e-6) ? 1e-6 : d;
      d = (d > 1e32) ? 1e32 : d;
      M[i * 6 + i] = Hs[i * 6 + i] + d / s.radius;
    }
    double db[6];
    bool ok = lm_chol_solve6(M, b, db);
    double mcc = 0.0;
    if (ok) {
      // model_cost_change = -delta^T (g + 0.5 H delta)   (scaled quantities)
#pragma unroll
      for (int i = 0; i < 6; ++i) {
        double hd = 0.0;
#pragma unroll
        for (int j = 0; j < 6; ++j) hd += Hs[i * 6 + j] * db[j];
        mcc -= db[i] * (gs[i] + 0.5 * hd);
      }
      ok = (mcc > 0.0);
    }
    if (!ok) {
      s.num_invalid += 1;
      if (s.num_invalid >= 5) { s.done = 1; s.term = LSB_LM_FAIL_INVALID; return; }
      s.radius = s.radius * 0.5;          // LevenbergMarquardtStrategy::StepIsInvalid: halve, decrease_factor untouched
      continue;
    }
    s.num_invalid = 0;
    s.model_cost_change = mcc;
    double sn = 0.0;
#pragma unroll
    for (int i = 0; i < 6; ++i) {
      const double d = db[i] * s.S[i];
      s.xc[i] = s.x[i] + d;
      const double e = s.x[i] - s.xc[i];
      sn += e * e;
    }
    s.step_norm = ::sqrt(sn);
    return;
  }
}

LSB_HD void lm_begin(LmState& s, const double* x0, int max_iter, int two_d)
{
  for (int i = 0; i < 6; ++i) { s.x[i] = x0[i]; s.xc[i] = x0[i]; s.g[i] = 0.0; s.S[i] = 1.0; }
  for (int i = 0; i < 21; ++i) s.H[i] = 0.0;
  s.cost = 0.0;
  s.radius = 1e4;
  s.decrease_factor = 2.0;
  s.model_cost_change = 0.0;
  s.step_norm = 0.0;
  s.iter = 0;
  s.max_iter = max_iter;
  s.num_successful = 0;
  s.num_invalid = 0;
  s.two_d = two_d;
  s.done = 0;
  s.term = LSB_LM_RUNNING;
}

LSB_HD double lm_grad_max(const LmState& s)
{
  double m = 0.0;
  for (int i = 0; i < 6; ++i) {
    if (!lm_free(s, i)) continue;
    const double a = (s.g[i] < 0.0) ? -s.g[i] : s.g[i];
    m = (a > m) ? a : m;
  }
  return m;
}

// E = the 28 sums evaluated at s.x (iteration 0).  Returns true when a first step has to be
// proposed (lm_propose), false when the solve is over.
LSB_HD bool lm_first_update(LmState& s, const double* E)
{
  for (int i = 0; i < 21; ++i) s.H[i] = E[i];
  for (int i = 0; i < 6; ++i) s.g[i] = E[21 + i];
  s.cost = E[27];
  for (int i = 0; i < 6; ++i) s.S[i] = 1.0 / (1.0 + ::sqrt(s.H[lm_hidx(i, i)]));
  if (lm_grad_max(s) <= 1e-10) { s.done = 1; s.term = LSB_LM_CONV_GRADIENT; return false; }
  return true;
}

// E = the 28 sums evaluated at s.xc: accepts or rejects the candidate.  Returns true when the
// next step has to be proposed, false when the solve is over.
LSB_HD bool lm_next_update(LmState& s, const double* E)
{
  const double cost_c = E[27];
  double xn = 0.0;
  for (int i = 0; i < 6; ++i) xn += s.x[i] * s.x[i];
  xn = ::sqrt(xn);
  if (s.step_norm <= 1e-8 * (xn + 1e-8)) { s.done = 1; s.term = LSB_LM_CONV_PARAMETER; return false; }
  const double cost_change = s.cost - cost_c;
  const double acc = (cost_change < 0.0) ? -cost_change : cost_change;
  if (acc <= 1e-6 * s.cost) { s.done = 1; s.term = LSB_LM_CONV_FUNCTION; return false; }
  const double rho = cost_change / s.model_cost_change;
  if (rho > 1e-3) {
    for (int i = 0; i < 6; ++i) { s.x[i] = s.xc[i]; s.g[i] = E[21 + i]; }
    for (int i = 0; i < 21; ++i) s.H[i] = E[i];
    s.cost = cost_c;
    s.num_successful += 1;
    const double q = 2.0 * rho - 1.0;
    double den = 1.0 - q * q * q;
    den = (den < 1.0 / 3.0) ? 1.0 / 3.0 : den;
    s.radius = s.radius / den;
    s.radius = (s.radius > 1e16) ? 1e16 : s.radius;
    s.decrease_factor = 2.0;
    if (lm_grad_max(s) <= 1e-10) { s.done = 1; s.term = LSB_LM_CONV_GRADIENT; return false; }
  } else {
    s.radius = s.radius / s.decrease_factor;
    s.decrease_factor *= 2.0;
  }
  return true;
}

// Extra residual blocks of the localization solve (Slam.cxx:1123-1131), at most one of each:
//  * wheel odometer (CeresCostFunctions.h:267-296): r = |t - prev| - dist (|.| := 0 with zero derivative when
//    |t - prev|^2 < 1e-6), loss ScaledLoss(NULL, weight) (SensorConstraints.cxx:51): rho(s) = weight * s;
//  * IMU gravity (CeresCostFunctions.h:311-341): r = R(rx, ry, rz) g_cur - g_ref, same loss.
// With rho' = weight and rho'' = 0 Ceres scales residual and Jacobian by sqrt(weight): the normal equations
// and the cost gain weight * (J^T J, J^T r, 0.5 r.r).  The terms are ADDED to the 28 reduced sums E of the
// keypoint records, after the reduction tree, wheel first (the order the reference adds the blocks).
struct LmSensorTerms {
  int use_wheel, use_gravity;
  double wheel_prev[3], wheel_dist, wheel_weight;
  double g_ref[3], g_cur[3], gravity_weight;
};

LSB_HD void lm_sensor_terms_clear(LmSensorTerms& c)
{
  c.use_wheel = c.use_gravity = 0;
  c.wheel_dist = c.wheel_weight = c.gravity_weight = 0.0;
  for (int i = 0; i < 3; ++i) c.wheel_prev[i] = c.g_ref[i] = c.g_cur[i] = 0.0;
}

LSB_HD void lm_add_sensor_terms(const LmSensorTerms& c, const LmPoint& pt, double* E)
{
  if (c.use_wheel) {
    const double d0 = pt.t.x - c.wheel_prev[0], d1 = pt.t.y - c.wheel_prev[1], d2 = pt.t.z - c.wheel_prev[2];
    const double sq = d0 * d0 + d1 * d1 + d2 * d2;
    double nrm = 0.0, j[3] = {0.0, 0.0, 0.0};
    if (!(sq < 1e-6)) {
      nrm = ::sqrt(sq);
      j[0] = d0 / nrm; j[1] = d1 / nrm; j[2] = d2 / nrm;
    }
    const double r = nrm - c.wheel_dist, w = c.wheel_weight;
    for (int a = 0; a < 3; ++a) {
      for (int b = a; b < 3; ++b) E[lm_hidx(a, b)] += w * (j[a] * j[b]);
      E[21 + a] += w * (j[a] * r);
    }
    E[27] += 0.5 * (w * (r * r));
  }
  if (c.use_gravity) {
    const Vec3d Y = pt.R * Vec3d(c.g_cur[0], c.g_cur[1], c.g_cur[2]);
    const double r0 = Y.x - c.g_ref[0], r1 = Y.y - c.g_ref[1], r2 = Y.z - c.g_ref[2];
    const double ax = pt.R.m[0], ay = pt.R.m[3], az = pt.R.m[6];
    // columns d(R g)/d(rx, ry, rz): the cross products of lm_eval_record
    const double J[3][3] = {{ay * Y.z - az * Y.y, az * Y.x - ax * Y.z, ax * Y.y - ay * Y.x},
                            {pt.cz * Y.z, pt.sz * Y.z, -(pt.sz * Y.y) - pt.cz * Y.x},
                            {-Y.y, Y.x, 0.0}};
    const double w = c.gravity_weight;
    for (int a = 0; a < 3; ++a) {
      for (int b = a; b < 3; ++b) E[lm_hidx(3 + a, 3 + b)] += w * (J[a][0] * J[b][0] + J[a][1] * J[b][1] + J[a][2] * J[b][2]);
      E[24 + a] += w * (J[a][0] * r0 + J[a][1] * r1 + J[a][2] * r2);
    }
    E[27] += 0.5 * (w * (r0 * r0 + r1 * r1 + r2 * r2));
  }
}

LSB_HD void lm_first_eval(LmState& s, const double* E) { if (lm_first_update(s, E)) lm_propose(s); }
LSB_HD void lm_next_eval(LmState& s, const double* E) { if (lm_next_update(s, E)) lm_propose(s); }

}  // namespace lsb
