// CPU ORACLE (test infrastructure) — LocalOptimizer restatement
// (reference slam_lib/src/LocalOptimizer.cxx:44-48, 74-102, 105-140).
// The LM state machine and per-record math come from common/lm_core.h (a
// restatement of Ceres, shared with the GPU on purpose); this file provides the
// sequential CPU evaluation in the canonical reduction order.
#pragma once
#include <array>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "orc_matcher.h"
#include "../lidarslam_b200/csrc/common/lm_core.h"
#include "../lidarslam_b200/csrc/common/lm_covariance.h"

namespace orc {

struct Optimizer {
  bool TwoDMode = false;
  unsigned LMMaxIter = 15;
  int NbThreads = 1;
  double Sat = 1.0;
  double PoseArray[6] = {0, 0, 0, 0, 0, 0};
  std::vector<lsb::LmRecord> Records;   // index-aligned with the matcher output, w < 0 = no residual
  lsb::LmState State;
  int NumEvaluations = 0;
  // one row per candidate evaluation: {iteration, cost at x, cost at the candidate, radius used for the step, accepted (1/0)}
  // (tests compare the accept / reject sequence with an independent numpy restatement of the trust-region loop)
  std::vector<std::array<double, 5>> Trace;
  lsb::LmSensorTerms SensorTerms;        // wheel-odometer / IMU gravity blocks (Slam.cxx:1123-1131), none by default
  Optimizer() { lsb::lm_sensor_terms_clear(SensorTerms); }

  void SetPosePrior(const Iso3d& T) { lsb::iso_to_xyzrpy(T, PoseArray); }   // LocalOptimizer.cxx:44-48
  void AddResiduals(const MatchingResults& m)                                // Slam.cxx:1120-1121
  {
    for (size_t i = 0; i < m.Residuals.size(); ++i) {
      lsb::LmRecord r;
      const MatchRecord& s = m.Residuals[i];
      lsb::lm_record_set(r, s.A, s.P, s.X, (m.Rejections[i] == LSB_MATCH_SUCCESS) ? s.weight : -1.0);
      Records.push_back(r);
    }
  }

  // fold32 of lm_core.h: v[0] receives the pairwise tree sum of v[0..31]
  static void Fold32(double v[32][LSB_LM_NSUM])
  {
    for (int off = 16; off >= 1; off >>= 1)
      for (int l = 0; l < off; ++l)
        for (int q = 0; q < LSB_LM_NSUM; ++q) v[l][q] += v[l + off][q];
  }

  // 28 sums at x in the canonical order of lm_core.h (radix-32 tree: records -> groups ->
  // chunks -> blocks, block sums added in order)
  void Evaluate(const double* x, double* E)
  {
    ++NumEvaluations;
    lsb::LmPoint pt;
    lsb::lm_make_point(x, pt);
    const size_t n = Records.size();
    const size_t nChunks = std::max<size_t>((n + LSB_LM_CHUNK - 1) / LSB_LM_CHUNK, 1);
    const size_t nBlocks = (nChunks + 31) / 32;
    std::vector<double> part(nBlocks * 32 * LSB_LM_NSUM, 0.0);   // missing chunks stay +0.0
#pragma omp parallel for num_threads(NbThreads) schedule(static)
    for (long c = 0; c < (long)nChunks; ++c) {
      double group[32][LSB_LM_NSUM];
      for (int g = 0; g < 32; ++g) {
        double lane[32][LSB_LM_NSUM];
        for (int l = 0; l < 32; ++l) {
          const size_t i = (size_t)c * LSB_LM_CHUNK + g * 32 + l;
          if (i < n) lsb::lm_eval_record(Records[i], pt, Sat, lane[l]);
          else for (int q = 0; q < LSB_LM_NSUM; ++q) lane[l][q] = 0.0;
        }
        Fold32(lane);
        for (int q = 0; q < LSB_LM_NSUM; ++q) group[g][q] = lane[0][q];
      }
      Fold32(group);
      for (int q = 0; q < LSB_LM_NSUM; ++q) part[c * LSB_LM_NSUM + q] = group[0][q];
    }
    for (size_t b = 0; b < nBlocks; ++b) {
      double chunk[32][LSB_LM_NSUM];
      for (int l = 0; l < 32; ++l)
        for (int q = 0; q < LSB_LM_NSUM; ++q) chunk[l][q] = part[(b * 32 + l) * LSB_LM_NSUM + q];
      Fold32(chunk);
      for (int q = 0; q < LSB_LM_NSUM; ++q) E[q] = (b == 0) ? chunk[0][q] : E[q] + chunk[0][q];
    }
    if (SensorTerms.use_wheel || SensorTerms.use_gravity) lsb::lm_add_sensor_terms(SensorTerms, pt, E);
  }

  // LocalOptimizer::Solve (:74-102). Returns Ceres' num_successful_steps
  // (iteration 0 counts as successful).
  int Solve()
  {
    double E[LSB_LM_NSUM];
    lsb::lm_begin(State, PoseArray, (int)LMMaxIter, TwoDMode ? 1 : 0);
    Evaluate(State.x, E);
    lsb::lm_first_eval(State, E);
    const bool trace = std::getenv("ORC_LM_TRACE") != nullptr;   // debugging aid: one line per evaluation on stderr
    Trace.clear();
    while (!State.done) {
      Evaluate(State.xc, E);
      Trace.push_back({(double)State.iter, State.cost, E[27], State.radius, 0.0});
      if (trace) std::fprintf(stderr, "[orc lm] iter %d cost %.12g -> %.12g radius %.3g step %.3g\n", State.iter, State.cost, E[27], State.radius, State.step_norm);
      const int accepted_before = State.num_successful;
      lsb::lm_next_eval(State, E);
      Trace.back()[4] = (State.num_successful > accepted_before) ? 1.0 : 0.0;
    }
    if (trace) std::fprintf(stderr, "[orc lm] done: iter %d accepted %d term %d n %zu\n", State.iter, State.num_successful, State.term, Records.size());
    for (int i = 0; i < 6; ++i) PoseArray[i] = State.x[i];
    return State.num_successful + 1;
  }

  Iso3d GetOptimizedPose() const { return lsb::xyzrpy_to_iso(PoseArray); }   // :105-109

  lsb_registration_error EstimateRegistrationError() const                   // :112-140
  {
    lsb_registration_error e;
    lsb::registration_error_from_normal_matrix(State.H, TwoDMode ? 1 : 0, &e);
    return e;
  }
};

}  // namespace orc
