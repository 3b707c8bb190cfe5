// TEST INFRASTRUCTURE -- CPU restatement of CeresScanMatcher2D,
// never linked into the product.
//
// Restates
//   cartographer/mapping/internal/2d/scan_matching/ceres_scan_matcher_2d.cc:64-110   (Match)
//   .../occupied_space_cost_function_2d.cc:30-96   (residuals, GridArrayAdapter, kPadding)
//   .../translation_delta_cost_functor_2d.h:30-60, rotation_delta_cost_functor_2d.h:30-54
// and, because the arithmetic lives in a dependency that is NOT under /root/reference, the
// published algorithm of Ceres Solver (bazel/repositories.bzl:137-141 pins commit
// 58c5edae2f7c4d2533fe8a975c1f5f0b892dfd3e, the 1.13/1.14 line):
//   include/ceres/cubic_interpolation.h  CubicHermiteSpline, BiCubicInterpolator::Evaluate
//   internal/ceres/trust_region_minimizer.cc      the iteration, its three tolerances
//   internal/ceres/levenberg_marquardt_strategy.cc  D = sqrt(clamp(diag(J'J)) / radius),
//                                                   radius update rules
//   internal/ceres/trust_region_step_evaluator.cc   (non)monotonic step quality
//   internal/ceres/dense_qr_solver.cc               least squares of [J; D] by Householder QR
// with Ceres' defaults where Cartographer does not set them (common/ceres_solver_options.cc:
// only use_nonmonotonic_steps, max_num_iterations, num_threads are set; Jacobi scaling on,
// initial radius 1e4, max radius 1e16, min radius 1e-32, min_relative_decrease 1e-3,
// function / gradient / parameter tolerance 1e-6 / 1e-10 / 1e-8, min/max LM diagonal 1e-6 / 1e32,
// max_consecutive_nonmonotonic_steps 5, max_num_consecutive_invalid_steps 5).
//
// PARITY UNPINNED at bit level: Ceres is absent from this image, so this restatement cannot be
// run against it.  It is pinned at the tolerance of the reference's own tests
// (ceres_scan_matcher_2d_test.cc:30-98: four start poses, pose within 1e-2, final cost ~ 0).
#ifndef ORACLE_CSM_CERES_H_
#define ORACLE_CSM_CERES_H_

#include <climits>

#include "csm_oracle.h"

namespace oracle {

struct CeresScanMatcherOptions2D {
  double occupied_space_weight = 1.;
  double translation_weight = 10.;
  double rotation_weight = 40.;
  bool use_nonmonotonic_steps = false;
  int max_num_iterations = 20;
};

struct CeresSummary {
  double initial_cost = 0., final_cost = 0.;
  int iterations = 0;  // iterations run after the initial evaluation
  int num_successful_steps = 0, num_unsuccessful_steps = 0;
  // 0 convergence (gradient), 1 convergence (parameter), 2 convergence (function),
  // 3 no convergence (max iterations), 4 no convergence (min radius), 5 failure
  int termination = 3;
};

// include/ceres/cubic_interpolation.h: CubicHermiteSpline<1>
inline void CubicHermiteSpline(const double p0, const double p1, const double p2, const double p3,
                               const double x, double* f, double* dfdx) {
  const double a = 0.5 * (-p0 + 3.0 * p1 - 3.0 * p2 + p3);
  const double b = 0.5 * (2.0 * p0 - 5.0 * p1 + 4.0 * p2 - p3);
  const double c = 0.5 * (-p0 + p2);
  const double d = p1;
  if (f != nullptr) *f = d + x * (c + x * (b + x * a));
  if (dfdx != nullptr) *dfdx = c + x * (2.0 * b + 3.0 * a * x);
}

class CeresScanMatcher2D {
 public:
  explicit CeresScanMatcher2D(const CeresScanMatcherOptions2D& options) : options_(options) {
    ORACLE_CHECK(options.occupied_space_weight > 0.);
    ORACLE_CHECK(options.translation_weight > 0.);
    ORACLE_CHECK(options.rotation_weight > 0.);
  }

  static constexpr int kPadding = INT_MAX / 4;  // occupied_space_cost_function_2d.cc:67

  // GridArrayAdapter::GetValue (occupied_space_cost_function_2d.cc:74-82)
  static double GetValue(const ProbabilityGrid& grid, const int row, const int column) {
    const int num_rows = grid.limits().cell_limits.num_y_cells + 2 * kPadding;
    const int num_cols = grid.limits().cell_limits.num_x_cells + 2 * kPadding;
    if (row < kPadding || column < kPadding || row >= num_rows - kPadding ||
        column >= num_cols - kPadding) {
      return static_cast<double>(kMaxCorrespondenceCost);
    }
    Array2i c;
    c.x = column - kPadding;
    c.y = row - kPadding;
    return static_cast<double>(grid.GetCorrespondenceCost(c));
  }

  // BiCubicInterpolator<GridArrayAdapter>::Evaluate(r, c, f, dfdr, dfdc)
  static void Interpolate(const ProbabilityGrid& grid, const double r, const double c, double* f,
                          double* dfdr, double* dfdc) {
    const int row = static_cast<int>(std::floor(r));
    const int col = static_cast<int>(std::floor(c));
    double fr[4], dfr[4];
    for (int k = 0; k < 4; ++k) {
      const double p0 = GetValue(grid, row - 1 + k, col - 1), p1 = GetValue(grid, row - 1 + k, col);
      const double p2 = GetValue(grid, row - 1 + k, col + 1), p3 = GetValue(grid, row - 1 + k, col + 2);
      CubicHermiteSpline(p0, p1, p2, p3, c - col, &fr[k], &dfr[k]);
    }
    CubicHermiteSpline(fr[0], fr[1], fr[2], fr[3], r - row, f, dfdr);
    CubicHermiteSpline(dfr[0], dfr[1], dfr[2], dfr[3], r - row, dfdc, nullptr);
  }

  // All N + 3 residuals and, optionally, the (N + 3) x 3 row-major Jacobian at `pose`:
  // what the three AutoDiffCostFunctions evaluate (Jet derivatives of the interpolator are
  // its analytic partials, cubic_interpolation.h BiCubicInterpolator::Evaluate<JetT>).
  bool Evaluate(const double pose[3], const double target[2], const double initial_angle,
                const PointCloud& cloud, const ProbabilityGrid& grid, std::vector<double>* residuals,
                std::vector<double>* jacobian) const {
    const size_t n = cloud.size();
    residuals->assign(n + 3, 0.);
    if (jacobian) jacobian->assign(3 * (n + 3), 0.);
    const double scaling = options_.occupied_space_weight / std::sqrt(static_cast<double>(n));
    const double cs = std::cos(pose[2]), sn = std::sin(pose[2]);
    const MapLimits& limits = grid.limits();
    for (size_t i = 0; i < n; ++i) {
      const double px = static_cast<double>(cloud[i].x), py = static_cast<double>(cloud[i].y);
      // transform * (px, py, 1): Eigen evaluates row * vector left to right.
      const double wx = cs * px + (-sn) * py + pose[0] * 1.;
      const double wy = sn * px + cs * py + pose[1] * 1.;
      const double r = (limits.max_x - wx) / limits.resolution - 0.5 + static_cast<double>(kPadding);
      const double c = (limits.max_y - wy) / limits.resolution - 0.5 + static_cast<double>(kPadding);
      double f, dfdr, dfdc;
      Interpolate(grid, r, c, &f, &dfdr, &dfdc);
      (*residuals)[i] = scaling * f;
      if (jacobian) {
        const double dr = -1. / limits.resolution;  // d r / d wx  ==  d c / d wy
        const double dwx_dth = -sn * px - cs * py, dwy_dth = cs * px - sn * py;
        (*jacobian)[3 * i + 0] = scaling * (dfdr * dr);
        (*jacobian)[3 * i + 1] = scaling * (dfdc * dr);
        (*jacobian)[3 * i + 2] = scaling * (dfdr * dr * dwx_dth + dfdc * dr * dwy_dth);
      }
    }
    (*residuals)[n + 0] = options_.translation_weight * (pose[0] - target[0]);
    (*residuals)[n + 1] = options_.translation_weight * (pose[1] - target[1]);
    (*residuals)[n + 2] = options_.rotation_weight * (pose[2] - initial_angle);
    if (jacobian) {
      (*jacobian)[3 * (n + 0) + 0] = options_.translation_weight;
      (*jacobian)[3 * (n + 1) + 1] = options_.translation_weight;
      (*jacobian)[3 * (n + 2) + 2] = options_.rotation_weight;
    }
    return true;
  }

  // --- TSDF2D grids: tsdf_match_cost_function_2d.cc:28-75 on interpolated_tsdf_2d.h:35-131 ----
  // Grid2D::GetCorrespondenceCost of a TSDF2D (grid_2d.h:53-57): the TSD itself, read through
  // the table GetConversionTable(max, min, max) -- unknown cells and cells outside the grid
  // read +truncation_distance.
  static float TsdfCost(const TSDF2D& g, const Array2i& c) {
    if (!g.limits().Contains(c)) return g.GetMaxCorrespondenceCost();
    const uint16_t v = g.tsd_cells()[g.limits().cell_limits.num_x_cells * c.y + c.x];
    return (v & 0x7FFF) == 0 ? g.GetMaxCorrespondenceCost() : g.value_to_tsd()[v];
  }
  static float TsdfWeight(const TSDF2D& g, const Array2i& c) { return g.GetTSDAndWeight(c).second; }

  struct Bilinear {
    float x1, y1, x2, y2;
    Array2i index1;
  };
  // ComputeInterpolationDataPoints / CenterOfLowerPixel (interpolated_tsdf_2d.h:85-93,110-124)
  static Bilinear LowerPixel(const TSDF2D& g, const double x, const double y) {
    const MapLimits& l = g.limits();
    const Array2i c = l.GetCellIndex(static_cast<float>(x), static_cast<float>(y));
    float cx = static_cast<float>(l.max_x - l.resolution * (c.y + 0.5));  // GetCellCenter, map_limits.h:79-82
    float cy = static_cast<float>(l.max_y - l.resolution * (c.x + 0.5));
    if (cx > x) cx = static_cast<float>(cx - l.resolution);
    if (cy > y) cy = static_cast<float>(cy - l.resolution);
    Bilinear b;
    b.x1 = cx;
    b.y1 = cy;
    b.x2 = cx + static_cast<float>(l.resolution);
    b.y2 = cy + static_cast<float>(l.resolution);
    b.index1 = l.GetCellIndex(b.x1, b.y1);
    return b;
  }
  // InterpolateBilinear (:95-104) and its partials (what the Jets carry).
  static void InterpolateBilinear(const Bilinear& b, const double x, const double y, const float q11,
                                  const float q12, const float q21, const float q22, double* f,
                                  double* dfdx, double* dfdy) {
    const double wx = static_cast<double>(b.x2 - b.x1), wy = static_cast<double>(b.y2 - b.y1);
    const double nx = (x - static_cast<double>(b.x1)) / wx;
    const double ny = (y - static_cast<double>(b.y1)) / wy;
    const double d1 = static_cast<double>(q12 - q11), d2 = static_cast<double>(q22 - q21);
    const double q1 = d1 * ny + static_cast<double>(q11);
    const double q2 = d2 * ny + static_cast<double>(q21);
    *f = (q2 - q1) * nx + q1;
    *dfdx = (q2 - q1) / wx;
    *dfdy = ((d2 - d1) * nx + d1) / wy;
  }
  // GetCorrespondenceCost(x, y) and GetWeight(x, y) with partials (:41-83).
  static void TsdfSample(const TSDF2D& g, const double x, const double y, double* cost, double* dcdx,
                         double* dcdy, double* weight, double* dwdx, double* dwdy) {
    const Bilinear b = LowerPixel(g, x, y);
    Array2i i12 = b.index1, i21 = b.index1, i22 = b.index1;
    i12.x -= 1;
    i21.y -= 1;
    i22.x -= 1;
    i22.y -= 1;
    const float w11 = TsdfWeight(g, b.index1), w12 = TsdfWeight(g, i12), w21 = TsdfWeight(g, i21),
                w22 = TsdfWeight(g, i22);
    if (w11 == 0.0 || w12 == 0.0 || w21 == 0.0 || w22 == 0.0) {
      *cost = static_cast<double>(g.GetMaxCorrespondenceCost());
      *dcdx = *dcdy = 0.;
    } else {
      InterpolateBilinear(b, x, y, TsdfCost(g, b.index1), TsdfCost(g, i12), TsdfCost(g, i21),
                          TsdfCost(g, i22), cost, dcdx, dcdy);
    }
    InterpolateBilinear(b, x, y, w11, w12, w21, w22, weight, dwdx, dwdy);
  }

  bool Evaluate(const double pose[3], const double target[2], const double initial_angle,
                const PointCloud& cloud, const TSDF2D& grid, std::vector<double>* residuals,
                std::vector<double>* jacobian) const {
    const size_t n = cloud.size();
    residuals->assign(n + 3, 0.);
    if (jacobian) jacobian->assign(3 * (n + 3), 0.);
    const double scaling = options_.occupied_space_weight / std::sqrt(static_cast<double>(n));
    const double cs = std::cos(pose[2]), sn = std::sin(pose[2]);
    std::vector<double> c(n), w(n), dc(3 * n), dw(3 * n);
    double summed = 0., dsummed[3] = {0., 0., 0.};
    for (size_t i = 0; i < n; ++i) {
      const double px = static_cast<double>(cloud[i].x), py = static_cast<double>(cloud[i].y);
      const double wx = cs * px + (-sn) * py + pose[0] * 1.;
      const double wy = sn * px + cs * py + pose[1] * 1.;
      double dcdx, dcdy, dwdx, dwdy;
      TsdfSample(grid, wx, wy, &c[i], &dcdx, &dcdy, &w[i], &dwdx, &dwdy);
      const double dwx_dth = -sn * px - cs * py, dwy_dth = cs * px - sn * py;
      dc[3 * i] = dcdx, dc[3 * i + 1] = dcdy, dc[3 * i + 2] = dcdx * dwx_dth + dcdy * dwy_dth;
      dw[3 * i] = dwdx, dw[3 * i + 1] = dwdy, dw[3 * i + 2] = dwdx * dwx_dth + dwdy * dwy_dth;
      summed += w[i];
      for (int k = 0; k < 3; ++k) dsummed[k] += dw[3 * i + k];
    }
    if (summed == 0.) return false;  // tsdf_match_cost_function_2d.cc:62
    const double ns = static_cast<double>(n) * scaling;
    for (size_t i = 0; i < n; ++i) {
      const double num = ns * c[i] * w[i];
      (*residuals)[i] = num / summed;
      if (jacobian)
        for (int k = 0; k < 3; ++k)
          (*jacobian)[3 * i + k] = ns * (dc[3 * i + k] * w[i] + c[i] * dw[3 * i + k]) / summed -
                                   num * dsummed[k] / (summed * summed);
    }
    (*residuals)[n + 0] = options_.translation_weight * (pose[0] - target[0]);
    (*residuals)[n + 1] = options_.translation_weight * (pose[1] - target[1]);
    (*residuals)[n + 2] = options_.rotation_weight * (pose[2] - initial_angle);
    if (jacobian) {
      (*jacobian)[3 * (n + 0) + 0] = options_.translation_weight;
      (*jacobian)[3 * (n + 1) + 1] = options_.translation_weight;
      (*jacobian)[3 * (n + 2) + 2] = options_.rotation_weight;
    }
    return true;
  }

  // ceres_scan_matcher_2d.cc:64-110 + ceres::Solve (trust region, Levenberg-Marquardt, DENSE_QR).
  template <typename Grid>
  void Match(const double target_x, const double target_y, const Rigid2d& initial,
             const PointCloud& cloud, const Grid& grid, Rigid2d* pose_estimate,
             CeresSummary* summary) const {
    ORACLE_CHECK(!cloud.empty());
    const double target[2] = {target_x, target_y};
    const double angle0 = initial.angle;
    const size_t m = cloud.size() + 3;
    double x[3] = {initial.tx, initial.ty, initial.angle};
    double x_min[3] = {x[0], x[1], x[2]};
    std::vector<double> res, jac, cand_res;
    // --- Init: evaluate, Jacobi scaling from the first Jacobian ------------------------------
    if (!Evaluate(x, target, angle0, cloud, grid, &res, &jac)) {
      // TrustRegionMinimizer::Init: "Residual and Jacobian evaluation failed" -> FAILURE, x untouched
      CeresSummary failed;
      failed.termination = 5;
      *pose_estimate = initial;
      if (summary) *summary = failed;
      return;
    }
    double x_cost = Cost(res);
    double gradient[3];
    Gradient(res, jac, gradient);
    double scale[3];
    for (int k = 0; k < 3; ++k) {
      double s = 0.;
      for (size_t i = 0; i < m; ++i) s += jac[3 * i + k] * jac[3 * i + k];
      scale[k] = 1. / (1. + std::sqrt(s));
    }
    ScaleColumns(&jac, scale);
    double x_norm = Norm3(x);
    CeresSummary sum;
    sum.initial_cost = x_cost;
    double minimum_cost = x_cost;
    // LM strategy state
    double radius = 1e4, decrease_factor = 2.;
    bool reuse_diagonal = false;
    double diagonal[3] = {0., 0., 0.};
    // step evaluator state
    const int max_nonmonotonic = options_.use_nonmonotonic_steps ? 5 : 0;
    double ev_minimum = x_cost, ev_current = x_cost, ev_reference = x_cost, ev_candidate = x_cost;
    double acc_reference = 0., acc_candidate = 0.;
    int num_nonmonotonic = 0, num_invalid = 0;
    int iteration = 0;
    sum.termination = 3;
    bool last_successful = false;
    for (;;) {
      // FinalizeIterationAndCheckIfMinimizerCanContinue
      if (last_successful) {
        ++sum.num_successful_steps;
        if (x_cost < minimum_cost) {
          minimum_cost = x_cost;
          x_min[0] = x[0], x_min[1] = x[1], x_min[2] = x[2];
        }
      } else if (iteration > 0) {
        ++sum.num_unsuccessful_steps;
      }
      if (iteration >= options_.max_num_iterations) { sum.termination = 3; break; }
      if (MaxAbs3(gradient) <= 1e-10) { sum.termination = 0; break; }
      if (radius <= 1e-32) { sum.termination = 4; break; }
      ++iteration;
      last_successful = false;
      // --- ComputeTrustRegionStep (LevenbergMarquardtStrategy::ComputeStep) -----------------
      if (!reuse_diagonal) {
        for (int k = 0; k < 3; ++k) {
          double s = 0.;
          for (size_t i = 0; i < m; ++i) s += jac[3 * i + k] * jac[3 * i + k];
          diagonal[k] = std::min(std::max(s, 1e-6), 1e32);
        }
      }
      double lm_diagonal[3];
      for (int k = 0; k < 3; ++k) lm_diagonal[k] = std::sqrt(diagonal[k] / radius);
      double step[3];
      const bool solved = SolveAugmented(jac, res, lm_diagonal, step);  // min |J s - r|^2 + |D s|^2
      for (int k = 0; k < 3; ++k) step[k] = -step[k];
      reuse_diagonal = true;
      double model_cost_change = 0.;
      if (solved) {
        for (size_t i = 0; i < m; ++i) {
          const double mr = jac[3 * i] * step[0] + jac[3 * i + 1] * step[1] + jac[3 * i + 2] * step[2];
          model_cost_change -= mr * (res[i] + mr / 2.);
        }
      }
      if (!solved || !(model_cost_change > 0.)) {  // HandleInvalidStep
        if (++num_invalid >= 5) { sum.termination = 5; break; }
        radius = radius / decrease_factor;  // StepIsInvalid == StepRejected
        decrease_factor *= 2.;
        reuse_diagonal = true;
        continue;
      }
      num_invalid = 0;
      const double delta[3] = {step[0] * scale[0], step[1] * scale[1], step[2] * scale[2]};
      const double cand[3] = {x[0] + delta[0], x[1] + delta[1], x[2] + delta[2]};
      // A failed evaluation makes the candidate infinitely expensive: the step is rejected.
      const bool cand_ok = Evaluate(cand, target, angle0, cloud, grid, &cand_res, nullptr);
      const double cand_cost = cand_ok ? Cost(cand_res) : std::numeric_limits<double>::max();
      // ParameterToleranceReached / FunctionToleranceReached (before the step is judged)
      const double step_norm = Norm3(delta);
      if (step_norm <= 1e-8 * (x_norm + 1e-8)) { sum.termination = 1; break; }
      if (std::abs(x_cost - cand_cost) <= 1e-6 * x_cost) { sum.termination = 2; break; }
      // TrustRegionStepEvaluator::StepQuality
      const double relative_decrease = (ev_current - cand_cost) / model_cost_change;
      const double historical = (ev_reference - cand_cost) / (acc_reference + model_cost_change);
      const double quality =
          cand_ok ? std::max(relative_decrease, historical) : std::numeric_limits<double>::lowest();
      if (quality > 1e-3) {  // HandleSuccessfulStep
        x[0] = cand[0], x[1] = cand[1], x[2] = cand[2];
        x_norm = Norm3(x);
        Evaluate(x, target, angle0, cloud, grid, &res, &jac);
        x_cost = Cost(res);
        Gradient(res, jac, gradient);
        ScaleColumns(&jac, scale);
        last_successful = true;
        // LevenbergMarquardtStrategy::StepAccepted
        radius = radius / std::max(1. / 3., 1. - std::pow(2. * quality - 1., 3));
        radius = std::min(1e16, radius);
        decrease_factor = 2.;
        reuse_diagonal = false;
        // TrustRegionStepEvaluator::StepAccepted
        ev_current = cand_cost;
        acc_candidate += model_cost_change;
        acc_reference += model_cost_change;
        if (ev_current < ev_minimum) {
          ev_minimum = ev_current;
          num_nonmonotonic = 0;
          ev_candidate = ev_current;
          acc_candidate = 0.;
        } else {
          ++num_nonmonotonic;
          if (ev_current > ev_candidate) {
            ev_candidate = ev_current;
            acc_candidate = 0.;
          }
        }
        if (num_nonmonotonic == max_nonmonotonic) {
          ev_reference = ev_candidate;
          acc_reference = acc_candidate;
        }
      } else {  // HandleUnsuccessfulStep -> StepRejected
        radius = radius / decrease_factor;
        decrease_factor *= 2.;
        reuse_diagonal = true;
      }
    }
    sum.iterations = iteration;
    sum.final_cost = minimum_cost;
    pose_estimate->tx = x_min[0];
    pose_estimate->ty = x_min[1];
    pose_estimate->angle = x_min[2];
    if (summary) *summary = sum;
  }

 private:
  static double Cost(const std::vector<double>& r) {
    double s = 0.;
    for (double v : r) s += v * v;
    return 0.5 * s;
  }
  static void Gradient(const std::vector<double>& r, const std::vector<double>& j, double g[3]) {
    g[0] = g[1] = g[2] = 0.;
    for (size_t i = 0; i < r.size(); ++i)
      for (int k = 0; k < 3; ++k) g[k] += j[3 * i + k] * r[i];
  }
  static void ScaleColumns(std::vector<double>* j, const double s[3]) {
    for (size_t i = 0; i < j->size(); i += 3)
      for (int k = 0; k < 3; ++k) (*j)[i + k] *= s[k];
  }
  static double Norm3(const double v[3]) { return std::sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]); }
  static double MaxAbs3(const double v[3]) {
    return std::max(std::abs(v[0]), std::max(std::abs(v[1]), std::abs(v[2])));
  }
  // DenseQRSolver: least squares of [J; diag(D)] s = [r; 0] by Householder QR.
  static bool SolveAugmented(const std::vector<double>& j, const std::vector<double>& r,
                             const double d[3], double s[3]) {
    const size_t m = r.size(), rows = m + 3;
    std::vector<double> a(rows * 3, 0.), b(rows, 0.);
    for (size_t i = 0; i < m; ++i) {
      for (int k = 0; k < 3; ++k) a[3 * i + k] = j[3 * i + k];
      b[i] = r[i];
    }
    for (int k = 0; k < 3; ++k) a[3 * (m + k) + k] = d[k];
    for (int k = 0; k < 3; ++k) {
      double norm = 0.;
      for (size_t i = k; i < rows; ++i) norm += a[3 * i + k] * a[3 * i + k];
      norm = std::sqrt(norm);
      if (!(norm > 0.)) return false;
      const double alpha = a[3 * k + k] > 0. ? -norm : norm;
      std::vector<double> v(rows, 0.);
      for (size_t i = k; i < rows; ++i) v[i] = a[3 * i + k];
      v[k] -= alpha;
      double vv = 0.;
      for (size_t i = k; i < rows; ++i) vv += v[i] * v[i];
      if (!(vv > 0.)) continue;
      for (int c = k; c < 3; ++c) {
        double dot = 0.;
        for (size_t i = k; i < rows; ++i) dot += v[i] * a[3 * i + c];
        const double f = 2. * dot / vv;
        for (size_t i = k; i < rows; ++i) a[3 * i + c] -= f * v[i];
      }
      double dot = 0.;
      for (size_t i = k; i < rows; ++i) dot += v[i] * b[i];
      const double f = 2. * dot / vv;
      for (size_t i = k; i < rows; ++i) b[i] -= f * v[i];
    }
    for (int k = 2; k >= 0; --k) {
      double t = b[k];
      for (int c = k + 1; c < 3; ++c) t -= a[3 * k + c] * s[c];
      if (a[3 * k + k] == 0.) return false;
      s[k] = t / a[3 * k + k];
    }
    return std::isfinite(s[0]) && std::isfinite(s[1]) && std::isfinite(s[2]);
  }

  CeresScanMatcherOptions2D options_;
};

}  // namespace oracle

#endif  // ORACLE_CSM_CERES_H_
