// CeresScanMatcher2D::Match as a batched Levenberg-Marquardt solve on the device
// (the step right after the branch-and-bound search inside
// ConstraintBuilder2D::ComputeConstraint, constraint_builder_2d.cc:245-249).
//
// Replaces /root/reference/cartographer/mapping/internal/2d/scan_matching/
//   ceres_scan_matcher_2d.cc:64-110          Match: one occupied-space block (N residuals),
//                                            two translation residuals, one rotation residual
//   occupied_space_cost_function_2d.cc:30-96 residual_i = w / sqrt(N) * bicubic(cost grid at T p_i),
//                                            kMaxCorrespondenceCost outside the grid, kPadding
// and the parts of Ceres Solver the call runs through (not in /root/reference; pinned by
// bazel/repositories.bzl:137-141): BiCubicInterpolator / CubicHermiteSpline
// (include/ceres/cubic_interpolation.h) and the trust-region Levenberg-Marquardt loop with
// Jacobi scaling (trust_region_minimizer.cc, levenberg_marquardt_strategy.cc,
// trust_region_step_evaluator.cc) with Ceres' defaults.
//
// One warp per problem.  The problem has 3 parameters, so each evaluation reduces the N
// residual rows to J'J (6 numbers), J'r (3) and |r|^2 over the warp in fp64 (lane = point,
// butterfly reduction, every lane ends with the same sums) and the 3 x 3 damped normal
// equations are solved by every lane redundantly -- no shared memory, no divergence.
// Ceres solves the same least-squares problem by QR of [J; D]; the normal equations give
// the same step up to rounding (the columns are Jacobi-scaled, so the system is well
// conditioned).  Parity with the CPU restatement is therefore by tolerance (1e-6 on the
// pose), not bitwise; see DESIGN.md.
#include <map>
#include <vector>

#include "csm_internal.cuh"

namespace {

constexpr int kPadding = 0x7FFFFFFF / 4;  // INT_MAX / 4, occupied_space_cost_function_2d.cc:67
constexpr unsigned kFullMask = 0xFFFFFFFFu;
constexpr int kCeresWarps = 4;

struct CeresJobDev {
  const uint16_t* cells;
  const float* cost_table;
  const uint16_t* weight_cells;  // TSDF only (nullptr: probability grid)
  const float* weight_table;
  float max_cc;                  // TSDF: truncation distance
  long long pts_off;
  int num_x, num_y, n;
  double max_x, max_y, res;
  double init[3], target[2];
  double w_occ, w_t, w_r;
  int max_iterations, nonmonotonic;
};

struct Sums {
  double rr;    // sum of squared residuals
  double g[3];  // J' r
  double h[6];  // J' J: 00 01 02 11 12 22
};

__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(kFullMask, v, o);
  return v;
}

__device__ __forceinline__ void hermite(double p0, double p1, double p2, double p3, double x, double* f,
                                        double* dfdx) {
  const double a = 0.5 * (-p0 + 3.0 * p1 - 3.0 * p2 + p3);
  const double b = 0.5 * (2.0 * p0 - 5.0 * p1 + 4.0 * p2 - p3);
  const double c = 0.5 * (-p0 + p2);
  const double d = p1;
  *f = d + x * (c + x * (b + x * a));
  if (dfdx != nullptr) *dfdx = c + x * (2.0 * b + 3.0 * a * x);
}

// GridArrayAdapter::GetValue: correspondence cost of cell (x = column - pad, y = row - pad),
// kMaxCorrespondenceCost outside the grid.
__device__ __forceinline__ double cost_at(const CeresJobDev& j, int row, int col) {
  const int y = row - kPadding, x = col - kPadding;
  if (static_cast<unsigned>(x) >= static_cast<unsigned>(j.num_x) ||
      static_cast<unsigned>(y) >= static_cast<unsigned>(j.num_y))
    return static_cast<double>(0.9f);
  return static_cast<double>(__ldg(j.cost_table + __ldg(j.cells + static_cast<size_t>(y) * j.num_x + x)));
}

// Residuals (and the Jacobian's normal-equation sums) at `pose`; every lane returns the totals.
__device__ void evaluate(const CeresJobDev& j, const float* __restrict__ pts, const double pose[3],
                         bool with_jacobian, int lane, Sums* out) {
  const double scaling = j.w_occ / sqrt(static_cast<double>(j.n));
  double sn, cs;
  sincos(pose[2], &sn, &cs);
  Sums s;
  s.rr = 0.;
  s.g[0] = s.g[1] = s.g[2] = 0.;
#pragma unroll
  for (int k = 0; k < 6; ++k) s.h[k] = 0.;
  const double dr = -1. / j.res;
  for (int p = lane; p < j.n; p += 32) {
    const double px = static_cast<double>(pts[3 * p]), py = static_cast<double>(pts[3 * p + 1]);
    const double wx = __dadd_rn(__dadd_rn(__dmul_rn(cs, px), __dmul_rn(-sn, py)), pose[0]);
    const double wy = __dadd_rn(__dadd_rn(__dmul_rn(sn, px), __dmul_rn(cs, py)), pose[1]);
    const double r = __dadd_rn(__dsub_rn(__ddiv_rn(__dsub_rn(j.max_x, wx), j.res), 0.5),
                               static_cast<double>(kPadding));
    const double c = __dadd_rn(__dsub_rn(__ddiv_rn(__dsub_rn(j.max_y, wy), j.res), 0.5),
                               static_cast<double>(kPadding));
    const int row = static_cast<int>(floor(r)), col = static_cast<int>(floor(c));
    double fr[4], dfr[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const double p0 = cost_at(j, row - 1 + k, col - 1), p1 = cost_at(j, row - 1 + k, col);
      const double p2 = cost_at(j, row - 1 + k, col + 1), p3 = cost_at(j, row - 1 + k, col + 2);
      hermite(p0, p1, p2, p3, c - col, &fr[k], &dfr[k]);
    }
    double f, dfdr, dfdc;
    hermite(fr[0], fr[1], fr[2], fr[3], r - row, &f, &dfdr);
    hermite(dfr[0], dfr[1], dfr[2], dfr[3], r - row, &dfdc, nullptr);
    const double res_i = scaling * f;
    s.rr += res_i * res_i;
    if (with_jacobian) {
      const double dwx = -sn * px - cs * py, dwy = cs * px - sn * py;
      const double j0 = scaling * (dfdr * dr), j1 = scaling * (dfdc * dr);
      const double j2 = scaling * (dfdr * dr * dwx + dfdc * dr * dwy);
      s.g[0] += j0 * res_i;
      s.g[1] += j1 * res_i;
      s.g[2] += j2 * res_i;
      s.h[0] += j0 * j0;
      s.h[1] += j0 * j1;
      s.h[2] += j0 * j2;
      s.h[3] += j1 * j1;
      s.h[4] += j1 * j2;
      s.h[5] += j2 * j2;
    }
  }
  s.rr = warp_sum_d(s.rr);
  if (with_jacobian) {
#pragma unroll
    for (int k = 0; k < 3; ++k) s.g[k] = warp_sum_d(s.g[k]);
#pragma unroll
    for (int k = 0; k < 6; ++k) s.h[k] = warp_sum_d(s.h[k]);
  }
  // translation_delta_cost_functor_2d.h:48-53, rotation_delta_cost_functor_2d.h:44-47
  const double r0 = j.w_t * (pose[0] - j.target[0]), r1 = j.w_t * (pose[1] - j.target[1]);
  const double r2 = j.w_r * (pose[2] - j.init[2]);
  s.rr += r0 * r0 + r1 * r1 + r2 * r2;
  if (with_jacobian) {
    s.g[0] += j.w_t * r0;
    s.g[1] += j.w_t * r1;
    s.g[2] += j.w_r * r2;
    s.h[0] += j.w_t * j.w_t;
    s.h[3] += j.w_t * j.w_t;
    s.h[5] += j.w_r * j.w_r;
  }
  *out = s;
}

// ---- TSDF2D: interpolated_tsdf_2d.h:35-131, tsdf_match_cost_function_2d.cc:28-75 ----------
// MapLimits::GetCellIndex (map_limits.h:69-76), lround.
__device__ __forceinline__ int2 cell_of(const CeresJobDev& j, float px, float py) {
  int2 c;
  c.x = static_cast<int>(llround(__dsub_rn(__ddiv_rn(__dsub_rn(j.max_y, static_cast<double>(py)), j.res), 0.5)));
  c.y = static_cast<int>(llround(__dsub_rn(__ddiv_rn(__dsub_rn(j.max_x, static_cast<double>(px)), j.res), 0.5)));
  return c;
}
__device__ __forceinline__ float tsdf_weight(const CeresJobDev& j, int x, int y) {
  if (static_cast<unsigned>(x) >= static_cast<unsigned>(j.num_x) ||
      static_cast<unsigned>(y) >= static_cast<unsigned>(j.num_y))
    return 0.f;
  return __ldg(j.weight_table + __ldg(j.weight_cells + static_cast<size_t>(y) * j.num_x + x));
}
__device__ __forceinline__ float tsdf_cost(const CeresJobDev& j, int x, int y) {
  if (static_cast<unsigned>(x) >= static_cast<unsigned>(j.num_x) ||
      static_cast<unsigned>(y) >= static_cast<unsigned>(j.num_y))
    return j.max_cc;
  const uint16_t v = __ldg(j.cells + static_cast<size_t>(y) * j.num_x + x);
  return (v & 0x7FFF) == 0 ? j.max_cc : __ldg(j.cost_table + v);  // unknown reads +truncation (grid_2d.h:53-57)
}
__device__ __forceinline__ void bilinear(float x1, float y1, float x2, float y2, double x, double y, float q11,
                                         float q12, float q21, float q22, double* f, double* dfdx,
                                         double* dfdy) {
  const double wx = static_cast<double>(__fsub_rn(x2, x1)), wy = static_cast<double>(__fsub_rn(y2, y1));
  const double nx = __ddiv_rn(__dsub_rn(x, static_cast<double>(x1)), wx);
  const double ny = __ddiv_rn(__dsub_rn(y, static_cast<double>(y1)), wy);
  const double d1 = static_cast<double>(__fsub_rn(q12, q11)), d2 = static_cast<double>(__fsub_rn(q22, q21));
  const double q1 = __dadd_rn(__dmul_rn(d1, ny), static_cast<double>(q11));
  const double q2 = __dadd_rn(__dmul_rn(d2, ny), static_cast<double>(q21));
  *f = __dadd_rn(__dmul_rn(__dsub_rn(q2, q1), nx), q1);
  *dfdx = (q2 - q1) / wx;
  *dfdy = ((d2 - d1) * nx + d1) / wy;
}
// Cost and weight at the world point (x, y) with their partials.
__device__ void tsdf_sample(const CeresJobDev& j, double x, double y, double* c, double* dcdx, double* dcdy,
                            double* w, double* dwdx, double* dwdy) {
  const int2 cell = cell_of(j, static_cast<float>(x), static_cast<float>(y));
  // GetCellCenter (map_limits.h:79-82) narrowed to float, then the next lower pixel centre
  float cx = static_cast<float>(__dsub_rn(j.max_x, __dmul_rn(j.res, cell.y + 0.5)));
  float cy = static_cast<float>(__dsub_rn(j.max_y, __dmul_rn(j.res, cell.x + 0.5)));
  if (static_cast<double>(cx) > x) cx = static_cast<float>(__dsub_rn(static_cast<double>(cx), j.res));
  if (static_cast<double>(cy) > y) cy = static_cast<float>(__dsub_rn(static_cast<double>(cy), j.res));
  const float x2 = __fadd_rn(cx, static_cast<float>(j.res)), y2 = __fadd_rn(cy, static_cast<float>(j.res));
  const int2 i1 = cell_of(j, cx, cy);
  const float w11 = tsdf_weight(j, i1.x, i1.y), w12 = tsdf_weight(j, i1.x - 1, i1.y);
  const float w21 = tsdf_weight(j, i1.x, i1.y - 1), w22 = tsdf_weight(j, i1.x - 1, i1.y - 1);
  if (w11 == 0.f || w12 == 0.f || w21 == 0.f || w22 == 0.f) {
    *c = static_cast<double>(j.max_cc);
    *dcdx = *dcdy = 0.;
  } else {
    bilinear(cx, cy, x2, y2, x, y, tsdf_cost(j, i1.x, i1.y), tsdf_cost(j, i1.x - 1, i1.y),
             tsdf_cost(j, i1.x, i1.y - 1), tsdf_cost(j, i1.x - 1, i1.y - 1), c, dcdx, dcdy);
  }
  bilinear(cx, cy, x2, y2, x, y, w11, w12, w21, w22, w, dwdx, dwdy);
}

// Two passes: the summed weight W and its gradient couple every residual,
// r_i = N s c_i w_i / W,  dr_i = N s ((dc_i w_i + c_i dw_i) / W - c_i w_i dW / W^2).
// Returns false (every lane alike) when W == 0.
__device__ bool evaluate_tsdf(const CeresJobDev& j, const float* __restrict__ pts, const double pose[3],
                              bool with_jacobian, int lane, Sums* out) {
  const double scaling = j.w_occ / sqrt(static_cast<double>(j.n));
  double sn, cs;
  sincos(pose[2], &sn, &cs);
  double W = 0., dW[3] = {0., 0., 0.};
  for (int p = lane; p < j.n; p += 32) {
    const double px = static_cast<double>(pts[3 * p]), py = static_cast<double>(pts[3 * p + 1]);
    const double wx = __dadd_rn(__dadd_rn(__dmul_rn(cs, px), __dmul_rn(-sn, py)), pose[0]);
    const double wy = __dadd_rn(__dadd_rn(__dmul_rn(sn, px), __dmul_rn(cs, py)), pose[1]);
    double c, dcdx, dcdy, w, dwdx, dwdy;
    tsdf_sample(j, wx, wy, &c, &dcdx, &dcdy, &w, &dwdx, &dwdy);
    W += w;
    dW[0] += dwdx;
    dW[1] += dwdy;
    dW[2] += dwdx * (-sn * px - cs * py) + dwdy * (cs * px - sn * py);
  }
  W = warp_sum_d(W);
#pragma unroll
  for (int k = 0; k < 3; ++k) dW[k] = warp_sum_d(dW[k]);
  if (W == 0.) return false;
  Sums s;
  s.rr = 0.;
  s.g[0] = s.g[1] = s.g[2] = 0.;
#pragma unroll
  for (int k = 0; k < 6; ++k) s.h[k] = 0.;
  const double ns = static_cast<double>(j.n) * scaling;
  for (int p = lane; p < j.n; p += 32) {
    const double px = static_cast<double>(pts[3 * p]), py = static_cast<double>(pts[3 * p + 1]);
    const double wx = __dadd_rn(__dadd_rn(__dmul_rn(cs, px), __dmul_rn(-sn, py)), pose[0]);
    const double wy = __dadd_rn(__dadd_rn(__dmul_rn(sn, px), __dmul_rn(cs, py)), pose[1]);
    double c, dcdx, dcdy, w, dwdx, dwdy;
    tsdf_sample(j, wx, wy, &c, &dcdx, &dcdy, &w, &dwdx, &dwdy);
    const double num = ns * c * w;
    const double res_i = num / W;
    s.rr += res_i * res_i;
    if (with_jacobian) {
      const double tx = -sn * px - cs * py, ty = cs * px - sn * py;
      const double dc[3] = {dcdx, dcdy, dcdx * tx + dcdy * ty};
      const double dw[3] = {dwdx, dwdy, dwdx * tx + dwdy * ty};
      double jr[3];
#pragma unroll
      for (int k = 0; k < 3; ++k) jr[k] = ns * (dc[k] * w + c * dw[k]) / W - num * dW[k] / (W * W);
      s.g[0] += jr[0] * res_i;
      s.g[1] += jr[1] * res_i;
      s.g[2] += jr[2] * res_i;
      s.h[0] += jr[0] * jr[0];
      s.h[1] += jr[0] * jr[1];
      s.h[2] += jr[0] * jr[2];
      s.h[3] += jr[1] * jr[1];
      s.h[4] += jr[1] * jr[2];
      s.h[5] += jr[2] * jr[2];
    }
  }
  s.rr = warp_sum_d(s.rr);
  if (with_jacobian) {
#pragma unroll
    for (int k = 0; k < 3; ++k) s.g[k] = warp_sum_d(s.g[k]);
#pragma unroll
    for (int k = 0; k < 6; ++k) s.h[k] = warp_sum_d(s.h[k]);
  }
  const double r0 = j.w_t * (pose[0] - j.target[0]), r1 = j.w_t * (pose[1] - j.target[1]);
  const double r2 = j.w_r * (pose[2] - j.init[2]);
  s.rr += r0 * r0 + r1 * r1 + r2 * r2;
  if (with_jacobian) {
    s.g[0] += j.w_t * r0;
    s.g[1] += j.w_t * r1;
    s.g[2] += j.w_r * r2;
    s.h[0] += j.w_t * j.w_t;
    s.h[3] += j.w_t * j.w_t;
    s.h[5] += j.w_r * j.w_r;
  }
  *out = s;
  return true;
}

__device__ __forceinline__ bool evaluate_any(const CeresJobDev& j, const float* __restrict__ pts,
                                             const double pose[3], int lane, Sums* out) {
  if (j.weight_cells != nullptr) return evaluate_tsdf(j, pts, pose, true, lane, out);
  evaluate(j, pts, pose, true, lane, out);
  return true;
}

// Solves the symmetric positive definite 3 x 3 system A s = b (A: 00 01 02 11 12 22) by Cholesky.
__device__ __forceinline__ bool solve_spd3(const double a[6], const double b[3], double s[3]) {
  const double l00 = sqrt(a[0]);
  if (!(l00 > 0.)) return false;
  const double l10 = a[1] / l00, l20 = a[2] / l00;
  const double d1 = a[3] - l10 * l10;
  if (!(d1 > 0.)) return false;
  const double l11 = sqrt(d1);
  const double l21 = (a[4] - l20 * l10) / l11;
  const double d2 = a[5] - l20 * l20 - l21 * l21;
  if (!(d2 > 0.)) return false;
  const double l22 = sqrt(d2);
  const double y0 = b[0] / l00;
  const double y1 = (b[1] - l10 * y0) / l11;
  const double y2 = (b[2] - l20 * y0 - l21 * y1) / l22;
  s[2] = y2 / l22;
  s[1] = (y1 - l21 * s[2]) / l11;
  s[0] = (y0 - l10 * s[1] - l20 * s[2]) / l00;
  return isfinite(s[0]) && isfinite(s[1]) && isfinite(s[2]);
}

__device__ __forceinline__ double norm3(const double v[3]) {
  return sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
}

__global__ void __launch_bounds__(kCeresWarps * 32)
k_ceres_lm(const CeresJobDev* __restrict__ jobs, const float* __restrict__ points, int num_jobs,
           csm_ceres_result2d* __restrict__ results) {
  const int job = blockIdx.x * kCeresWarps + (threadIdx.x >> 5);
  if (job >= num_jobs) return;
  const int lane = threadIdx.x & 31;
  const CeresJobDev& j = jobs[job];
  const float* pts = points + 3 * j.pts_off;
  double x[3] = {j.init[0], j.init[1], j.init[2]};
  double x_min[3] = {x[0], x[1], x[2]};
  Sums cur;
  if (!evaluate_any(j, pts, x, lane, &cur)) {  // "Residual and Jacobian evaluation failed": x untouched
    if (lane == 0) {
      csm_ceres_result2d r;
      r.pose[0] = x[0], r.pose[1] = x[1], r.pose[2] = x[2];
      r.initial_cost = r.final_cost = 0.;
      r.iterations = r.num_successful_steps = r.num_unsuccessful_steps = 0;
      r.termination = 5;
      results[job] = r;
    }
    return;
  }
  double x_cost = 0.5 * cur.rr;
  // Jacobi scaling from the first Jacobian: 1 / (1 + column norm)
  const double scale[3] = {1. / (1. + sqrt(cur.h[0])), 1. / (1. + sqrt(cur.h[3])),
                           1. / (1. + sqrt(cur.h[5]))};
  double x_norm = norm3(x);
  const double initial_cost = x_cost;
  double minimum_cost = x_cost;
  double radius = 1e4, decrease_factor = 2.;
  bool reuse_diagonal = false;
  double diagonal[3] = {0., 0., 0.};
  const int max_nonmonotonic = j.nonmonotonic ? 5 : 0;
  double ev_minimum = x_cost, ev_current = x_cost, ev_reference = x_cost, ev_candidate = x_cost;
  double acc_reference = 0., acc_candidate = 0.;
  int num_nonmonotonic = 0, num_invalid = 0, iteration = 0;
  int successful = 0, unsuccessful = 0, termination = 3;
  bool last_successful = false;
  for (;;) {
    if (last_successful) {
      ++successful;
      if (x_cost < minimum_cost) {
        minimum_cost = x_cost;
        x_min[0] = x[0], x_min[1] = x[1], x_min[2] = x[2];
      }
    } else if (iteration > 0) {
      ++unsuccessful;
    }
    if (iteration >= j.max_iterations) { termination = 3; break; }
    if (fmax(fabs(cur.g[0]), fmax(fabs(cur.g[1]), fabs(cur.g[2]))) <= 1e-10) { termination = 0; break; }
    if (radius <= 1e-32) { termination = 4; break; }
    ++iteration;
    last_successful = false;
    // scaled normal-equation terms: Hs = S H S, gs = S g
    const double hs[6] = {cur.h[0] * scale[0] * scale[0], cur.h[1] * scale[0] * scale[1],
                          cur.h[2] * scale[0] * scale[2], cur.h[3] * scale[1] * scale[1],
                          cur.h[4] * scale[1] * scale[2], cur.h[5] * scale[2] * scale[2]};
    const double gs[3] = {cur.g[0] * scale[0], cur.g[1] * scale[1], cur.g[2] * scale[2]};
    if (!reuse_diagonal) {
      diagonal[0] = fmin(fmax(hs[0], 1e-6), 1e32);
      diagonal[1] = fmin(fmax(hs[3], 1e-6), 1e32);
      diagonal[2] = fmin(fmax(hs[5], 1e-6), 1e32);
    }
    const double a[6] = {hs[0] + diagonal[0] / radius, hs[1], hs[2], hs[3] + diagonal[1] / radius, hs[4],
                         hs[5] + diagonal[2] / radius};
    double step[3];
    const bool solved = solve_spd3(a, gs, step);
    step[0] = -step[0], step[1] = -step[1], step[2] = -step[2];
    reuse_diagonal = true;
    double model_cost_change = 0.;
    if (solved) {
      const double hstep[3] = {hs[0] * step[0] + hs[1] * step[1] + hs[2] * step[2],
                               hs[1] * step[0] + hs[3] * step[1] + hs[4] * step[2],
                               hs[2] * step[0] + hs[4] * step[1] + hs[5] * step[2]};
      model_cost_change = -((step[0] * gs[0] + step[1] * gs[1] + step[2] * gs[2]) +
                            0.5 * (step[0] * hstep[0] + step[1] * hstep[1] + step[2] * hstep[2]));
    }
    if (!solved || !(model_cost_change > 0.)) {
      if (++num_invalid >= 5) { termination = 5; break; }
      radius = radius / decrease_factor;
      decrease_factor *= 2.;
      continue;
    }
    num_invalid = 0;
    const double delta[3] = {step[0] * scale[0], step[1] * scale[1], step[2] * scale[2]};
    const double cand[3] = {x[0] + delta[0], x[1] + delta[1], x[2] + delta[2]};
    Sums cs;
    // the Jacobian sums are kept if the step is accepted; a failed evaluation makes the
    // candidate infinitely expensive (the step is rejected)
    const bool cand_ok = evaluate_any(j, pts, cand, lane, &cs);
    const double cand_cost = cand_ok ? 0.5 * cs.rr : 1.7976931348623157e308;
    if (norm3(delta) <= 1e-8 * (x_norm + 1e-8)) { termination = 1; break; }
    if (fabs(x_cost - cand_cost) <= 1e-6 * x_cost) { termination = 2; break; }
    const double relative_decrease = (ev_current - cand_cost) / model_cost_change;
    const double historical = (ev_reference - cand_cost) / (acc_reference + model_cost_change);
    const double quality = cand_ok ? fmax(relative_decrease, historical) : -1.7976931348623157e308;
    if (quality > 1e-3) {
      x[0] = cand[0], x[1] = cand[1], x[2] = cand[2];
      x_norm = norm3(x);
      cur = cs;
      x_cost = cand_cost;
      last_successful = true;
      const double t = 2. * quality - 1.;
      radius = radius / fmax(1. / 3., 1. - t * t * t);
      radius = fmin(1e16, radius);
      decrease_factor = 2.;
      reuse_diagonal = false;
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
    } else {
      radius = radius / decrease_factor;
      decrease_factor *= 2.;
    }
  }
  if (lane == 0) {
    csm_ceres_result2d r;
    r.pose[0] = x_min[0];
    r.pose[1] = x_min[1];
    r.pose[2] = x_min[2];
    r.initial_cost = initial_cost;
    r.final_cost = minimum_cost;
    r.iterations = iteration;
    r.num_successful_steps = successful;
    r.num_unsuccessful_steps = unsuccessful;
    r.termination = termination;
    results[job] = r;
  }
}

}  // namespace

extern "C" int csm_ceres_match2d_batch(csm_ctx* ctx, const csm_ceres_job2d* jobs, int num_jobs,
                                       csm_ceres_result2d* results) {
  CSM_REQUIRE(ctx != nullptr && jobs != nullptr && results != nullptr, "NULL argument");
  CSM_REQUIRE(num_jobs >= 1, "num_jobs must be positive");
  CSM_CUDA(cudaSetDevice(ctx->device));
  auto up = [](size_t v) { return (v + 255) & ~static_cast<size_t>(255); };
  std::vector<CeresJobDev> dev(num_jobs);
  std::map<const void*, size_t> grid_at, table_at;
  std::map<const void*, size_t> grid_size;
  size_t n_pts = 0, grid_bytes = 0, table_bytes = 0;
  for (int i = 0; i < num_jobs; ++i) {
    const csm_ceres_job2d& in = jobs[i];
    CSM_REQUIRE((in.cells || in.cells_dev) && in.cost_table && in.points_xyz, "ceres job has a NULL array");
    CSM_REQUIRE(in.num_points >= 1 && in.num_x >= 1 && in.num_y >= 1 && in.resolution > 0.,
                "ceres job sizes out of range");
    CSM_REQUIRE(in.occupied_space_weight > 0. && in.translation_weight > 0. && in.rotation_weight > 0.,
                "ceres weights must be positive (ceres_scan_matcher_2d.cc:75,94,99)");
    CSM_REQUIRE(in.max_num_iterations >= 0, "max_num_iterations must not be negative");
    if (in.cells_dev == nullptr && !grid_at.count(in.cells)) {
      grid_at[in.cells] = grid_bytes;
      grid_size[in.cells] = sizeof(uint16_t) * static_cast<size_t>(in.num_x) * in.num_y;
      grid_bytes += up(grid_size[in.cells]);
    }
    if (!table_at.count(in.cost_table)) {
      table_at[in.cost_table] = table_bytes;
      table_bytes += up(sizeof(float) * 65536);
    }
    if (in.weight_cells != nullptr) {
      CSM_REQUIRE(in.weight_table != nullptr && in.truncation_distance > 0.,
                  "TSDF ceres job needs a weight table and a positive truncation distance");
      if (!grid_at.count(in.weight_cells)) {
        grid_at[in.weight_cells] = grid_bytes;
        grid_size[in.weight_cells] = sizeof(uint16_t) * static_cast<size_t>(in.num_x) * in.num_y;
        grid_bytes += up(grid_size[in.weight_cells]);
      }
      if (!table_at.count(in.weight_table)) {
        table_at[in.weight_table] = table_bytes;
        table_bytes += up(sizeof(float) * 65536);
      }
    }
    CeresJobDev& j = dev[i];
    j.pts_off = static_cast<long long>(n_pts);
    j.num_x = in.num_x;
    j.num_y = in.num_y;
    j.n = in.num_points;
    j.max_x = in.max_x;
    j.max_y = in.max_y;
    j.res = in.resolution;
    for (int k = 0; k < 3; ++k) j.init[k] = in.initial_pose[k];
    j.target[0] = in.target_translation[0];
    j.target[1] = in.target_translation[1];
    j.w_occ = in.occupied_space_weight;
    j.w_t = in.translation_weight;
    j.w_r = in.rotation_weight;
    j.max_iterations = in.max_num_iterations;
    j.nonmonotonic = in.use_nonmonotonic_steps;
    n_pts += in.num_points;
  }
  const size_t o_jobs = 0;
  const size_t o_pts = o_jobs + up(sizeof(CeresJobDev) * num_jobs);
  const size_t o_tab = o_pts + up(sizeof(float) * 3 * n_pts);
  const size_t o_grid = o_tab + table_bytes;
  const size_t o_res = o_grid + grid_bytes;
  const size_t total = o_res + up(sizeof(csm_ceres_result2d) * num_jobs);
  int rc = ctx->tmp2.reserve(total);
  if (rc == CSM_OK) rc = ctx->pin0.reserve(total);
  if (rc != CSM_OK) return rc;
  char* pin = ctx->pin0.as<char>();
  char* base = ctx->tmp2.as<char>();
  for (int i = 0; i < num_jobs; ++i) {
    const csm_ceres_job2d& in = jobs[i];
    CeresJobDev& j = dev[i];
    memcpy(pin + o_pts + sizeof(float) * 3 * j.pts_off, in.points_xyz, sizeof(float) * 3 * j.n);
    j.cells = in.cells_dev != nullptr ? in.cells_dev
                                      : reinterpret_cast<const uint16_t*>(base + o_grid + grid_at[in.cells]);
    j.cost_table = reinterpret_cast<const float*>(base + o_tab + table_at[in.cost_table]);
    j.weight_cells = in.weight_cells
                         ? reinterpret_cast<const uint16_t*>(base + o_grid + grid_at[in.weight_cells])
                         : nullptr;
    j.weight_table = in.weight_cells
                         ? reinterpret_cast<const float*>(base + o_tab + table_at[in.weight_table])
                         : nullptr;
    j.max_cc = static_cast<float>(in.truncation_distance);
  }
  for (const auto& kv : grid_at) memcpy(pin + o_grid + kv.second, kv.first, grid_size[kv.first]);
  for (const auto& kv : table_at) memcpy(pin + o_tab + kv.second, kv.first, sizeof(float) * 65536);
  memcpy(pin + o_jobs, dev.data(), sizeof(CeresJobDev) * num_jobs);
  cudaStream_t st = ctx->stream;
  CSM_CUDA(cudaMemcpyAsync(base, pin, o_res, cudaMemcpyHostToDevice, st));
  ++ctx->launches, k_ceres_lm<<<(num_jobs + kCeresWarps - 1) / kCeresWarps, kCeresWarps * 32, 0, st>>>(
      reinterpret_cast<const CeresJobDev*>(base + o_jobs), reinterpret_cast<const float*>(base + o_pts),
      num_jobs, reinterpret_cast<csm_ceres_result2d*>(base + o_res));
  CSM_CUDA(cudaGetLastError());
  CSM_CUDA(cudaMemcpyAsync(pin + o_res, base + o_res, sizeof(csm_ceres_result2d) * num_jobs,
                           cudaMemcpyDeviceToHost, st));
  CSM_CUDA(cudaStreamSynchronize(st));
  memcpy(results, pin + o_res, sizeof(csm_ceres_result2d) * num_jobs);
  return CSM_OK;
}
