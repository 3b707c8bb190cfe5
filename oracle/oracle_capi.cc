// TEST INFRASTRUCTURE ONLY (see csm_oracle.h).  Flat C entry points over the oracle
// for ctypes: used by tests/ as the checker and by bench.py as the timed CPU
// baseline.  Never linked into the product library.
#include <chrono>
#include <cstring>

#include "cpu_pool.h"
#include "csm_oracle.h"
#include "csm_ceres.h"
#include "csm_oracle3d.h"
#include "csm_tsdf_inserter.h"

using namespace oracle;

extern "C" {

struct oracle_result2d {
  int32_t found;
  int32_t scan_index, x_index_offset, y_index_offset;
  float score;
  int32_t num_scans;
  double pose_x, pose_y, pose_theta;
  uint64_t scored_top, scored_lower;
};

static PointCloud ToCloud(const float* xyz, int n) {
  PointCloud c(n);
  for (int i = 0; i < n; ++i) c[i] = Vec3f{xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]};
  return c;
}
static MapLimits MakeLimits(double res, double max_x, double max_y, int nx, int ny) {
  MapLimits l;
  l.resolution = res;
  l.max_x = max_x;
  l.max_y = max_y;
  l.cell_limits.num_x_cells = nx;
  l.cell_limits.num_y_cells = ny;
  return l;
}

// --- tables -----------------------------------------------------------------
void oracle_correspondence_cost_table(float* out65536) {
  std::memcpy(out65536, ValueToCorrespondenceCostTable().data(), 65536 * sizeof(float));
}
uint16_t oracle_probability_to_cost_value(float probability) {
  return CorrespondenceCostToValue(1.f - probability);
}

// --- grid / matcher handles ---------------------------------------------------
void* oracle_grid2d_create(const uint16_t* cells, int nx, int ny, double res, double max_x,
                           double max_y) {
  return new ProbabilityGrid(MakeLimits(res, max_x, max_y, nx, ny), cells);
}
void oracle_grid2d_destroy(void* g) { delete static_cast<ProbabilityGrid*>(g); }

void* oracle_fcsm2d_create(void* grid, double linear_window, double angular_window, int depth) {
  FastCorrelativeScanMatcherOptions2D o;
  o.linear_search_window = linear_window;
  o.angular_search_window = angular_window;
  o.branch_and_bound_depth = depth;
  return new FastCorrelativeScanMatcher2D(*static_cast<ProbabilityGrid*>(grid), o);
}
// FastCorrelativeScanMatcher2D over a TSDF2D handle (oracle_tsdf2d_create): the reference takes
// any Grid2D (fast_correlative_scan_matcher_2d.cc:91-113 uses 1.f - |GetCorrespondenceCost|).
void* oracle_fcsm2d_create_tsdf(void* grid, double linear_window, double angular_window, int depth) {
  FastCorrelativeScanMatcherOptions2D o;
  o.linear_search_window = linear_window;
  o.angular_search_window = angular_window;
  o.branch_and_bound_depth = depth;
  return new FastCorrelativeScanMatcher2D(*static_cast<TSDF2D*>(grid), o);
}
void oracle_fcsm2d_destroy(void* m) { delete static_cast<FastCorrelativeScanMatcher2D*>(m); }

// Copies precomputation level `level` (wide_x * wide_y bytes) if out != NULL.
int oracle_fcsm2d_level(void* m, int level, uint8_t* out, int* wide_x, int* wide_y) {
  const auto* fm = static_cast<FastCorrelativeScanMatcher2D*>(m);
  if (level < 0 || level > fm->stack().max_depth()) return -1;
  const PrecomputationGrid2D& pg = fm->stack().Get(level);
  *wide_x = pg.wide_x();
  *wide_y = pg.wide_y();
  if (out) std::memcpy(out, pg.cells().data(), pg.cells().size());
  return 0;
}

static void Fill(oracle_result2d* out, bool ok, const FastCorrelativeScanMatcher2D::Winner& w,
                 float score, const Rigid2d& pose, const MatchStats& st) {
  out->found = ok ? 1 : 0;
  out->scan_index = w.scan_index;
  out->x_index_offset = w.x_index_offset;
  out->y_index_offset = w.y_index_offset;
  out->score = ok ? score : w.score;
  out->num_scans = static_cast<int32_t>(st.num_scans);
  out->pose_x = pose.tx;
  out->pose_y = pose.ty;
  out->pose_theta = pose.angle;
  out->scored_top = st.scored_top;
  out->scored_lower = st.scored_lower;
}

int oracle_fcsm2d_match(void* m, int full_submap, double init_x, double init_y, double init_theta,
                        const float* xyz, int n, float min_score, oracle_result2d* out) {
  const auto* fm = static_cast<FastCorrelativeScanMatcher2D*>(m);
  const PointCloud cloud = ToCloud(xyz, n);
  float score = 0.f;
  Rigid2d pose;
  FastCorrelativeScanMatcher2D::Winner w;
  MatchStats st;
  bool ok;
  if (full_submap) {
    ok = fm->MatchFullSubmap(cloud, min_score, &score, &pose, &w, &st);
  } else {
    Rigid2d init;
    init.tx = init_x;
    init.ty = init_y;
    init.angle = init_theta;
    ok = fm->Match(init, cloud, min_score, &score, &pose, &w, &st);
  }
  Fill(out, ok, w, score, pose, st);
  return 0;
}

// --- search parameters + discretisation (for the bit-exact index test) --------
int oracle_search_parameters2d(double linear_window, double angular_window, const float* xyz, int n,
                               double res, int* num_angular_perturbations, double* step,
                               int* num_scans, int* num_linear_perturbations) {
  const SearchParameters sp(linear_window, angular_window, ToCloud(xyz, n), res);
  *num_angular_perturbations = sp.num_angular_perturbations;
  *step = sp.angular_perturbation_step_size;
  *num_scans = sp.num_scans;
  *num_linear_perturbations = sp.linear_bounds.empty() ? 0 : sp.linear_bounds[0].max_x;
  return 0;
}

// Rotates by init_theta, generates the rotated scans, discretises them and shrinks
// the bounds exactly like MatchWithSearchParameters does
// (fast_correlative_scan_matcher_2d.cc:234-246).  out_xy: S*n*2 int32 (x, y);
// out_bounds: S*4 int32 (min_x, max_x, min_y, max_y).
int oracle_discretize2d(double res, double max_x, double max_y, int nx, int ny, const float* xyz,
                        int n, double init_x, double init_y, double init_theta,
                        int num_linear_perturbations, int num_angular_perturbations, double step,
                        int shrink, int32_t* out_xy, int32_t* out_bounds) {
  const MapLimits limits = MakeLimits(res, max_x, max_y, nx, ny);
  SearchParameters sp(num_linear_perturbations, num_angular_perturbations, step, res);
  const PointCloud rotated = RotatePointCloudYaw(ToCloud(xyz, n), static_cast<float>(init_theta));
  const std::vector<PointCloud> scans = GenerateRotatedScans(rotated, sp);
  const std::vector<DiscreteScan2D> d =
      DiscretizeScans(limits, scans, static_cast<float>(init_x), static_cast<float>(init_y));
  if (shrink) sp.ShrinkToFit(d, limits.cell_limits);
  for (int s = 0; s < sp.num_scans; ++s) {
    for (int i = 0; i < n; ++i) {
      out_xy[(static_cast<size_t>(s) * n + i) * 2 + 0] = d[s][i].x;
      out_xy[(static_cast<size_t>(s) * n + i) * 2 + 1] = d[s][i].y;
    }
    out_bounds[4 * s + 0] = sp.linear_bounds[s].min_x;
    out_bounds[4 * s + 1] = sp.linear_bounds[s].max_x;
    out_bounds[4 * s + 2] = sp.linear_bounds[s].min_y;
    out_bounds[4 * s + 3] = sp.linear_bounds[s].max_y;
  }
  return 0;
}

// --- real-time matcher ---------------------------------------------------------
double oracle_rt_match2d(void* grid, double linear_window, double angular_window,
                         double translation_delta_cost_weight, double rotation_delta_cost_weight,
                         double init_x, double init_y, double init_theta, const float* xyz, int n,
                         oracle_result2d* out) {
  RealTimeCorrelativeScanMatcherOptions o;
  o.linear_search_window = linear_window;
  o.angular_search_window = angular_window;
  o.translation_delta_cost_weight = translation_delta_cost_weight;
  o.rotation_delta_cost_weight = rotation_delta_cost_weight;
  const RealTimeCorrelativeScanMatcher2D matcher(o);
  Rigid2d init, pose;
  init.tx = init_x;
  init.ty = init_y;
  init.angle = init_theta;
  FastCorrelativeScanMatcher2D::Winner w;
  const double score =
      matcher.Match(init, ToCloud(xyz, n), *static_cast<ProbabilityGrid*>(grid), &pose, &w);
  MatchStats st;
  Fill(out, true, w, static_cast<float>(score), pose, st);
  return score;
}

// --- ProbabilityGridRangeDataInserter2D -----------------------
// Inserts one range data into the grid handle (GrowLimits included, so the grid may
// change size: query it with oracle_grid2d_info / oracle_grid2d_cells afterwards).
void oracle_grid2d_insert(void* grid, double hit_probability, double miss_probability,
                          int insert_free_space, const float* origin_xy, const float* returns_xyz,
                          int num_returns, const float* misses_xyz, int num_misses) {
  RangeData rd;
  rd.origin = Vec3f{origin_xy[0], origin_xy[1], 0.f};
  rd.returns = ToCloud(returns_xyz, num_returns);
  rd.misses = ToCloud(misses_xyz, num_misses);
  ProbabilityGridRangeDataInserter2D(hit_probability, miss_probability, insert_free_space != 0)
      .Insert(rd, static_cast<ProbabilityGrid*>(grid));
}
void oracle_grid2d_info(void* grid, int* nx, int* ny, double* max_x, double* max_y, double* res) {
  const MapLimits& l = static_cast<ProbabilityGrid*>(grid)->limits();
  *nx = l.cell_limits.num_x_cells;
  *ny = l.cell_limits.num_y_cells;
  *max_x = l.max_x;
  *max_y = l.max_y;
  *res = l.resolution;
}
void oracle_grid2d_cells(void* grid, uint16_t* out) {
  const std::vector<uint16_t>& c = static_cast<ProbabilityGrid*>(grid)->cells();
  std::memcpy(out, c.data(), sizeof(uint16_t) * c.size());
}
// ComputeLookupTableToApplyCorrespondenceCostOdds(Odds(p)), 32768 entries.
void oracle_apply_odds_table(double probability, uint16_t* out32768) {
  const std::vector<uint16_t> t =
      ComputeLookupTableToApplyCorrespondenceCostOdds(Odds(static_cast<float>(probability)));
  std::memcpy(out32768, t.data(), sizeof(uint16_t) * t.size());
}

// --- TSDF grid + real-time matcher on it ----------------------------------------
void* oracle_tsdf2d_create(const uint16_t* tsd_cells, const uint16_t* weight_cells, int nx, int ny,
                           double res, double max_x, double max_y, float truncation_distance,
                           float max_weight) {
  return new TSDF2D(MakeLimits(res, max_x, max_y, nx, ny), truncation_distance, max_weight, tsd_cells,
                    weight_cells);
}
void oracle_tsdf2d_destroy(void* g) { delete static_cast<TSDF2D*>(g); }
// TSDFRangeDataInserter2D::Insert on a TSDF2D handle (GrowLimits included); options9 as in
// ref_capi.cc's entry of the same name.
void oracle_tsdf2d_insert(void* grid, const double* options9, const float* origin_xy, const float* returns_xyz,
                          int num_returns) {
  TSDFRangeDataInserterOptions2D o;
  o.truncation_distance = options9[0];
  o.maximum_weight = options9[1];
  o.update_free_space = options9[2] != 0.;
  o.num_normal_samples = static_cast<int>(options9[3]);
  o.sample_radius = static_cast<float>(options9[4]);
  o.project_sdf_distance_to_scan_normal = options9[5] != 0.;
  o.update_weight_range_exponent = static_cast<int>(options9[6]);
  o.update_weight_angle_scan_normal_to_ray_kernel_bandwidth = options9[7];
  o.update_weight_distance_cell_to_hit_kernel_bandwidth = options9[8];
  TSDFRangeDataInserter2D(o).Insert(Vec3f{origin_xy[0], origin_xy[1], 0.f}, ToCloud(returns_xyz, num_returns),
                                    static_cast<TSDF2D*>(grid));
}
void oracle_tsdf2d_info(void* grid, int* nx, int* ny, double* max_x, double* max_y, double* res) {
  const MapLimits& l = static_cast<TSDF2D*>(grid)->limits();
  *nx = l.cell_limits.num_x_cells;
  *ny = l.cell_limits.num_y_cells;
  *max_x = l.max_x;
  *max_y = l.max_y;
  *res = l.resolution;
}
void oracle_tsdf2d_cells(void* grid, uint16_t* tsd_out, uint16_t* weight_out) {
  const TSDF2D& t = *static_cast<TSDF2D*>(grid);
  std::memcpy(tsd_out, t.tsd_cells().data(), sizeof(uint16_t) * t.tsd_cells().size());
  std::memcpy(weight_out, t.weight_cells().data(), sizeof(uint16_t) * t.weight_cells().size());
}
// TSDValueConverter::TSDToValue / WeightToValue of n floats (for painting fixtures).
void oracle_tsdf2d_values(float truncation_distance, float max_weight, const float* tsd,
                          const float* weight, int n, uint16_t* tsd_values, uint16_t* weight_values) {
  const TSDF2D g(MakeLimits(1., 0., 0., 1, 1), truncation_distance, max_weight);
  for (int i = 0; i < n; ++i) {
    tsd_values[i] = g.TSDToValue(tsd[i]);
    weight_values[i] = g.WeightToValue(weight[i]);
  }
}
double oracle_rt_match2d_tsdf(void* grid, double linear_window, double angular_window,
                              double translation_delta_cost_weight, double rotation_delta_cost_weight,
                              double init_x, double init_y, double init_theta, const float* xyz, int n,
                              oracle_result2d* out) {
  RealTimeCorrelativeScanMatcherOptions o;
  o.linear_search_window = linear_window;
  o.angular_search_window = angular_window;
  o.translation_delta_cost_weight = translation_delta_cost_weight;
  o.rotation_delta_cost_weight = rotation_delta_cost_weight;
  const RealTimeCorrelativeScanMatcher2D matcher(o);
  Rigid2d init, pose;
  init.tx = init_x;
  init.ty = init_y;
  init.angle = init_theta;
  FastCorrelativeScanMatcher2D::Winner w;
  const double score = matcher.Match(init, ToCloud(xyz, n), *static_cast<TSDF2D*>(grid), &pose, &w);
  MatchStats st;
  Fill(out, true, w, static_cast<float>(score), pose, st);
  return score;
}
// ScoreCandidates of ONE candidate at zero rotation (the shape of the reference's
// real_time_correlative_scan_matcher_2d_test.cc:143-201).
float oracle_rt_score_candidate_tsdf(void* grid, const float* xyz, int n, int x_offset, int y_offset) {
  const TSDF2D& g = *static_cast<TSDF2D*>(grid);
  const SearchParameters sp(0, 0, 0., 0.);
  const std::vector<PointCloud> scans = GenerateRotatedScans(ToCloud(xyz, n), sp);
  const std::vector<DiscreteScan2D> discrete = DiscretizeScans(g.limits(), scans, 0.f, 0.f);
  RealTimeCorrelativeScanMatcherOptions o;
  o.translation_delta_cost_weight = 0.;
  o.rotation_delta_cost_weight = 0.;
  std::vector<Candidate2D> cands;
  cands.emplace_back(0, x_offset, y_offset, sp);
  RealTimeCorrelativeScanMatcher2D(o).ScoreCandidates(g, discrete, sp, &cands);
  return cands[0].score;
}

// --- CeresScanMatcher2D ------------------------------------
// options5 = occupied_space_weight, translation_weight, rotation_weight,
// use_nonmonotonic_steps, max_num_iterations; summary6 = initial_cost, final_cost,
// iterations, successful steps, unsuccessful steps, termination.
void oracle_ceres_match2d(void* grid, const double* options5, const double* target2,
                          const double* init3, const float* xyz, int n, double* pose3,
                          double* summary6) {
  CeresScanMatcherOptions2D o;
  o.occupied_space_weight = options5[0];
  o.translation_weight = options5[1];
  o.rotation_weight = options5[2];
  o.use_nonmonotonic_steps = options5[3] != 0.;
  o.max_num_iterations = static_cast<int>(options5[4]);
  Rigid2d init, pose;
  init.tx = init3[0];
  init.ty = init3[1];
  init.angle = init3[2];
  CeresSummary s;
  CeresScanMatcher2D(o).Match(target2[0], target2[1], init, ToCloud(xyz, n),
                              *static_cast<ProbabilityGrid*>(grid), &pose, &s);
  pose3[0] = pose.tx;
  pose3[1] = pose.ty;
  pose3[2] = pose.angle;
  summary6[0] = s.initial_cost;
  summary6[1] = s.final_cost;
  summary6[2] = s.iterations;
  summary6[3] = s.num_successful_steps;
  summary6[4] = s.num_unsuccessful_steps;
  summary6[5] = s.termination;
}
// The same on a TSDF2D handle (oracle_tsdf2d_create).
void oracle_ceres_match2d_tsdf(void* grid, const double* options5, const double* target2,
                               const double* init3, const float* xyz, int n, double* pose3,
                               double* summary6) {
  CeresScanMatcherOptions2D o;
  o.occupied_space_weight = options5[0];
  o.translation_weight = options5[1];
  o.rotation_weight = options5[2];
  o.use_nonmonotonic_steps = options5[3] != 0.;
  o.max_num_iterations = static_cast<int>(options5[4]);
  Rigid2d init, pose;
  init.tx = init3[0];
  init.ty = init3[1];
  init.angle = init3[2];
  CeresSummary s;
  CeresScanMatcher2D(o).Match(target2[0], target2[1], init, ToCloud(xyz, n), *static_cast<TSDF2D*>(grid),
                              &pose, &s);
  pose3[0] = pose.tx;
  pose3[1] = pose.ty;
  pose3[2] = pose.angle;
  summary6[0] = s.initial_cost;
  summary6[1] = s.final_cost;
  summary6[2] = s.iterations;
  summary6[3] = s.num_successful_steps;
  summary6[4] = s.num_unsuccessful_steps;
  summary6[5] = s.termination;
}
int oracle_ceres_evaluate2d_tsdf(void* grid, const double* options5, const double* target2,
                                 double initial_angle, const double* pose3, const float* xyz, int n,
                                 double* residuals, double* jacobian) {
  CeresScanMatcherOptions2D o;
  o.occupied_space_weight = options5[0];
  o.translation_weight = options5[1];
  o.rotation_weight = options5[2];
  std::vector<double> r, j;
  if (!CeresScanMatcher2D(o).Evaluate(pose3, target2, initial_angle, ToCloud(xyz, n),
                                      *static_cast<TSDF2D*>(grid), &r, &j))
    return 0;
  std::memcpy(residuals, r.data(), sizeof(double) * r.size());
  std::memcpy(jacobian, j.data(), sizeof(double) * j.size());
  return 1;
}
// Residuals (n + 3) and row-major Jacobian ((n + 3) x 3) at a pose, for derivative checks.
void oracle_ceres_evaluate2d(void* grid, const double* options5, const double* target2,
                             double initial_angle, const double* pose3, const float* xyz, int n,
                             double* residuals, double* jacobian) {
  CeresScanMatcherOptions2D o;
  o.occupied_space_weight = options5[0];
  o.translation_weight = options5[1];
  o.rotation_weight = options5[2];
  std::vector<double> r, j;
  CeresScanMatcher2D(o).Evaluate(pose3, target2, initial_angle, ToCloud(xyz, n),
                                 *static_cast<ProbabilityGrid*>(grid), &r, &j);
  std::memcpy(residuals, r.data(), sizeof(double) * r.size());
  std::memcpy(jacobian, j.data(), sizeof(double) * j.size());
}

// --- thread-pooled constraint builder (timed CPU baseline) ---------------------
// grids[num_submaps], submap_pose[num_submaps*3]; searches: submap index, node id,
// full flag, relative pose [3], cloud given by offsets into xyz (offsets has
// num_searches+1 entries, in points).  One NotifyEndOfNode per search (the caller
// models one node per search).  Results in submission order; found==0 rows are the
// reference's null constraint slots.  Returns wall seconds from the first
// MaybeAdd* call until the WhenDone callback has run (matcher construction
// included, exactly as in the reference where it is one more pool task).
double oracle_constraint_batch(int num_threads, void* const* grids, const double* submap_pose,
                               int num_submaps, int num_searches, const int32_t* search_submap,
                               const int32_t* search_full, const double* search_rel_pose,
                               const float* xyz, const int64_t* offsets, double sampling_ratio,
                               double max_constraint_distance, double min_score,
                               double global_min_score, double linear_window, double angular_window,
                               int depth, oracle_result2d* out, double* out_stats4) {
  ConstraintBuilderOptions o;
  o.sampling_ratio = sampling_ratio;
  o.max_constraint_distance = max_constraint_distance;
  o.min_score = min_score;
  o.global_localization_min_score = global_min_score;
  o.fast_correlative_scan_matcher_options.linear_search_window = linear_window;
  o.fast_correlative_scan_matcher_options.angular_search_window = angular_window;
  o.fast_correlative_scan_matcher_options.branch_and_bound_depth = depth;
  std::vector<ConstraintBuilder2D::Submap> submaps(num_submaps);
  for (int i = 0; i < num_submaps; ++i) {
    submaps[i].grid = static_cast<const ProbabilityGrid*>(grids[i]);
    submaps[i].local_pose.tx = submap_pose[3 * i];
    submaps[i].local_pose.ty = submap_pose[3 * i + 1];
    submaps[i].local_pose.angle = submap_pose[3 * i + 2];
  }
  std::vector<PointCloud> clouds(num_searches);
  for (int i = 0; i < num_searches; ++i)
    clouds[i] = ToCloud(xyz + 3 * offsets[i], static_cast<int>(offsets[i + 1] - offsets[i]));
  for (int i = 0; i < num_searches; ++i) out[i] = oracle_result2d{};

  ThreadPool pool(num_threads);
  ConstraintBuilder2D builder(o, &pool);
  std::mutex m;
  std::condition_variable cv;
  bool done = false;
  std::vector<OracleConstraint> result;
  const auto t0 = std::chrono::steady_clock::now();
  for (int i = 0; i < num_searches; ++i) {
    const int s = search_submap[i];
    if (search_full[i]) {
      builder.MaybeAddGlobalConstraint(s, &submaps[s], i, &clouds[i]);
    } else {
      Rigid2d rel;
      rel.tx = search_rel_pose[3 * i];
      rel.ty = search_rel_pose[3 * i + 1];
      rel.angle = search_rel_pose[3 * i + 2];
      builder.MaybeAddConstraint(s, &submaps[s], i, &clouds[i], rel);
    }
    builder.NotifyEndOfNode();
  }
  builder.WhenDone([&](const std::vector<OracleConstraint>& r) {
    std::lock_guard<std::mutex> l(m);
    result = r;
    done = true;
    cv.notify_all();
  });
  {
    std::unique_lock<std::mutex> l(m);
    cv.wait(l, [&] { return done; });
  }
  const auto t1 = std::chrono::steady_clock::now();
  for (const OracleConstraint& c : result) {
    oracle_result2d& r = out[c.node_id];
    r.found = 1;
    r.scan_index = c.winner.scan_index;
    r.x_index_offset = c.winner.x_index_offset;
    r.y_index_offset = c.winner.y_index_offset;
    r.score = c.score;
    r.pose_x = c.tx;
    r.pose_y = c.ty;
    r.pose_theta = c.angle;
  }
  const MatchStats st = builder.stats();
  if (out_stats4) {
    out_stats4[0] = static_cast<double>(st.scored_top);
    out_stats4[1] = static_cast<double>(st.scored_lower);
    out_stats4[2] = static_cast<double>(st.num_scans);
    out_stats4[3] = static_cast<double>(st.num_points);
  }
  for (int i = 0; i < num_submaps; ++i) builder.DeleteScanMatcher(i);
  return std::chrono::duration<double>(t1 - t0).count();
}

// --- 3D path ---------------------------------------------------------------------
struct oracle_result3d {
  int32_t found, scan_index, offset[3], num_scans;
  float score, rotational_score, low_resolution_score;
  float pose[7];  // tx, ty, tz, qw, qx, qy, qz
  uint64_t candidates_scored;
};

void* oracle_grid3d_create(float resolution, const int32_t* xyz, const uint16_t* values, int n) {
  HybridGrid* g = new HybridGrid(resolution);
  for (int i = 0; i < n; ++i)
    *g->mutable_value(Array3i{xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]}) = values[i];
  return g;
}
void oracle_grid3d_destroy(void* g) { delete static_cast<HybridGrid*>(g); }
// RangeDataInserter3D::Insert of one range data (test fixture builder).
void oracle_grid3d_insert(void* g, const float* origin3, const float* returns, int n, float hit_probability,
                          float miss_probability, int num_free_space_voxels) {
  std::vector<Vec3f> r;
  for (int i = 0; i < n; ++i) r.push_back(Vec3f{returns[3 * i], returns[3 * i + 1], returns[3 * i + 2]});
  RangeDataInserter3D(hit_probability, miss_probability, num_free_space_voxels)
      .Insert(Vec3f{origin3[0], origin3[1], origin3[2]}, r, static_cast<HybridGrid*>(g));
}
// Stored cells with a non-zero value, as the reference's HybridGrid::Iterator visits
// them (hybrid_grid.h:327-419 skips default values); NULL outputs: count only.
int oracle_grid3d_cells(void* g, int32_t* xyz, uint16_t* values) {
  int n = 0;
  static_cast<HybridGrid*>(g)->ForEach([&](const Array3i& i, uint16_t v) {
    if (v == 0) return;
    if (xyz != nullptr) {
      xyz[3 * n] = i.x;
      xyz[3 * n + 1] = i.y;
      xyz[3 * n + 2] = i.z;
      values[n] = v;
    }
    ++n;
  });
  return n;
}
int oracle_grid3d_size(void* g) { return static_cast<HybridGrid*>(g)->grid_size(); }

static FastCorrelativeScanMatcherOptions3D Options3D(const double* o7) {
  FastCorrelativeScanMatcherOptions3D o;
  o.branch_and_bound_depth = static_cast<int>(o7[0]);
  o.full_resolution_depth = static_cast<int>(o7[1]);
  o.min_rotational_score = o7[2];
  o.min_low_resolution_score = o7[3];
  o.linear_xy_search_window = o7[4];
  o.linear_z_search_window = o7[5];
  o.angular_search_window = o7[6];
  return o;
}

struct Matcher3DHolder {
  Histogram submap_histogram;
  std::unique_ptr<FastCorrelativeScanMatcher3D> matcher;
};

void* oracle_fcsm3d_create(void* grid, void* low_grid, const float* hist, int n_hist,
                           const double* options7) {
  Matcher3DHolder* h = new Matcher3DHolder;
  h->submap_histogram.assign(hist, hist + n_hist);
  h->matcher.reset(new FastCorrelativeScanMatcher3D(*static_cast<HybridGrid*>(grid),
                                                    static_cast<HybridGrid*>(low_grid),
                                                    &h->submap_histogram, Options3D(options7)));
  return h;
}
void oracle_fcsm3d_destroy(void* m) { delete static_cast<Matcher3DHolder*>(m); }

// Dense copy of precomputation level `level`: first call with out == NULL fills the
// bounding box (lo/hi inclusive over stored cells, all zero if the level is empty).
int oracle_fcsm3d_level(void* m, int level, int32_t* lo3, int32_t* hi3, uint8_t* out) {
  const PrecomputationGrid3D& g = static_cast<Matcher3DHolder*>(m)->matcher->stack().Get(level);
  if (out == nullptr) {
    bool first = true;
    g.ForEach([&](const Array3i& i, uint8_t) {
      const int v[3] = {i.x, i.y, i.z};
      for (int k = 0; k < 3; ++k) {
        lo3[k] = first ? v[k] : std::min(lo3[k], v[k]);
        hi3[k] = first ? v[k] : std::max(hi3[k], v[k]);
      }
      first = false;
    });
    if (first) for (int k = 0; k < 3; ++k) lo3[k] = hi3[k] = 0;
    return static_cast<int>(g.size());
  }
  const int64_t nx = hi3[0] - lo3[0] + 1, ny = hi3[1] - lo3[1] + 1, nz = hi3[2] - lo3[2] + 1;
  std::fill(out, out + nx * ny * nz, uint8_t(0));
  g.ForEach([&](const Array3i& i, uint8_t v) {
    if (i.x < lo3[0] || i.x > hi3[0] || i.y < lo3[1] || i.y > hi3[1] || i.z < lo3[2] || i.z > hi3[2]) return;
    out[((i.z - lo3[2]) * ny + (i.y - lo3[1])) * nx + (i.x - lo3[0])] = v;
  });
  return static_cast<int>(g.size());
}

static NodeData3D MakeNode(const float* hi, int n_hi, const float* lo, int n_lo, const float* hist,
                           int n_hist, const double* gravity4) {
  NodeData3D d;
  for (int i = 0; i < n_hi; ++i) d.high_resolution_point_cloud.push_back(Vec3f{hi[3 * i], hi[3 * i + 1], hi[3 * i + 2]});
  for (int i = 0; i < n_lo; ++i) d.low_resolution_point_cloud.push_back(Vec3f{lo[3 * i], lo[3 * i + 1], lo[3 * i + 2]});
  d.rotational_scan_matcher_histogram.assign(hist, hist + n_hist);
  for (int k = 0; k < 4; ++k) d.gravity_alignment[k] = gravity4[k];
  return d;
}
static Rigid3f Pose7(const float* p) {
  Rigid3f r;
  r.t = Vec3f{p[0], p[1], p[2]};
  r.q = Quat4f{p[3], p[4], p[5], p[6]};
  return r;
}

// full != 0: MatchFullSubmap with the rotations of node_pose7 / submap_pose7.
void oracle_fcsm3d_match(void* m, int full, const float* node_pose7, const float* submap_pose7,
                         const float* hi, int n_hi, const float* lo, int n_lo, const float* hist,
                         int n_hist, const double* gravity4, float min_score, oracle_result3d* out) {
  const FastCorrelativeScanMatcher3D& matcher = *static_cast<Matcher3DHolder*>(m)->matcher;
  const NodeData3D d = MakeNode(hi, n_hi, lo, n_lo, hist, n_hist, gravity4);
  const Result3D r = full ? matcher.MatchFullSubmap(Pose7(node_pose7).q, Pose7(submap_pose7).q, d, min_score)
                          : matcher.Match(Pose7(node_pose7), Pose7(submap_pose7), d, min_score);
  *out = oracle_result3d{};
  out->found = r.found ? 1 : 0;
  out->scan_index = r.scan_index;
  out->offset[0] = r.offset.x;
  out->offset[1] = r.offset.y;
  out->offset[2] = r.offset.z;
  out->num_scans = r.num_scans;
  out->score = r.score;
  out->rotational_score = r.rotational_score;
  out->low_resolution_score = r.low_resolution_score;
  const float p[7] = {r.pose_estimate.t.x, r.pose_estimate.t.y, r.pose_estimate.t.z, r.pose_estimate.q.w,
                      r.pose_estimate.q.x, r.pose_estimate.q.y, r.pose_estimate.q.z};
  for (int k = 0; k < 7; ++k) out->pose[k] = p[k];
  out->candidates_scored = r.candidates_scored;
}

// RealTimeCorrelativeScanMatcher3D::Match.  options4 = {linear_search_window,
// angular_search_window, translation_delta_cost_weight, rotation_delta_cost_weight}.
// Returns the score; pose7_out = winning pose, *best_index its generation index,
// *num_candidates the size of the search space.
float oracle_rt_match3d(void* grid, const double* options4, const float* initial_pose7, const float* pts,
                        int n, float* pose7_out, int* best_index, int* num_candidates) {
  RealTimeCorrelativeScanMatcherOptions o;
  o.linear_search_window = options4[0];
  o.angular_search_window = options4[1];
  o.translation_delta_cost_weight = options4[2];
  o.rotation_delta_cost_weight = options4[3];
  const RealTimeCorrelativeScanMatcher3D matcher(o);
  std::vector<Vec3f> cloud;
  for (int i = 0; i < n; ++i) cloud.push_back(Vec3f{pts[3 * i], pts[3 * i + 1], pts[3 * i + 2]});
  const HybridGrid& g = *static_cast<HybridGrid*>(grid);
  if (num_candidates)
    *num_candidates = static_cast<int>(matcher.GenerateExhaustiveSearchTransforms(g.resolution(), cloud).size());
  Rigid3f pose;
  const float score = matcher.Match(Pose7(initial_pose7), cloud, g, &pose, best_index);
  const float v[7] = {pose.t.x, pose.t.y, pose.t.z, pose.q.w, pose.q.x, pose.q.y, pose.q.z};
  for (int k = 0; k < 7; ++k) pose7_out[k] = v[k];
  return score;
}

// RotationalScanMatcher::Match (rotational_scan_matcher.cc:178-189).
void oracle_rotational_match(const float* submap_hist, const float* hist, int n, float initial_angle,
                             const float* angles, int n_angles, float* out) {
  const std::vector<float> scores =
      RotationalMatch(Histogram(submap_hist, submap_hist + n), Histogram(hist, hist + n), initial_angle,
                      std::vector<float>(angles, angles + n_angles));
  for (int i = 0; i < n_angles; ++i) out[i] = scores[i];
}

// Discrete scans of a local Match: returns the number of surviving scans; with non-NULL
// outputs fills poses (S x 7), rotational scores (S) and full-resolution cell indices
// (S x n_hi x 3).
int oracle_fcsm3d_discrete_scans(void* m, const double* options7, float resolution,
                                 const float* node_pose7, const float* submap_pose7, const float* hi,
                                 int n_hi, const float* hist, int n_hist, const double* gravity4,
                                 float* poses, float* rot_scores, int32_t* cells) {
  const FastCorrelativeScanMatcher3D& matcher = *static_cast<Matcher3DHolder*>(m)->matcher;
  const FastCorrelativeScanMatcherOptions3D o = Options3D(options7);
  const NodeData3D d = MakeNode(hi, n_hi, nullptr, 0, hist, n_hist, gravity4);
  const FastCorrelativeScanMatcher3D::SearchParameters sp{
      RoundToInt(o.linear_xy_search_window / resolution), RoundToInt(o.linear_z_search_window / resolution),
      o.angular_search_window};
  const auto scans = matcher.GenerateDiscreteScans(sp, d, Pose7(node_pose7), Pose7(submap_pose7));
  if (poses != nullptr) {
    for (size_t s = 0; s != scans.size(); ++s) {
      const Rigid3f& p = scans[s].pose;
      const float v[7] = {p.t.x, p.t.y, p.t.z, p.q.w, p.q.x, p.q.y, p.q.z};
      for (int k = 0; k < 7; ++k) poses[7 * s + k] = v[k];
      rot_scores[s] = scans[s].rotational_score;
      for (int i = 0; i < n_hi; ++i) {
        const Array3i& c = scans[s].cell_indices_per_depth[0][i];
        cells[(s * n_hi + i) * 3] = c.x;
        cells[(s * n_hi + i) * 3 + 1] = c.y;
        cells[(s * n_hi + i) * 3 + 2] = c.z;
      }
    }
  }
  return static_cast<int>(scans.size());
}

}  // extern "C"
