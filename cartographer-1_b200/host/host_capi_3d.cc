// Flat C entry points over the 3D host layer (FastCorrelativeScanMatcher3D), for
// ctypes-driven parity tests and the C5 bench.  Product code: reference-shaped
// classes + C ABI only; never touches oracle/.
#include <chrono>
#include <condition_variable>
#include <cstring>
#include <memory>
#include <mutex>

#include "cartographer/mapping/3d/range_data_inserter_3d.h"
#include "cartographer/mapping/internal/3d/scan_matching/fast_correlative_scan_matcher_3d.h"
#include "cartographer/mapping/internal/3d/scan_matching/real_time_correlative_scan_matcher_3d.h"
#include "cartographer/mapping/internal/constraints/constraint_builder_3d.h"

using namespace cartographer;
using mapping::scan_matching::DeviceContext;
using mapping::scan_matching::FastCorrelativeScanMatcher3D;

namespace {

struct Matcher3DHolder {
  std::unique_ptr<mapping::HybridGrid> grid, low_grid;
  Eigen::VectorXf histogram;
  std::unique_ptr<FastCorrelativeScanMatcher3D> matcher;
};

struct Node3DHolder {
  mapping::TrajectoryNode::Data data;
};

transform::Rigid3d Pose7(const double* p) {
  return transform::Rigid3d(transform::Vector3d(p[0], p[1], p[2]), transform::Quaterniond(p[3], p[4], p[5], p[6]));
}

struct Batch3DHolder {
  FastCorrelativeScanMatcher3D::QueryList queries;
  std::unique_ptr<FastCorrelativeScanMatcher3D::Batch> batch;
};

}  // namespace

extern "C" {

struct csmh_result3d {
  int32_t found;
  int32_t scan_index;
  int32_t offset[3];
  int32_t num_scans;
  float score, rotational_score, low_resolution_score;
  int32_t tie_resolution;
  double pose[7];  // tx, ty, tz, qw, qx, qy, qz
  uint64_t candidates_scored;
};

// Grids arrive as voxel lists (cell index, uint16 value), the form
// HybridGrid::Iterator yields.  options7: branch_and_bound_depth,
// full_resolution_depth, min_rotational_score, min_low_resolution_score,
// linear_xy_search_window, linear_z_search_window, angular_search_window.
void* csmh_fcsm3d_create(void* ctx, float resolution, const int32_t* xyz, const uint16_t* values, int n,
                         float low_resolution, const int32_t* low_xyz, const uint16_t* low_values,
                         int n_low, const float* histogram, int n_histogram, const double* options7) {
  Matcher3DHolder* h = new Matcher3DHolder;
  h->grid.reset(new mapping::HybridGrid(resolution));
  for (int i = 0; i < n; ++i)
    *h->grid->mutable_value(Eigen::Array3i(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2])) = values[i];
  h->low_grid.reset(new mapping::HybridGrid(low_resolution));
  for (int i = 0; i < n_low; ++i)
    *h->low_grid->mutable_value(Eigen::Array3i(low_xyz[3 * i], low_xyz[3 * i + 1], low_xyz[3 * i + 2])) =
        low_values[i];
  h->histogram.assign(histogram, histogram + n_histogram);
  mapping::scan_matching::proto::FastCorrelativeScanMatcherOptions3D o;
  o.set_branch_and_bound_depth(static_cast<int>(options7[0]));
  o.set_full_resolution_depth(static_cast<int>(options7[1]));
  o.set_min_rotational_score(options7[2]);
  o.set_min_low_resolution_score(options7[3]);
  o.set_linear_xy_search_window(options7[4]);
  o.set_linear_z_search_window(options7[5]);
  o.set_angular_search_window(options7[6]);
  h->matcher.reset(new FastCorrelativeScanMatcher3D(*h->grid, h->low_grid.get(), &h->histogram, o,
                                                    static_cast<DeviceContext*>(ctx)));
  return h;
}
void csmh_fcsm3d_destroy(void* m) { delete static_cast<Matcher3DHolder*>(m); }
int csmh_fcsm3d_grid_size(void* m) { return static_cast<Matcher3DHolder*>(m)->grid->grid_size(); }

int csmh_fcsm3d_level(void* ctx, void* m, int level, uint8_t* out, int32_t* lo3, int32_t* dims3) {
  return csm_stack3d_level(static_cast<DeviceContext*>(ctx)->get(),
                           static_cast<Matcher3DHolder*>(m)->matcher->stack(), level, out, lo3, dims3);
}

// TrajectoryNode::Data of one node.
void* csmh_node3d_create(const float* hi, int n_hi, const float* lo, int n_lo, const float* histogram,
                         int n_histogram, const double* gravity4) {
  Node3DHolder* h = new Node3DHolder;
  for (int i = 0; i < n_hi; ++i)
    h->data.high_resolution_point_cloud.push_back({sensor::Vector3f(hi[3 * i], hi[3 * i + 1], hi[3 * i + 2])});
  for (int i = 0; i < n_lo; ++i)
    h->data.low_resolution_point_cloud.push_back({sensor::Vector3f(lo[3 * i], lo[3 * i + 1], lo[3 * i + 2])});
  h->data.rotational_scan_matcher_histogram.assign(histogram, histogram + n_histogram);
  h->data.gravity_alignment = transform::Quaterniond(gravity4[0], gravity4[1], gravity4[2], gravity4[3]);
  return h;
}
void csmh_node3d_destroy(void* n) { delete static_cast<Node3DHolder*>(n); }

// A prepared batch: search i runs nodes[i] against matchers[i]; full[i] selects
// MatchFullSubmap; node_poses / submap_poses are n x 7 doubles.
void* csmh_batch3d_create(void* ctx, int n, void* const* matchers, void* const* nodes, const int32_t* full,
                          const double* node_poses, const double* submap_poses, const float* min_scores) {
  Batch3DHolder* h = new Batch3DHolder;
  h->queries.reserve(n);
  for (int i = 0; i < n; ++i) {
    FastCorrelativeScanMatcher3D::Query q;
    q.constant_data = &static_cast<Node3DHolder*>(nodes[i])->data;
    q.full_submap = full[i] != 0;
    q.global_node_pose = Pose7(node_poses + 7 * i);
    q.global_submap_pose = Pose7(submap_poses + 7 * i);
    q.min_score = min_scores[i];
    h->queries.emplace_back(static_cast<Matcher3DHolder*>(matchers[i])->matcher.get(), q);
  }
  h->batch.reset(new FastCorrelativeScanMatcher3D::Batch(static_cast<DeviceContext*>(ctx), h->queries));
  return h;
}
void csmh_batch3d_run(void* b) { static_cast<Batch3DHolder*>(b)->batch->Run(); }
void csmh_batch3d_fetch(void* b, csmh_result3d* out) {
  const std::vector<FastCorrelativeScanMatcher3D::Answer> answers = static_cast<Batch3DHolder*>(b)->batch->Fetch();
  for (size_t i = 0; i != answers.size(); ++i) {
    const FastCorrelativeScanMatcher3D::Answer& a = answers[i];
    csmh_result3d& r = out[i];
    std::memset(&r, 0, sizeof(r));
    r.found = a.found ? 1 : 0;
    r.scan_index = a.scan_index;
    for (int k = 0; k < 3; ++k) r.offset[k] = a.offset[k];
    r.num_scans = a.num_scans;
    r.score = a.result.score;
    r.rotational_score = a.result.rotational_score;
    r.low_resolution_score = a.result.low_resolution_score;
    r.tie_resolution = a.tie_resolution;
    const transform::Rigid3d& p = a.result.pose_estimate;
    const double v[7] = {p.translation().x(), p.translation().y(), p.translation().z(), p.rotation().w(),
                         p.rotation().x(),    p.rotation().y(),    p.rotation().z()};
    for (int k = 0; k < 7; ++k) r.pose[k] = v[k];
    r.candidates_scored = a.candidates_scored;
  }
}
void csmh_batch3d_destroy(void* b) { delete static_cast<Batch3DHolder*>(b); }

// The reference's single-search entry points (Match / MatchFullSubmap returning a
// unique_ptr<Result>): returns 1 and fills `out` iff a Result was returned.
int csmh_fcsm3d_match(void* m, void* node, int full, const double* node_pose7, const double* submap_pose7,
                      float min_score, csmh_result3d* out) {
  const FastCorrelativeScanMatcher3D& matcher = *static_cast<Matcher3DHolder*>(m)->matcher;
  const mapping::TrajectoryNode::Data& data = static_cast<Node3DHolder*>(node)->data;
  const transform::Rigid3d np = Pose7(node_pose7), sp = Pose7(submap_pose7);
  const std::unique_ptr<FastCorrelativeScanMatcher3D::Result> result =
      full ? matcher.MatchFullSubmap(np.rotation(), sp.rotation(), data, min_score)
           : matcher.Match(np, sp, data, min_score);
  std::memset(out, 0, sizeof(*out));
  if (result == nullptr) return 0;
  out->found = 1;
  out->score = result->score;
  out->rotational_score = result->rotational_score;
  out->low_resolution_score = result->low_resolution_score;
  const transform::Rigid3d& p = result->pose_estimate;
  const double v[7] = {p.translation().x(), p.translation().y(), p.translation().z(), p.rotation().w(),
                       p.rotation().x(),    p.rotation().y(),    p.rotation().z()};
  for (int k = 0; k < 7; ++k) out->pose[k] = v[k];
  return 1;
}

// RotationalScanMatcher::Match alone (host code; needs no device).
void csmh_rotational_match(const float* submap_hist, const float* hist, int n, float initial_angle,
                           const float* angles, int n_angles, float* out) {
  const Eigen::VectorXf submap(submap_hist, submap_hist + n);
  const mapping::scan_matching::RotationalScanMatcher matcher(&submap);
  const std::vector<float> scores = matcher.Match(Eigen::VectorXf(hist, hist + n), initial_angle,
                                                  std::vector<float>(angles, angles + n_angles));
  for (int i = 0; i < n_angles; ++i) out[i] = scores[i];
}

// Discrete scans of a local Match: kept yaw slices' poses (S x 7), rotational scores
// (S) and device-computed full-resolution cell indices (S x n_hi x 3).  Returns S;
// outputs may be NULL to query S.
int csmh_fcsm3d_discrete_scans(void* ctx, void* m, void* node, const double* node_pose7,
                               const double* submap_pose7, float* poses, float* rotational_scores,
                               int32_t* cells) {
  const Matcher3DHolder* h = static_cast<Matcher3DHolder*>(m);
  const mapping::TrajectoryNode::Data& data = static_cast<Node3DHolder*>(node)->data;
  std::vector<float> scores;
  const std::vector<float> p =
      h->matcher->DiscreteScanPosesForTesting(Pose7(node_pose7), Pose7(submap_pose7), data, &scores);
  const int S = static_cast<int>(scores.size());
  if (poses == nullptr || S == 0) return S;
  std::memcpy(poses, p.data(), sizeof(float) * p.size());
  std::memcpy(rotational_scores, scores.data(), sizeof(float) * scores.size());
  std::vector<float> pts;
  for (const sensor::RangefinderPoint& point : data.high_resolution_point_cloud)
    for (const float v : {point.position.x(), point.position.y(), point.position.z()}) pts.push_back(v);
  std::vector<float> leaf(4 * S, 0.f);
  const float one_low_point[3] = {0.f, 0.f, 0.f};
  csm_job3d job;
  std::memset(&job, 0, sizeof(job));
  job.stack = h->matcher->stack();
  job.points_xyz = pts.data();
  job.num_points = static_cast<int32_t>(data.high_resolution_point_cloud.size());
  job.low_points_xyz = one_low_point;
  job.num_low_points = 1;
  job.num_scans = S;
  job.scan_poses = p.data();
  job.scan_leaf_quat = leaf.data();
  job.resolution = h->grid->resolution();
  job.low_resolution = h->low_grid->resolution();
  job.linear_xy_window = 0;
  job.linear_z_window = 0;
  job.low_resolution_grid = h->matcher->low_resolution_grid();
  return csm_discretize3d(static_cast<DeviceContext*>(ctx)->get(), &job, cells) == CSM_OK ? S : -1;
}

// ConstraintBuilder3D driven like PoseGraph3D does (one NotifyEndOfNode per search) on
// `num_devices` GPUs.  Submaps are given as matcher holders (their grids and
// histogram are borrowed; the builder constructs its own matchers).  options6 =
// {sampling_ratio, max_constraint_distance, min_score, global_localization_min_score,
// loop_closure_translation_weight, loop_closure_rotation_weight}; fcsm_options7 as for
// csmh_fcsm3d_create.  Per search i: found[i] and the constraint transform
// submap <- node as 7 doubles.  Returns wall seconds from the first MaybeAdd* call to
// the WhenDone callback.
double csmh_constraint_builder3d_run2(int num_devices, const int* devices, int num_pool_threads,
                                      int flush_threshold, const double* options6,
                                      const double* fcsm_options7, int num_submaps, void* const* submap_holders,
                                      int num_searches, const int32_t* search_submap,
                                      const int32_t* search_full, void* const* nodes,
                                      const double* node_poses, const double* submap_poses, int32_t* found,
                                      double* transform7, double* stats4, int burst);
double csmh_constraint_builder3d_run(int num_devices, const int* devices, int num_pool_threads,
                                     int flush_threshold, const double* options6,
                                     const double* fcsm_options7, int num_submaps, void* const* submap_holders,
                                     int num_searches, const int32_t* search_submap,
                                     const int32_t* search_full, void* const* nodes,
                                     const double* node_poses, const double* submap_poses, int32_t* found,
                                     double* transform7, double* stats4) {
  return csmh_constraint_builder3d_run2(num_devices, devices, num_pool_threads, flush_threshold, options6,
                                        fcsm_options7, num_submaps, submap_holders, num_searches, search_submap,
                                        search_full, nodes, node_poses, submap_poses, found, transform7, stats4, 0);
}
// burst > 0: the searches enter through MaybeAddConstraints in bursts of that many requests
// (whole-node batches), followed by the burst's NotifyEndOfNode calls.
double csmh_constraint_builder3d_run2(int num_devices, const int* devices, int num_pool_threads,
                                      int flush_threshold, const double* options6,
                                      const double* fcsm_options7, int num_submaps, void* const* submap_holders,
                                      int num_searches, const int32_t* search_submap,
                                      const int32_t* search_full, void* const* nodes,
                                      const double* node_poses, const double* submap_poses, int32_t* found,
                                      double* transform7, double* stats4, int burst) {
  using mapping::constraints::ConstraintBuilder3D;
  mapping::constraints::proto::ConstraintBuilderOptions o;
  o.set_sampling_ratio(options6[0]);
  o.set_max_constraint_distance(options6[1]);
  o.set_min_score(options6[2]);
  o.set_global_localization_min_score(options6[3]);
  o.set_loop_closure_translation_weight(options6[4]);
  o.set_loop_closure_rotation_weight(options6[5]);
  auto* f = o.mutable_fast_correlative_scan_matcher_options_3d();
  f->set_branch_and_bound_depth(static_cast<int>(fcsm_options7[0]));
  f->set_full_resolution_depth(static_cast<int>(fcsm_options7[1]));
  f->set_min_rotational_score(fcsm_options7[2]);
  f->set_min_low_resolution_score(fcsm_options7[3]);
  f->set_linear_xy_search_window(fcsm_options7[4]);
  f->set_linear_z_search_window(fcsm_options7[5]);
  f->set_angular_search_window(fcsm_options7[6]);
  std::vector<mapping::Submap3D> submaps;
  for (int i = 0; i < num_submaps; ++i) {
    const Matcher3DHolder* h = static_cast<Matcher3DHolder*>(submap_holders[i]);
    submaps.emplace_back(transform::Rigid3d::Identity(), h->grid.get(), h->low_grid.get(), &h->histogram);
  }
  for (int i = 0; i < num_searches; ++i) found[i] = 0;
  double seconds = 0.;
  {
    common::ThreadPool pool(num_pool_threads);
    ConstraintBuilder3D builder(o, &pool, std::vector<int>(devices, devices + num_devices));
    builder.set_flush_threshold(flush_threshold);
    std::mutex m;
    std::condition_variable cv;
    bool done = false;
    ConstraintBuilder3D::Result result;
    const auto t0 = std::chrono::steady_clock::now();
    for (int begin = 0; burst > 0 && begin < num_searches; begin += burst) {
      const int end = std::min(num_searches, begin + burst);
      std::vector<ConstraintBuilder3D::ConstraintRequest> requests;
      for (int i = begin; i < end; ++i) {
        ConstraintBuilder3D::ConstraintRequest r;
        r.submap_id = mapping::SubmapId{0, search_submap[i]};
        r.submap = &submaps[search_submap[i]];
        r.node_id = mapping::NodeId{0, i};
        r.constant_data = &static_cast<Node3DHolder*>(nodes[i])->data;
        r.match_full_submap = search_full[i] != 0;
        r.global_node_pose = Pose7(node_poses + 7 * i);
        r.global_submap_pose = Pose7(submap_poses + 7 * i);
        requests.push_back(r);
      }
      builder.MaybeAddConstraints(requests);
      for (int i = begin; i < end; ++i) builder.NotifyEndOfNode();
    }
    for (int i = 0; burst <= 0 && i < num_searches; ++i) {
      const int s = search_submap[i];
      const mapping::TrajectoryNode::Data* data = &static_cast<Node3DHolder*>(nodes[i])->data;
      const transform::Rigid3d np = Pose7(node_poses + 7 * i), sp = Pose7(submap_poses + 7 * i);
      if (search_full[i]) {
        builder.MaybeAddGlobalConstraint(mapping::SubmapId{0, s}, &submaps[s], mapping::NodeId{0, i}, data,
                                         np.rotation(), sp.rotation());
      } else {
        builder.MaybeAddConstraint(mapping::SubmapId{0, s}, &submaps[s], mapping::NodeId{0, i}, data, np, sp);
      }
      builder.NotifyEndOfNode();
    }
    builder.WhenDone([&](const ConstraintBuilder3D::Result& r) {
      std::lock_guard<std::mutex> l(m);
      result = r;
      done = true;
      cv.notify_all();
    });
    {
      std::unique_lock<std::mutex> l(m);
      cv.wait(l, [&] { return done; });
    }
    seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    int last = -1;
    for (const auto& c : result) {
      const int i = c.node_id.node_index;
      CSM_CHECK(i > last);  // submission order
      last = i;
      found[i] = 1;
      const transform::Rigid3d& p = c.pose.zbar_ij;
      const double v[7] = {p.translation().x(), p.translation().y(), p.translation().z(), p.rotation().w(),
                           p.rotation().x(),    p.rotation().y(),    p.rotation().z()};
      for (int k = 0; k < 7; ++k) transform7[7 * i + k] = v[k];
    }
    const auto st = builder.stats();
    if (stats4) {
      stats4[0] = static_cast<double>(st.searches);
      stats4[1] = static_cast<double>(st.found);
      stats4[2] = static_cast<double>(st.candidates_scored);
      stats4[3] = static_cast<double>(st.flushes);
    }
    CSM_CHECK(builder.GetNumFinishedNodes() == num_searches);
    for (int s = 0; s < num_submaps; ++s) builder.DeleteScanMatcher(mapping::SubmapId{0, s});
  }
  return seconds;
}

// RealTimeCorrelativeScanMatcher3D::Match through the reference signature: the grid is
// given as a voxel list.  options4 = {linear_search_window, angular_search_window,
// translation_delta_cost_weight, rotation_delta_cost_weight}.  Returns the score.
float csmh_rt_match3d(float resolution, const int32_t* xyz, const uint16_t* values, int n_voxels,
                      const double* options4, const double* initial_pose7, const float* pts, int n,
                      double* pose7_out) {
  mapping::HybridGrid grid(resolution);
  for (int i = 0; i < n_voxels; ++i)
    *grid.mutable_value(Eigen::Array3i(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2])) = values[i];
  mapping::scan_matching::proto::RealTimeCorrelativeScanMatcherOptions o;
  o.set_linear_search_window(options4[0]);
  o.set_angular_search_window(options4[1]);
  o.set_translation_delta_cost_weight(options4[2]);
  o.set_rotation_delta_cost_weight(options4[3]);
  const mapping::scan_matching::RealTimeCorrelativeScanMatcher3D matcher(o);
  sensor::PointCloud cloud;
  for (int i = 0; i < n; ++i) cloud.push_back({sensor::Vector3f(pts[3 * i], pts[3 * i + 1], pts[3 * i + 2])});
  transform::Rigid3d pose;
  const float score = matcher.Match(Pose7(initial_pose7), cloud, grid, &pose);
  const double v[7] = {pose.translation().x(), pose.translation().y(), pose.translation().z(),
                       pose.rotation().w(),    pose.rotation().x(),    pose.rotation().y(),
                       pose.rotation().z()};
  for (int k = 0; k < 7; ++k) pose7_out[k] = v[k];
  return score;
}

// The same for a batch of clouds against one grid state resident on the device.
int csmh_rt_match3d_batch(void* ctx, float resolution, const int32_t* xyz, const uint16_t* values,
                          int n_voxels, const double* options4, int num, const double* initial_poses7,
                          const float* pts, const int64_t* offsets, float* scores, double* poses7_out,
                          int32_t* best_index, int32_t* num_candidates) {
  DeviceContext* context = static_cast<DeviceContext*>(ctx);
  mapping::HybridGrid grid(resolution);
  for (int i = 0; i < n_voxels; ++i)
    *grid.mutable_value(Eigen::Array3i(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2])) = values[i];
  const mapping::scan_matching::DeviceHybridGrid device_grid(context, grid);
  mapping::scan_matching::proto::RealTimeCorrelativeScanMatcherOptions o;
  o.set_linear_search_window(options4[0]);
  o.set_angular_search_window(options4[1]);
  o.set_translation_delta_cost_weight(options4[2]);
  o.set_rotation_delta_cost_weight(options4[3]);
  const mapping::scan_matching::RealTimeCorrelativeScanMatcher3D matcher(o);
  std::vector<sensor::PointCloud> clouds(num);
  std::vector<mapping::scan_matching::RealTimeCorrelativeScanMatcher3D::Query> queries(num);
  for (int i = 0; i < num; ++i) {
    for (int64_t p = offsets[i]; p < offsets[i + 1]; ++p)
      clouds[i].push_back({sensor::Vector3f(pts[3 * p], pts[3 * p + 1], pts[3 * p + 2])});
    queries[i].initial_pose_estimate = Pose7(initial_poses7 + 7 * i);
    queries[i].point_cloud = &clouds[i];
    queries[i].grid = &device_grid;
  }
  const auto answers = matcher.MatchBatch(context, queries);
  for (int i = 0; i < num; ++i) {
    scores[i] = answers[i].score;
    best_index[i] = answers[i].best_index;
    num_candidates[i] = answers[i].num_candidates;
    const transform::Rigid3d& p = answers[i].pose_estimate;
    const double v[7] = {p.translation().x(), p.translation().y(), p.translation().z(), p.rotation().w(),
                         p.rotation().x(),    p.rotation().y(),    p.rotation().z()};
    for (int k = 0; k < 7; ++k) poses7_out[7 * i + k] = v[k];
  }
  return num;
}

// A HybridGrid kept on the device between calls (local SLAM matches every new scan against
// the same live submap): create once, match many clouds, destroy.
void* csmh_device_grid3d_create(void* ctx, float resolution, const int32_t* xyz, const uint16_t* values,
                                int n_voxels) {
  mapping::HybridGrid grid(resolution);
  for (int i = 0; i < n_voxels; ++i)
    *grid.mutable_value(Eigen::Array3i(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2])) = values[i];
  return new mapping::scan_matching::DeviceHybridGrid(static_cast<DeviceContext*>(ctx), grid);
}
void csmh_device_grid3d_destroy(void* g) { delete static_cast<mapping::scan_matching::DeviceHybridGrid*>(g); }

int csmh_rt_match3d_resident(void* ctx, void* device_grid, const double* options4, int num,
                             const double* initial_poses7, const float* pts, const int64_t* offsets,
                             float* scores, double* poses7_out, int32_t* best_index,
                             int32_t* num_candidates) {
  DeviceContext* context = static_cast<DeviceContext*>(ctx);
  mapping::scan_matching::proto::RealTimeCorrelativeScanMatcherOptions o;
  o.set_linear_search_window(options4[0]);
  o.set_angular_search_window(options4[1]);
  o.set_translation_delta_cost_weight(options4[2]);
  o.set_rotation_delta_cost_weight(options4[3]);
  const mapping::scan_matching::RealTimeCorrelativeScanMatcher3D matcher(o);
  std::vector<sensor::PointCloud> clouds(num);
  std::vector<mapping::scan_matching::RealTimeCorrelativeScanMatcher3D::Query> queries(num);
  for (int i = 0; i < num; ++i) {
    for (int64_t p = offsets[i]; p < offsets[i + 1]; ++p)
      clouds[i].push_back({sensor::Vector3f(pts[3 * p], pts[3 * p + 1], pts[3 * p + 2])});
    queries[i].initial_pose_estimate = Pose7(initial_poses7 + 7 * i);
    queries[i].point_cloud = &clouds[i];
    queries[i].grid = static_cast<mapping::scan_matching::DeviceHybridGrid*>(device_grid);
  }
  const auto answers = matcher.MatchBatch(context, queries);
  for (int i = 0; i < num; ++i) {
    scores[i] = answers[i].score;
    best_index[i] = answers[i].best_index;
    num_candidates[i] = answers[i].num_candidates;
    const transform::Rigid3d& p = answers[i].pose_estimate;
    const double v[7] = {p.translation().x(), p.translation().y(), p.translation().z(), p.rotation().w(),
                         p.rotation().x(),    p.rotation().y(),    p.rotation().z()};
    for (int k = 0; k < 7; ++k) poses7_out[7 * i + k] = v[k];
  }
  return num;
}


// RangeDataInserter3D.  scans: origins[3 * num], returns given by offsets (num + 1 entries, in
// points) into returns_xyz.  options3 = {hit_probability, miss_probability, num_free_space_voxels}.
// resident == 0: one reference-signature Insert(range_data, HybridGrid*) per scan (grid up and down
// every call); resident != 0: one DeviceGrid, every scan inserted, exported once.  The grid's
// non-default voxels come back through csmh_grid3d_result_* (the handle owns them).
struct Grid3DResult {
  std::vector<int32_t> xyz;
  std::vector<uint16> values;
  int grid_size = 0;
};
void* csmh_grid3d_insert(void* ctx, float resolution, const int32_t* xyz, const uint16_t* values, int n_voxels,
                         const double* options3, int num, const float* origins, const float* returns_xyz,
                         const int64_t* offsets, int resident) {
  DeviceContext* context = static_cast<DeviceContext*>(ctx);
  mapping::HybridGrid grid(resolution);
  for (int i = 0; i < n_voxels; ++i)
    *grid.mutable_value(Eigen::Array3i(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2])) = values[i];
  mapping::proto::RangeDataInserterOptions3D o;
  o.set_hit_probability(options3[0]);
  o.set_miss_probability(options3[1]);
  o.set_num_free_space_voxels(static_cast<int>(options3[2]));
  const mapping::RangeDataInserter3D inserter(o, context);
  std::vector<sensor::RangeData> scans(num);
  for (int s = 0; s < num; ++s) {
    scans[s].origin = sensor::Vector3f(origins[3 * s], origins[3 * s + 1], origins[3 * s + 2]);
    for (int64_t i = offsets[s]; i < offsets[s + 1]; ++i)
      scans[s].returns.push_back({sensor::Vector3f(returns_xyz[3 * i], returns_xyz[3 * i + 1], returns_xyz[3 * i + 2])});
  }
  Grid3DResult* out = new Grid3DResult;
  if (resident) {
    mapping::RangeDataInserter3D::DeviceGrid device_grid(context, grid);
    std::vector<const sensor::RangeData*> list;
    for (const sensor::RangeData& rd : scans) list.push_back(&rd);
    inserter.Insert(list, &device_grid);
    device_grid.Download(&grid);
  } else {
    for (const sensor::RangeData& rd : scans) inserter.Insert(rd, &grid);
  }
  grid.Export(&out->xyz, &out->values);
  out->grid_size = grid.grid_size();
  return out;
}
int csmh_grid3d_result_size(void* r) { return static_cast<int>(static_cast<Grid3DResult*>(r)->values.size()); }
int csmh_grid3d_result_grid_size(void* r) { return static_cast<Grid3DResult*>(r)->grid_size; }
void csmh_grid3d_result_copy(void* r, int32_t* xyz, uint16_t* values) {
  const Grid3DResult* g = static_cast<Grid3DResult*>(r);
  std::memcpy(xyz, g->xyz.data(), sizeof(int32_t) * g->xyz.size());
  std::memcpy(values, g->values.data(), sizeof(uint16_t) * g->values.size());
}
void csmh_grid3d_result_destroy(void* r) { delete static_cast<Grid3DResult*>(r); }


// The resident 3D pipeline end to end: voxels -> DeviceGrid, every scan inserted, then the matcher's
// stack built (a) straight from the device grid (csm_stack3d_build_from_hgrid) and (b) from the
// exported voxel list (csm_stack3d_build).  Returns the number of level bytes that differ between
// the two (0 = identical boxes and contents), -1 on differing boxes.
int64_t csmh_grid3d_resident_stack_check(void* ctx, float resolution, const int32_t* xyz, const uint16_t* values,
                                         int n_voxels, const double* options3, int num, const float* origins,
                                         const float* returns_xyz, const int64_t* offsets, int depth,
                                         int full_resolution_depth) {
  DeviceContext* context = static_cast<DeviceContext*>(ctx);
  mapping::HybridGrid grid(resolution);
  for (int i = 0; i < n_voxels; ++i)
    *grid.mutable_value(Eigen::Array3i(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2])) = values[i];
  mapping::proto::RangeDataInserterOptions3D o;
  o.set_hit_probability(options3[0]);
  o.set_miss_probability(options3[1]);
  o.set_num_free_space_voxels(static_cast<int>(options3[2]));
  const mapping::RangeDataInserter3D inserter(o, context);
  std::vector<sensor::RangeData> scans(num);
  std::vector<const sensor::RangeData*> list;
  for (int s = 0; s < num; ++s) {
    scans[s].origin = sensor::Vector3f(origins[3 * s], origins[3 * s + 1], origins[3 * s + 2]);
    for (int64_t i = offsets[s]; i < offsets[s + 1]; ++i)
      scans[s].returns.push_back({sensor::Vector3f(returns_xyz[3 * i], returns_xyz[3 * i + 1], returns_xyz[3 * i + 2])});
    list.push_back(&scans[s]);
  }
  mapping::RangeDataInserter3D::DeviceGrid device_grid(context, grid);
  inserter.Insert(list, &device_grid);
  const std::vector<uint8_t> q = mapping::scan_matching::ComputeQuantizationTable3D();
  csm_stack3d *direct = nullptr, *listed = nullptr;
  if (csm_stack3d_build_from_hgrid(context->get(), device_grid.get(), q.data(), depth, full_resolution_depth,
                                   &direct) != CSM_OK)
    return -2;
  std::vector<int32_t> vx;
  std::vector<uint16_t> vv;
  device_grid.Export(&vx, &vv);
  if (csm_stack3d_build(context->get(), vx.data(), vv.data(), static_cast<int>(vv.size()), q.data(), depth,
                        full_resolution_depth, &listed) != CSM_OK)
    return -3;
  int64_t differing = 0;
  for (int level = 0; level < depth && differing >= 0; ++level) {
    int32_t lo_a[3], dim_a[3], lo_b[3], dim_b[3];
    csm_stack3d_level(context->get(), direct, level, nullptr, lo_a, dim_a);
    csm_stack3d_level(context->get(), listed, level, nullptr, lo_b, dim_b);
    if (std::memcmp(lo_a, lo_b, sizeof(lo_a)) != 0 || std::memcmp(dim_a, dim_b, sizeof(dim_a)) != 0) {
      differing = -1;
      break;
    }
    const size_t bytes = static_cast<size_t>(dim_a[0]) * dim_a[1] * dim_a[2];
    std::vector<uint8_t> a(bytes), b(bytes);
    csm_stack3d_level(context->get(), direct, level, a.data(), lo_a, dim_a);
    csm_stack3d_level(context->get(), listed, level, b.data(), lo_b, dim_b);
    for (size_t i = 0; i != bytes; ++i) differing += a[i] != b[i];
  }
  csm_stack3d_destroy(context->get(), direct);
  csm_stack3d_destroy(context->get(), listed);
  return differing;
}

}  // extern "C"
