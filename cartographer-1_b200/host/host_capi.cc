// Flat C entry points over the C++ host layer, for ctypes-driven parity tests and
// bench.py.  Product code: goes through the reference-shaped classes and the C ABI
// only; never touches oracle/.
#include <chrono>
#include <cmath>
#include <condition_variable>
#include <cstring>
#include <memory>
#include <mutex>

#include "cartographer/mapping/internal/2d/tsdf_range_data_inserter_2d.h"
#include "cartographer/io/grid_proto_reader.h"
#include "cartographer/mapping/internal/2d/scan_matching/fast_correlative_scan_matcher_2d.h"
#include "cartographer/mapping/2d/probability_grid_range_data_inserter_2d.h"
#include "cartographer/mapping/internal/2d/scan_matching/ceres_scan_matcher_2d.h"
#include "cartographer/mapping/internal/2d/scan_matching/real_time_correlative_scan_matcher_2d.h"
#include "cartographer/mapping/internal/constraints/constraint_builder_2d.h"

#include "host_capi_holders.h"
#include "parallel_for.h"

using namespace cartographer;
using mapping::scan_matching::DeviceContext;
using mapping::scan_matching::FastCorrelativeScanMatcher2D;

extern "C" {

struct csmh_result2d {
  int32_t found;
  int32_t scan_index, x_index_offset, y_index_offset;
  float score;
  int32_t tie_resolution;
  double pose_x, pose_y, pose_theta;
  uint64_t candidates_scored;
};

void* csmh_context_create(int device) {
  if (csm_device_count() <= device) return nullptr;
  return new DeviceContext(device);
}
void csmh_context_destroy(void* c) { delete static_cast<DeviceContext*>(c); }
void* csmh_context_raw(void* c) { return static_cast<DeviceContext*>(c)->get(); }


void* csmh_grid_create(const uint16_t* cells, int nx, int ny, double res, double max_x,
                       double max_y) {
  GridHolder* h = new GridHolder;
  h->grid.reset(new mapping::ProbabilityGrid(
      mapping::MapLimits(res, transform::Vector2d(max_x, max_y), mapping::CellLimits(nx, ny)),
      &h->tables));
  std::memcpy(h->grid->mutable_correspondence_cost_cells()->data(), cells,
              sizeof(uint16_t) * static_cast<size_t>(nx) * ny);
  return h;
}
// A TSDF2D from its two cell arrays (tsdf_2d.h): only the real-time matcher takes it.
void* csmh_tsdf_create(const uint16_t* tsd_cells, const uint16_t* weight_cells, int nx, int ny,
                       double res, double max_x, double max_y, float truncation_distance,
                       float max_weight) {
  GridHolder* h = new GridHolder;
  mapping::TSDF2D* tsdf = new mapping::TSDF2D(
      mapping::MapLimits(res, transform::Vector2d(max_x, max_y), mapping::CellLimits(nx, ny)),
      truncation_distance, max_weight, &h->tables);
  h->grid.reset(tsdf);
  const size_t bytes = sizeof(uint16_t) * static_cast<size_t>(nx) * ny;
  std::memcpy(tsdf->mutable_correspondence_cost_cells()->data(), tsd_cells, bytes);
  std::memcpy(tsdf->mutable_weight_cells()->data(), weight_cells, bytes);
  return h;
}
void csmh_grid_destroy(void* g) { delete static_cast<GridHolder*>(g); }
// Page-locks the grid's cells where they are (csm_host_register), so that stack builds copy them with
// an asynchronous transfer; released when the handle is destroyed.  The grid must not grow afterwards.
int csmh_grid_pin(void* g) {
  GridHolder* h = static_cast<GridHolder*>(g);
  if (h->locked != nullptr) return 0;
  const std::vector<uint16_t>& cells = h->grid->correspondence_cost_cells();
  if (cells.empty()) return 0;
  const int rc = csm_host_register(const_cast<uint16_t*>(cells.data()), sizeof(uint16_t) * cells.size());
  if (rc == 0) h->locked = cells.data();
  return rc;
}

// TSDFRangeDataInserter2D on a TSDF grid handle (csmh_tsdf_create / csmh_grid_from_proto); the grid
// may grow.  options9 = {truncation_distance, maximum_weight, update_free_space, num_normal_samples,
// sample_radius, project_sdf_distance_to_scan_normal, update_weight_range_exponent, angle kernel
// bandwidth, distance kernel bandwidth}.  resident == 0: one reference-signature Insert per scan;
// != 0: grown for every scan first, then one DeviceGrid, every scan inserted, downloaded once.
int csmh_tsdf_insert(void* ctx, void* grid, const double* options9, int num, const float* origins_xy,
                     const float* returns_xyz, const int64_t* offsets, int resident) {
  DeviceContext* context = static_cast<DeviceContext*>(ctx);
  mapping::Grid2D* g = static_cast<GridHolder*>(grid)->grid.get();
  if (g->GetGridType() != mapping::GridType::TSDF) return -1;
  mapping::TSDF2D* tsdf = static_cast<mapping::TSDF2D*>(g);
  mapping::proto::TSDFRangeDataInserterOptions2D o;
  o.set_truncation_distance(options9[0]);
  o.set_maximum_weight(options9[1]);
  o.set_update_free_space(options9[2] != 0.);
  o.mutable_normal_estimation_options()->set_num_normal_samples(static_cast<int>(options9[3]));
  o.mutable_normal_estimation_options()->set_sample_radius(static_cast<float>(options9[4]));
  o.set_project_sdf_distance_to_scan_normal(options9[5] != 0.);
  o.set_update_weight_range_exponent(static_cast<int>(options9[6]));
  o.set_update_weight_angle_scan_normal_to_ray_kernel_bandwidth(options9[7]);
  o.set_update_weight_distance_cell_to_hit_kernel_bandwidth(options9[8]);
  const mapping::TSDFRangeDataInserter2D inserter(o, context);
  std::vector<sensor::RangeData> scans(num);
  for (int s = 0; s < num; ++s) {
    scans[s].origin = sensor::Vector3f(origins_xy[2 * s], origins_xy[2 * s + 1], 0.f);
    for (int64_t i = offsets[s]; i < offsets[s + 1]; ++i)
      scans[s].returns.push_back({sensor::Vector3f(returns_xyz[3 * i], returns_xyz[3 * i + 1], returns_xyz[3 * i + 2])});
  }
  if (resident) {
    std::vector<const sensor::RangeData*> list;
    for (const sensor::RangeData& rd : scans) {
      mapping::TSDFRangeDataInserter2D::GrowAsNeeded(rd, static_cast<float>(o.truncation_distance()), tsdf);
      list.push_back(&rd);
    }
    mapping::TSDFRangeDataInserter2D::DeviceGrid device_grid(context, *tsdf);
    inserter.Insert(list, *tsdf, &device_grid);
    device_grid.Download(tsdf);
  } else {
    for (const sensor::RangeData& rd : scans) inserter.Insert(rd, tsdf);
  }
  return 0;
}

// Grid ingestion: a serialized proto::Grid2D message -> a grid handle (ProbabilityGrid or TSDF2D,
// whichever the message holds) every matcher entry point takes.  info6 = {num_x, num_y,
// resolution, max_x, max_y, is_tsdf}.
void* csmh_grid_from_proto(const uint8_t* bytes, int64_t size, double* info6) {
  GridHolder* h = new GridHolder;
  h->grid = io::ReadGrid2DProto(bytes, static_cast<size_t>(size), &h->tables);
  const mapping::MapLimits& l = h->grid->limits();
  info6[0] = l.cell_limits().num_x_cells;
  info6[1] = l.cell_limits().num_y_cells;
  info6[2] = l.resolution();
  info6[3] = l.max().x();
  info6[4] = l.max().y();
  info6[5] = h->grid->GetGridType() == mapping::GridType::TSDF ? 1. : 0.;
  return h;
}
void csmh_grid_weight_cells(void* grid, uint16_t* out) {
  const mapping::TSDF2D& t = static_cast<const mapping::TSDF2D&>(*static_cast<GridHolder*>(grid)->grid);
  std::memcpy(out, t.weight_cells().data(), sizeof(uint16_t) * t.weight_cells().size());
}
// A serialized proto::HybridGrid -> its non-default voxels (xyz / values may be NULL to count);
// *resolution and *grid_size are always filled.
int csmh_hybrid_grid_from_proto(const uint8_t* bytes, int64_t size, float* resolution, int* grid_size,
                                int32_t* xyz, uint16_t* values) {
  const std::unique_ptr<mapping::HybridGrid> grid = io::ReadHybridGridProto(bytes, static_cast<size_t>(size));
  *resolution = grid->resolution();
  *grid_size = grid->grid_size();
  std::vector<int32_t> cells;
  std::vector<uint16_t> vals;
  grid->Export(&cells, &vals);
  if (xyz != nullptr) {
    std::memcpy(xyz, cells.data(), sizeof(int32_t) * cells.size());
    std::memcpy(values, vals.data(), sizeof(uint16_t) * vals.size());
  }
  return static_cast<int>(vals.size());
}

void csmh_quantization_table(void* g, uint8_t* out65536) {
  const std::vector<uint8_t> t =
      mapping::scan_matching::ComputeQuantizationTable(*static_cast<GridHolder*>(g)->grid);
  std::memcpy(out65536, t.data(), 65536);
}

void* csmh_matcher_create(void* ctx, void* grid, double linear_window, double angular_window,
                          int depth) {
  mapping::scan_matching::proto::FastCorrelativeScanMatcherOptions2D o;
  o.set_linear_search_window(linear_window);
  o.set_angular_search_window(angular_window);
  o.set_branch_and_bound_depth(depth);
  return new FastCorrelativeScanMatcher2D(*static_cast<GridHolder*>(grid)->grid, o,
                                          static_cast<DeviceContext*>(ctx));
}
void csmh_matcher_destroy(void* m) { delete static_cast<FastCorrelativeScanMatcher2D*>(m); }
// FastCorrelativeScanMatcher2D::CreateMany: `n` matchers (handles as csmh_matcher_create's) in out[].
void csmh_matchers_create(void* ctx, int n, void* const* grids, double linear_window, double angular_window,
                          int depth, void** out) {
  mapping::scan_matching::proto::FastCorrelativeScanMatcherOptions2D o;
  o.set_linear_search_window(linear_window);
  o.set_angular_search_window(angular_window);
  o.set_branch_and_bound_depth(depth);
  std::vector<const mapping::Grid2D*> list(n);
  for (int i = 0; i < n; ++i) list[i] = static_cast<GridHolder*>(grids[i])->grid.get();
  auto matchers = FastCorrelativeScanMatcher2D::CreateMany(list, o, static_cast<DeviceContext*>(ctx));
  for (int i = 0; i < n; ++i) out[i] = matchers[i].release();
}
const void* csmh_matcher_stack(void* m) {
  return static_cast<FastCorrelativeScanMatcher2D*>(m)->stack();
}

int csmh_matcher_level(void* ctx, void* m, int level, uint8_t* out, int* wide_x, int* wide_y) {
  return csm_stack2d_level(static_cast<DeviceContext*>(ctx)->get(),
                           static_cast<FastCorrelativeScanMatcher2D*>(m)->stack(), level, out,
                           wide_x, wide_y);
}

struct BatchHolder {
  std::vector<sensor::PointCloud> clouds;
  std::unique_ptr<FastCorrelativeScanMatcher2D::Batch> batch;
};

static void Convert(const std::vector<FastCorrelativeScanMatcher2D::Answer>& answers,
                    csmh_result2d* out) {
  for (size_t i = 0; i < answers.size(); ++i) {
    const auto& a = answers[i];
    out[i].found = a.found ? 1 : 0;
    out[i].scan_index = a.best_candidate.scan_index;
    out[i].x_index_offset = a.best_candidate.x_index_offset;
    out[i].y_index_offset = a.best_candidate.y_index_offset;
    out[i].score = a.score;
    out[i].tie_resolution = a.tie_resolution;
    out[i].pose_x = a.pose_estimate.translation().x();
    out[i].pose_y = a.pose_estimate.translation().y();
    out[i].pose_theta = a.pose_estimate.rotation().angle();
    out[i].candidates_scored = a.candidates_scored;
  }
}

// n searches; matcher[i] handles, full[i] flag, init_pose[3i..] (x, y, theta) for
// local searches, clouds concatenated in xyz with offsets[n+1] (in points).
// Prepares (host libm work + upload) without running.
void* csmh_batch_create(void* ctx, int n, void* const* matchers, const int32_t* full,
                        const double* init_pose, const float* xyz, const int64_t* offsets,
                        const float* min_score) {
  BatchHolder* h = new BatchHolder;
  h->clouds.resize(n);
  FastCorrelativeScanMatcher2D::QueryList q(n);
  for (int i = 0; i < n; ++i) {
    for (int64_t p = offsets[i]; p < offsets[i + 1]; ++p)
      h->clouds[i].push_back({sensor::Vector3f(xyz[3 * p], xyz[3 * p + 1], xyz[3 * p + 2])});
    q[i].first = static_cast<const FastCorrelativeScanMatcher2D*>(matchers[i]);
    q[i].second.point_cloud = &h->clouds[i];
    q[i].second.full_submap = full[i] != 0;
    q[i].second.initial_pose_estimate = transform::Rigid2d(
        transform::Vector2d(init_pose[3 * i], init_pose[3 * i + 1]), init_pose[3 * i + 2]);
    q[i].second.min_score = min_score[i];
  }
  h->batch.reset(new FastCorrelativeScanMatcher2D::Batch(static_cast<DeviceContext*>(ctx), q));
  return h;
}
void csmh_batch_run(void* b) { static_cast<BatchHolder*>(b)->batch->Run(); }
void csmh_batch_fetch(void* b, csmh_result2d* out) {
  Convert(static_cast<BatchHolder*>(b)->batch->Fetch(), out);
}
void csmh_batch_destroy(void* b) { delete static_cast<BatchHolder*>(b); }

int csmh_match_batch(void* ctx, int n, void* const* matchers, const int32_t* full,
                     const double* init_pose, const float* xyz, const int64_t* offsets,
                     const float* min_score, csmh_result2d* out) {
  if (n == 0) return 0;
  // The public one-call path: FastCorrelativeScanMatcher2D::MatchBatch streams the device
  // batches of a large job through the device instead of uploading all of them first.
  std::vector<sensor::PointCloud> clouds(n);
  FastCorrelativeScanMatcher2D::QueryList q(n);
  common::ParallelFor(static_cast<size_t>(n), [&](const size_t i) {
    std::vector<sensor::RangefinderPoint> points(static_cast<size_t>(offsets[i + 1] - offsets[i]));
    const float* src = xyz + 3 * offsets[i];
    for (size_t p = 0; p != points.size(); ++p)
      points[p].position = sensor::Vector3f(src[3 * p], src[3 * p + 1], src[3 * p + 2]);
    clouds[i] = sensor::PointCloud(std::move(points));
  }, 64);
  for (int i = 0; i < n; ++i) {
    q[i].first = static_cast<const FastCorrelativeScanMatcher2D*>(matchers[i]);
    q[i].second.point_cloud = &clouds[i];
    q[i].second.full_submap = full[i] != 0;
    q[i].second.initial_pose_estimate = transform::Rigid2d(
        transform::Vector2d(init_pose[3 * i], init_pose[3 * i + 1]), init_pose[3 * i + 2]);
    q[i].second.min_score = min_score[i];
  }
  Convert(FastCorrelativeScanMatcher2D::MatchBatch(static_cast<DeviceContext*>(ctx), q), out);
  return 0;
}

// Rebuilds the matcher's precomputation stack from cells already in device memory
// (asynchronous on the context's stream).
int csmh_matcher_rebuild_dev(void* ctx, void* m, const uint16_t* cells_dev) {
  return csm_stack2d_rebuild_dev(
      static_cast<DeviceContext*>(ctx)->get(),
      const_cast<csm_stack2d*>(static_cast<FastCorrelativeScanMatcher2D*>(m)->stack()), cells_dev);
}

// ProbabilityGridRangeDataInserter2D over a grid holder.  The scans (origin_xy[2i..],
// returns / misses given by offsets in points) are inserted one after the other.
// resident = 0: the reference-signature Insert per scan (the grid grows as needed);
// resident = 1: the grid is grown for all scans first, uploaded once, every scan inserted on
// the device, and downloaded once.  Query the (possibly grown) grid with csmh_grid_info /
// csmh_grid_cells.
int csmh_grid_insert(void* ctx, void* grid, double hit_probability, double miss_probability,
                     int insert_free_space, int resident, int num_scans, const float* origin_xy,
                     const float* returns_xyz, const int64_t* return_offsets, const float* misses_xyz,
                     const int64_t* miss_offsets) {
  mapping::proto::ProbabilityGridRangeDataInserterOptions2D o;
  o.set_hit_probability(hit_probability);
  o.set_miss_probability(miss_probability);
  o.set_insert_free_space(insert_free_space != 0);
  DeviceContext* context = static_cast<DeviceContext*>(ctx);
  const mapping::ProbabilityGridRangeDataInserter2D inserter(o, context);
  mapping::Grid2D* g = static_cast<GridHolder*>(grid)->grid.get();
  std::vector<sensor::RangeData> scans(num_scans);
  for (int i = 0; i < num_scans; ++i) {
    scans[i].origin = sensor::Vector3f(origin_xy[2 * i], origin_xy[2 * i + 1], 0.f);
    for (int64_t p = return_offsets[i]; p < return_offsets[i + 1]; ++p)
      scans[i].returns.push_back({sensor::Vector3f(returns_xyz[3 * p], returns_xyz[3 * p + 1], returns_xyz[3 * p + 2])});
    for (int64_t p = miss_offsets[i]; p < miss_offsets[i + 1]; ++p)
      scans[i].misses.push_back({sensor::Vector3f(misses_xyz[3 * p], misses_xyz[3 * p + 1], misses_xyz[3 * p + 2])});
  }
  if (!resident) {
    for (const sensor::RangeData& rd : scans) inserter.Insert(rd, g);
    return 0;
  }
  for (const sensor::RangeData& rd : scans)  // GrowAsNeeded of every scan, up front
    mapping::ProbabilityGridRangeDataInserter2D::GrowAsNeeded(rd, g);
  mapping::ProbabilityGridRangeDataInserter2D::DeviceGrid device_grid(context, *g);
  std::vector<const sensor::RangeData*> ptrs;
  for (const sensor::RangeData& rd : scans) ptrs.push_back(&rd);
  inserter.Insert(ptrs, &device_grid);
  device_grid.Download(g);
  return 0;
}
void csmh_grid_info(void* grid, int* nx, int* ny, double* max_x, double* max_y) {
  const mapping::MapLimits& l = static_cast<GridHolder*>(grid)->grid->limits();
  *nx = l.cell_limits().num_x_cells;
  *ny = l.cell_limits().num_y_cells;
  *max_x = l.max().x();
  *max_y = l.max().y();
}
void csmh_grid_cells(void* grid, uint16_t* out) {
  const std::vector<uint16_t>& c = static_cast<GridHolder*>(grid)->grid->correspondence_cost_cells();
  std::memcpy(out, c.data(), sizeof(uint16_t) * c.size());
}

// CeresScanMatcher2D::MatchBatch.  options5 = occupied_space_weight, translation_weight,
// rotation_weight, use_nonmonotonic_steps, max_num_iterations; target[2i..], init_pose[3i..];
// out_pose[3i..]; out_summary[6i..] = initial_cost, final_cost, iterations, successful,
// unsuccessful steps, termination type (0 convergence, 1 no convergence, 2 failure).
int csmh_ceres_match_batch(void* ctx, int n, void* const* grids, const double* options5,
                           const double* target, const double* init_pose, const float* xyz,
                           const int64_t* offsets, double* out_pose, double* out_summary) {
  using mapping::scan_matching::CeresScanMatcher2D;
  mapping::scan_matching::proto::CeresScanMatcherOptions2D o;
  o.set_occupied_space_weight(options5[0]);
  o.set_translation_weight(options5[1]);
  o.set_rotation_weight(options5[2]);
  o.mutable_ceres_solver_options()->set_use_nonmonotonic_steps(options5[3] != 0.);
  o.mutable_ceres_solver_options()->set_max_num_iterations(static_cast<int>(options5[4]));
  const CeresScanMatcher2D matcher(o, static_cast<DeviceContext*>(ctx));
  std::vector<sensor::PointCloud> clouds(n);
  std::vector<CeresScanMatcher2D::Query> q(n);
  for (int i = 0; i < n; ++i) {
    for (int64_t p = offsets[i]; p < offsets[i + 1]; ++p)
      clouds[i].push_back({sensor::Vector3f(xyz[3 * p], xyz[3 * p + 1], xyz[3 * p + 2])});
    q[i].point_cloud = &clouds[i];
    q[i].grid = static_cast<GridHolder*>(grids[i])->grid.get();
    q[i].target_translation = transform::Vector2d(target[2 * i], target[2 * i + 1]);
    q[i].initial_pose_estimate = transform::Rigid2d(
        transform::Vector2d(init_pose[3 * i], init_pose[3 * i + 1]), init_pose[3 * i + 2]);
  }
  const auto answers = matcher.MatchBatch(q);
  for (int i = 0; i < n; ++i) {
    const auto& a = answers[i];
    out_pose[3 * i] = a.pose_estimate.translation().x();
    out_pose[3 * i + 1] = a.pose_estimate.translation().y();
    out_pose[3 * i + 2] = a.pose_estimate.rotation().angle();
    out_summary[6 * i] = a.summary.initial_cost;
    out_summary[6 * i + 1] = a.summary.final_cost;
    out_summary[6 * i + 2] = a.summary.num_iterations;
    out_summary[6 * i + 3] = a.summary.num_successful_steps;
    out_summary[6 * i + 4] = a.summary.num_unsuccessful_steps;
    out_summary[6 * i + 5] = static_cast<int>(a.summary.termination_type);
  }
  return 0;
}

// RealTimeCorrelativeScanMatcher2D::Match for n scans in one device batch.
// options4 = {linear_search_window, angular_search_window,
// translation_delta_cost_weight, rotation_delta_cost_weight}; grids[i] handles from
// csmh_grid_create; init_pose[3i..] = (x, y, theta).
int csmh_rt_match_batch(void* ctx, int n, void* const* grids, const double* options4,
                        const double* init_pose, const float* xyz, const int64_t* offsets,
                        csmh_result2d* out) {
  using mapping::scan_matching::RealTimeCorrelativeScanMatcher2D;
  mapping::scan_matching::proto::RealTimeCorrelativeScanMatcherOptions o;
  o.set_linear_search_window(options4[0]);
  o.set_angular_search_window(options4[1]);
  o.set_translation_delta_cost_weight(options4[2]);
  o.set_rotation_delta_cost_weight(options4[3]);
  const RealTimeCorrelativeScanMatcher2D matcher(o, static_cast<DeviceContext*>(ctx));
  std::vector<sensor::PointCloud> clouds(n);
  std::vector<RealTimeCorrelativeScanMatcher2D::Query> q(n);
  common::ParallelFor(static_cast<size_t>(n), [&](const size_t i) {
    std::vector<sensor::RangefinderPoint> points(static_cast<size_t>(offsets[i + 1] - offsets[i]));
    const float* src = xyz + 3 * offsets[i];
    for (size_t p = 0; p != points.size(); ++p)
      points[p].position = sensor::Vector3f(src[3 * p], src[3 * p + 1], src[3 * p + 2]);
    clouds[i] = sensor::PointCloud(std::move(points));
  }, 64);
  for (int i = 0; i < n; ++i) {
    q[i].point_cloud = &clouds[i];
    q[i].grid = static_cast<GridHolder*>(grids[i])->grid.get();
    q[i].initial_pose_estimate = transform::Rigid2d(
        transform::Vector2d(init_pose[3 * i], init_pose[3 * i + 1]), init_pose[3 * i + 2]);
  }
  const auto answers = matcher.MatchBatch(q);
  for (int i = 0; i < n; ++i) {
    const auto& a = answers[i];
    out[i].found = 1;
    out[i].scan_index = a.best_candidate.scan_index;
    out[i].x_index_offset = a.best_candidate.x_index_offset;
    out[i].y_index_offset = a.best_candidate.y_index_offset;
    out[i].score = a.best_candidate.score;
    out[i].tie_resolution = 0;
    out[i].pose_x = a.pose_estimate.translation().x();
    out[i].pose_y = a.pose_estimate.translation().y();
    out[i].pose_theta = a.pose_estimate.rotation().angle();
    out[i].candidates_scored = 0;
  }
  return 0;
}

// ConstraintBuilder2D driven like PoseGraph2D does (one NotifyEndOfNode per search),
// on `num_devices` GPUs.  options8 = {sampling_ratio, max_constraint_distance,
// min_score, global_localization_min_score, linear_search_window,
// angular_search_window, loop_closure_translation_weight,
// loop_closure_rotation_weight}.  Outputs per search i (node index i): found[i] and
// the constraint transform submap <- node as (x, y, yaw) in transform3[3i..].
// stats4 = {searches run, constraints found, candidates scored, flushes}.
// Returns wall seconds from the first MaybeAdd* call to the WhenDone callback.
// csmh_constraint_builder_run2 adds: entry_mode 0 = one MaybeAdd*Constraint + NotifyEndOfNode per
// search (the reference's caller), 1 = MaybeAddConstraints bursts of `burst` requests (whole-node
// batches) followed by the NotifyEndOfNode calls; an InMemoryFamilyFactory registered through
// RegisterMetrics; stats_ext (32 doubles): [0..3] as stats4, [4] local searched, [5] local found,
// [6] global searched, [7] global found, [8] replicated stacks, [9..12] the same four counts read
// back from the metrics families, [13] queue-length gauge after the callback, [14] observations in
// the two score histograms, [15] scores in common::Histogram, [16 + d] searches run on device d.
double csmh_constraint_builder_run2(int num_devices, const int* devices, int num_pool_threads,
                                    int flush_threshold, int depth, const double* options8,
                                    int num_submaps, void* const* grids, const double* submap_pose,
                                    int num_searches, const int32_t* search_submap,
                                    const int32_t* search_full, const double* search_rel_pose,
                                    const float* xyz, const int64_t* offsets, int32_t* found,
                                    double* transform3, double* stats4, int entry_mode, int burst,
                                    double* stats_ext);

double csmh_constraint_builder_run(int num_devices, const int* devices, int num_pool_threads,
                                   int flush_threshold, int depth, const double* options8,
                                   int num_submaps, void* const* grids, const double* submap_pose,
                                   int num_searches, const int32_t* search_submap,
                                   const int32_t* search_full, const double* search_rel_pose,
                                   const float* xyz, const int64_t* offsets, int32_t* found,
                                   double* transform3, double* stats4) {
  return csmh_constraint_builder_run2(num_devices, devices, num_pool_threads, flush_threshold, depth, options8,
                                      num_submaps, grids, submap_pose, num_searches, search_submap, search_full,
                                      search_rel_pose, xyz, offsets, found, transform3, stats4, 0, 0, nullptr);
}

double csmh_constraint_builder_run2(int num_devices, const int* devices, int num_pool_threads,
                                    int flush_threshold, int depth, const double* options8,
                                    int num_submaps, void* const* grids, const double* submap_pose,
                                    int num_searches, const int32_t* search_submap,
                                    const int32_t* search_full, const double* search_rel_pose,
                                    const float* xyz, const int64_t* offsets, int32_t* found,
                                    double* transform3, double* stats4, int entry_mode, int burst,
                                    double* stats_ext) {
  using mapping::constraints::ConstraintBuilder2D;
  mapping::constraints::proto::ConstraintBuilderOptions o;
  o.set_sampling_ratio(options8[0]);
  o.set_max_constraint_distance(options8[1]);
  o.set_min_score(options8[2]);
  o.set_global_localization_min_score(options8[3]);
  o.mutable_fast_correlative_scan_matcher_options()->set_linear_search_window(options8[4]);
  o.mutable_fast_correlative_scan_matcher_options()->set_angular_search_window(options8[5]);
  o.mutable_fast_correlative_scan_matcher_options()->set_branch_and_bound_depth(depth);
  o.set_loop_closure_translation_weight(options8[6]);
  o.set_loop_closure_rotation_weight(options8[7]);
  // options8[8..12]: ceres_scan_matcher_options (occupied_space_weight 0 = no refinement)
  o.mutable_ceres_scan_matcher_options()->set_occupied_space_weight(options8[8]);
  o.mutable_ceres_scan_matcher_options()->set_translation_weight(options8[9]);
  o.mutable_ceres_scan_matcher_options()->set_rotation_weight(options8[10]);
  o.mutable_ceres_scan_matcher_options()->mutable_ceres_solver_options()->set_use_nonmonotonic_steps(
      options8[11] != 0.);
  o.mutable_ceres_scan_matcher_options()->mutable_ceres_solver_options()->set_max_num_iterations(
      static_cast<int>(options8[12]));
  std::vector<mapping::Submap2D> submaps;
  for (int i = 0; i < num_submaps; ++i) {
    submaps.emplace_back(
        transform::Embed3D(transform::Rigid2d(
            transform::Vector2d(submap_pose[3 * i], submap_pose[3 * i + 1]), submap_pose[3 * i + 2])),
        static_cast<GridHolder*>(grids[i])->grid.get());
  }
  std::vector<mapping::TrajectoryNode::Data> nodes(num_searches);
  for (int i = 0; i < num_searches; ++i) {
    for (int64_t p = offsets[i]; p < offsets[i + 1]; ++p)
      nodes[i].filtered_gravity_aligned_point_cloud.push_back(
          {sensor::Vector3f(xyz[3 * p], xyz[3 * p + 1], xyz[3 * p + 2])});
    found[i] = 0;
  }
  double seconds = 0.;
  static metrics::InMemoryFamilyFactory metric_factory;  // the metric handles are process-wide statics
  static bool registered = false;
  if (!registered) {
    ConstraintBuilder2D::RegisterMetrics(&metric_factory);
    registered = true;
  }
  const std::string kCounts = "mapping_constraints_constraint_builder_2d_constraints";
  const std::string kScores = "mapping_constraints_constraint_builder_2d_scores";
  auto metric_snapshot = [&](double* v) {
    v[0] = metric_factory.Value(kCounts + "{matcher=searched,search_region=local}");
    v[1] = metric_factory.Value(kCounts + "{matcher=found,search_region=local}");
    v[2] = metric_factory.Value(kCounts + "{matcher=searched,search_region=global}");
    v[3] = metric_factory.Value(kCounts + "{matcher=found,search_region=global}");
    v[4] = 0.;
    for (const char* region : {"local", "global"})
      for (const double c : metric_factory.BucketCounts(kScores + "{search_region=" + region + "}")) v[4] += c;
  };
  double before[5];
  metric_snapshot(before);
  {
    common::ThreadPool pool(num_pool_threads);
    ConstraintBuilder2D builder(o, &pool, std::vector<int>(devices, devices + num_devices));
    builder.set_flush_threshold(flush_threshold);
    std::mutex m;
    std::condition_variable cv;
    bool done = false;
    ConstraintBuilder2D::Result result;
    const auto t0 = std::chrono::steady_clock::now();
    if (entry_mode == 1) {
      for (int begin = 0; begin < num_searches; begin += std::max(1, burst)) {
        const int end = std::min(num_searches, begin + std::max(1, burst));
        std::vector<ConstraintBuilder2D::ConstraintRequest> requests;
        for (int i = begin; i < end; ++i) {
          ConstraintBuilder2D::ConstraintRequest r;
          r.submap_id = mapping::SubmapId{0, search_submap[i]};
          r.submap = &submaps[search_submap[i]];
          r.node_id = mapping::NodeId{0, i};
          r.constant_data = &nodes[i];
          r.match_full_submap = search_full[i] != 0;
          r.initial_relative_pose = transform::Rigid2d(
              transform::Vector2d(search_rel_pose[3 * i], search_rel_pose[3 * i + 1]), search_rel_pose[3 * i + 2]);
          requests.push_back(r);
        }
        builder.MaybeAddConstraints(requests);
        for (int i = begin; i < end; ++i) builder.NotifyEndOfNode();
      }
    } else {
      for (int i = 0; i < num_searches; ++i) {
        const int s = search_submap[i];
        if (search_full[i]) {
          builder.MaybeAddGlobalConstraint(mapping::SubmapId{0, s}, &submaps[s],
                                           mapping::NodeId{0, i}, &nodes[i]);
        } else {
          builder.MaybeAddConstraint(
              mapping::SubmapId{0, s}, &submaps[s], mapping::NodeId{0, i}, &nodes[i],
              transform::Rigid2d(transform::Vector2d(search_rel_pose[3 * i], search_rel_pose[3 * i + 1]),
                                 search_rel_pose[3 * i + 2]));
        }
        builder.NotifyEndOfNode();
      }
    }
    builder.WhenDone([&](const ConstraintBuilder2D::Result& r) {
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
      transform3[3 * i] = c.pose.zbar_ij.translation().x();
      transform3[3 * i + 1] = c.pose.zbar_ij.translation().y();
      transform3[3 * i + 2] = 2. * std::atan2(c.pose.zbar_ij.rotation().z(), c.pose.zbar_ij.rotation().w());
    }
    const auto st = builder.stats();
    if (stats4) {
      stats4[0] = static_cast<double>(st.searches);
      stats4[1] = static_cast<double>(st.found);
      stats4[2] = static_cast<double>(st.candidates_scored);
      stats4[3] = static_cast<double>(st.flushes);
    }
    if (stats_ext) {
      double after[5];
      metric_snapshot(after);
      for (int k = 0; k < 32; ++k) stats_ext[k] = 0.;
      stats_ext[0] = static_cast<double>(st.searches);
      stats_ext[1] = static_cast<double>(st.found);
      stats_ext[2] = static_cast<double>(st.candidates_scored);
      stats_ext[3] = static_cast<double>(st.flushes);
      stats_ext[4] = static_cast<double>(st.local_searched);
      stats_ext[5] = static_cast<double>(st.local_found);
      stats_ext[6] = static_cast<double>(st.global_searched);
      stats_ext[7] = static_cast<double>(st.global_found);
      stats_ext[8] = static_cast<double>(st.replicated_stacks);
      for (int k = 0; k < 4; ++k) stats_ext[9 + k] = after[k] - before[k];
      stats_ext[13] = metric_factory.Value("mapping_constraints_constraint_builder_2d_queue_length{}");
      stats_ext[14] = after[4] - before[4];
      const std::string hist = builder.ScoreHistogram(10);
      stats_ext[15] = std::atof(hist.c_str() + 7);  // "Count: N ..."
      for (size_t d = 0; d < st.searches_per_device.size() && d < 16; ++d)
        stats_ext[16 + d] = static_cast<double>(st.searches_per_device[d]);
    }
    CSM_CHECK(builder.GetNumFinishedNodes() == num_searches);
    for (int s = 0; s < num_submaps; ++s) builder.DeleteScanMatcher(mapping::SubmapId{0, s});
  }
  return seconds;
}

// GenerateRotatedScans + DiscretizeScans (+ ShrinkToFit when `shrink`) on the device
// for one cloud, with the search parameters the reference would compute.  Call with
// out_xy == NULL to query num_scans.  out_xy: num_scans * n * 2 int32, out_bounds:
// num_scans * 4 int32 (min_x, max_x, min_y, max_y).
int csmh_discretize(void* ctx, double res, double max_x, double max_y, int nx, int ny,
                    const float* xyz, int n, double init_x, double init_y, double init_theta,
                    double linear_window, double angular_window, int shrink, int* num_scans,
                    int32_t* out_xy, int32_t* out_bounds) {
  sensor::PointCloud cloud;
  for (int p = 0; p < n; ++p)
    cloud.push_back({sensor::Vector3f(xyz[3 * p], xyz[3 * p + 1], xyz[3 * p + 2])});
  const mapping::scan_matching::SearchParameters sp(linear_window, angular_window, cloud, res);
  *num_scans = sp.num_scans;
  if (out_xy == nullptr) return 0;
  const std::vector<float> quats = sp.ScanQuaternionTable();
  csm_job2d job;
  std::memset(&job, 0, sizeof(job));
  job.points_xyz = xyz;
  job.num_points = n;
  job.num_scans = sp.num_scans;
  job.scan_quat = quats.data();
  const float half_yaw = 0.5f * static_cast<float>(init_theta);
  job.init_quat_w = std::cos(half_yaw);
  job.init_quat_s = std::sin(half_yaw);
  job.init_tx = static_cast<float>(init_x);
  job.init_ty = static_cast<float>(init_y);
  job.max_x = max_x;
  job.max_y = max_y;
  job.resolution = res;
  job.linear_bound = sp.linear_bounds[0].max_x;
  return csm_discretize2d(static_cast<DeviceContext*>(ctx)->get(), &job, nx, ny, shrink, out_xy,
                          out_bounds);
}

}  // extern "C"
