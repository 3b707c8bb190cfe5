#include "cartographer/mapping/internal/2d/scan_matching/fast_correlative_scan_matcher_2d.h"

#include <mutex>

#include <algorithm>
#include <cmath>
#include <cstring>

#include <map>
#include <memory>

#include "parallel_for.h"

namespace cartographer {
namespace mapping {
namespace scan_matching {
namespace {

// The reference aborts on violated preconditions (glog CHECK); device errors get
// the same treatment, with the C ABI's message.
void CheckOk(int status, const char* what) {
  if (status != CSM_OK) {
    std::fprintf(stderr, "%s failed: %s\n", what, csm_last_error());
    std::abort();
  }
}

}  // namespace

DeviceContext* DeviceContext::ThreadDefault() { return ThreadLocal(0); }

DeviceContext* DeviceContext::ThreadLocal(const int device) {
  thread_local std::map<int, std::unique_ptr<DeviceContext>> contexts;
  std::unique_ptr<DeviceContext>& slot = contexts[device];
  if (!slot) slot.reset(new DeviceContext(device));
  return slot.get();
}

DeviceContext::DeviceContext(int device) : device_(device) {
  CheckOk(csm_ctx_create(device, &ctx_), "csm_ctx_create");
}
DeviceContext::~DeviceContext() { csm_ctx_destroy(ctx_); }

std::vector<uint8> ComputeQuantizationTable(const Grid2D& grid) {
  const float min_score = 1.f - grid.GetMaxCorrespondenceCost();
  const float max_score = 1.f - grid.GetMinCorrespondenceCost();
  const std::vector<float>& cost = grid.value_to_correspondence_cost_table();
  CSM_CHECK(cost.size() == 65536);
  // 65536 lround calls are 0.4 ms -- as long as the whole stack build of a 1000 x 1000 grid -- and
  // the submaps of a map share one value table and one pair of bounds: the last few tables are
  // kept, keyed by the bounds and the FULL contents of the value table (compared, not hashed).
  struct Entry {
    float min_score, max_score;
    std::vector<float> cost;
    std::vector<uint8> table;
  };
  static std::mutex mutex;
  static std::vector<Entry> cache;
  {
    std::lock_guard<std::mutex> lock(mutex);
    for (const Entry& e : cache)
      if (e.min_score == min_score && e.max_score == max_score &&
          std::memcmp(e.cost.data(), cost.data(), sizeof(float) * cost.size()) == 0)
        return e.table;
  }
  const float scale = 255.f / (max_score - min_score);
  std::vector<uint8> table(65536);
  for (size_t v = 0; v != table.size(); ++v) {
    const float probability = 1.f - std::abs(cost[v]);
    const int q = common::RoundToInt((probability - min_score) * scale);
    CSM_CHECK(q >= 0);
    CSM_CHECK(q <= 255);
    table[v] = static_cast<uint8>(q);
  }
  std::lock_guard<std::mutex> lock(mutex);
  if (cache.size() >= 8) cache.erase(cache.begin());
  cache.push_back(Entry{min_score, max_score, cost, table});
  return table;
}

proto::FastCorrelativeScanMatcherOptions2D CreateFastCorrelativeScanMatcherOptions2D(
    common::LuaParameterDictionary* const parameter_dictionary) {
  proto::FastCorrelativeScanMatcherOptions2D options;
  options.set_linear_search_window(parameter_dictionary->GetDouble("linear_search_window"));
  options.set_angular_search_window(parameter_dictionary->GetDouble("angular_search_window"));
  options.set_branch_and_bound_depth(parameter_dictionary->GetInt("branch_and_bound_depth"));
  return options;
}

PrecomputationGrid2D::PrecomputationGrid2D(const Grid2D& grid, const CellLimits& limits, const int width,
                                           std::vector<float>* /*reusable_intermediate_grid*/)
    : offset_(-width + 1, -width + 1),
      wide_limits_(limits.num_x_cells + width - 1, limits.num_y_cells + width - 1),
      min_score_(1.f - grid.GetMaxCorrespondenceCost()),
      max_score_(1.f - grid.GetMinCorrespondenceCost()),
      cells_(static_cast<size_t>(wide_limits_.num_x_cells) * wide_limits_.num_y_cells) {
  CSM_CHECK(width >= 1);
  CSM_CHECK(limits.num_x_cells >= 1);
  CSM_CHECK(limits.num_y_cells >= 1);
  // The reference reads the grid through GetCorrespondenceCost(), i.e. cells outside the grid's own
  // limits count as unknown; `limits` is the grid's cell limits at every call site.
  CSM_CHECK(limits.num_x_cells == grid.limits().cell_limits().num_x_cells);
  CSM_CHECK(limits.num_y_cells == grid.limits().cell_limits().num_y_cells);
  const std::vector<uint8> q = ComputeQuantizationTable(grid);
  CheckOk(csm_precomputation_grid2d(DeviceContext::ThreadDefault()->get(), grid.correspondence_cost_cells().data(),
                                    limits.num_x_cells, limits.num_y_cells, q.data(), width, cells_.data()),
          "csm_precomputation_grid2d");
}

PrecomputationGrid2D::PrecomputationGrid2D(const CellLimits& limits, const int width, const float min_score,
                                           const float max_score, std::vector<uint8> cells)
    : offset_(-width + 1, -width + 1),
      wide_limits_(limits.num_x_cells + width - 1, limits.num_y_cells + width - 1),
      min_score_(min_score),
      max_score_(max_score),
      cells_(std::move(cells)) {
  CSM_CHECK(cells_.size() == static_cast<size_t>(wide_limits_.num_x_cells) * wide_limits_.num_y_cells);
}

PrecomputationGridStack2D::PrecomputationGridStack2D(const Grid2D& grid,
                                                     const proto::FastCorrelativeScanMatcherOptions2D& options)
    : limits_(grid.limits().cell_limits()),
      min_score_(1.f - grid.GetMaxCorrespondenceCost()),
      max_score_(1.f - grid.GetMinCorrespondenceCost()) {
  CSM_CHECK(options.branch_and_bound_depth() >= 1);
  const std::vector<uint8> q = ComputeQuantizationTable(grid);
  build_context_ = DeviceContext::ThreadDefault()->get();
  CheckOk(csm_stack2d_build(build_context_, grid.correspondence_cost_cells().data(), limits_.num_x_cells,
                            limits_.num_y_cells, q.data(), options.branch_and_bound_depth(), &stack_),
          "csm_stack2d_build");
  precomputation_grids_.resize(options.branch_and_bound_depth());
}

PrecomputationGridStack2D::~PrecomputationGridStack2D() { csm_stack2d_destroy(build_context_, stack_); }

const PrecomputationGrid2D& PrecomputationGridStack2D::Get(const int index) {
  CSM_CHECK(index >= 0 && index <= max_depth());
  if (!precomputation_grids_[index]) {
    int wide_x = 0, wide_y = 0;
    csm_ctx* ctx = DeviceContext::ThreadDefault()->get();
    CheckOk(csm_stack2d_level(ctx, stack_, index, nullptr, &wide_x, &wide_y), "csm_stack2d_level");
    std::vector<uint8> cells(static_cast<size_t>(wide_x) * wide_y);
    CheckOk(csm_stack2d_level(ctx, stack_, index, cells.data(), &wide_x, &wide_y), "csm_stack2d_level");
    precomputation_grids_[index].reset(
        new PrecomputationGrid2D(limits_, 1 << index, min_score_, max_score_, std::move(cells)));
  }
  return *precomputation_grids_[index];
}

FastCorrelativeScanMatcher2D::FastCorrelativeScanMatcher2D(
    const Grid2D& grid, const proto::FastCorrelativeScanMatcherOptions2D& options)
    : options_(options), limits_(grid.limits()), context_(DeviceContext::ThreadDefault()) {
  Build(grid);
}

FastCorrelativeScanMatcher2D::FastCorrelativeScanMatcher2D(
    const Grid2D& grid, const proto::FastCorrelativeScanMatcherOptions2D& options,
    DeviceContext* context)
    : options_(options), limits_(grid.limits()), context_(context) {
  Build(grid);
}

FastCorrelativeScanMatcher2D::FastCorrelativeScanMatcher2D(
    const Grid2D& grid, const proto::FastCorrelativeScanMatcherOptions2D& options, DeviceContext* context,
    Prebuilt)
    : options_(options), limits_(grid.limits()), context_(context) {
  CSM_CHECK(options_.branch_and_bound_depth() >= 1);
  min_score_ = 1.f - grid.GetMaxCorrespondenceCost();
  max_score_ = 1.f - grid.GetMinCorrespondenceCost();
  build_context_ = context_->get();
  device_ = context_->device();
}

std::vector<std::unique_ptr<FastCorrelativeScanMatcher2D>> FastCorrelativeScanMatcher2D::CreateMany(
    const std::vector<const Grid2D*>& grids, const proto::FastCorrelativeScanMatcherOptions2D& options,
    DeviceContext* context) {
  const size_t n = grids.size();
  std::vector<std::unique_ptr<FastCorrelativeScanMatcher2D>> matchers(n);
  if (n == 0) return matchers;
  std::vector<std::vector<uint8>> tables(n);
  std::vector<const uint16_t*> cells(n);
  std::vector<const uint8_t*> luts(n);
  std::vector<int> num_x(n), num_y(n);
  for (size_t i = 0; i != n; ++i) {
    CSM_CHECK(grids[i] != nullptr);
    matchers[i].reset(new FastCorrelativeScanMatcher2D(*grids[i], options, context, Prebuilt()));
    tables[i] = ComputeQuantizationTable(*grids[i]);
    cells[i] = grids[i]->correspondence_cost_cells().data();
    luts[i] = tables[i].data();
    num_x[i] = grids[i]->limits().cell_limits().num_x_cells;
    num_y[i] = grids[i]->limits().cell_limits().num_y_cells;
  }
  std::vector<csm_stack2d*> stacks(n, nullptr);
  CheckOk(csm_stack2d_build_many(context->get(), static_cast<int>(n), cells.data(), num_x.data(), num_y.data(),
                                 luts.data(), options.branch_and_bound_depth(), stacks.data()),
          "csm_stack2d_build_many");
  for (size_t i = 0; i != n; ++i) matchers[i]->stack_ = stacks[i];
  return matchers;
}

void FastCorrelativeScanMatcher2D::Build(const Grid2D& grid) {
  CSM_CHECK(options_.branch_and_bound_depth() >= 1);
  min_score_ = 1.f - grid.GetMaxCorrespondenceCost();
  max_score_ = 1.f - grid.GetMinCorrespondenceCost();
  const std::vector<uint8> q = ComputeQuantizationTable(grid);
  const CellLimits& cells = limits_.cell_limits();
  build_context_ = context_->get();
  device_ = context_->device();
  CheckOk(csm_stack2d_build(build_context_, grid.correspondence_cost_cells().data(),
                            cells.num_x_cells, cells.num_y_cells, q.data(),
                            options_.branch_and_bound_depth(), &stack_),
          "csm_stack2d_build");
}

FastCorrelativeScanMatcher2D::~FastCorrelativeScanMatcher2D() {
  // The context the stack was built on may already be gone (the ABI checks its registry
  // of live contexts before handing the block back to that context's pool).
  csm_stack2d_destroy(build_context_, stack_);
}

bool FastCorrelativeScanMatcher2D::Match(const transform::Rigid2d& initial_pose_estimate,
                                         const sensor::PointCloud& point_cloud,
                                         const float min_score, float* score,
                                         transform::Rigid2d* pose_estimate) const {
  CSM_CHECK(score != nullptr);
  CSM_CHECK(pose_estimate != nullptr);
  Query q;
  q.point_cloud = &point_cloud;
  q.full_submap = false;
  q.initial_pose_estimate = initial_pose_estimate;
  q.min_score = min_score;
  // Whatever context the matcher was built on, a call runs on the calling thread's own context
  // for the stack's device (a csm_ctx is single-threaded; Match is const and called concurrently).
  DeviceContext* ctx = DeviceContext::ThreadLocal(device_);
  const Answer a = MatchBatch(ctx, {{this, q}})[0];
  if (!a.found) return false;
  *score = a.score;
  *pose_estimate = a.pose_estimate;
  return true;
}

bool FastCorrelativeScanMatcher2D::MatchFullSubmap(const sensor::PointCloud& point_cloud,
                                                   const float min_score, float* score,
                                                   transform::Rigid2d* pose_estimate) const {
  CSM_CHECK(score != nullptr);
  CSM_CHECK(pose_estimate != nullptr);
  Query q;
  q.point_cloud = &point_cloud;
  q.full_submap = true;
  q.min_score = min_score;
  DeviceContext* ctx = DeviceContext::ThreadLocal(device_);
  const Answer a = MatchBatch(ctx, {{this, q}})[0];
  if (!a.found) return false;
  *score = a.score;
  *pose_estimate = a.pose_estimate;
  return true;
}

std::vector<FastCorrelativeScanMatcher2D::Answer> FastCorrelativeScanMatcher2D::MatchBatch(
    DeviceContext* context, const QueryList& queries) {
  if (queries.empty()) return {};
  Batch batch(context, queries, /*resident=*/false);
  return batch.RunStreaming(queries);
}

FastCorrelativeScanMatcher2D::Batch::Batch(DeviceContext* context, const QueryList& queries,
                                           const bool resident)
    : context_(context) {
  const size_t n = queries.size();
  initial_.resize(n);
  xyz_.resize(n);
  quats_.resize(n);
  jobs_.resize(n);
  std::vector<csm_job2d>& jobs = jobs_;
  std::vector<double> lattice_estimate(n);
  params_.assign(n, SearchParameters(0, 0, 0., 0.));
  // Host side of every search (search parameters, angle tables: libm): independent.
  common::ParallelFor(n, [&](const size_t i) {
    const FastCorrelativeScanMatcher2D& m = *queries[i].first;
    const Query& q = queries[i].second;
    CSM_CHECK(q.point_cloud != nullptr);
    const double res = m.limits_.resolution();
    if (q.full_submap) {
      // Window around the submap centre that includes it fully
      // (reference: fast_correlative_scan_matcher_2d.cc:215-222).
      params_[i] = SearchParameters(1e6 * res, M_PI, *q.point_cloud, res);
      const double half_res = 0.5 * res;
      initial_[i] = transform::Rigid2d::Translation(transform::Vector2d(
          m.limits_.max().x() - half_res * m.limits_.cell_limits().num_y_cells,
          m.limits_.max().y() - half_res * m.limits_.cell_limits().num_x_cells));
    } else {
      params_[i] = SearchParameters(m.options_.linear_search_window(), m.options_.angular_search_window(),
                                    *q.point_cloud, res);
      initial_[i] = q.initial_pose_estimate;
    }
    const SearchParameters& sp = params_[i];
    double reach = 0.;
    for (size_t p = 0; p != q.point_cloud->size(); ++p) {
      const sensor::Vector3f& v = (*q.point_cloud)[p].position;
      reach = std::max(reach, std::hypot(static_cast<double>(v.x()), static_cast<double>(v.y())));
    }
    csm_job2d& job = jobs[i];
    std::memset(&job, 0, sizeof(job));
    job.stack = m.stack_;
    job.num_points = static_cast<int32_t>(q.point_cloud->size());
    job.num_scans = sp.num_scans;  // points_xyz and scan_quat: Prepare()
    // Rotation2Dd -> Rotation2Df -> AngleAxisf -> Quaternionf (reference :235-239)
    const float yaw = static_cast<float>(initial_[i].rotation().angle());
    const float half_yaw = 0.5f * yaw;
    job.init_quat_w = std::cos(half_yaw);
    job.init_quat_s = std::sin(half_yaw);
    job.init_tx = static_cast<float>(initial_[i].translation().x());
    job.init_ty = static_cast<float>(initial_[i].translation().y());
    job.max_x = m.limits_.max().x();
    job.max_y = m.limits_.max().y();
    job.resolution = res;
    job.linear_bound = sp.linear_bounds.empty() ? 0 : sp.linear_bounds[0].max_x;
    job.min_score = q.min_score;
    job.score_offset = m.min_score_;
    job.score_scale = (m.max_score_ - m.min_score_) / 255.f;
    // Rough size of the lowest-resolution lattice, to bound the device footprint.
    const double step = static_cast<double>(1 << (m.options_.branch_and_bound_depth() - 1));
    const double ext = 2. * reach / res + 4.;
    const double wx = std::min(2. * job.linear_bound, m.limits_.cell_limits().num_x_cells + ext);
    const double wy = std::min(2. * job.linear_bound, m.limits_.cell_limits().num_y_cells + ext);
    lattice_estimate[i] = sp.num_scans * (wx / step + 1.) * (wy / step + 1.);
  });
  // Split into device batches of bounded footprint.
  constexpr double kLatticeBudget = 128e6;  // lowest-resolution candidates per device batch (4 B each)
  constexpr size_t kMaxJobs = 4096;
  size_t begin = 0;
  while (begin < n) {
    size_t end = begin;
    double sum = 0.;
    while (end < n && end - begin < kMaxJobs && (end == begin || sum + lattice_estimate[end] <= kLatticeBudget)) {
      sum += lattice_estimate[end];
      ++end;
    }
    chunks_.emplace_back(begin, end);
    if (resident) {
      Prepare(queries, begin, end);
      csm_batch2d* device_batch = nullptr;
      CheckOk(csm_batch2d_upload(context_->get(), jobs.data() + begin, static_cast<int>(end - begin),
                                 &device_batch),
              "csm_batch2d_upload");
      device_batches_.push_back(device_batch);
    }
    begin = end;
  }
}

void FastCorrelativeScanMatcher2D::Batch::Prepare(const QueryList& queries, const size_t begin, const size_t end) {
  common::ParallelFor(end - begin, [&](const size_t k) {
    const size_t i = begin + k;
    const sensor::PointCloud& cloud = *queries[i].second.point_cloud;
    xyz_[i].resize(3 * cloud.size());
    for (size_t p = 0; p != cloud.size(); ++p) {
      const sensor::Vector3f& v = cloud[p].position;
      xyz_[i][3 * p] = v.x();
      xyz_[i][3 * p + 1] = v.y();
      xyz_[i][3 * p + 2] = v.z();
    }
    quats_[i] = params_[i].ScanQuaternionTable();
    jobs_[i].points_xyz = xyz_[i].data();
    jobs_[i].scan_quat = quats_[i].data();
  });
}

FastCorrelativeScanMatcher2D::Batch::~Batch() {
  for (csm_batch2d* b : device_batches_) csm_batch2d_destroy(context_->get(), b);
}

void FastCorrelativeScanMatcher2D::Batch::Run() {
  for (csm_batch2d* b : device_batches_) CheckOk(csm_batch2d_run(context_->get(), b), "csm_batch2d_run");
}

std::vector<FastCorrelativeScanMatcher2D::Answer> FastCorrelativeScanMatcher2D::Batch::Fetch() {
  const size_t n = initial_.size();
  std::vector<csm_result2d> results(n);
  CSM_CHECK(device_batches_.size() == chunks_.size());  // resident form
  for (size_t c = 0; c != chunks_.size(); ++c)
    CheckOk(csm_batch2d_fetch(context_->get(), device_batches_[c], results.data() + chunks_[c].first),
            "csm_batch2d_fetch");
  return Convert(results);
}

std::vector<FastCorrelativeScanMatcher2D::Answer> FastCorrelativeScanMatcher2D::Batch::RunStreaming(
    const QueryList& queries) {
  // Device batch k+1 is prepared on the host (angle tables: libm), uploaded and its kernels queued
  // while k runs; a batch is fetched (its own event, not the stream) and its blocks go back to the
  // context's pool once kInFlight are queued.  Only the first device batch's host preparation is
  // not hidden behind device work.
  constexpr size_t kInFlight = 3;
  std::vector<csm_result2d> results(initial_.size());
  std::vector<std::pair<size_t, csm_batch2d*>> live;
  auto retire = [&]() {
    const size_t c = live.front().first;
    CheckOk(csm_batch2d_fetch(context_->get(), live.front().second, results.data() + chunks_[c].first),
            "csm_batch2d_fetch");
    csm_batch2d_destroy(context_->get(), live.front().second);
    live.erase(live.begin());
  };
  CSM_CHECK(queries.size() == initial_.size());
  if (!chunks_.empty()) Prepare(queries, chunks_[0].first, chunks_[0].second);
  for (size_t c = 0; c != chunks_.size(); ++c) {
    csm_batch2d* device_batch = nullptr;
    CheckOk(csm_batch2d_upload(context_->get(), jobs_.data() + chunks_[c].first,
                               static_cast<int>(chunks_[c].second - chunks_[c].first), &device_batch),
            "csm_batch2d_upload");
    CheckOk(csm_batch2d_run(context_->get(), device_batch), "csm_batch2d_run");
    live.emplace_back(c, device_batch);
    // the upload staged this batch's points and tables: the host copies are done with
    for (size_t i = chunks_[c].first; i != chunks_[c].second; ++i) {
      std::vector<float>().swap(xyz_[i]);
      std::vector<float>().swap(quats_[i]);
      jobs_[i].points_xyz = nullptr;
      jobs_[i].scan_quat = nullptr;
    }
    if (c + 1 != chunks_.size()) Prepare(queries, chunks_[c + 1].first, chunks_[c + 1].second);
    if (live.size() >= kInFlight) retire();
  }
  while (!live.empty()) retire();
  return Convert(results);
}

std::vector<FastCorrelativeScanMatcher2D::Answer> FastCorrelativeScanMatcher2D::Batch::Convert(
    const std::vector<csm_result2d>& results) const {
  const size_t n = initial_.size();
  std::vector<Answer> answers(n);
  for (size_t i = 0; i != n; ++i) {
    const csm_result2d& r = results[i];
    Answer& a = answers[i];
    a.found = r.found == 1;
    a.candidates_scored = r.candidates_scored;
    a.tie_resolution = r.tie_resolution;
    a.best_candidate = Candidate2D(r.scan_index, r.x_index_offset, r.y_index_offset, params_[i]);
    a.best_candidate.score = r.score;
    a.score = r.score;
    if (a.found) {  // reference :253-260
      a.pose_estimate = transform::Rigid2d(
          transform::Vector2d(initial_[i].translation().x() + a.best_candidate.x,
                              initial_[i].translation().y() + a.best_candidate.y),
          initial_[i].rotation() * transform::Rotation2Dd(a.best_candidate.orientation));
    }
  }
  return answers;
}

}  // namespace scan_matching
}  // namespace mapping
}  // namespace cartographer
