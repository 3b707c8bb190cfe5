#include "cartographer/mapping/internal/3d/scan_matching/fast_correlative_scan_matcher_3d.h"

#include <algorithm>
#include <cmath>
#include <cstring>

#include "parallel_for.h"

namespace cartographer {
namespace mapping {
namespace scan_matching {
namespace {

void CheckOk(int status, const char* what) {
  if (status != CSM_OK) {
    std::fprintf(stderr, "%s failed: %s\n", what, csm_last_error());
    std::abort();
  }
}

constexpr float kMinProbability = 0.1f;
constexpr float kMaxProbability = 1.f - kMinProbability;

}  // namespace

// value_conversion_tables.cc:29-67 for the probability range: value 0 is unknown
// (-> kMinProbability), the update-marker bit is ignored.
const std::vector<float>& ValueToProbabilityTable() {
  static const std::vector<float> table = [] {
    std::vector<float> t(65536);
    const float scale = (kMaxProbability - kMinProbability) / 32766.f;
    for (uint32_t v = 0; v != 65536; ++v) {
      const uint16 value = static_cast<uint16>(v) & static_cast<uint16>(~(1u << 15));
      t[v] = value == 0 ? kMinProbability : value * scale + (kMinProbability - scale);
    }
    return t;
  }();
  return table;
}

// precomputation_grid_3d.cc:49-61
std::vector<uint8> ComputeQuantizationTable3D() {
  // no inputs: computed once per process (function-local static, thread-safe initialisation)
  static const std::vector<uint8> kTable = [] {
    const std::vector<float>& probability = ValueToProbabilityTable();
    std::vector<uint8> table(65536);
    for (size_t v = 0; v != table.size(); ++v) {
      const int cell_value = common::RoundToInt((probability[v] - kMinProbability) *
                                                (255.f / (kMaxProbability - kMinProbability)));
      CSM_CHECK(cell_value >= 0);
      CSM_CHECK(cell_value <= 255);
      table[v] = static_cast<uint8>(cell_value);
    }
    return table;
  }();
  return kTable;
}

proto::FastCorrelativeScanMatcherOptions3D CreateFastCorrelativeScanMatcherOptions3D(
    common::LuaParameterDictionary* const parameter_dictionary) {
  proto::FastCorrelativeScanMatcherOptions3D options;
  options.set_branch_and_bound_depth(parameter_dictionary->GetInt("branch_and_bound_depth"));
  options.set_full_resolution_depth(parameter_dictionary->GetInt("full_resolution_depth"));
  options.set_min_rotational_score(parameter_dictionary->GetDouble("min_rotational_score"));
  options.set_min_low_resolution_score(parameter_dictionary->GetDouble("min_low_resolution_score"));
  options.set_linear_xy_search_window(parameter_dictionary->GetDouble("linear_xy_search_window"));
  options.set_linear_z_search_window(parameter_dictionary->GetDouble("linear_z_search_window"));
  options.set_angular_search_window(parameter_dictionary->GetDouble("angular_search_window"));
  return options;
}

PrecomputationGridStack3D::PrecomputationGridStack3D(const HybridGrid& hybrid_grid,
                                                     const proto::FastCorrelativeScanMatcherOptions3D& options) {
  CSM_CHECK(options.branch_and_bound_depth() >= 1);  // reference :60
  CSM_CHECK(options.full_resolution_depth() >= 1);   // reference :61
  std::vector<int32_t> xyz;
  std::vector<uint16> values;
  hybrid_grid.Export(&xyz, &values);
  const std::vector<uint8> q = ComputeQuantizationTable3D();
  csm_ctx* ctx = DeviceContext::ThreadDefault()->get();
  csm_stack3d* stack = nullptr;
  CheckOk(csm_stack3d_build(ctx, xyz.data(), values.data(), static_cast<int>(values.size()), q.data(),
                            options.branch_and_bound_depth(), options.full_resolution_depth(), &stack),
          "csm_stack3d_build");
  // Every level reports the source grid's resolution(), half-resolution levels included
  // (PrecomputeGrid constructs its result with grid.resolution(), precomputation_grid_3d.cc:66).
  const float resolution = hybrid_grid.resolution();
  for (int depth = 0; depth != options.branch_and_bound_depth(); ++depth) {
    int32_t lo[3] = {0, 0, 0}, dims[3] = {0, 0, 0};
    CheckOk(csm_stack3d_level(ctx, stack, depth, nullptr, lo, dims), "csm_stack3d_level");
    std::vector<uint8> bytes(static_cast<size_t>(dims[0]) * dims[1] * dims[2]);
    if (!bytes.empty()) CheckOk(csm_stack3d_level(ctx, stack, depth, bytes.data(), lo, dims), "csm_stack3d_level");
    precomputation_grids_.emplace_back(resolution, lo, dims, std::move(bytes));
  }
  csm_stack3d_destroy(ctx, stack);
}

FastCorrelativeScanMatcher3D::FastCorrelativeScanMatcher3D(
    const HybridGrid& hybrid_grid, const HybridGrid* const low_resolution_hybrid_grid,
    const Eigen::VectorXf* const rotational_scan_matcher_histogram,
    const proto::FastCorrelativeScanMatcherOptions3D& options)
    : FastCorrelativeScanMatcher3D(hybrid_grid, low_resolution_hybrid_grid,
                                   rotational_scan_matcher_histogram, options,
                                   DeviceContext::ThreadDefault()) {}

FastCorrelativeScanMatcher3D::FastCorrelativeScanMatcher3D(
    const HybridGrid& hybrid_grid, const HybridGrid* const low_resolution_hybrid_grid,
    const Eigen::VectorXf* const rotational_scan_matcher_histogram,
    const proto::FastCorrelativeScanMatcherOptions3D& options, DeviceContext* const context)
    : options_(options),
      resolution_(hybrid_grid.resolution()),
      width_in_voxels_(hybrid_grid.grid_size()),
      low_resolution_hybrid_grid_(low_resolution_hybrid_grid),
      rotational_scan_matcher_(rotational_scan_matcher_histogram),
      context_(context) {
  Build(hybrid_grid);
}

void FastCorrelativeScanMatcher3D::Build(const HybridGrid& hybrid_grid) {
  CSM_CHECK(options_.branch_and_bound_depth() >= 1);   // reference :60
  CSM_CHECK(options_.full_resolution_depth() >= 1);    // reference :61
  CSM_CHECK(low_resolution_hybrid_grid_ != nullptr);
  std::vector<int32_t> xyz;
  std::vector<uint16> values;
  hybrid_grid.Export(&xyz, &values);
  const std::vector<uint8> q = ComputeQuantizationTable3D();
  build_context_ = context_->get();
  device_ = context_->device();
  CheckOk(csm_stack3d_build(build_context_, xyz.data(), values.data(), static_cast<int>(values.size()),
                            q.data(), options_.branch_and_bound_depth(),
                            options_.full_resolution_depth(), &stack_),
          "csm_stack3d_build");
  low_resolution_hybrid_grid_->Export(&xyz, &values);
  CheckOk(csm_grid3d_build(context_->get(), xyz.data(), values.data(), static_cast<int>(values.size()),
                           ValueToProbabilityTable().data(), &low_grid_),
          "csm_grid3d_build");
}

FastCorrelativeScanMatcher3D::~FastCorrelativeScanMatcher3D() {
  // build_context_ may be stale by now; the ABI checks it against its live contexts.
  csm_stack3d_destroy(build_context_, stack_);
  csm_grid3d_destroy(build_context_, low_grid_);
}

std::unique_ptr<FastCorrelativeScanMatcher3D::Result> FastCorrelativeScanMatcher3D::Match(
    const transform::Rigid3d& global_node_pose, const transform::Rigid3d& global_submap_pose,
    const TrajectoryNode::Data& constant_data, const float min_score) const {
  Query q;
  q.constant_data = &constant_data;
  q.full_submap = false;
  q.global_node_pose = global_node_pose;
  q.global_submap_pose = global_submap_pose;
  q.min_score = min_score;
  DeviceContext* ctx = DeviceContext::ThreadLocal(device_);
  const Answer a = MatchBatch(ctx, {{this, q}})[0];
  if (!a.found) return nullptr;
  return std::unique_ptr<Result>(new Result(a.result));
}

std::unique_ptr<FastCorrelativeScanMatcher3D::Result> FastCorrelativeScanMatcher3D::MatchFullSubmap(
    const Eigen::Quaterniond& global_node_rotation, const Eigen::Quaterniond& global_submap_rotation,
    const TrajectoryNode::Data& constant_data, const float min_score) const {
  Query q;
  q.constant_data = &constant_data;
  q.full_submap = true;
  q.global_node_pose = transform::Rigid3d(transform::Vector3d(), global_node_rotation);
  q.global_submap_pose = transform::Rigid3d(transform::Vector3d(), global_submap_rotation);
  q.min_score = min_score;
  DeviceContext* ctx = DeviceContext::ThreadLocal(device_);
  const Answer a = MatchBatch(ctx, {{this, q}})[0];
  if (!a.found) return nullptr;
  return std::unique_ptr<Result>(new Result(a.result));
}

// reference :127-170
FastCorrelativeScanMatcher3D::SearchParameters FastCorrelativeScanMatcher3D::ParametersFor(
    const Query& query) const {
  if (!query.full_submap) {
    return SearchParameters{common::RoundToInt(options_.linear_xy_search_window() / resolution_),
                            common::RoundToInt(options_.linear_z_search_window() / resolution_),
                            options_.angular_search_window()};
  }
  float max_point_distance = 0.f;
  for (const sensor::RangefinderPoint& point : query.constant_data->high_resolution_point_cloud)
    max_point_distance = std::max(max_point_distance, transform::Norm(point.position));
  const int linear_window_size =
      (width_in_voxels_ + 1) / 2 + common::RoundToInt(max_point_distance / resolution_ + 0.5f);
  return SearchParameters{linear_window_size, linear_window_size, M_PI};
}

// reference :246-295 without the per-scan DiscretizeScan (device): angles, rotational
// prefilter, pose of each kept yaw slice.
void FastCorrelativeScanMatcher3D::GenerateDiscreteScanPoses(
    const SearchParameters& search_parameters, const TrajectoryNode::Data& constant_data,
    const transform::Rigid3f& global_node_pose, const transform::Rigid3f& global_submap_pose,
    std::vector<float>* const scan_poses, std::vector<float>* const scan_leaf_quat,
    std::vector<float>* const rotational_scores) const {
  scan_poses->clear();
  scan_leaf_quat->clear();
  rotational_scores->clear();
  // Something on the order of resolution keeps the acos() below defined.
  float max_scan_range = 3.f * resolution_;
  for (const sensor::RangefinderPoint& point : constant_data.high_resolution_point_cloud)
    max_scan_range = std::max(transform::Norm(point.position), max_scan_range);
  const float kSafetyMargin = 1.f - 1e-2f;
  const float angular_step_size =
      kSafetyMargin *
      std::acos(1.f - (resolution_ * resolution_) / (2.f * (max_scan_range * max_scan_range)));
  const int angular_window_size =
      common::RoundToInt(search_parameters.angular_search_window / angular_step_size);
  std::vector<float> angles;
  angles.reserve(2 * angular_window_size + 1);
  for (int rz = -angular_window_size; rz <= angular_window_size; ++rz)
    angles.push_back(rz * angular_step_size);
  const transform::Rigid3f node_to_submap = global_submap_pose.inverse() * global_node_pose;
  const std::vector<float> scores = rotational_scan_matcher_.Match(
      constant_data.rotational_scan_matcher_histogram,
      transform::GetYaw(node_to_submap.rotation() * transform::InverseAsFloat(constant_data.gravity_alignment)),
      angles);
  const transform::Quaternionf submap_rotation_inverse = global_submap_pose.rotation().inverse();
  for (size_t i = 0; i != angles.size(); ++i) {
    if (scores[i] < options_.min_rotational_score()) continue;
    // The yaw is applied between the translation and the rotation of the initial
    // pose: around the origin of the range data, in the map frame.
    const transform::Quaternionf rotation =
        submap_rotation_inverse *
        transform::AngleAxisVectorToRotationQuaternion(transform::Vector3f(0.f, 0.f, angles[i])) *
        global_node_pose.rotation();
    const transform::Vector3f& t = node_to_submap.translation();
    for (const float v : {t.x(), t.y(), t.z(), rotation.w(), rotation.x(), rotation.y(), rotation.z()})
      scan_poses->push_back(v);
    // GetPoseFromCandidate (:363-369) multiplies Translation(..) * pose: the product
    // of the identity rotation with this one, normalised.
    const transform::Quaternionf leaf = (transform::Quaternionf() * rotation).normalized();
    for (const float v : {leaf.w(), leaf.x(), leaf.y(), leaf.z()}) scan_leaf_quat->push_back(v);
    rotational_scores->push_back(scores[i]);
  }
}

std::vector<float> FastCorrelativeScanMatcher3D::DiscreteScanPosesForTesting(
    const transform::Rigid3d& global_node_pose, const transform::Rigid3d& global_submap_pose,
    const TrajectoryNode::Data& constant_data, std::vector<float>* rotational_scores) const {
  Query q;
  q.constant_data = &constant_data;
  std::vector<float> poses, leaf, scores;
  GenerateDiscreteScanPoses(ParametersFor(q), constant_data, transform::CastToFloat(global_node_pose),
                            transform::CastToFloat(global_submap_pose), &poses, &leaf, &scores);
  if (rotational_scores != nullptr) *rotational_scores = scores;
  return poses;
}

std::vector<FastCorrelativeScanMatcher3D::Answer> FastCorrelativeScanMatcher3D::MatchBatch(
    DeviceContext* context, const QueryList& queries) {
  if (queries.empty()) return {};
  Batch batch(context, queries);
  batch.Run();
  return batch.Fetch();
}

FastCorrelativeScanMatcher3D::Batch::Batch(DeviceContext* context, const QueryList& queries)
    : context_(context) {
  const size_t n = queries.size();
  prepared_.resize(n);
  std::vector<csm_job3d> jobs(n);
  std::vector<double> footprint(n);
  // Host side of every search (angles, rotational prefilter, slice poses): independent.
  common::ParallelFor(n, [&](const size_t i) {
    const FastCorrelativeScanMatcher3D& m = *queries[i].first;
    const Query& q = queries[i].second;
    CSM_CHECK(q.constant_data != nullptr);
    const TrajectoryNode::Data& data = *q.constant_data;
    Prepared& p = prepared_[i];
    const SearchParameters sp = m.ParametersFor(q);
    // MatchFullSubmap passes pure rotations (reference :165-167).
    transform::Rigid3f node = transform::CastToFloat(q.global_node_pose);
    transform::Rigid3f submap = transform::CastToFloat(q.global_submap_pose);
    if (q.full_submap) {
      node = transform::Rigid3f::Rotation(node.rotation());
      submap = transform::Rigid3f::Rotation(submap.rotation());
    }
    m.GenerateDiscreteScanPoses(sp, data, node, submap, &p.scan_poses, &p.scan_leaf_quat,
                                &p.rotational_scores);
    p.resolution = m.resolution_;
    p.points.reserve(3 * data.high_resolution_point_cloud.size());
    for (const sensor::RangefinderPoint& point : data.high_resolution_point_cloud)
      for (const float v : {point.position.x(), point.position.y(), point.position.z()}) p.points.push_back(v);
    p.low_points.reserve(3 * data.low_resolution_point_cloud.size());
    for (const sensor::RangefinderPoint& point : data.low_resolution_point_cloud)
      for (const float v : {point.position.x(), point.position.y(), point.position.z()})
        p.low_points.push_back(v);
    csm_job3d& job = jobs[i];
    std::memset(&job, 0, sizeof(job));
    job.stack = m.stack_;
    job.low_resolution_grid = m.low_grid_;
    job.points_xyz = p.points.data();
    job.num_points = static_cast<int32_t>(data.high_resolution_point_cloud.size());
    job.low_points_xyz = p.low_points.data();
    job.num_low_points = static_cast<int32_t>(data.low_resolution_point_cloud.size());
    job.num_scans = static_cast<int32_t>(p.rotational_scores.size());
    job.scan_poses = p.scan_poses.data();
    job.scan_leaf_quat = p.scan_leaf_quat.data();
    job.resolution = m.resolution_;
    job.low_resolution = m.low_resolution_hybrid_grid_->resolution();
    job.linear_xy_window = sp.linear_xy_window_size;
    job.linear_z_window = sp.linear_z_window_size;
    job.min_score = q.min_score;
    job.min_low_resolution_score = m.options_.min_low_resolution_score();
    footprint[i] = static_cast<double>(job.num_scans) * job.num_points;
  });
  constexpr double kCellBudget = 128e6;  // discrete cells (16 B each) per device batch
  constexpr size_t kMaxJobs = 1024;
  size_t begin = 0;
  while (begin < n) {
    size_t end = begin;
    double sum = 0.;
    while (end < n && end - begin < kMaxJobs && (end == begin || sum + footprint[end] <= kCellBudget)) {
      sum += footprint[end];
      ++end;
    }
    csm_batch3d* device_batch = nullptr;
    CheckOk(csm_batch3d_upload(context_->get(), jobs.data() + begin, static_cast<int>(end - begin),
                               &device_batch),
            "csm_batch3d_upload");
    chunks_.emplace_back(begin, end);
    device_batches_.push_back(device_batch);
    begin = end;
  }
}

FastCorrelativeScanMatcher3D::Batch::~Batch() {
  for (csm_batch3d* b : device_batches_) csm_batch3d_destroy(context_->get(), b);
}

void FastCorrelativeScanMatcher3D::Batch::Run() {
  for (csm_batch3d* b : device_batches_) CheckOk(csm_batch3d_run(context_->get(), b), "csm_batch3d_run");
}

std::vector<FastCorrelativeScanMatcher3D::Answer> FastCorrelativeScanMatcher3D::Batch::Fetch() {
  const size_t n = prepared_.size();
  std::vector<csm_result3d> results(n);
  for (size_t c = 0; c != chunks_.size(); ++c)
    CheckOk(csm_batch3d_fetch(context_->get(), device_batches_[c], results.data() + chunks_[c].first),
            "csm_batch3d_fetch");
  std::vector<Answer> answers(n);
  for (size_t i = 0; i != n; ++i) {
    const csm_result3d& r = results[i];
    const Prepared& p = prepared_[i];
    Answer& a = answers[i];
    a.found = r.found == 1;
    a.candidates_scored = r.candidates_scored;
    a.tie_resolution = r.tie_resolution;
    a.num_scans = static_cast<int>(p.rotational_scores.size());
    a.scan_index = r.scan_index;
    for (int k = 0; k < 3; ++k) a.offset[k] = r.offset[k];
    if (!a.found) continue;
    // GetPoseFromCandidate (:363-369): Translation(resolution * offset) * scan pose.
    const float* sp = p.scan_poses.data() + 7 * r.scan_index;
    const transform::Rigid3f scan_pose(transform::Vector3f(sp[0], sp[1], sp[2]),
                                       transform::Quaternionf(sp[3], sp[4], sp[5], sp[6]));
    const transform::Rigid3f pose =
        transform::Rigid3f::Translation(transform::Vector3f(p.resolution * static_cast<float>(r.offset[0]),
                                                            p.resolution * static_cast<float>(r.offset[1]),
                                                            p.resolution * static_cast<float>(r.offset[2]))) *
        scan_pose;
    a.result = Result{r.score, pose.cast_double(), p.rotational_scores[r.scan_index], r.low_resolution_score};
  }
  return answers;
}

}  // namespace scan_matching
}  // namespace mapping
}  // namespace cartographer
