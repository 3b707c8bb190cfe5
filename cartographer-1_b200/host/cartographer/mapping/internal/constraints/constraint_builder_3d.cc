#include "cartographer/mapping/internal/constraints/constraint_builder_3d.h"

#include <cmath>
#include <cstdio>
#include <thread>
#include <utility>

namespace cartographer {
namespace mapping {
namespace constraints {

using scan_matching::DeviceContext;
using scan_matching::FastCorrelativeScanMatcher3D;

ConstraintBuilder3D::ConstraintBuilder3D(const proto::ConstraintBuilderOptions& options,
                                         common::ThreadPoolInterface* const thread_pool)
    : ConstraintBuilder3D(options, thread_pool, std::vector<int>{0}) {}

ConstraintBuilder3D::ConstraintBuilder3D(const proto::ConstraintBuilderOptions& options,
                                         common::ThreadPoolInterface* const thread_pool,
                                         const std::vector<int>& devices)
    : options_(options), thread_pool_(thread_pool), when_done_task_(new common::Task) {
  CSM_CHECK(!devices.empty());
  for (const int device : devices) devices_.emplace_back(new DeviceContext(device));
}

ConstraintBuilder3D::~ConstraintBuilder3D() {
  std::lock_guard<std::mutex> locker(mutex_);
  CSM_CHECK(when_done_task_->GetState() == common::Task::NEW);
  CSM_CHECK(constraints_.size() == 0);  // WhenDone() was not called
  CSM_CHECK(num_started_nodes_ == num_finished_nodes_);
  CSM_CHECK(when_done_ == nullptr);
}

// reference: DispatchScanMatcherConstruction (.cc:177-198); the stack itself is built
// inside the flush, on the thread that drives the chosen device.
ConstraintBuilder3D::SubmapScanMatcher* ConstraintBuilder3D::GetOrCreateSlot(const SubmapId& submap_id,
                                                                             const Submap3D* const submap) {
  auto it = submap_scan_matchers_.find(submap_id);
  if (it == submap_scan_matchers_.end()) {
    SubmapScanMatcher m;
    m.high_resolution_hybrid_grid = &submap->high_resolution_hybrid_grid();
    m.low_resolution_hybrid_grid = &submap->low_resolution_hybrid_grid();
    m.rotational_scan_matcher_histogram = &submap->rotational_scan_matcher_histogram();
    m.device_slot = next_device_slot_;
    next_device_slot_ = (next_device_slot_ + 1) % static_cast<int>(devices_.size());
    it = submap_scan_matchers_.emplace(submap_id, std::move(m)).first;
  }
  return &it->second;
}

void ConstraintBuilder3D::Enqueue(const SubmapId& submap_id, const Submap3D* const submap,
                                  const NodeId& node_id, const bool match_full_submap,
                                  const TrajectoryNode::Data* const constant_data,
                                  const transform::Rigid3d& global_node_pose,
                                  const transform::Rigid3d& global_submap_pose) {
  if (when_done_) {
    std::fprintf(stderr, "W MaybeAdd(Global)Constraint was called while WhenDone was scheduled.\n");
  }
  constraints_.emplace_back();
  GetOrCreateSlot(submap_id, submap);
  pending_.push_back(PendingSearch{submap_id, submap, node_id, match_full_submap, constant_data,
                                   global_node_pose, global_submap_pose, &constraints_.back()});
}

void ConstraintBuilder3D::MaybeAddConstraint(const SubmapId& submap_id, const Submap3D* const submap,
                                             const NodeId& node_id,
                                             const TrajectoryNode::Data* const constant_data,
                                             const transform::Rigid3d& global_node_pose,
                                             const transform::Rigid3d& global_submap_pose) {
  const double dx = global_node_pose.translation().x() - global_submap_pose.translation().x();
  const double dy = global_node_pose.translation().y() - global_submap_pose.translation().y();
  const double dz = global_node_pose.translation().z() - global_submap_pose.translation().z();
  if (std::sqrt(dx * dx + dy * dy + dz * dz) > options_.max_constraint_distance()) return;
  std::lock_guard<std::mutex> locker(mutex_);
  if (!per_submap_sampler_
           .emplace(std::piecewise_construct, std::forward_as_tuple(submap_id),
                    std::forward_as_tuple(options_.sampling_ratio()))
           .first->second.Pulse()) {
    return;
  }
  Enqueue(submap_id, submap, node_id, false, constant_data, global_node_pose, global_submap_pose);
}

void ConstraintBuilder3D::MaybeAddGlobalConstraint(const SubmapId& submap_id, const Submap3D* const submap,
                                                   const NodeId& node_id,
                                                   const TrajectoryNode::Data* const constant_data,
                                                   const Eigen::Quaterniond& global_node_rotation,
                                                   const Eigen::Quaterniond& global_submap_rotation) {
  std::lock_guard<std::mutex> locker(mutex_);
  Enqueue(submap_id, submap, node_id, true, constant_data,
          transform::Rigid3d(transform::Vector3d(), global_node_rotation),
          transform::Rigid3d(transform::Vector3d(), global_submap_rotation));
}

void ConstraintBuilder3D::MaybeAddConstraints(const std::vector<ConstraintRequest>& requests) {
  std::lock_guard<std::mutex> locker(mutex_);
  for (const ConstraintRequest& r : requests) {
    if (r.match_full_submap) {
      Enqueue(r.submap_id, r.submap, r.node_id, true, r.constant_data,
              transform::Rigid3d(transform::Vector3d(), r.global_node_pose.rotation()),
              transform::Rigid3d(transform::Vector3d(), r.global_submap_pose.rotation()));
      continue;
    }
    const double dx = r.global_node_pose.translation().x() - r.global_submap_pose.translation().x();
    const double dy = r.global_node_pose.translation().y() - r.global_submap_pose.translation().y();
    const double dz = r.global_node_pose.translation().z() - r.global_submap_pose.translation().z();
    if (std::sqrt(dx * dx + dy * dy + dz * dz) > options_.max_constraint_distance()) continue;
    if (!per_submap_sampler_
             .emplace(std::piecewise_construct, std::forward_as_tuple(r.submap_id),
                      std::forward_as_tuple(options_.sampling_ratio()))
             .first->second.Pulse()) {
      continue;
    }
    Enqueue(r.submap_id, r.submap, r.node_id, false, r.constant_data, r.global_node_pose, r.global_submap_pose);
  }
  if (static_cast<int>(pending_.size()) >= flush_threshold_) ScheduleFlush();
}

std::weak_ptr<common::Task> ConstraintBuilder3D::ScheduleFlush() {
  auto searches = std::make_shared<std::vector<PendingSearch>>(std::move(pending_));
  pending_.clear();
  const int finished_after = num_started_nodes_;
  auto task = std::unique_ptr<common::Task>(new common::Task);
  task->SetWorkItem([this, searches, finished_after]() { RunFlush(*searches, finished_after); });
  task->AddDependency(last_flush_);  // one flush at a time, in order
  last_flush_ = thread_pool_->Schedule(std::move(task));
  when_done_task_->AddDependency(last_flush_);
  ++stats_.flushes;
  return last_flush_;
}

void ConstraintBuilder3D::NotifyEndOfNode() {
  std::lock_guard<std::mutex> locker(mutex_);
  ++num_started_nodes_;
  if (static_cast<int>(pending_.size()) >= flush_threshold_) ScheduleFlush();
}

void ConstraintBuilder3D::WhenDone(const std::function<void(const Result&)>& callback) {
  std::lock_guard<std::mutex> locker(mutex_);
  CSM_CHECK(when_done_ == nullptr);
  when_done_.reset(new std::function<void(const Result&)>(callback));
  CSM_CHECK(when_done_task_ != nullptr);
  ScheduleFlush();
  when_done_task_->SetWorkItem([this] { RunWhenDoneCallback(); });
  thread_pool_->Schedule(std::move(when_done_task_));
  when_done_task_.reset(new common::Task);
}

void ConstraintBuilder3D::RunFlush(const std::vector<PendingSearch>& searches, const int finished_nodes_after) {
  const int num_devices = static_cast<int>(devices_.size());
  std::vector<std::vector<size_t>> shard(num_devices);
  std::vector<SubmapScanMatcher> matcher_of(searches.size());
  {
    std::lock_guard<std::mutex> locker(mutex_);
    for (size_t i = 0; i != searches.size(); ++i) {
      const SubmapScanMatcher* slot = GetOrCreateSlot(searches[i].submap_id, searches[i].submap);
      matcher_of[i] = *slot;
      shard[slot->device_slot].push_back(i);
    }
  }
  std::vector<FastCorrelativeScanMatcher3D::Answer> answers(searches.size());
  auto run_shard = [&](const int slot) {
    DeviceContext* ctx = devices_[slot].get();
    std::map<SubmapId, std::shared_ptr<FastCorrelativeScanMatcher3D>> built;
    FastCorrelativeScanMatcher3D::QueryList queries;
    for (const size_t i : shard[slot]) {
      const PendingSearch& s = searches[i];
      std::shared_ptr<FastCorrelativeScanMatcher3D>& m = matcher_of[i].fast_correlative_scan_matcher;
      if (!m) {
        auto b = built.find(s.submap_id);
        if (b == built.end()) {
          b = built.emplace(s.submap_id,
                            std::make_shared<FastCorrelativeScanMatcher3D>(
                                *matcher_of[i].high_resolution_hybrid_grid,
                                matcher_of[i].low_resolution_hybrid_grid,
                                matcher_of[i].rotational_scan_matcher_histogram,
                                options_.fast_correlative_scan_matcher_options_3d(), ctx)).first;
        }
        m = b->second;
      }
      FastCorrelativeScanMatcher3D::Query q;
      q.constant_data = s.constant_data;
      q.full_submap = s.match_full_submap;
      q.global_node_pose = s.global_node_pose;
      q.global_submap_pose = s.global_submap_pose;
      q.min_score = static_cast<float>(s.match_full_submap ? options_.global_localization_min_score()
                                                           : options_.min_score());
      queries.emplace_back(m.get(), q);
    }
    if (!built.empty()) {
      std::lock_guard<std::mutex> locker(mutex_);
      for (auto& kv : built) {
        auto it = submap_scan_matchers_.find(kv.first);
        if (it != submap_scan_matchers_.end() && !it->second.fast_correlative_scan_matcher)
          it->second.fast_correlative_scan_matcher = kv.second;
      }
    }
    if (queries.empty()) return;
    const auto shard_answers = FastCorrelativeScanMatcher3D::MatchBatch(ctx, queries);
    for (size_t k = 0; k != shard[slot].size(); ++k) answers[shard[slot][k]] = shard_answers[k];
  };
  std::vector<std::thread> workers;
  for (int slot = 1; slot < num_devices; ++slot)
    if (!shard[slot].empty()) workers.emplace_back(run_shard, slot);
  run_shard(0);
  for (std::thread& t : workers) t.join();

  uint64_t found = 0, scored = 0;
  for (size_t i = 0; i != searches.size(); ++i) {
    const PendingSearch& s = searches[i];
    const auto& a = answers[i];
    scored += a.candidates_scored;
    if (!a.found) continue;  // below-threshold searches leave a null slot (reference :231,:249)
    ++found;
    // [Ceres hook] the reference refines match_result->pose_estimate over both grids here
    // (ceres_scan_matcher_.Match, :264-275); outside this path and absent from this image.
    const transform::Rigid3d constraint_transform = a.result.pose_estimate;
    s.constraint->reset(new Constraint{s.submap_id,
                                       s.node_id,
                                       {constraint_transform, options_.loop_closure_translation_weight(),
                                        options_.loop_closure_rotation_weight()},
                                       Constraint::INTER_SUBMAP});
    if (options_.log_matches()) {
      std::fprintf(stderr,
                   "I Node (%d, %d) with %zu points on submap (%d, %d) %s with score %.1f%% rotational "
                   "score %.1f%% low resolution score %.1f%%.\n",
                   s.node_id.trajectory_id, s.node_id.node_index,
                   s.constant_data->high_resolution_point_cloud.size(), s.submap_id.trajectory_id,
                   s.submap_id.submap_index, s.match_full_submap ? "matches" : "matches locally",
                   100. * a.result.score, 100. * a.result.rotational_score,
                   100. * a.result.low_resolution_score);
    }
  }
  std::lock_guard<std::mutex> locker(mutex_);
  stats_.searches += searches.size();
  stats_.found += found;
  stats_.candidates_scored += scored;
  if (finished_nodes_after > num_finished_nodes_) num_finished_nodes_ = finished_nodes_after;
}

void ConstraintBuilder3D::RunWhenDoneCallback() {
  Result result;
  std::unique_ptr<std::function<void(const Result&)>> callback;
  {
    std::lock_guard<std::mutex> locker(mutex_);
    CSM_CHECK(when_done_ != nullptr);
    for (const std::unique_ptr<Constraint>& constraint : constraints_) {
      if (constraint == nullptr) continue;
      result.push_back(*constraint);
    }
    if (options_.log_matches()) {
      std::fprintf(stderr, "I %zu computations resulted in %zu additional constraints.\n",
                   constraints_.size(), result.size());
    }
    constraints_.clear();
    callback = std::move(when_done_);
    when_done_.reset();
  }
  (*callback)(result);
}

int ConstraintBuilder3D::GetNumFinishedNodes() {
  std::lock_guard<std::mutex> locker(mutex_);
  return num_finished_nodes_;
}

void ConstraintBuilder3D::DeleteScanMatcher(const SubmapId& submap_id) {
  std::lock_guard<std::mutex> locker(mutex_);
  if (when_done_) {
    std::fprintf(stderr, "W DeleteScanMatcher was called while WhenDone was scheduled.\n");
  }
  submap_scan_matchers_.erase(submap_id);
  per_submap_sampler_.erase(submap_id);
}

void ConstraintBuilder3D::RegisterMetrics(metrics::FamilyFactory*) {}

ConstraintBuilder3D::Stats ConstraintBuilder3D::stats() {
  std::lock_guard<std::mutex> locker(mutex_);
  return stats_;
}

}  // namespace constraints
}  // namespace mapping
}  // namespace cartographer
