#include "cartographer/mapping/internal/constraints/constraint_builder_2d.h"

#include "cartographer/mapping/internal/2d/scan_matching/ceres_scan_matcher_2d.h"

#include <cmath>
#include <cstdio>
#include <thread>
#include <utility>

namespace cartographer {
namespace mapping {
namespace constraints {

using scan_matching::DeviceContext;
using scan_matching::FastCorrelativeScanMatcher2D;

transform::Rigid2d ComputeSubmapPose(const Submap2D& submap) {
  return transform::Project2D(submap.local_pose());
}

ConstraintBuilder2D::ConstraintBuilder2D(const proto::ConstraintBuilderOptions& options,
                                         common::ThreadPoolInterface* const thread_pool)
    : ConstraintBuilder2D(options, thread_pool, std::vector<int>{0}) {}

ConstraintBuilder2D::ConstraintBuilder2D(const proto::ConstraintBuilderOptions& options,
                                         common::ThreadPoolInterface* const thread_pool,
                                         const std::vector<int>& devices)
    : options_(options), thread_pool_(thread_pool), when_done_task_(new common::Task) {
  CSM_CHECK(!devices.empty());
  for (const int device : devices) devices_.emplace_back(new DeviceContext(device));
}

ConstraintBuilder2D::~ConstraintBuilder2D() {
  std::lock_guard<std::mutex> locker(mutex_);
  CSM_CHECK(when_done_task_->GetState() == common::Task::NEW);
  CSM_CHECK(constraints_.size() == 0);  // WhenDone() was not called
  CSM_CHECK(num_started_nodes_ == num_finished_nodes_);
  CSM_CHECK(when_done_ == nullptr);
}

void ConstraintBuilder2D::Enqueue(const SubmapId& submap_id, const Submap2D* const submap,
                                  const NodeId& node_id, const bool match_full_submap,
                                  const TrajectoryNode::Data* const constant_data,
                                  const transform::Rigid2d& initial_relative_pose) {
  if (when_done_) {
    std::fprintf(stderr, "W MaybeAdd(Global)Constraint was called while WhenDone was scheduled.\n");
  }
  CSM_CHECK(submap->grid() != nullptr);
  constraints_.emplace_back();
  // Get or create the matcher slot (reference: DispatchScanMatcherConstruction,
  // .cc:165-186).  The stack itself is built inside the flush, on the thread that
  // owns the chosen device's context.
  auto it = submap_scan_matchers_.find(submap_id);
  if (it == submap_scan_matchers_.end()) {
    SubmapScanMatcher m;
    m.grid = submap->grid();
    m.device_slot = next_device_slot_;
    next_device_slot_ = (next_device_slot_ + 1) % static_cast<int>(devices_.size());
    submap_scan_matchers_.emplace(submap_id, std::move(m));
  }
  pending_.push_back(PendingSearch{submap_id, submap, node_id, match_full_submap, constant_data,
                                   initial_relative_pose, &constraints_.back()});
}

void ConstraintBuilder2D::MaybeAddConstraint(const SubmapId& submap_id, const Submap2D* const submap,
                                             const NodeId& node_id,
                                             const TrajectoryNode::Data* const constant_data,
                                             const transform::Rigid2d& initial_relative_pose) {
  if (initial_relative_pose.translation().norm() > options_.max_constraint_distance()) return;
  std::lock_guard<std::mutex> locker(mutex_);
  if (!per_submap_sampler_
           .emplace(std::piecewise_construct, std::forward_as_tuple(submap_id),
                    std::forward_as_tuple(options_.sampling_ratio()))
           .first->second.Pulse()) {
    return;
  }
  Enqueue(submap_id, submap, node_id, false, constant_data, initial_relative_pose);
}

void ConstraintBuilder2D::MaybeAddGlobalConstraint(const SubmapId& submap_id,
                                                   const Submap2D* const submap,
                                                   const NodeId& node_id,
                                                   const TrajectoryNode::Data* const constant_data) {
  std::lock_guard<std::mutex> locker(mutex_);
  Enqueue(submap_id, submap, node_id, true, constant_data, transform::Rigid2d::Identity());
}

std::weak_ptr<common::Task> ConstraintBuilder2D::ScheduleFlush() {
  auto searches = std::make_shared<std::vector<PendingSearch>>(std::move(pending_));
  pending_.clear();
  const int finished_after = num_started_nodes_;
  auto task = std::unique_ptr<common::Task>(new common::Task);
  task->SetWorkItem([this, searches, finished_after]() { RunFlush(*searches, finished_after); });
  // Flushes run one at a time, in order: a device context serves one host thread.
  task->AddDependency(last_flush_);
  last_flush_ = thread_pool_->Schedule(std::move(task));
  when_done_task_->AddDependency(last_flush_);
  ++stats_.flushes;
  return last_flush_;
}

void ConstraintBuilder2D::NotifyEndOfNode() {
  std::lock_guard<std::mutex> locker(mutex_);
  ++num_started_nodes_;
  if (static_cast<int>(pending_.size()) >= flush_threshold_) ScheduleFlush();
}

void ConstraintBuilder2D::WhenDone(const std::function<void(const Result&)>& callback) {
  std::lock_guard<std::mutex> locker(mutex_);
  CSM_CHECK(when_done_ == nullptr);
  when_done_.reset(new std::function<void(const Result&)>(callback));
  CSM_CHECK(when_done_task_ != nullptr);
  ScheduleFlush();  // also advances num_finished_nodes_ when nothing is pending
  when_done_task_->SetWorkItem([this] { RunWhenDoneCallback(); });
  thread_pool_->Schedule(std::move(when_done_task_));
  when_done_task_.reset(new common::Task);
}

void ConstraintBuilder2D::RunFlush(const std::vector<PendingSearch>& searches,
                                   const int finished_nodes_after) {
  const int num_devices = static_cast<int>(devices_.size());
  // Shard by submap affinity; keep submission order inside a shard.
  std::vector<std::vector<size_t>> shard(num_devices);
  std::vector<SubmapScanMatcher> matcher_of(searches.size());
  {
    std::lock_guard<std::mutex> locker(mutex_);
    for (size_t i = 0; i != searches.size(); ++i) {
      auto it = submap_scan_matchers_.find(searches[i].submap_id);
      if (it == submap_scan_matchers_.end()) {  // deleted meanwhile: recreate the slot
        SubmapScanMatcher m;
        m.grid = searches[i].submap->grid();
        m.device_slot = next_device_slot_;
        next_device_slot_ = (next_device_slot_ + 1) % num_devices;
        it = submap_scan_matchers_.emplace(searches[i].submap_id, std::move(m)).first;
      }
      matcher_of[i] = it->second;
      shard[it->second.device_slot].push_back(i);
    }
  }
  std::vector<FastCorrelativeScanMatcher2D::Answer> answers(searches.size());
  std::vector<transform::Rigid2d> initial_pose(searches.size());
  auto run_shard = [&](const int slot) {
    DeviceContext* ctx = devices_[slot].get();
    // Build the stacks this shard needs and does not have yet (reference: one pool
    // task per submap, FastCorrelativeScanMatcher2D constructor).
    std::map<SubmapId, std::shared_ptr<FastCorrelativeScanMatcher2D>> built;
    FastCorrelativeScanMatcher2D::QueryList queries;
    for (const size_t i : shard[slot]) {
      const PendingSearch& s = searches[i];
      std::shared_ptr<FastCorrelativeScanMatcher2D>& m = matcher_of[i].fast_correlative_scan_matcher;
      if (!m) {
        auto b = built.find(s.submap_id);
        if (b == built.end()) {
          b = built.emplace(s.submap_id,
                            std::make_shared<FastCorrelativeScanMatcher2D>(
                                *matcher_of[i].grid, options_.fast_correlative_scan_matcher_options(),
                                ctx)).first;
        }
        m = b->second;
      }
      FastCorrelativeScanMatcher2D::Query q;
      q.point_cloud = &s.constant_data->filtered_gravity_aligned_point_cloud;
      q.full_submap = s.match_full_submap;
      // initial_pose = ComputeSubmapPose(*submap) * initial_relative_pose (.cc:197-198)
      initial_pose[i] = ComputeSubmapPose(*s.submap) * s.initial_relative_pose;
      q.initial_pose_estimate = initial_pose[i];
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
    const auto shard_answers = FastCorrelativeScanMatcher2D::MatchBatch(ctx, queries);
    for (size_t k = 0; k != shard[slot].size(); ++k) answers[shard[slot][k]] = shard_answers[k];
    // Refinement of every found pose, one more device batch on the same context
    // (reference .cc:245-249: "Use the CSM estimate as both the initial and previous pose").
    if (options_.ceres_scan_matcher_options().occupied_space_weight() > 0.) {
      std::vector<scan_matching::CeresScanMatcher2D::Query> refine;
      std::vector<size_t> which;
      for (const size_t i : shard[slot]) {
        if (!answers[i].found) continue;
        scan_matching::CeresScanMatcher2D::Query q;
        q.target_translation = answers[i].pose_estimate.translation();
        q.initial_pose_estimate = answers[i].pose_estimate;
        q.point_cloud = &searches[i].constant_data->filtered_gravity_aligned_point_cloud;
        q.grid = matcher_of[i].grid;
        refine.push_back(q);
        which.push_back(i);
      }
      if (!refine.empty()) {
        const scan_matching::CeresScanMatcher2D ceres_scan_matcher(options_.ceres_scan_matcher_options(),
                                                                   ctx);
        const auto refined = ceres_scan_matcher.MatchBatch(refine);
        for (size_t k = 0; k != which.size(); ++k) answers[which[k]].pose_estimate = refined[k].pose_estimate;
      }
    }
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
    if (!a.found) continue;  // below-threshold searches leave a null slot (.cc:225-241)
    ++found;
    // a.pose_estimate is already refined (run_shard) when the Ceres options are set.
    // constraint_transform = ComputeSubmapPose(*submap).inverse() * pose_estimate
    const transform::Rigid2d constraint_transform =
        ComputeSubmapPose(*s.submap).inverse() * a.pose_estimate;
    s.constraint->reset(new Constraint{s.submap_id,
                                       s.node_id,
                                       {transform::Embed3D(constraint_transform),
                                        options_.loop_closure_translation_weight(),
                                        options_.loop_closure_rotation_weight()},
                                       Constraint::INTER_SUBMAP});
    if (options_.log_matches()) {
      std::fprintf(stderr, "I Node (%d, %d) with %zu points on submap (%d, %d) %s with score %.1f%%.\n",
                   s.node_id.trajectory_id, s.node_id.node_index,
                   s.constant_data->filtered_gravity_aligned_point_cloud.size(),
                   s.submap_id.trajectory_id, s.submap_id.submap_index,
                   s.match_full_submap ? "matches" : "matches locally", 100. * a.score);
    }
  }
  std::lock_guard<std::mutex> locker(mutex_);
  stats_.searches += searches.size();
  stats_.found += found;
  stats_.candidates_scored += scored;
  if (finished_nodes_after > num_finished_nodes_) num_finished_nodes_ = finished_nodes_after;
}

void ConstraintBuilder2D::RunWhenDoneCallback() {
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

int ConstraintBuilder2D::GetNumFinishedNodes() {
  std::lock_guard<std::mutex> locker(mutex_);
  return num_finished_nodes_;
}

void ConstraintBuilder2D::DeleteScanMatcher(const SubmapId& submap_id) {
  std::lock_guard<std::mutex> locker(mutex_);
  if (when_done_) {
    std::fprintf(stderr, "W DeleteScanMatcher was called while WhenDone was scheduled.\n");
  }
  submap_scan_matchers_.erase(submap_id);
  per_submap_sampler_.erase(submap_id);
}

void ConstraintBuilder2D::RegisterMetrics(metrics::FamilyFactory*) {
  // Metrics families are owned by the embedding process; this standalone build keeps
  // only the counters behind stats().
}

ConstraintBuilder2D::Stats ConstraintBuilder2D::stats() {
  std::lock_guard<std::mutex> locker(mutex_);
  return stats_;
}

}  // namespace constraints
}  // namespace mapping
}  // namespace cartographer
