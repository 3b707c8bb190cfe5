#include "cartographer/mapping/internal/constraints/constraint_builder_2d.h"

#include "cartographer/mapping/internal/2d/scan_matching/ceres_scan_matcher_2d.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <numeric>
#include <thread>
#include <utility>

namespace cartographer {
namespace mapping {
namespace constraints {

using scan_matching::DeviceContext;
using scan_matching::FastCorrelativeScanMatcher2D;

// The reference's metric handles (constraint_builder_2d.cc:46-53): no-ops until
// RegisterMetrics is given a factory.
static auto* kConstraintsSearchedMetric = metrics::Counter::Null();
static auto* kConstraintsFoundMetric = metrics::Counter::Null();
static auto* kGlobalConstraintsSearchedMetric = metrics::Counter::Null();
static auto* kGlobalConstraintsFoundMetric = metrics::Counter::Null();
static auto* kQueueLengthMetric = metrics::Gauge::Null();
static auto* kConstraintScoresMetric = metrics::Histogram::Null();
static auto* kGlobalConstraintScoresMetric = metrics::Histogram::Null();
static auto* kNumSubmapScanMatchersMetric = metrics::Gauge::Null();

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
  stats_.searches_per_device.assign(devices.size(), 0);
  stats_.cost_per_device.assign(devices.size(), 0.);
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
    m.per_device.resize(devices_.size());
    submap_scan_matchers_.emplace(submap_id, std::move(m));
    kNumSubmapScanMatchersMetric->Set(submap_scan_matchers_.size());
  }
  pending_.push_back(PendingSearch{submap_id, submap, node_id, match_full_submap, constant_data,
                                   initial_relative_pose, &constraints_.back()});
  kQueueLengthMetric->Set(constraints_.size());
}

void ConstraintBuilder2D::MaybeAddConstraints(const std::vector<ConstraintRequest>& requests) {
  std::lock_guard<std::mutex> locker(mutex_);
  for (const ConstraintRequest& r : requests) {
    if (r.match_full_submap) {
      Enqueue(r.submap_id, r.submap, r.node_id, true, r.constant_data, transform::Rigid2d::Identity());
      continue;
    }
    if (r.initial_relative_pose.translation().norm() > options_.max_constraint_distance()) continue;
    if (!per_submap_sampler_
             .emplace(std::piecewise_construct, std::forward_as_tuple(r.submap_id),
                      std::forward_as_tuple(options_.sampling_ratio()))
             .first->second.Pulse()) {
      continue;
    }
    Enqueue(r.submap_id, r.submap, r.node_id, false, r.constant_data, r.initial_relative_pose);
  }
  // One flush decision for the whole burst: a burst larger than the threshold goes to the
  // devices right away instead of waiting for the caller's next NotifyEndOfNode.
  if (static_cast<int>(pending_.size()) >= flush_threshold_) ScheduleFlush();
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

double ConstraintBuilder2D::EstimateCost(const PendingSearch& s) const {
  // points x angular slices x lowest-resolution lattice positions (what k_score_top scores;
  // the sweep below it scales with the same product).
  const sensor::PointCloud& cloud = s.constant_data->filtered_gravity_aligned_point_cloud;
  const MapLimits& limits = s.submap->grid()->limits();
  const double res = limits.resolution();
  double reach = 3. * res;
  for (size_t p = 0; p != cloud.size(); ++p)
    reach = std::max(reach, std::hypot(static_cast<double>(cloud[p].position.x()),
                                       static_cast<double>(cloud[p].position.y())));
  const auto& o = options_.fast_correlative_scan_matcher_options();
  const double angular = s.match_full_submap ? M_PI : o.angular_search_window();
  const double slices = 2. * std::ceil(angular / std::max(1e-9, res / reach)) + 1.;
  const double step = static_cast<double>(1 << std::max(0, o.branch_and_bound_depth() - 1));
  const double ext = 2. * reach / res;
  const double lin = s.match_full_submap ? 1e9 : 2. * o.linear_search_window() / res;
  const double wx = std::min(lin, limits.cell_limits().num_x_cells + ext);
  const double wy = std::min(lin, limits.cell_limits().num_y_cells + ext);
  return std::max<double>(1., cloud.size()) * slices * (wx / step + 1.) * (wy / step + 1.);
}

std::vector<int> ConstraintBuilder2D::AssignDevices(const std::vector<PendingSearch>& searches,
                                                    std::vector<double>* cost_out) {
  const int num_devices = static_cast<int>(devices_.size());
  std::vector<int> slot(searches.size(), 0);
  std::vector<double>& cost = *cost_out;
  cost.resize(searches.size());
  for (size_t i = 0; i != searches.size(); ++i) cost[i] = EstimateCost(searches[i]);
  if (num_devices == 1) return slot;
  // Units = the submaps of this flush with their searches (submission order inside a unit).
  struct Unit {
    SubmapScanMatcher* matcher;
    std::vector<size_t> searches;
    double cost = 0.;
  };
  std::map<SubmapId, size_t> unit_of;
  std::vector<Unit> units;
  for (size_t i = 0; i != searches.size(); ++i) {
    auto found = unit_of.find(searches[i].submap_id);
    if (found == unit_of.end()) {
      found = unit_of.emplace(searches[i].submap_id, units.size()).first;
      units.emplace_back();
      units.back().matcher = &submap_scan_matchers_.at(searches[i].submap_id);
    }
    units[found->second].searches.push_back(i);
    units[found->second].cost += cost[i];
  }
  // Greedy longest-processing-time with affinity: the most expensive submap first, onto the
  // device that finishes it earliest, where a device that does not hold the stack yet pays for
  // building it (about a quarter of one search at the default depth).
  std::vector<size_t> order(units.size());
  std::iota(order.begin(), order.end(), size_t{0});
  std::stable_sort(order.begin(), order.end(), [&](size_t a, size_t b) { return units[a].cost > units[b].cost; });
  std::vector<double> load(num_devices, 0.);
  std::vector<int> unit_slot(units.size(), 0);
  for (const size_t u : order) {
    const double build = 0.25 * units[u].cost / units[u].searches.size();
    int best = 0;
    double best_finish = 0.;
    for (int d = 0; d < num_devices; ++d) {
      const double finish = load[d] + units[u].cost + (units[u].matcher->per_device[d] ? 0. : build);
      if (d == 0 || finish < best_finish) {
        best = d;
        best_finish = finish;
      }
    }
    unit_slot[u] = best;
    load[best] = best_finish;
    for (const size_t i : units[u].searches) slot[i] = best;
  }
  // A hot submap: the most loaded device hands the tail of one submap's node list to the
  // least loaded one (which replicates the stack) until the spread is within 5 % of the mean.
  const double mean = std::accumulate(load.begin(), load.end(), 0.) / num_devices;
  for (int round = 0; round < 4 * num_devices; ++round) {
    const int hi = static_cast<int>(std::max_element(load.begin(), load.end()) - load.begin());
    const int lo = static_cast<int>(std::min_element(load.begin(), load.end()) - load.begin());
    if (load[hi] - load[lo] <= 0.05 * mean) break;
    const double want = 0.5 * (load[hi] - load[lo]);
    size_t pick = units.size();
    double pick_cost = 0.;
    for (size_t u = 0; u != units.size(); ++u) {
      double on_hi = 0.;
      for (const size_t i : units[u].searches)
        if (slot[i] == hi) on_hi += cost[i];
      if (on_hi > pick_cost) {
        pick = u;
        pick_cost = on_hi;
      }
    }
    if (pick == units.size()) break;
    double moved = 0.;
    for (auto it = units[pick].searches.rbegin(); it != units[pick].searches.rend(); ++it) {
      if (slot[*it] != hi) continue;
      if (moved > 0. && moved + cost[*it] > want) break;
      if (moved + cost[*it] > 2. * want) break;
      slot[*it] = lo;
      moved += cost[*it];
    }
    if (moved == 0.) break;
    load[hi] -= moved;
    load[lo] += moved;
  }
  return slot;
}

void ConstraintBuilder2D::RunFlush(const std::vector<PendingSearch>& searches,
                                   const int finished_nodes_after) {
  const int num_devices = static_cast<int>(devices_.size());
  // Shard by submap affinity and estimated cost; keep submission order inside a shard.
  std::vector<std::vector<size_t>> shard(num_devices);
  std::vector<const Grid2D*> grid_of(searches.size());
  std::vector<std::shared_ptr<FastCorrelativeScanMatcher2D>> matcher_of(searches.size());
  std::vector<double> cost;
  {
    std::lock_guard<std::mutex> locker(mutex_);
    for (size_t i = 0; i != searches.size(); ++i) {
      auto it = submap_scan_matchers_.find(searches[i].submap_id);
      if (it == submap_scan_matchers_.end()) {  // deleted meanwhile: recreate the slot
        SubmapScanMatcher m;
        m.grid = searches[i].submap->grid();
        m.per_device.resize(num_devices);
        submap_scan_matchers_.emplace(searches[i].submap_id, std::move(m));
      }
    }
    const std::vector<int> slot = AssignDevices(searches, &cost);
    for (size_t i = 0; i != searches.size(); ++i) {
      const SubmapScanMatcher& m = submap_scan_matchers_.at(searches[i].submap_id);
      grid_of[i] = m.grid;
      matcher_of[i] = m.per_device[slot[i]];
      shard[slot[i]].push_back(i);
      stats_.searches_per_device[slot[i]] += 1;
      stats_.cost_per_device[slot[i]] += cost[i];
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
      std::shared_ptr<FastCorrelativeScanMatcher2D>& m = matcher_of[i];
      if (!m) {
        auto b = built.find(s.submap_id);
        if (b == built.end()) {
          b = built.emplace(s.submap_id,
                            std::make_shared<FastCorrelativeScanMatcher2D>(
                                *grid_of[i], options_.fast_correlative_scan_matcher_options(), ctx)).first;
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
        if (it == submap_scan_matchers_.end() || it->second.per_device[slot]) continue;
        for (const auto& other : it->second.per_device)
          if (other) {
            ++stats_.replicated_stacks;
            break;
          }
        it->second.per_device[slot] = kv.second;
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
        q.grid = grid_of[i];
        // The submap's stack keeps the grid's cells on this device: the refinement reads them there.
        if (matcher_of[i] && matcher_of[i]->device() == ctx->device()) q.device_cells = matcher_of[i]->device_cells();
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

  uint64_t found = 0, scored = 0, local_searched = 0, local_found = 0, global_searched = 0, global_found = 0;
  std::vector<float> found_scores;
  for (size_t i = 0; i != searches.size(); ++i) {
    const PendingSearch& s = searches[i];
    const auto& a = answers[i];
    scored += a.candidates_scored;
    // the reference's counters and score histograms (.cc:205-240)
    if (s.match_full_submap) {
      ++global_searched;
      kGlobalConstraintsSearchedMetric->Increment();
    } else {
      ++local_searched;
      kConstraintsSearchedMetric->Increment();
    }
    if (!a.found) continue;  // below-threshold searches leave a null slot (.cc:225-241)
    ++found;
    found_scores.push_back(a.score);
    if (s.match_full_submap) {
      ++global_found;
      kGlobalConstraintsFoundMetric->Increment();
      kGlobalConstraintScoresMetric->Observe(a.score);
    } else {
      ++local_found;
      kConstraintsFoundMetric->Increment();
      kConstraintScoresMetric->Observe(a.score);
    }
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
  stats_.local_searched += local_searched;
  stats_.local_found += local_found;
  stats_.global_searched += global_searched;
  stats_.global_found += global_found;
  for (const float score : found_scores) score_histogram_.Add(score);
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
      std::fprintf(stderr, "I Score histogram:\n%s\n", score_histogram_.ToString(10).c_str());
    }
    constraints_.clear();
    callback = std::move(when_done_);
    when_done_.reset();
    kQueueLengthMetric->Set(constraints_.size());
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
  kNumSubmapScanMatchersMetric->Set(submap_scan_matchers_.size());
}

// Same families, names and labels as the reference (constraint_builder_2d.cc:318-343).
void ConstraintBuilder2D::RegisterMetrics(metrics::FamilyFactory* factory) {
  auto* counts = factory->NewCounterFamily("mapping_constraints_constraint_builder_2d_constraints",
                                           "Constraints computed");
  kConstraintsSearchedMetric = counts->Add({{"search_region", "local"}, {"matcher", "searched"}});
  kConstraintsFoundMetric = counts->Add({{"search_region", "local"}, {"matcher", "found"}});
  kGlobalConstraintsSearchedMetric = counts->Add({{"search_region", "global"}, {"matcher", "searched"}});
  kGlobalConstraintsFoundMetric = counts->Add({{"search_region", "global"}, {"matcher", "found"}});
  auto* queue_length =
      factory->NewGaugeFamily("mapping_constraints_constraint_builder_2d_queue_length", "Queue length");
  kQueueLengthMetric = queue_length->Add({});
  auto boundaries = metrics::Histogram::FixedWidth(0.05, 20);
  auto* scores = factory->NewHistogramFamily("mapping_constraints_constraint_builder_2d_scores",
                                             "Constraint scores built", boundaries);
  kConstraintScoresMetric = scores->Add({{"search_region", "local"}});
  kGlobalConstraintScoresMetric = scores->Add({{"search_region", "global"}});
  auto* num_matchers =
      factory->NewGaugeFamily("mapping_constraints_constraint_builder_2d_num_submap_scan_matchers",
                              "Current number of constructed submap scan matchers");
  kNumSubmapScanMatchersMetric = num_matchers->Add({});
}

ConstraintBuilder2D::Stats ConstraintBuilder2D::stats() {
  std::lock_guard<std::mutex> locker(mutex_);
  Stats out = stats_;
  out.queue_length = constraints_.size();
  out.num_submap_scan_matchers = submap_scan_matchers_.size();
  return out;
}

std::string ConstraintBuilder2D::ScoreHistogram(const int buckets) {
  std::lock_guard<std::mutex> locker(mutex_);
  return score_histogram_.ToString(buckets);
}

}  // namespace constraints
}  // namespace mapping
}  // namespace cartographer
