// TEST INFRASTRUCTURE ONLY (see csm_oracle.h).
//
// Restatement of the reference's task DAG + thread pool (common/task.{h,cc},
// common/thread_pool.{h,cc}) on std::thread, and of ConstraintBuilder2D
// (mapping/internal/constraints/constraint_builder_2d.{h,cc}) on top of it.  Used
// by the timed CPU baseline so that the baseline has the reference's shape: one
// task per submap that builds its FastCorrelativeScanMatcher2D, one task per
// (node, submap) search depending on it, a finish-node barrier per node and a
// when-done barrier (constraint_builder_2d.cc:77-186,279-305).  The Ceres
// refinement (constraint_builder_2d.cc:245-249) is outside the oracle.
#ifndef ORACLE_CPU_POOL_H_
#define ORACLE_CPU_POOL_H_

#include <unistd.h>

#include <condition_variable>
#include <deque>
#include <functional>
#include <map>
#include <memory>
#include <mutex>
#include <set>
#include <thread>
#include <unordered_map>
#include <vector>

#include "csm_oracle.h"

namespace oracle {

class ThreadPool;

// common/task.{h,cc}
class Task {
 public:
  enum State { NEW, DISPATCHED, DEPENDENCIES_COMPLETED, RUNNING, COMPLETED };
  State GetState() {
    std::lock_guard<std::mutex> l(mu_);
    return state_;
  }
  void SetWorkItem(std::function<void()> w) {
    std::lock_guard<std::mutex> l(mu_);
    ORACLE_CHECK(state_ == NEW);
    work_ = std::move(w);
  }
  void AddDependency(std::weak_ptr<Task> dep) {
    std::shared_ptr<Task> d;
    {
      std::lock_guard<std::mutex> l(mu_);
      ORACLE_CHECK(state_ == NEW);
      if ((d = dep.lock())) ++pending_;
    }
    if (d) d->AddDependent(this);
  }

 private:
  friend class ThreadPool;
  void AddDependent(Task* t) {
    std::unique_lock<std::mutex> l(mu_);
    if (state_ == COMPLETED) {
      l.unlock();
      t->OnDependencyCompleted();
      return;
    }
    ORACLE_CHECK(dependents_.insert(t).second);
  }
  void OnDependencyCompleted();
  void SetThreadPool(ThreadPool* pool);
  void Execute() {
    {
      std::lock_guard<std::mutex> l(mu_);
      ORACLE_CHECK(state_ == DEPENDENCIES_COMPLETED);
      state_ = RUNNING;
    }
    if (work_) work_();
    std::set<Task*> deps;
    {
      std::lock_guard<std::mutex> l(mu_);
      state_ = COMPLETED;
      deps.swap(dependents_);
    }
    for (Task* t : deps) t->OnDependencyCompleted();
  }
  std::mutex mu_;
  State state_ = NEW;
  std::function<void()> work_;
  unsigned pending_ = 0;
  std::set<Task*> dependents_;
  ThreadPool* pool_ = nullptr;
};

// common/thread_pool.{h,cc}
class ThreadPool {
 public:
  explicit ThreadPool(int num_threads) {
    ORACLE_CHECK(num_threads > 0);
    for (int i = 0; i != num_threads; ++i) threads_.emplace_back([this] { DoWork(); });
  }
  ~ThreadPool() {
    {
      std::lock_guard<std::mutex> l(mu_);
      running_ = false;
    }
    cv_.notify_all();
    for (std::thread& t : threads_) t.join();
  }
  std::weak_ptr<Task> Schedule(std::unique_ptr<Task> task) {
    std::shared_ptr<Task> shared;
    {
      std::lock_guard<std::mutex> l(mu_);
      Task* key = task.get();
      auto r = not_ready_.emplace(key, std::shared_ptr<Task>(std::move(task)));
      ORACLE_CHECK(r.second);
      shared = r.first->second;
    }
    shared->SetThreadPool(this);
    return shared;
  }
  void NotifyDependenciesCompleted(Task* task) {
    {
      std::lock_guard<std::mutex> l(mu_);
      auto it = not_ready_.find(task);
      ORACLE_CHECK(it != not_ready_.end());
      queue_.push_back(it->second);
      not_ready_.erase(it);
    }
    cv_.notify_one();
  }

 private:
  void DoWork() {
    // thread_pool.cc:74-79: background work runs at nice(10).
    if (nice(10) == -1) { /* not fatal in a container */ }
    for (;;) {
      std::shared_ptr<Task> task;
      {
        std::unique_lock<std::mutex> l(mu_);
        cv_.wait(l, [this] { return !queue_.empty() || !running_; });
        if (!queue_.empty()) {
          task = std::move(queue_.front());
          queue_.pop_front();
        } else if (!running_) {
          return;
        }
      }
      task->Execute();
    }
  }
  std::mutex mu_;
  std::condition_variable cv_;
  bool running_ = true;
  std::vector<std::thread> threads_;
  std::deque<std::shared_ptr<Task>> queue_;
  std::unordered_map<Task*, std::shared_ptr<Task>> not_ready_;
};

inline void Task::OnDependencyCompleted() {
  ThreadPool* notify = nullptr;
  {
    std::lock_guard<std::mutex> l(mu_);
    ORACLE_CHECK(state_ == NEW || state_ == DISPATCHED);
    --pending_;
    if (pending_ == 0 && state_ == DISPATCHED) {
      state_ = DEPENDENCIES_COMPLETED;
      notify = pool_;
    }
  }
  if (notify) notify->NotifyDependenciesCompleted(this);
}
inline void Task::SetThreadPool(ThreadPool* pool) {
  bool ready = false;
  {
    std::lock_guard<std::mutex> l(mu_);
    ORACLE_CHECK(state_ == NEW);
    state_ = DISPATCHED;
    pool_ = pool;
    if (pending_ == 0) {
      state_ = DEPENDENCIES_COMPLETED;
      ready = true;
    }
  }
  if (ready) pool->NotifyDependenciesCompleted(this);
}

// common/fixed_ratio_sampler.cc:32-39
class FixedRatioSampler {
 public:
  explicit FixedRatioSampler(double ratio) : ratio_(ratio) {}
  bool Pulse() {
    ++num_pulses_;
    if (static_cast<double>(num_samples_) / num_pulses_ < ratio_) {
      ++num_samples_;
      return true;
    }
    return false;
  }

 private:
  double ratio_;
  int64_t num_pulses_ = 0, num_samples_ = 0;
};

struct ConstraintBuilderOptions {
  double sampling_ratio = 0.3;
  double max_constraint_distance = 15.;
  double min_score = 0.55;
  double global_localization_min_score = 0.6;
  double loop_closure_translation_weight = 1.1e4;
  double loop_closure_rotation_weight = 1e5;
  FastCorrelativeScanMatcherOptions2D fast_correlative_scan_matcher_options;
};

// What the reference emits per found constraint, minus the Ceres refinement:
// pose = ComputeSubmapPose(submap)^-1 * pose_estimate (constraint_builder_2d.cc:250-258).
struct OracleConstraint {
  int submap_id = 0, node_id = 0;
  double tx = 0., ty = 0., angle = 0.;  // constraint_transform (submap <- node)
  float score = 0.f;
  FastCorrelativeScanMatcher2D::Winner winner;
};

class ConstraintBuilder2D {
 public:
  ConstraintBuilder2D(const ConstraintBuilderOptions& options, ThreadPool* pool)
      : options_(options),
        pool_(pool),
        finish_node_task_(new Task),
        when_done_task_(new Task) {}

  struct Submap {
    const ProbabilityGrid* grid;
    Rigid2d local_pose;  // Project2D(submap.local_pose())
  };

  // constraint_builder_2d.cc:77-112
  void MaybeAddConstraint(int submap_id, const Submap* submap, int node_id,
                          const PointCloud* cloud, const Rigid2d& initial_relative_pose) {
    if (std::sqrt(initial_relative_pose.tx * initial_relative_pose.tx +
                  initial_relative_pose.ty * initial_relative_pose.ty) >
        options_.max_constraint_distance)
      return;
    if (!samplers_.emplace(submap_id, FixedRatioSampler(options_.sampling_ratio))
             .first->second.Pulse())
      return;
    Add(submap_id, submap, node_id, cloud, false, initial_relative_pose);
  }
  // constraint_builder_2d.cc:114-137
  void MaybeAddGlobalConstraint(int submap_id, const Submap* submap, int node_id,
                                const PointCloud* cloud) {
    Add(submap_id, submap, node_id, cloud, true, Rigid2d());
  }
  // constraint_builder_2d.cc:139-151
  void NotifyEndOfNode() {
    std::lock_guard<std::mutex> l(mu_);
    finish_node_task_->SetWorkItem([this] {
      std::lock_guard<std::mutex> l2(mu_);
      ++num_finished_nodes_;
    });
    auto h = pool_->Schedule(std::move(finish_node_task_));
    finish_node_task_.reset(new Task);
    when_done_task_->AddDependency(h);
    ++num_started_nodes_;
  }
  // constraint_builder_2d.cc:153-163,279-300
  void WhenDone(std::function<void(const std::vector<OracleConstraint>&)> cb) {
    std::lock_guard<std::mutex> l(mu_);
    ORACLE_CHECK(!when_done_);
    when_done_ = std::move(cb);
    when_done_task_->SetWorkItem([this] {
      std::vector<OracleConstraint> result;
      std::function<void(const std::vector<OracleConstraint>&)> cb2;
      {
        std::lock_guard<std::mutex> l2(mu_);
        for (const auto& c : constraints_)
          if (c) result.push_back(*c);
        constraints_.clear();
        cb2 = std::move(when_done_);
        when_done_ = nullptr;
      }
      cb2(result);
    });
    pool_->Schedule(std::move(when_done_task_));
    when_done_task_.reset(new Task);
  }
  int GetNumFinishedNodes() {
    std::lock_guard<std::mutex> l(mu_);
    return num_finished_nodes_;
  }
  // constraint_builder_2d.cc:307-316
  void DeleteScanMatcher(int submap_id) {
    std::lock_guard<std::mutex> l(mu_);
    matchers_.erase(submap_id);
    samplers_.erase(submap_id);
  }
  MatchStats stats() {
    std::lock_guard<std::mutex> l(mu_);
    return stats_;
  }

 private:
  struct SubmapScanMatcher {
    const ProbabilityGrid* grid = nullptr;
    std::unique_ptr<FastCorrelativeScanMatcher2D> matcher;
    std::weak_ptr<Task> creation;
  };
  void Add(int submap_id, const Submap* submap, int node_id, const PointCloud* cloud, bool full,
           const Rigid2d& rel) {
    std::lock_guard<std::mutex> l(mu_);
    constraints_.emplace_back();
    std::unique_ptr<OracleConstraint>* slot = &constraints_.back();
    const SubmapScanMatcher* m = Dispatch(submap_id, submap->grid);
    std::unique_ptr<Task> t(new Task);
    t->SetWorkItem([=] { Compute(submap_id, submap, node_id, full, cloud, rel, *m, slot); });
    t->AddDependency(m->creation);
    auto h = pool_->Schedule(std::move(t));
    finish_node_task_->AddDependency(h);
  }
  // constraint_builder_2d.cc:165-186
  const SubmapScanMatcher* Dispatch(int submap_id, const ProbabilityGrid* grid) {
    auto it = matchers_.find(submap_id);
    if (it != matchers_.end()) return &it->second;
    SubmapScanMatcher& m = matchers_[submap_id];
    m.grid = grid;
    std::unique_ptr<Task> t(new Task);
    SubmapScanMatcher* mp = &m;
    const FastCorrelativeScanMatcherOptions2D* o = &options_.fast_correlative_scan_matcher_options;
    t->SetWorkItem([mp, o] { mp->matcher.reset(new FastCorrelativeScanMatcher2D(*mp->grid, *o)); });
    m.creation = pool_->Schedule(std::move(t));
    return &m;
  }
  // constraint_builder_2d.cc:188-277 without the Ceres stage.
  void Compute(int submap_id, const Submap* submap, int node_id, bool full, const PointCloud* cloud,
               const Rigid2d& rel, const SubmapScanMatcher& m,
               std::unique_ptr<OracleConstraint>* slot) {
    ORACLE_CHECK(m.matcher);
    const Rigid2d& sp = submap->local_pose;
    // initial_pose = ComputeSubmapPose(*submap) * initial_relative_pose
    Rigid2d initial;
    {
      const double c = std::cos(sp.angle), s = std::sin(sp.angle);
      initial.tx = (c * rel.tx + -s * rel.ty) + sp.tx;
      initial.ty = (s * rel.tx + c * rel.ty) + sp.ty;
      initial.angle = sp.angle + rel.angle;
    }
    float score = 0.f;
    Rigid2d pose;
    FastCorrelativeScanMatcher2D::Winner w;
    MatchStats st;
    bool ok;
    if (full) {
      ok = m.matcher->MatchFullSubmap(*cloud, static_cast<float>(options_.global_localization_min_score),
                                      &score, &pose, &w, &st);
    } else {
      ok = m.matcher->Match(initial, *cloud, static_cast<float>(options_.min_score), &score, &pose,
                            &w, &st);
    }
    {
      std::lock_guard<std::mutex> l(mu_);
      stats_.scored_top += st.scored_top;
      stats_.scored_lower += st.scored_lower;
      stats_.num_scans += st.num_scans;
      stats_.num_points += st.num_points;
    }
    if (!ok) return;
    // constraint_transform = ComputeSubmapPose(*submap).inverse() * pose_estimate
    const double ia = -sp.angle;
    const double c = std::cos(ia), s = std::sin(ia);
    const double itx = -(c * sp.tx + -s * sp.ty), ity = -(s * sp.tx + c * sp.ty);
    std::unique_ptr<OracleConstraint> out(new OracleConstraint);
    out->submap_id = submap_id;
    out->node_id = node_id;
    out->tx = (c * pose.tx + -s * pose.ty) + itx;
    out->ty = (s * pose.tx + c * pose.ty) + ity;
    out->angle = ia + pose.angle;
    out->score = score;
    out->winner = w;
    *slot = std::move(out);
  }

  ConstraintBuilderOptions options_;
  ThreadPool* pool_;
  std::mutex mu_;
  std::function<void(const std::vector<OracleConstraint>&)> when_done_;
  int num_started_nodes_ = 0, num_finished_nodes_ = 0;
  std::unique_ptr<Task> finish_node_task_, when_done_task_;
  std::deque<std::unique_ptr<OracleConstraint>> constraints_;
  std::map<int, SubmapScanMatcher> matchers_;
  std::map<int, FixedRatioSampler> samplers_;
  MatchStats stats_;
};

}  // namespace oracle

#endif  // ORACLE_CPU_POOL_H_
