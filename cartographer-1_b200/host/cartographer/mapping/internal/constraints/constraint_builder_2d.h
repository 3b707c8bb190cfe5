// B200 implementation behind the reference's ConstraintBuilder2D interface
// (cartographer/mapping/internal/constraints/constraint_builder_2d.h:60-176).
//
// Same public contract: intermingle MaybeAddConstraint / MaybeAddGlobalConstraint /
// NotifyEndOfNode, then WhenDone once; the callback receives the non-null
// constraints in submission order.  What changes is how the searches execute: the
// reference schedules one thread-pool task per (node, submap) search, each running a
// depth-first branch-and-bound on a CPU core.  Here the searches are queued and
// flushed in batches: each flush is ONE thread-pool task that shards the queued
// searches over the configured GPUs by submap affinity (all searches against a
// submap go to a GPU that holds its precomputation stack; greedy longest-processing-time
// on an estimated cost, and a submap whose searches would unbalance the GPUs -- the
// "new submap x all old nodes" burst of pose_graph_2d.cc:383-393 -- is replicated and
// its node list split), runs one FastCorrelativeScanMatcher2D::MatchBatch per GPU
// concurrently, and fills the result slots.  There is no collective on the data path;
// results are gathered on the host.  Ceres refinement (reference :245-249) runs as one more device batch per flush
// (CeresScanMatcher2D::MatchBatch) when options.ceres_scan_matcher_options() is filled in;
// with an empty message the constraint carries the correlative scan matcher's pose estimate.
#ifndef CARTOGRAPHER_MAPPING_INTERNAL_CONSTRAINTS_CONSTRAINT_BUILDER_2D_H_
#define CARTOGRAPHER_MAPPING_INTERNAL_CONSTRAINTS_CONSTRAINT_BUILDER_2D_H_

#include <deque>
#include <functional>
#include <map>
#include <memory>
#include <mutex>
#include <vector>

#include "cartographer/common/fixed_ratio_sampler.h"
#include "cartographer/common/histogram.h"
#include "cartographer/common/thread_pool.h"
#include "cartographer/metrics/family_factory.h"
#include "cartographer/mapping/internal/2d/scan_matching/fast_correlative_scan_matcher_2d.h"
#include "cartographer_compat.h"

namespace cartographer {
namespace mapping {
namespace constraints {

// Returns (map <- submap) where 'submap' is a coordinate system at the origin of
// the Submap (reference: constraint_builder_2d.h:55, .cc:55-57).
transform::Rigid2d ComputeSubmapPose(const Submap2D& submap);

class ConstraintBuilder2D {
 public:
  using Constraint = PoseGraphInterface::Constraint;
  using Result = std::vector<Constraint>;

  ConstraintBuilder2D(const proto::ConstraintBuilderOptions& options,
                      common::ThreadPoolInterface* thread_pool);
  // Same, sharding over the given CUDA devices (default: device 0 only).
  ConstraintBuilder2D(const proto::ConstraintBuilderOptions& options,
                      common::ThreadPoolInterface* thread_pool, const std::vector<int>& devices);
  ~ConstraintBuilder2D();

  ConstraintBuilder2D(const ConstraintBuilder2D&) = delete;
  ConstraintBuilder2D& operator=(const ConstraintBuilder2D&) = delete;

  // Schedules exploring a new constraint between 'submap' identified by 'submap_id',
  // and the point cloud of 'node_id'.  'initial_relative_pose' is relative to the
  // 'submap'.  The pointees must stay valid until all computations are finished.
  void MaybeAddConstraint(const SubmapId& submap_id, const Submap2D* submap,
                          const NodeId& node_id, const TrajectoryNode::Data* const constant_data,
                          const transform::Rigid2d& initial_relative_pose);

  // Same with full-submap matching.
  void MaybeAddGlobalConstraint(const SubmapId& submap_id, const Submap2D* submap,
                                const NodeId& node_id,
                                const TrajectoryNode::Data* const constant_data);

  // Must be called after all computations related to one node have been added.
  void NotifyEndOfNode();

  // Registers the 'callback' to be called with the results, after all computations
  // triggered by 'MaybeAdd*Constraint' have finished.  'callback' is executed in the
  // 'ThreadPool'.
  void WhenDone(const std::function<void(const Result&)>& callback);

  // Returns the number of consecutive finished nodes.
  int GetNumFinishedNodes();

  // Delete data related to 'submap_id'.
  void DeleteScanMatcher(const SubmapId& submap_id);

  static void RegisterMetrics(metrics::FamilyFactory* family_factory);

  // --- additions of the B200 path (not in the reference) ---
  // Whole-node / whole-submap batches.  The reference's caller emits
  // one MaybeAdd*Constraint call at a time, each taking the builder's mutex
  // (PoseGraph2D::ComputeConstraintsForNode, pose_graph_2d.cc:312-402: a node against every
  // finished submap, and at :383-393 a newly finished submap against all old nodes).  One call
  // here carries the whole burst: same per-request semantics as MaybeAddConstraint
  // (match_full_submap false: distance gate, then the submap's sampler) and
  // MaybeAddGlobalConstraint (true), in order, under ONE lock and with ONE flush decision.
  // NotifyEndOfNode stays the caller's.
  struct ConstraintRequest {
    SubmapId submap_id{0, 0};
    const Submap2D* submap = nullptr;
    NodeId node_id{0, 0};
    const TrajectoryNode::Data* constant_data = nullptr;
    bool match_full_submap = false;
    transform::Rigid2d initial_relative_pose;  // ignored for full-submap requests
  };
  void MaybeAddConstraints(const std::vector<ConstraintRequest>& requests);

  // Searches queued before a flush is scheduled from NotifyEndOfNode (WhenDone always
  // flushes the rest).  Default 256.
  void set_flush_threshold(int searches) { flush_threshold_ = searches; }
  // Totals since construction (what the reference reports through its metrics families,
  // constraint_builder_2d.cc:46-53, plus the sharding of the flushes over the devices).
  struct Stats {
    uint64_t searches = 0, found = 0, candidates_scored = 0, flushes = 0;
    uint64_t local_searched = 0, local_found = 0, global_searched = 0, global_found = 0;
    uint64_t queue_length = 0;         // result slots not yet handed to a WhenDone callback
    uint64_t num_submap_scan_matchers = 0;
    uint64_t replicated_stacks = 0;    // stacks built on a second (third, ...) device
    std::vector<uint64_t> searches_per_device;
    std::vector<double> cost_per_device;  // estimated cost assigned (the LPT quantity)
  };
  Stats stats();
  // common::Histogram::ToString of the scores of the constraints found (the reference logs it
  // from RunWhenDoneCallback when log_matches is set, .cc:289-293).
  std::string ScoreHistogram(int buckets = 10);

 private:
  struct SubmapScanMatcher {
    const Grid2D* grid = nullptr;
    // One stack per device that has searched this submap so far (index = slot in devices_;
    // nullptr = not built there).  Searches prefer a device that holds the stack.
    std::vector<std::shared_ptr<scan_matching::FastCorrelativeScanMatcher2D>> per_device;
  };
  struct PendingSearch {
    SubmapId submap_id;
    const Submap2D* submap;
    NodeId node_id;
    bool match_full_submap;
    const TrajectoryNode::Data* constant_data;
    transform::Rigid2d initial_relative_pose;
    std::unique_ptr<Constraint>* constraint;  // slot in constraints_
  };

  void Enqueue(const SubmapId& submap_id, const Submap2D* submap, const NodeId& node_id,
               bool match_full_submap, const TrajectoryNode::Data* constant_data,
               const transform::Rigid2d& initial_relative_pose);  // mutex_ held
  std::weak_ptr<common::Task> ScheduleFlush();                     // mutex_ held
  void RunFlush(const std::vector<PendingSearch>& searches, int finished_nodes_after);
  // Estimated cost of a search (points x angular slices x lowest-resolution lattice): what the
  // device assignment balances.
  double EstimateCost(const PendingSearch& search) const;
  // device slot of every search of a flush (mutex_ held).
  std::vector<int> AssignDevices(const std::vector<PendingSearch>& searches, std::vector<double>* cost);
  void RunWhenDoneCallback();

  const proto::ConstraintBuilderOptions options_;
  common::ThreadPoolInterface* thread_pool_;
  std::mutex mutex_;

  std::unique_ptr<std::function<void(const Result&)>> when_done_;
  int num_started_nodes_ = 0;
  int num_finished_nodes_ = 0;
  std::unique_ptr<common::Task> when_done_task_;
  std::weak_ptr<common::Task> last_flush_;
  std::vector<PendingSearch> pending_;
  int flush_threshold_ = 256;

  // Result slots in submission order; below-threshold searches stay nullptr.
  std::deque<std::unique_ptr<Constraint>> constraints_;

  std::vector<std::unique_ptr<scan_matching::DeviceContext>> devices_;
  std::map<SubmapId, SubmapScanMatcher> submap_scan_matchers_;
  std::map<SubmapId, common::FixedRatioSampler> per_submap_sampler_;
  Stats stats_;
  common::Histogram score_histogram_;
};

}  // namespace constraints
}  // namespace mapping
}  // namespace cartographer

#endif  // CARTOGRAPHER_MAPPING_INTERNAL_CONSTRAINTS_CONSTRAINT_BUILDER_2D_H_
