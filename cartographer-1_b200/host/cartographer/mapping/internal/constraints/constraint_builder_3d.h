// B200 implementation behind the reference's ConstraintBuilder3D interface
// (cartographer/mapping/internal/constraints/constraint_builder_3d.h:59-171).
//
// Same public contract as the reference and the same execution scheme as
// ConstraintBuilder2D in this tree: searches are queued, flushed in batches as one
// thread-pool task, sharded over the configured GPUs by submap affinity and run as one
// FastCorrelativeScanMatcher3D::Batch per GPU.  Results are gathered on the host in
// submission order.  Ceres refinement over both grids (reference :264-275) is not part
// of this path: the constraint carries the correlative scan matcher's pose estimate.
#ifndef CARTOGRAPHER_MAPPING_INTERNAL_CONSTRAINTS_CONSTRAINT_BUILDER_3D_H_
#define CARTOGRAPHER_MAPPING_INTERNAL_CONSTRAINTS_CONSTRAINT_BUILDER_3D_H_

#include <deque>
#include <functional>
#include <map>
#include <memory>
#include <mutex>
#include <vector>

#include "cartographer/common/fixed_ratio_sampler.h"
#include "cartographer/common/thread_pool.h"
#include "cartographer/mapping/internal/3d/scan_matching/fast_correlative_scan_matcher_3d.h"
#include "cartographer_compat_3d.h"

namespace cartographer {
namespace mapping {
namespace constraints {

class ConstraintBuilder3D {
 public:
  using Constraint = PoseGraphInterface::Constraint;
  using Result = std::vector<Constraint>;

  ConstraintBuilder3D(const proto::ConstraintBuilderOptions& options,
                      common::ThreadPoolInterface* thread_pool);
  // Same, sharding over the given CUDA devices (default: device 0 only).
  ConstraintBuilder3D(const proto::ConstraintBuilderOptions& options,
                      common::ThreadPoolInterface* thread_pool, const std::vector<int>& devices);
  ~ConstraintBuilder3D();

  ConstraintBuilder3D(const ConstraintBuilder3D&) = delete;
  ConstraintBuilder3D& operator=(const ConstraintBuilder3D&) = delete;

  // 'global_node_pose' and 'global_submap_pose' are initial estimates of poses in the
  // global map frame, i.e. both are gravity aligned.  The pointees must stay valid
  // until all computations are finished.
  void MaybeAddConstraint(const SubmapId& submap_id, const Submap3D* submap, const NodeId& node_id,
                          const TrajectoryNode::Data* const constant_data,
                          const transform::Rigid3d& global_node_pose,
                          const transform::Rigid3d& global_submap_pose);

  // Full-submap matching; the rotations are initial estimates of roll and pitch, their
  // yaw is essentially ignored.
  void MaybeAddGlobalConstraint(const SubmapId& submap_id, const Submap3D* submap,
                                const NodeId& node_id,
                                const TrajectoryNode::Data* const constant_data,
                                const Eigen::Quaterniond& global_node_rotation,
                                const Eigen::Quaterniond& global_submap_rotation);

  void NotifyEndOfNode();
  void WhenDone(const std::function<void(const Result&)>& callback);
  int GetNumFinishedNodes();
  void DeleteScanMatcher(const SubmapId& submap_id);
  static void RegisterMetrics(metrics::FamilyFactory* family_factory);

  // --- additions of the B200 path (not in the reference) ---
  // Whole-node / whole-submap batches, as ConstraintBuilder2D's: the
  // reference's caller (PoseGraph3D::ComputeConstraintsForNode, pose_graph_3d.cc) emits one
  // MaybeAdd*Constraint call per (node, submap) pair, each taking the builder's mutex.  One call
  // carries the burst: per request the semantics of MaybeAddConstraint (match_full_submap false:
  // distance gate, then the submap's sampler) or MaybeAddGlobalConstraint (true: only the poses'
  // rotations are used), in order, under ONE lock and with ONE flush decision.
  struct ConstraintRequest {
    SubmapId submap_id{0, 0};
    const Submap3D* submap = nullptr;
    NodeId node_id{0, 0};
    const TrajectoryNode::Data* constant_data = nullptr;
    bool match_full_submap = false;
    transform::Rigid3d global_node_pose;
    transform::Rigid3d global_submap_pose;
  };
  void MaybeAddConstraints(const std::vector<ConstraintRequest>& requests);
  void set_flush_threshold(int searches) { flush_threshold_ = searches; }
  struct Stats {
    uint64_t searches = 0, found = 0, candidates_scored = 0, flushes = 0;
  };
  Stats stats();

 private:
  struct SubmapScanMatcher {
    const HybridGrid* high_resolution_hybrid_grid = nullptr;
    const HybridGrid* low_resolution_hybrid_grid = nullptr;
    const Eigen::VectorXf* rotational_scan_matcher_histogram = nullptr;
    int device_slot = 0;
    std::shared_ptr<scan_matching::FastCorrelativeScanMatcher3D> fast_correlative_scan_matcher;
  };
  struct PendingSearch {
    SubmapId submap_id;
    const Submap3D* submap;
    NodeId node_id;
    bool match_full_submap;
    const TrajectoryNode::Data* constant_data;
    transform::Rigid3d global_node_pose;
    transform::Rigid3d global_submap_pose;
    std::unique_ptr<Constraint>* constraint;
  };

  SubmapScanMatcher* GetOrCreateSlot(const SubmapId& submap_id, const Submap3D* submap);  // mutex_ held
  void Enqueue(const SubmapId& submap_id, const Submap3D* submap, const NodeId& node_id,
               bool match_full_submap, const TrajectoryNode::Data* constant_data,
               const transform::Rigid3d& global_node_pose,
               const transform::Rigid3d& global_submap_pose);  // mutex_ held
  std::weak_ptr<common::Task> ScheduleFlush();                 // mutex_ held
  void RunFlush(const std::vector<PendingSearch>& searches, int finished_nodes_after);
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
  std::deque<std::unique_ptr<Constraint>> constraints_;
  std::vector<std::unique_ptr<scan_matching::DeviceContext>> devices_;
  std::map<SubmapId, SubmapScanMatcher> submap_scan_matchers_;
  std::map<SubmapId, common::FixedRatioSampler> per_submap_sampler_;
  int next_device_slot_ = 0;
  Stats stats_;
};

}  // namespace constraints
}  // namespace mapping
}  // namespace cartographer

#endif  // CARTOGRAPHER_MAPPING_INTERNAL_CONSTRAINTS_CONSTRAINT_BUILDER_3D_H_
