// B200 implementation behind the reference's FastCorrelativeScanMatcher3D interface
// (cartographer/mapping/internal/3d/scan_matching/fast_correlative_scan_matcher_3d.h:66-151).
//
// Construction exports the HybridGrid's voxels and builds the precomputation stack and
// the dense low-resolution probability grid on the device (csm_stack3d_build,
// csm_grid3d_build).  Match / MatchFullSubmap keep the reference's signatures and
// semantics: the host computes the search parameters, the angular step, the rotational
// prefilter and the poses of the surviving yaw slices (libm-dependent, O(slices)),
// everything from DiscretizeScan to the best candidate runs on the device
// (csm_match3d_batch).  MatchBatch coalesces many searches into one launch sequence.
// There is no CPU fallback.
#ifndef CARTOGRAPHER_MAPPING_INTERNAL_3D_SCAN_MATCHING_FAST_CORRELATIVE_SCAN_MATCHER_3D_H_
#define CARTOGRAPHER_MAPPING_INTERNAL_3D_SCAN_MATCHING_FAST_CORRELATIVE_SCAN_MATCHER_3D_H_

#include <memory>
#include <vector>

#include "cartographer/mapping/internal/2d/scan_matching/fast_correlative_scan_matcher_2d.h"
#include "cartographer/mapping/internal/3d/scan_matching/rotational_scan_matcher.h"
#include "cartographer_compat_3d.h"
#include "csm_b200.h"

namespace cartographer {
namespace mapping {
namespace scan_matching {

// uint16 cell value -> uint8 precomputation value (ConvertToPrecomputationGrid,
// precomputation_grid_3d.cc:49-61) and -> fp32 probability (probability_values.cc:29-60).
std::vector<uint8> ComputeQuantizationTable3D();
const std::vector<float>& ValueToProbabilityTable();

// Reference :41-43, .cc:36-56.
proto::FastCorrelativeScanMatcherOptions3D CreateFastCorrelativeScanMatcherOptions3D(
    common::LuaParameterDictionary* parameter_dictionary);

// One level of the 3D precomputation stack as the reference's callers read it
// (precomputation_grid_3d.h:27-38: a HybridGridBase<uint8> with value() and ToProbability()).
// Here a dense byte box downloaded from the device; outside the box a voxel reads the sparse
// grid's default, 0.
class PrecomputationGrid3D {
 public:
  PrecomputationGrid3D(float resolution, const int32_t lo[3], const int32_t dims[3], std::vector<uint8> bytes)
      : resolution_(resolution), lo_{lo[0], lo[1], lo[2]}, dims_{dims[0], dims[1], dims[2]}, bytes_(std::move(bytes)) {}
  float resolution() const { return resolution_; }
  uint8 value(const Eigen::Array3i& index) const {
    const int x = index.x() - lo_[0], y = index.y() - lo_[1], z = index.z() - lo_[2];
    if (static_cast<unsigned>(x) >= static_cast<unsigned>(dims_[0]) ||
        static_cast<unsigned>(y) >= static_cast<unsigned>(dims_[1]) ||
        static_cast<unsigned>(z) >= static_cast<unsigned>(dims_[2]))
      return 0;
    return bytes_[(static_cast<size_t>(z) * dims_[1] + y) * dims_[0] + x];
  }
  // Maps values from [0, 255] to [kMinProbability, kMaxProbability].
  static float ToProbability(float value) {
    return kMinProbability + value * ((kMaxProbability - kMinProbability) / 255.f);
  }

 private:
  float resolution_;
  int32_t lo_[3], dims_[3];
  std::vector<uint8> bytes_;
};

// Reference :45-64, .cc:57-77.  One device build (csm_stack3d_build); every level is downloaded at
// construction (Get() is const in the reference).  Levels at or beyond full_resolution_depth hold
// half-resolution indices, like the reference's.
class PrecomputationGridStack3D {
 public:
  PrecomputationGridStack3D(const HybridGrid& hybrid_grid,
                            const proto::FastCorrelativeScanMatcherOptions3D& options);
  const PrecomputationGrid3D& Get(int depth) const { return precomputation_grids_.at(depth); }
  int max_depth() const { return static_cast<int>(precomputation_grids_.size()) - 1; }

 private:
  std::vector<PrecomputationGrid3D> precomputation_grids_;
};

class FastCorrelativeScanMatcher3D {
 public:
  struct Result {
    float score;
    transform::Rigid3d pose_estimate;
    float rotational_score;
    float low_resolution_score;
  };

  FastCorrelativeScanMatcher3D(const HybridGrid& hybrid_grid,
                               const HybridGrid* low_resolution_hybrid_grid,
                               const Eigen::VectorXf* rotational_scan_matcher_histogram,
                               const proto::FastCorrelativeScanMatcherOptions3D& options);
  // Same, on an explicit device context (multi-GPU sharding).
  FastCorrelativeScanMatcher3D(const HybridGrid& hybrid_grid,
                               const HybridGrid* low_resolution_hybrid_grid,
                               const Eigen::VectorXf* rotational_scan_matcher_histogram,
                               const proto::FastCorrelativeScanMatcherOptions3D& options,
                               DeviceContext* context);
  ~FastCorrelativeScanMatcher3D();

  FastCorrelativeScanMatcher3D(const FastCorrelativeScanMatcher3D&) = delete;
  FastCorrelativeScanMatcher3D& operator=(const FastCorrelativeScanMatcher3D&) = delete;

  // 'Result' is only returned if a score above 'min_score' (excluding equality) is
  // possible.
  std::unique_ptr<Result> Match(const transform::Rigid3d& global_node_pose,
                                const transform::Rigid3d& global_submap_pose,
                                const TrajectoryNode::Data& constant_data, float min_score) const;
  std::unique_ptr<Result> MatchFullSubmap(const Eigen::Quaterniond& global_node_rotation,
                                          const Eigen::Quaterniond& global_submap_rotation,
                                          const TrajectoryNode::Data& constant_data,
                                          float min_score) const;

  struct Query {
    const TrajectoryNode::Data* constant_data = nullptr;
    bool full_submap = false;
    transform::Rigid3d global_node_pose;    // rotation only is used when full_submap
    transform::Rigid3d global_submap_pose;
    float min_score = 0.f;
  };
  struct Answer {
    bool found = false;
    Result result{0.f, transform::Rigid3d(), 0.f, 0.f};
    int scan_index = 0;        // Candidate3D::scan_index among the kept yaw slices
    int offset[3] = {0, 0, 0};  // Candidate3D::offset
    int num_scans = 0;          // yaw slices that passed the rotational prefilter
    uint64_t candidates_scored = 0;
    int tie_resolution = 0;
  };
  using QueryList = std::vector<std::pair<const FastCorrelativeScanMatcher3D*, Query>>;
  static std::vector<Answer> MatchBatch(DeviceContext* context, const QueryList& queries);

  // Prepared batch kept resident on the device: the constructor does the host-side
  // per-search work and uploads; Run() enqueues the kernels; Fetch() waits and
  // converts candidates to poses.  Run() may be repeated.
  class Batch {
   public:
    Batch(DeviceContext* context, const QueryList& queries);
    ~Batch();
    Batch(const Batch&) = delete;
    Batch& operator=(const Batch&) = delete;
    void Run();
    std::vector<Answer> Fetch();
    size_t size() const { return prepared_.size(); }

   private:
    struct Prepared {
      std::vector<float> scan_poses;      // S x 7
      std::vector<float> scan_leaf_quat;  // S x 4
      std::vector<float> rotational_scores;
      std::vector<float> points, low_points;
      float resolution = 0.f;
    };
    DeviceContext* context_;
    std::vector<Prepared> prepared_;
    std::vector<std::pair<size_t, size_t>> chunks_;
    std::vector<csm_batch3d*> device_batches_;
  };

  const csm_stack3d* stack() const { return stack_; }
  const csm_grid3d* low_resolution_grid() const { return low_grid_; }
  DeviceContext* context() const { return context_; }

  // Discrete scan poses of a local Match (visible for testing): S x 7 floats.
  std::vector<float> DiscreteScanPosesForTesting(const transform::Rigid3d& global_node_pose,
                                                 const transform::Rigid3d& global_submap_pose,
                                                 const TrajectoryNode::Data& constant_data,
                                                 std::vector<float>* rotational_scores) const;

 private:
  struct SearchParameters {
    int linear_xy_window_size;     // voxels
    int linear_z_window_size;      // voxels
    double angular_search_window;  // radians
  };
  void Build(const HybridGrid& hybrid_grid);
  SearchParameters ParametersFor(const Query& query) const;
  void GenerateDiscreteScanPoses(const SearchParameters& search_parameters,
                                 const TrajectoryNode::Data& constant_data,
                                 const transform::Rigid3f& global_node_pose,
                                 const transform::Rigid3f& global_submap_pose,
                                 std::vector<float>* scan_poses, std::vector<float>* scan_leaf_quat,
                                 std::vector<float>* rotational_scores) const;

  const proto::FastCorrelativeScanMatcherOptions3D options_;
  const float resolution_;
  const int width_in_voxels_;
  const HybridGrid* const low_resolution_hybrid_grid_;
  RotationalScanMatcher rotational_scan_matcher_;
  DeviceContext* context_ = nullptr;
  csm_stack3d* stack_ = nullptr;
  csm_grid3d* low_grid_ = nullptr;
  csm_ctx* build_context_ = nullptr;  // raw handle the device blocks go back to
  int device_ = 0;                    // where the stack lives
};

}  // namespace scan_matching
}  // namespace mapping
}  // namespace cartographer

#endif  // CARTOGRAPHER_MAPPING_INTERNAL_3D_SCAN_MATCHING_FAST_CORRELATIVE_SCAN_MATCHER_3D_H_
