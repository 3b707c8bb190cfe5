// B200 implementation behind the reference's FastCorrelativeScanMatcher2D interface
// (cartographer/mapping/internal/2d/scan_matching/fast_correlative_scan_matcher_2d.h:112-164).
//
// Construction uploads the grid and builds the precomputation stack on the device
// (csm_stack2d_build); Match / MatchFullSubmap keep the reference's signatures and
// semantics and forward to csm_match2d_batch.  MatchBatch is the additional entry
// the constraint builder uses to coalesce many (node, submap) searches into one
// launch sequence.  There is no CPU fallback: construction aborts (CHECK semantics)
// when no CUDA device is usable.
#ifndef CARTOGRAPHER_MAPPING_INTERNAL_2D_SCAN_MATCHING_FAST_CORRELATIVE_SCAN_MATCHER_2D_H_
#define CARTOGRAPHER_MAPPING_INTERNAL_2D_SCAN_MATCHING_FAST_CORRELATIVE_SCAN_MATCHER_2D_H_

#include <memory>
#include <vector>

#include "cartographer/common/lua_parameter_dictionary.h"
#include "cartographer/mapping/internal/2d/scan_matching/correlative_scan_matcher_2d.h"
#include "cartographer_compat.h"
#include "csm_b200.h"

namespace cartographer {
namespace mapping {
namespace scan_matching {

// Owns one csm_ctx per device and hands it to matchers; thread-compatible (callers
// serialise use of one context, see include/csm_b200.h).
class DeviceContext {
 public:
  explicit DeviceContext(int device = 0);
  ~DeviceContext();
  DeviceContext(const DeviceContext&) = delete;
  DeviceContext& operator=(const DeviceContext&) = delete;
  csm_ctx* get() const { return ctx_; }
  // One lazily created context on device 0 per calling thread (the reference's
  // Match* are called concurrently from pool threads; a csm_ctx is single-threaded).
  static DeviceContext* ThreadDefault();
  // One lazily created context per (calling thread, device): what the const, concurrently
  // callable Match / MatchFullSubmap of a matcher living on `device` run on.
  static DeviceContext* ThreadLocal(int device);
  int device() const { return device_; }

 private:
  csm_ctx* ctx_ = nullptr;
  int device_ = 0;
};

// uint16 cell value -> uint8 precomputation value, i.e. the reference's
// ComputeCellValue(1.f - |GetCorrespondenceCost|) for every possible cell value
// (fast_correlative_scan_matcher_2d.cc:97-98,112-113,163-169).
std::vector<uint8> ComputeQuantizationTable(const Grid2D& grid);

// Reference :42-44, .cc:76-89.
proto::FastCorrelativeScanMatcherOptions2D CreateFastCorrelativeScanMatcherOptions2D(
    common::LuaParameterDictionary* parameter_dictionary);

// A precomputed grid that contains in each cell (x0, y0) the maximum probability in the
// width x width area x0 <= x < x0 + width, y0 <= y < y0 + width (reference :49-92, .cc:91-169).
// Built on the device for ANY width (csm_precomputation_grid2d); the bytes are kept on the host
// so that GetValue is the reference's inline lookup.  `reusable_intermediate_grid` is accepted
// for signature compatibility and left untouched: the device build has no float intermediate.
class PrecomputationGrid2D {
 public:
  PrecomputationGrid2D(const Grid2D& grid, const CellLimits& limits, int width,
                       std::vector<float>* reusable_intermediate_grid);
  // Adopts bytes that are already on the host (PrecomputationGridStack2D downloads its levels).
  PrecomputationGrid2D(const CellLimits& limits, int width, float min_score, float max_score,
                       std::vector<uint8> cells);

  // Returns a value between 0 and 255 to represent probabilities between min_score and max_score.
  int GetValue(const Eigen::Array2i& xy_index) const {
    const Eigen::Array2i local_xy_index = xy_index - offset_;
    if (static_cast<unsigned>(local_xy_index.x()) >= static_cast<unsigned>(wide_limits_.num_x_cells) ||
        static_cast<unsigned>(local_xy_index.y()) >= static_cast<unsigned>(wide_limits_.num_y_cells)) {
      return 0;
    }
    const int stride = wide_limits_.num_x_cells;
    return cells_[local_xy_index.x() + local_xy_index.y() * stride];
  }

  // Maps values from [0, 255] to [min_score, max_score].
  float ToScore(float value) const { return min_score_ + value * ((max_score_ - min_score_) / 255.f); }

 private:
  Eigen::Array2i offset_;
  CellLimits wide_limits_;
  float min_score_;
  float max_score_;
  std::vector<uint8> cells_;
};

// Reference :94-109, .cc:171-186: levels of width 1 << i.  One device build (the matcher's own
// csm_stack2d_build); a level is downloaded the first time Get() asks for it.
class PrecomputationGridStack2D {
 public:
  PrecomputationGridStack2D(const Grid2D& grid, const proto::FastCorrelativeScanMatcherOptions2D& options);
  ~PrecomputationGridStack2D();
  PrecomputationGridStack2D(const PrecomputationGridStack2D&) = delete;
  PrecomputationGridStack2D& operator=(const PrecomputationGridStack2D&) = delete;

  const PrecomputationGrid2D& Get(int index);
  int max_depth() const { return static_cast<int>(precomputation_grids_.size()) - 1; }

 private:
  CellLimits limits_;
  float min_score_ = 0.f, max_score_ = 0.f;  // the grid is borrowed during construction only
  csm_ctx* build_context_ = nullptr;
  csm_stack2d* stack_ = nullptr;
  std::vector<std::unique_ptr<PrecomputationGrid2D>> precomputation_grids_;
};

class FastCorrelativeScanMatcher2D {
 public:
  FastCorrelativeScanMatcher2D(const Grid2D& grid,
                               const proto::FastCorrelativeScanMatcherOptions2D& options);
  // Same, on an explicit device context (multi-GPU sharding).
  FastCorrelativeScanMatcher2D(const Grid2D& grid,
                               const proto::FastCorrelativeScanMatcherOptions2D& options,
                               DeviceContext* context);
  ~FastCorrelativeScanMatcher2D();

  // Several matchers at once (the submaps of a loaded map, a burst of finished submaps): the
  // same matchers as constructing one per grid, in the order of `grids`; their stacks are built by
  // csm_stack2d_build_many -- the host stages grid k + 1 while the device builds stack k.
  static std::vector<std::unique_ptr<FastCorrelativeScanMatcher2D>> CreateMany(
      const std::vector<const Grid2D*>& grids, const proto::FastCorrelativeScanMatcherOptions2D& options,
      DeviceContext* context);

  FastCorrelativeScanMatcher2D(const FastCorrelativeScanMatcher2D&) = delete;
  FastCorrelativeScanMatcher2D& operator=(const FastCorrelativeScanMatcher2D&) = delete;

  // Reference semantics: true iff a score above 'min_score' (strictly) exists; only
  // then 'score' and 'pose_estimate' are written.
  bool Match(const transform::Rigid2d& initial_pose_estimate,
             const sensor::PointCloud& point_cloud, float min_score, float* score,
             transform::Rigid2d* pose_estimate) const;
  bool MatchFullSubmap(const sensor::PointCloud& point_cloud, float min_score, float* score,
                       transform::Rigid2d* pose_estimate) const;

  // One search of a batch; 'full_submap' selects MatchFullSubmap semantics.
  struct Query {
    const sensor::PointCloud* point_cloud = nullptr;
    bool full_submap = false;
    transform::Rigid2d initial_pose_estimate;
    float min_score = 0.f;
  };
  struct Answer {
    bool found = false;
    float score = 0.f;
    transform::Rigid2d pose_estimate;
    Candidate2D best_candidate{0, 0, 0, SearchParameters(0, 0, 0., 0.)};
    uint64_t candidates_scored = 0;
    int tie_resolution = 0;
  };
  using QueryList = std::vector<std::pair<const FastCorrelativeScanMatcher2D*, Query>>;
  // All queries may target different matchers on the same device context.
  static std::vector<Answer> MatchBatch(DeviceContext* context, const QueryList& queries);

  // The same in three steps, so that a caller can keep a prepared batch resident on
  // the device and overlap host work: the constructor computes the search
  // parameters (libm) and uploads; Run() enqueues every kernel on the context's
  // stream without waiting; Fetch() waits, copies the results back and converts the
  // winning candidates to poses.  Run() may be repeated.
  //
  // A large batch is cut into device batches of bounded footprint.  With `resident` false
  // nothing is uploaded by the constructor and RunStreaming() moves the device batches
  // through the device a few at a time -- upload k+1 and its kernels are queued while k runs,
  // results are collected as each finishes -- so the device footprint stays at a few device
  // batches however large the job is (MatchBatch does this).
  class Batch {
   public:
    Batch(DeviceContext* context, const QueryList& queries, bool resident = true);
    ~Batch();
    Batch(const Batch&) = delete;
    Batch& operator=(const Batch&) = delete;
    void Run();
    std::vector<Answer> Fetch();
    // `queries`: the list the batch was constructed from (the streaming form reads the point
    // clouds of device batch k+1 while k runs).
    std::vector<Answer> RunStreaming(const QueryList& queries);
    size_t size() const { return initial_.size(); }

   private:
    std::vector<Answer> Convert(const std::vector<csm_result2d>& results) const;
    // The heavy host half of jobs [begin, end): point copies and the scans' angle tables (libm).
    void Prepare(const QueryList& queries, size_t begin, size_t end);

    DeviceContext* context_;
    std::vector<SearchParameters> params_;
    std::vector<transform::Rigid2d> initial_;
    std::vector<std::vector<float>> xyz_, quats_;    // host copies the job descriptors point into
    std::vector<csm_job2d> jobs_;
    std::vector<std::pair<size_t, size_t>> chunks_;  // [begin, end) job ranges
    std::vector<csm_batch2d*> device_batches_;       // resident form only
  };

  const csm_stack2d* stack() const { return stack_; }
  // The grid's uint16 cells, resident on device() beside the stack (for CeresScanMatcher2D::Query).
  const uint16* device_cells() const { return csm_stack2d_cells_dev(stack_); }
  int device() const { return device_; }
  const MapLimits& limits() const { return limits_; }
  DeviceContext* context() const { return context_; }

 private:
  struct Prebuilt {};  // CreateMany: every member but the stack
  FastCorrelativeScanMatcher2D(const Grid2D& grid, const proto::FastCorrelativeScanMatcherOptions2D& options,
                               DeviceContext* context, Prebuilt);
  void Build(const Grid2D& grid);

  const proto::FastCorrelativeScanMatcherOptions2D options_;
  MapLimits limits_;
  float min_score_ = 0.f;  // PrecomputationGrid2D::min_score_
  float max_score_ = 0.f;  // PrecomputationGrid2D::max_score_
  DeviceContext* context_ = nullptr;
  csm_stack2d* stack_ = nullptr;
  csm_ctx* build_context_ = nullptr;  // raw handle: DeviceContext may be deleted before us
  int device_ = 0;                    // where the stack lives (context_ may dangle once its thread exits)
};

}  // namespace scan_matching
}  // namespace mapping
}  // namespace cartographer

#endif  // CARTOGRAPHER_MAPPING_INTERNAL_2D_SCAN_MATCHING_FAST_CORRELATIVE_SCAN_MATCHER_2D_H_
