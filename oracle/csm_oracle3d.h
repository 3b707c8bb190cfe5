// TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference's 3D correlative scan
// matching path (juwangvsu/cartographer-1).  Only tests/, __graft_entry__.smoke() and
// bench.py's CPU legs may use it; the product never includes or links it.
//
// Follows, function for function (paths relative to /root/reference/cartographer):
//   mapping/3d/hybrid_grid.h:40-52,263-279,420-446,480-520   (HybridGrid semantics)
//   mapping/internal/3d/scan_matching/precomputation_grid_3d.{h,cc}
//   mapping/internal/3d/scan_matching/rotational_scan_matcher.cc:120-189
//   mapping/internal/3d/scan_matching/low_resolution_matcher.cc:23-35
//   mapping/internal/3d/scan_matching/fast_correlative_scan_matcher_3d.cc:57-440
//   transform/rigid_transform.h:116-196, transform/transform.h:57-115
//
// PARITY UNPINNED at bit level: the reference's float arithmetic goes through Eigen
// (quaternion products, normalized(), norm(), dot()), which is not available here.
// The formulas below are Eigen 3.3's scalar (non-vectorised) definitions; Eigen's SSE
// paths may order additions differently.  The integer parts (precomputation grids,
// candidate enumeration, branch and bound, tie order) do not depend on that.
#ifndef ORACLE_CSM_ORACLE3D_H_
#define ORACLE_CSM_ORACLE3D_H_

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <functional>
#include <limits>
#include <memory>
#include <unordered_map>
#include <vector>

#include "csm_oracle.h"

namespace oracle {

struct Array3i {
  int x = 0, y = 0, z = 0;
};
inline Array3i operator+(const Array3i& a, const Array3i& b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline Array3i operator-(const Array3i& a, const Array3i& b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }

struct Quat4f {  // Eigen::Quaternionf
  float w = 1.f, x = 0.f, y = 0.f, z = 0.f;
};
struct Rigid3f {
  Vec3f t{0.f, 0.f, 0.f};
  Quat4f q;
};

// Eigen 3.3 quat_product (generic): a * b.
inline Quat4f QuatMul(const Quat4f& a, const Quat4f& b) {
  Quat4f r;
  r.w = a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z;
  r.x = a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y;
  r.y = a.w * b.y + a.y * b.w + a.z * b.x - a.x * b.z;
  r.z = a.w * b.z + a.z * b.w + a.x * b.y - a.y * b.x;
  return r;
}
inline Quat4f QuatConj(const Quat4f& a) { return Quat4f{a.w, -a.x, -a.y, -a.z}; }
// Quaternion::normalized(): coeffs / norm, norm = sqrt(x^2 + y^2 + z^2 + w^2)
// (coefficient order x, y, z, w; summed left to right).
inline Quat4f QuatNormalized(const Quat4f& a) {
  const float n = std::sqrt(((a.x * a.x + a.y * a.y) + a.z * a.z) + a.w * a.w);
  return Quat4f{a.w / n, a.x / n, a.y / n, a.z / n};
}
// Quaternion::inverse(): conjugate / squaredNorm.
inline Quat4f QuatInverse(const Quat4f& a) {
  const float n2 = ((a.x * a.x + a.y * a.y) + a.z * a.z) + a.w * a.w;
  return Quat4f{a.w / n2, -a.x / n2, -a.y / n2, -a.z / n2};
}
// QuaternionBase::_transformVector: uv = 2 * (vec x v); v + w * uv + vec x uv.
inline Vec3f QuatRotate4(const Quat4f& q, const Vec3f& v) {
  return QuatRotate(Quatf{q.w, q.x, q.y, q.z}, v);
}
inline Vec3f Add(const Vec3f& a, const Vec3f& b) { return Vec3f{a.x + b.x, a.y + b.y, a.z + b.z}; }
inline float Norm3(const Vec3f& v) { return std::sqrt((v.x * v.x + v.y * v.y) + v.z * v.z); }

// transform/rigid_transform.h:160-164,186-190
inline Rigid3f Inverse(const Rigid3f& a) {
  Rigid3f r;
  r.q = QuatConj(a.q);
  const Vec3f t = QuatRotate4(r.q, a.t);
  r.t = Vec3f{-t.x, -t.y, -t.z};
  return r;
}
inline Rigid3f Mul(const Rigid3f& l, const Rigid3f& r) {
  Rigid3f o;
  o.t = Add(QuatRotate4(l.q, r.t), l.t);
  o.q = QuatNormalized(QuatMul(l.q, r.q));
  return o;
}
inline Vec3f Apply(const Rigid3f& a, const Vec3f& p) { return Add(QuatRotate4(a.q, p), a.t); }
// transform.h:41-47 with T = float.  The reference writes an UNQUALIFIED atan2(direction.y(),
// direction.x()): from namespace cartographer::transform, with <cmath> but not <math.h> in the
// translation unit, that resolves to the C library's ::atan2(double, double), so the yaw is
// computed in double and narrowed (it differs from atan2f in the last bit for ~16 % of inputs).
// This is what the reference's own sources do when compiled here (oracle/_ref; see
// tests/test_oracle_vs_ref3d.py); a build whose headers drag in libstdc++'s <math.h> wrapper would
// pick the float overload instead.
inline float GetYaw(const Quat4f& q) {
  const Vec3f d = QuatRotate4(q, Vec3f{1.f, 0.f, 0.f});
  return static_cast<float>(std::atan2(static_cast<double>(d.y), static_cast<double>(d.x)));
}
// transform.h:94-108 with T = float (the 2. literals promote to double)
inline Quat4f AngleAxisVectorToRotationQuaternion(const Vec3f& aa) {
  float scale = 0.5f, w = 1.f;
  const float sq = (aa.x * aa.x + aa.y * aa.y) + aa.z * aa.z;
  if (sq > 1e-8) {
    const float norm = std::sqrt(sq);
    scale = static_cast<float>(std::sin(norm / 2.) / norm);
    w = static_cast<float>(std::cos(norm / 2.));
  }
  return Quat4f{w, scale * aa.x, scale * aa.y, scale * aa.z};
}

// ---------------------------------------------------------------------------
// HybridGridBase<T>: a sparse map cell index -> value, default 0.  The reference's
// three-level tree (hybrid_grid.h:68-419) is an implementation detail; what the
// path observes is value(), the iteration over stored cells, and grid_size(), which
// starts at 128 and doubles while an inserted index does not fit (:255-301).
// ---------------------------------------------------------------------------
template <typename T>
class HybridGridBase3 {
 public:
  explicit HybridGridBase3(const float resolution) : resolution_(resolution) {}
  float resolution() const { return resolution_; }
  int grid_size() const { return grid_size_; }
  // hybrid_grid.h:428-433
  Array3i GetCellIndex(const Vec3f& p) const {
    return Array3i{RoundToInt(p.x / resolution_), RoundToInt(p.y / resolution_),
                   RoundToInt(p.z / resolution_)};
  }
  T value(const Array3i& i) const {
    const auto it = cells_.find(Key(i));
    return it == cells_.end() ? T(0) : it->second;
  }
  T* mutable_value(const Array3i& i) {
    while (!Fits(i)) grid_size_ <<= 1;
    return &cells_[Key(i)];
  }
  template <typename F>
  void ForEach(F f) const {
    for (const auto& kv : cells_) f(Unkey(kv.first), kv.second);
  }
  size_t size() const { return cells_.size(); }

 private:
  bool Fits(const Array3i& i) const {
    const int h = grid_size_ >> 1;
    return i.x >= -h && i.x < h && i.y >= -h && i.y < h && i.z >= -h && i.z < h;
  }
  static uint64_t Key(const Array3i& i) {
    return (static_cast<uint64_t>(static_cast<uint32_t>(i.x + (1 << 20))) << 42) |
           (static_cast<uint64_t>(static_cast<uint32_t>(i.y + (1 << 20))) << 21) |
           static_cast<uint64_t>(static_cast<uint32_t>(i.z + (1 << 20)));
  }
  static Array3i Unkey(uint64_t k) {
    return Array3i{static_cast<int>((k >> 42) & 0x1FFFFF) - (1 << 20),
                   static_cast<int>((k >> 21) & 0x1FFFFF) - (1 << 20),
                   static_cast<int>(k & 0x1FFFFF) - (1 << 20)};
  }
  float resolution_;
  int grid_size_ = 128;
  std::unordered_map<uint64_t, T> cells_;
};

// probability_values.cc:76-87
inline std::vector<uint16_t> ComputeLookupTableToApplyOdds(const float odds) {
  std::vector<uint16_t> result;
  result.reserve(32768);
  result.push_back(ProbabilityToValue(ProbabilityFromOdds(odds)) + kUpdateMarker);
  for (int cell = 1; cell != 32768; ++cell)
    result.push_back(
        ProbabilityToValue(ProbabilityFromOdds(odds * Odds(ValueToProbabilityTable()[cell]))) +
        kUpdateMarker);
  return result;
}

class HybridGrid : public HybridGridBase3<uint16_t> {
 public:
  explicit HybridGrid(const float resolution) : HybridGridBase3<uint16_t>(resolution) {}
  void SetProbability(const Array3i& i, const float p) { *mutable_value(i) = ProbabilityToValue(p); }
  float GetProbability(const Array3i& i) const { return ValueToProbabilityTable()[value(i)]; }
  // hybrid_grid.h:492-518 (std::unordered_map keeps element addresses stable)
  bool ApplyLookupTable(const Array3i& i, const std::vector<uint16_t>& table) {
    uint16_t* const cell = mutable_value(i);
    if (*cell >= kUpdateMarker) return false;
    update_indices_.push_back(cell);
    *cell = table[*cell];
    return true;
  }
  void FinishUpdate() {
    while (!update_indices_.empty()) {
      *update_indices_.back() -= kUpdateMarker;
      update_indices_.pop_back();
    }
  }

 private:
  std::vector<uint16_t*> update_indices_;
};

// RangeDataInserter3D::Insert without the intensity grid
// (cartographer/mapping/3d/range_data_inserter_3d.cc:47-140); test fixture builder.
class RangeDataInserter3D {
 public:
  RangeDataInserter3D(const float hit_probability, const float miss_probability,
                      const int num_free_space_voxels)
      : hit_table_(ComputeLookupTableToApplyOdds(Odds(hit_probability))),
        miss_table_(ComputeLookupTableToApplyOdds(Odds(miss_probability))),
        num_free_space_voxels_(num_free_space_voxels) {}
  void Insert(const Vec3f& origin, const std::vector<Vec3f>& returns, HybridGrid* grid) const {
    for (const Vec3f& hit : returns) grid->ApplyLookupTable(grid->GetCellIndex(hit), hit_table_);
    const Array3i origin_cell = grid->GetCellIndex(origin);
    for (const Vec3f& hit : returns) {
      const Array3i delta = grid->GetCellIndex(hit) - origin_cell;
      const int num_samples = std::max(std::abs(delta.x), std::max(std::abs(delta.y), std::abs(delta.z)));
      ORACLE_CHECK(num_samples < (1 << 15));
      for (int position = std::max(0, num_samples - num_free_space_voxels_); position < num_samples;
           ++position) {
        const Array3i miss_cell{origin_cell.x + delta.x * position / num_samples,
                                origin_cell.y + delta.y * position / num_samples,
                                origin_cell.z + delta.z * position / num_samples};
        grid->ApplyLookupTable(miss_cell, miss_table_);
      }
    }
    grid->FinishUpdate();
  }

 private:
  std::vector<uint16_t> hit_table_, miss_table_;
  int num_free_space_voxels_;
};

using PrecomputationGrid3D = HybridGridBase3<uint8_t>;
// precomputation_grid_3d.h:32-35
inline float ToProbability3D(const float value) {
  return kMinProbability + value * ((kMaxProbability - kMinProbability) / 255.f);
}

// precomputation_grid_3d.cc:49-61
inline PrecomputationGrid3D ConvertToPrecomputationGrid(const HybridGrid& grid) {
  PrecomputationGrid3D result(grid.resolution());
  grid.ForEach([&](const Array3i& idx, const uint16_t v) {
    const int cell_value = RoundToInt((ValueToProbabilityTable()[v] - kMinProbability) *
                                      (255.f / (kMaxProbability - kMinProbability)));
    ORACLE_CHECK(cell_value >= 0 && cell_value <= 255);
    *result.mutable_value(idx) = static_cast<uint8_t>(cell_value);
  });
  return result;
}

// precomputation_grid_3d.cc:63-81
inline PrecomputationGrid3D PrecomputeGrid(const PrecomputationGrid3D& grid, const bool half_resolution,
                                           const Array3i& shift) {
  PrecomputationGrid3D result(grid.resolution());
  grid.ForEach([&](const Array3i& idx, const uint8_t v) {
    for (int i = 0; i != 8; ++i) {
      Array3i c{idx.x - shift.x * ((i & 1) ? 1 : 0), idx.y - shift.y * ((i & 2) ? 1 : 0),
                idx.z - shift.z * ((i & 4) ? 1 : 0)};
      if (half_resolution) c = Array3i{c.x >> 1, c.y >> 1, c.z >> 1};
      uint8_t* cell = result.mutable_value(c);
      *cell = std::max(v, *cell);
    }
  });
  return result;
}

struct FastCorrelativeScanMatcherOptions3D {
  int branch_and_bound_depth = 8;
  int full_resolution_depth = 3;
  double min_rotational_score = 0.77;
  double min_low_resolution_score = 0.55;
  double linear_xy_search_window = 5.;
  double linear_z_search_window = 1.;
  double angular_search_window = 15. * M_PI / 180.;
};

// fast_correlative_scan_matcher_3d.cc:57-77
class PrecomputationGridStack3D {
 public:
  PrecomputationGridStack3D(const HybridGrid& grid, const FastCorrelativeScanMatcherOptions3D& o) {
    ORACLE_CHECK(o.branch_and_bound_depth >= 1);
    ORACLE_CHECK(o.full_resolution_depth >= 1);
    grids_.push_back(ConvertToPrecomputationGrid(grid));
    int last_width = 1;
    for (int depth = 1; depth != o.branch_and_bound_depth; ++depth) {
      const bool half_resolution = depth >= o.full_resolution_depth;
      const int next_width = 1 << depth;
      const int f = 1 << std::max(0, depth - o.full_resolution_depth);
      const int shift = (next_width - last_width + (f - 1)) / f;
      grids_.push_back(PrecomputeGrid(grids_.back(), half_resolution, Array3i{shift, shift, shift}));
      last_width = next_width;
    }
  }
  const PrecomputationGrid3D& Get(int depth) const { return grids_.at(depth); }
  int max_depth() const { return static_cast<int>(grids_.size()) - 1; }

 private:
  std::vector<PrecomputationGrid3D> grids_;
};

// ---------------------------------------------------------------------------
// RotationalScanMatcher::Match (rotational_scan_matcher.cc:120-189).  Histograms are
// Eigen::VectorXf; norm() and dot() are restated as left-to-right scalar sums.
// ---------------------------------------------------------------------------
using Histogram = std::vector<float>;

inline float HistNorm(const Histogram& h) {
  float s = 0.f;
  for (const float v : h) s += v * v;
  return std::sqrt(s);
}
inline float MatchHistograms(const Histogram& submap, const Histogram& scan) {
  const float scan_norm = HistNorm(scan);
  const float submap_norm = HistNorm(submap);
  const float normalization = scan_norm * submap_norm;
  if (normalization < 1e-3f) return 1.f;
  float dot = 0.f;
  for (size_t i = 0; i != submap.size(); ++i) dot += submap[i] * scan[i];
  return dot / normalization;
}
inline Histogram RotateHistogram(const Histogram& histogram, const float angle) {
  if (histogram.empty()) return histogram;
  const int n = static_cast<int>(histogram.size());
  // -angle * histogram.size() / M_PI : float * Index -> float, then / double -> double,
  // narrowed to float on assignment.
  const float rotate_by_buckets = static_cast<float>((-angle * static_cast<float>(n)) / M_PI);
  int full_buckets = RoundToInt(rotate_by_buckets - 0.5f);
  const float fraction = rotate_by_buckets - full_buckets;
  while (full_buckets < 0) full_buckets += n;
  Histogram out(n);
  for (int i = 0; i != n; ++i) {
    const float h0 = histogram[(i + full_buckets) % n];
    const float h1 = histogram[(i + 1 + full_buckets) % n];
    out[i] = fraction * h1 + (1.f - fraction) * h0;
  }
  return out;
}
inline std::vector<float> RotationalMatch(const Histogram& submap_histogram, const Histogram& histogram,
                                          const float initial_angle, const std::vector<float>& angles) {
  std::vector<float> result;
  for (const float angle : angles)
    result.push_back(MatchHistograms(submap_histogram, RotateHistogram(histogram, initial_angle + angle)));
  return result;
}

// gravity_alignment.inverse().cast<float>(): inverse in double (conjugate / squaredNorm),
// then narrowed (fast_correlative_scan_matcher_3d.cc:275-276).
inline Quat4f GravityInverseAsFloat(const double g[4]) {
  const double n2 = ((g[1] * g[1] + g[2] * g[2]) + g[3] * g[3]) + g[0] * g[0];
  return Quat4f{static_cast<float>(g[0] / n2), static_cast<float>(-g[1] / n2),
                static_cast<float>(-g[2] / n2), static_cast<float>(-g[3] / n2)};
}

// low_resolution_matcher.cc:23-35
inline float LowResolutionScore(const HybridGrid& grid, const std::vector<Vec3f>& points,
                                const Rigid3f& pose) {
  float score = 0.f;
  for (const Vec3f& p : points) score += grid.GetProbability(grid.GetCellIndex(Apply(pose, p)));
  return score / points.size();
}

// ---------------------------------------------------------------------------
// FastCorrelativeScanMatcher3D (fast_correlative_scan_matcher_3d.cc:79-440)
// ---------------------------------------------------------------------------
struct DiscreteScan3D {
  Rigid3f pose;
  std::vector<std::vector<Array3i>> cell_indices_per_depth;
  float rotational_score;
};

struct Candidate3D {
  Candidate3D(int scan_index, const Array3i& offset) : scan_index(scan_index), offset(offset) {}
  static Candidate3D Unsuccessful() { return Candidate3D(0, Array3i{0, 0, 0}); }
  int scan_index;
  Array3i offset;
  float score = -std::numeric_limits<float>::infinity();
  float low_resolution_score = 0.f;
  bool operator<(const Candidate3D& o) const { return score < o.score; }
  bool operator>(const Candidate3D& o) const { return score > o.score; }
};

struct NodeData3D {  // the fields of TrajectoryNode::Data this path reads
  std::vector<Vec3f> high_resolution_point_cloud;
  std::vector<Vec3f> low_resolution_point_cloud;
  Histogram rotational_scan_matcher_histogram;
  double gravity_alignment[4] = {1., 0., 0., 0.};  // Eigen::Quaterniond (w, x, y, z)
};

struct Result3D {
  bool found = false;
  float score = 0.f;
  Rigid3f pose_estimate;  // the reference widens this to Rigid3d on return
  float rotational_score = 0.f;
  float low_resolution_score = 0.f;
  int scan_index = 0;
  Array3i offset;
  uint64_t candidates_scored = 0;
  int num_scans = 0;
};

class FastCorrelativeScanMatcher3D {
 public:
  FastCorrelativeScanMatcher3D(const HybridGrid& grid, const HybridGrid* low_resolution_grid,
                               const Histogram* rotational_histogram,
                               const FastCorrelativeScanMatcherOptions3D& options)
      : options_(options),
        resolution_(grid.resolution()),
        width_in_voxels_(grid.grid_size()),
        stack_(grid, options),
        low_resolution_grid_(low_resolution_grid),
        submap_histogram_(rotational_histogram) {}

  const PrecomputationGridStack3D& stack() const { return stack_; }

  struct SearchParameters {
    int linear_xy_window_size;
    int linear_z_window_size;
    double angular_search_window;
  };

  // :127-147.  Poses arrive as the reference's Rigid3d narrowed to float.
  Result3D Match(const Rigid3f& global_node_pose, const Rigid3f& global_submap_pose,
                 const NodeData3D& data, const float min_score) const {
    const SearchParameters sp{RoundToInt(options_.linear_xy_search_window / resolution_),
                              RoundToInt(options_.linear_z_search_window / resolution_),
                              options_.angular_search_window};
    return MatchWithSearchParameters(sp, global_node_pose, global_submap_pose, data, min_score);
  }
  // :149-170
  Result3D MatchFullSubmap(const Quat4f& global_node_rotation, const Quat4f& global_submap_rotation,
                           const NodeData3D& data, const float min_score) const {
    float max_point_distance = 0.f;
    for (const Vec3f& p : data.high_resolution_point_cloud)
      max_point_distance = std::max(max_point_distance, Norm3(p));
    const int w = (width_in_voxels_ + 1) / 2 + RoundToInt(max_point_distance / resolution_ + 0.5f);
    const SearchParameters sp{w, w, M_PI};
    Rigid3f node, submap;
    node.q = global_node_rotation;
    submap.q = global_submap_rotation;
    return MatchWithSearchParameters(sp, node, submap, data, min_score);
  }

  // :246-295, exposed so that a device path can be checked stage by stage
  std::vector<DiscreteScan3D> GenerateDiscreteScans(const SearchParameters& sp, const NodeData3D& data,
                                                    const Rigid3f& global_node_pose,
                                                    const Rigid3f& global_submap_pose) const {
    std::vector<DiscreteScan3D> result;
    float max_scan_range = 3.f * resolution_;
    for (const Vec3f& p : data.high_resolution_point_cloud)
      max_scan_range = std::max(Norm3(p), max_scan_range);
    const float kSafetyMargin = 1.f - 1e-2f;
    const float angular_step_size =
        kSafetyMargin * std::acos(1.f - (resolution_ * resolution_) /
                                            (2.f * (max_scan_range * max_scan_range)));
    const int angular_window_size = RoundToInt(sp.angular_search_window / angular_step_size);
    std::vector<float> angles;
    for (int rz = -angular_window_size; rz <= angular_window_size; ++rz)
      angles.push_back(rz * angular_step_size);
    const Rigid3f node_to_submap = Mul(Inverse(global_submap_pose), global_node_pose);
    const std::vector<float> scores = RotationalMatch(
        *submap_histogram_, data.rotational_scan_matcher_histogram,
        GetYaw(QuatMul(node_to_submap.q, GravityInverseAsFloat(data.gravity_alignment))), angles);
    for (size_t i = 0; i != angles.size(); ++i) {
      if (scores[i] < options_.min_rotational_score) continue;
      Rigid3f pose;
      pose.t = node_to_submap.t;
      pose.q = QuatMul(QuatMul(QuatInverse(global_submap_pose.q),
                               AngleAxisVectorToRotationQuaternion(Vec3f{0.f, 0.f, angles[i]})),
                       global_node_pose.q);
      result.push_back(DiscretizeScan(sp, data.high_resolution_point_cloud, pose, scores[i]));
    }
    return result;
  }

 private:
  // :200-244
  DiscreteScan3D DiscretizeScan(const SearchParameters& sp, const std::vector<Vec3f>& cloud,
                                const Rigid3f& pose, const float rotational_score) const {
    std::vector<std::vector<Array3i>> per_depth;
    std::vector<Array3i> full;
    for (const Vec3f& p : cloud) full.push_back(stack_.Get(0).GetCellIndex(Apply(pose, p)));
    const int frd = std::min(options_.full_resolution_depth, options_.branch_and_bound_depth);
    ORACLE_CHECK(frd >= 1);
    for (int i = 0; i != frd; ++i) per_depth.push_back(full);
    const int low = options_.branch_and_bound_depth - frd;
    const Array3i start{-sp.linear_xy_window_size, -sp.linear_xy_window_size, -sp.linear_z_window_size};
    for (int i = 0; i != low; ++i) {
      const int e = i + 1;
      const Array3i low_start{start.x >> e, start.y >> e, start.z >> e};
      per_depth.emplace_back();
      for (const Array3i& c : full) {
        const Array3i at{c.x + start.x, c.y + start.y, c.z + start.z};
        per_depth.back().push_back(
            Array3i{(at.x >> e) - low_start.x, (at.y >> e) - low_start.y, (at.z >> e) - low_start.z});
      }
    }
    return DiscreteScan3D{pose, per_depth, rotational_score};
  }

  // :297-326
  std::vector<Candidate3D> GenerateLowestResolutionCandidates(const SearchParameters& sp,
                                                              const int num_scans) const {
    const int step = 1 << stack_.max_depth();
    std::vector<Candidate3D> out;
    for (int s = 0; s != num_scans; ++s)
      for (int z = -sp.linear_z_window_size; z <= sp.linear_z_window_size; z += step)
        for (int y = -sp.linear_xy_window_size; y <= sp.linear_xy_window_size; y += step)
          for (int x = -sp.linear_xy_window_size; x <= sp.linear_xy_window_size; x += step)
            out.emplace_back(s, Array3i{x, y, z});
    return out;
  }

  // :328-352
  void ScoreCandidates(const int depth, const std::vector<DiscreteScan3D>& scans,
                       std::vector<Candidate3D>* cands, uint64_t* scored) const {
    const int e = std::max(0, depth - options_.full_resolution_depth + 1);
    for (Candidate3D& c : *cands) {
      int sum = 0;
      const DiscreteScan3D& scan = scans[c.scan_index];
      const Array3i off{c.offset.x >> e, c.offset.y >> e, c.offset.z >> e};
      for (const Array3i& cell : scan.cell_indices_per_depth[depth])
        sum += stack_.Get(depth).value(cell + off);
      c.score = ToProbability3D(sum / static_cast<float>(scan.cell_indices_per_depth[depth].size()));
    }
    *scored += cands->size();
    std::sort(cands->begin(), cands->end(), std::greater<Candidate3D>());
  }

  // :363-369
  Rigid3f GetPoseFromCandidate(const std::vector<DiscreteScan3D>& scans, const Candidate3D& c) const {
    Rigid3f shift;
    shift.t = Vec3f{resolution_ * static_cast<float>(c.offset.x), resolution_ * static_cast<float>(c.offset.y),
                    resolution_ * static_cast<float>(c.offset.z)};
    return Mul(shift, scans[c.scan_index].pose);
  }

  // :371-440
  Candidate3D BranchAndBound(const SearchParameters& sp, const NodeData3D& data,
                             const std::vector<DiscreteScan3D>& scans,
                             const std::vector<Candidate3D>& cands, const int depth, float min_score,
                             uint64_t* scored) const {
    if (depth == 0) {
      for (const Candidate3D& c : cands) {
        if (c.score <= min_score) return Candidate3D::Unsuccessful();
        const float low = LowResolutionScore(*low_resolution_grid_, data.low_resolution_point_cloud,
                                             GetPoseFromCandidate(scans, c));
        if (low >= options_.min_low_resolution_score) {
          Candidate3D best = c;
          best.low_resolution_score = low;
          return best;
        }
      }
      return Candidate3D::Unsuccessful();
    }
    Candidate3D best = Candidate3D::Unsuccessful();
    best.score = min_score;
    for (const Candidate3D& c : cands) {
      if (c.score <= min_score) break;
      std::vector<Candidate3D> higher;
      const int half = 1 << (depth - 1);
      for (int z : {0, half}) {
        if (c.offset.z + z > sp.linear_z_window_size) break;
        for (int y : {0, half}) {
          if (c.offset.y + y > sp.linear_xy_window_size) break;
          for (int x : {0, half}) {
            if (c.offset.x + x > sp.linear_xy_window_size) break;
            higher.emplace_back(c.scan_index, Array3i{c.offset.x + x, c.offset.y + y, c.offset.z + z});
          }
        }
      }
      ScoreCandidates(depth - 1, scans, &higher, scored);
      best = std::max(best, BranchAndBound(sp, data, scans, higher, depth - 1, best.score, scored));
    }
    return best;
  }

  // :172-198
  Result3D MatchWithSearchParameters(const SearchParameters& sp, const Rigid3f& global_node_pose,
                                     const Rigid3f& global_submap_pose, const NodeData3D& data,
                                     const float min_score) const {
    Result3D r;
    const std::vector<DiscreteScan3D> scans =
        GenerateDiscreteScans(sp, data, global_node_pose, global_submap_pose);
    r.num_scans = static_cast<int>(scans.size());
    std::vector<Candidate3D> lowest = GenerateLowestResolutionCandidates(sp, static_cast<int>(scans.size()));
    ScoreCandidates(stack_.max_depth(), scans, &lowest, &r.candidates_scored);
    const Candidate3D best = BranchAndBound(sp, data, scans, lowest, stack_.max_depth(), min_score,
                                            &r.candidates_scored);
    if (best.score > min_score) {
      r.found = true;
      r.score = best.score;
      r.pose_estimate = GetPoseFromCandidate(scans, best);
      r.rotational_score = scans[best.scan_index].rotational_score;
      r.low_resolution_score = best.low_resolution_score;
      r.scan_index = best.scan_index;
      r.offset = best.offset;
    }
    return r;
  }

  FastCorrelativeScanMatcherOptions3D options_;
  float resolution_;
  int width_in_voxels_;
  PrecomputationGridStack3D stack_;
  const HybridGrid* low_resolution_grid_;
  const Histogram* submap_histogram_;
};

// ---------------------------------------------------------------------------
// RealTimeCorrelativeScanMatcher3D (real_time_correlative_scan_matcher_3d.cc:30-114)
// ---------------------------------------------------------------------------
// transform.h:33-37 with FloatType = float
inline float GetAngle(const Rigid3f& t) {
  return 2.f * std::atan2(Norm3(Vec3f{t.q.x, t.q.y, t.q.z}), std::abs(t.q.w));
}

class RealTimeCorrelativeScanMatcher3D {
 public:
  explicit RealTimeCorrelativeScanMatcher3D(const RealTimeCorrelativeScanMatcherOptions& o) : options_(o) {}

  // :55-98
  std::vector<Rigid3f> GenerateExhaustiveSearchTransforms(const float resolution,
                                                          const std::vector<Vec3f>& cloud) const {
    std::vector<Rigid3f> result;
    const int linear_window_size = RoundToInt(options_.linear_search_window / resolution);
    float max_scan_range = 3.f * resolution;
    for (const Vec3f& p : cloud) max_scan_range = std::max(Norm3(p), max_scan_range);
    const float kSafetyMargin = 1.f - 1e-3f;
    const float angular_step_size =
        kSafetyMargin *
        std::acos(1.f - (resolution * resolution) / (2.f * (max_scan_range * max_scan_range)));
    const int angular_window_size = RoundToInt(options_.angular_search_window / angular_step_size);
    for (int z = -linear_window_size; z <= linear_window_size; ++z)
      for (int y = -linear_window_size; y <= linear_window_size; ++y)
        for (int x = -linear_window_size; x <= linear_window_size; ++x)
          for (int rz = -angular_window_size; rz <= angular_window_size; ++rz)
            for (int ry = -angular_window_size; ry <= angular_window_size; ++ry)
              for (int rx = -angular_window_size; rx <= angular_window_size; ++rx) {
                Rigid3f t;
                t.t = Vec3f{x * resolution, y * resolution, z * resolution};
                t.q = AngleAxisVectorToRotationQuaternion(
                    Vec3f{rx * angular_step_size, ry * angular_step_size, rz * angular_step_size});
                result.push_back(t);
              }
    return result;
  }

  // :100-114; the cloud is transformed point by point (sensor/point_cloud.cc:56-64)
  float ScoreCandidate(const HybridGrid& grid, const std::vector<Vec3f>& cloud, const Rigid3f& candidate,
                       const Rigid3f& transform) const {
    float score = 0.f;
    for (const Vec3f& p : cloud) score += grid.GetProbability(grid.GetCellIndex(Apply(candidate, p)));
    score /= static_cast<float>(cloud.size());
    const float angle = GetAngle(transform);
    const double t = Norm3(transform.t) * options_.translation_delta_cost_weight +
                     angle * options_.rotation_delta_cost_weight;
    score *= std::exp(-(t * t));  // float *= double
    ORACLE_CHECK(score > 0.f);
    return score;
  }

  // :35-53.  best_index: position of the winner among the generated transforms.
  float Match(const Rigid3f& initial_pose_estimate, const std::vector<Vec3f>& cloud, const HybridGrid& grid,
              Rigid3f* pose_estimate, int* best_index = nullptr) const {
    ORACLE_CHECK(pose_estimate != nullptr);
    float best_score = -1.f;
    int index = 0;
    for (const Rigid3f& transform : GenerateExhaustiveSearchTransforms(grid.resolution(), cloud)) {
      const Rigid3f candidate = Mul(initial_pose_estimate, transform);
      const float score = ScoreCandidate(grid, cloud, candidate, transform);
      if (score > best_score) {
        best_score = score;
        *pose_estimate = candidate;
        if (best_index) *best_index = index;
      }
      ++index;
    }
    return best_score;
  }

 private:
  RealTimeCorrelativeScanMatcherOptions options_;
};

}  // namespace oracle

#endif  // ORACLE_CSM_ORACLE3D_H_
