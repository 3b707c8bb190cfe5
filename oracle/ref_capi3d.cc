// TEST INFRASTRUCTURE ONLY.  The 3D half of oracle/_ref/libref.so: flat C entry points over the
// reference's own HybridGrid, PrecomputationGridStack3D, FastCorrelativeScanMatcher3D,
// RotationalScanMatcher and RealTimeCorrelativeScanMatcher3D, compiled unmodified from
// /root/reference (see ref_capi.cc for what is and is not the reference's in this library).  Same
// entry-point names and arguments as oracle_capi.cc's 3D section.
//
// The reference's Result carries score, pose, rotational and low-resolution score, not the
// winning candidate: scan_index / offset are reported as INT32_MIN.  Floating point that goes
// through the Eigen stand-in (quaternion products, normalized(), VectorXf::norm()/dot()) is the
// scalar formula; see ref_shim/Eigen/Core on why that part has no single reference bit pattern.
#include <cstdint>
#include <cstring>
#include <memory>
#include <vector>

#include "cartographer/mapping/3d/hybrid_grid.h"
#include "cartographer/mapping/3d/range_data_inserter_3d.h"
#include "cartographer/mapping/internal/3d/scan_matching/fast_correlative_scan_matcher_3d.h"
#include "cartographer/mapping/internal/3d/scan_matching/precomputation_grid_3d.h"
#include "cartographer/mapping/internal/3d/scan_matching/real_time_correlative_scan_matcher_3d.h"
#include "cartographer/mapping/internal/3d/scan_matching/rotational_scan_matcher.h"
#include "cartographer/mapping/trajectory_node.h"
#include "cartographer/sensor/point_cloud.h"

namespace carto = ::cartographer;
using carto::mapping::HybridGrid;
using carto::mapping::scan_matching::FastCorrelativeScanMatcher3D;
using carto::mapping::scan_matching::PrecomputationGrid3D;
using carto::mapping::scan_matching::PrecomputationGridStack3D;
using carto::mapping::scan_matching::RealTimeCorrelativeScanMatcher3D;
using carto::mapping::scan_matching::RotationalScanMatcher;

namespace {

struct Matcher3DHolder {
  Eigen::VectorXf submap_histogram;
  carto::mapping::scan_matching::proto::FastCorrelativeScanMatcherOptions3D options;
  std::unique_ptr<FastCorrelativeScanMatcher3D> matcher;
  std::unique_ptr<PrecomputationGridStack3D> stack;  // the same construction, for level read-back
};

carto::sensor::PointCloud ToCloud3(const float* xyz, int n) {
  carto::sensor::PointCloud cloud;
  for (int i = 0; i < n; ++i) cloud.push_back({Eigen::Vector3f(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2])});
  return cloud;
}
Eigen::VectorXf ToVector(const float* v, int n) {
  Eigen::VectorXf out(n);
  for (int i = 0; i < n; ++i) out[i] = v[i];
  return out;
}
carto::transform::Rigid3d Pose7(const float* p) {  // x y z, qw qx qy qz
  return carto::transform::Rigid3d(Eigen::Vector3d(p[0], p[1], p[2]), Eigen::Quaterniond(p[3], p[4], p[5], p[6]));
}
void StorePose7(const carto::transform::Rigid3d& pose, float* out) {
  out[0] = static_cast<float>(pose.translation().x());
  out[1] = static_cast<float>(pose.translation().y());
  out[2] = static_cast<float>(pose.translation().z());
  out[3] = static_cast<float>(pose.rotation().w());
  out[4] = static_cast<float>(pose.rotation().x());
  out[5] = static_cast<float>(pose.rotation().y());
  out[6] = static_cast<float>(pose.rotation().z());
}

}  // namespace

extern "C" {

struct oracle_result3d {
  int32_t found;
  int32_t scan_index;
  int32_t offset[3];
  int32_t num_scans;
  float score, rotational_score, low_resolution_score;
  float pose[7];
  uint64_t candidates_scored;
};

// A HybridGrid holding the given uint16 values at the given cell indices (values written as they
// are, like a grid that was filled by the inserter and finished its update).
void* oracle_grid3d_create(float resolution, const int32_t* xyz, const uint16_t* values, int n) {
  HybridGrid* g = new HybridGrid(resolution);
  for (int i = 0; i < n; ++i) {
    *g->mutable_value(Eigen::Array3i(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2])) = values[i];
  }
  return g;
}
void oracle_grid3d_destroy(void* g) { delete static_cast<HybridGrid*>(g); }
int oracle_grid3d_size(void* g) { return static_cast<HybridGrid*>(g)->grid_size(); }

// Non-default voxels through the reference's own iterator; xyz / values may be NULL to count.
int oracle_grid3d_cells(void* g, int32_t* xyz, uint16_t* values) {
  int n = 0;
  for (auto it = HybridGrid::Iterator(*static_cast<HybridGrid*>(g)); !it.Done(); it.Next()) {
    if (xyz != nullptr) {
      const Eigen::Array3i c = it.GetCellIndex();
      xyz[3 * n] = c.x();
      xyz[3 * n + 1] = c.y();
      xyz[3 * n + 2] = c.z();
      values[n] = it.GetValue();
    }
    ++n;
  }
  return n;
}

// RangeDataInserter3D::Insert of one range data (no intensity grid).  The fork's debugging
// printout (hgrid_info, range_data_inserter_3d.cc:26-43) goes to stdout as in the reference.
void oracle_grid3d_insert(void* g, const float* origin3, const float* returns, int n, float hit_probability,
                          float miss_probability, int num_free_space_voxels) {
  carto::mapping::proto::RangeDataInserterOptions3D o;
  o.set_hit_probability(hit_probability);
  o.set_miss_probability(miss_probability);
  o.set_num_free_space_voxels(num_free_space_voxels);
  carto::sensor::RangeData rd;
  rd.origin = Eigen::Vector3f(origin3[0], origin3[1], origin3[2]);
  rd.returns = ToCloud3(returns, n);
  carto::mapping::RangeDataInserter3D(o).Insert(rd, static_cast<HybridGrid*>(g), nullptr);
}

void* oracle_fcsm3d_create(void* grid, void* low_grid, const float* hist, int n_hist, const double* o7) {
  Matcher3DHolder* h = new Matcher3DHolder;
  h->submap_histogram = ToVector(hist, n_hist);
  h->options.set_branch_and_bound_depth(static_cast<int>(o7[0]));
  h->options.set_full_resolution_depth(static_cast<int>(o7[1]));
  h->options.set_min_rotational_score(o7[2]);
  h->options.set_min_low_resolution_score(o7[3]);
  h->options.set_linear_xy_search_window(o7[4]);
  h->options.set_linear_z_search_window(o7[5]);
  h->options.set_angular_search_window(o7[6]);
  h->matcher = std::make_unique<FastCorrelativeScanMatcher3D>(
      *static_cast<HybridGrid*>(grid), static_cast<HybridGrid*>(low_grid), &h->submap_histogram, h->options);
  h->stack = std::make_unique<PrecomputationGridStack3D>(*static_cast<HybridGrid*>(grid), h->options);
  return h;
}
void oracle_fcsm3d_destroy(void* m) { delete static_cast<Matcher3DHolder*>(m); }

// Dense copy of precomputation level `level` (same protocol as oracle_capi.cc's): with out == NULL
// fills the bounding box of the stored cells and returns their number.
int oracle_fcsm3d_level(void* m, int level, int32_t* lo3, int32_t* hi3, uint8_t* out) {
  const PrecomputationGrid3D& g = static_cast<Matcher3DHolder*>(m)->stack->Get(level);
  int count = 0;
  if (out == nullptr) {
    bool first = true;
    for (auto it = PrecomputationGrid3D::Iterator(g); !it.Done(); it.Next()) {
      const Eigen::Array3i c = it.GetCellIndex();
      for (int k = 0; k < 3; ++k) {
        lo3[k] = first ? c[k] : std::min(lo3[k], c[k]);
        hi3[k] = first ? c[k] : std::max(hi3[k], c[k]);
      }
      first = false;
      ++count;
    }
    if (first) {
      for (int k = 0; k < 3; ++k) lo3[k] = hi3[k] = 0;
    }
    return count;
  }
  const int64_t nx = hi3[0] - lo3[0] + 1, ny = hi3[1] - lo3[1] + 1, nz = hi3[2] - lo3[2] + 1;
  std::fill(out, out + nx * ny * nz, uint8_t(0));
  for (auto it = PrecomputationGrid3D::Iterator(g); !it.Done(); it.Next()) {
    const Eigen::Array3i c = it.GetCellIndex();
    ++count;
    if (c.x() < lo3[0] || c.x() > hi3[0] || c.y() < lo3[1] || c.y() > hi3[1] || c.z() < lo3[2] || c.z() > hi3[2]) {
      continue;
    }
    out[((c.z() - lo3[2]) * ny + (c.y() - lo3[1])) * nx + (c.x() - lo3[0])] = it.GetValue();
  }
  return count;
}

void oracle_fcsm3d_match(void* m, int full, const float* node_pose7, const float* submap_pose7, const float* hi,
                         int n_hi, const float* lo, int n_lo, const float* hist, int n_hist,
                         const double* gravity4, float min_score, oracle_result3d* out) {
  const FastCorrelativeScanMatcher3D& matcher = *static_cast<Matcher3DHolder*>(m)->matcher;
  carto::mapping::TrajectoryNode::Data data;
  data.gravity_alignment = Eigen::Quaterniond(gravity4[0], gravity4[1], gravity4[2], gravity4[3]);
  data.high_resolution_point_cloud = ToCloud3(hi, n_hi);
  data.low_resolution_point_cloud = ToCloud3(lo, n_lo);
  data.rotational_scan_matcher_histogram = ToVector(hist, n_hist);
  data.local_pose = carto::transform::Rigid3d::Identity();
  const carto::transform::Rigid3d node_pose = Pose7(node_pose7), submap_pose = Pose7(submap_pose7);
  const std::unique_ptr<FastCorrelativeScanMatcher3D::Result> r =
      full ? matcher.MatchFullSubmap(node_pose.rotation(), submap_pose.rotation(), data, min_score)
           : matcher.Match(node_pose, submap_pose, data, min_score);
  std::memset(out, 0, sizeof(*out));
  out->scan_index = out->offset[0] = out->offset[1] = out->offset[2] = INT32_MIN;
  if (r == nullptr) return;
  out->found = 1;
  out->score = r->score;
  out->rotational_score = r->rotational_score;
  out->low_resolution_score = r->low_resolution_score;
  StorePose7(r->pose_estimate, out->pose);
}

// RealTimeCorrelativeScanMatcher3D::Match.  The reference exposes neither the winner's index nor
// the size of the search space: *best_index = -1, *num_candidates = -1.
float oracle_rt_match3d(void* grid, const double* options4, const float* initial_pose7, const float* pts, int n,
                        float* pose7_out, int* best_index, int* num_candidates) {
  carto::mapping::scan_matching::proto::RealTimeCorrelativeScanMatcherOptions o;
  o.set_linear_search_window(options4[0]);
  o.set_angular_search_window(options4[1]);
  o.set_translation_delta_cost_weight(options4[2]);
  o.set_rotation_delta_cost_weight(options4[3]);
  const RealTimeCorrelativeScanMatcher3D matcher(o);
  carto::transform::Rigid3d pose = carto::transform::Rigid3d::Identity();
  const float score = matcher.Match(Pose7(initial_pose7), ToCloud3(pts, n), *static_cast<HybridGrid*>(grid), &pose);
  StorePose7(pose, pose7_out);
  if (best_index) *best_index = -1;
  if (num_candidates) *num_candidates = -1;
  return score;
}

// RotationalScanMatcher::Match (rotational_scan_matcher.cc:178-189).
void oracle_rotational_match(const float* submap_hist, const float* hist, int n, float initial_angle,
                             const float* angles, int n_angles, float* out) {
  const Eigen::VectorXf submap_histogram = ToVector(submap_hist, n);
  const RotationalScanMatcher matcher(&submap_histogram);
  const std::vector<float> scores =
      matcher.Match(ToVector(hist, n), initial_angle, std::vector<float>(angles, angles + n_angles));
  for (int i = 0; i < n_angles; ++i) out[i] = scores[i];
}

}  // extern "C"
