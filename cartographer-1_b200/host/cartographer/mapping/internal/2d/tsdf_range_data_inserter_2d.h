// B200 implementation behind the reference's TSDFRangeDataInserter2D interface
// (cartographer/mapping/internal/2d/tsdf_range_data_inserter_2d.h:30-58) with its normal
// estimation (normal_estimation_2d.h:28-37).
//
// Insert(range_data, grid) keeps the reference's signature and result.  The host does what is per
// return and libm-bound, with the reference's expressions: GrowAsNeeded, the bearing sort,
// EstimateNormals, cosf / sinf of the normal, the range and angle weight factors.  Every per-cell
// step -- which return owns a cell, the projected distance, the distance kernel, the weighted
// running mean, quantisation -- runs on the device (csm_tsdfgrid2d_insert).  DeviceGrid keeps a
// live TSDF submap's two cell arrays in HBM between insertions.  No CPU fallback.
#ifndef CARTOGRAPHER_MAPPING_INTERNAL_2D_TSDF_RANGE_DATA_INSERTER_2D_H_
#define CARTOGRAPHER_MAPPING_INTERNAL_2D_TSDF_RANGE_DATA_INSERTER_2D_H_

#include <vector>

#include "cartographer/mapping/internal/2d/scan_matching/fast_correlative_scan_matcher_2d.h"
#include "cartographer_compat.h"
#include "csm_b200.h"

namespace cartographer {
namespace mapping {

namespace proto {
// mapping/proto/normal_estimation_options_2d.proto
class NormalEstimationOptions2D {
 public:
  int num_normal_samples() const { return num_normal_samples_; }
  float sample_radius() const { return sample_radius_; }
  void set_num_normal_samples(int v) { num_normal_samples_ = v; }
  void set_sample_radius(float v) { sample_radius_ = v; }

 private:
  int num_normal_samples_ = 0;
  float sample_radius_ = 0.f;
};
// mapping/proto/tsdf_range_data_inserter_options_2d.proto
class TSDFRangeDataInserterOptions2D {
 public:
  double truncation_distance() const { return truncation_distance_; }
  double maximum_weight() const { return maximum_weight_; }
  bool update_free_space() const { return update_free_space_; }
  const NormalEstimationOptions2D& normal_estimation_options() const { return normal_estimation_options_; }
  NormalEstimationOptions2D* mutable_normal_estimation_options() { return &normal_estimation_options_; }
  bool project_sdf_distance_to_scan_normal() const { return project_sdf_distance_to_scan_normal_; }
  int update_weight_range_exponent() const { return update_weight_range_exponent_; }
  double update_weight_angle_scan_normal_to_ray_kernel_bandwidth() const { return angle_bandwidth_; }
  double update_weight_distance_cell_to_hit_kernel_bandwidth() const { return distance_bandwidth_; }
  void set_truncation_distance(double v) { truncation_distance_ = v; }
  void set_maximum_weight(double v) { maximum_weight_ = v; }
  void set_update_free_space(bool v) { update_free_space_ = v; }
  void set_project_sdf_distance_to_scan_normal(bool v) { project_sdf_distance_to_scan_normal_ = v; }
  void set_update_weight_range_exponent(int v) { update_weight_range_exponent_ = v; }
  void set_update_weight_angle_scan_normal_to_ray_kernel_bandwidth(double v) { angle_bandwidth_ = v; }
  void set_update_weight_distance_cell_to_hit_kernel_bandwidth(double v) { distance_bandwidth_ = v; }

 private:
  double truncation_distance_ = 0., maximum_weight_ = 0., angle_bandwidth_ = 0., distance_bandwidth_ = 0.;
  bool update_free_space_ = false, project_sdf_distance_to_scan_normal_ = false;
  int update_weight_range_exponent_ = 0;
  NormalEstimationOptions2D normal_estimation_options_;
};
}  // namespace proto

// normal_estimation_2d.cc:68-78, tsdf_range_data_inserter_2d.cc:107-127
proto::NormalEstimationOptions2D CreateNormalEstimationOptions2D(common::LuaParameterDictionary* parameter_dictionary);
proto::TSDFRangeDataInserterOptions2D CreateTSDFRangeDataInserterOptions2D(
    common::LuaParameterDictionary* parameter_dictionary);

// Estimates the normal for each 'return' in 'range_data' (normal_estimation_2d.cc:81-110); the
// returns must be sorted by bearing around the origin.
std::vector<float> EstimateNormals(const sensor::RangeData& range_data,
                                   const proto::NormalEstimationOptions2D& normal_estimation_options);

using scan_matching::DeviceContext;

class TSDFRangeDataInserter2D {
 public:
  explicit TSDFRangeDataInserter2D(const proto::TSDFRangeDataInserterOptions2D& options);
  TSDFRangeDataInserter2D(const proto::TSDFRangeDataInserterOptions2D& options, DeviceContext* context);

  TSDFRangeDataInserter2D(const TSDFRangeDataInserter2D&) = delete;
  TSDFRangeDataInserter2D& operator=(const TSDFRangeDataInserter2D&) = delete;

  // Casts a ray from origin towards hit for each hit in range data (reference :131-165;
  // GridInterface* there).  `grid` must be a TSDF2D.
  void Insert(const sensor::RangeData& range_data, Grid2D* grid) const;

  // GrowAsNeeded (reference :33-48).
  static void GrowAsNeeded(const sensor::RangeData& range_data, float truncation_distance, TSDF2D* tsdf);

  // The two cell arrays of one TSDF2D kept on the device between insertions.
  class DeviceGrid {
   public:
    DeviceGrid(DeviceContext* context, const TSDF2D& tsdf);
    ~DeviceGrid();
    DeviceGrid(const DeviceGrid&) = delete;
    DeviceGrid& operator=(const DeviceGrid&) = delete;
    csm_tsdfgrid2d* get() const { return grid_; }
    const MapLimits& limits() const { return limits_; }
    void Download(TSDF2D* tsdf) const;  // same limits required

   private:
    DeviceContext* context_;
    MapLimits limits_;
    csm_tsdfgrid2d* grid_ = nullptr;
  };
  // Inserts every range data, in order, into the device-resident grid (which must already contain
  // them: call GrowAsNeeded on the host grid for each before creating the DeviceGrid).
  void Insert(const std::vector<const sensor::RangeData*>& range_data, const TSDF2D& layout, DeviceGrid* grid) const;

 private:
  const proto::TSDFRangeDataInserterOptions2D options_;
  DeviceContext* context_ = nullptr;
};

}  // namespace mapping
}  // namespace cartographer

#endif  // CARTOGRAPHER_MAPPING_INTERNAL_2D_TSDF_RANGE_DATA_INSERTER_2D_H_
