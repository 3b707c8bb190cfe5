// Host-side search-space set-up for the B200 scan matchers.
//
// Same names and meaning as the reference's
// cartographer/mapping/internal/2d/scan_matching/correlative_scan_matcher_2d.h:35-103
// (SearchParameters, Candidate2D).  These are the O(num_scans), libm-dependent
// pieces that stay on the host so that glibc's sqrtf/acos/ceil/cosf/sinf are the
// ones the reference itself would call; everything O(num_scans x num_points) and
// beyond runs on the device behind include/csm_b200.h.
#ifndef CARTOGRAPHER_MAPPING_INTERNAL_2D_SCAN_MATCHING_CORRELATIVE_SCAN_MATCHER_2D_H_
#define CARTOGRAPHER_MAPPING_INTERNAL_2D_SCAN_MATCHING_CORRELATIVE_SCAN_MATCHER_2D_H_

#include <vector>

#include "cartographer_compat.h"

namespace cartographer {
namespace mapping {
namespace scan_matching {

// Describes the search space (reference: correlative_scan_matcher_2d.h:35-60).
struct SearchParameters {
  struct LinearBounds {
    int min_x;
    int max_x;
    int min_y;
    int max_y;
  };

  SearchParameters(double linear_search_window, double angular_search_window,
                   const sensor::PointCloud& point_cloud, double resolution);
  // For testing.
  SearchParameters(int num_linear_perturbations, int num_angular_perturbations,
                   double angular_perturbation_step_size, double resolution);

  // (cosf(a/2), sinf(a/2)) of the fp32-narrowed rotation of every scan, with the
  // double accumulation of delta_theta of GenerateRotatedScans
  // (reference: correlative_scan_matcher_2d.cc:99-106).  2 * num_scans floats.
  std::vector<float> ScanQuaternionTable() const;

  int num_angular_perturbations;
  double angular_perturbation_step_size;
  double resolution;
  int num_scans;
  std::vector<LinearBounds> linear_bounds;  // per rotated scan, before ShrinkToFit
};

typedef std::vector<Eigen::Array2i> DiscreteScan2D;  // reference :33

// Generates a collection of rotated scans (reference :62-65, .cc:93-109).  Runs on the calling
// thread's default device context (csm_rotate_scans2d); the matchers do not call it -- their
// kernels rotate, discretise and score in one pass -- it exists for callers and tests that use
// the reference's public helper.
std::vector<sensor::PointCloud> GenerateRotatedScans(const sensor::PointCloud& point_cloud,
                                                     const SearchParameters& search_parameters);

// Translates and discretizes the rotated scans into a vector of integer indices (reference
// :67-72, .cc:111-127; csm_discretize_points2d).
std::vector<DiscreteScan2D> DiscretizeScans(const MapLimits& map_limits,
                                            const std::vector<sensor::PointCloud>& scans,
                                            const Eigen::Translation2f& initial_translation);

// A possible solution (reference: correlative_scan_matcher_2d.h:74-103).
struct Candidate2D {
  Candidate2D(const int init_scan_index, const int init_x_index_offset,
              const int init_y_index_offset, const SearchParameters& search_parameters)
      : scan_index(init_scan_index),
        x_index_offset(init_x_index_offset),
        y_index_offset(init_y_index_offset),
        x(-y_index_offset * search_parameters.resolution),
        y(-x_index_offset * search_parameters.resolution),
        orientation((scan_index - search_parameters.num_angular_perturbations) *
                    search_parameters.angular_perturbation_step_size) {}

  int scan_index = 0;
  int x_index_offset = 0;
  int y_index_offset = 0;
  double x = 0.;
  double y = 0.;
  double orientation = 0.;
  float score = 0.f;

  bool operator<(const Candidate2D& other) const { return score < other.score; }
  bool operator>(const Candidate2D& other) const { return score > other.score; }
};

}  // namespace scan_matching
}  // namespace mapping
}  // namespace cartographer

#endif  // CARTOGRAPHER_MAPPING_INTERNAL_2D_SCAN_MATCHING_CORRELATIVE_SCAN_MATCHER_2D_H_
