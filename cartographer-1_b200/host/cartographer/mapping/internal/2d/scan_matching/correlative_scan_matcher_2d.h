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

// The search space of one matcher call (the reference's struct of the same name, :35-60): how many
// rotated copies of the scan there are, how far apart in angle, and per copy the window of cell
// offsets that is searched.
struct SearchParameters {
  struct LinearBounds {
    int min_x, max_x, min_y, max_y;  // inclusive cell offsets
  };

  // public data, as in the reference
  int num_angular_perturbations;
  double angular_perturbation_step_size;
  double resolution;
  int num_scans;
  std::vector<LinearBounds> linear_bounds;  // one entry per rotated scan; ShrinkToFit runs on the device

  // From the matcher options and the cloud (its farthest point sets the angular step, .cc:27-53).
  SearchParameters(double linear_search_window, double angular_search_window,
                   const sensor::PointCloud& point_cloud, double resolution);
  // From explicit counts (the reference's test-only constructor, .cc:55-69).
  SearchParameters(int num_linear_perturbations, int num_angular_perturbations,
                   double angular_perturbation_step_size, double resolution);

  // What the device needs of the rotations: (cosf(a / 2), sinf(a / 2)) of every scan's angle a,
  // with a accumulated in double and narrowed to float exactly as GenerateRotatedScans does
  // (.cc:99-106).  2 * num_scans floats.
  std::vector<float> ScanQuaternionTable() const;
};

using DiscreteScan2D = std::vector<Eigen::Array2i>;  // :33

// The two public helpers of the reference header (:62-72, .cc:93-127), for callers and tests that
// use them: the matchers themselves rotate, discretise and score inside one kernel pass.  Both run
// on the calling thread's default device context (csm_rotate_scans2d, csm_discretize_points2d).
std::vector<sensor::PointCloud> GenerateRotatedScans(const sensor::PointCloud& point_cloud,
                                                     const SearchParameters& search_parameters);
std::vector<DiscreteScan2D> DiscretizeScans(const MapLimits& map_limits,
                                            const std::vector<sensor::PointCloud>& scans,
                                            const Eigen::Translation2f& initial_translation);

// One pose hypothesis (:74-103): which rotated scan, which cell offset, and the same as metres /
// radians relative to the initial pose estimate; ordered by score.
struct Candidate2D {
  int scan_index = 0;
  int x_index_offset = 0;
  int y_index_offset = 0;
  double x = 0.;  // metres; the grid's x axis runs along -y of the cell index
  double y = 0.;
  double orientation = 0.;
  float score = 0.f;

  Candidate2D(const int scan, const int x_offset, const int y_offset, const SearchParameters& sp)
      : scan_index(scan), x_index_offset(x_offset), y_index_offset(y_offset) {
    x = -y_index_offset * sp.resolution;
    y = -x_index_offset * sp.resolution;
    orientation = (scan_index - sp.num_angular_perturbations) * sp.angular_perturbation_step_size;
  }

  bool operator<(const Candidate2D& rhs) const { return score < rhs.score; }
  bool operator>(const Candidate2D& rhs) const { return score > rhs.score; }
};

}  // namespace scan_matching
}  // namespace mapping
}  // namespace cartographer

#endif  // CARTOGRAPHER_MAPPING_INTERNAL_2D_SCAN_MATCHING_CORRELATIVE_SCAN_MATCHER_2D_H_
