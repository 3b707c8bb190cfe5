// TEST INFRASTRUCTURE.  Compiles against the REPLACEMENT headers under cartographer-1_b200/host and
// exercises the public symbols the reference's own unit tests use, with the reference tests'
// inputs and expectations:
//   correlative_scan_matcher_test.cc:25-96        SearchParameters, Candidate2D, GenerateRotatedScans,
//                                                 DiscretizeScans (exact expected integers)
//   fast_correlative_scan_matcher_2d_test.cc:37-246  PrecomputationGrid2D{ctor, GetValue, ToScore} for
//                                                 widths 1, 2, 3, 8, 200; CreateFastCorrelativeScanMatcher
//                                                 Options2D; CorrectPose (50 poses) and FullSubmapMatching
//                                                 (20 poses) within the reference's 0.03 tolerance
//   real_time_correlative_scan_matcher_2d_test.cc:125-142  ScoreCandidates on a caller-made candidate,
//                                                 GenerateExhaustiveSearchCandidates
//   PrecomputationGridStack2D{Get, max_depth}; PrecomputationGridStack3D, CreateFastCorrelativeScanMatcher
//   Options3D (fast_correlative_scan_matcher_3d.h:41-64) against a brute-force window maximum.
// Everything runs on the device through libcsm_host.so / libcsm_b200.so; needs a GPU.
// Build and run: tests/test_gpu_boundary_cpp.py.
#include <cmath>
#include <cstdio>
#include <limits>
#include <random>
#include <string>

#include "cartographer/mapping/2d/probability_grid_range_data_inserter_2d.h"
#include "cartographer/mapping/internal/2d/scan_matching/fast_correlative_scan_matcher_2d.h"
#include "cartographer/mapping/internal/2d/scan_matching/real_time_correlative_scan_matcher_2d.h"
#include "cartographer/mapping/internal/3d/scan_matching/fast_correlative_scan_matcher_3d.h"

using namespace cartographer;
using namespace cartographer::mapping;
using namespace cartographer::mapping::scan_matching;

static int g_checks = 0, g_failures = 0;
#define EXPECT_TRUE(c)                                                              \
  do {                                                                              \
    ++g_checks;                                                                     \
    if (!(c)) {                                                                     \
      ++g_failures;                                                                 \
      if (g_failures < 20) std::fprintf(stderr, "FAILED %s:%d: %s\n", __FILE__, __LINE__, #c); \
    }                                                                               \
  } while (0)
#define EXPECT_EQ(a, b) EXPECT_TRUE((a) == (b))
#define EXPECT_LT(a, b) EXPECT_TRUE((a) < (b))
#define EXPECT_NEAR(a, b, tol) EXPECT_TRUE(std::abs(static_cast<double>(a) - static_cast<double>(b)) <= (tol))

static sensor::PointCloud Cloud(std::initializer_list<sensor::Vector3f> points) {
  sensor::PointCloud cloud;
  for (const sensor::Vector3f& p : points) cloud.push_back({p});
  return cloud;
}

static void TestSearchParametersAndCandidate() {
  const SearchParameters search_parameters(4, 5, 0.03, 0.05);
  EXPECT_EQ(5, search_parameters.num_angular_perturbations);
  EXPECT_NEAR(0.03, search_parameters.angular_perturbation_step_size, 1e-9);
  EXPECT_NEAR(0.05, search_parameters.resolution, 1e-9);
  EXPECT_EQ(11, search_parameters.num_scans);
  EXPECT_EQ(11u, search_parameters.linear_bounds.size());
  for (const SearchParameters::LinearBounds& b : search_parameters.linear_bounds) {
    EXPECT_EQ(-4, b.min_x);
    EXPECT_EQ(4, b.max_x);
    EXPECT_EQ(-4, b.min_y);
    EXPECT_EQ(4, b.max_y);
  }
  const Candidate2D candidate(3, 4, -5, search_parameters);
  EXPECT_EQ(3, candidate.scan_index);
  EXPECT_EQ(4, candidate.x_index_offset);
  EXPECT_EQ(-5, candidate.y_index_offset);
  EXPECT_NEAR(0.25, candidate.x, 1e-9);
  EXPECT_NEAR(-0.2, candidate.y, 1e-9);
  EXPECT_NEAR(-0.06, candidate.orientation, 1e-9);
  EXPECT_NEAR(0., candidate.score, 1e-9);
  Candidate2D bigger_candidate(3, 4, 5, search_parameters);
  bigger_candidate.score = 1.;
  EXPECT_LT(candidate, bigger_candidate);
}

static void TestGenerateRotatedScans() {
  const sensor::PointCloud point_cloud = Cloud({{-1.f, 1.f, 0.f}});
  const std::vector<sensor::PointCloud> scans =
      GenerateRotatedScans(point_cloud, SearchParameters(0, 1, M_PI / 2., 0.));
  EXPECT_EQ(3u, scans.size());
  EXPECT_NEAR(1., scans[0][0].position.x(), 1e-6);
  EXPECT_NEAR(1., scans[0][0].position.y(), 1e-6);
  EXPECT_NEAR(-1., scans[1][0].position.x(), 1e-6);
  EXPECT_NEAR(1., scans[1][0].position.y(), 1e-6);
  EXPECT_NEAR(-1., scans[2][0].position.x(), 1e-6);
  EXPECT_NEAR(-1., scans[2][0].position.y(), 1e-6);
}

static const sensor::PointCloud& SevenPoints() {
  static const sensor::PointCloud cloud =
      Cloud({{0.025f, 0.175f, 0.f}, {-0.025f, 0.175f, 0.f}, {-0.075f, 0.175f, 0.f}, {-0.125f, 0.175f, 0.f},
             {-0.125f, 0.125f, 0.f}, {-0.125f, 0.075f, 0.f}, {-0.125f, 0.025f, 0.f}});
  return cloud;
}

static void TestDiscretizeScans() {
  const MapLimits map_limits(0.05, transform::Vector2d(0.05, 0.25), CellLimits(6, 6));
  const std::vector<sensor::PointCloud> scans =
      GenerateRotatedScans(SevenPoints(), SearchParameters(0, 0, 0., 0.));
  const std::vector<DiscreteScan2D> discrete_scans =
      DiscretizeScans(map_limits, scans, Eigen::Translation2f::Identity());
  EXPECT_EQ(1u, discrete_scans.size());
  EXPECT_EQ(7u, discrete_scans[0].size());
  const int want[7][2] = {{1, 0}, {1, 1}, {1, 2}, {1, 3}, {2, 3}, {3, 3}, {4, 3}};
  for (int i = 0; i < 7; ++i)
    EXPECT_TRUE((Eigen::Array2i(want[i][0], want[i][1]) == discrete_scans[0][i]).all());
}

// PrecomputationGridTest.CorrectValues / TinyProbabilityGrid: every cell of a precomputation grid
// of `width` equals the brute-force maximum probability over its width x width window.
static void TestPrecomputationGrid(const MapLimits& limits, const Eigen::Array2i& first,
                                   std::initializer_list<int> widths) {
  std::mt19937 prng(42);
  std::uniform_int_distribution<int> distribution(0, 255);
  ValueConversionTables conversion_tables;
  ProbabilityGrid probability_grid(limits, &conversion_tables);
  std::vector<float> reusable_intermediate_grid;
  PrecomputationGrid2D precomputation_grid_dummy(probability_grid, probability_grid.limits().cell_limits(), 1,
                                                 &reusable_intermediate_grid);
  const CellLimits& cells = probability_grid.limits().cell_limits();
  for (const Eigen::Array2i& xy_index :
       XYIndexRangeIterator(first, Eigen::Array2i(cells.num_x_cells - 1, cells.num_y_cells - 1))) {
    probability_grid.SetProbability(xy_index, precomputation_grid_dummy.ToScore(distribution(prng)));
  }
  for (const int width : widths) {
    PrecomputationGrid2D precomputation_grid(probability_grid, cells, width, &reusable_intermediate_grid);
    for (const Eigen::Array2i& xy_index : XYIndexRangeIterator(cells)) {
      float max_score = -std::numeric_limits<float>::infinity();
      for (int y = 0; y != width; ++y)
        for (int x = 0; x != width; ++x)
          max_score = std::max<float>(max_score, probability_grid.GetProbability(xy_index + Eigen::Array2i(x, y)));
      EXPECT_NEAR(max_score, precomputation_grid.ToScore(precomputation_grid.GetValue(xy_index)), 1e-4);
    }
    // cells the reference's wide grid also holds: window starts left of / above the grid
    EXPECT_EQ(0, precomputation_grid.GetValue(Eigen::Array2i(-width, 0)));
    EXPECT_EQ(precomputation_grid.GetValue(Eigen::Array2i(0, 0)) >= 0, true);
  }
  // The stack hands out the same levels (width 1 << i).
  scan_matching::proto::FastCorrelativeScanMatcherOptions2D options;
  options.set_branch_and_bound_depth(4);
  PrecomputationGridStack2D stack(probability_grid, options);
  EXPECT_EQ(3, stack.max_depth());
  for (int i = 0; i <= stack.max_depth(); ++i) {
    PrecomputationGrid2D direct(probability_grid, cells, 1 << i, &reusable_intermediate_grid);
    const int w = 1 << i;
    for (const Eigen::Array2i& xy_index :
         XYIndexRangeIterator(Eigen::Array2i(-w, -w), Eigen::Array2i(cells.num_x_cells, cells.num_y_cells)))
      EXPECT_EQ(direct.GetValue(xy_index), stack.Get(i).GetValue(xy_index));
  }
}

static scan_matching::proto::FastCorrelativeScanMatcherOptions2D TestOptions2D(const int branch_and_bound_depth) {
  auto parameter_dictionary = common::MakeDictionary(R"text(
      return {
         linear_search_window = 3.,
         angular_search_window = 1.,
         branch_and_bound_depth = )text" + std::to_string(branch_and_bound_depth) + "}");
  return CreateFastCorrelativeScanMatcherOptions2D(parameter_dictionary.get());
}

static mapping::proto::ProbabilityGridRangeDataInserterOptions2D InserterOptions() {
  mapping::proto::ProbabilityGridRangeDataInserterOptions2D options;
  options.set_insert_free_space(true);
  options.set_hit_probability(0.7);
  options.set_miss_probability(0.4);
  return options;
}

struct Pose2f {
  float x, y, angle;
};
// sensor::TransformPointCloud(cloud, Embed3D(pose)) in fp32 (the tests compare within 0.03).
static sensor::PointCloud Transform(const sensor::PointCloud& cloud, const Pose2f& pose) {
  const float c = std::cos(pose.angle), s = std::sin(pose.angle);
  sensor::PointCloud out;
  for (const sensor::RangefinderPoint& p : cloud)
    out.push_back({sensor::Vector3f(c * p.position.x() - s * p.position.y() + pose.x,
                                    s * p.position.x() + c * p.position.y() + pose.y, p.position.z())});
  return out;
}
static Pose2f Compose(const Pose2f& a, const Pose2f& b) {
  const float c = std::cos(a.angle), s = std::sin(a.angle);
  return Pose2f{c * b.x - s * b.y + a.x, s * b.x + c * b.y + a.y, a.angle + b.angle};
}
static Pose2f Inverse(const Pose2f& a) {
  const float c = std::cos(-a.angle), s = std::sin(-a.angle);
  return Pose2f{-(c * a.x - s * a.y), -(s * a.x + c * a.y), -a.angle};
}
// transform::IsNearly (transform/rigid_transform_test_helpers.h:31-46): Eigen's isApprox on the 3x3
// affine matrices, |a - b|_F^2 <= tol^2 * min(|a|_F^2, |b|_F^2), in fp32 like the reference test.
static bool IsNearly(const Pose2f& want, const transform::Rigid2d& got, const float tol) {
  const float ga = static_cast<float>(got.rotation().angle());
  const float m0[9] = {std::cos(want.angle), -std::sin(want.angle), want.x, std::sin(want.angle),
                       std::cos(want.angle), want.y, 0.f, 0.f, 1.f};
  const float m1[9] = {std::cos(ga), -std::sin(ga), static_cast<float>(got.translation().x()), std::sin(ga),
                       std::cos(ga), static_cast<float>(got.translation().y()), 0.f, 0.f, 1.f};
  float d2 = 0.f, n0 = 0.f, n1 = 0.f;
  for (int i = 0; i < 9; ++i) {
    d2 += (m0[i] - m1[i]) * (m0[i] - m1[i]);
    n0 += m0[i] * m0[i];
    n1 += m1[i] * m1[i];
  }
  return d2 <= tol * tol * std::min(n0, n1);
}

static void TestCorrectPose() {
  std::mt19937 prng(42);
  std::uniform_real_distribution<float> distribution(-1.f, 1.f);
  const ProbabilityGridRangeDataInserter2D range_data_inserter(InserterOptions());
  constexpr float kMinScore = 0.1f;
  const auto options = TestOptions2D(3);
  EXPECT_EQ(3, options.branch_and_bound_depth());
  EXPECT_NEAR(3., options.linear_search_window(), 0.);
  const sensor::PointCloud point_cloud = Cloud({{-2.5f, 0.5f, 0.f}, {-2.f, 0.5f, 0.f}, {0.f, -0.5f, 0.f},
                                                {0.5f, -1.6f, 0.f}, {2.5f, 0.5f, 0.f}, {2.5f, 1.7f, 0.f}});
  for (int i = 0; i != 50; ++i) {
    Pose2f expected_pose;
    expected_pose.x = 2. * distribution(prng);
    expected_pose.y = 2. * distribution(prng);
    expected_pose.angle = 0.5 * distribution(prng);
    ValueConversionTables conversion_tables;
    ProbabilityGrid probability_grid(MapLimits(0.05, transform::Vector2d(5., 5.), CellLimits(200, 200)),
                                     &conversion_tables);
    range_data_inserter.Insert(sensor::RangeData{sensor::Vector3f(expected_pose.x, expected_pose.y, 0.f),
                                                 Transform(point_cloud, expected_pose),
                                                 {}},
                               &probability_grid);
    FastCorrelativeScanMatcher2D fast_correlative_scan_matcher(probability_grid, options);
    transform::Rigid2d pose_estimate;
    float score;
    EXPECT_TRUE(fast_correlative_scan_matcher.Match(transform::Rigid2d::Identity(), point_cloud, kMinScore, &score,
                                                    &pose_estimate));
    EXPECT_LT(kMinScore, score);
    EXPECT_TRUE(IsNearly(expected_pose, pose_estimate, 0.03f));
  }
}

static void TestFullSubmapMatching() {
  std::mt19937 prng(42);
  std::uniform_real_distribution<float> distribution(-1.f, 1.f);
  const ProbabilityGridRangeDataInserter2D range_data_inserter(InserterOptions());
  constexpr float kMinScore = 0.1f;
  const auto options = TestOptions2D(6);
  const sensor::PointCloud unperturbed = Cloud({{-2.5f, 0.5f, 0.f}, {-2.25f, 0.5f, 0.f}, {0.f, 0.5f, 0.f},
                                                {0.25f, 1.6f, 0.f}, {2.5f, 0.5f, 0.f}, {2.f, 1.8f, 0.f}});
  for (int i = 0; i != 20; ++i) {
    Pose2f perturbation;
    perturbation.x = 10. * distribution(prng);
    perturbation.y = 10. * distribution(prng);
    perturbation.angle = 1.6 * distribution(prng);
    const sensor::PointCloud point_cloud = Transform(unperturbed, perturbation);
    Pose2f base;
    base.x = 2. * distribution(prng);
    base.y = 2. * distribution(prng);
    base.angle = 0.5 * distribution(prng);
    const Pose2f expected_pose = Compose(base, Inverse(perturbation));
    const Pose2f origin = Compose(expected_pose, perturbation);
    ValueConversionTables conversion_tables;
    ProbabilityGrid probability_grid(MapLimits(0.05, transform::Vector2d(5., 5.), CellLimits(200, 200)),
                                     &conversion_tables);
    range_data_inserter.Insert(
        sensor::RangeData{sensor::Vector3f(origin.x, origin.y, 0.f), Transform(point_cloud, expected_pose), {}},
        &probability_grid);
    FastCorrelativeScanMatcher2D fast_correlative_scan_matcher(probability_grid, options);
    transform::Rigid2d pose_estimate;
    float score;
    EXPECT_TRUE(fast_correlative_scan_matcher.MatchFullSubmap(point_cloud, kMinScore, &score, &pose_estimate));
    EXPECT_LT(kMinScore, score);
    EXPECT_TRUE(IsNearly(expected_pose, pose_estimate, 0.03f));
  }
}

static void TestRealTimeScoreCandidates() {
  scan_matching::proto::RealTimeCorrelativeScanMatcherOptions options;
  options.set_linear_search_window(0.6);
  options.set_angular_search_window(0.16);
  options.set_translation_delta_cost_weight(0.);
  options.set_rotation_delta_cost_weight(0.);
  const RealTimeCorrelativeScanMatcher2D matcher(options);
  ValueConversionTables conversion_tables;
  ProbabilityGrid grid(MapLimits(0.05, transform::Vector2d(0.05, 0.25), CellLimits(6, 6)), &conversion_tables);
  const ProbabilityGridRangeDataInserter2D range_data_inserter(InserterOptions());
  range_data_inserter.Insert(sensor::RangeData{sensor::Vector3f::Zero(), SevenPoints(), {}}, &grid);
  const std::vector<sensor::PointCloud> scans = GenerateRotatedScans(SevenPoints(), SearchParameters(0, 0, 0., 0.));
  const std::vector<DiscreteScan2D> discrete_scans =
      DiscretizeScans(grid.limits(), scans, Eigen::Translation2f::Identity());
  std::vector<Candidate2D> candidates;
  candidates.emplace_back(0, 0, 0, SearchParameters(0, 0, 0., 0.));
  matcher.ScoreCandidates(grid, discrete_scans, SearchParameters(0, 0, 0., 0.), &candidates);
  EXPECT_EQ(0, candidates[0].scan_index);
  EXPECT_EQ(0, candidates[0].x_index_offset);
  EXPECT_EQ(0, candidates[0].y_index_offset);
  EXPECT_NEAR(0.7, candidates[0].score, 1e-2);  // every point aligns perfectly
  // ScoreCandidates over the exhaustive list agrees with Match (first maximum wins).
  const SearchParameters sp(0.1, 0.05, SevenPoints(), 0.05);
  const std::vector<sensor::PointCloud> all = GenerateRotatedScans(SevenPoints(), sp);
  const std::vector<DiscreteScan2D> all_discrete = DiscretizeScans(grid.limits(), all, Eigen::Translation2f::Identity());
  std::vector<Candidate2D> list = matcher.GenerateExhaustiveSearchCandidates(sp);
  EXPECT_EQ(static_cast<size_t>(sp.num_scans) * 25, list.size());
  matcher.ScoreCandidates(grid, all_discrete, sp, &list);
  size_t best = 0;
  for (size_t c = 1; c < list.size(); ++c)
    if (list[c].score > list[best].score) best = c;
  scan_matching::proto::RealTimeCorrelativeScanMatcherOptions narrow = options;
  narrow.set_linear_search_window(0.1);
  narrow.set_angular_search_window(0.05);
  const RealTimeCorrelativeScanMatcher2D narrow_matcher(narrow);
  transform::Rigid2d pose;
  const double score = narrow_matcher.Match(transform::Rigid2d::Identity(), SevenPoints(), grid, &pose);
  EXPECT_EQ(static_cast<float>(score), list[best].score);
  EXPECT_NEAR(pose.translation().x(), list[best].x, 1e-12);
  EXPECT_NEAR(pose.translation().y(), list[best].y, 1e-12);
}

static void TestPrecomputationGridStack3D() {
  HybridGrid hybrid_grid(2.f);
  std::mt19937 rng(23847);
  std::uniform_int_distribution<int> coordinate_distribution(-50, 49);
  std::uniform_real_distribution<float> value_distribution(kMinProbability, kMaxProbability);
  for (int i = 0; i < 1000; ++i) {
    const int x = coordinate_distribution(rng), y = coordinate_distribution(rng), z = coordinate_distribution(rng);
    hybrid_grid.SetProbability(Eigen::Array3i(x, y, z), value_distribution(rng));
  }
  auto parameter_dictionary = common::MakeDictionary(
      "return { branch_and_bound_depth = 4, full_resolution_depth = 4, min_rotational_score = 0.1, "
      "min_low_resolution_score = 0.15, linear_xy_search_window = 0.8, linear_z_search_window = 0.8, "
      "angular_search_window = 0.3 }");
  const auto options = CreateFastCorrelativeScanMatcherOptions3D(parameter_dictionary.get());
  EXPECT_EQ(4, options.branch_and_bound_depth());
  EXPECT_NEAR(0.15, options.min_low_resolution_score(), 0.);
  const PrecomputationGridStack3D stack(hybrid_grid, options);
  EXPECT_EQ(3, stack.max_depth());
  const std::vector<float>& to_probability = ValueToProbabilityTable();
  for (int depth = 0; depth <= 3; ++depth) {
    const int width = 1 << depth;
    for (int i = 0; i < 100; ++i) {
      const int x = coordinate_distribution(rng), y = coordinate_distribution(rng), z = coordinate_distribution(rng);
      float max_probability = 0.;
      for (int dx = 0; dx < width; ++dx)
        for (int dy = 0; dy < width; ++dy)
          for (int dz = 0; dz < width; ++dz)
            max_probability = std::max(
                max_probability, to_probability[hybrid_grid.value(Eigen::Array3i(x + dx, y + dy, z + dz))]);
      EXPECT_NEAR(max_probability,
                  PrecomputationGrid3D::ToProbability(stack.Get(depth).value(Eigen::Array3i(x, y, z))), 1e-2);
    }
  }
}

int main() {
  if (csm_device_count() < 1) {
    std::fprintf(stderr, "no CUDA device: the scan matchers have no CPU fallback\n");
    return 2;
  }
  TestSearchParametersAndCandidate();
  TestGenerateRotatedScans();
  TestDiscretizeScans();
  TestPrecomputationGrid(MapLimits(0.05, transform::Vector2d(5., 5.), CellLimits(250, 250)), Eigen::Array2i(50, 50),
                         {1, 2, 3, 8});
  TestPrecomputationGrid(MapLimits(0.05, transform::Vector2d(0.1, 0.1), CellLimits(4, 4)), Eigen::Array2i(0, 0),
                         {1, 2, 3, 8, 200});
  TestCorrectPose();
  TestFullSubmapMatching();
  TestRealTimeScoreCandidates();
  TestPrecomputationGridStack3D();
  std::printf("boundary tests: %d checks, %d failures\n", g_checks, g_failures);
  return g_failures == 0 ? 0 : 1;
}
