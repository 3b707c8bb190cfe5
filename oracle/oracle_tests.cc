// TEST INFRASTRUCTURE ONLY (see csm_oracle.h).
//
// Ports of the reference's own unit tests for the hot path onto the oracle.  The
// oracle is only trusted once every one of these known-answer tests passes.  Each
// test cites the reference test it restates (paths relative to
// /root/reference/cartographer/).  No gtest in this image: a tiny EXPECT harness is
// used instead, and the process exit code is the number of failed expectations.
#include <cstring>
#include <random>
#include <string>

#include "cpu_pool.h"
#include "csm_oracle.h"
#include "csm_oracle3d.h"

namespace {

int g_failures = 0;
int g_checks = 0;
const char* g_test = "";

#define EXPECT_TRUE(c)                                                        \
  do {                                                                        \
    ++g_checks;                                                               \
    if (!(c)) {                                                               \
      ++g_failures;                                                           \
      std::fprintf(stderr, "[%s] EXPECT_TRUE(%s) failed at line %d\n", g_test, #c, __LINE__); \
    }                                                                         \
  } while (0)
#define EXPECT_EQ(a, b) EXPECT_TRUE((a) == (b))
#define EXPECT_LT(a, b) EXPECT_TRUE((a) < (b))
#define EXPECT_GT(a, b) EXPECT_TRUE((a) > (b))
#define EXPECT_NEAR(a, b, tol)                                                  \
  do {                                                                          \
    ++g_checks;                                                                 \
    const double _a = (a), _b = (b);                                            \
    if (!(std::abs(_a - _b) <= (tol))) {                                        \
      ++g_failures;                                                             \
      std::fprintf(stderr, "[%s] EXPECT_NEAR(%s=%.9g, %s=%.9g, %g) failed at line %d\n", g_test, \
                   #a, _a, #b, _b, static_cast<double>(tol), __LINE__);         \
    }                                                                           \
  } while (0)

using namespace oracle;

MapLimits Limits(double res, double max_x, double max_y, int nx, int ny) {
  MapLimits l;
  l.resolution = res;
  l.max_x = max_x;
  l.max_y = max_y;
  l.cell_limits.num_x_cells = nx;
  l.cell_limits.num_y_cells = ny;
  return l;
}

// Fixture-only float rigid transform (transform/rigid_transform.h:34-103,
// transform/transform.h Embed3D) used to build the reference tests' inputs.
struct Rigid2f {
  float tx, ty, angle;
};
Rigid2f Inverse(const Rigid2f& t) {
  const float a = -t.angle;
  const float c = std::cos(a), s = std::sin(a);
  return Rigid2f{-(c * t.tx + -s * t.ty), -(s * t.tx + c * t.ty), a};
}
Rigid2f Mul(const Rigid2f& l, const Rigid2f& r) {
  const float c = std::cos(l.angle), s = std::sin(l.angle);
  return Rigid2f{(c * r.tx + -s * r.ty) + l.tx, (s * r.tx + c * r.ty) + l.ty, l.angle + r.angle};
}
PointCloud TransformEmbed3D(const PointCloud& cloud, const Rigid2f& t) {
  const Quatf q = QuatFromYaw(t.angle);
  PointCloud out;
  for (const Vec3f& p : cloud) {
    Vec3f r = QuatRotate(q, p);
    r.x += t.tx;
    r.y += t.ty;
    r.z += 0.f;
    out.push_back(r);
  }
  return out;
}

// fast_correlative_scan_matcher_2d_test.cc:37-77 and :79-117.
void PrecomputationGridAgainstBruteForce(const char* name, ProbabilityGrid* grid, int x0, int y0,
                                         std::initializer_list<int> widths) {
  g_test = name;
  std::mt19937 prng(42);
  std::uniform_int_distribution<int> distribution(0, 255);
  std::vector<float> scratch;
  const CellLimits cl = grid->limits().cell_limits;
  PrecomputationGrid2D dummy(*grid, cl, 1, &scratch);
  for (int y = y0; y < cl.num_y_cells; ++y)
    for (int x = x0; x < cl.num_x_cells; ++x) {
      Array2i c;
      c.x = x;
      c.y = y;
      grid->SetProbability(c, dummy.ToScore(distribution(prng)));
    }
  scratch.clear();
  for (const int width : widths) {
    PrecomputationGrid2D pg(*grid, cl, width, &scratch);
    for (int y = 0; y < cl.num_y_cells; ++y)
      for (int x = 0; x < cl.num_x_cells; ++x) {
        float max_score = -std::numeric_limits<float>::infinity();
        for (int dy = 0; dy != width; ++dy)
          for (int dx = 0; dx != width; ++dx) {
            Array2i c;
            c.x = x + dx;
            c.y = y + dy;
            max_score = std::max<float>(max_score, grid->GetProbability(c));
          }
        Array2i c;
        c.x = x;
        c.y = y;
        EXPECT_NEAR(max_score, pg.ToScore(pg.GetValue(c)), 1e-4);
      }
  }
}

void TestPrecomputationGridCorrectValues() {
  ProbabilityGrid grid(Limits(0.05, 5., 5., 250, 250));
  PrecomputationGridAgainstBruteForce("PrecomputationGridTest.CorrectValues", &grid, 50, 50,
                                      {1, 2, 3, 8});
}
void TestPrecomputationGridTiny() {
  ProbabilityGrid grid(Limits(0.05, 0.1, 0.1, 4, 4));
  PrecomputationGridAgainstBruteForce("PrecomputationGridTest.TinyProbabilityGrid", &grid, 0, 0,
                                      {1, 2, 3, 8, 200});
}

// correlative_scan_matcher_test.cc:26-40
void TestSearchParametersConstruction() {
  g_test = "SearchParameters.Construction";
  const SearchParameters sp(4, 5, 0.03, 0.05);
  EXPECT_EQ(5, sp.num_angular_perturbations);
  EXPECT_NEAR(0.03, sp.angular_perturbation_step_size, 1e-9);
  EXPECT_NEAR(0.05, sp.resolution, 1e-9);
  EXPECT_EQ(11, sp.num_scans);
  EXPECT_EQ(11u, sp.linear_bounds.size());
  for (const LinearBounds& lb : sp.linear_bounds) {
    EXPECT_EQ(-4, lb.min_x);
    EXPECT_EQ(4, lb.max_x);
    EXPECT_EQ(-4, lb.min_y);
    EXPECT_EQ(4, lb.max_y);
  }
}
// correlative_scan_matcher_test.cc:42-56
void TestCandidateConstruction() {
  g_test = "Candidate.Construction";
  const SearchParameters sp(4, 5, 0.03, 0.05);
  const Candidate2D c(3, 4, -5, sp);
  EXPECT_EQ(3, c.scan_index);
  EXPECT_EQ(4, c.x_index_offset);
  EXPECT_EQ(-5, c.y_index_offset);
  EXPECT_NEAR(0.25, c.x, 1e-9);
  EXPECT_NEAR(-0.2, c.y, 1e-9);
  EXPECT_NEAR(-0.06, c.orientation, 1e-9);
  EXPECT_NEAR(0., c.score, 1e-9);
  Candidate2D bigger(3, 4, 5, sp);
  bigger.score = 1.;
  EXPECT_TRUE(c < bigger);
}
// correlative_scan_matcher_test.cc:58-70
void TestGenerateRotatedScans() {
  g_test = "GenerateRotatedScans";
  PointCloud cloud;
  cloud.push_back(Vec3f{-1.f, 1.f, 0.f});
  const std::vector<PointCloud> scans =
      GenerateRotatedScans(cloud, SearchParameters(0, 1, M_PI / 2., 0.));
  EXPECT_EQ(3u, scans.size());
  EXPECT_NEAR(1., scans[0][0].x, 1e-6);
  EXPECT_NEAR(1., scans[0][0].y, 1e-6);
  EXPECT_NEAR(-1., scans[1][0].x, 1e-6);
  EXPECT_NEAR(1., scans[1][0].y, 1e-6);
  EXPECT_NEAR(-1., scans[2][0].x, 1e-6);
  EXPECT_NEAR(-1., scans[2][0].y, 1e-6);
}
PointCloud SevenPointCloud() {
  PointCloud cloud;
  cloud.push_back(Vec3f{0.025f, 0.175f, 0.f});
  cloud.push_back(Vec3f{-0.025f, 0.175f, 0.f});
  cloud.push_back(Vec3f{-0.075f, 0.175f, 0.f});
  cloud.push_back(Vec3f{-0.125f, 0.175f, 0.f});
  cloud.push_back(Vec3f{-0.125f, 0.125f, 0.f});
  cloud.push_back(Vec3f{-0.125f, 0.075f, 0.f});
  cloud.push_back(Vec3f{-0.125f, 0.025f, 0.f});
  return cloud;
}
// correlative_scan_matcher_test.cc:72-96 -- exact integer cell indices.
void TestDiscretizeScans() {
  g_test = "DiscretizeScans";
  const PointCloud cloud = SevenPointCloud();
  const MapLimits limits = Limits(0.05, 0.05, 0.25, 6, 6);
  const std::vector<PointCloud> scans = GenerateRotatedScans(cloud, SearchParameters(0, 0, 0., 0.));
  const std::vector<DiscreteScan2D> d = DiscretizeScans(limits, scans, 0.f, 0.f);
  EXPECT_EQ(1u, d.size());
  EXPECT_EQ(7u, d[0].size());
  const int expect[7][2] = {{1, 0}, {1, 1}, {1, 2}, {1, 3}, {2, 3}, {3, 3}, {4, 3}};
  for (int i = 0; i < 7; ++i) {
    EXPECT_EQ(expect[i][0], d[0][i].x);
    EXPECT_EQ(expect[i][1], d[0][i].y);
  }
}

bool IsNearly(const Rigid2f& expected, const Rigid2d& actual, float tol) {
  // transform/rigid_transform_test_helpers.h:31-46: Eigen isApprox on the 3x3
  // affine matrices, i.e. |a - b|_F^2 <= tol^2 * min(|a|_F^2, |b|_F^2).
  const float m0[9] = {std::cos(expected.angle), -std::sin(expected.angle), expected.tx,
                       std::sin(expected.angle), std::cos(expected.angle),  expected.ty,
                       0.f, 0.f, 1.f};
  const float aa = static_cast<float>(actual.angle);
  const float m1[9] = {std::cos(aa), -std::sin(aa), static_cast<float>(actual.tx),
                       std::sin(aa), std::cos(aa),  static_cast<float>(actual.ty),
                       0.f, 0.f, 1.f};
  float d2 = 0.f, n0 = 0.f, n1 = 0.f;
  for (int i = 0; i < 9; ++i) {
    d2 += (m0[i] - m1[i]) * (m0[i] - m1[i]);
    n0 += m0[i] * m0[i];
    n1 += m1[i] * m1[i];
  }
  return d2 <= tol * tol * std::min(n0, n1);
}

// fast_correlative_scan_matcher_2d_test.cc:144-192
void TestFastCorrectPose() {
  g_test = "FastCorrelativeScanMatcherTest.CorrectPose";
  std::mt19937 prng(42);
  std::uniform_real_distribution<float> distribution(-1.f, 1.f);
  const ProbabilityGridRangeDataInserter2D inserter(0.7, 0.4, true);
  constexpr float kMinScore = 0.1f;
  FastCorrelativeScanMatcherOptions2D options;
  options.linear_search_window = 3.;
  options.angular_search_window = 1.;
  options.branch_and_bound_depth = 3;
  PointCloud cloud;
  cloud.push_back(Vec3f{-2.5f, 0.5f, 0.f});
  cloud.push_back(Vec3f{-2.f, 0.5f, 0.f});
  cloud.push_back(Vec3f{0.f, -0.5f, 0.f});
  cloud.push_back(Vec3f{0.5f, -1.6f, 0.f});
  cloud.push_back(Vec3f{2.5f, 0.5f, 0.f});
  cloud.push_back(Vec3f{2.5f, 1.7f, 0.f});
  for (int i = 0; i != 50; ++i) {
    const float ex = static_cast<float>(2. * distribution(prng));
    const float ey = static_cast<float>(2. * distribution(prng));
    const float ea = static_cast<float>(0.5 * distribution(prng));
    const Rigid2f expected{ex, ey, ea};
    ProbabilityGrid grid(Limits(0.05, 5., 5., 200, 200));
    RangeData rd;
    rd.origin = Vec3f{expected.tx, expected.ty, 0.f};
    rd.returns = TransformEmbed3D(cloud, expected);
    inserter.Insert(rd, &grid);
    grid.FinishUpdate();
    FastCorrelativeScanMatcher2D matcher(grid, options);
    Rigid2d pose;
    float score = 0.f;
    EXPECT_TRUE(matcher.Match(Rigid2d(), cloud, kMinScore, &score, &pose));
    EXPECT_LT(kMinScore, score);
    EXPECT_TRUE(IsNearly(expected, pose, 0.03f));
    if (!IsNearly(expected, pose, 0.03f)) std::fprintf(stderr, "  exp %.4f %.4f %.4f got %.4f %.4f %.4f score %.4f\n", expected.tx, expected.ty, expected.angle, pose.tx, pose.ty, pose.angle, score);
  }
}

// fast_correlative_scan_matcher_2d_test.cc:194-246
void TestFastFullSubmapMatching() {
  g_test = "FastCorrelativeScanMatcherTest.FullSubmapMatching";
  std::mt19937 prng(42);
  std::uniform_real_distribution<float> distribution(-1.f, 1.f);
  const ProbabilityGridRangeDataInserter2D inserter(0.7, 0.4, true);
  constexpr float kMinScore = 0.1f;
  FastCorrelativeScanMatcherOptions2D options;
  options.linear_search_window = 3.;
  options.angular_search_window = 1.;
  options.branch_and_bound_depth = 6;
  PointCloud unperturbed;
  unperturbed.push_back(Vec3f{-2.5f, 0.5f, 0.f});
  unperturbed.push_back(Vec3f{-2.25f, 0.5f, 0.f});
  unperturbed.push_back(Vec3f{0.f, 0.5f, 0.f});
  unperturbed.push_back(Vec3f{0.25f, 1.6f, 0.f});
  unperturbed.push_back(Vec3f{2.5f, 0.5f, 0.f});
  unperturbed.push_back(Vec3f{2.f, 1.8f, 0.f});
  for (int i = 0; i != 20; ++i) {
    const float px = static_cast<float>(10. * distribution(prng));
    const float py = static_cast<float>(10. * distribution(prng));
    const float pa = static_cast<float>(1.6 * distribution(prng));
    const Rigid2f perturbation{px, py, pa};
    const PointCloud cloud = TransformEmbed3D(unperturbed, perturbation);
    const float qx = static_cast<float>(2. * distribution(prng));
    const float qy = static_cast<float>(2. * distribution(prng));
    const float qa = static_cast<float>(0.5 * distribution(prng));
    const Rigid2f expected = Mul(Rigid2f{qx, qy, qa}, Inverse(perturbation));
    ProbabilityGrid grid(Limits(0.05, 5., 5., 200, 200));
    const Rigid2f origin_pose = Mul(expected, perturbation);
    RangeData rd;
    rd.origin = Vec3f{origin_pose.tx, origin_pose.ty, 0.f};
    rd.returns = TransformEmbed3D(cloud, expected);
    inserter.Insert(rd, &grid);
    grid.FinishUpdate();
    FastCorrelativeScanMatcher2D matcher(grid, options);
    Rigid2d pose;
    float score = 0.f;
    EXPECT_TRUE(matcher.MatchFullSubmap(cloud, kMinScore, &score, &pose));
    EXPECT_LT(kMinScore, score);
    EXPECT_TRUE(IsNearly(expected, pose, 0.03f));
  }
}

// real_time_correlative_scan_matcher_2d_test.cc:125-142 and :162-180
void TestRealTimeScores() {
  g_test = "RealTimeCorrelativeScanMatcherTest.*ProbabilityGrid";
  const PointCloud cloud = SevenPointCloud();
  ProbabilityGrid grid(Limits(0.05, 0.05, 0.25, 6, 6));
  const ProbabilityGridRangeDataInserter2D inserter(0.7, 0.4, true);
  RangeData rd;
  rd.origin = Vec3f{0.f, 0.f, 0.f};
  rd.returns = cloud;
  inserter.Insert(rd, &grid);
  grid.FinishUpdate();
  RealTimeCorrelativeScanMatcherOptions o;
  o.linear_search_window = 0.6;
  o.angular_search_window = 0.16;
  o.translation_delta_cost_weight = 0.;
  o.rotation_delta_cost_weight = 0.;
  const RealTimeCorrelativeScanMatcher2D matcher(o);
  const SearchParameters sp(0, 0, 0., 0.);
  const std::vector<PointCloud> scans = GenerateRotatedScans(cloud, sp);
  const std::vector<DiscreteScan2D> d = DiscretizeScans(grid.limits(), scans, 0.f, 0.f);
  {
    std::vector<Candidate2D> cands;
    cands.emplace_back(0, 0, 0, sp);
    matcher.ScoreCandidates(grid, d, sp, &cands);
    EXPECT_EQ(0, cands[0].scan_index);
    EXPECT_EQ(0, cands[0].x_index_offset);
    EXPECT_EQ(0, cands[0].y_index_offset);
    EXPECT_NEAR(0.7, cands[0].score, 1e-2);
  }
  {
    std::vector<Candidate2D> cands;
    cands.emplace_back(0, 0, 1, sp);
    matcher.ScoreCandidates(grid, d, sp, &cands);
    EXPECT_EQ(1, cands[0].y_index_offset);
    EXPECT_LT(0.7 * 3. / 7., cands[0].score);
    EXPECT_GT(0.7, cands[0].score);
  }
}

// constraint_builder_2d_test.cc:58-112 (CallsBack, FindsConstraints): protocol test
// on an empty 100x110 @1.0 grid with min_score 0 and pose_graph.lua defaults ->
// every search yields a constraint, in submission order.
void TestConstraintBuilderProtocol() {
  g_test = "ConstraintBuilder2DTest";
  ThreadPool pool(1);
  ConstraintBuilderOptions o;  // defaults = configuration_files/pose_graph.lua:17-29
  o.sampling_ratio = 1.;
  o.min_score = 0.;
  o.global_localization_min_score = 0.;
  auto wait_done = [](ConstraintBuilder2D* b) {
    std::mutex m;
    std::condition_variable cv;
    bool done = false;
    size_t got = 0;
    b->WhenDone([&](const std::vector<OracleConstraint>& r) {
      std::lock_guard<std::mutex> l(m);
      got = r.size();
      done = true;
      cv.notify_all();
    });
    std::unique_lock<std::mutex> l(m);
    cv.wait(l, [&] { return done; });
    return got;
  };
  {  // CallsBack
    ConstraintBuilder2D builder(o, &pool);
    EXPECT_EQ(0, builder.GetNumFinishedNodes());
    builder.NotifyEndOfNode();
    EXPECT_EQ(0u, wait_done(&builder));
    EXPECT_EQ(1, builder.GetNumFinishedNodes());
  }
  ConstraintBuilder2D builder(o, &pool);
  ProbabilityGrid grid(Limits(1., 2., 3., 100, 110));
  Rigid2d local_pose;
  local_pose.tx = 4.;
  local_pose.ty = 5.;
  ConstraintBuilder2D::Submap submap{&grid, local_pose};
  PointCloud cloud;
  cloud.push_back(Vec3f{0.1f, 0.2f, 0.3f});
  int expected_nodes = 0;
  for (int i = 0; i < 2; ++i) {
    EXPECT_EQ(expected_nodes, builder.GetNumFinishedNodes());
    for (int j = 0; j < 2; ++j) builder.MaybeAddConstraint(1, &submap, 0, &cloud, Rigid2d());
    builder.MaybeAddGlobalConstraint(1, &submap, 0, &cloud);
    builder.NotifyEndOfNode();
    builder.NotifyEndOfNode();
    expected_nodes += 2;
    EXPECT_EQ(3u, wait_done(&builder));
    EXPECT_EQ(expected_nodes, builder.GetNumFinishedNodes());
    builder.DeleteScanMatcher(1);
  }
}

// Not a reference test: cross-check of the deque sliding maximum against the
// closed form the CUDA path relies on (SURVEY.md 8(a) dagger): level i equals the
// uint8 window max of Q[cell] with Q = quantise(1 - |LUT[cell]|), zero outside.
void TestStackEqualsQuantisedWindowMax() {
  g_test = "StackEqualsQuantisedWindowMax";
  std::mt19937 prng(7);
  for (int trial = 0; trial < 6; ++trial) {
    const int nx = 1 + prng() % 70, ny = 1 + prng() % 70;
    std::vector<uint16_t> cells(static_cast<size_t>(nx) * ny);
    for (auto& c : cells) c = (prng() % 4 == 0) ? 0 : static_cast<uint16_t>(1 + prng() % 32767);
    ProbabilityGrid grid(Limits(0.05, 1., 1., nx, ny), cells.data());
    const int depth = 1 + trial;
    PrecomputationGridStack2D stack(grid, depth);
    const std::vector<float>& lut = ValueToCorrespondenceCostTable();
    const float mn = 1.f - kMaxCorrespondenceCost, mx = 1.f - kMinCorrespondenceCost;
    auto q = [&](uint16_t v) {
      return RoundToInt(((1.f - std::abs(lut[v])) - mn) * (255.f / (mx - mn)));
    };
    for (int lvl = 0; lvl < depth; ++lvl) {
      const int w = 1 << lvl;
      const PrecomputationGrid2D& pg = stack.Get(lvl);
      for (int y = -w + 1; y < ny; ++y)
        for (int x = -w + 1; x < nx; ++x) {
          int best = 0;
          for (int dy = 0; dy < w; ++dy)
            for (int dx = 0; dx < w; ++dx) {
              const int xx = x + dx, yy = y + dy;
              if (xx < 0 || yy < 0 || xx >= nx || yy >= ny) continue;
              best = std::max(best, q(cells[xx + yy * nx]));
            }
          Array2i c;
          c.x = x;
          c.y = y;
          EXPECT_EQ(best, pg.GetValue(c));
        }
    }
  }
}

// ---- 3D ---------------------------------------------------------------------
// precomputation_grid_3d_test.cc:33-69 (TestAgainstNaiveAlgorithm)
void TestPrecomputeGrid3DAgainstNaive() {
  HybridGrid hybrid_grid(2.f);
  std::mt19937 rng(23847);
  std::uniform_int_distribution<int> coordinate_distribution(-50, 49);
  std::uniform_real_distribution<float> value_distribution(kMinProbability, kMaxProbability);
  for (int i = 0; i < 1000; ++i) {
    const auto x = coordinate_distribution(rng);
    const auto y = coordinate_distribution(rng);
    const auto z = coordinate_distribution(rng);
    hybrid_grid.SetProbability(Array3i{x, y, z}, value_distribution(rng));
  }
  std::vector<PrecomputationGrid3D> grids;
  grids.push_back(ConvertToPrecomputationGrid(hybrid_grid));
  for (int depth = 1; depth <= 3; ++depth) {
    const int s = 1 << (depth - 1);
    grids.push_back(PrecomputeGrid(grids.back(), false, Array3i{s, s, s}));
  }
  for (int depth = 0; depth <= 3; ++depth) {
    const int width = 1 << depth;
    for (int i = 0; i < 100; ++i) {
      const auto x = coordinate_distribution(rng);
      const auto y = coordinate_distribution(rng);
      const auto z = coordinate_distribution(rng);
      float max_probability = 0.f;
      for (int dx = 0; dx < width; ++dx)
        for (int dy = 0; dy < width; ++dy)
          for (int dz = 0; dz < width; ++dz)
            max_probability =
                std::max(max_probability, hybrid_grid.GetProbability(Array3i{x + dx, y + dy, z + dz}));
      EXPECT_NEAR(max_probability, ToProbability3D(grids[depth].value(Array3i{x, y, z})), 1e-2);
    }
  }
}

// rotational_scan_matcher_test.cc:29-43
void TestRotationalOnlySameHistogramIsScoreOne() {
  const Histogram histogram{1.f, 43.f, 0.5f, 0.3123f, 23.f, 42.f, 0.f};
  const std::vector<float> scores = RotationalMatch(histogram, histogram, 0.f, {0.f, 1.f});
  EXPECT_EQ(scores.size(), size_t{2});
  EXPECT_NEAR(1.f, scores[0], 1e-6);
  EXPECT_LT(scores[1], 1.f);
}

// rotational_scan_matcher_test.cc:45-66
void TestRotationalInterpolatesAsExpected() {
  constexpr int kNumBuckets = 10;
  constexpr float kAnglePerBucket = M_PI / kNumBuckets;
  constexpr float kNoInitialRotation = 0.f;
  Histogram unit3(kNumBuckets, 0.f), unit2(kNumBuckets, 0.f);
  unit3[3] = 1.f;
  unit2[2] = 1.f;
  for (float t = 0.f; t < 1.f; t += 0.1f) {
    // 't' is the fraction of overlap and we have to divide by the norm of the
    // histogram to get the expected score.
    const float expected_score = t / std::hypot(t, 1 - t);
    // We rotate the 't'-th fraction of a bucket into the matcher's histogram.
    const std::vector<float> scores =
        RotationalMatch(unit3, unit2, kNoInitialRotation, {t * kAnglePerBucket});
    EXPECT_EQ(scores.size(), size_t{1});
    EXPECT_NEAR(expected_score, scores[0], 1e-6);
  }
}

// low_resolution_matcher (low_resolution_matcher.cc:23-35): a point in a painted cell
// scores that cell's probability, unknown space kMinProbability.
void TestLowResolutionScore() {
  HybridGrid grid(0.5f);
  grid.SetProbability(Array3i{2, 0, 0}, 0.8f);
  Rigid3f identity;
  EXPECT_NEAR(LowResolutionScore(grid, {Vec3f{1.f, 0.f, 0.f}}, identity), 0.8f, 1e-4);
  EXPECT_NEAR(LowResolutionScore(grid, {Vec3f{1.f, 0.f, 0.f}, Vec3f{-3.f, 0.f, 0.f}}, identity),
              0.5f * (0.8f + kMinProbability), 1e-4);
}

// fast_correlative_scan_matcher_3d_test.cc:35-205
class Fcsm3DFixture {
 public:
  Fcsm3DFixture() : inserter_(0.7f, 0.4f, 5), histogram_(10, 0.f) {
    for (int axis = 0; axis != 3; ++axis)
      for (const float d : {4.f, 4.5f, 5.f, 5.5f}) {
        Vec3f p{0.f, 0.f, 0.f};
        (axis == 0 ? p.x : axis == 1 ? p.y : p.z) = d;
        point_cloud_.push_back(p);
      }
    options_.branch_and_bound_depth = 6;
    options_.full_resolution_depth = 6;
    options_.min_rotational_score = 0.1;
    options_.min_low_resolution_score = 0.15;
    options_.linear_xy_search_window = 0.8;
    options_.linear_z_search_window = 0.8;
    options_.angular_search_window = 0.3;
  }
  Rigid3f GetRandomPose() {
    const float x = 0.7f * distribution_(prng_);
    const float y = 0.7f * distribution_(prng_);
    const float z = 0.7f * distribution_(prng_);
    const float theta = 0.2f * distribution_(prng_);
    Rigid3f pose;
    pose.t = Vec3f{x, y, z};
    pose.q = Quat4f{std::cos(0.5f * theta), 0.f, 0.f, std::sin(0.5f * theta)};
    return pose;
  }
  std::unique_ptr<FastCorrelativeScanMatcher3D> GetMatcher(const Rigid3f& pose) {
    hybrid_grid_.reset(new HybridGrid(0.05f));
    std::vector<Vec3f> returns;
    for (const Vec3f& p : point_cloud_) returns.push_back(Apply(pose, p));
    inserter_.Insert(pose.t, returns, hybrid_grid_.get());
    return std::unique_ptr<FastCorrelativeScanMatcher3D>(
        new FastCorrelativeScanMatcher3D(*hybrid_grid_, hybrid_grid_.get(), &histogram_, options_));
  }
  NodeData3D CreateConstantData(const std::vector<Vec3f>& low_resolution_point_cloud) const {
    NodeData3D d;
    d.high_resolution_point_cloud = point_cloud_;
    d.low_resolution_point_cloud = low_resolution_point_cloud;
    d.rotational_scan_matcher_histogram = histogram_;
    return d;
  }
  const std::vector<Vec3f>& point_cloud() const { return point_cloud_; }

 private:
  std::mt19937 prng_ = std::mt19937(42);
  std::uniform_real_distribution<float> distribution_ =
      std::uniform_real_distribution<float>(-1.f, 1.f);
  RangeDataInserter3D inserter_;
  FastCorrelativeScanMatcherOptions3D options_;
  std::vector<Vec3f> point_cloud_;
  std::unique_ptr<HybridGrid> hybrid_grid_;
  Histogram histogram_;
};

// transform::IsNearly(rigid, 0.05f) of rigid_transform_test_helpers.h: translation and
// rotation (as the angle of the relative quaternion) both within the bound.
bool PoseIsNearly(const Rigid3f& a, const Rigid3f& b, const float eps) {
  const float dt = Norm3(Vec3f{a.t.x - b.t.x, a.t.y - b.t.y, a.t.z - b.t.z});
  const Quat4f d = QuatMul(QuatInverse(a.q), b.q);
  const float angle = 2.f * std::atan2(std::sqrt(d.x * d.x + d.y * d.y + d.z * d.z), std::abs(d.w));
  return dt < eps && angle < eps;
}

// The literal matcher: ToEigen(a).isApprox(ToEigen(b), eps) on the 4x4 isometry
// matrices, i.e. ||A - B||_F <= eps * min(||A||_F, ||B||_F)
// (rigid_transform_test_helpers.h:36-46).
bool IsometryIsApprox(const Rigid3f& a, const Rigid3f& b, const double eps) {
  auto matrix = [](const Rigid3f& r, double m[16]) {
    const double w = r.q.w, x = r.q.x, y = r.q.y, z = r.q.z;
    const double rot[9] = {1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y),
                           2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x),
                           2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)};
    const double t[3] = {r.t.x, r.t.y, r.t.z};
    for (int i = 0; i < 3; ++i) {
      for (int j = 0; j < 3; ++j) m[4 * i + j] = rot[3 * i + j];
      m[4 * i + 3] = t[i];
    }
    m[12] = m[13] = m[14] = 0.;
    m[15] = 1.;
  };
  double ma[16], mb[16], d2 = 0., na = 0., nb = 0.;
  matrix(a, ma);
  matrix(b, mb);
  for (int k = 0; k < 16; ++k) {
    d2 += (ma[k] - mb[k]) * (ma[k] - mb[k]);
    na += ma[k] * ma[k];
    nb += mb[k] * mb[k];
  }
  return d2 <= eps * eps * std::min(na, nb);
}

void TestFcsm3DCorrectPoseForMatch(const int iterations) {
  Fcsm3DFixture f;
  constexpr float kMinScore = 0.1f;
  Rigid3f identity;
  for (int i = 0; i != iterations; ++i) {
    const Rigid3f expected_pose = f.GetRandomPose();
    auto matcher = f.GetMatcher(expected_pose);
    const Result3D result =
        matcher->Match(identity, identity, f.CreateConstantData(f.point_cloud()), kMinScore);
    EXPECT_TRUE(result.found);
    EXPECT_LT(kMinScore, result.score);
    EXPECT_LT(0.09f, result.rotational_score);
    EXPECT_LT(0.14f, result.low_resolution_score);
    EXPECT_TRUE(PoseIsNearly(expected_pose, result.pose_estimate, 0.05f));
    const Result3D low_resolution_result =
        matcher->Match(identity, identity, f.CreateConstantData({Vec3f{42.f, 42.f, 42.f}}), kMinScore);
    EXPECT_TRUE(!low_resolution_result.found);
  }
}

void TestFcsm3DCorrectPoseForMatchFullSubmap() {
  Fcsm3DFixture f;
  constexpr float kMinScore = 0.1f;
  const Rigid3f expected_pose = f.GetRandomPose();
  auto matcher = f.GetMatcher(expected_pose);
  const Quat4f identity;
  const Result3D result =
      matcher->MatchFullSubmap(identity, identity, f.CreateConstantData(f.point_cloud()), kMinScore);
  EXPECT_TRUE(result.found);
  EXPECT_LT(kMinScore, result.score);
  EXPECT_LT(0.09f, result.rotational_score);
  EXPECT_LT(0.14f, result.low_resolution_score);
  EXPECT_TRUE(PoseIsNearly(expected_pose, result.pose_estimate, 0.05f));
  const Result3D low_resolution_result = matcher->MatchFullSubmap(
      identity, identity, f.CreateConstantData({Vec3f{42.f, 42.f, 42.f}}), kMinScore);
  EXPECT_TRUE(!low_resolution_result.found);
}

// real_time_correlative_scan_matcher_3d_test.cc:33-120: seven points painted into a
// 0.1 m grid at the expected pose; every initial pose must come back to it within 1e-3.
void TestRealTimeCorrelativeScanMatcher3D() {
  HybridGrid grid(0.1f);
  Rigid3f expected;
  expected.t = Vec3f{-1.f, 0.f, 0.f};
  std::vector<Vec3f> cloud;
  for (const Vec3f& p : {Vec3f{-3.f, 2.f, 0.f}, Vec3f{-4.f, 2.f, 0.f}, Vec3f{-5.f, 2.f, 0.f},
                         Vec3f{-6.f, 2.f, 0.f}, Vec3f{-6.f, 3.f, 1.f}, Vec3f{-6.f, 4.f, 2.f},
                         Vec3f{-7.f, 3.f, 1.f}}) {
    cloud.push_back(p);
    grid.SetProbability(grid.GetCellIndex(Apply(expected, p)), 1.f);
  }
  RealTimeCorrelativeScanMatcherOptions o;
  o.linear_search_window = 0.3;
  o.angular_search_window = 1. * M_PI / 180.;
  o.translation_delta_cost_weight = 1e-1;
  o.rotation_delta_cost_weight = 1.;
  const RealTimeCorrelativeScanMatcher3D matcher(o);
  const double a = 0.8 / 180. * M_PI;
  struct Case {
    Vec3f t;
    int axis;  // -1: no rotation, 0: x, 1: y, 2: the (unnormalised) axis (0, 1, 1) of RotationAroundYZ
  };
  for (const Case& c : {Case{{-1.f, 0.f, 0.f}, -1}, Case{{-0.8f, 0.f, 0.f}, -1}, Case{{-1.f, 0.f, -0.2f}, -1},
                        Case{{-0.9f, -0.2f, 0.2f}, -1}, Case{{-1.f, 0.f, 0.f}, 0}, Case{{-1.f, 0.f, 0.f}, 1},
                        Case{{-1.f, 0.f, 0.f}, 2}}) {
    Rigid3f initial;
    initial.t = c.t;
    if (c.axis >= 0) {
      // Eigen::Quaterniond(AngleAxisd(a, axis)).cast<float>()
      initial.q.w = static_cast<float>(std::cos(0.5 * a));
      const float s = static_cast<float>(std::sin(0.5 * a));
      if (c.axis == 0) initial.q.x = s;
      if (c.axis >= 1) initial.q.y = s;
      if (c.axis == 2) initial.q.z = s;
    }
    Rigid3f pose;
    const float score = matcher.Match(initial, cloud, grid, &pose);
    EXPECT_GT(score, 0.f);
    EXPECT_TRUE(IsometryIsApprox(pose, expected, 1e-3));
  }
}

}  // namespace

int main(int argc, char** argv) {
  const bool quick = argc > 1 && std::strcmp(argv[1], "--quick") == 0;
  TestSearchParametersConstruction();
  TestCandidateConstruction();
  TestGenerateRotatedScans();
  TestDiscretizeScans();
  TestPrecomputationGridTiny();
  TestRealTimeScores();
  TestStackEqualsQuantisedWindowMax();
  TestConstraintBuilderProtocol();
  TestPrecomputeGrid3DAgainstNaive();
  TestRotationalOnlySameHistogramIsScoreOne();
  TestRotationalInterpolatesAsExpected();
  TestLowResolutionScore();
  TestRealTimeCorrelativeScanMatcher3D();
  TestFcsm3DCorrectPoseForMatch(quick ? 2 : 20);
  if (!quick) {
    TestFcsm3DCorrectPoseForMatchFullSubmap();
    TestPrecomputationGridCorrectValues();
    TestFastCorrectPose();
    TestFastFullSubmapMatching();
  }
  std::printf("oracle_tests: %d checks, %d failures\n", g_checks, g_failures);
  return g_failures == 0 ? 0 : 1;
}
