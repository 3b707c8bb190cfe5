// Device entry points behind the PUBLIC helpers of the replaced headers that the reference's own
// unit tests call directly:
//   GenerateRotatedScans / DiscretizeScans as free functions
//     (/root/reference/cartographer/mapping/internal/2d/scan_matching/
//      correlative_scan_matcher_2d.h:63-72, .cc:93-127)
//   RealTimeCorrelativeScanMatcher2D::ScoreCandidates on a caller-supplied candidate list
//     (real_time_correlative_scan_matcher_2d.h:75-79, .cc:151-176)
// The matchers themselves do not use them: their kernels fuse rotation, discretisation and
// scoring (csm_match2d.cu, csm_rt2d.cu).  Same arithmetic, op for op.
#include "csm_match2d.cuh"

namespace csm {

// One thread per (scan, point): Eigen's quaternion * vector for the scan's yaw (CSM2D.cc:104-106).
__global__ void __launch_bounds__(256)
k_rotate_scans(const float* __restrict__ points, int n, const float* __restrict__ quats, int num_scans,
               float* __restrict__ out) {
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= static_cast<long long>(n) * num_scans) return;
  const int scan = static_cast<int>(i / n), p = static_cast<int>(i - static_cast<long long>(scan) * n);
  float vx = points[3 * p], vy = points[3 * p + 1];
  yaw_rotate(quats[2 * scan], quats[2 * scan + 1], vx, vy);
  out[3 * i] = vx;
  out[3 * i + 1] = vy;
  out[3 * i + 2] = points[3 * p + 2];  // a rotation about z leaves z alone
}

// One thread per point: Affine2f(Translation2f(tx, ty)) * p.head<2>(), then
// MapLimits::GetCellIndex in fp64 (CSM2D.cc:117-124, map_limits.h:69-76).
__global__ void __launch_bounds__(256)
k_discretize_points(const float* __restrict__ points, long long count, float tx, float ty, double max_x,
                    double max_y, double res, int2* __restrict__ out) {
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= count) return;
  const float wx = __fadd_rn(tx, points[3 * i]), wy = __fadd_rn(ty, points[3 * i + 1]);
  int cx, cy;
  cell_index(max_x, max_y, res, wx, wy, cx, cy);
  out[i] = make_int2(cx, cy);
}

struct RtScoreArgs {
  const uint16_t* cells;
  const float* table;
  const uint16_t* weight_cells;  // TSDF only
  const float* weight_table;
  int num_x, num_y, n;
  float oob, max_cc, oob_tsd;
};

// One thread per candidate (scan, x, y): ComputeCandidateScore (RT2D.cc:38-75) -- the fp32 sum
// runs over the points in order, exactly like the CPU loop.
__global__ void __launch_bounds__(128)
k_rt_score_list(const RtScoreArgs a, const int2* __restrict__ xy, const int* __restrict__ cand, int num_cand,
                float* __restrict__ out) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= num_cand) return;
  const int scan = cand[3 * c], x = cand[3 * c + 1], y = cand[3 * c + 2];
  const int2* pts = xy + static_cast<long long>(scan) * a.n;
  float sum = 0.f, wsum = 0.f;
  for (int p = 0; p < a.n; ++p) {
    const int px = pts[p].x + x, py = pts[p].y + y;
    const bool in = static_cast<unsigned>(px) < static_cast<unsigned>(a.num_x) &&
                    static_cast<unsigned>(py) < static_cast<unsigned>(a.num_y);
    const size_t flat = in ? static_cast<size_t>(py) * a.num_x + px : 0;
    if (a.weight_cells == nullptr) {
      const float prob = in ? __fsub_rn(1.f, __ldg(a.table + __ldg(a.cells + flat))) : a.oob;
      sum = __fadd_rn(sum, prob);
    } else {
      const float tsd = in ? __ldg(a.table + __ldg(a.cells + flat)) : a.oob_tsd;
      const float w = in ? __ldg(a.weight_table + __ldg(a.weight_cells + flat)) : 0.f;
      sum = __fadd_rn(sum, __fmul_rn(__fdiv_rn(__fsub_rn(a.max_cc, fabsf(tsd)), a.max_cc), w));
      wsum = __fadd_rn(wsum, w);
    }
  }
  if (a.weight_cells == nullptr) out[c] = __fdiv_rn(sum, static_cast<float>(a.n));
  else out[c] = wsum == 0.f ? 0.f : __fdiv_rn(sum, wsum);
}

}  // namespace csm

using namespace csm;

extern "C" {

int csm_rotate_scans2d(csm_ctx* ctx, const float* points_xyz, int num_points, const float* scan_quat,
                       int num_scans, float* out_xyz) {
  CSM_REQUIRE(ctx != nullptr && scan_quat != nullptr && out_xyz != nullptr, "NULL argument");
  CSM_REQUIRE(num_points >= 0 && num_scans >= 1 && (num_points == 0 || points_xyz != nullptr), "bad sizes");
  if (num_points == 0) return CSM_OK;
  CSM_CUDA(cudaSetDevice(ctx->device));
  const size_t b_pts = (sizeof(float) * 3 * num_points + 255) & ~static_cast<size_t>(255);
  const size_t b_quat = (sizeof(float) * 2 * num_scans + 255) & ~static_cast<size_t>(255);
  const size_t b_out = sizeof(float) * 3 * static_cast<size_t>(num_points) * num_scans;
  const int rc = ctx->tmp1.reserve(b_pts + b_quat + b_out);
  if (rc != CSM_OK) return rc;
  char* base = ctx->tmp1.as<char>();
  cudaStream_t st = ctx->stream;
  CSM_CUDA(cudaMemcpyAsync(base, points_xyz, sizeof(float) * 3 * num_points, cudaMemcpyHostToDevice, st));
  CSM_CUDA(cudaMemcpyAsync(base + b_pts, scan_quat, sizeof(float) * 2 * num_scans, cudaMemcpyHostToDevice, st));
  const long long total = static_cast<long long>(num_points) * num_scans;
  ++ctx->launches, k_rotate_scans<<<static_cast<unsigned int>((total + 255) / 256), 256, 0, st>>>(
      reinterpret_cast<const float*>(base), num_points, reinterpret_cast<const float*>(base + b_pts), num_scans,
      reinterpret_cast<float*>(base + b_pts + b_quat));
  CSM_CUDA(cudaGetLastError());
  CSM_CUDA(cudaMemcpyAsync(out_xyz, base + b_pts + b_quat, b_out, cudaMemcpyDeviceToHost, st));
  CSM_CUDA(cudaStreamSynchronize(st));
  return CSM_OK;
}

int csm_discretize_points2d(csm_ctx* ctx, const float* points_xyz, long long num_points, float tx, float ty,
                            double max_x, double max_y, double resolution, int32_t* out_xy) {
  CSM_REQUIRE(ctx != nullptr && out_xy != nullptr, "NULL argument");
  CSM_REQUIRE(num_points >= 0 && (num_points == 0 || points_xyz != nullptr) && resolution > 0., "bad sizes");
  if (num_points == 0) return CSM_OK;
  CSM_CUDA(cudaSetDevice(ctx->device));
  const size_t b_pts = (sizeof(float) * 3 * num_points + 255) & ~static_cast<size_t>(255);
  const int rc = ctx->tmp1.reserve(b_pts + sizeof(int2) * num_points);
  if (rc != CSM_OK) return rc;
  char* base = ctx->tmp1.as<char>();
  cudaStream_t st = ctx->stream;
  CSM_CUDA(cudaMemcpyAsync(base, points_xyz, sizeof(float) * 3 * num_points, cudaMemcpyHostToDevice, st));
  ++ctx->launches, k_discretize_points<<<static_cast<unsigned int>((num_points + 255) / 256), 256, 0, st>>>(
      reinterpret_cast<const float*>(base), num_points, tx, ty, max_x, max_y, resolution,
      reinterpret_cast<int2*>(base + b_pts));
  CSM_CUDA(cudaGetLastError());
  CSM_CUDA(cudaMemcpyAsync(out_xy, base + b_pts, sizeof(int2) * num_points, cudaMemcpyDeviceToHost, st));
  CSM_CUDA(cudaStreamSynchronize(st));
  return CSM_OK;
}

int csm_rt_score_candidates2d(csm_ctx* ctx, const csm_rt_job2d* grid, const int32_t* discrete_xy,
                              int num_scans, const int32_t* candidates, int num_candidates, float* out_scores) {
  CSM_REQUIRE(ctx != nullptr && grid != nullptr && discrete_xy != nullptr && candidates != nullptr &&
                  out_scores != nullptr,
              "NULL argument");
  CSM_REQUIRE(grid->cells && grid->cost_table && grid->num_x >= 1 && grid->num_y >= 1 && grid->num_points >= 1 &&
                  num_scans >= 1 && num_candidates >= 1,
              "bad sizes");
  CSM_REQUIRE(grid->weight_cells == nullptr || (grid->weight_table != nullptr && grid->max_correspondence_cost > 0.f),
              "a TSDF grid needs a weight table and a positive truncation distance");
  for (int c = 0; c < num_candidates; ++c)
    CSM_REQUIRE(candidates[3 * c] >= 0 && candidates[3 * c] < num_scans, "candidate scan index out of range");
  CSM_CUDA(cudaSetDevice(ctx->device));
  auto up = [](size_t v) { return (v + 255) & ~static_cast<size_t>(255); };
  const size_t cells_b = sizeof(uint16_t) * static_cast<size_t>(grid->num_x) * grid->num_y;
  const bool tsdf = grid->weight_cells != nullptr;
  const size_t o_cells = 0, o_table = o_cells + up(cells_b), o_wcells = o_table + up(sizeof(float) * 65536);
  const size_t o_wtable = o_wcells + (tsdf ? up(cells_b) : 0), o_xy = o_wtable + (tsdf ? up(sizeof(float) * 65536) : 0);
  const size_t xy_b = sizeof(int2) * static_cast<size_t>(num_scans) * grid->num_points;
  const size_t o_cand = o_xy + up(xy_b), o_out = o_cand + up(sizeof(int) * 3 * num_candidates);
  const int rc = ctx->tmp1.reserve(o_out + sizeof(float) * num_candidates);
  if (rc != CSM_OK) return rc;
  char* base = ctx->tmp1.as<char>();
  cudaStream_t st = ctx->stream;
  CSM_CUDA(cudaMemcpyAsync(base + o_cells, grid->cells, cells_b, cudaMemcpyHostToDevice, st));
  CSM_CUDA(cudaMemcpyAsync(base + o_table, grid->cost_table, sizeof(float) * 65536, cudaMemcpyHostToDevice, st));
  if (tsdf) {
    CSM_CUDA(cudaMemcpyAsync(base + o_wcells, grid->weight_cells, cells_b, cudaMemcpyHostToDevice, st));
    CSM_CUDA(cudaMemcpyAsync(base + o_wtable, grid->weight_table, sizeof(float) * 65536, cudaMemcpyHostToDevice, st));
  }
  CSM_CUDA(cudaMemcpyAsync(base + o_xy, discrete_xy, xy_b, cudaMemcpyHostToDevice, st));
  CSM_CUDA(cudaMemcpyAsync(base + o_cand, candidates, sizeof(int) * 3 * num_candidates, cudaMemcpyHostToDevice, st));
  RtScoreArgs a;
  a.cells = reinterpret_cast<const uint16_t*>(base + o_cells);
  a.table = reinterpret_cast<const float*>(base + o_table);
  a.weight_cells = tsdf ? reinterpret_cast<const uint16_t*>(base + o_wcells) : nullptr;
  a.weight_table = tsdf ? reinterpret_cast<const float*>(base + o_wtable) : nullptr;
  a.num_x = grid->num_x;
  a.num_y = grid->num_y;
  a.n = grid->num_points;
  a.oob = grid->out_of_bounds_probability;
  a.max_cc = grid->max_correspondence_cost;
  a.oob_tsd = grid->out_of_bounds_tsd;
  ++ctx->launches, k_rt_score_list<<<(num_candidates + 127) / 128, 128, 0, st>>>(
      a, reinterpret_cast<const int2*>(base + o_xy), reinterpret_cast<const int*>(base + o_cand), num_candidates,
      reinterpret_cast<float*>(base + o_out));
  CSM_CUDA(cudaGetLastError());
  CSM_CUDA(cudaMemcpyAsync(out_scores, base + o_out, sizeof(float) * num_candidates, cudaMemcpyDeviceToHost, st));
  CSM_CUDA(cudaStreamSynchronize(st));
  return CSM_OK;
}

}  // extern "C"
