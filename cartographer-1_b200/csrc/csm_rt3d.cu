// RealTimeCorrelativeScanMatcher3D on the device: exhaustive scoring of candidate poses.
//
// Reference: cartographer/mapping/internal/3d/scan_matching/
//   real_time_correlative_scan_matcher_3d.cc:35-53   Match (first maximum wins)
//   :100-114                                          ScoreCandidate
//   cartographer/sensor/point_cloud.cc:56-64          TransformPointCloud
//   cartographer/mapping/3d/hybrid_grid.h:428-433     GetCellIndex
#include <algorithm>
#include <cstring>
#include <vector>

#include "csm_match3d.cuh"

namespace csm {
namespace {

constexpr int kRt3Warps = 8;
constexpr unsigned int kFull = 0xFFFFFFFFu;

struct RtJob3Dev {
  const Grid3Dev* grid;
  long long pts_off;   // float3 index
  long long cand_off;  // candidate index
  int n;
  int K;
  float res;
};

__device__ __forceinline__ float3 rotate3(float qw, float qx, float qy, float qz, float3 v) {
  float ux = __fsub_rn(__fmul_rn(qy, v.z), __fmul_rn(qz, v.y));
  float uy = __fsub_rn(__fmul_rn(qz, v.x), __fmul_rn(qx, v.z));
  float uz = __fsub_rn(__fmul_rn(qx, v.y), __fmul_rn(qy, v.x));
  ux = __fadd_rn(ux, ux);
  uy = __fadd_rn(uy, uy);
  uz = __fadd_rn(uz, uz);
  const float cx = __fsub_rn(__fmul_rn(qy, uz), __fmul_rn(qz, uy));
  const float cy = __fsub_rn(__fmul_rn(qz, ux), __fmul_rn(qx, uz));
  const float cz = __fsub_rn(__fmul_rn(qx, uy), __fmul_rn(qy, ux));
  return make_float3(__fadd_rn(__fadd_rn(v.x, __fmul_rn(qw, ux)), cx),
                     __fadd_rn(__fadd_rn(v.y, __fmul_rn(qw, uy)), cy),
                     __fadd_rn(__fadd_rn(v.z, __fmul_rn(qw, uz)), cz));
}

// One warp per candidate: lanes transform and look up 32 points at a time, the fp32
// sum then runs through the 32 values in point order (the reference's `score +=` loop).
__global__ void __launch_bounds__(kRt3Warps * 32)
k_rt3_score(const RtJob3Dev* __restrict__ jobs, const float* __restrict__ points,
            const float* __restrict__ poses, const double* __restrict__ penalty,
            unsigned long long* __restrict__ best) {
  const RtJob3Dev& j = jobs[blockIdx.y];
  const int lane = threadIdx.x & 31;
  const Grid3Dev& g = *j.grid;
  const float* __restrict__ pts = points + 3 * j.pts_off;
  unsigned long long mine = 0ull;
  for (int c = blockIdx.x * kRt3Warps + (threadIdx.x >> 5); c < j.K; c += gridDim.x * kRt3Warps) {
    const float* p7 = poses + 7 * (j.cand_off + c);
    const float tx = p7[0], ty = p7[1], tz = p7[2], qw = p7[3], qx = p7[4], qy = p7[5], qz = p7[6];
    float score = 0.f;
    for (int base = 0; base < j.n; base += 32) {
      const int i = base + lane;
      float p = 0.f;
      if (i < j.n) {
        const float3 r = rotate3(qw, qx, qy, qz, make_float3(pts[3 * i], pts[3 * i + 1], pts[3 * i + 2]));
        const int x = static_cast<int>(lroundf(__fdiv_rn(__fadd_rn(r.x, tx), j.res))) - g.lo[0];
        const int y = static_cast<int>(lroundf(__fdiv_rn(__fadd_rn(r.y, ty), j.res))) - g.lo[1];
        const int z = static_cast<int>(lroundf(__fdiv_rn(__fadd_rn(r.z, tz), j.res))) - g.lo[2];
        const bool in = static_cast<unsigned>(x) < static_cast<unsigned>(g.dim[0]) &&
                        static_cast<unsigned>(y) < static_cast<unsigned>(g.dim[1]) &&
                        static_cast<unsigned>(z) < static_cast<unsigned>(g.dim[2]);
        p = in ? g.data[(static_cast<size_t>(z) * g.dim[1] + y) * g.dim[0] + x] : g.unknown;
      }
      const int count = min(32, j.n - base);
      for (int k = 0; k < count; ++k) score = __fadd_rn(score, __shfl_sync(kFull, p, k));
    }
    score = __fdiv_rn(score, static_cast<float>(j.n));
    // score *= std::exp(...): float times double, rounded back to float
    score = static_cast<float>(__dmul_rn(static_cast<double>(score), penalty[j.cand_off + c]));
    // first maximum in generation order: larger score, then smaller index
    const unsigned long long key = (static_cast<unsigned long long>(float_to_ordered(score)) << 32) |
                                   (0xFFFFFFFFull - static_cast<unsigned long long>(c));
    mine = max(mine, key);
  }
  if (lane == 0 && mine) atomicMax(&best[blockIdx.y], mine);
}

}  // namespace
}  // namespace csm

using namespace csm;

extern "C" int csm_rt_match3d_batch(csm_ctx* ctx, const csm_rt_job3d* jobs, int num_jobs,
                                    csm_rt_result3d* results) {
  CSM_REQUIRE(ctx != nullptr && jobs != nullptr && results != nullptr, "NULL argument");
  CSM_REQUIRE(num_jobs >= 1 && num_jobs <= 65535, "1 <= num_jobs <= 65535");
  CSM_CUDA(cudaSetDevice(ctx->device));
  std::vector<RtJob3Dev> host(num_jobs);
  long long total_pts = 0, total_cand = 0;
  int max_k = 0;
  for (int i = 0; i < num_jobs; ++i) {
    const csm_rt_job3d& in = jobs[i];
    CSM_REQUIRE(in.grid != nullptr && in.grid->device == ctx->device, "grid missing or on another device");
    // the reference divides by the cloud size and CHECKs score > 0 (:107,:113)
    CSM_REQUIRE(in.num_points >= 1 && in.points_xyz != nullptr, "empty point cloud");
    CSM_REQUIRE(in.num_candidates >= 1 && in.candidate_poses != nullptr && in.penalty != nullptr,
                "no candidates");
    CSM_REQUIRE(in.resolution > 0.f, "resolution");
    host[i] = RtJob3Dev{in.grid->dev, total_pts, total_cand, in.num_points, in.num_candidates, in.resolution};
    total_pts += in.num_points;
    total_cand += in.num_candidates;
    max_k = std::max(max_k, in.num_candidates);
  }
  auto al = [](size_t v) { return (v + 255) & ~static_cast<size_t>(255); };
  const size_t o_jobs = 0, o_pts = o_jobs + al(sizeof(RtJob3Dev) * num_jobs);
  const size_t o_pose = o_pts + al(sizeof(float) * 3 * total_pts);
  const size_t o_pen = o_pose + al(sizeof(float) * 7 * total_cand);
  const size_t o_best = o_pen + al(sizeof(double) * total_cand);
  const size_t total = o_best + al(sizeof(unsigned long long) * num_jobs);
  void *pin = nullptr, *dev = nullptr;
  size_t pin_cap = 0, dev_cap = 0;
  if (ctx->pool.take(true, total, &pin, &pin_cap) != CSM_OK) return CSM_ERR_CUDA;
  if (ctx->pool.take(false, total, &dev, &dev_cap) != CSM_OK) {
    ctx->pool.give(true, pin, pin_cap);
    return CSM_ERR_CUDA;
  }
  char* hp = static_cast<char*>(pin);
  char* dp = static_cast<char*>(dev);
  memcpy(hp + o_jobs, host.data(), sizeof(RtJob3Dev) * num_jobs);
  for (int i = 0; i < num_jobs; ++i) {
    const csm_rt_job3d& in = jobs[i];
    memcpy(hp + o_pts + sizeof(float) * 3 * host[i].pts_off, in.points_xyz, sizeof(float) * 3 * in.num_points);
    memcpy(hp + o_pose + sizeof(float) * 7 * host[i].cand_off, in.candidate_poses,
           sizeof(float) * 7 * in.num_candidates);
    memcpy(hp + o_pen + sizeof(double) * host[i].cand_off, in.penalty, sizeof(double) * in.num_candidates);
  }
  memset(hp + o_best, 0, sizeof(unsigned long long) * num_jobs);
  cudaStream_t st = ctx->stream;
  int rc = CSM_OK;
  auto finish = [&](int code) {
    cudaStreamSynchronize(st);
    ctx->pool.give(true, pin, pin_cap);
    ctx->pool.give(false, dev, dev_cap);
    return code;
  };
  if (cudaMemcpyAsync(dp, hp, total, cudaMemcpyHostToDevice, st) != cudaSuccess) {
    set_error("csm_rt_match3d_batch upload: %s", cudaGetErrorString(cudaGetLastError()));
    return finish(CSM_ERR_CUDA);
  }
  const int blocks = std::max(1, std::min((max_k + kRt3Warps - 1) / kRt3Warps, ctx->num_sms * 8));
  ++ctx->launches, k_rt3_score<<<dim3(blocks, num_jobs), kRt3Warps * 32, 0, st>>>(
      reinterpret_cast<const RtJob3Dev*>(dp + o_jobs), reinterpret_cast<const float*>(dp + o_pts),
      reinterpret_cast<const float*>(dp + o_pose), reinterpret_cast<const double*>(dp + o_pen),
      reinterpret_cast<unsigned long long*>(dp + o_best));
  if (cudaMemcpyAsync(hp + o_best, dp + o_best, sizeof(unsigned long long) * num_jobs, cudaMemcpyDeviceToHost,
                      st) != cudaSuccess ||
      cudaStreamSynchronize(st) != cudaSuccess || cudaGetLastError() != cudaSuccess) {
    set_error("csm_rt_match3d_batch: %s", cudaGetErrorString(cudaGetLastError()));
    return finish(CSM_ERR_CUDA);
  }
  const unsigned long long* best = reinterpret_cast<const unsigned long long*>(hp + o_best);
  for (int i = 0; i < num_jobs; ++i) {
    results[i].best_index = static_cast<int32_t>(0xFFFFFFFFull - (best[i] & 0xFFFFFFFFull));
    results[i].score = ordered_to_float(static_cast<unsigned int>(best[i] >> 32));
  }
  return finish(rc);
}
