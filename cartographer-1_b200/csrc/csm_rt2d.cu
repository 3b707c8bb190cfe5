// RealTimeCorrelativeScanMatcher2D::Match on the device: exhaustive search on the
// live uint16 grid, no precomputation.
//
// Replaces /root/reference/cartographer/mapping/internal/2d/scan_matching/
// real_time_correlative_scan_matcher_2d.cc:61-75 (ComputeCandidateScore on a
// ProbabilityGrid), :38-59 (the same on a TSDF2D: weighted mean of the normalised
// truncated distance, two cell arrays and two tables), :83-115 (GenerateExhaustiveSearchCandidates), :117-176 (Match,
// ScoreCandidates).  One block per (job, scan): the scan is rotated and discretised
// into shared memory once, then every thread scores whole candidates, summing the
// probabilities in point order with sequential fp32 adds exactly like the CPU loop
// (the sum is order-sensitive, so it is not parallelised across points).  The
// motion penalty exp(-(hypot*w_t + |theta|*w_r)^2) is evaluated by the host with
// glibc in fp64 and applied here as float(double(score) * penalty).  The winner is
// std::max_element's: the first maximum in generation order scan -> x -> y.
#include <map>

#include "csm_match2d.cuh"

namespace csm {

struct RtJobDev {
  const uint16_t* cells;
  const float* cost_table;
  const uint16_t* weight_cells;  // TSDF only (nullptr selects the probability score)
  const float* weight_table;
  // The grid decoded once per batch (k_rt_decode): probability 1 - cost_table[cell] for a
  // ProbabilityGrid, cost_table[cell] (the TSD) and weight_table[weight cell] for a TSDF2D.
  // Same fp32 values as the two dependent lookups per point they replace.
  const float* dec;
  const float* wdec;
  const double* penalty;
  int compact_nap;  // >= 0: `penalty` is the compact table indexed by (|scan - nap|, |x|, |y|)
  int num_x, num_y, n, S, lin;
  long long pts_off, quat_off;
  float q0w, q0s, tx, ty;
  double max_x, max_y, res;
  float oob;
  float max_cc, oob_tsd;  // TSDF: truncation distance, TSD outside the grid
};

// One warp per (job, scan); a block carries kRtWarps scans of one job.  The scan is
// rotated and discretised into the warp's slice of shared memory, then each lane
// scores candidates (lane, lane + 32, ...).  The fp32 sum runs in point order exactly
// like the CPU loop; the loads of 8 points are issued together ahead of their adds so
// that the two dependent reads (cell, then table) are pipelined, not serialised.
__global__ void __launch_bounds__(256)
k_rt_decode(const uint16_t* __restrict__ cells, const float* __restrict__ table, float* __restrict__ out,
            size_t n, int one_minus) {
  const size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float t = __ldg(table + __ldg(cells + i));
  out[i] = one_minus ? __fsub_rn(1.f, t) : t;
}

constexpr int kRtWarps = 8;
__global__ void __launch_bounds__(kRtWarps * 32)
k_rt_score(const RtJobDev* __restrict__ jobs, const float* __restrict__ points,
           const float* __restrict__ quats, unsigned long long* __restrict__ best, int max_n,
           int2* __restrict__ spill) {
  extern __shared__ int2 cell_all[];
  const RtJobDev& j = jobs[blockIdx.y];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int scan = blockIdx.x * kRtWarps + warp;
  if (scan >= j.S) return;
  // The discrete scan of a warp lives in shared memory; clouds too large for that (more than
  // 3200 points) keep it in a global scratch slice per (job, scan) instead -- the reference has
  // no limit on the cloud's size.
  int2* cell = spill != nullptr ? spill + static_cast<size_t>(j.quat_off + scan) * max_n
                                : cell_all + static_cast<size_t>(warp) * max_n;
  const float qw = quats[2 * (j.quat_off + scan)], qs = quats[2 * (j.quat_off + scan) + 1];
  for (int p = lane; p < j.n; p += 32) {
    float vx = points[3 * (j.pts_off + p)], vy = points[3 * (j.pts_off + p) + 1];
    yaw_rotate(j.q0w, j.q0s, vx, vy);  // RT2D.cc:124-127
    yaw_rotate(qw, qs, vx, vy);
    const float wx = __fadd_rn(j.tx, vx), wy = __fadd_rn(j.ty, vy);
    int cx, cy;
    cell_index(j.max_x, j.max_y, j.res, wx, wy, cx, cy);
    cell[p] = make_int2(cx, cy);
  }
  __syncwarp();
  const int side = 2 * j.lin + 1;
  const int count = side * side;
  const float n_float = static_cast<float>(j.n);
  unsigned long long local = 0ull;
  for (int c = lane; c < count; c += 32) {
    const int x = c / side - j.lin, y = c % side - j.lin;  // x outer, y inner (RT2D.cc:100-112)
    float score;
    if (j.weight_cells == nullptr) {
      float sum = 0.f;
      for (int p0 = 0; p0 < j.n; p0 += 8) {
        float prob[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          prob[k] = j.oob;
          if (p0 + k < j.n) {
            const int px = cell[p0 + k].x + x, py = cell[p0 + k].y + y;
            if (static_cast<unsigned>(px) < static_cast<unsigned>(j.num_x) &&
                static_cast<unsigned>(py) < static_cast<unsigned>(j.num_y)) {
              prob[k] = __ldg(j.dec + static_cast<size_t>(py) * j.num_x + px);
            }
          }
        }
#pragma unroll
        for (int k = 0; k < 8; ++k)
          if (p0 + k < j.n) sum = __fadd_rn(sum, prob[k]);
      }
      score = __fdiv_rn(sum, n_float);
    } else {
      // TSDF2D (RT2D.cc:38-59): per point, in order, score += ((max_cc - |tsd|) / max_cc) * w
      // and summed_weight += w, all fp32 without contraction.
      float sum = 0.f, wsum = 0.f;
      for (int p0 = 0; p0 < j.n; p0 += 8) {
        float tsd[8], wt[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          tsd[k] = j.oob_tsd;
          wt[k] = 0.f;
          if (p0 + k < j.n) {
            const int px = cell[p0 + k].x + x, py = cell[p0 + k].y + y;
            if (static_cast<unsigned>(px) < static_cast<unsigned>(j.num_x) &&
                static_cast<unsigned>(py) < static_cast<unsigned>(j.num_y)) {
              const size_t flat = static_cast<size_t>(py) * j.num_x + px;
              tsd[k] = __ldg(j.dec + flat);
              wt[k] = __ldg(j.wdec + flat);
            }
          }
        }
#pragma unroll
        for (int k = 0; k < 8; ++k)
          if (p0 + k < j.n) {
            const float norm = __fdiv_rn(__fsub_rn(j.max_cc, fabsf(tsd[k])), j.max_cc);
            sum = __fadd_rn(sum, __fmul_rn(norm, wt[k]));
            wsum = __fadd_rn(wsum, wt[k]);
          }
      }
      score = wsum == 0.f ? 0.f : __fdiv_rn(sum, wsum);
    }
    const unsigned int cand = static_cast<unsigned int>(scan) * count + c;
    const double factor = j.compact_nap >= 0
                              ? j.penalty[(abs(scan - j.compact_nap) * (j.lin + 1) + abs(x)) * (j.lin + 1) + abs(y)]
                              : j.penalty[cand];
    score = __double2float_rn(__dmul_rn(static_cast<double>(score), factor));
    const unsigned long long key =
        (static_cast<unsigned long long>(float_to_ordered(score)) << 32) | (0xFFFFFFFFu - cand);
    local = max(local, key);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) local = max(local, __shfl_xor_sync(0xFFFFFFFFu, local, o));
  if (lane == 0 && local != 0ull) atomicMax(&best[blockIdx.y], local);
}

}  // namespace csm

using namespace csm;

extern "C" int csm_rt_match2d_batch(csm_ctx* ctx, const csm_rt_job2d* jobs, int num_jobs,
                                    csm_result2d* results) {
  CSM_REQUIRE(ctx != nullptr && jobs != nullptr && results != nullptr, "NULL argument");
  CSM_REQUIRE(num_jobs >= 1 && num_jobs <= 65535, "num_jobs must be in [1, 65535]");
  CSM_CUDA(cudaSetDevice(ctx->device));
  auto up = [](size_t v) { return (v + 255) & ~static_cast<size_t>(255); };
  // Layout of one upload blob: [jobs | points | quats | penalties | tables | grids | best]
  std::vector<RtJobDev> dev(num_jobs);
  std::map<const void*, size_t> grid_at, table_at;
  size_t n_pts = 0, n_scans = 0, n_pen = 0, grid_bytes = 0, table_bytes = 0;
  int max_scans = 0, max_n = 0;
  for (int i = 0; i < num_jobs; ++i) {
    const csm_rt_job2d& in = jobs[i];
    CSM_REQUIRE(in.cells && in.cost_table && in.points_xyz && in.scan_quat && (in.penalty || in.penalty_table),
                "rt job has a NULL array");
    CSM_REQUIRE(in.penalty_table == nullptr ||
                    (in.num_angular_perturbations >= 0 && in.num_scans == 2 * in.num_angular_perturbations + 1),
                "compact penalty table needs num_scans == 2 * num_angular_perturbations + 1");
    CSM_REQUIRE(in.num_points >= 1 && in.num_scans >= 1 && in.num_x >= 1 && in.num_y >= 1 &&
                    in.linear_bound >= 0 && in.linear_bound <= 1024 && in.resolution > 0.,
                "rt job sizes out of range");
    const size_t side = 2 * static_cast<size_t>(in.linear_bound) + 1;
    CSM_REQUIRE(side * side * in.num_scans < 0xFFFFFFFFull, "too many rt candidates");
    if (!grid_at.count(in.cells)) {
      grid_at[in.cells] = grid_bytes;
      grid_bytes += up(sizeof(uint16_t) * static_cast<size_t>(in.num_x) * in.num_y);
    }
    if (!table_at.count(in.cost_table)) {
      table_at[in.cost_table] = table_bytes;
      table_bytes += up(sizeof(float) * 65536);
    }
    if (in.weight_cells != nullptr) {
      CSM_REQUIRE(in.weight_table != nullptr && in.max_correspondence_cost > 0.f,
                  "TSDF rt job needs a weight table and a positive truncation distance");
      if (!grid_at.count(in.weight_cells)) {
        grid_at[in.weight_cells] = grid_bytes;
        grid_bytes += up(sizeof(uint16_t) * static_cast<size_t>(in.num_x) * in.num_y);
      }
      if (!table_at.count(in.weight_table)) {
        table_at[in.weight_table] = table_bytes;
        table_bytes += up(sizeof(float) * 65536);
      }
    }
    RtJobDev& j = dev[i];
    j.num_x = in.num_x;
    j.num_y = in.num_y;
    j.n = in.num_points;
    j.S = in.num_scans;
    j.lin = in.linear_bound;
    j.pts_off = static_cast<long long>(n_pts);
    j.quat_off = static_cast<long long>(n_scans);
    j.q0w = in.init_quat_w;
    j.q0s = in.init_quat_s;
    j.tx = in.init_tx;
    j.ty = in.init_ty;
    j.max_x = in.max_x;
    j.max_y = in.max_y;
    j.res = in.resolution;
    j.oob = in.out_of_bounds_probability;
    j.max_cc = in.max_correspondence_cost;
    j.oob_tsd = in.out_of_bounds_tsd;
    n_pts += in.num_points;
    n_scans += in.num_scans;
    n_pen += in.penalty_table ? static_cast<size_t>(in.num_angular_perturbations + 1) * (in.linear_bound + 1) *
                                    (in.linear_bound + 1)
                              : side * side * in.num_scans;
    max_scans = std::max(max_scans, in.num_scans);
    max_n = std::max(max_n, in.num_points);
  }
  size_t smem = sizeof(int2) * static_cast<size_t>(max_n) * kRtWarps;
  const bool use_spill = smem > 200 * 1024;
  const size_t spill_bytes = use_spill ? up(sizeof(int2) * n_scans * static_cast<size_t>(max_n)) : 0;
  if (use_spill) smem = 0;
  const size_t o_jobs = 0;
  const size_t o_pts = o_jobs + up(sizeof(RtJobDev) * num_jobs);
  const size_t o_quat = o_pts + up(sizeof(float) * 3 * n_pts);
  const size_t o_pen = o_quat + up(sizeof(float) * 2 * n_scans);
  const size_t o_tab = o_pen + up(sizeof(double) * n_pen);
  const size_t o_grid = o_tab + table_bytes;
  const size_t o_best = o_grid + grid_bytes;
  const size_t total = o_best + up(sizeof(unsigned long long) * num_jobs);
  // Device only, behind the uploaded blob: one decoded float image per distinct (cells, table).
  struct Decoded {
    const void* cells;
    const void* table;
    size_t cells_at, table_at, out_at, n;
    int one_minus;
  };
  std::vector<Decoded> decoded;
  size_t dec_bytes = 0;
  auto decoded_at = [&](const void* cells, const void* table, size_t n, int one_minus) -> size_t {
    for (const Decoded& d : decoded)
      if (d.cells == cells && d.table == table && d.one_minus == one_minus) return d.out_at;
    Decoded d{cells, table, grid_at[cells], table_at[table], dec_bytes, n, one_minus};
    decoded.push_back(d);
    dec_bytes += up(sizeof(float) * n);
    return d.out_at;
  };
  std::vector<size_t> dec_of(num_jobs), wdec_of(num_jobs, 0);
  for (int i = 0; i < num_jobs; ++i) {
    const csm_rt_job2d& in = jobs[i];
    const size_t n = static_cast<size_t>(in.num_x) * in.num_y;
    dec_of[i] = decoded_at(in.cells, in.cost_table, n, in.weight_cells ? 0 : 1);
    if (in.weight_cells) wdec_of[i] = decoded_at(in.weight_cells, in.weight_table, n, 0);
  }
  const size_t o_dec = total;
  int rc = ctx->tmp2.reserve(total + dec_bytes + spill_bytes);
  if (rc == CSM_OK) rc = ctx->pin0.reserve(total);
  if (rc != CSM_OK) return rc;
  char* pin = ctx->pin0.as<char>();
  char* base = ctx->tmp2.as<char>();
  size_t pen_at = 0;
  for (int i = 0; i < num_jobs; ++i) {
    const csm_rt_job2d& in = jobs[i];
    RtJobDev& j = dev[i];
    const size_t side = 2 * static_cast<size_t>(in.linear_bound) + 1;
    const size_t pen_n = in.penalty_table ? static_cast<size_t>(in.num_angular_perturbations + 1) *
                                                (in.linear_bound + 1) * (in.linear_bound + 1)
                                          : side * side * in.num_scans;
    memcpy(pin + o_pts + sizeof(float) * 3 * j.pts_off, in.points_xyz, sizeof(float) * 3 * j.n);
    memcpy(pin + o_quat + sizeof(float) * 2 * j.quat_off, in.scan_quat, sizeof(float) * 2 * j.S);
    memcpy(pin + o_pen + sizeof(double) * pen_at, in.penalty_table ? in.penalty_table : in.penalty,
           sizeof(double) * pen_n);
    j.compact_nap = in.penalty_table ? in.num_angular_perturbations : -1;
    j.penalty = reinterpret_cast<const double*>(base + o_pen) + pen_at;
    j.cells = reinterpret_cast<const uint16_t*>(base + o_grid + grid_at[in.cells]);
    j.cost_table = reinterpret_cast<const float*>(base + o_tab + table_at[in.cost_table]);
    j.weight_cells = in.weight_cells
                         ? reinterpret_cast<const uint16_t*>(base + o_grid + grid_at[in.weight_cells])
                         : nullptr;
    j.weight_table = in.weight_cells
                         ? reinterpret_cast<const float*>(base + o_tab + table_at[in.weight_table])
                         : nullptr;
    j.dec = reinterpret_cast<const float*>(base + o_dec + dec_of[i]);
    j.wdec = in.weight_cells ? reinterpret_cast<const float*>(base + o_dec + wdec_of[i]) : nullptr;
    pen_at += pen_n;
  }
  for (const auto& kv : grid_at) {
    for (int i = 0; i < num_jobs; ++i)
      if (jobs[i].cells == kv.first || jobs[i].weight_cells == kv.first) {
        memcpy(pin + o_grid + kv.second, kv.first,
               sizeof(uint16_t) * static_cast<size_t>(jobs[i].num_x) * jobs[i].num_y);
        break;
      }
  }
  for (const auto& kv : table_at) memcpy(pin + o_tab + kv.second, kv.first, sizeof(float) * 65536);
  memcpy(pin + o_jobs, dev.data(), sizeof(RtJobDev) * num_jobs);
  memset(pin + o_best, 0, sizeof(unsigned long long) * num_jobs);
  cudaStream_t st = ctx->stream;
  CSM_CUDA(cudaMemcpyAsync(base, pin, total, cudaMemcpyHostToDevice, st));
  for (const Decoded& d : decoded)
    ++ctx->launches, k_rt_decode<<<static_cast<unsigned int>((d.n + 255) / 256), 256, 0, st>>>(
        reinterpret_cast<const uint16_t*>(base + o_grid + d.cells_at),
        reinterpret_cast<const float*>(base + o_tab + d.table_at),
        reinterpret_cast<float*>(base + o_dec + d.out_at), d.n, d.one_minus);
  CSM_CUDA(cudaFuncSetAttribute(k_rt_score, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                static_cast<int>(std::max<size_t>(smem, 1024))));
  ++ctx->launches, k_rt_score<<<dim3((max_scans + kRtWarps - 1) / kRtWarps, num_jobs), kRtWarps * 32, smem, st>>>(
      reinterpret_cast<const RtJobDev*>(base + o_jobs), reinterpret_cast<const float*>(base + o_pts),
      reinterpret_cast<const float*>(base + o_quat),
      reinterpret_cast<unsigned long long*>(base + o_best), max_n,
      use_spill ? reinterpret_cast<int2*>(base + o_dec + dec_bytes) : nullptr);
  CSM_CUDA(cudaGetLastError());
  unsigned long long* best = reinterpret_cast<unsigned long long*>(pin + o_best);
  CSM_CUDA(cudaMemcpyAsync(best, base + o_best, sizeof(unsigned long long) * num_jobs,
                           cudaMemcpyDeviceToHost, st));
  CSM_CUDA(cudaStreamSynchronize(st));
  for (int i = 0; i < num_jobs; ++i) {
    const int side = 2 * jobs[i].linear_bound + 1;
    const unsigned int cand = 0xFFFFFFFFu - static_cast<unsigned int>(best[i] & 0xFFFFFFFFull);
    csm_result2d& r = results[i];
    r.found = 1;
    r.scan_index = static_cast<int>(cand / (side * side));
    const int c = static_cast<int>(cand % (side * side));
    r.x_index_offset = c / side - jobs[i].linear_bound;
    r.y_index_offset = c % side - jobs[i].linear_bound;
    r.score = ordered_to_float(static_cast<unsigned int>(best[i] >> 32));
    r.tie_resolution = 0;
    r.candidates_scored = static_cast<uint64_t>(side) * side * jobs[i].num_scans;
  }
  return CSM_OK;
}
