// Batched FastCorrelativeScanMatcher3D::MatchWithSearchParameters on the device.
//
// Reference: cartographer/mapping/internal/3d/scan_matching/
//   fast_correlative_scan_matcher_3d.cc:172-198  MatchWithSearchParameters
//   :200-244  DiscretizeScan            -> k3_discretize (+ per-lookup index arithmetic)
//   :297-326  GenerateLowestResolutionCandidates (scan -> z -> y -> x)
//   :328-352  ScoreCandidates           -> node_sum / children_sums
//   :363-369  GetPoseFromCandidate      -> low_resolution_score (leaf pose)
//   :371-440  BranchAndBound            -> k3_dive, k3_seed, k3_expand, k3_filter,
//                                          k3_collect, k3_ordered_dfs, k3_pick
//   low_resolution_matcher.cc:23-35     -> low_resolution_score
//
// The reference's depth-first search returns, among the leaves whose score exceeds
// min_score and whose low-resolution score reaches min_low_resolution_score, the one
// with the highest score; among equals, the one its sorted traversal reaches first.
// Here:
//   1. greedy dives, one warp per discrete scan, give an incumbent S (a leaf that
//      passed the filter);
//   2. a level-synchronous sweep expands every node whose bound is > S, eight
//      children per warp, and sends surviving leaves through the filter; the maximum
//      is S*;
//   3. the traversal order is reproduced only over nodes with bound >= S*: one warp per
//      lowest-resolution candidate walks its subtree in the reference's sibling order
//      (std::sort of <= 8 elements is an insertion sort, hence stable) and stops at the
//      first leaf that attains S* and passes the filter.  Lowest-resolution candidates
//      are ordered by score; equal scores are resolved on the host by replaying the
//      reference's std::sort.
#include <algorithm>
#include <cstring>
#include <functional>

#include "csm_match3d.cuh"

namespace csm {
namespace {

constexpr int kWarp = 32;
constexpr unsigned int kFull = 0xFFFFFFFFu;
constexpr int kCntContenders = CSM_MAX_DEPTH;
constexpr int kCntOverflow = CSM_MAX_DEPTH + 1;
constexpr int kCntWinners = CSM_MAX_DEPTH + 2;
constexpr int kCntDeferred = CSM_MAX_DEPTH + 3;
constexpr int kNumCounters3 = CSM_MAX_DEPTH + 4;
constexpr unsigned int kDeferredPerJob = 1024;  // more than this: walk in order instead

// PrecomputationGrid3D::ToProbability (precomputation_grid_3d.h:32-35) of the mean.
__device__ __forceinline__ float to_score3(int sum, float n_float) {
  constexpr float kMin = 0.1f;
  constexpr float kMax = 1.f - kMin;
  constexpr float kScale = (kMax - kMin) / 255.f;
  return __fadd_rn(kMin, __fmul_rn(__fdiv_rn(static_cast<float>(sum), n_float), kScale));
}

// Eigen 3.3 QuaternionBase::_transformVector, fp32, no contraction
// (oracle/csm_oracle.h QuatRotate).
__device__ __forceinline__ float3 quat_rotate(float qw, float qx, float qy, float qz, float3 v) {
  float ux = __fsub_rn(__fmul_rn(qy, v.z), __fmul_rn(qz, v.y));
  float uy = __fsub_rn(__fmul_rn(qz, v.x), __fmul_rn(qx, v.z));
  float uz = __fsub_rn(__fmul_rn(qx, v.y), __fmul_rn(qy, v.x));
  ux = __fadd_rn(ux, ux);
  uy = __fadd_rn(uy, uy);
  uz = __fadd_rn(uz, uz);
  const float cx = __fsub_rn(__fmul_rn(qy, uz), __fmul_rn(qz, uy));
  const float cy = __fsub_rn(__fmul_rn(qz, ux), __fmul_rn(qx, uz));
  const float cz = __fsub_rn(__fmul_rn(qx, uy), __fmul_rn(qy, ux));
  float3 r;
  r.x = __fadd_rn(__fadd_rn(v.x, __fmul_rn(qw, ux)), cx);
  r.y = __fadd_rn(__fadd_rn(v.y, __fmul_rn(qw, uy)), cy);
  r.z = __fadd_rn(__fadd_rn(v.z, __fmul_rn(qw, uz)), cz);
  return r;
}

// HybridGrid::GetCellIndex (hybrid_grid.h:428-433): lround(p / resolution) per axis.
__device__ __forceinline__ int cell_of(float v, float res) {
  return static_cast<int>(lroundf(__fdiv_rn(v, res)));
}

__device__ __forceinline__ bool inside(int v, int dim) {
  return static_cast<unsigned int>(v) < static_cast<unsigned int>(dim);
}

// Sum of level-d bytes under the discrete scan shifted by (ox, oy, oz) voxels
// (ScoreCandidates, :328-352; per-depth indices of DiscretizeScan, :213-243).
__device__ int node_sum(const Job3Dev& j, const int4* __restrict__ cells, int d, int ox, int oy, int oz,
                        int lane) {
  const Level3& L = j.sd->level[d];
  const int e = max(0, d - j.frd + 1);
  const int sx = -j.wxy, sz = -j.wz;
  const int kx = (ox >> e) - (sx >> e) - L.lo[0];
  const int ky = (oy >> e) - (sx >> e) - L.lo[1];
  const int kz = (oz >> e) - (sz >> e) - L.lo[2];
  const int nx = L.dim[0], ny = L.dim[1], nz = L.dim[2];
  const uint8_t* __restrict__ data = L.data;
  int sum = 0;
  for (int i = lane; i < j.n; i += kWarp) {
    const int4 c = cells[i];
    const int x = ((c.x + sx) >> e) + kx, y = ((c.y + sx) >> e) + ky, z = ((c.z + sz) >> e) + kz;
    if (inside(x, nx) && inside(y, ny) && inside(z, nz)) sum += data[(z * ny + y) * nx + x];
  }
  return __reduce_add_sync(kFull, sum);
}

// Children of the node (ox, oy, oz) at depth d >= 1 in generation order
// k = 4*zbit + 2*ybit + xbit (:412-430).  Returns the valid-child mask.  One 8-byte
// pack load per point delivers all eight child values; two children are accumulated
// per 32-bit register as 16-bit halves and flushed before they can overflow.
__device__ unsigned int children_sums(const Job3Dev& j, const int4* __restrict__ cells, int d, int ox,
                                      int oy, int oz, int lane, int sums[8]) {
  const int cd = d - 1;
  const int half = 1 << cd;
  const Level3& L = j.sd->level[cd];
  const int e = max(0, cd - j.frd + 1);
  const int stride = L.stride;  // == half >> e
  const int sx = -j.wxy, sz = -j.wz;
  const int kx = (ox >> e) - (sx >> e) - L.lo[0] + stride;  // pack coordinates of the first child
  const int ky = (oy >> e) - (sx >> e) - L.lo[1] + stride;
  const int kz = (oz >> e) - (sz >> e) - L.lo[2] + stride;
  const int px = L.pdim[0], py = L.pdim[1], pz = L.pdim[2];
  const uint2* __restrict__ pack = L.pack;
  const bool vx = ox + half <= j.wxy, vy = oy + half <= j.wxy, vz = oz + half <= j.wz;
  unsigned int valid = 1u;
  for (int k = 1; k < 8; ++k)
    if ((!(k & 1) || vx) && (!(k & 2) || vy) && (!(k & 4) || vz)) valid |= 1u << k;
  unsigned int p01 = 0, p23 = 0, p45 = 0, p67 = 0;
  int acc[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) acc[k] = 0;
  // 256 points per lane between flushes (256 * 255 < 65536)
  for (int base = lane; base < j.n; base += 256 * kWarp) {
    const int end = min(j.n, base + 256 * kWarp);
#pragma unroll 4
    for (int i = base; i < end; i += kWarp) {
      const int4 c = cells[i];
      const int xp = ((c.x + sx) >> e) + kx, yp = ((c.y + sx) >> e) + ky, zp = ((c.z + sz) >> e) + kz;
      if (inside(xp, px) && inside(yp, py) && inside(zp, pz)) {
        const uint2 v = pack[(zp * py + yp) * px + xp];
        p01 += __byte_perm(v.x, 0u, 0x4140);
        p23 += __byte_perm(v.x, 0u, 0x4342);
        p45 += __byte_perm(v.y, 0u, 0x4140);
        p67 += __byte_perm(v.y, 0u, 0x4342);
      }
    }
    acc[0] += p01 & 0xFFFF;
    acc[1] += p01 >> 16;
    acc[2] += p23 & 0xFFFF;
    acc[3] += p23 >> 16;
    acc[4] += p45 & 0xFFFF;
    acc[5] += p45 >> 16;
    acc[6] += p67 & 0xFFFF;
    acc[7] += p67 >> 16;
    p01 = p23 = p45 = p67 = 0;
  }
#pragma unroll
  for (int k = 0; k < 8; ++k) sums[k] = ((valid >> k) & 1u) ? __reduce_add_sync(kFull, acc[k]) : 0;
  return valid;
}

// CreateLowResolutionMatcher's lambda at the pose of a leaf
// (low_resolution_matcher.cc:23-35; pose from GetPoseFromCandidate, :363-369):
// translation = scan translation + resolution * offset, rotation = the normalised
// product the host passes as scan_leaf_quat.  The fp32 sum runs in point order.
__device__ float low_resolution_score(const Job3Dev& j, const float* __restrict__ low_points,
                                      const float* __restrict__ pose7, const float* __restrict__ leaf_q,
                                      int ox, int oy, int oz, int lane) {
  const Grid3Dev& g = *j.low;
  const float tx = __fadd_rn(pose7[0], __fmul_rn(j.res, static_cast<float>(ox)));
  const float ty = __fadd_rn(pose7[1], __fmul_rn(j.res, static_cast<float>(oy)));
  const float tz = __fadd_rn(pose7[2], __fmul_rn(j.res, static_cast<float>(oz)));
  const float qw = leaf_q[0], qx = leaf_q[1], qy = leaf_q[2], qz = leaf_q[3];
  const float* __restrict__ pts = low_points + 3 * j.low_off;
  float score = 0.f;
  for (int base = 0; base < j.n_low; base += kWarp) {
    const int i = base + lane;
    float p = 0.f;
    if (i < j.n_low) {
      float3 r = quat_rotate(qw, qx, qy, qz, make_float3(pts[3 * i], pts[3 * i + 1], pts[3 * i + 2]));
      const int x = cell_of(__fadd_rn(r.x, tx), j.low_res) - g.lo[0];
      const int y = cell_of(__fadd_rn(r.y, ty), j.low_res) - g.lo[1];
      const int z = cell_of(__fadd_rn(r.z, tz), j.low_res) - g.lo[2];
      p = (inside(x, g.dim[0]) && inside(y, g.dim[1]) && inside(z, g.dim[2]))
              ? g.data[(z * g.dim[1] + y) * g.dim[0] + x]
              : g.unknown;
    }
    const int count = min(kWarp, j.n_low - base);
    for (int k = 0; k < count; ++k) score = __fadd_rn(score, __shfl_sync(kFull, p, k));
  }
  return __fdiv_rn(score, static_cast<float>(j.n_low));
}

// Pruning threshold: min_score until a leaf has passed the filter, then the best such
// leaf's score (both strict, :384 and :407).  The incumbent moves while a kernel runs;
// any value read is a valid bound.
__device__ __forceinline__ float threshold_raw(const Job3Dev& j, const Job3State* st) {
  const unsigned int b = *reinterpret_cast<const volatile unsigned int*>(&st->best);
  return b ? fmaxf(j.min_score, ordered_to_float(b)) : j.min_score;
}
// Warp-uniform variant (lane 0 reads), for conditions that guard warp collectives.
__device__ __forceinline__ float threshold_of(const Job3Dev& j, const Job3State* st) {
  return __shfl_sync(kFull, threshold_raw(j, st), 0);
}

// Stable insertion of child k into idx[0..n) by descending score.
__device__ __forceinline__ void insert_sorted(const float score[8], int idx[8], int& n, int k) {
  int pos = n++;
  while (pos > 0 && score[idx[pos - 1]] < score[k]) {
    idx[pos] = idx[pos - 1];
    --pos;
  }
  idx[pos] = k;
}

// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(128)
k3_discretize(const Job3Dev* __restrict__ jobs, const float* __restrict__ points,
              const float* __restrict__ poses, int4* __restrict__ cells,
              unsigned long long* __restrict__ scan_best) {
  const Job3Dev& j = jobs[blockIdx.y];
  const int scan = blockIdx.x;
  if (scan >= j.S) return;
  const float* pose = poses + 7 * (j.scan_off + scan);
  const float tx = pose[0], ty = pose[1], tz = pose[2];
  const float qw = pose[3], qx = pose[4], qy = pose[5], qz = pose[6];
  const float* __restrict__ pts = points + 3 * j.pts_off;
  int4* __restrict__ out = cells + j.cell_off + static_cast<long long>(scan) * j.n;
  for (int i = threadIdx.x; i < j.n; i += blockDim.x) {
    const float3 r = quat_rotate(qw, qx, qy, qz, make_float3(pts[3 * i], pts[3 * i + 1], pts[3 * i + 2]));
    out[i] = make_int4(cell_of(__fadd_rn(r.x, tx), j.res), cell_of(__fadd_rn(r.y, ty), j.res),
                       cell_of(__fadd_rn(r.z, tz), j.res), 0);
  }
  if (threadIdx.x == 0) scan_best[j.scan_off + scan] = 0ull;
}

constexpr int kTopWarps3 = 4;

// Lowest-resolution candidates in generation order scan -> z -> y -> x (:314-326).
__global__ void __launch_bounds__(kTopWarps3 * kWarp)
k3_score_top(const Job3Dev* __restrict__ jobs, const int4* __restrict__ cells,
             float* __restrict__ top_scores, Job3State* __restrict__ state,
             unsigned long long* __restrict__ scan_best) {
  const Job3Dev& j = jobs[blockIdx.y];
  const int lane = threadIdx.x & 31;
  const int per_scan = j.nz * j.nxy * j.nxy;
  const long long total = static_cast<long long>(j.S) * per_scan;
  const int step = 1 << (j.depth - 1);
  unsigned long long scored = 0;
  for (long long c = blockIdx.x * kTopWarps3 + (threadIdx.x >> 5); c < total;
       c += static_cast<long long>(gridDim.x) * kTopWarps3) {
    const int scan = static_cast<int>(c / per_scan);
    int r = static_cast<int>(c - static_cast<long long>(scan) * per_scan);
    const int iz = r / (j.nxy * j.nxy);
    r -= iz * j.nxy * j.nxy;
    const int iy = r / j.nxy, ix = r - iy * j.nxy;
    const int sum = node_sum(j, cells + j.cell_off + static_cast<long long>(scan) * j.n, j.depth - 1,
                             -j.wxy + ix * step, -j.wxy + iy * step, -j.wz + iz * step, lane);
    if (lane == 0) {
      const float score = to_score3(sum, j.n_float);
      top_scores[j.cand_off + c] = score;
      atomicMax(&scan_best[j.scan_off + scan],
                (static_cast<unsigned long long>(float_to_ordered(score)) << 32) |
                    (0xFFFFFFFFull - static_cast<unsigned long long>(c)));
      ++scored;
    }
  }
  if (lane == 0 && scored) atomicAdd(&state[blockIdx.y].scored, scored);
}

constexpr int kDiveWarps3 = 4;

// One warp per discrete scan: follow the best child down to a leaf; the first leaf
// (in sorted order) that passes the filter becomes an incumbent.
__global__ void __launch_bounds__(kDiveWarps3 * kWarp)
k3_dive(const Job3Dev* __restrict__ jobs, const int4* __restrict__ cells,
        const float* __restrict__ low_points, const float* __restrict__ poses,
        const float* __restrict__ leaf_q, const unsigned long long* __restrict__ scan_best,
        Job3State* __restrict__ state) {
  const Job3Dev& j = jobs[blockIdx.y];
  const int scan = blockIdx.x * kDiveWarps3 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (scan >= j.S) return;
  Job3State* st = &state[blockIdx.y];
  const unsigned long long key = scan_best[j.scan_off + scan];
  if (key == 0ull) return;
  float score = ordered_to_float(static_cast<unsigned int>(key >> 32));
  if (!(score > threshold_of(j, st))) return;
  const long long c = static_cast<long long>(0xFFFFFFFFull - (key & 0xFFFFFFFFull));
  const int per_scan = j.nz * j.nxy * j.nxy;
  int r = static_cast<int>(c - static_cast<long long>(scan) * per_scan);
  const int iz = r / (j.nxy * j.nxy);
  r -= iz * j.nxy * j.nxy;
  const int iy = r / j.nxy, ix = r - iy * j.nxy;
  const int step = 1 << (j.depth - 1);
  int ox = -j.wxy + ix * step, oy = -j.wxy + iy * step, oz = -j.wz + iz * step;
  const int4* sc = cells + j.cell_off + static_cast<long long>(scan) * j.n;
  const float* pose = poses + 7 * (j.scan_off + scan);
  const float* lq = leaf_q + 4 * (j.scan_off + scan);
  unsigned long long scored = 0;
  bool alive = true;
  for (int d = j.depth - 1; d >= 1 && alive; --d) {
    int sums[8];
    const unsigned int valid = children_sums(j, sc, d, ox, oy, oz, lane, sums);
    const int half = 1 << (d - 1);
    float cs[8];
    int idx[8], n = 0;
    const float thr = threshold_of(j, st);
    for (int k = 0; k < 8; ++k) {
      cs[k] = to_score3(sums[k], j.n_float);
      if (!((valid >> k) & 1u)) continue;
      ++scored;
      if (cs[k] > thr) insert_sorted(cs, idx, n, k);
    }
    if (n == 0) {
      alive = false;
      break;
    }
    if (d > 1) {
      const int k = idx[0];
      ox += (k & 1) ? half : 0;
      oy += (k & 2) ? half : 0;
      oz += (k & 4) ? half : 0;
      score = cs[k];
      continue;
    }
    alive = false;
    for (int t = 0; t < n; ++t) {
      const int k = idx[t];
      if (!(cs[k] > threshold_of(j, st))) break;
      const int lx = ox + ((k & 1) ? half : 0), ly = oy + ((k & 2) ? half : 0), lz = oz + ((k & 4) ? half : 0);
      const float low = low_resolution_score(j, low_points, pose, lq, lx, ly, lz, lane);
      if (static_cast<double>(low) >= j.min_low) {
        if (lane == 0) atomicMax(&st->best, float_to_ordered(cs[k]));
        break;
      }
    }
  }
  if (j.depth == 1) {  // the lowest-resolution candidate is itself a leaf
    const float low = low_resolution_score(j, low_points, pose, lq, ox, oy, oz, lane);
    if (static_cast<double>(low) >= j.min_low && lane == 0) atomicMax(&st->best, float_to_ordered(score));
  }
  if (lane == 0 && scored) atomicAdd(&st->scored, scored);
}

// The incumbent as (has, value): a node is kept while its bound is strictly above the
// incumbent (:407), a leaf while it ties or beats it -- the sweep has to see every leaf
// that attains S* to reproduce which of them the reference reaches first.
struct Incumbent {
  bool has;
  float thr;  // max(min_score, incumbent)
};
__device__ __forceinline__ Incumbent incumbent_raw(const Job3Dev& j, const Job3State* st) {
  const unsigned int b = *reinterpret_cast<const volatile unsigned int*>(&st->best);
  Incumbent inc;
  inc.has = b != 0u;
  inc.thr = b ? fmaxf(j.min_score, ordered_to_float(b)) : j.min_score;
  return inc;
}
__device__ __forceinline__ Incumbent incumbent_of(const Job3Dev& j, const Job3State* st) {
  const unsigned int b = __shfl_sync(kFull, *reinterpret_cast<const volatile unsigned int*>(&st->best), 0);
  Incumbent inc;
  inc.has = b != 0u;
  inc.thr = b ? fmaxf(j.min_score, ordered_to_float(b)) : j.min_score;
  return inc;
}
__device__ __forceinline__ bool keep_node(const Incumbent& inc, float score) { return score > inc.thr; }
__device__ __forceinline__ bool keep_leaf(const Incumbent& inc, const Job3Dev& j, float score) {
  return score > j.min_score && (!inc.has || score >= inc.thr);
}
// A bound exactly equal to the incumbent is pruned like the reference does, but an
// S*-leaf may hide below it if that value turns out to be S*.  Such nodes are deferred:
// after the sweep those whose bound equals the final S* are expanded with a non-strict
// threshold.  A job that defers too many (a world of equal scores) is instead resolved
// by walking the tree in the reference's order, which finds its winner immediately.
struct Deferred {
  Node3S* list;
  unsigned int cap;
  unsigned int* counter;
};
__device__ __forceinline__ void defer_if_equal(const Incumbent& inc, Job3State* st, const Deferred& df,
                                               const Node3S& node, int depth) {
  if (!(inc.has && node.score == inc.thr)) return;
  if (atomicAdd(&st->eq_count, 1u) >= kDeferredPerJob) {
    st->use_dfs = 1u;
    return;
  }
  const unsigned int slot = atomicAdd(df.counter, 1u);
  if (slot >= df.cap) {
    st->use_dfs = 1u;
    return;
  }
  Node3S n = node;
  n.path = (n.path & 0x0FFFFFFFFFFFFFFFull) | (static_cast<unsigned long long>(depth) << 60);
  df.list[slot] = n;
}

// Lowest-resolution candidates of the jobs whose top depth is d.
__global__ void __launch_bounds__(256)
k3_seed(const Job3Dev* __restrict__ jobs, const float* __restrict__ top_scores,
        Job3State* __restrict__ state, Node3S* __restrict__ out, unsigned int* __restrict__ counters,
        int d, unsigned long long cap, Deferred df) {
  const Job3Dev& j = jobs[blockIdx.y];
  if (j.depth - 1 != d) return;
  Job3State* st = &state[blockIdx.y];
  const Incumbent inc = incumbent_raw(j, st);
  const int per_scan = j.nz * j.nxy * j.nxy;
  const long long total = static_cast<long long>(j.S) * per_scan;
  const int step = 1 << d;
  for (long long c = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; c < total;
       c += static_cast<long long>(gridDim.x) * blockDim.x) {
    const float score = top_scores[j.cand_off + c];
    const bool keep = d == 0 ? keep_leaf(inc, j, score) : keep_node(inc, score);
    if (!keep && !(d != 0 && inc.has && score == inc.thr)) continue;
    const int scan = static_cast<int>(c / per_scan);
    int r = static_cast<int>(c - static_cast<long long>(scan) * per_scan);
    const int iz = r / (j.nxy * j.nxy);
    r -= iz * j.nxy * j.nxy;
    const int iy = r / j.nxy, ix = r - iy * j.nxy;
    Node3S nd;
    nd.job_scan = (blockIdx.y << 16) | scan;
    nd.ox = -j.wxy + ix * step;
    nd.oy = -j.wxy + iy * step;
    nd.oz = -j.wz + iz * step;
    nd.score = score;
    nd.top_cand = static_cast<unsigned int>(c);
    nd.path = 0ull;
    if (!keep) {
      defer_if_equal(inc, st, df, nd, d);
      continue;
    }
    const unsigned int slot = atomicAdd(&counters[d], 1u);
    if (slot >= cap) {
      atomicOr(&counters[kCntOverflow], 1u);
      continue;
    }
    out[slot] = nd;
  }
}

// Second sweep: deferred nodes at depth d whose bound equals the final S*.
__global__ void __launch_bounds__(256)
k3_seed_deferred(const Job3Dev* __restrict__ jobs, const Job3State* __restrict__ state,
                 const Node3S* __restrict__ deferred, unsigned int deferred_cap, Node3S* __restrict__ out,
                 unsigned int* __restrict__ counters, int d, unsigned long long cap) {
  const unsigned int count = min(counters[kCntDeferred], deferred_cap);
  for (unsigned int i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += gridDim.x * blockDim.x) {
    Node3S nd = deferred[i];
    if (static_cast<int>(nd.path >> 60) != d) continue;
    const Job3State& st = state[nd.job_scan >> 16];
    if (st.use_dfs || float_to_ordered(nd.score) != st.best) continue;
    nd.path &= 0x0FFFFFFFFFFFFFFFull;
    const unsigned int slot = atomicAdd(&counters[d], 1u);
    if (slot >= cap) {
      atomicOr(&counters[kCntOverflow], 1u);
      continue;
    }
    out[slot] = nd;
  }
}

constexpr int kExpandWarps3 = 8;

// One warp per node at depth d: score its (up to) eight children at depth d - 1, keep
// those whose bound beats the incumbent (leaves: ties it as well).
__global__ void __launch_bounds__(kExpandWarps3 * kWarp)
k3_expand(const Job3Dev* __restrict__ jobs, const int4* __restrict__ cells,
          Job3State* __restrict__ state, const Node3S* __restrict__ in, Node3S* __restrict__ out,
          unsigned int* __restrict__ counters, int d, unsigned long long cap, Deferred df, bool second) {
  const unsigned int count = static_cast<unsigned int>(min(static_cast<unsigned long long>(counters[d]), cap));
  const int lane = threadIdx.x & 31;
  const unsigned int warps = gridDim.x * kExpandWarps3;
  // the second sweep keeps whatever ties S*, like the first does for leaves
  const bool to_leaves = d == 1 || second;
  for (unsigned int i = blockIdx.x * kExpandWarps3 + (threadIdx.x >> 5); i < count; i += warps) {
    const Node3S nd = in[i];
    const int job = nd.job_scan >> 16, scan = nd.job_scan & 0xFFFF;
    const Job3Dev& j = jobs[job];
    Job3State* st = &state[job];
    Incumbent inc = incumbent_of(j, st);
    if (!(second ? keep_leaf(inc, j, nd.score) : keep_node(inc, nd.score))) {  // the incumbent improved
      if (lane == 0 && !second) defer_if_equal(inc, st, df, nd, d);
      continue;
    }
    int sums[8];
    const unsigned int valid = children_sums(j, cells + j.cell_off + static_cast<long long>(scan) * j.n, d,
                                             nd.ox, nd.oy, nd.oz, lane, sums);
    int mine = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) mine = (lane == k) ? sums[k] : mine;
    const float score = to_score3(mine, j.n_float);
    // rank among the valid siblings after the reference's stable descending sort
    int rank = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const float other = __shfl_sync(kFull, score, k);
      if (((valid >> k) & 1u) && (other > score || (other == score && k < lane))) ++rank;
    }
    inc = incumbent_of(j, st);
    const bool is_child = lane < 8 && ((valid >> lane) & 1u);
    const bool keep = is_child && (to_leaves ? keep_leaf(inc, j, score) : keep_node(inc, score));
    const int half = 1 << (d - 1);
    Node3S ch;
    ch.job_scan = nd.job_scan;
    ch.ox = nd.ox + ((lane & 1) ? half : 0);
    ch.oy = nd.oy + ((lane & 2) ? half : 0);
    ch.oz = nd.oz + ((lane & 4) ? half : 0);
    ch.score = score;
    ch.top_cand = nd.top_cand;
    ch.path = (nd.path << 3) | static_cast<unsigned long long>(rank);
    if (is_child && !keep && !to_leaves) defer_if_equal(inc, st, df, ch, d - 1);
    const unsigned int m = __ballot_sync(kFull, keep);
    if (m) {
      unsigned int base = 0;
      if (lane == 0) base = atomicAdd(&counters[d - 1], static_cast<unsigned int>(__popc(m)));
      base = __shfl_sync(kFull, base, 0);
      if (keep) {
        const unsigned int slot = base + __popc(m & ((1u << lane) - 1u));
        if (slot < cap) {
          out[slot] = ch;
        } else {
          atomicOr(&counters[kCntOverflow], 1u);
        }
      }
    }
    if (lane == 0) atomicAdd(&st->scored, static_cast<unsigned long long>(__popc(valid)));
  }
}

// Leaves that still tie or beat the incumbent go through the low-resolution filter;
// those that pass raise the incumbent and are recorded.
__global__ void __launch_bounds__(kExpandWarps3 * kWarp)
k3_filter(const Job3Dev* __restrict__ jobs, const float* __restrict__ low_points,
          const float* __restrict__ poses, const float* __restrict__ leaf_q, Job3State* __restrict__ state,
          const Node3S* __restrict__ leaves, unsigned int* __restrict__ counters, unsigned long long cap,
          Winner3* __restrict__ winners, unsigned int winner_cap) {
  const unsigned int count = static_cast<unsigned int>(min(static_cast<unsigned long long>(counters[0]), cap));
  const int lane = threadIdx.x & 31;
  const unsigned int warps = gridDim.x * kExpandWarps3;
  for (unsigned int i = blockIdx.x * kExpandWarps3 + (threadIdx.x >> 5); i < count; i += warps) {
    const Node3S nd = leaves[i];
    const int job = nd.job_scan >> 16, scan = nd.job_scan & 0xFFFF;
    const Job3Dev& j = jobs[job];
    Job3State* st = &state[job];
    if (!keep_leaf(incumbent_of(j, st), j, nd.score)) continue;
    const float low = low_resolution_score(j, low_points, poses + 7 * (j.scan_off + scan),
                                           leaf_q + 4 * (j.scan_off + scan), nd.ox, nd.oy, nd.oz, lane);
    if (static_cast<double>(low) >= j.min_low && lane == 0) {
      atomicMax(&st->best, float_to_ordered(nd.score));
      const unsigned int slot = atomicAdd(&counters[kCntWinners], 1u);
      if (slot < winner_cap) {
        Winner3 w;
        w.job = job;
        w.scan = scan;
        w.leaf[0] = nd.ox;
        w.leaf[1] = nd.oy;
        w.leaf[2] = nd.oz;
        w.cand = nd.top_cand;
        w.score = nd.score;
        w.low = low;
        w.path = nd.path;
        winners[slot] = w;
      } else {
        atomicOr(&counters[kCntOverflow], 8u);
      }
    }
  }
}

// Which recorded S*-leaf does the reference reach first?  The one below the best-
// ranked lowest-resolution candidate (score descending; equal scores are left to the
// host's std::sort replay), and there the one with the smallest path.  One block: the
// list is short, the four passes are separated by block barriers.
__global__ void __launch_bounds__(1024)
k3_pick_winner(const Job3Dev* __restrict__ jobs, const float* __restrict__ top_scores,
               Job3State* __restrict__ state, const Winner3* __restrict__ winners,
               const unsigned int* __restrict__ counters, unsigned int winner_cap) {
  const unsigned int count = min(counters[kCntWinners], winner_cap);
  for (int pass = 0; pass < 4; ++pass) {
    for (unsigned int i = threadIdx.x; i < count; i += blockDim.x) {
      const Winner3 w = winners[i];
      Job3State* st = &state[w.job];
      if (float_to_ordered(w.score) != st->best) continue;
      const unsigned int top = float_to_ordered(top_scores[jobs[w.job].cand_off + w.cand]);
      if (pass == 0) {
        atomicMax(&st->w_group, top);
        continue;
      }
      if (top != st->w_group) continue;
      if (pass == 1) {
        atomicMin(&st->w_cand_min, w.cand);
        atomicMax(&st->w_cand_max, w.cand);
        continue;
      }
      if (w.cand != st->w_cand_min) continue;
      if (pass == 2) {
        atomicMin(&st->w_path_min, w.path);
        continue;
      }
      if (w.path == st->w_path_min) atomicMin(&st->w_slot, i);
    }
    __syncthreads();
  }
}

// ---- phase 2: traversal order ------------------------------------------------
__global__ void __launch_bounds__(256)
k3_collect(const Job3Dev* __restrict__ jobs, const float* __restrict__ top_scores,
           const Job3State* __restrict__ state, Contender3* __restrict__ out,
           unsigned int* __restrict__ counters, unsigned int cap) {
  const Job3Dev& j = jobs[blockIdx.y];
  const unsigned int best = state[blockIdx.y].best;
  if (best == 0u || !state[blockIdx.y].use_dfs) return;  // the sweep saw every S*-leaf
  const float s_star = ordered_to_float(best);
  const int per_scan = j.nz * j.nxy * j.nxy;
  const long long total = static_cast<long long>(j.S) * per_scan;
  const int step = 1 << (j.depth - 1);
  for (long long c = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; c < total;
       c += static_cast<long long>(gridDim.x) * blockDim.x) {
    const float score = top_scores[j.cand_off + c];
    if (!(score >= s_star)) continue;
    const unsigned int slot = atomicAdd(&counters[kCntContenders], 1u);
    if (slot >= cap) {
      atomicOr(&counters[kCntOverflow], 4u);
      continue;
    }
    const int scan = static_cast<int>(c / per_scan);
    int r = static_cast<int>(c - static_cast<long long>(scan) * per_scan);
    const int iz = r / (j.nxy * j.nxy);
    r -= iz * j.nxy * j.nxy;
    const int iy = r / j.nxy, ix = r - iy * j.nxy;
    Contender3 ct;
    ct.job = blockIdx.y;
    ct.scan = scan;
    ct.ox = -j.wxy + ix * step;
    ct.oy = -j.wxy + iy * step;
    ct.oz = -j.wz + iz * step;
    ct.cand = static_cast<unsigned int>(c);
    ct.score = score;
    ct.leaf[0] = ct.leaf[1] = ct.leaf[2] = 0;
    ct.has_leaf = 0;
    ct.leaf_low = 0.f;
    out[slot] = ct;
  }
}

__global__ void __launch_bounds__(128)
k3_ordered_dfs(const Job3Dev* __restrict__ jobs, const int4* __restrict__ cells,
               const float* __restrict__ low_points, const float* __restrict__ poses,
               const float* __restrict__ leaf_q, Job3State* __restrict__ state,
               Contender3* __restrict__ list, const unsigned int* __restrict__ counters, unsigned int cap) {
  const unsigned int count = min(counters[kCntContenders], cap);
  const int lane = threadIdx.x & 31;
  const unsigned int warps = gridDim.x * (blockDim.x >> 5);
  for (unsigned int i = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); i < count; i += warps) {
    const Contender3 ct = list[i];
    const Job3Dev& j = jobs[ct.job];
    const float s_star = ordered_to_float(state[ct.job].best);
    const int4* sc = cells + j.cell_off + static_cast<long long>(ct.scan) * j.n;
    const float* pose = poses + 7 * (j.scan_off + ct.scan);
    const float* lq = leaf_q + 4 * (j.scan_off + ct.scan);
    const int top_depth = j.depth - 1;
    int fx[CSM_MAX_DEPTH], fy[CSM_MAX_DEPTH], fz[CSM_MAX_DEPTH];
    unsigned int forder[CSM_MAX_DEPTH];  // 8 x 3-bit child indices, low bits first
    int fcount[CSM_MAX_DEPTH], fnext[CSM_MAX_DEPTH];
    int lv = 0;
    fx[0] = ct.ox;
    fy[0] = ct.oy;
    fz[0] = ct.oz;
    fcount[0] = -1;
    bool found = false;
    int leaf[3] = {0, 0, 0};
    float leaf_low = 0.f;
    unsigned long long scored = 0;
    if (top_depth == 0) {
      const float low = low_resolution_score(j, low_points, pose, lq, ct.ox, ct.oy, ct.oz, lane);
      if (static_cast<double>(low) >= j.min_low) {
        found = true;
        leaf[0] = ct.ox;
        leaf[1] = ct.oy;
        leaf[2] = ct.oz;
        leaf_low = low;
      }
      lv = -1;
    }
    while (lv >= 0 && !found) {
      const int d = top_depth - lv;
      const int half = 1 << (d - 1);
      if (fcount[lv] < 0) {
        int sums[8];
        const unsigned int valid = children_sums(j, sc, d, fx[lv], fy[lv], fz[lv], lane, sums);
        float cs[8];
        int idx[8], n = 0;
        for (int k = 0; k < 8; ++k) {
          cs[k] = to_score3(sums[k], j.n_float);
          if (!((valid >> k) & 1u)) continue;
          ++scored;
          if (cs[k] >= s_star && cs[k] > j.min_score) insert_sorted(cs, idx, n, k);
        }
        if (d == 1) {
          for (int t = 0; t < n && !found; ++t) {
            const int k = idx[t];
            const int lx = fx[lv] + ((k & 1) ? half : 0), ly = fy[lv] + ((k & 2) ? half : 0),
                      lz = fz[lv] + ((k & 4) ? half : 0);
            const float low = low_resolution_score(j, low_points, pose, lq, lx, ly, lz, lane);
            if (static_cast<double>(low) >= j.min_low) {
              found = true;
              leaf[0] = lx;
              leaf[1] = ly;
              leaf[2] = lz;
              leaf_low = low;
            }
          }
          if (!found) --lv;
          continue;
        }
        unsigned int packed = 0;
        for (int k = 0; k < n; ++k) packed |= static_cast<unsigned int>(idx[k]) << (3 * k);
        forder[lv] = packed;
        fcount[lv] = n;
        fnext[lv] = 0;
      }
      if (fnext[lv] < fcount[lv]) {
        const int k = (forder[lv] >> (3 * fnext[lv])) & 7;
        ++fnext[lv];
        fx[lv + 1] = fx[lv] + ((k & 1) ? half : 0);
        fy[lv + 1] = fy[lv] + ((k & 2) ? half : 0);
        fz[lv + 1] = fz[lv] + ((k & 4) ? half : 0);
        fcount[lv + 1] = -1;
        ++lv;
      } else {
        --lv;
      }
    }
    if (lane == 0) {
      list[i].has_leaf = found ? 1 : 0;
      list[i].leaf[0] = leaf[0];
      list[i].leaf[1] = leaf[1];
      list[i].leaf[2] = leaf[2];
      list[i].leaf_low = leaf_low;
      if (found) atomicMax(&state[ct.job].group_best, float_to_ordered(ct.score));
      atomicAdd(&state[ct.job].scored, scored);
    }
  }
}

__global__ void __launch_bounds__(256)
k3_pick(Job3State* __restrict__ state, const Contender3* __restrict__ list,
        const unsigned int* __restrict__ counters, unsigned int cap) {
  const unsigned int count = min(counters[kCntContenders], cap);
  const unsigned int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= count) return;
  const Contender3 ct = list[i];
  if (!ct.has_leaf) return;
  Job3State* st = &state[ct.job];
  if (float_to_ordered(ct.score) != st->group_best) return;
  atomicAdd(&st->tie_count, 1u);
  atomicMin(&st->win, (static_cast<unsigned long long>(ct.cand) << 32) | i);
}

__global__ void __launch_bounds__(128)
k3_finalize(const Job3Dev* __restrict__ jobs, const Job3State* __restrict__ state,
            const Contender3* __restrict__ list, const Winner3* __restrict__ winners,
            csm_result3d* __restrict__ results, int num_jobs) {
  const int job = blockIdx.x * blockDim.x + threadIdx.x;
  if (job >= num_jobs) return;
  const Job3State st = state[job];
  csm_result3d r;
  r.found = 0;
  r.scan_index = 0;
  r.offset[0] = r.offset[1] = r.offset[2] = 0;
  r.score = jobs[job].min_score;
  r.low_resolution_score = 0.f;
  r.tie_resolution = 0;
  r.candidates_scored = st.scored;
  if (st.best != 0u) {
    r.score = ordered_to_float(st.best);
    if (st.use_dfs) {  // traversal-order fallback decided
      if (st.win == ~0ull) {
        r.found = -1;
      } else {
        const Contender3 ct = list[st.win & 0xFFFFFFFFull];
        r.found = 1;
        r.scan_index = ct.scan;
        r.offset[0] = ct.leaf[0];
        r.offset[1] = ct.leaf[1];
        r.offset[2] = ct.leaf[2];
        r.low_resolution_score = ct.leaf_low;
        r.tie_resolution = st.tie_count > 1u ? 2 : 0;  // 2: host replays std::sort over contenders
      }
    } else if (st.w_slot == 0xFFFFFFFFu) {
      r.found = -1;
    } else {
      const Winner3 w = winners[st.w_slot];
      r.found = 1;
      r.scan_index = w.scan;
      r.offset[0] = w.leaf[0];
      r.offset[1] = w.leaf[1];
      r.offset[2] = w.leaf[2];
      r.low_resolution_score = w.low;
      r.tie_resolution = st.w_cand_min != st.w_cand_max ? 3 : 0;  // 3: replay over recorded leaves
    }
  }
  results[job] = r;
}

__global__ void __launch_bounds__(128)
k3_state_init(Job3State* __restrict__ state, const Job3Dev* __restrict__ jobs, int num_jobs) {
  const int job = blockIdx.x * blockDim.x + threadIdx.x;
  if (job >= num_jobs) return;
  Job3State s;
  s.best = 0u;
  s.eq_count = 0u;
  s.w_group = 0u;
  s.w_cand_min = 0xFFFFFFFFu;
  s.w_cand_max = 0u;
  s.w_slot = 0xFFFFFFFFu;
  s.w_path_min = ~0ull;
  s.group_best = 0u;
  s.win = ~0ull;
  s.tie_count = 0u;
  s.use_dfs = 0u;
  const Job3Dev& j = jobs[job];
  s.total = static_cast<unsigned int>(static_cast<long long>(j.S) * j.nz * j.nxy * j.nxy);
  s.scored = 0ull;
  state[job] = s;
}

}  // namespace
}  // namespace csm

// ===========================================================================
// host side
// ===========================================================================
using namespace csm;

namespace {

size_t align_up3(size_t v) { return (v + 255) & ~static_cast<size_t>(255); }

struct Carve {
  size_t at = 0;
  size_t take(size_t bytes) {
    const size_t off = at;
    at += align_up3(bytes);
    return off;
  }
};

int allocate_batch3(csm_batch3d* b, unsigned int contender_cap, unsigned int winner_cap) {
  if (b->arena) b->owner->pool.give(false, b->arena, b->arena_cap);
  b->arena = nullptr;
  b->contender_cap = contender_cap;
  b->winner_cap = winner_cap;
  Carve c;
  const size_t o_jobs = c.take(sizeof(Job3Dev) * b->num_jobs);
  const size_t o_pts = c.take(sizeof(float) * 3 * b->total_pts);
  const size_t o_low = c.take(sizeof(float) * 3 * b->total_low);
  const size_t o_pose = c.take(sizeof(float) * 7 * b->total_scans);
  const size_t o_lq = c.take(sizeof(float) * 4 * b->total_scans);
  const size_t o_cells = c.take(sizeof(int4) * b->total_cells);
  const size_t o_sbest = c.take(sizeof(unsigned long long) * b->total_scans);
  const size_t o_top = c.take(sizeof(float) * b->total_cand);
  const size_t o_state = c.take(sizeof(Job3State) * b->num_jobs);
  const size_t o_cont = c.take(sizeof(Contender3) * static_cast<size_t>(contender_cap));
  const size_t o_win = c.take(sizeof(Winner3) * static_cast<size_t>(winner_cap));
  b->deferred_cap = static_cast<unsigned int>(
      std::min<unsigned long long>(static_cast<unsigned long long>(b->num_jobs) * kDeferredPerJob, 1ull << 20));
  const size_t o_def = c.take(sizeof(Node3S) * static_cast<size_t>(b->deferred_cap));
  const size_t o_cnt = c.take(sizeof(unsigned int) * kNumCounters3);
  const size_t o_res = c.take(sizeof(csm_result3d) * b->num_jobs);
  b->arena_bytes = std::max<size_t>(c.at, 256);
  if (b->owner->pool.take(false, b->arena_bytes, &b->arena, &b->arena_cap) != CSM_OK) {
    set_error("allocation of %zu bytes for a 3D batch failed: %s", b->arena_bytes,
              cudaGetErrorString(cudaGetLastError()));
    b->arena = nullptr;
    return CSM_ERR_CUDA;
  }
  char* base = static_cast<char*>(b->arena);
  b->jobs = reinterpret_cast<Job3Dev*>(base + o_jobs);
  b->points = reinterpret_cast<float*>(base + o_pts);
  b->low_points = reinterpret_cast<float*>(base + o_low);
  b->scan_poses = reinterpret_cast<float*>(base + o_pose);
  b->scan_leaf_q = reinterpret_cast<float*>(base + o_lq);
  b->cells = reinterpret_cast<int4*>(base + o_cells);
  b->scan_best = reinterpret_cast<unsigned long long*>(base + o_sbest);
  b->top_scores = reinterpret_cast<float*>(base + o_top);
  b->state = reinterpret_cast<Job3State*>(base + o_state);
  b->contenders = reinterpret_cast<Contender3*>(base + o_cont);
  b->winners = reinterpret_cast<Winner3*>(base + o_win);
  b->deferred = reinterpret_cast<Node3S*>(base + o_def);
  b->counters = reinterpret_cast<unsigned int*>(base + o_cnt);
  b->results_dev = reinterpret_cast<csm_result3d*>(base + o_res);
  return CSM_OK;
}

// Pinned staging layout: jobs | points | low points | poses | leaf quats | results.
struct PinLayout {
  size_t jobs, pts, low, pose, lq, res, total;
};
PinLayout pin_layout(const csm_batch3d* b) {
  Carve c;
  PinLayout p;
  p.jobs = c.take(sizeof(Job3Dev) * b->num_jobs);
  p.pts = c.take(sizeof(float) * 3 * b->total_pts);
  p.low = c.take(sizeof(float) * 3 * b->total_low);
  p.pose = c.take(sizeof(float) * 7 * b->total_scans);
  p.lq = c.take(sizeof(float) * 4 * b->total_scans);
  p.res = c.take(sizeof(csm_result3d) * b->num_jobs);
  p.total = std::max<size_t>(c.at, 256);
  return p;
}

int copy_inputs3(csm_ctx* ctx, csm_batch3d* b) {
  const PinLayout p = pin_layout(b);
  char* pin = static_cast<char*>(b->pinned);
  cudaStream_t st = ctx->stream;
  CSM_CUDA(cudaMemcpyAsync(b->jobs, pin + p.jobs, sizeof(Job3Dev) * b->num_jobs, cudaMemcpyHostToDevice, st));
  if (b->total_pts)
    CSM_CUDA(cudaMemcpyAsync(b->points, pin + p.pts, sizeof(float) * 3 * b->total_pts, cudaMemcpyHostToDevice, st));
  if (b->total_low)
    CSM_CUDA(cudaMemcpyAsync(b->low_points, pin + p.low, sizeof(float) * 3 * b->total_low, cudaMemcpyHostToDevice, st));
  if (b->total_scans) {
    CSM_CUDA(cudaMemcpyAsync(b->scan_poses, pin + p.pose, sizeof(float) * 7 * b->total_scans, cudaMemcpyHostToDevice, st));
    CSM_CUDA(cudaMemcpyAsync(b->scan_leaf_q, pin + p.lq, sizeof(float) * 4 * b->total_scans, cudaMemcpyHostToDevice, st));
  }
  return CSM_OK;
}

}  // namespace

extern "C" {

void csm_batch3d_destroy(csm_ctx* ctx, csm_batch3d* b) {
  if (!b) return;
  if (ctx) {
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
  }
  for (cudaEvent_t e : b->prof_events) cudaEventDestroy(e);
  if (b->arena) {
    if (b->owner) b->owner->pool.give(false, b->arena, b->arena_cap);
    else cudaFree(b->arena);
  }
  if (b->pinned) {
    if (b->owner) b->owner->pool.give(true, b->pinned, b->pinned_cap);
    else cudaFreeHost(b->pinned);
  }
  delete b;
}

int csm_batch3d_upload(csm_ctx* ctx, const csm_job3d* jobs, int num_jobs, csm_batch3d** out) {
  CSM_REQUIRE(ctx != nullptr && jobs != nullptr && out != nullptr, "NULL argument");
  CSM_REQUIRE(num_jobs >= 1 && num_jobs <= 16384, "1 <= num_jobs <= 16384 per batch");
  CSM_CUDA(cudaSetDevice(ctx->device));
  *out = nullptr;
  csm_batch3d* b = new csm_batch3d;
  b->owner = ctx;
  b->num_jobs = num_jobs;
  b->jobs_host.resize(num_jobs);
  auto fail = [&](int rc) {
    csm_batch3d_destroy(ctx, b);
    return rc;
  };
  for (int i = 0; i < num_jobs; ++i) {
    const csm_job3d& in = jobs[i];
    Job3Dev& j = b->jobs_host[i];
    memset(&j, 0, sizeof(j));
    if (in.stack == nullptr || in.low_resolution_grid == nullptr) {
      set_error("job %d: NULL stack or low-resolution grid", i);
      return fail(CSM_ERR_ARG);
    }
    if (in.stack->device != ctx->device || in.low_resolution_grid->device != ctx->device) {
      set_error("job %d: grids live on another device than the context", i);
      return fail(CSM_ERR_ARG);
    }
    if (in.num_points < 0 || in.num_low_points < 0 || in.num_scans < 0 || in.num_scans >= 65536 ||
        in.linear_xy_window < 0 || in.linear_z_window < 0 || !(in.resolution > 0.f) || !(in.low_resolution > 0.f) ||
        (in.num_points > 0 && in.points_xyz == nullptr) || (in.num_low_points > 0 && in.low_points_xyz == nullptr) ||
        (in.num_scans > 0 && (in.scan_poses == nullptr || in.scan_leaf_quat == nullptr))) {
      set_error("job %d: invalid sizes or NULL arrays", i);
      return fail(CSM_ERR_ARG);
    }
    j.sd = in.stack->dev;
    j.low = in.low_resolution_grid->dev;
    j.depth = in.stack->depth;
    j.frd = in.stack->frd;
    j.n = in.num_points;
    j.n_low = in.num_low_points;
    // An empty cloud scores 0/0 = NaN, which never compares greater than min_score:
    // the reference returns nullptr (:189-197).  No device work for such a job.
    j.S = (in.num_points == 0 || in.num_low_points == 0) ? 0 : in.num_scans;
    j.wxy = in.linear_xy_window;
    j.wz = in.linear_z_window;
    const int step = 1 << (j.depth - 1);
    j.nxy = 2 * j.wxy / step + 1;  // offsets -w, -w + step, ... <= w
    j.nz = 2 * j.wz / step + 1;
    const long long cand = static_cast<long long>(j.S) * j.nz * j.nxy * j.nxy;
    if (cand >= (1ll << 31)) {
      set_error("job %d: %lld lowest-resolution candidates exceed the 2^31 limit", i, cand);
      return fail(CSM_ERR_ARG);
    }
    j.pts_off = b->total_pts;
    j.low_off = b->total_low;
    j.scan_off = b->total_scans;
    j.cell_off = b->total_cells;
    j.cand_off = b->total_cand;
    j.res = in.resolution;
    j.low_res = in.low_resolution;
    j.min_score = in.min_score;
    j.n_float = static_cast<float>(in.num_points);
    j.min_low = in.min_low_resolution_score;
    b->total_pts += j.n;
    b->total_low += j.n_low;
    b->total_scans += j.S;
    b->total_cells += static_cast<long long>(j.S) * j.n;
    b->total_cand += cand;
    b->max_scans = std::max(b->max_scans, j.S);
    b->max_depth = std::max(b->max_depth, j.depth);
  }
  b->frontier_want = std::max<unsigned long long>(static_cast<unsigned long long>(b->total_cand), 1ull << 21);
  const PinLayout p = pin_layout(b);
  b->pinned_bytes = p.total;
  b->pin_results_off = p.res;
  if (ctx->pool.take(true, b->pinned_bytes, &b->pinned, &b->pinned_cap) != CSM_OK) {
    set_error("pinned allocation of %zu bytes failed", b->pinned_bytes);
    b->pinned = nullptr;
    return fail(CSM_ERR_CUDA);
  }
  char* pin = static_cast<char*>(b->pinned);
  memcpy(pin + p.jobs, b->jobs_host.data(), sizeof(Job3Dev) * num_jobs);
  for (int i = 0; i < num_jobs; ++i) {
    const csm_job3d& in = jobs[i];
    const Job3Dev& j = b->jobs_host[i];
    if (j.n) memcpy(pin + p.pts + sizeof(float) * 3 * j.pts_off, in.points_xyz, sizeof(float) * 3 * j.n);
    if (j.n_low) memcpy(pin + p.low + sizeof(float) * 3 * j.low_off, in.low_points_xyz, sizeof(float) * 3 * j.n_low);
    if (j.S) {
      memcpy(pin + p.pose + sizeof(float) * 7 * j.scan_off, in.scan_poses, sizeof(float) * 7 * j.S);
      memcpy(pin + p.lq + sizeof(float) * 4 * j.scan_off, in.scan_leaf_quat, sizeof(float) * 4 * j.S);
    }
  }
  const unsigned long long ccap = std::min<unsigned long long>(
      std::max<unsigned long long>(static_cast<unsigned long long>(b->total_cand) / 4 + 4096ull, 1ull << 16),
      0xFFFFFFF0ull);
  int rc = allocate_batch3(b, static_cast<unsigned int>(ccap), 1u << 16);
  if (rc == CSM_OK) rc = copy_inputs3(ctx, b);
  if (rc != CSM_OK) return fail(rc);
  b->uploaded = true;
  *out = b;
  return CSM_OK;
}

int csm_batch3d_run(csm_ctx* ctx, csm_batch3d* b) {
  CSM_REQUIRE(ctx != nullptr && b != nullptr && b->uploaded, "batch was not uploaded");
  CSM_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t st = ctx->stream;
  const size_t want_bytes = sizeof(Node3S) * b->frontier_want;
  if (ctx->frontier[0].cap < want_bytes || ctx->frontier[1].cap < want_bytes) {
    CSM_CUDA(cudaStreamSynchronize(st));
    int rc = ctx->frontier[0].reserve(want_bytes);
    if (rc == CSM_OK) rc = ctx->frontier[1].reserve(want_bytes);
    if (rc != CSM_OK) return rc;
  }
  Node3S* const fbuf[2] = {ctx->frontier[0].as<Node3S>(), ctx->frontier[1].as<Node3S>()};
  const unsigned long long fcap = std::min<unsigned long long>(
      std::min(ctx->frontier[0].cap, ctx->frontier[1].cap) / sizeof(Node3S), 0xFFFFFFF0ull);
  const int nj = b->num_jobs;
  CSM_CUDA(cudaMemsetAsync(b->counters, 0, sizeof(unsigned int) * kNumCounters3, st));
  for (cudaEvent_t e : b->prof_events) cudaEventDestroy(e);
  b->prof_events.clear();
  // csm_ctx_set_profiling phases: 0 discretize, 1 lowest-resolution scoring, 2 dives,
  // 3 sweep + leaf filter, 4 traversal-order phase.
  auto mark = [&]() {
    if (!ctx->profiling) return;
    cudaEvent_t e;
    cudaEventCreate(&e);
    cudaEventRecord(e, st);
    b->prof_events.push_back(e);
  };
  mark();
  ++ctx->launches, k3_state_init<<<(nj + 127) / 128, 128, 0, st>>>(b->state, b->jobs, nj);
  if (b->max_scans > 0) {
    const dim3 grid_sj(b->max_scans, nj);
    const int wide = std::max(1, std::min(ctx->num_sms * 8, 65535));
    ++ctx->launches, k3_discretize<<<grid_sj, 128, 0, st>>>(b->jobs, b->points, b->scan_poses, b->cells, b->scan_best);
    mark();
    // enough blocks per job to cover its candidates, bounded so the grid stays modest
    long long max_cand = 0;
    for (const Job3Dev& j : b->jobs_host)
      max_cand = std::max(max_cand, static_cast<long long>(j.S) * j.nz * j.nxy * j.nxy);
    const int top_blocks = static_cast<int>(std::min<long long>((max_cand + kTopWarps3 - 1) / kTopWarps3, 4096));
    ++ctx->launches, k3_score_top<<<dim3(top_blocks, nj), kTopWarps3 * kWarp, 0, st>>>(b->jobs, b->cells, b->top_scores,
                                                                       b->state, b->scan_best);
    mark();
    ++ctx->launches, k3_dive<<<dim3((b->max_scans + kDiveWarps3 - 1) / kDiveWarps3, nj), kDiveWarps3 * kWarp, 0, st>>>(
        b->jobs, b->cells, b->low_points, b->scan_poses, b->scan_leaf_q, b->scan_best, b->state);
    mark();
    const int seed_blocks = static_cast<int>(std::min<long long>((max_cand + 255) / 256, 1024));
    const Deferred df{b->deferred, b->deferred_cap, b->counters + kCntDeferred};
    for (int d = b->max_depth - 1; d >= 0; --d) {
      bool any = false;
      for (const Job3Dev& j : b->jobs_host) any = any || (j.depth - 1 == d && j.S > 0);
      if (any)
        ++ctx->launches, k3_seed<<<dim3(seed_blocks, nj), 256, 0, st>>>(b->jobs, b->top_scores, b->state, fbuf[d & 1],
                                                         b->counters, d, fcap, df);
      if (d >= 1)
        ++ctx->launches, k3_expand<<<wide, kExpandWarps3 * kWarp, 0, st>>>(b->jobs, b->cells, b->state, fbuf[d & 1],
                                                            fbuf[(d - 1) & 1], b->counters, d, fcap, df, false);
    }
    ++ctx->launches, k3_filter<<<wide, kExpandWarps3 * kWarp, 0, st>>>(b->jobs, b->low_points, b->scan_poses,
                                                        b->scan_leaf_q, b->state, fbuf[0], b->counters, fcap,
                                                        b->winners, b->winner_cap);
    // Second sweep over the nodes deferred at bound == S* (usually none: small grids).
    CSM_CUDA(cudaMemsetAsync(b->counters, 0, sizeof(unsigned int) * CSM_MAX_DEPTH, st));
    const int small = std::max(1, ctx->num_sms);
    for (int d = b->max_depth - 1; d >= 0; --d) {
      ++ctx->launches, k3_seed_deferred<<<32, 256, 0, st>>>(b->jobs, b->state, b->deferred, b->deferred_cap,
                                             fbuf[d & 1], b->counters, d, fcap);
      if (d >= 1)
        ++ctx->launches, k3_expand<<<small, kExpandWarps3 * kWarp, 0, st>>>(b->jobs, b->cells, b->state, fbuf[d & 1],
                                                             fbuf[(d - 1) & 1], b->counters, d, fcap, df, true);
    }
    ++ctx->launches, k3_filter<<<small, kExpandWarps3 * kWarp, 0, st>>>(b->jobs, b->low_points, b->scan_poses,
                                                         b->scan_leaf_q, b->state, fbuf[0], b->counters, fcap,
                                                         b->winners, b->winner_cap);
    ++ctx->launches, k3_pick_winner<<<1, 1024, 0, st>>>(b->jobs, b->top_scores, b->state, b->winners, b->counters,
                                         b->winner_cap);
    mark();
    ++ctx->launches, k3_collect<<<dim3(seed_blocks, nj), 256, 0, st>>>(b->jobs, b->top_scores, b->state, b->contenders,
                                                        b->counters, b->contender_cap);
    ++ctx->launches, k3_ordered_dfs<<<wide, 128, 0, st>>>(b->jobs, b->cells, b->low_points, b->scan_poses,
                                           b->scan_leaf_q, b->state, b->contenders, b->counters,
                                           b->contender_cap);
    ++ctx->launches, k3_pick<<<(b->contender_cap + 255) / 256, 256, 0, st>>>(b->state, b->contenders, b->counters,
                                                              b->contender_cap);
  } else {
    for (int k = 0; k < 4; ++k) mark();
  }
  ++ctx->launches, k3_finalize<<<(nj + 127) / 128, 128, 0, st>>>(b->jobs, b->state, b->contenders, b->winners,
                                                 b->results_dev, nj);
  mark();
  CSM_CUDA(cudaGetLastError());
  b->ran = true;
  return CSM_OK;
}

}  // extern "C"

namespace {

struct ScoreIdx3 {
  float score;
  unsigned int idx;
  bool operator>(const ScoreIdx3& o) const { return score > o.score; }
};

// Equal lowest-resolution scores: the reference's visiting order is whatever its
// std::sort (FCSM3D.cc:350-351) makes of the generation-ordered list; replay it.
int resolve_tie3(csm_ctx* ctx, csm_batch3d* b, int job, const std::vector<Contender3>& list,
                 csm_result3d* r) {
  const Job3Dev& j = b->jobs_host[job];
  Job3State st;
  CSM_CUDA(cudaMemcpyAsync(&st, b->state + job, sizeof(st), cudaMemcpyDeviceToHost, ctx->stream));
  CSM_CUDA(cudaStreamSynchronize(ctx->stream));
  std::vector<float> scores(st.total);
  CSM_CUDA(cudaMemcpyAsync(scores.data(), b->top_scores + j.cand_off, sizeof(float) * st.total,
                           cudaMemcpyDeviceToHost, ctx->stream));
  CSM_CUDA(cudaStreamSynchronize(ctx->stream));
  std::vector<ScoreIdx3> v(st.total);
  for (unsigned int i = 0; i < st.total; ++i) v[i] = ScoreIdx3{scores[i], i};
  std::sort(v.begin(), v.end(), std::greater<ScoreIdx3>());
  std::vector<unsigned int> rank(st.total);
  for (unsigned int i = 0; i < st.total; ++i) rank[v[i].idx] = i;
  const Contender3* best = nullptr;
  for (const Contender3& c : list) {
    if (c.job != job || !c.has_leaf || float_to_ordered(c.score) != st.group_best) continue;
    if (!best || rank[c.cand] < rank[best->cand]) best = &c;
  }
  if (!best) {
    set_error("tie replay found no contender for 3D job %d", job);
    return CSM_ERR_CUDA;
  }
  r->scan_index = best->scan;
  for (int a = 0; a < 3; ++a) r->offset[a] = best->leaf[a];
  r->low_resolution_score = best->leaf_low;
  r->tie_resolution = 1;
  return CSM_OK;
}

// The same over the leaves the sweep recorded: best-ranked candidate, then smallest path.
int resolve_tie3_winners(csm_ctx* ctx, csm_batch3d* b, int job, const std::vector<Winner3>& list,
                         csm_result3d* r) {
  const Job3Dev& j = b->jobs_host[job];
  Job3State st;
  CSM_CUDA(cudaMemcpyAsync(&st, b->state + job, sizeof(st), cudaMemcpyDeviceToHost, ctx->stream));
  CSM_CUDA(cudaStreamSynchronize(ctx->stream));
  std::vector<float> scores(st.total);
  CSM_CUDA(cudaMemcpyAsync(scores.data(), b->top_scores + j.cand_off, sizeof(float) * st.total,
                           cudaMemcpyDeviceToHost, ctx->stream));
  CSM_CUDA(cudaStreamSynchronize(ctx->stream));
  std::vector<ScoreIdx3> v(st.total);
  for (unsigned int i = 0; i < st.total; ++i) v[i] = ScoreIdx3{scores[i], i};
  std::sort(v.begin(), v.end(), std::greater<ScoreIdx3>());
  std::vector<unsigned int> rank(st.total);
  for (unsigned int i = 0; i < st.total; ++i) rank[v[i].idx] = i;
  const Winner3* best = nullptr;
  for (const Winner3& w : list) {
    if (w.job != job || float_to_ordered(w.score) != st.best ||
        float_to_ordered(scores[w.cand]) != st.w_group)
      continue;
    if (!best || rank[w.cand] < rank[best->cand] || (w.cand == best->cand && w.path < best->path)) best = &w;
  }
  if (!best) {
    set_error("tie replay found no recorded leaf for 3D job %d", job);
    return CSM_ERR_CUDA;
  }
  r->scan_index = best->scan;
  for (int a = 0; a < 3; ++a) r->offset[a] = best->leaf[a];
  r->low_resolution_score = best->low;
  r->tie_resolution = 2;
  return CSM_OK;
}

}  // namespace

extern "C" {

int csm_batch3d_fetch(csm_ctx* ctx, csm_batch3d* b, csm_result3d* results) {
  CSM_REQUIRE(ctx != nullptr && b != nullptr && results != nullptr, "NULL argument");
  CSM_REQUIRE(b->ran, "batch was not run");
  CSM_CUDA(cudaSetDevice(ctx->device));
  csm_result3d* pin_res = reinterpret_cast<csm_result3d*>(static_cast<char*>(b->pinned) + b->pin_results_off);
  unsigned int counters[kNumCounters3];
  for (int attempt = 0;; ++attempt) {
    CSM_CUDA(cudaMemcpyAsync(pin_res, b->results_dev, sizeof(csm_result3d) * b->num_jobs,
                             cudaMemcpyDeviceToHost, ctx->stream));
    CSM_CUDA(cudaMemcpyAsync(counters, b->counters, sizeof(counters), cudaMemcpyDeviceToHost, ctx->stream));
    CSM_CUDA(cudaStreamSynchronize(ctx->stream));
    const unsigned int ov = counters[kCntOverflow];
    if (ov == 0) break;
    if (attempt >= 4) {
      set_error("3D work list overflow (code %u) after %d regrowths", ov, attempt);
      return CSM_ERR_CAPACITY;
    }
    if (ov & 1u) {
      const unsigned long long have =
          std::min(ctx->frontier[0].cap, ctx->frontier[1].cap) / sizeof(Node3S);
      b->frontier_want = std::max(b->frontier_want, have) * 4ull;
    }
    unsigned long long ccap = (ov & 4u) ? static_cast<unsigned long long>(counters[kCntContenders]) + 1024ull
                                        : b->contender_cap;
    if (ccap > 0xFFFFFFF0ull) ccap = 0xFFFFFFF0ull;
    unsigned long long wcap = (ov & 8u) ? static_cast<unsigned long long>(counters[kCntWinners]) * 2ull + 1024ull
                                        : b->winner_cap;
    if (wcap > 0xFFFFFFF0ull) wcap = 0xFFFFFFF0ull;
    int rc = CSM_OK;
    if (ccap != b->contender_cap || wcap != b->winner_cap) {
      rc = allocate_batch3(b, static_cast<unsigned int>(ccap), static_cast<unsigned int>(wcap));
      if (rc == CSM_OK) rc = copy_inputs3(ctx, b);
    }
    if (rc == CSM_OK) rc = csm_batch3d_run(ctx, b);
    if (rc != CSM_OK) return rc;
  }
  if (ctx->profiling) {
    if (static_cast<int>(b->prof_events.size()) == csm_ctx::kPhases + 1) {
      const cudaEvent_t* ev = b->prof_events.data();
      for (int k = 0; k < csm_ctx::kPhases; ++k) {
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, ev[k], ev[k + 1]) == cudaSuccess) ctx->phase_ms[k] += ms;
      }
      for (int i = 0; i < b->num_jobs; ++i)
        ctx->phase_scored[3] += static_cast<double>(pin_res[i].candidates_scored);
      ++ctx->profiled_runs;
    }
    for (cudaEvent_t e : b->prof_events) cudaEventDestroy(e);
    b->prof_events.clear();
  }
  std::vector<Contender3> list;
  std::vector<Winner3> wlist;
  bool have_list = false, have_wlist = false;
  for (int i = 0; i < b->num_jobs; ++i) {
    csm_result3d r = pin_res[i];
    if (r.found < 0) {
      set_error("3D job %d: the best leaf was not recorded (internal error)", i);
      return CSM_ERR_CUDA;
    }
    if (r.found == 1 && r.tie_resolution == 2) {
      if (!have_list) {
        const unsigned int n = std::min(counters[kCntContenders], b->contender_cap);
        list.resize(n);
        CSM_CUDA(cudaMemcpyAsync(list.data(), b->contenders, sizeof(Contender3) * n, cudaMemcpyDeviceToHost,
                                 ctx->stream));
        CSM_CUDA(cudaStreamSynchronize(ctx->stream));
        have_list = true;
      }
      const int rc = resolve_tie3(ctx, b, i, list, &r);
      if (rc != CSM_OK) return rc;
    } else if (r.found == 1 && r.tie_resolution == 3) {
      if (!have_wlist) {
        const unsigned int n = std::min(counters[kCntWinners], b->winner_cap);
        wlist.resize(n);
        CSM_CUDA(cudaMemcpyAsync(wlist.data(), b->winners, sizeof(Winner3) * n, cudaMemcpyDeviceToHost,
                                 ctx->stream));
        CSM_CUDA(cudaStreamSynchronize(ctx->stream));
        have_wlist = true;
      }
      const int rc = resolve_tie3_winners(ctx, b, i, wlist, &r);
      if (rc != CSM_OK) return rc;
    } else {
      r.tie_resolution = 0;
    }
    results[i] = r;
  }
  return CSM_OK;
}

int csm_match3d_batch(csm_ctx* ctx, const csm_job3d* jobs, int num_jobs, csm_result3d* results) {
  CSM_REQUIRE(ctx != nullptr && jobs != nullptr && results != nullptr, "NULL argument");
  CSM_REQUIRE(num_jobs >= 0, "num_jobs must be >= 0");
  // Bound the device footprint: chunks of at most 1024 jobs and 2^27 discrete cells.
  for (int at = 0; at < num_jobs;) {
    int n = 0;
    long long cells = 0;
    while (at + n < num_jobs && n < 1024) {
      const long long c = static_cast<long long>(std::max(jobs[at + n].num_scans, 0)) *
                          std::max(jobs[at + n].num_points, 0);
      if (n > 0 && cells + c > (1ll << 27)) break;
      cells += c;
      ++n;
    }
    csm_batch3d* b = nullptr;
    int rc = csm_batch3d_upload(ctx, jobs + at, n, &b);
    if (rc == CSM_OK) rc = csm_batch3d_run(ctx, b);
    if (rc == CSM_OK) rc = csm_batch3d_fetch(ctx, b, results + at);
    csm_batch3d_destroy(ctx, b);
    if (rc != CSM_OK) return rc;
    at += n;
  }
  return CSM_OK;
}

int csm_discretize3d(csm_ctx* ctx, const csm_job3d* job, int32_t* out_xyz) {
  CSM_REQUIRE(ctx != nullptr && job != nullptr && out_xyz != nullptr, "NULL argument");
  csm_batch3d* b = nullptr;
  int rc = csm_batch3d_upload(ctx, job, 1, &b);
  if (rc != CSM_OK) return rc;
  const Job3Dev& j = b->jobs_host[0];
  if (j.S > 0) {
    ++ctx->launches, k3_discretize<<<dim3(j.S, 1), 128, 0, ctx->stream>>>(b->jobs, b->points, b->scan_poses, b->cells,
                                                         b->scan_best);
    std::vector<int4> host(static_cast<size_t>(j.S) * j.n);
    if (cudaMemcpyAsync(host.data(), b->cells, sizeof(int4) * host.size(), cudaMemcpyDeviceToHost, ctx->stream) !=
            cudaSuccess ||
        cudaStreamSynchronize(ctx->stream) != cudaSuccess) {
      set_error("csm_discretize3d: %s", cudaGetErrorString(cudaGetLastError()));
      rc = CSM_ERR_CUDA;
    } else {
      for (size_t i = 0; i < host.size(); ++i) {
        out_xyz[3 * i] = host[i].x;
        out_xyz[3 * i + 1] = host[i].y;
        out_xyz[3 * i + 2] = host[i].z;
      }
    }
  }
  csm_batch3d_destroy(ctx, b);
  return rc;
}

}  // extern "C"
