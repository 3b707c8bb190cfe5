// Batched FastCorrelativeScanMatcher2D::MatchWithSearchParameters on the device.
//
// Replaces, for a whole batch of (node, submap) searches at once,
// /root/reference/cartographer/mapping/internal/2d/scan_matching/
//   correlative_scan_matcher_2d.cc:73-127  (ShrinkToFit, GenerateRotatedScans, DiscretizeScans)
//   fast_correlative_scan_matcher_2d.cc:264-378 (lowest-resolution candidates,
//                                                ScoreCandidates, BranchAndBound)
//
// Reformulation.  The reference walks the branch-and-bound tree depth first, one
// search per CPU thread.  Here the tree is swept breadth first, level by level, for
// all searches of the batch together, and the unit of work is a TILE: kTW x kTH
// lattice-adjacent parents of one (job, scan) with a 32-bit alive mask, one per
// warp, lane = parent.  The precomputation stack is stored in the quad-phase layout
// (csm_internal.cuh): one 32-bit word holds the four child values of a parent, and
// parents that are neighbours on the search lattice sit in neighbouring words.  So
// for every scan point the warp issues ONE load instruction that touches kTH
// contiguous kTW-word segments and yields the 4 x 32 child values of the tile.
//   1. k_discretize     one block per (job, scan): rotate + discretise every point
//                       (fp32/fp64 op order of Eigen/MapLimits, no FMA contraction)
//                       and reduce the per-scan ShrinkToFit bounds.
//   2. k_scan_prefix    per job: lattice sizes and generation-order bases.
//   3. k_score_top      every lowest-resolution candidate, as children of virtual
//                       parents one level above the stack (all alive).
//   4. k_dive           one warp per job: greedy descent from the best bound to a
//                       leaf; a real leaf score is a valid pruning incumbent.
//   5. k_seed / k_expand  tiles of nodes with bound > max(min_score, incumbent);
//                       children that survive are re-tiled with warp votes; leaves
//                       feed a 64-bit atomicMax.  Strict pruning never discards a
//                       subtree holding a strictly better leaf, so the maximum leaf
//                       score S* equals the reference's.
//   6. phase 2          the reference returns the DFS-first leaf attaining S*, where
//                       DFS order = sorted order of siblings (stable for <= 4
//                       children, libstdc++ introsort order at the top level).  Only
//                       top-level candidates whose score is >= that of a known
//                       S*-leaf's ancestor can precede it: k_collect lists them,
//                       k_ordered_dfs replays the reference's sibling order inside
//                       each (one warp per contender, only nodes with bound >= S*),
//                       k_pick chooses the highest-scoring contender holding an
//                       S*-leaf.  If several tie at the top level, the host replays
//                       std::sort on the top-level scores (csm_batch2d_fetch).
// Scores are integer sums pushed through the reference's three fp32 ops, so every
// comparison above is bit-identical to the CPU's.
#include <algorithm>
#include <cstring>
#include <functional>

#include "csm_match2d.cuh"

namespace csm {

constexpr int kWarp = 32;
constexpr unsigned kFull = 0xFFFFFFFFu;
constexpr int kCntContenders = 32;
constexpr int kCntOverflow = 33;
constexpr int kCntLeaves = 35;        // path mode: leaves recorded
constexpr int kCntDeferred = 36;      // path mode: nodes set aside (bound == incumbent)
constexpr int kCntDeferCursor = 37;   // path mode: next deferred node to walk
constexpr int kNumCounters = 112;

// Tile shape: kTW x kTH parents, lane = lj * kTW + li.
#ifndef CSM_LOG_TW
#define CSM_LOG_TW 3
#endif
constexpr int kLogTW = CSM_LOG_TW, kLogTH = 5 - CSM_LOG_TW;
constexpr int kTW = 1 << kLogTW, kTH = 1 << kLogTH;
static_assert(kLogTW >= 1 && kLogTW <= 5, "a tile is one warp");

// ---------------------------------------------------------------------------
// device helpers
// ---------------------------------------------------------------------------
// PrecomputationGrid2D::ToScore(sum / float(n))  (FCSM2D.cc:326-329, FCSM2D.h:74-76)
__device__ __forceinline__ float to_score(const JobDev& j, int sum) {
  const float mean = __fdiv_rn(static_cast<float>(sum), j.n_float);
  return __fadd_rn(j.score_offset, __fmul_rn(mean, j.score_scale));
}

// One quad word of level `lvl` for the window start (x, y): the four child values of
// a parent at (x, y) whose children live on level `lvl` (see StackDev).  Zero outside.
__device__ __forceinline__ uint32_t quad_at(const StackDev& sd, int lvl, int x, int y) {
  const int sh = lvl + 1, two_s = 2 << lvl;
  const int u = x + two_s - 1, v = y + two_s - 1;
  const int i = u >> sh, jj = v >> sh;  // arithmetic shift = floor
  const int in = sd.in[lvl], jn = sd.jn[lvl];
  if (static_cast<unsigned>(i) >= static_cast<unsigned>(in) ||
      static_cast<unsigned>(jj) >= static_cast<unsigned>(jn))
    return 0u;
  const int pu = u & (two_s - 1), pv = v & (two_s - 1);
  const size_t idx = (static_cast<size_t>(pv * two_s + pu) * jn + jj) * in + i;
  return __ldg(sd.quad[lvl] + idx);
}

__device__ __forceinline__ int warp_sum(int v) { return __reduce_add_sync(kFull, v); }

// Scores the <= 4 children of ONE node (scan, x, y) at depth `d` (children live on
// level d-1, half = 2^(d-1)); whole warp cooperates, lanes stride the points.  Used by
// the greedy dive and the ordered DFS of phase 2, which follow single paths.
// Children outside the search window are not generated by the reference
// (FCSM2D.cc:353-367); their slot gets valid = false.
struct Children {
  float score[4];
  bool valid[4];
};
__device__ __forceinline__ Children score_children(const JobDev& j, const int2* __restrict__ xy,
                                                   const ScanInfo& si, int scan, int x, int y, int d,
                                                   int lane) {
  const int lvl = d - 1;
  const int half = 1 << lvl;
  const int2* pts = xy + j.xy_off + static_cast<long long>(scan) * j.n;
  const StackDev& sd = *j.sd;
  // Packed accumulation: two 16-bit lanes per register (bytes 0/2 and 1/3), flushed
  // before they can overflow (255 * 256 < 65536).
  int s0 = 0, s1 = 0, s2 = 0, s3 = 0;
  uint32_t lo = 0, hi = 0;
  int since = 0;
  // kPathPts points per lane and pass: the dependent load pairs (cell, then quad word) of a
  // pass are issued together, so a single-path visit waits for one L2 round trip per pass
  // instead of one per point (these visits are latency-bound: one warp follows one path).
  constexpr int kPathPts = 4;  // 8 was slower (phase 2 3.9 vs 2.9 ms: registers)
  for (int p0 = lane; p0 < j.n; p0 += kPathPts * kWarp) {
    int2 c[kPathPts];
#pragma unroll
    for (int k = 0; k < kPathPts; ++k) c[k] = p0 + k * kWarp < j.n ? pts[p0 + k * kWarp] : make_int2(0, 0);
    uint32_t w[kPathPts];
#pragma unroll
    for (int k = 0; k < kPathPts; ++k)
      w[k] = p0 + k * kWarp < j.n ? quad_at(sd, lvl, c[k].x + x, c[k].y + y) : 0u;
#pragma unroll
    for (int k = 0; k < kPathPts; ++k) {
      lo += w[k] & 0x00FF00FFu;
      hi += (w[k] >> 8) & 0x00FF00FFu;
    }
    since += kPathPts;
    if (since >= 252) {  // flush before a 16-bit field can overflow (255 * 256 < 65536)
      s0 += lo & 0xFFFF; s2 += lo >> 16; s1 += hi & 0xFFFF; s3 += hi >> 16;
      lo = hi = 0;
      since = 0;
    }
  }
  s0 += lo & 0xFFFF; s2 += lo >> 16; s1 += hi & 0xFFFF; s3 += hi >> 16;
  s0 = warp_sum(s0);
  s1 = warp_sum(s1);
  s2 = warp_sum(s2);
  s3 = warp_sum(s3);
  Children c;
  const bool xok = x + half <= si.max_x, yok = y + half <= si.max_y;
  // generation order: x offset outer {0, half}, y offset inner {0, half}
  c.valid[0] = true;
  c.valid[1] = yok;
  c.valid[2] = xok;
  c.valid[3] = xok && yok;
  c.score[0] = to_score(j, s0);
  c.score[1] = to_score(j, s1);
  c.score[2] = to_score(j, s2);
  c.score[3] = to_score(j, s3);
  return c;
}

// The dense kernel core.  The warp holds one tile of parents whose children live on
// level `lvl` (children lattice step s = 2^lvl, parents 2s apart).  Parent (li, lj)
// of the tile has the window start (X0 + 2s*li, Y0 + 2s*lj); bx = X0 + 2s - 1 and
// by = Y0 + 2s - 1 are passed in.  For a scan point c the phase of the window start
// is the same for every parent of the tile, so the words the 32 lanes need are
// quad[idx0 + lj*in + li]: kTH contiguous runs of kTW words.
// The per-point address terms are computed once per 32 points (lane = point) and
// broadcast through a 512-byte shared staging buffer private to the warp.
// Result: sum[k] = sum over points of child k's value, k = 2*cx + cy.
__device__ __forceinline__ uint32_t ldg_pred(const uint32_t* base, int word_offset, uint32_t on) {
  // Predicated 32-bit read-only load: no branch, no memory access when `on` is 0.
  uint32_t w;
  asm("{\n\t"
      ".reg .pred p;\n\t"
      ".reg .s64 a;\n\t"
      "setp.ne.u32 p, %3, 0;\n\t"
      "mad.wide.s32 a, %2, 4, %1;\n\t"
      "mov.u32 %0, 0;\n\t"
      "@p ld.global.nc.u32 %0, [a];\n\t"
      "}"
      : "=r"(w)
      : "l"(base), "r"(word_offset), "r"(on));
  return w;
}

__device__ __forceinline__ void tile_sums(const StackDev& sd, const int2* __restrict__ pts, int n,
                                          int lvl, int bx, int by, uint32_t alive_mask, int lane,
                                          int2* __restrict__ stage, int (&sum)[4]) {
  const int sh = lvl + 1, two_s = 2 << lvl, msk = two_s - 1;
  const int in = sd.in[lvl], jn = sd.jn[lvl];
  const uint32_t* __restrict__ quad = sd.quad[lvl];
  const int li = lane & (kTW - 1), lj = lane >> kLogTW;
  const uint32_t* __restrict__ qlane = quad + (lj * in + li);
  const uint32_t lane_bit = 1u << lane;
  int s0 = 0, s1 = 0, s2 = 0, s3 = 0;
  uint32_t lo = 0, hi = 0;
  int chunks = 0;
  for (int p0 = 0; p0 < n; p0 += kWarp) {
    // Lane = point: the word index of the tile's first parent for this point, and
    // the mask of tile lanes whose word lies inside the plane (others read as 0).
    int2 e = make_int2(0, 0);
    if (p0 + lane < n) {
      const int2 c = pts[p0 + lane];
      const int u = c.x + bx, v = c.y + by;
      const int ca = u >> sh, ra = v >> sh;  // arithmetic shift = floor
      e.x = ((((v & msk) << sh) + (u & msk)) * jn + ra) * in + ca;
      const int clo = max(0, -ca), chi = min(kTW, in - ca);
      uint32_t vmask = 0u;
      if (chi > clo) {
        const uint32_t cols = (chi >= 32 ? 0xFFFFFFFFu : (1u << chi) - 1u) & ~((1u << clo) - 1u);
#pragma unroll
        for (int r = 0; r < kTH; ++r)
          if (static_cast<unsigned>(ra + r) < static_cast<unsigned>(jn)) vmask |= cols << (r * kTW);
      }
      e.y = static_cast<int>(vmask & alive_mask);
    }
    __syncwarp();
    stage[lane] = e;
    __syncwarp();
    const int cnt = min(kWarp, n - p0);
    if (cnt == kWarp) {
#pragma unroll
      for (int k = 0; k < kWarp; ++k) {
        const int2 t = stage[k];
        const uint32_t w = ldg_pred(qlane, t.x, static_cast<uint32_t>(t.y) & lane_bit);
        lo += w & 0x00FF00FFu;
        hi += __byte_perm(w, 0u, 0x4341);  // (w >> 8) & 0x00FF00FF
      }
    } else {
#pragma unroll 4
      for (int k = 0; k < cnt; ++k) {
        const int2 t = stage[k];
        const uint32_t w = ldg_pred(qlane, t.x, static_cast<uint32_t>(t.y) & lane_bit);
        lo += w & 0x00FF00FFu;
        hi += __byte_perm(w, 0u, 0x4341);
      }
    }
    if (++chunks == 8) {  // 256 points: flush before a 16-bit field can overflow
      s0 += lo & 0xFFFF; s2 += lo >> 16; s1 += hi & 0xFFFF; s3 += hi >> 16;
      lo = hi = 0;
      chunks = 0;
    }
  }
  s0 += lo & 0xFFFF; s2 += lo >> 16; s1 += hi & 0xFFFF; s3 += hi >> 16;
  sum[0] = s0;
  sum[1] = s1;
  sum[2] = s2;
  sum[3] = s3;
}

// ---------------------------------------------------------------------------
// 1. rotate + discretise + ShrinkToFit bounds.  grid = (max_scans, num_jobs)
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(128)
k_discretize(const JobDev* __restrict__ jobs, const float* __restrict__ points,
             const float* __restrict__ quats, int2* __restrict__ xy, ScanInfo* __restrict__ scans,
             unsigned long long* __restrict__ scan_best, unsigned int* __restrict__ bucket_js,
             int shrink) {
  const JobDev& j = jobs[blockIdx.y];
  const int scan = blockIdx.x;
  if (scan >= j.S) return;
  if (bucket_js != nullptr && threadIdx.x == 0)
    bucket_js[j.scan_off + scan] = (blockIdx.y << 16) | static_cast<unsigned int>(scan);
  const float qw = quats[2 * (j.quat_off + scan)], qs = quats[2 * (j.quat_off + scan) + 1];
  int mnx = 0, mny = 0, mxx = 0, mxy = 0;  // Array2i::Zero() seeds (CSM2D.cc:78-79)
  for (int p = threadIdx.x; p < j.n; p += blockDim.x) {
    float vx = points[3 * (j.pts_off + p)], vy = points[3 * (j.pts_off + p) + 1];
    yaw_rotate(j.q0w, j.q0s, vx, vy);  // initial yaw (FCSM2D.cc:235-239)
    yaw_rotate(qw, qs, vx, vy);        // scan rotation (CSM2D.cc:104-106)
    const float wx = __fadd_rn(j.tx, vx), wy = __fadd_rn(j.ty, vy);  // CSM2D.cc:120-121
    int cx, cy;
    cell_index(j.max_x, j.max_y, j.res, wx, wy, cx, cy);
    xy[j.xy_off + static_cast<long long>(scan) * j.n + p] = make_int2(cx, cy);
    mnx = min(mnx, -cx);
    mny = min(mny, -cy);
    mxx = max(mxx, j.num_x - 1 - cx);
    mxy = max(mxy, j.num_y - 1 - cy);
  }
  __shared__ int red[4][4];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    mnx = min(mnx, __shfl_xor_sync(kFull, mnx, o));
    mny = min(mny, __shfl_xor_sync(kFull, mny, o));
    mxx = max(mxx, __shfl_xor_sync(kFull, mxx, o));
    mxy = max(mxy, __shfl_xor_sync(kFull, mxy, o));
  }
  const int warp = threadIdx.x >> 5;
  if ((threadIdx.x & 31) == 0) {
    red[warp][0] = mnx;
    red[warp][1] = mny;
    red[warp][2] = mxx;
    red[warp][3] = mxy;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < 4; ++w) {
      mnx = min(mnx, red[w][0]);
      mny = min(mny, red[w][1]);
      mxx = max(mxx, red[w][2]);
      mxy = max(mxy, red[w][3]);
    }
    ScanInfo si;
    si.min_x = -j.lin;
    si.max_x = j.lin;
    si.min_y = -j.lin;
    si.max_y = j.lin;
    if (shrink) {  // CSM2D.cc:86-89
      si.min_x = max(si.min_x, mnx);
      si.max_x = min(si.max_x, mxx);
      si.min_y = max(si.min_y, mny);
      si.max_y = min(si.max_y, mxy);
    }
    // FCSM2D.cc:275-283.  An inverted window yields no candidate for this scan.
    const int step = 1 << (j.depth - 1);
    si.nx = si.max_x >= si.min_x ? (si.max_x - si.min_x + step) / step : 0;
    si.ny = si.max_y >= si.min_y ? (si.max_y - si.min_y + step) / step : 0;
    if (si.nx == 0 || si.ny == 0) si.nx = si.ny = 0;
    si.base = 0;
    si.pad = 0;
    scans[j.scan_off + scan] = si;
    if (scan_best != nullptr) scan_best[j.scan_off + scan] = 0ull;
  }
}

// 2. generation-order base of every scan's lattice.  One block per job.
__global__ void __launch_bounds__(256)
k_scan_prefix(const JobDev* __restrict__ jobs, ScanInfo* __restrict__ scans,
              JobState* __restrict__ state, unsigned int* __restrict__ counters) {
  const JobDev& j = jobs[blockIdx.x];
  __shared__ int carry;
  __shared__ int part[256];
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (int s0 = 0; s0 < j.S; s0 += 256) {
    const int s = s0 + threadIdx.x;
    const int v = s < j.S ? scans[j.scan_off + s].nx * scans[j.scan_off + s].ny : 0;
    part[threadIdx.x] = v;
    __syncthreads();
    for (int o = 1; o < 256; o <<= 1) {  // Hillis-Steele inclusive scan
      const int t = threadIdx.x >= o ? part[threadIdx.x - o] : 0;
      __syncthreads();
      part[threadIdx.x] += t;
      __syncthreads();
    }
    if (s < j.S) scans[j.scan_off + s].base = carry + part[threadIdx.x] - v;
    __syncthreads();
    if (threadIdx.x == 255) carry += part[255];
    __syncthreads();
  }
  // A lattice larger than the host's bound would make k_score_top write past this job's
  // slice of top_scores: empty the job instead (the overflow flag fails the fetch).
  const bool too_big = static_cast<long long>(carry) > j.cand_cap;
  if (too_big)
    for (int s = threadIdx.x; s < j.S; s += blockDim.x) scans[j.scan_off + s].nx = scans[j.scan_off + s].ny = 0;
  if (threadIdx.x == 0) {
    JobState st;
    st.best = 0ull;
    st.top_best = 0ull;
    st.win = ~0ull;
    st.group_best = 0u;
    st.tie_count = 0u;
    st.total = too_big ? 0u : static_cast<unsigned int>(carry);
    st.cand_max = 0u;
    st.scored = 0ull;
    for (int k = 0; k < 4; ++k) st.scored_phase[k] = 0ull;
    state[blockIdx.x] = st;
    if (static_cast<long long>(carry) > j.cand_cap) atomicOr(&counters[kCntOverflow], 2u);
  }
}

// Decodes a generation index into (scan, x, y) by binary search over the bases.
__device__ inline void decode_candidate(const JobDev& j, const ScanInfo* __restrict__ scans,
                                        unsigned int cand, int& scan, int& x, int& y) {
  int lo = 0, hi = j.S - 1;
  while (lo < hi) {  // last scan with base <= cand and a non-empty lattice
    const int mid = (lo + hi + 1) >> 1;
    if (static_cast<unsigned int>(scans[j.scan_off + mid].base) <= cand) lo = mid; else hi = mid - 1;
  }
  while (lo > 0 && scans[j.scan_off + lo].nx * scans[j.scan_off + lo].ny == 0) --lo;
  const ScanInfo si = scans[j.scan_off + lo];
  const int c = static_cast<int>(cand) - si.base;
  const int step = 1 << (j.depth - 1);
  const int ix = c / si.ny, iy = c - ix * si.ny;
  scan = lo;
  x = si.min_x + ix * step;
  y = si.min_y + iy * step;
}

__device__ __forceinline__ unsigned long long leaf_key(float leaf_score, float top_score) {
  return (static_cast<unsigned long long>(float_to_ordered(leaf_score)) << 32) |
         float_to_ordered(top_score);
}

__device__ __forceinline__ void offer_leaf(JobState* st, float leaf_score, float top_score) {
  atomicMax(&st->best, leaf_key(leaf_score, top_score));
}

// Current pruning threshold of a job: max(min_score, best leaf so far).
__device__ __forceinline__ float threshold(const JobDev& j, const JobState* st) {
  const unsigned long long b = *reinterpret_cast<const volatile unsigned long long*>(&st->best);
  const unsigned int hi = static_cast<unsigned int>(b >> 32);
  if (hi == 0u) return j.min_score;
  return fmaxf(j.min_score, ordered_to_float(hi));
}

// Snapshot of a job's incumbent: the best real leaf score so far (has == false: none yet).
__device__ __forceinline__ float incumbent(const JobState* st, bool& has) {
  const unsigned long long b = *reinterpret_cast<const volatile unsigned long long*>(&st->best);
  const unsigned int hi = static_cast<unsigned int>(b >> 32);
  has = hi != 0u;
  return has ? ordered_to_float(hi) : 0.f;
}

// Path mode: appends one record to a bounded list (rare events: plain atomics).
template <typename T>
__device__ __forceinline__ void push_record(T* __restrict__ list, unsigned int* __restrict__ counter,
                                            unsigned int cap, unsigned int* __restrict__ overflow, const T& rec) {
  const unsigned int slot = atomicAdd(counter, 1u);
  if (slot < cap) list[slot] = rec; else atomicOr(overflow, 4u);
}

// ---------------------------------------------------------------------------
// 4. greedy dives: one warp per (job, scan), from the scan's best lowest-resolution
//    bound down to a leaf.  Any real leaf score above min_score is a valid
//    incumbent; one dive per angular slice makes it a strong one, which is what lets
//    the level-synchronous sweep prune as hard as the reference's depth-first search.
//    grid = (ceil(max_scans / 4), jobs)
// ---------------------------------------------------------------------------
constexpr int kDiveWarps = 4;
__global__ void __launch_bounds__(kDiveWarps * kWarp)
k_dive(const JobDev* __restrict__ jobs, const int2* __restrict__ xy,
       const ScanInfo* __restrict__ scans, const unsigned long long* __restrict__ scan_best,
       JobState* __restrict__ state) {
  const JobDev& j = jobs[blockIdx.y];
  const int scan = blockIdx.x * kDiveWarps + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (scan >= j.S) return;
  JobState* st = &state[blockIdx.y];
  const unsigned long long tb = scan_best[j.scan_off + scan];
  if (tb == 0ull) return;
  const unsigned int cand = 0xFFFFFFFFu - static_cast<unsigned int>(tb & 0xFFFFFFFFull);
  const float top_score = ordered_to_float(static_cast<unsigned int>(tb >> 32));
  if (!(top_score > j.min_score)) return;
  const ScanInfo si = scans[j.scan_off + scan];
  const int c = static_cast<int>(cand) - si.base;
  const int ix = c / si.ny, iy = c - ix * si.ny;
  int x = si.min_x + (ix << (j.depth - 1)), y = si.min_y + (iy << (j.depth - 1));
  float score = top_score;
  unsigned long long scored = 0;
  for (int d = j.depth - 1; d >= 1; --d) {
    // A bound that cannot beat the incumbent any more ends the dive early.
    float thr = threshold(j, st);
    thr = __shfl_sync(kFull, thr, 0);
    if (!(score > thr)) break;
    const Children ch = score_children(j, xy, si, scan, x, y, d, lane);
    const int half = 1 << (d - 1);
    int bi = 0;
    for (int k = 1; k < 4; ++k)
      if (ch.valid[k] && ch.score[k] > ch.score[bi]) bi = k;
    for (int k = 0; k < 4; ++k) scored += ch.valid[k] ? 1 : 0;
    x += (bi & 2) ? half : 0;
    y += (bi & 1) ? half : 0;
    score = ch.score[bi];
    if (d == 1 && lane == 0 && score > j.min_score) offer_leaf(st, score, top_score);
  }
  if (j.depth == 1 && lane == 0 && score > j.min_score) offer_leaf(st, score, top_score);
  if (lane == 0 && scored != 0ull) {
    atomicAdd(&st->scored, scored);
    atomicAdd(&st->scored_phase[1], scored);
  }
}

// Appends tiles with warp-aggregated atomics; sets the overflow flag when full.
__device__ __forceinline__ void push_tile(Tile* __restrict__ out, unsigned int* __restrict__ counter,
                                          unsigned long long cap, unsigned int* overflow, bool want,
                                          const Tile& tl) {
  const unsigned int mask = __ballot_sync(__activemask(), want);
  if (!want) return;
  const int leader = __ffs(mask) - 1;
  const int lane = threadIdx.x & 31;
  unsigned int base = 0;
  if (lane == leader) base = atomicAdd(counter, __popc(mask));
  base = __shfl_sync(mask, base, leader);
  const unsigned int slot = base + __popc(mask & ((1u << lane) - 1u));
  if (slot < cap) out[slot] = tl; else atomicOr(overflow, 1u);
}

// ---------------------------------------------------------------------------
// 5a. seed the frontier: tiles of lowest-resolution candidates above the threshold.
//     grid = (max_scans, jobs); only jobs whose top depth is `slot` take part.
// ---------------------------------------------------------------------------
constexpr int kSeedWarps = 4;
__global__ void __launch_bounds__(kSeedWarps * kWarp)
k_seed(const JobDev* __restrict__ jobs, const ScanInfo* __restrict__ scans,
       const float* __restrict__ top_scores, const JobState* __restrict__ state,
       Tile* __restrict__ out, unsigned int* __restrict__ counters, int slot,
       unsigned long long cap) {
  const JobDev& j = jobs[blockIdx.y];
  const int scan = blockIdx.x;
  if (scan >= j.S || j.depth - 1 != slot) return;
  const ScanInfo si = scans[j.scan_off + scan];
  if (si.nx == 0) return;
  const float thr = threshold(j, &state[blockIdx.y]);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int li = lane & (kTW - 1), lj = lane >> kLogTW;
  const int tiles_x = (si.nx + kTW - 1) >> kLogTW, tiles_y = (si.ny + kTH - 1) >> kLogTH;
  for (int t = warp; t < tiles_x * tiles_y; t += kSeedWarps) {
    const int ti = t % tiles_x, tj = t / tiles_x;
    const int ci = kTW * ti + li, cj = kTH * tj + lj;
    bool alive = false;
    if (ci < si.nx && cj < si.ny)
      alive = top_scores[j.cand_off + si.base + ci * si.ny + cj] > thr;
    Tile tl;
    tl.job_scan = (static_cast<int>(blockIdx.y) << 16) | scan;
    tl.tij = ti | (tj << 16);
    tl.mask = __ballot_sync(kFull, alive);
    tl.pad = 0;
    push_tile(out, &counters[slot], cap, &counters[kCntOverflow], lane == 0 && tl.mask != 0u, tl);
  }
}

// ---------------------------------------------------------------------------
// 5b. expand one level: one warp per tile of parents at depth d.
// ---------------------------------------------------------------------------
constexpr int kExpandWarps = 8;
__global__ void __launch_bounds__(kExpandWarps * kWarp)
k_expand(const JobDev* __restrict__ jobs, const int2* __restrict__ xy,
         const ScanInfo* __restrict__ scans, const float* __restrict__ top_scores,
         JobState* __restrict__ state, const Tile* __restrict__ in, Tile* __restrict__ out,
         unsigned int* __restrict__ counters, int in_slot, int out_slot, int d,
         unsigned long long cap, unsigned long long* __restrict__ hist) {
  __shared__ int2 stage_all[kExpandWarps][kWarp];
  const unsigned int count = min(static_cast<unsigned long long>(counters[in_slot]), cap);
  const int lane = threadIdx.x & 31;
  int2* stage = stage_all[threadIdx.x >> 5];
  const int li = lane & (kTW - 1), lj = lane >> kLogTW;
  const unsigned int warps = gridDim.x * kExpandWarps;
  const int lvl = d - 1;
  const int two_s = 2 << lvl;
  int acc_job = -1;
  unsigned long long acc_scored = 0ull;
  for (unsigned int i = blockIdx.x * kExpandWarps + (threadIdx.x >> 5); i < count; i += warps) {
    const Tile tl = in[i];
    const int job = static_cast<unsigned int>(tl.job_scan) >> 16, scan = tl.job_scan & 0xFFFF;
    const JobDev& j = jobs[job];
    JobState* st = &state[job];
    // The threshold moves under concurrent atomics: read once, keep the warp uniform.
    float thr = threshold(j, st);
    thr = __shfl_sync(kFull, thr, 0);
    const ScanInfo si = scans[j.scan_off + scan];
    const int ti = tl.tij & 0xFFFF, tj = static_cast<unsigned int>(tl.tij) >> 16;
    const bool alive = (tl.mask >> lane) & 1u;
    if (hist != nullptr && lane == 0) atomicAdd(&hist[d * 33 + __popc(tl.mask)], 1ull);
    int sum[4];
    tile_sums(*j.sd, xy + j.xy_off + static_cast<long long>(scan) * j.n, j.n, lvl,
              si.min_x + (two_s - 1) + two_s * kTW * ti, si.min_y + (two_s - 1) + two_s * kTH * tj,
              tl.mask, lane, stage, sum);
    // Children lattice of this scan on level lvl (FCSM2D.cc:353-367: a child beyond
    // max_x / max_y is not generated).
    const int nxl = ((si.max_x - si.min_x) >> lvl) + 1, nyl = ((si.max_y - si.min_y) >> lvl) + 1;
    const int ci0 = 2 * (kTW * ti + li), cj0 = 2 * (kTH * tj + lj);
    bool pass[4];
    float sc[4];
    int nvalid = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const bool valid = alive && ci0 + (k >> 1) < nxl && cj0 + (k & 1) < nyl;
      sc[k] = to_score(j, sum[k]);
      pass[k] = valid && sc[k] > thr;
      nvalid += valid ? 1 : 0;
    }
    const int tot = __reduce_add_sync(kFull, nvalid);
    if (job != acc_job) {
      if (lane == 0 && acc_job >= 0) {
        atomicAdd(&state[acc_job].scored, acc_scored);
        atomicAdd(&state[acc_job].scored_phase[2], acc_scored);
      }
      acc_job = job;
      acc_scored = 0ull;
    }
    acc_scored += static_cast<unsigned long long>(tot);
    if (lvl == 0) {
      // Leaves.  Rare: only a leaf above the current incumbent reaches the atomic.
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        if (!pass[k]) continue;
        const int ci = ci0 + (k >> 1), cj = cj0 + (k & 1);
        const int ix = ci >> (j.depth - 1), iy = cj >> (j.depth - 1);  // top-level ancestor
        const float ts = top_scores[j.cand_off + si.base + ix * si.ny + iy];
        const unsigned long long key = leaf_key(sc[k], ts);
        const unsigned long long cur =
            *reinterpret_cast<const volatile unsigned long long*>(&st->best);
        if (key > cur) atomicMax(&st->best, key);
      }
    } else {
      // Re-tile the surviving children: child (2*li + cx, 2*lj + cy) of the tile's
      // 2kTW x 2kTH child area goes to child tile (a, b) = (that index / kTW, / kTH).
      unsigned int mine = 0u;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        unsigned int m = 0u;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const int ci = 2 * li + (k >> 1), cj = 2 * lj + (k & 1);
          const int tile_of_k = (ci >> kLogTW) + 2 * (cj >> kLogTH);
          const int bit = ((cj & (kTH - 1)) << kLogTW) + (ci & (kTW - 1));
          if (tile_of_k == q && pass[k]) m |= 1u << bit;
        }
        const unsigned int mq = __reduce_or_sync(kFull, m);
        if (lane == q) mine = mq;
      }
      Tile ch;
      ch.job_scan = tl.job_scan;
      ch.tij = (2 * ti + (lane & 1)) | ((2 * tj + ((lane >> 1) & 1)) << 16);
      ch.mask = mine;
      ch.pad = 0;
      push_tile(out, &counters[out_slot], cap, &counters[kCntOverflow], lane < 4 && mine != 0u, ch);
    }
  }
  if (lane == 0 && acc_job >= 0) {
    atomicAdd(&state[acc_job].scored, acc_scored);
    atomicAdd(&state[acc_job].scored_phase[2], acc_scored);
  }
}

// ---------------------------------------------------------------------------
// 5'. Node-list sweep.  The tile sweep above keeps lane = lattice position, so a warp
//     with 4 alive nodes of 32 still issues every instruction for 32 (mean occupancy
//     on C3: 6.9 of 32).  Here the alive nodes of one (job, scan) -- a BUCKET -- are kept
//     contiguous in one list, 4 bytes per node (i | j << 16 on the depth's lattice), and
//     a warp takes 32 consecutive ones: lane = alive node.  For a scan point the word a
//     node needs is quad[P(point) + j*in + i] with P the same for every node of the
//     scan on that level (that is what the quad-phase layout buys), so the per-point
//     term is staged once per warp and the per-(node, point) work is one broadcast
//     shared-memory read, one address add, one load and the packed accumulate.
//     Points whose window lies inside the plane for EVERY node of the chunk (by the
//     chunk's bounding box) take the unchecked loop, points outside for every node are
//     dropped, the rest takes a bounds-checked loop.
//
//     Keeping buckets contiguous: k_expand_nodes / k_seed_nodes append their survivors
//     as BLOCKS to a scratch list (one atomic each on the scratch cursor, the bucket's
//     counter and the block list), k_bucket_prefix turns the bucket counters into list
//     offsets and k_place copies every block to its bucket's range and writes the work
//     items (32-node chunks) of the next level.  Children are emitted in (parent lane,
//     child) order, so lattice neighbours stay neighbours in the list.
// ---------------------------------------------------------------------------
struct NodeBlock {
  unsigned int src, count, bucket, rel;  // scratch position, nodes, bucket, offset inside it
};
constexpr int kCntNodes = 0;     // + depth: nodes on the sorted list
constexpr int kCntTmp = 48;      // + depth: scratch list fill
constexpr int kCntBlocks = 64;   // + depth: blocks appended
constexpr int kCntItems = 80;    // + depth: 32-node work items
constexpr int kCntCursor = 96;   // + depth: next work item to hand out (k_expand_nodes)
constexpr int kCntDfsCursor = 96;  // depth 0 is never swept: next contender (k_ordered_dfs)
constexpr int kNodeBlocksPerSm = 4;  // resident k_expand_nodes blocks per SM (register bound)
constexpr int kStagePoints = 128;

struct NodeStage {
  int in_p[kStagePoints];        // P of points inside for the whole chunk
  int4 mix[kStagePoints];        // (P, ca0, ra0, -) of points that need the bounds check
};

// Read-only 32-bit load that does not allocate an L1 line (CSM_EXPAND_LOAD=noalloc): the
// sweep re-uses almost nothing from L1 (ncu: 20-25 % sector hits, all of them inside one
// request), so the fills only compete with the loads for the L1 data pipe.
template <bool kNoAlloc>
__device__ __forceinline__ uint32_t ldg_w(const uint32_t* p) {
  if constexpr (kNoAlloc) {
    uint32_t w;
    asm("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(w) : "l"(p));
    return w;
  } else {
    return __ldg(p);
  }
}

// Sums of the four child values over all scan points for this lane's node (ni, nj).
// Lanes beyond the chunk's node count carry a copy of lane 0's node.
template <bool kNoAlloc = false>
__device__ __forceinline__ void node_sums(const StackDev& sd, const int2* __restrict__ pts, int n,
                                          int lvl, int x0, int y0, int ni, int nj, int lane,
                                          NodeStage* __restrict__ stage, int (&sum)[4]) {
  const int sh = lvl + 1, msk = (2 << lvl) - 1;
  const int in = sd.in[lvl], jn = sd.jn[lvl];
  const uint32_t* __restrict__ qb = sd.quad[lvl] + (static_cast<long long>(nj) * in + ni);
  const int bi0 = __reduce_min_sync(kFull, ni), bi1 = __reduce_max_sync(kFull, ni);
  const int bj0 = __reduce_min_sync(kFull, nj), bj1 = __reduce_max_sync(kFull, nj);
  const uint32_t lt = (1u << lane) - 1u;
  int s0 = 0, s1 = 0, s2 = 0, s3 = 0;
  for (int p0 = 0; p0 < n; p0 += kStagePoints) {
    int cin = 0, cmix = 0;
    __syncwarp();
#pragma unroll
    for (int r = 0; r < kStagePoints / kWarp; ++r) {
      const int p = p0 + r * kWarp + lane;
      bool inside = false, mixed = false;
      int P = 0, ca0 = 0, ra0 = 0;
      if (p < n) {
        const int2 c = pts[p];
        const int u0 = c.x + x0, v0 = c.y + y0;
        ca0 = u0 >> sh;  // arithmetic shift = floor
        ra0 = v0 >> sh;
        P = ((((v0 & msk) << sh) + (u0 & msk)) * jn + ra0) * in + ca0;
        const bool out = bi1 + ca0 < 0 || bi0 + ca0 >= in || bj1 + ra0 < 0 || bj0 + ra0 >= jn;
        inside = bi0 + ca0 >= 0 && bi1 + ca0 < in && bj0 + ra0 >= 0 && bj1 + ra0 < jn;
        mixed = !out && !inside;
      }
      const uint32_t m_in = __ballot_sync(kFull, inside), m_mix = __ballot_sync(kFull, mixed);
      if (inside) stage->in_p[cin + __popc(m_in & lt)] = P;
      if (mixed) stage->mix[cmix + __popc(m_mix & lt)] = make_int4(P, ca0, ra0, 0);
      cin += __popc(m_in);
      cmix += __popc(m_mix);
    }
    __syncwarp();
    uint32_t lo = 0, hi = 0;  // two 16-bit fields each; 128 * 255 < 65536
    int k = 0;
    for (; k + 8 <= cin; k += 8) {
      const int4 a = *reinterpret_cast<const int4*>(&stage->in_p[k]);
      const int4 b = *reinterpret_cast<const int4*>(&stage->in_p[k + 4]);
      const uint32_t w0 = ldg_w<kNoAlloc>(qb + a.x), w1 = ldg_w<kNoAlloc>(qb + a.y), w2 = ldg_w<kNoAlloc>(qb + a.z),
                     w3 = ldg_w<kNoAlloc>(qb + a.w), w4 = ldg_w<kNoAlloc>(qb + b.x), w5 = ldg_w<kNoAlloc>(qb + b.y),
                     w6 = ldg_w<kNoAlloc>(qb + b.z), w7 = ldg_w<kNoAlloc>(qb + b.w);
      lo += (w0 & 0x00FF00FFu) + (w1 & 0x00FF00FFu) + (w2 & 0x00FF00FFu) + (w3 & 0x00FF00FFu) +
            (w4 & 0x00FF00FFu) + (w5 & 0x00FF00FFu) + (w6 & 0x00FF00FFu) + (w7 & 0x00FF00FFu);
      hi += __byte_perm(w0, 0u, 0x4341) + __byte_perm(w1, 0u, 0x4341) +
            __byte_perm(w2, 0u, 0x4341) + __byte_perm(w3, 0u, 0x4341) +
            __byte_perm(w4, 0u, 0x4341) + __byte_perm(w5, 0u, 0x4341) +
            __byte_perm(w6, 0u, 0x4341) + __byte_perm(w7, 0u, 0x4341);
    }
    for (; k < cin; ++k) {
      const uint32_t w = ldg_w<kNoAlloc>(qb + stage->in_p[k]);
      lo += w & 0x00FF00FFu;
      hi += __byte_perm(w, 0u, 0x4341);
    }
    for (k = 0; k < cmix; ++k) {
      const int4 t = stage->mix[k];
      const bool ok = static_cast<unsigned>(t.y + ni) < static_cast<unsigned>(in) &&
                      static_cast<unsigned>(t.z + nj) < static_cast<unsigned>(jn);
      const uint32_t w = ldg_pred(qb, t.x, ok ? 1u : 0u);
      lo += w & 0x00FF00FFu;
      hi += __byte_perm(w, 0u, 0x4341);
    }
    s0 += lo & 0xFFFF; s2 += lo >> 16; s1 += hi & 0xFFFF; s3 += hi >> 16;
  }
  sum[0] = s0;
  sum[1] = s1;
  sum[2] = s2;
  sum[3] = s3;
}

// ---------------------------------------------------------------------------
// 3. lowest-resolution candidates (FCSM2D.cc:264-333).  grid = (max_scans, jobs).
//    The candidates of a scan form an nx x ny lattice with step 2^(depth-1); they are
//    scored as the children of virtual parents with step 2^depth.  The parents are
//    taken row by row, 32 per warp (lane = parent, node_sums): the lattice is small
//    (C3: about 10 x 10 parents per scan), so fixed 8 x 4 tiles left half the lanes idle.
//    Scores are stored in generation order (x outer, y inner).
// ---------------------------------------------------------------------------
constexpr int kTopWarps = 4;
__global__ void __launch_bounds__(kTopWarps * kWarp)
k_score_top(const JobDev* __restrict__ jobs, const int2* __restrict__ xy,
            const ScanInfo* __restrict__ scans, float* __restrict__ top_scores,
            JobState* __restrict__ state, unsigned long long* __restrict__ scan_best) {
  __shared__ NodeStage stage_all[kTopWarps];
  const JobDev& j = jobs[blockIdx.y];
  const int scan = blockIdx.x;
  if (scan >= j.S) return;
  const ScanInfo si = scans[j.scan_off + scan];
  const int count = si.nx * si.ny;
  if (count == 0) return;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  NodeStage* stage = &stage_all[warp];
  const int lvl = j.depth - 1;
  const int two_s = 2 << lvl;
  const int nvx = (si.nx + 1) >> 1, nvy = (si.ny + 1) >> 1;  // virtual parents
  const int parents = nvx * nvy;
  const int2* pts = xy + j.xy_off + static_cast<long long>(scan) * j.n;
  const StackDev& sd = *j.sd;
  unsigned long long local_best = 0ull;
  for (int f0 = warp * kWarp; f0 < parents; f0 += kTopWarps * kWarp) {
    const bool alive = f0 + lane < parents;
    const int f = alive ? f0 + lane : f0;  // idle lanes copy the item's first parent
    const int pj = f / nvx, pi = f - pj * nvx;
    int sum[4];
    node_sums(sd, pts, j.n, lvl, si.min_x + (two_s - 1), si.min_y + (two_s - 1), pi, pj, lane, stage,
              sum);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int ci = 2 * pi + (k >> 1), cj = 2 * pj + (k & 1);
      if (alive && ci < si.nx && cj < si.ny) {
        const float score = to_score(j, sum[k]);
        const unsigned int cand = static_cast<unsigned int>(si.base + ci * si.ny + cj);
        top_scores[j.cand_off + cand] = score;
        const unsigned long long key =
            (static_cast<unsigned long long>(float_to_ordered(score)) << 32) | (0xFFFFFFFFu - cand);
        local_best = max(local_best, key);
      }
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1)
    local_best = max(local_best, __shfl_xor_sync(kFull, local_best, o));
  if (lane == 0 && local_best != 0ull) {
    atomicMax(&state[blockIdx.y].top_best, local_best);
    atomicMax(&scan_best[j.scan_off + scan], local_best);
  }
  if (threadIdx.x == 0) {
    atomicAdd(&state[blockIdx.y].scored, static_cast<unsigned long long>(count));
    atomicAdd(&state[blockIdx.y].scored_phase[0], static_cast<unsigned long long>(count));
  }
}

// Appends `count` nodes of one bucket as a block: scratch range, bucket-relative offset
// and descriptor.  Called by one lane; returns the scratch position or ~0u on overflow.
__device__ __forceinline__ unsigned int open_block(unsigned int* __restrict__ counters,
                                                   unsigned int* __restrict__ bucket_cnt,
                                                   NodeBlock* __restrict__ blocks, int slot,
                                                   unsigned int bucket, unsigned int count,
                                                   unsigned long long cap,
                                                   unsigned long long block_cap) {
  const unsigned int src = atomicAdd(&counters[kCntTmp + slot], count);
  const unsigned int rel = atomicAdd(&bucket_cnt[bucket], count);
  const unsigned int q = atomicAdd(&counters[kCntBlocks + slot], 1u);
  if (static_cast<unsigned long long>(src) + count > cap || q >= block_cap) {
    atomicOr(&counters[kCntOverflow], 1u);
    return ~0u;
  }
  NodeBlock nb;
  nb.src = src;
  nb.count = count;
  nb.bucket = bucket;
  nb.rel = rel;
  blocks[q] = nb;
  return src;
}

// 5'a. seeds: lowest-resolution candidates above the threshold, one block of threads
//      per (job, scan), visited tile by tile; every warp batches up to 128 nodes.
__global__ void __launch_bounds__(kSeedWarps * kWarp)
k_seed_nodes(const JobDev* __restrict__ jobs, const ScanInfo* __restrict__ scans,
             const float* __restrict__ top_scores, const JobState* __restrict__ state,
             unsigned int* __restrict__ tmp, NodeBlock* __restrict__ blocks,
             unsigned int* __restrict__ bucket_cnt, unsigned int* __restrict__ counters, int slot,
             unsigned long long cap, unsigned long long block_cap, unsigned int* __restrict__ tmp_path,
             Deferred2* __restrict__ deferred, unsigned int record_cap) {
  __shared__ unsigned int buf_all[kSeedWarps][128];
  const JobDev& j = jobs[blockIdx.y];
  const int scan = blockIdx.x;
  if (scan >= j.S || j.depth - 1 != slot) return;
  const ScanInfo si = scans[j.scan_off + scan];
  if (si.nx == 0) return;
  const float thr = threshold(j, &state[blockIdx.y]);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  unsigned int* buf = buf_all[warp];
  const unsigned int bucket = static_cast<unsigned int>(j.scan_off + scan);
  // Row by row (32 x 1 lanes): every listed run has consecutive ni at one nj.
  const int tiles_x = (si.nx + kWarp - 1) / kWarp, tiles_y = si.ny;
  int held = 0;
  auto flush = [&]() {
    unsigned int src = 0u;
    if (lane == 0)
      src = open_block(counters, bucket_cnt, blocks, slot, bucket, held, cap, block_cap);
    src = __shfl_sync(kFull, src, 0);
    __syncwarp();
    if (src != ~0u)
      for (int t = lane; t < held; t += kWarp) {
        tmp[src + t] = buf[t];
        if (tmp_path != nullptr) tmp_path[src + t] = 0u;  // a top-level candidate: empty path
      }
    __syncwarp();
    held = 0;
  };
  bool has_inc = false;
  const float inc = incumbent(&state[blockIdx.y], has_inc);
  for (int t = warp; t < tiles_x * tiles_y; t += kSeedWarps) {
    const int ti = t % tiles_x, cj = t / tiles_x;
    const int ci = kWarp * ti + lane;
    bool alive = false;
    if (ci < si.nx) {
      const float top = top_scores[j.cand_off + si.base + ci * si.ny + cj];
      alive = top > thr;
      if (deferred != nullptr && !alive && has_inc && top == inc) {  // bound == incumbent: set aside
        Deferred2 d;
        d.job = blockIdx.y;
        d.scan = scan;
        d.x = si.min_x + (ci << slot);
        d.y = si.min_y + (cj << slot);
        d.depth = slot;
        d.cand = static_cast<unsigned int>(si.base + ci * si.ny + cj);
        d.path = 0u;
        d.score = top;
        push_record(deferred, &counters[kCntDeferred], record_cap, &counters[kCntOverflow], d);
      }
    }
    const uint32_t m = __ballot_sync(kFull, alive);
    if (alive) buf[held + __popc(m & ((1u << lane) - 1u))] = static_cast<unsigned int>(ci | (cj << 16));
    held += __popc(m);
    if (held > 96) flush();
  }
  if (held > 0) flush();
}

// 5'b. bucket counters -> list offsets and work-item offsets of depth `slot`.  Block q scans
//      the kPrefixChunk buckets [q * kPrefixChunk, ...) (kPrefixPer consecutive buckets per
//      thread) and leaves chunk-local offsets in bucket_start / bucket_istart plus its totals
//      in the tail of those arrays ([num_buckets + 1 + q]); the last block to finish turns
//      the totals into chunk bases in place.  Readers add the base of a bucket's chunk
//      (bucket_base).  A C3 step has 126 k buckets: one block needed 137 us per level.
constexpr int kPrefixPer = 8;
constexpr int kLogPrefixChunk = 13;  // 1024 threads x kPrefixPer buckets
constexpr int kPrefixChunk = 1 << kLogPrefixChunk;
constexpr int kCntTicket = 34;  // blocks of k_bucket_prefix that have finished
static_assert(kPrefixChunk == 1024 * kPrefixPer, "one pass per block");

__device__ __forceinline__ unsigned int bucket_base(const unsigned int* __restrict__ starts,
                                                    unsigned int num_buckets, unsigned int b) {
  // absolute offset of bucket b (b == num_buckets: the total)
  return b == num_buckets ? starts[num_buckets]
                          : starts[b] + starts[num_buckets + 1 + (b >> kLogPrefixChunk)];
}

__global__ void __launch_bounds__(1024)
k_bucket_prefix(unsigned int* __restrict__ bucket_cnt, unsigned int* __restrict__ bucket_start,
                unsigned int* __restrict__ bucket_istart, unsigned int* __restrict__ counters,
                int slot, unsigned int num_buckets, unsigned long long cap,
                unsigned long long item_cap) {
  __shared__ unsigned int wsum[2][32];
  __shared__ bool s_last;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  {
    const unsigned int first = blockIdx.x * kPrefixChunk + threadIdx.x * kPrefixPer;
    unsigned int c[kPrefixPer];
    unsigned int tc = 0u, ti = 0u;  // this thread's totals
#pragma unroll
    for (int k = 0; k < kPrefixPer; ++k) {
      c[k] = 0u;
      if (first + k < num_buckets) {
        c[k] = bucket_cnt[first + k];
        bucket_cnt[first + k] = 0u;  // ready for the next level
      }
      tc += c[k];
      ti += (c[k] + kWarp - 1) / kWarp;
    }
    unsigned int pc = tc, pi = ti;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const unsigned int uc = __shfl_up_sync(kFull, pc, o), ui = __shfl_up_sync(kFull, pi, o);
      if (lane >= o) {
        pc += uc;
        pi += ui;
      }
    }
    if (lane == 31) {
      wsum[0][warp] = pc;
      wsum[1][warp] = pi;
    }
    __syncthreads();
    if (warp == 0) {
      unsigned int a = wsum[0][lane], i2 = wsum[1][lane];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const unsigned int ta = __shfl_up_sync(kFull, a, o), tb = __shfl_up_sync(kFull, i2, o);
        if (lane >= o) {
          a += ta;
          i2 += tb;
        }
      }
      wsum[0][lane] = a;
      wsum[1][lane] = i2;
    }
    __syncthreads();
    unsigned int at_c = (warp ? wsum[0][warp - 1] : 0u) + pc - tc;
    unsigned int at_i = (warp ? wsum[1][warp - 1] : 0u) + pi - ti;
#pragma unroll
    for (int k = 0; k < kPrefixPer; ++k) {
      if (first + k < num_buckets) {
        bucket_start[first + k] = at_c;
        bucket_istart[first + k] = at_i;
      }
      at_c += c[k];
      at_i += (c[k] + kWarp - 1) / kWarp;
    }
    if (threadIdx.x == 0) {
      bucket_start[num_buckets + 1 + blockIdx.x] = wsum[0][31];   // chunk totals for now
      bucket_istart[num_buckets + 1 + blockIdx.x] = wsum[1][31];
    }
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) s_last = atomicAdd(&counters[kCntTicket], 1u) == gridDim.x - 1;
  __syncthreads();
  if (s_last && threadIdx.x == 0) {
    __threadfence();
    volatile unsigned int* vc = bucket_start + num_buckets + 1;
    volatile unsigned int* vi = bucket_istart + num_buckets + 1;
    unsigned int run_c = 0u, run_i = 0u;
    for (unsigned int q = 0; q < gridDim.x; ++q) {  // totals -> chunk bases, in place
      const unsigned int tcq = vc[q], tiq = vi[q];
      vc[q] = run_c;
      vi[q] = run_i;
      run_c += tcq;
      run_i += tiq;
    }
    bucket_start[num_buckets] = run_c;
    bucket_istart[num_buckets] = run_i;
    counters[kCntNodes + slot] = run_c;
    counters[kCntItems + slot] = run_i;
    counters[kCntTicket] = 0u;
    if (run_c > cap || run_i > item_cap) atomicOr(&counters[kCntOverflow], 1u);
  }
}

// 5'c. blocks -> bucket ranges of the sorted list; work items of depth `slot`.
//      item = (list position, bucket * 32 + nodes - 1)
__global__ void __launch_bounds__(256)
k_place(const unsigned int* __restrict__ tmp, const NodeBlock* __restrict__ blocks,
        const unsigned int* __restrict__ bucket_start, const unsigned int* __restrict__ bucket_istart,
        unsigned int* __restrict__ nodes, uint2* __restrict__ items,
        const unsigned int* __restrict__ counters, int slot, unsigned int num_buckets,
        unsigned long long cap, unsigned long long block_cap, unsigned long long item_cap,
        const unsigned int* __restrict__ tmp_path, unsigned int* __restrict__ node_path) {
  if (counters[kCntOverflow] & 1u) return;  // the batch is replayed with larger lists
  const int lane = threadIdx.x & 31;
  const unsigned int warp = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const unsigned int warps = gridDim.x * (blockDim.x >> 5);
  const unsigned int nblocks =
      static_cast<unsigned int>(min(static_cast<unsigned long long>(counters[kCntBlocks + slot]), block_cap));
  // A block is at most a few hundred bytes, so one copy is three dependent memory latencies
  // (descriptor, bucket base, data) for little traffic: two blocks per turn, their loads issued
  // together, and twice the resident warps keep more of them in flight (1.2 TB/s before).
  for (unsigned int q = warp; q < nblocks; q += 2 * warps) {
    const unsigned int q2 = q + warps;
    const bool two = q2 < nblocks;
    const NodeBlock na = blocks[q];
    const NodeBlock nb = two ? blocks[q2] : NodeBlock{};
    const unsigned int da = bucket_base(bucket_start, num_buckets, na.bucket) + na.rel;
    const unsigned int db = two ? bucket_base(bucket_start, num_buckets, nb.bucket) + nb.rel : 0u;
    const unsigned int cb = two ? nb.count : 0u;
    for (unsigned int t = lane; t < max(na.count, cb); t += kWarp) {
      unsigned int va = 0, vb = 0, pa = 0, pb = 0;
      const bool ha = t < na.count, hb = t < cb;
      if (ha) va = tmp[na.src + t];
      if (hb) vb = tmp[nb.src + t];
      if (tmp_path != nullptr) {
        if (ha) pa = tmp_path[na.src + t];
        if (hb) pb = tmp_path[nb.src + t];
      }
      if (ha) nodes[da + t] = va;
      if (hb) nodes[db + t] = vb;
      if (tmp_path != nullptr) {
        if (ha) node_path[da + t] = pa;
        if (hb) node_path[db + t] = pb;
      }
    }
  }
  // Work items.  A lane looks at one bucket: most of the 126 k buckets of a C3 step are empty or
  // hold a handful of items, and a lane writes the first two of its bucket itself; a bucket with
  // more (thousands at the dense depths, where one lane writing them all held its warp for the
  // whole launch) is then written by the 32 lanes together.
  const unsigned int group = blockIdx.x * blockDim.x + (threadIdx.x & ~31u);  // first bucket of this warp's pass
  const unsigned int threads = gridDim.x * blockDim.x;
  for (unsigned int b0 = group; b0 < num_buckets; b0 += threads) {
    const unsigned int b = b0 + lane;
    unsigned int first = 0, c = 0, ibase = 0;
    if (b < num_buckets) {
      first = bucket_base(bucket_start, num_buckets, b);
      c = bucket_base(bucket_start, num_buckets, b + 1) - first;
      if (c != 0u) ibase = bucket_base(bucket_istart, num_buckets, b);
    }
    const unsigned int n_items = (c + kWarp - 1) / kWarp;
    for (unsigned int t = 0; t < min(n_items, 2u); ++t)
      items[ibase + t] = make_uint2(first + t * kWarp, b * kWarp + (min(c - t * kWarp, 32u) - 1u));
    unsigned int big = __ballot_sync(kFull, n_items > 2u);
    while (big != 0u) {
      const int src = __ffs(big) - 1;
      big &= big - 1u;
      const unsigned int s_first = __shfl_sync(kFull, first, src), s_c = __shfl_sync(kFull, c, src);
      const unsigned int s_ibase = __shfl_sync(kFull, ibase, src), s_items = __shfl_sync(kFull, n_items, src);
      const unsigned int s_b = b0 + src;
      for (unsigned int t = 2u + lane; t < s_items; t += kWarp)
        items[s_ibase + t] = make_uint2(s_first + t * kWarp, s_b * kWarp + (min(s_c - t * kWarp, 32u) - 1u));
    }
  }
}

// 5'd. expand the nodes of depth d: one warp per work item, lane = alive node.  Work
//      items are handed out in list order through a cursor (one atomic per item, the
//      next one fetched while the current item is summed): an item costs between a few
//      and several hundred loads per lane, so a static split leaves SMs idle at the end
//      (ncu: l1tex busy 60 % of elapsed but 86 % of active cycles before this).
template <bool kNoAlloc>
__global__ void __launch_bounds__(kExpandWarps * kWarp, kNodeBlocksPerSm)
k_expand_nodes(const JobDev* __restrict__ jobs, const int2* __restrict__ xy,
               const ScanInfo* __restrict__ scans, const float* __restrict__ top_scores,
               JobState* __restrict__ state, const unsigned int* __restrict__ bucket_js,
               const unsigned int* __restrict__ nodes, const uint2* __restrict__ items,
               unsigned int* __restrict__ tmp, NodeBlock* __restrict__ blocks,
               unsigned int* __restrict__ bucket_cnt, unsigned int* __restrict__ counters, int d,
               unsigned long long cap, unsigned long long block_cap, unsigned long long item_cap,
               unsigned long long* __restrict__ hist, const unsigned int* __restrict__ node_path,
               unsigned int* __restrict__ tmp_path, Leaf2* __restrict__ leaves,
               Deferred2* __restrict__ deferred, unsigned int record_cap) {
  __shared__ NodeStage stage_all[kExpandWarps];
  if (counters[kCntOverflow] & 1u) return;
  const unsigned int count =
      static_cast<unsigned int>(min(static_cast<unsigned long long>(counters[kCntItems + d]), item_cap));
  const int lane = threadIdx.x & 31;
  NodeStage* stage = &stage_all[threadIdx.x >> 5];
  const int lvl = d - 1;
  const int two_s = 2 << lvl;
  int acc_job = -1;
  unsigned long long acc_scored = 0ull;
  auto next_item = [&]() -> unsigned int {
    unsigned int v = 0u;
    if (lane == 0) v = atomicAdd(&counters[kCntCursor + d], 1u);
    return v;
  };
  unsigned int fetched = next_item();
  for (;;) {
    const unsigned int it = __shfl_sync(kFull, fetched, 0);
    if (it >= count) break;
    fetched = next_item();
    const uint2 item = items[it];
    const unsigned int bucket = item.y >> 5;
    const int cnt = static_cast<int>(item.y & 31u) + 1;
    const unsigned int js = bucket_js[bucket];
    const int job = js >> 16, scan = js & 0xFFFF;
    const JobDev& j = jobs[job];
    JobState* st = &state[job];
    const ScanInfo si = scans[j.scan_off + scan];
    const bool alive = lane < cnt;
    const unsigned int node = nodes[item.x + (alive ? lane : 0)];
    const int ni = node & 0xFFFF, nj = node >> 16;
    const bool paths = node_path != nullptr;
    if (hist != nullptr && lane == 0) atomicAdd(&hist[d * 33 + cnt], 1ull);
    int sum[4];
    node_sums<kNoAlloc>(*j.sd, xy + j.xy_off + static_cast<long long>(scan) * j.n, j.n, lvl,
              si.min_x + (two_s - 1), si.min_y + (two_s - 1), ni, nj, lane, stage, sum);
    // One snapshot of the incumbent per item, warp-uniform (it moves under concurrent atomics),
    // taken after the sums: fresher, and nothing of it is live across the summing loop.
    unsigned long long snap = *reinterpret_cast<const volatile unsigned long long*>(&st->best);
    snap = __shfl_sync(kFull, snap, 0);
    const bool has_inc = (snap >> 32) != 0ull;
    const float inc = has_inc ? ordered_to_float(static_cast<unsigned int>(snap >> 32)) : 0.f;
    const float thr = has_inc ? fmaxf(j.min_score, inc) : j.min_score;
    const unsigned int ppath = paths ? node_path[item.x + (alive ? lane : 0)] : 0u;
    // Children lattice of this scan on level lvl (FCSM2D.cc:353-367: a child beyond
    // max_x / max_y is not generated).
    const int nxl = ((si.max_x - si.min_x) >> lvl) + 1, nyl = ((si.max_y - si.min_y) >> lvl) + 1;
    const int ci0 = 2 * ni, cj0 = 2 * nj;
    bool pass[4], valid[4];
    float sc[4];
    int nvalid = 0, npass = 0;
    bool any_tie = false;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      valid[k] = alive && ci0 + (k >> 1) < nxl && cj0 + (k & 1) < nyl;
      sc[k] = to_score(j, sum[k]);
      pass[k] = valid[k] && sc[k] > thr;
      any_tie = any_tie || (valid[k] && has_inc && sc[k] == inc);
      nvalid += valid[k] ? 1 : 0;
      npass += pass[k] ? 1 : 0;
    }
    // Path mode: rank of every child among its sorted siblings (score descending, generation
    // order among equals), packed two bits each.  A bound that EQUALS the incumbent is neither
    // expanded nor forgotten: it is set aside (interior node) or recorded (leaf).
    unsigned int ranks = 0u;
    if (paths && (npass != 0 || any_tie)) {
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        unsigned int r = 0u;
#pragma unroll
        for (int q = 0; q < 4; ++q)
          r += (q != k && valid[q] && (sc[q] > sc[k] || (sc[q] == sc[k] && q < k))) ? 1u : 0u;
        ranks |= r << (2 * k);
      }
    }
    if (paths && any_tie) {
      const int top_shift = j.depth - 1 - lvl;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        if (!(valid[k] && has_inc && sc[k] == inc)) continue;
        const int ci = ci0 + (k >> 1), cj = cj0 + (k & 1);
        const unsigned int cand = static_cast<unsigned int>(si.base + (ci >> top_shift) * si.ny + (cj >> top_shift));
        const unsigned int path = (ppath << 2) | ((ranks >> (2 * k)) & 3u);
        if (lvl == 0) {
          Leaf2 lf;
          lf.job = job; lf.scan = scan; lf.x = si.min_x + ci; lf.y = si.min_y + cj;
          lf.cand = cand; lf.path = path; lf.score = sc[k]; lf.pad = 0;
          push_record(leaves, &counters[kCntLeaves], record_cap, &counters[kCntOverflow], lf);
        } else {
          Deferred2 df;
          df.job = job; df.scan = scan; df.x = si.min_x + (ci << lvl); df.y = si.min_y + (cj << lvl);
          df.depth = lvl; df.cand = cand; df.path = path; df.score = sc[k];
          push_record(deferred, &counters[kCntDeferred], record_cap, &counters[kCntOverflow], df);
        }
      }
    }
    const int tot = __reduce_add_sync(kFull, nvalid);
    if (job != acc_job) {
      if (lane == 0 && acc_job >= 0) {
        atomicAdd(&state[acc_job].scored, acc_scored);
        atomicAdd(&state[acc_job].scored_phase[2], acc_scored);
      }
      acc_job = job;
      acc_scored = 0ull;
    }
    acc_scored += static_cast<unsigned long long>(tot);
    if (lvl == 0) {
      // Leaves.  Rare: only a leaf above the current incumbent reaches the atomic.
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        if (!pass[k]) continue;
        const int ci = ci0 + (k >> 1), cj = cj0 + (k & 1);
        const int ix = ci >> (j.depth - 1), iy = cj >> (j.depth - 1);  // top-level ancestor
        const float ts = top_scores[j.cand_off + si.base + ix * si.ny + iy];
        const unsigned long long key = leaf_key(sc[k], ts);
        const unsigned long long cur =
            *reinterpret_cast<const volatile unsigned long long*>(&st->best);
        if (key > cur) atomicMax(&st->best, key);
        if (paths) {
          Leaf2 lf;
          lf.job = job; lf.scan = scan; lf.x = si.min_x + ci; lf.y = si.min_y + cj;
          lf.cand = static_cast<unsigned int>(si.base + ix * si.ny + iy);
          lf.path = (ppath << 2) | ((ranks >> (2 * k)) & 3u); lf.score = sc[k]; lf.pad = 0;
          push_record(leaves, &counters[kCntLeaves], record_cap, &counters[kCntOverflow], lf);
        }
      }
    } else {
      // Survivors are emitted as two ROW RUNS per work item: the even-row children
      // (cj0) of every lane, then the odd-row ones (cj0 + 1), each in lane order.  A
      // plane row is contiguous in the quad layout (index = P + nj * in + ni), so when
      // the parents of an item lie along a row -- seeds are listed row by row, and by
      // induction every later item is a row run -- the 32 nodes of a child item share
      // their 32-byte sectors along ni instead of spreading over a 2-D Z-order block.
      const int ne = (pass[0] ? 1 : 0) + (pass[2] ? 1 : 0), no = npass - ne;
      int pre = ne | (no << 16);  // inclusive scan over lanes, both counts at once
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(kFull, pre, o);
        if (lane >= o) pre += t;
      }
      const int totals = __shfl_sync(kFull, pre, 31);
      const int total_e = totals & 0xFFFF, total = total_e + (totals >> 16);
      if (total > 0) {
        unsigned int src = 0u;
        if (lane == 0)
          src = open_block(counters, bucket_cnt, blocks, d - 1, bucket, total, cap, block_cap);
        src = __shfl_sync(kFull, src, 0);
        if (src != ~0u) {
          unsigned int at_e = src + (pre & 0xFFFF) - ne;
          unsigned int at_o = src + total_e + (pre >> 16) - no;
          if (paths) {
            const unsigned int pp = ppath << 2;
            unsigned int pe = at_e, po = at_o;
            if (pass[0]) tmp_path[pe++] = pp | (ranks & 3u);
            if (pass[2]) tmp_path[pe] = pp | ((ranks >> 4) & 3u);
            if (pass[1]) tmp_path[po++] = pp | ((ranks >> 2) & 3u);
            if (pass[3]) tmp_path[po] = pp | ((ranks >> 6) & 3u);
          }
          if (pass[0]) tmp[at_e++] = static_cast<unsigned int>(ci0 | (cj0 << 16));
          if (pass[2]) tmp[at_e] = static_cast<unsigned int>((ci0 + 1) | (cj0 << 16));
          if (pass[1]) tmp[at_o++] = static_cast<unsigned int>(ci0 | ((cj0 + 1) << 16));
          if (pass[3]) tmp[at_o] = static_cast<unsigned int>((ci0 + 1) | ((cj0 + 1) << 16));
        }
      }
    }
  }
  if (lane == 0 && acc_job >= 0) {
    atomicAdd(&state[acc_job].scored, acc_scored);
    atomicAdd(&state[acc_job].scored_phase[2], acc_scored);
  }
}

// Profiling only (csm_ctx_set_profiling(ctx, 2)): what the work items of depth d ask of the
// memory system.  One warp per item; per item and scan point the sweep issues ONE warp-wide
// 4-byte load whose lanes read quad[P + nj * in + ni].  The 32-byte sectors (128-byte lines)
// that request touches are the distinct (nj, (ni + a) >> 3) ((ni + a) >> 5) over the alive
// lanes, a = the point's alignment inside a sector (line): averaged over all alignments.
// stats[d * 6 ..] += items, requests (points; the few points the sweep drops at the plane's
// border are counted too), lane loads, sectors, lines, nodes.
__global__ void __launch_bounds__(256)
k_item_stats(const JobDev* __restrict__ jobs, const unsigned int* __restrict__ bucket_js,
             const unsigned int* __restrict__ nodes, const uint2* __restrict__ items,
             const unsigned int* __restrict__ counters, int d, unsigned long long item_cap,
             double* __restrict__ stats) {
  const unsigned int count =
      static_cast<unsigned int>(min(static_cast<unsigned long long>(counters[kCntItems + d]), item_cap));
  const int lane = threadIdx.x & 31;
  const unsigned int warps = gridDim.x * (blockDim.x >> 5);
  double acc[6] = {0., 0., 0., 0., 0., 0.};
  for (unsigned int it = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); it < count; it += warps) {
    const uint2 item = items[it];
    const int cnt = static_cast<int>(item.y & 31u) + 1;
    const unsigned int js = bucket_js[item.y >> 5];
    const int n = jobs[js >> 16].n;
    const bool alive = lane < cnt;
    const unsigned int mask = __ballot_sync(kFull, alive);
    int sectors = 0, lines = 0;
    if (alive) {
      const unsigned int node = nodes[item.x + lane];
      const unsigned int ni = node & 0xFFFFu, nj = node >> 16;
      for (unsigned int a = 0; a < 8; ++a) {
        const unsigned int m = __match_any_sync(mask, (nj << 16) | ((ni + a) >> 3));
        sectors += __popc(__ballot_sync(mask, lane == __ffs(m) - 1));
      }
      for (unsigned int a = 0; a < 32; a += 8) {
        const unsigned int m = __match_any_sync(mask, (nj << 16) | ((ni + a) >> 5));
        lines += __popc(__ballot_sync(mask, lane == __ffs(m) - 1));
      }
    }
    sectors = __shfl_sync(kFull, sectors, 0);
    lines = __shfl_sync(kFull, lines, 0);
    acc[0] += 1.;
    acc[1] += n;
    acc[2] += static_cast<double>(cnt) * n;
    acc[3] += n * (sectors / 8.);
    acc[4] += n * (lines / 4.);
    acc[5] += cnt;
  }
  if (lane == 0 && acc[0] > 0.)
    for (int k = 0; k < 6; ++k) atomicAdd(&stats[d * 6 + k], acc[k]);
}

// ---------------------------------------------------------------------------
// 6. phase 2: which leaf attaining S* does the reference's DFS reach first?
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
k_collect(const JobDev* __restrict__ jobs, const ScanInfo* __restrict__ scans,
          const float* __restrict__ top_scores, const JobState* __restrict__ state,
          Contender* __restrict__ out, unsigned int* __restrict__ counters, unsigned int cap) {
  const JobDev& j = jobs[blockIdx.y];
  const int scan = blockIdx.x;
  if (scan >= j.S) return;
  const unsigned long long best = state[blockIdx.y].best;
  if ((best >> 32) == 0ull) return;  // nothing above min_score
  const float v0 = ordered_to_float(static_cast<unsigned int>(best & 0xFFFFFFFFull));
  const ScanInfo si = scans[j.scan_off + scan];
  const int count = si.nx * si.ny;
  const int step = 1 << (j.depth - 1);
  for (int c = threadIdx.x; c < count; c += blockDim.x) {
    const float score = top_scores[j.cand_off + si.base + c];
    if (!(score >= v0)) continue;
    const unsigned int slot = atomicAdd(&counters[kCntContenders], 1u);
    if (slot >= cap) {
      atomicOr(&counters[kCntOverflow], 4u);
      continue;
    }
    const int ix = c / si.ny, iy = c - ix * si.ny;
    Contender ct;
    ct.job = blockIdx.y;
    ct.scan = scan;
    ct.x = si.min_x + ix * step;
    ct.y = si.min_y + iy * step;
    ct.cand = static_cast<unsigned int>(si.base + c);
    ct.score = score;
    ct.leaf_x = ct.leaf_y = 0;
    ct.has_leaf = 0;
    ct.pad[0] = ct.pad[1] = ct.pad[2] = 0;
    out[slot] = ct;
  }
}

// One warp per contender: depth-first in the reference's sibling order
// (std::sort of <= 4 children = insertion sort = stable: score descending, then
// generation order; FCSM2D.cc:331-332,353-376), visiting only nodes whose bound is
// >= S*, stopping at the first leaf that attains S*.
__global__ void __launch_bounds__(128)
k_ordered_dfs(const JobDev* __restrict__ jobs, const int2* __restrict__ xy,
              const ScanInfo* __restrict__ scans, JobState* __restrict__ state,
              Contender* __restrict__ list, unsigned int* __restrict__ counters,
              unsigned int cap) {
  const unsigned int count = min(counters[kCntContenders], cap);
  const int lane = threadIdx.x & 31;
  // Contenders are handed out through a cursor: most die after one visit, a few walk deep.
  for (;;) {
    unsigned int i = 0u;
    if (lane == 0) i = atomicAdd(&counters[kCntDfsCursor], 1u);
    i = __shfl_sync(kFull, i, 0);
    if (i >= count) break;
    Contender ct = list[i];
    const JobDev& j = jobs[ct.job];
    const ScanInfo si = scans[j.scan_off + ct.scan];
    const float s_star = ordered_to_float(static_cast<unsigned int>(state[ct.job].best >> 32));
    const int top_depth = j.depth - 1;
    int fx[CSM_MAX_DEPTH], fy[CSM_MAX_DEPTH];
    unsigned int forder[CSM_MAX_DEPTH];  // 4 x 2-bit child indices, low bits first
    int fcount[CSM_MAX_DEPTH], fnext[CSM_MAX_DEPTH];
    int lv = 0;
    fx[0] = ct.x;
    fy[0] = ct.y;
    fcount[0] = -1;  // not expanded yet
    bool found = false;
    int leaf_x = 0, leaf_y = 0;
    unsigned long long scored = 0;
    if (top_depth == 0) {  // depth-1 stack: the candidate is itself the leaf
      found = ct.score == s_star;
      leaf_x = ct.x;
      leaf_y = ct.y;
      lv = -1;
    }
    while (lv >= 0 && !found) {
      const int d = top_depth - lv;
      const int half = 1 << (d - 1);
      if (fcount[lv] < 0) {
        const Children c = score_children(j, xy, si, ct.scan, fx[lv], fy[lv], d, lane);
        int idx[4], n = 0;
        for (int k = 0; k < 4; ++k) {
          scored += c.valid[k] ? 1 : 0;
          if (!(c.valid[k] && c.score[k] >= s_star && c.score[k] > j.min_score)) continue;
          int pos = n++;
          while (pos > 0 && c.score[idx[pos - 1]] < c.score[k]) {  // stable insertion
            idx[pos] = idx[pos - 1];
            --pos;
          }
          idx[pos] = k;
        }
        if (d == 1) {
          if (n > 0) {  // leaves cannot exceed S*, so the first survivor attains it
            found = true;
            leaf_x = fx[lv] + ((idx[0] & 2) ? half : 0);
            leaf_y = fy[lv] + ((idx[0] & 1) ? half : 0);
          } else {
            --lv;
          }
          continue;
        }
        unsigned int packed = 0;
        for (int k = 0; k < n; ++k) packed |= static_cast<unsigned int>(idx[k]) << (2 * k);
        forder[lv] = packed;
        fcount[lv] = n;
        fnext[lv] = 0;
      }
      if (fnext[lv] < fcount[lv]) {
        const int k = (forder[lv] >> (2 * fnext[lv])) & 3;
        ++fnext[lv];
        fx[lv + 1] = fx[lv] + ((k & 2) ? half : 0);
        fy[lv + 1] = fy[lv] + ((k & 1) ? half : 0);
        fcount[lv + 1] = -1;
        ++lv;
      } else {
        --lv;
      }
    }
    if (lane == 0) {
      list[i].has_leaf = found ? 1 : 0;
      list[i].leaf_x = leaf_x;
      list[i].leaf_y = leaf_y;
      if (found) atomicMax(&state[ct.job].group_best, float_to_ordered(ct.score));
      atomicAdd(&state[ct.job].scored, scored);
      atomicAdd(&state[ct.job].scored_phase[3], scored);
    }
  }
}

__global__ void __launch_bounds__(256)
k_pick(JobState* __restrict__ state, const Contender* __restrict__ list,
       const unsigned int* __restrict__ counters, unsigned int cap) {
  const unsigned int count = min(counters[kCntContenders], cap);
  const unsigned int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= count) return;
  const Contender ct = list[i];
  if (!ct.has_leaf) return;
  JobState* st = &state[ct.job];
  if (float_to_ordered(ct.score) != st->group_best) return;
  atomicAdd(&st->tie_count, 1u);
  atomicMin(&st->win, (static_cast<unsigned long long>(ct.cand) << 32) | i);
}

// ---------------------------------------------------------------------------
// 6'. path mode: the winner from the leaves the sweep recorded.
// ---------------------------------------------------------------------------
// One warp per set-aside node whose bound equals the final S*: depth-first in the reference's
// sibling order over nodes with bound >= S*; the first leaf reached attains S* and is the first
// S*-leaf of that subtree.  Its path continues the node's.  Few nodes qualify (exact ties of a
// bound with the best leaf), so single-path latency does not matter here.
__global__ void __launch_bounds__(128)
k_deferred_dfs(const JobDev* __restrict__ jobs, const int2* __restrict__ xy,
               const ScanInfo* __restrict__ scans, JobState* __restrict__ state,
               const Deferred2* __restrict__ list, Leaf2* __restrict__ leaves,
               unsigned int* __restrict__ counters, unsigned int cap) {
  const unsigned int count = min(counters[kCntDeferred], cap);
  const int lane = threadIdx.x & 31;
  for (;;) {
    unsigned int i = 0u;
    if (lane == 0) i = atomicAdd(&counters[kCntDeferCursor], 1u);
    i = __shfl_sync(kFull, i, 0);
    if (i >= count) break;
    const Deferred2 df = list[i];
    const JobDev& j = jobs[df.job];
    const float s_star = ordered_to_float(static_cast<unsigned int>(state[df.job].best >> 32));
    if (!(df.score == s_star)) continue;  // its bound stayed below the best leaf: nothing to find
    const ScanInfo si = scans[j.scan_off + df.scan];
    int fx[CSM_MAX_DEPTH], fy[CSM_MAX_DEPTH];
    unsigned int forder[CSM_MAX_DEPTH], fpath[CSM_MAX_DEPTH];
    int fcount[CSM_MAX_DEPTH], fnext[CSM_MAX_DEPTH];
    int lv = 0;
    fx[0] = df.x;
    fy[0] = df.y;
    fpath[0] = df.path;
    fcount[0] = -1;
    bool found = false;
    int leaf_x = 0, leaf_y = 0;
    unsigned int leaf_path = 0u;
    unsigned long long scored = 0;
    while (lv >= 0 && !found) {
      const int d = df.depth - lv;
      const int half = 1 << (d - 1);
      if (fcount[lv] < 0) {
        const Children c = score_children(j, xy, si, df.scan, fx[lv], fy[lv], d, lane);
        int idx[4], n = 0;
        for (int k = 0; k < 4; ++k) {
          scored += c.valid[k] ? 1 : 0;
          if (!(c.valid[k] && c.score[k] >= s_star && c.score[k] > j.min_score)) continue;
          int pos = n++;
          while (pos > 0 && c.score[idx[pos - 1]] < c.score[k]) {  // stable insertion
            idx[pos] = idx[pos - 1];
            --pos;
          }
          idx[pos] = k;
        }
        if (d == 1) {
          if (n > 0) {  // leaves cannot exceed S*: the first survivor attains it, with rank 0
            found = true;
            leaf_x = fx[lv] + ((idx[0] & 2) ? half : 0);
            leaf_y = fy[lv] + ((idx[0] & 1) ? half : 0);
            leaf_path = fpath[lv] << 2;
          } else {
            --lv;
          }
          continue;
        }
        unsigned int packed = 0;
        for (int k = 0; k < n; ++k) packed |= static_cast<unsigned int>(idx[k]) << (2 * k);
        forder[lv] = packed;
        fcount[lv] = n;
        fnext[lv] = 0;
      }
      if (fnext[lv] < fcount[lv]) {
        // survivors (bound >= S*) precede every other sibling in sorted order, so the position
        // among them IS the rank among all siblings
        const int k = (forder[lv] >> (2 * fnext[lv])) & 3;
        fpath[lv + 1] = (fpath[lv] << 2) | static_cast<unsigned int>(fnext[lv]);
        ++fnext[lv];
        fx[lv + 1] = fx[lv] + ((k & 2) ? half : 0);
        fy[lv + 1] = fy[lv] + ((k & 1) ? half : 0);
        fcount[lv + 1] = -1;
        ++lv;
      } else {
        --lv;
      }
    }
    if (lane == 0) {
      if (found) {
        Leaf2 lf;
        lf.job = df.job; lf.scan = df.scan; lf.x = leaf_x; lf.y = leaf_y;
        lf.cand = df.cand; lf.path = leaf_path; lf.score = s_star; lf.pad = 0;
        push_record(leaves, &counters[kCntLeaves], cap, &counters[kCntOverflow], lf);
      }
      atomicAdd(&state[df.job].scored, scored);
      atomicAdd(&state[df.job].scored_phase[3], scored);
    }
  }
}

// Depth-1 stacks are not swept: their lowest-resolution candidates ARE the leaves, and k_dive has
// offered the best of every scan.  Every candidate that attains S* is recorded.
__global__ void __launch_bounds__(256)
k_leaves_depth1(const JobDev* __restrict__ jobs, const ScanInfo* __restrict__ scans,
                const float* __restrict__ top_scores, const JobState* __restrict__ state,
                Leaf2* __restrict__ leaves, unsigned int* __restrict__ counters, unsigned int cap) {
  const JobDev& j = jobs[blockIdx.y];
  const int scan = blockIdx.x;
  if (scan >= j.S || j.depth != 1) return;
  const unsigned long long best = state[blockIdx.y].best;
  if ((best >> 32) == 0ull) return;
  const float s_star = ordered_to_float(static_cast<unsigned int>(best >> 32));
  const ScanInfo si = scans[j.scan_off + scan];
  const int count = si.nx * si.ny;
  for (int c = threadIdx.x; c < count; c += blockDim.x) {
    const float score = top_scores[j.cand_off + si.base + c];
    if (!(score == s_star)) continue;
    const int ix = c / si.ny, iy = c - ix * si.ny;
    Leaf2 lf;
    lf.job = blockIdx.y; lf.scan = scan; lf.x = si.min_x + ix; lf.y = si.min_y + iy;
    lf.cand = static_cast<unsigned int>(si.base + c); lf.path = 0u; lf.score = score; lf.pad = 0;
    push_record(leaves, &counters[kCntLeaves], cap, &counters[kCntOverflow], lf);
  }
}

// stage 0: best top-level score among the ancestors of S*-leaves; stage 1: smallest (ancestor,
// path) inside that group, and the largest ancestor (more than one ancestor: the host ranks them
// with the reference's std::sort); stage 2: the winning leaf writes its offsets.
__global__ void __launch_bounds__(256)
k_pick_leaf(const JobDev* __restrict__ jobs, const float* __restrict__ top_scores,
            JobState* __restrict__ state, const Leaf2* __restrict__ leaves,
            const unsigned int* __restrict__ counters, unsigned int cap, int stage,
            csm_result2d* __restrict__ results) {
  const unsigned int count = min(counters[kCntLeaves], cap);
  for (unsigned int i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += gridDim.x * blockDim.x) {
    const Leaf2 lf = leaves[i];
    JobState* st = &state[lf.job];
    if (float_to_ordered(lf.score) != static_cast<unsigned int>(st->best >> 32)) continue;
    const unsigned int top = float_to_ordered(top_scores[jobs[lf.job].cand_off + lf.cand]);
    if (stage == 0) {
      atomicMax(&st->group_best, top);
      continue;
    }
    if (top != st->group_best) continue;
    const unsigned long long key = (static_cast<unsigned long long>(lf.cand) << 32) | lf.path;
    if (stage == 1) {
      atomicMin(&st->win, key);
      atomicMax(&st->cand_max, lf.cand);
    } else if (key == st->win) {
      results[lf.job].scan_index = lf.scan;
      results[lf.job].x_index_offset = lf.x;
      results[lf.job].y_index_offset = lf.y;
    }
  }
}

__global__ void __launch_bounds__(128)
k_finalize_paths(const JobDev* __restrict__ jobs, const JobState* __restrict__ state,
                 csm_result2d* __restrict__ results, int num_jobs) {
  const int job = blockIdx.x * blockDim.x + threadIdx.x;
  if (job >= num_jobs) return;
  const JobState st = state[job];
  csm_result2d r;
  r.found = 0;
  r.scan_index = r.x_index_offset = r.y_index_offset = 0;  // Candidate2D(0, 0, 0) (FCSM2D.cc:345)
  r.score = jobs[job].min_score;
  r.tie_resolution = 0;
  r.candidates_scored = st.scored;
  if ((st.best >> 32) != 0ull) {
    r.score = ordered_to_float(static_cast<unsigned int>(st.best >> 32));
    if (st.win == ~0ull) {
      r.found = -1;  // internal inconsistency: no recorded leaf attains the best score
    } else {
      r.found = 1;
      r.tie_resolution = st.cand_max != static_cast<unsigned int>(st.win >> 32) ? 2 : 1;  // 2: host replay
    }
  }
  results[job] = r;
}

__global__ void __launch_bounds__(128)
k_finalize(const JobDev* __restrict__ jobs, const JobState* __restrict__ state,
           const Contender* __restrict__ list, csm_result2d* __restrict__ results, int num_jobs) {
  const int job = blockIdx.x * blockDim.x + threadIdx.x;
  if (job >= num_jobs) return;
  const JobState st = state[job];
  csm_result2d r;
  r.found = 0;
  r.scan_index = r.x_index_offset = r.y_index_offset = 0;  // Candidate2D(0, 0, 0) (FCSM2D.cc:345)
  r.score = jobs[job].min_score;
  r.tie_resolution = 0;
  r.candidates_scored = st.scored;
  if ((st.best >> 32) != 0ull) {
    r.score = ordered_to_float(static_cast<unsigned int>(st.best >> 32));
    if (st.win == ~0ull) {
      r.found = -1;  // internal inconsistency: no contender holds the best leaf
    } else {
      const Contender ct = list[st.win & 0xFFFFFFFFull];
      r.found = 1;
      r.scan_index = ct.scan;
      r.x_index_offset = ct.leaf_x;
      r.y_index_offset = ct.leaf_y;
      r.tie_resolution = static_cast<int32_t>(st.tie_count);  // > 1: host must replay std::sort
    }
  }
  results[job] = r;
}

}  // namespace csm

// ===========================================================================
// host side
// ===========================================================================
using namespace csm;

namespace {

size_t align_up(size_t v) { return (v + 255) & ~static_cast<size_t>(255); }

void free_batch_device(csm_batch2d* b) {
  if (b->arena) {
    if (b->owner) b->owner->pool.give(false, b->arena, b->arena_cap);
    else cudaFree(b->arena);
  }
  b->arena = nullptr;
}

// Carves all device arrays of the batch out of one allocation.
int allocate_batch(csm_batch2d* b, unsigned int contender_cap) {
  free_batch_device(b);
  b->contender_cap = contender_cap;
  size_t off = 0;
  auto take = [&off](size_t bytes) {
    const size_t at = off;
    off += align_up(bytes);
    return at;
  };
  const size_t o_jobs = take(sizeof(JobDev) * b->num_jobs);
  const size_t o_pts = take(sizeof(float) * 3 * b->total_pts);
  const size_t o_quat = take(sizeof(float) * 2 * b->total_scans);
  const size_t o_xy = take(sizeof(int2) * b->total_xy);
  const size_t o_scans = take(sizeof(ScanInfo) * b->total_scans);
  const size_t o_sbest = take(sizeof(unsigned long long) * b->total_scans);
  const size_t o_top = take(sizeof(float) * b->total_cand);
  const size_t o_state = take(sizeof(JobState) * b->num_jobs);
  const bool dfs = b->owner->phase2_dfs;
  const size_t o_cont = take(dfs ? sizeof(Contender) * contender_cap : 0);
  const size_t o_leaf = take(dfs ? 0 : sizeof(Leaf2) * contender_cap);
  const size_t o_defer = take(dfs ? 0 : sizeof(Deferred2) * contender_cap);
  const size_t o_cnt = take(sizeof(unsigned int) * kNumCounters);
  const size_t o_res = take(sizeof(csm_result2d) * b->num_jobs);
  const size_t o_hist = take(sizeof(unsigned long long) * CSM_MAX_DEPTH * 33);
  const size_t o_dstats = take(sizeof(double) * CSM_MAX_DEPTH * 6);
  // per bucket (= scan of a job): owner, counter, list offset, work-item offset
  const size_t o_bjs = take(sizeof(unsigned int) * b->total_scans);
  const size_t o_bcnt = take(sizeof(unsigned int) * b->total_scans);
  // offsets | total | one base per k_bucket_prefix chunk
  const size_t prefix_len = static_cast<size_t>(b->total_scans) + 1 + b->total_scans / 8192 + 2;
  const size_t o_bstart = take(sizeof(unsigned int) * prefix_len);
  const size_t o_bistart = take(sizeof(unsigned int) * prefix_len);
  if (b->owner->pool.take(false, off, &b->arena, &b->arena_cap) != CSM_OK) {
    b->arena = nullptr;
    set_error("allocation of a %zu-byte batch arena failed: %s", off, cudaGetErrorString(cudaGetLastError()));
    return CSM_ERR_CUDA;
  }
  b->arena_bytes = off;
  char* base = static_cast<char*>(b->arena);
  b->jobs = reinterpret_cast<JobDev*>(base + o_jobs);
  b->points = reinterpret_cast<float*>(base + o_pts);
  b->quats = reinterpret_cast<float*>(base + o_quat);
  b->xy = reinterpret_cast<int2*>(base + o_xy);
  b->scans = reinterpret_cast<ScanInfo*>(base + o_scans);
  b->scan_best = reinterpret_cast<unsigned long long*>(base + o_sbest);
  b->top_scores = reinterpret_cast<float*>(base + o_top);
  b->state = reinterpret_cast<JobState*>(base + o_state);
  b->contenders = reinterpret_cast<Contender*>(base + o_cont);
  b->leaves = reinterpret_cast<Leaf2*>(base + o_leaf);
  b->deferred = reinterpret_cast<Deferred2*>(base + o_defer);
  b->counters = reinterpret_cast<unsigned int*>(base + o_cnt);
  b->results_dev = reinterpret_cast<csm_result2d*>(base + o_res);
  b->hist = reinterpret_cast<unsigned long long*>(base + o_hist);
  b->depth_stats = reinterpret_cast<double*>(base + o_dstats);
  b->bucket_js = reinterpret_cast<unsigned int*>(base + o_bjs);
  b->bucket_cnt = reinterpret_cast<unsigned int*>(base + o_bcnt);
  b->bucket_start = reinterpret_cast<unsigned int*>(base + o_bstart);
  b->bucket_istart = reinterpret_cast<unsigned int*>(base + o_bistart);
  return CSM_OK;
}

// The batch's pinned block: [jobs | points | quats | results | counters].
csm_result2d* pinned_results(const csm_batch2d* b) {
  const size_t s_jobs = align_up(sizeof(JobDev) * b->num_jobs);
  const size_t s_pts = align_up(sizeof(float) * 3 * b->total_pts);
  const size_t s_quat = align_up(sizeof(float) * 2 * b->total_scans);
  return reinterpret_cast<csm_result2d*>(static_cast<char*>(b->pinned) + s_jobs + s_pts + s_quat);
}
unsigned int* pinned_counters(const csm_batch2d* b) {
  return reinterpret_cast<unsigned int*>(reinterpret_cast<char*>(pinned_results(b)) +
                                         align_up(sizeof(csm_result2d) * b->num_jobs));
}

// Host -> device copy of job descriptors, points and quaternion tables through the
// batch's pinned staging buffer (filled by csm_batch2d_upload).
int copy_inputs(csm_ctx* ctx, csm_batch2d* b) {
  char* pin = static_cast<char*>(b->pinned);
  const size_t s_jobs = align_up(sizeof(JobDev) * b->num_jobs);
  const size_t s_pts = align_up(sizeof(float) * 3 * b->total_pts);
  const size_t s_quat = align_up(sizeof(float) * 2 * b->total_scans);
  memcpy(pin, b->jobs_host.data(), sizeof(JobDev) * b->num_jobs);
  // jobs, points and quats are adjacent and identically aligned on both sides
  CSM_CUDA(cudaMemcpyAsync(b->jobs, pin, s_jobs + s_pts + s_quat, cudaMemcpyHostToDevice,
                           ctx->stream));
  return CSM_OK;
}

}  // namespace

extern "C" {

void* csm_ctx_stream(csm_ctx* ctx) { return ctx ? static_cast<void*>(ctx->stream) : nullptr; }

int csm_ctx_set_profiling(csm_ctx* ctx, int enabled) {
  CSM_REQUIRE(ctx != nullptr, "ctx must not be NULL");
  ctx->profiling = enabled;
  for (int k = 0; k < csm_ctx::kPhases; ++k) ctx->phase_ms[k] = ctx->phase_scored[k] = 0.;
  for (double& v : ctx->tile_hist) v = 0.;
  for (double& v : ctx->depth_profile) v = 0.;
  ctx->profiled_runs = 0;
  return CSM_OK;
}

int csm_ctx_get_depth_profile(csm_ctx* ctx, double* out) {
  CSM_REQUIRE(ctx != nullptr && out != nullptr, "NULL argument");
  for (int k = 0; k < CSM_MAX_DEPTH * csm_ctx::kDepthFields; ++k) out[k] = ctx->depth_profile[k];
  return CSM_OK;
}

int csm_ctx_get_profile(csm_ctx* ctx, double* phase_ms5, double* phase_scored5, int* runs) {
  CSM_REQUIRE(ctx != nullptr && phase_ms5 != nullptr && phase_scored5 != nullptr, "NULL argument");
  for (int k = 0; k < csm_ctx::kPhases; ++k) {
    phase_ms5[k] = ctx->phase_ms[k];
    phase_scored5[k] = ctx->phase_scored[k];
  }
  if (runs) *runs = ctx->profiled_runs;
  return CSM_OK;
}

int csm_ctx_get_tile_histogram(csm_ctx* ctx, double* hist) {
  CSM_REQUIRE(ctx != nullptr && hist != nullptr, "NULL argument");
  for (int k = 0; k < CSM_MAX_DEPTH * 33; ++k) hist[k] = ctx->tile_hist[k];
  return CSM_OK;
}

void csm_batch2d_destroy(csm_ctx* ctx, csm_batch2d* b) {
  if (!b) return;
  if (ctx) {
    cudaSetDevice(ctx->device);
    // A batch that was fetched after its last run has nothing in flight: its blocks can go
    // back to the pool while later batches are still queued on the stream.
    if (!b->idle) cudaStreamSynchronize(ctx->stream);
  }
  free_batch_device(b);
  for (cudaEvent_t e : b->prof_events) cudaEventDestroy(e);
  for (cudaEvent_t e : b->depth_events) cudaEventDestroy(e);
  if (b->done) cudaEventDestroy(b->done);
  if (b->pinned) {
    if (b->owner) b->owner->pool.give(true, b->pinned, b->pinned_cap);
    else cudaFreeHost(b->pinned);
  }
  delete b;
}

int csm_batch2d_upload(csm_ctx* ctx, const csm_job2d* jobs, int num_jobs, csm_batch2d** out) {
  CSM_REQUIRE(ctx != nullptr && out != nullptr, "ctx/out must not be NULL");
  CSM_REQUIRE(num_jobs >= 1 && num_jobs <= 65535, "num_jobs must be in [1, 65535]");
  CSM_REQUIRE(jobs != nullptr, "jobs must not be NULL");
  *out = nullptr;
  CSM_CUDA(cudaSetDevice(ctx->device));
  csm_batch2d* b = new csm_batch2d;
  b->num_jobs = num_jobs;
  b->jobs_host.resize(num_jobs);
  b->empty_job.assign(num_jobs, 0);
  for (int i = 0; i < num_jobs; ++i) {
    const csm_job2d& in = jobs[i];
    JobDev& j = b->jobs_host[i];
    memset(&j, 0, sizeof(j));
    if (!(in.stack && in.stack->device == ctx->device && in.num_points >= 0 && in.num_scans >= 1 &&
          in.num_scans <= 65535 && in.resolution > 0. && in.linear_bound >= 0 &&
          (in.num_points == 0 || in.points_xyz) && in.scan_quat)) {
      set_error("invalid job %d (stack/points/scan table/limits)", i);
      delete b;
      return CSM_ERR_ARG;
    }
    const csm_stack2d* s = in.stack;
    j.sd = s->dev;
    j.num_x = s->num_x;
    j.num_y = s->num_y;
    j.depth = s->depth;
    j.n = in.num_points;
    j.S = in.num_scans;
    j.lin = std::min(in.linear_bound, 1 << 28);
    j.q0w = in.init_quat_w;
    j.q0s = in.init_quat_s;
    j.tx = in.init_tx;
    j.ty = in.init_ty;
    j.max_x = in.max_x;
    j.max_y = in.max_y;
    j.res = in.resolution;
    j.min_score = in.min_score;
    j.score_offset = in.score_offset;
    j.score_scale = in.score_scale;
    j.n_float = static_cast<float>(in.num_points);
    if (in.num_points == 0) {  // NaN scores in the reference: never above min_score
      b->empty_job[i] = 1;
      j.S = 0;
    }
    // Upper bound of the lowest-resolution lattice (FCSM2D.cc:275-283): after
    // ShrinkToFit the window spans at most num_cells - 1 + scan extent cells.
    double r2 = 0.;
    for (int p = 0; p < in.num_points; ++p) {
      const double px = in.points_xyz[3 * p], py = in.points_xyz[3 * p + 1];
      r2 = std::max(r2, px * px + py * py);
    }
    const long long ext = 2 * static_cast<long long>(std::ceil(std::sqrt(r2) / in.resolution)) + 4;
    const long long step = 1ll << (s->depth - 1);
    // ShrinkToFit seeds its bounds at 0 (CSM2D.cc:78-79): the window of a scan whose cells span
    // [a, b] is max(0, W - 1 - a) + max(0, b), clamped by the linear bound.  With the scan
    // centred on cell c (the initial translation) and reaching e cells, a >= c - e and
    // b <= c + e -- which also covers a scan that lies wholly outside the grid.
    const double e = 0.5 * static_cast<double>(ext) + 2.;
    const double cx = (in.max_y - static_cast<double>(in.init_ty)) / in.resolution;
    const double cy = (in.max_x - static_cast<double>(in.init_tx)) / in.resolution;
    auto window = [&](double c, int cells) {
      const double w = std::max(0., cells - 1 - c + e) + std::max(0., c + e);
      return static_cast<long long>(std::min(2. * j.lin, std::ceil(w)));
    };
    const long long wx = window(cx, s->num_x), wy = window(cy, s->num_y);
    const long long nx = wx / step + 1, ny = wy / step + 1;
    if (std::max(wx, wy) + ext > 32000) {
      set_error("job %d: search window of %lld cells exceeds the 16-bit offset range", i,
                std::max(wx, wy));
      delete b;
      return CSM_ERR_ARG;
    }
    j.pts_off = b->total_pts;
    j.quat_off = b->total_scans;
    j.scan_off = b->total_scans;
    j.xy_off = b->total_xy;
    j.cand_off = b->total_cand;
    j.cand_cap = static_cast<long long>(j.S) * nx * ny;
    b->total_pts += j.n;
    b->total_scans += j.S;
    b->total_xy += static_cast<long long>(j.S) * j.n;
    b->total_cand += j.cand_cap;
    b->max_scans = std::max(b->max_scans, j.S);
  }
  // pinned staging: [jobs | points | quats | results]
  const size_t s_jobs = align_up(sizeof(JobDev) * num_jobs);
  const size_t s_pts = align_up(sizeof(float) * 3 * b->total_pts);
  const size_t s_quat = align_up(sizeof(float) * 2 * b->total_scans);
  const size_t s_res = align_up(sizeof(csm_result2d) * num_jobs);
  b->pinned_bytes = s_jobs + s_pts + s_quat + s_res + align_up(sizeof(unsigned int) * kNumCounters);
  b->owner = ctx;
  if (ctx->pool.take(true, b->pinned_bytes, &b->pinned, &b->pinned_cap) != CSM_OK) {
    set_error("pinned allocation of %zu bytes failed: %s", b->pinned_bytes,
              cudaGetErrorString(cudaGetLastError()));
    b->pinned = nullptr;
    delete b;
    return CSM_ERR_CUDA;
  }
  char* pin = static_cast<char*>(b->pinned);
  float* ppts = reinterpret_cast<float*>(pin + s_jobs);
  float* pquat = reinterpret_cast<float*>(pin + s_jobs + s_pts);
  for (int i = 0; i < num_jobs; ++i) {
    const JobDev& j = b->jobs_host[i];
    if (j.n > 0) memcpy(ppts + 3 * j.pts_off, jobs[i].points_xyz, sizeof(float) * 3 * j.n);
    if (j.S > 0) memcpy(pquat + 2 * j.quat_off, jobs[i].scan_quat, sizeof(float) * 2 * j.S);
  }
  // Tiles per frontier buffer; an overflow regrows it and replays the batch (fetch).
  b->frontier_want = static_cast<unsigned long long>(ctx->frontier_per_candidate * static_cast<double>(b->total_cand)) +
                     (1ull << 20);
  const double want_contenders = ctx->contenders_per_candidate * static_cast<double>(b->total_cand);
  int rc = allocate_batch(b, static_cast<unsigned int>(std::min(4.0e9, std::max(1048576., want_contenders))));
  if (rc == CSM_OK) rc = copy_inputs(ctx, b);
  if (rc != CSM_OK) {
    csm_batch2d_destroy(nullptr, b);
    return rc;
  }
  b->uploaded = true;
  *out = b;
  return CSM_OK;
}

int csm_batch2d_run(csm_ctx* ctx, csm_batch2d* b) {
  CSM_REQUIRE(ctx != nullptr && b != nullptr && b->uploaded, "batch was not uploaded");
  CSM_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t st = ctx->stream;
  // Scratch lists of the sweep, owned by the context (runs are stream-ordered).
  //   node lists: frontier[0] = bucket-sorted nodes of the level being expanded,
  //               frontier[1] = survivors in arrival order; aux = blocks | work items
  //   tile lists (CSM_SWEEP=tiles): two ping-pong tile lists
  const bool tiles = ctx->sweep_tiles;
  const unsigned long long nbuckets = static_cast<unsigned long long>(b->total_scans);
  const unsigned long long node_cap = 2ull * b->frontier_want;
  const unsigned long long block_cap = node_cap / 16ull + 8ull * nbuckets + 4096ull;
  const unsigned long long item_cap = node_cap / 16ull + nbuckets + 4096ull;
  const bool paths = !tiles && !ctx->phase2_dfs;  // every node carries its sibling-rank path
  const size_t want_list =
      tiles ? sizeof(Tile) * b->frontier_want : sizeof(unsigned int) * node_cap * (paths ? 2 : 1);
  const size_t o_items = align_up(sizeof(NodeBlock) * block_cap);
  const size_t want_aux = tiles ? 0 : o_items + sizeof(uint2) * item_cap;
  if (ctx->frontier[0].cap < want_list || ctx->frontier[1].cap < want_list ||
      ctx->frontier_aux.cap < want_aux) {
    CSM_CUDA(cudaStreamSynchronize(st));  // earlier batches may still use the old buffers
    int rc = ctx->frontier[0].reserve(want_list);
    if (rc == CSM_OK) rc = ctx->frontier[1].reserve(want_list);
    if (rc == CSM_OK && want_aux) rc = ctx->frontier_aux.reserve(want_aux);
    if (rc != CSM_OK) return rc;
  }
  Tile* const fbuf[2] = {ctx->frontier[0].as<Tile>(), ctx->frontier[1].as<Tile>()};
  const unsigned long long fcap = b->frontier_want;
  unsigned int* const node_list = ctx->frontier[0].as<unsigned int>();
  unsigned int* const node_tmp = ctx->frontier[1].as<unsigned int>();
  unsigned int* const node_path = paths ? node_list + node_cap : nullptr;
  unsigned int* const tmp_path = paths ? node_tmp + node_cap : nullptr;
  Leaf2* const leaves = paths ? b->leaves : nullptr;
  Deferred2* const deferred = paths ? b->deferred : nullptr;
  const unsigned int record_cap = b->contender_cap;
  NodeBlock* const node_blocks = ctx->frontier_aux.as<NodeBlock>();
  uint2* const node_items =
      want_aux ? reinterpret_cast<uint2*>(ctx->frontier_aux.as<char>() + o_items) : nullptr;
  CSM_CUDA(cudaMemsetAsync(b->counters, 0, sizeof(unsigned int) * kNumCounters, st));
  if (ctx->profiling >= 2) {
    CSM_CUDA(cudaMemsetAsync(b->hist, 0, sizeof(unsigned long long) * CSM_MAX_DEPTH * 33, st));
    CSM_CUDA(cudaMemsetAsync(b->depth_stats, 0, sizeof(double) * CSM_MAX_DEPTH * 6, st));
  }
  for (cudaEvent_t e : b->depth_events) cudaEventDestroy(e);
  b->depth_events.clear();
  b->depth_of_events.clear();
  auto mark_depth = [&](int d, bool first) {
    if (!ctx->profiling) return;
    cudaEvent_t e;
    cudaEventCreate(&e);
    cudaEventRecord(e, st);
    b->depth_events.push_back(e);
    if (first) b->depth_of_events.push_back(d);
  };
  const int nj = b->num_jobs;
  for (cudaEvent_t e : b->prof_events) cudaEventDestroy(e);
  b->prof_events.clear();
  auto mark = [&]() {
    if (!ctx->profiling) return;
    cudaEvent_t e;
    cudaEventCreate(&e);
    cudaEventRecord(e, st);
    b->prof_events.push_back(e);
  };
  mark();
  if (b->max_scans > 0) {
    const dim3 grid_sj(b->max_scans, nj);
    ++ctx->launches, k_discretize<<<grid_sj, 128, 0, st>>>(b->jobs, b->points, b->quats, b->xy, b->scans,
                                          b->scan_best, b->bucket_js, 1);
    if (!tiles) CSM_CUDA(cudaMemsetAsync(b->bucket_cnt, 0, sizeof(unsigned int) * b->total_scans, st));
    ++ctx->launches, k_scan_prefix<<<nj, 256, 0, st>>>(b->jobs, b->scans, b->state, b->counters);
    mark();
    ++ctx->launches, k_score_top<<<grid_sj, kTopWarps * kWarp, 0, st>>>(b->jobs, b->xy, b->scans, b->top_scores,
                                                         b->state, b->scan_best);
    mark();
    ++ctx->launches, k_dive<<<dim3((b->max_scans + kDiveWarps - 1) / kDiveWarps, nj), kDiveWarps * kWarp, 0, st>>>(
        b->jobs, b->xy, b->scans, b->scan_best, b->state);
    mark();
    int max_depth = 0;
    for (const JobDev& j : b->jobs_host) max_depth = std::max(max_depth, j.depth);
    // Frontier slot d holds the nodes at depth d of every job (buffer d & 1).  Jobs
    // whose stack is shallower than the deepest one seed their own top depth just
    // before that depth is expanded.  Depth-1 stacks need no sweep: their
    // lowest-resolution candidates are leaves and k_dive already offered the best.
    for (int d = max_depth - 1; d >= 1; --d) {
      bool any = false;
      for (const JobDev& j : b->jobs_host) any = any || (j.depth - 1 == d && j.S > 0);
      unsigned long long* const hist = ctx->profiling >= 2 ? b->hist : nullptr;
      if (tiles) {
        if (any)
          ++ctx->launches, k_seed<<<grid_sj, kSeedWarps * kWarp, 0, st>>>(
              b->jobs, b->scans, b->top_scores, b->state, fbuf[d & 1], b->counters, d, fcap);
        ++ctx->launches, k_expand<<<ctx->num_sms * 8, kExpandWarps * kWarp, 0, st>>>(
            b->jobs, b->xy, b->scans, b->top_scores, b->state, fbuf[d & 1], fbuf[(d - 1) & 1],
            b->counters, d, d - 1, d, fcap, hist);
        continue;
      }
      // Survivors of depth d (children of depth d + 1 and, for stacks whose top is d,
      // the seeds) are gathered into bucket order, then expanded.
      mark_depth(d, true);
      if (any)
        ++ctx->launches, k_seed_nodes<<<grid_sj, kSeedWarps * kWarp, 0, st>>>(
            b->jobs, b->scans, b->top_scores, b->state, node_tmp, node_blocks, b->bucket_cnt,
            b->counters, d, node_cap, block_cap, tmp_path, deferred, record_cap);
      ++ctx->launches, k_bucket_prefix<<<static_cast<unsigned int>((nbuckets + kPrefixChunk - 1) / kPrefixChunk), 1024, 0, st>>>(
          b->bucket_cnt, b->bucket_start, b->bucket_istart, b->counters, d,
          static_cast<unsigned int>(nbuckets), node_cap, item_cap);
      ++ctx->launches, k_place<<<ctx->num_sms * 8, 256, 0, st>>>(
          node_tmp, node_blocks, b->bucket_start, b->bucket_istart, node_list, node_items,
          b->counters, d, static_cast<unsigned int>(nbuckets), node_cap, block_cap, item_cap, tmp_path, node_path);
      mark_depth(d, false);
      ++ctx->launches, (ctx->expand_noalloc ? k_expand_nodes<true> : k_expand_nodes<false>)<<<ctx->num_sms * kNodeBlocksPerSm, kExpandWarps * kWarp, 0, st>>>(
          b->jobs, b->xy, b->scans, b->top_scores, b->state, b->bucket_js, node_list, node_items,
          node_tmp, node_blocks, b->bucket_cnt, b->counters, d, node_cap, block_cap, item_cap, hist, node_path,
          tmp_path, leaves, deferred, record_cap);
      mark_depth(d, false);
      if (ctx->profiling >= 2)
        ++ctx->launches, k_item_stats<<<ctx->num_sms * 4, 256, 0, st>>>(
            b->jobs, b->bucket_js, node_list, node_items, b->counters, d, item_cap, b->depth_stats);
    }
    mark();
    if (paths) {
      bool any_depth1 = false;
      for (const JobDev& j : b->jobs_host) any_depth1 = any_depth1 || (j.depth == 1 && j.S > 0);
      ++ctx->launches, k_deferred_dfs<<<ctx->num_sms * 4, 128, 0, st>>>(b->jobs, b->xy, b->scans, b->state,
                                                                    b->deferred, b->leaves, b->counters, record_cap);
      if (any_depth1)
        ++ctx->launches, k_leaves_depth1<<<grid_sj, 256, 0, st>>>(b->jobs, b->scans, b->top_scores, b->state,
                                                                b->leaves, b->counters, record_cap);
    } else {
      ++ctx->launches, k_collect<<<grid_sj, 256, 0, st>>>(b->jobs, b->scans, b->top_scores, b->state, b->contenders,
                                         b->counters, b->contender_cap);
      ++ctx->launches, k_ordered_dfs<<<ctx->num_sms * 8, 128, 0, st>>>(b->jobs, b->xy, b->scans, b->state,
                                                      b->contenders, b->counters, b->contender_cap);
      ++ctx->launches, k_pick<<<(b->contender_cap + 255) / 256, 256, 0, st>>>(b->state, b->contenders, b->counters,
                                                             b->contender_cap);
    }
  } else {
    CSM_CUDA(cudaMemsetAsync(b->state, 0, sizeof(JobState) * nj, st));
    for (int k = 0; k < 4; ++k) mark();
  }
  if (paths) {
    // The number of recorded leaves is known on the device only; the grids cover a modest count
    // and the rest is strided over (ties of whole worlds are the rare case).
    const bool ran = b->max_scans > 0;
    const unsigned int leaf_blocks = std::min(record_cap, 1u << 18) / 256 + 1;
    for (int stage = 0; ran && stage < 2; ++stage)
      ++ctx->launches, k_pick_leaf<<<leaf_blocks, 256, 0, st>>>(b->jobs, b->top_scores, b->state, b->leaves,
                                                              b->counters, record_cap, stage, b->results_dev);
    ++ctx->launches, k_finalize_paths<<<(nj + 127) / 128, 128, 0, st>>>(b->jobs, b->state, b->results_dev, nj);
    if (ran)
      ++ctx->launches, k_pick_leaf<<<leaf_blocks, 256, 0, st>>>(b->jobs, b->top_scores, b->state, b->leaves,
                                                              b->counters, record_cap, 2, b->results_dev);
  } else {
    ++ctx->launches, k_finalize<<<(nj + 127) / 128, 128, 0, st>>>(b->jobs, b->state, b->contenders, b->results_dev, nj);
  }
  mark();
  CSM_CUDA(cudaGetLastError());
  // Results and counters follow the kernels into the pinned block; fetch() waits for this
  // batch's own event only, so a caller can keep later batches queued behind it.
  CSM_CUDA(cudaMemcpyAsync(pinned_results(b), b->results_dev, sizeof(csm_result2d) * b->num_jobs,
                           cudaMemcpyDeviceToHost, st));
  CSM_CUDA(cudaMemcpyAsync(pinned_counters(b), b->counters, sizeof(unsigned int) * kNumCounters,
                           cudaMemcpyDeviceToHost, st));
  if (!b->done) CSM_CUDA(cudaEventCreateWithFlags(&b->done, cudaEventDisableTiming));
  CSM_CUDA(cudaEventRecord(b->done, st));
  b->ran = true;
  b->idle = false;
  return CSM_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------
// fetch: results back, capacity regrowth, top-level tie replay
// ---------------------------------------------------------------------------
namespace {

// The reference sorts all lowest-resolution candidates with a non-stable std::sort
// (FCSM2D.cc:331-332), so among equal scores the visiting order is whatever
// libstdc++'s introsort produces.  std::sort's control flow depends only on the
// comparison results, so sorting (score, generation index) pairs with the same
// comparator reproduces the reference's permutation exactly.
struct ScoreIdx {
  float score;
  unsigned int idx;
  bool operator>(const ScoreIdx& o) const { return score > o.score; }
};

int resolve_tie(csm_ctx* ctx, csm_batch2d* b, int job, const std::vector<Contender>& list,
                const std::vector<Leaf2>& leaves, csm_result2d* r) {
  const JobDev& j = b->jobs_host[job];
  JobState st;
  CSM_CUDA(cudaMemcpyAsync(&st, b->state + job, sizeof(st), cudaMemcpyDeviceToHost, ctx->stream));
  CSM_CUDA(cudaStreamSynchronize(ctx->stream));
  std::vector<float> scores(st.total);
  CSM_CUDA(cudaMemcpyAsync(scores.data(), b->top_scores + j.cand_off, sizeof(float) * st.total,
                           cudaMemcpyDeviceToHost, ctx->stream));
  CSM_CUDA(cudaStreamSynchronize(ctx->stream));
  std::vector<ScoreIdx> v(st.total);
  for (unsigned int i = 0; i < st.total; ++i) v[i] = ScoreIdx{scores[i], i};
  std::sort(v.begin(), v.end(), std::greater<ScoreIdx>());
  std::vector<unsigned int> rank(st.total);
  for (unsigned int i = 0; i < st.total; ++i) rank[v[i].idx] = i;
  if (!ctx->phase2_dfs) {
    // Path mode: among the leaves attaining S* whose ancestor has the group's top-level score, the
    // reference reaches first the one whose ancestor sorts first, then the smallest path.
    const Leaf2* win = nullptr;
    for (const Leaf2& lf : leaves) {
      if (lf.job != job || float_to_ordered(lf.score) != float_to_ordered(r->score) ||
          float_to_ordered(scores[lf.cand]) != st.group_best)
        continue;
      if (!win || rank[lf.cand] < rank[win->cand] || (lf.cand == win->cand && lf.path < win->path)) win = &lf;
    }
    if (!win) {
      set_error("tie replay found no leaf for job %d", job);
      return CSM_ERR_CUDA;
    }
    r->scan_index = win->scan;
    r->x_index_offset = win->x;
    r->y_index_offset = win->y;
    r->tie_resolution = 1;
    return CSM_OK;
  }
  const Contender* best = nullptr;
  for (const Contender& c : list) {
    if (c.job != job || !c.has_leaf || float_to_ordered(c.score) != st.group_best) continue;
    if (!best || rank[c.cand] < rank[best->cand]) best = &c;
  }
  if (!best) {
    set_error("tie replay found no contender for job %d", job);
    return CSM_ERR_CUDA;
  }
  r->scan_index = best->scan;
  r->x_index_offset = best->leaf_x;
  r->y_index_offset = best->leaf_y;
  r->tie_resolution = 1;
  return CSM_OK;
}

}  // namespace

extern "C" {

int csm_batch2d_fetch(csm_ctx* ctx, csm_batch2d* b, csm_result2d* results) {
  CSM_REQUIRE(ctx != nullptr && b != nullptr && results != nullptr, "NULL argument");
  CSM_REQUIRE(b->ran, "batch was not run");
  CSM_CUDA(cudaSetDevice(ctx->device));
  csm_result2d* pin_res = pinned_results(b);
  const unsigned int* counters = pinned_counters(b);
  for (int attempt = 0;; ++attempt) {
    CSM_CUDA(cudaEventSynchronize(b->done));
    const unsigned int ov = counters[kCntOverflow];
    if (ov == 0) break;
    if ((ov & 2u) || attempt >= 4) {
      set_error("work list overflow (code %u) after %d regrowths", ov, attempt);
      return CSM_ERR_CAPACITY;
    }
    // A work list overflowed: regrow it and replay the whole batch.
    if (ov & 1u) {
      b->frontier_want *= 4ull;
      ctx->frontier_per_candidate = std::max(
          ctx->frontier_per_candidate, static_cast<double>(b->frontier_want) / std::max<double>(1., b->total_cand));
    }
    const unsigned int records = ctx->phase2_dfs ? counters[kCntContenders]
                                                 : std::max(counters[kCntLeaves], counters[kCntDeferred]);
    unsigned long long ccap = (ov & 4u) ? static_cast<unsigned long long>(records) + 1024ull : b->contender_cap;
    if (ov & 4u)
      ctx->contenders_per_candidate = std::max(
          ctx->contenders_per_candidate, 1.5 * static_cast<double>(ccap) / std::max<double>(1., b->total_cand));
    if (ccap > 0xFFFFFFF0ull) ccap = 0xFFFFFFF0ull;
    int rc = allocate_batch(b, static_cast<unsigned int>(ccap));
    if (rc == CSM_OK) rc = copy_inputs(ctx, b);
    if (rc == CSM_OK) rc = csm_batch2d_run(ctx, b);
    if (rc != CSM_OK) return rc;
  }
  if (ctx->profiling) {
    if (static_cast<int>(b->prof_events.size()) == csm_ctx::kPhases + 1) {
      const cudaEvent_t* ev = b->prof_events.data();
      for (int k = 0; k < csm_ctx::kPhases; ++k) {
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, ev[k], ev[k + 1]) == cudaSuccess) ctx->phase_ms[k] += ms;
      }
      std::vector<JobState> states(b->num_jobs);
      CSM_CUDA(cudaMemcpyAsync(states.data(), b->state, sizeof(JobState) * b->num_jobs,
                               cudaMemcpyDeviceToHost, ctx->stream));
      CSM_CUDA(cudaStreamSynchronize(ctx->stream));
      for (const JobState& s : states) {
        ctx->phase_scored[1] += static_cast<double>(s.scored_phase[0]);
        ctx->phase_scored[2] += static_cast<double>(s.scored_phase[1]);
        ctx->phase_scored[3] += static_cast<double>(s.scored_phase[2]);
        ctx->phase_scored[4] += static_cast<double>(s.scored_phase[3]);
      }
      for (size_t k = 0; k < b->depth_of_events.size() && 3 * k + 2 < b->depth_events.size(); ++k) {
        const int d = b->depth_of_events[k];
        float book = 0.f, expand = 0.f;
        cudaEventElapsedTime(&book, b->depth_events[3 * k], b->depth_events[3 * k + 1]);
        cudaEventElapsedTime(&expand, b->depth_events[3 * k + 1], b->depth_events[3 * k + 2]);
        ctx->depth_profile[d * csm_ctx::kDepthFields + 0] += expand;
        ctx->depth_profile[d * csm_ctx::kDepthFields + 1] += book;
      }
      if (ctx->profiling >= 2) {
        std::vector<double> dstats(CSM_MAX_DEPTH * 6);
        CSM_CUDA(cudaMemcpyAsync(dstats.data(), b->depth_stats, sizeof(double) * dstats.size(),
                                 cudaMemcpyDeviceToHost, ctx->stream));
        CSM_CUDA(cudaStreamSynchronize(ctx->stream));
        for (int d = 0; d < CSM_MAX_DEPTH; ++d)
          for (int k = 0; k < 6; ++k) ctx->depth_profile[d * csm_ctx::kDepthFields + 2 + k] += dstats[d * 6 + k];
      }
      if (ctx->profiling >= 2) {
        std::vector<unsigned long long> hist(CSM_MAX_DEPTH * 33);
        CSM_CUDA(cudaMemcpyAsync(hist.data(), b->hist, sizeof(unsigned long long) * hist.size(),
                                 cudaMemcpyDeviceToHost, ctx->stream));
        CSM_CUDA(cudaStreamSynchronize(ctx->stream));
        for (size_t k = 0; k < hist.size(); ++k) ctx->tile_hist[k] += static_cast<double>(hist[k]);
      }
      ++ctx->profiled_runs;
    }
    for (cudaEvent_t e : b->prof_events) cudaEventDestroy(e);
    b->prof_events.clear();
    for (cudaEvent_t e : b->depth_events) cudaEventDestroy(e);
    b->depth_events.clear();
    b->depth_of_events.clear();
  }
  std::vector<Contender> list;
  std::vector<Leaf2> leaf_list;
  bool have_list = false;
  for (int i = 0; i < b->num_jobs; ++i) {
    csm_result2d r = pin_res[i];
    if (r.found < 0) {
      set_error("job %d: no contender holds the best leaf (internal error)", i);
      return CSM_ERR_CUDA;
    }
    if (r.found == 1 && r.tie_resolution > 1) {
      if (!have_list && ctx->phase2_dfs) {
        const unsigned int n = std::min(counters[kCntContenders], b->contender_cap);
        list.resize(n);
        CSM_CUDA(cudaMemcpyAsync(list.data(), b->contenders, sizeof(Contender) * n,
                                 cudaMemcpyDeviceToHost, ctx->stream));
        CSM_CUDA(cudaStreamSynchronize(ctx->stream));
        have_list = true;
      }
      if (!have_list && !ctx->phase2_dfs) {
        const unsigned int n = std::min(counters[kCntLeaves], b->contender_cap);
        leaf_list.resize(n);
        CSM_CUDA(cudaMemcpyAsync(leaf_list.data(), b->leaves, sizeof(Leaf2) * n, cudaMemcpyDeviceToHost,
                                 ctx->stream));
        CSM_CUDA(cudaStreamSynchronize(ctx->stream));
        have_list = true;
      }
      const int rc = resolve_tie(ctx, b, i, list, leaf_list, &r);
      if (rc != CSM_OK) return rc;
    } else {
      r.tie_resolution = 0;
    }
    results[i] = r;
  }
  b->idle = true;
  return CSM_OK;
}

int csm_match2d_batch(csm_ctx* ctx, const csm_job2d* jobs, int num_jobs, csm_result2d* results) {
  CSM_REQUIRE(ctx != nullptr && jobs != nullptr && results != nullptr, "NULL argument");
  CSM_REQUIRE(num_jobs >= 0, "num_jobs must be >= 0");
  // Bound the device footprint: split into chunks of at most 1024 jobs.
  for (int at = 0; at < num_jobs;) {
    const int n = std::min(num_jobs - at, 1024);
    csm_batch2d* b = nullptr;
    int rc = csm_batch2d_upload(ctx, jobs + at, n, &b);
    if (rc == CSM_OK) rc = csm_batch2d_run(ctx, b);
    if (rc == CSM_OK) rc = csm_batch2d_fetch(ctx, b, results + at);
    csm_batch2d_destroy(ctx, b);
    if (rc != CSM_OK) return rc;
    at += n;
  }
  return CSM_OK;
}

int csm_discretize2d(csm_ctx* ctx, const csm_job2d* job, int num_x, int num_y, int shrink,
                     int32_t* out_xy, int32_t* out_bounds) {
  CSM_REQUIRE(ctx != nullptr && job != nullptr && out_xy != nullptr && out_bounds != nullptr,
              "NULL argument");
  CSM_REQUIRE(job->num_points >= 1 && job->num_scans >= 1 && job->points_xyz && job->scan_quat,
              "job needs points and a scan table");
  CSM_CUDA(cudaSetDevice(ctx->device));
  JobDev j;
  memset(&j, 0, sizeof(j));
  j.num_x = num_x;
  j.num_y = num_y;
  j.depth = 1;
  j.n = job->num_points;
  j.S = job->num_scans;
  j.lin = job->linear_bound;
  j.q0w = job->init_quat_w;
  j.q0s = job->init_quat_s;
  j.tx = job->init_tx;
  j.ty = job->init_ty;
  j.max_x = job->max_x;
  j.max_y = job->max_y;
  j.res = job->resolution;
  const size_t b_job = align_up(sizeof(JobDev));
  const size_t b_pts = align_up(sizeof(float) * 3 * j.n);
  const size_t b_quat = align_up(sizeof(float) * 2 * j.S);
  const size_t b_xy = align_up(sizeof(int2) * static_cast<size_t>(j.S) * j.n);
  const size_t b_scan = align_up(sizeof(ScanInfo) * j.S);
  int rc = ctx->tmp1.reserve(b_job + b_pts + b_quat + b_xy + b_scan);
  if (rc != CSM_OK) return rc;
  char* base = ctx->tmp1.as<char>();
  JobDev* d_job = reinterpret_cast<JobDev*>(base);
  float* d_pts = reinterpret_cast<float*>(base + b_job);
  float* d_quat = reinterpret_cast<float*>(base + b_job + b_pts);
  int2* d_xy = reinterpret_cast<int2*>(base + b_job + b_pts + b_quat);
  ScanInfo* d_scan = reinterpret_cast<ScanInfo*>(base + b_job + b_pts + b_quat + b_xy);
  cudaStream_t st = ctx->stream;
  CSM_CUDA(cudaMemcpyAsync(d_job, &j, sizeof(j), cudaMemcpyHostToDevice, st));
  CSM_CUDA(cudaMemcpyAsync(d_pts, job->points_xyz, sizeof(float) * 3 * j.n, cudaMemcpyHostToDevice, st));
  CSM_CUDA(cudaMemcpyAsync(d_quat, job->scan_quat, sizeof(float) * 2 * j.S, cudaMemcpyHostToDevice, st));
  ++ctx->launches, k_discretize<<<dim3(j.S, 1), 128, 0, st>>>(d_job, d_pts, d_quat, d_xy, d_scan, nullptr, nullptr,
                                                              shrink);
  CSM_CUDA(cudaGetLastError());
  std::vector<ScanInfo> scans(j.S);
  CSM_CUDA(cudaMemcpyAsync(out_xy, d_xy, sizeof(int2) * static_cast<size_t>(j.S) * j.n,
                           cudaMemcpyDeviceToHost, st));
  CSM_CUDA(cudaMemcpyAsync(scans.data(), d_scan, sizeof(ScanInfo) * j.S, cudaMemcpyDeviceToHost, st));
  CSM_CUDA(cudaStreamSynchronize(st));
  for (int s = 0; s < j.S; ++s) {
    out_bounds[4 * s + 0] = scans[s].min_x;
    out_bounds[4 * s + 1] = scans[s].max_x;
    out_bounds[4 * s + 2] = scans[s].min_y;
    out_bounds[4 * s + 3] = scans[s].max_y;
  }
  return CSM_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------
// csm_discretize2d_noalloc: csm_discretize2d for callers that must not allocate (code under CUDA
// graph capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the
// reference.  Same kernel (k_discretize), but its
// scratch -- job descriptor, points, scan table, discrete scans, per-scan windows -- lives in
// ONE file-scope device array of fixed size instead of the context's growable buffer:
//   num_points <= 4096, num_scans <= 4096, num_scans * num_points <= 524 288
// (a C3 full-submap search is 1259 x 360 = 453 240).  Every GPU that loads the library holds the
// array (about 4.2 MiB) until exit.  All calls share it: nothing in it has to start at zero (the
// kernel writes every element it later copies out), but two calls that overlap on the GPU --
// different contexts, streams or host threads -- race.  The entry point calls no cudaMalloc,
// cudaFree, cudaHostAlloc or any other allocating routine, and allocates nothing on the host.
// ---------------------------------------------------------------------------
namespace {

constexpr int kNoAllocMaxPoints = 4096;
constexpr int kNoAllocMaxScans = 4096;
constexpr long long kNoAllocMaxCells = 1ll << 19;  // num_scans * num_points
constexpr size_t kNaJob = 256;                                                   // JobDev
constexpr size_t kNaPts = sizeof(float) * 3 * kNoAllocMaxPoints;                 // 48 KiB
constexpr size_t kNaQuat = sizeof(float) * 2 * kNoAllocMaxScans;                 // 32 KiB
constexpr size_t kNaXy = sizeof(int2) * static_cast<size_t>(kNoAllocMaxCells);   // 4 MiB
constexpr size_t kNaScan = sizeof(csm::ScanInfo) * kNoAllocMaxScans;             // 128 KiB
constexpr size_t kNaBytes = kNaJob + kNaPts + kNaQuat + kNaXy + kNaScan;
static_assert(sizeof(csm::JobDev) <= kNaJob, "job descriptor outgrew its slot");
static_assert(kNaJob % 256 == 0 && kNaPts % 256 == 0 && kNaQuat % 256 == 0 && kNaXy % 256 == 0, "slots stay aligned");

// In host code this name is a host-side stand-in, NOT a device address: the address the kernel
// gets comes from cudaGetSymbolAddress, once per device, with that device current.
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNaBytes];

constexpr int kNoAllocMaxDevices = 64;
unsigned char* g_noalloc_ptr[kNoAllocMaxDevices] = {nullptr};
std::mutex g_noalloc_mu;

// Device address of the scratch array on `device` (which must be the current device).
int noalloc_scratch(int device, unsigned char** out) {
  CSM_REQUIRE(device >= 0 && device < kNoAllocMaxDevices, "device index out of range");
  std::lock_guard<std::mutex> lock(g_noalloc_mu);
  if (g_noalloc_ptr[device] == nullptr) {
    void* p = nullptr;
    CSM_CUDA(cudaGetSymbolAddress(&p, g_noalloc_scratch));
    if (p == nullptr) {
      set_error("cudaGetSymbolAddress returned a null scratch address on device %d", device);
      return CSM_ERR_CUDA;
    }
    g_noalloc_ptr[device] = static_cast<unsigned char*>(p);
  }
  *out = g_noalloc_ptr[device];
  return CSM_OK;
}

}  // namespace

extern "C" int csm_discretize2d_noalloc(csm_ctx* ctx, const csm_job2d* job, int num_x, int num_y, int shrink,
                                        int32_t* out_xy, int32_t* out_bounds) {
  CSM_REQUIRE(ctx != nullptr && job != nullptr && out_xy != nullptr && out_bounds != nullptr, "NULL argument");
  CSM_REQUIRE(job->num_points >= 1 && job->num_scans >= 1 && job->points_xyz && job->scan_quat,
              "job needs points and a scan table");
  if (job->num_points > kNoAllocMaxPoints || job->num_scans > kNoAllocMaxScans ||
      static_cast<long long>(job->num_points) * job->num_scans > kNoAllocMaxCells) {
    set_error("csm_discretize2d_noalloc: %d points x %d scans exceed the fixed scratch (at most %d points, %d "
              "scans, %lld points x scans); use csm_discretize2d",
              job->num_points, job->num_scans, kNoAllocMaxPoints, kNoAllocMaxScans, kNoAllocMaxCells);
    return CSM_ERR_ARG;
  }
  CSM_CUDA(cudaSetDevice(ctx->device));
  unsigned char* base = nullptr;
  const int rc = noalloc_scratch(ctx->device, &base);
  if (rc != CSM_OK) return rc;
  JobDev j;
  memset(&j, 0, sizeof(j));
  j.num_x = num_x;
  j.num_y = num_y;
  j.depth = 1;
  j.n = job->num_points;
  j.S = job->num_scans;
  j.lin = job->linear_bound;
  j.q0w = job->init_quat_w;
  j.q0s = job->init_quat_s;
  j.tx = job->init_tx;
  j.ty = job->init_ty;
  j.max_x = job->max_x;
  j.max_y = job->max_y;
  j.res = job->resolution;
  JobDev* d_job = reinterpret_cast<JobDev*>(base);
  float* d_pts = reinterpret_cast<float*>(base + kNaJob);
  float* d_quat = reinterpret_cast<float*>(base + kNaJob + kNaPts);
  int2* d_xy = reinterpret_cast<int2*>(base + kNaJob + kNaPts + kNaQuat);
  ScanInfo* d_scan = reinterpret_cast<ScanInfo*>(base + kNaJob + kNaPts + kNaQuat + kNaXy);
  cudaStream_t st = ctx->stream;
  // `j` is on this frame: the copy must have left it before we return (the stream is synchronised below).
  CSM_CUDA(cudaMemcpyAsync(d_job, &j, sizeof(j), cudaMemcpyHostToDevice, st));
  CSM_CUDA(cudaMemcpyAsync(d_pts, job->points_xyz, sizeof(float) * 3 * j.n, cudaMemcpyHostToDevice, st));
  CSM_CUDA(cudaMemcpyAsync(d_quat, job->scan_quat, sizeof(float) * 2 * j.S, cudaMemcpyHostToDevice, st));
  ++ctx->launches, k_discretize<<<dim3(j.S, 1), 128, 0, st>>>(d_job, d_pts, d_quat, d_xy, d_scan, nullptr, nullptr,
                                                              shrink);
  CSM_CUDA(cudaGetLastError());
  CSM_CUDA(cudaMemcpyAsync(out_xy, d_xy, sizeof(int2) * static_cast<size_t>(j.S) * j.n, cudaMemcpyDeviceToHost, st));
  // ScanInfo starts with (min_x, max_x, min_y, max_y): a strided copy drops the other four ints.
  CSM_CUDA(cudaMemcpy2DAsync(out_bounds, 4 * sizeof(int32_t), d_scan, sizeof(ScanInfo), 4 * sizeof(int32_t), j.S,
                             cudaMemcpyDeviceToHost, st));
  CSM_CUDA(cudaStreamSynchronize(st));
  return CSM_OK;
}
