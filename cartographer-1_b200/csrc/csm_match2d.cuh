// Device-side data structures of the batched 2D branch-and-bound matcher.
#ifndef CSM_MATCH2D_CUH_
#define CSM_MATCH2D_CUH_

#include "csm_internal.cuh"

namespace csm {

// One (node, submap) search as the kernels see it.  All offsets index the batch's
// shared device arrays.
struct JobDev {
  const StackDev* sd;                   // precomputation stack (device descriptor)
  int num_x, num_y;                     // grid cell limits
  int depth;                            // branch_and_bound_depth
  int n;                                // points per scan
  int S;                                // num_scans
  int lin;                              // initial +-linear bound (cells)
  long long pts_off;                    // float3 index into points
  long long quat_off;                   // float2 index into scan quaternions
  long long xy_off;                     // int2 index into discrete scans (S*n)
  long long scan_off;                   // index into per-scan arrays (S)
  long long cand_off;                   // index into top-level score array
  long long cand_cap;                   // reserved top-level candidates
  float q0w, q0s;                       // initial-yaw quaternion (w, z)
  float tx, ty;                         // fp32 initial translation
  double max_x, max_y, res;             // MapLimits
  float min_score;
  float score_offset, score_scale;      // ToScore
  float n_float;                        // float(n)
};

// Per-scan search window after ShrinkToFit plus the lowest-resolution lattice.
struct ScanInfo {
  int min_x, max_x, min_y, max_y;
  int nx, ny;        // lowest-resolution candidates per axis (0 if the window is empty)
  int base;          // exclusive prefix of nx*ny inside the job
  int pad;
};

// Mutable per-job search state.
struct JobState {
  unsigned long long best;      // (ordered leaf score bits << 32) | ordered top score of its ancestor
  unsigned long long top_best;  // (ordered top-level score bits << 32) | candidate index
  unsigned long long win;       // phase 2: min over tied contenders of (cand index << 32 | list slot)
  unsigned int group_best;      // phase 2: ordered top score of the winning group
  unsigned int tie_count;       // phase 2: contenders in the winning group that hold an S*-leaf
  unsigned int total;           // top-level candidates of this job
  unsigned int cand_max;        // path mode: largest ancestor index in the winning group (!= win >> 32: tie)
  unsigned long long scored;    // candidate scores evaluated (all phases)
  unsigned long long scored_phase[4];  // score_top, dive, expand, phase 2
};

// Branch-and-bound work item: a tile of kTW x kTH lattice-adjacent nodes of one
// (job, scan) at one depth, 16 bytes.  Node (li, lj) of tile (ti, tj) is the lattice
// point (kTW*ti + li, kTH*tj + lj), i.e. the offset (min_x + step*(kTW*ti + li),
// min_y + step*(kTH*tj + lj)) with step = 2^depth; bit lj*kTW + li of `mask` says
// whether its bound survived pruning.
struct Tile {
  int job_scan;       // job << 16 | scan   (job < 65536 per chunk, scan < 65536)
  int tij;            // ti | tj << 16
  unsigned int mask;  // alive nodes
  int pad;
};

// Phase-2 contender: a top-level candidate that may contain the winning leaf.
struct Contender {
  int job, scan, x, y;
  unsigned int cand;    // generation index inside the job
  float score;          // its top-level score
  int leaf_x, leaf_y;   // first S*-leaf in reference DFS order, valid if has_leaf
  int has_leaf;
  int pad[3];
};

// Traversal order without a second walk (node sweep).  The reference returns the FIRST leaf
// attaining the best score S* in its depth-first order: top-level candidates by descending score,
// then at every level the <= 4 siblings by descending score, generation order among equals
// (std::sort of <= 16 elements is an insertion sort, i.e. stable).  Every node of the sweep carries
// its `path`: two bits per level below the top = its rank among its sorted siblings, most
// significant first.  Among the leaves that attain S*, the winner is the one with the best-ranked
// top-level ancestor and then the smallest path.
//   Leaf2:     a leaf whose score was >= the incumbent when it was scored.
//   Deferred2: a node whose bound EQUALLED the incumbent: the reference prunes it only if the
//              incumbent was found earlier in ITS order, so it is set aside and, if its bound turns
//              out to equal S*, walked depth-first in sibling order for its first S*-leaf.
struct Leaf2 {
  int job, scan;
  int x, y;            // index offsets of the leaf (Candidate2D::x_index_offset / y_index_offset)
  unsigned int cand;   // generation index of its top-level ancestor inside the job
  unsigned int path;
  float score;
  int pad;
};
struct Deferred2 {
  int job, scan;
  int x, y;            // index offsets of the node
  int depth;           // its depth (children live on level depth - 1)
  unsigned int cand;   // top-level ancestor
  unsigned int path;   // ranks from the top down to this node
  float score;         // its bound
};

// Order-preserving map float -> uint32 (works for negative values too).
__host__ __device__ inline unsigned int float_to_ordered(float f) {
#ifdef __CUDA_ARCH__
  unsigned int u = __float_as_uint(f);
#else
  unsigned int u;
  memcpy(&u, &f, 4);
#endif
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__host__ __device__ inline float ordered_to_float(unsigned int o) {
  unsigned int u = (o & 0x80000000u) ? (o & 0x7FFFFFFFu) : ~o;
#ifdef __CUDA_ARCH__
  return __uint_as_float(u);
#else
  float f;
  memcpy(&f, &u, 4);
  return f;
#endif
}


#ifdef __CUDACC__
// Eigen's quaternion * vector for a pure yaw quaternion (w, 0, 0, s), fp32, in
// Eigen 3.3's op order uv = 2 * (vec x v); v' = (v + w*uv) + vec x uv, with the
// terms that are exactly zero dropped (see oracle/csm_oracle.h QuatRotate).
__device__ __forceinline__ void yaw_rotate(float w, float s, float& vx, float& vy) {
  float uvx = -__fmul_rn(s, vy);
  float uvy = __fmul_rn(s, vx);
  uvx = __fadd_rn(uvx, uvx);
  uvy = __fadd_rn(uvy, uvy);
  const float cx = -__fmul_rn(s, uvy);
  const float cy = __fmul_rn(s, uvx);
  const float rx = __fadd_rn(__fadd_rn(vx, __fmul_rn(w, uvx)), cx);
  const float ry = __fadd_rn(__fadd_rn(vy, __fmul_rn(w, uvy)), cy);
  vx = rx;
  vy = ry;
}

// MapLimits::GetCellIndex (cartographer/mapping/2d/map_limits.h:69-76), fp64; the
// cell x comes from the world y coordinate and vice versa.
__device__ __forceinline__ void cell_index(double max_x, double max_y, double res, float wx,
                                           float wy, int& cx, int& cy) {
  const double fx = __dsub_rn(__ddiv_rn(__dsub_rn(max_y, static_cast<double>(wy)), res), 0.5);
  const double fy = __dsub_rn(__ddiv_rn(__dsub_rn(max_x, static_cast<double>(wx)), res), 0.5);
  cx = static_cast<int>(lround(fx));
  cy = static_cast<int>(lround(fy));
}
#endif

}  // namespace csm

struct csm_batch2d {
  int num_jobs = 0;
  std::vector<csm::JobDev> jobs_host;
  std::vector<int> empty_job;          // jobs with zero points: not found, no device work
  int max_scans = 0;
  long long total_pts = 0, total_scans = 0, total_xy = 0, total_cand = 0;
  // device arrays
  csm::JobDev* jobs = nullptr;
  float* points = nullptr;             // total_pts * 3
  float* quats = nullptr;              // total_scans * 2
  int2* xy = nullptr;                  // total_xy
  csm::ScanInfo* scans = nullptr;      // total_scans
  unsigned long long* scan_best = nullptr;  // total_scans: best lowest-resolution key per scan
  float* top_scores = nullptr;         // total_cand
  csm::JobState* state = nullptr;      // num_jobs
  unsigned long long frontier_want = 0;  // nodes per frontier buffer this batch asks for
  csm::Contender* contenders = nullptr;
  unsigned int contender_cap = 0;        // also the capacity of `leaves` and `deferred`
  csm::Leaf2* leaves = nullptr;          // path mode
  csm::Deferred2* deferred = nullptr;    // path mode
  unsigned int* counters = nullptr;    // [0..CSM_MAX_DEPTH]: frontier sizes, [32]: contenders, [33]: overflow
  csm_result2d* results_dev = nullptr;
  unsigned long long* hist = nullptr;  // [depth][alive nodes per tile] tile counts (profiling)
  unsigned int* bucket_js = nullptr;   // total_scans: job << 16 | scan of every bucket
  unsigned int* bucket_cnt = nullptr;  // total_scans: survivors appended for the next level
  unsigned int* bucket_start = nullptr;   // total_scans + 1: offsets into the sorted node list
  unsigned int* bucket_istart = nullptr;  // total_scans + 1: offsets into the work items
  csm_ctx* owner = nullptr;            // context whose block pool backs arena / pinned
  void* arena = nullptr;               // single device block backing all of the above
  size_t arena_bytes = 0, arena_cap = 0;
  size_t pinned_cap = 0;
  void* pinned = nullptr;              // staging for uploads / results
  size_t pinned_bytes = 0;
  bool uploaded = false, ran = false;
  cudaEvent_t done = nullptr;            // recorded by run() after the results reached the pinned block
  bool idle = false;                     // fetched since the last run: nothing of it is in flight
  std::vector<cudaEvent_t> prof_events;  // kPhases + 1 events of the last profiled run
  std::vector<cudaEvent_t> depth_events;  // 3 per swept depth: lists, expand start, expand end
  std::vector<int> depth_of_events;       // the depth of every such triple
  double* depth_stats = nullptr;          // CSM_MAX_DEPTH x 6 (k_item_stats, profiling 2)
};

#endif  // CSM_MATCH2D_CUH_
