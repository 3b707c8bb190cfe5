// Device-side data structures of the 3D path (precomputation stack over a dense-ified
// HybridGrid, low-resolution probability grid, batched branch-and-bound matcher).
#ifndef CSM_MATCH3D_CUH_
#define CSM_MATCH3D_CUH_

#include "csm_internal.cuh"
#include "csm_match2d.cuh"  // float_to_ordered / ordered_to_float

namespace csm {

// One precomputation level over the box lo .. lo + dim - 1 (cell indices of that
// level's own resolution); everything outside the box is 0, the default of the
// reference's sparse grid (hybrid_grid.h:263-279).
//   data: bytes [z][y][x], byte-identical to the reference's level (parity tests, lowest-
//         resolution scoring, input of the next level's build).
//   pack: what the branch-and-bound sweep reads.  Element [zp][yp][xp] with
//         (xp, yp, zp) = (x, y, z) + stride in [0, pdim) holds the EIGHT values a parent
//         needs for its children, byte k = value(x + stride*(k&1), y + stride*((k>>1)&1),
//         z + stride*(k>>2)), k in the reference's generation order (FCSM3D.cc:412-430);
//         `stride` is the distance in cells between sibling children on this level
//         (2^level, divided by the per-level reduction of :335-342).  One 8-byte load per
//         (parent, point) instead of eight scattered byte loads.
struct Level3 {
  const uint8_t* data;
  const uint2* pack;
  int lo[3];
  int dim[3];
  int stride;
  int pdim[3];  // dim + stride
};

struct Stack3Dev {
  int depth;  // branch_and_bound_depth = number of levels
  int frd;    // full_resolution_depth
  Level3 level[CSM_MAX_DEPTH];
};

// Low-resolution grid: fp32 probabilities [z][y][x]; `unknown` outside the box.
struct Grid3Dev {
  const float* data;
  int lo[3];
  int dim[3];
  float unknown;
};

struct Job3Dev {
  const Stack3Dev* sd;
  const Grid3Dev* low;
  int depth, frd;
  int n;        // high-resolution points
  int n_low;    // low-resolution points
  int S;        // discrete scans
  int wxy, wz;  // linear windows in voxels
  int nxy, nz;  // lowest-resolution candidates per axis
  long long pts_off;    // float3 index into points
  long long low_off;    // float3 index into low points
  long long scan_off;   // index into per-scan arrays
  long long cell_off;   // int4 index into discrete scans (S * n)
  long long cand_off;   // index into top-level score array
  float res, low_res;
  float min_score;
  float n_float;
  double min_low;
};

struct Job3State {
  unsigned int best;        // ordered bits of S* (best leaf score that passed the filter), 0 = none
  unsigned int eq_count;    // nodes pruned while exactly equal to the incumbent (deferred)
  // winner selection over the leaves the sweep itself found (the common case)
  unsigned int w_group;     // ordered lowest-resolution score of the winning group
  unsigned int w_cand_min, w_cand_max;  // lowest-resolution candidates of that group holding an S*-leaf
  unsigned int w_slot;      // winners-list slot of the selected leaf
  unsigned long long w_path_min;
  // traversal-order fallback (only when use_dfs: too many deferred nodes, e.g. a world
  // of equal scores, where walking in order finds the winner at once)
  unsigned int group_best;  // ordered top score of the winning group
  unsigned int tie_count;
  unsigned long long win;   // min over tied contenders of (cand index << 32 | list slot)
  unsigned int total;       // top-level candidates
  unsigned int use_dfs;
  unsigned long long scored;
};

// Branch-and-bound work item, 32 bytes.  `path` is the node's position in the
// reference's traversal below its lowest-resolution ancestor `top_cand`: three bits
// per level, the rank of the node among its (stably sorted) siblings, most significant
// level first -- so smaller paths are visited earlier.  Bits 60..63 hold the node's
// depth while it waits in the deferred list.
struct Node3S {
  int job_scan;  // job << 16 | scan
  int ox, oy, oz;
  float score;
  unsigned int top_cand;
  unsigned long long path;
};

// A leaf that exceeded min_score, tied or beat the incumbent and passed the filter.
struct Winner3 {
  int job, scan;
  int leaf[3];
  unsigned int cand;
  float score;
  float low;
  unsigned long long path;
};

struct Contender3 {
  int job, scan;
  int ox, oy, oz;
  unsigned int cand;
  float score;
  int leaf[3];
  int has_leaf;
  float leaf_low;
};

}  // namespace csm

struct csm_stack3d {
  int device = 0;
  int depth = 0, frd = 0;
  uint8_t* base = nullptr;
  size_t level_off[CSM_MAX_DEPTH] = {0};
  size_t pack_off[CSM_MAX_DEPTH] = {0};
  csm::Stack3Dev host_view;
  csm::Stack3Dev* dev = nullptr;
  size_t bytes = 0, cap = 0;  // cap: capacity of the pooled block behind base
};

struct csm_grid3d {
  int device = 0;
  float* data = nullptr;
  csm::Grid3Dev host_view;
  csm::Grid3Dev* dev = nullptr;
  size_t bytes = 0, cap = 0;
};

struct csm_batch3d {
  int num_jobs = 0;
  std::vector<csm::Job3Dev> jobs_host;
  int max_scans = 0;
  int max_depth = 0;
  long long total_pts = 0, total_low = 0, total_scans = 0, total_cells = 0, total_cand = 0;
  csm::Job3Dev* jobs = nullptr;
  float* points = nullptr;       // total_pts * 3
  float* low_points = nullptr;   // total_low * 3
  float* scan_poses = nullptr;   // total_scans * 7
  float* scan_leaf_q = nullptr;  // total_scans * 4
  int4* cells = nullptr;         // total_cells
  unsigned long long* scan_best = nullptr;  // total_scans
  float* top_scores = nullptr;   // total_cand
  csm::Job3State* state = nullptr;
  csm::Contender3* contenders = nullptr;
  unsigned int contender_cap = 0;
  csm::Winner3* winners = nullptr;
  unsigned int winner_cap = 0;
  csm::Node3S* deferred = nullptr;   // nodes pruned at bound == incumbent
  unsigned int deferred_cap = 0;
  unsigned long long frontier_want = 0;
  unsigned int* counters = nullptr;
  csm_result3d* results_dev = nullptr;
  csm_ctx* owner = nullptr;  // context whose block pool backs arena / pinned
  void* arena = nullptr;
  size_t arena_bytes = 0, arena_cap = 0;
  void* pinned = nullptr;
  size_t pinned_bytes = 0, pinned_cap = 0;
  size_t pin_results_off = 0;
  bool uploaded = false, ran = false;
  std::vector<cudaEvent_t> prof_events;  // kPhases + 1 events of the last profiled run
};

#endif  // CSM_MATCH3D_CUH_
