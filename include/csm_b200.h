/* csm_b200.h -- C ABI of the B200-native correlative scan matcher.
 *
 * This is the drop-in boundary: everything the reference's scan-matching classes
 * compute on the CPU between "grid + point cloud in" and "best candidate out" is
 * behind these entry points, implemented as hand-written sm_100a CUDA kernels.
 * There is no CPU fallback: every entry point fails with CSM_ERR_CUDA when no
 * usable device exists.
 *
 * The reference (juwangvsu/cartographer-1) has no FFI layer; the interface each
 * entry point replaces is a C++ member function.  Citations are relative to
 * cartographer/mapping/internal/2d/scan_matching/ unless a path is given:
 *   FCSM2D = fast_correlative_scan_matcher_2d.{h,cc}
 *   CSM2D  = correlative_scan_matcher_2d.{h,cc}
 *   RT2D   = real_time_correlative_scan_matcher_2d.{h,cc}
 *
 * What stays on the host side of the boundary (libm-dependent, O(num_scans)):
 * SearchParameters construction (CSM2D.cc:27-55: sqrtf/acos/ceil), the per-scan
 * half-angle (cosf, sinf) table for GenerateRotatedScans (CSM2D.cc:99-106), the
 * uint16 -> uint8 quantisation table of PrecomputationGrid2D::ComputeCellValue
 * (FCSM2D.cc:163-169) and the candidate -> pose conversion (CSM2D.h:81-87,
 * FCSM2D.cc:253-260).  cartographer-1_b200/host/ holds that host code behind the
 * reference's own class signatures.
 *
 * All pointers are plain host pointers unless the name says "_dev".  All functions
 * return 0 on success and a negative csm_status otherwise; csm_last_error() gives a
 * thread-local message.  "No match above min_score" is a normal result
 * (found == 0), not an error (FCSM2D.cc:253-261).
 *
 * Thread safety: a csm_ctx owns one CUDA stream and its scratch memory and must be
 * used by one host thread at a time; create one ctx per worker thread.  Stacks are
 * immutable after construction and may be shared by any number of contexts on the
 * same device, matching FastCorrelativeScanMatcher2D::Match being const and called
 * concurrently (constraints/constraint_builder_2d.cc:102-111).
 */
#ifndef CSM_B200_H_
#define CSM_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CSM_MAX_DEPTH 16

typedef enum {
  CSM_OK = 0,
  CSM_ERR_ARG = -1,      /* invalid argument (the reference would CHECK-fail) */
  CSM_ERR_CUDA = -2,     /* CUDA runtime error or no device */
  CSM_ERR_CAPACITY = -3  /* an internal work list overflowed even after regrowth */
} csm_status;

typedef struct csm_ctx csm_ctx;
typedef struct csm_stack2d csm_stack2d;
typedef struct csm_batch2d csm_batch2d;

const char* csm_last_error(void);
int csm_device_count(void);

/* One context = one device + one stream + scratch arenas. */
int csm_ctx_create(int device, csm_ctx** out);
void csm_ctx_destroy(csm_ctx* ctx);
int csm_ctx_synchronize(csm_ctx* ctx);
/* The context's cudaStream_t, so a caller can record its own CUDA events on it. */
void* csm_ctx_stream(csm_ctx* ctx);
/* Per-phase CUDA-event timing of csm_batch2d_run on the context's own stream
 * (measurement aid for bench.py).  Phases: 0 rotate/discretise, 1 lowest-resolution
 * scoring, 2 greedy dive, 3 level-synchronous expansion, 4 tie-order phase.
 * phase_ms5 / phase_scored5: accumulated milliseconds and candidate scores evaluated
 * per phase since profiling was switched on; runs: number of profiled batch runs.
 * enabled: 0 off, 1 phase timing, 2 phase timing + tile histogram (adds atomics). */
int csm_ctx_set_profiling(csm_ctx* ctx, int enabled);
int csm_ctx_get_profile(csm_ctx* ctx, double* phase_ms5, double* phase_scored5, int* runs);
/* Expanded branch-and-bound tiles since profiling was switched on, as a histogram
 * hist[depth * 33 + alive] (CSM_MAX_DEPTH x 33 doubles): tiles of parents at `depth`
 * that held `alive` surviving nodes. */
int csm_ctx_get_tile_histogram(csm_ctx* ctx, double* hist);

/* ---- PrecomputationGridStack2D (FCSM2D.cc:91-186, FCSM2D.h:49-109) ----------
 * cells:  the Grid2D correspondence-cost cells, row-major, flat = num_x*y + x
 *         (cartographer/mapping/2d/grid_2d.h:113-116).
 * q_lut:  65536 bytes, q_lut[v] = ComputeCellValue(1.f - |table[v]|) with `table`
 *         the grid's value-to-correspondence-cost table
 *         (cartographer/mapping/value_conversion_tables.cc:29-67); computed by the
 *         host with the reference's fp32 ops so the device work is integer-exact.
 * depth:  branch_and_bound_depth (>= 1); level i has width 1 << i.
 * Level i holds (num_x + 2^i - 1) x (num_y + 2^i - 1) bytes, byte-identical to
 * PrecomputationGrid2D::cells_ of the reference.
 */
int csm_stack2d_build(csm_ctx* ctx, const uint16_t* cells, int num_x, int num_y,
                      const uint8_t* q_lut, int depth, csm_stack2d** out);
/* `count` stacks in one call, with the semantics of `count` csm_stack2d_build calls (cells[k] is
 * num_x[k] x num_y[k], q_lut[k] its table; one depth): the reference builds one
 * PrecomputationGridStack2D per FastCorrelativeScanMatcher2D constructor
 * (fast_correlative_scan_matcher_2d.cc:171-186, 198-202); a caller that has several finished submaps at
 * once (a loaded map, a burst of constraint searches) hands them over together and the host's staging of
 * stack k + 1 overlaps the device's build of stack k.  Synchronises once, at the end; on failure every
 * out[k] is NULL.  csm_stack2d_build is the count = 1 case. */
int csm_stack2d_build_many(csm_ctx* ctx, int count, const uint16_t* const* cells, const int* num_x,
                           const int* num_y, const uint8_t* const* q_lut, int depth, csm_stack2d** out);
/* Page-locks (cudaHostRegister) / releases a host range the caller owns, e.g. a finished submap's
 * cells: csm_stack2d_build_many then copies the cells straight from the caller's memory with an
 * asynchronous transfer instead of through its staging blocks.  Plain pageable memory keeps working. */
int csm_host_register(void* ptr, size_t bytes);
int csm_host_unregister(void* ptr);
/* Same, with `cells_dev` already resident in device memory (benchmark path). */
int csm_stack2d_build_dev(csm_ctx* ctx, const uint16_t* cells_dev, int num_x, int num_y,
                          const uint8_t* q_lut, int depth, csm_stack2d** out);
/* ---- A new submap's grid to every replica (no counterpart in the reference,
 * whose submaps live in one address space: constraint_builder_2d.cc:165-186 builds the matcher
 * from the Grid2D it is handed).  One process per GPU: the owner of a finished submap broadcasts
 * its uint16 cells over NCCL (NVLink) and every rank builds its own stack from its copy.
 * nccl_comm is the caller's ncclComm_t (one rank per process); NCCL is reached through dlsym on the
 * libnccl.so.2 the process already loaded, so this header and the library carry no NCCL dependency.
 * Collective: every rank of the communicator calls it with the same root, sizes and depth.
 * csm_grid2d_broadcast_dev: in place on a device buffer, asynchronous on the context's stream.
 * csm_stack2d_broadcast: root passes is_root != 0 and its host cells (others may pass NULL). */
int csm_grid2d_broadcast_dev(csm_ctx* ctx, void* nccl_comm, int root, uint16_t* cells_dev, size_t num_cells);
int csm_stack2d_broadcast(csm_ctx* ctx, void* nccl_comm, int root, int is_root, const uint16_t* cells_host,
                          int num_x, int num_y, const uint8_t* q_lut, int depth, csm_stack2d** out);
/* Rebuilds an existing stack in place from device-resident cells; asynchronous on
 * the context's stream (used to time the build kernels alone). */
int csm_stack2d_rebuild_dev(csm_ctx* ctx, csm_stack2d* stack, const uint16_t* cells_dev);
/* Copies level `level` to `out` (may be NULL to query the size only). */
int csm_stack2d_level(csm_ctx* ctx, const csm_stack2d* stack, int level, uint8_t* out,
                      int* wide_x, int* wide_y);
/* Copies the matcher's re-layout of level `level` (the quad-phase words, DESIGN.md §2;
 * the reference has no counterpart, its PrecomputationGrid2D::GetValue reads cells_
 * directly, fast_correlative_scan_matcher_2d.h:56-70) to `out`: 4 * s * s * in * jn
 * words, s = 1 << level.  `out` may be NULL to query in / jn only.  Test support. */
int csm_stack2d_quad(csm_ctx* ctx, const csm_stack2d* stack, int level, uint32_t* out, int* in,
                     int* jn);
int csm_stack2d_depth(const csm_stack2d* stack);
/* The grid's uint16 cells as given to the last build, resident beside the levels (what the
 * refinement after a found match interpolates: csm_ceres_job2d::cells_dev). */
const uint16_t* csm_stack2d_cells_dev(const csm_stack2d* stack);
/* Stack builds on `ctx` write that copy from their level-0 pass (default on).  Off: builds do
 * exactly the work of PrecomputationGridStack2D and csm_stack2d_cells_dev is stale or unset --
 * what bench.py's C4 record times. */
int csm_ctx_set_keep_cells(csm_ctx* ctx, int enabled);
size_t csm_stack2d_device_bytes(const csm_stack2d* stack);
/* `ctx` may be NULL or a context that no longer exists: the stack's device block goes
 * back to the block pool of `ctx` when that is still a live context on the stack's
 * device (the next build reuses it without a cudaMalloc), and is freed otherwise. */
void csm_stack2d_destroy(csm_ctx* ctx, csm_stack2d* stack);

/* ---- PrecomputationGrid2D of any width (the public class of FCSM2D.h:49-92; ctor FCSM2D.cc:91-161)
 * out: (num_x + width - 1) x (num_y + width - 1) bytes, dense, byte-identical to the reference's
 * cells_ (wide index (x0 + width - 1, y0 + width - 1) = maximum over the width x width cells from
 * (x0, y0), 0 outside the grid).  Used by the host class PrecomputationGrid2D::GetValue. */
int csm_precomputation_grid2d(csm_ctx* ctx, const uint16_t* cells, int num_x, int num_y,
                              const uint8_t* q_lut, int width, uint8_t* out);

/* ---- FastCorrelativeScanMatcher2D::MatchWithSearchParameters (FCSM2D.cc:227-262)
 * One job = one (node, submap) search.  The host supplies the search parameters it
 * computed with libm; the device does GenerateRotatedScans, DiscretizeScans,
 * ShrinkToFit, ComputeLowestResolutionCandidates and BranchAndBound.
 */
typedef struct {
  const csm_stack2d* stack;   /* precomputation stack of the submap               */
  const float* points_xyz;    /* num_points x 3 fp32, sensor frame                */
  int32_t num_points;
  int32_t num_scans;          /* SearchParameters::num_scans (CSM2D.cc:45)         */
  const float* scan_quat;     /* num_scans x 2: (cosf(a/2), sinf(a/2)) of the     */
                              /* fp32-narrowed delta_theta of each scan            */
  float init_quat_w;          /* cosf/sinf of half the fp32-narrowed initial yaw   */
  float init_quat_s;          /* (FCSM2D.cc:235-239)                               */
  float init_tx, init_ty;     /* initial translation narrowed to fp32 (:244-245)   */
  double max_x, max_y;        /* MapLimits::max()                                  */
  double resolution;          /* MapLimits::resolution()                           */
  int32_t linear_bound;       /* initial +-bound in cells (CSM2D.cc:48-54)         */
  float min_score;            /* strict threshold (FCSM2D.cc:253)                  */
  float score_offset;         /* PrecomputationGrid2D::min_score_ (FCSM2D.cc:97)   */
  float score_scale;          /* (max_score_ - min_score_) / 255.f (FCSM2D.h:74-76)*/
} csm_job2d;

typedef struct {
  int32_t found;              /* best.score > min_score                            */
  int32_t scan_index;         /* Candidate2D::scan_index of the winner             */
  int32_t x_index_offset;     /* Candidate2D::x_index_offset                       */
  int32_t y_index_offset;     /* Candidate2D::y_index_offset                       */
  float score;                /* Candidate2D::score (bit-identical fp32)           */
  int32_t tie_resolution;     /* 0: unique on device, 1: host replay of std::sort  */
  uint64_t candidates_scored; /* candidate scores evaluated for this job           */
} csm_result2d;

/* Synchronous convenience: upload + run + fetch. */
int csm_match2d_batch(csm_ctx* ctx, const csm_job2d* jobs, int num_jobs, csm_result2d* results);

/* Split form, so a caller can keep inputs resident and overlap work:
 *   upload : host -> device copy of points / tables (asynchronous)
 *   run    : all kernels, asynchronous on the context's stream
 *   fetch  : device -> host copy of results, waits, resolves rare top-level ties
 *            with the reference's own std::sort order on the host
 */
int csm_batch2d_upload(csm_ctx* ctx, const csm_job2d* jobs, int num_jobs, csm_batch2d** out);
int csm_batch2d_run(csm_ctx* ctx, csm_batch2d* batch);
int csm_batch2d_fetch(csm_ctx* ctx, csm_batch2d* batch, csm_result2d* results);
void csm_batch2d_destroy(csm_ctx* ctx, csm_batch2d* batch);

/* ---- DiscretizeScans + ShrinkToFit alone (CSM2D.cc:73-127), for the bit-exact
 * cell-index parity test.  out_xy: num_scans*num_points*2 int32 (x, y);
 * out_bounds: num_scans*4 int32 (min_x, max_x, min_y, max_y).
 */
int csm_discretize2d(csm_ctx* ctx, const csm_job2d* job, int num_x, int num_y, int shrink,
                     int32_t* out_xy, int32_t* out_bounds);

/* csm_discretize2d for callers that must not allocate.  NO COUNTERPART IN THE REFERENCE: added
 * for code under CUDA graph capture, tight serving loops and harnesses
 * that forbid cudaMalloc.  Same kernel, same arguments and outputs as csm_discretize2d, but the scratch
 * (job descriptor, points, scan table, discrete scans, per-scan windows) is one fixed device array
 * inside the library instead of the context's growable buffer, and the call makes no cudaMalloc /
 * cudaFree / cudaHostAlloc and no host allocation.
 * Limits: num_points <= 4096, num_scans <= 4096, num_scans * num_points <= 524288; larger inputs
 * are rejected with CSM_ERR_ARG and a message naming the limits (use csm_discretize2d then).
 * Every GPU that loads the library holds the array (about 4.2 MiB) until exit.  All calls share it
 * and it keeps the last call's contents (the kernel overwrites every element it copies out, nothing
 * has to be cleared): calls that overlap on the GPU -- other contexts, streams or threads -- race. */
int csm_discretize2d_noalloc(csm_ctx* ctx, const csm_job2d* job, int num_x, int num_y, int shrink,
                             int32_t* out_xy, int32_t* out_bounds);

/* ---- the free functions of CSM2D.h:63-72 (the matchers fuse them; the reference's tests and
 * any other caller reach them here).
 * csm_rotate_scans2d: GenerateRotatedScans (CSM2D.cc:93-109): out_xyz = num_scans x num_points x 3,
 *   scan s = the cloud rotated about z by the angle whose half-angle (cosf, sinf) is scan_quat[2s..].
 * csm_discretize_points2d: DiscretizeScans for num_points points of any number of scans laid end
 *   to end (CSM2D.cc:111-127): out_xy[2i..] = GetCellIndex(Translation2f(tx, ty) * p_i.head<2>()). */
int csm_rotate_scans2d(csm_ctx* ctx, const float* points_xyz, int num_points, const float* scan_quat,
                       int num_scans, float* out_xyz);
int csm_discretize_points2d(csm_ctx* ctx, const float* points_xyz, long long num_points, float tx, float ty,
                            double max_x, double max_y, double resolution, int32_t* out_xy);

/* ---- RealTimeCorrelativeScanMatcher2D::Match (RT2D.cc:117-176) -----------------
 * Exhaustive search directly on the live uint16 grid (no precomputation).
 * cost_table: the grid's 65536-entry value-to-correspondence-cost table.
 * penalty:    num_scans x (2*linear_bound+1)^2 doubles, the host-computed
 *             exp(-(hypot(x,y)*w_t + |theta|*w_r)^2) factor per candidate in
 *             generation order scan -> x -> y (RT2D.cc:170-174).
 */
typedef struct {
  const uint16_t* cells;      /* live grid cells, host                             */
  int32_t num_x, num_y;
  const float* cost_table;    /* 65536 floats                                      */
  const float* points_xyz;
  int32_t num_points;
  int32_t num_scans;
  const float* scan_quat;
  float init_quat_w, init_quat_s;
  float init_tx, init_ty;
  double max_x, max_y, resolution;
  int32_t linear_bound;
  const double* penalty;
  float out_of_bounds_probability; /* kMinProbability for a ProbabilityGrid            */
                                   /* (cartographer/mapping/2d/probability_grid.cc:78-82)*/
  /* TSDF2D grids (RT2D.cc:38-59, mapping/internal/2d/tsdf_2d.cc:88-98): a non-NULL
   * weight_cells selects the TSDF score sum(w * (max_cc - |tsd|) / max_cc) / sum(w),
   * 0 when sum(w) == 0.  `cells` / `cost_table` are then the TSD cells and the
   * value-to-TSD table; a cell outside the grid reads (out_of_bounds_tsd, weight 0). */
  const uint16_t* weight_cells;    /* NULL for a ProbabilityGrid                       */
  const float* weight_table;       /* 65536 floats, value-to-weight                    */
  float max_correspondence_cost;   /* the truncation distance (tsdf_2d.cc:26)          */
  float out_of_bounds_tsd;         /* TSDValueConverter::getMinTSD()                   */
  /* Compact form of `penalty` (then `penalty` may be NULL): the factor depends only on
   * (|scan - num_angular_perturbations|, |x offset|, |y offset|), so
   * penalty_table[(a * (linear_bound + 1) + |x|) * (linear_bound + 1) + |y|], a = 0 ..
   * num_angular_perturbations, holds every distinct value once (5 KB instead of 28 KB per scan at
   * the default options) and the device indexes it per candidate. */
  const double* penalty_table;     /* NULL: use `penalty`                              */
  int32_t num_angular_perturbations;
  int32_t pad;
} csm_rt_job2d;

int csm_rt_match2d_batch(csm_ctx* ctx, const csm_rt_job2d* jobs, int num_jobs,
                         csm_result2d* results);
/* RealTimeCorrelativeScanMatcher2D::ScoreCandidates on a caller-supplied list ("visible for
 * testing", RT2D.h:75-79; RT2D.cc:151-176 up to the motion penalty, which the host applies in
 * double): out_scores[c] = ComputeCandidateScore(grid, discrete_scans[scan_c], x_c, y_c).
 * `grid` supplies cells / tables / limits / num_points (its point, scan-table and penalty fields are
 * not read); discrete_xy = num_scans x num_points x 2 int32; candidates = num_candidates x
 * (scan_index, x_index_offset, y_index_offset). */
int csm_rt_score_candidates2d(csm_ctx* ctx, const csm_rt_job2d* grid, const int32_t* discrete_xy,
                              int num_scans, const int32_t* candidates, int num_candidates, float* out_scores);

/* ---- CeresScanMatcher2D::Match, batched ------------------------
 * Replaces cartographer/mapping/internal/2d/scan_matching/ceres_scan_matcher_2d.cc:64-110
 * (with occupied_space_cost_function_2d.cc:30-96 and the translation / rotation delta
 * functors) for ProbabilityGrid: minimises over (x, y, theta)
 *   sum_i [w_occ / sqrt(N) * C(T p_i)]^2 + [w_t (x - tx)]^2 + [w_t (y - ty)]^2 + [w_r (theta - theta0)]^2
 * with C the bicubic interpolation of the correspondence-cost grid (kMaxCorrespondenceCost
 * outside), by the trust-region Levenberg-Marquardt iteration ceres::Solve runs for
 * Cartographer's options (Jacobi scaling, Ceres' default radii and tolerances).  One warp per
 * problem, fp64.  Results agree with the CPU restatement to ~1e-9 on typical problems; the
 * parity tests allow 1e-6 on the pose (normal equations here, QR there).
 * termination: 0 gradient tolerance, 1 parameter tolerance, 2 function tolerance,
 *              3 max_num_iterations, 4 minimum trust-region radius, 5 failure.
 */
typedef struct {
  const uint16_t* cells;      /* grid cells, host                                    */
  int32_t num_x, num_y;
  const float* cost_table;    /* 65536 floats, value-to-correspondence-cost          */
  const float* points_xyz;
  int32_t num_points;
  int32_t max_num_iterations;       /* ceres_solver_options.max_num_iterations        */
  int32_t use_nonmonotonic_steps;   /* ceres_solver_options.use_nonmonotonic_steps    */
  int32_t pad;
  double max_x, max_y, resolution;
  double initial_pose[3];           /* x, y, theta                                    */
  double target_translation[2];
  double occupied_space_weight, translation_weight, rotation_weight;
  /* TSDF2D grids (tsdf_match_cost_function_2d.cc:28-75 on interpolated_tsdf_2d.h:35-131): a
   * non-NULL weight_cells selects residual_i = N * w_occ / sqrt(N) * tsd(T p_i) * weight(T p_i) /
   * sum_j weight(T p_j) with bilinear interpolation; `cells` / `cost_table` are then the TSD
   * cells and the value-to-TSD table.  A pose whose points see no weight at all fails the
   * evaluation like the reference's functor (termination 5 at the initial pose, a rejected
   * step later). */
  const uint16_t* weight_cells;
  const float* weight_table;        /* 65536 floats                                   */
  double truncation_distance;       /* Grid2D::GetMaxCorrespondenceCost of a TSDF2D   */
  /* The same cells already resident on the context's device (csm_stack2d_cells_dev of the submap's
   * stack, csm_grid2d_cells_dev of a live grid): then `cells` may be NULL and nothing is uploaded
   * for the grid (2 MB per 1000 x 1000 submap and batch otherwise). */
  const uint16_t* cells_dev;
} csm_ceres_job2d;

typedef struct {
  double pose[3];
  double initial_cost, final_cost;  /* ceres::Solver::Summary                         */
  int32_t iterations, num_successful_steps, num_unsuccessful_steps, termination;
} csm_ceres_result2d;

int csm_ceres_match2d_batch(csm_ctx* ctx, const csm_ceres_job2d* jobs, int num_jobs,
                            csm_ceres_result2d* results);

/* ---- ProbabilityGridRangeDataInserter2D::Insert on a device-resident grid
 * Replaces cartographer/mapping/2d/probability_grid_range_data_inserter_2d.cc:47-100 (CastRays)
 * with mapping/internal/2d/ray_to_pixel_mask.cc:25-170 and Grid2D::FinishUpdate
 * (mapping/2d/grid_2d.cc:98-106).  GrowLimits (:30-45, grid_2d.cc:131-166) is the caller's:
 * the grid handle has a fixed size and every ray must lie inside it.  A csm_grid2d keeps the
 * uint16 cells of a live submap in HBM: scans are inserted as they arrive and the finished
 * submap's stack is built from csm_grid2d_cells_dev() with csm_stack2d_build_dev -- the cells
 * never cross PCIe.  Cells are bit-identical to the reference's after every insertion. */
typedef struct csm_grid2d csm_grid2d;
typedef struct {
  double max_x, max_y, resolution;  /* MapLimits of the grid                              */
  float origin_x, origin_y;         /* range_data.origin                                  */
  const float* returns_xy;          /* num_returns x (x, y)                               */
  int32_t num_returns;
  int32_t num_misses;
  const float* misses_xy;           /* num_misses x (x, y)                                */
  const uint16_t* hit_table;        /* ComputeLookupTableToApplyCorrespondenceCostOdds    */
  const uint16_t* miss_table;       /* (probability_values.cc:89-105), 32768 entries each */
  int32_t insert_free_space;
  int32_t pad;
} csm_range_data2d;

/* cells may be NULL: every cell unknown (0). */
int csm_grid2d_create(csm_ctx* ctx, const uint16_t* cells, int num_x, int num_y, csm_grid2d** out);
/* Inserts the scans one after the other (each is one reference Insert + FinishUpdate). */
int csm_grid2d_insert(csm_ctx* ctx, csm_grid2d* grid, const csm_range_data2d* scans, int num_scans);
int csm_grid2d_download(csm_ctx* ctx, const csm_grid2d* grid, uint16_t* cells_out);
const uint16_t* csm_grid2d_cells_dev(const csm_grid2d* grid);
void csm_grid2d_destroy(csm_ctx* ctx, csm_grid2d* grid);

/* ---- TSDFRangeDataInserter2D::Insert on a device-resident TSDF2D ------
 * Replaces cartographer/mapping/internal/2d/tsdf_range_data_inserter_2d.cc:167-239 (InsertHit,
 * UpdateCell) with ray_to_pixel_mask.cc:25-170 and TSDF2D::SetCell / GetTSDAndWeight /
 * CellIsUpdated / FinishUpdate (mapping/internal/2d/tsdf_2d.cc:44-98).  The host keeps what is per
 * return and goes through libm -- the bearing sort (:71-90,146-150), EstimateNormals
 * (normal_estimation_2d.cc), cosf / sinf of the normal, the range and angle weight factors
 * (:189-203) -- and GrowAsNeeded (:33-48); the device owns every per-cell step: a cell is updated by
 * the first return, in the given order, whose ray crosses it with a non-zero weight.  Cells are
 * bit-identical to the reference's up to the per-cell double-precision exp of the distance kernel
 * (see csrc/csm_insert_tsdf2d.cu). */
typedef struct csm_tsdfgrid2d csm_tsdfgrid2d;
typedef struct {
  double max_x, max_y, resolution;   /* MapLimits of the grid                                  */
  float origin_x, origin_y;          /* range_data.origin                                      */
  int32_t num_returns;
  int32_t update_free_space;         /* TSDFRangeDataInserterOptions2D                         */
  const float* hits_xy;              /* num_returns x (x, y), in insertion order               */
  const float* normal_cos_sin;       /* num_returns x (cosf(normal), sinf(normal)); may be NULL */
                                     /* unless project_sdf_distance_to_scan_normal             */
  const float* hit_weight;           /* weight_factor_range * weight_factor_angle_ray_normal   */
  float truncation_distance, maximum_weight;
  float distance_kernel_bandwidth;   /* update_weight_distance_cell_to_hit_kernel_bandwidth; 0: off */
  int32_t project_sdf_distance_to_scan_normal;
  float min_tsd, max_tsd, tsd_resolution;      /* TSDValueConverter (tsd_value_converter.h)     */
  float max_weight_value, weight_resolution;   /* its max weight and 32766 / max weight          */
  int32_t pad;
} csm_tsdf_range_data2d;

/* tsd_cells / weight_cells may be NULL (all unknown); the tables are the converter's 65536-entry
 * value-to-TSD and value-to-weight tables. */
int csm_tsdfgrid2d_create(csm_ctx* ctx, const uint16_t* tsd_cells, const uint16_t* weight_cells, int num_x,
                          int num_y, const float* value_to_tsd, const float* value_to_weight,
                          csm_tsdfgrid2d** out);
int csm_tsdfgrid2d_insert(csm_ctx* ctx, csm_tsdfgrid2d* grid, const csm_tsdf_range_data2d* scans, int num_scans);
int csm_tsdfgrid2d_download(csm_ctx* ctx, const csm_tsdfgrid2d* grid, uint16_t* tsd_out, uint16_t* weight_out);
const uint16_t* csm_tsdfgrid2d_tsd_dev(const csm_tsdfgrid2d* grid);
const uint16_t* csm_tsdfgrid2d_weight_dev(const csm_tsdfgrid2d* grid);
void csm_tsdfgrid2d_destroy(csm_ctx* ctx, csm_tsdfgrid2d* grid);

/* ==== 3D path ======================================================================
 * Citations relative to cartographer/mapping/internal/3d/scan_matching/:
 *   FCSM3D = fast_correlative_scan_matcher_3d.{h,cc}
 *   PG3D   = precomputation_grid_3d.{h,cc}
 *   LRM    = low_resolution_matcher.{h,cc}
 * Host side of the boundary (libm-dependent, O(num_scans)): the search parameters
 * (FCSM3D.cc:127-170), the angular step (acos, :255-264), RotationalScanMatcher::Match
 * and the surviving DiscreteScan3D poses (:266-290; rotational_scan_matcher.cc:121-189)
 * and the candidate -> pose conversion (:363-369).
 */
typedef struct csm_stack3d csm_stack3d;
typedef struct csm_grid3d csm_grid3d;
typedef struct csm_batch3d csm_batch3d;

/* ---- PrecomputationGridStack3D (FCSM3D.cc:57-77, PG3D.cc:49-81) -------------------
 * The HybridGrid arrives as the list of its non-default voxels, as
 * HybridGrid::Iterator visits them (cartographer/mapping/3d/hybrid_grid.h:327-419):
 * voxel_xyz num_voxels x 3 int32 cell indices, voxel_values the uint16 cells.
 * q_lut: 65536 bytes, q_lut[v] = lround((ValueToProbability(v) - 0.1f) * (255.f / 0.8f))
 *        (ConvertToPrecomputationGrid, PG3D.cc:49-61), host-computed in fp32.
 * Level k is stored dense over a box that contains every voxel PrecomputeGrid would
 * create; levels k >= full_resolution_depth are at half the resolution of level k-1.
 */
int csm_stack3d_build(csm_ctx* ctx, const int32_t* voxel_xyz, const uint16_t* voxel_values,
                      int num_voxels, const uint8_t* q_lut, int depth, int full_resolution_depth,
                      csm_stack3d** out);
/* Copies level `level` ([z][y][x] bytes, dims3 = nx, ny, nz; lo3 = cell index of element
 * 0) to `out` (may be NULL to query the box only). */
int csm_stack3d_level(csm_ctx* ctx, const csm_stack3d* stack, int level, uint8_t* out, int32_t* lo3,
                      int32_t* dims3);
int csm_stack3d_depth(const csm_stack3d* stack);
size_t csm_stack3d_device_bytes(const csm_stack3d* stack);
void csm_stack3d_destroy(csm_ctx* ctx, csm_stack3d* stack);

/* ---- low-resolution HybridGrid for the leaf filter (LRM.cc:23-35) ------------------
 * Dense fp32 probabilities over the voxels' bounding box; probability_table is the
 * 65536-entry ValueToProbability table, probability_table[0] the value of unknown
 * space (cartographer/mapping/probability_values.cc:29-60). */
int csm_grid3d_build(csm_ctx* ctx, const int32_t* voxel_xyz, const uint16_t* voxel_values,
                     int num_voxels, const float* probability_table, csm_grid3d** out);
void csm_grid3d_destroy(csm_ctx* ctx, csm_grid3d* grid);

/* ---- FastCorrelativeScanMatcher3D::MatchWithSearchParameters (FCSM3D.cc:172-198) ---
 * One job = one (node, submap) search over the discrete scans that survived the
 * rotational prefilter.  The device does DiscretizeScan (:200-244),
 * GenerateLowestResolutionCandidates (:297-326), ScoreCandidates (:328-352) and
 * BranchAndBound (:371-440) including the low-resolution leaf filter (:382-402).
 */
typedef struct {
  const csm_stack3d* stack;
  const csm_grid3d* low_resolution_grid;
  const float* points_xyz;        /* high_resolution_point_cloud, num_points x 3        */
  int32_t num_points;
  const float* low_points_xyz;    /* low_resolution_point_cloud, num_low_points x 3     */
  int32_t num_low_points;
  int32_t num_scans;              /* discrete scans kept by the rotational prefilter    */
  const float* scan_poses;        /* num_scans x 7 fp32 (tx,ty,tz,qw,qx,qy,qz):         */
                                  /* DiscreteScan3D::pose (FCSM3D.cc:282-290)            */
  const float* scan_leaf_quat;    /* num_scans x 4 (w,x,y,z): rotation of               */
                                  /* Translation(..) * scan.pose, i.e. the normalised   */
                                  /* product GetPoseFromCandidate yields (:363-369)     */
  float resolution;               /* HybridGrid::resolution() of the high-res grid      */
  float low_resolution;           /* of the low-resolution grid                         */
  int32_t linear_xy_window;       /* SearchParameters, in voxels (:131-134, :158-164)   */
  int32_t linear_z_window;
  float min_score;                /* strict threshold                                   */
  double min_low_resolution_score;/* options.min_low_resolution_score() (:392)          */
} csm_job3d;

typedef struct {
  int32_t found;
  int32_t scan_index;             /* Candidate3D::scan_index (index among kept scans)   */
  int32_t offset[3];              /* Candidate3D::offset                                */
  float score;                    /* Candidate3D::score (bit-identical fp32)            */
  float low_resolution_score;     /* Candidate3D::low_resolution_score                  */
  int32_t tie_resolution;         /* 0: unique on device; host replay of std::sort over  */
                                  /* equal lowest-resolution scores: 1 among traversal-  */
                                  /* order contenders, 2 among the sweep's recorded leaves*/
  uint64_t candidates_scored;
} csm_result3d;

int csm_match3d_batch(csm_ctx* ctx, const csm_job3d* jobs, int num_jobs, csm_result3d* results);
int csm_batch3d_upload(csm_ctx* ctx, const csm_job3d* jobs, int num_jobs, csm_batch3d** out);
int csm_batch3d_run(csm_ctx* ctx, csm_batch3d* batch);
int csm_batch3d_fetch(csm_ctx* ctx, csm_batch3d* batch, csm_result3d* results);
void csm_batch3d_destroy(csm_ctx* ctx, csm_batch3d* batch);

/* DiscretizeScan alone for the bit-exact cell-index parity test: full-resolution cell
 * indices, num_scans x num_points x 3 int32 (FCSM3D.cc:205-211). */
int csm_discretize3d(csm_ctx* ctx, const csm_job3d* job, int32_t* out_xyz);

/* ---- RealTimeCorrelativeScanMatcher3D::Match
 * (.../3d/scan_matching/real_time_correlative_scan_matcher_3d.cc:35-114) -------------
 * Exhaustive search over the transforms of GenerateExhaustiveSearchTransforms (:55-98).
 * The host generates them (AngleAxis -> quaternion needs libm), composes each with the
 * initial pose and evaluates the motion penalty exp(-(|t| w_t + angle w_r)^2) in double
 * (:108-112); the device transforms the cloud by every candidate pose, looks the
 * probabilities up (grid = csm_grid3d_build of the live HybridGrid), sums them in point
 * order in fp32 (:104-107) and returns the first maximum in generation order (:47-50).
 */
typedef struct {
  const csm_grid3d* grid;
  const float* points_xyz;
  int32_t num_points;
  int32_t num_candidates;
  const float* candidate_poses;   /* num_candidates x 7 (tx,ty,tz,qw,qx,qy,qz): initial * transform */
  const double* penalty;          /* num_candidates                                     */
  float resolution;               /* HybridGrid::resolution()                           */
} csm_rt_job3d;

typedef struct {
  int32_t best_index;             /* generation index of the winner                     */
  float score;                    /* its score (bit-identical fp32)                     */
} csm_rt_result3d;

int csm_rt_match3d_batch(csm_ctx* ctx, const csm_rt_job3d* jobs, int num_jobs, csm_rt_result3d* results);

/* ---- RangeDataInserter3D::Insert on a device-resident HybridGrid (SURVEY 8(f) rank 3, 3D half)
 * Replaces cartographer/mapping/3d/range_data_inserter_3d.cc:44-76 (InsertMissesIntoGrid) and
 * :112-131 (Insert) with HybridGrid::GetCellIndex / ApplyLookupTable / FinishUpdate
 * (mapping/3d/hybrid_grid.h:428-434,492-518).  A csm_hgrid3d keeps a live submap's uint16 voxels in
 * HBM as a dense box [z][y][x] that grows when a range data reaches outside it (the reference's
 * sparse grid has no box; its hard limit of +-8192 cells per axis is kept).  What the path
 * observes -- the set of non-default voxels the reference's HybridGrid::Iterator visits, with their
 * values -- is bit-identical after every insertion; csm_hgrid3d_export returns it (unordered) and
 * is what csm_stack3d_build / csm_grid3d_build take.  The intensity grid of this fork
 * (range_data_inserter_3d.cc:78-92) is not carried. */
typedef struct csm_hgrid3d csm_hgrid3d;
typedef struct {
  float origin[3];                 /* range_data.origin                                   */
  int32_t num_returns;
  const float* returns_xyz;        /* num_returns x (x, y, z)                             */
  const uint16_t* hit_table;       /* ComputeLookupTableToApplyOdds(Odds(p)),             */
  const uint16_t* miss_table;      /* probability_values.cc:72-87, 32768 entries each     */
  int32_t num_free_space_voxels;   /* RangeDataInserterOptions3D                          */
  int32_t pad;
} csm_range_data3d;

/* A grid holding the given voxels (num_voxels may be 0: an empty grid). */
int csm_hgrid3d_create(csm_ctx* ctx, float resolution, const int32_t* voxel_xyz, const uint16_t* voxel_values,
                       int num_voxels, csm_hgrid3d** out);
/* Inserts the scans one after the other (each is one reference Insert incl. FinishUpdate). */
int csm_hgrid3d_insert(csm_ctx* ctx, csm_hgrid3d* grid, const csm_range_data3d* scans, int num_scans);
/* Non-default voxels.  capacity == 0: count only.  *num_voxels is the total even if it exceeds
 * capacity (then only `capacity` voxels were written). */
int csm_hgrid3d_export(csm_ctx* ctx, const csm_hgrid3d* grid, int32_t* voxel_xyz, uint16_t* voxel_values,
                       int capacity, int* num_voxels);
int csm_hgrid3d_box(const csm_hgrid3d* grid, int32_t* lo3, int32_t* dims3);
/* The matcher's stack straight from the device-resident grid (same result as csm_stack3d_build of
 * csm_hgrid3d_export's list): the voxels never cross PCIe. */
int csm_stack3d_build_from_hgrid(csm_ctx* ctx, const csm_hgrid3d* grid, const uint8_t* q_lut, int depth,
                                 int full_resolution_depth, csm_stack3d** out);
void csm_hgrid3d_destroy(csm_ctx* ctx, csm_hgrid3d* grid);

/* ---- measurement aids (bench.py) ------------------------------------------------
 * csm_ctx_launch_count: kernels launched through `ctx` since it was created.
 * csm_measure_gather_rate: the L2 gather roofline G_L2 of the scoring loop
 * (FCSM2D.cc:314-333): a micro-kernel with one warp per candidate and one lane per
 * scan point gathers single bytes from a `table_bytes` table (the size of a
 * precomputation stack, L2-resident) at table[candidate_base + y*row_pitch + x] for
 * the given discretised scan `point_xy` (num_points x 2 int32).  Writes the
 * sustained rate in lookups per second.
 */
unsigned long long csm_ctx_launch_count(csm_ctx* ctx);
/* csm_measure_gather_ceiling: the ceiling of the ACCESS SHAPE of the 2D branch-and-bound sweep
 * (k_expand_nodes): per (work item, scan point) one warp-wide 4-byte read-only load from an
 * L2-resident table, base broadcast from shared memory, lanes spread evenly over
 * `sectors_per_request` consecutive 32-byte sectors (4 = perfectly coalesced, 32 = one sector
 * per lane), the packed accumulate, and nothing else.  Writes warp load requests per second.
 * bench.py evaluates it at the sectors / request the run's own work items have
 * (csm_ctx_get_depth_profile) -- the roofline denominator of the dominant kernel. */
int csm_measure_gather_ceiling(csm_ctx* ctx, size_t table_bytes, int sectors_per_request,
                               unsigned int num_items, int points_per_item, int repeats,
                               double* requests_per_second);
/* Per-depth profile of the 2D sweep since profiling was switched on: CSM_MAX_DEPTH x 8 doubles,
 * per depth d (the depth of the PARENTS a launch expands): [0] k_expand_nodes milliseconds,
 * [1] list bookkeeping milliseconds (seed, bucket prefix, placement), and -- profiling level 2
 * only -- [2] work items, [3] warp load requests (items x points), [4] lane loads, [5] 32-byte
 * sectors and [6] 128-byte lines those requests touch, [7] nodes. */
int csm_ctx_get_depth_profile(csm_ctx* ctx, double* out);
int csm_measure_gather_rate(csm_ctx* ctx, size_t table_bytes, int row_pitch, const int32_t* point_xy,
                            int num_points, unsigned int num_candidates, int repeats,
                            double* lookups_per_second);

#ifdef __cplusplus
}
#endif

#endif /* CSM_B200_H_ */
