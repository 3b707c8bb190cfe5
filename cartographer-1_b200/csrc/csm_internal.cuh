// Internal declarations shared by the CUDA translation units of libcsm_b200.so.
// Not part of the ABI (include/csm_b200.h is).
#ifndef CSM_INTERNAL_CUH_
#define CSM_INTERNAL_CUH_

#include <cuda_runtime.h>
#include <stdint.h>

#include <cstdio>
#include <mutex>
#include <string>
#include <vector>

#include "csm_b200.h"

namespace csm {

void set_error(const char* fmt, ...);

#define CSM_CUDA(call)                                                                 \
  do {                                                                                 \
    cudaError_t _e = (call);                                                           \
    if (_e != cudaSuccess) {                                                           \
      csm::set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(_e), __FILE__, \
                     __LINE__);                                                        \
      return CSM_ERR_CUDA;                                                             \
    }                                                                                  \
  } while (0)

#define CSM_REQUIRE(cond, msg)                                        \
  do {                                                                \
    if (!(cond)) {                                                    \
      csm::set_error("invalid argument: %s (%s)", msg, #cond);        \
      return CSM_ERR_ARG;                                             \
    }                                                                 \
  } while (0)

// Grow-only device / pinned-host scratch buffer owned by a context.
struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
  int reserve(size_t bytes);  // keeps contents only if no regrowth is needed
  void release();
  template <typename T>
  T* as() const {
    return static_cast<T*>(p);
  }
};
struct PinBuf {
  void* p = nullptr;
  size_t cap = 0;
  int reserve(size_t bytes);
  void release();
  template <typename T>
  T* as() const {
    return static_cast<T*>(p);
  }
};

// Free lists of device / pinned blocks a context keeps between batches and stacks: a batch's
// arena and staging buffer are a cudaMalloc + cudaHostAlloc + two frees otherwise (several
// milliseconds per batch end to end), and a loop-closure job that rebuilds the stacks of 100
// submaps per flush would free and re-allocate 3.6 GB every time (cudaFree synchronises the
// device: 77 - 640 ms per 100 stacks were measured before the pool kept them).  Blocks are reused
// when at most 4x oversized.  The pool keeps up to `device_budget` / `pinned_budget` bytes and
// kMaxBlocks blocks per kind (CSM_POOL_DEVICE_MB / CSM_POOL_PINNED_MB override the budgets: 16 GiB
// and 1 GiB by default, of 180 GB of HBM); beyond that the smallest block is released.  take/give
// are internally locked: an object built on a context may be destroyed from another thread than
// the one driving it.
struct BlockPool {
  struct Block {
    void* p;
    size_t cap;
  };
  static constexpr size_t kMaxBlocks = 1024;
  size_t device_budget = size_t(16) << 30, pinned_budget = size_t(1) << 30;
  size_t device_bytes = 0, pinned_bytes = 0;
  std::vector<Block> device, pinned;
  std::mutex mu;
  int take(bool pin, size_t bytes, void** out, size_t* cap);
  void give(bool pin, void* p, size_t cap);
  void release();
};

// True while `ctx` is a live context on `device` (registry of csm_ctx_create/destroy):
// destroy entry points use it to decide whether a block can go back to a pool.
bool ctx_alive_on(const csm_ctx* ctx, int device);
// csm_insert3d.cu: writes the non-default voxels of a device-resident HybridGrid into ctx->tmp0
// (xyz, int32) / ctx->tmp1 (values, uint16) -- where csm_stack3d_build stages a host list -- and
// returns their number and bounding box (inclusive).  Synchronises the context's stream.
int hgrid3d_stage(::csm_ctx* ctx, const ::csm_hgrid3d* grid, int* num_voxels, int* lo3, int* hi3);
// Hands a device block back to the pool of `ctx` if it is a live context on `device` (under the
// registry lock); false: the caller frees the block.
bool give_if_alive(::csm_ctx* ctx, int device, void* p, size_t cap);

}  // namespace csm

struct csm_ctx {
  int device = 0;
  int num_sms = 0;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  // scratch for csm_discretize2d / stack builds
  csm::DevBuf tmp0, tmp1, tmp2;
  csm::PinBuf pin0;
  csm::BlockPool pool;
  // Branch-and-bound frontier ping-pong buffers, shared by every batch that runs on
  // this context (runs are stream-ordered, the frontier is pure scratch).
  csm::DevBuf frontier[2];
  csm::DevBuf frontier_aux;             // 2D node sweep: block descriptors | work items
  bool sweep_tiles = false;             // CSM_SWEEP=tiles: the tile sweep instead of node lists
  bool expand_noalloc = false;          // CSM_EXPAND_LOAD=noalloc: L1::no_allocate loads in the sweep
  bool phase2_dfs = false;              // CSM_PHASE2=dfs: the ordered walk over top-level contenders
                                        // instead of sibling-rank paths carried through the sweep
  unsigned long long frontier_cap = 0;  // unused (capacities come from the buffers)
  // Work-list sizes learned from earlier batches on this context, per lowest-resolution
  // candidate: a batch that overflowed a list (and was replayed) makes the next ones ask for
  // enough up front.
  double frontier_per_candidate = 0.5, contenders_per_candidate = 0.;
  // Optional per-phase CUDA-event timing of csm_batch2d_run (csm_ctx_set_profiling).
  static constexpr int kPhases = 5;  // discretize, score_top, dive, expand, phase2
  int profiling = 0;  // 0 off, 1 phase events, 2 + tile histogram
  int keep_cells = 1;  // stack builds leave the grid's cells beside the levels (csm_ctx_set_keep_cells)
  double phase_ms[kPhases] = {0, 0, 0, 0, 0};
  double phase_scored[kPhases] = {0, 0, 0, 0, 0};  // candidate scores evaluated per phase
  int profiled_runs = 0;
  unsigned long long launches = 0;  // kernels launched through this context so far
  double tile_hist[CSM_MAX_DEPTH * 33] = {0};  // expanded tiles by [depth][alive nodes]
  // Per-depth profile of the 2D node sweep (csm_ctx_get_depth_profile): kDepthFields doubles
  // per depth = expand kernel ms, list bookkeeping ms (seed / prefix / place), and -- with
  // profiling 2 -- work items, warp load requests, lane loads, 32-byte sectors and 128-byte
  // lines those requests touch, nodes.
  static constexpr int kDepthFields = 8;
  double depth_profile[CSM_MAX_DEPTH * kDepthFields] = {0};
};

namespace csm {
// Device-resident description of one precomputation stack (what the kernels read).
//
// level[l]: the reference's PrecomputationGrid2D::cells_ of width 2^l, row-major bytes
//           with rows pitched to 16 bytes (host side: csm_stack2d::level_pitch).
// quad[l]:  the same level re-laid out for the matcher ("quad-phase" layout): one
//           uint32 per window start (x, y), x in [-2s+1, num_x-1], s = 2^l, holding the
//           four values a branch-and-bound parent needs for its children in generation
//           order: byte0 = G(x, y), byte1 = G(x, y+s), byte2 = G(x+s, y), byte3 =
//           G(x+s, y+s).  Words are grouped by phase (u mod 2s, v mod 2s) with
//           u = x + 2s - 1, v = y + 2s - 1, so that parents 2s cells apart -- neighbours
//           on the search lattice -- sit in neighbouring words:
//           index = ((v%2s * 2s + u%2s) * jn + v/2s) * in + u/2s.
struct StackDev {
  int num_x, num_y, depth, pad;
  const uint8_t* level[CSM_MAX_DEPTH];
  const uint32_t* quad[CSM_MAX_DEPTH];
  int in[CSM_MAX_DEPTH], jn[CSM_MAX_DEPTH];
};
}  // namespace csm

struct csm_stack2d {
  int device = 0;
  int num_x = 0, num_y = 0, depth = 0;
  uint8_t* base = nullptr;                // one allocation: levels, quads, descriptor, LUT
  size_t base_cap = 0;                    // capacity of that block (pooled)
  size_t level_off[CSM_MAX_DEPTH] = {0};  // byte offset of level i inside base
  int level_pitch[CSM_MAX_DEPTH] = {0};   // row pitch of level i in bytes (multiple of 16)
  size_t quad_off[CSM_MAX_DEPTH] = {0};   // byte offset of quad level i inside base
  csm::StackDev host_view;                // copy of *dev
  csm::StackDev* dev = nullptr;           // device-resident descriptor
  size_t bytes = 0;
  uint8_t* q_lut_dev = nullptr;  // 65536 bytes
  uint16_t* cells_dev = nullptr;  // the grid's uint16 cells as given to the build (csm_stack2d_cells_dev)
};

#endif  // CSM_INTERNAL_CUH_
