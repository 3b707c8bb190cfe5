// ProbabilityGridRangeDataInserter2D::Insert on a device-resident grid (SURVEY.md section
// 8(f) rank 3: submap grids born on the GPU, so that the precomputation stack of a finished
// submap is built without the cells ever crossing PCIe).
//
// Replaces /root/reference/cartographer/mapping/2d/probability_grid_range_data_inserter_2d.cc:
// 47-100 (CastRays: hits first, then the free-space rays of the returns, then of the misses)
// and /root/reference/cartographer/mapping/internal/2d/ray_to_pixel_mask.cc:25-170 (the exact
// set of pixels a sub-pixel line segment crosses, integer arithmetic at kSubpixelScale = 1000),
// plus Grid2D::FinishUpdate (grid_2d.cc:98-106).  GrowLimits (probability_grid_range_data_
// inserter_2d.cc:30-45, grid_2d.cc:131-166) stays with the host class: it only re-allocates.
//
// The reference applies a lookup table to a cell at most once per insertion (the update
// marker), hits before misses.  The result therefore does not depend on the order of the
// rays: a cell ends as hit_table[cell] if any return falls into it, else miss_table[cell] if
// any ray crosses it.  On the device every ray is walked by one thread that only ORs a flag
// byte per crossed cell (fire-and-forget reductions, no compare-and-swap loops), and one
// pass over the grid applies the tables, drops the marker and clears the flags.  Integer
// work end to end: cells are bit-identical to the reference's.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <vector>

#include "csm_internal.cuh"

struct csm_grid2d {
  int device = 0;
  int num_x = 0, num_y = 0;
  uint16_t* cells = nullptr;     // num_x * num_y
  uint32_t* flags = nullptr;     // one byte per cell (bit 0 hit, bit 1 miss), padded to words
  uint16_t* tables = nullptr;    // hit table | miss table, 32768 entries each
  float* points = nullptr;       // staging for ray end points
  size_t points_cap = 0;
};

namespace {

constexpr int kSubpixelScale = 1000;  // probability_grid_range_data_inserter_2d.cc:31
constexpr uint16_t kUpdateMarker = 1u << 15;

struct InsertArgs {
  double max_x, max_y, super_res;
  float origin_x, origin_y;
  int num_x, num_y, num_returns, num_rays, insert_free_space;
};

// MapLimits::GetCellIndex on the superscaled limits (map_limits.h:69-76): lround, fp64.
__device__ __forceinline__ int2 super_cell(const InsertArgs& a, float px, float py) {
  int2 c;  // common::RoundToInt == lround: half away from zero
  c.x = static_cast<int>(llround(__dsub_rn(__ddiv_rn(__dsub_rn(a.max_y, static_cast<double>(py)), a.super_res), 0.5)));
  c.y = static_cast<int>(llround(__dsub_rn(__ddiv_rn(__dsub_rn(a.max_x, static_cast<double>(px)), a.super_res), 0.5)));
  return c;
}

__device__ __forceinline__ void mark(uint32_t* __restrict__ flags, const InsertArgs& a, int x, int y,
                                     uint32_t bit) {
  if (static_cast<unsigned>(x) >= static_cast<unsigned>(a.num_x) ||
      static_cast<unsigned>(y) >= static_cast<unsigned>(a.num_y))
    return;  // the reference DCHECKs containment; the host grew the grid beforehand
  const size_t flat = static_cast<size_t>(y) * a.num_x + x;
  atomicOr(&flags[flat >> 2], bit << (8 * (flat & 3)));
}

// ray_to_pixel_mask.cc:25-170 with the pixels marked instead of listed.
__device__ void mark_ray(uint32_t* __restrict__ flags, const InsertArgs& a, int2 b, int2 e) {
  if (b.x > e.x) {  // "For simplicity, we order 'scaled_begin' and 'scaled_end' by their x coordinate."
    const int2 t = b;
    b = e;
    e = t;
  }
  if (b.x < 0 || b.y < 0 || e.y < 0) return;  // CHECK_GE in the reference
  const int scale = kSubpixelScale;
  if (b.x / scale == e.x / scale) {  // vertical line in full pixels
    const int x = b.x / scale;
    const int y1 = max(b.y, e.y) / scale;
    for (int y = min(b.y, e.y) / scale; y <= y1; ++y) mark(flags, a, x, y, 2u);
    return;
  }
  const long long dx = e.x - b.x, dy = e.y - b.y;
  const long long denom = 2ll * scale * dx;
  int cx = b.x / scale, cy = b.y / scale;
  mark(flags, a, cx, cy, 2u);
  long long sub_y = (2ll * (b.y % scale) + 1) * dx;
  const int first_pixel = 2 * scale - 2 * (b.x % scale) - 1;
  const int last_pixel = 2 * (e.x % scale) + 1;
  const int end_x = max(b.x, e.x) / scale;
  sub_y += dy * first_pixel;
  const int dir = dy > 0 ? 1 : -1;
  // Ascending rays step while sub_y exceeds the denominator, the others while it is
  // negative; on an exact pixel corner the step happens after moving to the next column.
  for (;;) {
    mark(flags, a, cx, cy, 2u);
    while (dir > 0 ? sub_y > denom : sub_y < 0) {
      sub_y -= dir * denom;
      cy += dir;
      mark(flags, a, cx, cy, 2u);
    }
    ++cx;
    if (dir > 0 ? sub_y == denom : sub_y == 0) {
      sub_y -= dir * denom;
      cy += dir;
    }
    if (cx == end_x) break;
    sub_y += dy * 2 * scale;
  }
  sub_y += dy * last_pixel;
  mark(flags, a, cx, cy, 2u);
  while (dir > 0 ? sub_y > denom : sub_y < 0) {
    sub_y -= dir * denom;
    cy += dir;
    mark(flags, a, cx, cy, 2u);
  }
}

// One thread per ray: rays [0, num_returns) end in a hit, the rest are misses.
__global__ void __launch_bounds__(128)
k_insert_mark(const InsertArgs a, const float* __restrict__ ends_xy, uint32_t* __restrict__ flags) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= a.num_rays) return;
  const int2 end = super_cell(a, ends_xy[2 * r], ends_xy[2 * r + 1]);
  if (r < a.num_returns) mark(flags, a, end.x / kSubpixelScale, end.y / kSubpixelScale, 1u);
  if (!a.insert_free_space) return;
  mark_ray(flags, a, super_cell(a, a.origin_x, a.origin_y), end);
}

// cell = table[cell] - kUpdateMarker for every flagged cell (ApplyLookupTable + FinishUpdate);
// flags are cleared for the next insertion.  4 cells per thread, over the flag words
// [word0, word1) of the rows the scan can have touched.
__global__ void __launch_bounds__(256)
k_insert_apply(uint16_t* __restrict__ cells, uint32_t* __restrict__ flags,
               const uint16_t* __restrict__ tables, size_t num_cells, size_t word0, size_t word1) {
  const size_t w = word0 + static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (w >= word1 || 4 * w >= num_cells) return;
  const uint32_t f = flags[w];
  if (f == 0u) return;
  flags[w] = 0u;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const uint32_t fk = (f >> (8 * k)) & 3u;
    const size_t i = 4 * w + k;
    if (fk == 0u || i >= num_cells) continue;
    const uint16_t old = cells[i];
    if (old >= kUpdateMarker) continue;  // cannot happen between insertions (grid_2d.cc:98-106)
    const uint16_t* t = tables + ((fk & 1u) ? 0 : 32768);
    cells[i] = static_cast<uint16_t>(t[old] - kUpdateMarker);
  }
}

}  // namespace

extern "C" {

int csm_grid2d_create(csm_ctx* ctx, const uint16_t* cells, int num_x, int num_y, csm_grid2d** out) {
  CSM_REQUIRE(ctx != nullptr && out != nullptr, "ctx/out must not be NULL");
  CSM_REQUIRE(num_x >= 1 && num_y >= 1, "cell limits must be positive");
  CSM_CUDA(cudaSetDevice(ctx->device));
  csm_grid2d* g = new csm_grid2d;
  g->device = ctx->device;
  g->num_x = num_x;
  g->num_y = num_y;
  const size_t n = static_cast<size_t>(num_x) * num_y;
  const size_t flag_words = (n + 3) / 4;
  if (cudaMalloc(&g->cells, sizeof(uint16_t) * n + 16) != cudaSuccess ||
      cudaMalloc(&g->flags, sizeof(uint32_t) * flag_words) != cudaSuccess ||
      cudaMalloc(&g->tables, sizeof(uint16_t) * 65536) != cudaSuccess) {
    csm::set_error("allocation of a %d x %d grid failed: %s", num_x, num_y,
                   cudaGetErrorString(cudaGetLastError()));
    cudaFree(g->cells);
    cudaFree(g->flags);
    cudaFree(g->tables);
    delete g;
    return CSM_ERR_CUDA;
  }
  if (cells != nullptr) {
    CSM_CUDA(cudaMemcpyAsync(g->cells, cells, sizeof(uint16_t) * n, cudaMemcpyHostToDevice, ctx->stream));
  } else {
    CSM_CUDA(cudaMemsetAsync(g->cells, 0, sizeof(uint16_t) * n, ctx->stream));  // kUnknownCorrespondenceValue
  }
  CSM_CUDA(cudaMemsetAsync(g->flags, 0, sizeof(uint32_t) * flag_words, ctx->stream));
  CSM_CUDA(cudaStreamSynchronize(ctx->stream));
  *out = g;
  return CSM_OK;
}

int csm_grid2d_download(csm_ctx* ctx, const csm_grid2d* g, uint16_t* cells_out) {
  CSM_REQUIRE(ctx != nullptr && g != nullptr && cells_out != nullptr, "NULL argument");
  CSM_CUDA(cudaSetDevice(g->device));
  CSM_CUDA(cudaMemcpyAsync(cells_out, g->cells, sizeof(uint16_t) * static_cast<size_t>(g->num_x) * g->num_y,
                           cudaMemcpyDeviceToHost, ctx->stream));
  CSM_CUDA(cudaStreamSynchronize(ctx->stream));
  return CSM_OK;
}

const uint16_t* csm_grid2d_cells_dev(const csm_grid2d* g) { return g ? g->cells : nullptr; }

void csm_grid2d_destroy(csm_ctx* ctx, csm_grid2d* g) {
  if (!g) return;
  cudaSetDevice(g->device);
  if (csm::ctx_alive_on(ctx, g->device)) cudaStreamSynchronize(ctx->stream);
  else cudaDeviceSynchronize();
  cudaFree(g->cells);
  cudaFree(g->flags);
  cudaFree(g->tables);
  cudaFree(g->points);
  delete g;
}

int csm_grid2d_insert(csm_ctx* ctx, csm_grid2d* g, const csm_range_data2d* scans, int num_scans) {
  CSM_REQUIRE(ctx != nullptr && g != nullptr && scans != nullptr, "NULL argument");
  CSM_REQUIRE(num_scans >= 1, "num_scans must be positive");
  CSM_REQUIRE(g->device == ctx->device, "grid lives on another device");
  CSM_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t st = ctx->stream;
  const size_t n = static_cast<size_t>(g->num_x) * g->num_y;
  size_t total_rays = 0;
  for (int s = 0; s < num_scans; ++s) {
    const csm_range_data2d& in = scans[s];
    CSM_REQUIRE(in.hit_table != nullptr && in.miss_table != nullptr, "lookup tables must not be NULL");
    CSM_REQUIRE(in.num_returns >= 0 && in.num_misses >= 0 && in.resolution > 0., "range data sizes");
    CSM_REQUIRE((in.num_returns == 0 || in.returns_xy) && (in.num_misses == 0 || in.misses_xy),
                "range data arrays must not be NULL");
    total_rays += static_cast<size_t>(in.num_returns) + in.num_misses;
  }
  if (total_rays == 0) return CSM_OK;
  // The end points of every scan go up in ONE copy from pinned memory; the kernels of the
  // scans then run back to back on the stream (scan s + 1 must see the cells scan s left).
  if (g->points_cap < total_rays) {
    CSM_CUDA(cudaStreamSynchronize(st));
    cudaFree(g->points);
    g->points = nullptr;
    g->points_cap = 0;
    CSM_CUDA(cudaMalloc(&g->points, sizeof(float) * 2 * (total_rays + 4096)));
    g->points_cap = total_rays + 4096;
  }
  int rc = ctx->pin0.reserve(sizeof(float) * 2 * total_rays);
  if (rc != CSM_OK) return rc;
  float* pin = ctx->pin0.as<float>();
  {
    size_t at = 0;
    for (int s = 0; s < num_scans; ++s) {
      const csm_range_data2d& in = scans[s];
      if (in.num_returns) memcpy(pin + 2 * at, in.returns_xy, sizeof(float) * 2 * in.num_returns);
      at += in.num_returns;
      if (in.num_misses) memcpy(pin + 2 * at, in.misses_xy, sizeof(float) * 2 * in.num_misses);
      at += in.num_misses;
    }
  }
  CSM_CUDA(cudaMemcpyAsync(g->points, pin, sizeof(float) * 2 * total_rays, cudaMemcpyHostToDevice, st));
  const uint16_t* loaded_hit = nullptr;
  const uint16_t* loaded_miss = nullptr;
  size_t at = 0;
  for (int s = 0; s < num_scans; ++s) {
    const csm_range_data2d& in = scans[s];
    const int rays = in.num_returns + in.num_misses;
    if (rays == 0) continue;
    if (in.hit_table != loaded_hit || in.miss_table != loaded_miss) {  // pageable: synchronous copies
      CSM_CUDA(cudaMemcpyAsync(g->tables, in.hit_table, sizeof(uint16_t) * 32768, cudaMemcpyHostToDevice, st));
      CSM_CUDA(cudaMemcpyAsync(g->tables + 32768, in.miss_table, sizeof(uint16_t) * 32768,
                               cudaMemcpyHostToDevice, st));
      loaded_hit = in.hit_table;
      loaded_miss = in.miss_table;
    }
    InsertArgs a;
    a.max_x = in.max_x;
    a.max_y = in.max_y;
    a.super_res = in.resolution / kSubpixelScale;
    a.origin_x = in.origin_x;
    a.origin_y = in.origin_y;
    a.num_x = g->num_x;
    a.num_y = g->num_y;
    a.num_returns = in.num_returns;
    a.num_rays = rays;
    a.insert_free_space = in.insert_free_space;
    // Rows the rays can touch: between the rows of the origin and of every end point (the
    // cell row comes from the world x, map_limits.h:69-76), one row of margin.
    double lo_x = in.origin_x, hi_x = in.origin_x;
    const float* ends = pin + 2 * at;
    for (int r = 0; r < rays; ++r) {
      lo_x = std::min<double>(lo_x, ends[2 * r]);
      hi_x = std::max<double>(hi_x, ends[2 * r]);
    }
    long long row0 = static_cast<long long>(std::floor((in.max_x - hi_x) / in.resolution)) - 2;
    long long row1 = static_cast<long long>(std::floor((in.max_x - lo_x) / in.resolution)) + 2;
    row0 = std::max<long long>(row0, 0);
    row1 = std::min<long long>(row1, g->num_y - 1);
    ++ctx->launches, k_insert_mark<<<(rays + 127) / 128, 128, 0, st>>>(a, g->points + 2 * at, g->flags);
    if (row1 >= row0) {
      const size_t w0 = static_cast<size_t>(row0) * g->num_x / 4;
      const size_t w1 = (static_cast<size_t>(row1 + 1) * g->num_x + 3) / 4;  // exclusive
      ++ctx->launches, k_insert_apply<<<static_cast<unsigned int>((w1 - w0 + 255) / 256), 256, 0, st>>>(
          g->cells, g->flags, g->tables, n, w0, w1);
    }
    at += rays;
  }
  CSM_CUDA(cudaGetLastError());
  CSM_CUDA(cudaStreamSynchronize(st));
  return CSM_OK;
}

}  // extern "C"
