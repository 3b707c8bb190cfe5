// TSDFRangeDataInserter2D::Insert on a device-resident TSDF2D (SURVEY.md 8(f) rank 3 remainder).
//
// Reference: cartographer/mapping/internal/2d/tsdf_range_data_inserter_2d.cc:131-239 with
// ray_to_pixel_mask.cc:25-170, TSDF2D::SetCell / GetTSDAndWeight / CellIsUpdated
// (mapping/internal/2d/tsdf_2d.cc:44-98) and TSDValueConverter (tsd_value_converter.h:38-66).
//
// What the reference does: the returns are sorted by bearing, a normal is estimated per return
// (normal_estimation_2d.cc), and then the returns are inserted ONE AFTER THE OTHER; every return
// walks the cells of a short ray around it and updates each cell that no earlier return of this
// insertion has updated (CellIsUpdated, :207; a zero update weight leaves the cell open, :228).
// So a cell belongs to the FIRST return, in sorted order, whose ray crosses it with a non-zero
// weight.  On the device:
//   pass 0  one thread per return walks its ray and, for every cell it would update, takes
//           owner[cell] = min(owner[cell], index of the return)
//   pass 1  the same walk; the thread that finds its own index in owner[cell] swaps it back to
//           "free" and writes the cell (weighted running mean of the TSD, capped weight)
// Sorting, normals and everything else that is per RETURN and goes through libm (atan2f, cosf,
// sinf, powf, the angle kernel's exp) is computed by the host class with the reference's
// expressions and arrives as per-return arrays.  Per CELL the arithmetic is IEEE float / double
// operations in the reference's order (division, sqrt, multiply, add: correctly rounded on both
// sides, no contraction) -- with ONE exception: the distance kernel
// exp(-0.5 x^2 / sigma^2) (GaussianKernel, :51-53) is evaluated per cell in double.  CUDA's exp
// and glibc's are both faithful to about half an ulp but not bit-identical; the factor is
// narrowed to float right away, so a difference survives only when the double lands within one
// ulp of a float rounding boundary (about 2^-29 of the evaluations) and then has to move a 15-bit
// quantised cell value.  glibc itself picks different exp kernels on CPUs with and without FMA,
// so the reference is not bit-stable there either.  Tests compare every cell bit for bit.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <vector>

#include "csm_internal.cuh"

using namespace csm;

struct csm_tsdfgrid2d {
  int device = 0;
  int num_x = 0, num_y = 0;
  uint16_t* tsd = nullptr;     // num_x * num_y
  uint16_t* weight = nullptr;  // num_x * num_y
  int* owner = nullptr;        // num_x * num_y, kFree between insertions
  float* tables = nullptr;     // value -> TSD (65536), value -> weight (65536)
  float* staging = nullptr;    // per-return arrays of one insert call
  size_t staging_cap = 0;
};

namespace {

constexpr int kSubpixelScale = 1000;  // tsdf_range_data_inserter_2d.cc:29
constexpr int kFree = 0x7FFFFFFF;

struct TsdfArgs {
  double max_x, max_y, resolution, super_res;
  double kernel_scale;     // 1.0 / (kSqrtTwoPi * sigma), fp64 of an fp32 product
  double kernel_sigma2;    // sigma * sigma, fp32 product widened
  float origin_x, origin_y;
  float truncation_distance, maximum_weight;
  float min_tsd, max_tsd, tsd_resolution;  // TSDValueConverter
  float max_weight_value, weight_resolution;
  int num_x, num_y, num_returns;
  int update_free_space, project, use_distance_kernel;
};

__device__ __forceinline__ int2 super_cell(const TsdfArgs& a, float px, float py) {
  int2 c;  // MapLimits::GetCellIndex on the superscaled limits (map_limits.h:69-76)
  c.x = static_cast<int>(llround(__dsub_rn(__ddiv_rn(__dsub_rn(a.max_y, static_cast<double>(py)), a.super_res), 0.5)));
  c.y = static_cast<int>(llround(__dsub_rn(__ddiv_rn(__dsub_rn(a.max_x, static_cast<double>(px)), a.super_res), 0.5)));
  return c;
}

// ray_to_pixel_mask.cc:25-170: visit(x, y) for every pixel the segment crosses (a pixel may be
// visited twice in a row; the visitors below are idempotent).
template <typename F>
__device__ void walk_ray(int2 b, int2 e, F&& visit) {
  if (b.x > e.x) {
    const int2 t = b;
    b = e;
    e = t;
  }
  if (b.x < 0 || b.y < 0 || e.y < 0) return;  // CHECK_GE in the reference
  const int scale = kSubpixelScale;
  if (b.x / scale == e.x / scale) {
    const int x = b.x / scale;
    const int y1 = max(b.y, e.y) / scale;
    for (int y = min(b.y, e.y) / scale; y <= y1; ++y) visit(x, y);
    return;
  }
  const long long dx = e.x - b.x, dy = e.y - b.y;
  const long long denom = 2ll * scale * dx;
  int cx = b.x / scale, cy = b.y / scale;
  visit(cx, cy);
  long long sub_y = (2ll * (b.y % scale) + 1) * dx;
  const int first_pixel = 2 * scale - 2 * (b.x % scale) - 1;
  const int last_pixel = 2 * (e.x % scale) + 1;
  const int end_x = max(b.x, e.x) / scale;
  sub_y += dy * first_pixel;
  const int dir = dy > 0 ? 1 : -1;
  for (;;) {
    visit(cx, cy);
    while (dir > 0 ? sub_y > denom : sub_y < 0) {
      sub_y -= dir * denom;
      cy += dir;
      visit(cx, cy);
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
  visit(cx, cy);
  while (dir > 0 ? sub_y > denom : sub_y < 0) {
    sub_y -= dir * denom;
    cy += dir;
    visit(cx, cy);
  }
}

__device__ __forceinline__ float clampf(float v, float lo, float hi) {  // common::Clamp (math.h:31-40)
  if (v > hi) return hi;
  if (v < lo) return lo;
  return v;
}

// One thread per return; hits are in the reference's insertion order.
__global__ void __launch_bounds__(128)
k_tsdf_insert(const TsdfArgs a, const float* __restrict__ hits_xy, const float* __restrict__ normal_cs,
              const float* __restrict__ hit_weight, uint16_t* __restrict__ tsd_cells,
              uint16_t* __restrict__ weight_cells, int* __restrict__ owner, const float* __restrict__ tables,
              int pass) {
  const int h = blockIdx.x * blockDim.x + threadIdx.x;
  if (h >= a.num_returns) return;
  const float hx = hits_xy[2 * h], hy = hits_xy[2 * h + 1];
  // InsertHit, :171-186
  const float rx = __fsub_rn(hx, a.origin_x), ry = __fsub_rn(hy, a.origin_y);
  const float range = __fsqrt_rn(__fadd_rn(__fmul_rn(rx, rx), __fmul_rn(ry, ry)));
  if (range < a.truncation_distance) return;
  const float ratio = __fdiv_rn(a.truncation_distance, range);
  float bx = a.origin_x, by = a.origin_y;
  if (!a.update_free_space) {
    const float f = __fsub_rn(1.0f, ratio);
    bx = __fadd_rn(a.origin_x, __fmul_rn(f, rx));
    by = __fadd_rn(a.origin_y, __fmul_rn(f, ry));
  }
  const float g = __fadd_rn(1.0f, ratio);
  const float ex = __fadd_rn(a.origin_x, __fmul_rn(g, rx)), ey = __fadd_rn(a.origin_y, __fmul_rn(g, ry));
  const float ncos = a.project ? normal_cs[2 * h] : 0.f, nsin = a.project ? normal_cs[2 * h + 1] : 0.f;
  const float base_weight = hit_weight[h];  // weight_factor_range * weight_factor_angle_ray_normal
  walk_ray(super_cell(a, bx, by), super_cell(a, ex, ey), [&](int x, int y) {
    if (static_cast<unsigned>(x) >= static_cast<unsigned>(a.num_x) ||
        static_cast<unsigned>(y) >= static_cast<unsigned>(a.num_y))
      return;  // the host grew the grid beforehand (GrowAsNeeded, :33-48)
    const size_t flat = static_cast<size_t>(y) * a.num_x + x;
    if (pass == 1 && owner[flat] != h) return;
    // :208-223.  MapLimits::GetCellCenter (map_limits.h:79-83): fp64, narrowed.
    const float cx = __double2float_rn(__dsub_rn(a.max_x, __dmul_rn(a.resolution, static_cast<double>(y) + 0.5)));
    const float cy = __double2float_rn(__dsub_rn(a.max_y, __dmul_rn(a.resolution, static_cast<double>(x) + 0.5)));
    float update_tsd;
    if (a.project) {
      update_tsd = __fadd_rn(__fmul_rn(__fsub_rn(cx, hx), ncos), __fmul_rn(__fsub_rn(cy, hy), nsin));
    } else {
      const float ox = __fsub_rn(cx, a.origin_x), oy = __fsub_rn(cy, a.origin_y);
      update_tsd = __fsub_rn(range, __fsqrt_rn(__fadd_rn(__fmul_rn(ox, ox), __fmul_rn(oy, oy))));
    }
    update_tsd = clampf(update_tsd, -a.truncation_distance, a.truncation_distance);
    float update_weight = base_weight;
    if (a.use_distance_kernel) {
      // GaussianKernel(update_tsd, sigma): 1.0 / (kSqrtTwoPi * sigma) * exp(-0.5 * x * x / (sigma * sigma))
      const double xd = static_cast<double>(update_tsd);
      const double arg = __ddiv_rn(__dmul_rn(__dmul_rn(-0.5, xd), xd), a.kernel_sigma2);
      update_weight = __fmul_rn(update_weight, __double2float_rn(__dmul_rn(a.kernel_scale, exp(arg))));
    }
    if (update_weight == 0.f) return;  // UpdateCell, :228: the cell stays open for a later return
    if (pass == 0) {
      atomicMin(owner + flat, h);
      return;
    }
    if (atomicCAS(owner + flat, h, kFree) != h) return;  // a second visit of the same pixel
    // UpdateCell, :229-238
    const float old_tsd = tables[tsd_cells[flat]], old_weight = tables[65536 + weight_cells[flat]];
    float updated_weight = __fadd_rn(old_weight, update_weight);
    const float updated_sdf =
        __fdiv_rn(__fadd_rn(__fmul_rn(old_tsd, old_weight), __fmul_rn(update_tsd, update_weight)), updated_weight);
    updated_weight = fminf(updated_weight, a.maximum_weight);
    // TSDF2D::SetCell + FinishUpdate (tsdf_2d.cc:51-68): the quantised values without the marker.
    tsd_cells[flat] = static_cast<uint16_t>(
        static_cast<int>(roundf(__fmul_rn(__fsub_rn(clampf(updated_sdf, a.min_tsd, a.max_tsd), a.min_tsd),
                                          a.tsd_resolution))) + 1);
    weight_cells[flat] = static_cast<uint16_t>(
        static_cast<int>(roundf(__fmul_rn(__fsub_rn(clampf(updated_weight, 0.f, a.max_weight_value), 0.f),
                                          a.weight_resolution))) + 1);
  });
}

__global__ void k_fill_int(int* p, size_t n, int v) {
  const size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}

}  // namespace

extern "C" {

int csm_tsdfgrid2d_create(csm_ctx* ctx, const uint16_t* tsd_cells, const uint16_t* weight_cells, int num_x,
                          int num_y, const float* value_to_tsd, const float* value_to_weight,
                          csm_tsdfgrid2d** out) {
  CSM_REQUIRE(ctx != nullptr && out != nullptr && value_to_tsd != nullptr && value_to_weight != nullptr,
              "NULL argument");
  CSM_REQUIRE(num_x >= 1 && num_y >= 1, "cell limits must be positive");
  CSM_CUDA(cudaSetDevice(ctx->device));
  csm_tsdfgrid2d* g = new csm_tsdfgrid2d;
  g->device = ctx->device;
  g->num_x = num_x;
  g->num_y = num_y;
  const size_t n = static_cast<size_t>(num_x) * num_y;
  if (cudaMalloc(&g->tsd, sizeof(uint16_t) * n) != cudaSuccess ||
      cudaMalloc(&g->weight, sizeof(uint16_t) * n) != cudaSuccess ||
      cudaMalloc(&g->owner, sizeof(int) * n) != cudaSuccess ||
      cudaMalloc(&g->tables, sizeof(float) * 2 * 65536) != cudaSuccess) {
    set_error("allocation of a %d x %d TSDF2D failed: %s", num_x, num_y, cudaGetErrorString(cudaGetLastError()));
    cudaFree(g->tsd);
    cudaFree(g->weight);
    cudaFree(g->owner);
    cudaFree(g->tables);
    delete g;
    return CSM_ERR_CUDA;
  }
  cudaStream_t st = ctx->stream;
  if (tsd_cells != nullptr) {
    CSM_CUDA(cudaMemcpyAsync(g->tsd, tsd_cells, sizeof(uint16_t) * n, cudaMemcpyHostToDevice, st));
  } else {
    CSM_CUDA(cudaMemsetAsync(g->tsd, 0, sizeof(uint16_t) * n, st));
  }
  if (weight_cells != nullptr) {
    CSM_CUDA(cudaMemcpyAsync(g->weight, weight_cells, sizeof(uint16_t) * n, cudaMemcpyHostToDevice, st));
  } else {
    CSM_CUDA(cudaMemsetAsync(g->weight, 0, sizeof(uint16_t) * n, st));
  }
  CSM_CUDA(cudaMemcpyAsync(g->tables, value_to_tsd, sizeof(float) * 65536, cudaMemcpyHostToDevice, st));
  CSM_CUDA(cudaMemcpyAsync(g->tables + 65536, value_to_weight, sizeof(float) * 65536, cudaMemcpyHostToDevice, st));
  ++ctx->launches, k_fill_int<<<static_cast<unsigned int>((n + 255) / 256), 256, 0, st>>>(g->owner, n, kFree);
  CSM_CUDA(cudaStreamSynchronize(st));
  *out = g;
  return CSM_OK;
}

void csm_tsdfgrid2d_destroy(csm_ctx* ctx, csm_tsdfgrid2d* g) {
  if (!g) return;
  cudaSetDevice(g->device);
  if (ctx_alive_on(ctx, g->device)) cudaStreamSynchronize(ctx->stream);
  else cudaDeviceSynchronize();
  cudaFree(g->tsd);
  cudaFree(g->weight);
  cudaFree(g->owner);
  cudaFree(g->tables);
  cudaFree(g->staging);
  delete g;
}

int csm_tsdfgrid2d_download(csm_ctx* ctx, const csm_tsdfgrid2d* g, uint16_t* tsd_out, uint16_t* weight_out) {
  CSM_REQUIRE(ctx != nullptr && g != nullptr && tsd_out != nullptr && weight_out != nullptr, "NULL argument");
  CSM_CUDA(cudaSetDevice(g->device));
  const size_t bytes = sizeof(uint16_t) * static_cast<size_t>(g->num_x) * g->num_y;
  CSM_CUDA(cudaMemcpyAsync(tsd_out, g->tsd, bytes, cudaMemcpyDeviceToHost, ctx->stream));
  CSM_CUDA(cudaMemcpyAsync(weight_out, g->weight, bytes, cudaMemcpyDeviceToHost, ctx->stream));
  CSM_CUDA(cudaStreamSynchronize(ctx->stream));
  return CSM_OK;
}

const uint16_t* csm_tsdfgrid2d_tsd_dev(const csm_tsdfgrid2d* g) { return g ? g->tsd : nullptr; }
const uint16_t* csm_tsdfgrid2d_weight_dev(const csm_tsdfgrid2d* g) { return g ? g->weight : nullptr; }

int csm_tsdfgrid2d_insert(csm_ctx* ctx, csm_tsdfgrid2d* g, const csm_tsdf_range_data2d* scans, int num_scans) {
  CSM_REQUIRE(ctx != nullptr && g != nullptr && scans != nullptr, "NULL argument");
  CSM_REQUIRE(num_scans >= 1, "num_scans must be positive");
  CSM_REQUIRE(g->device == ctx->device, "grid lives on another device");
  CSM_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t st = ctx->stream;
  size_t total = 0;
  for (int s = 0; s < num_scans; ++s) {
    const csm_tsdf_range_data2d& in = scans[s];
    CSM_REQUIRE(in.num_returns >= 0 && in.resolution > 0. && in.truncation_distance > 0.f && in.maximum_weight > 0.f,
                "TSDF range data sizes");
    CSM_REQUIRE(in.num_returns == 0 || (in.hits_xy != nullptr && in.hit_weight != nullptr), "per-return arrays");
    CSM_REQUIRE(!in.project_sdf_distance_to_scan_normal || in.num_returns == 0 || in.normal_cos_sin != nullptr,
                "projection onto the scan normal needs the normals");
    total += in.num_returns;
  }
  if (total == 0) return CSM_OK;
  // [hits xy | normal cos sin | weight] per scan, one pinned staging copy for the whole call
  const size_t floats = 5 * total;
  if (g->staging_cap < floats) {
    CSM_CUDA(cudaStreamSynchronize(st));
    cudaFree(g->staging);
    g->staging = nullptr;
    g->staging_cap = 0;
    CSM_CUDA(cudaMalloc(&g->staging, sizeof(float) * (floats + 4096)));
    g->staging_cap = floats + 4096;
  }
  int rc = ctx->pin0.reserve(sizeof(float) * floats);
  if (rc != CSM_OK) return rc;
  float* pin = ctx->pin0.as<float>();
  size_t at = 0;
  for (int s = 0; s < num_scans; ++s) {
    const csm_tsdf_range_data2d& in = scans[s];
    const size_t n = in.num_returns;
    if (n == 0) continue;
    std::memcpy(pin + at, in.hits_xy, sizeof(float) * 2 * n);
    if (in.normal_cos_sin != nullptr) std::memcpy(pin + at + 2 * n, in.normal_cos_sin, sizeof(float) * 2 * n);
    else std::memset(pin + at + 2 * n, 0, sizeof(float) * 2 * n);
    std::memcpy(pin + at + 4 * n, in.hit_weight, sizeof(float) * n);
    at += 5 * n;
  }
  CSM_CUDA(cudaMemcpyAsync(g->staging, pin, sizeof(float) * floats, cudaMemcpyHostToDevice, st));
  at = 0;
  for (int s = 0; s < num_scans; ++s) {
    const csm_tsdf_range_data2d& in = scans[s];
    const size_t n = in.num_returns;
    if (n == 0) continue;
    TsdfArgs a;
    a.max_x = in.max_x;
    a.max_y = in.max_y;
    a.resolution = in.resolution;
    a.super_res = in.resolution / kSubpixelScale;
    a.use_distance_kernel = in.distance_kernel_bandwidth != 0.f ? 1 : 0;
    const float sqrt_two_pi = static_cast<float>(std::sqrt(2.0 * M_PI));  // const float kSqrtTwoPi, :31
    a.kernel_scale = a.use_distance_kernel ? 1.0 / (sqrt_two_pi * in.distance_kernel_bandwidth) : 0.;
    a.kernel_sigma2 = in.distance_kernel_bandwidth * in.distance_kernel_bandwidth;  // fp32 product, widened
    a.origin_x = in.origin_x;
    a.origin_y = in.origin_y;
    a.truncation_distance = in.truncation_distance;
    a.maximum_weight = in.maximum_weight;
    a.min_tsd = in.min_tsd;
    a.max_tsd = in.max_tsd;
    a.tsd_resolution = in.tsd_resolution;
    a.max_weight_value = in.max_weight_value;
    a.weight_resolution = in.weight_resolution;
    a.num_x = g->num_x;
    a.num_y = g->num_y;
    a.num_returns = in.num_returns;
    a.update_free_space = in.update_free_space;
    a.project = in.project_sdf_distance_to_scan_normal;
    const int blocks = (in.num_returns + 127) / 128;
    for (int pass = 0; pass < 2; ++pass) {
      ++ctx->launches, k_tsdf_insert<<<blocks, 128, 0, st>>>(a, g->staging + at, g->staging + at + 2 * n,
                                                            g->staging + at + 4 * n, g->tsd, g->weight, g->owner,
                                                            g->tables, pass);
    }
    at += 5 * n;
  }
  CSM_CUDA(cudaStreamSynchronize(st));
  CSM_CUDA(cudaGetLastError());
  return CSM_OK;
}

}  // extern "C"
