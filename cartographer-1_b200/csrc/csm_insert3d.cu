// RangeDataInserter3D on a live HybridGrid that stays in HBM (the 3D half).
//
// Reference: cartographer/mapping/3d/range_data_inserter_3d.cc:44-131 and
// cartographer/mapping/3d/hybrid_grid.h:423-445 (GetCellIndex), :492-518 (FinishUpdate,
// ApplyLookupTable).  What the reference does per range data:
//   1. every return:  ApplyLookupTable(GetCellIndex(hit), hit_table)           (:119-122)
//   2. every return:  the last num_free_space_voxels samples of the line from the origin's cell
//      to the hit's cell, miss_cell = origin_cell + delta * position / num_samples in integer
//      arithmetic, ApplyLookupTable(miss_cell, miss_table)                       (:44-76)
//   3. FinishUpdate() drops the update markers                                   (:492-498)
// ApplyLookupTable touches a cell at most once between two FinishUpdate calls (the marker), so a
// cell that holds any hit takes the hit table once, a cell crossed only by free-space samples takes
// the miss table once, whatever the order and multiplicity.  That is a per-cell flag: pass 1 marks
// (one thread per return: the hit bit on its cell, the miss bit on its free-space samples), pass 2
// walks the same cells again and the thread that clears a cell's flags applies its table -- hit
// wins -- and removes the marker.  Everything is integer arithmetic except GetCellIndex, which is
// one IEEE division and a round-half-away per component, identical on the device.
//
// The grid is a dense uint16 box [z][y][x] with its own origin; it grows (on the stream, by a copy
// kernel) when a range data reaches outside.  The sparse reference grid has no box; what the path
// observes is the set of non-default voxels, which csm_hgrid3d_export returns.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <vector>

#include "csm_internal.cuh"

using namespace csm;

struct csm_hgrid3d {
  int device = 0;
  float resolution = 0.f;
  int lo[3] = {0, 0, 0};
  int dim[3] = {0, 0, 0};
  uint16_t* cells = nullptr;      // dim[0] * dim[1] * dim[2]
  unsigned int* flags = nullptr;  // one byte per cell, addressed as words
  size_t cells_cap = 0, flags_cap = 0;
  uint16_t* tables = nullptr;  // hit table, miss table (32768 entries each)
  float* points = nullptr;
  size_t points_cap = 0;
  unsigned int* counter = nullptr;
};

namespace {

constexpr uint16_t kUpdateMarker = 1u << 15;

struct Box3 {
  int lo[3];
  int dim[3];
};

__host__ __device__ inline int round_half_away(float x) {
#ifdef __CUDA_ARCH__
  return static_cast<int>(roundf(x));  // std::lround: halfway cases away from zero
#else
  return static_cast<int>(std::lround(x));
#endif
}

__host__ __device__ inline void cell_index(const float* p, float resolution, int* c) {
#ifdef __CUDA_ARCH__
  for (int a = 0; a < 3; ++a) c[a] = round_half_away(__fdiv_rn(p[a], resolution));
#else
  for (int a = 0; a < 3; ++a) c[a] = round_half_away(p[a] / resolution);
#endif
}

__device__ inline size_t flat(const Box3& b, int x, int y, int z) {
  return (static_cast<size_t>(z - b.lo[2]) * b.dim[1] + (y - b.lo[1])) * b.dim[0] + (x - b.lo[0]);
}

__device__ inline void mark(unsigned int* flags, size_t i, unsigned int bit) {
  atomicOr(flags + (i >> 2), bit << (8 * (i & 3)));
}

// Clears the flag byte of cell i; the caller that sees a non-zero old value owns the update.
__device__ inline unsigned int claim(unsigned int* flags, size_t i) {
  const unsigned int shift = 8 * (i & 3);
  const unsigned int old = atomicAnd(flags + (i >> 2), ~(0xFFu << shift));
  return (old >> shift) & 0xFFu;
}

__device__ inline void apply(uint16_t* cells, const uint16_t* tables, size_t i, unsigned int f) {
  if (f == 0u) return;
  const uint16_t old = cells[i];
  if (old >= kUpdateMarker) return;  // cannot happen between insertions
  const uint16_t* t = tables + ((f & 1u) ? 0 : 32768);
  cells[i] = static_cast<uint16_t>(t[old] - kUpdateMarker);
}

// One thread per return.  pass 0: mark; pass 1: claim and apply.
__global__ void k_insert3d(const float* __restrict__ returns, int num_returns, float ox, float oy, float oz,
                           float resolution, int num_free_space_voxels, Box3 box, uint16_t* cells,
                           unsigned int* flags, const uint16_t* __restrict__ tables, int pass) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= num_returns) return;
  const float origin[3] = {ox, oy, oz};
  const float hit[3] = {returns[3 * i], returns[3 * i + 1], returns[3 * i + 2]};
  int oc[3], hc[3];
  cell_index(origin, resolution, oc);
  cell_index(hit, resolution, hc);
  const size_t at = flat(box, hc[0], hc[1], hc[2]);
  if (pass == 0) {
    mark(flags, at, 1u);
  } else {
    apply(cells, tables, at, claim(flags, at));
  }
  const int d[3] = {hc[0] - oc[0], hc[1] - oc[1], hc[2] - oc[2]};
  const int num_samples = max(abs(d[0]), max(abs(d[1]), abs(d[2])));
  for (int position = max(0, num_samples - num_free_space_voxels); position < num_samples; ++position) {
    // origin_cell + delta * position / num_samples, component-wise, truncating division
    const size_t m = flat(box, oc[0] + d[0] * position / num_samples, oc[1] + d[1] * position / num_samples,
                          oc[2] + d[2] * position / num_samples);
    if (pass == 0) {
      mark(flags, m, 2u);
    } else {
      apply(cells, tables, m, claim(flags, m));
    }
  }
}

__global__ void k_copy_box(const uint16_t* __restrict__ in, Box3 from, uint16_t* __restrict__ out, Box3 to) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  const int y = blockIdx.y, z = blockIdx.z;
  if (x >= from.dim[0]) return;
  const uint16_t v = in[(static_cast<size_t>(z) * from.dim[1] + y) * from.dim[0] + x];
  if (v != 0) out[flat(to, from.lo[0] + x, from.lo[1] + y, from.lo[2] + z)] = v;
}

__global__ void k_scatter_voxels(const int* __restrict__ xyz, const uint16_t* __restrict__ values, int n, Box3 box,
                                 uint16_t* cells) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  cells[flat(box, xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2])] = values[i];
}

// Non-default voxels, in [z][y][x] order within a block and arbitrary order between blocks
// (the host sorts when it needs an order).  out == nullptr: count only.
__global__ void k_export(const uint16_t* __restrict__ cells, Box3 box, size_t n, unsigned int* counter, int* xyz,
                         uint16_t* values, unsigned int cap) {
  const size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  const uint16_t v = i < n ? cells[i] : 0;
  const unsigned int ballot = __ballot_sync(0xFFFFFFFFu, v != 0);
  if (ballot == 0u) return;
  const int lane = threadIdx.x & 31;
  unsigned int base = 0;
  if (lane == 0) base = atomicAdd(counter, __popc(ballot));
  base = __shfl_sync(0xFFFFFFFFu, base, 0);
  if (v == 0 || xyz == nullptr) return;
  const unsigned int at = base + __popc(ballot & ((1u << lane) - 1u));
  if (at >= cap) return;
  const size_t plane = static_cast<size_t>(box.dim[0]) * box.dim[1];
  const int z = static_cast<int>(i / plane);
  const int y = static_cast<int>((i - z * plane) / box.dim[0]);
  const int x = static_cast<int>(i - z * plane - static_cast<size_t>(y) * box.dim[0]);
  xyz[3 * at] = box.lo[0] + x;
  xyz[3 * at + 1] = box.lo[1] + y;
  xyz[3 * at + 2] = box.lo[2] + z;
  values[at] = v;
}

// lo/hi of the n exported voxels: out6 = {min x, y, z, max x, y, z}, pre-set to +-INT_MAX.
__global__ void k_voxel_box(const int* __restrict__ xyz, unsigned int n, int* out6) {
  const unsigned int i = blockIdx.x * blockDim.x + threadIdx.x;
  int v[3] = {0, 0, 0};
  const bool have = i < n;
  if (have) {
    v[0] = xyz[3 * i];
    v[1] = xyz[3 * i + 1];
    v[2] = xyz[3 * i + 2];
  }
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    const int mn = __reduce_min_sync(0xFFFFFFFFu, have ? v[a] : 0x7FFFFFFF);
    const int mx = __reduce_max_sync(0xFFFFFFFFu, have ? v[a] : -0x7FFFFFFF);
    if ((threadIdx.x & 31) == 0) {
      atomicMin(out6 + a, mn);
      atomicMax(out6 + 3 + a, mx);
    }
  }
}

Box3 box_of(const csm_hgrid3d* g) {
  Box3 b;
  for (int a = 0; a < 3; ++a) {
    b.lo[a] = g->lo[a];
    b.dim[a] = g->dim[a];
  }
  return b;
}

size_t volume(const int* dim) { return static_cast<size_t>(dim[0]) * dim[1] * dim[2]; }

// Makes the box contain [lo, hi] (inclusive); keeps the contents.  Grows with slack so that a
// trajectory that creeps outward does not re-allocate every scan.
int ensure_box(csm_ctx* ctx, csm_hgrid3d* g, const int* lo, const int* hi) {
  bool fits = g->cells != nullptr;
  for (int a = 0; a < 3 && fits; ++a) fits = lo[a] >= g->lo[a] && hi[a] < g->lo[a] + g->dim[a];
  if (fits) return CSM_OK;
  int nlo[3], ndim[3];
  for (int a = 0; a < 3; ++a) {
    int l = lo[a], h = hi[a];
    if (g->cells != nullptr) {
      l = std::min(l, g->lo[a]);
      h = std::max(h, g->lo[a] + g->dim[a] - 1);
    }
    const int slack = std::max(8, (h - l + 1) / 8);
    // The hard limit of the reference's grid is +-8192 cells around the origin (hybrid_grid.h:467).
    nlo[a] = std::max(-8192, l - slack);
    ndim[a] = std::min(8191, h + slack) - nlo[a] + 1;
    CSM_REQUIRE(l >= -8192 && h <= 8191, "cell index outside the +-8192 range of a HybridGrid");
  }
  const size_t n = volume(ndim);
  CSM_REQUIRE(n < (size_t(1) << 33), "a dense box of this HybridGrid would exceed 16 GB");
  cudaStream_t st = ctx->stream;
  uint16_t* cells = nullptr;
  unsigned int* flags = nullptr;
  const size_t flag_words = (n + 3) / 4;
  if (cudaMalloc(&cells, sizeof(uint16_t) * n) != cudaSuccess ||
      cudaMalloc(&flags, sizeof(unsigned int) * flag_words) != cudaSuccess) {
    set_error("allocation of a %d x %d x %d HybridGrid box failed: %s", ndim[0], ndim[1], ndim[2],
              cudaGetErrorString(cudaGetLastError()));
    cudaFree(cells);
    return CSM_ERR_CUDA;
  }
  CSM_CUDA(cudaMemsetAsync(cells, 0, sizeof(uint16_t) * n, st));
  CSM_CUDA(cudaMemsetAsync(flags, 0, sizeof(unsigned int) * flag_words, st));
  Box3 to;
  for (int a = 0; a < 3; ++a) {
    to.lo[a] = nlo[a];
    to.dim[a] = ndim[a];
  }
  if (g->cells != nullptr) {
    const Box3 from = box_of(g);
    ++ctx->launches, k_copy_box<<<dim3((from.dim[0] + 127) / 128, from.dim[1], from.dim[2]), 128, 0, st>>>(
        g->cells, from, cells, to);
    CSM_CUDA(cudaStreamSynchronize(st));
    cudaFree(g->cells);
    cudaFree(g->flags);
  }
  g->cells = cells;
  g->flags = flags;
  for (int a = 0; a < 3; ++a) {
    g->lo[a] = nlo[a];
    g->dim[a] = ndim[a];
  }
  return CSM_OK;
}

}  // namespace

namespace csm {
int hgrid3d_stage(::csm_ctx* ctx, const ::csm_hgrid3d* g, int* num_voxels, int* lo3, int* hi3) {
  CSM_REQUIRE(g->device == ctx->device, "grid lives on another device");
  CSM_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t st = ctx->stream;
  *num_voxels = 0;
  const size_t n = g->cells ? volume(g->dim) : 0;
  if (n == 0) return CSM_OK;
  // pass 1: count; pass 2: write (the order between blocks is arbitrary, a set is all the build needs)
  unsigned int count = 0;
  CSM_CUDA(cudaMemsetAsync(g->counter, 0, sizeof(unsigned int), st));
  ++ctx->launches, k_export<<<static_cast<unsigned int>((n + 255) / 256), 256, 0, st>>>(g->cells, box_of(g), n, g->counter,
                                                                                       nullptr, nullptr, 0u);
  CSM_CUDA(cudaMemcpyAsync(&count, g->counter, sizeof(count), cudaMemcpyDeviceToHost, st));
  CSM_CUDA(cudaStreamSynchronize(st));
  if (count == 0) return CSM_OK;
  int rc = ctx->tmp0.reserve(sizeof(int) * 3 * static_cast<size_t>(count) + 64);
  if (rc == CSM_OK) rc = ctx->tmp1.reserve(sizeof(uint16_t) * static_cast<size_t>(count) + 64);
  if (rc == CSM_OK) rc = ctx->tmp2.reserve(64);
  if (rc != CSM_OK) return rc;
  CSM_CUDA(cudaMemsetAsync(g->counter, 0, sizeof(unsigned int), st));
  ++ctx->launches, k_export<<<static_cast<unsigned int>((n + 255) / 256), 256, 0, st>>>(
      g->cells, box_of(g), n, g->counter, ctx->tmp0.as<int>(), ctx->tmp1.as<uint16_t>(), count);
  const int init[6] = {0x7FFFFFFF, 0x7FFFFFFF, 0x7FFFFFFF, -0x7FFFFFFF, -0x7FFFFFFF, -0x7FFFFFFF};
  int box[6];
  CSM_CUDA(cudaMemcpyAsync(ctx->tmp2.p, init, sizeof(init), cudaMemcpyHostToDevice, st));
  ++ctx->launches, k_voxel_box<<<(count + 255) / 256, 256, 0, st>>>(ctx->tmp0.as<int>(), count, ctx->tmp2.as<int>());
  CSM_CUDA(cudaMemcpyAsync(box, ctx->tmp2.p, sizeof(box), cudaMemcpyDeviceToHost, st));
  CSM_CUDA(cudaStreamSynchronize(st));
  CSM_CUDA(cudaGetLastError());
  for (int a = 0; a < 3; ++a) {
    lo3[a] = box[a];
    hi3[a] = box[3 + a];
  }
  *num_voxels = static_cast<int>(count);
  return CSM_OK;
}
}  // namespace csm

extern "C" {

int csm_hgrid3d_create(csm_ctx* ctx, float resolution, const int32_t* voxel_xyz, const uint16_t* voxel_values,
                       int num_voxels, csm_hgrid3d** out) {
  CSM_REQUIRE(ctx != nullptr && out != nullptr, "ctx/out must not be NULL");
  CSM_REQUIRE(resolution > 0.f, "resolution must be positive");
  CSM_REQUIRE(num_voxels >= 0 && (num_voxels == 0 || (voxel_xyz && voxel_values)), "voxel arrays");
  CSM_CUDA(cudaSetDevice(ctx->device));
  csm_hgrid3d* g = new csm_hgrid3d;
  g->device = ctx->device;
  g->resolution = resolution;
  auto fail = [&](int rc) {
    cudaFree(g->cells);
    cudaFree(g->flags);
    cudaFree(g->tables);
    cudaFree(g->counter);
    delete g;
    return rc;
  };
  if (cudaMalloc(&g->tables, sizeof(uint16_t) * 65536) != cudaSuccess ||
      cudaMalloc(&g->counter, sizeof(unsigned int)) != cudaSuccess) {
    set_error("allocation failed: %s", cudaGetErrorString(cudaGetLastError()));
    return fail(CSM_ERR_CUDA);
  }
  if (num_voxels > 0) {
    int lo[3], hi[3];
    for (int a = 0; a < 3; ++a) lo[a] = hi[a] = voxel_xyz[a];
    for (int i = 1; i < num_voxels; ++i)
      for (int a = 0; a < 3; ++a) {
        lo[a] = std::min(lo[a], voxel_xyz[3 * i + a]);
        hi[a] = std::max(hi[a], voxel_xyz[3 * i + a]);
      }
    int rc = ensure_box(ctx, g, lo, hi);
    if (rc != CSM_OK) return fail(rc);
    rc = ctx->tmp0.reserve(sizeof(int) * 3 * static_cast<size_t>(num_voxels));
    if (rc == CSM_OK) rc = ctx->tmp1.reserve(sizeof(uint16_t) * static_cast<size_t>(num_voxels));
    if (rc != CSM_OK) return fail(rc);
    cudaStream_t st = ctx->stream;
    if (cudaMemcpyAsync(ctx->tmp0.p, voxel_xyz, sizeof(int) * 3 * static_cast<size_t>(num_voxels),
                        cudaMemcpyHostToDevice, st) != cudaSuccess ||
        cudaMemcpyAsync(ctx->tmp1.p, voxel_values, sizeof(uint16_t) * static_cast<size_t>(num_voxels),
                        cudaMemcpyHostToDevice, st) != cudaSuccess) {
      set_error("voxel upload failed: %s", cudaGetErrorString(cudaGetLastError()));
      return fail(CSM_ERR_CUDA);
    }
    ++ctx->launches, k_scatter_voxels<<<(num_voxels + 255) / 256, 256, 0, st>>>(
        ctx->tmp0.as<int>(), ctx->tmp1.as<uint16_t>(), num_voxels, box_of(g), g->cells);
    if (cudaStreamSynchronize(st) != cudaSuccess) {
      set_error("voxel scatter failed: %s", cudaGetErrorString(cudaGetLastError()));
      return fail(CSM_ERR_CUDA);
    }
  }
  *out = g;
  return CSM_OK;
}

void csm_hgrid3d_destroy(csm_ctx* ctx, csm_hgrid3d* g) {
  if (!g) return;
  cudaSetDevice(g->device);
  if (ctx_alive_on(ctx, g->device)) cudaStreamSynchronize(ctx->stream);
  else cudaDeviceSynchronize();
  cudaFree(g->cells);
  cudaFree(g->flags);
  cudaFree(g->tables);
  cudaFree(g->points);
  cudaFree(g->counter);
  delete g;
}

int csm_hgrid3d_insert(csm_ctx* ctx, csm_hgrid3d* g, const csm_range_data3d* scans, int num_scans) {
  CSM_REQUIRE(ctx != nullptr && g != nullptr && scans != nullptr, "NULL argument");
  CSM_REQUIRE(num_scans >= 1, "num_scans must be positive");
  CSM_REQUIRE(g->device == ctx->device, "grid lives on another device");
  CSM_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t st = ctx->stream;
  // Bounding box of everything the scans touch: hit cells and the origin's cell (free-space
  // samples lie between them component-wise).  GetCellIndex on the host is the same IEEE
  // division and rounding as on the device.
  size_t total = 0;
  int lo[3] = {0, 0, 0}, hi[3] = {0, 0, 0};
  bool any = false;
  for (int s = 0; s < num_scans; ++s) {
    const csm_range_data3d& in = scans[s];
    CSM_REQUIRE(in.hit_table != nullptr && in.miss_table != nullptr, "lookup tables must not be NULL");
    CSM_REQUIRE(in.num_returns >= 0 && (in.num_returns == 0 || in.returns_xyz != nullptr), "returns array");
    CSM_REQUIRE(in.num_free_space_voxels >= 0, "num_free_space_voxels must not be negative");
    if (in.num_returns == 0) continue;
    int oc[3];
    cell_index(in.origin, g->resolution, oc);
    for (int i = -1; i < in.num_returns; ++i) {
      int c[3];
      if (i < 0) {
        std::memcpy(c, oc, sizeof(c));
      } else {
        cell_index(in.returns_xyz + 3 * i, g->resolution, c);
        int num_samples = 0;
        for (int a = 0; a < 3; ++a) num_samples = std::max(num_samples, std::abs(c[a] - oc[a]));
        CSM_REQUIRE(num_samples < (1 << 15), "a ray spans 2^15 or more voxels (reference CHECK)");
      }
      for (int a = 0; a < 3; ++a) {
        lo[a] = any ? std::min(lo[a], c[a]) : c[a];
        hi[a] = any ? std::max(hi[a], c[a]) : c[a];
      }
      any = true;
    }
    total += in.num_returns;
  }
  if (!any) return CSM_OK;
  int rc = ensure_box(ctx, g, lo, hi);
  if (rc != CSM_OK) return rc;
  if (g->points_cap < total) {
    CSM_CUDA(cudaStreamSynchronize(st));
    cudaFree(g->points);
    g->points = nullptr;
    g->points_cap = 0;
    CSM_CUDA(cudaMalloc(&g->points, sizeof(float) * 3 * (total + 4096)));
    g->points_cap = total + 4096;
  }
  rc = ctx->pin0.reserve(sizeof(float) * 3 * total);
  if (rc != CSM_OK) return rc;
  float* pin = ctx->pin0.as<float>();
  size_t at = 0;
  for (int s = 0; s < num_scans; ++s) {
    if (scans[s].num_returns) std::memcpy(pin + 3 * at, scans[s].returns_xyz, sizeof(float) * 3 * scans[s].num_returns);
    at += scans[s].num_returns;
  }
  CSM_CUDA(cudaMemcpyAsync(g->points, pin, sizeof(float) * 3 * total, cudaMemcpyHostToDevice, st));
  const uint16_t* loaded_hit = nullptr;
  const uint16_t* loaded_miss = nullptr;
  const Box3 box = box_of(g);
  at = 0;
  for (int s = 0; s < num_scans; ++s) {
    const csm_range_data3d& in = scans[s];
    if (in.num_returns == 0) continue;
    if (in.hit_table != loaded_hit || in.miss_table != loaded_miss) {  // pageable: synchronous copies
      CSM_CUDA(cudaMemcpyAsync(g->tables, in.hit_table, sizeof(uint16_t) * 32768, cudaMemcpyHostToDevice, st));
      CSM_CUDA(cudaMemcpyAsync(g->tables + 32768, in.miss_table, sizeof(uint16_t) * 32768, cudaMemcpyHostToDevice, st));
      loaded_hit = in.hit_table;
      loaded_miss = in.miss_table;
    }
    const int blocks = (in.num_returns + 127) / 128;
    for (int pass = 0; pass < 2; ++pass) {
      ++ctx->launches, k_insert3d<<<blocks, 128, 0, st>>>(g->points + 3 * at, in.num_returns, in.origin[0], in.origin[1],
                                                       in.origin[2], g->resolution, in.num_free_space_voxels, box,
                                                       g->cells, g->flags, g->tables, pass);
    }
    at += in.num_returns;
  }
  CSM_CUDA(cudaStreamSynchronize(st));  // the staging buffer and the callers' tables are free again
  CSM_CUDA(cudaGetLastError());
  return CSM_OK;
}

int csm_hgrid3d_box(const csm_hgrid3d* g, int32_t* lo3, int32_t* dims3) {
  CSM_REQUIRE(g != nullptr, "NULL grid");
  for (int a = 0; a < 3; ++a) {
    if (lo3) lo3[a] = g->lo[a];
    if (dims3) dims3[a] = g->dim[a];
  }
  return CSM_OK;
}

int csm_hgrid3d_export(csm_ctx* ctx, const csm_hgrid3d* g, int32_t* voxel_xyz, uint16_t* voxel_values,
                       int capacity, int* num_voxels) {
  CSM_REQUIRE(ctx != nullptr && g != nullptr && num_voxels != nullptr, "NULL argument");
  CSM_REQUIRE(capacity >= 0 && (capacity == 0 || (voxel_xyz && voxel_values)), "output arrays");
  CSM_CUDA(cudaSetDevice(g->device));
  cudaStream_t st = ctx->stream;
  *num_voxels = 0;
  const size_t n = g->cells ? volume(g->dim) : 0;
  if (n == 0) return CSM_OK;
  int rc = CSM_OK;
  if (capacity > 0) {
    rc = ctx->tmp0.reserve(sizeof(int) * 3 * static_cast<size_t>(capacity));
    if (rc == CSM_OK) rc = ctx->tmp1.reserve(sizeof(uint16_t) * static_cast<size_t>(capacity));
    if (rc != CSM_OK) return rc;
  }
  CSM_CUDA(cudaMemsetAsync(g->counter, 0, sizeof(unsigned int), st));
  ++ctx->launches, k_export<<<static_cast<unsigned int>((n + 255) / 256), 256, 0, st>>>(
      g->cells, box_of(g), n, g->counter, capacity > 0 ? ctx->tmp0.as<int>() : nullptr,
      capacity > 0 ? ctx->tmp1.as<uint16_t>() : nullptr, static_cast<unsigned int>(capacity));
  unsigned int count = 0;
  CSM_CUDA(cudaMemcpyAsync(&count, g->counter, sizeof(count), cudaMemcpyDeviceToHost, st));
  CSM_CUDA(cudaStreamSynchronize(st));
  *num_voxels = static_cast<int>(count);
  if (capacity > 0) {
    const size_t k = std::min<size_t>(count, capacity);
    CSM_CUDA(cudaMemcpyAsync(voxel_xyz, ctx->tmp0.p, sizeof(int) * 3 * k, cudaMemcpyDeviceToHost, st));
    CSM_CUDA(cudaMemcpyAsync(voxel_values, ctx->tmp1.p, sizeof(uint16_t) * k, cudaMemcpyDeviceToHost, st));
    CSM_CUDA(cudaStreamSynchronize(st));
  }
  return CSM_OK;
}

}  // extern "C"
