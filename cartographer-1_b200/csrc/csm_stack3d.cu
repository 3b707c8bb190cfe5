// PrecomputationGridStack3D on the device.
//
// Reference: cartographer/mapping/internal/3d/scan_matching/
//   fast_correlative_scan_matcher_3d.cc:57-77   (stack constructor: widths, shifts)
//   precomputation_grid_3d.cc:49-61             (ConvertToPrecomputationGrid)
//   precomputation_grid_3d.cc:63-81             (PrecomputeGrid: 8-octant max scatter,
//                                                optional index >> 1)
//
// The reference scatters every stored voxel of level k-1 into (up to) eight voxels of
// a growing sparse tree.  Here each level is a dense byte box and the same values are
// produced in gather form: a voxel c of a full-resolution level is
//     max over octants o of in(c + shift * o),
// and a voxel c of a half-resolution level is the maximum over its pre-image under
// (idx - shift * o) >> 1, i.e. over c' in {2c, 2c+1}^3 and o of in(c' + shift * o).
// The boxes are computed on the host so that they contain every voxel the scatter
// would create; everything outside is the sparse grid's default 0.
#include <algorithm>
#include <cstring>

#include "csm_match3d.cuh"

namespace csm {
namespace {

__global__ void __launch_bounds__(256)
k3_level0(const int* __restrict__ xyz, const uint16_t* __restrict__ values, int n,
          const uint8_t* __restrict__ q_lut, uint8_t* __restrict__ out, int lox, int loy, int loz,
          int nx, int ny) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint16_t v = values[i];
  if (v == 0) return;  // the reference's iterator skips default cells (hybrid_grid.h:386-399)
  const int x = xyz[3 * i] - lox, y = xyz[3 * i + 1] - loy, z = xyz[3 * i + 2] - loz;
  out[(static_cast<size_t>(z) * ny + y) * nx + x] = q_lut[v];
}

struct UpArgs {
  const uint8_t* in;
  uint8_t* out;
  int in_lo[3], in_dim[3];
  int out_lo[3], out_dim[3];
  int shift;
  int half;  // 1: half-resolution level
};

__device__ __forceinline__ unsigned int in_at(const UpArgs& a, int gx, int gy, int gz) {
  const int x = gx - a.in_lo[0], y = gy - a.in_lo[1], z = gz - a.in_lo[2];
  if (static_cast<unsigned int>(x) >= static_cast<unsigned int>(a.in_dim[0]) ||
      static_cast<unsigned int>(y) >= static_cast<unsigned int>(a.in_dim[1]) ||
      static_cast<unsigned int>(z) >= static_cast<unsigned int>(a.in_dim[2]))
    return 0u;
  return a.in[(static_cast<size_t>(z) * a.in_dim[1] + y) * a.in_dim[0] + x];
}

// One thread per output voxel, x fastest so that a warp reads runs of input bytes.
__global__ void __launch_bounds__(256)
k3_level_up(const UpArgs a) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  const int y = blockIdx.y, z = blockIdx.z;
  if (x >= a.out_dim[0]) return;
  const int cx = a.out_lo[0] + x, cy = a.out_lo[1] + y, cz = a.out_lo[2] + z;
  unsigned int m = 0;
  if (!a.half) {
    for (int o = 0; o < 8; ++o)
      m = max(m, in_at(a, cx + ((o & 1) ? a.shift : 0), cy + ((o & 2) ? a.shift : 0),
                       cz + ((o & 4) ? a.shift : 0)));
  } else {
    // per axis the source positions are {2c, 2c+1, 2c+shift, 2c+1+shift}
    int px[4], py[4], pz[4];
    for (int k = 0; k < 4; ++k) {
      const int d = (k & 1) + ((k & 2) ? a.shift : 0);
      px[k] = 2 * cx + d;
      py[k] = 2 * cy + d;
      pz[k] = 2 * cz + d;
    }
    for (int kz = 0; kz < 4; ++kz)
      for (int ky = 0; ky < 4; ++ky)
        for (int kx = 0; kx < 4; ++kx) m = max(m, in_at(a, px[kx], py[ky], pz[kz]));
  }
  a.out[(static_cast<size_t>(z) * a.out_dim[1] + y) * a.out_dim[0] + x] = static_cast<uint8_t>(m);
}

// The sweep's layout: element (xp, yp, zp) gathers the eight child values of the parent
// whose first child sits at (xp, yp, zp) - stride.  One thread per element, x fastest.
struct PackArgs {
  const uint8_t* in;
  uint2* out;
  int dim[3], pdim[3];
  int stride;
};
__global__ void __launch_bounds__(256)
k3_pack(const PackArgs a) {
  const int xp = blockIdx.x * blockDim.x + threadIdx.x;
  const int yp = blockIdx.y, zp = blockIdx.z;
  if (xp >= a.pdim[0]) return;
  unsigned int w[2] = {0u, 0u};
  for (int k = 0; k < 8; ++k) {
    const int x = xp - a.stride + ((k & 1) ? a.stride : 0);
    const int y = yp - a.stride + ((k & 2) ? a.stride : 0);
    const int z = zp - a.stride + ((k & 4) ? a.stride : 0);
    unsigned int v = 0;
    if (static_cast<unsigned int>(x) < static_cast<unsigned int>(a.dim[0]) &&
        static_cast<unsigned int>(y) < static_cast<unsigned int>(a.dim[1]) &&
        static_cast<unsigned int>(z) < static_cast<unsigned int>(a.dim[2]))
      v = a.in[(static_cast<size_t>(z) * a.dim[1] + y) * a.dim[0] + x];
    w[k >> 2] |= v << (8 * (k & 3));
  }
  a.out[(static_cast<size_t>(zp) * a.pdim[1] + yp) * a.pdim[0] + xp] = make_uint2(w[0], w[1]);
}

__global__ void __launch_bounds__(256)
k3_grid_fill(float* __restrict__ out, size_t count, float v) {
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < count;
       i += static_cast<size_t>(gridDim.x) * blockDim.x)
    out[i] = v;
}

__global__ void __launch_bounds__(256)
k3_grid_scatter(const int* __restrict__ xyz, const uint16_t* __restrict__ values, int n,
                const float* __restrict__ table, float* __restrict__ out, int lox, int loy, int loz,
                int nx, int ny) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int x = xyz[3 * i] - lox, y = xyz[3 * i + 1] - loy, z = xyz[3 * i + 2] - loz;
  out[(static_cast<size_t>(z) * ny + y) * nx + x] = table[values[i]];
}

// Bounding box of the non-default voxels; false if there are none.
bool voxel_box(const int32_t* xyz, const uint16_t* values, int n, bool skip_zero, int lo[3], int hi[3]) {
  bool any = false;
  for (int i = 0; i < n; ++i) {
    if (skip_zero && values[i] == 0) continue;
    for (int k = 0; k < 3; ++k) {
      const int v = xyz[3 * i + k];
      lo[k] = any ? std::min(lo[k], v) : v;
      hi[k] = any ? std::max(hi[k], v) : v;
    }
    any = true;
  }
  return any;
}

size_t align256(size_t v) { return (v + 255) & ~static_cast<size_t>(255); }

// Stages the voxel list in ctx->tmp0 (xyz) / ctx->tmp1 (values) through pinned memory.
int upload_voxels(csm_ctx* ctx, const int32_t* xyz, const uint16_t* values, int n) {
  const size_t bx = sizeof(int32_t) * 3 * static_cast<size_t>(n), bv = sizeof(uint16_t) * static_cast<size_t>(n);
  int rc = ctx->tmp0.reserve(std::max<size_t>(bx, 16));
  if (rc == CSM_OK) rc = ctx->tmp1.reserve(std::max<size_t>(bv, 16));
  if (rc == CSM_OK) rc = ctx->pin0.reserve(std::max<size_t>(align256(bx) + bv, 16));
  if (rc != CSM_OK) return rc;
  if (n == 0) return CSM_OK;
  char* pin = ctx->pin0.as<char>();
  memcpy(pin, xyz, bx);
  memcpy(pin + align256(bx), values, bv);
  CSM_CUDA(cudaMemcpyAsync(ctx->tmp0.p, pin, bx, cudaMemcpyHostToDevice, ctx->stream));
  CSM_CUDA(cudaMemcpyAsync(ctx->tmp1.p, pin + align256(bx), bv, cudaMemcpyHostToDevice, ctx->stream));
  return CSM_OK;
}

}  // namespace
}  // namespace csm

using namespace csm;

extern "C" {

// staged_box != nullptr: the voxel list already sits in ctx->tmp0 / tmp1 on the device (put there by
// hgrid3d_stage) and staged_box = {lo[3], hi[3]} is its bounding box; else it is taken from the host
// arrays.
static int build_stack3d(csm_ctx* ctx, const int32_t* voxel_xyz, const uint16_t* voxel_values, int num_voxels,
                         const int* staged_box, const uint8_t* q_lut, int depth, int full_resolution_depth,
                         csm_stack3d** out);

int csm_stack3d_build(csm_ctx* ctx, const int32_t* voxel_xyz, const uint16_t* voxel_values,
                      int num_voxels, const uint8_t* q_lut, int depth, int full_resolution_depth,
                      csm_stack3d** out) {
  CSM_REQUIRE(num_voxels >= 0 && (num_voxels == 0 || (voxel_xyz != nullptr && voxel_values != nullptr)),
              "voxel list");
  return build_stack3d(ctx, voxel_xyz, voxel_values, num_voxels, nullptr, q_lut, depth, full_resolution_depth, out);
}

int csm_stack3d_build_from_hgrid(csm_ctx* ctx, const csm_hgrid3d* grid, const uint8_t* q_lut, int depth,
                                 int full_resolution_depth, csm_stack3d** out) {
  CSM_REQUIRE(ctx != nullptr && grid != nullptr, "NULL argument");
  int n = 0, box[6] = {0, 0, 0, -1, -1, -1};
  const int rc = hgrid3d_stage(ctx, grid, &n, box, box + 3);
  if (rc != CSM_OK) return rc;
  return build_stack3d(ctx, nullptr, nullptr, n, box, q_lut, depth, full_resolution_depth, out);
}

static int build_stack3d(csm_ctx* ctx, const int32_t* voxel_xyz, const uint16_t* voxel_values, int num_voxels,
                         const int* staged_box, const uint8_t* q_lut, int depth, int full_resolution_depth,
                         csm_stack3d** out) {
  CSM_REQUIRE(ctx != nullptr && out != nullptr && q_lut != nullptr, "NULL argument");
  CSM_REQUIRE(depth >= 1 && depth <= CSM_MAX_DEPTH, "branch_and_bound_depth out of range");
  CSM_REQUIRE(full_resolution_depth >= 1, "full_resolution_depth must be >= 1");  // FCSM3D.cc:61
  CSM_CUDA(cudaSetDevice(ctx->device));
  *out = nullptr;
  csm_stack3d* s = new csm_stack3d;
  s->device = ctx->device;
  s->depth = depth;
  s->frd = full_resolution_depth;
  memset(&s->host_view, 0, sizeof(s->host_view));
  s->host_view.depth = depth;
  s->host_view.frd = full_resolution_depth;
  int lo[3] = {0, 0, 0}, hi[3] = {-1, -1, -1};
  bool any;
  if (staged_box != nullptr) {
    any = num_voxels > 0;  // a staged list holds non-default voxels only
    for (int a = 0; a < 3; ++a) {
      lo[a] = staged_box[a];
      hi[a] = staged_box[3 + a];
    }
  } else {
    any = voxel_box(voxel_xyz, voxel_values, num_voxels, true, lo, hi);
  }
  // Boxes of all levels and the shift between them (FCSM3D.cc:63-76).
  int shifts[CSM_MAX_DEPTH] = {0};
  int halfs[CSM_MAX_DEPTH] = {0};
  size_t total = 0;
  {
    int l[3] = {lo[0], lo[1], lo[2]}, h[3] = {hi[0], hi[1], hi[2]};
    int last_width = 1;
    for (int k = 0; k < depth; ++k) {
      if (k > 0) {
        const bool half = k >= full_resolution_depth;
        const int next_width = 1 << k;
        const int f = 1 << std::max(0, k - full_resolution_depth);
        const int shift = (next_width - last_width + (f - 1)) / f;
        shifts[k] = shift;
        halfs[k] = half ? 1 : 0;
        for (int a = 0; a < 3; ++a) {
          l[a] -= shift;
          if (half) {
            l[a] >>= 1;
            h[a] >>= 1;
          }
        }
        last_width = next_width;
      }
      Level3& L = s->host_view.level[k];
      for (int a = 0; a < 3; ++a) {
        L.lo[a] = any ? l[a] : 0;
        L.dim[a] = any ? h[a] - l[a] + 1 : 0;
      }
      // distance between sibling children of a parent whose children live on level k
      L.stride = (1 << k) >> std::max(0, k - full_resolution_depth + 1);
      size_t pack_elems = 1;
      for (int a = 0; a < 3; ++a) {
        L.pdim[a] = any ? L.dim[a] + L.stride : 0;
        pack_elems *= static_cast<size_t>(L.pdim[a]);
      }
      if (pack_elems >= (1ull << 31) || L.pdim[1] > 65535 || L.pdim[2] > 65535) {
        set_error("3D level %d is too large (%d x %d x %d)", k, L.dim[0], L.dim[1], L.dim[2]);
        delete s;
        return CSM_ERR_ARG;
      }
      s->level_off[k] = total;
      total += align256(static_cast<size_t>(L.dim[0]) * L.dim[1] * L.dim[2]);
      // the top level is never a child level: no pack
      s->pack_off[k] = total;
      if (k + 1 < depth) total += align256(sizeof(uint2) * pack_elems);
    }
  }
  s->bytes = total + align256(sizeof(Stack3Dev)) + 65536;
  auto fail = [&](int rc) {
    ctx->pool.give(false, s->base, s->cap);
    delete s;
    return rc;
  };
  {
    void* block = nullptr;
    if (ctx->pool.take(false, s->bytes, &block, &s->cap) != CSM_OK) {
      set_error("allocation of %zu bytes for a 3D precomputation stack failed", s->bytes);
      return fail(CSM_ERR_CUDA);
    }
    s->base = static_cast<uint8_t*>(block);
  }
  uint8_t* lut_dev = s->base + total + align256(sizeof(Stack3Dev));
  s->dev = reinterpret_cast<Stack3Dev*>(s->base + total);
  for (int k = 0; k < depth; ++k) {
    s->host_view.level[k].data = s->base + s->level_off[k];
    s->host_view.level[k].pack =
        k + 1 < depth ? reinterpret_cast<const uint2*>(s->base + s->pack_off[k]) : nullptr;
  }
  cudaStream_t st = ctx->stream;
  int rc = staged_box != nullptr ? CSM_OK : upload_voxels(ctx, voxel_xyz, voxel_values, num_voxels);
  if (rc != CSM_OK) return fail(rc);
  if (cudaMemcpyAsync(lut_dev, q_lut, 65536, cudaMemcpyHostToDevice, st) != cudaSuccess ||
      cudaMemcpyAsync(s->dev, &s->host_view, sizeof(Stack3Dev), cudaMemcpyHostToDevice, st) != cudaSuccess ||
      (total > 0 && cudaMemsetAsync(s->base, 0, total, st) != cudaSuccess)) {
    set_error("3D stack upload failed: %s", cudaGetErrorString(cudaGetLastError()));
    return fail(CSM_ERR_CUDA);
  }
  if (any) {
    const Level3& L0 = s->host_view.level[0];
    ++ctx->launches, k3_level0<<<(num_voxels + 255) / 256, 256, 0, st>>>(
        ctx->tmp0.as<int>(), ctx->tmp1.as<uint16_t>(), num_voxels, lut_dev, s->base + s->level_off[0],
        L0.lo[0], L0.lo[1], L0.lo[2], L0.dim[0], L0.dim[1]);
    for (int k = 1; k < depth; ++k) {
      const Level3& in = s->host_view.level[k - 1];
      const Level3& o = s->host_view.level[k];
      UpArgs a;
      a.in = in.data;
      a.out = s->base + s->level_off[k];
      for (int c = 0; c < 3; ++c) {
        a.in_lo[c] = in.lo[c];
        a.in_dim[c] = in.dim[c];
        a.out_lo[c] = o.lo[c];
        a.out_dim[c] = o.dim[c];
      }
      a.shift = shifts[k];
      a.half = halfs[k];
      if (o.dim[1] > 65535 || o.dim[2] > 65535) {
        set_error("3D level %d is too large (%d x %d x %d)", k, o.dim[0], o.dim[1], o.dim[2]);
        return fail(CSM_ERR_ARG);
      }
      ++ctx->launches, k3_level_up<<<dim3((o.dim[0] + 255) / 256, o.dim[1], o.dim[2]), 256, 0, st>>>(a);
    }
    for (int k = 0; k + 1 < depth; ++k) {
      const Level3& L = s->host_view.level[k];
      PackArgs p;
      p.in = L.data;
      p.out = reinterpret_cast<uint2*>(s->base + s->pack_off[k]);
      for (int c = 0; c < 3; ++c) {
        p.dim[c] = L.dim[c];
        p.pdim[c] = L.pdim[c];
      }
      p.stride = L.stride;
      ++ctx->launches, k3_pack<<<dim3((L.pdim[0] + 255) / 256, L.pdim[1], L.pdim[2]), 256, 0, st>>>(p);
    }
  }
  if (cudaStreamSynchronize(st) != cudaSuccess || cudaGetLastError() != cudaSuccess) {
    set_error("3D stack build failed: %s", cudaGetErrorString(cudaGetLastError()));
    return fail(CSM_ERR_CUDA);
  }
  *out = s;
  return CSM_OK;
}

int csm_stack3d_level(csm_ctx* ctx, const csm_stack3d* s, int level, uint8_t* out, int32_t* lo3,
                      int32_t* dims3) {
  CSM_REQUIRE(ctx != nullptr && s != nullptr, "NULL argument");
  CSM_REQUIRE(level >= 0 && level < s->depth, "level out of range");
  const Level3& L = s->host_view.level[level];
  for (int a = 0; a < 3; ++a) {
    if (lo3) lo3[a] = L.lo[a];
    if (dims3) dims3[a] = L.dim[a];
  }
  const size_t bytes = static_cast<size_t>(L.dim[0]) * L.dim[1] * L.dim[2];
  if (out != nullptr && bytes > 0) {
    CSM_CUDA(cudaSetDevice(ctx->device));
    CSM_CUDA(cudaMemcpyAsync(out, L.data, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    CSM_CUDA(cudaStreamSynchronize(ctx->stream));
  }
  return CSM_OK;
}

int csm_stack3d_depth(const csm_stack3d* s) { return s ? s->depth : 0; }
size_t csm_stack3d_device_bytes(const csm_stack3d* s) { return s ? s->bytes : 0; }

void csm_stack3d_destroy(csm_ctx* ctx, csm_stack3d* s) {
  if (!s) return;
  cudaSetDevice(s->device);
  cudaDeviceSynchronize();  // see csm_stack2d_destroy
  if (!give_if_alive(ctx, s->device, s->base, s->cap) && s->base) cudaFree(s->base);
  delete s;
}

int csm_grid3d_build(csm_ctx* ctx, const int32_t* voxel_xyz, const uint16_t* voxel_values,
                     int num_voxels, const float* probability_table, csm_grid3d** out) {
  CSM_REQUIRE(ctx != nullptr && out != nullptr && probability_table != nullptr, "NULL argument");
  CSM_REQUIRE(num_voxels >= 0 && (num_voxels == 0 || (voxel_xyz != nullptr && voxel_values != nullptr)),
              "voxel list");
  CSM_CUDA(cudaSetDevice(ctx->device));
  *out = nullptr;
  csm_grid3d* g = new csm_grid3d;
  g->device = ctx->device;
  memset(&g->host_view, 0, sizeof(g->host_view));
  int lo[3] = {0, 0, 0}, hi[3] = {-1, -1, -1};
  const bool any = voxel_box(voxel_xyz, voxel_values, num_voxels, false, lo, hi);
  size_t count = 1;
  for (int a = 0; a < 3; ++a) {
    g->host_view.lo[a] = any ? lo[a] : 0;
    g->host_view.dim[a] = any ? hi[a] - lo[a] + 1 : 0;
    count *= static_cast<size_t>(g->host_view.dim[a]);
  }
  g->host_view.unknown = probability_table[0];
  const size_t s_data = align256(sizeof(float) * count), s_desc = align256(sizeof(Grid3Dev));
  g->bytes = s_data + s_desc + sizeof(float) * 65536;
  auto fail = [&](int rc) {
    ctx->pool.give(false, g->data, g->cap);
    delete g;
    return rc;
  };
  {
    void* block = nullptr;
    if (ctx->pool.take(false, g->bytes, &block, &g->cap) != CSM_OK) {
      set_error("allocation of %zu bytes for a low-resolution grid failed", g->bytes);
      return fail(CSM_ERR_CUDA);
    }
    g->data = static_cast<float*>(block);
  }
  char* base = reinterpret_cast<char*>(g->data);
  g->dev = reinterpret_cast<Grid3Dev*>(base + s_data);
  float* table_dev = reinterpret_cast<float*>(base + s_data + s_desc);
  g->host_view.data = g->data;
  cudaStream_t st = ctx->stream;
  int rc = upload_voxels(ctx, voxel_xyz, voxel_values, num_voxels);
  if (rc != CSM_OK) return fail(rc);
  if (cudaMemcpyAsync(table_dev, probability_table, sizeof(float) * 65536, cudaMemcpyHostToDevice, st) !=
          cudaSuccess ||
      cudaMemcpyAsync(g->dev, &g->host_view, sizeof(Grid3Dev), cudaMemcpyHostToDevice, st) != cudaSuccess) {
    set_error("low-resolution grid upload failed: %s", cudaGetErrorString(cudaGetLastError()));
    return fail(CSM_ERR_CUDA);
  }
  if (any) {
    ++ctx->launches, k3_grid_fill<<<std::min<size_t>((count + 255) / 256, 4096), 256, 0, st>>>(
        g->data, count, g->host_view.unknown);
    ++ctx->launches, k3_grid_scatter<<<(num_voxels + 255) / 256, 256, 0, st>>>(
        ctx->tmp0.as<int>(), ctx->tmp1.as<uint16_t>(), num_voxels, table_dev, g->data, lo[0], lo[1], lo[2],
        g->host_view.dim[0], g->host_view.dim[1]);
  }
  if (cudaStreamSynchronize(st) != cudaSuccess || cudaGetLastError() != cudaSuccess) {
    set_error("low-resolution grid build failed: %s", cudaGetErrorString(cudaGetLastError()));
    return fail(CSM_ERR_CUDA);
  }
  *out = g;
  return CSM_OK;
}

void csm_grid3d_destroy(csm_ctx* ctx, csm_grid3d* g) {
  if (!g) return;
  cudaSetDevice(g->device);
  cudaDeviceSynchronize();  // see csm_stack2d_destroy
  if (!give_if_alive(ctx, g->device, g->data, g->cap) && g->data) cudaFree(g->data);
  delete g;
}

}  // extern "C"
