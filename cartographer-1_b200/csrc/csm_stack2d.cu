// PrecomputationGridStack2D on the device.
//
// Replaces PrecomputationGrid2D::PrecomputationGrid2D / PrecomputationGridStack2D
// (/root/reference/cartographer/mapping/internal/2d/scan_matching/
// fast_correlative_scan_matcher_2d.cc:91-186).
//
// The reference computes, per level i (width w = 2^i), a float sliding-window
// maximum of 1 - |cost| over w x w cells and quantises it to uint8.  The quantiser
// is monotone, so quantise(max) == max(quantise): every level is the uint8 window
// maximum of Q[cell] with Q = q_lut (host-computed with the reference's fp32 ops),
// zero outside the grid.  In "window-end" coordinates (the reference's wide index
// X is the original x of the window's last cell) level i+1 is the 4-tap maximum of
// level i at offsets {0, 2^i}^2, which is what the kernels below evaluate.  Pure
// integer work => bit-identical to the reference's cells_.
//
// Data layout in HBM: one allocation per stack; level i is a dense row-major
// (num_x + 2^i - 1) x (num_y + 2^i - 1) byte image at a 256-byte aligned offset.
#include <cstring>

#include "csm_internal.cuh"

namespace {

constexpr int kThreads = 256;

// Level 0: out[i] = q_lut[cells[i]].  4 cells per thread, 8-byte loads / 4-byte stores.
__global__ void __launch_bounds__(kThreads)
k_stack_level0(const uint16_t* __restrict__ cells, const uint8_t* __restrict__ q_lut,
               uint8_t* __restrict__ out, size_t n) {
  __shared__ uint8_t lut[65536 / 2];  // values 0..32767; the update marker bit is masked off
  for (int i = threadIdx.x; i < 32768 / 4; i += kThreads)
    reinterpret_cast<uint32_t*>(lut)[i] = reinterpret_cast<const uint32_t*>(q_lut)[i];
  __syncthreads();
  const size_t n4 = n / 4;
  const size_t stride = static_cast<size_t>(gridDim.x) * kThreads;
  for (size_t i = static_cast<size_t>(blockIdx.x) * kThreads + threadIdx.x; i < n4; i += stride) {
    const uint2 v = reinterpret_cast<const uint2*>(cells)[i];
    const uint32_t b0 = lut[v.x & 0x7FFFu], b1 = lut[(v.x >> 16) & 0x7FFFu];
    const uint32_t b2 = lut[v.y & 0x7FFFu], b3 = lut[(v.y >> 16) & 0x7FFFu];
    reinterpret_cast<uint32_t*>(out)[i] = b0 | (b1 << 8) | (b2 << 16) | (b3 << 24);
  }
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
    const size_t i = n4 * 4 + threadIdx.x;
    out[i] = lut[cells[i] & 0x7FFFu];
  }
}

// Level i -> i+1 (h = 2^i): out(X, Y) = max of in at (X, Y), (X-h, Y), (X, Y-h),
// (X-h, Y-h); taps outside the input image are 0.
__global__ void __launch_bounds__(kThreads)
k_stack_level_up(const uint8_t* __restrict__ in, int in_w, int in_h, uint8_t* __restrict__ out,
                 int out_w, int out_h, int h) {
  const int X = blockIdx.x * blockDim.x + threadIdx.x;
  const int Y = blockIdx.y * blockDim.y + threadIdx.y;
  if (X >= out_w || Y >= out_h) return;
  auto tap = [&](int x, int y) -> unsigned {
    if (static_cast<unsigned>(x) >= static_cast<unsigned>(in_w) ||
        static_cast<unsigned>(y) >= static_cast<unsigned>(in_h))
      return 0u;
    return in[static_cast<size_t>(y) * in_w + x];
  };
  const unsigned a = max(tap(X, Y), tap(X - h, Y));
  const unsigned b = max(tap(X, Y - h), tap(X - h, Y - h));
  out[static_cast<size_t>(Y) * out_w + X] = static_cast<uint8_t>(max(a, b));
}

// Byte level -> quad-phase layout (see StackDev).  One thread per output word, so the
// 4-byte stores are coalesced; the four byte taps are strided gathers served by L2.
__global__ void __launch_bounds__(kThreads)
k_stack_quads(const uint8_t* __restrict__ lvl, int wide_x, int wide_y, uint32_t* __restrict__ out,
              int log2s, int in, int jn) {
  const int two_s = 2 << log2s, s = 1 << log2s;
  const size_t total = static_cast<size_t>(two_s) * two_s * in * jn;
  const size_t stride = static_cast<size_t>(gridDim.x) * kThreads;
  for (size_t idx = static_cast<size_t>(blockIdx.x) * kThreads + threadIdx.x; idx < total;
       idx += stride) {
    const int i = static_cast<int>(idx % in);
    size_t r = idx / in;
    const int j = static_cast<int>(r % jn);
    r /= jn;
    const int pu = static_cast<int>(r % two_s), pv = static_cast<int>(r / two_s);
    // u = x + 2s - 1; the level's wide index of window start x is X = x + s - 1 = u - s.
    const int X = i * two_s + pu - s, Y = j * two_s + pv - s;
    auto tap = [&](int x, int y) -> uint32_t {
      if (static_cast<unsigned>(x) >= static_cast<unsigned>(wide_x) ||
          static_cast<unsigned>(y) >= static_cast<unsigned>(wide_y))
        return 0u;
      return lvl[static_cast<size_t>(y) * wide_x + x];
    };
    out[idx] = tap(X, Y) | (tap(X, Y + s) << 8) | (tap(X + s, Y) << 16) | (tap(X + s, Y + s) << 24);
  }
}

int launch_build(csm_ctx* ctx, csm_stack2d* s, const uint16_t* cells_dev) {
  const size_t n = static_cast<size_t>(s->num_x) * s->num_y;
  int blocks = static_cast<int>(std::min<size_t>((n / 4 + kThreads - 1) / kThreads + 1,
                                                 static_cast<size_t>(ctx->num_sms) * 8));
  ++ctx->launches, k_stack_level0<<<blocks, kThreads, 0, ctx->stream>>>(cells_dev, s->q_lut_dev,
                                                       s->base + s->level_off[0], n);
  for (int i = 0; i + 1 < s->depth; ++i) {
    const int h = 1 << i;
    const int in_w = s->num_x + h - 1, in_h = s->num_y + h - 1;
    const int out_w = s->num_x + 2 * h - 1, out_h = s->num_y + 2 * h - 1;
    dim3 block(64, 4);
    dim3 grid((out_w + block.x - 1) / block.x, (out_h + block.y - 1) / block.y);
    ++ctx->launches, k_stack_level_up<<<grid, block, 0, ctx->stream>>>(s->base + s->level_off[i], in_w, in_h,
                                                      s->base + s->level_off[i + 1], out_w, out_h, h);
  }
  for (int i = 0; i < s->depth; ++i) {
    const int w = s->num_x + (1 << i) - 1, hgt = s->num_y + (1 << i) - 1;
    const int in = s->host_view.in[i], jn = s->host_view.jn[i];
    const size_t total = static_cast<size_t>(4) * (1u << (2 * i)) * in * jn;
    const int qblocks = static_cast<int>(
        std::min<size_t>((total + kThreads - 1) / kThreads, static_cast<size_t>(ctx->num_sms) * 16));
    ++ctx->launches, k_stack_quads<<<qblocks, kThreads, 0, ctx->stream>>>(
        s->base + s->level_off[i], w, hgt, reinterpret_cast<uint32_t*>(s->base + s->quad_off[i]), i,
        in, jn);
  }
  CSM_CUDA(cudaGetLastError());
  return CSM_OK;
}

int create_stack(csm_ctx* ctx, int num_x, int num_y, const uint8_t* q_lut, int depth,
                 csm_stack2d** out) {
  CSM_REQUIRE(ctx != nullptr && out != nullptr, "ctx/out must not be NULL");
  CSM_REQUIRE(q_lut != nullptr, "q_lut must not be NULL");
  CSM_REQUIRE(depth >= 1 && depth <= CSM_MAX_DEPTH, "branch_and_bound_depth out of range");
  CSM_REQUIRE(num_x >= 1 && num_y >= 1, "cell limits must be positive");
  CSM_REQUIRE(num_x <= 30000 && num_y <= 30000, "grid too large for 16-bit offsets");
  CSM_CUDA(cudaSetDevice(ctx->device));
  csm_stack2d* s = new csm_stack2d;
  s->device = ctx->device;
  s->num_x = num_x;
  s->num_y = num_y;
  s->depth = depth;
  size_t off = 0;
  for (int i = 0; i < depth; ++i) {
    const size_t w = static_cast<size_t>(num_x) + (1u << i) - 1;
    const size_t hgt = static_cast<size_t>(num_y) + (1u << i) - 1;
    s->level_off[i] = off;
    off += (w * hgt + 255) & ~static_cast<size_t>(255);
  }
  memset(&s->host_view, 0, sizeof(s->host_view));
  for (int i = 0; i < depth; ++i) {
    const int two_s = 2 << i;
    const int in = (num_x + two_s - 1 + two_s - 1) / two_s;  // ceil((num_x + 2s - 1) / 2s)
    const int jn = (num_y + two_s - 1 + two_s - 1) / two_s;
    s->host_view.in[i] = in;
    s->host_view.jn[i] = jn;
    s->quad_off[i] = off;
    off += (static_cast<size_t>(two_s) * two_s * in * jn * 4 + 255) & ~static_cast<size_t>(255);
  }
  s->bytes = off;
  cudaError_t e = cudaMalloc(&s->base, s->bytes);
  if (e == cudaSuccess) e = cudaMalloc(&s->q_lut_dev, 65536);
  if (e == cudaSuccess) e = cudaMalloc(&s->dev, sizeof(csm::StackDev));
  if (e != cudaSuccess) {
    csm::set_error("cudaMalloc of a %zu-byte stack failed: %s", s->bytes, cudaGetErrorString(e));
    if (s->base) cudaFree(s->base);
    delete s;
    return CSM_ERR_CUDA;
  }
  s->host_view.num_x = num_x;
  s->host_view.num_y = num_y;
  s->host_view.depth = depth;
  for (int i = 0; i < depth; ++i) {
    s->host_view.level[i] = s->base + s->level_off[i];
    s->host_view.quad[i] = reinterpret_cast<const uint32_t*>(s->base + s->quad_off[i]);
  }
  CSM_CUDA(cudaMemcpyAsync(s->dev, &s->host_view, sizeof(csm::StackDev), cudaMemcpyHostToDevice,
                           ctx->stream));
  CSM_CUDA(cudaMemcpyAsync(s->q_lut_dev, q_lut, 65536, cudaMemcpyHostToDevice, ctx->stream));
  CSM_CUDA(cudaStreamSynchronize(ctx->stream));  // host_view/q_lut are read from pageable memory
  *out = s;
  return CSM_OK;
}

}  // namespace

extern "C" {

int csm_stack2d_build_dev(csm_ctx* ctx, const uint16_t* cells_dev, int num_x, int num_y,
                          const uint8_t* q_lut, int depth, csm_stack2d** out) {
  CSM_REQUIRE(cells_dev != nullptr, "cells must not be NULL");
  int rc = create_stack(ctx, num_x, num_y, q_lut, depth, out);
  if (rc != CSM_OK) return rc;
  rc = launch_build(ctx, *out, cells_dev);
  if (rc != CSM_OK) return rc;
  CSM_CUDA(cudaStreamSynchronize(ctx->stream));
  return CSM_OK;
}

int csm_stack2d_build(csm_ctx* ctx, const uint16_t* cells, int num_x, int num_y,
                      const uint8_t* q_lut, int depth, csm_stack2d** out) {
  CSM_REQUIRE(cells != nullptr, "cells must not be NULL");
  int rc = create_stack(ctx, num_x, num_y, q_lut, depth, out);
  if (rc != CSM_OK) return rc;
  const size_t bytes = static_cast<size_t>(num_x) * num_y * sizeof(uint16_t);
  rc = ctx->tmp0.reserve(bytes + 16);
  if (rc != CSM_OK) return rc;
  CSM_CUDA(cudaMemcpyAsync(ctx->tmp0.p, cells, bytes, cudaMemcpyHostToDevice, ctx->stream));
  rc = launch_build(ctx, *out, ctx->tmp0.as<uint16_t>());
  if (rc != CSM_OK) return rc;
  CSM_CUDA(cudaStreamSynchronize(ctx->stream));
  return CSM_OK;
}

int csm_stack2d_rebuild_dev(csm_ctx* ctx, csm_stack2d* stack, const uint16_t* cells_dev) {
  CSM_REQUIRE(ctx != nullptr && stack != nullptr && cells_dev != nullptr, "NULL argument");
  CSM_REQUIRE(stack->device == ctx->device, "stack lives on another device");
  CSM_CUDA(cudaSetDevice(ctx->device));
  return launch_build(ctx, stack, cells_dev);
}

int csm_stack2d_level(csm_ctx* ctx, const csm_stack2d* s, int level, uint8_t* out, int* wide_x,
                      int* wide_y) {
  CSM_REQUIRE(ctx != nullptr && s != nullptr, "NULL argument");
  CSM_REQUIRE(level >= 0 && level < s->depth, "level out of range");
  const int w = s->num_x + (1 << level) - 1, h = s->num_y + (1 << level) - 1;
  if (wide_x) *wide_x = w;
  if (wide_y) *wide_y = h;
  if (out) {
    CSM_CUDA(cudaSetDevice(ctx->device));
    CSM_CUDA(cudaMemcpyAsync(out, s->base + s->level_off[level], static_cast<size_t>(w) * h,
                             cudaMemcpyDeviceToHost, ctx->stream));
    CSM_CUDA(cudaStreamSynchronize(ctx->stream));
  }
  return CSM_OK;
}

int csm_stack2d_depth(const csm_stack2d* s) { return s ? s->depth : 0; }
size_t csm_stack2d_device_bytes(const csm_stack2d* s) { return s ? s->bytes + 65536 : 0; }

void csm_stack2d_destroy(csm_ctx* ctx, csm_stack2d* s) {
  if (!s) return;
  cudaSetDevice(s->device);
  if (ctx) cudaStreamSynchronize(ctx->stream); else cudaDeviceSynchronize();
  if (s->base) cudaFree(s->base);
  if (s->q_lut_dev) cudaFree(s->q_lut_dev);
  if (s->dev) cudaFree(s->dev);
  delete s;
}

}  // extern "C"
