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
// Data layout in HBM: one allocation per stack; level i is a row-major
// (num_x + 2^i - 1) x (num_y + 2^i - 1) byte image at a 256-byte aligned offset whose
// rows are PITCHED to a multiple of 16 bytes (pad bytes are zero), so every kernel
// moves whole words; csm_stack2d_level() hands back the dense image.
#include <algorithm>
#include <cstdlib>
#include <cstring>

#include "csm_internal.cuh"

namespace {

constexpr int kThreads = 256;

// Level 0: out(x, y) = q_lut[cells(x, y)], zero in the row padding.  A thread makes 4
// cells = one output word; the table (32 K useful entries, one byte each) is read
// through L1 -- staging it per block cost more than the rows a block converts.
// blockDim = (words per pass, rows per block); grid = (passes, row groups).
__global__ void __launch_bounds__(kThreads)
k_stack_level0(const uint16_t* __restrict__ cells, const uint8_t* __restrict__ q_lut,
               uint8_t* __restrict__ out, int num_x, int num_y, int pitch, uint16_t* __restrict__ keep) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  const int y = blockIdx.y * blockDim.y + threadIdx.y;
  if (y >= num_y || k >= (pitch >> 2)) return;
  const size_t row = static_cast<size_t>(y) * num_x;
  const uint16_t* __restrict__ crow = cells + row;
  const int x = 4 * k;
  uint32_t w = 0u;
  if ((row & 3) == 0 && x + 3 < num_x) {
    const uint2 v = __ldg(reinterpret_cast<const uint2*>(crow + x));
    if (keep != nullptr) *reinterpret_cast<uint2*>(keep + row + x) = v;  // csm_stack2d_cells_dev
    const uint32_t b0 = __ldg(q_lut + (v.x & 0x7FFFu)), b1 = __ldg(q_lut + ((v.x >> 16) & 0x7FFFu));
    const uint32_t b2 = __ldg(q_lut + (v.y & 0x7FFFu)), b3 = __ldg(q_lut + ((v.y >> 16) & 0x7FFFu));
    w = b0 | (b1 << 8) | (b2 << 16) | (b3 << 24);
  } else {
    uint32_t c[4];
#pragma unroll
    for (int b = 0; b < 4; ++b) {
      c[b] = 0x10000u;
      if (x + b < num_x) {
        const uint16_t raw = __ldg(crow + x + b);
        if (keep != nullptr) keep[row + x + b] = raw;
        c[b] = raw & 0x7FFFu;
      }
    }
#pragma unroll
    for (int b = 0; b < 4; ++b)
      if (c[b] != 0x10000u) w |= static_cast<uint32_t>(__ldg(q_lut + c[b])) << (8 * b);
  }
  reinterpret_cast<uint32_t*>(out + static_cast<size_t>(y) * pitch)[k] = w;
}

// Level i -> i+1 (h = 2^i): out(X, Y) = max of in at (X, Y), (X-h, Y), (X, Y-h),
// (X-h, Y-h); taps outside the input image are 0.
__global__ void __launch_bounds__(kThreads)
k_stack_level_up(const uint8_t* __restrict__ in, int in_w, int in_h, int in_pitch,
                 uint8_t* __restrict__ out, int out_w, int out_h, int out_pitch, int h) {
  const int X = blockIdx.x * blockDim.x + threadIdx.x;
  const int Y = blockIdx.y * blockDim.y + threadIdx.y;
  if (X >= out_w || Y >= out_h) return;
  auto tap = [&](int x, int y) -> unsigned {
    if (static_cast<unsigned>(x) >= static_cast<unsigned>(in_w) ||
        static_cast<unsigned>(y) >= static_cast<unsigned>(in_h))
      return 0u;
    return in[static_cast<size_t>(y) * in_pitch + x];
  };
  const unsigned a = max(tap(X, Y), tap(X - h, Y));
  const unsigned b = max(tap(X, Y - h), tap(X - h, Y - h));
  out[static_cast<size_t>(Y) * out_pitch + X] = static_cast<uint8_t>(max(a, b));
}

// Byte level -> quad-phase layout (see StackDev).  One thread per output word, so the
// 4-byte stores are coalesced; the four byte taps are strided gathers served by L2.
__global__ void __launch_bounds__(kThreads)
k_stack_quads(const uint8_t* __restrict__ lvl, int wide_x, int wide_y, int pitch,
              uint32_t* __restrict__ out, int log2s, int in, int jn) {
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
      return lvl[static_cast<size_t>(y) * pitch + x];
    };
    out[idx] = tap(X, Y) | (tap(X, Y + s) << 8) | (tap(X + s, Y) << 16) | (tap(X + s, Y + s) << 24);
  }
}

// ---------------------------------------------------------------------------
// Fused build: every byte level AND every quad-phase level in one launch.
//
// All levels share the "window end" coordinate X (cell x sits at X = x): level i at X
// is the maximum of level 0 over [X - (2^i - 1), X], so level i+1 at X is the 4-tap
// maximum of level i at {X, X - 2^i}^2, and the quad word of level i at (u, v) = (X, Y)
// is exactly those four taps packed (byte0 (X-s, Y-s), byte1 (X-s, Y), byte2 (X, Y-s),
// byte3 (X, Y); see StackDev).  One block owns a kFusedTile^2 tile of output
// coordinates and stages the cells of [X0 - P, X0 + T), P = 2^depth - 1, as level-0
// bytes in shared memory (row staging: coalesced 2-byte loads, LUT through L1), then
// doubles level by level between two shared buffers, 4 bytes per thread with
// __vmaxu4, writing the tile's bytes of every level and its quad words as it goes.
// HBM traffic: the cells once (plus halo re-reads served by L2) and every output once.
// ---------------------------------------------------------------------------
#ifndef CSM_FUSED_LOG_TILE
#define CSM_FUSED_LOG_TILE 6
#endif
constexpr int kFusedLogTile = CSM_FUSED_LOG_TILE;
constexpr int kFusedTile = 1 << kFusedLogTile;
constexpr int kFusedMaxDepth = 7;
constexpr int kFusedThreads = 512;

// Byte-wise maximum of four words with the 16x2 SIMD min/max unit (VIMNMX3.U16x2):
// even and odd bytes are widened to 16-bit lanes, reduced, and re-interleaved.
__device__ __forceinline__ uint32_t max4_u8x4(uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  const uint32_t me = __vmaxu2(__vimax3_u16x2(a & 0x00FF00FFu, b & 0x00FF00FFu, c & 0x00FF00FFu),
                               d & 0x00FF00FFu);
  const uint32_t mo = __vmaxu2(__vimax3_u16x2(__byte_perm(a, 0u, 0x4341), __byte_perm(b, 0u, 0x4341),
                                              __byte_perm(c, 0u, 0x4341)),
                               __byte_perm(d, 0u, 0x4341));
  return __byte_perm(me, mo, 0x6240);
}

// Quad-phase layout from the byte levels, all levels in one launch.  A block owns one
// output row coordinate v of one level: it stages the two byte rows it needs (Y = v
// and Y = v - s, zero outside the level) in shared memory with coalesced loads, then
// writes, phase by phase, the contiguous runs of `in` words that v contributes to
// every phase plane -- full-sector stores, no gather from global memory.
struct QuadArgs {
  const uint8_t* level[CSM_MAX_DEPTH];
  uint32_t* quad[CSM_MAX_DEPTH];
  int in[CSM_MAX_DEPTH], jn[CSM_MAX_DEPTH], pitch[CSM_MAX_DEPTH];
  int num_x, num_y, depth;
};

__global__ void __launch_bounds__(kThreads)
k_stack_quad_rows(const QuadArgs a) {
  extern __shared__ uint8_t rows8[];
  const int i = blockIdx.y;
  const int s = 1 << i, two_s = 2 << i, sh = i + 1;
  const int in = a.in[i], jn = a.jn[i];
  const int v = blockIdx.x;
  if (v >= jn * two_s) return;
  const int wide_x = a.num_x + s - 1, wide_y = a.num_y + s - 1;
  const int span = in * two_s;          // u range of the plane rows
  const int pitch = (span + s + 3) & ~3;  // staged row: s zero cells, then X = 0 .. span-1
  uint8_t* cur = rows8;                 // Y = v
  uint8_t* up = rows8 + pitch;          // Y = v - s
  const uint8_t* lvl = a.level[i];
  for (int r = 0; r < 2; ++r) {
    const int Y = r == 0 ? v : v - s;
    uint8_t* dst = r == 0 ? cur : up;
    const bool row_ok = static_cast<unsigned>(Y) < static_cast<unsigned>(wide_y);
    const uint8_t* src = lvl + static_cast<size_t>(row_ok ? Y : 0) * a.pitch[i];
    for (int x = threadIdx.x; x < pitch; x += kThreads) {
      const int X = x - s;
      dst[x] = (row_ok && static_cast<unsigned>(X) < static_cast<unsigned>(wide_x)) ? __ldg(src + X)
                                                                                    : static_cast<uint8_t>(0);
    }
  }
  __syncthreads();
  uint32_t* out = a.quad[i];
  const size_t rowbase = (static_cast<size_t>((v & (two_s - 1)) << sh) * jn + (v >> sh)) * in;
  const size_t plane = static_cast<size_t>(in) * jn;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (i >= 2) {
    // Four adjacent phases per thread from four aligned shared-memory words: a 4x4 byte
    // transpose gives the four quad words; a warp's lanes walk ii, so every store is one
    // contiguous run inside a phase plane.
    const uint32_t* cur32 = reinterpret_cast<const uint32_t*>(cur);
    const uint32_t* up32 = reinterpret_cast<const uint32_t*>(up);
    const int sw = s >> 2, pitch_ok = pitch;  // pitch is a multiple of 4 => `up` is word-aligned
    (void)pitch_ok;
    for (int pg = warp; pg < (two_s >> 2); pg += kThreads / 32) {  // group of 4 phases
      uint32_t* o = out + rowbase + static_cast<size_t>(pg << 2) * plane;
      for (int ii = lane; ii < in; ii += 32) {
        const int xw = ((ii << sh) + (pg << 2) + s) >> 2;  // staged word index of X = u
        const uint32_t c0 = cur32[xw], l0 = cur32[xw - sw], c1 = up32[xw], l1 = up32[xw - sw];
        // q_j = l1.b_j | l0.b_j << 8 | c1.b_j << 16 | c0.b_j << 24
        const uint32_t a_lo = __byte_perm(l1, l0, 0x5140), a_hi = __byte_perm(l1, l0, 0x7362);
        const uint32_t b_lo = __byte_perm(c1, c0, 0x5140), b_hi = __byte_perm(c1, c0, 0x7362);
        o[ii] = __byte_perm(a_lo, b_lo, 0x5410);
        o[plane + ii] = __byte_perm(a_lo, b_lo, 0x7632);
        o[2 * plane + ii] = __byte_perm(a_hi, b_hi, 0x5410);
        o[3 * plane + ii] = __byte_perm(a_hi, b_hi, 0x7632);
      }
    }
  } else {
    for (int pu = 0; pu < two_s; ++pu) {
      for (int ii = threadIdx.x; ii < in; ii += kThreads) {
        const int x = (ii << sh) + pu + s;  // staged index of X = u
        const uint32_t w = static_cast<uint32_t>(up[x - s]) | (static_cast<uint32_t>(cur[x - s]) << 8) |
                           (static_cast<uint32_t>(up[x]) << 16) | (static_cast<uint32_t>(cur[x]) << 24);
        out[rowbase + pu * plane + ii] = w;
      }
    }
  }
}

struct FusedArgs {
  const uint16_t* cells;
  const uint8_t* q_lut;
  uint8_t* level[kFusedMaxDepth];
  uint32_t* quad[kFusedMaxDepth];
  int in[kFusedMaxDepth], jn[kFusedMaxDepth], pitch[kFusedMaxDepth];
  int num_x, num_y, depth;
};

__global__ void __launch_bounds__(kFusedThreads)
k_stack_fused(const FusedArgs a) {
  extern __shared__ uint32_t smem32[];
  // Region = tile plus a low-side halo of P = 2^(depth-1) cells: one more than the
  // widest window needs, which keeps the tile word-aligned inside the region.
  const int P = a.depth >= 3 ? (1 << (a.depth - 1)) : 4;
  const int R = kFusedTile + P;  // region side in bytes, a multiple of 4
  const int Rw = R >> 2;         // words per region row
  uint32_t* bufA = smem32;
  uint32_t* bufB = smem32 + Rw * R;
  const int X0 = blockIdx.x * kFusedTile, Y0 = blockIdx.y * kFusedTile;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  constexpr int kWarps = kFusedThreads / 32;

  // level 0 of the region: LUT[cell], 0 outside the grid.  Row staging: a warp takes a
  // region row, issues all its cell loads, then all LUT loads, then the stores, so
  // the two dependent global loads are pipelined across the row instead of serialised.
  {
    uint8_t* A8 = reinterpret_cast<uint8_t*>(bufA);
    constexpr int kMaxPerLane = (kFusedTile + (1 << (kFusedMaxDepth - 1))) / 32;
    for (int ry = warp; ry < R; ry += kWarps) {
      const int y = Y0 - P + ry;
      const bool row_ok = static_cast<unsigned>(y) < static_cast<unsigned>(a.num_y);
      const uint16_t* crow = a.cells + static_cast<size_t>(row_ok ? y : 0) * a.num_x;
      uint16_t c[kMaxPerLane];
      bool ok[kMaxPerLane];
#pragma unroll
      for (int k = 0; k < kMaxPerLane; ++k) {
        const int rx = lane + 32 * k;
        const int x = X0 - P + rx;
        ok[k] = row_ok && rx < R && static_cast<unsigned>(x) < static_cast<unsigned>(a.num_x);
        c[k] = ok[k] ? __ldg(crow + x) : static_cast<uint16_t>(0);
      }
      uint8_t v[kMaxPerLane];
#pragma unroll
      for (int k = 0; k < kMaxPerLane; ++k) v[k] = ok[k] ? __ldg(a.q_lut + c[k]) : static_cast<uint8_t>(0);
#pragma unroll
      for (int k = 0; k < kMaxPerLane; ++k) {
        const int rx = lane + 32 * k;
        if (rx < R) A8[ry * R + rx] = v[k];
      }
    }
  }
  __syncthreads();

  for (int i = 0; i < a.depth; ++i) {
    const int h = 1 << i;
    // (a) the tile's bytes of level i
    {
      const uint8_t* A8 = reinterpret_cast<const uint8_t*>(bufA);
      const int wide_x = a.num_x + h - 1, wide_y = a.num_y + h - 1;
      uint8_t* out = a.level[i];
      for (int e = tid; e < kFusedTile * kFusedTile; e += kFusedThreads) {
        const int lx = e & (kFusedTile - 1), ly = e >> kFusedLogTile;
        const int X = X0 + lx, Y = Y0 + ly;
        if (X < wide_x && Y < wide_y)
          out[static_cast<size_t>(Y) * a.pitch[i] + X] = A8[(P + ly) * R + P + lx];
      }
    }
    // Tap words of level i for the word k of region row ry: c = (X, .), l = (X - h, .),
    // rows ry and ry - h.  Reads that fall before the row start land in the previous
    // row's tail: they only feed cells inside the halo's unusable low margin (rx < h).
    const int hw = h >> 2;  // tap distance in words when h >= 4
    auto taps = [&](int ry, int k, uint32_t& c0, uint32_t& l0, uint32_t& c1, uint32_t& l1) {
      const uint32_t* row = bufA + ry * Rw + k;
      const uint32_t* up = row - h * Rw;
      c0 = row[0];
      c1 = up[0];
      if (h >= 4) {
        l0 = row[-hw];
        l1 = up[-hw];
      } else {
        l0 = __funnelshift_l(row[-1], c0, 8 * h);
        l1 = __funnelshift_l(up[-1], c1, 8 * h);
      }
    };
    const bool make_next = i + 1 < a.depth;
    if (make_next) {
      // (c) level i+1 of the whole region: byte-wise max of the tap words
      for (int ry = h + warp; ry < R; ry += kWarps) {
        for (int k = lane; k < Rw; k += 32) {
          if (ry == h && k < (hw > 0 ? hw : 1)) continue;  // would read before the buffer
          uint32_t c0, l0, c1, l1;
          taps(ry, k, c0, l0, c1, l1);
          bufB[ry * Rw + k] = max4_u8x4(c0, l0, c1, l1);
        }
      }
    }
    __syncthreads();
    uint32_t* t = bufA;
    bufA = bufB;
    bufB = t;
  }
}

int launch_build_fused(csm_ctx* ctx, csm_stack2d* s, const uint16_t* cells_dev) {
  FusedArgs a;
  memset(&a, 0, sizeof(a));
  a.cells = cells_dev;
  a.q_lut = s->q_lut_dev;
  a.num_x = s->num_x;
  a.num_y = s->num_y;
  a.depth = s->depth;
  for (int i = 0; i < s->depth; ++i) {
    a.level[i] = s->base + s->level_off[i];
    a.quad[i] = reinterpret_cast<uint32_t*>(s->base + s->quad_off[i]);
    a.in[i] = s->host_view.in[i];
    a.jn[i] = s->host_view.jn[i];
    a.pitch[i] = s->level_pitch[i];
  }
  // Tiles cover the widest byte level.
  const int ext_x = s->num_x + (1 << (s->depth - 1)) - 1, ext_y = s->num_y + (1 << (s->depth - 1)) - 1;
  const int R = kFusedTile + (s->depth >= 3 ? (1 << (s->depth - 1)) : 4);
  const size_t smem = static_cast<size_t>(2) * (R / 4) * R * sizeof(uint32_t);
  {  // per device, cheap: allow the 128 KB of dynamic shared memory
    constexpr int kMaxR = kFusedTile + (1 << (kFusedMaxDepth - 1));
    CSM_CUDA(cudaFuncSetAttribute(k_stack_fused, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  2 * (kMaxR / 4) * kMaxR * 4));
  }
  dim3 grid((ext_x + kFusedTile - 1) / kFusedTile, (ext_y + kFusedTile - 1) / kFusedTile);
  ++ctx->launches, k_stack_fused<<<grid, kFusedThreads, smem, ctx->stream>>>(a);
  {  // every quad-phase level from the byte levels
    QuadArgs q;
    memset(&q, 0, sizeof(q));
    q.num_x = s->num_x;
    q.num_y = s->num_y;
    q.depth = s->depth;
    int max_v = 0;
    size_t max_pitch = 0;
    for (int i = 0; i < s->depth; ++i) {
      q.level[i] = a.level[i];
      q.quad[i] = a.quad[i];
      q.in[i] = a.in[i];
      q.jn[i] = a.jn[i];
      q.pitch[i] = a.pitch[i];
      max_v = std::max(max_v, a.jn[i] << (i + 1));
      max_pitch = std::max<size_t>(max_pitch, (static_cast<size_t>(a.in[i]) << (i + 1)) + (1 << i) + 4);
    }
    const size_t qsmem = 2 * max_pitch;
    CSM_CUDA(cudaFuncSetAttribute(k_stack_quad_rows, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  static_cast<int>(std::max<size_t>(qsmem, 48 * 1024))));
    ++ctx->launches, k_stack_quad_rows<<<dim3(max_v, s->depth), kThreads, qsmem, ctx->stream>>>(q);
  }
  CSM_CUDA(cudaGetLastError());
  return CSM_OK;
}

// ---------------------------------------------------------------------------
// Row build (default): one launch per level i makes BOTH the quad-phase words of level
// i and the bytes of level i+1 from the bytes of level i -- they are the same four
// taps: quad(X, Y) packs level i at {X - s, X} x {Y - s, Y} and level i+1 at (X, Y) is
// their maximum.  A block owns one row Y: it stages rows Y and Y - s of level i in
// shared memory with whole-word coalesced loads (rows are pitched; s zero bytes in
// front stand for the taps left of the image), then
//   (1) next-level bytes: lane = consecutive word, 4 cells per thread, one 4-byte
//       coalesced store;
//   (2) quad words: a thread takes 4 adjacent phases pu (one staged word and its tap
//       s bytes to the left, rows Y and Y - s), transposes the 4 x 4 bytes into 4 quad
//       words and stores them to 4 phase planes at column ii; lanes walk ii, so every
//       store instruction writes contiguous runs of a plane row.
// Lanes that walk ii read staged words 2s/4 apart: the staged rows carry one pad word
// per 32 (index w + w/32), which makes every power-of-two stride conflict-free.
// No halo is recomputed (the tile-fused kernel recomputes 4x at depth 7) and no byte
// is moved one at a time; per level the kernel reads level i once (L2-resident) and
// writes each output once.
// ---------------------------------------------------------------------------
struct RowsArgs {
  const uint8_t* lvl;  // level i, pitched rows
  uint8_t* next;       // level i + 1, pitched rows (nullptr on the last level)
  uint32_t* quad;      // quad-phase words of level i
  int pitch, next_pitch, wide_y, next_wide_y;
  int in, jn, log2s;
  unsigned int in_magic;  // ceil(2^32 / in): t / in == __umulhi(t, in_magic) for t < 2^16; 0 when in == 1
};

__device__ __forceinline__ int skew(int w) { return w + (w >> 5); }

__global__ void __launch_bounds__(kThreads)
k_stack_rows(const RowsArgs a) {
  extern __shared__ uint32_t rows32[];
  // Programmatic dependent launch: this grid may be scheduled while the kernel that makes
  // level i is still draining; nothing of level i is read before the dependency resolves.
  cudaGridDependencySynchronize();
  const int s = 1 << a.log2s, two_s = 2 << a.log2s, sh = a.log2s + 1;
  const int v = blockIdx.x;  // window-end row Y of level i == row of level i + 1
  const int nthr = blockDim.x;  // sized by the launcher so that a row's words fill whole passes
  const int padw = s >= 4 ? (s >> 2) : 1;           // zero words in front of X = 0
  const int spanw = (a.in * two_s + 3) >> 2;        // staged words from X = 0
  const int nw = padw + spanw;
  uint32_t* cur = rows32;                 // Y = v
  uint32_t* up = rows32 + skew(nw) + 1;   // Y = v - s
  const int pw = a.pitch >> 2;
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    const int Y = r == 0 ? v : v - s;
    uint32_t* dst = r == 0 ? cur : up;
    const bool row_ok = static_cast<unsigned>(Y) < static_cast<unsigned>(a.wide_y);
    const uint32_t* __restrict__ src =
        reinterpret_cast<const uint32_t*>(a.lvl + static_cast<size_t>(row_ok ? Y : 0) * a.pitch);
    for (int w = threadIdx.x; w < nw; w += nthr) {
      const int k = w - padw;
      dst[skew(w)] = (row_ok && static_cast<unsigned>(k) < static_cast<unsigned>(pw)) ? __ldg(src + k) : 0u;
    }
  }
  __syncthreads();
  // The four tap words of the staged word xw: c = (X.., .), l = (X - s.., .).
  auto taps = [&](int xw, uint32_t& c0, uint32_t& l0, uint32_t& c1, uint32_t& l1) {
    const int ic = skew(xw);
    c0 = cur[ic];
    c1 = up[ic];
    if (s >= 4) {
      const int il = skew(xw - (s >> 2));
      l0 = cur[il];
      l1 = up[il];
    } else {
      const int il = skew(xw - 1);
      l0 = __funnelshift_l(cur[il], c0, 8 * s);
      l1 = __funnelshift_l(up[il], c1, 8 * s);
    }
  };
  // (1) bytes of level i + 1, row v (pad words come out zero: their taps lie outside).
  if (a.next != nullptr && v < a.next_wide_y) {
    uint32_t* __restrict__ orow = reinterpret_cast<uint32_t*>(a.next + static_cast<size_t>(v) * a.next_pitch);
    const int npw = a.next_pitch >> 2;
    for (int k = threadIdx.x; k < npw; k += nthr) {
      uint32_t m = 0u;
      if (k < spanw) {
        uint32_t c0, l0, c1, l1;
        taps(k + padw, c0, l0, c1, l1);
        m = max4_u8x4(c0, l0, c1, l1);
      }
      orow[k] = m;
    }
  }
  // (2) quad words of row v.  q_j = l1.b_j | l0.b_j << 8 | c1.b_j << 16 | c0.b_j << 24.
  uint32_t* __restrict__ out = a.quad;
  const size_t plane = static_cast<size_t>(a.in) * a.jn;
  const size_t rowbase = (static_cast<size_t>((v & (two_s - 1)) << sh) * a.jn + (v >> sh)) * a.in;
  if (a.log2s == 0) {
    // 2 phases: the word X = 4k .. 4k + 3 holds (pu, ii) = (0, 2k), (1, 2k), (0, 2k + 1), (1, 2k + 1).
    for (int k = threadIdx.x; 2 * k < a.in; k += nthr) {
      uint32_t c0, l0, c1, l1;
      taps(k + padw, c0, l0, c1, l1);
      const uint32_t a_lo = __byte_perm(l1, l0, 0x5140), a_hi = __byte_perm(l1, l0, 0x7362);
      const uint32_t b_lo = __byte_perm(c1, c0, 0x5140), b_hi = __byte_perm(c1, c0, 0x7362);
      uint32_t* o = out + rowbase + 2 * k;
      o[0] = __byte_perm(a_lo, b_lo, 0x5410);
      o[plane] = __byte_perm(a_lo, b_lo, 0x7632);
      if (2 * k + 1 < a.in) {
        o[1] = __byte_perm(a_hi, b_hi, 0x5410);
        o[plane + 1] = __byte_perm(a_hi, b_hi, 0x7632);
      }
    }
  } else {
    const int items = (two_s >> 2) * a.in;  // (group of 4 phases, ii)
    for (int t = threadIdx.x; t < items; t += nthr) {
      const int pg = a.in_magic ? static_cast<int>(__umulhi(static_cast<unsigned int>(t), a.in_magic)) : t;
      const int ii = t - pg * a.in;
      uint32_t c0, l0, c1, l1;
      taps(((ii << sh) >> 2) + pg + padw, c0, l0, c1, l1);
      const uint32_t a_lo = __byte_perm(l1, l0, 0x5140), a_hi = __byte_perm(l1, l0, 0x7362);
      const uint32_t b_lo = __byte_perm(c1, c0, 0x5140), b_hi = __byte_perm(c1, c0, 0x7362);
      uint32_t* o = out + rowbase + static_cast<size_t>(pg << 2) * plane + ii;
      o[0] = __byte_perm(a_lo, b_lo, 0x5410);
      o[plane] = __byte_perm(a_lo, b_lo, 0x7632);
      o[2 * plane] = __byte_perm(a_hi, b_hi, 0x5410);
      o[3 * plane] = __byte_perm(a_hi, b_hi, 0x7632);
    }
  }
}

// ---------------------------------------------------------------------------
// Row build, pipelined ("rows2"): the same per-level work as k_stack_rows, but a block owns
// GROUPS of kGroupRows consecutive rows, walks them with a stride of the grid, and brings the rows of
// its NEXT group into a second shared-memory buffer with cp.async while it works on the current
// one -- one row per block left the SMs waiting on a load -> barrier -> store chain (2000 x 2000:
// 11 us per level for 21 MB).  4-byte cp.async lets every word land at its skewed position.
// ---------------------------------------------------------------------------
#ifndef CSM_GROUP_ROWS
#define CSM_GROUP_ROWS 1
#endif
constexpr int kGroupRows = CSM_GROUP_ROWS;

struct Rows2Args {
  RowsArgs r;
  int rows_total;             // jn << (log2s + 1): rows of the quad planes (>= wide_y)
  int nw, rw;                 // staged words per row, row pitch in shared memory (words)
  int npw, items;             // words of a next-level row, quad work items of a row
  unsigned int npw_magic, items_magic;  // ceil(2^32 / x): t / x == __umulhi(t, magic) for t < 2^16
};

__device__ __forceinline__ void cp_async4(uint32_t* smem_dst, const uint32_t* gsrc) {
  const unsigned int d = static_cast<unsigned int>(__cvta_generic_to_shared(smem_dst));
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d), "l"(gsrc) : "memory");
}

__global__ void __launch_bounds__(kThreads)
k_stack_rows2(const Rows2Args p) {
  extern __shared__ uint32_t rows32[];
  cudaGridDependencySynchronize();
  const RowsArgs& a = p.r;
  const int s = 1 << a.log2s, two_s = 2 << a.log2s, sh = a.log2s + 1;
  const int padw = s >= 4 ? (s >> 2) : 1;
  const int pw = a.pitch >> 2;
  const int buf_words = 2 * kGroupRows * p.rw;
  const int groups = (p.rows_total + kGroupRows - 1) / kGroupRows;
  // rows 0 .. R-1 of a buffer: Y = v0 + r; rows R .. 2R-1: Y = v0 + r - s
  auto issue = [&](int g, uint32_t* buf) {
    const int v0 = g * kGroupRows;
    for (int t = threadIdx.x; t < 2 * kGroupRows * p.nw; t += kThreads) {
      const int row = t / p.nw, w = t - row * p.nw;
      const int Y = row < kGroupRows ? v0 + row : v0 + row - kGroupRows - s;
      const int k = w - padw;
      uint32_t* dst = buf + row * p.rw + skew(w);
      if (static_cast<unsigned>(Y) < static_cast<unsigned>(a.wide_y) && static_cast<unsigned>(k) < static_cast<unsigned>(pw))
        cp_async4(dst, reinterpret_cast<const uint32_t*>(a.lvl + static_cast<size_t>(Y) * a.pitch) + k);
      else
        *dst = 0u;
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  int g = blockIdx.x;
  if (g >= groups) return;
  issue(g, rows32);
  int which = 0;
  uint32_t* __restrict__ out = a.quad;
  const size_t plane = static_cast<size_t>(a.in) * a.jn;
  for (; g < groups; g += gridDim.x, which ^= 1) {
    uint32_t* buf = rows32 + which * buf_words;
    const int gn = g + gridDim.x;
    if (gn < groups) {
      issue(gn, rows32 + (which ^ 1) * buf_words);
      asm volatile("cp.async.wait_group 1;" ::: "memory");
    } else {
      asm volatile("cp.async.wait_group 0;" ::: "memory");
    }
    __syncthreads();
    const int v0 = g * kGroupRows;
    auto taps = [&](const uint32_t* cur, const uint32_t* up, int xw, uint32_t& c0, uint32_t& l0, uint32_t& c1,
                    uint32_t& l1) {
      const int ic = skew(xw);
      c0 = cur[ic];
      c1 = up[ic];
      if (s >= 4) {
        const int il = skew(xw - (s >> 2));
        l0 = cur[il];
        l1 = up[il];
      } else {
        const int il = skew(xw - 1);
        l0 = __funnelshift_l(cur[il], c0, 8 * s);
        l1 = __funnelshift_l(up[il], c1, 8 * s);
      }
    };
    // (1) bytes of level i + 1
    if (a.next != nullptr) {
      const int spanw = p.nw - padw;
      for (int t = threadIdx.x; t < kGroupRows * p.npw; t += kThreads) {
        const int row = static_cast<int>(__umulhi(static_cast<unsigned int>(t), p.npw_magic));
        const int k = t - row * p.npw, v = v0 + row;
        if (v >= a.next_wide_y) continue;
        uint32_t m = 0u;
        if (k < spanw) {
          uint32_t c0, l0, c1, l1;
          taps(buf + row * p.rw, buf + (kGroupRows + row) * p.rw, k + padw, c0, l0, c1, l1);
          m = max4_u8x4(c0, l0, c1, l1);
        }
        reinterpret_cast<uint32_t*>(a.next + static_cast<size_t>(v) * a.next_pitch)[k] = m;
      }
    }
    // (2) quad words
    for (int t = threadIdx.x; t < kGroupRows * p.items; t += kThreads) {
      const int row = static_cast<int>(__umulhi(static_cast<unsigned int>(t), p.items_magic));
      const int e = t - row * p.items, v = v0 + row;
      if (v >= p.rows_total) continue;
      const uint32_t* cur = buf + row * p.rw;
      const uint32_t* up = buf + (kGroupRows + row) * p.rw;
      const size_t rowbase = (static_cast<size_t>((v & (two_s - 1)) << sh) * a.jn + (v >> sh)) * a.in;
      uint32_t c0, l0, c1, l1;
      if (a.log2s == 0) {
        if (2 * e >= a.in) continue;
        taps(cur, up, e + padw, c0, l0, c1, l1);
        const uint32_t a_lo = __byte_perm(l1, l0, 0x5140), a_hi = __byte_perm(l1, l0, 0x7362);
        const uint32_t b_lo = __byte_perm(c1, c0, 0x5140), b_hi = __byte_perm(c1, c0, 0x7362);
        uint32_t* o = out + rowbase + 2 * e;
        o[0] = __byte_perm(a_lo, b_lo, 0x5410);
        o[plane] = __byte_perm(a_lo, b_lo, 0x7632);
        if (2 * e + 1 < a.in) {
          o[1] = __byte_perm(a_hi, b_hi, 0x5410);
          o[plane + 1] = __byte_perm(a_hi, b_hi, 0x7632);
        }
      } else {
        const int pg = a.in_magic ? static_cast<int>(__umulhi(static_cast<unsigned int>(e), a.in_magic)) : e;
        const int ii = e - pg * a.in;
        taps(cur, up, ((ii << sh) >> 2) + pg + padw, c0, l0, c1, l1);
        const uint32_t a_lo = __byte_perm(l1, l0, 0x5140), a_hi = __byte_perm(l1, l0, 0x7362);
        const uint32_t b_lo = __byte_perm(c1, c0, 0x5140), b_hi = __byte_perm(c1, c0, 0x7362);
        uint32_t* o = out + rowbase + static_cast<size_t>(pg << 2) * plane + ii;
        o[0] = __byte_perm(a_lo, b_lo, 0x5410);
        o[plane] = __byte_perm(a_lo, b_lo, 0x7632);
        o[2 * plane] = __byte_perm(a_hi, b_hi, 0x5410);
        o[3 * plane] = __byte_perm(a_hi, b_hi, 0x7632);
      }
    }
    __syncthreads();  // the buffer is refilled two iterations from now, by the next issue()
  }
}

// ---------------------------------------------------------------------------
// Row build, lean ("rows3", the default).  ncu on k_stack_rows (profiles/r02_ncu): one level of a
// 2000 x 2000 build issues 7.06 M warp instructions -- ~30 per staged word, ~70 per next-level
// word, ~50 per quad item -- and the issue slots, not memory, bound it.  Same work, fewer
// instructions:
//   * rows are staged with 16-byte loads into TWO copies: a linear one (16-byte aligned, for the
//     vector taps of the next level) and a word-skewed one (conflict-free for the strided reads of
//     the quad words);
//   * the next level is computed four words per thread from 16-byte shared-memory loads and
//     stored with one 16-byte store;
//   * the quad loop keeps its shape (4 adjacent phases per thread, lanes walk ii) with the
//     address arithmetic hoisted.
// PAD zero words in front of X = 0 stand for the taps left of the image: s / 4 for s >= 16, else 4
// (a multiple of four keeps the data 16-byte aligned).
// ---------------------------------------------------------------------------
struct Rows3Args {
  RowsArgs r;
  int pad, spanw, nw4;        // zero words in front, data words, staged 16-byte chunks per row
  int lin_pitch, skw_pitch;   // words per staged row (linear / skewed copy)
  int npw4, items;            // 16-byte chunks of a next-level row, quad work items of a row
  int rows_total, rows_per_block;  // a block builds rows_per_block consecutive rows
  int nq;                     // kVec: in / 4, groups of four consecutive ii
  unsigned int nq_magic;      // ceil(2^32 / nq), 0 when nq == 1
};

__device__ __forceinline__ uint4 max4_u8x16(const uint4& a, const uint4& b, const uint4& c, const uint4& d) {
  return make_uint4(max4_u8x4(a.x, b.x, c.x, d.x), max4_u8x4(a.y, b.y, c.y, d.y), max4_u8x4(a.z, b.z, c.z, d.z),
                    max4_u8x4(a.w, b.w, c.w, d.w));
}

// kVec: the quad words leave as uint4s -- a thread makes the words of four consecutive ii for four
// adjacent phases (16 words from 16 staged taps) and stores one 16-byte run per phase plane, instead
// of four 4-byte stores with their own 64-bit addresses (the quad loop was ~80 instructions per four
// words: more than half of the kernel's issue slots).  Needs in % 4 == 0 (create_stack pads it).
template <bool kVec>
__global__ void __launch_bounds__(kThreads, 8)
k_stack_rows3(const Rows3Args p) {
  extern __shared__ uint4 rows128[];
  cudaGridDependencySynchronize();
  const RowsArgs& a = p.r;
  const int s = 1 << a.log2s, two_s = 2 << a.log2s, sh = a.log2s + 1;
  const int nthr = blockDim.x;
  uint32_t* const lin_cur = reinterpret_cast<uint32_t*>(rows128);
  uint32_t* const lin_up = lin_cur + p.lin_pitch;
  uint32_t* const skw_cur = lin_up + p.lin_pitch;
  uint32_t* const skw_up = skw_cur + p.skw_pitch;
  const int pw = a.pitch >> 2;
  const size_t plane = static_cast<size_t>(a.in) * a.jn;
  // quad work items (pg, ii) = (t / in, t % in) of this thread: one division here, then the
  // loop over t += nthr advances them with a carry
  const int pg_first = a.in_magic ? static_cast<int>(__umulhi(threadIdx.x, a.in_magic)) : static_cast<int>(threadIdx.x);
  const int ii_first = static_cast<int>(threadIdx.x) - pg_first * a.in;
  const int dpg = a.in_magic ? static_cast<int>(__umulhi(static_cast<unsigned int>(nthr), a.in_magic)) : nthr;
  const int dii = nthr - dpg * a.in;
  const int v_end = min(p.rows_total, (static_cast<int>(blockIdx.x) + 1) * p.rows_per_block);
  for (int v = blockIdx.x * p.rows_per_block; v < v_end; ++v) {
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    const int Y = r == 0 ? v : v - s;
    uint32_t* const lin = r == 0 ? lin_cur : lin_up;
    uint32_t* const skw = r == 0 ? skw_cur : skw_up;
    const bool row_ok = static_cast<unsigned>(Y) < static_cast<unsigned>(a.wide_y);
    const uint4* __restrict__ src =
        reinterpret_cast<const uint4*>(a.lvl + static_cast<size_t>(row_ok ? Y : 0) * a.pitch);
    for (int q = threadIdx.x; q < p.nw4; q += nthr) {
      const int w0 = 4 * q, k0 = w0 - p.pad;  // k0 is a multiple of 4; pw too: a chunk is all in or all out
      uint4 x = make_uint4(0u, 0u, 0u, 0u);
      if (row_ok && k0 >= 0 && k0 < pw) x = __ldg(src + (k0 >> 2));
      reinterpret_cast<uint4*>(lin)[q] = x;
      const int sk = skew(w0);  // w0 % 4 == 0: the four words stay inside one 32-word run
      skw[sk] = x.x;
      skw[sk + 1] = x.y;
      skw[sk + 2] = x.z;
      skw[sk + 3] = x.w;
    }
  }
  __syncthreads();
  // (1) bytes of level i + 1, row v: four words per thread.
  if (a.next != nullptr && v < a.next_wide_y) {
    uint4* __restrict__ orow = reinterpret_cast<uint4*>(a.next + static_cast<size_t>(v) * a.next_pitch);
    const uint4* c4 = reinterpret_cast<const uint4*>(lin_cur);
    const uint4* u4 = reinterpret_cast<const uint4*>(lin_up);
    const int padq = p.pad >> 2, shiftq = s >> 4;  // left tap in 16-byte chunks when s >= 16
    for (int q = threadIdx.x; q < p.npw4; q += nthr) {
      const int k0 = 4 * q;
      uint4 m = make_uint4(0u, 0u, 0u, 0u);
      if (k0 < p.spanw) {
        const uint4 C0 = c4[q + padq], C1 = u4[q + padq];
        uint4 L0, L1;
        if (s >= 16) {
          L0 = c4[q + padq - shiftq];
          L1 = u4[q + padq - shiftq];
        } else {
          const uint4 P0 = c4[q + padq - 1], P1 = u4[q + padq - 1];
          if (s == 8) {
            L0 = make_uint4(P0.z, P0.w, C0.x, C0.y);
            L1 = make_uint4(P1.z, P1.w, C1.x, C1.y);
          } else if (s == 4) {
            L0 = make_uint4(P0.w, C0.x, C0.y, C0.z);
            L1 = make_uint4(P1.w, C1.x, C1.y, C1.z);
          } else {  // the tap is s bytes to the left: shift across neighbouring words
            const int b = 8 * s;
            L0 = make_uint4(__funnelshift_l(P0.w, C0.x, b), __funnelshift_l(C0.x, C0.y, b),
                            __funnelshift_l(C0.y, C0.z, b), __funnelshift_l(C0.z, C0.w, b));
            L1 = make_uint4(__funnelshift_l(P1.w, C1.x, b), __funnelshift_l(C1.x, C1.y, b),
                            __funnelshift_l(C1.y, C1.z, b), __funnelshift_l(C1.z, C1.w, b));
          }
        }
        m = max4_u8x16(C0, L0, C1, L1);
        // words at and beyond spanw are row padding of the next level: zero
        if (k0 + 1 >= p.spanw) m.y = 0u;
        if (k0 + 2 >= p.spanw) m.z = 0u;
        if (k0 + 3 >= p.spanw) m.w = 0u;
      }
      orow[q] = m;
    }
  }
  // (2) quad words of row v.  q_j = l1.b_j | l0.b_j << 8 | c1.b_j << 16 | c0.b_j << 24.
  auto taps = [&](int xw, uint32_t& c0, uint32_t& l0, uint32_t& c1, uint32_t& l1) {
    const int ic = skew(xw);
    c0 = skw_cur[ic];
    c1 = skw_up[ic];
    if (s >= 4) {
      const int il = skew(xw - (s >> 2));
      l0 = skw_cur[il];
      l1 = skw_up[il];
    } else {
      const int il = skew(xw - 1);
      l0 = __funnelshift_l(skw_cur[il], c0, 8 * s);
      l1 = __funnelshift_l(skw_up[il], c1, 8 * s);
    }
  };
  uint32_t* __restrict__ const orow =
      a.quad + (static_cast<size_t>((v & (two_s - 1)) << sh) * a.jn + (v >> sh)) * a.in;
  auto quad4 = [](uint32_t c0, uint32_t l0, uint32_t c1, uint32_t l1, uint32_t (&q)[4]) {
    const uint32_t a_lo = __byte_perm(l1, l0, 0x5140), a_hi = __byte_perm(l1, l0, 0x7362);
    const uint32_t b_lo = __byte_perm(c1, c0, 0x5140), b_hi = __byte_perm(c1, c0, 0x7362);
    q[0] = __byte_perm(a_lo, b_lo, 0x5410);
    q[1] = __byte_perm(a_lo, b_lo, 0x7632);
    q[2] = __byte_perm(a_hi, b_hi, 0x5410);
    q[3] = __byte_perm(a_hi, b_hi, 0x7632);
  };
  if (kVec && a.log2s == 0) {
    // 2 phases: staged words 2t and 2t + 1 hold X = 8t .. 8t + 7 = (pu, ii) = (0, 4t), (1, 4t), (0, 4t + 1), ...
    for (int t = threadIdx.x; t < p.nq; t += nthr) {
      const int xw = 2 * t + p.pad;
      const uint32_t P0 = skw_cur[skew(xw - 1)], A0 = skw_cur[skew(xw)], B0 = skw_cur[skew(xw + 1)];
      const uint32_t P1 = skw_up[skew(xw - 1)], A1 = skw_up[skew(xw)], B1 = skw_up[skew(xw + 1)];
      uint32_t qa[4], qb[4];
      quad4(A0, __funnelshift_l(P0, A0, 8), A1, __funnelshift_l(P1, A1, 8), qa);
      quad4(B0, __funnelshift_l(A0, B0, 8), B1, __funnelshift_l(A1, B1, 8), qb);
      uint32_t* o = orow + 4 * t;
      *reinterpret_cast<uint4*>(o) = make_uint4(qa[0], qa[2], qb[0], qb[2]);
      *reinterpret_cast<uint4*>(o + plane) = make_uint4(qa[1], qa[3], qb[1], qb[3]);
    }
  } else if (kVec) {
    const int step_words = two_s >> 2, groups = two_s >> 2;
    const int t0 = threadIdx.x;
    int pg = p.nq_magic ? static_cast<int>(__umulhi(static_cast<unsigned int>(t0), p.nq_magic)) : t0;
    int q = t0 - pg * p.nq;
    const int dpg = p.nq_magic ? static_cast<int>(__umulhi(static_cast<unsigned int>(nthr), p.nq_magic)) : nthr;
    const int dq = nthr - dpg * p.nq;
    while (pg < groups) {
      uint32_t w[4][4];  // [phase within the group][ii within the four]
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        uint32_t c0, l0, c1, l1, qk[4];
        taps((4 * q + k) * step_words + pg + p.pad, c0, l0, c1, l1);
        quad4(c0, l0, c1, l1, qk);
        w[0][k] = qk[0];
        w[1][k] = qk[1];
        w[2][k] = qk[2];
        w[3][k] = qk[3];
      }
      uint32_t* o = orow + (static_cast<size_t>(pg << 2) * plane + 4 * q);
#pragma unroll
      for (int ph = 0; ph < 4; ++ph)
        *reinterpret_cast<uint4*>(o + ph * plane) = make_uint4(w[ph][0], w[ph][1], w[ph][2], w[ph][3]);
      pg += dpg;
      q += dq;
      if (q >= p.nq) {
        q -= p.nq;
        ++pg;
      }
    }
  } else if (a.log2s == 0) {
    for (int k = threadIdx.x; 2 * k < a.in; k += nthr) {
      uint32_t c0, l0, c1, l1;
      taps(k + p.pad, c0, l0, c1, l1);
      const uint32_t a_lo = __byte_perm(l1, l0, 0x5140), a_hi = __byte_perm(l1, l0, 0x7362);
      const uint32_t b_lo = __byte_perm(c1, c0, 0x5140), b_hi = __byte_perm(c1, c0, 0x7362);
      uint32_t* o = orow + 2 * k;
      o[0] = __byte_perm(a_lo, b_lo, 0x5410);
      o[plane] = __byte_perm(a_lo, b_lo, 0x7632);
      if (2 * k + 1 < a.in) {
        o[1] = __byte_perm(a_hi, b_hi, 0x5410);
        o[plane + 1] = __byte_perm(a_hi, b_hi, 0x7632);
      }
    }
  } else {
    const int step_words = two_s >> 2;  // staged words between consecutive ii
    const int groups = two_s >> 2;      // groups of 4 phases
    int pg = pg_first, ii = ii_first;
    while (pg < groups) {
      uint32_t c0, l0, c1, l1;
      taps(ii * step_words + pg + p.pad, c0, l0, c1, l1);
      const uint32_t a_lo = __byte_perm(l1, l0, 0x5140), a_hi = __byte_perm(l1, l0, 0x7362);
      const uint32_t b_lo = __byte_perm(c1, c0, 0x5140), b_hi = __byte_perm(c1, c0, 0x7362);
      uint32_t* o = orow + static_cast<size_t>(pg << 2) * plane + ii;
      o[0] = __byte_perm(a_lo, b_lo, 0x5410);
      o[plane] = __byte_perm(a_lo, b_lo, 0x7632);
      o[2 * plane] = __byte_perm(a_hi, b_hi, 0x5410);
      o[3 * plane] = __byte_perm(a_hi, b_hi, 0x7632);
      pg += dpg;
      ii += dii;
      if (ii >= a.in) {
        ii -= a.in;
        ++pg;
      }
    }
  }
  __syncthreads();  // the staged rows are overwritten by the next row of this block
  }
}

void launch_level0(csm_ctx* ctx, csm_stack2d* s, const uint16_t* cells_dev) {
  const int pw = s->level_pitch[0] >> 2;
  const int passes = (pw + kThreads - 1) / kThreads;
  const int bx = std::min(kThreads, (((pw + passes - 1) / passes) + 31) & ~31);
  const int by = std::max(1, kThreads / bx);
  const dim3 grid((pw + bx - 1) / bx, (s->num_y + by - 1) / by);
  // The level-0 pass reads every cell anyway: it also leaves the copy the stack keeps.
  ++ctx->launches, k_stack_level0<<<grid, dim3(bx, by), 0, ctx->stream>>>(
      cells_dev, s->q_lut_dev, s->base + s->level_off[0], s->num_x, s->num_y, s->level_pitch[0],
      ctx->keep_cells && cells_dev != s->cells_dev ? s->cells_dev : nullptr);
}

int launch_build_rows(csm_ctx* ctx, csm_stack2d* s, const uint16_t* cells_dev) {
  launch_level0(ctx, s, cells_dev);
  size_t max_smem = 0;
  for (int i = 0; i < s->depth; ++i) {
    const int padw = (1 << i) >= 4 ? (1 << i) >> 2 : 1;
    const int nw = padw + ((s->host_view.in[i] * (2 << i) + 3) >> 2);
    max_smem = std::max<size_t>(max_smem, 2 * static_cast<size_t>(nw + (nw >> 5) + 2) * 4);
  }
  if (max_smem > 48 * 1024)
    CSM_CUDA(cudaFuncSetAttribute(k_stack_rows, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  static_cast<int>(max_smem)));
  for (int i = 0; i < s->depth; ++i) {
    RowsArgs a;
    a.lvl = s->base + s->level_off[i];
    a.next = i + 1 < s->depth ? s->base + s->level_off[i + 1] : nullptr;
    a.quad = reinterpret_cast<uint32_t*>(s->base + s->quad_off[i]);
    a.pitch = s->level_pitch[i];
    a.next_pitch = i + 1 < s->depth ? s->level_pitch[i + 1] : 0;
    a.wide_y = s->num_y + (1 << i) - 1;
    a.next_wide_y = s->num_y + (2 << i) - 1;
    a.in = s->host_view.in[i];
    a.jn = s->host_view.jn[i];
    a.log2s = i;
    a.in_magic = a.in > 1 ? static_cast<unsigned int>((0x100000000ull + a.in - 1) / a.in) : 0u;
    const int padw = (1 << i) >= 4 ? (1 << i) >> 2 : 1;
    const int nw = padw + ((a.in * (2 << i) + 3) >> 2);
    const size_t smem = 2 * static_cast<size_t>(nw + (nw >> 5) + 2) * 4;
    // A row is ~nw words; 256 threads would run ceil(nw / 256) passes with the last one
    // mostly empty (1000 cells: 283 words = 2 passes at 55 %).  Keep the pass count, shrink
    // the block to the warps that fill it.
    const int passes = (nw + kThreads - 1) / kThreads;
    const int threads = std::min(kThreads, (((nw + passes - 1) / passes) + 31) & ~31);
    // Launched with programmatic stream serialization: the launch latency of level i + 1
    // overlaps the tail of level i (the kernel waits in cudaGridDependencySynchronize).
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(a.jn << (i + 1));
    cfg.blockDim = dim3(threads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = ctx->stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    ++ctx->launches;
    CSM_CUDA(cudaLaunchKernelEx(&cfg, k_stack_rows, a));
  }
  CSM_CUDA(cudaGetLastError());
  return CSM_OK;
}

int launch_build_rows3(csm_ctx* ctx, csm_stack2d* s, const uint16_t* cells_dev, bool vec = true) {
  launch_level0(ctx, s, cells_dev);
  std::vector<Rows3Args> args(s->depth);
  size_t max_smem = 0;
  for (int i = 0; i < s->depth; ++i) {
    Rows3Args& p = args[i];
    RowsArgs& a = p.r;
    a.lvl = s->base + s->level_off[i];
    a.next = i + 1 < s->depth ? s->base + s->level_off[i + 1] : nullptr;
    a.quad = reinterpret_cast<uint32_t*>(s->base + s->quad_off[i]);
    a.pitch = s->level_pitch[i];
    a.next_pitch = i + 1 < s->depth ? s->level_pitch[i + 1] : 0;
    a.wide_y = s->num_y + (1 << i) - 1;
    a.next_wide_y = s->num_y + (2 << i) - 1;
    a.in = s->host_view.in[i];
    a.jn = s->host_view.jn[i];
    a.log2s = i;
    a.in_magic = a.in > 1 ? static_cast<unsigned int>((0x100000000ull + a.in - 1) / a.in) : 0u;
    const int sw = 1 << i;
    p.pad = sw >= 16 ? sw >> 2 : 4;
    p.spanw = (a.in * (2 << i) + 3) >> 2;
    p.nw4 = (p.pad + p.spanw + 3) >> 2;
    p.lin_pitch = 4 * p.nw4;
    const int nw = 4 * p.nw4;
    p.skw_pitch = ((nw + (nw >> 5) + 2) + 3) & ~3;
    p.npw4 = a.next_pitch >> 4;
    p.items = i == 0 ? (a.in + 1) / 2 : ((2 << i) >> 2) * a.in;
    p.nq = a.in >> 2;
    p.nq_magic = p.nq > 1 ? static_cast<unsigned int>((0x100000000ull + p.nq - 1) / p.nq) : 0u;
    if (vec) p.items = i == 0 ? p.nq : ((2 << i) >> 2) * p.nq;
    p.rows_total = a.jn << (i + 1);
    // one wave: as many rows per block as it takes to fit the grid into the resident block slots
    // (k_stack_rows3 is held to 32 registers: 16 blocks of 128 threads per SM)
    const int slots = ctx->num_sms * (vec ? 16 : 8);
    p.rows_per_block = std::max(1, (p.rows_total + slots - 1) / slots);
    max_smem = std::max<size_t>(max_smem, (2 * static_cast<size_t>(p.lin_pitch) + 2 * p.skw_pitch) * 4);
  }
  for (const Rows3Args& p : args)  // t / in via __umulhi needs t < 2^16
    if (p.items >= 65536) return launch_build_rows(ctx, s, cells_dev);
  if (max_smem > 200 * 1024) return launch_build_rows(ctx, s, cells_dev);
  if (max_smem > 48 * 1024)
    CSM_CUDA(cudaFuncSetAttribute(vec ? k_stack_rows3<true> : k_stack_rows3<false>,
                                  cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(max_smem)));
  for (int i = 0; i < s->depth; ++i) {
    const Rows3Args& p = args[i];
    const size_t smem = (2 * static_cast<size_t>(p.lin_pitch) + 2 * p.skw_pitch) * 4;
    // the widest loop is the quad loop (items); size the block so that its passes are full
    const int widest = std::max(p.items, std::max(p.nw4, p.npw4));
    const int passes = (widest + kThreads - 1) / kThreads;
    const int threads = std::min(kThreads, (((widest + passes - 1) / passes) + 31) & ~31);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((p.rows_total + p.rows_per_block - 1) / p.rows_per_block);
    cfg.blockDim = dim3(std::max(32, threads));
    cfg.dynamicSmemBytes = smem;
    cfg.stream = ctx->stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    ++ctx->launches;
    CSM_CUDA(cudaLaunchKernelEx(&cfg, vec ? k_stack_rows3<true> : k_stack_rows3<false>, p));
  }
  CSM_CUDA(cudaGetLastError());
  return CSM_OK;
}

int launch_build_rows2(csm_ctx* ctx, csm_stack2d* s, const uint16_t* cells_dev) {
  launch_level0(ctx, s, cells_dev);
  auto magic = [](int x) { return x > 1 ? static_cast<unsigned int>((0x100000000ull + x - 1) / x) : 0u; };
  size_t max_smem = 0;
  std::vector<Rows2Args> args(s->depth);
  for (int i = 0; i < s->depth; ++i) {
    Rows2Args& p = args[i];
    RowsArgs& a = p.r;
    a.lvl = s->base + s->level_off[i];
    a.next = i + 1 < s->depth ? s->base + s->level_off[i + 1] : nullptr;
    a.quad = reinterpret_cast<uint32_t*>(s->base + s->quad_off[i]);
    a.pitch = s->level_pitch[i];
    a.next_pitch = i + 1 < s->depth ? s->level_pitch[i + 1] : 0;
    a.wide_y = s->num_y + (1 << i) - 1;
    a.next_wide_y = s->num_y + (2 << i) - 1;
    a.in = s->host_view.in[i];
    a.jn = s->host_view.jn[i];
    a.log2s = i;
    a.in_magic = magic(a.in);
    const int padw = (1 << i) >= 4 ? (1 << i) >> 2 : 1;
    p.nw = padw + ((a.in * (2 << i) + 3) >> 2);
    p.rw = p.nw + (p.nw >> 5) + 2;
    p.rows_total = a.jn << (i + 1);
    p.npw = std::max(1, a.next_pitch >> 2);
    p.items = i == 0 ? (a.in + 1) / 2 : ((2 << i) >> 2) * a.in;
    p.npw_magic = p.npw > 1 ? magic(p.npw) : 0u;
    p.items_magic = p.items > 1 ? magic(p.items) : 0u;
    max_smem = std::max<size_t>(max_smem, static_cast<size_t>(4) * kGroupRows * p.rw * 4);
  }
  // t / x via __umulhi needs t < 2^16; one-word rows (x == 1) divide by themselves
  for (const Rows2Args& p : args)
    if (kGroupRows * std::max(p.npw, p.items) >= 65536 || p.npw == 1 || p.items == 1)
      return launch_build_rows(ctx, s, cells_dev);
  if (max_smem > 200 * 1024) return launch_build_rows(ctx, s, cells_dev);
  if (max_smem > 48 * 1024)
    CSM_CUDA(cudaFuncSetAttribute(k_stack_rows2, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  static_cast<int>(max_smem)));
  for (int i = 0; i < s->depth; ++i) {
    const Rows2Args& p = args[i];
    const size_t smem = static_cast<size_t>(4) * kGroupRows * p.rw * 4;
    const int groups = (p.rows_total + kGroupRows - 1) / kGroupRows;
    const int per_sm = static_cast<int>(std::max<size_t>(1, std::min<size_t>(8, (200 * 1024) / smem)));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(std::min(groups, ctx->num_sms * per_sm));
    cfg.blockDim = dim3(kThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = ctx->stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    ++ctx->launches;
    CSM_CUDA(cudaLaunchKernelEx(&cfg, k_stack_rows2, p));
  }
  CSM_CUDA(cudaGetLastError());
  return CSM_OK;
}

int launch_build_levelwise(csm_ctx* ctx, csm_stack2d* s, const uint16_t* cells_dev);

int launch_build(csm_ctx* ctx, csm_stack2d* s, const uint16_t* cells_dev) {
  // CSM_STACK_BUILD selects the path: "rows" (default, one launch per level),
  // "fused" (tile-fused byte levels + quad re-layout, depth <= 7) or "levelwise" (the
  // plain gather kernels, kept as a second opinion for the parity tests).
  static const int mode = [] {
    const char* e = getenv("CSM_STACK_BUILD");
    if (e != nullptr && strcmp(e, "levelwise") == 0) return 2;
    if (e != nullptr && strcmp(e, "fused") == 0) return 1;
    if (e != nullptr && strcmp(e, "rows2") == 0) return 3;
    if (e != nullptr && strcmp(e, "rows1") == 0) return 4;  // round 1's row kernel
    if (e != nullptr && strcmp(e, "rows3") == 0) return 5;  // 4-byte quad stores (before the uint4 ones)
    return 0;
  }();
  if (mode == 3) return launch_build_rows2(ctx, s, cells_dev);
  if (mode == 0) return launch_build_rows3(ctx, s, cells_dev, true);
  if (mode == 5) return launch_build_rows3(ctx, s, cells_dev, false);
  if (mode == 1 && s->depth <= kFusedMaxDepth) return launch_build_fused(ctx, s, cells_dev);
  if (mode == 2) return launch_build_levelwise(ctx, s, cells_dev);
  return launch_build_rows(ctx, s, cells_dev);
}

int launch_build_levelwise(csm_ctx* ctx, csm_stack2d* s, const uint16_t* cells_dev) {
  launch_level0(ctx, s, cells_dev);
  for (int i = 0; i + 1 < s->depth; ++i) {
    const int h = 1 << i;
    const int in_w = s->num_x + h - 1, in_h = s->num_y + h - 1;
    const int out_w = s->num_x + 2 * h - 1, out_h = s->num_y + 2 * h - 1;
    dim3 block(64, 4);
    dim3 grid((out_w + block.x - 1) / block.x, (out_h + block.y - 1) / block.y);
    ++ctx->launches, k_stack_level_up<<<grid, block, 0, ctx->stream>>>(
        s->base + s->level_off[i], in_w, in_h, s->level_pitch[i], s->base + s->level_off[i + 1], out_w,
        out_h, s->level_pitch[i + 1], h);
  }
  for (int i = 0; i < s->depth; ++i) {
    const int w = s->num_x + (1 << i) - 1, hgt = s->num_y + (1 << i) - 1;
    const int in = s->host_view.in[i], jn = s->host_view.jn[i];
    const size_t total = static_cast<size_t>(4) * (1u << (2 * i)) * in * jn;
    const int qblocks = static_cast<int>(
        std::min<size_t>((total + kThreads - 1) / kThreads, static_cast<size_t>(ctx->num_sms) * 16));
    ++ctx->launches, k_stack_quads<<<qblocks, kThreads, 0, ctx->stream>>>(
        s->base + s->level_off[i], w, hgt, s->level_pitch[i],
        reinterpret_cast<uint32_t*>(s->base + s->quad_off[i]), i, in, jn);
  }
  CSM_CUDA(cudaGetLastError());
  return CSM_OK;
}

// upload false: the descriptor and the table are left for the caller to copy (csm_stack2d_build_many
// sends them with the cells in one transfer); o_dev_out receives the descriptor's offset in the block.
int create_stack(csm_ctx* ctx, int num_x, int num_y, const uint8_t* q_lut, int depth,
                 csm_stack2d** out, bool upload = true, size_t* o_dev_out = nullptr) {
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
    const size_t pitch = (w + 15) & ~static_cast<size_t>(15);
    s->level_off[i] = off;
    s->level_pitch[i] = static_cast<int>(pitch);
    off += (pitch * hgt + 255) & ~static_cast<size_t>(255);
  }
  memset(&s->host_view, 0, sizeof(s->host_view));
  for (int i = 0; i < depth; ++i) {
    const int two_s = 2 << i;
    // ceil((num_x + 2s - 1) / 2s), rounded up to a multiple of 4 words: every row of every phase plane
    // starts on a 16-byte boundary, so the build stores whole uint4s (k_stack_rows3<true>); the
    // extra words lie beyond the level and hold zeros like everything outside it.
    const int in = (((num_x + two_s - 1 + two_s - 1) / two_s) + 3) & ~3;
    const int jn = (num_y + two_s - 1 + two_s - 1) / two_s;
    s->host_view.in[i] = in;
    s->host_view.jn[i] = jn;
    s->quad_off[i] = off;
    off += (static_cast<size_t>(two_s) * two_s * in * jn * 4 + 255) & ~static_cast<size_t>(255);
  }
  s->bytes = off;
  // levels | quads | descriptor | quantisation table, one pooled block
  const size_t o_dev = off, o_lut = o_dev + ((sizeof(csm::StackDev) + 255) & ~static_cast<size_t>(255));
  void* block = nullptr;
  // ... | the grid's own uint16 cells: the refinement that follows a found match
  // (CeresScanMatcher2D) interpolates them, and they are already here for the build.
  const size_t o_cells = o_lut + 65536;
  const size_t cells_bytes = (sizeof(uint16_t) * static_cast<size_t>(num_x) * num_y + 255) & ~static_cast<size_t>(255);
  if (ctx->pool.take(false, o_cells + cells_bytes, &block, &s->base_cap) != CSM_OK) {
    csm::set_error("allocation of a %zu-byte stack failed: %s", s->bytes,
                   cudaGetErrorString(cudaGetLastError()));
    delete s;
    return CSM_ERR_CUDA;
  }
  s->base = static_cast<uint8_t*>(block);
  s->dev = reinterpret_cast<csm::StackDev*>(s->base + o_dev);
  s->q_lut_dev = s->base + o_lut;
  s->cells_dev = reinterpret_cast<uint16_t*>(s->base + o_cells);
  s->host_view.num_x = num_x;
  s->host_view.num_y = num_y;
  s->host_view.depth = depth;
  for (int i = 0; i < depth; ++i) {
    s->host_view.level[i] = s->base + s->level_off[i];
    s->host_view.quad[i] = reinterpret_cast<const uint32_t*>(s->base + s->quad_off[i]);
  }
  if (o_dev_out != nullptr) *o_dev_out = o_dev;
  if (upload) {
    CSM_CUDA(cudaMemcpyAsync(s->dev, &s->host_view, sizeof(csm::StackDev), cudaMemcpyHostToDevice,
                             ctx->stream));
    CSM_CUDA(cudaMemcpyAsync(s->q_lut_dev, q_lut, 65536, cudaMemcpyHostToDevice, ctx->stream));
    CSM_CUDA(cudaStreamSynchronize(ctx->stream));  // host_view/q_lut are read from pageable memory
  }
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
  if (rc == CSM_OK && cudaStreamSynchronize(ctx->stream) != cudaSuccess) {
    csm::set_error("stack build failed: %s", cudaGetErrorString(cudaGetLastError()));
    rc = CSM_ERR_CUDA;
  }
  if (rc != CSM_OK) {  // no half-built stack escapes
    csm_stack2d_destroy(ctx, *out);
    *out = nullptr;
  }
  return rc;
}

int csm_stack2d_build_many(csm_ctx* ctx, int count, const uint16_t* const* cells, const int* num_x,
                           const int* num_y, const uint8_t* const* q_lut, int depth, csm_stack2d** out);

int csm_stack2d_build(csm_ctx* ctx, const uint16_t* cells, int num_x, int num_y,
                      const uint8_t* q_lut, int depth, csm_stack2d** out) {
  CSM_REQUIRE(cells != nullptr, "cells must not be NULL");
  CSM_REQUIRE(q_lut != nullptr, "q_lut must not be NULL");
  CSM_REQUIRE(out != nullptr, "ctx/out must not be NULL");
  return csm_stack2d_build_many(ctx, 1, &cells, &num_x, &num_y, &q_lut, depth, out);
}

// `count` stacks, pipelined: while the device builds stack k the host stages stack k + 1.  A
// cudaMemcpyAsync from pageable memory first waits for the stream (CUDA's API synchronisation rules),
// so the cells, the table and the descriptor of a stack are copied by the host into one of a few pinned
// staging blocks laid out like the tail of the stack's device block (descriptor | table | cells) and
// travel in ONE asynchronous transfer; the build reads the cells where they land (the copy the stack
// keeps, csm_stack2d_cells_dev).  One synchronisation at the end; on failure no stack survives.
int csm_stack2d_build_many(csm_ctx* ctx, int count, const uint16_t* const* cells, const int* num_x,
                           const int* num_y, const uint8_t* const* q_lut, int depth, csm_stack2d** out) {
  CSM_REQUIRE(ctx != nullptr && count >= 0, "ctx must not be NULL, count not negative");
  if (count == 0) return CSM_OK;
  CSM_REQUIRE(out != nullptr, "out must not be NULL");
  for (int k = 0; k < count; ++k) out[k] = nullptr;
  CSM_REQUIRE(cells != nullptr && num_x != nullptr && num_y != nullptr && q_lut != nullptr, "NULL argument");
  for (int k = 0; k < count; ++k) CSM_REQUIRE(cells[k] != nullptr && q_lut[k] != nullptr, "cells / q_lut must not be NULL");
  CSM_CUDA(cudaSetDevice(ctx->device));
  constexpr int kSlots = 3;
  const size_t head = ((sizeof(csm::StackDev) + 255) & ~static_cast<size_t>(255)) + 65536;  // descriptor | table
  size_t slot_bytes = 0;
  for (int k = 0; k < count; ++k) {
    CSM_REQUIRE(num_x[k] >= 1 && num_y[k] >= 1 && num_x[k] <= 30000 && num_y[k] <= 30000, "cell limits out of range");
    slot_bytes = std::max(slot_bytes, head + sizeof(uint16_t) * static_cast<size_t>(num_x[k]) * num_y[k]);
  }
  const int slots = std::min(kSlots, count);
  void* slot[kSlots] = {nullptr, nullptr, nullptr};
  size_t slot_cap[kSlots] = {0, 0, 0};
  cudaEvent_t slot_free[kSlots] = {nullptr, nullptr, nullptr};
  int rc = CSM_OK;
  for (int i = 0; i < slots && rc == CSM_OK; ++i) {
    rc = ctx->pool.take(true, slot_bytes, &slot[i], &slot_cap[i]);
    if (rc == CSM_OK && cudaEventCreateWithFlags(&slot_free[i], cudaEventDisableTiming) != cudaSuccess) rc = CSM_ERR_CUDA;
  }
  if (rc != CSM_OK) csm::set_error("staging for %d stack builds failed: %s", count, cudaGetErrorString(cudaGetLastError()));
  for (int k = 0; k < count && rc == CSM_OK; ++k) {
    size_t o_dev = 0;
    rc = create_stack(ctx, num_x[k], num_y[k], q_lut[k], depth, &out[k], /*upload=*/false, &o_dev);
    if (rc != CSM_OK) break;
    csm_stack2d* s = out[k];
    const int i = k % slots;
    if (k >= slots && cudaEventSynchronize(slot_free[i]) != cudaSuccess) {  // the transfer that used the slot is done
      csm::set_error("stack build failed: %s", cudaGetErrorString(cudaGetLastError()));
      rc = CSM_ERR_CUDA;
      break;
    }
    uint8_t* stage = static_cast<uint8_t*>(slot[i]);
    const size_t cell_bytes = sizeof(uint16_t) * static_cast<size_t>(num_x[k]) * num_y[k];
    memcpy(stage, &s->host_view, sizeof(csm::StackDev));
    memcpy(stage + head - 65536, q_lut[k], 65536);
    // cells the caller page-locked (csm_host_register, cudaHostAlloc) need no staging copy
    cudaPointerAttributes where;
    const bool locked = cudaPointerGetAttributes(&where, cells[k]) == cudaSuccess && where.type == cudaMemoryTypeHost;
    cudaGetLastError();  // an unregistered pointer is reported as an error by older drivers
    if (!locked) memcpy(stage + head, cells[k], cell_bytes);
    // descriptor, table and cells are contiguous in the stack's block, in this order (create_stack)
    if (cudaMemcpyAsync(s->base + o_dev, stage, locked ? head : head + cell_bytes, cudaMemcpyHostToDevice,
                        ctx->stream) != cudaSuccess ||
        (locked && cudaMemcpyAsync(s->cells_dev, cells[k], cell_bytes, cudaMemcpyHostToDevice, ctx->stream) !=
                       cudaSuccess) ||
        cudaEventRecord(slot_free[i], ctx->stream) != cudaSuccess) {
      csm::set_error("grid upload failed: %s", cudaGetErrorString(cudaGetLastError()));
      rc = CSM_ERR_CUDA;
      break;
    }
    rc = launch_build(ctx, s, s->cells_dev);
  }
  if (rc == CSM_OK && cudaStreamSynchronize(ctx->stream) != cudaSuccess) {
    csm::set_error("stack build failed: %s", cudaGetErrorString(cudaGetLastError()));
    rc = CSM_ERR_CUDA;
  }
  if (rc != CSM_OK) cudaStreamSynchronize(ctx->stream);  // nothing may still read a staging block
  for (int i = 0; i < slots; ++i) {
    if (slot_free[i] != nullptr) cudaEventDestroy(slot_free[i]);
    if (slot[i] != nullptr) ctx->pool.give(true, slot[i], slot_cap[i]);
  }
  if (rc != CSM_OK) {  // no half-built stack escapes
    for (int k = 0; k < count; ++k) {
      if (out[k] != nullptr) csm_stack2d_destroy(ctx, out[k]);
      out[k] = nullptr;
    }
  }
  return rc;
}

// PrecomputationGrid2D of ANY width (the public class takes one, FCSM2D.h:49-54; the stack only
// uses powers of two): the window maximum over `width` cells is the maximum of two windows of
// p = the largest power of two <= width cells, `width - p` apart, so the levels are doubled up
// to p (k_stack_level_up with h = 1, 2, 4, ...) and one more pass with h = width - p finishes.
int csm_precomputation_grid2d(csm_ctx* ctx, const uint16_t* cells, int num_x, int num_y,
                              const uint8_t* q_lut, int width, uint8_t* out) {
  CSM_REQUIRE(ctx != nullptr && cells != nullptr && q_lut != nullptr && out != nullptr, "NULL argument");
  CSM_REQUIRE(num_x >= 1 && num_y >= 1 && num_x <= 30000 && num_y <= 30000, "cell limits out of range");
  CSM_REQUIRE(width >= 1 && width <= 30000, "width out of range");
  CSM_CUDA(cudaSetDevice(ctx->device));
  const size_t wide_x = static_cast<size_t>(num_x) + width - 1, wide_y = static_cast<size_t>(num_y) + width - 1;
  const size_t pitch = (wide_x + 15) & ~static_cast<size_t>(15);
  const size_t image = pitch * wide_y;
  const size_t cell_bytes = sizeof(uint16_t) * static_cast<size_t>(num_x) * num_y;
  const size_t lut_at = (cell_bytes + 255) & ~static_cast<size_t>(255);
  int rc = ctx->tmp0.reserve(lut_at + 65536 + 256);
  if (rc == CSM_OK) rc = ctx->tmp1.reserve(image);
  if (rc == CSM_OK) rc = ctx->tmp2.reserve(image);
  if (rc != CSM_OK) return rc;
  cudaStream_t st = ctx->stream;
  uint16_t* d_cells = ctx->tmp0.as<uint16_t>();
  uint8_t* d_lut = ctx->tmp0.as<uint8_t>() + lut_at;
  CSM_CUDA(cudaMemcpyAsync(d_cells, cells, cell_bytes, cudaMemcpyHostToDevice, st));
  CSM_CUDA(cudaMemcpyAsync(d_lut, q_lut, 65536, cudaMemcpyHostToDevice, st));
  uint8_t* buf[2] = {ctx->tmp1.as<uint8_t>(), ctx->tmp2.as<uint8_t>()};
  {  // level 0 into buf[0] at the final pitch (rows beyond num_y stay zero)
    const int pw = static_cast<int>(pitch >> 2);
    CSM_CUDA(cudaMemsetAsync(buf[0], 0, image, st));
    const int bx = std::min(kThreads, (pw + 31) & ~31), by = std::max(1, kThreads / bx);
    const dim3 grid((pw + bx - 1) / bx, (num_y + by - 1) / by);
    ++ctx->launches, k_stack_level0<<<grid, dim3(bx, by), 0, st>>>(d_cells, d_lut, buf[0], num_x, num_y,
                                                                   static_cast<int>(pitch), nullptr);
  }
  int have = 1, cur = 0;  // window width of buf[cur]
  auto widen = [&](int h) {  // width `have` -> `have + h` (h <= have)
    const int in_w = num_x + have - 1, in_h = num_y + have - 1;
    const int out_w = in_w + h, out_h = in_h + h;
    dim3 block(64, 4);
    dim3 grid((out_w + block.x - 1) / block.x, (out_h + block.y - 1) / block.y);
    ++ctx->launches, k_stack_level_up<<<grid, block, 0, st>>>(buf[cur], in_w, in_h, static_cast<int>(pitch),
                                                            buf[cur ^ 1], out_w, out_h, static_cast<int>(pitch), h);
    cur ^= 1;
    have += h;
  };
  while (2 * have <= width) widen(have);
  if (have < width) widen(width - have);
  CSM_CUDA(cudaGetLastError());
  CSM_CUDA(cudaMemcpy2DAsync(out, wide_x, buf[cur], pitch, wide_x, wide_y, cudaMemcpyDeviceToHost, st));
  CSM_CUDA(cudaStreamSynchronize(st));
  return CSM_OK;
}

int csm_stack2d_rebuild_dev(csm_ctx* ctx, csm_stack2d* stack, const uint16_t* cells_dev) {
  CSM_REQUIRE(ctx != nullptr && stack != nullptr && cells_dev != nullptr, "NULL argument");
  CSM_REQUIRE(stack->device == ctx->device, "stack lives on another device");
  CSM_CUDA(cudaSetDevice(ctx->device));
  return launch_build(ctx, stack, cells_dev);  // level 0 refreshes csm_stack2d_cells_dev too
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
    CSM_CUDA(cudaMemcpy2DAsync(out, static_cast<size_t>(w), s->base + s->level_off[level],
                               static_cast<size_t>(s->level_pitch[level]), static_cast<size_t>(w),
                               static_cast<size_t>(h), cudaMemcpyDeviceToHost, ctx->stream));
    CSM_CUDA(cudaStreamSynchronize(ctx->stream));
  }
  return CSM_OK;
}

int csm_stack2d_quad(csm_ctx* ctx, const csm_stack2d* s, int level, uint32_t* out, int* in, int* jn) {
  CSM_REQUIRE(ctx != nullptr && s != nullptr, "NULL argument");
  CSM_REQUIRE(level >= 0 && level < s->depth, "level out of range");
  if (in) *in = s->host_view.in[level];
  if (jn) *jn = s->host_view.jn[level];
  if (out) {
    const size_t words = (static_cast<size_t>(4) << (2 * level)) * s->host_view.in[level] * s->host_view.jn[level];
    CSM_CUDA(cudaSetDevice(ctx->device));
    CSM_CUDA(cudaMemcpyAsync(out, s->base + s->quad_off[level], words * 4, cudaMemcpyDeviceToHost,
                             ctx->stream));
    CSM_CUDA(cudaStreamSynchronize(ctx->stream));
  }
  return CSM_OK;
}

int csm_host_register(void* ptr, size_t bytes) {
  CSM_REQUIRE(ptr != nullptr && bytes > 0, "nothing to register");
  CSM_CUDA(cudaHostRegister(ptr, bytes, cudaHostRegisterDefault));
  return CSM_OK;
}
int csm_host_unregister(void* ptr) {
  CSM_REQUIRE(ptr != nullptr, "nothing to unregister");
  CSM_CUDA(cudaHostUnregister(ptr));
  return CSM_OK;
}

int csm_ctx_set_keep_cells(csm_ctx* ctx, int enabled) {
  CSM_REQUIRE(ctx != nullptr, "ctx must not be NULL");
  ctx->keep_cells = enabled != 0;
  return CSM_OK;
}

int csm_stack2d_depth(const csm_stack2d* s) { return s ? s->depth : 0; }
const uint16_t* csm_stack2d_cells_dev(const csm_stack2d* s) { return s ? s->cells_dev : nullptr; }
size_t csm_stack2d_device_bytes(const csm_stack2d* s) { return s ? s->bytes + 65536 : 0; }

void csm_stack2d_destroy(csm_ctx* ctx, csm_stack2d* s) {
  if (!s) return;
  cudaSetDevice(s->device);
  // `ctx` may be stale (a matcher can outlive the context it was built on): only a
  // live context of the stack's device gets the block back into its pool.
  // A stack may be searched from any context of its device (csm_b200.h), so the whole device is
  // drained, not only the building context's stream, before the block can be handed out again.
  cudaDeviceSynchronize();
  if (!csm::give_if_alive(ctx, s->device, s->base, s->base_cap) && s->base) cudaFree(s->base);
  delete s;
}

}  // extern "C"
