// Measurement aids: launch counter and the L2 gather-rate micro-benchmark that gives
// the roofline denominator G_L2 of SURVEY.md 8(d) ("sustained random 1-byte gather
// rate of a micro-kernel reading an L2-resident table of the same size as the stack
// with a scan-shaped offset pattern").  Not on the product path.
#include "csm_internal.cuh"

namespace {

// One warp per candidate, lane = scan point: the access pattern of the reference's
// ScoreCandidates inner loop (FCSM2D.cc:314-333) mapped naively onto a warp -- every
// lane gathers one byte at table[candidate_base + point_offset].
__global__ void __launch_bounds__(256)
k_gather_bytes(const uint8_t* __restrict__ table, unsigned int table_bytes,
               const unsigned int* __restrict__ point_offsets, int num_points,
               unsigned int num_candidates, unsigned int* __restrict__ sink) {
  const int lane = threadIdx.x & 31;
  const unsigned int warps = gridDim.x * (blockDim.x >> 5);
  unsigned int acc = 0;
  for (unsigned int c = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); c < num_candidates;
       c += warps) {
    // Candidate bases hop pseudo-randomly over the table (different offsets and levels).
    const unsigned int base = (c * 2654435761u) % table_bytes;
    for (int p = lane; p < num_points; p += 32) {
      unsigned int a = base + point_offsets[p];
      if (a >= table_bytes) a -= table_bytes;
      acc += __ldg(table + a);
    }
  }
  if (acc == 0xFFFFFFFFu) *sink = acc;  // keeps the loads alive
}

// The ceiling of the sweep's access shape (k_expand_nodes / node_sums in csm_match2d.cu): a warp
// takes an "item", and for every "scan point" issues ONE warp-wide 4-byte read-only load whose
// per-point base comes from a broadcast shared-memory read (128 bases staged per pass, eight
// loads in flight, the same packed accumulate) -- and nothing else: no node lists, no
// thresholds, no bounds checks, no survivors.  The 32 lanes of a request spread evenly over
// `sectors` consecutive 32-byte sectors (4 = 32 adjacent words, 32 = one sector per lane), the
// footprint of a row run of alive nodes with alive fraction 4 / sectors; bases hop
// pseudo-randomly over an L2-resident table.  Same launch shape as the sweep (256 threads,
// 4 blocks per SM).
constexpr int kCeilStage = 128;
__global__ void __launch_bounds__(256, 4)
k_gather_ceiling(const uint32_t* __restrict__ table, unsigned int span_words, int sectors,
                 unsigned int num_items, int points_per_item, unsigned int* __restrict__ sink) {
  __shared__ int stage_all[8][kCeilStage];
  const int lane = threadIdx.x & 31;
  int* stage = stage_all[threadIdx.x >> 5];
  const unsigned int warps = gridDim.x * (blockDim.x >> 5);
  const uint32_t* __restrict__ qlane = table + (((lane * sectors) >> 5) * 8 + (lane & 7));
  uint32_t lo = 0, hi = 0;
  for (unsigned int it = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); it < num_items; it += warps) {
    for (int p0 = 0; p0 < points_per_item; p0 += kCeilStage) {
      __syncwarp();
#pragma unroll
      for (int r = 0; r < kCeilStage / 32; ++r) {
        const unsigned int h = (it * 2654435761u) ^ ((p0 + r * 32 + lane) * 2246822519u);
        stage[r * 32 + lane] = static_cast<int>((static_cast<unsigned long long>(h * 3266489917u) * span_words) >> 32);
      }
      __syncwarp();
      const int cnt = min(kCeilStage, points_per_item - p0);
      int k = 0;
      for (; k + 8 <= cnt; k += 8) {
        const int4 a = *reinterpret_cast<const int4*>(&stage[k]);
        const int4 b = *reinterpret_cast<const int4*>(&stage[k + 4]);
        const uint32_t w0 = __ldg(qlane + a.x), w1 = __ldg(qlane + a.y), w2 = __ldg(qlane + a.z),
                       w3 = __ldg(qlane + a.w), w4 = __ldg(qlane + b.x), w5 = __ldg(qlane + b.y),
                       w6 = __ldg(qlane + b.z), w7 = __ldg(qlane + b.w);
        lo += (w0 & 0x00FF00FFu) + (w1 & 0x00FF00FFu) + (w2 & 0x00FF00FFu) + (w3 & 0x00FF00FFu) +
              (w4 & 0x00FF00FFu) + (w5 & 0x00FF00FFu) + (w6 & 0x00FF00FFu) + (w7 & 0x00FF00FFu);
        hi += __byte_perm(w0, 0u, 0x4341) + __byte_perm(w1, 0u, 0x4341) + __byte_perm(w2, 0u, 0x4341) +
              __byte_perm(w3, 0u, 0x4341) + __byte_perm(w4, 0u, 0x4341) + __byte_perm(w5, 0u, 0x4341) +
              __byte_perm(w6, 0u, 0x4341) + __byte_perm(w7, 0u, 0x4341);
      }
      for (; k < cnt; ++k) {
        const uint32_t w = __ldg(qlane + stage[k]);
        lo += w & 0x00FF00FFu;
        hi += __byte_perm(w, 0u, 0x4341);
      }
    }
  }
  if (lo + hi == 0xFFFFFFFFu) *sink = lo;  // keeps the loads alive
}

}  // namespace

extern "C" {

int csm_measure_gather_ceiling(csm_ctx* ctx, size_t table_bytes, int sectors_per_request,
                               unsigned int num_items, int points_per_item, int repeats,
                               double* requests_per_second) {
  CSM_REQUIRE(ctx != nullptr && requests_per_second != nullptr, "NULL argument");
  CSM_REQUIRE(table_bytes >= (1u << 16) && table_bytes < (1ull << 32) && sectors_per_request >= 1 &&
                  sectors_per_request <= 32 && num_items >= 1 && points_per_item >= 1 && repeats >= 1,
              "ceiling benchmark sizes out of range");
  CSM_CUDA(cudaSetDevice(ctx->device));
  int rc = ctx->tmp0.reserve(table_bytes + 256);
  if (rc != CSM_OK) return rc;
  uint32_t* table = ctx->tmp0.as<uint32_t>();
  cudaStream_t st = ctx->stream;
  CSM_CUDA(cudaMemsetAsync(table, 1, table_bytes, st));
  // bases leave room for the request's footprint (32 sectors x 8 words at most)
  const unsigned int span = static_cast<unsigned int>(table_bytes / 4) - 32 * 8 - 8;
  unsigned int* sink = reinterpret_cast<unsigned int*>(ctx->tmp0.as<char>() + table_bytes);
  const int blocks = ctx->num_sms * 4;
  // warm-up: brings the table into L2
  ++ctx->launches, k_gather_ceiling<<<blocks, 256, 0, st>>>(table, span, sectors_per_request, num_items,
                                                           points_per_item, sink);
  CSM_CUDA(cudaEventRecord(ctx->ev0, st));
  for (int r = 0; r < repeats; ++r)
    ++ctx->launches, k_gather_ceiling<<<blocks, 256, 0, st>>>(table, span, sectors_per_request, num_items,
                                                             points_per_item, sink);
  CSM_CUDA(cudaEventRecord(ctx->ev1, st));
  CSM_CUDA(cudaStreamSynchronize(st));
  CSM_CUDA(cudaGetLastError());
  float ms = 0.f;
  CSM_CUDA(cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1));
  *requests_per_second = static_cast<double>(repeats) * num_items * static_cast<double>(points_per_item) / (ms * 1e-3);
  return CSM_OK;
}

unsigned long long csm_ctx_launch_count(csm_ctx* ctx) { return ctx ? ctx->launches : 0ull; }

int csm_measure_gather_rate(csm_ctx* ctx, size_t table_bytes, int row_pitch, const int32_t* point_xy,
                            int num_points, unsigned int num_candidates, int repeats,
                            double* lookups_per_second) {
  CSM_REQUIRE(ctx != nullptr && point_xy != nullptr && lookups_per_second != nullptr, "NULL argument");
  CSM_REQUIRE(table_bytes >= 4096 && table_bytes < (1ull << 31) && num_points >= 1 && row_pitch >= 1 &&
                  num_candidates >= 1 && repeats >= 1,
              "gather benchmark sizes out of range");
  CSM_CUDA(cudaSetDevice(ctx->device));
  std::vector<unsigned int> offs(num_points);
  for (int p = 0; p < num_points; ++p) {
    const long long o = static_cast<long long>(point_xy[2 * p + 1]) * row_pitch + point_xy[2 * p];
    offs[p] = static_cast<unsigned int>(((o % static_cast<long long>(table_bytes)) +
                                         static_cast<long long>(table_bytes)) %
                                        static_cast<long long>(table_bytes));
  }
  int rc = ctx->tmp0.reserve(table_bytes + 256);
  if (rc == CSM_OK) rc = ctx->tmp1.reserve(sizeof(unsigned int) * (num_points + 64));
  if (rc != CSM_OK) return rc;
  uint8_t* table = ctx->tmp0.as<uint8_t>();
  unsigned int* d_offs = ctx->tmp1.as<unsigned int>();
  unsigned int* sink = d_offs + num_points;
  cudaStream_t st = ctx->stream;
  CSM_CUDA(cudaMemsetAsync(table, 1, table_bytes, st));
  CSM_CUDA(cudaMemcpyAsync(d_offs, offs.data(), sizeof(unsigned int) * num_points,
                           cudaMemcpyHostToDevice, st));
  CSM_CUDA(cudaStreamSynchronize(st));
  const int blocks = ctx->num_sms * 8;
  // warm-up: brings the table into L2
  ++ctx->launches, k_gather_bytes<<<blocks, 256, 0, st>>>(table, static_cast<unsigned int>(table_bytes),
                                                        d_offs, num_points, num_candidates, sink);
  CSM_CUDA(cudaEventRecord(ctx->ev0, st));
  for (int r = 0; r < repeats; ++r)
    ++ctx->launches, k_gather_bytes<<<blocks, 256, 0, st>>>(
        table, static_cast<unsigned int>(table_bytes), d_offs, num_points, num_candidates, sink);
  CSM_CUDA(cudaEventRecord(ctx->ev1, st));
  CSM_CUDA(cudaStreamSynchronize(st));
  CSM_CUDA(cudaGetLastError());
  float ms = 0.f;
  CSM_CUDA(cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1));
  *lookups_per_second =
      static_cast<double>(repeats) * num_candidates * static_cast<double>(num_points) / (ms * 1e-3);
  return CSM_OK;
}

}  // extern "C"
