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

}  // namespace

extern "C" {

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
