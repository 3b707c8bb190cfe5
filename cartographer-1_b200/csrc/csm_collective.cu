// A new submap's grid reaching every GPU that will search it ("NCCL only for
// broadcasting new submap grids").  The reference has no counterpart -- its submaps live in one
// address space (mapping/internal/constraints/constraint_builder_2d.cc:165-186 builds a matcher
// from the Grid2D it is handed); with one process per GPU the owner of a finished submap sends
// its uint16 cells once over NVLink and every replica builds the precomputation stack from its
// own copy, instead of each replica pulling 2 MB over PCIe from host memory (measured in
// bench.py: 24 / 46 us against 67 / 218 us for 2 / 8 MB).
//
// NCCL is reached through dlsym on the library the process already has (torch's bundled
// libnccl.so.2, else the system's): the C ABI stays free of NCCL headers and link-time
// dependencies, and the communicator is the caller's.
#include <dlfcn.h>

#include <mutex>

#include "csm_internal.cuh"

using namespace csm;

namespace {

// ncclResult_t ncclBroadcast(const void* sendbuff, void* recvbuff, size_t count,
//                            ncclDataType_t datatype, int root, ncclComm_t comm, cudaStream_t stream)
typedef int (*NcclBroadcastFn)(const void*, void*, size_t, int, int, void*, cudaStream_t);
typedef const char* (*NcclGetErrorStringFn)(int);
constexpr int kNcclUint8 = 1;  // ncclDataType_t: ncclInt8 = 0, ncclUint8 = 1 (nccl.h)

struct Nccl {
  NcclBroadcastFn broadcast = nullptr;
  NcclGetErrorStringFn error_string = nullptr;
};

const Nccl* nccl() {
  static Nccl api;
  static std::once_flag once;
  std::call_once(once, [] {
    void* lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);  // the one torch.distributed loaded
    if (lib == nullptr) lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (lib == nullptr) lib = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (lib == nullptr) return;
    api.broadcast = reinterpret_cast<NcclBroadcastFn>(dlsym(lib, "ncclBroadcast"));
    api.error_string = reinterpret_cast<NcclGetErrorStringFn>(dlsym(lib, "ncclGetErrorString"));
  });
  return api.broadcast != nullptr ? &api : nullptr;
}

}  // namespace

extern "C" {

int csm_grid2d_broadcast_dev(csm_ctx* ctx, void* nccl_comm, int root, uint16_t* cells_dev, size_t num_cells) {
  CSM_REQUIRE(ctx != nullptr && nccl_comm != nullptr && cells_dev != nullptr, "NULL argument");
  CSM_REQUIRE(num_cells > 0 && root >= 0, "num_cells must be positive and root non-negative");
  const Nccl* api = nccl();
  if (api == nullptr) {
    set_error("libnccl.so.2 with ncclBroadcast is not loadable in this process");
    return CSM_ERR_CUDA;
  }
  CSM_CUDA(cudaSetDevice(ctx->device));
  const int rc = api->broadcast(cells_dev, cells_dev, num_cells * sizeof(uint16_t), kNcclUint8, root, nccl_comm,
                                ctx->stream);
  if (rc != 0) {
    set_error("ncclBroadcast failed: %s", api->error_string ? api->error_string(rc) : "unknown NCCL error");
    return CSM_ERR_CUDA;
  }
  return CSM_OK;  // ordered on ctx->stream: what is queued there next sees the cells
}

int csm_stack2d_broadcast(csm_ctx* ctx, void* nccl_comm, int root, int is_root, const uint16_t* cells_host,
                          int num_x, int num_y, const uint8_t* q_lut, int depth, csm_stack2d** out) {
  CSM_REQUIRE(ctx != nullptr && nccl_comm != nullptr && out != nullptr && q_lut != nullptr, "NULL argument");
  CSM_REQUIRE(num_x >= 1 && num_y >= 1, "cell limits must be positive");
  CSM_REQUIRE(!is_root || cells_host != nullptr, "the root rank supplies the cells");
  CSM_CUDA(cudaSetDevice(ctx->device));
  const size_t n = static_cast<size_t>(num_x) * num_y;
  int rc = ctx->tmp2.reserve(sizeof(uint16_t) * n + 16);
  if (rc != CSM_OK) return rc;
  uint16_t* cells_dev = ctx->tmp2.as<uint16_t>();
  if (is_root) {
    CSM_CUDA(cudaMemcpyAsync(cells_dev, cells_host, sizeof(uint16_t) * n, cudaMemcpyHostToDevice, ctx->stream));
  }
  rc = csm_grid2d_broadcast_dev(ctx, nccl_comm, root, cells_dev, n);
  if (rc != CSM_OK) return rc;
  return csm_stack2d_build_dev(ctx, cells_dev, num_x, num_y, q_lut, depth, out);
}

}  // extern "C"
