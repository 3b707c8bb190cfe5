// Context, error reporting and scratch-memory helpers of libcsm_b200.so.
#include <cstdarg>
#include <cstdlib>
#include <cstring>
#include <set>

#include "csm_internal.cuh"

namespace csm {

static thread_local char g_error[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_error, sizeof(g_error), fmt, ap);
  va_end(ap);
}

int DevBuf::reserve(size_t bytes) {
  if (bytes <= cap) return CSM_OK;
  if (p) {
    CSM_CUDA(cudaFree(p));
    p = nullptr;
    cap = 0;
  }
  size_t want = bytes + bytes / 4 + 256;
  CSM_CUDA(cudaMalloc(&p, want));
  cap = want;
  return CSM_OK;
}
void DevBuf::release() {
  if (p) cudaFree(p);
  p = nullptr;
  cap = 0;
}
int PinBuf::reserve(size_t bytes) {
  if (bytes <= cap) return CSM_OK;
  if (p) {
    CSM_CUDA(cudaFreeHost(p));
    p = nullptr;
    cap = 0;
  }
  size_t want = bytes + bytes / 4 + 256;
  CSM_CUDA(cudaHostAlloc(&p, want, cudaHostAllocDefault));
  cap = want;
  return CSM_OK;
}
void PinBuf::release() {
  if (p) cudaFreeHost(p);
  p = nullptr;
  cap = 0;
}

namespace {
std::mutex g_live_mu;
std::set<const csm_ctx*> g_live;
}  // namespace

// Returns the block to `ctx`'s pool if `ctx` is a live context on `device`; the registry lock is
// held across the hand-over so that a concurrent csm_ctx_destroy cannot free the pool in between.
bool give_if_alive(::csm_ctx* ctx, int device, void* p, size_t cap) {
  if (ctx == nullptr || p == nullptr) return false;
  std::lock_guard<std::mutex> l(g_live_mu);
  if (g_live.count(ctx) == 0 || ctx->device != device) return false;
  ctx->pool.give(false, p, cap);
  return true;
}

bool ctx_alive_on(const csm_ctx* ctx, int device) {
  if (ctx == nullptr) return false;
  std::lock_guard<std::mutex> l(g_live_mu);
  return g_live.count(ctx) != 0 && ctx->device == device;
}

int BlockPool::take(bool pin, size_t bytes, void** out, size_t* cap) {
  std::lock_guard<std::mutex> l(mu);
  std::vector<Block>& list = pin ? pinned : device;
  size_t& held = pin ? pinned_bytes : device_bytes;
  int best = -1;
  for (size_t i = 0; i != list.size(); ++i)
    if (list[i].cap >= bytes && list[i].cap <= 4 * bytes + (1u << 20) &&
        (best < 0 || list[i].cap < list[best].cap))
      best = static_cast<int>(i);
  if (best >= 0) {
    *out = list[best].p;
    *cap = list[best].cap;
    held -= list[best].cap;
    list.erase(list.begin() + best);
    return CSM_OK;
  }
  const size_t want = bytes + bytes / 8 + 4096;
  cudaError_t e = pin ? cudaHostAlloc(out, want, cudaHostAllocDefault) : cudaMalloc(out, want);
  if (e != cudaSuccess && !list.empty()) {
    // Out of memory with blocks parked in the pool: release them and try once more.
    cudaGetLastError();
    for (Block& b : list) {
      if (pin) cudaFreeHost(b.p); else cudaFree(b.p);
    }
    list.clear();
    held = 0;
    e = pin ? cudaHostAlloc(out, want, cudaHostAllocDefault) : cudaMalloc(out, want);
  }
  CSM_CUDA(e);
  *cap = want;
  return CSM_OK;
}
void BlockPool::give(bool pin, void* p, size_t cap) {
  if (!p) return;
  std::lock_guard<std::mutex> l(mu);
  std::vector<Block>& list = pin ? pinned : device;
  size_t& held = pin ? pinned_bytes : device_bytes;
  const size_t budget = pin ? pinned_budget : device_budget;
  list.push_back(Block{p, cap});
  held += cap;
  while (!list.empty() && (list.size() > kMaxBlocks || held > budget)) {  // drop the smallest
    size_t k = 0;
    for (size_t i = 1; i != list.size(); ++i)
      if (list[i].cap < list[k].cap) k = i;
    if (pin) cudaFreeHost(list[k].p); else cudaFree(list[k].p);
    held -= list[k].cap;
    list.erase(list.begin() + k);
  }
}
void BlockPool::release() {
  std::lock_guard<std::mutex> l(mu);
  for (Block& b : device) cudaFree(b.p);
  for (Block& b : pinned) cudaFreeHost(b.p);
  device.clear();
  pinned.clear();
  device_bytes = pinned_bytes = 0;
}

}  // namespace csm

extern "C" {

const char* csm_last_error(void) { return csm::g_error; }

int csm_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return n;
}

int csm_ctx_create(int device, csm_ctx** out) {
  CSM_REQUIRE(out != nullptr, "out must not be NULL");
  *out = nullptr;
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n == 0) {
    cudaGetLastError();
    csm::set_error("no CUDA device available (%s); this library has no CPU fallback",
                   e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
    return CSM_ERR_CUDA;
  }
  CSM_REQUIRE(device >= 0 && device < n, "device index out of range");
  CSM_CUDA(cudaSetDevice(device));
  cudaDeviceProp prop;
  CSM_CUDA(cudaGetDeviceProperties(&prop, device));
  if (prop.major < 10) {
    csm::set_error("device %d is sm_%d%d; this library is built for sm_100a only", device,
                   prop.major, prop.minor);
    return CSM_ERR_CUDA;
  }
  csm_ctx* c = new csm_ctx;
  c->device = device;
  c->num_sms = prop.multiProcessorCount;
  {
    const char* sweep = getenv("CSM_SWEEP");  // diagnostic: "tiles" selects the tile sweep
    c->sweep_tiles = sweep != nullptr && strcmp(sweep, "tiles") == 0;
    const char* ld = getenv("CSM_EXPAND_LOAD");  // experiment: "noalloc" = L1::no_allocate loads
    c->expand_noalloc = ld != nullptr && strcmp(ld, "noalloc") == 0;
    if (const char* mb = getenv("CSM_POOL_DEVICE_MB")) c->pool.device_budget = static_cast<size_t>(atoll(mb)) << 20;
    if (const char* mb = getenv("CSM_POOL_PINNED_MB")) c->pool.pinned_budget = static_cast<size_t>(atoll(mb)) << 20;
    const char* p2 = getenv("CSM_PHASE2");  // second opinion: "dfs" = the ordered walk of round 1
    c->phase2_dfs = p2 != nullptr && strcmp(p2, "dfs") == 0;
  }
  CSM_CUDA(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
  CSM_CUDA(cudaEventCreate(&c->ev0));
  CSM_CUDA(cudaEventCreate(&c->ev1));
  {
    std::lock_guard<std::mutex> l(csm::g_live_mu);
    csm::g_live.insert(c);
  }
  *out = c;
  return CSM_OK;
}

void csm_ctx_destroy(csm_ctx* c) {
  if (!c) return;
  {
    std::lock_guard<std::mutex> l(csm::g_live_mu);
    csm::g_live.erase(c);
  }
  cudaSetDevice(c->device);
  if (c->stream) cudaStreamSynchronize(c->stream);
  c->tmp0.release();
  c->tmp1.release();
  c->tmp2.release();
  c->pin0.release();
  c->pool.release();
  c->frontier[0].release();
  c->frontier[1].release();
  c->frontier_aux.release();
  if (c->ev0) cudaEventDestroy(c->ev0);
  if (c->ev1) cudaEventDestroy(c->ev1);
  if (c->stream) cudaStreamDestroy(c->stream);
  delete c;
}

int csm_ctx_synchronize(csm_ctx* c) {
  CSM_REQUIRE(c != nullptr, "ctx must not be NULL");
  CSM_CUDA(cudaSetDevice(c->device));
  CSM_CUDA(cudaStreamSynchronize(c->stream));
  return CSM_OK;
}

}  // extern "C"
