// topn_static.cu -- oryx_topn_device_static: the single-query top-N scan for callers that must not
// allocate.
//
// Who needs this.  oryx_topn / oryx_topn_device take their scratch (the per-CTA candidate lists, the
// scan work counter and the query's LSH index) from a query context that the library creates with
// cudaMalloc / cudaMallocHost the first time a context is needed.  Some hosts cannot afford that:
//   * a serving thread on a latency budget -- cudaMalloc may synchronise the whole device and stall
//     every other query in flight;
//   * a process that shares the GPU with an allocator which has already claimed the memory (a
//     framework's caching pool, a fixed reservation handed to a sidecar), where a late cudaMalloc can
//     simply fail;
//   * code that runs in a phase where the CUDA runtime refuses allocation calls.
// For them this entry point runs the same kernel with its scratch in a fixed array that is part of
// the loaded module, so the call performs no device, pinned or pooled allocation at all.
//
// Cost and limits.  The array (about 598 KB) is resident on every GPU that loads the library, from
// load until the process exits.  It bounds the request: how_many <= ORYX_STATIC_MAX_HOW_MANY (56).
// There is ONE array per device, shared by all callers: it still holds whatever the previous call left
// in it (this call resets the work counter before launching; candidate slots are fully rewritten by every
// launch), and two calls whose kernels overlap on the same GPU -- different streams, different
// threads, different models -- race on it and return garbage.  Serialise them (one stream, or
// events).  The first call on each device looks up the array's device address
// (cudaGetSymbolAddress); make that first call before starting any stream capture.
#include "common.cuh"

namespace oryx {

namespace {

constexpr int kStaticMaxHowMany = ORYX_STATIC_MAX_HOW_MANY;  // 56
constexpr int kStaticK = kStaticMaxHowMany + kTopnMargin;    // 64 candidate entries per CTA
constexpr size_t kCandBytes = (size_t)kMaxGrid * kStaticK * sizeof(uint2);
constexpr size_t kCounterOffset = kCandBytes;          // unsigned int work counter
constexpr size_t kMainIndexOffset = kCandBytes + 64;   // int32 LSH index of the query
constexpr size_t kListMaxOffset = kCandBytes + 256;    // uint32 best key per CTA list
constexpr size_t kScratchBytes = kListMaxOffset + (size_t)kMaxGrid * sizeof(uint32_t);

// In host code this name is only a handle for the runtime's symbol table, never a device address.
static __device__ __align__(256) unsigned char g_topn_static_scratch[kScratchBytes];

constexpr int kMaxDevices = 64;
std::mutex g_setup_mu;
unsigned char* g_scratch_addr[kMaxDevices] = {nullptr};

// Once per device, with that device current: resolve the array's address on that device.
int32_t scratch_for_current_device(int device, unsigned char** out) {
  if (device < 0 || device >= kMaxDevices) {
    set_error("device ordinal %d outside the static-scratch table (0..%d)", device, kMaxDevices - 1);
    return ORYX_E_INVALID;
  }
  std::lock_guard<std::mutex> lk(g_setup_mu);
  if (g_scratch_addr[device] == nullptr) {
    unsigned char* ptr = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void**)&ptr, g_topn_static_scratch);
    if (e != cudaSuccess || ptr == nullptr) {
      cudaGetLastError();
      set_error("cudaGetSymbolAddress(static top-N scratch) failed on device %d: %s", device,
                cudaGetErrorString(e));
      return ORYX_E_CUDA;
    }
    g_scratch_addr[device] = ptr;
  }
  *out = g_scratch_addr[device];
  return ORYX_OK;
}

}  // namespace
}  // namespace oryx

using namespace oryx;

extern "C" int32_t oryx_topn_device_static(oryx_model* mh, int32_t kind, const float* d_query, int32_t how_many,
                                           int32_t* d_out_rows, double* d_out_scores, int32_t* d_out_n,
                                           void* stream) {
  ORYX_ARG(mh && d_query && d_out_rows && d_out_scores && d_out_n, "null argument");
  ORYX_ARG(kind == ORYX_DOT || kind == ORYX_COSINE, "kind must be ORYX_DOT or ORYX_COSINE");
  ORYX_ARG(how_many >= 1, "how_many must be >= 1");
  if (how_many > kStaticMaxHowMany) {
    set_error("how_many=%d exceeds what the static scratch holds (ORYX_STATIC_MAX_HOW_MANY=%d); "
              "use oryx_topn_device, which allocates its scratch", how_many, kStaticMaxHowMany);
    return ORYX_E_INVALID;
  }
  Model* m = reinterpret_cast<Model*>(mh);
  int prev = -1;
  if (cudaGetDevice(&prev) != cudaSuccess) { cudaGetLastError(); prev = -1; }
  if (cudaSetDevice(m->device) != cudaSuccess) {
    cudaGetLastError();
    set_error("cannot select CUDA device %d (no CPU fallback exists)", m->device);
    return ORYX_E_CUDA;
  }
  struct Restore { int p; ~Restore() { if (p >= 0) cudaSetDevice(p); } } restore{prev};

  unsigned char* base = nullptr;
  ORYX_TRY(scratch_for_current_device(m->device, &base));
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);

  // A context that owns nothing: it only points the existing launch path at the static array.
  QueryCtx ctx;
  ctx.d_cand = reinterpret_cast<uint2*>(base);
  ctx.d_counter = reinterpret_cast<unsigned int*>(base + kCounterOffset);
  ctx.d_list_max = reinterpret_cast<uint32_t*>(base + kListMaxOffset);
  ctx.d_main_index = reinterpret_cast<int32_t*>(base + kMainIndexOffset);
  // The work counter must start at zero; the array keeps the last call's contents (a call that
  // failed midway could have left it non-zero).
  ORYX_CUDA(cudaMemsetAsync(base + kCounterOffset, 0, 128, st));

  TopnRequest rq{};
  rq.kind = kind;
  rq.d_query = d_query;
  rq.how_many = how_many;
  rq.d_excl = nullptr;
  rq.n_excl = 0;
  rq.d_out_rows = d_out_rows;
  rq.d_out_scores = d_out_scores;
  rq.d_out_n = d_out_n;
  std::shared_lock<std::shared_mutex> lk(m->mu);
  ORYX_TRY(scan_topn_launch(m, &ctx, rq, st));
  return async_mark(m, st);  // records an existing event: mutators wait for this scan before moving rows
}
