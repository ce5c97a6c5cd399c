// exchange.cu -- the one exchange step of a row-sharded query, as ONE kernel over peer memory.
//
// Reference counterpart: the tail of ALSServingModel.topN (app/oryx-app-serving/src/main/java/com/
// cloudera/oryx/app/serving/als/model/ALSServingModel.java:316-334): every partition task hands back its
// own top-N and the caller merges them.  Here a "partition" is a GPU's shard of Y.  After the local scan
// and finishing kernel, `exchange_merge_kernel`
//   1. stores this GPU's packed result record (scores | rows | count, 128 bytes at how_many = 10)
//      straight into every peer's inbox over NVLink (plain stores to peer-mapped memory), then a
//      sequence flag with system-scope release;
//   2. waits (acquire loads on its OWN inbox, bounded by a timeout) until all `world` records of this
//      sequence number have landed;
//   3. merges them by (score desc, global row asc) -- same ordering as merge_topn_kernel.
// No NCCL call, no host round trip; every GPU ends up with the merged list.  The kernel releases its
// programmatic dependents first, so the next query's scan starts while this one is still waiting for
// slower peers.  Inboxes are double-buffered by sequence parity and each rank runs its exchange kernels
// in order, which keeps a fast rank from overwriting a record a slow rank has not merged yet.
#include <algorithm>
#include <cstdlib>
#include <mutex>

#include "common.cuh"

struct oryx_exchange {
  int device = 0, world = 1, rank = 0, max_how_many = 0;
  size_t rec_bytes = 0;                       // per-rank record slot in an inbox (multiple of 128)
  unsigned char* inbox = nullptr;             // [2][world][rec_bytes], written by peers
  unsigned char* peer_inbox[ORYX_MAX_WORLD] = {nullptr};
  bool opened_ipc[ORYX_MAX_WORLD] = {false};
  unsigned char* d_local = nullptr;           // [2][rec_bytes] this rank's own packed records
  unsigned long long* d_done = nullptr;       // [0] last sequence this rank finished merging, [1] last local record copied out
  unsigned long long seq = 0;
  bool connected = false;
  // Gram all-reduce (oryx_gram_allreduce_device): a second region of the same inbox allocation, so the one
  // handle the ranks exchange covers it too.  [2 parities][world][kGramSlotBytes]
  size_t gram_off = 0;
  unsigned long long gram_seq = 0;
  int32_t* d_gram_status = nullptr;
  std::mutex gram_mu;
};

namespace oryx {

namespace {

constexpr int kExchangeThreads = 256;
constexpr int kStageBytes = 40 * 1024;
constexpr int kLocalBytes = 13 * 1024;  // tail of the staging buffer: snapshot of the local record (how_many <= 1016)

// record slot: [scores f64 x hm | rows i32 x hm | count i32 | pad ... | seq u64 at rec_bytes - 8]
size_t record_slot_bytes(int how_many) {
  size_t payload = ((size_t)how_many * 12 + 4 + 15) / 16 * 16;  // copied in 16-byte units
  return (payload + 16 + 127) / 128 * 128;
}

struct XParams {
  const unsigned char* local;  // this rank's record for this sequence (scores | rows | count)
  unsigned char* peers[ORYX_MAX_WORLD];  // base of each rank's inbox (own rank: local memory)
  unsigned char* inbox;        // own inbox base
  unsigned long long* done;   // [0] merged, [1] snapshot taken
  unsigned long long seq;
  unsigned long long timeout_ns;  // a peer that never arrives is an error, not a hang
  size_t rec_bytes;
  int world, rank, how_many, parity;
  int64_t row_stride;
  int64_t* out_rows;
  double* out_scores;
  int32_t* out_n;
};

__device__ __forceinline__ unsigned long long global_ns_x() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

__global__ void __launch_bounds__(kExchangeThreads, 1) exchange_merge_kernel(const __grid_constant__ XParams p) {
  __shared__ __align__(16) unsigned char s_stage[kStageBytes];
  __shared__ int s_timeout;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, n_warps = kExchangeThreads / 32;
  // The next query's scan does not depend on this exchange: let it start.  This kernel itself was
  // launched ahead of time and must wait for the finishing kernel that writes the local record.
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  asm volatile("griddepcontrol.wait;" ::: "memory");
  // Take the local record out of global memory at once: the finishing kernel of query seq+2 reuses
  // the buffer and is gated on this snapshot (done[1]), never on the peers.
  const int payload16 = (int)(((size_t)p.how_many * 12 + 4 + 15) / 16);
  uint4* s_local = reinterpret_cast<uint4*>(s_stage + kStageBytes - kLocalBytes);
  const bool local_in_smem = (size_t)payload16 * 16 <= (size_t)kLocalBytes;
  if (local_in_smem) {
    for (int i = tid; i < payload16; i += kExchangeThreads) s_local[i] = __ldcg(reinterpret_cast<const uint4*>(p.local) + i);
  }
  __syncthreads();
  if (tid == 0) {
    if (local_in_smem) st_release_sys(p.done + 1, p.seq);
    s_timeout = 0;
    // exchanges of one rank complete in order (the stream does not guarantee it once kernels are
    // released early)
    const unsigned long long t0 = global_ns_x();
    while (ld_acquire_sys(p.done) + 1 < p.seq) {
      if (global_ns_x() - t0 > p.timeout_ns) { s_timeout = 1; break; }
    }
  }
  __syncthreads();
  if (s_timeout) {
    // The previous exchange of this rank never finished.  Publishing now would be out of order and could
    // overwrite a parity slot a slow peer has not merged yet: fail this call cleanly instead (the peers
    // time out on their side); the host must re-create the exchange before using it again.
    if (!local_in_smem && tid == 0) st_release_sys(p.done + 1, p.seq);
    if (tid == 0) {
      *p.out_n = -2;
      __threadfence();
      st_release_sys(p.done, p.seq);
    }
    return;
  }

  // 1. publish: one warp per peer copies the payload, then the flag
  const size_t slot_off = ((size_t)p.parity * p.world + p.rank) * p.rec_bytes;
  // payload = scores | rows | count of this call's how_many; the slot's last 16 bytes hold the flag
  for (int peer = warp; peer < p.world; peer += n_warps) {
    unsigned char* dst = p.peers[peer] + slot_off;
    const uint4* src = reinterpret_cast<const uint4*>(p.local);
    for (int i = lane; i < payload16; i += 32) reinterpret_cast<uint4*>(dst)[i] = local_in_smem ? s_local[i] : __ldcg(src + i);
    __threadfence_system();
    __syncwarp();
    if (lane == 0) st_release_sys(reinterpret_cast<unsigned long long*>(dst + p.rec_bytes - 8), p.seq);
  }

  if (!local_in_smem) {  // very large how_many: the record stayed in global memory until published
    __syncthreads();
    if (tid == 0) st_release_sys(p.done + 1, p.seq);
  }

  // 2. wait for every rank's record of this sequence in our own inbox
  const unsigned char* in = p.inbox + (size_t)p.parity * p.world * p.rec_bytes;
  if (tid < p.world) {
    const unsigned long long* flag = reinterpret_cast<const unsigned long long*>(in + (size_t)tid * p.rec_bytes + p.rec_bytes - 8);
    const unsigned long long t0 = global_ns_x();
    while (ld_acquire_sys(flag) != p.seq) {
      if (global_ns_x() - t0 > p.timeout_ns) { atomicExch(&s_timeout, 1); break; }
    }
  }
  __syncthreads();
  if (s_timeout) {
    if (tid == 0) {
      *p.out_n = -2;  // a peer never arrived
      __threadfence();
      st_release_sys(p.done, p.seq);
    }
    return;
  }

  // 3. merge (rank sort by score desc, global row asc)
  const int hm = p.how_many;
  const bool staged = (size_t)p.world * p.rec_bytes <= (size_t)(kStageBytes - kLocalBytes);
  const unsigned char* base = in;
  if (staged) {
    const int n16 = (int)((size_t)p.world * p.rec_bytes / 16);
    for (int i = tid; i < n16; i += kExchangeThreads)
      reinterpret_cast<uint4*>(s_stage)[i] = __ldcg(reinterpret_cast<const uint4*>(in) + i);
    __syncthreads();
    base = s_stage;
  }
  auto score_at = [&](int l, int j) {
    const double* ptr = reinterpret_cast<const double*>(base + (size_t)l * p.rec_bytes) + j;
    return staged ? *ptr : __ldcg(ptr);
  };
  auto row_at = [&](int l, int j) {
    const int32_t* ptr = reinterpret_cast<const int32_t*>(base + (size_t)l * p.rec_bytes + (size_t)hm * 8) + j;
    return staged ? *ptr : __ldcg(ptr);
  };
  auto count_at = [&](int l) {
    const int32_t* ptr = reinterpret_cast<const int32_t*>(base + (size_t)l * p.rec_bytes + (size_t)hm * 12);
    return min(max(staged ? *ptr : __ldcg(ptr), 0), hm);
  };
  const int total = p.world * hm;
  int valid_total = 0;
  for (int l = 0; l < p.world; l++) valid_total += count_at(l);
  const int n_out = min(valid_total, hm);
  for (int i = tid; i < total; i += kExchangeThreads) {
    const int li = i / hm, pi = i - li * hm;
    if (pi >= count_at(li)) continue;
    const double si = score_at(li, pi);
    const int64_t ri = (int64_t)li * p.row_stride + row_at(li, pi);
    int rank = 0;
    for (int j = 0; j < total; j++) {
      const int lj = j / hm, pj = j - lj * hm;
      if (pj >= count_at(lj)) continue;
      const double sj = score_at(lj, pj);
      const int64_t rj = (int64_t)lj * p.row_stride + row_at(lj, pj);
      rank += ((sj > si) || (sj == si && rj < ri)) ? 1 : 0;
    }
    if (rank < n_out) { p.out_rows[rank] = ri; p.out_scores[rank] = si; }
  }
  __syncthreads();
  if (tid == 0) {
    *p.out_n = n_out;
    __threadfence();
    st_release_sys(p.done, p.seq);
  }
}

// ---- YtY = sum of the shards' Grams (ALSServingModel.java:362-368) as ONE kernel over peer memory.
// Every rank's F x F binary64 partial is cut into slices; CTA c stores slice c of the local matrix into every
// rank's inbox (plain stores over NVLink), fences, raises flag (parity, rank, c) there, waits for the world's flags
// of slice c in its OWN inbox and adds the slices in rank order -- so every rank ends up with the same bits.
// No NCCL call and no host round trip; CTAs depend on nothing but their own slice.  A rank's all-reduce kernels run
// in stream order and never release dependents early, so when a rank has seen every peer's flags of sequence s,
// every peer has finished reading sequence s-1: two parities suffice.
constexpr int kGramMaxFeatures = 256;
constexpr int kGramMaxSlices = 64;
constexpr size_t kGramPayloadBytes = (size_t)kGramMaxFeatures * kGramMaxFeatures * sizeof(double);
constexpr size_t kGramSlotBytes = kGramPayloadBytes + 1024;  // payload | [kGramMaxSlices] u64 flags | pad
constexpr int kGramThreads = 256;

struct GParams {
  double* local;                        // in: this rank's partial; out: the sum
  unsigned char* peers[ORYX_MAX_WORLD];  // base of each rank's Gram region
  unsigned long long seq, timeout_ns;
  int world, rank, parity, n_elems, per_slice;
  int32_t* status;                      // 0 ok, -2 a peer never arrived
};

__global__ void __launch_bounds__(kGramThreads) gram_allreduce_kernel(const __grid_constant__ GParams p) {
  __shared__ int s_timeout;
  const int tid = threadIdx.x, c = blockIdx.x;
  const int e0 = c * p.per_slice, e1 = min(p.n_elems, e0 + p.per_slice);
  const size_t slot = ((size_t)p.parity * p.world + p.rank) * kGramSlotBytes;
  if (tid == 0) s_timeout = 0;
  // 1. publish slice c to every rank (the own inbox included: the reduction reads all slots alike)
  for (int r = 0; r < p.world; r++) {
    double* dst = reinterpret_cast<double*>(p.peers[r] + slot);
    for (int e = e0 + tid; e < e1; e += kGramThreads) dst[e] = p.local[e];
  }
  __threadfence_system();
  __syncthreads();
  if (tid < p.world) {
    unsigned long long* flag = reinterpret_cast<unsigned long long*>(p.peers[tid] + slot + kGramPayloadBytes) + c;
    st_release_sys(flag, p.seq);
  }
  // 2. wait for slice c of every rank in the own inbox
  const unsigned char* in = p.peers[p.rank] + (size_t)p.parity * p.world * kGramSlotBytes;
  if (tid < p.world) {
    const unsigned long long* flag = reinterpret_cast<const unsigned long long*>(in + (size_t)tid * kGramSlotBytes + kGramPayloadBytes) + c;
    const unsigned long long t0 = global_ns_x();
    while (ld_acquire_sys(flag) != p.seq) {
      if (global_ns_x() - t0 > p.timeout_ns) { atomicExch(&s_timeout, 1); break; }
    }
  }
  __syncthreads();
  if (s_timeout) {
    if (tid == 0) *p.status = -2;
    return;
  }
  // 3. add the slices in rank order
  for (int e = e0 + tid; e < e1; e += kGramThreads) {
    double acc = 0.0;
    for (int r = 0; r < p.world; r++) acc += __ldcg(reinterpret_cast<const double*>(in + (size_t)r * kGramSlotBytes) + e);
    p.local[e] = acc;
  }
}

// Selects a device for the duration of a call and puts the caller's device back afterwards.
struct ScopedDevice {
  int prev = -1;
  bool ok = false;
  explicit ScopedDevice(int device) {
    if (cudaGetDevice(&prev) != cudaSuccess) { cudaGetLastError(); prev = -1; }
    ok = cudaSetDevice(device) == cudaSuccess;
    if (!ok) cudaGetLastError();
  }
  ~ScopedDevice() { if (prev >= 0) cudaSetDevice(prev); }
};
#define ORYX_SCOPED_DEVICE(dev)                                                         \
  ScopedDevice scoped_device__(dev);                                                    \
  if (!scoped_device__.ok) {                                                            \
    set_error("cannot select CUDA device %d (no CPU fallback exists)", (int)(dev));     \
    return ORYX_E_CUDA;                                                                 \
  }

}  // namespace
}  // namespace oryx

using namespace oryx;

extern "C" {

int32_t oryx_exchange_create(int32_t device, int32_t world, int32_t rank, int32_t max_how_many, oryx_exchange** out) {
  ORYX_ARG(out, "null output");
  ORYX_ARG(world >= 1 && world <= ORYX_MAX_WORLD, "world out of range [1, ORYX_MAX_WORLD]");
  ORYX_ARG(rank >= 0 && rank < world, "rank out of range");
  ORYX_ARG(max_how_many >= 1 && max_how_many <= ORYX_MAX_HOW_MANY, "max_how_many out of range");
  *out = nullptr;
  ORYX_SCOPED_DEVICE(device);
  auto x = std::make_unique<oryx_exchange>();
  x->device = device;
  x->world = world;
  x->rank = rank;
  x->max_how_many = max_how_many;
  x->rec_bytes = record_slot_bytes(max_how_many);
  x->gram_off = (2 * (size_t)world * x->rec_bytes + 255) & ~(size_t)255;
  const size_t inbox_bytes = x->gram_off + 2 * (size_t)world * kGramSlotBytes;
  ORYX_CUDA(cudaMalloc(&x->inbox, inbox_bytes));
  ORYX_CUDA(cudaMemset(x->inbox, 0, inbox_bytes));
  ORYX_CUDA(cudaMalloc(&x->d_gram_status, sizeof(int32_t)));
  ORYX_CUDA(cudaMemset(x->d_gram_status, 0, sizeof(int32_t)));
  ORYX_CUDA(cudaMalloc(&x->d_local, 2 * x->rec_bytes));
  ORYX_CUDA(cudaMemset(x->d_local, 0, 2 * x->rec_bytes));
  ORYX_CUDA(cudaMalloc(&x->d_done, 2 * sizeof(unsigned long long)));
  ORYX_CUDA(cudaMemset(x->d_done, 0, 2 * sizeof(unsigned long long)));
  ORYX_CUDA(cudaDeviceSynchronize());
  x->peer_inbox[rank] = x->inbox;
  x->connected = world == 1;
  *out = x.release();
  return ORYX_OK;
}

int32_t oryx_exchange_handle(oryx_exchange* x, void* out_handle) {
  ORYX_ARG(x && out_handle, "null argument");
  static_assert(sizeof(cudaIpcMemHandle_t) == ORYX_EXCHANGE_HANDLE_BYTES, "handle size");
  ORYX_SCOPED_DEVICE(x->device);
  cudaIpcMemHandle_t h;
  ORYX_CUDA(cudaIpcGetMemHandle(&h, x->inbox));
  std::memcpy(out_handle, &h, sizeof(h));
  return ORYX_OK;
}

int32_t oryx_exchange_connect(oryx_exchange* x, const void* handles) {
  ORYX_ARG(x && handles, "null argument");
  ORYX_SCOPED_DEVICE(x->device);
  const unsigned char* hb = reinterpret_cast<const unsigned char*>(handles);
  for (int r = 0; r < x->world; r++) {
    if (r == x->rank || x->peer_inbox[r]) continue;
    cudaIpcMemHandle_t h;
    std::memcpy(&h, hb + (size_t)r * ORYX_EXCHANGE_HANDLE_BYTES, sizeof(h));
    void* ptr = nullptr;
    cudaError_t e = cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) {
      cudaGetLastError();
      set_error("cannot map the inbox of rank %d into rank %d (cudaIpcOpenMemHandle: %s)", r, x->rank,
                cudaGetErrorString(e));
      return ORYX_E_CUDA;
    }
    x->peer_inbox[r] = reinterpret_cast<unsigned char*>(ptr);
    x->opened_ipc[r] = true;
  }
  x->connected = true;
  return ORYX_OK;
}

int32_t oryx_exchange_connect_local(oryx_exchange* const* all, int32_t world) {
  ORYX_ARG(all && world >= 1 && world <= ORYX_MAX_WORLD, "bad argument");
  for (int r = 0; r < world; r++) ORYX_ARG(all[r] && all[r]->world == world && all[r]->rank == r, "exchanges must be given in rank order");
  for (int r = 0; r < world; r++) {
    oryx_exchange* x = all[r];
    ORYX_SCOPED_DEVICE(x->device);
    for (int q = 0; q < world; q++) {
      if (q == r) continue;
      if (all[q]->device != x->device) {
        int can = 0;
        ORYX_CUDA(cudaDeviceCanAccessPeer(&can, x->device, all[q]->device));
        if (!can) {
          set_error("device %d cannot access device %d as a peer", x->device, all[q]->device);
          return ORYX_E_UNSUPPORTED;
        }
        cudaError_t e = cudaDeviceEnablePeerAccess(all[q]->device, 0);
        if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) {
          set_error("cudaDeviceEnablePeerAccess(%d -> %d): %s", x->device, all[q]->device, cudaGetErrorString(e));
          cudaGetLastError();
          return ORYX_E_CUDA;
        }
        cudaGetLastError();
      }
      x->peer_inbox[q] = all[q]->inbox;
    }
    x->connected = true;
  }
  return ORYX_OK;
}

int32_t oryx_exchange_destroy(oryx_exchange* x) {
  if (!x) return ORYX_OK;
  ScopedDevice scoped(x->device);
  cudaDeviceSynchronize();
  for (int r = 0; r < x->world; r++)
    if (x->opened_ipc[r]) cudaIpcCloseMemHandle(x->peer_inbox[r]);
  cudaFree(x->inbox);
  cudaFree(x->d_local);
  cudaFree(x->d_done);
  cudaFree(x->d_gram_status);
  cudaGetLastError();
  delete x;
  return ORYX_OK;
}

int32_t oryx_topn_sharded_device(oryx_model* mh, oryx_exchange* x, int32_t kind, const float* d_query,
                                 int32_t how_many, int64_t row_stride, int64_t* d_out_rows, double* d_out_scores,
                                 int32_t* d_out_n, void* stream) {
  ORYX_ARG(mh && x && d_query && d_out_rows && d_out_scores && d_out_n, "null argument");
  ORYX_ARG(x->connected, "exchange is not connected to its peers yet");
  ORYX_ARG(how_many >= 1 && how_many <= x->max_how_many, "how_many exceeds the exchange's max_how_many");
  ORYX_ARG(row_stride >= 0, "row_stride must be >= 0");
  ORYX_SCOPED_DEVICE(x->device);
  static const unsigned long long timeout_ns = [] {
    const char* e = std::getenv("ORYX_EXCHANGE_TIMEOUT_MS");
    long long ms = e ? std::atoll(e) : 20000;
    if (ms < 1) ms = 1;
    return (unsigned long long)ms * 1000000ull;
  }();
  const unsigned long long seq = ++x->seq;
  const int parity = (int)(seq & 1);
  // local scan + finishing kernel into this sequence's local record (scores | rows | count)
  unsigned char* rec = x->d_local + (size_t)parity * x->rec_bytes;
  // the finishing kernel may reuse this buffer only after the exchange of sequence seq-2 copied it out
  int32_t rc = topn_device_impl(reinterpret_cast<Model*>(mh), kind, d_query, how_many,
                                reinterpret_cast<int32_t*>(rec + (size_t)how_many * 8), reinterpret_cast<double*>(rec),
                                reinterpret_cast<int32_t*>(rec + (size_t)how_many * 12),
                                reinterpret_cast<cudaStream_t>(stream), x->d_done + 1, seq >= 2 ? seq - 2 : 0);
  if (rc != ORYX_OK) return rc;
  XParams p{};
  p.local = rec;
  for (int r = 0; r < x->world; r++) p.peers[r] = x->peer_inbox[r];
  p.inbox = x->inbox;
  p.done = x->d_done;
  p.seq = seq;
  p.timeout_ns = timeout_ns;
  // every rank must lay its slots out identically: the slot size comes from max_how_many, the payload
  // layout from this call's how_many
  p.rec_bytes = x->rec_bytes;
  p.world = x->world;
  p.rank = x->rank;
  p.how_many = how_many;
  p.parity = parity;
  p.row_stride = row_stride;
  p.out_rows = d_out_rows;
  p.out_scores = d_out_scores;
  p.out_n = d_out_n;
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(1);
  cfg.blockDim = dim3(kExchangeThreads);
  cfg.dynamicSmemBytes = 0;
  cfg.stream = reinterpret_cast<cudaStream_t>(stream);
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  ORYX_CUDA(cudaLaunchKernelEx(&cfg, exchange_merge_kernel, p));
  reinterpret_cast<Model*>(mh)->launches++;
  return ORYX_OK;
}

int32_t oryx_gram_allreduce_device(oryx_exchange* x, double* d_gram, int32_t features, int32_t* d_status, void* stream) {
  ORYX_ARG(x && d_gram && d_status, "null argument");
  ORYX_ARG(x->connected, "exchange is not connected to its peers yet");
  ORYX_ARG(features >= 1 && features <= kGramMaxFeatures, "features beyond the peer-memory Gram all-reduce (256): use NCCL");
  ORYX_SCOPED_DEVICE(x->device);
  static const unsigned long long timeout_ns = [] {
    const char* e = std::getenv("ORYX_EXCHANGE_TIMEOUT_MS");
    long long ms = e ? std::atoll(e) : 20000;
    if (ms < 1) ms = 1;
    return (unsigned long long)ms * 1000000ull;
  }();
  std::lock_guard<std::mutex> lk(x->gram_mu);
  GParams p{};
  p.local = d_gram;
  for (int r = 0; r < x->world; r++) p.peers[r] = x->peer_inbox[r] + x->gram_off;
  p.seq = ++x->gram_seq;
  p.timeout_ns = timeout_ns;
  p.world = x->world;
  p.rank = x->rank;
  p.parity = (int)(p.seq & 1);
  p.n_elems = features * features;
  const int n_slices = std::max(1, std::min(kGramMaxSlices, (p.n_elems + 255) / 256));
  p.per_slice = (p.n_elems + n_slices - 1) / n_slices;
  p.status = d_status;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  ORYX_CUDA(cudaMemsetAsync(d_status, 0, sizeof(int32_t), st));
  gram_allreduce_kernel<<<n_slices, kGramThreads, 0, st>>>(p);
  ORYX_CUDA(cudaGetLastError());
  return ORYX_OK;
}

int32_t oryx_gram_sharded(oryx_model* mh, oryx_exchange* x, int32_t which, double* out, int64_t* out_rows_used) {
  ORYX_ARG(mh && x && out, "null argument");
  Model* m = reinterpret_cast<Model*>(mh);
  ORYX_ARG(m->device == x->device, "model and exchange live on different devices");
  const int F = m->cfg.features;
  // this rank's partial (tensor-core kernel and its error check included); a copy stays in m->d_gram
  ORYX_TRY(oryx_gram(mh, which, out, out_rows_used));
  ORYX_SCOPED_DEVICE(x->device);
  QueryCtx* c = nullptr;
  ORYX_TRY(ctx_acquire(m, &c));
  int32_t rc = ORYX_OK;
  {
    std::lock_guard<std::mutex> gram_lock(m->gram_mu);  // m->d_gram is this model's Gram buffer
    cudaStream_t st = c->stream;
    int32_t status = 0;
    // (an empty shard's Gram is all zeros: the reference skips null partitions, ASM:365-366)
    rc = oryx_gram_allreduce_device(x, m->d_gram, F, x->d_gram_status, st);
    if (rc == ORYX_OK &&
        (cudaMemcpyAsync(out, m->d_gram, (size_t)F * F * sizeof(double), cudaMemcpyDeviceToHost, st) != cudaSuccess ||
         cudaMemcpyAsync(&status, x->d_gram_status, sizeof(status), cudaMemcpyDeviceToHost, st) != cudaSuccess ||
         cudaStreamSynchronize(st) != cudaSuccess)) {
      set_error("Gram all-reduce failed: %s", cudaGetErrorString(cudaGetLastError()));
      rc = ORYX_E_CUDA;
    }
    if (rc == ORYX_OK && status != 0) {
      set_error("Gram all-reduce: a peer did not arrive within ORYX_EXCHANGE_TIMEOUT_MS");
      rc = ORYX_E_CUDA;
    }
  }
  if (rc != ORYX_OK) cudaStreamSynchronize(c->stream);
  ctx_release(m, c);
  return rc;
}

}  // extern "C"
