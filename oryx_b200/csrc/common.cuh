// common.cuh -- shared declarations of the oryx_b200 native library (sm_100a only).
#pragma once

#include <cuda_runtime.h>

#include <atomic>
#include <condition_variable>
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <memory>
#include <mutex>
#include <shared_mutex>
#include <vector>

#include "oryx_b200.h"

namespace oryx {

// ---- error plumbing: status codes + thread-local message, no exceptions across the ABI.
void set_error(const char* fmt, ...);
const char* last_error();

#define ORYX_CUDA(expr)                                                                     \
  do {                                                                                      \
    cudaError_t e__ = (expr);                                                               \
    if (e__ != cudaSuccess) {                                                               \
      ::oryx::set_error("CUDA: %s -> %s (%s:%d)", #expr, cudaGetErrorString(e__), __FILE__, \
                        __LINE__);                                                          \
      return ORYX_E_CUDA;                                                                   \
    }                                                                                       \
  } while (0)

#define ORYX_ARG(cond, msg)                                           \
  do {                                                                \
    if (!(cond)) {                                                    \
      ::oryx::set_error("invalid argument: %s (%s)", msg, #cond);     \
      return ORYX_E_INVALID;                                          \
    }                                                                 \
  } while (0)

#define ORYX_TRY(expr)             \
  do {                             \
    int32_t s__ = (expr);          \
    if (s__ != ORYX_OK) return s__; \
  } while (0)

inline int round_up(int v, int m) { return (v + m - 1) / m * m; }

// ---- device-side helpers
#ifdef __CUDACC__
// 128-bit streaming load through the read-only path, no L1 allocation (Y is read once per scan).
__device__ __forceinline__ float4 ldg_stream(const float4* p) {
  float4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
               : "l"(p));
  return r;
}

// Order-preserving map float -> uint32 (larger float <=> larger key).  Key 0 is reserved as the
// "empty" sentinel: it corresponds to a negative NaN, which is never admitted to a top-N list.
__device__ __forceinline__ uint32_t float_key(float f) {
  uint32_t b = __float_as_uint(f);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float key_float(uint32_t k) {
  uint32_t b = (k & 0x80000000u) ? (k & 0x7fffffffu) : ~k;
  return __uint_as_float(b);
}

// The reference's VectorMath.dot arithmetic, exactly: binary32 product, widened, binary64 add, in
// index order (reference VectorMath.java:38-45).  Intrinsics forbid FMA contraction.
__device__ __forceinline__ double exact_step(double acc, float a, float b) {
  return __dadd_rn(acc, (double)__fmul_rn(a, b));
}

// The same value -- VectorMath.dot(a, b) over n terms, bit for bit -- computed by a whole warp WITHOUT walking
// the chain, whenever the chain provably never rounds.  (A dependent binary64 add per term costs ~45 ns on this
// part: 12 us per 256-feature row in the finishing kernel, most of its run time.)
//   t_j = fl32(a_j b_j) = +-m_j 2^(e_j - 150), m_j < 2^24: lane l takes the contiguous terms [l c, (l+1) c),
//   scales them to integers v_j = +-m_j 2^(e_j - e_min) (e_min: smallest exponent of a non-zero term) and the warp
//   forms every prefix sum P_j = v_0 + ... + v_j in int64, exactly (refused when the exponents spread too far for
//   that).  If every P_j fits a binary64 significand (bit length minus trailing zeros <= 53), then by induction
//   the reference's running sum equals P_j 2^(e_min - 150) after every step -- no addition ever rounds -- and the
//   result is (double)P_{n-1} scaled by a power of two.  For Gaussian data all but ~1 row in 10^4 qualify.
// Returns false (all lanes) when it cannot tell: NaN / infinite terms, too wide a spread, a prefix that would
// round.  The caller then walks the chain with exact_step.  `a`, `b`: any memory space; n <= 2048.
__device__ __forceinline__ bool warp_exact_dot(const float* __restrict__ a, const float* __restrict__ b, int n, int lane,
                                               double* out) {
  const int c = (n + 31) >> 5;
  const int j0 = lane * c, j1 = min(n, j0 + c);
  // pass 1: exponent range of the non-zero terms
  int emin = 1 << 20, emax = 0;
  bool bad = n > 2048;
  for (int j = j0; j < j1; j++) {
    const uint32_t bits = __float_as_uint(__fmul_rn(a[j], b[j]));
    const int ex = (int)((bits >> 23) & 255u);
    if (ex == 255) bad = true;
    if ((bits & 0x7fffffffu) != 0u) {
      const int e = ex == 0 ? 1 : ex;
      emin = min(emin, e);
      emax = max(emax, e);
    }
  }
  emin = __reduce_min_sync(0xffffffffu, emin);
  emax = __reduce_max_sync(0xffffffffu, emax);
  if (__any_sync(0xffffffffu, bad)) return false;
  if (emax == 0) {  // every term is zero: 0.0 + (+-0.0) + ... = +0.0
    *out = 0.0;
    return true;
  }
  // int64 headroom: |v_j| < 2^(24 + spread), n of them
  if (emax - emin > 62 - 24 - (32 - __clz(n))) return false;
  auto term = [&](int jj) -> long long {
    const uint32_t bits = __float_as_uint(__fmul_rn(a[jj], b[jj]));
    if ((bits & 0x7fffffffu) == 0u) return 0;
    const int ex = (int)((bits >> 23) & 255u);
    const long long m = (long long)((bits & 0x7fffffu) | (ex == 0 ? 0u : 0x800000u));
    const long long v = m << ((ex == 0 ? 1 : ex) - emin);
    return (bits & 0x80000000u) ? -v : v;
  };
  // pass 2: lane totals and their warp scan; pass 3: every prefix must fit a binary64 significand
  long long run = 0;
  for (int j = j0; j < j1; j++) run += term(j);
  long long incl = run;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const long long t = __shfl_up_sync(0xffffffffu, incl, d);
    if (lane >= d) incl += t;
  }
  long long p = incl - run;  // sum of the lanes in front
  bool ok = true;
  for (int j = j0; j < j1; j++) {
    p += term(j);
    const unsigned long long m = (unsigned long long)(p < 0 ? -p : p);
    ok = ok && (p == 0 || (64 - __clzll((long long)m)) - (__ffsll((long long)m) - 1) <= 53);
  }
  if (!__all_sync(0xffffffffu, ok)) return false;
  const long long total = __shfl_sync(0xffffffffu, incl, 31);
  // 2^(emin - 150) >= 2^-149: a normal binary64; the product is exact
  *out = (double)total * __longlong_as_double((long long)(1023 + emin - 150) << 52);
  return true;
}

// The same test and result for terms the lanes already hold in registers: lane l owns the cnt <= 8 consecutive terms
// t[0..cnt) of a chain of n <= 256 terms (lane l's terms come before lane l + 1's).  Used where the row is loaded once
// and needed again (fold_prep_kernel).
__device__ __forceinline__ bool warp_exact_sum_regs(const float (&t)[8], int cnt, int n, int lane, double* out) {
  int emin = 1 << 20, emax = 0;
  bool bad = false;
#pragma unroll
  for (int i = 0; i < 8; i++) {
    if (i < cnt) {
      const uint32_t bits = __float_as_uint(t[i]);
      const int ex = (int)((bits >> 23) & 255u);
      if (ex == 255) bad = true;
      if ((bits & 0x7fffffffu) != 0u) {
        const int e = ex == 0 ? 1 : ex;
        emin = min(emin, e);
        emax = max(emax, e);
      }
    }
  }
  emin = __reduce_min_sync(0xffffffffu, emin);
  emax = __reduce_max_sync(0xffffffffu, emax);
  if (__any_sync(0xffffffffu, bad)) return false;
  if (emax == 0) {
    *out = 0.0;
    return true;
  }
  if (emax - emin > 62 - 24 - (32 - __clz(n))) return false;
  long long v[8];
  long long run = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) {
    v[i] = 0;
    if (i < cnt) {
      const uint32_t bits = __float_as_uint(t[i]);
      if ((bits & 0x7fffffffu) != 0u) {
        const int ex = (int)((bits >> 23) & 255u);
        const long long m = (long long)((bits & 0x7fffffu) | (ex == 0 ? 0u : 0x800000u));
        const long long x = m << ((ex == 0 ? 1 : ex) - emin);
        v[i] = (bits & 0x80000000u) ? -x : x;
      }
    }
    run += v[i];
  }
  long long incl = run;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const long long up = __shfl_up_sync(0xffffffffu, incl, d);
    if (lane >= d) incl += up;
  }
  long long p = incl - run;
  bool ok = true;
#pragma unroll
  for (int i = 0; i < 8; i++) {
    if (i < cnt) {
      p += v[i];
      const unsigned long long m = (unsigned long long)(p < 0 ? -p : p);
      ok = ok && (p == 0 || (64 - __clzll((long long)m)) - (__ffsll((long long)m) - 1) <= 53);
    }
  }
  if (!__all_sync(0xffffffffu, ok)) return false;
  const long long total = __shfl_sync(0xffffffffu, incl, 31);
  *out = (double)total * __longlong_as_double((long long)(1023 + emin - 150) << 52);
  return true;
}

// The same for a chain of n <= 64 terms held by a GROUP of eight lanes (lane g of the group owns the cnt <= 8
// consecutive terms from 8-lane-aligned position g * ceil(n / 8) on): four short chains per warp, so that short rows do
// not leave three quarters of a warp idle.  All 32 lanes must call it; results are per group.
__device__ __forceinline__ bool group8_exact_sum_regs(const float (&t)[8], int cnt, int n, int lane, double* out) {
  const int g = lane & 7;
  const unsigned gmask = 0xffu << (lane & 24);
  int emin = 1 << 20, emax = 0;
  bool bad = false;
#pragma unroll
  for (int i = 0; i < 8; i++) {
    if (i < cnt) {
      const uint32_t bits = __float_as_uint(t[i]);
      const int ex = (int)((bits >> 23) & 255u);
      if (ex == 255) bad = true;
      if ((bits & 0x7fffffffu) != 0u) {
        const int e = ex == 0 ? 1 : ex;
        emin = min(emin, e);
        emax = max(emax, e);
      }
    }
  }
#pragma unroll
  for (int o = 4; o > 0; o >>= 1) {
    emin = min(emin, __shfl_xor_sync(0xffffffffu, emin, o));
    emax = max(emax, __shfl_xor_sync(0xffffffffu, emax, o));
  }
  const bool any_bad = (__ballot_sync(0xffffffffu, bad) & gmask) != 0u;
  const bool all_zero = emax == 0;
  const bool too_wide = !all_zero && emax - emin > 62 - 24 - (32 - __clz(n));
  long long v[8];
  long long run = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) {
    v[i] = 0;
    if (i < cnt && !all_zero && !too_wide) {
      const uint32_t bits = __float_as_uint(t[i]);
      if ((bits & 0x7fffffffu) != 0u) {
        const int ex = (int)((bits >> 23) & 255u);
        const long long m = (long long)((bits & 0x7fffffu) | (ex == 0 ? 0u : 0x800000u));
        const long long x = m << ((ex == 0 ? 1 : ex) - emin);
        v[i] = (bits & 0x80000000u) ? -x : x;
      }
    }
    run += v[i];
  }
  long long incl = run;
#pragma unroll
  for (int d = 1; d < 8; d <<= 1) {
    const long long up = __shfl_up_sync(0xffffffffu, incl, d, 8);
    if (g >= d) incl += up;
  }
  long long p = incl - run;
  bool ok = true;
#pragma unroll
  for (int i = 0; i < 8; i++) {
    if (i < cnt) {
      p += v[i];
      const unsigned long long m = (unsigned long long)(p < 0 ? -p : p);
      ok = ok && (p == 0 || (64 - __clzll((long long)m)) - (__ffsll((long long)m) - 1) <= 53);
    }
  }
  const bool all_ok = (__ballot_sync(0xffffffffu, ok) & gmask) == gmask;
  const long long total = __shfl_sync(0xffffffffu, incl, 7, 8);
  if (any_bad || too_wide || !all_ok) return false;
  *out = all_zero ? 0.0 : (double)total * __longlong_as_double((long long)(1023 + emin - 150) << 52);
  return true;
}
#endif

// ---- the device-resident feature-vector store (one per X / Y).
//
// Layout in HBM:
//   vec       [cap_slots x fpad] fp32, row-major, fpad = features rounded up to 4 so every row is
//             16-byte aligned and can be read with 128-bit loads; padding floats are zero.
//   slot_row  [cap_slots] int32: caller's row id living in the slot, -1 = empty / tombstone.
//   slot_part [cap_slots] int32: LSH partition of the row (0 when the store is not hashed), -1 = empty.
//   block_part [cap_slots/256] int32: partition summary of each 256-slot block (see below).
// The host keeps row -> slot and a free list; removed slots (tombstones) are NaN-filled and reused.
struct Store {
  int features = 0;
  int fpad = 0;
  bool hashed = false;  // Y store with num_hashes > 0

  float* d_vec = nullptr;
  int32_t* d_slot_row = nullptr;
  int32_t* d_slot_part = nullptr;
  // per 256-slot block: the partition shared by all its live rows, -1 if it has no live row, -2 if
  // mixed.  Lets an LSH-filtered scan skip or stream whole blocks without per-row metadata.
  int32_t* d_block_part = nullptr;
  // Bit pattern of the largest |x| ever written into the store (non-negative floats order like their bit
  // patterns).  Only grows: removals and overwrites leave it in place, which keeps it an upper bound --
  // all the tensor-core Gram needs to pick its fixed-point scale (gram_tc.cu).  Written by the row
  // scatter kernel; NaN never enters it, +inf does (and sends the Gram to the floating-point kernels).
  unsigned int* d_absmax_bits = nullptr;
  int64_t cap_slots = 0;
  int64_t n_slots = 0;  // high-water mark
  int64_t live = 0;
  uint64_t version = 0;  // bumped by every mutation; lets derived buffers (the bf16 shadow) be reused

  std::vector<int32_t> row2slot;
  std::vector<int32_t> free_slots;
  // Device copy of row -> slot (-1 = absent): lets a query translate a device-resident list of known-item
  // ROWS into the slots the scan filters on (known.cu), and lets a peer process gather rows by row id
  // (peer_stores.cu).
  bool track_row_slot = false;
  int32_t* d_row_slot = nullptr;
  int64_t row_slot_cap = 0;
  // Bumped whenever d_vec or d_row_slot is reallocated; mirrored in Model::d_share_desc once the stores
  // have been exported to peer processes, so that a peer holding a stale mapping fails instead of reading it.
  unsigned long long layout_gen = 1;

  int64_t device_bytes() const { return cap_slots * ((int64_t)fpad * 4 + 8) + row_slot_cap * 4; }
};

// ---- known items (ALSServingModel.java:75-76 knownItems, :189-258): per user row, the item rows the
// user has interacted with.  The host keeps the authoritative sorted lists (set semantics of
// addKnownItems); a mirror of every list lives in one device pool so that a /recommend query takes
// its exclusion list from HBM instead of receiving it from the caller.
// Grow-only staging of the (single) updater thread of a model.
struct UpdateStaging {
  float* d_vec = nullptr; size_t vec_bytes = 0;
  int32_t* d_i32 = nullptr; size_t i32_count = 0;  // [3 x chunk]: dst, rows, parts
  int32_t* h_i32 = nullptr; size_t h_count = 0;    // pinned
  int64_t* d_known_dst = nullptr; int32_t* d_known_val = nullptr; size_t known_cap = 0;
};

struct KnownItems {
  std::vector<std::vector<int32_t>> lists;  // [user row] sorted unique item rows
  std::vector<int64_t> seg_off;             // [user row] start of the user's segment in the pool
  std::vector<int32_t> seg_cap;             // [user row] capacity of that segment (0 = none yet)
  int32_t* d_pool = nullptr;
  int64_t pool_cap = 0, pool_used = 0, pool_garbage = 0;
  int64_t total = 0;                        // sum of list sizes
};

// ---- per-query resources: a stream, device staging, scratch for the per-CTA candidate lists and
// pinned host staging.  A small pool of these serves concurrent callers.
struct QueryCtx {
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  float* d_query = nullptr;       // [fpad]
  int32_t* d_excl = nullptr;      // [excl_cap]
  int32_t excl_cap = 0;
  // Scan scratch.  When `scratch_sets` == 2 the launcher alternates between two sets so the finishing
  // kernel of one query can overlap the scan of the next (programmatic dependent launch).
  uint2* d_cand = nullptr;        // [scratch_sets x max_grid x max_k] (key, slot)
  uint32_t* d_list_max = nullptr; // [scratch_sets x max_grid] best key of each CTA's list
  unsigned int* d_counter = nullptr;  // [scratch_sets x 16] dynamic tile counters (64 B apart)
  int scratch_sets = 1;
  int flip = 0;
  // Finishing kernels of one context complete in order: kernel number n releases its dependents
  // (the scan of query n+1, which reuses the scratch set of query n-1) only after number n-1 is done.
  unsigned long long* d_finish_seq = nullptr;
  unsigned long long finish_seq = 0;
  int32_t* d_main_index = nullptr;
  // LSH-filtered scans: the candidate (or mixed) 256-slot blocks of the current query, built on the
  // device right after the query is hashed; [0] = count, [1..] = block ids.  Null in a context that
  // owns no memory (the scan then tests every block itself).  Two lists of worklist_cap entries alternate
  // per query (worklist_flip): the kernel that fills one zeroes the other's count.
  int32_t* d_worklist = nullptr;
  int64_t worklist_cap = 0;
  int worklist_flip = 0;
  unsigned int* d_worklist_ctl = nullptr;  // [0] CTAs of this context's list kernels that have completed, ever
  unsigned int worklist_seq = 0;           // ... and how many were launched (mod 2^32)
  int32_t* d_out_rows = nullptr;  // [max_k]
  double* d_out_scores = nullptr; // [max_k]
  int32_t* d_out_n = nullptr;
  // grow-only device scratch of the point-lookup entry points (/estimate, /similarityToItem, LSH index):
  // kept with the context so those calls do not go through cudaMalloc / cudaFree per request
  void* d_tmp[2] = {nullptr, nullptr};
  size_t tmp_bytes[2] = {0, 0};
  // pinned host staging
  float* h_query = nullptr;
  int32_t* h_excl = nullptr;
  int32_t* h_out_rows = nullptr;
  double* h_out_scores = nullptr;
  int32_t* h_out_n = nullptr;
};

struct BatchTimings {
  double convert_ms = 0, score_ms = 0, threshold_ms = 0, refine_ms = 0, flops = 0;
  double cand_mean = 0, cand_max = 0;  // candidate sub-blocks per query handed to the refine step
  double runs = 0, shadow_rebuilds = 0;  // batches answered by the tensor-core path / full conversions of Y
};

// Grow-only device scratch of the tensor-core batch path, kept across calls (one batch at a time).
struct BatchScratch {
  std::mutex mu;
  void* buf[14] = {nullptr};
  size_t bytes[14] = {0};
  cudaEvent_t ev[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};  // phase timing, created once
  uint64_t y_version = ~0ull;  // store version the bf16 shadow of Y was converted from
  int y_normalized = -1;       // ... and whether it holds y/||y|| (cosine) or y (dot)
  int64_t y_rows = -1;      // slots the shadow covers (== n_slots of the store while it is current)
  int64_t y_cap_rows = 0;   // rows the shadow buffer has room for (the store's slot capacity when it was built)
};

// Grow-only buffers, streams and events of one solver's fold-in pipeline (foldin.cu): allocated the first
// time a batch of that size runs, kept until the solver is destroyed, so a steady stream of micro-batches
// (ALSSpeedModelManager.buildUpdates every batch interval) never touches the device allocator.
struct FoldScratch {
  void* d[8] = {nullptr};
  size_t d_bytes[8] = {0};
  void* h[4] = {nullptr};  // pinned
  size_t h_bytes[4] = {0};
  cudaStream_t copy_stream = nullptr;
  cudaEvent_t ev_done[2] = {nullptr, nullptr};    // chunk's kernel finished (compute stream)
  cudaEvent_t ev_copied[2] = {nullptr, nullptr};  // chunk's results are in pinned host memory (copy stream)
  cudaEvent_t ev_k0[2] = {nullptr, nullptr}, ev_k1[2] = {nullptr, nullptr};  // around fold_in_kernel
  // last call: device time of fold_in_kernel summed over chunks, of the gathers, wall time, sizes
  double last_kernel_ms = 0, last_wall_ms = 0;
  int64_t last_n = 0;
  int last_chunks = 0;
  int gemm_ctas_per_sm = 0;  // occupancy of fold_gemm_async_kernel on this device (development aid)
};
void fold_scratch_free(FoldScratch* fs);

struct Coalescer;  // coalesce.cu
Coalescer* coalescer_create();
void coalescer_destroy(Coalescer* c);

struct Model {
  oryx_config cfg{};
  int device = 0;
  int num_sms = 0;
  int num_hashes = 0;
  int max_bits = 0;
  float* d_hash_vectors = nullptr;  // [num_hashes x fpad]
  std::vector<float> h_hash_vectors;
  Store store[2];
  KnownItems known;
  UpdateStaging staging;

  std::shared_mutex mu;  // queries shared, mutations exclusive
  cudaStream_t update_stream = nullptr;

  // query context pool
  std::mutex pool_mu;
  std::condition_variable pool_cv;
  std::vector<QueryCtx*> pool_free;
  std::vector<std::unique_ptr<QueryCtx>> pool_all;
  int pool_limit = 8;
  BatchScratch batch;
  BatchTimings last_batch;  // phase timings of the most recent tensor-core batch (oryx_debug_batch_timings)
  unsigned long long* d_debug_ts = nullptr;  // development aid, see oryx_debug_scan_timestamps
  unsigned long long* d_share_desc = nullptr;  // [2] layout_gen of X and Y, read by peer processes (peer_stores.cu)
  QueryCtx* async_ctx = nullptr;  // dedicated to oryx_topn_device (stream-ordered by the caller)
  // The stream-ordered entry points return with their kernels still queued on the CALLER's stream.  Each
  // records this event behind its launches; every mutator makes update_stream wait for it before touching
  // the store (async_fence), so an upsert / removal / compaction never runs under a queued scan.
  cudaEvent_t async_done = nullptr;
  std::atomic<bool> async_pending{false};
  Coalescer* coalescer = nullptr;  // request combiner of oryx_topn_coalesced

  // YtY / XtX (gram.cu): one persistent F x F binary64 output buffer plus whatever scratch the kernel needs,
  // allocated by the first oryx_gram call; one Gram at a time per model.
  std::mutex gram_mu;
  double* d_gram = nullptr;
  void* d_gram_scratch = nullptr;
  size_t gram_scratch_bytes = 0;
  std::atomic<double> last_gram_ms{0.0};
  std::atomic<int> last_gram_tc{0};  // 1 when the last Gram came from the tensor-core kernel

  std::atomic<int64_t> queries{0};
  std::atomic<int64_t> launches{0};
  std::atomic<double> last_query_ms{0.0};
};

constexpr int kBlockRows = 256;                    // slots per block_part entry
constexpr int kTopnMargin = 8;                      // extra candidates re-scored exactly
constexpr int kMaxK = ORYX_MAX_HOW_MANY + kTopnMargin;  // 1024
constexpr int kMaxGrid = 148 * 8;

// store.cu
int32_t store_init(Model* m, int which, int64_t capacity_hint);
void store_free(Store* s);
int32_t store_upsert(Model* m, int which, const int32_t* rows, const float* vecs, int32_t n,
                     int32_t* out_partitions);
int32_t store_upsert_device(Model* m, int which, int32_t first_row, const float* d_vecs, int64_t n);
int32_t store_remove(Model* m, int which, const int32_t* rows, int32_t n);
int32_t store_get(Model* m, int which, int32_t row, float* out, int32_t* found);
int32_t store_compact(Model* m, int which);

// known.cu
void known_free(KnownItems* k);
int32_t known_add(Model* m, const int32_t* user_rows, const int64_t* begin, const int32_t* item_rows, int32_t n_users);
int32_t known_retain(Model* m, const uint8_t* keep_users, int32_t n_user_rows, const uint8_t* keep_items,
                     int32_t n_item_rows);
int32_t known_counts(Model* m, int which, int32_t* out_counts, int32_t n_rows);
// Writes the sorted exclusion SLOTS of `user_row` (padded with INT32_MAX) into ctx->d_excl on `stream`;
// *n_excl = entries written.  Lists longer than the kernel's shared-memory sort fall back to a host
// translation of the authoritative list.
int32_t known_exclusions_device(Model* m, QueryCtx* ctx, int32_t user_row, int32_t* n_excl, cudaStream_t stream);
int32_t ensure_excl(QueryCtx* c, int32_t n);

// lsh.cu
int32_t lsh_plan(double sample_rate, int num_cores, int* hashes, int* bits);
// Hash n packed rows (row stride `stride` floats) on `stream`; writes int32 partition per row.
int32_t lsh_hash_rows_device(Model* m, const float* d_rows, int64_t n, int stride, int32_t* d_out,
                             cudaStream_t stream, int32_t* d_zero_counter = nullptr, bool early_start = false);
int32_t lsh_worklist_device(Model* m, const int32_t* d_main_index, int32_t* d_worklist, cudaStream_t stream);
int32_t lsh_query_worklist_device(Model* m, const float* d_query, int32_t* d_main_index, int32_t* d_worklist,
                                  int32_t* d_other_count, unsigned int* d_done_ctas, unsigned int* expected_total,
                                  cudaStream_t stream);
int32_t lsh_candidates_device(Model* m, const int32_t* d_main_index, int32_t* d_out, int32_t* d_out_n,
                              cudaStream_t stream);

// scan_topn.cu
struct TopnRequest {
  int kind;
  const float* d_query;  // [fpad] device
  int how_many;
  const int32_t* d_excl;  // sorted SLOT ids, device
  int n_excl;
  int32_t* d_out_rows;
  double* d_out_scores;
  int32_t* d_out_n;
  // optional: the finishing kernel holds its output writes until *gate >= gate_value (exchange.cu)
  const unsigned long long* gate = nullptr;
  unsigned long long gate_value = 0;
};
int32_t scan_topn_launch(Model* m, QueryCtx* ctx, const TopnRequest& rq, cudaStream_t stream);
// topn_deep.cu: how_many beyond the fused scan's limit
size_t deep_scratch_bytes(int64_t n_slots, int how_many);
int32_t topn_deep_launch(Model* m, QueryCtx* ctx, int kind, const float* d_query, int how_many, const int32_t* d_excl,
                         int n_excl, void* d_scratch, int32_t* d_out_rows, double* d_out_scores, int32_t* d_out_n,
                         cudaStream_t stream);
int32_t scores_all_launch(Model* m, int kind, const float* d_query, int32_t n_rows, double* d_out,
                          cudaStream_t stream);

int32_t scores_rows_launch(Model* m, int kind, const float* d_query, const int32_t* d_slots, int n,
                           double* d_out, cudaStream_t stream);

int32_t pair_scores_launch(Model* m, const int32_t* d_x_slots, const int32_t* d_y_slots, int64_t n, int wide_products,
                           double* d_out, cudaStream_t stream);

int32_t merge_topn_launch(const double* d_scores, const int32_t* d_rows, const int32_t* d_n, int n_lists,
                          int how_many, int64_t row_stride, int64_t* d_out_rows, double* d_out_scores,
                          int32_t* d_out_n, cudaStream_t stream);

int32_t merge_topn_batch_launch(const double* d_scores, const int32_t* d_rows, const int32_t* d_n, int n_lists, int nq,
                                int how_many, int64_t row_stride, int64_t* d_out_rows, double* d_out_scores,
                                int32_t* d_out_n, cudaStream_t stream);

int32_t merge_topn_packed_launch(const void* d_records, int n_lists, int record_bytes, int how_many,
                                 int64_t row_stride, int64_t* d_out_rows, double* d_out_scores, int32_t* d_out_n,
                                 cudaStream_t stream);

// batch_topn.cu
bool batch_topn_supported(Model* m, int nq);
// The conditions of batch_topn_supported that do not depend on the page: false means no page can ever ride
// on the tensor-core path for this model as it stands (the request combiner then stays out of the way).
bool batch_topn_possible(Model* m);
// Set by a caller that will keep sending small pages (the request combiner): lets a page of fewer than 32
// queries (re)build the bf16 shadow instead of falling back to one scan per query.
extern thread_local bool tl_batch_small_pages_ok;
bool batch_shadow_rows_changed(Model* m, const int32_t* d_slots, int64_t n, uint64_t version_before, cudaStream_t st);
// `queries` is a HOST [nq x features] array -- or, when `query_slots` is given, ignored: query i is then
// the row in slot query_slots[i] of store `query_store` (a negative slot gives a zero query).
int32_t batch_topn_run(Model* m, int kind, const float* queries, int nq, int how_many, const int32_t* excl_offsets,
                       const int32_t* excl_slots, int32_t* out_rows, double* out_scores, int32_t* out_n,
                       cudaStream_t st, BatchTimings* timings, const int32_t* query_slots = nullptr,
                       int query_store = ORYX_X);

// gram.cu
int32_t gram_launch(Model* m, int which, double* d_out, cudaStream_t stream, bool allow_tc = true,
                    float* tc_absmax = nullptr);
// gram_tc.cu: expected largest entry error of the tensor-core Gram relative to its largest entry
double gram_tc_error_estimate(float absmax, double max_diag);

// cudaFuncSetAttribute is per device: raises the kernel's dynamic shared-memory limit on the CURRENT
// device if it has not been raised that far yet (api.cu).
int32_t ensure_max_dynamic_smem(const void* kernel, size_t bytes);

// oryx_topn_device with an optional output gate (see TopnRequest); api.cu
int32_t topn_device_impl(Model* m, int kind, const float* d_query, int how_many, int32_t* d_out_rows,
                         double* d_out_scores, int32_t* d_out_n, cudaStream_t stream,
                         const unsigned long long* gate, unsigned long long gate_value);

// Called after a stream-ordered entry point queued its kernels on `stream` (still under the shared lock).
int32_t async_mark(Model* m, cudaStream_t stream);
// Called by every mutator right after it took the exclusive lock.
int32_t async_fence(Model* m);

// pool
int32_t ctx_acquire(Model* m, QueryCtx** out);
void ctx_release(Model* m, QueryCtx* ctx);
void ctx_destroy(QueryCtx* c);

}  // namespace oryx
