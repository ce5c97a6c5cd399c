// scan_topn.cu -- single-query scoring scan with a fused per-CTA top-N candidate list, and the
// finishing kernel that merges the lists and re-scores exactly (kernels K1/K2/K6/K7 of SURVEY.md
// section 2.3).
//
// Replaces the hot loop of ALSServingModel.topN (reference: app/oryx-app-serving/src/main/java/com/
// cloudera/oryx/app/serving/als/model/ALSServingModel.java:289-335): per candidate partition a
// TopNConsumer (TopNConsumer.java:56-74) scores every vector with DotsFunction or
// CosineAverageFunction (VectorMath.dot, VectorMath.java:38-45) and keeps the howMany best; the
// per-partition lists are concatenated, sorted descending and cut (ALSServingModel.java:329-334).
//
// Device algorithm (two launches per query):
//   scan_topn_kernel -- HBM-bound.
//     * A group of G lanes owns one row and reads it with coalesced 128-bit streaming loads; fp32
//       FMA partial sums are combined with a transposed xor-shuffle reduction.  Rows of non-candidate
//       LSH partitions are skipped before their loads are issued (candidate(p) <=>
//       popcount(p ^ mainIndex) <= maxBitsDiffering, exactly the set
//       LocalitySensitiveHash.getCandidateIndices enumerates).
//     * Work distribution: SMs do not stream at the same speed (distance to the memory partitions),
//       and a static split left the fastest CTAs idle for the last ~20 % of the launch.  Three
//       quarters of the tiles are therefore assigned statically (interleaved by warp), the rest is
//       handed out in small chunks from a global counter that each warp fetches one chunk ahead.
//     * Scores above the CTA's current k-th best (k = howMany + 8) take a rare slow path: the
//       allowed-predicate (slot not in the exclusion list) and a lock-free replace-the-minimum in a
//       shared-memory candidate array.  Strict '>' admission, NaN never enters
//       (TopNConsumer.java:60-64); tombstoned slots are NaN-filled so they never qualify.
//     * Each CTA publishes its k candidates (order-preserving key, slot) to global scratch.
//   topn_finish_kernel -- one 1024-thread CTA.
//     * Selects the k best of all lists (for k <= 32: the k-th largest list maximum prunes all but
//       ~k lists, the survivors are rank-sorted in shared memory; otherwise a radix select), re-scores those k rows with the reference's exact arithmetic (binary32 product,
//       binary64 accumulate, index order; cosine adds the exact norm and a binary64 sqrt/divide),
//       ranks by (exact score desc, row asc) and writes howMany results.
// Returned scores are therefore bit-identical to the Java path's; the fp32 scan only decides
// which k rows get re-scored.
#include <algorithm>
#include <cstdlib>

#include "common.cuh"

namespace oryx {

namespace {

constexpr int kScanThreads = 256;
constexpr int kFinishThreads = 1024;
constexpr int kDynChunkMax = 16;       // largest guided chunk, in tiles
constexpr int kExclSmemMax = 4096;     // exclusion entries staged in shared memory (16 KB)
constexpr int kRowStageFloats = 16384; // floats of row staging in the finishing CTA (64 KB)

struct ScanParams {
  const float4* y;
  const int32_t* slot_row;
  const int32_t* slot_part;
  const int32_t* block_part;  // per 256-slot block: shared partition, -1 empty, -2 mixed
  const int32_t* worklist;    // LSH scans: [0] = number of candidate blocks, [1..] their ids (may be null)
  const float* q;  // [fpad]
  const int32_t* excl;  // sorted SLOT ids the allowed-predicate rejects
  int n_excl;
  int excl_in_smem;     // the list was small enough to stage in shared memory
  int64_t n_slots;
  int features;
  int f4;  // fpad / 4
  int k;   // candidates per list = how_many + margin
  int how_many;
  int max_bits;
  const int32_t* main_index;
  uint2* cand;            // scratch [n_lists x k]
  uint32_t* list_max;     // scratch [n_lists]: largest key of each published list (0 = empty list)
  unsigned int* counter;  // dynamic tile counter; zero at launch, reset by the finishing kernel
  const unsigned int* worklist_flag;  // CTAs of the context's list kernels completed so far ...
  unsigned int worklist_seq;          // ... and the total once this query's list is complete
  int32_t* out_rows;
  double* out_scores;
  int32_t* out_n;
  unsigned long long* debug_ts;  // optional [8] timestamps (ns): see oryx_debug_scan_timestamps
  // Optional gate: the finishing kernel does not overwrite the output record before *gate >= gate_value
  // (the sharded path's exchange kernel may still have to read the record of an earlier query).
  const unsigned long long* gate;
  unsigned long long gate_value;
  // Optional in-order token of the context's finishing kernels: this is kernel number order_value;
  // it releases its dependents only once *order >= order_value - 1 and stores order_value at its end.
  unsigned long long* order;
  unsigned long long order_value;
};

// Programmatic dependent launch: let the next kernel in the stream start / wait for the previous one.
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait_primary() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

__device__ __forceinline__ bool excluded(const int32_t* excl, int n, int32_t slot) {
  int lo = 0, hi = n - 1;
  while (lo <= hi) {
    int mid = (lo + hi) >> 1;
    int32_t v = excl[mid];
    if (v == slot) return true;
    if (v < slot) lo = mid + 1; else hi = mid - 1;
  }
  return false;
}

// ------------------------------------------------------------------------------------------------
// The CTA's candidate list: k packed words (score key << 32 | slot) in shared memory, unordered,
// all starting as the empty sentinel 0.  An offer replaces the current minimum if the new score is
// strictly greater (TopNConsumer.java:60-64: strict '>', so equal scores never displace), using a
// 64-bit compare-and-swap -- no lock, so the warps of a CTA never queue behind one another, and the
// whole warp takes part in the search for the minimum.  The published threshold only ever rises
// (atomicMax on the key); a stale, lower value merely causes a redundant offer.
//
// This slow path runs for ~k*ln(rows per CTA / k) rows per CTA and touches shared memory only: the
// memory system is saturated by the scan, so one dependent global load here would stall the warp
// for microseconds.
__device__ __forceinline__ void warp_min_entry(const unsigned long long* arr, int k, int lane,
                                               unsigned long long& best, int& best_idx) {
  best = ~0ull;
  best_idx = 0;
  for (int i = lane; i < k; i += 32) {
    unsigned long long v = *(const volatile unsigned long long*)(arr + i);
    if (v < best) { best = v; best_idx = i; }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    unsigned long long ov = __shfl_xor_sync(0xffffffffu, best, o);
    int oi = __shfl_xor_sync(0xffffffffu, best_idx, o);
    if (ov < best || (ov == best && oi < best_idx)) { best = ov; best_idx = oi; }
  }
}

__device__ __noinline__ void warp_offer(const ScanParams& p, unsigned long long* arr, const int32_t* s_excl,
                                        unsigned int* s_thr_key, uint32_t key, int32_t slot, int lane) {
  // TopNConsumer.java:57 allowedPredicate: the row must not be in the exclusion list.  Excluded rows
  // are translated to slots by the host, so the test needs no slot -> row lookup.
  if (p.n_excl > 0 && excluded(s_excl, p.n_excl, slot)) return;
  const unsigned long long x = ((unsigned long long)key << 32) | (unsigned long long)(uint32_t)slot;
  unsigned long long best;
  int best_idx;
  for (;;) {
    warp_min_entry(arr, p.k, lane, best, best_idx);
    if (key <= (uint32_t)(best >> 32)) {  // not strictly above the current minimum: rejected
      if (lane == 0) atomicMax(s_thr_key, (uint32_t)(best >> 32));
      return;
    }
    unsigned long long old = 0ull;
    if (lane == 0) old = atomicCAS(arr + best_idx, best, x);
    old = __shfl_sync(0xffffffffu, old, 0);
    if (old == best) break;  // replaced the minimum; otherwise another warp got there first: retry
  }
  warp_min_entry(arr, p.k, lane, best, best_idx);
  if (lane == 0) atomicMax(s_thr_key, (uint32_t)(best >> 32));
}

// Scores of one warp tile.  G lanes share a row, each lane holds NV float4 of U rows (rows
// row0 + u*RPI).  `vals` are per-lane partial sums; the reduction is "transposed": log2(U) exchange
// steps hand each lane one row to own while halving the lanes per row, then log2(G/U) plain xor
// steps finish -- log2(G) shuffles for U rows instead of U*log2(G).  On return lane gl < U holds
// the total of row u = owned_row<U>(gl) in vals[0].
template <int G, int U>
__device__ __forceinline__ void reduce_rows(float (&vals)[U], int gl) {
  static_assert(U <= G, "rows per group must not exceed lanes per group");
  int bit = 1;
#pragma unroll
  for (int half = U / 2; half >= 1; half /= 2) {
    const bool upper = (gl & bit) != 0;
#pragma unroll
    for (int j = 0; j < half; j++) {
      const float keep = upper ? vals[j + half] : vals[j];
      const float send = upper ? vals[j] : vals[j + half];
      vals[j] = keep + __shfl_xor_sync(0xffffffffu, send, bit);
    }
    bit <<= 1;
  }
#pragma unroll
  for (int o = U; o < G; o <<= 1) vals[0] += __shfl_xor_sync(0xffffffffu, vals[0], o);
}

// Which of the U rows of its group lane `gl` (< U) owns after reduce_rows.
template <int U>
__device__ __forceinline__ int owned_row(int gl) {
  int u = 0, bit = 1;
#pragma unroll
  for (int half = U / 2; half >= 1; half /= 2) {
    if (gl & bit) u += half;
    bit <<= 1;
  }
  return u;
}

// G lanes per row, NV float4 per lane, U rows per group in flight.
template <int G, int NV, int U, int KIND, bool LSH>
__global__ void __launch_bounds__(kScanThreads, (NV * U <= 8) ? 4 : 2)
scan_topn_kernel(const __grid_constant__ ScanParams p) {
  constexpr int T = kScanThreads;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ unsigned int s_thr_key;

  unsigned long long* cand_arr = reinterpret_cast<unsigned long long*>(smem_raw);  // [k]
  int32_t* s_excl_buf = reinterpret_cast<int32_t*>(smem_raw + (size_t)p.k * 8);    // [n_excl] if staged

  const int tid = threadIdx.x;
  const int lane = tid & 31;
  const int gl = lane % G;   // lane within the row group
  const int grp = lane / G;  // which row of the warp's row set
  constexpr int RPI = 32 / G;       // rows covered by one load instruction of a warp
  constexpr int RPT = RPI * U;      // rows per warp tile

  // A scan released early must still see what its immediate predecessors wrote: the exclusion list
  // (a device-resident known-items list is prepared by a kernel) and, below, the LSH index / work list.
  // An owning context's LSH scan (p.worklist != nullptr) does NOT wait here: griddepcontrol.wait would wait for
  // the whole chain in front of it -- the previous query's finishing kernel included (measured, round 2: 29 us
  // between two pipelined filtered scans = that kernel's duration).  Its inputs are complete anyway: the kernel
  // that hashes the query and builds the list never releases its dependents early, so this grid is launched
  // only after every thread of it has fenced its writes and exited.  Every CTA of the list kernel counts itself
  // with a release reduction; reading the expected total here (acquire) makes all of that grid's writes visible,
  // and a shortfall -- never observed -- sends the CTA to griddepcontrol.wait.  List and index are read with
  // ld.global.cg.
  __shared__ unsigned int s_list_seen;
  const bool guarded = LSH && p.worklist != nullptr && p.n_excl == 0;
  if (!guarded && (LSH || p.n_excl > 0)) pdl_wait_primary();
  if (guarded && tid == 0) {  // one acquire load per CTA, in flight while the CTA sets itself up
    unsigned int seen;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(seen) : "l"(p.worklist_flag) : "memory");
    s_list_seen = seen;
  }
  const int32_t* excl = p.excl_in_smem ? s_excl_buf : p.excl;
  if (p.excl_in_smem) {
    for (int i = tid; i < p.n_excl; i += T) s_excl_buf[i] = __ldg(p.excl + i);
  }
  if (p.debug_ts && tid == 0) atomicMin(p.debug_ts + 0, global_ns());  // earliest CTA start
  for (int i = tid; i < p.k; i += T) cand_arr[i] = 0ull;  // empty sentinels
  if (tid == 0) s_thr_key = 0u;
  __syncthreads();
  if (guarded && s_list_seen != p.worklist_seq) {  // not expected; the formal guarantee
    if (p.debug_ts && tid == 0) atomicAdd(p.debug_ts + 11, 1ull);
    pdl_wait_primary();
  }

  const int f4 = p.f4;
  float4 qv[NV];
#pragma unroll
  for (int v = 0; v < NV; v++) {
    int idx = gl + v * G;
    qv[v] = idx < f4 ? __ldg(reinterpret_cast<const float4*>(p.q) + idx) : make_float4(0.f, 0.f, 0.f, 0.f);
  }
  const int main_idx = LSH ? __ldcg(p.main_index) : 0;
  const bool full_width = (f4 == G * NV);  // every lane's every float4 is inside the row

  const int warps_total = (int)gridDim.x * (T / 32);
  const int gw = (int)blockIdx.x * (T / 32) + (tid >> 5);
  const int n_slots = (int)p.n_slots;  // slots are int32 row handles
  const int n_tiles = (n_slots + RPT - 1) / RPT;
  const int n_full_tiles = n_slots / RPT;
  const int my_u = owned_row<U>(gl);

  // Offer the rows whose reduced score beats the CTA threshold (slow path, rare).  The whole warp
  // takes part in each offer.
  auto offer = [&](float score, bool hit, int tile) {
    unsigned ball = __ballot_sync(0xffffffffu, hit);
    if (ball == 0u) return;
    const int my_slot = tile * RPT + grp + my_u * RPI;
    const uint32_t my_key = float_key(score);
    while (ball != 0u) {
      const int src = __ffs(ball) - 1;
      ball &= ball - 1;
      const uint32_t key = __shfl_sync(0xffffffffu, my_key, src);
      const int slot = __shfl_sync(0xffffffffu, my_slot, src);
      warp_offer(p, cand_arr, excl, &s_thr_key, key, slot, lane);
    }
  };

  auto score_tile = [&](const float4 (&d)[U][NV], float (&acc)[U], float (&nrm)[U]) {
#pragma unroll
    for (int u = 0; u < U; u++) {
      float a = 0.f, n = 0.f;
#pragma unroll
      for (int v = 0; v < NV; v++) {
        a = fmaf(d[u][v].x, qv[v].x, a);
        a = fmaf(d[u][v].y, qv[v].y, a);
        a = fmaf(d[u][v].z, qv[v].z, a);
        a = fmaf(d[u][v].w, qv[v].w, a);
        if (KIND == ORYX_COSINE) {
          n = fmaf(d[u][v].x, d[u][v].x, n);
          n = fmaf(d[u][v].y, d[u][v].y, n);
          n = fmaf(d[u][v].z, d[u][v].z, n);
          n = fmaf(d[u][v].w, d[u][v].w, n);
        }
      }
      acc[u] = a;
      nrm[u] = n;
    }
    reduce_rows<G, U>(acc, gl);
    if (KIND == ORYX_COSINE) reduce_rows<G, U>(nrm, gl);
  };

  // A full tile of an unfiltered scan: every load unconditional.
  auto fast_tile = [&](int tile) {
    const float4* ptr = p.y + ((int64_t)tile * RPT + grp) * f4 + gl;
    const int row_step = RPI * f4;
    float4 d[U][NV];
#pragma unroll
    for (int u = 0; u < U; u++)
#pragma unroll
      for (int v = 0; v < NV; v++) d[u][v] = ldg_stream(ptr + u * row_step + v * G);
    const uint32_t thr_key = *(volatile unsigned int*)&s_thr_key;
    float acc[U], nrm[U];
    score_tile(d, acc, nrm);
    const float score = (KIND == ORYX_COSINE) ? acc[0] * rsqrtf(nrm[0]) : acc[0];
    offer(score, gl < U && score == score && float_key(score) > thr_key, tile);
  };

  // Any tile: LSH-filtered scans, ragged widths, the final partial tile.  `check_part` = test each
  // row's partition (mixed blocks); uniform candidate blocks skip that.
  auto general_tile = [&](int tile, bool check_part) {
    const int row0 = tile * RPT + grp;
    float4 d[U][NV];
    bool act[U];
#pragma unroll
    for (int u = 0; u < U; u++) {
      const int row = row0 + u * RPI;
      act[u] = row < n_slots;
      if (LSH && check_part) {
        int part = act[u] ? __ldg(p.slot_part + row) : -1;
        act[u] = part >= 0 && __popc(part ^ main_idx) <= p.max_bits;
      }
    }
#pragma unroll
    for (int u = 0; u < U; u++) {
      const float4* src = p.y + (int64_t)(row0 + u * RPI) * f4;
#pragma unroll
      for (int v = 0; v < NV; v++) {
        int idx = gl + v * G;
        d[u][v] = (act[u] && idx < f4) ? ldg_stream(src + idx) : make_float4(0.f, 0.f, 0.f, 0.f);
      }
    }
    const uint32_t thr_key = *(volatile unsigned int*)&s_thr_key;
    float acc[U], nrm[U];
    score_tile(d, acc, nrm);
    const float score = (KIND == ORYX_COSINE) ? acc[0] * rsqrtf(nrm[0]) : acc[0];
    // the owning lane decides; its row's activity flag sits at index my_u
    bool mine = false;
#pragma unroll
    for (int u = 0; u < U; u++) mine = mine || (u == my_u && act[u]);
    offer(score, gl < U && mine && score == score && float_key(score) > thr_key, tile);
  };

  // Hybrid schedule over tiles [0, total): 3/4 statically interleaved by warp, the rest handed out
  // from the global counter in guided chunks (large while much work remains, single tiles at the
  // end).  Every fetch is issued one chunk ahead -- the first one before the static phase -- so the
  // atomic's latency (and the burst of all warps hitting one address) hides behind useful work.
  auto for_each_tile = [&](int total, auto&& body) {
    const int iters_static = (total / warps_total) * 3 / 4;
    const int static_tiles = iters_static * warps_total;
    const int dyn_total = total - static_tiles;
    auto guided = [&](int remaining) { return max(1, min(kDynChunkMax, remaining / (2 * warps_total))); };
    int size = guided(dyn_total);
    int nxt = 0;
    if (lane == 0) nxt = (int)atomicAdd(p.counter, (unsigned)size);
    for (int i = 0; i < iters_static; i++) body(gw + i * warps_total);
    for (;;) {
      const int cur = __shfl_sync(0xffffffffu, nxt, 0);
      if (cur >= dyn_total) break;
      const int end = min(cur + size, dyn_total);
      size = guided(dyn_total - end);
      if (lane == 0) nxt = (int)atomicAdd(p.counter, (unsigned)size);
      for (int t = cur; t < end; t++) body(static_tiles + t);
    }
  };

  if (LSH) {
    // Filtered scan, scheduled by 256-slot blocks: a block whose live rows all sit in one partition is
    // skipped outright or streamed like an unfiltered tile run; only mixed blocks pay for per-row
    // metadata.  After oryx_items_compact nearly every block is uniform.
    // Work is handed out in 32-row units (8 per block) so that the few candidate blocks spread evenly
    // over all warps.
    constexpr int kUnitRows = 32;
    constexpr int tiles_per_unit = kUnitRows / RPT;
    constexpr int units_per_block = kBlockRows / kUnitRows;
    auto scan_unit = [&](int unit, int bp) {
      const bool uniform = bp >= 0;
      const int t0 = unit * tiles_per_unit, t1 = min(n_tiles, t0 + tiles_per_unit);
      for (int t = t0; t < t1; t++) {
        if (uniform && full_width && t < n_full_tiles) fast_tile(t);
        else general_tile(t, !uniform);
      }
    };
    if (p.worklist != nullptr) {
      // the candidate blocks were listed on the device after the query was hashed: every work item
      // below is real work, so the hybrid schedule balances it like an unfiltered scan
      const int n_items = __ldcg(p.worklist) * units_per_block;
      for_each_tile(n_items, [&](int w) {
        const int b = __ldcg(p.worklist + 1 + w / units_per_block);
        scan_unit(b * units_per_block + (w % units_per_block), __ldg(p.block_part + b));
      });
    } else {
      const int n_units = (n_slots + kUnitRows - 1) / kUnitRows;
      for_each_tile(n_units, [&](int unit) {
        const int bp = __ldg(p.block_part + unit / units_per_block);
        if (bp == -1) return;
        if (bp >= 0 && __popc(bp ^ main_idx) > p.max_bits) return;
        scan_unit(unit, bp);
      });
    }
  } else if (full_width) {
    for_each_tile(n_full_tiles, fast_tile);
    if (n_tiles > n_full_tiles && gw == n_full_tiles % warps_total) general_tile(n_full_tiles, false);
  } else {
    for_each_tile(n_tiles, [&](int t) { general_tile(t, false); });
  }

  // publish this CTA's candidates (key, slot); unused entries still carry the empty sentinel key 0.
  __syncthreads();
  if (p.debug_ts && tid == 0) {
    unsigned long long t = global_ns();
    atomicMin(p.debug_ts + 1, t);
    atomicMax(p.debug_ts + 2, t);
  }
  uint2* mine_out = p.cand + (size_t)blockIdx.x * p.k;
  for (int i = tid; i < p.k; i += T) {
    const unsigned long long v = cand_arr[i];
    uint2 e;
    e.x = (uint32_t)(v >> 32);
    e.y = (uint32_t)v;
    mine_out[i] = e;
  }
  if (tid < 32) {  // the list's best key lets the finishing kernel discard whole lists unread
    uint32_t mx = 0u;
    for (int i = lane; i < p.k; i += 32) mx = max(mx, (uint32_t)(cand_arr[i] >> 32));
    mx = __reduce_max_sync(0xffffffffu, mx);
    if (lane == 0) p.list_max[blockIdx.x] = mx;
  }
}

// ------------------------------------------------------------------------------------------------
// Finishing kernel: one CTA of 1024 threads.

// ---- selection for small k (k <= 32) -----------------------------------------------------------
// The k-th largest of the lists' maxima, M, is a lower bound of the global k-th best score (k
// different lists each hold an entry >= M).  So only lists whose maximum reaches M can contribute,
// and only with entries >= M: at most ~k lists x k entries instead of n_lists x k.  Those few
// entries are gathered into shared memory and rank-sorted, one thread per entry.
constexpr int kSmallKMax = 32;
constexpr int kGatherCap = 1024;  // = kFinishThreads: one thread per gathered entry

// k-th largest of n 32-bit keys in shared memory, by the whole CTA: radix select, four passes of eight bits,
// most significant first, a 256-bin histogram in shared memory per pass.  Returns (to every thread) the largest t
// with count(key >= t) >= k, or 1 if fewer than k keys are non-zero.  (A 32-step bisection by one warp took
// 9.7 us of the finishing kernel's 27: device timestamps, round 2.)
__device__ __forceinline__ uint32_t cta_kth_largest_key(const uint32_t* keys, int n, int k, unsigned int* hist /*[256]*/) {
  constexpr int T = kFinishThreads;
  const int tid = threadIdx.x;
  __shared__ unsigned int s_prefix, s_mask, s_remaining;
  if (tid == 0) { s_prefix = 0u; s_mask = 0u; s_remaining = (unsigned)k; }
  for (int pass = 0; pass < 4; pass++) {
    const int shift = 24 - 8 * pass;
    for (int b = tid; b < 256; b += T) hist[b] = 0u;
    __syncthreads();
    const unsigned prefix = s_prefix, mask = s_mask;
    for (int i = tid; i < n; i += T) {
      const unsigned key = keys[i];
      if ((key & mask) == prefix) atomicAdd(&hist[(key >> shift) & 255u], 1u);
    }
    __syncthreads();
    if (tid < 32) {
      // lane l owns bins [255-8l-7, 255-8l]; walk from the top bin down
      unsigned local[8];
      unsigned sum = 0;
#pragma unroll
      for (int q = 0; q < 8; q++) { local[q] = hist[255 - 8 * tid - q]; sum += local[q]; }
      unsigned incl = sum;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const unsigned t = __shfl_up_sync(0xffffffffu, incl, d);
        if (tid >= d) incl += t;
      }
      const unsigned excl_sum = incl - sum;
      const unsigned remaining = s_remaining;
      // exactly one lane owns the crossing: the keys matching the prefix number >= remaining >= 1 (n >= k)
      if (excl_sum < remaining && incl >= remaining) {
        unsigned cum = excl_sum;
        int digit = 0;
#pragma unroll
        for (int q = 0; q < 8; q++) {
          if (cum + local[q] >= remaining) { digit = 255 - 8 * tid - q; break; }
          cum += local[q];
        }
        s_prefix = prefix | ((unsigned)digit << shift);
        s_mask = mask | (255u << shift);
        s_remaining = remaining - cum;
      }
    }
    __syncthreads();
  }
  return max(s_prefix, 1u);  // the k-th largest key itself; 0 = fewer than k non-zero keys
}

// Returns the number selected (written to out[0..) in descending order), or -1 if the gather buffer
// overflowed (pathological ties) and the caller must use the general path.
__device__ int select_small_k(const ScanParams& p, int n_lists, uint32_t* s_max /*[n_lists]*/,
                              unsigned long long* s_ent /*[kGatherCap]*/, unsigned int* hist /*[256]*/,
                              unsigned long long* out) {
  constexpr int T = kFinishThreads;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int k = p.k;
  __shared__ unsigned int s_count;
  for (int i = tid; i < n_lists; i += T) s_max[i] = __ldcg(p.list_max + i);
  if (tid == 0) s_count = 0u;
  __syncthreads();
  if (p.debug_ts && tid == 0) p.debug_ts[8] = global_ns();
  const uint32_t m = n_lists >= k ? cta_kth_largest_key(s_max, n_lists, k, hist) : 1u;
  if (p.debug_ts && tid == 0) p.debug_ts[9] = global_ns();
  // one warp per list (k <= 32: a lane per entry); a list whose maximum misses the cut is never read
  for (int list = warp; list < n_lists; list += T / 32) {
    if (s_max[list] < m) continue;
    if (lane < k) {
      const uint2 e = __ldcg(&p.cand[(int64_t)list * k + lane]);
      if (e.x != 0u && e.x >= m) {
        const unsigned pos = atomicAdd(&s_count, 1u);
        if (pos < (unsigned)kGatherCap) s_ent[pos] = ((unsigned long long)e.x << 32) | (unsigned long long)e.y;
      }
    }
  }
  __syncthreads();
  if (p.debug_ts && tid == 0) p.debug_ts[10] = global_ns();
  const int n = (int)s_count;
  if (n > kGatherCap) return -1;
  // rank-sort: words are unique, rank = number of larger words
  if (tid < n) {
    const unsigned long long mine = s_ent[tid];
    int rank = 0;
    for (int j = 0; j < n; j++) rank += s_ent[j] > mine ? 1 : 0;
    if (rank < k) out[rank] = mine;
  }
  __syncthreads();
  return min(n, k);
}

// Radix select (4 passes x 8 bits, most significant first) of the k largest keys among `total`
// entries read through `key_at` / `entry_at`.  Used when k is large.  Writes packed words to out
// (unordered) and returns how many.
template <typename KeyAt, typename EntryAt>
__device__ int select_radix(int64_t total, int k, unsigned int* hist, unsigned long long* out, KeyAt key_at,
                            EntryAt entry_at) {
  constexpr int T = kFinishThreads;
  const int tid = threadIdx.x;
  __shared__ unsigned int s_prefix, s_mask, s_remaining, s_nsel, s_neq;
  if (tid == 0) { s_prefix = 0; s_mask = 0; s_remaining = (unsigned)k; s_nsel = 0; s_neq = 0; }
  __syncthreads();
  for (int pass = 0; pass < 4; pass++) {
    const int shift = 24 - 8 * pass;
    for (int b = tid; b < 256; b += T) hist[b] = 0;
    __syncthreads();
    const unsigned prefix = s_prefix, mask = s_mask;
    for (int64_t i = tid; i < total; i += T) {
      unsigned key = key_at(i);
      if ((key & mask) == prefix) atomicAdd(&hist[(key >> shift) & 255u], 1u);
    }
    __syncthreads();
    if (tid < 32) {
      // lane l owns bins [255-8l-7, 255-8l]; walk from the top bin down.
      unsigned local[8];
      unsigned sum = 0;
#pragma unroll
      for (int j = 0; j < 8; j++) { local[j] = hist[255 - 8 * tid - j]; sum += local[j]; }
      unsigned incl = sum;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        unsigned t = __shfl_up_sync(0xffffffffu, incl, d);
        if (tid >= d) incl += t;
      }
      unsigned excl_sum = incl - sum;
      unsigned remaining = s_remaining;
      // Exactly one lane owns the crossing: the matching keys always number >= remaining >= 1
      // (pass 0 sees all entries, sentinels included).
      if ((excl_sum < remaining) && (incl >= remaining)) {
        unsigned cum = excl_sum;
        int digit = 0;
#pragma unroll
        for (int j = 0; j < 8; j++) {
          if (cum + local[j] >= remaining) { digit = 255 - 8 * tid - j; break; }
          cum += local[j];
        }
        s_prefix = prefix | ((unsigned)digit << shift);
        s_mask = mask | (255u << shift);
        s_remaining = remaining - cum;
      }
    }
    __syncthreads();
  }
  const unsigned thr_key = s_prefix;
  const unsigned need_eq = s_remaining;
  __syncthreads();
  for (int64_t i = tid; i < total; i += T) {
    unsigned long long e = entry_at(i);
    unsigned key = (unsigned)(e >> 32);
    if (key == 0u) continue;  // empty sentinel
    bool take = key > thr_key;
    if (!take && key == thr_key && thr_key != 0u) take = atomicAdd(&s_neq, 1u) < need_eq;
    if (take) {
      unsigned pos = atomicAdd(&s_nsel, 1u);
      if (pos < (unsigned)k) out[pos] = e;
    }
  }
  __syncthreads();
  return min((int)s_nsel, k);
}

// Byte offsets of the finishing CTA's shared-memory regions (host and device agree through this).
struct FinishLayout {
  size_t sel, exact, rows, hist, q, rowbuf, lmax, gather, total;
};
__host__ __device__ inline FinishLayout finish_layout(int k, int fpad, int rescore_warps, int n_lists) {
  FinishLayout L;
  size_t o = 0;
  L.sel = o;    o += (size_t)k * 8;
  L.exact = o;  o += (size_t)k * 8;
  L.rows = o;   o += (size_t)k * 4;
  o = (o + 15) & ~(size_t)15;
  L.hist = o;   o += 256 * 4;
  L.q = o;      o += (size_t)fpad * 4;
  L.rowbuf = o; o += (size_t)rescore_warps * fpad * 4;
  o = (o + 15) & ~(size_t)15;
  L.gather = o; o += (k <= 32) ? (size_t)1024 * 8 : 0;
  L.lmax = o;   o += (k <= 32) ? (size_t)n_lists * 4 : 0;
  L.total = (o + 15) & ~(size_t)15;
  return L;
}

__global__ void __launch_bounds__(kFinishThreads, 1)
topn_finish_kernel(const __grid_constant__ ScanParams p, int n_lists, int kind, int rescore_warps) {
  constexpr int T = kFinishThreads;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int k = p.k;
  const int64_t total = (int64_t)n_lists * k;
  const int fpad = p.f4 * 4;
  const FinishLayout L = finish_layout(k, fpad, rescore_warps, n_lists);
  unsigned long long* sel = reinterpret_cast<unsigned long long*>(smem_raw + L.sel);
  double* sel_exact = reinterpret_cast<double*>(smem_raw + L.exact);
  int32_t* sel_row = reinterpret_cast<int32_t*>(smem_raw + L.rows);
  unsigned int* hist = reinterpret_cast<unsigned int*>(smem_raw + L.hist);
  float* s_q = reinterpret_cast<float*>(smem_raw + L.q);
  float* rowbuf = reinterpret_cast<float*>(smem_raw + L.rowbuf);  // [rescore_warps x fpad]
  unsigned long long* s_ent = reinterpret_cast<unsigned long long*>(smem_raw + L.gather);
  uint32_t* s_max = reinterpret_cast<uint32_t*>(smem_raw + L.lmax);

  // The next query's scan does not depend on this kernel (it uses the other scratch set): let it
  // start now -- but not before the PREVIOUS finishing kernel is done with that other set (a tiny
  // scan can complete, and release this kernel, while its predecessor is still selecting).  This
  // kernel itself was launched ahead of time and must wait for its own scan.
  if (p.order) {
    if (tid == 0) {
      const unsigned long long t0 = global_ns();
      unsigned long long v;
      do {
        asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p.order) : "memory");
      } while (v + 1 < p.order_value && global_ns() - t0 < 20000000000ull);
    }
    __syncthreads();
  }
  pdl_launch_dependents();
  for (int j = tid; j < fpad; j += T) s_q[j] = __ldg(p.q + j);
  pdl_wait_primary();
  if (p.debug_ts && tid == 0) p.debug_ts[3] = global_ns();

  int nsel = -1;
  if (k <= kSmallKMax) nsel = select_small_k(p, n_lists, s_max, s_ent, hist, sel);
  if (nsel < 0) {
    __syncthreads();
    nsel = select_radix(total, k, hist, sel,
                        [&](int64_t i) { return __ldcg(&p.cand[i].x); },
                        [&](int64_t i) {
                          uint2 e = __ldcg(&p.cand[i]);
                          return ((unsigned long long)e.x << 32) | (unsigned long long)e.y;
                        });
  }
  if (p.debug_ts && tid == 0) p.debug_ts[4] = global_ns();

  // Exact re-score: VectorMath.dot / CosineAverageFunction.applyAsDouble arithmetic.  One warp per
  // selected row: the lanes stage the row in shared memory with coalesced loads (one round trip).
  if (warp < rescore_warps) {
    for (int i = warp; i < nsel; i += rescore_warps) {
      const int32_t slot = (int32_t)(uint32_t)sel[i];
      const float4* y4 = p.y + (int64_t)slot * p.f4;
      float4* rb4 = reinterpret_cast<float4*>(rowbuf + (size_t)warp * fpad);
      for (int j = lane; j < p.f4; j += 32) rb4[j] = y4[j];
      __syncwarp();
      const int32_t row_id = lane == 0 ? __ldg(p.slot_row + slot) : 0;  // issued before the sums, used after them
      const float* y = rowbuf + (size_t)warp * fpad;
      // the whole warp computes the chain's value when the chain provably never rounds (common.cuh) ...
      double dot = 0.0, tot = 0.0;
      bool fast = warp_exact_dot(s_q, y, p.features, lane, &dot);
      if (kind == ORYX_COSINE) fast = warp_exact_dot(y, y, p.features, lane, &tot) && fast;
      if (lane == 0) {
        if (!fast) {  // ... otherwise lane 0 walks it
          dot = 0.0;
          tot = 0.0;
#pragma unroll 8
          for (int j = 0; j < p.features; j++) {
            const float yv = y[j];
            dot = exact_step(dot, s_q[j], yv);
            if (kind == ORYX_COSINE) tot = exact_step(tot, yv, yv);
          }
        }
        sel_exact[i] = (kind == ORYX_COSINE) ? __ddiv_rn(dot, __dsqrt_rn(tot)) : dot;
        sel_row[i] = row_id;
      }
      __syncwarp();
    }
  }
  __syncthreads();
  if (p.debug_ts && tid == 0) p.debug_ts[5] = global_ns();

  // Rank by (exact score desc, row asc) -- Pairs.orderBySecond(DESCENDING) (Pairs.java:48-54), ties
  // made deterministic.  A NaN exact score (possible only if fp32 and fp64 disagree about a zero
  // norm) ranks last.
  const int n_out = min(nsel, p.how_many);
  if (p.gate) {
    if (tid == 0) {
      const unsigned long long t0 = global_ns();
      unsigned long long v;
      do {
        asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p.gate) : "memory");
      } while (v < p.gate_value && global_ns() - t0 < 20000000000ull);
    }
    __syncthreads();
  }
  for (int i = tid; i < nsel; i += T) {
    const double si = sel_exact[i];
    const int32_t ri = sel_row[i];
    const bool nan_i = si != si;
    int rank = 0;
    for (int j = 0; j < nsel; j++) {
      const double sj = sel_exact[j];
      const bool nan_j = sj != sj;
      const int32_t rj = sel_row[j];
      bool better;  // does j rank before i?
      if (nan_j) better = nan_i && rj < ri;
      else if (nan_i) better = true;
      else better = (sj > si) || (sj == si && rj < ri);
      rank += better ? 1 : 0;
    }
    if (rank < n_out) {
      p.out_rows[rank] = ri;
      p.out_scores[rank] = si;
    }
  }
  __syncthreads();
  if (tid == 0) {
    *p.out_n = n_out;
    *p.counter = 0u;  // the dynamic tile counter starts at zero for the next query on this context
    if (p.debug_ts) p.debug_ts[6] = global_ns();
    if (p.order) {
      __threadfence();
      asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(p.order), "l"(p.order_value) : "memory");
    }
  }
}

// ------------------------------------------------------------------------------------------------

// cudaOccupancyMaxActiveBlocksPerMultiprocessor costs microseconds; remember it per (kernel, smem).
int occupancy_cached(const void* kern, int threads, size_t smem, size_t max_dynamic_smem) {
  struct Entry { int device; const void* kern; size_t smem; int per_sm; };
  static std::mutex mu;
  static std::vector<Entry> cache;
  int device = 0;
  if (cudaGetDevice(&device) != cudaSuccess) { cudaGetLastError(); device = -1; }
  std::lock_guard<std::mutex> lk(mu);
  for (auto& e : cache) if (e.device == device && e.kern == kern && e.smem == smem) return e.per_sm;
  int per_sm = 0;
  // function attributes are per device: raise the limit on this one before asking about occupancy
  if (ensure_max_dynamic_smem(kern, max_dynamic_smem) != ORYX_OK ||
      cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, threads, smem) != cudaSuccess) {
    cudaGetLastError();
    per_sm = 1;
  }
  if (per_sm < 1) per_sm = 1;
  cache.push_back({device, kern, smem, per_sm});
  return per_sm;
}

// Dynamic shared memory of a scan CTA: the candidate array (8 B per entry) plus the staged exclusion list.
size_t scan_smem_bytes(int k, int n_excl_staged) {
  return (size_t)k * 8 + (size_t)n_excl_staged * sizeof(int32_t);
}

template <int G, int NV, int U>
int32_t launch_variant(Model* m, const ScanParams& p, int kind, bool lsh, bool overlap, cudaStream_t stream,
                       int* grid_out) {
  // 256-thread CTAs; the register allocation is bounded for 4 CTAs per SM (<= 64 registers) when a
  // thread keeps at most 8 float4 in flight.
  constexpr int T = kScanThreads;
  auto pick = [&](auto kern) -> int32_t {
    constexpr int RPT = (32 / G) * U;
    int64_t n_tiles = (p.n_slots + RPT - 1) / RPT;
    int64_t want = (n_tiles + (T / 32) - 1) / (T / 32);
    const size_t smem = scan_smem_bytes(p.k, p.excl_in_smem ? p.n_excl : 0);
    int per_sm = occupancy_cached(reinterpret_cast<const void*>(kern), T, smem, scan_smem_bytes(kMaxK, kExclSmemMax));
    const int sms = std::max(1, m->num_sms - std::max(0, m->cfg.reserved_sms));
    int grid = (int)std::min<int64_t>(std::min<int64_t>((int64_t)sms * per_sm, kMaxGrid), std::max<int64_t>(want, 1));
    *grid_out = grid;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(T);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = overlap ? 1 : 0;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    ORYX_CUDA(cudaLaunchKernelEx(&cfg, kern, p));
    return ORYX_OK;
  };
  if (kind == ORYX_COSINE) {
    return lsh ? pick(scan_topn_kernel<G, NV, U, ORYX_COSINE, true>)
               : pick(scan_topn_kernel<G, NV, U, ORYX_COSINE, false>);
  }
  return lsh ? pick(scan_topn_kernel<G, NV, U, ORYX_DOT, true>)
             : pick(scan_topn_kernel<G, NV, U, ORYX_DOT, false>);
}

int32_t launch_finish(const ScanParams& p, int n_lists, int kind, cudaStream_t stream) {
  const int fpad = p.f4 * 4;
  const int rescore_warps = std::max(1, std::min(kFinishThreads / 32, kRowStageFloats / fpad));
  const FinishLayout L = finish_layout(p.k, fpad, rescore_warps, n_lists);
  ORYX_TRY(ensure_max_dynamic_smem(reinterpret_cast<const void*>(topn_finish_kernel), 224 * 1024));
  if (L.total > 224 * 1024) {
    set_error("finishing kernel needs %zu bytes of shared memory (features=%d, how_many=%d)", L.total, p.features,
              p.how_many);
    return ORYX_E_UNSUPPORTED;
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(1);
  cfg.blockDim = dim3(kFinishThreads);
  cfg.dynamicSmemBytes = L.total;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;  // pre-launched; waits for its scan on the device
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  ORYX_CUDA(cudaLaunchKernelEx(&cfg, topn_finish_kernel, p, n_lists, kind, rescore_warps));
  return ORYX_OK;
}

}  // namespace

int32_t scan_topn_launch(Model* m, QueryCtx* ctx, const TopnRequest& rq, cudaStream_t stream) {
  Store* s = &m->store[ORYX_Y];
  ORYX_ARG(rq.how_many >= 1 && rq.how_many <= ORYX_MAX_HOW_MANY, "how_many out of range [1, ORYX_MAX_HOW_MANY]");
  ORYX_ARG(rq.kind == ORYX_DOT || rq.kind == ORYX_COSINE, "kind must be ORYX_DOT or ORYX_COSINE");
  if (s->n_slots == 0) {  // ASM:307-310 no tasks -> Stream.empty()
    ORYX_CUDA(cudaMemsetAsync(rq.d_out_n, 0, sizeof(int32_t), stream));
    return ORYX_OK;
  }
  const bool lsh = s->hashed && m->num_hashes != m->max_bits;  // LSH.java:160-162: all partitions otherwise
  int32_t* worklist = nullptr;
  if (lsh) {
    const bool owns = ctx->scratch_sets == 2;  // a context that owns memory: list the candidate blocks
    if (owns) {
      const int64_t n_blocks = (s->n_slots + kBlockRows - 1) / kBlockRows;
      if (ctx->worklist_cap < n_blocks + 1) {
        cudaFree(ctx->d_worklist);
        ctx->d_worklist = nullptr;
        ctx->worklist_cap = 0;
        const int64_t cap = (s->cap_slots + kBlockRows - 1) / kBlockRows + 2;
        ORYX_CUDA(cudaMalloc(&ctx->d_worklist, 2 * (size_t)cap * sizeof(int32_t)));
        ORYX_CUDA(cudaMemsetAsync(ctx->d_worklist, 0, 2 * (size_t)cap * sizeof(int32_t), stream));  // both counts start at zero
        ctx->worklist_cap = cap;
      }
      if (ctx->d_worklist_ctl == nullptr) {
        ORYX_CUDA(cudaMalloc(&ctx->d_worklist_ctl, 4 * sizeof(unsigned int)));
        ORYX_CUDA(cudaMemsetAsync(ctx->d_worklist_ctl, 0, 4 * sizeof(unsigned int), stream));
      }
      // One launch hashes the query and lists its candidate blocks.  It depends on nothing earlier in the
      // stream, so it starts while the previous query's finishing kernel still runs; the previous scan (the
      // only reader of main_index and of the work lists) has completed by then, because the finishing kernel
      // that releases us is itself only launched when its scan has exited.
      worklist = ctx->d_worklist + (size_t)(ctx->worklist_flip ^= 1) * ctx->worklist_cap;
      ORYX_TRY(lsh_query_worklist_device(m, rq.d_query, ctx->d_main_index, worklist,
                                         ctx->d_worklist + (size_t)(ctx->worklist_flip ^ 1) * ctx->worklist_cap,
                                         ctx->d_worklist_ctl, &ctx->worklist_seq, stream));
    } else {
      ORYX_TRY(lsh_hash_rows_device(m, rq.d_query, 1, s->fpad, ctx->d_main_index, stream, nullptr, false));
    }
  }
  ScanParams p{};
  p.y = reinterpret_cast<const float4*>(s->d_vec);
  p.slot_row = s->d_slot_row;
  p.slot_part = s->d_slot_part;
  p.block_part = s->d_block_part;
  p.worklist = worklist;
  p.worklist_flag = worklist ? ctx->d_worklist_ctl : nullptr;
  p.worklist_seq = ctx->worklist_seq;
  p.q = rq.d_query;
  p.excl = rq.d_excl;
  p.n_excl = rq.n_excl;
  p.excl_in_smem = rq.n_excl > 0 && rq.n_excl <= kExclSmemMax;
  p.n_slots = s->n_slots;
  p.features = s->features;
  p.f4 = s->fpad / 4;
  p.k = rq.how_many + kTopnMargin;
  p.how_many = rq.how_many;
  p.max_bits = m->max_bits;
  p.main_index = ctx->d_main_index;
  // Two scratch sets alternate so that this query's scan may start while the previous query's
  // finishing kernel (which reads the other set) is still running (an LSH-filtered scan waits, on the
  // device, for the kernels that hash its query and list its candidate blocks).
  const bool overlap = ctx->scratch_sets == 2;
  const int set = ctx->scratch_sets == 2 ? (ctx->flip ^= 1) : 0;
  p.cand = ctx->d_cand + (size_t)set * kMaxGrid * kMaxK;
  p.list_max = ctx->d_list_max + (size_t)set * kMaxGrid;
  p.counter = ctx->d_counter + set * 16;
  p.out_rows = rq.d_out_rows;
  p.out_scores = rq.d_out_scores;
  p.out_n = rq.d_out_n;
  p.debug_ts = m->d_debug_ts;
  p.gate = rq.gate;
  p.gate_value = rq.gate_value;
  p.order = (ctx->scratch_sets == 2 && ctx->d_finish_seq) ? ctx->d_finish_seq : nullptr;
  p.order_value = p.order ? ++ctx->finish_seq : 0;
  int grid = 0;
  int32_t rc;
  const int f4 = p.f4;
  if (f4 <= 8) rc = launch_variant<8, 1, 8>(m, p, rq.kind, lsh, overlap, stream, &grid);
  else if (f4 <= 16) rc = launch_variant<8, 2, 4>(m, p, rq.kind, lsh, overlap, stream, &grid);
  else if (f4 <= 24) rc = launch_variant<8, 3, 2>(m, p, rq.kind, lsh, overlap, stream, &grid);
  else if (f4 <= 32) rc = launch_variant<8, 4, 2>(m, p, rq.kind, lsh, overlap, stream, &grid);
  else if (f4 <= 48) rc = launch_variant<8, 6, 1>(m, p, rq.kind, lsh, overlap, stream, &grid);
  else if (f4 <= 64) rc = launch_variant<8, 8, 1>(m, p, rq.kind, lsh, overlap, stream, &grid);
  else if (f4 <= 128) rc = launch_variant<32, 4, 2>(m, p, rq.kind, lsh, overlap, stream, &grid);
  else if (f4 <= 256) rc = launch_variant<32, 8, 1>(m, p, rq.kind, lsh, overlap, stream, &grid);
  else if (f4 <= 512) rc = launch_variant<32, 16, 1>(m, p, rq.kind, lsh, overlap, stream, &grid);
  else {
    set_error("features=%d beyond the scan kernel's envelope (2048)", s->features);
    return ORYX_E_UNSUPPORTED;
  }
  ORYX_TRY(rc);
  ORYX_TRY(launch_finish(p, grid, rq.kind, stream));
  m->launches += 2;
  m->queries++;
  return ORYX_OK;
}

}  // namespace oryx
