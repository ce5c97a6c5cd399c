// scan_topn.cu -- single-query scoring scan with a fused per-CTA top-N heap (kernels K1/K2/K7 of
// SURVEY.md section 2.3).
//
// Replaces the hot loop of ALSServingModel.topN (reference: app/oryx-app-serving/src/main/java/com/
// cloudera/oryx/app/serving/als/model/ALSServingModel.java:289-335): per candidate partition a
// TopNConsumer (TopNConsumer.java:56-74) scores every vector with DotsFunction or
// CosineAverageFunction (VectorMath.dot, VectorMath.java:38-45) and keeps the howMany best; the
// per-partition lists are concatenated, sorted descending and cut (ALSServingModel.java:329-334).
//
// Device algorithm (one launch per query):
//   1. scan     -- HBM-bound.  A group of G lanes owns one row and reads it with coalesced 128-bit
//                  streaming loads; fp32 FMA partial sums are combined with xor-shuffles.  Rows of
//                  non-candidate LSH partitions are skipped before their loads are issued
//                  (candidate(p) <=> popcount(p ^ mainIndex) <= maxBitsDiffering, which is exactly
//                  the set LocalitySensitiveHash.getCandidateIndices enumerates).
//   2. per-CTA heap -- scores above the CTA's current k-th best (k = howMany + 8) take a rare slow
//                  path: the allowed-predicate (row is live, not in the exclusion list) is tested and
//                  the row is pushed into a shared-memory binary min-heap under a CTA-local lock.
//                  Strict '>' admission, NaN never enters (TopNConsumer.java:60-64).
//   3. merge    -- every CTA publishes its heap to global scratch; the last CTA to finish (atomic
//                  ticket) radix-selects the k best of all lists,
//   4. finish   -- re-scores those k rows with the reference's exact arithmetic (binary32 product,
//                  binary64 accumulate, index order; cosine adds the exact norm and a binary64
//                  sqrt/divide), sorts by (exact score desc, row asc) and writes howMany results.
// Returned scores are therefore bit-identical to the Java path's; the fp32 scan only decides
// which k rows get re-scored.
#include <algorithm>

#include "common.cuh"

namespace oryx {

namespace {

constexpr int kThreads = 256;

struct ScanParams {
  const float4* y;
  const int32_t* slot_row;
  const int32_t* slot_part;
  const float* q;  // [fpad]
  const int32_t* excl;
  int n_excl;
  int64_t n_slots;
  int features;
  int f4;  // fpad / 4
  int k;   // heap size = how_many + margin
  int how_many;
  int max_bits;
  const int32_t* main_index;
  uint2* cand;  // scratch [gridDim.x x k]
  unsigned int* counter;
  int32_t* out_rows;
  double* out_scores;
  int32_t* out_n;
};

struct HeapView {
  float* score;
  int32_t* slot;
};

__device__ __forceinline__ void heap_sift_up(HeapView h, int i) {
  float s = h.score[i];
  int v = h.slot[i];
  while (i > 0) {
    int p = (i - 1) >> 1;
    if (!(s < h.score[p])) break;
    h.score[i] = h.score[p];
    h.slot[i] = h.slot[p];
    i = p;
  }
  h.score[i] = s;
  h.slot[i] = v;
}

__device__ __forceinline__ void heap_replace_root(HeapView h, int n, float s, int v) {
  int i = 0;
  int half = n >> 1;
  while (i < half) {
    int c = 2 * i + 1;
    if (c + 1 < n && h.score[c + 1] < h.score[c]) c++;
    if (!(h.score[c] < s)) break;
    h.score[i] = h.score[c];
    h.slot[i] = h.slot[c];
    i = c;
  }
  h.score[i] = s;
  h.slot[i] = v;
}

__device__ __forceinline__ bool excluded(const int32_t* excl, int n, int32_t row) {
  int lo = 0, hi = n - 1;
  while (lo <= hi) {
    int mid = (lo + hi) >> 1;
    int32_t v = __ldg(excl + mid);
    if (v == row) return true;
    if (v < row) lo = mid + 1; else hi = mid - 1;
  }
  return false;
}

// Slow path, executed by one lane of a warp for one qualifying row.
__device__ __noinline__ void heap_offer(const ScanParams& p, HeapView h, int* s_cnt, int* s_lock,
                                        volatile float* s_thr, float score, int32_t slot) {
  // TopNConsumer.java:57 allowedPredicate: the slot must hold a live row that is not excluded.
  int32_t row = __ldg(p.slot_row + slot);
  if (row < 0) return;
  if (p.n_excl > 0 && excluded(p.excl, p.n_excl, row)) return;
  while (atomicCAS(s_lock, 0, 1) != 0) {
  }
  __threadfence_block();
  int n = *(volatile int*)s_cnt;
  if (n < p.k) {
    h.score[n] = score;
    h.slot[n] = slot;
    heap_sift_up(h, n);
    n++;
    *(volatile int*)s_cnt = n;
    if (n == p.k) *s_thr = h.score[0];
  } else if (score > h.score[0]) {  // re-check against the live minimum (TopNConsumer.java:63)
    heap_replace_root(h, n, score, slot);
    *s_thr = h.score[0];
  }
  __threadfence_block();
  atomicExch(s_lock, 0);
}

// ---- the last CTA's work: select, exact re-score, order, write.
__device__ void final_merge(const ScanParams& p, int kind, unsigned char* smem_raw, int n_lists) {
  const int tid = threadIdx.x;
  const int k = p.k;
  const int64_t total = (int64_t)n_lists * k;
  // shared layout: sel_key[k] u32 | sel_row[k] i32 | sel_exact[k] f64 | hist[256] | misc
  uint32_t* sel_key = reinterpret_cast<uint32_t*>(smem_raw);
  int32_t* sel_slot = reinterpret_cast<int32_t*>(smem_raw + (size_t)k * 4);
  double* sel_exact = reinterpret_cast<double*>(smem_raw + (size_t)k * 8);
  unsigned int* hist = reinterpret_cast<unsigned int*>(smem_raw + (size_t)k * 16);
  __shared__ unsigned int s_prefix, s_mask, s_remaining, s_nsel, s_neq;
  if (tid == 0) { s_prefix = 0; s_mask = 0; s_remaining = (unsigned)k; s_nsel = 0; s_neq = 0; }
  __syncthreads();

  // Radix select (4 passes x 8 bits, most significant first) of the k-th largest key.
  for (int pass = 0; pass < 4; pass++) {
    const int shift = 24 - 8 * pass;
    for (int b = tid; b < 256; b += kThreads) hist[b] = 0;
    __syncthreads();
    const unsigned prefix = s_prefix, mask = s_mask;
    for (int64_t i = tid; i < total; i += kThreads) {
      unsigned key = __ldcg(&p.cand[i].x);
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
      bool mine = (excl_sum < remaining) && (incl >= remaining);
      // Exactly one lane owns the crossing: the matching keys always number >= remaining >= 1
      // (pass 0 sees all gridDim.x * k entries, sentinels included).
      if (mine) {
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
  for (int64_t i = tid; i < total; i += kThreads) {
    uint2 e = __ldcg(&p.cand[i]);
    if (e.x == 0u) continue;  // empty sentinel
    bool take = e.x > thr_key;
    if (!take && e.x == thr_key && thr_key != 0u) take = atomicAdd(&s_neq, 1u) < need_eq;
    if (take) {
      unsigned pos = atomicAdd(&s_nsel, 1u);
      if (pos < (unsigned)k) { sel_key[pos] = e.x; sel_slot[pos] = (int32_t)e.y; }
    }
  }
  __syncthreads();
  const int nsel = min((int)s_nsel, k);

  // Exact re-score: VectorMath.dot / CosineAverageFunction.applyAsDouble arithmetic, one thread per row.
  for (int i = tid; i < nsel; i += kThreads) {
    const int32_t slot = sel_slot[i];
    const float* y = reinterpret_cast<const float*>(p.y) + (int64_t)slot * p.f4 * 4;
    double dot = 0.0, tot = 0.0;
    for (int j = 0; j < p.features; j++) {
      float yv = y[j];
      dot = exact_step(dot, __ldg(p.q + j), yv);
      if (kind == ORYX_COSINE) tot = exact_step(tot, yv, yv);
    }
    sel_exact[i] = (kind == ORYX_COSINE) ? __ddiv_rn(dot, __dsqrt_rn(tot)) : dot;
    sel_slot[i] = __ldg(p.slot_row + slot);  // from here on: row ids
  }
  __syncthreads();
  // Rank by (exact score desc, row asc) -- Pairs.orderBySecond(DESCENDING) (Pairs.java:48-54), ties
  // made deterministic.  A NaN exact score (possible only if fp32 and fp64 disagree about a zero
  // norm) ranks last.
  const int n_out = min(nsel, p.how_many);
  for (int i = tid; i < nsel; i += kThreads) {
    const double si = sel_exact[i];
    const int32_t ri = sel_slot[i];
    const bool nan_i = si != si;
    int rank = 0;
    for (int j = 0; j < nsel; j++) {
      const double sj = sel_exact[j];
      const bool nan_j = sj != sj;
      const int32_t rj = sel_slot[j];
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
  if (tid == 0) *p.out_n = n_out;
}

// G lanes per row, NV float4 per lane, U rows per group in flight.
template <int G, int NV, int U, int KIND, bool LSH>
__global__ void __launch_bounds__(kThreads) scan_topn_kernel(const __grid_constant__ ScanParams p) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ float s_thr;
  __shared__ int s_cnt, s_lock, s_last;

  HeapView heap;
  heap.score = reinterpret_cast<float*>(smem_raw);
  heap.slot = reinterpret_cast<int32_t*>(smem_raw + (size_t)p.k * 4);

  const int tid = threadIdx.x;
  const int lane = tid & 31;
  const int gl = lane % G;   // lane within the row group
  const int grp = lane / G;  // which row of the warp's row set
  constexpr int RPI = 32 / G;       // rows covered by one load instruction of a warp
  constexpr int RPT = RPI * U;      // rows per warp tile

  if (tid == 0) { s_thr = -INFINITY; s_cnt = 0; s_lock = 0; }
  __syncthreads();

  float4 qv[NV];
#pragma unroll
  for (int v = 0; v < NV; v++) {
    int idx = gl + v * G;
    qv[v] = idx < p.f4 ? __ldg(reinterpret_cast<const float4*>(p.q) + idx) : make_float4(0.f, 0.f, 0.f, 0.f);
  }
  const int main_idx = LSH ? __ldg(p.main_index) : 0;

  const int64_t warps_total = (int64_t)gridDim.x * (kThreads / 32);
  const int64_t gw = (int64_t)blockIdx.x * (kThreads / 32) + (tid >> 5);
  const int64_t n_tiles = (p.n_slots + RPT - 1) / RPT;

  for (int64_t tile = gw; tile < n_tiles; tile += warps_total) {
    const int64_t row0 = tile * RPT + grp;
    float4 d[U][NV];
    bool act[U];
#pragma unroll
    for (int u = 0; u < U; u++) {
      const int64_t row = row0 + u * RPI;
      act[u] = row < p.n_slots;
      if (LSH) {
        int part = act[u] ? __ldg(p.slot_part + row) : -1;
        act[u] = part >= 0 && __popc(part ^ main_idx) <= p.max_bits;
      }
    }
#pragma unroll
    for (int u = 0; u < U; u++) {
      const int64_t row = row0 + u * RPI;
      const float4* src = p.y + row * p.f4;
#pragma unroll
      for (int v = 0; v < NV; v++) {
        int idx = gl + v * G;
        d[u][v] = (act[u] && idx < p.f4) ? ldg_stream(src + idx) : make_float4(0.f, 0.f, 0.f, 0.f);
      }
    }
    float thr = *(volatile float*)&s_thr;
#pragma unroll
    for (int u = 0; u < U; u++) {
      float acc = 0.f, nrm = 0.f;
#pragma unroll
      for (int v = 0; v < NV; v++) {
        acc = fmaf(d[u][v].x, qv[v].x, acc);
        acc = fmaf(d[u][v].y, qv[v].y, acc);
        acc = fmaf(d[u][v].z, qv[v].z, acc);
        acc = fmaf(d[u][v].w, qv[v].w, acc);
        if (KIND == ORYX_COSINE) {
          nrm = fmaf(d[u][v].x, d[u][v].x, nrm);
          nrm = fmaf(d[u][v].y, d[u][v].y, nrm);
          nrm = fmaf(d[u][v].z, d[u][v].z, nrm);
          nrm = fmaf(d[u][v].w, d[u][v].w, nrm);
        }
      }
#pragma unroll
      for (int o = G / 2; o > 0; o >>= 1) {
        acc += __shfl_xor_sync(0xffffffffu, acc, o);
        if (KIND == ORYX_COSINE) nrm += __shfl_xor_sync(0xffffffffu, nrm, o);
      }
      float score = (KIND == ORYX_COSINE) ? acc * rsqrtf(nrm) : acc;
      bool hit = act[u] && gl == 0 && score > thr;
      unsigned ball = __ballot_sync(0xffffffffu, hit);
      if (ball != 0u) {
        while (ball != 0u) {
          int src = __ffs(ball) - 1;
          ball &= ball - 1;
          float s = __shfl_sync(0xffffffffu, score, src);
          int64_t r = row0 + u * RPI;
          int32_t slot = __shfl_sync(0xffffffffu, (int32_t)r, src);
          if (lane == 0) heap_offer(p, heap, &s_cnt, &s_lock, &s_thr, s, slot);
        }
        __syncwarp();
        thr = *(volatile float*)&s_thr;
      }
    }
  }

  // publish this CTA's heap (key, slot); unused entries carry the empty sentinel key 0.
  __syncthreads();
  const int cnt = s_cnt;
  uint2* mine = p.cand + (size_t)blockIdx.x * p.k;
  for (int i = tid; i < p.k; i += kThreads) {
    uint2 e;
    if (i < cnt) { e.x = float_key(heap.score[i]); e.y = (unsigned)heap.slot[i]; }
    else { e.x = 0u; e.y = 0u; }
    __stcg(mine + i, e);
  }
  __threadfence();
  __syncthreads();
  if (tid == 0) {
    unsigned ticket = atomicAdd(p.counter, 1u);
    s_last = (ticket == gridDim.x - 1);
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  final_merge(p, KIND, smem_raw, (int)gridDim.x);
  if (tid == 0) *p.counter = 0u;  // self-cleaning for the next query on this context
}

// cudaOccupancyMaxActiveBlocksPerMultiprocessor costs microseconds; remember it per (kernel, smem).
int occupancy_cached(const void* kern, size_t smem) {
  static std::mutex mu;
  static std::vector<std::pair<std::pair<const void*, size_t>, int>> cache;
  std::lock_guard<std::mutex> lk(mu);
  for (auto& e : cache) if (e.first.first == kern && e.first.second == smem) return e.second;
  int per_sm = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kThreads, smem) != cudaSuccess) {
    cudaGetLastError();
    per_sm = 1;
  }
  cache.push_back({{kern, smem}, per_sm});
  return per_sm;
}

template <int G, int NV, int U>
int32_t launch_variant(Model* m, const ScanParams& p, int kind, bool lsh, size_t smem, cudaStream_t stream,
                       int* grid_out) {
  auto pick = [&](auto kern) -> int32_t {
    int per_sm = occupancy_cached(reinterpret_cast<const void*>(kern), smem);
    if (per_sm < 1) per_sm = 1;
    constexpr int RPT = (32 / G) * U;
    int64_t n_tiles = (p.n_slots + RPT - 1) / RPT;
    int64_t want = (n_tiles + (kThreads / 32) - 1) / (kThreads / 32);
    int grid = (int)std::min<int64_t>(std::min<int64_t>((int64_t)m->num_sms * per_sm, kMaxGrid), std::max<int64_t>(want, 1));
    *grid_out = grid;
    kern<<<grid, kThreads, smem, stream>>>(p);
    ORYX_CUDA(cudaGetLastError());
    return ORYX_OK;
  };
  if (kind == ORYX_COSINE) {
    return lsh ? pick(scan_topn_kernel<G, NV, U, ORYX_COSINE, true>)
               : pick(scan_topn_kernel<G, NV, U, ORYX_COSINE, false>);
  }
  return lsh ? pick(scan_topn_kernel<G, NV, U, ORYX_DOT, true>)
             : pick(scan_topn_kernel<G, NV, U, ORYX_DOT, false>);
}

// One thread per slot, exact arithmetic, result scattered to out[row].
__global__ void scores_all_kernel(const float* __restrict__ y, const int32_t* __restrict__ slot_row,
                                  int64_t n_slots, int fpad, int features, const float* __restrict__ q,
                                  int kind, int32_t n_rows, double* __restrict__ out) {
  extern __shared__ float s_tile[];  // [blockDim.x x (features+1)]
  const int ld = features + 1;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  for (int64_t base = (int64_t)blockIdx.x * blockDim.x; base < n_slots; base += (int64_t)gridDim.x * blockDim.x) {
    __syncthreads();
    for (int r = warp; r < (int)blockDim.x; r += nwarps) {
      int64_t slot = base + r;
      if (slot < n_slots) {
        const float* src = y + slot * fpad;
        for (int j = lane; j < features; j += 32) s_tile[r * ld + j] = src[j];
      }
    }
    __syncthreads();
    int64_t slot = base + threadIdx.x;
    if (slot < n_slots) {
      int32_t row = slot_row[slot];
      if (row >= 0 && row < n_rows) {
        const float* v = s_tile + threadIdx.x * ld;
        double dot = 0.0, tot = 0.0;
        for (int j = 0; j < features; j++) {
          float yv = v[j];
          dot = exact_step(dot, __ldg(q + j), yv);
          if (kind == ORYX_COSINE) tot = exact_step(tot, yv, yv);
        }
        out[row] = (kind == ORYX_COSINE) ? __ddiv_rn(dot, __dsqrt_rn(tot)) : dot;
      }
    }
  }
}

// Point lookups (kernel K10): one thread per requested slot, exact arithmetic.
__global__ void scores_rows_kernel(const float* __restrict__ y, int fpad, int features,
                                   const float* __restrict__ q, int kind, const int32_t* __restrict__ slots,
                                   int n, double* __restrict__ out) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int32_t slot = slots[i];
  if (slot < 0) { out[i] = 0.0; return; }
  const float* v = y + (int64_t)slot * fpad;
  double dot = 0.0, tot = 0.0, qq = 0.0;
  for (int j = 0; j < features; j++) {
    float yv = v[j], qv = q[j];
    dot = exact_step(dot, qv, yv);
    if (kind == ORYX_COSINE) { tot = exact_step(tot, yv, yv); qq = exact_step(qq, qv, qv); }
  }
  // SimilarityToItem.java:62-63: dot / (toItemNorm * norm(item))
  out[i] = (kind == ORYX_COSINE) ? __ddiv_rn(dot, __dmul_rn(__dsqrt_rn(qq), __dsqrt_rn(tot))) : dot;
}

// Cross-shard merge (kernel K7, multi-GPU leg): rank-sort of n_lists*how_many exact (score,row) pairs
// by (score desc, global row asc) in one CTA.
__global__ void merge_topn_kernel(const double* __restrict__ scores, const int32_t* __restrict__ rows,
                                  const int32_t* __restrict__ counts, int n_lists, int how_many,
                                  int64_t row_stride, int64_t* __restrict__ out_rows,
                                  double* __restrict__ out_scores, int32_t* __restrict__ out_n) {
  const int total = n_lists * how_many;
  int valid_total = 0;
  for (int l = 0; l < n_lists; l++) valid_total += min(max(counts[l], 0), how_many);
  const int n_out = min(valid_total, how_many);
  for (int i = threadIdx.x; i < total; i += blockDim.x) {
    const int li = i / how_many, pi = i - li * how_many;
    if (pi >= counts[li]) continue;
    const double si = scores[i];
    const int64_t ri = (int64_t)li * row_stride + rows[i];
    int rank = 0;
    for (int j = 0; j < total; j++) {
      const int lj = j / how_many, pj = j - lj * how_many;
      if (pj >= counts[lj]) continue;
      const double sj = scores[j];
      const int64_t rj = (int64_t)lj * row_stride + rows[j];
      rank += ((sj > si) || (sj == si && rj < ri)) ? 1 : 0;
    }
    if (rank < n_out) { out_rows[rank] = ri; out_scores[rank] = si; }
  }
  if (threadIdx.x == 0) *out_n = n_out;
}

__global__ void fill_nan_kernel(double* p, int64_t n) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = __longlong_as_double(0x7ff8000000000000LL);
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
  if (lsh) {
    ORYX_TRY(lsh_hash_rows_device(m, rq.d_query, 1, s->fpad, ctx->d_main_index, stream));
  }
  ScanParams p{};
  p.y = reinterpret_cast<const float4*>(s->d_vec);
  p.slot_row = s->d_slot_row;
  p.slot_part = s->d_slot_part;
  p.q = rq.d_query;
  p.excl = rq.d_excl;
  p.n_excl = rq.n_excl;
  p.n_slots = s->n_slots;
  p.features = s->features;
  p.f4 = s->fpad / 4;
  p.k = rq.how_many + kTopnMargin;
  p.how_many = rq.how_many;
  p.max_bits = m->max_bits;
  p.main_index = ctx->d_main_index;
  p.cand = ctx->d_cand;
  p.counter = ctx->d_counter;
  p.out_rows = rq.d_out_rows;
  p.out_scores = rq.d_out_scores;
  p.out_n = rq.d_out_n;
  size_t smem = (size_t)p.k * 16 + 256 * sizeof(unsigned int);
  int grid = 0;
  int32_t rc;
  const int f4 = p.f4;
  if (f4 <= 8) rc = launch_variant<8, 1, 8>(m, p, rq.kind, lsh, smem, stream, &grid);
  else if (f4 <= 16) rc = launch_variant<8, 2, 4>(m, p, rq.kind, lsh, smem, stream, &grid);
  else if (f4 <= 24) rc = launch_variant<8, 3, 2>(m, p, rq.kind, lsh, smem, stream, &grid);
  else if (f4 <= 32) rc = launch_variant<8, 4, 2>(m, p, rq.kind, lsh, smem, stream, &grid);
  else if (f4 <= 48) rc = launch_variant<8, 6, 1>(m, p, rq.kind, lsh, smem, stream, &grid);
  else if (f4 <= 64) rc = launch_variant<8, 8, 1>(m, p, rq.kind, lsh, smem, stream, &grid);
  else if (f4 <= 128) rc = launch_variant<32, 4, 2>(m, p, rq.kind, lsh, smem, stream, &grid);
  else if (f4 <= 256) rc = launch_variant<32, 8, 1>(m, p, rq.kind, lsh, smem, stream, &grid);
  else if (f4 <= 512) rc = launch_variant<32, 16, 1>(m, p, rq.kind, lsh, smem, stream, &grid);
  else {
    set_error("features=%d beyond the scan kernel's envelope (2048)", s->features);
    return ORYX_E_UNSUPPORTED;
  }
  ORYX_TRY(rc);
  m->launches++;
  m->queries++;
  return ORYX_OK;
}

int32_t scores_all_launch(Model* m, int kind, const float* d_query, int32_t n_rows, double* d_out,
                          cudaStream_t stream) {
  Store* s = &m->store[ORYX_Y];
  int threads = 256;
  if (n_rows > 0) {
    fill_nan_kernel<<<(n_rows + threads - 1) / threads, threads, 0, stream>>>(d_out, n_rows);
    m->launches++;
  }
  if (s->n_slots == 0 || n_rows <= 0) return ORYX_OK;
  int tile = 128;
  while (tile > 32 && (size_t)tile * (s->features + 1) * sizeof(float) > 200 * 1024) tile -= 32;
  size_t smem = (size_t)tile * (s->features + 1) * sizeof(float);
  if (smem > 200 * 1024) {
    set_error("features=%d too large for the exact scoring kernel", s->features);
    return ORYX_E_UNSUPPORTED;
  }
  static thread_local size_t configured = 0;
  if (smem > 48 * 1024 && smem > configured) {
    ORYX_CUDA(cudaFuncSetAttribute(scores_all_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    configured = smem;
  }
  int64_t tiles = (s->n_slots + tile - 1) / tile;
  int grid = (int)std::min<int64_t>(tiles, (int64_t)m->num_sms * 8);
  scores_all_kernel<<<grid, tile, smem, stream>>>(s->d_vec, s->d_slot_row, s->n_slots, s->fpad, s->features,
                                                  d_query, kind, n_rows, d_out);
  m->launches++;
  ORYX_CUDA(cudaGetLastError());
  return ORYX_OK;
}

int32_t scores_rows_launch(Model* m, int kind, const float* d_query, const int32_t* d_slots, int n,
                           double* d_out, cudaStream_t stream) {
  Store* s = &m->store[ORYX_Y];
  scores_rows_kernel<<<(n + 127) / 128, 128, 0, stream>>>(s->d_vec, s->fpad, s->features, d_query, kind,
                                                         d_slots, n, d_out);
  m->launches++;
  ORYX_CUDA(cudaGetLastError());
  return ORYX_OK;
}

int32_t merge_topn_launch(const double* d_scores, const int32_t* d_rows, const int32_t* d_n, int n_lists,
                          int how_many, int64_t row_stride, int64_t* d_out_rows, double* d_out_scores,
                          int32_t* d_out_n, cudaStream_t stream) {
  merge_topn_kernel<<<1, 256, 0, stream>>>(d_scores, d_rows, d_n, n_lists, how_many, row_stride, d_out_rows,
                                           d_out_scores, d_out_n);
  ORYX_CUDA(cudaGetLastError());
  return ORYX_OK;
}

}  // namespace oryx
