// topn_deep.cu -- top-N for large N (how_many + offset beyond the fused scan's 1016, up to 65536).
//
// The reference puts no bound on howMany + offset (Recommend.java:81-82,108-112; paging in
// AbstractALSResource.java:37-43): its TopNConsumer heap simply grows (TopNConsumer.java:43-53).  The
// fused scan keeps its candidate lists in shared memory and stops at 1016; beyond that this path runs:
//
//   1. keys     every slot's EXACT score (the reference's float-multiply / double-add chain, as in
//               scores_all) mapped to an order-preserving 64-bit key; 0 for a slot that does not compete
//               (tombstone, excluded row, partition outside the query's LSH candidates, NaN score).  The
//               tie-break of the ranking rule (score descending, row ascending) becomes a 32-bit low word
//               ~row, so the 96-bit composite keys are all distinct.
//   2. select   most-significant-byte-first radix select of the how_many-th largest composite key: twelve
//               passes, each a 256-bin histogram of the keys that match the prefix found so far.
//   3. collect  the slots at or above that key -- exactly min(how_many, competing rows) of them.
//   4. sort     bitonic sort of the collected records, descending.
//   5. emit     rows and scores.
//
// Selection and ranking both run on exact values, so the result is the reference's list (ties between
// equal scores by ascending row, as everywhere in this library).  Cost is dominated by step 1 (binary64
// chains over every row, the same work as oryx_scores_all) -- tens of milliseconds at 10 M rows instead of
// an 80 MB copy of scores to the host; deep paging is not the hot path, the fused scan is.
#include <algorithm>

#include "common.cuh"

namespace oryx {

namespace {

struct DeepRec {
  unsigned long long hi;  // order-preserving key of the exact score; 0 = does not compete
  uint32_t lo;            // ~row: larger = smaller row
  int32_t slot;
};

struct DeepState {
  unsigned long long prefix_hi;
  uint32_t prefix_lo;
  uint32_t k_rem;     // rank still to resolve inside the current prefix
  uint32_t k_eff;     // min(how_many, competing rows)
  uint32_t n_collected;
  uint32_t hist[256];
};

__device__ __forceinline__ unsigned long long score_key(double s) {
  if (s != s) return 0ull;     // NaN never enters a top-N list (TopNConsumer.java:66-68 compares with >)
  if (s == 0.0) s = 0.0;       // -0.0 and +0.0 rank equal
  const unsigned long long b = (unsigned long long)__double_as_longlong(s);
  return (b & 0x8000000000000000ull) ? ~b : (b | 0x8000000000000000ull);
}
__device__ __forceinline__ double key_score(unsigned long long k) {
  const unsigned long long b = (k & 0x8000000000000000ull) ? (k & 0x7fffffffffffffffull) : ~k;
  return __longlong_as_double((long long)b);
}

__device__ __forceinline__ bool slot_excluded(const int32_t* excl, int n, int32_t slot) {
  int lo = 0, hi = n - 1;
  while (lo <= hi) {
    const int mid = (lo + hi) >> 1;
    const int32_t v = __ldg(excl + mid);
    if (v == slot) return true;
    if (v < slot) lo = mid + 1; else hi = mid - 1;
  }
  return false;
}

// One thread per slot; the row tile is staged in shared memory so global reads stay coalesced.
__global__ void deep_keys_kernel(const float* __restrict__ y, const int32_t* __restrict__ slot_row,
                                 const int32_t* __restrict__ slot_part, int64_t n_slots, int fpad, int features,
                                 const float* __restrict__ q, int kind, const int32_t* __restrict__ excl, int n_excl,
                                 const int32_t* __restrict__ main_index, int max_bits, int lsh,
                                 unsigned long long* __restrict__ keys) {
  extern __shared__ float s_tile[];  // [blockDim.x x (features+1)]
  const int ld = features + 1;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  const int main_idx = lsh ? *main_index : 0;
  for (int64_t base = (int64_t)blockIdx.x * blockDim.x; base < n_slots; base += (int64_t)gridDim.x * blockDim.x) {
    __syncthreads();
    for (int r = warp; r < (int)blockDim.x; r += nwarps) {
      const int64_t slot = base + r;
      if (slot < n_slots) {
        const float* src = y + slot * fpad;
        for (int j = lane; j < features; j += 32) s_tile[r * ld + j] = src[j];
      }
    }
    __syncthreads();
    const int64_t slot = base + threadIdx.x;
    if (slot >= n_slots) continue;
    unsigned long long key = 0ull;
    const int32_t row = slot_row[slot];
    bool competes = row >= 0;
    if (competes && lsh) {
      const int part = slot_part[slot];
      competes = part >= 0 && __popc(part ^ main_idx) <= max_bits;
    }
    if (competes && n_excl > 0) competes = !slot_excluded(excl, n_excl, (int32_t)slot);
    if (competes) {
      const float* v = s_tile + threadIdx.x * ld;
      double dot = 0.0, tot = 0.0;
      for (int j = 0; j < features; j++) {
        const float yv = v[j];
        dot = exact_step(dot, __ldg(q + j), yv);
        if (kind == ORYX_COSINE) tot = exact_step(tot, yv, yv);
      }
      key = score_key(kind == ORYX_COSINE ? __ddiv_rn(dot, __dsqrt_rn(tot)) : dot);
    }
    keys[slot] = key;
  }
}

// Byte `pass` (0 = most significant) of the 96-bit composite (hi, lo).
__device__ __forceinline__ uint32_t composite_byte(unsigned long long hi, uint32_t lo, int pass) {
  return pass < 8 ? (uint32_t)(hi >> (56 - 8 * pass)) & 255u : (lo >> (24 - 8 * (pass - 8))) & 255u;
}
// Do the bytes above `pass` equal the prefix?
__device__ __forceinline__ bool prefix_matches(unsigned long long hi, uint32_t lo, unsigned long long p_hi, uint32_t p_lo,
                                               int pass) {
  if (pass == 0) return true;
  if (pass <= 8) {
    const int sh = 64 - 8 * pass;
    return sh == 0 ? hi == p_hi : (hi >> sh) == (p_hi >> sh);
  }
  if (hi != p_hi) return false;
  const int sh = 32 - 8 * (pass - 8);
  return (lo >> sh) == (p_lo >> sh);
}

__global__ void deep_hist_kernel(const unsigned long long* __restrict__ keys, const int32_t* __restrict__ slot_row,
                                 int64_t n_slots, int pass, DeepState* __restrict__ st) {
  __shared__ uint32_t s_hist[256];
  for (int i = threadIdx.x; i < 256; i += blockDim.x) s_hist[i] = 0;
  __syncthreads();
  const unsigned long long p_hi = st->prefix_hi;
  const uint32_t p_lo = st->prefix_lo;
  for (int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; s < n_slots; s += (int64_t)gridDim.x * blockDim.x) {
    const unsigned long long hi = keys[s];
    if (hi == 0ull) continue;
    const uint32_t lo = pass >= 8 ? ~(uint32_t)__ldg(slot_row + s) : 0u;
    if (!prefix_matches(hi, lo, p_hi, p_lo, pass)) continue;
    atomicAdd(&s_hist[composite_byte(hi, lo, pass)], 1u);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < 256; i += blockDim.x)
    if (s_hist[i]) atomicAdd(&st->hist[i], s_hist[i]);
}

// One thread: pick the byte in which the k_rem-th largest key of the current prefix lies.
__global__ void deep_pick_kernel(int pass, uint32_t how_many, DeepState* __restrict__ st) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  if (pass == 0) {
    unsigned long long total = 0;
    for (int d = 0; d < 256; d++) total += st->hist[d];
    st->k_eff = (uint32_t)min((unsigned long long)how_many, total);
    st->k_rem = st->k_eff;
  }
  uint32_t k = st->k_rem;
  if (st->k_eff > 0) {
    for (int d = 255; d >= 0; d--) {
      const uint32_t c = st->hist[d];
      if (k <= c) {
        if (pass < 8) st->prefix_hi |= (unsigned long long)d << (56 - 8 * pass);
        else st->prefix_lo |= (uint32_t)d << (24 - 8 * (pass - 8));
        break;
      }
      k -= c;
    }
    st->k_rem = k;
  }
  for (int d = 0; d < 256; d++) st->hist[d] = 0;
}

// Every competing slot whose composite key is >= the selected one.
__global__ void deep_collect_kernel(const unsigned long long* __restrict__ keys, const int32_t* __restrict__ slot_row,
                                    int64_t n_slots, DeepState* __restrict__ st, DeepRec* __restrict__ out, uint32_t cap) {
  const unsigned long long t_hi = st->prefix_hi;
  const uint32_t t_lo = st->prefix_lo;
  if (st->k_eff == 0) return;
  for (int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; s < n_slots; s += (int64_t)gridDim.x * blockDim.x) {
    const unsigned long long hi = keys[s];
    if (hi == 0ull || hi < t_hi) continue;
    const uint32_t lo = ~(uint32_t)__ldg(slot_row + s);
    if (hi == t_hi && lo < t_lo) continue;
    const uint32_t pos = atomicAdd(&st->n_collected, 1u);
    if (pos < cap) out[pos] = DeepRec{hi, lo, (int32_t)s};
  }
}

__device__ __forceinline__ bool rec_less(const DeepRec& a, const DeepRec& b) {
  return a.hi < b.hi || (a.hi == b.hi && a.lo < b.lo);
}

// One compare-exchange step of a bitonic sort, descending.
__global__ void deep_bitonic_kernel(DeepRec* __restrict__ recs, uint32_t n, uint32_t k2, uint32_t j) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t ixj = i ^ j;
  if (ixj <= i) return;
  const DeepRec a = recs[i], b = recs[ixj];
  const bool desc = (i & k2) == 0;
  if (rec_less(a, b) == desc) { recs[i] = b; recs[ixj] = a; }
}

__global__ void deep_pad_kernel(DeepRec* __restrict__ recs, const DeepState* __restrict__ st, uint32_t n) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n && i >= st->k_eff) recs[i] = DeepRec{0ull, 0u, -1};
}

__global__ void deep_emit_kernel(const DeepRec* __restrict__ recs, const DeepState* __restrict__ st, int32_t* __restrict__ out_rows,
                                 double* __restrict__ out_scores, int32_t* __restrict__ out_n) {
  const uint32_t n = st->k_eff;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    out_rows[i] = (int32_t)~recs[i].lo;
    out_scores[i] = key_score(recs[i].hi);
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) *out_n = (int32_t)n;
}

}  // namespace

// d_query: [fpad] device; d_excl: sorted exclusion SLOTS (device); outputs: device arrays of how_many entries.
// `d_scratch` must hold deep_scratch_bytes(n_slots, how_many).  Caller holds the model's shared lock.
size_t deep_scratch_bytes(int64_t n_slots, int how_many) {
  uint32_t p2 = 1;
  while ((int)p2 < how_many) p2 <<= 1;
  return 256 + (size_t)n_slots * 8 + (size_t)p2 * sizeof(DeepRec) + 1024;
}

int32_t topn_deep_launch(Model* m, QueryCtx* ctx, int kind, const float* d_query, int how_many, const int32_t* d_excl,
                         int n_excl, void* d_scratch, int32_t* d_out_rows, double* d_out_scores, int32_t* d_out_n,
                         cudaStream_t stream) {
  Store* s = &m->store[ORYX_Y];
  ORYX_ARG(how_many >= 1 && how_many <= ORYX_MAX_HOW_MANY_DEEP, "how_many out of range [1, ORYX_MAX_HOW_MANY_DEEP]");
  if (s->n_slots == 0) {
    ORYX_CUDA(cudaMemsetAsync(d_out_n, 0, sizeof(int32_t), stream));
    return ORYX_OK;
  }
  uint32_t p2 = 1;
  while ((int)p2 < how_many) p2 <<= 1;
  unsigned char* base = reinterpret_cast<unsigned char*>(d_scratch);
  DeepState* st = reinterpret_cast<DeepState*>(base);
  unsigned long long* keys = reinterpret_cast<unsigned long long*>(base + ((sizeof(DeepState) + 255) & ~(size_t)255));
  DeepRec* recs = reinterpret_cast<DeepRec*>(reinterpret_cast<unsigned char*>(keys) + (((size_t)s->n_slots * 8 + 255) & ~(size_t)255));
  ORYX_CUDA(cudaMemsetAsync(st, 0, sizeof(DeepState), stream));

  const bool lsh = s->hashed && m->num_hashes != m->max_bits;
  if (lsh) ORYX_TRY(lsh_hash_rows_device(m, d_query, 1, s->fpad, ctx->d_main_index, stream));
  int tile = 128;
  while (tile > 32 && (size_t)tile * (s->features + 1) * sizeof(float) > 200 * 1024) tile -= 32;
  const size_t smem = (size_t)tile * (s->features + 1) * sizeof(float);
  if (smem > 200 * 1024) {
    set_error("features=%d too large for the exact scoring kernel", s->features);
    return ORYX_E_UNSUPPORTED;
  }
  if (smem > 48 * 1024) ORYX_TRY(ensure_max_dynamic_smem(reinterpret_cast<const void*>(deep_keys_kernel), smem));
  const int64_t tiles = (s->n_slots + tile - 1) / tile;
  const int kgrid = (int)std::min<int64_t>(tiles, (int64_t)m->num_sms * 8);
  deep_keys_kernel<<<kgrid, tile, smem, stream>>>(s->d_vec, s->d_slot_row, s->d_slot_part, s->n_slots, s->fpad, s->features,
                                                  d_query, kind, d_excl, n_excl, ctx->d_main_index, m->max_bits, lsh ? 1 : 0,
                                                  keys);
  ORYX_CUDA(cudaGetLastError());
  const int sgrid = (int)std::min<int64_t>((s->n_slots + 255) / 256, (int64_t)m->num_sms * 8);
  for (int pass = 0; pass < 12; pass++) {
    deep_hist_kernel<<<sgrid, 256, 0, stream>>>(keys, s->d_slot_row, s->n_slots, pass, st);
    deep_pick_kernel<<<1, 32, 0, stream>>>(pass, (uint32_t)how_many, st);
  }
  deep_collect_kernel<<<sgrid, 256, 0, stream>>>(keys, s->d_slot_row, s->n_slots, st, recs, p2);
  deep_pad_kernel<<<(p2 + 255) / 256, 256, 0, stream>>>(recs, st, p2);
  for (uint32_t k2 = 2; k2 <= p2; k2 <<= 1)
    for (uint32_t j = k2 >> 1; j > 0; j >>= 1) deep_bitonic_kernel<<<(p2 + 255) / 256, 256, 0, stream>>>(recs, p2, k2, j);
  deep_emit_kernel<<<std::max(1u, std::min(64u, (p2 + 255) / 256)), 256, 0, stream>>>(recs, st, d_out_rows, d_out_scores, d_out_n);
  ORYX_CUDA(cudaGetLastError());
  int sort_steps = 0;
  for (uint32_t k2 = 2; k2 <= p2; k2 <<= 1)
    for (uint32_t j = k2 >> 1; j > 0; j >>= 1) sort_steps++;
  m->launches += 1 + 24 + 3 + sort_steps + (lsh ? 1 : 0);
  m->queries++;
  return ORYX_OK;
}

}  // namespace oryx
