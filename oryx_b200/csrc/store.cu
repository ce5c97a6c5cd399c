// store.cu -- device-resident feature-vector store: the B200 counterpart of the reference's
// FeatureVectors (app/oryx-app-common/.../als/FeatureVectors.java:36-154) and of the partitioned
// `FeatureVectors[] Y` of ALSServingModel (.../serving/als/model/ALSServingModel.java:69,161-183).
//
// The reference keeps one heap-allocated float[] per ID in a hash map per LSH partition and moves an
// item between maps when its hash changes.  Here all rows of a store live in one row-major fp32
// matrix in HBM; partition membership is a per-slot int32 (`slot_part`) that the scan kernel tests,
// so "moving" a row between partitions is an in-place update of that integer.  oryx_items_compact
// re-orders slots so rows of one partition are contiguous, which only affects scan locality.
#include <algorithm>
#include <cstdlib>
#include <numeric>

#include "common.cuh"

namespace oryx {

namespace {

constexpr int64_t kChunkRows = 1 << 18;

// One warp per incoming row: copy `features` floats to the destination slot, zero the padding,
// record the row id and partition.
__global__ void scatter_rows_kernel(const float* __restrict__ in, int64_t in_stride, int features,
                                    int fpad, const int32_t* __restrict__ dst_slot,
                                    const int32_t* __restrict__ rows, const int32_t* __restrict__ parts,
                                    int64_t n, float* __restrict__ vec, int32_t* __restrict__ slot_row,
                                    int32_t* __restrict__ slot_part, int32_t* __restrict__ row_slot,
                                    unsigned int* __restrict__ absmax_bits) {
  int64_t w = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  int lane = threadIdx.x & 31;
  if (w >= n) return;
  int32_t slot = dst_slot[w];
  if (slot < 0) return;  // superseded by a later occurrence of the same row in this batch
  const float* src = in + w * in_stride;
  float* dst = vec + (int64_t)slot * fpad;
  float amax = 0.f;
  for (int j = lane; j < fpad; j += 32) {
    const float v = j < features ? src[j] : 0.0f;
    dst[j] = v;
    amax = fmaxf(amax, fabsf(v));  // fmaxf drops NaN
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
  // the running maximum settles after the first few rows: nearly every warp only reads it
  if (lane == 0 && __float_as_uint(amax) > *(volatile unsigned int*)absmax_bits) atomicMax(absmax_bits, __float_as_uint(amax));
  if (lane == 0) {
    slot_row[slot] = rows[w];
    slot_part[slot] = parts ? parts[w] : 0;
    if (row_slot) row_slot[rows[w]] = slot;
  }
}

__global__ void clear_slots_kernel(const int32_t* __restrict__ slots, int64_t n, int fpad,
                                   float* __restrict__ vec, int32_t* __restrict__ slot_row,
                                   int32_t* __restrict__ slot_part, int32_t* __restrict__ row_slot) {
  int64_t w = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  int lane = threadIdx.x & 31;
  if (w >= n) return;
  int32_t slot = slots[w];
  // A tombstone is all-NaN: whatever the query, its score is NaN and "score > threshold" is false, so
  // the scan never has to look at slot_row on its hot path (TopNConsumer never admits NaN either).
  float* dst = vec + (int64_t)slot * fpad;
  const float qnan = __int_as_float(0x7fc00000);
  for (int j = lane; j < fpad; j += 32) dst[j] = qnan;
  if (lane == 0) {
    const int32_t row = slot_row[slot];
    if (row_slot && row >= 0) row_slot[row] = -1;
    slot_row[slot] = -1;
    slot_part[slot] = -1;
  }
}

// One warp per 256-slot block: summarise the partitions of its live rows.
__global__ void block_part_kernel(const int32_t* __restrict__ slot_part, int64_t n_slots, int64_t n_blocks,
                                  int32_t* __restrict__ block_part) {
  const int64_t b = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (b >= n_blocks) return;
  int mn = 0x7fffffff, mx = -1;
  for (int i = lane; i < kBlockRows; i += 32) {
    const int64_t slot = b * kBlockRows + i;
    const int part = slot < n_slots ? slot_part[slot] : -1;
    if (part >= 0) { mn = min(mn, part); mx = max(mx, part); }
  }
  mn = __reduce_min_sync(0xffffffffu, mn);
  mx = __reduce_max_sync(0xffffffffu, mx);
  if (lane == 0) block_part[b] = mx < 0 ? -1 : (mn == mx ? mn : -2);
}

__global__ void fill_i32_kernel(int32_t* p, int64_t n, int32_t v) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}

// row_slot[slot_row[i]] = i for every live slot (after compaction)
__global__ void rebuild_row_slot_kernel(const int32_t* __restrict__ slot_row, int64_t n, int32_t* __restrict__ row_slot) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n && slot_row[i] >= 0) row_slot[slot_row[i]] = (int32_t)i;
}

__global__ void iota_rows_kernel(int32_t* p, int64_t n, int32_t first) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = first + (int32_t)i;
}

// Gather for compaction: new slot i takes old slot src[i].
__global__ void gather_slots_kernel(const int32_t* __restrict__ src, int64_t n, int fpad,
                                    const float4* __restrict__ old_vec,
                                    const int32_t* __restrict__ old_row,
                                    const int32_t* __restrict__ old_part, float4* __restrict__ new_vec,
                                    int32_t* __restrict__ new_row, int32_t* __restrict__ new_part) {
  int64_t w = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  int lane = threadIdx.x & 31;
  if (w >= n) return;
  int32_t s = src[w];
  int f4 = fpad >> 2;
  for (int j = lane; j < f4; j += 32) new_vec[w * f4 + j] = old_vec[(int64_t)s * f4 + j];
  if (lane == 0) {
    new_row[w] = old_row[s];
    new_part[w] = old_part[s];
  }
}

// ---- in-place compaction (store_compact when a second copy of the store does not fit) ----
// The permutation "slot src[i] moves to slot i" decomposes into paths and cycles; each is cut into
// segments of at most kSegMoves moves.  pos[] lists the slots of every segment in move order
// (pos[j] moves to pos[j+1]); seg_begin[k] .. seg_begin[k+1] are segment k's entries (moves + 1 of them).
// Pass 1 saves the row in each segment's FIRST slot (the previous segment overwrites it); pass 2 walks
// every segment backwards, so a slot is always read before the one write that lands in it, and the
// segment's first row comes from the saved copy.  One warp per segment.
constexpr int kSegMoves = 32;

__global__ void compact_save_kernel(const float4* __restrict__ vec, int f4, const int32_t* __restrict__ pos,
                                    const int64_t* __restrict__ seg_begin, int64_t n_seg, float4* __restrict__ saved) {
  const int64_t k = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (k >= n_seg) return;
  const int64_t first = pos[seg_begin[k]];
  for (int j = lane; j < f4; j += 32) saved[k * f4 + j] = vec[first * f4 + j];
}

__global__ void compact_move_kernel(float4* __restrict__ vec, int f4, const int32_t* __restrict__ pos,
                                    const int64_t* __restrict__ seg_begin, int64_t n_seg,
                                    const float4* __restrict__ saved) {
  const int64_t k = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (k >= n_seg) return;
  const int64_t b = seg_begin[k], e = seg_begin[k + 1];  // entries b .. e-1, moves b->b+1 ... e-2->e-1
  for (int64_t j = e - 2; j >= b; j--) {
    const int64_t from = pos[j], to = pos[j + 1];
    // each lane handles the same columns at every step: its read of `from` here precedes, in program
    // order, its write to `from` at the next step
    for (int c = lane; c < f4; c += 32) vec[to * f4 + c] = (j == b) ? saved[k * f4 + c] : vec[from * f4 + c];
  }
}

// The store's device arrays moved (or went away): peers that mapped them (peer_stores.cu) must notice.
void layout_changed(Model* m, Store* s) {
  s->layout_gen++;
  if (m->d_share_desc) {
    const int which = (int)(s - &m->store[0]);
    cudaMemcpy(m->d_share_desc + which, &s->layout_gen, sizeof(unsigned long long), cudaMemcpyHostToDevice);
    cudaGetLastError();
  }
}

int32_t grow(Model* m, Store* s, int64_t need) {
  if (need <= s->cap_slots) return ORYX_OK;
  int64_t cap = std::max<int64_t>(need, s->cap_slots + s->cap_slots / 2);
  cap = std::max<int64_t>(cap, 1024);
  float* nv = nullptr;
  int32_t *nr = nullptr, *np = nullptr, *nbp = nullptr;
  cudaError_t e = cudaMalloc(&nv, (size_t)cap * s->fpad * sizeof(float));
  if (e == cudaSuccess) e = cudaMalloc(&nr, (size_t)cap * sizeof(int32_t));
  if (e == cudaSuccess) e = cudaMalloc(&np, (size_t)cap * sizeof(int32_t));
  if (e == cudaSuccess) e = cudaMalloc(&nbp, (size_t)(cap / kBlockRows + 1) * sizeof(int32_t));
  if (e != cudaSuccess) {
    cudaFree(nv); cudaFree(nr); cudaFree(np); cudaFree(nbp);
    cudaGetLastError();
    set_error("device allocation of %lld slots x %d floats failed: %s", (long long)cap, s->fpad,
              cudaGetErrorString(e));
    return ORYX_E_NOMEM;
  }
  cudaStream_t st = m->update_stream;
  if (s->n_slots > 0) {
    ORYX_CUDA(cudaMemcpyAsync(nv, s->d_vec, (size_t)s->n_slots * s->fpad * sizeof(float),
                              cudaMemcpyDeviceToDevice, st));
    ORYX_CUDA(cudaMemcpyAsync(nr, s->d_slot_row, (size_t)s->n_slots * sizeof(int32_t),
                              cudaMemcpyDeviceToDevice, st));
    ORYX_CUDA(cudaMemcpyAsync(np, s->d_slot_part, (size_t)s->n_slots * sizeof(int32_t),
                              cudaMemcpyDeviceToDevice, st));
  }
  int64_t tail = cap - s->n_slots;
  if (tail > 0) {
    int threads = 256;
    int64_t blocks = (tail + threads - 1) / threads;
    fill_i32_kernel<<<(unsigned)blocks, threads, 0, st>>>(nr + s->n_slots, tail, -1);
    fill_i32_kernel<<<(unsigned)blocks, threads, 0, st>>>(np + s->n_slots, tail, -1);
    m->launches += 2;
    ORYX_CUDA(cudaMemsetAsync(nv + s->n_slots * s->fpad, 0, (size_t)tail * s->fpad * sizeof(float), st));
  }
  ORYX_CUDA(cudaStreamSynchronize(st));
  cudaFree(s->d_vec); cudaFree(s->d_slot_row); cudaFree(s->d_slot_part); cudaFree(s->d_block_part);
  s->d_vec = nv; s->d_slot_row = nr; s->d_slot_part = np; s->d_block_part = nbp;
  s->cap_slots = cap;
  layout_changed(m, s);
  return ORYX_OK;
}

// Recompute the per-block partition summaries after a mutation (cheap: reads 4 bytes per slot).
int32_t refresh_block_parts(Model* m, Store* s) {
  if (s->n_slots == 0) return ORYX_OK;
  const int64_t n_blocks = (s->n_slots + kBlockRows - 1) / kBlockRows;
  const int threads = 256;
  const int64_t blocks = (n_blocks * 32 + threads - 1) / threads;
  block_part_kernel<<<(unsigned)blocks, threads, 0, m->update_stream>>>(s->d_slot_part, s->n_slots, n_blocks,
                                                                        s->d_block_part);
  m->launches++;
  ORYX_CUDA(cudaGetLastError());
  return ORYX_OK;
}

// Grow-only staging used by the (single) updater thread of the model.
using Staging = UpdateStaging;

int32_t staging_reserve(Model* m, int64_t rows, int features, bool need_vec) {
  Staging& g = m->staging;
  size_t vb = need_vec ? (size_t)rows * features * sizeof(float) : 0;
  if (vb > g.vec_bytes) {
    cudaFree(g.d_vec);
    g.d_vec = nullptr; g.vec_bytes = 0;
    ORYX_CUDA(cudaMalloc(&g.d_vec, vb));
    g.vec_bytes = vb;
  }
  size_t ic = (size_t)rows * 3;
  if (ic > g.i32_count) {
    cudaFree(g.d_i32);
    g.d_i32 = nullptr; g.i32_count = 0;
    ORYX_CUDA(cudaMalloc(&g.d_i32, ic * sizeof(int32_t)));
    g.i32_count = ic;
  }
  if (ic > g.h_count) {
    cudaFreeHost(g.h_i32);
    g.h_i32 = nullptr; g.h_count = 0;
    ORYX_CUDA(cudaMallocHost(&g.h_i32, ic * sizeof(int32_t)));
    g.h_count = ic;
  }
  return ORYX_OK;
}

// Host-side slot assignment for rows[0..n): writes dst slots into h_dst (or -1 for an occurrence
// superseded later in the same chunk).  Returns the new high-water mark through the store.
int32_t assign_slots(Model* m, Store* s, const int32_t* rows, int64_t n, int32_t first_row,
                     int32_t* h_dst) {
  int32_t max_row = -1;
  for (int64_t k = 0; k < n; k++) {
    int32_t r = rows ? rows[k] : first_row + (int32_t)k;
    if (r < 0) {
      set_error("negative row id %d at position %lld", r, (long long)k);
      return ORYX_E_INVALID;
    }
    max_row = std::max(max_row, r);
  }
  if ((int64_t)max_row >= (int64_t)s->row2slot.size()) {
    s->row2slot.resize((size_t)max_row + 1 + (size_t)max_row / 4, -1);
  }
  if (s->track_row_slot && (int64_t)s->row2slot.size() > s->row_slot_cap) {
    const int64_t cap = (int64_t)s->row2slot.size();
    int32_t* nrs = nullptr;
    if (cudaMalloc(&nrs, (size_t)cap * sizeof(int32_t)) != cudaSuccess) {
      cudaGetLastError();
      set_error("device allocation of the row->slot map (%lld rows) failed", (long long)cap);
      return ORYX_E_NOMEM;
    }
    cudaStream_t st = m->update_stream;
    if (s->row_slot_cap > 0)
      ORYX_CUDA(cudaMemcpyAsync(nrs, s->d_row_slot, (size_t)s->row_slot_cap * sizeof(int32_t), cudaMemcpyDeviceToDevice, st));
    const int64_t tail = cap - s->row_slot_cap;
    fill_i32_kernel<<<(unsigned)((tail + 255) / 256), 256, 0, st>>>(nrs + s->row_slot_cap, tail, -1);
    m->launches++;
    ORYX_CUDA(cudaStreamSynchronize(st));
    cudaFree(s->d_row_slot);
    s->d_row_slot = nrs;
    s->row_slot_cap = cap;
    layout_changed(m, s);
  }
  // Count new slots needed first so the device buffers can grow before any kernel runs.
  int64_t fresh = 0;
  for (int64_t k = 0; k < n; k++) {
    int32_t r = rows ? rows[k] : first_row + (int32_t)k;
    if (s->row2slot[r] < 0) fresh++;
  }
  int64_t from_free = std::min<int64_t>(fresh, (int64_t)s->free_slots.size());
  ORYX_TRY(grow(m, s, s->n_slots + (fresh - from_free)));
  // last occurrence wins inside a chunk: remember where each slot was last written
  std::vector<int64_t> slot_writer;  // lazily sized only when a duplicate shows up
  for (int64_t k = 0; k < n; k++) {
    int32_t r = rows ? rows[k] : first_row + (int32_t)k;
    int32_t slot = s->row2slot[r];
    if (slot < 0) {
      if (!s->free_slots.empty()) {
        slot = s->free_slots.back();
        s->free_slots.pop_back();
      } else {
        slot = (int32_t)s->n_slots++;
      }
      s->row2slot[r] = slot;
      s->live++;
    }
    h_dst[k] = slot;
  }
  if (rows) {
    // Detect duplicates: same slot targeted twice in this chunk.
    std::vector<int32_t> order((size_t)n);
    std::iota(order.begin(), order.end(), 0);
    std::stable_sort(order.begin(), order.end(), [&](int32_t a, int32_t b) { return h_dst[a] < h_dst[b]; });
    for (int64_t i = 0; i + 1 < n; i++) {
      if (h_dst[order[i]] == h_dst[order[i + 1]]) h_dst[order[i]] = -1;  // earlier occurrence loses
    }
  }
  return ORYX_OK;
}

}  // namespace

int32_t store_init(Model* m, int which, int64_t capacity_hint) {
  Store* s = &m->store[which];
  s->features = m->cfg.features;
  s->fpad = round_up(m->cfg.features, 4);
  s->hashed = (which == ORYX_Y) && m->num_hashes > 0;
  s->track_row_slot = true;  // Y: known-items exclusion lists; both: rows gathered by peers (peer_stores.cu)
  ORYX_CUDA(cudaMalloc(&s->d_absmax_bits, sizeof(unsigned int)));
  ORYX_CUDA(cudaMemset(s->d_absmax_bits, 0, sizeof(unsigned int)));
  if (capacity_hint > 0) ORYX_TRY(grow(m, s, capacity_hint));
  return ORYX_OK;
}

void store_free(Store* s) {
  cudaFree(s->d_vec); cudaFree(s->d_slot_row); cudaFree(s->d_slot_part); cudaFree(s->d_block_part);
  cudaFree(s->d_row_slot); cudaFree(s->d_absmax_bits);
  s->d_absmax_bits = nullptr;
  s->d_vec = nullptr; s->d_slot_row = nullptr; s->d_slot_part = nullptr; s->d_block_part = nullptr;
  s->d_row_slot = nullptr; s->row_slot_cap = 0;
  s->cap_slots = s->n_slots = s->live = 0;
  s->row2slot.clear(); s->free_slots.clear();
}

int32_t store_upsert(Model* m, int which, const int32_t* rows, const float* vecs, int32_t n,
                     int32_t* out_partitions) {
  Store* s = &m->store[which];
  const uint64_t version_before = s->version;
  s->version++;
  bool shadow_ok = which == ORYX_Y;  // keep the batch path's bf16 shadow current if it was (batch_topn.cu)
  cudaStream_t st = m->update_stream;
  const int F = s->features;
  for (int64_t off = 0; off < n; off += kChunkRows) {
    int64_t c = std::min<int64_t>(kChunkRows, n - off);
    ORYX_TRY(staging_reserve(m, c, F, true));
    Staging& g = m->staging;
    int32_t* h_dst = g.h_i32;
    int32_t* h_rows = g.h_i32 + c;
    int32_t* h_parts = g.h_i32 + 2 * c;
    int32_t* d_dst = g.d_i32;
    int32_t* d_rows = g.d_i32 + c;
    int32_t* d_parts = g.d_i32 + 2 * c;
    ORYX_TRY(assign_slots(m, s, rows + off, c, 0, h_dst));
    std::memcpy(h_rows, rows + off, (size_t)c * sizeof(int32_t));
    ORYX_CUDA(cudaMemcpyAsync(g.d_vec, vecs + off * F, (size_t)c * F * sizeof(float),
                              cudaMemcpyHostToDevice, st));
    ORYX_CUDA(cudaMemcpyAsync(d_dst, h_dst, (size_t)c * 2 * sizeof(int32_t), cudaMemcpyHostToDevice, st));
    if (s->hashed) {
      ORYX_TRY(lsh_hash_rows_device(m, g.d_vec, c, F, d_parts, st));
      if (out_partitions) {
        ORYX_CUDA(cudaMemcpyAsync(h_parts, d_parts, (size_t)c * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
      }
    }
    int threads = 256;
    int64_t blocks = (c * 32 + threads - 1) / threads;
    scatter_rows_kernel<<<(unsigned)blocks, threads, 0, st>>>(
        g.d_vec, F, F, s->fpad, d_dst, d_rows, s->hashed ? d_parts : nullptr, c, s->d_vec,
        s->d_slot_row, s->d_slot_part, s->d_row_slot, s->d_absmax_bits);
    m->launches++;
    ORYX_CUDA(cudaGetLastError());
    if (shadow_ok) shadow_ok = batch_shadow_rows_changed(m, d_dst, c, version_before, st);
    ORYX_CUDA(cudaStreamSynchronize(st));
    if (out_partitions) {
      if (s->hashed) std::memcpy(out_partitions + off, h_parts, (size_t)c * sizeof(int32_t));
      else std::fill(out_partitions + off, out_partitions + off + c, 0);
    }
  }
  if (s->hashed) {
    ORYX_TRY(refresh_block_parts(m, s));
    ORYX_CUDA(cudaStreamSynchronize(st));
  }
  if (shadow_ok) m->batch.y_version = s->version;
  return ORYX_OK;
}

int32_t store_upsert_device(Model* m, int which, int32_t first_row, const float* d_vecs, int64_t n) {
  Store* s = &m->store[which];
  const uint64_t version_before = s->version;
  s->version++;
  bool shadow_ok = which == ORYX_Y;
  cudaStream_t st = m->update_stream;
  const int F = s->features;
  const int64_t chunk = 1 << 20;
  for (int64_t off = 0; off < n; off += chunk) {
    int64_t c = std::min<int64_t>(chunk, n - off);
    ORYX_TRY(staging_reserve(m, c, F, false));
    Staging& g = m->staging;
    int32_t* h_dst = g.h_i32;
    int32_t* d_dst = g.d_i32;
    int32_t* d_rows = g.d_i32 + c;
    int32_t* d_parts = g.d_i32 + 2 * c;
    ORYX_TRY(assign_slots(m, s, nullptr, c, first_row + (int32_t)off, h_dst));
    ORYX_CUDA(cudaMemcpyAsync(d_dst, h_dst, (size_t)c * sizeof(int32_t), cudaMemcpyHostToDevice, st));
    int threads = 256;
    iota_rows_kernel<<<(unsigned)((c + threads - 1) / threads), threads, 0, st>>>(d_rows, c, first_row + (int32_t)off);
    m->launches++;
    const float* src = d_vecs + off * F;
    if (s->hashed) ORYX_TRY(lsh_hash_rows_device(m, src, c, F, d_parts, st));
    int64_t blocks = (c * 32 + threads - 1) / threads;
    scatter_rows_kernel<<<(unsigned)blocks, threads, 0, st>>>(
        src, F, F, s->fpad, d_dst, d_rows, s->hashed ? d_parts : nullptr, c, s->d_vec,
        s->d_slot_row, s->d_slot_part, s->d_row_slot, s->d_absmax_bits);
    m->launches++;
    ORYX_CUDA(cudaGetLastError());
    if (shadow_ok) shadow_ok = batch_shadow_rows_changed(m, d_dst, c, version_before, st);
    ORYX_CUDA(cudaStreamSynchronize(st));
  }
  if (s->hashed) {
    ORYX_TRY(refresh_block_parts(m, s));
    ORYX_CUDA(cudaStreamSynchronize(st));
  }
  if (shadow_ok) m->batch.y_version = s->version;
  return ORYX_OK;
}

int32_t store_remove(Model* m, int which, const int32_t* rows, int32_t n) {
  Store* s = &m->store[which];
  const uint64_t version_before = s->version;
  s->version++;
  cudaStream_t st = m->update_stream;
  ORYX_TRY(staging_reserve(m, std::max(n, 1), s->features, false));
  Staging& g = m->staging;
  int32_t cnt = 0;
  for (int32_t k = 0; k < n; k++) {
    int32_t r = rows[k];
    if (r < 0 || (size_t)r >= s->row2slot.size()) continue;
    int32_t slot = s->row2slot[r];
    if (slot < 0) continue;  // FeatureVectors.removeVector on an absent id is a no-op (FV:84-89)
    s->row2slot[r] = -1;
    s->free_slots.push_back(slot);
    s->live--;
    g.h_i32[cnt++] = slot;
  }
  if (cnt == 0) {
    if (which == ORYX_Y && m->batch.y_version == version_before) m->batch.y_version = s->version;  // nothing changed
    return ORYX_OK;
  }
  ORYX_CUDA(cudaMemcpyAsync(g.d_i32, g.h_i32, (size_t)cnt * sizeof(int32_t), cudaMemcpyHostToDevice, st));
  int threads = 256;
  int64_t blocks = ((int64_t)cnt * 32 + threads - 1) / threads;
  clear_slots_kernel<<<(unsigned)blocks, threads, 0, st>>>(g.d_i32, cnt, s->fpad, s->d_vec,
                                                          s->d_slot_row, s->d_slot_part, s->d_row_slot);
  m->launches++;
  ORYX_CUDA(cudaGetLastError());
  // tombstones are NaN rows in the shadow as well
  const bool shadow_ok = which == ORYX_Y && batch_shadow_rows_changed(m, g.d_i32, cnt, version_before, st);
  if (s->hashed) ORYX_TRY(refresh_block_parts(m, s));
  ORYX_CUDA(cudaStreamSynchronize(st));
  if (shadow_ok) m->batch.y_version = s->version;
  return ORYX_OK;
}

int32_t store_get(Model* m, int which, int32_t row, float* out, int32_t* found) {
  Store* s = &m->store[which];
  *found = 0;
  if (row < 0 || (size_t)row >= s->row2slot.size()) return ORYX_OK;
  int32_t slot = s->row2slot[row];
  if (slot < 0) return ORYX_OK;
  ORYX_CUDA(cudaMemcpy(out, s->d_vec + (int64_t)slot * s->fpad, (size_t)s->features * sizeof(float),
                       cudaMemcpyDeviceToHost));
  *found = 1;
  return ORYX_OK;
}

namespace {

// Moves the rows of `s` so that slot i holds what slot src[i] held (i < live), without a second copy of
// the matrix: see compact_save_kernel / compact_move_kernel.  Extra device memory: one row per segment
// (1/32 of the matrix) plus 4 bytes per slot.
int32_t compact_rows_in_place(Model* m, Store* s, const std::vector<int32_t>& src, int64_t live,
                              const std::vector<uint8_t>& is_live) {
  const int64_t n = s->n_slots;
  std::vector<int32_t> dst((size_t)n, -1);
  for (int64_t i = 0; i < live; i++) dst[(size_t)src[(size_t)i]] = (int32_t)i;
  std::vector<int32_t> pos;
  std::vector<int64_t> seg_begin{0};
  pos.reserve((size_t)(n + n / kSegMoves + 16));
  int moves_in_seg = 0;
  auto begin_chain = [&](int32_t p0) { pos.push_back(p0); moves_in_seg = 0; };
  auto push = [&](int32_t p) {
    pos.push_back(p);
    if (++moves_in_seg == kSegMoves) {  // close the segment; the next one starts at the same slot
      seg_begin.push_back((int64_t)pos.size());
      pos.push_back(p);
      moves_in_seg = 0;
    }
  };
  auto end_chain = [&]() {
    if (moves_in_seg > 0) seg_begin.push_back((int64_t)pos.size());
    else pos.pop_back();  // a segment start without moves
  };
  std::vector<uint8_t> visited((size_t)n, 0);
  // paths: start at a live slot beyond the compacted range, end in a hole of that range
  for (int64_t start = live; start < n; start++) {
    if (!is_live[(size_t)start]) continue;
    begin_chain((int32_t)start);
    int32_t cur = (int32_t)start;
    for (;;) {
      visited[(size_t)cur] = 1;
      const int32_t d = dst[(size_t)cur];
      push(d);
      if (!is_live[(size_t)d]) break;  // a hole: the path ends, nothing has to leave d
      cur = d;
    }
    end_chain();
  }
  // cycles among the live slots of the compacted range that no path went through
  for (int64_t start = 0; start < live; start++) {
    if (!is_live[(size_t)start] || visited[(size_t)start] || dst[(size_t)start] == (int32_t)start) continue;
    begin_chain((int32_t)start);
    int32_t cur = (int32_t)start;
    do {
      visited[(size_t)cur] = 1;
      cur = dst[(size_t)cur];
      push(cur);
    } while (cur != (int32_t)start);
    end_chain();
  }
  const int64_t n_seg = (int64_t)seg_begin.size() - 1;
  if (n_seg == 0) return ORYX_OK;  // already in order
  const int f4 = s->fpad / 4;
  int32_t* d_pos = nullptr;
  int64_t* d_seg = nullptr;
  float4* d_saved = nullptr;
  cudaError_t e = cudaMalloc(&d_pos, pos.size() * sizeof(int32_t));
  if (e == cudaSuccess) e = cudaMalloc(&d_seg, seg_begin.size() * sizeof(int64_t));
  if (e == cudaSuccess) e = cudaMalloc(&d_saved, (size_t)n_seg * s->fpad * sizeof(float));
  if (e != cudaSuccess) {
    cudaFree(d_pos); cudaFree(d_seg); cudaFree(d_saved);
    cudaGetLastError();
    set_error("in-place compaction needs %.1f MB of scratch: %s",
              ((double)n_seg * s->fpad * 4 + pos.size() * 4.0) / 1e6, cudaGetErrorString(e));
    return ORYX_E_NOMEM;
  }
  cudaStream_t st = m->update_stream;
  int32_t rc = ORYX_OK;
  do {
    if (cudaMemcpyAsync(d_pos, pos.data(), pos.size() * sizeof(int32_t), cudaMemcpyHostToDevice, st) != cudaSuccess ||
        cudaMemcpyAsync(d_seg, seg_begin.data(), seg_begin.size() * sizeof(int64_t), cudaMemcpyHostToDevice, st) != cudaSuccess) {
      set_error("H2D copy failed: %s", cudaGetErrorString(cudaGetLastError())); rc = ORYX_E_CUDA; break;
    }
    const unsigned blocks = (unsigned)((n_seg * 32 + 255) / 256);
    compact_save_kernel<<<blocks, 256, 0, st>>>(reinterpret_cast<const float4*>(s->d_vec), f4, d_pos, d_seg, n_seg, d_saved);
    compact_move_kernel<<<blocks, 256, 0, st>>>(reinterpret_cast<float4*>(s->d_vec), f4, d_pos, d_seg, n_seg, d_saved);
    m->launches += 2;
    if (cudaGetLastError() != cudaSuccess || cudaStreamSynchronize(st) != cudaSuccess) {
      set_error("in-place compaction failed: %s", cudaGetErrorString(cudaGetLastError())); rc = ORYX_E_CUDA; break;
    }
  } while (0);
  cudaFree(d_pos); cudaFree(d_seg); cudaFree(d_saved);
  return rc;
}

}  // namespace

int32_t store_compact(Model* m, int which) {
  Store* s = &m->store[which];
  s->version++;
  cudaStream_t st = m->update_stream;
  int64_t n = s->n_slots;
  if (n == 0) return ORYX_OK;
  std::vector<int32_t> part((size_t)n), rowid((size_t)n);
  ORYX_CUDA(cudaMemcpy(part.data(), s->d_slot_part, (size_t)n * sizeof(int32_t), cudaMemcpyDeviceToHost));
  ORYX_CUDA(cudaMemcpy(rowid.data(), s->d_slot_row, (size_t)n * sizeof(int32_t), cudaMemcpyDeviceToHost));
  // stable counting sort of live slots by partition
  int n_parts = 1 << m->num_hashes;
  if (!s->hashed) n_parts = 1;
  std::vector<int64_t> start((size_t)n_parts + 1, 0);
  int64_t live = 0;
  for (int64_t i = 0; i < n; i++) {
    if (rowid[i] >= 0 && part[i] >= 0 && part[i] < n_parts) { start[part[i] + 1]++; live++; }
  }
  for (int p = 0; p < n_parts; p++) start[p + 1] += start[p];
  std::vector<int32_t> src((size_t)std::max<int64_t>(live, 1));
  {
    std::vector<int64_t> fill(start.begin(), start.end() - 1);
    for (int64_t i = 0; i < n; i++) {
      if (rowid[i] >= 0 && part[i] >= 0 && part[i] < n_parts) src[fill[part[i]]++] = (int32_t)i;
    }
  }
  int64_t cap = std::max<int64_t>(s->cap_slots, 1024);
  float* nv = nullptr; int32_t *nr = nullptr, *np = nullptr, *d_src = nullptr;
  const bool force_in_place = std::getenv("ORYX_COMPACT_IN_PLACE") != nullptr;  // tests
  cudaError_t e = force_in_place ? cudaErrorMemoryAllocation : cudaMalloc(&nv, (size_t)cap * s->fpad * sizeof(float));
  if (e == cudaSuccess) e = cudaMalloc(&nr, (size_t)cap * sizeof(int32_t));
  if (e == cudaSuccess) e = cudaMalloc(&np, (size_t)cap * sizeof(int32_t));
  if (e == cudaSuccess) e = cudaMalloc(&d_src, (size_t)std::max<int64_t>(live, 1) * sizeof(int32_t));
  if (e != cudaSuccess) {
    // No room for a second copy of the matrix (a shard that fills most of the GPU): permute it in place.
    cudaFree(nv); cudaFree(nr); cudaFree(np); cudaFree(d_src);
    cudaGetLastError();
    std::vector<uint8_t> is_live((size_t)n);
    for (int64_t i = 0; i < n; i++) is_live[(size_t)i] = rowid[(size_t)i] >= 0 && part[(size_t)i] >= 0 && part[(size_t)i] < n_parts;
    ORYX_TRY(compact_rows_in_place(m, s, src, live, is_live));
    // the small per-slot arrays are rebuilt on the host and uploaded into the existing allocations
    std::vector<int32_t> new_row((size_t)n, -1), new_part((size_t)n, -1);
    for (int64_t i = 0; i < live; i++) {
      new_row[(size_t)i] = rowid[(size_t)src[(size_t)i]];
      new_part[(size_t)i] = part[(size_t)src[(size_t)i]];
    }
    ORYX_CUDA(cudaMemcpyAsync(s->d_slot_row, new_row.data(), (size_t)n * sizeof(int32_t), cudaMemcpyHostToDevice, st));
    ORYX_CUDA(cudaMemcpyAsync(s->d_slot_part, new_part.data(), (size_t)n * sizeof(int32_t), cudaMemcpyHostToDevice, st));
    if (s->d_row_slot) {
      fill_i32_kernel<<<(unsigned)((s->row_slot_cap + 255) / 256), 256, 0, st>>>(s->d_row_slot, s->row_slot_cap, -1);
      if (live > 0) rebuild_row_slot_kernel<<<(unsigned)((live + 255) / 256), 256, 0, st>>>(s->d_slot_row, live, s->d_row_slot);
      m->launches += 2;
    }
    ORYX_CUDA(cudaGetLastError());
    ORYX_CUDA(cudaStreamSynchronize(st));
    s->n_slots = live;
    s->free_slots.clear();
    for (int64_t i = 0; i < live; i++) s->row2slot[(size_t)rowid[(size_t)src[(size_t)i]]] = (int32_t)i;
    if (s->hashed) {
      ORYX_TRY(refresh_block_parts(m, s));
      ORYX_CUDA(cudaStreamSynchronize(st));
    }
    return ORYX_OK;
  }
  int threads = 256;
  if (live > 0) {
    ORYX_CUDA(cudaMemcpyAsync(d_src, src.data(), (size_t)live * sizeof(int32_t), cudaMemcpyHostToDevice, st));
    int64_t blocks = (live * 32 + threads - 1) / threads;
    gather_slots_kernel<<<(unsigned)blocks, threads, 0, st>>>(
        d_src, live, s->fpad, reinterpret_cast<const float4*>(s->d_vec), s->d_slot_row, s->d_slot_part,
        reinterpret_cast<float4*>(nv), nr, np);
    m->launches++;
  }
  int64_t tail = cap - live;
  if (tail > 0) {
    int64_t blocks = (tail + threads - 1) / threads;
    fill_i32_kernel<<<(unsigned)blocks, threads, 0, st>>>(nr + live, tail, -1);
    fill_i32_kernel<<<(unsigned)blocks, threads, 0, st>>>(np + live, tail, -1);
    m->launches += 2;
    ORYX_CUDA(cudaMemsetAsync(nv + live * s->fpad, 0, (size_t)tail * s->fpad * sizeof(float), st));
  }
  if (s->d_row_slot) {
    fill_i32_kernel<<<(unsigned)((s->row_slot_cap + 255) / 256), 256, 0, st>>>(s->d_row_slot, s->row_slot_cap, -1);
    if (live > 0) rebuild_row_slot_kernel<<<(unsigned)((live + 255) / 256), 256, 0, st>>>(nr, live, s->d_row_slot);
    m->launches += 2;
  }
  ORYX_CUDA(cudaGetLastError());
  ORYX_CUDA(cudaStreamSynchronize(st));
  cudaFree(s->d_vec); cudaFree(s->d_slot_row); cudaFree(s->d_slot_part); cudaFree(d_src);
  s->d_vec = nv; s->d_slot_row = nr; s->d_slot_part = np;
  layout_changed(m, s);
  s->cap_slots = cap;  // unchanged, so d_block_part keeps its size
  s->n_slots = live;
  s->free_slots.clear();
  for (int64_t i = 0; i < live; i++) s->row2slot[rowid[src[i]]] = (int32_t)i;
  if (s->hashed) {
    ORYX_TRY(refresh_block_parts(m, s));
    ORYX_CUDA(cudaStreamSynchronize(st));
  }
  return ORYX_OK;
}

}  // namespace oryx
