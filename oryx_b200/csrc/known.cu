// known.cu -- device-resident known items.
//
// Replaces the knownItems map of ALSServingModel (app/oryx-app-serving/src/main/java/com/cloudera/oryx/
// app/serving/als/model/ALSServingModel.java:75-76) and its accessors: getKnownItems :189-201,
// getUserCounts / getItemCounts :212-239, addKnownItems :241-258, retainRecentAndKnownItems :418-441;
// and the allowed-predicate Recommend.java:92-99 builds from it for every /recommend request.
//
// One int32 pool in HBM holds every user's list of item ROWS in its own segment; the host keeps the
// authoritative sorted lists (set semantics) and the segment table.  At query time one small kernel
// turns the user's rows into the sorted SLOT list the scan's allowed-predicate expects (row -> slot
// through the store's device map, absent rows become a sentinel that matches nothing), so a
// /recommend call moves no exclusion list over PCIe.
#include <algorithm>

#include "common.cuh"

namespace oryx {

namespace {

constexpr int kPrepareMax = 4096;   // longest list the prepare kernel sorts in shared memory
constexpr int kPrepareThreads = 1024;

__global__ void known_scatter_kernel(int32_t* __restrict__ pool, const int64_t* __restrict__ dst,
                                     const int32_t* __restrict__ vals, int64_t n) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) pool[dst[i]] = vals[i];
}

// One CTA: rows -> slots (INT32_MAX for rows without a live slot), bitonic sort ascending in shared
// memory, first `n` entries out.
__global__ void __launch_bounds__(kPrepareThreads)
known_prepare_kernel(const int32_t* __restrict__ list, int n, int n_pow2, const int32_t* __restrict__ row_slot,
                     int64_t row_slot_cap, int32_t* __restrict__ out) {
  __shared__ int32_t s[kPrepareMax];
  for (int i = threadIdx.x; i < n_pow2; i += blockDim.x) {
    int32_t v = 0x7fffffff;
    if (i < n) {
      const int32_t row = list[i];
      if (row >= 0 && row < row_slot_cap) {
        const int32_t slot = row_slot[row];
        if (slot >= 0) v = slot;
      }
    }
    s[i] = v;
  }
  __syncthreads();
  for (int k = 2; k <= n_pow2; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int i = threadIdx.x; i < n_pow2; i += blockDim.x) {
        const int ixj = i ^ j;
        if (ixj > i) {
          const int32_t a = s[i], b = s[ixj];
          const bool up = (i & k) == 0;
          if ((a > b) == up) { s[i] = b; s[ixj] = a; }
        }
      }
      __syncthreads();
    }
  }
  for (int i = threadIdx.x; i < n; i += blockDim.x) out[i] = s[i];
}

// One warp per user: counts[item] += 1 for every known item (getItemCounts, ASM:227-239).
__global__ void known_item_counts_kernel(const int32_t* __restrict__ pool, const int64_t* __restrict__ seg_off,
                                         const int32_t* __restrict__ seg_len, int64_t n_users,
                                         int32_t* __restrict__ counts, int32_t n_items) {
  const int64_t u = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (u >= n_users) return;
  const int32_t* list = pool + seg_off[u];
  const int len = seg_len[u];
  for (int i = lane; i < len; i += 32) {
    const int32_t item = list[i];
    if (item >= 0 && item < n_items) atomicAdd(counts + item, 1);
  }
}

int32_t staging_reserve(Model* m, size_t n) {
  UpdateStaging& g = m->staging;
  if (n <= g.known_cap) return ORYX_OK;
  cudaFree(g.d_known_dst); cudaFree(g.d_known_val);
  g.d_known_dst = nullptr; g.d_known_val = nullptr; g.known_cap = 0;
  size_t cap = std::max<size_t>(n, 1024);
  ORYX_CUDA(cudaMalloc(&g.d_known_dst, cap * sizeof(int64_t)));
  ORYX_CUDA(cudaMalloc(&g.d_known_val, cap * sizeof(int32_t)));
  g.known_cap = cap;
  return ORYX_OK;
}

int32_t pool_reserve(Model* m, KnownItems* k, int64_t need) {
  if (need <= k->pool_cap) return ORYX_OK;
  int64_t cap = std::max<int64_t>(need, k->pool_cap * 2);
  cap = std::max<int64_t>(cap, 1 << 16);
  int32_t* np = nullptr;
  if (cudaMalloc(&np, (size_t)cap * sizeof(int32_t)) != cudaSuccess) {
    cudaGetLastError();
    set_error("device allocation of the known-items pool (%lld entries) failed", (long long)cap);
    return ORYX_E_NOMEM;
  }
  cudaStream_t st = m->update_stream;
  if (k->pool_cap > 0)
    ORYX_CUDA(cudaMemcpyAsync(np, k->d_pool, (size_t)k->pool_cap * sizeof(int32_t), cudaMemcpyDeviceToDevice, st));
  ORYX_CUDA(cudaStreamSynchronize(st));
  cudaFree(k->d_pool);
  k->d_pool = np;
  k->pool_cap = cap;
  return ORYX_OK;
}

int32_t segment_capacity(size_t size) {
  size_t cap = 8;
  while (cap < size + size / 2) cap <<= 1;
  return (int32_t)cap;
}

int32_t upload(Model* m, KnownItems* k, const std::vector<int64_t>& dst, const std::vector<int32_t>& val) {
  if (dst.empty()) return ORYX_OK;
  ORYX_TRY(staging_reserve(m, dst.size()));
  UpdateStaging& g = m->staging;
  cudaStream_t st = m->update_stream;
  ORYX_CUDA(cudaMemcpyAsync(g.d_known_dst, dst.data(), dst.size() * sizeof(int64_t), cudaMemcpyHostToDevice, st));
  ORYX_CUDA(cudaMemcpyAsync(g.d_known_val, val.data(), val.size() * sizeof(int32_t), cudaMemcpyHostToDevice, st));
  const int threads = 256;
  known_scatter_kernel<<<(unsigned)((dst.size() + threads - 1) / threads), threads, 0, st>>>(k->d_pool, g.d_known_dst, g.d_known_val,
                                                                                           (int64_t)dst.size());
  m->launches++;
  ORYX_CUDA(cudaGetLastError());
  ORYX_CUDA(cudaStreamSynchronize(st));
  return ORYX_OK;
}

// Re-lay every list compactly into a fresh pool (after retain, or when relocations left the pool
// more than half garbage).
int32_t rebuild(Model* m, KnownItems* k) {
  int64_t used = 0;
  for (size_t u = 0; u < k->lists.size(); u++) {
    if (k->lists[u].empty()) { k->seg_cap[u] = 0; k->seg_off[u] = 0; continue; }
    k->seg_cap[u] = segment_capacity(k->lists[u].size());
    k->seg_off[u] = used;
    used += k->seg_cap[u];
  }
  k->pool_used = used;
  k->pool_garbage = 0;
  ORYX_TRY(pool_reserve(m, k, used));
  std::vector<int64_t> dst;
  std::vector<int32_t> val;
  dst.reserve((size_t)k->total);
  val.reserve((size_t)k->total);
  for (size_t u = 0; u < k->lists.size(); u++) {
    for (size_t i = 0; i < k->lists[u].size(); i++) {
      dst.push_back(k->seg_off[u] + (int64_t)i);
      val.push_back(k->lists[u][i]);
    }
  }
  return upload(m, k, dst, val);
}

}  // namespace

void known_free(KnownItems* k) {
  cudaFree(k->d_pool);
  k->d_pool = nullptr;
  k->pool_cap = k->pool_used = k->pool_garbage = k->total = 0;
  k->lists.clear();
  k->seg_off.clear();
  k->seg_cap.clear();
}

int32_t known_add(Model* m, const int32_t* user_rows, const int64_t* begin, const int32_t* item_rows, int32_t n_users) {
  KnownItems* k = &m->known;
  int32_t max_user = -1;
  for (int32_t i = 0; i < n_users; i++) {
    if (user_rows[i] < 0) { set_error("negative user row %d", user_rows[i]); return ORYX_E_INVALID; }
    max_user = std::max(max_user, user_rows[i]);
    for (int64_t j = begin[i]; j < begin[i + 1]; j++) {
      if (item_rows[j] < 0) { set_error("negative item row %d", item_rows[j]); return ORYX_E_INVALID; }
    }
  }
  if ((size_t)max_user >= k->lists.size()) {
    const size_t n = (size_t)max_user + 1 + (size_t)max_user / 4;
    k->lists.resize(n);
    k->seg_off.resize(n, 0);
    k->seg_cap.resize(n, 0);
  }
  std::vector<int64_t> dst;
  std::vector<int32_t> val;
  std::vector<int32_t> fresh;
  for (int32_t i = 0; i < n_users; i++) {
    const int32_t u = user_rows[i];
    std::vector<int32_t>& list = k->lists[(size_t)u];
    fresh.assign(item_rows + begin[i], item_rows + begin[i + 1]);
    std::sort(fresh.begin(), fresh.end());
    fresh.erase(std::unique(fresh.begin(), fresh.end()), fresh.end());
    // keep only items the set does not hold yet (addKnownItems is Set.addAll, ASM:241-258)
    fresh.erase(std::remove_if(fresh.begin(), fresh.end(),
                               [&](int32_t v) { return std::binary_search(list.begin(), list.end(), v); }),
                fresh.end());
    if (fresh.empty()) continue;
    const size_t old_size = list.size(), new_size = old_size + fresh.size();
    if ((size_t)k->seg_cap[(size_t)u] >= new_size) {
      // the device segment is an unordered mirror of the set: append
      for (size_t j = 0; j < fresh.size(); j++) {
        dst.push_back(k->seg_off[(size_t)u] + (int64_t)(old_size + j));
        val.push_back(fresh[j]);
      }
      // the host list stays sorted, so the segment must be rewritten in the host's order only when
      // relocated; appended entries sit at the tail of the segment in arrival order
    } else {
      k->pool_garbage += k->seg_cap[(size_t)u];
      k->seg_cap[(size_t)u] = segment_capacity(new_size);
      k->seg_off[(size_t)u] = k->pool_used;
      k->pool_used += k->seg_cap[(size_t)u];
      // relocate: old entries (device order does not matter) then the new ones
      for (size_t j = 0; j < old_size; j++) {
        dst.push_back(k->seg_off[(size_t)u] + (int64_t)j);
        val.push_back(list[j]);
      }
      for (size_t j = 0; j < fresh.size(); j++) {
        dst.push_back(k->seg_off[(size_t)u] + (int64_t)(old_size + j));
        val.push_back(fresh[j]);
      }
    }
    const size_t mid = list.size();
    list.insert(list.end(), fresh.begin(), fresh.end());
    std::inplace_merge(list.begin(), list.begin() + (std::ptrdiff_t)mid, list.end());
    k->total += (int64_t)fresh.size();
  }
  if (k->pool_garbage > (1 << 20) && k->pool_garbage > k->pool_used / 2) return rebuild(m, k);
  ORYX_TRY(pool_reserve(m, k, k->pool_used));
  return upload(m, k, dst, val);
}

int32_t known_retain(Model* m, const uint8_t* keep_users, int32_t n_user_rows, const uint8_t* keep_items,
                     int32_t n_item_rows) {
  KnownItems* k = &m->known;
  int64_t total = 0;
  for (size_t u = 0; u < k->lists.size(); u++) {
    std::vector<int32_t>& list = k->lists[u];
    if (list.empty()) continue;
    if (keep_users && !((int64_t)u < n_user_rows && keep_users[u])) {  // ASM:423-425
      std::vector<int32_t>().swap(list);
      continue;
    }
    if (keep_items) {  // ASM:434-440
      list.erase(std::remove_if(list.begin(), list.end(),
                                [&](int32_t v) { return !(v < n_item_rows && keep_items[v]); }),
                 list.end());
    }
    total += (int64_t)list.size();
  }
  k->total = total;
  return rebuild(m, k);
}

int32_t known_counts(Model* m, int which, int32_t* out_counts, int32_t n_rows) {
  KnownItems* k = &m->known;
  std::fill(out_counts, out_counts + n_rows, 0);
  if (which == ORYX_X) {  // getUserCounts ASM:212-225: the size of each user's set
    for (size_t u = 0; u < k->lists.size() && (int64_t)u < n_rows; u++) out_counts[u] = (int32_t)k->lists[u].size();
    return ORYX_OK;
  }
  if (k->total == 0) return ORYX_OK;
  const int64_t n_users = (int64_t)k->lists.size();
  std::vector<int32_t> len((size_t)n_users);
  for (int64_t u = 0; u < n_users; u++) len[(size_t)u] = (int32_t)k->lists[(size_t)u].size();
  int64_t* d_off = nullptr;
  int32_t *d_len = nullptr, *d_counts = nullptr;
  cudaStream_t st = m->update_stream;
  cudaError_t e = cudaMalloc(&d_off, (size_t)n_users * sizeof(int64_t));
  if (e == cudaSuccess) e = cudaMalloc(&d_len, (size_t)n_users * sizeof(int32_t));
  if (e == cudaSuccess) e = cudaMalloc(&d_counts, (size_t)n_rows * sizeof(int32_t));
  int32_t rc = ORYX_OK;
  do {
    if (e != cudaSuccess) { cudaGetLastError(); set_error("device allocation for known-item counts failed"); rc = ORYX_E_NOMEM; break; }
    if (cudaMemcpyAsync(d_off, k->seg_off.data(), (size_t)n_users * sizeof(int64_t), cudaMemcpyHostToDevice, st) != cudaSuccess ||
        cudaMemcpyAsync(d_len, len.data(), (size_t)n_users * sizeof(int32_t), cudaMemcpyHostToDevice, st) != cudaSuccess ||
        cudaMemsetAsync(d_counts, 0, (size_t)n_rows * sizeof(int32_t), st) != cudaSuccess) {
      set_error("CUDA copy failed: %s", cudaGetErrorString(cudaGetLastError())); rc = ORYX_E_CUDA; break;
    }
    const int threads = 256;
    known_item_counts_kernel<<<(unsigned)((n_users * 32 + threads - 1) / threads), threads, 0, st>>>(
        k->d_pool, d_off, d_len, n_users, d_counts, n_rows);
    m->launches++;
    if (cudaMemcpyAsync(out_counts, d_counts, (size_t)n_rows * sizeof(int32_t), cudaMemcpyDeviceToHost, st) != cudaSuccess ||
        cudaStreamSynchronize(st) != cudaSuccess) {
      set_error("known-item count kernel failed: %s", cudaGetErrorString(cudaGetLastError())); rc = ORYX_E_CUDA; break;
    }
  } while (0);
  cudaFree(d_off); cudaFree(d_len); cudaFree(d_counts);
  return rc;
}

int32_t known_exclusions_device(Model* m, QueryCtx* ctx, int32_t user_row, int32_t* n_excl, cudaStream_t stream) {
  KnownItems* k = &m->known;
  *n_excl = 0;
  if (user_row < 0 || (size_t)user_row >= k->lists.size()) return ORYX_OK;
  const std::vector<int32_t>& list = k->lists[(size_t)user_row];
  const int n = (int)list.size();
  if (n == 0) return ORYX_OK;
  ORYX_TRY(ensure_excl(ctx, n));
  Store* s = &m->store[ORYX_Y];
  if (n <= kPrepareMax) {
    int n_pow2 = 1;
    while (n_pow2 < n) n_pow2 <<= 1;
    known_prepare_kernel<<<1, kPrepareThreads, 0, stream>>>(k->d_pool + k->seg_off[(size_t)user_row], n, n_pow2,
                                                           s->d_row_slot, s->row_slot_cap, ctx->d_excl);
    m->launches++;
    ORYX_CUDA(cudaGetLastError());
    *n_excl = n;
    return ORYX_OK;
  }
  // very long lists: translate the authoritative host list (same as the caller-supplied path)
  int32_t cnt = 0;
  for (int32_t r : list) {
    if ((size_t)r < s->row2slot.size() && s->row2slot[(size_t)r] >= 0) ctx->h_excl[cnt++] = s->row2slot[(size_t)r];
  }
  std::sort(ctx->h_excl, ctx->h_excl + cnt);
  if (cnt > 0)
    ORYX_CUDA(cudaMemcpyAsync(ctx->d_excl, ctx->h_excl, (size_t)cnt * sizeof(int32_t), cudaMemcpyHostToDevice, stream));
  *n_excl = cnt;
  return ORYX_OK;
}

}  // namespace oryx
