// aggregate.cu -- the speed layer's aggregation step on the device.
//
// Replaces, for interactions whose ids have been mapped to integer rows, the part of
// ALSSpeedModelManager.buildUpdates between parsing and fold-in (app/oryx-app/.../speed/als/
// ALSSpeedModelManager.java:141-171): order by timestamp, group by (user, item), then
//   implicit:  MLFunctions.SUM_WITH_NAN over the group in timestamp order (app/oryx-app-common/.../common/fn/
//              MLFunctions.java:67-78): tally = NaN; for each s: tally = isNaN(tally) ? s : tally + s -- a NaN
//              (a delete) poisons the tally and the next value starts it afresh;
//   explicit:  foldByKey(NaN, (current, next) -> next): the last value wins (:164);
// drop groups whose result is NaN (:167-168) and narrow the strength to float (:171; widened again at :184).
//
// One bitonic sort of (user << 32 | item, timestamp, input position) -- all three, so the order is total and the
// sort need not be stable -- then one thread per group head walks its group in order (binary64 adds in the
// reference's sequence), an exclusive scan compacts the surviving groups, and they come out ordered by (user,
// item).  The reference's own output order is whatever Spark's hash partitioning produces; callers treat it
// as a set.
#include <algorithm>

#include "common.cuh"

namespace oryx {

namespace {

struct AggRec {
  unsigned long long key;  // user << 32 | item; ~0 = padding
  long long ts;
  uint32_t idx;            // position in the input (ties in timestamp keep input order)
  uint32_t pad;
};

__device__ __forceinline__ bool agg_less(const AggRec& a, const AggRec& b) {
  if (a.key != b.key) return a.key < b.key;
  if (a.ts != b.ts) return a.ts < b.ts;
  return a.idx < b.idx;
}

__global__ void agg_build_kernel(const int32_t* __restrict__ users, const int32_t* __restrict__ items,
                                 const long long* __restrict__ ts, int64_t n, int64_t padded, AggRec* __restrict__ recs) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= padded) return;
  AggRec r;
  if (i < n) {
    r.key = ((unsigned long long)(uint32_t)users[i] << 32) | (unsigned long long)(uint32_t)items[i];
    r.ts = ts[i];
    r.idx = (uint32_t)i;
  } else {
    r.key = ~0ull;
    r.ts = 0x7fffffffffffffffLL;
    r.idx = 0xffffffffu;
  }
  r.pad = 0;
  recs[i] = r;
}

// One compare-exchange step of an ascending bitonic sort.
__global__ void agg_bitonic_kernel(AggRec* __restrict__ recs, int64_t n, int64_t k2, int64_t j) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int64_t ixj = i ^ j;
  if (ixj <= i) return;
  const AggRec a = recs[i], b = recs[ixj];
  const bool asc = (i & k2) == 0;
  if (agg_less(b, a) == asc) { recs[i] = b; recs[ixj] = a; }
}

// All steps with j < 1024 of one merge stage, inside shared memory (1024 records per CTA).
__global__ void __launch_bounds__(512) agg_bitonic_smem_kernel(AggRec* __restrict__ recs, int64_t k2, int j_start) {
  __shared__ AggRec s[1024];
  const int64_t base = (int64_t)blockIdx.x * 1024;
  for (int e = threadIdx.x; e < 1024; e += 512) s[e] = recs[base + e];
  __syncthreads();
  for (int j = j_start; j > 0; j >>= 1) {
    for (int e = threadIdx.x; e < 1024; e += 512) {
      const int ixj = e ^ j;
      if (ixj > e) {
        const bool asc = ((base + e) & k2) == 0;
        const AggRec a = s[e], b = s[ixj];
        if (agg_less(b, a) == asc) { s[e] = b; s[ixj] = a; }
      }
    }
    __syncthreads();
  }
  for (int e = threadIdx.x; e < 1024; e += 512) recs[base + e] = s[e];
}

// Group heads reduce their group; flag = the group survives (result is not NaN).
__global__ void agg_reduce_kernel(const AggRec* __restrict__ recs, const double* __restrict__ strengths, int64_t n,
                                  int implicit, double* __restrict__ value, uint32_t* __restrict__ flag) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const unsigned long long key = recs[i].key;
  uint32_t f = 0;
  if (i == 0 || recs[i - 1].key != key) {
    double tally = __longlong_as_double(0x7ff8000000000000LL);
    for (int64_t j = i; j < n && recs[j].key == key; j++) {
      const double s = strengths[recs[j].idx];
      if (implicit) tally = (tally != tally) ? s : __dadd_rn(tally, s);  // SUM_WITH_NAN
      else tally = s;                                                     // last wins
    }
    if (tally == tally) {
      f = 1;
      value[i] = (double)(float)tally;  // strength.floatValue(), widened again by getStrength()
    }
  }
  flag[i] = f;
}

// Exclusive scan of flags, three steps: per-CTA totals, scan of the totals (one CTA), positions.
__global__ void agg_scan_totals_kernel(const uint32_t* __restrict__ flag, int64_t n, uint32_t* __restrict__ totals) {
  __shared__ uint32_t s_sum;
  if (threadIdx.x == 0) s_sum = 0;
  __syncthreads();
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint32_t v = i < n ? flag[i] : 0u;
  const uint32_t w = __reduce_add_sync(0xffffffffu, v);
  if ((threadIdx.x & 31) == 0 && w) atomicAdd(&s_sum, w);
  __syncthreads();
  if (threadIdx.x == 0) totals[blockIdx.x] = s_sum;
}
__global__ void agg_scan_blocks_kernel(uint32_t* __restrict__ totals, int n_blocks, uint32_t* __restrict__ grand_total) {
  uint32_t run = 0;  // one thread: a few thousand entries at most
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    for (int b = 0; b < n_blocks; b++) { const uint32_t t = totals[b]; totals[b] = run; run += t; }
    *grand_total = run;
  }
}
__global__ void agg_scatter_kernel(const AggRec* __restrict__ recs, const double* __restrict__ value,
                                   const uint32_t* __restrict__ flag, const uint32_t* __restrict__ totals, int64_t n,
                                   int32_t* __restrict__ out_users, int32_t* __restrict__ out_items,
                                   double* __restrict__ out_strengths) {
  __shared__ uint32_t s_warp[32];
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t v = i < n ? flag[i] : 0u;
  const uint32_t ball = __ballot_sync(0xffffffffu, v != 0);
  if (lane == 0) s_warp[warp] = __popc(ball);
  __syncthreads();
  uint32_t before = totals[blockIdx.x];
  for (int w = 0; w < warp; w++) before += s_warp[w];
  before += __popc(ball & ((1u << lane) - 1u));
  if (v) {
    out_users[before] = (int32_t)(recs[i].key >> 32);
    out_items[before] = (int32_t)(recs[i].key & 0xffffffffu);
    out_strengths[before] = value[i];
  }
}

}  // namespace
}  // namespace oryx

using namespace oryx;

extern "C" int32_t oryx_aggregate_interactions(oryx_model* mh, const int32_t* user_rows, const int32_t* item_rows,
                                               const double* strengths, const int64_t* timestamps, int64_t n,
                                               int32_t implicit, int32_t* out_user_rows, int32_t* out_item_rows,
                                               double* out_strengths, int64_t* out_n) {
  ORYX_ARG(mh && out_n, "null argument");
  ORYX_ARG(n >= 0 && n <= ((int64_t)1 << 26), "n out of range [0, 2^26]");
  *out_n = 0;
  if (n == 0) return ORYX_OK;
  ORYX_ARG(user_rows && item_rows && strengths && timestamps && out_user_rows && out_item_rows && out_strengths,
           "null argument");
  for (int64_t i = 0; i < n; i++) ORYX_ARG(user_rows[i] >= 0 && item_rows[i] >= 0, "row ids must be >= 0");
  Model* m = reinterpret_cast<Model*>(mh);
  int prev = -1;
  if (cudaGetDevice(&prev) != cudaSuccess) { cudaGetLastError(); prev = -1; }
  if (cudaSetDevice(m->device) != cudaSuccess) {
    cudaGetLastError();
    set_error("cannot select CUDA device %d (no CPU fallback exists)", m->device);
    return ORYX_E_CUDA;
  }
  struct Restore { int p; ~Restore() { if (p >= 0) cudaSetDevice(p); } } restore{prev};
  int64_t padded = 1024;
  while (padded < n) padded <<= 1;
  const int64_t n_blocks = (n + 255) / 256;
  // scratch: recs | users | items | ts | strengths | value | flag | totals (+ grand total) | outputs
  size_t off = 0;
  auto take = [&](size_t bytes) { const size_t o = off; off += (bytes + 255) & ~(size_t)255; return o; };
  const size_t o_recs = take((size_t)padded * sizeof(AggRec)), o_users = take((size_t)n * 4), o_items = take((size_t)n * 4);
  const size_t o_ts = take((size_t)n * 8), o_str = take((size_t)n * 8), o_val = take((size_t)n * 8), o_flag = take((size_t)n * 4);
  const size_t o_tot = take((size_t)(n_blocks + 1) * 4), o_ou = take((size_t)n * 4), o_oi = take((size_t)n * 4), o_os = take((size_t)n * 8);
  std::lock_guard<std::mutex> lk(m->gram_mu);  // shares the model's Gram scratch buffer: one user at a time
  if (m->gram_scratch_bytes < off) {
    cudaFree(m->d_gram_scratch);
    m->d_gram_scratch = nullptr;
    m->gram_scratch_bytes = 0;
    if (cudaMalloc(&m->d_gram_scratch, off) != cudaSuccess) {
      cudaGetLastError();
      set_error("device allocation of %zu bytes failed (interaction aggregation)", off);
      return ORYX_E_NOMEM;
    }
    m->gram_scratch_bytes = off;
  }
  unsigned char* base = reinterpret_cast<unsigned char*>(m->d_gram_scratch);
  AggRec* recs = reinterpret_cast<AggRec*>(base + o_recs);
  int32_t* d_users = reinterpret_cast<int32_t*>(base + o_users);
  int32_t* d_items = reinterpret_cast<int32_t*>(base + o_items);
  long long* d_ts = reinterpret_cast<long long*>(base + o_ts);
  double* d_str = reinterpret_cast<double*>(base + o_str);
  double* d_val = reinterpret_cast<double*>(base + o_val);
  uint32_t* d_flag = reinterpret_cast<uint32_t*>(base + o_flag);
  uint32_t* d_tot = reinterpret_cast<uint32_t*>(base + o_tot);
  cudaStream_t st = m->update_stream;
  ORYX_CUDA(cudaMemcpyAsync(d_users, user_rows, (size_t)n * 4, cudaMemcpyHostToDevice, st));
  ORYX_CUDA(cudaMemcpyAsync(d_items, item_rows, (size_t)n * 4, cudaMemcpyHostToDevice, st));
  ORYX_CUDA(cudaMemcpyAsync(d_ts, timestamps, (size_t)n * 8, cudaMemcpyHostToDevice, st));
  ORYX_CUDA(cudaMemcpyAsync(d_str, strengths, (size_t)n * 8, cudaMemcpyHostToDevice, st));
  agg_build_kernel<<<(unsigned)((padded + 255) / 256), 256, 0, st>>>(d_users, d_items, d_ts, n, padded, recs);
  int launches = 1;
  for (int64_t k2 = 2; k2 <= padded; k2 <<= 1) {
    int64_t j = k2 >> 1;
    for (; j >= 1024; j >>= 1) {
      agg_bitonic_kernel<<<(unsigned)((padded + 255) / 256), 256, 0, st>>>(recs, padded, k2, j);
      launches++;
    }
    agg_bitonic_smem_kernel<<<(unsigned)(padded / 1024), 512, 0, st>>>(recs, k2, (int)j);
    launches++;
  }
  agg_reduce_kernel<<<(unsigned)n_blocks, 256, 0, st>>>(recs, d_str, n, implicit ? 1 : 0, d_val, d_flag);
  agg_scan_totals_kernel<<<(unsigned)n_blocks, 256, 0, st>>>(d_flag, n, d_tot);
  agg_scan_blocks_kernel<<<1, 32, 0, st>>>(d_tot, (int)n_blocks, d_tot + n_blocks);
  agg_scatter_kernel<<<(unsigned)n_blocks, 256, 0, st>>>(recs, d_val, d_flag, d_tot, n, reinterpret_cast<int32_t*>(base + o_ou),
                                                        reinterpret_cast<int32_t*>(base + o_oi), reinterpret_cast<double*>(base + o_os));
  ORYX_CUDA(cudaGetLastError());
  m->launches += launches + 4;
  uint32_t kept = 0;
  ORYX_CUDA(cudaMemcpyAsync(&kept, d_tot + n_blocks, 4, cudaMemcpyDeviceToHost, st));
  ORYX_CUDA(cudaStreamSynchronize(st));
  if (kept > 0) {
    ORYX_CUDA(cudaMemcpyAsync(out_user_rows, base + o_ou, (size_t)kept * 4, cudaMemcpyDeviceToHost, st));
    ORYX_CUDA(cudaMemcpyAsync(out_item_rows, base + o_oi, (size_t)kept * 4, cudaMemcpyDeviceToHost, st));
    ORYX_CUDA(cudaMemcpyAsync(out_strengths, base + o_os, (size_t)kept * 8, cudaMemcpyDeviceToHost, st));
    ORYX_CUDA(cudaStreamSynchronize(st));
  }
  *out_n = kept;
  return ORYX_OK;
}
