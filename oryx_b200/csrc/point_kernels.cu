// point_kernels.cu -- the small exact-arithmetic kernels around the scan: scores of every row (the
// arbitrary-Rescorer path, TopNConsumer.java:56-58), scores of a handful of rows (Estimate /
// SimilarityToItem / Because / MostSurprising point lookups, kernel K10) and the cross-shard merge of
// per-GPU top-N lists (ALSServingModel.java:334 when Y is row-sharded over several GPUs).
#include <algorithm>

#include "common.cuh"

namespace oryx {

namespace {

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

// Many (user, item) pairs at once: one thread per pair, the exact sequential arithmetic.
//   wide_products == 0: VectorMath.dot (Estimate.java:62-71): binary32 product, binary64 sum;
//   wide_products == 1: MLlib's MatrixFactorizationModel.predict, which the batch layer's Evaluation uses
//                       (features widened to double first, binary64 product and sum).
__global__ void pair_scores_kernel(const float* __restrict__ x, const float* __restrict__ y, int fpad, int features,
                                   const int32_t* __restrict__ x_slots, const int32_t* __restrict__ y_slots,
                                   int64_t n, int wide_products, double* __restrict__ out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int32_t xs = x_slots[i], ys = y_slots[i];
  if (xs < 0 || ys < 0) { out[i] = 0.0; return; }
  const float* xv = x + (int64_t)xs * fpad;
  const float* yv = y + (int64_t)ys * fpad;
  double acc = 0.0;
  if (wide_products) {
    for (int j = 0; j < features; j++) acc = __dadd_rn(acc, __dmul_rn((double)xv[j], (double)yv[j]));
  } else {
    for (int j = 0; j < features; j++) acc = exact_step(acc, xv[j], yv[j]);
  }
  out[i] = acc;
}

// Cross-shard merge (kernel K7, multi-GPU leg): rank-sort of n_lists*how_many exact (score,row) pairs
// by (score desc, global row asc) in one CTA.
__global__ void merge_topn_kernel(const unsigned char* __restrict__ scores_b, const unsigned char* __restrict__ rows_b,
                                  const unsigned char* __restrict__ counts_b, int64_t scores_stride,
                                  int64_t rows_stride, int64_t counts_stride, int n_lists, int how_many,
                                  int64_t row_stride, int64_t* __restrict__ out_rows,
                                  double* __restrict__ out_scores, int32_t* __restrict__ out_n,
                                  int64_t q_scores_stride = 0, int64_t q_rows_stride = 0, int64_t q_counts_stride = 0) {
  // one CTA per query (blockIdx.x): a page of queries answered by every shard is merged in one launch
  scores_b += (int64_t)blockIdx.x * q_scores_stride;
  rows_b += (int64_t)blockIdx.x * q_rows_stride;
  counts_b += (int64_t)blockIdx.x * q_counts_stride;
  out_rows += (int64_t)blockIdx.x * how_many;
  out_scores += (int64_t)blockIdx.x * how_many;
  out_n += blockIdx.x;
  // list l: scores at scores_b + l*scores_stride, rows at rows_b + l*rows_stride, count at counts_b + l*counts_stride
  auto score_at = [&](int l, int j) { return reinterpret_cast<const double*>(scores_b + l * scores_stride)[j]; };
  auto row_at = [&](int l, int j) { return reinterpret_cast<const int32_t*>(rows_b + l * rows_stride)[j]; };
  auto count_at = [&](int l) { return *reinterpret_cast<const int32_t*>(counts_b + l * counts_stride); };
  const int total = n_lists * how_many;
  int valid_total = 0;
  for (int l = 0; l < n_lists; l++) valid_total += min(max(count_at(l), 0), how_many);
  const int n_out = min(valid_total, how_many);
  for (int i = threadIdx.x; i < total; i += blockDim.x) {
    const int li = i / how_many, pi = i - li * how_many;
    if (pi >= count_at(li)) continue;
    const double si = score_at(li, pi);
    const int64_t ri = (int64_t)li * row_stride + row_at(li, pi);
    int rank = 0;
    for (int j = 0; j < total; j++) {
      const int lj = j / how_many, pj = j - lj * how_many;
      if (pj >= count_at(lj)) continue;
      const double sj = score_at(lj, pj);
      const int64_t rj = (int64_t)lj * row_stride + row_at(lj, pj);
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
  if (smem > 48 * 1024) ORYX_TRY(ensure_max_dynamic_smem(reinterpret_cast<const void*>(scores_all_kernel), smem));
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

int32_t pair_scores_launch(Model* m, const int32_t* d_x_slots, const int32_t* d_y_slots, int64_t n, int wide_products,
                           double* d_out, cudaStream_t stream) {
  const Store* sx = &m->store[ORYX_X];
  const Store* sy = &m->store[ORYX_Y];
  pair_scores_kernel<<<(unsigned)((n + 127) / 128), 128, 0, stream>>>(sx->d_vec, sy->d_vec, sy->fpad, sy->features,
                                                                     d_x_slots, d_y_slots, n, wide_products, d_out);
  m->launches++;
  ORYX_CUDA(cudaGetLastError());
  return ORYX_OK;
}

int32_t merge_topn_launch(const double* d_scores, const int32_t* d_rows, const int32_t* d_n, int n_lists,
                          int how_many, int64_t row_stride, int64_t* d_out_rows, double* d_out_scores,
                          int32_t* d_out_n, cudaStream_t stream) {
  merge_topn_kernel<<<1, 256, 0, stream>>>(reinterpret_cast<const unsigned char*>(d_scores),
                                           reinterpret_cast<const unsigned char*>(d_rows),
                                           reinterpret_cast<const unsigned char*>(d_n), (int64_t)how_many * 8,
                                           (int64_t)how_many * 4, 4, n_lists, how_many, row_stride, d_out_rows,
                                           d_out_scores, d_out_n);
  ORYX_CUDA(cudaGetLastError());
  return ORYX_OK;
}

// d_scores [n_lists][nq][how_many] f64, d_rows [n_lists][nq][how_many] i32, d_n [n_lists][nq] i32 (what every shard's
// oryx_topn_batch returned, gathered); outputs [nq][how_many] / [nq].
int32_t merge_topn_batch_launch(const double* d_scores, const int32_t* d_rows, const int32_t* d_n, int n_lists, int nq,
                                int how_many, int64_t row_stride, int64_t* d_out_rows, double* d_out_scores,
                                int32_t* d_out_n, cudaStream_t stream) {
  merge_topn_kernel<<<nq, 256, 0, stream>>>(reinterpret_cast<const unsigned char*>(d_scores),
                                            reinterpret_cast<const unsigned char*>(d_rows),
                                            reinterpret_cast<const unsigned char*>(d_n), (int64_t)nq * how_many * 8,
                                            (int64_t)nq * how_many * 4, (int64_t)nq * 4, n_lists, how_many, row_stride,
                                            d_out_rows, d_out_scores, d_out_n, (int64_t)how_many * 8, (int64_t)how_many * 4, 4);
  ORYX_CUDA(cudaGetLastError());
  return ORYX_OK;
}

int32_t merge_topn_packed_launch(const void* d_records, int n_lists, int record_bytes, int how_many,
                                 int64_t row_stride, int64_t* d_out_rows, double* d_out_scores, int32_t* d_out_n,
                                 cudaStream_t stream) {
  const unsigned char* base = reinterpret_cast<const unsigned char*>(d_records);
  merge_topn_kernel<<<1, 256, 0, stream>>>(base, base + (size_t)how_many * 8, base + (size_t)how_many * 12,
                                           record_bytes, record_bytes, record_bytes, n_lists, how_many,
                                           row_stride, d_out_rows, d_out_scores, d_out_n);
  ORYX_CUDA(cudaGetLastError());
  return ORYX_OK;
}

}  // namespace oryx
