// gram.cu -- G = V^T V over a device store (kernel K4 of SURVEY.md section 2.3).
//
// Replaces VectorMath.transposeTimesSelf (reference framework/oryx-common/.../math/VectorMath.java:
// 64-83) as called through FeatureVectors.getVTV (app/oryx-app-common/.../als/FeatureVectors.java:
// 150-154) and summed over LSH partitions (ALSServingModel.java:362-368; ALSSpeedModel.java:118-126).
//
// Reference arithmetic: G[r][c] += (double)(v[r]*v[c]) -- a binary32 product widened and added into
// a binary64 entry, for every row v.  gram_exact_kernel performs exactly those operations
// (__fmul_rn, widen, __dadd_rn); only the order in which rows are summed differs (the reference's
// own order is hash-map iteration order, i.e. unspecified), so entries agree to ~1e-16 relative.
// Tombstoned slots (NaN-filled by the store) are skipped via slot_row.
#include <algorithm>

#include "common.cuh"

namespace oryx {

namespace {

constexpr int kTile = 64;   // output block is kTile x kTile
constexpr int kKC = 32;     // rows staged per step

// grid.x = upper-triangular block index, grid.y = row chunk.
__global__ void __launch_bounds__(256) gram_exact_kernel(const float* __restrict__ vec,
                                                        const int32_t* __restrict__ slot_row, int64_t n_slots,
                                                        int fpad, int features, int nb,
                                                        int64_t rows_per_chunk, double* __restrict__ out) {
  __shared__ __align__(16) float As[kKC][kTile];
  __shared__ __align__(16) float Bs[kKC][kTile];
  // decode block (bi <= bj) from the linear upper-triangular index
  int t = blockIdx.x, bi = 0;
  while (t >= nb - bi) { t -= nb - bi; bi++; }
  const int bj = bi + t;
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const int64_t row_begin = (int64_t)blockIdx.y * rows_per_chunk;
  const int64_t row_end = min(n_slots, row_begin + rows_per_chunk);

  double acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; i++)
#pragma unroll
    for (int j = 0; j < 4; j++) acc[i][j] = 0.0;

  for (int64_t r0 = row_begin; r0 < row_end; r0 += kKC) {
    __syncthreads();
    // stage kKC rows x 64 columns of both column blocks (coalesced along columns)
    for (int e = threadIdx.x; e < kKC * kTile; e += 256) {
      int k = e / kTile, c = e - k * kTile;
      int64_t row = r0 + k;
      int ca = bi * kTile + c, cb = bj * kTile + c;
      bool ok = row < row_end && slot_row[row] >= 0;  // tombstones are NaN-filled: skip them
      As[k][c] = (ok && ca < features) ? vec[row * fpad + ca] : 0.0f;
      Bs[k][c] = (ok && cb < features) ? vec[row * fpad + cb] : 0.0f;
    }
    __syncthreads();
#pragma unroll 4
    for (int k = 0; k < kKC; k++) {
      float4 a = *reinterpret_cast<const float4*>(&As[k][ty * 4]);
      float4 b = *reinterpret_cast<const float4*>(&Bs[k][tx * 4]);
      const float av[4] = {a.x, a.y, a.z, a.w};
      const float bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
      for (int i = 0; i < 4; i++)
#pragma unroll
        for (int j = 0; j < 4; j++) acc[i][j] = exact_step(acc[i][j], av[i], bv[j]);
    }
  }
#pragma unroll
  for (int i = 0; i < 4; i++) {
#pragma unroll
    for (int j = 0; j < 4; j++) {
      int r = bi * kTile + ty * 4 + i, c = bj * kTile + tx * 4 + j;
      if (r < features && c < features) {
        atomicAdd(out + (size_t)r * features + c, acc[i][j]);
        if (bi != bj) atomicAdd(out + (size_t)c * features + r, acc[i][j]);
      }
    }
  }
}

// ---- large stores: fp32 FMA tiles with binary64 flushes ------------------------------------------
// The exact kernel above is bound by the float->double conversion pipe (one convert per term).  For
// large stores the per-term fp32 rounding the reference applies averages out: over n rows it perturbs
// an entry by ~2^-24/sqrt(n) relative, far below anything the solver's float output can see.  This
// kernel therefore accumulates fma(a, b, acc) in fp32 for kFlushRows rows at a time (rounding error
// ~sqrt(kFlushRows) 2^-24 of one chunk's partial sum, unbiased), adds each chunk into binary64
// accumulators in shared memory, and reduces CTAs with binary64 atomics.  Agreement with the exact
// kernel improves with n: ~1e-8 relative at 7e4 rows, ~1e-9 at 1e7.  128 x 128 output block per CTA, 8 x 8 per thread.
constexpr int kFastTile = 128;
constexpr int kFastKC = 16;
constexpr int kFlushRows = 64;

// T threads: 256 (8 x 8 outputs per thread) or 512 (4 x 8; twice the warps to hide shared-memory latency).
template <int T>
__global__ void __launch_bounds__(T, 1) gram_fast_kernel(const float* __restrict__ vec,
                                                           const int32_t* __restrict__ slot_row, int64_t n_slots,
                                                           int fpad, int features, int nb, int64_t rows_per_chunk,
                                                           double* __restrict__ out) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double* acc64 = reinterpret_cast<double*>(smem_raw);                       // [128][128]
  float* As = reinterpret_cast<float*>(smem_raw + (size_t)kFastTile * kFastTile * 8);  // [2][KC][128]
  float* Bs = As + 2 * kFastKC * kFastTile;                                  // [2][KC][128]
  int t = blockIdx.x, bi = 0;
  while (t >= nb - bi) { t -= nb - bi; bi++; }
  const int bj = bi + t;
  constexpr int RH = T == 256 ? 2 : 1;  // 64-row halves of the tile a thread covers
  constexpr int RI = 4 * RH;            // output rows per thread
  constexpr int kFastStage = kFastKC * 32 / T;  // float4 of A (and of B) each thread stages per step
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int64_t row_begin = (int64_t)blockIdx.y * rows_per_chunk;
  const int64_t row_end = min(n_slots, row_begin + rows_per_chunk);
  for (int e = tid; e < kFastTile * kFastTile; e += T) acc64[e] = 0.0;

  float acc[RI][8];
#pragma unroll
  for (int i = 0; i < RI; i++)
#pragma unroll
    for (int j = 0; j < 8; j++) acc[i][j] = 0.f;

  // each thread stages kFastStage float4 of A and of B per step: row k = e / 32, float4 column = e % 32.
  // fetch() only issues the global loads (into registers; the tombstone test and the row load are
  // independent, one latency instead of two), commit() stores them to shared memory after the math of
  // the current step -- so the loads are in flight for a whole step instead of stalling it.
  float4 fa[kFastStage], fb[kFastStage];
  int32_t fr[kFastStage];
  bool fin[kFastStage];
  auto fetch = [&](int64_t r0) {
#pragma unroll
    for (int h = 0; h < kFastStage; h++) {
      const int e = tid + h * T;
      const int k = e >> 5, c4 = e & 31;
      const int64_t row = r0 + k;
      fin[h] = row < row_end;
      const int ca = bi * kFastTile + c4 * 4, cb = bj * kFastTile + c4 * 4;
      fa[h] = make_float4(0.f, 0.f, 0.f, 0.f);
      fb[h] = fa[h];
      fr[h] = -1;
      if (fin[h]) {
        fr[h] = __ldg(slot_row + row);
        if (ca < fpad) fa[h] = __ldg(reinterpret_cast<const float4*>(vec + row * fpad + ca));
        if (cb < fpad) fb[h] = __ldg(reinterpret_cast<const float4*>(vec + row * fpad + cb));
      }
    }
  };
  auto commit = [&](int buf) {
#pragma unroll
    for (int h = 0; h < kFastStage; h++) {
      const int e = tid + h * T;
      const int k = e >> 5, c4 = e & 31;
      const bool ok = fin[h] && fr[h] >= 0;  // tombstones are NaN-filled: contribute zeros
      const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
      *reinterpret_cast<float4*>(As + ((size_t)buf * kFastKC + k) * kFastTile + c4 * 4) = ok ? fa[h] : z;
      *reinterpret_cast<float4*>(Bs + ((size_t)buf * kFastKC + k) * kFastTile + c4 * 4) = ok ? fb[h] : z;
    }
  };
  auto flush = [&]() {
#pragma unroll
    for (int i = 0; i < RI; i++)
#pragma unroll
      for (int j = 0; j < 8; j++) {
        // accumulator (i, j) of thread tid lives at [(i*8+j)][tid]: a warp touches 32 consecutive
        // doubles, conflict-free (a [row][col] layout made this read-modify-write a 4-way bank conflict
        // and the largest user of shared-memory bandwidth in the kernel)
        double* d = acc64 + (size_t)(i * 8 + j) * T + tid;
        *d += (double)acc[i][j];
        acc[i][j] = 0.f;
      }
  };

  __syncthreads();
  int buf = 0, since_flush = 0;
  if (row_begin < row_end) { fetch(row_begin); commit(0); }
  __syncthreads();
  for (int64_t r0 = row_begin; r0 < row_end; r0 += kFastKC) {
    const bool more = r0 + kFastKC < row_end;
    if (more) fetch(r0 + kFastKC);  // the next step's loads fly during this step's math
    const float* a_s = As + (size_t)buf * kFastKC * kFastTile;
    const float* b_s = Bs + (size_t)buf * kFastKC * kFastTile;
#pragma unroll
    for (int k = 0; k < kFastKC; k++) {
      // lanes read consecutive float4 (16-byte stride): conflict-free; second half 64 columns on
      const float4 a0 = *reinterpret_cast<const float4*>(a_s + k * kFastTile + ty * 4);
      const float4 a1 = RH == 2 ? *reinterpret_cast<const float4*>(a_s + k * kFastTile + 64 + ty * 4) : a0;
      const float4 b0 = *reinterpret_cast<const float4*>(b_s + k * kFastTile + tx * 4);
      const float4 b1 = *reinterpret_cast<const float4*>(b_s + k * kFastTile + 64 + tx * 4);
      const float av[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
      const float bv[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
      for (int i = 0; i < RI; i++)
#pragma unroll
        for (int j = 0; j < 8; j++) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    if (more) commit(buf ^ 1);
    since_flush += kFastKC;
    if (since_flush >= kFlushRows) { flush(); since_flush = 0; }
    __syncthreads();
    buf ^= 1;
  }
  flush();
  __syncthreads();
  for (int e = tid; e < kFastTile * kFastTile; e += T) {
    // e = (i*8+j)*T + owner thread  ->  the tile coordinates that thread's accumulator (i, j) stands for
    const int ij = e / T, owner = e % T;
    const int ai = ij >> 3, aj = ij & 7, otx = owner & 15, oty = owner >> 4;
    const int i = (ai < 4 ? 0 : 64) + oty * 4 + (ai & 3), j = (aj < 4 ? 0 : 64) + otx * 4 + (aj & 3);
    const int r = bi * kFastTile + i, c = bj * kFastTile + j;
    if (r < features && c < features) {
      const double v = acc64[e];
      atomicAdd(out + (size_t)r * features + c, v);
      if (bi != bj) atomicAdd(out + (size_t)c * features + r, v);
    }
  }
}

}  // namespace

// Both kernels above add every row chunk's partial into G with binary64 atomics, entry (r, c) and its mirror (c, r)
// separately: the two receive the same partials but not necessarily in the same order, so they could differ in the
// last bit (seen once in ~20 runs of the test suite).  G is symmetric by definition: the lower triangle is copied
// from the upper one once the sums are complete.
__global__ void gram_symmetrise_kernel(double* __restrict__ g, int features) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= features * features) return;
  const int r = e / features, c = e - r * features;
  if (r > c) g[e] = g[(size_t)c * features + r];
}

int32_t gram_symmetrise(double* d_out, int F, cudaStream_t stream) {
  gram_symmetrise_kernel<<<(unsigned)((F * F + 255) / 256), 256, 0, stream>>>(d_out, F);
  ORYX_CUDA(cudaGetLastError());
  return ORYX_OK;
}

constexpr int64_t kGramFastMinRows = 1 << 16;

int32_t gram_fast_launch(Model* m, Store* s, double* d_out, cudaStream_t stream) {
  const int F = s->features;
  const int nb = (F + kFastTile - 1) / kFastTile;
  const int n_blocks = nb * (nb + 1) / 2;
  // one CTA per SM (160 KB of shared memory): size the grid to one full wave, never a wave plus a few
  int64_t chunks = std::max<int64_t>(1, (int64_t)m->num_sms / n_blocks);
  int64_t rows_per_chunk = (s->n_slots + chunks - 1) / chunks;
  rows_per_chunk = (rows_per_chunk + kFlushRows - 1) / kFlushRows * kFlushRows;
  chunks = (s->n_slots + rows_per_chunk - 1) / rows_per_chunk;
  const size_t smem = (size_t)kFastTile * kFastTile * 8 + (size_t)4 * kFastKC * kFastTile * 4;
  constexpr int kThreads = 256;  // 512 threads (4 x 8 outputs each) measured slower: 30.8 ms vs 25.0 ms at 10 M x 256
  ORYX_TRY(ensure_max_dynamic_smem(reinterpret_cast<const void*>(gram_fast_kernel<kThreads>), smem));
  dim3 grid((unsigned)n_blocks, (unsigned)chunks);
  gram_fast_kernel<kThreads><<<grid, kThreads, smem, stream>>>(s->d_vec, s->d_slot_row, s->n_slots, s->fpad, F, nb, rows_per_chunk, d_out);
  m->launches++;
  ORYX_CUDA(cudaGetLastError());
  return ORYX_OK;
}

// gram_tc.cu
bool gram_tc_applicable(const Store* s);
int32_t gram_tc_absmax(const Store* s, cudaStream_t stream, float* out);
int32_t gram_tc_launch(Model* m, Store* s, double* d_out, cudaStream_t stream);

// `allow_tc`: let large stores take the exact-sum tensor-core kernel (gram_tc.cu).  `*tc_absmax` (optional)
// receives the store's largest |x| when that kernel ran and 0 otherwise, so the caller can check the
// error estimate that goes with its fixed-point scale.
int32_t gram_launch(Model* m, int which, double* d_out, cudaStream_t stream, bool allow_tc, float* tc_absmax) {
  Store* s = &m->store[which];
  const int F = s->features;
  if (tc_absmax) *tc_absmax = 0.f;
  ORYX_CUDA(cudaMemsetAsync(d_out, 0, (size_t)F * F * sizeof(double), stream));
  if (s->n_slots == 0) return ORYX_OK;
  if (allow_tc && gram_tc_applicable(s)) {
    float absmax = 0.f;
    ORYX_TRY(gram_tc_absmax(s, stream, &absmax));
    if (absmax > 0.f && absmax < 3.0e38f) {  // a store holding +-inf goes to the floating-point kernels
      if (tc_absmax) *tc_absmax = absmax;
      return gram_tc_launch(m, s, d_out, stream);
    }
  }
  if (s->n_slots >= kGramFastMinRows) {
    ORYX_TRY(gram_fast_launch(m, s, d_out, stream));
    return gram_symmetrise(d_out, F, stream);
  }
  const int nb = (F + kTile - 1) / kTile;
  const int n_blocks = nb * (nb + 1) / 2;
  // enough row chunks to fill the GPU a few times over, but chunks no shorter than 256 rows
  int64_t want_chunks = std::max<int64_t>(1, ((int64_t)m->num_sms * 4 + n_blocks - 1) / n_blocks);
  int64_t max_chunks = std::max<int64_t>(1, s->n_slots / 256);
  int64_t chunks = std::min<int64_t>(std::min(want_chunks, max_chunks), 65535);
  int64_t rows_per_chunk = (s->n_slots + chunks - 1) / chunks;
  rows_per_chunk = (rows_per_chunk + kKC - 1) / kKC * kKC;
  chunks = (s->n_slots + rows_per_chunk - 1) / rows_per_chunk;
  dim3 grid((unsigned)n_blocks, (unsigned)chunks);
  gram_exact_kernel<<<grid, 256, 0, stream>>>(s->d_vec, s->d_slot_row, s->n_slots, s->fpad, F, nb, rows_per_chunk, d_out);
  m->launches++;
  ORYX_CUDA(cudaGetLastError());
  return gram_symmetrise(d_out, F, stream);
}

}  // namespace oryx
