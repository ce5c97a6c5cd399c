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

}  // namespace

int32_t gram_launch(Model* m, int which, double* d_out, cudaStream_t stream) {
  Store* s = &m->store[which];
  const int F = s->features;
  ORYX_CUDA(cudaMemsetAsync(d_out, 0, (size_t)F * F * sizeof(double), stream));
  if (s->n_slots == 0) return ORYX_OK;
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
  return ORYX_OK;
}

}  // namespace oryx
