// lsh.cu -- locality-sensitive hashing of feature vectors and candidate-partition enumeration.
//
// Replaces, bit-exactly, LocalitySensitiveHash.getIndexFor / getCandidateIndices and the sizing loop
// of its constructor (reference: app/oryx-app-serving/src/main/java/com/cloudera/oryx/app/serving/
// als/model/LocalitySensitiveHash.java:45-74, 110-117, 142-150, 156-177).
//
// Bit-exactness: bit i of the index is [dot(h_i, v) > 0.0] where dot() is VectorMath.dot
// (binary32 product, binary64 accumulate, index order; VectorMath.java:38-45).  The kernel performs
// exactly that operation sequence per (row, hash) with __fmul_rn / __dadd_rn.
#include <algorithm>

#include "common.cuh"

namespace oryx {

namespace {

int64_t binom(int n, int k) {
  if (k < 0 || k > n) return 0;
  if (k > n - k) k = n - k;
  int64_t r = 1;
  for (int i = 1; i <= k; i++) r = r * (n - k + i) / i;
  return r;
}

// One thread per row; the row tile and the hash vectors are staged in shared memory so global
// reads stay coalesced while each thread walks its own row sequentially (stride features+1 floats
// => conflict-free).  Hashes are processed four at a time for ILP across the dependent DADD chains.
__global__ void lsh_hash_kernel(const float* __restrict__ rows, int64_t n, int64_t stride, int features,
                                const float* __restrict__ hash_vectors, int hv_stride, int num_hashes,
                                int tile_rows, int32_t* __restrict__ out, int32_t* __restrict__ zero_counter) {
  extern __shared__ float smem[];
  float* s_hv = smem;                                  // [num_hashes x features]
  float* s_rows = smem + (size_t)num_hashes * features;  // [tile_rows x (features+1)]
  const int ld = features + 1;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  // a query's hashing kernel also resets the count of its work list (built by the next kernel)
  if (zero_counter && blockIdx.x == 0 && threadIdx.x == 0) *zero_counter = 0;

  for (int e = threadIdx.x; e < num_hashes * features; e += blockDim.x) {
    int h = e / features, j = e - h * features;
    s_hv[e] = hash_vectors[(size_t)h * hv_stride + j];
  }
  for (int64_t row0 = (int64_t)blockIdx.x * tile_rows; row0 < n; row0 += (int64_t)gridDim.x * tile_rows) {
    __syncthreads();
    for (int r = warp; r < tile_rows; r += nwarps) {
      int64_t row = row0 + r;
      if (row < n) {
        const float* src = rows + row * stride;
        for (int j = lane; j < features; j += 32) s_rows[r * ld + j] = src[j];
      }
    }
    __syncthreads();
    int t = threadIdx.x;
    if (t < tile_rows && row0 + t < n) {
      const float* y = s_rows + t * ld;
      int index = 0;
      for (int h0 = 0; h0 < num_hashes; h0 += 4) {
        double acc[4] = {0.0, 0.0, 0.0, 0.0};
        const float* hv[4];
#pragma unroll
        for (int k = 0; k < 4; k++) hv[k] = s_hv + (size_t)min(h0 + k, num_hashes - 1) * features;
        for (int j = 0; j < features; j++) {
          float v = y[j];
#pragma unroll
          for (int k = 0; k < 4; k++) acc[k] = exact_step(acc[k], hv[k][j], v);
        }
#pragma unroll
        for (int k = 0; k < 4; k++) {
          if (h0 + k < num_hashes && acc[k] > 0.0) index |= 1 << (h0 + k);
        }
      }
      out[row0 + t] = index;
    }
  }
}

// LSH.java:156-177 given the main index: one thread per integer in [0, 2^H).
//   H == maxBits  -> allIndices: natural order, NOT xor-ed (LSH.java:160-162)
//   maxBits == 0  -> { main }
//   otherwise     -> prototype order (ascending popcount, ascending value inside a band;
//                    LSH.java:110-117) truncated to popcount <= maxBits, each xor main.
__global__ void lsh_candidates_kernel(int num_hashes, int max_bits, const int32_t* __restrict__ main_index,
                                      int32_t* __restrict__ out, int32_t* __restrict__ out_n) {
  __shared__ int s_binom[ORYX_MAX_HASHES + 1][ORYX_MAX_HASHES + 1];
  for (int e = threadIdx.x; e < (ORYX_MAX_HASHES + 1) * (ORYX_MAX_HASHES + 1); e += blockDim.x) {
    int nn = e / (ORYX_MAX_HASHES + 1), kk = e % (ORYX_MAX_HASHES + 1);
    // Pascal's triangle entry computed directly (small n)
    long long r = (kk <= nn) ? 1 : 0;
    int k2 = kk > nn - kk ? nn - kk : kk;
    if (kk <= nn) for (int i = 1; i <= k2; i++) r = r * (nn - k2 + i) / i;
    s_binom[nn][kk] = (int)r;
  }
  __syncthreads();
  const int total = 1 << num_hashes;
  const int main_idx = *main_index;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    if (num_hashes == max_bits) {
      out[i] = i;
      if (i == 0) *out_n = total;
      continue;
    }
    if (max_bits == 0) {
      if (i == 0) { out[0] = main_idx; *out_n = 1; }
      continue;
    }
    int c = __popc(i);
    if (i == 0) {
      int how_many = 0;
      for (int b = 0; b <= max_bits; b++) how_many += s_binom[num_hashes][b];
      *out_n = how_many;
    }
    if (c > max_bits) continue;
    int offset = 0;
    for (int b = 0; b < c; b++) offset += s_binom[num_hashes][b];
    // rank of i among same-popcount integers in ascending order (combinatorial number system)
    int rank = 0, k = c;
    for (int b = num_hashes - 1; b >= 0 && k > 0; b--) {
      if (i & (1 << b)) { rank += s_binom[b][k]; k--; }
    }
    out[offset + rank] = i ^ main_idx;
  }
}

// Work list of an LSH-filtered scan: ids of the 256-slot blocks that hold (or may hold) rows of a
// candidate partition.  worklist[0] is the count (zeroed by the launcher), worklist[1..] the ids, in
// no particular order.
__global__ void lsh_worklist_kernel(const int32_t* __restrict__ block_part, int64_t n_blocks,
                                    const int32_t* __restrict__ main_index, int max_bits,
                                    int32_t* __restrict__ worklist) {
  const int main_idx = *main_index;
  for (int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; b < n_blocks; b += (int64_t)gridDim.x * blockDim.x) {
    const int bp = block_part[b];
    const bool want = bp == -2 || (bp >= 0 && __popc(bp ^ main_idx) <= max_bits);
    // warp-aggregated append
    const unsigned ball = __ballot_sync(__activemask(), want);
    if (want) {
      const int lane = threadIdx.x & 31;
      const int leader = __ffs(ball) - 1;
      int base = 0;
      if (lane == leader) base = atomicAdd(worklist, __popc(ball));
      base = __shfl_sync(ball, base, leader);
      worklist[1 + base + __popc(ball & ((1u << lane) - 1u))] = (int32_t)b;
    }
  }
}

}  // namespace

// LSH.java:45-74.  Configuration-time integer arithmetic (runs once per model), kept on the host.
int32_t lsh_plan(double sample_rate, int num_cores, int* hashes, int* bits) {
  int num_hashes = 0, bits_differing = 0;
  for (; num_hashes < ORYX_MAX_HASHES; num_hashes++) {
    bits_differing = 0;
    int64_t to_try = 1;
    while (bits_differing < num_hashes && to_try < num_cores) {
      bits_differing++;
      to_try += binom(num_hashes, bits_differing);
    }
    if (bits_differing == num_hashes && to_try < num_cores) continue;
    if ((double)to_try <= sample_rate * (double)((int64_t)1 << num_hashes)) break;
  }
  *hashes = num_hashes;
  *bits = bits_differing;
  return ORYX_OK;
}

int32_t lsh_hash_rows_device(Model* m, const float* d_rows, int64_t n, int stride, int32_t* d_out,
                             cudaStream_t stream, int32_t* d_zero_counter, bool early_start) {
  if (n <= 0) return ORYX_OK;
  const int F = m->cfg.features;
  const int H = m->num_hashes;
  if (H == 0) {
    ORYX_CUDA(cudaMemsetAsync(d_out, 0, (size_t)n * sizeof(int32_t), stream));
    if (d_zero_counter) ORYX_CUDA(cudaMemsetAsync(d_zero_counter, 0, sizeof(int32_t), stream));
    return ORYX_OK;
  }
  size_t hv_bytes = (size_t)H * F * sizeof(float);
  size_t budget = 200 * 1024 - hv_bytes;
  int tile_rows = (int)std::min<size_t>(128, budget / ((size_t)(F + 1) * sizeof(float)));
  tile_rows = tile_rows / 32 * 32;
  if (tile_rows < 32) {
    set_error("features=%d too large for the LSH hashing kernel's shared-memory tile", F);
    return ORYX_E_UNSUPPORTED;
  }
  size_t smem = hv_bytes + (size_t)tile_rows * (F + 1) * sizeof(float);
  ORYX_TRY(ensure_max_dynamic_smem(reinterpret_cast<const void*>(lsh_hash_kernel), smem));
  int64_t tiles = (n + tile_rows - 1) / tile_rows;
  int grid = (int)std::min<int64_t>(tiles, (int64_t)m->num_sms * 8);
  // `early_start`: the kernel depends on nothing the stream's previous kernels write (a query vector
  // and the hash vectors), so it may begin as soon as they release their programmatic dependents.
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3((unsigned)grid);
  cfg.blockDim = dim3((unsigned)tile_rows);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = early_start ? 1 : 0;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  ORYX_CUDA(cudaLaunchKernelEx(&cfg, lsh_hash_kernel, d_rows, n, (int64_t)stride, F, (const float*)m->d_hash_vectors,
                               m->store[ORYX_Y].fpad, H, tile_rows, d_out, d_zero_counter));
  m->launches++;
  return ORYX_OK;
}

// The count d_worklist[0] must already be zero (the query's hashing kernel resets it).
int32_t lsh_worklist_device(Model* m, const int32_t* d_main_index, int32_t* d_worklist, cudaStream_t stream) {
  Store* s = &m->store[ORYX_Y];
  const int64_t n_blocks = (s->n_slots + kBlockRows - 1) / kBlockRows;
  const int threads = 256;
  const int grid = (int)std::min<int64_t>((n_blocks + threads - 1) / threads, (int64_t)m->num_sms * 4);
  lsh_worklist_kernel<<<grid, threads, 0, stream>>>(s->d_block_part, n_blocks, d_main_index, m->max_bits, d_worklist);
  m->launches++;
  ORYX_CUDA(cudaGetLastError());
  return ORYX_OK;
}

int32_t lsh_candidates_device(Model* m, const int32_t* d_main_index, int32_t* d_out, int32_t* d_out_n,
                              cudaStream_t stream) {
  int total = 1 << m->num_hashes;
  int threads = 256;
  int grid = std::min((total + threads - 1) / threads, 64);
  lsh_candidates_kernel<<<grid, threads, 0, stream>>>(m->num_hashes, m->max_bits, d_main_index, d_out, d_out_n);
  m->launches++;
  ORYX_CUDA(cudaGetLastError());
  return ORYX_OK;
}

}  // namespace oryx
