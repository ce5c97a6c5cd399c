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

// One thread per row; the row tile and the hash vectors are staged in shared memory so global reads stay
// coalesced while each thread walks its own row (row stride a multiple of 4 floats with an odd number of
// 16-byte units: 128-bit accesses of a quarter warp then cover all 32 banks).
//
// Only the SIGN of each dot product enters the index, so the kernel first evaluates all hashes of a row
// with fp32 FMAs (H + 1 independent chains: the H dot products and |v|^2) and accepts a sign whenever the
// fp32 value is further from zero than a rigorous bound on its distance from the reference's value:
//   reference  R = sum_j fl32(h_j v_j), accumulated in binary64      |R - sum h_j v_j| <= (2^-24 + F 2^-53) sum |h_j v_j|
//   fast       D = fp32 FMA chain                                    |D - sum h_j v_j| <= F 2^-24 / (1 - F 2^-24) sum |h_j v_j|
//   sum |h_j v_j| <= |h| |v| (Cauchy-Schwarz), |v|^2 from its own fp32 chain (relative error <= F 2^-24)
//   =>  |D - R| <= bound = 1.02 (F + 2) 2^-24 |h|_up sqrt(|v|^2_fp32)        (|h|_up: the norm rounded up, per hash)
// |D| > bound  =>  sign(R) = sign(D) and R != 0.  Otherwise (about one pair in 10^4 for Gaussian data, plus
// zero, tiny, NaN and infinite cases) the thread repeats that hash with the reference's exact operation
// sequence (__fmul_rn, widen, __dadd_rn, index order).  The resulting index is bit-identical either way.
constexpr int kLshGroup = 8;  // hashes evaluated per pass over a row

__device__ __forceinline__ void cp_async_16(float* smem_dst, const float* gmem_src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_4(float* smem_dst, const float* gmem_src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)), "l"(gmem_src) : "memory");
}

template <int G>
__device__ __forceinline__ int lsh_group_bits(const float* __restrict__ y, const float* __restrict__ s_hv, int ld,
                                              const float* __restrict__ s_hnorm, int h0, int f4, int features,
                                              float bound_scale) {
  float dot[G], v2 = 0.f;
#pragma unroll
  for (int k = 0; k < G; k++) dot[k] = 0.f;
  const float4* y4 = reinterpret_cast<const float4*>(y);
  const float4* hp[G];  // row pointers hoisted: the loop body is loads and FMAs only
#pragma unroll
  for (int k = 0; k < G; k++) hp[k] = reinterpret_cast<const float4*>(s_hv + (h0 + k) * ld);
  // all loads of a quad first, then the FMAs component by component across the G chains: consecutive
  // instructions are independent (issuing the four FMAs of one chain back to back left the warp waiting on
  // the FMA latency), and the unrolled loop has the next quad's loads in flight meanwhile
#pragma unroll 4
  for (int j = 0; j < f4; j++) {
    const float4 v = y4[j];
    float4 h[G];
#pragma unroll
    for (int k = 0; k < G; k++) h[k] = hp[k][j];  // same address for the whole warp
    v2 = fmaf(v.x, v.x, v2);
#pragma unroll
    for (int k = 0; k < G; k++) dot[k] = fmaf(v.x, h[k].x, dot[k]);
    v2 = fmaf(v.y, v.y, v2);
#pragma unroll
    for (int k = 0; k < G; k++) dot[k] = fmaf(v.y, h[k].y, dot[k]);
    v2 = fmaf(v.z, v.z, v2);
#pragma unroll
    for (int k = 0; k < G; k++) dot[k] = fmaf(v.z, h[k].z, dot[k]);
    v2 = fmaf(v.w, v.w, v2);
#pragma unroll
    for (int k = 0; k < G; k++) dot[k] = fmaf(v.w, h[k].w, dot[k]);
  }
  const float vnorm = sqrtf(v2) * bound_scale;
  int bits = 0;
#pragma unroll
  for (int k = 0; k < G; k++) {
    const float bound = vnorm * s_hnorm[h0 + k];
    bool positive;
    if (fabsf(dot[k]) > bound && fabsf(dot[k]) > 1e-30f && bound < 3.0e38f) {
      positive = dot[k] > 0.f;
    } else {
      // too close to zero to call (or NaN / inf / denormal range): the reference's arithmetic, exactly
      const float* hv = s_hv + (size_t)(h0 + k) * ld;
      double acc = 0.0;
      for (int j = 0; j < features; j++) acc = exact_step(acc, hv[j], y[j]);
      positive = acc > 0.0;
    }
    if (positive) bits |= 1 << (h0 + k);
  }
  return bits;
}

__global__ void lsh_hash_kernel(const float* __restrict__ rows, int64_t n, int64_t stride, int features,
                                const float* __restrict__ hash_vectors, int hv_stride, int num_hashes,
                                int tile_rows, int ld, int32_t* __restrict__ out, int32_t* __restrict__ zero_counter) {
  extern __shared__ __align__(16) float smem[];
  float* s_hv = smem;                                  // [num_hashes x ld], zero padded
  float* s_hnorm = s_hv + (size_t)num_hashes * ld;     // [16] norms, rounded up
  float* s_rows = s_hnorm + 16;                        // [tile_rows x ld], zero padded
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  const int f4 = (features + 3) >> 2;
  // a query's hashing kernel also resets the count of its work list (built by the next kernel)
  if (zero_counter && blockIdx.x == 0 && threadIdx.x == 0) *zero_counter = 0;

  for (int e = threadIdx.x; e < num_hashes * ld; e += blockDim.x) {
    int h = e / ld, j = e - h * ld;
    s_hv[e] = j < features ? hash_vectors[(size_t)h * hv_stride + j] : 0.f;
  }
  __syncthreads();
  if (threadIdx.x < num_hashes) {
    double ss = 0.0;
    for (int j = 0; j < features; j++) { const double v = (double)s_hv[(size_t)threadIdx.x * ld + j]; ss += v * v; }
    s_hnorm[threadIdx.x] = __double2float_ru(sqrt(ss) * (1.0 + 1e-12));
  }
  const float bound_scale = 1.02f * (float)(features + 2) * 5.9604645e-08f;  // 1.02 (F + 2) 2^-24
  const bool vec4 = (stride & 3) == 0 && ((uintptr_t)rows & 15) == 0;
  for (int64_t row0 = (int64_t)blockIdx.x * tile_rows; row0 < n; row0 += (int64_t)gridDim.x * tile_rows) {
    __syncthreads();
    // asynchronous global -> shared copies: every row of the tile is in flight at once (a load-then-store
    // loop waited for each row's load in turn and left the kernel bound by memory latency)
    const int rows_here = (int)min((int64_t)tile_rows, n - row0);
    if (vec4 && (features & 3) == 0) {
      // the common case (rows 16-byte aligned, features a multiple of 4): one flat loop over the tile's
      // 16-byte units, a handful of instructions per copy
      const int q = features >> 2;
      const float* src0 = rows + row0 * stride;
      if ((q & (q - 1)) == 0) {
        const int sh = 31 - __clz(q);
        for (int e = threadIdx.x; e < rows_here * q; e += blockDim.x) {
          const int r = e >> sh, j = e & (q - 1);
          cp_async_16(s_rows + r * ld + 4 * j, src0 + (int64_t)r * stride + 4 * j);
        }
      } else {
        for (int e = threadIdx.x; e < rows_here * q; e += blockDim.x) {
          const int r = e / q, j = e - r * q;
          cp_async_16(s_rows + r * ld + 4 * j, src0 + (int64_t)r * stride + 4 * j);
        }
      }
    } else {
      for (int r = warp; r < rows_here; r += nwarps) {
        const float* src = rows + (row0 + r) * stride;
        float* dst = s_rows + (size_t)r * ld;
        for (int j = lane; j < 4 * f4; j += 32) {
          if (j < features) cp_async_4(dst + j, src + j); else dst[j] = 0.f;
        }
      }
    }
    asm volatile("cp.async.wait_all;" ::: "memory");
    __syncthreads();
    const int t = threadIdx.x;
    if (t < tile_rows && row0 + t < n) {
      const float* y = s_rows + (size_t)t * ld;
      int index = 0;
      for (int h0 = 0; h0 < num_hashes; h0 += kLshGroup) {
        switch (min(kLshGroup, num_hashes - h0)) {
          case 1: index |= lsh_group_bits<1>(y, s_hv, ld, s_hnorm, h0, f4, features, bound_scale); break;
          case 2: index |= lsh_group_bits<2>(y, s_hv, ld, s_hnorm, h0, f4, features, bound_scale); break;
          case 3: index |= lsh_group_bits<3>(y, s_hv, ld, s_hnorm, h0, f4, features, bound_scale); break;
          case 4: index |= lsh_group_bits<4>(y, s_hv, ld, s_hnorm, h0, f4, features, bound_scale); break;
          case 5: index |= lsh_group_bits<5>(y, s_hv, ld, s_hnorm, h0, f4, features, bound_scale); break;
          case 6: index |= lsh_group_bits<6>(y, s_hv, ld, s_hnorm, h0, f4, features, bound_scale); break;
          case 7: index |= lsh_group_bits<7>(y, s_hv, ld, s_hnorm, h0, f4, features, bound_scale); break;
          default: index |= lsh_group_bits<8>(y, s_hv, ld, s_hnorm, h0, f4, features, bound_scale); break;
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

// A query's LSH index AND the work list of its scan, in one launch (owning contexts, scan_topn.cu).  Every
// CTA hashes the query for itself -- one warp per hash, lanes striding over the features, the same
// fast-path/bound/exact-fallback rule as lsh_group_bits (the bound holds for any summation order of depth
// <= F; the norms are reduced the same way) -- and then appends its share of the candidate blocks.
// Two separately launched kernels (hash on one warp, then the list) left the GPU idle for ~30 us between
// two filtered scans (round 2: 0.319 ms per query on a 12.5 M x 256 shard against 0.285 ms for an unfiltered
// store of the candidate rows alone).
//
// The list's count must be zero at launch: lists alternate per query and this kernel zeroes the count of the
// OTHER list (its scan has completed: this kernel is released by that scan's finishing kernel), which the
// next query appends to.  The scan that follows is launched as a programmatic dependent and normally does NOT
// call griddepcontrol.wait (that would make it wait for the previous query's finishing kernel as well): this
// kernel never releases its dependents early, every thread fences its writes, and every CTA then adds one to a
// per-context counter with a release reduction (fire and forget: nobody waits for its round trip).  The scan's
// CTAs -- launched after the last CTA of this grid has exited -- read the counter with an acquire load: the
// expected total (all CTAs of all list kernels this context ever launched) means every write of this grid is
// visible to them; should it ever fall short (nothing in the launch order itself promises otherwise), they fall
// back to griddepcontrol.wait.
__global__ void __launch_bounds__(256)
lsh_query_worklist_kernel(const float* __restrict__ q, int features, const float* __restrict__ hash_vectors,
                          int hv_stride, int num_hashes, const int32_t* __restrict__ block_part, int64_t n_blocks,
                          int max_bits, int32_t* __restrict__ main_index_out, int32_t* __restrict__ worklist,
                          int32_t* __restrict__ other_count, unsigned int* __restrict__ done_ctas) {
  __shared__ int s_bits;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  if (threadIdx.x == 0) s_bits = 0;
  __syncthreads();
  const float bound_scale = 1.02f * (float)(features + 2) * 5.9604645e-08f;  // 1.02 (F + 2) 2^-24
  for (int h = warp; h < num_hashes; h += nwarps) {
    const float* hv = hash_vectors + (size_t)h * hv_stride;
    float dot = 0.f, v2 = 0.f;
    double h2 = 0.0;
    for (int j = lane; j < features; j += 32) {
      const float a = __ldg(hv + j), b = __ldg(q + j);
      dot = fmaf(a, b, dot);
      v2 = fmaf(b, b, v2);
      h2 += (double)a * (double)a;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      dot += __shfl_xor_sync(0xffffffffu, dot, o);
      v2 += __shfl_xor_sync(0xffffffffu, v2, o);
      h2 += __shfl_xor_sync(0xffffffffu, h2, o);
    }
    if (lane == 0) {
      const float bound = sqrtf(v2) * bound_scale * __double2float_ru(sqrt(h2) * (1.0 + 1e-12));
      bool positive;
      if (fabsf(dot) > bound && fabsf(dot) > 1e-30f && bound < 3.0e38f) {
        positive = dot > 0.f;
      } else {  // too close to call: the reference's arithmetic, exactly
        double acc = 0.0;
        for (int j = 0; j < features; j++) acc = exact_step(acc, __ldg(hv + j), __ldg(q + j));
        positive = acc > 0.0;
      }
      if (positive) atomicOr(&s_bits, 1 << h);
    }
  }
  __syncthreads();
  const int main_idx = s_bits;
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    *main_index_out = main_idx;
    *other_count = 0;
  }
  for (int64_t b0 = (int64_t)blockIdx.x * blockDim.x; b0 < n_blocks; b0 += (int64_t)gridDim.x * blockDim.x) {
    const int64_t b = b0 + threadIdx.x;
    const int bp = b < n_blocks ? __ldg(block_part + b) : -1;
    const bool want = bp == -2 || (bp >= 0 && __popc(bp ^ main_idx) <= max_bits);
    const unsigned ball = __ballot_sync(0xffffffffu, want);  // warp-aggregated append
    if (ball != 0u) {
      int base = 0;
      if (lane == 0) base = atomicAdd(worklist, __popc(ball));
      base = __shfl_sync(0xffffffffu, base, 0);
      if (want) worklist[1 + base + __popc(ball & ((1u << lane) - 1u))] = (int32_t)b;
    }
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(done_ctas) : "memory");
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
  // shared-memory row stride: features rounded up to 4, then to an odd number of 16-byte units
  int ld = round_up(F, 4);
  if (((ld >> 2) & 1) == 0) ld += 4;
  const size_t hv_bytes = ((size_t)H * ld + 16) * sizeof(float);
  const size_t budget = 200 * 1024 - hv_bytes;
  // up to 128 rows per tile, but small enough that three CTAs fit on an SM (their load and compute phases
  // then overlap); never below one warp
  int tile_rows = (int)std::min<size_t>(128, std::min<size_t>(budget, 70 * 1024) / ((size_t)ld * sizeof(float)));
  tile_rows = tile_rows / 32 * 32;
  if (tile_rows < 32 && budget / ((size_t)ld * sizeof(float)) >= 32) tile_rows = 32;
  if (tile_rows < 32) {
    set_error("features=%d too large for the LSH hashing kernel's shared-memory tile", F);
    return ORYX_E_UNSUPPORTED;
  }
  ORYX_TRY(ensure_max_dynamic_smem(reinterpret_cast<const void*>(lsh_hash_kernel), hv_bytes + (size_t)tile_rows * ld * sizeof(float)));
  if (n < tile_rows) tile_rows = (int)round_up((int)n, 32);  // a query: one warp is enough
  const size_t smem = hv_bytes + (size_t)tile_rows * ld * sizeof(float);
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
                               m->store[ORYX_Y].fpad, H, tile_rows, ld, d_out, d_zero_counter));
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

// Hash the query at d_query and list the candidate blocks of Y in `d_worklist` (count at [0], which must be
// zero: see lsh_query_worklist_kernel); `d_other_count` is the count of the context's other list.
int32_t lsh_query_worklist_device(Model* m, const float* d_query, int32_t* d_main_index, int32_t* d_worklist,
                                  int32_t* d_other_count, unsigned int* d_done_ctas, unsigned int* expected_total,
                                  cudaStream_t stream) {
  Store* s = &m->store[ORYX_Y];
  const int64_t n_blocks = (s->n_slots + kBlockRows - 1) / kBlockRows;
  const int threads = 256;
  const int grid = (int)std::max<int64_t>(1, std::min<int64_t>((n_blocks + threads - 1) / threads, (int64_t)m->num_sms * 4));
  // The kernel depends on nothing the stream's previous kernels write (a query vector, the hash vectors, the
  // block table; both lists' previous readers have completed once the previous finishing kernel is running),
  // so it may begin as soon as they release their programmatic dependents.
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3((unsigned)grid);
  cfg.blockDim = dim3((unsigned)threads);
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  ORYX_CUDA(cudaLaunchKernelEx(&cfg, lsh_query_worklist_kernel, d_query, m->cfg.features,
                               (const float*)m->d_hash_vectors, s->fpad, m->num_hashes,
                               (const int32_t*)s->d_block_part, n_blocks, m->max_bits, d_main_index, d_worklist,
                               d_other_count, d_done_ctas));
  *expected_total += (unsigned int)grid;
  m->launches++;
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
