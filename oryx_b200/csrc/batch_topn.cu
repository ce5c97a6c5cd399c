// batch_topn.cu -- batched top-N over the whole item store on the 5th-generation tensor cores
// (kernel K3 of SURVEY.md section 2.3).
//
// The reference has no batched call: concurrent /recommend requests, /recommendToMany-style fan-outs
// and the PopularRepresentativeItems loop each run ALSServingModel.topN once per query
// (ALSServingModel.java:289-335, PopularRepresentativeItems.java:52-57).  oryx_topn_batch answers nq
// such calls at once; per query the result is the same list the single-query path returns.
//
// Y is read once per batch instead of once per query, so the work is a dense contraction S = Q Y^T
// (2*nq*N*F flops) and belongs on tcgen05.  The 4096 x 10M score matrix is never materialised:
//
//   1. convert   Y (fp32 store) and Q -> bf16 scratch, K padded to 64; the conversions also MEASURE what the
//                rounding took away, which gives a rigorous bound on what it can do to a score:
//                s_bf16 - s = bf(q).(bf(y) - y) + (bf(q) - q).y, hence
//                |s_bf16 - s| <= delta_q = ||bf(q)|| max_j ||y_j - bf(y_j)|| + ||q - bf(q)|| max_j ||y_j||
//                (+ fp32 accumulation), about 2.5 x below the worst case 2^-7 ||q|| max ||y|| of round 1.
//   2. score     batch_score_kernel: persistent, warp-specialised.  One warp drives TMA
//                (cp.async.bulk.tensor, 128-byte swizzle), one thread issues tcgen05.mma
//                (cta_group::1, kind::f16, M=128 queries x N=256 items x K=16, accumulators in TMEM,
//                two accumulator stages), eight warps run the epilogue: tcgen05.ld the accumulator
//                and keep only the MAXIMUM of every 32-item sub-block per query -> bmax (1/32 of the
//                scores, tile-major [tile][column half][query][4] so that a warp stores 512 contiguous
//                bytes) plus the maximum of the whole tile -> tmax ([tile / 32][query][32]: the threshold
//                step reads whole 128-byte lines per query).  The Y tile stays resident in shared memory
//                while all query tiles stream past it.  Measured (round 2, 4096 x 10 M x 128): 6.8 ms;
//                the same kernel with the epilogue removed 6.06 ms, with query-major outputs 7.45 ms.
//   3. threshold per query: T' = k-th largest sub-block maximum (k sub-blocks each hold an item with
//                bf16 score >= T', so the true k-th best exact score is >= T' - delta).  Every item of
//                the true top k therefore lives in a sub-block whose maximum is >= T' - 2 delta: those
//                sub-blocks are the candidates (typically ~1.2 k of them).
//   4. refine    one CTA per query re-scores the candidate sub-blocks in fp32 straight from the fp32
//                store, keeps the k best, re-scores those with the reference's exact arithmetic and
//                ranks them -- the same finishing pass as the single-query path, so both paths return
//                identical rows and bit-identical scores.
//
// Candidate pruning is exact (the bound is rigorous), so unlike a plain reduced-precision scorer
// this path needs no recall caveat: recall@N against the fp32 path is 1.0 by construction, and the
// tests check it.
#include <cuda.h>
#include <cuda_bf16.h>

#include <algorithm>
#include <cstdlib>

#include "common.cuh"

namespace oryx {

namespace {

constexpr int BM = 128;        // queries per Q tile (UMMA M)
constexpr int BN = 256;        // items per Y tile (UMMA N)
constexpr int BK = 64;         // bf16 elements per K chunk = one 128-byte swizzle row
constexpr int UMMA_K = 16;
constexpr int QSTAGES = 4;     // Q chunk ring
constexpr int SUB = 32;        // items per sub-block whose maximum is kept
constexpr int SUBS_PER_TILE = BN / SUB;  // 8
constexpr int kScoreThreads = 320;       // warp 0: TMA, warp 1: MMA + TMEM alloc, warps 2-9: epilogue
constexpr int kEpiWarps = 8;             // two per TMEM lane quarter: each reads half of the accumulator's columns
constexpr uint32_t Y_CHUNK_BYTES = BN * BK * 2;  // 32 KB
constexpr uint32_t Q_CHUNK_BYTES = BM * BK * 2;  // 16 KB

// ---------------------------------------------------------------- PTX wrappers (sm_100a)
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_LOOP:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra WAIT_DONE;\n"
      "bra WAIT_LOOP;\n"
      "WAIT_DONE:\n"
      "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* map, int c0, int c1, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
      ::"r"(smem_u32(dst)), "l"(map), "r"(c0), "r"(c1), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tcgen05_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tcgen05_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tcgen05_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// D[tmem] (+)= A[smem] * B[smem], bf16 inputs, fp32 accumulate, issued by one thread.
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
      "}\n" ::"r"(tmem_d), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate) : "memory");
}
// K-major operand tile, 128-byte swizzle, 8-row atoms packed 1024 B apart.
__device__ __forceinline__ uint64_t umma_desc_k_sw128(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr >> 4) & 0x3FFF);   // start address, 16-byte units
  d |= (uint64_t)1 << 16;                       // leading byte offset (unused for swizzled K-major)
  d |= (uint64_t)(1024 >> 4) << 32;             // stride byte offset between 8-row atoms
  d |= (uint64_t)1 << 46;                       // descriptor version (Blackwell)
  d |= (uint64_t)2 << 61;                       // SWIZZLE_128B
  return d;
}
// kind::f16 instruction descriptor: D=f32, A=B=bf16, both K-major, M=128, N=256.
constexpr uint32_t kIdesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);

__device__ __forceinline__ void tmem_ld_32x32b_x32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
        "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
        "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// ---------------------------------------------------------------- step 1: conversions
// Y fp32 [n x fpad] -> bf16 [n x kpad] (zero padded), and the largest row norm.  One warp per row.
// For the cosine score (dot(q, y) / ||y||, CosineAverageFunction.java:54-56) rows are divided by their
// norm first, so the tensor-core pass scores q . (y/||y||) and every "row norm" is 1.
__global__ void convert_y_kernel(const float* __restrict__ y, int64_t n, int fpad, int features, int kpad,
                                 int normalize, __nv_bfloat16* __restrict__ out,
                                 unsigned int* __restrict__ max_norm_bits) {
  const int lane = threadIdx.x & 31;
  const int64_t w0 = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int64_t stride = (int64_t)gridDim.x * (blockDim.x >> 5);
  float local_max = 0.f, local_res = 0.f;
  for (int64_t row = w0; row < n; row += stride) {
    const float* src = y + row * fpad;
    __nv_bfloat16* dst = out + row * kpad;
    float ss = 0.f;
    for (int j = lane; j < features; j += 32) { float v = src[j]; ss = fmaf(v, v, ss); }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
    const float scale = normalize ? rsqrtf(ss) : 1.0f;  // zero row: inf -> NaN entries, never selected
    float rs = 0.f;  // squared norm of what the rounding to bf16 took away from this row
    for (int j = lane; j < kpad; j += 32) {
      float v = j < features ? src[j] * scale : 0.f;
      if (normalize && j >= features) v = 0.f;
      const __nv_bfloat16 b = __float2bfloat16_rn(v);
      dst[j] = b;
      const float d = v - __bfloat162float(b);  // exact in binary32
      rs = fmaf(d, d, rs);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) rs += __shfl_xor_sync(0xffffffffu, rs, o);
    if (ss == ss) {  // tombstones are NaN rows: ignore
      local_max = fmaxf(local_max, normalize ? 1.0002f : ss);
      local_res = fmaxf(local_res, rs);
    }
  }
  // sqrt is monotone: keep the max of squared norms (non-negative floats order like their bit patterns)
  if (lane == 0) {
    atomicMax(max_norm_bits, __float_as_uint(local_max));
    atomicMax(max_norm_bits + 1, __float_as_uint(local_res));
  }
}

// The same conversion for a list of slots that just changed (upsert / remove): keeps the shadow current
// under a trickle of updates instead of rebuilding 2.6 GB per change.  max_norm_bits only ever grows,
// which keeps the error bound valid (a removed row's norm may linger in it: looser, never wrong).
__global__ void convert_rows_kernel(const float* __restrict__ y, const int32_t* __restrict__ slots, int64_t n_changed,
                                    int64_t shadow_rows, int fpad, int features, int kpad, int normalize,
                                    __nv_bfloat16* __restrict__ out, unsigned int* __restrict__ max_norm_bits) {
  const int lane = threadIdx.x & 31;
  const int64_t w = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (w >= n_changed) return;
  const int64_t row = slots[w];
  if (row < 0 || row >= shadow_rows) return;
  const float* src = y + row * fpad;
  __nv_bfloat16* dst = out + row * kpad;
  float ss = 0.f;
  for (int j = lane; j < features; j += 32) { float v = src[j]; ss = fmaf(v, v, ss); }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
  const float scale = normalize ? rsqrtf(ss) : 1.0f;
  float rs = 0.f;
  for (int j = lane; j < kpad; j += 32) {
    const float v = j < features ? src[j] * scale : 0.f;
    const __nv_bfloat16 b = __float2bfloat16_rn(v);
    dst[j] = b;
    const float d = v - __bfloat162float(b);
    rs = fmaf(d, d, rs);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) rs += __shfl_xor_sync(0xffffffffu, rs, o);
  if (lane == 0 && ss == ss) {
    atomicMax(max_norm_bits, __float_as_uint(normalize ? 1.0002f : ss));
    atomicMax(max_norm_bits + 1, __float_as_uint(rs));
  }
}

// One warp per query: the row in slots[i] of a store (zeros for a negative slot) -> packed [nq x f].
__global__ void gather_queries_kernel(const float* __restrict__ vec, int fpad, int f, const int32_t* __restrict__ slots,
                                      int nq, float* __restrict__ out) {
  const int w = (int)(((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5);
  const int lane = threadIdx.x & 31;
  if (w >= nq) return;
  const int32_t slot = slots[w];
  for (int j = lane; j < f; j += 32) out[(size_t)w * f + j] = slot >= 0 ? vec[(int64_t)slot * fpad + j] : 0.f;
}

// Q fp32 [nq x features] -> bf16 [nq_pad x kpad] (zero padded rows and columns), and per row the three norms
// the error bound needs: qnorm[row] = ||q||, qnorm[nq_pad + row] = ||q - bf16(q)||, qnorm[2 nq_pad + row] = ||bf16(q)||.
__global__ void convert_q_kernel(const float* __restrict__ q, int nq, int nq_pad, int features, int kpad,
                                 __nv_bfloat16* __restrict__ out, float* __restrict__ qnorm) {
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= nq_pad) return;
  float ss = 0.f, rs = 0.f, bs = 0.f;
  for (int j = lane; j < kpad; j += 32) {
    float v = (row < nq && j < features) ? q[(size_t)row * features + j] : 0.f;
    ss = fmaf(v, v, ss);
    const __nv_bfloat16 b = __float2bfloat16_rn(v);
    out[(size_t)row * kpad + j] = b;
    const float bf = __bfloat162float(b), d = v - bf;  // d: exact in binary32
    rs = fmaf(d, d, rs);
    bs = fmaf(bf, bf, bs);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    ss += __shfl_xor_sync(0xffffffffu, ss, o);
    rs += __shfl_xor_sync(0xffffffffu, rs, o);
    bs += __shfl_xor_sync(0xffffffffu, bs, o);
  }
  if (lane == 0) {
    qnorm[row] = sqrtf(ss);
    qnorm[nq_pad + row] = sqrtf(rs);
    qnorm[2 * nq_pad + row] = sqrtf(bs);
  }
}

// ---------------------------------------------------------------- step 2: tensor-core scoring
struct ScoreSmem {
  // dynamic shared memory, 1024-byte aligned: [YB][NCH] Y chunks, [QSTAGES] Q chunks, then barriers
  static __host__ __device__ size_t bytes(int nch, int yb) {
    return (size_t)yb * nch * Y_CHUNK_BYTES + (size_t)QSTAGES * Q_CHUNK_BYTES + 1280 + 1024;
  }
};

// LSH filtering (template flag): query q only competes for rows whose partition p satisfies
// popcount(p ^ main_q) <= max_bits (LocalitySensitiveHash.getCandidateIndices, LSH.java:156-177).  Y tiles are
// the store's 256-slot blocks, so block_part says per tile whether all live rows share one partition
// (one test per query and tile; non-candidate tiles are not even read back from TMEM), hold no live row, or
// are mixed (-2: per-row test against slot_part).  After oryx_items_compact nearly every tile is pure.
struct BatchLsh {
  const int32_t* block_part;
  const int32_t* slot_part;
  const int32_t* qmain;  // [nq_pad] LSH index of every query
  int max_bits;
};

template <int NCH, bool LSH>
__global__ void __launch_bounds__(kScoreThreads, 1)
batch_score_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_y,
                   int n_qtiles, int n_ytiles, int64_t n_slots, float* __restrict__ bmax, int64_t bmax_stride,
                   float* __restrict__ tmax, int64_t tmax_stride, const BatchLsh lsh) {
  constexpr int YB = (NCH <= 2) ? 2 : 1;  // double-buffer the resident Y tile when it fits
  extern __shared__ unsigned char smem_dyn[];
  unsigned char* base = reinterpret_cast<unsigned char*>(((uintptr_t)smem_dyn + 1023) & ~(uintptr_t)1023);
  unsigned char* ybuf = base;                                            // [YB][NCH][32 KB]
  unsigned char* qbuf = base + (size_t)YB * NCH * Y_CHUNK_BYTES;         // [QSTAGES][16 KB]
  uint64_t* bars = reinterpret_cast<uint64_t*>(qbuf + (size_t)QSTAGES * Q_CHUNK_BYTES);
  uint64_t* y_full = bars;                 // [2]
  uint64_t* y_empty = bars + 2;            // [2]
  uint64_t* q_full = bars + 4;             // [QSTAGES]
  uint64_t* q_empty = bars + 4 + QSTAGES;  // [QSTAGES]
  uint64_t* t_full = bars + 4 + 2 * QSTAGES;   // [2]
  uint64_t* t_empty = bars + 6 + 2 * QSTAGES;  // [2]
  uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(bars + 8 + 2 * QSTAGES);
  float* half_max = reinterpret_cast<float*>(tmem_holder + 2);  // [2][BM]: upper column half's maximum, per accumulator

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tm_q) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tm_y) : "memory");
    for (int i = 0; i < 2; i++) { mbar_init(&y_full[i], 1); mbar_init(&y_empty[i], 1); }
    for (int i = 0; i < QSTAGES; i++) { mbar_init(&q_full[i], 1); mbar_init(&q_empty[i], 1); }
    for (int i = 0; i < 2; i++) { mbar_init(&t_full[i], 1); mbar_init(&t_empty[i], kEpiWarps); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {  // TMEM: all 512 columns = two 128 x 256 fp32 accumulators
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(tmem_holder)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_holder;

  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      uint32_t qs = 0, qphase = 0;
      int it = 0;
      for (int t = blockIdx.x; t < n_ytiles; t += gridDim.x, it++) {
        const int yb = it % YB;
        const uint32_t yph = (uint32_t)(it / YB) & 1u;
        mbar_wait(&y_empty[yb], yph ^ 1u);
        mbar_expect_tx(&y_full[yb], NCH * Y_CHUNK_BYTES);
#pragma unroll
        for (int c = 0; c < NCH; c++)
          tma_load_2d(ybuf + ((size_t)yb * NCH + c) * Y_CHUNK_BYTES, &tm_y, c * BK, t * BN, &y_full[yb]);
        for (int qt = 0; qt < n_qtiles; qt++) {
#pragma unroll
          for (int c = 0; c < NCH; c++) {
            mbar_wait(&q_empty[qs], qphase ^ 1u);
            mbar_expect_tx(&q_full[qs], Q_CHUNK_BYTES);
            tma_load_2d(qbuf + (size_t)qs * Q_CHUNK_BYTES, &tm_q, c * BK, qt * BM, &q_full[qs]);
            if (++qs == QSTAGES) { qs = 0; qphase ^= 1u; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer =====
    if (lane == 0) {
      uint32_t qs = 0, qphase = 0;
      int it = 0, acc_it = 0;
      for (int t = blockIdx.x; t < n_ytiles; t += gridDim.x, it++) {
        const int yb = it % YB;
        const uint32_t yph = (uint32_t)(it / YB) & 1u;
        mbar_wait(&y_full[yb], yph);
        for (int qt = 0; qt < n_qtiles; qt++, acc_it++) {
          const int acc = acc_it & 1;
          const uint32_t aph = (uint32_t)(acc_it >> 1) & 1u;
          mbar_wait(&t_empty[acc], aph ^ 1u);
          tcgen05_fence_after();
          const uint32_t d_tmem = tmem_base + (uint32_t)acc * BN;
#pragma unroll
          for (int c = 0; c < NCH; c++) {
            mbar_wait(&q_full[qs], qphase);
            tcgen05_fence_after();
            const uint32_t a_addr = smem_u32(qbuf + (size_t)qs * Q_CHUNK_BYTES);
            const uint32_t b_addr = smem_u32(ybuf + ((size_t)yb * NCH + c) * Y_CHUNK_BYTES);
#pragma unroll
            for (int k = 0; k < BK / UMMA_K; k++) {
              const uint64_t a_desc = umma_desc_k_sw128(a_addr + k * UMMA_K * 2);
              const uint64_t b_desc = umma_desc_k_sw128(b_addr + k * UMMA_K * 2);
              umma_bf16(d_tmem, a_desc, b_desc, kIdesc, (uint32_t)((c | k) != 0));
            }
            tcgen05_commit(&q_empty[qs]);  // the Q stage is free once these MMAs have read it
            if (++qs == QSTAGES) { qs = 0; qphase ^= 1u; }
          }
          tcgen05_commit(&t_full[acc]);    // accumulator ready for the epilogue
        }
        tcgen05_commit(&y_empty[yb]);      // every MMA that reads this Y tile has been issued before this commit
      }
    }
  } else {
    // ===== epilogue: TMEM -> registers -> sub-block maxima =====
    // Eight warps: a warp may only read the TMEM lane quarter warp % 4, so two warps share each quarter and
    // split the accumulator's 256 columns (4 sub-blocks each).  With one warp per quarter the read-out of an
    // accumulator took as long as the MMAs that fill it and the MMA issuer spun on t_empty (ncu, round 2); two
    // warps per scheduler hide each other's TMEM-load and FMNMX latency.
    constexpr int HSUBS = SUBS_PER_TILE / 2;  // sub-blocks per warp
    const int w4 = warp & 3;                  // TMEM lane quarter this warp may read
    const int half = (warp - 2) >> 2;         // column half
    int acc_it = 0;
    for (int t = blockIdx.x; t < n_ytiles; t += gridDim.x) {
      const int64_t col0 = (int64_t)t * BN + half * (BN / 2);
      const bool ragged = col0 + BN / 2 > n_slots;
      const int bp = LSH ? __ldg(lsh.block_part + t) : 0;
      for (int qt = 0; qt < n_qtiles; qt++, acc_it++) {
        const int acc = acc_it & 1;
        const uint32_t aph = (uint32_t)(acc_it >> 1) & 1u;
        mbar_wait(&t_full[acc], aph);
        tcgen05_fence_after();
        float mx[HSUBS];
        int main_q = 0;
        bool q_cand = true;  // may this thread's query see this tile at all?
        if (LSH) {
          main_q = __ldg(lsh.qmain + (int64_t)qt * BM + w4 * 32 + lane);
          q_cand = bp == -2 || (bp >= 0 && __popc(bp ^ main_q) <= lsh.max_bits);
        }
        if (LSH && !__any_sync(0xffffffffu, q_cand)) {
          // no query of this warp has the tile's partition among its candidates: nothing to read back
#pragma unroll
          for (int j = 0; j < HSUBS; j++) mx[j] = -INFINITY;
        } else {
        // TMEM loads are software-pipelined: the load of sub-block j+1 is in flight while the 32 values of
        // sub-block j are reduced
        uint32_t r[2][32];
        const uint32_t taddr0 = tmem_base + ((uint32_t)(w4 * 32) << 16) + (uint32_t)(acc * BN + half * (BN / 2));
        tmem_ld_32x32b_x32(taddr0, r[0]);
#pragma unroll
        for (int j = 0; j < HSUBS; j++) {
          tmem_ld_wait();
          if (j + 1 < HSUBS) tmem_ld_32x32b_x32(taddr0 + (uint32_t)((j + 1) * SUB), r[(j + 1) & 1]);
          const uint32_t* rr = r[j & 1];
          float m = -INFINITY;
          if (LSH && bp == -2) {
            // mixed tile: row by row (the partition load has the same address for the whole warp)
#pragma unroll
            for (int i = 0; i < 32; i++) {
              const int64_t slot = col0 + j * SUB + i;
              const int part = slot < n_slots ? __ldg(lsh.slot_part + slot) : -1;
              const float v = __uint_as_float(rr[i]);
              if (part >= 0 && __popc(part ^ main_q) <= lsh.max_bits) m = fmaxf(m, v);
            }
          } else if (!ragged) {
#pragma unroll
            for (int i = 0; i < 32; i++) m = fmaxf(m, __uint_as_float(rr[i]));  // fmaxf drops NaN (tombstones)
          } else {
#pragma unroll
            for (int i = 0; i < 32; i++) {
              const float v = __uint_as_float(rr[i]);
              if (col0 + j * SUB + i < n_slots) m = fmaxf(m, v);
            }
          }
          mx[j] = (LSH && !q_cand) ? -INFINITY : m;
        }
        }
        tcgen05_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&t_empty[acc]);  // this warp's half is drained (8 warps arrive)
        const int64_t q = (int64_t)qt * BM + w4 * 32 + lane;
        // bmax is tile-major, [tile][column half][query][4]: one warp's store is 512 contiguous bytes (query-major
        // rows put every lane in a different row: 32 half-written sectors per store)
        *reinterpret_cast<float4*>(bmax + (((int64_t)t * 2 + half) * bmax_stride + q) * 4) = make_float4(mx[0], mx[1], mx[2], mx[3]);
        // coarse level: the maximum of the whole 256-item tile (the threshold step reads this one).  The warp of
        // the upper half hands its maximum over through shared memory (double-buffered per accumulator; the
        // named barrier of the two warps of this lane quarter orders the hand-over and the buffer's reuse).
        const float hm = fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3]));
        if (half == 1) half_max[acc * BM + w4 * 32 + lane] = hm;
        asm volatile("bar.sync %0, 64;" ::"r"(1 + w4) : "memory");
        if (half == 0) tmax[(((int64_t)(t >> 5)) * tmax_stride + q) * 32 + (t & 31)] = fmaxf(hm, half_max[acc * BM + w4 * 32 + lane]);
      }
    }
  }

  tcgen05_fence_before();
  __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem_base) : "memory");
  }
}

// ---------------------------------------------------------------- step 3: thresholds and candidates
constexpr int kThrThreads = 256;
constexpr int kTopPerThread = 4;     // 256 x 4 = 1024 >= largest k
constexpr int kCollectCap = 4096;    // sub-blocks above the first cut, staged in shared memory
constexpr int kTileCap = 2048;       // tiles above the first cut

// k-th largest of n floats in shared memory (n >= k >= 1), by the whole CTA: radix select over order-preserving
// 32-bit keys, four passes of eight bits, one 256-bin histogram per pass.  Returns the value to every thread.
// (The threshold step used a 55-step bitonic sort for the first cut and an n x n rank count for the second.)
__device__ float cta_kth_largest_float(const float* vals, int n, int k, unsigned int* hist /*[256]*/) {
  const int tid = threadIdx.x, T = blockDim.x;
  __shared__ unsigned int s_prefix, s_mask, s_remaining;
  __syncthreads();  // callers' writes to vals[] and earlier users of hist[] / the selection state
  if (tid == 0) { s_prefix = 0u; s_mask = 0u; s_remaining = (unsigned)k; }
  for (int pass = 0; pass < 4; pass++) {
    const int shift = 24 - 8 * pass;
    for (int b = tid; b < 256; b += T) hist[b] = 0u;
    __syncthreads();
    const unsigned prefix = s_prefix, mask = s_mask;
    for (int i = tid; i < n; i += T) {
      const unsigned key = float_key(vals[i]);
      if ((key & mask) == prefix) atomicAdd(&hist[(key >> shift) & 255u], 1u);
    }
    __syncthreads();
    if (tid < 32) {
      unsigned local[8];
      unsigned sum = 0;
#pragma unroll
      for (int q = 0; q < 8; q++) { local[q] = hist[255 - 8 * tid - q]; sum += local[q]; }
      unsigned incl = sum;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const unsigned t = __shfl_up_sync(0xffffffffu, incl, d);
        if (tid >= d) incl += t;
      }
      const unsigned excl_sum = incl - sum;
      const unsigned remaining = s_remaining;
      if (excl_sum < remaining && incl >= remaining) {  // exactly one lane owns the crossing
        unsigned cum = excl_sum;
        int digit = 0;
#pragma unroll
        for (int q = 0; q < 8; q++) {
          if (cum + local[q] >= remaining) { digit = 255 - 8 * tid - q; break; }
          cum += local[q];
        }
        s_prefix = prefix | ((unsigned)digit << shift);
        s_mask = mask | (255u << shift);
        s_remaining = remaining - cum;
      }
    }
    __syncthreads();
  }
  return key_float(s_prefix);
}

// One CTA per query.  cand[q][0..cand_cap) receives candidate sub-block ids, cand_n[q] their number
// (or -1 when too many qualified: the host re-runs that query unbatched).
//
// Two levels keep this step off the 4-bytes-per-sub-block array: the per-TILE maxima (256 items,
// 1/8 of the data) decide which tiles can matter, and only those tiles' 8 sub-block maxima are read.
//   Tc  = k-th largest tile maximum: at least k tiles, hence k items, have a bf16 score >= Tc, so the
//         k-th best exact score is >= Tc - delta and every top-k item has a bf16 score >= Tc - 2 delta;
//         it lives in a tile (and a sub-block) whose maximum is >= Tc - 2 delta.
//   T'  = k-th largest SUB-BLOCK maximum (>= Tc); same argument, tighter: candidates are the sub-blocks
//         with maximum >= T' - 2 delta.  Every sub-block >= T' - 2 delta sits in a tile >= Tc - 2 delta.
__global__ void __launch_bounds__(kThrThreads)
batch_threshold_kernel(const float* __restrict__ bmax, int64_t bmax_stride, const float* __restrict__ tmax,
                       int64_t tmax_stride, int n_tiles, int k_base, const int32_t* __restrict__ excl_offsets,
                       const float* __restrict__ qnorm, int nq_pad,
                       const unsigned int* __restrict__ max_norm_sq_bits, int32_t* __restrict__ cand, int cand_cap,
                       int32_t* __restrict__ cand_n, float* __restrict__ cut_out) {
  __shared__ float s_top[kThrThreads * kTopPerThread];
  __shared__ float s_val[kCollectCap];
  __shared__ int32_t s_id[kCollectCap];
  __shared__ int32_t s_tile[kTileCap];
  __shared__ unsigned int s_count, s_out, s_ntile;
  __shared__ unsigned int s_hist[256];
  const int q = blockIdx.x, tid = threadIdx.x;
  // The maxima include rows the query excludes (exclusions are applied row by row in the refine step),
  // so the rank that guarantees k ALLOWED rows at or above the cut is k + the number of excluded rows:
  // of any k + n_excl sub-blocks, at most n_excl owe their maximum to an excluded row.
  const int k = k_base + (excl_offsets ? excl_offsets[q + 1] - excl_offsets[q] : 0);
  if (k > kThrThreads * kTopPerThread) {  // beyond what pass A can rank: the host answers this query unbatched
    if (tid == 0) cand_n[q] = -1;
    return;
  }
  // rigorous bound on |bf16 score - exact score| for this query (see file header)
  // s_bf - s = bf(q) . (bf(y) - y) + (bf(q) - q) . y, so |s_bf - s| <= ||bf(q)|| R + ||q - bf(q)|| Ymax with
  // R = the largest ||y - bf(y)|| and Ymax the largest ||y|| over the shadow's rows (both measured when the
  // shadow was built) -- about 2.5 x tighter than the worst case 2^-7 ||q|| Ymax; plus 2^-13 ||q|| Ymax for the
  // fp32 accumulation of up to 256 terms (4 x the 2^-23-per-add figure, truncating adds included) and 0.2 % for
  // the fp32 norms themselves.  The refine step's bf16 pre-filter needs ||q|| R + its own fp32 accumulation
  // <= delta, which holds because ||q|| <= ||bf(q)|| + ||q - bf(q)|| and R <= Ymax.
  const float ymax = sqrtf(__uint_as_float(max_norm_sq_bits[0])), rmax = sqrtf(__uint_as_float(max_norm_sq_bits[1]));
  const float delta = 1.002f * (qnorm[nq_pad + q] * ymax + qnorm[2 * nq_pad + q] * (rmax + 1e-6f * ymax)) +
                      1.2207e-4f * qnorm[q] * ymax;

  // pass A (tiles): every thread keeps the 4 largest values of its strided share
  float top[kTopPerThread];
#pragma unroll
  for (int j = 0; j < kTopPerThread; j++) top[j] = -INFINITY;
  // 128-bit loads, four of them in flight per thread (a dependent load per element would leave the
  // pass latency-bound)
  // tmax is [tile / 32][query][32]: float4 i of this query's tile maxima (tiles 4i .. 4i+3)
  auto trow4 = [&](int i) { return reinterpret_cast<const float4*>(tmax + (((int64_t)(i >> 3)) * tmax_stride + q) * 32) + (i & 7); };
  const int n4 = (n_tiles + 3) >> 2;
  auto offer = [&](float v) {
    if (v > top[kTopPerThread - 1]) {
      top[kTopPerThread - 1] = v;
#pragma unroll
      for (int j = kTopPerThread - 1; j > 0; j--) {
        if (top[j] > top[j - 1]) { float t = top[j]; top[j] = top[j - 1]; top[j - 1] = t; }
      }
    }
  };
  for (int i0 = tid; i0 < n4; i0 += 4 * kThrThreads) {
    float4 v[4];
#pragma unroll
    for (int u = 0; u < 4; u++) {
      const int i = i0 + u * kThrThreads;
      v[u] = i < n4 ? __ldg(trow4(i)) : make_float4(-INFINITY, -INFINITY, -INFINITY, -INFINITY);
    }
#pragma unroll
    for (int u = 0; u < 4; u++) {
      const int base = (i0 + u * kThrThreads) * 4;
      if (base + 0 < n_tiles) offer(v[u].x);
      if (base + 1 < n_tiles) offer(v[u].y);
      if (base + 2 < n_tiles) offer(v[u].z);
      if (base + 3 < n_tiles) offer(v[u].w);
    }
  }
#pragma unroll
  for (int j = 0; j < kTopPerThread; j++) s_top[tid * kTopPerThread + j] = top[j];
  if (tid == 0) { s_count = 0u; s_out = 0u; s_ntile = 0u; }
  __syncthreads();
  // lower bound of Tc: the k-th largest of these 1024 values (radix select)
  const float lb = cta_kth_largest_float(s_top, kThrThreads * kTopPerThread, min(k, kThrThreads * kTopPerThread), s_hist);
  // (-inf when fewer than k finite tile maxima exist: then everything is collected)
  const float cut1 = lb - 2.0f * delta;

  // pass B (tiles): the tiles at or above the first cut ...
  auto collect = [&](float v, int i) {
    if (i < n_tiles && v >= cut1 && v > -INFINITY) {
      unsigned pos = atomicAdd(&s_ntile, 1u);
      if (pos < (unsigned)kTileCap) s_tile[pos] = i;
    }
  };
  for (int i0 = tid; i0 < n4; i0 += 4 * kThrThreads) {
    float4 v[4];
#pragma unroll
    for (int u = 0; u < 4; u++) {
      const int i = i0 + u * kThrThreads;
      v[u] = i < n4 ? __ldg(trow4(i)) : make_float4(-INFINITY, -INFINITY, -INFINITY, -INFINITY);
    }
#pragma unroll
    for (int u = 0; u < 4; u++) {
      const int base = (i0 + u * kThrThreads) * 4;
      collect(v[u].x, base + 0);
      collect(v[u].y, base + 1);
      collect(v[u].z, base + 2);
      collect(v[u].w, base + 3);
    }
  }
  __syncthreads();
  const int nt = (int)s_ntile;
  if (nt > kTileCap) {
    if (tid == 0) cand_n[q] = -1;
    return;
  }
  // ... and their sub-blocks at or above it
  for (int e = tid; e < nt * SUBS_PER_TILE; e += kThrThreads) {
    const int tl = s_tile[e / SUBS_PER_TILE], j = e % SUBS_PER_TILE;
    const float v = __ldg(bmax + (((int64_t)tl * 2 + (j >> 2)) * bmax_stride + q) * 4 + (j & 3));
    if (v >= cut1 && v > -INFINITY) {
      unsigned pos = atomicAdd(&s_count, 1u);
      if (pos < (unsigned)kCollectCap) { s_val[pos] = v; s_id[pos] = tl * SUBS_PER_TILE + j; }
    }
  }
  __syncthreads();
  const int n = (int)s_count;
  if (n > kCollectCap) {
    if (tid == 0) cand_n[q] = -1;
    return;
  }
  // exact k-th largest sub-block maximum T' among the collected ones (they include everything >= lb)
  const float t_prime = n >= k ? cta_kth_largest_float(s_val, n, k, s_hist) : -INFINITY;  // n, k are CTA-uniform
  const float cut2 = (n >= k) ? t_prime - 2.0f * delta : -INFINITY;
  int32_t* out = cand + (int64_t)q * cand_cap;
  for (int e = tid; e < n; e += kThrThreads) {
    if (s_val[e] >= cut2) {
      unsigned pos = atomicAdd(&s_out, 1u);
      if (pos < (unsigned)cand_cap) out[pos] = s_id[e];
    }
  }
  __syncthreads();
  if (tid == 0) {
    cand_n[q] = (int)s_out > cand_cap ? -1 : (int)s_out;
    cut_out[q] = cut2;  // the refine step applies the same cut row by row
  }
}

// ---------------------------------------------------------------- step 4: refine
constexpr int kRefineThreads = 256;

__device__ __forceinline__ bool excluded_slot(const int32_t* excl, int n, int32_t slot) {
  int lo = 0, hi = n - 1;
  while (lo <= hi) {
    int mid = (lo + hi) >> 1;
    int32_t v = __ldg(excl + mid);
    if (v == slot) return true;
    if (v < slot) lo = mid + 1; else hi = mid - 1;
  }
  return false;
}

// One CTA per query, one warp per candidate sub-block, two stages:
//   A. score the sub-block's 32 rows against the bf16 shadow of Y (the very rows the tensor cores saw,
//      256 bytes each at 128 features; fp32 query, fp32 accumulate) and keep the rows whose score
//      reaches the query's cut T' - 2 delta.  |q . bf16(y) - q . y| <= ||q|| max ||y - bf(y)|| <= delta, so every
//      row with an exact score >= T' - delta -- every possible top-k member -- passes;
//   B. the survivors (one or two per sub-block) are scored from the fp32 rows, the whole warp on one
//      row, and offered to the shared top-k array.
// Then the exact finishing pass: same selection rule and the same exact arithmetic as scan_topn.cu,
// so results are identical to the single-query path.  Reading 8 KB of bf16 per sub-block instead of
// 16 KB of fp32 nearly halves this step's traffic.
template <int NCH>
__global__ void __launch_bounds__(kRefineThreads)
batch_refine_kernel(const float* __restrict__ y, const __nv_bfloat16* __restrict__ ybf, int kpad,
                    const int32_t* __restrict__ slot_row, int64_t n_slots, int fpad,
                    int features, const float* __restrict__ queries, int kind, int k, int how_many,
                    const int32_t* __restrict__ cand, int cand_cap, const int32_t* __restrict__ cand_n,
                    const float* __restrict__ cutv,
                    const int32_t* __restrict__ excl_offsets, const int32_t* __restrict__ excl_slots,
                    const BatchLsh lsh,  // slot_part == nullptr: no LSH filtering
                    int32_t* __restrict__ out_rows, double* __restrict__ out_scores, int32_t* __restrict__ out_n) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  // layout: q[kpad] f32 | arr[k] u64 | exact[k] f64 | rows[k] i32
  float* s_q = reinterpret_cast<float*>(smem_raw);
  unsigned long long* arr = reinterpret_cast<unsigned long long*>(smem_raw + (((size_t)kpad * 4 + 15) & ~(size_t)15));
  double* s_exact = reinterpret_cast<double*>(arr + k);
  int32_t* s_rows = reinterpret_cast<int32_t*>(s_exact + k);
  __shared__ unsigned int s_thr_key;
  const int qi = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int nc = cand_n[qi];
  if (nc < 0) {  // overflow in the threshold step: the host answers this query unbatched
    if (tid == 0) out_n[qi] = -1;
    return;
  }
  for (int j = tid; j < kpad; j += kRefineThreads) s_q[j] = j < features ? queries[(size_t)qi * features + j] : 0.f;
  for (int i = tid; i < k; i += kRefineThreads) arr[i] = 0ull;
  if (tid == 0) s_thr_key = 0u;
  __syncthreads();
  const int32_t* excl = excl_offsets ? excl_slots + excl_offsets[qi] : nullptr;
  const int n_excl = excl_offsets ? excl_offsets[qi + 1] - excl_offsets[qi] : 0;
  const int32_t* my_cand = cand + (int64_t)qi * cand_cap;
  const float cut = cutv[qi];
  const int main_q = lsh.slot_part ? __ldg(lsh.qmain + qi) : 0;

  const int gl = lane & 7, grp = lane >> 3;
  const int f4 = fpad >> 2;
  const float4* q4 = reinterpret_cast<const float4*>(s_q);
  // offer one scored row (all lanes hold the same score/slot) to the shared top-k array
  auto offer_row = [&](float score, int32_t slot) {
    const uint32_t ky = float_key(score);
    if (!(score == score) || ky <= *(volatile unsigned int*)&s_thr_key) return;
    if (n_excl > 0 && excluded_slot(excl, n_excl, slot)) return;
    // replace-the-minimum with a 64-bit CAS, whole warp searches (as in scan_topn.cu)
    const unsigned long long x = ((unsigned long long)ky << 32) | (unsigned long long)(uint32_t)slot;
    for (;;) {
      unsigned long long best = ~0ull;
      int best_idx = 0;
      for (int i = lane; i < k; i += 32) {
        unsigned long long v = *(const volatile unsigned long long*)(arr + i);
        if (v < best) { best = v; best_idx = i; }
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        unsigned long long ov = __shfl_xor_sync(0xffffffffu, best, o);
        int oi = __shfl_xor_sync(0xffffffffu, best_idx, o);
        if (ov < best || (ov == best && oi < best_idx)) { best = ov; best_idx = oi; }
      }
      if (ky <= (uint32_t)(best >> 32)) {
        if (lane == 0) atomicMax(&s_thr_key, (uint32_t)(best >> 32));
        break;
      }
      unsigned long long old = 0ull;
      if (lane == 0) old = atomicCAS(arr + best_idx, best, x);
      old = __shfl_sync(0xffffffffu, old, 0);
      if (old == best) break;
    }
  };
  const int u4_per_row = kpad >> 3;  // 8 bf16 per 128-bit load; 8 lanes share a row: NCH loads per lane
  for (int c = warp; c < nc; c += kRefineThreads / 32) {
    const int64_t sub0 = (int64_t)my_cand[c] * SUB;
    // ---- stage A: bf16 scores of the 32 rows, 4 passes (16 rows) of loads in flight at a time
    unsigned pass_mask = 0u;  // bit (pass * 4 + grp) = row sub0 + pass*4 + grp reaches the cut
#pragma unroll
    for (int half = 0; half < 2; half++) {
      uint4 a[4][NCH];
#pragma unroll
      for (int h = 0; h < 4; h++) {
        const int64_t slot = sub0 + (half * 4 + h) * 4 + grp;
        const uint4* r = reinterpret_cast<const uint4*>(ybf + slot * kpad);
#pragma unroll
        for (int u = 0; u < NCH; u++)
          a[h][u] = slot < n_slots ? __ldg(r + gl + u * 8) : make_uint4(0u, 0u, 0u, 0u);
      }
#pragma unroll
      for (int h = 0; h < 4; h++) {
        float acc = 0.f;
#pragma unroll
        for (int u = 0; u < NCH; u++) {
          const int j = gl + u * 8;  // 128-bit unit -> q elements 8j .. 8j+7
          const float4 b0 = q4[2 * j], b1 = q4[2 * j + 1];
          const uint4 v = a[h][u];
          acc = fmaf(__uint_as_float(v.x << 16), b0.x, acc); acc = fmaf(__uint_as_float(v.x & 0xffff0000u), b0.y, acc);
          acc = fmaf(__uint_as_float(v.y << 16), b0.z, acc); acc = fmaf(__uint_as_float(v.y & 0xffff0000u), b0.w, acc);
          acc = fmaf(__uint_as_float(v.z << 16), b1.x, acc); acc = fmaf(__uint_as_float(v.z & 0xffff0000u), b1.y, acc);
          acc = fmaf(__uint_as_float(v.w << 16), b1.z, acc); acc = fmaf(__uint_as_float(v.w & 0xffff0000u), b1.w, acc);
        }
#pragma unroll
        for (int o = 4; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
        const int64_t slot = sub0 + (half * 4 + h) * 4 + grp;
        bool keep = gl == 0 && slot < n_slots && acc >= cut;  // NaN (tombstone) compares false
        if (keep && lsh.slot_part) {  // rows outside the query's candidate partitions do not compete
          const int part = __ldg(lsh.slot_part + slot);
          keep = part >= 0 && __popc(part ^ main_q) <= lsh.max_bits;
        }
        const unsigned ball = __ballot_sync(0xffffffffu, keep);     // bits 0, 8, 16, 24 <-> grp 0..3
        pass_mask |= (((ball >> 0) & 1u) | (((ball >> 8) & 1u) << 1) | (((ball >> 16) & 1u) << 2) |
                      (((ball >> 24) & 1u) << 3))
                     << ((half * 4 + h) * 4);
      }
    }
    (void)u4_per_row;
    // ---- stage B: fp32 score of each surviving row, the whole warp on one row
    while (pass_mask != 0u) {
      const int r = __ffs(pass_mask) - 1;
      pass_mask &= pass_mask - 1;
      const int64_t slot = sub0 + r;
      const float4* r4 = reinterpret_cast<const float4*>(y + slot * fpad);
      float acc = 0.f, nrm = 0.f;
      for (int j = lane; j < f4; j += 32) {
        const float4 v = __ldg(r4 + j);
        const float4 b = q4[j];
        acc = fmaf(v.x, b.x, acc); acc = fmaf(v.y, b.y, acc); acc = fmaf(v.z, b.z, acc); acc = fmaf(v.w, b.w, acc);
        if (kind == ORYX_COSINE) {
          nrm = fmaf(v.x, v.x, nrm); nrm = fmaf(v.y, v.y, nrm); nrm = fmaf(v.z, v.z, nrm); nrm = fmaf(v.w, v.w, nrm);
        }
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        acc += __shfl_xor_sync(0xffffffffu, acc, o);
        if (kind == ORYX_COSINE) nrm += __shfl_xor_sync(0xffffffffu, nrm, o);
      }
      offer_row((kind == ORYX_COSINE) ? acc * rsqrtf(nrm) : acc, (int32_t)slot);
    }
  }
  __syncthreads();

  // exact re-score (one thread per kept row here: k rows, short chains run side by side)
  __shared__ int s_nsel;
  if (tid == 0) s_nsel = 0;
  __syncthreads();
  for (int i = tid; i < k; i += kRefineThreads) {
    const unsigned long long v = arr[i];
    if ((v >> 32) == 0ull) continue;
    const int32_t slot = (int32_t)(uint32_t)v;
    const float* r = y + (int64_t)slot * fpad;
    double dot = 0.0, tot = 0.0;
    for (int j = 0; j < features; j++) {
      const float yv = r[j];
      dot = exact_step(dot, s_q[j], yv);
      if (kind == ORYX_COSINE) tot = exact_step(tot, yv, yv);
    }
    const int pos = atomicAdd(&s_nsel, 1);
    s_exact[pos] = (kind == ORYX_COSINE) ? __ddiv_rn(dot, __dsqrt_rn(tot)) : dot;
    s_rows[pos] = __ldg(slot_row + slot);
  }
  __syncthreads();
  const int nsel = s_nsel;
  const int n_out = min(nsel, how_many);
  for (int i = tid; i < nsel; i += kRefineThreads) {
    const double si = s_exact[i];
    const int32_t ri = s_rows[i];
    const bool nan_i = si != si;
    int rank = 0;
    for (int j = 0; j < nsel; j++) {
      const double sj = s_exact[j];
      const bool nan_j = sj != sj;
      const int32_t rj = s_rows[j];
      bool better;
      if (nan_j) better = nan_i && rj < ri;
      else if (nan_i) better = true;
      else better = (sj > si) || (sj == si && rj < ri);
      rank += better ? 1 : 0;
    }
    if (rank < n_out) {
      out_rows[(size_t)qi * how_many + rank] = ri;
      out_scores[(size_t)qi * how_many + rank] = si;
    }
  }
  if (tid == 0) out_n[qi] = n_out;
}

// ---------------------------------------------------------------- host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) != cudaSuccess ||
        qres != cudaDriverEntryPointSuccess) {
      cudaGetLastError();
      p = nullptr;
    }
    return reinterpret_cast<EncodeTiledFn>(p);
  }();
  return fn;
}

// bf16 row-major [rows x kpad] matrix, box = {64 elements (128 B), box_rows}, 128-byte swizzle.
int32_t make_map(CUtensorMap* map, void* base, uint64_t rows, uint64_t kpad, uint32_t box_rows) {
  EncodeTiledFn fn = encode_tiled_fn();
  if (!fn) {
    set_error("cuTensorMapEncodeTiled is not available from this driver");
    return ORYX_E_CUDA;
  }
  cuuint64_t dims[2] = {kpad, rows};
  cuuint64_t strides[1] = {kpad * 2};
  cuuint32_t box[2] = {(cuuint32_t)BK, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, base, dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
    return ORYX_E_CUDA;
  }
  return ORYX_OK;
}

template <int NCH, bool LSH>
int32_t launch_score(const CUtensorMap& tq, const CUtensorMap& ty, int n_qtiles, int n_ytiles, int64_t n_slots,
                     float* bmax, int64_t bmax_stride, float* tmax, int64_t tmax_stride, const BatchLsh& lsh, int num_sms,
                     cudaStream_t st) {
  constexpr int YB = (NCH <= 2) ? 2 : 1;
  const size_t smem = ScoreSmem::bytes(NCH, YB);
  ORYX_TRY(ensure_max_dynamic_smem(reinterpret_cast<const void*>(batch_score_kernel<NCH, LSH>), smem));
  const int grid = std::min(num_sms, n_ytiles);
  batch_score_kernel<NCH, LSH><<<grid, kScoreThreads, smem, st>>>(tq, ty, n_qtiles, n_ytiles, n_slots, bmax, bmax_stride,
                                                                  tmax, tmax_stride, lsh);
  ORYX_CUDA(cudaGetLastError());
  return ORYX_OK;
}

// Slot `i` of the model's grow-only batch scratch, at least `bytes` large.
int32_t scratch(BatchScratch& bs, int i, size_t bytes, void** out, bool* grew = nullptr) {
  if (grew) *grew = false;
  if (bs.bytes[i] < bytes) {
    cudaFree(bs.buf[i]);
    bs.buf[i] = nullptr;
    bs.bytes[i] = 0;
    if (cudaMalloc(&bs.buf[i], bytes) != cudaSuccess) {
      cudaGetLastError();
      set_error("device allocation of %zu bytes failed (batched top-N scratch)", bytes);
      return ORYX_E_NOMEM;
    }
    bs.bytes[i] = bytes;
    if (grew) *grew = true;
  }
  *out = bs.buf[i];
  return ORYX_OK;
}

}  // namespace

// Called by the store (under the model's exclusive lock) after the rows in d_slots[0..n) of Y changed.
// Returns true when the shadow was current before the mutation and has been brought up to date.
bool batch_shadow_rows_changed(Model* m, const int32_t* d_slots, int64_t n, uint64_t version_before, cudaStream_t st) {
  BatchScratch& bs = m->batch;
  Store* s = &m->store[ORYX_Y];
  // the slot array may have grown inside its capacity: the new slots are in d_slots like any other change
  if (bs.buf[0] == nullptr || bs.buf[4] == nullptr || bs.y_version != version_before || bs.y_rows > s->n_slots ||
      s->n_slots > bs.y_cap_rows || bs.y_normalized < 0)
    return false;
  bs.y_rows = s->n_slots;
  if (n > 0) {
    const int kpad = round_up(s->features, BK);
    convert_rows_kernel<<<(unsigned)((n * 32 + 255) / 256), 256, 0, st>>>(
        s->d_vec, d_slots, n, bs.y_rows, s->fpad, s->features, kpad, bs.y_normalized, (__nv_bfloat16*)bs.buf[0],
        (unsigned int*)bs.buf[4]);
    m->launches++;
    if (cudaGetLastError() != cudaSuccess) return false;
  }
  return true;
}

thread_local bool tl_batch_small_pages_ok = false;

bool batch_topn_possible(Model* m) {
  Store* s = &m->store[ORYX_Y];
  return round_up(s->features, BK) <= 256 && s->n_slots >= 4096 && encode_tiled_fn() != nullptr;
}

bool batch_topn_supported(Model* m, int nq) {
  Store* s = &m->store[ORYX_Y];
  const int kpad = round_up(s->features, BK);
  // Below 32 queries the path only pays off while the bf16 shadow of Y is current (it then reads half the
  // bytes of an fp32 scan even for a single query); building the shadow for a handful of queries does not.
  const bool shadow_current = m->batch.y_version == s->version && m->batch.y_rows == s->n_slots;
  return kpad <= 256 && (nq >= 32 || (nq >= 1 && (shadow_current || tl_batch_small_pages_ok))) && s->n_slots >= 4096 &&
         encode_tiled_fn() != nullptr;
}

static int32_t batch_topn_run_pass(Model* m, int kind, const float* queries, int nq, int how_many, const int32_t* excl_offsets,
                                   const int32_t* excl_slots, int32_t* out_rows, double* out_scores, int32_t* out_n,
                                   cudaStream_t st, BatchTimings* timings, const int32_t* query_slots, int query_store);

// queries: HOST [nq x features].  Outputs: HOST arrays.  out_n[i] == -1 marks a query the caller must
// answer unbatched (candidate overflow; pathological score distributions only).
//
// The sub-block maxima take nq x (N/32 + N/256) x 4 bytes of scratch (5.8 GB for 4096 queries over 10 M items).
// That grows with nq x N, so a page is answered in passes of as many queries as fit a scratch budget (8 GB, or
// ORYX_BATCH_SCRATCH_MB): 4096 queries over 100 M items run as eight passes of 512 instead of asking for 57 GB.
int32_t batch_topn_run(Model* m, int kind, const float* queries, int nq, int how_many, const int32_t* excl_offsets,
                       const int32_t* excl_slots, int32_t* out_rows, double* out_scores, int32_t* out_n,
                       cudaStream_t st, BatchTimings* timings, const int32_t* query_slots, int query_store) {
  Store* s = &m->store[ORYX_Y];
  const char* budget_env = std::getenv("ORYX_BATCH_SCRATCH_MB");  // read per call: tests lower it to force passes
  const size_t budget = (size_t)std::max<long long>(budget_env ? std::atoll(budget_env) : 8192, 1) << 20;
  const int64_t n_tiles = (s->n_slots + BN - 1) / BN;
  const size_t per_query = (size_t)(n_tiles * SUBS_PER_TILE + (n_tiles + 3) / 4 * 4) * 4;
  int pass = (int)std::min<size_t>((size_t)nq, std::max<size_t>(BM, budget / std::max<size_t>(per_query, 1) / BM * BM));
  if (pass >= nq)
    return batch_topn_run_pass(m, kind, queries, nq, how_many, excl_offsets, excl_slots, out_rows, out_scores, out_n, st,
                               timings, query_slots, query_store);
  const int F = s->features;
  BatchTimings total{};
  std::vector<int32_t> eo;
  for (int off = 0; off < nq; off += pass) {
    const int c = std::min(pass, nq - off);
    const int32_t* eo_p = nullptr;
    const int32_t* es_p = nullptr;
    if (excl_offsets) {  // offsets of this pass, rebased to its first exclusion
      eo.assign((size_t)c + 1, 0);
      for (int i = 0; i <= c; i++) eo[(size_t)i] = excl_offsets[off + i] - excl_offsets[off];
      eo_p = eo.data();
      es_p = excl_slots + excl_offsets[off];
    }
    BatchTimings t{};
    ORYX_TRY(batch_topn_run_pass(m, kind, queries ? queries + (size_t)off * F : nullptr, c, how_many, eo_p, es_p,
                                 out_rows + (size_t)off * how_many, out_scores + (size_t)off * how_many, out_n + off, st,
                                 timings ? &t : nullptr, query_slots ? query_slots + off : nullptr, query_store));
    total.convert_ms += t.convert_ms; total.score_ms += t.score_ms; total.threshold_ms += t.threshold_ms;
    total.refine_ms += t.refine_ms; total.flops += t.flops; total.shadow_rebuilds += t.shadow_rebuilds;
    total.cand_mean += t.cand_mean * c / nq; total.cand_max = std::max(total.cand_max, t.cand_max);
  }
  if (timings) {
    total.runs = timings->runs + 1;
    total.shadow_rebuilds += timings->shadow_rebuilds;
    *timings = total;
  }
  return ORYX_OK;
}

static int32_t batch_topn_run_pass(Model* m, int kind, const float* queries, int nq, int how_many, const int32_t* excl_offsets,
                                   const int32_t* excl_slots, int32_t* out_rows, double* out_scores, int32_t* out_n,
                                   cudaStream_t st, BatchTimings* timings, const int32_t* query_slots, int query_store) {
  Store* s = &m->store[ORYX_Y];
  const int F = s->features, fpad = s->fpad;
  const int kpad = round_up(F, BK), nch = kpad / BK;
  const int k = how_many + kTopnMargin;
  const int nq_pad = round_up(nq, BM);
  const int n_qtiles = nq_pad / BM;
  const int64_t n = s->n_slots;
  const int n_ytiles = (int)((n + BN - 1) / BN);
  const int n_sub = n_ytiles * SUBS_PER_TILE;
  const int64_t bmax_stride = nq_pad;  // tile-major: [tile][column half][nq_pad][4]
  const int64_t tmax_stride = nq_pad;  // [tile / 32][nq_pad][32]
  const int64_t tmax_floats = (int64_t)((n_ytiles + 31) / 32) * 32 * nq_pad;
  const int cand_cap = 4096;

  BatchScratch& bs = m->batch;
  std::lock_guard<std::mutex> batch_lock(bs.mu);
  struct P { void* p = nullptr; } d_ybf, d_qbf, d_q, d_qnorm, d_maxn, d_bmax, d_tmax, d_cut, d_cand, d_candn, d_orows, d_oscores, d_on, d_eo, d_es;
  bool y_grew = false;
  // room for the store's whole slot capacity, so that the shadow can follow the store as it fills up
  const int64_t shadow_rows_cap = std::max<int64_t>(n, s->cap_slots);
  ORYX_TRY(scratch(bs, 0, (size_t)shadow_rows_cap * kpad * 2, &d_ybf.p, &y_grew));
  bs.y_cap_rows = (int64_t)(bs.bytes[0] / ((size_t)kpad * 2));
  ORYX_TRY(scratch(bs, 1, (size_t)nq_pad * kpad * 2, &d_qbf.p));
  ORYX_TRY(scratch(bs, 2, (size_t)nq * F * 4, &d_q.p));
  ORYX_TRY(scratch(bs, 3, (size_t)nq_pad * 12, &d_qnorm.p));
  ORYX_TRY(scratch(bs, 4, 8, &d_maxn.p));
  ORYX_TRY(scratch(bs, 5, ((size_t)nq_pad * n_sub + (size_t)tmax_floats) * 4, &d_bmax.p));
  d_tmax.p = reinterpret_cast<float*>(d_bmax.p) + (size_t)nq_pad * n_sub;
  ORYX_TRY(scratch(bs, 6, (size_t)nq * cand_cap * 4, &d_cand.p));
  ORYX_TRY(scratch(bs, 7, (size_t)nq * 8, &d_candn.p));
  d_cut.p = reinterpret_cast<int32_t*>(d_candn.p) + nq;
  ORYX_TRY(scratch(bs, 8, (size_t)nq * how_many * 4, &d_orows.p));
  ORYX_TRY(scratch(bs, 9, (size_t)nq * how_many * 8, &d_oscores.p));
  ORYX_TRY(scratch(bs, 10, (size_t)nq * 4, &d_on.p));
  const int n_excl_total = excl_offsets ? excl_offsets[nq] : 0;
  if (excl_offsets) {
    ORYX_TRY(scratch(bs, 11, (size_t)(nq + 1) * 4 + (size_t)std::max(n_excl_total, 1) * 4, &d_eo.p));
    d_es.p = reinterpret_cast<unsigned char*>(d_eo.p) + (size_t)(nq + 1) * 4;
    ORYX_CUDA(cudaMemcpyAsync(d_eo.p, excl_offsets, (size_t)(nq + 1) * 4, cudaMemcpyHostToDevice, st));
    if (n_excl_total > 0)
      ORYX_CUDA(cudaMemcpyAsync(d_es.p, excl_slots, (size_t)n_excl_total * 4, cudaMemcpyHostToDevice, st));
  }
  // the bf16 shadow of Y (and the norm bound that goes with it) is reused while the store is unchanged
  const int normalized = kind == ORYX_COSINE ? 1 : 0;
  const bool y_fresh = !y_grew && bs.y_version == s->version && bs.y_normalized == normalized && bs.y_rows == n;
  cudaEvent_t* ev = bs.ev;  // under bs.mu
  for (int i = 0; i < 5; i++)
    if (ev[i] == nullptr) ORYX_CUDA(cudaEventCreate(&ev[i]));

  if (query_slots) {
    // the queries are rows of a device-resident store: gather them, nothing but slot numbers is uploaded
    void* d_qs = nullptr;
    ORYX_TRY(scratch(bs, 12, (size_t)nq * 4, &d_qs));
    ORYX_CUDA(cudaMemcpyAsync(d_qs, query_slots, (size_t)nq * 4, cudaMemcpyHostToDevice, st));
    const Store* qs = &m->store[query_store];
    gather_queries_kernel<<<(unsigned)(((int64_t)nq * 32 + 255) / 256), 256, 0, st>>>(
        qs->d_vec, qs->fpad, F, (const int32_t*)d_qs, nq, (float*)d_q.p);
    ORYX_CUDA(cudaGetLastError());
  } else {
    ORYX_CUDA(cudaMemcpyAsync(d_q.p, queries, (size_t)nq * F * 4, cudaMemcpyHostToDevice, st));
  }
  ORYX_CUDA(cudaEventRecord(ev[0], st));
  if (!y_fresh) {
    ORYX_CUDA(cudaMemsetAsync(d_maxn.p, 0, 8, st));
    convert_y_kernel<<<m->num_sms * 8, 256, 0, st>>>(s->d_vec, n, fpad, F, kpad, normalized,
                                                    (__nv_bfloat16*)d_ybf.p, (unsigned int*)d_maxn.p);
    bs.y_version = s->version;
    bs.y_normalized = normalized;
    bs.y_rows = n;
    if (timings) timings->shadow_rebuilds += 1;
  }
  convert_q_kernel<<<(nq_pad + 7) / 8, 256, 0, st>>>((const float*)d_q.p, nq, nq_pad, F, kpad,
                                                    (__nv_bfloat16*)d_qbf.p, (float*)d_qnorm.p);
  ORYX_CUDA(cudaGetLastError());
  ORYX_CUDA(cudaEventRecord(ev[1], st));

  CUtensorMap tq, ty;
  ORYX_TRY(make_map(&tq, d_qbf.p, (uint64_t)nq_pad, (uint64_t)kpad, BM));
  ORYX_TRY(make_map(&ty, d_ybf.p, (uint64_t)n, (uint64_t)kpad, BN));
  // LSH filtering on: every query's partition index, by the same kernel that hashes rows at ingest
  const bool lsh_on = s->hashed && m->num_hashes != m->max_bits;
  BatchLsh lsh{};
  if (lsh_on) {
    void* d_qmain = nullptr;
    ORYX_TRY(scratch(bs, 13, (size_t)nq_pad * 4, &d_qmain));
    ORYX_CUDA(cudaMemsetAsync(d_qmain, 0, (size_t)nq_pad * 4, st));
    ORYX_TRY(lsh_hash_rows_device(m, (const float*)d_q.p, nq, F, (int32_t*)d_qmain, st));
    lsh.block_part = s->d_block_part;
    lsh.slot_part = s->d_slot_part;
    lsh.qmain = (const int32_t*)d_qmain;
    lsh.max_bits = m->max_bits;
  }
  int32_t rc;
#define ORYX_SCORE(NCH_)                                                                                              \
  (lsh_on ? launch_score<NCH_, true>(tq, ty, n_qtiles, n_ytiles, n, (float*)d_bmax.p, bmax_stride, (float*)d_tmax.p, \
                                     tmax_stride, lsh, m->num_sms, st)                                                \
          : launch_score<NCH_, false>(tq, ty, n_qtiles, n_ytiles, n, (float*)d_bmax.p, bmax_stride, (float*)d_tmax.p, \
                                      tmax_stride, lsh, m->num_sms, st))
  switch (nch) {
    case 1: rc = ORYX_SCORE(1); break;
    case 2: rc = ORYX_SCORE(2); break;
    case 3: rc = ORYX_SCORE(3); break;
    default: rc = ORYX_SCORE(4); break;
  }
#undef ORYX_SCORE
  ORYX_TRY(rc);
  ORYX_CUDA(cudaEventRecord(ev[2], st));

  batch_threshold_kernel<<<nq, kThrThreads, 0, st>>>((const float*)d_bmax.p, bmax_stride, (const float*)d_tmax.p,
                                                    tmax_stride, n_ytiles, k, excl_offsets ? (const int32_t*)d_eo.p : nullptr,
                                                    (const float*)d_qnorm.p, nq_pad,
                                                    (const unsigned int*)d_maxn.p, (int32_t*)d_cand.p, cand_cap,
                                                    (int32_t*)d_candn.p, (float*)d_cut.p);
  ORYX_CUDA(cudaGetLastError());
  ORYX_CUDA(cudaEventRecord(ev[3], st));

  const size_t rsmem = (((size_t)kpad * 4 + 15) & ~(size_t)15) + (size_t)k * 20 + 16;
  auto launch_refine = [&](auto kern) -> int32_t {
    ORYX_TRY(ensure_max_dynamic_smem(reinterpret_cast<const void*>(kern), 64 * 1024));
    kern<<<nq, kRefineThreads, rsmem, st>>>(
        s->d_vec, (const __nv_bfloat16*)d_ybf.p, kpad, s->d_slot_row, n, fpad, F, (const float*)d_q.p, kind, k,
        how_many, (const int32_t*)d_cand.p, cand_cap, (const int32_t*)d_candn.p, (const float*)d_cut.p,
        excl_offsets ? (const int32_t*)d_eo.p : nullptr, excl_offsets ? (const int32_t*)d_es.p : nullptr, lsh,
        (int32_t*)d_orows.p, (double*)d_oscores.p, (int32_t*)d_on.p);
    return ORYX_OK;
  };
  int32_t rrc;
  switch (nch) {  // 128-bit loads per lane per bf16 row (8 lanes share a row)
    case 1: rrc = launch_refine(batch_refine_kernel<1>); break;
    case 2: rrc = launch_refine(batch_refine_kernel<2>); break;
    case 3: rrc = launch_refine(batch_refine_kernel<3>); break;
    default: rrc = launch_refine(batch_refine_kernel<4>); break;
  }
  ORYX_TRY(rrc);
  ORYX_CUDA(cudaGetLastError());
  ORYX_CUDA(cudaEventRecord(ev[4], st));
  ORYX_CUDA(cudaMemcpyAsync(out_rows, d_orows.p, (size_t)nq * how_many * 4, cudaMemcpyDeviceToHost, st));
  ORYX_CUDA(cudaMemcpyAsync(out_scores, d_oscores.p, (size_t)nq * how_many * 8, cudaMemcpyDeviceToHost, st));
  ORYX_CUDA(cudaMemcpyAsync(out_n, d_on.p, (size_t)nq * 4, cudaMemcpyDeviceToHost, st));
  ORYX_CUDA(cudaStreamSynchronize(st));
  m->launches += 5;
  if (timings) {
    timings->runs += 1;
    float ms;
    cudaEventElapsedTime(&ms, ev[0], ev[1]); timings->convert_ms = ms;
    cudaEventElapsedTime(&ms, ev[1], ev[2]); timings->score_ms = ms;
    cudaEventElapsedTime(&ms, ev[2], ev[3]); timings->threshold_ms = ms;
    cudaEventElapsedTime(&ms, ev[3], ev[4]); timings->refine_ms = ms;
    timings->flops = 2.0 * (double)nq_pad * (double)n_ytiles * BN * (double)kpad;
    std::vector<int32_t> cn((size_t)nq);
    if (cudaMemcpy(cn.data(), d_candn.p, (size_t)nq * 4, cudaMemcpyDeviceToHost) == cudaSuccess) {
      double sum = 0, mx = 0;
      for (int32_t v : cn) { sum += std::max(v, 0); mx = std::max(mx, (double)v); }
      timings->cand_mean = sum / std::max(nq, 1);
      timings->cand_max = mx;
    } else {
      cudaGetLastError();
    }
  }
  return ORYX_OK;
}

}  // namespace oryx
