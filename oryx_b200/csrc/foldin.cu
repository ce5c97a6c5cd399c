// foldin.cu -- batched speed-layer fold-in (kernel K8 of SURVEY.md section 2.3) and the Gram entry
// point.
//
// Replaces n independent calls of ALSUtils.computeUpdatedXu (reference app/oryx-app-common/.../als/
// ALSUtils.java:74-106) as issued by ALSSpeedModelManager.buildUpdates (app/oryx-app/.../speed/als/
// ALSSpeedModelManager.java:182-204) and by EstimateForAnonymous.buildTemporaryUserVector
// (app/oryx-app-serving/.../als/EstimateForAnonymous.java:73-87):
//   Qui    = Xu == null ? 0 : dot(Xu, Yi)                    (binary32 products, binary64 sum)
//   target = computeTargetQui(implicit, value, Xu == null ? 0.5 : Qui)    (ALSUtils.java:37-59)
//   dQui   = target - Qui;  b[i] = (float)((double)Yi[i] * dQui)          (float *= double)
//   dXu    = solveFToF(b) = (float)(G^-1 b)                  (binary64 inside, Solver.java:33-44)
//   result = Xu == null ? dXu : Xu + dXu                     (binary32 add)
// The per-interaction solve is done as one binary64 GEMM against the inverse that solver.cu
// derives from the same QR factors; everything else is the reference's operation sequence.
#include <algorithm>
#include <chrono>
#include <mutex>
#include <thread>

#include "common.cuh"

namespace oryx {

struct Solver;
const double* solver_inverse(const Solver* s);
int solver_n(const Solver* s);
int solver_device(const Solver* s);
cudaStream_t solver_stream(const Solver* s);
std::mutex& solver_mutex(Solver* s);
FoldScratch* solver_fold_scratch(Solver* s);

namespace {

constexpr int kRowsPerCta = 128;  // interactions per CTA

// One warp per interaction: copy the row living in `slots[i]` of a store (or zeros when the slot is
// negative) into a packed [n x f] buffer.
__global__ void gather_rows_kernel(const float* __restrict__ vec, int fpad, int f, const int32_t* __restrict__ slots,
                                   int64_t n, float* __restrict__ out, uint8_t* __restrict__ has) {
  const int64_t w = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (w >= n) return;
  const int32_t slot = slots[w];
  float* dst = out + w * f;
  if (slot >= 0) {
    const float* src = vec + (int64_t)slot * fpad;
    for (int j = lane; j < f; j += 32) dst[j] = src[j];
  } else {
    for (int j = lane; j < f; j += 32) dst[j] = 0.f;
  }
  if (lane == 0 && has) has[w] = slot >= 0 ? 1 : 0;
}

// The same over a store that is sharded across the GPUs of one process: interaction i takes its row
// from slot slots[i] of the shard owners[i]; bases[] are the shards' vector arrays, peers' ones read
// over NVLink (peer access enabled by the caller).
struct ShardBases {
  const float* base[ORYX_MAX_WORLD];
};
__global__ void gather_rows_sharded_kernel(const __grid_constant__ ShardBases shards, int fpad, int f,
                                           const int32_t* __restrict__ owners, const int32_t* __restrict__ slots,
                                           int64_t n, float* __restrict__ out, uint8_t* __restrict__ has) {
  const int64_t w = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (w >= n) return;
  const int32_t slot = slots[w];
  float* dst = out + w * f;
  if (slot >= 0) {
    const float* src = shards.base[owners[w]] + (int64_t)slot * fpad;
    for (int j = lane; j < f; j += 32) dst[j] = src[j];
  } else {
    for (int j = lane; j < f; j += 32) dst[j] = 0.f;
  }
  if (lane == 0 && has) has[w] = slot >= 0 ? 1 : 0;
}

// The same over shards that live in other PROCESSES (peer_stores below): interaction i names a GLOBAL row;
// its owner is row / stride, the slot comes from the owner's device row -> slot table and the row from the
// owner's vector array -- both mapped through CUDA IPC, read with plain loads over NVLink.  One warp per
// interaction: lane 0 looks the slot up, every lane copies its share of the row once.
struct PeerView {
  const float* vec[ORYX_MAX_WORLD];
  const int32_t* row_slot[ORYX_MAX_WORLD];
  long long row_slot_cap[ORYX_MAX_WORLD];
  long long cap_slots[ORYX_MAX_WORLD];
  int world;
};
__global__ void gather_rows_peers_kernel(const __grid_constant__ PeerView pv, long long stride, int fpad, int f,
                                         const long long* __restrict__ rows, int64_t n, float* __restrict__ out,
                                         uint8_t* __restrict__ has) {
  const int64_t w = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (w >= n) return;
  const long long row = rows[w];
  int owner = 0;
  int32_t slot = -1;
  if (lane == 0 && row >= 0) {
    const long long r = row / stride;
    if (r < pv.world && pv.row_slot[r] != nullptr) {
      const long long local = row - r * stride;
      if (local < pv.row_slot_cap[r]) {
        const int32_t sl = pv.row_slot[r][local];
        if (sl >= 0 && sl < pv.cap_slots[r]) { slot = sl; owner = (int)r; }
      }
    }
  }
  slot = __shfl_sync(0xffffffffu, slot, 0);
  owner = __shfl_sync(0xffffffffu, owner, 0);
  float* dst = out + w * f;
  if (slot >= 0) {
    const float* src = pv.vec[owner] + (int64_t)slot * fpad;
    for (int j = lane; j < f; j += 32) dst[j] = src[j];
  } else {
    for (int j = lane; j < f; j += 32) dst[j] = 0.f;
  }
  if (lane == 0 && has) has[w] = slot >= 0 ? 1 : 0;
}

// ALSUtils.java:37-59
__device__ __forceinline__ double target_qui(bool implicit, double value, double current) {
  if (implicit) {
    if (value > 0.0 && current < 1.0) {
      double diff = 1.0 - fmax(0.0, current);
      return current + (value / (1.0 + value)) * diff;
    }
    if (value < 0.0 && current > 0.0) {
      double diff = -fmin(1.0, current);
      return current + (value / (value - 1.0)) * diff;
    }
    return __longlong_as_double(0x7ff8000000000000LL);
  }
  return value;
}

// Step 1 of a fold-in chunk, one warp per interaction (memory-bound): Qui = dot(Xu, Yi) with the reference's exact
// arithmetic (the whole warp when the chain provably never rounds, lane 0 otherwise: common.cuh warp_exact_dot),
// computeTargetQui, dQui, the null verdict, and b = (float)((double)Yi * dQui) (ALSUtils.java:93-95) written over Yi
// in the chunk's scratch.  Keeping this inside the GEMM kernel cost it a tenth of its time in per-thread binary64
// chains and made every CTA redo the b tile -- a DMUL and three conversions per element -- once per output block.
__global__ void __launch_bounds__(256) fold_prep_kernel(int f, int implicit, const float* __restrict__ xu,
                                                        const uint8_t* __restrict__ has_xu, float* __restrict__ yi_to_b,
                                                        const uint8_t* __restrict__ has_yi,
                                                        const double* __restrict__ values, int64_t n,
                                                        uint8_t* __restrict__ out_valid) {
  const int lane = threadIdx.x & 31;
  if (f <= 64) {
    // short rows: four interactions per warp, eight lanes each (lane g of a group holds the c = ceil(f / 8)
    // consecutive elements from g c on).  Every lane stays in the warp-wide shuffles; lanes past the end idle.
    const int64_t kq = ((int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * 4 + (lane >> 3);
    const bool live = kq < n;
    const int g = lane & 7, c = (f + 7) >> 3, j0 = g * c;
    const int cnt = live ? max(0, min(c, f - j0)) : 0;
    const uint8_t has = live ? (has_xu ? has_xu[kq] : 1) : 0;
    float* y = yi_to_b + (live ? kq : 0) * f;
    const float* x = xu + (live ? kq : 0) * f;
    float yr[8], t[8];
#pragma unroll
    for (int i = 0; i < 8; i++) {
      yr[i] = i < cnt ? y[j0 + i] : 0.f;
      t[i] = (has && i < cnt) ? __fmul_rn(x[j0 + i], yr[i]) : 0.f;
    }
    double qui = 0.0;
    const bool fast = group8_exact_sum_regs(t, cnt, f, lane, &qui);
    if (has && !fast) {  // the group's first lane walks the chain (lanes of one group agree on `fast`)
      if (g == 0) {
        qui = 0.0;
        for (int j = 0; j < f; j++) qui = exact_step(qui, x[j], y[j]);
      }
    }
    {
      const double q0 = __shfl_sync(0xffffffffu, qui, 0, 8);  // all lanes execute the shuffle
      if (has && !fast) qui = q0;
    }
    if (!has) qui = 0.0;
    if (!live) return;
    const double target = target_qui(implicit != 0, values[kq], has ? qui : 0.5);
    // computeUpdatedXu returns null when Yi is null (ALSUtils.java:79-81) or the target is NaN (:90-92)
    const bool valid = target == target && (has_yi == nullptr || has_yi[kq]);
    const double dq = valid ? target - qui : 0.0;
#pragma unroll
    for (int i = 0; i < 8; i++)
      if (i < cnt) y[j0 + i] = (float)((double)yr[i] * dq);
    if (g == 0) out_valid[kq] = valid ? 1 : 0;
    return;
  }
  const int64_t k = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (k >= n) return;
  const uint8_t has = has_xu ? has_xu[k] : 1;
  float* y = yi_to_b + k * f;
  const float* x = xu + k * f;
  if (f <= 256) {
    // rows of up to 256 features live in registers: lane l holds the c = ceil(f / 32) consecutive elements from l c on,
    // loaded once for the dot product and for b
    const int c = (f + 31) >> 5, j0 = lane * c;
    const int cnt = max(0, min(c, f - j0));
    float yr[8], t[8];
#pragma unroll
    for (int i = 0; i < 8; i++) {
      yr[i] = i < cnt ? y[j0 + i] : 0.f;
      t[i] = (has && i < cnt) ? __fmul_rn(x[j0 + i], yr[i]) : 0.f;
    }
    double qui = 0.0;
    if (has && !warp_exact_sum_regs(t, cnt, f, lane, &qui)) {
      if (lane == 0) {
        qui = 0.0;
        for (int j = 0; j < f; j++) qui = exact_step(qui, x[j], y[j]);
      }
      qui = __shfl_sync(0xffffffffu, qui, 0);
    }
    const double target = target_qui(implicit != 0, values[k], has ? qui : 0.5);
    // computeUpdatedXu returns null when Yi is null (ALSUtils.java:79-81) or the target is NaN (:90-92)
    const bool valid = target == target && (has_yi == nullptr || has_yi[k]);
    const double dq = valid ? target - qui : 0.0;
#pragma unroll
    for (int i = 0; i < 8; i++)
      if (i < cnt) y[j0 + i] = (float)((double)yr[i] * dq);
    if (lane == 0) out_valid[k] = valid ? 1 : 0;
    return;
  }
  double qui = 0.0;
  if (has) {
    if (!warp_exact_dot(x, y, f, lane, &qui)) {
      if (lane == 0) {
        qui = 0.0;
        for (int j = 0; j < f; j++) qui = exact_step(qui, x[j], y[j]);
      }
      qui = __shfl_sync(0xffffffffu, qui, 0);
    }
  }
  const double target = target_qui(implicit != 0, values[k], has ? qui : 0.5);
  // computeUpdatedXu returns null when Yi is null (ALSUtils.java:79-81) or the target is NaN (:90-92)
  const bool valid = target == target && (has_yi == nullptr || has_yi[k]);
  const double dq = valid ? target - qui : 0.0;
  for (int j = lane; j < f; j += 32) y[j] = (float)((double)y[j] * dq);
  if (lane == 0) out_valid[k] = valid ? 1 : 0;
}

// Step 2.  One CTA = 128 interactions x all F outputs: the binary64 product
// X[k][r] = sum_c b[k][c] * inv[r][c] on the FP64 tensor cores (mma.sync m8n8k4, round 2): CTA tile 128
// (interactions) x 64 (outputs) x 32 (c), a warp owns 16 interactions x 64 outputs = 2 x 8 accumulator tiles
// (32 binary64 registers per thread) and reads one A and one B element per tile and k-step from shared memory.
// Both operand tiles sit in shared memory the way they sit in global memory ([interaction][c] and [output][c]),
// with a row stride of kFoldK + 4 doubles (= 4 mod 16): tile stores are contiguous per warp and the 32 lanes of a
// fragment load hit the 16 bank pairs exactly twice, the minimum.
// The scalar-FMA version issued 32 DFMA + 6 LDS.128 per thread for 1024 multiply-adds and kept the binary64 pipe
// 37 % busy; a DMMA carries 256 multiply-adds.  Sums over c run in index order, four terms per instruction.
constexpr int kFoldK = 32;                 // c per tile
constexpr int kFoldOut = 64;               // outputs per tile
constexpr int kFoldLd = kFoldK + 4;        // padded row stride (doubles), = 4 mod 16

__device__ __forceinline__ void dmma_m8n8k4(double (&d)[2], double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0, %1}, {%2}, {%3}, {%0, %1};"
               : "+d"(d[0]), "+d"(d[1]) : "d"(a), "d"(b));
}

__global__ void __launch_bounds__(256, 2) fold_in_kernel(const double* __restrict__ inv, int f,
                                                         const float* __restrict__ xu,
                                                         const uint8_t* __restrict__ has_xu,
                                                         const float* __restrict__ bmat,  // b of fold_prep_kernel
                                                         int64_t n, float* __restrict__ out,
                                                         const uint8_t* __restrict__ valid_in) {
  __shared__ uint8_t s_valid[kRowsPerCta];
  __shared__ uint8_t s_has[kRowsPerCta];
  extern __shared__ __align__(16) unsigned char fold_smem[];
  // { b tile [k][c], inverse tile [r][c] }  (a second buffer, so that tile t + 1 could be stored while other warps
  // still read tile t, measured slower: 8.2-8.4 ms against 8.0 ms at 1 M x 256)
  constexpr size_t kBufDoubles = (size_t)(kRowsPerCta + kFoldOut) * kFoldLd;
  auto b_tile = [&](int buf) { return reinterpret_cast<double (*)[kFoldLd]>(fold_smem + sizeof(double) * kBufDoubles * buf); };
  auto i_tile = [&](int buf) {
    return reinterpret_cast<double (*)[kFoldLd]>(fold_smem + sizeof(double) * (kBufDoubles * buf + (size_t)kRowsPerCta * kFoldLd));
  };
  const int tid = threadIdx.x;
  const int64_t k0 = (int64_t)blockIdx.x * kRowsPerCta;

  if (tid < kRowsPerCta) {
    const int64_t k = k0 + tid;
    s_valid[tid] = k < n ? valid_in[k] : 0;
    s_has[tid] = k < n ? (has_xu ? has_xu[k] : 1) : 0;
  }
  __syncthreads();

  const int warp = tid >> 5, lane = tid & 31;
  const int fr = lane >> 2, fc = lane & 3;  // fragment coordinates: row (or column) 0..7, k index 0..3
  // Operand tiles travel global -> registers -> shared memory, one tile ahead: the loads of tile t + 1 are issued
  // (all 24 of a thread at once) before the DMMAs of tile t and waited for only when tile t + 1 is stored.
  constexpr int kRowsPerPass = 256 / kFoldK;                     // tile rows covered by one pass of the CTA's threads
  constexpr int kBLoads = kRowsPerCta / kRowsPerPass, kILoads = kFoldOut / kRowsPerPass;
  const int lc = tid % kFoldK, lr = tid / kFoldK;  // this thread's column of a tile and its first row (rows lr + kRowsPerPass u)
  float yreg[kBLoads];
  double ireg[kILoads];
  auto load_tile = [&](int r0, int c0) {
    const bool c_ok = c0 + lc < f;
#pragma unroll
    for (int u = 0; u < kBLoads; u++) {
      const int64_t k = k0 + lr + kRowsPerPass * u;
      yreg[u] = (k < n && c_ok) ? __ldg(bmat + k * f + c0 + lc) : 0.f;
    }
#pragma unroll
    for (int u = 0; u < kILoads; u++) {
      const int r = r0 + lr + kRowsPerPass * u;
      ireg[u] = (r < f && c_ok) ? __ldg(inv + (size_t)r * f + c0 + lc) : 0.0;
    }
  };
  auto store_tile = [&](int buf) {
    double (*Bs)[kFoldLd] = b_tile(buf);
    double (*Is)[kFoldLd] = i_tile(buf);
#pragma unroll
    for (int u = 0; u < kBLoads; u++) {
      const int kk = lr + kRowsPerPass * u;
      Bs[kk][lc] = (double)yreg[u];
    }
#pragma unroll
    for (int u = 0; u < kILoads; u++) Is[lr + kRowsPerPass * u][lc] = ireg[u];  // inverse tile: inv[r][c] (row-major)
  };
  load_tile(0, 0);
  for (int r0 = 0; r0 < f; r0 += kFoldOut) {
    double acc[2][8][2];
#pragma unroll
    for (int mt = 0; mt < 2; mt++)
#pragma unroll
      for (int nt = 0; nt < 8; nt++) acc[mt][nt][0] = acc[mt][nt][1] = 0.0;
    for (int c0 = 0; c0 < f; c0 += kFoldK) {
      __syncthreads();  // every warp is done with the previous tile
      store_tile(0);
      __syncthreads();
      // next tile of this output block, or the first tile of the next one: its global loads fly during the DMMAs
      const bool wrap = c0 + kFoldK >= f;
      const int nr0 = wrap ? r0 + kFoldOut : r0, nc0 = wrap ? 0 : c0 + kFoldK;
      if (nr0 < f) load_tile(nr0, nc0);
      double (*Bs)[kFoldLd] = b_tile(0);
      double (*Is)[kFoldLd] = i_tile(0);
#pragma unroll
      for (int kc = 0; kc < kFoldK / 4; kc++) {
        const double a0 = Bs[warp * 16 + fr][4 * kc + fc];      // A (row-major 8 x 4): lane holds A[row fr][k fc]
        const double a1 = Bs[warp * 16 + 8 + fr][4 * kc + fc];
        double bfrag[8];
#pragma unroll
        for (int nt = 0; nt < 8; nt++) bfrag[nt] = Is[8 * nt + fr][4 * kc + fc];  // B (col-major 4 x 8): lane holds B[k fc][col fr]
#pragma unroll
        for (int nt = 0; nt < 8; nt++) {
          dmma_m8n8k4(acc[0][nt], a0, bfrag[nt]);
          dmma_m8n8k4(acc[1][nt], a1, bfrag[nt]);
        }
      }
    }
    // C fragment (8 x 8): lane holds C[row fr][columns 2 fc, 2 fc + 1].  The old vectors are fetched first, all
    // sixteen of an accumulator row at once (one at a time, every add waited out its own global load).
#pragma unroll
    for (int mt = 0; mt < 2; mt++) {
      const int kk = warp * 16 + 8 * mt + fr;
      const int64_t k = k0 + kk;
      if (k >= n) continue;
      const bool valid = s_valid[kk] != 0, has = s_has[kk] != 0;
      float old[8][2];
#pragma unroll
      for (int nt = 0; nt < 8; nt++)
#pragma unroll
        for (int j = 0; j < 2; j++) {
          const int r = r0 + 8 * nt + 2 * fc + j;
          old[nt][j] = (valid && has && r < f) ? __ldg(xu + k * f + r) : 0.f;
        }
#pragma unroll
      for (int nt = 0; nt < 8; nt++) {
#pragma unroll
        for (int j = 0; j < 2; j++) {
          const int r = r0 + 8 * nt + 2 * fc + j;
          if (r >= f) continue;
          const float dxu = (float)acc[mt][nt][j];  // Solver.java:40-42
          float res = 0.0f;
          if (valid) res = has ? __fadd_rn(old[nt][j], dxu) : dxu;  // ALSUtils.java:98-104
          out[k * f + r] = res;
        }
      }
    }
  }
}

// The same product when rows are 16-byte aligned (features a multiple of 4: every production size): operand tiles
// go global -> shared memory with cp.async, three stages deep, so a tile costs its threads eight copy instructions and
// ONE barrier; b stays binary32 in shared memory (row stride 36 floats = 4 mod 32: a fragment load touches 32
// distinct banks) and is widened when the fragment is loaded.  Out-of-range rows and columns are zero-filled by the
// copy itself (src-size 0).
constexpr int kAsyncStages = 3;
constexpr int kFoldBLdF = kFoldK + 4;  // floats
constexpr size_t kAsyncStageBytes = (size_t)kRowsPerCta * kFoldBLdF * sizeof(float) + (size_t)kFoldOut * kFoldLd * sizeof(double);
constexpr size_t kFoldAsyncSmem = kAsyncStages * kAsyncStageBytes;

__device__ __forceinline__ void cp_async_16_zfill(void* smem_dst, const void* gmem_src, int src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)), "l"(gmem_src),
               "r"(src_bytes) : "memory");
}

__global__ void __launch_bounds__(256, 2) fold_gemm_async_kernel(const double* __restrict__ inv, int f,
                                                                 const float* __restrict__ xu,
                                                                 const uint8_t* __restrict__ has_xu,
                                                                 const float* __restrict__ bmat, int64_t n,
                                                                 float* __restrict__ out,
                                                                 const uint8_t* __restrict__ valid_in) {
  __shared__ uint8_t s_valid[kRowsPerCta];
  __shared__ uint8_t s_has[kRowsPerCta];
  extern __shared__ __align__(16) unsigned char fold_smem[];
  auto b_tile = [&](int st) { return reinterpret_cast<float (*)[kFoldBLdF]>(fold_smem + kAsyncStageBytes * st); };
  auto i_tile = [&](int st) {
    return reinterpret_cast<double (*)[kFoldLd]>(fold_smem + kAsyncStageBytes * st + (size_t)kRowsPerCta * kFoldBLdF * sizeof(float));
  };
  const int tid = threadIdx.x;
  const int64_t k0 = (int64_t)blockIdx.x * kRowsPerCta;
  if (tid < kRowsPerCta) {
    const int64_t k = k0 + tid;
    s_valid[tid] = k < n ? valid_in[k] : 0;
    s_has[tid] = k < n ? (has_xu ? has_xu[k] : 1) : 0;
  }
  const int warp = tid >> 5, lane = tid & 31;
  const int fr = lane >> 2, fc = lane & 3;
  const int n_ctiles = (f + kFoldK - 1) / kFoldK, n_rblocks = (f + kFoldOut - 1) / kFoldOut;
  const int n_tiles = n_ctiles * n_rblocks;
  // tile t = (output block t / n_ctiles, c block t % n_ctiles)
  auto issue = [&](int t) {
    if (t < n_tiles) {
      const int r0 = (t / n_ctiles) * kFoldOut, c0 = (t % n_ctiles) * kFoldK;
      float (*Bs)[kFoldBLdF] = b_tile(t % kAsyncStages);
      double (*Is)[kFoldLd] = i_tile(t % kAsyncStages);
#pragma unroll
      for (int u = 0; u < 4; u++) {  // b: 128 rows x 8 chunks of 4 floats
        const int e = tid + 256 * u, row = e >> 3, ch = e & 7;
        const int64_t k = k0 + row;
        const int c = c0 + 4 * ch;
        const int bytes = (k < n && c < f) ? min(16, (f - c) * 4) : 0;
        cp_async_16_zfill(&Bs[row][4 * ch], bmat + (bytes ? k * f + c : 0), bytes);
      }
#pragma unroll
      for (int u = 0; u < 4; u++) {  // inverse: 64 rows x 16 chunks of 2 doubles
        const int e = tid + 256 * u, row = e >> 4, ch = e & 15;
        const int r = r0 + row, c = c0 + 2 * ch;
        const int bytes = (r < f && c < f) ? min(16, (f - c) * 8) : 0;
        cp_async_16_zfill(&Is[row][2 * ch], inv + (bytes ? (size_t)r * f + c : 0), bytes);
      }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");  // an empty group past the end keeps the counting uniform
  };
  issue(0);
  issue(1);
  double acc[2][8][2];
  for (int t = 0; t < n_tiles; t++) {
    const int cb = t % n_ctiles;
    if (cb == 0) {
#pragma unroll
      for (int mt = 0; mt < 2; mt++)
#pragma unroll
        for (int nt = 0; nt < 8; nt++) acc[mt][nt][0] = acc[mt][nt][1] = 0.0;
    }
    asm volatile("cp.async.wait_group 1;" ::: "memory");  // tile t has landed (this thread's share)
    __syncthreads();                                       // ... everybody's; and tile t - 1 has been read by all
    issue(t + 2);                                          // into the stage tile t - 1 used
    float (*Bs)[kFoldBLdF] = b_tile(t % kAsyncStages);
    double (*Is)[kFoldLd] = i_tile(t % kAsyncStages);
#pragma unroll
    for (int kc = 0; kc < kFoldK / 4; kc++) {
      const double a0 = (double)Bs[warp * 16 + fr][4 * kc + fc];
      const double a1 = (double)Bs[warp * 16 + 8 + fr][4 * kc + fc];
      double bfrag[8];
#pragma unroll
      for (int nt = 0; nt < 8; nt++) bfrag[nt] = Is[8 * nt + fr][4 * kc + fc];
#pragma unroll
      for (int nt = 0; nt < 8; nt++) {
        dmma_m8n8k4(acc[0][nt], a0, bfrag[nt]);
        dmma_m8n8k4(acc[1][nt], a1, bfrag[nt]);
      }
    }
    if (cb == n_ctiles - 1) {
      const int r0 = (t / n_ctiles) * kFoldOut;
#pragma unroll
      for (int mt = 0; mt < 2; mt++) {
        const int kk = warp * 16 + 8 * mt + fr;
        const int64_t k = k0 + kk;
        if (k >= n) continue;
        const bool valid = s_valid[kk] != 0, has = s_has[kk] != 0;
        float old[8][2];
#pragma unroll
        for (int nt = 0; nt < 8; nt++)
#pragma unroll
          for (int j = 0; j < 2; j++) {
            const int r = r0 + 8 * nt + 2 * fc + j;
            old[nt][j] = (valid && has && r < f) ? __ldg(xu + k * f + r) : 0.f;
          }
#pragma unroll
        for (int nt = 0; nt < 8; nt++) {
#pragma unroll
          for (int j = 0; j < 2; j++) {
            const int r = r0 + 8 * nt + 2 * fc + j;
            if (r >= f) continue;
            const float dxu = (float)acc[mt][nt][j];  // Solver.java:40-42
            float res = 0.0f;
            if (valid) res = has ? __fadd_rn(old[nt][j], dxu) : dxu;  // ALSUtils.java:98-104
            out[k * f + r] = res;
          }
        }
      }
    }
  }
  asm volatile("cp.async.wait_group 0;" ::: "memory");
}

constexpr size_t kFoldSmem = sizeof(double) * kFoldLd * (kRowsPerCta + kFoldOut);
// Interactions per pipeline stage: whole waves of the kernel (2 CTAs of 128 interactions per SM), four of them --
// a power-of-two chunk left the last wave of every launch 46 % full on 148 SMs.
int64_t fold_chunk() {
  static const int64_t chunk = [] {
    int dev = 0, sms = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms <= 0) {
      cudaGetLastError();
      sms = 148;
    }
    return (int64_t)sms * 2 * kRowsPerCta * 4;
  }();
  return chunk;
}

int32_t fold_buf(FoldScratch* fs, int i, size_t bytes, void** out) {
  if (fs->d_bytes[i] < bytes) {
    cudaFree(fs->d[i]);
    fs->d[i] = nullptr;
    fs->d_bytes[i] = 0;
    if (cudaMalloc(&fs->d[i], bytes) != cudaSuccess) {
      cudaGetLastError();
      set_error("device allocation of %zu bytes failed (fold-in scratch)", bytes);
      return ORYX_E_NOMEM;
    }
    fs->d_bytes[i] = bytes;
  }
  *out = fs->d[i];
  return ORYX_OK;
}
int32_t fold_pinned(FoldScratch* fs, int i, size_t bytes, void** out) {
  if (fs->h_bytes[i] < bytes) {
    cudaFreeHost(fs->h[i]);
    fs->h[i] = nullptr;
    fs->h_bytes[i] = 0;
    if (cudaMallocHost(&fs->h[i], bytes) != cudaSuccess) {
      cudaGetLastError();
      set_error("pinned allocation of %zu bytes failed (fold-in staging)", bytes);
      return ORYX_E_NOMEM;
    }
    fs->h_bytes[i] = bytes;
  }
  *out = fs->h[i];
  return ORYX_OK;
}

struct FoldBuffers {
  float *d_xu, *d_yi, *d_out[2];
  double* d_val;
  uint8_t *d_has_xu, *d_has_yi, *d_valid[2];
  int32_t* d_idx;          // [4 x c_max]
  float* h_out[2];         // pinned
  uint8_t* h_valid[2];     // pinned
  int32_t* h_idx[2];       // pinned, [4 x c_max] each
  int64_t c_max;
};

int32_t fold_prepare(FoldScratch* fs, int f, int64_t n, bool need_idx, FoldBuffers* b) {
  const int64_t c_max = std::min<int64_t>(fold_chunk(), n);
  b->c_max = c_max;
  void* p = nullptr;
  ORYX_TRY(fold_buf(fs, 0, (size_t)c_max * f * 4, &p)); b->d_xu = (float*)p;
  ORYX_TRY(fold_buf(fs, 1, (size_t)c_max * f * 4, &p)); b->d_yi = (float*)p;
  ORYX_TRY(fold_buf(fs, 2, (size_t)c_max * f * 4, &p)); b->d_out[0] = (float*)p;
  ORYX_TRY(fold_buf(fs, 3, (size_t)c_max * f * 4, &p)); b->d_out[1] = (float*)p;
  ORYX_TRY(fold_buf(fs, 4, (size_t)c_max * 8, &p)); b->d_val = (double*)p;
  ORYX_TRY(fold_buf(fs, 5, (size_t)c_max * 2, &p)); b->d_has_xu = (uint8_t*)p; b->d_has_yi = b->d_has_xu + c_max;
  ORYX_TRY(fold_buf(fs, 6, (size_t)c_max * 2, &p)); b->d_valid[0] = (uint8_t*)p; b->d_valid[1] = b->d_valid[0] + c_max;
  b->d_idx = nullptr;
  b->h_idx[0] = b->h_idx[1] = nullptr;
  if (need_idx) {
    ORYX_TRY(fold_buf(fs, 7, (size_t)c_max * 16, &p)); b->d_idx = (int32_t*)p;
    ORYX_TRY(fold_pinned(fs, 3, (size_t)c_max * 32, &p)); b->h_idx[0] = (int32_t*)p; b->h_idx[1] = b->h_idx[0] + 4 * c_max;
  }
  ORYX_TRY(fold_pinned(fs, 0, (size_t)c_max * f * 4, &p)); b->h_out[0] = (float*)p;
  ORYX_TRY(fold_pinned(fs, 1, (size_t)c_max * f * 4, &p)); b->h_out[1] = (float*)p;
  ORYX_TRY(fold_pinned(fs, 2, (size_t)c_max * 2, &p)); b->h_valid[0] = (uint8_t*)p; b->h_valid[1] = b->h_valid[0] + c_max;
  if (fs->copy_stream == nullptr) ORYX_CUDA(cudaStreamCreateWithFlags(&fs->copy_stream, cudaStreamNonBlocking));
  for (int i = 0; i < 2; i++) {
    if (!fs->ev_done[i]) ORYX_CUDA(cudaEventCreateWithFlags(&fs->ev_done[i], cudaEventDisableTiming));
    if (!fs->ev_copied[i]) ORYX_CUDA(cudaEventCreateWithFlags(&fs->ev_copied[i], cudaEventDisableTiming));
    if (!fs->ev_k0[i]) ORYX_CUDA(cudaEventCreate(&fs->ev_k0[i]));
    if (!fs->ev_k1[i]) ORYX_CUDA(cudaEventCreate(&fs->ev_k1[i]));
  }
  return ORYX_OK;
}

// The pipeline shared by the three entry points.  stage(off, c, parity) puts chunk [off, off + c) of the
// inputs into b.d_xu / d_has_xu / d_yi / d_has_yi / d_val on `st` (host-to-device copies or gathers from
// the device stores).  Results leave through two pinned buffers on a second stream, so the copy-out of
// one chunk and the host's slot translation for the next overlap the kernel of the chunk in between.
template <class Stage>
int32_t fold_pipeline(Solver* s, const FoldBuffers& b, int f, int implicit, int64_t n, bool use_has_yi, float* out,
                      uint8_t* out_valid, Stage stage) {
  FoldScratch* fs = solver_fold_scratch(s);
  cudaStream_t st = solver_stream(s);
  ORYX_TRY(ensure_max_dynamic_smem(reinterpret_cast<const void*>(fold_in_kernel), kFoldSmem));
  ORYX_TRY(ensure_max_dynamic_smem(reinterpret_cast<const void*>(fold_gemm_async_kernel), kFoldAsyncSmem));
  {
    // two CTAs of 108 KB each per SM need the largest shared-memory carve-out; ask for it once per device
    static std::once_flag once[64];
    const int dev = solver_device(s);
    if (dev >= 0 && dev < 64) {
      std::call_once(once[dev], [fs] {
        cudaFuncSetAttribute(reinterpret_cast<const void*>(fold_gemm_async_kernel),
                             cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
        int per_sm = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fold_gemm_async_kernel, 256, kFoldAsyncSmem) == cudaSuccess)
          fs->gemm_ctas_per_sm = per_sm;
        cudaGetLastError();
      });
    }
  }
  // rows of b, of the inverse and of the outputs are 16-byte aligned when the row length is a multiple of 4
  const bool async_gemm = (f & 3) == 0 && std::getenv("ORYX_FOLD_NO_ASYNC") == nullptr;
  const auto wall0 = std::chrono::steady_clock::now();
  double kernel_ms = 0;
  int chunks = 0;
  // A caller whose result array is page-locked (cudaHostAlloc / cudaHostRegister: a JNI direct buffer registered
  // once, a torch pinned tensor) gets the rows copied straight into it; pageable arrays go through the pinned
  // staging buffers and a multi-threaded memcpy.
  bool direct = false;
  {
    cudaPointerAttributes attr{};
    if (cudaPointerGetAttributes(&attr, out) == cudaSuccess) direct = attr.type == cudaMemoryTypeHost;
    else cudaGetLastError();
  }
  auto finalize = [&](int64_t off, int64_t c, int par) -> cudaError_t {
    cudaError_t e = cudaEventSynchronize(fs->ev_copied[par]);
    if (e != cudaSuccess) return e;
    if (direct) {
      std::memcpy(out_valid + off, b.h_valid[par], (size_t)c);
      float ms = 0.f;
      if (cudaEventElapsedTime(&ms, fs->ev_k0[par], fs->ev_k1[par]) == cudaSuccess) kernel_ms += ms;
      else cudaGetLastError();
      return cudaSuccess;
    }
    // pinned staging -> the caller's (pageable, often freshly allocated) array: a few threads, because one
    // core's memcpy and first-touch page faults were what the whole call waited for
    const size_t bytes = (size_t)c * f * sizeof(float);
    const int n_thr = bytes >= ((size_t)64 << 20) ? 8 : bytes >= ((size_t)16 << 20) ? 4 : 1;
    if (n_thr == 1) {
      std::memcpy(out + off * f, b.h_out[par], bytes);
    } else {
      std::thread workers[8];
      const size_t piece = (bytes / n_thr + 4095) & ~(size_t)4095;
      for (int t = 0; t < n_thr; t++) {
        const size_t lo = std::min(bytes, piece * t), hi = std::min(bytes, piece * (t + 1));
        workers[t] = std::thread([=, &b] {
          std::memcpy(reinterpret_cast<char*>(out + off * f) + lo, reinterpret_cast<const char*>(b.h_out[par]) + lo, hi - lo);
        });
      }
      for (int t = 0; t < n_thr; t++) workers[t].join();
    }
    std::memcpy(out_valid + off, b.h_valid[par], (size_t)c);
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, fs->ev_k0[par], fs->ev_k1[par]) == cudaSuccess) kernel_ms += ms;
    else cudaGetLastError();
    return cudaSuccess;
  };
  int64_t prev_off = -1, prev_c = 0;
  cudaError_t e = cudaSuccess;
  const int64_t chunk = fold_chunk();
  for (int64_t off = 0; off < n; off += chunk, chunks++) {
    const int64_t c = std::min<int64_t>(chunk, n - off);
    const int par = chunks & 1;
    e = stage(off, c, par);
    if (e != cudaSuccess) break;
    const unsigned grid = (unsigned)((c + kRowsPerCta - 1) / kRowsPerCta);
    if ((e = cudaEventRecord(fs->ev_k0[par], st)) != cudaSuccess) break;
    // one warp per interaction, or per four interactions when rows are short (f <= 64)
    fold_prep_kernel<<<(unsigned)(((f <= 64 ? (c + 3) / 4 : c) * 32 + 255) / 256), 256, 0, st>>>(f, implicit, b.d_xu, b.d_has_xu, b.d_yi,
                                                                      use_has_yi ? b.d_has_yi : nullptr, b.d_val, c,
                                                                      b.d_valid[par]);
    if (async_gemm) {
      fold_gemm_async_kernel<<<grid, 256, kFoldAsyncSmem, st>>>(solver_inverse(s), f, b.d_xu, b.d_has_xu, b.d_yi, c,
                                                               b.d_out[par], b.d_valid[par]);
    } else {
      fold_in_kernel<<<grid, 256, kFoldSmem, st>>>(solver_inverse(s), f, b.d_xu, b.d_has_xu, b.d_yi, c, b.d_out[par],
                                                   b.d_valid[par]);
    }
    if ((e = cudaGetLastError()) != cudaSuccess) break;
    if ((e = cudaEventRecord(fs->ev_k1[par], st)) != cudaSuccess) break;
    if ((e = cudaEventRecord(fs->ev_done[par], st)) != cudaSuccess) break;
    if ((e = cudaStreamWaitEvent(fs->copy_stream, fs->ev_done[par], 0)) != cudaSuccess) break;
    if ((e = cudaMemcpyAsync(direct ? out + off * f : b.h_out[par], b.d_out[par], (size_t)c * f * sizeof(float), cudaMemcpyDeviceToHost, fs->copy_stream)) != cudaSuccess) break;
    if ((e = cudaMemcpyAsync(b.h_valid[par], b.d_valid[par], (size_t)c, cudaMemcpyDeviceToHost, fs->copy_stream)) != cudaSuccess) break;
    if ((e = cudaEventRecord(fs->ev_copied[par], fs->copy_stream)) != cudaSuccess) break;
    if (prev_off >= 0 && (e = finalize(prev_off, prev_c, par ^ 1)) != cudaSuccess) break;
    prev_off = off;
    prev_c = c;
  }
  if (e == cudaSuccess && prev_off >= 0) e = finalize(prev_off, prev_c, (chunks - 1) & 1);
  if (e != cudaSuccess) {
    cudaStreamSynchronize(st);
    cudaStreamSynchronize(fs->copy_stream);
    cudaGetLastError();
    set_error("fold-in batch failed: %s", cudaGetErrorString(e));
    return ORYX_E_CUDA;
  }
  fs->last_kernel_ms = kernel_ms;
  fs->last_chunks = chunks;
  fs->last_n = n;
  fs->last_wall_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - wall0).count();
  return ORYX_OK;
}

struct ScopedDevice {
  int prev = -1;
  bool ok = false;
  explicit ScopedDevice(int dev) {
    if (cudaGetDevice(&prev) != cudaSuccess) { cudaGetLastError(); prev = -1; }
    ok = cudaSetDevice(dev) == cudaSuccess;
    if (!ok) { cudaGetLastError(); set_error("cannot select CUDA device %d", dev); }
  }
  ~ScopedDevice() { if (prev >= 0) cudaSetDevice(prev); }
};

}  // namespace

void fold_scratch_free(FoldScratch* fs) {
  for (auto& p : fs->d) { cudaFree(p); p = nullptr; }
  for (auto& p : fs->h) { cudaFreeHost(p); p = nullptr; }
  for (int i = 0; i < 2; i++) {
    if (fs->ev_done[i]) cudaEventDestroy(fs->ev_done[i]);
    if (fs->ev_copied[i]) cudaEventDestroy(fs->ev_copied[i]);
    if (fs->ev_k0[i]) cudaEventDestroy(fs->ev_k0[i]);
    if (fs->ev_k1[i]) cudaEventDestroy(fs->ev_k1[i]);
  }
  if (fs->copy_stream) cudaStreamDestroy(fs->copy_stream);
  *fs = FoldScratch{};
}

}  // namespace oryx

using namespace oryx;

// ---- stores of other processes (oryx_peer_stores_*): what one rank knows about every rank's X and Y.
namespace {

constexpr uint32_t kPeerMagic = 0x4f525950u;  // "ORYP"

struct PeerStoreWire {
  uint32_t present;               // 0: the store holds nothing that could be mapped (no rows yet)
  uint32_t pad;
  unsigned long long gen;         // Store::layout_gen when the handle was made
  long long row_slot_cap, cap_slots;
  cudaIpcMemHandle_t vec, row_slot;
};
struct PeerWire {
  uint32_t magic, version;
  int32_t device, features, fpad, rank;
  PeerStoreWire st[2];
  cudaIpcMemHandle_t desc;        // the owner's [2] layout generations
};
static_assert(sizeof(PeerWire) <= ORYX_PEER_STORES_HANDLE_BYTES, "handle size");

struct PeerRank {
  bool connected = false;
  void* vec[2] = {nullptr, nullptr};
  void* row_slot[2] = {nullptr, nullptr};
  unsigned long long* desc = nullptr;
  unsigned long long gen[2] = {0, 0};
  long long row_slot_cap[2] = {0, 0}, cap_slots[2] = {0, 0};
  bool opened_vec[2] = {false, false}, opened_rs[2] = {false, false}, opened_desc = false;
};

void peer_rank_close(PeerRank* r) {
  for (int w = 0; w < 2; w++) {
    if (r->opened_vec[w]) cudaIpcCloseMemHandle(r->vec[w]);
    if (r->opened_rs[w]) cudaIpcCloseMemHandle(r->row_slot[w]);
  }
  if (r->opened_desc) cudaIpcCloseMemHandle(r->desc);
  cudaGetLastError();
  *r = PeerRank{};
}

}  // namespace

struct oryx_peer_stores {
  Model* m = nullptr;
  int world = 1, rank = 0;
  PeerRank peer[ORYX_MAX_WORLD];
  unsigned long long* h_gens = nullptr;  // pinned [world x 2]: the owners' current generations, read per call
};

extern "C" {

int32_t oryx_peer_stores_create(oryx_model* mh, int32_t world, int32_t rank, oryx_peer_stores** out) {
  ORYX_ARG(mh && out, "null argument");
  ORYX_ARG(world >= 1 && world <= ORYX_MAX_WORLD, "world out of range [1, ORYX_MAX_WORLD]");
  ORYX_ARG(rank >= 0 && rank < world, "rank out of range");
  Model* m = reinterpret_cast<Model*>(mh);
  ScopedDevice dev(m->device);
  if (!dev.ok) return ORYX_E_CUDA;
  auto* p = new (std::nothrow) oryx_peer_stores();
  if (!p) { set_error("host allocation failed"); return ORYX_E_NOMEM; }
  p->m = m;
  p->world = world;
  p->rank = rank;
  if (cudaMallocHost(&p->h_gens, sizeof(unsigned long long) * 2 * ORYX_MAX_WORLD) != cudaSuccess) {
    cudaGetLastError();
    delete p;
    set_error("pinned allocation failed");
    return ORYX_E_NOMEM;
  }
  *out = p;
  return ORYX_OK;
}

int32_t oryx_peer_stores_handle(oryx_peer_stores* p, void* out_handle) {
  ORYX_ARG(p && out_handle, "null argument");
  Model* m = p->m;
  ScopedDevice dev(m->device);
  if (!dev.ok) return ORYX_E_CUDA;
  std::unique_lock<std::shared_mutex> lk(m->mu);  // the stores must not move while they are described
  if (m->d_share_desc == nullptr) {
    ORYX_CUDA(cudaMalloc(&m->d_share_desc, 256));
    ORYX_CUDA(cudaMemset(m->d_share_desc, 0, 256));
  }
  unsigned long long gens[2] = {m->store[0].layout_gen, m->store[1].layout_gen};
  ORYX_CUDA(cudaMemcpy(m->d_share_desc, gens, sizeof(gens), cudaMemcpyHostToDevice));
  PeerWire w{};
  w.magic = kPeerMagic;
  w.version = 1;
  w.device = m->device;
  w.features = m->cfg.features;
  w.fpad = m->store[ORYX_Y].fpad;
  w.rank = p->rank;
  ORYX_CUDA(cudaIpcGetMemHandle(&w.desc, m->d_share_desc));
  for (int which = 0; which < 2; which++) {
    const Store* s = &m->store[which];
    PeerStoreWire& sw = w.st[which];
    sw.gen = s->layout_gen;
    if (s->d_vec == nullptr || s->d_row_slot == nullptr) continue;  // nothing stored yet: every row is absent
    sw.present = 1;
    sw.row_slot_cap = s->row_slot_cap;
    sw.cap_slots = s->cap_slots;
    ORYX_CUDA(cudaIpcGetMemHandle(&sw.vec, s->d_vec));
    ORYX_CUDA(cudaIpcGetMemHandle(&sw.row_slot, s->d_row_slot));
  }
  std::memset(out_handle, 0, ORYX_PEER_STORES_HANDLE_BYTES);
  std::memcpy(out_handle, &w, sizeof(w));
  return ORYX_OK;
}

int32_t oryx_peer_stores_connect(oryx_peer_stores* p, const void* handles) {
  ORYX_ARG(p && handles, "null argument");
  Model* m = p->m;
  ScopedDevice dev(m->device);
  if (!dev.ok) return ORYX_E_CUDA;
  for (int r = 0; r < p->world; r++) peer_rank_close(&p->peer[r]);
  for (int r = 0; r < p->world; r++) {
    PeerRank& pr = p->peer[r];
    if (r == p->rank) {  // own stores: used directly, re-read on every call
      pr.connected = true;
      continue;
    }
    PeerWire w;
    std::memcpy(&w, reinterpret_cast<const unsigned char*>(handles) + (size_t)r * ORYX_PEER_STORES_HANDLE_BYTES, sizeof(w));
    if (w.magic != kPeerMagic || w.version != 1 || w.rank != r) {
      set_error("entry %d of the handle table is not rank %d's oryx_peer_stores_handle", r, r);
      return ORYX_E_INVALID;
    }
    if (w.features != m->cfg.features || w.fpad != m->store[ORYX_Y].fpad) {
      set_error("rank %d stores %d features, this rank %d", r, w.features, m->cfg.features);
      return ORYX_E_INVALID;
    }
    auto open = [&](const cudaIpcMemHandle_t& h, void** ptr, const char* what) -> int32_t {
      cudaError_t e = cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess);
      if (e != cudaSuccess) {
        cudaGetLastError();
        set_error("cannot map %s of rank %d into rank %d (cudaIpcOpenMemHandle: %s)", what, r, p->rank, cudaGetErrorString(e));
        return ORYX_E_CUDA;
      }
      return ORYX_OK;
    };
    void* d = nullptr;
    ORYX_TRY(open(w.desc, &d, "the layout descriptor"));
    pr.desc = reinterpret_cast<unsigned long long*>(d);
    pr.opened_desc = true;
    for (int which = 0; which < 2; which++) {
      pr.gen[which] = w.st[which].gen;
      if (!w.st[which].present) continue;
      ORYX_TRY(open(w.st[which].vec, &pr.vec[which], "a vector array"));
      pr.opened_vec[which] = true;
      ORYX_TRY(open(w.st[which].row_slot, &pr.row_slot[which], "a row->slot table"));
      pr.opened_rs[which] = true;
      pr.row_slot_cap[which] = w.st[which].row_slot_cap;
      pr.cap_slots[which] = w.st[which].cap_slots;
    }
    pr.connected = true;
  }
  return ORYX_OK;
}

int32_t oryx_peer_stores_destroy(oryx_peer_stores* p) {
  if (!p) return ORYX_OK;
  ScopedDevice dev(p->m->device);
  for (int r = 0; r < p->world; r++) peer_rank_close(&p->peer[r]);
  cudaFreeHost(p->h_gens);
  cudaGetLastError();
  delete p;
  return ORYX_OK;
}

int32_t oryx_fold_in_rows_peers(oryx_solver* sh, oryx_peer_stores* p, int64_t xu_row_stride, int64_t yi_row_stride,
                                int32_t implicit, int32_t xu_which, const int64_t* xu_rows, const int64_t* yi_rows,
                                const double* values, int64_t n, float* out, uint8_t* out_valid) {
  ORYX_ARG(sh && p && xu_rows && yi_rows && values && out && out_valid, "null argument");
  ORYX_ARG(xu_which == ORYX_X || xu_which == ORYX_Y, "xu_which must be ORYX_X or ORYX_Y");
  ORYX_ARG(xu_row_stride > 0 && yi_row_stride > 0, "row strides must be positive");
  ORYX_ARG(n >= 0, "n must be >= 0");
  if (n == 0) return ORYX_OK;
  Solver* s = reinterpret_cast<Solver*>(sh);
  Model* m = p->m;
  const int f = solver_n(s);
  ORYX_ARG(f == m->cfg.features, "solver and model disagree on the number of features");
  ORYX_ARG(solver_device(s) == m->device, "solver and model live on different devices");
  for (int r = 0; r < p->world; r++) ORYX_ARG(p->peer[r].connected, "oryx_peer_stores_connect has not succeeded");
  ScopedDevice dev(m->device);
  if (!dev.ok) return ORYX_E_CUDA;
  std::lock_guard<std::mutex> lk(solver_mutex(s));
  std::shared_lock<std::shared_mutex> model_lock(m->mu);  // the own stores must not move while rows are gathered
  cudaStream_t st = solver_stream(s);
  // a peer whose arrays moved since the handles were exchanged must not be read through the old mapping
  for (int r = 0; r < p->world; r++) {
    if (r == p->rank) continue;
    ORYX_CUDA(cudaMemcpyAsync(p->h_gens + 2 * r, p->peer[r].desc, 2 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
  }
  ORYX_CUDA(cudaStreamSynchronize(st));
  for (int r = 0; r < p->world; r++) {
    if (r == p->rank) continue;
    for (int which = 0; which < 2; which++) {
      if (p->h_gens[2 * r + which] != p->peer[r].gen[which]) {
        set_error("the %s store of rank %d moved after its handle was exchanged: exchange fresh oryx_peer_stores handles",
                  which == ORYX_X ? "X" : "Y", r);
        return ORYX_E_INVALID;
      }
    }
  }
  const int yi_which = xu_which == ORYX_X ? ORYX_Y : ORYX_X;
  auto view = [&](int which) {
    PeerView v{};
    v.world = p->world;
    for (int r = 0; r < p->world; r++) {
      if (r == p->rank) {
        const Store* own = &m->store[which];
        v.vec[r] = own->d_vec;
        v.row_slot[r] = own->d_vec ? own->d_row_slot : nullptr;
        v.row_slot_cap[r] = own->row_slot_cap;
        v.cap_slots[r] = own->cap_slots;
      } else {
        v.vec[r] = reinterpret_cast<const float*>(p->peer[r].vec[which]);
        v.row_slot[r] = reinterpret_cast<const int32_t*>(p->peer[r].row_slot[which]);
        v.row_slot_cap[r] = p->peer[r].row_slot_cap[which];
        v.cap_slots[r] = p->peer[r].cap_slots[which];
      }
    }
    return v;
  };
  const PeerView vx = view(xu_which), vy = view(yi_which);
  const int fpad = m->store[ORYX_Y].fpad;
  FoldBuffers b{};
  ORYX_TRY(fold_prepare(solver_fold_scratch(s), f, n, true, &b));
  const int64_t cm = b.c_max;
  return fold_pipeline(s, b, f, implicit, n, true, out, out_valid, [&](int64_t off, int64_t c, int par) -> cudaError_t {
    // the index buffers hold [2 x c_max] 64-bit global rows here: xu rows, then yi rows
    long long* h = reinterpret_cast<long long*>(b.h_idx[par]);
    std::memcpy(h, xu_rows + off, (size_t)c * sizeof(long long));
    std::memcpy(h + cm, yi_rows + off, (size_t)c * sizeof(long long));
    long long* d = reinterpret_cast<long long*>(b.d_idx);
    cudaError_t e = cudaMemcpyAsync(d, h, (size_t)cm * 2 * sizeof(long long), cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(b.d_val, values + off, (size_t)c * sizeof(double), cudaMemcpyHostToDevice, st);
    if (e != cudaSuccess) return e;
    const unsigned g = (unsigned)((c * 32 + 255) / 256);
    gather_rows_peers_kernel<<<g, 256, 0, st>>>(vx, (long long)xu_row_stride, fpad, f, d, c, b.d_xu, b.d_has_xu);
    gather_rows_peers_kernel<<<g, 256, 0, st>>>(vy, (long long)yi_row_stride, fpad, f, d + cm, c, b.d_yi, b.d_has_yi);
    m->launches += 2;
    return cudaGetLastError();
  });
}

int32_t oryx_fold_in_batch(oryx_solver* sh, int32_t implicit, const float* xu, const uint8_t* has_xu,
                           const float* yi, const double* values, int64_t n, float* out, uint8_t* out_valid) {
  ORYX_ARG(sh && yi && values && out && out_valid, "null argument");
  ORYX_ARG(n >= 0, "n must be >= 0");
  if (n == 0) return ORYX_OK;
  Solver* s = reinterpret_cast<Solver*>(sh);
  const int f = solver_n(s);
  ScopedDevice dev(solver_device(s));
  if (!dev.ok) return ORYX_E_CUDA;
  std::lock_guard<std::mutex> lk(solver_mutex(s));
  cudaStream_t st = solver_stream(s);
  FoldBuffers b{};
  ORYX_TRY(fold_prepare(solver_fold_scratch(s), f, n, false, &b));
  return fold_pipeline(s, b, f, implicit, n, false, out, out_valid, [&](int64_t off, int64_t c, int) -> cudaError_t {
    cudaError_t e;
    if (xu) e = cudaMemcpyAsync(b.d_xu, xu + off * f, (size_t)c * f * sizeof(float), cudaMemcpyHostToDevice, st);
    else e = cudaMemsetAsync(b.d_xu, 0, (size_t)c * f * sizeof(float), st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(b.d_yi, yi + off * f, (size_t)c * f * sizeof(float), cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(b.d_val, values + off, (size_t)c * sizeof(double), cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) {
      if (has_xu) e = cudaMemcpyAsync(b.d_has_xu, has_xu + off, (size_t)c, cudaMemcpyHostToDevice, st);
      else e = cudaMemsetAsync(b.d_has_xu, xu ? 1 : 0, (size_t)c, st);
    }
    return e;
  });
}

int32_t oryx_fold_in_rows(oryx_solver* sh, oryx_model* mh, int32_t implicit, int32_t xu_which,
                          const int32_t* xu_rows, const int32_t* yi_rows, const double* values, int64_t n,
                          float* out, uint8_t* out_valid) {
  ORYX_ARG(sh && mh && xu_rows && yi_rows && values && out && out_valid, "null argument");
  ORYX_ARG(xu_which == ORYX_X || xu_which == ORYX_Y, "xu_which must be ORYX_X or ORYX_Y");
  ORYX_ARG(n >= 0, "n must be >= 0");
  if (n == 0) return ORYX_OK;
  Solver* s = reinterpret_cast<Solver*>(sh);
  Model* m = reinterpret_cast<Model*>(mh);
  const int f = solver_n(s);
  ORYX_ARG(f == m->cfg.features, "solver and model disagree on the number of features");
  ORYX_ARG(solver_device(s) == m->device, "solver and model live on different devices");
  ScopedDevice dev(m->device);
  if (!dev.ok) return ORYX_E_CUDA;
  std::lock_guard<std::mutex> lk(solver_mutex(s));
  std::shared_lock<std::shared_mutex> model_lock(m->mu);  // the stores must not move while rows are gathered
  const Store* sx = &m->store[xu_which];
  const Store* sy = &m->store[xu_which == ORYX_X ? ORYX_Y : ORYX_X];
  cudaStream_t st = solver_stream(s);
  FoldBuffers b{};
  ORYX_TRY(fold_prepare(solver_fold_scratch(s), f, n, true, &b));
  auto slot_of = [](const Store* store, int32_t row) -> int32_t {
    return (row >= 0 && (size_t)row < store->row2slot.size()) ? store->row2slot[(size_t)row] : -1;
  };
  const int64_t cm = b.c_max;
  return fold_pipeline(s, b, f, implicit, n, true, out, out_valid, [&](int64_t off, int64_t c, int par) -> cudaError_t {
    int32_t* h = b.h_idx[par];
    for (int64_t i = 0; i < c; i++) {
      h[i] = slot_of(sx, xu_rows[off + i]);
      h[cm + i] = slot_of(sy, yi_rows[off + i]);
    }
    cudaError_t e = cudaMemcpyAsync(b.d_idx, h, (size_t)cm * 2 * sizeof(int32_t), cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(b.d_val, values + off, (size_t)c * sizeof(double), cudaMemcpyHostToDevice, st);
    if (e != cudaSuccess) return e;
    const unsigned g = (unsigned)((c * 32 + 255) / 256);
    gather_rows_kernel<<<g, 256, 0, st>>>(sx->d_vec, sx->fpad, f, b.d_idx, c, b.d_xu, b.d_has_xu);
    gather_rows_kernel<<<g, 256, 0, st>>>(sy->d_vec, sy->fpad, f, b.d_idx + cm, c, b.d_yi, b.d_has_yi);
    m->launches += 3;
    return cudaGetLastError();
  });
}

int32_t oryx_fold_in_rows_sharded(oryx_solver* sh, oryx_model* const* models, int32_t world, int64_t xu_row_stride,
                                  int64_t yi_row_stride, int32_t implicit, int32_t xu_which, const int64_t* xu_rows,
                                  const int64_t* yi_rows, const double* values, int64_t n, float* out,
                                  uint8_t* out_valid) {
  ORYX_ARG(sh && models && xu_rows && yi_rows && values && out && out_valid, "null argument");
  ORYX_ARG(world >= 1 && world <= ORYX_MAX_WORLD, "world out of range [1, ORYX_MAX_WORLD]");
  ORYX_ARG(xu_which == ORYX_X || xu_which == ORYX_Y, "xu_which must be ORYX_X or ORYX_Y");
  ORYX_ARG(xu_row_stride > 0 && yi_row_stride > 0, "row strides must be positive");
  ORYX_ARG(n >= 0, "n must be >= 0");
  if (n == 0) return ORYX_OK;
  Solver* s = reinterpret_cast<Solver*>(sh);
  const int f = solver_n(s);
  const int dev = solver_device(s);
  for (int r = 0; r < world; r++) {
    ORYX_ARG(models[r] != nullptr, "null model in the shard list");
    ORYX_ARG(reinterpret_cast<Model*>(models[r])->cfg.features == f, "solver and shards disagree on the number of features");
  }
  ScopedDevice scoped(dev);
  if (!scoped.ok) return ORYX_E_CUDA;
  // the kernels run on the solver's GPU and read the other shards' rows over NVLink
  for (int r = 0; r < world; r++) {
    const int peer = reinterpret_cast<Model*>(models[r])->device;
    if (peer == dev) continue;
    int can = 0;
    ORYX_CUDA(cudaDeviceCanAccessPeer(&can, dev, peer));
    if (!can) {
      set_error("device %d cannot read device %d as a peer", dev, peer);
      return ORYX_E_UNSUPPORTED;
    }
    cudaError_t e = cudaDeviceEnablePeerAccess(peer, 0);
    if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) {
      cudaGetLastError();
      set_error("cudaDeviceEnablePeerAccess(%d -> %d): %s", dev, peer, cudaGetErrorString(e));
      return ORYX_E_CUDA;
    }
    cudaGetLastError();
  }
  std::lock_guard<std::mutex> lk(solver_mutex(s));
  // the stores must not move while rows are gathered: shared locks on every shard, in shard order
  std::vector<std::shared_lock<std::shared_mutex>> locks;
  locks.reserve((size_t)world);
  for (int r = 0; r < world; r++) locks.emplace_back(reinterpret_cast<Model*>(models[r])->mu);
  const int yi_which = xu_which == ORYX_X ? ORYX_Y : ORYX_X;
  ShardBases bx{}, by{};
  int fpad = 0;
  for (int r = 0; r < world; r++) {
    Model* m = reinterpret_cast<Model*>(models[r]);
    bx.base[r] = m->store[xu_which].d_vec;
    by.base[r] = m->store[yi_which].d_vec;
    fpad = m->store[ORYX_Y].fpad;
  }
  cudaStream_t st = solver_stream(s);
  FoldBuffers b{};
  ORYX_TRY(fold_prepare(solver_fold_scratch(s), f, n, true, &b));
  const int64_t cm = b.c_max;
  // global row -> (owning shard, slot in that shard's store); -1 slot when the row has no vector
  auto locate = [&](int which, int64_t stride, int64_t row, int32_t* owner, int32_t* slot) {
    *owner = 0;
    *slot = -1;
    if (row < 0) return;
    const int64_t r = row / stride;
    if (r >= world) return;
    const Store* store = &reinterpret_cast<Model*>(models[r])->store[which];
    const int64_t local = row - r * stride;
    *owner = (int32_t)r;
    if ((size_t)local < store->row2slot.size()) *slot = store->row2slot[(size_t)local];
  };
  return fold_pipeline(s, b, f, implicit, n, true, out, out_valid, [&](int64_t off, int64_t c, int par) -> cudaError_t {
    int32_t* h = b.h_idx[par];  // [4 x c_max]: xu owner, xu slot, yi owner, yi slot
    for (int64_t i = 0; i < c; i++) {
      locate(xu_which, xu_row_stride, xu_rows[off + i], &h[i], &h[cm + i]);
      locate(yi_which, yi_row_stride, yi_rows[off + i], &h[2 * cm + i], &h[3 * cm + i]);
    }
    cudaError_t e = cudaMemcpyAsync(b.d_idx, h, (size_t)cm * 4 * sizeof(int32_t), cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(b.d_val, values + off, (size_t)c * sizeof(double), cudaMemcpyHostToDevice, st);
    if (e != cudaSuccess) return e;
    const unsigned g = (unsigned)((c * 32 + 255) / 256);
    gather_rows_sharded_kernel<<<g, 256, 0, st>>>(bx, fpad, f, b.d_idx, b.d_idx + cm, c, b.d_xu, b.d_has_xu);
    gather_rows_sharded_kernel<<<g, 256, 0, st>>>(by, fpad, f, b.d_idx + 2 * cm, b.d_idx + 3 * cm, c, b.d_yi, b.d_has_yi);
    return cudaGetLastError();
  });
}

// Development aid (not in the public header): device time of fold_in_kernel summed over the chunks of the
// solver's last fold-in call (ms), number of chunks, interactions, wall time of the call (ms).
int32_t oryx_debug_fold_in_timings(oryx_solver* sh, double* out4) {
  Solver* s = reinterpret_cast<Solver*>(sh);
  FoldScratch* fs = solver_fold_scratch(s);
  out4[0] = fs->last_kernel_ms;
  out4[1] = fs->last_chunks;
  out4[2] = (double)fs->last_n;
  out4[3] = fs->last_wall_ms;
  out4[4] = (double)fs->gemm_ctas_per_sm;
  return ORYX_OK;
}

int32_t oryx_gram(oryx_model* mh, int32_t which, double* out, int64_t* out_rows_used) {
  ORYX_ARG(mh && out, "null argument");
  ORYX_ARG(which == ORYX_X || which == ORYX_Y, "which must be ORYX_X or ORYX_Y");
  Model* m = reinterpret_cast<Model*>(mh);
  ScopedDevice dev(m->device);
  if (!dev.ok) return ORYX_E_CUDA;
  const int F = m->cfg.features;
  QueryCtx* c = nullptr;
  ORYX_TRY(ctx_acquire(m, &c));
  int32_t rc = ORYX_OK;
  {
    std::lock_guard<std::mutex> gram_lock(m->gram_mu);
    do {
      if (m->d_gram == nullptr && cudaMalloc(&m->d_gram, (size_t)F * F * sizeof(double)) != cudaSuccess) {
        cudaGetLastError(); set_error("device allocation failed"); rc = ORYX_E_NOMEM; break;
      }
      std::shared_lock<std::shared_mutex> lk(m->mu);
      if (out_rows_used) *out_rows_used = m->store[which].live;
      // Large stores take the exact-sum tensor-core kernel first; its only error is the 24-bit fixed-point
      // quantisation of the inputs, whose size follows from the store's largest |x| and the result's
      // diagonal.  Should that estimate exceed kGramTcTolerance (heavy-tailed rows), the floating-point
      // kernel recomputes the Gram.
      constexpr double kGramTcTolerance = 4e-9;
      for (int attempt = 0; attempt < 2 && rc == ORYX_OK; attempt++) {
        float tc_absmax = 0.f;
        cudaEventRecord(c->ev0, c->stream);
        rc = gram_launch(m, which, m->d_gram, c->stream, attempt == 0, &tc_absmax);
        if (rc != ORYX_OK) break;
        cudaEventRecord(c->ev1, c->stream);
        if (cudaMemcpyAsync(out, m->d_gram, (size_t)F * F * sizeof(double), cudaMemcpyDeviceToHost, c->stream) != cudaSuccess ||
            cudaStreamSynchronize(c->stream) != cudaSuccess) {
          set_error("gram failed: %s", cudaGetErrorString(cudaGetLastError()));
          rc = ORYX_E_CUDA;
          break;
        }
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, c->ev0, c->ev1) == cudaSuccess) m->last_gram_ms.store((double)ms);
        else cudaGetLastError();
        m->last_gram_tc.store(tc_absmax > 0.f ? 1 : 0);
        if (tc_absmax == 0.f) break;  // a floating-point kernel ran: nothing to check
        double max_diag = 0.0;
        for (int i = 0; i < F; i++) max_diag = std::max(max_diag, out[(size_t)i * F + i]);
        if (gram_tc_error_estimate(tc_absmax, max_diag) <= kGramTcTolerance) break;
      }
    } while (0);
  }
  if (rc != ORYX_OK) cudaStreamSynchronize(c->stream);
  ctx_release(m, c);
  return rc;
}

// Development aid (not in the public header): device time (ms) of the model's last oryx_gram, memset included.
int32_t oryx_debug_gram_ms(oryx_model* mh, double* out_ms) {  // out_ms holds 2 doubles: ms, 1 if the tcgen05 kernel produced it
  out_ms[0] = reinterpret_cast<Model*>(mh)->last_gram_ms.load();
  out_ms[1] = (double)reinterpret_cast<Model*>(mh)->last_gram_tc.load();
  return ORYX_OK;
}

}  // extern "C"
