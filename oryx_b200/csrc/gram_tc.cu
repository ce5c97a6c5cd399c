// gram_tc.cu -- G = V^T V on the 5th-generation tensor cores, exact-sum fixed-point version (kernel K4 of
// SURVEY.md section 2.3 for large stores).
//
// Replaces VectorMath.transposeTimesSelf (reference framework/oryx-common/.../math/VectorMath.java:64-83)
// as summed over partitions by ALSServingModel.getYTYSolver (ALSServingModel.java:357-373) and
// ALSSpeedModel.getXTXSolver/getYTYSolver (ALSSpeedModel.java:118-126).
//
// Why not a plain bf16 / tf32 tensor-core SYRK: the fp32 accumulator of tcgen05.mma kind::f16 TRUNCATES
// (measured on B200, tools/tc_probe.cu -> profiles/r02_tc_probe.txt: 2^24 + 15 x 1 gives 2^24 + 14, three
// later "+3" steps each add 2).  A Gram's diagonal is a sum of squares, so every truncation pulls it the
// same way and the error grows with the number of rows instead of averaging out.  Integer accumulation
// has no rounding at all, so the contraction is done in fixed point:
//
//   x  ~=  g * (d0 * 2^16 + s1 * 2^8 + s2)        d0, s1, s2 signed 8-bit digits, g = absmax / (2^23 - 2^16)
//
// (24 significant bits at the scale of the store's largest |x|; balanced digits, so what is left out below
// is zero-mean.)  x_r * x_c then splits into digit products that tcgen05.mma kind::i8 sums EXACTLY into
// int32 TMEM accumulators, one accumulator per weight: 2^32 (d0 d0), 2^24 (d0 s1), 2^16 (d0 s2, s1 s1),
// 2^8 (s1 s2); only s2 s2 (2^-30 of the leading term) is dropped.  Every <= 32768 rows the four int32 tiles
// are combined in int64 and added to binary64 partials.  The only approximation is the 24-bit quantisation
// of the inputs -- the same order as the binary32 rounding of every product that the reference itself
// performs -- and it is unbiased; DESIGN.md section 4.4 has the error model and oryx_gram checks the bound.
//
// Work split (features > 128): three kinds of persistent CTA, one per upper-triangular 128 x 128 block of
// G -- (0,0), (1,1), (0,1) -- because the four int32 accumulators of one block are exactly the 512 TMEM
// columns of an SM.  Each CTA pulls 2048-row units from its role's counter; the roles keep within a
// window of each other so a row range is fetched from HBM once and found in L2 by the other two.
// Inside a CTA: one warp drives TMA (fp32 tiles of 32 rows x 128 columns into a shared-memory ring), eight
// warps turn each tile into the three digit planes, written in the K-major 128-byte-swizzled layout the
// tensor core reads (rows of the store = K), one thread issues the MMAs, and the converter warps drain
// TMEM.  Diagonal blocks use symmetry: d0 s1 and s1 s2 are accumulated one way round only and mirrored
// when the partials are reduced (gram_tc_finish_kernel), which also makes the result exactly symmetric.
#include <cuda.h>

#include <algorithm>
#include <cmath>
#include <cstdlib>

#include "common.cuh"

namespace oryx {

namespace {

constexpr int kKStep = 32;        // rows per tcgen05.mma kind::i8 instruction (K = 32 bytes)
constexpr int kBlk = 128;         // features per block: UMMA M, and the N of one block
constexpr int kUnitRows = 2048;   // rows per scheduled work unit
constexpr int kDrainRows = 32768; // rows accumulated in int32 between drains (<= 2^31 / (3 * 2^14) = 43690)
constexpr int kConvGroups = 4;     // converter groups: group k converts tiles k, k + 4, ... into operand slot k
constexpr int kGroupWarps = 4;
constexpr int kConvWarps = kConvGroups * kGroupWarps;
constexpr int kThreads = 64 + kConvWarps * 32;  // warp 0: TMA, warp 1: MMA + TMEM, warps 2..17: convert / drain
constexpr int kHalfStageBytes = kKStep * kBlk * 4;  // 16 KB: 32 rows x 128 fp32
constexpr int kTileBytes = kBlk * 128;              // 16 KB: 128 feature rows x 128 K-bytes (4 K-steps)
constexpr int kOpRing = 4;                          // K-step slots in an operand tile
constexpr int kMaxStages = 8;
constexpr uint32_t kOperandBytes = 6 * kTileBytes;  // 96 KB (diagonal roles use the first 48 KB)
constexpr uint32_t kStagingBytes = 8 * kHalfStageBytes;  // 128 KB: 4 stages x 2 halves, or 8 stages x 1 half
constexpr size_t kSmemBytes = 1024 + kOperandBytes + kStagingBytes + 1024;
constexpr float kScaleTop = 8388608.0f - 65536.0f;  // |x| * scale <= 2^23 - 2^16
constexpr int kDigitOffset = 32768 + 128;           // makes the two low bytes balanced digits after ^ 0x80

// kFlagDrain: after this tile's MMAs the accumulators are drained (issuer and converters); kFlagJoin: the
// converters of this tile take part in a drain flagged on one of the three tiles before it.
// kFlagIdle (on end markers): nothing has been accumulated since the last drain, so there is no final drain.
constexpr int kFlagDrain = 1, kFlagEnd = 2, kFlagJoin = 4, kFlagIdle = 8;
enum { kRoleD0 = 0, kRoleD1 = 1, kRoleO = 2 };

struct GramTcParams {
  const int32_t* slot_row;
  int64_t n_slots;
  int n_roles;        // 1 (one 128-column block) or 3
  int grid;
  // Role of every CTA.  Roles are interleaved over the CTA indices so that any prefix of the grid holds all
  // three in proportion: should fewer CTAs than launched be resident for a while (SMs busy with queries),
  // every role still advances and nobody waits at the window for a role that has not started.
  uint8_t role_of[256];
  unsigned int n_units;
  unsigned int window_units;
  int drain_rows;          // rows accumulated in int32 between drains (kDrainRows; tests lower it)
  unsigned int* counters;  // [4], zeroed before the launch
  const unsigned int* absmax_bits;
  double* partials;        // [grid][2][128 x 128], element (i, j) at [j * 128 + i]
  int32_t* contributed;    // [grid]
};

struct StageMeta {
  int64_t row0;
  uint32_t alive;  // bit r: row row0 + r is a live slot
  int32_t flags;
};

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
// D[tmem] (+)= A[smem] * B[smem]^T, signed 8-bit inputs, int32 accumulate, issued by one thread.
__device__ __forceinline__ void umma_i8(uint32_t tmem_d, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n"
      "}\n" ::"r"(tmem_d), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate) : "memory");
}
// K-major operand tile, 128-byte swizzle, 8-row atoms 1024 B apart (same layout as batch_topn.cu).
__device__ __forceinline__ uint64_t umma_desc_k_sw128(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr >> 4) & 0x3FFF);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
// kind::i8 instruction descriptor: D = s32, A = B = s8, both K-major, M = 128.
constexpr uint32_t idesc_i8(int n) {
  return (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(kBlk >> 4) << 24);
}
__device__ __forceinline__ void tmem_ld_32x32b_x16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// x * scale rounded to an integer, offset so that the two low bytes become balanced digits:
// result = d0 * 65536 + (s1 + 128) * 256 + (s2 + 128) with d0 the (signed) third byte.
__device__ __forceinline__ uint32_t quantise(float x, float scale) {
  return (uint32_t)__float2int_rn(fmaf(x, scale, (float)kDigitOffset));
}

__device__ __forceinline__ float gram_scale(const unsigned int* absmax_bits) {
  const float absmax = __uint_as_float(*absmax_bits);
  return absmax > 0.f ? __fdiv_rd(kScaleTop, absmax) : 0.f;
}

__global__ void __launch_bounds__(kThreads, 1)  // 18 warps = 5 on one scheduler: 96 registers per thread at most
gram_tc_kernel(const __grid_constant__ CUtensorMap tm, const __grid_constant__ GramTcParams p) {
  extern __shared__ unsigned char smem_dyn[];
  unsigned char* base = reinterpret_cast<unsigned char*>(((uintptr_t)smem_dyn + 1023) & ~(uintptr_t)1023);
  unsigned char* op = base;                          // operand tiles: [A d0 | A s1 | A s2 | B d0 | B s1 | B s2]
  unsigned char* staging = base + kOperandBytes;     // fp32 ring
  uint64_t* bars = reinterpret_cast<uint64_t*>(staging + kStagingBytes);
  uint64_t* stage_full = bars;                       // [kMaxStages]
  uint64_t* stage_empty = bars + kMaxStages;         // [kMaxStages]
  uint64_t* op_full = bars + 2 * kMaxStages;         // [kOpRing]
  uint64_t* op_empty = op_full + kOpRing;            // [kOpRing]
  uint64_t* acc_full = op_empty + kOpRing;
  uint64_t* acc_empty = acc_full + 1;
  StageMeta* meta = reinterpret_cast<StageMeta*>(acc_empty + 1);  // [kMaxStages]
  int32_t* op_meta = reinterpret_cast<int32_t*>(meta + kMaxStages);  // [kOpRing]
  uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(op_meta + kOpRing);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int bid = blockIdx.x;
  const int role = p.role_of[bid];
  const int halves = role == kRoleO ? 2 : 1;
  const int n_stages = role == kRoleO ? 4 : 8;
  const int col_a = role == kRoleD1 ? kBlk : 0;  // first column of the A-side block (the B side of (0,1) is 128)

  if (threadIdx.x == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tm) : "memory");
    for (int i = 0; i < kMaxStages; i++) { mbar_init(&stage_full[i], 1); mbar_init(&stage_empty[i], kGroupWarps); }
    for (int i = 0; i < kOpRing; i++) { mbar_init(&op_full[i], kGroupWarps); mbar_init(&op_empty[i], 1); }
    mbar_init(acc_full, 1);
    mbar_init(acc_empty, kConvWarps);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {  // all 512 TMEM columns: four 128 x 128 int32 accumulators
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(tmem_holder)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_holder;

  if (warp == 0) {
    // ===== producer: work units, liveness masks, TMA =====
    int stage = 0;
    uint32_t sphase = 0;
    int rows_since_drain = 0, join_left = 0;
    for (;;) {
      unsigned int u = 0;
      if (lane == 0) u = atomicAdd(&p.counters[role], 1u);
      u = __shfl_sync(0xffffffffu, u, 0);
      if (u >= p.n_units) break;
      if (p.n_roles > 1 && lane == 0) {
        // stay within a window of the slowest other role: the rows this CTA is about to fetch are then
        // either still in L2 (fetched by a role ahead of us) or will be found there by the roles behind
        for (;;) {
          unsigned int mn = 0xffffffffu;
          for (int r = 0; r < 3; r++)
            if (r != role) mn = min(mn, *(volatile unsigned int*)&p.counters[r]);
          if (mn + p.window_units >= u) break;
          __nanosleep(500);
        }
      }
      __syncwarp();
      const int64_t unit_row0 = (int64_t)u * kUnitRows;
      auto load_flags = [&](int g, int32_t (&f)[4]) {
#pragma unroll
        for (int t = 0; t < 4; t++) {
          const int64_t r = unit_row0 + (int64_t)g * 128 + t * 32 + lane;
          f[t] = r < p.n_slots ? __ldg(p.slot_row + r) : -1;
        }
      };
      int32_t cur[4], nxt[4];
      load_flags(0, cur);
      for (int g = 0; g < kUnitRows / 128; g++) {
        if (unit_row0 + (int64_t)g * 128 >= p.n_slots) break;
        if (g + 1 < kUnitRows / 128) load_flags(g + 1, nxt);
#pragma unroll
        for (int t = 0; t < 4; t++) {
          const uint32_t alive = __ballot_sync(0xffffffffu, cur[t] >= 0);
          const int64_t row0 = unit_row0 + (int64_t)g * 128 + t * 32;
          if (row0 >= p.n_slots) continue;
          if (lane == 0) {
            mbar_wait(&stage_empty[stage], sphase ^ 1u);
            rows_since_drain += kKStep;
            int flags = 0;
            if (join_left > 0) { flags = kFlagJoin; join_left--; }
            if (rows_since_drain >= p.drain_rows) { flags |= kFlagDrain; rows_since_drain = 0; join_left = kConvGroups - 1; }
            meta[stage].row0 = row0;
            meta[stage].alive = alive;
            meta[stage].flags = flags;
            mbar_expect_tx(&stage_full[stage], (uint32_t)halves * kHalfStageBytes);
            unsigned char* dst = staging + (size_t)stage * halves * kHalfStageBytes;
            tma_load_2d(dst, &tm, col_a, (int)row0, &stage_full[stage]);
            if (halves == 2) tma_load_2d(dst + kHalfStageBytes, &tm, kBlk, (int)row0, &stage_full[stage]);
          }
          if (++stage == n_stages) { stage = 0; sphase ^= 1u; }
        }
#pragma unroll
        for (int t = 0; t < 4; t++) cur[t] = nxt[t];
      }
    }
    if (lane == 0) {  // end markers, one per converter group: no data behind them
      for (int k = 0; k < kConvGroups; k++) {
        mbar_wait(&stage_empty[stage], sphase ^ 1u);
        meta[stage].row0 = 0;
        meta[stage].alive = 0;
        meta[stage].flags = kFlagEnd | (join_left > 0 ? kFlagJoin : 0) | (rows_since_drain == 0 ? kFlagIdle : 0);
        if (join_left > 0) join_left--;
        mbar_arrive(&stage_full[stage]);
        if (++stage == n_stages) { stage = 0; sphase ^= 1u; }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer =====
    if (lane == 0) {
      constexpr uint32_t kI256 = idesc_i8(256), kI128 = idesc_i8(128);
      const uint32_t c0 = tmem_base, c1 = tmem_base + 128, c2 = tmem_base + 256, c3 = tmem_base + 384;
      const uint32_t a_d0 = smem_u32(op), a_s1 = a_d0 + kTileBytes, a_s2 = a_d0 + 2 * kTileBytes;
      const uint32_t b_d0 = role == kRoleO ? a_d0 + 3 * kTileBytes : a_d0, b_s1 = b_d0 + kTileBytes, b_s2 = b_d0 + 2 * kTileBytes;
      int slot = 0;
      uint32_t ophase = 0, dphase = 0;
      bool fresh = true;  // accumulators hold nothing yet (start, or just drained)
      for (;;) {
        mbar_wait(&op_full[slot], ophase);
        tcgen05_fence_after();
        const int fl = *(volatile int32_t*)&op_meta[slot];
        if (fl & kFlagEnd) {
          tcgen05_commit(acc_full);  // everything issued so far has landed in TMEM when this arrives
          break;
        }
        const uint32_t k = (uint32_t)slot * kKStep;  // byte offset of this K-step inside the 128-byte rows
        const uint32_t acc = fresh ? 0u : 1u;
        if (role == kRoleO) {
          // block (0,1): every ordered digit pair (i, j) with i + j <= 3
          umma_i8(c0, umma_desc_k_sw128(a_d0 + k), umma_desc_k_sw128(b_d0 + k), kI256, acc);  // (0,0)->c0 (0,1)->c1
          umma_i8(c2, umma_desc_k_sw128(a_s2 + k), umma_desc_k_sw128(b_d0 + k), kI256, acc);  // (2,0)->c2 (2,1)->c3
          umma_i8(c1, umma_desc_k_sw128(a_s1 + k), umma_desc_k_sw128(b_d0 + k), kI256, 1u);   // (1,0)->c1 (1,1)->c2
          umma_i8(c2, umma_desc_k_sw128(a_d0 + k), umma_desc_k_sw128(b_s2 + k), kI128, 1u);   // (0,2)->c2
          umma_i8(c3, umma_desc_k_sw128(a_s1 + k), umma_desc_k_sw128(b_s2 + k), kI128, 1u);   // (1,2)->c3
        } else {
          // diagonal block: (0,1) and (1,2) one way round only (mirrored by the finishing kernel)
          umma_i8(c0, umma_desc_k_sw128(a_d0 + k), umma_desc_k_sw128(b_d0 + k), kI256, acc);  // (0,0)->c0 (0,1)->c1
          umma_i8(c2, umma_desc_k_sw128(a_s1 + k), umma_desc_k_sw128(b_s1 + k), kI256, acc);  // (1,1)->c2 (1,2)->c3
          umma_i8(c2, umma_desc_k_sw128(a_d0 + k), umma_desc_k_sw128(b_s2 + k), kI128, 1u);   // (0,2)->c2
          umma_i8(c2, umma_desc_k_sw128(a_s2 + k), umma_desc_k_sw128(b_d0 + k), kI128, 1u);   // (2,0)->c2
        }
        tcgen05_commit(&op_empty[slot]);  // this K-step slot of the operand tiles may be rewritten
        fresh = false;
        if (fl & kFlagDrain) {
          tcgen05_commit(acc_full);
          mbar_wait(acc_empty, dphase);
          dphase ^= 1u;
          tcgen05_fence_after();
          fresh = true;
        }
        if (++slot == kOpRing) { slot = 0; ophase ^= 1u; }
      }
    }
  } else {
    // ===== converters: fp32 tile -> digit planes in the tensor core's operand layout; TMEM drains =====
    // Four groups of four warps; group k converts tiles k, k + 4, ... into K-step slot k of the operand tiles,
    // so four tiles are in flight and the wait -> load -> convert -> store -> fence -> arrive latency of one
    // tile overlaps the others'.  A thread owns one column of the 32-row x 128-column tile: its 32 values
    // become 8 words per digit plane (4 rows = one 32-bit word of K bytes), i.e. two 16-byte chunks of that
    // feature's swizzled operand row.
    const int cw = warp - 2;                  // 0..15
    const int group = cw >> 2, wig = cw & 3;
    const int col = 32 * wig + lane;          // column of the 128-column half this thread converts
    const uint32_t r8 = (uint32_t)col & 7u;
    const uint32_t row_off = ((uint32_t)col >> 3) * 1024u + r8 * 128u;  // start of this feature's 128-byte row
    const float scale = gram_scale(p.absmax_bits);
    double* part = p.partials + (size_t)bid * 2 * kBlk * kBlk;
    const int slot = group;
    uint32_t aphase = 0;
    bool first_drain = true;

    auto drain = [&]() {
      mbar_wait(acc_full, aphase);
      aphase ^= 1u;
      tcgen05_fence_after();
      const int lq = warp & 3;                 // TMEM lane quarter this warp may read
      const int i = 32 * lq + lane;            // row of the block (A-side feature)
      const int j_base = 32 * (cw >> 2);       // four warps share a quarter: 32 columns each
      const uint32_t t0 = tmem_base + ((uint32_t)(32 * lq) << 16);
#pragma unroll 1
      for (int jc = 0; jc < 2; jc++) {
        const int j0 = j_base + 16 * jc;
        uint32_t a0[16], a1[16], a2[16], a3[16];
        tmem_ld_32x32b_x16(t0 + 0 + (uint32_t)j0, a0);
        tmem_ld_32x32b_x16(t0 + 128 + (uint32_t)j0, a1);
        tmem_ld_32x32b_x16(t0 + 256 + (uint32_t)j0, a2);
        tmem_ld_32x32b_x16(t0 + 384 + (uint32_t)j0, a3);
        tmem_ld_wait();
#pragma unroll
        for (int jj = 0; jj < 16; jj++) {
          const long long s = (long long)(int)a0[jj] * 4294967296LL + (long long)(int)a2[jj] * 65536LL;
          const long long uu = (long long)(int)a1[jj] * 16777216LL + (long long)(int)a3[jj] * 256LL;
          double* d = part + (size_t)(j0 + jj) * kBlk + i;
          if (role == kRoleO) {
            const double v = (double)(s + uu);
            d[0] = first_drain ? v : d[0] + v;
          } else {
            const double vs = (double)s, vu = (double)uu;
            d[0] = first_drain ? vs : d[0] + vs;
            d[kBlk * kBlk] = first_drain ? vu : d[kBlk * kBlk] + vu;
          }
        }
      }
      first_drain = false;
      tcgen05_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(acc_empty);
    };

    for (int i = group;; i += kConvGroups) {
      const int stage = i % n_stages;
      mbar_wait(&stage_full[stage], (uint32_t)(i / n_stages) & 1u);
      const int flags = *(volatile int32_t*)&meta[stage].flags;
      const uint32_t alive = *(volatile uint32_t*)&meta[stage].alive;
      mbar_wait(&op_empty[slot], ((uint32_t)(i / kOpRing) & 1u) ^ 1u);
      if (flags & kFlagEnd) {
        if (wig == 0 && lane == 0) *(volatile int32_t*)&op_meta[slot] = flags;
        __syncwarp();
        if (lane == 0) mbar_arrive(&op_full[slot]);
        if (flags & kFlagJoin) drain();
        // the issuer answers its first end marker with a last commit on acc_full; when nothing was issued
        // since the last drain the accumulators still hold what that drain already took: leave them alone
        if (!(flags & kFlagIdle)) drain();
        break;
      }
      // 16-byte chunks of this K-step inside the swizzled 128-byte row: rows 0..15 and rows 16..31
      const uint32_t chunk0 = ((((uint32_t)(2 * slot)) ^ r8) & 7u) << 4, chunk1 = ((((uint32_t)(2 * slot + 1)) ^ r8) & 7u) << 4;
      for (int h = 0; h < halves; h++) {
        unsigned char* t_d0 = op + (size_t)(h * 3) * kTileBytes + row_off;
        const float* src = reinterpret_cast<const float*>(staging + ((size_t)stage * halves + h) * kHalfStageBytes) + col;
#pragma unroll
        for (int hh = 0; hh < 2; hh++) {  // rows 0..15, rows 16..31 of this thread's column
          float x[16];
#pragma unroll
          for (int r = 0; r < 16; r++) x[r] = src[(16 * hh + r) * kBlk];  // lanes = consecutive columns: one wavefront
          const uint32_t alive16 = (alive >> (16 * hh)) & 0xffffu;
          if (alive16 != 0xffffu) {  // tombstones are NaN rows: they contribute nothing
#pragma unroll
            for (int r = 0; r < 16; r++) x[r] = ((alive16 >> r) & 1u) ? x[r] : 0.f;
          }
          uint32_t wd[4], w1[4], w2[4];
#pragma unroll
          for (int g4 = 0; g4 < 4; g4++) {
            const uint32_t v0 = quantise(x[4 * g4 + 0], scale), v1 = quantise(x[4 * g4 + 1], scale);
            const uint32_t v2 = quantise(x[4 * g4 + 2], scale), v3 = quantise(x[4 * g4 + 3], scale);
            // bytes of four consecutive rows, per digit: byte 2 -> d0, byte 1 -> s1 + 128, byte 0 -> s2 + 128
            const uint32_t lo01 = __byte_perm(v0, v1, 0x5140), lo23 = __byte_perm(v2, v3, 0x5140);
            const uint32_t hi01 = __byte_perm(v0, v1, 0x0062), hi23 = __byte_perm(v2, v3, 0x0062);
            wd[g4] = __byte_perm(hi01, hi23, 0x5410);
            w1[g4] = __byte_perm(lo01, lo23, 0x7632) ^ 0x80808080u;
            w2[g4] = __byte_perm(lo01, lo23, 0x5410) ^ 0x80808080u;
          }
          // 128-bit stores: a quarter warp (8 consecutive features = one swizzle atom) covers all 32 banks
          const uint32_t chunk = hh == 0 ? chunk0 : chunk1;
          *reinterpret_cast<uint4*>(t_d0 + chunk) = make_uint4(wd[0], wd[1], wd[2], wd[3]);
          *reinterpret_cast<uint4*>(t_d0 + kTileBytes + chunk) = make_uint4(w1[0], w1[1], w1[2], w1[3]);
          *reinterpret_cast<uint4*>(t_d0 + 2 * kTileBytes + chunk) = make_uint4(w2[0], w2[1], w2[2], w2[3]);
        }
      }
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // these stores are read by the tensor core
      if (wig == 0 && lane == 0) *(volatile int32_t*)&op_meta[slot] = flags;
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(&op_full[slot]);
        mbar_arrive(&stage_empty[stage]);
      }
      if (flags & kFlagJoin) drain();
      if (flags & kFlagDrain) drain();
    }
    if (cw == 0 && lane == 0) p.contributed[bid] = first_drain ? 0 : 1;
  }

  tcgen05_fence_before();
  __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem_base) : "memory");
  }
}

// Sums the CTAs' binary64 partials per block, mirrors what the diagonal blocks accumulated one way round
// only, applies 1 / scale^2 and writes both triangles of the F x F result from the same value.
__global__ void __launch_bounds__(256) gram_tc_finish_kernel(const __grid_constant__ GramTcParams p, int features,
                                                            double* __restrict__ out) {
  const int role = blockIdx.y;
  const int e = blockIdx.x * 256 + threadIdx.x;  // = j * 128 + i
  const int j = e >> 7, i = e & 127;
  double s = 0.0, u = 0.0, ut = 0.0;
  for (int cta = 0; cta < p.grid; cta++) {
    if (p.role_of[cta] != role || !p.contributed[cta]) continue;
    const double* part = p.partials + (size_t)cta * 2 * kBlk * kBlk;
    s += part[e];
    if (role != kRoleO) {
      u += part[kBlk * kBlk + e];
      ut += part[kBlk * kBlk + i * kBlk + j];
    }
  }
  const double scale = (double)gram_scale(p.absmax_bits);
  const double inv = scale > 0.0 ? 1.0 / (scale * scale) : 0.0;
  const double v = (s + (u + ut)) * inv;
  const int r = (role == kRoleD1 ? kBlk : 0) + i;
  const int c = (role == kRoleD0 ? 0 : kBlk) + j;
  if (r < features && c < features) {
    out[(size_t)r * features + c] = v;
    if (role == kRoleO) out[(size_t)c * features + r] = v;
  }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn gram_encode_fn() {
  static EncodeTiledFn fn = [] {
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) != cudaSuccess ||
        qres != cudaDriverEntryPointSuccess) {
      cudaGetLastError();
      ptr = nullptr;
    }
    return reinterpret_cast<EncodeTiledFn>(ptr);
  }();
  return fn;
}

}  // namespace

constexpr int64_t kGramTcMinRows = 1 << 20;

// True when the tensor-core Gram applies: 128..256 padded features (one or three 128-column blocks), a
// store large enough for the 24-bit input quantisation to vanish in the sum, TMA available.
bool gram_tc_applicable(const Store* s) {
  static const bool disabled = [] { const char* e = std::getenv("ORYX_GRAM_TC"); return e && e[0] == '0'; }();
  return !disabled && s->fpad >= 128 && s->fpad <= 256 && s->n_slots >= kGramTcMinRows && gram_encode_fn() != nullptr;
}

// The store's largest |x| (host copy, synchronises `stream`).
int32_t gram_tc_absmax(const Store* s, cudaStream_t stream, float* out) {
  unsigned int bits = 0;
  ORYX_CUDA(cudaMemcpyAsync(&bits, s->d_absmax_bits, sizeof(bits), cudaMemcpyDeviceToHost, stream));
  ORYX_CUDA(cudaStreamSynchronize(stream));
  std::memcpy(out, &bits, sizeof(bits));
  return ORYX_OK;
}

// Expected largest error of a tensor-core Gram entry relative to the largest entry, from the 24-bit input
// quantisation (uniform in +-g/2 per element, independent): rms 0.29 g sqrt(G_rr + G_cc), the largest of
// F^2 entries about 4.5 rms.  `max_diag` is the largest diagonal entry of the result.
double gram_tc_error_estimate(float absmax, double max_diag) {
  if (!(max_diag > 0.0)) return 0.0;
  const double g = (double)absmax / (double)kScaleTop;
  return 4.5 * 0.29 * g * std::sqrt(2.0 * max_diag) / max_diag;
}

int32_t gram_tc_launch(Model* m, Store* s, double* d_out, cudaStream_t stream) {
  const int F = s->features;
  const int grid = m->num_sms;
  const int n_roles = s->fpad > kBlk ? 3 : 1;
  GramTcParams p{};
  p.slot_row = s->d_slot_row;
  p.n_slots = s->n_slots;
  p.n_roles = n_roles;
  p.grid = grid;
  if (grid > (int)sizeof(p.role_of)) {
    set_error("device has %d SMs; the tensor-core Gram's role table holds %d", grid, (int)sizeof(p.role_of));
    return ORYX_E_UNSUPPORTED;
  }
  if (n_roles == 1) {
    std::memset(p.role_of, kRoleD0, sizeof(p.role_of));
  } else {
    // CTAs per diagonal role; the (0,1) block gets the rest (it converts twice the columns and issues five
    // instructions per K-step to the diagonal roles' four).  ORYX_GRAM_TC_DIAG_CTAS overrides for tuning.
    static const int env_diag = [] { const char* e = std::getenv("ORYX_GRAM_TC_DIAG_CTAS"); return e ? std::atoi(e) : 0; }();
    int n_diag = env_diag > 0 ? env_diag : (int)std::lround(grid * 0.27);
    n_diag = std::max(1, std::min(n_diag, (grid - 1) / 2));
    const int want[3] = {n_diag, n_diag, grid - 2 * n_diag};
    int have[3] = {0, 0, 0};
    for (int b = 0; b < grid; b++) {  // the role furthest behind its share of the first b + 1 CTAs
      int best = 0;
      double best_deficit = -1e30;
      for (int r = 0; r < 3; r++) {
        const double deficit = (double)want[r] * (b + 1) / grid - have[r];
        if (have[r] < want[r] && deficit > best_deficit) { best_deficit = deficit; best = r; }
      }
      p.role_of[b] = (uint8_t)best;
      have[best]++;
    }
  }
  p.n_units = (unsigned int)((s->n_slots + kUnitRows - 1) / kUnitRows);
  p.window_units = 24;  // 24 x 2048 rows x 1 KB = 48 MB of the 126 MB L2
  p.drain_rows = kDrainRows;
  if (const char* e = std::getenv("ORYX_GRAM_TC_DRAIN_ROWS")) {  // test hook: exercise the drain protocol on small stores
    const int v = std::atoi(e);
    if (v >= kKStep && v <= kDrainRows) p.drain_rows = v / kKStep * kKStep;
  }
  const size_t head = 256 + (size_t)grid * sizeof(int32_t);
  const size_t head_pad = (head + 255) & ~(size_t)255;
  const size_t need = head_pad + (size_t)grid * 2 * kBlk * kBlk * sizeof(double);
  if (m->gram_scratch_bytes < need) {
    cudaFree(m->d_gram_scratch);
    m->d_gram_scratch = nullptr;
    m->gram_scratch_bytes = 0;
    if (cudaMalloc(&m->d_gram_scratch, need) != cudaSuccess) {
      cudaGetLastError();
      set_error("device allocation of %zu bytes failed (Gram partials)", need);
      return ORYX_E_NOMEM;
    }
    m->gram_scratch_bytes = need;
  }
  unsigned char* sc = reinterpret_cast<unsigned char*>(m->d_gram_scratch);
  p.counters = reinterpret_cast<unsigned int*>(sc);
  p.contributed = reinterpret_cast<int32_t*>(sc + 256);
  p.partials = reinterpret_cast<double*>(sc + head_pad);
  p.absmax_bits = s->d_absmax_bits;
  ORYX_CUDA(cudaMemsetAsync(sc, 0, 256, stream));

  CUtensorMap tm;
  {
    cuuint64_t dims[2] = {(cuuint64_t)s->fpad, (cuuint64_t)s->n_slots};
    cuuint64_t strides[1] = {(cuuint64_t)s->fpad * sizeof(float)};
    cuuint32_t box[2] = {(cuuint32_t)kBlk, (cuuint32_t)kKStep};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = gram_encode_fn()(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, s->d_vec, dims, strides, box, estr,
                                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
      set_error("cuTensorMapEncodeTiled failed with CUresult %d (Gram)", (int)r);
      return ORYX_E_CUDA;
    }
  }
  ORYX_TRY(ensure_max_dynamic_smem(reinterpret_cast<const void*>(gram_tc_kernel), kSmemBytes));
  gram_tc_kernel<<<grid, kThreads, kSmemBytes, stream>>>(tm, p);
  ORYX_CUDA(cudaGetLastError());
  dim3 fgrid(kBlk * kBlk / 256, n_roles);
  gram_tc_finish_kernel<<<fgrid, 256, 0, stream>>>(p, F, d_out);
  ORYX_CUDA(cudaGetLastError());
  m->launches += 2;
  return ORYX_OK;
}

}  // namespace oryx
