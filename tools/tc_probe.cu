// tc_probe.cu -- small hardware probes behind the design of the tensor-core Gram (gram_tc.cu):
//   1. tcgen05.mma kind::i8 (s8 x s8 -> s32) on hand-swizzled K-major operand tiles: exactness against a
//      CPU product (validates the instruction descriptor, the 128-byte swizzle formula used by the in-kernel
//      operand conversion, and the generic-proxy -> async-proxy fence).
//   2. how the fp32 accumulator of kind::f16 rounds (inside one K=16 instruction and between instructions).
//   3. issue rate of kind::i8 at N=128 and N=256 (is the 128 B/clk shared-memory port the limit at N=128?).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/tc_probe tools/tc_probe.cu
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n.reg .pred p;\nWAIT_LOOP:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra WAIT_DONE;\nbra WAIT_LOOP;\nWAIT_DONE:\n}\n"
      ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void tcgen05_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ uint64_t desc_k_sw128(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr >> 4) & 0x3FFF);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
template <int KIND>  // 0: i8, 1: f16
__device__ __forceinline__ void umma(uint32_t tmem_d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
  if (KIND == 0)
    asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n}\n"
                 ::"r"(tmem_d), "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
  else
    asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}\n"
                 ::"r"(tmem_d), "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
        "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
        "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// byte offset of (row r, byte b) in a K-major tile of 128-byte rows under the 128-byte swizzle
__host__ __device__ inline uint32_t sw128(uint32_t r, uint32_t b) {
  return (r >> 3) * 1024u + (r & 7u) * 128u + ((((b >> 4) ^ (r & 7u)) & 7u) << 4) + (b & 15u);
}

// One CTA of 128 threads.  a_g: [M=128][kbytes] row-major bytes, b_g: [N][kbytes] row-major bytes
// (kbytes = 128 per tile; `ktiles` tiles along K).  D[128][N] 32-bit out.  iters > 1 repeats the MMAs
// (accumulating) for the timing probe; cycles_out gets clock64 ticks from first issue to commit arrival.
template <int KIND>
__global__ void __launch_bounds__(128, 1) probe_kernel(const uint8_t* a_g, const uint8_t* b_g, int n, int ktiles, uint32_t idesc,
                                                        int kstep_bytes, int iters, uint32_t* d_out, long long* cycles_out) {
  extern __shared__ unsigned char smem_dyn[];
  unsigned char* base = reinterpret_cast<unsigned char*>(((uintptr_t)smem_dyn + 1023) & ~(uintptr_t)1023);
  unsigned char* a_s = base;                          // [ktiles][128 rows x 128 B]
  unsigned char* b_s = base + (size_t)ktiles * 16384;  // [ktiles][n rows x 128 B]
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_holder;
  const int tid = threadIdx.x, warp = tid >> 5;
  // operands -> shared memory in the canonical swizzled layout, written by ordinary stores
  for (int t = 0; t < ktiles; t++) {
    for (int e = tid; e < 128 * 128; e += 128) {
      const int r = e >> 7, b = e & 127;
      a_s[(size_t)t * 16384 + sw128(r, b)] = a_g[(size_t)r * ktiles * 128 + t * 128 + b];
    }
    for (int e = tid; e < n * 128; e += 128) {
      const int r = e >> 7, b = e & 127;
      b_s[(size_t)t * n * 128 + sw128(r, b)] = b_g[(size_t)r * ktiles * 128 + t * 128 + b];
    }
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy writes -> visible to the tensor core
  if (tid == 0) {
    mbar_init(&bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 256;" ::"r"(smem_u32(&tmem_holder)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = tmem_holder;
  if (tid == 0) {
    const long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
      for (int t = 0; t < ktiles; t++) {
        const uint32_t a_addr = smem_u32(a_s + (size_t)t * 16384), b_addr = smem_u32(b_s + (size_t)t * n * 128);
        for (int k = 0; k < 128 / kstep_bytes; k++)
          umma<KIND>(tmem, desc_k_sw128(a_addr + k * kstep_bytes), desc_k_sw128(b_addr + k * kstep_bytes), idesc,
                     (uint32_t)((it | t | k) != 0));
      }
    }
    tcgen05_commit(&bar);
    mbar_wait(&bar, 0);
    if (cycles_out) *cycles_out = clock64() - t0;
  }
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  for (int c0 = 0; c0 < n; c0 += 32) {
    uint32_t r[32];
    tmem_ld32(tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0, r);
    for (int i = 0; i < 32; i++) d_out[(size_t)tid * n + c0 + i] = r[i];
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 256;" ::"r"(tmem) : "memory");
}

// ---- no-swizzle K-major layout: 8-row x 16-byte core matrices (128 B each); per 32-byte K-step tile the two
// core matrices along K are `lbo` bytes apart, 8-row groups `sbo` bytes apart.
__host__ __device__ inline uint32_t nosw(uint32_t r, uint32_t b, uint32_t lbo, uint32_t sbo) {
  return (r >> 3) * sbo + (b >> 4) * lbo + (r & 7u) * 16u + (b & 15u);
}
__device__ __forceinline__ uint64_t desc_k_nosw(uint32_t smem_addr, uint32_t lbo, uint32_t sbo) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  return d;
}
// a_g: [128][ksteps*32] bytes, b_g: [n][ksteps*32]; every K-step is its own tile: A tile 128 rows (4 KB), B tile n rows.
__global__ void __launch_bounds__(128, 1) probe_nosw_kernel(const uint8_t* a_g, const uint8_t* b_g, int n, int ksteps, uint32_t idesc,
                                                             uint32_t lbo, uint32_t sbo, uint32_t* d_out) {
  extern __shared__ unsigned char smem_dyn[];
  unsigned char* base = reinterpret_cast<unsigned char*>(((uintptr_t)smem_dyn + 1023) & ~(uintptr_t)1023);
  const uint32_t a_tile = 128 / 8 * sbo, b_tile = (uint32_t)n / 8 * sbo;
  unsigned char* a_s = base;
  unsigned char* b_s = base + (size_t)ksteps * a_tile;
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_holder;
  const int tid = threadIdx.x, warp = tid >> 5;
  const int kb = ksteps * 32;
  for (int t = 0; t < ksteps; t++) {
    for (int e = tid; e < 128 * 32; e += 128) { const int r = e >> 5, b = e & 31; a_s[(size_t)t * a_tile + nosw(r, b, lbo, sbo)] = a_g[(size_t)r * kb + t * 32 + b]; }
    for (int e = tid; e < n * 32; e += 128) { const int r = e >> 5, b = e & 31; b_s[(size_t)t * b_tile + nosw(r, b, lbo, sbo)] = b_g[(size_t)r * kb + t * 32 + b]; }
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  if (tid == 0) { mbar_init(&bar, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 256;" ::"r"(smem_u32(&tmem_holder)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = tmem_holder;
  if (tid == 0) {
    for (int t = 0; t < ksteps; t++)
      umma<0>(tmem, desc_k_nosw(smem_u32(a_s + (size_t)t * a_tile), lbo, sbo), desc_k_nosw(smem_u32(b_s + (size_t)t * b_tile), lbo, sbo), idesc, (uint32_t)(t != 0));
    tcgen05_commit(&bar);
    mbar_wait(&bar, 0);
  }
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  for (int c0 = 0; c0 < n; c0 += 32) {
    uint32_t r[32];
    tmem_ld32(tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0, r);
    for (int i = 0; i < 32; i++) d_out[(size_t)tid * n + c0 + i] = r[i];
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 256;" ::"r"(tmem) : "memory");
}

static uint32_t idesc_i8(int m, int n, int a_signed, int b_signed) {
  return (2u << 4) | ((uint32_t)a_signed << 7) | ((uint32_t)b_signed << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}
static uint32_t idesc_bf16(int m, int n) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}

int main() {
  const int M = 128;
  uint8_t *d_a, *d_b;
  uint32_t* d_d;
  long long* d_cyc;
  CK(cudaMalloc(&d_a, 128 * 512));
  CK(cudaMalloc(&d_b, 256 * 512));
  CK(cudaMalloc(&d_d, 128 * 256 * 4));
  CK(cudaMalloc(&d_cyc, 8));
  const size_t smem = 1024 + 4 * 16384 + 4 * 256 * 128 + 64;
  CK(cudaFuncSetAttribute(probe_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  CK(cudaFuncSetAttribute(probe_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));

  // ---- 1. kind::i8 exactness
  for (int variant = 0; variant < 4; variant++) {
    const int n = (variant & 1) ? 256 : 128, ktiles = (variant & 2) ? 3 : 1, kb = ktiles * 128;
    const int b_signed = variant == 3 ? 0 : 1;
    std::vector<uint8_t> a((size_t)M * kb), b((size_t)n * kb);
    srand(17 + variant);
    for (auto& v : a) v = (uint8_t)(rand() & 255);
    for (auto& v : b) v = (uint8_t)(rand() & 255);
    CK(cudaMemcpy(d_a, a.data(), a.size(), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(d_b, b.data(), b.size(), cudaMemcpyHostToDevice));
    probe_kernel<0><<<1, 128, smem>>>(d_a, d_b, n, ktiles, idesc_i8(M, n, 1, b_signed), 32, 1, d_d, d_cyc);
    CK(cudaDeviceSynchronize());
    std::vector<int32_t> d((size_t)M * n);
    CK(cudaMemcpy(d.data(), d_d, d.size() * 4, cudaMemcpyDeviceToHost));
    long long bad = 0;
    for (int i = 0; i < M; i++)
      for (int j = 0; j < n; j++) {
        int32_t s = 0;
        for (int k = 0; k < kb; k++) {
          const int av = (int8_t)a[(size_t)i * kb + k];
          const int bv = b_signed ? (int)(int8_t)b[(size_t)j * kb + k] : (int)b[(size_t)j * kb + k];
          s += av * bv;
        }
        if (s != d[(size_t)i * n + j]) { if (bad < 3) printf("  mismatch (%d,%d): got %d want %d\n", i, j, d[(size_t)i * n + j], s); bad++; }
      }
    printf("i8 probe: M=128 N=%d K=%d A=s8 B=%s8 -> %s (%lld mismatches)\n", n, kb, b_signed ? "s" : "u", bad ? "WRONG" : "exact", bad);
  }

  // ---- 1b. kind::i8 on the no-swizzle K-major layout (independent 32-byte K-step tiles)
  CK(cudaFuncSetAttribute(probe_nosw_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  for (int variant = 0; variant < 4; variant++) {
    const int n = (variant & 1) ? 256 : 128, ksteps = (variant & 2) ? 5 : 1, kb = ksteps * 32;
    std::vector<uint8_t> a((size_t)M * kb), b((size_t)n * kb);
    srand(99 + variant);
    for (auto& v : a) v = (uint8_t)(rand() & 255);
    for (auto& v : b) v = (uint8_t)(rand() & 255);
    CK(cudaMemcpy(d_a, a.data(), a.size(), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(d_b, b.data(), b.size(), cudaMemcpyHostToDevice));
    CK(cudaMemset(d_d, 0, 128 * 256 * 4));
    probe_nosw_kernel<<<1, 128, smem>>>(d_a, d_b, n, ksteps, idesc_i8(M, n, 1, 1), 128, 256, d_d);
    CK(cudaDeviceSynchronize());
    std::vector<int32_t> d((size_t)M * n);
    CK(cudaMemcpy(d.data(), d_d, d.size() * 4, cudaMemcpyDeviceToHost));
    long long bad = 0;
    for (int i = 0; i < M; i++)
      for (int j = 0; j < n; j++) {
        int32_t sacc = 0;
        for (int k = 0; k < kb; k++) sacc += (int)(int8_t)a[(size_t)i * kb + k] * (int)(int8_t)b[(size_t)j * kb + k];
        if (sacc != d[(size_t)i * n + j]) { if (bad < 3) printf("  mismatch (%d,%d): got %d want %d\n", i, j, d[(size_t)i * n + j], sacc); bad++; }
      }
    printf("i8 probe, no-swizzle layout (LBO=128, SBO=256): M=128 N=%d K=%d -> %s (%lld mismatches)\n", n, kb, bad ? "WRONG" : "exact", bad);
  }

  // ---- 2. fp32 accumulator rounding of kind::f16 (bf16 operands)
  {
    const int n = 128, kb = 128;  // 64 bf16 per row = 4 K-steps of 16
    auto run = [&](const std::vector<float>& arow, const std::vector<float>& bcol, const char* what) {
      std::vector<__nv_bfloat16> a((size_t)M * 64, __float2bfloat16(0.f)), b((size_t)n * 64, __float2bfloat16(0.f));
      for (int k = 0; k < 64; k++) { a[k] = __float2bfloat16(arow[k]); b[k] = __float2bfloat16(bcol[k]); }
      CK(cudaMemcpy(d_a, a.data(), a.size() * 2, cudaMemcpyHostToDevice));
      CK(cudaMemcpy(d_b, b.data(), b.size() * 2, cudaMemcpyHostToDevice));
      probe_kernel<1><<<1, 128, smem>>>(d_a, d_b, n, 1, idesc_bf16(M, n), 32, 1, d_d, d_cyc);
      CK(cudaDeviceSynchronize());
      float got;
      CK(cudaMemcpy(&got, d_d, 4, cudaMemcpyDeviceToHost));
      double exact = 0;
      for (int k = 0; k < 64; k++) exact += (double)arow[k] * (double)bcol[k];
      printf("f16-kind accumulate probe [%s]: got %.1f, exact %.1f (delta %+.1f)\n", what, got, exact, got - exact);
    };
    std::vector<float> a(64, 0.f), b(64, 0.f);
    // (a) inside one K=16 instruction: 2^24 + 15 x 1  (fp32 ulp at 2^24 is 2)
    a[0] = 4096.f; b[0] = 4096.f;
    for (int k = 1; k < 16; k++) { a[k] = 1.f; b[k] = 1.f; }
    run(a, b, "one instruction: 2^24 + 15*1");
    // (b) across instructions: first instruction leaves 2^24, the next three add 3 each
    a.assign(64, 0.f); b.assign(64, 0.f);
    a[0] = 4096.f; b[0] = 4096.f;
    for (int s = 1; s < 4; s++) for (int k = 0; k < 3; k++) { a[16 * s + k] = 1.f; b[16 * s + k] = 1.f; }
    run(a, b, "2^24, then +3, +3, +3 in later instructions");
    // (c) negative side
    a.assign(64, 0.f); b.assign(64, 0.f);
    a[0] = -4096.f; b[0] = 4096.f;
    for (int s = 1; s < 4; s++) for (int k = 0; k < 3; k++) { a[16 * s + k] = -1.f; b[16 * s + k] = 1.f; }
    run(a, b, "-2^24, then -3, -3, -3");
    // (d) +1 per later instruction (ties / below half an ulp)
    a.assign(64, 0.f); b.assign(64, 0.f);
    a[0] = 4096.f; b[0] = 4096.f;
    for (int s = 1; s < 4; s++) { a[16 * s] = 1.f; b[16 * s] = 1.f; }
    run(a, b, "2^24, then +1, +1, +1");
    // (e) small terms first inside one instruction
    a.assign(64, 0.f); b.assign(64, 0.f);
    for (int k = 0; k < 15; k++) { a[k] = 1.f; b[k] = 1.f; }
    a[15] = 4096.f; b[15] = 4096.f;
    run(a, b, "one instruction: 15*1 + 2^24 (large term last)");
  }

  // ---- 3. issue rate of kind::i8: cycles per instruction at N=128 / N=256 (one CTA, operands resident)
  for (int n : {128, 256}) {
    const int iters = 256, ktiles = 4;
    CK(cudaMemset(d_a, 1, 128 * 512));
    CK(cudaMemset(d_b, 1, 256 * 512));
    probe_kernel<0><<<1, 128, smem>>>(d_a, d_b, n, ktiles, idesc_i8(M, n, 1, 1), 32, iters, d_d, d_cyc);
    CK(cudaDeviceSynchronize());
    long long cyc;
    CK(cudaMemcpy(&cyc, d_cyc, 8, cudaMemcpyDeviceToHost));
    const double per = (double)cyc / (iters * ktiles * 4);
    printf("i8 issue rate: M=128 N=%d K=32: %.1f cycles per instruction (%.0f MAC/clk/SM)\n", n, per, 128.0 * n * 32 / per);
  }
  for (int n : {128, 256}) {
    const int iters = 256, ktiles = 4;
    probe_kernel<1><<<1, 128, smem>>>(d_a, d_b, n, ktiles, idesc_bf16(M, n), 32, iters, d_d, d_cyc);
    CK(cudaDeviceSynchronize());
    long long cyc;
    CK(cudaMemcpy(&cyc, d_cyc, 8, cudaMemcpyDeviceToHost));
    const double per = (double)cyc / (iters * ktiles * 4);
    printf("bf16 issue rate: M=128 N=%d K=16: %.1f cycles per instruction (%.0f MAC/clk/SM)\n", n, per, 128.0 * n * 16 / per);
  }
  return 0;
}
