// bw_probe.cu -- read-bandwidth ceiling probe for the scan kernel's design space (development tool,
// not part of the product).  Variants: (a) 128-bit streaming loads with U loads in flight per thread
// at several occupancies, (b) cp.async.bulk global->shared ring (TMA engine) with S stages per CTA.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <vector>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA %s at %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)

__device__ __forceinline__ float4 ldg_stream(const float4* p) {
  float4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
  return r;
}

template <int U>
__global__ void ldg_kernel(const float4* __restrict__ p, size_t n4, float* sink) {
  size_t stride = (size_t)gridDim.x * blockDim.x;
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  float acc = 0.f;
  for (; i + (U - 1) * stride < n4; i += U * stride) {
    float4 v[U];
#pragma unroll
    for (int u = 0; u < U; u++) v[u] = ldg_stream(p + i + u * stride);
#pragma unroll
    for (int u = 0; u < U; u++) acc += v[u].x + v[u].y + v[u].z + v[u].w;
  }
  if (acc == 123.456f) *sink = acc;
}

__device__ __forceinline__ void mbar_init(uint64_t* b, int c) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"((uint32_t)__cvta_generic_to_shared(b)), "r"(c)); }
__device__ __forceinline__ void mbar_expect(uint64_t* b, uint32_t bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"((uint32_t)__cvta_generic_to_shared(b)), "r"(bytes)); }
__device__ __forceinline__ void mbar_arrive(uint64_t* b) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"((uint32_t)__cvta_generic_to_shared(b))); }
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t parity) {
  asm volatile("{\n.reg .pred p;\nW: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra D;\nbra W;\nD:\n}" :: "r"((uint32_t)__cvta_generic_to_shared(b)), "r"(parity));
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               :: "r"((uint32_t)__cvta_generic_to_shared(dst)), "l"(src), "r"(bytes), "r"((uint32_t)__cvta_generic_to_shared(bar)) : "memory");
}

// One producer thread keeps S bulk copies of CH bytes in flight; all warps consume (sum) each stage.
template <int S>
__global__ void bulk_kernel(const char* __restrict__ p, size_t bytes, int ch, float* sink) {
  extern __shared__ __align__(128) unsigned char smem[];
  __shared__ uint64_t full[S], empty[S];
  const int nthreads = blockDim.x;
  if (threadIdx.x == 0) {
    for (int s = 0; s < S; s++) { mbar_init(&full[s], 1); mbar_init(&empty[s], nthreads / 32); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  size_t n_chunks = bytes / ch;
  size_t per = (n_chunks + gridDim.x - 1) / gridDim.x;
  // interleave chunks across CTAs: chunk c -> CTA c % grid
  size_t my = 0;
  for (size_t c = blockIdx.x; c < n_chunks; c += gridDim.x) my++;
  (void)per;
  float acc = 0.f;
  if (threadIdx.x == 0) {
    // prologue
    size_t issued = 0;
    for (; issued < (size_t)S && issued < my; issued++) {
      mbar_expect(&full[issued], ch);
      bulk_g2s(smem + issued * ch, p + (blockIdx.x + issued * gridDim.x) * (size_t)ch, ch, &full[issued]);
    }
  }
  for (size_t k = 0; k < my; k++) {
    int s = k % S;
    uint32_t par = (k / S) & 1;
    mbar_wait(&full[s], par);
    const float4* v = reinterpret_cast<const float4*>(smem + (size_t)s * ch);
    for (int i = threadIdx.x; i < ch / 16; i += nthreads) { float4 t = v[i]; acc += t.x + t.y + t.z + t.w; }
    __syncwarp();
    if ((threadIdx.x & 31) == 0) mbar_arrive(&empty[s]);
    if (threadIdx.x == 0) {
      size_t nxt = k + S;
      if (nxt < my) {
        mbar_wait(&empty[s], par);
        mbar_expect(&full[s], ch);
        bulk_g2s(smem + (size_t)s * ch, p + (blockIdx.x + nxt * gridDim.x) * (size_t)ch, ch, &full[s]);
      }
    }
  }
  if (acc == 123.456f) *sink = acc;
}


// ---- scan-pattern probes: the product kernel's access pattern with features added one by one.
// MODE 0: loads + fp32 dot only;  1: + xor-shuffle reduce;  2: + ballot on score > thr (thr in smem)
template <int MODE, int U>
__global__ void __launch_bounds__(256) scanpat_kernel(const float4* __restrict__ y, long n_rows, const float4* __restrict__ q, float* sink) {
  __shared__ float s_thr;
  if (threadIdx.x == 0) s_thr = 1e30f;
  __syncthreads();
  constexpr int G = 8, NV = 4, RPI = 4, RPT = RPI * U, F4 = 32;
  const int lane = threadIdx.x & 31, gl = lane % G, grp = lane / G;
  float4 qv[NV];
#pragma unroll
  for (int v = 0; v < NV; v++) qv[v] = q[gl + v * G];
  const long warps_total = (long)gridDim.x * 8, gw = (long)blockIdx.x * 8 + (threadIdx.x >> 5);
  const long n_tiles = n_rows / RPT;
  float keep = 0.f;
  for (long tile = gw; tile < n_tiles; tile += warps_total) {
    const long row0 = tile * RPT + grp;
    float4 d[U][NV];
#pragma unroll
    for (int u = 0; u < U; u++)
#pragma unroll
      for (int v = 0; v < NV; v++) d[u][v] = ldg_stream(y + (row0 + u * RPI) * F4 + gl + v * G);
    float thr = MODE >= 2 ? *(volatile float*)&s_thr : 0.f;
#pragma unroll
    for (int u = 0; u < U; u++) {
      float acc = 0.f;
#pragma unroll
      for (int v = 0; v < NV; v++) {
        acc = fmaf(d[u][v].x, qv[v].x, acc); acc = fmaf(d[u][v].y, qv[v].y, acc);
        acc = fmaf(d[u][v].z, qv[v].z, acc); acc = fmaf(d[u][v].w, qv[v].w, acc);
      }
      if (MODE >= 1) {
#pragma unroll
        for (int o = G / 2; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
      }
      if (MODE >= 2) {
        unsigned ball = __ballot_sync(0xffffffffu, gl == 0 && acc > thr);
        if (ball) keep += 1.f;
      } else keep += acc;
    }
  }
  if (keep == 123.456f) *sink = keep;
}

__global__ void fill_random(uint32_t* p, size_t n, int mode) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  size_t stride = (size_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) {
    uint32_t x = (uint32_t)(i * 2654435761u) ^ (uint32_t)(i >> 32);
    x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
    if (mode == 1) {  // random float in [-2, 2): random mantissa + sign, narrow exponent range
      uint32_t sign = x & 0x80000000u, man = x & 0x007fffffu, e = 126 + ((x >> 23) & 1);
      x = sign | (e << 23) | man;
    }
    p[i] = x;
  }
}

int main(int argc, char** argv) {
  int fill_mode = argc > 1 ? atoi(argv[1]) : -1;
  size_t bytes = 5120000000ull;
  char* d; float* sink;
  CK(cudaMalloc(&d, bytes)); CK(cudaMalloc(&sink, 4));
  CK(cudaMemset(d, 1, bytes));
  if (fill_mode >= 0) { fill_random<<<148 * 8, 256>>>((uint32_t*)d, bytes / 4, fill_mode); CK(cudaDeviceSynchronize()); }
  printf("fill mode %d (-1 = constant bytes, 0 = random bits, 1 = random floats)\n", fill_mode);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  int sms = 148;
  auto time_it = [&](auto launch, const char* name) {
    for (int i = 0; i < 3; i++) launch();
    cudaEventRecord(e0);
    for (int i = 0; i < 10; i++) launch();
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    cudaError_t e = cudaGetLastError();
    printf("%-40s %8.1f us  %7.1f GB/s %s\n", name, ms * 100, bytes / (ms / 10 * 1e-3) / 1e9, e == cudaSuccess ? "" : cudaGetErrorString(e));
  };
  size_t n4 = bytes / 16;
  char name[128];
  float4* dq; cudaMalloc(&dq, 512); cudaMemset(dq, 0, 512);
  long n_rows = bytes / 512;
  for (int bps : {4}) {
    snprintf(name, 128, "scanpat mode0 U=2 x %d/SM", bps); time_it([&] { scanpat_kernel<0, 2><<<sms * bps, 256>>>((const float4*)d, n_rows, dq, sink); }, name);
    snprintf(name, 128, "scanpat mode1 U=2 x %d/SM", bps); time_it([&] { scanpat_kernel<1, 2><<<sms * bps, 256>>>((const float4*)d, n_rows, dq, sink); }, name);
    snprintf(name, 128, "scanpat mode2 U=2 x %d/SM", bps); time_it([&] { scanpat_kernel<2, 2><<<sms * bps, 256>>>((const float4*)d, n_rows, dq, sink); }, name);
    snprintf(name, 128, "scanpat mode2 U=1 x %d/SM", bps); time_it([&] { scanpat_kernel<2, 1><<<sms * bps, 256>>>((const float4*)d, n_rows, dq, sink); }, name);
    snprintf(name, 128, "scanpat mode2 U=4 x %d/SM", bps); time_it([&] { scanpat_kernel<2, 4><<<sms * bps, 256>>>((const float4*)d, n_rows, dq, sink); }, name);
  }
  for (int bps : {8}) {
    snprintf(name, 128, "scanpat mode2 U=1 x %d/SM", bps); time_it([&] { scanpat_kernel<2, 1><<<sms * bps, 256>>>((const float4*)d, n_rows, dq, sink); }, name);
    snprintf(name, 128, "scanpat mode2 U=2 x %d/SM", bps); time_it([&] { scanpat_kernel<2, 2><<<sms * bps, 256>>>((const float4*)d, n_rows, dq, sink); }, name);
  }
  for (int bps : {3, 4}) {
    snprintf(name, 128, "ldg U=4  256thr x %d/SM", bps); time_it([&] { ldg_kernel<4><<<sms * bps, 256>>>((const float4*)d, n4, sink); }, name);
    snprintf(name, 128, "ldg U=8  256thr x %d/SM", bps); time_it([&] { ldg_kernel<8><<<sms * bps, 256>>>((const float4*)d, n4, sink); }, name);
    snprintf(name, 128, "ldg U=16 256thr x %d/SM", bps); time_it([&] { ldg_kernel<16><<<sms * bps, 256>>>((const float4*)d, n4, sink); }, name);
  }
  for (int ch : {32768}) {
    for (int thr : {128, 256}) {
      {
        constexpr int S = 4; size_t sm = (size_t)S * ch;
        cudaFuncSetAttribute(bulk_kernel<S>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
        for (int bps : {1, 2}) { if (sm * bps > 220 * 1024) continue;
          snprintf(name, 128, "bulk S=4 ch=%d thr=%d x %d/SM", ch, thr, bps); time_it([&] { bulk_kernel<S><<<sms * bps, thr, sm>>>(d, bytes, ch, sink); }, name); }
      }
      {
        constexpr int S = 6; size_t sm = (size_t)S * ch;
        cudaFuncSetAttribute(bulk_kernel<S>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
        for (int bps : {1, 2}) { if (sm * bps > 220 * 1024) continue;
          snprintf(name, 128, "bulk S=6 ch=%d thr=%d x %d/SM", ch, thr, bps); time_it([&] { bulk_kernel<S><<<sms * bps, thr, sm>>>(d, bytes, ch, sink); }, name); }
      }
      {
        constexpr int S = 12; size_t sm = (size_t)S * ch;
        if (sm <= 220 * 1024) {
          cudaFuncSetAttribute(bulk_kernel<S>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
          snprintf(name, 128, "bulk S=12 ch=%d thr=%d x 1/SM", ch, thr); time_it([&] { bulk_kernel<S><<<sms, thr, sm>>>(d, bytes, ch, sink); }, name);
        }
      }
    }
  }
  return 0;
}
