// Binary64 throughput probe for B200: scalar DFMA vs FP64 tensor-core DMMA (mma.sync m8n8k4), registers only.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a tools/fp64_probe.cu -o tools/fp64_probe && ./tools/fp64_probe
#include <cstdio>
#include <cuda_runtime.h>

__global__ void dfma_kernel(double* out, int iters) {
  double a[16];
#pragma unroll
  for (int i = 0; i < 16; i++) a[i] = threadIdx.x * 1e-3 + i;
  const double b = 1.0000001, c = 1e-9;
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int i = 0; i < 16; i++) a[i] = fma(a[i], b, c);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 16; i++) s += a[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void dmma_kernel(double* out, int iters) {
  double d[8][2];
#pragma unroll
  for (int i = 0; i < 8; i++) d[i][0] = d[i][1] = threadIdx.x * 1e-3 + i;
  const double a = 1.0000001, b = 1e-3;
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int i = 0; i < 8; i++)
      asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0, %1}, {%2}, {%3}, {%0, %1};"
                   : "+d"(d[i][0]), "+d"(d[i][1]) : "d"(a), "d"(b));
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) s += d[i][0] + d[i][1];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

int main() {
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  double* out;
  cudaMalloc(&out, sizeof(double) * sms * 8 * 256);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int iters = 20000;
  for (int warps_per_sm : {8, 16, 32, 64}) {
    const int ctas = sms * warps_per_sm / 8;
    float ms;
    dfma_kernel<<<ctas, 256>>>(out, 100);
    cudaEventRecord(e0); dfma_kernel<<<ctas, 256>>>(out, iters); cudaEventRecord(e1); cudaEventSynchronize(e1);
    cudaEventElapsedTime(&ms, e0, e1);
    const double dfma_tf = 2.0 * 16 * iters * (double)ctas * 256 / (ms * 1e-3) / 1e12;
    dmma_kernel<<<ctas, 256>>>(out, 100);
    cudaEventRecord(e0); dmma_kernel<<<ctas, 256>>>(out, iters); cudaEventRecord(e1); cudaEventSynchronize(e1);
    cudaEventElapsedTime(&ms, e0, e1);
    const double dmma_tf = 2.0 * 256 * 8 * iters * (double)ctas * 8 / (ms * 1e-3) / 1e12;
    printf("%2d warps/SM: DFMA %.2f TFLOP/s, DMMA m8n8k4 %.2f TFLOP/s\n", warps_per_sm, dfma_tf, dmma_tf);
  }
  printf("status %s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
