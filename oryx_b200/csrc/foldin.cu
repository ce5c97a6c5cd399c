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

#include "common.cuh"

namespace oryx {

struct Solver;
const double* solver_inverse(const Solver* s);
int solver_n(const Solver* s);
int solver_device(const Solver* s);
cudaStream_t solver_stream(const Solver* s);
std::mutex& solver_mutex(Solver* s);
int32_t gram_launch(Model* m, int which, double* d_out, cudaStream_t stream);

namespace {

constexpr int kRowsPerCta = 64;  // interactions per CTA

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

// One CTA = 64 interactions x all F outputs.  Phase 1: one thread per interaction computes Qui with
// the exact sequential arithmetic and the scalar dQui.  Phase 2: binary64 GEMM
// X[k][r] = sum_c inv[r][c] * b[k][c], tiled 64 (interactions) x 64 (outputs) x 16 (c).
__global__ void __launch_bounds__(256) fold_in_kernel(const double* __restrict__ inv, int f, int implicit,
                                                      const float* __restrict__ xu,
                                                      const uint8_t* __restrict__ has_xu,
                                                      const float* __restrict__ yi,
                                                      const double* __restrict__ values, int64_t n,
                                                      float* __restrict__ out, uint8_t* __restrict__ out_valid) {
  __shared__ double s_dq[kRowsPerCta];
  __shared__ uint8_t s_valid[kRowsPerCta];
  __shared__ uint8_t s_has[kRowsPerCta];
  __shared__ double Bs[16][kRowsPerCta + 1];  // b tile: [c][k]
  __shared__ double Is[16][64 + 1];           // inverse tile: [c][r]
  const int tid = threadIdx.x;
  const int64_t k0 = (int64_t)blockIdx.x * kRowsPerCta;

  if (tid < kRowsPerCta) {
    int64_t k = k0 + tid;
    double dq = 0.0;
    uint8_t valid = 0, has = 0;
    if (k < n) {
      has = has_xu ? has_xu[k] : 1;
      const float* y = yi + k * f;
      double qui = 0.0;
      if (has) {
        const float* x = xu + k * f;
        for (int j = 0; j < f; j++) qui = exact_step(qui, x[j], y[j]);
      }
      double target = target_qui(implicit != 0, values[k], has ? qui : 0.5);
      if (target == target) {  // not NaN
        valid = 1;
        dq = target - qui;
      }
    }
    s_dq[tid] = dq;
    s_valid[tid] = valid;
    s_has[tid] = has;
    if (k < n) out_valid[k] = valid;
  }
  __syncthreads();

  const int tx = tid & 15, ty = tid >> 4;  // ty -> interactions (4 each), tx -> outputs (4 each)
  for (int r0 = 0; r0 < f; r0 += 64) {
    double acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
      for (int j = 0; j < 4; j++) acc[i][j] = 0.0;
    for (int c0 = 0; c0 < f; c0 += 16) {
      __syncthreads();
      // b tile: b[k][c] = (float)((double)yi[k][c] * dQui[k])  (ALSUtils.java:93-95)
      for (int e = tid; e < 16 * kRowsPerCta; e += 256) {
        int kk = e >> 4, c = e & 15;
        int64_t k = k0 + kk;
        double v = 0.0;
        if (k < n && c0 + c < f) {
          float bf = (float)((double)yi[k * f + c0 + c] * s_dq[kk]);
          v = (double)bf;
        }
        Bs[c][kk] = v;
      }
      // inverse tile: inv[r][c] (row-major)
      for (int e = tid; e < 16 * 64; e += 256) {
        int r = e >> 4, c = e & 15;
        Is[c][r] = (r0 + r < f && c0 + c < f) ? inv[(size_t)(r0 + r) * f + c0 + c] : 0.0;
      }
      __syncthreads();
#pragma unroll
      for (int c = 0; c < 16; c++) {
        double bv[4], iv[4];
#pragma unroll
        for (int i = 0; i < 4; i++) bv[i] = Bs[c][ty * 4 + i];
#pragma unroll
        for (int j = 0; j < 4; j++) iv[j] = Is[c][tx * 4 + j];
#pragma unroll
        for (int i = 0; i < 4; i++)
#pragma unroll
          for (int j = 0; j < 4; j++) acc[i][j] = fma(bv[i], iv[j], acc[i][j]);
      }
    }
#pragma unroll
    for (int i = 0; i < 4; i++) {
      int kk = ty * 4 + i;
      int64_t k = k0 + kk;
      if (k >= n) continue;
#pragma unroll
      for (int j = 0; j < 4; j++) {
        int r = r0 + tx * 4 + j;
        if (r >= f) continue;
        float dxu = (float)acc[i][j];  // Solver.java:40-42
        float res = 0.0f;
        if (s_valid[kk]) res = s_has[kk] ? __fadd_rn(xu[k * f + r], dxu) : dxu;  // ALSUtils.java:98-104
        out[k * f + r] = res;
      }
    }
  }
}

}  // namespace
}  // namespace oryx

using namespace oryx;

extern "C" {

int32_t oryx_fold_in_batch(oryx_solver* sh, int32_t implicit, const float* xu, const uint8_t* has_xu,
                           const float* yi, const double* values, int64_t n, float* out, uint8_t* out_valid) {
  ORYX_ARG(sh && yi && values && out && out_valid, "null argument");
  ORYX_ARG(n >= 0, "n must be >= 0");
  if (n == 0) return ORYX_OK;
  Solver* s = reinterpret_cast<Solver*>(sh);
  const int f = solver_n(s);
  int prev = -1;
  if (cudaGetDevice(&prev) != cudaSuccess) { cudaGetLastError(); prev = -1; }
  if (cudaSetDevice(solver_device(s)) != cudaSuccess) {
    cudaGetLastError();
    set_error("cannot select CUDA device %d", solver_device(s));
    return ORYX_E_CUDA;
  }
  struct Restore { int p; ~Restore() { if (p >= 0) cudaSetDevice(p); } } restore{prev};
  std::lock_guard<std::mutex> lk(solver_mutex(s));
  cudaStream_t st = solver_stream(s);
  const int64_t chunk = 1 << 18;  // interactions per staging round
  float *d_xu = nullptr, *d_yi = nullptr, *d_out = nullptr;
  double* d_val = nullptr;
  uint8_t *d_has = nullptr, *d_valid = nullptr;
  int64_t c_max = std::min<int64_t>(chunk, n);
  int32_t rc = ORYX_OK;
  auto cleanup = [&]() { cudaFree(d_xu); cudaFree(d_yi); cudaFree(d_out); cudaFree(d_val); cudaFree(d_has); cudaFree(d_valid); };
  if (cudaMalloc(&d_xu, (size_t)c_max * f * sizeof(float)) != cudaSuccess ||
      cudaMalloc(&d_yi, (size_t)c_max * f * sizeof(float)) != cudaSuccess ||
      cudaMalloc(&d_out, (size_t)c_max * f * sizeof(float)) != cudaSuccess ||
      cudaMalloc(&d_val, (size_t)c_max * sizeof(double)) != cudaSuccess ||
      cudaMalloc(&d_has, (size_t)c_max) != cudaSuccess || cudaMalloc(&d_valid, (size_t)c_max) != cudaSuccess) {
    cudaGetLastError();
    cleanup();
    set_error("device allocation failed for a fold-in batch of %lld", (long long)c_max);
    return ORYX_E_NOMEM;
  }
  for (int64_t off = 0; off < n && rc == ORYX_OK; off += chunk) {
    int64_t c = std::min<int64_t>(chunk, n - off);
    cudaError_t e = cudaSuccess;
    if (xu) e = cudaMemcpyAsync(d_xu, xu + off * f, (size_t)c * f * sizeof(float), cudaMemcpyHostToDevice, st);
    else e = cudaMemsetAsync(d_xu, 0, (size_t)c * f * sizeof(float), st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_yi, yi + off * f, (size_t)c * f * sizeof(float), cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_val, values + off, (size_t)c * sizeof(double), cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) {
      if (has_xu) e = cudaMemcpyAsync(d_has, has_xu + off, (size_t)c, cudaMemcpyHostToDevice, st);
      else e = cudaMemsetAsync(d_has, xu ? 1 : 0, (size_t)c, st);
    }
    if (e == cudaSuccess) {
      unsigned grid = (unsigned)((c + kRowsPerCta - 1) / kRowsPerCta);
      fold_in_kernel<<<grid, 256, 0, st>>>(solver_inverse(s), f, implicit, d_xu, d_has, d_yi, d_val, c, d_out, d_valid);
      e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(out + off * f, d_out, (size_t)c * f * sizeof(float), cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(out_valid + off, d_valid, (size_t)c, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) {
      set_error("fold-in batch failed: %s", cudaGetErrorString(e));
      rc = ORYX_E_CUDA;
    }
  }
  cleanup();
  return rc;
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
  int prev = -1;
  if (cudaGetDevice(&prev) != cudaSuccess) { cudaGetLastError(); prev = -1; }
  if (cudaSetDevice(m->device) != cudaSuccess) {
    cudaGetLastError();
    set_error("cannot select CUDA device %d", m->device);
    return ORYX_E_CUDA;
  }
  struct Restore { int p; ~Restore() { if (p >= 0) cudaSetDevice(p); } } restore{prev};
  std::lock_guard<std::mutex> lk(solver_mutex(s));
  std::shared_lock<std::shared_mutex> model_lock(m->mu);  // the stores must not move while rows are gathered
  const Store* sx = &m->store[xu_which];
  const Store* sy = &m->store[xu_which == ORYX_X ? ORYX_Y : ORYX_X];
  cudaStream_t st = solver_stream(s);
  const int64_t chunk = 1 << 18;
  const int64_t c_max = std::min<int64_t>(chunk, n);
  float *d_xu = nullptr, *d_yi = nullptr, *d_out = nullptr;
  double* d_val = nullptr;
  uint8_t *d_has = nullptr, *d_valid = nullptr;
  int32_t* d_slots = nullptr;  // [2 x c_max]
  int32_t* h_slots = nullptr;  // pinned
  auto cleanup = [&]() {
    cudaFree(d_xu); cudaFree(d_yi); cudaFree(d_out); cudaFree(d_val); cudaFree(d_has); cudaFree(d_valid);
    cudaFree(d_slots); cudaFreeHost(h_slots);
  };
  if (cudaMalloc(&d_xu, (size_t)c_max * f * sizeof(float)) != cudaSuccess ||
      cudaMalloc(&d_yi, (size_t)c_max * f * sizeof(float)) != cudaSuccess ||
      cudaMalloc(&d_out, (size_t)c_max * f * sizeof(float)) != cudaSuccess ||
      cudaMalloc(&d_val, (size_t)c_max * sizeof(double)) != cudaSuccess ||
      cudaMalloc(&d_has, (size_t)c_max) != cudaSuccess || cudaMalloc(&d_valid, (size_t)c_max) != cudaSuccess ||
      cudaMalloc(&d_slots, (size_t)c_max * 2 * sizeof(int32_t)) != cudaSuccess ||
      cudaMallocHost(&h_slots, (size_t)c_max * 2 * sizeof(int32_t)) != cudaSuccess) {
    cudaGetLastError();
    cleanup();
    set_error("allocation failed for a fold-in batch of %lld", (long long)c_max);
    return ORYX_E_NOMEM;
  }
  auto slot_of = [](const Store* store, int32_t row) -> int32_t {
    return (row >= 0 && (size_t)row < store->row2slot.size()) ? store->row2slot[(size_t)row] : -1;
  };
  int32_t rc = ORYX_OK;
  for (int64_t off = 0; off < n && rc == ORYX_OK; off += chunk) {
    const int64_t c = std::min<int64_t>(chunk, n - off);
    for (int64_t i = 0; i < c; i++) {
      h_slots[i] = slot_of(sx, xu_rows[off + i]);
      h_slots[c_max + i] = slot_of(sy, yi_rows[off + i]);
    }
    cudaError_t e = cudaMemcpyAsync(d_slots, h_slots, (size_t)c_max * 2 * sizeof(int32_t), cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_val, values + off, (size_t)c * sizeof(double), cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) {
      const unsigned g = (unsigned)((c * 32 + 255) / 256);
      gather_rows_kernel<<<g, 256, 0, st>>>(sx->d_vec, sx->fpad, f, d_slots, c, d_xu, d_has);
      gather_rows_kernel<<<g, 256, 0, st>>>(sy->d_vec, sy->fpad, f, d_slots + c_max, c, d_yi, nullptr);
      const unsigned grid = (unsigned)((c + kRowsPerCta - 1) / kRowsPerCta);
      fold_in_kernel<<<grid, 256, 0, st>>>(solver_inverse(s), f, implicit, d_xu, d_has, d_yi, d_val, c, d_out, d_valid);
      m->launches += 3;
      e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(out + off * f, d_out, (size_t)c * f * sizeof(float), cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(out_valid + off, d_valid, (size_t)c, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) {
      set_error("fold-in batch failed: %s", cudaGetErrorString(e));
      rc = ORYX_E_CUDA;
      break;
    }
    // computeUpdatedXu returns null when Yi is null (ALSUtils.java:79-81)
    for (int64_t i = 0; i < c; i++) if (h_slots[c_max + i] < 0) out_valid[off + i] = 0;
  }
  cleanup();
  return rc;
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
  int prev = -1;
  if (cudaGetDevice(&prev) != cudaSuccess) { cudaGetLastError(); prev = -1; }
  if (cudaSetDevice(dev) != cudaSuccess) {
    cudaGetLastError();
    set_error("cannot select CUDA device %d", dev);
    return ORYX_E_CUDA;
  }
  struct Restore { int p; ~Restore() { if (p >= 0) cudaSetDevice(p); } } restore{prev};
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
  const int64_t chunk = 1 << 18;
  const int64_t c_max = std::min<int64_t>(chunk, n);
  float *d_xu = nullptr, *d_yi = nullptr, *d_out = nullptr;
  double* d_val = nullptr;
  uint8_t *d_has = nullptr, *d_valid = nullptr;
  int32_t* d_idx = nullptr;  // [4 x c_max]: xu owner, xu slot, yi owner, yi slot
  int32_t* h_idx = nullptr;  // pinned
  auto cleanup = [&]() {
    cudaFree(d_xu); cudaFree(d_yi); cudaFree(d_out); cudaFree(d_val); cudaFree(d_has); cudaFree(d_valid);
    cudaFree(d_idx); cudaFreeHost(h_idx);
  };
  if (cudaMalloc(&d_xu, (size_t)c_max * f * sizeof(float)) != cudaSuccess ||
      cudaMalloc(&d_yi, (size_t)c_max * f * sizeof(float)) != cudaSuccess ||
      cudaMalloc(&d_out, (size_t)c_max * f * sizeof(float)) != cudaSuccess ||
      cudaMalloc(&d_val, (size_t)c_max * sizeof(double)) != cudaSuccess ||
      cudaMalloc(&d_has, (size_t)c_max) != cudaSuccess || cudaMalloc(&d_valid, (size_t)c_max) != cudaSuccess ||
      cudaMalloc(&d_idx, (size_t)c_max * 4 * sizeof(int32_t)) != cudaSuccess ||
      cudaMallocHost(&h_idx, (size_t)c_max * 4 * sizeof(int32_t)) != cudaSuccess) {
    cudaGetLastError();
    cleanup();
    set_error("allocation failed for a fold-in batch of %lld", (long long)c_max);
    return ORYX_E_NOMEM;
  }
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
  int32_t rc = ORYX_OK;
  for (int64_t off = 0; off < n && rc == ORYX_OK; off += chunk) {
    const int64_t c = std::min<int64_t>(chunk, n - off);
    for (int64_t i = 0; i < c; i++) {
      locate(xu_which, xu_row_stride, xu_rows[off + i], &h_idx[i], &h_idx[c_max + i]);
      locate(yi_which, yi_row_stride, yi_rows[off + i], &h_idx[2 * c_max + i], &h_idx[3 * c_max + i]);
    }
    cudaError_t e = cudaMemcpyAsync(d_idx, h_idx, (size_t)c_max * 4 * sizeof(int32_t), cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_val, values + off, (size_t)c * sizeof(double), cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) {
      const unsigned g = (unsigned)((c * 32 + 255) / 256);
      gather_rows_sharded_kernel<<<g, 256, 0, st>>>(bx, fpad, f, d_idx, d_idx + c_max, c, d_xu, d_has);
      gather_rows_sharded_kernel<<<g, 256, 0, st>>>(by, fpad, f, d_idx + 2 * c_max, d_idx + 3 * c_max, c, d_yi, nullptr);
      const unsigned grid = (unsigned)((c + kRowsPerCta - 1) / kRowsPerCta);
      fold_in_kernel<<<grid, 256, 0, st>>>(solver_inverse(s), f, implicit, d_xu, d_has, d_yi, d_val, c, d_out, d_valid);
      e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(out + off * f, d_out, (size_t)c * f * sizeof(float), cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(out_valid + off, d_valid, (size_t)c, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) {
      set_error("sharded fold-in batch failed: %s", cudaGetErrorString(e));
      rc = ORYX_E_CUDA;
      break;
    }
    for (int64_t i = 0; i < c; i++) if (h_idx[3 * c_max + i] < 0) out_valid[off + i] = 0;  // Yi == null -> null
  }
  cleanup();
  return rc;
}

int32_t oryx_gram(oryx_model* mh, int32_t which, double* out, int64_t* out_rows_used) {
  ORYX_ARG(mh && out, "null argument");
  ORYX_ARG(which == ORYX_X || which == ORYX_Y, "which must be ORYX_X or ORYX_Y");
  Model* m = reinterpret_cast<Model*>(mh);
  int prev = -1;
  if (cudaGetDevice(&prev) != cudaSuccess) { cudaGetLastError(); prev = -1; }
  if (cudaSetDevice(m->device) != cudaSuccess) {
    cudaGetLastError();
    set_error("cannot select CUDA device %d", m->device);
    return ORYX_E_CUDA;
  }
  struct Restore { int p; ~Restore() { if (p >= 0) cudaSetDevice(p); } } restore{prev};
  const int F = m->cfg.features;
  QueryCtx* c = nullptr;
  ORYX_TRY(ctx_acquire(m, &c));
  double* d_out = nullptr;
  int32_t rc = ORYX_OK;
  do {
    if (cudaMalloc(&d_out, (size_t)F * F * sizeof(double)) != cudaSuccess) {
      cudaGetLastError(); set_error("device allocation failed"); rc = ORYX_E_NOMEM; break;
    }
    std::shared_lock<std::shared_mutex> lk(m->mu);
    if (out_rows_used) *out_rows_used = m->store[which].live;
    rc = gram_launch(m, which, d_out, c->stream);
    if (rc != ORYX_OK) break;
    if (cudaMemcpyAsync(out, d_out, (size_t)F * F * sizeof(double), cudaMemcpyDeviceToHost, c->stream) != cudaSuccess ||
        cudaStreamSynchronize(c->stream) != cudaSuccess) {
      set_error("gram failed: %s", cudaGetErrorString(cudaGetLastError()));
      rc = ORYX_E_CUDA;
    }
  } while (0);
  cudaFree(d_out);
  ctx_release(m, c);
  return rc;
}

}  // extern "C"
