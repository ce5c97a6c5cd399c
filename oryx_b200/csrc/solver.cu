// solver.cu -- rank-revealing QR factorisation and solve of the F x F Gram matrix on the device
// (kernel K9 of SURVEY.md section 2.3), in binary64 throughout.
//
// Replaces LinearSystemSolver.getSolver / isNonSingular (reference framework/oryx-common/.../math/
// LinearSystemSolver.java:35-63) and Solver.solveFToF (.../math/Solver.java:33-44).  Those delegate to
// commons-math3 3.6 RRQRDecomposition, which is not in the reference tree; its published algorithm
// is restated here: Householder QR on the transposed storage qrt[col][row], pivoting at every step
// on the remaining column of greatest squared 2-norm; the system is declared singular when any
// |rDiag| <= threshold, threshold = (max absolute column sum of M) * 1e-5; the reported apparent
// rank walks Frobenius norms of trailing blocks of R with drop threshold 0.01.
//
// One CTA factorises (F <= 1024 is tiny); reductions are warp-shuffle trees, so sums are ordered
// differently from the sequential Java loops -- results agree to binary64 rounding and are
// narrowed to float by solveFToF exactly as the reference does.
#include <algorithm>

#include "common.cuh"

namespace oryx {

struct Solver {
  int n = 0;
  int device = 0;
  double* d_qrt = nullptr;    // [n x n], qrt[col*n + row]
  double* d_rdiag = nullptr;  // [n]
  int32_t* d_perm = nullptr;  // [n]
  double* d_inv = nullptr;    // [n x n] row-major inverse of M (for the batched fold-in GEMM)
  double* d_work = nullptr;   // [2n] scratch for single solves
  float* d_bx = nullptr;      // [2n] staging for single solves
  cudaStream_t stream = nullptr;
  std::mutex mu;
  FoldScratch fold;
};

namespace {

constexpr int kQrThreads = 512;

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Block-wide sum, result broadcast to all threads.  `red` has >= 32 doubles.
__device__ double block_sum(double v, double* red) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  v = warp_sum(v);
  __syncthreads();
  if (lane == 0) red[warp] = v;
  __syncthreads();
  double t = (lane < nw) ? red[lane] : 0.0;
  t = warp_sum(t);
  return t;
}

// info[0] = 1 if non-singular else 0; info[1] = apparent rank (valid when singular)
__global__ void __launch_bounds__(kQrThreads) rrqr_kernel(const double* __restrict__ m, int n,
                                                          double* __restrict__ qrt,
                                                          double* __restrict__ rdiag,
                                                          int32_t* __restrict__ perm,
                                                          double* __restrict__ colnorm,
                                                          int32_t* __restrict__ info) {
  __shared__ double red[32];
  __shared__ int red_i[32];
  __shared__ int s_best;
  __shared__ double s_a, s_threshold;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nw = kQrThreads >> 5;

  // qrt = M^T; threshold = max absolute column sum * 1e-5 (LinearSystemSolver.java:31,38-39)
  for (int e = tid; e < n * n; e += kQrThreads) {
    int c = e / n, r = e - c * n;
    qrt[e] = m[(size_t)r * n + c];
  }
  for (int c = tid; c < n; c += kQrThreads) perm[c] = c;
  __syncthreads();
  {
    double best = 0.0;
    for (int c = warp; c < n; c += nw) {
      double s = 0.0;
      for (int r = lane; r < n; r += 32) s += fabs(qrt[(size_t)c * n + r]);
      s = warp_sum(s);
      best = fmax(best, s);
    }
    if (lane == 0) red[warp] = best;
    __syncthreads();
    if (tid == 0) {
      double b = 0.0;
      for (int w = 0; w < nw; w++) b = fmax(b, red[w]);
      s_threshold = b * 1.0e-5;
    }
    __syncthreads();
  }

  for (int minor = 0; minor < n; minor++) {
    // (1) pivot: remaining column with the greatest squared 2-norm (first such on ties)
    for (int c = minor + warp; c < n; c += nw) {
      double s = 0.0;
      for (int r = lane; r < n; r += 32) { double v = qrt[(size_t)c * n + r]; s += v * v; }
      s = warp_sum(s);
      if (lane == 0) colnorm[c] = s;
    }
    __syncthreads();
    {
      double bv = 0.0;
      int bi = minor;
      for (int c = minor + tid; c < n; c += kQrThreads) {
        double v = colnorm[c];
        if (v > bv) { bv = v; bi = c; }
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        double ov = __shfl_xor_sync(0xffffffffu, bv, o);
        int oi = __shfl_xor_sync(0xffffffffu, bi, o);
        if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
      }
      if (lane == 0) { red[warp] = bv; red_i[warp] = bi; }
      __syncthreads();
      if (tid == 0) {
        double b = 0.0;
        int idx = minor;
        for (int w = 0; w < nw; w++) {
          if (red[w] > b || (red[w] == b && red[w] > 0.0 && red_i[w] < idx)) { b = red[w]; idx = red_i[w]; }
        }
        s_best = idx;
      }
      __syncthreads();
    }
    const int best = s_best;
    if (best != minor) {
      for (int r = tid; r < n; r += kQrThreads) {
        double t = qrt[(size_t)minor * n + r];
        qrt[(size_t)minor * n + r] = qrt[(size_t)best * n + r];
        qrt[(size_t)best * n + r] = t;
      }
      if (tid == 0) { int t = perm[minor]; perm[minor] = perm[best]; perm[best] = t; }
    }
    __syncthreads();
    // (2) Householder reflection of column `minor`
    double* qm = qrt + (size_t)minor * n;
    double part = 0.0;
    for (int r = minor + tid; r < n; r += kQrThreads) part += qm[r] * qm[r];
    double x_norm_sqr = block_sum(part, red);
    if (tid == 0) {
      double a = (qm[minor] > 0) ? -sqrt(x_norm_sqr) : sqrt(x_norm_sqr);
      rdiag[minor] = a;
      s_a = a;
      if (a != 0.0) qm[minor] -= a;
    }
    __syncthreads();
    const double a = s_a;
    if (a != 0.0) {
      const double denom = a * qm[minor];
      for (int c = minor + 1 + warp; c < n; c += nw) {
        double* qc = qrt + (size_t)c * n;
        double alpha = 0.0;
        for (int r = minor + lane; r < n; r += 32) alpha -= qc[r] * qm[r];
        alpha = warp_sum(alpha);
        alpha /= denom;
        for (int r = minor + lane; r < n; r += 32) qc[r] -= alpha * qm[r];
      }
    }
    __syncthreads();
  }

  // singularity verdict and apparent rank
  if (tid == 0) {
    int ok = 1;
    for (int i = 0; i < n; i++) if (fabs(rdiag[i]) <= s_threshold) ok = 0;
    info[0] = ok;
  }
  __syncthreads();
  if (info[0] == 0) {
    // RRQRDecomposition.getRank(0.01): R[row][col] = rdiag on the diagonal, qrt[col][row] above it.
    // suffix[k] = squared Frobenius norm of R[k.., k..]
    for (int k = tid; k < n; k += kQrThreads) {
      // contribution of "shell" k: row k (cols >= k) and column k (rows > k): only row k is non-zero
      double s = rdiag[k] * rdiag[k];
      for (int c = k + 1; c < n; c++) { double v = qrt[(size_t)c * n + k]; s += v * v; }
      colnorm[k] = s;
    }
    __syncthreads();
    if (tid == 0) {
      for (int k = n - 2; k >= 0; k--) colnorm[k] += colnorm[k + 1];
      double r_norm = sqrt(colnorm[0]);
      double last_norm = r_norm;
      int rank = 1;
      while (rank < n) {
        double this_norm = sqrt(colnorm[rank]);
        if (this_norm == 0 || (this_norm / last_norm) * r_norm < 0.01) break;
        last_norm = this_norm;
        rank++;
      }
      info[1] = rank;
    }
  }
}

// Solve M x = b for `nrhs` right-hand sides held as columns of the identity (inverse) or given
// explicitly.  One CTA per right-hand side; follows QRDecomposition.Solver.solve then the RRQR
// un-permutation.  y: [nrhs x n] work (in: b, out: permuted x), out: [n x nrhs] or [nrhs x n].
__global__ void __launch_bounds__(256) qr_solve_kernel(const double* __restrict__ qrt,
                                                       const double* __restrict__ rdiag,
                                                       const int32_t* __restrict__ perm, int n,
                                                       const double* __restrict__ b, int identity_rhs,
                                                       double* __restrict__ work, double* __restrict__ out,
                                                       int out_col_major) {
  __shared__ double red[32];
  __shared__ double s_dp;
  const int rhs = blockIdx.x, tid = threadIdx.x;
  double* y = work + (size_t)rhs * n;
  for (int r = tid; r < n; r += blockDim.x) y[r] = identity_rhs ? (r == rhs ? 1.0 : 0.0) : b[(size_t)rhs * n + r];
  __syncthreads();
  for (int minor = 0; minor < n; minor++) {
    const double* qm = qrt + (size_t)minor * n;
    double part = 0.0;
    for (int r = minor + tid; r < n; r += blockDim.x) part += y[r] * qm[r];
    double dp = block_sum(part, red);
    if (tid == 0) s_dp = dp / (rdiag[minor] * qm[minor]);
    __syncthreads();
    dp = s_dp;
    for (int r = minor + tid; r < n; r += blockDim.x) y[r] += dp * qm[r];
    __syncthreads();
  }
  for (int row = n - 1; row >= 0; --row) {
    if (tid == 0) y[row] /= rdiag[row];
    __syncthreads();
    const double y_row = y[row];
    const double* qr = qrt + (size_t)row * n;
    for (int i = tid; i < row; i += blockDim.x) y[i] -= y_row * qr[i];
    __syncthreads();
  }
  for (int i = tid; i < n; i += blockDim.x) {
    int dst = perm[i];
    if (out_col_major) out[(size_t)dst * gridDim.x + rhs] = y[i];   // out[row dst][col rhs]
    else out[(size_t)rhs * n + dst] = y[i];
  }
}

__global__ void f2d_kernel(const float* __restrict__ in, double* __restrict__ out, int n) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = (double)in[i];
}
__global__ void d2f_kernel(const double* __restrict__ in, float* __restrict__ out, int n) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = (float)in[i];  // Solver.java:40-42 narrowing, round-to-nearest
}

void solver_free(Solver* s) {
  if (!s) return;
  cudaFree(s->d_qrt); cudaFree(s->d_rdiag); cudaFree(s->d_perm); cudaFree(s->d_inv);
  cudaFree(s->d_work); cudaFree(s->d_bx);
  fold_scratch_free(&s->fold);
  if (s->stream) cudaStreamDestroy(s->stream);
  delete s;
}

}  // namespace

const double* solver_inverse(const Solver* s) { return s->d_inv; }
int solver_n(const Solver* s) { return s->n; }
int solver_device(const Solver* s) { return s->device; }
cudaStream_t solver_stream(const Solver* s) { return s->stream; }
std::mutex& solver_mutex(Solver* s) { return s->mu; }
FoldScratch* solver_fold_scratch(Solver* s) { return &s->fold; }

}  // namespace oryx

using namespace oryx;

extern "C" {

int32_t oryx_solver_create(const double* gram, int32_t features, int32_t device, oryx_solver** out,
                           int32_t* apparent_rank) {
  ORYX_ARG(gram && out, "null argument");
  ORYX_ARG(features > 0 && features <= 1024, "features must be in [1, 1024]");
  *out = nullptr;
  int prev = -1;
  if (cudaGetDevice(&prev) != cudaSuccess) { cudaGetLastError(); prev = -1; }
  if (cudaSetDevice(device) != cudaSuccess) {
    cudaGetLastError();
    set_error("cannot select CUDA device %d (no CPU fallback exists)", device);
    return ORYX_E_CUDA;
  }
  struct Restore { int p; ~Restore() { if (p >= 0) cudaSetDevice(p); } } restore{prev};
  const int n = features;
  Solver* s = new Solver();
  s->n = n;
  s->device = device;
  double* d_m = nullptr;
  double* d_colnorm = nullptr;
  int32_t* d_info = nullptr;
  int32_t rc = ORYX_OK;
  int32_t h_info[2] = {0, 0};
  auto fail = [&](int32_t code) { cudaFree(d_m); cudaFree(d_colnorm); cudaFree(d_info); solver_free(s); return code; };
#define SOLVER_CUDA(expr)                                                        \
  do {                                                                           \
    cudaError_t e__ = (expr);                                                    \
    if (e__ != cudaSuccess) {                                                    \
      set_error("CUDA: %s -> %s (%s:%d)", #expr, cudaGetErrorString(e__), __FILE__, __LINE__); \
      return fail(ORYX_E_CUDA);                                                  \
    }                                                                            \
  } while (0)
  SOLVER_CUDA(cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking));
  SOLVER_CUDA(cudaMalloc(&d_m, (size_t)n * n * sizeof(double)));
  SOLVER_CUDA(cudaMalloc(&d_colnorm, (size_t)n * sizeof(double)));
  SOLVER_CUDA(cudaMalloc(&d_info, 2 * sizeof(int32_t)));
  SOLVER_CUDA(cudaMalloc(&s->d_qrt, (size_t)n * n * sizeof(double)));
  SOLVER_CUDA(cudaMalloc(&s->d_rdiag, (size_t)n * sizeof(double)));
  SOLVER_CUDA(cudaMalloc(&s->d_perm, (size_t)n * sizeof(int32_t)));
  SOLVER_CUDA(cudaMalloc(&s->d_inv, (size_t)n * n * sizeof(double)));
  SOLVER_CUDA(cudaMalloc(&s->d_work, (size_t)n * std::max(n, 3) * sizeof(double)));
  SOLVER_CUDA(cudaMalloc(&s->d_bx, (size_t)2 * n * sizeof(float)));
  SOLVER_CUDA(cudaMemcpyAsync(d_m, gram, (size_t)n * n * sizeof(double), cudaMemcpyHostToDevice, s->stream));
  SOLVER_CUDA(cudaMemsetAsync(d_info, 0, 2 * sizeof(int32_t), s->stream));
  rrqr_kernel<<<1, kQrThreads, 0, s->stream>>>(d_m, n, s->d_qrt, s->d_rdiag, s->d_perm, d_colnorm, d_info);
  SOLVER_CUDA(cudaGetLastError());
  SOLVER_CUDA(cudaMemcpyAsync(h_info, d_info, 2 * sizeof(int32_t), cudaMemcpyDeviceToHost, s->stream));
  SOLVER_CUDA(cudaStreamSynchronize(s->stream));
  if (h_info[0] == 0) {
    if (apparent_rank) *apparent_rank = h_info[1];
    set_error("%d x %d matrix is near-singular; apparent rank: %d", n, n, h_info[1]);
    return fail(ORYX_E_SINGULAR);
  }
  // inverse, one CTA per column of the identity; stored row-major (out[row][col])
  qr_solve_kernel<<<n, 256, 0, s->stream>>>(s->d_qrt, s->d_rdiag, s->d_perm, n, nullptr, 1, s->d_work, s->d_inv, 1);
  SOLVER_CUDA(cudaGetLastError());
  SOLVER_CUDA(cudaStreamSynchronize(s->stream));
#undef SOLVER_CUDA
  cudaFree(d_m); cudaFree(d_colnorm); cudaFree(d_info);
  if (apparent_rank) *apparent_rank = n;
  *out = reinterpret_cast<oryx_solver*>(s);
  return rc;
}

int32_t oryx_solver_destroy(oryx_solver* sh) {
  if (!sh) return ORYX_OK;
  Solver* s = reinterpret_cast<Solver*>(sh);
  int prev = -1;
  if (cudaGetDevice(&prev) != cudaSuccess) { cudaGetLastError(); prev = -1; }
  cudaSetDevice(s->device);
  solver_free(s);
  if (prev >= 0) cudaSetDevice(prev);
  cudaGetLastError();
  return ORYX_OK;
}

int32_t oryx_solve_f2f(oryx_solver* sh, const float* b, float* x) {
  ORYX_ARG(sh && b && x, "null argument");
  Solver* s = reinterpret_cast<Solver*>(sh);
  int prev = -1;
  if (cudaGetDevice(&prev) != cudaSuccess) { cudaGetLastError(); prev = -1; }
  if (cudaSetDevice(s->device) != cudaSuccess) {
    cudaGetLastError();
    set_error("cannot select CUDA device %d", s->device);
    return ORYX_E_CUDA;
  }
  struct Restore { int p; ~Restore() { if (p >= 0) cudaSetDevice(p); } } restore{prev};
  std::lock_guard<std::mutex> lk(s->mu);
  const int n = s->n;
  double* d_b = s->d_work;          // [n] widened b
  double* d_y = s->d_work + n;      // [n] work
  double* d_x = s->d_work + 2 * n;  // [n] un-permuted solution
  ORYX_CUDA(cudaMemcpyAsync(s->d_bx, b, (size_t)n * sizeof(float), cudaMemcpyHostToDevice, s->stream));
  f2d_kernel<<<(n + 255) / 256, 256, 0, s->stream>>>(s->d_bx, d_b, n);
  qr_solve_kernel<<<1, 256, 0, s->stream>>>(s->d_qrt, s->d_rdiag, s->d_perm, n, d_b, 0, d_y, d_x, 0);
  d2f_kernel<<<(n + 255) / 256, 256, 0, s->stream>>>(d_x, s->d_bx + n, n);
  ORYX_CUDA(cudaGetLastError());
  ORYX_CUDA(cudaMemcpyAsync(x, s->d_bx + n, (size_t)n * sizeof(float), cudaMemcpyDeviceToHost, s->stream));
  ORYX_CUDA(cudaStreamSynchronize(s->stream));
  return ORYX_OK;
}

}  // extern "C"
