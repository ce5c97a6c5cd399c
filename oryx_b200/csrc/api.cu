// api.cu -- the extern "C" surface declared in include/oryx_b200.h: handle lifetime, locking,
// host<->device staging around the kernels.  Each entry point's reference counterpart is cited in
// the header.
#include <algorithm>
#include <cmath>

#include "common.cuh"

namespace oryx {

namespace {
thread_local char g_err[512] = {0};
}

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
const char* last_error() { return g_err; }

void ctx_destroy(QueryCtx* c) {
  if (!c) return;
  if (c->stream) cudaStreamDestroy(c->stream);
  if (c->ev0) cudaEventDestroy(c->ev0);
  if (c->ev1) cudaEventDestroy(c->ev1);
  cudaFree(c->d_query); cudaFree(c->d_excl); cudaFree(c->d_cand); cudaFree(c->d_list_max); cudaFree(c->d_counter); cudaFree(c->d_finish_seq);
  cudaFree(c->d_main_index); cudaFree(c->d_worklist); cudaFree(c->d_worklist_ctl); cudaFree(c->d_out_n);
  cudaFree(c->d_tmp[0]); cudaFree(c->d_tmp[1]);
  cudaFreeHost(c->h_query); cudaFreeHost(c->h_excl); cudaFreeHost(c->h_out_n);
}

static int32_t ctx_create(Model* m, QueryCtx* c) {
  const int fpad = m->store[ORYX_Y].fpad;
  ORYX_CUDA(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
  ORYX_CUDA(cudaEventCreate(&c->ev0));
  ORYX_CUDA(cudaEventCreate(&c->ev1));
  ORYX_CUDA(cudaMalloc(&c->d_query, (size_t)fpad * sizeof(float)));
  c->excl_cap = 4096;
  ORYX_CUDA(cudaMalloc(&c->d_excl, (size_t)c->excl_cap * sizeof(int32_t)));
  c->scratch_sets = 2;
  ORYX_CUDA(cudaMalloc(&c->d_cand, (size_t)2 * kMaxGrid * kMaxK * sizeof(uint2)));
  ORYX_CUDA(cudaMalloc(&c->d_list_max, (size_t)2 * kMaxGrid * sizeof(uint32_t)));
  ORYX_CUDA(cudaMalloc(&c->d_counter, 32 * sizeof(unsigned int)));
  ORYX_CUDA(cudaMemset(c->d_counter, 0, 32 * sizeof(unsigned int)));
  ORYX_CUDA(cudaMalloc(&c->d_finish_seq, sizeof(unsigned long long)));
  ORYX_CUDA(cudaMemset(c->d_finish_seq, 0, sizeof(unsigned long long)));
  ORYX_CUDA(cudaMalloc(&c->d_main_index, sizeof(int32_t)));
  ORYX_CUDA(cudaMemset(c->d_main_index, 0, sizeof(int32_t)));
  // one packed result record: [count i32, pad][rows i32 x how_many, padded to 8][scores f64 x how_many]
  // so a query's result comes back in a single device-to-host copy
  ORYX_CUDA(cudaMalloc(&c->d_out_n, (size_t)kMaxK * 12 + 16));
  c->d_out_rows = nullptr;
  c->d_out_scores = nullptr;
  ORYX_CUDA(cudaMallocHost(&c->h_query, (size_t)fpad * sizeof(float)));
  ORYX_CUDA(cudaMallocHost(&c->h_excl, (size_t)c->excl_cap * sizeof(int32_t)));
  ORYX_CUDA(cudaMallocHost(&c->h_out_n, (size_t)kMaxK * 12 + 16));
  c->h_out_rows = nullptr;
  c->h_out_scores = nullptr;
  return ORYX_OK;
}

int32_t ctx_acquire(Model* m, QueryCtx** out) {
  std::unique_lock<std::mutex> lk(m->pool_mu);
  for (;;) {
    if (!m->pool_free.empty()) {
      *out = m->pool_free.back();
      m->pool_free.pop_back();
      return ORYX_OK;
    }
    if ((int)m->pool_all.size() < m->pool_limit) {
      auto c = std::make_unique<QueryCtx>();
      int32_t rc = ctx_create(m, c.get());
      if (rc != ORYX_OK) { ctx_destroy(c.get()); return rc; }
      *out = c.get();
      m->pool_all.push_back(std::move(c));
      return ORYX_OK;
    }
    m->pool_cv.wait(lk);
  }
}

void ctx_release(Model* m, QueryCtx* ctx) {
  {
    std::lock_guard<std::mutex> lk(m->pool_mu);
    m->pool_free.push_back(ctx);
  }
  m->pool_cv.notify_one();
}

int32_t ensure_max_dynamic_smem(const void* kernel, size_t bytes) {
  struct Entry { int device; const void* kernel; size_t bytes; };
  static std::mutex mu;
  static std::vector<Entry> done;
  int device = 0;
  ORYX_CUDA(cudaGetDevice(&device));
  std::lock_guard<std::mutex> lk(mu);
  for (auto& e : done) {
    if (e.device == device && e.kernel == kernel) {
      if (e.bytes >= bytes) return ORYX_OK;
      ORYX_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
      e.bytes = bytes;
      return ORYX_OK;
    }
  }
  ORYX_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
  done.push_back({device, kernel, bytes});
  return ORYX_OK;
}

int32_t ensure_excl(QueryCtx* c, int32_t n) {
  if (n <= c->excl_cap) return ORYX_OK;
  int32_t cap = std::max(n, c->excl_cap * 2);
  cudaFree(c->d_excl); cudaFreeHost(c->h_excl);
  c->d_excl = nullptr; c->h_excl = nullptr; c->excl_cap = 0;
  ORYX_CUDA(cudaMalloc(&c->d_excl, (size_t)cap * sizeof(int32_t)));
  ORYX_CUDA(cudaMallocHost(&c->h_excl, (size_t)cap * sizeof(int32_t)));
  c->excl_cap = cap;
  return ORYX_OK;
}

// Scratch slot `i` of a context, at least `bytes` large.  Buffers up to kTmpKeepBytes stay with the context
// (the point lookups of /estimate and /similarityToItem reuse them request after request); a larger one
// (oryx_scores_all over millions of rows) is released again by ctx_tmp_trim at the end of the call.
constexpr size_t kTmpKeepBytes = (size_t)32 << 20;
static int32_t ctx_tmp(QueryCtx* c, int i, size_t bytes, void** out) {
  if (c->tmp_bytes[i] < bytes) {
    cudaFree(c->d_tmp[i]);
    c->d_tmp[i] = nullptr;
    c->tmp_bytes[i] = 0;
    if (cudaMalloc(&c->d_tmp[i], bytes) != cudaSuccess) {
      cudaGetLastError();
      set_error("device allocation of %zu bytes failed", bytes);
      return ORYX_E_NOMEM;
    }
    c->tmp_bytes[i] = bytes;
  }
  *out = c->d_tmp[i];
  return ORYX_OK;
}
static void ctx_tmp_trim(QueryCtx* c) {
  for (int i = 0; i < 2; i++) {
    if (c->tmp_bytes[i] > kTmpKeepBytes) {
      cudaFree(c->d_tmp[i]);
      c->d_tmp[i] = nullptr;
      c->tmp_bytes[i] = 0;
    }
  }
}

struct DeviceGuard {
  int prev = -1;
  bool ok = false;
  explicit DeviceGuard(int dev) {
    if (cudaGetDevice(&prev) != cudaSuccess) { cudaGetLastError(); prev = -1; }
    ok = cudaSetDevice(dev) == cudaSuccess;
    if (!ok) cudaGetLastError();
  }
  ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

#define ORYX_DEVICE(m)                                                   \
  DeviceGuard guard__((m)->device);                                      \
  if (!guard__.ok) {                                                     \
    set_error("cannot select CUDA device %d (no CPU fallback exists)", (m)->device); \
    return ORYX_E_CUDA;                                                  \
  }

// Second half of a query: the query vector and the sorted exclusion slots already sit in device
// memory (d_query; c->d_excl[0..n_excl_slots)).  Launch, copy the packed record back, wait.
static int32_t ctx_tmp(QueryCtx* c, int i, size_t bytes, void** out);
static void ctx_tmp_trim(QueryCtx* c);

// how_many beyond the fused scan's limit: exact scores of every row, radix select, sort (topn_deep.cu).
static int32_t run_topn_deep(Model* m, QueryCtx* c, int32_t kind, const float* d_query, int32_t n_excl_slots,
                             int32_t how_many, int32_t* out_rows, double* out_scores, int32_t* out_n) {
  Store* s = &m->store[ORYX_Y];
  const size_t work = (deep_scratch_bytes(s->n_slots, how_many) + 255) & ~(size_t)255;
  const size_t scores_off = work, rows_off = scores_off + (size_t)how_many * 8, n_off = rows_off + (((size_t)how_many * 4 + 15) & ~(size_t)15);
  unsigned char* buf = nullptr;
  ORYX_TRY(ctx_tmp(c, 0, n_off + 16, (void**)&buf));
  ORYX_CUDA(cudaEventRecord(c->ev0, c->stream));
  int32_t rc = topn_deep_launch(m, c, kind, d_query, how_many, c->d_excl, n_excl_slots, buf,
                                reinterpret_cast<int32_t*>(buf + rows_off), reinterpret_cast<double*>(buf + scores_off),
                                reinterpret_cast<int32_t*>(buf + n_off), c->stream);
  if (rc == ORYX_OK) {
    int32_t n = 0;
    cudaEventRecord(c->ev1, c->stream);
    if (cudaMemcpyAsync(&n, buf + n_off, sizeof(n), cudaMemcpyDeviceToHost, c->stream) != cudaSuccess ||
        cudaStreamSynchronize(c->stream) != cudaSuccess || n < 0 || n > how_many) {
      set_error("deep top-N failed: %s", cudaGetErrorString(cudaGetLastError()));
      rc = ORYX_E_CUDA;
    } else {
      if (n > 0 && (cudaMemcpy(out_rows, buf + rows_off, (size_t)n * 4, cudaMemcpyDeviceToHost) != cudaSuccess ||
                    cudaMemcpy(out_scores, buf + scores_off, (size_t)n * 8, cudaMemcpyDeviceToHost) != cudaSuccess)) {
        set_error("deep top-N copy failed: %s", cudaGetErrorString(cudaGetLastError()));
        rc = ORYX_E_CUDA;
      }
      *out_n = n;
      float ms = 0.f;
      if (cudaEventElapsedTime(&ms, c->ev0, c->ev1) == cudaSuccess) m->last_query_ms.store((double)ms);
      else cudaGetLastError();
    }
  }
  ctx_tmp_trim(c);
  return rc;
}

static int32_t run_topn_staged(Model* m, QueryCtx* c, int32_t kind, const float* d_query, int32_t n_excl_slots,
                               int32_t how_many, int32_t* out_rows, double* out_scores, int32_t* out_n) {
  if (how_many > ORYX_MAX_HOW_MANY) return run_topn_deep(m, c, kind, d_query, n_excl_slots, how_many, out_rows, out_scores, out_n);
  // packed result record (see ctx_create)
  const size_t rows_off = 8, scores_off = 8 + (((size_t)how_many * 4 + 7) & ~(size_t)7);
  const size_t rec_bytes = scores_off + (size_t)how_many * 8;
  unsigned char* d_rec = reinterpret_cast<unsigned char*>(c->d_out_n);
  unsigned char* h_rec = reinterpret_cast<unsigned char*>(c->h_out_n);
  TopnRequest rq{};
  rq.kind = kind;
  rq.d_query = d_query;
  rq.how_many = how_many;
  rq.d_excl = c->d_excl;
  rq.n_excl = n_excl_slots;
  rq.d_out_n = reinterpret_cast<int32_t*>(d_rec);
  rq.d_out_rows = reinterpret_cast<int32_t*>(d_rec + rows_off);
  rq.d_out_scores = reinterpret_cast<double*>(d_rec + scores_off);
  ORYX_CUDA(cudaEventRecord(c->ev0, c->stream));
  ORYX_TRY(scan_topn_launch(m, c, rq, c->stream));
  ORYX_CUDA(cudaEventRecord(c->ev1, c->stream));
  ORYX_CUDA(cudaMemcpyAsync(h_rec, d_rec, rec_bytes, cudaMemcpyDeviceToHost, c->stream));
  ORYX_CUDA(cudaStreamSynchronize(c->stream));
  int32_t n = *reinterpret_cast<int32_t*>(h_rec);
  if (n < 0 || n > how_many) {
    set_error("device returned an impossible result count %d", n);
    return ORYX_E_CUDA;
  }
  std::memcpy(out_rows, h_rec + rows_off, (size_t)n * sizeof(int32_t));
  std::memcpy(out_scores, h_rec + scores_off, (size_t)n * sizeof(double));
  *out_n = n;
  float ms = 0.f;
  if (cudaEventElapsedTime(&ms, c->ev0, c->ev1) == cudaSuccess) m->last_query_ms.store((double)ms);
  else cudaGetLastError();
  return ORYX_OK;
}

// One query through a context: stage, launch, copy back, wait.
static int32_t run_topn(Model* m, QueryCtx* c, int32_t kind, const float* query, int32_t how_many,
                        const int32_t* excluded_rows, int32_t n_excluded, int32_t* out_rows,
                        double* out_scores, int32_t* out_n) {
  Store* s = &m->store[ORYX_Y];
  const int fpad = s->fpad, F = s->features;
  std::memcpy(c->h_query, query, (size_t)F * sizeof(float));
  for (int j = F; j < fpad; j++) c->h_query[j] = 0.0f;
  ORYX_CUDA(cudaMemcpyAsync(c->d_query, c->h_query, (size_t)fpad * sizeof(float), cudaMemcpyHostToDevice, c->stream));
  int32_t n_excl_slots = 0;
  if (n_excluded > 0) {
    // The allowed-predicate arrives as row ids; the kernel filters by slot, so translate here (the
    // caller holds the model's shared lock, row2slot is stable) and sort for the binary search.
    ORYX_TRY(ensure_excl(c, n_excluded));
    for (int32_t i = 0; i < n_excluded; i++) {
      int32_t r = excluded_rows[i];
      if (r >= 0 && (size_t)r < s->row2slot.size() && s->row2slot[r] >= 0) c->h_excl[n_excl_slots++] = s->row2slot[r];
    }
    std::sort(c->h_excl, c->h_excl + n_excl_slots);
    if (n_excl_slots > 0) {
      ORYX_CUDA(cudaMemcpyAsync(c->d_excl, c->h_excl, (size_t)n_excl_slots * sizeof(int32_t), cudaMemcpyHostToDevice, c->stream));
    }
  }
  return run_topn_staged(m, c, kind, c->d_query, n_excl_slots, how_many, out_rows, out_scores, out_n);
}

int32_t async_mark(Model* m, cudaStream_t stream) {
  ORYX_CUDA(cudaEventRecord(m->async_done, stream));  // the event exists since oryx_model_create
  m->async_pending.store(true, std::memory_order_release);
  return ORYX_OK;
}

int32_t async_fence(Model* m) {
  if (!m->async_pending.load(std::memory_order_acquire)) return ORYX_OK;
  // exclusive lock held: no stream-ordered launch can be in progress, the event is the latest one
  ORYX_CUDA(cudaStreamWaitEvent(m->update_stream, m->async_done, 0));
  m->async_pending.store(false, std::memory_order_release);
  return ORYX_OK;
}

int32_t topn_device_impl(Model* m, int kind, const float* d_query, int how_many, int32_t* d_out_rows,
                         double* d_out_scores, int32_t* d_out_n, cudaStream_t stream,
                         const unsigned long long* gate, unsigned long long gate_value) {
  // The asynchronous form uses one dedicated context (scratch + ticket counter): the caller orders
  // these calls on `stream`, so they never overlap each other.
  {
    std::unique_lock<std::mutex> lk(m->pool_mu);
    if (m->async_ctx == nullptr) {
      auto c = std::make_unique<QueryCtx>();
      int32_t rc = ctx_create(m, c.get());
      if (rc != ORYX_OK) { ctx_destroy(c.get()); return rc; }
      m->async_ctx = c.get();
      m->pool_all.push_back(std::move(c));
    }
  }
  QueryCtx* async_ctx = m->async_ctx;
  TopnRequest rq{};
  rq.kind = kind;
  rq.d_query = d_query;
  rq.how_many = how_many;
  rq.d_excl = nullptr;
  rq.n_excl = 0;
  rq.d_out_rows = d_out_rows;
  rq.d_out_scores = d_out_scores;
  rq.d_out_n = d_out_n;
  rq.gate = gate;
  rq.gate_value = gate_value;
  std::shared_lock<std::shared_mutex> lk(m->mu);
  ORYX_TRY(scan_topn_launch(m, async_ctx, rq, stream));
  return async_mark(m, stream);
}

}  // namespace oryx

using namespace oryx;

extern "C" {

int32_t oryx_last_error(char* buf, int32_t buf_len) {
  if (!buf || buf_len <= 0) return ORYX_E_INVALID;
  std::snprintf(buf, (size_t)buf_len, "%s", last_error());
  return ORYX_OK;
}

int32_t oryx_lsh_plan(double sample_rate, int32_t num_cores, int32_t* num_hashes, int32_t* max_bits_differing) {
  ORYX_ARG(num_hashes && max_bits_differing, "null output");
  ORYX_ARG(sample_rate > 0.0 && sample_rate <= 1.0, "sampleRate must be in (0,1]");
  ORYX_ARG(num_cores >= 1, "numCores must be >= 1");
  int h = 0, b = 0;
  lsh_plan(sample_rate, num_cores, &h, &b);
  *num_hashes = h;
  *max_bits_differing = b;
  return ORYX_OK;
}

int32_t oryx_model_create(const oryx_config* config, oryx_model** out) {
  ORYX_ARG(config && out, "null argument");
  ORYX_ARG(config->features > 0, "features must be positive");
  ORYX_ARG(config->num_hashes >= 0 && config->num_hashes <= ORYX_MAX_HASHES, "num_hashes out of range");
  ORYX_ARG(config->max_bits_differing >= 0 && config->max_bits_differing <= config->num_hashes, "max_bits_differing out of range");
  ORYX_ARG(config->num_hashes == 0 || config->hash_vectors != nullptr, "hash_vectors required when num_hashes > 0");
  *out = nullptr;
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count == 0) {
    cudaGetLastError();
    set_error("no usable CUDA device (%s); this library has no CPU fallback",
              e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
    return ORYX_E_CUDA;
  }
  ORYX_ARG(config->device >= 0 && config->device < count, "device ordinal out of range");
  auto m = std::make_unique<Model>();
  m->cfg = *config;
  m->cfg.hash_vectors = nullptr;
  m->device = config->device;
  m->num_hashes = config->num_hashes;
  m->max_bits = config->max_bits_differing;
  ORYX_DEVICE(m.get());
  cudaDeviceProp prop{};
  ORYX_CUDA(cudaGetDeviceProperties(&prop, m->device));
  if (prop.major != 10) {
    set_error("device %d is sm_%d%d; this library is built for sm_100a only", m->device, prop.major, prop.minor);
    return ORYX_E_CUDA;
  }
  m->num_sms = prop.multiProcessorCount;
  ORYX_CUDA(cudaStreamCreateWithFlags(&m->update_stream, cudaStreamNonBlocking));
  ORYX_CUDA(cudaEventCreateWithFlags(&m->async_done, cudaEventDisableTiming));
  const int F = config->features, fpad = round_up(F, 4);
  if (m->num_hashes > 0) {
    m->h_hash_vectors.assign((size_t)m->num_hashes * fpad, 0.0f);
    for (int h = 0; h < m->num_hashes; h++) {
      std::memcpy(&m->h_hash_vectors[(size_t)h * fpad], config->hash_vectors + (size_t)h * F, (size_t)F * sizeof(float));
    }
    ORYX_CUDA(cudaMalloc(&m->d_hash_vectors, m->h_hash_vectors.size() * sizeof(float)));
    ORYX_CUDA(cudaMemcpy(m->d_hash_vectors, m->h_hash_vectors.data(), m->h_hash_vectors.size() * sizeof(float), cudaMemcpyHostToDevice));
  }
  ORYX_TRY(store_init(m.get(), ORYX_X, 0));
  ORYX_TRY(store_init(m.get(), ORYX_Y, config->capacity_hint));
  m->coalescer = coalescer_create();
  *out = reinterpret_cast<oryx_model*>(m.release());
  return ORYX_OK;
}

int32_t oryx_model_destroy(oryx_model* mh) {
  if (!mh) return ORYX_OK;
  Model* m = reinterpret_cast<Model*>(mh);
  {
    DeviceGuard g(m->device);
    cudaDeviceSynchronize();
    for (auto& c : m->pool_all) ctx_destroy(c.get());
    store_free(&m->store[0]);
    store_free(&m->store[1]);
    known_free(&m->known);
    coalescer_destroy(m->coalescer);
    m->coalescer = nullptr;
    cudaFree(m->staging.d_vec); cudaFree(m->staging.d_i32); cudaFreeHost(m->staging.h_i32);
    cudaFree(m->staging.d_known_dst); cudaFree(m->staging.d_known_val);
    cudaFree(m->d_hash_vectors);
    cudaFree(m->d_gram); cudaFree(m->d_gram_scratch);
    for (auto& b : m->batch.buf) cudaFree(b);
    for (auto& e : m->batch.ev) if (e) cudaEventDestroy(e);
    cudaFree(m->d_debug_ts);
    cudaFree(m->d_share_desc);
    if (m->async_done) cudaEventDestroy(m->async_done);
    if (m->update_stream) cudaStreamDestroy(m->update_stream);
    cudaGetLastError();
  }
  delete m;
  return ORYX_OK;
}

int32_t oryx_items_upsert(oryx_model* mh, int32_t which, const int32_t* rows, const float* vecs, int32_t n,
                          int32_t* out_partitions) {
  ORYX_ARG(mh, "null model");
  ORYX_ARG(which == ORYX_X || which == ORYX_Y, "which must be ORYX_X or ORYX_Y");
  ORYX_ARG(n >= 0, "n must be >= 0");
  if (n == 0) return ORYX_OK;
  ORYX_ARG(rows && vecs, "null rows/vecs");
  Model* m = reinterpret_cast<Model*>(mh);
  ORYX_DEVICE(m);
  std::unique_lock<std::shared_mutex> lk(m->mu);
  ORYX_TRY(async_fence(m));
  return store_upsert(m, which, rows, vecs, n, out_partitions);
}

int32_t oryx_items_upsert_device(oryx_model* mh, int32_t which, int32_t first_row, const float* d_vecs, int64_t n) {
  ORYX_ARG(mh, "null model");
  ORYX_ARG(which == ORYX_X || which == ORYX_Y, "which must be ORYX_X or ORYX_Y");
  ORYX_ARG(n >= 0 && first_row >= 0, "n and first_row must be >= 0");
  ORYX_ARG((int64_t)first_row + n <= 0x7fffffffLL, "row ids must fit int32");
  if (n == 0) return ORYX_OK;
  ORYX_ARG(d_vecs, "null d_vecs");
  Model* m = reinterpret_cast<Model*>(mh);
  ORYX_DEVICE(m);
  std::unique_lock<std::shared_mutex> lk(m->mu);
  ORYX_TRY(async_fence(m));
  return store_upsert_device(m, which, first_row, d_vecs, n);
}

int32_t oryx_items_remove(oryx_model* mh, int32_t which, const int32_t* rows, int32_t n) {
  ORYX_ARG(mh, "null model");
  ORYX_ARG(which == ORYX_X || which == ORYX_Y, "which must be ORYX_X or ORYX_Y");
  ORYX_ARG(n >= 0, "n must be >= 0");
  if (n == 0) return ORYX_OK;
  ORYX_ARG(rows, "null rows");
  Model* m = reinterpret_cast<Model*>(mh);
  ORYX_DEVICE(m);
  std::unique_lock<std::shared_mutex> lk(m->mu);
  ORYX_TRY(async_fence(m));
  return store_remove(m, which, rows, n);
}

int32_t oryx_items_get(oryx_model* mh, int32_t which, int32_t row, float* out_vec, int32_t* found) {
  ORYX_ARG(mh && out_vec && found, "null argument");
  ORYX_ARG(which == ORYX_X || which == ORYX_Y, "which must be ORYX_X or ORYX_Y");
  Model* m = reinterpret_cast<Model*>(mh);
  ORYX_DEVICE(m);
  std::shared_lock<std::shared_mutex> lk(m->mu);
  return store_get(m, which, row, out_vec, found);
}

int32_t oryx_items_compact(oryx_model* mh, int32_t which) {
  ORYX_ARG(mh, "null model");
  ORYX_ARG(which == ORYX_X || which == ORYX_Y, "which must be ORYX_X or ORYX_Y");
  Model* m = reinterpret_cast<Model*>(mh);
  ORYX_DEVICE(m);
  std::unique_lock<std::shared_mutex> lk(m->mu);
  ORYX_TRY(async_fence(m));
  return store_compact(m, which);
}

int32_t oryx_lsh_index_rows(oryx_model* mh, const float* vecs, int32_t n, int32_t* out_indices) {
  ORYX_ARG(mh && out_indices, "null argument");
  ORYX_ARG(n >= 0, "n must be >= 0");
  if (n == 0) return ORYX_OK;
  ORYX_ARG(vecs, "null vecs");
  Model* m = reinterpret_cast<Model*>(mh);
  ORYX_DEVICE(m);
  const int F = m->cfg.features;
  float* d_in = nullptr;
  int32_t* d_out = nullptr;
  QueryCtx* c = nullptr;
  ORYX_TRY(ctx_acquire(m, &c));
  int32_t rc = ORYX_OK;
  do {
    if ((rc = ctx_tmp(c, 0, (size_t)n * F * sizeof(float), (void**)&d_in)) != ORYX_OK ||
        (rc = ctx_tmp(c, 1, (size_t)n * sizeof(int32_t), (void**)&d_out)) != ORYX_OK)
      break;
    if (cudaMemcpyAsync(d_in, vecs, (size_t)n * F * sizeof(float), cudaMemcpyHostToDevice, c->stream) != cudaSuccess) { rc = ORYX_E_CUDA; set_error("H2D copy failed"); break; }
    rc = lsh_hash_rows_device(m, d_in, n, F, d_out, c->stream);
    if (rc != ORYX_OK) break;
    if (cudaMemcpyAsync(out_indices, d_out, (size_t)n * sizeof(int32_t), cudaMemcpyDeviceToHost, c->stream) != cudaSuccess ||
        cudaStreamSynchronize(c->stream) != cudaSuccess) {
      set_error("LSH hashing failed: %s", cudaGetErrorString(cudaGetLastError()));
      rc = ORYX_E_CUDA;
    }
  } while (0);
  ctx_tmp_trim(c);
  ctx_release(m, c);
  return rc;
}

int32_t oryx_lsh_index(oryx_model* mh, const float* vec, int32_t* out_index) {
  return oryx_lsh_index_rows(mh, vec, 1, out_index);
}

int32_t oryx_lsh_candidates(oryx_model* mh, const float* vec, int32_t* out, int32_t* out_n) {
  ORYX_ARG(mh && vec && out && out_n, "null argument");
  Model* m = reinterpret_cast<Model*>(mh);
  ORYX_DEVICE(m);
  const int F = m->cfg.features, fpad = round_up(F, 4);
  const int total = 1 << m->num_hashes;
  QueryCtx* c = nullptr;
  ORYX_TRY(ctx_acquire(m, &c));
  int32_t* d_out = nullptr;
  int32_t rc = ORYX_OK;
  do {
    if ((rc = ctx_tmp(c, 0, (size_t)(total + 1) * sizeof(int32_t), (void**)&d_out)) != ORYX_OK) break;
    std::memcpy(c->h_query, vec, (size_t)F * sizeof(float));
    for (int j = F; j < fpad; j++) c->h_query[j] = 0.f;
    if (cudaMemcpyAsync(c->d_query, c->h_query, (size_t)fpad * sizeof(float), cudaMemcpyHostToDevice, c->stream) != cudaSuccess) { rc = ORYX_E_CUDA; set_error("H2D copy failed"); break; }
    rc = lsh_hash_rows_device(m, c->d_query, 1, fpad, c->d_main_index, c->stream);
    if (rc != ORYX_OK) break;
    rc = lsh_candidates_device(m, c->d_main_index, d_out, d_out + total, c->stream);
    if (rc != ORYX_OK) break;
    std::vector<int32_t> host((size_t)total + 1);
    if (cudaMemcpyAsync(host.data(), d_out, (size_t)(total + 1) * sizeof(int32_t), cudaMemcpyDeviceToHost, c->stream) != cudaSuccess ||
        cudaStreamSynchronize(c->stream) != cudaSuccess) {
      set_error("candidate enumeration failed: %s", cudaGetErrorString(cudaGetLastError()));
      rc = ORYX_E_CUDA;
      break;
    }
    int32_t n = host[(size_t)total];
    std::memcpy(out, host.data(), (size_t)n * sizeof(int32_t));
    *out_n = n;
  } while (0);
  ctx_release(m, c);
  return rc;
}

int32_t oryx_topn(oryx_model* mh, int32_t kind, const float* query, int32_t how_many,
                  const int32_t* excluded_rows, int32_t n_excluded,
                  int32_t* out_rows, double* out_scores, int32_t* out_n) {
  ORYX_ARG(mh && query && out_rows && out_scores && out_n, "null argument");
  ORYX_ARG(n_excluded >= 0 && (n_excluded == 0 || excluded_rows), "bad exclusion list");
  ORYX_ARG(how_many >= 1 && how_many <= ORYX_MAX_HOW_MANY_DEEP, "how_many out of range [1, ORYX_MAX_HOW_MANY_DEEP]");
  Model* m = reinterpret_cast<Model*>(mh);
  ORYX_DEVICE(m);
  QueryCtx* c = nullptr;
  ORYX_TRY(ctx_acquire(m, &c));
  int32_t rc;
  {
    std::shared_lock<std::shared_mutex> lk(m->mu);
    rc = run_topn(m, c, kind, query, how_many, excluded_rows, n_excluded, out_rows, out_scores, out_n);
  }
  if (rc != ORYX_OK) cudaStreamSynchronize(c->stream);
  ctx_release(m, c);
  return rc;
}

// ---- known items on the device (known.cu)

int32_t oryx_known_items_add(oryx_model* mh, const int32_t* user_rows, const int64_t* begin, const int32_t* item_rows,
                             int32_t n_users) {
  ORYX_ARG(mh, "null model");
  ORYX_ARG(n_users >= 0, "n_users must be >= 0");
  if (n_users == 0) return ORYX_OK;
  ORYX_ARG(user_rows && begin, "null argument");
  ORYX_ARG(begin[n_users] == begin[0] || item_rows, "null item_rows");
  for (int32_t i = 0; i < n_users; i++) ORYX_ARG(begin[i + 1] >= begin[i], "begin[] must be non-decreasing");
  Model* m = reinterpret_cast<Model*>(mh);
  ORYX_DEVICE(m);
  std::unique_lock<std::shared_mutex> lk(m->mu);
  ORYX_TRY(async_fence(m));
  return known_add(m, user_rows, begin, item_rows, n_users);
}

int32_t oryx_known_items_get(oryx_model* mh, int32_t user_row, int32_t* out_rows, int32_t cap, int32_t* out_n) {
  ORYX_ARG(mh && out_n, "null argument");
  ORYX_ARG(cap >= 0 && (cap == 0 || out_rows), "bad output buffer");
  Model* m = reinterpret_cast<Model*>(mh);
  std::shared_lock<std::shared_mutex> lk(m->mu);
  *out_n = 0;
  if (user_row < 0 || (size_t)user_row >= m->known.lists.size()) return ORYX_OK;
  const std::vector<int32_t>& list = m->known.lists[(size_t)user_row];
  *out_n = (int32_t)list.size();
  if ((int32_t)list.size() > cap) {
    set_error("known-items list of user row %d has %d entries, buffer holds %d", user_row, (int)list.size(), cap);
    return ORYX_E_NOMEM;
  }
  std::copy(list.begin(), list.end(), out_rows);
  return ORYX_OK;
}

int32_t oryx_known_items_retain(oryx_model* mh, const uint8_t* keep_users, int32_t n_user_rows,
                                const uint8_t* keep_items, int32_t n_item_rows) {
  ORYX_ARG(mh, "null model");
  ORYX_ARG(n_user_rows >= 0 && n_item_rows >= 0, "row counts must be >= 0");
  Model* m = reinterpret_cast<Model*>(mh);
  ORYX_DEVICE(m);
  std::unique_lock<std::shared_mutex> lk(m->mu);
  ORYX_TRY(async_fence(m));
  return known_retain(m, keep_users, n_user_rows, keep_items, n_item_rows);
}

int32_t oryx_known_counts(oryx_model* mh, int32_t which, int32_t* out_counts, int32_t n_rows) {
  ORYX_ARG(mh, "null model");
  ORYX_ARG(which == ORYX_X || which == ORYX_Y, "which must be ORYX_X or ORYX_Y");
  ORYX_ARG(n_rows >= 0, "n_rows must be >= 0");
  if (n_rows == 0) return ORYX_OK;
  ORYX_ARG(out_counts, "null output");
  Model* m = reinterpret_cast<Model*>(mh);
  ORYX_DEVICE(m);
  // the count kernel runs on the update stream: exclusive, like any other use of that stream
  std::unique_lock<std::shared_mutex> lk(m->mu);
  ORYX_TRY(async_fence(m));
  return known_counts(m, which, out_counts, n_rows);
}

int32_t oryx_recommend(oryx_model* mh, int32_t user_row, int32_t how_many, int32_t consider_known_items,
                       int32_t* out_rows, double* out_scores, int32_t* out_n, int32_t* found) {
  ORYX_ARG(mh && out_rows && out_scores && out_n && found, "null argument");
  ORYX_ARG(how_many >= 1 && how_many <= ORYX_MAX_HOW_MANY_DEEP, "how_many out of range [1, ORYX_MAX_HOW_MANY_DEEP]");
  Model* m = reinterpret_cast<Model*>(mh);
  ORYX_DEVICE(m);
  *out_n = 0;
  *found = 0;
  QueryCtx* c = nullptr;
  ORYX_TRY(ctx_acquire(m, &c));
  int32_t rc = ORYX_OK;
  {
    std::shared_lock<std::shared_mutex> lk(m->mu);
    Store* x = &m->store[ORYX_X];
    const int32_t slot = (user_row >= 0 && (size_t)user_row < x->row2slot.size()) ? x->row2slot[(size_t)user_row] : -1;
    if (slot >= 0) {
      *found = 1;
      // the query vector is the user's row of the device-resident X store (Recommend.java:84-90);
      // both stores pad rows to the same fpad
      const float* d_query = x->d_vec + (int64_t)slot * x->fpad;
      int32_t n_excl = 0;
      if (!consider_known_items) rc = known_exclusions_device(m, c, user_row, &n_excl, c->stream);
      if (rc == ORYX_OK) rc = run_topn_staged(m, c, ORYX_DOT, d_query, n_excl, how_many, out_rows, out_scores, out_n);
    }
  }
  if (rc != ORYX_OK) cudaStreamSynchronize(c->stream);
  ctx_release(m, c);
  return rc;
}

int32_t oryx_topn_batch(oryx_model* mh, int32_t kind, const float* queries, int32_t nq, int32_t how_many,
                        const int32_t* excl_offsets, const int32_t* excl_rows,
                        int32_t* out_rows, double* out_scores, int32_t* out_n) {
  ORYX_ARG(mh && out_rows && out_scores && out_n, "null argument");
  ORYX_ARG(nq >= 0, "nq must be >= 0");
  if (nq == 0) return ORYX_OK;
  ORYX_ARG(queries, "null queries");
  ORYX_ARG(how_many >= 1 && how_many <= ORYX_MAX_HOW_MANY, "how_many out of range [1, ORYX_MAX_HOW_MANY]");
  Model* m = reinterpret_cast<Model*>(mh);
  ORYX_DEVICE(m);
  const int F = m->cfg.features;
  QueryCtx* c = nullptr;
  ORYX_TRY(ctx_acquire(m, &c));
  int32_t rc = ORYX_OK;
  {
    std::shared_lock<std::shared_mutex> lk(m->mu);
    Store* s = &m->store[ORYX_Y];
    bool batched = false;
    if (batch_topn_supported(m, nq)) {
      // exclusion rows -> sorted slots, CSR per query
      std::vector<int32_t> eo, es;
      if (excl_offsets && excl_rows) {
        eo.resize((size_t)nq + 1, 0);
        for (int32_t i = 0; i < nq; i++) {
          size_t begin = es.size();
          for (int32_t j = excl_offsets[i]; j < excl_offsets[i + 1]; j++) {
            int32_t r = excl_rows[j];
            if (r >= 0 && (size_t)r < s->row2slot.size() && s->row2slot[r] >= 0) es.push_back(s->row2slot[r]);
          }
          std::sort(es.begin() + begin, es.end());
          eo[i + 1] = (int32_t)es.size();
        }
      }
      rc = batch_topn_run(m, kind, queries, nq, how_many, eo.empty() ? nullptr : eo.data(),
                          es.empty() ? nullptr : es.data(), out_rows, out_scores, out_n, c->stream, &m->last_batch);
      batched = rc == ORYX_OK;
      if (rc == ORYX_E_NOMEM) rc = ORYX_OK;  // not enough scratch for the tensor-core path: answer one by one
    }
    for (int32_t i = 0; i < nq && rc == ORYX_OK; i++) {
      if (batched && out_n[i] >= 0) continue;  // -1 marks a query the batched path handed back
      const int32_t* ex = nullptr;
      int32_t nex = 0;
      if (excl_offsets && excl_rows) {
        ex = excl_rows + excl_offsets[i];
        nex = excl_offsets[i + 1] - excl_offsets[i];
      }
      rc = run_topn(m, c, kind, queries + (size_t)i * F, how_many, ex, nex, out_rows + (size_t)i * how_many,
                    out_scores + (size_t)i * how_many, out_n + i);
    }
  }
  if (rc != ORYX_OK) cudaStreamSynchronize(c->stream);
  ctx_release(m, c);
  return rc;
}

int32_t oryx_recommend_batch(oryx_model* mh, const int32_t* user_rows, int32_t n_users, int32_t how_many,
                             int32_t consider_known_items, int32_t* out_rows, double* out_scores, int32_t* out_n) {
  ORYX_ARG(mh && out_rows && out_scores && out_n, "null argument");
  ORYX_ARG(n_users >= 0, "n_users must be >= 0");
  if (n_users == 0) return ORYX_OK;
  ORYX_ARG(user_rows, "null user_rows");
  ORYX_ARG(how_many >= 1 && how_many <= ORYX_MAX_HOW_MANY, "how_many out of range [1, ORYX_MAX_HOW_MANY]");
  Model* m = reinterpret_cast<Model*>(mh);
  ORYX_DEVICE(m);
  QueryCtx* c = nullptr;
  ORYX_TRY(ctx_acquire(m, &c));
  int32_t rc = ORYX_OK;
  {
    std::shared_lock<std::shared_mutex> lk(m->mu);
    Store* x = &m->store[ORYX_X];
    Store* y = &m->store[ORYX_Y];
    std::vector<int32_t> qslots((size_t)n_users);
    for (int32_t i = 0; i < n_users; i++) {
      const int32_t r = user_rows[i];
      qslots[(size_t)i] = (r >= 0 && (size_t)r < x->row2slot.size()) ? x->row2slot[(size_t)r] : -1;
    }
    bool batched = false;
    if (batch_topn_supported(m, n_users)) {
      // the allowed-predicate of every user as sorted slots, CSR per query, from the authoritative lists
      std::vector<int32_t> eo, es;
      if (!consider_known_items) {
        eo.assign((size_t)n_users + 1, 0);
        for (int32_t i = 0; i < n_users; i++) {
          const int32_t r = user_rows[i];
          const size_t begin = es.size();
          if (r >= 0 && (size_t)r < m->known.lists.size()) {
            for (int32_t item : m->known.lists[(size_t)r]) {
              if ((size_t)item < y->row2slot.size() && y->row2slot[(size_t)item] >= 0) es.push_back(y->row2slot[(size_t)item]);
            }
          }
          std::sort(es.begin() + (std::ptrdiff_t)begin, es.end());
          eo[(size_t)i + 1] = (int32_t)es.size();
        }
        if (es.empty()) eo.clear();
      }
      rc = batch_topn_run(m, ORYX_DOT, nullptr, n_users, how_many, eo.empty() ? nullptr : eo.data(),
                          es.empty() ? nullptr : es.data(), out_rows, out_scores, out_n, c->stream, &m->last_batch,
                          qslots.data(), ORYX_X);
      batched = rc == ORYX_OK;
      if (rc == ORYX_E_NOMEM) rc = ORYX_OK;  // not enough scratch for the tensor-core path: answer one by one
    }
    for (int32_t i = 0; i < n_users && rc == ORYX_OK; i++) {
      if (qslots[(size_t)i] < 0) { out_n[i] = -1; continue; }  // no such user (the reference's 404)
      if (batched && out_n[i] >= 0) continue;                  // -1 marks a query the batched path handed back
      const float* d_query = x->d_vec + (int64_t)qslots[(size_t)i] * x->fpad;
      int32_t n_excl = 0;
      if (!consider_known_items) rc = known_exclusions_device(m, c, user_rows[i], &n_excl, c->stream);
      if (rc == ORYX_OK)
        rc = run_topn_staged(m, c, ORYX_DOT, d_query, n_excl, how_many, out_rows + (size_t)i * how_many,
                             out_scores + (size_t)i * how_many, out_n + i);
    }
  }
  if (rc != ORYX_OK) cudaStreamSynchronize(c->stream);
  ctx_release(m, c);
  return rc;
}

// Development aid (not in the public header): phase timings (ms) and flops of the last tensor-core batch.
int32_t oryx_debug_batch_timings(oryx_model* mh, double* out5) {  // out5 holds 9 doubles
  Model* m = reinterpret_cast<Model*>(mh);
  out5[0] = m->last_batch.convert_ms;
  out5[1] = m->last_batch.score_ms;
  out5[2] = m->last_batch.threshold_ms;
  out5[3] = m->last_batch.refine_ms;
  out5[4] = m->last_batch.flops;
  out5[5] = m->last_batch.cand_mean;
  out5[6] = m->last_batch.cand_max;
  out5[7] = m->last_batch.runs;
  out5[8] = m->last_batch.shadow_rebuilds;
  return ORYX_OK;
}

int32_t oryx_scores_all(oryx_model* mh, int32_t kind, const float* query, int32_t n_rows, double* out_scores) {
  ORYX_ARG(mh && query && out_scores, "null argument");
  ORYX_ARG(kind == ORYX_DOT || kind == ORYX_COSINE, "kind must be ORYX_DOT or ORYX_COSINE");
  ORYX_ARG(n_rows >= 0, "n_rows must be >= 0");
  if (n_rows == 0) return ORYX_OK;
  Model* m = reinterpret_cast<Model*>(mh);
  ORYX_DEVICE(m);
  Store* s = &m->store[ORYX_Y];
  QueryCtx* c = nullptr;
  ORYX_TRY(ctx_acquire(m, &c));
  double* d_out = nullptr;
  int32_t rc = ORYX_OK;
  do {
    if ((rc = ctx_tmp(c, 0, (size_t)n_rows * sizeof(double), (void**)&d_out)) != ORYX_OK) break;
    std::memcpy(c->h_query, query, (size_t)s->features * sizeof(float));
    for (int j = s->features; j < s->fpad; j++) c->h_query[j] = 0.f;
    std::shared_lock<std::shared_mutex> lk(m->mu);
    if (cudaMemcpyAsync(c->d_query, c->h_query, (size_t)s->fpad * sizeof(float), cudaMemcpyHostToDevice, c->stream) != cudaSuccess) { rc = ORYX_E_CUDA; set_error("H2D copy failed"); break; }
    rc = scores_all_launch(m, kind, c->d_query, n_rows, d_out, c->stream);
    if (rc != ORYX_OK) break;
    if (cudaMemcpyAsync(out_scores, d_out, (size_t)n_rows * sizeof(double), cudaMemcpyDeviceToHost, c->stream) != cudaSuccess ||
        cudaStreamSynchronize(c->stream) != cudaSuccess) {
      set_error("scores_all failed: %s", cudaGetErrorString(cudaGetLastError()));
      rc = ORYX_E_CUDA;
    }
  } while (0);
  ctx_tmp_trim(c);
  ctx_release(m, c);
  return rc;
}

int32_t oryx_estimate_pairs(oryx_model* mh, const int32_t* user_rows, const int32_t* item_rows, int64_t n,
                            int32_t wide_products, double* out_scores, uint8_t* out_found) {
  ORYX_ARG(mh && out_scores, "null argument");
  ORYX_ARG(n >= 0, "n must be >= 0");
  if (n == 0) return ORYX_OK;
  ORYX_ARG(user_rows && item_rows, "null rows");
  Model* m = reinterpret_cast<Model*>(mh);
  ORYX_DEVICE(m);
  QueryCtx* c = nullptr;
  ORYX_TRY(ctx_acquire(m, &c));
  int32_t* d_slots = nullptr;
  double* d_out = nullptr;
  int32_t rc = ORYX_OK;
  do {
    std::shared_lock<std::shared_mutex> lk(m->mu);
    const Store* sx = &m->store[ORYX_X];
    const Store* sy = &m->store[ORYX_Y];
    std::vector<int32_t> slots((size_t)n * 2);
    for (int64_t i = 0; i < n; i++) {
      const int32_t u = user_rows[i], it = item_rows[i];
      slots[(size_t)i] = (u >= 0 && (size_t)u < sx->row2slot.size()) ? sx->row2slot[(size_t)u] : -1;
      slots[(size_t)(n + i)] = (it >= 0 && (size_t)it < sy->row2slot.size()) ? sy->row2slot[(size_t)it] : -1;
      if (out_found) out_found[i] = slots[(size_t)i] >= 0 && slots[(size_t)(n + i)] >= 0;
    }
    if ((rc = ctx_tmp(c, 0, (size_t)n * 2 * sizeof(int32_t), (void**)&d_slots)) != ORYX_OK ||
        (rc = ctx_tmp(c, 1, (size_t)n * sizeof(double), (void**)&d_out)) != ORYX_OK)
      break;
    if (cudaMemcpyAsync(d_slots, slots.data(), (size_t)n * 2 * sizeof(int32_t), cudaMemcpyHostToDevice, c->stream) != cudaSuccess) {
      set_error("H2D copy failed"); rc = ORYX_E_CUDA; break;
    }
    rc = pair_scores_launch(m, d_slots, d_slots + n, n, wide_products ? 1 : 0, d_out, c->stream);
    if (rc != ORYX_OK) break;
    if (cudaMemcpyAsync(out_scores, d_out, (size_t)n * sizeof(double), cudaMemcpyDeviceToHost, c->stream) != cudaSuccess ||
        cudaStreamSynchronize(c->stream) != cudaSuccess) {
      set_error("estimate_pairs failed: %s", cudaGetErrorString(cudaGetLastError()));
      rc = ORYX_E_CUDA;
    }
  } while (0);
  ctx_tmp_trim(c);
  ctx_release(m, c);
  return rc;
}

int32_t oryx_scores_rows(oryx_model* mh, int32_t kind, const float* query, const int32_t* rows, int32_t n,
                         double* out_scores, uint8_t* found) {
  ORYX_ARG(mh && query && out_scores, "null argument");
  ORYX_ARG(kind == ORYX_DOT || kind == ORYX_COSINE, "kind must be ORYX_DOT or ORYX_COSINE");
  ORYX_ARG(n >= 0, "n must be >= 0");
  if (n == 0) return ORYX_OK;
  ORYX_ARG(rows, "null rows");
  Model* m = reinterpret_cast<Model*>(mh);
  ORYX_DEVICE(m);
  Store* s = &m->store[ORYX_Y];
  QueryCtx* c = nullptr;
  ORYX_TRY(ctx_acquire(m, &c));
  int32_t* d_slots = nullptr;
  double* d_out = nullptr;
  int32_t rc = ORYX_OK;
  do {
    std::shared_lock<std::shared_mutex> lk(m->mu);
    std::vector<int32_t> slots((size_t)n);
    for (int32_t i = 0; i < n; i++) {
      int32_t r = rows[i];
      slots[i] = (r >= 0 && (size_t)r < s->row2slot.size()) ? s->row2slot[r] : -1;
      if (found) found[i] = slots[i] >= 0;
    }
    if ((rc = ctx_tmp(c, 0, (size_t)n * sizeof(int32_t), (void**)&d_slots)) != ORYX_OK ||
        (rc = ctx_tmp(c, 1, (size_t)n * sizeof(double), (void**)&d_out)) != ORYX_OK)
      break;
    std::memcpy(c->h_query, query, (size_t)s->features * sizeof(float));
    for (int j = s->features; j < s->fpad; j++) c->h_query[j] = 0.f;
    if (cudaMemcpyAsync(c->d_query, c->h_query, (size_t)s->fpad * sizeof(float), cudaMemcpyHostToDevice, c->stream) != cudaSuccess ||
        cudaMemcpyAsync(d_slots, slots.data(), (size_t)n * sizeof(int32_t), cudaMemcpyHostToDevice, c->stream) != cudaSuccess) {
      set_error("H2D copy failed"); rc = ORYX_E_CUDA; break;
    }
    rc = scores_rows_launch(m, kind, c->d_query, d_slots, n, d_out, c->stream);
    if (rc != ORYX_OK) break;
    if (cudaMemcpyAsync(out_scores, d_out, (size_t)n * sizeof(double), cudaMemcpyDeviceToHost, c->stream) != cudaSuccess ||
        cudaStreamSynchronize(c->stream) != cudaSuccess) {
      set_error("scores_rows failed: %s", cudaGetErrorString(cudaGetLastError()));
      rc = ORYX_E_CUDA;
    }
  } while (0);
  ctx_tmp_trim(c);
  ctx_release(m, c);
  return rc;
}

int32_t oryx_topn_device(oryx_model* mh, int32_t kind, const float* d_query, int32_t how_many,
                         int32_t* d_out_rows, double* d_out_scores, int32_t* d_out_n, void* stream) {
  ORYX_ARG(mh && d_query && d_out_rows && d_out_scores && d_out_n, "null argument");
  ORYX_ARG(how_many >= 1 && how_many <= ORYX_MAX_HOW_MANY, "how_many out of range [1, ORYX_MAX_HOW_MANY]");
  Model* m = reinterpret_cast<Model*>(mh);
  ORYX_DEVICE(m);
  return topn_device_impl(m, kind, d_query, how_many, d_out_rows, d_out_scores, d_out_n,
                          reinterpret_cast<cudaStream_t>(stream), nullptr, 0);
}

int32_t oryx_merge_topn_device(const double* d_scores, const int32_t* d_rows, const int32_t* d_n, int32_t n_lists,
                               int32_t how_many, int64_t row_stride, int64_t* d_out_rows, double* d_out_scores,
                               int32_t* d_out_n, void* stream) {
  ORYX_ARG(d_scores && d_rows && d_n && d_out_rows && d_out_scores && d_out_n, "null argument");
  ORYX_ARG(n_lists >= 1 && n_lists <= 1024, "n_lists out of range [1, 1024]");
  ORYX_ARG(how_many >= 1 && how_many <= ORYX_MAX_HOW_MANY, "how_many out of range [1, ORYX_MAX_HOW_MANY]");
  return merge_topn_launch(d_scores, d_rows, d_n, n_lists, how_many, row_stride, d_out_rows, d_out_scores,
                           d_out_n, reinterpret_cast<cudaStream_t>(stream));
}

int32_t oryx_merge_topn_batch_device(const double* d_scores, const int32_t* d_rows, const int32_t* d_n, int32_t n_lists,
                                     int32_t nq, int32_t how_many, int64_t row_stride, int64_t* d_out_rows,
                                     double* d_out_scores, int32_t* d_out_n, void* stream) {
  ORYX_ARG(d_scores && d_rows && d_n && d_out_rows && d_out_scores && d_out_n, "null argument");
  ORYX_ARG(n_lists >= 1 && n_lists <= 1024, "n_lists out of range [1, 1024]");
  ORYX_ARG(nq >= 0, "nq must be >= 0");
  ORYX_ARG(how_many >= 1 && how_many <= ORYX_MAX_HOW_MANY, "how_many out of range [1, ORYX_MAX_HOW_MANY]");
  if (nq == 0) return ORYX_OK;
  return merge_topn_batch_launch(d_scores, d_rows, d_n, n_lists, nq, how_many, row_stride, d_out_rows, d_out_scores, d_out_n,
                                 reinterpret_cast<cudaStream_t>(stream));
}

int32_t oryx_merge_topn_packed(const void* d_records, int32_t n_lists, int32_t record_bytes, int32_t how_many,
                               int64_t row_stride, int64_t* d_out_rows, double* d_out_scores, int32_t* d_out_n,
                               void* stream) {
  ORYX_ARG(d_records && d_out_rows && d_out_scores && d_out_n, "null argument");
  ORYX_ARG(n_lists >= 1 && n_lists <= 1024, "n_lists out of range [1, 1024]");
  ORYX_ARG(how_many >= 1 && how_many <= ORYX_MAX_HOW_MANY, "how_many out of range [1, ORYX_MAX_HOW_MANY]");
  ORYX_ARG(record_bytes >= how_many * 12 + 4 && record_bytes % 8 == 0, "record_bytes too small or not a multiple of 8");
  return merge_topn_packed_launch(d_records, n_lists, record_bytes, how_many, row_stride, d_out_rows, d_out_scores,
                                  d_out_n, reinterpret_cast<cudaStream_t>(stream));
}

// Development aid (not part of the public header): arm == 1 allocates/re-arms the timestamp buffer the
// scan kernel fills; arm == 0 synchronises and copies out 8 globaltimer values (ns):
// [0] earliest CTA start, [1]/[2] earliest/latest end of a CTA's scan, [3] merge start,
// [4] selection done, [5] exact re-score done, [6] merge end; [8]-[10] inside the small-k selection: list maxima
// loaded, cut found, candidates gathered.  `out` holds 16 values.
int32_t oryx_debug_scan_timestamps(oryx_model* mh, int32_t arm, unsigned long long* out) {
  Model* m = reinterpret_cast<Model*>(mh);
  ORYX_DEVICE(m);
  if (arm) {
    if (!m->d_debug_ts) ORYX_CUDA(cudaMalloc(&m->d_debug_ts, 16 * sizeof(unsigned long long)));
    unsigned long long init[16] = {~0ull, ~0ull, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
    ORYX_CUDA(cudaDeviceSynchronize());
    ORYX_CUDA(cudaMemcpy(m->d_debug_ts, init, sizeof(init), cudaMemcpyHostToDevice));
    return ORYX_OK;
  }
  ORYX_ARG(out && m->d_debug_ts, "not armed");
  ORYX_CUDA(cudaDeviceSynchronize());
  ORYX_CUDA(cudaMemcpy(out, m->d_debug_ts, 16 * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
  return ORYX_OK;
}

int32_t oryx_stats_get(oryx_model* mh, oryx_stats* out) {
  ORYX_ARG(mh && out, "null argument");
  Model* m = reinterpret_cast<Model*>(mh);
  std::shared_lock<std::shared_mutex> lk(m->mu);
  std::memset(out, 0, sizeof(*out));
  for (int w = 0; w < 2; w++) {
    out->rows[w] = m->store[w].live;
    out->slots[w] = m->store[w].n_slots;
    out->device_bytes += m->store[w].device_bytes();
  }
  out->queries = m->queries.load();
  out->kernel_launches = m->launches.load();
  out->last_query_ms = m->last_query_ms.load();
  return ORYX_OK;
}

}  // extern "C"
