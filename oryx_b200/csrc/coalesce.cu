// coalesce.cu -- oryx_topn_coalesced: concurrent topN callers answered as one tensor-core batch.
//
// The reference serves up to 400 request threads at once (framework/oryx-lambda-serving/.../
// ServingLayer.java:225), each of which calls ALSServingModel.topN (ALSServingModel.java:289-335) on its
// own.  One fp32 scan answers one of them in 0.73 ms; one pass of the batched path over the bf16 shadow
// of Y (batch_topn.cu) answers a whole page of them in less than that, with the same exact results.
// This entry point has the signature and the semantics of oryx_topn; callers that arrive while a batch
// is in flight queue up and are answered together by the next batch ("group commit"): the first caller
// to find no batch running becomes the leader, takes every compatible request in the queue, runs
// oryx_topn_batch, hands the results back and wakes the others.  Nobody waits for a timer: an isolated
// caller runs at once, as a batch of one.
#include <algorithm>
#include <deque>

#include "common.cuh"

namespace oryx {

struct CoalesceRequest {
  int32_t kind, how_many;
  int32_t user_row = -1;        // >= 0: a /recommend by user row (oryx_recommend), query / exclusions unused
  int32_t consider_known = 0;
  int32_t* found = nullptr;
  const float* query;
  const int32_t* excluded_rows;
  int32_t n_excluded;
  int32_t* out_rows;
  double* out_scores;
  int32_t* out_n;
  int32_t status = ORYX_OK;
  char error[256] = {0};
  bool done = false;
};

struct Coalescer {
  std::mutex mu;
  std::condition_variable cv;
  std::deque<CoalesceRequest*> queue;
  bool leader_active = false;
};

namespace {
constexpr int kMaxPage = 1024;  // requests answered by one batch

// Answers `page` (same kind and how_many) with one oryx_topn_batch call; fills status / outputs.
void run_page(oryx_model* mh, const std::vector<CoalesceRequest*>& page) {
  Model* m = reinterpret_cast<Model*>(mh);
  // pages keep coming: even a short one may (re)build the bf16 shadow rather than scan once per query
  struct SmallPages { SmallPages() { tl_batch_small_pages_ok = true; } ~SmallPages() { tl_batch_small_pages_ok = false; } } small_pages;
  const int F = m->cfg.features;
  const int nq = (int)page.size();
  const int hm = page[0]->how_many;
  if (page[0]->user_row >= 0) {  // a page of /recommend requests: oryx_recommend_batch
    std::vector<int32_t> users((size_t)nq), rows((size_t)nq * hm), counts((size_t)nq, 0);
    std::vector<double> scores((size_t)nq * hm);
    for (int i = 0; i < nq; i++) users[(size_t)i] = page[i]->user_row;
    const int32_t st = oryx_recommend_batch(mh, users.data(), nq, hm, page[0]->consider_known, rows.data(),
                                            scores.data(), counts.data());
    char err[256] = {0};
    if (st != ORYX_OK) oryx_last_error(err, sizeof(err));
    for (int i = 0; i < nq; i++) {
      CoalesceRequest* r = page[i];
      r->status = st;
      if (st != ORYX_OK) { std::memcpy(r->error, err, sizeof(err)); continue; }
      const int n = counts[(size_t)i];
      *r->found = n >= 0 ? 1 : 0;
      *r->out_n = std::max(n, 0);
      if (n > 0) {
        std::memcpy(r->out_rows, &rows[(size_t)i * hm], (size_t)n * sizeof(int32_t));
        std::memcpy(r->out_scores, &scores[(size_t)i * hm], (size_t)n * sizeof(double));
      }
    }
    return;
  }
  std::vector<float> queries((size_t)nq * F);
  std::vector<int32_t> offsets((size_t)nq + 1, 0), excl;
  bool any_excl = false;
  for (int i = 0; i < nq; i++) {
    std::memcpy(&queries[(size_t)i * F], page[i]->query, (size_t)F * sizeof(float));
    if (page[i]->n_excluded > 0) {
      any_excl = true;
      excl.insert(excl.end(), page[i]->excluded_rows, page[i]->excluded_rows + page[i]->n_excluded);
    }
    offsets[(size_t)i + 1] = (int32_t)excl.size();
  }
  std::vector<int32_t> rows((size_t)nq * hm), counts((size_t)nq, 0);
  std::vector<double> scores((size_t)nq * hm);
  static const int32_t none = 0;
  const int32_t st = oryx_topn_batch(mh, page[0]->kind, queries.data(), nq, hm, any_excl ? offsets.data() : nullptr,
                                     any_excl ? (excl.empty() ? &none : excl.data()) : nullptr, rows.data(),
                                     scores.data(), counts.data());
  char err[256] = {0};
  if (st != ORYX_OK) oryx_last_error(err, sizeof(err));  // the leader's thread-local message
  for (int i = 0; i < nq; i++) {
    CoalesceRequest* r = page[i];
    r->status = st;
    if (st == ORYX_OK) {
      const int n = counts[(size_t)i];
      std::memcpy(r->out_rows, &rows[(size_t)i * hm], (size_t)n * sizeof(int32_t));
      std::memcpy(r->out_scores, &scores[(size_t)i * hm], (size_t)n * sizeof(double));
      *r->out_n = n;
    } else {
      std::memcpy(r->error, err, sizeof(err));
    }
  }
}
}  // namespace

Coalescer* coalescer_create() { return new Coalescer(); }
void coalescer_destroy(Coalescer* c) { delete c; }

// Queue the request; lead a page if nobody does; return once the request has been answered.
int32_t coalesce_submit(oryx_model* mh, Coalescer* c, CoalesceRequest* req) {
  std::unique_lock<std::mutex> lk(c->mu);
  c->queue.push_back(req);
  for (;;) {
    if (req->done) break;
    if (c->leader_active) {  // a batch is in flight: sleep until it (or a later one) has answered us
      c->cv.wait(lk);
      continue;
    }
    // become the leader for one page: everything queued that is compatible with the oldest request
    c->leader_active = true;
    std::vector<CoalesceRequest*> page;
    const CoalesceRequest* first = c->queue.front();
    const bool by_user = first->user_row >= 0;
    for (auto it = c->queue.begin(); it != c->queue.end() && (int)page.size() < kMaxPage;) {
      const CoalesceRequest* r = *it;
      const bool same = r->how_many == first->how_many && ((r->user_row >= 0) == by_user) &&
                        (by_user ? r->consider_known == first->consider_known : r->kind == first->kind);
      if (same) {
        page.push_back(*it);
        it = c->queue.erase(it);
      } else {
        ++it;
      }
    }
    lk.unlock();
    run_page(mh, page);
    lk.lock();
    for (CoalesceRequest* r : page) r->done = true;
    c->leader_active = false;
    c->cv.notify_all();  // the answered ones leave; one of the others leads the next page
  }
  lk.unlock();
  if (req->status != ORYX_OK) set_error("%s", req->error);
  return req->status;
}

// Queueing behind a leader only pays when the leader can answer the page in one tensor-core pass.  When
// it cannot (LSH filtering on, more than 256 features, a small store) the page would be answered one
// scan at a time on a single context -- slower than the same threads scanning side by side on the
// context pool -- so those callers bypass the queue.
static bool combining_pays(Model* m) {
  std::shared_lock<std::shared_mutex> lk(m->mu);
  return batch_topn_possible(m);
}

}  // namespace oryx

using namespace oryx;

extern "C" int32_t oryx_topn_coalesced(oryx_model* mh, int32_t kind, const float* query, int32_t how_many,
                                       const int32_t* excluded_rows, int32_t n_excluded, int32_t* out_rows,
                                       double* out_scores, int32_t* out_n) {
  ORYX_ARG(mh && query && out_rows && out_scores && out_n, "null argument");
  ORYX_ARG(kind == ORYX_DOT || kind == ORYX_COSINE, "kind must be ORYX_DOT or ORYX_COSINE");
  ORYX_ARG(n_excluded >= 0 && (n_excluded == 0 || excluded_rows), "bad exclusion list");
  ORYX_ARG(how_many >= 1 && how_many <= ORYX_MAX_HOW_MANY, "how_many out of range [1, ORYX_MAX_HOW_MANY]");
  Model* m = reinterpret_cast<Model*>(mh);
  if (!combining_pays(m))  // no page could ride on the tensor cores: concurrent callers use the context pool instead
    return oryx_topn(mh, kind, query, how_many, excluded_rows, n_excluded, out_rows, out_scores, out_n);
  CoalesceRequest req;
  req.kind = kind;
  req.how_many = how_many;
  req.query = query;
  req.excluded_rows = excluded_rows;
  req.n_excluded = n_excluded;
  req.out_rows = out_rows;
  req.out_scores = out_scores;
  req.out_n = out_n;
  return coalesce_submit(mh, m->coalescer, &req);
}

extern "C" int32_t oryx_recommend_coalesced(oryx_model* mh, int32_t user_row, int32_t how_many,
                                            int32_t consider_known_items, int32_t* out_rows, double* out_scores,
                                            int32_t* out_n, int32_t* found) {
  ORYX_ARG(mh && out_rows && out_scores && out_n && found, "null argument");
  ORYX_ARG(how_many >= 1 && how_many <= ORYX_MAX_HOW_MANY, "how_many out of range [1, ORYX_MAX_HOW_MANY]");
  Model* m = reinterpret_cast<Model*>(mh);
  *out_n = 0;
  *found = 0;
  if (user_row < 0) return ORYX_OK;  // no such user
  if (!combining_pays(m)) return oryx_recommend(mh, user_row, how_many, consider_known_items, out_rows, out_scores, out_n, found);
  CoalesceRequest req;
  req.kind = ORYX_DOT;
  req.how_many = how_many;
  req.user_row = user_row;
  req.consider_known = consider_known_items ? 1 : 0;
  req.found = found;
  req.query = nullptr;
  req.excluded_rows = nullptr;
  req.n_excluded = 0;
  req.out_rows = out_rows;
  req.out_scores = out_scores;
  req.out_n = out_n;
  return coalesce_submit(mh, m->coalescer, &req);
}
