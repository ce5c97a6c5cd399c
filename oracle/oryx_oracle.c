/*
 * oryx_oracle.c -- CPU restatement of Oryx 2's ALS serving/speed hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oryx_b200/ (the product) may link,
 * import or call this file; only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs use it, and there only as the checker or
 * the timed CPU arm.
 *
 * Parity status: PINNED.  Every function below restates one reference function
 * (cited file:line, relative to the reference root) and tests/test_oracle_golden.py
 * replays the reference's own golden vectors against it (tests/golden/*.json,
 * transcribed from the reference's JUnit tests, each with its source citation).
 *
 * The reference is Java.  Java's `float * float` is one IEEE-754 binary32
 * multiply and `double += float` is one binary64 add, evaluated strictly in
 * source order (strictfp semantics are the default for these scalar loops on
 * every 64-bit JVM).  Compile this file with -ffp-contract=off and without
 * -ffast-math so gcc emits exactly that sequence; then results are bit-identical.
 *
 * Third-party arithmetic that is NOT in the reference tree:
 *   org.apache.commons:commons-math3:3.6 (reference pom.xml:214-215) --
 *   RRQRDecomposition / QRDecomposition.Solver / AbstractRealMatrix.getNorm /
 *   CombinatoricsUtils.binomialCoefficient.  Their published algorithms are
 *   restated below (Householder QR with column pivoting on the largest column
 *   2-norm, max-absolute-column-sum norm).  The results at that boundary are
 *   pinned by the reference's own LinearSystemSolverTest / ALSUtilsTest /
 *   EstimateForAnonymousTest goldens.
 *
 * Short path names used in citations:
 *   VM   = framework/oryx-common/src/main/java/com/cloudera/oryx/common/math/VectorMath.java
 *   LSS  = framework/oryx-common/src/main/java/com/cloudera/oryx/common/math/LinearSystemSolver.java
 *   SOL  = framework/oryx-common/src/main/java/com/cloudera/oryx/common/math/Solver.java
 *   LSH  = app/oryx-app-serving/src/main/java/com/cloudera/oryx/app/serving/als/model/LocalitySensitiveHash.java
 *   TNC  = app/oryx-app-serving/src/main/java/com/cloudera/oryx/app/serving/als/model/TopNConsumer.java
 *   ASM  = app/oryx-app-serving/src/main/java/com/cloudera/oryx/app/serving/als/model/ALSServingModel.java
 *   DOTS = app/oryx-app-serving/src/main/java/com/cloudera/oryx/app/serving/als/DotsFunction.java
 *   COS  = app/oryx-app-serving/src/main/java/com/cloudera/oryx/app/serving/als/CosineAverageFunction.java
 *   ALSU = app/oryx-app-common/src/main/java/com/cloudera/oryx/app/als/ALSUtils.java
 *   PAIRS= framework/oryx-common/src/main/java/com/cloudera/oryx/common/collection/Pairs.java
 */
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define ORO_MAX_HASHES 16 /* LSH:33 MAX_HASHES */

/* ------------------------------------------------------------------ VectorMath */

/* VM:38-45  dot(): float*float product (binary32), widened, added to a double in index order. */
double oro_dot(const float* x, const float* y, int n) {
  double dot = 0.0;
  for (int i = 0; i < n; i++) {
    float p = x[i] * y[i];
    dot += (double)p;
  }
  return dot;
}

/* VM:52-58  norm(): sqrt of sum of float squares accumulated in double. */
double oro_norm(const float* x, int n) {
  double total = 0.0;
  for (int i = 0; i < n; i++) {
    float p = x[i] * x[i];
    total += (double)p;
  }
  return sqrt(total);
}

/* VM:64-83  transposeTimesSelf(): for each row vector v, G[r][c] += (double)(v[r]*v[c]),
 * rows visited in the given order.  `out` is f*f doubles, row-major, zeroed here. */
void oro_transpose_times_self(const float* m, int64_t rows, int f, double* out) {
  memset(out, 0, sizeof(double) * (size_t)f * (size_t)f);
  for (int64_t i = 0; i < rows; i++) {
    const float* v = m + i * (int64_t)f;
    for (int r = 0; r < f; r++) {
      float rv = v[r];
      double* o = out + (size_t)r * f;
      for (int c = 0; c < f; c++) {
        float p = rv * v[c];
        o[c] += (double)p;
      }
    }
  }
}

/* ------------------------------------------------------------------ LSH */

/* commons-math3 CombinatoricsUtils.binomialCoefficient (exact for the small n used here). */
static int64_t binom(int n, int k) {
  if (k < 0 || k > n) return 0;
  if (k > n - k) k = n - k;
  int64_t r = 1;
  for (int i = 1; i <= k; i++) r = r * (n - k + i) / i;
  return r;
}

/* LSH:45-74  sizing loop of the LocalitySensitiveHash constructor. */
void oro_lsh_plan(double sample_rate, int num_cores, int* out_hashes, int* out_bits) {
  int num_hashes = 0;
  int bits_differing = 0;
  for (; num_hashes < ORO_MAX_HASHES; num_hashes++) {
    bits_differing = 0;
    int64_t num_partitions_to_try = 1;
    while (bits_differing < num_hashes && num_partitions_to_try < num_cores) {
      bits_differing++;
      num_partitions_to_try += binom(num_hashes, bits_differing);
    }
    if (bits_differing == num_hashes && num_partitions_to_try < num_cores) {
      continue;
    }
    if ((double)num_partitions_to_try <= sample_rate * (double)((int64_t)1 << num_hashes)) {
      break;
    }
  }
  *out_hashes = num_hashes;
  *out_bits = bits_differing;
}

/* LSH:142-150  getIndexFor(): bit i set iff dot(hash_i, v) > 0.0 (exact zero / NaN -> 0). */
int oro_lsh_index(const float* hash_vectors, int num_hashes, int f, const float* v) {
  int index = 0;
  for (int i = 0; i < num_hashes; i++) {
    if (oro_dot(hash_vectors + (size_t)i * f, v, f) > 0.0) {
      index |= 1 << i;
    }
  }
  return index;
}

/* LSH:110-117  candidateIndicesPrototype: all 2^H ints ordered by popcount, ascending inside a band. */
void oro_lsh_prototype(int num_hashes, int* out) {
  int n = 1 << num_hashes;
  int offset[ORO_MAX_HASHES + 2];
  offset[0] = 0;
  for (int i = 1; i <= num_hashes; i++) {
    offset[i] = offset[i - 1] + (int)binom(num_hashes, i - 1);
  }
  for (int i = 0; i < n; i++) {
    out[offset[__builtin_popcount((unsigned)i)]++] = i;
  }
}

/* LSH:156-177  getCandidateIndices() given the main index.  Returns count written to `out`
 * (capacity 2^H). */
int oro_lsh_candidates(int num_hashes, int max_bits_differing, int main_index, int* out) {
  int n = 1 << num_hashes;
  if (num_hashes == max_bits_differing) { /* :160-162 allIndices */
    for (int i = 0; i < n; i++) out[i] = i;
    return n;
  }
  if (max_bits_differing == 0) { /* :163-165 */
    out[0] = main_index;
    return 1;
  }
  int how_many = 0;
  for (int i = 0; i <= max_bits_differing; i++) how_many += (int)binom(num_hashes, i);
  int* proto = (int*)malloc(sizeof(int) * (size_t)n);
  oro_lsh_prototype(num_hashes, proto);
  for (int i = 0; i < how_many; i++) out[i] = proto[i] ^ main_index;
  free(proto);
  return how_many;
}

/* Bulk form of getIndexFor over a row-major matrix (what setItemVector does per row, ASM:163). */
void oro_lsh_index_rows(const float* hash_vectors, int num_hashes, int f, const float* rows,
                        int64_t n, int32_t* out) {
  for (int64_t i = 0; i < n; i++) {
    out[i] = oro_lsh_index(hash_vectors, num_hashes, f, rows + i * (int64_t)f);
  }
}

/* ------------------------------------------------------------------ score functions */

/* DOTS:34-45  mean of user vectors: float += float, then float /= int (int widened to float). */
void oro_dots_mean(const float* vecs, int n, int f, float* out) {
  for (int i = 0; i < f; i++) out[i] = 0.0f;
  for (int v = 0; v < n; v++) {
    for (int i = 0; i < f; i++) out[i] += vecs[(size_t)v * f + i];
  }
  for (int i = 0; i < f; i++) out[i] /= (float)n;
}

/* COS:29-37 (n==1, no final divide) and COS:39-51 (n>1): each vec[i]/norm is float/double ->
 * double; `float += double` compound assignment computes (float)((double)acc + quotient).
 * The single-vector constructor does NOT divide by the count. */
void oro_cosine_average(const float* vecs, int n, int f, int single_ctor, float* out) {
  for (int i = 0; i < f; i++) out[i] = 0.0f;
  for (int v = 0; v < n; v++) {
    const float* vec = vecs + (size_t)v * f;
    double vec_norm = oro_norm(vec, f);
    for (int i = 0; i < f; i++) {
      out[i] = (float)((double)out[i] + (double)vec[i] / vec_norm);
    }
  }
  if (!single_ctor) {
    for (int i = 0; i < f; i++) out[i] /= (float)n;
  }
}

/* COS:54-56  applyAsDouble(): dot(q, y) / norm(y), double divide. */
double oro_cosine_score(const float* q, const float* y, int f) {
  return oro_dot(q, y, f) / oro_norm(y, f);
}

enum { ORO_DOT = 0, ORO_COSINE = 1 };

static inline double score_of(int kind, const float* q, const float* y, int f) {
  return kind == ORO_COSINE ? oro_cosine_score(q, y, f) : oro_dot(q, y, f);
}

/* All scores for a row-major matrix (the DotsFunction / CosineAverageFunction applied per row). */
void oro_scores_all(int kind, const float* q, const float* y, int64_t n, int f, double* out) {
  for (int64_t i = 0; i < n; i++) out[i] = score_of(kind, q, y + i * (int64_t)f, f);
}

/* ------------------------------------------------------------------ TopNConsumer */

typedef struct {
  double score;
  int64_t row;
  int64_t seq; /* arrival order; only used to make the final sort stable */
} oro_pair;

typedef struct {
  oro_pair* heap; /* binary min-heap on score, as java.util.PriorityQueue with ASCENDING order (TNC:49) */
  int size;
  int how_many;
  double lower_bound; /* TNC:41 topScoreLowerBound */
} oro_topn;

static void heap_sift_up(oro_pair* h, int k) {
  oro_pair x = h[k];
  while (k > 0) {
    int parent = (k - 1) >> 1;
    if (!(x.score < h[parent].score)) break;
    h[k] = h[parent];
    k = parent;
  }
  h[k] = x;
}

static void heap_sift_down(oro_pair* h, int n, int k) {
  oro_pair x = h[k];
  int half = n >> 1;
  while (k < half) {
    int child = 2 * k + 1;
    int right = child + 1;
    if (right < n && h[right].score < h[child].score) child = right;
    if (!(h[child].score < x.score)) break;
    h[k] = h[child];
    k = child;
  }
  h[k] = x;
}

static void topn_init(oro_topn* t, int how_many) {
  t->heap = (oro_pair*)malloc(sizeof(oro_pair) * (size_t)(how_many > 0 ? how_many : 1));
  t->size = 0;
  t->how_many = how_many;
  t->lower_bound = -INFINITY; /* TNC:53 */
}

/* TNC:56-74  accept() after the allowed-predicate and rescore steps: strict '>' against the cached
 * lower bound, re-check against the live minimum when full, replace the minimum.  NaN never enters. */
static void topn_accept(oro_topn* t, int64_t row, double score, int64_t seq) {
  if (score > t->lower_bound) {
    if (t->size >= t->how_many) {
      t->lower_bound = t->heap[0].score;
      if (score > t->lower_bound) {
        t->heap[0].score = score;
        t->heap[0].row = row;
        t->heap[0].seq = seq;
        heap_sift_down(t->heap, t->size, 0);
      }
    } else {
      t->heap[t->size].score = score;
      t->heap[t->size].row = row;
      t->heap[t->size].seq = seq;
      heap_sift_up(t->heap, t->size);
      t->size++;
    }
  }
}

/* PAIRS:48-54 orderBySecond(DESCENDING) via Double.compareTo, made total with arrival order
 * (Java's sort is stable over an unspecified concatenation order; ours is partition order). */
static int cmp_desc(const void* a, const void* b) {
  const oro_pair* x = (const oro_pair*)a;
  const oro_pair* y = (const oro_pair*)b;
  if (x->score > y->score) return -1;
  if (x->score < y->score) return 1;
  if (x->seq < y->seq) return -1;
  if (x->seq > y->seq) return 1;
  return 0;
}

static int is_excluded(const int32_t* excl, int n_excl, int64_t row) {
  int lo = 0, hi = n_excl - 1;
  while (lo <= hi) {
    int mid = (lo + hi) >> 1;
    if (excl[mid] == row) return 1;
    if (excl[mid] < row) lo = mid + 1; else hi = mid - 1;
  }
  return 0;
}

/*
 * ASM:289-335  topN(): one TopNConsumer per non-empty candidate partition, visited in candidate
 * order; concat; sort DESC; limit howMany.
 *   y          row-major [n x f]
 *   part       LSH partition of each row (NULL => everything in partition 0), < 0 => row absent
 *   cand/n_cand candidate partition list (NULL => all rows are candidates)
 *   excl       sorted row ids the allowedPredicate rejects (known items), may be NULL
 * Returns the number of results written (<= how_many).
 */
int oro_topn_rows(int kind, const float* q, const float* y, int64_t n, int f,
                  const int32_t* part, const int32_t* cand, int n_cand,
                  const int32_t* excl, int n_excl, int how_many,
                  int64_t* out_rows, double* out_scores) {
  if (how_many <= 0 || n <= 0) return 0;
  int n_parts = 1;
  if (part != NULL) {
    for (int64_t i = 0; i < n; i++) if (part[i] + 1 > n_parts) n_parts = part[i] + 1;
  }
  int one = 0;
  const int32_t* cands = cand;
  int nc = n_cand;
  if (cands == NULL) {
    if (part == NULL) { cands = &one; nc = 1; }
  }
  /* Bucket rows by partition (what the per-partition FeatureVectors maps hold, ASM:69). */
  int64_t* start = (int64_t*)calloc((size_t)n_parts + 1, sizeof(int64_t));
  int64_t* order = (int64_t*)malloc(sizeof(int64_t) * (size_t)n);
  for (int64_t i = 0; i < n; i++) {
    int p = part ? part[i] : 0;
    if (p >= 0) start[p + 1]++;
  }
  for (int p = 0; p < n_parts; p++) start[p + 1] += start[p];
  int64_t* fill = (int64_t*)malloc(sizeof(int64_t) * (size_t)n_parts);
  memcpy(fill, start, sizeof(int64_t) * (size_t)n_parts);
  for (int64_t i = 0; i < n; i++) {
    int p = part ? part[i] : 0;
    if (p >= 0) order[fill[p]++] = i;
  }
  free(fill);

  size_t cap = 0;
  oro_pair* all = NULL;
  size_t n_all = 0;
  int64_t seq = 0;
  int total_parts = (cands == NULL) ? n_parts : nc;
  for (int ci = 0; ci < total_parts; ci++) {
    int p = (cands == NULL) ? ci : cands[ci];
    if (p < 0 || p >= n_parts) continue;
    if (start[p + 1] == start[p]) continue; /* ASM:298 skip empty partitions */
    oro_topn t;
    topn_init(&t, how_many);
    for (int64_t k = start[p]; k < start[p + 1]; k++) {
      int64_t row = order[k];
      if (excl != NULL && is_excluded(excl, n_excl, row)) continue; /* TNC:57 allowedPredicate */
      double s = score_of(kind, q, y + row * (int64_t)f, f);
      topn_accept(&t, row, s, seq++);
    }
    if (n_all + (size_t)t.size > cap) {
      cap = (n_all + (size_t)t.size) * 2 + 16;
      all = (oro_pair*)realloc(all, sizeof(oro_pair) * cap);
    }
    memcpy(all + n_all, t.heap, sizeof(oro_pair) * (size_t)t.size);
    n_all += (size_t)t.size;
    free(t.heap);
  }
  qsort(all, n_all, sizeof(oro_pair), cmp_desc); /* ASM:334 sorted(DESC) */
  int n_out = n_all < (size_t)how_many ? (int)n_all : how_many; /* ASM:334 limit */
  for (int i = 0; i < n_out; i++) {
    out_rows[i] = all[i].row;
    out_scores[i] = all[i].score;
  }
  free(all);
  free(order);
  free(start);
  return n_out;
}

/* TopNConsumer fed with explicit (row, score) pairs (TopNConsumerTest's shape, TNC:56-78). */
int oro_topn_values(const double* scores, int64_t n, int how_many, int64_t* out_rows,
                    double* out_scores) {
  oro_topn t;
  topn_init(&t, how_many);
  for (int64_t i = 0; i < n; i++) topn_accept(&t, i, scores[i], i);
  qsort(t.heap, (size_t)t.size, sizeof(oro_pair), cmp_desc);
  for (int i = 0; i < t.size; i++) {
    out_rows[i] = t.heap[i].row;
    out_scores[i] = t.heap[i].score;
  }
  int r = t.size;
  free(t.heap);
  return r;
}

/* ------------------------------------------------------------------ LinearSystemSolver / Solver */

/*
 * commons-math3 3.6 RRQRDecomposition (Householder QR with column pivoting), restated from the
 * published algorithm.  Storage follows QRDecomposition: qrt = A^T, i.e. qrt[col][row].
 */
typedef struct oro_solver {
  int n;
  double* qrt;    /* n x n, qrt[col*n + row] */
  double* r_diag; /* n */
  int* p;         /* column permutation */
  double threshold;
} oro_solver;

static void householder(oro_solver* s, int minor) {
  int n = s->n;
  double* qrt = s->qrt;
  /* RRQRDecomposition.performHouseholderReflection: pick the remaining column with the greatest
   * squared 2-norm (norm taken over the whole stored column), swap it into place. */
  double best = 0.0;
  int best_idx = minor;
  for (int i = minor; i < n; i++) {
    double l2 = 0.0;
    for (int j = 0; j < n; j++) l2 += qrt[(size_t)i * n + j] * qrt[(size_t)i * n + j];
    if (l2 > best) { best = l2; best_idx = i; }
  }
  if (best_idx != minor) {
    for (int j = 0; j < n; j++) {
      double t = qrt[(size_t)minor * n + j];
      qrt[(size_t)minor * n + j] = qrt[(size_t)best_idx * n + j];
      qrt[(size_t)best_idx * n + j] = t;
    }
    int tp = s->p[minor]; s->p[minor] = s->p[best_idx]; s->p[best_idx] = tp;
  }
  /* QRDecomposition.performHouseholderReflection */
  double* qm = qrt + (size_t)minor * n;
  double x_norm_sqr = 0.0;
  for (int row = minor; row < n; row++) x_norm_sqr += qm[row] * qm[row];
  double a = (qm[minor] > 0) ? -sqrt(x_norm_sqr) : sqrt(x_norm_sqr);
  s->r_diag[minor] = a;
  if (a != 0.0) {
    qm[minor] -= a;
    for (int col = minor + 1; col < n; col++) {
      double* qc = qrt + (size_t)col * n;
      double alpha = 0.0;
      for (int row = minor; row < n; row++) alpha -= qc[row] * qm[row];
      alpha /= a * qm[minor];
      for (int row = minor; row < n; row++) qc[row] -= alpha * qm[row];
    }
  }
}

/* AbstractRealMatrix.getNorm(): maximum absolute column sum (LSS:38 calls it infNorm). */
static double max_abs_col_sum(const double* m, int n) {
  double best = 0.0;
  for (int c = 0; c < n; c++) {
    double s = 0.0;
    for (int r = 0; r < n; r++) s += fabs(m[(size_t)r * n + c]);
    if (s > best) best = s;
  }
  return best;
}

static int solver_non_singular(const oro_solver* s) {
  /* QRDecomposition.Solver.isNonSingular(): every |rDiag| must exceed the threshold. */
  for (int i = 0; i < s->n; i++) {
    if (fabs(s->r_diag[i]) <= s->threshold) return 0;
  }
  return 1;
}

/* RRQRDecomposition.getRank(dropThreshold): walk Frobenius norms of trailing blocks of R. */
static int solver_rank(const oro_solver* s, double drop_threshold) {
  int n = s->n;
  /* R[row][col] = rDiag[row] on the diagonal, qrt[col][row] above it. */
  #define R_AT(row, col) ((row) == (col) ? s->r_diag[(row)] : ((col) > (row) ? s->qrt[(size_t)(col) * n + (row)] : 0.0))
  double fro = 0.0;
  for (int r = 0; r < n; r++) for (int c = 0; c < n; c++) fro += R_AT(r, c) * R_AT(r, c);
  double last_norm = sqrt(fro);
  double r_norm = last_norm;
  int rank = 1;
  while (rank < n) {
    double sub = 0.0;
    for (int r = rank; r < n; r++) for (int c = rank; c < n; c++) sub += R_AT(r, c) * R_AT(r, c);
    double this_norm = sqrt(sub);
    if (this_norm == 0 || (this_norm / last_norm) * r_norm < drop_threshold) break;
    last_norm = this_norm;
    rank++;
  }
  #undef R_AT
  return rank;
}

void oro_solver_destroy(oro_solver* s) {
  if (!s) return;
  free(s->qrt); free(s->r_diag); free(s->p); free(s);
}

/*
 * LSS:35-55  getSolver(): threshold = getNorm() * 1e-5; RRQR; non-singular => solver, else report
 * getRank(0.01).  Returns 0 and *out on success; 1 (singular) with *apparent_rank otherwise.
 * `m` is n x n row-major.
 */
int oro_solver_create(const double* m, int n, oro_solver** out, int* apparent_rank) {
  oro_solver* s = (oro_solver*)malloc(sizeof(oro_solver));
  s->n = n;
  s->qrt = (double*)malloc(sizeof(double) * (size_t)n * n);
  s->r_diag = (double*)calloc((size_t)n, sizeof(double));
  s->p = (int*)malloc(sizeof(int) * (size_t)n);
  for (int c = 0; c < n; c++) {
    s->p[c] = c;
    for (int r = 0; r < n; r++) s->qrt[(size_t)c * n + r] = m[(size_t)r * n + c];
  }
  s->threshold = max_abs_col_sum(m, n) * 1.0e-5; /* LSS:31,38-39 */
  for (int minor = 0; minor < n; minor++) householder(s, minor);
  if (solver_non_singular(s)) {
    *out = s;
    if (apparent_rank) *apparent_rank = n;
    return 0;
  }
  if (apparent_rank) *apparent_rank = solver_rank(s, 0.01); /* LSS:47 */
  oro_solver_destroy(s);
  *out = NULL;
  return 1;
}

/* LSS:57-63 isNonSingular(). */
int oro_is_non_singular(const double* m, int n) {
  oro_solver* s = NULL;
  int rc = oro_solver_create(m, n, &s, NULL);
  oro_solver_destroy(s);
  return rc == 0;
}

/* QRDecomposition.Solver.solve(RealVector) followed by RRQR's P * x un-permutation. */
void oro_solver_solve(const oro_solver* s, const double* b, double* x_out) {
  int n = s->n;
  double* y = (double*)malloc(sizeof(double) * (size_t)n);
  double* x = (double*)malloc(sizeof(double) * (size_t)n);
  memcpy(y, b, sizeof(double) * (size_t)n);
  for (int minor = 0; minor < n; minor++) {
    const double* qm = s->qrt + (size_t)minor * n;
    double dp = 0.0;
    for (int row = minor; row < n; row++) dp += y[row] * qm[row];
    dp /= s->r_diag[minor] * qm[minor];
    for (int row = minor; row < n; row++) y[row] += dp * qm[row];
  }
  for (int row = n - 1; row >= 0; --row) {
    y[row] /= s->r_diag[row];
    double y_row = y[row];
    const double* qr = s->qrt + (size_t)row * n;
    x[row] = y_row;
    for (int i = 0; i < row; i++) y[i] -= y_row * qr[i];
  }
  for (int i = 0; i < n; i++) x_out[s->p[i]] = x[i];
  free(y); free(x);
}

/* SOL:33-44  solveFToF(): widen b to double, solve, narrow each entry to float. */
void oro_solve_f2f(const oro_solver* s, const float* b, float* out) {
  int n = s->n;
  double* bd = (double*)malloc(sizeof(double) * (size_t)n);
  double* xd = (double*)malloc(sizeof(double) * (size_t)n);
  for (int i = 0; i < n; i++) bd[i] = (double)b[i];
  oro_solver_solve(s, bd, xd);
  for (int i = 0; i < n; i++) out[i] = (float)xd[i];
  free(bd); free(xd);
}

/* ------------------------------------------------------------------ ALSUtils */

/* ALSU:37-59  computeTargetQui(). */
double oro_compute_target_qui(int implicit, double value, double current_value) {
  if (implicit) {
    if (value > 0.0f && current_value < 1.0) {
      double diff = 1.0 - fmax(0.0, current_value);
      return current_value + (value / (1.0 + value)) * diff;
    }
    if (value < 0.0f && current_value > 0.0) {
      double diff = -fmin(1.0, current_value);
      return current_value + (value / (value - 1.0)) * diff;
    }
    return NAN;
  }
  return value;
}

/*
 * ALSU:74-106  computeUpdatedXu().  xu may be NULL (no existing vector); yi may be NULL.
 * Returns 1 and writes f floats to `out` when an update is produced, 0 when the reference
 * returns null.
 */
int oro_compute_updated_xu(const oro_solver* solver, double value, const float* xu,
                           const float* yi, int f, int implicit, float* out) {
  if (yi == NULL) return 0; /* ALSU:79-81 */
  double qui = xu == NULL ? 0.0 : oro_dot(xu, yi, f);
  double target = oro_compute_target_qui(implicit, value, xu == NULL ? 0.5 : qui);
  if (isnan(target)) return 0;
  double d_qui = target - qui;
  float* d_qui_yi = (float*)malloc(sizeof(float) * (size_t)f);
  for (int i = 0; i < f; i++) {
    /* `float *= double`: widen, multiply in double, narrow (ALSU:93-95). */
    d_qui_yi[i] = (float)((double)yi[i] * d_qui);
  }
  oro_solve_f2f(solver, d_qui_yi, out); /* ALSU:96 dXu */
  free(d_qui_yi);
  if (xu != NULL) {
    for (int i = 0; i < f; i++) out[i] = xu[i] + out[i]; /* ALSU:101-104 float add */
  }
  return 1;
}

/*
 * Batched form of the speed layer's hot section (ALSSpeedModelManager.java:182-204, one side):
 * for each k, out[k] = computeUpdatedXu(solver, values[k], xu[k] (if has_xu[k]), yi[k]).
 */
void oro_fold_in_batch(const oro_solver* solver, int implicit, int f, int64_t n,
                       const float* xu, const uint8_t* has_xu, const float* yi,
                       const double* values, float* out, uint8_t* out_valid) {
  for (int64_t k = 0; k < n; k++) {
    const float* x = (has_xu == NULL || has_xu[k]) ? xu + k * (int64_t)f : NULL;
    out_valid[k] = (uint8_t)oro_compute_updated_xu(solver, values[k], x, yi + k * (int64_t)f, f,
                                                   implicit, out + k * (int64_t)f);
  }
}

/* ------------------------------------------------------------------ timed CPU baseline */

/*
 * Multi-threaded "best-effort" arm of the CPU baseline: the same arithmetic and the same
 * TopNConsumer logic as above over a contiguous row-major Y, one task per slice of rows on a
 * pool of `threads` workers (the reference runs one task per LSH partition on a pool of
 * availableProcessors() threads, ASM:61-63,296-333).  Results are merged like ASM:329-334.
 */
typedef struct {
  int kind; const float* q; const float* y; int f;
  int64_t lo, hi; int how_many;
  const int32_t* excl; int n_excl;
  oro_topn t;
} oro_task;

static void* topn_worker(void* arg) {
  oro_task* k = (oro_task*)arg;
  topn_init(&k->t, k->how_many);
  for (int64_t row = k->lo; row < k->hi; row++) {
    if (k->excl != NULL && is_excluded(k->excl, k->n_excl, row)) continue;
    double s = score_of(k->kind, k->q, k->y + row * (int64_t)k->f, k->f);
    topn_accept(&k->t, row, s, row);
  }
  return NULL;
}

int oro_topn_mt(int kind, const float* q, const float* y, int64_t n, int f,
                const int32_t* excl, int n_excl, int how_many, int threads,
                int64_t* out_rows, double* out_scores) {
  if (threads < 1) threads = 1;
  oro_task* tasks = (oro_task*)calloc((size_t)threads, sizeof(oro_task));
  pthread_t* tids = (pthread_t*)malloc(sizeof(pthread_t) * (size_t)threads);
  int64_t per = (n + threads - 1) / threads;
  for (int i = 0; i < threads; i++) {
    oro_task* k = &tasks[i];
    k->kind = kind; k->q = q; k->y = y; k->f = f; k->how_many = how_many;
    k->excl = excl; k->n_excl = n_excl;
    k->lo = (int64_t)i * per; k->hi = k->lo + per; if (k->hi > n) k->hi = n; if (k->lo > n) k->lo = n;
    pthread_create(&tids[i], NULL, topn_worker, k);
  }
  size_t n_all = 0;
  oro_pair* all = (oro_pair*)malloc(sizeof(oro_pair) * (size_t)threads * (size_t)(how_many > 0 ? how_many : 1));
  for (int i = 0; i < threads; i++) {
    pthread_join(tids[i], NULL);
    memcpy(all + n_all, tasks[i].t.heap, sizeof(oro_pair) * (size_t)tasks[i].t.size);
    n_all += (size_t)tasks[i].t.size;
    free(tasks[i].t.heap);
  }
  qsort(all, n_all, sizeof(oro_pair), cmp_desc);
  int n_out = n_all < (size_t)how_many ? (int)n_all : how_many;
  for (int i = 0; i < n_out; i++) { out_rows[i] = all[i].row; out_scores[i] = all[i].score; }
  free(all); free(tasks); free(tids);
  return n_out;
}

/* ASM:357-373 getYTYSolver(): one transposeTimesSelf per partition (FV:150-154 getVTV), the partial
 * matrices added together in partition order (`YTY = YTY == null ? YTYpartial : YTY.add(YTYpartial)`).
 * Here a "partition" is a contiguous slice of rows, one worker thread per slice; each partial is the
 * exact VM:64-83 loop, and the partials are summed in slice order.  (Which rows fall in which partition
 * is an LSH detail the sum does not depend on beyond the order of binary64 additions.) */
typedef struct { const float* m; int64_t lo, hi; int f; double* out; } oro_gram_task;

static void* gram_worker(void* arg) {
  oro_gram_task* k = (oro_gram_task*)arg;
  oro_transpose_times_self(k->m + k->lo * (int64_t)k->f, k->hi - k->lo, k->f, k->out);
  return NULL;
}

void oro_transpose_times_self_mt(const float* m, int64_t rows, int f, int threads, double* out) {
  if (threads < 1) threads = 1;
  size_t ff = (size_t)f * (size_t)f;
  oro_gram_task* tasks = (oro_gram_task*)calloc((size_t)threads, sizeof(oro_gram_task));
  pthread_t* tids = (pthread_t*)malloc(sizeof(pthread_t) * (size_t)threads);
  double* partial = (double*)malloc(sizeof(double) * ff * (size_t)threads);
  int64_t per = (rows + threads - 1) / threads;
  for (int i = 0; i < threads; i++) {
    oro_gram_task* k = &tasks[i];
    k->m = m; k->f = f; k->out = partial + ff * (size_t)i;
    k->lo = (int64_t)i * per; k->hi = k->lo + per; if (k->hi > rows) k->hi = rows; if (k->lo > rows) k->lo = rows;
    pthread_create(&tids[i], NULL, gram_worker, k);
  }
  memset(out, 0, sizeof(double) * ff);
  for (int i = 0; i < threads; i++) {
    pthread_join(tids[i], NULL);
    const double* p = partial + ff * (size_t)i;
    if (tasks[i].hi > tasks[i].lo)
      for (size_t e = 0; e < ff; e++) out[e] += p[e];
  }
  free(partial); free(tasks); free(tids);
}

/* ------------------------------------------------------------------ "faithful" CPU baseline
 *
 * The second arm of the timed CPU baseline (variant i): not only the
 * reference's arithmetic but its data layout and control flow --
 *   - one hash map per LSH partition, String id -> separately heap-allocated float[F]
 *     (FV:36-46 ObjObjMap<String,float[]>; ASM:69,97-124 `FeatureVectors[] Y`);
 *   - the allowed-predicate is a hash-set lookup on the String id for every item
 *     (Recommend.java:92-99 `knownItems.contains(id)`; TNC:57-59);
 *   - one task per candidate partition on a pool of `threads` workers, each with its own
 *     TopNConsumer, results concatenated and sorted (ASM:295-334).
 * Ids carry their cached hash, as a java.lang.String does.  Iteration is in table order, so ties
 * between equal scores may resolve differently from the row-ordered arms (as in the reference,
 * where Koloboke's iteration order decides).
 */
typedef struct { char* id; uint32_t hash; float* vec; int64_t row; } oro_fentry;
typedef struct { oro_fentry* slots; size_t cap, n; } oro_fmap;
typedef struct { int f, parts; oro_fmap* maps; } oro_faithful;

static uint32_t str_hash(const char* s) { /* String.hashCode(): s[0]*31^(n-1) + ... */
  uint32_t h = 0;
  for (; *s; s++) h = 31u * h + (uint32_t)(unsigned char)*s;
  return h;
}
static uint32_t mix_hash(uint32_t h) { h ^= h >> 16; h *= 0x9E3779B1u; return h ^ (h >> 15); }

static void fmap_grow(oro_fmap* m) {
  size_t ncap = m->cap ? m->cap * 2 : 64;
  oro_fentry* ns = (oro_fentry*)calloc(ncap, sizeof(oro_fentry));
  for (size_t i = 0; i < m->cap; i++) {
    if (!m->slots[i].id) continue;
    size_t p = mix_hash(m->slots[i].hash) & (ncap - 1);
    while (ns[p].id) p = (p + 1) & (ncap - 1);
    ns[p] = m->slots[i];
  }
  free(m->slots);
  m->slots = ns;
  m->cap = ncap;
}

void* oro_faithful_create(int features, int partitions) {
  oro_faithful* h = (oro_faithful*)calloc(1, sizeof(oro_faithful));
  h->f = features;
  h->parts = partitions < 1 ? 1 : partitions;
  h->maps = (oro_fmap*)calloc((size_t)h->parts, sizeof(oro_fmap));
  return h;
}

/* setItemVector (ASM:161-183): new float[] per item, keyed by id, in the partition's map. */
void oro_faithful_put(void* hh, int partition, const char* id, int64_t row, const float* vec) {
  oro_faithful* h = (oro_faithful*)hh;
  oro_fmap* m = &h->maps[partition];
  if ((m->n + 1) * 2 > m->cap) fmap_grow(m);
  uint32_t hs = str_hash(id);
  size_t p = mix_hash(hs) & (m->cap - 1);
  while (m->slots[p].id) {
    if (m->slots[p].hash == hs && strcmp(m->slots[p].id, id) == 0) {
      memcpy(m->slots[p].vec, vec, sizeof(float) * (size_t)h->f);
      return;
    }
    p = (p + 1) & (m->cap - 1);
  }
  {
    size_t len = strlen(id) + 1;
    m->slots[p].id = (char*)malloc(len);
    memcpy(m->slots[p].id, id, len);
  }
  m->slots[p].hash = hs;
  m->slots[p].row = row;
  m->slots[p].vec = (float*)malloc(sizeof(float) * (size_t)h->f);
  memcpy(m->slots[p].vec, vec, sizeof(float) * (size_t)h->f);
  m->n++;
}

/* Bulk load: rows [0, n) of y with ids "<prefix><row>" (LoadTestALSModelFactory.java:62-95 ids "I<j>"),
 * partition part[row] (NULL: partition 0). */
void oro_faithful_put_rows(void* hh, const int32_t* part, const char* prefix, const float* y, int64_t n) {
  oro_faithful* h = (oro_faithful*)hh;
  char id[64];
  for (int64_t row = 0; row < n; row++) {
    snprintf(id, sizeof(id), "%s%lld", prefix, (long long)row);
    oro_faithful_put(hh, part ? part[row] : 0, id, row, y + row * (int64_t)h->f);
  }
}

void oro_faithful_destroy(void* hh) {
  oro_faithful* h = (oro_faithful*)hh;
  if (!h) return;
  for (int p = 0; p < h->parts; p++) {
    for (size_t i = 0; i < h->maps[p].cap; i++) { free(h->maps[p].slots[i].id); free(h->maps[p].slots[i].vec); }
    free(h->maps[p].slots);
  }
  free(h->maps);
  free(h);
}

typedef struct { char** ids; uint32_t* hashes; size_t cap; } oro_strset;
static int strset_contains(const oro_strset* s, const char* id, uint32_t hs) {
  if (s->cap == 0) return 0;
  size_t p = mix_hash(hs) & (s->cap - 1);
  while (s->ids[p]) {
    if (s->hashes[p] == hs && strcmp(s->ids[p], id) == 0) return 1;
    p = (p + 1) & (s->cap - 1);
  }
  return 0;
}

typedef struct {
  const oro_faithful* h; int kind; const float* q; int how_many;
  const int32_t* cand; int n_cand; int next; pthread_mutex_t* mu;
  const oro_strset* known;
  oro_topn* per_task; /* one TopNConsumer per candidate partition */
} oro_fjob;

static void* faithful_worker(void* arg) {
  oro_fjob* j = (oro_fjob*)arg;
  for (;;) {
    pthread_mutex_lock(j->mu);
    int t = j->next < j->n_cand ? j->next++ : -1;
    pthread_mutex_unlock(j->mu);
    if (t < 0) return NULL;
    const oro_fmap* m = &j->h->maps[j->cand[t]];
    oro_topn* top = &j->per_task[t];
    topn_init(top, j->how_many);
    for (size_t i = 0; i < m->cap; i++) { /* FeatureVectors.forEach (FV:141-145) */
      const oro_fentry* e = &m->slots[i];
      if (!e->id) continue;
      if (j->known && strset_contains(j->known, e->id, e->hash)) continue; /* TNC:57 */
      double s = score_of(j->kind, j->q, e->vec, j->h->f);
      topn_accept(top, e->row, s, e->row);
    }
  }
}

/* ASM:289-335 over the candidate partitions `cand`; known_ids is the user's known-items set. */
int oro_faithful_topn(void* hh, int kind, const float* q, int how_many, const int32_t* cand, int n_cand,
                      const char* const* known_ids, int n_known, int threads, int64_t* out_rows,
                      double* out_scores) {
  oro_faithful* h = (oro_faithful*)hh;
  oro_strset known = {0};
  if (n_known > 0) {
    size_t cap = 16;
    while (cap < (size_t)n_known * 2) cap <<= 1;
    known.cap = cap;
    known.ids = (char**)calloc(cap, sizeof(char*));
    known.hashes = (uint32_t*)calloc(cap, sizeof(uint32_t));
    for (int i = 0; i < n_known; i++) {
      uint32_t hs = str_hash(known_ids[i]);
      size_t p = mix_hash(hs) & (cap - 1);
      while (known.ids[p]) p = (p + 1) & (cap - 1);
      known.ids[p] = (char*)known_ids[i];
      known.hashes[p] = hs;
    }
  }
  if (threads < 1) threads = 1;
  pthread_mutex_t mu = PTHREAD_MUTEX_INITIALIZER;
  oro_fjob job;
  job.h = h; job.kind = kind; job.q = q; job.how_many = how_many;
  job.cand = cand; job.n_cand = n_cand; job.next = 0; job.mu = &mu;
  job.known = n_known > 0 ? &known : NULL;
  job.per_task = (oro_topn*)calloc((size_t)(n_cand > 0 ? n_cand : 1), sizeof(oro_topn));
  pthread_t* tids = (pthread_t*)malloc(sizeof(pthread_t) * (size_t)threads);
  for (int i = 0; i < threads; i++) pthread_create(&tids[i], NULL, faithful_worker, &job);
  for (int i = 0; i < threads; i++) pthread_join(tids[i], NULL);
  size_t n_all = 0;
  oro_pair* all = (oro_pair*)malloc(sizeof(oro_pair) * (size_t)(n_cand > 0 ? n_cand : 1) * (size_t)(how_many > 0 ? how_many : 1));
  for (int t = 0; t < n_cand; t++) {
    if (!job.per_task[t].heap) continue;
    memcpy(all + n_all, job.per_task[t].heap, sizeof(oro_pair) * (size_t)job.per_task[t].size);
    n_all += (size_t)job.per_task[t].size;
    free(job.per_task[t].heap);
  }
  qsort(all, n_all, sizeof(oro_pair), cmp_desc);
  int n_out = n_all < (size_t)how_many ? (int)n_all : how_many;
  for (int i = 0; i < n_out; i++) { out_rows[i] = all[i].row; out_scores[i] = all[i].score; }
  free(all); free(job.per_task); free(tids); free(known.ids); free(known.hashes);
  return n_out;
}
