/*
 * oryx_b200_jni.c -- JNI glue between com.cloudera.oryx.app.als.gpu.OryxB200 (jni/OryxB200.java) and the C
 * ABI of include/oryx_b200.h.  One function per native method; nothing here computes anything.
 *
 * Build (on a machine with a JDK; this repository's image has none, so the file ships as source and is
 * only syntax-checked against a stand-in header by tests/test_jni_source.py):
 *   gcc -shared -fPIC -I"$JAVA_HOME/include" -I"$JAVA_HOME/include/linux" -Iinclude \
 *       jni/oryx_b200_jni.c -Loryx_b200 -loryx_b200 -o liboryx_b200_jni.so
 *
 * Conventions
 *   - handles travel as jlong; arrays are borrowed with GetPrimitiveArrayCritical for the duration of the
 *     call only (the library never keeps a caller buffer), so no copies are made on the way in;
 *   - a non-zero status becomes a Java exception (throw_for_status) and the
 *     function returns null / 0;
 *   - "no such row / user" is not an error: the Java method returns null, as the reference's model does.
 */
#include <jni.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "oryx_b200.h"

#define PKG "com/cloudera/oryx/app/als/gpu/"
#define FN(name) Java_com_cloudera_oryx_app_als_gpu_OryxB200_##name

/* ---------------------------------------------------------------- helpers */

static void throw_named(JNIEnv* env, const char* cls_name, const char* msg) {
  jclass cls = (*env)->FindClass(env, cls_name);
  if (cls) (*env)->ThrowNew(env, cls, msg);
}

/* ORYX_E_INVALID -> IllegalArgumentException (Preconditions.checkArgument, ALSServingModel.java:98-99,154,162);
 * ORYX_E_CUDA / ORYX_E_NOMEM -> IllegalStateException (what topN throws on task failure, :316-331);
 * ORYX_E_UNSUPPORTED -> UnsupportedOperationException.  ORYX_E_SINGULAR is handled by solverCreate. */
static void throw_for_status(JNIEnv* env, int32_t status) {
  char msg[512];
  msg[0] = 0;
  oryx_last_error(msg, (int32_t)sizeof(msg));
  switch (status) {
    case ORYX_E_INVALID: throw_named(env, "java/lang/IllegalArgumentException", msg); break;
    case ORYX_E_UNSUPPORTED: throw_named(env, "java/lang/UnsupportedOperationException", msg); break;
    default: throw_named(env, "java/lang/IllegalStateException", msg); break;
  }
}

static void* pin(JNIEnv* env, jarray a) { return a ? (*env)->GetPrimitiveArrayCritical(env, a, NULL) : NULL; }
static void unpin(JNIEnv* env, jarray a, void* p) { if (a && p) (*env)->ReleasePrimitiveArrayCritical(env, a, p, JNI_ABORT); }
static jsize len(JNIEnv* env, jarray a) { return a ? (*env)->GetArrayLength(env, a) : 0; }

static jintArray new_ints(JNIEnv* env, const int32_t* v, jsize n) {
  jintArray a = (*env)->NewIntArray(env, n);
  if (a && n > 0) (*env)->SetIntArrayRegion(env, a, 0, n, (const jint*)v);
  return a;
}
static jdoubleArray new_doubles(JNIEnv* env, const double* v, jsize n) {
  jdoubleArray a = (*env)->NewDoubleArray(env, n);
  if (a && n > 0) (*env)->SetDoubleArrayRegion(env, a, 0, n, v);
  return a;
}
static jfloatArray new_floats(JNIEnv* env, const float* v, jsize n) {
  jfloatArray a = (*env)->NewFloatArray(env, n);
  if (a && n > 0) (*env)->SetFloatArrayRegion(env, a, 0, n, v);
  return a;
}
/* {int[] rows, double[] scores} */
static jobjectArray rows_and_scores(JNIEnv* env, const int32_t* rows, const double* scores, jsize n) {
  jclass obj = (*env)->FindClass(env, "java/lang/Object");
  jobjectArray out = obj ? (*env)->NewObjectArray(env, 2, obj, NULL) : NULL;
  if (!out) return NULL;
  (*env)->SetObjectArrayElement(env, out, 0, new_ints(env, rows, n));
  (*env)->SetObjectArrayElement(env, out, 1, new_doubles(env, scores, n));
  return out;
}

/* ---------------------------------------------------------------- model */

JNIEXPORT jintArray JNICALL FN(lshPlan)(JNIEnv* env, jclass cls, jdouble sample_rate, jint num_cores) {
  (void)cls;
  int32_t hb[2] = {0, 0};
  int32_t st = oryx_lsh_plan(sample_rate, num_cores, &hb[0], &hb[1]);
  if (st != ORYX_OK) { throw_for_status(env, st); return NULL; }
  return new_ints(env, hb, 2);
}

JNIEXPORT jlong JNICALL FN(modelCreate)(JNIEnv* env, jclass cls, jint features, jboolean implicit, jint num_hashes,
                                        jint max_bits, jfloatArray hash_vectors, jint device, jlong capacity_hint) {
  (void)cls;
  oryx_config cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.features = features;
  cfg.implicit = implicit ? 1 : 0;
  cfg.num_hashes = num_hashes;
  cfg.max_bits_differing = max_bits;
  cfg.device = device;
  cfg.capacity_hint = capacity_hint;
  float* hv = (float*)pin(env, hash_vectors);   /* copied by the library during the call */
  cfg.hash_vectors = hv;
  oryx_model* m = NULL;
  int32_t st = oryx_model_create(&cfg, &m);
  unpin(env, hash_vectors, hv);
  if (st != ORYX_OK) { throw_for_status(env, st); return 0; }
  return (jlong)(intptr_t)m;
}

JNIEXPORT void JNICALL FN(modelDestroy)(JNIEnv* env, jclass cls, jlong model) {
  (void)env; (void)cls;
  oryx_model_destroy((oryx_model*)(intptr_t)model);
}

JNIEXPORT jintArray JNICALL FN(itemsUpsert)(JNIEnv* env, jclass cls, jlong model, jint which, jintArray rows,
                                            jfloatArray vecs) {
  (void)cls;
  const jsize n = len(env, rows);
  int32_t* parts = (int32_t*)malloc(sizeof(int32_t) * (size_t)(n > 0 ? n : 1));
  if (!parts) { throw_named(env, "java/lang/OutOfMemoryError", "itemsUpsert"); return NULL; }
  int32_t* r = (int32_t*)pin(env, rows);
  float* v = (float*)pin(env, vecs);
  int32_t st = oryx_items_upsert((oryx_model*)(intptr_t)model, which, r, v, n, parts);
  unpin(env, vecs, v);
  unpin(env, rows, r);
  jintArray out = NULL;
  if (st != ORYX_OK) throw_for_status(env, st);
  else out = new_ints(env, parts, n);
  free(parts);
  return out;
}

JNIEXPORT void JNICALL FN(itemsRemove)(JNIEnv* env, jclass cls, jlong model, jint which, jintArray rows) {
  (void)cls;
  int32_t* r = (int32_t*)pin(env, rows);
  int32_t st = oryx_items_remove((oryx_model*)(intptr_t)model, which, r, len(env, rows));
  unpin(env, rows, r);
  if (st != ORYX_OK) throw_for_status(env, st);
}

JNIEXPORT jfloatArray JNICALL FN(itemsGet)(JNIEnv* env, jclass cls, jlong model, jint which, jint row, jint features) {
  (void)cls;
  if (features < 1) { throw_named(env, "java/lang/IllegalArgumentException", "features must be >= 1"); return NULL; }
  float* buf = (float*)malloc(sizeof(float) * (size_t)features);
  if (!buf) { throw_named(env, "java/lang/OutOfMemoryError", "itemsGet"); return NULL; }
  int32_t found = 0;
  int32_t st = oryx_items_get((oryx_model*)(intptr_t)model, which, row, buf, &found);
  jfloatArray out = NULL;
  if (st != ORYX_OK) throw_for_status(env, st);
  else if (found) out = new_floats(env, buf, features);   /* null: the row has no vector */
  free(buf);
  return out;
}

JNIEXPORT void JNICALL FN(itemsCompact)(JNIEnv* env, jclass cls, jlong model, jint which) {
  (void)cls;
  int32_t st = oryx_items_compact((oryx_model*)(intptr_t)model, which);
  if (st != ORYX_OK) throw_for_status(env, st);
}

/* ---------------------------------------------------------------- LSH */

JNIEXPORT jint JNICALL FN(lshIndex)(JNIEnv* env, jclass cls, jlong model, jfloatArray vec) {
  (void)cls;
  int32_t idx = 0;
  float* v = (float*)pin(env, vec);
  int32_t st = oryx_lsh_index((oryx_model*)(intptr_t)model, v, &idx);
  unpin(env, vec, v);
  if (st != ORYX_OK) { throw_for_status(env, st); return 0; }
  return idx;
}

JNIEXPORT jintArray JNICALL FN(lshCandidates)(JNIEnv* env, jclass cls, jlong model, jfloatArray vec) {
  (void)cls;
  int32_t* out = (int32_t*)malloc(sizeof(int32_t) * ((size_t)1 << ORYX_MAX_HASHES));
  if (!out) { throw_named(env, "java/lang/OutOfMemoryError", "lshCandidates"); return NULL; }
  int32_t n = 0;
  float* v = (float*)pin(env, vec);
  int32_t st = oryx_lsh_candidates((oryx_model*)(intptr_t)model, v, out, &n);
  unpin(env, vec, v);
  jintArray res = NULL;
  if (st != ORYX_OK) throw_for_status(env, st);
  else res = new_ints(env, out, n);
  free(out);
  return res;
}

/* ---------------------------------------------------------------- queries */

typedef int32_t (*topn_fn)(oryx_model*, int32_t, const float*, int32_t, const int32_t*, int32_t, int32_t*, double*,
                           int32_t*);

static jobjectArray topn_common(JNIEnv* env, topn_fn fn, jlong model, jint kind, jfloatArray query, jint how_many,
                                jintArray excluded) {
  if (how_many < 1) { throw_named(env, "java/lang/IllegalArgumentException", "howMany must be >= 1"); return NULL; }
  int32_t* rows = (int32_t*)malloc(sizeof(int32_t) * (size_t)how_many);
  double* scores = (double*)malloc(sizeof(double) * (size_t)how_many);
  if (!rows || !scores) { free(rows); free(scores); throw_named(env, "java/lang/OutOfMemoryError", "topN"); return NULL; }
  int32_t n = 0;
  float* q = (float*)pin(env, query);
  int32_t* ex = (int32_t*)pin(env, excluded);
  int32_t st = fn((oryx_model*)(intptr_t)model, kind, q, how_many, ex, len(env, excluded), rows, scores, &n);
  unpin(env, excluded, ex);
  unpin(env, query, q);
  jobjectArray out = NULL;
  if (st != ORYX_OK) throw_for_status(env, st);
  else out = rows_and_scores(env, rows, scores, n);
  free(rows);
  free(scores);
  return out;
}

JNIEXPORT jobjectArray JNICALL FN(topN)(JNIEnv* env, jclass cls, jlong model, jint kind, jfloatArray query, jint how_many,
                                        jintArray excluded) {
  (void)cls;
  return topn_common(env, oryx_topn, model, kind, query, how_many, excluded);
}

JNIEXPORT jobjectArray JNICALL FN(topNCoalesced)(JNIEnv* env, jclass cls, jlong model, jint kind, jfloatArray query,
                                                 jint how_many, jintArray excluded) {
  (void)cls;
  return topn_common(env, oryx_topn_coalesced, model, kind, query, how_many, excluded);
}

JNIEXPORT jobjectArray JNICALL FN(topNBatch)(JNIEnv* env, jclass cls, jlong model, jint kind, jfloatArray queries, jint nq,
                                             jint how_many, jintArray excl_offsets, jintArray excl_rows) {
  (void)cls;
  if (nq < 0 || how_many < 1) { throw_named(env, "java/lang/IllegalArgumentException", "nq >= 0 and howMany >= 1 required"); return NULL; }
  if (excl_offsets && len(env, excl_offsets) != nq + 1) {
    throw_named(env, "java/lang/IllegalArgumentException", "exclOffsets must hold nq + 1 entries");
    return NULL;
  }
  const size_t cells = (size_t)(nq > 0 ? nq : 1) * (size_t)how_many;
  int32_t* rows = (int32_t*)malloc(sizeof(int32_t) * cells);
  double* scores = (double*)malloc(sizeof(double) * cells);
  int32_t* cnt = (int32_t*)calloc((size_t)(nq > 0 ? nq : 1), sizeof(int32_t));
  jobjectArray out = NULL;
  if (!rows || !scores || !cnt) {
    throw_named(env, "java/lang/OutOfMemoryError", "topNBatch");
  } else {
    float* q = (float*)pin(env, queries);
    int32_t* eo = (int32_t*)pin(env, excl_offsets);
    int32_t* er = (int32_t*)pin(env, excl_rows);
    const int has_excl = eo != NULL && er != NULL && eo[nq] > 0;
    int32_t st = oryx_topn_batch((oryx_model*)(intptr_t)model, kind, q, nq, how_many, has_excl ? eo : NULL,
                                 has_excl ? er : NULL, rows, scores, cnt);
    unpin(env, excl_rows, er);
    unpin(env, excl_offsets, eo);
    unpin(env, queries, q);
    if (st != ORYX_OK) {
      throw_for_status(env, st);
    } else {
      jclass obj = (*env)->FindClass(env, "java/lang/Object");
      out = obj ? (*env)->NewObjectArray(env, 3, obj, NULL) : NULL;
      if (out) {
        (*env)->SetObjectArrayElement(env, out, 0, new_ints(env, rows, nq * how_many));
        (*env)->SetObjectArrayElement(env, out, 1, new_doubles(env, scores, nq * how_many));
        (*env)->SetObjectArrayElement(env, out, 2, new_ints(env, cnt, nq));
      }
    }
  }
  free(rows); free(scores); free(cnt);
  return out;
}

JNIEXPORT jdoubleArray JNICALL FN(scoresAll)(JNIEnv* env, jclass cls, jlong model, jint kind, jfloatArray query,
                                             jint n_rows) {
  (void)cls;
  if (n_rows < 0) { throw_named(env, "java/lang/IllegalArgumentException", "nRows must be >= 0"); return NULL; }
  jdoubleArray out = (*env)->NewDoubleArray(env, n_rows);
  if (!out) return NULL;
  /* the result array itself is the output buffer: no second copy of what may be tens of megabytes */
  float* q = (float*)pin(env, query);
  double* o = (double*)pin(env, out);
  int32_t st = oryx_scores_all((oryx_model*)(intptr_t)model, kind, q, n_rows, o);
  if (out && o) (*env)->ReleasePrimitiveArrayCritical(env, out, o, 0);   /* 0: copy back if it was a copy */
  unpin(env, query, q);
  if (st != ORYX_OK) { throw_for_status(env, st); return NULL; }
  return out;
}

JNIEXPORT jdoubleArray JNICALL FN(scoresRows)(JNIEnv* env, jclass cls, jlong model, jint kind, jfloatArray query,
                                              jintArray rows) {
  (void)cls;
  const jsize n = len(env, rows);
  double* scores = (double*)malloc(sizeof(double) * (size_t)(n > 0 ? n : 1));
  uint8_t* found = (uint8_t*)malloc((size_t)(n > 0 ? n : 1));
  jdoubleArray out = NULL;
  if (!scores || !found) {
    throw_named(env, "java/lang/OutOfMemoryError", "scoresRows");
  } else {
    float* q = (float*)pin(env, query);
    int32_t* r = (int32_t*)pin(env, rows);
    int32_t st = oryx_scores_rows((oryx_model*)(intptr_t)model, kind, q, r, n, scores, found);
    unpin(env, rows, r);
    unpin(env, query, q);
    if (st != ORYX_OK) {
      throw_for_status(env, st);
    } else {
      for (jsize i = 0; i < n; i++) if (!found[i]) scores[i] = 0.0 / 0.0;   /* NaN: the row has no vector */
      out = new_doubles(env, scores, n);
    }
  }
  free(scores); free(found);
  return out;
}

JNIEXPORT jdoubleArray JNICALL FN(estimatePairs)(JNIEnv* env, jclass cls, jlong model, jintArray user_rows,
                                                 jintArray item_rows) {
  (void)cls;
  const jsize n = len(env, user_rows);
  if (len(env, item_rows) != n) { throw_named(env, "java/lang/IllegalArgumentException", "row arrays differ in length"); return NULL; }
  double* scores = (double*)malloc(sizeof(double) * (size_t)(n > 0 ? n : 1));
  if (!scores) { throw_named(env, "java/lang/OutOfMemoryError", "estimatePairs"); return NULL; }
  int32_t* u = (int32_t*)pin(env, user_rows);
  int32_t* it = (int32_t*)pin(env, item_rows);
  int32_t st = oryx_estimate_pairs((oryx_model*)(intptr_t)model, u, it, n, 0, scores, NULL);
  unpin(env, item_rows, it);
  unpin(env, user_rows, u);
  jdoubleArray out = NULL;
  if (st != ORYX_OK) throw_for_status(env, st);
  else out = new_doubles(env, scores, n);
  free(scores);
  return out;
}

/* ---------------------------------------------------------------- known items, /recommend from device state */

JNIEXPORT void JNICALL FN(knownItemsAdd)(JNIEnv* env, jclass cls, jlong model, jintArray user_rows, jlongArray begin,
                                         jintArray item_rows) {
  (void)cls;
  int32_t* u = (int32_t*)pin(env, user_rows);
  int64_t* b = (int64_t*)pin(env, begin);
  int32_t* it = (int32_t*)pin(env, item_rows);
  int32_t st = oryx_known_items_add((oryx_model*)(intptr_t)model, u, b, it, len(env, user_rows));
  unpin(env, item_rows, it);
  unpin(env, begin, b);
  unpin(env, user_rows, u);
  if (st != ORYX_OK) throw_for_status(env, st);
}

JNIEXPORT jintArray JNICALL FN(knownItemsGet)(JNIEnv* env, jclass cls, jlong model, jint user_row) {
  (void)cls;
  int32_t n = 0;
  /* first call sizes the list (ORYX_E_NOMEM with *out_n set when the buffer is too small) */
  int32_t st = oryx_known_items_get((oryx_model*)(intptr_t)model, user_row, NULL, 0, &n);
  if (st != ORYX_OK && st != ORYX_E_NOMEM) { throw_for_status(env, st); return NULL; }
  int32_t* rows = (int32_t*)malloc(sizeof(int32_t) * (size_t)(n > 0 ? n : 1));
  if (!rows) { throw_named(env, "java/lang/OutOfMemoryError", "knownItemsGet"); return NULL; }
  st = oryx_known_items_get((oryx_model*)(intptr_t)model, user_row, rows, n, &n);
  jintArray out = NULL;
  if (st != ORYX_OK) throw_for_status(env, st);
  else out = new_ints(env, rows, n);
  free(rows);
  return out;
}

JNIEXPORT void JNICALL FN(knownItemsRetain)(JNIEnv* env, jclass cls, jlong model, jbyteArray keep_users,
                                            jbyteArray keep_items) {
  (void)cls;
  uint8_t* ku = (uint8_t*)pin(env, keep_users);
  uint8_t* ki = (uint8_t*)pin(env, keep_items);
  int32_t st = oryx_known_items_retain((oryx_model*)(intptr_t)model, ku, len(env, keep_users), ki, len(env, keep_items));
  unpin(env, keep_items, ki);
  unpin(env, keep_users, ku);
  if (st != ORYX_OK) throw_for_status(env, st);
}

JNIEXPORT jintArray JNICALL FN(knownCounts)(JNIEnv* env, jclass cls, jlong model, jint which, jint n_rows) {
  (void)cls;
  if (n_rows < 0) { throw_named(env, "java/lang/IllegalArgumentException", "nRows must be >= 0"); return NULL; }
  int32_t* counts = (int32_t*)calloc((size_t)(n_rows > 0 ? n_rows : 1), sizeof(int32_t));
  if (!counts) { throw_named(env, "java/lang/OutOfMemoryError", "knownCounts"); return NULL; }
  int32_t st = oryx_known_counts((oryx_model*)(intptr_t)model, which, counts, n_rows);
  jintArray out = NULL;
  if (st != ORYX_OK) throw_for_status(env, st);
  else out = new_ints(env, counts, n_rows);
  free(counts);
  return out;
}

typedef int32_t (*recommend_fn)(oryx_model*, int32_t, int32_t, int32_t, int32_t*, double*, int32_t*, int32_t*);

static jobjectArray recommend_common(JNIEnv* env, recommend_fn fn, jlong model, jint user_row, jint how_many,
                                     jboolean consider_known) {
  if (how_many < 1) { throw_named(env, "java/lang/IllegalArgumentException", "howMany must be >= 1"); return NULL; }
  int32_t* rows = (int32_t*)malloc(sizeof(int32_t) * (size_t)how_many);
  double* scores = (double*)malloc(sizeof(double) * (size_t)how_many);
  if (!rows || !scores) { free(rows); free(scores); throw_named(env, "java/lang/OutOfMemoryError", "recommend"); return NULL; }
  int32_t n = 0, found = 0;
  int32_t st = fn((oryx_model*)(intptr_t)model, user_row, how_many, consider_known ? 1 : 0, rows, scores, &n, &found);
  jobjectArray out = NULL;
  if (st != ORYX_OK) throw_for_status(env, st);
  else if (found) out = rows_and_scores(env, rows, scores, n);   /* null: no such user (HTTP 404) */
  free(rows);
  free(scores);
  return out;
}

JNIEXPORT jobjectArray JNICALL FN(recommend)(JNIEnv* env, jclass cls, jlong model, jint user_row, jint how_many,
                                             jboolean consider_known) {
  (void)cls;
  return recommend_common(env, oryx_recommend, model, user_row, how_many, consider_known);
}

JNIEXPORT jobjectArray JNICALL FN(recommendCoalesced)(JNIEnv* env, jclass cls, jlong model, jint user_row, jint how_many,
                                                      jboolean consider_known) {
  (void)cls;
  return recommend_common(env, oryx_recommend_coalesced, model, user_row, how_many, consider_known);
}

JNIEXPORT jobjectArray JNICALL FN(recommendBatch)(JNIEnv* env, jclass cls, jlong model, jintArray user_rows, jint how_many,
                                                  jboolean consider_known) {
  (void)cls;
  const jsize n = len(env, user_rows);
  if (how_many < 1) { throw_named(env, "java/lang/IllegalArgumentException", "howMany must be >= 1"); return NULL; }
  int32_t* rows = (int32_t*)malloc(sizeof(int32_t) * (size_t)(n > 0 ? n : 1) * (size_t)how_many);
  double* scores = (double*)malloc(sizeof(double) * (size_t)(n > 0 ? n : 1) * (size_t)how_many);
  int32_t* cnt = (int32_t*)calloc((size_t)(n > 0 ? n : 1), sizeof(int32_t));
  jobjectArray out = NULL;
  if (!rows || !scores || !cnt) {
    throw_named(env, "java/lang/OutOfMemoryError", "recommendBatch");
  } else {
    int32_t* u = (int32_t*)pin(env, user_rows);
    int32_t st = oryx_recommend_batch((oryx_model*)(intptr_t)model, u, n, how_many, consider_known ? 1 : 0, rows, scores, cnt);
    unpin(env, user_rows, u);
    if (st != ORYX_OK) {
      throw_for_status(env, st);
    } else {
      jclass obj = (*env)->FindClass(env, "java/lang/Object");
      out = obj ? (*env)->NewObjectArray(env, 3, obj, NULL) : NULL;
      if (out) {
        (*env)->SetObjectArrayElement(env, out, 0, new_ints(env, rows, n * how_many));
        (*env)->SetObjectArrayElement(env, out, 1, new_doubles(env, scores, n * how_many));
        (*env)->SetObjectArrayElement(env, out, 2, new_ints(env, cnt, n));   /* -1: no such user */
      }
    }
  }
  free(rows); free(scores); free(cnt);
  return out;
}

/* ---------------------------------------------------------------- Gram, solver, fold-in */

JNIEXPORT jdoubleArray JNICALL FN(gram)(JNIEnv* env, jclass cls, jlong model, jint which, jint features) {
  (void)cls;
  if (features < 1) { throw_named(env, "java/lang/IllegalArgumentException", "features must be >= 1"); return NULL; }
  jdoubleArray out = (*env)->NewDoubleArray(env, features * features);
  if (!out) return NULL;
  int64_t used = 0;
  double* o = (double*)pin(env, out);
  int32_t st = oryx_gram((oryx_model*)(intptr_t)model, which, o, &used);
  if (o) (*env)->ReleasePrimitiveArrayCritical(env, out, o, 0);
  if (st != ORYX_OK) { throw_for_status(env, st); return NULL; }
  return used > 0 ? out : NULL;   /* FeatureVectors.getVTV returns null for an empty store */
}

/* The global Gram of a store sharded over the GPUs of this box: the shard's Gram, then ONE kernel over peer memory
 * that adds every rank's matrix (ALSServingModel.java:362-368 across GPUs).  Every rank calls it; never null. */
JNIEXPORT jdoubleArray JNICALL FN(gramSharded)(JNIEnv* env, jclass cls, jlong model, jlong exchange, jint which, jint features) {
  (void)cls;
  if (features < 1) { throw_named(env, "java/lang/IllegalArgumentException", "features must be >= 1"); return NULL; }
  jdoubleArray out = (*env)->NewDoubleArray(env, features * features);
  if (!out) return NULL;
  int64_t used = 0;
  double* o = (double*)pin(env, out);
  int32_t st = oryx_gram_sharded((oryx_model*)(intptr_t)model, (oryx_exchange*)(intptr_t)exchange, which, o, &used);
  if (o) (*env)->ReleasePrimitiveArrayCritical(env, out, o, 0);
  if (st != ORYX_OK) { throw_for_status(env, st); return NULL; }
  return out;
}

JNIEXPORT jlong JNICALL FN(solverCreate)(JNIEnv* env, jclass cls, jdoubleArray gram, jint features, jint device) {
  (void)cls;
  oryx_solver* s = NULL;
  int32_t rank = 0;
  double* g = (double*)pin(env, gram);
  int32_t st = oryx_solver_create(g, features, device, &s, &rank);
  unpin(env, gram, g);
  if (st == ORYX_E_SINGULAR) {
    /* SingularMatrixSolverException(int apparentRank, String message), LinearSystemSolver.java:47-54 */
    char msg[512];
    msg[0] = 0;
    oryx_last_error(msg, (int32_t)sizeof(msg));
    jclass ex = (*env)->FindClass(env, "com/cloudera/oryx/common/math/SingularMatrixSolverException");
    jmethodID ctor = ex ? (*env)->GetMethodID(env, ex, "<init>", "(ILjava/lang/String;)V") : NULL;
    jstring jmsg = (*env)->NewStringUTF(env, msg);
    jobject exc = ctor ? (*env)->NewObject(env, ex, ctor, (jint)rank, jmsg) : NULL;
    if (exc) (*env)->Throw(env, (jthrowable)exc);
    return 0;
  }
  if (st != ORYX_OK) { throw_for_status(env, st); return 0; }
  return (jlong)(intptr_t)s;
}

JNIEXPORT void JNICALL FN(solverDestroy)(JNIEnv* env, jclass cls, jlong solver) {
  (void)env; (void)cls;
  oryx_solver_destroy((oryx_solver*)(intptr_t)solver);
}

JNIEXPORT jfloatArray JNICALL FN(solveFToF)(JNIEnv* env, jclass cls, jlong solver, jfloatArray b) {
  (void)cls;
  const jsize n = len(env, b);
  jfloatArray out = (*env)->NewFloatArray(env, n);
  if (!out) return NULL;
  float* bp = (float*)pin(env, b);
  float* xp = (float*)pin(env, out);
  int32_t st = oryx_solve_f2f((oryx_solver*)(intptr_t)solver, bp, xp);
  if (xp) (*env)->ReleasePrimitiveArrayCritical(env, out, xp, 0);
  unpin(env, b, bp);
  if (st != ORYX_OK) { throw_for_status(env, st); return NULL; }
  return out;
}

/* float[n][f] (rows may be null) -> packed float[n * f] + has[n] */
static int pack_rows(JNIEnv* env, jobjectArray rows, jsize n, int f, float* packed, uint8_t* has) {
  for (jsize i = 0; i < n; i++) {
    jfloatArray r = rows ? (jfloatArray)(*env)->GetObjectArrayElement(env, rows, i) : NULL;
    if (r) {
      if (len(env, r) != f) { (*env)->DeleteLocalRef(env, r); return -1; }
      (*env)->GetFloatArrayRegion(env, r, 0, f, packed + (size_t)i * (size_t)f);
      (*env)->DeleteLocalRef(env, r);
      if (has) has[i] = 1;
    } else {
      memset(packed + (size_t)i * (size_t)f, 0, sizeof(float) * (size_t)f);
      if (has) has[i] = 0;
    }
  }
  return 0;
}

/* packed float[n * f] + valid[n] -> float[n][] with null where valid[i] == 0 */
static jobjectArray unpack_rows(JNIEnv* env, const float* packed, const uint8_t* valid, jsize n, int f) {
  jclass fa = (*env)->FindClass(env, "[F");
  jobjectArray out = fa ? (*env)->NewObjectArray(env, n, fa, NULL) : NULL;
  if (!out) return NULL;
  for (jsize i = 0; i < n; i++) {
    if (!valid[i]) continue;
    jfloatArray r = new_floats(env, packed + (size_t)i * (size_t)f, f);
    (*env)->SetObjectArrayElement(env, out, i, r);
    (*env)->DeleteLocalRef(env, r);
  }
  return out;
}

JNIEXPORT jobjectArray JNICALL FN(foldInBatch)(JNIEnv* env, jclass cls, jlong solver, jboolean implicit, jobjectArray xu,
                                               jobjectArray yi, jdoubleArray values) {
  (void)cls;
  const jsize n = len(env, values);
  if (n == 0) return unpack_rows(env, NULL, NULL, 0, 0);
  jfloatArray first = (jfloatArray)(*env)->GetObjectArrayElement(env, yi, 0);
  const int f = (int)len(env, first);
  if (first) (*env)->DeleteLocalRef(env, first);
  float* pxu = (float*)malloc(sizeof(float) * (size_t)n * (size_t)f);
  float* pyi = (float*)malloc(sizeof(float) * (size_t)n * (size_t)f);
  float* res = (float*)malloc(sizeof(float) * (size_t)n * (size_t)f);
  uint8_t* has = (uint8_t*)malloc((size_t)n);
  uint8_t* valid = (uint8_t*)malloc((size_t)n);
  jobjectArray out = NULL;
  if (!pxu || !pyi || !res || !has || !valid) {
    throw_named(env, "java/lang/OutOfMemoryError", "foldInBatch");
  } else if (pack_rows(env, xu, n, f, pxu, has) != 0 || pack_rows(env, yi, n, f, pyi, NULL) != 0) {
    throw_named(env, "java/lang/IllegalArgumentException", "vectors differ in length");
  } else {
    double* v = (double*)pin(env, values);
    int32_t st = oryx_fold_in_batch((oryx_solver*)(intptr_t)solver, implicit ? 1 : 0, pxu, has, pyi, v, n, res, valid);
    unpin(env, values, v);
    if (st != ORYX_OK) throw_for_status(env, st);
    else out = unpack_rows(env, res, valid, n, f);
  }
  free(pxu); free(pyi); free(res); free(has); free(valid);
  return out;
}

JNIEXPORT jobjectArray JNICALL FN(foldInRows)(JNIEnv* env, jclass cls, jlong solver, jlong model, jint features,
                                               jboolean implicit, jint xu_which, jintArray xu_rows, jintArray yi_rows,
                                               jdoubleArray values) {
  (void)cls;
  const jsize n = len(env, values);
  if (len(env, xu_rows) != n || len(env, yi_rows) != n) {
    throw_named(env, "java/lang/IllegalArgumentException", "row / value arrays differ in length");
    return NULL;
  }
  float* res = (float*)malloc(sizeof(float) * (size_t)(n > 0 ? n : 1) * (size_t)features);
  uint8_t* valid = (uint8_t*)calloc((size_t)(n > 0 ? n : 1), 1);
  jobjectArray out = NULL;
  if (!res || !valid) {
    throw_named(env, "java/lang/OutOfMemoryError", "foldInRows");
  } else {
    int32_t* xr = (int32_t*)pin(env, xu_rows);
    int32_t* yr = (int32_t*)pin(env, yi_rows);
    double* v = (double*)pin(env, values);
    int32_t st = oryx_fold_in_rows((oryx_solver*)(intptr_t)solver, (oryx_model*)(intptr_t)model, implicit ? 1 : 0, xu_which,
                                   xr, yr, v, n, res, valid);
    unpin(env, values, v);
    unpin(env, yi_rows, yr);
    unpin(env, xu_rows, xr);
    if (st != ORYX_OK) throw_for_status(env, st);
    else out = unpack_rows(env, res, valid, n, features);
  }
  free(res); free(valid);
  return out;
}

/* ---------------------------------------------------------------- several GPUs in one JVM */

JNIEXPORT jlong JNICALL FN(exchangeCreate)(JNIEnv* env, jclass cls, jint device, jint world, jint rank, jint max_how_many) {
  (void)cls;
  oryx_exchange* x = NULL;
  int32_t st = oryx_exchange_create(device, world, rank, max_how_many, &x);
  if (st != ORYX_OK) { throw_for_status(env, st); return 0; }
  return (jlong)(intptr_t)x;
}

JNIEXPORT void JNICALL FN(exchangeConnectLocal)(JNIEnv* env, jclass cls, jlongArray exchanges) {
  (void)cls;
  const jsize world = len(env, exchanges);
  if (world < 1 || world > ORYX_MAX_WORLD) { throw_named(env, "java/lang/IllegalArgumentException", "world out of range"); return; }
  jlong handles[ORYX_MAX_WORLD];
  oryx_exchange* all[ORYX_MAX_WORLD];
  (*env)->GetLongArrayRegion(env, exchanges, 0, world, handles);
  for (jsize i = 0; i < world; i++) all[i] = (oryx_exchange*)(intptr_t)handles[i];
  int32_t st = oryx_exchange_connect_local(all, world);
  if (st != ORYX_OK) throw_for_status(env, st);
}

JNIEXPORT void JNICALL FN(exchangeDestroy)(JNIEnv* env, jclass cls, jlong exchange) {
  (void)env; (void)cls;
  oryx_exchange_destroy((oryx_exchange*)(intptr_t)exchange);
}

/* ---- stores of other processes (one JVM per GPU): oryx_peer_stores_* + oryx_fold_in_rows_peers */

JNIEXPORT jlong JNICALL FN(peerStoresCreate)(JNIEnv* env, jclass cls, jlong model, jint world, jint rank) {
  (void)cls;
  oryx_peer_stores* p = NULL;
  int32_t st = oryx_peer_stores_create((oryx_model*)(intptr_t)model, world, rank, &p);
  if (st != ORYX_OK) { throw_for_status(env, st); return 0; }
  return (jlong)(intptr_t)p;
}

JNIEXPORT jbyteArray JNICALL FN(peerStoresHandle)(JNIEnv* env, jclass cls, jlong peers) {
  (void)cls;
  jbyte buf[ORYX_PEER_STORES_HANDLE_BYTES];
  int32_t st = oryx_peer_stores_handle((oryx_peer_stores*)(intptr_t)peers, buf);
  if (st != ORYX_OK) { throw_for_status(env, st); return NULL; }
  jbyteArray out = (*env)->NewByteArray(env, ORYX_PEER_STORES_HANDLE_BYTES);
  if (out) (*env)->SetByteArrayRegion(env, out, 0, ORYX_PEER_STORES_HANDLE_BYTES, buf);
  return out;
}

JNIEXPORT void JNICALL FN(peerStoresConnect)(JNIEnv* env, jclass cls, jlong peers, jbyteArray handles_in_rank_order) {
  (void)cls;
  void* h = pin(env, handles_in_rank_order);
  int32_t st = h ? oryx_peer_stores_connect((oryx_peer_stores*)(intptr_t)peers, h) : ORYX_E_INVALID;
  unpin(env, handles_in_rank_order, h);
  if (!h) throw_named(env, "java/lang/IllegalArgumentException", "null handle table");
  else if (st != ORYX_OK) throw_for_status(env, st);
}

JNIEXPORT void JNICALL FN(peerStoresDestroy)(JNIEnv* env, jclass cls, jlong peers) {
  (void)env; (void)cls;
  oryx_peer_stores_destroy((oryx_peer_stores*)(intptr_t)peers);
}

JNIEXPORT jobjectArray JNICALL FN(foldInRowsPeers)(JNIEnv* env, jclass cls, jlong solver, jlong peers, jint features,
                                                    jlong xu_row_stride, jlong yi_row_stride, jboolean implicit,
                                                    jint xu_which, jlongArray xu_rows, jlongArray yi_rows,
                                                    jdoubleArray values) {
  (void)cls;
  const jsize n = len(env, values);
  if (len(env, xu_rows) != n || len(env, yi_rows) != n) {
    throw_named(env, "java/lang/IllegalArgumentException", "row / value arrays differ in length");
    return NULL;
  }
  float* res = (float*)malloc(sizeof(float) * (size_t)(n > 0 ? n : 1) * (size_t)features);
  uint8_t* valid = (uint8_t*)calloc((size_t)(n > 0 ? n : 1), 1);
  jobjectArray out = NULL;
  if (!res || !valid) {
    throw_named(env, "java/lang/OutOfMemoryError", "foldInRowsPeers");
  } else {
    int64_t* xr = (int64_t*)pin(env, xu_rows);
    int64_t* yr = (int64_t*)pin(env, yi_rows);
    double* v = (double*)pin(env, values);
    int32_t st = oryx_fold_in_rows_peers((oryx_solver*)(intptr_t)solver, (oryx_peer_stores*)(intptr_t)peers, xu_row_stride,
                                         yi_row_stride, implicit ? 1 : 0, xu_which, xr, yr, v, n, res, valid);
    unpin(env, values, v);
    unpin(env, yi_rows, yr);
    unpin(env, xu_rows, xr);
    if (st != ORYX_OK) throw_for_status(env, st);
    else out = unpack_rows(env, res, valid, n, features);
  }
  free(res); free(valid);
  return out;
}

/* ---- ALSSpeedModelManager.buildUpdates aggregation (SMM:141-171) on the model's GPU.  Returns the number of
 * surviving (user, item) groups; the three output arrays (length >= n) receive them ordered by (user, item). */
JNIEXPORT jlong JNICALL FN(aggregateInteractions)(JNIEnv* env, jclass cls, jlong model, jintArray user_rows,
                                                   jintArray item_rows, jdoubleArray strengths, jlongArray timestamps,
                                                   jboolean implicit, jintArray out_user_rows, jintArray out_item_rows,
                                                   jdoubleArray out_strengths) {
  (void)cls;
  const jsize n = len(env, strengths);
  if (len(env, user_rows) != n || len(env, item_rows) != n || len(env, timestamps) != n || len(env, out_user_rows) < n ||
      len(env, out_item_rows) < n || len(env, out_strengths) < n) {
    throw_named(env, "java/lang/IllegalArgumentException", "input arrays differ in length or outputs are too short");
    return 0;
  }
  int64_t out_n = 0;
  int32_t* ur = (int32_t*)pin(env, user_rows);
  int32_t* ir = (int32_t*)pin(env, item_rows);
  double* sv = (double*)pin(env, strengths);
  int64_t* ts = (int64_t*)pin(env, timestamps);
  int32_t* our = (int32_t*)(*env)->GetPrimitiveArrayCritical(env, out_user_rows, NULL);
  int32_t* oir = (int32_t*)(*env)->GetPrimitiveArrayCritical(env, out_item_rows, NULL);
  double* osv = (double*)(*env)->GetPrimitiveArrayCritical(env, out_strengths, NULL);
  int32_t st = ORYX_E_NOMEM;
  if (our && oir && osv && (n == 0 || (ur && ir && sv && ts)))
    st = oryx_aggregate_interactions((oryx_model*)(intptr_t)model, ur, ir, sv, ts, n, implicit ? 1 : 0, our, oir, osv, &out_n);
  if (osv) (*env)->ReleasePrimitiveArrayCritical(env, out_strengths, osv, 0);  /* 0: copy back */
  if (oir) (*env)->ReleasePrimitiveArrayCritical(env, out_item_rows, oir, 0);
  if (our) (*env)->ReleasePrimitiveArrayCritical(env, out_user_rows, our, 0);
  unpin(env, timestamps, ts);
  unpin(env, strengths, sv);
  unpin(env, item_rows, ir);
  unpin(env, user_rows, ur);
  if (st != ORYX_OK) { throw_for_status(env, st); return 0; }
  return (jlong)out_n;
}

JNIEXPORT jlongArray JNICALL FN(stats)(JNIEnv* env, jclass cls, jlong model) {
  (void)cls;
  oryx_stats s;
  int32_t st = oryx_stats_get((oryx_model*)(intptr_t)model, &s);
  if (st != ORYX_OK) { throw_for_status(env, st); return NULL; }
  jlong v[7] = {s.rows[0], s.rows[1], s.slots[0], s.slots[1], s.device_bytes, s.queries, s.kernel_launches};
  jlongArray out = (*env)->NewLongArray(env, 7);
  if (out) (*env)->SetLongArrayRegion(env, out, 0, 7, v);
  return out;
}
