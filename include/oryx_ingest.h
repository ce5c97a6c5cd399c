/*
 * oryx_ingest.h -- C ABI of the update-stream side of the ALS path (liboryx_host.so): the text
 * formats that feed and leave the hot path, and the two model managers that turn the update topic
 * into bulk device upserts.  Same conventions as oryx_b200.h:
 * int32 status, caller-owned host buffers, no exceptions across the boundary,
 * oryx_ingest_last_error() for the calling thread's last failure.
 *
 * Citations, relative to the reference root:
 *   SMGR = app/oryx-app-serving/src/main/java/com/cloudera/oryx/app/serving/als/model/ALSServingModelManager.java
 *   SMM  = app/oryx-app/src/main/java/com/cloudera/oryx/app/speed/als/ALSSpeedModelManager.java
 *   TU   = framework/oryx-common/src/main/java/com/cloudera/oryx/common/text/TextUtils.java
 *   ALSUP= app/oryx-app-mllib/src/main/java/com/cloudera/oryx/app/batch/mllib/als/ALSUpdate.java
 *
 * A "message array" is n strings laid back to back in `text`; string i is text[begin[i] .. begin[i+1]).
 */
#ifndef ORYX_INGEST_H_
#define ORYX_INGEST_H_

#include <stdint.h>

#include "oryx_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct oryx_serving_manager oryx_serving_manager;
typedef struct oryx_speed_manager oryx_speed_manager;

/* Replaces TU:170-176 readJSON(message, List.class) + TU:186-188 convertViaJSON(.., float[].class) as
 * SMGR:76-79 / SMM:87-90 apply them to "UP" payloads ["X"|"Y", id, [v...], [ids...]?] (tagged = 1), and
 * ALSUP:465-467 to part-file lines [id, [v...]] (tagged = 0); parsed on `threads` host threads
 * (0 = all).  Outputs for message i < *n_parsed: matrix[i] ('X'/'Y', or 0 untagged); its vector
 * values[vec_begin[i] .. vec_begin[i+1]); its id ids[id_begin[i] .. id_begin[i+1]); has_others[i];
 * its other ids k in [other_begin[i], other_begin[i+1]) as other_ids[other_id_begin[k] .. other_id_begin[k+1]).
 * *n_parsed < n means message *n_parsed is malformed (status ORYX_E_INVALID, reason in last_error).
 * Capacities are in elements; ORYX_E_NOMEM when one is too small. */
int32_t oryx_parse_feature_messages(const char* text, const int64_t* begin, int64_t n, int32_t tagged,
                                    int32_t threads, uint8_t* matrix, int64_t* vec_begin, float* values,
                                    int64_t values_cap, int64_t* id_begin, char* ids, int64_t ids_cap,
                                    uint8_t* has_others, int64_t* other_begin, int64_t* other_id_begin,
                                    int64_t other_count_cap, char* other_ids, int64_t other_ids_cap,
                                    int64_t* n_parsed);
/* Replaces SMM:207-215 toUpdateJSON (other == NULL <=> oryx.als.no-known-items) and, with matrix == 0,
 * the line ALSUP:420-426 writes.  Returns the length written (excluding the NUL), or -1 if cap is too small. */
int64_t oryx_format_update(int32_t matrix, const char* id, const float* vector, int32_t features,
                           const char* other, char* out, int64_t cap);
/* Float.toString / Double.toString as Jackson emits them.  `out` needs 32 bytes. */
int32_t oryx_float_to_string(float v, char* out);
int32_t oryx_double_to_string(double v, char* out);

/* --- ALSServingModelManager (SMGR:46-61 constructor, :64-129 consume, :131-134 getModel) --- */
int32_t oryx_serving_manager_create(double sample_rate, int32_t num_cores, int32_t device, int32_t parse_threads,
                                    oryx_serving_manager** out);
void oryx_serving_manager_destroy(oryx_serving_manager* mgr);
/* consume(): n (key, message) pairs in arrival order. */
int32_t oryx_serving_manager_consume(oryx_serving_manager* mgr, const char* keys, const int64_t* key_begin,
                                     const char* messages, const int64_t* message_begin, int64_t n);
/* A run of "UP" payloads only (the bulk of a model load); ignored while there is no model (SMGR:73-75). */
int32_t oryx_serving_manager_consume_updates(oryx_serving_manager* mgr, const char* messages,
                                             const int64_t* begin, int64_t n);
/* Bulk load of part-file lines (ALSUP:415-428 format) into store `which` (ORYX_X / ORYX_Y). */
int32_t oryx_serving_manager_load_features(oryx_serving_manager* mgr, int32_t which, const char* lines,
                                           const int64_t* begin, int64_t n);
/* getModel(): *model is NULL until the first MODEL message; it stays owned by the manager. */
int32_t oryx_serving_manager_model(oryx_serving_manager* mgr, oryx_model** model, int32_t* features,
                                   int32_t* implicit, int64_t* n_users, int64_t* n_items, float* fraction_loaded);
/* The String<->row map the manager keeps for the Y store: *row = -1 when the item is unknown. */
int32_t oryx_serving_manager_item_row(oryx_serving_manager* mgr, const char* item, int32_t* row);

/* --- ALSSpeedModelManager (SMM:60-64 constructor, :66-132 consume, :135-205 buildUpdates) --- */
int32_t oryx_speed_manager_create(int32_t no_known_items, double min_model_load_fraction, int32_t device,
                                  int32_t parse_threads, oryx_speed_manager** out);
void oryx_speed_manager_destroy(oryx_speed_manager* mgr);
int32_t oryx_speed_manager_consume(oryx_speed_manager* mgr, const char* keys, const int64_t* key_begin,
                                   const char* messages, const int64_t* message_begin, int64_t n);
/* buildUpdates(): n input lines "user,item,strength,timestamp" (CSV or JSON array) -> "UP" payloads,
 * written back to back into `out` with out_begin[0..*n_out] (capacities in elements; ORYX_E_NOMEM
 * reports the needed sizes in *n_out / *out_len). */
int32_t oryx_speed_manager_build_updates(oryx_speed_manager* mgr, const char* lines, const int64_t* begin,
                                         int64_t n, char* out, int64_t out_cap, int64_t* out_begin,
                                         int64_t out_begin_cap, int64_t* n_out, int64_t* out_len);

int32_t oryx_ingest_last_error(char* buf, int32_t cap);

#ifdef __cplusplus
}
#endif
#endif /* ORYX_INGEST_H_ */
