// ingest_c.cpp -- the extern "C" face of update_stream.hpp (include/oryx_ingest.h).
#include <cstring>
#include <string>

#include "oryx_ingest.h"
#include "update_stream.hpp"

using namespace oryx::host;

namespace {
thread_local std::string g_err;

template <typename Fn>
int32_t guarded(Fn fn) {
  try {
    return fn();
  } catch (const BadMessage& e) {
    g_err = e.what();
    return ORYX_E_INVALID;
  } catch (const std::invalid_argument& e) {
    g_err = e.what();
    return ORYX_E_INVALID;
  } catch (const std::out_of_range& e) {
    g_err = e.what();
    return ORYX_E_INVALID;
  } catch (const std::bad_alloc&) {
    g_err = "out of host memory";
    return ORYX_E_NOMEM;
  } catch (const std::exception& e) {
    g_err = e.what();
    return ORYX_E_CUDA;
  }
}

std::vector<KeyMessage> toKeyMessages(const char* keys, const int64_t* kb, const char* msgs, const int64_t* mb, int64_t n) {
  std::vector<KeyMessage> v((size_t)n);
  for (int64_t i = 0; i < n; i++) {
    v[(size_t)i].key.assign(keys + kb[i], (size_t)(kb[i + 1] - kb[i]));
    v[(size_t)i].message.assign(msgs + mb[i], (size_t)(mb[i + 1] - mb[i]));
  }
  return v;
}
}  // namespace

struct oryx_serving_manager {
  ALSServingModelManager impl;
};
struct oryx_speed_manager {
  ALSSpeedModelManager impl;
};

extern "C" {

int32_t oryx_parse_feature_messages(const char* text, const int64_t* begin, int64_t n, int32_t tagged,
                                    int32_t threads, uint8_t* matrix, int64_t* vec_begin, float* values,
                                    int64_t values_cap, int64_t* id_begin, char* ids, int64_t ids_cap,
                                    uint8_t* has_others, int64_t* other_begin, int64_t* other_id_begin,
                                    int64_t other_count_cap, char* other_ids, int64_t other_ids_cap,
                                    int64_t* n_parsed) {
  return guarded([&]() -> int32_t {
    if (!text || !begin || n < 0 || !n_parsed) { g_err = "null argument"; return ORYX_E_INVALID; }
    FeatureBatch b = parseFeatureMessages(text, begin, n, tagged != 0, threads);
    const int64_t m = (int64_t)b.size();
    *n_parsed = m;
    if ((int64_t)b.values.size() > values_cap || (int64_t)b.others.size() > other_count_cap) {
      g_err = "output capacity too small";
      return ORYX_E_NOMEM;
    }
    int64_t idBytes = 0, otherBytes = 0;
    for (auto& s : b.ids) idBytes += (int64_t)s.size();
    for (auto& s : b.others) otherBytes += (int64_t)s.size();
    if (idBytes > ids_cap || otherBytes > other_ids_cap) { g_err = "output capacity too small"; return ORYX_E_NOMEM; }
    std::memcpy(values, b.values.data(), b.values.size() * sizeof(float));
    int64_t off = 0;
    for (int64_t i = 0; i < m; i++) {
      matrix[i] = (uint8_t)b.matrix[(size_t)i];
      has_others[i] = b.hasOthers[(size_t)i];
      vec_begin[i] = b.vecBegin[(size_t)i];
      other_begin[i] = b.otherBegin[(size_t)i];
      id_begin[i] = off;
      std::memcpy(ids + off, b.ids[(size_t)i].data(), b.ids[(size_t)i].size());
      off += (int64_t)b.ids[(size_t)i].size();
    }
    vec_begin[m] = b.vecBegin[(size_t)m];
    other_begin[m] = b.otherBegin[(size_t)m];
    id_begin[m] = off;
    off = 0;
    for (size_t k = 0; k < b.others.size(); k++) {
      other_id_begin[k] = off;
      std::memcpy(other_ids + off, b.others[k].data(), b.others[k].size());
      off += (int64_t)b.others[k].size();
    }
    other_id_begin[b.others.size()] = off;
    if (b.firstBad >= 0) {
      g_err = "message " + std::to_string(b.firstBad) + ": " + b.firstBadWhy;
      return ORYX_E_INVALID;
    }
    return ORYX_OK;
  });
}

int64_t oryx_format_update(int32_t matrix, const char* id, const float* vector, int32_t features,
                           const char* other, char* out, int64_t cap) {
  std::string s;
  if (matrix == 0) {
    s = toFeatureLine(id, vector, features);
  } else {
    std::string o = other ? other : "";
    s = toUpdateJSON((char)matrix, id, vector, features, other ? &o : nullptr);
  }
  if ((int64_t)s.size() + 1 > cap) return -1;
  std::memcpy(out, s.c_str(), s.size() + 1);
  return (int64_t)s.size();
}

int32_t oryx_float_to_string(float v, char* out) {
  std::string s = TextUtils::floatToString(v);
  std::memcpy(out, s.c_str(), s.size() + 1);
  return (int32_t)s.size();
}
int32_t oryx_double_to_string(double v, char* out) {
  std::string s = TextUtils::doubleToString(v);
  std::memcpy(out, s.c_str(), s.size() + 1);
  return (int32_t)s.size();
}

int32_t oryx_serving_manager_create(double sample_rate, int32_t num_cores, int32_t device, int32_t parse_threads,
                                    oryx_serving_manager** out) {
  return guarded([&]() -> int32_t {
    if (!out) { g_err = "null argument"; return ORYX_E_INVALID; }
    *out = new oryx_serving_manager{ALSServingModelManager(sample_rate, num_cores, device, parse_threads)};
    return ORYX_OK;
  });
}
void oryx_serving_manager_destroy(oryx_serving_manager* mgr) { delete mgr; }

int32_t oryx_serving_manager_consume(oryx_serving_manager* mgr, const char* keys, const int64_t* key_begin,
                                     const char* messages, const int64_t* message_begin, int64_t n) {
  return guarded([&]() -> int32_t {
    if (!mgr || !keys || !key_begin || !messages || !message_begin || n < 0) { g_err = "null argument"; return ORYX_E_INVALID; }
    mgr->impl.consume(toKeyMessages(keys, key_begin, messages, message_begin, n));
    return ORYX_OK;
  });
}

int32_t oryx_serving_manager_consume_updates(oryx_serving_manager* mgr, const char* messages,
                                             const int64_t* begin, int64_t n) {
  return guarded([&]() -> int32_t {
    if (!mgr || !messages || !begin || n < 0) { g_err = "null argument"; return ORYX_E_INVALID; }
    mgr->impl.consumeUpdates(messages, begin, n);
    return ORYX_OK;
  });
}

int32_t oryx_serving_manager_load_features(oryx_serving_manager* mgr, int32_t which, const char* lines,
                                           const int64_t* begin, int64_t n) {
  return guarded([&]() -> int32_t {
    if (!mgr || !lines || !begin || n < 0) { g_err = "null argument"; return ORYX_E_INVALID; }
    if (which != ORYX_X && which != ORYX_Y) { g_err = "which must be ORYX_X or ORYX_Y"; return ORYX_E_INVALID; }
    mgr->impl.loadFeatureLines(which == ORYX_X ? 'X' : 'Y', lines, begin, n);
    return ORYX_OK;
  });
}

int32_t oryx_serving_manager_model(oryx_serving_manager* mgr, oryx_model** model, int32_t* features,
                                   int32_t* implicit, int64_t* n_users, int64_t* n_items, float* fraction_loaded) {
  return guarded([&]() -> int32_t {
    if (!mgr || !model) { g_err = "null argument"; return ORYX_E_INVALID; }
    ALSServingModel* m = mgr->impl.getModel();
    *model = m ? m->handle() : nullptr;
    if (features) *features = m ? m->getFeatures() : 0;
    if (implicit) *implicit = m && m->isImplicit() ? 1 : 0;
    if (n_users) *n_users = m ? m->getNumUsers() : 0;
    if (n_items) *n_items = m ? m->getNumItems() : 0;
    if (fraction_loaded) *fraction_loaded = m ? m->getFractionLoaded() : 0.0f;
    return ORYX_OK;
  });
}

int32_t oryx_serving_manager_item_row(oryx_serving_manager* mgr, const char* item, int32_t* row) {
  return guarded([&]() -> int32_t {
    if (!mgr || !item || !row) { g_err = "null argument"; return ORYX_E_INVALID; }
    ALSServingModel* m = mgr->impl.getModel();
    *row = m ? m->itemRow(item) : -1;
    return ORYX_OK;
  });
}

int32_t oryx_speed_manager_create(int32_t no_known_items, double min_model_load_fraction, int32_t device,
                                  int32_t parse_threads, oryx_speed_manager** out) {
  return guarded([&]() -> int32_t {
    if (!out) { g_err = "null argument"; return ORYX_E_INVALID; }
    *out = new oryx_speed_manager{ALSSpeedModelManager(no_known_items != 0, min_model_load_fraction, device, parse_threads)};
    return ORYX_OK;
  });
}
void oryx_speed_manager_destroy(oryx_speed_manager* mgr) { delete mgr; }

int32_t oryx_speed_manager_consume(oryx_speed_manager* mgr, const char* keys, const int64_t* key_begin,
                                   const char* messages, const int64_t* message_begin, int64_t n) {
  return guarded([&]() -> int32_t {
    if (!mgr || !keys || !key_begin || !messages || !message_begin || n < 0) { g_err = "null argument"; return ORYX_E_INVALID; }
    mgr->impl.consume(toKeyMessages(keys, key_begin, messages, message_begin, n));
    return ORYX_OK;
  });
}

int32_t oryx_speed_manager_build_updates(oryx_speed_manager* mgr, const char* lines, const int64_t* begin,
                                         int64_t n, char* out, int64_t out_cap, int64_t* out_begin,
                                         int64_t out_begin_cap, int64_t* n_out, int64_t* out_len) {
  return guarded([&]() -> int32_t {
    if (!mgr || !lines || !begin || n < 0 || !n_out || !out_len) { g_err = "null argument"; return ORYX_E_INVALID; }
    std::vector<std::string> in((size_t)n);
    for (int64_t i = 0; i < n; i++) in[(size_t)i].assign(lines + begin[i], (size_t)(begin[i + 1] - begin[i]));
    auto ups = mgr->impl.buildUpdates(in);
    int64_t total = 0;
    for (auto& s : ups) total += (int64_t)s.size();
    *n_out = (int64_t)ups.size();
    *out_len = total;
    if (total > out_cap || (int64_t)ups.size() + 1 > out_begin_cap || !out || !out_begin) {
      g_err = "output capacity too small";
      return ORYX_E_NOMEM;
    }
    int64_t off = 0;
    for (size_t i = 0; i < ups.size(); i++) {
      out_begin[i] = off;
      std::memcpy(out + off, ups[i].data(), ups[i].size());
      off += (int64_t)ups[i].size();
    }
    out_begin[ups.size()] = off;
    return ORYX_OK;
  });
}

int32_t oryx_ingest_last_error(char* buf, int32_t cap) {
  if (!buf || cap <= 0) return ORYX_E_INVALID;
  std::strncpy(buf, g_err.c_str(), (size_t)cap - 1);
  buf[cap - 1] = 0;
  return ORYX_OK;
}

}  // extern "C"
