// update_stream.hpp -- the text formats either side of the ALS hot path, and the two model managers
// that turn the update topic into batched device upserts.
//
// Mirrors, by name and behaviour:
//   TextUtils      framework/oryx-common/src/main/java/com/cloudera/oryx/common/text/TextUtils.java
//                  parseDelimited :56-58, parsePMMLDelimited :64-75, parseJSONArray :148-150,
//                  joinJSON :156-162, readJSON :170-176 + convertViaJSON :186-188 as used on "UP" messages
//   MLFunctions    app/oryx-app-common/src/main/java/com/cloudera/oryx/app/common/fn/MLFunctions.java:39-78
//   AppPMMLUtils   app/oryx-app-common/src/main/java/com/cloudera/oryx/app/pmml/AppPMMLUtils.java:62-84, 256-283
//   ALSServingModelManager.consume  app/oryx-app-serving/.../als/model/ALSServingModelManager.java:64-129
//   ALSSpeedModelManager.consume / buildUpdates / toUpdateJSON
//                  app/oryx-app/src/main/java/com/cloudera/oryx/app/speed/als/ALSSpeedModelManager.java:66-132,
//                  135-205, 207-215
//   feature part files ("[id,[v0,v1,...]]" per line)
//                  app/oryx-app-mllib/.../batch/mllib/als/ALSUpdate.java:415-428 (writer), :461-470 (reader)
//
// Third-party behaviour restated here (not under /root/reference): Jackson databind 2.x
// (ObjectMapper.readValue / convertValue / writeValueAsString) for the JSON grammar, the
// number -> Double -> float narrowing and Float.toString number output; Apache commons-csv
// (CSVFormat.RFC4180.withEscape('\\')) for the delimited formats.
#pragma once

#include <cstdint>
#include <memory>
#include <stdexcept>
#include <string>
#include <string_view>
#include <unordered_map>
#include <vector>

#include "oryx_host.hpp"

namespace oryx {
namespace host {

// What Java reports as IOException / IllegalStateException / IllegalArgumentException("Bad message").
class BadMessage : public std::runtime_error {
 public:
  using std::runtime_error::runtime_error;
};

namespace TextUtils {
std::vector<std::string> parseDelimited(std::string_view delimited, char delimiter);  // TextUtils.java:56-58
std::vector<std::string> parsePMMLDelimited(std::string_view delimited);              // :64-75
std::vector<std::string> parseJSONArray(std::string_view json);                       // :148-150
std::string joinJSON(const std::vector<std::string>& elements);                       // :156-162 (strings)
// Java's Float.toString / Double.toString (what Jackson emits for float / double values).
std::string floatToString(float v);
std::string doubleToString(double v);
}  // namespace TextUtils

namespace MLFunctions {
std::vector<std::string> parse(std::string_view line);  // PARSE_FN, MLFunctions.java:39-53
int64_t toTimestamp(std::string_view line);             // TO_TIMESTAMP_FN :58-65 (throws std::out_of_range / BadMessage)
double sumWithNaN(const std::vector<double>& orderedStrengths);  // SUM_WITH_NAN :67-78
}  // namespace MLFunctions

// The name -> value / name -> content view AppPMMLUtils gives of a model's top-level <Extension>s.
struct PMMLExtensions {
  std::unordered_map<std::string, std::string> values;                 // getExtensionValue, AppPMMLUtils.java:62-65
  std::unordered_map<std::string, std::vector<std::string>> contents;  // getExtensionContent, :73-84
  bool hasValue(const std::string& name) const { return values.count(name) != 0; }
  static PMMLExtensions fromString(std::string_view pmmlXML);
  // readPMMLFromUpdateKeyMessage :256-283: "MODEL" carries the document, "MODEL-REF" a path to it
  // (a local file here; no Hadoop FileSystem exists on this side).
  static PMMLExtensions fromUpdateKeyMessage(const std::string& key, const std::string& message);
};

// n feature-vector messages parsed into flat arrays.  `tagged` messages are update-topic "UP" payloads
// ["X"|"Y", id, [v...], [other ids...]?]; untagged ones are part-file lines [id, [v...]].
struct FeatureBatch {
  std::vector<char> matrix;          // 'X' / 'Y' (tagged) or 0
  std::vector<std::string> ids;      // update.get(1).toString()
  std::vector<int64_t> vecBegin;     // n+1 offsets into `values`
  std::vector<float> values;         // all vectors back to back
  std::vector<uint8_t> hasOthers;    // update.size() > 3
  std::vector<int64_t> otherBegin;   // n+1 offsets into `others`
  std::vector<std::string> others;   // known item / user ids
  int64_t firstBad = -1;             // index of the first message that failed to parse (later ones untouched)
  std::string firstBadWhy;
  size_t size() const { return ids.size(); }
  int vecLen(size_t i) const { return (int)(vecBegin[i + 1] - vecBegin[i]); }
  const float* vec(size_t i) const { return values.data() + vecBegin[i]; }
};

// Parses messages [begin[i], begin[i+1]) of `text`, i < n, on `threads` host threads (0 = all cores).
// Messages after the first malformed one are not reported (consume() stops there, as the reference's
// loop does when readJSON throws).
FeatureBatch parseFeatureMessages(const char* text, const int64_t* begin, int64_t n, bool tagged, int threads = 0);
// The same without the final concatenation: one batch per parsing thread, in message order (what the
// managers consume; avoids a serial copy of every vector).
std::vector<FeatureBatch> parseFeatureMessageParts(const char* text, const int64_t* begin, int64_t n, bool tagged,
                                                   int threads = 0);
// One message; throws BadMessage.
void parseFeatureMessage(std::string_view message, bool tagged, char* matrix, std::string* id,
                         std::vector<float>* vector, bool* hasOthers, std::vector<std::string>* others);
// ALSSpeedModelManager.toUpdateJSON :207-215 (other == nullptr <=> oryx.als.no-known-items).
std::string toUpdateJSON(char matrix, const std::string& id, const float* vector, int features,
                         const std::string* other);
// ALSUpdate.saveFeaturesRDD line :420-426 with a string key.
std::string toFeatureLine(const std::string& id, const float* vector, int features);

struct KeyMessage {
  std::string key, message;
};

// ALSServingModelManager.java.  consume() gives the same final model state as feeding the messages to
// the reference one by one; runs of "UP" messages are parsed on all cores and reach the device as bulk
// upserts (one C ABI call per run of up to `maxBatchRows` item vectors).
class ALSServingModelManager {
 public:
  ALSServingModelManager(double sampleRate, int numCores = 0, int device = 0, int parseThreads = 0,
                         int64_t maxBatchRows = 1 << 20);
  void consume(const std::vector<KeyMessage>& updates);                     // :64-129
  // Same, for a run of "UP" payloads laid out back to back (what a JNI caller hands over).
  void consumeUpdates(const char* text, const int64_t* begin, int64_t n);
  // Bulk load of a features part file's lines into X ('X') or Y ('Y').
  void loadFeatureLines(char matrix, const char* text, const int64_t* begin, int64_t n);
  ALSServingModel* getModel() { return model_.get(); }                      // :131-134
  bool isReadOnly() const { return false; }

 private:
  void applyBatch(const FeatureBatch& b, char forceMatrix);
  double sampleRate_;
  int numCores_, device_, parseThreads_;
  int64_t maxBatchRows_;
  std::unique_ptr<ALSServingModel> model_;
};

// ALSSpeedModelManager.java
class ALSSpeedModelManager {
 public:
  ALSSpeedModelManager(bool noKnownItems, double minModelLoadFraction = 0.0, int device = 0, int parseThreads = 0);
  void consume(const std::vector<KeyMessage>& updates);                     // :66-132
  // newData values (CSV or JSON-array lines "user,item,strength,timestamp") -> "UP" payloads, :135-205.
  std::vector<std::string> buildUpdates(const std::vector<std::string>& newDataLines);
  // The aggregation half on its own (:141-171): sort by timestamp, per-(user,item) sum-with-NaN for an
  // implicit model / last-wins for an explicit one, NaN results dropped.  Order of the result: first
  // appearance of the pair in timestamp order.
  std::vector<UserItemStrength> aggregate(const std::vector<std::string>& newDataLines, bool implicit) const;
  ALSSpeedModel* getModel() { return model_.get(); }

 private:
  bool noKnownItems_;
  double minModelLoadFraction_;
  int device_, parseThreads_;
  std::unique_ptr<ALSSpeedModel> model_;
};

}  // namespace host
}  // namespace oryx
