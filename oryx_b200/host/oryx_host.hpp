// oryx_host.hpp -- host-side mirror, in C++, of the reference's Java model classes on the ALS hot
// path.  The reference is Java and no JVM exists in this build image, so the layer that sits above
// the C ABI (include/oryx_b200.h) is written in C++ with the reference's names, argument meaning and
// error behaviour; a maintainer of the Java tree replaces the method BODIES of the cited classes with
// the same calls through JNI.  String ids, known-item sets, the user store, paging
// and callbacks stay on the host exactly as in the reference; every score, hash, Gram, solve and
// fold-in goes to the device through the C ABI.
//
//   LocalitySensitiveHash   app/oryx-app-serving/.../serving/als/model/LocalitySensitiveHash.java
//   DotsFunction            app/oryx-app-serving/.../serving/als/DotsFunction.java
//   CosineAverageFunction   app/oryx-app-serving/.../serving/als/CosineAverageFunction.java
//   ALSServingModel         app/oryx-app-serving/.../serving/als/model/ALSServingModel.java
//   Solver / LinearSystemSolver  framework/oryx-common/.../common/math/{Solver,LinearSystemSolver}.java
//   ALSUtils                app/oryx-app-common/.../app/als/ALSUtils.java
//   ALSSpeedModel / ALSSpeedModelManager (hot section)  app/oryx-app/.../app/speed/als/
#pragma once

#include <cstdint>
#include <functional>
#include <memory>
#include <mutex>
#include <stdexcept>
#include <string>
#include <unordered_map>
#include <unordered_set>
#include <utility>
#include <vector>

#include "oryx_b200.h"

namespace oryx {
namespace host {

// com.cloudera.oryx.common.math.SingularMatrixSolverException
class SingularMatrixSolverException : public std::runtime_error {
 public:
  SingularMatrixSolverException(int apparent_rank, const std::string& msg)
      : std::runtime_error(msg), apparent_rank_(apparent_rank) {}
  int getApparentRank() const { return apparent_rank_; }

 private:
  int apparent_rank_;
};

// ORYX_E_INVALID -> std::invalid_argument (IllegalArgumentException); ORYX_E_CUDA / others ->
// std::runtime_error (IllegalStateException).
void check_status(int32_t status);

using FloatVec = std::vector<float>;
using IDValue = std::pair<std::string, double>;

// LocalitySensitiveHash.java:41-124.  The sizing rule is the reference's; hash vectors are chosen by
// the same greedy "most orthogonal of >= 1000 Gaussian candidates" procedure from a seeded
// std::mt19937_64 (the reference draws from commons-math3 Well19937c, which is not available here).
class LocalitySensitiveHash {
 public:
  LocalitySensitiveHash(double sampleRate, int numFeatures, int numCores, uint64_t seed = 1234567890123456789ULL);
  int getNumHashes() const { return num_hashes_; }
  int getNumPartitions() const { return 1 << num_hashes_; }
  int getMaxBitsDiffering() const { return max_bits_; }
  const std::vector<float>& hashVectors() const { return hash_vectors_; }  // [numHashes x numFeatures]

 private:
  int num_hashes_ = 0, max_bits_ = 0;
  std::vector<float> hash_vectors_;
};

// CosineDistanceSensitiveFunction: a target vector plus the score kind the device applies.
struct ScoreFunction {
  FloatVec target;
  int kind;  // ORYX_DOT or ORYX_COSINE
  const FloatVec& getTargetVector() const { return target; }
};
// DotsFunction.java:29-45
ScoreFunction DotsFunction(const FloatVec& userVector);
ScoreFunction DotsFunction(const std::vector<FloatVec>& userFeaturesVectors);
// CosineAverageFunction.java:29-51
ScoreFunction CosineAverageFunction(const FloatVec& itemFeatureVector);
ScoreFunction CosineAverageFunction(const std::vector<FloatVec>& itemFeatureVectors);

// app/oryx-app-api Rescorer: arbitrary host callbacks.
struct Rescorer {
  std::function<double(const std::string&, double)> rescore;  // NaN = exclude
  std::function<bool(const std::string&)> isFiltered;
};

// Solver.java / LinearSystemSolver.java
class Solver {
 public:
  Solver(const std::vector<double>& gramRowMajor, int features, int device);
  ~Solver();
  Solver(const Solver&) = delete;
  Solver& operator=(const Solver&) = delete;
  FloatVec solveFToF(const FloatVec& b) const;
  oryx_solver* handle() const { return h_; }
  int features() const { return features_; }

 private:
  oryx_solver* h_ = nullptr;
  int features_ = 0;
};

namespace ALSUtils {
// ALSUtils.java:37-59
double computeTargetQui(bool implicit, double value, double currentValue);
// ALSUtils.java:74-106; nullptr arguments / empty result mirror Java's null.
std::unique_ptr<FloatVec> computeUpdatedXu(const Solver& solver, double value, const FloatVec* Xu,
                                           const FloatVec* Yi, bool implicit);
}  // namespace ALSUtils

// ALSServingModel.java
class ALSServingModel {
 public:
  ALSServingModel(int features, bool implicit, double sampleRate, int numCores = 0, int device = 0);
  ~ALSServingModel();
  ALSServingModel(const ALSServingModel&) = delete;
  ALSServingModel& operator=(const ALSServingModel&) = delete;

  int getFeatures() const { return features_; }
  bool isImplicit() const { return implicit_; }
  const LocalitySensitiveHash& lsh() const { return lsh_; }

  std::unique_ptr<FloatVec> getUserVector(const std::string& user) const;           // :138-140
  std::unique_ptr<FloatVec> getItemVector(const std::string& item) const;           // :142-151
  void setUserVector(const std::string& user, const FloatVec& vector);              // :153-159
  void setUserVectors(const std::string* users, const float* vectorsRowMajor, size_t n);  // bulk form, one C ABI call
  void setItemVector(const std::string& item, const FloatVec& vector);              // :161-183
  // bulk form of setItemVector for model load (one C ABI call)
  void setItemVectors(const std::vector<std::string>& items, const std::vector<float>& vectorsRowMajor);
  void setItemVectors(const std::string* items, const float* vectorsRowMajor, size_t n);
  std::unordered_set<std::string> getKnownItems(const std::string& user) const;     // :189-201
  void addKnownItems(const std::string& user, const std::vector<std::string>& items);  // :241-258
  // bulk form: user i gains items[begin[i] .. begin[i+1]); one C ABI call for the device copy
  void addKnownItems(const std::string* users, const int64_t* begin, const std::string* items, size_t nUsers);
  std::unordered_map<std::string, int> getUserCounts() const;                       // :212-225
  std::unordered_map<std::string, int> getItemCounts() const;                       // :227-239
  // Recommend.get without a rescorer (serving/als/Recommend.java:74-114): query vector and known items
  // are taken from device memory (oryx_recommend).  nullptr when the user is unknown (404).
  std::unique_ptr<std::vector<IDValue>> recommend(const std::string& user, int howMany, bool considerKnownItems) const;
  std::unique_ptr<std::vector<std::pair<std::string, FloatVec>>> getKnownItemVectorsForUser(const std::string& user) const;  // :264-287

  // :289-335.  allowed: IDs the predicate rejects (known items); extraPredicate / rescorer: arbitrary
  // host callbacks (may be null).  Returns at most howMany (id, score) pairs, best first.
  std::vector<IDValue> topN(const ScoreFunction& scoreFn, const Rescorer* rescorer, int howMany,
                            const std::unordered_set<std::string>* excludedIDs) const;

  std::shared_ptr<Solver> getYTYSolver();                                           // :357-373
  // IDs set since the previous retain call count as "recent" and survive (FeatureVectors.java:72-79, 131-136)
  void retainRecentAndUserIDs(const std::vector<std::string>& users);               // :382-389
  void retainRecentAndItemIDs(const std::vector<std::string>& items);               // :398-409
  void retainRecentAndKnownItems(const std::vector<std::string>& users, const std::vector<std::string>& items);  // :418-441
  std::vector<std::string> getAllItemIDs() const;                                   // :337-343
  int getNumUsers() const;
  int getNumItems() const;
  float getFractionLoaded() const;                                                  // :462-475
  int itemRow(const std::string& item) const;  // dense device row of an item, -1 when unknown
  int partitionOf(const std::string& item) const;  // LSH partition the device assigned (diagnostics/tests)
  oryx_model* handle() const { return h_; }

 private:
  int rowOf(const std::string& item) const;             // row of an item that HAS a vector, else -1
  int itemRowFor(const std::string& item);              // row of any mentioned item id (created on demand)
  int userRowFor(const std::string& user);
  int features_;
  bool implicit_;
  int device_;
  LocalitySensitiveHash lsh_;
  oryx_model* h_ = nullptr;
  mutable std::mutex mu_;
  std::unordered_map<std::string, FloatVec> X_;
  // Rows are permanent per id (never re-used), so the device-resident known-item lists can refer to
  // them; item_has_vec_ tells whether the row currently holds a vector.
  std::unordered_map<std::string, int> item_row_;
  std::vector<std::string> row_item_;
  std::vector<uint8_t> item_has_vec_;
  std::vector<uint8_t> item_recent_;   // by row: vector set since the last retainRecentAndItemIDs (FeatureVectors recentIDs)
  std::vector<int32_t> item_part_;     // by row: LSH partition the device assigned
  int n_item_vecs_ = 0;
  std::unordered_map<std::string, int> user_row_;
  std::vector<std::string> row_user_;
  std::unordered_map<std::string, std::unordered_set<std::string>> known_items_;
  std::unordered_set<std::string> expected_users_, expected_items_;
  std::unordered_set<std::string> recent_users_;
  bool itemIsRecent(const std::string& item) const;
  std::shared_ptr<Solver> cached_yty_;
};

// One (user, item, strength) interaction after aggregation (ALSSpeedModelManager.java:141-171).
struct UserItemStrength {
  std::string user, item;
  double strength;
};
// One "UP" message: ["X"|"Y", id, vector, [otherID]] (ALSSpeedModelManager.java:196-214).
struct Update {
  char matrix;  // 'X' or 'Y'
  std::string id;
  FloatVec vector;
  std::string other;
};

// ALSSpeedModel.java + the hot section of ALSSpeedModelManager.buildUpdates (:173-204).
class ALSSpeedModel {
 public:
  ALSSpeedModel(int features, bool implicit, int device = 0);
  ~ALSSpeedModel();
  int getFeatures() const { return features_; }
  bool isImplicit() const { return implicit_; }
  void setUserVector(const std::string& user, const FloatVec& vector);
  void setItemVector(const std::string& item, const FloatVec& vector);
  // bulk form for the update stream: which = 'X' or 'Y'; one C ABI call
  void setVectors(char which, const std::string* ids, const float* vectorsRowMajor, size_t n);
  void retainRecentAndUserIDs(const std::vector<std::string>& users);  // ALSSpeedModel.java:98-105
  void retainRecentAndItemIDs(const std::vector<std::string>& items);  // :107-114
  float getFractionLoaded() const;                                      // :127-141
  int getNumUsers() const { return (int)x_.row.size(); }
  oryx_model* handle() const { return h_; }
  int getNumItems() const { return (int)y_.row.size(); }
  std::unique_ptr<FloatVec> getUserVector(const std::string& user) const;
  std::unique_ptr<FloatVec> getItemVector(const std::string& item) const;
  std::shared_ptr<Solver> getXTXSolver();  // ALSSpeedModel.java:118-121
  std::shared_ptr<Solver> getYTYSolver();  // :123-126
  // All interactions read the same model snapshot; the two computeUpdatedXu calls per interaction run
  // as two device batches.  Throws SingularMatrixSolverException like the reference's solver getters.
  std::vector<Update> buildUpdates(const std::vector<UserItemStrength>& input);

 private:
  struct Side {
    std::unordered_map<std::string, int> row;
    std::unordered_map<std::string, FloatVec> vec;  // host copy for fold-in gathers
    std::vector<int> free_rows;
    int next_row = 0;
    std::unordered_set<std::string> recent, expected;
  };
  void set(Side& side, int which, const std::string& id, const FloatVec& v);
  void retain(Side& side, int which, const std::vector<std::string>& ids);
  std::shared_ptr<Solver> solver(int which);
  int features_;
  bool implicit_;
  int device_;
  oryx_model* h_ = nullptr;
  Side x_, y_;
};

// Evaluation.java (app/oryx-app-mllib/.../batch/mllib/als/Evaluation.java) over a model whose X and Y
// live on the device; users and products are the dense rows of those stores.  Predictions use MLlib's
// arithmetic (binary64 products), through oryx_estimate_pairs; pairs whose user or product has no vector
// yield no prediction, as with MatrixFactorizationModel.predict.
struct Rating {
  int user, product;
  double rating;
};
namespace Evaluation {
double rmse(oryx_model* model, const std::vector<Rating>& testData);  // :49-63
// :95-120: for every user about as many random non-positive items as positive ones, drawn from the distinct
// items of the data (the reference draws with its own RandomGenerator; here a seeded std::mt19937_64)
std::vector<std::pair<int, int>> sampleNegatives(const std::vector<Rating>& positiveData, uint64_t seed);
// :70-135 "mean AUC": per user, the share of (positive, negative) pairs ranked correctly
double areaUnderCurve(oryx_model* model, const std::vector<Rating>& positiveData,
                      const std::vector<std::pair<int, int>>& negativeUserProducts);
}  // namespace Evaluation

}  // namespace host
}  // namespace oryx
