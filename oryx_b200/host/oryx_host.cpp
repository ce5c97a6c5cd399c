// oryx_host.cpp -- see oryx_host.hpp.  Host bookkeeping only; all arithmetic on vectors of the model
// goes through the C ABI to the device.
#include "oryx_host.hpp"

#include <algorithm>
#include <cmath>
#include <limits>
#include <random>
#include <thread>

namespace oryx {
namespace host {

void check_status(int32_t status) {
  if (status == ORYX_OK) return;
  char buf[512];
  oryx_last_error(buf, sizeof(buf));
  if (status == ORYX_E_INVALID) throw std::invalid_argument(buf);  // IllegalArgumentException
  throw std::runtime_error(buf);                                   // IllegalStateException
}

namespace {
// VectorMath.dot / norm (VectorMath.java:38-58) for the handful of host-side uses (query-vector
// construction, hash-vector selection): float product, double accumulate, index order.
double vm_dot(const float* x, const float* y, size_t n) {
  double d = 0.0;
  for (size_t i = 0; i < n; i++) {
    volatile float p = x[i] * y[i];
    d += (double)p;
  }
  return d;
}
double vm_norm(const float* x, size_t n) { return std::sqrt(vm_dot(x, x, n)); }
}  // namespace

// ---------------------------------------------------------------- LocalitySensitiveHash
LocalitySensitiveHash::LocalitySensitiveHash(double sampleRate, int numFeatures, int numCores, uint64_t seed) {
  int32_t h = 0, b = 0;
  check_status(oryx_lsh_plan(sampleRate, numCores, &h, &b));  // LocalitySensitiveHash.java:45-74
  num_hashes_ = h;
  max_bits_ = b;
  hash_vectors_.assign((size_t)h * numFeatures, 0.0f);
  std::mt19937_64 rng(seed);
  std::normal_distribution<double> gauss(0.0, 1.0);
  std::vector<float> cand((size_t)numFeatures), best((size_t)numFeatures);
  for (int i = 0; i < h; i++) {  // :80-104: most-orthogonal next random vector
    double bestTotal = std::numeric_limits<double>::infinity();
    int sinceBest = 0;
    bool have = false;
    while (sinceBest < 1000) {
      for (int j = 0; j < numFeatures; j++) cand[j] = (float)gauss(rng);  // VectorMath.randomVectorF
      double newNorm = vm_norm(cand.data(), numFeatures), score = 0.0;
      for (int e = 0; e < i; e++) {  // totalAbsCos :179-186
        const float* ex = &hash_vectors_[(size_t)e * numFeatures];
        score += std::fabs(vm_dot(ex, cand.data(), numFeatures)) / vm_norm(ex, numFeatures) / newNorm;
      }
      if (score < bestTotal) {
        best = cand;
        have = true;
        if (score == 0.0) break;
        bestTotal = score;
        sinceBest = 0;
      } else {
        sinceBest++;
      }
    }
    if (have) std::copy(best.begin(), best.end(), hash_vectors_.begin() + (size_t)i * numFeatures);
  }
}

// ---------------------------------------------------------------- score functions
ScoreFunction DotsFunction(const FloatVec& userVector) { return ScoreFunction{userVector, ORYX_DOT}; }

ScoreFunction DotsFunction(const std::vector<FloatVec>& vecs) {  // DotsFunction.java:34-45
  if (vecs.empty()) throw std::invalid_argument("need at least one user vector");
  FloatVec t(vecs[0].size(), 0.0f);
  for (const auto& v : vecs)
    for (size_t i = 0; i < v.size(); i++) t[i] += v[i];
  for (auto& x : t) x /= (float)vecs.size();
  return ScoreFunction{t, ORYX_DOT};
}

ScoreFunction CosineAverageFunction(const FloatVec& v) {  // CosineAverageFunction.java:29-37
  FloatVec t(v.size(), 0.0f);
  double n = vm_norm(v.data(), v.size());
  for (size_t i = 0; i < v.size(); i++) t[i] = (float)((double)t[i] + (double)v[i] / n);
  return ScoreFunction{t, ORYX_COSINE};
}

ScoreFunction CosineAverageFunction(const std::vector<FloatVec>& vecs) {  // :39-51
  if (vecs.empty()) throw std::invalid_argument("need at least one item vector");
  FloatVec t(vecs[0].size(), 0.0f);
  for (const auto& v : vecs) {
    double n = vm_norm(v.data(), v.size());
    for (size_t i = 0; i < v.size(); i++) t[i] = (float)((double)t[i] + (double)v[i] / n);
  }
  for (auto& x : t) x /= (float)vecs.size();
  return ScoreFunction{t, ORYX_COSINE};
}

// ---------------------------------------------------------------- Solver / ALSUtils
Solver::Solver(const std::vector<double>& gram, int features, int device) : features_(features) {
  int32_t rank = 0;
  int32_t st = oryx_solver_create(gram.data(), features, device, &h_, &rank);
  if (st == ORYX_E_SINGULAR) {
    throw SingularMatrixSolverException(rank, "Apparent rank: " + std::to_string(rank));  // LinearSystemSolver.java:54
  }
  check_status(st);
}
Solver::~Solver() { oryx_solver_destroy(h_); }

FloatVec Solver::solveFToF(const FloatVec& b) const {
  FloatVec x((size_t)features_);
  check_status(oryx_solve_f2f(h_, b.data(), x.data()));
  return x;
}

namespace ALSUtils {
double computeTargetQui(bool implicit, double value, double currentValue) {
  if (implicit) {
    if (value > 0.0f && currentValue < 1.0) {
      double diff = 1.0 - std::max(0.0, currentValue);
      return currentValue + (value / (1.0 + value)) * diff;
    }
    if (value < 0.0f && currentValue > 0.0) {
      double diff = -std::min(1.0, currentValue);
      return currentValue + (value / (value - 1.0)) * diff;
    }
    return std::numeric_limits<double>::quiet_NaN();
  }
  return value;
}

std::unique_ptr<FloatVec> computeUpdatedXu(const Solver& solver, double value, const FloatVec* Xu,
                                           const FloatVec* Yi, bool implicit) {
  if (Yi == nullptr) return nullptr;  // ALSUtils.java:79-81
  const int f = solver.features();
  uint8_t has = Xu ? 1 : 0, valid = 0;
  FloatVec zero((size_t)f, 0.0f), out((size_t)f);
  check_status(oryx_fold_in_batch(solver.handle(), implicit ? 1 : 0, Xu ? Xu->data() : zero.data(), &has,
                                  Yi->data(), &value, 1, out.data(), &valid));
  if (!valid) return nullptr;
  return std::make_unique<FloatVec>(std::move(out));
}
}  // namespace ALSUtils

// ---------------------------------------------------------------- ALSServingModel
static int default_cores(int numCores) {
  if (numCores > 0) return numCores;
  unsigned hc = std::thread::hardware_concurrency();  // Runtime.availableProcessors()
  return hc ? (int)hc : 1;
}

ALSServingModel::ALSServingModel(int features, bool implicit, double sampleRate, int numCores, int device)
    : features_(features), implicit_(implicit), device_(device),
      lsh_((features > 0 && sampleRate > 0.0 && sampleRate <= 1.0)
               ? LocalitySensitiveHash(sampleRate, features, default_cores(numCores))
               : throw std::invalid_argument("features must be > 0 and sampleRate in (0,1]")) {  // :98-99
  oryx_config cfg{};
  cfg.features = features;
  cfg.implicit = implicit ? 1 : 0;
  cfg.num_hashes = lsh_.getNumHashes();
  cfg.max_bits_differing = lsh_.getMaxBitsDiffering();
  cfg.hash_vectors = lsh_.hashVectors().empty() ? nullptr : lsh_.hashVectors().data();
  cfg.device = device;
  check_status(oryx_model_create(&cfg, &h_));
}

ALSServingModel::~ALSServingModel() { oryx_model_destroy(h_); }

int ALSServingModel::rowOf(const std::string& item) const {
  auto it = item_row_.find(item);
  return (it == item_row_.end() || !item_has_vec_[(size_t)it->second]) ? -1 : it->second;
}

int ALSServingModel::itemRowFor(const std::string& item) {
  auto it = item_row_.find(item);
  if (it != item_row_.end()) return it->second;
  const int row = (int)row_item_.size();
  row_item_.push_back(item);
  item_has_vec_.push_back(0);
  item_recent_.push_back(0);
  item_part_.push_back(-1);
  item_row_.emplace(item, row);
  return row;
}

bool ALSServingModel::itemIsRecent(const std::string& item) const {
  auto it = item_row_.find(item);
  return it != item_row_.end() && item_recent_[(size_t)it->second];
}

int ALSServingModel::userRowFor(const std::string& user) {
  auto it = user_row_.find(user);
  if (it != user_row_.end()) return it->second;
  const int row = (int)row_user_.size();
  row_user_.push_back(user);
  user_row_.emplace(user, row);
  return row;
}

std::unique_ptr<FloatVec> ALSServingModel::getUserVector(const std::string& user) const {
  std::lock_guard<std::mutex> lk(mu_);
  auto it = X_.find(user);
  return it == X_.end() ? nullptr : std::make_unique<FloatVec>(it->second);
}

std::unique_ptr<FloatVec> ALSServingModel::getItemVector(const std::string& item) const {
  int row;
  {
    std::lock_guard<std::mutex> lk(mu_);
    row = rowOf(item);
  }
  if (row < 0) return nullptr;
  FloatVec v((size_t)features_);
  int32_t found = 0;
  check_status(oryx_items_get(h_, ORYX_Y, row, v.data(), &found));
  return found ? std::make_unique<FloatVec>(std::move(v)) : nullptr;
}

void ALSServingModel::setUserVector(const std::string& user, const FloatVec& vector) {
  if ((int)vector.size() != features_) throw std::invalid_argument("vector.length != features");  // :154
  setUserVectors(&user, vector.data(), 1);
}

void ALSServingModel::setUserVectors(const std::string* users, const float* vecs, size_t n) {
  if (n == 0) return;
  const size_t f = (size_t)features_;
  std::vector<int32_t> rows(n);
  {
    std::lock_guard<std::mutex> lk(mu_);
    for (size_t i = 0; i < n; i++) {
      rows[i] = userRowFor(users[i]);
      auto it = X_.find(users[i]);
      if (it == X_.end()) {
        X_.emplace(users[i], FloatVec(vecs + i * f, vecs + (i + 1) * f));
        recent_users_.insert(users[i]);  // FeatureVectors.java:74-77: ID was actually new
      } else {
        it->second.assign(vecs + i * f, vecs + (i + 1) * f);
      }
      if (!expected_users_.empty()) expected_users_.erase(users[i]);
    }
  }
  // the device copy of X serves oryx_recommend (query vector read from HBM)
  check_status(oryx_items_upsert(h_, ORYX_X, rows.data(), vecs, (int32_t)n, nullptr));
}

void ALSServingModel::setItemVector(const std::string& item, const FloatVec& vector) {
  setItemVectors({item}, vector);
}

void ALSServingModel::setItemVectors(const std::vector<std::string>& items, const std::vector<float>& vecs) {
  if (vecs.size() != items.size() * (size_t)features_) throw std::invalid_argument("vector.length != features");  // :162
  setItemVectors(items.data(), vecs.data(), items.size());
}

void ALSServingModel::setItemVectors(const std::string* items, const float* vecs, size_t n) {
  if (n == 0) return;
  std::vector<int32_t> rows(n), parts(n);
  {
    std::lock_guard<std::mutex> lk(mu_);
    for (size_t i = 0; i < n; i++) {
      const int row = itemRowFor(items[i]);
      if (!item_has_vec_[(size_t)row]) {
        item_has_vec_[(size_t)row] = 1;
        n_item_vecs_++;
        item_recent_[(size_t)row] = 1;  // FeatureVectors.java:74-77: ID was actually new
      }
      rows[i] = row;
    }
  }
  // hashing (getIndexFor), partition move and storage happen on the device (:163-178)
  check_status(oryx_items_upsert(h_, ORYX_Y, rows.data(), vecs, (int32_t)n, parts.data()));
  std::lock_guard<std::mutex> lk(mu_);
  for (size_t i = 0; i < n; i++) {
    item_part_[(size_t)rows[i]] = parts[i];
    if (!expected_items_.empty()) expected_items_.erase(items[i]);
  }
  cached_yty_.reset();  // :182 cachedYTYSolver.set(null)
}

std::unordered_set<std::string> ALSServingModel::getKnownItems(const std::string& user) const {
  std::lock_guard<std::mutex> lk(mu_);
  auto it = known_items_.find(user);
  return it == known_items_.end() ? std::unordered_set<std::string>{} : it->second;
}

void ALSServingModel::addKnownItems(const std::string& user, const std::vector<std::string>& items) {
  const int64_t begin[2] = {0, (int64_t)items.size()};
  addKnownItems(&user, begin, items.data(), 1);
}

void ALSServingModel::addKnownItems(const std::string* users, const int64_t* begin, const std::string* items,
                                    size_t nUsers) {
  if (nUsers == 0) return;
  std::vector<int32_t> userRows(nUsers), itemRows((size_t)(begin[nUsers] - begin[0]));
  std::vector<int64_t> rel(nUsers + 1);
  {
    std::lock_guard<std::mutex> lk(mu_);
    for (size_t i = 0; i < nUsers; i++) {
      userRows[i] = userRowFor(users[i]);
      rel[i] = begin[i] - begin[0];
      auto& s = known_items_[users[i]];
      for (int64_t j = begin[i]; j < begin[i + 1]; j++) {
        s.insert(items[j]);
        itemRows[(size_t)(j - begin[0])] = itemRowFor(items[j]);
      }
    }
    rel[nUsers] = begin[nUsers] - begin[0];
  }
  static const int32_t none = 0;
  check_status(oryx_known_items_add(h_, userRows.data(), rel.data(), itemRows.empty() ? &none : itemRows.data(),
                                    (int32_t)nUsers));
}

std::unordered_map<std::string, int> ALSServingModel::getUserCounts() const {
  std::vector<int32_t> counts;
  {
    std::lock_guard<std::mutex> lk(mu_);
    counts.resize(row_user_.size());
  }
  if (!counts.empty()) check_status(oryx_known_counts(h_, ORYX_X, counts.data(), (int32_t)counts.size()));
  std::unordered_map<std::string, int> out;
  std::lock_guard<std::mutex> lk(mu_);
  // the reference's map has an entry per user with a known-items set, empty sets included
  for (size_t r = 0; r < counts.size(); r++)
    if (counts[r] > 0 || known_items_.count(row_user_[r])) out[row_user_[r]] = counts[r];
  return out;
}

std::unordered_map<std::string, int> ALSServingModel::getItemCounts() const {
  std::vector<int32_t> counts;
  {
    std::lock_guard<std::mutex> lk(mu_);
    counts.resize(row_item_.size());
  }
  if (!counts.empty()) check_status(oryx_known_counts(h_, ORYX_Y, counts.data(), (int32_t)counts.size()));
  std::unordered_map<std::string, int> out;
  std::lock_guard<std::mutex> lk(mu_);
  for (size_t r = 0; r < counts.size(); r++)
    if (counts[r] > 0) out[row_item_[r]] = counts[r];
  return out;
}

std::unique_ptr<std::vector<IDValue>> ALSServingModel::recommend(const std::string& user, int howMany,
                                                                 bool considerKnownItems) const {
  int row;
  {
    std::lock_guard<std::mutex> lk(mu_);
    auto it = user_row_.find(user);
    row = it == user_row_.end() ? -1 : it->second;
  }
  if (row < 0) return nullptr;
  std::vector<int32_t> rows((size_t)howMany);
  std::vector<double> scores((size_t)howMany);
  int32_t n = 0, found = 0;
  check_status(oryx_recommend(h_, row, howMany, considerKnownItems ? 1 : 0, rows.data(), scores.data(), &n, &found));
  if (!found) return nullptr;
  auto out = std::make_unique<std::vector<IDValue>>();
  std::lock_guard<std::mutex> lk(mu_);
  for (int i = 0; i < n; i++) out->emplace_back(row_item_[(size_t)rows[i]], scores[i]);
  return out;
}

std::unique_ptr<std::vector<std::pair<std::string, FloatVec>>> ALSServingModel::getKnownItemVectorsForUser(
    const std::string& user) const {
  if (!getUserVector(user)) return nullptr;
  auto known = getKnownItems(user);
  if (known.empty()) return nullptr;
  auto out = std::make_unique<std::vector<std::pair<std::string, FloatVec>>>();
  for (const auto& id : known) {
    auto v = getItemVector(id);
    if (v) out->emplace_back(id, *v);
  }
  if (out->empty()) return nullptr;
  return out;
}

std::vector<IDValue> ALSServingModel::topN(const ScoreFunction& scoreFn, const Rescorer* rescorer, int howMany,
                                           const std::unordered_set<std::string>* excludedIDs) const {
  if ((int)scoreFn.target.size() != features_) throw std::invalid_argument("target vector length != features");
  std::vector<int32_t> excl;
  int n_rows;
  {
    std::lock_guard<std::mutex> lk(mu_);
    n_rows = (int)row_item_.size();
    if (excludedIDs) {
      for (const auto& id : *excludedIDs) {
        int r = rowOf(id);
        if (r >= 0) excl.push_back(r);
      }
    }
  }
  std::vector<IDValue> out;
  if (rescorer == nullptr) {
    // allowedPredicate = "not a known item": the device applies it (TopNConsumer.java:57)
    std::vector<int32_t> rows((size_t)howMany);
    std::vector<double> scores((size_t)howMany);
    int32_t n = 0;
    // many request threads call this at once (ServingLayer.java:225): the combining entry point answers
    // whoever is waiting with one batch pass, with oryx_topn's exact results
    check_status(oryx_topn_coalesced(h_, scoreFn.kind, scoreFn.target.data(), howMany,
                                     excl.empty() ? nullptr : excl.data(), (int32_t)excl.size(), rows.data(),
                                     scores.data(), &n));
    std::lock_guard<std::mutex> lk(mu_);
    for (int i = 0; i < n; i++) out.emplace_back(row_item_[rows[i]], scores[i]);
    return out;
  }
  // Arbitrary Rescorer callbacks: exact scores of every row from the device, then the reference's
  // TopNConsumer logic on the host (TopNConsumer.java:56-74; ALSServingModel.java:334).
  std::vector<double> all((size_t)std::max(n_rows, 1));
  if (n_rows > 0) check_status(oryx_scores_all(h_, scoreFn.kind, scoreFn.target.data(), n_rows, all.data()));
  std::unordered_set<int32_t> excl_set(excl.begin(), excl.end());
  std::vector<std::pair<double, int>> heap;  // min-heap on score
  auto cmp = [](const std::pair<double, int>& a, const std::pair<double, int>& b) { return a.first > b.first; };
  double lower = -std::numeric_limits<double>::infinity();
  {
    std::lock_guard<std::mutex> lk(mu_);
    for (int r = 0; r < n_rows; r++) {
      if (std::isnan(all[r]) || !item_has_vec_[(size_t)r] || excl_set.count(r)) continue;  // absent rows score NaN
      const std::string& id = row_item_[r];
      if (rescorer->isFiltered && rescorer->isFiltered(id)) continue;
      double s = rescorer->rescore ? rescorer->rescore(id, all[r]) : all[r];
      if (s > lower) {
        if ((int)heap.size() >= howMany) {
          lower = heap.front().first;
          if (s > lower) {
            std::pop_heap(heap.begin(), heap.end(), cmp);
            heap.back() = {s, r};
            std::push_heap(heap.begin(), heap.end(), cmp);
          }
        } else {
          heap.emplace_back(s, r);
          std::push_heap(heap.begin(), heap.end(), cmp);
        }
      }
    }
    std::stable_sort(heap.begin(), heap.end(), [](const auto& a, const auto& b) { return a.first > b.first; });
    for (const auto& e : heap) out.emplace_back(row_item_[e.second], e.first);
  }
  return out;
}

std::shared_ptr<Solver> ALSServingModel::getYTYSolver() {
  {
    std::lock_guard<std::mutex> lk(mu_);
    if (cached_yty_) return cached_yty_;
  }
  std::vector<double> g((size_t)features_ * features_);
  int64_t used = 0;
  check_status(oryx_gram(h_, ORYX_Y, g.data(), &used));
  if (used == 0) return nullptr;  // LinearSystemSolver.getSolver(null) == null
  auto s = std::make_shared<Solver>(g, features_, device_);
  std::lock_guard<std::mutex> lk(mu_);
  cached_yty_ = s;
  return s;
}

void ALSServingModel::retainRecentAndUserIDs(const std::vector<std::string>& users) {
  std::unordered_set<std::string> keep(users.begin(), users.end());
  std::vector<int32_t> drop;
  {
    std::lock_guard<std::mutex> lk(mu_);
    for (auto it = X_.begin(); it != X_.end();) {
      if (keep.count(it->first) || recent_users_.count(it->first)) { ++it; continue; }
      drop.push_back(user_row_.at(it->first));
      it = X_.erase(it);
    }
    recent_users_.clear();
    expected_users_.clear();
    for (const auto& u : users) if (!X_.count(u)) expected_users_.insert(u);
  }
  if (!drop.empty()) check_status(oryx_items_remove(h_, ORYX_X, drop.data(), (int32_t)drop.size()));
}

void ALSServingModel::retainRecentAndItemIDs(const std::vector<std::string>& items) {
  std::unordered_set<std::string> keep(items.begin(), items.end());
  std::vector<int32_t> drop;
  {
    std::lock_guard<std::mutex> lk(mu_);
    for (size_t r = 0; r < row_item_.size(); r++) {
      if (!item_has_vec_[r]) continue;
      const std::string& id = row_item_[r];
      if (item_recent_[r] || keep.count(id)) continue;
      drop.push_back((int32_t)r);
      item_has_vec_[r] = 0;
      n_item_vecs_--;
      item_part_[r] = -1;
    }
    std::fill(item_recent_.begin(), item_recent_.end(), 0);
    expected_items_.clear();
    for (const auto& i : items) if (rowOf(i) < 0) expected_items_.insert(i);
    cached_yty_.reset();
  }
  if (!drop.empty()) check_status(oryx_items_remove(h_, ORYX_Y, drop.data(), (int32_t)drop.size()));
}

void ALSServingModel::retainRecentAndKnownItems(const std::vector<std::string>& users,
                                                const std::vector<std::string>& items) {
  std::unordered_set<std::string> keepUsers(users.begin(), users.end()), keepItems(items.begin(), items.end());
  std::vector<uint8_t> userMask, itemMask;
  {
    std::lock_guard<std::mutex> lk(mu_);
    for (auto it = known_items_.begin(); it != known_items_.end();)
      it = (keepUsers.count(it->first) || recent_users_.count(it->first)) ? std::next(it) : known_items_.erase(it);
    for (auto& kv : known_items_) {
      auto& set = kv.second;
      for (auto it = set.begin(); it != set.end();)
        it = (keepItems.count(*it) || itemIsRecent(*it)) ? std::next(it) : set.erase(it);
    }
    // the same predicate as row masks for the device copy
    userMask.resize(row_user_.size());
    for (size_t r = 0; r < row_user_.size(); r++)
      userMask[r] = keepUsers.count(row_user_[r]) || recent_users_.count(row_user_[r]);
    itemMask.resize(row_item_.size());
    for (size_t r = 0; r < row_item_.size(); r++)
      itemMask[r] = item_recent_[r] || keepItems.count(row_item_[r]);
  }
  static const uint8_t none = 0;
  check_status(oryx_known_items_retain(h_, userMask.empty() ? &none : userMask.data(), (int32_t)userMask.size(),
                                       itemMask.empty() ? &none : itemMask.data(), (int32_t)itemMask.size()));
}

std::vector<std::string> ALSServingModel::getAllItemIDs() const {
  std::lock_guard<std::mutex> lk(mu_);
  std::vector<std::string> out;
  out.reserve((size_t)n_item_vecs_);
  for (size_t r = 0; r < row_item_.size(); r++) if (item_has_vec_[r]) out.push_back(row_item_[r]);
  return out;
}

int ALSServingModel::getNumUsers() const { std::lock_guard<std::mutex> lk(mu_); return (int)X_.size(); }
int ALSServingModel::getNumItems() const { std::lock_guard<std::mutex> lk(mu_); return n_item_vecs_; }

float ALSServingModel::getFractionLoaded() const {
  std::lock_guard<std::mutex> lk(mu_);
  int expected = (int)expected_users_.size() + (int)expected_items_.size();
  if (expected == 0) return 1.0f;
  float loaded = (float)X_.size() + (float)n_item_vecs_;
  return loaded / (loaded + expected);
}

int ALSServingModel::itemRow(const std::string& item) const {
  std::lock_guard<std::mutex> lk(mu_);
  return rowOf(item);
}

int ALSServingModel::partitionOf(const std::string& item) const {
  std::lock_guard<std::mutex> lk(mu_);
  auto it = item_row_.find(item);
  return it == item_row_.end() ? -1 : item_part_[(size_t)it->second];
}

// ---------------------------------------------------------------- ALSSpeedModel
ALSSpeedModel::ALSSpeedModel(int features, bool implicit, int device)
    : features_(features), implicit_(implicit), device_(device) {
  if (features <= 0) throw std::invalid_argument("features must be > 0");  // ALSSpeedModel.java:57
  oryx_config cfg{};
  cfg.features = features;
  cfg.implicit = implicit ? 1 : 0;
  cfg.device = device;
  check_status(oryx_model_create(&cfg, &h_));
}
ALSSpeedModel::~ALSSpeedModel() { oryx_model_destroy(h_); }

void ALSSpeedModel::set(Side& side, int which, const std::string& id, const FloatVec& v) {
  if ((int)v.size() != features_) throw std::invalid_argument("vector.length != features");
  (void)side;
  setVectors(which == ORYX_X ? 'X' : 'Y', &id, v.data(), 1);
}

void ALSSpeedModel::setVectors(char whichChar, const std::string* ids, const float* vecs, size_t n) {
  if (n == 0) return;
  Side& side = whichChar == 'X' ? x_ : y_;
  const int which = whichChar == 'X' ? ORYX_X : ORYX_Y;
  const size_t f = (size_t)features_;
  std::vector<int32_t> rows(n);
  for (size_t i = 0; i < n; i++) {
    auto it = side.row.find(ids[i]);
    int32_t row;
    if (it == side.row.end()) {
      if (!side.free_rows.empty()) { row = side.free_rows.back(); side.free_rows.pop_back(); }
      else row = side.next_row++;
      side.row.emplace(ids[i], row);
      side.recent.insert(ids[i]);  // FeatureVectors.java:74-77
    } else {
      row = it->second;
    }
    rows[i] = row;
    side.vec[ids[i]].assign(vecs + i * f, vecs + (i + 1) * f);
    if (!side.expected.empty()) side.expected.erase(ids[i]);
  }
  check_status(oryx_items_upsert(h_, which, rows.data(), vecs, (int32_t)n, nullptr));
}

void ALSSpeedModel::retain(Side& side, int which, const std::vector<std::string>& ids) {
  std::unordered_set<std::string> keep(ids.begin(), ids.end());
  std::vector<int32_t> drop;
  for (auto it = side.row.begin(); it != side.row.end();) {
    if (keep.count(it->first) || side.recent.count(it->first)) { ++it; continue; }
    drop.push_back(it->second);
    side.free_rows.push_back(it->second);
    side.vec.erase(it->first);
    it = side.row.erase(it);
  }
  side.recent.clear();
  side.expected.clear();
  for (const auto& id : ids) if (!side.row.count(id)) side.expected.insert(id);
  if (!drop.empty()) check_status(oryx_items_remove(h_, which, drop.data(), (int32_t)drop.size()));
}
void ALSSpeedModel::retainRecentAndUserIDs(const std::vector<std::string>& users) { retain(x_, ORYX_X, users); }
void ALSSpeedModel::retainRecentAndItemIDs(const std::vector<std::string>& items) { retain(y_, ORYX_Y, items); }

float ALSSpeedModel::getFractionLoaded() const {
  int expected = (int)x_.expected.size() + (int)y_.expected.size();
  if (expected == 0) return 1.0f;
  float loaded = (float)x_.row.size() + (float)y_.row.size();
  return loaded / (loaded + expected);
}
void ALSSpeedModel::setUserVector(const std::string& user, const FloatVec& v) { set(x_, ORYX_X, user, v); }
void ALSSpeedModel::setItemVector(const std::string& item, const FloatVec& v) { set(y_, ORYX_Y, item, v); }

std::unique_ptr<FloatVec> ALSSpeedModel::getUserVector(const std::string& user) const {
  auto it = x_.vec.find(user);
  return it == x_.vec.end() ? nullptr : std::make_unique<FloatVec>(it->second);
}
std::unique_ptr<FloatVec> ALSSpeedModel::getItemVector(const std::string& item) const {
  auto it = y_.vec.find(item);
  return it == y_.vec.end() ? nullptr : std::make_unique<FloatVec>(it->second);
}

std::shared_ptr<Solver> ALSSpeedModel::solver(int which) {
  std::vector<double> g((size_t)features_ * features_);
  int64_t used = 0;
  check_status(oryx_gram(h_, which, g.data(), &used));  // uncached, as ALSSpeedModel.java:118-126
  if (used == 0) return nullptr;
  return std::make_shared<Solver>(g, features_, device_);
}
std::shared_ptr<Solver> ALSSpeedModel::getXTXSolver() { return solver(ORYX_X); }
std::shared_ptr<Solver> ALSSpeedModel::getYTYSolver() { return solver(ORYX_Y); }

std::vector<Update> ALSSpeedModel::buildUpdates(const std::vector<UserItemStrength>& input) {
  std::vector<Update> out;
  if (input.empty()) return out;
  auto xtx = getXTXSolver();  // ALSSpeedModelManager.java:173-180
  auto yty = getYTYSolver();
  if (!xtx || !yty) return out;
  const size_t n = input.size(), f = (size_t)features_;
  // Every interaction reads the same model snapshot (:186-187); the vectors never leave the device:
  //   newXu = computeUpdatedXu(YTYsolver, value, Xu, Yi)   :192
  //   newYi = computeUpdatedXu(XTXsolver, value, Yi, Xu)   :194
  std::vector<int32_t> xr(n), yr(n);
  std::vector<uint8_t> has_x(n, 0), has_y(n, 0);
  std::vector<double> values(n);
  for (size_t k = 0; k < n; k++) {
    values[k] = input[k].strength;
    auto ix = x_.row.find(input[k].user);
    auto iy = y_.row.find(input[k].item);
    xr[k] = ix == x_.row.end() ? -1 : ix->second;
    yr[k] = iy == y_.row.end() ? -1 : iy->second;
    has_x[k] = xr[k] >= 0;
    has_y[k] = yr[k] >= 0;
  }
  std::vector<float> new_x(n * f), new_y(n * f);
  std::vector<uint8_t> ok_x(n), ok_y(n);
  check_status(oryx_fold_in_rows(yty->handle(), h_, implicit_ ? 1 : 0, ORYX_X, xr.data(), yr.data(), values.data(),
                                 (int64_t)n, new_x.data(), ok_x.data()));
  check_status(oryx_fold_in_rows(xtx->handle(), h_, implicit_ ? 1 : 0, ORYX_Y, yr.data(), xr.data(), values.data(),
                                 (int64_t)n, new_y.data(), ok_y.data()));
  for (size_t k = 0; k < n; k++) {
    // computeUpdatedXu returns null when its "Yi" argument is null (ALSUtils.java:79-81)
    if (ok_x[k] && has_y[k]) out.push_back(Update{'X', input[k].user, FloatVec(new_x.begin() + k * f, new_x.begin() + (k + 1) * f), input[k].item});
    if (ok_y[k] && has_x[k]) out.push_back(Update{'Y', input[k].item, FloatVec(new_y.begin() + k * f, new_y.begin() + (k + 1) * f), input[k].user});
  }
  return out;
}

// ---------------------------------------------------------------- Evaluation
namespace Evaluation {

namespace {
// predict() for pairs; result[i] is meaningful where found[i]
void predict(oryx_model* model, const std::vector<std::pair<int, int>>& pairs, std::vector<double>* scores,
             std::vector<uint8_t>* found) {
  std::vector<int32_t> users(pairs.size()), items(pairs.size());
  for (size_t i = 0; i < pairs.size(); i++) { users[i] = pairs[i].first; items[i] = pairs[i].second; }
  scores->assign(pairs.size(), 0.0);
  found->assign(pairs.size(), 0);
  if (pairs.empty()) return;
  check_status(oryx_estimate_pairs(model, users.data(), items.data(), (int64_t)pairs.size(), /*wide_products=*/1,
                                   scores->data(), found->data()));
}
}  // namespace

double rmse(oryx_model* model, const std::vector<Rating>& testData) {
  std::vector<std::pair<int, int>> pairs;
  pairs.reserve(testData.size());
  for (const auto& r : testData) pairs.emplace_back(r.user, r.product);
  std::vector<double> pred;
  std::vector<uint8_t> found;
  predict(model, pairs, &pred, &found);
  double sum = 0.0;
  size_t n = 0;
  for (size_t i = 0; i < testData.size(); i++) {
    if (!found[i]) continue;  // no prediction: the join drops the pair (:57-62)
    const double diff = pred[i] - testData[i].rating;
    sum += diff * diff;
    n++;
  }
  return n == 0 ? std::numeric_limits<double>::quiet_NaN() : std::sqrt(sum / (double)n);
}

std::vector<std::pair<int, int>> sampleNegatives(const std::vector<Rating>& positiveData, uint64_t seed) {
  std::vector<int> allItemIDs;  // distinct items, first-appearance order (:93)
  {
    std::unordered_set<int> seen;
    for (const auto& r : positiveData) if (seen.insert(r.product).second) allItemIDs.push_back(r.product);
  }
  std::unordered_map<int, std::unordered_set<int>> positives;
  std::vector<int> userOrder;
  for (const auto& r : positiveData) {
    auto ins = positives.emplace(r.user, std::unordered_set<int>{});
    if (ins.second) userOrder.push_back(r.user);
    ins.first->second.insert(r.product);
  }
  std::mt19937_64 rng(seed);
  const int numItems = (int)allItemIDs.size();
  std::vector<std::pair<int, int>> negative;
  for (int user : userOrder) {
    const auto& pos = positives[user];
    const size_t numPositive = pos.size();
    size_t added = 0;
    // "Sample about as many negative examples as positive" (:108-114)
    for (int i = 0; i < numItems && added < numPositive; i++) {
      const int item = allItemIDs[(size_t)(rng() % (uint64_t)numItems)];
      if (!pos.count(item)) {
        negative.emplace_back(user, item);
        added++;
      }
    }
  }
  return negative;
}

double areaUnderCurve(oryx_model* model, const std::vector<Rating>& positiveData,
                      const std::vector<std::pair<int, int>>& negativeUserProducts) {
  std::vector<std::pair<int, int>> posPairs;
  posPairs.reserve(positiveData.size());
  for (const auto& r : positiveData) posPairs.emplace_back(r.user, r.product);
  std::vector<double> ps, ns;
  std::vector<uint8_t> pf, nf;
  predict(model, posPairs, &ps, &pf);
  predict(model, negativeUserProducts, &ns, &nf);
  std::unordered_map<int, std::pair<std::vector<double>, std::vector<double>>> byUser;
  for (size_t i = 0; i < posPairs.size(); i++) if (pf[i]) byUser[posPairs[i].first].first.push_back(ps[i]);
  for (size_t i = 0; i < negativeUserProducts.size(); i++) if (nf[i]) byUser[negativeUserProducts[i].first].second.push_back(ns[i]);
  double sum = 0.0;
  size_t users = 0;
  for (const auto& kv : byUser) {
    const auto& pos = kv.second.first;
    const auto& neg = kv.second.second;
    if (pos.empty() || neg.empty()) continue;  // the join keeps users with both (:122)
    long correct = 0, total = 0;
    for (double p : pos) for (double q : neg) { if (p > q) correct++; total++; }
    sum += total == 0 ? 0.0 : (double)correct / (double)total;
    users++;
  }
  return users == 0 ? std::numeric_limits<double>::quiet_NaN() : sum / (double)users;
}

}  // namespace Evaluation

}  // namespace host
}  // namespace oryx
