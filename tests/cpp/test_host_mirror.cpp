// test_host_mirror.cpp -- the reference's serving/speed unit tests, replayed in C++ against the
// host-side mirror (oryx_b200/host) which calls the CUDA library through its C ABI.  Each block names
// the reference test it transcribes (paths relative to the reference root,
// T = app/oryx-app-serving/src/test/java/com/cloudera/oryx/app/serving/als).
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <random>
#include <set>
#include <string>

#include "oryx_host.hpp"
#include "update_stream.hpp"

using namespace oryx::host;

static int g_checks = 0, g_fail = 0;
#define CHECK(cond)                                                                  \
  do {                                                                               \
    g_checks++;                                                                      \
    if (!(cond)) { g_fail++; std::printf("FAIL %s:%d: %s\n", __FILE__, __LINE__, #cond); } \
  } while (0)
#define CHECK_NEAR(a, b, eps) CHECK(std::fabs((double)(a) - (double)(b)) <= (eps))

static const double FLOAT_EPSILON = 1.0e-6, DOUBLE_EPSILON = 1.0e-12;  // OryxTest.java:41-42

// T/model/TestALSModelFactory.java:67-99
static void buildTestModel(ALSServingModel& model) {
  const float X[7][2] = {{-0.35837504f, 0.60391283f}, {-0.7757129f, 0.4327127f}, {-0.7757129f, -0.4327127f},
                         {-0.35837504f, -0.60391283f}, {-1.1340879f, 1.0366255f}, {-1.5514258f, -3.5398357e-16f},
                         {-1.1340879f, -1.0366255f}};
  const float Y[9][2] = {{-0.23176478f, 0.504302f}, {-0.53749436f, 0.45167503f}, {-0.53749436f, -0.45167503f},
                         {-0.23176478f, -0.504302f}, {-0.7692591f, 0.9559771f}, {-1.0749887f, 1.0619507e-16f},
                         {-0.7692591f, -0.9559771f}, {-1.3067534f, 0.504302f}, {-1.3067534f, -0.504302f}};
  const int A[7][9] = {{1, 0, 0, 0, 1, 0, 0, 1, 0}, {0, 1, 0, 0, 1, 1, 0, 1, 1}, {0, 0, 1, 0, 0, 1, 1, 1, 1},
                       {0, 0, 0, 1, 0, 0, 1, 0, 1}, {1, 1, 0, 0, 2, 1, 0, 2, 1}, {0, 1, 1, 0, 1, 2, 1, 2, 2},
                       {0, 0, 1, 1, 0, 1, 2, 1, 2}};
  for (int u = 0; u < 7; u++) model.setUserVector("U" + std::to_string(u), {X[u][0], X[u][1]});
  for (int i = 0; i < 9; i++) model.setItemVector("I" + std::to_string(i), {Y[i][0], Y[i][1]});
  for (int u = 0; u < 7; u++) {
    std::vector<std::string> known;
    for (int i = 0; i < 9; i++) if (A[u][i] > 0) known.push_back("I" + std::to_string(i));
    model.addKnownItems("U" + std::to_string(u), known);
  }
}

// T/TestALSRescorerProvider.java:25-35
static Rescorer testRescorer() {
  Rescorer r;
  r.isFiltered = [](const std::string& id) { return id.back() % 2 == 0; };
  r.rescore = [](const std::string& id, double s) { return (id.back() % 2 == 0) ? std::nan("") : s * 2.0; };
  return r;
}

// Recommend.java:74-114
static std::vector<IDValue> recommend(ALSServingModel& m, const std::string& user, int howMany = 10, int offset = 0,
                                      bool considerKnownItems = false, const Rescorer* rescorer = nullptr) {
  auto uv = m.getUserVector(user);
  if (!uv) throw std::out_of_range(user);  // checkExists -> 404
  std::unordered_set<std::string> known;
  if (!considerKnownItems) known = m.getKnownItems(user);
  auto top = m.topN(DotsFunction(*uv), rescorer, howMany + offset, known.empty() ? nullptr : &known);
  if ((int)top.size() <= offset) return {};
  return std::vector<IDValue>(top.begin() + offset, top.end());  // AbstractALSResource.toIDValueResponse :37-43
}

// EstimateForAnonymous.buildTemporaryUserVector, EstimateForAnonymous.java:73-87
static std::unique_ptr<FloatVec> buildTemporaryUserVector(ALSServingModel& m,
                                                          const std::vector<std::pair<std::string, double>>& ctx,
                                                          std::unique_ptr<FloatVec> Xu) {
  auto solver = m.getYTYSolver();
  for (const auto& iv : ctx) {
    auto Yi = m.getItemVector(iv.first);
    auto newXu = ALSUtils::computeUpdatedXu(*solver, iv.second, Xu.get(), Yi.get(), m.isImplicit());
    if (newXu) Xu = std::move(newXu);
  }
  return Xu;
}

static double dotExact(ALSServingModel& m, const FloatVec& q, const std::string& item) {
  // Estimate.java:62-71 point lookup through the C ABI
  auto v = m.getItemVector(item);
  if (!v) return 0.0;
  double d = 0.0;
  for (size_t i = 0; i < q.size(); i++) { volatile float p = q[i] * (*v)[i]; d += (double)p; }
  return d;
}

static void testServingGoldens() {
  ALSServingModel model(2, true, 1.0, /*numCores=*/8);
  buildTestModel(model);
  CHECK(model.getNumUsers() == 7 && model.getNumItems() == 9);
  CHECK(model.getFractionLoaded() == 1.0f);

  {  // T/RecommendTest.java:31-37
    auto recs = recommend(model, "U0");
    CHECK(recs.size() == 6);
    for (size_t i = 0; i + 1 < recs.size(); i++) CHECK(recs[i].second >= recs[i + 1].second);
    CHECK(recs[0].first == "I1");
    CHECK_NEAR(0.4653969f, recs[0].second, FLOAT_EPSILON);
  }
  {  // T/RecommendTest.java:46-50 testHowMany, :58-63 testOffset
    CHECK(recommend(model, "U5", 10).size() == 2);
    CHECK(recommend(model, "U5", 2).size() == 2);
    CHECK(recommend(model, "U5", 1).size() == 1);
    CHECK(recommend(model, "U6", 2, 1).size() == 2);
    CHECK(recommend(model, "U6", 3, 1).size() == 2);
    CHECK(recommend(model, "U6", 1, 1).size() == 1);
    CHECK(recommend(model, "U6", 3, 3).size() == 0);
  }
  {  // T/RecommendTest.java:71-84 testConsiderKnownItems
    auto normal = recommend(model, "U4");
    CHECK(normal.size() == 3 && normal[0].first == "I2");
    CHECK_NEAR(0.14134796f, normal[0].second, FLOAT_EPSILON);
    auto with = recommend(model, "U4", 10, 0, true);
    CHECK(with.size() == 9 && with[0].first == "I7");
    CHECK_NEAR(2.0047457f, with[0].second, FLOAT_EPSILON);
  }
  {  // T/RecommendTest.java:92-98 testRescorer
    Rescorer r = testRescorer();
    auto recs = recommend(model, "U4", 10, 0, false, &r);
    CHECK(recs.size() == 1 && recs[0].first == "I3");
    CHECK_NEAR(2.0f * -0.2599307f, recs[0].second, FLOAT_EPSILON);
  }
  {  // T/RecommendToManyTest.java:31-37 (DotsFunction(float[][]) mean vector)
    std::vector<FloatVec> vs = {*model.getUserVector("U0"), *model.getUserVector("U2")};
    std::unordered_set<std::string> known = model.getKnownItems("U0");
    for (const auto& k : model.getKnownItems("U2")) known.insert(k);
    auto recs = model.topN(DotsFunction(vs), nullptr, 10, &known);
    CHECK(recs.size() == 2 && recs[0].first == "I1");
    CHECK_NEAR(0.34344634f, recs[0].second, FLOAT_EPSILON);
  }
  {  // T/SimilarityTest.java:35-41
    std::vector<FloatVec> vs = {*model.getItemVector("I0"), *model.getItemVector("I4"), *model.getItemVector("I6")};
    std::unordered_set<std::string> known = {"I0", "I4", "I6"};
    auto recs = model.topN(CosineAverageFunction(vs), nullptr, 10, &known);
    CHECK(recs.size() == 6);
    CHECK(recs[1].first == "I1");
    CHECK_NEAR(0.5571406537227921, recs[2].second, DOUBLE_EPSILON);
  }
  {  // T/EstimateTest.java:29-34
    auto uv = model.getUserVector("U0");
    CHECK_NEAR(0.38761318f, dotExact(model, *uv, "I0"), FLOAT_EPSILON);
    CHECK_NEAR(0.4653969f, dotExact(model, *uv, "I1"), FLOAT_EPSILON);
    CHECK_NEAR(-0.0801478f, dotExact(model, *uv, "I2"), FLOAT_EPSILON);
    CHECK(dotExact(model, *uv, "I10") == 0.0);  // :50-53 unknown item
  }
  {  // T/EstimateForAnonymousTest.java:28-46
    auto xu = buildTemporaryUserVector(model, {{"I4", 1.0}, {"I5", 2.0}}, nullptr);
    CHECK(xu != nullptr);
    CHECK_NEAR(0.35964763164520264, dotExact(model, *xu, "I7"), DOUBLE_EPSILON);
    auto xu2 = buildTemporaryUserVector(model, {{"foo", 1.0}, {"I4", 1.0}, {"I5", 2.0}}, nullptr);
    CHECK_NEAR(-0.06707492843270302, dotExact(model, *xu2, "I3"), DOUBLE_EPSILON);
    auto xu3 = buildTemporaryUserVector(model, {{"foo", 1.0}}, nullptr);
    CHECK(xu3 == nullptr);  // -> estimate 0.0
  }
  {  // T/RecommendToAnonymousTest.java:32-36 and T/RecommendWithContextTest.java:109-113
    auto xu = buildTemporaryUserVector(model, {{"I4", 1.0}, {"I5", 2.0}}, nullptr);
    std::unordered_set<std::string> known = {"I4", "I5"};
    auto recs = model.topN(DotsFunction(*xu), nullptr, 10, &known);
    CHECK(recs.size() == 7 && recs[0].first == "I7");
    CHECK_NEAR(0.35964763f, recs[0].second, FLOAT_EPSILON);
    auto xu2 = buildTemporaryUserVector(model, {{"I4", 1.0}, {"I5", 2.0}}, model.getUserVector("U0"));
    for (const auto& k : model.getKnownItems("U0")) known.insert(k);
    auto recs2 = model.topN(DotsFunction(*xu2), nullptr, 10, &known);
    CHECK(recs2.size() == 5 && recs2[0].first == "I1");
    CHECK_NEAR(0.51607955f, recs2[0].second, FLOAT_EPSILON);
  }
  {  // the same /recommend goldens served from device-resident state (oryx_recommend): RecommendTest.java:31-37, 71-84
    auto r0 = model.recommend("U0", 10, false);
    CHECK(r0 && r0->size() == 6 && (*r0)[0].first == "I1");
    if (r0 && !r0->empty()) CHECK_NEAR(0.4653969f, (*r0)[0].second, FLOAT_EPSILON);
    auto r4 = model.recommend("U4", 10, false);
    CHECK(r4 && r4->size() == 3 && (*r4)[0].first == "I2");
    if (r4 && !r4->empty()) CHECK_NEAR(0.14134796f, (*r4)[0].second, FLOAT_EPSILON);
    auto r4k = model.recommend("U4", 10, true);
    CHECK(r4k && r4k->size() == 9 && (*r4k)[0].first == "I7");
    if (r4k && !r4k->empty()) CHECK_NEAR(2.0047457f, (*r4k)[0].second, FLOAT_EPSILON);
    CHECK(model.recommend("nobody", 10, false) == nullptr);
    // identical to the path that ships the exclusion list with the call
    auto viaTopN = recommend(model, "U0");
    CHECK(r0 && viaTopN.size() == r0->size());
    for (size_t i = 0; r0 && i < viaTopN.size(); i++) CHECK(viaTopN[i].first == (*r0)[i].first && viaTopN[i].second == (*r0)[i].second);
    // T/MostActiveUsersTest / MostPopularItemsTest inputs: counts of the known-items matrix (TestALSModelFactory.java:80-96)
    auto uc = model.getUserCounts();
    auto ic = model.getItemCounts();
    CHECK(uc["U0"] == 3 && uc["U1"] == 5 && uc["U4"] == 6 && uc["U5"] == 7);
    CHECK(ic["I0"] == 2 && ic["I4"] == 4 && ic["I7"] == 6 && ic["I3"] == 2);
  }
  {  // T/PopularRepresentativeItemsTest.java:30-36
    std::vector<std::string> items;
    FloatVec unit(2, 0.0f);
    for (int i = 0; i < 2; i++) {
      unit[i] = 1.0f;
      auto top = model.topN(DotsFunction(unit), nullptr, 1, nullptr);
      items.push_back(top.empty() ? "" : top[0].first);
      unit[i] = 0.0f;
    }
    CHECK((items[0] == "I0" || items[0] == "I3") && items[1] == "I4");
  }
  {  // model bookkeeping on the golden model: rows set since the last retain are "recent" and survive one call
    const std::vector<std::string> keep = {"I0", "I1", "I2", "I3", "I4", "I5", "I6", "I7", "I99"};
    model.retainRecentAndItemIDs(keep);
    CHECK(model.getNumItems() == 9 && model.getItemVector("I8") != nullptr);  // protected: recent
    model.retainRecentAndItemIDs(keep);
    CHECK(model.getNumItems() == 8);
    CHECK(model.getItemVector("I8") == nullptr);
    CHECK(model.getFractionLoaded() < 1.0f);  // I99 is expected but absent
    model.setItemVector("I99", {0.1f, 0.2f});
    CHECK(model.getFractionLoaded() == 1.0f);
    auto recs = recommend(model, "U0", 10, 0, true);
    for (const auto& r : recs) CHECK(r.first != "I8");
  }
  {  // T/model/ALSServingModelTest.java:75-103 testRetainUsersItems
    ALSServingModel m(2, true, 1.0);
    m.setUserVector("U0", {1.0f, 1.0f});
    m.retainRecentAndUserIDs({});
    CHECK(m.getUserVector("U0") != nullptr);  // protected because of recent users/items
    m.retainRecentAndUserIDs({});
    CHECK(m.getUserVector("U0") == nullptr);
    m.setUserVector("U0", {1.0f, 1.0f});
    m.retainRecentAndUserIDs({"U0"});
    CHECK(m.getUserVector("U0") != nullptr);
    m.retainRecentAndUserIDs({"U0"});
    CHECK(m.getUserVector("U0") != nullptr);
    m.setItemVector("I0", {1.0f, 1.0f});
    m.retainRecentAndItemIDs({});
    CHECK(m.getItemVector("I0") != nullptr);
    m.retainRecentAndItemIDs({});
    CHECK(m.getItemVector("I0") == nullptr);
    m.setItemVector("I0", {1.0f, 1.0f});
    m.retainRecentAndItemIDs({"I0"});
    CHECK(m.getItemVector("I0") != nullptr);
    m.retainRecentAndItemIDs({"I0"});
    CHECK(m.getItemVector("I0") != nullptr);
  }
  {  // T/model/ALSServingModelTest.java:105-132 testRetainKnown (populateKnownItems :142-152)
    ALSServingModel m(2, true, 1.0);
    for (int i = 0; i < 10; i++)
      for (int j = 0; j < 10; j++)
        if (std::abs(i - j) <= 1) m.addKnownItems("U" + std::to_string(i), {"I" + std::to_string(j)});
    for (int i = 0; i < 10; i++) {
      m.setUserVector("U" + std::to_string(i), {0.0f, 0.0f});
      m.setItemVector("I" + std::to_string(i), {0.0f, 0.0f});
    }
    m.retainRecentAndKnownItems({"U4", "U5", "U6"}, {"I4", "I5", "I6"});
    CHECK(m.getKnownItems("U3").count("I4") && m.getKnownItems("U4").count("I4"));
    CHECK(m.getKnownItems("U6").count("I6") && m.getKnownItems("U6").count("I7"));
    CHECK(m.getKnownItems("U2").count("I2"));  // protected because of recent users/items
    m.retainRecentAndUserIDs({});              // clears recent users/items
    m.retainRecentAndItemIDs({});
    m.retainRecentAndKnownItems({"U4", "U5", "U6"}, {"I4", "I5", "I6"});
    CHECK(m.getKnownItems("U3").empty());
    CHECK(m.getKnownItems("U4").count("I4") && m.getKnownItems("U6").count("I6"));
    CHECK(!m.getKnownItems("U6").count("I7") && m.getKnownItems("U2").empty());
  }
  {  // T/model/ALSServingModelTest.java:222-234 testFractionLoaded
    ALSServingModel m(2, true, 1.0);
    CHECK(m.getFractionLoaded() == 1.0f);
    m.retainRecentAndUserIDs({"U1"});
    m.retainRecentAndItemIDs({"I0"});
    CHECK(m.getFractionLoaded() == 0.0f);
    m.setUserVector("U1", {1.5f, -2.5f});
    CHECK(m.getFractionLoaded() == 0.5f);
    m.setItemVector("I0", {0.5f, 0.0f});
    CHECK(m.getFractionLoaded() == 1.0f);
  }
  {  // argument checking: ALSServingModel.java:98-99,154,162 -> IllegalArgumentException
    bool threw = false;
    try { ALSServingModel bad(0, true, 1.0); } catch (const std::invalid_argument&) { threw = true; }
    CHECK(threw);
    threw = false;
    try { model.setItemVector("I0", {1.0f}); } catch (const std::invalid_argument&) { threw = true; }
    CHECK(threw);
  }
}

// LocalitySensitiveHashTest.java:36-70,85-139 and ALSServingModelTest.testLSHEffect-style recall (:154-220)
static void testLSH() {
  struct { double rate; int cores, hashes, bits; } table[] = {
      {1.0, 1, 0, 0}, {0.5, 1, 1, 0}, {0.1, 1, 4, 0}, {1.0, 2, 1, 1}, {0.75, 3, 2, 1}, {0.5, 3, 3, 1},
      {0.1, 8, 7, 1}, {0.01, 8, 11, 1}, {0.001, 8, 14, 1}, {0.0001, 8, 16, 1}, {0.00001, 8, 16, 1}, {0.5, 32, 7, 3}};
  for (auto& t : table) {
    LocalitySensitiveHash lsh(t.rate, 10, t.cores);
    CHECK(lsh.getNumHashes() == t.hashes && lsh.getMaxBitsDiffering() == t.bits);
  }
  // recall: sample-rate 0.5 vs 1.0 on a 20k x 20f model, mean matching prefix of the top 10
  const int n = 20000, f = 20;
  ALSServingModel full(f, true, 1.0, 8), sampled(f, true, 0.5, 8);
  std::mt19937_64 rng(42);
  std::normal_distribution<float> g(0.f, 1.f);
  std::vector<std::string> ids(n);
  std::vector<float> vecs((size_t)n * f);
  for (int i = 0; i < n; i++) { ids[i] = "I" + std::to_string(i); for (int j = 0; j < f; j++) vecs[(size_t)i * f + j] = g(rng); }
  full.setItemVectors(ids, vecs);
  sampled.setItemVectors(ids, vecs);
  double prefix_sum = 0;
  const int trials = 50;
  for (int t = 0; t < trials; t++) {
    FloatVec q(f);
    for (auto& x : q) x = g(rng);
    auto a = full.topN(DotsFunction(q), nullptr, 10, nullptr);
    auto b = sampled.topN(DotsFunction(q), nullptr, 10, nullptr);
    size_t p = 0;
    while (p < a.size() && p < b.size() && a[p].first == b[p].first) p++;
    prefix_sum += (double)p;
    // everything the sampled model returns must carry the exact score the full model would give it
    for (const auto& r : b) for (const auto& s : a) if (r.first == s.first) CHECK(r.second == s.second);
  }
  std::printf("LSH recall: mean matching prefix %.2f of 10 at sample-rate 0.5\n", prefix_sum / trials);
  CHECK(prefix_sum / trials >= 4.0);  // ALSServingModelTest.java:154-220 expects >= 4.0 for dots
}

// SPEED-T/ALSSpeedIT.java:80-113 with MockALSModelUpdateGenerator.java:67-79 / MockALSInputGenerator.java:26-37
static void testSpeedFoldIn() {
  ALSSpeedModel model(2, true);
  const char* xid[4] = {"G6", "H7", "I8", "J9"};
  const float X[4][2] = {{-0.6790019f, 0.1732324f}, {-0.8232442f, -0.9200852f}, {-1.1865344f, 0.44631857f}, {-0.20789514f, 0.5303508f}};
  const char* yid[5] = {"B1", "C2", "D3", "E4", "F5"};
  const float Y[5][2] = {{-0.7203235f, 0.45654634f}, {-0.77601856f, -0.34911805f}, {-0.5384191f, 0.7197065f},
                         {-1.0381957f, -0.22146331f}, {-0.31787223f, -0.6780096f}};
  for (int i = 0; i < 4; i++) model.setUserVector(xid[i], {X[i][0], X[i][1]});
  for (int i = 0; i < 5; i++) model.setItemVector(yid[i], {Y[i][0], Y[i][1]});
  std::vector<UserItemStrength> input = {{"W100", "B1", 1.0}, {"X101", "C2", 1.0}, {"Y102", "D3", 1.0}, {"Z103", "E4", 1.0},
                                         {"A104", "F5", 1.0}, {"G6", "B105", 1.0}, {"H7", "C106", 1.0}, {"I8", "D107", 1.0},
                                         {"J9", "E108", 1.0}};
  auto updates = model.buildUpdates(input);
  CHECK(updates.size() == 9);  // "9 generate just 1 update since user or item is new"
  const float EX[5][2] = {{-0.20859924f, 0.25232133f}, {-0.22472803f, -0.1929485f}, {-0.15592135f, 0.3977631f},
                          {-0.3006522f, -0.12239703f}, {-0.09205295f, -0.37471837f}};
  const float EY[4][2] = {{-0.19663288f, 0.09574106f}, {-0.23840417f, -0.50850725f}, {-0.34360975f, 0.2466687f},
                          {-0.060204573f, 0.29311115f}};
  const char* newx[5] = {"W100", "X101", "Y102", "Z103", "A104"};
  const char* newy[4] = {"B105", "C106", "D107", "E108"};
  for (const auto& u : updates) {
    bool matched = false;
    if (u.matrix == 'X') {
      for (int i = 0; i < 5; i++) if (u.id == newx[i]) { matched = true; CHECK_NEAR(EX[i][0], u.vector[0], 1.0e-5); CHECK_NEAR(EX[i][1], u.vector[1], 1.0e-5); }
    } else {
      for (int i = 0; i < 4; i++) if (u.id == newy[i]) { matched = true; CHECK_NEAR(EY[i][0], u.vector[0], 1.0e-5); CHECK_NEAR(EY[i][1], u.vector[1], 1.0e-5); }
    }
    CHECK(matched);
  }
}

// COMMON-T/math/LinearSystemSolverTest.java:32-82 and APPCOMMON-T/als/ALSUtilsTest.java:31-84
static void testSolverAndALSUtils() {
  Solver s({1.3, -2.0, 3.0, 2.0, 0.0, 5.0, 0.0, -1.5, 5.5}, 3, 0);
  auto y = s.solveFToF({1.0f, 2.0f, 6.5f});
  CHECK(y[0] == -1.956044f && y[1] == 0.0021978023f && y[2] == 1.1824176f);
  int rank = -1;
  try { Solver bad({1.3001, -2.0, 3.0, 2.6, -4.0001, 6.0001, 0.0, -1.5, 5.5}, 3, 0); }
  catch (const SingularMatrixSolverException& e) { rank = e.getApparentRank(); }
  CHECK(rank == 2);
  rank = -1;
  try { Solver bad({1.3001, -2.0, 3.0, 2.6, -4.0001, 6.0001, 1.3, -2.0002, 3.0002}, 3, 0); }
  catch (const SingularMatrixSolverException& e) { rank = e.getApparentRank(); }
  CHECK(rank == 1);
  CHECK(std::isnan(ALSUtils::computeTargetQui(true, 0.0, 1.0)) && std::isnan(ALSUtils::computeTargetQui(true, 0.5, 1.0)));
  CHECK(ALSUtils::computeTargetQui(true, 1.0, 0.5) == 0.75 && ALSUtils::computeTargetQui(true, -1.0, 0.5) == 0.25);
  CHECK(ALSUtils::computeTargetQui(false, 2.0, 0.0) == 2.0);
  // rows {1,2},{3,0},{0,1}: transposeTimesSelf = {{10,2},{2,5}}
  Solver g({10.0, 2.0, 2.0, 5.0}, 2, 0);
  FloatVec yi = {2.0f, 1.0f}, xu = {0.1f, 0.1f};
  CHECK(ALSUtils::computeUpdatedXu(g, 1.0, nullptr, nullptr, true) == nullptr);
  auto a = ALSUtils::computeUpdatedXu(g, 1.0, nullptr, &yi, true);
  CHECK(a && (*a)[0] == 0.13043478f && (*a)[1] == 0.097826086f);
  auto b = ALSUtils::computeUpdatedXu(g, 0.5, nullptr, &yi, true);
  CHECK(b && (*b)[0] == 0.11594203f && (*b)[1] == 0.08695652f);
  auto c = ALSUtils::computeUpdatedXu(g, 1.0, &xu, &yi, true);
  CHECK(c && (*c)[0] == 0.16086957f && (*c)[1] == 0.14565217f);
}

// ---- the update topic as MockALSModelUpdateGenerator.java:67-100 produces it: message 0 is the MODEL
// skeleton, messages 1..9 are "UP" payloads (ids 1-5 are Y rows with known users, 6-9 X rows with known items)
static std::string idToStringID(int id) { return std::string(1, (char)('A' + id % 26)) + std::to_string(id); }  // ALSUtilsTest.java:50-52
static const float GEN_X[4][2] = {{-0.6790019f, 0.1732324f}, {-0.8232442f, -0.9200852f}, {-1.1865344f, 0.44631857f}, {-0.20789514f, 0.5303508f}};
static const float GEN_Y[5][2] = {{-0.7203235f, 0.45654634f}, {-0.77601856f, -0.34911805f}, {-0.5384191f, 0.7197065f},
                                  {-1.0381957f, -0.22146331f}, {-0.31787223f, -0.6780096f}};
static const std::vector<std::vector<std::string>> GEN_A = {{"1", "4"}, {"2", "4", "5"}, {"1", "2", "3", "4"}, {"3"}};        // users 6..9
static const std::vector<std::vector<std::string>> GEN_At = {{"6", "8"}, {"7", "8"}, {"8", "9"}, {"6", "7", "8"}, {"7"}};     // items 1..5
static std::vector<KeyMessage> mockUpdateTopic() {
  std::vector<KeyMessage> out;
  std::string xids, yids;
  for (int i = 6; i <= 9; i++) xids += (xids.empty() ? "" : " ") + idToStringID(i);
  for (int i = 1; i <= 5; i++) yids += (yids.empty() ? "" : " ") + idToStringID(i);
  out.push_back({"MODEL",
                 "<?xml version=\"1.0\" encoding=\"UTF-8\" standalone=\"yes\"?>\n<PMML version=\"4.3\" "
                 "xmlns=\"http://www.dmg.org/PMML-4_3\">\n    <Header>\n        <Application name=\"Oryx\"/>\n"
                 "        <Timestamp>2026-01-01T00:00:00+0000</Timestamp>\n    </Header>\n"
                 "    <Extension name=\"features\" value=\"2\"/>\n    <Extension name=\"implicit\" value=\"true\"/>\n"
                 "    <Extension name=\"XIDs\">" + xids + "</Extension>\n    <Extension name=\"YIDs\">" + yids +
                 "</Extension>\n</PMML>\n"});
  auto others = [](const std::vector<std::string>& v) {
    std::string s = "[";
    for (size_t i = 0; i < v.size(); i++) s += (i ? ",\"" : "\"") + v[i] + "\"";
    return s + "]";
  };
  for (int id = 1; id <= 9; id++) {
    std::string msg;
    if (id >= 6) {
      msg = toUpdateJSON('X', idToStringID(id), GEN_X[id - 6], 2, nullptr);
      msg.insert(msg.size() - 1, "," + others(GEN_A[id - 6]));
    } else {
      msg = toUpdateJSON('Y', idToStringID(id), GEN_Y[id - 1], 2, nullptr);
      msg.insert(msg.size() - 1, "," + others(GEN_At[id - 1]));
    }
    out.push_back({"UP", msg});
  }
  return out;
}

// SERVING-T/als/model/ALSServingModelManagerIT.java:35-69
static void testServingModelManager() {
  ALSServingModelManager manager(1.0, 4);
  auto topic = mockUpdateTopic();
  manager.consume({topic[5]});  // an "UP" before any MODEL is skipped (ALSServingModelManager.java:73-75)
  CHECK(manager.getModel() == nullptr);
  manager.consume(topic);
  ALSServingModel* model = manager.getModel();
  CHECK(model != nullptr);
  if (!model) return;
  CHECK(model->getFeatures() == 2 && model->isImplicit());
  auto ids = model->getAllItemIDs();
  std::set<std::string> got(ids.begin(), ids.end()), want;
  for (int i = 1; i <= 5; i++) want.insert(idToStringID(i));
  CHECK(got == want);
  CHECK(model->getYTYSolver() != nullptr);
  for (int i = 6; i <= 9; i++) {
    auto v = model->getUserVector(idToStringID(i));
    CHECK(v && (*v)[0] == GEN_X[i - 6][0] && (*v)[1] == GEN_X[i - 6][1]);
    auto known = model->getKnownItems(idToStringID(i));
    CHECK(known == std::unordered_set<std::string>(GEN_A[i - 6].begin(), GEN_A[i - 6].end()));
  }
  for (int i = 1; i <= 5; i++) {
    auto v = model->getItemVector(idToStringID(i));
    CHECK(v && (*v)[0] == GEN_Y[i - 1][0] && (*v)[1] == GEN_Y[i - 1][1]);
  }
  CHECK(model->getFractionLoaded() == 1.0f);
  // a second MODEL naming fewer ids: ids set since the previous MODEL are "recent" and survive once
  // (FeatureVectors.java:131-136), then go
  KeyMessage smaller = topic[0];
  size_t at = smaller.message.find(idToStringID(5));
  smaller.message.erase(at - 1, idToStringID(5).size() + 1);
  manager.consume({smaller});
  CHECK(model->getItemVector(idToStringID(5)) != nullptr);
  manager.consume({smaller});
  CHECK(model->getItemVector(idToStringID(5)) == nullptr && model->getNumItems() == 4);
  // malformed traffic: the stream stops at the bad message, earlier ones are applied
  bool threw = false;
  try {
    manager.consume({{"UP", "[\"Y\",\"N1\",[1.0,2.0]]"}, {"UP", "[\"Q\",\"N2\",[1.0,2.0]]"}, {"UP", "[\"Y\",\"N3\",[1.0,2.0]]"}});
  } catch (const BadMessage&) { threw = true; }
  CHECK(threw && model->getItemVector("N1") != nullptr && model->getItemVector("N3") == nullptr);
  threw = false;
  try { manager.consume({{"UP", "[\"Y\",\"N4\",[1.0,2.0,3.0]]"}}); } catch (const std::invalid_argument&) { threw = true; }
  CHECK(threw);  // checkArgument(vector.length == features), ALSServingModel.java:162
  threw = false;
  try { manager.consume({{"WHAT", "x"}}); } catch (const std::invalid_argument&) { threw = true; }
  CHECK(threw);  // ALSServingModelManager.java:125-126
  // part-file bulk load (ALSUpdate.java:415-428 format)
  std::string text;
  std::vector<int64_t> begin{0};
  for (int i = 0; i < 3000; i++) {
    float v[2] = {(float)i * 0.25f, -1.0f / (float)(i + 1)};
    text += toFeatureLine("bulk" + std::to_string(i), v, 2);
    begin.push_back((int64_t)text.size());
  }
  manager.loadFeatureLines('Y', text.data(), begin.data(), 3000);
  auto v = model->getItemVector("bulk2999");
  CHECK(v && (*v)[0] == 2999 * 0.25f && (*v)[1] == -1.0f / 3000.0f && model->getNumItems() == 4 + 1 + 3000);
}

// SPEED-T/ALSSpeedIT.java:50-113 end to end through the text formats: update topic in, input lines in,
// "UP" payloads out
static void testSpeedModelManager() {
  ALSSpeedModelManager manager(/*noKnownItems=*/false);
  CHECK(manager.buildUpdates({"a,b,1,1"}).empty());  // no model yet (ALSSpeedModelManager.java:136-138)
  manager.consume(mockUpdateTopic());
  CHECK(manager.getModel() && manager.getModel()->getFeatures() == 2 && manager.getModel()->isImplicit());
  std::vector<std::string> lines;  // MockALSInputGenerator.java:26-37, ids 0..8
  for (int id = 0; id < 9; id++) {
    std::string large = idToStringID(100 + id), small = idToStringID(1 + id);
    lines.push_back((id < 5 ? large + "," + small : small + "," + large) + ",1," + std::to_string(1000 + id));
  }
  auto updates = manager.buildUpdates(lines);
  CHECK(updates.size() == 9);
  const float EX[5][2] = {{-0.20859924f, 0.25232133f}, {-0.22472803f, -0.1929485f}, {-0.15592135f, 0.3977631f},
                          {-0.3006522f, -0.12239703f}, {-0.09205295f, -0.37471837f}};
  const float EY[4][2] = {{-0.19663288f, 0.09574106f}, {-0.23840417f, -0.50850725f}, {-0.34360975f, 0.2466687f},
                          {-0.060204573f, 0.29311115f}};
  for (const auto& msg : updates) {
    char matrix;
    std::string id;
    std::vector<float> vec;
    bool has;
    std::vector<std::string> others;
    parseFeatureMessage(msg, true, &matrix, &id, &vec, &has, &others);
    int n = std::atoi(id.c_str() + 1);
    const float* e = matrix == 'X' ? EX[n - 100] : EY[n - 105];
    CHECK((matrix == 'X') == (n < 105) && vec.size() == 2);
    CHECK_NEAR(e[0], vec[0], 1.0e-5);
    CHECK_NEAR(e[1], vec[1], 1.0e-5);
    CHECK(has && others.size() == 1 && others[0] == idToStringID(n - 99));  // ALSSpeedIT.java:107-111
  }
  // aggregation (ALSSpeedModelManager.java:141-171): timestamp order, implicit sums with NaN as a
  // delete marker, explicit last-wins
  std::vector<std::string> agg = {"u,i,2,30", "u,i,1,10", "u,j,NaN,5", "[\"u\",\"i\",\"4\",20]", "v,i,1,7", "v,i,NaN,8", "w,k,3,1", "w,k,NaN,2", "w,k,5,3"};
  auto imp = manager.aggregate(agg, true);
  CHECK(imp.size() == 2 && imp[0].user == "w" && imp[0].strength == 5.0 && imp[1].user == "u" && imp[1].item == "i" && imp[1].strength == 7.0);
  auto exp = manager.aggregate(agg, false);
  CHECK(exp.size() == 2 && exp[0].user == "w" && exp[0].strength == 5.0 && exp[1].user == "u" && exp[1].strength == 2.0);
  bool threw = false;
  try { manager.aggregate({"u,i,,1"}, true); } catch (const BadMessage&) { threw = true; }
  CHECK(threw);  // Double.valueOf("") -> NumberFormatException
  threw = false;
  try { manager.aggregate({"u,i,1"}, true); } catch (const BadMessage&) { threw = true; }
  CHECK(threw);  // tokens[3] -> ArrayIndexOutOfBoundsException
  ALSSpeedModelManager quiet(/*noKnownItems=*/true);
  quiet.consume(mockUpdateTopic());
  auto q = quiet.buildUpdates({lines[0]});
  CHECK(q.size() == 1 && q[0].find("],[") == std::string::npos);  // no trailing known-ids element (:209-210)
}

// Evaluation.java:49-145 (batch layer's model evaluation) on device-resident X / Y
static void testEvaluation() {
  const int F = 6, NU = 40, NI = 60;
  ALSSpeedModel model(F, true);
  std::mt19937_64 rng(77);
  std::normal_distribution<float> g(0.f, 1.f);
  std::vector<FloatVec> X(NU, FloatVec(F)), Y(NI, FloatVec(F));
  for (int u = 0; u < NU; u++) { for (auto& v : X[u]) v = g(rng); model.setUserVector("u" + std::to_string(u), X[u]); }
  for (int i = 0; i < NI; i++) { for (auto& v : Y[i]) v = g(rng); model.setItemVector("i" + std::to_string(i), Y[i]); }
  // rows were handed out in insertion order: user u -> row u, item i -> row i
  auto predict = [&](int u, int i) { double s = 0; for (int j = 0; j < F; j++) s += (double)X[u][j] * (double)Y[i][j]; return s; };
  std::vector<Rating> test;
  double sum = 0;
  for (int k = 0; k < 200; k++) {
    int u = (int)(rng() % NU), i = (int)(rng() % NI);
    double r = (double)(rng() % 5);
    test.push_back({u, i, r});
    double d = predict(u, i) - r;
    sum += d * d;
  }
  test.push_back({NU + 3, 1, 4.0});  // unknown user: no prediction, dropped by the join
  CHECK_NEAR(Evaluation::rmse(model.handle(), test), std::sqrt(sum / 200.0), 1e-12);
  // AUC: positives = each user's 5 best items -> every sampled negative ranks below every positive... not
  // necessarily (a negative can outrank the 5th best), so compare with a direct recomputation instead
  std::vector<Rating> positives;
  for (int u = 0; u < NU; u++) {
    std::vector<std::pair<double, int>> s;
    for (int i = 0; i < NI; i++) s.push_back({predict(u, i), i});
    std::sort(s.rbegin(), s.rend());
    for (int k = 0; k < 5; k++) positives.push_back({u, s[k].second, 1.0});
  }
  auto negatives = Evaluation::sampleNegatives(positives, 1234);
  CHECK(!negatives.empty() && negatives.size() <= positives.size());
  std::set<std::pair<int, int>> posSet;
  for (auto& r : positives) posSet.insert({r.user, r.product});
  for (auto& n : negatives) CHECK(!posSet.count(n));
  double aucSum = 0; int aucUsers = 0;
  for (int u = 0; u < NU; u++) {
    long correct = 0, total = 0;
    for (auto& r : positives) if (r.user == u)
      for (auto& n : negatives) if (n.first == u) { if (predict(u, r.product) > predict(u, n.second)) correct++; total++; }
    if (total) { aucSum += (double)correct / total; aucUsers++; }
  }
  const double auc = Evaluation::areaUnderCurve(model.handle(), positives, negatives);
  CHECK_NEAR(auc, aucSum / aucUsers, 1e-12);
  CHECK(auc > 0.9);  // positives are each user's best items
}

int main() {
  try {
    testSolverAndALSUtils();
    testServingGoldens();
    testLSH();
    testSpeedFoldIn();
    testServingModelManager();
    testSpeedModelManager();
    testEvaluation();
  } catch (const std::exception& e) {
    std::printf("EXCEPTION: %s\n", e.what());
    return 2;
  }
  std::printf("%d checks, %d failed\n", g_checks, g_fail);
  return g_fail == 0 ? 0 : 1;
}
