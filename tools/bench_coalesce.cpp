// bench_coalesce.cpp -- throughput of oryx_topn_coalesced under many native request threads (what a JVM's
// worker pool looks like to the library), next to oryx_topn from the same threads.
// usage: bench_coalesce [items] [features] [threads] [queries_per_thread]
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <thread>
#include <vector>

#include "oryx_b200.h"

using Clock = std::chrono::steady_clock;

int main(int argc, char** argv) {
  const int64_t n = argc > 1 ? std::atoll(argv[1]) : 10000000;
  const int F = argc > 2 ? std::atoi(argv[2]) : 128;
  const int T = argc > 3 ? std::atoi(argv[3]) : 128;
  const int per = argc > 4 ? std::atoi(argv[4]) : 200;
  const int HM = 10;
  oryx_config cfg{};
  cfg.features = F;
  cfg.implicit = 1;
  cfg.capacity_hint = n;
  oryx_model* m = nullptr;
  if (oryx_model_create(&cfg, &m) != ORYX_OK) { std::puts("model_create failed"); return 2; }
  // N(0,1) item vectors, loaded in chunks
  {
    const int64_t chunk = 1 << 20;
    const int gen_threads = (int)std::max(1u, std::thread::hardware_concurrency());
    std::vector<float> buf((size_t)chunk * F);
    std::vector<int32_t> rows((size_t)chunk);
    for (int64_t off = 0; off < n; off += chunk) {
      const int64_t c = std::min(chunk, n - off);
      std::vector<std::thread> pool;
      for (int t = 0; t < gen_threads; t++) pool.emplace_back([&, t] {
        std::mt19937_64 rng(0xB20000E0ULL + (uint64_t)off * 131 + (uint64_t)t);
        std::normal_distribution<float> g(0.f, 1.f);
        for (int64_t r = c * t / gen_threads; r < c * (t + 1) / gen_threads; r++)
          for (int j = 0; j < F; j++) buf[(size_t)r * F + j] = g(rng);
      });
      for (auto& th : pool) th.join();
      for (int64_t r = 0; r < c; r++) rows[(size_t)r] = (int32_t)(off + r);
      if (oryx_items_upsert(m, ORYX_Y, rows.data(), buf.data(), (int32_t)c, nullptr) != ORYX_OK) { std::puts("upsert failed"); return 2; }
    }
  }
  const int NQ = 256;
  std::vector<float> q((size_t)NQ * F);
  {
    std::mt19937_64 rng(0xB20000E1ULL);
    std::normal_distribution<float> g(0.f, 1.f);
    for (auto& v : q) v = g(rng);
  }
  // build the bf16 shadow once (a page of 32 through the batch entry point)
  {
    std::vector<int32_t> r(32 * HM), c(32);
    std::vector<double> s(32 * HM);
    oryx_topn_batch(m, ORYX_DOT, q.data(), 32, HM, nullptr, nullptr, r.data(), s.data(), c.data());
  }
  auto run = [&](bool coalesced, int per) {
    std::atomic<int> bad{0};
    auto t0 = Clock::now();
    std::vector<std::thread> pool;
    for (int t = 0; t < T; t++) pool.emplace_back([&, t] {
      int32_t rows[HM], cnt = 0;
      double scores[HM];
      for (int i = 0; i < per; i++) {
        const float* qq = &q[(size_t)((t * per + i) % NQ) * F];
        int32_t st = coalesced ? oryx_topn_coalesced(m, ORYX_DOT, qq, HM, nullptr, 0, rows, scores, &cnt)
                               : oryx_topn(m, ORYX_DOT, qq, HM, nullptr, 0, rows, scores, &cnt);
        if (st != ORYX_OK || cnt != HM) bad++;
      }
    });
    for (auto& th : pool) th.join();
    double s = std::chrono::duration<double>(Clock::now() - t0).count();
    return std::make_pair((double)T * per / s, bad.load());
  };
  // identical answers
  int mismatches = 0;
  for (int i = 0; i < 16; i++) {
    int32_t r1[HM], r2[HM], c1 = 0, c2 = 0;
    double s1[HM], s2[HM];
    oryx_topn(m, ORYX_DOT, &q[(size_t)i * F], HM, nullptr, 0, r1, s1, &c1);
    oryx_topn_coalesced(m, ORYX_DOT, &q[(size_t)i * F], HM, nullptr, 0, r2, s2, &c2);
    if (c1 != c2 || std::memcmp(r1, r2, sizeof(int32_t) * (size_t)c1) || std::memcmp(s1, s2, sizeof(double) * (size_t)c1)) mismatches++;
  }
  // the plain path is device-bound at ~1.4 k queries/s: a few calls per thread are enough to time it
  const int per_plain = std::min(per, 8);
  auto plain = run(false, per_plain);
  auto coal = run(true, per);
  std::printf("{\"workload\": \"%lld items x %d f, top-%d, %d native request threads, %d (oryx_topn) / %d (coalesced) single-query calls each\", "
              "\"oryx_topn_qps\": %.0f, \"oryx_topn_coalesced_qps\": %.0f, \"failed_calls\": %d, \"answers_differ\": %d}\n",
              (long long)n, F, HM, T, per_plain, per, plain.first, coal.first, plain.second + coal.second, mismatches);
  oryx_model_destroy(m);
  return (plain.second + coal.second + mismatches) ? 1 : 0;
}
