// bench_ingest.cpp -- model-load throughput of the update-stream path:
// "UP" payloads ["Y", id, [F floats]] -> native parse on all cores -> bulk oryx_items_upsert (H2D +
// on-device LSH hashing).  Prints one JSON line.  usage: bench_ingest [rows] [features] [threads]
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <thread>

#include "update_stream.hpp"

using namespace oryx::host;
using Clock = std::chrono::steady_clock;
static double secs(Clock::time_point a, Clock::time_point b) { return std::chrono::duration<double>(b - a).count(); }

int main(int argc, char** argv) {
  const int64_t rows = argc > 1 ? std::atoll(argv[1]) : 1000000;
  const int F = argc > 2 ? std::atoi(argv[2]) : 128;
  int threads = argc > 3 ? std::atoi(argv[3]) : 0;
  if (threads <= 0) threads = (int)std::thread::hardware_concurrency();
  // 1. synthetic update topic: N(0,1) components (LoadTestALSModelFactory.java:62-95), ids "I<j>"
  std::vector<std::string> chunks((size_t)threads);
  std::vector<std::vector<int64_t>> lens((size_t)threads);
  std::vector<float> truth((size_t)rows * F);
  {
    std::vector<std::thread> pool;
    for (int t = 0; t < threads; t++) {
      pool.emplace_back([&, t] {
        std::mt19937_64 rng(0xB2000020ULL + (uint64_t)t);
        std::normal_distribution<float> gauss(0.0f, 1.0f);
        const int64_t lo = rows * t / threads, hi = rows * (t + 1) / threads;
        for (int64_t r = lo; r < hi; r++) {
          float* v = &truth[(size_t)r * F];
          for (int j = 0; j < F; j++) v[j] = gauss(rng);
          std::string m = toUpdateJSON('Y', "I" + std::to_string(r), v, F, nullptr);
          lens[(size_t)t].push_back((int64_t)m.size());
          chunks[(size_t)t] += m;
        }
      });
    }
    for (auto& th : pool) th.join();
  }
  std::string text;
  std::vector<int64_t> begin{0};
  begin.reserve((size_t)rows + 1);
  for (int t = 0; t < threads; t++) {
    text += chunks[(size_t)t];
    chunks[(size_t)t].clear();
    chunks[(size_t)t].shrink_to_fit();
    for (int64_t l : lens[(size_t)t]) begin.push_back(begin.back() + l);
  }
  const double text_bytes = (double)text.size();

  // 2. parse only
  auto t0 = Clock::now();
  int64_t parsed_rows = 0;
  for (int64_t lo = 0; lo < rows; lo += 1 << 20) {
    const int64_t n = std::min<int64_t>(rows - lo, 1 << 20);
    for (auto& part : parseFeatureMessageParts(text.data(), begin.data() + lo, n, true, threads)) parsed_rows += (int64_t)part.size();
  }
  auto t1 = Clock::now();

  // 3. the full consume(): MODEL skeleton, then the "UP" run
  ALSServingModelManager manager(1.0, 32, 0, threads);
  manager.consume({{"MODEL",
                    "<PMML version=\"4.3\"><Header/><Extension name=\"features\" value=\"" + std::to_string(F) +
                        "\"/><Extension name=\"implicit\" value=\"true\"/><Extension name=\"XIDs\">U0</Extension>"
                        "<Extension name=\"YIDs\">I0</Extension></PMML>"}});
  // warm the device path (context creation, first allocation) outside the timed region
  manager.consume({{"UP", "[\"Y\",\"warm\",[" + [&] { std::string s; for (int j = 0; j < F; j++) s += j ? ",0.5" : "0.5"; return s; }() + "]]"}});
  auto t2 = Clock::now();
  manager.consumeUpdates(text.data(), begin.data(), rows);
  auto t3 = Clock::now();

  // 4. check: every 9973rd row reads back with the bits it was written with
  ALSServingModel* model = manager.getModel();
  int64_t checked = 0, wrong = 0;
  for (int64_t r = 0; r < rows; r += 9973) {
    auto v = model->getItemVector("I" + std::to_string(r));
    checked++;
    if (!v || std::memcmp(v->data(), &truth[(size_t)r * F], sizeof(float) * (size_t)F) != 0) wrong++;
  }
  const double parse_s = secs(t0, t1), load_s = secs(t2, t3);
  std::printf("{\"workload\": \"update-topic model load: %lld UP messages x %d f\", \"threads\": %d, \"text_bytes\": %.0f, "
              "\"parse_rows_per_sec\": %.0f, \"parse_MB_per_sec\": %.1f, \"load_rows_per_sec\": %.0f, "
              "\"load_MB_per_sec\": %.1f, \"load_seconds\": %.3f, \"items_in_model\": %d, \"readback_checked\": %lld, "
              "\"readback_wrong\": %lld}\n",
              (long long)rows, F, threads, text_bytes, parsed_rows / parse_s, text_bytes / parse_s / 1e6, rows / load_s,
              text_bytes / load_s / 1e6, load_s, model->getNumItems(), (long long)checked, (long long)wrong);
  return wrong == 0 && parsed_rows == rows ? 0 : 1;
}
