// Timed CPU baseline: multi-threaded C++ restatements of the benchmark queries' operator
// semantics (SURVEY §7.2 — the reference's LocalDebug / vertex runtime is C# and cannot run here).
//   cpu_wordcount        Split(' ') per line + hash GroupBy/Count (samples/WordCount.cs.pp:72-75;
//                        per-thread tables merged at the end, like ParallelHashGroupByPartialAccumulate,
//                        LinqToDryad/DryadLinqVertex.cs:5718-6032)
//   cpu_sort_rows        OrderBy on fixed-width rows: chunked std::sort + k-way merge
//                        (ParallelSort, DryadLinqVertex.cs:9321-9818)
//   cpu_groupby_sum_count  decomposable Sum/Count per int64 key
#include "../../include/dryad_host.h"

#include <string.h>

#include <algorithm>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

namespace {
struct SV {
  const char* p; size_t n;
  bool operator==(const SV& o) const { return n == o.n && memcmp(p, o.p, n) == 0; }
};
struct SVHash {
  size_t operator()(const SV& s) const {
    uint64_t h = 0xcbf29ce484222325ULL;
    for (size_t i = 0; i < s.n; ++i) { h ^= (unsigned char)s.p[i]; h *= 0x100000001b3ULL; }
    return (size_t)h;
  }
};
typedef std::unordered_map<SV, int64_t, SVHash> WordMap;
}  // namespace

extern "C" {

// keep_empty = 1: String.Split(' ') semantics.  Returns distinct words; *total_tokens out.
int64_t cpu_wordcount(const char* text, int64_t n, int threads, int keep_empty, int64_t* total_tokens) {
  if (threads < 1) threads = 1;
  std::vector<WordMap> maps(threads);
  std::vector<int64_t> totals(threads, 0);
  std::vector<std::thread> th;
  // split at line boundaries so no token straddles two workers
  std::vector<int64_t> cut(threads + 1, n);
  cut[0] = 0;
  for (int t = 1; t < threads; ++t) {
    int64_t c = n * t / threads;
    while (c < n && text[c] != '\n') ++c;
    cut[t] = c < n ? c + 1 : n;
  }
  for (int t = 0; t < threads; ++t)
    th.emplace_back([&, t] {
      WordMap& m = maps[t];
      int64_t i = cut[t], e = cut[t + 1], tot = 0;
      while (i < e) {
        int64_t le = i;
        while (le < e && text[le] != '\n') ++le;      // one line [i, le)
        int64_t s = i;
        for (int64_t j = i; j <= le; ++j) {
          if (j == le || text[j] == ' ') {
            if (keep_empty || j > s) { m[SV{text + s, (size_t)(j - s)}]++; ++tot; }
            s = j + 1;
          }
        }
        i = le + 1;
      }
      totals[t] = tot;
    });
  for (auto& x : th) x.join();
  WordMap& all = maps[0];
  int64_t tot = totals[0];
  for (int t = 1; t < threads; ++t) { for (auto& kv : maps[t]) all[kv.first] += kv.second; tot += totals[t]; }
  if (total_tokens) *total_tokens = tot;
  return (int64_t)all.size();
}

// Sorts row indices by the first key_len bytes (bytewise), then permutes rows into out.
int cpu_sort_rows(const uint8_t* rows, int64_t n, int width, int key_len, int threads, uint8_t* out) {
  if (threads < 1) threads = 1;
  std::vector<uint32_t> idx(n);
  for (int64_t i = 0; i < n; ++i) idx[i] = (uint32_t)i;
  auto less = [&](uint32_t a, uint32_t b) {
    int c = memcmp(rows + (size_t)a * width, rows + (size_t)b * width, (size_t)key_len);
    return c < 0 || (c == 0 && a < b);
  };
  std::vector<int64_t> cut(threads + 1);
  for (int t = 0; t <= threads; ++t) cut[t] = n * t / threads;
  std::vector<std::thread> th;
  for (int t = 0; t < threads; ++t) th.emplace_back([&, t] { std::sort(idx.begin() + cut[t], idx.begin() + cut[t + 1], less); });
  for (auto& x : th) x.join();
  for (int step = 1; step < threads; step *= 2) {           // pairwise merges of sorted runs
    std::vector<std::thread> mt;
    for (int t = 0; t + step < threads; t += 2 * step)
      mt.emplace_back([&, t, step] {
        int hi = std::min(threads, t + 2 * step);
        std::inplace_merge(idx.begin() + cut[t], idx.begin() + cut[t + step], idx.begin() + cut[hi], less);
      });
    for (auto& x : mt) x.join();
  }
  std::vector<std::thread> pt;
  for (int t = 0; t < threads; ++t)
    pt.emplace_back([&, t] { for (int64_t i = cut[t]; i < cut[t + 1]; ++i) memcpy(out + (size_t)i * width, rows + (size_t)idx[i] * width, (size_t)width); });
  for (auto& x : pt) x.join();
  return 0;
}

// Returns the number of groups; sums/counts are not returned (timing baseline), checksum out.
int64_t cpu_groupby_sum_count(const int64_t* keys, const int64_t* vals, int64_t n, int threads, int64_t* checksum) {
  if (threads < 1) threads = 1;
  typedef std::unordered_map<int64_t, std::pair<int64_t, int64_t>> M;
  std::vector<M> maps(threads);
  std::vector<std::thread> th;
  for (int t = 0; t < threads; ++t)
    th.emplace_back([&, t] {
      M& m = maps[t];
      for (int64_t i = n * t / threads; i < n * (t + 1) / threads; ++i) { auto& e = m[keys[i]]; e.first += vals[i]; e.second++; }
    });
  for (auto& x : th) x.join();
  M& all = maps[0];
  for (int t = 1; t < threads; ++t) for (auto& kv : maps[t]) { auto& e = all[kv.first]; e.first += kv.second.first; e.second += kv.second.second; }
  int64_t cs = 0;
  for (auto& kv : all) cs += kv.second.first + kv.second.second;
  if (checksum) *checksum = cs;
  return (int64_t)all.size();
}

}  // extern "C"
