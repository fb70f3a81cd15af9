// Native host vertex operators for traced string pipelines.
//
// The reference compiles every vertex into a .NET method over typed records (DryadLinqCodeGen.cs:
// 1910-2138) and runs hash GroupBy with partial accumulation on worker threads
// (ParallelHashGroupByPartialAccumulate, LinqToDryad/DryadLinqVertex.cs:5718-6032).  The host backend
// of this engine interprets vertices over Python records; for the pipelines the expression tracer
// understands, the whole vertex body runs here instead:
//
//   hv_split_count   SelectMany(line => line.Split(c)) -> GroupBy(word => word) with Count() as the
//                    only accumulator, straight from the partition's raw text bytes: per-thread
//                    open-addressing tables over string VIEWS (no per-token allocation), merged at
//                    the end, exported as (word bytes, offsets, counts).
#include "../../include/dryad_host.h"

#include <string.h>

#include <algorithm>
#include <thread>
#include <vector>

namespace {

struct Entry { const char* p; uint32_t n; uint64_t h; int64_t c; };

struct Table {
  std::vector<Entry> slots;
  size_t used = 0, mask = 0;
  explicit Table(size_t cap = 1 << 12) { slots.assign(cap, Entry{nullptr, 0, 0, 0}); mask = cap - 1; }
  static uint64_t hash(const char* p, size_t n) {
    uint64_t h = 0xcbf29ce484222325ULL;
    for (size_t i = 0; i < n; ++i) { h ^= (unsigned char)p[i]; h *= 0x100000001b3ULL; }
    h ^= h >> 32;
    return h | 1;                                   // never 0: 0 marks an empty slot
  }
  void grow() {
    std::vector<Entry> old;
    old.swap(slots);
    slots.assign(old.size() * 2, Entry{nullptr, 0, 0, 0});
    mask = slots.size() - 1;
    for (const Entry& e : old)
      if (e.h) { size_t i = e.h & mask; while (slots[i].h) i = (i + 1) & mask; slots[i] = e; }
  }
  void add(const char* p, size_t n, uint64_t h, int64_t c) {
    size_t i = h & mask;
    for (;;) {
      Entry& e = slots[i];
      if (!e.h) {
        e = Entry{p, (uint32_t)n, h, c};
        if (++used * 2 > slots.size()) grow();
        return;
      }
      if (e.h == h && e.n == n && memcmp(e.p, p, n) == 0) { e.c += c; return; }
      i = (i + 1) & mask;
    }
  }
};

struct Result {
  Table table;
  int64_t total = 0;
  int64_t word_bytes = 0;
};

}  // namespace

extern "C" {

void* hv_split_count(const char* text, int64_t n, int delim, int keep_empty, int threads) {
  if (threads < 1) threads = 1;
  if (n < (1 << 16)) threads = 1;
  std::vector<Table> tabs(threads);
  std::vector<int64_t> totals(threads, 0);
  std::vector<int64_t> cut(threads + 1, n);
  cut[0] = 0;
  for (int t = 1; t < threads; ++t) {                // split at line boundaries: no token straddles two workers
    int64_t c = n * t / threads;
    while (c < n && text[c] != '\n') ++c;
    cut[t] = c < n ? c + 1 : n;
  }
  const char d = (char)delim;
  auto work = [&](int t) {
    Table& m = tabs[t];
    int64_t i = cut[t], e = cut[t + 1], tot = 0;
    while (i < e) {
      int64_t le = i;
      while (le < e && text[le] != '\n') ++le;        // one line [i, le)
      int64_t s = i;
      for (int64_t j = i; j <= le; ++j) {
        if (j == le || text[j] == d) {
          if (keep_empty || j > s) { m.add(text + s, (size_t)(j - s), Table::hash(text + s, (size_t)(j - s)), 1); ++tot; }
          s = j + 1;
        }
      }
      i = le + 1;
    }
    totals[t] = tot;
  };
  if (threads == 1) work(0);
  else {
    std::vector<std::thread> th;
    for (int t = 0; t < threads; ++t) th.emplace_back(work, t);
    for (auto& x : th) x.join();
  }
  Result* r = new Result();
  r->table = std::move(tabs[0]);
  r->total = totals[0];
  for (int t = 1; t < threads; ++t) {
    for (const Entry& e : tabs[t].slots) if (e.h) r->table.add(e.p, e.n, e.h, e.c);
    r->total += totals[t];
  }
  for (const Entry& e : r->table.slots) if (e.h) r->word_bytes += e.n;
  return r;
}

int64_t hv_groups(void* h) { return (int64_t)((Result*)h)->table.used; }
int64_t hv_word_bytes(void* h) { return ((Result*)h)->word_bytes; }
int64_t hv_total(void* h) { return ((Result*)h)->total; }

// words: hv_word_bytes bytes; offsets: groups + 1 entries; counts: groups entries.
void hv_export(void* h, char* words, int64_t* offsets, int64_t* counts) {
  Result* r = (Result*)h;
  int64_t g = 0, pos = 0;
  for (const Entry& e : r->table.slots)
    if (e.h) {
      offsets[g] = pos;
      memcpy(words + pos, e.p, e.n);
      pos += e.n;
      counts[g] = e.c;
      ++g;
    }
  offsets[g] = pos;
}

void hv_free(void* h) { delete (Result*)h; }

}  // extern "C"
