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

// A slot keeps the word's first 8 bytes and its length next to the hash: words of up to 8 bytes are
// compared without touching the text they were first seen in (that text is cold by then — comparing
// through the pointer cost one DRAM miss per token).
struct Entry { const char* p; uint64_t h; int64_t c; uint64_t prefix; uint32_t n; };

inline uint64_t load_prefix(const char* p, size_t n, const char* end) {
  uint64_t v = 0;
  if (p + 8 <= end) memcpy(&v, p, 8); else memcpy(&v, p, (size_t)(end - p));
  return n >= 8 ? v : (v & ((1ULL << (8 * n)) - 1));
}

inline uint64_t mix64(uint64_t x) {
  x ^= x >> 33; x *= 0xff51afd7ed558ccdULL; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL; x ^= x >> 33;
  return x;
}

struct Table {
  std::vector<Entry> slots;
  size_t used = 0, mask = 0;
  explicit Table(size_t cap = 1 << 12) { slots.assign(cap, Entry{nullptr, 0, 0, 0, 0}); mask = cap - 1; }
  // 8 bytes at a time; `prefix` is the (masked) first word, already loaded by the caller
  static uint64_t hash(const char* p, size_t n, uint64_t prefix) {
    uint64_t h = mix64(prefix ^ (0x9E3779B97F4A7C15ULL * (uint64_t)(n + 1)));
    for (size_t i = 8; i < n; i += 8) {
      uint64_t v = 0;
      memcpy(&v, p + i, n - i >= 8 ? 8 : n - i);
      h = mix64(h ^ v);
    }
    return h | 1;                                   // never 0: 0 marks an empty slot
  }
  void grow() {
    std::vector<Entry> old;
    old.swap(slots);
    slots.assign(old.size() * 2, Entry{nullptr, 0, 0, 0, 0});
    mask = slots.size() - 1;
    for (const Entry& e : old)
      if (e.h) { size_t i = e.h & mask; while (slots[i].h) i = (i + 1) & mask; slots[i] = e; }
  }
  void add(const char* p, size_t n, uint64_t h, uint64_t prefix, int64_t c) {
    size_t i = h & mask;
    for (;;) {
      Entry& e = slots[i];
      if (!e.h) {
        e = Entry{p, h, c, prefix, (uint32_t)n};
        if (++used * 2 > slots.size()) grow();
        return;
      }
      if (e.h == h && e.n == n && e.prefix == prefix && (n <= 8 || memcmp(e.p + 8, p + 8, n - 8) == 0)) { e.c += c; return; }
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
  const char* const text_end = text + n;
  auto work = [&](int t) {
    Table& m = tabs[t];
    int64_t e = cut[t + 1], tot = 0;
    int64_t s = cut[t];                               // start of the current token
    for (int64_t j = cut[t]; j < e; ++j) {            // one pass: a token ends at the delimiter or at the end of its line
      const char ch = text[j];
      if (ch == d || ch == '\n') {
        if (keep_empty || j > s) {
          const size_t len = (size_t)(j - s);
          const uint64_t pre = load_prefix(text + s, len, text_end);
          m.add(text + s, len, Table::hash(text + s, len, pre), pre, 1);
          ++tot;
        }
        s = j + 1;
      }
    }
    if (s < e || (keep_empty && s == e && e > cut[t] && text[e - 1] == d)) {   // last line without a newline
      const size_t len = (size_t)(e - s);
      const uint64_t pre = load_prefix(text + s, len, text_end);
      m.add(text + s, len, Table::hash(text + s, len, pre), pre, 1);
      ++tot;
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
    for (const Entry& e : tabs[t].slots) if (e.h) r->table.add(e.p, e.n, e.h, e.prefix, e.c);
    r->total += totals[t];
  }
  for (const Entry& e : r->table.slots) if (e.h) r->word_bytes += e.n;
  return r;
}

// Occurrences of one byte (line count of a text partition for the vertex statistics): memchr speed.
int64_t hv_count_byte(const char* text, int64_t n, int byte) {
  int64_t c = 0;
  const char* p = text;
  const char* end = text + n;
  while (p < end) {
    const char* q = (const char*)memchr(p, byte, (size_t)(end - p));
    if (!q) break;
    ++c;
    p = q + 1;
  }
  return c;
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
