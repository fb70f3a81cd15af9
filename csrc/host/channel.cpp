// Host channel primitives: bounded FIFO of data blocks, block cache with spill, Rabin fingerprint.
//
// Reference behaviour (re-implemented natively):
//  * DryadVertex/VertexHost/system/channel/src/channelfifo.cpp + .h:25-241 — in-process FIFO
//    channels between co-located vertices: blocking with a depth (fifo://32/...,
//    GraphManager/vertex/DrOutputGenerator.cpp:236-241) or non-blocking (unbounded); end-of-stream
//    marker items (include/channelitem.h:95-130).
//  * ProcessService/Cache.cs:58-300 + SpillMachine.cs:39-282 — in-memory block cache for channel
//    data with low/high water marks that spills the oldest blocks to disk files.
//  * DryadVertex/VertexHost/system/classlib/DrFPrint.cpp, ms_fprint.cpp and LinqToDryad/Hash64.cs
//    — table-driven 64-bit Rabin fingerprint (polynomial 0x911498ae0e66bad6).
#include "../../include/dryad_host.h"

#include <stdio.h>
#include <string.h>

#include <chrono>
#include <condition_variable>
#include <deque>
#include <map>
#include <mutex>
#include <string>
#include <vector>

namespace {

struct Fifo {
  std::mutex mu;
  std::condition_variable not_full, not_empty;
  std::deque<std::string> q;
  size_t depth;               // 0 = unbounded (non-blocking writer)
  bool eos = false, aborted = false;
  uint64_t bytes_written = 0, bytes_read = 0;
};

struct CacheBlock { std::string data; bool spilled = false; int64_t size = 0; uint64_t tick = 0; };
struct Cache {
  std::mutex mu;
  std::map<std::string, CacheBlock> blocks;
  int64_t resident = 0, high, low;
  std::string dir;
  uint64_t tick = 0, spills = 0, loads = 0;
};

uint64_t g_fp_table[256];
bool g_fp_init = false;
const uint64_t kPoly = 0x911498ae0e66bad6ULL;

void fp_init() {
  // T[b] = (b * x^64) mod P over GF(2), bit-reflected form with the polynomial's low 64 bits.
  for (int b = 0; b < 256; ++b) {
    uint64_t v = (uint64_t)b;
    for (int i = 0; i < 8; ++i) v = (v & 1) ? (v >> 1) ^ kPoly : (v >> 1);
    g_fp_table[b] = v;
  }
  g_fp_init = true;
}

std::string spill_path(const Cache* c, const std::string& key) {
  std::string safe = key;
  for (auto& ch : safe) if (ch == '/' || ch == '\\' || ch == ':') ch = '_';
  return c->dir + "/" + safe + ".spill";
}

void enforce(Cache* c) {
  if (c->resident <= c->high) return;
  // spill least-recently-used resident blocks until below the low water mark
  while (c->resident > c->low) {
    CacheBlock* victim = nullptr; const std::string* vkey = nullptr;
    for (auto& kv : c->blocks)
      if (!kv.second.spilled && (!victim || kv.second.tick < victim->tick)) { victim = &kv.second; vkey = &kv.first; }
    if (!victim) break;
    FILE* f = fopen(spill_path(c, *vkey).c_str(), "wb");
    if (!f) break;
    fwrite(victim->data.data(), 1, victim->data.size(), f);
    fclose(f);
    c->resident -= victim->size;
    victim->data.clear(); victim->data.shrink_to_fit();
    victim->spilled = true;
    c->spills++;
  }
}

}  // namespace

extern "C" {

// ---- FIFO ---------------------------------------------------------------------------------------------------
void* fifo_create(int depth) { Fifo* f = new Fifo(); f->depth = depth > 0 ? (size_t)depth : 0; return f; }
void fifo_destroy(void* h) { delete (Fifo*)h; }
// 0 ok, -1 timeout, -2 aborted/closed
int fifo_write(void* h, const void* data, int64_t n, double timeout_seconds) {
  Fifo* f = (Fifo*)h;
  std::unique_lock<std::mutex> l(f->mu);
  if (f->depth) {
    bool ok = f->not_full.wait_for(l, std::chrono::duration<double>(timeout_seconds),
                                   [&] { return f->q.size() < f->depth || f->aborted; });
    if (!ok) return -1;
  }
  if (f->aborted || f->eos) return -2;
  f->q.emplace_back((const char*)data, (size_t)n);
  f->bytes_written += (uint64_t)n;
  f->not_empty.notify_one();
  return 0;
}
void fifo_close_writer(void* h) { Fifo* f = (Fifo*)h; { std::lock_guard<std::mutex> l(f->mu); f->eos = true; } f->not_empty.notify_all(); }
void fifo_abort(void* h) { Fifo* f = (Fifo*)h; { std::lock_guard<std::mutex> l(f->mu); f->aborted = true; } f->not_empty.notify_all(); f->not_full.notify_all(); }
// Returns the block size (>= 0) and copies it; -1 timeout, -2 end of stream, -3 aborted, -4 buffer too small
int64_t fifo_read(void* h, void* out, int64_t cap, double timeout_seconds) {
  Fifo* f = (Fifo*)h;
  std::unique_lock<std::mutex> l(f->mu);
  bool ok = f->not_empty.wait_for(l, std::chrono::duration<double>(timeout_seconds),
                                  [&] { return !f->q.empty() || f->eos || f->aborted; });
  if (!ok) return -1;
  if (f->aborted) return -3;
  if (f->q.empty()) return -2;
  std::string& b = f->q.front();
  if ((int64_t)b.size() > cap) return -4;
  int64_t n = (int64_t)b.size();
  memcpy(out, b.data(), (size_t)n);
  f->q.pop_front();
  f->bytes_read += (uint64_t)n;
  f->not_full.notify_one();
  return n;
}
int fifo_depth_now(void* h) { Fifo* f = (Fifo*)h; std::lock_guard<std::mutex> l(f->mu); return (int)f->q.size(); }

// ---- cache + spill ---------------------------------------------------------------------------------------
void* cache_create(const char* spill_dir, int64_t low_water, int64_t high_water) {
  Cache* c = new Cache(); c->dir = spill_dir ? spill_dir : "/tmp"; c->low = low_water; c->high = high_water; return c;
}
void cache_destroy(void* h) {
  Cache* c = (Cache*)h;
  for (auto& kv : c->blocks) if (kv.second.spilled) remove(spill_path(c, kv.first).c_str());
  delete c;
}
int cache_put(void* h, const char* key, const void* data, int64_t n) {
  Cache* c = (Cache*)h;
  std::lock_guard<std::mutex> l(c->mu);
  CacheBlock& b = c->blocks[key];
  if (!b.spilled) c->resident -= b.size;
  b.data.assign((const char*)data, (size_t)n); b.size = n; b.spilled = false; b.tick = ++c->tick;
  c->resident += n;
  enforce(c);
  return 0;
}
// Returns size, or -1 if absent, -4 if cap too small.  Spilled blocks are read back from disk.
int64_t cache_get(void* h, const char* key, void* out, int64_t cap) {
  Cache* c = (Cache*)h;
  std::lock_guard<std::mutex> l(c->mu);
  auto it = c->blocks.find(key);
  if (it == c->blocks.end()) return -1;
  CacheBlock& b = it->second;
  if (b.size > cap) return -4;
  if (b.spilled) {
    FILE* f = fopen(spill_path(c, key).c_str(), "rb");
    if (!f) return -1;
    size_t got = fread(out, 1, (size_t)b.size, f);
    fclose(f);
    c->loads++;
    return got == (size_t)b.size ? b.size : -1;
  }
  memcpy(out, b.data.data(), (size_t)b.size);
  b.tick = ++c->tick;
  return b.size;
}
int cache_stats(void* h, int64_t* resident, int64_t* spills, int64_t* loads, int64_t* blocks) {
  Cache* c = (Cache*)h; std::lock_guard<std::mutex> l(c->mu);
  *resident = c->resident; *spills = (int64_t)c->spills; *loads = (int64_t)c->loads; *blocks = (int64_t)c->blocks.size();
  return 0;
}

// ---- Rabin fingerprint -----------------------------------------------------------------------------------
uint64_t fprint64(uint64_t seed, const void* data, int64_t n) {
  if (!g_fp_init) fp_init();
  const unsigned char* p = (const unsigned char*)data;
  uint64_t fp = seed;
  for (int64_t i = 0; i < n; ++i) fp = (fp >> 8) ^ g_fp_table[(fp ^ p[i]) & 0xFF];
  return fp;
}

}  // extern "C"
