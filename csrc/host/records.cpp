// Binary record wire format, block streams, gzip channel transform and partfile metadata.
//
// Reference behaviour (re-implemented, not copied):
//  * LinqToDryad/DryadLinqBinaryWriter.cs / DryadLinqBinaryReader.cs — block-buffered streams
//    (default block 256 KiB, :39); a record never straddles two blocks (FlushDataBlock :203-236;
//    an oversized record grows the block); little-endian fixed-width primitives (:336-480); compact
//    int = 1 byte below 0x80, else 4 bytes big-endian with the top bit set (WriteCompact :355-372,
//    ReadCompactInt32 reader :341); string = compact(char count) + compact(byte count) + UTF-8
//    (:523-546).
//  * LinqToDryad/NativeBlockStream.cs, DryadLinqBlockStream.cs — Read/Write/Allocate/Release block
//    API over files.
//  * DryadVertex/.../GzipCompressionChannelTransform.cpp:124-253 — zlib deflate/inflate of channel
//    blocks (used here for file outputs; intermediate compression defaults to None).
//  * LinqToDryad/DataProvider.cs:400-537 + GraphManager/filesystem/DrPartitionFile.cpp:393-402,
//    561-580 — the partfile text format and the rename-on-commit of successful part versions.
#include "../../include/dryad_host.h"

#include <errno.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <zlib.h>

#include <string>
#include <vector>

namespace {

constexpr size_t kDefaultBlock = 256 * 1024;
constexpr uint32_t kBlockMagic = 0x4B4C4244u;  // "DBLK": [magic][raw_len][stored_len][flags]

struct Writer {
  FILE* f = nullptr;
  std::vector<uint8_t> block;
  size_t block_size = kDefaultBlock;
  size_t record_start = 0;
  int compress = 0;  // 0 none, 1 gzip(deflate)
  uint64_t total_raw = 0, total_stored = 0, records = 0;
  int err = 0;
  std::string path;
};

struct Reader {
  FILE* f = nullptr;
  std::vector<uint8_t> block;
  size_t pos = 0;
  uint64_t total_raw = 0;
  int err = 0;
  bool eof = false;
};

int flush_block(Writer* w, size_t upto) {
  if (upto == 0) return 0;
  const uint8_t* src = w->block.data();
  std::vector<uint8_t> z;
  uint32_t hdr[4] = {kBlockMagic, (uint32_t)upto, (uint32_t)upto, 0};
  const uint8_t* out = src;
  size_t out_len = upto;
  if (w->compress) {
    uLongf zl = compressBound((uLong)upto);
    z.resize(zl);
    if (compress2(z.data(), &zl, src, (uLong)upto, Z_DEFAULT_COMPRESSION) != Z_OK) return w->err = -5;
    out = z.data(); out_len = zl;
    hdr[2] = (uint32_t)zl; hdr[3] = 1;
  }
  if (fwrite(hdr, 1, sizeof(hdr), w->f) != sizeof(hdr)) return w->err = -6;
  if (fwrite(out, 1, out_len, w->f) != out_len) return w->err = -6;
  w->total_raw += upto;
  w->total_stored += sizeof(hdr) + out_len;
  // keep the partial record (bytes after upto) at the front of the next block
  size_t rest = w->block.size() - upto;
  if (rest) memmove(w->block.data(), w->block.data() + upto, rest);
  w->block.resize(rest);
  w->record_start = 0;
  return 0;
}

inline void put(Writer* w, const void* p, size_t n) {
  const uint8_t* b = (const uint8_t*)p;
  w->block.insert(w->block.end(), b, b + n);
}

int fill_block(Reader* r) {
  uint32_t hdr[4];
  size_t got = fread(hdr, 1, sizeof(hdr), r->f);
  if (got == 0) { r->eof = true; return 1; }
  if (got != sizeof(hdr) || hdr[0] != kBlockMagic) return r->err = -7;
  std::vector<uint8_t> stored(hdr[2]);
  if (hdr[2] && fread(stored.data(), 1, hdr[2], r->f) != hdr[2]) return r->err = -7;
  if (hdr[3] & 1) {
    r->block.resize(hdr[1]);
    uLongf raw = hdr[1];
    if (uncompress(r->block.data(), &raw, stored.data(), hdr[2]) != Z_OK || raw != hdr[1]) return r->err = -8;
  } else {
    r->block.swap(stored);
  }
  r->pos = 0;
  r->total_raw += hdr[1];
  return 0;
}

inline int need(Reader* r, size_t n) {
  // Records never straddle blocks, so a primitive is always inside the current block.
  while (r->pos + n > r->block.size()) {
    if (r->pos < r->block.size()) return r->err = -9;  // torn record
    int rc = fill_block(r);
    if (rc) return rc < 0 ? rc : (r->err = -10);       // unexpected end of stream
  }
  return 0;
}

}  // namespace

extern "C" {

// ---- writer ------------------------------------------------------------------------------------
void* dh_writer_open(const char* path, int compress, int64_t block_size) {
  FILE* f = fopen(path, "wb");
  if (!f) return nullptr;
  Writer* w = new Writer();
  w->f = f; w->compress = compress; w->path = path;
  if (block_size > 0) w->block_size = (size_t)block_size;
  w->block.reserve(w->block_size + 4096);
  return w;
}

void dh_write_bytes(void* h, const void* p, int64_t n) { put((Writer*)h, p, (size_t)n); }
void dh_write_u8(void* h, uint8_t v) { put((Writer*)h, &v, 1); }
void dh_write_i32(void* h, int32_t v) { put((Writer*)h, &v, 4); }
void dh_write_i64(void* h, int64_t v) { put((Writer*)h, &v, 8); }
void dh_write_f64(void* h, double v) { put((Writer*)h, &v, 8); }

void dh_write_compact(void* h, int32_t v) {
  Writer* w = (Writer*)h;
  if (v >= 0 && v < 0x80) { uint8_t b = (uint8_t)v; put(w, &b, 1); return; }
  uint8_t b[4] = {(uint8_t)(((uint32_t)v >> 24) | 0x80), (uint8_t)(v >> 16), (uint8_t)(v >> 8), (uint8_t)v};
  put(w, b, 4);
}

void dh_write_string(void* h, const char* utf8, int32_t nbytes, int32_t nchars) {
  dh_write_compact(h, nchars);
  dh_write_compact(h, nbytes);
  put((Writer*)h, utf8, (size_t)nbytes);
}

// Marks the end of one record: if the block is over its budget, everything up to the START of
// this record is flushed, so the record stays whole in the next block.
int dh_writer_end_record(void* h) {
  Writer* w = (Writer*)h;
  w->records++;
  size_t sz = w->block.size();
  if (sz >= w->block_size) {
    size_t upto = w->record_start > 0 ? w->record_start : sz;  // oversized single record: own block
    int rc = flush_block(w, upto);
    if (rc) return rc;
  }
  w->record_start = w->block.size();
  return 0;
}

// Bulk path for variable-length records: n records back to back in `data`, record i = data[offs[i], offs[i+1]).
// Same effect as dh_write_bytes + dh_writer_end_record per record (records never straddle blocks).
int dh_write_records(void* h, const void* data, const int64_t* offs, int64_t n) {
  const char* p = (const char*)data;
  for (int64_t i = 0; i < n; ++i) {
    dh_write_bytes(h, p + offs[i], offs[i + 1] - offs[i]);
    int rc = dh_writer_end_record(h);
    if (rc) return rc;
  }
  return 0;
}

// Bulk path: n fixed-width records already laid out row-major.
int dh_writer_write_rows(void* h, const void* rows, int64_t n, int32_t width) {
  Writer* w = (Writer*)h;
  const uint8_t* p = (const uint8_t*)rows;
  for (int64_t i = 0; i < n; ++i) {
    put(w, p + (size_t)i * width, (size_t)width);
    int rc = dh_writer_end_record(h);
    if (rc) return rc;
  }
  return 0;
}

int dh_writer_close(void* h, int64_t* raw_bytes, int64_t* stored_bytes, int64_t* records) {
  Writer* w = (Writer*)h;
  int rc = w->err;
  if (!rc && !w->block.empty()) rc = flush_block(w, w->block.size());
  if (raw_bytes) *raw_bytes = (int64_t)w->total_raw;
  if (stored_bytes) *stored_bytes = (int64_t)w->total_stored;
  if (records) *records = (int64_t)w->records;
  if (fclose(w->f) != 0 && !rc) rc = -6;
  delete w;
  return rc;
}

// ---- reader ------------------------------------------------------------------------------------
void* dh_reader_open(const char* path) {
  FILE* f = fopen(path, "rb");
  if (!f) return nullptr;
  Reader* r = new Reader();
  r->f = f;
  return r;
}

// 1 = at end of stream, 0 = more data, <0 error
int dh_reader_at_end(void* h) {
  Reader* r = (Reader*)h;
  while (r->pos >= r->block.size()) {
    int rc = fill_block(r);
    if (rc == 1) return 1;
    if (rc < 0) return rc;
  }
  return 0;
}

int dh_read_bytes(void* h, void* out, int64_t n) {
  Reader* r = (Reader*)h;
  int rc = need(r, (size_t)n); if (rc) return rc;
  memcpy(out, r->block.data() + r->pos, (size_t)n); r->pos += (size_t)n; return 0;
}
int dh_read_u8(void* h, uint8_t* v) { return dh_read_bytes(h, v, 1); }
int dh_read_i32(void* h, int32_t* v) { return dh_read_bytes(h, v, 4); }
int dh_read_i64(void* h, int64_t* v) { return dh_read_bytes(h, v, 8); }
int dh_read_f64(void* h, double* v) { return dh_read_bytes(h, v, 8); }

int dh_read_compact(void* h, int32_t* v) {
  Reader* r = (Reader*)h;
  int rc = need(r, 1); if (rc) return rc;
  uint8_t b0 = r->block[r->pos];
  if (b0 < 0x80) { r->pos += 1; *v = b0; return 0; }
  rc = need(r, 4); if (rc) return rc;
  const uint8_t* b = r->block.data() + r->pos;
  *v = (int32_t)(((uint32_t)(b[0] & 0x7F) << 24) | ((uint32_t)b[1] << 16) | ((uint32_t)b[2] << 8) | b[3]);
  r->pos += 4;
  return 0;
}

// Current block access for bulk decoders: pointer + remaining bytes; advance with dh_reader_skip.
int dh_reader_peek(void* h, const uint8_t** p, int64_t* avail) {
  Reader* r = (Reader*)h;
  int e = dh_reader_at_end(h);
  if (e) { *p = nullptr; *avail = 0; return e < 0 ? e : 0; }
  *p = r->block.data() + r->pos; *avail = (int64_t)(r->block.size() - r->pos); return 0;
}
int dh_reader_skip(void* h, int64_t n) {
  Reader* r = (Reader*)h;
  if (r->pos + (size_t)n > r->block.size()) return -9;
  r->pos += (size_t)n; return 0;
}

int dh_reader_close(void* h) {
  Reader* r = (Reader*)h;
  int rc = r->err;
  fclose(r->f);
  delete r;
  return rc;
}

// ---- one-shot gzip transforms (channel transform role) --------------------------------------------
int64_t dh_gzip_bound(int64_t n) { return (int64_t)compressBound((uLong)n); }
int64_t dh_gzip_compress(const void* src, int64_t n, void* dst, int64_t cap) {
  uLongf zl = (uLongf)cap;
  if (compress2((Bytef*)dst, &zl, (const Bytef*)src, (uLong)n, Z_DEFAULT_COMPRESSION) != Z_OK) return -1;
  return (int64_t)zl;
}
int64_t dh_gzip_decompress(const void* src, int64_t n, void* dst, int64_t cap) {
  uLongf raw = (uLongf)cap;
  if (uncompress((Bytef*)dst, &raw, (const Bytef*)src, (uLong)n) != Z_OK) return -1;
  return (int64_t)raw;
}

// ---- partfile metadata -------------------------------------------------------------------------------
// Text: line 1 base path, line 2 partition count, then "index,size[,machine[:path]]" per
// partition; data files are <base>.%08X.
int dh_partfile_write(const char* pt_path, const char* base, int32_t nparts, const int64_t* sizes,
                      const char* machine) {
  std::string tmp = std::string(pt_path) + ".tmp";
  FILE* f = fopen(tmp.c_str(), "w");
  if (!f) return -errno;
  fprintf(f, "%s\n%d\n", base, nparts);
  for (int i = 0; i < nparts; ++i) {
    if (machine && machine[0]) fprintf(f, "%d,%lld,%s\n", i, (long long)sizes[i], machine);
    else fprintf(f, "%d,%lld\n", i, (long long)sizes[i]);
  }
  if (fclose(f) != 0) return -errno;
  if (rename(tmp.c_str(), pt_path) != 0) return -errno;   // atomic publish of the table
  return 0;
}

// Returns the partition count (>=0) or <0; fills base (cap bytes) and up to max sizes.
int dh_partfile_read(const char* pt_path, char* base, int32_t cap, int64_t* sizes, int32_t max_parts) {
  FILE* f = fopen(pt_path, "r");
  if (!f) return -errno;
  char line[4096];
  if (!fgets(line, sizeof line, f)) { fclose(f); return -1; }
  size_t l = strlen(line);
  while (l && (line[l - 1] == '\n' || line[l - 1] == '\r')) line[--l] = 0;
  if ((int32_t)l + 1 > cap) { fclose(f); return -2; }
  memcpy(base, line, l + 1);
  if (!fgets(line, sizeof line, f)) { fclose(f); return -1; }
  int n = atoi(line);
  for (int i = 0; i < n; ++i) {
    if (!fgets(line, sizeof line, f)) { fclose(f); return -3; }
    int idx = 0; long long sz = 0;
    if (sscanf(line, "%d,%lld", &idx, &sz) != 2 || idx != i) { fclose(f); return -4; }
    if (i < max_parts) sizes[i] = sz;
  }
  fclose(f);
  return n;
}

void dh_partfile_part_path(const char* base, int32_t index, char* out, int32_t cap) {
  snprintf(out, (size_t)cap, "%s.%08X", base, (unsigned)index);
}

// Commit: rename "<part>.v<version>.tmp" (a successful attempt) to the final part path.
int dh_commit_part(const char* tmp_path, const char* final_path) {
  return rename(tmp_path, final_path) == 0 ? 0 : -errno;
}

}  // extern "C"
