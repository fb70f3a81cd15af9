/* dryad_b200 — C ABI of the C++ host runtime (libdryad_host.so).
 * Handles are opaque pointers; functions return 0 on success and a negative code on failure
 * unless stated otherwise.  Each group cites the reference code whose role it takes. */
#ifndef DRYAD_HOST_H
#define DRYAD_HOST_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* ---- binary record streams: LinqToDryad/DryadLinqBinaryWriter.cs, DryadLinqBinaryReader.cs,
 *      NativeBlockStream.cs; gzip: DryadVertex/.../GzipCompressionChannelTransform.cpp --------- */
void* dh_writer_open(const char* path, int compress, int64_t block_size);
void dh_write_bytes(void* h, const void* p, int64_t n);
void dh_write_u8(void* h, uint8_t v);
void dh_write_i32(void* h, int32_t v);
void dh_write_i64(void* h, int64_t v);
void dh_write_f64(void* h, double v);
void dh_write_compact(void* h, int32_t v);
void dh_write_string(void* h, const char* utf8, int32_t nbytes, int32_t nchars);
int dh_writer_end_record(void* h);
int dh_writer_write_rows(void* h, const void* rows, int64_t n, int32_t width);
int dh_writer_close(void* h, int64_t* raw_bytes, int64_t* stored_bytes, int64_t* records);
void* dh_reader_open(const char* path);
int dh_reader_at_end(void* h);
int dh_read_bytes(void* h, void* out, int64_t n);
int dh_read_u8(void* h, uint8_t* v);
int dh_read_i32(void* h, int32_t* v);
int dh_read_i64(void* h, int64_t* v);
int dh_read_f64(void* h, double* v);
int dh_read_compact(void* h, int32_t* v);
int dh_reader_peek(void* h, const uint8_t** p, int64_t* avail);
int dh_reader_skip(void* h, int64_t n);
int dh_reader_close(void* h);
int64_t dh_gzip_bound(int64_t n);
int64_t dh_gzip_compress(const void* src, int64_t n, void* dst, int64_t cap);
int64_t dh_gzip_decompress(const void* src, int64_t n, void* dst, int64_t cap);

/* ---- partfile: LinqToDryad/DataProvider.cs:400-537, GraphManager/filesystem/DrPartitionFile.cpp */
int dh_partfile_write(const char* pt_path, const char* base, int32_t nparts, const int64_t* sizes,
                      const char* machine);
int dh_partfile_read(const char* pt_path, char* base, int32_t cap, int64_t* sizes, int32_t max_parts);
void dh_partfile_part_path(const char* base, int32_t index, char* out, int32_t cap);
int dh_commit_part(const char* tmp_path, const char* final_path);

#ifdef __cplusplus
}
#endif
#endif
