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
int dh_write_records(void* h, const void* data, const int64_t* offs, int64_t n);   /* n records back to back, offs[n + 1] */
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

/* ---- graph manager core: GraphManager/vertex/DrGraph.cpp, DrVertex.cpp, DrCohort.cpp,
 *      stagemanager/DrStageStatistics.cpp, DrDefaultManager.cpp, DrDynamic*.{h,cpp} -------------- */
void* gm_create(int max_failures, int speculative);
void gm_destroy(void* h);
void gm_set_outlier_params(void* h, double default_s, double min_s, double nonparam_fraction);
int gm_add_stage(void* h, const char* name, int nparts);
int gm_connect(void* h, int src_stage, int dst_stage, int kind);      /* 0 pointwise, 1 cross product */
/* channel 0 = file (output persists), 1 = FIFO / pipe: producer and consumer form a start clique
 * (DrClique): they become runnable together and restart together. */
int gm_connect_ex(void* h, int src_stage, int dst_stage, int kind, int channel);
/* Dynamic resize of a not-yet-started stage; follow != 0 propagates the new width to the pointwise
 * stages downstream of equal width (DrPipelineSplitManager).  Returns stages resized or < 0. */
int gm_resize_stage(void* h, int stage, int new_parts, int follow);
/* run-time rewrite: one new vertex in `stage` takes over the edges inputs -> consumers; returns its id */
int gm_splice(void* h, int stage, const int* inputs, int n_in, const int* consumers, int n_cons);
int gm_vertex_inputs(void* h, int vertex, int* out, int cap);
int gm_stage_vertices(void* h, int stage, int* out, int cap);
int gm_vertex_clique(void* h, int vertex, int* out, int cap);
int gm_live_vertex_count(void* h);
int gm_gang_restarts(void* h);
int gm_start(void* h);
int gm_next_ready(void* h, double now, int* vertex, int* version);
int gm_report_completed(void* h, int vertex, int version, double seconds, double bytes_in);
int gm_report_failed(void* h, int vertex, int version, int bad_input, const char* message);
int gm_check_duplicates(void* h, double now);
int gm_job_state(void* h);                                            /* 0 not started,1 running,3 ok,4 failed */
int gm_vertex_count(void* h);
int gm_vertex_info(void* h, int vertex, int* stage, int* part, int* state, int* completed_version, int* failures);
int gm_counters(void* h, int* completed, int* failures, int* duplicates, int* invalidations);
int gm_error(void* h, char* buf, int cap);
int gm_stage_model(void* h, int stage, double* startup, double* per_byte, double* sigma,
                   double* threshold_for_bytes, double bytes);
int dyn_aggregate_groups(int n, const int64_t* bytes, const int32_t* machine, const int32_t* pod,
                         int max_inputs, int64_t max_bytes, int64_t skip_bytes, int32_t* group_out);
int dyn_consumer_count(int64_t total_bytes, int64_t data_per_vertex, int min_count, int max_count);
int dyn_broadcast_tree(int consumers, int32_t* copy_of);
/* DrDynamicRangeDistributor: consumers from sample volume / sampling rate. */
int dyn_range_consumer_count(int64_t sample_bytes, double sampling_rate, int64_t data_per_vertex,
                             int min_count, int max_count);

/* ---- message pump: GraphManager/kernel/DrMessagePump.h ------------------------------------------ */
typedef void (*dh_pump_handler)(int64_t listener, int64_t message, void* user);
void* pump_create(int threads, dh_pump_handler handler, void* user);
void pump_post(void* h, int64_t listener, int64_t message, double delay_seconds);
int64_t pump_wait_delivered(void* h, int64_t count, double timeout_seconds);
/* CUDA stream completion -> pump message: pass pump_notifier_fn() and pump_make_notifier(...) to
 * cudaLaunchHostFunc (dryad_stream_host_func in libdryad_kernels).  The notifier frees itself. */
void* pump_make_notifier(void* h, int64_t listener, int64_t message);
void* pump_notifier_fn(void);
void pump_destroy(void* h);

/* ---- versioned mailbox: ProcessService/ProcessService.cs:674-746 ---------------------------------- */
void* mbox_create(void);
void mbox_destroy(void* h);
uint64_t mbox_set(void* h, const char* key, const char* short_status, const void* data, int64_t n);
uint64_t mbox_block_on_status(void* h, const char* key, uint64_t last_seen, double timeout_seconds,
                              char* status_out, int32_t status_cap, void* data_out, int64_t cap, int64_t* n_out);

/* ---- delay scheduler: LocalScheduler/LocalScheduler.cs, Queues.cs ---------------------------------- */
void* ls_create(int ncomputers, const int32_t* rack_of, double rack_delay, double cluster_delay);
void ls_destroy(void* h);
void ls_submit(void* h, int64_t proc_id, const int32_t* computers, int n, int hard);
int64_t ls_computer_idle(void* h, int computer, double age_bias);
int ls_waiting(void* h);

/* ---- FIFO channels, block cache + spill, Rabin fingerprint: channelfifo.cpp, ProcessService/Cache.cs,
 *      SpillMachine.cs, classlib/DrFPrint.cpp, LinqToDryad/Hash64.cs ------------------------------- */
void* fifo_create(int depth);
void fifo_destroy(void* h);
int fifo_write(void* h, const void* data, int64_t n, double timeout_seconds);
void fifo_close_writer(void* h);
void fifo_abort(void* h);
int64_t fifo_read(void* h, void* out, int64_t cap, double timeout_seconds);
int fifo_depth_now(void* h);
void* cache_create(const char* spill_dir, int64_t low_water, int64_t high_water);
void cache_destroy(void* h);
int cache_put(void* h, const char* key, const void* data, int64_t n);
int64_t cache_get(void* h, const char* key, void* out, int64_t cap);
int cache_stats(void* h, int64_t* resident, int64_t* spills, int64_t* loads, int64_t* blocks);
uint64_t fprint64(uint64_t seed, const void* data, int64_t n);

/* ---- vertex command / status blocks (tagged TLV) and affinity merging:
 *      DryadVertex/VertexHost/system/common/include/dvertexcommand.h:48-243, GraphManager/vertex/DrVertexCommand.cpp,
 *      GraphManager/gang/DrProperty.cpp, GraphManager/kernel/DrResources.cpp ------------------------------ */
void* vc_create(int command, int vertex_id, int version);
void vc_destroy(void* h);
void vc_add_arg(void* h, const char* a);
void vc_add_input(void* h, const char* uri, int64_t length);
void vc_add_output(void* h, const char* uri, int64_t size_hint);
void vc_set_status(void* h, int state, int exit_code, const char* message);
void vc_add_channel_status(void* h, int64_t bytes, int error);
int64_t vc_serialize(void* h, void* out, int64_t cap);
void* vc_parse(const void* data, int64_t n);
int vc_get_header(void* h, int* command, int* vertex_id, int* version, int* state, int* exit_code);
int vc_count(void* h, int what);
int64_t vc_get_item(void* h, int what, int i, char* buf, int cap);
int affinity_merge(int n, const int32_t* computer, const int64_t* weight, const int32_t* rack_of, int ncomputers,
                   double threshold, int32_t* choice);

/* ---- timed CPU baselines (multi-threaded restatements of the benchmark queries) ------------------ */
/* Native host vertex operators for traced string pipelines (csrc/host/vertex_native.cpp):
 * SelectMany(line.Split(delim)) -> GroupBy(word) with Count(), from raw text bytes (LinqToDryad/
 * DryadLinqVertex.cs:5718-6032, samples/WordCount.cs.pp:72-75).  The handle owns the group table. */
void* hv_split_count(const char* text, int64_t n, int delim, int keep_empty, int threads);
int64_t hv_count_byte(const char* text, int64_t n, int byte);
int64_t hv_groups(void* h);
int64_t hv_word_bytes(void* h);
int64_t hv_total(void* h);
void hv_export(void* h, char* words, int64_t* offsets, int64_t* counts);
void hv_free(void* h);

int64_t cpu_wordcount(const char* text, int64_t n, int threads, int keep_empty, int64_t* total_tokens);
int cpu_sort_rows(const uint8_t* rows, int64_t n, int width, int key_len, int threads, uint8_t* out);
int64_t cpu_groupby_sum_count(const int64_t* keys, const int64_t* vals, int64_t n, int threads, int64_t* checksum);

#ifdef __cplusplus
}
#endif
#endif
