/* C ABI of libdtf_b200_host.so — host-side native I/O of the framework: crc32c, TFRecord /
 * TensorBoard event files, LevelDB-format tables and TF1 "tensor bundle" (Saver V2)
 * checkpoints. The reference gets all of these from inside the TensorFlow wheel
 * (pywrap_tensorflow.EventsWriter at logger.py:148-155; Saver V2 through
 * MonitoredTrainingSession(checkpoint_dir=...) at
 * templates/00_between_graph_replication_async_mnist.py:43-45). Functions return 0 on success
 * unless stated; dtf_io_last_error() describes the last failure of the calling thread. */
#ifndef DTF_B200_IO_H_
#define DTF_B200_IO_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

const char* dtf_io_last_error(void);

/* crc32c (Castagnoli), hardware accelerated when SSE4.2 is present. */
uint32_t dtf_crc32c(const void* data, size_t n);
uint32_t dtf_crc32c_extend(uint32_t crc, const void* data, size_t n);
uint32_t dtf_crc32c_mask(uint32_t crc);    /* ((crc >> 15) | (crc << 17)) + 0xa282ead8 */
uint32_t dtf_crc32c_unmask(uint32_t masked);
int dtf_crc32c_is_accelerated(void);
/* crc32c(A || B) from crc32c(A), crc32c(B) and len(B) (both unmasked). */
uint32_t dtf_crc32c_combine(uint32_t crc_a, uint32_t crc_b, uint64_t len_b);

/* TensorBoard event file writer: <prefix>.out.tfevents.<unix time>.<hostname>. The first record
 * is Event{file_version: "brain.Event:2"}. */
typedef struct dtf_events_writer dtf_events_writer;
dtf_events_writer* dtf_events_open(const char* path_prefix, const char* hostname /*or NULL*/);
const char* dtf_events_filename(dtf_events_writer* w);
int dtf_events_write_scalars(dtf_events_writer* w, double wall_time, int64_t step, int n,
                             const char* const* tags, const float* values);
int dtf_events_write_raw(dtf_events_writer* w, const void* event_proto, size_t n);
int dtf_events_flush(dtf_events_writer* w);
int dtf_events_close(dtf_events_writer* w);

/* TFRecord reader (inverse of the framing above; verifies both masked crcs). */
typedef struct dtf_tfrecord_reader dtf_tfrecord_reader;
dtf_tfrecord_reader* dtf_tfrecord_open(const char* path);
/* 1 = a record was returned (valid until the next call), 0 = clean EOF, <0 = error */
int dtf_tfrecord_next(dtf_tfrecord_reader* r, const void** data, size_t* n);
int dtf_tfrecord_close(dtf_tfrecord_reader* r);

/* TF1 tensor-bundle checkpoint. dtype uses TensorFlow's DataType enum
 * (DT_FLOAT = 1, DT_INT32 = 3, DT_INT64 = 9, DT_BFLOAT16 = 14). */
typedef struct dtf_bundle_writer dtf_bundle_writer;
dtf_bundle_writer* dtf_bundle_writer_open(const char* prefix);
int dtf_bundle_writer_add(dtf_bundle_writer* w, const char* name, int dtype, int ndim,
                          const int64_t* dims, const void* data, size_t nbytes);
int dtf_bundle_writer_finish(dtf_bundle_writer* w); /* writes <prefix>.index, frees w */
/* MergeV2Checkpoints: N single-shard bundles -> <merged_prefix>.index +
 * <merged_prefix>.data-0000i-of-0000N (data files are renamed, indices re-written). */
int dtf_bundle_merge(const char* const* part_prefixes, int n, const char* merged_prefix,
                     int delete_old_dirs);

/* Index-only writer for a checkpoint whose data shards are written in parallel by the PS tasks
 * themselves (each pwrite()s the bytes it owns into <prefix>.data-0000i-of-0000N at offsets the
 * chief planned): the chief adds one entry per tensor - shard, offset, size, UNMASKED crc32c of the
 * whole tensor (combined from the writers' per-slice checksums) - and writes <prefix>.index.
 * Reference: the sharded Scaffold saver behind
 * templates/00_between_graph_replication_async_mnist.py:43-45. */
typedef struct dtf_bundle_index_writer dtf_bundle_index_writer;
dtf_bundle_index_writer* dtf_bundle_index_open(const char* prefix, int num_shards);
int dtf_bundle_index_add(dtf_bundle_index_writer* w, const char* name, int dtype, int ndim,
                         const int64_t* dims, int shard_id, int64_t offset, int64_t size,
                         uint32_t crc32c);
int dtf_bundle_index_data_path(dtf_bundle_index_writer* w, int shard_id, char* buf, size_t cap);
int dtf_bundle_index_finish(dtf_bundle_index_writer* w); /* writes <prefix>.index, frees w */

typedef struct dtf_bundle_reader dtf_bundle_reader;
dtf_bundle_reader* dtf_bundle_reader_open(const char* prefix);
int dtf_bundle_reader_num_shards(dtf_bundle_reader* r);
int dtf_bundle_reader_num_entries(dtf_bundle_reader* r);
#define DTF_BUNDLE_MAX_DIMS 8
int dtf_bundle_reader_entry(dtf_bundle_reader* r, int index, char* name_buf, size_t name_cap,
                            int* dtype, int* ndim, int64_t* dims, int64_t* nbytes, int* shard_id);
int dtf_bundle_reader_read(dtf_bundle_reader* r, const char* name, void* dst, size_t nbytes,
                           int verify_crc);
int dtf_bundle_reader_close(dtf_bundle_reader* r);

/* LevelDB-format table, exposed for tests (golden bytes) and tooling. */
typedef struct dtf_table_builder dtf_table_builder;
dtf_table_builder* dtf_table_builder_open(const char* path, int block_size);
int dtf_table_builder_add(dtf_table_builder* b, const void* key, size_t klen, const void* value,
                          size_t vlen); /* keys must be added in strictly increasing order */
int dtf_table_builder_finish(dtf_table_builder* b);

#ifdef __cplusplus
}
#endif
#endif /* DTF_B200_IO_H_ */
