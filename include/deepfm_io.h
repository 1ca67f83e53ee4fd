/*
 * deepfm_io.h -- C ABI of libdfm_io.so, the host-side data layer:
 * CRC32C, TFRecord framing (reader, writer, streaming reader for FIFOs),
 * tf.train.Example codec for the three DeepFM keys (label / ids / values),
 * a multi-threaded sharding batch loader that fills caller-owned (pinned)
 * buffers, the libsvm -> TFRecord converter and a Criteo-shaped generator.
 *
 * It replaces, natively, what the reference gets from TensorFlow's C++ runtime:
 *   tf.data.TFRecordDataset / PipeModeDataset      HVD.py:128,136
 *   tf.parse_example                               HVD.py:109-118
 *   shard / batch(drop_remainder) / repeat / prefetch  HVD.py:130-159
 *   tf.python_io.TFRecordWriter + tf.train.Example TOOL.py:29,42-57
 * (HVD.py = 2-hvd-gpu/DeepFM-hvd-tfrecord-vectorized-map.py,
 *  TOOL.py = tools/libsvm_to_tfrecord.py in the reference tree).
 *
 * All functions return >= 0 on success and a negative value on error;
 * dfio_last_error() returns the message for the calling thread.
 */
#ifndef DEEPFM_IO_H_
#define DEEPFM_IO_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

const char* dfio_last_error(void);

/* CRC-32C (Castagnoli) and TFRecord's masked form. */
uint32_t dfio_crc32c(const void* data, size_t n);
uint32_t dfio_masked_crc32c(const void* data, size_t n);

/* ---- raw record access ------------------------------------------------ */
/* path may be a regular file (mmap) or a FIFO / "-" for stdin (streamed). */
void* dfio_reader_open(const char* path, int verify_crc);
/* 1 = record returned (*data valid until the next call), 0 = end of stream. */
int dfio_reader_next(void* reader, const uint8_t** data, uint64_t* len);
void dfio_reader_close(void* reader);
int64_t dfio_count_records(const char* path, int verify_crc);

void* dfio_writer_open(const char* path);
int dfio_writer_write(void* writer, const void* data, uint64_t len);
int dfio_writer_write_example(void* writer, float label, const int64_t* ids, const float* vals, int n);
int dfio_writer_close(void* writer);

/* ---- Example codec ------------------------------------------------------ */
/* Encodes {label, ids, values}; returns bytes written (<= cap) or <0. */
int64_t dfio_encode_example(float label, const int64_t* ids, const float* vals, int n, uint8_t* out,
                            uint64_t cap, int packed);
/* Decodes one serialized Example with exactly field_size ids and values.
 * ids_out is int64 when ids_i32 == 0, else int32. */
int dfio_parse_example(const void* buf, uint64_t len, int field_size, void* ids_out, int ids_i32,
                       float* vals_out, float* label_out);

/* Reads a whole file (after record-level sharding index % shard_n == shard_i);
 * returns the number of records written (<= capacity). */
int64_t dfio_read_all(const char* path, int field_size, int shard_n, int shard_i, int64_t capacity,
                      void* ids_out, int ids_i32, float* vals_out, float* labels_out, int verify_crc);

/* ---- batch loader ------------------------------------------------------- */
typedef struct {
  int32_t field_size;
  int32_t batch_size;
  int32_t num_epochs;      /* >= 1 */
  int32_t drop_remainder;  /* 1: drop the tail of every epoch (reference behaviour) */
  int32_t shard_n;         /* record-level shard: keep records with index % shard_n == shard_i */
  int32_t shard_i;
  int32_t num_workers;     /* parser threads */
  int32_t num_slots;       /* ring depth == number of caller buffers */
  int32_t ids_i32;         /* 1: ids written as int32, else int64 */
  int32_t verify_crc;
  int32_t streaming;       /* 1: paths are FIFOs / stdin, read sequentially without mmap */
  int32_t feature_size;    /* > 0: every id must lie in [0, feature_size); a record that violates it fails the stream
                              with the file and byte offset (TF's embedding_lookup raises InvalidArgument, HVD.py:197) */
} dfio_loader_config;

/* slot_*[i] are the caller's buffers for ring slot i:
 *   ids    [batch_size * field_size] int64|int32
 *   vals   [batch_size * field_size] float
 *   labels [batch_size] float                                   */
void* dfio_loader_create(const dfio_loader_config* cfg, const char* const* paths, int n_paths,
                         void* const* slot_ids, void* const* slot_vals, void* const* slot_labels);
/* Blocks for the next batch in stream order.  1 = batch (slot, count set), 0 = end, <0 = error. */
int dfio_loader_next(void* loader, int* slot, int* count);
int dfio_loader_release(void* loader, int slot);
/* records parsed, batches produced, bytes read so far */
int dfio_loader_stats(void* loader, int64_t* records, int64_t* batches, int64_t* bytes);
void dfio_loader_destroy(void* loader);

/* ---- tools ---------------------------------------------------------------- */
/* "<label> <id>:<val> ..." text -> TFRecord; expect_fields > 0 enforces the pair count.
 * Returns the number of records written. */
int64_t dfio_libsvm_to_tfrecord(const char* in_path, const char* out_path, int expect_fields);
/* Criteo-shaped synthetic data: n_dense numeric fields with constant ids 1..n_dense and
 * real values, the remaining fields categorical (value 1.0) with Zipf(zipf_s)-distributed
 * ids inside per-field vocab ranges that partition [n_dense+1, feature_size). */
int64_t dfio_generate_synthetic(const char* out_path, int64_t n_records, int field_size,
                                int64_t feature_size, int n_dense, uint64_t seed, double pos_rate,
                                double zipf_s);

#ifdef __cplusplus
}
#endif
#endif /* DEEPFM_IO_H_ */
