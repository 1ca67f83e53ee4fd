// Internal C++ interfaces of libdfm_io.so (see include/deepfm_io.h for the C ABI).
#pragma once

#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <nmmintrin.h>
#include <memory>
#include <string>
#include <vector>

#include "deepfm_io.h"

namespace dfio {

void set_error(const char* fmt, ...);

uint32_t crc32c(const void* data, size_t n);
inline uint32_t mask_crc(uint32_t c) { return ((c >> 15) | (c << 17)) + 0xa282ead8u; }
// CRC32C of the eight bytes of a record's length field (little endian): one instruction, no call
inline uint32_t crc32c_u64(uint64_t v) { return static_cast<uint32_t>(_mm_crc32_u64(0xffffffffu, v)) ^ 0xffffffffu; }

// ---------------------------------------------------------------------------
// Record sources.  next() yields a pointer valid until the following next()
// (streaming) or for the lifetime of the source (mapped file).
// ---------------------------------------------------------------------------
class RecordSource {
 public:
  virtual ~RecordSource() = default;
  // 1 = record, 0 = clean end, -1 = error (message set)
  virtual int next(const uint8_t** data, uint64_t* len) = 0;
  virtual bool stable_pointers() const = 0;
  virtual uint64_t bytes_consumed() const = 0;
  // mapped files only: the whole file as one byte range (lets several threads walk the framing of disjoint ranges)
  virtual bool mapping(const uint8_t** base, size_t* size) const { return false; }
  // streams only: the descriptor, for a reader that takes the bytes in large blocks itself (the batch loader does;
  // it must then never call next()).  -1 for everything else.
  virtual int raw_fd() const { return -1; }
};

// Walks the TFRecord framing of base[pos, limit_start) (records that START before limit_start): appends
// {payload offset, payload length}; returns the position after the last record walked, or (size_t)-1 on a framing
// error (message set).  verify_len_crc: check the CRC of every length field.
struct RecSpan { uint64_t off; uint32_t len; };
size_t walk_framing(const uint8_t* base, size_t size, size_t pos, size_t limit_start, bool verify_len_crc,
                    std::vector<RecSpan>* out, const char* path);
// First position >= from (and < to) that looks like a record start: its length CRC checks out and so does the next
// record's (or the file ends there).  (size_t)-1 if none.  A guess: the caller confirms it against the true chain.
size_t find_record_start(const uint8_t* base, size_t size, size_t from, size_t to);

// verify_mapped / verify_stream: check both CRCs of every record for mapped files / streams.
std::unique_ptr<RecordSource> open_source(const std::string& path, bool verify_mapped, bool verify_stream,
                                          bool force_stream);

// ---------------------------------------------------------------------------
// Example codec
// ---------------------------------------------------------------------------
// Decodes label / ids / values from a serialized tf.train.Example.  Keys may
// come in any order; lists may be packed or unpacked.  Returns 0 or -1.
// max_id > 0: ids must lie in [0, max_id) (what TF's embedding_lookup enforces on CPU); a violation is an error.
int parse_example(const uint8_t* p, uint64_t len, int field_size, void* ids_out, bool ids_i32,
                  float* vals_out, float* label_out, uint64_t max_id = 0);
// Appends the serialized Example to `out`.
void encode_example(float label, const int64_t* ids, const float* vals, int n, bool packed,
                    std::vector<uint8_t>& out);

class RecordWriter {
 public:
  explicit RecordWriter(FILE* f) : f_(f) {}
  ~RecordWriter();
  bool write(const void* data, uint64_t len);
  bool close();
 private:
  FILE* f_;
};

}  // namespace dfio
