// Internal C++ interfaces of libdfm_io.so (see include/deepfm_io.h for the C ABI).
#pragma once

#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

#include "deepfm_io.h"

namespace dfio {

void set_error(const char* fmt, ...);

uint32_t crc32c(const void* data, size_t n);
inline uint32_t mask_crc(uint32_t c) { return ((c >> 15) | (c << 17)) + 0xa282ead8u; }

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
};

// verify_mapped / verify_stream: check both CRCs of every record for mapped files / streams.
std::unique_ptr<RecordSource> open_source(const std::string& path, bool verify_mapped, bool verify_stream,
                                          bool force_stream);

// ---------------------------------------------------------------------------
// Example codec
// ---------------------------------------------------------------------------
// Decodes label / ids / values from a serialized tf.train.Example.  Keys may
// come in any order; lists may be packed or unpacked.  Returns 0 or -1.
int parse_example(const uint8_t* p, uint64_t len, int field_size, void* ids_out, bool ids_i32,
                  float* vals_out, float* label_out);
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
