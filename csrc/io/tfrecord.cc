// CRC32C, TFRecord framing (mapped-file and streaming sources, writer) and the
// tf.train.Example codec specialised to the DeepFM schema.
//
// Record framing: u64le length | u32le masked_crc32c(length) | payload |
// u32le masked_crc32c(payload); mask(c) = rotr(c,15) + 0xa282ead8.
// Payload: Example{features(1): Features{feature(1): map<string, Feature>}} with
// Feature = oneof{bytes_list(1), float_list(2), int64_list(3)}.
#include <errno.h>
#include <fcntl.h>
#include <nmmintrin.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include "io_internal.h"

namespace dfio {

static thread_local char g_err[1024] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

// ------------------------------- CRC32C ------------------------------------
// The crc32 instruction has a latency of 3 cycles and a throughput of one per cycle, so one dependent chain runs at
// a third of what the unit can do.  Three chains over three adjacent lanes of L bytes fill it; the lanes are joined
// with crc(A|B) = shift_L(crc(A)) ^ crc(B), where shift_L -- "append L zero bytes" -- is linear in the 32 state
// bits and is tabulated per byte of the state (4 x 256 words for every supported L, built once with the
// instruction itself).  A Criteo record (~300 bytes) takes one pass of 3 x 96 bytes plus a short tail.
namespace {

constexpr int kLaneSizes[] = {1024, 256, 96, 32};
constexpr int kNumLanes = sizeof(kLaneSizes) / sizeof(kLaneSizes[0]);

struct CrcShiftTables {
  uint32_t t[kNumLanes][4][256];
  CrcShiftTables() {
    for (int l = 0; l < kNumLanes; ++l)
      for (int k = 0; k < 4; ++k)
        for (uint32_t b = 0; b < 256; ++b) {
          uint64_t s = uint64_t(b) << (8 * k);
          for (int i = 0; i < kLaneSizes[l] / 8; ++i) s = _mm_crc32_u64(s, 0);
          t[l][k][b] = static_cast<uint32_t>(s);
        }
  }
};

inline uint32_t crc_shift(const uint32_t (&t)[4][256], uint32_t c) {
  return t[0][c & 0xff] ^ t[1][(c >> 8) & 0xff] ^ t[2][(c >> 16) & 0xff] ^ t[3][c >> 24];
}

inline uint64_t load64(const uint8_t* p) {
  uint64_t v;
  memcpy(&v, p, 8);
  return v;
}

}  // namespace

uint32_t crc32c(const void* data, size_t n) {
  static const CrcShiftTables tables;
  const uint8_t* p = static_cast<const uint8_t*>(data);
  uint64_t c = 0xffffffffu;
  for (int l = 0; l < kNumLanes; ++l) {
    const size_t L = kLaneSizes[l];
    while (n >= 3 * L) {
      uint64_t c1 = 0, c2 = 0;
      for (size_t i = 0; i < L; i += 8) {
        c = _mm_crc32_u64(c, load64(p + i));
        c1 = _mm_crc32_u64(c1, load64(p + L + i));
        c2 = _mm_crc32_u64(c2, load64(p + 2 * L + i));
      }
      c = crc_shift(tables.t[l], static_cast<uint32_t>(c)) ^ c1;
      c = crc_shift(tables.t[l], static_cast<uint32_t>(c)) ^ c2;
      p += 3 * L;
      n -= 3 * L;
    }
  }
  while (n >= 8) {
    uint64_t v;
    memcpy(&v, p, 8);
    c = _mm_crc32_u64(c, v);
    p += 8;
    n -= 8;
  }
  while (n) {
    c = _mm_crc32_u8(static_cast<uint32_t>(c), *p++);
    --n;
  }
  return static_cast<uint32_t>(c) ^ 0xffffffffu;
}

// ------------------------------ sources -------------------------------------
namespace {

class MappedSource final : public RecordSource {
 public:
  MappedSource(const uint8_t* base, size_t size, bool verify, std::string path)
      : base_(base), size_(size), verify_(verify), path_(std::move(path)) {}
  ~MappedSource() override {
    if (base_ && size_) munmap(const_cast<uint8_t*>(base_), size_);
  }
  int next(const uint8_t** data, uint64_t* len) override {
    if (pos_ == size_) return 0;
    if (size_ - pos_ < 12) {
      set_error("%s: truncated record header at offset %zu", path_.c_str(), pos_);
      return -1;
    }
    uint64_t n;
    uint32_t lcrc;
    memcpy(&n, base_ + pos_, 8);
    memcpy(&lcrc, base_ + pos_ + 8, 4);
    if (verify_ && mask_crc(crc32c(base_ + pos_, 8)) != lcrc) {
      set_error("%s: corrupt length CRC at offset %zu", path_.c_str(), pos_);
      return -1;
    }
    if (n > size_ - pos_ - 12 || size_ - pos_ - 12 - n < 4) {
      set_error("%s: truncated record body at offset %zu (len %llu)", path_.c_str(), pos_,
                (unsigned long long)n);
      return -1;
    }
    const uint8_t* body = base_ + pos_ + 12;
    if (verify_) {
      uint32_t dcrc;
      memcpy(&dcrc, body + n, 4);
      if (mask_crc(crc32c(body, n)) != dcrc) {
        set_error("%s: corrupt data CRC at offset %zu", path_.c_str(), pos_);
        return -1;
      }
    }
    *data = body;
    *len = n;
    pos_ += 12 + n + 4;
    return 1;
  }
  bool stable_pointers() const override { return true; }
  uint64_t bytes_consumed() const override { return pos_; }
  bool mapping(const uint8_t** base, size_t* size) const override {
    *base = base_;
    *size = size_;
    return true;
  }

 private:
  const uint8_t* base_;
  size_t size_;
  size_t pos_ = 0;
  bool verify_;
  std::string path_;
};

}  // namespace

static inline bool header_ok(const uint8_t* base, size_t size, size_t pos, uint64_t* n) {
  if (size - pos < 16) return false;
  uint32_t lcrc;
  memcpy(n, base + pos, 8);
  memcpy(&lcrc, base + pos + 8, 4);
  if (*n > size - pos - 16) return false;
  return mask_crc(crc32c_u64(*n)) == lcrc;
}

size_t find_record_start(const uint8_t* base, size_t size, size_t from, size_t to) {
  for (size_t pos = from; pos < to && pos + 16 <= size; ++pos) {
    uint64_t n, n2;
    if (!header_ok(base, size, pos, &n)) continue;
    const size_t nxt = pos + 16 + n;
    if (nxt == size || header_ok(base, size, nxt, &n2)) return pos;
  }
  return static_cast<size_t>(-1);
}

size_t walk_framing(const uint8_t* base, size_t size, size_t pos, size_t limit_start, bool verify_len_crc,
                    std::vector<RecSpan>* out, const char* path) {
  while (pos < limit_start && pos < size) {
    if (size - pos < 12) {
      set_error("%s: truncated record header at offset %zu", path, pos);
      return static_cast<size_t>(-1);
    }
    uint64_t n;
    memcpy(&n, base + pos, 8);
    if (verify_len_crc) {
      uint32_t lcrc;
      memcpy(&lcrc, base + pos + 8, 4);
      if (mask_crc(crc32c_u64(n)) != lcrc) {
        set_error("%s: corrupt length CRC at offset %zu", path, pos);
        return static_cast<size_t>(-1);
      }
    }
    if (n > size - pos - 12 || size - pos - 12 - n < 4) {
      set_error("%s: truncated record body at offset %zu (len %llu)", path, pos, (unsigned long long)n);
      return static_cast<size_t>(-1);
    }
    out->push_back(RecSpan{pos + 12, static_cast<uint32_t>(n)});
    // The walk is a pointer chase -- the place of the next header is known only once this one has arrived -- and it
    // touches one cache line in five, which no hardware prefetcher follows: left alone it pays one DRAM latency per
    // record.  Records of one dataset are nearly the same size, so the header kAhead records on lies within a line
    // or two of pos + kAhead * (this record's size): ask for those lines now.
    constexpr size_t kAhead = 12;
    const size_t guess = pos + kAhead * (n + 16);
    if (guess + 64 <= size) {
      __builtin_prefetch(base + guess - 32);
      __builtin_prefetch(base + guess + 32);
    }
    pos += 12 + n + 4;
  }
  return pos;
}

namespace {

// Sequential reader for FIFOs / stdin / anything that cannot be mapped
// (the SageMaker "pipe mode" transport, HVD.py:136).
class StreamSource final : public RecordSource {
 public:
  StreamSource(int fd, bool verify, std::string path, bool owns)
      : fd_(fd), verify_(verify), path_(std::move(path)), owns_(owns), buf_(1 << 20) {}
  ~StreamSource() override {
    if (owns_ && fd_ >= 0) close(fd_);
  }
  int next(const uint8_t** data, uint64_t* len) override {
    uint8_t hdr[12];
    int r = read_exact(hdr, 12, /*eof_ok=*/true);
    if (r <= 0) return r;
    uint64_t n;
    uint32_t lcrc;
    memcpy(&n, hdr, 8);
    memcpy(&lcrc, hdr + 8, 4);
    if (verify_ && mask_crc(crc32c(hdr, 8)) != lcrc) {
      set_error("%s: corrupt length CRC at stream offset %llu", path_.c_str(),
                (unsigned long long)consumed_);
      return -1;
    }
    if (n > (1ull << 31)) {
      set_error("%s: implausible record length %llu", path_.c_str(), (unsigned long long)n);
      return -1;
    }
    if (rec_.size() < n + 4) rec_.resize(n + 4);
    if (read_exact(rec_.data(), n + 4, false) <= 0) {
      set_error("%s: truncated record body", path_.c_str());
      return -1;
    }
    if (verify_) {
      uint32_t dcrc;
      memcpy(&dcrc, rec_.data() + n, 4);
      if (mask_crc(crc32c(rec_.data(), n)) != dcrc) {
        set_error("%s: corrupt data CRC at stream offset %llu", path_.c_str(),
                  (unsigned long long)consumed_);
        return -1;
      }
    }
    *data = rec_.data();
    *len = n;
    return 1;
  }
  bool stable_pointers() const override { return false; }
  uint64_t bytes_consumed() const override { return consumed_; }
  int raw_fd() const override { return lo_ == hi_ && consumed_ == 0 ? fd_ : -1; }

 private:
  // 1 = ok, 0 = clean EOF before any byte (only if eof_ok), -1 = error
  int read_exact(uint8_t* dst, size_t n, bool eof_ok) {
    size_t got = 0;
    while (got < n) {
      if (lo_ == hi_) {
        ssize_t r = ::read(fd_, buf_.data(), buf_.size());
        if (r < 0) {
          if (errno == EINTR) continue;
          set_error("%s: read failed: %s", path_.c_str(), strerror(errno));
          return -1;
        }
        if (r == 0) {
          if (got == 0 && eof_ok) return 0;
          set_error("%s: unexpected end of stream", path_.c_str());
          return -1;
        }
        lo_ = 0;
        hi_ = static_cast<size_t>(r);
      }
      size_t take = std::min(n - got, hi_ - lo_);
      memcpy(dst + got, buf_.data() + lo_, take);
      lo_ += take;
      got += take;
    }
    consumed_ += n;
    return 1;
  }
  int fd_;
  bool verify_;
  std::string path_;
  bool owns_;
  std::vector<uint8_t> buf_;
  size_t lo_ = 0, hi_ = 0;
  std::vector<uint8_t> rec_;
  uint64_t consumed_ = 0;
};

}  // namespace

std::unique_ptr<RecordSource> open_source(const std::string& path, bool verify_mapped, bool verify_stream,
                                          bool force_stream) {
  if (path == "-") return std::make_unique<StreamSource>(0, verify_stream, "<stdin>", false);
  int fd = ::open(path.c_str(), O_RDONLY);
  if (fd < 0) {
    set_error("cannot open %s: %s", path.c_str(), strerror(errno));
    return nullptr;
  }
  struct stat st;
  if (fstat(fd, &st) != 0) {
    set_error("cannot stat %s: %s", path.c_str(), strerror(errno));
    close(fd);
    return nullptr;
  }
  if (force_stream || !S_ISREG(st.st_mode)) return std::make_unique<StreamSource>(fd, verify_stream, path, true);
  if (st.st_size == 0) {
    close(fd);
    return std::make_unique<MappedSource>(nullptr, 0, verify_mapped, path);
  }
  void* m = mmap(nullptr, st.st_size, PROT_READ, MAP_PRIVATE, fd, 0);
  close(fd);
  if (m == MAP_FAILED) {
    set_error("cannot mmap %s: %s", path.c_str(), strerror(errno));
    return nullptr;
  }
  madvise(m, st.st_size, MADV_SEQUENTIAL);
  return std::make_unique<MappedSource>(static_cast<const uint8_t*>(m), st.st_size, verify_mapped, path);
}

// ------------------------------- writer -------------------------------------
RecordWriter::~RecordWriter() { close(); }

bool RecordWriter::write(const void* data, uint64_t len) {
  uint8_t hdr[12];
  memcpy(hdr, &len, 8);
  uint32_t c = mask_crc(crc32c(hdr, 8));
  memcpy(hdr + 8, &c, 4);
  uint32_t d = mask_crc(crc32c(data, len));
  if (fwrite(hdr, 1, 12, f_) != 12 || fwrite(data, 1, len, f_) != len || fwrite(&d, 1, 4, f_) != 4) {
    set_error("record write failed: %s", strerror(errno));
    return false;
  }
  return true;
}

bool RecordWriter::close() {
  if (!f_) return true;
  bool ok = fclose(f_) == 0;
  f_ = nullptr;
  if (!ok) set_error("close failed: %s", strerror(errno));
  return ok;
}

// --------------------------- protobuf helpers --------------------------------
namespace {

inline bool get_varint(const uint8_t*& p, const uint8_t* end, uint64_t& v) {
  uint64_t r = 0;
  int shift = 0;
  while (p < end && shift < 70) {
    uint8_t b = *p++;
    r |= static_cast<uint64_t>(b & 0x7f) << shift;
    if (!(b & 0x80)) {
      v = r;
      return true;
    }
    shift += 7;
  }
  return false;
}

inline bool get_len(const uint8_t*& p, const uint8_t* end, const uint8_t*& sub_end) {
  uint64_t n;
  if (!get_varint(p, end, n) || n > static_cast<uint64_t>(end - p)) return false;
  sub_end = p + n;
  return true;
}

inline bool skip_field(const uint8_t*& p, const uint8_t* end, uint32_t wire) {
  uint64_t v;
  switch (wire) {
    case 0: return get_varint(p, end, v);
    case 1: if (end - p < 8) return false; p += 8; return true;
    case 2: { const uint8_t* e; if (!get_len(p, end, e)) return false; p = e; return true; }
    case 5: if (end - p < 4) return false; p += 4; return true;
    default: return false;
  }
}

inline void put_varint(std::vector<uint8_t>& o, uint64_t v) {
  while (v >= 0x80) {
    o.push_back(static_cast<uint8_t>(v) | 0x80);
    v >>= 7;
  }
  o.push_back(static_cast<uint8_t>(v));
}

inline size_t varint_size(uint64_t v) {
  size_t n = 1;
  while (v >= 0x80) { v >>= 7; ++n; }
  return n;
}

enum Key { kLabel = 0, kIds = 1, kValues = 2, kOther = 3 };

inline Key classify(const uint8_t* k, size_t n) {
  if (n == 5 && memcmp(k, "label", 5) == 0) return kLabel;
  if (n == 3 && memcmp(k, "ids", 3) == 0) return kIds;
  if (n == 6 && memcmp(k, "values", 6) == 0) return kValues;
  return kOther;
}

// FloatList body -> out[0..cap); returns count or -1
int read_float_list(const uint8_t* p, const uint8_t* end, float* out, int cap) {
  int n = 0;
  while (p < end) {
    uint64_t tag;
    if (!get_varint(p, end, tag)) return -1;
    if (tag == 0x0A) {  // packed
      const uint8_t* e;
      if (!get_len(p, end, e) || ((e - p) & 3)) return -1;
      int cnt = static_cast<int>((e - p) / 4);
      if (n + cnt > cap) return -1;
      memcpy(out + n, p, size_t(cnt) * 4);
      n += cnt;
      p = e;
    } else if (tag == 0x0D) {  // unpacked fixed32
      if (end - p < 4 || n + 1 > cap) return -1;
      memcpy(out + n, p, 4);
      ++n;
      p += 4;
    } else if (!skip_field(p, end, tag & 7)) {
      return -1;
    }
  }
  return n;
}

// max_id > 0: every id must lie in [0, max_id) (checked on the 64-bit value, before any narrowing); a violation
// returns -2 with the offending value in *bad.
template <typename T>
int read_int64_list(const uint8_t* p, const uint8_t* end, T* out, int cap, uint64_t max_id, uint64_t* bad) {
  int n = 0;
  while (p < end) {
    uint64_t tag;
    if (!get_varint(p, end, tag)) return -1;
    if (tag == 0x0A) {  // packed varints
      const uint8_t* e;
      if (!get_len(p, end, e)) return -1;
      while (p < e) {
        uint64_t v;
        // fast paths: 1-, 2- and 3-byte varints cover every id < 2^21 (Criteo-sized vocabularies); the rest go the
        // general way.  The three-byte read stays inside the list (p + 3 <= e) or falls through.
        const uint32_t b0 = p[0];
        if (b0 < 0x80) {
          v = b0;
          p += 1;
        } else if (e - p >= 3) {
          const uint32_t b1 = p[1], b2 = p[2];
          if (b1 < 0x80) {
            v = (b0 & 0x7f) | (b1 << 7);
            p += 2;
          } else if (b2 < 0x80) {
            v = (b0 & 0x7f) | ((b1 & 0x7f) << 7) | (b2 << 14);
            p += 3;
          } else if (!get_varint(p, e, v)) {
            return -1;
          }
        } else if (!get_varint(p, e, v)) {
          return -1;
        }
        if (n >= cap) return -1;
        if (max_id && v >= max_id) { *bad = v; return -2; }
        out[n++] = static_cast<T>(v);
      }
    } else if (tag == 0x08) {  // unpacked
      uint64_t v;
      if (!get_varint(p, end, v) || n >= cap) return -1;
      if (max_id && v >= max_id) { *bad = v; return -2; }
      out[n++] = static_cast<T>(v);
    } else if (!skip_field(p, end, tag & 7)) {
      return -1;
    }
  }
  return n;
}

// length prefix of at most two bytes (< 16384): all a record of the DeepFM schema ever needs
inline bool get_len2(const uint8_t*& p, const uint8_t* end, uint32_t& n) {
  if (end - p < 2) return false;
  const uint32_t b0 = p[0], b1 = p[1];
  if (b0 < 0x80) {
    n = b0;
    p += 1;
    return true;
  }
  if (b1 >= 0x80) return false;
  n = (b0 & 0x7f) | (b1 << 7);
  p += 2;
  return true;
}

// The layout every stock writer produces (TFRecordWriter + tf.train.Example, TOOL.py:42-57, and encode_example
// below): one Features message filling the record, one map entry per key with `key` before `value`, one packed list
// per feature, nothing else.  Decoded by expectation instead of by a generic field walk: no skip logic, no 10-byte
// varints, at most three bytes per id.  Returns true when the record had exactly that shape and was decoded; false
// means "not handled" (another layout, an id of four or more bytes, or any error) and the caller runs the generic
// walker, which decodes it or names what is wrong with it.  Entries may come in any order.
template <typename T>
bool parse_canonical(const uint8_t* p, uint64_t len, int F, T* ids_out, float* vals_out, float* label_out,
                     uint64_t max_id) {
  const uint8_t* const end = p + len;
  uint32_t n;
  if (len < 2 || *p++ != 0x0A || !get_len2(p, end, n) || p + n != end) return false;
  unsigned have = 0;
  while (p < end) {
    if (*p++ != 0x0A || !get_len2(p, end, n) || n > static_cast<uint32_t>(end - p) || n < 8) return false;
    const uint8_t* const ee = p + n;                 // map entry: at least 0A kl k.. 12 l T l 0A l
    if (p[0] != 0x0A) return false;
    const uint32_t kl = p[1];
    const uint8_t* const k = p + 2;
    if (kl > 6 || k + kl + 6 > ee) return false;
    p = k + kl;
    if (*p++ != 0x12 || !get_len2(p, ee, n) || p + n != ee) return false;      // value: Feature
    if (ee - p < 4) return false;
    const uint32_t kind = *p++;                                                // 0x12 float_list, 0x1A int64_list
    if (!get_len2(p, ee, n) || p + n != ee) return false;
    if (ee - p < 2 || *p++ != 0x0A || !get_len2(p, ee, n) || p + n != ee) return false;   // one packed run
    if (kl == 3 && kind == 0x1A && k[0] == 'i' && k[1] == 'd' && k[2] == 's') {
      if (have & 1u) return false;
      int c = 0;
      // ids < 2^21 always pass a bound of 2^31 or more, so the vector compare below can be a signed 32-bit one
      const __m128i vmax = _mm_set1_epi32(max_id && max_id < 0x80000000ull ? static_cast<int>(max_id) : 0x7fffffff);
      while (p < ee) {
        if (end - p >= 16 && c + 4 <= F) {      // the load may run past this list, never past the record
          // Two shapes cover a Criteo record almost entirely: a run of one-byte ids (the 13 numeric fields) and a
          // run of three-byte ids (categorical fields of a vocabulary above 16383).  The sign bits of the next 16
          // bytes tell which one starts here; four ids are decoded per step.
          const __m128i raw = _mm_loadu_si128(reinterpret_cast<const __m128i*>(p));
          const unsigned cont = static_cast<unsigned>(_mm_movemask_epi8(raw));
          __m128i v4;
          int used = 0;
          if ((cont & 0xfffu) == 0x6dbu && ee - p >= 12) {          // 110 110 110 110, lowest bit = first byte
            const __m128i g = _mm_shuffle_epi8(raw, _mm_setr_epi8(0, 1, 2, -1, 3, 4, 5, -1, 6, 7, 8, -1, 9, 10, 11, -1));
            v4 = _mm_or_si128(_mm_or_si128(_mm_and_si128(g, _mm_set1_epi32(0x7f)),
                                           _mm_srli_epi32(_mm_and_si128(g, _mm_set1_epi32(0x7f00)), 1)),
                              _mm_srli_epi32(_mm_and_si128(g, _mm_set1_epi32(0x7f0000)), 2));
            used = 12;
          } else if ((cont & 0xfu) == 0u && ee - p >= 4) {
            v4 = _mm_cvtepu8_epi32(raw);
            used = 4;
          }
          if (used) {
            if (_mm_movemask_epi8(_mm_cmplt_epi32(v4, vmax)) != 0xffff) return false;
            if (sizeof(T) == 4) {
              _mm_storeu_si128(reinterpret_cast<__m128i*>(ids_out + c), v4);
            } else {
              _mm_storeu_si128(reinterpret_cast<__m128i*>(ids_out + c), _mm_cvtepu32_epi64(v4));
              _mm_storeu_si128(reinterpret_cast<__m128i*>(ids_out + c + 2), _mm_cvtepu32_epi64(_mm_srli_si128(v4, 8)));
            }
            c += 4;
            p += used;
            continue;
          }
        }
        uint32_t v;
        const uint32_t b0 = p[0];
        if (b0 < 0x80) {
          v = b0;
          p += 1;
        } else {
          if (ee - p < 2) return false;
          const uint32_t b1 = p[1];
          if (b1 < 0x80) {
            v = (b0 & 0x7f) | (b1 << 7);
            p += 2;
          } else {
            if (ee - p < 3 || p[2] >= 0x80) return false;
            v = (b0 & 0x7f) | ((b1 & 0x7f) << 7) | (uint32_t(p[2]) << 14);
            p += 3;
          }
        }
        if (c >= F || (max_id && v >= max_id)) return false;
        ids_out[c++] = static_cast<T>(v);
      }
      if (c != F) return false;
      have |= 1u;
    } else if (kl == 6 && kind == 0x12 && memcmp(k, "values", 6) == 0) {
      if ((have & 2u) || n != static_cast<uint32_t>(F) * 4u) return false;
      memcpy(vals_out, p, n);
      have |= 2u;
    } else if (kl == 5 && kind == 0x12 && memcmp(k, "label", 5) == 0) {
      if ((have & 4u) || n != 4) return false;
      memcpy(label_out, p, 4);
      have |= 4u;
    } else {
      return false;
    }
    p = ee;
  }
  return have == 7u;
}

}  // namespace

int parse_example(const uint8_t* p, uint64_t len, int F, void* ids_out, bool ids_i32, float* vals_out,
                  float* label_out, uint64_t max_id) {
  if (ids_i32 ? parse_canonical<int32_t>(p, len, F, static_cast<int32_t*>(ids_out), vals_out, label_out, max_id)
              : parse_canonical<int64_t>(p, len, F, static_cast<int64_t*>(ids_out), vals_out, label_out, max_id))
    return 0;
  const uint8_t* end = p + len;
  bool have[3] = {false, false, false};
  while (p < end) {
    uint64_t tag;
    if (!get_varint(p, end, tag)) goto malformed;
    if (tag != 0x0A) {  // not Example.features
      if (!skip_field(p, end, tag & 7)) goto malformed;
      continue;
    }
    const uint8_t* fe;
    if (!get_len(p, end, fe)) goto malformed;
    // Features: repeated map entries
    while (p < fe) {
      uint64_t t2;
      if (!get_varint(p, fe, t2)) goto malformed;
      if (t2 != 0x0A) {
        if (!skip_field(p, fe, t2 & 7)) goto malformed;
        continue;
      }
      const uint8_t* ee;
      if (!get_len(p, fe, ee)) goto malformed;
      Key key = kOther;
      const uint8_t* val = nullptr;
      const uint8_t* val_end = nullptr;
      while (p < ee) {
        uint64_t t3;
        if (!get_varint(p, ee, t3)) goto malformed;
        if (t3 == 0x0A) {
          const uint8_t* ke;
          if (!get_len(p, ee, ke)) goto malformed;
          key = classify(p, ke - p);
          p = ke;
        } else if (t3 == 0x12) {
          if (!get_len(p, ee, val_end)) goto malformed;
          val = p;
          p = val_end;
        } else if (!skip_field(p, ee, t3 & 7)) {
          goto malformed;
        }
      }
      if (key == kOther || !val) continue;
      // Feature: oneof list
      const uint8_t* q = val;
      while (q < val_end) {
        uint64_t t4;
        if (!get_varint(q, val_end, t4)) goto malformed;
        const uint8_t* le;
        if ((t4 & 7) != 2) {
          if (!skip_field(q, val_end, t4 & 7)) goto malformed;
          continue;
        }
        if (!get_len(q, val_end, le)) goto malformed;
        if (t4 == 0x12 && key == kLabel) {
          float tmp[1];
          if (read_float_list(q, le, tmp, 1) != 1) {
            set_error("Example: key 'label' must hold exactly one float");
            return -1;
          }
          *label_out = tmp[0];
          have[kLabel] = true;
        } else if (t4 == 0x12 && key == kValues) {
          int n = read_float_list(q, le, vals_out, F);
          if (n != F) {
            set_error("Example: key 'values' holds %d floats, field_size is %d", n, F);
            return -1;
          }
          have[kValues] = true;
        } else if (t4 == 0x1A && key == kIds) {
          uint64_t bad = 0;
          int n = ids_i32 ? read_int64_list<int32_t>(q, le, static_cast<int32_t*>(ids_out), F, max_id, &bad)
                          : read_int64_list<int64_t>(q, le, static_cast<int64_t*>(ids_out), F, max_id, &bad);
          if (n == -2) {
            // what tf.nn.embedding_lookup answers with InvalidArgument on CPU (HVD.py:197,201)
            set_error("Example: feature id %lld is outside [0, feature_size = %llu)", (long long)bad,
                      (unsigned long long)max_id);
            return -1;
          }
          if (n != F) {
            set_error("Example: key 'ids' holds %d ints, field_size is %d", n, F);
            return -1;
          }
          have[kIds] = true;
        }
        q = le;
      }
    }
  }
  if (!have[kLabel] || !have[kIds] || !have[kValues]) {
    set_error("Example: missing key(s):%s%s%s", have[kLabel] ? "" : " label", have[kIds] ? "" : " ids",
              have[kValues] ? "" : " values");
    return -1;
  }
  return 0;
malformed:
  set_error("Example: malformed protobuf");
  return -1;
}

void encode_example(float label, const int64_t* ids, const float* vals, int n, bool packed,
                    std::vector<uint8_t>& out) {
  // Build the three Feature bodies, then wrap: entry{key,value} -> Features -> Example.
  std::vector<uint8_t> feat[3];
  {  // label: FloatList
    std::vector<uint8_t> fl;
    if (packed) { fl.push_back(0x0A); put_varint(fl, 4); }
    else fl.push_back(0x0D);
    const uint8_t* b = reinterpret_cast<const uint8_t*>(&label);
    fl.insert(fl.end(), b, b + 4);
    feat[0].push_back(0x12); put_varint(feat[0], fl.size());
    feat[0].insert(feat[0].end(), fl.begin(), fl.end());
  }
  {  // ids: Int64List
    std::vector<uint8_t> il;
    if (packed) {
      size_t body = 0;
      for (int i = 0; i < n; ++i) body += varint_size(static_cast<uint64_t>(ids[i]));
      il.push_back(0x0A); put_varint(il, body);
      for (int i = 0; i < n; ++i) put_varint(il, static_cast<uint64_t>(ids[i]));
    } else {
      for (int i = 0; i < n; ++i) { il.push_back(0x08); put_varint(il, static_cast<uint64_t>(ids[i])); }
    }
    feat[1].push_back(0x1A); put_varint(feat[1], il.size());
    feat[1].insert(feat[1].end(), il.begin(), il.end());
  }
  {  // values: FloatList
    std::vector<uint8_t> fl;
    if (packed) {
      fl.push_back(0x0A); put_varint(fl, size_t(n) * 4);
      const uint8_t* b = reinterpret_cast<const uint8_t*>(vals);
      fl.insert(fl.end(), b, b + size_t(n) * 4);
    } else {
      for (int i = 0; i < n; ++i) {
        fl.push_back(0x0D);
        const uint8_t* b = reinterpret_cast<const uint8_t*>(vals + i);
        fl.insert(fl.end(), b, b + 4);
      }
    }
    feat[2].push_back(0x12); put_varint(feat[2], fl.size());
    feat[2].insert(feat[2].end(), fl.begin(), fl.end());
  }
  static const char* names[3] = {"label", "ids", "values"};
  std::vector<uint8_t> features;
  for (int k = 0; k < 3; ++k) {
    std::vector<uint8_t> entry;
    size_t kn = strlen(names[k]);
    entry.push_back(0x0A); put_varint(entry, kn);
    entry.insert(entry.end(), names[k], names[k] + kn);
    entry.push_back(0x12); put_varint(entry, feat[k].size());
    entry.insert(entry.end(), feat[k].begin(), feat[k].end());
    features.push_back(0x0A); put_varint(features, entry.size());
    features.insert(features.end(), entry.begin(), entry.end());
  }
  out.push_back(0x0A); put_varint(out, features.size());
  out.insert(out.end(), features.begin(), features.end());
}

}  // namespace dfio

// ------------------------------- C ABI ---------------------------------------
using namespace dfio;

extern "C" const char* dfio_last_error(void) { return g_err; }
extern "C" uint32_t dfio_crc32c(const void* d, size_t n) { return crc32c(d, n); }
extern "C" uint32_t dfio_masked_crc32c(const void* d, size_t n) { return mask_crc(crc32c(d, n)); }

extern "C" void* dfio_reader_open(const char* path, int verify_crc) {
  auto s = open_source(path, verify_crc != 0, verify_crc != 0, false);
  return s ? s.release() : nullptr;
}
extern "C" int dfio_reader_next(void* r, const uint8_t** data, uint64_t* len) {
  return static_cast<RecordSource*>(r)->next(data, len);
}
extern "C" void dfio_reader_close(void* r) { delete static_cast<RecordSource*>(r); }

extern "C" int64_t dfio_count_records(const char* path, int verify_crc) {
  auto s = open_source(path, verify_crc != 0, verify_crc != 0, false);
  if (!s) return -1;
  const uint8_t* d;
  uint64_t n;
  int64_t cnt = 0;
  int r;
  while ((r = s->next(&d, &n)) == 1) ++cnt;
  return r < 0 ? -1 : cnt;
}

extern "C" void* dfio_writer_open(const char* path) {
  FILE* f = fopen(path, "wb");
  if (!f) {
    set_error("cannot create %s: %s", path, strerror(errno));
    return nullptr;
  }
  setvbuf(f, nullptr, _IOFBF, 1 << 20);
  return new RecordWriter(f);
}
extern "C" int dfio_writer_write(void* w, const void* data, uint64_t len) {
  return static_cast<RecordWriter*>(w)->write(data, len) ? 0 : -1;
}
extern "C" int dfio_writer_write_example(void* w, float label, const int64_t* ids, const float* vals, int n) {
  std::vector<uint8_t> buf;
  buf.reserve(64 + size_t(n) * 14);
  encode_example(label, ids, vals, n, true, buf);
  return static_cast<RecordWriter*>(w)->write(buf.data(), buf.size()) ? 0 : -1;
}
extern "C" int dfio_writer_close(void* w) {
  auto* rw = static_cast<RecordWriter*>(w);
  bool ok = rw->close();
  delete rw;
  return ok ? 0 : -1;
}

extern "C" int64_t dfio_encode_example(float label, const int64_t* ids, const float* vals, int n,
                                       uint8_t* out, uint64_t cap, int packed) {
  std::vector<uint8_t> buf;
  encode_example(label, ids, vals, n, packed != 0, buf);
  if (buf.size() > cap) {
    set_error("dfio_encode_example: need %zu bytes, have %llu", buf.size(), (unsigned long long)cap);
    return -1;
  }
  memcpy(out, buf.data(), buf.size());
  return static_cast<int64_t>(buf.size());
}

extern "C" int dfio_parse_example(const void* buf, uint64_t len, int field_size, void* ids_out, int ids_i32,
                                  float* vals_out, float* label_out) {
  return parse_example(static_cast<const uint8_t*>(buf), len, field_size, ids_out, ids_i32 != 0, vals_out,
                       label_out);
}

extern "C" int64_t dfio_read_all(const char* path, int F, int shard_n, int shard_i, int64_t capacity,
                                 void* ids_out, int ids_i32, float* vals_out, float* labels_out,
                                 int verify_crc) {
  auto s = open_source(path, verify_crc != 0, verify_crc != 0, false);
  if (!s) return -1;
  if (shard_n < 1) shard_n = 1;
  const uint8_t* d;
  uint64_t n;
  int64_t idx = 0, kept = 0;
  int r;
  const size_t id_sz = ids_i32 ? 4 : 8;
  while ((r = s->next(&d, &n)) == 1) {
    if (idx++ % shard_n != shard_i) continue;
    if (kept >= capacity) break;
    if (parse_example(d, n, F, static_cast<uint8_t*>(ids_out) + size_t(kept) * F * id_sz, ids_i32 != 0,
                      vals_out + size_t(kept) * F, labels_out + kept) != 0)
      return -1;
    ++kept;
  }
  return r < 0 ? -1 : kept;
}
