// Offline tools: libsvm text -> TFRecord (the job of the reference's
// tools/libsvm_to_tfrecord.py, TOOL.py:22-61) and a Criteo-shaped synthetic
// generator with the statistics of the bundled data/val.tfrecords (13 numeric
// fields with constant ids 1..13, 26 categorical fields with value 1.0 and
// heavy-tailed ids, ~25% positives).
#include <cerrno>
#include <cmath>
#include <cstdlib>

#include "io_internal.h"

namespace dfio {
namespace {

struct SplitMix64 {
  uint64_t s;
  explicit SplitMix64(uint64_t seed) : s(seed) {}
  uint64_t next() {
    uint64_t z = (s += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
  }
  double uniform() { return (next() >> 11) * (1.0 / 9007199254740992.0); }
};

// Approximate Zipf(s) over [0, n): inverse-CDF of the continuous power law.
inline int64_t zipf(SplitMix64& rng, int64_t n, double s) {
  if (n <= 1) return 0;
  const double u = rng.uniform();
  double x;
  if (std::fabs(s - 1.0) < 1e-9) {
    x = std::exp(u * std::log(double(n) + 1.0)) - 1.0;
  } else {
    const double a = 1.0 - s;
    x = std::pow(u * (std::pow(double(n) + 1.0, a) - 1.0) + 1.0, 1.0 / a) - 1.0;
  }
  int64_t k = static_cast<int64_t>(x);
  return k < 0 ? 0 : (k >= n ? n - 1 : k);
}

}  // namespace
}  // namespace dfio

using namespace dfio;

extern "C" int64_t dfio_libsvm_to_tfrecord(const char* in_path, const char* out_path, int expect_fields) {
  FILE* in = strcmp(in_path, "-") == 0 ? stdin : fopen(in_path, "r");
  if (!in) {
    set_error("cannot open %s: %s", in_path, strerror(errno));
    return -1;
  }
  FILE* out = fopen(out_path, "wb");
  if (!out) {
    set_error("cannot create %s: %s", out_path, strerror(errno));
    if (in != stdin) fclose(in);
    return -1;
  }
  setvbuf(out, nullptr, _IOFBF, 1 << 20);
  RecordWriter w(out);
  char* line = nullptr;
  size_t cap = 0;
  ssize_t n;
  int64_t count = 0, lineno = 0;
  std::vector<int64_t> ids;
  std::vector<float> vals;
  std::vector<uint8_t> buf;
  int64_t rc = 0;
  while ((n = getline(&line, &cap, in)) >= 0) {
    ++lineno;
    char* p = line;
    while (*p == ' ' || *p == '\t') ++p;
    if (*p == '\n' || *p == '\r' || *p == 0 || *p == '#') continue;
    char* e;
    float label = strtof(p, &e);
    if (e == p) {
      set_error("%s:%lld: cannot parse label", in_path, (long long)lineno);
      rc = -1;
      break;
    }
    p = e;
    ids.clear();
    vals.clear();
    for (;;) {
      while (*p == ' ' || *p == '\t') ++p;
      if (*p == '\n' || *p == '\r' || *p == 0) break;
      long long id = strtoll(p, &e, 10);
      if (e == p || *e != ':') {
        set_error("%s:%lld: expected <id>:<val> near '%.16s'", in_path, (long long)lineno, p);
        rc = -1;
        break;
      }
      p = e + 1;
      float v = strtof(p, &e);
      if (e == p) {
        set_error("%s:%lld: cannot parse value of id %lld", in_path, (long long)lineno, id);
        rc = -1;
        break;
      }
      p = e;
      ids.push_back(id);
      vals.push_back(v);
    }
    if (rc < 0) break;
    if (expect_fields > 0 && static_cast<int>(ids.size()) != expect_fields) {
      set_error("%s:%lld: %zu id:val pairs, expected %d", in_path, (long long)lineno, ids.size(), expect_fields);
      rc = -1;
      break;
    }
    buf.clear();
    encode_example(label, ids.data(), vals.data(), static_cast<int>(ids.size()), true, buf);
    if (!w.write(buf.data(), buf.size())) {
      rc = -1;
      break;
    }
    ++count;
  }
  free(line);
  if (in != stdin) fclose(in);
  if (!w.close() && rc == 0) rc = -1;
  return rc < 0 ? rc : count;
}

extern "C" int64_t dfio_generate_synthetic(const char* out_path, int64_t n_records, int F, int64_t V, int n_dense,
                                           uint64_t seed, double pos_rate, double zipf_s) {
  if (F <= 0 || n_dense < 0 || n_dense > F || V <= n_dense + (F - n_dense)) {
    set_error("dfio_generate_synthetic: need feature_size > field_size and 0 <= n_dense <= field_size");
    return -1;
  }
  FILE* out = fopen(out_path, "wb");
  if (!out) {
    set_error("cannot create %s: %s", out_path, strerror(errno));
    return -1;
  }
  setvbuf(out, nullptr, _IOFBF, 1 << 20);
  RecordWriter w(out);
  SplitMix64 rng(seed ? seed : 0x5EEDull);
  const int n_cat = F - n_dense;
  // categorical field c owns ids [lo_c, lo_c + size_c) inside [n_dense + 1, V)
  std::vector<int64_t> lo(n_cat), size(n_cat);
  const int64_t span = V - (n_dense + 1);
  for (int c = 0; c < n_cat; ++c) {
    lo[c] = n_dense + 1 + span * c / n_cat;
    size[c] = n_dense + 1 + span * (c + 1) / n_cat - lo[c];
    if (size[c] < 1) size[c] = 1;
  }
  // a fixed hidden logistic model gives the labels some learnable signal
  std::vector<float> wd(n_dense);
  for (auto& x : wd) x = static_cast<float>(rng.uniform() * 2.0 - 1.0);
  const double bias = std::log(pos_rate / (1.0 - pos_rate));
  std::vector<int64_t> ids(F);
  std::vector<float> vals(F);
  std::vector<uint8_t> buf;
  for (int64_t r = 0; r < n_records; ++r) {
    double z = bias;
    for (int f = 0; f < n_dense; ++f) {
      ids[f] = f + 1;
      // mostly small values with an occasional large one, like the log-scaled Criteo counters
      double u = rng.uniform();
      float v = static_cast<float>(u < 0.3 ? 0.0 : (u < 0.95 ? rng.uniform() * 0.5 : rng.uniform() * 8.0));
      vals[f] = v;
      z += wd[f] * (v - 0.2);
    }
    for (int c = 0; c < n_cat; ++c) {
      int64_t k = zipf(rng, size[c], zipf_s);
      ids[n_dense + c] = lo[c] + k;
      vals[n_dense + c] = 1.0f;
      // hash of the id nudges the logit so categorical embeddings matter too
      uint64_t h = (static_cast<uint64_t>(ids[n_dense + c]) * 0x9E3779B97F4A7C15ull) >> 40;
      z += ((h & 0xFF) / 255.0 - 0.5) * 0.6;
    }
    const float label = rng.uniform() < 1.0 / (1.0 + std::exp(-z)) ? 1.0f : 0.0f;
    buf.clear();
    encode_example(label, ids.data(), vals.data(), F, true, buf);
    if (!w.write(buf.data(), buf.size())) return -1;
  }
  return w.close() ? n_records : -1;
}
