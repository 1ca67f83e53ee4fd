// Address/UB-sanitizer fuzz of the tf.train.Example decoder (csrc/io/tfrecord.cc): canonical-layout fast path, the
// four-ids-per-step vector decode and the generic walker behind them.  Every payload is copied into a heap block of
// EXACTLY its length, so any read past the record is reported.  Records mix varint widths in the shapes that steer
// the decoder (random, small-then-large like Criteo, all three-byte, three-byte with stray two-byte ids); a third
// of them are damaged (random byte flips, truncation).  Intact records must round-trip; damaged ones must be
// decoded or rejected -- identically for int32 and int64 output -- and never crash.
//
//   g++ -std=c++17 -O1 -g -fsanitize=address,undefined -fno-sanitize-recover=undefined -msse4.2 -pthread -Iinclude \
//       csrc/bench/io_fuzz_driver.cc csrc/io/*.cc -o build/io_fuzz && build/io_fuzz
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <vector>
#include "deepfm_io.h"
int main() {
  std::mt19937_64 rng(7);
  long ok = 0, bad = 0;
  for (int it = 0; it < 400000; ++it) {
    int F = 1 + rng() % 48;
    std::vector<int64_t> ids(F); std::vector<float> vals(F, 1.f);
    int mode = rng() % 4;
    for (int i = 0; i < F; ++i) {
      int cls = mode == 0 ? rng() % 4 : mode == 1 ? (i < F / 3 ? 0 : 2) : mode == 2 ? 2 : (rng() % 16 ? 2 : 1);
      int64_t lo[] = {0, 128, 16384, 1 << 21}, hi[] = {128, 16384, 1 << 21, 1 << 28};
      ids[i] = lo[cls] + rng() % (hi[cls] - lo[cls]);
    }
    uint8_t tmp[2048];
    int64_t n = dfio_encode_example(1.f, ids.data(), vals.data(), F, tmp, sizeof(tmp), 1);
    if (n <= 0) return 2;
    int flips = rng() % 4;            // 0 = intact
    for (int k = 0; k < flips; ++k) tmp[rng() % n] = (uint8_t)rng();
    int64_t cut = (rng() % 8 == 0) ? rng() % n : 0;
    size_t len = n - cut;
    uint8_t* exact = (uint8_t*)malloc(len ? len : 1);   // exact size: ASAN flags any overread
    memcpy(exact, tmp, len);
    std::vector<int32_t> o32(F); std::vector<int64_t> o64(F); std::vector<float> ov(F); float lab;
    int r1 = dfio_parse_example(exact, len, F, o32.data(), 1, ov.data(), &lab);
    int r2 = dfio_parse_example(exact, len, F, o64.data(), 0, ov.data(), &lab);
    if (r1 != r2) { printf("dtype disagreement at it %d\n", it); return 3; }
    if (r1 == 0) { ++ok; for (int i = 0; i < F; ++i) if ((int32_t)o64[i] != o32[i]) { printf("value disagreement\n"); return 4; }
      if (!flips && !cut) for (int i = 0; i < F; ++i) if (o64[i] != ids[i]) { printf("roundtrip mismatch it %d i %d\n", it, i); return 5; } }
    else { ++bad; if (!flips && !cut) { printf("intact record rejected: %s\n", dfio_last_error()); return 6; } }
    free(exact);
  }
  printf("fuzz ok: %ld parsed, %ld rejected\n", ok, bad);
}
