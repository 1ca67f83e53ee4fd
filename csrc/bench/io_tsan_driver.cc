// ThreadSanitizer driver for the host data layer: the batcher thread, the worker pool (framing-index and parse tasks) and the
// consumer of libdfm_io's batch loader run over a synthetic file with a small ring, twice (eager release and
// release lagging two batches behind, the way Trainer.train_batches holds slots), and the stream is checked.  Both
// phases then run again in stream mode (the file read through read(2) in blocks that the batches pin and recycle).
//   g++ -std=c++17 -O1 -g -fsanitize=thread -msse4.2 -pthread -Iinclude csrc/bench/io_tsan_driver.cc csrc/io/*.cc -o build/io_tsan && build/io_tsan
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <vector>
#include "deepfm_io.h"

int main() {
  const char* path = "/tmp/dfm_tsan.tfrecords";
  const int F = 39, B = 64, slots = 4;
  const int64_t n = 64 * 200 + 17;   // ~4 MB: four framing-index ranges walked in parallel
  if (dfio_generate_synthetic(path, n, F, 5000, 13, 7, 0.25, 1.05) != n) { printf("generate: %s\n", dfio_last_error()); return 1; }
  for (int streaming = 0; streaming <= 1; ++streaming)
  for (int lag = 0; lag <= 2; lag += 2) {
    std::vector<std::vector<int32_t>> ids(slots, std::vector<int32_t>(B * F));
    std::vector<std::vector<float>> vals(slots, std::vector<float>(B * F)), labels(slots, std::vector<float>(B));
    std::vector<void*> pi, pv, pl;
    for (int s = 0; s < slots; ++s) { pi.push_back(ids[s].data()); pv.push_back(vals[s].data()); pl.push_back(labels[s].data()); }
    dfio_loader_config cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.field_size = F; cfg.batch_size = B; cfg.num_epochs = 2; cfg.drop_remainder = 1; cfg.shard_n = 1; cfg.shard_i = 0;
    cfg.num_workers = 4; cfg.num_slots = slots; cfg.ids_i32 = 1; cfg.verify_crc = 1; cfg.streaming = streaming;
    const char* paths[1] = {path};
    void* L = dfio_loader_create(&cfg, paths, 1, pi.data(), pv.data(), pl.data());
    if (!L) { printf("create: %s\n", dfio_last_error()); return 1; }
    std::deque<int> held;
    int slot, count, batches = 0;
    double checksum = 0;
    while (true) {
      const int r = dfio_loader_next(L, &slot, &count);
      if (r == 0) break;
      if (r < 0) { printf("next: %s\n", dfio_last_error()); return 1; }
      for (int i = 0; i < count; ++i) checksum += labels[slot][i] + ids[slot][i * F + 20] * 1e-6;
      ++batches;
      held.push_back(slot);
      while (int(held.size()) > lag) { dfio_loader_release(L, held.front()); held.pop_front(); }
    }
    while (!held.empty()) { dfio_loader_release(L, held.front()); held.pop_front(); }
    dfio_loader_destroy(L);
    printf("%s, lag %d: %d batches, checksum %.3f\n", streaming ? "stream" : "mapped", lag, batches, checksum);
    if (batches != 400) { printf("expected 400 batches\n"); return 1; }
  }
  // early destroy: the batcher abandons framing-index ranges a worker may still be walking (the mapping must outlive them)
  for (int it = 0; it < 40; ++it) {
    std::vector<std::vector<int32_t>> ids(slots, std::vector<int32_t>(B * F));
    std::vector<std::vector<float>> vals(slots, std::vector<float>(B * F)), labels(slots, std::vector<float>(B));
    std::vector<void*> pi, pv, pl;
    for (int s = 0; s < slots; ++s) { pi.push_back(ids[s].data()); pv.push_back(vals[s].data()); pl.push_back(labels[s].data()); }
    dfio_loader_config cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.field_size = F; cfg.batch_size = B; cfg.num_epochs = 1000; cfg.drop_remainder = 1; cfg.shard_n = 1;
    cfg.num_workers = 4; cfg.num_slots = slots; cfg.ids_i32 = 1; cfg.verify_crc = 1; cfg.streaming = it >= 20;
    const char* paths[1] = {path};
    void* L = dfio_loader_create(&cfg, paths, 1, pi.data(), pv.data(), pl.data());
    int slot, count;
    for (int k = 0; k < 1 + it % 3; ++k) {
      if (dfio_loader_next(L, &slot, &count) != 1) { printf("early: %s\n", dfio_last_error()); return 1; }
      dfio_loader_release(L, slot);
    }
    dfio_loader_destroy(L);
  }
  printf("40 early destroys ok (20 mapped, 20 stream)\n");
  printf("io tsan driver ok\n");
  return 0;
}
