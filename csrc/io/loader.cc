// Sharding batch loader: shard -> batch(drop_remainder) -> parse(batch) -> repeat ->
// prefetch, the order of the reference's input_fn (HVD.py:127-159), as a native
// pipeline:
//
//   reader thread   walks the record framing of every file of every epoch, applies the
//                   record-level shard (index % n == i), groups kept records into batches
//                   and hands each batch to the workers in row ranges
//   worker threads  verify + decode Examples straight into the caller's (pinned) slot
//   consumer        dfio_loader_next() returns slots strictly in stream order
//
// Slots form a ring: batch b lives in slot b % num_slots, so "prefetch depth" is
// num_slots - 1 and the consumer's release() is what lets the reader run ahead.
#include <atomic>
#include <condition_variable>
#include <deque>
#include <mutex>
#include <thread>

#include "io_internal.h"

namespace dfio {
namespace {

constexpr uint64_t kNoOffset = ~0ull;

struct Rec {
  const uint8_t* p;   // payload (resolved from `off` once the arena is final)
  uint32_t len;
  uint64_t off;       // payload offset inside BatchWork::arena, or kNoOffset for mapped records
};

// One batch worth of records.  For mapped files the records point into the
// mapping (kept alive by `maps`); for streams they point into `arena`.
struct BatchWork {
  int slot = 0;
  int count = 0;
  std::vector<Rec> recs;
  std::vector<uint8_t> arena;
  std::vector<std::shared_ptr<RecordSource>> maps;
  std::atomic<int> remaining{0};
};

struct Task {
  std::shared_ptr<BatchWork> bw;
  int row0, row1;
};

enum SlotState { kFree = 0, kFilling = 1, kReady = 2, kInUse = 3 };

class Loader {
 public:
  Loader(const dfio_loader_config& c, std::vector<std::string> paths, void* const* ids, void* const* vals,
         void* const* labels)
      : cfg_(c), paths_(std::move(paths)) {
    if (cfg_.num_workers < 1) cfg_.num_workers = 1;
    if (cfg_.num_slots < 2) cfg_.num_slots = 2;
    if (cfg_.shard_n < 1) cfg_.shard_n = 1;
    if (cfg_.num_epochs < 1) cfg_.num_epochs = 1;
    slots_.resize(cfg_.num_slots);
    for (int i = 0; i < cfg_.num_slots; ++i) {
      slots_[i].ids = ids[i];
      slots_[i].vals = static_cast<float*>(vals[i]);
      slots_[i].labels = static_cast<float*>(labels[i]);
    }
    reader_ = std::thread([this] { reader_main(); });
    for (int i = 0; i < cfg_.num_workers; ++i) workers_.emplace_back([this] { worker_main(); });
  }

  ~Loader() {
    {
      std::lock_guard<std::mutex> l(mu_);
      stop_ = true;
    }
    cv_slot_.notify_all();
    cv_task_.notify_all();
    cv_ready_.notify_all();
    if (reader_.joinable()) reader_.join();
    for (auto& w : workers_) w.join();
  }

  int next(int* slot, int* count) {
    std::unique_lock<std::mutex> l(mu_);
    const int s = static_cast<int>(next_batch_ % cfg_.num_slots);
    cv_ready_.wait(l, [&] {
      return stop_ || failed_ || (slots_[s].state == kReady && slots_[s].batch_no == next_batch_) ||
             (reader_done_ && next_batch_ >= produced_);
    });
    if (failed_) {
      set_error("%s", err_.c_str());
      return -1;
    }
    if (slots_[s].state == kReady && slots_[s].batch_no == next_batch_) {
      slots_[s].state = kInUse;
      *slot = s;
      *count = slots_[s].count;
      ++next_batch_;
      return 1;
    }
    return 0;
  }

  int release(int s) {
    {
      std::lock_guard<std::mutex> l(mu_);
      if (s < 0 || s >= cfg_.num_slots || slots_[s].state != kInUse) {
        set_error("dfio_loader_release: slot %d is not held by the consumer", s);
        return -1;
      }
      slots_[s].state = kFree;
    }
    cv_slot_.notify_all();
    return 0;
  }

  void stats(int64_t* rec, int64_t* bat, int64_t* bytes) {
    if (rec) *rec = records_.load();
    if (bat) *bat = batches_.load();
    if (bytes) *bytes = bytes_.load();
  }

 private:
  struct Slot {
    void* ids = nullptr;
    float* vals = nullptr;
    float* labels = nullptr;
    int state = kFree;
    int count = 0;
    int64_t batch_no = -1;
  };

  void fail(const std::string& msg) {
    {
      std::lock_guard<std::mutex> l(mu_);
      if (!failed_) {
        failed_ = true;
        err_ = msg;
      }
    }
    cv_ready_.notify_all();
    cv_slot_.notify_all();
    cv_task_.notify_all();
  }

  // Blocks until the ring slot of batch `produced_` is free, then queues its tasks.
  bool submit(std::shared_ptr<BatchWork> bw) {
    std::unique_lock<std::mutex> l(mu_);
    const int s = static_cast<int>(produced_ % cfg_.num_slots);
    cv_slot_.wait(l, [&] { return stop_ || failed_ || slots_[s].state == kFree; });
    if (stop_ || failed_) return false;
    slots_[s].state = kFilling;
    slots_[s].count = bw->count;
    slots_[s].batch_no = produced_;
    bw->slot = s;
    // split into row ranges so several workers share one batch
    const int chunk = std::max(64, (bw->count + cfg_.num_workers - 1) / cfg_.num_workers);
    int ntasks = (bw->count + chunk - 1) / chunk;
    bw->remaining.store(ntasks);
    for (int r = 0; r < bw->count; r += chunk) tasks_.push_back(Task{bw, r, std::min(bw->count, r + chunk)});
    ++produced_;
    l.unlock();
    cv_task_.notify_all();
    return true;
  }

  void reader_main() {
    const int B = cfg_.batch_size;
    for (int epoch = 0; epoch < cfg_.num_epochs; ++epoch) {
      auto bw = std::make_shared<BatchWork>();
      bw->recs.reserve(B);
      int64_t idx = 0;  // record index within the epoch's concatenated stream
      for (const auto& path : paths_) {
        std::shared_ptr<RecordSource> src(open_source(path, /*verify_mapped=*/false, /*verify_stream=*/cfg_.verify_crc != 0,
                                                        cfg_.streaming != 0).release());
        if (!src) {
          fail(dfio_last_error());
          return;
        }
        const bool stable = src->stable_pointers();
        if (stable) bw->maps.push_back(src);
        const uint8_t* d;
        uint64_t n;
        int r;
        while ((r = src->next(&d, &n)) == 1) {
          if (idx++ % cfg_.shard_n != cfg_.shard_i) continue;
          if (stable) {
            bw->recs.push_back(Rec{d, static_cast<uint32_t>(n), kNoOffset});
          } else {
            // Streamed payloads are copied (their CRCs were checked by the stream source);
            // offsets now, pointers once the arena stops growing.
            const size_t off = bw->arena.size();
            bw->arena.insert(bw->arena.end(), d, d + n);
            bw->recs.push_back(Rec{nullptr, static_cast<uint32_t>(n), off});
          }
          if (static_cast<int>(bw->recs.size()) == B) {
            if (!finish_and_submit(bw)) return;
            bw = std::make_shared<BatchWork>();
            bw->recs.reserve(B);
            if (stable) bw->maps.push_back(src);
          }
          // (stop / failure are noticed in submit(), once per batch: no lock per record on the reader's path)
        }
        bytes_ += static_cast<int64_t>(src->bytes_consumed());
        if (r < 0) {
          fail(dfio_last_error());
          return;
        }
      }
      if (!bw->recs.empty() && !cfg_.drop_remainder) {
        if (!finish_and_submit(bw)) return;
      }
    }
    {
      std::lock_guard<std::mutex> l(mu_);
      reader_done_ = true;
    }
    cv_ready_.notify_all();
  }

  bool finish_and_submit(std::shared_ptr<BatchWork>& bw) {
    bw->count = static_cast<int>(bw->recs.size());
    const uint8_t* base = bw->arena.data();
    for (auto& r : bw->recs)
      if (r.off != kNoOffset) r.p = base + r.off;
    return submit(bw);
  }

  void worker_main() {
    const int F = cfg_.field_size;
    const size_t id_sz = cfg_.ids_i32 ? 4 : 8;
    for (;;) {
      Task t;
      {
        std::unique_lock<std::mutex> l(mu_);
        cv_task_.wait(l, [&] { return stop_ || failed_ || !tasks_.empty(); });
        if (stop_ || failed_) return;
        t = std::move(tasks_.front());
        tasks_.pop_front();
      }
      Slot& s = slots_[t.bw->slot];
      bool ok = true;
      for (int r = t.row0; r < t.row1 && ok; ++r) {
        const Rec& rec = t.bw->recs[r];
        if (cfg_.verify_crc && rec.off == kNoOffset) {  // mapped record: framing sits around the payload
          uint32_t lcrc, dcrc;
          memcpy(&lcrc, rec.p - 4, 4);
          memcpy(&dcrc, rec.p + rec.len, 4);
          if (mask_crc(crc32c(rec.p - 12, 8)) != lcrc || mask_crc(crc32c(rec.p, rec.len)) != dcrc) {
            fail("TFRecord CRC mismatch (corrupt input)");
            ok = false;
            break;
          }
        }
        if (parse_example(rec.p, rec.len, F, static_cast<uint8_t*>(s.ids) + size_t(r) * F * id_sz,
                          cfg_.ids_i32 != 0, s.vals + size_t(r) * F, s.labels + r) != 0) {
          fail(dfio_last_error());
          ok = false;
        }
      }
      if (!ok) return;
      records_ += t.row1 - t.row0;
      if (t.bw->remaining.fetch_sub(1) == 1) {
        {
          std::lock_guard<std::mutex> l(mu_);
          s.state = kReady;
        }
        ++batches_;
        cv_ready_.notify_all();
      }
    }
  }

  dfio_loader_config cfg_;
  std::vector<std::string> paths_;
  std::vector<Slot> slots_;
  std::mutex mu_;
  std::condition_variable cv_slot_, cv_task_, cv_ready_;
  std::deque<Task> tasks_;
  int64_t produced_ = 0;    // batches handed to workers
  int64_t next_batch_ = 0;  // next batch the consumer gets
  bool reader_done_ = false, stop_ = false, failed_ = false;
  std::string err_;
  std::thread reader_;
  std::vector<std::thread> workers_;
  std::atomic<int64_t> records_{0}, batches_{0}, bytes_{0};
};

}  // namespace
}  // namespace dfio

using namespace dfio;

extern "C" void* dfio_loader_create(const dfio_loader_config* cfg, const char* const* paths, int n_paths,
                                    void* const* slot_ids, void* const* slot_vals, void* const* slot_labels) {
  if (!cfg || cfg->field_size <= 0 || cfg->batch_size <= 0) {
    set_error("dfio_loader_create: field_size and batch_size must be positive");
    return nullptr;
  }
  if (cfg->shard_n > 0 && (cfg->shard_i < 0 || cfg->shard_i >= cfg->shard_n)) {
    set_error("dfio_loader_create: shard index %d outside [0,%d)", cfg->shard_i, cfg->shard_n);
    return nullptr;
  }
  std::vector<std::string> p;
  for (int i = 0; i < n_paths; ++i) p.emplace_back(paths[i]);
  return new Loader(*cfg, std::move(p), slot_ids, slot_vals, slot_labels);
}
extern "C" int dfio_loader_next(void* h, int* slot, int* count) { return static_cast<Loader*>(h)->next(slot, count); }
extern "C" int dfio_loader_release(void* h, int slot) { return static_cast<Loader*>(h)->release(slot); }
extern "C" int dfio_loader_stats(void* h, int64_t* r, int64_t* b, int64_t* y) {
  static_cast<Loader*>(h)->stats(r, b, y);
  return 0;
}
extern "C" void dfio_loader_destroy(void* h) { delete static_cast<Loader*>(h); }
