// Sharding batch loader: shard -> batch(drop_remainder) -> parse(batch) -> repeat ->
// prefetch, the order of the reference's input_fn (HVD.py:127-159), as a native
// pipeline:
//
//   index tasks     (mapped files) the record framing of disjoint 1 MiB byte ranges is walked by the worker pool in
//                   parallel: a range that does not start at byte 0 first looks for a plausible record start (two
//                   consecutive length CRCs that check out) and the batcher confirms it against the true chain --
//                   the previous range's end -- re-walking the range itself on the rare false guess
//   batcher thread  consumes the ranges in order, applies the record-level shard (index % n == i), groups kept
//                   records into batches and hands each batch to the workers in row ranges.  FIFOs / stdin
//                   (SageMaker pipe mode) cannot be indexed ahead: this thread reads them in 4 MiB blocks, checks
//                   the length fields and cuts the block into records IN PLACE -- the batch keeps the block alive,
//                   nothing is copied after read(2), and the payload CRCs are left to the workers
//   parse tasks     verify payload CRCs + decode Examples straight into the caller's (pinned) slot, check the id
//                   range
//   consumer        dfio_loader_next() returns slots strictly in stream order
//
// Slots form a ring: batch b lives in slot b % num_slots, so "prefetch depth" is
// num_slots - 1 and the consumer's release() is what lets the batcher run ahead.
#include <errno.h>
#include <pthread.h>
#include <unistd.h>

#include <atomic>
#include <condition_variable>
#include <deque>
#include <mutex>
#include <new>
#include <thread>

#include "io_internal.h"

namespace dfio {
namespace {

constexpr size_t kIndexChunk = 1u << 20;
constexpr size_t kStreamBlock = 4u << 20;
constexpr uint64_t kMaxStreamRecord = 1ull << 31;

struct Rec {
  const uint8_t* p;   // payload; the record's 4-byte payload CRC follows it
  uint32_t len;
  uint32_t seg;       // index into BatchWork::segs (what the record points into)
};

// What the records of a batch point into: a mapped file, or one block of a stream.
struct FileSeg {
  std::shared_ptr<const void> keep;    // the mapping / the block: alive while the batch points into it
  const uint8_t* base = nullptr;       // its first byte ...
  uint64_t base_off = 0;               // ... and where that byte sits in the file or stream (for error messages)
  std::string path;
};

// A stretch of a stream as read(2) delivered it; recycled through Loader::free_blocks_ so its pages stay resident.
struct StreamBlock {
  std::vector<uint8_t> buf;
  uint64_t stream_off = 0;             // stream position of buf[0]
};

// One batch worth of records (they point into what `segs` keeps alive).
struct BatchWork {
  int slot = 0;
  int count = 0;
  std::vector<Rec> recs;
  std::vector<FileSeg> segs;
  std::atomic<int> remaining{0};
};

// A byte range of a mapped file whose framing one worker walks.
struct Chunk {
  std::shared_ptr<RecordSource> src;        // keeps the mapping alive: on stop / failure the batcher abandons ranges that
                                            // a worker may still be walking
  const uint8_t* base = nullptr;
  size_t size = 0, begin = 0, end = 0;
  bool verify = false;
  const char* path = "";
  size_t start = static_cast<size_t>(-1);   // first record start found in [begin, end) (guess unless begin == 0)
  size_t stop = 0;                          // position after the last record walked
  std::vector<RecSpan> recs;
  std::string err;
  bool done = false;                        // guarded by Loader::mu_
};

struct Task {
  std::shared_ptr<BatchWork> bw;            // parse task: rows [row0, row1) of this batch
  int row0 = 0, row1 = 0;
  std::shared_ptr<Chunk> chunk;             // index task
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
    cv_chunk_.notify_all();
    if (reader_.joinable()) reader_.join();
    for (auto& w : workers_) w.join();
    parse_tasks_.clear();        // batches still queued give their stream blocks back while the pool exists
    index_tasks_.clear();
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
    cv_chunk_.notify_all();
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
    // split into row ranges so several workers share one batch -- but not finer than ~30 us of decoding: with many
    // workers the ring holds several batches in flight, and hand-offs, not rows, are what finer tasks multiply
    const int chunk = std::max(256, (bw->count + cfg_.num_workers - 1) / cfg_.num_workers);
    int ntasks = (bw->count + chunk - 1) / chunk;
    bw->remaining.store(ntasks);
    for (int r = 0; r < bw->count; r += chunk) parse_tasks_.push_back(Task{bw, r, std::min(bw->count, r + chunk), nullptr});
    ++produced_;
    l.unlock();
    cv_task_.notify_all();
    return true;
  }

  // ---- batcher -------------------------------------------------------------------------------
  struct EpochState {
    std::shared_ptr<BatchWork> bw;
    int64_t idx = 0;          // record index within the epoch's concatenated stream
  };

  void new_batch(EpochState& es) {
    es.bw = std::make_shared<BatchWork>();
    es.bw->recs.reserve(cfg_.batch_size);
  }

  // the batch's reference to the mapping / block the next records lie in
  uint32_t seg_of(EpochState& es, const std::shared_ptr<const void>& keep, const uint8_t* base, uint64_t base_off,
                  const std::string& path) {
    auto& segs = es.bw->segs;
    if (segs.empty() || segs.back().keep != keep) segs.push_back(FileSeg{keep, base, base_off, path});
    return static_cast<uint32_t>(segs.size() - 1);
  }

  bool add_record(EpochState& es, const Rec& r) {
    es.bw->recs.push_back(r);
    if (static_cast<int>(es.bw->recs.size()) < cfg_.batch_size) return true;
    if (!finish_and_submit(es.bw)) return false;
    new_batch(es);
    return true;
  }

  void index_chunk(Chunk& c) {
    size_t pos = c.begin == 0 ? 0 : find_record_start(c.base, c.size, c.begin, c.end);
    c.start = pos;
    if (pos == static_cast<size_t>(-1)) {
      c.stop = c.end;
      return;
    }
    c.recs.reserve((c.end - pos) / 256 + 16);
    const size_t stop = walk_framing(c.base, c.size, pos, c.end, c.verify, &c.recs, c.path);
    if (stop == static_cast<size_t>(-1)) c.err = dfio_last_error();
    c.stop = stop;
  }

  // Mapped file: ranges are indexed by the worker pool ahead of the batcher, which consumes them in order.
  bool read_mapped(EpochState& es, const std::shared_ptr<RecordSource>& src, const uint8_t* base, size_t size,
                   const std::string& path) {
    const size_t n_chunks = (size + kIndexChunk - 1) / kIndexChunk;
    const size_t window = static_cast<size_t>(std::max(2, cfg_.num_workers));
    std::deque<std::shared_ptr<Chunk>> inflight;
    const std::shared_ptr<const void> keep = src;
    size_t next_chunk = 0, expected = 0;      // `expected` = true start of the next record (the chain)
    for (size_t ci = 0; ci < n_chunks; ++ci) {
      while (next_chunk < n_chunks && inflight.size() < window) {
        auto c = std::make_shared<Chunk>();
        c->src = src; c->base = base; c->size = size; c->verify = cfg_.verify_crc != 0; c->path = path.c_str();
        c->begin = next_chunk * kIndexChunk;
        c->end = std::min(size, c->begin + kIndexChunk);
        {
          std::lock_guard<std::mutex> l(mu_);
          index_tasks_.push_back(Task{nullptr, 0, 0, c});
        }
        cv_task_.notify_one();
        inflight.push_back(c);
        ++next_chunk;
      }
      std::shared_ptr<Chunk> c = inflight.front();
      inflight.pop_front();
      {
        std::unique_lock<std::mutex> l(mu_);
        cv_chunk_.wait(l, [&] { return stop_ || failed_ || c->done; });
        if (stop_ || failed_) return false;
      }
      // the true first record of this range starts at `expected` if that lies inside it, else the range has none
      const size_t truth = expected < c->end ? expected : static_cast<size_t>(-1);
      if (c->start != truth) {                // wrong guess (or a record spanning the whole range): walk it here
        c->recs.clear();
        c->err.clear();
        c->stop = expected;
        if (truth != static_cast<size_t>(-1)) {
          const size_t stop = walk_framing(base, size, truth, c->end, cfg_.verify_crc != 0, &c->recs, path.c_str());
          if (stop == static_cast<size_t>(-1)) c->err = dfio_last_error();
          c->stop = stop;
        }
      }
      if (!c->err.empty()) {
        fail(c->err);
        return false;
      }
      expected = std::max(expected, c->stop);
      if (!add_spans(es, c->recs, keep, base, path)) return false;
    }
    bytes_ += static_cast<int64_t>(size);
    return true;
  }

  // The records of one indexed range, after the record-level shard, appended to the open batch(es).  This loop is
  // the batcher thread's whole per-record cost and the one serial stage of the loader, so it is kept to a copy:
  // no division per record (the shard phase is carried along), the batch's reference to the mapping is looked up
  // once per run of records, and runs go in with one resize.
  bool add_spans(EpochState& es, const std::vector<RecSpan>& spans, const std::shared_ptr<const void>& keep,
                 const uint8_t* base, const std::string& path) {
    const size_t B = static_cast<size_t>(cfg_.batch_size);
    const int64_t N = cfg_.shard_n, I = cfg_.shard_i;
    size_t i = 0;
    const size_t n = spans.size();
    while (i < n) {
      std::vector<Rec>& v = es.bw->recs;
      const uint32_t seg = seg_of(es, keep, base, 0, path);
      if (N == 1) {
        const size_t take = std::min(B - v.size(), n - i);
        const size_t o = v.size();
        v.resize(o + take);
        for (size_t k = 0; k < take; ++k) v[o + k] = Rec{base + spans[i + k].off, spans[i + k].len, seg};
        i += take;
        es.idx += static_cast<int64_t>(take);
      } else {
        int64_t phase = es.idx % N;
        for (; i < n && v.size() < B; ++i, ++es.idx) {
          if (phase == I) v.push_back(Rec{base + spans[i].off, spans[i].len, seg});
          if (++phase == N) phase = 0;
        }
      }
      if (v.size() == B) {
        if (!finish_and_submit(es.bw)) return false;
        new_batch(es);
      }
    }
    return true;
  }

  // ---- streams ----------------------------------------------------------------------------------
  std::shared_ptr<StreamBlock> get_block(size_t cap) {
    std::unique_ptr<StreamBlock> b;
    {
      std::lock_guard<std::mutex> l(pool_mu_);
      if (!free_blocks_.empty()) {
        b = std::move(free_blocks_.back());
        free_blocks_.pop_back();
      }
    }
    if (!b) b = std::make_unique<StreamBlock>();
    if (b->buf.size() < cap) b->buf.resize(cap);
    return std::shared_ptr<StreamBlock>(b.release(), [this](StreamBlock* x) {
      std::lock_guard<std::mutex> l(pool_mu_);
      if (free_blocks_.size() < 16 && x->buf.size() <= 4 * kStreamBlock) free_blocks_.emplace_back(x);
      else delete x;       // (a block grown for one huge record is not kept around)
    });
  }

  // FIFO / stdin: read in blocks, cut into records in place.  A record that straddles the end of a block is carried
  // to the front of the next one (a record larger than a block gets a block of its own size).
  bool read_stream(EpochState& es, int fd, const std::string& path) {
    const bool verify = cfg_.verify_crc != 0;
    std::shared_ptr<StreamBlock> blk = get_block(kStreamBlock);
    std::shared_ptr<const void> keep = blk;
    blk->stream_off = 0;
    size_t fill = 0, pos = 0;       // bytes of the block that hold data / that are already cut into records
    uint64_t total = 0;
    for (bool eof = false; !eof;) {
      if (fill == blk->buf.size()) {
        const size_t tail = fill - pos;
        size_t need = kStreamBlock;
        if (tail >= 12) {           // its length field passed the checks below when it was first seen
          uint64_t n;
          memcpy(&n, blk->buf.data() + pos, 8);
          need = std::max<size_t>(need, n + 16);
        }
        std::shared_ptr<StreamBlock> nb;
        try {
          nb = get_block(need);
        } catch (const std::bad_alloc&) {
          fail(path + ": cannot buffer a record of " + std::to_string(need - 16) + " bytes (stream offset " +
               std::to_string(blk->stream_off + pos) + ")");
          return false;
        }
        memcpy(nb->buf.data(), blk->buf.data() + pos, tail);
        nb->stream_off = blk->stream_off + pos;
        blk = std::move(nb);
        keep = blk;
        fill = tail;
        pos = 0;
      }
      const ssize_t r = ::read(fd, blk->buf.data() + fill, blk->buf.size() - fill);
      if (r < 0) {
        if (errno == EINTR) continue;
        fail(path + ": read failed: " + strerror(errno));
        return false;
      }
      eof = r == 0;
      fill += static_cast<size_t>(r);
      total += static_cast<uint64_t>(r);
      const uint8_t* b = blk->buf.data();
      while (fill - pos >= 12) {
        uint64_t n;
        uint32_t lcrc;
        memcpy(&n, b + pos, 8);
        memcpy(&lcrc, b + pos + 8, 4);
        if (verify && mask_crc(crc32c_u64(n)) != lcrc) {
          fail(path + ": corrupt length CRC at stream offset " + std::to_string(blk->stream_off + pos));
          return false;
        }
        if (n > kMaxStreamRecord) {
          fail(path + ": implausible record length " + std::to_string(n) + " at stream offset " +
               std::to_string(blk->stream_off + pos));
          return false;
        }
        if (fill - pos < 16 + n) break;     // the rest of it has not arrived yet
        if (es.idx++ % cfg_.shard_n == cfg_.shard_i &&
            !add_record(es, Rec{b + pos + 12, static_cast<uint32_t>(n), seg_of(es, keep, b, blk->stream_off, path)}))
          return false;
        pos += 16 + n;
      }
    }
    bytes_ += static_cast<int64_t>(total);
    if (pos != fill) {
      fail(path + ": stream ends inside a record (stream offset " + std::to_string(blk->stream_off + pos) + ")");
      return false;
    }
    return true;
  }

  void reader_main() {
    pthread_setname_np(pthread_self(), "dfio-batcher");     // names show up in /proc/<pid>/task/*/comm, top -H
    for (int epoch = 0; epoch < cfg_.num_epochs; ++epoch) {
      EpochState es;
      new_batch(es);
      for (const auto& path : paths_) {
        std::shared_ptr<RecordSource> src(open_source(path, /*verify_mapped=*/false, /*verify_stream=*/cfg_.verify_crc != 0,
                                                        cfg_.streaming != 0).release());
        if (!src) {
          fail(dfio_last_error());
          return;
        }
        const uint8_t* base = nullptr;
        size_t size = 0;
        bool ok;
        if (src->mapping(&base, &size)) {
          ok = read_mapped(es, src, base, size, path);
        } else {
          const int fd = src->raw_fd();      // `src` owns the descriptor; the loader takes the bytes itself
          ok = fd >= 0 && read_stream(es, fd, path);
          if (fd < 0) fail(path + ": not a mapped file and not a readable stream");
        }
        if (!ok) return;
      }
      if (!es.bw->recs.empty() && !cfg_.drop_remainder) {
        if (!finish_and_submit(es.bw)) return;
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
    return submit(bw);
  }

  // ---- workers --------------------------------------------------------------------------------
  std::string where(const BatchWork& bw, const Rec& rec) const {
    const FileSeg& sg = bw.segs[rec.seg];
    char buf[64] = "";
    if (sg.base)
      snprintf(buf, sizeof(buf), ", record payload at byte offset %llu", (unsigned long long)(sg.base_off + (rec.p - sg.base)));
    return " (" + sg.path + buf + ")";
  }

  void worker_main() {
    pthread_setname_np(pthread_self(), "dfio-worker");
    const int F = cfg_.field_size;
    const size_t id_sz = cfg_.ids_i32 ? 4 : 8;
    const uint64_t max_id = cfg_.feature_size > 0 ? static_cast<uint64_t>(cfg_.feature_size) : 0;
    for (;;) {
      Task t;
      {
        std::unique_lock<std::mutex> l(mu_);
        cv_task_.wait(l, [&] { return stop_ || failed_ || !parse_tasks_.empty() || !index_tasks_.empty(); });
        if (stop_ || failed_) return;
        // Index ranges first: there are at most `window` of them, each ~70 us, and the batcher can build no further
        // batch until the next range is walked.  Served last, they waited until the parse queue ran dry, and the
        // workers then idled while the batcher caught up -- a bubble every half dozen batches.
        auto& q = !index_tasks_.empty() ? index_tasks_ : parse_tasks_;
        t = std::move(q.front());
        q.pop_front();
      }
      if (t.chunk) {
        index_chunk(*t.chunk);
        {
          std::lock_guard<std::mutex> l(mu_);
          t.chunk->done = true;
        }
        cv_chunk_.notify_all();
        continue;
      }
      Slot& s = slots_[t.bw->slot];
      bool ok = true;
      for (int r = t.row0; r < t.row1 && ok; ++r) {
        const Rec& rec = t.bw->recs[r];
        if (cfg_.verify_crc) {      // the framing sits around the payload, in a mapping and in a stream block alike
          uint32_t dcrc;
          memcpy(&dcrc, rec.p + rec.len, 4);
          if (mask_crc(crc32c(rec.p, rec.len)) != dcrc) {
            fail("TFRecord CRC mismatch (corrupt input)" + where(*t.bw, rec));
            ok = false;
            break;
          }
        }
        if (parse_example(rec.p, rec.len, F, static_cast<uint8_t*>(s.ids) + size_t(r) * F * id_sz,
                          cfg_.ids_i32 != 0, s.vals + size_t(r) * F, s.labels + r, max_id) != 0) {
          fail(std::string(dfio_last_error()) + where(*t.bw, rec));
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
  std::mutex pool_mu_;                                        // declared first: blocks come back here while the
  std::vector<std::unique_ptr<StreamBlock>> free_blocks_;     // members below are being destroyed
  std::vector<Slot> slots_;
  std::mutex mu_;
  std::condition_variable cv_slot_, cv_task_, cv_ready_, cv_chunk_;
  std::deque<Task> parse_tasks_, index_tasks_;
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
