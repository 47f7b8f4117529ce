// Several threads parsing ONE plain FastQ file — SURVEY.md §8(f) item 1 ("multi-threaded ... it bounds end-to-end
// reads/s long before the GPU does").  One block-parsing thread per file (host/fastq_block.h) reads ~2.5 GB/s; the
// device takes a batch in a fraction of the time the host needs to parse it.
//
// The file is cut into ranges of 32 MiB.  A pool takes the ranges in order:
//   parse    (side by side) a range's first record is guessed — the first line at or behind the range's first byte
//            that starts with '@' and whose second-next line starts with '+' (a quality line may start with '@', but then
//            the second-next line is a sequence) — and the records that START inside the range are parsed with the
//            ordinary BlockReader; per record a fixed-size row (the caller's extract: the word's letters from this
//            file) is kept, with the record's offset and stored length;
//   commit   (one range at a time, in file order) the parser has no state between records but the offset it stands
//            at, so a range's rows are the sequential parser's rows exactly if the range was entered where the
//            range before it says the next record starts.  If it was not (junk between records, '\r' line ends, a
//            record longer than the look-ahead, a wrong guess) the committing thread parses the range again from
//            the right offset.  Rows then go to the caller in record order.
// A malformed record (sequence and quality of different length) ends the file where the sequential reader would
// end it, and its message is printed when — and only if — the commit reaches it.
#pragma once
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <atomic>
#include <condition_variable>
#include <functional>
#include <memory>
#include <mutex>
#include <thread>

#include "fastq_block.h"

namespace humid_host {

class ParallelPlainParser {
 public:
  static constexpr uint64_t kNone = ~0ull;
  struct Range {  // what parsing one range gives
    uint64_t entered = kNone;  // offset of its first record's name line (kNone: no record start found)
    uint64_t next = kNone;     // where the record behind its last one starts (kNone: the file ends / a bad record ended it)
    std::vector<uint8_t> rows;
    std::vector<uint64_t> at;
    std::vector<uint32_t> stored;  // bytes the record takes in the file if stored as written back, else 0
    std::string error;             // message of the malformed record that ended the file inside this range
    bool done = false;
  };
  // extract(record, row): fills row_bytes bytes for one record (runs on the pool's threads, no shared state)
  using Extract = std::function<void(const RecordView &, uint8_t *)>;
  // row(t, row, offset, stored_length): record t of the file, in order (one thread at a time); false = stop reading
  using Row = std::function<bool(uint64_t, const uint8_t *, uint64_t, uint32_t)>;

  ParallelPlainParser(const std::string &path, size_t row_bytes, Extract extract, uint64_t range_bytes = range_size())
      : path_(path), row_bytes_(row_bytes), extract_(std::move(extract)), range_(range_bytes) {
    const int fd = open(path.c_str(), O_RDONLY);
    if (fd < 0) error_exit("Failed to open file: " + path);
    struct stat st;
    if (fstat(fd, &st) != 0) error_exit("Failed to open file: " + path);
    size_ = (uint64_t)st.st_size;
    // one read-only mapping for all the pool's readers: the file is parsed where the page cache holds it
    if (S_ISREG(st.st_mode) && size_ > 0 && !std::getenv("HUMID_NO_MMAP")) {
      void *p = mmap(nullptr, (size_t)size_, PROT_READ, MAP_PRIVATE, fd, 0);
      if (p != MAP_FAILED) map_ = static_cast<const char *>(p);
    }
    close(fd);
    n_ranges_ = std::max<uint64_t>(1, (size_ + range_ - 1) / range_);
  }
  ~ParallelPlainParser() {
    if (map_) munmap(const_cast<char *>(map_), (size_t)size_);
  }
  ParallelPlainParser(const ParallelPlainParser &) = delete;
  ParallelPlainParser &operator=(const ParallelPlainParser &) = delete;
  // HUMID_INGEST_RANGE: bytes per range (the tests use small ones)
  static uint64_t range_size() {
    const char *e = std::getenv("HUMID_INGEST_RANGE");
    const long long v = e ? std::atoll(e) : 0;
    return v >= 64 ? (uint64_t)v : (32ull << 20);
  }

  // commit(first, range, last): the range's records are records first, first + 1, ... of the file; called in file
  //   order, one range at a time — keep it short; last: the file ends with this range; false = stop reading.
  // scatter(first, range): the same range once more, after its turn, side by side with other ranges (may be empty).
  using Commit = std::function<bool(uint64_t, const Range &, bool)>;
  using Scatter = std::function<void(uint64_t, const Range &)>;

  // Parses the file with n_threads threads; returns the number of records accounted for.
  uint64_t run(unsigned n_threads, const Commit &commit, const Scatter &scatter) {
    std::vector<std::thread> pool;
    for (unsigned i = 0; i < std::max(1u, n_threads); i++) pool.emplace_back([&] { work(commit, scatter); });
    for (std::thread &t : pool) t.join();
    return delivered_;
  }
  // One call per record, in order.
  uint64_t run(unsigned n_threads, const Row &row) {
    return run(n_threads, [&](uint64_t first, const Range &r, bool) {
      for (size_t i = 0; i < r.at.size(); i++)
        if (!row(first + i, r.rows.data() + i * row_bytes_, r.at[i], r.stored[i])) return false;
      return true;
    }, nullptr);
  }
  uint64_t reparsed_ranges() const { return reparsed_; }
  uint64_t ranges() const { return n_ranges_; }

 private:
  // First line at or behind `from` that looks like a record's name line; kNone if none within the look-ahead.
  uint64_t guess_start(uint64_t from) const {
    if (from == 0) return 0;
    constexpr size_t kLook = 1u << 20;
    std::vector<char> w;
    const char *b;
    ssize_t got;
    if (map_) {  // one byte earlier: is `from` a line start?
      b = map_ + from - 1;
      got = (ssize_t)std::min<uint64_t>(kLook + 1, size_ - (from - 1));
    } else {
      w.resize(kLook + 1);
      const int fd = open(path_.c_str(), O_RDONLY);
      if (fd < 0) return kNone;
      got = pread(fd, w.data(), kLook + 1, (off_t)(from - 1));
      close(fd);
      b = w.data();
    }
    if (got <= 1) return kNone;
    const char *e = b + got;
    const char *line = b + 1;  // == from, if the byte in front of it ends a line
    if (b[0] != '\n') {
      const char *nl = (const char *)std::memchr(b + 1, '\n', (size_t)(e - b - 1));
      line = nl ? nl + 1 : nullptr;
    }
    while (line && line < e) {
      const char *l1 = (const char *)std::memchr(line, '\n', (size_t)(e - line));
      if (!l1) break;
      const char *l2 = (const char *)std::memchr(l1 + 1, '\n', (size_t)(e - l1 - 1));
      if (!l2 || l2 + 1 >= e) break;
      if (line[0] == '@' && l2[1] == '+') return from - 1 + (uint64_t)(line - b);
      line = l1 + 1;
    }
    return kNone;
  }

  // Records that start in [entered, hi), parsed from `entered`.
  void parse(uint64_t entered, uint64_t hi, Range &r) const {
    r.entered = entered;
    r.next = kNone;
    r.rows.clear();
    r.at.clear();
    r.stored.clear();
    r.error.clear();
    if (entered == kNone || entered >= size_) return;
    std::unique_ptr<BlockReader> own(map_ ? new BlockReader(map_, (size_t)size_, entered, path_) : new BlockReader(path_, 4u << 20, entered));
    BlockReader &reader = *own;
    reader.hold_errors(&r.error);
    RecordView rec;
    while (reader.read(rec)) {
      const uint64_t at = reader.offset_of(rec);
      if (at >= hi) {
        r.next = at;
        return;
      }
      r.at.push_back(at);
      r.stored.push_back((uint32_t)reader.raw_length(rec));
      r.rows.resize(r.rows.size() + row_bytes_);
      extract_(rec, r.rows.data() + r.rows.size() - row_bytes_);
    }
  }

  void work(const Commit &commit, const Scatter &scatter) {
    while (true) {
      uint64_t j;
      {
        std::unique_lock<std::mutex> lk(mu_);
        // not too far ahead of the commit: a range's rows stay in memory until its turn
        cv_.wait(lk, [&] { return stop_ || next_range_ >= n_ranges_ || next_range_ < turn_ + 2 * 8; });
        if (stop_ || next_range_ >= n_ranges_) return;
        j = next_range_++;
      }
      Range r;
      const uint64_t lo = j * range_, hi = std::min(size_, (j + 1) * range_);
      if (!speculation_off_) parse(guess_start(lo), hi, r);
      std::unique_lock<std::mutex> lk(mu_);
      cv_.wait(lk, [&] { return stop_ || turn_ == j; });
      if (stop_) return;
      lk.unlock();
      // ---- this range's turn: the records before it are accounted for, expect_ is where the next one starts ----
      uint64_t first = delivered_;
      bool mine = false;
      if (expect_ != kNone && expect_ < hi) {
        if (r.entered != expect_) {
          if (j > 0) {
            reparsed_++;
            if (++misses_in_a_row_ >= 4) speculation_off_ = true;  // not a file to guess in: the rest is parsed in turn
          }
          parse(expect_, hi, r);
        } else {
          misses_in_a_row_ = 0;
        }
        const bool last = r.next == kNone;
        const bool go_on = commit(first, r, last);
        delivered_ += r.at.size();
        last_signalled_ |= last;
        mine = go_on;
        if (go_on && !r.error.empty()) std::cerr << r.error << std::flush;
        expect_ = go_on ? r.next : kNone;  // (kNone: the file, or what can be read of it, ends in this range)
      } else if (j + 1 == n_ranges_ && expect_ != kNone && !last_signalled_) {
        // the record before reaches to the end of the file (or the file is empty): nothing starts in the last range
        r = Range();
        commit(first, r, true);
        last_signalled_ = true;
        expect_ = kNone;
      }
      lk.lock();
      turn_ = j + 1;
      if (expect_ == kNone) stop_ = true;
      cv_.notify_all();
      lk.unlock();
      if (mine && scatter) scatter(first, r);  // side by side with the other ranges' parsing and scattering
    }
  }

  std::string path_;
  size_t row_bytes_;
  Extract extract_;
  uint64_t range_, size_ = 0, n_ranges_ = 0;
  const char *map_ = nullptr;
  std::mutex mu_;
  std::condition_variable cv_;
  uint64_t next_range_ = 0, turn_ = 0;
  bool stop_ = false;
  // the commit's state (one thread at a time: turn_)
  uint64_t expect_ = 0, delivered_ = 0, reparsed_ = 0;
  unsigned misses_in_a_row_ = 0;
  bool last_signalled_ = false;
  std::atomic<bool> speculation_off_{false};
};

}  // namespace humid_host
