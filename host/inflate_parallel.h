// Several threads inflating ONE ordinary gzip member — SURVEY.md §8(f) item 3, second half.
//
// A deflate stream has no index: a block can start at any bit, and its matches copy from the 32 KiB in front of it.
// The reference (isa-l behind lib/fastp/src/fastqreader.cpp:79-133) and host/inflate_fast.h therefore read a file
// front to back on one thread.  This reader cuts the file into spans of 1 MiB of compressed data and lets a pool
// work on them side by side, in two passes:
//
//   speculate   (any thread, any order) find the first block header at or behind the span's first byte — every bit
//               position is tried: not final, dynamic codes, a complete code-length code, complete literal/length
//               and distance codes (Inflate::probe_dynamic_header) — and decode from there until the first block
//               boundary at or behind the span's end.  The 32 KiB in front are unknown, so the output is 16-bit:
//               32768 place holders stand for them, matches copy place holders like bytes (Inflate::run_marked);
//   chain       (one thread at a time, in file order) the stretch that ends at bit position P is followed by the span
//               whose speculation STARTED at P — anything else (no header found, a false positive, a span that
//               begins with stored or fixed-code blocks) is decoded again from P with the real window, on the
//               spot.  The last 32 KiB of a stretch, place holders replaced, are the next one's window;
//   resolve     (any thread) place holders -> bytes through a 64 KiB table, CRC-32 of the stretch.
//
// read() hands the stretches out in file order and checks every member's CRC-32 (zlib's crc32_combine over the
// stretches) and ISIZE, like the one-thread reader.  Member ends, further members (each starts with a known, empty
// window), trailing garbage and damage behave as in GzFastReader: same bytes, same error exits — wrong guesses
// cost time, never correctness, because a speculation is only used when it starts exactly where the verified
// stretch before it ended and that stretch's window is what resolves it.
#pragma once
#include <fcntl.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <condition_variable>
#include <deque>
#include <memory>
#include <mutex>
#include <thread>

#include "inflate_fast.h"

namespace humid_host {

class ParallelGzReader {
 public:
  static constexpr uint64_t kDefaultSpan = 1u << 20;
  // Files below four spans are not worth the threads.
  static bool worth_it(const std::string &path, unsigned n_threads, uint64_t span = span_bytes()) {
    struct stat st;
    return n_threads > 1 && stat(path.c_str(), &st) == 0 && S_ISREG(st.st_mode) && (uint64_t)st.st_size >= 4 * span;
  }
  // HUMID_INFLATE_SPAN: compressed bytes per span (the tests use small ones)
  static uint64_t span_bytes() {
    const char *e = std::getenv("HUMID_INFLATE_SPAN");
    const long long v = e ? std::atoll(e) : 0;
    return v >= 4096 ? (uint64_t)v : kDefaultSpan;
  }

  ParallelGzReader(const std::string &path, unsigned n_threads) : path_(path), span_(span_bytes()) {
    fd_ = open(path.c_str(), O_RDONLY);
    if (fd_ < 0) error_exit("Failed to open file: " + path);
    struct stat st;
    if (fstat(fd_, &st) != 0) error_exit("Failed to open file: " + path);
    file_size_ = (uint64_t)st.st_size;
    n_spans_ = (file_size_ + span_ - 1) / span_;
    specs_.resize(n_spans_);
    max_ahead_ = 2 * n_threads + 2;
    uint64_t at = 0;
    if (file_size_ == 0) {
      chain_done_ = true;
    } else if (!gzip_header(0, at)) {
      error_exit("igzip: Error invalid gzip header found: " + path_);
    }
    pos_bit_ = 8 * at;
    chain_span_locked_ = chain_span();
    for (unsigned i = 0; i < n_threads; i++) pool_.emplace_back([this] { work(); });
  }
  ~ParallelGzReader() {
    {
      std::lock_guard<std::mutex> lk(mu_);
      stop_ = true;
    }
    cv_.notify_all();
    for (std::thread &t : pool_) t.join();
    close(fd_);
    if (std::getenv("HUMID_INFLATE_STATS"))
      std::cerr << "inflate " << path_ << ": " << n_spans_ << " spans, " << used_ << " speculations used, " << redone_
                << " decoded again" << std::endl;
  }
  ParallelGzReader(const ParallelGzReader &) = delete;
  ParallelGzReader &operator=(const ParallelGzReader &) = delete;

  // Like gzread: up to `room` inflated bytes into dst; 0 at the end of the file.
  long read(char *dst, size_t room) {
    while (true) {
      std::shared_ptr<Segment> seg;
      {
        std::unique_lock<std::mutex> lk(mu_);
        cv_.wait(lk, [&] { return (!segments_.empty() && segments_.front()->ready) || (chain_done_ && segments_.empty()); });
        if (segments_.empty()) return 0;
        seg = segments_.front();
      }
      if (!seg->error.empty()) error_exit(seg->error);
      const size_t n = std::min<size_t>(room, seg->n - seg->taken);
      std::memcpy(dst, seg->bytes.data() + seg->begin + seg->taken, n);
      seg->taken += n;
      if (seg->taken == seg->n) {
        member_crc_ = crc32_combine(member_crc_, seg->crc, (z_off_t)seg->n);
        member_bytes_ += seg->n;
        if (seg->ends_member) {
          if ((uint32_t)member_crc_ != seg->want_crc || (uint32_t)member_bytes_ != seg->want_isize) corrupt("incorrect data check");
          member_crc_ = crc32(0L, Z_NULL, 0);
          member_bytes_ = 0;
        }
        {
          std::lock_guard<std::mutex> lk(mu_);
          segments_.pop_front();
        }
        give_back(spare_bytes_, seg->bytes);
        cv_.notify_all();
      }
      if (n) return (long)n;
    }
  }

 private:
  static constexpr size_t kHistory = 32768;
  static constexpr size_t kMaxElements = 256u << 20;  // output of one stretch: beyond it a speculation is given up

  struct Spec {  // speculative decode of one span
    enum State { kPending, kRunning, kDone, kFailed } state = kPending;
    uint64_t start_bit = 0, end_bit = 0;
    bool stream_end = false;
    std::vector<uint16_t> marked;  // kHistory place holders, then the output
    size_t n = 0;
  };
  struct Segment {  // one stretch of output, in file order
    std::vector<uint8_t> bytes;
    size_t begin = 0, n = 0, taken = 0;
    std::vector<uint16_t> marked;  // still to be resolved with `window`
    std::vector<uint8_t> window;
    bool ready = false, resolving = false;
    unsigned long crc = 0;
    bool ends_member = false;
    uint32_t want_crc = 0, want_isize = 0;
    std::string error;
  };

  // A window of the file in memory, Inflate::kInPad zero bytes behind it.
  struct Input {
    int fd = -1;
    uint64_t file_size = 0, base = 0;
    size_t len = 0;
    std::vector<uint8_t> data;
    bool ok = true;
    void fetch(size_t at, size_t n) {
      for (size_t done = 0; done < n;) {
        const ssize_t got = pread(fd, data.data() + at + done, n - done, (off_t)(base + at + done));
        if (got <= 0) {
          ok = false;
          std::memset(data.data() + at + done, 0, n - done);
          break;
        }
        done += (size_t)got;
      }
    }
    void load(uint64_t off, size_t n) {
      ok = true;
      base = std::min(off, file_size);
      n = (size_t)std::min<uint64_t>(n, file_size - base);
      data.assign(n + Inflate::kInPad, 0);
      len = n;
      fetch(0, n);
    }
    // drop what lies in front of file offset `keep_from`, read `more` bytes behind the rest
    void slide(uint64_t keep_from, size_t more) {
      const size_t drop = (size_t)(keep_from - base);
      std::memmove(data.data(), data.data() + drop, len - drop);
      len -= drop;
      base = keep_from;
      const size_t n = (size_t)std::min<uint64_t>(more, file_size - (base + len));
      data.resize(len + n + Inflate::kInPad);
      fetch(len, n);
      len += n;
      std::memset(data.data() + len, 0, Inflate::kInPad);
    }
    bool eof() const { return base + len >= file_size; }
    const uint8_t *at(uint64_t off) const { return data.data() + (off - base); }
    const uint8_t *end() const { return data.data() + len; }
    uint64_t offset_of(const uint8_t *p) const { return base + (uint64_t)(p - data.data()); }
  };

  [[noreturn]] void corrupt(const std::string &what) {
    error_exit("igzip: encountered while decompressing file: " + path_ + " (" + what + ")");
  }

  // Output buffers are handed round instead of freed and allocated again: tens of megabytes per span going through
  // mmap / munmap and first-touch page faults hold up everything else in the process that maps memory — the CUDA
  // context that is coming up beside the readers most of all (0.3 s alone, up to 1.1 s next to eight inflate threads).
  template <class V>
  void give_back(std::vector<V> &pool, V &buffer) {
    if (buffer.capacity() == 0) return;
    std::lock_guard<std::mutex> lk(pool_mu_);
    if (pool.size() < 2 * max_ahead_ && !no_pool_) pool.push_back(std::move(buffer));
    buffer = V();
  }
  template <class V>
  V take(std::vector<V> &pool) {
    std::lock_guard<std::mutex> lk(pool_mu_);
    if (pool.empty()) return V();
    V v = std::move(pool.back());
    pool.pop_back();
    return v;
  }

  static Inflate::Status step(Inflate &inf, const uint8_t *&ip, const uint8_t *end, bool eof, std::vector<uint8_t> &out, uint8_t *&op,
                              const uint8_t *boundary) {
    return inf.run(ip, end, eof, out.data(), op, out.data() + out.size(), boundary);
  }
  static Inflate::Status step(Inflate &inf, const uint8_t *&ip, const uint8_t *end, bool eof, std::vector<uint16_t> &out, uint16_t *&op,
                              const uint8_t *boundary) {
    return inf.run_marked(ip, end, eof, out.data(), op, out.data() + out.size(), boundary);
  }

  // Blocks from bit position start_bit (a block header) up to the first block boundary at or behind file offset
  // boundary_off, or the end of the deflate stream.  out holds `prefix` elements of history already; the output goes
  // behind them.  false: damaged data (why in `err`).
  template <class T>
  bool decode_blocks(Input &in, Inflate &inf, uint64_t start_bit, uint64_t boundary_off, std::vector<T> &out, size_t prefix,
                     size_t max_elements, uint64_t &end_bit, bool &stream_end, size_t &n_out, std::string &err) {
    if (!in.ok || start_bit / 8 < in.base || start_bit / 8 >= in.base + in.len) in.load(start_bit / 8, (size_t)span_ + (256u << 10));
    const uint8_t *ip = in.at(start_bit / 8);
    inf.start_at(ip, (unsigned)(start_bit & 7));
    if (out.size() < prefix + (1u << 16)) out.resize(prefix + (size_t)(4 * span_) + Inflate::kOutSlack);
    T *op = out.data() + prefix;
    while (true) {
      if (!in.ok) {
        err = "read error";
        return false;
      }
      const uint8_t *boundary;
      if ((size_t)(op - out.data()) - prefix >= max_elements || boundary_off <= in.base) boundary = in.data.data();  // the next one will do
      else if (boundary_off >= in.base + in.len) boundary = in.end() + Inflate::kInPad;                              // not in sight
      else boundary = in.at(boundary_off);
      const Inflate::Status st = step(inf, ip, in.end(), in.eof(), out, op, boundary);
      if (st == Inflate::kError) {
        err = inf.error();
        return false;
      }
      if (st == Inflate::kNeedInput) {
        if (in.eof()) {
          err = "unexpected end of file";
          return false;
        }
        in.slide(in.offset_of(ip), 1u << 20);
        ip = in.data.data();
      } else if (st == Inflate::kNeedOutput) {
        const size_t have = (size_t)(op - out.data());
        if (have - prefix > 2 * max_elements) {
          err = "a block inflates to more than this reader holds";
          return false;
        }
        out.resize(out.size() * 2);
        op = out.data() + have;
      } else {  // kBoundary, kDone
        end_bit = 8 * in.offset_of(ip) - inf.pending_bits();
        stream_end = st == Inflate::kDone;
        n_out = (size_t)(op - out.data()) - prefix;
        return true;
      }
    }
  }

  // ---- speculate ----
  void speculate(uint64_t j, Spec &sp) {
    const uint64_t lo = j * span_, hi = std::min((j + 1) * span_, file_size_);
    Input in;
    in.fd = fd_;
    in.file_size = file_size_;
    Inflate inf;
    int tries = 0;
    sp.marked = take(spare_marked_);
    bool holders = false;
    for (uint64_t bit = 8 * lo; bit < 8 * hi && tries < 6; bit++) {
      if (!in.ok) break;
      if (bit / 8 < in.base || bit / 8 + 1200 > in.base + in.len) {
        if (!(in.eof() && bit / 8 >= in.base && bit / 8 < in.base + in.len)) in.load(bit / 8, (size_t)span_ + (256u << 10));
      }
      const uint8_t *p = in.at(bit / 8);
      if ((size_t)(in.end() - p) < 8) break;  // no block fits any more
      if (!inf.probe_dynamic_header(p, std::min(p + 1100, in.end()), (unsigned)(bit & 7))) continue;
      tries++;
      if (!holders) {
        if (sp.marked.size() < kHistory + (size_t)(4 * span_) + Inflate::kOutSlack) sp.marked.resize(kHistory + (size_t)(4 * span_) + Inflate::kOutSlack);
        for (size_t i = 0; i < kHistory; i++) sp.marked[i] = (uint16_t)(0x8000u | i);
        holders = true;
      }
      std::string err;
      if (decode_blocks(in, inf, bit, hi, sp.marked, kHistory, kMaxElements, sp.end_bit, sp.stream_end, sp.n, err) &&
          (sp.stream_end || sp.end_bit >= 8 * hi)) {
        sp.start_bit = bit;
        return;  // (state set by the caller, under the lock)
      }
    }
    give_back(spare_marked_, sp.marked);
    sp.n = 0;
    sp.start_bit = ~0ull;
  }

  // ---- chain ----
  // RFC 1952 header at file offset `at`; false: not one.  `behind` = first byte of the deflate data.
  bool gzip_header(uint64_t at, uint64_t &behind) {
    std::vector<uint8_t> h;
    for (size_t want = 1u << 12;; want *= 16) {
      const size_t n = (size_t)std::min<uint64_t>(want, file_size_ - at);
      h.resize(n);
      for (size_t done = 0; done < n;) {
        const ssize_t got = pread(fd_, h.data() + done, n - done, (off_t)(at + done));
        if (got <= 0) return false;
        done += (size_t)got;
      }
      size_t p = 10;
      bool cut = n < p;
      if (!cut) {
        if (h[0] != 0x1f || h[1] != 0x8b || h[2] != 8 || (h[3] & 0xe0)) return false;
        const uint8_t flg = h[3];
        if (flg & 4) {
          if (n < p + 2) cut = true;
          else p += 2 + (size_t)(h[p] | (h[p + 1] << 8));
        }
        for (int field = 0; field < 2 && !cut; field++) {
          if (!(flg & (field ? 16 : 8))) continue;
          const void *z = p < n ? std::memchr(h.data() + p, 0, n - p) : nullptr;
          if (!z) cut = true;
          else p = (size_t)((const uint8_t *)z - h.data()) + 1;
        }
        if (!cut && (flg & 2)) p += 2;
        if (p > n) cut = true;
      }
      if (!cut) {
        behind = at + p;
        return true;
      }
      if (n == file_size_ - at || want > (1u << 20)) return false;  // cut by the end of the file / absurdly long
    }
  }

  uint64_t chain_span() const { return spec_valid_ ? next_spec_ : pos_bit_ / 8 / span_ + 1; }

  void publish(const std::shared_ptr<Segment> &seg) {
    std::lock_guard<std::mutex> lk(mu_);
    segments_.push_back(seg);
  }
  void chain_error(const std::string &what) {
    auto seg = std::make_shared<Segment>();
    seg->error = what;
    seg->ready = true;
    publish(seg);
    std::lock_guard<std::mutex> lk(mu_);
    chain_done_ = true;
  }

  // The deflate stream ended at pos_bit_: trailer, then another member or the end of the file.
  void member_end(Segment &last) {
    const uint64_t at = (pos_bit_ + 7) / 8;
    uint8_t t[8];
    if (at + 8 > file_size_ || pread(fd_, t, 8, (off_t)at) != 8) {
      last.error = "igzip: encountered while decompressing file: " + path_ + " (unexpected end of file)";
      finished_ = true;
      return;
    }
    last.ends_member = true;
    last.want_crc = t[0] | (t[1] << 8) | (t[2] << 16) | ((uint32_t)t[3] << 24);
    last.want_isize = t[4] | (t[5] << 8) | (t[6] << 16) | ((uint32_t)t[7] << 24);
    if (at + 8 == file_size_) {
      finished_ = true;
      return;
    }
    uint64_t behind = 0;
    if (!gzip_header(at + 8, behind)) {
      pending_error_ = "igzip: invalid gzip header found";
      finished_ = true;
      return;
    }
    pos_bit_ = 8 * behind;
    window_.clear();
    spec_valid_ = false;
    members_started_++;
  }

  // From pos_bit_ with the real window up to the first block boundary at or behind boundary_off.
  void exact(uint64_t boundary_off) {
    auto seg = std::make_shared<Segment>();
    seg->bytes = take(spare_bytes_);
    if (seg->bytes.size() < window_.size()) seg->bytes.resize(window_.size());
    std::copy(window_.begin(), window_.end(), seg->bytes.begin());
    seg->begin = window_.size();
    uint64_t end_bit = 0;
    bool stream_end = false;
    std::string err;
    chain_input_.fd = fd_;
    chain_input_.file_size = file_size_;
    if (seg->bytes.size() < seg->begin + (size_t)(4 * span_) + Inflate::kOutSlack) seg->bytes.resize(seg->begin + (size_t)(4 * span_) + Inflate::kOutSlack);
    if (!decode_blocks(chain_input_, chain_inflate_, pos_bit_, boundary_off, seg->bytes, seg->begin, kMaxElements, end_bit, stream_end,
                       seg->n, err)) {
      chain_error("igzip: encountered while decompressing file: " + path_ + " (" + err + ")");
      finished_ = true;
      return;
    }
    pos_bit_ = end_bit;
    const uint8_t *first = seg->bytes.data() + seg->begin;
    seg->crc = crc32(crc32(0L, Z_NULL, 0), first, (uInt)0);
    for (size_t done = 0; done < seg->n;) {  // (uInt lengths)
      const size_t n = std::min<size_t>(seg->n - done, 1u << 30);
      seg->crc = crc32(seg->crc, first + done, (uInt)n);
      done += n;
    }
    const size_t keep = std::min(kHistory, seg->begin + seg->n);
    window_.assign(first + seg->n - keep, first + seg->n);
    seg->ready = true;
    if (stream_end) member_end(*seg);
    publish(seg);
  }

  static uint8_t resolved(uint16_t v, const std::vector<uint8_t> &window) {
    if (v < 0x8000u) return (uint8_t)v;
    const size_t i = v & 0x7FFFu, gap = kHistory - window.size();
    return i >= gap ? window[i - gap] : 0;  // (in front of the member's first byte: damaged data, the CRC will tell)
  }

  // A speculation that starts where the chain stands.
  void accept(Spec &sp) {
    auto seg = std::make_shared<Segment>();
    seg->marked = std::move(sp.marked);
    seg->n = sp.n;
    seg->window = window_;
    // the next window: the last 32 KiB of (window ++ this stretch)
    std::vector<uint8_t> next;
    const size_t from_new = std::min(kHistory, seg->n), from_old = std::min(kHistory - from_new, window_.size());
    next.reserve(from_old + from_new);
    next.insert(next.end(), window_.end() - from_old, window_.end());
    const uint16_t *tail = seg->marked.data() + kHistory + seg->n - from_new;
    for (size_t i = 0; i < from_new; i++) next.push_back(resolved(tail[i], window_));
    window_.swap(next);
    pos_bit_ = sp.end_bit;
    if (sp.stream_end) member_end(*seg);
    publish(seg);
  }

  // One step of the chain.  false: it has to wait for a speculation that is still running.
  bool chain_step() {
    if (finished_) {
      if (!pending_error_.empty()) chain_error(pending_error_);
      std::lock_guard<std::mutex> lk(mu_);
      chain_done_ = true;
      return true;
    }
    const uint64_t members = members_started_;
    if (!spec_valid_) {  // behind a gzip header (or a stretch cut short): to the next span's first boundary with the real window
      const uint64_t j = pos_bit_ / 8 / span_ + 1;
      exact(j * span_);
      if (!finished_ && members == members_started_ && pos_bit_ >= 8 * j * span_) {
        next_spec_ = j;
        spec_valid_ = true;
      }
      return true;
    }
    if (next_spec_ >= n_spans_) {  // the file's last blocks: nothing speculative lies behind them
      exact(~0ull);
      return true;
    }
    Spec &sp = specs_[next_spec_];
    {
      std::lock_guard<std::mutex> lk(mu_);
      if (sp.state == Spec::kPending || sp.state == Spec::kRunning) return false;
    }
    const uint64_t want = 8 * (next_spec_ + 1) * span_;
    if (sp.state == Spec::kDone && sp.start_bit == pos_bit_) {
      accept(sp);
      used_++;
    } else {
      const uint64_t before = pos_bit_;
      exact((next_spec_ + 1) * span_);
      if (pos_bit_ != before) redone_++;  // (a block that reaches over the whole span leaves nothing to decode)
    }
    give_back(spare_marked_, sp.marked);
    if (!finished_ && members == members_started_) {
      if (pos_bit_ >= want) next_spec_++;
      else spec_valid_ = false;
    }
    return true;
  }

  // ---- resolve ----
  void resolve(Segment &seg) {
    std::vector<uint8_t> lut(65536, 0);
    for (unsigned v = 0; v < 256; v++) lut[v] = (uint8_t)v;
    const size_t gap = kHistory - seg.window.size();
    for (size_t i = gap; i < kHistory; i++) lut[0x8000u + i] = seg.window[i - gap];
    seg.bytes = take(spare_bytes_);
    if (seg.bytes.size() < seg.n) seg.bytes.resize(seg.n);
    const uint16_t *from = seg.marked.data() + kHistory;
    uint8_t *to = seg.bytes.data();
    for (size_t i = 0; i < seg.n; i++) to[i] = lut[from[i]];
    seg.begin = 0;
    seg.crc = crc32(0L, Z_NULL, 0);
    for (size_t done = 0; done < seg.n;) {
      const size_t n = std::min<size_t>(seg.n - done, 1u << 30);
      seg.crc = crc32(seg.crc, to + done, (uInt)n);
      done += n;
    }
    give_back(spare_marked_, seg.marked);
    seg.window = std::vector<uint8_t>();
  }

  // ---- the pool ----
  void work() {
    std::unique_lock<std::mutex> lk(mu_);
    while (!stop_) {
      std::shared_ptr<Segment> todo;
      for (const auto &seg : segments_)
        if (!seg->ready && !seg->resolving) {
          todo = seg;
          break;
        }
      if (todo) {
        todo->resolving = true;
        lk.unlock();
        resolve(*todo);
        lk.lock();
        todo->ready = true;
        cv_.notify_all();
        continue;
      }
      if (!chain_busy_ && !chain_done_ && segments_.size() < max_ahead_) {
        chain_busy_ = true;
        lk.unlock();
        const bool moved = chain_step();
        lk.lock();
        chain_busy_ = false;
        chain_span_locked_ = chain_span();
        for (; freed_ < chain_span_locked_ && freed_ < n_spans_; freed_++)  // speculations the chain has passed
          if (specs_[freed_].state == Spec::kDone || specs_[freed_].state == Spec::kFailed) give_back(spare_marked_, specs_[freed_].marked);
        if (moved) {
          cv_.notify_all();
          continue;
        }
      }
      if (!chain_done_) {
        const uint64_t first = chain_span_locked_;
        const uint64_t j = std::max(next_speculation_, first);
        if (j < n_spans_ && j < first + max_ahead_) {
          next_speculation_ = j + 1;
          specs_[j].state = Spec::kRunning;
          lk.unlock();
          speculate(j, specs_[j]);
          lk.lock();
          specs_[j].state = specs_[j].start_bit == ~0ull ? Spec::kFailed : Spec::kDone;
          if (j < freed_) give_back(spare_marked_, specs_[j].marked);  // too late: the chain has gone past
          cv_.notify_all();
          continue;
        }
      }
      cv_.wait(lk);
    }
  }

  std::string path_;
  int fd_ = -1;
  uint64_t file_size_ = 0, span_, n_spans_ = 0;
  std::mutex mu_;
  std::condition_variable cv_;
  std::vector<std::thread> pool_;
  bool stop_ = false;

  // under mu_
  std::vector<Spec> specs_;  // (a running speculation is written by its thread alone; state changes under mu_)
  std::deque<std::shared_ptr<Segment>> segments_;
  uint64_t next_speculation_ = 1, chain_span_locked_ = 1, freed_ = 0;
  size_t max_ahead_ = 4;
  bool chain_busy_ = false, chain_done_ = false;

  // the chain (one thread at a time: chain_busy_)
  uint64_t pos_bit_ = 0;         // a block header of the current member starts here
  std::vector<uint8_t> window_;  // the member's last <= 32 KiB of output in front of it
  bool spec_valid_ = false;      // pos_bit_ is the first block boundary at or behind span next_spec_'s first byte
  uint64_t next_spec_ = 0;
  bool finished_ = false;
  uint64_t members_started_ = 1;
  std::string pending_error_;
  Input chain_input_;
  Inflate chain_inflate_;
  uint64_t used_ = 0, redone_ = 0;

  const bool no_pool_ = std::getenv("HUMID_INFLATE_NO_POOL") != nullptr;  // (comparisons)
  std::mutex pool_mu_;
  std::vector<std::vector<uint16_t>> spare_marked_;
  std::vector<std::vector<uint8_t>> spare_bytes_;

  // the reader
  unsigned long member_crc_ = crc32(0L, Z_NULL, 0);
  uint64_t member_bytes_ = 0;
};

}  // namespace humid_host
