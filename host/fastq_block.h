// Block-wise FastQ reading for the `humid` executable: SURVEY.md §8(f) item 1, the host ingest in
// front of the device path.  One BlockReader per input file runs in its own thread
// (host/humid_main.cc); it reads 16 MiB blocks (plain, gzip through host/inflate_fast.h or — several threads on one
// member — host/inflate_parallel.h, or BGZF inflated by several threads: host/bgzf.h), finds line ends with memchr and
// hands out records as views into its buffer — no allocation and no copy per record.
//
// The observable rules are those of host/fastq_io.h's FastqReader (i.e. of fastp's reader as the
// reference uses it, lib/fastp/src/fastqreader.cpp:206-327):
//   * lines end at '\n', '\r' or "\r\n";
//   * before a record, empty lines and lines not starting with '@' are skipped; the name keeps
//     its '@';
//   * a record whose sequence and quality lengths differ is reported on stderr and ends the input;
//   * missing lines at the end of the data read as empty.
#pragma once
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>
#include <zlib.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <memory>
#include <string>
#include <thread>
#include <vector>

#include "bgzf.h"
#include "fastq_io.h"
#include "inflate_fast.h"
#include "inflate_parallel.h"

namespace humid_host {

// Inflate threads of one compressed input (HUMID_INFLATE_THREADS overrides; 1 or 2 = one thread front to back for
// ordinary members).  Without a hint from the program: min(4, cores / 4).
inline unsigned &inflate_threads_hint() {
  static unsigned hint = 0;  // 0: not told how many compressed inputs share the machine
  return hint;
}
inline unsigned default_inflate_threads() {
  if (const char *e = std::getenv("HUMID_INFLATE_THREADS")) return (unsigned)std::max(1, std::atoi(e));
  if (inflate_threads_hint()) return inflate_threads_hint();
  const unsigned hw = std::thread::hardware_concurrency();
  return std::max(1u, std::min(4u, hw / 4));
}
// The program reads n_files inputs side by side: the cores are shared out among them (one parser thread each beside
// the inflate threads), at most 8 inflate threads per file.
inline void share_inflate_threads(size_t n_files) {
  const unsigned hw = std::max(1u, std::thread::hardware_concurrency());
  const unsigned per_file = hw / (unsigned)std::max<size_t>(1, n_files);
  inflate_threads_hint() = per_file > 2 ? std::min(8u, per_file - 1) : 1u;
}

// HUMID_INFLATE=zlib: ordinary .gz files through zlib's gzread instead of host/inflate_fast.h (for comparisons).
inline bool zlib_inflate_wanted() {
  const char *e = std::getenv("HUMID_INFLATE");
  return e && std::string(e) == "zlib";
}

struct Span {
  const char *p = nullptr;
  size_t n = 0;
};

struct RecordView {
  Span name, seq, strand, qual;
};

class BlockReader {
 public:
  // start: byte offset to begin at (plain files only; it must be the start of a line)
  // inflate_threads > 1: a .gz that is BGZF is inflated by that many threads (host/bgzf.h)
  explicit BlockReader(const std::string &path, size_t block = 16u << 20, uint64_t start = 0,
                       unsigned inflate_threads = default_inflate_threads())
      : path_(path), base_(start) {
    // a plain file is parsed where the page cache holds it (no copy into a buffer); small blocks — the tests' way of
    // making every line cross a buffer end — and HUMID_NO_MMAP=1 keep the buffered reads
    if (!ends_with(path, ".gz") && block >= (1u << 20) && !std::getenv("HUMID_NO_MMAP") && map_file(path)) {
      mark_ = (size_t)std::min<uint64_t>(start, end_);
      base_ = 0;
      return;
    }
    buf_.resize(block);
    if (ends_with(path, ".gz") && inflate_threads > 1 && !start && is_bgzf(path)) {
      bgzf_.reset(new BgzfInflater(path, inflate_threads));
    } else if (ends_with(path, ".gz") && !zlib_inflate_wanted()) {
      if (start) error_exit("cannot seek in a compressed file: " + path);
      // three threads or more decode an ordinary member faster than one does front to back (a speculative decode runs at
      // about half the speed of a plain one)
      if (inflate_threads >= 3 && ParallelGzReader::worth_it(path, inflate_threads)) parallel_.reset(new ParallelGzReader(path, inflate_threads));
      else fast_.reset(new GzFastReader(path));
    } else if (ends_with(path, ".gz")) {
      if (start) error_exit("cannot seek in a compressed file: " + path);
      gz_ = gzopen(path.c_str(), "rb");
      if (!gz_) error_exit("Failed to open file: " + path);
      gzbuffer(gz_, 1u << 20);
    } else {
      fp_ = std::fopen(path.c_str(), "rb");
      if (!fp_) error_exit("Failed to open file: " + path);
      if (start && fseeko(fp_, (off_t)start, SEEK_SET) != 0) error_exit("Failed to seek in file: " + path);
    }
  }
  // Over bytes that are in memory already ([mem, mem + size): e.g. a mapping shared by several readers); `start` as above.
  BlockReader(const char *mem, size_t size, uint64_t start, const std::string &path) : path_(path) {
    view_ = mem;
    end_ = size;
    mark_ = (size_t)std::min<uint64_t>(start, size);
    finished_ = drained_ = true;
  }
  ~BlockReader() {
    if (gz_) gzclose(gz_);
    if (fp_) std::fclose(fp_);
    if (map_) munmap(map_, map_size_);
  }
  BlockReader(const BlockReader &) = delete;
  BlockReader &operator=(const BlockReader &) = delete;

  // Byte offset (in the uncompressed data) of the name line of the record read last: a reader started
  // there sees the same records from that one on.
  size_t buffer_bytes() const { return buf_.size(); }
  bool mapped() const { return view_ != nullptr; }
  // The message of a malformed record goes into *to instead of stderr (a reader that runs ahead of the program's order:
  // host/ingest_parallel.h prints it when its turn comes).
  void hold_errors(std::string *to) { held_errors_ = to; }

  // Bytes the record read last takes in the file when it is stored exactly as the executable writes records back —
  // four lines, each ended by one '\n', nothing in between (Read::toString, lib/fastp/src/read.cpp:116-118) — else 0.
  size_t raw_length(const RecordView &r) const {
    const char *e = data() + end_;
    auto ended = [&](const Span &s) { return s.p + s.n < e && s.p[s.n] == '\n'; };
    if (!ended(r.name) || !ended(r.seq) || !ended(r.strand) || !ended(r.qual)) return 0;
    if (r.seq.p != r.name.p + r.name.n + 1 || r.strand.p != r.seq.p + r.seq.n + 1 || r.qual.p != r.strand.p + r.strand.n + 1) return 0;
    return (size_t)(r.qual.p + r.qual.n + 1 - r.name.p);
  }
  uint64_t offset_of(const RecordView &r) const { return base_ + (uint64_t)(r.name.p - data()); }

  // false at end of input (or at a malformed record).  The views stay valid until the next call.
  bool read(RecordView &r) {
    while (true) {
      size_t pos = mark_;
      int st = parse(pos, r);
      if (st == kNeedMore) {
        if (!more()) drained_ = true;  // parse again: lines now end where the data ends
        continue;
      }
      if (st == kEnd) return false;
      mark_ = pos;
      if (r.qual.n != r.seq.n) {
        const std::string what = "ERROR: sequence and quality have different length:\n" + std::string(r.name.p, r.name.n) + "\n" +
                                 std::string(r.seq.p, r.seq.n) + "\n" + std::string(r.strand.p, r.strand.n) + "\n" +
                                 std::string(r.qual.p, r.qual.n) + "\n";
        if (held_errors_) *held_errors_ += what;
        else std::cerr << what << std::flush;
        return false;
      }
      return true;
    }
  }

 private:
  enum { kOk, kNeedMore, kEnd };

  // One line starting at pos.  kNeedMore: nothing in the buffer ends it and the file has more.
  int line(size_t &pos, Span &s) {
    const char *b = data() + pos;
    const size_t avail = end_ - pos;
    // [nl_free_from_, nl_free_to_) is known to hold no '\n' (a file with '\r' line ends only would otherwise be
    // searched to the end of the buffer once per line)
    const size_t from = (pos >= nl_free_from_ && pos < nl_free_to_) ? nl_free_to_ : pos;
    const char *nl = from < end_ ? (const char *)std::memchr(data() + from, '\n', end_ - from) : nullptr;
    if (!nl) {
      if (from == pos) nl_free_from_ = pos;
      nl_free_to_ = end_;
      if (!drained_) {
        // a '\r' with data behind it ends the line where it stands (that byte is not a '\n': there is none); as the
        // last byte in the buffer it may be the first half of "\r\n"
        const char *cr = avail ? (const char *)std::memchr(b, '\r', avail) : nullptr;
        if (!cr || (size_t)(cr - b) + 1 >= avail) return kNeedMore;
        s.p = b;
        s.n = (size_t)(cr - b);
        pos += s.n + 1;
        return kOk;
      }
    }
    size_t len = nl ? (size_t)(nl - b) : avail;
    size_t next = nl ? len + 1 : len;
    if (const char *cr = len ? (const char *)std::memchr(b, '\r', len) : nullptr) {
      size_t at = (size_t)(cr - b);
      if (at + 1 < len || !nl) next = at + 1;  // a lone '\r' ends the line where it stands
      len = at;                                // ("\r\n" keeps next = behind the '\n')
    }
    s.p = b;
    s.n = len;
    pos += next;
    return kOk;
  }

  int parse(size_t &pos, RecordView &r) {
    if (pos >= end_ && drained_) return kEnd;
    // name: skip empty lines while data remains, and lines that do not start with '@'
    while (true) {
      if (int st = line(pos, r.name)) return st;
      bool at_end = pos >= end_ && drained_;
      if ((r.name.n == 0 && !at_end) || (r.name.n != 0 && r.name.p[0] != '@')) {
        if (pos >= end_ && !drained_) return kNeedMore;
        mark_ = pos;  // skipped for good
        continue;
      }
      break;
    }
    if (r.name.n == 0) return kEnd;
    if (int st = line(pos, r.seq)) return st;
    if (int st = line(pos, r.strand)) return st;
    if (int st = line(pos, r.qual)) return st;
    return kOk;
  }

  // Keep the unconsumed tail [mark_, end_) at the front and read behind it.  false: nothing came.
  bool more() {
    if (finished_) return false;
    if (mark_ > 0) {
      std::memmove(buf_.data(), buf_.data() + mark_, end_ - mark_);
      end_ -= mark_;
      base_ += mark_;
      nl_free_from_ = nl_free_from_ > mark_ ? nl_free_from_ - mark_ : 0;
      nl_free_to_ = nl_free_to_ > mark_ ? nl_free_to_ - mark_ : 0;
      mark_ = 0;
    }
    if (end_ == buf_.size()) buf_.resize(buf_.size() * 2);  // one record longer than the block
    size_t room = buf_.size() - end_;
    long got;
    if (bgzf_) {
      got = bgzf_->read(buf_.data() + end_, room);
    } else if (parallel_) {
      got = parallel_->read(buf_.data() + end_, room);
    } else if (fast_) {
      got = fast_->read(buf_.data() + end_, room);
    } else if (gz_) {
      got = gzread(gz_, buf_.data() + end_, (unsigned)std::min<size_t>(room, 1u << 30));
      if (got < 0) error_exit("igzip: encountered while decompressing file: " + path_);
    } else {
      got = (long)std::fread(buf_.data() + end_, 1, room, fp_);
    }
    if (got <= 0) {
      finished_ = true;
      return false;
    }
    end_ += (size_t)got;
    return true;
  }

  const char *data() const { return view_ ? view_ : buf_.data(); }

  // The whole file mapped read-only; false: not a regular file / empty / mmap refused (the caller reads it in blocks).
  bool map_file(const std::string &path) {
    const int fd = open(path.c_str(), O_RDONLY);
    if (fd < 0) error_exit("Failed to open file: " + path);
    struct stat st;
    if (fstat(fd, &st) != 0 || !S_ISREG(st.st_mode) || st.st_size <= 0) {
      close(fd);
      return false;
    }
    void *p = mmap(nullptr, (size_t)st.st_size, PROT_READ, MAP_PRIVATE, fd, 0);
    close(fd);
    if (p == MAP_FAILED) return false;
    madvise(p, (size_t)st.st_size, MADV_SEQUENTIAL);
    map_ = p;
    map_size_ = (size_t)st.st_size;
    view_ = static_cast<const char *>(p);
    end_ = map_size_;
    finished_ = drained_ = true;
    return true;
  }

  std::string path_;
  const char *view_ = nullptr;  // the data when it is not in buf_: the mapping, or the caller's memory
  void *map_ = nullptr;
  size_t map_size_ = 0;
  std::vector<char> buf_;
  uint64_t base_ = 0;          // offset of buf_[0] in the (uncompressed) data
  size_t mark_ = 0, end_ = 0;  // [mark_, end_) = bytes not yet handed out
  size_t nl_free_from_ = 0, nl_free_to_ = 0;
  bool finished_ = false;      // the file has no more bytes
  bool drained_ = false;       // ... and the parser knows: the last line ends at end_
  gzFile gz_ = nullptr;
  FILE *fp_ = nullptr;
  std::string *held_errors_ = nullptr;
  std::unique_ptr<BgzfInflater> bgzf_;
  std::unique_ptr<GzFastReader> fast_;
  std::unique_ptr<ParallelGzReader> parallel_;
};

}  // namespace humid_host
