// Pass 2 of the `humid` executable: writeFiltered / writeAnnotated (reference src/humid.cc:203-292) —
// the kept (or all, annotated) records are written out as name + suffix, sequence, strand, quality, each
// followed by '\n' (Read::toString).
//
// SURVEY.md §8(f) item 2.  The reference opens and parses the whole input a second time on one thread
// (src/humid.cc:220, :268).  Here pass 1 leaves a RecordMap per plain (uncompressed) file: the byte offset of
// every 65 536th record, the length in bytes of every record, and per chunk of 65 536 records whether all of
// them are stored exactly as they would be written back.  Pass 2 cuts the file into those chunks; several
// threads per file take them side by side.  A chunk stored as written is not parsed again: it is read with
// one pread and the kept records — runs of them as one copy — are gathered by their lengths (annotated
// output: the id goes in behind the name line's end, the only byte that is looked for).  Any other chunk
// (line ends with '\r', junk between records, a last line without '\n', a record of 64 KiB or more) goes
// through the parser as before.  A chunk's place in the output is known as soon as the chunks before it are
// done, so the buffers are written with pwrite concurrently.  The bytes written are those of the one-thread
// pass either way.
//
// Compressed files: a .gz input cannot be entered in the middle, so one thread inflates and parses it and hands
// the text of every chunk to the pool; a .gz output is written as one gzip member per chunk (the reference's
// writer produces multi-member files too, lib/fastp/src/writer.cpp:105-128), the members deflated side by
// side at the reference's level 4 — the inflated bytes are those of the one-thread pass.
#pragma once
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <zlib.h>

#include <atomic>
#include <condition_variable>
#include <deque>
#include <mutex>
#include <thread>

#include "fastq_block.h"
#include "fastq_io.h"

namespace humid_host {

#ifndef HUMID_REWRITE_CHUNK
#define HUMID_REWRITE_CHUNK (1u << 16)
#endif
constexpr uint64_t REWRITE_CHUNK = HUMID_REWRITE_CHUNK;  // records between two noted offsets (the tests build with a tiny one)

// What pass 1 notes about one input file for pass 2.
struct RecordMap {
  std::vector<uint64_t> chunk_offset;   // byte offset of record c * REWRITE_CHUNK
  std::vector<uint16_t> raw_length;     // per record (only when `lengths`)
  std::vector<uint8_t> chunk_verbatim;  // per chunk: every record stored as written back, one behind the other
  bool lengths = false;                 // plain files: keep raw_length / chunk_verbatim
  uint64_t expect_ = 0;
  // record t (0, 1, 2 ... in order) has just been read
  void note(uint64_t t, const BlockReader &reader, const RecordView &rec) {
    note_values(t, reader.offset_of(rec), lengths ? reader.raw_length(rec) : 0);
  }
  // the same from what a parser thread kept: offset of the name line, BlockReader::raw_length
  void note_values(uint64_t t, uint64_t at, size_t len) {
    if (t % REWRITE_CHUNK == 0) {
      chunk_offset.push_back(at);
      if (lengths) chunk_verbatim.push_back(1);
      expect_ = at;
    }
    if (!lengths) return;
    if (len == 0 || len > 0xFFFFu || at != expect_) chunk_verbatim.back() = 0;
    raw_length.push_back((uint16_t)len);
    expect_ = at + len;
  }
};

inline void append_record(std::string &out, const RecordView &rec, const uint32_t *ids, uint64_t t) {
  out.append(rec.name.p, rec.name.n);
  if (ids) out.append(":").append(std::to_string(ids[t]));
  out.push_back('\n');
  out.append(rec.seq.p, rec.seq.n).push_back('\n');
  out.append(rec.strand.p, rec.strand.n).push_back('\n');
  out.append(rec.qual.p, rec.qual.n).push_back('\n');
}

// One thread: tuples [0, total) of the file; keep == nullptr keeps everything, ids != nullptr annotates.
inline void rewrite_file(const std::string &in_path, const std::string &out_path, uint64_t total, const uint64_t *keep,
                         const uint32_t *ids) {
  BlockReader reader(in_path);
  Writer out(out_path);
  RecordView rec;
  std::string buf;
  for (uint64_t t = 0; t < total && reader.read(rec); t++) {
    if (keep && !(keep[t >> 6] >> (t & 63) & 1)) continue;
    buf.clear();
    append_record(buf, rec, ids, t);
    out.write(buf.data(), buf.size());
  }
}

// One gzip member holding `in` (level 4: lib/fastp/src/options.cpp:15).
inline void gzip_member(const std::string &in, std::string &out) {
  z_stream zs{};
  if (deflateInit2(&zs, 4, Z_DEFLATED, 15 + 16, 8, Z_DEFAULT_STRATEGY) != Z_OK) error_exit("deflateInit2 failed");
  out.resize(deflateBound(&zs, in.size()) + 32);
  zs.next_in = reinterpret_cast<Bytef *>(const_cast<char *>(in.data()));
  zs.avail_in = (uInt)in.size();
  zs.next_out = reinterpret_cast<Bytef *>(&out[0]);
  zs.avail_out = (uInt)out.size();
  if (deflate(&zs, Z_FINISH) != Z_STREAM_END) error_exit("deflate failed");
  out.resize(zs.total_out);
  deflateEnd(&zs);
}

// Output file whose pieces arrive from several threads: piece c goes behind pieces [0, c).  A piece's place is
// known once the pieces before it have been placed, so only the bookkeeping is ordered, not the writes.
class OrderedOutput {
 public:
  explicit OrderedOutput(const std::string &path) : path_(path), gz_(ends_with(path, ".gz")) {
    fd_ = open(path.c_str(), O_WRONLY | O_CREAT | O_TRUNC, 0666);
    if (fd_ < 0) error_exit("Failed to write: " + path);
  }
  OrderedOutput(const OrderedOutput &) = delete;
  OrderedOutput &operator=(const OrderedOutput &) = delete;
  // text of piece c (deflated here when the file is a .gz); `scratch` is the caller's reusable buffer
  void put(uint64_t c, const std::string &text, std::string &scratch) {
    const std::string *data = &text;
    if (gz_) {
      if (text.empty()) scratch.clear();  // no empty members in the middle of the file
      else gzip_member(text, scratch);
      data = &scratch;
    }
    uint64_t at;
    {
      std::unique_lock<std::mutex> lk(mu_);
      cv_.wait(lk, [&] { return placed_ == c; });
      at = end_;
      end_ += data->size();
      placed_ = c + 1;
    }
    cv_.notify_all();
    write_at(*data, at);
  }
  void finish() {  // after every put() has returned
    if (gz_ && end_ == 0) {  // an empty .gz is one empty member, as gzclose would leave it
      std::string member;
      gzip_member(std::string(), member);
      write_at(member, 0);
    }
    if (close(fd_) != 0) error_exit("Failed to write: " + path_);
  }

 private:
  void write_at(const std::string &data, uint64_t at) {
    for (size_t done = 0; done < data.size();) {
      const ssize_t w = pwrite(fd_, data.data() + done, data.size() - done, (off_t)(at + done));
      if (w <= 0) error_exit("Failed to write: " + path_);
      done += (size_t)w;
    }
  }
  std::string path_;
  bool gz_;
  int fd_;
  std::mutex mu_;
  std::condition_variable cv_;
  uint64_t placed_ = 0, end_ = 0;  // pieces [0, placed_) have their place; the next one starts at end_
};

// A .gz on either side.  Seekable input: the chunk workers below parse and deflate.  .gz input: this thread
// inflates + parses, the pool deflates (or just writes) the chunks it is handed; at most 2 chunks per pool
// thread are in flight.
inline void rewrite_file_piped(const std::string &in_path, const std::string &out_path, uint64_t total,
                               const uint64_t *keep, const uint32_t *ids, unsigned n_threads) {
  OrderedOutput out(out_path);
  std::mutex mu;
  std::condition_variable cv;
  std::deque<std::pair<uint64_t, std::string>> queue;
  std::vector<std::string> spare;
  bool closed = false;
  const size_t max_queued = 2 * n_threads;
  auto worker = [&] {
    std::string scratch;
    while (true) {
      std::pair<uint64_t, std::string> job;
      {
        std::unique_lock<std::mutex> lk(mu);
        cv.wait(lk, [&] { return closed || !queue.empty(); });
        if (queue.empty()) return;
        job = std::move(queue.front());
        queue.pop_front();
      }
      cv.notify_all();
      out.put(job.first, job.second, scratch);
      job.second.clear();
      std::lock_guard<std::mutex> lk(mu);
      spare.push_back(std::move(job.second));  // keeps its capacity for a later chunk
    }
  };
  std::vector<std::thread> pool;
  for (unsigned i = 0; i < n_threads; i++) pool.emplace_back(worker);
  {
    BlockReader reader(in_path);
    RecordView rec;
    std::string buf;
    uint64_t c = 0;
    auto hand_over = [&] {
      std::unique_lock<std::mutex> lk(mu);
      cv.wait(lk, [&] { return queue.size() < max_queued; });
      queue.emplace_back(c++, std::move(buf));
      if (!spare.empty()) {
        buf = std::move(spare.back());
        spare.pop_back();
      } else {
        buf = std::string();
      }
      lk.unlock();
      cv.notify_all();
    };
    for (uint64_t t = 0; t < total && reader.read(rec); t++) {
      if (!keep || (keep[t >> 6] >> (t & 63) & 1)) append_record(buf, rec, ids, t);
      if ((t + 1) % REWRITE_CHUNK == 0) hand_over();
    }
    if (!buf.empty()) hand_over();
  }
  {
    std::lock_guard<std::mutex> lk(mu);
    closed = true;
  }
  cv.notify_all();
  for (std::thread &t : pool) t.join();
  out.finish();
}

// Records [t0, t1) of a chunk that is stored as written back: `raw` holds the chunk's bytes, the records lie one
// behind the other with the lengths noted in pass 1.  No line of it is parsed.
inline void gather_records(const char *raw, const uint16_t *len, uint64_t t0, uint64_t t1, const uint64_t *keep, const uint32_t *ids,
                           std::string &out) {
  size_t at = 0;
  if (ids) {  // every record, ":<id>" behind its name
    for (uint64_t t = t0; t < t1; t++) {
      const size_t n = len[t - t0];
      const char *nl = (const char *)std::memchr(raw + at, '\n', n);
      const size_t name = (size_t)(nl - (raw + at));
      out.append(raw + at, name).append(":").append(std::to_string(ids[t]));
      out.append(raw + at + name, n - name);
      at += n;
    }
    return;
  }
  size_t run_from = 0, run = 0;  // kept records that follow one another leave as one copy
  for (uint64_t t = t0; t < t1; t++) {
    const size_t n = len[t - t0];
    if (!keep || (keep[t >> 6] >> (t & 63) & 1)) {
      if (!run) run_from = at;
      run += n;
    } else if (run) {
      out.append(raw + run_from, run);
      run = 0;
    }
    at += n;
  }
  if (run) out.append(raw + run_from, run);
}

// Several threads over the chunks noted in pass 1.
inline void rewrite_file_parallel(const std::string &in_path, const std::string &out_path, uint64_t total,
                                  const uint64_t *keep, const uint32_t *ids, const RecordMap &map, unsigned n_threads) {
  const std::vector<uint64_t> &offsets = map.chunk_offset;
  const uint64_t n_chunks = (total + REWRITE_CHUNK - 1) / REWRITE_CHUNK;
  const bool gather = map.lengths && !ends_with(in_path, ".gz") && offsets.size() >= n_chunks && map.raw_length.size() >= total &&
                      !std::getenv("HUMID_REWRITE_PARSE");  // (set: every chunk through the parser, for comparisons)
  if ((n_threads < 2 || n_chunks < 2) && !gather) {
    rewrite_file(in_path, out_path, total, keep, ids);
    return;
  }
  if (ends_with(in_path, ".gz") || offsets.size() < n_chunks) {
    rewrite_file_piped(in_path, out_path, total, keep, ids, n_threads);
    return;
  }
  OrderedOutput out(out_path);
  std::atomic<uint64_t> next{0};
  int fd = -1;
  const char *mapped = nullptr;  // the records are gathered out of the page cache when the file can be mapped
  size_t mapped_bytes = 0;
  if (gather) {
    fd = open(in_path.c_str(), O_RDONLY);
    if (fd < 0) error_exit("Failed to open file: " + in_path);
    struct stat st;
    if (fstat(fd, &st) == 0 && S_ISREG(st.st_mode) && st.st_size > 0 && !std::getenv("HUMID_NO_MMAP")) {
      void *p = mmap(nullptr, (size_t)st.st_size, PROT_READ, MAP_PRIVATE, fd, 0);
      if (p != MAP_FAILED) {
        mapped = static_cast<const char *>(p);
        mapped_bytes = (size_t)st.st_size;
      }
    }
  }
  auto worker = [&] {
    std::string buf, scratch;
    std::vector<char> raw;
    RecordView rec;
    while (true) {
      const uint64_t c = next.fetch_add(1);
      if (c >= n_chunks) return;
      const uint64_t t0 = c * REWRITE_CHUNK, t1 = std::min(total, t0 + REWRITE_CHUNK);
      buf.clear();
      if (gather && map.chunk_verbatim[c]) {
        // about 1 MiB of records at a time: read and gathered while they are still in the cache
        uint64_t at = offsets[c];
        for (uint64_t t = t0; t < t1;) {
          size_t bytes = 0;
          uint64_t upto = t;
          while (upto < t1 && bytes < (1u << 20)) bytes += map.raw_length[upto++];
          const char *text = mapped + at;
          if (!mapped || at + bytes > mapped_bytes) {
            if (raw.size() < bytes) raw.resize(bytes);
            for (size_t done = 0; done < bytes;) {
              const ssize_t got = pread(fd, raw.data() + done, bytes - done, (off_t)(at + done));
              if (got <= 0) error_exit("Failed to read: " + in_path);
              done += (size_t)got;
            }
            text = raw.data();
          }
          gather_records(text, map.raw_length.data() + t, t, upto, keep, ids, buf);
          at += bytes;
          t = upto;
        }
      } else {
        BlockReader reader(in_path, 4u << 20, offsets[c]);
        for (uint64_t t = t0; t < t1 && reader.read(rec); t++) {
          if (keep && !(keep[t >> 6] >> (t & 63) & 1)) continue;
          append_record(buf, rec, ids, t);
        }
      }
      out.put(c, buf, scratch);
    }
  };
  std::vector<std::thread> pool;
  for (unsigned i = 0; i < std::max(1u, n_threads); i++) pool.emplace_back(worker);
  for (std::thread &t : pool) t.join();
  if (mapped) munmap(const_cast<char *>(mapped), mapped_bytes);
  if (fd >= 0) close(fd);
  out.finish();
}

}  // namespace humid_host
