// Pass 2 of the `humid` executable: writeFiltered / writeAnnotated (reference src/humid.cc:203-292) —
// every input file is read again and the kept (or all, annotated) records are written out as
// name + suffix, sequence, strand, quality, each followed by '\n' (Read::toString).
//
// SURVEY.md §8(f) item 2.  The reference parses the whole input a second time on one thread.  Here pass 1
// notes the byte offset of every 65 536th record of a plain (uncompressed) file; pass 2 then cuts the file
// into those chunks and several threads per file parse them side by side, each into its own buffer.  A
// chunk's place in the output is known as soon as the chunks before it have been parsed, so the buffers
// are written with pwrite concurrently.  The bytes written are those of the one-thread pass (same parser,
// same serialisation); compressed inputs (no seeking) and compressed outputs keep the one-thread pass.
#pragma once
#include <fcntl.h>
#include <unistd.h>

#include <atomic>
#include <condition_variable>
#include <mutex>
#include <thread>

#include "fastq_block.h"
#include "fastq_io.h"

namespace humid_host {

#ifndef HUMID_REWRITE_CHUNK
#define HUMID_REWRITE_CHUNK (1u << 16)
#endif
constexpr uint64_t REWRITE_CHUNK = HUMID_REWRITE_CHUNK;  // records between two noted offsets (the tests build with a tiny one)

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

// Several threads: offsets[c] = byte offset of record c * REWRITE_CHUNK (noted in pass 1).
inline void rewrite_file_parallel(const std::string &in_path, const std::string &out_path, uint64_t total,
                                  const uint64_t *keep, const uint32_t *ids, const std::vector<uint64_t> &offsets,
                                  unsigned n_threads) {
  const uint64_t n_chunks = (total + REWRITE_CHUNK - 1) / REWRITE_CHUNK;
  if (ends_with(in_path, ".gz") || ends_with(out_path, ".gz") || n_threads < 2 || n_chunks < 2 ||
      offsets.size() < n_chunks) {
    rewrite_file(in_path, out_path, total, keep, ids);
    return;
  }
  const int fd = open(out_path.c_str(), O_WRONLY | O_CREAT | O_TRUNC, 0666);
  if (fd < 0) error_exit("Failed to write: " + out_path);
  std::atomic<uint64_t> next{0};
  std::mutex mu;
  std::condition_variable cv;
  uint64_t parsed = 0, out_end = 0;  // chunks [0, parsed) have their place; the next one starts at out_end
  auto worker = [&] {
    std::string buf;
    RecordView rec;
    while (true) {
      const uint64_t c = next.fetch_add(1);
      if (c >= n_chunks) return;
      const uint64_t t0 = c * REWRITE_CHUNK, t1 = std::min(total, t0 + REWRITE_CHUNK);
      buf.clear();
      {
        BlockReader reader(in_path, 4u << 20, offsets[c]);
        for (uint64_t t = t0; t < t1 && reader.read(rec); t++) {
          if (keep && !(keep[t >> 6] >> (t & 63) & 1)) continue;
          append_record(buf, rec, ids, t);
        }
      }
      uint64_t at;
      {
        std::unique_lock<std::mutex> lk(mu);
        cv.wait(lk, [&] { return parsed == c; });
        at = out_end;
        out_end += buf.size();
        parsed = c + 1;
      }
      cv.notify_all();
      for (size_t done = 0; done < buf.size();) {
        const ssize_t w = pwrite(fd, buf.data() + done, buf.size() - done, (off_t)(at + done));
        if (w <= 0) error_exit("Failed to write: " + out_path);
        done += (size_t)w;
      }
    }
  };
  std::vector<std::thread> pool;
  for (unsigned i = 0; i < n_threads; i++) pool.emplace_back(worker);
  for (std::thread &t : pool) t.join();
  if (close(fd) != 0) error_exit("Failed to write: " + out_path);
}

}  // namespace humid_host
