// Parallel inflate of BGZF files (bgzip / htslib "blocked gzip": a series of gzip members of at most 64 KiB, each
// carrying its own compressed size in a 'BC' extra field) — SURVEY.md §8(f) item 3.
//
// The reference reads every .gz through one inflate stream per file (lib/fastp/src/fastqreader.cpp:79-133, which
// already walks multi-member files member by member).  A BGZF file can be cut without inflating it: the size
// in the header leads from one member to the next, and the last four bytes of a member give its inflated size.
// So one feeder thread walks the members and packs them into jobs of a few MiB, a pool inflates the jobs side by
// side with this repo's decoder (host/inflate_fast.h; every member into its known place of the job's output, CRC-32 checked as
// gzread would), and read()
// hands the outputs out in file order.  Anything that is not BGZF keeps the one-stream path (host/fastq_block.h).
#pragma once
#include <zlib.h>

#include <condition_variable>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <deque>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "fastq_io.h"
#include "inflate_fast.h"

namespace humid_host {

// EVERY member of the file: gzip magic, deflate, FEXTRA set, a 'B','C' subfield of length 2 giving the member's size,
// and at most 64 KiB of data (the BGZF bound; a trailer claiming more is a corrupt or foreign member).  A file that
// starts with BGZF members and goes on with ordinary gzip members (`cat a.bgz b.gz`) is valid gzip — the reference
// reads it (lib/fastp/src/fastqreader.cpp:79-133) — but cannot be cut at member headers: it, like any file with a
// member that fails the test, is left to the one-stream reader.  One small read + seek per member.
inline bool is_bgzf(const std::string &path) {
  FILE *fp = std::fopen(path.c_str(), "rb");
  if (!fp) return false;
  bool yes = false;
  std::vector<unsigned char> extra;
  while (true) {
    unsigned char h[12];
    const size_t got = std::fread(h, 1, 12, fp);
    if (got == 0) break;  // clean end of file after >= 1 member (or an empty file: yes stays false)
    yes = false;
    if (got != 12 || h[0] != 0x1f || h[1] != 0x8b || h[2] != 8 || !(h[3] & 4)) break;
    const unsigned xlen = h[10] | (h[11] << 8);
    extra.resize(xlen);
    if (xlen < 6 || std::fread(extra.data(), 1, xlen, fp) != xlen) break;
    long bsize = -1;
    for (unsigned i = 0; i + 4 <= xlen;) {
      const unsigned slen = extra[i + 2] | (extra[i + 3] << 8);
      if (extra[i] == 'B' && extra[i + 1] == 'C' && slen == 2 && i + 6 <= xlen) bsize = extra[i + 4] | (extra[i + 5] << 8);
      i += 4 + slen;
    }
    const long rest = bsize + 1 - 12 - (long)xlen;  // deflate data + CRC-32 + ISIZE
    unsigned char t[4];
    if (bsize < 0 || rest < 8 || std::fseek(fp, rest - 4, SEEK_CUR) != 0 || std::fread(t, 1, 4, fp) != 4) break;
    const uint32_t isize = t[0] | (t[1] << 8) | (t[2] << 16) | ((uint32_t)t[3] << 24);
    if (isize > 65536u) break;
    yes = true;
  }
  std::fclose(fp);
  return yes;
}

class BgzfInflater {
 public:
  BgzfInflater(const std::string &path, unsigned n_threads) : path_(path) {
    fp_ = std::fopen(path.c_str(), "rb");
    if (!fp_) error_exit("Failed to open file: " + path);
    std::setvbuf(fp_, nullptr, _IOFBF, 4u << 20);
    max_jobs_ = 2 * n_threads + 2;
    feeder_ = std::thread([this] { feed(); });
    for (unsigned i = 0; i < n_threads; i++) pool_.emplace_back([this] { work(); });
  }
  ~BgzfInflater() {
    {
      std::lock_guard<std::mutex> lk(mu_);
      stop_ = true;
    }
    cv_.notify_all();
    feeder_.join();
    for (std::thread &t : pool_) t.join();
    std::fclose(fp_);
  }
  BgzfInflater(const BgzfInflater &) = delete;
  BgzfInflater &operator=(const BgzfInflater &) = delete;

  // Like gzread: up to `room` inflated bytes into dst; 0 at the end of the file.
  long read(char *dst, size_t room) {
    size_t got = 0;
    while (got == 0) {
      std::shared_ptr<Job> job;
      {
        std::unique_lock<std::mutex> lk(mu_);
        cv_.wait(lk, [&] { return !error_.empty() || (!jobs_.empty() && jobs_.front()->done) || (fed_all_ && jobs_.empty()); });
        if (!error_.empty()) error_exit(error_);
        if (jobs_.empty()) return 0;
        job = jobs_.front();
      }
      const size_t n = std::min(room, job->out_bytes - job->taken);
      std::memcpy(dst, job->out.data() + job->taken, n);
      job->taken += n;
      got = n;
      if (job->taken == job->out_bytes) {
        {
          std::lock_guard<std::mutex> lk(mu_);
          jobs_.pop_front();
          claimed_--;
        }
        cv_.notify_all();
      }
    }
    return (long)got;
  }

 private:
  struct Member {
    size_t raw_at, raw_len;  // deflate data inside Job::raw
    size_t out_at;
    uint32_t isize, crc;
  };
  struct Job {
    std::vector<unsigned char> raw;
    std::vector<Member> members;
    std::vector<char> out;
    size_t out_bytes = 0;
    size_t taken = 0;
    bool done = false;
  };
  static constexpr size_t JOB_BYTES = 4u << 20;  // inflated bytes per job

  void fail(const std::string &what) {
    {
      std::lock_guard<std::mutex> lk(mu_);
      if (error_.empty()) error_ = what;
    }
    cv_.notify_all();
  }

  // Walks the members; false at the end of the file (or on an error, which is recorded).
  bool next_member(Job &job, size_t &out_bytes) {
    unsigned char h[12];
    const size_t got = std::fread(h, 1, 12, fp_);
    if (got == 0) return false;
    if (got != 12 || h[0] != 0x1f || h[1] != 0x8b || h[2] != 8 || !(h[3] & 4)) {
      fail("igzip: encountered while decompressing file: " + path_ + " (not a BGZF member)");
      return false;
    }
    const unsigned xlen = h[10] | (h[11] << 8);
    unsigned char extra[65536];
    if (std::fread(extra, 1, xlen, fp_) != xlen) {
      fail("igzip: encountered while decompressing file: " + path_ + " (truncated)");
      return false;
    }
    long bsize = -1;
    for (unsigned i = 0; i + 4 <= xlen;) {
      const unsigned slen = extra[i + 2] | (extra[i + 3] << 8);
      if (extra[i] == 'B' && extra[i + 1] == 'C' && slen == 2 && i + 6 <= xlen) bsize = extra[i + 4] | (extra[i + 5] << 8);
      i += 4 + slen;
    }
    const long rest = bsize + 1 - 12 - (long)xlen;  // deflate data + CRC-32 + ISIZE
    if (bsize < 0 || rest < 8) {
      fail("igzip: encountered while decompressing file: " + path_ + " (not a BGZF member)");
      return false;
    }
    const size_t at = job.raw.size();
    job.raw.resize(at + (size_t)rest);
    if (std::fread(job.raw.data() + at, 1, (size_t)rest, fp_) != (size_t)rest) {
      fail("igzip: encountered while decompressing file: " + path_ + " (truncated)");
      return false;
    }
    const unsigned char *t = job.raw.data() + at + rest - 8;
    Member m;
    m.raw_at = at;
    m.raw_len = (size_t)rest - 8;
    m.crc = t[0] | (t[1] << 8) | (t[2] << 16) | ((uint32_t)t[3] << 24);
    m.isize = t[4] | (t[5] << 8) | (t[6] << 16) | ((uint32_t)t[7] << 24);
    if (m.isize > 65536u) {
      fail("igzip: encountered while decompressing file: " + path_ + " (BGZF member larger than 64 KiB)");
      return false;
    }
    m.out_at = out_bytes;
    out_bytes += m.isize;
    if (m.isize) job.members.push_back(m);  // (the end-of-file marker is an empty member)
    else job.raw.resize(at);
    return true;
  }

  void feed() {
    while (true) {
      auto job = std::make_shared<Job>();
      size_t out_bytes = 0;
      bool more = true;
      while (out_bytes < JOB_BYTES && (more = next_member(*job, out_bytes))) {
      }
      job->out.resize(out_bytes + Inflate::kOutSlack);  // (the decoder wants room to overshoot; handed out: out_bytes)
      job->out_bytes = out_bytes;
      job->raw.resize(job->raw.size() + Inflate::kInPad);  // readable bytes behind the last member's data
      std::unique_lock<std::mutex> lk(mu_);
      if (out_bytes) {
        cv_.wait(lk, [&] { return stop_ || jobs_.size() < max_jobs_; });
        if (stop_) return;
        jobs_.push_back(job);
      }
      if (!more || stop_) {
        fed_all_ = true;
        lk.unlock();
        cv_.notify_all();
        return;
      }
      lk.unlock();
      cv_.notify_all();
    }
  }

  void work() {
    Inflate inflate;  // (host/inflate_fast.h: about twice zlib's speed on FastQ text)
    while (true) {
      std::shared_ptr<Job> job;
      {
        std::unique_lock<std::mutex> lk(mu_);
        cv_.wait(lk, [&] { return stop_ || claimed_ < jobs_.size() || (fed_all_ && claimed_ >= jobs_.size()); });
        if (stop_ || claimed_ >= jobs_.size()) break;
        job = jobs_[claimed_++];
      }
      bool ok = true;
      for (const Member &m : job->members) {  // in file order: what a decode writes past its member's end is overwritten by the next
        inflate.reset();
        const uint8_t *in = job->raw.data() + m.raw_at;
        uint8_t *const first = reinterpret_cast<uint8_t *>(job->out.data() + m.out_at);
        uint8_t *out = first;
        const Inflate::Status st = inflate.run(in, job->raw.data() + m.raw_at + m.raw_len, true, first, out, first + m.isize + Inflate::kOutSlack);
        if (st != Inflate::kDone || (size_t)(out - first) != m.isize ||
            crc32(crc32(0L, Z_NULL, 0), reinterpret_cast<const Bytef *>(first), m.isize) != m.crc) {
          ok = false;
          break;
        }
      }
      if (!ok) fail("igzip: encountered while decompressing file: " + path_);
      std::vector<unsigned char>().swap(job->raw);
      {
        std::lock_guard<std::mutex> lk(mu_);
        job->done = true;
      }
      cv_.notify_all();
    }
  }

  std::string path_;
  FILE *fp_ = nullptr;
  std::mutex mu_;
  std::condition_variable cv_;
  std::deque<std::shared_ptr<Job>> jobs_;  // in file order; [0, claimed_) taken by the pool
  size_t claimed_ = 0, max_jobs_ = 4;
  bool fed_all_ = false, stop_ = false;
  std::string error_;
  std::thread feeder_;
  std::vector<std::thread> pool_;
};

}  // namespace humid_host
