// Host-side FastQ reading / writing for the `humid` executable (host/humid_main.cc).
//
// This is deliberately plain CPU code: the north star keeps FastQ parsing and gzip on the
// host.  Behaviour follows what the reference observes through fastp's FastqReader / Writer
// (lib/fastp/src/fastqreader.cpp:206-327, lib/fastp/src/writer.cpp:105-128):
//   * `.gz` is decided by the file name suffix; multi-member gzip streams are read through;
//   * lines end at '\n', '\r' or "\r\n"; lines before a record that do not start with '@'
//     are skipped; the name keeps its leading '@';
//   * a record whose sequence and quality lengths differ is reported on stderr and ends the
//     input (fastqreader.cpp:314-321 returns NULL, which src/fastq.cc:41-43 treats as EOF);
//   * a record is written back as name, sequence, strand, quality, each followed by '\n'
//     (Read::toString, lib/fastp/src/read.cpp:116-118).
// Compressed OUTPUT bytes depend on the deflate library (the reference uses libdeflate level 4,
// one gzip member per 4 MiB flush); parity is on the decompressed content.
#pragma once
#include <zlib.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <string>
#include <vector>

namespace humid_host {

inline bool ends_with(const std::string &s, const char *suffix) {
  size_t n = std::strlen(suffix);
  return s.size() >= n && s.compare(s.size() - n, n, suffix) == 0;
}

[[noreturn]] inline void error_exit(const std::string &msg) {  // lib/fastp/src/util.h:270-273
  std::cerr << "ERROR: " << msg << std::endl;
  std::exit(-1);
}

struct Record {
  std::string name, seq, strand, qual;
};

class FastqReader {
 public:
  explicit FastqReader(const std::string &path, size_t buffer_bytes = 1u << 23) : path_(path), buf_(buffer_bytes) {
    if (ends_with(path, ".gz")) {
      gz_ = gzopen(path.c_str(), "rb");
      if (!gz_) error_exit("Failed to open file: " + path);
      gzbuffer(gz_, 1u << 20);
    } else {
      fp_ = std::fopen(path.c_str(), "rb");
      if (!fp_) error_exit("Failed to open file: " + path);
    }
  }
  ~FastqReader() {
    if (gz_) gzclose(gz_);
    if (fp_) std::fclose(fp_);
  }
  FastqReader(const FastqReader &) = delete;
  FastqReader &operator=(const FastqReader &) = delete;

  // false at end of input (or at a malformed record, like the reference)
  bool read(Record &r) {
    if (at_end()) return false;
    get_line(r.name);
    while ((r.name.empty() && !at_end()) || (!r.name.empty() && r.name[0] != '@')) get_line(r.name);
    if (r.name.empty()) return false;
    get_line(r.seq);
    get_line(r.strand);
    get_line(r.qual);
    if (r.qual.size() != r.seq.size()) {
      std::cerr << "ERROR: sequence and quality have different length:\n"
                << r.name << "\n" << r.seq << "\n" << r.strand << "\n" << r.qual << std::endl;
      return false;
    }
    return true;
  }

 private:
  bool fill() {
    if (finished_) return false;
    int got;
    if (gz_) {
      got = gzread(gz_, buf_.data(), (unsigned)buf_.size());
      if (got < 0) error_exit("igzip: encountered while decompressing file: " + path_);
    } else {
      got = (int)std::fread(buf_.data(), 1, buf_.size(), fp_);
    }
    pos_ = 0;
    len_ = got > 0 ? (size_t)got : 0;
    if (len_ == 0) finished_ = true;
    return len_ > 0;
  }
  bool at_end() {
    if (pos_ < len_) return false;
    return !fill();
  }
  // one line without its terminator; "\r\n" counts as one terminator even across buffers
  void get_line(std::string &line) {
    line.clear();
    while (true) {
      if (pos_ >= len_ && !fill()) return;
      if (pending_lf_) {
        pending_lf_ = false;
        if (buf_[pos_] == '\n') { pos_++; continue; }
      }
      size_t e = pos_;
      while (e < len_ && buf_[e] != '\n' && buf_[e] != '\r') e++;
      line.append(buf_.data() + pos_, e - pos_);
      if (e < len_) {
        pending_lf_ = buf_[e] == '\r';
        pos_ = e + 1;
        if (pending_lf_ && pos_ < len_) {
          if (buf_[pos_] == '\n') pos_++;
          pending_lf_ = false;
        }
        return;
      }
      pos_ = len_;
    }
  }

  std::string path_;
  std::vector<char> buf_;
  size_t pos_ = 0, len_ = 0;
  bool finished_ = false, pending_lf_ = false;
  gzFile gz_ = nullptr;
  FILE *fp_ = nullptr;
};

class Writer {
 public:
  explicit Writer(const std::string &path) : path_(path) {
    if (ends_with(path, ".gz")) {
      gz_ = gzopen(path.c_str(), "wb4");  // compression level 4 (lib/fastp/src/options.cpp:15)
      if (!gz_) error_exit("Failed to write: " + path);
      gzbuffer(gz_, 1u << 20);
    } else {
      fp_ = std::fopen(path.c_str(), "wb");
      if (!fp_) error_exit("Failed to write: " + path);
      std::setvbuf(fp_, nullptr, _IOFBF, 1u << 22);
    }
  }
  ~Writer() {
    if (gz_) gzclose(gz_);
    if (fp_) std::fclose(fp_);
  }
  Writer(const Writer &) = delete;
  Writer &operator=(const Writer &) = delete;

  void write(const char *p, size_t n) {
    if (gz_) {
      if (gzwrite(gz_, p, (unsigned)n) != (int)n) error_exit("Failed to write: " + path_);
    } else if (std::fwrite(p, 1, n, fp_) != n) {
      error_exit("Failed to write: " + path_);
    }
  }
  // name + suffix, sequence, strand, quality, each followed by '\n' (Read::toString)
  template <class SpanT>
  void write_view(const SpanT &name, const SpanT &seq, const SpanT &strand, const SpanT &qual,
                  const std::string &name_suffix) {
    line_.clear();
    line_.append(name.p, name.n).append(name_suffix).push_back('\n');
    line_.append(seq.p, seq.n).push_back('\n');
    line_.append(strand.p, strand.n).push_back('\n');
    line_.append(qual.p, qual.n).push_back('\n');
    write(line_.data(), line_.size());
  }
  void write_record(const Record &r, const std::string &name_suffix = std::string()) {
    line_.clear();
    line_.append(r.name).append(name_suffix).push_back('\n');
    line_.append(r.seq).push_back('\n');
    line_.append(r.strand).push_back('\n');
    line_.append(r.qual).push_back('\n');
    write(line_.data(), line_.size());
  }

 private:
  std::string path_, line_;
  gzFile gz_ = nullptr;
  FILE *fp_ = nullptr;
};

// extractUMI_ (src/fastq.cc:72-93): text before the first space; the field after the last
// '_' if it is non-empty upper-case ACGT, else the field after the last ':' under the same
// test, else nothing.  Returns offset/length inside `header` (which includes the '@').
inline bool valid_umi(const char *s, size_t n) {
  if (n == 0) return false;
  for (size_t i = 0; i < n; i++)
    if (s[i] != 'A' && s[i] != 'C' && s[i] != 'G' && s[i] != 'T') return false;
  return true;
}
inline void extract_umi(const char *header, size_t size, size_t &off, size_t &len) {
  const char *sp = size ? (const char *)std::memchr(header, ' ', size) : nullptr;
  const size_t end = sp ? (size_t)(sp - header) : size;
  for (char sep : {'_', ':'}) {
    size_t last = end;  // position of the last `sep` before `end`, or `end` if there is none
    while (last > 0 && header[last - 1] != sep) last--;
    if (last > 0 && valid_umi(header + last, end - last)) {
      off = last;
      len = end - last;
      return;
    }
  }
  off = 0;
  len = 0;
}
inline void extract_umi(const std::string &header, size_t &off, size_t &len) {
  extract_umi(header.data(), header.size(), off, len);
}

}  // namespace humid_host
