// `humid` — the reference's command line (src/humid.cc:413-429) on top of libhumid_gpu.
//
//   humid [-n 24] [-m 1] [-l /dev/stderr] [-d .] [-s] [-q] [-a] [-e] [-x] files...
//
// Same options, defaults and toggling flags (-q is ON by default and switches the
// deduplicated output OFF, lib/commandIO/src/args.tcc:138-146), same output files
// (<name>_dedup.<ext>, <name>_annotated.<ext>, stats.dat, counts.dat, neigh.dat, clusters.dat)
// and the same phase lines in the log.  The host keeps FastQ parsing, gzip, header-UMI
// location and file naming; everything from word construction onwards is a call into the C
// ABI (include/humid_gpu.h) and runs on the GPU.  There is no CPU implementation of those
// steps here: if the library reports an error the program prints it and exits like the
// reference's error_exit does.
//
// Ingest: every input file is parsed in blocks (host/fastq_block.h) — a compressed one by
// one thread behind its inflate threads, a plain one by several threads side by side (host/ingest_parallel.h) — and
// its share of every word goes straight into a page-locked batch buffer; the main thread hands full batches to humid_gpu_add_batch, which uploads them while the
// next batch is being parsed and the previous one is counted on the device.  The output passes
// (host/fastq_rewrite.h) do not parse plain files again: ingest notes where every record
// lies, and several threads per file gather the kept ones by pread.  Compressed inputs are inflated a second
// time (host/inflate_fast.h, host/inflate_parallel.h, host/bgzf.h) and parsed by one thread per file.
#include <sys/stat.h>

#include <chrono>
#include <condition_variable>
#include <cstdlib>
#include <ctime>
#include <fstream>
#include <functional>
#include <memory>
#include <mutex>
#include <thread>

#include "../include/humid_gpu.h"
#include "../include/humid_gpu_shard.h"
#include "fastq_block.h"
#include "fastq_io.h"
#include "fastq_rewrite.h"
#include "ingest_parallel.h"

using namespace humid_host;
using std::string;
using std::vector;

namespace {

struct Options {
  size_t word_length = 24, distance = 1;
  string log_name = "/dev/stderr", dir_name = ".";
  bool stats = false, filter = true, annotate = false, edit = false, maximum = false;
  size_t gpus = 0;  // 0: $HUMID_GPUS, else 1
  vector<string> files;
};

void usage(const char *prog) {
  std::cout << "usage: " << prog << " [-n WORD_LENGTH] [-m MISMATCHES] [-l LOG] [-d DIR] [-s] [-q] [-a] [-e] [-x] files...\n\n"
            << "Deduplicate a dataset.\n\n"
            << "positional arguments:\n  files\t\tFastQ files\n\n"
            << "optional arguments:\n"
            << "  -n\t\tword length (default: 24)\n"
            << "  -m\t\tallowed mismatches (default: 1)\n"
            << "  -l\t\tlog file name (default: /dev/stderr)\n"
            << "  -d\t\toutput directory (default: .)\n"
            << "  -s\t\tcalculate statistics\n"
            << "  -q\t\twrite deduplicated FastQ files (default: on; the flag switches it off)\n"
            << "  -a\t\twrite annotated FastQ files\n"
            << "  -e\t\tuse edit distance\n"
            << "  -x\t\tuse maximum clustering method\n"
            << "\nlimits of this implementation (the reference has none): word length 1.." << HUMID_MAX_WORD_LENGTH
            << "; Hamming distance <= 8;\n  edit distance <= 3, and <= 1 for words of more than " << HUMID_MAX_NARROW_WORD_LENGTH
            << " letters; fewer than 2^32 - 1 read tuples; at most " << HUMID_MAX_SEGMENTS - 1 << " files\n"
            << "\nnot in the reference:\n"
            << "  -g\t\tnumber of GPUs to share the word table over (default: $HUMID_GPUS, else 1);\n"
            << "    \t\twords of more than 31 letters and edit distances above 1 run on one GPU\n";
}

bool parse(int argc, char **argv, Options &o) {
  for (int i = 1; i < argc; i++) {
    string a = argv[i];
    auto value = [&](string &dst) { if (i + 1 < argc) dst = argv[++i]; };
    string v;
    if (a == "-h" || a == "--help") return false;
    else if (a == "-n") { value(v); o.word_length = std::strtoull(v.c_str(), nullptr, 10); }
    else if (a == "-m") { value(v); o.distance = std::strtoull(v.c_str(), nullptr, 10); }
    else if (a == "-g") { value(v); o.gpus = std::strtoull(v.c_str(), nullptr, 10); }
    else if (a == "-l") value(o.log_name);
    else if (a == "-d") value(o.dir_name);
    else if (a == "-s") o.stats = !o.stats;
    else if (a == "-q") o.filter = !o.filter;
    else if (a == "-a") o.annotate = !o.annotate;
    else if (a == "-e") o.edit = !o.edit;
    else if (a == "-x") o.maximum = !o.maximum;
    else o.files.push_back(a);  // unknown tokens fall through to the positional list (eval.tcc:123-131)
  }
  return !o.files.empty();
}

// HUMID_TIMING=1: phase times in milliseconds on stderr (the log keeps the reference's whole seconds)
std::chrono::steady_clock::time_point phase_start;
const char *phase_name = "";
time_t start_message(std::ofstream &log, const char *msg) {  // src/log.cc:4-9
  log << msg << "... ";
  log.flush();
  phase_name = msg;
  phase_start = std::chrono::steady_clock::now();
  return time(nullptr);
}
void end_message(std::ofstream &log, time_t start) {  // src/log.cc:11-15
  if (std::getenv("HUMID_TIMING"))
    std::cerr << phase_name << ": "
              << std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - phase_start).count() << " ms\n";
  time_t s = static_cast<time_t>(difftime(time(nullptr), start));
  log << "done. (" << s / 60 << 'm' << s % 60 << "s)\n";
  log.flush();
}

// src/fastq.cc:170-181 with fastp's basename (lib/fastp/src/util.h:126-134)
string make_file_name(const string &filename, const string &dir, const string &suffix) {
  string name = filename;
  size_t slash = filename.find_last_of('/');
  if (slash != string::npos) name = slash == filename.size() - 1 ? "" : filename.substr(slash + 1);
  size_t pos = name.find('.');
  string out = name.substr(0, pos) + '_' + suffix + (pos == string::npos ? "" : name.substr(pos));
  return dir + '/' + out;
}

void make_dirs(const string &dir) {  // std::filesystem::create_directories
  string cur;
  for (size_t i = 0; i <= dir.size(); i++) {
    if (i == dir.size() || dir[i] == '/') {
      if (!cur.empty()) mkdir(cur.c_str(), 0777);
    }
    if (i < dir.size()) cur.push_back(dir[i]);
  }
}

struct Layout {  // preCompute, src/humid.cc:38-59 (+ ntFromFile, src/fastq.cc:220-230)
  size_t header_umi = 0;
  vector<size_t> nt_to_take;
};

Layout pre_compute(const vector<string> &files, size_t word_length) {
  Layout L;
  size_t umi = 0;
  {
    FastqReader reader(files.front(), 1u << 16);  // one record is all that is read
    Record r;
    if (reader.read(r)) {
      size_t off, len;
      extract_umi(r.name, off, len);
      umi = len;
    }
  }
  size_t from_file = word_length > umi ? word_length - umi : 0;
  size_t div = from_file / files.size();
  L.nt_to_take.assign(files.size(), div);
  L.nt_to_take.back() += from_file % files.size();
  L.header_umi = word_length < umi ? word_length : umi;
  return L;
}

// ---- pass 1: tuples of all files in lock step (readFiles / readFastq_, src/fastq.cc:37-47,
// :96-114): the tuple stream ends as soon as any file runs out --------------------------------
#ifndef HUMID_BATCH
#define HUMID_BATCH (1u << 20)
#endif
constexpr size_t BATCH = HUMID_BATCH;  // tuples per upload (a multiple of REWRITE_CHUNK; the tests build with a small one)
static_assert(BATCH % REWRITE_CHUNK == 0, "a batch must hold whole output chunks");
constexpr int N_SLOTS = 8;  // batches the readers may run ahead: they start while the device is still being brought up

struct Slot {
  uint8_t *chars = nullptr;     // BATCH x word_length, row-major; file i fills its own columns
  uint16_t *seg_len = nullptr;  // BATCH x n_segments
  uint64_t batch = 0;           // which batch this slot is collecting
  size_t done = 0;              // files that have finished it
  vector<size_t> rows;          // records each file contributed (< BATCH only at its end)
  vector<size_t> filled;        // several parser threads per file: rows of the batch written so far
  bool locked = false;          // page-locked for the uploads (main thread, at first use)
};

struct Ingest {
  std::mutex mu;
  std::condition_variable cv;
  Slot slots[N_SLOTS];
  bool stop = false;
  size_t n = 0, S = 0;          // word length, segments
  vector<size_t> seg_off;       // first column of a segment
  Layout L;
  vector<RecordMap> maps;       // per file: what pass 2 needs to know about its records (host/fastq_rewrite.h)
};

// file i: segment (header UMI ? i + 1 : i); file 0 also fills segment 0 from its headers
void ingest_file(Ingest &g, const string &path, size_t file) {
  BlockReader reader(path);
  RecordView rec;
  const size_t seg = g.L.header_umi ? file + 1 : file;
  const size_t take = g.L.nt_to_take[file], off = g.seg_off[seg];
  bool ended = false;
  for (uint64_t k = 0; !ended; k++) {
    Slot &s = g.slots[k % N_SLOTS];
    {
      std::unique_lock<std::mutex> lk(g.mu);
      g.cv.wait(lk, [&] { return g.stop || s.batch == k; });
      if (g.stop) return;
    }
    size_t m = 0;
    for (; m < BATCH; m++) {
      if (!reader.read(rec)) { ended = true; break; }
      g.maps[file].note(k * BATCH + m, reader, rec);  // (BATCH is a multiple of REWRITE_CHUNK)
      // getNucleotides (src/fastq.cc:116-144) WITHOUT the padding / validation: the device applies
      // makeStringSize_ and the ACGT filter from (chars, seg_len)
      uint8_t *row = s.chars + m * g.n;
      uint16_t *sl = s.seg_len + m * g.S;
      auto put = [&](const char *p, size_t have, size_t want, size_t at, size_t which) {
        size_t c = have < want ? have : want;
        std::memcpy(row + at, p, c);
        std::memset(row + at + c, 'N', want - c);
        sl[which] = (uint16_t)c;
      };
      if (file == 0 && g.L.header_umi) {
        size_t uo, ul;
        extract_umi(rec.name.p, rec.name.n, uo, ul);
        put(rec.name.p + uo, ul, g.L.header_umi, 0, 0);
      }
      put(rec.seq.p, rec.seq.n, take, off, seg);
    }
    std::lock_guard<std::mutex> lk(g.mu);
    s.rows[file] = m;
    s.done++;
    g.cv.notify_all();
  }
}

// The same for a plain file with several parser threads (host/ingest_parallel.h): the threads keep this file's letters
// of every record they parse; once a range's place in the file is known (commit, in file order: only the record map
// is written there) its thread copies them into the batch buffers, side by side with the other ranges.
// row: [take letters, 'N'-padded][u16 letters present] (+ [header UMI letters][u16 present] for file 0)
void ingest_file_parallel(Ingest &g, const string &path, size_t file, unsigned n_threads) {
  const size_t seg = g.L.header_umi ? file + 1 : file;
  const size_t take = g.L.nt_to_take[file], off = g.seg_off[seg];
  const size_t umi = file == 0 ? g.L.header_umi : 0;
  const size_t row_bytes = take + 2 + (umi ? umi + 2 : 0);
  auto put = [](uint8_t *dst, const char *p, size_t have, size_t want) {
    const size_t c = have < want ? have : want;
    std::memcpy(dst, p, c);
    std::memset(dst + c, 'N', want - c);
    const uint16_t n = (uint16_t)c;
    std::memcpy(dst + want, &n, 2);
  };
  ParallelPlainParser parser(path, row_bytes, [&](const RecordView &rec, uint8_t *row) {
    put(row, rec.seq.p, rec.seq.n, take);
    if (umi) {
      size_t uo, ul;
      extract_umi(rec.name.p, rec.name.n, uo, ul);
      put(row + take + 2, rec.name.p + uo, ul, umi);
    }
  });
  // the buffer of batch k once the main thread has handed it out again; nullptr: the tuple stream has ended
  auto slot_of = [&](uint64_t k) -> Slot * {
    Slot &s = g.slots[k % N_SLOTS];
    std::unique_lock<std::mutex> lk(g.mu);
    g.cv.wait(lk, [&] { return g.stop || s.batch == k; });
    return g.stop ? nullptr : &s;
  };
  auto publish = [&](Slot &s, size_t rows) {  // (g.mu held)
    s.rows[file] = rows;
    s.filled[file] = 0;
    s.done++;
    g.cv.notify_all();
  };
  const uint64_t n = parser.run(
      n_threads,
      [&](uint64_t first, const ParallelPlainParser::Range &r, bool) {
        for (size_t i = 0; i < r.at.size(); i++) g.maps[file].note_values(first + i, r.at[i], r.stored[i]);
        std::lock_guard<std::mutex> lk(g.mu);
        return !g.stop;
      },
      [&](uint64_t first, const ParallelPlainParser::Range &r) {
        const size_t total = r.at.size();
        for (size_t i = 0; i < total;) {
          const uint64_t t = first + i, k = t / BATCH;
          const size_t m0 = (size_t)(t % BATCH), cnt = std::min(total - i, BATCH - m0);
          Slot *s = slot_of(k);
          if (!s) return;
          for (size_t j = 0; j < cnt; j++) {
            const uint8_t *row = r.rows.data() + (i + j) * row_bytes;
            uint8_t *chars = s->chars + (m0 + j) * g.n;
            uint16_t *sl = s->seg_len + (m0 + j) * g.S;
            std::memcpy(chars + off, row, take);
            std::memcpy(&sl[seg], row + take, 2);
            if (umi) {
              std::memcpy(chars, row + take + 2, umi);
              std::memcpy(&sl[0], row + take + 2 + umi, 2);
            }
          }
          {
            std::lock_guard<std::mutex> lk(g.mu);
            s->filled[file] += cnt;
            if (s->filled[file] == BATCH) publish(*s, BATCH);  // (a short last batch is published below)
          }
          i += cnt;
        }
      });
  // the file's last batch: short, or empty if the file ended on a batch boundary (what tells the main thread)
  if (Slot *s = slot_of(n / BATCH)) {
    std::lock_guard<std::mutex> lk(g.mu);
    publish(*s, (size_t)(n % BATCH));
  }
  if (std::getenv("HUMID_TIMING"))
    std::cerr << "ingest " << path << ": " << parser.ranges() << " ranges, " << parser.reparsed_ranges() << " parsed again\n";
}

// Parser threads per plain input file (HUMID_INGEST_THREADS overrides; 1 = one block reader per file).
unsigned ingest_threads(size_t n_files) {
  if (const char *e = std::getenv("HUMID_INGEST_THREADS")) return (unsigned)std::max(1, std::atoi(e));
  const unsigned hw = std::max(1u, std::thread::hardware_concurrency());
  const unsigned per_file = hw / (unsigned)n_files;
  return per_file > 2 ? std::min(6u, per_file - 1) : 1u;
}

// Capacity hint for the word table (lets the device count batches while the next one uploads): size
// of the first file / size of its first record, for plain files.  A wrong guess only costs a recount
// (too small) or memory (too large), never the result.
uint64_t estimate_tuples(const string &path) {
  if (ends_with(path, ".gz")) return 0;
  struct stat st;
  if (stat(path.c_str(), &st) != 0 || st.st_size <= 0) return 0;
  BlockReader reader(path, 1u << 16);
  RecordView r;
  if (!reader.read(r)) return 0;
  size_t rec = r.name.n + r.seq.n + r.strand.n + r.qual.n + 4;
  uint64_t est = (uint64_t)st.st_size / rec;
  est += est / 16 + 1024;
  return est < 0xF0000000ull ? est : 0;
}

void check(humid_gpu *h, int rc, const char *what) {
  if (rc != HUMID_GPU_OK) error_exit(string(what) + ": " + humid_gpu_last_error(h));
}

}  // namespace

int main(int argc, char **argv) {
  Options opt;
  if (!parse(argc, argv, opt)) {
    usage(argv[0]);
    return argc > 1 && (string(argv[1]) == "-h" || string(argv[1]) == "--help") ? 0 : 1;
  }
  // what the device path does not cover is refused here, before any file is read (the reference takes any -n / -m)
  if (opt.word_length == 0 || opt.word_length > HUMID_MAX_WORD_LENGTH)
    error_exit("word length (-n) must be 1.." + std::to_string(HUMID_MAX_WORD_LENGTH) + " in this implementation");
  if (!opt.edit && opt.distance > 8) error_exit("Hamming distance (-m) above 8 is not supported by this implementation");
  if (opt.edit && opt.distance > 3) error_exit("edit distance (-e -m) above 3 is not supported by this implementation");
  if (opt.edit && opt.distance >= 2 && opt.word_length > HUMID_MAX_NARROW_WORD_LENGTH)
    error_exit("edit distance above 1 needs a word length (-n) of at most " + std::to_string(HUMID_MAX_NARROW_WORD_LENGTH) +
               " in this implementation");
  if (opt.files.size() > HUMID_MAX_SEGMENTS - 1) error_exit("too many input files");
  std::ofstream log(opt.log_name.c_str(), std::ios::out | std::ios::binary);
  const size_t F = opt.files.size();

  // ---- readData (src/humid.cc:70-103) ------------------------------------------------------
  time_t t0 = start_message(log, "Determing nucleotides to take");
  Layout L = pre_compute(opt.files, opt.word_length);
  end_message(log, t0);
  log << "  header: " << L.header_umi;
  for (size_t i = 0; i < F; i++) log << "\n  " << opt.files[i] << ": " << L.nt_to_take[i];
  log << "\n";

  humid_gpu_config cfg{};
  cfg.word_length = (uint32_t)opt.word_length;
  cfg.distance = (uint32_t)opt.distance;
  cfg.edit = opt.edit;
  cfg.maximum = opt.maximum;
  cfg.device = -1;
  vector<size_t> seg_take;
  if (L.header_umi) seg_take.push_back(L.header_umi);
  for (size_t t : L.nt_to_take) seg_take.push_back(t);
  if (seg_take.size() > HUMID_MAX_SEGMENTS) error_exit("too many input files");
  cfg.n_segments = (uint32_t)seg_take.size();
  for (size_t i = 0; i < seg_take.size(); i++) cfg.seg_take[i] = (uint32_t)seg_take[i];
  cfg.expected_tuples = estimate_tuples(opt.files.front());
  // Bringing the device up (CUDA context, first allocations) takes about a second: it runs beside the readers,
  // which fill ordinary page-aligned buffers meanwhile; the buffers are page-locked once the handle exists.
  // Several GPUs (not in the reference: `-g N` / $HUMID_GPUS; $HUMID_GPU_DEVICES = "0,1,..." picks the ordinals, the same
  // one may be named twice): one handle per GPU, joined into a group of the library (include/humid_gpu_shard.h).  Batches
  // go to the handles in turn under their position in the job; the word table is range-partitioned over the GPUs by
  // humid_gpu_finalize, and the per-read results of the whole job come back through the first handle.
  size_t n_gpus = opt.gpus ? opt.gpus : (std::getenv("HUMID_GPUS") ? std::strtoull(std::getenv("HUMID_GPUS"), nullptr, 10) : 1);
  if (n_gpus < 1) n_gpus = 1;
  if (n_gpus > HUMID_MAX_PARTS) error_exit("at most " + std::to_string(HUMID_MAX_PARTS) + " GPUs");
  if (n_gpus > 1 && (opt.word_length > HUMID_MAX_NARROW_WORD_LENGTH || (opt.edit && opt.distance >= 2))) n_gpus = 1;
  vector<int> devices;
  if (const char *list = std::getenv("HUMID_GPU_DEVICES")) {
    for (const char *p = list; *p;) {
      devices.push_back((int)std::strtol(p, const_cast<char **>(&p), 10));
      while (*p == ',' || *p == ' ') p++;
    }
  }
  vector<humid_gpu *> gpus(n_gpus, nullptr);
  humid_gpu *&gpu = gpus[0];  // counts, histograms and (in a group) every per-read result
  humid_gpu_group *group = n_gpus > 1 ? humid_gpu_group_create((uint32_t)n_gpus) : nullptr;
  if (n_gpus > 1 && !group) error_exit("humid_gpu_group_create failed");
  int create_rc = HUMID_GPU_OK;
  string create_msg;
  std::mutex create_mu;
  // run `body(rank)` on one thread per GPU (the group's calls are collective) and wait for all of them
  auto on_every_gpu = [&](const std::function<void(size_t)> &body) {
    if (n_gpus == 1) { body(0); return; }
    vector<std::thread> th;
    for (size_t r = 0; r < n_gpus; r++) th.emplace_back(body, r);
    for (std::thread &t : th) t.join();
  };
  std::thread creator([&] {
    const auto c0 = std::chrono::steady_clock::now();
    on_every_gpu([&](size_t r) {
      humid_gpu_config c = cfg;
      if (n_gpus > 1) {
        c.device = !devices.empty() ? devices[r % devices.size()] : (int)r;
        c.expected_tuples = 0;  // words are counted by the owners of their key range
      }
      int rc = humid_gpu_create(&c, &gpus[r]);
      if (rc == HUMID_GPU_OK && group)
        rc = humid_gpu_group_join(gpus[r], group, nullptr, (uint32_t)r, (uint32_t)n_gpus, HUMID_GROUP_BATCHES);
      if (rc != HUMID_GPU_OK) {
        std::lock_guard<std::mutex> lk(create_mu);
        create_rc = rc;
        create_msg = humid_gpu_last_error(gpus[r]);
      }
    });
    if (std::getenv("HUMID_TIMING"))
      std::cerr << "humid_gpu_create (CUDA context, beside the readers): "
                << std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - c0).count() << " ms\n";
  });

  t0 = start_message(log, "Reading data");
  uint64_t total = 0;
  vector<RecordMap> maps;
  {
    Ingest g;
    g.maps.assign(F, RecordMap());
    // plain files: the length of every record, so that the output passes gather instead of parsing again
    for (size_t i = 0; i < F; i++) g.maps[i].lengths = (opt.filter || opt.annotate) && !ends_with(opt.files[i], ".gz");
    g.n = opt.word_length;
    g.S = seg_take.size();
    g.L = L;
    size_t at = 0;
    for (size_t t : seg_take) { g.seg_off.push_back(at); at += t; }
    auto page_aligned = [](size_t bytes) {
      void *p = nullptr;
      if (posix_memalign(&p, 4096, (bytes + 4095) / 4096 * 4096) != 0) error_exit("out of memory");
      return p;
    };
    for (int j = 0; j < N_SLOTS; j++) {
      g.slots[j].chars = static_cast<uint8_t *>(page_aligned(BATCH * g.n));
      g.slots[j].seg_len = static_cast<uint16_t *>(page_aligned(BATCH * g.S * sizeof(uint16_t)));
      g.slots[j].batch = (uint64_t)j;
      g.slots[j].rows.assign(F, 0);
      g.slots[j].filled.assign(F, 0);
    }
    vector<std::thread> readers;
    const unsigned parsers = ingest_threads(F);
    share_inflate_threads(F);
    for (size_t i = 0; i < F; i++) {
      struct stat st;
      const bool big_plain = !ends_with(opt.files[i], ".gz") && stat(opt.files[i].c_str(), &st) == 0 && S_ISREG(st.st_mode) &&
                             (uint64_t)st.st_size >= 2 * ParallelPlainParser::range_size();
      if (parsers > 1 && big_plain) readers.emplace_back(ingest_file_parallel, std::ref(g), opt.files[i], i, parsers);
      else readers.emplace_back(ingest_file, std::ref(g), opt.files[i], i);
    }
    for (uint64_t k = 0;; k++) {
      Slot &s = g.slots[k % N_SLOTS];
      {
        std::unique_lock<std::mutex> lk(g.mu);
        g.cv.wait(lk, [&] { return s.done == F; });
      }
      if (creator.joinable()) {  // first batch: the device must be there now
        creator.join();
        if (create_rc != HUMID_GPU_OK) error_exit("humid_gpu_create: " + create_msg);
      }
      if (!s.locked) {
        // page-locked when first used (a small job never touches most of the buffers: 30 MB each, ~0.2 ms per MB to pin);
        // page-locking is an optimisation: without it the uploads are staged by the driver
        s.locked = true;
        const bool ok = humid_gpu_host_register(gpu, s.chars, BATCH * g.n) == HUMID_GPU_OK &&
                        humid_gpu_host_register(gpu, s.seg_len, BATCH * g.S * sizeof(uint16_t)) == HUMID_GPU_OK;
        if (!ok && std::getenv("HUMID_TIMING")) std::cerr << "batch buffers could not be page-locked: " << humid_gpu_last_error(gpu) << "\n";
      }
      size_t m = BATCH;
      for (size_t r : s.rows) m = r < m ? r : m;
      humid_gpu *dst = gpus[k % n_gpus];  // in a group `total` is the batch's position in the job
      if (m) check(dst, humid_gpu_add_batch(dst, s.chars, s.seg_len, total, m), "humid_gpu_add_batch");
      total += m;
      std::lock_guard<std::mutex> lk(g.mu);
      if (m < BATCH) g.stop = true;  // some file has ended: so has the tuple stream
      s.done = 0;
      s.batch = k + N_SLOTS;
      g.cv.notify_all();
      if (g.stop) break;
    }
    for (std::thread &t : readers) t.join();
    maps.swap(g.maps);
    for (int j = 0; j < N_SLOTS; j++) {
      if (g.slots[j].locked) {
        humid_gpu_host_unregister(gpu, g.slots[j].chars);
        humid_gpu_host_unregister(gpu, g.slots[j].seg_len);
      }
      std::free(g.slots[j].chars);
      std::free(g.slots[j].seg_len);
    }
  }
  end_message(log, t0);

  // ---- neighbours + clusters (src/humid.cc:381-388) ---------------------------------------------
  // The device computes both in one call; the two phase lines are kept for log compatibility.
  t0 = start_message(log, opt.edit ? "Calculating neighbours using Levenshtein distance"
                                   : "Calculating neighbours using Hamming distance");
  on_every_gpu([&](size_t r) { check(gpus[r], humid_gpu_finalize(gpus[r]), "humid_gpu_finalize"); });
  end_message(log, t0);
  t0 = start_message(log, opt.maximum ? "Calculating maximum clusters" : "Calculating directional clusters");
  end_message(log, t0);

  make_dirs(opt.dir_name);
  // pass 2 (host/fastq_rewrite.h): one writer per file, each with its share of the host's threads
  const unsigned hw = std::max(1u, std::thread::hardware_concurrency());
  const unsigned threads_per_file = std::max(1u, std::min(8u, hw / (unsigned)F));

  // ---- writeFiltered (src/humid.cc:203-241): the keep bit of a tuple decides for all its files --
  if (opt.filter) {
    t0 = start_message(log, "Writing filtered results");
    vector<uint64_t> keep((total + 63) / 64 + 1);
    check(gpu, humid_gpu_keep_mask(gpu, 0, total, keep.data()), "humid_gpu_keep_mask");
    vector<std::thread> writers;
    for (size_t i = 0; i < F; i++)
      writers.emplace_back(rewrite_file_parallel, opt.files[i], make_file_name(opt.files[i], opt.dir_name, "dedup"), total,
                           keep.data(), nullptr, std::cref(maps[i]), threads_per_file);
    for (std::thread &t : writers) t.join();
    end_message(log, t0);
  }

  // ---- writeAnnotated (:251-292) ------------------------------------------------------------
  if (opt.annotate) {
    t0 = start_message(log, "Writing annotated results");
    vector<uint32_t> ids(total + 1);
    if (group)  // the owners of the words send every read's id to the first handle
      on_every_gpu([&](size_t r) { check(gpus[r], humid_gpu_group_collect_ids(gpus[r]), "humid_gpu_group_collect_ids"); });
    for (uint64_t lo = 0; lo < total; lo += 16 * BATCH) {
      uint64_t m = std::min<uint64_t>(16 * BATCH, total - lo);
      check(gpu, humid_gpu_cluster_ids(gpu, lo, m, ids.data() + lo), "humid_gpu_cluster_ids");
    }
    vector<std::thread> writers;
    for (size_t i = 0; i < F; i++)
      writers.emplace_back(rewrite_file_parallel, opt.files[i], make_file_name(opt.files[i], opt.dir_name, "annotated"),
                           total, nullptr, ids.data(), std::cref(maps[i]), threads_per_file);
    for (std::thread &t : writers) t.join();
    end_message(log, t0);
  }

  // ---- runStatistics / clusterStats / writeStatistics (:301-357) ------------------------------
  if (opt.stats) {
    t0 = start_message(log, "Calculating count and neighbour stats");
    uint64_t tot, usable, unique, clusters;
    check(gpu, humid_gpu_counts(gpu, &tot, &usable, &unique, &clusters), "humid_gpu_counts");
    const char *names[3] = {"counts.dat", "neigh.dat", "clusters.dat"};
    vector<uint64_t> k[3], v[3];
    for (int w = 0; w < 3; w++) {
      uint64_t m = 0;
      check(gpu, humid_gpu_histogram(gpu, w, nullptr, nullptr, &m), "humid_gpu_histogram");
      k[w].resize(m + 1);
      v[w].resize(m + 1);
      check(gpu, humid_gpu_histogram(gpu, w, k[w].data(), v[w].data(), &m), "humid_gpu_histogram");
      k[w].resize(m);
      v[w].resize(m);
    }
    end_message(log, t0);
    for (int w = 0; w < 3; w++) {
      std::ofstream o(opt.dir_name + '/' + names[w], std::ios::out | std::ios::binary);
      for (size_t i = 0; i < k[w].size(); i++) o << k[w][i] << ' ' << v[w][i] << '\n';
    }
    std::ofstream o(opt.dir_name + "/stats.dat", std::ios::out | std::ios::binary);
    o << "total: " << tot << '\n' << "usable: " << usable << '\n' << "unique: " << unique << '\n'
      << "clusters: " << clusters << '\n';
  }

  log.close();
  phase_start = std::chrono::steady_clock::now();
  for (humid_gpu *g : gpus) humid_gpu_destroy(g);
  if (group) humid_gpu_group_destroy(group);
  if (std::getenv("HUMID_TIMING"))
    std::cerr << "humid_gpu_destroy: "
              << std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - phase_start).count() << " ms\n";
  // Every output file and the log are closed, the handles are destroyed.  What is left is the CUDA runtime's own
  // teardown at exit (its context, the page-locked batch buffers): ~0.1 s that a finished program does not need —
  // the driver releases everything with the process.  HUMID_FULL_EXIT=1 returns normally (sanitizer / leak checks).
  std::cout.flush();
  std::cerr.flush();
  if (!std::getenv("HUMID_FULL_EXIT")) std::_Exit(0);
  return 0;
}
