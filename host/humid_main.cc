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
#include <sys/stat.h>

#include <ctime>
#include <fstream>
#include <memory>

#include "../include/humid_gpu.h"
#include "fastq_io.h"

using namespace humid_host;
using std::string;
using std::vector;

namespace {

struct Options {
  size_t word_length = 24, distance = 1;
  string log_name = "/dev/stderr", dir_name = ".";
  bool stats = false, filter = true, annotate = false, edit = false, maximum = false;
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
            << "  -x\t\tuse maximum clustering method\n";
}

bool parse(int argc, char **argv, Options &o) {
  for (int i = 1; i < argc; i++) {
    string a = argv[i];
    auto value = [&](string &dst) { if (i + 1 < argc) dst = argv[++i]; };
    string v;
    if (a == "-h" || a == "--help") return false;
    else if (a == "-n") { value(v); o.word_length = std::strtoull(v.c_str(), nullptr, 10); }
    else if (a == "-m") { value(v); o.distance = std::strtoull(v.c_str(), nullptr, 10); }
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

time_t start_message(std::ofstream &log, const char *msg) {  // src/log.cc:4-9
  log << msg << "... ";
  log.flush();
  return time(nullptr);
}
void end_message(std::ofstream &log, time_t start) {  // src/log.cc:11-15
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
    FastqReader reader(files.front());
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

// lock-step reader over all input files (readFiles / readFastq_, src/fastq.cc:37-47, :96-114):
// the tuple stream ends as soon as any file runs out.
class TupleReader {
 public:
  explicit TupleReader(const vector<string> &files) {
    for (const string &f : files) readers_.emplace_back(new FastqReader(f));
    recs.resize(files.size());
  }
  bool next() {
    bool ok = true;
    for (size_t i = 0; i < readers_.size(); i++)
      if (!readers_[i]->read(recs[i])) ok = false;
    return ok;
  }
  vector<Record> recs;

 private:
  vector<std::unique_ptr<FastqReader>> readers_;
};

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
  humid_gpu *gpu = nullptr;
  check(nullptr, humid_gpu_create(&cfg, &gpu), "humid_gpu_create");

  t0 = start_message(log, "Reading data");
  const size_t n = opt.word_length, S = seg_take.size(), BATCH = 1u << 20;
  vector<uint8_t> chars(BATCH * n);
  vector<uint16_t> seg_len(BATCH * S);
  uint64_t total = 0, pending = 0, first = 0;
  auto flush = [&] {
    if (!pending) return;
    check(gpu, humid_gpu_add_batch(gpu, chars.data(), seg_len.data(), first, pending), "humid_gpu_add_batch");
    first += pending;
    pending = 0;
  };
  {
    TupleReader in(opt.files);
    while (in.next()) {
      // getNucleotides (src/fastq.cc:116-144) WITHOUT the padding / validation: the device
      // applies makeStringSize_ and the ACGT filter from (chars, seg_len)
      uint8_t *dst = chars.data() + pending * n;
      uint16_t *sl = seg_len.data() + pending * S;
      size_t s = 0;
      auto put = [&](const char *p, size_t have, size_t take) {
        size_t m = have < take ? have : take;
        std::memcpy(dst, p, m);
        std::memset(dst + m, 'N', take - m);
        dst += take;
        sl[s++] = (uint16_t)m;
      };
      if (L.header_umi) {
        size_t off, len;
        extract_umi(in.recs[0].name, off, len);
        put(in.recs[0].name.data() + off, len, L.header_umi);
      }
      for (size_t i = 0; i < F; i++) put(in.recs[i].seq.data(), in.recs[i].seq.size(), L.nt_to_take[i]);
      total++;
      if (++pending == BATCH) flush();
    }
    flush();
  }
  end_message(log, t0);

  // ---- neighbours + clusters (src/humid.cc:381-388) ---------------------------------------------
  // The device computes both in one call; the two phase lines are kept for log compatibility.
  t0 = start_message(log, opt.edit ? "Calculating neighbours using Levenshtein distance"
                                   : "Calculating neighbours using Hamming distance");
  check(gpu, humid_gpu_finalize(gpu), "humid_gpu_finalize");
  end_message(log, t0);
  t0 = start_message(log, opt.maximum ? "Calculating maximum clusters" : "Calculating directional clusters");
  end_message(log, t0);

  make_dirs(opt.dir_name);

  // ---- writeFiltered (src/humid.cc:203-241) ---------------------------------------------------
  if (opt.filter) {
    t0 = start_message(log, "Writing filtered results");
    vector<uint64_t> keep((total + 63) / 64 + 1);
    check(gpu, humid_gpu_keep_mask(gpu, 0, total, keep.data()), "humid_gpu_keep_mask");
    vector<std::unique_ptr<Writer>> out;
    for (const string &f : opt.files) out.emplace_back(new Writer(make_file_name(f, opt.dir_name, "dedup")));
    TupleReader in(opt.files);
    for (uint64_t t = 0; t < total && in.next(); t++)
      if (keep[t >> 6] >> (t & 63) & 1)
        for (size_t i = 0; i < F; i++) out[i]->write_record(in.recs[i]);
    end_message(log, t0);
  }

  // ---- writeAnnotated (:251-292) ------------------------------------------------------------
  if (opt.annotate) {
    t0 = start_message(log, "Writing annotated results");
    vector<std::unique_ptr<Writer>> out;
    for (const string &f : opt.files) out.emplace_back(new Writer(make_file_name(f, opt.dir_name, "annotated")));
    TupleReader in(opt.files);
    vector<uint32_t> ids(BATCH);
    for (uint64_t lo = 0; lo < total; lo += BATCH) {
      uint64_t m = std::min<uint64_t>(BATCH, total - lo);
      check(gpu, humid_gpu_cluster_ids(gpu, lo, m, ids.data()), "humid_gpu_cluster_ids");
      for (uint64_t k = 0; k < m && in.next(); k++) {
        string suffix = ":" + std::to_string(ids[k]);
        for (size_t i = 0; i < F; i++) out[i]->write_record(in.recs[i], suffix);
      }
    }
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
  humid_gpu_destroy(gpu);
  return 0;
}
