// TEST / BASELINE INFRASTRUCTURE ONLY (oracle/_ref build) — never linked into the product.
//
// The reference's hot path timed from HOST MEMORY: this file #includes the UNMODIFIED src/humid.cc from
// /root/reference where it lies (its `main` renamed away) and calls the reference's own functions —
// makeWord (src/fastq.cc:146-161), Trie::add, findHammingNeighbours / findEditNeighbours (src/humid.cc:113-158),
// findClusters (:167-193), Trie::find + the visited / maxLeaf test of writeFiltered (:223-231, the only lines
// repeated here, because the reference interleaves them with file output), runStatistics (:301-315) and
// clusterStats (src/cluster.cc:89-95).  What it leaves out is exactly what the C ABI of this repository leaves to
// the host: FastQ parsing, gzip and writing files.  So its rate is the like-for-like CPU figure for bench.py's
// `e2e` (host buffers in -> keep mask + histograms out).
//
// usage: hotpath_probe <chars.bin> <n_tuples> <word_length> <distance> <edit 0|1> <maximum 0|1> <header_umi>
//                      <out_prefix> <nt_to_take>...
//   chars.bin   n_tuples x word_length ASCII bytes, the layout humid_gpu_add_batch takes (header UMI first, then the
//               share of every file)
//   out_prefix  writes <out_prefix>.keep (bit t = tuple t is written to the *_dedup files, LSB first) and
//               <out_prefix>.{counts,neigh,clusters}.dat in writeStatistics' format
// stdout: one JSON line with the counts and the seconds per phase.
#include <chrono>
#include <cstdio>

#define main humid_reference_cli_main
#include "src/humid.cc"
#undef main

namespace {
double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

struct Tuples {
  const char* chars;
  size_t n, wordLength, headerUMI;
  vector<size_t> ntToTake;
  // Read objects of tuples [lo, hi) as readFiles would hand them to makeWord (fastp Read: name, seq, strand, quality)
  void build(size_t lo, size_t hi, vector<vector<Read*>>& out) const {
    out.assign(hi - lo, {});
    string qual;
    for (size_t t = lo; t < hi; t++) {
      const char* row = chars + t * wordLength;
      size_t at = headerUMI;
      for (size_t f = 0; f < ntToTake.size(); f++) {
        string name = "@t";
        if (f == 0 && headerUMI) name += "_" + string(row, headerUMI);
        string seq(row + at, ntToTake[f]);
        qual.assign(seq.size(), 'I');
        out[t - lo].push_back(new Read(name.c_str(), seq.c_str(), "+", qual.c_str()));
        at += ntToTake[f];
      }
    }
  }
  static void release(vector<vector<Read*>>& v) {
    for (auto& reads : v) for (Read* r : reads) delete r;
    v.clear();
  }
};

void writeMap(string const& path, map<size_t, size_t> const& m) {
  ofstream o(path, ios::out | ios::binary);
  for (auto const& kv : m) o << kv.first << ' ' << kv.second << '\n';
}
}  // namespace

int main(int argc, char** argv) {
  if (argc < 10) {
    fprintf(stderr, "usage: %s chars.bin n_tuples word_length distance edit maximum header_umi out_prefix nt_to_take...\n", argv[0]);
    return 2;
  }
  Tuples T;
  T.n = strtoull(argv[2], nullptr, 10);
  T.wordLength = strtoull(argv[3], nullptr, 10);
  size_t const distance = strtoull(argv[4], nullptr, 10);
  bool const edit = atoi(argv[5]) != 0, maximum = atoi(argv[6]) != 0;
  T.headerUMI = strtoull(argv[7], nullptr, 10);
  string const prefix = argv[8];
  for (int i = 9; i < argc; i++) T.ntToTake.push_back(strtoull(argv[i], nullptr, 10));
  vector<char> buf(T.n * T.wordLength);
  {
    FILE* f = fopen(argv[1], "rb");
    if (!f || fread(buf.data(), 1, buf.size(), f) != buf.size()) { fprintf(stderr, "cannot read %s\n", argv[1]); return 2; }
    fclose(f);
  }
  T.chars = buf.data();
  ofstream log("/dev/null", ios::out | ios::binary);
  size_t const CHUNK = 1 << 16;
  vector<vector<Read*>> reads;

  // readData's loop (src/humid.cc:92-99)
  Trie<4, NLeaf> trie;
  size_t total = 0, usable = 0;
  double tWords = 0;
  for (size_t lo = 0; lo < T.n; lo += CHUNK) {
    size_t hi = std::min(T.n, lo + CHUNK);
    T.build(lo, hi, reads);
    double t0 = now();
    for (auto const& r : reads) {
      Word word {makeWord(r, T.ntToTake, T.headerUMI)};
      if (not word.filtered) { trie.add(word.data); usable++; }
      total++;
    }
    tWords += now() - t0;
    Tuples::release(reads);
  }
  double t0 = now();
  size_t unique = edit ? findEditNeighbours(trie, distance, log) : findHammingNeighbours(trie, distance, log);
  double tNeigh = now() - t0;
  t0 = now();
  vector<Cluster*> clusters {findClusters(trie, maximum, log)};
  double tClus = now() - t0;

  // writeFiltered's decision per tuple (src/humid.cc:220-234) without the file output
  vector<uint8_t> keep((T.n + 7) / 8, 0);
  double tPass2 = 0;
  for (size_t lo = 0; lo < T.n; lo += CHUNK) {
    size_t hi = std::min(T.n, lo + CHUNK);
    T.build(lo, hi, reads);
    t0 = now();
    for (size_t t = lo; t < hi; t++) {
      Word word {makeWord(reads[t - lo], T.ntToTake, T.headerUMI)};
      if (not word.filtered) {
        Node<4, NLeaf>* node {trie.find(word.data)};
        if (!node->leaf->cluster->visited && node->leaf->cluster->maxLeaf == node->leaf) {
          keep[t >> 3] |= (uint8_t)(1u << (t & 7));
          node->leaf->cluster->visited = true;
        }
      }
    }
    tPass2 += now() - t0;
    Tuples::release(reads);
  }
  t0 = now();
  tuple<map<size_t, size_t>, map<size_t, size_t>> stats {runStatistics(trie, log)};
  map<size_t, size_t> cStats {clusterStats(clusters)};
  double tStats = now() - t0;

  {
    FILE* f = fopen((prefix + ".keep").c_str(), "wb");
    if (f) { fwrite(keep.data(), 1, keep.size(), f); fclose(f); }
  }
  writeMap(prefix + ".counts.dat", get<0>(stats));
  writeMap(prefix + ".neigh.dat", get<1>(stats));
  writeMap(prefix + ".clusters.dat", cStats);
  double const sum = tWords + tNeigh + tClus + tPass2 + tStats;
  printf("{\"total\": %zu, \"usable\": %zu, \"unique\": %zu, \"clusters\": %zu, \"seconds\": {\"words_trie\": %.6f, "
         "\"neighbours\": %.6f, \"clusters\": %.6f, \"second_pass\": %.6f, \"stats\": %.6f, \"hot_path\": %.6f}}\n",
         total, usable, unique, clusters.size(), tWords, tNeigh, tClus, tPass2, tStats, sum);
  size_t const nClusters = clusters.size();
  (void)nClusters;
  freeClusters(clusters);
  return 0;
}
