// Measurement helper: how fast do N threads parse one plain FastQ file (host/ingest_parallel.h) when the commit does next to
// nothing?  Usage: parse_bench <file.fastq> <threads> [<file2> ...]  (several files: one parser pool per file, side by side)
#include <chrono>
#include <cstdio>
#include "../host/ingest_parallel.h"
using namespace humid_host;
static void one(const char *path, unsigned threads) {
  for (int rep = 0; rep < 3; rep++) {
    auto t0 = std::chrono::steady_clock::now();
    size_t bases = 0;
    ParallelPlainParser p(path, 26, [](const RecordView &r, uint8_t *row) { size_t c = r.seq.n < 24 ? r.seq.n : 24; memcpy(row, r.seq.p, c); uint16_t n = c; memcpy(row + 24, &n, 2); });
    uint64_t n = p.run(threads, [&](uint64_t, const uint8_t *row, uint64_t, uint32_t) { bases += row[0]; return true; });
    double s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    printf("%s threads %u: %llu records %.3f s (ranges %llu reparsed %llu)\n", path, threads, (unsigned long long)n, s, (unsigned long long)p.ranges(), (unsigned long long)p.reparsed_ranges());
  }
}
int main(int argc, char **argv) {
  unsigned threads = (unsigned)atoi(argv[2]);
  std::vector<std::thread> files;
  files.emplace_back(one, argv[1], threads);
  for (int i = 3; i < argc; i++) files.emplace_back(one, argv[i], threads);
  for (auto &t : files) t.join();
}
