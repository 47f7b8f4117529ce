// Test helper (tests/test_host_reader.py): prints the records of a FastQ file as seen by the
// line-by-line reader (mode "old", host/fastq_io.h) or by the block reader (mode "block <bytes>",
// host/fastq_block.h), one field per line, so that the two can be compared byte for byte.
#include <cstdlib>
#include <cstring>
#include <iostream>

#include <chrono>

#include "../host/fastq_block.h"
#include "../host/fastq_rewrite.h"
#include "../host/ingest_parallel.h"

using namespace humid_host;

int main(int argc, char **argv) {
  if (argc < 3) return 2;
  std::string mode = argv[1], path = argv[2];
  size_t n = 0;
  if (mode == "rewrite") {  // rewrite <in> <out one thread> <out threads> <seed> <annotate 0|1> <threads>
    // pass 1 as the executable does it: count the records, note the offset of every REWRITE_CHUNK-th
    RecordMap map;
    map.lengths = path.size() < 3 || path.compare(path.size() - 3, 3, ".gz") != 0;
    uint64_t total = 0;
    {
      BlockReader reader(path, 64);
      RecordView r;
      while (reader.read(r)) map.note(total++, reader, r);
    }
    uint64_t x = std::strtoull(argv[5], nullptr, 10) * 0x9E3779B97F4A7C15ull + 1;
    std::vector<uint64_t> keep((total + 63) / 64 + 1);
    std::vector<uint32_t> ids(total + 1);
    for (uint64_t t = 0; t < total; t++) {
      x ^= x << 13; x ^= x >> 7; x ^= x << 17;
      if (x % 3 == 0) keep[t >> 6] |= 1ull << (t & 63);
      ids[t] = (uint32_t)(x >> 40);
    }
    const bool annotate = std::atoi(argv[6]) != 0;
    if (std::string(argv[3]) != "-") rewrite_file(path, argv[3], total, annotate ? nullptr : keep.data(), annotate ? ids.data() : nullptr);
    const auto t0 = std::chrono::steady_clock::now();
    rewrite_file_parallel(path, argv[4], total, annotate ? nullptr : keep.data(), annotate ? ids.data() : nullptr, map,
                          (unsigned)std::atoi(argv[7]));
    std::cerr << "parallel rewrite " << std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() << " s\n";
    size_t gathered = 0;
    for (uint8_t v : map.chunk_verbatim) gathered += v;
    std::cout << "records " << total << " chunks " << map.chunk_offset.size() << " gathered " << gathered << "\n";
    return 0;
  }
  if (mode == "hash_seq" || mode == "hash_par") {  // hash_seq <in> | hash_par <in> <threads>: one line per record
    auto fnv = [](const Span &s) {
      uint32_t h = 2166136261u;
      for (size_t i = 0; i < s.n; i++) h = (h ^ (uint8_t)s.p[i]) * 16777619u;
      return h;
    };
    auto extract = [&](const RecordView &r, uint8_t *row) {
      const uint32_t h[4] = {fnv(r.name), fnv(r.seq), fnv(r.strand), fnv(r.qual)};
      std::memcpy(row, h, 16);
    };
    auto print = [&](uint64_t t, const uint8_t *row, uint64_t at, uint32_t stored) {
      uint32_t h[4];
      std::memcpy(h, row, 16);
      std::cout << t << " @" << at << " +" << stored << " " << h[0] << " " << h[1] << " " << h[2] << " " << h[3] << "\n";
      return true;
    };
    if (mode == "hash_seq") {
      BlockReader reader(path);
      RecordView r;
      uint8_t row[16];
      while (reader.read(r)) {
        extract(r, row);
        print(n++, row, reader.offset_of(r), (uint32_t)reader.raw_length(r));
      }
    } else {
      ParallelPlainParser parser(path, 16, extract);
      n = parser.run((unsigned)std::atoi(argv[3]), print);
      std::cerr << "ranges " << parser.ranges() << " reparsed " << parser.reparsed_ranges() << "\n";
    }
    std::cout << "records " << n << "\n";
    return 0;
  }
  if (mode == "old") {
    FastqReader reader(path);
    Record r;
    while (reader.read(r)) {
      size_t off, len;
      extract_umi(r.name, off, len);
      std::cout << "[" << r.name << "][" << r.seq << "][" << r.strand << "][" << r.qual << "][" << off << "," << len << "]\n";
      n++;
    }
  } else {
    BlockReader reader(path, argc > 3 ? std::strtoull(argv[3], nullptr, 10) : (16u << 20));
    RecordView r;
    while (reader.read(r)) {
      size_t off, len;
      extract_umi(r.name.p, r.name.n, off, len);
      std::cout << "[" << std::string(r.name.p, r.name.n) << "][" << std::string(r.seq.p, r.seq.n) << "]["
                << std::string(r.strand.p, r.strand.n) << "][" << std::string(r.qual.p, r.qual.n) << "][" << off << ","
                << len << "]\n";
      n++;
    }
    std::cerr << "buffer " << reader.buffer_bytes() << "\n";
  }
  std::cout << "records " << n << "\n";
  return 0;
}
