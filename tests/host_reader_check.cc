// Test helper (tests/test_host_reader.py): prints the records of a FastQ file as seen by the
// line-by-line reader (mode "old", host/fastq_io.h) or by the block reader (mode "block <bytes>",
// host/fastq_block.h), one field per line, so that the two can be compared byte for byte.
#include <cstdlib>
#include <iostream>

#include "../host/fastq_block.h"

using namespace humid_host;

int main(int argc, char **argv) {
  if (argc < 3) return 2;
  std::string mode = argv[1], path = argv[2];
  size_t n = 0;
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
  }
  std::cout << "records " << n << "\n";
  return 0;
}
