// Test / measurement helper for host/inflate_fast.h (not part of the product):
//   inflate_check fast|zlib|parN <file.gz> [read_size]  -> "<bytes> <crc32 hex> <seconds>" on stdout, exit 0
//                                                         (parN: host/inflate_parallel.h with N threads, N = 1..64)
//   inflate_check dump <file.gz> <out>                 -> the inflated bytes written to <out> through the fast reader
// Errors leave through the reader's error_exit (message on stderr, non-zero exit status).
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

#include "../host/inflate_parallel.h"

int main(int argc, char **argv) {
  if (argc < 3) return 2;
  const std::string mode = argv[1], path = argv[2];
  const auto t0 = std::chrono::steady_clock::now();
  unsigned long crc = crc32(0L, Z_NULL, 0);
  uint64_t total = 0;
  if (mode == "dump") {
    FILE *o = std::fopen(argv[3], "wb");
    if (!o) return 2;
    humid_host::GzFastReader r(path);
    std::vector<char> buf(1 << 16);
    long got;
    while ((got = r.read(buf.data(), buf.size())) > 0) std::fwrite(buf.data(), 1, (size_t)got, o);
    std::fclose(o);
    return 0;
  }
  const size_t read_size = argc > 3 ? (size_t)std::atol(argv[3]) : (16u << 20);
  std::vector<char> buf(read_size);
  double crc_seconds = 0;
  auto account = [&](long got) {
    const auto c0 = std::chrono::steady_clock::now();
    crc = crc32(crc, reinterpret_cast<const Bytef *>(buf.data()), (uInt)got);
    crc_seconds += std::chrono::duration<double>(std::chrono::steady_clock::now() - c0).count();
    total += (uint64_t)got;
  };
  if (mode == "fast") {
    humid_host::GzFastReader r(path);
    long got;
    while ((got = r.read(buf.data(), buf.size())) > 0) account(got);
  } else if (mode.compare(0, 3, "par") == 0) {
    humid_host::ParallelGzReader r(path, (unsigned)std::max(1, std::atoi(mode.c_str() + 3)));
    long got;
    while ((got = r.read(buf.data(), buf.size())) > 0) account(got);
  } else {
    gzFile gz = gzopen(path.c_str(), "rb");
    if (!gz) return 2;
    gzbuffer(gz, 1u << 20);
    int got;
    while ((got = gzread(gz, buf.data(), (unsigned)buf.size())) > 0) account(got);
    if (got < 0) {
      std::fprintf(stderr, "zlib error\n");
      return 1;
    }
    gzclose(gz);
  }
  const double s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() - crc_seconds;
  std::printf("%llu %08lx %.4f\n", (unsigned long long)total, crc, s);
  return 0;
}
