/* TEST INFRASTRUCTURE ONLY (oracle/_ref build).
 *
 * The reference links the system libdeflate (only lib/fastp/src/libdeflate.h is vendored;
 * used at lib/fastp/src/writer.cpp:66,110,112,133). libdeflate is absent from this image,
 * so the four entry points the reference's Writer calls are provided on top of zlib.
 * This only changes the compressed *encoding* of `.gz` outputs; parity is always checked
 * on decompressed bytes.
 */
#include <stdlib.h>
#include <string.h>
#include <zlib.h>

struct libdeflate_compressor { int level; };

struct libdeflate_compressor *libdeflate_alloc_compressor(int level) {
  struct libdeflate_compressor *c = malloc(sizeof *c);
  if (c) c->level = level;
  return c;
}

void libdeflate_free_compressor(struct libdeflate_compressor *c) { free(c); }

size_t libdeflate_gzip_compress_bound(struct libdeflate_compressor *c, size_t n) {
  (void)c;
  return compressBound(n) + 64;
}

size_t libdeflate_gzip_compress(struct libdeflate_compressor *c, const void *in, size_t n,
                                void *out, size_t avail) {
  z_stream zs;
  memset(&zs, 0, sizeof zs);
  if (deflateInit2(&zs, c->level, Z_DEFLATED, 15 + 16 /* gzip wrapper */, 8,
                   Z_DEFAULT_STRATEGY) != Z_OK)
    return 0;
  zs.next_in = (Bytef *)in;
  zs.avail_in = (uInt)n;
  zs.next_out = (Bytef *)out;
  zs.avail_out = (uInt)avail;
  int rc = deflate(&zs, Z_FINISH);
  size_t written = zs.total_out;
  deflateEnd(&zs);
  return rc == Z_STREAM_END ? written : 0;
}
