// A deflate decoder for the `humid` executable's .gz inputs — SURVEY.md §8(f) item 3.
//
// The reference inflates every .gz with isa-l's igzip (lib/fastp/src/fastqreader.cpp:79-133, :151-181: gzip header
// read by isal_read_gzip_header, crc_flag = ISAL_GZIP_NO_HDR_VER, member after member until the file ends, anything
// that is not a gzip header behind a member is an error).  zlib's inflate, which this host used so far, decodes one
// symbol per table walk from a 32-bit hold; on FastQ text it is what bounds both passes of the executable.  This
// decoder is written for that text (RFC 1951, nothing else assumed):
//   * a 64-bit hold refilled by one unaligned 8-byte load per symbol, never byte by byte;
//   * one look-up in a 2^11-entry table resolves every literal / length code of up to 11 bits together with its
//     extra-bit count and base value (longer codes: one more look-up in a sub-table), 2^9 entries for distances;
//   * matches are copied 8 bytes at a time (distance >= 8), as a byte fill (distance 1) or byte-wise;
//   * the loop runs over the caller's whole input / output windows and can be suspended at any symbol, so the
//     file is streamed through two fixed buffers whatever the size of a deflate block.
// GzFastReader (below) wraps it for gzip files: header, members one after the other, CRC-32 and ISIZE checked
// like the reference does.  Errors end the program through error_exit with the reference's message.
//
// Nothing here is taken from zlib, libdeflate or isa-l; the table layout is this file's own.
#pragma once
#include <zlib.h>  // crc32 only

#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "fastq_io.h"

namespace humid_host {

class Inflate {
 public:
  enum Status { kDone, kNeedInput, kNeedOutput, kError, kBoundary };

  Inflate() { reset(); }

  // Before the first block of a deflate stream.
  void reset() {
    hold_ = 0;
    bits_ = 0;
    state_ = kHeader;
    final_ = false;
    stored_left_ = 0;
    error_ = nullptr;
  }
  const char *error() const { return error_ ? error_ : ""; }

  // Whole bytes sitting unused in the hold after kDone (a gzip member ends on a byte boundary): the caller moves
  // its input pointer back by this many.
  unsigned unread() {
    const unsigned n = bits_ >> 3;
    bits_ = 0;
    hold_ = 0;
    return n;
  }

  // Decodes from [in, in_end) into [out, out_end).
  //   * kInPad readable bytes must follow in_end (their value does not matter);
  //   * eof: no input will follow in_end.  Without eof the call returns kNeedInput while fewer than kInSlack bytes
  //     are left, so that no symbol and no block header is ever cut by the end of the window;
  //   * [window, out) is the output produced so far that is still in memory: matches reach back into it, a
  //     distance beyond it is an error.  kNeedOutput comes back when fewer than kOutSlack bytes of room are left.
  static constexpr size_t kInPad = 32;
  static constexpr size_t kInSlack = 1024;  // > a dynamic block header (at most 14 + 57 + 320 * 14 bits)
  static constexpr size_t kOutSlack = 320;  // three look-ups of literals (4-byte stores) + a match of 258 (+ 7)
  //   * boundary (may be null): the call ends with kBoundary in front of the first block header that starts at or
  //     behind this input position; the stream goes on from there with the next call.
  Status run(const uint8_t *&in, const uint8_t *in_end, bool eof, const uint8_t *window, uint8_t *&out, uint8_t *out_end,
             const uint8_t *boundary = nullptr) {
    return run_t<uint8_t>(in, in_end, eof, window, out, out_end, boundary);
  }
  // The same with 16-bit output elements (host/inflate_parallel.h): a deflate stream entered in the middle copies
  // from 32 KiB it has not seen, so the caller puts 32768 place holders (values >= 0x8000) in front of `out`, they
  // travel through the matches like bytes, and whoever knows the real bytes later replaces them.
  Status run_marked(const uint8_t *&in, const uint8_t *in_end, bool eof, const uint16_t *window, uint16_t *&out,
                    uint16_t *out_end, const uint8_t *boundary = nullptr) {
    return run_t<uint16_t>(in, in_end, eof, window, out, out_end, boundary);
  }

  // Enter a stream at bit `bit` (0..7, counted from the low end) of *in: a block header must start there.
  void start_at(const uint8_t *&in, unsigned bit) {
    reset();
    if (bit) {
      hold_ = (uint64_t)*in++ >> bit;
      bits_ = 8 - bit;
    }
  }
  // Bits loaded from in front of the caller's `in` and not yet used: the stream stands 8 * offset(in) - pending_bits().
  unsigned pending_bits() const { return bits_; }
  bool stream_finished() const { return state_ == kFinished; }

  // Does a dynamic-code block header that a compressor would write start at bit `bit` of p?  (Block finder of the
  // parallel reader: not final, both code sets complete.)  The header is read up to `end`; 16 bytes behind `end`
  // must be readable.
  bool probe_dynamic_header(const uint8_t *p, const uint8_t *end, unsigned bit);

 private:
  // Table entries, 32 bits:
  //   [3:0]   bits this look-up uses up
  //   [5:4]   literals: how many (1..3)
  //   [7:6]   0 literals | kBase | kSub | kOther
  //   [31:8]  literals: the bytes, first one lowest
  //           kBase:  [23:8] base length / distance, [27:24] number of extra bits behind the code;
  //                   [4:0] counts the code AND its extra bits, [31:28] the code alone
  //           kSub:   [23:8] first entry of the sub-table, [27:24] its index bits
  //           kOther: kEndOfBlock or kInvalid
  // One look-up of the literal/length table yields up to three literals when their codes fit its index together
  // (FastQ text is literals in short codes — A, C, G, T and a few quality values — so this is the common case).
  enum : uint32_t {
    kBase = 1u << 6,
    kSub = 2u << 6,
    kOther = 3u << 6,
    kKindMask = 3u << 6,
    kEndOfBlock = kOther | (1u << 8),
    kInvalid = kOther | 1u,  // (uses up a bit so that a careless loop would still advance)
  };
  static uint32_t literal_entry(uint32_t byte) { return (1u << 4) | (byte << 8); }
  static uint32_t base_entry(uint32_t base, uint32_t extra) { return kBase | (base << 8) | (extra << 24); }
  static uint32_t with_length(uint32_t payload, uint32_t code_bits) {
    if ((payload & kKindMask) == kBase) return payload | (code_bits + ((payload >> 24) & 15u)) | (code_bits << 28);
    return payload | code_bits;
  }
  static constexpr int kLitBits = 11, kDistBits = 9, kPreBits = 7;
  static constexpr int kLitCap = 1 << 12, kDistCap = 1 << 11;

  enum State { kHeader, kStored, kCodes, kFinished };

  static uint64_t load64(const uint8_t *p) {
    uint64_t v;
    std::memcpy(&v, p, 8);
    return v;
  }
  static void store32(uint8_t *p, uint32_t v) { std::memcpy(p, &v, 4); }
  static void store64(uint8_t *p, uint64_t v) { std::memcpy(p, &v, 8); }

  Status fail(const char *what) {
    error_ = what;
    return kError;
  }

  // `n` (<= 32) bits into the hold byte by byte, never past `end`: used for block headers only.
  bool need(const uint8_t *&in, const uint8_t *end, unsigned n) {
    while (bits_ < n) {
      if (in >= end) return false;
      hold_ |= (uint64_t)*in++ << bits_;
      bits_ += 8;
    }
    return true;
  }
  uint32_t take(unsigned n) {
    const uint32_t v = (uint32_t)(hold_ & ((1ull << n) - 1));
    hold_ >>= n;
    bits_ -= n;
    return v;
  }

  static uint32_t reverse_bits(uint32_t code, int len) {
    uint32_t r = 0;
    for (int i = 0; i < len; i++) r |= ((code >> i) & 1u) << (len - 1 - i);
    return r;
  }

  // Canonical Huffman code of lens[0..n) -> look-up table, one symbol per entry.  payload[s] = the entry of symbol
  // s without its bit count.  An over-subscribed set of lengths is an error; an incomplete one leaves kInvalid
  // entries (an encoder may send a single distance code, RFC 1951 3.2.7).
  bool build(const uint8_t *lens, int n, const uint32_t *payload, int index_bits, uint32_t *table, int cap,
             bool *complete = nullptr) {
    int count[16] = {0};
    for (int s = 0; s < n; s++) count[lens[s]]++;
    count[0] = 0;
    int left = 1;
    for (int len = 1; len <= 15; len++) {
      left = 2 * left - count[len];
      if (left < 0) return false;
    }
    if (complete) *complete = left == 0;
    uint32_t next_code[16];
    uint32_t code = 0;
    for (int len = 1; len <= 15; len++) {
      code = (code + (uint32_t)count[len - 1]) << 1;
      next_code[len] = code;
    }
    const int main_size = 1 << index_bits;
    for (int i = 0; i < main_size; i++) table[i] = kInvalid;
    uint8_t deepest[1 << kLitBits];  // per index of the main table: the longest code that starts with it
    bool any_long = false;
    uint32_t reversed[320];
    for (int s = 0; s < n; s++) {
      const int len = lens[s];
      if (!len) continue;
      const uint32_t r = reverse_bits(next_code[len]++, len);
      reversed[s] = r;
      if (len <= index_bits) {
        const uint32_t e = with_length(payload[s], (uint32_t)len);
        for (uint32_t i = r; i < (uint32_t)main_size; i += 1u << len) table[i] = e;
      } else {
        if (!any_long) std::memset(deepest, 0, (size_t)main_size);
        any_long = true;
        uint8_t &d = deepest[r & (main_size - 1)];
        if (len > d) d = (uint8_t)len;
      }
    }
    if (!any_long) return true;
    int next_free = main_size;
    for (int i = 0; i < main_size; i++) {
      if (!deepest[i]) continue;
      const int sub_bits = deepest[i] - index_bits;
      if (next_free + (1 << sub_bits) > cap) return false;
      table[i] = kSub | ((uint32_t)next_free << 8) | ((uint32_t)sub_bits << 24);
      for (int k = 0; k < (1 << sub_bits); k++) table[next_free + k] = kInvalid;
      next_free += 1 << sub_bits;
    }
    for (int s = 0; s < n; s++) {
      const int len = lens[s];
      if (len <= index_bits) continue;
      const uint32_t ptr = table[reversed[s] & (main_size - 1)];
      const uint32_t first = (ptr >> 8) & 0xFFFFu, sub_size = 1u << ((ptr >> 24) & 15u);
      const uint32_t e = with_length(payload[s], (uint32_t)(len - index_bits));
      for (uint32_t i = reversed[s] >> index_bits; i < sub_size; i += 1u << (len - index_bits)) table[first + i] = e;
    }
    return true;
  }

  // Second and third literal into the entries of the literal/length table whose index bits hold them whole.
  void pack_literals() {
    constexpr uint32_t size = 1u << kLitBits;
    uint32_t single[size];
    std::memcpy(single, lit_, sizeof(single));
    for (uint32_t i = 0; i < size; i++) {
      uint32_t e = single[i];
      if (e & kKindMask) continue;
      uint32_t used = e & 15u, bytes = e >> 8, n = 1;
      while (n < 3 && used < (uint32_t)kLitBits) {
        const uint32_t next = single[i >> used];  // the index bits above `used`; the bits above those read as 0
        if ((next & kKindMask) || used + (next & 15u) > (uint32_t)kLitBits) break;
        bytes |= (next >> 8) << (8 * n);
        used += next & 15u;
        n++;
      }
      lit_[i] = used | (n << 4) | (bytes << 8);
    }
  }

  struct Payloads {
    uint32_t litlen[288], dist[32], pre[19];
    Payloads() {
      static const uint16_t len_base[29] = {3,  4,  5,  6,  7,  8,  9,  10, 11,  13,  15,  17,  19,  23, 27,
                                            31, 35, 43, 51, 59, 67, 83, 99, 115, 131, 163, 195, 227, 258};
      static const uint8_t len_extra[29] = {0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 0};
      for (uint32_t s = 0; s < 256; s++) litlen[s] = literal_entry(s);
      litlen[256] = kEndOfBlock;
      for (uint32_t s = 0; s < 29; s++) litlen[257 + s] = base_entry(len_base[s], len_extra[s]);
      litlen[286] = litlen[287] = kInvalid & ~15u;
      uint32_t base = 1;
      for (uint32_t s = 0; s < 30; s++) {
        const uint32_t extra = s < 2 ? 0 : s / 2 - 1;
        dist[s] = base_entry(base, extra);
        base += 1u << extra;
      }
      dist[30] = dist[31] = kInvalid & ~15u;
      for (uint32_t s = 0; s < 19; s++) pre[s] = literal_entry(s);
    }
  };
  static const Payloads &payloads() {
    static const Payloads p;
    return p;
  }

  bool fixed_tables() {
    uint8_t lens[288];
    for (int s = 0; s < 288; s++) lens[s] = s < 144 ? 8 : s < 256 ? 9 : s < 280 ? 7 : 8;
    if (!build(lens, 288, payloads().litlen, kLitBits, lit_, kLitCap)) return false;
    pack_literals();
    for (int s = 0; s < 32; s++) lens[s] = 5;
    return build(lens, 32, payloads().dist, kDistBits, dist_, kDistCap);
  }

  Status block_header(const uint8_t *&in, const uint8_t *in_end);
  template <class T>
  Status stored(const uint8_t *&in, const uint8_t *in_end, bool eof, T *&out, T *out_end);
  template <class T>
  Status run_t(const uint8_t *&in, const uint8_t *in_end, bool eof, const T *window, T *&out, T *out_end, const uint8_t *boundary);
  // n literals of a table entry / `length` elements from `distance` back, for both element types
  static void put_literals(uint8_t *op, uint32_t bytes) { store32(op, bytes); }
  static void put_literals(uint16_t *op, uint32_t bytes) {
    op[0] = (uint16_t)(bytes & 0xFFu);
    op[1] = (uint16_t)((bytes >> 8) & 0xFFu);
    op[2] = (uint16_t)((bytes >> 16) & 0xFFu);
  }
  static void copy_match(uint8_t *op, uint32_t distance, uint32_t length);
  static void copy_match(uint16_t *op, uint32_t distance, uint32_t length);

  uint64_t hold_;
  unsigned bits_;
  State state_;
  bool final_;
  uint32_t stored_left_;
  const char *error_;
  bool lit_complete_ = false, dist_complete_ = false;
  uint32_t lit_[kLitCap], dist_[kDistCap];
};

// BTYPE / BFINAL and, for a dynamic block, the two code-length sets (RFC 1951 3.2.3 - 3.2.7).  Read with the
// careful byte-wise loader; the caller made sure the header cannot be cut by the end of the window unless the
// file ends there.
inline Inflate::Status Inflate::block_header(const uint8_t *&in, const uint8_t *in_end) {
  hold_ &= bits_ < 64 ? (1ull << bits_) - 1 : ~0ull;  // drop look-ahead bits of the symbol loop's wide loads
  if (!need(in, in_end, 3)) return fail("deflate data ends inside a block header");
  final_ = take(1) != 0;
  const uint32_t type = take(2);
  if (type == 0) {
    take(bits_ & 7);  // to the next byte boundary
    if (!need(in, in_end, 32)) return fail("deflate data ends inside a stored block's length");
    const uint32_t len = take(16), nlen = take(16);
    if ((len ^ nlen) != 0xFFFFu) return fail("stored block length check failed");
    stored_left_ = len;
    state_ = kStored;
    return kDone;
  }
  if (type == 1) {
    if (!fixed_tables()) return fail("internal: fixed code tables");
    state_ = kCodes;
    return kDone;
  }
  if (type == 3) return fail("reserved block type");
  if (!need(in, in_end, 14)) return fail("deflate data ends inside a block header");
  const uint32_t n_lit = take(5) + 257, n_dist = take(5) + 1, n_pre = take(4) + 4;
  if (n_lit > 286 || n_dist > 30) return fail("too many length or distance codes");
  static const uint8_t order[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};
  uint8_t pre_lens[19] = {0};
  for (uint32_t i = 0; i < n_pre; i++) {
    if (!need(in, in_end, 3)) return fail("deflate data ends inside a block header");
    pre_lens[order[i]] = (uint8_t)take(3);
  }
  uint32_t pre[1 << kPreBits];
  if (!build(pre_lens, 19, payloads().pre, kPreBits, pre, 1 << kPreBits)) return fail("invalid code lengths set");
  uint8_t lens[286 + 30 + 138];
  uint32_t have = 0;
  const uint32_t want = n_lit + n_dist;
  while (have < want) {
    if (!need(in, in_end, 7 + 7) && bits_ == 0) return fail("deflate data ends inside a block header");
    const uint32_t e = pre[hold_ & ((1u << kPreBits) - 1)];
    if ((e & kKindMask) || (e & 15u) > bits_) return fail("invalid code lengths set");
    take(e & 15u);
    const uint32_t sym = (e >> 8) & 0xFFu;
    if (sym < 16) {
      lens[have++] = (uint8_t)sym;
      continue;
    }
    uint32_t repeat;
    uint8_t value = 0;
    const unsigned xbits = sym == 16 ? 2 : sym == 17 ? 3 : 7;
    if (bits_ < xbits && !need(in, in_end, xbits)) return fail("deflate data ends inside a block header");
    if (sym == 16) {
      if (!have) return fail("invalid bit length repeat");
      value = lens[have - 1];
      repeat = 3 + take(2);
    } else if (sym == 17) {
      repeat = 3 + take(3);
    } else {
      repeat = 11 + take(7);
    }
    if (have + repeat > want) return fail("invalid bit length repeat");
    std::memset(lens + have, value, repeat);
    have += repeat;
  }
  if (!lens[256]) return fail("invalid code -- missing end-of-block");
  if (!build(lens, (int)n_lit, payloads().litlen, kLitBits, lit_, kLitCap, &lit_complete_)) return fail("invalid literal/lengths set");
  pack_literals();
  if (!build(lens + n_lit, (int)n_dist, payloads().dist, kDistBits, dist_, kDistCap, &dist_complete_)) return fail("invalid distances set");
  state_ = kCodes;
  return kDone;
}

inline bool Inflate::probe_dynamic_header(const uint8_t *p, const uint8_t *end, unsigned bit) {
  if (end - p < 10) return false;
  // cheap tests on the first 17 bits: BFINAL = 0, BTYPE = 2, at most 286 / 30 codes
  const uint64_t v = load64(p) >> bit;
  if ((v & 7u) != 4u) return false;
  if (((v >> 3) & 31u) > 29u || ((v >> 8) & 31u) > 29u) return false;
  // the code-length code must be complete: sum of 2^(7 - len) over its 4..19 three-bit lengths == 2^7
  const unsigned n_pre = (unsigned)((v >> 13) & 15u) + 4;
  unsigned at = bit + 17, kraft = 0;
  for (unsigned i = 0; i < n_pre; i += 16) {
    uint64_t w = load64(p + (at >> 3)) >> (at & 7);
    const unsigned n = n_pre - i < 16 ? n_pre - i : 16;
    for (unsigned k = 0; k < n; k++, w >>= 3)
      if (w & 7u) kraft += 128u >> (w & 7u);
    at += 3 * n;
  }
  if (kraft != 128u) return false;
  // the whole header through the ordinary parser
  const uint8_t *in = p;
  start_at(in, bit);
  if (block_header(in, end) != kDone || state_ != kCodes) return false;
  return lit_complete_ && dist_complete_;
}

// The bytes of a stored block: from the hold first (whole bytes only: the length fields ended on a byte
// boundary), then straight from the window.
template <class T>
inline Inflate::Status Inflate::stored(const uint8_t *&in, const uint8_t *in_end, bool eof, T *&out, T *out_end) {
  hold_ &= bits_ < 64 ? (1ull << bits_) - 1 : ~0ull;
  while (stored_left_ && bits_ >= 8) {
    if (out >= out_end) return kNeedOutput;
    *out++ = (T)take(8);
    stored_left_--;
  }
  if (stored_left_) {
    bits_ = 0;
    hold_ = 0;
    const size_t n = std::min<size_t>(std::min<size_t>(stored_left_, (size_t)(in_end - in)), (size_t)(out_end - out));
    if (sizeof(T) == 1) std::memcpy(out, in, n);
    else
      for (size_t i = 0; i < n; i++) out[i] = (T)in[i];
    in += n;
    out += n;
    stored_left_ -= (uint32_t)n;
    if (stored_left_) {
      if (out >= out_end) return kNeedOutput;
      return eof ? fail("deflate data ends inside a stored block") : kNeedInput;
    }
  }
  state_ = final_ ? kFinished : kHeader;
  return kDone;
}

inline void Inflate::copy_match(uint8_t *op, uint32_t distance, uint32_t length) {
  const uint8_t *from = op - distance;
  uint8_t *const stop = op + length;
  if (distance >= 8) {
    // 16 bytes whatever the length (most matches of FastQ text are shorter), 8 at a time so that a distance of
    // 8..15 repeats its pattern; up to 15 bytes past `stop` are written: inside kOutSlack
    store64(op, load64(from));
    store64(op + 8, load64(from + 8));
    if (length > 16) {
      op += 16;
      from += 16;
      do {
        store64(op, load64(from));
        op += 8;
        from += 8;
      } while (op < stop);
    }
  } else if (distance == 1) {  // a run of one byte: quality strings are full of them
    const uint64_t v = 0x0101010101010101ull * *from;
    do {
      store64(op, v);
      op += 8;
    } while (op < stop);
  } else {
    do *op++ = *from++;
    while (op < stop);
  }
}

inline void Inflate::copy_match(uint16_t *op, uint32_t distance, uint32_t length) {
  const uint16_t *from = op - distance;
  uint16_t *const stop = op + length;
  if (distance >= 8) {  // sixteen elements whatever the length, eight (16 bytes) at a time; up to 15 past `stop`
    std::memcpy(op, from, 16);
    std::memcpy(op + 8, from + 8, 16);
    if (length > 16) {
      op += 16;
      from += 16;
      do {
        std::memcpy(op, from, 16);
        op += 8;
        from += 8;
      } while (op < stop);
    }
  } else if (distance >= 4) {  // four elements at a time; up to 3 past `stop`
    do {
      std::memcpy(op, from, 8);
      op += 4;
      from += 4;
    } while (op < stop);
  } else {
    do *op++ = *from++;
    while (op < stop);
  }
}

template <class T>
inline Inflate::Status Inflate::run_t(const uint8_t *&in, const uint8_t *in_end, bool eof, const T *window, T *&out, T *out_end,
                                      const uint8_t *boundary) {
  while (true) {
    if (state_ == kFinished) return kDone;
    if (state_ == kHeader) {
      if (boundary && in - ((bits_ + 7) >> 3) >= boundary) return kBoundary;
      if (!eof && (size_t)(in_end - in) < kInSlack) return kNeedInput;
      if (block_header(in, in_end) != kDone) return kError;
      continue;
    }
    if (state_ == kStored) {
      const Status st = stored<T>(in, in_end, eof, out, out_end);
      if (st != kDone) return st;
      continue;
    }
    // ---- the symbol loop of a block with Huffman codes ----
    if ((size_t)(in_end - in) < 16 && !eof) return kNeedInput;
    if ((size_t)(out_end - out) < kOutSlack) return kNeedOutput;
    uint64_t hold = hold_;
    unsigned bits = bits_;
    const uint8_t *ip = in;
    T *op = out;
    // one round of the loop loads 8 bytes twice, at ip <= in_limit + 14, and writes less than kOutSlack bytes
    const uint8_t *const in_limit = eof ? in_end : in_end - 24;
    T *const out_limit = out_end - kOutSlack;
    constexpr uint64_t lit_mask = (1u << kLitBits) - 1, dist_mask = (1u << kDistBits) - 1;
    Status st = kDone;
    bool end_of_block = false;
    while (true) {
      if (ip > in_limit) {
        // at the end of the file the padding behind in_end is what gets loaded; data that is complete ends
        // with the end-of-block code before more bits are USED than the file had
        if (!eof) {
          st = kNeedInput;
          break;
        }
        if (ip - (bits >> 3) > in_end) {
          st = fail("deflate data ends inside a block");
          break;
        }
        // (not yet: so ip <= in_end + 7, and both loads below stay inside the kInPad readable bytes)
      }
      if (op > out_limit) {
        st = kNeedOutput;
        break;
      }
      // the look-up needs the low kLitBits bits only, which a round nearly always leaves in the hold: it is started
      // before the load so that the two do not wait for each other
      uint32_t e = lit_[hold & lit_mask];
      const unsigned had = bits;
      hold |= load64(ip) << bits;
      ip += (63 - bits) >> 3;
      bits |= 56;
      if (had < (unsigned)kLitBits) e = lit_[hold & lit_mask];
      // up to three look-ups of literals on one load: each uses at most kLitBits bits (15 through a sub-table,
      // which ends the run), 56 were there
      if (!(e & kKindMask)) {
        put_literals(op, e >> 8);
        op += (e >> 4) & 3u;
        hold >>= e & 15u;
        bits -= e & 15u;
        e = lit_[hold & lit_mask];
        if (!(e & kKindMask)) {
          put_literals(op, e >> 8);
          op += (e >> 4) & 3u;
          hold >>= e & 15u;
          bits -= e & 15u;
          e = lit_[hold & lit_mask];
          if (!(e & kKindMask)) {
            put_literals(op, e >> 8);
            op += (e >> 4) & 3u;
            hold >>= e & 15u;
            bits -= e & 15u;
            continue;
          }
        }
        // something else follows the literals: its entry stays right when the hold is topped up
        hold |= load64(ip) << bits;
        ip += (63 - bits) >> 3;
        bits |= 56;
      }
      if ((e & kKindMask) == kSub) {
        hold >>= kLitBits;
        bits -= kLitBits;
        e = lit_[((e >> 8) & 0xFFFFu) + (hold & ((1u << ((e >> 24) & 15u)) - 1))];
        if (!(e & kKindMask)) {  // a literal with a long code
          *op++ = (T)((e >> 8) & 0xFFu);
          hold >>= e & 15u;
          bits -= e & 15u;
          continue;
        }
      }
      if ((e & kKindMask) != kBase) {
        hold >>= e & 15u;
        bits -= e & 15u;
        if (e & (1u << 8)) end_of_block = true;
        else st = fail("invalid literal/length code");
        break;
      }
      // code and extra bits leave the hold in one shift
      const uint32_t length = ((e >> 8) & 0xFFFFu) + (uint32_t)((hold & ((1u << (e & 31u)) - 1)) >> (e >> 28));
      hold >>= e & 31u;
      bits -= e & 31u;
      uint32_t d = dist_[hold & dist_mask];
      if ((d & kKindMask) == kSub) {
        hold >>= kDistBits;
        bits -= kDistBits;
        d = dist_[((d >> 8) & 0xFFFFu) + (hold & ((1u << ((d >> 24) & 15u)) - 1))];
      }
      if ((d & kKindMask) != kBase) {
        st = fail("invalid distance code");
        break;
      }
      const uint32_t distance = ((d >> 8) & 0xFFFFu) + (uint32_t)((hold & ((1u << (d & 31u)) - 1)) >> (d >> 28));
      hold >>= d & 31u;
      bits -= d & 31u;
      if (distance > (size_t)(op - window)) {
        st = fail("invalid distance too far back");
        break;
      }
      copy_match(op, distance, length);
      op += length;
    }
    hold_ = hold;
    bits_ = bits;
    in = ip;
    out = op;
    if (end_of_block) {
      if (eof && ip - (bits >> 3) > in_end) return fail("deflate data ends inside a block");
      state_ = final_ ? kFinished : kHeader;
      continue;
    }
    return st;
  }
}

// gzread for a .gz file, on the decoder above.  Members follow one another until the file ends; like the reference
// (lib/fastp/src/fastqreader.cpp:100-131) whatever follows a member must be another gzip header.  CRC-32 and ISIZE
// of every member are checked (the reference sets ISAL_GZIP_NO_HDR_VER, :161).
class GzFastReader {
 public:
  explicit GzFastReader(const std::string &path) : path_(path), in_(kInBuf + Inflate::kInPad), out_(kWindow + kOutBuf) {
    fp_ = std::fopen(path.c_str(), "rb");
    if (!fp_) error_exit("Failed to open file: " + path);
    std::setvbuf(fp_, nullptr, _IONBF, 0);
    window_ = op_ = taken_ = out_.data() + kWindow;
    fill();
    if (in_len_ == 0) {  // an empty file: nothing to read (gzread agrees; the reference stops at its first header read)
      at_end_ = true;
      return;
    }
    if (!member_header()) error_exit("igzip: Error invalid gzip header found: " + path_);
  }
  ~GzFastReader() {
    if (fp_) std::fclose(fp_);
  }
  GzFastReader(const GzFastReader &) = delete;
  GzFastReader &operator=(const GzFastReader &) = delete;

  // Up to `room` inflated bytes into dst; 0 at the end of the file.
  long read(char *dst, size_t room) {
    size_t got = 0;
    while (got < room) {
      if (taken_ == op_) {
        if (at_end_) break;
        if (got) break;  // hand over what there is before decoding more
        decode();
        continue;
      }
      const size_t n = std::min<size_t>(room - got, (size_t)(op_ - taken_));
      std::memcpy(dst + got, taken_, n);
      taken_ += n;
      got += n;
    }
    return (long)got;
  }

 private:
  static constexpr size_t kInBuf = 1u << 20, kWindow = 32768, kOutBuf = (1u << 19) + Inflate::kOutSlack;

  [[noreturn]] void corrupt(const char *what) {
    error_exit("igzip: encountered while decompressing file: " + path_ + " (" + what + ")");
  }

  // Unused input to the front of the buffer, the file's next bytes behind it.
  void fill() {
    if (file_done_) return;
    const size_t left = in_len_ - in_pos_;
    if (in_pos_) std::memmove(in_.data(), in_.data() + in_pos_, left);
    in_pos_ = 0;
    in_len_ = left;
    while (in_len_ < kInBuf) {
      const size_t got = std::fread(in_.data() + in_len_, 1, kInBuf - in_len_, fp_);
      if (got == 0) {
        file_done_ = true;
        break;
      }
      in_len_ += got;
    }
    std::memset(in_.data() + in_len_, 0, Inflate::kInPad);
  }

  // RFC 1952 2.3: magic, CM = 8, FLG, MTIME, XFL, OS, then the optional fields FLG announces.  false: not a header.
  bool member_header() {
    while (true) {
      const uint8_t *p = in_.data() + in_pos_;
      const size_t n = in_len_ - in_pos_;
      size_t at = 10;
      bool cut = n < at;
      if (!cut) {
        if (p[0] != 0x1f || p[1] != 0x8b || p[2] != 8 || (p[3] & 0xe0)) return false;
        const uint8_t flg = p[3];
        if (flg & 4) {  // FEXTRA
          if (n < at + 2) cut = true;
          else at += 2 + (size_t)(p[at] | (p[at + 1] << 8));
        }
        for (int field = 0; field < 2 && !cut; field++) {  // FNAME, FCOMMENT: zero-terminated
          if (!(flg & (field ? 16 : 8))) continue;
          const void *z = at < n ? std::memchr(p + at, 0, n - at) : nullptr;
          if (!z) cut = true;
          else at = (size_t)((const uint8_t *)z - p) + 1;
        }
        if (!cut && (flg & 2)) at += 2;  // FHCRC
        if (at > n) cut = true;
      }
      if (!cut) {
        in_pos_ += at;
        inflate_.reset();
        crc_ = crc32(0L, Z_NULL, 0);
        member_bytes_ = 0;
        window_ = op_;  // matches do not reach into the member before
        return true;
      }
      if (file_done_ || (in_pos_ == 0 && in_len_ == kInBuf)) return false;  // cut by the end of the file / larger than the buffer
      fill();
    }
  }

  // Decodes until there is something to hand out or the file ends.
  void decode() {
    // slide: the last 32 KiB of output stay in front of the new data
    if (op_ != out_.data() + kWindow) {
      const size_t keep = std::min<size_t>((size_t)(op_ - window_), kWindow);
      std::memmove(out_.data() + kWindow - keep, op_ - keep, keep);
      window_ = out_.data() + kWindow - keep;
      op_ = taken_ = out_.data() + kWindow;
    }
    uint8_t *const out_end = out_.data() + out_.size();
    while (op_ == taken_ && !at_end_) {
      const uint8_t *ip = in_.data() + in_pos_;
      uint8_t *const before = op_;
      const Inflate::Status st = inflate_.run(ip, in_.data() + in_len_, file_done_, window_, op_, out_end);
      in_pos_ = (size_t)(ip - in_.data());
      if (op_ != before) {
        crc_ = crc32(crc_, before, (uInt)(op_ - before));
        member_bytes_ += (uint64_t)(op_ - before);
      }
      if (st == Inflate::kError) corrupt(inflate_.error());
      if (st == Inflate::kNeedInput) {
        if (file_done_) corrupt("unexpected end of file");
        fill();
      } else if (st == Inflate::kDone) {
        in_pos_ -= std::min<size_t>(inflate_.unread(), in_pos_);
        if (in_len_ - in_pos_ < 8) fill();
        if (in_len_ - in_pos_ < 8) corrupt("unexpected end of file");
        const uint8_t *t = in_.data() + in_pos_;
        const uint32_t crc = t[0] | (t[1] << 8) | (t[2] << 16) | ((uint32_t)t[3] << 24);
        const uint32_t isize = t[4] | (t[5] << 8) | (t[6] << 16) | ((uint32_t)t[7] << 24);
        if (crc != (uint32_t)crc_ || isize != (uint32_t)member_bytes_) corrupt("incorrect data check");
        in_pos_ += 8;
        if (in_pos_ == in_len_) fill();
        if (in_pos_ == in_len_) at_end_ = true;
        else if (!member_header()) error_exit("igzip: invalid gzip header found");
      }
      // kNeedOutput: op_ != taken_ now (the buffer is far larger than kOutSlack)
    }
  }

  std::string path_;
  FILE *fp_ = nullptr;
  std::vector<uint8_t> in_, out_;
  size_t in_pos_ = 0, in_len_ = 0;
  bool file_done_ = false, at_end_ = false;
  const uint8_t *window_;  // first byte matches may reach
  uint8_t *op_, *taken_;   // [taken_, op_) is decoded and not yet handed out
  Inflate inflate_;
  unsigned long crc_ = 0;
  uint64_t member_bytes_ = 0;
};

}  // namespace humid_host
