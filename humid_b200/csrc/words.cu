// K1/K2 — word construction, distinct-word table with counts, rank order.
//
// Replaces makeWord/getNucleotides/makeStringSize_ (reference src/fastq.cc:57-66, :116-161),
// Trie::add (lib/trie/src/trieFunctions.tcc:45-59) and Trie::walk order (:129-143).
#include <cstdlib>
#include <algorithm>
#include <string>

#include "state.cuh"

namespace humid {

struct PackParams {
  uint32_t n, n_segs;
  uint32_t take[HUMID_MAX_SEGMENTS];
};

constexpr int PACK_THREADS = 256;
#ifndef HUMID_PACK_ROWS
#define HUMID_PACK_ROWS 1024
#endif
constexpr int PACK_ROWS = HUMID_PACK_ROWS;
#ifndef HUMID_PACK_LOADS
#define HUMID_PACK_LOADS 6
#endif
constexpr int PACK_LOADS = HUMID_PACK_LOADS;  // tuples per CTA: 4 x 16-byte loads in flight per thread while staging
#ifndef HUMID_PACK_DEFAULT_BULK
#define HUMID_PACK_DEFAULT_BULK 0
#endif
constexpr int PACK_ROWS_WIDE = 512;  // words of 32..63 letters: the tile still fits 48 KB of shared memory

// Four characters at a time (SIMD inside a 32-bit register).  A=0 C=1 G=2 T=3
// (src/fastq.cc:12) falls out of bits 1-2 of the ASCII code; a character is valid iff the
// letter that its code maps back to is the character itself (anything else — N, lower case,
// padding — filters the word, :151-158).  Returns the four 2-bit codes packed into 8 bits,
// first character most significant.
__device__ __forceinline__ uint32_t pack4(uint32_t x, bool &bad) {
  uint32_t c = (x >> 1) & 0x03030303u;  // A:0 C:1 G:3 T:2
  c ^= (c >> 1) & 0x01010101u;          // A:0 C:1 G:2 T:3
#ifdef HUMID_PACK4_PRMT
  const uint32_t sel = __byte_perm(c | (c >> 4), 0u, 0x4420u);
#else
  const uint32_t sel = (c & 0x3u) | ((c >> 4) & 0x30u) | ((c >> 8) & 0x300u) | ((c >> 12) & 0x3000u);
#endif
  bad |= __byte_perm(0x54474341u /* "ACGT" */, 0u, sel) != x;
  return (c * 0x40100401u) >> 24;  // b0<<6 | b1<<4 | b2<<2 | b3
}

// makeWord for one tuple whose n characters lie at `row` in shared memory; returns 1 when the tuple is usable.
template <bool WIDE>
__device__ __forceinline__ uint32_t pack_row(const uint8_t *row, const uint16_t *__restrict__ seg_len, uint64_t t,
                                             const PackParams &pp, uint64_t *__restrict__ keys_out,
                                             uint64_t *__restrict__ keys_hi_out) {
  uint64_t key = 0, key_hi = 0;
  bool bad = false;
  if (seg_len) {
    const uint16_t *sl = seg_len + t * pp.n_segs;
    for (uint32_t s = 0; s < pp.n_segs; s++) bad |= sl[s] < pp.take[s];
  }
  const uint32_t full = pp.n & ~3u;
  if ((pp.n & 3u) == 0) {  // rows are 4-byte aligned in shared memory
    const uint32_t *row4 = reinterpret_cast<const uint32_t *>(row);
    for (uint32_t q = 0; q < full; q += 4) {
      if (WIDE) key_hi = (key_hi << 8) | (key >> 56);
      key = (key << 8) | pack4(row4[q >> 2], bad);
    }
  } else {
    for (uint32_t q = 0; q < full; q += 4) {
      uint32_t x = row[q] | (row[q + 1] << 8) | (row[q + 2] << 16) | ((uint32_t)row[q + 3] << 24);
      if (WIDE) key_hi = (key_hi << 8) | (key >> 56);
      key = (key << 8) | pack4(x, bad);
    }
  }
  for (uint32_t q = full; q < pp.n; q++) {
    uint32_t one = pack4(row[q] | 0x41414100u, bad);  // pad with 'A' (code 0, valid)
    if (WIDE) key_hi = (key_hi << 2) | (key >> 62);
    key = (key << 2) | ((one >> 6) & 3u);
  }
  keys_out[t] = bad ? KEY_INVALID : key;
  if (WIDE) keys_hi_out[t] = bad ? KEY_INVALID : key_hi;
  return bad ? 0u : 1u;
}

// One CTA packs PACK_ROWS consecutive tuples: their n*PACK_ROWS characters are staged into
// shared memory with coalesced 16-byte loads, then thread t encodes tuple t from its row.
// The cut / pad rule (makeStringSize_, src/fastq.cc:57-66) only decides `filtered`: a segment
// shorter than its share would be padded with 'N', so the tuple is dropped and its characters
// no longer matter.
// WIDE: the word is hi << 64 | lo (its last 32 letters in lo); a filtered tuple is all-ones in both.
template <bool WIDE, int ROWS>
__global__ void __launch_bounds__(PACK_THREADS)
    pack_words_kernel(const uint8_t *__restrict__ chars, const uint16_t *__restrict__ seg_len, uint64_t n_tuples,
                      PackParams pp, uint64_t *__restrict__ keys_out, uint64_t *__restrict__ keys_hi_out,
                      unsigned long long *__restrict__ usable) {
  extern __shared__ __align__(16) uint8_t tile[];
  const uint64_t t0 = (uint64_t)blockIdx.x * ROWS;
  const uint32_t rows = (uint32_t)min((uint64_t)ROWS, n_tuples - t0);
  const uint32_t bytes = rows * pp.n;
  const uint8_t *src = chars + t0 * pp.n;
  if ((((uintptr_t)src) & 15u) == 0) {
    // PACK_LOADS 16-byte loads in flight per thread whatever the tile holds (a plain unrolled loop leaves the
    // last vecs % (PACK_LOADS * PACK_THREADS) to a one-load-at-a-time remainder: 2x the time when that is all of them)
    const uint32_t vecs = bytes >> 4;
    for (uint32_t base = 0; base < vecs; base += PACK_LOADS * PACK_THREADS) {
      uint4 v[PACK_LOADS];
#pragma unroll
      for (int k = 0; k < PACK_LOADS; k++) {
        const uint32_t i = base + k * PACK_THREADS + threadIdx.x;
        if (i < vecs) v[k] = ld_stream_u128(src + 16ull * i);
      }
#pragma unroll
      for (int k = 0; k < PACK_LOADS; k++) {
        const uint32_t i = base + k * PACK_THREADS + threadIdx.x;
        if (i < vecs) reinterpret_cast<uint4 *>(tile)[i] = v[k];
      }
    }
    for (uint32_t i = (vecs << 4) + threadIdx.x; i < bytes; i += PACK_THREADS) tile[i] = src[i];
  } else {
    for (uint32_t i = threadIdx.x; i < bytes; i += PACK_THREADS) tile[i] = src[i];
  }
  __syncthreads();

  uint32_t n_good = 0;
  for (uint32_t rr = threadIdx.x; rr < rows; rr += PACK_THREADS)
    n_good += pack_row<WIDE>(tile + rr * pp.n, seg_len, t0 + rr, pp, keys_out, keys_hi_out);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) n_good += __shfl_xor_sync(0xffffffffu, n_good, o);
  if (lane_id() == 0 && n_good) atomicAdd(usable, (unsigned long long)n_good);
}

// ---- the same with the tile brought in by the bulk-copy engine (sm_90+: cp.async.bulk, UBLKCP in SASS) -----------
// One elected thread asks for a whole tile (ROWS x n bytes, one contiguous piece of the input) with ONE instruction;
// the bytes land in shared memory without passing through registers and complete an mbarrier's transaction count.
// Persistent CTAs, two stages: the next tile of a CTA is in flight while the current one is encoded.
__device__ __forceinline__ uint32_t smem_addr(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
  uint32_t done;
  do {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t"
        "}"
        : "=r"(done)
        : "r"(smem_addr(bar)), "r"(parity)
        : "memory");
  } while (!done);
}
// bytes: a multiple of 16; both addresses 16-byte aligned
__device__ __forceinline__ void bulk_load(void *dst_smem, const void *src_gmem, uint32_t bytes, uint64_t *bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_addr(dst_smem)),
               "l"(src_gmem), "r"(bytes), "r"(smem_addr(bar))
               : "memory");
}

template <bool WIDE, int ROWS>
__global__ void __launch_bounds__(PACK_THREADS)
    pack_words_bulk_kernel(const uint8_t *__restrict__ chars, const uint16_t *__restrict__ seg_len, uint64_t n_tuples,
                           uint32_t n_tiles, uint32_t stage_bytes, PackParams pp, uint64_t *__restrict__ keys_out,
                           uint64_t *__restrict__ keys_hi_out, unsigned long long *__restrict__ usable) {
  extern __shared__ __align__(128) uint8_t stages[];  // 2 x stage_bytes
  __shared__ __align__(8) uint64_t full[2];
  if (threadIdx.x == 0) {
    mbar_init(&full[0], 1);
    mbar_init(&full[1], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  // tile `tl` -> stage `st`: the part that is a multiple of 16 bytes by the copy engine, a last tile's odd tail by hand
  auto issue = [&](uint32_t tl, uint32_t st) {
    const uint64_t t0 = (uint64_t)tl * ROWS;
    const uint32_t bytes = (uint32_t)min((uint64_t)ROWS, n_tuples - t0) * pp.n;
    const uint32_t bulk = bytes & ~15u;
    uint8_t *dst = stages + (size_t)st * stage_bytes;
    const uint8_t *src = chars + t0 * pp.n;
    if (bulk) {
      mbar_expect_tx(&full[st], bulk);
      bulk_load(dst, src, bulk, &full[st]);
    } else {
      asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_addr(&full[st])) : "memory");
    }
    for (uint32_t i = bulk; i < bytes; i++) dst[i] = src[i];
  };
  uint32_t tile = blockIdx.x, it = 0, n_good = 0;
  if (threadIdx.x == 0 && tile < n_tiles) issue(tile, 0);
  for (; tile < n_tiles; tile += gridDim.x, it++) {
    const uint32_t st = it & 1u, next = tile + gridDim.x;
    // stage st ^ 1 was read in the previous iteration, which ended with a __syncthreads
    if (threadIdx.x == 0 && next < n_tiles) issue(next, st ^ 1u);
    mbar_wait(&full[st], (it >> 1) & 1u);
    __syncthreads();  // the hand-copied tail of a last tile (written by thread 0) is visible to all
    const uint64_t t0 = (uint64_t)tile * ROWS;
    const uint32_t rows = (uint32_t)min((uint64_t)ROWS, n_tuples - t0);
    const uint8_t *src = stages + (size_t)st * stage_bytes;
    for (uint32_t rr = threadIdx.x; rr < rows; rr += PACK_THREADS)
      n_good += pack_row<WIDE>(src + rr * pp.n, seg_len, t0 + rr, pp, keys_out, keys_hi_out);
    __syncthreads();
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) n_good += __shfl_xor_sync(0xffffffffu, n_good, o);
  if (lane_id() == 0 && n_good) atomicAdd(usable, (unsigned long long)n_good);
}

template <bool WIDE, int ROWS>
static void launch_pack_bulk(State &s, const uint8_t *d_chars, const uint16_t *d_seg_len, uint64_t first, uint64_t n,
                             const PackParams &pp) {
  const uint32_t stage_bytes = ((uint32_t)ROWS * pp.n + 127u) & ~127u;
  const size_t smem = 2 * (size_t)stage_bytes;
  static int per_sm = 0;
  auto *fn = pack_words_bulk_kernel<WIDE, ROWS>;
  if (!per_sm) {
    HUMID_CUDA(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * (((int)ROWS * 64 + 127) & ~127)));
    per_sm = 3;
  }
  const uint32_t n_tiles = (uint32_t)((n + ROWS - 1) / ROWS);
  const unsigned grid = std::min<unsigned>(n_tiles, SM_COUNT_B200 * per_sm);
  fn<<<grid, PACK_THREADS, smem, s.stream>>>(d_chars, d_seg_len, n, n_tiles, stage_bytes, pp, s.tuple_key.p + first,
                                             WIDE ? s.tuple_hi.p + first : nullptr, s.counters.p + 0);
}

// HUMID_GPU_PACK=bulk|plain picks the tile movement of the pack kernel.  Measured on B200, 50 M tuples of 24 letters
// (profiles/r02_sass_grep.md): register-staged 16-byte loads 0.395 ms, bulk copy + mbarrier, two stages, 3 persistent
// CTAs per SM 0.422 ms — the plain loads stay the default, the bulk path stays tested.
static bool pack_by_bulk_copy() {
  const char *e = std::getenv("HUMID_GPU_PACK");  // read per batch: the tests switch it inside one process
  return e ? std::string(e) == "bulk" : HUMID_PACK_DEFAULT_BULK != 0;
}

int stage_pack(State &s, const uint8_t *d_chars, const uint16_t *d_seg_len, uint64_t first, uint64_t n) {
  if (n == 0) return 0;
  PackParams pp{};
  pp.n = s.cfg.word_length;
  pp.n_segs = s.cfg.n_segments;
  for (uint32_t i = 0; i < pp.n_segs; i++) pp.take[i] = s.cfg.seg_take[i];
  // makeWord: n ASCII bytes in, one 8-byte key (two for a word of more than 31 letters) out
  Timed t(s, "pack", n, n * (pp.n + (s.wide() ? 16 : 8)));
  // the copy engine wants 16-byte aligned tiles: true for every tile when the input is and a tile's size is a multiple
  if (pack_by_bulk_copy() && (((uintptr_t)d_chars) & 15u) == 0 &&
      (((uint64_t)(s.wide() ? PACK_ROWS_WIDE : PACK_ROWS) * pp.n) & 15u) == 0) {
    if (s.wide()) launch_pack_bulk<true, PACK_ROWS_WIDE>(s, d_chars, d_seg_len, first, n, pp);
    else launch_pack_bulk<false, PACK_ROWS>(s, d_chars, d_seg_len, first, n, pp);
    HUMID_CUDA(cudaGetLastError());
    return 1;
  }
  if (s.wide())
    pack_words_kernel<true, PACK_ROWS_WIDE><<<grid_for(n, PACK_ROWS_WIDE), PACK_THREADS, (size_t)PACK_ROWS_WIDE * pp.n,
                                              s.stream>>>(d_chars, d_seg_len, n, pp, s.tuple_key.p + first,
                                                          s.tuple_hi.p + first, s.counters.p + 0);
  else
    pack_words_kernel<false, PACK_ROWS><<<grid_for(n, PACK_ROWS), PACK_THREADS, (size_t)PACK_ROWS * pp.n, s.stream>>>(
        d_chars, d_seg_len, n, pp, s.tuple_key.p + first, nullptr, s.counters.p + 0);
  HUMID_CUDA(cudaGetLastError());
  return 1;
}

// ---- the 4-bit wire format (humid_gpu_add_batch_packed4) ---------------------------------------------------------
// Half the bytes of the ASCII form on the host -> device link: one nibble per character, 0..3 = A C G T
// (src/fastq.cc:12), anything else = not a nucleotide; two characters per byte, the earlier one in the HIGH nibble;
// (word_length + 1) / 2 bytes per tuple.  The cut / pad rule and the filter still run here (makeStringSize_,
// makeWord): a short segment or a nibble above 3 filters the tuple.
// Eight characters at a time: one little-endian u32 = bytes 0..3 = characters 0..7.
__device__ __forceinline__ uint32_t pack8_nibbles(uint32_t x, bool &bad) {
  x = __byte_perm(x, 0u, 0x0123);                             // character 0 into bits 31..28
  bad |= (x & 0xCCCCCCCCu) != 0u;                             // a code above 3
  x = ((x >> 2) & 0x0C0C0C0Cu) | (x & 0x03030303u);           // per byte: two characters in bits 3..0
  x = ((x >> 4) & 0x00F000F0u) | (x & 0x000F000Fu);           // per half: four characters in bits 7..0
  return ((x >> 8) & 0xFF00u) | (x & 0xFFu);                  // eight characters, the first most significant
}

// One thread per tuple: the 12 bytes of a 24-letter word are three u32 loads, a warp reads 384 consecutive bytes.
template <bool WIDE>
__global__ void __launch_bounds__(256)
    pack_words4_kernel(const uint8_t *__restrict__ nibbles, const uint16_t *__restrict__ seg_len, uint64_t n_tuples,
                       PackParams pp, uint64_t *__restrict__ keys_out, uint64_t *__restrict__ keys_hi_out,
                       unsigned long long *__restrict__ usable) {
  const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint32_t row_bytes = (pp.n + 1u) >> 1;
  bool bad = false;
  uint64_t key = 0, key_hi = 0;
  if (t < n_tuples) {
    if (seg_len) {
      const uint16_t *sl = seg_len + t * pp.n_segs;
      for (uint32_t s = 0; s < pp.n_segs; s++) bad |= sl[s] < pp.take[s];
    }
    const uint8_t *row = nibbles + t * row_bytes;
    uint32_t q = 0;
    if ((row_bytes & 3u) == 0u && (((uintptr_t)nibbles) & 3u) == 0u) {  // rows are 4-byte aligned
      const uint32_t *row4 = reinterpret_cast<const uint32_t *>(row);
      for (; q + 8 <= pp.n; q += 8) {
        const uint32_t v = pack8_nibbles(__ldg(row4 + (q >> 3)), bad);
        if (WIDE) key_hi = (key_hi << 16) | (key >> 48);
        key = (key << 16) | v;
      }
    }
    for (; q < pp.n; q++) {
      const uint32_t b = row[q >> 1];
      const uint32_t v = (q & 1u) ? (b & 15u) : (b >> 4);
      bad |= v > 3u;
      if (WIDE) key_hi = (key_hi << 2) | (key >> 62);
      key = (key << 2) | (v & 3u);
    }
    keys_out[t] = bad ? KEY_INVALID : key;
    if (WIDE) keys_hi_out[t] = bad ? KEY_INVALID : key_hi;
  }
  uint32_t n_good = (t < n_tuples && !bad) ? 1u : 0u;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) n_good += __shfl_xor_sync(0xffffffffu, n_good, o);
  if (lane_id() == 0 && n_good) atomicAdd(usable, (unsigned long long)n_good);
}

int stage_pack4(State &s, const uint8_t *d_nibbles, const uint16_t *d_seg_len, uint64_t first, uint64_t n) {
  if (n == 0) return 0;
  PackParams pp{};
  pp.n = s.cfg.word_length;
  pp.n_segs = s.cfg.n_segments;
  for (uint32_t i = 0; i < pp.n_segs; i++) pp.take[i] = s.cfg.seg_take[i];
  // makeWord from the 4-bit form: (n + 1) / 2 bytes in, one 8-byte key (two for a word of more than 31 letters) out
  Timed t(s, "pack", n, n * ((pp.n + 1) / 2 + (s.wide() ? 16 : 8)));
  if (s.wide())
    pack_words4_kernel<true><<<grid_for(n, 256), 256, 0, s.stream>>>(d_nibbles, d_seg_len, n, pp, s.tuple_key.p + first,
                                                                      s.tuple_hi.p + first, s.counters.p + 0);
  else
    pack_words4_kernel<false><<<grid_for(n, 256), 256, 0, s.stream>>>(d_nibbles, d_seg_len, n, pp, s.tuple_key.p + first,
                                                                       nullptr, s.counters.p + 0);
  HUMID_CUDA(cudaGetLastError());
  return 1;
}

// ------------------------------------------------------------------------------------
// Distinct-word table: open addressing, linear probing, 16-B slots (two per 32-B sector).
// ------------------------------------------------------------------------------------
__global__ void table_clear_kernel(Slot *slots, uint64_t cap) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < cap) {
    uint4 v;
    v.x = v.y = 0xFFFFFFFFu;  // key = KEY_INVALID
    v.z = 0u;                 // count
    v.w = 0xFFFFFFFFu;        // first
    reinterpret_cast<uint4 *>(slots)[i] = v;
  }
}

// Whole-slot compare-and-swap (ATOMG.CAS.128): returns the slot as it was.
__device__ __forceinline__ void slot_cas(Slot *p, uint64_t cmp_lo, uint64_t cmp_hi, uint64_t new_lo, uint64_t new_hi,
                                         uint64_t &old_lo, uint64_t &old_hi) {
  asm volatile(
      "{\n\t"
      ".reg .b128 c, n, o;\n\t"
      "mov.b128 c, {%3, %4};\n\t"
      "mov.b128 n, {%5, %6};\n\t"
      "atom.global.relaxed.gpu.cas.b128 o, [%2], c, n;\n\t"
      "mov.b128 {%0, %1}, o;\n\t"
      "}"
      : "=l"(old_lo), "=l"(old_hi)
      : "l"(p), "l"(cmp_lo), "l"(cmp_hi), "l"(new_lo), "l"(new_hi)
      : "memory");
}

// Trie::add of one tuple per thread.  The whole-slot CAS against the empty pattern either claims the slot (key, count 1,
// first = this tuple: one L2 operation for a new word, where reading the slot, claiming the key and two more atomics
// for count and first took four: 1.17 -> 0.93 ms at 50 M tuples) or returns what the slot holds.
#ifndef HUMID_INSERT_THREADS
#define HUMID_INSERT_THREADS 128  // 0.887 ms at 50 M tuples; 256: 0.934, 512: 1.044, 64: 0.889
#endif
__global__ void __launch_bounds__(HUMID_INSERT_THREADS)
    table_insert_kernel(const uint64_t *__restrict__ keys, const uint32_t *__restrict__ tuple_index,
                        uint64_t n_tuples, Slot *slots, uint64_t mask, uint32_t index_base) {
  uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n_tuples) return;
  uint64_t key = ld_stream_u64(keys + t);
  if (key == KEY_INVALID) return;
  // position of the tuple in the job's input order (shard mode: tuples arrive from other ranks)
  const uint32_t ti = tuple_index ? tuple_index[t] : index_base + (uint32_t)t;
  uint64_t idx = mix64(key) & mask;
  constexpr uint64_t EMPTY_HI = 0xFFFFFFFF00000000ull;  // count 0, first 0xFFFFFFFF (table_clear_kernel)
  const uint64_t fresh_hi = ((uint64_t)ti << 32) | 1ull;  // count 1, first = this tuple
  uint64_t cur, hi;
  while (true) {
    slot_cas(slots + idx, KEY_INVALID, EMPTY_HI, key, fresh_hi, cur, hi);
    if (cur == KEY_INVALID && hi == EMPTY_HI) return;  // the slot was empty: it is this word's now, and counted
    if (cur == key) break;
    idx = (idx + 1) & mask;
  }
  atomicAdd(&slots[idx].count, 1u);  // Leaf::count++
  // first tuple in input order; `first` only ever decreases, so a stale larger value in `hi`
  // merely costs the atomic, a smaller one proves it is not needed
  if (ti < (uint32_t)(hi >> 32)) atomicMin(&slots[idx].first, ti);
}

// Occupied slots -> dense columns in any order (they are sorted next): the key's 32 leading bits (what is sorted; the
// whole key for the 64-bit sort, keys_out != nullptr), the whole 16-byte slot {key, count, first} (what the gather
// behind the sort reads: ONE random access per word gives its key and its columns), slot number.  The whole
// 16-byte slot is read (the key alone pulls the same sectors in), so count and first come along for free and the
// rank-ordered columns are later gathered from these dense arrays instead of from the table (which is 3-5x larger and
// mostly empty: 2.6 GB of DRAM fills for 0.5 GB of useful bytes in round 1's gather).  One atomic per CTA:
// warp ballots -> per-warp counts -> CTA prefix in shared memory.
constexpr int COMPACT_ITEMS = 8;
__global__ void __launch_bounds__(256)
    table_compact_kernel(const Slot *__restrict__ slots, uint64_t cap, uint64_t *__restrict__ keys_out,
                         uint32_t *__restrict__ prefix_out, int low_bits, uint4 *__restrict__ rec_out,
                         uint32_t *__restrict__ slot_out, unsigned long long *__restrict__ counter) {
  __shared__ uint32_t warp_cnt[COMPACT_ITEMS][8];
  __shared__ unsigned long long cta_base;
  const uint64_t base = (uint64_t)blockIdx.x * (256 * COMPACT_ITEMS);
  const uint32_t lane = lane_id(), w = threadIdx.x >> 5;
  uint4 slot[COMPACT_ITEMS];
  unsigned bal[COMPACT_ITEMS];
#pragma unroll
  for (int j = 0; j < COMPACT_ITEMS; j++) {
    uint64_t i = base + (uint64_t)j * 256 + threadIdx.x;
    slot[j] = i < cap ? ld_stream_u128(slots + i) : make_uint4(0xFFFFFFFFu, 0xFFFFFFFFu, 0u, 0u);
    bal[j] = __ballot_sync(0xffffffffu, (slot[j].x & slot[j].y) != 0xFFFFFFFFu);
    if (lane == 0) warp_cnt[j][w] = __popc(bal[j]);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t run = 0;
    for (int j = 0; j < COMPACT_ITEMS; j++)
      for (int k = 0; k < 8; k++) {
        uint32_t c = warp_cnt[j][k];
        warp_cnt[j][k] = run;
        run += c;
      }
    cta_base = run ? atomicAdd(counter, (unsigned long long)run) : 0ull;
  }
  __syncthreads();
#pragma unroll
  for (int j = 0; j < COMPACT_ITEMS; j++) {
    if ((slot[j].x & slot[j].y) != 0xFFFFFFFFu) {  // a key never has bit 63 set, so only KEY_INVALID is all ones
      uint64_t pos = cta_base + warp_cnt[j][w] + __popc(bal[j] & ((1u << lane) - 1u));
      const uint64_t key = ((uint64_t)slot[j].y << 32) | slot[j].x;
      if (keys_out) keys_out[pos] = key;
      else prefix_out[pos] = (uint32_t)(key >> low_bits);
      rec_out[pos] = slot[j];
      slot_out[pos] = (uint32_t)(base + (uint64_t)j * 256 + threadIdx.x);
    }
  }
}

// Rank-ordered columns from the compaction's dense ones: word_index[r] = position of rank r's word in compaction
// order (the value the sort carried).  One random 8-byte read per word out of a dense array (two random reads — the
// slot number as well — were slower than round 1's single 16-byte read out of the table: past L2 a random gather costs
// per access, not per byte).  The slot numbers are only needed for key -> rank look-ups and translated on demand.
__global__ void __launch_bounds__(256)
    gather_columns_kernel(const uint4 *__restrict__ compact_rec, const uint32_t *__restrict__ word_index, uint64_t U,
                          uint32_t *__restrict__ word_count, uint32_t *__restrict__ word_first) {
  uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= U) return;
  const uint4 rec = __ldg(compact_rec + word_index[r]);
  word_count[r] = rec.z;
  word_first[r] = rec.w;
}

// ---- sort_words by 32 leading key bits + this gather (the default; HUMID_GPU_SORT=full takes the 64-bit sort above) ----
// The sort (radix_sort_u32_indexed) ordered the words by the first 32 of their key bits only: 8 bytes per element and
// pass instead of 12, four passes instead of six for 24-letter words.  Words whose 32 leading bits agree (16 leading
// letters: rare in a library, common in none) stand next to each other in compaction order.  This kernel fetches
// every word's slot record — one random 16-byte access per word gives the key and the rank columns; the key's low bits
// and {count, first} from two smaller arrays were two accesses and slower (0.73 against 0.51 ms, call r02ag) — and
// finishes the order:
//   alone with its prefix        write key, count, first
//   first of a run of 2..64      the thread fetches the run, orders it by the whole key (insertion sort in place) and
//                                writes it, permuting word_index alike
//   first of a longer run        noted in long_runs for sort_long_runs_kernel
constexpr uint32_t RUN_SERIAL_MAX = 64;
__device__ __forceinline__ uint64_t rec_key(const uint4 &r) { return ((uint64_t)r.y << 32) | r.x; }

__global__ void __launch_bounds__(256)
    gather_sorted_words_kernel(const uint32_t *__restrict__ prefix, uint32_t *word_index, const uint4 *__restrict__ compact_rec,
                               uint64_t U, uint64_t *__restrict__ word_key, uint32_t *__restrict__ word_count,
                               uint32_t *__restrict__ word_first, uint32_t *__restrict__ long_runs,
                               unsigned long long *__restrict__ n_long) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= U) return;
  const uint32_t p = prefix[i];
  const bool after_same = i > 0 && prefix[i - 1] == p, before_same = i + 1 < U && prefix[i + 1] == p;
  if (after_same) return;  // the run's first thread writes this word
  if (!before_same) {
    const uint4 rec = __ldg(compact_rec + word_index[i]);
    word_key[i] = rec_key(rec);
    word_count[i] = rec.z;
    word_first[i] = rec.w;
    return;
  }
  uint32_t len = 2;
  while (i + len < U && len <= RUN_SERIAL_MAX && prefix[i + len] == p) len++;
  if (len > RUN_SERIAL_MAX) {
    long_runs[atomicAdd(n_long, 1ull)] = (uint32_t)i;
    return;
  }
  for (uint32_t k = 0; k < len; k++) {  // (stable insertion by the whole key; the keys of a table are distinct)
    const uint32_t id = word_index[i + k];
    const uint4 rec = __ldg(compact_rec + id);
    const uint64_t key = rec_key(rec);
    uint32_t j = k;
    while (j > 0 && word_key[i + j - 1] > key) {
      word_key[i + j] = word_key[i + j - 1];
      word_count[i + j] = word_count[i + j - 1];
      word_first[i + j] = word_first[i + j - 1];
      word_index[i + j] = word_index[i + j - 1];
      j--;
    }
    word_key[i + j] = key;
    word_count[i + j] = rec.z;
    word_first[i + j] = rec.w;
    word_index[i + j] = id;
  }
}

// Runs of more than RUN_SERIAL_MAX words with one prefix, one CTA at a time each (persistent CTAs over the list).
//   up to LONG_RUN_SMEM words   (key, index) pairs ordered by a bitonic network in shared memory
//   more                        stable counting sort on 4-bit digits of the key bits BELOW the prefix (the others agree),
//                               thread t owning a contiguous piece of the run; scratch = the same positions of the
//                               alternate buffers (key_alt, index_alt)
constexpr uint32_t LONG_RUN_SMEM = 2048;
__global__ void __launch_bounds__(256)
    sort_long_runs_kernel(const uint32_t *__restrict__ prefix, uint32_t *word_index, uint32_t *index_alt,
                          const uint4 *__restrict__ compact_rec, uint64_t U, int low_bits, uint64_t *word_key, uint64_t *key_alt,
                          uint32_t *__restrict__ word_count, uint32_t *__restrict__ word_first,
                          const uint32_t *__restrict__ long_runs, const unsigned long long *__restrict__ n_long) {
  __shared__ uint64_t skey[LONG_RUN_SMEM];
  __shared__ uint32_t sidx[LONG_RUN_SMEM];
  __shared__ uint32_t counts[16 * 256];
  __shared__ uint32_t run_len;
  const unsigned long long runs = *n_long;
  for (unsigned long long r = blockIdx.x; r < runs; r += gridDim.x) {
    const uint64_t first = long_runs[r];
    const uint32_t p = prefix[first];
    if (threadIdx.x == 0) run_len = 0xFFFFFFFFu;
    __syncthreads();
    // the run's length: the first position behind `first` with another prefix (strided search, smallest hit wins)
    for (uint64_t base = first;; base += blockDim.x) {
      const uint64_t at = base + threadIdx.x;
      if (at >= U || prefix[at] != p) atomicMin(&run_len, (uint32_t)((at < U ? at : U) - first));
      __syncthreads();
      const uint32_t seen = run_len;
      __syncthreads();
      if (seen != 0xFFFFFFFFu) break;
    }
    const uint32_t len = run_len;
    if (len <= LONG_RUN_SMEM) {
      uint32_t p2 = 1;
      while (p2 < len) p2 <<= 1;
      for (uint32_t k = threadIdx.x; k < p2; k += blockDim.x) {
        const uint32_t id = k < len ? word_index[first + k] : 0u;
        skey[k] = k < len ? rec_key(__ldg(compact_rec + id)) : ~0ull;
        sidx[k] = id;
      }
      __syncthreads();
      for (uint32_t size = 2; size <= p2; size <<= 1)
        for (uint32_t stride = size >> 1; stride > 0; stride >>= 1) {
          for (uint32_t k = threadIdx.x; k < p2; k += blockDim.x) {
            const uint32_t l = k ^ stride;
            if (l > k) {
              const uint64_t a = skey[k], b = skey[l];
              if ((a > b) == ((k & size) == 0)) {
                skey[k] = b;
                skey[l] = a;
                const uint32_t t = sidx[k];
                sidx[k] = sidx[l];
                sidx[l] = t;
              }
            }
          }
          __syncthreads();
        }
      for (uint32_t k = threadIdx.x; k < len; k += blockDim.x) {
        const uint4 rec = __ldg(compact_rec + sidx[k]);
        word_key[first + k] = skey[k];
        word_count[first + k] = rec.z;
        word_first[first + k] = rec.w;
        word_index[first + k] = sidx[k];
      }
    } else {
      uint64_t *ka = word_key + first, *kb = key_alt + first;
      uint32_t *ia = word_index + first, *ib = index_alt + first;
      for (uint32_t k = threadIdx.x; k < len; k += blockDim.x) ka[k] = rec_key(__ldg(compact_rec + ia[k]));
      __syncthreads();
      const uint32_t piece = (len + blockDim.x - 1) / blockDim.x;
      const uint32_t lo = threadIdx.x * piece < len ? threadIdx.x * piece : len, hi = lo + piece < len ? lo + piece : len;
      for (int shift = 0; shift < low_bits; shift += 4) {
        for (uint32_t d = 0; d < 16; d++) counts[d * 256 + threadIdx.x] = 0;
        for (uint32_t k = lo; k < hi; k++) counts[((uint32_t)(ka[k] >> shift) & 15u) * 256 + threadIdx.x]++;
        __syncthreads();
        if (threadIdx.x == 0) {  // digit-major, thread-minor: 4096 running sums
          uint32_t run = 0;
          for (uint32_t c = 0; c < 16 * 256; c++) {
            const uint32_t v = counts[c];
            counts[c] = run;
            run += v;
          }
        }
        __syncthreads();
        for (uint32_t k = lo; k < hi; k++) {
          const uint32_t at = counts[((uint32_t)(ka[k] >> shift) & 15u) * 256 + threadIdx.x]++;
          kb[at] = ka[k];
          ib[at] = ia[k];
        }
        __syncthreads();
        uint64_t *tk = ka;
        ka = kb;
        kb = tk;
        uint32_t *ti = ia;
        ia = ib;
        ib = ti;
      }
      // (an odd number of passes leaves the result in the alternate buffers)
      for (uint32_t k = threadIdx.x; k < len; k += blockDim.x) {
        const uint64_t key = ka[k];
        const uint32_t id = ia[k];
        const uint4 rec = __ldg(compact_rec + id);
        word_key[first + k] = key;
        word_index[first + k] = id;
        word_count[first + k] = rec.z;
        word_first[first + k] = rec.w;
      }
    }
    __syncthreads();
  }
}

__global__ void __launch_bounds__(256)
    index_to_slot_kernel(const uint32_t *__restrict__ compact_slot, uint32_t *word_index_then_slot, uint64_t U) {
  uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r < U) word_index_then_slot[r] = __ldg(compact_slot + word_index_then_slot[r]);
}

__global__ void __launch_bounds__(256)
    scatter_ranks_kernel(Slot *slots, const uint32_t *__restrict__ word_slot, uint64_t U) {
  uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r < U) slots[word_slot[r]].count = (uint32_t)r;
}

// Fingerprint filter over the distinct words (layout: common.cuh).  One CAS per word into an
// L2-resident array.
__global__ void __launch_bounds__(256)
    filter_build_kernel(const uint64_t *__restrict__ word_key, uint64_t U, uint32_t *words, uint32_t n_buckets) {
  uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= U) return;
  const uint64_t h = filter_hash(word_key[r]);
  uint32_t *w = words + 4ull * filter_bucket(h, n_buckets);
  const uint32_t fp = filter_fp(h);
  const uint32_t start = (uint32_t)(h >> 40) & 3u;
  for (uint32_t t = 0; t < 4; t++) {
    const uint32_t wi = (start + t) & 3u;
    uint32_t old = *reinterpret_cast<volatile uint32_t *>(w + wi);
    while (true) {
      int b = -1;
      for (int q = wi == 0 ? 1 : 0; q < 4; q++)
        if (((old >> (8 * q)) & 0xFFu) == 0u) { b = q; break; }
      if (b < 0) break;  // word full
      uint32_t prev = atomicCAS(w + wi, old, old | (fp << (8 * b)));
      if (prev == old) return;
      old = prev;
    }
  }
  atomicOr(w, 1u);  // bucket overflowed: queries falling here always go to the table
}

int stage_build_filter(State &s) {
  const uint64_t U = s.unique;
  s.filter_buckets = 0;
  if (U == 0) return 0;
  uint64_t nb = std::max<uint64_t>(1024, U / 7 + 1);  // ~7 words per 15-slot bucket: 43 MB at 21 M words
  if (nb > 0xFFFFFFF0ull) return 0;
  s.filter.ensure(nb);
  Timed t(s, "filter_build", U, U * 9 + nb * 16);
  HUMID_CUDA(cudaMemsetAsync(s.filter.p, 0, nb * sizeof(uint4), s.stream));
  filter_build_kernel<<<grid_for(U, 256), 256, 0, s.stream>>>(s.word_key.p, U, reinterpret_cast<uint32_t *>(s.filter.p),
                                                              (uint32_t)nb);
  s.filter_buckets = (uint32_t)nb;
  return 1;
}

static uint64_t table_capacity_for(uint64_t tuples) {
  uint64_t cap = 1024;
  while (4 * cap < 5 * tuples) cap <<= 1;  // load factor <= 0.8 if every tuple were distinct, ~0.3 typically
  if (cap > (1ull << 32)) throw CudaError{"word table would exceed 2^32 slots"};
  return cap;
}

static int table_clear(State &s, uint64_t cap) {
  s.table_cap = cap;
  s.slots.ensure(cap);
  Timed t(s, "table_clear", cap, cap * 16);
  table_clear_kernel<<<grid_for(cap, 256), 256, 0, s.stream>>>(s.slots.p, cap);
  return 1;
}

// Trie::add for n tuples.  Large inputs are first grouped by the 1/256 of the table their home slot
// lies in, so the random slot updates run against L2 instead of paying one DRAM line fill each.
constexpr uint64_t PARTITION_MIN_TUPLES = 1ull << 21;
static int table_insert(State &s, const uint64_t *keys, const uint32_t *index, uint64_t n, uint32_t index_base,
                        uint64_t algo_bytes) {
  int k = 0;
  if (n >= PARTITION_MIN_TUPLES && s.table_cap * sizeof(Slot) > (64ull << 20)) {
    s.part_key.ensure(n);
    s.part_index.ensure(n);
    Timed t(s, "table_partition", n, n * (8 + 8 + 12));  // histogram read, pass read, (key, index) written
    int launches = 0;
    partition_by_slot_u64(keys, index, index_base, s.part_key.p, s.part_index.p, n, s.table_cap - 1, s.scratch,
                          s.stream, &launches);
    t.launches = launches;
    k += launches;
    keys = s.part_key.p;
    index = s.part_index.p;
  }
  Timed t(s, "table_insert", n, algo_bytes);
  table_insert_kernel<<<grid_for(n, HUMID_INSERT_THREADS), HUMID_INSERT_THREADS, 0, s.stream>>>(keys, index, n, s.slots.p, s.table_cap - 1, index_base);
  return k + 1;
}

int stage_early_insert(State &s, uint64_t first, uint64_t n) {
  if (s.cfg.defer_table || s.cfg.expected_tuples == 0 || s.wide()) return 0;
  int k = 0;
  if (first == 0) {
    k += table_clear(s, table_capacity_for(s.cfg.expected_tuples));
    s.early_table = true;
    s.early_inserted = 0;
  }
  if (!s.early_table || first != s.early_inserted) return k;
  if (first + n > s.cfg.expected_tuples) {  // outgrew the hint: finalize recounts into a larger table
    s.early_table = false;
    return k;
  }
  // key in (8) + slot key (8) + count RMW (8) + first RMW (8)
  k += table_insert(s, s.tuple_key.p + first, nullptr, n, (uint32_t)first, n * 32);
  s.early_inserted = first + n;
  return k;
}

int stage_build_table(State &s, bool with_lookup) {
  int k = 0;
  const uint64_t *keys = s.tuple_key.p;
  const uint32_t *index = nullptr;
  uint64_t n_in = s.n_tuples;
  if (s.src_keys) {  // shard mode: tuples received from all ranks, every one usable
    keys = s.src_keys;
    index = s.src_index;
    n_in = s.src_n;
    s.usable = n_in;
  } else {
    HUMID_CUDA(cudaMemcpyAsync(&s.usable, s.counters.p + 0, sizeof(uint64_t), cudaMemcpyDeviceToHost, s.stream));
    HUMID_CUDA(cudaStreamSynchronize(s.stream));
  }
  const bool counted = !s.src_keys && s.early_table && s.early_inserted == s.n_tuples && s.n_tuples > 0;
  if (!counted) {
    k += table_clear(s, table_capacity_for(s.usable));
    if (n_in) k += table_insert(s, keys, index, n_in, 0u, n_in * 8 + s.usable * 24);
  }
  const uint64_t cap = s.table_cap;
  // distinct words cannot outnumber usable tuples
  s.word_key.ensure(s.usable + 1);
  s.key_alt.ensure(s.usable + 1);
  s.word_slot.ensure(s.usable + 1);
  s.slot_alt.ensure(s.usable + 1);
  // HUMID_GPU_SORT=full: the 64-bit sort + plain gather (round 1's path, kept for comparisons)
  const char *sort_env = std::getenv("HUMID_GPU_SORT");
  const bool by_prefix = !(sort_env && std::string(sort_env) == "full");
  const int key_bits = 2 * (int)s.cfg.word_length;
  const int prefix_bits = std::min(32, key_bits), low_bits = key_bits - prefix_bits;
  if (by_prefix) s.compact_prefix.ensure(s.usable + 1);
  else s.compact_key.ensure(s.usable + 1);
  s.compact_rec.ensure(s.usable + 1);
  s.compact_slot.ensure(s.usable + 1);
  HUMID_CUDA(cudaMemsetAsync(s.counters.p + 1, 0, sizeof(unsigned long long), s.stream));
  {
    Timed t(s, "table_compact", cap, cap * 16 + s.usable * 20);
    table_compact_kernel<<<grid_for(cap, 256, COMPACT_ITEMS), 256, 0, s.stream>>>(s.slots.p, cap, by_prefix ? nullptr : s.compact_key.p,
                                                                                 s.compact_prefix.p, low_bits, s.compact_rec.p,
                                                                                 s.compact_slot.p,
                                                                                 s.counters.p + 1);
  }
  k++;
  HUMID_CUDA(cudaMemcpyAsync(&s.unique, s.counters.p + 1, sizeof(uint64_t), cudaMemcpyDeviceToHost, s.stream));
  HUMID_CUDA(cudaStreamSynchronize(s.stream));
  const uint64_t U = s.unique;
  s.word_count.ensure(U + 1);
  s.word_first.ensure(U + 1);
  if (U && by_prefix) {
    // the 32 leading key bits order nearly all words; the gather behind the sort finishes the rest (see there)
    s.prefix_a.ensure(s.usable + 1);
    s.prefix_b.ensure(s.usable + 1);
    uint32_t *pk[2] = {s.prefix_a.p, s.prefix_b.p};
    uint32_t *vb[2] = {s.word_slot.p, s.slot_alt.p};
    int launches = 0;
    int cur;
    {
      // 8 bytes per element: histograms read the prefixes once, the first pass reads them and writes (prefix, index), the
      // others read and write both
      const uint64_t passes = ((uint64_t)prefix_bits + 7) / 8;
      Timed t(s, "sort_words", U, U * (4 + 4 + 8 + (passes - 1) * 16));
      cur = radix_sort_u32_indexed(s.compact_prefix.p, pk, vb, U, prefix_bits, s.scratch, s.stream, &launches);
      t.launches = launches;
    }
    k += launches;
    if (cur == 1) std::swap(s.word_slot, s.slot_alt);
    uint32_t *long_runs = pk[cur ^ 1];  // (the sort's other prefix buffer is free now)
    unsigned long long *n_long = s.counters.p + 14;
    HUMID_CUDA(cudaMemsetAsync(n_long, 0, sizeof(unsigned long long), s.stream));
    Timed t(s, "rank_columns", U, U * (4 + 4 + 16 + 8 + 4 + 4));
    gather_sorted_words_kernel<<<grid_for(U, 256), 256, 0, s.stream>>>(pk[cur], s.word_slot.p, s.compact_rec.p, U, s.word_key.p,
                                                                       s.word_count.p, s.word_first.p, long_runs, n_long);
    k++;
    if (low_bits) {  // (no low bits: the prefix is the whole key and no two words share it)
      sort_long_runs_kernel<<<2 * 148, 256, 0, s.stream>>>(pk[cur], s.word_slot.p, s.slot_alt.p, s.compact_rec.p, U, low_bits,
                                                           s.word_key.p, s.key_alt.p, s.word_count.p, s.word_first.p, long_runs,
                                                           n_long);
      k++;
    }
    t.launches = low_bits ? 2 : 1;
  } else if (U) {
    uint64_t *kb[2] = {s.word_key.p, s.key_alt.p};
    uint32_t *vb[2] = {s.word_slot.p, s.slot_alt.p};
    int launches = 0;
    int cur;
    {
      // per 8-bit pass the (key, index) pairs are read once and written once (the first pass reads keys only: the index
      // of an element is its position); one more read of the keys for the histograms
      const uint64_t passes = (2 * (uint64_t)s.cfg.word_length + 7) / 8;
      Timed t(s, "sort_words", U, U * (8 + passes * 2 * (8 + 4) - 4));
      cur = radix_sort_bits_indexed(s.compact_key.p, kb, vb, U, 0, 2 * (int)s.cfg.word_length, s.scratch, s.stream, &launches);
      t.launches = launches;
    }
    k += launches;
    if (cur == 1) {
      std::swap(s.word_key, s.key_alt);
      std::swap(s.word_slot, s.slot_alt);
    }
    Timed t(s, "rank_columns", U, U * (4 + 8 + 4 + 4));
    gather_columns_kernel<<<grid_for(U, 256), 256, 0, s.stream>>>(s.compact_rec.p, s.word_slot.p, U, s.word_count.p,
                                                                  s.word_first.p);
    k++;
  }
  s.word_slot_is_index = true;  // translated to slot numbers when look-ups are set up (stage_ensure_lookup)
  s.ranks_in_table = false;
  s.filter_buckets = 0;
  if (with_lookup) k += stage_ensure_lookup(s);
  HUMID_CUDA(cudaGetLastError());
  return k;
}

// key -> rank look-ups (Lookup in common.cuh) need the ranks in the table slots and the filter in
// front of them.  The Hamming block join never asks, so this is done on first use.
int stage_ensure_lookup(State &s) {
  int k = 0;
  const uint64_t U = s.unique;
  if (!s.ranks_in_table && U) {
    Timed t(s, "table_ranks", U, U * (4 + 4 + 4 + 4));
    if (s.word_slot_is_index) {
      index_to_slot_kernel<<<grid_for(U, 256), 256, 0, s.stream>>>(s.compact_slot.p, s.word_slot.p, U);
      s.word_slot_is_index = false;
      k++;
    }
    scatter_ranks_kernel<<<grid_for(U, 256), 256, 0, s.stream>>>(s.slots.p, s.word_slot.p, U);
    t.launches = 2;
    k++;
  }
  s.ranks_in_table = true;
  if (!s.filter_buckets) k += stage_build_filter(s);
  HUMID_CUDA(cudaGetLastError());
  return k;
}

}  // namespace humid
