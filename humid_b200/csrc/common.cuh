// Shared device/host helpers for libhumid_gpu (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>

namespace humid {

constexpr uint64_t KEY_INVALID = ~0ull;  // filtered tuple / empty table slot (bit 63 is never
                                         // set in a real key: word_length <= 31 => 62 bits)
constexpr uint32_t NONE32 = 0xFFFFFFFFu;
constexpr int SM_COUNT_B200 = 148;

struct CudaError {
  std::string msg;
};

#define HUMID_CUDA(expr)                                                                       \
  do {                                                                                         \
    cudaError_t _e = (expr);                                                                   \
    if (_e != cudaSuccess)                                                                     \
      throw ::humid::CudaError{std::string(#expr) + ": " + cudaGetErrorString(_e) + " (" +     \
                               __FILE__ + ":" + std::to_string(__LINE__) + ")"};               \
  } while (0)

// One open-addressing slot of the distinct-word table (16 B => two slots per 32-B sector).
// After the rank pass `count` is re-used to hold the word's rank (count/first have then been
// gathered into rank-ordered columns).
struct __align__(16) Slot {
  uint64_t key;
  uint32_t count;  // Leaf::count, later: rank
  uint32_t first;  // smallest tuple index carrying the word
};

__host__ __device__ __forceinline__ uint64_t mix64(uint64_t x) {
  // splitmix64 finaliser: full avalanche, so the low bits index the table uniformly even
  // though neighbouring words differ in a single 2-bit field.
  x ^= x >> 30;
  x *= 0xbf58476d1ce4e5b9ull;
  x ^= x >> 27;
  x *= 0x94d049bb133111ebull;
  x ^= x >> 31;
  return x;
}

// ---- key -> rank lookup: L2-resident fingerprint filter in front of the HBM slot table ------
// Neighbour search asks for ~1.5 n (d=1) to thousands (edit d=2) of variants per word and
// almost all of them do not exist.  A random probe of the 16-B slot table costs one HBM sector
// fill (measured ceiling: ~32 G random sectors/s); a probe of a 16-B filter bucket that stays
// in the 126 MB L2 runs at ~227 G/s.  Bucket = 16 bytes: byte 0 = overflow flag, bytes 1..15 =
// 8-bit fingerprints (2..255; 0 = free).  ~5 words per bucket on average => false-positive rate
// ~2 %; only filter hits (and overflowed buckets) touch the table.
struct Lookup {
  const Slot *slots;
  uint64_t mask;
  const uint4 *filter;  // may be nullptr (table only)
  uint32_t n_buckets;
  // optional first level for searches that ask for thousands of variants per word (edit distance): one bit per
  // hash value, ~3 % of them set, L2-resident — a miss costs one 4-byte load and a handful of instructions
  const uint32_t *bitmap;  // may be nullptr
  int bitmap_shift;        // bit index = (key * BITMAP_MUL) >> bitmap_shift
};
constexpr uint64_t BITMAP_MUL = 0xD6E8FEB86659FD93ull;

// The filter uses its own, cheaper hash: hf(key) = key * FILTER_MUL (Fibonacci hashing, high
// bits).  It is linear, so the hash of a variant w + delta is hf(w) + delta * FILTER_MUL: one
// multiply per word, one add per probed variant.  (The slot table keeps the avalanche hash
// mix64: its index is taken from the LOW bits.)
constexpr uint64_t FILTER_MUL = 0x9E3779B97F4A7C15ull;
__host__ __device__ __forceinline__ uint64_t filter_hash(uint64_t key) { return key * FILTER_MUL; }
__host__ __device__ __forceinline__ uint32_t filter_bucket(uint64_t hf, uint32_t n_buckets) {
#ifdef __CUDA_ARCH__
  return __umulhi((uint32_t)(hf >> 32), n_buckets);  // the 64-bit spelling below was once miscompiled (primitives.cu: DestDigit)
#else
  return (uint32_t)(((hf >> 32) * (uint64_t)n_buckets) >> 32);
#endif
}
__host__ __device__ __forceinline__ uint32_t filter_fp(uint64_t hf) {
  uint32_t b = (uint32_t)(hf >> 32) & 0xFFu;  // low byte of the high word: not used by the bucket index
  return b < 2u ? b + 2u : b;
}

#ifdef __CUDACC__
__device__ __forceinline__ bool filter_match(const uint4 &b, uint64_t hf) {
  const uint32_t f4 = filter_fp(hf) * 0x01010101u;
  const uint32_t m = __vcmpeq4(b.x, f4) | __vcmpeq4(b.y, f4) | __vcmpeq4(b.z, f4) | __vcmpeq4(b.w, f4);
  return m != 0u || (b.x & 1u) != 0u;
}

// slot table only: rank of `key` or NONE32 (after stage_ensure_lookup `count` holds the rank)
__device__ __forceinline__ uint32_t table_rank(const Lookup &lk, uint64_t key) {
  uint64_t idx = mix64(key) & lk.mask;
  while (true) {
    const uint4 v = __ldg(reinterpret_cast<const uint4 *>(lk.slots + idx));
    const uint64_t k = ((uint64_t)v.y << 32) | v.x;
    if (k == key) return v.z;
    if (k == KEY_INVALID) return NONE32;
    idx = (idx + 1) & lk.mask;
  }
}

// rank of `key`, or NONE32: filter first, table only on a fingerprint match
__device__ __forceinline__ uint32_t lookup_rank(const Lookup &lk, uint64_t key) {
  if (lk.bitmap) {
    const uint64_t bit = (key * BITMAP_MUL) >> lk.bitmap_shift;
    if (((__ldg(lk.bitmap + (bit >> 5)) >> (bit & 31u)) & 1u) == 0u) return NONE32;
  }
  if (lk.filter) {
    const uint64_t hf = filter_hash(key);
    const uint4 b = __ldg(lk.filter + filter_bucket(hf, lk.n_buckets));
    if (!filter_match(b, hf)) return NONE32;
  }
  return table_rank(lk, key);
}
#endif

__device__ __forceinline__ uint32_t lane_id() { return threadIdx.x & 31u; }

// Grid size helper: enough CTAs to fill the 148 SMs a few times over, never more than needed.
// Reserve `each` consecutive output slots per calling lane at a global cursor.  All lanes that
// reach the call together share ONE atomic: a cursor is a single address, and returning atomics on
// one address complete at well under 1 G/s, so a per-thread atomicAdd caps a kernel at that rate.
__device__ __forceinline__ unsigned long long reserve_slots(unsigned long long *cursor, uint32_t each = 1u) {
  const unsigned m = __activemask();
  const uint32_t lane = threadIdx.x & 31u;
  const int leader = __ffs(m) - 1;
  unsigned long long base = 0;
  if ((int)lane == leader) base = atomicAdd(cursor, (unsigned long long)each * __popc(m));
  return __shfl_sync(m, base, leader) + (unsigned long long)each * __popc(m & ((1u << lane) - 1u));
}

inline unsigned grid_for(uint64_t items, unsigned block, unsigned per_thread = 1) {
  uint64_t per_block = (uint64_t)block * per_thread;
  uint64_t g = (items + per_block - 1) / per_block;
  if (g == 0) g = 1;
  if (g > 0x7FFFFFFFull) g = 0x7FFFFFFFull;
  return (unsigned)g;
}

// Streaming (read-once) loads / stores that do not pollute L1.
__device__ __forceinline__ uint64_t ld_stream_u64(const uint64_t *p) {
  uint64_t v;
  asm volatile("ld.global.nc.L1::no_allocate.u64 %0, [%1];" : "=l"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ uint4 ld_stream_u128(const void *p) {
  uint4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(p));
  return v;
}

}  // namespace humid
