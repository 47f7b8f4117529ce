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

__device__ __forceinline__ uint32_t lane_id() { return threadIdx.x & 31u; }

// Grid size helper: enough CTAs to fill the 148 SMs a few times over, never more than needed.
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
