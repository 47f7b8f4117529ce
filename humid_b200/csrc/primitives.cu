// Exclusive scan + stable LSD radix sort (see primitives.cuh).
#include <cstdlib>
#include <mutex>

#include "../../include/humid_gpu.h"
#include "primitives.cuh"

namespace humid {

// ------------------------------------------------------------------------------------
// Exclusive scan: reduce per tile -> scan of tile sums (recursive) -> scan per tile.
// ------------------------------------------------------------------------------------
constexpr int SCAN_THREADS = 256;
constexpr int SCAN_ITEMS = 16;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;

__device__ __forceinline__ uint32_t warp_inclusive_scan(uint32_t v) {
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t t = __shfl_up_sync(0xffffffffu, v, o);
    if ((int)lane_id() >= o) v += t;
  }
  return v;
}

// Block-wide exclusive scan of one value per thread; returns the exclusive prefix and the
// block total (valid in all threads).
__device__ __forceinline__ uint32_t block_exclusive_scan(uint32_t v, uint32_t *total) {
  __shared__ uint32_t warp_sums[SCAN_THREADS / 32];
  __shared__ uint32_t block_total;
  uint32_t inc = warp_inclusive_scan(v);
  int w = threadIdx.x >> 5;
  if (lane_id() == 31) warp_sums[w] = inc;
  __syncthreads();
  if (w == 0) {
    uint32_t s = lane_id() < SCAN_THREADS / 32 ? warp_sums[lane_id()] : 0;
    uint32_t si = warp_inclusive_scan(s);
    if (lane_id() < SCAN_THREADS / 32) warp_sums[lane_id()] = si - s;
    if (lane_id() == SCAN_THREADS / 32 - 1) block_total = si;
  }
  __syncthreads();
  uint32_t r = inc - v + warp_sums[w];
  *total = block_total;
  __syncthreads();
  return r;
}

__global__ void __launch_bounds__(SCAN_THREADS) scan_reduce_kernel(const uint32_t *__restrict__ in,
                                                                   uint64_t n, uint32_t *__restrict__ sums) {
  uint64_t base = (uint64_t)blockIdx.x * SCAN_TILE;
  uint32_t s = 0;
#pragma unroll
  for (int j = 0; j < SCAN_ITEMS; j++) {
    uint64_t i = base + (uint64_t)j * SCAN_THREADS + threadIdx.x;
    if (i < n) s += in[i];
  }
  uint32_t total;
  block_exclusive_scan(s, &total);
  if (threadIdx.x == 0) sums[blockIdx.x] = total;
}

// Thread t owns items [t*ITEMS, (t+1)*ITEMS) of the tile (blocked), staged through shared
// memory so that global accesses stay coalesced; the skew k + k/32 makes the blocked
// (stride-16) shared-memory accesses conflict-free.
__device__ __forceinline__ uint32_t skew(uint32_t k) { return k + (k >> 5); }

__global__ void __launch_bounds__(SCAN_THREADS) scan_tile_kernel(const uint32_t *in, uint32_t *out, uint64_t n,
                                                                 const uint32_t *__restrict__ tile_offsets,
                                                                 uint32_t *total_dev) {
  __shared__ uint32_t stage[SCAN_TILE + SCAN_TILE / 32];
  uint64_t base = (uint64_t)blockIdx.x * SCAN_TILE;
#pragma unroll
  for (int j = 0; j < SCAN_ITEMS; j++) {
    uint32_t k = j * SCAN_THREADS + threadIdx.x;
    uint64_t i = base + k;
    stage[skew(k)] = i < n ? in[i] : 0u;
  }
  __syncthreads();
  uint32_t run[SCAN_ITEMS];
  uint32_t s = 0;
#pragma unroll
  for (int j = 0; j < SCAN_ITEMS; j++) {
    run[j] = stage[skew(threadIdx.x * SCAN_ITEMS + j)];
    s += run[j];
  }
  uint32_t total;
  uint32_t off = block_exclusive_scan(s, &total) + (tile_offsets ? tile_offsets[blockIdx.x] : 0u);
#pragma unroll
  for (int j = 0; j < SCAN_ITEMS; j++) {
    stage[skew(threadIdx.x * SCAN_ITEMS + j)] = off;
    off += run[j];
  }
  __syncthreads();
#pragma unroll
  for (int j = 0; j < SCAN_ITEMS; j++) {
    uint32_t k = j * SCAN_THREADS + threadIdx.x;
    uint64_t i = base + k;
    if (i < n) out[i] = stage[skew(k)];
  }
  if (total_dev && blockIdx.x == gridDim.x - 1 && threadIdx.x == 0)
    *total_dev = (tile_offsets ? tile_offsets[blockIdx.x] : 0u) + total;
}

static int scan_rec(const uint32_t *in, uint32_t *out, uint64_t n, uint32_t *total_dev, uint32_t *scratch,
                    cudaStream_t stream) {
  uint64_t tiles = (n + SCAN_TILE - 1) / SCAN_TILE;
  if (tiles <= 1) {
    scan_tile_kernel<<<1, SCAN_THREADS, 0, stream>>>(in, out, n, nullptr, total_dev);
    return 1;
  }
  uint32_t *sums = scratch;
  scan_reduce_kernel<<<(unsigned)tiles, SCAN_THREADS, 0, stream>>>(in, n, sums);
  int k = 1 + scan_rec(sums, sums, tiles, nullptr, scratch + ((tiles + 63) & ~63ull), stream);
  scan_tile_kernel<<<(unsigned)tiles, SCAN_THREADS, 0, stream>>>(in, out, n, sums, total_dev);
  return k + 1;
}

int exclusive_scan_u32(const uint32_t *in, uint32_t *out, uint64_t n, uint32_t *total_dev, Scratch &scratch,
                       cudaStream_t stream) {
  if (n == 0) {
    if (total_dev) HUMID_CUDA(cudaMemsetAsync(total_dev, 0, sizeof(uint32_t), stream));
    return 0;
  }
  size_t words = 0;
  for (uint64_t t = (n + SCAN_TILE - 1) / SCAN_TILE; t > 1; t = (t + SCAN_TILE - 1) / SCAN_TILE)
    words += (t + 63) & ~63ull;
  uint32_t *s = (uint32_t *)scratch.get((words + 64) * sizeof(uint32_t));
  int k = scan_rec(in, out, n, total_dev, s, stream);
  HUMID_CUDA(cudaGetLastError());
  return k;
}

// ------------------------------------------------------------------------------------
// Stable LSD radix sort, 8 bits per pass, one read and one write of the data per pass:
//   radix_digit_hist_kernel   ONE pass over the keys: the 256-bin histogram of every digit
//   radix_digit_bases_kernel  exclusive scan of each histogram -> first output position of a digit
//   radix_pass_kernel         per digit pass.  A tile (RS_TILE consecutive elements) ranks its
//                             elements by digit (warp ballots + per-warp counters), learns how many
//                             elements of each digit precede it from the tiles before it through a
//                             chained look-back over `status` (aggregate / inclusive words, tiles
//                             ordered by a ticket), reorders the tile in shared memory and writes
//                             every digit run as consecutive addresses.
// Warp w owns a contiguous sub-tile and walks it in strips of 32 consecutive elements, so
// "element order" == (warp, strip, lane) and the in-tile ranking is stable.
// ------------------------------------------------------------------------------------
constexpr int RS_THREADS = 256;
constexpr int RS_WARPS = RS_THREADS / 32;
// Tile shape (A/B builds override these).  Measured on sort_words, 21.3 M (key, value) pairs, 6 passes: 256 threads x 16
// elements x 3 CTAs per SM 1.09-1.12 ms; 256 x 24 x 2 1.03-1.07 (longer digit runs per tile, fewer look-backs);
// 256 x 12 x 4 1.22; 256 x 32 x 1 1.18; 512 x 16 x 1 1.08; 512 x 12 x 2 1.17; 1024 x 8 x 1 1.17; 1024 x 6 x 1 1.24.
#ifndef HUMID_RS_ITEMS
#define HUMID_RS_ITEMS 24
#endif
#ifndef HUMID_RS_CTAS
#define HUMID_RS_CTAS 2
#endif
constexpr int RS_ITEMS = HUMID_RS_ITEMS;
#ifndef HUMID_RS_THREADS
#define HUMID_RS_THREADS 256
#endif
constexpr int RS_PASS_THREADS = HUMID_RS_THREADS;  // threads of a radix_pass_kernel CTA (the histogram kernels keep RS_THREADS)
constexpr int RS_PASS_WARPS = RS_PASS_THREADS / 32;
constexpr int RS_TILE = RS_PASS_THREADS * RS_ITEMS;
constexpr int RS_BINS = 256;
constexpr int RS_MAX_PASSES = 8;
constexpr int RS_LOOK = 8;
constexpr uint64_t RS_AGGREGATE = 1ull << 62;  // status word: this tile's own count
constexpr uint64_t RS_INCLUSIVE = 1ull << 63;  // status word: count of this tile and all before it

// The pass works on 64-bit keys (every sort but one) or on 32-bit ones (radix_sort_u32_indexed).
__device__ __forceinline__ uint64_t ld_stream_key(const uint64_t *p) { return ld_stream_u64(p); }
__device__ __forceinline__ uint32_t ld_stream_key(const uint32_t *p) { return __ldcs(p); }

__device__ __forceinline__ uint32_t pass_mask(int begin_bit, int end_bit, int p) {
  const int left = end_bit - begin_bit - 8 * p;
  return left >= 8 ? 255u : (1u << left) - 1u;
}

template <class K>
__global__ void __launch_bounds__(RS_THREADS) radix_digit_hist_kernel(const K *__restrict__ keys, uint64_t n,
                                                                      int begin_bit, int end_bit, int passes,
                                                                      uint32_t *__restrict__ hist) {
  __shared__ uint32_t h[RS_MAX_PASSES][RS_BINS];
  for (int p = 0; p < passes; p++) h[p][threadIdx.x] = 0;
  __syncthreads();
  const uint64_t stride = (uint64_t)gridDim.x * RS_THREADS;
  for (uint64_t i = (uint64_t)blockIdx.x * RS_THREADS + threadIdx.x; i < n; i += stride) {
    const K key = ld_stream_key(keys + i);
    for (int p = 0; p < passes; p++)
      atomicAdd(&h[p][(uint32_t)(key >> (begin_bit + 8 * p)) & pass_mask(begin_bit, end_bit, p)], 1u);
  }
  __syncthreads();
  for (int p = 0; p < passes; p++) {
    uint32_t c = h[p][threadIdx.x];
    if (c) atomicAdd(&hist[p * RS_BINS + threadIdx.x], c);
  }
}

// exclusive prefix over the 256 values held one per thread
__device__ __forceinline__ uint32_t block_exclusive_256(uint32_t v, uint32_t *warp_sums) {
  const uint32_t lane = lane_id();
  uint32_t inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= (uint32_t)o) inc += t;
  }
  if (lane == 31) warp_sums[threadIdx.x >> 5] = inc;
  __syncthreads();
  uint32_t before = 0;
  for (uint32_t k = 0; k < (threadIdx.x >> 5); k++) before += warp_sums[k];
  return before + inc - v;
}

__global__ void __launch_bounds__(RS_THREADS) radix_digit_bases_kernel(uint32_t *hist) {
  __shared__ uint32_t warp_sums[RS_WARPS];
  uint32_t *h = hist + blockIdx.x * RS_BINS;
  h[threadIdx.x] = block_exclusive_256(h[threadIdx.x], warp_sums);
}

// digit of a key in a sort pass / in the slot partition of the word table (top 8 bits of the home slot)
struct ShiftDigit {
  int shift;
  uint32_t mask;
  __device__ __forceinline__ uint32_t operator()(uint64_t key) const { return (uint32_t)(key >> shift) & mask; }
};
struct SlotDigit {
  uint64_t mask;
  int shift;
  __device__ __forceinline__ uint32_t operator()(uint64_t key) const {
    return (uint32_t)((mix64(key) & mask) >> shift) & 255u;
  }
};

template <class Digit>
__global__ void __launch_bounds__(RS_THREADS) digit_hist_kernel(const uint64_t *__restrict__ keys, uint64_t n,
                                                                Digit digit_of, uint32_t *__restrict__ hist) {
  __shared__ uint32_t h[RS_BINS];
  h[threadIdx.x] = 0;
  __syncthreads();
  const uint64_t stride = (uint64_t)gridDim.x * RS_THREADS;
  for (uint64_t i = (uint64_t)blockIdx.x * RS_THREADS + threadIdx.x; i < n; i += stride)
    atomicAdd(&h[digit_of(ld_stream_u64(keys + i))], 1u);
  __syncthreads();
  if (h[threadIdx.x]) atomicAdd(&hist[threadIdx.x], h[threadIdx.x]);
}

// HAS_VALS with vals_in == nullptr: the value of element i is index_base + i.
// Persistent CTAs taking tiles by ticket.  A CTA draws the ticket of its NEXT tile when it starts
// on the current one and pulls that tile into L2 meanwhile, which halves the exposed load latency
// of a kernel that is otherwise a chain of dependent phases (load, rank, look-back, reorder,
// store).  A ticket is only ever held by a running CTA that works through its (at most two)
// tickets in order, so every tile a look-back waits for makes progress whatever else shares the
// GPU.
// Dynamic shared memory: RS_TILE keys (+ RS_TILE values).
__device__ __forceinline__ void prefetch_l2(const void *p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }

// Shared memory of one CTA of radix_pass_kernel (behind the dynamic part: RS_TILE keys + RS_TILE values).
struct RadixShared {
  uint32_t cnt[RS_PASS_WARPS][RS_BINS];
  uint32_t gbase[RS_BINS];   // output position of tile-local position 0 of a digit
  uint32_t warp_sums[RS_PASS_WARPS];
  uint32_t next;
};

// Lanes of the (converged) warp whose `dig` agrees with this lane's in the bit `mask`.  Written out in PTX so that the
// bit test, the ballot and the complement for lanes holding a 0 share one predicate: four instructions per bit
// (test, VOTE, SEL, LOP3) where the compiler made six of the C++ form.
__device__ __forceinline__ uint32_t same_bit_lanes(uint32_t dig, uint32_t mask) {
  uint32_t m, flip;
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      ".reg .b32 t;\n\t"
      "and.b32 t, %2, %3;\n\t"
      "setp.ne.u32 p, t, 0;\n\t"
      "vote.sync.ballot.b32 %0, p, 0xffffffff;\n\t"
      "selp.b32 %1, 0, 0xffffffff, p;\n\t"
      "}"
      : "=r"(m), "=r"(flip)
      : "r"(dig), "r"(mask));
  return m ^ flip;
}

// One tile.  FULL: all RS_TILE elements exist (every tile but the last), no bounds tests in the loops.
template <bool HAS_VALS, bool FULL, class Digit, class K>
__device__ __forceinline__ uint32_t radix_tile(const K *__restrict__ keys_in, const uint32_t *__restrict__ vals_in,
                                               K *__restrict__ keys_out, uint32_t *__restrict__ vals_out,
                                               uint64_t n, uint32_t n_tiles, const Digit &digit_of, uint32_t my_base,
                                               volatile uint64_t *status, uint32_t *ticket, uint32_t index_base,
                                               uint32_t tile, K *stage, uint32_t *svals, RadixShared &sh) {
  const int w = threadIdx.x >> 5;
  const uint32_t lane = lane_id();
  const uint32_t lt = (1u << lane) - 1u;
  const uint32_t d = threadIdx.x;  // the digit this thread looks after
  if (threadIdx.x == 0) sh.next = atomicAdd(ticket, 1u);
#pragma unroll
  for (int i = 0; i < RS_PASS_WARPS * RS_BINS / RS_PASS_THREADS; i++) (&sh.cnt[0][0])[i * RS_PASS_THREADS + threadIdx.x] = 0;
  const uint64_t tile_base = (uint64_t)tile * RS_TILE;
  const uint32_t in_tile = FULL ? (uint32_t)RS_TILE : (uint32_t)(n - tile_base);
  const uint32_t warp_first = (uint32_t)w * (32 * RS_ITEMS) + lane;  // tile-local index of this lane's first element
  const K *kin = keys_in + tile_base + warp_first;
  K key[RS_ITEMS];
  uint32_t val[HAS_VALS ? RS_ITEMS : 1];
  uint16_t lpos[RS_ITEMS];
#pragma unroll
  for (int j = 0; j < RS_ITEMS; j++) key[j] = (FULL || warp_first + j * 32 < in_tile) ? ld_stream_key(kin + j * 32) : (K)0;
  if (HAS_VALS) {
    if (vals_in) {
      const uint32_t *vin = vals_in + tile_base + warp_first;
#pragma unroll
      for (int j = 0; j < RS_ITEMS; j++) val[j] = (FULL || warp_first + j * 32 < in_tile) ? __ldcs(vin + j * 32) : 0u;
    } else {
      const uint32_t v0 = index_base + (uint32_t)tile_base + warp_first;
#pragma unroll
      for (int j = 0; j < RS_ITEMS; j++) val[j] = v0 + j * 32;
    }
  }
  __syncthreads();  // counters cleared, next ticket known
  const uint32_t next_tile = sh.next;
  if (next_tile < n_tiles) {  // -> L2 (128-B lines)
    const uint64_t next = (uint64_t)next_tile * RS_TILE;
    constexpr uint32_t per_line = 128 / sizeof(K);
#pragma unroll
    for (uint32_t l = threadIdx.x; l < RS_TILE / per_line; l += RS_PASS_THREADS)
      if (next + (uint64_t)l * per_line < n) prefetch_l2(keys_in + next + (uint64_t)l * per_line);
    if (HAS_VALS && vals_in) {
#pragma unroll
      for (uint32_t l = threadIdx.x; l < RS_TILE / 32; l += RS_PASS_THREADS)
        if (next + (uint64_t)l * 32 < n) prefetch_l2(vals_in + next + (uint64_t)l * 32);
    }
  }
  // rank of every element among the elements of its digit in this warp's sub-tile
  uint32_t *const my_cnt = sh.cnt[w];
#pragma unroll
  for (int j = 0; j < RS_ITEMS; j++) {
    const bool ok = FULL || warp_first + j * 32 < in_tile;
    const uint32_t dig = ok ? digit_of(key[j]) : 0u;
    // lanes holding the same digit: eight ballots (match.any is a slow-pipe instruction: the
    // first version of this kernel was bound by it, ncu r01f)
    uint32_t peers = FULL ? 0xffffffffu : __ballot_sync(0xffffffffu, ok);
#pragma unroll
    for (int bit = 0; bit < 8; bit++) peers &= same_bit_lanes(dig, 1u << bit);
    const uint32_t before = ok ? my_cnt[dig] : 0u;
    __syncwarp();
    if (ok && (peers & lt) == 0) my_cnt[dig] = before + __popc(peers);  // one leader per digit
    __syncwarp();
    lpos[j] = (uint16_t)(before + __popc(peers & lt));
  }
  __syncthreads();
  // thread d (of the first RS_BINS threads): digit d of this tile
  const bool digit_thread = RS_PASS_THREADS == RS_BINS || threadIdx.x < RS_BINS;
  uint32_t total = 0;
  volatile uint64_t *mine = nullptr;
  if (digit_thread) {
#pragma unroll
    for (int k = 0; k < RS_PASS_WARPS; k++) {
      uint32_t c = sh.cnt[k][d];
      sh.cnt[k][d] = total;
      total += c;
    }
    mine = status + (uint64_t)tile * RS_BINS + d;
    *mine = (tile == 0 ? RS_INCLUSIVE : RS_AGGREGATE) | total;
  }
  const uint32_t local = block_exclusive_256(total, sh.warp_sums);  // first tile-local position of digit d
  if (digit_thread) {
#pragma unroll
    for (int k = 0; k < RS_PASS_WARPS; k++) sh.cnt[k][d] += local;  // -> tile-local position of warp k's first element of d
  }
  __syncthreads();
  // reorder in shared memory: every digit run becomes consecutive
#pragma unroll
  for (int j = 0; j < RS_ITEMS; j++) {
    if (FULL || warp_first + j * 32 < in_tile) {
      uint32_t dig = digit_of(key[j]);
      uint32_t p = my_cnt[dig] + lpos[j];
      stage[p] = key[j];
      if (HAS_VALS) svals[p] = val[j];
    }
  }
  // how many elements of digit d precede this tile: chained look-back over the earlier tiles,
  // RS_LOOK status words per step (independent loads)
  if (digit_thread) {
    uint32_t preceding = 0;
    if (tile > 0) {
      uint32_t t = tile;  // next tile to examine is t - 1
      bool done = false;
      while (!done) {
        uint64_t v[RS_LOOK];
#pragma unroll
        for (int q = 0; q < RS_LOOK; q++)
          v[q] = t > (uint32_t)q ? (uint64_t)status[(uint64_t)(t - 1 - q) * RS_BINS + d] : (1ull << 63);
#pragma unroll
        for (int q = 0; q < RS_LOOK; q++) {
          if (done) break;
          uint64_t x = v[q];
          while (x == 0) x = status[(uint64_t)(t - 1 - q) * RS_BINS + d];  // not published yet
          preceding += (uint32_t)x;
          done = (x & RS_INCLUSIVE) != 0;
        }
        t = t > RS_LOOK ? t - RS_LOOK : 0;
      }
      *mine = RS_INCLUSIVE | (uint64_t)(preceding + total);
    }
    sh.gbase[d] = my_base + preceding - local;
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < RS_ITEMS; k++) {
    const uint32_t idx = k * RS_PASS_THREADS + threadIdx.x;
    if (FULL || idx < in_tile) {
      const K kk = stage[idx];
      const uint32_t out = sh.gbase[digit_of(kk)] + idx;
      keys_out[out] = kk;
      if (HAS_VALS) vals_out[out] = svals[idx];
    }
  }
  __syncthreads();  // stage / gbase / cnt / next are reused by the next tile
  return next_tile;
}

template <bool HAS_VALS, class Digit, class K>
__global__ void __launch_bounds__(RS_PASS_THREADS, HUMID_RS_CTAS)
    radix_pass_kernel(const K *__restrict__ keys_in, const uint32_t *__restrict__ vals_in,
                      K *__restrict__ keys_out, uint32_t *__restrict__ vals_out, uint64_t n, uint32_t n_tiles,
                      Digit digit_of, const uint32_t *__restrict__ digit_base, volatile uint64_t *status,
                      uint32_t *ticket, uint32_t index_base) {
  extern __shared__ __align__(16) unsigned char stage_bytes[];  // [RS_TILE] keys, then [RS_TILE] u32 values
  __shared__ RadixShared sh;
  K *const stage = reinterpret_cast<K *>(stage_bytes);
  uint32_t *const svals = reinterpret_cast<uint32_t *>(stage + RS_TILE);
  const uint32_t my_base = threadIdx.x < RS_BINS ? digit_base[threadIdx.x] : 0u;

  if (threadIdx.x == 0) sh.next = atomicAdd(ticket, 1u);
  __syncthreads();
  uint32_t tile = sh.next;
  __syncthreads();
  while (tile < n_tiles) {
    if ((uint64_t)(tile + 1) * RS_TILE <= n)
      tile = radix_tile<HAS_VALS, true>(keys_in, vals_in, keys_out, vals_out, n, n_tiles, digit_of, my_base, status,
                                        ticket, index_base, tile, stage, svals, sh);
    else
      tile = radix_tile<HAS_VALS, false>(keys_in, vals_in, keys_out, vals_out, n, n_tiles, digit_of, my_base, status,
                                         ticket, index_base, tile, stage, svals, sh);
  }
}

template <bool HAS_VALS, class Digit, class K>
static void launch_radix_pass(uint32_t n_tiles, cudaStream_t stream, const K *keys_in, const uint32_t *vals_in,
                              K *keys_out, uint32_t *vals_out, uint64_t n, Digit dg, const uint32_t *digit_base,
                              uint64_t *status, uint32_t *ticket, uint32_t index_base) {
  const size_t smem = (size_t)RS_TILE * (sizeof(K) + (HAS_VALS ? 4 : 0));
  static int resident = 0;  // per instantiation: CTAs of this kernel that fit on the device at once
  if (!resident) {
    auto *fn = radix_pass_kernel<HAS_VALS, Digit, K>;
    HUMID_CUDA(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int dev = 0, sms = 0, per_sm = 0;
    HUMID_CUDA(cudaGetDevice(&dev));
    HUMID_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    HUMID_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, RS_PASS_THREADS, smem));
    if (per_sm < 1) throw CudaError{"radix_pass_kernel does not fit on an SM"};
    resident = sms * per_sm;
  }
  const unsigned grid = std::min<unsigned>(n_tiles, (unsigned)resident);
  radix_pass_kernel<HAS_VALS, Digit, K><<<grid, RS_PASS_THREADS, smem, stream>>>(keys_in, vals_in, keys_out, vals_out, n, n_tiles, dg,
                                                                          digit_base, status, ticket, index_base);
}

// Passes over bits [begin_bit, end_bit).  The first pass reads (in_keys, in_vals) and writes buffer
// `first_out`; later passes ping-pong between the two buffers.  in_vals == nullptr with value
// buffers present: the value of input element i is i.  Returns the buffer holding the result.
static int radix_sort_impl(const uint64_t *in_keys, const uint32_t *in_vals, uint64_t *keys[2], uint32_t *vals[2],
                           int first_out, uint64_t n, int begin_bit, int end_bit, Scratch &scratch,
                           cudaStream_t stream, int *launches) {
  if (n >= 0xFFFFFFFFull) throw CudaError{"radix sort: more than 2^32-1 elements"};
  const int passes = (end_bit - begin_bit + 7) / 8;
  if (passes > RS_MAX_PASSES) throw CudaError{"radix sort: more than 64 key bits"};
  const uint32_t n_tiles = (uint32_t)((n + RS_TILE - 1) / RS_TILE);
  // scratch: [status: n_tiles*256 u64][tickets: 8 u32 (one per pass)][hist: passes*256 u32]
  const size_t status_bytes = (size_t)n_tiles * RS_BINS * sizeof(uint64_t);
  const size_t tail_bytes = (RS_MAX_PASSES + RS_MAX_PASSES * RS_BINS) * sizeof(uint32_t);
  char *base = (char *)scratch.get(status_bytes + tail_bytes);
  uint64_t *status = (uint64_t *)base;
  uint32_t *tickets = (uint32_t *)(base + status_bytes);
  uint32_t *hist = tickets + RS_MAX_PASSES;
  HUMID_CUDA(cudaMemsetAsync(tickets, 0, tail_bytes, stream));
  unsigned hist_grid = (unsigned)std::min<uint64_t>((n + RS_THREADS - 1) / RS_THREADS, 148 * 8);
  radix_digit_hist_kernel<<<hist_grid, RS_THREADS, 0, stream>>>(in_keys, n, begin_bit, end_bit, passes, hist);
  radix_digit_bases_kernel<<<passes, RS_THREADS, 0, stream>>>(hist);
  *launches += 2;
  const bool has_vals = vals && vals[0];
  const uint64_t *src_k = in_keys;
  const uint32_t *src_v = in_vals;
  int out = first_out;
  for (int p = 0; p < passes; p++) {
    const int shift = begin_bit + 8 * p;
    const int left = end_bit - shift;
    const ShiftDigit dg{shift, left >= 8 ? 255u : (1u << left) - 1u};
    HUMID_CUDA(cudaMemsetAsync(status, 0, status_bytes, stream));
    if (has_vals)
      launch_radix_pass<true>(n_tiles, stream, src_k, src_v, keys[out], vals[out], n, dg, hist + p * RS_BINS, status,
                              tickets + p, 0u);
    else
      launch_radix_pass<false>(n_tiles, stream, src_k, nullptr, keys[out], nullptr, n, dg, hist + p * RS_BINS, status,
                               tickets + p, 0u);
    *launches += 1;
    src_k = keys[out];
    src_v = has_vals ? vals[out] : nullptr;
    out ^= 1;
  }
  HUMID_CUDA(cudaGetLastError());
  return out ^ 1;
}

// Sort of (32-bit key, position) pairs: see primitives.cuh.
int radix_sort_u32_indexed(const uint32_t *keys_in, uint32_t *keys[2], uint32_t *index[2], uint64_t n, int bits, Scratch &scratch,
                           cudaStream_t stream, int *launches) {
  if (n == 0 || bits <= 0 || bits > 32) throw CudaError{"radix_sort_u32_indexed: 1..32 bits"};
  if (n >= 0xFFFFFFFFull) throw CudaError{"radix sort: more than 2^32-1 elements"};
  const int passes = (bits + 7) / 8;
  const uint32_t n_tiles = (uint32_t)((n + RS_TILE - 1) / RS_TILE);
  const size_t status_bytes = (size_t)n_tiles * RS_BINS * sizeof(uint64_t);
  const size_t tail_bytes = (RS_MAX_PASSES + RS_MAX_PASSES * RS_BINS) * sizeof(uint32_t);
  char *base = (char *)scratch.get(status_bytes + tail_bytes);
  uint64_t *status = (uint64_t *)base;
  uint32_t *tickets = (uint32_t *)(base + status_bytes);
  uint32_t *hist = tickets + RS_MAX_PASSES;
  HUMID_CUDA(cudaMemsetAsync(tickets, 0, tail_bytes, stream));
  unsigned hist_grid = (unsigned)std::min<uint64_t>((n + RS_THREADS - 1) / RS_THREADS, 148 * 8);
  radix_digit_hist_kernel<<<hist_grid, RS_THREADS, 0, stream>>>(keys_in, n, 0, bits, passes, hist);
  radix_digit_bases_kernel<<<passes, RS_THREADS, 0, stream>>>(hist);
  *launches += 2;
  const uint32_t *src_k = keys_in, *src_v = nullptr;  // (no values yet: the value of element i is i)
  int out = 0;
  for (int p = 0; p < passes; p++) {
    const int left = bits - 8 * p;
    const ShiftDigit dg{8 * p, left >= 8 ? 255u : (1u << left) - 1u};
    HUMID_CUDA(cudaMemsetAsync(status, 0, status_bytes, stream));
    launch_radix_pass<true>(n_tiles, stream, src_k, src_v, keys[out], index[out], n, dg, hist + p * RS_BINS, status, tickets + p, 0u);
    *launches += 1;
    src_k = keys[out];
    src_v = index[out];
    out ^= 1;
  }
  HUMID_CUDA(cudaGetLastError());
  return out ^ 1;
}

int radix_sort_u64(uint64_t *keys[2], uint32_t *vals[2], uint64_t n, int end_bit, Scratch &scratch,
                   Scratch &scan_scratch, cudaStream_t stream, int *launches) {
  (void)scan_scratch;
  if (n <= 1 || end_bit <= 0) return 0;
  return radix_sort_impl(keys[0], vals ? vals[0] : nullptr, keys, vals, 1, n, 0, end_bit, scratch, stream, launches);
}

int radix_sort_bits_indexed(const uint64_t *keys_in, uint64_t *keys[2], uint32_t *index[2], uint64_t n, int begin_bit,
                            int end_bit, Scratch &scratch, cudaStream_t stream, int *launches) {
  if (n == 0 || end_bit <= begin_bit) throw CudaError{"radix_sort_bits_indexed: nothing to sort by"};
  return radix_sort_impl(keys_in, nullptr, keys, index, 0, n, begin_bit, end_bit, scratch, stream, launches);
}

// digit = destination part of a word in the exchange of one join block: hash of the block's letters
struct DestDigit {
  uint64_t mask;
  uint32_t n_parts;
  __device__ __forceinline__ uint32_t operator()(uint64_t key) const {
    // (high word of the hash * n_parts) >> 32.  Written with the intrinsic: nvcc 12.9 folds the
    // 64-bit form "((h >> 32) * n) >> 32" into the shared-memory address arithmetic of the
    // histogram kernel as h_hi * (n >> 30), i.e. always part 0 (seen in the SASS and on the device).
    return __umulhi((uint32_t)(mix64(key & mask) >> 32), n_parts);
  }
};
// digit = owner of the row of a directed entry (row << 32 | column), rows are global ranks
struct OwnerDigit {
  uint32_t n_parts;
  uint32_t bases[17];
  __device__ __forceinline__ uint32_t operator()(uint64_t entry) const {
    const uint32_t row = (uint32_t)(entry >> 32);
    uint32_t p = 0;
#pragma unroll
    for (int q = 1; q < 16; q++) p += (q < (int)n_parts && row >= bases[q]) ? 1u : 0u;
    return p;
  }
};

// One stable partition pass by `dg`; host_counts (may be nullptr) receives the size of the first
// n_counts parts.
template <bool HAS_VALS, class Digit>
static int partition_pass(const uint64_t *keys_in, const uint32_t *vals_in, uint32_t index_base, uint64_t *keys_out,
                          uint32_t *vals_out, uint64_t n, Digit dg, uint64_t *host_counts, uint32_t n_counts,
                          Scratch &scratch, cudaStream_t stream) {
  for (uint32_t p = 0; host_counts && p < n_counts; p++) host_counts[p] = 0;
  if (n == 0) return 0;
  if (n >= 0xFFFFFFFFull) throw CudaError{"partition: more than 2^32-1 elements"};
  const uint32_t n_tiles = (uint32_t)((n + RS_TILE - 1) / RS_TILE);
  const size_t status_bytes = (size_t)n_tiles * RS_BINS * sizeof(uint64_t);
  const size_t tail_bytes = (RS_MAX_PASSES + 2 * RS_BINS) * sizeof(uint32_t);
  char *base = (char *)scratch.get(status_bytes + tail_bytes);
  uint32_t *tickets = (uint32_t *)(base + status_bytes);
  uint32_t *hist = tickets + RS_MAX_PASSES;
  uint32_t *counts = hist + RS_BINS;
  HUMID_CUDA(cudaMemsetAsync(base, 0, status_bytes + tail_bytes, stream));
  unsigned hist_grid = (unsigned)std::min<uint64_t>((n + RS_THREADS - 1) / RS_THREADS, 148 * 8);
  digit_hist_kernel<<<hist_grid, RS_THREADS, 0, stream>>>(keys_in, n, dg, hist);
  if (host_counts) HUMID_CUDA(cudaMemcpyAsync(counts, hist, RS_BINS * sizeof(uint32_t), cudaMemcpyDeviceToDevice, stream));
  radix_digit_bases_kernel<<<1, RS_THREADS, 0, stream>>>(hist);
  launch_radix_pass<HAS_VALS>(n_tiles, stream, keys_in, vals_in, keys_out, vals_out, n, dg, hist, (uint64_t *)base,
                              tickets, index_base);
  HUMID_CUDA(cudaGetLastError());
  if (host_counts) {
    uint32_t hc[RS_BINS];
    HUMID_CUDA(cudaMemcpyAsync(hc, counts, n_counts * sizeof(uint32_t), cudaMemcpyDeviceToHost, stream));
    HUMID_CUDA(cudaStreamSynchronize(stream));
    for (uint32_t p = 0; p < n_counts; p++) host_counts[p] = hc[p];
  }
  return 3;
}

int partition_by_slot_u64(const uint64_t *keys_in, const uint32_t *vals_in, uint32_t index_base, uint64_t *keys_out,
                          uint32_t *vals_out, uint64_t n, uint64_t slot_mask, Scratch &scratch, cudaStream_t stream,
                          int *launches) {
  int bits = 0;
  while ((slot_mask >> bits) != 0) bits++;
  const SlotDigit dg{slot_mask, bits > 8 ? bits - 8 : 0};
  int k = partition_pass<true>(keys_in, vals_in, index_base, keys_out, vals_out, n, dg, nullptr, 0, scratch, stream);
  *launches += k;
  return k;
}

int partition_by_block_hash(const uint64_t *keys_in, uint32_t index_base, uint64_t *keys_out, uint32_t *index_out,
                            uint64_t n, uint64_t block_mask, uint32_t n_parts, uint64_t *host_counts, Scratch &scratch,
                            cudaStream_t stream) {
  return partition_pass<true>(keys_in, nullptr, index_base, keys_out, index_out, n, DestDigit{block_mask, n_parts},
                              host_counts, n_parts, scratch, stream);
}

int partition_by_row_owner(const uint64_t *entries_in, uint64_t *entries_out, uint64_t n, const uint64_t *bases,
                           uint32_t n_parts, uint64_t *host_counts, Scratch &scratch, cudaStream_t stream) {
  OwnerDigit dg{};
  dg.n_parts = n_parts;
  for (uint32_t p = 0; p <= n_parts && p < 17; p++) dg.bases[p] = (uint32_t)bases[p];
  return partition_pass<false>(entries_in, nullptr, 0u, entries_out, nullptr, n, dg, host_counts, n_parts, scratch,
                               stream);
}

int radix_sort_bits_pairs(const uint64_t *keys_in, const uint32_t *vals_in, uint64_t *keys[2], uint32_t *vals[2],
                          uint64_t n, int begin_bit, int end_bit, Scratch &scratch, cudaStream_t stream, int *launches) {
  if (n == 0 || end_bit <= begin_bit) throw CudaError{"radix_sort_bits_pairs: nothing to sort by"};
  return radix_sort_impl(keys_in, vals_in, keys, vals, 0, n, begin_bit, end_bit, scratch, stream, launches);
}

// ------------------------------------------------------------------------------------
// The slot partition for callers that must not allocate (include/humid_gpu.h: humid_gpu_partition_by_slot_noalloc;
// no counterpart in the reference).  Same kernels, same result as
// partition_by_slot_u64; the scratch (tile status words, ticket, histogram) lives in a file-scope device array
// instead of the handle's growable arena, so nothing here can reach cudaMalloc / cudaFree.  Every GPU that loads the
// library holds the array until the process exits.  All calls on a device share it and it keeps the last call's
// contents: the part a call uses is cleared on its stream first, and two calls that overlap on the GPU race.
// ------------------------------------------------------------------------------------
constexpr uint64_t NOALLOC_MAX_TILES = (HUMID_NOALLOC_MAX_ELEMENTS + RS_TILE - 1) / RS_TILE;
constexpr size_t NOALLOC_TAIL_BYTES = (RS_MAX_PASSES + 2 * RS_BINS) * sizeof(uint32_t);
constexpr size_t NOALLOC_SCRATCH_BYTES = NOALLOC_MAX_TILES * RS_BINS * sizeof(uint64_t) + NOALLOC_TAIL_BYTES;
static __device__ __align__(256) unsigned char noalloc_scratch[NOALLOC_SCRATCH_BYTES];

// In host code `noalloc_scratch` is only a stand-in for the symbol, not a device address: the address is asked for
// once per device, with that device current, and cached.
constexpr int NOALLOC_MAX_DEVICES = 64;
static std::mutex noalloc_mutex;
static void *noalloc_address[NOALLOC_MAX_DEVICES] = {};

static void *noalloc_scratch_on_current_device() {
  int device = -1;
  HUMID_CUDA(cudaGetDevice(&device));
  if (device < 0 || device >= NOALLOC_MAX_DEVICES) throw CudaError{"no-allocation scratch: device ordinal out of range"};
  std::lock_guard<std::mutex> lock(noalloc_mutex);
  if (!noalloc_address[device]) {
    void *p = nullptr;
    HUMID_CUDA(cudaGetSymbolAddress(&p, noalloc_scratch));
    if (!p) throw CudaError{"no-allocation scratch: cudaGetSymbolAddress returned a null address"};
    noalloc_address[device] = p;
  }
  return noalloc_address[device];
}

size_t partition_noalloc_scratch_bytes() { return NOALLOC_SCRATCH_BYTES; }

int partition_by_slot_u64_noalloc(const uint64_t *keys_in, uint32_t index_base, uint64_t *keys_out, uint32_t *vals_out,
                                  uint64_t n, uint64_t slot_mask, cudaStream_t stream) {
  if (n == 0) return 0;
  if (n > HUMID_NOALLOC_MAX_ELEMENTS) throw CudaError{"no-allocation partition: more elements than the static scratch holds"};
  int bits = 0;
  while ((slot_mask >> bits) != 0) bits++;
  const SlotDigit dg{slot_mask, bits > 8 ? bits - 8 : 0};
  const uint32_t n_tiles = (uint32_t)((n + RS_TILE - 1) / RS_TILE);
  // [status: n_tiles * 256 u64][tickets: 8 u32][hist: 256 u32][unused: 256 u32] — the layout of partition_pass
  const size_t status_bytes = (size_t)n_tiles * RS_BINS * sizeof(uint64_t);
  char *base = static_cast<char *>(noalloc_scratch_on_current_device());
  uint32_t *tickets = reinterpret_cast<uint32_t *>(base + status_bytes);
  uint32_t *hist = tickets + RS_MAX_PASSES;
  HUMID_CUDA(cudaMemsetAsync(base, 0, status_bytes + NOALLOC_TAIL_BYTES, stream));  // the last call's contents
  const unsigned hist_grid = (unsigned)std::min<uint64_t>((n + RS_THREADS - 1) / RS_THREADS, 148 * 8);
  digit_hist_kernel<<<hist_grid, RS_THREADS, 0, stream>>>(keys_in, n, dg, hist);
  radix_digit_bases_kernel<<<1, RS_THREADS, 0, stream>>>(hist);
  launch_radix_pass<true>(n_tiles, stream, keys_in, nullptr, keys_out, vals_out, n, dg, hist,
                          reinterpret_cast<uint64_t *>(base), tickets, index_base);
  HUMID_CUDA(cudaGetLastError());
  return 3;
}

}  // namespace humid
