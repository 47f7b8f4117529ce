// Exclusive scan + stable LSD radix sort (see primitives.cuh).
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
// Stable LSD radix sort, 8 bits per pass.  Per pass:
//   radix_hist_kernel     per-tile digit histogram            hist[digit][tile]
//   exclusive_scan_u32    over hist in (digit, tile) order    -> global offset of every
//                                                                (digit, tile) run
//   radix_scatter_kernel  stable in-tile ranking with warp match + per-warp counters
// A tile is RS_THREADS*RS_ITEMS consecutive elements; warp w owns a contiguous sub-tile and
// walks it in strips of 32 consecutive elements, so "element order" == (warp, strip, lane).
// ------------------------------------------------------------------------------------
constexpr int RS_THREADS = 256;
constexpr int RS_WARPS = RS_THREADS / 32;
constexpr int RS_ITEMS = 16;
constexpr int RS_TILE = RS_THREADS * RS_ITEMS;
constexpr int RS_BINS = 256;

__global__ void __launch_bounds__(RS_THREADS) radix_hist_kernel(const uint64_t *__restrict__ keys, uint64_t n,
                                                                int shift, uint32_t *__restrict__ hist,
                                                                uint32_t n_tiles) {
  __shared__ uint32_t h[RS_BINS];
  h[threadIdx.x] = 0;
  __syncthreads();
  uint64_t base = (uint64_t)blockIdx.x * RS_TILE;
#pragma unroll
  for (int j = 0; j < RS_ITEMS; j++) {
    uint64_t i = base + (uint64_t)j * RS_THREADS + threadIdx.x;
    if (i < n) atomicAdd(&h[(uint32_t)(ld_stream_u64(keys + i) >> shift) & 255u], 1u);
  }
  __syncthreads();
  hist[(uint64_t)threadIdx.x * n_tiles + blockIdx.x] = h[threadIdx.x];
}

template <bool HAS_VALS>
__global__ void __launch_bounds__(RS_THREADS)
    radix_scatter_kernel(const uint64_t *__restrict__ keys_in, const uint32_t *__restrict__ vals_in,
                         uint64_t *__restrict__ keys_out, uint32_t *__restrict__ vals_out, uint64_t n, int shift,
                         const uint32_t *__restrict__ offsets, uint32_t n_tiles) {
  __shared__ uint32_t cnt[RS_WARPS][RS_BINS];
  const int w = threadIdx.x >> 5;
  const uint32_t lane = lane_id();
  const uint32_t lt = (1u << lane) - 1u;
#pragma unroll
  for (int k = 0; k < RS_WARPS; k++) cnt[k][threadIdx.x] = 0;
  __syncthreads();

  const uint64_t warp_base = (uint64_t)blockIdx.x * RS_TILE + (uint64_t)w * (32 * RS_ITEMS);
  uint64_t key[RS_ITEMS];
  uint32_t val[RS_ITEMS];
  // phase 1: load, per-warp digit totals
#pragma unroll
  for (int j = 0; j < RS_ITEMS; j++) {
    uint64_t i = warp_base + (uint64_t)j * 32 + lane;
    bool ok = i < n;
    key[j] = ok ? ld_stream_u64(keys_in + i) : 0ull;
    if (HAS_VALS) val[j] = ok ? vals_in[i] : 0u;
    uint32_t dig = ok ? ((uint32_t)(key[j] >> shift) & 255u) : 256u;
    uint32_t peers = __match_any_sync(0xffffffffu, dig);
    if (ok && (peers & lt) == 0) cnt[w][dig] += __popc(peers);  // one leader per digit
    __syncwarp();
  }
  __syncthreads();
  // per digit: global base of this tile's run, then exclusive prefix over the warps
  {
    uint32_t d = threadIdx.x;
    uint32_t run = offsets[(uint64_t)d * n_tiles + blockIdx.x];
#pragma unroll
    for (int k = 0; k < RS_WARPS; k++) {
      uint32_t c = cnt[k][d];
      cnt[k][d] = run;
      run += c;
    }
  }
  __syncthreads();
  // phase 2: stable rank inside the warp strip, scatter
#pragma unroll
  for (int j = 0; j < RS_ITEMS; j++) {
    uint64_t i = warp_base + (uint64_t)j * 32 + lane;
    bool ok = i < n;
    uint32_t dig = ok ? ((uint32_t)(key[j] >> shift) & 255u) : 256u;
    uint32_t peers = __match_any_sync(0xffffffffu, dig);
    uint32_t rank = __popc(peers & lt);
    uint32_t pos = 0;
    if (ok) {
      pos = cnt[w][dig] + rank;
    }
    __syncwarp();
    if (ok && rank == 0) cnt[w][dig] += __popc(peers);
    __syncwarp();
    if (ok) {
      keys_out[pos] = key[j];
      if (HAS_VALS) vals_out[pos] = val[j];
    }
  }
}

int radix_sort_u64(uint64_t *keys[2], uint32_t *vals[2], uint64_t n, int end_bit, Scratch &scratch,
                   Scratch &scan_scratch, cudaStream_t stream, int *launches) {
  if (n <= 1 || end_bit <= 0) return 0;
  if (n >= 0xFFFFFFFFull) throw CudaError{"radix_sort_u64: more than 2^32-1 elements"};
  uint32_t n_tiles = (uint32_t)((n + RS_TILE - 1) / RS_TILE);
  uint64_t hist_n = (uint64_t)RS_BINS * n_tiles;
  uint32_t *hist = (uint32_t *)scratch.get(hist_n * sizeof(uint32_t));
  int cur = 0;
  const bool has_vals = vals && vals[0];
  for (int shift = 0; shift < end_bit; shift += 8) {
    radix_hist_kernel<<<n_tiles, RS_THREADS, 0, stream>>>(keys[cur], n, shift, hist, n_tiles);
    *launches += 1 + exclusive_scan_u32(hist, hist, hist_n, nullptr, scan_scratch, stream);
    if (has_vals)
      radix_scatter_kernel<true><<<n_tiles, RS_THREADS, 0, stream>>>(keys[cur], vals[cur], keys[cur ^ 1],
                                                                    vals[cur ^ 1], n, shift, hist, n_tiles);
    else
      radix_scatter_kernel<false><<<n_tiles, RS_THREADS, 0, stream>>>(keys[cur], nullptr, keys[cur ^ 1], nullptr, n,
                                                                     shift, hist, n_tiles);
    *launches += 1;
    cur ^= 1;
  }
  HUMID_CUDA(cudaGetLastError());
  return cur;
}

}  // namespace humid
