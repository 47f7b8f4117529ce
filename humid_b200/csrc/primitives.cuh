// Device-wide primitives written for this library (no CUB/Thrust): exclusive scan and a
// stable LSD radix sort.  Both are plain HBM-streaming kernels: coalesced loads, shared
// memory staging, warp match/ballot/shuffle for the in-tile ranking.
#pragma once
#include "common.cuh"

namespace humid {

// Scratch arena owned by the handle; grows on demand, never shrinks.
struct Scratch {
  void *ptr = nullptr;
  size_t bytes = 0;
  Scratch() = default;
  Scratch(const Scratch &) = delete;
  Scratch &operator=(const Scratch &) = delete;
  ~Scratch() { release(); }
  void *get(size_t need) {
    if (need > bytes) {
      if (ptr) cudaFree(ptr);
      ptr = nullptr;
      bytes = 0;
      size_t want = need + need / 4 + 4096;
      HUMID_CUDA(cudaMalloc(&ptr, want));
      bytes = want;
    }
    return ptr;
  }
  void release() {
    if (ptr) cudaFree(ptr);
    ptr = nullptr;
    bytes = 0;
  }
};

// out[i] = sum(in[0..i)), n up to 2^32-1 elements, sums must fit uint32_t.
// If total_dev != nullptr the grand total is written there.  in may alias out.
// Returns the number of kernels launched.
int exclusive_scan_u32(const uint32_t *in, uint32_t *out, uint64_t n, uint32_t *total_dev,
                       Scratch &scratch, cudaStream_t stream);

// Stable LSD radix sort on bits [0, end_bit) of 64-bit keys, 8 bits per pass.
// keys/vals are double buffers: input in index 0; the function returns the index (0 or 1)
// of the buffer pair holding the sorted result.  vals may be {nullptr,nullptr} (keys only).
// *launches is incremented by the number of kernels launched.
// `scratch` holds the per-tile digit histograms, `scan_scratch` is handed to the scan.
int radix_sort_u64(uint64_t *keys[2], uint32_t *vals[2], uint64_t n, int end_bit, Scratch &scratch,
                   Scratch &scan_scratch, cudaStream_t stream, int *launches);

// Stable sort of (key, position) pairs by bits [begin_bit, end_bit) of the key.  keys_in is only
// read; element i carries the value i.  Returns the index (0 or 1) of the (keys, index) buffer
// pair holding the result.
int radix_sort_bits_indexed(const uint64_t *keys_in, uint64_t *keys[2], uint32_t *index[2], uint64_t n, int begin_bit,
                            int end_bit, Scratch &scratch, cudaStream_t stream, int *launches);

// Stable sort of (32-bit key, position) pairs by the low `bits` (1..32) bits of the key: 8 bytes per element and pass
// where the 64-bit sorts move 12.  keys_in is only read; element i carries the value i.  Returns the index (0 or 1) of
// the (keys, index) buffer pair holding the result.
int radix_sort_u32_indexed(const uint32_t *keys_in, uint32_t *keys[2], uint32_t *index[2], uint64_t n, int bits, Scratch &scratch,
                           cudaStream_t stream, int *launches);

// One stable partition pass of (key, value) pairs by the top 8 bits of the key's home slot in an
// open-addressing table with `slot_mask + 1` slots (slot = mix64(key) & slot_mask): afterwards
// consecutive elements touch one 1/256 of the table, which then stays in L2 while it is worked on.
// vals_in == nullptr: the value of element i is index_base + i.
int partition_by_slot_u64(const uint64_t *keys_in, const uint32_t *vals_in, uint32_t index_base, uint64_t *keys_out,
                          uint32_t *vals_out, uint64_t n, uint64_t slot_mask, Scratch &scratch, cudaStream_t stream,
                          int *launches);

// Multi-GPU block join (shard.cu).  Stable partition of the words by the part that joins their
// block (hash of the letters under block_mask): (key, index_base + i) pairs grouped by part,
// host_counts[p] = words for part p.  Synchronises the stream.
int partition_by_block_hash(const uint64_t *keys_in, uint32_t index_base, uint64_t *keys_out, uint32_t *index_out,
                            uint64_t n, uint64_t block_mask, uint32_t n_parts, uint64_t *host_counts, Scratch &scratch,
                            cudaStream_t stream);
// Directed entries (row << 32 | column, rows = global ranks) grouped by the part owning the row
// (bases[p] <= row < bases[p+1]).  Synchronises the stream.
int partition_by_row_owner(const uint64_t *entries_in, uint64_t *entries_out, uint64_t n, const uint64_t *bases,
                           uint32_t n_parts, uint64_t *host_counts, Scratch &scratch, cudaStream_t stream);
// Stable sort of (key, value) pairs by bits [begin_bit, end_bit); inputs are only read.
int radix_sort_bits_pairs(const uint64_t *keys_in, const uint32_t *vals_in, uint64_t *keys[2], uint32_t *vals[2],
                          uint64_t n, int begin_bit, int end_bit, Scratch &scratch, cudaStream_t stream, int *launches);

// partition_by_slot_u64 (values = index_base + i) with its scratch in a file-scope device array of primitives.cu: no
// allocation of any kind, at most HUMID_NOALLOC_MAX_ELEMENTS (include/humid_gpu.h) elements, enqueued on `stream`
// without synchronising.  Calls on one device share the scratch: they must not overlap on the GPU.
int partition_by_slot_u64_noalloc(const uint64_t *keys_in, uint32_t index_base, uint64_t *keys_out, uint32_t *vals_out,
                                  uint64_t n, uint64_t slot_mask, cudaStream_t stream);
size_t partition_noalloc_scratch_bytes();

}  // namespace humid
