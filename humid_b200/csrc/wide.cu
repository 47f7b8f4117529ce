// Words of 32..63 letters (`-n 36` for 12-nt UMIs, reference docs/usage.rst:18-20): the key is 128
// bits, kept as two 64-bit halves in separate columns (lo = the last 32 letters, hi = the letters
// before them, right-aligned; bit 63 of hi is never set, so all-ones still marks a filtered tuple).
//
// Same reference functions as words.cu / neighbours.cu — Trie::add, Trie::walk order
// (lib/trie/src/trieFunctions.tcc:45-59, :129-143), asymmetricHamming (:156-181) — with one change
// of method: there is no 128-bit word table.  The tuples are sorted by (hi, lo) with the stable
// radix sort (two key halves = two sorts, the tuple index rides along), runs of equal keys are the
// distinct words: run length = Leaf::count, first element of the run = the first tuple in input
// order (stability), run number = rank.  The sorted tuple indices and the rank of every sorted
// position are kept: scattered into a tuple -> rank column on demand, they answer what the table
// look-up of the narrow path answers (cluster ids).  The neighbour search is the
// same block join; a block other than the prefix is extracted into a 64-bit column, the (value,
// rank) pairs are sorted and the key halves gathered into that order.
#include <algorithm>

#include "state.cuh"

namespace humid {

__global__ void __launch_bounds__(256)
    gather_u64_kernel(const uint64_t *__restrict__ src, const uint32_t *__restrict__ index, uint64_t n,
                      uint64_t *__restrict__ out) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = src[index[i]];
}

// head[i] = 1 iff sorted position i starts a run of equal words (positions >= usable hold the filtered tuples)
__global__ void __launch_bounds__(256)
    wide_heads_kernel(const uint64_t *__restrict__ hi, const uint64_t *__restrict__ lo, uint64_t usable,
                      uint32_t *__restrict__ head) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= usable) return;
  head[i] = i == 0 || hi[i] != hi[i - 1] || lo[i] != lo[i - 1];
}

// rank of position i = heads before it (+ its own) - 1; it replaces the head flag in place.  Heads write the
// word columns and where their run starts.
__global__ void __launch_bounds__(256)
    wide_words_kernel(const uint64_t *__restrict__ hi, const uint64_t *__restrict__ lo, const uint32_t *__restrict__ index,
                      uint32_t *__restrict__ head_then_rank, const uint32_t *__restrict__ head_scan, uint64_t usable,
                      uint64_t *__restrict__ word_hi, uint64_t *__restrict__ word_lo, uint32_t *__restrict__ word_first,
                      uint32_t *__restrict__ run_start) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= usable) return;
  const uint32_t h = head_then_rank[i];
  const uint32_t r = head_scan[i] + h - 1u;
  head_then_rank[i] = r;
  if (h) {
    word_hi[r] = hi[i];
    word_lo[r] = lo[i];
    word_first[r] = index[i];
    run_start[r] = (uint32_t)i;
  }
}

// tuple -> rank of its word (NONE32 for the filtered tuples, which sort behind the usable ones): only
// humid_gpu_cluster_ids asks, so the scatter waits for the first such call
__global__ void __launch_bounds__(256)
    wide_tuple_ranks_kernel(const uint32_t *__restrict__ index, const uint32_t *__restrict__ pos_rank, uint64_t usable,
                            uint64_t n_tuples, uint32_t *__restrict__ tuple_rank) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_tuples) return;
  tuple_rank[index[i]] = i < usable ? pos_rank[i] : NONE32;
}

__global__ void __launch_bounds__(256)
    wide_counts_kernel(const uint32_t *__restrict__ run_start, uint64_t U, uint64_t usable,
                       uint32_t *__restrict__ word_count) {
  const uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= U) return;
  const uint32_t end = r + 1 < U ? run_start[r + 1] : (uint32_t)usable;
  word_count[r] = end - run_start[r];
}

int stage_build_words_wide(State &s) {
  const uint64_t R = s.n_tuples;
  const uint32_t n = s.cfg.word_length;
  int k = 0;
  HUMID_CUDA(cudaMemcpyAsync(&s.usable, s.counters.p + 0, sizeof(uint64_t), cudaMemcpyDeviceToHost, s.stream));
  HUMID_CUDA(cudaStreamSynchronize(s.stream));
  s.unique = 0;
  s.ranks_in_table = false;
  s.filter_buckets = 0;
  s.word_key.ensure(s.usable + 1);
  s.word_hi.ensure(s.usable + 1);
  s.word_count.ensure(s.usable + 1);
  s.word_first.ensure(s.usable + 1);
  s.wide_ranks_ready = false;
  if (R == 0) return 0;
  s.wide_k0.ensure(R);
  s.wide_k1.ensure(R);
  s.wide_v0.ensure(R);
  s.wide_v1.ensure(R);
  const unsigned gR = grid_for(R, 256);
  uint64_t *k2[2];
  uint32_t *v2[2];
  int c2;
  {
    // 8 passes over the low half, then the bits of the high half (+ the bit above them, set only in
    // the all-ones filtered keys, which therefore end up last); (key half, tuple index) read and written per pass
    const int hi_bits = 2 * (int)(n - 32) + 1;
    const uint64_t passes = 8 + (uint64_t)(hi_bits + 7) / 8;
    Timed t(s, "sort_tuples", R, R * (passes * 2 * (8 + 4) + 2 * (4 + 8 + 8)));
    int launches = 0;
    uint64_t *k1[2] = {s.wide_k0.p, s.wide_k1.p};
    uint32_t *v1[2] = {s.wide_v0.p, s.wide_v1.p};
    const int c1 = radix_sort_bits_indexed(s.tuple_key.p, k1, v1, R, 0, 64, s.scratch, s.stream, &launches);
    // the high halves in that order go where the first sort no longer needs room
    gather_u64_kernel<<<gR, 256, 0, s.stream>>>(s.tuple_hi.p, v1[c1], R, k1[1 - c1]);
    k2[0] = k1[c1];      // first pass of the second sort reads (k1[1-c1], v1[c1]) and writes (k2[0], v2[0])
    k2[1] = k1[1 - c1];
    v2[0] = v1[1 - c1];
    v2[1] = v1[c1];
    c2 = radix_sort_bits_pairs(k1[1 - c1], v1[c1], k2, v2, R, 0, hi_bits, s.scratch, s.stream, &launches);
    gather_u64_kernel<<<gR, 256, 0, s.stream>>>(s.tuple_key.p, v2[c2], R, k2[1 - c2]);
    t.launches = launches + 2;
    k += launches + 2;
  }
  const uint64_t *hi = k2[c2], *lo = k2[1 - c2];
  const uint32_t *index = v2[c2];
  uint32_t *head = v2[1 - c2];
  {
    Timed t(s, "word_runs", s.usable, s.usable * (16 + 4 + 4 + 4 + 4));
    s.work_a.ensure(s.usable + 1);  // scan of the heads
    s.work_b.ensure(s.usable + 1);  // where each run starts
    uint32_t *total = reinterpret_cast<uint32_t *>(s.counters.p + 1);
    HUMID_CUDA(cudaMemsetAsync(s.counters.p + 1, 0, sizeof(unsigned long long), s.stream));
    int launches = 0;
    if (s.usable) {
      wide_heads_kernel<<<grid_for(s.usable, 256), 256, 0, s.stream>>>(hi, lo, s.usable, head);
      launches += 1 + exclusive_scan_u32(head, s.work_a.p, s.usable, total, s.scan_scratch, s.stream);
    }
    if (s.usable) {
      wide_words_kernel<<<grid_for(s.usable, 256), 256, 0, s.stream>>>(hi, lo, index, head, s.work_a.p, s.usable,
                                                                      s.word_hi.p, s.word_key.p, s.word_first.p, s.work_b.p);
      launches++;
    }
    s.wide_sorted_index = index;  // both stay in wide_v0 / wide_v1 until the next job
    s.wide_pos_rank = head;
    HUMID_CUDA(cudaMemcpyAsync(&s.unique, s.counters.p + 1, sizeof(uint64_t), cudaMemcpyDeviceToHost, s.stream));
    HUMID_CUDA(cudaStreamSynchronize(s.stream));
    if (s.unique) {
      wide_counts_kernel<<<grid_for(s.unique, 256), 256, 0, s.stream>>>(s.work_b.p, s.unique, s.usable, s.word_count.p);
      launches++;
    }
    t.launches = launches;
    k += launches;
  }
  HUMID_CUDA(cudaGetLastError());
  return k;
}

// ---- neighbours ---------------------------------------------------------------------------------
// value of the letters [p0, p1) of the word, right-aligned (at most 32 letters)
__global__ void __launch_bounds__(256)
    wide_block_values_kernel(const uint64_t *__restrict__ hi, const uint64_t *__restrict__ lo, uint64_t U, int lo_bit,
                             int width, uint64_t *__restrict__ out) {
  const uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= U) return;
  const uint64_t h = hi[r], l = lo[r];
  uint64_t v;
  if (lo_bit >= 64) v = h >> (lo_bit - 64);
  else if (lo_bit == 0) v = l;
  else v = (l >> lo_bit) | (h << (64 - lo_bit));
  if (width < 64) v &= (1ull << width) - 1ull;
  out[r] = v;
}

__global__ void __launch_bounds__(256)
    gather_halves_kernel(const uint64_t *__restrict__ hi, const uint64_t *__restrict__ lo,
                         const uint32_t *__restrict__ rank, uint64_t U, uint64_t *__restrict__ out_hi,
                         uint64_t *__restrict__ out_lo) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= U) return;
  const uint32_t r = rank[i];
  out_hi[i] = hi[r];
  out_lo[i] = lo[r];
}

// bits [lo_bit, hi_bit) of the 128-bit key, split over the halves
static void wide_mask(int lo_bit, int hi_bit, uint64_t &mask_hi, uint64_t &mask_lo) {
  mask_hi = mask_lo = 0;
  for (int b = lo_bit; b < hi_bit; b++) {
    if (b >= 64) mask_hi |= 1ull << (b - 64);
    else mask_lo |= 1ull << b;
  }
}

int stage_neighbours_hamming_wide(State &s) {
  const uint64_t U = s.unique;
  const uint32_t n = s.cfg.word_length, d = s.cfg.distance;
  s.n_entries = s.n_pairs = 0;
  s.n_flagged = 0;
  s.deg.ensure(U + 1);
  if (U == 0 || d == 0) {
    if (U) HUMID_CUDA(cudaMemsetAsync(s.deg.p, 0, U * sizeof(uint32_t), s.stream));
    return 0;
  }
  if (d > HAM_MAX_D) throw CudaError{"Hamming distance > 8 is not supported"};
  int k = 0;
  WideBlocks B{};
  B.n_blocks = d + 1;
  int lo_bit[HAM_MAX_D + 1], hi_bit[HAM_MAX_D + 1];
  // Any cut of the letters into d+1 blocks finds every pair.  The prefix block costs nothing (rank order
  // groups by it), every other block one sort of its bits: those get at most 12 letters each (16 beyond
  // 64 M words) — enough for groups of a word or two — and the prefix takes the rest.
  const uint32_t other = std::min<uint32_t>(n / (d + 1), U > (64ull << 20) ? 16u : 12u);
  for (uint32_t b = 0; b <= d; b++) {
    const uint32_t p0 = b == 0 ? 0u : n - (d + 1 - b) * other, p1 = n - (d - b) * other;  // letters [p0, p1)
    lo_bit[b] = 2 * (int)(n - p1);
    hi_bit[b] = 2 * (int)(n - p0);
    wide_mask(lo_bit[b], hi_bit[b], B.mask_hi[b], B.mask_lo[b]);
  }
  uint64_t cap = s.entries.cap ? s.entries.cap : std::max<uint64_t>(2 * U, 1u << 20);
  s.key_alt.ensure(U + 1);
  s.wide_k0.ensure(U + 1);
  s.wide_k1.ensure(U + 1);
  s.wide_k2.ensure(U + 1);
  s.entry_val.ensure(U + 1);
  s.entry_val_alt.ensure(U + 1);
  unsigned long long *counter = s.counters.p + 2;
  const unsigned gU = grid_for(U, 256);
  for (int attempt = 0; attempt < 2; attempt++) {
    s.entries.ensure(cap);
    cap = s.entries.cap;
    HUMID_CUDA(cudaMemsetAsync(counter, 0, sizeof(unsigned long long), s.stream));
    HUMID_CUDA(cudaMemsetAsync(s.deg.p, 0, U * sizeof(uint32_t), s.stream));
    for (uint32_t b = 0; b <= d; b++) {
      const uint64_t *keys_hi = s.word_hi.p, *keys_lo = s.word_key.p;
      const uint32_t *rank = nullptr;
      if (b > 0) {  // never empty: n >= 32 and d <= 8
        const int width = hi_bit[b] - lo_bit[b];
        const uint64_t passes = (uint64_t)(width + 7) / 8;
        Timed t(s, "join_sort", U, U * (16 + 8 + 8 + (8 + 12) + (passes - 1) * 24 + 4 + 32));
        int launches = 0;
        wide_block_values_kernel<<<gU, 256, 0, s.stream>>>(s.word_hi.p, s.word_key.p, U, lo_bit[b], width, s.key_alt.p);
        uint64_t *kb[2] = {s.wide_k0.p, s.wide_k1.p};
        uint32_t *vb[2] = {s.entry_val.p, s.entry_val_alt.p};
        const int cur = radix_sort_bits_indexed(s.key_alt.p, kb, vb, U, 0, width, s.scratch, s.stream, &launches);
        rank = vb[cur];
        gather_halves_kernel<<<gU, 256, 0, s.stream>>>(s.word_hi.p, s.word_key.p, rank, U, s.wide_k2.p, s.key_alt.p);
        keys_hi = s.wide_k2.p;
        keys_lo = s.key_alt.p;
        t.launches = launches + 2;
        k += launches + 2;
      }
      static const char *const join_names[HAM_MAX_D + 1] = {"join_hamming_b0", "join_hamming_b1", "join_hamming_b2",
                                                            "join_hamming_b3", "join_hamming_b4", "join_hamming_b5",
                                                            "join_hamming_b6", "join_hamming_b7", "join_hamming_b8"};
      Timed t(s, join_names[b], U, U * (16 + 16 + (rank ? 4 : 0)));
      launch_hamming_join_wide(s, keys_hi, keys_lo, rank, U, B, b, s.entries.p, cap, counter, s.deg.p);
      k++;
    }
    HUMID_CUDA(cudaMemcpyAsync(&s.n_pairs, counter, sizeof(uint64_t), cudaMemcpyDeviceToHost, s.stream));
    HUMID_CUDA(cudaStreamSynchronize(s.stream));
    if (s.n_pairs <= cap) break;
    cap = s.n_pairs;  // the counter kept counting: exact size is now known, run once more
  }
  if (2 * s.n_pairs >= 0xFFFFFFFFull) throw CudaError{"more than 2^32-2 neighbour entries"};
  HUMID_CUDA(cudaGetLastError());
  return k;
}

// ---- writeAnnotated: cluster id of the tuple's word (src/humid.cc:268-285), 0 for filtered tuples ----
__global__ void __launch_bounds__(256)
    cluster_ids_by_rank_kernel(const uint32_t *__restrict__ tuple_rank, uint64_t first, uint64_t n,
                               const uint32_t *__restrict__ label, const uint32_t *__restrict__ seed_index,
                               uint32_t *__restrict__ out) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t r = tuple_rank[first + i];
  out[i] = r == NONE32 ? 0u : seed_index[label[r]] + 1u;
}

int stage_cluster_ids_wide(State &s, uint64_t first, uint64_t n, uint32_t *d_out) {
  if (n == 0) return 0;
  if (!s.wide_ranks_ready) {
    const uint64_t R = s.n_tuples;
    s.tuple_rank.ensure(R + 1);
    Timed t(s, "tuple_ranks", R, R * (4 + 4 + 4));
    wide_tuple_ranks_kernel<<<grid_for(R, 256), 256, 0, s.stream>>>(s.wide_sorted_index, s.wide_pos_rank, s.usable, R,
                                                                    s.tuple_rank.p);
    s.wide_ranks_ready = true;
  }
  Timed t(s, "cluster_ids", n, n * (4 + 4 + 4 + 4));
  cluster_ids_by_rank_kernel<<<grid_for(n, 256), 256, 0, s.stream>>>(s.tuple_rank.p, first, n, s.label.p, s.seed_index.p,
                                                                     d_out);
  HUMID_CUDA(cudaGetLastError());
  return 1;
}

}  // namespace humid
