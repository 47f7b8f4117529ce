// K3 — Hamming neighbour search by block join (variant probing only for pathological groups),
// and CSR construction.
//
// Replaces Trie::asymmetricHamming -> hamming_<4,NLeaf,false>
// (lib/trie/src/trieFunctions.tcc:156-181) as called from findHammingNeighbours
// (src/humid.cc:113-130).  The asymmetric search of query w yields every stored v with
// v[p] >= w[p] at EVERY position and ham(w,v) <= d (component-wise dominance, not
// lexicographic order); the caller links both directions.  So a pair of words is linked iff
// they differ in 1..d positions and one of them dominates the other, and list(L) ends up as all
// linked words in ascending rank, without duplicates.
//
// Block join: cut the n letters into d+1 blocks.  Two words that differ in at most d positions
// agree on at least one whole block, so every linked pair sits inside one group of the words
// ordered by that block ("first" = lowest block they agree on, which makes the pair unique).
// Block 0 is the word's prefix: rank order already groups by it.  For every other block the
// (key, rank) pairs are stably radix-sorted by the block's bits, which keeps rank order inside a
// group.  One thread per word then walks forward through its group and tests each later member
// with XOR / popcount.  Nothing is looked up: no hash probes, no false positives.
//
// A group larger than the cap (low-complexity data: thousands of words sharing a block) would
// make the walk quadratic.  Such groups are skipped whole; their members are flagged per block and
// resolved afterwards by enumerating their upward variants against the word table, keeping only
// the hits whose first common block is a flagged one — the pair set stays exactly the same.
#include <algorithm>
#include <cstdlib>

#include "state.cuh"

namespace humid {

// Undirected pair (a < b in rank) + the two degrees.
__device__ __forceinline__ void emit_pair(uint64_t *pairs, uint64_t cap, unsigned long long *counter, uint32_t *deg,
                                          uint32_t a, uint32_t b) {
  const unsigned long long pos = reserve_slots(counter);
  if (pos < cap) {
    pairs[pos] = ((uint64_t)a << 32) | b;
    atomicAdd(deg + a, 1u);
    atomicAdd(deg + b, 1u);
  }
}

constexpr uint64_t LETTER_LOW = 0x5555555555555555ull;

// lowest block on which two words (x = a ^ b) agree; n_blocks if none.  Unrolled with constant
// indices: B is a kernel parameter and must stay in the constant bank, not be copied to the stack.
__device__ __forceinline__ uint32_t first_common_block(const Blocks &B, uint64_t x) {
  uint32_t first = B.n_blocks;
#pragma unroll
  for (int b = HAM_MAX_D; b >= 0; b--)
    if ((uint32_t)b < B.n_blocks && (x & B.mask[b]) == 0) first = (uint32_t)b;
  return first;
}

// do two words (x = a ^ b) agree on a block before `block`?  `block` is uniform: the unrolled loop
// leaves by a uniform branch, so block 0 costs nothing and block 1 one test.
__device__ __forceinline__ bool shares_earlier_block(const Blocks &B, uint32_t block, uint64_t x) {
  bool shared = false;
#pragma unroll
  for (int b = 0; b < HAM_MAX_D; b++) {
    if ((uint32_t)b >= block) break;
    shared |= (x & B.mask[b]) == 0;
  }
  return shared;
}

// a precedes b in rank, 1..d letters differ: linked iff a[p] <= b[p] wherever they differ
// DOMINANCE false: any pair within Hamming distance d (candidates of the edit-distance search)
template <bool DOMINANCE>
__device__ __forceinline__ bool linked(uint64_t a, uint64_t b, uint32_t d) {
  const uint64_t x = a ^ b;
  const uint64_t diff = (x | (x >> 1)) & LETTER_LOW;
  if ((uint32_t)__popcll(diff) > d) return false;
  if (!DOMINANCE || d == 1) return true;  // d == 1: the only differing letter decides the order, and a < b
  const uint64_t ah = (a >> 1) & LETTER_LOW, bh = (b >> 1) & LETTER_LOW, al = a & LETTER_LOW, bl = b & LETTER_LOW;
  const uint64_t a_greater = (ah & ~bh) | (~(ah ^ bh) & al & ~bl);
  return (a_greater & diff) == 0;
}

// One thread per position i of the block order.  keys: the words in that order; rank == nullptr
// means the order IS rank order.  A group is skipped (and its members flagged with bit `block`)
// iff it has more than group_cap members; every member reaches the same verdict by looking at
// most group_cap positions back and forth.
//
// EMIT (state.cuh): JOIN_PAIRS_DEG   pairs[k] = a << 32 | b, deg[a]++, deg[b]++        (single GPU)
//                   JOIN_DIRECTED    out[2k] = a << 32 | (col_base + b), out[2k+1] = b << 32 | (col_base + a),
//                                    deg (if not null) counted per row                  (shards)
//                   JOIN_CANDIDATES  as JOIN_DIRECTED, but every pair within Hamming distance d counts
//                                    (no dominance rule)                                 (edit-distance candidates)
// group_cap == JOIN_NO_CAP: every group is walked whatever its size, nothing is flagged.
#ifndef HUMID_JOIN_HITS
#define HUMID_JOIN_HITS 4
#endif
#ifndef HUMID_JOIN_CTAS
#define HUMID_JOIN_CTAS 4  // walk kernel at 32 registers: 4 CTAs of 512 per SM (0.44 + 0.47 -> 0.39 + 0.40 ms on the bench workload)
#endif
#ifndef HUMID_JOIN_WIDE_CTAS
#define HUMID_JOIN_WIDE_CTAS 3  // 40 registers instead of 64: 0.46 + 0.67 -> 0.36 + 0.52 ms at -n 36 on the bench workload
#endif
constexpr int JOIN_LOCAL_HITS = HUMID_JOIN_HITS;
#ifndef HUMID_JOIN_THREADS
#define HUMID_JOIN_THREADS 512
#endif
constexpr int JOIN_THREADS = HUMID_JOIN_THREADS;

template <int EMIT>
__device__ __forceinline__ void join_emit(uint64_t *out, uint64_t pos, uint32_t a, uint32_t b, uint32_t col_base,
                                          uint32_t *deg) {
  if (EMIT == JOIN_PAIRS_DEG) {
    out[pos] = ((uint64_t)a << 32) | b;
  } else {
    out[2 * pos] = ((uint64_t)a << 32) | (col_base + b);
    out[2 * pos + 1] = ((uint64_t)b << 32) | (col_base + a);
  }
  if (deg) atomicAdd(deg + b, 1u);
}

// A thread's hits stay in registers (offsets from its own position); when the buffer is full it is
// appended at once (rare), at the end the whole CTA appends with ONE atomic on the pair counter — a
// per-pair or even per-warp atomic on that single address would serialise the kernel.
struct HitBuf {
  uint32_t hits[JOIN_LOCAL_HITS];
  uint32_t nh = 0;
};

template <int EMIT>
__device__ __forceinline__ void hit_push(HitBuf &hb, uint32_t t, uint64_t i, uint32_t ri, const uint32_t *rank,
                                         uint64_t *pairs, uint64_t cap, unsigned long long *counter, uint32_t *deg,
                                         uint32_t col_base) {
  if (hb.nh == JOIN_LOCAL_HITS) {
    unsigned long long pos = atomicAdd(counter, (unsigned long long)hb.nh);
#pragma unroll
    for (int k = 0; k < JOIN_LOCAL_HITS; k++, pos++) {
      const uint32_t rb = rank ? rank[i + hb.hits[k]] : (uint32_t)(i + hb.hits[k]);
      if (pos < cap) join_emit<EMIT>(pairs, pos, ri, rb, col_base, deg);
    }
    if (pos <= cap && deg) atomicAdd(deg + ri, hb.nh);
    hb.nh = 0;
  }
#pragma unroll
  for (int k = 0; k < JOIN_LOCAL_HITS; k++)  // constant indices keep the buffer in registers
    if (k == (int)hb.nh) hb.hits[k] = t;
  hb.nh++;
}

// every thread of the CTA must call this
template <int EMIT>
__device__ __forceinline__ void hits_flush_cta(const HitBuf &hb, uint64_t i, uint32_t ri, const uint32_t *rank,
                                               uint64_t *pairs, uint64_t cap, unsigned long long *counter,
                                               uint32_t *deg, uint32_t col_base) {
  const uint32_t lane = lane_id();
  uint32_t inc = hb.nh;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
    if ((int)lane >= o) inc += t;
  }
  __shared__ uint32_t warp_total[JOIN_THREADS / 32];
  __shared__ unsigned long long cta_base;
  const uint32_t wid = threadIdx.x >> 5;
  if (lane == 31) warp_total[wid] = inc;
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t sum = 0;
    for (int k = 0; k < JOIN_THREADS / 32; k++) {
      uint32_t c = warp_total[k];
      warp_total[k] = sum;
      sum += c;
    }
    cta_base = sum ? atomicAdd(counter, (unsigned long long)sum) : 0ull;
  }
  __syncthreads();
  if (hb.nh) {
    unsigned long long pos = cta_base + warp_total[wid] + (inc - hb.nh);
    if (pos + hb.nh <= cap) {
#pragma unroll
      for (int k = 0; k < JOIN_LOCAL_HITS; k++) {
        if (k < (int)hb.nh) {
          const uint32_t rb = rank ? rank[i + hb.hits[k]] : (uint32_t)(i + hb.hits[k]);
          join_emit<EMIT>(pairs, pos + k, ri, rb, col_base, deg);
        }
      }
      if (deg) atomicAdd(deg + ri, hb.nh);
    }
  }
}

// Small groups (a word or two on average, the usual case at d = 1): one thread per position i of the
// block order walks forward through its group.  A group is skipped (and its members flagged with
// bit `block`) iff it has more than group_cap members; every member reaches the same verdict by
// looking at most group_cap positions back and forth.
template <int EMIT>
__global__ void __launch_bounds__(JOIN_THREADS, HUMID_JOIN_CTAS)
    hamming_join_walk_kernel(const uint64_t *__restrict__ keys, const uint32_t *__restrict__ rank, uint64_t U,
                             uint32_t d, const Blocks B, uint32_t block, uint64_t mask, uint32_t group_cap,
                             uint64_t *__restrict__ pairs, uint64_t cap, unsigned long long *__restrict__ counter,
                             uint32_t *__restrict__ deg, uint32_t col_base, uint32_t *__restrict__ flags,
                             unsigned long long *__restrict__ n_flagged) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const bool live = i < U;
  const uint64_t w = live ? keys[i] : 0ull;
  const uint64_t g = w & mask;
  HitBuf hb;
  uint32_t ri = 0;
  if (live) {
    // members before i (stop as soon as the group is known to be too large)
    uint32_t before = 0;
    if (group_cap != JOIN_NO_CAP)
      while (before < group_cap && before < i && (keys[i - 1 - before] & mask) == g) before++;
    uint32_t after = 0;
    if (before < group_cap) {
      const uint32_t room = group_cap - before;  // the group fits iff after < room
      while (after < room && i + 1 + after < U && (keys[i + 1 + after] & mask) == g) after++;
    }
    ri = after || before >= group_cap ? (rank ? rank[i] : (uint32_t)i) : 0u;
    if (before + after >= group_cap) {
      atomicOr(flags + ri, 1u << block);
      atomicAdd(n_flagged, 1ull);
    } else {
      for (uint32_t t = 1; t <= after; t++) {
        const uint64_t v = keys[i + t];
        if (!linked<EMIT != JOIN_CANDIDATES>(w, v, d)) continue;
        if (shares_earlier_block(B, block, v ^ w)) continue;  // found in an earlier block's order
        hit_push<EMIT>(hb, t, i, ri, rank, pairs, cap, counter, deg, col_base);
      }
    }
  }
  hits_flush_cta<EMIT>(hb, i, ri, rank, pairs, cap, counter, deg, col_base);
}

// Large groups: one warp per tile of 32 consecutive positions of the block order; lane L owns position t0 + L.
// The warp learns the extent of every group that touches its tile (chunked scans backwards from
// the tile for lane 0's group and forwards for the last lane's, head ballots inside), then streams
// the keys from the tile onwards in chunks of 32: one coalesced load per chunk, every key
// broadcast by shuffle, every lane testing it against its own word.  All lanes work whatever the
// group sizes, and a key is loaded once per tile instead of once per pair.
template <int EMIT>
__global__ void __launch_bounds__(JOIN_THREADS)
    hamming_join_tile_kernel(const uint64_t *__restrict__ keys, const uint32_t *__restrict__ rank, uint64_t U, uint32_t d,
                        const Blocks B, uint32_t block, uint64_t mask, uint32_t group_cap,
                        uint64_t *__restrict__ pairs,
                        uint64_t cap, unsigned long long *__restrict__ counter, uint32_t *__restrict__ deg,
                        uint32_t col_base, uint32_t *__restrict__ flags, unsigned long long *__restrict__ n_flagged) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint32_t lane = lane_id();
  const uint64_t t0 = i - lane;
  const bool live = i < U;
  const uint64_t w = live ? keys[i] : 0ull;
  const uint64_t g = w & mask;
  HitBuf hb;
  uint32_t ri = 0;
  if (t0 < U) {  // warp-uniform
    const uint32_t last = (uint32_t)min((uint64_t)31, U - 1 - t0);
    const uint64_t g0 = __shfl_sync(0xffffffffu, g, 0), gl = __shfl_sync(0xffffffffu, g, last);
    const uint64_t reach = group_cap == JOIN_NO_CAP ? ~0ull : (uint64_t)group_cap;
    // Chunked scans.  The result is exact, or (after `reach` positions) merely known to exceed the cap.
    uint64_t back = 0;
    while (true) {
      const uint64_t off = back + lane + 1;
      const bool same = off <= t0 && (keys[t0 - off] & mask) == g0;
      const unsigned b = __ballot_sync(0xffffffffu, same);
      if (b != 0xffffffffu) { back += __ffs(~b) - 1; break; }
      back += 32;
      if (back > reach) break;
    }
    uint64_t fwd = 0;
    const uint64_t behind = t0 + last + 1;
    while (true) {
      const uint64_t p = behind + fwd + lane;
      const bool same = p < U && (keys[p] & mask) == gl;
      const unsigned b = __ballot_sync(0xffffffffu, same);
      if (b != 0xffffffffu) { fwd += __ffs(~b) - 1; break; }
      fwd += 32;
      if (fwd > reach) break;
    }
    // this lane's group [start, end)
    const uint64_t gprev = __shfl_up_sync(0xffffffffu, g, 1);
    const unsigned heads = __ballot_sync(0xffffffffu, live && (lane == 0 || g != gprev));
    const int h = 31 - __clz((int)(heads & (0xffffffffu >> (31 - lane))));
    const unsigned above = lane == 31 ? 0u : (heads & (0xffffffffu << (lane + 1)));
    const uint64_t start = h <= 0 ? t0 - back : t0 + (uint64_t)h;
    const uint64_t end = above ? t0 + (uint64_t)(__ffs(above) - 1) : behind + fwd;
    const bool oversized = live && group_cap != JOIN_NO_CAP && end - start > (uint64_t)group_cap;
    const bool testing = live && !oversized && end > i + 1;
    if (testing || oversized) ri = rank ? rank[i] : (uint32_t)i;
    if (oversized) {
      atomicOr(flags + ri, 1u << block);
      atomicAdd(n_flagged, 1ull);
    }
    const uint64_t my_end = testing ? end : 0ull;
    uint64_t max_end = my_end;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) max_end = max(max_end, __shfl_xor_sync(0xffffffffu, max_end, o));
    for (uint64_t c = t0; c < max_end; c += 32) {
      const uint64_t kc = c + lane < U ? keys[c + lane] : 0ull;
      const int m = (int)min((uint64_t)32, max_end - c);
      for (int q = 0; q < m; q++) {
        const uint64_t v = __shfl_sync(0xffffffffu, kc, q);
        const uint64_t j = c + q;
        if (j > i && j < my_end && linked<EMIT != JOIN_CANDIDATES>(w, v, d) && !shares_earlier_block(B, block, v ^ w))
          hit_push<EMIT>(hb, (uint32_t)(j - i), i, ri, rank, pairs, cap, counter, deg, col_base);
      }
    }
  }
  hits_flush_cta<EMIT>(hb, i, ri, rank, pairs, cap, counter, deg, col_base);
}

// Flagged words only: one thread per (word, first modified position); iterative enumeration of
// the upward variants with at most d modified positions, positions strictly increasing.  A hit
// counts iff the first block the two words share is one the word is flagged for (the join
// skipped exactly those groups).
__global__ void __launch_bounds__(128)
    hamming_flagged_probe_kernel(const uint64_t *__restrict__ word_key, uint64_t U, uint32_t n, uint32_t d,
                                 const Blocks B, const uint32_t *__restrict__ flags, const Lookup lk,
                                 uint64_t *__restrict__ pairs, uint64_t cap, unsigned long long *__restrict__ counter,
                                 uint32_t *__restrict__ deg) {
  uint64_t gid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  uint64_t r = gid / n;
  if (r >= U) return;
  const uint32_t fl = flags[r];
  if (fl == 0) return;
  const uint64_t w = word_key[r];
  uint32_t pos[HAM_MAX_D];
  uint32_t let[HAM_MAX_D];
  uint64_t cur = w;
  int depth = 0;
  pos[0] = (uint32_t)(gid - r * n);
  let[0] = (uint32_t)(w >> (2 * (n - 1 - pos[0]))) & 3u;
  while (true) {
    uint32_t sh = 2 * (n - 1 - pos[depth]);
    uint32_t orig = (uint32_t)(w >> sh) & 3u;
    if (++let[depth] > 3) {
      cur = (cur & ~(3ull << sh)) | ((uint64_t)orig << sh);  // restore this position
      if (depth == 0) break;
      if (++pos[depth] >= n) { depth--; continue; }
      let[depth] = (uint32_t)(w >> (2 * (n - 1 - pos[depth]))) & 3u;
      continue;
    }
    cur = (cur & ~(3ull << sh)) | ((uint64_t)let[depth] << sh);
    const uint32_t fb = first_common_block(B, cur ^ w);
    if (fb < B.n_blocks && ((fl >> fb) & 1u)) {
      uint32_t hit = lookup_rank(lk, cur);
      if (hit != NONE32) emit_pair(pairs, cap, counter, deg, (uint32_t)r, hit);
    }
    if (depth + 1 < (int)d && pos[depth] + 1 < n) {
      depth++;
      pos[depth] = pos[depth - 1] + 1;
      let[depth] = (uint32_t)(w >> (2 * (n - 1 - pos[depth]))) & 3u;
    }
  }
}

// ---- words of 32..63 letters: the same join on (hi, lo) key halves ---------------------------
// Same group walk as hamming_join_walk_kernel, every group walked whatever its size (blocks are at
// least 16 letters long here, so groups are tiny unless the data is degenerate); pairs + degrees out.
template <bool DOMINANCE>
__device__ __forceinline__ bool linked_wide(uint64_t a_hi, uint64_t a_lo, uint64_t b_hi, uint64_t b_lo, uint32_t d) {
  const uint64_t xh = a_hi ^ b_hi, xl = a_lo ^ b_lo;
  const uint64_t dh = (xh | (xh >> 1)) & LETTER_LOW, dl = (xl | (xl >> 1)) & LETTER_LOW;
  if ((uint32_t)(__popcll(dh) + __popcll(dl)) > d) return false;
  if (!DOMINANCE || d == 1) return true;
  const uint64_t ahh = (a_hi >> 1) & LETTER_LOW, bhh = (b_hi >> 1) & LETTER_LOW, ahl = a_hi & LETTER_LOW,
                 bhl = b_hi & LETTER_LOW;
  const uint64_t alh = (a_lo >> 1) & LETTER_LOW, blh = (b_lo >> 1) & LETTER_LOW, all_ = a_lo & LETTER_LOW,
                 bll = b_lo & LETTER_LOW;
  const uint64_t gh = (ahh & ~bhh) | (~(ahh ^ bhh) & ahl & ~bhl);
  const uint64_t gl = (alh & ~blh) | (~(alh ^ blh) & all_ & ~bll);
  return ((gh & dh) | (gl & dl)) == 0;
}

__device__ __forceinline__ bool shares_earlier_block_wide(const WideBlocks &B, uint32_t block, uint64_t xh, uint64_t xl) {
  bool shared = false;
#pragma unroll
  for (int b = 0; b < HAM_MAX_D; b++) {
    if ((uint32_t)b >= block) break;
    shared |= ((xh & B.mask_hi[b]) | (xl & B.mask_lo[b])) == 0;
  }
  return shared;
}

__global__ void __launch_bounds__(JOIN_THREADS, HUMID_JOIN_WIDE_CTAS)
    hamming_join_wide_kernel(const uint64_t *__restrict__ keys_hi, const uint64_t *__restrict__ keys_lo,
                             const uint32_t *__restrict__ rank, uint64_t U, uint32_t d, const WideBlocks B, uint32_t block,
                             uint64_t *__restrict__ pairs, uint64_t cap, unsigned long long *__restrict__ counter,
                             uint32_t *__restrict__ deg) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const bool live = i < U;
  const uint64_t mh = B.mask_hi[block], ml = B.mask_lo[block];
  const uint64_t wh = live ? keys_hi[i] : 0ull, wl = live ? keys_lo[i] : 0ull;
  const uint64_t gh = wh & mh, gl = wl & ml;
  HitBuf hb;
  uint32_t ri = 0;
  if (live) {
    bool have_rank = false;
    for (uint64_t j = i + 1; j < U; j++) {
      const uint64_t vh = keys_hi[j], vl = keys_lo[j];
      if ((vh & mh) != gh || (vl & ml) != gl) break;
      if (!linked_wide<true>(wh, wl, vh, vl, d)) continue;
      if (shares_earlier_block_wide(B, block, vh ^ wh, vl ^ wl)) continue;  // found in an earlier block's order
      if (!have_rank) { ri = rank ? rank[i] : (uint32_t)i; have_rank = true; }
      hit_push<JOIN_PAIRS_DEG>(hb, (uint32_t)(j - i), i, ri, rank, pairs, cap, counter, deg, 0u);
    }
  }
  hits_flush_cta<JOIN_PAIRS_DEG>(hb, i, ri, rank, pairs, cap, counter, deg, 0u);
}

void launch_hamming_join_wide(State &s, const uint64_t *keys_hi, const uint64_t *keys_lo, const uint32_t *rank,
                              uint64_t n_words, const WideBlocks &B, uint32_t block, uint64_t *out, uint64_t cap,
                              unsigned long long *counter, uint32_t *deg) {
  hamming_join_wide_kernel<<<grid_for(n_words, JOIN_THREADS), JOIN_THREADS, 0, s.stream>>>(
      keys_hi, keys_lo, rank, n_words, s.cfg.distance, B, block, out, cap, counter, deg);
}

void launch_hamming_join(State &s, int emit, const uint64_t *keys, const uint32_t *rank, uint64_t n_words, const Blocks &B,
                         uint32_t block, uint32_t group_cap, uint64_t *out, uint64_t cap, unsigned long long *counter,
                         uint32_t *deg, uint32_t col_base, uint32_t *flags, unsigned long long *n_flagged) {
  const uint32_t d = s.cfg.distance;
  const unsigned grid = grid_for(n_words, JOIN_THREADS);
  // expected group size if the block values were uniform: words / 4^letters.  Measured: at 1.3 and at 13
  // words per group the per-thread walks win (0.43 vs 1.06 ms, 9.3 vs 13.5 ms), at 300 the tile kernel
  // does (5.7 vs 8.9 ms); the switch sits in between.
  const int letters = __builtin_popcountll(B.mask[block]) / 2;
  const double expected = letters >= 32 ? 0.0 : (double)n_words / (double)(1ull << (2 * letters));
  bool tile = expected >= 48.0;
  if (const char *e = std::getenv("HUMID_GPU_JOIN_KERNEL"))  // testing knob: "tile" or "walk"
    tile = e[0] == 't';
#define HUMID_JOIN(KERNEL, MODE)                                                                              \
  KERNEL<MODE><<<grid, JOIN_THREADS, 0, s.stream>>>(keys, rank, n_words, d, B, block, B.mask[block], group_cap, out, \
                                                    cap, counter, deg, col_base, flags, n_flagged)
  if (emit == JOIN_PAIRS_DEG) {
    if (tile) HUMID_JOIN(hamming_join_tile_kernel, JOIN_PAIRS_DEG);
    else HUMID_JOIN(hamming_join_walk_kernel, JOIN_PAIRS_DEG);
  } else if (emit == JOIN_DIRECTED) {
    if (tile) HUMID_JOIN(hamming_join_tile_kernel, JOIN_DIRECTED);
    else HUMID_JOIN(hamming_join_walk_kernel, JOIN_DIRECTED);
  } else {
    if (tile) HUMID_JOIN(hamming_join_tile_kernel, JOIN_CANDIDATES);
    else HUMID_JOIN(hamming_join_walk_kernel, JOIN_CANDIDATES);
  }
#undef HUMID_JOIN
}

Blocks make_blocks(uint32_t n, uint32_t d, int lo_bit[], int hi_bit[]) {
  Blocks B{};
  B.n_blocks = d + 1;
  for (uint32_t b = 0; b <= d; b++) {
    const uint32_t p0 = b * n / (d + 1), p1 = (b + 1) * n / (d + 1);  // letters [p0, p1); may be empty if n <= d
    lo_bit[b] = 2 * (int)(n - p1);
    hi_bit[b] = 2 * (int)(n - p0);
    B.mask[b] = p1 > p0 ? ((~0ull >> (64 - 2 * (p1 - p0))) << lo_bit[b]) : 0ull;
  }
  return B;
}

// What the single-GPU Hamming search joins on.  Entry e: the words are grouped by the letters under
// B.mask[e] (obtained from rank order by up to two stable sorts over contiguous bit ranges, least significant
// range first; a group inside the prefix needs none), and a pair is reported by the first entry both words
// agree on.
//   * d + 1 blocks (make_blocks): one block per entry.
//   * d == 2: FOUR blocks, and an entry per pair of blocks — two words within distance 2 agree on at least two
//     of four blocks.  Six entries instead of three, but a group is now the words sharing half of their letters
//     (about one word at 20 M words of 24 letters) instead of a third (about 300): at 50 M tuples of
//     configs[2] the three joins took 5.7 ms each.  A pair of blocks that contains the prefix block only
//     needs the other block sorted: inside a group of that block, rank order keeps equal prefixes together.
JoinPlan make_join_plan(uint32_t n, uint32_t d) {
  JoinPlan P{};
  bool combos = d == 2 && n >= 8;
  if (const char *e = std::getenv("HUMID_GPU_JOIN_SCHEME"))  // testing knob: "blocks" or "combos"
    combos = combos && e[0] == 'c';
  if (!combos) {
    int lo[HAM_MAX_D + 1], hi[HAM_MAX_D + 1];
    P.B = make_blocks(n, d, lo, hi);
    for (uint32_t b = 0; b <= d; b++) {
      P.n_sorts[b] = b > 0 && hi[b] > lo[b] ? 1 : 0;
      P.lo_bit[b][0] = lo[b];
      P.hi_bit[b][0] = hi[b];
    }
    return P;
  }
  int lo[4], hi[4];
  uint64_t mask[4];
  for (uint32_t b = 0; b < 4; b++) {
    const uint32_t p0 = b * n / 4, p1 = (b + 1) * n / 4;  // letters [p0, p1), never empty (n >= 8)
    lo[b] = 2 * (int)(n - p1);
    hi[b] = 2 * (int)(n - p0);
    mask[b] = (~0ull >> (64 - 2 * (p1 - p0))) << lo[b];
  }
  uint32_t e = 0;
  for (int a = 0; a < 4; a++)
    for (int b = a + 1; b < 4; b++, e++) {
      P.B.mask[e] = mask[a] | mask[b];
      if (a == 0 && b == 1) {
        P.n_sorts[e] = 0;  // the prefix: rank order
      } else if (a == 0 || b == a + 1) {  // the prefix block + one other, or two adjacent blocks: one bit range
        P.n_sorts[e] = 1;
        P.lo_bit[e][0] = lo[b];
        P.hi_bit[e][0] = a == 0 ? hi[b] : hi[a];
      } else {  // two separate ranges: the less significant one first
        P.n_sorts[e] = 2;
        P.lo_bit[e][0] = lo[b];
        P.hi_bit[e][0] = hi[b];
        P.lo_bit[e][1] = lo[a];
        P.hi_bit[e][1] = hi[a];
      }
    }
  P.B.n_blocks = e;  // 6
  return P;
}

// The words in the order of entry b of the plan: keys (and their ranks; nullptr = the order is rank order).
// Uses key_alt / entries_alt / entry_val / entry_val_alt (U + 1 elements each, ensured by the caller).
int join_plan_order(State &s, const JoinPlan &P, uint32_t b, uint64_t U, const uint64_t *&keys, const uint32_t *&rank) {
  keys = s.word_key.p;
  rank = nullptr;
  if (P.n_sorts[b] == 0) return 0;
  uint64_t *kb[2] = {s.key_alt.p, s.entries_alt.p};
  uint32_t *vb[2] = {s.entry_val.p, s.entry_val_alt.p};
  int launches = 0;
  uint64_t passes = (uint64_t)(P.hi_bit[b][0] - P.lo_bit[b][0] + 7) / 8;  // first pass reads keys only (rank = position)
  if (P.n_sorts[b] > 1) passes += (uint64_t)(P.hi_bit[b][1] - P.lo_bit[b][1] + 7) / 8 + 1;  // + its own histogram read
  Timed t(s, "join_sort", U, U * (8 + (8 + 12) + (passes - 1) * 24));
  const int cur = radix_sort_bits_indexed(s.word_key.p, kb, vb, U, P.lo_bit[b][0], P.hi_bit[b][0], s.scratch, s.stream,
                                          &launches);
  keys = kb[cur];
  rank = vb[cur];
  if (P.n_sorts[b] > 1) {  // then, stably, by the more significant range: the first pass writes the free pair
    uint64_t *kb2[2] = {kb[1 - cur], kb[cur]};
    uint32_t *vb2[2] = {vb[1 - cur], vb[cur]};
    const int cur2 = radix_sort_bits_pairs(kb[cur], vb[cur], kb2, vb2, U, P.lo_bit[b][1], P.hi_bit[b][1], s.scratch,
                                           s.stream, &launches);
    keys = kb2[cur2];
    rank = vb2[cur2];
  }
  t.launches = launches;
  return launches;
}

// HUMID_GPU_JOIN_CAP: group size above which the join hands a group to the probe path (testing knob)
static uint32_t join_group_cap(uint32_t d) {
  if (const char *e = std::getenv("HUMID_GPU_JOIN_CAP")) {
    long v = std::strtol(e, nullptr, 10);
    if (v >= 1) return (uint32_t)v;
  }
  return d == 1 ? 64u : 2048u;
}

int stage_neighbours_hamming(State &s) {
  const uint64_t U = s.unique;
  const uint32_t n = s.cfg.word_length, d = s.cfg.distance;
  s.n_entries = s.n_pairs = 0;
  s.deg.ensure(U + 1);
  if (U == 0 || d == 0) {
    if (U) HUMID_CUDA(cudaMemsetAsync(s.deg.p, 0, U * sizeof(uint32_t), s.stream));
    return 0;
  }
  if (d > HAM_MAX_D) throw CudaError{"Hamming distance > 8 is not supported"};
  int k = 0;
  const JoinPlan P = make_join_plan(n, d);
  const Blocks &B = P.B;
  const uint32_t group_cap = join_group_cap(d);
  uint64_t cap = s.entries.cap ? s.entries.cap : std::max<uint64_t>(2 * U, 1u << 20);
  s.word_flags.ensure(U + 1);
  s.key_alt.ensure(U + 1);
  s.entries_alt.ensure(U + 1);
  s.entry_val.ensure(U + 1);
  s.entry_val_alt.ensure(U + 1);
  unsigned long long *counter = s.counters.p + 2, *n_flagged = s.counters.p + 40;
  for (int attempt = 0; attempt < 2; attempt++) {
    s.entries.ensure(cap);
    cap = s.entries.cap;
    HUMID_CUDA(cudaMemsetAsync(counter, 0, sizeof(unsigned long long), s.stream));
    HUMID_CUDA(cudaMemsetAsync(n_flagged, 0, sizeof(unsigned long long), s.stream));
    HUMID_CUDA(cudaMemsetAsync(s.deg.p, 0, U * sizeof(uint32_t), s.stream));
    HUMID_CUDA(cudaMemsetAsync(s.word_flags.p, 0, U * sizeof(uint32_t), s.stream));
    for (uint32_t b = 0; b < B.n_blocks; b++) {
      const uint64_t *keys;
      const uint32_t *rank;
      k += join_plan_order(s, P, b, U, keys, rank);
      // the word, its successor in the group (the walk ends there for almost every word) and the pair out
      static const char *const join_names[HAM_MAX_D + 1] = {"join_hamming_b0", "join_hamming_b1", "join_hamming_b2",
                                                            "join_hamming_b3", "join_hamming_b4", "join_hamming_b5",
                                                            "join_hamming_b6", "join_hamming_b7", "join_hamming_b8"};
      Timed t(s, join_names[b], U, U * (8 + 8 + (rank ? 4 : 0)));
      launch_hamming_join(s, JOIN_PAIRS_DEG, keys, rank, U, B, b, group_cap, s.entries.p, cap, counter, s.deg.p, 0u,
                          s.word_flags.p, n_flagged);
      k++;
    }
    HUMID_CUDA(cudaMemcpyAsync(&s.n_flagged, n_flagged, sizeof(uint64_t), cudaMemcpyDeviceToHost, s.stream));
    HUMID_CUDA(cudaStreamSynchronize(s.stream));
    if (s.n_flagged) {  // rare: groups beyond the cap go through the word table instead
      k += stage_ensure_lookup(s);
      Timed t(s, "probe_flagged", s.n_flagged, s.n_flagged * 8);
      hamming_flagged_probe_kernel<<<grid_for(U * n, 128), 128, 0, s.stream>>>(s.word_key.p, U, n, d, B, s.word_flags.p,
                                                                              s.lookup(), s.entries.p, cap, counter,
                                                                              s.deg.p);
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

// ------------------------------------------------------------------------------------
// CSR from undirected pairs + degrees (Hamming mode): row_ptr = scan(deg); every pair drops
// its two entries at an atomic cursor of its rows; rows are then put in ascending order
// (the reference's list order for Hamming lists, no duplicates possible).
// ------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
    csr_rows_kernel(const uint32_t *__restrict__ excl, uint64_t U, uint64_t total, uint64_t *__restrict__ row_ptr) {
  uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r < U) row_ptr[r] = excl[r];
  if (r == U) row_ptr[U] = total;
}

__global__ void __launch_bounds__(256)
    csr_fill_kernel(const uint64_t *__restrict__ pairs, uint64_t P, const uint32_t *__restrict__ excl,
                    uint32_t *__restrict__ cursor, uint32_t *__restrict__ col) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P) return;
  uint64_t e = ld_stream_u64(pairs + i);
  uint32_t a = (uint32_t)(e >> 32), b = (uint32_t)e;
  col[excl[a] + atomicAdd(cursor + a, 1u)] = b;
  col[excl[b] + atomicAdd(cursor + b, 1u)] = a;
}

// one thread per row: insertion sort (rows are short: mean < 1 entry at d = 1)
constexpr uint32_t ROW_SORT_SERIAL_MAX = 48;
__global__ void __launch_bounds__(256)
    csr_sort_rows_kernel(const uint32_t *__restrict__ excl, const uint32_t *__restrict__ deg, uint64_t U,
                         uint32_t *__restrict__ col, uint32_t *__restrict__ big_rows, unsigned long long *n_big) {
  uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= U) return;
  const uint32_t m = deg[r];
  if (m < 2) return;
  if (m > ROW_SORT_SERIAL_MAX) {  // rare: handed to the cooperative kernel
    big_rows[atomicAdd(n_big, 1ull)] = (uint32_t)r;
    return;
  }
  uint32_t *row = col + excl[r];
  for (uint32_t i = 1; i < m; i++) {
    uint32_t x = row[i];
    uint32_t j = i;
    while (j > 0 && row[j - 1] > x) { row[j] = row[j - 1]; j--; }
    row[j] = x;
  }
}

// one CTA per long row: bitonic sort in shared memory for rows up to 4096 entries, otherwise
// an odd-even transposition sort in global memory (quadratic, but such rows are pathological)
constexpr uint32_t ROW_SORT_SMEM = 4096;
__global__ void __launch_bounds__(256)
    csr_sort_big_rows_kernel(const uint32_t *__restrict__ excl, const uint32_t *__restrict__ deg,
                             const uint32_t *__restrict__ big_rows, uint32_t *__restrict__ col) {
  __shared__ uint32_t buf[ROW_SORT_SMEM];
  const uint32_t r = big_rows[blockIdx.x];
  const uint32_t m = deg[r];
  uint32_t *row = col + excl[r];
  if (m <= ROW_SORT_SMEM) {
    uint32_t p2 = 1;
    while (p2 < m) p2 <<= 1;
    for (uint32_t i = threadIdx.x; i < p2; i += blockDim.x) buf[i] = i < m ? row[i] : 0xFFFFFFFFu;
    __syncthreads();
    for (uint32_t k = 2; k <= p2; k <<= 1)
      for (uint32_t j = k >> 1; j > 0; j >>= 1) {
        for (uint32_t i = threadIdx.x; i < p2; i += blockDim.x) {
          uint32_t l = i ^ j;
          if (l > i) {
            uint32_t x = buf[i], y = buf[l];
            bool up = (i & k) == 0;
            if ((x > y) == up) { buf[i] = y; buf[l] = x; }
          }
        }
        __syncthreads();
      }
    for (uint32_t i = threadIdx.x; i < m; i += blockDim.x) row[i] = buf[i];
  } else {
    for (uint32_t pass = 0; pass < m; pass++) {
      for (uint32_t i = (pass & 1u) + 2 * threadIdx.x; i + 1 < m; i += 2 * blockDim.x) {
        uint32_t x = row[i], y = row[i + 1];
        if (x > y) { row[i] = y; row[i + 1] = x; }
      }
      __syncthreads();
    }
  }
}

// Directed entries (row << 32 | column) with the list lengths already counted in `deg`
// (shard mode: rows are shard-local ranks, columns global ranks).
__global__ void __launch_bounds__(256)
    csr_fill_directed_kernel(const uint64_t *__restrict__ entries, uint64_t E, const uint32_t *__restrict__ excl,
                             uint32_t *__restrict__ cursor, uint32_t *__restrict__ col) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= E) return;
  uint64_t e = ld_stream_u64(entries + i);
  uint32_t a = (uint32_t)(e >> 32);
  col[excl[a] + atomicAdd(cursor + a, 1u)] = (uint32_t)e;
}

static int csr_from_degrees(State &s, uint64_t E, bool pairs) {
  const uint64_t U = s.unique;
  s.row_ptr.ensure(U + 2);
  s.col.ensure(E + 1);
  s.work_a.ensure(U + 1);  // exclusive scan of deg
  s.work_b.ensure(U + 1);  // fill cursors, then the list of long rows
  int k = 0;
  Timed t(s, "csr_build", E, E * 12 + U * 16);
  if (U) k += exclusive_scan_u32(s.deg.p, s.work_a.p, U, nullptr, s.scan_scratch, s.stream);
  csr_rows_kernel<<<grid_for(U + 1, 256), 256, 0, s.stream>>>(s.work_a.p, U, E, s.row_ptr.p);
  k++;
  if (E) {
    HUMID_CUDA(cudaMemsetAsync(s.work_b.p, 0, U * sizeof(uint32_t), s.stream));
    if (pairs)
      csr_fill_kernel<<<grid_for(E / 2, 256), 256, 0, s.stream>>>(s.entries.p, E / 2, s.work_a.p, s.work_b.p, s.col.p);
    else
      csr_fill_directed_kernel<<<grid_for(E, 256), 256, 0, s.stream>>>(s.entries.p, E, s.work_a.p, s.work_b.p, s.col.p);
    HUMID_CUDA(cudaMemsetAsync(s.counters.p + 7, 0, sizeof(unsigned long long), s.stream));
    csr_sort_rows_kernel<<<grid_for(U, 256), 256, 0, s.stream>>>(s.work_a.p, s.deg.p, U, s.col.p, s.work_b.p,
                                                                 s.counters.p + 7);
    k += 2;
    uint64_t n_big = 0;
    HUMID_CUDA(cudaMemcpyAsync(&n_big, s.counters.p + 7, sizeof(uint64_t), cudaMemcpyDeviceToHost, s.stream));
    HUMID_CUDA(cudaStreamSynchronize(s.stream));
    if (n_big) {
      csr_sort_big_rows_kernel<<<(unsigned)n_big, 256, 0, s.stream>>>(s.work_a.p, s.deg.p, s.work_b.p, s.col.p);
      k++;
    }
  }
  HUMID_CUDA(cudaGetLastError());
  t.launches = k;
  return k;
}

int stage_csr_pairs(State &s) {
  s.n_entries = 2 * s.n_pairs;
  s.cluster_pairs = s.entries.p;  // the union-find takes the pairs as they are
  s.n_cluster_pairs = s.n_pairs;
  s.cluster_entries = nullptr;
  return csr_from_degrees(s, s.n_entries, true);
}

int stage_csr_directed(State &s) {
  s.cluster_pairs = nullptr;
  s.cluster_entries = s.entries.p;  // csr_fill_directed_kernel only reads them
  s.n_cluster_entries = s.n_entries;
  return csr_from_degrees(s, s.n_entries, false);
}

// ------------------------------------------------------------------------------------
// CSR: sort the directed entries by (source rank, order key), then split.
// ------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
    csr_split_kernel(const uint64_t *__restrict__ entries, const uint32_t *__restrict__ vals, uint64_t E, uint64_t U,
                     uint64_t *__restrict__ row_ptr, uint32_t *__restrict__ col) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i > E) return;
  // rows (prev_src, src] start at i; the virtual entry E closes all remaining rows
  uint64_t src = i < E ? (entries[i] >> 32) : U;
  uint64_t prev = i == 0 ? 0 : (entries[i - 1] >> 32) + 1;
  for (uint64_t r = prev; r <= src; r++) row_ptr[r] = i;
  if (i < E) col[i] = vals ? vals[i] : (uint32_t)entries[i];
}

int stage_csr(State &s) {
  const uint64_t U = s.unique, E = s.n_entries;
  s.cluster_pairs = nullptr;
  s.cluster_entries = nullptr;
  int k = 0;
  s.row_ptr.ensure(U + 2);
  s.col.ensure(E + 1);
  const bool pairs = s.cfg.edit && s.cfg.distance >= 2;
  int ubits = 1;
  while ((1ull << ubits) < U) ubits++;
  const uint32_t *vals = nullptr;
  if (E) {
    s.entries_alt.ensure(s.entries.cap);
    uint64_t *kb[2] = {s.entries.p, s.entries_alt.p};
    uint32_t *vb[2] = {nullptr, nullptr};
    if (pairs) {
      s.entry_val_alt.ensure(s.entry_val.cap);
      vb[0] = s.entry_val.p;
      vb[1] = s.entry_val_alt.p;
    }
    int launches = 0;
    int cur;
    {
      Timed t(s, "sort_entries", E, E * 16);
      cur = radix_sort_u64(kb, vb, E, 32 + ubits, s.scratch, s.scan_scratch, s.stream, &launches);
      t.launches = launches;
    }
    k += launches;
    if (cur == 1) {
      std::swap(s.entries, s.entries_alt);
      if (pairs) std::swap(s.entry_val, s.entry_val_alt);
    }
    if (pairs) vals = s.entry_val.p;
  }
  {
    Timed t(s, "csr_split", E, E * 12 + U * 8);
    csr_split_kernel<<<grid_for(E + 1, 256), 256, 0, s.stream>>>(s.entries.p, vals, E, U, s.row_ptr.p, s.col.p);
  }
  k++;
  HUMID_CUDA(cudaGetLastError());
  return k;
}

}  // namespace humid
