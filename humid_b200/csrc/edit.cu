// K6 — Levenshtein neighbour search with the reference's exact multiplicities and order.
//
// Replaces Trie::asymmetricLevenshtein -> levenshtein_<4,NLeaf,false>
// (lib/trie/src/trieFunctions.tcc:196-233) as called from findEditNeighbours
// (src/humid.cc:140-158).  That recursion yields a stored word v once per distinct move
// sequence (deletion / substitution-or-match / insertion, including "phantom" deletions past
// the end of the query), every yield != self pushes both directions, and the order of a
// word's own yields is the depth-first order of the recursion.  Distance 1 is handled by the
// Hamming-1 kernel (identical result for fixed-length words); this file is d >= 2.
//
// Two phases, both exact:
//   1. candidates: every yielded v satisfies Levenshtein(w, v) <= d (a move sequence is an
//      edit script of cost <= d), so for each word all same-length words within that distance
//      are enumerated as variants (substitutions only; one deletion + one insertion + further
//      substitutions) and looked up through the fingerprint filter + table.  Hits (rank pairs
//      (q, v), q != v) are sorted and deduplicated.
//   2. replay: the order of the yields of query q is a function of the move strings only
//      (deletion < substitution/insertion by child letter < own leaf), so running the
//      recursion on the tiny trie made of q's candidates alone gives the same yield sequence,
//      order and multiplicity, as the reference's walk of the whole trie.  One thread per
//      query walks the implicit trie over its sorted candidate keys with an explicit stack.
// Own yields are appended in yield order (key = q<<32|q), the mirrored pushes as (v<<32|q);
// the stable sort of stage_csr then produces list(L) = [pushes by queries q < L] ++ [own
// yields] ++ [pushes by queries q > L], which is the reference's list order.
#include <algorithm>
#include <cstdlib>

#include "state.cuh"

namespace humid {

constexpr int EDIT_MAX_D = 3;  // one deletion+insertion pair plus substitutions

// v is a candidate of query q and q one of query v (the relation is enumerated from one side only)
__device__ __forceinline__ void emit_candidate(uint64_t *cand, uint64_t cap, unsigned long long *counter, uint32_t q,
                                               uint32_t v) {
  const unsigned long long pos = reserve_slots(counter, 2u);
  if (pos + 1 < cap) {
    cand[pos] = ((uint64_t)q << 32) | v;
    cand[pos + 1] = ((uint64_t)v << 32) | q;
  }
}

// All words obtained from `base` by 0..budget further substitutions (positions strictly
// increasing, any other letter), each looked up once.
__device__ void probe_substitutions(uint64_t base, uint32_t n, int budget, uint32_t first_pos, uint64_t self,
                                    uint32_t q, const Lookup &lk, uint64_t *cand, uint64_t cap,
                                    unsigned long long *counter, bool probe_base) {
  if (probe_base && base != self) {
    uint32_t hit = lookup_rank(lk, base);
    if (hit != NONE32) emit_candidate(cand, cap, counter, q, hit);
  }
  if (budget <= 0) return;
  uint32_t pos[EDIT_MAX_D];
  uint32_t add[EDIT_MAX_D];  // letter offset 1..3 applied at pos (mod 4)
  uint64_t cur = base;
  int depth = 0;
  pos[0] = first_pos;
  add[0] = 0;
  if (first_pos >= n) return;
  while (true) {
    const uint32_t sh = 2 * (n - 1 - pos[depth]);
    const uint64_t orig = (base >> sh) & 3ull;
    if (++add[depth] > 3) {
      cur = (cur & ~(3ull << sh)) | (orig << sh);
      if (++pos[depth] >= n) {
        if (depth == 0) break;
        depth--;
        continue;
      }
      add[depth] = 0;
      continue;
    }
    cur = (cur & ~(3ull << sh)) | (((orig + add[depth]) & 3ull) << sh);
    if (cur != self) {
      uint32_t hit = lookup_rank(lk, cur);
      if (hit != NONE32) emit_candidate(cand, cap, counter, q, hit);
    }
    if (depth + 1 < budget && pos[depth] + 1 < n) {
      depth++;
      pos[depth] = pos[depth - 1] + 1;
      add[depth] = 0;
    }
  }
}

// Deletion of letter i + insertion of letter c so that it lands at position q (a one-letter
// block shift), then up to d-2 substitutions.  Only "right" shifts (q < i) are enumerated: v is a right
// shift of w exactly when w is a left shift of v (undo the insertion, redo the deletion), so every hit
// is recorded in both directions and the other half of the (i, q) square is never probed.  Left out as
// well, because another probe or the block join finds the same word:
//   * i - q == 1: the shift changes two positions; with the d-2 substitutions still allowed that is
//     within Hamming distance d, and the block join has every such pair;
//   * a deleted letter that is followed by an equal letter: deleting any letter of a run gives the
//     same word, only the last letter of a run is deleted;
//   * inserting the letter w[q]: in front of w[q] it is the same word as behind it, i.e. the shift
//     (i, q + 1), or a word within Hamming distance d when q + 1 == i - 1.
// One CTA takes SHIFT_WPB words.  The (i, q) pairs come from a table in shared memory; the live (word,
// pair) combinations are first compacted into a shared work list, so that the probing loop runs with
// full warps whatever the words look like.
constexpr int SHIFT_WPB = 32;
constexpr int SHIFT_THREADS = 256;
constexpr int SHIFT_MAX_PAIRS = 30 * 29 / 2;  // n <= 31: (n - 1)(n - 2) / 2 pairs with i - q >= 2
template <bool D2>  // D2: d == 2 and the bitmap is there — nothing but the shift itself, probed inline
__global__ void __launch_bounds__(SHIFT_THREADS)
    edit_candidates_shift_kernel(const uint64_t *__restrict__ word_key, uint64_t U, uint32_t n, uint32_t d,
                                 const Lookup lk, uint64_t *__restrict__ cand, uint64_t cap,
                                 unsigned long long *__restrict__ counter, uint32_t n_pairs, uint32_t inv_pairs) {
  __shared__ uint16_t pair_tab[SHIFT_MAX_PAIRS];             // i << 8 | q
  __shared__ uint16_t work[SHIFT_WPB * SHIFT_MAX_PAIRS];     // word in the CTA << 9 | pair
  __shared__ uint64_t words[SHIFT_WPB];
  __shared__ uint32_t n_work;
  const uint64_t first = (uint64_t)blockIdx.x * SHIFT_WPB;
  const uint32_t n_words = (uint32_t)min((uint64_t)SHIFT_WPB, U - first);
  const uint32_t lane = threadIdx.x & 31u;
  if (threadIdx.x == 0) n_work = 0;
  if (threadIdx.x < n_words) words[threadIdx.x] = word_key[first + threadIdx.x];
  for (uint32_t t = threadIdx.x; t < n_pairs; t += SHIFT_THREADS) {  // t = j (j - 1) / 2 + q with j = i - 1, q < j
    uint32_t j = (uint32_t)((1.0f + sqrtf(1.0f + 8.0f * (float)t)) * 0.5f);
    while (j * (j - 1) / 2 > t) j--;
    while ((j + 1) * j / 2 <= t) j++;
    pair_tab[t] = (uint16_t)(((j + 1) << 8) | (t - j * (j - 1) / 2));
  }
  __syncthreads();
  const uint32_t slots = n_words * n_pairs;
  for (uint32_t base = 0; base < slots; base += SHIFT_THREADS) {
    const uint32_t t = base + threadIdx.x;
    bool live = t < slots;
    uint32_t wl = 0, pi = 0;
    if (live) {
      wl = __umulhi(t, inv_pairs);  // t / n_pairs (exact: t < 2^16)
      pi = t - wl * n_pairs;
      const uint32_t i = pair_tab[pi] >> 8;
      const uint64_t w = words[wl];
      const uint32_t sh_i = 2 * (n - 1 - i);
      live = !(i + 1 < n && ((w >> sh_i) & 3ull) == ((w >> (sh_i - 2)) & 3ull));  // only the last letter of a run
    }
    const unsigned m = __ballot_sync(0xffffffffu, live);
    uint32_t at = 0;
    if (lane == 0 && m) at = atomicAdd(&n_work, (uint32_t)__popc(m));
    at = __shfl_sync(0xffffffffu, at, 0);
    if (live) work[at + __popc(m & ((1u << lane) - 1u))] = (uint16_t)((wl << 9) | pi);
  }
  __syncthreads();
  const uint32_t total = n_work;
  Lookup table_only = lk;  // for the words that passed the bitmap here
  table_only.bitmap = nullptr;
  if (D2) {
    // two work items per turn: six independent bitmap loads in flight per thread (the loop waits on L2)
    for (uint32_t k0 = threadIdx.x; k0 < total; k0 += 2 * SHIFT_THREADS) {
      uint64_t base0[2], hb0[2], wq[2];
      uint32_t sh_q[2], r[2];
      bool have[2];
#pragma unroll
      for (int u = 0; u < 2; u++) {
        const uint32_t k = k0 + u * SHIFT_THREADS;
        have[u] = k < total;
        const uint32_t e = have[u] ? work[k] : 0u;
        const uint32_t wl = e >> 9, pr = pair_tab[e & 511u];
        const uint32_t i = pr >> 8, q = pr & 255u;
        const uint64_t w = words[wl];
        r[u] = (uint32_t)(first + wl);
        // affected positions [q, i]: letters q..i-1 move one place to the right, the new letter lands at q
        sh_q[u] = 2 * (n - 1 - q);
        const uint32_t sh_i = 2 * (n - 1 - i);
        const uint64_t seg_mask = (sh_q[u] + 2 >= 64 ? ~0ull : ((1ull << (sh_q[u] + 2)) - 1ull)) & ~((1ull << sh_i) - 1ull);
        const uint64_t moved = ((w & seg_mask) >> 2) & seg_mask;
        base0[u] = (w & ~seg_mask) | (moved & ~(3ull << sh_q[u]));
        wq[u] = (w >> sh_q[u]) & 3ull;
        hb0[u] = base0[u] * BITMAP_MUL;  // the hash is linear: one multiply per shift, one add per letter
      }
      uint32_t bits[2][3];
#pragma unroll
      for (int u = 0; u < 2; u++)
#pragma unroll
        for (uint64_t c = 1; c < 4; c++) {
          const uint64_t bit = (hb0[u] + ((wq[u] + c) & 3ull) * (BITMAP_MUL << sh_q[u])) >> lk.bitmap_shift;
          bits[u][c - 1] = (__ldg(lk.bitmap + (bit >> 5)) >> (bit & 31u)) & 1u;
        }
#pragma unroll
      for (int u = 0; u < 2; u++)
#pragma unroll
        for (uint64_t c = 1; c < 4; c++) {
          if (!have[u] || bits[u][c - 1] == 0u) continue;
          const uint32_t hit = lookup_rank(table_only, base0[u] | (((wq[u] + c) & 3ull) << sh_q[u]));
          if (hit != NONE32) emit_candidate(cand, cap, counter, r[u], hit);
        }
    }
  } else {
    for (uint32_t k = threadIdx.x; k < total; k += SHIFT_THREADS) {
      const uint32_t e = work[k];
      const uint32_t wl = e >> 9, pr = pair_tab[e & 511u];
      const uint32_t i = pr >> 8, q = pr & 255u;
      const uint64_t w = words[wl];
      const uint32_t r = (uint32_t)(first + wl);
      const uint32_t sh_q = 2 * (n - 1 - q), sh_i = 2 * (n - 1 - i);
      const uint64_t seg_mask = (sh_q + 2 >= 64 ? ~0ull : ((1ull << (sh_q + 2)) - 1ull)) & ~((1ull << sh_i) - 1ull);
      const uint64_t moved = ((w & seg_mask) >> 2) & seg_mask;
      const uint64_t base0 = (w & ~seg_mask) | (moved & ~(3ull << sh_q));
      const uint64_t wq = (w >> sh_q) & 3ull;
#pragma unroll
      for (uint64_t c = 1; c < 4; c++) {
        const uint64_t v = base0 | (((wq + c) & 3ull) << sh_q);
        probe_substitutions(v, n, (int)d - 2, 0, w, r, lk, cand, cap, counter, true);
      }
    }
  }
}

// One bit per hash value of a stored word (Lookup::bitmap)
__global__ void __launch_bounds__(256)
    bitmap_build_kernel(const uint64_t *__restrict__ word_key, uint64_t U, uint32_t *__restrict__ bits, int shift) {
  const uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= U) return;
  const uint64_t bit = (word_key[r] * BITMAP_MUL) >> shift;
  atomicOr(bits + (bit >> 5), 1u << (bit & 31u));
}

// ---- distinct candidates per query -----------------------------------------------------------
__global__ void __launch_bounds__(256)
    mark_unique_kernel(const uint64_t *__restrict__ sorted, uint64_t m, uint32_t *__restrict__ flag) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < m) flag[i] = (i == 0 || sorted[i] != sorted[i - 1]) ? 1u : 0u;
}

__global__ void __launch_bounds__(256)
    compact_candidates_kernel(const uint64_t *__restrict__ sorted, const uint32_t *__restrict__ flag,
                              const uint32_t *__restrict__ index, uint64_t m, const uint64_t *__restrict__ word_key,
                              uint32_t *__restrict__ cand_q, uint32_t *__restrict__ cand_v,
                              uint64_t *__restrict__ cand_key) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m || !flag[i]) return;
  uint32_t o = index[i];
  uint64_t e = sorted[i];
  cand_q[o] = (uint32_t)(e >> 32);
  cand_v[o] = (uint32_t)e;
  cand_key[o] = word_key[(uint32_t)e];
}

__global__ void __launch_bounds__(256)
    candidate_rows_kernel(const uint32_t *__restrict__ cand_q, uint64_t m, uint64_t U, uint32_t *__restrict__ row) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i > m) return;
  uint64_t src = i < m ? cand_q[i] : U;
  uint64_t prev = i == 0 ? 0 : (uint64_t)cand_q[i - 1] + 1;
  for (uint64_t r = prev; r <= src; r++) row[r] = (uint32_t)i;
}

// ---- replay of levenshtein_<4,T,false> on the query's own candidates ---------------------------
struct Frame {
  uint32_t lo, hi;   // candidate sub-range == trie node
  uint8_t depth;     // depth of the node
  uint8_t pos;       // position in the query word
  int8_t dist;       // remaining budget
  uint8_t stage;     // next move to try: 0 deletion, 1+2c substitution by child c, 2+2c insertion, 9 leaf
};
constexpr int EDIT_STACK = 48;

__device__ __forceinline__ uint32_t lower_letter(const uint64_t *__restrict__ keys, uint32_t lo, uint32_t hi,
                                                 uint32_t sh, uint32_t c) {
  while (lo < hi) {
    uint32_t mid = (lo + hi) >> 1;
    if (((uint32_t)(keys[mid] >> sh) & 3u) < c) lo = mid + 1; else hi = mid;
  }
  return lo;
}

// Number of times levenshtein_ yields `v` for query `w` when v is the only other word on its paths:
// the number of move sequences (frames of the recursion) that arrive at v's leaf with the query used
// up.  State = (letters of v consumed t, query position p, budget k); p - t stays within +-d, so a
// level is a (2 D + 1) x (D + 1) band of counters.  Moves exactly as in the replay below: deletion
// (t, p+1, k-1) with no bound on p; the child v[t] only if v[t] >= w[p] (w[p] = 0 past the end);
// match / substitution (t+1, p+1, k - [v[t] != w[p]]) needs p < n; insertion (t+1, p, k-1).
// The count can start anywhere: from the frame (t0 letters of v consumed, query position p0, budget
// k0) it is the number of yields of the subtree of the recursion below that frame, which is what the
// replay needs once a node of the candidate trie holds a single word.
constexpr int DPD = 3;  // EDIT_MAX_D
__device__ uint32_t candidate_yields_from(uint64_t w, uint64_t v, uint32_t n, uint32_t t0, uint32_t p0, int k0) {
  uint32_t cur[2 * DPD + 1][DPD + 1];
  const int o0 = (int)p0 - (int)t0 + DPD;  // |p - t| never exceeds the budget spent: 0 <= o0 <= 2 DPD
#pragma unroll
  for (int o = 0; o <= 2 * DPD; o++)
#pragma unroll
    for (int k = 0; k <= DPD; k++) cur[o][k] = (o == o0 && k == k0) ? 1u : 0u;
  for (uint32_t t = t0;; t++) {
    // deletions inside the level, in increasing p so that chains are counted
#pragma unroll
    for (int o = 0; o < 2 * DPD; o++)
#pragma unroll
      for (int k = 1; k <= DPD; k++) cur[o + 1][k - 1] += cur[o][k];
    if (t == n) break;
    const uint32_t c = (uint32_t)(v >> (2 * (n - 1 - t))) & 3u;
    uint32_t nxt[2 * DPD + 1][DPD + 1];
#pragma unroll
    for (int o = 0; o <= 2 * DPD; o++)
#pragma unroll
      for (int k = 0; k <= DPD; k++) nxt[o][k] = 0;
#pragma unroll
    for (int o = 0; o <= 2 * DPD; o++) {
      const int p = (int)t + o - DPD;
      if (p < 0) continue;
      const bool inside = (uint32_t)p < n;
      const uint32_t wl = inside ? (uint32_t)(w >> (2 * (n - 1 - (uint32_t)p))) & 3u : 0u;
      if (c < wl) continue;
      const int cost = c != wl ? 1 : 0;
#pragma unroll
      for (int k = 0; k <= DPD; k++) {
        const uint32_t ways = cur[o][k];
        if (inside) {  // match / substitution: same p - t (constant indices keep the band in registers)
          if (cost == 0) nxt[o][k] += ways;
          else if (k >= 1) nxt[o][k - 1] += ways;
        }
        if (k >= 1 && o >= 1) nxt[o - 1][k - 1] += ways;         // insertion: p stays, t advances
      }
    }
#pragma unroll
    for (int o = 0; o <= 2 * DPD; o++)
#pragma unroll
      for (int k = 0; k <= DPD; k++) cur[o][k] = nxt[o][k];
  }
  uint32_t yields = 0;  // level n, p >= n
#pragma unroll
  for (int o = DPD; o <= 2 * DPD; o++)
#pragma unroll
    for (int k = 0; k <= DPD; k++) yields += cur[o][k];
  return yields;
}

__device__ __forceinline__ uint32_t single_candidate_yields(uint64_t w, uint64_t v, uint32_t n, int d) {
  return candidate_yields_from(w, v, n, 0u, 0u, d);
}

// Queries with one candidate are finished here; those with more are listed for the walk below, so
// that its warps are full of work.
__global__ void __launch_bounds__(128)
    edit_replay_single_kernel(const uint64_t *__restrict__ word_key, uint64_t U, uint32_t n, int d,
                              const uint32_t *__restrict__ cand_row, const uint32_t *__restrict__ cand_v,
                              const uint64_t *__restrict__ cand_key, uint64_t *__restrict__ entries,
                              uint32_t *__restrict__ entry_val, uint64_t cap, unsigned long long *__restrict__ counter,
                              uint32_t *__restrict__ multi_list, unsigned long long *__restrict__ n_multi,
                              uint32_t heavy_from, uint32_t *__restrict__ heavy_list,
                              unsigned long long *__restrict__ n_heavy) {
  uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const bool live = r < U;
  const uint32_t seg = live ? cand_row[r] : 0u, m = live ? cand_row[r + 1] - seg : 0u;
  const uint64_t w = live ? word_key[r] : 0ull;
  // One candidate (most queries): every yield is that word, only their number matters — counted by
  // the path-counting recurrence instead of walking the paths.  The warp reserves its output once.
  const uint32_t lane = threadIdx.x & 31u;
  const uint32_t yields = m == 1 ? single_candidate_yields(w, cand_key[seg], n, d) : 0u;
  uint32_t inc = yields;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
    if ((int)lane >= o) inc += t;
  }
  const uint32_t total = __shfl_sync(0xffffffffu, inc, 31);
  if (total) {
    unsigned long long base = 0;
    if (lane == 31) base = atomicAdd(counter, 2ull * total);
    base = __shfl_sync(0xffffffffu, base, 31) + 2ull * (inc - yields);
    if (yields && base + 2ull * yields <= cap) {
      const uint32_t v = cand_v[seg];
      for (uint32_t y = 0; y < yields; y++, base += 2) {
        entries[base] = (r << 32) | (uint32_t)r;
        entry_val[base] = v;
        entries[base + 1] = ((uint64_t)v << 32) | (uint32_t)r;
        entry_val[base + 1] = (uint32_t)r;
      }
    }
  }
  // two lists: queries with many candidates (heavy_from and more) are walked first, so that the longest
  // walks start when the replay does and not at its end
  const bool is_heavy = m >= 2 && m >= heavy_from, is_multi = m >= 2 && !is_heavy;
  const unsigned many = __ballot_sync(0xffffffffu, is_multi);
  if (many) {
    unsigned long long at = 0;
    const int leader = __ffs(many) - 1;
    if ((int)lane == leader) at = atomicAdd(n_multi, (unsigned long long)__popc(many));
    at = __shfl_sync(0xffffffffu, at, leader);
    if (is_multi) multi_list[at + __popc(many & ((1u << lane) - 1u))] = (uint32_t)r;
  }
  const unsigned heavy = __ballot_sync(0xffffffffu, is_heavy);
  if (heavy) {
    unsigned long long at = 0;
    const int leader = __ffs(heavy) - 1;
    if ((int)lane == leader) at = atomicAdd(n_heavy, (unsigned long long)__popc(heavy));
    at = __shfl_sync(0xffffffffu, at, leader);
    if (is_heavy) heavy_list[at + __popc(heavy & ((1u << lane) - 1u))] = (uint32_t)r;
  }
}

// Two exact shortcuts of the recursion, shared by both replay kernels.
//
// frame_alive: can the frame (t letters of the stored word consumed, query position p, budget k) still
// reach a yield?  Every stored word has the query's length n.  A substitution / match needs p < n and the
// yield needs p >= n at t = n.  So a frame behind the diagonal (p < t) takes t - p deletions, and before the
// leaf the n - t letters still to come take an insertion each except for the min(n - t, n - p) that a
// substitution / match can cover — every one of these moves costs one unit of budget.
// (Checked exhaustively against the recursion for short words: every frame with a yield below it is alive.)
__device__ __forceinline__ bool frame_alive(uint32_t n, uint32_t t, uint32_t p, int k) {
  if (p < t) return k >= (int)(t - p);
  if (t < n) {
    const uint32_t rem = n - t, subs = p < n ? min(rem, n - p) : 0u;
    return k >= (int)(rem - subs);
  }
  return true;
}
// A live frame without budget (so p == t before the leaf) can only match letter by letter: it yields the one
// word of its range [lo, hi) whose remaining letters are the query's, once — or nothing.  Returns that
// candidate's index or NONE32.  (The words of the range share their first t letters.)
__device__ __forceinline__ uint32_t zero_budget_match(const uint64_t *__restrict__ keys, uint32_t lo, uint32_t hi,
                                                      uint64_t w, uint32_t n, uint32_t t) {
  if (t >= n) return lo;  // a leaf: the range is one word, the query is used up (frame_alive)
  const uint32_t end = hi;
  const uint64_t low = (1ull << (2 * (n - t))) - 1ull;
  const uint64_t target = (keys[lo] & ~low) | (w & low);
  while (lo < hi) {
    const uint32_t mid = (lo + hi) >> 1;
    if (keys[mid] < target) lo = mid + 1; else hi = mid;
  }
  return lo < end && keys[lo] == target ? lo : NONE32;
}

// The recursion of one query, walked frame by frame with an explicit stack.
__device__ __forceinline__ void edit_walk_query(const uint64_t r, const uint64_t *__restrict__ word_key, uint32_t n, int d,
                                                const uint32_t *__restrict__ cand_row, const uint32_t *__restrict__ cand_v,
                                                const uint64_t *__restrict__ cand_key, uint64_t *__restrict__ entries,
                                                uint32_t *__restrict__ entry_val, uint64_t cap,
                                                unsigned long long *__restrict__ counter) {
  const uint32_t seg = cand_row[r], m = cand_row[r + 1] - seg;
  const uint64_t w = word_key[r];
  const uint64_t *keys = cand_key + seg;
  const uint32_t *ranks = cand_v + seg;
  Frame stk[EDIT_STACK];
  int sp = 0;
  auto yield = [&](uint32_t v) {
    const unsigned long long p = reserve_slots(counter, 2u);
    if (p + 2 <= cap) {
      entries[p] = (r << 32) | (uint32_t)r;       // own yield, in yield order
      entry_val[p] = v;
      entries[p + 1] = ((uint64_t)v << 32) | (uint32_t)r;  // r.leaf->neighbours.push_back(walk leaf)
      entry_val[p + 1] = (uint32_t)r;
    }
  };
  stk[sp++] = Frame{0u, m, 0, 0, (int8_t)d, 0};
  while (sp > 0) {
    Frame &f = stk[sp - 1];
    const uint32_t st = f.stage++;
    if (st == 0) {  // deletion: same node, next query position (also past the end, :202-203)
      if (f.dist >= 1 && frame_alive(n, f.depth, f.pos + 1u, f.dist - 1)) {
        if (f.dist == 1) {  // no budget left below: at most one word, once
          const uint32_t ci = zero_budget_match(keys, f.lo, f.hi, w, n, f.depth);
          if (ci != NONE32) yield(ranks[ci]);
        } else {
          stk[sp] = Frame{f.lo, f.hi, f.depth, (uint8_t)(f.pos + 1), (int8_t)(f.dist - 1), 0};
          sp++;
        }
      }
    } else if (st <= 8) {
      const uint32_t c = (st - 1) >> 1;
      const bool insertion = ((st - 1) & 1u) != 0;
      if (f.depth >= n) { f.stage = 9; continue; }  // a leaf has no children
      const uint32_t wl = f.pos < n ? (uint32_t)(w >> (2 * (n - 1 - f.pos))) & 3u : 0u;
      if (c < wl) { f.stage = (uint8_t)(1 + 2 * wl); continue; }  // children start at word[position]
      if (!insertion && f.pos >= n) continue;                      // substitution needs a query letter
      const int nd = insertion ? f.dist - 1 : f.dist - (int)(c != wl);
      if (nd < 0) continue;
      const uint32_t npos = insertion ? f.pos : f.pos + 1u;
      if (!frame_alive(n, f.depth + 1u, npos, nd)) continue;
      const uint32_t sh = 2 * (n - 1 - f.depth);
      const uint32_t clo = lower_letter(keys, f.lo, f.hi, sh, c);
      const uint32_t chi = lower_letter(keys, clo, f.hi, sh, c + 1);
      if (clo >= chi) { if (!insertion) f.stage++; continue; }  // no such child: skip its insertion too
      if (nd == 0) {
        const uint32_t ci = zero_budget_match(keys, clo, chi, w, n, f.depth + 1u);
        if (ci != NONE32) yield(ranks[ci]);
        continue;
      }
      stk[sp] = Frame{clo, chi, (uint8_t)(f.depth + 1), (uint8_t)npos, (int8_t)nd, 0};
      sp++;
    } else if (st == 9) {
      if (f.pos >= n && f.depth == n && f.lo < f.hi) yield(ranks[f.lo]);  // yield node->leaf
    } else {
      sp--;
    }
  }
}

__global__ void __launch_bounds__(128)
    edit_replay_kernel(const uint64_t *__restrict__ word_key, const uint32_t *__restrict__ multi_list, uint64_t n_multi,
                       uint32_t n, int d, const uint32_t *__restrict__ cand_row, const uint32_t *__restrict__ cand_v,
                       const uint64_t *__restrict__ cand_key, uint64_t *__restrict__ entries,
                       uint32_t *__restrict__ entry_val, uint64_t cap, unsigned long long *__restrict__ counter) {
  const uint64_t gid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid < n_multi) edit_walk_query(multi_list[gid], word_key, n, d, cand_row, cand_v, cand_key, entries, entry_val, cap, counter);
}

// Queries with two or more candidates, in two converged phases instead of one divergent walk.
// Phase A walks the recursion only through the nodes of the candidate trie that still hold two or more
// words; a child holding ONE word v is not entered: all yields below it are v, they are consecutive in
// the depth-first order, and their number is the path count from that frame — so the frame is noted as a
// task (candidate, depth, query position, budget), in order.  Phase B runs the recurrence for task j of
// every lane side by side, then the warp reserves its output with one atomic and every lane writes
// its runs.  A query with more tasks than the list holds (or more than 1023 candidates) is left to the
// walking kernel above; queries with many candidates are walked by the first blocks of this one.
constexpr int EDIT_TASKS = 96;
constexpr uint32_t EDIT_HEAVY = 16;
__global__ void __launch_bounds__(128)
    edit_replay_tasks_kernel(const uint64_t *__restrict__ word_key, const uint32_t *__restrict__ multi_list,
                             uint64_t n_multi, uint32_t n, int d, uint32_t task_cap, const uint32_t *__restrict__ cand_row,
                             const uint32_t *__restrict__ cand_v, const uint64_t *__restrict__ cand_key,
                             uint64_t *__restrict__ entries, uint32_t *__restrict__ entry_val, uint64_t cap,
                             unsigned long long *__restrict__ counter, uint32_t *__restrict__ walk_list,
                             unsigned long long *__restrict__ n_walk, const uint32_t *__restrict__ heavy_list,
                             uint32_t n_heavy) {
  // the first blocks walk the queries with many candidates (block-uniform branch)
  const uint32_t heavy_blocks = (n_heavy + blockDim.x - 1) / blockDim.x;
  if (blockIdx.x < heavy_blocks) {
    const uint32_t h = blockIdx.x * blockDim.x + threadIdx.x;
    if (h < n_heavy) edit_walk_query(heavy_list[h], word_key, n, d, cand_row, cand_v, cand_key, entries, entry_val, cap, counter);
    return;
  }
  const uint64_t gid = (uint64_t)(blockIdx.x - heavy_blocks) * blockDim.x + threadIdx.x;
  const bool live = gid < n_multi;
  const uint32_t lane = threadIdx.x & 31u;
  const uint64_t r = live ? multi_list[gid] : 0ull;
  const uint32_t seg = live ? cand_row[r] : 0u, m = live ? cand_row[r + 1] - seg : 0u;
  const uint64_t w = live ? word_key[r] : 0ull;
  const uint64_t *keys = cand_key + seg;
  const uint32_t *ranks = cand_v + seg;
  uint32_t tasks[EDIT_TASKS];   // candidate << 14 | depth << 8 | position << 2 | budget
  uint32_t counts[EDIT_TASKS];
  uint32_t nt = 0;
  bool over = live && m > 1023u;
  if (live && !over) {
    Frame stk[EDIT_STACK];
    int sp = 0;
    stk[sp++] = Frame{0u, m, 0, 0, (int8_t)d, 0};
    while (sp > 0) {
      Frame &f = stk[sp - 1];
      const uint32_t st = f.stage++;
      if (st == 0) {  // deletion
        if (f.dist >= 1 && frame_alive(n, f.depth, f.pos + 1u, f.dist - 1)) {
          if (f.dist == 1) {  // no budget left below: at most one word, once
            const uint32_t ci = zero_budget_match(keys, f.lo, f.hi, w, n, f.depth);
            if (ci != NONE32) {
              if (nt == task_cap) { over = true; break; }
              tasks[nt++] = (ci << 14) | ((uint32_t)f.depth << 8) | ((f.pos + 1u) << 2);
            }
          } else {
            stk[sp] = Frame{f.lo, f.hi, f.depth, (uint8_t)(f.pos + 1), (int8_t)(f.dist - 1), 0};
            sp++;
          }
        }
      } else if (st <= 8) {
        const uint32_t c = (st - 1) >> 1;
        const bool insertion = ((st - 1) & 1u) != 0;
        if (f.depth >= n) { f.stage = 9; continue; }
        const uint32_t wl = f.pos < n ? (uint32_t)(w >> (2 * (n - 1 - f.pos))) & 3u : 0u;
        if (c < wl) { f.stage = (uint8_t)(1 + 2 * wl); continue; }
        if (!insertion && f.pos >= n) continue;
        const int nd = insertion ? f.dist - 1 : f.dist - (int)(c != wl);
        if (nd < 0) continue;
        const uint32_t npos = insertion ? f.pos : f.pos + 1u;
        if (!frame_alive(n, f.depth + 1u, npos, nd)) continue;
        const uint32_t sh = 2 * (n - 1 - f.depth);
        const uint32_t clo = lower_letter(keys, f.lo, f.hi, sh, c);
        const uint32_t chi = lower_letter(keys, clo, f.hi, sh, c + 1);
        if (clo >= chi) { if (!insertion) f.stage++; continue; }
        if (nd == 0) {
          const uint32_t ci = zero_budget_match(keys, clo, chi, w, n, f.depth + 1u);
          if (ci != NONE32) {
            if (nt == task_cap) { over = true; break; }
            tasks[nt++] = (ci << 14) | ((uint32_t)(f.depth + 1) << 8) | (npos << 2);
          }
        } else if (chi - clo == 1u) {  // a single word below: count instead of walking
          if (nt == task_cap) { over = true; break; }
          tasks[nt++] = (clo << 14) | ((uint32_t)(f.depth + 1) << 8) | (npos << 2) | (uint32_t)nd;
        } else {
          stk[sp] = Frame{clo, chi, (uint8_t)(f.depth + 1), (uint8_t)npos, (int8_t)nd, 0};
          sp++;
        }
      } else {
        sp--;  // (a node holding two words is never a leaf: the candidates are distinct)
      }
    }
  }
  {  // queries handed to the walking kernel
    const unsigned ov = __ballot_sync(0xffffffffu, over);
    if (ov) {
      unsigned long long at = 0;
      const int leader = __ffs(ov) - 1;
      if ((int)lane == leader) at = atomicAdd(n_walk, (unsigned long long)__popc(ov));
      at = __shfl_sync(0xffffffffu, at, leader);
      if (over) walk_list[at + __popc(ov & ((1u << lane) - 1u))] = (uint32_t)r;
    }
    if (over) nt = 0;
  }
  uint32_t total = 0;
  for (uint32_t j = 0; j < nt; j++) {
    const uint32_t tk = tasks[j];
    // a task without budget was matched when it was noted: one yield
    const uint32_t c = (tk & 3u) == 0u ? 1u
                                       : candidate_yields_from(w, keys[tk >> 14], n, (tk >> 8) & 63u, (tk >> 2) & 63u,
                                                               (int)(tk & 3u));
    counts[j] = c;
    total += c;
  }
  uint32_t inc = total;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
    if ((int)lane >= o) inc += t;
  }
  const uint32_t warp_total = __shfl_sync(0xffffffffu, inc, 31);
  if (warp_total == 0) return;
  unsigned long long base = 0;
  if (lane == 31) base = atomicAdd(counter, 2ull * warp_total);
  base = __shfl_sync(0xffffffffu, base, 31) + 2ull * (inc - total);
  if (total && base + 2ull * total <= cap) {
    for (uint32_t j = 0; j < nt; j++) {
      const uint32_t v = ranks[tasks[j] >> 14];
      for (uint32_t y = 0; y < counts[j]; y++, base += 2) {
        entries[base] = (r << 32) | (uint32_t)r;  // own yield, in yield order
        entry_val[base] = v;
        entries[base + 1] = ((uint64_t)v << 32) | (uint32_t)r;  // r.leaf->neighbours.push_back(walk leaf)
        entry_val[base + 1] = (uint32_t)r;
      }
    }
  }
}

// ---- d == 2: the shift candidates by a join on deletion variants --------------------------------
// v is one deletion + one insertion away from w exactly when the two words share a word of n - 1 letters
// obtained by deleting one letter from each.  So every word lists its distinct deletion variants (one per
// run of equal letters: deleting any letter of a run gives the same variant), the list is sorted by a
// 32-bit hash of the variant, and words meeting in a group — with truly equal variants — are candidates
// of each other.  No look-ups: ~n entries per word are written, sorted in four radix passes and read once,
// against 3 (n-1)(n-2)/2 probes per word through L2 (23 ms -> see DESIGN.md).  A group also holds the pairs
// that differ by one substitution; they are candidates as well (the block join lists them too, and the
// list is deduplicated afterwards).
// entry = (position of the deleted letter << 27 | rank) << 32 | hash of the variant
constexpr uint32_t DELJOIN_MAX_WORDS = 1u << 27;

__device__ __forceinline__ uint64_t delete_letter(uint64_t w, uint32_t n, uint32_t p) {
  const uint32_t sh = 2 * (n - 1 - p);
  return ((w >> (sh + 2)) << sh) | (w & ((1ull << sh) - 1ull));
}
__device__ __forceinline__ bool run_ends_at(uint64_t w, uint32_t n, uint32_t p) {
  return p + 1 == n || (((w >> (2 * (n - 1 - p))) ^ (w >> (2 * (n - 2 - p)))) & 3ull) != 0ull;
}

__global__ void __launch_bounds__(256)
    deletion_counts_kernel(const uint64_t *__restrict__ word_key, uint64_t U, uint32_t n, uint32_t *__restrict__ count) {
  const uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= U) return;
  const uint64_t w = word_key[r];
  uint32_t c = 0;
  for (uint32_t p = 0; p < n; p++) c += run_ends_at(w, n, p) ? 1u : 0u;
  count[r] = c;
}

__global__ void __launch_bounds__(256)
    deletion_variants_kernel(const uint64_t *__restrict__ word_key, uint64_t U, uint32_t n,
                             const uint32_t *__restrict__ offset, uint64_t *__restrict__ out) {
  const uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= U) return;
  const uint64_t w = word_key[r];
  uint64_t o = offset[r];
  for (uint32_t p = 0; p < n; p++)
    if (run_ends_at(w, n, p))
      out[o++] = ((uint64_t)((p << 27) | (uint32_t)r) << 32) | (mix64(delete_letter(w, n, p)) >> 32);
}

// one thread per sorted entry: forward through the entries with the same hash
__global__ void __launch_bounds__(256)
    deletion_join_kernel(const uint64_t *__restrict__ sorted, uint64_t D, const uint64_t *__restrict__ word_key, uint32_t n,
                         uint64_t *__restrict__ cand, uint64_t cap, unsigned long long *__restrict__ counter) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i + 1 >= D) return;
  const uint64_t e = ld_stream_u64(sorted + i);
  const uint32_t h = (uint32_t)e;
  uint64_t f = sorted[i + 1];
  if ((uint32_t)f != h) return;  // almost every entry: alone with its hash
  const uint32_t ra = (uint32_t)(e >> 32) & (DELJOIN_MAX_WORDS - 1u);
  const uint64_t va = delete_letter(word_key[ra], n, (uint32_t)(e >> 59));
  for (uint64_t j = i + 1;;) {
    const uint32_t rb = (uint32_t)(f >> 32) & (DELJOIN_MAX_WORDS - 1u);
    if (rb != ra && delete_letter(word_key[rb], n, (uint32_t)(f >> 59)) == va) emit_candidate(cand, cap, counter, ra, rb);
    if (++j >= D) break;
    f = sorted[j];
    if ((uint32_t)f != h) break;
  }
}

int stage_neighbours_edit(State &s) {
  const uint64_t U = s.unique;
  const uint32_t n = s.cfg.word_length, d = s.cfg.distance;
  s.n_entries = 0;
  if (U == 0) return 0;
  if (d > EDIT_MAX_D) throw CudaError{"edit distance > 3 is not supported (two deletion+insertion pairs)"};
  if (n + d + 2 > EDIT_STACK) throw CudaError{"word too long for the edit replay stack"};
  // the shift candidates (b) come from a join on deletion variants at d == 2, from variant probing otherwise
  // (or when the entries would not fit their 27 + 5 bit payload / 2^32 entries)
  bool by_join = d == 2 && U < DELJOIN_MAX_WORDS && (uint64_t)U * n < 0xFFFFFFFFull && n >= 2;
  if (const char *e = std::getenv("HUMID_GPU_EDIT_SHIFTS"))  // testing knob: "probe" or "join"
    by_join = by_join && e[0] == 'j';
  int k = 0;
  Lookup lk{};
  if (!by_join) {  // key -> rank look-ups: table + fingerprint filter + a first-level bitmap of 16 to 32 bits per word
    k += stage_ensure_lookup(s);
    lk = s.lookup();
    int log2_bits = 20;
    while (log2_bits < 33 && (1ull << log2_bits) < 16 * U) log2_bits++;
    const uint64_t words32 = (1ull << log2_bits) / 32;
    s.comp.ensure(words32);  // free until the clustering stage
    Timed t(s, "edit_bitmap", U, U * 12 + words32 * 4);
    HUMID_CUDA(cudaMemsetAsync(s.comp.p, 0, words32 * sizeof(uint32_t), s.stream));
    bitmap_build_kernel<<<grid_for(U, 256), 256, 0, s.stream>>>(s.word_key.p, U, s.comp.p, 64 - log2_bits);
    lk.bitmap = s.comp.p;
    lk.bitmap_shift = 64 - log2_bits;
    k++;
  }

  // phase 1: candidate pairs (with duplicates)
  uint64_t cap = std::max<uint64_t>(s.entries.cap, std::max<uint64_t>(8 * U, 1u << 20));
  uint64_t n_cand = 0;
  for (int attempt = 0; attempt < 4; attempt++) {
    s.entries.ensure(cap);
    cap = s.entries.cap;
    unsigned long long *counter = s.counters.p + 5;
    HUMID_CUDA(cudaMemsetAsync(counter, 0, sizeof(unsigned long long), s.stream));
    // (a) substitutions only = every pair within Hamming distance d: block join without the
    //     dominance rule (neighbours.cu), both directions of a pair become candidates
    {
      const JoinPlan P = make_join_plan(n, d);  // d + 1 blocks, or pairs of four blocks at d == 2
      s.key_alt.ensure(U + 1);
      s.entries_alt.ensure(U + 1);
      s.entry_val.ensure(U + 1);
      s.entry_val_alt.ensure(U + 1);
      for (uint32_t b = 0; b < P.B.n_blocks; b++) {
        const uint64_t *keys;
        const uint32_t *rank;
        k += join_plan_order(s, P, b, U, keys, rank);
        Timed t(s, "edit_candidates_join", U, U * (8 + 8 + (rank ? 4 : 0)));
        launch_hamming_join(s, JOIN_CANDIDATES, keys, rank, U, P.B, b, JOIN_NO_CAP, s.entries.p, cap / 2, counter, nullptr,
                            0u, nullptr, nullptr);
        k++;
      }
      uint64_t pairs = 0;
      HUMID_CUDA(cudaMemcpyAsync(&pairs, counter, sizeof(uint64_t), cudaMemcpyDeviceToHost, s.stream));
      HUMID_CUDA(cudaStreamSynchronize(s.stream));
      if (2 * pairs > cap) {  // the counter kept counting: come back with room for these and as many again
        cap = 4 * pairs;
        n_cand = cap + 1;
        continue;
      }
      pairs *= 2;  // from here on the cursor counts candidates
      HUMID_CUDA(cudaMemcpyAsync(counter, &pairs, sizeof(uint64_t), cudaMemcpyHostToDevice, s.stream));
      HUMID_CUDA(cudaStreamSynchronize(s.stream));
    }
    // (b) one deletion + one insertion (a one-letter block shift) and up to d-2 substitutions: join on deletion
    //     variants, or variants probed against the word table, 3 (n-1)(n-2)/2 (x (1 + 3 n) at d = 3) per word
    if (by_join) {
      s.seed_flag.ensure(U + 1);   // variants per word
      s.seed_index.ensure(U + 1);  // ... and where they go
      uint32_t *total = reinterpret_cast<uint32_t *>(s.counters.p + 46);
      HUMID_CUDA(cudaMemsetAsync(s.counters.p + 46, 0, sizeof(unsigned long long), s.stream));
      size_t t_gen = s.timer_begin("edit_deletion_variants", U, U * (8 + 4 + 4) + U * (uint64_t)n * 8);
      deletion_counts_kernel<<<grid_for(U, 256), 256, 0, s.stream>>>(s.word_key.p, U, n, s.seed_flag.p);
      int launches = 1 + exclusive_scan_u32(s.seed_flag.p, s.seed_index.p, U, total, s.scan_scratch, s.stream);
      uint64_t D = 0;
      HUMID_CUDA(cudaMemcpyAsync(&D, s.counters.p + 46, sizeof(uint64_t), cudaMemcpyDeviceToHost, s.stream));
      HUMID_CUDA(cudaStreamSynchronize(s.stream));
      s.join_entries.ensure(D + 1);
      s.join_entries_out.ensure(D + 1);
      deletion_variants_kernel<<<grid_for(U, 256), 256, 0, s.stream>>>(s.word_key.p, U, n, s.seed_index.p, s.join_entries.p);
      s.timer_end(t_gen, launches + 1);
      k += launches + 1;
      uint64_t *kb[2] = {s.join_entries.p, s.join_entries_out.p};
      uint32_t *vb[2] = {nullptr, nullptr};
      int cur;
      {
        Timed t(s, "edit_deletion_sort", D, D * (8 + 4 * 16));
        launches = 0;
        cur = radix_sort_u64(kb, vb, D, 32, s.scratch, s.scan_scratch, s.stream, &launches);
        t.launches = launches;
        k += launches;
      }
      Timed t(s, "edit_deletion_join", D, D * 8);
      if (D > 1)
        deletion_join_kernel<<<grid_for(D, 256), 256, 0, s.stream>>>(kb[cur], D, s.word_key.p, n, s.entries.p, cap, counter);
    } else {
      double shifts = 1.5 * (n - 1) * (n - 2) * (d >= 3 ? 1.0 + 3.0 * n : 1.0);
      Timed t(s, "edit_candidates", U, (uint64_t)(U * 8.0 * (1.0 + shifts)));
      const uint32_t n_pairs = n >= 3 ? (n - 1) * (n - 2) / 2 : 0u;  // (i, q) with i - q >= 2
      const uint32_t inv_pairs = n_pairs ? (uint32_t)((1ull << 32) / n_pairs + 1) : 0u;
      if (n_pairs && d == 2 && lk.bitmap)
        edit_candidates_shift_kernel<true><<<grid_for(U, SHIFT_WPB), SHIFT_THREADS, 0, s.stream>>>(
            s.word_key.p, U, n, d, lk, s.entries.p, cap, counter, n_pairs, inv_pairs);
      else if (n_pairs)
        edit_candidates_shift_kernel<false><<<grid_for(U, SHIFT_WPB), SHIFT_THREADS, 0, s.stream>>>(
            s.word_key.p, U, n, d, lk, s.entries.p, cap, counter, n_pairs, inv_pairs);
    }
    k++;
    HUMID_CUDA(cudaMemcpyAsync(&n_cand, s.counters.p + 5, sizeof(uint64_t), cudaMemcpyDeviceToHost, s.stream));
    HUMID_CUDA(cudaStreamSynchronize(s.stream));
    if (n_cand <= cap) break;
    cap = n_cand;
  }
  if (n_cand > cap) throw CudaError{"edit-distance candidates: the list kept outgrowing its buffer"};
  if (n_cand >= 0xFFFFFFFFull) throw CudaError{"too many edit-distance candidate pairs"};

  // sort + unique -> per-query candidate lists in rank (== lexicographic) order
  s.work_a.ensure(U + 2);  // cand_row
  uint64_t n_distinct = 0;
  if (n_cand) {
    size_t t_sort = s.timer_begin("edit_candidates_unique", n_cand, n_cand * 28);
    s.entries_alt.ensure(s.entries.cap);
    uint64_t *kb[2] = {s.entries.p, s.entries_alt.p};
    uint32_t *vb[2] = {nullptr, nullptr};
    int ubits = 1;
    while ((1ull << ubits) < U) ubits++;
    int launches = 0;
    int cur = radix_sort_u64(kb, vb, n_cand, 32 + ubits, s.scratch, s.scan_scratch, s.stream, &launches);
    if (cur == 1) std::swap(s.entries, s.entries_alt);
    s.seed_flag.ensure(n_cand);
    s.seed_index.ensure(n_cand);
    mark_unique_kernel<<<grid_for(n_cand, 256), 256, 0, s.stream>>>(s.entries.p, n_cand, s.seed_flag.p);
    uint32_t *total = reinterpret_cast<uint32_t *>(s.counters.p + 6);
    launches += 1 + exclusive_scan_u32(s.seed_flag.p, s.seed_index.p, n_cand, total, s.scan_scratch, s.stream);
    uint32_t nd32 = 0;
    HUMID_CUDA(cudaMemcpyAsync(&nd32, total, sizeof(uint32_t), cudaMemcpyDeviceToHost, s.stream));
    HUMID_CUDA(cudaStreamSynchronize(s.stream));
    n_distinct = nd32;
    s.work_b.ensure(n_distinct + 1);   // cand_q
    s.work_c.ensure(n_distinct + 1);   // cand_v
    s.key_alt.ensure(n_distinct + 1);  // cand_key
    compact_candidates_kernel<<<grid_for(n_cand, 256), 256, 0, s.stream>>>(
        s.entries.p, s.seed_flag.p, s.seed_index.p, n_cand, s.word_key.p, s.work_b.p, s.work_c.p, s.key_alt.p);
    launches++;
    s.timer_end(t_sort, launches);
    k += launches;
  } else {
    s.work_b.ensure(1);
    s.work_c.ensure(1);
    s.key_alt.ensure(1);
  }
  candidate_rows_kernel<<<grid_for(n_distinct + 1, 256), 256, 0, s.stream>>>(s.work_b.p, n_distinct, U, s.work_a.p);
  k++;

  // phase 2: replay
  uint64_t ecap = std::max<uint64_t>(std::max<uint64_t>(s.entries_alt.cap, s.entry_val.cap),
                                     std::max<uint64_t>(8 * n_distinct, 1u << 20));
  for (int attempt = 0; attempt < 2; attempt++) {
    // entries_alt is free again (the sorted candidates were compacted out of `entries`)
    s.entries_alt.ensure(ecap);
    s.entry_val.ensure(ecap);
    ecap = std::min<uint64_t>(s.entries_alt.cap, s.entry_val.cap);
    HUMID_CUDA(cudaMemsetAsync(s.counters.p + 2, 0, sizeof(unsigned long long), s.stream));
    HUMID_CUDA(cudaMemsetAsync(s.counters.p + 43, 0, 3 * sizeof(unsigned long long), s.stream));  // 43, 44, 45
    s.word_flags.ensure(U + 1);  // list of the queries with two or more candidates
    s.seed_index.ensure(U + 1);  // ... and of those with EDIT_HEAVY or more
    // testing knobs.  HUMID_GPU_EDIT_TASKS: task list length of the counting replay, 0 = walk every query;
    // HUMID_GPU_EDIT_HEAVY: number of candidates from which a query is walked straight away
    uint32_t task_cap = EDIT_TASKS, heavy_from = EDIT_HEAVY;
    if (const char *e = std::getenv("HUMID_GPU_EDIT_TASKS"))
      task_cap = (uint32_t)std::min<long>(std::max<long>(std::strtol(e, nullptr, 10), 0), EDIT_TASKS);
    if (const char *e = std::getenv("HUMID_GPU_EDIT_HEAVY"))
      heavy_from = (uint32_t)std::max<long>(std::strtol(e, nullptr, 10), 2);
    if (task_cap == 0) heavy_from = 0xFFFFFFFFu;  // one list, walked
    uint64_t n_multi = 0, n_heavy = 0;
    {
      Timed t(s, "edit_replay_single", U, U * 16 + n_distinct * 12);
      edit_replay_single_kernel<<<grid_for(U, 128), 128, 0, s.stream>>>(s.word_key.p, U, n, (int)d, s.work_a.p, s.work_c.p,
                                                                        s.key_alt.p, s.entries_alt.p, s.entry_val.p, ecap,
                                                                        s.counters.p + 2, s.word_flags.p,
                                                                        s.counters.p + 43, heavy_from, s.seed_index.p,
                                                                        s.counters.p + 45);
    }
    HUMID_CUDA(cudaMemcpyAsync(&n_multi, s.counters.p + 43, sizeof(uint64_t), cudaMemcpyDeviceToHost, s.stream));
    HUMID_CUDA(cudaMemcpyAsync(&n_heavy, s.counters.p + 45, sizeof(uint64_t), cudaMemcpyDeviceToHost, s.stream));
    HUMID_CUDA(cudaStreamSynchronize(s.stream));
    const uint32_t *walk_list = s.word_flags.p;
    uint64_t n_walk = n_multi;
    if ((n_multi || n_heavy) && task_cap) {
      s.seed_flag.ensure(n_multi + 1);  // the queries left to the walk
      HUMID_CUDA(cudaMemsetAsync(s.counters.p + 44, 0, sizeof(unsigned long long), s.stream));
      {
        Timed t(s, "edit_replay_count", n_multi + n_heavy, n_distinct * 12 + (n_multi + n_heavy) * 16);
        const unsigned blocks = (unsigned)((n_heavy + 127) / 128 + (n_multi + 127) / 128);
        edit_replay_tasks_kernel<<<blocks, 128, 0, s.stream>>>(
            s.word_key.p, s.word_flags.p, n_multi, n, (int)d, task_cap, s.work_a.p, s.work_c.p, s.key_alt.p, s.entries_alt.p,
            s.entry_val.p, ecap, s.counters.p + 2, s.seed_flag.p, s.counters.p + 44, s.seed_index.p, (uint32_t)n_heavy);
      }
      k++;
      HUMID_CUDA(cudaMemcpyAsync(&n_walk, s.counters.p + 44, sizeof(uint64_t), cudaMemcpyDeviceToHost, s.stream));
      HUMID_CUDA(cudaStreamSynchronize(s.stream));
      walk_list = s.seed_flag.p;
    }
    if (n_walk) {
      Timed t(s, "edit_replay", n_walk, n_walk * 16);
      edit_replay_kernel<<<grid_for(n_walk, 128), 128, 0, s.stream>>>(s.word_key.p, walk_list, n_walk, n, (int)d,
                                                                      s.work_a.p, s.work_c.p, s.key_alt.p,
                                                                      s.entries_alt.p, s.entry_val.p, ecap,
                                                                      s.counters.p + 2);
    }
    k += 2;
    HUMID_CUDA(cudaMemcpyAsync(&s.n_entries, s.counters.p + 2, sizeof(uint64_t), cudaMemcpyDeviceToHost, s.stream));
    HUMID_CUDA(cudaStreamSynchronize(s.stream));
    if (s.n_entries <= ecap) break;
    ecap = s.n_entries;
  }
  std::swap(s.entries, s.entries_alt);  // stage_csr sorts `entries` / `entry_val`
  HUMID_CUDA(cudaGetLastError());
  return k;
}

}  // namespace humid
