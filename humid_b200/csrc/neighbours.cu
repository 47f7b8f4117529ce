// K3 — Hamming neighbour search by variant enumeration + hash probe, and CSR construction.
//
// Replaces Trie::asymmetricHamming -> hamming_<4,NLeaf,false>
// (lib/trie/src/trieFunctions.tcc:156-181) as called from findHammingNeighbours
// (src/humid.cc:113-130).  The asymmetric search of query w yields every stored v with
// v[p] >= w[p] at EVERY position and ham(w,v) <= d (component-wise dominance, not
// lexicographic order); the caller links both directions.  So each unordered linked pair is
// found exactly once, from its dominated member, by enumerating only UPWARD variants
// (letters greater than w[p]).  list(L) ends up as all linked words in ascending rank
// (dominated ones are pushed while L is the result of earlier queries, dominating ones by
// L's own query, each group ascending), without duplicates.
#include "state.cuh"

namespace humid {

__device__ __forceinline__ void emit_pair(uint64_t *entries, uint64_t cap, unsigned long long *counter, uint32_t a,
                                          uint32_t b) {
  unsigned long long pos = atomicAdd(counter, 2ull);
  if (pos + 2 <= cap) {
    entries[pos] = ((uint64_t)a << 32) | b;
    entries[pos + 1] = ((uint64_t)b << 32) | a;
  }
}

// d == 1: one thread per (word, position); all lanes run the same 3 iterations so hits can
// be appended with one atomic per warp.
__global__ void __launch_bounds__(256)
    hamming1_probe_kernel(const uint64_t *__restrict__ word_key, uint64_t U, uint32_t n,
                          const Lookup lk, uint64_t *__restrict__ entries,
                          uint64_t cap, unsigned long long *__restrict__ counter) {
  uint64_t gid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  uint64_t r = gid / n;
  uint32_t p = (uint32_t)(gid - r * n);
  bool live = r < U;
  uint64_t w = live ? word_key[r] : 0ull;
  uint32_t sh = 2 * (n - 1 - p);
  uint32_t wl = (uint32_t)(w >> sh) & 3u;
  const uint32_t lane = lane_id();
#pragma unroll
  for (uint32_t c = 1; c < 4; c++) {
    uint32_t hit = NONE32;
    if (live && c > wl) hit = lookup_rank(lk, w + ((uint64_t)(c - wl) << sh));
    unsigned m = __ballot_sync(0xffffffffu, hit != NONE32);
    if (m) {
      unsigned long long base = 0;
      int leader = __ffs(m) - 1;
      if ((int)lane == leader) base = atomicAdd(counter, 2ull * __popc(m));
      base = __shfl_sync(0xffffffffu, base, leader);
      if (hit != NONE32) {
        unsigned long long pos = base + 2ull * __popc(m & ((1u << lane) - 1u));
        if (pos + 2 <= cap) {
          entries[pos] = (r << 32) | hit;
          entries[pos + 1] = ((uint64_t)hit << 32) | (uint32_t)r;
        }
      }
    }
  }
}

// d == 2: one thread per (word, first modified position p1).
__global__ void __launch_bounds__(256)
    hamming2_probe_kernel(const uint64_t *__restrict__ word_key, uint64_t U, uint32_t n,
                          const Lookup lk, uint64_t *__restrict__ entries,
                          uint64_t cap, unsigned long long *__restrict__ counter) {
  uint64_t gid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  uint64_t r = gid / n;
  if (r >= U) return;
  uint32_t p1 = (uint32_t)(gid - r * n);
  uint64_t w = word_key[r];
  uint32_t sh1 = 2 * (n - 1 - p1);
  uint32_t w1 = (uint32_t)(w >> sh1) & 3u;
  for (uint32_t c1 = w1 + 1; c1 < 4; c1++) {
    uint64_t v1 = w + ((uint64_t)(c1 - w1) << sh1);
    uint32_t hit = lookup_rank(lk, v1);
    if (hit != NONE32) emit_pair(entries, cap, counter, (uint32_t)r, hit);
    for (uint32_t p2 = p1 + 1; p2 < n; p2++) {
      uint32_t sh2 = 2 * (n - 1 - p2);
      uint32_t w2 = (uint32_t)(w >> sh2) & 3u;
      for (uint32_t c2 = w2 + 1; c2 < 4; c2++) {
        uint32_t h2 = lookup_rank(lk, v1 + ((uint64_t)(c2 - w2) << sh2));
        if (h2 != NONE32) emit_pair(entries, cap, counter, (uint32_t)r, h2);
      }
    }
  }
}

// any d >= 1: one thread per (word, first modified position); iterative enumeration of the
// upward variants with at most d modified positions, positions strictly increasing.
constexpr int HAM_MAX_D = 8;
__global__ void __launch_bounds__(128)
    hammingd_probe_kernel(const uint64_t *__restrict__ word_key, uint64_t U, uint32_t n, uint32_t d,
                          const Lookup lk, uint64_t *__restrict__ entries,
                          uint64_t cap, unsigned long long *__restrict__ counter) {
  uint64_t gid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  uint64_t r = gid / n;
  if (r >= U) return;
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
    uint32_t hit = lookup_rank(lk, cur);
    if (hit != NONE32) emit_pair(entries, cap, counter, (uint32_t)r, hit);
    if (depth + 1 < (int)d && pos[depth] + 1 < n) {
      depth++;
      pos[depth] = pos[depth - 1] + 1;
      let[depth] = (uint32_t)(w >> (2 * (n - 1 - pos[depth]))) & 3u;
    }
  }
}

int stage_neighbours_hamming(State &s) {
  const uint64_t U = s.unique;
  const uint32_t n = s.cfg.word_length, d = s.cfg.distance;
  s.n_entries = 0;
  if (U == 0 || d == 0) return 0;
  if (d > HAM_MAX_D) throw CudaError{"Hamming distance > 8 is not supported by the variant-probe path"};
  int k = 0;
  uint64_t cap = s.entries.cap ? s.entries.cap : std::max<uint64_t>(4 * U, 1u << 20);
  for (int attempt = 0; attempt < 2; attempt++) {
    s.entries.ensure(cap);
    cap = s.entries.cap;
    HUMID_CUDA(cudaMemsetAsync(s.counters.p + 2, 0, sizeof(unsigned long long), s.stream));
    uint64_t threads = U * n;
    // 8 B per probed upward variant (expected 1.5 n at d=1, + 2.25 C(n,2) at d=2) + the word
    double probes = d == 1 ? 1.5 * n : d == 2 ? 1.5 * n + 2.25 * n * (n - 1) / 2.0 : 0.0;
    Timed t(s, d == 1 ? "probe_hamming1" : d == 2 ? "probe_hamming2" : "probe_hammingd", U,
            (uint64_t)(U * 8.0 * (1.0 + probes)));
    if (d == 1)
      hamming1_probe_kernel<<<grid_for(threads, 256), 256, 0, s.stream>>>(
          s.word_key.p, U, n, s.lookup(), s.entries.p, cap, s.counters.p + 2);
    else if (d == 2)
      hamming2_probe_kernel<<<grid_for(threads, 256), 256, 0, s.stream>>>(
          s.word_key.p, U, n, s.lookup(), s.entries.p, cap, s.counters.p + 2);
    else
      hammingd_probe_kernel<<<grid_for(threads, 128), 128, 0, s.stream>>>(
          s.word_key.p, U, n, d, s.lookup(), s.entries.p, cap, s.counters.p + 2);
    k++;
    HUMID_CUDA(cudaMemcpyAsync(&s.n_entries, s.counters.p + 2, sizeof(uint64_t), cudaMemcpyDeviceToHost, s.stream));
    HUMID_CUDA(cudaStreamSynchronize(s.stream));
    if (s.n_entries <= cap) break;
    cap = s.n_entries;  // the counter kept counting: exact size is now known, run once more
  }
  HUMID_CUDA(cudaGetLastError());
  return k;
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
