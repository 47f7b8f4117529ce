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
#include <algorithm>

#include "state.cuh"

namespace humid {

// Undirected pair (a < b in rank: an upward variant is numerically larger) + the two degrees.
__device__ __forceinline__ void emit_pair(uint64_t *pairs, uint64_t cap, unsigned long long *counter, uint32_t *deg,
                                          uint32_t a, uint32_t b) {
  unsigned long long pos = atomicAdd(counter, 1ull);
  if (pos < cap) {
    pairs[pos] = ((uint64_t)a << 32) | b;
    atomicAdd(deg + a, 1u);
    atomicAdd(deg + b, 1u);
  }
}

// d == 1: one thread per word, looping over the positions.  Per position the (up to) three
// filter buckets are fetched back to back (independent L2 requests, evict_last so that the
// filter stays resident), only fingerprint matches go on to the HBM table (evict_first).
// The filter hash is linear, so a variant costs one 64-bit add.  A thread keeps its hits in
// registers; the warp appends all of them with one global atomic.
constexpr int H1_LOCAL_HITS = 6;
constexpr uint32_t H1_SCAN_CAP = 32;
__global__ void __launch_bounds__(128)
    hamming1_probe_kernel(const uint64_t *__restrict__ word_key, uint64_t U, uint32_t n, uint32_t scan_from,
                          const Lookup lk, uint64_t *__restrict__ pairs, uint64_t cap,
                          unsigned long long *__restrict__ counter, uint32_t *__restrict__ deg) {
  const uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const bool live = r < U;
  const uint64_t w = live ? word_key[r] : 0ull;
  const uint64_t hw = filter_hash(w);
  const uint64_t pol_keep = l2_policy_evict_last();
  uint32_t hits[H1_LOCAL_HITS];
  uint32_t nh = 0;
  const uint32_t lane = lane_id();
  // Variants at positions >= scan_from share the word's first scan_from letters, so in rank
  // order they sit right behind it: walk forward while the prefix matches (rank == index, no
  // look-up at all).  Only if that run is unusually long are those positions probed instead.
  uint32_t probe_end = n;
  if (live && scan_from < n) {
    const uint32_t low = 2 * (n - scan_from);
    const uint64_t prefix = w >> low;
    probe_end = scan_from;
    for (uint64_t j = r + 1; j < U; j++) {
      const uint64_t k = word_key[j];
      if ((k >> low) != prefix) break;
      if (j - r > H1_SCAN_CAP) { probe_end = n; break; }
      const uint64_t x = k ^ w;
      if (__popcll((x | (x >> 1)) & 0x5555555555555555ull) == 1) {  // one letter differs (and k > w)
        if (nh == H1_LOCAL_HITS) { probe_end = n; break; }  // nothing emitted yet: safe to start over
        hits[nh++] = (uint32_t)j;
      }
    }
    if (probe_end == n) nh = 0;  // run too long: everything is found by probing below instead
  }
  for (uint32_t p = 0; p < probe_end; p++) {
    const uint32_t sh = 2 * (n - 1 - p);
    const uint32_t wl = (uint32_t)(w >> sh) & 3u;
    const uint64_t step = FILTER_MUL << sh;  // hash of "+1 at position p"
    uint4 bkt[3];
    uint64_t hf[3];
#pragma unroll
    for (uint32_t c = 1; c < 4; c++) {
      hf[c - 1] = hw + step * c;  // variants with letter wl + c (valid while wl + c <= 3)
      if (live && wl + c <= 3u) bkt[c - 1] = ld_hint_u128(lk.filter + filter_bucket(hf[c - 1], lk.n_buckets), pol_keep);
    }
#pragma unroll
    for (uint32_t c = 1; c < 4; c++) {
      if (live && wl + c <= 3u && filter_match(bkt[c - 1], hf[c - 1])) {
        const uint32_t hit = table_rank(lk, w + ((uint64_t)c << sh));
        if (hit != NONE32) {
          if (nh == H1_LOCAL_HITS) {  // rare: local buffer full, spill what we have
            unsigned long long pos = atomicAdd(counter, (unsigned long long)nh);
            for (uint32_t k = 0; k < nh; k++, pos++)
              if (pos < cap) { pairs[pos] = (r << 32) | hits[k]; atomicAdd(deg + hits[k], 1u); }
            if (pos <= cap) atomicAdd(deg + r, nh);
            nh = 0;
          }
          hits[nh++] = hit;
        }
      }
    }
  }
  uint32_t inc = nh;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
    if ((int)lane >= o) inc += t;
  }
  const uint32_t total = __shfl_sync(0xffffffffu, inc, 31);
  if (total == 0) return;
  unsigned long long base = 0;
  if (lane == 31) base = atomicAdd(counter, (unsigned long long)total);
  base = __shfl_sync(0xffffffffu, base, 31);
  if (nh) {
    unsigned long long pos = base + (inc - nh);
    if (pos + nh <= cap) {
      for (uint32_t k = 0; k < nh; k++) {
        pairs[pos + k] = (r << 32) | hits[k];
        atomicAdd(deg + hits[k], 1u);
      }
      atomicAdd(deg + r, nh);
    }
  }
}

// d == 2: one thread per (word, first modified position p1).
__global__ void __launch_bounds__(256)
    hamming2_probe_kernel(const uint64_t *__restrict__ word_key, uint64_t U, uint32_t n, const Lookup lk,
                          uint64_t *__restrict__ pairs, uint64_t cap, unsigned long long *__restrict__ counter,
                          uint32_t *__restrict__ deg) {
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
    if (hit != NONE32) emit_pair(pairs, cap, counter, deg, (uint32_t)r, hit);
    for (uint32_t p2 = p1 + 1; p2 < n; p2++) {
      uint32_t sh2 = 2 * (n - 1 - p2);
      uint32_t w2 = (uint32_t)(w >> sh2) & 3u;
      for (uint32_t c2 = w2 + 1; c2 < 4; c2++) {
        uint32_t h2 = lookup_rank(lk, v1 + ((uint64_t)(c2 - w2) << sh2));
        if (h2 != NONE32) emit_pair(pairs, cap, counter, deg, (uint32_t)r, h2);
      }
    }
  }
}

// any d >= 1: one thread per (word, first modified position); iterative enumeration of the
// upward variants with at most d modified positions, positions strictly increasing.
constexpr int HAM_MAX_D = 8;
__global__ void __launch_bounds__(128)
    hammingd_probe_kernel(const uint64_t *__restrict__ word_key, uint64_t U, uint32_t n, uint32_t d, const Lookup lk,
                          uint64_t *__restrict__ pairs, uint64_t cap, unsigned long long *__restrict__ counter,
                          uint32_t *__restrict__ deg) {
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
    if (hit != NONE32) emit_pair(pairs, cap, counter, deg, (uint32_t)r, hit);
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
  s.n_entries = s.n_pairs = 0;
  s.deg.ensure(U + 1);
  if (U == 0 || d == 0) {
    if (U) HUMID_CUDA(cudaMemsetAsync(s.deg.p, 0, U * sizeof(uint32_t), s.stream));
    return 0;
  }
  if (d > HAM_MAX_D) throw CudaError{"Hamming distance > 8 is not supported by the variant-probe path"};
  int k = 0;
  uint64_t cap = s.entries.cap ? s.entries.cap : std::max<uint64_t>(2 * U, 1u << 20);
  const Lookup lk = s.lookup();
  for (int attempt = 0; attempt < 2; attempt++) {
    s.entries.ensure(cap);
    cap = s.entries.cap;
    HUMID_CUDA(cudaMemsetAsync(s.counters.p + 2, 0, sizeof(unsigned long long), s.stream));
    HUMID_CUDA(cudaMemsetAsync(s.deg.p, 0, U * sizeof(uint32_t), s.stream));
    uint64_t threads = U * n;
    // 8 B per probed upward variant (expected 1.5 n at d=1, + 2.25 C(n,2) at d=2) + the word
    double probes = d == 1 ? 1.5 * n : d == 2 ? 1.5 * n + 2.25 * n * (n - 1) / 2.0 : 0.0;
    {
      Timed t(s, d == 1 ? "probe_hamming1" : d == 2 ? "probe_hamming2" : "probe_hammingd", U,
              (uint64_t)(U * 8.0 * (1.0 + probes)));
      if (d == 1 && lk.filter)
        hamming1_probe_kernel<<<grid_for(U, 128), 128, 0, s.stream>>>(s.word_key.p, U, n, scan_from_position(U, n), lk,
                                                                      s.entries.p, cap, s.counters.p + 2, s.deg.p);
      else if (d == 2)
        hamming2_probe_kernel<<<grid_for(threads, 256), 256, 0, s.stream>>>(s.word_key.p, U, n, lk, s.entries.p, cap,
                                                                            s.counters.p + 2, s.deg.p);
      else
        hammingd_probe_kernel<<<grid_for(threads, 128), 128, 0, s.stream>>>(s.word_key.p, U, n, d, lk, s.entries.p,
                                                                            cap, s.counters.p + 2, s.deg.p);
    }
    k++;
    HUMID_CUDA(cudaMemcpyAsync(&s.n_pairs, s.counters.p + 2, sizeof(uint64_t), cudaMemcpyDeviceToHost, s.stream));
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
  return csr_from_degrees(s, s.n_entries, true);
}

int stage_csr_directed(State &s) { return csr_from_degrees(s, s.n_entries, false); }

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
