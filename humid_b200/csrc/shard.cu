// Multi-GPU phases of the hot path (include/humid_gpu_shard.h).
//
// Words are RANGE-partitioned by key: shard p owns the keys in [splitters[p-1], splitters[p]),
// so its sorted words are a contiguous block of the global walk order
// (lib/trie/src/trieFunctions.tcc:129-143) and global rank = bases[p] + shard-local rank.
// A variant that differs from its word at nucleotide >= 2 almost always stays in the same
// range, so most look-ups of asymmetricHamming (trieFunctions.tcc:156-181) are answered in
// place; the others travel as {variant key, source rank} and come back as rank-or-miss.
#include <algorithm>
#include <vector>

#include "../../include/humid_gpu_shard.h"
#include "abi_internal.cuh"

namespace humid {

struct Layout {
  uint32_t part, n_parts;
  uint64_t splitters[HUMID_MAX_PARTS];  // n_parts-1 used
  uint32_t base;                        // global rank of this shard's first word
  uint64_t hi;                          // first key above this shard's range (upward variants never fall below it)
};

__device__ __forceinline__ uint32_t owner_of(const Layout &L, uint64_t key) {
  uint32_t p = 0;
  for (uint32_t i = 0; i + 1 < L.n_parts; i++) p += (key >= L.splitters[i]) ? 1u : 0u;
  return p;
}

// ---- tuples to their owners ------------------------------------------------------------------
// up to n evenly spaced usable tuple keys (tuples are in random order w.r.t. their keys)
__global__ void sample_tuple_keys_kernel(const uint64_t *__restrict__ keys, uint64_t R, uint32_t n, uint64_t *out,
                                         unsigned long long *counter) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint64_t k = keys[(uint64_t)(((unsigned __int128)i * R + R / 2) / n) % R];
  if (k != KEY_INVALID) out[atomicAdd(counter, 1ull)] = k;
}

// MODE 0: usable tuples per destination part.  MODE 1: scatter {key, global tuple index} into
// the destination's region (cursors = exclusive prefix of the counts).  One global atomic per
// CTA and part; positions inside a part are arbitrary (the owner does not depend on them).
template <int MODE>
__global__ void __launch_bounds__(256)
    partition_tuples_kernel(const uint64_t *__restrict__ keys, uint64_t R, uint32_t tuple_base, const Layout L,
                            unsigned long long *__restrict__ part_counts, uint64_t *__restrict__ out_keys,
                            uint32_t *__restrict__ out_index) {
  __shared__ uint32_t cnt[HUMID_MAX_PARTS];
  __shared__ unsigned long long base[HUMID_MAX_PARTS];
  constexpr int ITEMS = 4;
  if (threadIdx.x < HUMID_MAX_PARTS) cnt[threadIdx.x] = 0;
  __syncthreads();
  const uint64_t t0 = (uint64_t)blockIdx.x * (256 * ITEMS);
  uint64_t key[ITEMS];
  uint32_t own[ITEMS], slot[ITEMS];
#pragma unroll
  for (int j = 0; j < ITEMS; j++) {
    uint64_t t = t0 + (uint64_t)j * 256 + threadIdx.x;
    key[j] = t < R ? ld_stream_u64(keys + t) : KEY_INVALID;
    own[j] = 0;
    if (key[j] != KEY_INVALID) {
      own[j] = owner_of(L, key[j]);
      slot[j] = atomicAdd(&cnt[own[j]], 1u);
    }
  }
  __syncthreads();
  if (threadIdx.x < L.n_parts) {
    uint32_t c = cnt[threadIdx.x];
    base[threadIdx.x] = c ? atomicAdd(part_counts + threadIdx.x, (unsigned long long)c) : 0ull;
  }
  if (MODE == 0) return;
  __syncthreads();
#pragma unroll
  for (int j = 0; j < ITEMS; j++) {
    if (key[j] != KEY_INVALID) {
      unsigned long long pos = base[own[j]] + slot[j];
      out_keys[pos] = key[j];
      out_index[pos] = tuple_base + (uint32_t)(t0 + (uint64_t)j * 256 + threadIdx.x);
    }
  }
}

// ---- block join across shards ---------------------------------------------------------------
// Directed entries (global row << 32 | global column) that arrived for this shard's rows: rows
// become shard-local, the list lengths are counted.
__global__ void __launch_bounds__(256)
    import_entries_kernel(const uint64_t *__restrict__ in, uint64_t n, uint32_t base, uint64_t U,
                          uint64_t *__restrict__ entries, uint32_t *__restrict__ deg, unsigned long long *misrouted) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint64_t e = in[i];
  const uint32_t row = (uint32_t)(e >> 32) - base;
  if (row >= U) {  // an entry for a row this shard does not own: the host mixed up the exchange
    atomicAdd(misrouted, 1ull);
    entries[i] = 0;
    return;
  }
  entries[i] = ((uint64_t)row << 32) | (uint32_t)e;
  atomicAdd(deg + row, 1u);
}

// ---- graph ---------------------------------------------------------------------------------
__global__ void degree_kernel(const uint64_t *__restrict__ row_ptr, uint64_t U, uint32_t *deg) {
  uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r < U) deg[r] = (uint32_t)(row_ptr[r + 1] - row_ptr[r]);
}

__global__ void widen_rows_kernel(const uint32_t *__restrict__ excl, uint64_t U, uint64_t total, uint64_t *row_ptr) {
  uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r < U) row_ptr[r] = excl[r];
  if (r == U) row_ptr[U] = total;
}

// ---- clustering across shards ------------------------------------------------------------------
// A component of the neighbour graph that has no entry pointing outside the shard ("closed") is
// complete here and is replayed by stage_cluster as on one GPU.  The others ("open") are
// exported whole — every member with its count, first tuple and list — gathered on all shards
// and replayed there on a compact copy of that sub-graph; every shard keeps the results for
// the words it owns.
__global__ void open_flags_kernel(const uint32_t *__restrict__ comp, const uint32_t *__restrict__ comp_open,
                                  const uint64_t *__restrict__ row_ptr, uint64_t U, uint32_t *__restrict__ is_open,
                                  uint32_t *__restrict__ open_deg) {
  uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= U) return;
  uint32_t o = comp_open[comp[r]];
  is_open[r] = o;
  open_deg[r] = o ? (uint32_t)(row_ptr[r + 1] - row_ptr[r]) : 0u;
}

__global__ void __launch_bounds__(256)
    export_open_kernel(const uint32_t *__restrict__ is_open, const uint32_t *__restrict__ open_pos,
                       const uint32_t *__restrict__ open_ent, const uint64_t *__restrict__ row_ptr,
                       const uint32_t *__restrict__ col, const uint32_t *__restrict__ count,
                       const uint32_t *__restrict__ first, uint64_t U, uint32_t base, uint32_t *__restrict__ o_rank,
                       uint32_t *__restrict__ o_count, uint32_t *__restrict__ o_first, uint32_t *__restrict__ o_deg,
                       uint32_t *__restrict__ o_col) {
  uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= U || !is_open[r]) return;
  const uint32_t i = open_pos[r];
  const uint64_t lo = row_ptr[r], hi = row_ptr[r + 1];
  o_rank[i] = base + (uint32_t)r;
  o_count[i] = count[r];
  o_first[i] = first[r];
  o_deg[i] = (uint32_t)(hi - lo);
  uint32_t *dst = o_col + open_ent[r];
  for (uint64_t e = lo; e < hi; e++) dst[e - lo] = col[e];
}

// columns of the gathered open sub-graph: global rank -> position in the (ascending) rank list
__global__ void __launch_bounds__(256)
    remap_open_cols_kernel(const uint32_t *__restrict__ ranks, uint64_t n, const uint32_t *__restrict__ col_in,
                           uint64_t E, uint32_t *__restrict__ col_out) {
  uint64_t e = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= E) return;
  const uint32_t v = col_in[e];
  uint64_t lo = 0, hi = n;
  while (lo < hi) {
    uint64_t mid = (lo + hi) >> 1;
    if (ranks[mid] < v) lo = mid + 1; else hi = mid;
  }
  col_out[e] = (uint32_t)lo;
}

// closed seeds: maxLeaf as a global rank + its first tuple
__global__ void closed_seeds_kernel(const uint32_t *__restrict__ seed_flag, uint32_t *__restrict__ seed_max,
                                    const uint32_t *__restrict__ word_first, uint64_t U, uint32_t base,
                                    uint32_t *__restrict__ seed_first) {
  uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= U || !seed_flag[r]) return;
  seed_first[r] = word_first[seed_max[r]];
  seed_max[r] += base;
}

// results of the open sub-graph for the words this shard owns (positions [off, off + n))
__global__ void merge_open_kernel(uint64_t off, uint64_t n, const uint32_t *__restrict__ ranks,
                                  const uint32_t *__restrict__ firsts, const uint32_t *__restrict__ o_flag,
                                  const uint64_t *__restrict__ o_size, const uint32_t *__restrict__ o_max, uint32_t base,
                                  uint32_t *__restrict__ seed_flag, uint64_t *__restrict__ seed_size,
                                  uint32_t *__restrict__ seed_max, uint32_t *__restrict__ seed_first) {
  uint64_t k = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  const uint64_t i = off + k;
  const uint32_t r = ranks[i] - base;
  const uint32_t f = o_flag[i];
  seed_flag[r] = f;
  if (f) {
    seed_size[r] = o_size[i];
    seed_max[r] = ranks[o_max[i]];
    seed_first[r] = firsts[o_max[i]];
  }
}

__global__ void cluster_first_kernel(const uint32_t *__restrict__ seed_flag, const uint32_t *__restrict__ seed_index,
                                     const uint32_t *__restrict__ seed_first, uint64_t U,
                                     uint32_t *__restrict__ cluster_first) {
  uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r < U && seed_flag[r]) cluster_first[seed_index[r]] = seed_first[r];
}

__global__ void open_seed_ids_kernel(uint64_t off, uint64_t n, const uint32_t *__restrict__ ranks,
                                     const uint32_t *__restrict__ o_flag, const uint32_t *__restrict__ seed_index,
                                     uint32_t base, uint32_t id_offset, uint32_t *__restrict__ open_ids) {
  uint64_t k = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  const uint64_t i = off + k;
  if (o_flag[i]) open_ids[i] = id_offset + seed_index[ranks[i] - base] + 1u;
}

__global__ void word_ids_kernel(uint64_t U, const uint32_t *__restrict__ is_open, const uint32_t *__restrict__ open_pos,
                                uint64_t off, const uint32_t *__restrict__ o_label,
                                const uint32_t *__restrict__ open_ids, const uint32_t *__restrict__ label,
                                const uint32_t *__restrict__ seed_index, uint32_t id_offset,
                                uint32_t *__restrict__ word_cid) {
  uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= U) return;
  word_cid[r] = is_open[r] ? open_ids[o_label[off + open_pos[r]]] : id_offset + seed_index[label[r]] + 1u;
}

__global__ void keep_bits_kernel(const uint32_t *__restrict__ cluster_first, uint64_t C, uint32_t *__restrict__ bits) {
  uint64_t c = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  uint32_t t = cluster_first[c];
  atomicOr(bits + (t >> 5), 1u << (t & 31u));
}

__global__ void __launch_bounds__(256)
    shard_word_ids_for_tuples_kernel(const uint64_t *__restrict__ keys, uint64_t n, const Lookup lk,
                                     const uint32_t *__restrict__ word_cid, uint32_t *__restrict__ out) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  out[i] = word_cid[table_rank(lk, keys[i])];  // always present: the tuple was inserted here
}

}  // namespace humid

using namespace humid;

static int require_shard(humid_gpu *h, bool want) {
  if (!h) return HUMID_GPU_ERR_ARG;
  if (h->s.shard_mode != want) return fail(h, HUMID_GPU_ERR_ARG, "call order: not in the expected shard phase");
  return HUMID_GPU_OK;
}

extern "C" {

int humid_gpu_shard_begin(humid_gpu *h, uint64_t tuple_offset, uint64_t *n_usable) {
  if (!h) return HUMID_GPU_ERR_ARG;
  State &s = h->s;
  if (s.finalized || s.shard_mode) return fail(h, HUMID_GPU_ERR_ARG, "shard_begin: wrong call order");
  if (s.cfg.edit && s.cfg.distance >= 2)
    return fail(h, HUMID_GPU_ERR_UNSUPPORTED, "edit distance >= 2 is not sharded (use one GPU)");
  if (s.cfg.distance > (uint32_t)HAM_MAX_D) return fail(h, HUMID_GPU_ERR_UNSUPPORTED, "distance > 8");
  if (s.wide()) return fail(h, HUMID_GPU_ERR_UNSUPPORTED, "words of more than 31 letters are not sharded (use one GPU)");
  if (tuple_offset + s.n_tuples >= 0xFFFFFFFFull) return fail(h, HUMID_GPU_ERR_UNSUPPORTED, "more than 2^32-2 tuples");
  return guarded(h, [&] {
    HUMID_CUDA(cudaSetDevice(s.device));
    HUMID_CUDA(cudaMemcpyAsync(&s.usable, s.counters.p + 0, sizeof(uint64_t), cudaMemcpyDeviceToHost, s.stream));
    HUMID_CUDA(cudaStreamSynchronize(s.stream));
    s.tuple_base = tuple_offset;
    s.shard_mode = true;
    if (n_usable) *n_usable = s.usable;
  });
}

int humid_gpu_shard_sample_keys(humid_gpu *h, uint32_t n_samples, uint64_t *keys_out, uint32_t *n_out) {
  if (int rc = require_shard(h, true)) return rc;
  State &s = h->s;
  if (!keys_out || !n_out) return fail(h, HUMID_GPU_ERR_ARG, "sample_keys: null argument");
  *n_out = 0;
  if (n_samples == 0 || s.n_tuples == 0) return HUMID_GPU_OK;
  return guarded(h, [&] {
    HUMID_CUDA(cudaSetDevice(s.device));
    uint64_t *tmp = (uint64_t *)s.scratch.get(n_samples * sizeof(uint64_t));
    HUMID_CUDA(cudaMemsetAsync(s.counters.p + 9, 0, sizeof(unsigned long long), s.stream));
    sample_tuple_keys_kernel<<<(n_samples + 255) / 256, 256, 0, s.stream>>>(s.tuple_key.p, s.n_tuples, n_samples, tmp,
                                                                            s.counters.p + 9);
    uint64_t got = 0;
    HUMID_CUDA(cudaMemcpyAsync(&got, s.counters.p + 9, sizeof(uint64_t), cudaMemcpyDeviceToHost, s.stream));
    HUMID_CUDA(cudaStreamSynchronize(s.stream));
    if (got) HUMID_CUDA(cudaMemcpy(keys_out, tmp, got * sizeof(uint64_t), cudaMemcpyDeviceToHost));
    *n_out = (uint32_t)got;
  });
}

static void run_partition(State &s, const uint64_t *splitters, uint32_t n_parts, int mode, const uint64_t *offsets,
                          uint64_t *d_keys, uint32_t *d_index, uint64_t *counts_out) {
  Layout L{};
  L.n_parts = n_parts;
  for (uint32_t i = 0; i + 1 < n_parts; i++) L.splitters[i] = splitters[i];
  unsigned long long *pc = s.counters.p + 16;
  unsigned long long init[HUMID_MAX_PARTS] = {0};
  if (offsets) for (uint32_t p = 0; p < n_parts; p++) init[p] = offsets[p];
  HUMID_CUDA(cudaMemcpyAsync(pc, init, sizeof(init), cudaMemcpyHostToDevice, s.stream));
  if (s.n_tuples) {
    Timed t(s, mode ? "shard_partition_tuples" : "shard_count_tuples", s.n_tuples, s.n_tuples * (mode ? 20 : 8));
    unsigned grid = grid_for(s.n_tuples, 256, 4);
    if (mode)
      partition_tuples_kernel<1><<<grid, 256, 0, s.stream>>>(s.tuple_key.p, s.n_tuples, (uint32_t)s.tuple_base, L, pc,
                                                             d_keys, d_index);
    else
      partition_tuples_kernel<0><<<grid, 256, 0, s.stream>>>(s.tuple_key.p, s.n_tuples, (uint32_t)s.tuple_base, L, pc,
                                                             nullptr, nullptr);
  }
  HUMID_CUDA(cudaGetLastError());
  if (counts_out) {
    unsigned long long host[HUMID_MAX_PARTS];
    HUMID_CUDA(cudaMemcpyAsync(host, pc, sizeof(host), cudaMemcpyDeviceToHost, s.stream));
    HUMID_CUDA(cudaStreamSynchronize(s.stream));
    for (uint32_t p = 0; p < n_parts; p++) counts_out[p] = host[p];
  } else {
    HUMID_CUDA(cudaStreamSynchronize(s.stream));
  }
}

int humid_gpu_shard_partition(humid_gpu *h, const uint64_t *splitters, uint32_t n_parts, uint64_t *counts) {
  if (int rc = require_shard(h, true)) return rc;
  State &s = h->s;
  if (!counts || n_parts == 0 || n_parts > HUMID_MAX_PARTS || (n_parts > 1 && !splitters))
    return fail(h, HUMID_GPU_ERR_ARG, "partition: bad arguments");
  return guarded(h, [&] {
    HUMID_CUDA(cudaSetDevice(s.device));
    s.n_parts = n_parts;
    for (uint32_t i = 0; i + 1 < n_parts; i++) s.splitters[i] = splitters[i];
    run_partition(s, splitters, n_parts, 0, nullptr, nullptr, nullptr, counts);
  });
}

int humid_gpu_shard_export_tuples(humid_gpu *h, const uint64_t *offsets, uint64_t *d_keys, uint32_t *d_index) {
  if (int rc = require_shard(h, true)) return rc;
  State &s = h->s;
  if (!offsets) return fail(h, HUMID_GPU_ERR_ARG, "export_tuples: null offsets");
  return guarded(h, [&] {
    HUMID_CUDA(cudaSetDevice(s.device));
    run_partition(s, s.splitters, s.n_parts, 1, offsets, d_keys, d_index, nullptr);
  });
}

int humid_gpu_shard_import_tuples(humid_gpu *h, const uint64_t *d_keys, const uint32_t *d_index, uint64_t n,
                                  uint64_t *n_shard_words) {
  if (int rc = require_shard(h, true)) return rc;
  State &s = h->s;
  if (!n_shard_words || (n && (!d_keys || !d_index))) return fail(h, HUMID_GPU_ERR_ARG, "import_tuples: null buffer");
  if (n >= 0xFFFFFFFFull) return fail(h, HUMID_GPU_ERR_UNSUPPORTED, "too many tuples in one shard");
  return guarded(h, [&] {
    HUMID_CUDA(cudaSetDevice(s.device));
    // keep the received tuples: they are this shard's share of the job from here on
    if (s.alias_imports) {  // group mode: the tuples arrived in a buffer of the handle that stays as it is
      s.src_keys = d_keys;
      s.src_index = d_index;
    } else {
      s.shard_keys.ensure(n + 1);
      s.shard_index.ensure(n + 1);
      if (n) {
        HUMID_CUDA(cudaMemcpyAsync(s.shard_keys.p, d_keys, n * 8, cudaMemcpyDeviceToDevice, s.stream));
        HUMID_CUDA(cudaMemcpyAsync(s.shard_index.p, d_index, n * 4, cudaMemcpyDeviceToDevice, s.stream));
      }
      s.src_keys = s.shard_keys.p;
      s.src_index = s.shard_index.p;
    }
    s.src_n = n;
    stage_build_table(s, false);  // the single-GPU table / sort / column stages on this key range
    s.n_shard_words = s.unique;
    HUMID_CUDA(cudaStreamSynchronize(s.stream));
    *n_shard_words = s.unique;
  });
}

int humid_gpu_shard_set_layout(humid_gpu *h, uint32_t part, uint32_t n_parts, const uint64_t *splitters,
                               const uint64_t *bases) {
  if (int rc = require_shard(h, true)) return rc;
  State &s = h->s;
  if (n_parts == 0 || n_parts > HUMID_MAX_PARTS || part >= n_parts || !bases || (n_parts > 1 && !splitters))
    return fail(h, HUMID_GPU_ERR_ARG, "set_layout: bad arguments");
  if (bases[n_parts] >= 0xFFFFFFFFull) return fail(h, HUMID_GPU_ERR_UNSUPPORTED, "more than 2^32-2 distinct words");
  s.part = part;
  s.n_parts = n_parts;
  for (uint32_t i = 0; i + 1 < n_parts; i++) s.splitters[i] = splitters[i];
  for (uint32_t i = 0; i <= n_parts; i++) s.bases[i] = bases[i];
  return HUMID_GPU_OK;
}

// Block geometry of the job (neighbours.cu); in shard mode the joins never cap a group.
static Blocks shard_blocks(const State &s, int lo_bit[], int hi_bit[]) {
  return make_blocks(s.cfg.word_length, s.cfg.distance, lo_bit, hi_bit);
}

int humid_gpu_shard_join_local(humid_gpu *h, uint64_t *n_entries) {
  if (int rc = require_shard(h, true)) return rc;
  State &s = h->s;
  return guarded(h, [&] {
    HUMID_CUDA(cudaSetDevice(s.device));
    const uint64_t U = s.n_shard_words;
    const uint32_t d = s.cfg.distance;
    s.n_entries = 0;
    s.deg.ensure(U + 1);
    if (U) HUMID_CUDA(cudaMemsetAsync(s.deg.p, 0, U * sizeof(uint32_t), s.stream));
    if (n_entries) *n_entries = 0;
    if (U == 0 || d == 0) return;
    if (d > HAM_MAX_D) throw CudaError{"Hamming distance > 8 is not supported"};
    int lo_bit[HAM_MAX_D + 1], hi_bit[HAM_MAX_D + 1];
    const Blocks B = shard_blocks(s, lo_bit, hi_bit);
    unsigned long long *counter = s.counters.p + 2;
    uint64_t cap = std::max<uint64_t>(s.entries.cap, std::max<uint64_t>(4 * U, 1u << 20));
    uint64_t pairs = 0;
    for (int attempt = 0; attempt < 2; attempt++) {
      s.entries.ensure(cap);
      cap = s.entries.cap;
      HUMID_CUDA(cudaMemsetAsync(counter, 0, sizeof(unsigned long long), s.stream));
      HUMID_CUDA(cudaMemsetAsync(s.deg.p, 0, U * sizeof(uint32_t), s.stream));
      {
        Timed t(s, "shard_join_local", U, U * 16);
        launch_hamming_join(s, JOIN_DIRECTED, s.word_key.p, nullptr, U, B, 0, JOIN_NO_CAP, s.entries.p, cap / 2, counter,
                            s.deg.p, (uint32_t)s.bases[s.part], nullptr, nullptr);
      }
      HUMID_CUDA(cudaGetLastError());
      HUMID_CUDA(cudaMemcpyAsync(&pairs, counter, sizeof(uint64_t), cudaMemcpyDeviceToHost, s.stream));
      HUMID_CUDA(cudaStreamSynchronize(s.stream));
      if (2 * pairs <= cap) break;
      cap = 2 * pairs;
    }
    s.n_entries = 2 * pairs;
    if (n_entries) *n_entries = s.n_entries;
  });
}

int humid_gpu_shard_join_export(humid_gpu *h, uint32_t block, uint64_t *counts, uint64_t *d_keys, uint32_t *d_ranks) {
  if (int rc = require_shard(h, true)) return rc;
  State &s = h->s;
  if (!counts || block == 0 || block > s.cfg.distance) return fail(h, HUMID_GPU_ERR_ARG, "join_export: bad arguments");
  return guarded(h, [&] {
    HUMID_CUDA(cudaSetDevice(s.device));
    const uint64_t U = s.n_shard_words;
    for (uint32_t p = 0; p < s.n_parts; p++) counts[p] = 0;
    if (U == 0) return;
    if (!d_keys || !d_ranks) throw CudaError{"join_export: null buffer"};
    int lo_bit[HAM_MAX_D + 1], hi_bit[HAM_MAX_D + 1];
    const Blocks B = shard_blocks(s, lo_bit, hi_bit);
    Timed t(s, "shard_join_export", U, U * (8 + 12));
    t.launches = partition_by_block_hash(s.word_key.p, (uint32_t)s.bases[s.part], d_keys, d_ranks, U, B.mask[block],
                                         s.n_parts, counts, s.scratch, s.stream);
  });
}

int humid_gpu_shard_join_match(humid_gpu *h, uint32_t block, const uint64_t *d_keys, const uint32_t *d_ranks, uint64_t n,
                               uint64_t *counts) {
  if (int rc = require_shard(h, true)) return rc;
  State &s = h->s;
  if (!counts || block == 0 || block > s.cfg.distance) return fail(h, HUMID_GPU_ERR_ARG, "join_match: bad arguments");
  if (n >= 0xFFFFFFFFull) return fail(h, HUMID_GPU_ERR_UNSUPPORTED, "too many words in one join");
  return guarded(h, [&] {
    HUMID_CUDA(cudaSetDevice(s.device));
    for (uint32_t p = 0; p < s.n_parts; p++) counts[p] = 0;
    s.n_join_entries = 0;
    if (n < 2) return;
    if (!d_keys || !d_ranks) throw CudaError{"join_match: null buffer"};
    int lo_bit[HAM_MAX_D + 1], hi_bit[HAM_MAX_D + 1];
    const Blocks B = shard_blocks(s, lo_bit, hi_bit);
    const uint64_t *keys = d_keys;
    const uint32_t *rank = d_ranks;
    if (hi_bit[block] > lo_bit[block]) {  // words arrive in rank order: a stable sort keeps that inside a group
      s.key_alt.ensure(n + 1);
      s.entries_alt.ensure(n + 1);
      s.entry_val.ensure(n + 1);
      s.entry_val_alt.ensure(n + 1);
      uint64_t *kb[2] = {s.key_alt.p, s.entries_alt.p};
      uint32_t *vb[2] = {s.entry_val.p, s.entry_val_alt.p};
      int launches = 0;
      Timed t(s, "shard_join_sort", n, n * (8 + 4) * 2);
      int cur = radix_sort_bits_pairs(d_keys, d_ranks, kb, vb, n, lo_bit[block], hi_bit[block], s.scratch, s.stream,
                                      &launches);
      t.launches = launches;
      keys = kb[cur];
      rank = vb[cur];
    }
    unsigned long long *counter = s.counters.p + 41;
    uint64_t cap = std::max<uint64_t>(s.join_entries.cap, std::max<uint64_t>(n, 1u << 20));
    uint64_t pairs = 0;
    for (int attempt = 0; attempt < 2; attempt++) {
      s.join_entries.ensure(cap);
      cap = s.join_entries.cap;
      HUMID_CUDA(cudaMemsetAsync(counter, 0, sizeof(unsigned long long), s.stream));
      {
        Timed t(s, "shard_join_match", n, n * 16);
        launch_hamming_join(s, JOIN_DIRECTED, keys, rank, n, B, block, JOIN_NO_CAP, s.join_entries.p, cap / 2, counter,
                            nullptr, 0u, nullptr, nullptr);
      }
      HUMID_CUDA(cudaGetLastError());
      HUMID_CUDA(cudaMemcpyAsync(&pairs, counter, sizeof(uint64_t), cudaMemcpyDeviceToHost, s.stream));
      HUMID_CUDA(cudaStreamSynchronize(s.stream));
      if (2 * pairs <= cap) break;
      cap = 2 * pairs;
    }
    s.n_join_entries = 2 * pairs;
    if (s.n_join_entries == 0) return;
    s.join_entries_out.ensure(s.n_join_entries);
    Timed t(s, "shard_join_route", s.n_join_entries, s.n_join_entries * 16);
    t.launches = partition_by_row_owner(s.join_entries.p, s.join_entries_out.p, s.n_join_entries, s.bases, s.n_parts,
                                        counts, s.scratch, s.stream);
  });
}

int humid_gpu_shard_join_entries(humid_gpu *h, uint64_t *d_entries) {
  if (int rc = require_shard(h, true)) return rc;
  State &s = h->s;
  return guarded(h, [&] {
    HUMID_CUDA(cudaSetDevice(s.device));
    if (s.n_join_entries == 0) return;
    if (!d_entries) throw CudaError{"join_entries: null buffer"};
    HUMID_CUDA(cudaMemcpyAsync(d_entries, s.join_entries_out.p, s.n_join_entries * 8, cudaMemcpyDeviceToDevice, s.stream));
    HUMID_CUDA(cudaStreamSynchronize(s.stream));
  });
}

int humid_gpu_shard_join_import(humid_gpu *h, const uint64_t *d_entries, uint64_t n) {
  if (int rc = require_shard(h, true)) return rc;
  State &s = h->s;
  if (n && !d_entries) return fail(h, HUMID_GPU_ERR_ARG, "join_import: null buffer");
  return guarded(h, [&] {
    HUMID_CUDA(cudaSetDevice(s.device));
    if (n == 0) return;
    if (s.n_entries + n >= 0xFFFFFFFFull) throw CudaError{"more than 2^32-2 neighbour entries"};
    s.entries.grow(s.n_entries + n, s.n_entries, s.stream);
    Timed t(s, "shard_join_import", n, n * 20);
    unsigned long long *misrouted = s.counters.p + 42;
    HUMID_CUDA(cudaMemsetAsync(misrouted, 0, sizeof(unsigned long long), s.stream));
    import_entries_kernel<<<grid_for(n, 256), 256, 0, s.stream>>>(d_entries, n, (uint32_t)s.bases[s.part], s.n_shard_words,
                                                                  s.entries.p + s.n_entries, s.deg.p, misrouted);
    HUMID_CUDA(cudaGetLastError());
    s.n_entries += n;
    uint64_t bad = 0;
    HUMID_CUDA(cudaMemcpyAsync(&bad, misrouted, sizeof(uint64_t), cudaMemcpyDeviceToHost, s.stream));
    HUMID_CUDA(cudaStreamSynchronize(s.stream));
    if (bad) throw CudaError{"join_import: " + std::to_string(bad) + " entries for rows outside this shard"};
  });
}

int humid_gpu_shard_build_rows(humid_gpu *h, uint64_t *n_entries) {
  if (int rc = require_shard(h, true)) return rc;
  State &s = h->s;
  return guarded(h, [&] {
    HUMID_CUDA(cudaSetDevice(s.device));
    stage_csr_directed(s);  // rows: the shard's words, columns: global ranks
    HUMID_CUDA(cudaStreamSynchronize(s.stream));
    if (n_entries) *n_entries = s.n_entries;
  });
}

int humid_gpu_shard_cluster_local(humid_gpu *h, uint64_t *n_open_words, uint64_t *n_open_entries) {
  if (int rc = require_shard(h, true)) return rc;
  State &s = h->s;
  if (!n_open_words || !n_open_entries) return HUMID_GPU_ERR_ARG;
  return guarded(h, [&] {
    HUMID_CUDA(cudaSetDevice(s.device));
    const uint64_t U = s.n_shard_words;
    s.cluster_col_base = (uint32_t)s.bases[s.part];
    s.cluster_skip_open = true;
    s.n_open_local = s.n_open_entries_local = 0;
    stage_cluster(s);  // closed components only, no ids yet
    if (U) {
      Timed t(s, "shard_open_words", U, U * 24);
      s.is_open.ensure(U);
      s.open_pos.ensure(U + 1);
      s.open_ent.ensure(U + 1);
      s.work_c.ensure(U + 1);
      open_flags_kernel<<<grid_for(U, 256), 256, 0, s.stream>>>(s.comp.p, s.comp_open.p, s.row_ptr.p, U, s.is_open.p,
                                                                s.work_c.p);
      uint32_t *tot = reinterpret_cast<uint32_t *>(s.counters.p + 10);
      int l = exclusive_scan_u32(s.is_open.p, s.open_pos.p, U, tot, s.scan_scratch, s.stream);
      l += exclusive_scan_u32(s.work_c.p, s.open_ent.p, U, tot + 2, s.scan_scratch, s.stream);
      uint32_t host[4] = {0, 0, 0, 0};
      HUMID_CUDA(cudaMemcpyAsync(host, tot, sizeof(host), cudaMemcpyDeviceToHost, s.stream));
      HUMID_CUDA(cudaStreamSynchronize(s.stream));
      s.n_open_local = host[0];
      s.n_open_entries_local = host[2];
      t.launches = l + 1;
    }
    *n_open_words = s.n_open_local;
    *n_open_entries = s.n_open_entries_local;
  });
}

int humid_gpu_shard_export_open(humid_gpu *h, uint32_t *d_rank, uint32_t *d_count, uint32_t *d_first, uint32_t *d_deg,
                                uint32_t *d_col) {
  if (int rc = require_shard(h, true)) return rc;
  State &s = h->s;
  return guarded(h, [&] {
    HUMID_CUDA(cudaSetDevice(s.device));
    const uint64_t U = s.n_shard_words;
    if (U && s.n_open_local) {
      Timed t(s, "shard_export_open", s.n_open_local, s.n_open_local * 20 + s.n_open_entries_local * 8);
      export_open_kernel<<<grid_for(U, 256), 256, 0, s.stream>>>(s.is_open.p, s.open_pos.p, s.open_ent.p, s.row_ptr.p,
                                                                 s.col.p, s.word_count.p, s.word_first.p, U,
                                                                 (uint32_t)s.bases[s.part], d_rank, d_count, d_first,
                                                                 d_deg, d_col);
    }
    HUMID_CUDA(cudaGetLastError());
    HUMID_CUDA(cudaStreamSynchronize(s.stream));
  });
}

int humid_gpu_shard_import_open(humid_gpu *h, uint64_t open_offset, uint64_t n_open, uint64_t n_entries,
                                const uint32_t *d_rank, const uint32_t *d_count, const uint32_t *d_first,
                                const uint32_t *d_deg, const uint32_t *d_col, uint64_t *n_clusters_local) {
  if (int rc = require_shard(h, true)) return rc;
  State &s = h->s;
  if (!n_clusters_local || open_offset + s.n_open_local > n_open) return fail(h, HUMID_GPU_ERR_ARG, "import_open: bad arguments");
  return guarded(h, [&] {
    HUMID_CUDA(cudaSetDevice(s.device));
    const uint64_t U = s.n_shard_words;
    const uint32_t base = (uint32_t)s.bases[s.part];
    if (!s.open) {
      s.open = new State();
      s.open->cfg = s.cfg;
      s.open->device = s.device;
      s.open->stream = s.stream;
      s.open->counters.ensure(64);
    }
    State &o = *s.open;
    o.stats.clear();
    o.unique = n_open;
    o.n_entries = n_entries;
    s.open_offset = open_offset;
    s.n_open_total = n_open;
    s.open_rank.ensure(n_open + 1);
    s.open_first.ensure(n_open + 1);
    if (n_open) {
      Timed t(s, "shard_open_graph", n_open, n_open * 28 + n_entries * 12);
      HUMID_CUDA(cudaMemcpyAsync(s.open_rank.p, d_rank, n_open * 4, cudaMemcpyDeviceToDevice, s.stream));
      HUMID_CUDA(cudaMemcpyAsync(s.open_first.p, d_first, n_open * 4, cudaMemcpyDeviceToDevice, s.stream));
      o.word_count.ensure(n_open + 1);
      HUMID_CUDA(cudaMemcpyAsync(o.word_count.p, d_count, n_open * 4, cudaMemcpyDeviceToDevice, s.stream));
      o.row_ptr.ensure(n_open + 2);
      o.col.ensure(n_entries + 1);
      o.seed_index.ensure(n_open + 1);
      int l = exclusive_scan_u32(d_deg, o.seed_index.p, n_open, nullptr, s.scan_scratch, s.stream);
      widen_rows_kernel<<<grid_for(n_open + 1, 256), 256, 0, s.stream>>>(o.seed_index.p, n_open, n_entries, o.row_ptr.p);
      if (n_entries)
        remap_open_cols_kernel<<<grid_for(n_entries, 256), 256, 0, s.stream>>>(s.open_rank.p, n_open, d_col, n_entries,
                                                                               o.col.p);
      t.launches = l + 2;
    }
    HUMID_CUDA(cudaMemsetAsync(o.counters.p, 0, 64 * sizeof(unsigned long long), s.stream));
    o.cluster_no_ids = true;  // only flags / sizes / maxLeaf per seed are merged below
    o.cluster_owner_mod = s.open_owner_mod;  // a group shares the open components out among its ranks
    o.cluster_owner_me = s.open_owner_me;
    stage_cluster(o);  // the same replay, on the gathered open components
    if (s.after_open_cluster) s.after_open_cluster(o);  // results of the components other ranks own arrive here
    for (auto &st : o.stats) s.stats.push_back(st);
    o.stats.clear();
    s.seed_first.ensure(U + 1);
    if (U) {
      closed_seeds_kernel<<<grid_for(U, 256), 256, 0, s.stream>>>(s.seed_flag.p, s.seed_max.p, s.word_first.p, U, base,
                                                                  s.seed_first.p);
      if (s.n_open_local)
        merge_open_kernel<<<grid_for(s.n_open_local, 256), 256, 0, s.stream>>>(
            open_offset, s.n_open_local, s.open_rank.p, s.open_first.p, o.seed_flag.p, o.seed_size.p, o.seed_max.p, base,
            s.seed_flag.p, s.seed_size.p, s.seed_max.p, s.seed_first.p);
    }
    stage_cluster_number(s);  // ids of this shard's seeds, relative to the shard
    s.cluster_first.ensure(s.n_clusters + 1);
    if (U)
      cluster_first_kernel<<<grid_for(U, 256), 256, 0, s.stream>>>(s.seed_flag.p, s.seed_index.p, s.seed_first.p, U,
                                                                   s.cluster_first.p);
    HUMID_CUDA(cudaGetLastError());
    HUMID_CUDA(cudaStreamSynchronize(s.stream));
    *n_clusters_local = s.n_clusters;
  });
}

int humid_gpu_shard_assign_ids(humid_gpu *h, uint64_t id_offset, uint32_t *d_open_ids) {
  if (int rc = require_shard(h, true)) return rc;
  State &s = h->s;
  if (!s.open) return fail(h, HUMID_GPU_ERR_ARG, "assign_ids before import_open");
  return guarded(h, [&] {
    HUMID_CUDA(cudaSetDevice(s.device));
    s.id_offset = id_offset;
    if (s.n_open_local)
      open_seed_ids_kernel<<<grid_for(s.n_open_local, 256), 256, 0, s.stream>>>(
          s.open_offset, s.n_open_local, s.open_rank.p, s.open->seed_flag.p, s.seed_index.p, (uint32_t)s.bases[s.part],
          (uint32_t)id_offset, d_open_ids);
    HUMID_CUDA(cudaGetLastError());
    HUMID_CUDA(cudaStreamSynchronize(s.stream));
  });
}

int humid_gpu_shard_finish(humid_gpu *h, const uint32_t *d_open_ids, uint64_t n_tuples, uint64_t n_usable,
                           uint64_t n_words, uint64_t n_clusters) {
  if (int rc = require_shard(h, true)) return rc;
  State &s = h->s;
  if (!s.open) return fail(h, HUMID_GPU_ERR_ARG, "finish before import_open");
  if (n_tuples >= 0xFFFFFFFFull) return fail(h, HUMID_GPU_ERR_UNSUPPORTED, "more than 2^32-2 tuples");
  return guarded(h, [&] {
    HUMID_CUDA(cudaSetDevice(s.device));
    const uint64_t U = s.n_shard_words;
    s.word_cid.ensure(U + 1);
    if (U) {
      Timed t(s, "shard_word_ids", U, U * 20);
      word_ids_kernel<<<grid_for(U, 256), 256, 0, s.stream>>>(U, s.is_open.p, s.open_pos.p, s.open_offset,
                                                              s.open->label.p, d_open_ids, s.label.p, s.seed_index.p,
                                                              (uint32_t)s.id_offset, s.word_cid.p);
    }
    HUMID_CUDA(cudaGetLastError());
    HUMID_CUDA(cudaStreamSynchronize(s.stream));
    s.report_tuples = n_tuples;
    s.report_usable = n_usable;
    s.report_unique = n_words;
    s.report_clusters = n_clusters;
    resolve_stats(s);
    s.finalized = true;
  });
}

int humid_gpu_shard_keep_bits(humid_gpu *h, uint32_t *d_bits, uint64_t n_words32) {
  if (int rc = require_shard(h, true)) return rc;
  State &s = h->s;
  if (!s.finalized) return fail(h, HUMID_GPU_ERR_ARG, "not finalized");
  if (n_words32 * 32 < s.report_tuples) return fail(h, HUMID_GPU_ERR_ARG, "keep_bits: buffer too small");
  return guarded(h, [&] {
    HUMID_CUDA(cudaSetDevice(s.device));
    HUMID_CUDA(cudaMemsetAsync(d_bits, 0, n_words32 * 4, s.stream));
    if (s.n_clusters) {
      Timed t(s, "keep", s.n_clusters, s.n_clusters * 8);
      keep_bits_kernel<<<grid_for(s.n_clusters, 256), 256, 0, s.stream>>>(s.cluster_first.p, s.n_clusters, d_bits);
    }
    HUMID_CUDA(cudaGetLastError());
    HUMID_CUDA(cudaStreamSynchronize(s.stream));
    resolve_stats(s);
  });
}

int humid_gpu_shard_tuple_ids(humid_gpu *h, uint32_t *d_ids) {
  if (int rc = require_shard(h, true)) return rc;
  State &s = h->s;
  if (!s.finalized) return fail(h, HUMID_GPU_ERR_ARG, "not finalized");
  if (s.src_n == 0) return HUMID_GPU_OK;
  if (!d_ids) return HUMID_GPU_ERR_ARG;
  return guarded(h, [&] {
    HUMID_CUDA(cudaSetDevice(s.device));
    stage_ensure_lookup(s);  // key -> local rank
    {
      Timed t(s, "cluster_ids", s.src_n, s.src_n * 36);
      shard_word_ids_for_tuples_kernel<<<grid_for(s.src_n, 256), 256, 0, s.stream>>>(s.src_keys, s.src_n, s.lookup(),
                                                                                     s.word_cid.p, d_ids);
    }
    HUMID_CUDA(cudaGetLastError());
    HUMID_CUDA(cudaStreamSynchronize(s.stream));
    resolve_stats(s);
  });
}

}  // extern "C"
