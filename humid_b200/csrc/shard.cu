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
};

__device__ __forceinline__ uint32_t owner_of(const Layout &L, uint64_t key) {
  uint32_t p = 0;
  for (uint32_t i = 0; i + 1 < L.n_parts; i++) p += (key >= L.splitters[i]) ? 1u : 0u;
  return p;
}

// ---- words ---------------------------------------------------------------------------------
__global__ void add_offset_kernel(uint32_t *a, uint64_t n, uint32_t off) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) a[i] += off;
}

__global__ void sample_keys_kernel(const uint64_t *__restrict__ keys, uint64_t U, uint32_t n, uint64_t *out) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = keys[(uint64_t)(((unsigned __int128)i * U + U / 2) / n) % U];
}

__global__ void split_kernel(const uint64_t *__restrict__ keys, uint64_t U, const Layout L, uint64_t *bounds) {
  uint32_t i = threadIdx.x;  // bounds[i] = first index with key >= splitters[i]
  if (i + 1 >= L.n_parts) return;
  uint64_t lo = 0, hi = U;
  while (lo < hi) {
    uint64_t mid = (lo + hi) >> 1;
    if (keys[mid] < L.splitters[i]) lo = mid + 1; else hi = mid;
  }
  bounds[i] = lo;
}

__global__ void iota32_kernel(uint32_t *a, uint64_t n) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) a[i] = (uint32_t)i;
}

__global__ void mark_heads_kernel(const uint64_t *__restrict__ sorted, uint64_t n, uint32_t *flag) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) flag[i] = (i > 0 && sorted[i] != sorted[i - 1]) ? 1u : 0u;  // inclusive scan => word index
}

// sorted record i (original position idx[i]) belongs to shard word widx = scan[i] + flag[i]
__global__ void __launch_bounds__(256)
    merge_words_kernel(const uint64_t *__restrict__ sorted, const uint32_t *__restrict__ idx,
                       const uint32_t *__restrict__ flag, const uint32_t *__restrict__ scan, uint64_t n,
                       const uint32_t *__restrict__ in_count, const uint32_t *__restrict__ in_first,
                       uint64_t *__restrict__ word_key, uint32_t *__restrict__ word_count,
                       uint32_t *__restrict__ word_first, uint32_t *__restrict__ local_rank_out) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint32_t w = scan[i] + flag[i];
  uint32_t o = idx[i];
  if (i == 0 || flag[i]) word_key[w] = sorted[i];
  atomicAdd(word_count + w, in_count[o]);   // Leaf::count summed over the ranks
  atomicMin(word_first + w, in_first[o]);   // first tuple in global input order
  local_rank_out[o] = w;
}

__global__ void __launch_bounds__(256)
    word_table_insert_kernel(const uint64_t *__restrict__ word_key, uint64_t U, Slot *slots, uint64_t mask) {
  uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= U) return;
  uint64_t key = word_key[r];
  uint64_t idx = mix64(key) & mask;
  while (true) {
    unsigned long long prev = atomicCAS(reinterpret_cast<unsigned long long *>(&slots[idx].key),
                                        (unsigned long long)KEY_INVALID, (unsigned long long)key);
    if (prev == KEY_INVALID) break;  // keys are distinct
    idx = (idx + 1) & mask;
  }
  slots[idx].count = (uint32_t)r;  // rank
}

__global__ void table_clear2_kernel(Slot *slots, uint64_t cap) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < cap) {
    uint4 v;
    v.x = v.y = 0xFFFFFFFFu;
    v.z = 0u;
    v.w = 0xFFFFFFFFu;
    reinterpret_cast<uint4 *>(slots)[i] = v;
  }
}

// ---- probe ---------------------------------------------------------------------------------
// One thread per (word, first modified position); upward variants with at most d modified
// positions (strictly increasing), exactly the enumeration of hammingd_probe_kernel.
// MODE 0: count the queries per destination part.  MODE 1: look up the variants this shard
// owns, append the others to the destination's region of the query buffers.
constexpr int SHARD_MAX_D = 8;

template <typename F>
__device__ __forceinline__ void for_each_upward_variant(uint64_t w, uint32_t n, uint32_t d, uint32_t p1, F &&f) {
  uint32_t pos[SHARD_MAX_D];
  uint32_t let[SHARD_MAX_D];
  uint64_t cur = w;
  int depth = 0;
  pos[0] = p1;
  let[0] = (uint32_t)(w >> (2 * (n - 1 - p1))) & 3u;
  while (true) {
    uint32_t sh = 2 * (n - 1 - pos[depth]);
    uint32_t orig = (uint32_t)(w >> sh) & 3u;
    if (++let[depth] > 3) {
      cur = (cur & ~(3ull << sh)) | ((uint64_t)orig << sh);
      if (depth == 0) break;
      if (++pos[depth] >= n) { depth--; continue; }
      let[depth] = (uint32_t)(w >> (2 * (n - 1 - pos[depth]))) & 3u;
      continue;
    }
    cur = (cur & ~(3ull << sh)) | ((uint64_t)let[depth] << sh);
    f(cur);
    if (depth + 1 < (int)d && pos[depth] + 1 < n) {
      depth++;
      pos[depth] = pos[depth - 1] + 1;
      let[depth] = (uint32_t)(w >> (2 * (n - 1 - pos[depth]))) & 3u;
    }
  }
}

template <int MODE>
__global__ void __launch_bounds__(256)
    shard_probe_kernel(const uint64_t *__restrict__ word_key, uint64_t U, uint32_t n, uint32_t d, const Layout L,
                       const Lookup lk, unsigned long long *__restrict__ part_counts /* MODE 0: totals; MODE 1: cursors */,
                       uint64_t *__restrict__ q_keys, uint32_t *__restrict__ q_src, uint64_t *__restrict__ entries,
                       uint64_t cap, unsigned long long *__restrict__ entry_counter) {
  __shared__ uint32_t cnt[HUMID_MAX_PARTS];
  __shared__ unsigned long long base[HUMID_MAX_PARTS];
  if (threadIdx.x < HUMID_MAX_PARTS) cnt[threadIdx.x] = 0;
  __syncthreads();
  uint64_t gid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  uint64_t r = gid / n;
  const bool live = r < U;
  const uint32_t p1 = live ? (uint32_t)(gid - r * n) : 0;
  const uint64_t w = live ? word_key[r] : 0ull;
  if (live)
    for_each_upward_variant(w, n, d, p1, [&](uint64_t v) {
      uint32_t o = owner_of(L, v);
      if (o != L.part) atomicAdd(&cnt[o], 1u);
    });
  __syncthreads();
  if (threadIdx.x < L.n_parts) {
    uint32_t c = cnt[threadIdx.x];
    base[threadIdx.x] = c ? atomicAdd(part_counts + threadIdx.x, (unsigned long long)c) : 0ull;
    cnt[threadIdx.x] = 0;
  }
  if (MODE == 0) return;
  __syncthreads();
  if (!live) return;
  const uint32_t me = L.base + (uint32_t)r;
  for_each_upward_variant(w, n, d, p1, [&](uint64_t v) {
    uint32_t o = owner_of(L, v);
    if (o == L.part) {
      uint32_t hit = lookup_rank(lk, v);
      if (hit != NONE32) {
        unsigned long long pos = atomicAdd(entry_counter, 2ull);
        if (pos + 2 <= cap) {
          // rows are shard-local, columns global
          entries[pos] = (r << 32) | (uint64_t)(L.base + hit);
          entries[pos + 1] = ((uint64_t)hit << 32) | me;
        }
      }
    } else {
      unsigned long long pos = base[o] + atomicAdd(&cnt[o], 1u);
      q_keys[pos] = v;
      q_src[pos] = me;
    }
  });
}

__global__ void __launch_bounds__(256)
    shard_answer_kernel(const uint64_t *__restrict__ q_keys, const uint32_t *__restrict__ q_src, uint64_t n,
                        const Layout L, const Lookup lk, uint32_t *__restrict__ answers,
                        uint64_t *__restrict__ entries, uint64_t cap, unsigned long long *__restrict__ entry_counter) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint32_t hit = lookup_rank(lk, q_keys[i]);
  answers[i] = hit == NONE32 ? NONE32 : L.base + hit;
  if (hit != NONE32) {
    unsigned long long pos = atomicAdd(entry_counter, 1ull);
    if (pos < cap) entries[pos] = ((uint64_t)hit << 32) | q_src[i];  // result leaf <- walk leaf
  }
}

__global__ void __launch_bounds__(256)
    shard_apply_kernel(const uint32_t *__restrict__ q_src, const uint32_t *__restrict__ answers, uint64_t n,
                       const Layout L, uint64_t *__restrict__ entries, uint64_t cap,
                       unsigned long long *__restrict__ entry_counter) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint32_t a = answers[i];
  if (a != NONE32) {
    unsigned long long pos = atomicAdd(entry_counter, 1ull);
    if (pos < cap) entries[pos] = ((uint64_t)(q_src[i] - L.base) << 32) | a;  // walk leaf <- result leaf
  }
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

__global__ void __launch_bounds__(256)
    shard_cluster_ids_kernel(const uint64_t *__restrict__ tuple_key, uint64_t first, uint64_t n,
                             const Slot *__restrict__ slots, uint64_t mask, const uint32_t *__restrict__ grank,
                             const uint32_t *__restrict__ label, const uint32_t *__restrict__ seed_index,
                             uint32_t *__restrict__ out) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint64_t key = tuple_key[first + i];
  uint32_t id = 0;
  if (key != KEY_INVALID) {
    uint64_t idx = mix64(key) & mask;
    while (true) {
      uint4 v = __ldg(reinterpret_cast<const uint4 *>(slots + idx));
      uint64_t k = ((uint64_t)v.y << 32) | v.x;
      if (k == key) { id = seed_index[label[grank[v.z]]] + 1u; break; }
      idx = (idx + 1) & mask;
    }
  }
  out[i] = id;
}

static Layout layout_of(const State &s) {
  Layout L{};
  L.part = s.part;
  L.n_parts = s.n_parts;
  for (uint32_t i = 0; i + 1 < s.n_parts; i++) L.splitters[i] = s.splitters[i];
  L.base = (uint32_t)s.bases[s.part];
  return L;
}

}  // namespace humid

using namespace humid;

static int require_shard(humid_gpu *h, bool want) {
  if (!h) return HUMID_GPU_ERR_ARG;
  if (h->s.shard_mode != want) return fail(h, HUMID_GPU_ERR_ARG, "call order: not in the expected shard phase");
  return HUMID_GPU_OK;
}

extern "C" {

int humid_gpu_shard_local_words(humid_gpu *h, uint64_t tuple_offset, uint64_t *n_local_words, uint64_t *n_usable) {
  if (!h || !n_local_words) return HUMID_GPU_ERR_ARG;
  State &s = h->s;
  if (s.finalized || s.shard_mode) return fail(h, HUMID_GPU_ERR_ARG, "shard_local_words: wrong call order");
  if (s.cfg.edit && s.cfg.distance >= 2)
    return fail(h, HUMID_GPU_ERR_UNSUPPORTED, "edit distance >= 2 is not sharded (use one GPU)");
  if (s.cfg.distance > SHARD_MAX_D) return fail(h, HUMID_GPU_ERR_UNSUPPORTED, "distance > 8");
  if (tuple_offset + s.n_tuples >= 0xFFFFFFFFull) return fail(h, HUMID_GPU_ERR_UNSUPPORTED, "more than 2^32-2 tuples");
  return guarded(h, [&] {
    HUMID_CUDA(cudaSetDevice(s.device));
    stage_build_table(s, /*with_filter=*/false);
    s.tuple_base = tuple_offset;
    s.n_local_words = s.unique;
    if (tuple_offset && s.unique)
      add_offset_kernel<<<grid_for(s.unique, 256), 256, 0, s.stream>>>(s.word_first.p, s.unique, (uint32_t)tuple_offset);
    HUMID_CUDA(cudaStreamSynchronize(s.stream));
    s.shard_mode = true;
    *n_local_words = s.unique;
    if (n_usable) *n_usable = s.usable;
  });
}

int humid_gpu_shard_sample_keys(humid_gpu *h, uint32_t n_samples, uint64_t *keys_out) {
  if (int rc = require_shard(h, true)) return rc;
  State &s = h->s;
  if (!keys_out || n_samples == 0 || s.n_local_words == 0) return fail(h, HUMID_GPU_ERR_ARG, "sample_keys: nothing to sample");
  return guarded(h, [&] {
    HUMID_CUDA(cudaSetDevice(s.device));
    uint64_t *tmp = (uint64_t *)s.scratch.get(n_samples * sizeof(uint64_t));
    sample_keys_kernel<<<(n_samples + 255) / 256, 256, 0, s.stream>>>(s.word_key.p, s.n_local_words, n_samples, tmp);
    HUMID_CUDA(cudaMemcpyAsync(keys_out, tmp, n_samples * sizeof(uint64_t), cudaMemcpyDeviceToHost, s.stream));
    HUMID_CUDA(cudaStreamSynchronize(s.stream));
  });
}

int humid_gpu_shard_split(humid_gpu *h, const uint64_t *splitters, uint32_t n_parts, uint64_t *counts) {
  if (int rc = require_shard(h, true)) return rc;
  State &s = h->s;
  if (!counts || n_parts == 0 || n_parts > HUMID_MAX_PARTS || (n_parts > 1 && !splitters))
    return fail(h, HUMID_GPU_ERR_ARG, "split: bad arguments");
  return guarded(h, [&] {
    HUMID_CUDA(cudaSetDevice(s.device));
    Layout L{};
    L.n_parts = n_parts;
    for (uint32_t i = 0; i + 1 < n_parts; i++) L.splitters[i] = splitters[i];
    std::vector<uint64_t> b(n_parts + 1, 0);
    if (n_parts > 1 && s.n_local_words) {
      uint64_t *tmp = (uint64_t *)s.scratch.get(HUMID_MAX_PARTS * sizeof(uint64_t));
      split_kernel<<<1, HUMID_MAX_PARTS, 0, s.stream>>>(s.word_key.p, s.n_local_words, L, tmp);
      HUMID_CUDA(cudaMemcpyAsync(b.data() + 1, tmp, (n_parts - 1) * sizeof(uint64_t), cudaMemcpyDeviceToHost, s.stream));
      HUMID_CUDA(cudaStreamSynchronize(s.stream));
    }
    b[n_parts] = s.n_local_words;
    for (uint32_t p = 0; p < n_parts; p++) counts[p] = b[p + 1] - b[p];
  });
}

int humid_gpu_shard_export_words(humid_gpu *h, uint64_t *d_keys, uint32_t *d_counts, uint32_t *d_first) {
  if (int rc = require_shard(h, true)) return rc;
  State &s = h->s;
  const uint64_t U = s.n_local_words;
  if (U && (!d_keys || !d_counts || !d_first)) return fail(h, HUMID_GPU_ERR_ARG, "export_words: null buffer");
  return guarded(h, [&] {
    HUMID_CUDA(cudaSetDevice(s.device));
    if (U) {
      Timed t(s, "shard_export_words", U, U * 32);
      HUMID_CUDA(cudaMemcpyAsync(d_keys, s.word_key.p, U * 8, cudaMemcpyDeviceToDevice, s.stream));
      HUMID_CUDA(cudaMemcpyAsync(d_counts, s.word_count.p, U * 4, cudaMemcpyDeviceToDevice, s.stream));
      HUMID_CUDA(cudaMemcpyAsync(d_first, s.word_first.p, U * 4, cudaMemcpyDeviceToDevice, s.stream));
      t.launches = 0;
    }
    HUMID_CUDA(cudaStreamSynchronize(s.stream));
  });
}

int humid_gpu_shard_import_words(humid_gpu *h, const uint64_t *d_keys, const uint32_t *d_counts,
                                 const uint32_t *d_first, uint64_t n, uint32_t *d_local_rank, uint64_t *n_shard_words) {
  if (int rc = require_shard(h, true)) return rc;
  State &s = h->s;
  if (!n_shard_words || (n && (!d_keys || !d_counts || !d_first || !d_local_rank)))
    return fail(h, HUMID_GPU_ERR_ARG, "import_words: null buffer");
  if (n >= 0xFFFFFFFFull) return fail(h, HUMID_GPU_ERR_UNSUPPORTED, "too many records");
  return guarded(h, [&] {
    HUMID_CUDA(cudaSetDevice(s.device));
    uint64_t U = 0;
    if (n) {
      Timed t(s, "shard_merge_words", n, n * (2 * 12 + 24));
      s.key_alt.ensure(n + 1);
      s.entries_alt.ensure(n + 1);
      s.word_slot.ensure(n + 1);
      s.slot_alt.ensure(n + 1);
      HUMID_CUDA(cudaMemcpyAsync(s.key_alt.p, d_keys, n * 8, cudaMemcpyDeviceToDevice, s.stream));
      iota32_kernel<<<grid_for(n, 256), 256, 0, s.stream>>>(s.word_slot.p, n);
      uint64_t *kb[2] = {s.key_alt.p, s.entries_alt.p};
      uint32_t *vb[2] = {s.word_slot.p, s.slot_alt.p};
      int launches = 2;
      int cur = radix_sort_u64(kb, vb, n, 2 * (int)s.cfg.word_length, s.scratch, s.scan_scratch, s.stream, &launches);
      s.seed_flag.ensure(n);
      s.seed_index.ensure(n);
      mark_heads_kernel<<<grid_for(n, 256), 256, 0, s.stream>>>(kb[cur], n, s.seed_flag.p);
      uint32_t *total = reinterpret_cast<uint32_t *>(s.counters.p + 6);
      launches += 1 + exclusive_scan_u32(s.seed_flag.p, s.seed_index.p, n, total, s.scan_scratch, s.stream);
      uint32_t heads = 0;
      HUMID_CUDA(cudaMemcpyAsync(&heads, total, sizeof(uint32_t), cudaMemcpyDeviceToHost, s.stream));
      HUMID_CUDA(cudaStreamSynchronize(s.stream));
      U = (uint64_t)heads + 1;
      s.word_key.ensure(U + 1);
      s.word_count.ensure(U + 1);
      s.word_first.ensure(U + 1);
      HUMID_CUDA(cudaMemsetAsync(s.word_count.p, 0, U * 4, s.stream));
      HUMID_CUDA(cudaMemsetAsync(s.word_first.p, 0xFF, U * 4, s.stream));
      merge_words_kernel<<<grid_for(n, 256), 256, 0, s.stream>>>(kb[cur], vb[cur], s.seed_flag.p, s.seed_index.p, n,
                                                                 d_counts, d_first, s.word_key.p, s.word_count.p,
                                                                 s.word_first.p, d_local_rank);
      t.launches = launches + 1;
    }
    s.unique = s.n_shard_words = U;
    // table of the shard's words (key -> shard-local rank) + fingerprint filter
    uint64_t cap = 1024;
    while (cap < 2 * U) cap <<= 1;
    s.shard_cap = cap;
    s.shard_slots.ensure(cap);
    {
      Timed t(s, "shard_word_table", U, cap * 16 + U * 24);
      table_clear2_kernel<<<grid_for(cap, 256), 256, 0, s.stream>>>(s.shard_slots.p, cap);
      if (U) word_table_insert_kernel<<<grid_for(U, 256), 256, 0, s.stream>>>(s.word_key.p, U, s.shard_slots.p, cap - 1);
      t.launches = 2;
    }
    stage_build_filter(s);
    HUMID_CUDA(cudaGetLastError());
    HUMID_CUDA(cudaStreamSynchronize(s.stream));
    *n_shard_words = U;
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

int humid_gpu_shard_set_word_ranks(humid_gpu *h, const uint32_t *d_global_rank, uint64_t n_local_words) {
  if (int rc = require_shard(h, true)) return rc;
  State &s = h->s;
  if (n_local_words != s.n_local_words) return fail(h, HUMID_GPU_ERR_ARG, "set_word_ranks: wrong count");
  return guarded(h, [&] {
    HUMID_CUDA(cudaSetDevice(s.device));
    s.local_word_grank.ensure(n_local_words + 1);
    if (n_local_words)
      HUMID_CUDA(cudaMemcpyAsync(s.local_word_grank.p, d_global_rank, n_local_words * 4, cudaMemcpyDeviceToDevice, s.stream));
    HUMID_CUDA(cudaStreamSynchronize(s.stream));
  });
}

int humid_gpu_shard_probe_count(humid_gpu *h, uint64_t *counts) {
  if (int rc = require_shard(h, true)) return rc;
  State &s = h->s;
  if (!counts) return HUMID_GPU_ERR_ARG;
  return guarded(h, [&] {
    HUMID_CUDA(cudaSetDevice(s.device));
    const uint64_t U = s.n_shard_words;
    const uint32_t n = s.cfg.word_length, d = s.cfg.distance;
    unsigned long long *pc = s.counters.p + 16;
    HUMID_CUDA(cudaMemsetAsync(pc, 0, HUMID_MAX_PARTS * sizeof(unsigned long long), s.stream));
    if (U && d) {
      Timed t(s, "shard_probe_count", U, U * 8);
      shard_probe_kernel<0><<<grid_for(U * n, 256), 256, 0, s.stream>>>(s.word_key.p, U, n, d, layout_of(s), s.lookup(),
                                                                        pc, nullptr, nullptr, nullptr, 0, nullptr);
    }
    HUMID_CUDA(cudaGetLastError());
    unsigned long long host[HUMID_MAX_PARTS];
    HUMID_CUDA(cudaMemcpyAsync(host, pc, sizeof(host), cudaMemcpyDeviceToHost, s.stream));
    HUMID_CUDA(cudaStreamSynchronize(s.stream));
    for (uint32_t p = 0; p < s.n_parts; p++) counts[p] = host[p];
  });
}

int humid_gpu_shard_probe(humid_gpu *h, uint64_t *d_query_keys, uint32_t *d_query_src, const uint64_t *offsets) {
  if (int rc = require_shard(h, true)) return rc;
  State &s = h->s;
  if (!offsets) return HUMID_GPU_ERR_ARG;
  return guarded(h, [&] {
    HUMID_CUDA(cudaSetDevice(s.device));
    const uint64_t U = s.n_shard_words;
    const uint32_t n = s.cfg.word_length, d = s.cfg.distance;
    s.n_entries = 0;
    if (U == 0 || d == 0) return;
    unsigned long long *pc = s.counters.p + 16;
    unsigned long long cursors[HUMID_MAX_PARTS] = {0};
    for (uint32_t p = 0; p < s.n_parts; p++) cursors[p] = offsets[p];
    uint64_t cap = std::max<uint64_t>(s.entries.cap, std::max<uint64_t>(4 * U, 1u << 20));
    for (int attempt = 0; attempt < 2; attempt++) {
      s.entries.ensure(cap);
      cap = s.entries.cap;
      HUMID_CUDA(cudaMemcpyAsync(pc, cursors, sizeof(cursors), cudaMemcpyHostToDevice, s.stream));
      HUMID_CUDA(cudaMemsetAsync(s.counters.p + 2, 0, sizeof(unsigned long long), s.stream));
      {
        double probes = d == 1 ? 1.5 * n : d == 2 ? 1.5 * n + 2.25 * n * (n - 1) / 2.0 : 0.0;
        Timed t(s, "shard_probe", U, (uint64_t)(U * 8.0 * (1.0 + probes)));
        shard_probe_kernel<1><<<grid_for(U * n, 256), 256, 0, s.stream>>>(s.word_key.p, U, n, d, layout_of(s), s.lookup(),
                                                                          pc, d_query_keys, d_query_src, s.entries.p, cap,
                                                                          s.counters.p + 2);
      }
      HUMID_CUDA(cudaGetLastError());
      HUMID_CUDA(cudaMemcpyAsync(&s.n_entries, s.counters.p + 2, sizeof(uint64_t), cudaMemcpyDeviceToHost, s.stream));
      HUMID_CUDA(cudaStreamSynchronize(s.stream));
      if (s.n_entries <= cap) break;
      cap = s.n_entries;
    }
  });
}

int humid_gpu_shard_answer(humid_gpu *h, const uint64_t *d_query_keys, const uint32_t *d_query_src, uint64_t n,
                           uint32_t *d_answers) {
  if (int rc = require_shard(h, true)) return rc;
  State &s = h->s;
  if (n && (!d_query_keys || !d_query_src || !d_answers)) return fail(h, HUMID_GPU_ERR_ARG, "answer: null buffer");
  return guarded(h, [&] {
    HUMID_CUDA(cudaSetDevice(s.device));
    if (n == 0) return;
    s.entries.grow(s.n_entries + n, s.n_entries, s.stream);
    Timed t(s, "shard_answer", n, n * 16);
    shard_answer_kernel<<<grid_for(n, 256), 256, 0, s.stream>>>(d_query_keys, d_query_src, n, layout_of(s), s.lookup(),
                                                                d_answers, s.entries.p, s.entries.cap, s.counters.p + 2);
    HUMID_CUDA(cudaGetLastError());
    HUMID_CUDA(cudaMemcpyAsync(&s.n_entries, s.counters.p + 2, sizeof(uint64_t), cudaMemcpyDeviceToHost, s.stream));
    HUMID_CUDA(cudaStreamSynchronize(s.stream));
  });
}

int humid_gpu_shard_apply_answers(humid_gpu *h, const uint32_t *d_query_src, const uint32_t *d_answers, uint64_t n) {
  if (int rc = require_shard(h, true)) return rc;
  State &s = h->s;
  if (n && (!d_query_src || !d_answers)) return fail(h, HUMID_GPU_ERR_ARG, "apply_answers: null buffer");
  return guarded(h, [&] {
    HUMID_CUDA(cudaSetDevice(s.device));
    if (n == 0) return;
    s.entries.grow(s.n_entries + n, s.n_entries, s.stream);
    Timed t(s, "shard_apply_answers", n, n * 8);
    shard_apply_kernel<<<grid_for(n, 256), 256, 0, s.stream>>>(d_query_src, d_answers, n, layout_of(s), s.entries.p,
                                                               s.entries.cap, s.counters.p + 2);
    HUMID_CUDA(cudaGetLastError());
    HUMID_CUDA(cudaMemcpyAsync(&s.n_entries, s.counters.p + 2, sizeof(uint64_t), cudaMemcpyDeviceToHost, s.stream));
    HUMID_CUDA(cudaStreamSynchronize(s.stream));
  });
}

int humid_gpu_shard_build_rows(humid_gpu *h, uint64_t *n_entries) {
  if (int rc = require_shard(h, true)) return rc;
  State &s = h->s;
  return guarded(h, [&] {
    HUMID_CUDA(cudaSetDevice(s.device));
    stage_csr(s);  // rows: the shard's words; sort key (row << 32 | global column)
    HUMID_CUDA(cudaStreamSynchronize(s.stream));
    if (n_entries) *n_entries = s.n_entries;
  });
}

int humid_gpu_shard_export_graph(humid_gpu *h, uint64_t *d_keys, uint32_t *d_counts, uint32_t *d_first,
                                 uint32_t *d_degree, uint32_t *d_col) {
  if (int rc = require_shard(h, true)) return rc;
  State &s = h->s;
  const uint64_t U = s.n_shard_words, E = s.n_entries;
  return guarded(h, [&] {
    HUMID_CUDA(cudaSetDevice(s.device));
    if (U) {
      HUMID_CUDA(cudaMemcpyAsync(d_keys, s.word_key.p, U * 8, cudaMemcpyDeviceToDevice, s.stream));
      HUMID_CUDA(cudaMemcpyAsync(d_counts, s.word_count.p, U * 4, cudaMemcpyDeviceToDevice, s.stream));
      HUMID_CUDA(cudaMemcpyAsync(d_first, s.word_first.p, U * 4, cudaMemcpyDeviceToDevice, s.stream));
      degree_kernel<<<grid_for(U, 256), 256, 0, s.stream>>>(s.row_ptr.p, U, d_degree);
    }
    if (E) HUMID_CUDA(cudaMemcpyAsync(d_col, s.col.p, E * 4, cudaMemcpyDeviceToDevice, s.stream));
    HUMID_CUDA(cudaGetLastError());
    HUMID_CUDA(cudaStreamSynchronize(s.stream));
  });
}

int humid_gpu_shard_import_graph(humid_gpu *h, uint64_t n_words, uint64_t n_entries, uint64_t n_tuples,
                                 uint64_t n_usable, const uint64_t *d_keys, const uint32_t *d_counts,
                                 const uint32_t *d_first, const uint32_t *d_degree, const uint32_t *d_col) {
  if (int rc = require_shard(h, true)) return rc;
  State &s = h->s;
  if (n_words >= 0xFFFFFFFFull || n_entries >= 0xFFFFFFFFull || n_tuples >= 0xFFFFFFFFull)
    return fail(h, HUMID_GPU_ERR_UNSUPPORTED, "job too large for 32-bit ranks");
  return guarded(h, [&] {
    HUMID_CUDA(cudaSetDevice(s.device));
    const uint64_t U = n_words, E = n_entries;
    s.word_key.ensure(U + 1);
    s.word_count.ensure(U + 1);
    s.word_first.ensure(U + 1);
    s.row_ptr.ensure(U + 2);
    s.col.ensure(E + 1);
    {
      Timed t(s, "shard_import_graph", U, U * 28 + E * 8);
      if (U) {
        HUMID_CUDA(cudaMemcpyAsync(s.word_key.p, d_keys, U * 8, cudaMemcpyDeviceToDevice, s.stream));
        HUMID_CUDA(cudaMemcpyAsync(s.word_count.p, d_counts, U * 4, cudaMemcpyDeviceToDevice, s.stream));
        HUMID_CUDA(cudaMemcpyAsync(s.word_first.p, d_first, U * 4, cudaMemcpyDeviceToDevice, s.stream));
      }
      if (E) HUMID_CUDA(cudaMemcpyAsync(s.col.p, d_col, E * 4, cudaMemcpyDeviceToDevice, s.stream));
      s.seed_index.ensure(U + 1);
      int launches = 0;
      if (U) launches += exclusive_scan_u32(d_degree, s.seed_index.p, U, nullptr, s.scan_scratch, s.stream);
      widen_rows_kernel<<<grid_for(U + 1, 256), 256, 0, s.stream>>>(s.seed_index.p, U, E, s.row_ptr.p);
      t.launches = launches + 1;
    }
    s.unique = U;
    s.n_entries = E;
    s.report_tuples = n_tuples;
    s.report_usable = n_usable;
    stage_cluster(s);
    stage_keep(s);
    HUMID_CUDA(cudaGetLastError());
    HUMID_CUDA(cudaStreamSynchronize(s.stream));
    resolve_stats(s);
    s.finalized = true;
  });
}

int humid_gpu_shard_cluster_ids(humid_gpu *h, uint64_t first_local_tuple, uint64_t n, uint32_t *ids) {
  if (int rc = require_shard(h, true)) return rc;
  State &s = h->s;
  if (!s.finalized) return fail(h, HUMID_GPU_ERR_ARG, "not finalized");
  if (first_local_tuple + n > s.n_tuples) return fail(h, HUMID_GPU_ERR_ARG, "tuple range out of bounds");
  if (n == 0) return HUMID_GPU_OK;
  if (!ids) return HUMID_GPU_ERR_ARG;
  return guarded(h, [&] {
    HUMID_CUDA(cudaSetDevice(s.device));
    s.out_ids.ensure(n);
    {
      Timed t(s, "cluster_ids", n, n * 40);
      shard_cluster_ids_kernel<<<grid_for(n, 256), 256, 0, s.stream>>>(s.tuple_key.p, first_local_tuple, n, s.slots.p,
                                                                       s.table_cap - 1, s.local_word_grank.p, s.label.p,
                                                                       s.seed_index.p, s.out_ids.p);
    }
    HUMID_CUDA(cudaGetLastError());
    HUMID_CUDA(cudaMemcpyAsync(ids, s.out_ids.p, n * sizeof(uint32_t), cudaMemcpyDeviceToHost, s.stream));
    HUMID_CUDA(cudaStreamSynchronize(s.stream));
    resolve_stats(s);
  });
}

}  // extern "C"
