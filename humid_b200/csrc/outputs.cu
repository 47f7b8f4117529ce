// K5 + statistics — per-tuple results and the three histograms.
//
// Replaces the second pass of the reference: Trie::find + the visited/maxLeaf test of
// writeFiltered (src/humid.cc:220-234), the id lookup of writeAnnotated (:268-285),
// runStatistics (:301-315) and clusterStats (src/cluster.cc:89-95).
#include <algorithm>

#include "state.cuh"

namespace humid {

// writeFiltered keeps, per cluster, the FIRST tuple in input order whose word is the
// cluster's maxLeaf; the table recorded that tuple index at insert time (word_first), so
// the keep set is one scatter per cluster instead of a lookup per tuple.
__global__ void __launch_bounds__(256)
    keep_scatter_kernel(const uint32_t *__restrict__ cluster_max, const uint32_t *__restrict__ word_first, uint64_t C,
                        uint32_t *__restrict__ bits) {
  uint64_t c = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  uint32_t t = word_first[cluster_max[c]];
  atomicOr(bits + (t >> 5), 1u << (t & 31u));
}

int stage_keep(State &s) {
  uint64_t words = (s.report_tuples + 31) / 32 + 2;  // +2: the D2H view is in 64-bit units
  s.keep_bits.ensure(words);
  HUMID_CUDA(cudaMemsetAsync(s.keep_bits.p, 0, words * sizeof(uint32_t), s.stream));
  if (s.n_clusters == 0) return 0;
  Timed t(s, "keep", s.n_clusters, s.n_clusters * 12);
  keep_scatter_kernel<<<grid_for(s.n_clusters, 256), 256, 0, s.stream>>>(s.cluster_max.p, s.word_first.p,
                                                                         s.n_clusters, s.keep_bits.p);
  HUMID_CUDA(cudaGetLastError());
  return 1;
}

// writeAnnotated: cluster id of the tuple's word, 0 for filtered tuples.
__global__ void __launch_bounds__(256)
    cluster_ids_kernel(const uint64_t *__restrict__ tuple_key, uint64_t first, uint64_t n,
                       const Slot *__restrict__ slots, uint64_t mask, const uint32_t *__restrict__ label,
                       const uint32_t *__restrict__ seed_index, uint32_t *__restrict__ out) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint64_t key = ld_stream_u64(tuple_key + first + i);
  uint32_t id = 0;
  if (key != KEY_INVALID) {
    uint64_t idx = mix64(key) & mask;
    while (true) {
      uint4 v = __ldg(reinterpret_cast<const uint4 *>(slots + idx));
      uint64_t k = ((uint64_t)v.y << 32) | v.x;
      if (k == key) { id = seed_index[label[v.z]] + 1u; break; }
      idx = (idx + 1) & mask;  // the key is always present (it was inserted in pass 1)
    }
  }
  out[i] = id;
}

int stage_cluster_ids(State &s, uint64_t first, uint64_t n, uint32_t *d_out) {
  if (n == 0) return 0;
  if (!s.ranks_in_table) stage_ensure_lookup(s);  // slot.count -> rank
  // K5: key (8) + slot (16) + label (4) + seed index (4) + id out (4)
  Timed t(s, "cluster_ids", n, n * 36);
  cluster_ids_kernel<<<grid_for(n, 256), 256, 0, s.stream>>>(s.tuple_key.p, first, n, s.slots.p, s.table_cap - 1,
                                                             s.label.p, s.seed_index.p, d_out);
  HUMID_CUDA(cudaGetLastError());
  return 1;
}

// ---- histograms: sort the values, emit (value, first index) at every run start ------------
__global__ void hist_values_kernel(int which, uint64_t m, const uint32_t *__restrict__ word_count,
                                   const uint64_t *__restrict__ row_ptr, const uint64_t *__restrict__ cluster_size,
                                   uint64_t *__restrict__ out) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m) return;
  out[i] = which == HUMID_HIST_COUNTS ? (uint64_t)word_count[i]
           : which == HUMID_HIST_NEIGHBOURS ? row_ptr[i + 1] - row_ptr[i]
                                            : cluster_size[i];
}

__global__ void __launch_bounds__(256)
    run_starts_kernel(const uint64_t *__restrict__ sorted, uint64_t m, uint64_t *__restrict__ run_val,
                      uint64_t *__restrict__ run_start, uint64_t cap, unsigned long long *__restrict__ counter) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m) return;
  uint64_t v = sorted[i];
  if (i == 0 || sorted[i - 1] != v) {
    const unsigned long long pos = reserve_slots(counter);
    if (pos < cap) {
      run_val[pos] = v;
      run_start[pos] = i;
    }
  }
}

int stage_histogram(State &s, int which) {
  const uint64_t m = which == HUMID_HIST_CLUSTERS ? s.n_clusters : s.unique;
  s.hist_k[which].clear();
  s.hist_v[which].clear();
  s.hist_done[which] = true;
  if (m == 0) return 0;
  int k = 0;
  static const char *names[3] = {"hist_counts", "hist_neigh", "hist_clusters"};
  Timed t(s, names[which], m, m * 8);
  s.key_alt.ensure(m + 1);
  s.entries_alt.ensure(m + 1);
  uint64_t *kb[2] = {s.key_alt.p, s.entries_alt.p};
  uint32_t *vb[2] = {nullptr, nullptr};
  hist_values_kernel<<<grid_for(m, 256), 256, 0, s.stream>>>(which, m, s.word_count.p, s.row_ptr.p, s.cluster_size.p,
                                                             kb[0]);
  k++;
  // counts and cluster sizes are bounded by the number of tuples, list lengths by the entries
  uint64_t bound = which == HUMID_HIST_NEIGHBOURS ? s.n_entries : s.report_tuples;
  int bits = 1;
  while (bits < 64 && (1ull << bits) <= bound) bits++;
  int launches = 0;
  int cur = radix_sort_u64(kb, vb, m, bits, s.scratch, s.scan_scratch, s.stream, &launches);
  k += launches;
  uint64_t cap = 1 << 16;
  std::vector<uint64_t> hv, hs;
  for (int attempt = 0; attempt < 2; attempt++) {
    uint64_t *runs = (uint64_t *)s.scratch.get(2 * cap * sizeof(uint64_t));
    HUMID_CUDA(cudaMemsetAsync(s.counters.p + 4, 0, sizeof(unsigned long long), s.stream));
    run_starts_kernel<<<grid_for(m, 256), 256, 0, s.stream>>>(kb[cur], m, runs, runs + cap, cap, s.counters.p + 4);
    k++;
    uint64_t n_runs = 0;
    HUMID_CUDA(cudaMemcpyAsync(&n_runs, s.counters.p + 4, sizeof(uint64_t), cudaMemcpyDeviceToHost, s.stream));
    HUMID_CUDA(cudaStreamSynchronize(s.stream));
    if (n_runs > cap) { cap = n_runs; continue; }
    hv.resize(n_runs);
    hs.resize(n_runs);
    HUMID_CUDA(cudaMemcpyAsync(hv.data(), runs, n_runs * sizeof(uint64_t), cudaMemcpyDeviceToHost, s.stream));
    HUMID_CUDA(cudaMemcpyAsync(hs.data(), runs + cap, n_runs * sizeof(uint64_t), cudaMemcpyDeviceToHost, s.stream));
    HUMID_CUDA(cudaStreamSynchronize(s.stream));
    break;
  }
  // runs arrive in arbitrary order: order them by start index, lengths are the differences
  std::vector<size_t> order(hv.size());
  for (size_t i = 0; i < order.size(); i++) order[i] = i;
  std::sort(order.begin(), order.end(), [&](size_t a, size_t b) { return hs[a] < hs[b]; });
  for (size_t j = 0; j < order.size(); j++) {
    uint64_t end = j + 1 < order.size() ? hs[order[j + 1]] : m;
    s.hist_k[which].push_back(hv[order[j]]);
    s.hist_v[which].push_back(end - hs[order[j]]);
  }
  t.launches = k;
  return k;
}

}  // namespace humid
