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

// ---- histograms: values below HIST_SMALL are binned in shared memory per CTA, the few larger ones
// are listed and counted on the host ---------------------------------------------------------
constexpr uint32_t HIST_SMALL = 4096;
constexpr int HIST_ITEMS = 16;
__global__ void __launch_bounds__(256)
    hist_small_kernel(int which, uint64_t m, const uint32_t *__restrict__ word_count,
                      const uint64_t *__restrict__ row_ptr, const uint64_t *__restrict__ cluster_size,
                      unsigned long long *__restrict__ bins, uint64_t *__restrict__ large, uint64_t large_cap,
                      unsigned long long *__restrict__ n_large) {
  __shared__ uint32_t h[HIST_SMALL];
  for (uint32_t k = threadIdx.x; k < HIST_SMALL; k += 256) h[k] = 0;
  __syncthreads();
  const uint64_t base = (uint64_t)blockIdx.x * (256 * HIST_ITEMS);
#pragma unroll 4
  for (int j = 0; j < HIST_ITEMS; j++) {
    const uint64_t i = base + (uint64_t)j * 256 + threadIdx.x;
    if (i >= m) break;
    const uint64_t v = which == HUMID_HIST_COUNTS ? (uint64_t)word_count[i]
                       : which == HUMID_HIST_NEIGHBOURS ? row_ptr[i + 1] - row_ptr[i]
                                                        : cluster_size[i];
    if (v < HIST_SMALL) {
      atomicAdd(&h[(uint32_t)v], 1u);
    } else {
      const unsigned long long pos = reserve_slots(n_large);
      if (pos < large_cap) large[pos] = v;
    }
  }
  __syncthreads();
  for (uint32_t k = threadIdx.x; k < HIST_SMALL; k += 256)
    if (h[k]) atomicAdd(bins + k, (unsigned long long)h[k]);
}

// bins + listed large values of the three histograms -> host vectors in ascending key order (std::map order of the
// reference, src/humid.cc:328-349)
static void hist_to_vectors(State &s, int which, const unsigned long long *bins, uint64_t *large, uint64_t n_large) {
  s.hist_k[which].clear();
  s.hist_v[which].clear();
  for (uint32_t v = 0; v < HIST_SMALL; v++)
    if (bins[v]) {
      s.hist_k[which].push_back(v);
      s.hist_v[which].push_back(bins[v]);
    }
  std::sort(large, large + n_large);
  for (uint64_t i = 0; i < n_large;) {
    uint64_t j = i;
    while (j < n_large && large[j] == large[i]) j++;
    s.hist_k[which].push_back(large[i]);
    s.hist_v[which].push_back(j - i);
    i = j;
  }
  s.hist_done[which] = true;
}

static const char *const HIST_NAMES[3] = {"hist_counts", "hist_neigh", "hist_clusters"};

// runStatistics + clusterStats as part of finalize: the three kernels are enqueued behind the clustering, their bins
// (and up to HIST_EAGER_LARGE listed values each) travel to page-locked host memory with the keep bitmap, and
// stage_histograms_collect turns them into the (key, value) lists once finalize has synchronised — no extra host
// round trip.  A histogram with more large values than that is recomputed on demand (stage_histogram).
constexpr uint64_t HIST_EAGER_LARGE = 1 << 13;
constexpr uint64_t HIST_EAGER_STRIDE = HIST_SMALL + HIST_EAGER_LARGE + 1;  // bins, n_large, large values

int stage_histograms_enqueue(State &s) {
  int k = 0;
  s.hist_dev.ensure(3 * HIST_EAGER_STRIDE);
  s.hist_host.ensure(3 * HIST_EAGER_STRIDE);
  HUMID_CUDA(cudaMemsetAsync(s.hist_dev.p, 0, 3 * HIST_EAGER_STRIDE * sizeof(unsigned long long), s.stream));
  for (int which = 0; which < 3; which++) {
    s.hist_done[which] = false;
    const uint64_t m = which == HUMID_HIST_CLUSTERS ? s.n_clusters : s.unique;
    if (m == 0) continue;
    unsigned long long *d = s.hist_dev.p + which * HIST_EAGER_STRIDE;
    Timed t(s, HIST_NAMES[which], m, m * (which == HUMID_HIST_COUNTS ? 4 : 8));
    hist_small_kernel<<<grid_for(m, 256, HIST_ITEMS), 256, 0, s.stream>>>(
        which, m, s.word_count.p, s.row_ptr.p, s.cluster_size.p, d, reinterpret_cast<uint64_t *>(d + HIST_SMALL + 1),
        HIST_EAGER_LARGE, d + HIST_SMALL);
    k++;
  }
  HUMID_CUDA(cudaGetLastError());
  HUMID_CUDA(cudaMemcpyAsync(s.hist_host.p, s.hist_dev.p, 3 * HIST_EAGER_STRIDE * sizeof(unsigned long long),
                             cudaMemcpyDeviceToHost, s.stream));
  s.hist_pending = true;
  return k;
}

void stage_histograms_collect(State &s) {
  if (!s.hist_pending) return;
  s.hist_pending = false;
  for (int which = 0; which < 3; which++) {
    unsigned long long *hst = s.hist_host.p + which * HIST_EAGER_STRIDE;
    const uint64_t n_large = hst[HIST_SMALL];
    if (n_large > HIST_EAGER_LARGE) continue;  // stays !hist_done: recomputed when asked for
    hist_to_vectors(s, which, hst, reinterpret_cast<uint64_t *>(hst + HIST_SMALL + 1), n_large);
  }
}

int stage_histogram(State &s, int which) {
  const uint64_t m = which == HUMID_HIST_CLUSTERS ? s.n_clusters : s.unique;
  s.hist_k[which].clear();
  s.hist_v[which].clear();
  s.hist_done[which] = true;
  if (m == 0) return 0;
  int k = 0;
  Timed t(s, HIST_NAMES[which], m, m * (which == HUMID_HIST_COUNTS ? 4 : 8));
  std::vector<unsigned long long> bins(HIST_SMALL);
  std::vector<uint64_t> large;
  uint64_t cap = 1 << 16;
  for (int attempt = 0; attempt < 2; attempt++) {
    // scratch: [bins: HIST_SMALL u64][large values: cap u64]
    unsigned long long *d_bins = (unsigned long long *)s.scratch.get((HIST_SMALL + cap) * sizeof(uint64_t));
    uint64_t *d_large = (uint64_t *)(d_bins + HIST_SMALL);
    HUMID_CUDA(cudaMemsetAsync(d_bins, 0, HIST_SMALL * sizeof(uint64_t), s.stream));
    HUMID_CUDA(cudaMemsetAsync(s.counters.p + 4, 0, sizeof(unsigned long long), s.stream));
    hist_small_kernel<<<grid_for(m, 256, HIST_ITEMS), 256, 0, s.stream>>>(which, m, s.word_count.p, s.row_ptr.p,
                                                                         s.cluster_size.p, d_bins, d_large, cap,
                                                                         s.counters.p + 4);
    k++;
    uint64_t n_large = 0;
    HUMID_CUDA(cudaMemcpyAsync(&n_large, s.counters.p + 4, sizeof(uint64_t), cudaMemcpyDeviceToHost, s.stream));
    HUMID_CUDA(cudaMemcpyAsync(bins.data(), d_bins, HIST_SMALL * sizeof(uint64_t), cudaMemcpyDeviceToHost, s.stream));
    HUMID_CUDA(cudaStreamSynchronize(s.stream));
    if (n_large > cap) { cap = n_large; continue; }
    large.resize(n_large);
    if (n_large) {
      HUMID_CUDA(cudaMemcpyAsync(large.data(), d_large, n_large * sizeof(uint64_t), cudaMemcpyDeviceToHost, s.stream));
      HUMID_CUDA(cudaStreamSynchronize(s.stream));
    }
    break;
  }
  hist_to_vectors(s, which, bins.data(), large.data(), large.size());
  t.launches = k;
  return k;
}

}  // namespace humid
