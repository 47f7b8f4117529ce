// K4 — clustering.  Replaces findClusters (reference src/humid.cc:167-193) with
// assignMaxCluster / assignDirectionalCluster (src/cluster.cc:39-87).
//
// The reference walks the words in rank order and grows one cluster per still-unassigned
// word by recursive traversal of the neighbour lists; the result depends on walk order and
// on neighbour-list order, but never crosses a connected component of the neighbour graph.
// So: (1) lock-free union-find gives the components (root = smallest rank = the word the
// walk meets first); (2) the members of a component are needed in rank order: components of
// up to SMALL_COMP words (virtually all of them) are collected by their root's thread with a
// breadth-first walk of the lists and sorted in local memory, only the words of larger
// components go through a global sort; (3) every component is replayed exactly,
// independently of all others; (4) cluster ids are an
// exclusive scan over the seed flags in rank order (ids are handed out in walk order of
// the seeding word, src/humid.cc:180).
#include <algorithm>
#include <cstdlib>

#include "cluster.cuh"

namespace humid {

// ---- (1) components -----------------------------------------------------------------------
__global__ void iota_kernel(uint32_t *a, uint64_t n) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) a[i] = (uint32_t)i;
}

__device__ __forceinline__ uint32_t uf_find(uint32_t *parent, uint32_t x) {
  // path halving; parents only ever decrease, so racing writers stay consistent
  while (true) {
    uint32_t p = parent[x];
    if (p == x) return x;
    uint32_t g = parent[p];
    if (g != p) parent[x] = g;
    x = g;
  }
}

__global__ void __launch_bounds__(256)
    uf_union_kernel(const uint64_t *__restrict__ row_ptr, const uint32_t *__restrict__ col, uint64_t E, uint64_t U,
                    uint32_t base, uint32_t *__restrict__ open_flag, uint32_t *parent) {
  // one thread per directed entry; the source rank is recovered lazily only for a < b
  // entries via binary search would be wasteful, so sources come from a per-row loop instead
  uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= U) return;
  for (uint64_t e = row_ptr[r]; e < row_ptr[r + 1]; e++) {
    uint32_t b = col[e] - base;
    if (b >= U) {  // shard mode: the neighbour lives on another shard
      if (open_flag) open_flag[r] = 1u;
      continue;
    }
    if (b <= r) continue;  // every undirected pair is present in both rows; take it once
    uint32_t x = (uint32_t)r, y = b;
    while (true) {
      x = uf_find(parent, x);
      y = uf_find(parent, y);
      if (x == y) break;
      if (x > y) { uint32_t t = x; x = y; y = t; }
      // hook the larger root under the smaller one => the final root is the minimum rank
      if (atomicCAS(parent + y, y, x) == y) break;
    }
  }
}

// The same over the undirected pair list the Hamming join leaves behind (one thread per pair: every lane has an
// edge, where the row form above keeps a lane per word and two words in three have no neighbour).
__global__ void __launch_bounds__(256)
    uf_union_pairs_kernel(const uint64_t *__restrict__ pairs, uint64_t P, uint32_t *parent) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P) return;
  const uint64_t e = ld_stream_u64(pairs + i);
  uint32_t x = (uint32_t)(e >> 32), y = (uint32_t)e;
  while (true) {
    x = uf_find(parent, x);
    y = uf_find(parent, y);
    if (x == y) break;
    if (x > y) { uint32_t t = x; x = y; y = t; }
    if (atomicCAS(parent + y, y, x) == y) break;
  }
}

// The same over a list of DIRECTED entries (local row << 32 | column + base): shard mode, where a column outside
// [0, U) belongs to a word of another shard and marks the row's component as open.
__global__ void __launch_bounds__(256)
    uf_union_entries_kernel(const uint64_t *__restrict__ entries, uint64_t E, uint64_t U, uint32_t base,
                            uint32_t *__restrict__ open_flag, uint32_t *parent) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= E) return;
  const uint64_t e = ld_stream_u64(entries + i);
  uint32_t x = (uint32_t)(e >> 32), y = (uint32_t)e - base;
  if (y >= U) {
    if (open_flag) open_flag[x] = 1u;
    return;
  }
  if (y <= x) return;  // every undirected pair is present in both directions; take it once
  while (true) {
    x = uf_find(parent, x);
    y = uf_find(parent, y);
    if (x == y) break;
    if (x > y) { uint32_t t = x; x = y; y = t; }
    if (atomicCAS(parent + y, y, x) == y) break;
  }
}

// comp[r] = root; csize[root] = number of members (singletons are not counted: csize stays 0)
__global__ void uf_flatten_kernel(uint32_t *parent, uint32_t *comp, uint32_t *csize, uint64_t U,
                                  const uint32_t *__restrict__ open_flag, uint32_t *__restrict__ comp_open) {
  uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= U) return;
  uint32_t root = uf_find(parent, (uint32_t)r);
  comp[r] = root;
  if (open_flag && open_flag[r]) comp_open[root] = 1u;
  if (root != r) {
    atomicAdd(csize + root, 1u);  // members besides the root itself
  }
}

// ---- (2) every word by the size of its component ---------------------------------------------
// singletons are closed in place; the roots of components of 2..SMALL_COMP words go to a dense list (one busy thread
// per lane in the replay, where one thread per WORD left 27 of 32 lanes idle); the words of larger components are
// listed as (root, rank) keys and grouped by a sort.
// Components this pass leaves alone: (shard mode) those that continue on another shard, and — when the open components
// are shared out among the ranks of a group — those whose root another rank owns (root mod `mod` != `me`).
struct RootSkip {
  const uint32_t *flags;
  uint32_t mod, me;
  __device__ __forceinline__ bool operator()(uint32_t root) const {
    return (flags && flags[root]) || (mod && root % mod != me);
  }
};

template <bool MAXIMUM>
__global__ void __launch_bounds__(256)
    classify_words_kernel(const uint32_t *__restrict__ comp, const uint32_t *__restrict__ csize,
                          const uint32_t *__restrict__ count, uint64_t U, uint32_t small_max,
                          const RootSkip skip_root, uint32_t *__restrict__ label,
                          uint32_t *__restrict__ seed_flag, uint64_t *__restrict__ seed_size,
                          uint32_t *__restrict__ seed_max, uint32_t *__restrict__ small_roots,
                          unsigned long long *__restrict__ n_small, unsigned long long *__restrict__ n_big_roots) {
  const uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  bool small = false, big = false;
  if (r < U) {
    const uint32_t root = comp[r];
    // only roots look at their component (csize[r]: coalesced); the other words are listed later, and only when
    // some component is large
    uint32_t lab = MAXIMUM ? root : NONE32, flag = 0;
    if (root == r && !skip_root(root)) {
      const uint32_t m = csize[r] + 1u;
      if (MAXIMUM) flag = 1;
      if (m == 1) {  // no neighbours: its own cluster
        lab = (uint32_t)r;
        flag = 1;
        seed_size[r] = count[r];
        seed_max[r] = (uint32_t)r;
      } else if (m <= small_max) {
        small = true;
      } else {
        big = true;
      }
    }
    label[r] = lab;
    seed_flag[r] = flag;
  }
  // dense list of the small components' roots: one cursor bump per CTA (a cursor is one address, and returning
  // atomics on one address complete at well under 1 G/s: one per warp made this kernel 0.5 ms long)
  __shared__ uint32_t warp_small[8], warp_big[8];
  __shared__ unsigned long long cta_base;
  const uint32_t lane = lane_id(), lt = (1u << lane) - 1u, w = threadIdx.x >> 5;
  const unsigned ms = __ballot_sync(0xffffffffu, small), mb = __ballot_sync(0xffffffffu, big);
  if (lane == 0) {
    warp_small[w] = __popc(ms);
    warp_big[w] = __popc(mb);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t run = 0, nb = 0;
    for (int k = 0; k < 8; k++) {
      const uint32_t c = warp_small[k];
      warp_small[k] = run;
      run += c;
      nb += warp_big[k];
    }
    cta_base = run ? atomicAdd(n_small, (unsigned long long)run) : 0ull;
    if (nb) atomicAdd(n_big_roots, (unsigned long long)nb);
  }
  __syncthreads();
  if (small) small_roots[cta_base + warp_small[w] + __popc(ms & lt)] = (uint32_t)r;
}

// (root, rank) keys of the words of the large components (run only when there is one)
__global__ void __launch_bounds__(256)
    big_member_keys_kernel(const uint32_t *__restrict__ comp, const uint32_t *__restrict__ csize, uint64_t U,
                           uint32_t small_max, const RootSkip skip_root, uint64_t *__restrict__ keys,
                           unsigned long long *__restrict__ counter) {
  uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  bool big = false;
  uint32_t root = 0;
  if (r < U) {
    root = comp[r];
    big = csize[root] + 1u > small_max && !skip_root(root);
  }
  unsigned m = __ballot_sync(0xffffffffu, big);
  if (m == 0) return;
  unsigned long long base = 0;
  int leader = __ffs(m) - 1;
  if ((int)lane_id() == leader) base = atomicAdd(counter, (unsigned long long)__popc(m));
  base = __shfl_sync(0xffffffffu, base, leader);
  if (big) keys[base + __popc(m & ((1u << lane_id()) - 1u))] = ((uint64_t)root << 32) | (uint32_t)r;
}

__global__ void big_segments_kernel(const uint64_t *__restrict__ sorted, uint64_t n, uint32_t *__restrict__ members,
                                    uint32_t *__restrict__ seg_start) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint32_t c = (uint32_t)(sorted[i] >> 32);
  members[i] = (uint32_t)sorted[i];
  if (i == 0 || (uint32_t)(sorted[i - 1] >> 32) != c) seg_start[c] = (uint32_t)i;
}

// Members of the component of `root` (m of them) in rank order, collected by a breadth-first
// walk of the neighbour lists into `arr` (capacity SMALL_COMP) and sorted there.
__device__ __forceinline__ void gather_small_component(const Graph &g, uint32_t root, uint32_t m, uint32_t *arr) {
  arr[0] = root;
  uint32_t cnt = 1;
  for (uint32_t i = 0; i < cnt && cnt < m; i++) {
    const uint32_t u = arr[i];
    for (uint64_t e = g.row_ptr[u], hi = g.row_ptr[u + 1]; e < hi && cnt < m; e++) {
      const uint32_t v = gcol(g, e);
      bool seen = false;
      for (uint32_t j = 0; j < cnt; j++) seen |= arr[j] == v;
      if (!seen) arr[cnt++] = v;
    }
  }
  for (uint32_t i = 1; i < cnt; i++) {
    uint32_t x = arr[i], j = i;
    while (j > 0 && arr[j - 1] > x) { arr[j] = arr[j - 1]; j--; }
    arr[j] = x;
  }
}

// position of word v inside the sorted member array (the per-member work arrays are indexed by it)
__device__ __forceinline__ uint32_t member_index(const uint32_t *mem, uint32_t m, uint32_t v) {
  uint32_t lo = 0, hi = m;
  while (lo < hi) {
    uint32_t mid = (lo + hi) >> 1;
    if (mem[mid] < v) lo = mid + 1; else hi = mid;
  }
  return lo;
}

// ---- (3) exact replay per component ------------------------------------------------------
// assignMaxCluster (src/cluster.cc:72-80): cluster == component, size = sum of counts,
// maxLeaf = first word in DFS pre-order (neighbour-list order) holding the strictly largest
// count.  The DFS is only replayed when several members tie for the largest count.
// mem[0..m) = members in rank order (mem[0] == root); stk_* have room for m frames.
__device__ void replay_maximum(const Graph &g, const uint32_t *mem, uint32_t m, uint32_t *stk_node,
                               uint32_t *stk_edge, uint32_t *visited, uint64_t *size_out, uint32_t *max_out) {
  const uint32_t root = mem[0];
  uint64_t size = 0;
  uint32_t best = 0, best_rank = NONE32, ties = 0;
  for (uint32_t i = 0; i < m; i++) {
    const uint32_t w = mem[i];
    const uint32_t c = g.count[w];
    size += c;
    if (c > best) { best = c; best_rank = w; ties = 1; }
    else if (c == best) ties++;
  }
  if (ties > 1) {
    if (g.count[root] == best) {
      best_rank = root;
    } else {
      uint32_t sp = 0;
      visited[root] = 1;
      stk_node[sp] = root; stk_edge[sp] = 0; sp++;
      bool found = false;
      while (sp > 0 && !found) {
        const uint32_t u = stk_node[sp - 1];
        const uint64_t lo = g.row_ptr[u], hi = g.row_ptr[u + 1];
        const uint64_t j = lo + stk_edge[sp - 1];
        if (j < hi) {
          stk_edge[sp - 1]++;
          const uint32_t v = gcol(g, j);
          if (!visited[v]) {
            visited[v] = 1;
            if (g.count[v] == best) { best_rank = v; found = true; }
            else { stk_node[sp] = v; stk_edge[sp] = 0; sp++; }
          }
        } else {
          sp--;
        }
      }
    }
  }
  *size_out = size;
  *max_out = best_rank;
}

// assignDirectionalCluster (src/cluster.cc:82-87): for every still-unassigned member in rank
// order: climb (maxNeighbour_, :39-51) to the first-in-list unassigned neighbour with
// count >= 2x, restarting the scan after every move; the top word is maxLeaf; then take the
// closure over unassigned neighbours with count <= half (assignDirectionalCluster_, :58-69;
// a closure, so breadth-first order gives the same set as the reference's recursion).
__device__ void replay_directional(const Graph &g, const uint32_t *mem, uint32_t m, uint32_t *queue, uint32_t *label,
                                   uint32_t *seed_flag, uint64_t *seed_size, uint32_t *seed_max) {
  for (uint32_t i = 0; i < m; i++) {
    const uint32_t seed = mem[i];
    if (label[seed] != NONE32) continue;
    uint32_t node = seed;
    uint64_t j = g.row_ptr[node], hi = g.row_ptr[node + 1];
    uint64_t cn = g.count[node];
    while (j < hi) {
      const uint32_t nb = gcol(g, j++);
      if (label[nb] == NONE32 && (uint64_t)g.count[nb] >= 2 * cn) {
        node = nb;
        j = g.row_ptr[node];
        hi = g.row_ptr[node + 1];
        cn = g.count[node];
      }
    }
    uint32_t head = 0, tail = 0;
    label[node] = seed;
    uint64_t size = cn;
    queue[tail++] = node;
    while (head < tail) {
      const uint32_t u = queue[head++];
      const uint64_t cu = g.count[u];
      for (uint64_t k = g.row_ptr[u], e = g.row_ptr[u + 1]; k < e; k++) {
        const uint32_t nb = gcol(g, k);
        if (label[nb] == NONE32) {
          const uint64_t cb = g.count[nb];
          if (cu >= 2 * cb) {
            label[nb] = seed;
            size += cb;
            queue[tail++] = nb;
          }
        }
      }
    }
    seed_flag[seed] = 1;
    seed_size[seed] = size;
    seed_max[seed] = node;
  }
}

// One thread per component of 2..SMALL_COMP words (dense list of their roots): members gathered and replayed in
// local memory.
template <bool MAXIMUM>
#ifndef HUMID_REPLAY_THREADS
#define HUMID_REPLAY_THREADS 128
#endif
__global__ void __launch_bounds__(HUMID_REPLAY_THREADS)
    replay_small_kernel(Graph g, const uint32_t *__restrict__ roots, uint64_t n_roots, const uint32_t *__restrict__ csize,
                        uint32_t *visited, uint32_t *label, uint32_t *__restrict__ seed_flag,
                        uint64_t *__restrict__ seed_size, uint32_t *__restrict__ seed_max) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_roots) return;
  const uint32_t r = roots[i];
  const uint32_t m = csize[r] + 1u;
  uint32_t mem[SMALL_COMP], wa[SMALL_COMP], wb[SMALL_COMP];
  gather_small_component(g, r, m, mem);
  if (MAXIMUM) replay_maximum(g, mem, m, wa, wb, visited, seed_size + r, seed_max + r);
  else replay_directional(g, mem, m, wa, label, seed_flag, seed_size, seed_max);
}

// ---- (4) ids ----------------------------------------------------------------------------
__global__ void compact_clusters_kernel(const uint32_t *__restrict__ seed_flag, const uint32_t *__restrict__ seed_index,
                                        const uint64_t *__restrict__ seed_size, const uint32_t *__restrict__ seed_max,
                                        uint64_t U, uint64_t *__restrict__ cluster_size,
                                        uint32_t *__restrict__ cluster_max) {
  uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r < U && seed_flag[r]) {
    cluster_size[seed_index[r]] = seed_size[r];
    cluster_max[seed_index[r]] = seed_max[r];
  }
}

int stage_cluster(State &s) {
  const uint64_t U = s.unique;
  s.n_clusters = 0;
  if (U == 0) return 0;
  int k = 0;
  const unsigned g256 = grid_for(U, 256);
  s.comp.ensure(U);
  s.label.ensure(U);
  s.work_a.ensure(U);
  s.seg_end.ensure(U);  // csize
  uint32_t *csize = s.seg_end.p;
  {
    // 8 (entry) + 8 (two parent reads) per entry; 12 per word
    Timed t(s, "union_find", s.n_entries, s.n_entries * 16 + U * 12);
    HUMID_CUDA(cudaMemsetAsync(csize, 0, U * sizeof(uint32_t), s.stream));
    iota_kernel<<<g256, 256, 0, s.stream>>>(s.work_a.p, U);  // parent
    uint32_t *open_flag = nullptr;
    if (s.cluster_skip_open) {
      s.open_flag.ensure(U);
      s.comp_open.ensure(U);
      HUMID_CUDA(cudaMemsetAsync(s.open_flag.p, 0, U * sizeof(uint32_t), s.stream));
      HUMID_CUDA(cudaMemsetAsync(s.comp_open.p, 0, U * sizeof(uint32_t), s.stream));
      open_flag = s.open_flag.p;
    }
    if (s.cluster_pairs && !s.cluster_skip_open && s.cluster_col_base == 0) {
      if (s.n_cluster_pairs)
        uf_union_pairs_kernel<<<grid_for(s.n_cluster_pairs, 256), 256, 0, s.stream>>>(s.cluster_pairs, s.n_cluster_pairs,
                                                                                       s.work_a.p);
    } else if (s.cluster_entries) {
      if (s.n_cluster_entries)
        uf_union_entries_kernel<<<grid_for(s.n_cluster_entries, 256), 256, 0, s.stream>>>(
            s.cluster_entries, s.n_cluster_entries, U, s.cluster_col_base, open_flag, s.work_a.p);
    } else {
      uf_union_kernel<<<g256, 256, 0, s.stream>>>(s.row_ptr.p, s.col.p, s.n_entries, U, s.cluster_col_base, open_flag,
                                                  s.work_a.p);
    }
    uf_flatten_kernel<<<g256, 256, 0, s.stream>>>(s.work_a.p, s.comp.p, csize, U, open_flag, s.comp_open.p);
    t.launches = 3;
  }
  k += 3;

  s.seed_flag.ensure(U);
  s.seed_index.ensure(U);
  s.seed_size.ensure(U);
  s.seed_max.ensure(U);
  s.seg_start.ensure(U);
  s.key_alt.ensure(U + 1);
  s.work_b.ensure(U + 1);  // roots of the small components
  Graph g{s.row_ptr.p, s.col.p, s.word_count.p, s.cluster_col_base};
  const RootSkip skip_root{s.cluster_skip_open ? s.comp_open.p : nullptr, s.cluster_owner_mod, s.cluster_owner_me};
  // components of up to SMALL_COMP words are replayed by one thread each; HUMID_GPU_SMALL_COMP lowers the limit (tests:
  // 1 sends every component through cluster_big.cu)
  uint32_t small_max = SMALL_COMP;
  if (const char *e = std::getenv("HUMID_GPU_SMALL_COMP")) small_max = std::min<uint32_t>(SMALL_COMP, std::max(1, atoi(e)));
  uint64_t host_n[2] = {0, 0};  // roots of large components, roots of small components
  {
    Timed t(s, "classify_components", U, U * 16);
    HUMID_CUDA(cudaMemsetAsync(s.counters.p + 8, 0, 2 * sizeof(unsigned long long), s.stream));
    if (s.cfg.maximum)
      classify_words_kernel<true><<<g256, 256, 0, s.stream>>>(s.comp.p, csize, s.word_count.p, U, small_max, skip_root,
                                                              s.label.p, s.seed_flag.p, s.seed_size.p, s.seed_max.p,
                                                              s.work_b.p, s.counters.p + 9, s.counters.p + 8);
    else
      classify_words_kernel<false><<<g256, 256, 0, s.stream>>>(s.comp.p, csize, s.word_count.p, U, small_max, skip_root,
                                                               s.label.p, s.seed_flag.p, s.seed_size.p, s.seed_max.p,
                                                               s.work_b.p, s.counters.p + 9, s.counters.p + 8);
  }
  HUMID_CUDA(cudaMemcpyAsync(host_n, s.counters.p + 8, sizeof(host_n), cudaMemcpyDeviceToHost, s.stream));
  HUMID_CUDA(cudaStreamSynchronize(s.stream));
  k++;
  const uint64_t n_small = host_n[1];
  uint64_t n_big = 0;
  if (host_n[0]) {  // some component is large: list the words of all of them
    HUMID_CUDA(cudaMemsetAsync(s.counters.p + 8, 0, sizeof(unsigned long long), s.stream));
    big_member_keys_kernel<<<g256, 256, 0, s.stream>>>(s.comp.p, csize, U, small_max, skip_root, s.key_alt.p,
                                                       s.counters.p + 8);
    HUMID_CUDA(cudaMemcpyAsync(&n_big, s.counters.p + 8, sizeof(uint64_t), cudaMemcpyDeviceToHost, s.stream));
    HUMID_CUDA(cudaStreamSynchronize(s.stream));
    k++;
  }

  if (n_small) {
    Timed t(s, s.cfg.maximum ? "replay_maximum" : "replay_directional", n_small, s.n_entries * 8 + U * 24);
    const unsigned gs = grid_for(n_small, HUMID_REPLAY_THREADS);
    if (s.cfg.maximum) {
      s.work_c.ensure(U);  // visited flags of the depth-first search (only when counts tie)
      HUMID_CUDA(cudaMemsetAsync(s.work_c.p, 0, U * sizeof(uint32_t), s.stream));
      replay_small_kernel<true><<<gs, HUMID_REPLAY_THREADS, 0, s.stream>>>(g, s.work_b.p, n_small, csize, s.work_c.p,
                                                                           s.label.p, s.seed_flag.p, s.seed_size.p,
                                                                           s.seed_max.p);
    } else {
      replay_small_kernel<false><<<gs, HUMID_REPLAY_THREADS, 0, s.stream>>>(g, s.work_b.p, n_small, csize, nullptr,
                                                                            s.label.p, s.seed_flag.p, s.seed_size.p,
                                                                            s.seed_max.p);
    }
    k++;
  }

  // words of large components: grouped by component in rank order, then clustered all at once (cluster_big.cu)
  if (n_big) {
    int launches = 0;
    {
      Timed t(s, "group_big_components", n_big, n_big * 40);
      s.members.ensure(n_big + 1);
      s.entries_alt.ensure(n_big + 1);
      uint64_t *kb[2] = {s.key_alt.p, s.entries_alt.p};
      uint32_t *vb[2] = {nullptr, nullptr};
      int ubits = 1;
      while ((1ull << ubits) < U) ubits++;
      int cur = radix_sort_u64(kb, vb, n_big, 32 + ubits, s.scratch, s.scan_scratch, s.stream, &launches);
      big_segments_kernel<<<grid_for(n_big, 256), 256, 0, s.stream>>>(kb[cur], n_big, s.members.p, s.seg_start.p);
      launches++;
      t.launches = launches;
    }
    k += launches;
    k += s.cfg.maximum ? stage_big_maximum(s, g, n_big) : stage_big_directional(s, g, n_big);
  }
  HUMID_CUDA(cudaGetLastError());
  if (s.cluster_skip_open || s.cluster_no_ids) return k;  // shard mode: ids are handed out after the open components are merged in
  return k + stage_cluster_number(s);
}

// ids in walk order of the seeding word (src/humid.cc:180): exclusive scan of the seed flags
int stage_cluster_number(State &s) {
  const uint64_t U = s.unique;
  const unsigned g256 = grid_for(U, 256);
  int k = 0;
  uint32_t *total = reinterpret_cast<uint32_t *>(s.counters.p + 3);
  size_t t_ids = s.timer_begin("cluster_ids_scan", U, U * 8);
  int scan_launches = exclusive_scan_u32(s.seed_flag.p, s.seed_index.p, U, total, s.scan_scratch, s.stream);
  s.timer_end(t_ids, scan_launches);
  k += scan_launches;
  uint32_t C = 0;
  HUMID_CUDA(cudaMemcpyAsync(&C, total, sizeof(uint32_t), cudaMemcpyDeviceToHost, s.stream));
  HUMID_CUDA(cudaStreamSynchronize(s.stream));
  s.n_clusters = C;
  s.cluster_size.ensure(C + 1);
  s.cluster_max.ensure(C + 1);
  {
    Timed t(s, "compact_clusters", U, U * 8 + (uint64_t)C * 24);
    compact_clusters_kernel<<<g256, 256, 0, s.stream>>>(s.seed_flag.p, s.seed_index.p, s.seed_size.p, s.seed_max.p,
                                                        U, s.cluster_size.p, s.cluster_max.p);
  }
  k++;
  HUMID_CUDA(cudaGetLastError());
  return k;
}

}  // namespace humid
