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

#include "state.cuh"

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

#ifndef HUMID_SMALL_COMP
#define HUMID_SMALL_COMP 32
#endif
constexpr uint32_t SMALL_COMP = HUMID_SMALL_COMP;

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

// ---- (2) members of the LARGE components, grouped by a global sort ---------------------------
__global__ void __launch_bounds__(256)
    big_member_keys_kernel(const uint32_t *__restrict__ comp, const uint32_t *__restrict__ csize, uint64_t U,
                           uint64_t *__restrict__ keys, unsigned long long *__restrict__ counter) {
  uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  bool big = r < U && csize[comp[r]] + 1u > SMALL_COMP;
  unsigned m = __ballot_sync(0xffffffffu, big);
  if (m == 0) return;
  unsigned long long base = 0;
  int leader = __ffs(m) - 1;
  if ((int)lane_id() == leader) base = atomicAdd(counter, (unsigned long long)__popc(m));
  base = __shfl_sync(0xffffffffu, base, leader);
  if (big) keys[base + __popc(m & ((1u << lane_id()) - 1u))] = ((uint64_t)comp[r] << 32) | (uint32_t)r;
}

__global__ void big_segments_kernel(const uint64_t *__restrict__ sorted, uint64_t n, uint32_t *__restrict__ members,
                                    uint32_t *__restrict__ seg_start) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint32_t c = (uint32_t)(sorted[i] >> 32);
  members[i] = (uint32_t)sorted[i];
  if (i == 0 || (uint32_t)(sorted[i - 1] >> 32) != c) seg_start[c] = (uint32_t)i;
}

struct Graph {
  const uint64_t *row_ptr;
  const uint32_t *col;
  const uint32_t *count;
  uint32_t base;  // columns are stored as rank + base (shard mode: global ranks), rows are 0-based
};
__device__ __forceinline__ uint32_t gcol(const Graph &g, uint64_t e) { return g.col[e] - g.base; }

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

// One thread per word; only component roots work.  Singletons are closed in place, small
// components are gathered and replayed in local memory, large ones use the globally sorted
// member array (big_members / big_start) and global work arrays of the same length.
template <bool MAXIMUM>
#ifndef HUMID_REPLAY_THREADS
#define HUMID_REPLAY_THREADS 128
#endif
__global__ void __launch_bounds__(HUMID_REPLAY_THREADS)
    replay_kernel(Graph g, uint64_t U, const uint32_t *__restrict__ comp, const uint32_t *__restrict__ csize,
                  const uint32_t *__restrict__ big_members, const uint32_t *__restrict__ big_start,
                  uint32_t *__restrict__ big_work_a, uint32_t *__restrict__ big_work_b, uint32_t *visited,
                  uint32_t *label, uint32_t *__restrict__ seed_flag, uint64_t *__restrict__ seed_size,
                  uint32_t *__restrict__ seed_max, const uint32_t *__restrict__ skip_root) {
  const uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= U) return;
  const uint32_t root = comp[r];
  if (skip_root && skip_root[root]) {  // shard mode: component continues on another shard, replayed later
    if (MAXIMUM) seed_flag[r] = 0;
    return;
  }
  if (MAXIMUM) {
    label[r] = root;
    seed_flag[r] = root == r ? 1u : 0u;
  }
  if (root != r) return;
  const uint32_t m = csize[r] + 1u;
  if (m == 1) {  // no neighbours: its own cluster
    if (!MAXIMUM) { label[r] = (uint32_t)r; seed_flag[r] = 1; }
    seed_size[r] = g.count[r];
    seed_max[r] = (uint32_t)r;
    return;
  }
  if (m <= SMALL_COMP) {
    uint32_t mem[SMALL_COMP], wa[SMALL_COMP], wb[SMALL_COMP];
    gather_small_component(g, (uint32_t)r, m, mem);
    if (MAXIMUM) replay_maximum(g, mem, m, wa, wb, visited, seed_size + r, seed_max + r);
    else replay_directional(g, mem, m, wa, label, seed_flag, seed_size, seed_max);
  } else {
    const uint32_t s0 = big_start[r];
    if (MAXIMUM) replay_maximum(g, big_members + s0, m, big_work_a + s0, big_work_b + s0, visited, seed_size + r, seed_max + r);
    else replay_directional(g, big_members + s0, m, big_work_a + s0, label, seed_flag, seed_size, seed_max);
  }
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
  const unsigned g256 = grid_for(U, 256), g128 = grid_for(U, HUMID_REPLAY_THREADS);
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
    uf_union_kernel<<<g256, 256, 0, s.stream>>>(s.row_ptr.p, s.col.p, s.n_entries, U, s.cluster_col_base, open_flag,
                                                s.work_a.p);
    uf_flatten_kernel<<<g256, 256, 0, s.stream>>>(s.work_a.p, s.comp.p, csize, U, open_flag, s.comp_open.p);
    t.launches = 3;
  }
  k += 3;

  // words of large components, grouped by component in rank order (usually none)
  uint64_t n_big = 0;
  s.seg_start.ensure(U);
  {
    Timed t(s, "group_big_components", U, U * 8);
    s.key_alt.ensure(U + 1);
    HUMID_CUDA(cudaMemsetAsync(s.counters.p + 8, 0, sizeof(unsigned long long), s.stream));
    big_member_keys_kernel<<<g256, 256, 0, s.stream>>>(s.comp.p, csize, U, s.key_alt.p, s.counters.p + 8);
    HUMID_CUDA(cudaMemcpyAsync(&n_big, s.counters.p + 8, sizeof(uint64_t), cudaMemcpyDeviceToHost, s.stream));
    HUMID_CUDA(cudaStreamSynchronize(s.stream));
    int launches = 1;
    s.members.ensure(n_big + 1);
    s.work_b.ensure(std::max<uint64_t>(n_big + 1, 1));
    s.work_c.ensure(std::max<uint64_t>(n_big + 1, U));  // also `visited` in maximum mode
    if (n_big) {
      s.entries_alt.ensure(n_big + 1);
      uint64_t *kb[2] = {s.key_alt.p, s.entries_alt.p};
      uint32_t *vb[2] = {nullptr, nullptr};
      int ubits = 1;
      while ((1ull << ubits) < U) ubits++;
      int cur = radix_sort_u64(kb, vb, n_big, 32 + ubits, s.scratch, s.scan_scratch, s.stream, &launches);
      big_segments_kernel<<<grid_for(n_big, 256), 256, 0, s.stream>>>(kb[cur], n_big, s.members.p, s.seg_start.p);
      launches++;
    }
    t.launches = launches;
    k += launches;
  }

  s.seed_flag.ensure(U);
  s.seed_index.ensure(U);
  s.seed_size.ensure(U);
  s.seed_max.ensure(U);
  Graph g{s.row_ptr.p, s.col.p, s.word_count.p, s.cluster_col_base};
  const uint32_t *skip_root = s.cluster_skip_open ? s.comp_open.p : nullptr;
  size_t t_replay = s.timer_begin(s.cfg.maximum ? "replay_maximum" : "replay_directional", U,
                                  s.n_entries * 8 + U * 24);
  if (s.cfg.maximum) {
    // big components: DFS stack (node, edge) in work_b / slot_alt, visited flags per word in work_c
    s.slot_alt.ensure(n_big + 1);
    HUMID_CUDA(cudaMemsetAsync(s.work_c.p, 0, U * sizeof(uint32_t), s.stream));
    replay_kernel<true><<<g128, HUMID_REPLAY_THREADS, 0, s.stream>>>(g, U, s.comp.p, csize, s.members.p, s.seg_start.p, s.work_b.p,
                                                    s.slot_alt.p, s.work_c.p, s.label.p, s.seed_flag.p, s.seed_size.p,
                                                    s.seed_max.p, skip_root);
  } else {
    HUMID_CUDA(cudaMemsetAsync(s.label.p, 0xFF, U * sizeof(uint32_t), s.stream));
    HUMID_CUDA(cudaMemsetAsync(s.seed_flag.p, 0, U * sizeof(uint32_t), s.stream));
    replay_kernel<false><<<g128, HUMID_REPLAY_THREADS, 0, s.stream>>>(g, U, s.comp.p, csize, s.members.p, s.seg_start.p, s.work_b.p,
                                                     nullptr, nullptr, s.label.p, s.seed_flag.p, s.seed_size.p,
                                                     s.seed_max.p, skip_root);
  }
  s.timer_end(t_replay, 1);
  k++;
  HUMID_CUDA(cudaGetLastError());
  if (s.cluster_skip_open) return k;  // shard mode: ids are handed out after the open components are merged in
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
