// K4 — clustering.  Replaces findClusters (reference src/humid.cc:167-193) with
// assignMaxCluster / assignDirectionalCluster (src/cluster.cc:39-87).
//
// The reference walks the words in rank order and grows one cluster per still-unassigned
// word by recursive traversal of the neighbour lists; the result depends on walk order and
// on neighbour-list order, but never crosses a connected component of the neighbour graph.
// So: (1) lock-free union-find gives the components (root = smallest rank = the word the
// walk meets first); (2) members are grouped per component in rank order; (3) every
// component is replayed exactly, independently of all others; (4) cluster ids are an
// exclusive scan over the seed flags in rank order (ids are handed out in walk order of
// the seeding word, src/humid.cc:180).
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
                    uint32_t *parent) {
  // one thread per directed entry; the source rank is recovered lazily only for a < b
  // entries via binary search would be wasteful, so sources come from a per-row loop instead
  uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= U) return;
  for (uint64_t e = row_ptr[r]; e < row_ptr[r + 1]; e++) {
    uint32_t b = col[e];
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

__global__ void uf_flatten_kernel(uint32_t *parent, uint32_t *comp, uint64_t U) {
  uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r < U) comp[r] = uf_find(parent, (uint32_t)r);
}

// ---- (2) members per component -------------------------------------------------------------
__global__ void member_keys_kernel(const uint32_t *__restrict__ comp, uint64_t U, uint64_t *keys, uint32_t *vals) {
  uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r < U) {
    keys[r] = comp[r];
    vals[r] = (uint32_t)r;
  }
}

__global__ void segment_bounds_kernel(const uint64_t *__restrict__ sorted_comp, uint64_t U, uint32_t *seg_start,
                                      uint32_t *seg_end) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= U) return;
  uint32_t c = (uint32_t)sorted_comp[i];
  if (i == 0 || (uint32_t)sorted_comp[i - 1] != c) seg_start[c] = (uint32_t)i;
  if (i + 1 == U || (uint32_t)sorted_comp[i + 1] != c) seg_end[c] = (uint32_t)i + 1;
}

// ---- (3) exact replay per component ------------------------------------------------------
struct Graph {
  const uint64_t *row_ptr;
  const uint32_t *col;
  const uint32_t *count;
};

// assignMaxCluster (src/cluster.cc:72-80): cluster == component, size = sum of counts,
// maxLeaf = first word in DFS pre-order (neighbour-list order) holding the strictly largest
// count.  The DFS is only replayed when several members tie for the largest count.
__global__ void __launch_bounds__(128)
    replay_maximum_kernel(Graph g, uint64_t U, const uint32_t *__restrict__ comp,
                          const uint32_t *__restrict__ members, const uint32_t *__restrict__ seg_start,
                          const uint32_t *__restrict__ seg_end, uint32_t *__restrict__ stk_node,
                          uint32_t *__restrict__ stk_edge_lo, uint32_t *__restrict__ visited,
                          uint32_t *__restrict__ label, uint32_t *__restrict__ seed_flag,
                          uint64_t *__restrict__ seed_size, uint32_t *__restrict__ seed_max) {
  uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= U) return;
  uint32_t root = comp[r];
  label[r] = root;
  if (root != r) { seed_flag[r] = 0; return; }
  seed_flag[r] = 1;
  const uint32_t s = seg_start[r], e = seg_end[r];
  uint64_t size = 0;
  uint32_t best = 0, best_rank = NONE32, ties = 0;
  for (uint32_t i = s; i < e; i++) {
    uint32_t m = members[i];
    uint32_t c = g.count[m];
    size += c;
    if (c > best) { best = c; best_rank = m; ties = 1; }
    else if (c == best) ties++;
  }
  seed_size[r] = size;
  if (ties > 1) {
    // DFS pre-order from the root until the first word with count == best is entered.
    // Frames: (node, next edge index relative to row start) in the component's own slice.
    uint32_t sp = s;
    visited[r] = 1;
    if (g.count[r] == best) { best_rank = (uint32_t)r; }
    else {
      stk_node[sp] = (uint32_t)r; stk_edge_lo[sp] = 0; sp++;
      bool found = false;
      while (sp > s && !found) {
        uint32_t u = stk_node[sp - 1];
        uint64_t lo = g.row_ptr[u], hi = g.row_ptr[u + 1];
        uint64_t j = lo + stk_edge_lo[sp - 1];
        if (j < hi) {
          stk_edge_lo[sp - 1]++;
          uint32_t v = g.col[j];
          if (!visited[v]) {
            visited[v] = 1;
            if (g.count[v] == best) { best_rank = v; found = true; }
            else { stk_node[sp] = v; stk_edge_lo[sp] = 0; sp++; }
          }
        } else {
          sp--;
        }
      }
    }
  }
  seed_max[r] = best_rank;
}

// assignDirectionalCluster (src/cluster.cc:82-87): for every still-unassigned member in rank
// order: climb (maxNeighbour_, :39-51) to the first-in-list unassigned neighbour with
// count >= 2x, restarting the scan after every move; the top word is maxLeaf; then take the
// closure over unassigned neighbours with count <= half (assignDirectionalCluster_, :58-69;
// a closure, so breadth-first order gives the same set as the reference's recursion).
__global__ void __launch_bounds__(128)
    replay_directional_kernel(Graph g, uint64_t U, const uint32_t *__restrict__ comp,
                              const uint32_t *__restrict__ members, const uint32_t *__restrict__ seg_start,
                              const uint32_t *__restrict__ seg_end, uint32_t *__restrict__ queue,
                              uint32_t *label, uint32_t *__restrict__ seed_flag,
                              uint64_t *__restrict__ seed_size, uint32_t *__restrict__ seed_max) {
  uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= U) return;
  if (comp[r] != r) return;
  const uint32_t s = seg_start[r], e = seg_end[r];
  for (uint32_t i = s; i < e; i++) {
    uint32_t seed = members[i];
    if (label[seed] != NONE32) continue;
    uint32_t node = seed;
    uint64_t j = g.row_ptr[node], hi = g.row_ptr[node + 1];
    uint64_t cn = g.count[node];
    while (j < hi) {
      uint32_t nb = g.col[j++];
      if (label[nb] == NONE32 && (uint64_t)g.count[nb] >= 2 * cn) {
        node = nb;
        j = g.row_ptr[node];
        hi = g.row_ptr[node + 1];
        cn = g.count[node];
      }
    }
    uint32_t head = s, tail = s;
    label[node] = seed;
    uint64_t size = cn;
    queue[tail++] = node;
    while (head < tail) {
      uint32_t u = queue[head++];
      uint64_t cu = g.count[u];
      for (uint64_t k = g.row_ptr[u]; k < g.row_ptr[u + 1]; k++) {
        uint32_t nb = g.col[k];
        if (label[nb] == NONE32) {
          uint64_t cb = g.count[nb];
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
  const unsigned g256 = grid_for(U, 256), g128 = grid_for(U, 128);
  s.comp.ensure(U);
  s.label.ensure(U);
  s.work_a.ensure(U);
  {
    // 8 (entry) + 8 (two parent reads) per entry; 12 per word
    Timed t(s, "union_find", s.n_entries, s.n_entries * 16 + U * 12);
    iota_kernel<<<g256, 256, 0, s.stream>>>(s.work_a.p, U);  // parent
    uf_union_kernel<<<g256, 256, 0, s.stream>>>(s.row_ptr.p, s.col.p, s.n_entries, U, s.work_a.p);
    uf_flatten_kernel<<<g256, 256, 0, s.stream>>>(s.work_a.p, s.comp.p, U);
    t.launches = 3;
  }
  k += 3;

  // members grouped by component (stable sort keeps rank order inside a component)
  s.key_alt.ensure(U + 1);
  s.entries_alt.ensure(U + 1);
  s.members.ensure(U);
  s.slot_alt.ensure(U + 1);
  uint64_t *kb[2] = {s.key_alt.p, s.entries_alt.p};
  uint32_t *vb[2] = {s.members.p, s.slot_alt.p};
  int ubits = 1;
  while ((1ull << ubits) < U) ubits++;
  int launches = 0;
  int cur;
  s.seg_start.ensure(U);
  s.seg_end.ensure(U);
  {
    Timed t(s, "group_members", U, U * (4 + 2 * 12 + 8));
    member_keys_kernel<<<g256, 256, 0, s.stream>>>(s.comp.p, U, kb[0], vb[0]);
    cur = radix_sort_u64(kb, vb, U, ubits, s.scratch, s.scan_scratch, s.stream, &launches);
    if (cur == 1) std::swap(s.members, s.slot_alt);
    segment_bounds_kernel<<<g256, 256, 0, s.stream>>>(kb[cur], U, s.seg_start.p, s.seg_end.p);
    t.launches = launches + 2;
  }
  k += launches + 2;

  s.seed_flag.ensure(U);
  s.seed_index.ensure(U);
  s.seed_size.ensure(U);
  s.seed_max.ensure(U);
  s.work_b.ensure(U);
  s.work_c.ensure(U);
  Graph g{s.row_ptr.p, s.col.p, s.word_count.p};
  size_t t_replay = s.timer_begin(s.cfg.maximum ? "replay_maximum" : "replay_directional", U,
                                  s.n_entries * 8 + U * 24);
  if (s.cfg.maximum) {
    HUMID_CUDA(cudaMemsetAsync(s.work_c.p, 0, U * sizeof(uint32_t), s.stream));  // visited
    replay_maximum_kernel<<<g128, 128, 0, s.stream>>>(g, U, s.comp.p, s.members.p, s.seg_start.p, s.seg_end.p,
                                                      s.work_a.p, s.work_b.p, s.work_c.p, s.label.p, s.seed_flag.p,
                                                      s.seed_size.p, s.seed_max.p);
  } else {
    HUMID_CUDA(cudaMemsetAsync(s.label.p, 0xFF, U * sizeof(uint32_t), s.stream));
    HUMID_CUDA(cudaMemsetAsync(s.seed_flag.p, 0, U * sizeof(uint32_t), s.stream));
    replay_directional_kernel<<<g128, 128, 0, s.stream>>>(g, U, s.comp.p, s.members.p, s.seg_start.p, s.seg_end.p,
                                                          s.work_a.p, s.label.p, s.seed_flag.p, s.seed_size.p,
                                                          s.seed_max.p);
  }
  s.timer_end(t_replay, 1);
  k++;
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
