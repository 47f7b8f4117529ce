// K4, components of more than SMALL_COMP words — all of them at once, one thread per word, whatever their size
// (the reference's failure mode: docs/troubleshooting.rst:6-18, recursion depth = cluster diameter; one thread
// replaying a 10^5..10^6-word component of low-complexity reads would hold the GPU for seconds).
//
// Directional clusters (src/cluster.cc:39-69, :82-87; seeds in rank order, src/humid.cc:178-189) without the
// sequential replay.  Let D be the directed graph u -> v iff u ~ v and count[u] >= 2 count[v]; counts strictly fall
// along D, so D is acyclic and at most 32 levels deep.  Between two clusters
//   (A) an assigned word has only assigned D-descendants (assignDirectionalCluster_ took them),
//   (B) so an unassigned word has only unassigned D-ancestors and maxNeighbour_ never meets an assigned candidate:
//       the climb follows the STATIC pointer up(x) = first neighbour in list order with count >= 2 count[x];
//   (C) the tops (Cluster::maxLeaf) are exactly the sources of D, each the top of one cluster;
//   (D) a word belongs to the first-processed source among its ancestor sources: with T(t) = rank of the seed that
//       climbs to t first, label(v) = M(v) = min over ancestor sources a of T(a);
//   (E) T(t) = min { s : top(s) = t and M(s) >= s }  (s is still unassigned when its turn comes).
// (E) is solved in rounds over all sources: an unconfirmed source holds a candidate (the smallest not yet rejected
// word of its up-tree: a lower bound of T); M_lo / M_hi = min over ancestor sources of (confirmed T or candidate) /
// (confirmed T or infinity) are pushed down D level by level; a candidate s with M_hi(s) < s is rejected, one with
// M_lo(s) >= s confirmed.  The source with the smallest candidate always moves; random graphs of 10^4..10^5 words
// need 3-5 rounds.  tests/directional_fixpoint.py restates this in Python and is checked against the recursion.
//
// Maximum clusters (src/cluster.cc:72-80): cluster = component, size = sum of counts, maxLeaf = first word in
// depth-first pre-order with the largest count: a parallel reduction unless several words tie for it, then one warp
// per component walks the depth-first order (lists scanned 32 entries at a time) until it meets the first of them.
#include <algorithm>

#include "cluster.cuh"

namespace humid {

// ---- directional ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
    big_up_kernel(Graph g, const uint32_t *__restrict__ members, uint64_t n, uint32_t *__restrict__ up) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t x = members[i], cx = g.count[x];
  uint32_t u = NONE32;
  for (uint64_t e = g.row_ptr[x], hi = g.row_ptr[x + 1]; e < hi; e++) {
    const uint32_t nb = gcol(g, e);
    if (at_least_double(g.count[nb], cx)) { u = nb; break; }
  }
  up[x] = u;
}

// (top, word) keys of the up-trees; the sources are listed and start unconfirmed
__global__ void __launch_bounds__(256)
    big_top_keys_kernel(const uint32_t *__restrict__ members, uint64_t n, const uint32_t *__restrict__ up,
                        uint64_t *__restrict__ keys, uint32_t *__restrict__ T, uint32_t *__restrict__ level,
                        uint32_t *__restrict__ src, unsigned long long *__restrict__ n_src) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  bool is_src = false;
  uint32_t x = 0;
  if (i < n) {
    x = members[i];
    uint32_t t = x;
    while (up[t] != NONE32) t = up[t];
    keys[i] = ((uint64_t)t << 32) | x;
    is_src = t == x;
    T[x] = NONE32;
    level[x] = 0;
  }
  const unsigned m = __ballot_sync(0xffffffffu, is_src);
  if (m) {
    const uint32_t lane = lane_id();
    unsigned long long base = 0;
    const int leader = __ffs(m) - 1;
    if ((int)lane == leader) base = atomicAdd(n_src, (unsigned long long)__popc(m));
    base = __shfl_sync(0xffffffffu, base, leader);
    if (is_src) src[base + __popc(m & ((1u << lane) - 1u))] = x;
  }
}

__global__ void __launch_bounds__(256)
    big_tree_start_kernel(const uint64_t *__restrict__ sorted, uint64_t n, uint32_t *__restrict__ cursor) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t t = (uint32_t)(sorted[i] >> 32);
  if (i == 0 || (uint32_t)(sorted[i - 1] >> 32) != t) cursor[t] = (uint32_t)i;
}

// level(v) = longest D-path from a source; one Jacobi sweep, `changed` counts the words that moved
__global__ void __launch_bounds__(256)
    big_level_kernel(Graph g, const uint32_t *__restrict__ members, uint64_t n, const uint32_t *__restrict__ up,
                     uint32_t *level, unsigned long long *__restrict__ changed /* [0] words that moved, [1] deepest level */) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t x = members[i];
  if (up[x] == NONE32) return;
  const uint32_t cx = g.count[x];
  uint32_t l = 0;
  for (uint64_t e = g.row_ptr[x], hi = g.row_ptr[x + 1]; e < hi; e++) {
    const uint32_t u = gcol(g, e);
    if (at_least_double(g.count[u], cx)) l = max(l, level[u] + 1u);
  }
  if (l != level[x]) {
    level[x] = l;
    atomicAdd(changed, 1ull);
    atomicMax(changed + 1, (unsigned long long)l);
  }
}

// start of a propagation: sources carry their value, everything else infinity.  FINAL: lo = T only.
template <bool FINAL>
__global__ void __launch_bounds__(256)
    big_values_kernel(const uint32_t *__restrict__ members, uint64_t n, const uint32_t *__restrict__ up,
                      const uint32_t *__restrict__ T, const uint32_t *__restrict__ cursor,
                      const uint64_t *__restrict__ sorted, uint32_t *__restrict__ lo, uint32_t *__restrict__ hi) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t x = members[i];
  uint32_t a = NONE32, b = NONE32;
  if (up[x] == NONE32) {
    b = T[x];
    a = (FINAL || b != NONE32) ? b : (uint32_t)sorted[cursor[x]];
  }
  lo[x] = a;
  if (!FINAL) hi[x] = b;
}

// one sweep of min over the D-in-neighbours (in place: values only fall towards the fixed point, so any
// interleaving is safe; `sweeps` >= deepest level makes it exact)
template <bool FINAL>
__global__ void __launch_bounds__(256)
    big_sweep_kernel(Graph g, const uint32_t *__restrict__ members, uint64_t n, const uint32_t *__restrict__ up,
                     uint32_t *lo, uint32_t *hi) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t x = members[i];
  if (up[x] == NONE32) return;
  const uint32_t cx = g.count[x];
  uint32_t a = lo[x], b = FINAL ? 0u : hi[x];
  for (uint64_t e = g.row_ptr[x], end = g.row_ptr[x + 1]; e < end; e++) {
    const uint32_t u = gcol(g, e);
    if (at_least_double(g.count[u], cx)) {
      a = min(a, lo[u]);
      if (!FINAL) b = min(b, hi[u]);
    }
  }
  lo[x] = a;
  if (!FINAL) hi[x] = b;
}

__global__ void __launch_bounds__(256)
    big_round_kernel(const uint32_t *__restrict__ src, uint64_t n_src, const uint64_t *__restrict__ sorted,
                     const uint32_t *__restrict__ lo, const uint32_t *__restrict__ hi, uint32_t *__restrict__ T,
                     uint32_t *__restrict__ cursor, unsigned long long *__restrict__ open) {
  const uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n_src) return;
  const uint32_t t = src[j];
  if (T[t] != NONE32) return;
  uint32_t c = cursor[t];
  while (true) {
    const uint32_t s = (uint32_t)sorted[c];  // t's own tree never runs out: t itself is always a valid seed
    if (hi[s] < s) { c++; continue; }        // a source confirmed for an earlier time takes s first
    if (lo[s] >= s) T[t] = s;                // nothing can take s before its turn: s seeds t's cluster
    else atomicAdd(open, 1ull);
    break;
  }
  cursor[t] = c;
}

__global__ void __launch_bounds__(256)
    big_seeds_kernel(const uint32_t *__restrict__ src, uint64_t n_src, const uint32_t *__restrict__ T,
                     uint32_t *__restrict__ seed_flag, uint64_t *__restrict__ seed_size, uint32_t *__restrict__ seed_max) {
  const uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n_src) return;
  const uint32_t t = src[j], s = T[t];
  seed_flag[s] = 1;
  seed_size[s] = 0;
  seed_max[s] = t;
}

__global__ void __launch_bounds__(256)
    big_labels_kernel(Graph g, const uint32_t *__restrict__ members, uint64_t n, const uint32_t *__restrict__ lo,
                      uint32_t *__restrict__ label, uint64_t *seed_size) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t x = members[i], s = lo[x];
  label[x] = s;
  atomicAdd(reinterpret_cast<unsigned long long *>(seed_size + s), (unsigned long long)g.count[x]);
}

int stage_big_directional(State &s, const Graph &g, uint64_t n) {
  const uint64_t U = s.unique;
  const unsigned gn = grid_for(n, 256);
  int k = 0;
  Timed t(s, "cluster_big_directional", n, n * 64);
  s.big_up.ensure(U);
  s.big_t.ensure(U);
  s.big_cursor.ensure(U);
  s.big_lo.ensure(U);
  s.big_hi.ensure(U);
  s.big_level.ensure(U);
  s.big_src.ensure(n + 1);
  s.big_keys.ensure(n + 1);
  s.big_keys_alt.ensure(n + 1);
  unsigned long long *n_src_dev = s.counters.p + 12, *flag = s.counters.p + 13;
  HUMID_CUDA(cudaMemsetAsync(n_src_dev, 0, 3 * sizeof(unsigned long long), s.stream));
  big_up_kernel<<<gn, 256, 0, s.stream>>>(g, s.members.p, n, s.big_up.p);
  big_top_keys_kernel<<<gn, 256, 0, s.stream>>>(s.members.p, n, s.big_up.p, s.big_keys.p, s.big_t.p, s.big_level.p,
                                                s.big_src.p, n_src_dev);
  k += 2;
  uint64_t *kb[2] = {s.big_keys.p, s.big_keys_alt.p};
  uint32_t *vb[2] = {nullptr, nullptr};
  int ubits = 1;
  while ((1ull << ubits) < U) ubits++;
  const int cur = radix_sort_u64(kb, vb, n, 32 + ubits, s.scratch, s.scan_scratch, s.stream, &k);
  const uint64_t *sorted = kb[cur];
  big_tree_start_kernel<<<gn, 256, 0, s.stream>>>(sorted, n, s.big_cursor.p);
  k++;
  // deepest level of D over these words: sweeps until nothing moves (<= 32: counts double along every edge)
  uint64_t host[3] = {0, 0, 0};  // sources, words that moved in the last sweep, deepest level
  for (int depth = 0;; depth++) {
    HUMID_CUDA(cudaMemsetAsync(flag, 0, sizeof(unsigned long long), s.stream));
    big_level_kernel<<<gn, 256, 0, s.stream>>>(g, s.members.p, n, s.big_up.p, s.big_level.p, flag);
    k++;
    HUMID_CUDA(cudaMemcpyAsync(host, n_src_dev, sizeof(host), cudaMemcpyDeviceToHost, s.stream));
    HUMID_CUDA(cudaStreamSynchronize(s.stream));
    if (host[1] == 0) break;
    if (depth > 40) throw CudaError{"directional clustering: the count order of a large component does not settle"};
  }
  const uint64_t n_src = host[0];
  const int sweeps = std::max<int>((int)host[2], 1);
  const unsigned gs = grid_for(n_src, 256);
  for (int round = 0;; round++) {
    big_values_kernel<false><<<gn, 256, 0, s.stream>>>(s.members.p, n, s.big_up.p, s.big_t.p, s.big_cursor.p, sorted,
                                                       s.big_lo.p, s.big_hi.p);
    for (int q = 0; q < sweeps; q++)
      big_sweep_kernel<false><<<gn, 256, 0, s.stream>>>(g, s.members.p, n, s.big_up.p, s.big_lo.p, s.big_hi.p);
    HUMID_CUDA(cudaMemsetAsync(flag, 0, sizeof(unsigned long long), s.stream));
    big_round_kernel<<<gs, 256, 0, s.stream>>>(s.big_src.p, n_src, sorted, s.big_lo.p, s.big_hi.p, s.big_t.p,
                                               s.big_cursor.p, flag);
    k += 2 + sweeps;
    HUMID_CUDA(cudaMemcpyAsync(host + 1, flag, sizeof(uint64_t), cudaMemcpyDeviceToHost, s.stream));
    HUMID_CUDA(cudaStreamSynchronize(s.stream));
    if (host[1] == 0) break;
    if ((uint64_t)round > n_src + 8) throw CudaError{"directional clustering: seeds of a large component do not settle"};
  }
  big_values_kernel<true><<<gn, 256, 0, s.stream>>>(s.members.p, n, s.big_up.p, s.big_t.p, s.big_cursor.p, sorted,
                                                    s.big_lo.p, s.big_hi.p);
  for (int q = 0; q < sweeps; q++)
    big_sweep_kernel<true><<<gn, 256, 0, s.stream>>>(g, s.members.p, n, s.big_up.p, s.big_lo.p, s.big_hi.p);
  big_seeds_kernel<<<gs, 256, 0, s.stream>>>(s.big_src.p, n_src, s.big_t.p, s.seed_flag.p, s.seed_size.p, s.seed_max.p);
  big_labels_kernel<<<gn, 256, 0, s.stream>>>(g, s.members.p, n, s.big_lo.p, s.label.p, s.seed_size.p);
  k += 3 + sweeps;
  HUMID_CUDA(cudaGetLastError());
  t.launches = k;
  return k;
}

// ---- maximum ----------------------------------------------------------------------------------------------
// per component (indexed by its root): acc[root] = {size in seed_size, best count in lo, ties in hi, smallest tied
// rank in T}
__global__ void __launch_bounds__(256)
    big_max_clear_kernel(const uint32_t *__restrict__ members, uint64_t n, const uint32_t *__restrict__ comp,
                         uint64_t *__restrict__ seed_size, uint32_t *__restrict__ best, uint32_t *__restrict__ ties,
                         uint32_t *__restrict__ first_tied) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t x = members[i];
  if (comp[x] != x) return;
  seed_size[x] = 0;
  best[x] = 0;
  ties[x] = 0;
  first_tied[x] = NONE32;
}

__global__ void __launch_bounds__(256)
    big_max_sum_kernel(Graph g, const uint32_t *__restrict__ members, uint64_t n, const uint32_t *__restrict__ comp,
                       uint64_t *seed_size, uint32_t *best) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t x = members[i], root = comp[x], c = g.count[x];
  atomicAdd(reinterpret_cast<unsigned long long *>(seed_size + root), (unsigned long long)c);
  atomicMax(best + root, c);
}

__global__ void __launch_bounds__(256)
    big_max_ties_kernel(Graph g, const uint32_t *__restrict__ members, uint64_t n, const uint32_t *__restrict__ comp,
                        const uint32_t *__restrict__ best, uint32_t *ties, uint32_t *first_tied) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t x = members[i], root = comp[x];
  if (g.count[x] != best[root]) return;
  atomicAdd(ties + root, 1u);
  atomicMin(first_tied + root, x);
}

// maxLeaf of every large component that needs no search; the others are listed
__global__ void __launch_bounds__(256)
    big_max_roots_kernel(Graph g, const uint32_t *__restrict__ members, uint64_t n, const uint32_t *__restrict__ comp,
                         const uint32_t *__restrict__ best, const uint32_t *__restrict__ ties,
                         const uint32_t *__restrict__ first_tied, uint32_t *__restrict__ seed_max,
                         uint32_t *__restrict__ todo, unsigned long long *__restrict__ n_todo) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t x = members[i];
  if (comp[x] != x) return;
  if (ties[x] == 1) seed_max[x] = first_tied[x];
  else if (g.count[x] == best[x]) seed_max[x] = x;  // the walk starts at the root: it is the first word of the pre-order
  else todo[atomicAdd(n_todo, 1ull)] = x;
}

// One warp per component with tied counts: the depth-first walk of assignMaxCluster (src/cluster.cc:72-80) from the
// root in list order, until the first word holding the largest count.  The lanes scan a list 32 entries at a time;
// the stack (word, position in its list) lives behind the component's members.
__global__ void __launch_bounds__(128)
    big_max_dfs_kernel(Graph g, const uint32_t *__restrict__ todo, uint64_t n_todo, const uint32_t *__restrict__ seg_start,
                       const uint32_t *__restrict__ best, uint32_t *stk_node, uint32_t *stk_edge, uint32_t *visited,
                       uint32_t *__restrict__ seed_max) {
  const uint64_t w = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (w >= n_todo) return;
  const uint32_t lane = lane_id();
  const uint32_t root = todo[w], want = best[root];
  uint32_t *sn = stk_node + seg_start[root], *se = stk_edge + seg_start[root];
  uint32_t sp = 1, u = root, cur = 0, answer = root;
  if (lane == 0) {
    visited[root] = 1;
    sn[0] = root;
  }
  __syncwarp();
  while (sp > 0) {
    const uint64_t lo = g.row_ptr[u];
    const uint32_t deg = (uint32_t)(g.row_ptr[u + 1] - lo);
    uint32_t found = NONE32;
    while (cur < deg) {
      const uint32_t idx = cur + lane;
      const uint32_t v = idx < deg ? gcol(g, lo + idx) : NONE32;
      const bool fresh = v != NONE32 && reinterpret_cast<volatile uint32_t *>(visited)[v] == 0u;
      const unsigned m = __ballot_sync(0xffffffffu, fresh);
      if (m) {
        const int f = __ffs(m) - 1;
        found = __shfl_sync(0xffffffffu, v, f);
        cur += (uint32_t)f + 1u;
        break;
      }
      cur += 32;
    }
    if (found == NONE32) {  // list exhausted: back to the parent
      sp--;
      if (sp > 0) {
        u = reinterpret_cast<volatile uint32_t *>(sn)[sp - 1];
        cur = reinterpret_cast<volatile uint32_t *>(se)[sp - 1];
      }
      continue;
    }
    if (lane == 0) {
      visited[found] = 1;
      se[sp - 1] = cur;  // where to go on in u's list
    }
    if (g.count[found] == want) { answer = found; break; }
    if (lane == 0) sn[sp] = found;
    sp++;
    u = found;
    cur = 0;
    __syncwarp();
  }
  if (lane == 0) seed_max[root] = answer;
}

int stage_big_maximum(State &s, const Graph &g, uint64_t n) {
  const uint64_t U = s.unique;
  const unsigned gn = grid_for(n, 256);
  int k = 0;
  Timed t(s, "cluster_big_maximum", n, n * 40);
  s.big_lo.ensure(U);
  s.big_hi.ensure(U);
  s.big_t.ensure(U);
  s.big_src.ensure(n + 1);
  s.big_up.ensure(n + 1);      // stack: words
  s.big_cursor.ensure(n + 1);  // stack: list positions
  unsigned long long *n_todo_dev = s.counters.p + 12;
  HUMID_CUDA(cudaMemsetAsync(n_todo_dev, 0, sizeof(unsigned long long), s.stream));
  big_max_clear_kernel<<<gn, 256, 0, s.stream>>>(s.members.p, n, s.comp.p, s.seed_size.p, s.big_lo.p, s.big_hi.p, s.big_t.p);
  big_max_sum_kernel<<<gn, 256, 0, s.stream>>>(g, s.members.p, n, s.comp.p, s.seed_size.p, s.big_lo.p);
  big_max_ties_kernel<<<gn, 256, 0, s.stream>>>(g, s.members.p, n, s.comp.p, s.big_lo.p, s.big_hi.p, s.big_t.p);
  big_max_roots_kernel<<<gn, 256, 0, s.stream>>>(g, s.members.p, n, s.comp.p, s.big_lo.p, s.big_hi.p, s.big_t.p,
                                                 s.seed_max.p, s.big_src.p, n_todo_dev);
  k += 4;
  uint64_t n_todo = 0;
  HUMID_CUDA(cudaMemcpyAsync(&n_todo, n_todo_dev, sizeof(uint64_t), cudaMemcpyDeviceToHost, s.stream));
  HUMID_CUDA(cudaStreamSynchronize(s.stream));
  if (n_todo) {
    s.work_c.ensure(U);  // visited
    HUMID_CUDA(cudaMemsetAsync(s.work_c.p, 0, U * sizeof(uint32_t), s.stream));
    big_max_dfs_kernel<<<grid_for(n_todo * 32, 128), 128, 0, s.stream>>>(g, s.big_src.p, n_todo, s.seg_start.p, s.big_lo.p,
                                                                        s.big_up.p, s.big_cursor.p, s.work_c.p,
                                                                        s.seed_max.p);
    k++;
  }
  HUMID_CUDA(cudaGetLastError());
  t.launches = k;
  return k;
}

}  // namespace humid
