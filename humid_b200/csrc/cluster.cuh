// Shared between cluster.cu (components, small components) and cluster_big.cu (components of more than SMALL_COMP
// words: data-parallel over all their words at once).
#pragma once
#include "state.cuh"

namespace humid {

#ifndef HUMID_SMALL_COMP
#define HUMID_SMALL_COMP 32
#endif
constexpr uint32_t SMALL_COMP = HUMID_SMALL_COMP;

struct Graph {
  const uint64_t *row_ptr;
  const uint32_t *col;
  const uint32_t *count;
  uint32_t base;  // columns are stored as rank + base (shard mode: global ranks), rows are 0-based
};
__device__ __forceinline__ uint32_t gcol(const Graph &g, uint64_t e) { return g.col[e] - g.base; }

// atLeastDouble_ (src/cluster.cc:31-33) on 32-bit counts
__device__ __forceinline__ bool at_least_double(uint32_t a, uint32_t b) { return (uint64_t)a >= 2ull * b; }

// The words of the large components: members[0..n_big) grouped by component in rank order, seg_start[root] = first
// position of the component of `root`.  Both write label / seed_flag / seed_size / seed_max like the small replay.
int stage_big_directional(State &s, const Graph &g, uint64_t n_big);
int stage_big_maximum(State &s, const Graph &g, uint64_t n_big);

}  // namespace humid
