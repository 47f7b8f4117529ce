// Device-resident state of one humid_gpu handle and the stage launchers that operate on it.
// Data layout in HBM (all structure-of-arrays, indices are 32-bit):
//   per tuple t:    tuple_key[t]                     packed word or KEY_INVALID        8 B
//   table:          slots[cap]                       {key, count|rank, first}         16 B
//   per rank r:     word_key[r], word_count[r], word_first[r]                        8+4+4 B
//                   row_ptr[r] (u64), label[r] (cluster seed rank), comp[r]
//   per entry:      col[e]                           neighbour rank                    4 B
//   per cluster:    indexed by seed rank (sparse) then compacted by id
#pragma once
#include <functional>
#include <vector>

#include "../../include/humid_gpu.h"
#include "primitives.cuh"

namespace humid {

// Device array owned by the handle: freed when the handle goes away, movable (std::swap of two
// columns), never copied.
template <typename T>
struct DevBuf {
  T *p = nullptr;
  size_t cap = 0;  // elements
  DevBuf() = default;
  DevBuf(const DevBuf &) = delete;
  DevBuf &operator=(const DevBuf &) = delete;
  DevBuf(DevBuf &&o) noexcept : p(o.p), cap(o.cap) { o.p = nullptr; o.cap = 0; }
  DevBuf &operator=(DevBuf &&o) noexcept {
    if (this != &o) {
      release();
      p = o.p;
      cap = o.cap;
      o.p = nullptr;
      o.cap = 0;
    }
    return *this;
  }
  ~DevBuf() { release(); }
  // Make room for n elements; contents are NOT preserved when reallocating.
  T *ensure(size_t n) {
    if (n > cap) {
      release();
      size_t want = n + n / 8 + 64;
      HUMID_CUDA(cudaMalloc((void **)&p, want * sizeof(T)));
      cap = want;
    }
    return p;
  }
  // Make room for n elements, preserving the first `keep` ones.
  T *grow(size_t n, size_t keep, cudaStream_t stream) {
    if (n > cap) {
      size_t want = n + n / 2 + 64;
      T *q = nullptr;
      HUMID_CUDA(cudaMalloc((void **)&q, want * sizeof(T)));
      if (keep) HUMID_CUDA(cudaMemcpyAsync(q, p, keep * sizeof(T), cudaMemcpyDeviceToDevice, stream));
      HUMID_CUDA(cudaStreamSynchronize(stream));
      if (p) cudaFree(p);
      p = q;
      cap = want;
    }
    return p;
  }
  void release() {
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
  }
};

// Page-locked host array owned by the handle (results that the host will read right after finalize).
template <typename T>
struct HostBuf {
  T *p = nullptr;
  size_t cap = 0;
  HostBuf() = default;
  HostBuf(const HostBuf &) = delete;
  HostBuf &operator=(const HostBuf &) = delete;
  ~HostBuf() { if (p) cudaFreeHost(p); }
  T *ensure(size_t n) {
    if (n > cap) {
      if (p) cudaFreeHost(p);
      p = nullptr;
      cap = 0;
      HUMID_CUDA(cudaHostAlloc((void **)&p, (n + n / 8 + 64) * sizeof(T), cudaHostAllocDefault));
      cap = n + n / 8 + 64;
    }
    return p;
  }
};

struct StageStat {
  const char *name;
  float ms = 0.f;
  uint32_t launches = 0;
  uint64_t units = 0, algo_bytes = 0;
  cudaEvent_t e0 = nullptr, e1 = nullptr;
};

struct State {
  humid_gpu_config cfg{};
  int device = 0;
  cudaStream_t stream = nullptr;
  bool finalized = false;

  // pass 1
  uint64_t n_tuples = 0;
  DevBuf<uint64_t> tuple_key;
  // words of 32..63 letters (wide.cu): key = hi << 64 | lo, lo in tuple_key / word_key, hi here;
  // the word of a tuple is remembered by rank (there is no table to ask)
  bool wide() const { return cfg.word_length > HUMID_MAX_NARROW_WORD_LENGTH; }
  DevBuf<uint64_t> tuple_hi, word_hi, wide_k0, wide_k1, wide_k2;
  DevBuf<uint32_t> tuple_rank, wide_v0, wide_v1;
  const uint32_t *wide_sorted_index = nullptr, *wide_pos_rank = nullptr;  // inside wide_v0 / wide_v1
  bool wide_ranks_ready = false;                                         // tuple_rank has been filled for this job
  // H2D staging for add_batch: two buffers, uploads on their own stream so that batch i+1
  // travels while batch i is packed and counted
  DevBuf<uint8_t> stage_chars[2];
  DevBuf<uint16_t> stage_seg_len[2];
  cudaStream_t copy_stream = nullptr;
  cudaEvent_t ev_copied[2] = {nullptr, nullptr}, ev_consumed[2] = {nullptr, nullptr};
  bool stage_busy[2] = {false, false};
  uint64_t n_batches = 0;
  // words counted during add_batch (expected_tuples given): tuples [0, early_inserted) are in the table
  bool early_table = false;
  uint64_t early_inserted = 0;
  DevBuf<unsigned long long> counters;  // [0] usable, [1] unique, [2] entries, [3..] misc
  uint64_t usable = 0, unique = 0, n_entries = 0, n_clusters = 0;

  // word table
  DevBuf<Slot> slots;
  DevBuf<uint64_t> part_key;    // tuples grouped by table region (words.cu: table_insert)
  DevBuf<uint32_t> part_index;
  uint64_t table_cap = 0;  // power of two
  DevBuf<uint4> filter;    // fingerprint buckets in front of the table (common.cuh)
  uint32_t filter_buckets = 0;
  DevBuf<uint32_t> word_flags;  // Hamming join: bit b = the word's block-b group was too large to walk
  uint64_t n_flagged = 0;
  bool ranks_in_table = false;  // slot.count holds the word's rank (stage_ensure_lookup)
  Lookup lookup() const {
    return Lookup{slots.p, table_cap - 1, filter_buckets ? filter.p : nullptr, filter_buckets, nullptr, 0};
  }

  // what humid_gpu_counts reports / what keep_bits covers (the whole job in shard mode)
  uint64_t report_tuples = 0, report_usable = 0;

  // shard mode (include/humid_gpu_shard.h): the table is built from the tuples this shard
  // RECEIVED (keys of its range from every rank, with their global input positions) instead
  // of the rank's own tuples; the word columns then hold the shard's words.
  bool shard_mode = false;
  uint64_t tuple_base = 0, n_shard_words = 0;
  DevBuf<uint64_t> shard_keys;
  DevBuf<uint32_t> shard_index;
  bool alias_imports = false;           // import_tuples keeps the caller's buffers instead of copying them
  struct GroupLink *group = nullptr;    // group.cu: the handle is one rank of a multi-GPU job run by the library
  const uint64_t *src_keys = nullptr;   // table input override (nullptr: tuple_key / implicit index)
  const uint32_t *src_index = nullptr;
  uint64_t src_n = 0;
  // clustering across shards (shard.cu): open components are replayed in a second State
  State *open = nullptr;
  uint64_t n_open_local = 0, n_open_entries_local = 0, open_offset = 0, n_open_total = 0, id_offset = 0;
  uint64_t report_unique = 0, report_clusters = 0;
  DevBuf<uint32_t> is_open, open_pos, open_ent, open_rank, open_first, seed_first, cluster_first, word_cid;
  uint32_t cluster_col_base = 0;     // stage_cluster: columns are rank + base
  bool cluster_skip_open = false;    // ... and components with a neighbour outside [0, unique) are left for later
  DevBuf<uint32_t> open_flag, comp_open;
  DevBuf<uint64_t> join_entries, join_entries_out;  // join_match: directed entries, then grouped by owner
  uint64_t n_join_entries = 0;
  uint32_t part = 0, n_parts = 1;
  uint64_t splitters[16] = {0};
  uint64_t bases[17] = {0};

  // rank-ordered columns
  DevBuf<uint64_t> word_key, key_alt;
  DevBuf<uint32_t> word_slot, slot_alt;
  DevBuf<uint64_t> compact_key;   // the table's words in compaction (slot) order: key ...
  DevBuf<uint4> compact_rec;      // ... the whole slot {key lo, key hi, Leaf::count, first tuple} ...
  DevBuf<uint32_t> compact_prefix;  // ... the key's 32 leading bits (what sort_words sorts) ...
  DevBuf<uint32_t> compact_slot;  // ... and slot number
  DevBuf<uint32_t> prefix_a, prefix_b;  // sort_words: 32 leading key bits of every word, double buffer
  bool word_slot_is_index = false;  // word_slot still holds compaction positions (translated on first look-up set-up)
  DevBuf<uint32_t> word_count, word_first;

  // neighbour lists
  DevBuf<uint64_t> entries, entries_alt;  // (src rank << 32 | payload), see neighbours.cu
  DevBuf<uint32_t> entry_val, entry_val_alt;
  DevBuf<uint32_t> deg;                   // [U] list lengths (Hamming mode: counted by the probe kernels)
  uint64_t n_pairs = 0;                   // undirected pairs in `entries` (Hamming mode)
  DevBuf<uint64_t> row_ptr;               // [U+1]
  DevBuf<uint32_t> col;                   // [n_entries]

  // clustering
  const uint64_t *cluster_pairs = nullptr;  // undirected pairs (rank << 32 | rank) behind the CSR, when the search left them
  uint64_t n_cluster_pairs = 0;
  bool cluster_no_ids = false;  // stage_cluster stops before numbering the clusters (the gathered open components)
  uint32_t cluster_owner_mod = 0, cluster_owner_me = 0;  // != 0: only components with root % mod == me are clustered here
  uint32_t open_owner_mod = 0, open_owner_me = 0;        // shard mode: the same for the gathered open components ...
  std::function<void(State &)> after_open_cluster;       // ... whose results the group then exchanges (group.cu)
  const uint64_t *cluster_entries = nullptr;  // directed entries (local row << 32 | column + base) behind the CSR (shard mode)
  uint64_t n_cluster_entries = 0;
  DevBuf<uint32_t> big_up, big_t, big_cursor, big_lo, big_hi, big_src, big_level;  // cluster_big.cu
  DevBuf<uint64_t> big_keys, big_keys_alt;
  DevBuf<uint32_t> comp, label, members, seg_start, seg_end, work_a, work_b, work_c;
  DevBuf<uint32_t> seed_flag, seed_index;     // seed_index = exclusive scan of seed_flag
  DevBuf<uint64_t> seed_size;                 // Cluster::size, indexed by seed rank
  DevBuf<uint32_t> seed_max;                  // rank of Cluster::maxLeaf, indexed by seed rank
  DevBuf<uint64_t> cluster_size;              // compacted by id-1
  DevBuf<uint32_t> cluster_max;

  // outputs
  DevBuf<uint32_t> keep_bits;  // ceil(R/32) words
  HostBuf<uint32_t> keep_host; // the same bitmap, copied at the end of finalize (page-locked: full PCIe rate)
  uint64_t keep_host_words = 0;
  DevBuf<uint32_t> out_ids;
  std::vector<uint64_t> hist_k[3], hist_v[3];
  bool hist_done[3] = {false, false, false};
  DevBuf<unsigned long long> hist_dev;    // finalize: bins + listed large values of the three histograms
  HostBuf<unsigned long long> hist_host;  // ... on their way to the host with the keep bitmap
  bool hist_pending = false;

  Scratch scratch, scan_scratch;
  std::vector<StageStat> stats;
  std::string last_error;

  State() = default;
  State(const State &) = delete;
  State &operator=(const State &) = delete;
  ~State() { delete open; }  // every DevBuf / Scratch member frees itself; streams and events belong to the ABI layer

  // Per-kernel(-family) device timing: CUDA events on `stream` around the launches.
  size_t timer_begin(const char *name, uint64_t units, uint64_t algo_bytes) {
    StageStat st;
    st.name = name;
    st.units = units;
    st.algo_bytes = algo_bytes;
    HUMID_CUDA(cudaEventCreate(&st.e0));
    HUMID_CUDA(cudaEventCreate(&st.e1));
    HUMID_CUDA(cudaEventRecord(st.e0, stream));
    stats.push_back(st);
    return stats.size() - 1;
  }
  void timer_end(size_t idx, uint32_t launches) {
    stats[idx].launches = launches;
    HUMID_CUDA(cudaEventRecord(stats[idx].e1, stream));
  }
};

// RAII form: `{ Timed t(s, "probe", U, bytes); kernel<<<...>>>(); t.launches = 1; }`
struct Timed {
  State &s;
  size_t idx;
  uint32_t launches = 1;
  Timed(State &st, const char *name, uint64_t units, uint64_t algo_bytes)
      : s(st), idx(st.timer_begin(name, units, algo_bytes)) {}
  ~Timed() { cudaEventRecord(s.stats[idx].e1, s.stream); s.stats[idx].launches = launches; }
};

// ---- stage launchers (each returns the number of kernels it launched) --------------------
int stage_pack(State &s, const uint8_t *d_chars, const uint16_t *d_seg_len, uint64_t first, uint64_t n);
int stage_pack4(State &s, const uint8_t *d_nibbles, const uint16_t *d_seg_len, uint64_t first, uint64_t n);  // 4-bit wire format
int stage_build_table(State &s, bool with_lookup = true);  // insert + compact + sort + columns (+ look-up)
int stage_ensure_lookup(State &s);    // ranks into the table slots + fingerprint filter, once per job
int stage_early_insert(State &s, uint64_t first, uint64_t n);  // add_batch: count tuples [first, first+n)
int stage_build_filter(State &s);      // fingerprint filter over word_key[0..unique)
// Hamming block join (neighbours.cu), shared with the shard phases
constexpr int HAM_MAX_D = 8;
struct Blocks {
  uint32_t n_blocks;
  uint64_t mask[HAM_MAX_D + 1];  // bits of block b (block 0 = first letters = highest bits)
};
// what the single-GPU searches join on (neighbours.cu: make_join_plan)
struct JoinPlan {
  Blocks B;
  int n_sorts[HAM_MAX_D + 1];
  int lo_bit[HAM_MAX_D + 1][2], hi_bit[HAM_MAX_D + 1][2];
};

JoinPlan make_join_plan(uint32_t n, uint32_t d);
int join_plan_order(State &s, const JoinPlan &P, uint32_t b, uint64_t U, const uint64_t *&keys, const uint32_t *&rank);
constexpr int JOIN_PAIRS_DEG = 0, JOIN_DIRECTED = 1, JOIN_CANDIDATES = 2;
constexpr uint32_t JOIN_NO_CAP = 0xFFFFFFFFu;
Blocks make_blocks(uint32_t n, uint32_t d, int lo_bit[], int hi_bit[]);  // arrays of HAM_MAX_D + 1
void launch_hamming_join(State &s, int emit, const uint64_t *keys, const uint32_t *rank, uint64_t n_words, const Blocks &B,
                         uint32_t block, uint32_t group_cap, uint64_t *out, uint64_t cap, unsigned long long *counter,
                         uint32_t *deg, uint32_t col_base, uint32_t *flags, unsigned long long *n_flagged);
int stage_neighbours_hamming(State &s);
// words of 32..63 letters: bits of block b in the two key halves (block 0 = first letters)
struct WideBlocks {
  uint32_t n_blocks;
  uint64_t mask_hi[HAM_MAX_D + 1], mask_lo[HAM_MAX_D + 1];
};
void launch_hamming_join_wide(State &s, const uint64_t *keys_hi, const uint64_t *keys_lo, const uint32_t *rank,
                              uint64_t n_words, const WideBlocks &B, uint32_t block, uint64_t *out, uint64_t cap,
                              unsigned long long *counter, uint32_t *deg);
// wide.cu — the same stages for words of 32..63 letters
int stage_build_words_wide(State &s);        // distinct words, counts, first tuples, tuple -> rank (by sorting)
int stage_neighbours_hamming_wide(State &s);
int stage_cluster_ids_wide(State &s, uint64_t first, uint64_t n, uint32_t *d_out);
int stage_neighbours_edit(State &s);
int stage_csr(State &s);               // sort directed entries, row_ptr / col (edit and shard modes)
int stage_csr_pairs(State &s);         // Hamming mode: undirected pairs + degrees -> row_ptr / col
int stage_csr_directed(State &s);      // shard mode: directed entries + degrees -> row_ptr / col
int stage_cluster(State &s);           // components, replay, ids, sizes, maxLeaf
int stage_cluster_number(State &s);    // seed flags -> cluster ids, compacted size / maxLeaf
int stage_keep(State &s);              // keep bitmap
int stage_cluster_ids(State &s, uint64_t first, uint64_t n, uint32_t *d_out);
int stage_histogram(State &s, int which);        // one histogram, synchronously (shard mode, overflow of the eager form)
int stage_histograms_enqueue(State &s);          // all three behind the clustering, results to page-locked host memory
void stage_histograms_collect(State &s);         // ... turned into (key, value) lists after the stream has been synchronised

}  // namespace humid
