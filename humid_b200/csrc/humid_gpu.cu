// C ABI of libhumid_gpu (include/humid_gpu.h): handle management, argument checks, stage
// sequencing and timing.  All compute lives in the stage_* launchers (CUDA kernels); there is
// deliberately no host implementation of any stage — without a CUDA device every call fails.
#include <cstring>
#include <functional>

#include "state.cuh"

using namespace humid;

#include "../../include/humid_gpu_shard.h"
#include "abi_internal.cuh"

void humid::resolve_stats(State &s) {
  for (auto &st : s.stats) {
    if (!st.e0) continue;
    HUMID_CUDA(cudaEventSynchronize(st.e1));
    HUMID_CUDA(cudaEventElapsedTime(&st.ms, st.e0, st.e1));
    cudaEventDestroy(st.e0);
    cudaEventDestroy(st.e1);
    st.e0 = st.e1 = nullptr;
  }
}

static void clear_stats(State &s) {
  for (auto &st : s.stats) {
    if (st.e0) cudaEventDestroy(st.e0);
    if (st.e1) cudaEventDestroy(st.e1);
  }
  s.stats.clear();
}

extern "C" {

int humid_gpu_abi_version(void) { return HUMID_GPU_ABI_VERSION; }

void *humid_gpu_stream(humid_gpu *h) { return h ? (void *)h->s.stream : nullptr; }

const char *humid_gpu_last_error(const humid_gpu *h) { return h ? h->s.last_error.c_str() : create_error().c_str(); }

int humid_gpu_create(const humid_gpu_config *cfg, humid_gpu **out) {
  if (!cfg || !out) return fail(nullptr, HUMID_GPU_ERR_ARG, "humid_gpu_create: null argument");
  *out = nullptr;
  if (cfg->word_length == 0 || cfg->word_length > HUMID_MAX_WORD_LENGTH)
    return fail(nullptr, HUMID_GPU_ERR_UNSUPPORTED,
                "word length must be 1.." + std::to_string(HUMID_MAX_WORD_LENGTH) + " (at most two 64-bit keys per word)");
  if (cfg->word_length > HUMID_MAX_NARROW_WORD_LENGTH && cfg->edit && cfg->distance >= 2)
    return fail(nullptr, HUMID_GPU_ERR_UNSUPPORTED,
                "edit distance >= 2 needs a word length of at most " + std::to_string(HUMID_MAX_NARROW_WORD_LENGTH));
  if (cfg->n_segments == 0 || cfg->n_segments > HUMID_MAX_SEGMENTS)
    return fail(nullptr, HUMID_GPU_ERR_ARG, "n_segments out of range");
  uint64_t sum = 0;
  for (uint32_t i = 0; i < cfg->n_segments; i++) sum += cfg->seg_take[i];
  if (sum != cfg->word_length) return fail(nullptr, HUMID_GPU_ERR_ARG, "seg_take must sum to word_length");
  humid_gpu *h = new (std::nothrow) humid_gpu();
  if (!h) return fail(nullptr, HUMID_GPU_ERR_MEMORY, "host allocation failed");
  h->s.cfg = *cfg;
  int rc = guarded(h, [&] {
    State &s = h->s;
    int count = 0;
    HUMID_CUDA(cudaGetDeviceCount(&count));
    if (count == 0) throw CudaError{"no CUDA device (there is no CPU fallback)"};
    if (cfg->device >= 0) HUMID_CUDA(cudaSetDevice(cfg->device));
    HUMID_CUDA(cudaGetDevice(&s.device));
    // HUMID_GPU_L2_FETCH=32|64|128: the device's L2 fetch granularity (cudaLimitMaxL2FetchGranularity), for experiments with
    // the random-access stages; unset = whatever the context has
    if (const char *e = std::getenv("HUMID_GPU_L2_FETCH")) {
      const long v = std::strtol(e, nullptr, 10);
      if (v == 32 || v == 64 || v == 128) HUMID_CUDA(cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, (size_t)v));
    }
    HUMID_CUDA(cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking));
    HUMID_CUDA(cudaStreamCreateWithFlags(&s.copy_stream, cudaStreamNonBlocking));
    for (int b = 0; b < 2; b++) {
      HUMID_CUDA(cudaEventCreateWithFlags(&s.ev_copied[b], cudaEventDisableTiming));
      HUMID_CUDA(cudaEventCreateWithFlags(&s.ev_consumed[b], cudaEventDisableTiming));
    }
    s.counters.ensure(64);
    HUMID_CUDA(cudaMemsetAsync(s.counters.p, 0, 64 * sizeof(unsigned long long), s.stream));
    if (cfg->expected_tuples) s.tuple_key.ensure(cfg->expected_tuples);
    if (cfg->expected_tuples && s.wide()) s.tuple_hi.ensure(cfg->expected_tuples);
    HUMID_CUDA(cudaStreamSynchronize(s.stream));
  });
  if (rc != HUMID_GPU_OK) {
    create_error() = h->s.last_error;
    delete h;
    return rc;
  }
  *out = h;
  return HUMID_GPU_OK;
}

void humid_gpu_destroy(humid_gpu *h) {
  if (!h) return;
  State &s = h->s;
  cudaSetDevice(s.device);
  if (s.copy_stream) cudaStreamSynchronize(s.copy_stream);
  if (s.stream) cudaStreamSynchronize(s.stream);
  clear_stats(s);
  if (s.open) clear_stats(*s.open);
  group_release_handle(s);
  for (int b = 0; b < 2; b++) {
    if (s.ev_copied[b]) cudaEventDestroy(s.ev_copied[b]);
    if (s.ev_consumed[b]) cudaEventDestroy(s.ev_consumed[b]);
  }
  cudaStream_t stream = s.stream, copy_stream = s.copy_stream;
  delete h;  // device memory: every DevBuf / Scratch of the State (and of its open-component State) frees itself
  if (copy_stream) cudaStreamDestroy(copy_stream);
  if (stream) cudaStreamDestroy(stream);
}

int humid_gpu_reset(humid_gpu *h) {
  if (!h) return HUMID_GPU_ERR_ARG;
  return guarded(h, [&] {
    State &s = h->s;
    HUMID_CUDA(cudaSetDevice(s.device));
    HUMID_CUDA(cudaStreamSynchronize(s.copy_stream));
    HUMID_CUDA(cudaStreamSynchronize(s.stream));
    clear_stats(s);
    s.early_table = false;
    s.early_inserted = 0;
    s.stage_busy[0] = s.stage_busy[1] = false;
    s.n_tuples = s.usable = s.unique = s.n_entries = s.n_clusters = 0;
    s.finalized = false;
    s.shard_mode = false;
    s.cluster_skip_open = false;
    s.cluster_col_base = 0;
    s.cluster_pairs = nullptr;
    s.n_cluster_pairs = 0;
    s.cluster_entries = nullptr;
    s.n_cluster_entries = 0;
    s.report_tuples = s.report_usable = s.tuple_base = s.n_shard_words = s.src_n = 0;
    s.src_keys = nullptr;
    s.src_index = nullptr;
    s.part = 0;
    s.n_parts = 1;
    for (int i = 0; i < 3; i++) s.hist_done[i] = false;
    s.hist_pending = false;
    group_reset(s);
    HUMID_CUDA(cudaMemsetAsync(s.counters.p, 0, 64 * sizeof(unsigned long long), s.stream));
  });
}

int humid_gpu_host_alloc(humid_gpu *h, size_t bytes, void **out) {
  if (!h || !out) return HUMID_GPU_ERR_ARG;
  *out = nullptr;
  return guarded(h, [&] {
    HUMID_CUDA(cudaSetDevice(h->s.device));
    HUMID_CUDA(cudaHostAlloc(out, bytes ? bytes : 1, cudaHostAllocDefault));
  });
}

int humid_gpu_host_free(humid_gpu *h, void *p) {
  if (!h) return HUMID_GPU_ERR_ARG;
  if (!p) return HUMID_GPU_OK;
  return guarded(h, [&] { HUMID_CUDA(cudaFreeHost(p)); });
}

int humid_gpu_host_register(humid_gpu *h, void *p, size_t bytes) {
  if (!h || !p) return HUMID_GPU_ERR_ARG;
  return guarded(h, [&] {
    HUMID_CUDA(cudaSetDevice(h->s.device));
    HUMID_CUDA(cudaHostRegister(p, bytes, cudaHostRegisterDefault));
  });
}

int humid_gpu_host_unregister(humid_gpu *h, void *p) {
  if (!h) return HUMID_GPU_ERR_ARG;
  if (!p) return HUMID_GPU_OK;
  return guarded(h, [&] { HUMID_CUDA(cudaHostUnregister(p)); });
}

static int add_batch_impl(humid_gpu *h, const uint8_t *chars, const uint16_t *seg_len, uint64_t first, uint64_t n,
                          bool on_device, bool packed4 = false) {
  if (!h) return HUMID_GPU_ERR_ARG;
  State &s = h->s;
  if (s.finalized) return fail(h, HUMID_GPU_ERR_ARG, "add_batch after finalize (call humid_gpu_reset first)");
  // HUMID_GROUP_BATCHES: `first` is the batch's position in the JOB; this rank stores it behind its earlier batches
  const bool global_first = group_batches_numbering(s);
  const uint64_t global_pos = first;
  if (global_first) first = s.n_tuples;
  if (first != s.n_tuples) return fail(h, HUMID_GPU_ERR_ARG, "batches must be added in input order, without gaps");
  if (n && !chars) return fail(h, HUMID_GPU_ERR_ARG, "chars is null");
  if (s.n_tuples + n >= 0xFFFFFFFFull) return fail(h, HUMID_GPU_ERR_UNSUPPORTED, "more than 2^32-2 tuples");
  if (n == 0) return HUMID_GPU_OK;
  return guarded(h, [&] {
    HUMID_CUDA(cudaSetDevice(s.device));
    // bytes of one tuple on the wire: one per character, or two characters per byte
    const uint32_t wl = packed4 ? (s.cfg.word_length + 1) / 2 : s.cfg.word_length, ns = s.cfg.n_segments;
    s.tuple_key.grow(s.n_tuples + n, s.n_tuples, s.stream);
    if (s.wide()) s.tuple_hi.grow(s.n_tuples + n, s.n_tuples, s.stream);
    const uint8_t *d_chars = chars;
    const uint16_t *d_len = seg_len;
    const int b = (int)(s.n_batches & 1);
    if (!on_device) {
      // upload on the copy stream into staging buffer b; the kernels that read that buffer two
      // batches ago must be done, and this batch's kernels wait for the upload
      if (s.stage_busy[b]) HUMID_CUDA(cudaStreamWaitEvent(s.copy_stream, s.ev_consumed[b], 0));
      if (s.stage_chars[b].cap < n * wl || (seg_len && s.stage_seg_len[b].cap < n * ns)) {
        HUMID_CUDA(cudaStreamSynchronize(s.stream));  // reallocation: nothing may still read the old buffer
        s.stage_chars[b].ensure(n * wl);
        if (seg_len) s.stage_seg_len[b].ensure(n * ns);
      }
      HUMID_CUDA(cudaMemcpyAsync(s.stage_chars[b].p, chars, n * wl, cudaMemcpyHostToDevice, s.copy_stream));
      d_chars = s.stage_chars[b].p;
      if (seg_len) {
        HUMID_CUDA(cudaMemcpyAsync(s.stage_seg_len[b].p, seg_len, n * ns * sizeof(uint16_t), cudaMemcpyHostToDevice,
                                   s.copy_stream));
        d_len = s.stage_seg_len[b].p;
      }
      HUMID_CUDA(cudaEventRecord(s.ev_copied[b], s.copy_stream));
      HUMID_CUDA(cudaStreamWaitEvent(s.stream, s.ev_copied[b], 0));
    }
    // makeWord: n ASCII bytes in, one 8-byte key out; then Trie::add when the table is already sized
    if (packed4) stage_pack4(s, d_chars, d_len, first, n);
    else stage_pack(s, d_chars, d_len, first, n);
    stage_early_insert(s, first, n);
    if (!on_device) {
      HUMID_CUDA(cudaEventRecord(s.ev_consumed[b], s.stream));
      s.stage_busy[b] = true;
      s.n_batches++;
    }
    if (global_first) group_note_batch(s, s.n_tuples, global_pos, n);
    s.n_tuples += n;
    if (!on_device) HUMID_CUDA(cudaEventSynchronize(s.ev_copied[b]));  // the caller may reuse its buffers
  });
}

int humid_gpu_add_batch(humid_gpu *h, const uint8_t *chars, const uint16_t *seg_len, uint64_t first_tuple_index,
                        uint64_t n_tuples) {
  return add_batch_impl(h, chars, seg_len, first_tuple_index, n_tuples, false);
}

int humid_gpu_add_batch_packed4(humid_gpu *h, const uint8_t *nibbles, const uint16_t *seg_len, uint64_t first_tuple_index,
                                uint64_t n_tuples) {
  return add_batch_impl(h, nibbles, seg_len, first_tuple_index, n_tuples, false, true);
}

int humid_gpu_add_batch_device(humid_gpu *h, const uint8_t *d_chars, const uint16_t *d_seg_len,
                               uint64_t first_tuple_index, uint64_t n_tuples) {
  return add_batch_impl(h, d_chars, d_seg_len, first_tuple_index, n_tuples, true);
}

int humid_gpu_finalize(humid_gpu *h) {
  if (!h) return HUMID_GPU_ERR_ARG;
  State &s = h->s;
  if (s.finalized) return fail(h, HUMID_GPU_ERR_ARG, "finalize called twice");
  if (group_active(s)) return humid_gpu_group_finalize(h);  // a rank of a group: the collective form
  return guarded(h, [&] {
    HUMID_CUDA(cudaSetDevice(s.device));
    const bool edit2 = s.cfg.edit && s.cfg.distance >= 2;
    if (s.shard_mode) throw CudaError{"finalize on a handle that is in shard mode"};
    if (s.wide()) stage_build_words_wide(s);
    else stage_build_table(s, false);  // key -> rank look-ups are set up by the stages that need them
    s.report_tuples = s.n_tuples;
    s.report_usable = s.usable;
    if (s.wide()) {
      stage_neighbours_hamming_wide(s);
      stage_csr_pairs(s);
    } else if (edit2) {
      stage_neighbours_edit(s);
      stage_csr(s);
    } else {
      stage_neighbours_hamming(s);
      stage_csr_pairs(s);
    }
    stage_cluster(s);
    stage_keep(s);
    stage_histograms_enqueue(s);  // runStatistics / clusterStats (src/humid.cc:301-315, src/cluster.cc:89-95)
    // writeFiltered asks for the whole bitmap next: it travels while the stream drains
    s.keep_host_words = (s.report_tuples + 31) / 32 + 2;
    s.keep_host.ensure(s.keep_host_words);
    HUMID_CUDA(cudaMemcpyAsync(s.keep_host.p, s.keep_bits.p, s.keep_host_words * sizeof(uint32_t), cudaMemcpyDeviceToHost,
                               s.stream));
    HUMID_CUDA(cudaStreamSynchronize(s.stream));
    stage_histograms_collect(s);
    resolve_stats(s);
    s.finalized = true;
  });
}

int humid_gpu_counts(humid_gpu *h, uint64_t *total, uint64_t *usable, uint64_t *unique, uint64_t *clusters) {
  if (!h) return HUMID_GPU_ERR_ARG;
  if (!h->s.finalized) return fail(h, HUMID_GPU_ERR_ARG, "not finalized");
  if (total) *total = h->s.report_tuples;
  if (usable) *usable = h->s.report_usable;
  if (unique) *unique = h->s.shard_mode ? h->s.report_unique : h->s.unique;
  if (clusters) *clusters = h->s.shard_mode ? h->s.report_clusters : h->s.n_clusters;
  return HUMID_GPU_OK;
}

int humid_gpu_histogram(humid_gpu *h, int which, uint64_t *keys, uint64_t *vals, uint64_t *n) {
  if (!h || !n) return HUMID_GPU_ERR_ARG;
  State &s = h->s;
  if (!s.finalized) return fail(h, HUMID_GPU_ERR_ARG, "not finalized");
  if (which < 0 || which > 2) return fail(h, HUMID_GPU_ERR_ARG, "unknown histogram");
  int rc = HUMID_GPU_OK;
  if (!s.hist_done[which]) rc = guarded(h, [&] {
    HUMID_CUDA(cudaSetDevice(s.device));
    stage_histogram(s, which);
    resolve_stats(s);
  });
  if (rc != HUMID_GPU_OK) return rc;
  uint64_t m = s.hist_k[which].size();
  if (keys && vals) {
    if (*n < m) { *n = m; return fail(h, HUMID_GPU_ERR_ARG, "histogram buffer too small"); }
    std::memcpy(keys, s.hist_k[which].data(), m * sizeof(uint64_t));
    std::memcpy(vals, s.hist_v[which].data(), m * sizeof(uint64_t));
  }
  *n = m;
  return HUMID_GPU_OK;
}

int humid_gpu_keep_mask(humid_gpu *h, uint64_t first_tuple, uint64_t n, uint64_t *bits) {
  if (!h || (n && !bits)) return HUMID_GPU_ERR_ARG;
  State &s = h->s;
  if (!s.finalized) return fail(h, HUMID_GPU_ERR_ARG, "not finalized");
  if (s.shard_mode && !group_active(s)) return fail(h, HUMID_GPU_ERR_ARG, "use humid_gpu_shard_keep_bits in shard mode");
  if (first_tuple + n > (group_active(s) ? group_result_tuples(s) : s.report_tuples))
    return fail(h, HUMID_GPU_ERR_ARG, "tuple range out of bounds");
  if (first_tuple % 64) return fail(h, HUMID_GPU_ERR_ARG, "first_tuple must be a multiple of 64");
  if (n == 0) return HUMID_GPU_OK;
  const uint64_t words = (n + 63) / 64;
  std::memcpy(bits, s.keep_host.p + first_tuple / 32, words * 8);  // copied to the host by finalize
  if (n % 64) bits[words - 1] &= (~0ull) >> (64 - n % 64);
  return HUMID_GPU_OK;
}

int humid_gpu_cluster_ids(humid_gpu *h, uint64_t first_tuple, uint64_t n, uint32_t *ids) {
  if (!h || (n && !ids)) return HUMID_GPU_ERR_ARG;
  State &s = h->s;
  if (!s.finalized) return fail(h, HUMID_GPU_ERR_ARG, "not finalized");
  if (group_active(s)) {  // the ids of this rank's tuples were sent here by the owners of their words
    const uint32_t *d_ids = group_result_ids(s);
    if (!d_ids) return fail(h, HUMID_GPU_ERR_ARG, "call humid_gpu_group_collect_ids first (collective)");
    if (first_tuple + n > group_result_tuples(s)) return fail(h, HUMID_GPU_ERR_ARG, "tuple range out of bounds");
    if (n == 0) return HUMID_GPU_OK;
    return guarded(h, [&] {
      HUMID_CUDA(cudaSetDevice(s.device));
      HUMID_CUDA(cudaMemcpyAsync(ids, d_ids + first_tuple, n * sizeof(uint32_t), cudaMemcpyDeviceToHost, s.stream));
      HUMID_CUDA(cudaStreamSynchronize(s.stream));
    });
  }
  if (s.shard_mode) return fail(h, HUMID_GPU_ERR_ARG, "use humid_gpu_shard_cluster_ids in shard mode");
  if (first_tuple + n > s.n_tuples) return fail(h, HUMID_GPU_ERR_ARG, "tuple range out of bounds");
  if (n == 0) return HUMID_GPU_OK;
  return guarded(h, [&] {
    HUMID_CUDA(cudaSetDevice(s.device));
    s.out_ids.ensure(n);
    // K5: key (8) + slot (16) + label (4) + seed index (4) + id out (4)
    if (s.wide()) stage_cluster_ids_wide(s, first_tuple, n, s.out_ids.p);
    else stage_cluster_ids(s, first_tuple, n, s.out_ids.p);
    HUMID_CUDA(cudaMemcpyAsync(ids, s.out_ids.p, n * sizeof(uint32_t), cudaMemcpyDeviceToHost, s.stream));
    HUMID_CUDA(cudaStreamSynchronize(s.stream));
    resolve_stats(s);
  });
}

// ---- introspection for the parity tests ---------------------------------------------------
__global__ void word_cluster_ids_kernel(const uint32_t *label, const uint32_t *seed_index, uint64_t U, uint32_t *out) {
  uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r < U) out[r] = seed_index[label[r]] + 1u;
}

int humid_gpu_debug_words(humid_gpu *h, uint64_t cap, uint64_t *keys, uint32_t *counts, uint32_t *first_tuple,
                          uint32_t *cluster_ids) {
  if (!h) return HUMID_GPU_ERR_ARG;
  State &s = h->s;
  if (!s.finalized) return fail(h, HUMID_GPU_ERR_ARG, "not finalized");
  if (cap < s.unique) return fail(h, HUMID_GPU_ERR_ARG, "buffers too small");
  const uint64_t U = s.unique;
  if (U == 0) return HUMID_GPU_OK;
  return guarded(h, [&] {
    HUMID_CUDA(cudaSetDevice(s.device));
    if (keys) HUMID_CUDA(cudaMemcpyAsync(keys, s.word_key.p, U * 8, cudaMemcpyDeviceToHost, s.stream));
    if (counts) HUMID_CUDA(cudaMemcpyAsync(counts, s.word_count.p, U * 4, cudaMemcpyDeviceToHost, s.stream));
    if (first_tuple) HUMID_CUDA(cudaMemcpyAsync(first_tuple, s.word_first.p, U * 4, cudaMemcpyDeviceToHost, s.stream));
    if (cluster_ids && s.shard_mode) {
      HUMID_CUDA(cudaMemcpyAsync(cluster_ids, s.word_cid.p, U * 4, cudaMemcpyDeviceToHost, s.stream));
    } else if (cluster_ids) {
      s.out_ids.ensure(U);
      word_cluster_ids_kernel<<<grid_for(U, 256), 256, 0, s.stream>>>(s.label.p, s.seed_index.p, U, s.out_ids.p);
      HUMID_CUDA(cudaMemcpyAsync(cluster_ids, s.out_ids.p, U * 4, cudaMemcpyDeviceToHost, s.stream));
    }
    HUMID_CUDA(cudaStreamSynchronize(s.stream));
  });
}

int humid_gpu_debug_words_hi(humid_gpu *h, uint64_t cap, uint64_t *keys_hi) {
  if (!h || !keys_hi) return HUMID_GPU_ERR_ARG;
  State &s = h->s;
  if (!s.finalized) return fail(h, HUMID_GPU_ERR_ARG, "not finalized");
  if (cap < s.unique) return fail(h, HUMID_GPU_ERR_ARG, "buffers too small");
  const uint64_t U = s.unique;
  if (U == 0) return HUMID_GPU_OK;
  if (!s.wide()) {
    std::memset(keys_hi, 0, U * sizeof(uint64_t));
    return HUMID_GPU_OK;
  }
  return guarded(h, [&] {
    HUMID_CUDA(cudaSetDevice(s.device));
    HUMID_CUDA(cudaMemcpyAsync(keys_hi, s.word_hi.p, U * 8, cudaMemcpyDeviceToHost, s.stream));
    HUMID_CUDA(cudaStreamSynchronize(s.stream));
  });
}

int humid_gpu_debug_neighbours(humid_gpu *h, uint64_t *row_ptr, uint32_t *col, uint64_t *n_entries) {
  if (!h || !n_entries) return HUMID_GPU_ERR_ARG;
  State &s = h->s;
  if (!s.finalized) return fail(h, HUMID_GPU_ERR_ARG, "not finalized");
  if (col && *n_entries < s.n_entries) return fail(h, HUMID_GPU_ERR_ARG, "col buffer too small");
  *n_entries = s.n_entries;
  return guarded(h, [&] {
    HUMID_CUDA(cudaSetDevice(s.device));
    if (row_ptr) HUMID_CUDA(cudaMemcpyAsync(row_ptr, s.row_ptr.p, (s.unique + 1) * 8, cudaMemcpyDeviceToHost, s.stream));
    if (col && s.n_entries)
      HUMID_CUDA(cudaMemcpyAsync(col, s.col.p, s.n_entries * 4, cudaMemcpyDeviceToHost, s.stream));
    HUMID_CUDA(cudaStreamSynchronize(s.stream));
  });
}

int humid_gpu_debug_clusters(humid_gpu *h, uint64_t cap, uint64_t *sizes, uint32_t *max_leaf_rank) {
  if (!h) return HUMID_GPU_ERR_ARG;
  State &s = h->s;
  if (!s.finalized) return fail(h, HUMID_GPU_ERR_ARG, "not finalized");
  if (cap < s.n_clusters) return fail(h, HUMID_GPU_ERR_ARG, "buffers too small");
  if (s.n_clusters == 0) return HUMID_GPU_OK;
  return guarded(h, [&] {
    HUMID_CUDA(cudaSetDevice(s.device));
    if (sizes) HUMID_CUDA(cudaMemcpyAsync(sizes, s.cluster_size.p, s.n_clusters * 8, cudaMemcpyDeviceToHost, s.stream));
    if (max_leaf_rank)
      HUMID_CUDA(cudaMemcpyAsync(max_leaf_rank, s.cluster_max.p, s.n_clusters * 4, cudaMemcpyDeviceToHost, s.stream));
    HUMID_CUDA(cudaStreamSynchronize(s.stream));
  });
}

int humid_gpu_debug_sort_pairs(humid_gpu *h, const uint64_t *d_keys_in, const uint32_t *d_vals_in, uint64_t n,
                               int begin_bit, int end_bit, uint64_t *d_keys_out, uint32_t *d_vals_out) {
  if (!h) return HUMID_GPU_ERR_ARG;
  State &s = h->s;
  if (n == 0) return HUMID_GPU_OK;
  if (!d_keys_in || !d_keys_out || !d_vals_out || end_bit <= begin_bit || end_bit > 64)
    return fail(h, HUMID_GPU_ERR_ARG, "debug_sort_pairs: bad arguments");
  return guarded(h, [&] {
    HUMID_CUDA(cudaSetDevice(s.device));
    s.key_alt.ensure(n);
    s.entries_alt.ensure(n);
    s.entry_val.ensure(n);
    s.entry_val_alt.ensure(n);
    uint64_t *kb[2] = {s.key_alt.p, s.entries_alt.p};
    uint32_t *vb[2] = {s.entry_val.p, s.entry_val_alt.p};
    int launches = 0;
    int cur = d_vals_in ? radix_sort_bits_pairs(d_keys_in, d_vals_in, kb, vb, n, begin_bit, end_bit, s.scratch, s.stream,
                                                &launches)
                        : radix_sort_bits_indexed(d_keys_in, kb, vb, n, begin_bit, end_bit, s.scratch, s.stream,
                                                  &launches);
    HUMID_CUDA(cudaMemcpyAsync(d_keys_out, kb[cur], n * 8, cudaMemcpyDeviceToDevice, s.stream));
    HUMID_CUDA(cudaMemcpyAsync(d_vals_out, vb[cur], n * 4, cudaMemcpyDeviceToDevice, s.stream));
    HUMID_CUDA(cudaStreamSynchronize(s.stream));
  });
}

int humid_gpu_debug_partition(humid_gpu *h, int kind, const uint64_t *d_keys_in, uint64_t n, uint64_t arg,
                              uint32_t n_parts, const uint64_t *bases, uint32_t index_base, uint64_t *d_keys_out,
                              uint32_t *d_vals_out, uint64_t *counts) {
  if (!h) return HUMID_GPU_ERR_ARG;
  State &s = h->s;
  if (kind < 0 || kind > 2 || (n && (!d_keys_in || !d_keys_out)) || (kind != 2 && n && !d_vals_out) ||
      (kind != 0 && (!counts || n_parts == 0 || n_parts > 16)) || (kind == 2 && !bases))
    return fail(h, HUMID_GPU_ERR_ARG, "debug_partition: bad arguments");
  return guarded(h, [&] {
    HUMID_CUDA(cudaSetDevice(s.device));
    int launches = 0;
    if (kind == 0)
      partition_by_slot_u64(d_keys_in, nullptr, index_base, d_keys_out, d_vals_out, n, arg, s.scratch, s.stream, &launches);
    else if (kind == 1)
      partition_by_block_hash(d_keys_in, index_base, d_keys_out, d_vals_out, n, arg, n_parts, counts, s.scratch, s.stream);
    else
      partition_by_row_owner(d_keys_in, d_keys_out, n, bases, n_parts, counts, s.scratch, s.stream);
    HUMID_CUDA(cudaStreamSynchronize(s.stream));
  });
}

// No counterpart in the reference (include/humid_gpu.h).  Nothing on
// this path allocates: no DevBuf / Scratch / HostBuf is touched, failures are reported through the handle's message.
int humid_gpu_partition_by_slot_noalloc(humid_gpu *h, const uint64_t *d_keys_in, uint64_t n, uint64_t slot_mask,
                                        uint32_t index_base, uint64_t *d_keys_out, uint32_t *d_vals_out) {
  if (!h) return HUMID_GPU_ERR_ARG;
  State &s = h->s;
  if (n == 0) return HUMID_GPU_OK;
  if (!d_keys_in || !d_keys_out || !d_vals_out) return fail(h, HUMID_GPU_ERR_ARG, "partition_by_slot_noalloc: null buffer");
  if (n > HUMID_NOALLOC_MAX_ELEMENTS)
    return fail(h, HUMID_GPU_ERR_UNSUPPORTED,
                "partition_by_slot_noalloc: " + std::to_string(n) + " elements, the static scratch holds at most " +
                    std::to_string(HUMID_NOALLOC_MAX_ELEMENTS) + " (use humid_gpu_debug_partition, which may allocate)");
  try {
    HUMID_CUDA(cudaSetDevice(s.device));
    partition_by_slot_u64_noalloc(d_keys_in, index_base, d_keys_out, d_vals_out, n, slot_mask, s.stream);
    return HUMID_GPU_OK;
  } catch (const CudaError &e) {
    return fail(h, HUMID_GPU_ERR_CUDA, e.msg);
  }
}

int humid_gpu_stage_stats(humid_gpu *h, humid_gpu_stage_stat *out, uint32_t *n) {
  if (!h || !n) return HUMID_GPU_ERR_ARG;
  State &s = h->s;
  int rc = guarded(h, [&] { resolve_stats(s); });
  if (rc != HUMID_GPU_OK) return rc;
  uint32_t m = (uint32_t)s.stats.size();
  if (out) {
    if (*n < m) { *n = m; return fail(h, HUMID_GPU_ERR_ARG, "stats buffer too small"); }
    for (uint32_t i = 0; i < m; i++) {
      std::memset(&out[i], 0, sizeof(out[i]));
      std::strncpy(out[i].name, s.stats[i].name, sizeof(out[i].name) - 1);
      out[i].ms = s.stats[i].ms;
      out[i].launches = s.stats[i].launches;
      out[i].units = s.stats[i].units;
      out[i].algo_bytes = s.stats[i].algo_bytes;
    }
  }
  *n = m;
  return HUMID_GPU_OK;
}

}  // extern "C"
