/* humid_gpu_shard.h — phases of the deduplication hot path for one-process-per-GPU runs.
 *
 * The reference (src/humid.cc:369-409) is one process with one global trie; nothing in it
 * corresponds to these calls.  They cut humid_gpu_finalize() (include/humid_gpu.h) at the
 * points where data has to move between GPUs, so that a host (humid_b200/sharded.py over
 * torch.distributed / NCCL) can do the exchanges in between:
 *
 *   add_batch*            each rank packs its own slice of the read tuples          (makeWord)
 *   local_words           per-rank distinct words with counts / first tuple         (Trie::add)
 *   sample / split        splitters for a range partition of the word space
 *   export_words  ==all-to-all-v==>  import_words     owner merges counts, min(first)
 *   set_layout            global rank = bases[part] + local rank                    (Trie::walk order)
 *   probe_count / probe   variants owned by this shard are looked up in place, the
 *                         others become queries  ==all-to-all-v==>  answer  ==back==>  apply_answers
 *   build_rows            neighbour lists of the shard's words (columns = global ranks)
 *   export_graph  ==all-gather==>  import_graph       clustering, keep mask, histograms
 *
 * All pointers named d_* are device memory of the handle's GPU, allocated by the caller
 * (sizes are known from the preceding call); everything else is host memory.  Same error
 * conventions as humid_gpu.h.  Only Hamming mode (any -m) and edit distance 1 are sharded.
 */
#ifndef HUMID_GPU_SHARD_H
#define HUMID_GPU_SHARD_H

#include "humid_gpu.h"

#ifdef __cplusplus
extern "C" {
#endif

#define HUMID_MAX_PARTS 16

/* Distinct words of the tuples added so far, in key order.  tuple_offset = global index of
 * this rank's first tuple (the `first` column is global). */
int humid_gpu_shard_local_words(humid_gpu *h, uint64_t tuple_offset, uint64_t *n_local_words,
                                uint64_t *n_usable);
/* n_samples evenly spaced keys of the local words (n_local_words > 0). */
int humid_gpu_shard_sample_keys(humid_gpu *h, uint32_t n_samples, uint64_t *keys_out);
/* counts[p] = local words with splitters[p-1] <= key < splitters[p] (n_parts-1 splitters). */
int humid_gpu_shard_split(humid_gpu *h, const uint64_t *splitters, uint32_t n_parts, uint64_t *counts);
int humid_gpu_shard_export_words(humid_gpu *h, uint64_t *d_keys, uint32_t *d_counts, uint32_t *d_first);
/* Merge n received records into this shard's words; d_local_rank[i] = the shard-local rank
 * of record i (returned to its sender so that tuples can be resolved later). */
int humid_gpu_shard_import_words(humid_gpu *h, const uint64_t *d_keys, const uint32_t *d_counts,
                                 const uint32_t *d_first, uint64_t n, uint32_t *d_local_rank,
                                 uint64_t *n_shard_words);
/* bases has n_parts+1 entries (exclusive prefix sums of the shard sizes). */
int humid_gpu_shard_set_layout(humid_gpu *h, uint32_t part, uint32_t n_parts, const uint64_t *splitters,
                               const uint64_t *bases);
/* d_global_rank[i] = global rank of exported local word i (from the owners' replies). */
int humid_gpu_shard_set_word_ranks(humid_gpu *h, const uint32_t *d_global_rank, uint64_t n_local_words);

/* Number of variant queries this shard will send to every part (counts[part] = 0). */
int humid_gpu_shard_probe_count(humid_gpu *h, uint64_t *counts);
/* Local look-ups + fill of the query buffers; offsets (n_parts+1) = exclusive prefix of counts. */
int humid_gpu_shard_probe(humid_gpu *h, uint64_t *d_query_keys, uint32_t *d_query_src, const uint64_t *offsets);
/* d_answers[i] = global rank of query i's word, or 0xFFFFFFFF. */
int humid_gpu_shard_answer(humid_gpu *h, const uint64_t *d_query_keys, const uint32_t *d_query_src, uint64_t n,
                           uint32_t *d_answers);
int humid_gpu_shard_apply_answers(humid_gpu *h, const uint32_t *d_query_src, const uint32_t *d_answers, uint64_t n);
int humid_gpu_shard_build_rows(humid_gpu *h, uint64_t *n_entries);
int humid_gpu_shard_export_graph(humid_gpu *h, uint64_t *d_keys, uint32_t *d_counts, uint32_t *d_first,
                                 uint32_t *d_degree, uint32_t *d_col);
/* The gathered graph (all shards concatenated in part order) -> clusters, keep mask,
 * histograms; afterwards the handle answers humid_gpu_counts / histogram / keep_mask /
 * debug_* for the WHOLE job (tuple indices are global). */
int humid_gpu_shard_import_graph(humid_gpu *h, uint64_t n_words, uint64_t n_entries, uint64_t n_tuples,
                                 uint64_t n_usable, const uint64_t *d_keys, const uint32_t *d_counts,
                                 const uint32_t *d_first, const uint32_t *d_degree, const uint32_t *d_col);
/* Cluster ids of this rank's own tuples (local tuple indices), 0 for filtered tuples. */
int humid_gpu_shard_cluster_ids(humid_gpu *h, uint64_t first_local_tuple, uint64_t n, uint32_t *ids);

#ifdef __cplusplus
}
#endif
#endif /* HUMID_GPU_SHARD_H */
