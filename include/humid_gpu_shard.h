/* humid_gpu_shard.h — phases of the deduplication hot path for one-process-per-GPU runs.
 *
 * The reference (src/humid.cc:369-409) is one process with one global trie; nothing in it
 * corresponds to these calls.  They cut humid_gpu_finalize() (include/humid_gpu.h) at the
 * points where data has to move between GPUs, so that a host (humid_b200/sharded.py over
 * torch.distributed / NCCL) can do the exchanges in between:
 *
 *   add_batch*            each rank packs its own slice of the read tuples          (makeWord)
 *   begin / sample_keys   splitters for a range partition of the word space
 *   partition / export_tuples  ==all-to-all-v==>  import_tuples
 *                         the owner of a key range counts its words                 (Trie::add)
 *   set_layout            global rank = bases[part] + local rank                    (Trie::walk order)
 *   join_local            block join of block 0 inside the shard     (asymmetricHamming)
 *   per block b = 1..d:   join_export  ==all-to-all-v==>  join_match (the part that owns a block value joins
 *                         its group)  /  join_entries  ==all-to-all-v==>  join_import (entries back to the row owners)
 *   build_rows            neighbour lists of the shard's words (columns = global ranks)
 *   cluster_local         components inside the shard are clustered in place        (findClusters)
 *   export_open  ==all-gather==>  import_open   components that cross shards, replayed on every shard
 *   assign_ids  ==all-reduce==>  finish         cluster ids in global walk order
 *   keep_bits  ==all-reduce==>                  the job's keep mask
 *
 * All pointers named d_* are device memory of the handle's GPU, allocated by the caller
 * (sizes are known from the preceding call); everything else is host memory.  Same error
 * conventions as humid_gpu.h.  Only Hamming mode (any -m) and edit distance 1 are sharded.
 * Groups of the join are walked whatever their size here (no variant-probe fallback across shards).
 */
#ifndef HUMID_GPU_SHARD_H
#define HUMID_GPU_SHARD_H

#include "humid_gpu.h"

#ifdef __cplusplus
extern "C" {
#endif

#define HUMID_MAX_PARTS 16

/* Enter shard mode after the add_batch* calls.  tuple_offset = global input position of this
 * rank's first tuple (rank order == input order). */
int humid_gpu_shard_begin(humid_gpu *h, uint64_t tuple_offset, uint64_t *n_usable);
/* Up to n_samples evenly spaced usable tuple keys (splitter candidates); *n_out of them. */
int humid_gpu_shard_sample_keys(humid_gpu *h, uint32_t n_samples, uint64_t *keys_out, uint32_t *n_out);
/* counts[p] = usable local tuples whose key lies in [splitters[p-1], splitters[p])
 * (n_parts-1 ascending splitters).  The splitters are remembered for export_tuples. */
int humid_gpu_shard_partition(humid_gpu *h, const uint64_t *splitters, uint32_t n_parts, uint64_t *counts);
/* {key, global tuple index} of every usable local tuple, grouped by part; offsets (n_parts+1)
 * = exclusive prefix of the counts; order inside a part is unspecified. */
int humid_gpu_shard_export_tuples(humid_gpu *h, const uint64_t *offsets, uint64_t *d_keys, uint32_t *d_index);
/* The tuples of this shard's key range received from all ranks: distinct words with counts
 * and first (global) tuple, rank order, look-up table and filter — the single-GPU stages
 * restricted to the range. */
int humid_gpu_shard_import_tuples(humid_gpu *h, const uint64_t *d_keys, const uint32_t *d_index, uint64_t n,
                                  uint64_t *n_shard_words);
/* bases has n_parts+1 entries (exclusive prefix sums of the shard sizes). */
int humid_gpu_shard_set_layout(humid_gpu *h, uint32_t part, uint32_t n_parts, const uint64_t *splitters,
                               const uint64_t *bases);
/* Block 0 (the leading letters) of the Hamming block join, inside the shard: the splitters given to
 * partition / set_layout MUST be multiples of 4^(word_length - word_length / (distance + 1)), i.e. have all
 * letters after block 0 equal to 'A', so that no block-0 group straddles two shards.  Starts the
 * shard's entry list (local row, global column) and list lengths. */
int humid_gpu_shard_join_local(humid_gpu *h, uint64_t *n_entries);
/* Block b >= 1: every word of the shard as {key, global rank}, grouped by the part that joins its
 * block-b value (a hash of those letters); inside a part the words stay in rank order.  d_keys /
 * d_ranks hold n_shard_words elements; counts (n_parts) = words per part. */
int humid_gpu_shard_join_export(humid_gpu *h, uint32_t block, uint64_t *counts, uint64_t *d_keys, uint32_t *d_ranks);
/* The words received for block b from all parts, concatenated in part order (== ascending global
 * rank): joins them and keeps the directed entries {global row << 32 | global column} grouped by
 * the part owning the row; counts (n_parts) = entries per part. */
int humid_gpu_shard_join_match(humid_gpu *h, uint32_t block, const uint64_t *d_keys, const uint32_t *d_ranks, uint64_t n,
                               uint64_t *counts);
/* Copies the entries of the last join_match (sum of its counts) to d_entries. */
int humid_gpu_shard_join_entries(humid_gpu *h, uint64_t *d_entries);
/* Entries received for this shard's rows: appended to its entry list. */
int humid_gpu_shard_join_import(humid_gpu *h, const uint64_t *d_entries, uint64_t n);
int humid_gpu_shard_build_rows(humid_gpu *h, uint64_t *n_entries);
/* Clusters of the components that lie entirely inside this shard; the others ("open": some
 * member has a neighbour on another shard) are counted for the export below. */
int humid_gpu_shard_cluster_local(humid_gpu *h, uint64_t *n_open_words, uint64_t *n_open_entries);
/* Every word of the open components: global rank, count, first tuple, list length and list
 * (global ranks), words in rank order. */
int humid_gpu_shard_export_open(humid_gpu *h, uint32_t *d_rank, uint32_t *d_count, uint32_t *d_first, uint32_t *d_deg,
                                uint32_t *d_col);
/* The open words of ALL shards (concatenated in part order == rank order; this shard's own
 * ones start at open_offset): replays them and merges the result into the shard's clusters.
 * Returns the number of clusters seeded by words of this shard. */
int humid_gpu_shard_import_open(humid_gpu *h, uint64_t open_offset, uint64_t n_open, uint64_t n_entries,
                                const uint32_t *d_rank, const uint32_t *d_count, const uint32_t *d_first,
                                const uint32_t *d_deg, const uint32_t *d_col, uint64_t *n_clusters_local);
/* id_offset = clusters seeded on lower parts.  Writes the ids of the open seeds this shard owns
 * into d_open_ids (n_open entries, zero-initialised by the caller, summed over the shards). */
int humid_gpu_shard_assign_ids(humid_gpu *h, uint64_t id_offset, uint32_t *d_open_ids);
/* d_open_ids = the summed array.  The totals are what humid_gpu_counts reports afterwards;
 * humid_gpu_histogram / debug_* then describe THIS SHARD's words, lists and clusters. */
int humid_gpu_shard_finish(humid_gpu *h, const uint32_t *d_open_ids, uint64_t n_tuples, uint64_t n_usable,
                           uint64_t n_words, uint64_t n_clusters);
/* Bit t of d_bits (n_words32 32-bit words, global tuple positions) is set for the first tuple
 * of the maxLeaf word of every cluster seeded on this shard; the job's keep mask is the sum /
 * OR over the shards. */
int humid_gpu_shard_keep_bits(humid_gpu *h, uint32_t *d_bits, uint64_t n_words32);
/* After finish: d_ids[i] = cluster id of the i-th tuple handed to import_tuples (the
 * host routes them back to the ranks the tuples came from; filtered tuples never travelled
 * and keep id 0). */
int humid_gpu_shard_tuple_ids(humid_gpu *h, uint32_t *d_ids);

/* ---- Groups: the same phases run by the library itself, exchanges over peer memory -------------------------
 *
 * Every rank of a job — one GPU each; host threads of one process or processes of one node — creates its handle,
 * joins the group, adds its share of the read tuples and then calls humid_gpu_group_finalize(): a COLLECTIVE call that
 * runs everything between readData and the outputs (the phases above) with the exchanges done by the kernels
 * themselves: a rank's partition pass stores its tuples straight into the owners' buffers, words and entries of the
 * join and the open components travel as device-to-device copies, keep bits and cluster ids are scattered into the
 * buffer of the rank that fed the tuple — all through peer mappings (cudaDeviceEnablePeerAccess inside a process,
 * CUDA IPC handles between processes) over NVLink, with counts, splitters and barriers in a small shared host
 * segment.  No NCCL on the data path, no Python between the phases.
 *
 * Tuple numbering: with HUMID_GROUP_SLICES rank r's tuples are the r-th contiguous slice of the job's input order
 * (add_batch numbers them from 0 on every rank; the library adds the offsets); with HUMID_GROUP_BATCHES
 * first_tuple_index of humid_gpu_add_batch is the GLOBAL position of the batch (any rank may take any batch, as
 * long as a rank's own batches ascend) and the per-tuple results of the whole job are collected on rank 0. */
enum { HUMID_GROUP_SLICES = 0, HUMID_GROUP_BATCHES = 1 };

typedef struct humid_gpu_group humid_gpu_group; /* ranks that are threads of one process share this object */
humid_gpu_group *humid_gpu_group_create(uint32_t world);
void humid_gpu_group_destroy(humid_gpu_group *g);

/* g != NULL: in-process group (shm_name ignored).  g == NULL: ranks are processes; shm_name names a POSIX shared
 * memory segment (e.g. "/humid_<job id>") that rank 0 creates and the others attach to.  Collective. */
int humid_gpu_group_join(humid_gpu *h, humid_gpu_group *g, const char *shm_name, uint32_t rank, uint32_t world,
                         int tuple_numbering);
/* Collective; replaces humid_gpu_finalize.  Afterwards humid_gpu_counts reports the whole job; humid_gpu_keep_mask
 * and humid_gpu_cluster_ids answer for this rank's own tuples (HUMID_GROUP_SLICES: positions relative to the rank's
 * first tuple; HUMID_GROUP_BATCHES: global positions, on rank 0); humid_gpu_histogram returns the merged histograms. */
int humid_gpu_group_finalize(humid_gpu *h);
/* Collective: prepares humid_gpu_cluster_ids (the owners of the words send every tuple's id to the rank that fed it). */
int humid_gpu_group_collect_ids(humid_gpu *h);
/* Milliseconds of the phases of the last humid_gpu_group_finalize on this rank (host clock around each phase, device
 * synchronised): names[i] / ms[i], *n in: capacity, out: count. */
int humid_gpu_group_phases(humid_gpu *h, const char **names, float *ms, uint32_t *n);

#ifdef __cplusplus
}
#endif
#endif /* HUMID_GPU_SHARD_H */
