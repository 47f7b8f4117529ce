/* humid_gpu.h — C ABI of the B200-native HUMID deduplication hot path.
 *
 * The reference (jfjlaros/HUMID, a single-threaded C++ program) has no plugin / FFI seam:
 * `humid()` (src/humid.cc:369-409) calls the header-only `Trie<4,NLeaf>` and the free
 * functions of src/cluster.h directly.  The seam this library fills is exactly the set of
 * calls src/humid.cc makes on that trie; each entry point below names the reference call
 * sites it replaces.  Everything from word construction onwards runs on the GPU; there is
 * NO CPU fallback: every call fails with HUMID_GPU_ERR_CUDA when no usable device exists.
 *
 * Conventions: plain C, opaque handle, caller-owned host buffers unless a name ends in
 * `_device`; return 0 on success, a negative HUMID_GPU_ERR_* otherwise, with a message
 * from humid_gpu_last_error().  One host thread per handle.  The library owns its CUDA
 * stream and all device memory.  A word of n nucleotides is packed 2 bits per nucleotide
 * (A=0 C=1 G=2 T=3, src/fastq.cc:12), first nucleotide most significant, so unsigned key
 * order == the trie's pre-order walk order (lib/trie/src/trieFunctions.tcc:129-143);
 * "rank" below always means the position of a distinct word in that order.
 */
#ifndef HUMID_GPU_H
#define HUMID_GPU_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HUMID_GPU_ABI_VERSION 1
#define HUMID_MAX_SEGMENTS 16 /* header UMI + up to 15 input files */
#define HUMID_MAX_NARROW_WORD_LENGTH 31 /* one 64-bit key per word (2 bits/nt + a spare sentinel bit) */
#define HUMID_MAX_WORD_LENGTH 63 /* longer words take two 64-bit keys: Hamming (and edit distance 1) only,
                                    one GPU only; everything else about the calls is the same */

enum {
  HUMID_GPU_OK = 0,
  HUMID_GPU_ERR_ARG = -1,    /* bad argument / call order */
  HUMID_GPU_ERR_CUDA = -2,   /* CUDA failure (incl. no device): never falls back to the CPU */
  HUMID_GPU_ERR_MEMORY = -3, /* device or host allocation failed */
  HUMID_GPU_ERR_UNSUPPORTED = -4
};

enum { HUMID_HIST_COUNTS = 0, HUMID_HIST_NEIGHBOURS = 1, HUMID_HIST_CLUSTERS = 2 };

typedef struct humid_gpu humid_gpu;

/* Options that reach the device: `-n`, `-m`, `-e`, `-x` (src/humid.cc:419-428) plus the
 * word layout computed once by preCompute (src/humid.cc:38-59, ntFromFile
 * src/fastq.cc:220-230).  Segments are, in word order: the header UMI (only when
 * header_umi_size > 0) followed by one segment per input file; seg_take[] are their sizes
 * and must sum to word_length. */
typedef struct {
  uint32_t word_length;                  /* -n */
  uint32_t distance;                     /* -m */
  uint32_t n_segments;
  uint32_t seg_take[HUMID_MAX_SEGMENTS];
  uint8_t edit;                          /* -e: Levenshtein instead of Hamming */
  uint8_t maximum;                       /* -x: maximum instead of directional clusters */
  uint8_t defer_table;                   /* 1: never count words during add_batch (shard hosts) */
  uint8_t reserved;
  int32_t device;                        /* CUDA device ordinal, -1 = current */
  uint64_t expected_tuples;              /* capacity hint, 0 = grow on demand.  With a hint the word
                                            table is sized up front and every batch is counted into it
                                            as it arrives (Trie::add per read, src/humid.cc:69-77), so
                                            that work overlaps the next batch's upload; a job that
                                            outgrows the hint is recounted at finalize. */
} humid_gpu_config;

int humid_gpu_abi_version(void);

int humid_gpu_create(const humid_gpu_config *cfg, humid_gpu **out);
void humid_gpu_destroy(humid_gpu *h);
/* Forget all words and results, keep allocations (so a handle can run job after job). */
int humid_gpu_reset(humid_gpu *h);
/* Message of the last failing call on `h` (or of the last failing create when h == NULL). */
const char *humid_gpu_last_error(const humid_gpu *h);

/* The cudaStream_t all of h's device work is issued on (so a caller can record its own
 * events around calls, or order device-resident inputs before humid_gpu_add_batch_device). */
void *humid_gpu_stream(humid_gpu *h);

/* Page-locked host memory for the buffers handed to humid_gpu_add_batch: uploads from it run at
 * the full PCIe rate and overlap the device work of the previous batch.  Plain malloc'ed memory
 * works too, only slower.  Nothing in the reference corresponds to this. */
int humid_gpu_host_alloc(humid_gpu *h, size_t bytes, void **out);
int humid_gpu_host_free(humid_gpu *h, void *p);
/* The same for memory the caller already owns (page-aligned): page-locks it in place, so a host can fill its
 * buffers while humid_gpu_create is still bringing the device up and lock them afterwards. */
int humid_gpu_host_register(humid_gpu *h, void *p, size_t bytes);
int humid_gpu_host_unregister(humid_gpu *h, void *p);

/* Pass 1 — replaces `makeWord` + `trie.add` in readData (src/humid.cc:92-99; makeWord
 * src/fastq.cc:146-161, getNucleotides :116-144, makeStringSize_ :57-66, Trie::add
 * lib/trie/src/trieFunctions.tcc:45-59).
 *   chars    n_tuples x word_length raw ASCII bytes, tuple-major; inside a tuple segment s
 *            occupies seg_take[s] bytes at offset sum(seg_take[0..s)).  The host copies
 *            min(actual length, seg_take[s]) characters of the header UMI / read s there
 *            and does NOT validate them.
 *   seg_len  n_tuples x n_segments actual lengths (already clamped to seg_take), or NULL
 *            when every segment is full.  The cut / 'N'-pad rule and the ACGT-only filter
 *            run on the device: a short segment or any byte outside "ACGT" filters the
 *            tuple (it counts in `total` only and gets cluster id 0).
 * Tuples are numbered in call order; first_tuple_index must equal the number of tuples
 * added so far (the reference processes tuples strictly in input order). */
int humid_gpu_add_batch(humid_gpu *h, const uint8_t *chars, const uint16_t *seg_len,
                        uint64_t first_tuple_index, uint64_t n_tuples);
/* Same, with `chars` / `seg_len` already resident in device memory of h's device. */
int humid_gpu_add_batch_device(humid_gpu *h, const uint8_t *d_chars, const uint16_t *d_seg_len,
                               uint64_t first_tuple_index, uint64_t n_tuples);

/* humid_gpu_add_batch with the characters in the 4-bit wire format — half the bytes on the host -> device link,
 * which is what bounds the path end to end.  Same reference call sites as humid_gpu_add_batch (makeWord + trie.add,
 * src/humid.cc:92-99); the cut / pad rule, the ACGT-only filter and the 2-bit packing still run on the device.
 *   nibbles  n_tuples x ((word_length + 1) / 2) bytes, tuple-major; character q of a tuple (same order and segment
 *            layout as `chars` above) is the HIGH nibble of byte q / 2 when q is even, the low nibble when odd;
 *            code 0..3 = A C G T (src/fastq.cc:12), any other code (HUMID_NIBBLE_OTHER) = any other byte; the unused
 *            low nibble of the last byte of an odd-length word is ignored.
 *   seg_len  as for humid_gpu_add_batch. */
#define HUMID_NIBBLE_OTHER 4
int humid_gpu_add_batch_packed4(humid_gpu *h, const uint8_t *nibbles, const uint16_t *seg_len,
                                uint64_t first_tuple_index, uint64_t n_tuples);

/* Replaces findHammingNeighbours / findEditNeighbours (src/humid.cc:113-158; hamming_ /
 * levenshtein_ lib/trie/src/trieFunctions.tcc:156-233), findClusters (:167-193;
 * src/cluster.cc), runStatistics (:301-315) and clusterStats (src/cluster.cc:89-95). */
int humid_gpu_finalize(humid_gpu *h);

/* stats.dat scalars (writeStatistics, src/humid.cc:351-356). */
int humid_gpu_counts(humid_gpu *h, uint64_t *total, uint64_t *usable, uint64_t *unique,
                     uint64_t *clusters);
/* counts.dat / neigh.dat / clusters.dat: (key, value) pairs in ascending key order, like
 * the std::map the reference prints (src/humid.cc:328-349).  In: *n = capacity of keys /
 * vals; out: *n = number of pairs (call with keys == NULL to query it). */
int humid_gpu_histogram(humid_gpu *h, int which, uint64_t *keys, uint64_t *vals, uint64_t *n);

/* Pass 2 — replaces `trie.find` + the visited/maxLeaf test of writeFiltered
 * (src/humid.cc:220-234): bit (t - first_tuple) of `bits` (LSB-first inside each
 * uint64_t) is set iff tuple t is written to the *_dedup files.  `bits` holds
 * ceil(n/64) words.  (finalize has already brought the bitmap to the host: no device work here.) */
int humid_gpu_keep_mask(humid_gpu *h, uint64_t first_tuple, uint64_t n, uint64_t *bits);
/* … and of writeAnnotated (:268-285): the cluster id appended to each header, 0 for
 * filtered tuples. */
int humid_gpu_cluster_ids(humid_gpu *h, uint64_t first_tuple, uint64_t n, uint32_t *ids);

/* Introspection for the parity tests (the reference binary never prints these).
 * Words in rank order; any output pointer may be NULL.  `cap` = capacity of each array. */
int humid_gpu_debug_words(humid_gpu *h, uint64_t cap, uint64_t *keys, uint32_t *counts,
                          uint32_t *first_tuple, uint32_t *cluster_ids);
/* Words of more than HUMID_MAX_NARROW_WORD_LENGTH letters: humid_gpu_debug_words returns the low 64
 * bits of each word (its last 32 letters), this call the bits above them. */
int humid_gpu_debug_words_hi(humid_gpu *h, uint64_t cap, uint64_t *keys_hi);
/* Neighbour lists (NLeaf::neighbours, src/leaf.h:6-9) as CSR over ranks, in the
 * reference's list order and multiplicity.  row_ptr has unique+1 entries; call with
 * col == NULL to obtain *n_entries first. */
int humid_gpu_debug_neighbours(humid_gpu *h, uint64_t *row_ptr, uint32_t *col, uint64_t *n_entries);
/* Per cluster id (1-based; arrays indexed id-1): Cluster::size and rank of Cluster::maxLeaf. */
int humid_gpu_debug_clusters(humid_gpu *h, uint64_t cap, uint64_t *sizes, uint32_t *max_leaf_rank);
/* Test hooks for the device primitives the stages are built from (humid_b200/csrc/primitives.cu);
 * all pointers are device memory of n elements unless stated, the call synchronises.
 * sort_pairs: stable radix sort of (key, value) by key bits [begin_bit, end_bit); d_vals_in NULL =
 * the value of element i is i.
 * partition: one stable partition pass.  kind 0: by the top 8 bits of the home slot in a table of
 * arg + 1 slots; kind 1: by the part (of n_parts) joining the letters under mask arg; kind 2: 64-bit
 * entries by the part owning row = entry >> 32 (host array bases[n_parts + 1]); values as above + index_base;
 * counts (host, n_parts entries; NULL for kind 0) = elements per part. */
int humid_gpu_debug_sort_pairs(humid_gpu *h, const uint64_t *d_keys_in, const uint32_t *d_vals_in, uint64_t n,
                               int begin_bit, int end_bit, uint64_t *d_keys_out, uint32_t *d_vals_out);
int humid_gpu_debug_partition(humid_gpu *h, int kind, const uint64_t *d_keys_in, uint64_t n, uint64_t arg,
                              uint32_t n_parts, const uint64_t *bases, uint32_t index_base, uint64_t *d_keys_out,
                              uint32_t *d_vals_out, uint64_t *counts);

/* The slot partition of humid_gpu_debug_partition (kind 0) for callers that cannot allocate: code under CUDA graph
 * capture, tight serving loops, harnesses that forbid cudaMalloc.  NO COUNTERPART IN THE
 * REFERENCE.
 *   d_keys_in   n keys (device memory of h's device), only read
 *   d_keys_out  n keys, d_vals_out n values: the (key, index_base + i) pairs of the input stably grouped by the top 8
 *               bits of the key's home slot in a table of slot_mask + 1 slots — element for element what
 *               humid_gpu_debug_partition(h, 0, d_keys_in, n, slot_mask, 0, NULL, index_base, ...) writes.
 * The call allocates nothing, neither device nor page-locked memory, on any path: the scratch of the pass (tile status
 * words, ticket, digit histogram) is a static device array inside the library, sized for at most
 * HUMID_NOALLOC_MAX_ELEMENTS elements (about 1.4 MB, held by every GPU that loads the library until the process
 * exits).  n > HUMID_NOALLOC_MAX_ELEMENTS fails with HUMID_GPU_ERR_UNSUPPORTED and a message.  The work is ENQUEUED on
 * humid_gpu_stream(h) and the call returns without synchronising.  All calls on one device, whatever the handle,
 * share the one scratch array; it keeps the last call's contents and the part a call uses is cleared first on the
 * call's stream, so calls that follow each other on a stream are independent — but two calls that overlap on the GPU
 * (different handles / streams) race and give undefined results: order them yourself. */
#define HUMID_NOALLOC_MAX_ELEMENTS (1u << 22)
int humid_gpu_partition_by_slot_noalloc(humid_gpu *h, const uint64_t *d_keys_in, uint64_t n, uint64_t slot_mask,
                                        uint32_t index_base, uint64_t *d_keys_out, uint32_t *d_vals_out);

/* Per-stage device timings of the last finalize (CUDA events on the library's stream). */
typedef struct {
  char name[32];
  float ms;               /* device time of the stage                         */
  uint32_t launches;      /* kernels launched by the stage                    */
  uint64_t units;         /* work items (tuples, words, entries …)            */
  uint64_t algo_bytes;    /* algorithmic bytes moved (DESIGN.md per-unit formula) */
} humid_gpu_stage_stat;
int humid_gpu_stage_stats(humid_gpu *h, humid_gpu_stage_stat *out, uint32_t *n /* in: cap, out: count */);

#ifdef __cplusplus
}
#endif
#endif /* HUMID_GPU_H */
