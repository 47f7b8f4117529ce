/* TEST INFRASTRUCTURE ONLY — never built by humid_b200/build.py, never shipped, never loaded by the product.
 *
 * A stand-in for libhumid_gpu.so that answers the calls host/humid_main.cc makes (include/humid_gpu.h) from the CPU
 * oracle (oracle/humid_oracle.c), so that the HOST side of the `humid` executable — FastQ block parsing, gzip / BGZF
 * input, batch assembly across files, the parallel output pass, file naming, the statistics files — can be checked
 * against the reference binary in the `-m "not gpu"` suite, on a machine without a GPU (tests/test_cli_cpu.py).
 * The device path itself is what tests/test_gpu_cli.py checks, through the real library. */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../include/humid_gpu.h"

typedef struct oracle oracle;
void oracle_pack_words(const uint8_t *, const uint16_t *, uint64_t, uint32_t, uint32_t, const uint32_t *, uint64_t *);
void oracle_pack_words_wide(const uint8_t *, const uint16_t *, uint64_t, uint32_t, uint32_t, const uint32_t *, uint64_t *,
                            uint64_t *);
oracle *oracle_run(const uint64_t *, uint64_t, uint32_t, uint32_t, int, int, int);
oracle *oracle_run_wide(const uint64_t *, const uint64_t *, uint64_t, uint32_t, uint32_t, int, int, int);
void oracle_free(oracle *);
uint64_t oracle_total(const oracle *), oracle_usable(const oracle *), oracle_unique(const oracle *),
    oracle_clusters(const oracle *);
const uint8_t *oracle_keep(const oracle *);
const uint32_t *oracle_cid(const oracle *);
uint64_t oracle_histogram(const oracle *, int, uint64_t *, uint64_t *, uint64_t);

struct humid_gpu {
  humid_gpu_config cfg;
  uint64_t n, cap;
  uint64_t *lo, *hi;
  oracle *o;
  char err[128];
};

int humid_gpu_abi_version(void) { return HUMID_GPU_ABI_VERSION; }

int humid_gpu_create(const humid_gpu_config *cfg, humid_gpu **out) {
  humid_gpu *h = (humid_gpu *)calloc(1, sizeof(humid_gpu));
  h->cfg = *cfg;
  *out = h;
  return HUMID_GPU_OK;
}
void humid_gpu_destroy(humid_gpu *h) {
  if (!h) return;
  if (h->o) oracle_free(h->o);
  free(h->lo);
  free(h->hi);
  free(h);
}
const char *humid_gpu_last_error(const humid_gpu *h) { return h ? h->err : "stub"; }
int humid_gpu_host_register(humid_gpu *h, void *p, size_t bytes) { (void)h; (void)p; (void)bytes; return HUMID_GPU_OK; }
int humid_gpu_host_unregister(humid_gpu *h, void *p) { (void)h; (void)p; return HUMID_GPU_OK; }

int humid_gpu_add_batch(humid_gpu *h, const uint8_t *chars, const uint16_t *seg_len, uint64_t first, uint64_t n) {
  if (first != h->n) {
    snprintf(h->err, sizeof h->err, "add_batch: first_tuple_index %llu, expected %llu", (unsigned long long)first,
             (unsigned long long)h->n);
    return HUMID_GPU_ERR_ARG;
  }
  if (h->n + n > h->cap) {
    h->cap = (h->n + n) * 2;
    h->lo = (uint64_t *)realloc(h->lo, h->cap * 8);
    h->hi = (uint64_t *)realloc(h->hi, h->cap * 8);
  }
  const humid_gpu_config *c = &h->cfg;
  if (c->word_length > HUMID_MAX_NARROW_WORD_LENGTH)
    oracle_pack_words_wide(chars, seg_len, n, c->word_length, c->n_segments, c->seg_take, h->lo + h->n, h->hi + h->n);
  else
    oracle_pack_words(chars, seg_len, n, c->word_length, c->n_segments, c->seg_take, h->lo + h->n);
  h->n += n;
  return HUMID_GPU_OK;
}

int humid_gpu_finalize(humid_gpu *h) {
  const humid_gpu_config *c = &h->cfg;
  uint64_t dummy = 0;
  const uint64_t *lo = h->lo ? h->lo : &dummy, *hi = h->hi ? h->hi : &dummy;
  h->o = c->word_length > HUMID_MAX_NARROW_WORD_LENGTH
             ? oracle_run_wide(lo, hi, h->n, c->word_length, c->distance, c->edit, c->maximum, 0)
             : oracle_run(lo, h->n, c->word_length, c->distance, c->edit, c->maximum, 0);
  return h->o ? HUMID_GPU_OK : HUMID_GPU_ERR_MEMORY;
}

int humid_gpu_counts(humid_gpu *h, uint64_t *total, uint64_t *usable, uint64_t *unique, uint64_t *clusters) {
  if (total) *total = oracle_total(h->o);
  if (usable) *usable = oracle_usable(h->o);
  if (unique) *unique = oracle_unique(h->o);
  if (clusters) *clusters = oracle_clusters(h->o);
  return HUMID_GPU_OK;
}

int humid_gpu_histogram(humid_gpu *h, int which, uint64_t *keys, uint64_t *vals, uint64_t *n) {
  if (!keys || !vals) {
    *n = oracle_histogram(h->o, which, NULL, NULL, 0);
    return HUMID_GPU_OK;
  }
  *n = oracle_histogram(h->o, which, keys, vals, *n);
  return HUMID_GPU_OK;
}

int humid_gpu_keep_mask(humid_gpu *h, uint64_t first, uint64_t n, uint64_t *bits) {
  const uint8_t *keep = oracle_keep(h->o);
  memset(bits, 0, (n + 63) / 64 * 8);
  for (uint64_t t = 0; t < n; t++)
    if (keep[first + t]) bits[t >> 6] |= 1ull << (t & 63);
  return HUMID_GPU_OK;
}

int humid_gpu_cluster_ids(humid_gpu *h, uint64_t first, uint64_t n, uint32_t *ids) {
  memcpy(ids, oracle_cid(h->o) + first, n * sizeof(uint32_t));
  return HUMID_GPU_OK;
}

/* Groups (include/humid_gpu_shard.h) need real GPUs: the stand-in refuses them, so `humid -g 2` fails loudly here. */
struct humid_gpu_group;
struct humid_gpu_group *humid_gpu_group_create(uint32_t world) { (void)world; return NULL; }
void humid_gpu_group_destroy(struct humid_gpu_group *g) { (void)g; }
int humid_gpu_group_join(humid_gpu *h, struct humid_gpu_group *g, const char *shm, uint32_t rank, uint32_t world, int numbering) {
  (void)h; (void)g; (void)shm; (void)rank; (void)world; (void)numbering;
  return HUMID_GPU_ERR_UNSUPPORTED;
}
int humid_gpu_group_collect_ids(humid_gpu *h) { (void)h; return HUMID_GPU_ERR_UNSUPPORTED; }
