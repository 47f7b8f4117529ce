/* TEST INFRASTRUCTURE ONLY — CPU restatement of HUMID's deduplication hot path.
 *
 * Nothing in the product (humid_b200/, include/, host/) may call, link or import this
 * file; it is the checker used by tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline leg.  Parity status: PINNED — checked in tests/test_oracle.py against
 *   (1) the known-answer vectors of the reference's own unit tests
 *       (tests/test_fastq.cc, tests/test_cluster.cc, lib/trie/tests/test_trie.cc),
 *   (2) oracle/_ref/trie_probe (the unmodified reference trie + cluster code) on random
 *       word sets for every mode, and
 *   (3) oracle/_ref/humid (the unmodified reference binary) on synthetic FastQ files,
 *       byte-for-byte on all output files; golden copies live in tests/golden/.
 *
 * Representation: a word of n nucleotides is one 128-bit integer (okey), 2 bits per nucleotide
 * (A=0,C=1,G=2,T=3, src/fastq.cc:12), first nucleotide in the most significant used bits,
 * so unsigned key order == trie pre-order walk order (lib/trie/src/trieFunctions.tcc:129-143).
 * n <= 63 here; the entry points take either one uint64_t per word (n <= 31, the external form
 * of the one-key library path) or a (lo, hi) pair of uint64_t (n <= 63).  The pointer trie is restated as an *implicit trie over the sorted array
 * of distinct keys*: the node reached by a prefix of `depth` letters is the index range
 * [lo,hi) of keys carrying that prefix, child c is the sub-range whose next letter is c
 * (binary search).  With that substitution the recursions below are line-for-line
 * restatements of hamming_ / levenshtein_ / walk_, so yield order and multiplicity agree
 * by construction.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define ORACLE_INVALID UINT64_MAX
typedef unsigned __int128 okey;
#define OKEY_INVALID (~(okey)0)
#define ORACLE_MAX_N 63u

typedef struct {
  uint32_t *v;
  uint64_t n, cap;
} vec32;

static void vec_push(vec32 *a, uint32_t x) {
  if (a->n == a->cap) {
    a->cap = a->cap ? a->cap * 2 : 4;
    a->v = (uint32_t *)realloc(a->v, a->cap * sizeof(uint32_t));
  }
  a->v[a->n++] = x;
}

typedef struct oracle {
  uint32_t n, d;
  int edit, maximum;
  uint64_t R, usable, U, C;
  okey *keys;        /* [U] sorted distinct words (right-aligned, the external form) */
  uint64_t *keys_lo, *keys_hi; /* [U] the same, split for the getters               */
  okey *lk;          /* [U] same words left-aligned (internal trie form)             */
  uint8_t *len;      /* [U] word lengths (all n in the pipeline)                     */
  uint64_t *count;   /* [U] Leaf::count (lib/trie/src/node.tcc:14-16)               */
  uint64_t *first;   /* [U] first tuple index carrying the word (visited logic)     */
  vec32 *nbr;        /* [U] NLeaf::neighbours (src/leaf.h:6-9), as ranks            */
  uint32_t *cluster; /* [U] Cluster::id of the leaf (1-based)                       */
  uint64_t *csize;   /* [C] Cluster::size                                           */
  uint32_t *cmax;    /* [C] rank of Cluster::maxLeaf                                */
  uint64_t *cmaxcount;
  uint8_t *keep;     /* [R] tuple written to *_dedup                                */
  uint32_t *cid;     /* [R] annotated cluster id, 0 for filtered tuples             */
} oracle;

/* ------------------------------------------------------------------------------------
 * Word construction (src/fastq.cc)
 * ---------------------------------------------------------------------------------- */

/* ntFromFile, src/fastq.cc:220-230: length/files each, remainder to the LAST file. */
void oracle_nt_from_file(uint64_t files, uint64_t length, uint64_t *out) {
  uint64_t div = length / files;
  for (uint64_t i = 0; i + 1 < files; i++) out[i] = div;
  out[files - 1] = div + length % files;
}

/* preCompute, src/humid.cc:38-59. umi_size_first = extractUMI(first record).size(). */
void oracle_pre_compute(uint64_t umi_size_first, uint64_t word_length, uint64_t files,
                        uint64_t *header_umi_size, uint64_t *nt_to_take) {
  uint64_t from_file = 0;
  if (word_length > umi_size_first) from_file = word_length - umi_size_first;
  oracle_nt_from_file(files, from_file, nt_to_take);
  *header_umi_size = word_length < umi_size_first ? word_length : umi_size_first;
}

static int is_nuc(char c) { return c == 'A' || c == 'C' || c == 'G' || c == 'T'; }

/* validUMI, src/fastq.cc:201-214 */
static int valid_umi(const char *s, uint64_t len) {
  if (len == 0) return 0;
  for (uint64_t i = 0; i < len; i++)
    if (!is_nuc(s[i])) return 0;
  return 1;
}

/* extractUMI_, src/fastq.cc:72-93 (+ extractLastField :192-199).  `header` is the whole
 * name line including the leading '@' (lib/fastp/src/fastqreader.cpp:301).  Returns the
 * offset of the UMI inside `header` and its length (0 when there is none). */
void oracle_extract_umi(const char *header, uint64_t len, uint64_t *off, uint64_t *ulen) {
  uint64_t end = len;
  for (uint64_t i = 0; i < len; i++)
    if (header[i] == ' ') { end = i; break; }
  const char seps[2] = {'_', ':'};
  for (int k = 0; k < 2; k++) {
    int64_t last = -1;
    for (uint64_t i = 0; i < end; i++)
      if (header[i] == seps[k]) last = (int64_t)i;
    if (last >= 0 && valid_umi(header + last + 1, end - (uint64_t)last - 1)) {
      *off = (uint64_t)last + 1;
      *ulen = end - (uint64_t)last - 1;
      return;
    }
  }
  *off = 0;
  *ulen = 0;
}

/* getNucleotides + makeWord, src/fastq.cc:116-161, on the boundary layout of
 * include/humid_gpu.h: per tuple `n` bytes, segment s (header UMI first when
 * header_umi_size > 0, then one per file) occupies seg_take[s] bytes; seg_len (may be
 * NULL = all full) holds min(actual length, take) per (tuple, segment).  Bytes beyond
 * seg_len are padding 'N' (makeStringSize_, :57-66) => the word is filtered, like any byte
 * outside ACGT (:151-158).  Filtered => ORACLE_INVALID. */
static void pack_words128(const uint8_t *chars, const uint16_t *seg_len, uint64_t n_tuples,
                          uint32_t n, uint32_t n_segs, const uint32_t *seg_take, okey *keys_out) {
  for (uint64_t t = 0; t < n_tuples; t++) {
    const uint8_t *c = chars + t * n;
    okey key = 0;
    int filtered = 0;
    uint32_t p = 0;
    for (uint32_t s = 0; s < n_segs; s++) {
      uint32_t have = seg_len ? seg_len[t * n_segs + s] : seg_take[s];
      for (uint32_t q = 0; q < seg_take[s]; q++, p++) {
        char ch = q < have ? (char)c[p] : 'N';
        okey code;
        switch (ch) {
          case 'A': code = 0; break;
          case 'C': code = 1; break;
          case 'G': code = 2; break;
          case 'T': code = 3; break;
          default: code = 2; filtered = 1; break; /* nuc['G'] pushed, filtered=true */
        }
        key = (key << 2) | code;
      }
    }
    keys_out[t] = filtered ? OKEY_INVALID : key;
  }
}
void oracle_pack_words(const uint8_t *chars, const uint16_t *seg_len, uint64_t n_tuples,
                       uint32_t n, uint32_t n_segs, const uint32_t *seg_take,
                       uint64_t *keys_out) { /* n <= 31 */
  okey *k = (okey *)malloc((n_tuples ? n_tuples : 1) * sizeof(okey));
  pack_words128(chars, seg_len, n_tuples, n, n_segs, seg_take, k);
  for (uint64_t t = 0; t < n_tuples; t++) keys_out[t] = (uint64_t)k[t];
  free(k);
}
/* n <= 63: lo = the last 32 letters, hi = the letters before them; filtered => both all-ones */
void oracle_pack_words_wide(const uint8_t *chars, const uint16_t *seg_len, uint64_t n_tuples,
                            uint32_t n, uint32_t n_segs, const uint32_t *seg_take,
                            uint64_t *lo_out, uint64_t *hi_out) {
  okey *k = (okey *)malloc((n_tuples ? n_tuples : 1) * sizeof(okey));
  pack_words128(chars, seg_len, n_tuples, n, n_segs, seg_take, k);
  for (uint64_t t = 0; t < n_tuples; t++) { lo_out[t] = (uint64_t)k[t]; hi_out[t] = (uint64_t)(k[t] >> 64); }
  free(k);
}

/* ------------------------------------------------------------------------------------
 * Implicit trie over the sorted distinct keys
 *
 * Internally words are LEFT-aligned (letter p at bits 125-2p..124-2p, p <= 62) with an
 * explicit length, sorted by (left-aligned key, length), i.e. lexicographically with a
 * word before its extensions == walk_ order (lib/trie/src/trieFunctions.tcc:129-143).
 * HUMID only ever stores words of one length n; variable lengths are supported so that the
 * mixed-length golden vectors of lib/trie/tests/test_trie.cc:129-191 can pin the recursion.
 * ---------------------------------------------------------------------------------- */

static inline uint32_t letter_at(okey lkey, uint32_t p) {
  return (uint32_t)(lkey >> (2 * (ORACLE_MAX_N - 1 - p))) & 3u;
}

/* node->leaf: the (single) word of the range whose length equals the node depth. */
static inline int leaf_of(const oracle *o, uint64_t lo, uint64_t hi, uint32_t depth, uint64_t *leaf) {
  if (lo < hi && o->len[lo] == depth) { *leaf = lo; return 1; }
  return 0;
}

/* first index in [lo,hi) whose letter at `depth` is >= c (all keys of the range are longer
 * than depth and share the first `depth` letters, so they are ordered by that letter). */
static uint64_t lower_letter(const oracle *o, uint64_t lo, uint64_t hi, uint32_t depth,
                             uint32_t c) {
  while (lo < hi) {
    uint64_t mid = lo + (hi - lo) / 2;
    if (letter_at(o->lk[mid], depth) < c) lo = mid + 1; else hi = mid;
  }
  return lo;
}

/* node->child[c] (lib/trie/src/node.tcc:21-28): non-empty sub-range or "null". */
static int child_of(const oracle *o, uint64_t lo, uint64_t hi, uint32_t depth, uint32_t c,
                    uint64_t *clo, uint64_t *chi) {
  uint64_t leaf;
  if (leaf_of(o, lo, hi, depth, &leaf)) lo++; /* the node's own leaf is not below a child */
  if (lo >= hi || depth > ORACLE_MAX_N - 1) return 0;
  *clo = lower_letter(o, lo, hi, depth, c);
  *chi = lower_letter(o, *clo, hi, depth, c + 1);
  return *clo < *chi;
}

typedef void (*yield_fn)(oracle *o, void *ctx, uint64_t leaf);

/* hamming_<4,T,full>, lib/trie/src/trieFunctions.tcc:156-181.  (The reference yields a
 * Result with a null leaf when the path ends on an inner node; with HUMID's fixed-length
 * words that never happens and such results are skipped here.) */
static void hamming_rec(oracle *o, uint64_t lo, uint64_t hi, okey word, uint32_t wlen,
                        uint32_t position, int distance, int full, yield_fn yield, void *ctx) {
  if (distance >= 0) {
    if (position < wlen) {
      uint32_t wl = letter_at(word, position);
      for (uint32_t i = full ? 0 : wl; i < 4; i++) { /* asymmetric: start = word[position] */
        uint64_t clo, chi;
        if (child_of(o, lo, hi, position, i, &clo, &chi))
          hamming_rec(o, clo, chi, word, wlen, position + 1, distance - (int)(i != wl), full,
                      yield, ctx);
      }
    } else {
      uint64_t leaf;
      if (leaf_of(o, lo, hi, position, &leaf)) yield(o, ctx, leaf);
    }
  }
}

/* levenshtein_<4,T,full>, lib/trie/src/trieFunctions.tcc:196-233.  `depth` is the depth of
 * `node` in the trie (implicit in the reference's node pointer). */
static void levenshtein_rec(oracle *o, uint64_t lo, uint64_t hi, uint32_t depth, okey word,
                            uint32_t wlen, uint32_t position, int distance, int full,
                            yield_fn yield, void *ctx) {
  if (distance >= 0) {
    /* Deletion (no bound on position: also past the end of the word, :202-203). */
    levenshtein_rec(o, lo, hi, depth, word, wlen, position + 1, distance - 1, full, yield, ctx);

    uint32_t start = 0;
    if (!full && position < wlen) start = letter_at(word, position);
    for (uint32_t i = start; i < 4; i++) {
      uint64_t clo, chi;
      if (child_of(o, lo, hi, depth, i, &clo, &chi)) {
        /* Substitution. */
        if (position < wlen)
          levenshtein_rec(o, clo, chi, depth + 1, word, wlen, position + 1,
                          distance - (int)(i != letter_at(word, position)), full, yield, ctx);
        /* Insertion. */
        levenshtein_rec(o, clo, chi, depth + 1, word, wlen, position, distance - 1, full, yield,
                        ctx);
      }
    }
    uint64_t leaf;
    if (position >= wlen && leaf_of(o, lo, hi, depth, &leaf)) yield(o, ctx, leaf);
  }
}

/* findHammingNeighbours / findEditNeighbours loop body, src/humid.cc:120-123, :148-151 */
static void on_yield(oracle *o, void *ctx, uint64_t result_leaf) {
  uint32_t walk_leaf = *(uint32_t *)ctx;
  if (walk_leaf != (uint32_t)result_leaf) {
    vec_push(&o->nbr[walk_leaf], (uint32_t)result_leaf);
    vec_push(&o->nbr[result_leaf], walk_leaf);
  }
}

/* ------------------------------------------------------------------------------------
 * Clustering (src/cluster.cc), recursion unrolled onto an explicit stack
 * ---------------------------------------------------------------------------------- */

typedef struct { uint32_t node; uint64_t i; } frame;

/* assignLeaf_ :10-13 */
static void assign_leaf(oracle *o, uint32_t leaf, uint32_t cl) {
  o->cluster[leaf] = cl;
  o->csize[cl - 1] += o->count[leaf];
}
/* updateMaxCount_ :20-25 (strict >) */
static void update_max(oracle *o, uint32_t leaf, uint32_t cl) {
  if (o->count[leaf] > o->cmaxcount[cl - 1]) {
    o->cmax[cl - 1] = leaf;
    o->cmaxcount[cl - 1] = o->count[leaf];
  }
}
/* atLeastDouble_ :31-33 */
static int at_least_double(uint64_t a, uint64_t b) { return a >= 2 * b; }

/* assignMaxCluster :72-80 */
static void assign_max_cluster(oracle *o, uint32_t leaf, uint32_t cl, frame **stk, uint64_t *cap) {
  uint64_t sp = 0;
  assign_leaf(o, leaf, cl);
  update_max(o, leaf, cl);
  (*stk)[sp++] = (frame){leaf, 0};
  while (sp) {
    frame *f = &(*stk)[sp - 1];
    if (f->i < o->nbr[f->node].n) {
      uint32_t nb = o->nbr[f->node].v[f->i++];
      if (!o->cluster[nb]) {
        assign_leaf(o, nb, cl);
        update_max(o, nb, cl);
        if (sp == *cap) { *cap *= 2; *stk = (frame *)realloc(*stk, *cap * sizeof(frame)); }
        (*stk)[sp++] = (frame){nb, 0};
      }
    } else {
      sp--;
    }
  }
}

/* maxNeighbour_ :39-51 */
static uint32_t max_neighbour(const oracle *o, uint32_t leaf) {
  uint64_t i = 0;
  while (i < o->nbr[leaf].n) {
    uint32_t nb = o->nbr[leaf].v[i++];
    if (!o->cluster[nb] && at_least_double(o->count[nb], o->count[leaf])) {
      leaf = nb;
      i = 0;
    }
  }
  return leaf;
}

/* assignDirectionalCluster :82-87 with assignDirectionalCluster_ :58-69 */
static void assign_directional_cluster(oracle *o, uint32_t leaf, uint32_t cl, frame **stk,
                                       uint64_t *cap) {
  uint32_t node = max_neighbour(o, leaf);
  update_max(o, node, cl);
  uint64_t sp = 0;
  assign_leaf(o, node, cl);
  (*stk)[sp++] = (frame){node, 0};
  while (sp) {
    frame *f = &(*stk)[sp - 1];
    if (f->i < o->nbr[f->node].n) {
      uint32_t nb = o->nbr[f->node].v[f->i++];
      if (!o->cluster[nb] && at_least_double(o->count[f->node], o->count[nb])) {
        assign_leaf(o, nb, cl);
        if (sp == *cap) { *cap *= 2; *stk = (frame *)realloc(*stk, *cap * sizeof(frame)); }
        (*stk)[sp++] = (frame){nb, 0};
      }
    } else {
      sp--;
    }
  }
}

/* ------------------------------------------------------------------------------------
 * Pipeline (src/humid.cc:369-409)
 * ---------------------------------------------------------------------------------- */

static int cmp_u64(const void *a, const void *b) {
  uint64_t x = *(const uint64_t *)a, y = *(const uint64_t *)b;
  return x < y ? -1 : x > y;
}

static int cmp_okey(const void *a, const void *b) {
  okey x = *(const okey *)a, y = *(const okey *)b;
  return x < y ? -1 : x > y;
}

static uint64_t rank_of(const oracle *o, okey key) { /* Trie::find, :109-119 */
  uint64_t lo = 0, hi = o->U;
  while (lo < hi) {
    uint64_t mid = lo + (hi - lo) / 2;
    if (o->keys[mid] < key) lo = mid + 1; else hi = mid;
  }
  return lo;
}

void oracle_free(oracle *o) {
  if (!o) return;
  if (o->nbr) for (uint64_t i = 0; i < o->U; i++) free(o->nbr[i].v);
  free(o->nbr); free(o->keys); free(o->keys_lo); free(o->keys_hi); free(o->lk); free(o->len); free(o->count); free(o->first); free(o->cluster);
  free(o->csize); free(o->cmax); free(o->cmaxcount); free(o->keep); free(o->cid);
  free(o);
}

/* keys[t] = packed word of input tuple t, or ORACLE_INVALID when filtered.
 * stop_after: 0 = everything, 1 = stop after the word table (readData),
 *             2 = stop after neighbours. */
static oracle *oracle_run128(const okey *keys, uint64_t R, uint32_t n, uint32_t d, int edit,
                             int maximum, int stop_after) {
  oracle *o = (oracle *)calloc(1, sizeof(oracle));
  o->n = n; o->d = d; o->edit = edit; o->maximum = maximum; o->R = R;

  /* readData, src/humid.cc:92-99: total++, usable++ and trie.add for unfiltered tuples */
  okey *tmp = (okey *)malloc((R ? R : 1) * sizeof(okey));
  for (uint64_t t = 0; t < R; t++)
    if (keys[t] != OKEY_INVALID) tmp[o->usable++] = keys[t];
  qsort(tmp, o->usable, sizeof(okey), cmp_okey);
  o->keys = (okey *)malloc((o->usable ? o->usable : 1) * sizeof(okey));
  o->count = (uint64_t *)calloc(o->usable ? o->usable : 1, sizeof(uint64_t));
  for (uint64_t i = 0; i < o->usable; i++) {
    if (o->U == 0 || o->keys[o->U - 1] != tmp[i]) o->keys[o->U++] = tmp[i];
    o->count[o->U - 1]++;
  }
  free(tmp);
  uint64_t U = o->U;
  o->first = (uint64_t *)malloc((U ? U : 1) * sizeof(uint64_t));
  for (uint64_t i = 0; i < U; i++) o->first[i] = UINT64_MAX;
  for (uint64_t t = 0; t < R; t++)
    if (keys[t] != OKEY_INVALID) {
      uint64_t r = rank_of(o, keys[t]);
      if (o->first[r] == UINT64_MAX) o->first[r] = t;
    }
  o->lk = (okey *)malloc((U ? U : 1) * sizeof(okey));
  o->len = (uint8_t *)malloc(U ? U : 1);
  o->keys_lo = (uint64_t *)malloc((U ? U : 1) * sizeof(uint64_t));
  o->keys_hi = (uint64_t *)malloc((U ? U : 1) * sizeof(uint64_t));
  for (uint64_t i = 0; i < U; i++) {
    o->lk[i] = o->keys[i] << (2 * (ORACLE_MAX_N - n));
    o->len[i] = (uint8_t)n;
    o->keys_lo[i] = (uint64_t)o->keys[i];
    o->keys_hi[i] = (uint64_t)(o->keys[i] >> 64);
  }
  o->nbr = (vec32 *)calloc(U ? U : 1, sizeof(vec32));
  o->cluster = (uint32_t *)calloc(U ? U : 1, sizeof(uint32_t));
  o->keep = (uint8_t *)calloc(R ? R : 1, 1);
  o->cid = (uint32_t *)calloc(R ? R : 1, sizeof(uint32_t));
  if (stop_after == 1) return o;

  /* findHammingNeighbours :113-130 / findEditNeighbours :140-158; trie.walk() == rank order */
  for (uint64_t r = 0; r < U; r++) {
    uint32_t walk_leaf = (uint32_t)r;
    if (edit) levenshtein_rec(o, 0, U, 0, o->lk[r], n, 0, (int)d, 0, on_yield, &walk_leaf);
    else      hamming_rec(o, 0, U, o->lk[r], n, 0, (int)d, 0, on_yield, &walk_leaf);
  }
  if (stop_after == 2) return o;

  /* findClusters :167-193 */
  o->csize = (uint64_t *)calloc(U ? U : 1, sizeof(uint64_t));
  o->cmax = (uint32_t *)calloc(U ? U : 1, sizeof(uint32_t));
  o->cmaxcount = (uint64_t *)calloc(U ? U : 1, sizeof(uint64_t));
  uint64_t cap = 1024;
  frame *stk = (frame *)malloc(cap * sizeof(frame));
  uint32_t id = 1;
  for (uint64_t r = 0; r < U; r++) {
    if (!o->cluster[r]) {
      uint32_t cl = id++;
      if (maximum) assign_max_cluster(o, (uint32_t)r, cl, &stk, &cap);
      else         assign_directional_cluster(o, (uint32_t)r, cl, &stk, &cap);
    }
  }
  free(stk);
  o->C = id - 1;

  /* writeFiltered :220-234 (visited / maxLeaf test) and writeAnnotated :268-285 */
  uint8_t *visited = (uint8_t *)calloc(o->C ? o->C : 1, 1);
  for (uint64_t t = 0; t < R; t++) {
    if (keys[t] == OKEY_INVALID) continue; /* cid stays 0 */
    uint64_t r = rank_of(o, keys[t]);
    uint32_t cl = o->cluster[r];
    o->cid[t] = cl;
    if (!visited[cl - 1] && o->cmax[cl - 1] == r) {
      o->keep[t] = 1;
      visited[cl - 1] = 1;
    }
  }
  free(visited);
  return o;
}

oracle *oracle_run(const uint64_t *keys, uint64_t R, uint32_t n, uint32_t d, int edit,
                   int maximum, int stop_after) { /* n <= 31, one uint64_t per word */
  okey *k = (okey *)malloc((R ? R : 1) * sizeof(okey));
  for (uint64_t t = 0; t < R; t++) k[t] = keys[t] == ORACLE_INVALID ? OKEY_INVALID : (okey)keys[t];
  oracle *o = oracle_run128(k, R, n, d, edit, maximum, stop_after);
  free(k);
  return o;
}
/* n <= 63: word = hi << 64 | lo; hi == UINT64_MAX marks a filtered tuple */
oracle *oracle_run_wide(const uint64_t *lo, const uint64_t *hi, uint64_t R, uint32_t n, uint32_t d,
                        int edit, int maximum, int stop_after) {
  okey *k = (okey *)malloc((R ? R : 1) * sizeof(okey));
  for (uint64_t t = 0; t < R; t++) k[t] = hi[t] == ORACLE_INVALID ? OKEY_INVALID : ((okey)hi[t] << 64) | lo[t];
  oracle *o = oracle_run128(k, R, n, d, edit, maximum, stop_after);
  free(k);
  return o;
}

/* ---- getters (ctypes) ---- */
uint64_t oracle_total(const oracle *o) { return o->R; }
uint64_t oracle_usable(const oracle *o) { return o->usable; }
uint64_t oracle_unique(const oracle *o) { return o->U; }
uint64_t oracle_clusters(const oracle *o) { return o->C; }
const uint64_t *oracle_keys(const oracle *o) { return o->keys_lo; }
const uint64_t *oracle_keys_hi(const oracle *o) { return o->keys_hi; }
const uint64_t *oracle_count(const oracle *o) { return o->count; }
const uint64_t *oracle_first(const oracle *o) { return o->first; }
const uint32_t *oracle_cluster_of(const oracle *o) { return o->cluster; }
const uint64_t *oracle_cluster_size(const oracle *o) { return o->csize; }
const uint32_t *oracle_cluster_max(const oracle *o) { return o->cmax; }
const uint8_t *oracle_keep(const oracle *o) { return o->keep; }
const uint32_t *oracle_cid(const oracle *o) { return o->cid; }
uint64_t oracle_degree(const oracle *o, uint64_t r) { return o->nbr[r].n; }
const uint32_t *oracle_neighbours(const oracle *o, uint64_t r) { return o->nbr[r].v; }
uint64_t oracle_total_degree(const oracle *o) {
  uint64_t s = 0;
  for (uint64_t i = 0; i < o->U; i++) s += o->nbr[i].n;
  return s;
}
/* CSR copy of all neighbour lists: row_ptr[U+1], col[total_degree] */
void oracle_csr(const oracle *o, uint64_t *row_ptr, uint32_t *col) {
  uint64_t s = 0;
  for (uint64_t i = 0; i < o->U; i++) {
    row_ptr[i] = s;
    memcpy(col + s, o->nbr[i].v, o->nbr[i].n * sizeof(uint32_t));
    s += o->nbr[i].n;
  }
  row_ptr[o->U] = s;
}

/* runStatistics :301-315 / clusterStats src/cluster.cc:89-95: std::map<size_t,size_t>
 * histograms, ascending keys.  which: 0 counts, 1 neighbours, 2 cluster sizes.
 * Returns the number of (key,value) pairs; writes up to cap of them. */
uint64_t oracle_histogram(const oracle *o, int which, uint64_t *hk, uint64_t *hv, uint64_t cap) {
  uint64_t m = which == 2 ? o->C : o->U;
  uint64_t *v = (uint64_t *)malloc((m ? m : 1) * sizeof(uint64_t));
  for (uint64_t i = 0; i < m; i++)
    v[i] = which == 0 ? o->count[i] : which == 1 ? o->nbr[i].n : o->csize[i];
  qsort(v, m, sizeof(uint64_t), cmp_u64);
  uint64_t k = 0;
  for (uint64_t i = 0; i < m;) {
    uint64_t j = i;
    while (j < m && v[j] == v[i]) j++;
    if (k < cap) { hk[k] = v[i]; hv[k] = j - i; }
    k++;
    i = j;
  }
  free(v);
  return k;
}

/* ------------------------------------------------------------------------------------
 * Stand-alone trie queries on an explicit (possibly mixed-length) word set — used only to
 * pin the recursions against lib/trie/tests/test_trie.cc.  Words are given as letter
 * codes 0..3, concatenated, with lengths; kind: 0 hamming, 1 asymmetricHamming,
 * 2 levenshtein, 3 asymmetricLevenshtein.  Writes the yielded word indices (indices into
 * the caller's word list) in yield order; returns the number of yields.
 * ---------------------------------------------------------------------------------- */
typedef struct { uint64_t *out; uint64_t n, cap; const uint64_t *orig; } qctx;
static void q_yield(oracle *o, void *ctx, uint64_t leaf) {
  (void)o;
  qctx *q = (qctx *)ctx;
  if (q->n < q->cap) q->out[q->n] = q->orig[leaf];
  q->n++;
}
typedef struct { okey lk; uint8_t len; uint64_t orig; } qword;
static int cmp_qword(const void *a, const void *b) {
  const qword *x = (const qword *)a, *y = (const qword *)b;
  if (x->lk != y->lk) return x->lk < y->lk ? -1 : 1;
  return (int)x->len - (int)y->len;
}
static okey left_pack(const uint8_t *letters, uint32_t len) {
  okey k = 0;
  for (uint32_t p = 0; p < len; p++) k |= (okey)(letters[p] & 3) << (2 * (ORACLE_MAX_N - 1 - p));
  return k;
}
uint64_t oracle_trie_query(const uint8_t *letters, const uint32_t *lens, uint64_t n_words,
                           const uint8_t *query, uint32_t qlen, int distance, int kind,
                           uint64_t *out, uint64_t cap) {
  qword *w = (qword *)malloc((n_words ? n_words : 1) * sizeof(qword));
  uint64_t off = 0;
  for (uint64_t i = 0; i < n_words; i++) {
    w[i].lk = left_pack(letters + off, lens[i]);
    w[i].len = (uint8_t)lens[i];
    w[i].orig = i;
    off += lens[i];
  }
  qsort(w, n_words, sizeof(qword), cmp_qword);
  oracle o;
  memset(&o, 0, sizeof o);
  o.U = n_words;
  o.lk = (okey *)malloc((n_words ? n_words : 1) * sizeof(okey));
  o.len = (uint8_t *)malloc(n_words ? n_words : 1);
  uint64_t *orig = (uint64_t *)malloc((n_words ? n_words : 1) * sizeof(uint64_t));
  for (uint64_t i = 0; i < n_words; i++) { o.lk[i] = w[i].lk; o.len[i] = w[i].len; orig[i] = w[i].orig; }
  qctx q = {out, 0, cap, orig};
  okey qk = left_pack(query, qlen);
  if (kind < 2) hamming_rec(&o, 0, n_words, qk, qlen, 0, distance, kind == 0, q_yield, &q);
  else levenshtein_rec(&o, 0, n_words, 0, qk, qlen, 0, distance, kind == 2, q_yield, &q);
  free(o.lk); free(o.len); free(orig); free(w);
  return q.n;
}

/* Stand-alone clustering on an explicit graph (counts + neighbour lists in CSR form) —
 * used to pin the clustering restatement against tests/test_cluster.cc.  `seeds` lists the
 * leaves on which assign*Cluster is called, in order (each gets a new cluster id 1,2,…
 * even if already assigned, like the tests do by hand).  pre_assigned[i] != 0 marks leaves
 * that already belong to some other cluster (id UINT32_MAX).  Outputs cluster id per leaf,
 * size and maxLeaf per seed. */
void oracle_cluster_graph(uint64_t U, const uint64_t *count, const uint64_t *row_ptr,
                          const uint32_t *col, int maximum, const uint32_t *seeds,
                          uint64_t n_seeds, const uint8_t *pre_assigned, uint32_t *cluster_out,
                          uint64_t *size_out, uint32_t *max_out) {
  oracle o;
  memset(&o, 0, sizeof o);
  o.U = U;
  o.count = (uint64_t *)count;
  o.nbr = (vec32 *)calloc(U ? U : 1, sizeof(vec32));
  for (uint64_t i = 0; i < U; i++) {
    o.nbr[i].v = (uint32_t *)(col + row_ptr[i]);
    o.nbr[i].n = row_ptr[i + 1] - row_ptr[i];
  }
  uint64_t nc = n_seeds + 1;
  o.cluster = cluster_out;
  for (uint64_t i = 0; i < U; i++) o.cluster[i] = pre_assigned && pre_assigned[i] ? UINT32_MAX : 0;
  o.csize = (uint64_t *)calloc(nc, sizeof(uint64_t));
  o.cmax = (uint32_t *)calloc(nc, sizeof(uint32_t));
  o.cmaxcount = (uint64_t *)calloc(nc, sizeof(uint64_t));
  uint64_t cap = 64;
  frame *stk = (frame *)malloc(cap * sizeof(frame));
  for (uint64_t s = 0; s < n_seeds; s++) {
    uint32_t cl = (uint32_t)s + 1;
    o.cmax[s] = UINT32_MAX;
    if (maximum) assign_max_cluster(&o, seeds[s], cl, &stk, &cap);
    else assign_directional_cluster(&o, seeds[s], cl, &stk, &cap);
    size_out[s] = o.csize[s];
    max_out[s] = o.cmax[s];
  }
  free(stk); free(o.nbr); free(o.csize); free(o.cmax); free(o.cmaxcount);
}

/* maxNeighbour_ alone (tests/test_cluster.cc:23-71) */
uint32_t oracle_max_neighbour(uint64_t U, const uint64_t *count, const uint64_t *row_ptr,
                              const uint32_t *col, const uint8_t *assigned, uint32_t leaf) {
  oracle o;
  memset(&o, 0, sizeof o);
  o.U = U;
  o.count = (uint64_t *)count;
  o.nbr = (vec32 *)calloc(U ? U : 1, sizeof(vec32));
  o.cluster = (uint32_t *)calloc(U ? U : 1, sizeof(uint32_t));
  for (uint64_t i = 0; i < U; i++) {
    o.nbr[i].v = (uint32_t *)(col + row_ptr[i]);
    o.nbr[i].n = row_ptr[i + 1] - row_ptr[i];
    o.cluster[i] = assigned && assigned[i] ? 1 : 0;
  }
  uint32_t r = max_neighbour(&o, leaf);
  free(o.nbr); free(o.cluster);
  return r;
}
int oracle_at_least_double(uint64_t a, uint64_t b) { return at_least_double(a, b); }

/* getNucleotides, src/fastq.cc:116-144, on the boundary layout: writes the n characters
 * (after cut / 'N' padding) of one tuple — the string the reference's tests compare. */
void oracle_expand_word(const uint8_t *chars, const uint16_t *seg_len, uint32_t n_segs,
                        const uint32_t *seg_take, uint8_t *out) {
  uint32_t p = 0;
  for (uint32_t s = 0; s < n_segs; s++)
    for (uint32_t q = 0; q < seg_take[s]; q++, p++)
      out[p] = q < seg_len[s] ? chars[p] : (uint8_t)'N';
}

/* findClusters (src/humid.cc:167-193) on an explicit neighbour graph (counts + neighbour
 * lists in CSR form, in the reference's list order): walk order seeds, ids from 1.  Used by
 * the CPU shard engine of tests/ to finish a job whose graph was assembled from shards.
 * size_out / max_out need room for U entries; returns the number of clusters. */
uint64_t oracle_cluster_all(uint64_t U, const uint64_t *count, const uint64_t *row_ptr, const uint32_t *col,
                            int maximum, uint32_t *cluster_out, uint64_t *size_out, uint32_t *max_out) {
  oracle o;
  memset(&o, 0, sizeof o);
  o.U = U;
  o.count = (uint64_t *)count;
  o.nbr = (vec32 *)calloc(U ? U : 1, sizeof(vec32));
  for (uint64_t i = 0; i < U; i++) {
    o.nbr[i].v = (uint32_t *)(col + row_ptr[i]);
    o.nbr[i].n = row_ptr[i + 1] - row_ptr[i];
  }
  o.cluster = cluster_out;
  memset(cluster_out, 0, U * sizeof(uint32_t));
  o.csize = size_out;
  o.cmax = max_out;
  memset(size_out, 0, U * sizeof(uint64_t));
  memset(max_out, 0, U * sizeof(uint32_t));
  o.cmaxcount = (uint64_t *)calloc(U ? U : 1, sizeof(uint64_t));
  uint64_t cap = 1024;
  frame *stk = (frame *)malloc(cap * sizeof(frame));
  uint32_t id = 1;
  for (uint64_t r = 0; r < U; r++) {
    if (!o.cluster[r]) {
      uint32_t cl = id++;
      if (maximum) assign_max_cluster(&o, (uint32_t)r, cl, &stk, &cap);
      else assign_directional_cluster(&o, (uint32_t)r, cl, &stk, &cap);
    }
  }
  free(stk); free(o.nbr); free(o.cmaxcount);
  return (uint64_t)id - 1;
}
