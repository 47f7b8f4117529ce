// Groups (include/humid_gpu_shard.h, second half): the multi-GPU phases run by the library itself.
//
// The reference is one process with one global trie (src/humid.cc:369-409); nothing in it corresponds to this file.
// One rank = one GPU = one handle; ranks are host threads of one process or processes of one node.  What the ranks
// share is (1) a small host segment — heap memory inside a process, POSIX shared memory between processes — with a
// barrier, a few counters per rank and the IPC handles of (2) three device buffers per rank that the OTHER ranks'
// kernels and copies write into over NVLink: the tuple arena (tuples of the rank's key range), the transfer arena
// (words / entries of the distributed join, open components, ids) and the result buffer (keep bits, cluster ids of
// the rank's own tuples).  Every exchange is: counts into the segment, barrier, stores into the peers' buffers,
// stream synchronise, barrier.  There is no NCCL and no Python between the phases.
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <map>
#include <new>
#include <thread>
#include <vector>

#include "../../include/humid_gpu_shard.h"
#include "abi_internal.cuh"

namespace humid {

constexpr uint32_t G_MAX = HUMID_MAX_PARTS;
constexpr uint32_t G_SAMPLES = 1024;       // splitter candidates per rank
constexpr uint32_t G_HIST_CAP = 8192;      // (key, value) pairs per histogram and rank in the segment
constexpr uint32_t G_MAGIC = 0x48554d47u;  // "HUMG"
constexpr int G_BUFS = 3;                  // tuple arena, transfer arena, results
enum { BUF_TUPLES = 0, BUF_XFER = 1, BUF_RESULT = 2 };

struct GroupShared {
  std::atomic<uint32_t> magic;
  std::atomic<uint32_t> arrived, generation, failed;
  uint32_t world;
  struct alignas(64) Rank {
    uint64_t slot[4 * G_MAX + 32];
    uint64_t samples[G_SAMPLES];
    uint32_t n_samples;
    int32_t device;
    uint64_t ptr[G_BUFS], bytes[G_BUFS];
    cudaIpcMemHandle_t ipc[G_BUFS];
    uint32_t hist_n[3];
    uint64_t hist_k[3][G_HIST_CAP], hist_v[3][G_HIST_CAP];
  } r[G_MAX];
};

struct GroupLink {
  GroupShared *sh = nullptr;
  bool inproc = true;
  uint32_t rank = 0, world = 1;
  int numbering = HUMID_GROUP_SLICES;
  std::string shm_name;
  DevBuf<uint8_t> buf[G_BUFS];
  void *peer[G_BUFS][G_MAX] = {};
  DevBuf<uint64_t> exp_keys;  // join words on their way out
  DevBuf<uint32_t> exp_ranks;
  uint64_t tuple_off[G_MAX + 1] = {0};  // HUMID_GROUP_SLICES: first global tuple of every rank
  uint64_t n_total = 0;
  // results: global tuple t belongs to range q with res_start[q] <= t < res_start[q+1], held by rank res_rank[q]
  uint32_t n_res = 0;
  uint64_t res_start[G_MAX + 1] = {0};
  uint32_t res_rank[G_MAX] = {0};
  uint64_t my_res_n = 0;  // tuples of the range this rank holds (keep bits / ids are relative to its start)
  size_t my_bits_bytes = 0;
  bool ids_collected = false;
  struct Batch { uint64_t local, global, n; };  // HUMID_GROUP_BATCHES
  std::vector<Batch> batches;
  std::vector<std::pair<const char *, float>> phases;
  double wait_ms = 0.0;  // time spent inside barriers since the last phase mark (waiting for the slowest rank)
};

}  // namespace humid

struct humid_gpu_group {
  humid::GroupShared sh;
};

namespace humid {

static void group_barrier(GroupLink &g) {
  GroupShared &sh = *g.sh;
  if (g.world == 1) return;
  struct Waited {
    GroupLink &g;
    std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
    ~Waited() { g.wait_ms += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count(); }
  } waited{g};
  const uint32_t gen = sh.generation.load(std::memory_order_acquire);
  if (sh.arrived.fetch_add(1, std::memory_order_acq_rel) + 1 == g.world) {
    sh.arrived.store(0, std::memory_order_relaxed);
    sh.generation.store(gen + 1, std::memory_order_release);
    return;
  }
  const auto t0 = std::chrono::steady_clock::now();
  uint32_t spins = 0;
  while (sh.generation.load(std::memory_order_acquire) == gen) {
    if (sh.failed.load(std::memory_order_relaxed)) throw CudaError{"another rank of the group failed"};
    if (++spins > 2000) {
      std::this_thread::yield();
      if ((spins & 0xFFFF) == 0 &&
          std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() > 120.0)
        throw CudaError{"group barrier: a rank did not arrive within 120 s"};
    }
  }
  if (sh.failed.load(std::memory_order_relaxed)) throw CudaError{"another rank of the group failed"};
}

static inline size_t align_up(size_t x, size_t a = 256) { return (x + a - 1) / a * a; }

// Every rank calls this with the SAME `need` table (bytes wanted in buffer b of every rank, computed from exchanged
// counts).  Nothing happens when all buffers are large enough; otherwise importers drop their mappings of the buffers
// that will move, owners reallocate and publish, importers map again (barriers only in the first steps of a job).
static void group_ensure(State &s, GroupLink &g, int b, const uint64_t need[]) {
  GroupShared &sh = *g.sh;
  bool any = false, grows[G_MAX];
  for (uint32_t r = 0; r < g.world; r++) {
    grows[r] = need[r] > sh.r[r].bytes[b];
    any |= grows[r];
  }
  if (!any) return;
  HUMID_CUDA(cudaStreamSynchronize(s.stream));
  for (uint32_t r = 0; r < g.world; r++)
    if (grows[r] && r != g.rank && g.peer[b][r]) {
      if (!g.inproc) cudaIpcCloseMemHandle(g.peer[b][r]);
      g.peer[b][r] = nullptr;
    }
  group_barrier(g);  // nobody maps or uses the buffers that are about to move
  if (grows[g.rank]) {
    g.buf[b].release();
    g.buf[b].ensure(align_up((size_t)(need[g.rank] + need[g.rank] / 4 + 4096)));
    auto &me = sh.r[g.rank];
    me.ptr[b] = (uint64_t)(uintptr_t)g.buf[b].p;
    if (!g.inproc) HUMID_CUDA(cudaIpcGetMemHandle(&me.ipc[b], g.buf[b].p));
    g.peer[b][g.rank] = g.buf[b].p;
  }
  group_barrier(g);  // every new address / handle is published; `bytes` still holds the old sizes for everyone's test
  if (grows[g.rank]) sh.r[g.rank].bytes[b] = g.buf[b].cap;
  for (uint32_t r = 0; r < g.world; r++) {
    if (r == g.rank || !grows[r]) continue;
    if (g.inproc) {
      g.peer[b][r] = (void *)(uintptr_t)sh.r[r].ptr[b];
    } else {
      void *p = nullptr;
      HUMID_CUDA(cudaIpcOpenMemHandle(&p, sh.r[r].ipc[b], cudaIpcMemLazyEnablePeerAccess));
      g.peer[b][r] = p;
    }
  }
  group_barrier(g);  // sizes are published before anybody computes `grows` again
}

void group_release_handle(State &s) {
  GroupLink *g = s.group;
  if (!g) return;
  for (int b = 0; b < G_BUFS; b++)
    for (uint32_t r = 0; r < g->world; r++)
      if (r != g->rank && g->peer[b][r] && !g->inproc) cudaIpcCloseMemHandle(g->peer[b][r]);
  if (!g->inproc && g->sh) {
    munmap(g->sh, sizeof(GroupShared));
    if (g->rank == 0) shm_unlink(g->shm_name.c_str());
  }
  delete g;
  s.group = nullptr;
}

// ---- kernels that store into other ranks' buffers ---------------------------------------------------------
struct PushTable {
  uint32_t n_parts;
  uint64_t splitters[G_MAX];  // n_parts - 1 used: part p owns [splitters[p-1], splitters[p])
  uint64_t *keys[G_MAX];      // tuple arena of part p: keys ...
  uint32_t *index[G_MAX];     // ... and global tuple positions
};

// The exchange of the read tuples fused into the partition pass: a tuple's {key, global position} is stored
// straight into the arena of the rank that owns its key range (peer memory over NVLink), at a cursor that starts
// where this rank's share of that arena begins.  A CTA first groups its PUSH_TILE tuples by destination in shared
// memory, so that what crosses NVLink are runs of consecutive addresses written by consecutive threads (the first
// version stored every tuple where its thread happened to be: 8-byte stores scattered over eight peers took 2.6 ms
// for 50 M tuples on 8 GPUs).  One global atomic per CTA and part; order inside a share is arbitrary (Trie::add does
// not depend on it: counts add up, the first tuple is a minimum).
constexpr int PUSH_THREADS = 256, PUSH_ITEMS = 8, PUSH_TILE = PUSH_THREADS * PUSH_ITEMS;
__global__ void __launch_bounds__(PUSH_THREADS)
    push_tuples_kernel(const uint64_t *__restrict__ keys, uint64_t R, uint32_t tuple_base, const PushTable T,
                       unsigned long long *__restrict__ cursors) {
  __shared__ uint64_t skey[PUSH_TILE];
  __shared__ uint32_t sidx[PUSH_TILE];
  __shared__ uint32_t cnt[G_MAX + 1], loff[G_MAX + 1];
  __shared__ unsigned long long gbase[G_MAX];
  if (threadIdx.x <= G_MAX) cnt[threadIdx.x] = 0;
  __syncthreads();
  const uint64_t t0 = (uint64_t)blockIdx.x * PUSH_TILE;
  const uint32_t lane = lane_id(), lt = (1u << lane) - 1u;
  uint64_t key[PUSH_ITEMS];
  uint32_t own[PUSH_ITEMS], slot[PUSH_ITEMS];
#pragma unroll
  for (int j = 0; j < PUSH_ITEMS; j++) {
    const uint64_t t = t0 + (uint64_t)j * PUSH_THREADS + threadIdx.x;
    key[j] = t < R ? ld_stream_u64(keys + t) : KEY_INVALID;
    uint32_t p = T.n_parts;  // filtered tuples do not travel
    if (key[j] != KEY_INVALID) {
      p = 0;
      for (uint32_t i = 0; i + 1 < T.n_parts; i++) p += (key[j] >= T.splitters[i]) ? 1u : 0u;
    }
    own[j] = p;
    // lanes with the same destination share one shared-memory atomic
    const uint32_t peers = __match_any_sync(0xffffffffu, p);
    const int leader = __ffs(peers) - 1;
    uint32_t first = 0;
    if ((int)lane == leader) first = atomicAdd(&cnt[p], (uint32_t)__popc(peers));
    slot[j] = __shfl_sync(0xffffffffu, first, leader) + __popc(peers & lt);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t run = 0;
    for (uint32_t p = 0; p < T.n_parts; p++) {
      loff[p] = run;
      run += cnt[p];
    }
    loff[T.n_parts] = run;
  }
  if (threadIdx.x < T.n_parts) {
    const uint32_t c = cnt[threadIdx.x];
    gbase[threadIdx.x] = c ? atomicAdd(cursors + threadIdx.x, (unsigned long long)c) : 0ull;
  }
  __syncthreads();
#pragma unroll
  for (int j = 0; j < PUSH_ITEMS; j++) {
    if (own[j] < T.n_parts) {
      const uint32_t at = loff[own[j]] + slot[j];
      skey[at] = key[j];
      sidx[at] = tuple_base + (uint32_t)(t0 + (uint64_t)j * PUSH_THREADS + threadIdx.x);
    }
  }
  __syncthreads();
  const uint32_t total = loff[T.n_parts];
  for (uint32_t i = threadIdx.x; i < total; i += PUSH_THREADS) {
    uint32_t p = 0;
    while (i >= loff[p + 1]) p++;
    const unsigned long long pos = gbase[p] + (i - loff[p]);
    T.keys[p][pos] = skey[i];
    T.index[p][pos] = sidx[i];
  }
}

// Several device-to-device copies (to other ranks' buffers as well as local) as ONE kernel.  A dozen small
// cudaMemcpyAsync peer copies cost their launch + copy-engine latency each (35 of them made the gathering of the open
// components 1.2 ms long on 8 GPUs); stores from the SMs go over NVLink at full width when consecutive threads write
// consecutive 16-byte units.  The segments are INTERLEAVED over the grid (block b works on segment b mod n) and listed
// starting with the rank after the sender, so that all destinations are in flight at once and the ranks of a group do
// not all write to the same receiver at the same time: with one segment after the other, in the same order on every
// rank, the effective bandwidth fell from 440 GB/s on 2 GPUs to 180 GB/s on 8.  Sizes and addresses: multiples of 4 bytes.
constexpr int COPY_MAX_SEGMENTS = 96;
struct CopySegments {
  uint32_t n;
  void *dst[COPY_MAX_SEGMENTS];
  const void *src[COPY_MAX_SEGMENTS];
  uint64_t bytes[COPY_MAX_SEGMENTS];
};
__global__ void __launch_bounds__(256) multi_copy_kernel(const CopySegments S, uint32_t blocks_per_segment) {
  const uint32_t k = blockIdx.x % S.n, bx = blockIdx.x / S.n;
  const uint64_t bytes = S.bytes[k];
  char *dst = static_cast<char *>(S.dst[k]);
  const char *src = static_cast<const char *>(S.src[k]);
  const uint64_t stride = (uint64_t)blocks_per_segment * blockDim.x, tid = (uint64_t)bx * blockDim.x + threadIdx.x;
  if (((((uintptr_t)dst) | ((uintptr_t)src)) & 15u) == 0) {
    const uint64_t n16 = bytes >> 4;
    for (uint64_t i = tid; i < n16; i += stride) reinterpret_cast<uint4 *>(dst)[i] = reinterpret_cast<const uint4 *>(src)[i];
    for (uint64_t i = (n16 << 2) + tid; i < (bytes >> 2); i += stride)
      reinterpret_cast<uint32_t *>(dst)[i] = reinterpret_cast<const uint32_t *>(src)[i];
  } else {
    for (uint64_t i = tid; i < (bytes >> 2); i += stride)
      reinterpret_cast<uint32_t *>(dst)[i] = reinterpret_cast<const uint32_t *>(src)[i];
  }
}

struct CopyList {
  struct Seg {
    void *dst;
    const void *src;
    uint64_t bytes;
    uint32_t to;  // destination rank
  };
  std::vector<Seg> segs;
  void add(void *dst, const void *src, size_t bytes, uint32_t to) {
    if (bytes) segs.push_back({dst, src, bytes, to});
  }
  // me / world: this rank and the group's size (destinations are taken in the order me + 1, me + 2, ..., me)
  void run(cudaStream_t st, uint32_t me, uint32_t world) {
    if (segs.empty()) return;
    if (segs.size() > (size_t)COPY_MAX_SEGMENTS) throw CudaError{"group: too many copy segments"};
    std::stable_sort(segs.begin(), segs.end(), [&](const Seg &a, const Seg &b) {
      return (a.to + world - me - 1) % world < (b.to + world - me - 1) % world;
    });
    CopySegments S{};
    uint64_t largest = 0;
    for (const Seg &g : segs) {
      S.dst[S.n] = g.dst;
      S.src[S.n] = g.src;
      S.bytes[S.n] = g.bytes;
      S.n++;
      largest = std::max<uint64_t>(largest, g.bytes);
    }
    const uint32_t per = (uint32_t)std::min<uint64_t>(std::max<uint64_t>((largest / 16 + 255) / 256, 1), 148 * 2);
    multi_copy_kernel<<<per * S.n, 256, 0, st>>>(S, per);
    HUMID_CUDA(cudaGetLastError());
  }
};

struct ResultTable {
  uint32_t n;
  uint64_t start[G_MAX + 1];  // global tuple t lies in range q: start[q] <= t < start[q+1]
  uint32_t *bits[G_MAX];      // keep bitmap of the rank holding range q (bit t - start[q])
  uint32_t *ids[G_MAX];       // cluster ids, likewise
};

__device__ __forceinline__ uint32_t result_range(const ResultTable &T, uint32_t t) {
  uint32_t q = 0;
  for (uint32_t i = 1; i < T.n; i++) q += (t >= T.start[i]) ? 1u : 0u;
  return q;
}

// writeFiltered's decision (src/humid.cc:223-231): one bit per cluster seeded here, for the first tuple of the
// cluster's maxLeaf word.  The bits are set in LOCAL bitmaps, one per result range (bit t - start[q] of the bitmap
// at range_off[q]: the layout of the rank that holds the range), which then travel to those ranks as plain copies and
// are OR-ed there — scattering single bits into a peer's memory costs a remote atomic each (0.33 ms for 8 M clusters).
struct LocalBits {
  uint32_t n;
  uint64_t start[G_MAX + 1];
  uint64_t range_off[G_MAX];  // in 32-bit words
};
__global__ void __launch_bounds__(256)
    local_keep_bits_kernel(const uint32_t *__restrict__ cluster_first, uint64_t C, const LocalBits L,
                           uint32_t *__restrict__ bits) {
  const uint64_t c = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const uint32_t t = cluster_first[c];
  uint32_t q = 0;
  for (uint32_t i = 1; i < L.n; i++) q += (t >= L.start[i]) ? 1u : 0u;
  const uint32_t l = t - (uint32_t)L.start[q];
  atomicOr(bits + L.range_off[q] + (l >> 5), 1u << (l & 31u));
}

__global__ void __launch_bounds__(256)
    or_slices_kernel(const uint32_t *__restrict__ slices, uint32_t n_slices, uint64_t words, uint64_t stride_words,
                     uint32_t *__restrict__ out) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= words) return;
  uint32_t v = 0;
  for (uint32_t k = 0; k < n_slices; k++) v |= slices[k * stride_words + i];
  out[i] = v;
}

// writeAnnotated's id (src/humid.cc:276-277) of every tuple this shard counted, stored where its feeder reads it
__global__ void __launch_bounds__(256)
    push_tuple_ids_kernel(const uint64_t *__restrict__ keys, const uint32_t *__restrict__ index, uint64_t n,
                          const Lookup lk, const uint32_t *__restrict__ word_cid, const ResultTable T) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t id = word_cid[table_rank(lk, keys[i])];
  const uint32_t t = index[i];
  const uint32_t q = result_range(T, t);
  T.ids[q][t - (uint32_t)T.start[q]] = id;
}

// Results of the open components a rank has clustered (those with root % mod == me), stored where the rank that owns
// each word reads them: position i of the gathered open list belongs to rank d with off[d] <= i < off[d + 1].
struct OpenResults {
  uint32_t n;
  uint64_t off[G_MAX + 1];
  uint32_t *label[G_MAX], *flag[G_MAX], *mx[G_MAX];
  uint64_t *size[G_MAX];
};
__global__ void __launch_bounds__(256)
    scatter_open_results_kernel(const uint32_t *__restrict__ comp, const uint32_t *__restrict__ label,
                                const uint32_t *__restrict__ flag, const uint64_t *__restrict__ size,
                                const uint32_t *__restrict__ mx, uint64_t n_open, uint32_t mod, uint32_t me,
                                const OpenResults R) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_open || comp[i] % mod != me) return;
  uint32_t d = 0;
  for (uint32_t k = 1; k < R.n; k++) d += (i >= R.off[k]) ? 1u : 0u;
  R.label[d][i] = label[i];
  R.flag[d][i] = flag[i];
  R.mx[d][i] = mx[i];
  R.size[d][i] = size[i];
}

// ---- helpers of the collective finalize ---------------------------------------------------------------------
struct PhaseClock {
  GroupLink &g;
  std::chrono::steady_clock::time_point t;
  explicit PhaseClock(GroupLink &gl) : g(gl), t(std::chrono::steady_clock::now()) {
    g.phases.clear();
    g.wait_ms = 0.0;
  }
  void mark(const char *name) {  // called where the stream has just been synchronised
    const auto now = std::chrono::steady_clock::now();
    g.phases.emplace_back(name, std::chrono::duration<float, std::milli>(now - t).count());
    g.phases.emplace_back("  of which waiting at barriers", (float)g.wait_ms);
    g.wait_ms = 0.0;
    t = now;
  }
};

// counts[dst] of this rank go into the segment; after the barrier every rank knows the whole matrix:
// recv[r] = elements rank r receives, off_at[dst] = where this rank's share starts inside dst's buffer
static void exchange_counts(GroupLink &g, const uint64_t counts[], uint64_t recv[], uint64_t off_at[]) {
  GroupShared &sh = *g.sh;
  for (uint32_t p = 0; p < g.world; p++) sh.r[g.rank].slot[8 + p] = counts[p];
  group_barrier(g);
  for (uint32_t dst = 0; dst < g.world; dst++) {
    recv[dst] = 0;
    off_at[dst] = 0;
    for (uint32_t src = 0; src < g.world; src++) {
      if (src < g.rank) off_at[dst] += sh.r[src].slot[8 + dst];
      recv[dst] += sh.r[src].slot[8 + dst];
    }
  }
}

static size_t bits_bytes(uint64_t tuples) { return align_up(((tuples + 31) / 32 + 2) * 4); }

static ResultTable result_table(const GroupLink &g) {
  ResultTable RT{};
  RT.n = g.n_res;
  for (uint32_t q = 0; q <= g.n_res; q++) RT.start[q] = g.res_start[q];
  for (uint32_t q = 0; q < g.n_res; q++) {
    char *rb = reinterpret_cast<char *>(g.peer[BUF_RESULT][g.res_rank[q]]);
    RT.bits[q] = reinterpret_cast<uint32_t *>(rb);
    RT.ids[q] = reinterpret_cast<uint32_t *>(rb + bits_bytes(g.res_start[q + 1] - g.res_start[q]));
  }
  return RT;
}

// ---- the collective finalize --------------------------------------------------------------------------------
static void group_finalize_body(humid_gpu *h) {
  State &s = h->s;
  GroupLink &g = *s.group;
  GroupShared &sh = *g.sh;
  const uint32_t W = g.world, me = g.rank;
  auto ck = [&](int rc) {
    if (rc != HUMID_GPU_OK) throw CudaError{s.last_error};
  };
  auto &mine = sh.r[me];
  HUMID_CUDA(cudaSetDevice(s.device));
  PhaseClock clock(g);
  s.alias_imports = true;
  g.ids_collected = false;

  // ---- who has what --------------------------------------------------------------------------------------
  uint64_t usable = 0;
  ck(humid_gpu_shard_begin(h, 0, &usable));
  uint32_t ns = 0;
  ck(humid_gpu_shard_sample_keys(h, G_SAMPLES, mine.samples, &ns));
  mine.n_samples = ns;
  mine.slot[0] = s.n_tuples;
  mine.slot[1] = usable;
  uint64_t max_end = 0;
  for (const auto &b : g.batches) max_end = std::max(max_end, b.global + b.n);
  mine.slot[6] = max_end;
  group_barrier(g);
  uint64_t usable_total = 0, batches_end = 0;
  g.tuple_off[0] = 0;
  for (uint32_t r = 0; r < W; r++) {
    g.tuple_off[r + 1] = g.tuple_off[r] + sh.r[r].slot[0];
    usable_total += sh.r[r].slot[1];
    batches_end = std::max<uint64_t>(batches_end, sh.r[r].slot[6]);
  }
  g.n_total = g.tuple_off[W];
  if (g.n_total >= 0xFFFFFFFFull) throw CudaError{"more than 2^32-2 tuples in one job"};
  if (g.numbering == HUMID_GROUP_SLICES) {
    g.n_res = W;
    for (uint32_t r = 0; r <= W; r++) g.res_start[r] = g.tuple_off[r];
    for (uint32_t r = 0; r < W; r++) g.res_rank[r] = r;
    g.my_res_n = s.n_tuples;
    s.tuple_base = g.tuple_off[me];
  } else {
    if (batches_end != g.n_total) throw CudaError{"group batches do not cover the job's tuples without gaps"};
    g.n_res = 1;
    g.res_start[0] = 0;
    g.res_start[1] = g.n_total;
    g.res_rank[0] = 0;
    g.my_res_n = me == 0 ? g.n_total : 0;
  }
  g.my_bits_bytes = bits_bytes(g.my_res_n);
  // result buffers: keep bits, then the ids
  uint64_t need[G_MAX] = {0};
  for (uint32_t r = 0; r < W; r++) need[r] = 256;
  for (uint32_t q = 0; q < g.n_res; q++) {
    const uint64_t m = g.res_start[q + 1] - g.res_start[q];
    need[g.res_rank[q]] = bits_bytes(m) + align_up(m * 4) + (uint64_t)W * bits_bytes(m) + 256;  // bits, ids, W slices
  }
  group_ensure(s, g, BUF_RESULT, need);

  // splitters: quantiles of the pooled samples, cut back to a boundary of block 0 of the join (no group of that
  // block may straddle two shards, include/humid_gpu_shard.h: join_local)
  uint64_t splitters[G_MAX] = {0};
  {
    std::vector<uint64_t> pool;
    for (uint32_t r = 0; r < W; r++) pool.insert(pool.end(), sh.r[r].samples, sh.r[r].samples + sh.r[r].n_samples);
    std::sort(pool.begin(), pool.end());
    const uint32_t n = s.cfg.word_length, d = s.cfg.distance;
    const int tail_bits = 2 * (int)(n - n / (d + 1));
    for (uint32_t p = 1; p < W && !pool.empty(); p++) {
      const uint64_t v = pool[pool.size() * p / W];
      splitters[p - 1] = tail_bits >= 64 ? 0 : (v >> tail_bits) << tail_bits;
    }
  }
  clock.mark("offsets + splitters");

  // ---- tuples to the owners of their key range: stored by the partition pass itself ---------------------------
  uint64_t counts[G_MAX] = {0}, recv[G_MAX], off_at[G_MAX];
  ck(humid_gpu_shard_partition(h, splitters, W, counts));
  exchange_counts(g, counts, recv, off_at);
  for (uint32_t r = 0; r < W; r++) need[r] = align_up(recv[r] * 8) + recv[r] * 4 + 256;
  group_ensure(s, g, BUF_TUPLES, need);
  if (s.n_tuples) {
    PushTable T{};
    T.n_parts = W;
    for (uint32_t p = 0; p + 1 < W; p++) T.splitters[p] = splitters[p];
    for (uint32_t p = 0; p < W; p++) {
      char *arena = reinterpret_cast<char *>(g.peer[BUF_TUPLES][p]);
      T.keys[p] = reinterpret_cast<uint64_t *>(arena);
      T.index[p] = reinterpret_cast<uint32_t *>(arena + align_up(recv[p] * 8));
    }
    unsigned long long *cursors = s.counters.p + 16;
    unsigned long long init[G_MAX] = {0};
    for (uint32_t p = 0; p < W; p++) init[p] = off_at[p];
    HUMID_CUDA(cudaMemcpyAsync(cursors, init, sizeof(init), cudaMemcpyHostToDevice, s.stream));
    Timed t(s, "group_push_tuples", s.n_tuples, s.n_tuples * 20);
    uint32_t launches = 0;
    if (g.numbering == HUMID_GROUP_SLICES) {
      push_tuples_kernel<<<grid_for(s.n_tuples, PUSH_TILE), PUSH_THREADS, 0, s.stream>>>(s.tuple_key.p, s.n_tuples,
                                                                                         (uint32_t)g.tuple_off[me], T, cursors);
      launches = 1;
    } else {
      for (const auto &b : g.batches) {
        push_tuples_kernel<<<grid_for(b.n, PUSH_TILE), PUSH_THREADS, 0, s.stream>>>(s.tuple_key.p + b.local, b.n,
                                                                                   (uint32_t)b.global, T, cursors);
        launches++;
      }
    }
    t.launches = launches;
    HUMID_CUDA(cudaGetLastError());
  }
  HUMID_CUDA(cudaStreamSynchronize(s.stream));
  group_barrier(g);
  clock.mark("tuples: count + push");

  // ---- the owner counts its words (Trie::add) and orders them (Trie::walk) ------------------------------------
  uint64_t n_words = 0;
  {
    char *base = reinterpret_cast<char *>(g.buf[BUF_TUPLES].p);
    ck(humid_gpu_shard_import_tuples(h, reinterpret_cast<uint64_t *>(base),
                                     reinterpret_cast<uint32_t *>(base + align_up(recv[me] * 8)), recv[me], &n_words));
  }
  mine.slot[2] = n_words;
  group_barrier(g);
  uint64_t bases[G_MAX + 1] = {0};
  for (uint32_t r = 0; r < W; r++) bases[r + 1] = bases[r] + sh.r[r].slot[2];
  ck(humid_gpu_shard_set_layout(h, me, W, splitters, bases));
  clock.mark("word table + order");

  // ---- neighbours: block 0 inside the shard, every other block at the part that owns the block value ---------
  ck(humid_gpu_shard_join_local(h, nullptr));
  clock.mark("join block 0");
  const uint32_t d = s.cfg.distance;
  for (uint32_t b = 1; b <= d; b++) {
    g.exp_keys.ensure(n_words + 1);
    g.exp_ranks.ensure(n_words + 1);
    ck(humid_gpu_shard_join_export(h, b, counts, g.exp_keys.p, g.exp_ranks.p));
    exchange_counts(g, counts, recv, off_at);
    for (uint32_t r = 0; r < W; r++) need[r] = align_up(recv[r] * 8) + recv[r] * 4 + 256;
    group_ensure(s, g, BUF_XFER, need);
    uint64_t seg = 0;
    CopyList words;
    for (uint32_t p = 0; p < W; p++) {
      char *dst = reinterpret_cast<char *>(g.peer[BUF_XFER][p]);
      words.add(dst + off_at[p] * 8, g.exp_keys.p + seg, counts[p] * 8, p);
      words.add(dst + align_up(recv[p] * 8) + off_at[p] * 4, g.exp_ranks.p + seg, counts[p] * 4, p);
      seg += counts[p];
    }
    words.run(s.stream, me, W);
    HUMID_CUDA(cudaStreamSynchronize(s.stream));
    group_barrier(g);
    clock.mark("join: words to block owners");
    uint64_t ecounts[G_MAX] = {0}, erecv[G_MAX], eoff[G_MAX];
    {
      char *base = reinterpret_cast<char *>(g.buf[BUF_XFER].p);
      ck(humid_gpu_shard_join_match(h, b, reinterpret_cast<uint64_t *>(base),
                                    reinterpret_cast<uint32_t *>(base + align_up(recv[me] * 8)), recv[me], ecounts));
    }
    exchange_counts(g, ecounts, erecv, eoff);  // behind this barrier nobody reads the words any more
    for (uint32_t r = 0; r < W; r++) need[r] = erecv[r] * 8 + 256;
    group_ensure(s, g, BUF_XFER, need);
    seg = 0;
    CopyList ents;
    for (uint32_t p = 0; p < W; p++) {
      ents.add(reinterpret_cast<char *>(g.peer[BUF_XFER][p]) + eoff[p] * 8, s.join_entries_out.p + seg, ecounts[p] * 8, p);
      seg += ecounts[p];
    }
    ents.run(s.stream, me, W);
    HUMID_CUDA(cudaStreamSynchronize(s.stream));
    group_barrier(g);
    ck(humid_gpu_shard_join_import(h, reinterpret_cast<uint64_t *>(g.buf[BUF_XFER].p), erecv[me]));
    clock.mark("join: sort + match + entries home");
  }
  ck(humid_gpu_shard_build_rows(h, nullptr));
  clock.mark("rows");

  // ---- clusters: closed components here, open ones gathered on every rank ------------------------------------
  uint64_t n_open = 0, n_open_e = 0;
  ck(humid_gpu_shard_cluster_local(h, &n_open, &n_open_e));
  clock.mark("clusters inside the shard");
  mine.slot[3] = n_open;
  mine.slot[4] = n_open_e;
  group_barrier(g);  // also: every rank is past join_import, the transfer arena is free again
  uint64_t NO = 0, EO = 0, open_off = 0, eo_off = 0;
  for (uint32_t r = 0; r < W; r++) {
    if (r < me) { open_off += sh.r[r].slot[3]; eo_off += sh.r[r].slot[4]; }
    NO += sh.r[r].slot[3];
    EO += sh.r[r].slot[4];
  }
  const size_t col4 = align_up(NO * 4);
  const size_t o_rank = 0, o_count = col4, o_first = 2 * col4, o_deg = 3 * col4, o_col = 4 * col4,
               o_ids = 4 * col4 + align_up(EO * 4), r_label = o_ids + col4, r_flag = r_label + col4, r_max = r_flag + col4,
               r_size = r_max + col4;
  for (uint32_t r = 0; r < W; r++) need[r] = r_size + align_up(NO * 8) + 256;
  group_ensure(s, g, BUF_XFER, need);
  {
    char *base = reinterpret_cast<char *>(g.buf[BUF_XFER].p);
    auto u32 = [&](size_t off) { return reinterpret_cast<uint32_t *>(base + off); };
    ck(humid_gpu_shard_export_open(h, u32(o_rank) + open_off, u32(o_count) + open_off, u32(o_first) + open_off,
                                   u32(o_deg) + open_off, u32(o_col) + eo_off));
    HUMID_CUDA(cudaMemsetAsync(base + o_ids, 0, col4, s.stream));
    const size_t cols[4] = {o_rank, o_count, o_first, o_deg};
    CopyList open;
    for (uint32_t p = 0; p < W; p++) {
      if (p == me) continue;
      char *dst = reinterpret_cast<char *>(g.peer[BUF_XFER][p]);
      for (size_t off : cols) open.add(dst + off + open_off * 4, base + off + open_off * 4, n_open * 4, p);
      open.add(dst + o_col + eo_off * 4, base + o_col + eo_off * 4, n_open_e * 4, p);
    }
    open.run(s.stream, me, W);
    HUMID_CUDA(cudaStreamSynchronize(s.stream));
    group_barrier(g);
    clock.mark("open components gathered");
    // every open component is clustered by ONE rank (root mod W); its words' results go to the ranks that own them
    s.open_owner_mod = W;
    s.open_owner_me = me;
    s.after_open_cluster = [&](State &o) {
      if (NO) {
        OpenResults R{};
        R.n = W;
        for (uint32_t r = 0; r < W; r++) R.off[r + 1] = R.off[r] + sh.r[r].slot[3];
        for (uint32_t p = 0; p < W; p++) {
          char *xb = reinterpret_cast<char *>(g.peer[BUF_XFER][p]);
          R.label[p] = reinterpret_cast<uint32_t *>(xb + r_label);
          R.flag[p] = reinterpret_cast<uint32_t *>(xb + r_flag);
          R.mx[p] = reinterpret_cast<uint32_t *>(xb + r_max);
          R.size[p] = reinterpret_cast<uint64_t *>(xb + r_size);
        }
        scatter_open_results_kernel<<<grid_for(NO, 256), 256, 0, s.stream>>>(o.comp.p, o.label.p, o.seed_flag.p,
                                                                            o.seed_size.p, o.seed_max.p, NO, W, me, R);
        HUMID_CUDA(cudaGetLastError());
      }
      HUMID_CUDA(cudaStreamSynchronize(s.stream));
      group_barrier(g);
      CopyList back;  // this rank's own words: what their components' owners found
      back.add(o.label.p + open_off, base + r_label + open_off * 4, n_open * 4, me);
      back.add(o.seed_flag.p + open_off, base + r_flag + open_off * 4, n_open * 4, me);
      back.add(o.seed_max.p + open_off, base + r_max + open_off * 4, n_open * 4, me);
      back.add(o.seed_size.p + open_off, base + r_size + open_off * 8, n_open * 8, me);
      back.run(s.stream, me, W);
    };
    struct ClearHook {
      State &s;
      ~ClearHook() { s.after_open_cluster = nullptr; }
    } clear_hook{s};
    uint64_t n_cl = 0;
    ck(humid_gpu_shard_import_open(h, open_off, NO, EO, u32(o_rank), u32(o_count), u32(o_first), u32(o_deg), u32(o_col),
                                   &n_cl));
    mine.slot[5] = n_cl;
    group_barrier(g);
    uint64_t id_off = 0, clusters_total = 0;
    for (uint32_t r = 0; r < W; r++) {
      if (r < me) id_off += sh.r[r].slot[5];
      clusters_total += sh.r[r].slot[5];
    }
    clock.mark("open components clustered");
    // ids of the open seeds: every rank fills in the ones it owns, in every rank's copy
    ck(humid_gpu_shard_assign_ids(h, id_off, u32(o_ids)));
    CopyList ids;
    for (uint32_t p = 0; p < W; p++)
      if (p != me)
        ids.add(reinterpret_cast<char *>(g.peer[BUF_XFER][p]) + o_ids + open_off * 4, base + o_ids + open_off * 4, n_open * 4, p);
    ids.run(s.stream, me, W);
    HUMID_CUDA(cudaStreamSynchronize(s.stream));
    group_barrier(g);
    ck(humid_gpu_shard_finish(h, u32(o_ids), g.n_total, usable_total, bases[W], clusters_total));
    clock.mark("cluster ids");
  }

  // ---- outputs: keep bits to the ranks that fed the tuples, histograms merged --------------------------------
  {
    LocalBits L{};
    L.n = g.n_res;
    uint64_t off_words = 0;
    for (uint32_t q = 0; q < g.n_res; q++) {
      L.start[q] = g.res_start[q];
      L.range_off[q] = off_words;
      off_words += bits_bytes(g.res_start[q + 1] - g.res_start[q]) / 4;
    }
    L.start[g.n_res] = g.res_start[g.n_res];
    s.keep_bits.ensure(off_words + 2);
    HUMID_CUDA(cudaMemsetAsync(s.keep_bits.p, 0, off_words * 4, s.stream));
    if (s.n_clusters) {
      Timed t(s, "keep", s.n_clusters, s.n_clusters * 8);
      local_keep_bits_kernel<<<grid_for(s.n_clusters, 256), 256, 0, s.stream>>>(s.cluster_first.p, s.n_clusters, L,
                                                                              s.keep_bits.p);
      HUMID_CUDA(cudaGetLastError());
    }
    CopyList slices;
    for (uint32_t q = 0; q < g.n_res; q++) {  // my bitmap of range q -> slice `me` in the result buffer of its holder
      const uint64_t m = g.res_start[q + 1] - g.res_start[q];
      const size_t bb = bits_bytes(m);
      char *dst = reinterpret_cast<char *>(g.peer[BUF_RESULT][g.res_rank[q]]) + bb + align_up(m * 4) + (size_t)me * bb;
      slices.add(dst, s.keep_bits.p + L.range_off[q], bb, g.res_rank[q]);
    }
    slices.run(s.stream, me, W);
  }
  stage_histograms_enqueue(s);  // this shard's words / lists / clusters, on their way to page-locked memory
  HUMID_CUDA(cudaStreamSynchronize(s.stream));
  stage_histograms_collect(s);
  for (int w = 0; w < 3; w++)
    if (!s.hist_done[w]) stage_histogram(s, w);  // more large values than the eager form lists
  for (int w = 0; w < 3; w++) {
    const size_t m = s.hist_k[w].size();
    if (m > G_HIST_CAP) throw CudaError{"a histogram of one shard has more than " + std::to_string(G_HIST_CAP) + " keys"};
    mine.hist_n[w] = (uint32_t)m;
    std::copy(s.hist_k[w].begin(), s.hist_k[w].end(), mine.hist_k[w]);
    std::copy(s.hist_v[w].begin(), s.hist_v[w].end(), mine.hist_v[w]);
  }
  group_barrier(g);  // every slice has landed, every histogram is in the segment
  {
    const size_t bb = g.my_bits_bytes;
    char *rb = reinterpret_cast<char *>(g.buf[BUF_RESULT].p);
    if (g.my_res_n)
      or_slices_kernel<<<grid_for(bb / 4, 256), 256, 0, s.stream>>>(
          reinterpret_cast<const uint32_t *>(rb + bb + align_up(g.my_res_n * 4)), W, bb / 4, bb / 4,
          reinterpret_cast<uint32_t *>(rb));
    HUMID_CUDA(cudaGetLastError());
    s.keep_host_words = (g.my_res_n + 31) / 32 + 2;
    s.keep_host.ensure(s.keep_host_words);
    if (g.my_res_n)
      HUMID_CUDA(cudaMemcpyAsync(s.keep_host.p, rb, s.keep_host_words * 4, cudaMemcpyDeviceToHost, s.stream));
  }
  for (int w = 0; w < 3; w++) {
    std::map<uint64_t, uint64_t> merged;
    for (uint32_t r = 0; r < W; r++)
      for (uint32_t i = 0; i < sh.r[r].hist_n[w]; i++) merged[sh.r[r].hist_k[w][i]] += sh.r[r].hist_v[w][i];
    s.hist_k[w].clear();
    s.hist_v[w].clear();
    for (const auto &kv : merged) {
      s.hist_k[w].push_back(kv.first);
      s.hist_v[w].push_back(kv.second);
    }
    s.hist_done[w] = true;
  }
  HUMID_CUDA(cudaStreamSynchronize(s.stream));
  group_barrier(g);  // the segment and the slices may be written again
  resolve_stats(s);
  clock.mark("keep bits + histograms");
}

static void group_collect_ids_body(humid_gpu *h) {
  State &s = h->s;
  GroupLink &g = *s.group;
  HUMID_CUDA(cudaSetDevice(s.device));
  // filtered tuples never travelled: their id stays 0
  HUMID_CUDA(cudaMemsetAsync(g.buf[BUF_RESULT].p + g.my_bits_bytes, 0, g.my_res_n * 4, s.stream));
  HUMID_CUDA(cudaStreamSynchronize(s.stream));
  group_barrier(g);
  if (s.src_n) {
    stage_ensure_lookup(s);  // key -> local rank
    Timed t(s, "cluster_ids", s.src_n, s.src_n * 36);
    push_tuple_ids_kernel<<<grid_for(s.src_n, 256), 256, 0, s.stream>>>(s.src_keys, s.src_index, s.src_n, s.lookup(),
                                                                        s.word_cid.p, result_table(g));
    HUMID_CUDA(cudaGetLastError());
  }
  HUMID_CUDA(cudaStreamSynchronize(s.stream));
  group_barrier(g);
  resolve_stats(s);
  g.ids_collected = true;
}

// ---- what humid_gpu.cu asks about a handle that is a rank of a group ---------------------------------------------
bool group_active(const State &s) { return s.group != nullptr; }
bool group_batches_numbering(const State &s) { return s.group && s.group->numbering == HUMID_GROUP_BATCHES; }
void group_note_batch(State &s, uint64_t local, uint64_t global, uint64_t n) {
  if (s.group) s.group->batches.push_back({local, global, n});
}
void group_reset(State &s) {
  if (!s.group) return;
  s.group->batches.clear();
  s.group->ids_collected = false;
}
uint64_t group_result_tuples(const State &s) { return s.group ? s.group->my_res_n : 0; }
// cluster ids of the tuples this rank holds results for (device memory), or nullptr before collect_ids
const uint32_t *group_result_ids(const State &s) {
  const GroupLink *g = s.group;
  if (!g || !g->ids_collected) return nullptr;
  return reinterpret_cast<const uint32_t *>(g->buf[BUF_RESULT].p + g->my_bits_bytes);
}

}  // namespace humid

using namespace humid;

// a failing rank tells the others (they leave their barriers with an error instead of waiting)
static int group_guarded(humid_gpu *h, void (*body)(humid_gpu *)) {
  int rc = guarded(h, [&] { body(h); });
  if (rc != HUMID_GPU_OK && h->s.group && h->s.group->sh) h->s.group->sh->failed.store(1, std::memory_order_relaxed);
  return rc;
}

extern "C" {

humid_gpu_group *humid_gpu_group_create(uint32_t world) {
  if (world == 0 || world > G_MAX) return nullptr;
  humid_gpu_group *g = new (std::nothrow) humid_gpu_group();
  if (!g) return nullptr;
  std::memset(static_cast<void *>(&g->sh), 0, sizeof(g->sh));
  g->sh.world = world;
  g->sh.magic.store(G_MAGIC, std::memory_order_release);
  return g;
}

void humid_gpu_group_destroy(humid_gpu_group *g) { delete g; }

int humid_gpu_group_join(humid_gpu *h, humid_gpu_group *grp, const char *shm_name, uint32_t rank, uint32_t world,
                         int tuple_numbering) {
  if (!h) return HUMID_GPU_ERR_ARG;
  State &s = h->s;
  if (s.group) return fail(h, HUMID_GPU_ERR_ARG, "group_join: the handle is already in a group");
  if (world == 0 || world > G_MAX || rank >= world) return fail(h, HUMID_GPU_ERR_ARG, "group_join: bad rank / world");
  if (tuple_numbering != HUMID_GROUP_SLICES && tuple_numbering != HUMID_GROUP_BATCHES)
    return fail(h, HUMID_GPU_ERR_ARG, "group_join: unknown tuple numbering");
  if (!grp && (!shm_name || shm_name[0] != '/')) return fail(h, HUMID_GPU_ERR_ARG, "group_join: shm_name must start with '/'");
  if (s.n_tuples || s.finalized) return fail(h, HUMID_GPU_ERR_ARG, "group_join: join before the first batch");
  if (s.wide()) return fail(h, HUMID_GPU_ERR_UNSUPPORTED, "words of more than 31 letters are not sharded (use one GPU)");
  if (s.cfg.edit && s.cfg.distance >= 2) return fail(h, HUMID_GPU_ERR_UNSUPPORTED, "edit distance >= 2 is not sharded (use one GPU)");
  int rc = guarded(h, [&] {
    HUMID_CUDA(cudaSetDevice(s.device));
    GroupLink *g = new GroupLink();
    s.group = g;
    g->rank = rank;
    g->world = world;
    g->numbering = tuple_numbering;
    g->inproc = grp != nullptr;
    if (grp) {
      if (grp->sh.world != world) throw CudaError{"group_join: world differs from the group's"};
      g->sh = &grp->sh;
    } else {
      g->shm_name = shm_name;
      int fd = -1;
      if (rank == 0) {
        shm_unlink(shm_name);
        fd = shm_open(shm_name, O_CREAT | O_EXCL | O_RDWR, 0600);
        if (fd < 0 || ftruncate(fd, sizeof(GroupShared)) != 0) {
          if (fd >= 0) close(fd);
          throw CudaError{std::string("group_join: cannot create shared memory segment ") + shm_name};
        }
      } else {
        const auto t0 = std::chrono::steady_clock::now();
        while ((fd = shm_open(shm_name, O_RDWR, 0600)) < 0) {
          if (std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() > 60.0)
            throw CudaError{std::string("group_join: shared memory segment ") + shm_name + " did not appear"};
          std::this_thread::sleep_for(std::chrono::milliseconds(2));
        }
        struct stat st;
        while (fstat(fd, &st) == 0 && (size_t)st.st_size < sizeof(GroupShared)) {  // rank 0 is still sizing it
          if (std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() > 60.0) {
            close(fd);
            throw CudaError{"group_join: shared memory segment was never sized"};
          }
          std::this_thread::sleep_for(std::chrono::milliseconds(2));
        }
      }
      void *p = mmap(nullptr, sizeof(GroupShared), PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
      close(fd);
      if (p == MAP_FAILED) throw CudaError{"group_join: mmap of the shared segment failed"};
      g->sh = static_cast<GroupShared *>(p);
      if (rank == 0) {  // a fresh segment is all zeroes
        g->sh->world = world;
        g->sh->magic.store(G_MAGIC, std::memory_order_release);
      } else {
        const auto t0 = std::chrono::steady_clock::now();
        while (g->sh->magic.load(std::memory_order_acquire) != G_MAGIC) {
          if (std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() > 60.0)
            throw CudaError{"group_join: rank 0 never initialised the shared segment"};
          std::this_thread::yield();
        }
        if (g->sh->world != world) throw CudaError{"group_join: world differs from rank 0's"};
      }
    }
    g->sh->r[rank].device = s.device;
    for (int b = 0; b < G_BUFS; b++) {
      g->sh->r[rank].bytes[b] = 0;
      g->sh->r[rank].ptr[b] = 0;
    }
    group_barrier(*g);
    if (g->inproc) {  // between processes the IPC mapping enables peer access itself
      for (uint32_t r = 0; r < world; r++) {
        const int peer = g->sh->r[r].device;
        if (peer == s.device) continue;
        int can = 0;
        HUMID_CUDA(cudaDeviceCanAccessPeer(&can, s.device, peer));
        if (!can) throw CudaError{"group_join: no peer access between devices " + std::to_string(s.device) + " and " + std::to_string(peer)};
        cudaError_t e = cudaDeviceEnablePeerAccess(peer, 0);
        if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
        else HUMID_CUDA(e);
      }
    }
    group_barrier(*g);
    // the (small, for now) result buffers are mapped here already: a box without CUDA IPC between the ranks fails in
    // join, where the caller can still choose another transport, not in the middle of a job
    uint64_t need[G_MAX];
    for (uint32_t r = 0; r < world; r++) need[r] = 4096;
    group_ensure(s, *g, BUF_RESULT, need);
    s.cfg.defer_table = 1;  // words are counted by the owners of their key range, not as batches arrive
  });
  if (rc != HUMID_GPU_OK && s.group) {
    if (s.group->sh) s.group->sh->failed.store(1, std::memory_order_relaxed);
    if (s.group->inproc || !s.group->sh) { delete s.group; s.group = nullptr; }
    else group_release_handle(s);
  }
  return rc;
}

int humid_gpu_group_finalize(humid_gpu *h) {
  if (!h) return HUMID_GPU_ERR_ARG;
  if (!h->s.group) return fail(h, HUMID_GPU_ERR_ARG, "group_finalize: the handle is not in a group");
  if (h->s.finalized) return fail(h, HUMID_GPU_ERR_ARG, "finalize called twice");
  return group_guarded(h, group_finalize_body);
}

int humid_gpu_group_collect_ids(humid_gpu *h) {
  if (!h) return HUMID_GPU_ERR_ARG;
  if (!h->s.group || !h->s.finalized) return fail(h, HUMID_GPU_ERR_ARG, "group_collect_ids: call humid_gpu_group_finalize first");
  if (h->s.group->ids_collected) return HUMID_GPU_OK;
  return group_guarded(h, group_collect_ids_body);
}

int humid_gpu_group_phases(humid_gpu *h, const char **names, float *ms, uint32_t *n) {
  if (!h || !n) return HUMID_GPU_ERR_ARG;
  if (!h->s.group) return fail(h, HUMID_GPU_ERR_ARG, "group_phases: the handle is not in a group");
  const auto &ph = h->s.group->phases;
  const uint32_t cap = *n;
  *n = (uint32_t)ph.size();
  if (names && ms) {
    if (cap < ph.size()) return fail(h, HUMID_GPU_ERR_ARG, "group_phases: buffers too small");
    for (size_t i = 0; i < ph.size(); i++) {
      names[i] = ph[i].first;
      ms[i] = ph[i].second;
    }
  }
  return HUMID_GPU_OK;
}

}  // extern "C"
