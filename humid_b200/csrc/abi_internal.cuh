// Shared by the translation units that implement the C ABI (humid_gpu.cu, shard.cu).
#pragma once
#include <cstring>
#include <functional>

#include "state.cuh"

using namespace humid;

struct humid_gpu {
  State s;
};

inline std::string &create_error() {
  static thread_local std::string e;
  return e;
}

inline int fail(humid_gpu *h, int code, const std::string &msg) {
  if (h) h->s.last_error = msg; else create_error() = msg;
  return code;
}

// Runs `body`, mapping C++ failures onto the C error codes.
inline int guarded(humid_gpu *h, const std::function<void()> &body) {
  try {
    body();
    return HUMID_GPU_OK;
  } catch (const CudaError &e) {
    bool oom = e.msg.find("out of memory") != std::string::npos;
    return fail(h, oom ? HUMID_GPU_ERR_MEMORY : HUMID_GPU_ERR_CUDA, e.msg);
  } catch (const std::bad_alloc &) {
    return fail(h, HUMID_GPU_ERR_MEMORY, "host allocation failed");
  } catch (const std::exception &e) {
    return fail(h, HUMID_GPU_ERR_CUDA, e.what());
  }
}


namespace humid {
void resolve_stats(State &s);
// group.cu: what the single-handle entry points need to know about a handle that is a rank of a group
void group_release_handle(State &s);
bool group_active(const State &s);
bool group_batches_numbering(const State &s);
void group_note_batch(State &s, uint64_t local, uint64_t global, uint64_t n);
void group_reset(State &s);
uint64_t group_result_tuples(const State &s);
const uint32_t *group_result_ids(const State &s);
}
