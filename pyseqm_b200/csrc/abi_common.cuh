// Host-side helpers shared by the extern "C" entry points.
#pragma once
#include <cuda_runtime.h>
#include <atomic>
#include <cstdio>
#include <string>

#include "../../include/seqm_b200.h"
#include "common.cuh"

namespace seqm {

void set_error(const std::string& s);
// launch accounting and the optional event profiler (seqm_profile_*): prof_mark(name) opens a region that
// lasts until the next mark on the same stream
extern std::atomic<unsigned long long> g_launches;      // entry points may be called from several host threads
void prof_mark(const char* name, cudaStream_t st);
// 512 bytes of page-locked host memory owned by the calling thread (allocated on first use, kept for the life of the
// thread's library use): the status word a host-driven loop reads back; nullptr if the allocation fails
unsigned long long* host_status_slot();

// SP2 sharding hook of the SCF driver (sp2_kernels.cu, seqm_scf_sp2_comm)
extern void* g_scf_sp2_comm;
extern int g_scf_sp2_nranks;

inline int fail(int code, const char* what) {
  set_error(what);
  return code;
}

#define SEQM_REQUIRE(cond, msg) \
  do { if (!(cond)) return ::seqm::fail(SEQM_EINVAL, msg); } while (0)

#define SEQM_CHECK_LAUNCH(name)                                                            \
  do {                                                                                     \
    ++::seqm::g_launches;                                                                  \
    cudaError_t e_ = cudaGetLastError();                                                   \
    if (e_ != cudaSuccess) {                                                               \
      ::seqm::set_error(std::string(name) + ": " + cudaGetErrorString(e_));                \
      return SEQM_ECUDA;                                                                   \
    }                                                                                      \
  } while (0)

#define SEQM_CUDA(call)                                                                    \
  do {                                                                                     \
    cudaError_t e_ = (call);                                                               \
    if (e_ != cudaSuccess) {                                                               \
      ::seqm::set_error(std::string(#call) + ": " + cudaGetErrorString(e_));               \
      return SEQM_ECUDA;                                                                   \
    }                                                                                      \
  } while (0)

inline unsigned cdiv(long long a, long long b) { return (unsigned)((a + b - 1) / b); }

// class-tiled storage of the two-centre integrals: [tile][element][32 pairs]
__device__ __forceinline__ size_t wtile_index(int slot, int e, int nelem) {
  return ((size_t)(slot >> 5) * nelem + e) * 32 + (slot & 31);
}

}  // namespace seqm
