// common.cuh -- shared host/device helpers for libsystemml_b200.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "../../include/sysml_b200.h"

typedef long long i64;

// ------------------------------------------------------------------------------------------------
// Context: the native half of GPUContext (J/runtime/instructions/gpu/context/GPUContext.java:93-184).
// One device, one stream, a grow-only scratch workspace, a block of device counters used by the
// "last block done" reductions, pinned staging for the host-vector entry points and two events.
// ------------------------------------------------------------------------------------------------
struct sysml_b200_ctx {
  int device;
  int sm_count;
  int max_smem_optin;
  cudaStream_t stream;
  bool own_stream;
  void* ws;          // scratch (device)
  size_t ws_bytes;
  bool ws_external;  // the block belongs to the caller (sysml_b200_ctx_set_workspace): never grown or freed here
  size_t ws_needed;  // size of the last request an external block could not satisfy
  size_t ws_prefix;  // bytes at the start of the block that the current operation keeps for itself (nested requests start after them)
  unsigned int* counters;  // 64 zero-initialised counters; every kernel that uses one resets it
  void* pinned;      // pinned host staging
  size_t pinned_bytes;
  void* vec_stage;   // small device vectors for the host-vector entry points
  size_t vec_stage_bytes;
  cudaEvent_t ev0, ev1;
  i64 launches;
  char err[512];
};

extern thread_local char g_sysml_tls_err[512];

int sysml_set_error(sysml_b200_ctx* ctx, int code, const char* fmt, ...);
// Returns a device scratch pointer of at least `bytes` (grows with cudaMalloc when needed).
int sysml_workspace(sysml_b200_ctx* ctx, size_t bytes, void** out);
int sysml_pinned(sysml_b200_ctx* ctx, size_t bytes, void** out);

#define SYSML_CHECK_CTX(ctx)                                   \
  do {                                                         \
    if (!(ctx)) return sysml_set_error(nullptr, SYSML_E_INVALID, "null context"); \
  } while (0)

#define SYSML_CUDA(ctx, call)                                                                      \
  do {                                                                                             \
    cudaError_t _e = (call);                                                                       \
    if (_e != cudaSuccess)                                                                         \
      return sysml_set_error((ctx), _e == cudaErrorMemoryAllocation ? SYSML_E_OOM : SYSML_E_CUDA, \
                             "%s failed: %s (%s:%d)", #call, cudaGetErrorString(_e), __FILE__, __LINE__); \
  } while (0)

#define SYSML_LAUNCH_CHECK(ctx, name)                                                             \
  do {                                                                                             \
    cudaError_t _e = cudaGetLastError();                                                           \
    if (_e != cudaSuccess)                                                                         \
      return sysml_set_error((ctx), SYSML_E_CUDA, "launch of %s failed: %s", name, cudaGetErrorString(_e)); \
    (ctx)->launches++;                                                                             \
  } while (0)

static inline int sysml_set_device(sysml_b200_ctx* ctx) {
  int cur = -1;
  cudaGetDevice(&cur);
  if (cur != ctx->device) {
    cudaError_t e = cudaSetDevice(ctx->device);
    if (e != cudaSuccess) return sysml_set_error(ctx, SYSML_E_CUDA, "cudaSetDevice(%d): %s", ctx->device, cudaGetErrorString(e));
  }
  return SYSML_OK;
}
#define SYSML_ENTER(ctx)              \
  SYSML_CHECK_CTX(ctx);               \
  do {                                \
    int _rc = sysml_set_device(ctx);  \
    if (_rc) return _rc;              \
  } while (0)

// ------------------------------------------------------------------------------------------------
// Device helpers
// ------------------------------------------------------------------------------------------------
#ifdef __CUDACC__

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// ---- mbarrier / bulk-copy (TMA engine, 1-D) wrappers.  PTX ISA: mbarrier.*, cp.async.bulk ----
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  while (!mbar_try_wait(bar, parity)) {
  }
}
__device__ __forceinline__ uint64_t l2_policy_evict_first() {
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
// global -> shared bulk copy completing on an mbarrier (SASS: UBLKCP).  dst/src 16B aligned, bytes % 16 == 0.
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar, uint64_t policy) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(
          smem_u32(dst_smem)),
      "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar)), "l"(policy)
      : "memory");
}
// asks the memory system to bring [src, src+bytes) into L2 (no shared-memory destination, no completion to wait for);
// src 16B aligned, bytes % 16 == 0
__device__ __forceinline__ void bulk_prefetch_l2(const void* src_gmem, uint32_t bytes) {
  asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(src_gmem), "r"(bytes) : "memory");
}
__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// ---- programmatic dependent launch (PTX griddepcontrol).  Both are no-ops in a kernel that was launched without the
// programmatic-stream-serialization attribute.  launch_dependents: the next kernel of the stream may start its CTAs
// as soon as every CTA of this grid has executed it (or exited); wait: blocks until the PREVIOUS kernel of the stream
// has completed and its memory is visible -- nothing that kernel wrote may be read, and nothing it reads may be
// overwritten, before this returns. ----
__device__ __forceinline__ void griddep_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void griddep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

// ---- streaming loads that do not pollute L1 ----
__device__ __forceinline__ double ldg_stream(const double* p) {
  double v;
  asm volatile("ld.global.nc.L1::no_allocate.f64 %0, [%1];" : "=d"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ double2 ldg_stream2(const double2* p) {
  double2 v;
  asm volatile("ld.global.nc.L1::no_allocate.v2.f64 {%0,%1}, [%2];" : "=d"(v.x), "=d"(v.y) : "l"(p));
  return v;
}

// ---- warp reductions ----
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// ---- error-free transformation: (s, e) with s = fl(a+b), a+b = s+e exactly (Knuth two-sum) ----
__device__ __forceinline__ void two_sum(double a, double b, double& s, double& e) {
  s = __dadd_rn(a, b);
  double bb = __dsub_rn(s, a);
  e = __dadd_rn(__dsub_rn(a, __dsub_rn(s, bb)), __dsub_rn(b, bb));
}
// compensated accumulator: value = hi + lo
struct CompSum {
  double hi, lo;
  __device__ __forceinline__ CompSum() : hi(0.0), lo(0.0) {}
  __device__ __forceinline__ void add(double x) {
    double s, e;
    two_sum(hi, x, s, e);
    hi = s;
    lo = __dadd_rn(lo, e);
  }
  __device__ __forceinline__ void add(const CompSum& o) {
    add(o.hi);
    lo = __dadd_rn(lo, o.lo);
  }
  __device__ __forceinline__ double value() const { return __dadd_rn(hi, lo); }
};

#endif  // __CUDACC__
