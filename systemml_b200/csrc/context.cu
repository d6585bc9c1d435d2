// context.cu -- context, device memory and timers of libsystemml_b200.so.
// Native counterpart of GPUContext / GPUContextPool / GPUMemoryManager / GPUObject copies
// (J/runtime/instructions/gpu/context/*.java); see include/sysml_b200.h for the per-function citations.
#include <stdarg.h>

#include "common.cuh"

thread_local char g_sysml_tls_err[512] = {0};

int sysml_set_error(sysml_b200_ctx* ctx, int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_sysml_tls_err, sizeof(g_sysml_tls_err), fmt, ap);
  va_end(ap);
  if (ctx) memcpy(ctx->err, g_sysml_tls_err, sizeof(ctx->err));
  return code;
}

int sysml_workspace(sysml_b200_ctx* ctx, size_t bytes, void** out) {
  // an operation that keeps data of its own at the start of the block (ws_prefix) sized the block for its nested
  // requests beforehand: those are served after the prefix and must never move the block
  const size_t need = bytes + ctx->ws_prefix;
  if (need > ctx->ws_bytes && ctx->ws_external) {
    ctx->ws_needed = need;
    return sysml_set_error(ctx, SYSML_E_WORKSPACE, "scratch: this operation needs %zu bytes, the caller's workspace holds %zu (sysml_b200_ctx_set_workspace)",
                           need, ctx->ws_bytes);
  }
  if (need > ctx->ws_bytes) {
    if (ctx->ws_prefix) return sysml_set_error(ctx, SYSML_E_CUDA, "scratch: nested request of %zu bytes does not fit behind a %zu-byte prefix", bytes, ctx->ws_prefix);
    // grow-only; growing synchronises the stream because in-flight kernels may still use the old block
    SYSML_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (ctx->ws) SYSML_CUDA(ctx, cudaFree(ctx->ws));
    ctx->ws = nullptr;
    ctx->ws_bytes = 0;
    size_t want = need < (size_t(8) << 20) ? (size_t(8) << 20) : need + need / 4;
    SYSML_CUDA(ctx, cudaMalloc(&ctx->ws, want));
    ctx->ws_bytes = want;
  }
  *out = (char*)ctx->ws + ctx->ws_prefix;
  return SYSML_OK;
}

int sysml_pinned(sysml_b200_ctx* ctx, size_t bytes, void** out) {
  if (bytes > ctx->pinned_bytes) {
    SYSML_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (ctx->pinned) SYSML_CUDA(ctx, cudaFreeHost(ctx->pinned));
    ctx->pinned = nullptr;
    ctx->pinned_bytes = 0;
    size_t want = bytes < (size_t(1) << 20) ? (size_t(1) << 20) : bytes + bytes / 4;
    SYSML_CUDA(ctx, cudaMallocHost(&ctx->pinned, want));
    ctx->pinned_bytes = want;
  }
  *out = ctx->pinned;
  return SYSML_OK;
}

extern "C" {

const char* sysml_b200_version(void) { return "systemml_b200 0.1 (sm_100a)"; }

int sysml_b200_device_count(int* count) {
  if (!count) return sysml_set_error(nullptr, SYSML_E_INVALID, "count is null");
  cudaError_t e = cudaGetDeviceCount(count);
  if (e != cudaSuccess) {
    *count = 0;
    return sysml_set_error(nullptr, SYSML_E_CUDA, "cudaGetDeviceCount: %s", cudaGetErrorString(e));
  }
  return SYSML_OK;
}

static int ctx_create_impl(int device, cudaStream_t stream, bool own, sysml_b200_ctx** out) {
  if (!out) return sysml_set_error(nullptr, SYSML_E_INVALID, "out is null");
  *out = nullptr;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess) return sysml_set_error(nullptr, SYSML_E_CUDA, "cudaGetDeviceCount: %s", cudaGetErrorString(e));
  if (device < 0 || device >= ndev) return sysml_set_error(nullptr, SYSML_E_INVALID, "device %d out of range [0,%d)", device, ndev);
  sysml_b200_ctx* ctx = new sysml_b200_ctx();
  memset(ctx, 0, sizeof(*ctx));
  ctx->device = device;
#define CTX_TRY(call)                                                                         \
  do {                                                                                        \
    cudaError_t _e = (call);                                                                  \
    if (_e != cudaSuccess) {                                                                  \
      int rc = sysml_set_error(nullptr, SYSML_E_CUDA, "%s: %s", #call, cudaGetErrorString(_e)); \
      delete ctx;                                                                             \
      return rc;                                                                              \
    }                                                                                         \
  } while (0)
  CTX_TRY(cudaSetDevice(device));
  cudaDeviceProp prop;
  CTX_TRY(cudaGetDeviceProperties(&prop, device));
  if (prop.major < 10) {
    int rc = sysml_set_error(nullptr, SYSML_E_UNSUPPORTED, "device %d is sm_%d%d; this library is built for sm_100a only", device, prop.major, prop.minor);
    delete ctx;
    return rc;
  }
  ctx->sm_count = prop.multiProcessorCount;
  ctx->max_smem_optin = (int)prop.sharedMemPerBlockOptin;
  if (own) {
    CTX_TRY(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
  } else {
    ctx->stream = stream;
  }
  ctx->own_stream = own;
  CTX_TRY(cudaMalloc((void**)&ctx->counters, 64 * sizeof(unsigned int)));
  CTX_TRY(cudaMemset(ctx->counters, 0, 64 * sizeof(unsigned int)));
  CTX_TRY(cudaEventCreate(&ctx->ev0));
  CTX_TRY(cudaEventCreate(&ctx->ev1));
#undef CTX_TRY
  *out = ctx;
  return SYSML_OK;
}

int sysml_b200_ctx_create(int device, sysml_b200_ctx** out) { return ctx_create_impl(device, nullptr, true, out); }
int sysml_b200_ctx_create_on_stream(int device, void* cuda_stream, sysml_b200_ctx** out) {
  return ctx_create_impl(device, (cudaStream_t)cuda_stream, false, out);
}

int sysml_b200_ctx_destroy(sysml_b200_ctx* ctx) {
  SYSML_ENTER(ctx);
  cudaStreamSynchronize(ctx->stream);
  if (ctx->ws && !ctx->ws_external) cudaFree(ctx->ws);
  if (ctx->counters) cudaFree(ctx->counters);
  if (ctx->pinned) cudaFreeHost(ctx->pinned);
  if (ctx->vec_stage) cudaFree(ctx->vec_stage);
  cudaEventDestroy(ctx->ev0);
  cudaEventDestroy(ctx->ev1);
  if (ctx->own_stream) cudaStreamDestroy(ctx->stream);
  delete ctx;
  return SYSML_OK;
}

int sysml_b200_ctx_set_workspace(sysml_b200_ctx* ctx, void* d_ptr, int64_t bytes) {
  SYSML_ENTER(ctx);
  if (bytes < 0 || (bytes > 0 && !d_ptr)) return sysml_set_error(ctx, SYSML_E_INVALID, "set_workspace: bad arguments");
  if (d_ptr && (reinterpret_cast<uintptr_t>(d_ptr) & 255)) return sysml_set_error(ctx, SYSML_E_INVALID, "set_workspace: the block must be 256-byte aligned");
  SYSML_CUDA(ctx, cudaStreamSynchronize(ctx->stream));  // queued kernels may still use the current block
  if (ctx->ws && !ctx->ws_external) SYSML_CUDA(ctx, cudaFree(ctx->ws));
  ctx->ws = d_ptr;
  ctx->ws_bytes = d_ptr ? (size_t)bytes : 0;
  ctx->ws_external = d_ptr != nullptr;
  ctx->ws_needed = 0;
  return SYSML_OK;
}

int64_t sysml_b200_ctx_workspace_needed(sysml_b200_ctx* ctx) { return ctx ? (int64_t)ctx->ws_needed : 0; }
int64_t sysml_b200_ctx_workspace_bytes(sysml_b200_ctx* ctx) { return ctx ? (int64_t)ctx->ws_bytes : 0; }

int sysml_b200_sync(sysml_b200_ctx* ctx) {
  SYSML_ENTER(ctx);
  SYSML_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return SYSML_OK;
}

const char* sysml_b200_last_error(sysml_b200_ctx* ctx) { return ctx ? ctx->err : g_sysml_tls_err; }
void* sysml_b200_ctx_stream(sysml_b200_ctx* ctx) { return ctx ? (void*)ctx->stream : nullptr; }
int sysml_b200_ctx_device(sysml_b200_ctx* ctx) { return ctx ? ctx->device : -1; }
int sysml_b200_ctx_sm_count(sysml_b200_ctx* ctx) { return ctx ? ctx->sm_count : 0; }
int64_t sysml_b200_launch_count(sysml_b200_ctx* ctx) { return ctx ? ctx->launches : 0; }

int sysml_b200_mem_info(sysml_b200_ctx* ctx, int64_t* free_bytes, int64_t* total_bytes) {
  SYSML_ENTER(ctx);
  size_t f = 0, t = 0;
  SYSML_CUDA(ctx, cudaMemGetInfo(&f, &t));
  if (free_bytes) *free_bytes = (int64_t)f;
  if (total_bytes) *total_bytes = (int64_t)t;
  return SYSML_OK;
}

int sysml_b200_malloc(sysml_b200_ctx* ctx, int64_t bytes, void** d_ptr) {
  SYSML_ENTER(ctx);
  if (!d_ptr || bytes < 0) return sysml_set_error(ctx, SYSML_E_INVALID, "malloc: bad arguments (bytes=%lld)", (long long)bytes);
  *d_ptr = nullptr;
  if (bytes == 0) return SYSML_OK;
  cudaError_t e = cudaMalloc(d_ptr, (size_t)bytes);
  if (e != cudaSuccess) {
    cudaGetLastError();
    return sysml_set_error(ctx, SYSML_E_OOM, "cudaMalloc(%lld bytes): %s", (long long)bytes, cudaGetErrorString(e));
  }
  return SYSML_OK;
}

int sysml_b200_free(sysml_b200_ctx* ctx, void* d_ptr) {
  SYSML_ENTER(ctx);
  if (!d_ptr) return SYSML_OK;
  SYSML_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  SYSML_CUDA(ctx, cudaFree(d_ptr));
  return SYSML_OK;
}

int sysml_b200_memset0(sysml_b200_ctx* ctx, void* d_ptr, int64_t bytes) {
  SYSML_ENTER(ctx);
  if (bytes < 0 || (!d_ptr && bytes)) return sysml_set_error(ctx, SYSML_E_INVALID, "memset0: bad arguments");
  if (bytes) SYSML_CUDA(ctx, cudaMemsetAsync(d_ptr, 0, (size_t)bytes, ctx->stream));
  return SYSML_OK;
}

int sysml_b200_h2d(sysml_b200_ctx* ctx, void* d_dst, const void* h_src, int64_t bytes) {
  SYSML_ENTER(ctx);
  if (bytes < 0 || ((!d_dst || !h_src) && bytes)) return sysml_set_error(ctx, SYSML_E_INVALID, "h2d: bad arguments");
  if (bytes) {
    SYSML_CUDA(ctx, cudaMemcpyAsync(d_dst, h_src, (size_t)bytes, cudaMemcpyHostToDevice, ctx->stream));
    SYSML_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  }
  return SYSML_OK;
}

int sysml_b200_d2h(sysml_b200_ctx* ctx, void* h_dst, const void* d_src, int64_t bytes) {
  SYSML_ENTER(ctx);
  if (bytes < 0 || ((!h_dst || !d_src) && bytes)) return sysml_set_error(ctx, SYSML_E_INVALID, "d2h: bad arguments");
  if (bytes) {
    SYSML_CUDA(ctx, cudaMemcpyAsync(h_dst, d_src, (size_t)bytes, cudaMemcpyDeviceToHost, ctx->stream));
    SYSML_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  }
  return SYSML_OK;
}

int sysml_b200_d2d(sysml_b200_ctx* ctx, void* d_dst, const void* d_src, int64_t bytes) {
  SYSML_ENTER(ctx);
  if (bytes < 0 || ((!d_dst || !d_src) && bytes)) return sysml_set_error(ctx, SYSML_E_INVALID, "d2d: bad arguments");
  if (bytes) SYSML_CUDA(ctx, cudaMemcpyAsync(d_dst, d_src, (size_t)bytes, cudaMemcpyDeviceToDevice, ctx->stream));
  return SYSML_OK;
}

int sysml_b200_host_alloc(sysml_b200_ctx* ctx, int64_t bytes, void** h_ptr) {
  SYSML_ENTER(ctx);
  if (!h_ptr || bytes <= 0) return sysml_set_error(ctx, SYSML_E_INVALID, "host_alloc: bad arguments");
  SYSML_CUDA(ctx, cudaMallocHost(h_ptr, (size_t)bytes));
  return SYSML_OK;
}

int sysml_b200_host_free(sysml_b200_ctx* ctx, void* h_ptr) {
  SYSML_ENTER(ctx);
  if (h_ptr) SYSML_CUDA(ctx, cudaFreeHost(h_ptr));
  return SYSML_OK;
}

int sysml_b200_timer_start(sysml_b200_ctx* ctx) {
  SYSML_ENTER(ctx);
  SYSML_CUDA(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
  return SYSML_OK;
}

int sysml_b200_timer_stop_ms(sysml_b200_ctx* ctx, double* ms) {
  SYSML_ENTER(ctx);
  if (!ms) return sysml_set_error(ctx, SYSML_E_INVALID, "ms is null");
  SYSML_CUDA(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
  SYSML_CUDA(ctx, cudaEventSynchronize(ctx->ev1));
  float f = 0;
  SYSML_CUDA(ctx, cudaEventElapsedTime(&f, ctx->ev0, ctx->ev1));
  *ms = (double)f;
  return SYSML_OK;
}

}  // extern "C"
