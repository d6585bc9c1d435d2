// comm.cu -- sum-exchange of small per-GPU partial results over NVLink peer memory, fused into the
// finishing step of the producing kernel.
//
// What it replaces.  SystemML's distributed tsmm / mmchain plan maps every row block locally and then adds
// the m-sized partial results (RDDAggregateUtils.sumStable, J/runtime/instructions/spark/utils/
// RDDAggregateUtils.java:58-70, called by MapmmChainSPInstruction / TsmmSPInstruction).  Here a row block
// lives on one GPU and the sum is an all-reduce of m doubles (8 KB at m = 1000): pure latency.  Instead
// of launching NCCL after the kernel, the kernel that folds the per-CTA partials of the local shard
// also performs the exchange:
//   1. each CTA reduces 32 columns of the local partials (32 row slices x 32 columns, compensated),
//   2. stores the 32 values into the mailbox of EVERY peer (plain P2P stores through NVLink / NVSwitch),
//   3. publishes a per-CTA flag with release semantics at system scope,
//   4. waits (acquire) for the same CTA's flag from every peer and adds the world's vectors in rank order,
//   5. (CG step only) the CTA that finishes last applies the m-vector tail of the conjugate-gradient
//      iteration to the summed vector, so a CG step is two launches, not three,
// so all ranks end with bit-identical results and no second launch, no host involvement, no NCCL call.
// Mailboxes are double-buffered by call parity; flags carry the call epoch, so nothing is ever reset.
//
// Ranks are either processes (one per GPU, as torchrun launches them: the mailbox is exported with
// cudaIpcGetMemHandle and opened by the peers, comm_connect) or contexts of ONE process (the way the
// reference drives several GPUs: every parfor worker thread of one JVM owns a GPUContext and calls
// cudaSetDevice, GPUContext.java:181-184; comm_connect_local enables peer access and hands out raw pointers).
//
// Failure detection.  A flag is a 64-bit word {signature, epoch}: the signature packs (count, combiner, kind)
// of the call, so two ranks that issue different calls notice instead of adding unrelated vectors.  Every
// wait is bounded by a time budget (%globaltimer); when it runs out, or a signature differs, the kernel
// writes an error record to pinned host memory and ends; the next API call on the communicator (or
// comm_check) returns SYSML_E_CUDA with the peer and the epoch, and the communicator stays failed.
#include "common.cuh"
#include "cellwise_ops.cuh"
#include "cg_tail.cuh"

struct sysml_b200_comm {
  sysml_b200_ctx* ctx;
  int rank, world;
  int m_pad;     // doubles per slot
  int ncta_max;  // flags per slot
  unsigned int epoch[2];
  unsigned long long calls;
  unsigned long long timeout_ns;
  void* local;         // this rank's mailbox block
  void* peer_base[8];  // mapped base of every rank's block (own entry = local)
  bool opened[8];      // peer_base[p] came from cudaIpcOpenMemHandle
  unsigned long long* host_err;  // 2 words of mapped pinned host memory written by the kernel on failure
  unsigned long long* dev_err;   // the same words as the device sees them
  bool failed;
};

namespace {

constexpr int MAXW = 8;
enum { COMM_ERR_TIMEOUT = 1, COMM_ERR_MISMATCH = 2 };

struct PeerArgs {
  double* data[MAXW];               // data[p]  : peer p's mailbox data   [2][world][m_pad]
  unsigned long long* flags[MAXW];  // flags[p] : peer p's mailbox flags  [2][world][ncta_max]
  int world, rank, parity, m_pad, ncta_max;
  unsigned int epoch, sig;
  unsigned long long timeout_ns;
  unsigned long long* err;  // mapped host memory
};

struct CgArgs {  // m-vector tail of a CG iteration, applied by the last CTA (cg_tail.cuh)
  double *p, *r, *beta, *nr2;
  double lambda;
  unsigned int* ticket;
};

__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ double ld_relaxed_sys(const double* p) {
  double v;
  asm volatile("ld.relaxed.sys.global.f64 %0, [%1];" : "=d"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ void report_failure(const PeerArgs& pa, int code, int peer, unsigned int theirs) {
  volatile unsigned long long* e = pa.err;
  e[1] = ((unsigned long long)theirs << 32) | pa.sig;
  __threadfence_system();
  e[0] = (unsigned long long)code | ((unsigned long long)peer << 8) | ((unsigned long long)blockIdx.x << 16) | ((unsigned long long)pa.epoch << 32);
}

// partial: [P][m] local per-CTA partial rows (or, with P == 1, the local vector itself).
// Every rank launches this kernel with the same m and grid; CTA c of every rank exchanges columns [32c, 32c+32).
// NS row slices x 32 columns per CTA.  The kernel may be launched as the programmatic dependent of the kernel
// that writes `partial`: everything before griddep_wait() overlaps that kernel's tail.
template <int ROP, int NS, bool CG>  // ROP: 0 sum, 1 max, 2 min (Math.max/min semantics) over the ranks; the local P partial rows are always summed
__global__ void __launch_bounds__(NS * 32) finish_exchange_kernel(const double* partial, int P, int m, double* out,
                                                                   const PeerArgs pa, const CgArgs cg) {
  __shared__ double sh_hi[NS][33], sh_lo[NS][33];
  __shared__ int sh_last;
  const int c = threadIdx.x & 31, sl = threadIdx.x >> 5;
  const int j = blockIdx.x * 32 + c;
  griddep_launch_dependents();
  griddep_wait();
  CompSum a0, a1, a2, a3;
  if (j < m && P > 1) {
    int p = sl;
    for (; p + 3 * NS < P; p += 4 * NS) {
      const double x0 = __ldcg(partial + (size_t)p * m + j);
      const double x1 = __ldcg(partial + (size_t)(p + NS) * m + j);
      const double x2 = __ldcg(partial + (size_t)(p + 2 * NS) * m + j);
      const double x3 = __ldcg(partial + (size_t)(p + 3 * NS) * m + j);
      a0.add(x0);
      a1.add(x1);
      a2.add(x2);
      a3.add(x3);
    }
    double y0 = 0.0, y1 = 0.0, y2 = 0.0;
    const bool h0 = p < P, h1 = p + NS < P, h2 = p + 2 * NS < P;
    if (h0) y0 = __ldcg(partial + (size_t)p * m + j);
    if (h1) y1 = __ldcg(partial + (size_t)(p + NS) * m + j);
    if (h2) y2 = __ldcg(partial + (size_t)(p + 2 * NS) * m + j);
    if (h0) a0.add(y0);
    if (h1) a1.add(y1);
    if (h2) a2.add(y2);
    a0.add(a1);
    a2.add(a3);
    a0.add(a2);
  }
  sh_hi[sl][c] = a0.hi;
  sh_lo[sl][c] = a0.lo;
  __syncthreads();
  if (sl == 0) {  // warp 0 finishes the local sum and does the exchange
    double local = 0.0;
    if (j < m) {
#pragma unroll 4
      for (int q = 1; q < NS; q++) {
        CompSum t;
        t.hi = sh_hi[q][c];
        t.lo = sh_lo[q][c];
        a0.add(t);
      }
      local = (P == 1) ? __ldcg(partial + j) : a0.value();  // a single row passes through untouched (keeps -0.0 for max/min)
    }
    double s = local;
    if (pa.world > 1) {
      const size_t slot = (size_t)(pa.parity * pa.world + pa.rank);
      // 2. push this rank's 32 values into every mailbox (own one included)
      if (j < m) {
        for (int p = 0; p < pa.world; p++) pa.data[p][slot * pa.m_pad + j] = local;
      }
      __threadfence_system();
      __syncwarp();
      // 3. publish: lane p raises this CTA's flag at peer p
      const unsigned long long word = ((unsigned long long)pa.sig << 32) | pa.epoch;
      if (c < pa.world) st_release_sys(pa.flags[c] + slot * pa.ncta_max + blockIdx.x, word);
      // 4. wait for the same CTA of every rank (bounded), then add in rank order
      if (c < pa.world) {
        const unsigned long long* f = pa.flags[pa.rank] + (size_t)(pa.parity * pa.world + c) * pa.ncta_max + blockIdx.x;
        const unsigned long long t0 = globaltimer_ns();
        unsigned long long v;
        unsigned int spins = 0;
        for (;;) {
          v = ld_acquire_sys(f);
          if ((unsigned int)v == pa.epoch) break;
          if ((++spins & 255u) == 0 && globaltimer_ns() - t0 > pa.timeout_ns) break;
        }
        if ((unsigned int)v != pa.epoch) report_failure(pa, COMM_ERR_TIMEOUT, c, (unsigned int)v);
        else if ((unsigned int)(v >> 32) != pa.sig) report_failure(pa, COMM_ERR_MISMATCH, c, (unsigned int)(v >> 32));
      }
      __syncwarp();
      if (j < m) {
        const double* mine = pa.data[pa.rank];
        s = ld_relaxed_sys(mine + (size_t)(pa.parity * pa.world) * pa.m_pad + j);
        for (int src = 1; src < pa.world; src++) {
          const double x = ld_relaxed_sys(mine + (size_t)(pa.parity * pa.world + src) * pa.m_pad + j);
          s = (ROP == 0) ? __dadd_rn(s, x) : (ROP == 1) ? sysml_ops::jmax(s, x) : sysml_ops::jmin(s, x);
        }
      }
    }
    if (j < m) out[j] = s;
  }
  if (CG) {
    // 5. the CTA that finishes last owns the whole (summed) vector: apply the CG tail to it
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
      const unsigned int t = atomicAdd(cg.ticket, 1u);
      sh_last = (t == gridDim.x - 1) ? 1 : 0;
      if (sh_last) *cg.ticket = 0;  // everybody has arrived: leave the counter at zero for the next call
    }
    __syncthreads();
    if (sh_last) {
      __threadfence();
      cg_vector_update_body(cg.p, out, cg.r, cg.beta, cg.nr2, (i64)m, cg.lambda, &sh_hi[0][0]);
    }
  }
}

int decode_failure(sysml_b200_comm* comm) {
  const unsigned long long e0 = ((volatile unsigned long long*)comm->host_err)[0];
  if (!e0) return SYSML_OK;
  const unsigned long long e1 = ((volatile unsigned long long*)comm->host_err)[1];
  comm->failed = true;
  const int code = (int)(e0 & 0xff), peer = (int)((e0 >> 8) & 0xff), cta = (int)((e0 >> 16) & 0xffff);
  const unsigned int epoch = (unsigned int)(e0 >> 32);
  if (code == COMM_ERR_MISMATCH) {
    const unsigned int theirs = (unsigned int)(e1 >> 32), mine = (unsigned int)e1;
    return sysml_set_error(comm->ctx, SYSML_E_CUDA,
                           "allreduce: rank %d and peer %d issued different calls at epoch %u (this rank: %u values, op %u, kind %u; peer: %u values, op %u, kind %u)",
                           comm->rank, peer, epoch, mine & 0xffffffu, (mine >> 24) & 0xf, mine >> 28, theirs & 0xffffffu, (theirs >> 24) & 0xf, theirs >> 28);
  }
  return sysml_set_error(comm->ctx, SYSML_E_CUDA, "allreduce: peer %d did not arrive at epoch %u within %.0f ms (rank %d, CTA %d); the communicator is unusable",
                         peer, epoch, comm->timeout_ns / 1e6, comm->rank, cta);
}

template <int ROP, int NS, bool CG>
int launch_finish(sysml_b200_ctx* ctx, int ncta, bool pdl, const double* partial, int P, int m, double* out, const PeerArgs& pa, const CgArgs& cg) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3((unsigned)ncta);
  cfg.blockDim = dim3(NS * 32);
  cfg.dynamicSmemBytes = 0;
  cfg.stream = ctx->stream;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = pdl ? 1 : 0;
  SYSML_CUDA(ctx, cudaLaunchKernelEx(&cfg, finish_exchange_kernel<ROP, NS, CG>, partial, P, m, out, pa, cg));
  ctx->launches++;
  return SYSML_OK;
}

// With lazy module loading (the CUDA 12 default) the first launch of a kernel loads it, and loading may wait for work
// already running on the device.  An exchange kernel that is spinning for a peer on the same device -- ranks that
// share a GPU, or a peer whose first call needs a variant this rank has not used yet -- would then never see that
// peer arrive.  So every variant is made resident when a communicator is created, before any exchange can be in flight.
template <int ROP, int NS, bool CG>
cudaError_t preload_one() {
  cudaFuncAttributes fa;
  return cudaFuncGetAttributes(&fa, finish_exchange_kernel<ROP, NS, CG>);
}
cudaError_t preload_exchange_kernels() {
  cudaError_t e = preload_one<0, 32, true>();
  if (e == cudaSuccess) e = preload_one<0, 32, false>();
  if (e == cudaSuccess) e = preload_one<1, 32, false>();
  if (e == cudaSuccess) e = preload_one<2, 32, false>();
  if (e == cudaSuccess) e = preload_one<0, 1, false>();
  if (e == cudaSuccess) e = preload_one<1, 1, false>();
  if (e == cudaSuccess) e = preload_one<2, 1, false>();
  return e;
}

}  // namespace

// used by mmchain.cu (fused finish) and by sysml_b200_allreduce below.  comm may be null when cg is given
// (single GPU: local finish + CG tail).  pdl: launch as the programmatic dependent of the previous kernel.
int sysml_launch_finish_exchange(sysml_b200_ctx* ctx, sysml_b200_comm* comm, const double* partial, int P, int m, double* out, int rop,
                                 bool pdl, double* cg_p, double* cg_r, double* cg_beta, double* cg_nr2, double cg_lambda) {
  const bool with_cg = cg_p != nullptr;
  PeerArgs pa{};
  pa.world = 1;
  if (comm) {
    if (comm->ctx != ctx) return sysml_set_error(ctx, SYSML_E_INVALID, "allreduce: communicator belongs to another context");
    if (comm->failed) return sysml_set_error(ctx, SYSML_E_CUDA, "allreduce: the communicator failed earlier (%s)", ctx->err);
    int rc = decode_failure(comm);
    if (rc) return rc;
    if (m > comm->m_pad) return sysml_set_error(ctx, SYSML_E_LIMIT, "allreduce: %d values exceed the communicator's capacity of %d", m, comm->m_pad);
    for (int p = 0; p < comm->world; p++)
      if (!comm->peer_base[p]) return sysml_set_error(ctx, SYSML_E_INVALID, "allreduce: communicator is not connected (peer %d)", p);
    pa.world = comm->world;
    pa.rank = comm->rank;
    pa.m_pad = comm->m_pad;
    pa.ncta_max = comm->ncta_max;
    pa.parity = (int)(comm->calls & 1ull);
    pa.epoch = ++comm->epoch[pa.parity];
    pa.sig = ((unsigned int)m & 0xffffffu) | ((unsigned int)rop << 24) | ((with_cg ? 1u : 0u) << 28);
    pa.timeout_ns = comm->timeout_ns;
    pa.err = comm->dev_err;
    comm->calls++;
    const size_t data_bytes = (size_t)2 * comm->world * comm->m_pad * sizeof(double);
    for (int p = 0; p < comm->world; p++) {
      pa.data[p] = (double*)comm->peer_base[p];
      pa.flags[p] = (unsigned long long*)((char*)comm->peer_base[p] + data_bytes);
    }
  } else if (!with_cg) {
    return sysml_set_error(ctx, SYSML_E_INVALID, "allreduce: null communicator");
  }
  const int ncta = (m + 31) / 32;
  CgArgs cg{cg_p, cg_r, cg_beta, cg_nr2, cg_lambda, ctx->counters + 63};
  if (with_cg) return launch_finish<0, 32, true>(ctx, ncta, pdl, partial, P, m, out, pa, cg);
  if (P == 1) {  // plain vector exchange: one warp per CTA is all that works
    if (rop == SYSML_REDUCE_MAX) return launch_finish<1, 1, false>(ctx, ncta, pdl, partial, P, m, out, pa, cg);
    if (rop == SYSML_REDUCE_MIN) return launch_finish<2, 1, false>(ctx, ncta, pdl, partial, P, m, out, pa, cg);
    return launch_finish<0, 1, false>(ctx, ncta, pdl, partial, P, m, out, pa, cg);
  }
  if (rop == SYSML_REDUCE_MAX) return launch_finish<1, 32, false>(ctx, ncta, pdl, partial, P, m, out, pa, cg);
  if (rop == SYSML_REDUCE_MIN) return launch_finish<2, 32, false>(ctx, ncta, pdl, partial, P, m, out, pa, cg);
  return launch_finish<0, 32, false>(ctx, ncta, pdl, partial, P, m, out, pa, cg);
}

int sysml_launch_allreduce_finish(sysml_b200_ctx* ctx, sysml_b200_comm* comm, const double* partial, int P, int m, double* out, int rop) {
  if (!comm) return sysml_set_error(ctx, SYSML_E_INVALID, "allreduce: null communicator");
  return sysml_launch_finish_exchange(ctx, comm, partial, P, m, out, rop, false, nullptr, nullptr, nullptr, nullptr, 0.0);
}

extern "C" {

int sysml_b200_comm_create(sysml_b200_ctx* ctx, int rank, int world, int64_t max_values, sysml_b200_comm** out, void* handle_out64) {
  SYSML_ENTER(ctx);
  if (!out) return sysml_set_error(ctx, SYSML_E_INVALID, "comm_create: null output");
  *out = nullptr;
  if (world < 1 || world > MAXW || rank < 0 || rank >= world) return sysml_set_error(ctx, SYSML_E_INVALID, "comm_create: rank %d / world %d (at most %d ranks)", rank, world, MAXW);
  if (max_values < 1 || max_values > (1 << 20)) return sysml_set_error(ctx, SYSML_E_LIMIT, "comm_create: max_values must be in [1, 2^20]");
  sysml_b200_comm* c = new sysml_b200_comm();
  memset(c, 0, sizeof(*c));
  c->ctx = ctx;
  c->rank = rank;
  c->world = world;
  c->m_pad = (int)((max_values + 31) / 32) * 32;
  c->ncta_max = c->m_pad / 32;
  c->timeout_ns = 60ull * 1000000000ull;
  const size_t bytes = (size_t)2 * world * c->m_pad * sizeof(double) + (size_t)2 * world * c->ncta_max * sizeof(unsigned long long);
  cudaError_t e = cudaMalloc(&c->local, bytes);
  if (e != cudaSuccess) {
    delete c;
    return sysml_set_error(ctx, SYSML_E_OOM, "comm_create: cudaMalloc(%zu): %s", bytes, cudaGetErrorString(e));
  }
  auto fail = [&](int code, const char* what, cudaError_t err) {
    if (c->host_err) cudaFreeHost(c->host_err);
    cudaFree(c->local);
    delete c;
    return sysml_set_error(ctx, code, "comm_create: %s: %s", what, cudaGetErrorString(err));
  };
  // zero the mailbox on the context's own stream and wait: no peer may see (or be handed) the block before that
  e = cudaMemsetAsync(c->local, 0, bytes, ctx->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
  if (e != cudaSuccess) return fail(SYSML_E_CUDA, "clearing the mailbox", e);
  e = cudaHostAlloc((void**)&c->host_err, 2 * sizeof(unsigned long long), cudaHostAllocMapped | cudaHostAllocPortable);
  if (e != cudaSuccess) {
    c->host_err = nullptr;
    return fail(SYSML_E_OOM, "cudaHostAlloc(error record)", e);
  }
  c->host_err[0] = c->host_err[1] = 0;
  e = cudaHostGetDevicePointer((void**)&c->dev_err, c->host_err, 0);
  if (e != cudaSuccess) return fail(SYSML_E_CUDA, "cudaHostGetDevicePointer", e);
  e = preload_exchange_kernels();
  if (e != cudaSuccess) return fail(SYSML_E_CUDA, "loading the exchange kernels", e);
  if (handle_out64) {
    cudaIpcMemHandle_t h;
    e = cudaIpcGetMemHandle(&h, c->local);
    if (e != cudaSuccess) return fail(SYSML_E_CUDA, "cudaIpcGetMemHandle", e);
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    memcpy(handle_out64, &h, 64);
  }
  c->peer_base[rank] = c->local;
  *out = c;
  return SYSML_OK;
}

int sysml_b200_comm_connect(sysml_b200_comm* comm, const void* all_handles) {
  if (!comm) return sysml_set_error(nullptr, SYSML_E_INVALID, "comm_connect: null communicator");
  sysml_b200_ctx* ctx = comm->ctx;
  SYSML_ENTER(ctx);
  if (!all_handles) return sysml_set_error(ctx, SYSML_E_INVALID, "comm_connect: null handles");
  for (int p = 0; p < comm->world; p++) {
    if (p == comm->rank) continue;
    cudaIpcMemHandle_t h;
    memcpy(&h, (const char*)all_handles + (size_t)p * 64, 64);
    void* base = nullptr;
    cudaError_t e = cudaIpcOpenMemHandle(&base, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) return sysml_set_error(ctx, SYSML_E_CUDA, "comm_connect: cudaIpcOpenMemHandle(rank %d): %s", p, cudaGetErrorString(e));
    comm->peer_base[p] = base;
    comm->opened[p] = true;
  }
  return SYSML_OK;
}

// Ranks that are contexts of ONE process (threads of one JVM, each bound to its GPU): no IPC, the peers' mailboxes
// are reachable through their ordinary device pointers once peer access is enabled.  comms[r] must be rank r.
int sysml_b200_comm_connect_local(sysml_b200_comm** comms, int n) {
  if (!comms || n < 1 || n > MAXW) return sysml_set_error(nullptr, SYSML_E_INVALID, "comm_connect_local: need 1..%d communicators", MAXW);
  for (int r = 0; r < n; r++) {
    if (!comms[r]) return sysml_set_error(nullptr, SYSML_E_INVALID, "comm_connect_local: communicator %d is null", r);
    if (comms[r]->rank != r || comms[r]->world != n)
      return sysml_set_error(comms[r]->ctx, SYSML_E_INVALID, "comm_connect_local: entry %d is rank %d of %d, expected rank %d of %d", r, comms[r]->rank, comms[r]->world, r, n);
    if (comms[r]->m_pad != comms[0]->m_pad) return sysml_set_error(comms[r]->ctx, SYSML_E_INVALID, "comm_connect_local: capacities differ between ranks");
  }
  int prev = -1;
  cudaGetDevice(&prev);
  for (int r = 0; r < n; r++) {
    sysml_b200_ctx* ctx = comms[r]->ctx;
    cudaError_t e = cudaSetDevice(ctx->device);
    if (e != cudaSuccess) return sysml_set_error(ctx, SYSML_E_CUDA, "comm_connect_local: cudaSetDevice(%d): %s", ctx->device, cudaGetErrorString(e));
    for (int p = 0; p < n; p++) {
      const int pd = comms[p]->ctx->device;
      if (pd != ctx->device) {
        int can = 0;
        cudaDeviceCanAccessPeer(&can, ctx->device, pd);
        if (!can) return sysml_set_error(ctx, SYSML_E_UNSUPPORTED, "comm_connect_local: device %d cannot access device %d", ctx->device, pd);
        e = cudaDeviceEnablePeerAccess(pd, 0);
        if (e == cudaErrorPeerAccessAlreadyEnabled) {
          cudaGetLastError();
        } else if (e != cudaSuccess) {
          return sysml_set_error(ctx, SYSML_E_CUDA, "comm_connect_local: cudaDeviceEnablePeerAccess(%d -> %d): %s", ctx->device, pd, cudaGetErrorString(e));
        }
      }
      comms[r]->peer_base[p] = comms[p]->local;
      comms[r]->opened[p] = false;
    }
  }
  if (prev >= 0) cudaSetDevice(prev);
  return SYSML_OK;
}

int sysml_b200_comm_set_timeout_ms(sysml_b200_comm* comm, int64_t ms) {
  if (!comm) return sysml_set_error(nullptr, SYSML_E_INVALID, "comm_set_timeout_ms: null communicator");
  if (ms < 1) return sysml_set_error(comm->ctx, SYSML_E_INVALID, "comm_set_timeout_ms: the budget must be at least 1 ms");
  comm->timeout_ns = (unsigned long long)ms * 1000000ull;
  return SYSML_OK;
}

int sysml_b200_comm_check(sysml_b200_comm* comm) {
  if (!comm) return sysml_set_error(nullptr, SYSML_E_INVALID, "comm_check: null communicator");
  sysml_b200_ctx* ctx = comm->ctx;
  SYSML_ENTER(ctx);
  SYSML_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  if (comm->failed) return sysml_set_error(ctx, SYSML_E_CUDA, "allreduce: the communicator failed earlier");
  return decode_failure(comm);
}

int sysml_b200_comm_destroy(sysml_b200_comm* comm) {
  if (!comm) return SYSML_OK;
  sysml_b200_ctx* ctx = comm->ctx;
  SYSML_ENTER(ctx);
  cudaStreamSynchronize(ctx->stream);
  for (int p = 0; p < comm->world; p++)
    if (comm->opened[p]) cudaIpcCloseMemHandle(comm->peer_base[p]);
  if (comm->local) cudaFree(comm->local);
  if (comm->host_err) cudaFreeHost(comm->host_err);
  delete comm;
  return SYSML_OK;
}

int sysml_b200_allreduce(sysml_b200_ctx* ctx, sysml_b200_comm* comm, double* buf, int64_t len, int op) {
  SYSML_ENTER(ctx);
  if (len < 0 || (len > 0 && !buf)) return sysml_set_error(ctx, SYSML_E_INVALID, "allreduce: bad arguments");
  if (op != SYSML_REDUCE_SUM && op != SYSML_REDUCE_MAX && op != SYSML_REDUCE_MIN) return sysml_set_error(ctx, SYSML_E_INVALID, "allreduce: unknown op %d", op);
  if (!comm) return sysml_set_error(ctx, SYSML_E_INVALID, "allreduce: null communicator");
  // longer buffers go through the mailbox in capacity-sized pieces (tsmm results: m*m values)
  for (int64_t off = 0; off < len; off += comm->m_pad) {
    const int64_t n = (len - off < comm->m_pad) ? len - off : comm->m_pad;
    int rc = sysml_launch_allreduce_finish(ctx, comm, buf + off, 1, (int)n, buf + off, op);
    if (rc) return rc;
  }
  return SYSML_OK;
}

int sysml_b200_allreduce_sum(sysml_b200_ctx* ctx, sysml_b200_comm* comm, double* buf, int64_t len) {
  return sysml_b200_allreduce(ctx, comm, buf, len, SYSML_REDUCE_SUM);
}

}  // extern "C"
