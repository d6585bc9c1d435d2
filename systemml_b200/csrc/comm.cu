// comm.cu -- sum-exchange of small per-GPU partial results over NVLink peer memory, fused into the
// finishing step of the producing kernel.
//
// What it replaces.  SystemML's distributed tsmm / mmchain plan maps every row block locally and then adds
// the m-sized partial results (RDDAggregateUtils.sumStable, J/runtime/instructions/spark/utils/
// RDDAggregateUtils.java:58-70, called by MapmmChainSPInstruction / TsmmSPInstruction).  Here a row block
// lives on one GPU and the sum is an all-reduce of m doubles (8 KB at m = 1000): pure latency.  Instead
// of launching NCCL after the kernel, the kernel that folds the per-CTA partials of the local shard
// (colpartial_reduce in mmchain.cu) also performs the exchange:
//   1. each CTA reduces 32 columns of the local partials,
//   2. stores the 32 values into the mailbox of EVERY peer (plain P2P stores through NVLink / NVSwitch),
//   3. publishes a per-CTA flag with release semantics at system scope,
//   4. waits (acquire) for the same CTA's flag from every peer and adds the world's vectors in rank order,
// so all ranks end with bit-identical results and no second launch, no host involvement, no NCCL call.
// Mailboxes are double-buffered by call parity; flags carry the call epoch, so nothing is ever reset.
// The mailbox of a rank is one cudaMalloc block exported with cudaIpcGetMemHandle and opened by the
// peers (one process per GPU, as torchrun launches them).
#include "common.cuh"
#include "cellwise_ops.cuh"

struct sysml_b200_comm {
  sysml_b200_ctx* ctx;
  int rank, world;
  int m_pad;     // doubles per slot
  int ncta_max;  // flags per slot
  unsigned int epoch[2];
  unsigned long long calls;
  void* local;         // this rank's mailbox block
  void* peer_base[8];  // mapped base of every rank's block (own entry = local)
  bool opened[8];
};

namespace {

constexpr int MAXW = 8;

struct PeerArgs {
  double* data[MAXW];         // data[p]  : peer p's mailbox data   [2][world][m_pad]
  unsigned int* flags[MAXW];  // flags[p] : peer p's mailbox flags  [2][world][ncta_max]
  int world, rank, parity, m_pad, ncta_max;
  unsigned int epoch;
};

__device__ __forceinline__ void st_release_sys(unsigned int* p, unsigned int v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned int ld_acquire_sys(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ double ld_relaxed_sys(const double* p) {
  double v;
  asm volatile("ld.relaxed.sys.global.f64 %0, [%1];" : "=d"(v) : "l"(p) : "memory");
  return v;
}

// partial: [P][m] local per-CTA partial rows (or, with P == 1, the local vector itself).
// Every rank launches this kernel with the same m and grid; CTA c of every rank exchanges columns
// [32c, 32c+32).
template <int ROP>  // 0 sum, 1 max, 2 min (Math.max/min semantics) over the ranks; the local P partial rows are always summed
__global__ void __launch_bounds__(256) colpartial_allreduce_kernel(const double* __restrict__ partial, int P, int m,
                                                                   double* __restrict__ out, const PeerArgs pa) {
  __shared__ double sh_hi[8][33], sh_lo[8][33];
  const int c = threadIdx.x & 31, sl = threadIdx.x >> 5;
  const int j = blockIdx.x * 32 + c;
  CompSum a0, a1, a2, a3;
  if (j < m) {
    int p = sl;
    for (; p + 24 < P; p += 32) {
      const double x0 = __ldcg(partial + (size_t)p * m + j);
      const double x1 = __ldcg(partial + (size_t)(p + 8) * m + j);
      const double x2 = __ldcg(partial + (size_t)(p + 16) * m + j);
      const double x3 = __ldcg(partial + (size_t)(p + 24) * m + j);
      a0.add(x0);
      a1.add(x1);
      a2.add(x2);
      a3.add(x3);
    }
    for (; p < P; p += 8) a0.add(__ldcg(partial + (size_t)p * m + j));
    a0.add(a1);
    a2.add(a3);
    a0.add(a2);
  }
  sh_hi[sl][c] = a0.hi;
  sh_lo[sl][c] = a0.lo;
  __syncthreads();
  if (sl != 0) return;  // warp 0 finishes the local sum and does the exchange
  double local = 0.0;
  if (j < m) {
#pragma unroll
    for (int q = 1; q < 8; q++) {
      CompSum t;
      t.hi = sh_hi[q][c];
      t.lo = sh_lo[q][c];
      a0.add(t);
    }
    local = (P == 1) ? __ldcg(partial + j) : a0.value();  // a single row passes through untouched (keeps -0.0 for max/min)
  }
  const size_t slot = (size_t)(pa.parity * pa.world + pa.rank);
  // 2. push this rank's 32 values into every mailbox (own one included)
  if (j < m) {
    for (int p = 0; p < pa.world; p++) pa.data[p][slot * pa.m_pad + j] = local;
  }
  __threadfence_system();
  __syncwarp();
  // 3. publish: lane p raises this CTA's flag at peer p
  if (c < pa.world) st_release_sys(pa.flags[c] + slot * pa.ncta_max + blockIdx.x, pa.epoch);
  // 4. wait for the same CTA of every rank, then add in rank order
  if (c < pa.world) {
    const unsigned int* f = pa.flags[pa.rank] + (size_t)(pa.parity * pa.world + c) * pa.ncta_max + blockIdx.x;
    while (ld_acquire_sys(f) != pa.epoch) {
    }
  }
  __syncwarp();
  if (j < m) {
    const double* mine = pa.data[pa.rank];
    double s = ld_relaxed_sys(mine + (size_t)(pa.parity * pa.world) * pa.m_pad + j);
    for (int src = 1; src < pa.world; src++) {
      const double x = ld_relaxed_sys(mine + (size_t)(pa.parity * pa.world + src) * pa.m_pad + j);
      s = (ROP == 0) ? __dadd_rn(s, x) : (ROP == 1) ? sysml_ops::jmax(s, x) : sysml_ops::jmin(s, x);
    }
    out[j] = s;
  }
}

}  // namespace

// used by mmchain.cu (fused finish) and by sysml_b200_allreduce_sum below
int sysml_launch_allreduce_finish(sysml_b200_ctx* ctx, sysml_b200_comm* comm, const double* partial, int P, int m, double* out, int rop) {
  if (!comm) return sysml_set_error(ctx, SYSML_E_INVALID, "allreduce: null communicator");
  if (comm->ctx != ctx) return sysml_set_error(ctx, SYSML_E_INVALID, "allreduce: communicator belongs to another context");
  if (m > comm->m_pad) return sysml_set_error(ctx, SYSML_E_LIMIT, "allreduce: %d values exceed the communicator's capacity of %d", m, comm->m_pad);
  const int ncta = (m + 31) / 32;
  for (int p = 0; p < comm->world; p++)
    if (!comm->peer_base[p]) return sysml_set_error(ctx, SYSML_E_INVALID, "allreduce: communicator is not connected (peer %d)", p);
  PeerArgs pa{};
  pa.world = comm->world;
  pa.rank = comm->rank;
  pa.m_pad = comm->m_pad;
  pa.ncta_max = comm->ncta_max;
  pa.parity = (int)(comm->calls & 1ull);
  pa.epoch = ++comm->epoch[pa.parity];
  comm->calls++;
  const size_t data_bytes = (size_t)2 * comm->world * comm->m_pad * sizeof(double);
  for (int p = 0; p < comm->world; p++) {
    pa.data[p] = (double*)comm->peer_base[p];
    pa.flags[p] = (unsigned int*)((char*)comm->peer_base[p] + data_bytes);
  }
  if (rop == SYSML_REDUCE_MAX) colpartial_allreduce_kernel<1><<<ncta, 256, 0, ctx->stream>>>(partial, P, m, out, pa);
  else if (rop == SYSML_REDUCE_MIN) colpartial_allreduce_kernel<2><<<ncta, 256, 0, ctx->stream>>>(partial, P, m, out, pa);
  else colpartial_allreduce_kernel<0><<<ncta, 256, 0, ctx->stream>>>(partial, P, m, out, pa);
  SYSML_LAUNCH_CHECK(ctx, "colpartial_allreduce_kernel");
  return SYSML_OK;
}

extern "C" {

int sysml_b200_comm_create(sysml_b200_ctx* ctx, int rank, int world, int64_t max_values, sysml_b200_comm** out, void* handle_out64) {
  SYSML_ENTER(ctx);
  if (!out || !handle_out64) return sysml_set_error(ctx, SYSML_E_INVALID, "comm_create: null output");
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
  const size_t bytes = (size_t)2 * world * c->m_pad * sizeof(double) + (size_t)2 * world * c->ncta_max * sizeof(unsigned int);
  cudaError_t e = cudaMalloc(&c->local, bytes);
  if (e != cudaSuccess) {
    delete c;
    return sysml_set_error(ctx, SYSML_E_OOM, "comm_create: cudaMalloc(%zu): %s", bytes, cudaGetErrorString(e));
  }
  cudaMemset(c->local, 0, bytes);
  cudaIpcMemHandle_t h;
  e = cudaIpcGetMemHandle(&h, c->local);
  if (e != cudaSuccess) {
    cudaFree(c->local);
    delete c;
    return sysml_set_error(ctx, SYSML_E_CUDA, "comm_create: cudaIpcGetMemHandle: %s", cudaGetErrorString(e));
  }
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  memcpy(handle_out64, &h, 64);
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

int sysml_b200_comm_destroy(sysml_b200_comm* comm) {
  if (!comm) return SYSML_OK;
  sysml_b200_ctx* ctx = comm->ctx;
  SYSML_ENTER(ctx);
  cudaStreamSynchronize(ctx->stream);
  for (int p = 0; p < comm->world; p++)
    if (comm->opened[p]) cudaIpcCloseMemHandle(comm->peer_base[p]);
  if (comm->local) cudaFree(comm->local);
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
