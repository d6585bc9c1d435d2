// mmchain.cu -- single-pass row-streaming kernels over a dense row-major X (n x m, FP64):
//
//   CHAIN : out = t(X) %*% (w * (X %*% v))   (and the XtXv / XtXvy variants)   -- mmchain
//   MV    : out = X %*% v                                                       -- dense matrix-vector
//   VM    : out = t(X) %*% y  (= t(t(y) %*% X))                                 -- dense vector-matrix
//
// Reference semantics: LibMatrixMult.matrixMultChainDense (LibMatrixMult.java:1594-1653),
// matrixMultDenseDenseMVShortRHS/TallRHS (:950-970), matrixMultDenseDenseVM (:972-993).
//
// Design (B200): all three are HBM-bound, X must cross HBM exactly once.  A persistent grid of
// (SMs x CTAs/SM) CTAs walks row tiles.  Full rows are contiguous, so a tile of R rows is ONE
// contiguous span of R*m*8 bytes: a producer warp moves it global->shared with a 1-D bulk copy
// (cp.async.bulk, the TMA engine; SASS UBLKCP) completing on an mbarrier, through a STAGES-deep
// ring so that several tiles per SM are in flight.  Consumer threads pull their columns of the
// tile into registers ONCE, use them for the X*v partial dot products (warp-shuffle reduction),
// release the stage, and reuse the same registers for the t(X)*tmp update, accumulating their
// columns of the result in registers across all tiles of the CTA.  Per-CTA column partials go to
// a scratch buffer; a small second kernel adds them up with compensated (two-sum) arithmetic in
// a fixed order, so the result is deterministic run to run.
//
// Algorithmic bytes per launch: 8*n*m (X once) + 8*m (v) + 8*m (out) [+ 8*n (w)];  flops 4*n*m.
#include <stdlib.h>

#include "common.cuh"
#include "cellwise_rt.cuh"

struct sysml_b200_comm;
int sysml_launch_finish_exchange(sysml_b200_ctx* ctx, sysml_b200_comm* comm, const double* partial, int P, int m, double* out, int rop,
                                 bool pdl, double* cg_p, double* cg_r, double* cg_beta, double* cg_nr2, double cg_lambda);

namespace {

enum { MODE_CHAIN = 0, MODE_MV = 1, MODE_VM = 2, MODE_FN = 3 };  // FN: a chain whose row function is a postfix program

struct StreamArgs {
  const double* X;
  const double* v;   // m-vector (CHAIN, MV)
  const double* w;   // n-vector: weights / y (CHAIN ct 1,2; VM: the left vector)
  double* out;       // CHAIN, VM: m ; MV: n
  double* partial;   // [gridDim.x * G][m] scratch (CHAIN, VM)
  i64 n;
  int m;
  int ct;
  i64 total_tiles;
  int use_tma;
  int w_smem;  // w rides the stage ring too (R doubles per stage after the row tiles) instead of being read by the consumers
  // host side only:
  sysml_b200_comm* comm;  // non-null -> the finishing kernel also does the cross-GPU sum
  int pdl;                // launch this kernel and its finishing kernel as programmatic dependents of their predecessors
  double *cg_r, *cg_beta, *cg_nr2;  // non-null -> the finishing kernel also applies the CG m-vector tail (v is p, out is q)
  double cg_lambda;
  // MODE_FN: t[i] = fn(eta[i], w[i], y2[i]); a second result goes to side[i]
  const double* y2;
  double* side;
  sysml_rt::RowProgram fn;
};

// NTC consumer threads (+32 producer), TC threads cooperate on one row, CPT columns per thread,
// RG rows per thread-group per tile, STAGES-deep ring.  G = NTC/TC groups, R = G*RG rows per tile.
// XSMEM (wide rows, 4096 < m <= 8192): a row no longer fits the consumers' registers next to its accumulators, so
// the tile is read from shared memory in BOTH phases (v and the accumulators stay in registers, one row per tile);
// the stage is released after the second phase.  X is still read from HBM once.
template <int MODE, int NTC, int TC, int CPT, int RG, int STAGES, bool XSMEM = false>
__global__ void __launch_bounds__(NTC + 32) rowstream_kernel(const StreamArgs a) {
  constexpr int G = NTC / TC;
  constexpr int R = G * RG;
  constexpr int NWG = TC / 32;  // warps per group
  extern __shared__ __align__(128) unsigned char smem_raw[];
  uint64_t* full = reinterpret_cast<uint64_t*>(smem_raw);
  uint64_t* empty = full + STAGES;
  double* red = reinterpret_cast<double*>(smem_raw + 128);          // [2][G][NWG][RG]
  constexpr int FNB = (MODE == MODE_FN) ? 2 * G * RG * 8 : 0;       // MODE_FN: [2][G][RG] results of the row function
  double* fnbuf = red + 2 * G * NWG * RG;
  double* tiles = reinterpret_cast<double*>(smem_raw + 128 + 2 * G * NWG * RG * 8 + FNB);
  // tiles must be 16B aligned: 128 + multiple of 16 (2*G*NWG*RG*8 is a multiple of 16)
  const int tid = threadIdx.x;
  const int m = a.m;
  const i64 n = a.n;
  const size_t stage_elems = (size_t)R * m;

  if (tid == 0) {
    for (int s = 0; s < STAGES; s++) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], NTC);
    }
    mbar_fence_init();
  }
  // launched as a programmatic dependent (CG step): the set-up above overlaps the previous kernel's tail; nothing in
  // global memory is touched before that kernel has completed.  Its own dependent (the finishing kernel) may be
  // made resident as soon as an SM has room; it waits for this grid to complete before it reads the partials.
  griddep_launch_dependents();
  __syncthreads();
  griddep_wait();

  if (tid >= NTC) {
    // ===================== producer warp =====================
    const int lane = tid - NTC;
    const uint64_t pol = l2_policy_evict_first();
    int it = 0;
    for (i64 tile = blockIdx.x; tile < a.total_tiles; tile += gridDim.x, ++it) {
      const int s = it % STAGES;
      const uint32_t ph = (uint32_t)(it / STAGES) & 1u;
      mbar_wait(&empty[s], ph ^ 1u);
      const i64 r0 = tile * R;
      const i64 rows = (n - r0 < R) ? (n - r0) : R;
      double* dst = tiles + (size_t)s * stage_elems;
      double* wdst = tiles + (size_t)STAGES * stage_elems + (size_t)s * R;
      const double* src = a.X + r0 * (i64)m;
      if (a.use_tma && rows == R) {
        if (lane == 0) {
          const uint32_t bytes = (uint32_t)(stage_elems * sizeof(double));
          mbar_expect_tx(&full[s], bytes + (a.w_smem ? R * 8u : 0u));
          bulk_g2s(dst, src, bytes, &full[s], pol);
          if (a.w_smem) bulk_g2s(wdst, a.w + r0, R * 8u, &full[s], pol);
        }
      } else {
        // ragged last tile / unaligned base: warp copy with zero padding of rows >= n
        const i64 valid = rows * (i64)m;
        if (a.w_smem)
          for (int i = lane; i < R; i += 32) wdst[i] = (i < rows) ? a.w[r0 + i] : 0.0;
        for (i64 i = lane; i < (i64)stage_elems; i += 32) dst[i] = (i < valid) ? ldg_stream(src + i) : 0.0;
        __syncwarp();
        if (lane == 0) mbar_arrive(&full[s]);
      }
    }
    return;
  }

  // ===================== consumer threads =====================
  const int g = tid / TC;
  const int tt = tid % TC;
  const int lane = tid & 31;
  const int wg = tt >> 5;

  double vreg[CPT];
  double c[CPT];
#pragma unroll
  for (int k = 0; k < CPT; k++) {
    const int j = tt + TC * k;
    vreg[k] = (MODE != MODE_VM && j < m) ? a.v[j] : 0.0;
    c[k] = 0.0;
  }

  int it = 0;
  for (i64 tile = blockIdx.x; tile < a.total_tiles; tile += gridDim.x, ++it) {
    const int s = it % STAGES;
    const uint32_t ph = (uint32_t)(it / STAGES) & 1u;
    const i64 r0 = tile * R + (i64)g * RG;

    double wv[RG];
    const bool need_w = MODE == MODE_VM || (MODE == MODE_CHAIN && a.ct != SYSML_CHAIN_XtXv) || (MODE == MODE_FN && a.w != nullptr);
    if (need_w && !a.w_smem) {
#pragma unroll
      for (int rr = 0; rr < RG; rr++) wv[rr] = (r0 + rr < n) ? __ldg(a.w + r0 + rr) : 0.0;
    } else {
#pragma unroll
      for (int rr = 0; rr < RG; rr++) wv[rr] = 0.0;
    }

    mbar_wait(&full[s], ph);
    const double* T = tiles + (size_t)s * stage_elems + (size_t)g * RG * m;
    if (need_w && a.w_smem) {
      // w arrived with the tile: a global load here sits on the critical path of every (short) tile
      const double* ws = tiles + (size_t)STAGES * stage_elems + (size_t)s * R + g * RG;
#pragma unroll
      for (int rr = 0; rr < RG; rr++) wv[rr] = ws[rr];
    }

    double x[XSMEM ? 1 : RG][XSMEM ? 1 : CPT];
    if (!XSMEM) {
#pragma unroll
      for (int rr = 0; rr < RG; rr++) {
#pragma unroll
        for (int k = 0; k < CPT; k++) {
          const int j = tt + TC * k;
          x[rr][k] = (j < m) ? T[rr * m + j] : 0.0;
        }
      }
    }

    double t[RG];
    if (MODE != MODE_VM) {
#pragma unroll
      for (int rr = 0; rr < RG; rr++) {
        double p = 0.0;
        if (XSMEM) {
#pragma unroll
          for (int k = 0; k < CPT; k++) {
            const int j = tt + TC * k;
            if (j < m) p = fma(T[rr * m + j], vreg[k], p);
          }
        } else {
#pragma unroll
          for (int k = 0; k < CPT; k++) p = fma(x[rr][k], vreg[k], p);
        }
        t[rr] = p;
      }
    }
    // every x value has been consumed (or, for VM, is ordered by the release below): free the stage
    if (!XSMEM) mbar_arrive(&empty[s]);

    if (MODE != MODE_VM) {
#pragma unroll
      for (int rr = 0; rr < RG; rr++) t[rr] = warp_sum(t[rr]);
      if (NWG > 1) {
        double* rbuf = red + ((size_t)(it & 1) * G + g) * NWG * RG;
        if (lane == 0) {
#pragma unroll
          for (int rr = 0; rr < RG; rr++) rbuf[wg * RG + rr] = t[rr];
        }
        named_bar_sync(1 + g, TC);
#pragma unroll
        for (int rr = 0; rr < RG; rr++) {
          double sacc = 0.0;
#pragma unroll
          for (int q = 0; q < NWG; q++) sacc += rbuf[q * RG + rr];
          t[rr] = sacc;
        }
      }
    }

    if (MODE == MODE_MV) {
      if (tt == 0) {
#pragma unroll
        for (int rr = 0; rr < RG; rr++)
          if (r0 + rr < n) a.out[r0 + rr] = t[rr];
      }
      if (XSMEM) mbar_arrive(&empty[s]);
      continue;
    }
    if (MODE == MODE_CHAIN) {
      if (a.ct == SYSML_CHAIN_XtwXv) {
#pragma unroll
        for (int rr = 0; rr < RG; rr++) t[rr] *= wv[rr];
      } else if (a.ct == SYSML_CHAIN_XtXvy) {
#pragma unroll
        for (int rr = 0; rr < RG; rr++) t[rr] -= wv[rr];
      }
    } else if (MODE == MODE_FN) {
      // the row function, ONCE per row: thread rr of the group evaluates row rr (it sits on the tile's critical path), the group picks the
      // results up from shared memory after a barrier of its own
      double* fb = fnbuf + ((size_t)(it & 1) * G + g) * RG;
      if (tt < RG) {  // threads 0 .. RG-1 of the group: one row each, side by side (one pass through the interpreter)
        double e_in = t[0], w_in = wv[0];
#pragma unroll
        for (int rr = 1; rr < RG; rr++) {
          e_in = (tt == rr) ? t[rr] : e_in;
          w_in = (tt == rr) ? wv[rr] : w_in;
        }
        const bool live = r0 + tt < n;
        double res0, res1;
        sysml_rt::row_eval(a.fn, e_in, w_in, (a.y2 && live) ? __ldg(a.y2 + r0 + tt) : 0.0, res0, res1);
        fb[tt] = live ? res0 : 0.0;  // rows past the end are zero-filled in the tile: they must add nothing
        if (a.side && live) a.side[r0 + tt] = res1;
      }
      named_bar_sync(1 + g, TC);
#pragma unroll
      for (int rr = 0; rr < RG; rr++) t[rr] = fb[rr];
    } else {
#pragma unroll
      for (int rr = 0; rr < RG; rr++) t[rr] = wv[rr];
    }
    if (XSMEM) {
#pragma unroll
      for (int k = 0; k < CPT; k++) {
        const int j = tt + TC * k;
        if (j < m) {
#pragma unroll
          for (int rr = 0; rr < RG; rr++) c[k] = fma(t[rr], T[rr * m + j], c[k]);
        }
      }
      mbar_arrive(&empty[s]);
    } else {
#pragma unroll
      for (int k = 0; k < CPT; k++) {
#pragma unroll
        for (int rr = 0; rr < RG; rr++) c[k] = fma(t[rr], x[rr][k], c[k]);
      }
    }
  }

  if (MODE == MODE_MV) return;

  // ---- per-CTA column partials -> scratch; colpartial_reduce_kernel adds them up in a fixed order ----
  {
    double* prow = a.partial + ((size_t)blockIdx.x * G + g) * m;
#pragma unroll
    for (int k = 0; k < CPT; k++) {
      const int j = tt + TC * k;
      if (j < m) prow[j] = c[k];
    }
  }
}

// out[j] = sum over the P partial rows of partial[p][j], compensated (two-sum) and in a fixed order.
// 32 columns x 8 row-slices per CTA: 8 x (P/8) loads deep instead of one thread walking all P rows, so the
// reduction costs a few microseconds instead of P dependent L2 round trips inside the streaming kernel.
__global__ void __launch_bounds__(256) colpartial_reduce_kernel(const double* __restrict__ partial, int P, int m, double* __restrict__ out) {
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
  if (sl == 0 && j < m) {
#pragma unroll
    for (int q = 1; q < 8; q++) {
      CompSum t;
      t.hi = sh_hi[q][c];
      t.lo = sh_lo[q][c];
      a0.add(t);
    }
    out[j] = a0.value();
  }
}

template <int MODE, int NTC, int TC, int CPT, int RG, int STAGES, bool XSMEM = false>
int launch_cfg(sysml_b200_ctx* ctx, StreamArgs a, int ctas_per_sm) {
  constexpr int G = NTC / TC;
  constexpr int R = G * RG;
  constexpr int NWG = TC / 32;
  auto kern = rowstream_kernel<MODE, NTC, TC, CPT, RG, STAGES, XSMEM>;
  const size_t smem = 128 + 2 * G * NWG * RG * 8 + (MODE == MODE_FN ? 2 * G * RG * 8 : 0) + (size_t)STAGES * R * a.m * 8 + (size_t)STAGES * R * 8;
  if (smem > (size_t)ctx->max_smem_optin) return sysml_set_error(ctx, SYSML_E_UNSUPPORTED, "rowstream: smem %zu too large", smem);
  SYSML_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  a.total_tiles = (a.n + R - 1) / R;
  const size_t stage_bytes = (size_t)R * a.m * 8;
  a.use_tma = ((reinterpret_cast<uintptr_t>(a.X) & 15) == 0) && (stage_bytes % 16 == 0) && (stage_bytes < (1u << 20));
  const bool need_w = MODE == MODE_VM || (MODE == MODE_CHAIN && a.ct != SYSML_CHAIN_XtXv) || (MODE == MODE_FN && a.w != nullptr);
  // narrow rows only: there a tile is consumed in a few hundred cycles and a global load of w sits on its critical
  // path (10M x 100 XtwXv: 1.73 -> 1.18 ms); with 8 KB rows the consumers' own loads are hidden and the extra
  // 32-byte copy per tile costs more than it saves (10M x 1000: 11.2 -> 11.8 ms)
  bool w_ring = a.m <= 256;
  if (const char* e = getenv("SYSML_B200_MMCHAIN_W_RING")) w_ring = atoi(e) != 0;  // A/B switch for measurements
  a.w_smem = (need_w && a.w && w_ring && (reinterpret_cast<uintptr_t>(a.w) & 15) == 0 && (R % 2 == 0)) ? 1 : 0;
  i64 grid = (i64)ctx->sm_count * ctas_per_sm;
  if (grid > a.total_tiles) grid = a.total_tiles;
  if (grid < 1) grid = 1;
  if (MODE != MODE_MV) {
    void* ws = nullptr;
    int rc = sysml_workspace(ctx, (size_t)grid * G * a.m * sizeof(double), &ws);
    if (rc) return rc;
    a.partial = (double*)ws;
  }
  {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(NTC + 32);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = ctx->stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = a.pdl ? 1 : 0;
    SYSML_CUDA(ctx, cudaLaunchKernelEx(&cfg, kern, a));
    ctx->launches++;
  }
  if (MODE != MODE_MV) {
    if (a.comm || a.cg_r)
      return sysml_launch_finish_exchange(ctx, a.comm, a.partial, (int)(grid * G), a.m, a.out, SYSML_REDUCE_SUM, a.pdl != 0,
                                          a.cg_r ? const_cast<double*>(a.v) : nullptr, a.cg_r, a.cg_beta, a.cg_nr2, a.cg_lambda);
    colpartial_reduce_kernel<<<(unsigned)((a.m + 31) / 32), 256, 0, ctx->stream>>>(a.partial, (int)(grid * G), a.m, a.out);
    SYSML_LAUNCH_CHECK(ctx, "colpartial_reduce_kernel");
  }
  return SYSML_OK;
}

// Shape -> configuration.  Tile bytes are kept near 32 KB (64 KB for wide rows) so that, with a
// 3-deep ring and 2 CTAs per SM (1 for wide rows), 128-190 KB per SM are in flight.
template <int MODE>
int launch_rowstream(sysml_b200_ctx* ctx, const StreamArgs& a) {
  const int m = a.m;
  if (m <= 64) return launch_cfg<MODE, 128, 32, 2, 8, 4>(ctx, a, 2);
  if (m <= 128) return launch_cfg<MODE, 128, 32, 4, 8, 3>(ctx, a, 2);
  if (m <= 256) return launch_cfg<MODE, 128, 32, 8, 4, 3>(ctx, a, 2);
  if (m <= 512) return launch_cfg<MODE, 128, 64, 8, 4, 3>(ctx, a, 2);
  if (m <= 1024) return launch_cfg<MODE, 128, 128, 8, 4, 3>(ctx, a, 2);
  if (m <= 2048) return launch_cfg<MODE, 256, 256, 8, 4, 3>(ctx, a, 1);
  if (m <= 4096) return launch_cfg<MODE, 256, 256, 16, 2, 3>(ctx, a, 1);
  // wide rows: the tile stays in shared memory for both phases, one row per tile, 4 / 3 stages
  if (m <= 5600) return launch_cfg<MODE, 256, 256, 22, 1, 4, true>(ctx, a, 1);
  if (m <= 8192) return launch_cfg<MODE, 256, 256, 32, 1, 3, true>(ctx, a, 1);
  return SYSML_E_UNSUPPORTED;
}

// ------------------------------------------------------------------------------------------------
// Wide-row fallbacks (above 8192 columns; mmchain then reads X twice: MV then VM).
// ------------------------------------------------------------------------------------------------
// one warp per row, grid-stride over rows
__global__ void mv_wide_kernel(const double* __restrict__ X, const double* __restrict__ v, double* __restrict__ out,
                               i64 n, i64 m) {
  const int lane = threadIdx.x & 31;
  const i64 warp = ((i64)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const i64 nwarps = ((i64)gridDim.x * blockDim.x) >> 5;
  for (i64 i = warp; i < n; i += nwarps) {
    const double* row = X + i * m;
    double p0 = 0.0, p1 = 0.0;
    i64 j = lane;
    for (; j + 32 < m; j += 64) {
      p0 = fma(ldg_stream(row + j), __ldg(v + j), p0);
      p1 = fma(ldg_stream(row + j + 32), __ldg(v + j + 32), p1);
    }
    if (j < m) p0 = fma(ldg_stream(row + j), __ldg(v + j), p0);
    double p = warp_sum(p0 + p1);
    if (lane == 0) out[i] = p;
  }
}

// out[j] = sum_i y[i] * X[i,j].  grid.x tiles columns (blockDim.x each), grid.y splits rows;
// row-split partials are summed (compensated, fixed order) by vm_wide_reduce_kernel.
__global__ void vm_wide_kernel(const double* __restrict__ X, const double* __restrict__ y,
                               double* __restrict__ partial, i64 n, i64 m) {
  const i64 j = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= m) return;
  const i64 rows_per = (n + gridDim.y - 1) / gridDim.y;
  const i64 rl = (i64)blockIdx.y * rows_per;
  const i64 ru = (rl + rows_per < n) ? rl + rows_per : n;
  double c0 = 0.0, c1 = 0.0, c2 = 0.0, c3 = 0.0;
  i64 i = rl;
  for (; i + 3 < ru; i += 4) {
    c0 = fma(__ldg(y + i), ldg_stream(X + i * m + j), c0);
    c1 = fma(__ldg(y + i + 1), ldg_stream(X + (i + 1) * m + j), c1);
    c2 = fma(__ldg(y + i + 2), ldg_stream(X + (i + 2) * m + j), c2);
    c3 = fma(__ldg(y + i + 3), ldg_stream(X + (i + 3) * m + j), c3);
  }
  for (; i < ru; i++) c0 = fma(__ldg(y + i), ldg_stream(X + i * m + j), c0);
  partial[(size_t)blockIdx.y * m + j] = (c0 + c1) + (c2 + c3);
}

__global__ void vm_wide_reduce_kernel(const double* __restrict__ partial, double* __restrict__ out, int parts, i64 m) {
  const i64 j = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= m) return;
  CompSum acc;
  for (int b = 0; b < parts; b++) acc.add(partial[(size_t)b * m + j]);
  out[j] = acc.value();
}

__global__ void apply_weights_kernel(double* __restrict__ t, const double* __restrict__ w, i64 n, int ct) {
  const i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (ct == SYSML_CHAIN_XtwXv)
    t[i] *= w[i];
  else if (ct == SYSML_CHAIN_XtXvy)
    t[i] -= w[i];
}

int mv_wide(sysml_b200_ctx* ctx, const double* X, const double* v, double* out, i64 n, i64 m) {
  const int threads = 256;
  i64 blocks = (n + 7) / 8;
  const i64 maxb = (i64)ctx->sm_count * 8;
  if (blocks > maxb) blocks = maxb;
  if (blocks < 1) blocks = 1;
  mv_wide_kernel<<<(unsigned)blocks, threads, 0, ctx->stream>>>(X, v, out, n, m);
  SYSML_LAUNCH_CHECK(ctx, "mv_wide_kernel");
  return SYSML_OK;
}

int vm_wide(sysml_b200_ctx* ctx, const double* X, const double* y, double* out, i64 n, i64 m, size_t ws_offset_bytes) {
  const int threads = 128;
  const i64 ctiles = (m + threads - 1) / threads;
  i64 splits = ((i64)ctx->sm_count * 8 + ctiles - 1) / ctiles;
  if (splits > (n + 63) / 64) splits = (n + 63) / 64;
  if (splits < 1) splits = 1;
  if (splits > 65535) splits = 65535;
  void* ws = nullptr;
  int rc = sysml_workspace(ctx, ws_offset_bytes + (size_t)splits * m * sizeof(double), &ws);
  if (rc) return rc;
  double* partial = (double*)((char*)ws + ws_offset_bytes);
  dim3 grid((unsigned)ctiles, (unsigned)splits);
  vm_wide_kernel<<<grid, threads, 0, ctx->stream>>>(X, y, partial, n, m);
  SYSML_LAUNCH_CHECK(ctx, "vm_wide_kernel");
  if (m <= 0x7fffffffLL) {
    colpartial_reduce_kernel<<<(unsigned)((m + 31) / 32), 256, 0, ctx->stream>>>(partial, (int)splits, (int)m, out);
  } else {
    vm_wide_reduce_kernel<<<(unsigned)ctiles, threads, 0, ctx->stream>>>(partial, out, (int)splits, m);
  }
  SYSML_LAUNCH_CHECK(ctx, "vm_wide reduce");
  return SYSML_OK;
}

}  // namespace

// sums P partial rows of length m (fixed order, compensated); used by sparse.cu for its replicated outputs
int sysml_reduce_partial_rows(sysml_b200_ctx* ctx, const double* partial, int P, int m, double* out) {
  colpartial_reduce_kernel<<<(unsigned)((m + 31) / 32), 256, 0, ctx->stream>>>(partial, P, m, out);
  SYSML_LAUNCH_CHECK(ctx, "colpartial_reduce_kernel");
  return SYSML_OK;
}

// internal entry points used by matmult.cu as well
int sysml_dense_mv(sysml_b200_ctx* ctx, const double* X, const double* v, double* out, i64 n, i64 m) {
  if (m <= 8192) {
    StreamArgs a{};
    a.X = X; a.v = v; a.w = nullptr; a.out = out; a.n = n; a.m = (int)m; a.ct = 0;
    return launch_rowstream<MODE_MV>(ctx, a);
  }
  return mv_wide(ctx, X, v, out, n, m);
}

int sysml_dense_vm(sysml_b200_ctx* ctx, const double* X, const double* y, double* out, i64 n, i64 m) {
  if (m <= 8192) {
    StreamArgs a{};
    a.X = X; a.v = nullptr; a.w = y; a.out = out; a.n = n; a.m = (int)m; a.ct = 0;
    return launch_rowstream<MODE_VM>(ctx, a);
  }
  return vm_wide(ctx, X, y, out, n, m, 0);
}

static int validate_chain(sysml_b200_ctx* ctx, const void* X, const void* v, const void* w, const void* out, i64 n, i64 m, int ct) {
  if (ct < 0 || ct > 2) return sysml_set_error(ctx, SYSML_E_INVALID, "mmchain: unknown chain type %d", ct);
  if (n < 0 || m < 0) return sysml_set_error(ctx, SYSML_E_DIM, "mmchain: negative dimension (%lld x %lld)", n, m);
  if (m > 0 && !out) return sysml_set_error(ctx, SYSML_E_INVALID, "mmchain: out is null");
  if (n > 0 && m > 0 && (!X || !v)) return sysml_set_error(ctx, SYSML_E_INVALID, "mmchain: X or v is null");
  if (ct != SYSML_CHAIN_XtXv && n > 0 && !w) return sysml_set_error(ctx, SYSML_E_INVALID, "mmchain: chain type %d needs w", ct);
  if (m > 0x7fffffffLL) return sysml_set_error(ctx, SYSML_E_LIMIT, "mmchain: more than 2^31-1 columns");
  return SYSML_OK;
}

extern "C" int sysml_b200_mmchain_dense(sysml_b200_ctx* ctx, const double* X, const double* v, const double* w,
                                        double* out, int64_t n, int64_t m, int chain_type) {
  SYSML_ENTER(ctx);
  int rc = validate_chain(ctx, X, v, w, out, n, m, chain_type);
  if (rc) return rc;
  if (m == 0) return SYSML_OK;
  if (n == 0) {  // empty X: result is the zero vector (LibMatrixMult.java:245-248)
    SYSML_CUDA(ctx, cudaMemsetAsync(out, 0, (size_t)m * sizeof(double), ctx->stream));
    return SYSML_OK;
  }
  if (m <= 8192) {
    StreamArgs a{};
    a.X = X; a.v = v; a.w = w; a.out = out; a.n = n; a.m = (int)m; a.ct = chain_type;
    return launch_rowstream<MODE_CHAIN>(ctx, a);
  }
  // very wide rows (> 8192 columns): t = X v ; t = w*t | t-w ; out = t(X) t   (X read twice)
  void* ws = nullptr;
  const size_t tbytes = (((size_t)n * sizeof(double)) + 255) & ~(size_t)255;
  // reserve the partial area up front so the second workspace call cannot move the block
  {
    const i64 ctiles = (m + 127) / 128;
    i64 splits = ((i64)ctx->sm_count * 8 + ctiles - 1) / ctiles;
    if (splits > (n + 63) / 64) splits = (n + 63) / 64;
    if (splits < 1) splits = 1;
    rc = sysml_workspace(ctx, tbytes + (size_t)splits * m * sizeof(double), &ws);
    if (rc) return rc;
  }
  double* t = (double*)ws;
  rc = mv_wide(ctx, X, v, t, n, m);
  if (rc) return rc;
  if (chain_type != SYSML_CHAIN_XtXv) {
    apply_weights_kernel<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(t, w, n, chain_type);
    SYSML_LAUNCH_CHECK(ctx, "apply_weights_kernel");
  }
  return vm_wide(ctx, X, t, out, n, m, tbytes);
}

// mmchain on this rank's row shard + sum over all ranks of the communicator, in one finishing kernel
// (no NCCL call).  Every rank must make the call; out (m doubles) ends bit-identical on all of them.
extern "C" int sysml_b200_mmchain_dense_allreduce(sysml_b200_ctx* ctx, sysml_b200_comm* comm, const double* X, const double* v,
                                                  const double* w, double* out, int64_t n, int64_t m, int chain_type) {
  SYSML_ENTER(ctx);
  int rc = validate_chain(ctx, X, v, w, out, n, m, chain_type);
  if (rc) return rc;
  if (!comm) return sysml_set_error(ctx, SYSML_E_INVALID, "mmchain_allreduce: null communicator");
  if (m == 0) return SYSML_OK;
  if (m > 8192) return sysml_set_error(ctx, SYSML_E_UNSUPPORTED, "mmchain_allreduce: more than 8192 columns; run mmchain_dense then allreduce_sum");
  if (n == 0) {  // an empty shard still has to take part in the exchange
    SYSML_CUDA(ctx, cudaMemsetAsync(out, 0, (size_t)m * sizeof(double), ctx->stream));
    return sysml_launch_finish_exchange(ctx, comm, out, 1, (int)m, out, SYSML_REDUCE_SUM, false, nullptr, nullptr, nullptr, nullptr, 0.0);
  }
  StreamArgs a{};
  a.X = X; a.v = v; a.w = w; a.out = out; a.n = n; a.m = (int)m; a.ct = chain_type; a.comm = comm;
  a.pdl = 1;
  return launch_rowstream<MODE_CHAIN>(ctx, a);
}

// out = t(X) %*% f(X %*% v, w, y) with the row function f given as a postfix program, X read ONCE; a second value left by
// the program is written to side_out (one per row).  The chain types of mmchain_dense are the programs "in0", "in0 * in1",
// "in0 - in1"; this is the same kernel with the row function interpreted once per row (cellwise_rt.cuh).
extern "C" int sysml_b200_mmchain_rowfn_dense(sysml_b200_ctx* ctx, const double* X, const double* v, const double* w, const double* y,
                                              const int32_t* program, int n_tokens, const double* consts, int n_consts, int n_results,
                                              double* side_out, double* out, int64_t n, int64_t m) {
  SYSML_ENTER(ctx);
  if (n < 0 || m < 0) return sysml_set_error(ctx, SYSML_E_DIM, "mmchain_rowfn: negative dimension (%lld x %lld)", (long long)n, (long long)m);
  if (!program || (n_consts > 0 && !consts)) return sysml_set_error(ctx, SYSML_E_INVALID, "mmchain_rowfn: null program / constants");
  if (const char* why = sysml_rt::validate(program, n_tokens, 3, n_consts, n_results)) return sysml_set_error(ctx, SYSML_E_INVALID, "mmchain_rowfn: %s", why);
  for (int t = 0; t < n_tokens; t++) {
    const int kind = (program[t] >> 24) & 0xff, arg = program[t] & 0xffffff;
    if (kind == SYSML_FUSED_IN && n > 0 && ((arg == 1 && !w) || (arg == 2 && !y)))
      return sysml_set_error(ctx, SYSML_E_INVALID, "mmchain_rowfn: the program reads input %d, which is null", arg);
  }
  if (n_results == 2 && n > 0 && !side_out) return sysml_set_error(ctx, SYSML_E_INVALID, "mmchain_rowfn: two results but side_out is null");
  if (m > 0 && !out) return sysml_set_error(ctx, SYSML_E_INVALID, "mmchain_rowfn: out is null");
  if (m == 0) return SYSML_OK;
  if (n == 0) {
    SYSML_CUDA(ctx, cudaMemsetAsync(out, 0, (size_t)m * sizeof(double), ctx->stream));
    return SYSML_OK;
  }
  if (!X || !v) return sysml_set_error(ctx, SYSML_E_INVALID, "mmchain_rowfn: X or v is null");
  if (m > 8192)
    return sysml_set_error(ctx, SYSML_E_UNSUPPORTED, "mmchain_rowfn: more than 8192 columns: a row no longer fits on chip between the two products; run matmult, the cellwise chain and matmult");
  StreamArgs a{};
  a.X = X; a.v = v; a.w = w; a.out = out; a.n = n; a.m = (int)m; a.ct = 0; a.comm = nullptr;
  a.pdl = 0;
  a.y2 = y;
  a.side = n_results == 2 ? side_out : nullptr;
  a.fn.n_tokens = n_tokens;
  a.fn.n_results = n_results;
  for (int t = 0; t < n_tokens; t++) a.fn.tok[t] = program[t];
  for (int c = 0; c < n_consts; c++) a.fn.k[c] = consts[c];
  return launch_rowstream<MODE_FN>(ctx, a);
}

// One conjugate-gradient iteration of LinearRegCG / GLM on this rank's row shard in TWO launches:
//   q = t(X) %*% (w * (X %*% p))   [summed over the ranks of comm when it is not null]
//   q += lambda*p ; a = nr2/sum(p*q) ; beta += a*p ; r += a*q ; nr2' = sum(r^2) ; p = -r + (nr2'/nr2)*p ; nr2 = nr2'
// The streaming kernel is followed by one finishing kernel that folds the per-CTA partials, exchanges them with the
// peers and, in its last CTA, applies the m-vector tail (cg_tail.cuh; same arithmetic as sysml_b200_cg_vector_update).
// Both kernels are programmatic dependent launches: their set-up overlaps the tail of the kernel before them.
extern "C" int sysml_b200_cg_step_dense(sysml_b200_ctx* ctx, sysml_b200_comm* comm, const double* X, const double* w, double* p,
                                        double* q, double* r, double* beta, double* nr2, int64_t n, int64_t m, int chain_type,
                                        double lambda) {
  SYSML_ENTER(ctx);
  int rc = validate_chain(ctx, X, p, w, q, n, m, chain_type);
  if (rc) return rc;
  if (m <= 0 || !p || !q || !r || !beta || !nr2) return sysml_set_error(ctx, SYSML_E_INVALID, "cg_step: bad arguments");
  if (m > 8192) return sysml_set_error(ctx, SYSML_E_UNSUPPORTED, "cg_step: more than 8192 columns; run mmchain_dense (+ allreduce_sum) and cg_vector_update");
  if (n == 0) {
    SYSML_CUDA(ctx, cudaMemsetAsync(q, 0, (size_t)m * sizeof(double), ctx->stream));
    return sysml_launch_finish_exchange(ctx, comm, q, 1, (int)m, q, SYSML_REDUCE_SUM, false, p, r, beta, nr2, lambda);
  }
  StreamArgs a{};
  a.X = X; a.v = p; a.w = w; a.out = q; a.n = n; a.m = (int)m; a.ct = chain_type; a.comm = comm;
  a.pdl = 1;
  a.cg_r = r; a.cg_beta = beta; a.cg_nr2 = nr2; a.cg_lambda = lambda;
  return launch_rowstream<MODE_CHAIN>(ctx, a);
}

extern "C" int sysml_b200_mmchain_dense_hostvec(sysml_b200_ctx* ctx, const double* X, const double* h_v,
                                                const double* h_w, const double* d_w, double* h_out, int64_t n,
                                                int64_t m, int chain_type) {
  SYSML_ENTER(ctx);
  if (!h_v || !h_out) return sysml_set_error(ctx, SYSML_E_INVALID, "mmchain_hostvec: h_v or h_out is null");
  if (n < 0 || m < 0) return sysml_set_error(ctx, SYSML_E_DIM, "mmchain_hostvec: negative dimension");
  const bool need_w = chain_type != SYSML_CHAIN_XtXv;
  if (need_w && !h_w && !d_w) return sysml_set_error(ctx, SYSML_E_INVALID, "mmchain_hostvec: chain type %d needs w", chain_type);
  const bool copy_w = need_w && !d_w;
  // small device staging vectors owned by the context (v, out, and w when it comes from the host)
  const size_t mpad = (size_t)((m + 31) / 32) * 32;
  const size_t need = (2 * mpad + (copy_w ? (size_t)n : 0)) * sizeof(double);
  if (ctx->vec_stage_bytes < need) {
    SYSML_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (ctx->vec_stage) SYSML_CUDA(ctx, cudaFree(ctx->vec_stage));
    ctx->vec_stage = nullptr;
    ctx->vec_stage_bytes = 0;
    SYSML_CUDA(ctx, cudaMalloc(&ctx->vec_stage, need));
    ctx->vec_stage_bytes = need;
  }
  double* s_dev = (double*)ctx->vec_stage;
  double* d_v = s_dev;
  double* d_out = s_dev + mpad;
  double* d_wbuf = d_out + mpad;
  void* pin = nullptr;
  int rc = sysml_pinned(ctx, (size_t)(2 * m) * sizeof(double), &pin);
  if (rc) return rc;
  double* p_v = (double*)pin;
  double* p_out = p_v + m;
  memcpy(p_v, h_v, (size_t)m * sizeof(double));
  SYSML_CUDA(ctx, cudaMemcpyAsync(d_v, p_v, (size_t)m * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
  if (copy_w) SYSML_CUDA(ctx, cudaMemcpyAsync(d_wbuf, h_w, (size_t)n * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
  rc = sysml_b200_mmchain_dense(ctx, X, d_v, need_w ? (copy_w ? d_wbuf : d_w) : nullptr, d_out, n, m, chain_type);
  if (rc) return rc;
  SYSML_CUDA(ctx, cudaMemcpyAsync(p_out, d_out, (size_t)m * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
  SYSML_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  memcpy(h_out, p_out, (size_t)m * sizeof(double));
  return SYSML_OK;
}

// ================================================================================================
// Allocation-free mmchain.
//
// sysml_b200_mmchain_dense does its cross-CTA reduction through scratch that the context obtains with
// cudaMalloc the first time it is needed (and again whenever a larger request arrives).  Some callers
// cannot tolerate that.  The typical one is a host runtime that has handed the whole device to its own
// memory manager -- SystemML's GPUMemoryManager budgets the card up to sysml.gpu.memory.util.factor and
// evicts matrices according to its own books -- so a cudaMalloc issued behind its back can fail on a
// "full" device, or silently break the accounting it relies on to decide what to evict.  The same goes
// for callers that run inside a region where the CUDA runtime must not be asked for memory at all
// (cudaMalloc synchronises the whole device and is refused while a stream is being captured).
//
// For them this entry point keeps its scratch in a fixed array that is part of the library's device code:
// once that code is resident on a device the memory stays until the process exits, and nothing is
// allocated or freed by a call.  "Once resident" matters: CUDA loads device code lazily by default
// (CUDA_MODULE_LOADING=LAZY since CUDA 12), so the array -- and the kernels -- are placed on a device by the
// FIRST call there (the symbol lookup below triggers it).  That first call therefore does take device memory
// (about 2.4 MiB plus code) and can fail on a card that is already full: make one small warm-up call per
// device while memory is still plentiful (right after creating the context), or run the process with
// CUDA_MODULE_LOADING=EAGER, in which case loading the library places everything at once.  Every later
// call leaves the amount of free device memory exactly as it found it (tests/test_zz_no_alloc.py checks).
//   - limit: cols <= SYSML_B200_NOALLOC_MAX_COLS (1024); larger inputs are rejected with SYSML_E_LIMIT;
//   - footprint: NOALLOC_DOUBLES*8 + 256 bytes (about 2.4 MiB) of device memory per GPU, permanently;
//   - the array is shared by every call on that GPU and keeps whatever the previous call left in it; a
//     call overwrites every partial-sum row it later reads, so nothing in it needs to be cleared first;
//   - two calls that overlap on the same GPU (different streams / contexts / host threads) race on the
//     array and produce garbage: serialise them;
//   - the very first call on each device looks up the device address of the array
//     (cudaGetSymbolAddress); make that first call before any stream capture begins.
// ================================================================================================
namespace {

constexpr int NOALLOC_MAX_COLS = SYSML_B200_NOALLOC_MAX_COLS;
constexpr size_t NOALLOC_DOUBLES = (size_t)304 * 1024;  // (CTAs x row-groups) x cols partial sums
// [0, NOALLOC_DOUBLES): per-CTA column partials (the 256 bytes after them are padding)
static __device__ __align__(256) double g_noalloc_scratch[NOALLOC_DOUBLES + 32];

constexpr int NOALLOC_MAX_DEVICES = 64;
double* g_noalloc_ptr[NOALLOC_MAX_DEVICES] = {nullptr};  // device address of the array, per device

// `g_noalloc_scratch` named in host code is only a host-side handle for the symbol, not a device
// address: the address a kernel can use has to be asked for, once per device, with that device current.
int noalloc_scratch(sysml_b200_ctx* ctx, double** out) {
  const int dev = ctx->device;
  if (dev < 0 || dev >= NOALLOC_MAX_DEVICES) return sysml_set_error(ctx, SYSML_E_LIMIT, "mmchain_noalloc: device index %d not supported", dev);
  if (!g_noalloc_ptr[dev]) {
    double* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void**)&p, g_noalloc_scratch);
    if (e != cudaSuccess || !p)
      return sysml_set_error(ctx, SYSML_E_CUDA, "mmchain_noalloc: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
    g_noalloc_ptr[dev] = p;
  }
  *out = g_noalloc_ptr[dev];
  return SYSML_OK;
}

template <int NTC, int TC, int CPT, int RG, int STAGES>
int launch_cfg_noalloc(sysml_b200_ctx* ctx, StreamArgs a, int ctas_per_sm, double* scratch) {
  constexpr int G = NTC / TC;
  constexpr int R = G * RG;
  constexpr int NWG = TC / 32;
  auto kern = rowstream_kernel<MODE_CHAIN, NTC, TC, CPT, RG, STAGES>;
  const size_t smem = 128 + 2 * G * NWG * RG * 8 + (size_t)STAGES * R * a.m * 8;
  if (smem > (size_t)ctx->max_smem_optin) return sysml_set_error(ctx, SYSML_E_UNSUPPORTED, "mmchain_noalloc: smem %zu too large", smem);
  SYSML_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  a.total_tiles = (a.n + R - 1) / R;
  const size_t stage_bytes = (size_t)R * a.m * 8;
  a.use_tma = ((reinterpret_cast<uintptr_t>(a.X) & 15) == 0) && (stage_bytes % 16 == 0) && (stage_bytes < (1u << 20));
  i64 grid = (i64)ctx->sm_count * ctas_per_sm;
  const i64 fit = (i64)(NOALLOC_DOUBLES / ((size_t)G * a.m));  // CTAs whose partials fit in the fixed array
  if (grid > fit) grid = fit;
  if (grid > a.total_tiles) grid = a.total_tiles;
  if (grid < 1) grid = 1;
  a.partial = scratch;
  kern<<<(unsigned)grid, NTC + 32, smem, ctx->stream>>>(a);
  SYSML_LAUNCH_CHECK(ctx, "rowstream_kernel (noalloc)");
  // every partial row that is read below was written by the launch above: nothing has to start at zero
  colpartial_reduce_kernel<<<(unsigned)((a.m + 31) / 32), 256, 0, ctx->stream>>>(a.partial, (int)(grid * G), a.m, a.out);
  SYSML_LAUNCH_CHECK(ctx, "colpartial_reduce_kernel (noalloc)");
  return SYSML_OK;
}

}  // namespace

extern "C" int sysml_b200_mmchain_dense_noalloc(sysml_b200_ctx* ctx, const double* X, const double* v, const double* w,
                                                double* out, int64_t n, int64_t m, int chain_type) {
  SYSML_ENTER(ctx);
  int rc = validate_chain(ctx, X, v, w, out, n, m, chain_type);
  if (rc) return rc;
  if (m > NOALLOC_MAX_COLS)
    return sysml_set_error(ctx, SYSML_E_LIMIT,
                           "mmchain_noalloc: %lld columns exceed the fixed-scratch limit of %d; use sysml_b200_mmchain_dense",
                           (long long)m, NOALLOC_MAX_COLS);
  if (m == 0) return SYSML_OK;
  if (n == 0) {
    SYSML_CUDA(ctx, cudaMemsetAsync(out, 0, (size_t)m * sizeof(double), ctx->stream));
    return SYSML_OK;
  }
  double* scratch = nullptr;
  rc = noalloc_scratch(ctx, &scratch);
  if (rc) return rc;
  StreamArgs a{};
  a.X = X; a.v = v; a.w = w; a.out = out; a.n = n; a.m = (int)m; a.ct = chain_type;
  if (m <= 64) return launch_cfg_noalloc<128, 32, 2, 8, 4>(ctx, a, 2, scratch);
  if (m <= 128) return launch_cfg_noalloc<128, 32, 4, 8, 3>(ctx, a, 2, scratch);
  if (m <= 256) return launch_cfg_noalloc<128, 32, 8, 4, 3>(ctx, a, 2, scratch);
  if (m <= 512) return launch_cfg_noalloc<128, 64, 8, 4, 3>(ctx, a, 2, scratch);
  return launch_cfg_noalloc<128, 128, 8, 4, 3>(ctx, a, 2, scratch);
}
