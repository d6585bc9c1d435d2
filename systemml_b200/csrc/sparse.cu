// sparse.cu -- CSR kernels: SpMV, SpMM (CSR x dense), t(CSR) x dense, sparse mmchain, sparse tsmm,
// sparse aggregates, dense<->CSR conversion and nnz counting.
//
// CSR layout = CSRPointer.java:76-100: double val[nnz], int32 rowptr[rows+1], int32 colind[nnz], 0-based.
//
// Reference semantics: LibMatrixMult.matrixMultSparseDense (LibMatrixMult.java:1170-1395),
// matrixMultChainSparse (:1655-1684), matrixMultTransposeSelfSparse (:1839-1960);
// LibMatrixAgg s_ua* (LibMatrixAgg.java:2137-2195, 2426-2614); GPU twins cusparseDcsrmv / Dcsrmm2
// (LibMatrixCuMatMult.java:313-341), cusparseDnnz / Ddense2csr / Dcsr2dense (GPUObject.java:199-245,508).
//
// SpMV: L lanes cooperate on one row, L chosen from mean nnz/row (2..32) -- "vector-per-row" for long
// rows, "sub-warp per row" for short ones; consecutive rows sit in consecutive lane groups so val/colind
// reads of a warp are contiguous.  HBM-bound: 12*nnz + 4*(rows+1) + 8*cols + 8*rows bytes.
// Scatter-type outputs (t(X)%*%y, sparse mmchain, sparse tsmm, column sums) use fp64 RED atomics into
// an L2-resident output; their summation order is not fixed (run-to-run differences ~1 ulp).
#include <math.h>
#include <math_constants.h>

#include "common.cuh"

int sysml_reduce_partial_rows(sysml_b200_ctx* ctx, const double* partial, int P, int m, double* out);

namespace {

__device__ __forceinline__ double jmax(double a, double b) {
  if (a != a) return a;
  if (b != b) return b;
  if (a == 0.0 && b == 0.0) return __longlong_as_double(__double_as_longlong(a) & __double_as_longlong(b));
  return (a >= b) ? a : b;
}
__device__ __forceinline__ double jmin(double a, double b) {
  if (a != a) return a;
  if (b != b) return b;
  if (a == 0.0 && b == 0.0) return __longlong_as_double(__double_as_longlong(a) | __double_as_longlong(b));
  return (a <= b) ? a : b;
}
template <int L>
__device__ __forceinline__ double group_sum(double v) {
#pragma unroll
  for (int o = L >> 1; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

inline int pick_lanes(i64 nnz, i64 rows) {
  const double mean = rows > 0 ? (double)nnz / (double)rows : 0.0;
  int L = 2;  // largest power of two <= mean (at least 2): short rows share a warp, few lanes idle
  while (L < 32 && 2 * L <= mean) L <<= 1;
  return L;
}
inline unsigned rows_grid(sysml_b200_ctx* ctx, i64 rows, int L, int threads) {
  const i64 rows_per_block = (threads / 32) * (32 / L);  // per pass; kernels that unroll passes just loop less
  i64 blocks = (rows + rows_per_block - 1) / rows_per_block;
  const i64 cap = (i64)ctx->sm_count * 16;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return (unsigned)blocks;
}

// MODE 0: out[r] = A[r,:] . v                      (SpMV)
// MODE 1: c[col] += (w-adjusted A[r,:].v) * A[r,col] (sparse mmchain), ct as in SYSML_CHAIN_*
// MODE 2: c[col] += y[r] * A[r,col]                 (t(A) %*% y)
// MODE 3 (streamed kernel only): out[r] = sum / sum of squares / max / min of row r (ct = SYSML_AGG_*, implicit zeros
//         join max/min when the row has fewer than out_len = cols entries); colind is never read
template <int L, int MODE>
__global__ void __launch_bounds__(256, 8) csr_row_kernel(const int* __restrict__ rp, const int* __restrict__ ci,
                                                      const double* __restrict__ va, const double* __restrict__ v,
                                                      const double* __restrict__ w, double* __restrict__ out, i64 rows, int ct,
                                                      int replicas, i64 out_len) {
  // scatter modes: the L2 serialises atomics that hit the same address, and the output vector is small and
  // hot -- each CTA therefore adds into one of `replicas` copies of it (summed afterwards in a fixed order)
  if (MODE != 0 && replicas > 1) out += (i64)(blockIdx.x % replicas) * out_len;
  const int lane = threadIdx.x & 31;
  const int sub = lane & (L - 1);
  constexpr int RPW = 32 / L;
  const i64 warp = ((i64)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const i64 nwarps = ((i64)gridDim.x * blockDim.x) >> 5;
  for (i64 base = warp * RPW; base < rows; base += nwarps * RPW) {
    const i64 r = base + lane / L;
    int lo = 0, hi = 0;
    if (r < rows) {
      lo = __ldg(rp + r);
      hi = __ldg(rp + r + 1);
    }
    double p = 0.0;
    if (MODE != 2) {
      for (int k = lo + sub; k < hi; k += L) p = fma(ldg_stream(va + k), __ldg(v + __ldg(ci + k)), p);
      p = group_sum<L>(p);
    }
    if (MODE == 0) {
      if (r < rows && sub == 0) out[r] = p;
      continue;
    }
    if (r >= rows || hi == lo) continue;
    double t;
    if (MODE == 1) {
      const double wr = (ct != SYSML_CHAIN_XtXv) ? __ldg(w + r) : 0.0;
      if (ct == SYSML_CHAIN_XtwXv && wr == 0.0) continue;  // LibMatrixMult.java:1667
      t = (ct == SYSML_CHAIN_XtwXv) ? p * wr : ((ct == SYSML_CHAIN_XtXvy) ? p - wr : p);
    } else {
      t = __ldg(w + r);
    }
    if (t == 0.0) continue;  // LibMatrixMult.java:1680 / VM skips zero entries of the left vector (:978)
    for (int k = lo + sub; k < hi; k += L) atomicAdd(out + __ldg(ci + k), t * va[k]);
  }
}

// ---- streamed CSR row kernel (large matrices, short rows) ------------------------------------------------
// The row-at-a-time kernel above is latency-bound on short rows (three dependent loads per row: rowptr ->
// colind/val -> v; measured 0.41 of the HBM copy rate at 10 nnz/row).  rowptr, colind and val of a range of
// rows are CONTIGUOUS in CSR, so here a producer thread streams them with 1-D bulk copies (TMA engine) into a
// ring of shared-memory stages -- tens of KB in flight per SM whatever the consumers do -- and the 16 consumer
// warps work out of shared memory: L lanes per row (L ~ mean/4, so a lane holds up to four independent gathers
// of v), every consumer thread owns one (row, lane) slot of the tile, warps never wait for each other.
// Tiles that do not fit a stage (a run of very long rows), the ragged last tile, and tiles whose aligned
// copy would run past the arrays are flagged "slow" and processed straight from global memory, one warp per row.
// NC consumer threads (896 = 28 warps, or 512 when only the smaller tiles fit beside v / out); a tile has NC / L rows

struct CsrStreamArgs {
  const int* rp;
  const int* ci;
  const double* va;
  const double* v;
  const double* w;
  double* out;
  i64 rows, nnz, out_len;
  int ct, replicas, cap, nst, ntiles;
  int v_smem;    // > 0: v (that many doubles) is copied into shared memory once per CTA and gathered from there
  int out_smem;  // > 0: scatter modes accumulate into a per-CTA shared-memory copy of out, written to partial row blockIdx.x at the end
  // Shared-memory fp64 atomics are CAS loops (~0.8 per clock per SM measured) and L2 fp64 reductions top out near
  // 150 G/s for the whole chip; they are different units, so red_eighths of every 8 tiles scatter with L2
  // reductions into red_out (red_replicas zeroed copies of out) while the others use the shared-memory copy.
  double* red_out;
  int red_replicas, red_eighths;
};

template <int L, int MODE, int NC>
__global__ void __launch_bounds__(NC + 32, 1) csr_stream_kernel(const CsrStreamArgs a) {
  constexpr int TR = NC / L;     // rows per tile
  constexpr int RPB = (TR + 4) * 4;   // bytes of a stage's rowptr slice (TR+1 entries, padded to 16 B)
  extern __shared__ __align__(16) unsigned char cs_smem[];
  uint64_t* full = reinterpret_cast<uint64_t*>(cs_smem);
  uint64_t* empty = full + 8;
  int4* meta = reinterpret_cast<int4*>(cs_smem + 128);  // per stage: {aligned first k, fast?, -, -}
  const int cap = a.cap, nst = a.nst;
  const size_t stage_bytes = (size_t)RPB + TR * 8 + (size_t)cap * 12;
  unsigned char* stages = cs_smem + 256;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) {
    for (int s = 0; s < nst; s++) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], NC / 32);
    }
    mbar_fence_init();
  }
  __syncthreads();
  const bool have_w = (MODE == 2) || (MODE == 1 && a.ct != SYSML_CHAIN_XtXv);

  if (warp == NC / 32) {
    // ===================== producer =====================
    if (lane != 0) return;
    const uint64_t pol = l2_policy_evict_first();
    int tile = blockIdx.x;
    int nk0 = 0, nk1 = 0;
    if (tile < a.ntiles) {
      const i64 r0 = (i64)tile * TR;
      const i64 r1 = (r0 + TR < a.rows) ? r0 + TR : a.rows;
      nk0 = __ldg(a.rp + r0);
      nk1 = __ldg(a.rp + r1);
    }
    int s = 0;
    uint32_t ph = 0;  // stage index and phase advance incrementally (no division by the runtime stage count)
    for (; tile < a.ntiles; tile += gridDim.x, s = (s + 1 == nst) ? 0 : s + 1, ph ^= (s == 0) ? 1u : 0u) {
      const int k0 = nk0, k1 = nk1;
      const int nt = tile + gridDim.x;
      if (nt < a.ntiles) {  // boundaries of the next tile: in flight while this one waits for its stage
        const i64 q0 = (i64)nt * TR;
        const i64 q1 = (q0 + TR < a.rows) ? q0 + TR : a.rows;
        nk0 = __ldg(a.rp + q0);
        nk1 = __ldg(a.rp + q1);
      }
      mbar_wait(&empty[s], ph ^ 1u);
      const i64 r0 = (i64)tile * TR;
      const int ka = k0 & ~3, kb = (k1 + 3) & ~3;
      const bool fast = (r0 + TR + 3 <= a.rows) && ((i64)kb <= a.nnz) && (kb - ka <= cap);
      meta[s] = make_int4(ka, fast ? 1 : 0, 0, 0);
      if (fast) {
        unsigned char* st = stages + (size_t)s * stage_bytes;
        const uint32_t nk = (uint32_t)(kb - ka);
        mbar_expect_tx(&full[s], RPB + (have_w ? TR * 8 : 0) + nk * (MODE == 3 ? 8u : 12u));
        bulk_g2s(st, a.rp + r0, RPB, &full[s], pol);
        if (have_w) bulk_g2s(st + RPB, a.w + r0, TR * 8, &full[s], pol);
        if (nk) {
          bulk_g2s(st + RPB + TR * 8, a.va + ka, nk * 8u, &full[s], pol);
          if (MODE != 3) bulk_g2s(st + RPB + TR * 8 + (size_t)cap * 8, a.ci + ka, nk * 4u, &full[s], pol);
        }
      } else {
        mbar_arrive(&full[s]);
      }
    }
    return;
  }

  // ===================== consumers =====================
  double* out = a.out;
  if (MODE != 0 && a.replicas > 1) out += (i64)(blockIdx.x % a.replicas) * a.out_len;
  // Random 8-byte gathers of v from L2 cost a 32-byte sector each -- more L2 traffic than the matrix stream
  // itself -- and the stages leave little L1.  A v that fits is therefore staged in shared memory, and so is
  // the scatter target of modes 1/2 (shared-memory fp64 atomics instead of L2 atomics).
  double* v_s = reinterpret_cast<double*>(stages + (size_t)nst * stage_bytes);
  double* out_s = v_s + a.v_smem;
  const double* vsrc = a.v;
  if (MODE != 2 && a.v_smem > 0) {
    for (int i = tid; i < a.v_smem; i += NC) v_s[i] = __ldg(a.v + i);
    vsrc = v_s;
  }
  double* const out_g = out;  // global target (replica) of the scatter modes
  if (MODE != 0 && a.out_smem > 0) {
    for (int i = tid; i < a.out_smem; i += NC) out_s[i] = 0.0;
  }
  if ((MODE != 2 && a.v_smem > 0) || (MODE != 0 && a.out_smem > 0)) named_bar_sync(1, NC);
  const int sub = lane & (L - 1);
  const int row = tid / L;  // this thread's row of every tile
  int s = 0;
  uint32_t ph = 0, it = 0;
  for (int tile = blockIdx.x; tile < a.ntiles; tile += gridDim.x, it++, s = (s + 1 == nst) ? 0 : s + 1, ph ^= (s == 0) ? 1u : 0u) {
    mbar_wait(&full[s], ph);
    const int4 mt = meta[s];
    const i64 r0 = (i64)tile * TR;
    if (mt.y) {
      const unsigned char* st = stages + (size_t)s * stage_bytes;
      const int* rp_s = reinterpret_cast<const int*>(st);
      const double* w_s = reinterpret_cast<const double*>(st + RPB);
      const double* va_s = reinterpret_cast<const double*>(st + RPB + TR * 8);
      const int* ci_s = reinterpret_cast<const int*>(st + RPB + TR * 8 + (size_t)cap * 8);
      const int lo = rp_s[row] - mt.x, hi = rp_s[row + 1] - mt.x;
      if (MODE == 3) {
        if (a.ct <= SYSML_AGG_SUMSQ) {
          CompSum cs;
          for (int k = lo + sub; k < hi; k += L) {
            const double x = va_s[k];
            cs.add(a.ct == SYSML_AGG_SUMSQ ? __dmul_rn(x, x) : x);
          }
#pragma unroll
          for (int o = L >> 1; o > 0; o >>= 1) {
            CompSum t;
            t.hi = __shfl_xor_sync(0xffffffffu, cs.hi, o);
            t.lo = __shfl_xor_sync(0xffffffffu, cs.lo, o);
            cs.add(t);
          }
          if (sub == 0) a.out[r0 + row] = (isinf(cs.hi) || cs.lo != cs.lo) ? cs.hi : cs.value();
        } else {
          const bool ismax = a.ct == SYSML_AGG_MAX;
          double t = ismax ? -CUDART_INF : CUDART_INF;
          for (int k = lo + sub; k < hi; k += L) t = ismax ? jmax(t, va_s[k]) : jmin(t, va_s[k]);
#pragma unroll
          for (int o = L >> 1; o > 0; o >>= 1) {
            const double u = __shfl_xor_sync(0xffffffffu, t, o);
            t = ismax ? jmax(t, u) : jmin(t, u);
          }
          if ((i64)(hi - lo) < a.out_len) t = ismax ? jmax(t, 0.0) : jmin(t, 0.0);
          if (sub == 0) a.out[r0 + row] = t;
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[s]);
        continue;
      }
      double p = 0.0;
      if (MODE != 2) {
        if (a.v_smem > 0) {  // shared-memory gathers: short latency, a plain loop is cheapest
          for (int k = lo + sub; k < hi; k += L) p = fma(va_s[k], v_s[ci_s[k]], p);
        } else {
          for (int k = lo + sub; k < hi; k += 4 * L) {  // up to four independent L1/L2 gathers in flight per lane
            const int k1 = k + L, k2 = k + 2 * L, k3 = k + 3 * L;
            const double x0 = __ldg(a.v + ci_s[k]);
            const double x1 = (k1 < hi) ? __ldg(a.v + ci_s[k1]) : 0.0;
            const double x2 = (k2 < hi) ? __ldg(a.v + ci_s[k2]) : 0.0;
            const double x3 = (k3 < hi) ? __ldg(a.v + ci_s[k3]) : 0.0;
            p = fma(va_s[k], x0, p);
            if (k1 < hi) p = fma(va_s[k1], x1, p);
            if (k2 < hi) p = fma(va_s[k2], x2, p);
            if (k3 < hi) p = fma(va_s[k3], x3, p);
          }
        }
        p = group_sum<L>(p);
      }
      if (MODE == 0) {
        if (sub == 0) out[r0 + row] = p;
      } else if (hi > lo) {
        double t;
        bool skip = false;
        if (MODE == 1) {
          const double wr = have_w ? w_s[row] : 0.0;
          skip = (a.ct == SYSML_CHAIN_XtwXv && wr == 0.0);  // LibMatrixMult.java:1667
          t = (a.ct == SYSML_CHAIN_XtwXv) ? p * wr : ((a.ct == SYSML_CHAIN_XtXvy) ? p - wr : p);
        } else {
          t = w_s[row];
        }
        if (!skip && t != 0.0) {
          // two loops so that each atomic has a known address space (a generic-address atomic is much slower)
          if (a.out_smem > 0 && (int)(it & 7u) >= a.red_eighths) {
            for (int k = lo + sub; k < hi; k += L) atomicAdd(out_s + ci_s[k], t * va_s[k]);
          } else {
            double* o = (a.out_smem > 0) ? a.red_out + (i64)(blockIdx.x % a.red_replicas) * a.out_len : out_g;
            for (int k = lo + sub; k < hi; k += L) atomicAdd(o + ci_s[k], t * va_s[k]);
          }
        }
      }
    } else {
      // slow tile: straight from global memory, one warp per row
      const i64 rend = (r0 + TR < a.rows) ? r0 + TR : a.rows;
      for (i64 r = r0 + warp; r < rend; r += NC / 32) {
        const int lo = __ldg(a.rp + r), hi = __ldg(a.rp + r + 1);
        if (MODE == 3) {
          if (a.ct <= SYSML_AGG_SUMSQ) {
            CompSum cs;
            for (int k = lo + lane; k < hi; k += 32) {
              const double x = ldg_stream(a.va + k);
              cs.add(a.ct == SYSML_AGG_SUMSQ ? __dmul_rn(x, x) : x);
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
              CompSum t;
              t.hi = __shfl_xor_sync(0xffffffffu, cs.hi, o);
              t.lo = __shfl_xor_sync(0xffffffffu, cs.lo, o);
              cs.add(t);
            }
            if (lane == 0) a.out[r] = (isinf(cs.hi) || cs.lo != cs.lo) ? cs.hi : cs.value();
          } else {
            const bool ismax = a.ct == SYSML_AGG_MAX;
            double t = ismax ? -CUDART_INF : CUDART_INF;
            for (int k = lo + lane; k < hi; k += 32) t = ismax ? jmax(t, ldg_stream(a.va + k)) : jmin(t, ldg_stream(a.va + k));
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
              const double u = __shfl_xor_sync(0xffffffffu, t, o);
              t = ismax ? jmax(t, u) : jmin(t, u);
            }
            if ((i64)(hi - lo) < a.out_len) t = ismax ? jmax(t, 0.0) : jmin(t, 0.0);
            if (lane == 0) a.out[r] = t;
          }
          continue;
        }
        double p = 0.0;
        if (MODE != 2) {
          for (int k = lo + lane; k < hi; k += 32) p = fma(ldg_stream(a.va + k), vsrc[__ldg(a.ci + k)], p);
          p = group_sum<32>(p);
        }
        if (MODE == 0) {
          if (lane == 0) out[r] = p;
          continue;
        }
        if (hi == lo) continue;
        double t;
        if (MODE == 1) {
          const double wr = have_w ? __ldg(a.w + r) : 0.0;
          if (a.ct == SYSML_CHAIN_XtwXv && wr == 0.0) continue;
          t = (a.ct == SYSML_CHAIN_XtwXv) ? p * wr : ((a.ct == SYSML_CHAIN_XtXvy) ? p - wr : p);
        } else {
          t = __ldg(a.w + r);
        }
        if (t == 0.0) continue;
        if (a.out_smem > 0) {
          for (int k = lo + lane; k < hi; k += 32) atomicAdd(out_s + __ldg(a.ci + k), t * a.va[k]);
        } else {
          for (int k = lo + lane; k < hi; k += 32) atomicAdd(out_g + __ldg(a.ci + k), t * a.va[k]);
        }
      }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(&empty[s]);
  }
  if (MODE != 0 && a.out_smem > 0) {
    named_bar_sync(1, NC);
    double* dst = a.out + (i64)blockIdx.x * a.out_len;
    for (int i = tid; i < a.out_smem; i += NC) dst[i] = out_s[i];
  }
}

// C[r, 0:n) = sum_k A[r,k] * B[k, 0:n): one warp per row, lanes over output columns (n <= 64: two per lane)
template <int NPL>
__global__ void __launch_bounds__(256) csr_spmm_skinny_kernel(const int* __restrict__ rp, const int* __restrict__ ci,
                                                              const double* __restrict__ va, const double* __restrict__ B,
                                                              double* __restrict__ C, i64 rows, int n) {
  const int lane = threadIdx.x & 31;
  const i64 warp = ((i64)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const i64 nwarps = ((i64)gridDim.x * blockDim.x) >> 5;
  for (i64 r = warp; r < rows; r += nwarps) {
    const int lo = __ldg(rp + r), hi = __ldg(rp + r + 1);
    double acc[NPL];
#pragma unroll
    for (int q = 0; q < NPL; q++) acc[q] = 0.0;
    for (int k0 = lo; k0 < hi; k0 += 32) {
      // lanes fetch 32 non-zeros at once, then broadcast them one by one
      const int kk = k0 + lane;
      const double a_l = (kk < hi) ? ldg_stream(va + kk) : 0.0;
      const int c_l = (kk < hi) ? __ldg(ci + kk) : 0;
      const int cnt = (hi - k0 < 32) ? (hi - k0) : 32;
      for (int q = 0; q < cnt; q++) {
        const double a = __shfl_sync(0xffffffffu, a_l, q);
        const i64 brow = (i64)__shfl_sync(0xffffffffu, c_l, q) * n;
#pragma unroll
        for (int u = 0; u < NPL; u++) {
          const int j = lane + 32 * u;
          if (j < n) acc[u] = fma(a, __ldg(B + brow + j), acc[u]);
        }
      }
    }
#pragma unroll
    for (int u = 0; u < NPL; u++) {
      const int j = lane + 32 * u;
      if (j < n) C[r * n + j] = acc[u];
    }
  }
}

// general n: grid.y tiles the output columns by 128 (4 per lane)
__global__ void __launch_bounds__(256) csr_spmm_wide_kernel(const int* __restrict__ rp, const int* __restrict__ ci,
                                                            const double* __restrict__ va, const double* __restrict__ B,
                                                            double* __restrict__ C, i64 rows, i64 n) {
  const int lane = threadIdx.x & 31;
  const i64 warp = ((i64)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const i64 nwarps = ((i64)gridDim.x * blockDim.x) >> 5;
  const i64 j0 = (i64)blockIdx.y * 128;
  for (i64 r = warp; r < rows; r += nwarps) {
    const int lo = __ldg(rp + r), hi = __ldg(rp + r + 1);
    double acc[4] = {0.0, 0.0, 0.0, 0.0};
    for (int k = lo; k < hi; k++) {
      const double a = __ldg(va + k);
      const i64 brow = (i64)__ldg(ci + k) * n;
#pragma unroll
      for (int u = 0; u < 4; u++) {
        const i64 j = j0 + lane + 32 * u;
        if (j < n) acc[u] = fma(a, __ldg(B + brow + j), acc[u]);
      }
    }
#pragma unroll
    for (int u = 0; u < 4; u++) {
      const i64 j = j0 + lane + 32 * u;
      if (j < n) C[r * n + j] = acc[u];
    }
  }
}

// C (a_cols x n) += t(A) %*% B : row r of A scatters a * B[r, :] into C[col, :]
__global__ void __launch_bounds__(256) csr_tspmm_kernel(const int* __restrict__ rp, const int* __restrict__ ci,
                                                        const double* __restrict__ va, const double* __restrict__ B,
                                                        double* __restrict__ C, i64 rows, i64 n) {
  const int lane = threadIdx.x & 31;
  const i64 warp = ((i64)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const i64 nwarps = ((i64)gridDim.x * blockDim.x) >> 5;
  for (i64 r = warp; r < rows; r += nwarps) {
    const int lo = __ldg(rp + r), hi = __ldg(rp + r + 1);
    for (int k = lo; k < hi; k++) {
      const double a = __ldg(va + k);
      if (a == 0.0) continue;
      double* crow = C + (i64)__ldg(ci + k) * n;
      for (i64 j = lane; j < n; j += 32) atomicAdd(crow + j, a * __ldg(B + r * n + j));
    }
  }
}

// sparse tsmm LEFT: for each row, C[ci[i], ci[j]] += v[i]*v[j] for j >= i (upper triangle)
__global__ void __launch_bounds__(256) csr_tsmm_left_kernel(const int* __restrict__ rp, const int* __restrict__ ci,
                                                            const double* __restrict__ va, double* __restrict__ C, i64 rows, i64 n) {
  const int lane = threadIdx.x & 31;
  const i64 warp = ((i64)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const i64 nwarps = ((i64)gridDim.x * blockDim.x) >> 5;
  for (i64 r = warp; r < rows; r += nwarps) {
    const int lo = __ldg(rp + r), hi = __ldg(rp + r + 1);
    for (int i = lo; i < hi; i++) {
      const double a = __ldg(va + i);
      if (a == 0.0) continue;
      double* crow = C + (i64)__ldg(ci + i) * n;
      for (int j = i + lane; j < hi; j += 32) atomicAdd(crow + __ldg(ci + j), a * __ldg(va + j));
    }
  }
}

// mirror upper -> lower and count non-zeros (diagonal 1, off-diagonal 2): copyUpperToLowerTriangle
__global__ void mirror_count_kernel(double* __restrict__ C, i64 n, unsigned long long* __restrict__ nnz) {
  const i64 total = n * n;
  const i64 stride = (i64)gridDim.x * blockDim.x;
  unsigned long long cnt = 0;
  for (i64 idx = (i64)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += stride) {
    const i64 i = idx / n, j = idx - i * n;
    if (j < i) continue;
    const double v = C[idx];
    if (j > i) C[j * n + i] = v;
    if (v != 0.0) cnt += (j == i) ? 1ull : 2ull;
  }
  for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
  if ((threadIdx.x & 31) == 0 && cnt) atomicAdd(nnz, cnt);
}


// ---------------- remaining sparsity combinations (dense output) ----------------
// C[i, :] = sum_k A[i,k] * B[k, :] with B in CSR: one warp per output row, k sequential, lanes over the
// non-zeros of B's row k (distinct columns within a row, so no atomics; __syncwarp orders the k steps).
// LibMatrixMult.matrixMultDenseSparse (LibMatrixMult.java:1105-1168).
__global__ void __launch_bounds__(256) dense_csr_mm_kernel(const double* __restrict__ A, const int* __restrict__ rp, const int* __restrict__ ci,
                                                           const double* __restrict__ va, double* __restrict__ C, i64 m, i64 k, i64 n) {
  const int lane = threadIdx.x & 31;
  const i64 warp = ((i64)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const i64 nwarps = ((i64)gridDim.x * blockDim.x) >> 5;
  for (i64 i = warp; i < m; i += nwarps) {
    double* crow = C + i * n;
    for (i64 j = lane; j < n; j += 32) crow[j] = 0.0;
    __syncwarp();
    for (i64 kk = 0; kk < k; kk++) {
      const double a = __ldg(A + i * k + kk);
      if (a == 0.0) continue;  // warp-uniform: every lane reads the same a
      const int lo = __ldg(rp + kk), hi = __ldg(rp + kk + 1);
      for (int q = lo + lane; q < hi; q += 32) crow[ci[q]] += a * va[q];
      __syncwarp();
    }
  }
}
// C[i, :] = sum over non-zeros (k, a) of A's row i of a * B[k, :], A and B in CSR.
// LibMatrixMult.matrixMultSparseSparse (LibMatrixMult.java:1397-1500); the result is written dense.
__global__ void __launch_bounds__(256) csr_csr_mm_kernel(const int* __restrict__ arp, const int* __restrict__ aci, const double* __restrict__ ava,
                                                         const int* __restrict__ brp, const int* __restrict__ bci, const double* __restrict__ bva,
                                                         double* __restrict__ C, i64 m, i64 n) {
  const int lane = threadIdx.x & 31;
  const i64 warp = ((i64)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const i64 nwarps = ((i64)gridDim.x * blockDim.x) >> 5;
  for (i64 i = warp; i < m; i += nwarps) {
    double* crow = C + i * n;
    for (i64 j = lane; j < n; j += 32) crow[j] = 0.0;
    __syncwarp();
    const int alo = __ldg(arp + i), ahi = __ldg(arp + i + 1);
    for (int p = alo; p < ahi; p++) {
      const double a = __ldg(ava + p);
      const int kk = __ldg(aci + p);
      const int lo = __ldg(brp + kk), hi = __ldg(brp + kk + 1);
      for (int q = lo + lane; q < hi; q += 32) crow[bci[q]] += a * bva[q];
      __syncwarp();
    }
  }
}

// ---------------- nnz / conversions ----------------
__global__ void count_nnz_kernel(const double* __restrict__ X, i64 len, unsigned long long* __restrict__ out) {
  const i64 stride = (i64)gridDim.x * blockDim.x;
  unsigned long long cnt = 0;
  for (i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x; i < len; i += stride) cnt += (ldg_stream(X + i) != 0.0) ? 1ull : 0ull;
  for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
  if ((threadIdx.x & 31) == 0 && cnt) atomicAdd(out, cnt);
}

// one warp per row: rowcnt[r] = #non-zeros
__global__ void row_nnz_kernel(const double* __restrict__ X, i64 rows, i64 cols, int* __restrict__ rowcnt) {
  const int lane = threadIdx.x & 31;
  const i64 warp = ((i64)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const i64 nwarps = ((i64)gridDim.x * blockDim.x) >> 5;
  for (i64 r = warp; r < rows; r += nwarps) {
    int cnt = 0;
    for (i64 j = lane; j < cols; j += 32) cnt += (ldg_stream(X + r * cols + j) != 0.0) ? 1 : 0;
    for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
    if (lane == 0) rowcnt[r + 1] = cnt;
  }
}

// single-block exclusive scan of rowcnt[1..rows] into rowptr (rowptr[0] = 0); 64-bit running total
__global__ void __launch_bounds__(1024) scan_rowptr_kernel(int* __restrict__ rowptr, i64 rows, unsigned long long* __restrict__ total) {
  __shared__ unsigned long long wsum[32];
  __shared__ unsigned long long carry;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) {
    carry = 0;
    rowptr[0] = 0;
  }
  __syncthreads();
  for (i64 base = 0; base < rows; base += 1024) {
    const i64 i = base + threadIdx.x;
    unsigned long long v = (i < rows) ? (unsigned long long)rowptr[i + 1] : 0ull;
    unsigned long long x = v;
    for (int o = 1; o < 32; o <<= 1) {
      const unsigned long long y = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= o) x += y;
    }
    if (lane == 31) wsum[warp] = x;
    __syncthreads();
    if (warp == 0) {
      unsigned long long s = wsum[lane];
      for (int o = 1; o < 32; o <<= 1) {
        const unsigned long long y = __shfl_up_sync(0xffffffffu, s, o);
        if (lane >= o) s += y;
      }
      wsum[lane] = s;
    }
    __syncthreads();
    const unsigned long long incl = x + (warp ? wsum[warp - 1] : 0ull) + carry;
    if (i < rows) rowptr[i + 1] = (int)incl;
    __syncthreads();
    if (threadIdx.x == 1023) carry = incl;
    __syncthreads();
  }
  if (threadIdx.x == 0) *total = carry;
}

// one warp per row: ordered compaction with ballot
__global__ void fill_csr_kernel(const double* __restrict__ X, i64 rows, i64 cols, const int* __restrict__ rowptr,
                                int* __restrict__ ci, double* __restrict__ va) {
  const int lane = threadIdx.x & 31;
  const i64 warp = ((i64)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const i64 nwarps = ((i64)gridDim.x * blockDim.x) >> 5;
  for (i64 r = warp; r < rows; r += nwarps) {
    int pos = rowptr[r];
    for (i64 j0 = 0; j0 < cols; j0 += 32) {
      const i64 j = j0 + lane;
      const double x = (j < cols) ? ldg_stream(X + r * cols + j) : 0.0;
      const unsigned m = __ballot_sync(0xffffffffu, x != 0.0);
      if (x != 0.0) {
        const int off = __popc(m & ((1u << lane) - 1));
        ci[pos + off] = (int)j;
        va[pos + off] = x;
      }
      pos += __popc(m);
    }
  }
}

__global__ void csr_to_dense_kernel(const int* __restrict__ rp, const int* __restrict__ ci, const double* __restrict__ va,
                                    i64 rows, i64 cols, double* __restrict__ X) {
  const int lane = threadIdx.x & 31;
  const i64 warp = ((i64)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const i64 nwarps = ((i64)gridDim.x * blockDim.x) >> 5;
  for (i64 r = warp; r < rows; r += nwarps) {
    const int lo = rp[r], hi = rp[r + 1];
    for (int k = lo + lane; k < hi; k += 32) X[r * cols + ci[k]] = va[k];
  }
}

// ---------------- sparse aggregates ----------------
// per-row: sum / sumsq (op 0/1), max / min (2/3) with the implicit-zero correction, rowIndexMax/Min (4/5).
// L lanes per row (L chosen like SpMV from the mean nnz/row), consecutive rows in consecutive lane groups.
template <int L>
__global__ void __launch_bounds__(256) csr_agg_row_kernel(const int* __restrict__ rp, const int* __restrict__ ci,
                                                          const double* __restrict__ va, i64 rows, i64 cols, int op,
                                                          double* __restrict__ out) {
  const int lane = threadIdx.x & 31;
  const int sub = lane & (L - 1);
  constexpr int RPW = 32 / L;
  const i64 warp = ((i64)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const i64 nwarps = ((i64)gridDim.x * blockDim.x) >> 5;
  for (i64 base = warp * RPW; base < rows; base += nwarps * RPW) {
    const i64 r = base + lane / L;
    int lo = 0, hi = 0;
    if (r < rows) {
      lo = __ldg(rp + r);
      hi = __ldg(rp + r + 1);
    }
    const int alen = hi - lo;
    if (op <= 1) {
      CompSum s;
      for (int k = lo + sub; k < hi; k += L) {
        const double x = ldg_stream(va + k);
        s.add(op == 1 ? __dmul_rn(x, x) : x);
      }
#pragma unroll
      for (int o = L >> 1; o > 0; o >>= 1) {
        CompSum t;
        t.hi = __shfl_xor_sync(0xffffffffu, s.hi, o);
        t.lo = __shfl_xor_sync(0xffffffffu, s.lo, o);
        s.add(t);
      }
      if (r < rows && sub == 0) out[r] = (isinf(s.hi) || s.lo != s.lo) ? s.hi : s.value();
    } else if (op <= 3) {
      double t = (op == 2) ? -CUDART_INF : CUDART_INF;
      for (int k = lo + sub; k < hi; k += L) t = (op == 2) ? jmax(t, ldg_stream(va + k)) : jmin(t, ldg_stream(va + k));
#pragma unroll
      for (int o = L >> 1; o > 0; o >>= 1) {
        const double u = __shfl_xor_sync(0xffffffffu, t, o);
        t = (op == 2) ? jmax(t, u) : jmin(t, u);
      }
      if (alen < cols) t = (op == 2) ? jmax(t, 0.0) : jmin(t, 0.0);
      if (r < rows && sub == 0) out[r] = t;
    } else {
      // LibMatrixAgg.s_uarimxx / s_uarimin (LibMatrixAgg.java:2536-2614)
      const bool ismax = (op == 4);
      double best = ismax ? -CUDART_INF : CUDART_INF;
      int idx = -1;  // position within the row
      for (int k = lo + sub; k < hi; k += L) {
        const double x = ldg_stream(va + k);
        const bool take = ismax ? (x >= best) : (x <= best);
        idx = take ? (k - lo) : idx;
        best = take ? x : best;
      }
#pragma unroll
      for (int o = L >> 1; o > 0; o >>= 1) {
        const double ob = __shfl_xor_sync(0xffffffffu, best, o);
        const int oi = __shfl_xor_sync(0xffffffffu, idx, o);
        const bool other = (oi >= 0) && ((idx < 0) || (ismax ? (ob > best) : (ob < best)) || (ob == best && oi > idx));
        best = other ? ob : best;
        idx = other ? oi : idx;
      }
      if (r < rows && sub == 0) {
        double res;
        if (alen == 0) {
          res = (double)cols;  // empty row: last column, value 0 (:2559-2563)
        } else {
          res = (idx >= 0) ? (double)ci[lo + idx] + 1.0 : 0.0;
          const double bv = (idx >= 0) ? best : CUDART_NAN;
          const bool zero_wins = ismax ? (0.0 > bv) : (0.0 < bv);  // strictly: Builtin.execute(0,v)==1
          if (alen < cols && zero_wins) {
            // last column that holds an implicit zero: walk back while the tail is densely occupied
            i64 ix = cols - 1;
            for (int j = hi - 1; j >= lo; j--, ix--)
              if (ci[j] != ix) break;
            res = (double)(ix + 1);
          }
        }
        out[r] = res;
      }
    }
  }
}

// ALL over the value array (sum, sumsq): reuse the dense kernel through the public entry; max/min here
__global__ void csr_agg_all_minmax_kernel(const double* __restrict__ va, i64 nnz, int ismax, double* __restrict__ partial) {
  double t = ismax ? -CUDART_INF : CUDART_INF;
  const i64 stride = (i64)gridDim.x * blockDim.x;
  for (i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x; i < nnz; i += stride) t = ismax ? jmax(t, va[i]) : jmin(t, va[i]);
  for (int o = 16; o > 0; o >>= 1) {
    const double u = __shfl_xor_sync(0xffffffffu, t, o);
    t = ismax ? jmax(t, u) : jmin(t, u);
  }
  __shared__ double sh[32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) sh[warp] = t;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int q = 1; q < (int)(blockDim.x >> 5); q++) t = ismax ? jmax(t, sh[q]) : jmin(t, sh[q]);
    partial[blockIdx.x] = t;
  }
}
__global__ void csr_agg_all_minmax_finish(const double* __restrict__ partial, int nparts, int ismax, int has_zero, double* __restrict__ out) {
  if (threadIdx.x || blockIdx.x) return;
  double t = ismax ? -CUDART_INF : CUDART_INF;
  for (int i = 0; i < nparts; i++) t = ismax ? jmax(t, partial[i]) : jmin(t, partial[i]);
  if (has_zero) t = ismax ? jmax(t, 0.0) : jmin(t, 0.0);
  out[0] = t;
}

// per-column sum/sumsq via atomics into a zeroed output; max/min with counts
__global__ void csr_agg_col_sum_kernel(const int* __restrict__ ci, const double* __restrict__ va, i64 nnz, int sq, double* __restrict__ out) {
  const i64 stride = (i64)gridDim.x * blockDim.x;
  for (i64 k = (i64)blockIdx.x * blockDim.x + threadIdx.x; k < nnz; k += stride) {
    const double x = va[k];
    atomicAdd(out + ci[k], sq ? x * x : x);
  }
}
__device__ __forceinline__ void atomic_minmax_double(double* addr, double v, int ismax) {
  unsigned long long* a = reinterpret_cast<unsigned long long*>(addr);
  unsigned long long old = *a, assumed;
  do {
    assumed = old;
    const double cur = __longlong_as_double((long long)assumed);
    const double nv = ismax ? jmax(cur, v) : jmin(cur, v);
    if (__double_as_longlong(nv) == (long long)assumed) break;
    old = atomicCAS(a, assumed, (unsigned long long)__double_as_longlong(nv));
  } while (assumed != old);
}
__global__ void csr_agg_col_minmax_kernel(const int* __restrict__ ci, const double* __restrict__ va, i64 nnz, int ismax,
                                          double* __restrict__ out, int* __restrict__ cnt) {
  const i64 stride = (i64)gridDim.x * blockDim.x;
  for (i64 k = (i64)blockIdx.x * blockDim.x + threadIdx.x; k < nnz; k += stride) {
    atomic_minmax_double(out + ci[k], va[k], ismax);
    atomicAdd(cnt + ci[k], 1);
  }
}
__global__ void csr_agg_col_minmax_finish(double* __restrict__ out, const int* __restrict__ cnt, i64 cols, i64 rows, int ismax) {
  const i64 j = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= cols) return;
  if ((i64)cnt[j] < rows) out[j] = ismax ? jmax(out[j], 0.0) : jmin(out[j], 0.0);
}

inline unsigned flat_grid(sysml_b200_ctx* ctx, i64 items, int threads) {
  i64 blocks = (items + threads - 1) / threads;
  const i64 cap = (i64)ctx->sm_count * 16;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return (unsigned)blocks;
}
inline unsigned warp_rows_grid(sysml_b200_ctx* ctx, i64 rows) {
  i64 blocks = (rows + 7) / 8;
  const i64 cap = (i64)ctx->sm_count * 16;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return (unsigned)blocks;
}

inline int sysml_reduce_partial_rows_fwd(sysml_b200_ctx* ctx, const double* partial, int P, int m, double* out) {
  return ::sysml_reduce_partial_rows(ctx, partial, P, m, out);
}

constexpr int SYSML_NOT_STREAMED = 12345;  // internal: launch_csr_row<3> found the matrix not eligible for the streamed kernel

template <int MODE>
int launch_csr_row(sysml_b200_ctx* ctx, const int* rp, const int* ci, const double* va, const double* v, const double* w,
                   double* out, i64 rows, i64 nnz, int ct, i64 out_len, i64 vlen) {
  const int L = pick_lanes(nnz, rows);
  unsigned g = rows_grid(ctx, rows, L, 256);
  // streamed kernel: big matrices with short rows and 16-byte aligned arrays (bulk copies)
  CsrStreamArgs sa{};
  bool stream = false;
  size_t stream_smem = 0;
  int SL = 1;  // lanes per row in the streamed kernel
  int NC = 512;
  {
    const bool has_w = (MODE == 2) || (MODE == 1 && ct != SYSML_CHAIN_XtXv);
    const uintptr_t al = reinterpret_cast<uintptr_t>(rp) | reinterpret_cast<uintptr_t>(ci) | reinterpret_cast<uintptr_t>(va) |
                         (has_w ? reinterpret_cast<uintptr_t>(w) : 0);
    const double mean = (double)nnz / (double)rows;
    // ~8 entries per lane for the gather / reduce modes, ~4 for the scatter modes (measured on 20M x 10k, ~10 per row:
    // SpMV 0.55 -> 0.53 ms, rowSums 0.66 -> 0.46 ms with 2 lanes; t(X)y and mmchain are faster with 4)
    const int per_lane = (MODE == 0 || MODE == 3) ? 8 : 4;
    while (SL < 32 && per_lane * SL < mean) SL <<= 1;
    if (const char* e = getenv("SYSML_B200_CSR_STREAM_LANES")) {  // A/B switch for measurements
      const int v_ = atoi(e);
      if (v_ == 1 || v_ == 2 || v_ == 4 || v_ == 8 || v_ == 16 || v_ == 32) SL = v_;
    }
    if ((al & 15) == 0 && nnz >= (1 << 20) && rows >= (i64)(896 / SL) * ctx->sm_count && mean <= 256.0 &&
        getenv("SYSML_B200_CSR_NO_STREAM") == nullptr) {
      const size_t all = (size_t)ctx->max_smem_optin;
      auto stage_of = [&](int nc, int* cap) {
        const int TR = nc / SL;
        *cap = (int)(((i64)(TR * mean * 1.25) + 64 + 3) & ~3ll);
        return (size_t)(TR + 4) * 4 + (size_t)TR * 8 + (size_t)*cap * 12;
      };
      // Streaming pays only with v (modes 0/1) and the scatter target (modes 1/2) resident in shared memory
      // (measured: without them the row-at-a-time kernel, which has the whole L1 for v, is faster): one CTA per
      // SM with 896 consumers, or 512 when only the smaller tiles leave room for three stages.
      const size_t need = ((MODE == 0 || MODE == 1) ? (size_t)vlen * 8 : 0) + ((MODE == 1 || MODE == 2) ? (size_t)out_len * 8 : 0);
      const size_t fixed = 256 + need;
      int cap = 0, nst = 0;
      size_t stage = stage_of(896, &cap);
      NC = 896;
      int min_st = 3;
      if (fixed + 3 * stage > all && MODE == 1) {  // sparse mmchain with v AND out resident: 24 consumer warps on two stages
                                                  // beat 16 on three (the shared-memory atomics want warps: 1.47 -> 1.34 ms)
        stage = stage_of(768, &cap);
        NC = 768;
        min_st = 2;
      }
      if (fixed + min_st * stage > all) {
        stage = stage_of(512, &cap);
        NC = 512;
        min_st = 3;
      }
      if (fixed + min_st * stage <= all) {
        nst = (int)((all - fixed) / stage);
        if (MODE == 0 || MODE == 1) sa.v_smem = (int)vlen;
        if (MODE == 1 || MODE == 2) sa.out_smem = (int)out_len;
      }
      if (nst > 6) nst = 6;
      if (nst >= 2) {
        stream = true;
        stream_smem = fixed + (size_t)nst * stage;
        const int TR = NC / SL;
        sa.rp = rp; sa.ci = ci; sa.va = va; sa.v = v; sa.w = w; sa.rows = rows; sa.nnz = nnz; sa.out_len = out_len;
        sa.ct = ct; sa.cap = cap; sa.nst = nst; sa.ntiles = (int)((rows + TR - 1) / TR);
        g = (unsigned)ctx->sm_count;
        if ((i64)g > sa.ntiles) g = (unsigned)sa.ntiles;
      }
    }
  }
  int replicas = 1;
  double* target = out;
  if (MODE == 3 && !stream) return SYSML_NOT_STREAMED;
  if (stream && MODE != 0 && sa.out_smem > 0) {
    // one partial row per CTA (every cell written by the kernel) + 32 zeroed rows for the L2-reduction share
    sa.red_replicas = 32;
    sa.red_eighths = 3;
    if (const char* e = getenv("SYSML_B200_CSR_RED_EIGHTHS")) sa.red_eighths = atoi(e);
    replicas = (int)g + sa.red_replicas;
    void* ws = nullptr;
    int rc = sysml_workspace(ctx, (size_t)replicas * out_len * sizeof(double), &ws);
    if (rc) return rc;
    target = (double*)ws;
    sa.red_out = target + (size_t)g * out_len;
    SYSML_CUDA(ctx, cudaMemsetAsync(sa.red_out, 0, (size_t)sa.red_replicas * out_len * sizeof(double), ctx->stream));
  } else if ((MODE == 1 || MODE == 2) && out_len <= (1 << 20) && nnz > 64 * out_len) {
    replicas = 32;
    while (replicas > 1 && (i64)replicas * out_len * 8 > (64ll << 20)) replicas >>= 1;
    if ((unsigned)replicas > g) replicas = (int)g;
  }
  if (replicas > 1 && target == out) {
    void* ws = nullptr;
    int rc = sysml_workspace(ctx, (size_t)replicas * out_len * sizeof(double), &ws);
    if (rc) return rc;
    target = (double*)ws;
    SYSML_CUDA(ctx, cudaMemsetAsync(target, 0, (size_t)replicas * out_len * sizeof(double), ctx->stream));
  }
  if (stream) {
    sa.out = target;
    sa.replicas = replicas;
#define CS_LAUNCH2(LL, NCC)                                                                                             \
  {                                                                                                                     \
    auto k = csr_stream_kernel<LL, MODE, NCC>;                                                                          \
    SYSML_CUDA(ctx, cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)stream_smem));            \
    k<<<g, NCC + 32, stream_smem, ctx->stream>>>(sa);                                                                   \
  }
#define CS_LAUNCH(LL)              \
  if (NC == 896) CS_LAUNCH2(LL, 896) \
  else if (NC == 768 && MODE == 1) CS_LAUNCH2(LL, (MODE == 1 ? 768 : 512)) \
  else CS_LAUNCH2(LL, 512)
    switch (SL) {
      case 1: CS_LAUNCH(1) break;
      case 2: CS_LAUNCH(2) break;
      case 4: CS_LAUNCH(4) break;
      case 8: CS_LAUNCH(8) break;
      case 16: CS_LAUNCH(16) break;
      default: CS_LAUNCH(32) break;
    }
#undef CS_LAUNCH2
#undef CS_LAUNCH
    SYSML_LAUNCH_CHECK(ctx, "csr_stream_kernel");
    if (replicas > 1) return sysml_reduce_partial_rows_fwd(ctx, target, replicas, (int)out_len, out);
    return SYSML_OK;
  }
  if constexpr (MODE != 3) {
  switch (L) {
    case 2: csr_row_kernel<2, MODE><<<g, 256, 0, ctx->stream>>>(rp, ci, va, v, w, target, rows, ct, replicas, out_len); break;
    case 4: csr_row_kernel<4, MODE><<<g, 256, 0, ctx->stream>>>(rp, ci, va, v, w, target, rows, ct, replicas, out_len); break;
    case 8: csr_row_kernel<8, MODE><<<g, 256, 0, ctx->stream>>>(rp, ci, va, v, w, target, rows, ct, replicas, out_len); break;
    case 16: csr_row_kernel<16, MODE><<<g, 256, 0, ctx->stream>>>(rp, ci, va, v, w, target, rows, ct, replicas, out_len); break;
    default: csr_row_kernel<32, MODE><<<g, 256, 0, ctx->stream>>>(rp, ci, va, v, w, target, rows, ct, replicas, out_len); break;
  }
  }
  SYSML_LAUNCH_CHECK(ctx, "csr_row_kernel");
  if (replicas > 1) return sysml_reduce_partial_rows_fwd(ctx, target, replicas, (int)out_len, out);
  return SYSML_OK;
}

// row sums / sums of squares / max / min of a CSR matrix on the streamed kernel; *done = false when it does not apply
int launch_csr_row_agg_streamed(sysml_b200_ctx* ctx, const int* rp, const int* ci, const double* va, double* out, i64 rows, i64 cols,
                                i64 nnz, int op, bool* done) {
  const int rc = launch_csr_row<3>(ctx, rp, ci, va, nullptr, nullptr, out, rows, nnz, op, cols, 0);
  *done = (rc == SYSML_OK);
  return rc == SYSML_NOT_STREAMED ? SYSML_OK : rc;
}

int fetch_nnz(sysml_b200_ctx* ctx, const int32_t* rowptr, i64 rows, i64* nnz) {
  int last = 0;
  SYSML_CUDA(ctx, cudaMemcpyAsync(&last, rowptr + rows, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
  SYSML_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  *nnz = last;
  return SYSML_OK;
}

}  // namespace

extern "C" int sysml_b200_agg_unary_dense(sysml_b200_ctx* ctx, const double* X, int64_t rows, int64_t cols, int op, int dir, double* out);

extern "C" int sysml_b200_mmchain_csr(sysml_b200_ctx* ctx, const int32_t* rowptr, const int32_t* colind, const double* val,
                                      const double* v, const double* w, double* out, int64_t n, int64_t m, int chain_type) {
  SYSML_ENTER(ctx);
  if (chain_type < 0 || chain_type > 2) return sysml_set_error(ctx, SYSML_E_INVALID, "mmchain_csr: unknown chain type %d", chain_type);
  if (n < 0 || m < 0) return sysml_set_error(ctx, SYSML_E_DIM, "mmchain_csr: negative dimension");
  if (m == 0) return SYSML_OK;
  if (!out) return sysml_set_error(ctx, SYSML_E_INVALID, "mmchain_csr: out is null");
  SYSML_CUDA(ctx, cudaMemsetAsync(out, 0, (size_t)m * sizeof(double), ctx->stream));
  if (n == 0) return SYSML_OK;
  if (!rowptr || !v) return sysml_set_error(ctx, SYSML_E_INVALID, "mmchain_csr: null pointer");
  if (chain_type != SYSML_CHAIN_XtXv && !w) return sysml_set_error(ctx, SYSML_E_INVALID, "mmchain_csr: chain type %d needs w", chain_type);
  i64 nnz = 0;
  int rc = fetch_nnz(ctx, rowptr, n, &nnz);
  if (rc) return rc;
  if (nnz == 0) return SYSML_OK;
  return launch_csr_row<1>(ctx, rowptr, colind, val, v, w, out, n, nnz, chain_type, m, m);
}

extern "C" int sysml_b200_matmult_csr_dense(sysml_b200_ctx* ctx, const int32_t* rowptr, const int32_t* colind,
                                            const double* val, const double* B, double* C, int64_t a_rows, int64_t a_cols,
                                            int64_t b_rows, int64_t b_cols, int trans_a) {
  SYSML_ENTER(ctx);
  if (a_rows < 0 || a_cols < 0 || b_rows < 0 || b_cols < 0) return sysml_set_error(ctx, SYSML_E_DIM, "matmult_csr_dense: negative dimension");
  const i64 inner = trans_a ? a_rows : a_cols;
  const i64 m = trans_a ? a_cols : a_rows;
  if (inner != b_rows) return sysml_set_error(ctx, SYSML_E_DIM, "matmult_csr_dense: inner dimensions %lld vs %lld", (long long)inner, (long long)b_rows);
  const i64 n = b_cols;
  if (m * n == 0) return SYSML_OK;
  if (!C) return sysml_set_error(ctx, SYSML_E_INVALID, "matmult_csr_dense: C is null");
  if (inner == 0) {
    SYSML_CUDA(ctx, cudaMemsetAsync(C, 0, (size_t)(m * n) * sizeof(double), ctx->stream));
    return SYSML_OK;
  }
  if (!rowptr || !B) return sysml_set_error(ctx, SYSML_E_INVALID, "matmult_csr_dense: null pointer");
  i64 nnz = 0;
  int rc = fetch_nnz(ctx, rowptr, a_rows, &nnz);
  if (rc) return rc;
  if (trans_a) {
    SYSML_CUDA(ctx, cudaMemsetAsync(C, 0, (size_t)(m * n) * sizeof(double), ctx->stream));
    if (nnz == 0) return SYSML_OK;
    if (n == 1) return launch_csr_row<2>(ctx, rowptr, colind, val, nullptr, B, C, a_rows, nnz, 0, m, 0);
    csr_tspmm_kernel<<<warp_rows_grid(ctx, a_rows), 256, 0, ctx->stream>>>(rowptr, colind, val, B, C, a_rows, n);
    SYSML_LAUNCH_CHECK(ctx, "csr_tspmm_kernel");
    return SYSML_OK;
  }
  if (n == 1) return launch_csr_row<0>(ctx, rowptr, colind, val, B, nullptr, C, a_rows, nnz, 0, a_rows, a_cols);
  if (n <= 32) {
    csr_spmm_skinny_kernel<1><<<warp_rows_grid(ctx, a_rows), 256, 0, ctx->stream>>>(rowptr, colind, val, B, C, a_rows, (int)n);
  } else if (n <= 64) {
    csr_spmm_skinny_kernel<2><<<warp_rows_grid(ctx, a_rows), 256, 0, ctx->stream>>>(rowptr, colind, val, B, C, a_rows, (int)n);
  } else {
    const i64 ctiles = (n + 127) / 128;
    if (ctiles > 65535) return sysml_set_error(ctx, SYSML_E_LIMIT, "matmult_csr_dense: more than 8M output columns");
    i64 gx = (a_rows + 7) / 8;
    const i64 cap = ((i64)ctx->sm_count * 16 + ctiles - 1) / ctiles;
    if (gx > cap) gx = cap;
    if (gx < 1) gx = 1;
    csr_spmm_wide_kernel<<<dim3((unsigned)gx, (unsigned)ctiles), 256, 0, ctx->stream>>>(rowptr, colind, val, B, C, a_rows, n);
  }
  SYSML_LAUNCH_CHECK(ctx, "csr_spmm kernel");
  return SYSML_OK;
}

extern "C" int sysml_b200_tsmm_csr_left(sysml_b200_ctx* ctx, const int32_t* rowptr, const int32_t* colind, const double* val,
                                        double* out, int64_t rows, int64_t cols, int64_t* h_nnz_out) {
  SYSML_ENTER(ctx);
  if (rows < 0 || cols < 0) return sysml_set_error(ctx, SYSML_E_DIM, "tsmm_csr: negative dimension");
  if (h_nnz_out) *h_nnz_out = 0;
  if (cols == 0) return SYSML_OK;
  if (!out) return sysml_set_error(ctx, SYSML_E_INVALID, "tsmm_csr: out is null");
  SYSML_CUDA(ctx, cudaMemsetAsync(out, 0, (size_t)(cols * cols) * sizeof(double), ctx->stream));
  if (rows == 0) return SYSML_OK;
  if (!rowptr) return sysml_set_error(ctx, SYSML_E_INVALID, "tsmm_csr: rowptr is null");
  csr_tsmm_left_kernel<<<warp_rows_grid(ctx, rows), 256, 0, ctx->stream>>>(rowptr, colind, val, out, rows, cols);
  SYSML_LAUNCH_CHECK(ctx, "csr_tsmm_left_kernel");
  void* ws = nullptr;
  int rc = sysml_workspace(ctx, 256, &ws);
  if (rc) return rc;
  unsigned long long* d_nnz = (unsigned long long*)ws;
  SYSML_CUDA(ctx, cudaMemsetAsync(d_nnz, 0, sizeof(unsigned long long), ctx->stream));
  mirror_count_kernel<<<flat_grid(ctx, cols * cols, 256), 256, 0, ctx->stream>>>(out, cols, d_nnz);
  SYSML_LAUNCH_CHECK(ctx, "mirror_count_kernel");
  if (h_nnz_out) {
    unsigned long long h = 0;
    SYSML_CUDA(ctx, cudaMemcpyAsync(&h, d_nnz, sizeof(h), cudaMemcpyDeviceToHost, ctx->stream));
    SYSML_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    *h_nnz_out = (int64_t)h;
  }
  return SYSML_OK;
}

extern "C" int sysml_b200_count_nnz(sysml_b200_ctx* ctx, const double* X, int64_t len, int64_t* h_nnz) {
  SYSML_ENTER(ctx);
  if (len < 0 || !h_nnz || (len > 0 && !X)) return sysml_set_error(ctx, SYSML_E_INVALID, "count_nnz: bad arguments");
  *h_nnz = 0;
  if (len == 0) return SYSML_OK;
  void* ws = nullptr;
  int rc = sysml_workspace(ctx, 256, &ws);
  if (rc) return rc;
  unsigned long long* d = (unsigned long long*)ws;
  SYSML_CUDA(ctx, cudaMemsetAsync(d, 0, sizeof(unsigned long long), ctx->stream));
  count_nnz_kernel<<<flat_grid(ctx, len, 256), 256, 0, ctx->stream>>>(X, len, d);
  SYSML_LAUNCH_CHECK(ctx, "count_nnz_kernel");
  unsigned long long h = 0;
  SYSML_CUDA(ctx, cudaMemcpyAsync(&h, d, sizeof(h), cudaMemcpyDeviceToHost, ctx->stream));
  SYSML_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  *h_nnz = (int64_t)h;
  return SYSML_OK;
}

extern "C" int sysml_b200_dense_to_csr(sysml_b200_ctx* ctx, const double* X, int64_t rows, int64_t cols, int32_t* rowptr,
                                       int32_t* colind, double* val, int64_t* h_nnz) {
  SYSML_ENTER(ctx);
  if (rows < 0 || cols < 0 || !rowptr || !h_nnz) return sysml_set_error(ctx, SYSML_E_INVALID, "dense_to_csr: bad arguments");
  if (rows > 0 && cols > 0 && !X) return sysml_set_error(ctx, SYSML_E_INVALID, "dense_to_csr: X is null");
  void* ws = nullptr;
  int rc = sysml_workspace(ctx, 256, &ws);
  if (rc) return rc;
  unsigned long long* d_total = (unsigned long long*)ws;
  SYSML_CUDA(ctx, cudaMemsetAsync(rowptr, 0, (size_t)(rows + 1) * sizeof(int), ctx->stream));
  if (rows > 0 && cols > 0) {
    row_nnz_kernel<<<warp_rows_grid(ctx, rows), 256, 0, ctx->stream>>>(X, rows, cols, rowptr);
    SYSML_LAUNCH_CHECK(ctx, "row_nnz_kernel");
  }
  scan_rowptr_kernel<<<1, 1024, 0, ctx->stream>>>(rowptr, rows, d_total);
  SYSML_LAUNCH_CHECK(ctx, "scan_rowptr_kernel");
  unsigned long long h = 0;
  SYSML_CUDA(ctx, cudaMemcpyAsync(&h, d_total, sizeof(h), cudaMemcpyDeviceToHost, ctx->stream));
  SYSML_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  *h_nnz = (int64_t)h;
  if (h >= (1ull << 31)) return sysml_set_error(ctx, SYSML_E_LIMIT, "dense_to_csr: nnz %llu does not fit int32 row pointers", h);
  if (colind && val && h > 0) {
    fill_csr_kernel<<<warp_rows_grid(ctx, rows), 256, 0, ctx->stream>>>(X, rows, cols, rowptr, colind, val);
    SYSML_LAUNCH_CHECK(ctx, "fill_csr_kernel");
  }
  return SYSML_OK;
}

extern "C" int sysml_b200_csr_to_dense(sysml_b200_ctx* ctx, const int32_t* rowptr, const int32_t* colind, const double* val,
                                       int64_t rows, int64_t cols, double* X) {
  SYSML_ENTER(ctx);
  if (rows < 0 || cols < 0) return sysml_set_error(ctx, SYSML_E_DIM, "csr_to_dense: negative dimension");
  if (rows * cols == 0) return SYSML_OK;
  if (!rowptr || !X) return sysml_set_error(ctx, SYSML_E_INVALID, "csr_to_dense: null pointer");
  SYSML_CUDA(ctx, cudaMemsetAsync(X, 0, (size_t)(rows * cols) * sizeof(double), ctx->stream));
  csr_to_dense_kernel<<<warp_rows_grid(ctx, rows), 256, 0, ctx->stream>>>(rowptr, colind, val, rows, cols, X);
  SYSML_LAUNCH_CHECK(ctx, "csr_to_dense_kernel");
  return SYSML_OK;
}

extern "C" int sysml_b200_agg_unary_csr(sysml_b200_ctx* ctx, const int32_t* rowptr, const int32_t* colind, const double* val,
                                        int64_t rows, int64_t cols, int op, int dir, double* out) {
  SYSML_ENTER(ctx);
  if (rows < 0 || cols < 0) return sysml_set_error(ctx, SYSML_E_DIM, "agg_unary_csr: negative dimension");
  if (dir < 0 || dir > 2) return sysml_set_error(ctx, SYSML_E_INVALID, "agg_unary_csr: bad direction %d", dir);
  if (!out || !rowptr) return sysml_set_error(ctx, SYSML_E_INVALID, "agg_unary_csr: null pointer");
  if (rows * cols == 0) return sysml_set_error(ctx, SYSML_E_DIM, "agg_unary_csr: empty input");
  i64 nnz = 0;
  int rc = fetch_nnz(ctx, rowptr, rows, &nnz);
  if (rc) return rc;
  const bool is_sum = (op == SYSML_AGG_SUM || op == SYSML_AGG_SUMSQ);
  const bool is_mm = (op == SYSML_AGG_MAX || op == SYSML_AGG_MIN);
  const bool is_idx = (op == SYSML_AGG_MAXINDEX || op == SYSML_AGG_MININDEX);
  if (!is_sum && !is_mm && !is_idx) return sysml_set_error(ctx, SYSML_E_UNSUPPORTED, "agg_unary_csr: opcode %d on CSR input", op);
  if (is_idx && dir != SYSML_DIR_ROW) return sysml_set_error(ctx, SYSML_E_UNSUPPORTED, "rowIndexMax/Min exist only per row");
  if (dir == SYSML_DIR_ROW) {
    if (!is_idx && nnz > 0) {  // big matrices with short rows: streamed (falls through when not eligible)
      bool done = false;
      rc = launch_csr_row_agg_streamed(ctx, rowptr, colind, val, out, rows, cols, nnz, op, &done);
      if (rc || done) return rc;
    }
    const int L = pick_lanes(nnz, rows);
    const unsigned g = rows_grid(ctx, rows, L, 256);
    switch (L) {
      case 2: csr_agg_row_kernel<2><<<g, 256, 0, ctx->stream>>>(rowptr, colind, val, rows, cols, op, out); break;
      case 4: csr_agg_row_kernel<4><<<g, 256, 0, ctx->stream>>>(rowptr, colind, val, rows, cols, op, out); break;
      case 8: csr_agg_row_kernel<8><<<g, 256, 0, ctx->stream>>>(rowptr, colind, val, rows, cols, op, out); break;
      case 16: csr_agg_row_kernel<16><<<g, 256, 0, ctx->stream>>>(rowptr, colind, val, rows, cols, op, out); break;
      default: csr_agg_row_kernel<32><<<g, 256, 0, ctx->stream>>>(rowptr, colind, val, rows, cols, op, out); break;
    }
    SYSML_LAUNCH_CHECK(ctx, "csr_agg_row_kernel");
    return SYSML_OK;
  }
  if (dir == SYSML_DIR_ALL) {
    if (is_sum) {
      if (nnz == 0) {
        SYSML_CUDA(ctx, cudaMemsetAsync(out, 0, sizeof(double), ctx->stream));
        return SYSML_OK;
      }
      return sysml_b200_agg_unary_dense(ctx, val, 1, nnz, op, SYSML_DIR_ALL, out);  // s_uakp: sum over values()
    }
    const int nparts = (int)flat_grid(ctx, nnz > 0 ? nnz : 1, 256);
    void* ws = nullptr;
    rc = sysml_workspace(ctx, (size_t)nparts * sizeof(double), &ws);
    if (rc) return rc;
    csr_agg_all_minmax_kernel<<<nparts, 256, 0, ctx->stream>>>(val, nnz, op == SYSML_AGG_MAX, (double*)ws);
    SYSML_LAUNCH_CHECK(ctx, "csr_agg_all_minmax_kernel");
    csr_agg_all_minmax_finish<<<1, 32, 0, ctx->stream>>>((double*)ws, nparts, op == SYSML_AGG_MAX, nnz < rows * cols, out);
    SYSML_LAUNCH_CHECK(ctx, "csr_agg_all_minmax_finish");
    return SYSML_OK;
  }
  // COL
  if (is_sum) {
    SYSML_CUDA(ctx, cudaMemsetAsync(out, 0, (size_t)cols * sizeof(double), ctx->stream));
    if (nnz == 0) return SYSML_OK;
    csr_agg_col_sum_kernel<<<flat_grid(ctx, nnz, 256), 256, 0, ctx->stream>>>(colind, val, nnz, op == SYSML_AGG_SUMSQ, out);
    SYSML_LAUNCH_CHECK(ctx, "csr_agg_col_sum_kernel");
    return SYSML_OK;
  }
  void* ws = nullptr;
  rc = sysml_workspace(ctx, (size_t)cols * sizeof(int), &ws);
  if (rc) return rc;
  SYSML_CUDA(ctx, cudaMemsetAsync(ws, 0, (size_t)cols * sizeof(int), ctx->stream));
  {
    // initialise the output with -+Inf
    rc = sysml_b200_fill(ctx, out, cols, op == SYSML_AGG_MAX ? -HUGE_VAL : HUGE_VAL);
    if (rc) return rc;
  }
  if (nnz > 0) {
    csr_agg_col_minmax_kernel<<<flat_grid(ctx, nnz, 256), 256, 0, ctx->stream>>>(colind, val, nnz, op == SYSML_AGG_MAX, out, (int*)ws);
    SYSML_LAUNCH_CHECK(ctx, "csr_agg_col_minmax_kernel");
  }
  csr_agg_col_minmax_finish<<<(unsigned)((cols + 255) / 256), 256, 0, ctx->stream>>>(out, (int*)ws, cols, rows, op == SYSML_AGG_MAX);
  SYSML_LAUNCH_CHECK(ctx, "csr_agg_col_minmax_finish");
  return SYSML_OK;
}

extern "C" int sysml_b200_transpose(sysml_b200_ctx* ctx, const double* A, double* C, int64_t rows, int64_t cols);
extern "C" int sysml_b200_tsmm_dense(sysml_b200_ctx* ctx, const double* X, double* out, int64_t rows, int64_t cols, int left, int64_t* h_nnz_out);

extern "C" int sysml_b200_matmult_dense_csr(sysml_b200_ctx* ctx, const double* A, const int32_t* rowptr, const int32_t* colind,
                                            const double* val, double* C, int64_t a_rows, int64_t a_cols, int64_t b_rows,
                                            int64_t b_cols, int trans_a) {
  SYSML_ENTER(ctx);
  if (a_rows < 0 || a_cols < 0 || b_rows < 0 || b_cols < 0) return sysml_set_error(ctx, SYSML_E_DIM, "matmult_dense_csr: negative dimension");
  const i64 m = trans_a ? a_cols : a_rows, k = trans_a ? a_rows : a_cols;
  if (k != b_rows) return sysml_set_error(ctx, SYSML_E_DIM, "matmult_dense_csr: inner dimensions %lld vs %lld", (long long)k, (long long)b_rows);
  const i64 n = b_cols;
  if (m * n == 0) return SYSML_OK;
  if (!C) return sysml_set_error(ctx, SYSML_E_INVALID, "matmult_dense_csr: C is null");
  if (k == 0) {
    SYSML_CUDA(ctx, cudaMemsetAsync(C, 0, (size_t)(m * n) * sizeof(double), ctx->stream));
    return SYSML_OK;
  }
  if (!A || !rowptr) return sysml_set_error(ctx, SYSML_E_INVALID, "matmult_dense_csr: null operand");
  const double* Ause = A;
  if (trans_a) {  // materialise t(A) once in scratch (the reference's r' + ba+*, LibMatrixCUDA.java:1723-1729)
    void* ws = nullptr;
    int rc = sysml_workspace(ctx, (size_t)(m * k) * sizeof(double), &ws);
    if (rc) return rc;
    rc = sysml_b200_transpose(ctx, A, (double*)ws, a_rows, a_cols);
    if (rc) return rc;
    Ause = (const double*)ws;
  }
  dense_csr_mm_kernel<<<warp_rows_grid(ctx, m), 256, 0, ctx->stream>>>(Ause, rowptr, colind, val, C, m, k, n);
  SYSML_LAUNCH_CHECK(ctx, "dense_csr_mm_kernel");
  return SYSML_OK;
}

extern "C" int sysml_b200_matmult_csr_csr(sysml_b200_ctx* ctx, const int32_t* a_rowptr, const int32_t* a_colind, const double* a_val,
                                          const int32_t* b_rowptr, const int32_t* b_colind, const double* b_val, double* C,
                                          int64_t a_rows, int64_t a_cols, int64_t b_rows, int64_t b_cols) {
  SYSML_ENTER(ctx);
  if (a_rows < 0 || a_cols < 0 || b_rows < 0 || b_cols < 0) return sysml_set_error(ctx, SYSML_E_DIM, "matmult_csr_csr: negative dimension");
  if (a_cols != b_rows) return sysml_set_error(ctx, SYSML_E_DIM, "matmult_csr_csr: inner dimensions %lld vs %lld", (long long)a_cols, (long long)b_rows);
  if (a_rows * b_cols == 0) return SYSML_OK;
  if (!C || !a_rowptr || !b_rowptr) return sysml_set_error(ctx, SYSML_E_INVALID, "matmult_csr_csr: null pointer");
  csr_csr_mm_kernel<<<warp_rows_grid(ctx, a_rows), 256, 0, ctx->stream>>>(a_rowptr, a_colind, a_val, b_rowptr, b_colind, b_val, C, a_rows, b_cols);
  SYSML_LAUNCH_CHECK(ctx, "csr_csr_mm_kernel");
  return SYSML_OK;
}

// X %*% t(X) for CSR X: the reference transposes first and runs the LEFT kernel on t(X)
// (prepMatrixMultTransposeSelfInput, LibMatrixMult.java:3679-3685); here the (rows x rows) result is
// formed from a dense copy of X in scratch, which is only sensible while rows*cols fits comfortably.
extern "C" int sysml_b200_tsmm_csr_right(sysml_b200_ctx* ctx, const int32_t* rowptr, const int32_t* colind, const double* val,
                                         double* out, int64_t rows, int64_t cols, int64_t* h_nnz_out) {
  SYSML_ENTER(ctx);
  if (rows < 0 || cols < 0) return sysml_set_error(ctx, SYSML_E_DIM, "tsmm_csr_right: negative dimension");
  if (h_nnz_out) *h_nnz_out = 0;
  if (rows == 0) return SYSML_OK;
  if (!out || !rowptr) return sysml_set_error(ctx, SYSML_E_INVALID, "tsmm_csr_right: null pointer");
  if (cols == 0) {
    SYSML_CUDA(ctx, cudaMemsetAsync(out, 0, (size_t)(rows * rows) * sizeof(double), ctx->stream));
    return SYSML_OK;
  }
  if ((double)rows * (double)cols * 8.0 > 32e9) return sysml_set_error(ctx, SYSML_E_LIMIT, "tsmm_csr_right: dense copy of X would need more than 32 GB");
  double* dense = nullptr;
  SYSML_CUDA(ctx, cudaMalloc((void**)&dense, (size_t)(rows * cols) * sizeof(double)));
  int rc = sysml_b200_csr_to_dense(ctx, rowptr, colind, val, rows, cols, dense);
  if (!rc) rc = sysml_b200_tsmm_dense(ctx, dense, out, rows, cols, 0, h_nnz_out);
  cudaStreamSynchronize(ctx->stream);
  cudaFree(dense);
  return rc;
}
