// agg.cu -- aggregate-unary reductions over a dense row-major matrix: sum, sum of squares, max, min,
// product, mean in the three directions (ALL / per-ROW / per-COLUMN) and rowIndexMax / rowIndexMin.
//
// Reference semantics: LibMatrixAgg.aggregateUnaryMatrix (LibMatrixAgg.java:194-282) with the dense
// kernels d_uakp/d_uarkp/d_uackp (:1624-1669), d_ua*sqkp (:1685-1750), d_uamxx/d_uarmxx/d_uacmxx
// (:1843-1889), d_uarimxx/d_uarimin (:1902-1927, indexmax/indexmin :3081-3099), d_uam* (:2100-2135),
// d_ua*mean (:1940-2010).  GPU twin being replaced: LibMatrixCUDA.unaryAggregate
// (LibMatrixCUDA.java:573-894; kernels reduce / reduce_row / reduce_col, SystemML.cu:916-1540).
//
// The reference sums with Kahan compensation in strict element order.  A parallel reduction cannot
// keep that order, so every partial here is carried as an error-free (hi, lo) pair (two-sum) and the
// pairs are merged in a fixed order: the result is deterministic and tracks the Kahan sum to ~1 ulp.
// +-Inf inputs give +-Inf like KahanPlus' fast path (functionobjects/KahanPlus.java:92-95); a mix of
// +Inf and -Inf gives NaN here whereas the reference returns the LAST infinity seen.
//
// All kernels are HBM-bound: algorithmic bytes = 8*rows*cols + output.
#include <math_constants.h>

#include "common.cuh"

namespace {

enum { K_SUM = 0, K_SUMSQ = 1, K_MAX = 2, K_MIN = 3, K_PROD = 7 };

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

// Accumulator abstraction: OP-specific state, element update, merge and finalisation.
template <int OP>
struct Acc {
  double hi, lo;
  __device__ __forceinline__ void init() {
    hi = (OP == K_MAX) ? -CUDART_INF : (OP == K_MIN) ? CUDART_INF : (OP == K_PROD) ? 1.0 : 0.0;
    lo = 0.0;
  }
  __device__ __forceinline__ void push(double x) {
    if (OP == K_SUM || OP == K_SUMSQ) {
      const double v = (OP == K_SUMSQ) ? __dmul_rn(x, x) : x;
      double s, e;
      two_sum(hi, v, s, e);
      hi = s;
      lo = __dadd_rn(lo, e);
    } else if (OP == K_MAX) {
      hi = jmax(hi, x);
    } else if (OP == K_MIN) {
      hi = jmin(hi, x);
    } else {
      hi = __dmul_rn(hi, x);
    }
  }
  __device__ __forceinline__ void merge(double ohi, double olo) {
    if (OP == K_SUM || OP == K_SUMSQ) {
      double s, e;
      two_sum(hi, ohi, s, e);
      hi = s;
      lo = __dadd_rn(__dadd_rn(lo, olo), e);
    } else if (OP == K_MAX) {
      hi = jmax(hi, ohi);
    } else if (OP == K_MIN) {
      hi = jmin(hi, ohi);
    } else {
      hi = __dmul_rn(hi, ohi);
    }
  }
  __device__ __forceinline__ double result() const {
    if (OP == K_SUM || OP == K_SUMSQ) {
      // an infinite hi leaves NaN in lo (inf - inf inside two_sum): the sum is the infinity itself
      return (isinf(hi) || lo != lo) ? hi : __dadd_rn(hi, lo);
    }
    return hi;
  }
};

template <int OP>
__device__ __forceinline__ void warp_merge(Acc<OP>& a) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const double ohi = __shfl_xor_sync(0xffffffffu, a.hi, o);
    const double olo = __shfl_xor_sync(0xffffffffu, a.lo, o);
    a.merge(ohi, olo);
  }
}
// merge within aligned lane groups of width W (power of two <= 32)
template <int OP>
__device__ __forceinline__ void group_merge(Acc<OP>& a, int W) {
  for (int o = W >> 1; o > 0; o >>= 1) {
    const double ohi = __shfl_xor_sync(0xffffffffu, a.hi, o);
    const double olo = __shfl_xor_sync(0xffffffffu, a.lo, o);
    a.merge(ohi, olo);
  }
}

// ---------------- ALL: grid-stride + last-block finish ----------------
template <int OP>
__global__ void __launch_bounds__(512) agg_all_kernel(const double* __restrict__ X, i64 len, int vec2, double* __restrict__ partial,
                                                      unsigned int* counter, double* __restrict__ out, double scale) {
  Acc<OP> a[4];
#pragma unroll
  for (int q = 0; q < 4; q++) a[q].init();
  const i64 stride = (i64)gridDim.x * blockDim.x;
  const i64 tid = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (vec2) {
    const double2* X2 = reinterpret_cast<const double2*>(X);
    const i64 n2 = len >> 1;
    i64 i = tid;
    for (; i + stride < n2; i += 2 * stride) {
      const double2 u = ldg_stream2(X2 + i), w = ldg_stream2(X2 + i + stride);
      a[0].push(u.x);
      a[1].push(u.y);
      a[2].push(w.x);
      a[3].push(w.y);
    }
    if (i < n2) {
      const double2 u = ldg_stream2(X2 + i);
      a[0].push(u.x);
      a[1].push(u.y);
    }
    if (tid == 0 && (len & 1)) a[2].push(X[len - 1]);
  } else {
    for (i64 i = tid; i < len; i += stride) a[0].push(ldg_stream(X + i));
  }
  a[0].merge(a[1].hi, a[1].lo);
  a[2].merge(a[3].hi, a[3].lo);
  a[0].merge(a[2].hi, a[2].lo);
  warp_merge(a[0]);
  __shared__ double sh[2][32];
  __shared__ unsigned int s_ticket;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  if (lane == 0) {
    sh[0][warp] = a[0].hi;
    sh[1][warp] = a[0].lo;
  }
  __syncthreads();
  if (warp == 0) {
    Acc<OP> b;
    b.init();
    if (lane < nwarps) {
      b.hi = sh[0][lane];
      b.lo = sh[1][lane];
    }
    warp_merge(b);
    if (lane == 0) {
      partial[2 * blockIdx.x] = b.hi;
      partial[2 * blockIdx.x + 1] = b.lo;
      __threadfence();
      s_ticket = atomicAdd(counter, 1u);
    }
  }
  __syncthreads();
  if (s_ticket != gridDim.x - 1) return;
  __threadfence();
  if (warp == 0) {
    Acc<OP> b;
    b.init();
    for (int i = lane; i < (int)gridDim.x; i += 32) b.merge(__ldcg(partial + 2 * i), __ldcg(partial + 2 * i + 1));
    warp_merge(b);
    if (lane == 0) {
      double r = b.result();
      if (scale != 1.0) r = __ddiv_rn(r, scale);
      out[0] = r;
      *counter = 0u;
    }
  }
}

// ---------------- ROW: W lanes per row (W = 1..32, power of two) ----------------
template <int OP>
__global__ void agg_row_kernel(const double* __restrict__ X, i64 rows, i64 cols, int W, double* __restrict__ out, double scale) {
  const int lane = threadIdx.x & 31;
  const int sub = lane & (W - 1);
  const int rows_per_warp = 32 / W;
  const i64 warp = ((i64)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const i64 nwarps = ((i64)gridDim.x * blockDim.x) >> 5;
  for (i64 base = warp * rows_per_warp; base < rows; base += nwarps * rows_per_warp) {
    const i64 r = base + lane / W;
    Acc<OP> a, b;
    a.init();
    b.init();
    if (r < rows) {
      const double* row = X + r * cols;
      i64 j = sub;
      // four independent loads in flight per lane, two accumulator chains
      for (; j + 3 * W < cols; j += 4 * W) {
        const double x0 = ldg_stream(row + j), x1 = ldg_stream(row + j + W);
        const double x2 = ldg_stream(row + j + 2 * W), x3 = ldg_stream(row + j + 3 * W);
        a.push(x0);
        b.push(x1);
        a.push(x2);
        b.push(x3);
      }
      for (; j < cols; j += W) a.push(ldg_stream(row + j));
      a.merge(b.hi, b.lo);
    }
    group_merge(a, W);
    if (r < rows && sub == 0) {
      double v = a.result();
      if (scale != 1.0) v = __ddiv_rn(v, scale);
      out[r] = v;
    }
  }
}

// rowIndexMax / rowIndexMin: >= / <= so the LAST extremum wins; NaN never wins; all-NaN row -> 0
template <int ISMAX>
__global__ void agg_rowindex_kernel(const double* __restrict__ X, i64 rows, i64 cols, int W, double* __restrict__ out) {
  const int lane = threadIdx.x & 31;
  const int sub = lane & (W - 1);
  const int rows_per_warp = 32 / W;
  const i64 warp = ((i64)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const i64 nwarps = ((i64)gridDim.x * blockDim.x) >> 5;
  for (i64 base = warp * rows_per_warp; base < rows; base += nwarps * rows_per_warp) {
    const i64 r = base + lane / W;
    double best = ISMAX ? -CUDART_INF : CUDART_INF;
    i64 idx = -1;
    if (r < rows) {
      const double* row = X + r * cols;
      for (i64 j = sub; j < cols; j += W) {
        const double x = ldg_stream(row + j);
        const bool take = ISMAX ? (x >= best) : (x <= best);
        idx = take ? j : idx;
        best = take ? x : best;
      }
    }
    for (int o = W >> 1; o > 0; o >>= 1) {
      const double ob = __shfl_xor_sync(0xffffffffu, best, o);
      const i64 oi = __shfl_xor_sync(0xffffffffu, idx, o);
      // a candidate exists only if idx >= 0; ties go to the larger column index
      const bool other_better = (oi >= 0) && ((idx < 0) || (ISMAX ? (ob > best) : (ob < best)) || (ob == best && oi > idx));
      best = other_better ? ob : best;
      idx = other_better ? oi : idx;
    }
    if (r < rows && sub == 0) out[r] = (double)(idx + 1);
  }
}

// ---------------- COL: block = CPB columns x RPB row-lanes; grid = col tiles x row splits ----------------
template <int OP>
__global__ void __launch_bounds__(256) agg_col_kernel(const double* __restrict__ X, i64 rows, i64 cols, int CPB, int RPB,
                                                      double* __restrict__ partial) {
  extern __shared__ double sh[];  // [2][RPB][CPB]
  const int t = threadIdx.x;
  const int cl = t % CPB, rl = t / CPB;
  const i64 c = (i64)blockIdx.x * CPB + cl;
  const bool active = (rl < RPB) && (c < cols);
  Acc<OP> a0, a1;
  a0.init();
  a1.init();
  if (active) {
    const i64 rstep = (i64)RPB * gridDim.y;
    i64 r = (i64)blockIdx.y * RPB + rl;
    Acc<OP> a2, a3;
    a2.init();
    a3.init();
    for (; r + 3 * rstep < rows; r += 4 * rstep) {
      const double u0 = ldg_stream(X + r * cols + c);
      const double u1 = ldg_stream(X + (r + rstep) * cols + c);
      const double u2 = ldg_stream(X + (r + 2 * rstep) * cols + c);
      const double u3 = ldg_stream(X + (r + 3 * rstep) * cols + c);
      a0.push(u0);
      a1.push(u1);
      a2.push(u2);
      a3.push(u3);
    }
    for (; r < rows; r += rstep) a0.push(ldg_stream(X + r * cols + c));
    a0.merge(a1.hi, a1.lo);
    a2.merge(a3.hi, a3.lo);
    a0.merge(a2.hi, a2.lo);
  }
  if (rl < RPB) {
    sh[rl * CPB + cl] = a0.hi;
    sh[(RPB + rl) * CPB + cl] = a0.lo;
  }
  __syncthreads();
  if (rl == 0 && c < cols) {
    for (int q = 1; q < RPB; q++) a0.merge(sh[q * CPB + cl], sh[(RPB + q) * CPB + cl]);
    partial[((size_t)blockIdx.y * 2) * cols + c] = a0.hi;
    partial[((size_t)blockIdx.y * 2 + 1) * cols + c] = a0.lo;
  }
}
template <int OP>
__global__ void __launch_bounds__(256) agg_col_finish_kernel(const double* __restrict__ partial, int splits, i64 cols, double* __restrict__ out, double scale) {
  __shared__ double sh_hi[8][33], sh_lo[8][33];
  const int cl = threadIdx.x & 31, sl = threadIdx.x >> 5;
  const i64 c = (i64)blockIdx.x * 32 + cl;
  Acc<OP> a, b;
  a.init();
  b.init();
  if (c < cols) {
    int s = sl;
    for (; s + 8 < splits; s += 16) {
      const double h0 = __ldcg(partial + ((size_t)s * 2) * cols + c), l0 = __ldcg(partial + ((size_t)s * 2 + 1) * cols + c);
      const double h1 = __ldcg(partial + ((size_t)(s + 8) * 2) * cols + c), l1 = __ldcg(partial + ((size_t)(s + 8) * 2 + 1) * cols + c);
      a.merge(h0, l0);
      b.merge(h1, l1);
    }
    for (; s < splits; s += 8) a.merge(__ldcg(partial + ((size_t)s * 2) * cols + c), __ldcg(partial + ((size_t)s * 2 + 1) * cols + c));
    a.merge(b.hi, b.lo);
  }
  sh_hi[sl][cl] = a.hi;
  sh_lo[sl][cl] = a.lo;
  __syncthreads();
  if (sl == 0 && c < cols) {
    for (int q = 1; q < 8; q++) a.merge(sh_hi[q][cl], sh_lo[q][cl]);
    double v = a.result();
    if (scale != 1.0) v = __ddiv_rn(v, scale);
    out[c] = v;
  }
}

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }
// lanes cooperating on one row: about 8 elements per lane, so narrow matrices put several rows in a warp
inline int lanes_per_row(i64 cols) {
  int W = 1;
  while (W < 32 && (i64)W * 8 < cols) W <<= 1;
  return W;
}


// ------------------------------------------------------------------------------------------------
// Streamed variants for matrices with short rows (cols <= 1024): a tile of R full rows is ONE contiguous
// span, so a producer warp feeds a 3-deep shared-memory ring with 1-D bulk copies (cp.async.bulk, the
// TMA engine) exactly like the mmchain kernel; bytes in flight no longer depend on how many loads the
// reducing threads can issue, which is what held the direct-load kernels at ~50 % of HBM bandwidth on
// 10M x 50 (ncu: long-scoreboard stalls, 37-43 % DRAM utilisation).
//   ROWS : L lanes per row reduce out of shared memory (sum / sumsq / min / max / prod, rowIndexMax/Min)
//   COLS : thread (column, row-lane) accumulators live in registers across all tiles of the CTA
// ------------------------------------------------------------------------------------------------
constexpr int AS_THREADS = 256;  // consumer threads (+32 producer)
constexpr int AS_STAGES = 3;

// Per-row accumulator of the streamed kernel.  Sums/products reuse Acc; max/min use a branch-free form: the
// loop keeps a plain compare-select plus two sticky flags (a NaN was seen; a zero of the winning sign was seen)
// and the Math.max/min corner cases (NaN propagates, -0.0 < +0.0; Builtin.java:356-361) are settled once per row.
template <int OP>
struct RowAcc {
  Acc<OP> a;
  __device__ __forceinline__ void init() { a.init(); }
  __device__ __forceinline__ void push(double x) { a.push(x); }
  __device__ __forceinline__ RowAcc from_lane_xor(int o) const {
    RowAcc r;
    r.a.hi = __shfl_xor_sync(0xffffffffu, a.hi, o);
    r.a.lo = __shfl_xor_sync(0xffffffffu, a.lo, o);
    return r;
  }
  __device__ __forceinline__ void merge(const RowAcc& b) { a.merge(b.a.hi, b.a.lo); }
  __device__ __forceinline__ double result() const { return a.result(); }
};
template <bool ISMAX>
struct RowAccMinMax {
  double v;
  int fl;  // bit 0: NaN seen, bit 1: +0.0 (max) / -0.0 (min) seen
  __device__ __forceinline__ void init() {
    v = ISMAX ? -CUDART_INF : CUDART_INF;
    fl = 0;
  }
  __device__ __forceinline__ void push(double x) {
    const long long b = __double_as_longlong(x);
    fl |= (x != x) ? 1 : 0;
    fl |= (b == (ISMAX ? 0ll : (long long)0x8000000000000000ull)) ? 2 : 0;
    v = (ISMAX ? (x > v) : (x < v)) ? x : v;
  }
  __device__ __forceinline__ RowAccMinMax from_lane_xor(int o) const {
    RowAccMinMax r;
    r.v = __shfl_xor_sync(0xffffffffu, v, o);
    r.fl = __shfl_xor_sync(0xffffffffu, fl, o);
    return r;
  }
  __device__ __forceinline__ void merge(const RowAccMinMax& b) {
    fl |= b.fl;
    v = (ISMAX ? (b.v > v) : (b.v < v)) ? b.v : v;
  }
  __device__ __forceinline__ double result() const {
    if (fl & 1) return CUDART_NAN;
    if (v == 0.0 && (fl & 2)) return ISMAX ? 0.0 : -0.0;
    return v;
  }
};
template <>
struct RowAcc<K_MAX> : RowAccMinMax<true> {};
template <>
struct RowAcc<K_MIN> : RowAccMinMax<false> {};

// Reduces two per-lane accumulators (rows r0, r1 of an L-lane group, L >= 2) with half the shuffles of two
// separate butterflies: in the first step even lanes keep row r0 and odd lanes row r1, each receiving the
// partner's share of the row it keeps; the remaining steps stay within lanes of equal parity.  Afterwards
// x0 of lane sub==0 is row r0 and x0 of lane sub==1 is row r1.
template <class A>
__device__ __forceinline__ void pair_group_merge(A& x0, A& x1, int lane, int L) {
  const bool odd = lane & 1;
  A keep = odd ? x1 : x0;
  const A send = odd ? x0 : x1;
  keep.merge(send.from_lane_xor(1));
  for (int o = L >> 1; o > 1; o >>= 1) keep.merge(keep.from_lane_xor(o));
  x0 = keep;
}

// kept out of line: the compiler otherwise if-converts the rarely taken division (mean only) into every row
__device__ __noinline__ double div_by_scale(double v, double s) { return __ddiv_rn(v, s); }

__device__ __forceinline__ double lds_f64(uint32_t addr) {
  double v;
  asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(addr));
  return v;
}

struct AggStreamArgs {
  const double* X;
  i64 rows;
  int cols;
  int R;            // rows per tile
  int use_tma;
  i64 total_tiles;
  double* out;      // ROWS: rows results; COLS: unused
  double* partial;  // COLS: [gridDim.x][2][cols]
  double scale;
  int L;            // ROWS: lanes per row
};

template <int OP, int KIND, int KC>  // KIND 0: per-row value, 1: per-row index (OP = K_MAX / K_MIN), 2: per-column (KC column chunks per thread)
__global__ void __launch_bounds__(AS_THREADS + 32) aggstream_kernel(const AggStreamArgs a) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  uint64_t* full = reinterpret_cast<uint64_t*>(smem_raw);
  uint64_t* empty = full + AS_STAGES;
  double* tiles = reinterpret_cast<double*>(smem_raw + 128);
  const int tid = threadIdx.x;
  const int cols = a.cols, R = a.R;
  const size_t stage_elems = (size_t)R * cols;
  if (tid == 0) {
    for (int s = 0; s < AS_STAGES; s++) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], AS_THREADS);
    }
    mbar_fence_init();
  }
  __syncthreads();

  if (tid >= AS_THREADS) {  // producer warp
    const int lane = tid - AS_THREADS;
    const uint64_t pol = l2_policy_evict_first();
    int it = 0;
    for (i64 tile = blockIdx.x; tile < a.total_tiles; tile += gridDim.x, ++it) {
      const int s = it % AS_STAGES;
      mbar_wait(&empty[s], ((uint32_t)(it / AS_STAGES) & 1u) ^ 1u);
      const i64 r0 = tile * R;
      const i64 nr = (a.rows - r0 < R) ? (a.rows - r0) : R;
      double* dst = tiles + (size_t)s * stage_elems;
      const double* src = a.X + r0 * (i64)cols;
      if (a.use_tma && nr == R) {
        if (lane == 0) {
          const uint32_t bytes = (uint32_t)(stage_elems * sizeof(double));
          mbar_expect_tx(&full[s], bytes);
          bulk_g2s(dst, src, bytes, &full[s], pol);
        }
      } else {
        const i64 valid = nr * (i64)cols;  // rows past the end are never read by the consumers
        for (i64 i = lane; i < valid; i += 32) dst[i] = ldg_stream(src + i);
        __syncwarp();
        if (lane == 0) mbar_arrive(&full[s]);
      }
    }
    return;
  }

  const int lane = tid & 31, warp = tid >> 5;
  if (KIND == 2) {
    // ---- per-column: thread owns column chunk(s) cl + CPB*k, row lane rl ----
    const int CPB = cols < AS_THREADS ? cols : AS_THREADS;
    const int RPB = AS_THREADS / CPB;
    const int cl = tid % CPB, rl = tid / CPB;
    Acc<OP> acc[KC], acc2[KC];  // two chains per column: even / odd rows of the thread
#pragma unroll
    for (int k = 0; k < KC; k++) {
      acc[k].init();
      acc2[k].init();
    }
    int it = 0;
    for (i64 tile = blockIdx.x; tile < a.total_tiles; tile += gridDim.x, ++it) {
      const int s = it % AS_STAGES;
      const i64 r0 = tile * R;
      const int nr = (int)((a.rows - r0 < R) ? (a.rows - r0) : R);
      mbar_wait(&full[s], (uint32_t)(it / AS_STAGES) & 1u);
      const double* T = tiles + (size_t)s * stage_elems;
      if (rl < RPB) {
        int rr = rl;
        for (; rr + RPB < nr; rr += 2 * RPB) {
#pragma unroll
          for (int k = 0; k < KC; k++) {
            const int c = cl + CPB * k;
            if (KC == 1 || c < cols) {
              const double u = T[rr * cols + c], w = T[(rr + RPB) * cols + c];
              acc[k].push(u);
              acc2[k].push(w);
            }
          }
        }
        if (rr < nr) {
#pragma unroll
          for (int k = 0; k < KC; k++) {
            const int c = cl + CPB * k;
            if (KC == 1 || c < cols) acc[k].push(T[rr * cols + c]);
          }
        }
      }
      mbar_arrive(&empty[s]);
    }
#pragma unroll
    for (int k = 0; k < KC; k++) acc[k].merge(acc2[k].hi, acc2[k].lo);
    // combine the RPB row lanes through shared memory (reuse stage 0 after everyone is done with it)
    named_bar_sync(1, AS_THREADS);
    double* sh = tiles;  // [2][RPB][cols]
    if (rl < RPB) {
#pragma unroll
      for (int k = 0; k < KC; k++) {
        const int c = cl + CPB * k;
        if (c < cols) {
          sh[(size_t)rl * cols + c] = acc[k].hi;
          sh[(size_t)(RPB + rl) * cols + c] = acc[k].lo;
        }
      }
    }
    named_bar_sync(1, AS_THREADS);
    if (rl == 0) {
#pragma unroll
      for (int k = 0; k < KC; k++) {
        const int c = cl + CPB * k;
        if (c < cols) {
          for (int q = 1; q < RPB; q++) acc[k].merge(sh[(size_t)q * cols + c], sh[(size_t)(RPB + q) * cols + c]);
          a.partial[((size_t)blockIdx.x * 2) * cols + c] = acc[k].hi;
          a.partial[((size_t)blockIdx.x * 2 + 1) * cols + c] = acc[k].lo;
        }
      }
    }
    return;
  }

  // ---- per-row: L lanes per row ----
  const int L = a.L;
  const int sub = lane & (L - 1);
  const int rpw = 32 / L;                      // rows per warp pass
  const int rows_per_pass = rpw * (AS_THREADS / 32);
  const int n_it = (cols - sub + L - 1) / L;  // elements of a row this lane visits
  const uint32_t step = (uint32_t)L * 8u;
  const bool scaled = a.scale != 1.0;
  int it = 0;
  for (i64 tile = blockIdx.x; tile < a.total_tiles; tile += gridDim.x, ++it) {
    const int s = it % AS_STAGES;
    const i64 r0 = tile * R;
    const int nr = (int)((a.rows - r0 < R) ? (a.rows - r0) : R);
    mbar_wait(&full[s], (uint32_t)(it / AS_STAGES) & 1u);
    const double* T = tiles + (size_t)s * stage_elems;
    for (int rb = warp * rpw; rb < nr; rb += 2 * rows_per_pass) {
      // two rows per lane group per pass (independent chains hide the FP64 latency)
      const int rr0 = rb + lane / L, rr1 = rr0 + rows_per_pass;
      const bool live0 = rr0 < nr, live1 = rr1 < nr;
      const double* row0 = T + (size_t)(live0 ? rr0 : 0) * cols;
      const double* row1 = T + (size_t)(live1 ? rr1 : 0) * cols;
      if (KIND == 0) {
        // shared-window addresses and the trip count are hoisted: the loop body is two LDS and the two updates
        // (rows past the end of the tile read row 0 and are discarded)
        RowAcc<OP> x0, x1;
        x0.init();
        x1.init();
        uint32_t p0 = smem_u32(row0 + sub), p1 = smem_u32(row1 + sub);
#pragma unroll 2
        for (int i = 0; i < n_it; i++, p0 += step, p1 += step) {
          x0.push(lds_f64(p0));
          x1.push(lds_f64(p1));
        }
        if (L >= 2) {
          pair_group_merge(x0, x1, lane, L);
          const bool mine = (sub == 0) ? live0 : live1;
          if (sub < 2 && mine) {
            double v = x0.result();
            if (scaled) v = div_by_scale(v, a.scale);
            a.out[r0 + (sub == 0 ? rr0 : rr1)] = v;
          }
        } else {
          if (live0) {
            double v = x0.result();
            if (scaled) v = div_by_scale(v, a.scale);
            a.out[r0 + rr0] = v;
          }
          if (live1) {
            double v = x1.result();
            if (scaled) v = div_by_scale(v, a.scale);
            a.out[r0 + rr1] = v;
          }
        }
      } else {
        constexpr bool ISMAX = (OP == K_MAX);
        double best0 = ISMAX ? -CUDART_INF : CUDART_INF, best1 = best0;
        int idx0 = -1, idx1 = -1;
        for (int j = sub; j < cols; j += L) {
          const double u = row0[j], w = row1[j];
          const bool t0 = live0 && (ISMAX ? (u >= best0) : (u <= best0));
          const bool t1 = live1 && (ISMAX ? (w >= best1) : (w <= best1));
          idx0 = t0 ? j : idx0;
          best0 = t0 ? u : best0;
          idx1 = t1 ? j : idx1;
          best1 = t1 ? w : best1;
        }
        for (int o = L >> 1; o > 0; o >>= 1) {
          const double ob0 = __shfl_xor_sync(0xffffffffu, best0, o), ob1 = __shfl_xor_sync(0xffffffffu, best1, o);
          const int oi0 = __shfl_xor_sync(0xffffffffu, idx0, o), oi1 = __shfl_xor_sync(0xffffffffu, idx1, o);
          const bool b0 = (oi0 >= 0) && ((idx0 < 0) || (ISMAX ? (ob0 > best0) : (ob0 < best0)) || (ob0 == best0 && oi0 > idx0));
          const bool b1 = (oi1 >= 0) && ((idx1 < 0) || (ISMAX ? (ob1 > best1) : (ob1 < best1)) || (ob1 == best1 && oi1 > idx1));
          best0 = b0 ? ob0 : best0;
          idx0 = b0 ? oi0 : idx0;
          best1 = b1 ? ob1 : best1;
          idx1 = b1 ? oi1 : idx1;
        }
        if (sub == 0) {
          if (live0) a.out[r0 + rr0] = (double)(idx0 + 1);
          if (live1) a.out[r0 + rr1] = (double)(idx1 + 1);
        }
      }
    }
    mbar_arrive(&empty[s]);
  }
}

inline bool aggstream_ok(const double* X, i64 rows, i64 cols, bool per_row) {
  // per-row results need enough rows per tile to keep all consumer warps busy: short rows only
  return cols >= 1 && cols <= (per_row ? 128 : 2048) && rows * cols >= (i64)1 << 16 && (reinterpret_cast<uintptr_t>(X) & 7) == 0;
}

template <int OP, int KIND>
int run_aggstream(sysml_b200_ctx* ctx, const double* X, i64 rows, i64 cols, double* out, double scale) {
  AggStreamArgs a{};
  a.X = X; a.rows = rows; a.cols = (int)cols; a.out = out; a.scale = scale;
  // ~32 KB tiles, an even number of rows (16-byte multiples for the bulk copy), at least one pass of rows
  int R = (int)((32 * 1024) / (cols * 8));
  if (R < 2) R = 2;
  R &= ~1;
  if ((i64)R > rows) R = (int)((rows + 1) & ~(i64)1);
  a.L = lanes_per_row(cols);
  if (KIND != 2) {
    // per-row kinds: every warp handles 2 * (32 / L) rows per pass; make the tile a whole number of passes
    const int pass = 2 * (32 / a.L) * (AS_THREADS / 32);
    if (R >= pass) R = (R / pass) * pass;
    else if ((i64)pass * cols * 8 <= 64 * 1024) R = pass;
  }
  a.R = R;
  a.total_tiles = (rows + R - 1) / R;
  const size_t stage_bytes = (size_t)R * cols * 8;
  a.use_tma = ((reinterpret_cast<uintptr_t>(X) & 15) == 0) && (stage_bytes % 16 == 0) && stage_bytes < (1u << 20);
  const int CPB = (int)(cols < AS_THREADS ? cols : AS_THREADS), RPB = AS_THREADS / CPB;
  size_t smem = 128 + AS_STAGES * stage_bytes;
  if (KIND == 2 && (size_t)2 * RPB * cols * 8 > stage_bytes * AS_STAGES) smem = 128 + (size_t)2 * RPB * cols * 8;
  const int kc_needed = (int)((cols + CPB - 1) / CPB);
  void (*kern)(const AggStreamArgs) = aggstream_kernel<OP, KIND, 1>;
  if (KIND == 2) {
    if (kc_needed > 4) kern = aggstream_kernel<OP, 2, 8>;
    else if (kc_needed > 2) kern = aggstream_kernel<OP, 2, 4>;
    else if (kc_needed > 1) kern = aggstream_kernel<OP, 2, 2>;
    else kern = aggstream_kernel<OP, 2, 1>;
  }
  SYSML_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  i64 grid = (i64)ctx->sm_count * 2;
  if (grid > a.total_tiles) grid = a.total_tiles;
  if (grid < 1) grid = 1;
  if (KIND == 2) {
    void* ws = nullptr;
    int rc = sysml_workspace(ctx, (size_t)grid * 2 * cols * sizeof(double), &ws);
    if (rc) return rc;
    a.partial = (double*)ws;
  }
  kern<<<(unsigned)grid, AS_THREADS + 32, smem, ctx->stream>>>(a);
  SYSML_LAUNCH_CHECK(ctx, "aggstream_kernel");
  if (KIND == 2) {
    agg_col_finish_kernel<OP><<<(unsigned)((cols + 31) / 32), 256, 0, ctx->stream>>>(a.partial, (int)grid, cols, out, scale);
    SYSML_LAUNCH_CHECK(ctx, "agg_col_finish_kernel");
  }
  return SYSML_OK;
}

template <int OP>
int run_agg(sysml_b200_ctx* ctx, const double* X, i64 rows, i64 cols, int dir, double* out, double scale) {
  const i64 len = rows * cols;
  if (dir == SYSML_DIR_ALL) {
    i64 blocks = (len + 512 * 8 - 1) / (512 * 8);
    const i64 cap = (i64)ctx->sm_count * 4;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    void* ws = nullptr;
    int rc = sysml_workspace(ctx, (size_t)blocks * 2 * sizeof(double), &ws);
    if (rc) return rc;
    agg_all_kernel<OP><<<(unsigned)blocks, 512, 0, ctx->stream>>>(X, len, aligned16(X) ? 1 : 0, (double*)ws, ctx->counters + 1, out, scale);
    SYSML_LAUNCH_CHECK(ctx, "agg_all_kernel");
    return SYSML_OK;
  }
  if (dir == SYSML_DIR_ROW && aggstream_ok(X, rows, cols, true)) return run_aggstream<OP, 0>(ctx, X, rows, cols, out, scale);
  if (dir == SYSML_DIR_COL && aggstream_ok(X, rows, cols, false)) return run_aggstream<OP, 2>(ctx, X, rows, cols, out, scale);
  if (dir == SYSML_DIR_ROW) {
    const int W = lanes_per_row(cols);
    const i64 rows_per_block = (256 / 32) * (32 / W);
    i64 blocks = (rows + rows_per_block - 1) / rows_per_block;
    const i64 cap = (i64)ctx->sm_count * 16;
    if (blocks > cap) blocks = cap;
    agg_row_kernel<OP><<<(unsigned)blocks, 256, 0, ctx->stream>>>(X, rows, cols, W, out, scale);
    SYSML_LAUNCH_CHECK(ctx, "agg_row_kernel");
    return SYSML_OK;
  }
  // COL
  const int CPB = (int)(cols < 256 ? cols : 256);
  const int RPB = 256 / CPB;
  const i64 ctiles = (cols + CPB - 1) / CPB;
  i64 splits = ((i64)ctx->sm_count * 8 + ctiles - 1) / ctiles;
  const i64 max_splits = (rows + (i64)RPB * 8 - 1) / ((i64)RPB * 8);
  if (splits > max_splits) splits = max_splits;
  if (splits < 1) splits = 1;
  if (splits > 65535) splits = 65535;
  void* ws = nullptr;
  int rc = sysml_workspace(ctx, (size_t)splits * 2 * cols * sizeof(double), &ws);
  if (rc) return rc;
  dim3 grid((unsigned)ctiles, (unsigned)splits);
  agg_col_kernel<OP><<<grid, 256, 2 * RPB * CPB * sizeof(double), ctx->stream>>>(X, rows, cols, CPB, RPB, (double*)ws);
  SYSML_LAUNCH_CHECK(ctx, "agg_col_kernel");
  agg_col_finish_kernel<OP><<<(unsigned)((cols + 31) / 32), 256, 0, ctx->stream>>>((double*)ws, (int)splits, cols, out, scale);
  SYSML_LAUNCH_CHECK(ctx, "agg_col_finish_kernel");
  return SYSML_OK;
}


// ------------------------------------------------------------------------------------------------
// Variance (uavar / uarvar / uacvar): two passes -- the compensated mean first, then the compensated sum
// of squared deviations, divided by n-1 (0 for n == 1, CM_COV_Object.java:169-170).  The reference runs a
// one-pass Welford/Kahan update (functionobjects/CM.java:166-176); both are stable and agree to ~1e-15.
// mode: 0 one mean for everything, 1 mean per row, 2 mean per column.
// ------------------------------------------------------------------------------------------------
__global__ void var_all_kernel(const double* __restrict__ X, i64 len, const double* __restrict__ mean, double* __restrict__ partial) {
  const double mu = *mean;
  Acc<K_SUM> a;
  a.init();
  const i64 stride = (i64)gridDim.x * blockDim.x;
  for (i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x; i < len; i += stride) {
    const double d = __dsub_rn(ldg_stream(X + i), mu);
    a.push(__dmul_rn(d, d));
  }
  warp_merge(a);
  __shared__ double sh[2][32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) {
    sh[0][warp] = a.hi;
    sh[1][warp] = a.lo;
  }
  __syncthreads();
  if (warp == 0) {
    Acc<K_SUM> b;
    b.init();
    if (lane < (int)(blockDim.x >> 5)) {
      b.hi = sh[0][lane];
      b.lo = sh[1][lane];
    }
    warp_merge(b);
    if (lane == 0) {
      partial[2 * blockIdx.x] = b.hi;
      partial[2 * blockIdx.x + 1] = b.lo;
    }
  }
}
__global__ void var_all_finish_kernel(const double* __restrict__ partial, int nparts, double denom, double* __restrict__ out) {
  Acc<K_SUM> b;
  b.init();
  for (int i = threadIdx.x; i < nparts; i += 32) b.merge(partial[2 * i], partial[2 * i + 1]);
  warp_merge(b);
  if (threadIdx.x == 0) out[0] = denom > 0.0 ? __ddiv_rn(b.result(), denom) : 0.0;
}
__global__ void var_row_kernel(const double* __restrict__ X, i64 rows, i64 cols, int W, const double* __restrict__ mean, double* __restrict__ out) {
  const int lane = threadIdx.x & 31;
  const int sub = lane & (W - 1);
  const int rows_per_warp = 32 / W;
  const i64 warp = ((i64)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const i64 nwarps = ((i64)gridDim.x * blockDim.x) >> 5;
  for (i64 base = warp * rows_per_warp; base < rows; base += nwarps * rows_per_warp) {
    const i64 r = base + lane / W;
    Acc<K_SUM> a;
    a.init();
    if (r < rows) {
      const double mu = mean[r];
      for (i64 j = sub; j < cols; j += W) {
        const double d = __dsub_rn(ldg_stream(X + r * cols + j), mu);
        a.push(__dmul_rn(d, d));
      }
    }
    group_merge(a, W);
    if (r < rows && sub == 0) out[r] = cols > 1 ? __ddiv_rn(a.result(), (double)(cols - 1)) : 0.0;
  }
}
__global__ void var_col_kernel(const double* __restrict__ X, i64 rows, i64 cols, const double* __restrict__ mean, double* __restrict__ partial) {
  const i64 c = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= cols) return;
  const double mu = mean[c];
  Acc<K_SUM> a;
  a.init();
  for (i64 r = blockIdx.y; r < rows; r += gridDim.y) {
    const double d = __dsub_rn(ldg_stream(X + r * cols + c), mu);
    a.push(__dmul_rn(d, d));
  }
  partial[((size_t)blockIdx.y * 2) * cols + c] = a.hi;
  partial[((size_t)blockIdx.y * 2 + 1) * cols + c] = a.lo;
}

int run_var(sysml_b200_ctx* ctx, const double* X, i64 rows, i64 cols, int dir, double* out) {
  // pass 1: means, into the output buffer itself (overwritten by pass 2 only after it has been read)
  const i64 outlen = dir == SYSML_DIR_ALL ? 1 : (dir == SYSML_DIR_ROW ? rows : cols);
  int rc = SYSML_OK;
  double* mean = nullptr;
  SYSML_CUDA(ctx, cudaMalloc((void**)&mean, (size_t)outlen * sizeof(double)));
  const double cnt = dir == SYSML_DIR_ALL ? (double)rows * (double)cols : (dir == SYSML_DIR_ROW ? (double)cols : (double)rows);
  rc = run_agg<K_SUM>(ctx, X, rows, cols, dir, mean, cnt);
  if (!rc) {
    if (dir == SYSML_DIR_ALL) {
      i64 blocks = (rows * cols + 2047) / 2048;
      const i64 cap = (i64)ctx->sm_count * 8;
      if (blocks > cap) blocks = cap;
      void* ws = nullptr;
      rc = sysml_workspace(ctx, (size_t)blocks * 2 * sizeof(double), &ws);
      if (!rc) {
        var_all_kernel<<<(unsigned)blocks, 256, 0, ctx->stream>>>(X, rows * cols, mean, (double*)ws);
        var_all_finish_kernel<<<1, 32, 0, ctx->stream>>>((double*)ws, (int)blocks, cnt - 1.0, out);
        ctx->launches += 2;
      }
    } else if (dir == SYSML_DIR_ROW) {
      const int W = lanes_per_row(cols);
      const i64 rows_per_block = (256 / 32) * (32 / W);
      i64 blocks = (rows + rows_per_block - 1) / rows_per_block;
      const i64 cap = (i64)ctx->sm_count * 16;
      if (blocks > cap) blocks = cap;
      var_row_kernel<<<(unsigned)blocks, 256, 0, ctx->stream>>>(X, rows, cols, W, mean, out);
      ctx->launches++;
    } else {
      const i64 ctiles = (cols + 127) / 128;
      i64 splits = ((i64)ctx->sm_count * 8 + ctiles - 1) / ctiles;
      if (splits > (rows + 63) / 64) splits = (rows + 63) / 64;
      if (splits < 1) splits = 1;
      if (splits > 65535) splits = 65535;
      void* ws = nullptr;
      rc = sysml_workspace(ctx, (size_t)splits * 2 * cols * sizeof(double), &ws);
      if (!rc) {
        var_col_kernel<<<dim3((unsigned)ctiles, (unsigned)splits), 128, 0, ctx->stream>>>(X, rows, cols, mean, (double*)ws);
        agg_col_finish_kernel<K_SUM><<<(unsigned)((cols + 31) / 32), 256, 0, ctx->stream>>>((double*)ws, (int)splits, cols, out, rows > 1 ? (double)(rows - 1) : 1.0);
        ctx->launches += 2;
      }
    }
  }
  cudaError_t e = cudaGetLastError();
  cudaStreamSynchronize(ctx->stream);
  cudaFree(mean);
  if (rc) return rc;
  if (e != cudaSuccess) return sysml_set_error(ctx, SYSML_E_CUDA, "variance kernels: %s", cudaGetErrorString(e));
  return SYSML_OK;
}

}  // namespace

extern "C" int sysml_b200_agg_unary_dense(sysml_b200_ctx* ctx, const double* X, int64_t rows, int64_t cols, int op, int dir,
                                          double* out) {
  SYSML_ENTER(ctx);
  if (rows < 0 || cols < 0) return sysml_set_error(ctx, SYSML_E_DIM, "agg_unary: negative dimension");
  if (dir < 0 || dir > 2) return sysml_set_error(ctx, SYSML_E_INVALID, "agg_unary: bad direction %d", dir);
  if (!out) return sysml_set_error(ctx, SYSML_E_INVALID, "agg_unary: out is null");
  if (rows * cols == 0) return sysml_set_error(ctx, SYSML_E_DIM, "agg_unary: empty input (%lld x %lld)", (long long)rows, (long long)cols);
  if (!X) return sysml_set_error(ctx, SYSML_E_INVALID, "agg_unary: X is null");
  const double cnt = (dir == SYSML_DIR_ALL) ? (double)rows * (double)cols : (dir == SYSML_DIR_ROW ? (double)cols : (double)rows);
  switch (op) {
    case SYSML_AGG_SUM: return run_agg<K_SUM>(ctx, X, rows, cols, dir, out, 1.0);
    case SYSML_AGG_SUMSQ: return run_agg<K_SUMSQ>(ctx, X, rows, cols, dir, out, 1.0);
    case SYSML_AGG_MAX: return run_agg<K_MAX>(ctx, X, rows, cols, dir, out, 1.0);
    case SYSML_AGG_MIN: return run_agg<K_MIN>(ctx, X, rows, cols, dir, out, 1.0);
    case SYSML_AGG_PROD: return run_agg<K_PROD>(ctx, X, rows, cols, dir, out, 1.0);
    case SYSML_AGG_MEAN: return run_agg<K_SUM>(ctx, X, rows, cols, dir, out, cnt);
    case SYSML_AGG_MAXINDEX:
    case SYSML_AGG_MININDEX: {
      if (dir != SYSML_DIR_ROW)
        return sysml_set_error(ctx, SYSML_E_UNSUPPORTED, "rowIndexMax/Min exist only per row (LibMatrixAgg.java:1318-1329)");
      if (aggstream_ok(X, rows, cols, true))
        return op == SYSML_AGG_MAXINDEX ? run_aggstream<K_MAX, 1>(ctx, X, rows, cols, out, 1.0) : run_aggstream<K_MIN, 1>(ctx, X, rows, cols, out, 1.0);
      const int W = lanes_per_row(cols);
      const i64 rows_per_block = (256 / 32) * (32 / W);
      i64 blocks = (rows + rows_per_block - 1) / rows_per_block;
      const i64 cap = (i64)ctx->sm_count * 16;
      if (blocks > cap) blocks = cap;
      if (op == SYSML_AGG_MAXINDEX)
        agg_rowindex_kernel<1><<<(unsigned)blocks, 256, 0, ctx->stream>>>(X, rows, cols, W, out);
      else
        agg_rowindex_kernel<0><<<(unsigned)blocks, 256, 0, ctx->stream>>>(X, rows, cols, W, out);
      SYSML_LAUNCH_CHECK(ctx, "agg_rowindex_kernel");
      return SYSML_OK;
    }
    case SYSML_AGG_VAR: return run_var(ctx, X, rows, cols, dir, out);
  }
  return sysml_set_error(ctx, SYSML_E_INVALID, "agg_unary: unknown opcode %d", op);
}
