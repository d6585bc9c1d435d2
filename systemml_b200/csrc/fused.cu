// fused.cu -- fused cellwise chains: C = f(A1..Ak, s1..sj) evaluated in ONE pass over HBM, optionally ending in
// sum / rowSums / colSums.  
//
// What it replaces: the reference runs a chain like  sum(out * Y * Xd)  or  1 - Y * (X %*% w)  as one launch
// per operator (LibMatrixCUDA.matrixMatrixArithmetic / matrixScalarArithmetic / unaryOp, LibMatrixCUDA.java:
// 1126-1270, 1969-2290) plus one for the aggregate (unaryAggregate, :573-894), each a full HBM round trip
// through a temporary; its CPU-side answer to that is the code generator's cellwise template
// (runtime/codegen/SpoofCellwise.java: execute() with CellType NO_AGG / FULL_AGG / ROW_AGG / COL_AGG, one
// generated genexec() per fused operator).  This file is the GPU twin of that template: the chain arrives as a
// postfix program, is turned into CUDA source whose value functions are the SAME text the unfused kernels are
// compiled from (cellwise_ops.cuh, embedded at build time), compiled once with NVRTC for sm_100a, cached, and
// launched through the driver API.  Because every operator is the same correctly-rounded function either way
// (and none of them is contracted: they are written with __dadd_rn / __dmul_rn ...), a fused chain without an
// aggregate is bit-identical to the unfused sequence; sums use the same compensated accumulation as agg.cu.
//
// libnvrtc and libcuda are opened with dlopen on first use: the library itself must load on machines without a
// driver (the CPU test tier checks its exports), and there is no CPU fallback -- without NVRTC or a driver the
// call fails with SYSML_E_UNSUPPORTED.
#include <dlfcn.h>
#include <stdarg.h>
#include <stdlib.h>

#include <map>
#include <mutex>
#include <string>
#include <vector>

#include <math_constants.h>

#include "common.cuh"
#include "cellwise_ops.cuh"

namespace {

const char* kOpsHeaderText =
#include "cellwise_ops_embed.inc"
    ;

// ---- the kernel template handed to NVRTC (after the generated #defines, the ops header and fused_eval) ----
const char* kKernelTemplate = R"FUSEDSRC(
struct FusedArgs {
  const double* in[8];
  double k[8];
  double* out;
  double* partial;
  long long rows, cols;
};
__device__ __forceinline__ double ld_stream(const double* p) {
  double v;
  asm volatile("ld.global.nc.L1::no_allocate.f64 %0, [%1];" : "=d"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ double2 ld_stream2(const double2* p) {
  double2 v;
  asm volatile("ld.global.nc.L1::no_allocate.v2.f64 {%0,%1}, [%2];" : "=d"(v.x), "=d"(v.y) : "l"(p));
  return v;
}
__device__ __forceinline__ void two_sum(double a, double b, double& s, double& e) {
  s = __dadd_rn(a, b);
  double bb = __dsub_rn(s, a);
  e = __dadd_rn(__dsub_rn(a, __dsub_rn(s, bb)), __dsub_rn(b, bb));
}
// accumulator of the aggregate: FUSED_AOP 0 = compensated sum (value = hi + lo), 1 = min, 2 = max (Math.min/max
// semantics of the unfused kernels: NaN propagates, -0.0 < +0.0; lo unused)
#define FUSED_ACC_INIT {(FUSED_AOP == 1) ? __longlong_as_double(0x7ff0000000000000LL) : (FUSED_AOP == 2) ? __longlong_as_double(0xfff0000000000000LL) : 0.0, 0.0}
struct Comp {
  double hi, lo;
  __device__ __forceinline__ void add(double x) {
    if (FUSED_AOP == 0) {
      double s, e;
      two_sum(hi, x, s, e);
      hi = s;
      lo = __dadd_rn(lo, e);
    } else {
      // branch-free: plain compare-select, and two sticky flags kept as the bit pattern of lo (bit 0: a NaN was
      // seen, bit 1: a zero of the winning sign -- -0.0 for min, +0.0 for max -- was seen); settled in result()
      long long f = __double_as_longlong(lo);
      f |= (x != x) ? 1ll : 0ll;
      f |= (__double_as_longlong(x) == ((FUSED_AOP == 1) ? (long long)0x8000000000000000ull : 0ll)) ? 2ll : 0ll;
      lo = __longlong_as_double(f);
      hi = ((FUSED_AOP == 1) ? (x < hi) : (x > hi)) ? x : hi;
    }
  }
  __device__ __forceinline__ void merge(double ohi, double olo) {
    if (FUSED_AOP == 0) {
      double s, e;
      two_sum(hi, ohi, s, e);
      hi = s;
      lo = __dadd_rn(__dadd_rn(lo, olo), e);
    } else {
      lo = __longlong_as_double(__double_as_longlong(lo) | __double_as_longlong(olo));
      hi = ((FUSED_AOP == 1) ? (ohi < hi) : (ohi > hi)) ? ohi : hi;
    }
  }
  __device__ __forceinline__ double result() const {
    if (FUSED_AOP != 0) {
      const long long f = __double_as_longlong(lo);
      if (f & 1) return __longlong_as_double(0x7ff8000000000000LL);
      if (hi == 0.0 && (f & 2)) return (FUSED_AOP == 1) ? -0.0 : 0.0;
      return hi;
    }
    return (isinf(hi) || lo != lo) ? hi : __dadd_rn(hi, lo);
  }
};
__device__ __forceinline__ void group_merge(Comp& a, int W) {
  for (int o = W >> 1; o > 0; o >>= 1) {
    const double ohi = __shfl_xor_sync(0xffffffffu, a.hi, o);
    const double olo = __shfl_xor_sync(0xffffffffu, a.lo, o);
    a.merge(ohi, olo);
  }
}

#if FUSED_SHAPE == 0
// ---- flat: every input is a full matrix (or the chain runs over vectors); no (row, col) bookkeeping ----
extern "C" __global__ void __launch_bounds__(256, 4) fused_kernel(const FusedArgs a) {
  const long long len = a.rows * a.cols;
  const long long stride = (long long)gridDim.x * blockDim.x;
  const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
#if FUSED_AGG == 1
  Comp acc FUSED_ACC_INIT, acc2 FUSED_ACC_INIT;
#endif
#if FUSED_VEC2
  const long long n2 = len >> 1;
  // two strides per trip, all loads first: twice the bytes in flight per thread (a 50M-cell vector is a 0.2 ms kernel and
  // latency-bound otherwise); the additions happen in the order of the one-stride loop, so the sums are the same bits
  for (long long i = tid; i < n2; i += 2 * stride) {
    double x[FUSED_NIN + 1], y[FUSED_NIN + 1], xb[FUSED_NIN + 1], yb[FUSED_NIN + 1];
    const long long ib = i + stride;
    FUSED_LOAD2(i)
    if (ib < n2) { FUSED_LOAD2B(ib) }
    const double u = fused_eval(x, a.k), w = fused_eval(y, a.k);
#if FUSED_AGG == 1
    acc.add(u);
    acc2.add(w);
#else
    reinterpret_cast<double2*>(a.out)[i] = make_double2(u, w);
#endif
    if (ib < n2) {
      const double ub = fused_eval(xb, a.k), wb = fused_eval(yb, a.k);
#if FUSED_AGG == 1
      acc.add(ub);
      acc2.add(wb);
#else
      reinterpret_cast<double2*>(a.out)[ib] = make_double2(ub, wb);
#endif
    }
  }
#else
  for (long long i = tid; i < len; i += stride) {
    double x[FUSED_NIN + 1];
    FUSED_LOAD_FLAT(i)
    const double u = fused_eval(x, a.k);
#if FUSED_AGG == 1
    acc.add(u);
#else
    a.out[i] = u;
#endif
  }
#endif
#if FUSED_AGG == 1
  acc.merge(acc2.hi, acc2.lo);
  group_merge(acc, 32);
  __shared__ double sh[2][8];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) {
    sh[0][warp] = acc.hi;
    sh[1][warp] = acc.lo;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    Comp t{sh[0][0], sh[1][0]};
    for (int q = 1; q < 8; q++) t.merge(sh[0][q], sh[1][q]);
    a.partial[2 * blockIdx.x] = t.hi;
    a.partial[2 * blockIdx.x + 1] = t.lo;
  }
#endif
}
#elif FUSED_SHAPE == 1
// ---- broadcast, no aggregate or full sum: flat index with incremental (row, col) ----
extern "C" __global__ void __launch_bounds__(256, 4) fused_kernel(const FusedArgs a) {
  const long long len = a.rows * a.cols, cols = a.cols;
  const long long stride = (long long)gridDim.x * blockDim.x;
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
#if FUSED_AGG == 1
  Comp acc FUSED_ACC_INIT;
#endif
  if (i < len) {
    long long r = i / cols, c = i - r * cols;
    const long long dr = stride / cols, dc = stride - dr * cols;
    for (; i < len; i += stride) {
      double x[FUSED_NIN + 1];
      FUSED_LOAD_RC(i, r, c)
      const double u = fused_eval(x, a.k);
#if FUSED_AGG == 1
      acc.add(u);
#else
      a.out[i] = u;
#endif
      r += dr;
      c += dc;
      if (c >= cols) {
        c -= cols;
        r++;
      }
    }
  }
#if FUSED_AGG == 1
  group_merge(acc, 32);
  __shared__ double sh[2][8];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) {
    sh[0][warp] = acc.hi;
    sh[1][warp] = acc.lo;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    Comp t{sh[0][0], sh[1][0]};
    for (int q = 1; q < 8; q++) t.merge(sh[0][q], sh[1][q]);
    a.partial[2 * blockIdx.x] = t.hi;
    a.partial[2 * blockIdx.x + 1] = t.lo;
  }
#endif
}
#elif FUSED_SHAPE == 2
// ---- row sums: W lanes per row (W = FUSED_W), lanes stride over the columns ----
extern "C" __global__ void __launch_bounds__(256, 4) fused_kernel(const FusedArgs a) {
  const int lane = threadIdx.x & 31, sub = lane & (FUSED_W - 1);
  const long long rows = a.rows, cols = a.cols;
  const long long gw = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const long long nw = ((long long)gridDim.x * blockDim.x) >> 5;
  constexpr int RPW = 32 / FUSED_W;
  for (long long base = gw * RPW; base < rows; base += nw * RPW) {
    const long long r = base + lane / FUSED_W;
    Comp acc FUSED_ACC_INIT, accb FUSED_ACC_INIT;  // two chains per lane: twice the loads in flight
    if (r < rows) {
      const long long off = r * cols;
      long long c = sub;
      for (; c + FUSED_W < cols; c += 2 * FUSED_W) {
        double x[FUSED_NIN + 1], y[FUSED_NIN + 1];
        FUSED_LOAD_RC(off + c, r, c)
        FUSED_LOAD_RC_Y(off + c + FUSED_W, r, c + FUSED_W)
        acc.add(fused_eval(x, a.k));
        accb.add(fused_eval(y, a.k));
      }
      if (c < cols) {
        double x[FUSED_NIN + 1];
        FUSED_LOAD_RC(off + c, r, c)
        acc.add(fused_eval(x, a.k));
      }
    }
    acc.merge(accb.hi, accb.lo);
    group_merge(acc, FUSED_W);
    if (r < rows && sub == 0) a.out[r] = acc.result();
  }
}
#elif FUSED_SHAPE == 4
// ---- row aggregates of a NARROW matrix (specialised on the column count): a warp takes FUSED_RW adjacent rows,
// which are contiguous in memory, evaluates the chain over them with flat, fully coalesced, fully unrolled loads
// (dozens of independent loads per lane instead of a handful per row), parks the values in shared memory and
// then reduces every row with 32/FUSED_RW lanes ----
extern "C" __global__ void __launch_bounds__(256, 4) fused_kernel(const FusedArgs a) {
  constexpr int COLS = FUSED_COLS, RW = FUSED_RW, TILE = RW * COLS, LPR = 32 / RW;
  __shared__ double sh[8][TILE];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const long long rows = a.rows;
  const long long gw = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const long long nw = ((long long)gridDim.x * blockDim.x) >> 5;
  for (long long rb = gw * RW; rb < rows; rb += nw * RW) {
    const int nvalid = (int)((rows - rb < RW ? rows - rb : RW) * COLS);
#pragma unroll
    for (int e0 = 0; e0 < TILE; e0 += 32) {
      const int e = e0 + lane;
      if (e < nvalid) {
        const int rr = e / COLS, c = e - rr * COLS;
        const long long r = rb + rr;
        double x[FUSED_NIN + 1];
        FUSED_LOAD_RC(rb * COLS + e, r, c)
        sh[w][e] = fused_eval(x, a.k);
      }
    }
    __syncwarp();
    const int row = lane / LPR, sub = lane - row * LPR;
    Comp acc FUSED_ACC_INIT;
    if (rb + row < rows) {
      const double* p = &sh[w][row * COLS];
      for (int c = sub; c < COLS; c += LPR) acc.add(p[c]);
    }
    group_merge(acc, LPR);
    if (rb + row < rows && sub == 0) a.out[rb + row] = acc.result();
    __syncwarp();
  }
}
#elif FUSED_SHAPE == 5
// ---- several full sums over the same operands in one pass (the code generator's multi-aggregate template):
// fused_eval_multi leaves FUSED_NOUT values per cell, each with its own compensated accumulator;
// partial[out][block][2] ----
extern "C" __global__ void __launch_bounds__(256, 4) fused_kernel(const FusedArgs a) {
  const long long len = a.rows * a.cols;
  const long long stride = (long long)gridDim.x * blockDim.x;
  const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  Comp acc[FUSED_NOUT], acc2[FUSED_NOUT];
#pragma unroll
  for (int j = 0; j < FUSED_NOUT; j++) {
    acc[j].hi = acc[j].lo = 0.0;
    acc2[j].hi = acc2[j].lo = 0.0;
  }
#if FUSED_VEC2
  const long long n2 = len >> 1;
  for (long long i = tid; i < n2; i += 2 * stride) {  // two strides per trip, as in the single-result kernel
    double x[FUSED_NIN + 1], y[FUSED_NIN + 1], xb[FUSED_NIN + 1], yb[FUSED_NIN + 1], u[FUSED_NOUT], w[FUSED_NOUT];
    const long long ib = i + stride;
    FUSED_LOAD2(i)
    if (ib < n2) { FUSED_LOAD2B(ib) }
    fused_eval_multi(x, a.k, u);
    fused_eval_multi(y, a.k, w);
#pragma unroll
    for (int j = 0; j < FUSED_NOUT; j++) {
      acc[j].add(u[j]);
      acc2[j].add(w[j]);
    }
    if (ib < n2) {
      fused_eval_multi(xb, a.k, u);
      fused_eval_multi(yb, a.k, w);
#pragma unroll
      for (int j = 0; j < FUSED_NOUT; j++) {
        acc[j].add(u[j]);
        acc2[j].add(w[j]);
      }
    }
  }
#else
  {
    const long long cols = a.cols;
    long long i = tid;
    if (i < len) {
      long long r = i / cols, c = i - r * cols;
      const long long dr = stride / cols, dc = stride - dr * cols;
      for (; i < len; i += stride) {
        double x[FUSED_NIN + 1], u[FUSED_NOUT];
        FUSED_LOAD_RC(i, r, c)
        fused_eval_multi(x, a.k, u);
#pragma unroll
        for (int j = 0; j < FUSED_NOUT; j++) acc[j].add(u[j]);
        r += dr;
        c += dc;
        if (c >= cols) {
          c -= cols;
          r++;
        }
      }
    }
  }
#endif
  __shared__ double sh[FUSED_NOUT][2][8];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int j = 0; j < FUSED_NOUT; j++) {
    acc[j].merge(acc2[j].hi, acc2[j].lo);
    group_merge(acc[j], 32);
    if (lane == 0) {
      sh[j][0][warp] = acc[j].hi;
      sh[j][1][warp] = acc[j].lo;
    }
  }
  __syncthreads();
  if (threadIdx.x < FUSED_NOUT) {
    const int j = threadIdx.x;
    Comp t{sh[j][0][0], sh[j][1][0]};
    for (int q = 1; q < 8; q++) t.merge(sh[j][0][q], sh[j][1][q]);
    a.partial[((size_t)j * gridDim.x + blockIdx.x) * 2] = t.hi;
    a.partial[((size_t)j * gridDim.x + blockIdx.x) * 2 + 1] = t.lo;
  }
}
#else
// ---- column sums: a thread owns one column chunk, CTAs stride over row blocks; partial[grid][2][cols] ----
extern "C" __global__ void __launch_bounds__(256, 4) fused_kernel(const FusedArgs a) {
  const long long rows = a.rows, cols = a.cols;
  const int CPB = cols < 256 ? (int)cols : 256;
  const int RPB = 256 / CPB;
  const int cl = threadIdx.x % CPB, rl = threadIdx.x / CPB;
  __shared__ double sh[2][256];
  for (long long c0 = 0; c0 < cols; c0 += CPB) {
    const long long c = c0 + cl;
    Comp acc FUSED_ACC_INIT;
    if (rl < RPB && c < cols) {
      for (long long r = (long long)blockIdx.x * RPB + rl; r < rows; r += (long long)gridDim.x * RPB) {
        double x[FUSED_NIN + 1];
        FUSED_LOAD_RC(r * cols + c, r, c)
        acc.add(fused_eval(x, a.k));
      }
    }
    sh[0][threadIdx.x] = acc.hi;
    sh[1][threadIdx.x] = acc.lo;
    __syncthreads();
    if (rl == 0 && c < cols) {
      for (int q = 1; q < RPB; q++) acc.merge(sh[0][q * CPB + cl], sh[1][q * CPB + cl]);
      a.partial[((size_t)blockIdx.x * 2) * cols + c] = acc.hi;
      a.partial[((size_t)blockIdx.x * 2 + 1) * cols + c] = acc.lo;
    }
    __syncthreads();
  }
}
#endif
)FUSEDSRC";

// ---- precompiled finishers (fixed order => run-to-run reproducible) ----
template <int AOP>
__device__ __forceinline__ void fin_add(CompSum& s, double hi, double lo) {
  if (AOP == 0) {
    s.add(hi);
    s.lo = __dadd_rn(s.lo, lo);
  } else {  // (value, flag bits) pairs as the generated kernels write them
    s.lo = __longlong_as_double(__double_as_longlong(s.lo) | __double_as_longlong(lo));
    s.hi = ((AOP == 1) ? (hi < s.hi) : (hi > s.hi)) ? hi : s.hi;
  }
}
template <int AOP>
__device__ __forceinline__ CompSum fin_init() {
  CompSum s;
  if (AOP == 1) s.hi = CUDART_INF;
  if (AOP == 2) s.hi = -CUDART_INF;
  return s;
}
template <int AOP>
__device__ __forceinline__ double fin_result(const CompSum& s) {
  if (AOP != 0) {
    const long long f = __double_as_longlong(s.lo);
    if (f & 1) return CUDART_NAN;
    if (s.hi == 0.0 && (f & 2)) return (AOP == 1) ? -0.0 : 0.0;
    return s.hi;
  }
  return (isinf(s.hi) || s.lo != s.lo) ? s.hi : s.value();
}
template <int AOP>
__global__ void fused_all_finish(const double* __restrict__ partial, int n, double* __restrict__ out) {
  // one warp per result (blockIdx.x: the multi-aggregate form has several): lane-strided merge in a fixed order, then a butterfly
  partial += (size_t)blockIdx.x * 2 * n;
  out += blockIdx.x;
  CompSum s = fin_init<AOP>();
  for (int i = threadIdx.x; i < n; i += 32) fin_add<AOP>(s, partial[2 * i], partial[2 * i + 1]);
  for (int o = 16; o > 0; o >>= 1) {
    const double ohi = __shfl_xor_sync(0xffffffffu, s.hi, o);
    const double olo = __shfl_xor_sync(0xffffffffu, s.lo, o);
    if (AOP == 0) {
      CompSum t;
      t.hi = ohi;
      t.lo = olo;
      s.add(t);
    } else {
      fin_add<AOP>(s, ohi, olo);
    }
  }
  if (threadIdx.x == 0) out[0] = fin_result<AOP>(s);
}
template <int AOP>
__global__ void fused_col_finish(const double* __restrict__ partial, int P, i64 cols, double* __restrict__ out) {
  const i64 c = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= cols) return;
  CompSum s = fin_init<AOP>();
  for (int p = 0; p < P; p++) fin_add<AOP>(s, partial[((size_t)p * 2) * cols + c], partial[((size_t)p * 2 + 1) * cols + c]);
  out[c] = fin_result<AOP>(s);
}

// ---- dynamically loaded NVRTC + driver entry points ----
typedef int (*nvrtcCreateProgram_t)(void**, const char*, const char*, int, const char* const*, const char* const*);
typedef int (*nvrtcCompileProgram_t)(void*, int, const char* const*);
typedef int (*nvrtcGetSize_t)(void*, size_t*);
typedef int (*nvrtcGetData_t)(void*, char*);
typedef int (*nvrtcDestroyProgram_t)(void**);
typedef int (*cuModuleLoadData_t)(void**, const void*);
typedef int (*cuModuleGetFunction_t)(void**, void*, const char*);
typedef int (*cuLaunchKernel_t)(void*, unsigned, unsigned, unsigned, unsigned, unsigned, unsigned, unsigned, void*, void**, void**);
typedef int (*cuGetErrorString_t)(int, const char**);

struct Jit {
  bool tried = false, ok = false;
  std::string why;
  nvrtcCreateProgram_t createProgram;
  nvrtcCompileProgram_t compileProgram;
  nvrtcGetSize_t getCubinSize, getLogSize;
  nvrtcGetData_t getCubin, getLog;
  nvrtcDestroyProgram_t destroyProgram;
  cuModuleLoadData_t moduleLoadData;
  cuModuleGetFunction_t moduleGetFunction;
  cuLaunchKernel_t launchKernel;
  cuGetErrorString_t getErrorString;
};
Jit g_jit;
std::mutex g_jit_mu;
std::map<std::string, void*> g_kernels;  // (device, program key) -> CUfunction

void* open_first(const char* const* names) {
  for (; *names; names++)
    if (void* h = dlopen(*names, RTLD_NOW | RTLD_LOCAL)) return h;
  return nullptr;
}

bool jit_init() {  // g_jit_mu held
  if (g_jit.tried) return g_jit.ok;
  g_jit.tried = true;
  static const char* const rtc_names[] = {"libnvrtc.so.12", "/usr/local/cuda/lib64/libnvrtc.so.12", "libnvrtc.so", nullptr};
  static const char* const drv_names[] = {"libcuda.so.1", "libcuda.so", nullptr};
  void* rtc = open_first(rtc_names);
  if (!rtc) {
    g_jit.why = "libnvrtc.so.12 not found (fused cellwise kernels are compiled at run time)";
    return false;
  }
  void* drv = open_first(drv_names);
  if (!drv) {
    g_jit.why = "libcuda.so.1 not found (no NVIDIA driver)";
    return false;
  }
#define JIT_SYM(lib, field, type, name)                   \
  g_jit.field = (type)dlsym(lib, name);                   \
  if (!g_jit.field) {                                     \
    g_jit.why = std::string("missing symbol ") + name;    \
    return false;                                         \
  }
  JIT_SYM(rtc, createProgram, nvrtcCreateProgram_t, "nvrtcCreateProgram")
  JIT_SYM(rtc, compileProgram, nvrtcCompileProgram_t, "nvrtcCompileProgram")
  JIT_SYM(rtc, getCubinSize, nvrtcGetSize_t, "nvrtcGetCUBINSize")
  JIT_SYM(rtc, getCubin, nvrtcGetData_t, "nvrtcGetCUBIN")
  JIT_SYM(rtc, getLogSize, nvrtcGetSize_t, "nvrtcGetProgramLogSize")
  JIT_SYM(rtc, getLog, nvrtcGetData_t, "nvrtcGetProgramLog")
  JIT_SYM(rtc, destroyProgram, nvrtcDestroyProgram_t, "nvrtcDestroyProgram")
  JIT_SYM(drv, moduleLoadData, cuModuleLoadData_t, "cuModuleLoadData")
  JIT_SYM(drv, moduleGetFunction, cuModuleGetFunction_t, "cuModuleGetFunction")
  JIT_SYM(drv, launchKernel, cuLaunchKernel_t, "cuLaunchKernel")
  JIT_SYM(drv, getErrorString, cuGetErrorString_t, "cuGetErrorString")
#undef JIT_SYM
  g_jit.ok = true;
  return true;
}

struct FusedArgsHost {
  const double* in[8];
  double k[8];
  double* out;
  double* partial;
  long long rows, cols;
};

void appendf(std::string& s, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  s += buf;
}

// postfix program -> "static double fused_eval(const double* x, const double* k)"; returns "" and sets err on a malformed program
// (n_out > 0: "static void fused_eval_multi(const double* x, const double* k, double* o)", the program leaves n_out values)
std::string gen_eval(const int32_t* prog, int n_tokens, int n_inputs, int n_consts, std::string& err, int n_out = 0) {
  std::string src = n_out > 0 ? "__device__ __forceinline__ void fused_eval_multi(const double* x, const double* k, double* o) {\n"
                              : "__device__ __forceinline__ double fused_eval(const double* x, const double* k) {\n";
  std::vector<int> stack;
  int nt = 0;
  for (int t = 0; t < n_tokens; t++) {
    const int kind = (prog[t] >> 24) & 0xff, arg = prog[t] & 0xffffff;
    switch (kind) {
      case SYSML_FUSED_IN:
        if (arg >= n_inputs) { err = "input index out of range"; return ""; }
        appendf(src, "  const double t%d = x[%d];\n", nt, arg);
        stack.push_back(nt++);
        break;
      case SYSML_FUSED_CONST:
        if (arg >= n_consts) { err = "constant index out of range"; return ""; }
        appendf(src, "  const double t%d = k[%d];\n", nt, arg);
        stack.push_back(nt++);
        break;
      case SYSML_FUSED_BINARY: {
        if (arg > SYSML_OP_INTDIV) { err = "unknown binary opcode"; return ""; }
        if (stack.size() < 2) { err = "stack underflow at a binary operator"; return ""; }
        const int y = stack.back(); stack.pop_back();
        const int x = stack.back(); stack.pop_back();
        appendf(src, "  const double t%d = sysml_ops::binop<%d>(t%d, t%d);\n", nt, arg, x, y);
        stack.push_back(nt++);
        break;
      }
      case SYSML_FUSED_UNARY: {
        if (arg > SYSML_UN_TANH) { err = "unknown unary opcode"; return ""; }
        if (stack.empty()) { err = "stack underflow at a unary operator"; return ""; }
        const int x = stack.back(); stack.pop_back();
        appendf(src, "  const double t%d = sysml_ops::unop<%d>(t%d);\n", nt, arg, x);
        stack.push_back(nt++);
        break;
      }
      case SYSML_FUSED_PICK:
        if ((size_t)arg >= stack.size()) { err = "pick reaches below the stack"; return ""; }
        stack.push_back(stack[stack.size() - 1 - (size_t)arg]);
        break;
      case SYSML_FUSED_NIP:
        if (stack.size() < 2) { err = "stack underflow at nip"; return ""; }
        stack.erase(stack.end() - 2);
        break;
      default:
        err = "unknown token kind";
        return "";
    }
  }
  if (n_out > 0) {
    if ((int)stack.size() != n_out) { err = "program must leave exactly one value per result on the stack"; return ""; }
    for (int j = 0; j < n_out; j++) appendf(src, "  o[%d] = t%d;\n", j, stack[(size_t)j]);
    src += "}\n";
    return src;
  }
  if (stack.size() != 1) { err = "program must leave exactly one value on the stack"; return ""; }
  appendf(src, "  return t%d;\n}\n", stack.back());
  return src;
}

std::string gen_defines() {
  std::string d;
#define D(name) appendf(d, "#define " #name " %d\n", name);
  D(SYSML_OP_PLUS) D(SYSML_OP_MINUS) D(SYSML_OP_MULTIPLY) D(SYSML_OP_DIVIDE) D(SYSML_OP_POWER) D(SYSML_OP_LESS)
  D(SYSML_OP_LESSEQUAL) D(SYSML_OP_GREATER) D(SYSML_OP_GREATEREQUAL) D(SYSML_OP_EQUAL) D(SYSML_OP_NOTEQUAL)
  D(SYSML_OP_MIN) D(SYSML_OP_MAX) D(SYSML_OP_AND) D(SYSML_OP_OR) D(SYSML_OP_MINUS1MULTIPLY) D(SYSML_OP_MINUSNZ)
  D(SYSML_OP_MODULUS) D(SYSML_OP_INTDIV)
  D(SYSML_UN_EXP) D(SYSML_UN_SQRT) D(SYSML_UN_ROUND) D(SYSML_UN_ABS) D(SYSML_UN_LOG) D(SYSML_UN_FLOOR) D(SYSML_UN_CEIL)
  D(SYSML_UN_SIGN) D(SYSML_UN_SIGMOID) D(SYSML_UN_SIN) D(SYSML_UN_COS) D(SYSML_UN_TAN) D(SYSML_UN_TANH)
#undef D
  return d;
}

}  // namespace

// n_out == 0: one chain, result per `agg_in`;  n_out > 0: the program leaves n_out values per cell, each summed over all cells
static int fused_run(sysml_b200_ctx* ctx, const double* const* inputs, const int* input_kinds, int n_inputs, const double* consts,
                     int n_consts, const int32_t* program, int n_tokens, int64_t rows, int64_t cols, int agg_in, int n_out, double* out) {
  SYSML_ENTER(ctx);
  int agg = agg_in;
  if (n_inputs < 0 || n_inputs > SYSML_FUSED_MAX_INPUTS || n_consts < 0 || n_consts > SYSML_FUSED_MAX_CONSTS || n_tokens < 1 ||
      n_tokens > SYSML_FUSED_MAX_TOKENS)
    return sysml_set_error(ctx, SYSML_E_LIMIT, "fused_cellwise: at most %d inputs, %d constants, %d tokens", SYSML_FUSED_MAX_INPUTS,
                           SYSML_FUSED_MAX_CONSTS, SYSML_FUSED_MAX_TOKENS);
  if (rows < 0 || cols < 0) return sysml_set_error(ctx, SYSML_E_DIM, "fused_cellwise: negative dimension");
  if (agg < SYSML_FUSED_AGG_NONE || agg > SYSML_FUSED_AGG_COLMAXS) return sysml_set_error(ctx, SYSML_E_INVALID, "fused_cellwise: unknown aggregate %d", agg_in);
  // split into the accumulator (0 sum, 1 min, 2 max) and the direction, numbered like the three sum codes
  const int aop = agg <= SYSML_FUSED_AGG_COLSUMS ? 0 : (agg <= SYSML_FUSED_AGG_COLMINS ? 1 : 2);
  if (agg > SYSML_FUSED_AGG_COLSUMS) agg = (agg - 1) % 3 + 1;
  if (!program || (n_inputs && (!inputs || !input_kinds)) || (n_consts && !consts)) return sysml_set_error(ctx, SYSML_E_INVALID, "fused_cellwise: null pointer");
  const i64 len = rows * cols;
  const i64 out_len = n_out > 0 ? n_out : agg == SYSML_FUSED_AGG_NONE ? len : agg == SYSML_FUSED_AGG_SUM ? 1 : agg == SYSML_FUSED_AGG_ROWSUMS ? rows : cols;
  if (out_len == 0) return SYSML_OK;
  if (!out) return sysml_set_error(ctx, SYSML_E_INVALID, "fused_cellwise: out is null");
  if (len == 0) {  // an empty operand: every sum is 0 (LibMatrixAgg: empty input -> zero output)
    if (aop != 0) return sysml_set_error(ctx, SYSML_E_DIM, "fused_cellwise: min / max of an empty operand");
    SYSML_CUDA(ctx, cudaMemsetAsync(out, 0, (size_t)out_len * sizeof(double), ctx->stream));
    return SYSML_OK;
  }
  bool any_bcast = false;
  bool aligned16 = (reinterpret_cast<uintptr_t>(out) & 15) == 0;
  for (int i = 0; i < n_inputs; i++) {
    if (!inputs[i]) return sysml_set_error(ctx, SYSML_E_INVALID, "fused_cellwise: input %d is null", i);
    const int kd = input_kinds[i];
    if (kd != SYSML_BCAST_NONE && kd != SYSML_BCAST_COL && kd != SYSML_BCAST_ROW) return sysml_set_error(ctx, SYSML_E_INVALID, "fused_cellwise: input %d has unknown kind %d", i, kd);
    // a column vector over a one-column chain (or a row vector over a one-row chain) is just a full operand
    const bool full = kd == SYSML_BCAST_NONE || (kd == SYSML_BCAST_COL && cols == 1) || (kd == SYSML_BCAST_ROW && rows == 1);
    any_bcast |= !full;
    aligned16 &= (reinterpret_cast<uintptr_t>(inputs[i]) & 15) == 0;
  }
  // kernel shape
  int shape, vec2 = 0, W = 32;
  int RW = 0;
  if (n_out > 0) {
    shape = 5;
    vec2 = (!any_bcast && aligned16 && (len % 2 == 0)) ? 1 : 0;
  } else if (agg == SYSML_FUSED_AGG_ROWSUMS && cols > 1 && cols <= 96 && getenv("SYSML_B200_FUSED_NO_NARROW") == nullptr) {
    shape = 4;  // narrow: rows per warp = largest power of two with RW * cols <= 768 cells (6 KB per warp, 48 KB per CTA)
    RW = 32;
    while (RW > 1 && (i64)RW * cols > 768) RW >>= 1;
  } else if (agg == SYSML_FUSED_AGG_ROWSUMS && cols > 1) {
    shape = 2;
    W = 1;  // lanes per row: ~8 cells per lane, so a warp covers several adjacent (contiguous) rows of a narrow matrix
    if (const char* e = getenv("SYSML_B200_FUSED_ROW_LANES")) W = atoi(e);  // A/B switch for measurements
    else
      while (W < 32 && 8 * W < cols) W <<= 1;
    if (W != 1 && W != 2 && W != 4 && W != 8 && W != 16 && W != 32) W = 32;
  } else if (agg == SYSML_FUSED_AGG_COLSUMS && rows > 1) {
    shape = 3;
  } else {
    shape = any_bcast ? 1 : 0;
    vec2 = (shape == 0 && aligned16 && (len % 2 == 0)) ? 1 : 0;
  }
  // row sums over one column / column sums over one row are the chain itself
  const int kagg = ((shape <= 1 && agg == SYSML_FUSED_AGG_SUM) || shape == 5) ? 1 : ((shape == 2 || shape == 4) ? 2 : (shape == 3 ? 3 : 0));

  // ---- find or build the kernel ----
  std::string key;
  appendf(key, "d%d|s%d|v%d|w%d|a%d|o%d|n%d|c%lld|m%d|", ctx->device, shape, vec2, W, kagg, aop, n_inputs, shape == 4 ? (long long)cols : 0ll, n_out);
  for (int i = 0; i < n_inputs; i++) {
    const int kd = input_kinds[i];
    const bool full = kd == SYSML_BCAST_NONE || (kd == SYSML_BCAST_COL && cols == 1) || (kd == SYSML_BCAST_ROW && rows == 1);
    appendf(key, "%d,", full ? 0 : kd);
  }
  key += "|";
  for (int t = 0; t < n_tokens; t++) appendf(key, "%x,", program[t]);
  void* fn = nullptr;
  {
    std::lock_guard<std::mutex> lk(g_jit_mu);
    if (!jit_init()) return sysml_set_error(ctx, SYSML_E_UNSUPPORTED, "fused_cellwise: %s", g_jit.why.c_str());
    auto it = g_kernels.find(key);
    if (it != g_kernels.end()) {
      fn = it->second;
    } else {
      std::string err;
      const std::string eval = gen_eval(program, n_tokens, n_inputs, n_consts, err, n_out);
      if (eval.empty()) return sysml_set_error(ctx, SYSML_E_INVALID, "fused_cellwise: %s", err.c_str());
      std::string src = gen_defines();
      appendf(src, "#define FUSED_NIN %d\n#define FUSED_SHAPE %d\n#define FUSED_VEC2 %d\n#define FUSED_W %d\n#define FUSED_AGG %d\n#define FUSED_AOP %d\n",
              n_inputs, shape, vec2, W, kagg, aop);
      appendf(src, "#define FUSED_COLS %d\n#define FUSED_RW %d\n#define FUSED_NOUT %d\n", shape == 4 ? (int)cols : 1, shape == 4 ? RW : 1, n_out > 0 ? n_out : 1);
      // per-input loads: matrices stream past L1, broadcast vectors are cached reads
      std::string flat = "#define FUSED_LOAD_FLAT(i) ", rc = "#define FUSED_LOAD_RC(i, r, c) ", rcy = "#define FUSED_LOAD_RC_Y(i, r, c) ",
                  l2 = "#define FUSED_LOAD2(i) ", l2b = "#define FUSED_LOAD2B(i) ";
      for (int i = 0; i < n_inputs; i++) {
        const int kd = input_kinds[i];
        const bool full = kd == SYSML_BCAST_NONE || (kd == SYSML_BCAST_COL && cols == 1) || (kd == SYSML_BCAST_ROW && rows == 1);
        appendf(flat, "x[%d] = ld_stream(a.in[%d] + (i)); ", i, i);
        if (full) appendf(rc, "x[%d] = ld_stream(a.in[%d] + (i)); ", i, i);
        else if (kd == SYSML_BCAST_COL) appendf(rc, "x[%d] = __ldg(a.in[%d] + (r)); ", i, i);
        else appendf(rc, "x[%d] = __ldg(a.in[%d] + (c)); ", i, i);
        if (full) appendf(rcy, "y[%d] = ld_stream(a.in[%d] + (i)); ", i, i);
        else if (kd == SYSML_BCAST_COL) appendf(rcy, "y[%d] = __ldg(a.in[%d] + (r)); ", i, i);
        else appendf(rcy, "y[%d] = __ldg(a.in[%d] + (c)); ", i, i);
        appendf(l2, "{ const double2 v_ = ld_stream2(reinterpret_cast<const double2*>(a.in[%d]) + (i)); x[%d] = v_.x; y[%d] = v_.y; } ", i, i, i);
        appendf(l2b, "{ const double2 v_ = ld_stream2(reinterpret_cast<const double2*>(a.in[%d]) + (i)); xb[%d] = v_.x; yb[%d] = v_.y; } ", i, i, i);
      }
      src += flat + "\n" + rc + "\n" + rcy + "\n" + l2 + "\n" + l2b + "\n";
      src += kOpsHeaderText;
      src += eval;
      src += kKernelTemplate;
      void* prog = nullptr;
      if (g_jit.createProgram(&prog, src.c_str(), "sysml_fused.cu", 0, nullptr, nullptr) != 0)
        return sysml_set_error(ctx, SYSML_E_CUDA, "fused_cellwise: nvrtcCreateProgram failed");
      const char* opts[] = {"--gpu-architecture=sm_100a", "-std=c++17", "-lineinfo"};
      const int crc = g_jit.compileProgram(prog, 3, opts);
      if (crc != 0) {
        size_t ls = 0;
        g_jit.getLogSize(prog, &ls);
        std::string log(ls ? ls : 1, '\0');
        if (ls) g_jit.getLog(prog, &log[0]);
        g_jit.destroyProgram(&prog);
        return sysml_set_error(ctx, SYSML_E_CUDA, "fused_cellwise: NVRTC compile failed (%d): %.380s", crc, log.c_str());
      }
      size_t cs = 0;
      g_jit.getCubinSize(prog, &cs);
      std::vector<char> cubin(cs);
      g_jit.getCubin(prog, cubin.data());
      g_jit.destroyProgram(&prog);
      SYSML_CUDA(ctx, cudaFree(0));  // makes sure the runtime's primary context is current for the driver calls
      void* mod = nullptr;
      int drc = g_jit.moduleLoadData(&mod, cubin.data());
      if (drc == 0) drc = g_jit.moduleGetFunction(&fn, mod, "fused_kernel");
      if (drc != 0) {
        const char* es = nullptr;
        g_jit.getErrorString(drc, &es);
        return sysml_set_error(ctx, SYSML_E_CUDA, "fused_cellwise: loading the compiled kernel failed: %s", es ? es : "?");
      }
      g_kernels[key] = fn;  // the module stays loaded for the life of the process
    }
  }

  // ---- launch ----
  FusedArgsHost fa{};
  for (int i = 0; i < n_inputs; i++) fa.in[i] = inputs[i];
  for (int i = 0; i < n_consts; i++) fa.k[i] = consts[i];
  fa.out = out;
  fa.rows = rows;
  fa.cols = cols;
  i64 grid;
  const i64 cap = (i64)ctx->sm_count * 8;
  if (shape == 2) grid = (rows + (256 / W) - 1) / (256 / W);
  else if (shape == 4) grid = (rows + 8 * RW - 1) / (8 * RW);
  else if (shape == 3) grid = (rows + 7) / 8;
  else grid = ((vec2 ? len / 2 : len) + 255) / 256;
  if (grid > cap) grid = cap;
  if (grid < 1) grid = 1;
  if (shape == 3 && grid > 2 * ctx->sm_count) grid = 2 * ctx->sm_count;
  if (kagg == 1 || kagg == 3) {
    void* ws = nullptr;
    int rc = sysml_workspace(ctx, (size_t)grid * 2 * (kagg == 3 ? (size_t)cols : (n_out > 0 ? (size_t)n_out : 1)) * sizeof(double), &ws);
    if (rc) return rc;
    fa.partial = (double*)ws;
  }
  void* params[] = {&fa};
  const int lrc = g_jit.launchKernel(fn, (unsigned)grid, 1, 1, 256, 1, 1, 0, (void*)ctx->stream, params, nullptr);
  if (lrc != 0) {
    const char* es = nullptr;
    g_jit.getErrorString(lrc, &es);
    return sysml_set_error(ctx, SYSML_E_CUDA, "fused_cellwise: launch failed: %s", es ? es : "?");
  }
  ctx->launches++;
  if (kagg == 1) {
    if (aop == 0) fused_all_finish<0><<<n_out > 0 ? n_out : 1, 32, 0, ctx->stream>>>(fa.partial, (int)grid, out);
    else if (aop == 1) fused_all_finish<1><<<1, 32, 0, ctx->stream>>>(fa.partial, (int)grid, out);
    else fused_all_finish<2><<<1, 32, 0, ctx->stream>>>(fa.partial, (int)grid, out);
    SYSML_LAUNCH_CHECK(ctx, "fused_all_finish");
  } else if (kagg == 3) {
    const unsigned fg = (unsigned)((cols + 255) / 256);
    if (aop == 0) fused_col_finish<0><<<fg, 256, 0, ctx->stream>>>(fa.partial, (int)grid, cols, out);
    else if (aop == 1) fused_col_finish<1><<<fg, 256, 0, ctx->stream>>>(fa.partial, (int)grid, cols, out);
    else fused_col_finish<2><<<fg, 256, 0, ctx->stream>>>(fa.partial, (int)grid, cols, out);
    SYSML_LAUNCH_CHECK(ctx, "fused_col_finish");
  }
  return SYSML_OK;
}

extern "C" int sysml_b200_fused_cellwise(sysml_b200_ctx* ctx, const double* const* inputs, const int* input_kinds, int n_inputs,
                                         const double* consts, int n_consts, const int32_t* program, int n_tokens, int64_t rows,
                                         int64_t cols, int agg_in, double* out) {
  return fused_run(ctx, inputs, input_kinds, n_inputs, consts, n_consts, program, n_tokens, rows, cols, agg_in, 0, out);
}

extern "C" int sysml_b200_fused_multi_agg(sysml_b200_ctx* ctx, const double* const* inputs, const int* input_kinds, int n_inputs,
                                          const double* consts, int n_consts, const int32_t* program, int n_tokens, int64_t rows,
                                          int64_t cols, int n_out, double* out) {
  if (ctx && (n_out < 1 || n_out > SYSML_FUSED_MAX_RESULTS))
    return sysml_set_error(ctx, SYSML_E_LIMIT, "fused_multi_agg: between 1 and %d results", SYSML_FUSED_MAX_RESULTS);
  return fused_run(ctx, inputs, input_kinds, n_inputs, consts, n_consts, program, n_tokens, rows, cols, SYSML_FUSED_AGG_SUM, n_out, out);
}
