// gemm.cu -- FP64 tensor-core (DMMA) contractions: tsmm t(X)%*%X / X%*%t(X) and general dense matmult.
//
// Reference semantics: LibMatrixMult.matrixMultTransposeSelfDense (LibMatrixMult.java:1686-1837) +
// copyUpperToLowerTriangle (:3633-3672, incl. the exact nnz count) for tsmm; matrixMultDenseDense
// (:880-1103) for matmult.  GPU twins being replaced: cublasDsyrk + copy_u2l_dense
// (LibMatrixCUDA.java:477-524; SystemML.cu:413) and cublasDgemm/Dgemv/Ddot (LibMatrixCuMatMult.java:362-408).
//
// Design (B200).  tcgen05 / TMEM have no f64 kind; FP64 tensor work on sm_100a is the warp-level
// mma.sync m8n8k4 path (SASS DMMA.8x8x4).  FP64 is so slow relative to shared-memory and L2 bandwidth
// (ridge ~6 flop/B) that the kernel is DMMA-issue bound; the job of everything else is to stay out of
// the way: 4-stage cp.async (LDGSTS) pipeline, padded shared-memory tiles with leading dimensions
// = 4 (mod 16) doubles so that every fragment load is bank-conflict free for all four operand layouts,
// one __syncthreads per 16-deep k block, 128 DMMA per warp between barriers.
//   - tsmm computes only tiles with tj >= ti (the upper triangle); the finishing kernel sums the split-K
//     partials in a fixed order, writes both triangles (replaces copy_u2l_dense) and counts non-zeros.
//   - the contraction dimension (rows of X for tsmm: 10^7) is split across CTAs so that the grid is a
//     multiple of the SM count; partials are reduced deterministically, not with atomics.
// Algorithmic flops: tsmm n*m*(m+1); matmult 2*M*N*K.
#include <stdlib.h>

#include <type_traits>

#include "common.cuh"

int sysml_dense_mv(sysml_b200_ctx* ctx, const double* X, const double* v, double* out, i64 n, i64 m);
int sysml_dense_vm(sysml_b200_ctx* ctx, const double* X, const double* y, double* out, i64 n, i64 m);

namespace {

constexpr int BK = 16;
constexpr int STAGES = 4;

struct GemmArgs {
  const double* A;
  const double* B;
  double* out;  // C (splits == 1 && !sym) or partial base
  i64 M, N, K;
  i64 lda, ldb;
  i64 k_per_split;
  int tiles_m, tiles_n;
  int sym;      // only tiles tj >= ti
  int vecA, vecB;
  int diag_last;  // ws kernel, sym: 1-D grid, the (cheaper) diagonal tiles come last (see dgemm_ws_kernel)
  int nsplits;
  int edge_diag;  // 1 when the last diagonal tile hangs over the edge of the matrix
};

__device__ __forceinline__ void cp_async16(void* dst, const void* src, int src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_u32(dst)), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async8(void* dst, const void* src, int src_bytes) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;" ::"r"(smem_u32(dst)), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
// Copies a logical [R][Ccols] tile whose rows are contiguous in global memory (row r at src + r*ld)
// into shared memory with leading dimension LD; rows >= rvalid and columns >= cvalid are zero-filled.
template <int R, int CCOLS, int LD, int NT>
__device__ __forceinline__ void load_tile(double* s, const double* src, i64 ld, i64 rvalid, i64 cvalid, int vec, int tid) {
  if (vec) {
    constexpr int CH = CCOLS / 2;
    for (int c = tid; c < R * CH; c += NT) {
      const int r = c / CH, cc = (c - r * CH) * 2;
      i64 nb = (r < rvalid) ? (cvalid - cc) * 8 : 0;
      nb = nb < 0 ? 0 : (nb > 16 ? 16 : nb);
      cp_async16(s + r * LD + cc, nb > 0 ? (const void*)(src + (i64)r * ld + cc) : (const void*)src, (int)nb);
    }
  } else {
    for (int c = tid; c < R * CCOLS; c += NT) {
      const int r = c / CCOLS, cc = c - r * CCOLS;
      const bool ok = (r < rvalid) && (cc < cvalid);
      cp_async8(s + r * LD + cc, ok ? (const void*)(src + (i64)r * ld + cc) : (const void*)src, ok ? 8 : 0);
    }
  }
}

template <bool TA, bool TB, int BM, int BN, int WARPS_M, int WARPS_N, int BKT = BK, int NST = STAGES>
struct GemmCfg {
  static constexpr int NT = WARPS_M * WARPS_N * 32;
  static constexpr int WM = BM / WARPS_M, WN = BN / WARPS_N;
  static constexpr int MT = WM / 8, NTL = WN / 8;
  static constexpr int LDA = TA ? (BM + 4) : (BKT + 4);
  static constexpr int LDB = TB ? (BKT + 4) : (BN + 4);
  static constexpr int A_ELEMS = TA ? BKT * LDA : BM * LDA;
  static constexpr int B_ELEMS = TB ? BN * LDB : BKT * LDB;
  static constexpr size_t SMEM = (size_t)NST * (A_ELEMS + B_ELEMS) * sizeof(double);
};

template <bool TA, bool TB, int BM, int BN, int WARPS_M, int WARPS_N>
__global__ void __launch_bounds__(WARPS_M* WARPS_N * 32) dgemm_kernel(const GemmArgs g) {
  using Cfg = GemmCfg<TA, TB, BM, BN, WARPS_M, WARPS_N>;
  constexpr int NT = Cfg::NT, MT = Cfg::MT, NTL = Cfg::NTL, LDA = Cfg::LDA, LDB = Cfg::LDB;
  extern __shared__ __align__(16) double smem[];
  double* As = smem;
  double* Bs = smem + STAGES * Cfg::A_ELEMS;

  // ---- which tile / which slice of K ----
  int ti, tj;
  if (g.sym) {
    int t = blockIdx.x;
    ti = 0;
    while (t >= g.tiles_n - ti) {
      t -= g.tiles_n - ti;
      ti++;
    }
    tj = ti + t;
  } else {
    ti = blockIdx.x / g.tiles_n;
    tj = blockIdx.x - ti * g.tiles_n;
  }
  const i64 i0 = (i64)ti * BM, j0 = (i64)tj * BN;
  const i64 kbeg = (i64)blockIdx.y * g.k_per_split;
  const i64 kend = (kbeg + g.k_per_split < g.K) ? kbeg + g.k_per_split : g.K;
  const int nkb = (int)((kend - kbeg + BK - 1) / BK);

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int wm0 = (warp / WARPS_N) * Cfg::WM, wn0 = (warp % WARPS_N) * Cfg::WN;
  const int gid = lane >> 2, tig = lane & 3;

  auto issue = [&](int kb) {
    const int s = kb % STAGES;
    const i64 k0 = kbeg + (i64)kb * BK;
    const i64 kvalid = kend - k0;  // may exceed BK
    double* as = As + s * Cfg::A_ELEMS;
    double* bs = Bs + s * Cfg::B_ELEMS;
    if (TA)
      load_tile<BK, BM, LDA, NT>(as, g.A + k0 * g.lda + i0, g.lda, kvalid, g.M - i0, g.vecA, tid);
    else
      load_tile<BM, BK, LDA, NT>(as, g.A + i0 * g.lda + k0, g.lda, g.M - i0, kvalid, g.vecA, tid);
    if (TB)
      load_tile<BN, BK, LDB, NT>(bs, g.B + j0 * g.ldb + k0, g.ldb, g.N - j0, kvalid, g.vecB, tid);
    else
      load_tile<BK, BN, LDB, NT>(bs, g.B + k0 * g.ldb + j0, g.ldb, kvalid, g.N - j0, g.vecB, tid);
  };

  double acc[MT][NTL][2];
#pragma unroll
  for (int a = 0; a < MT; a++)
#pragma unroll
    for (int b = 0; b < NTL; b++) acc[a][b][0] = acc[a][b][1] = 0.0;

#pragma unroll
  for (int s = 0; s < STAGES - 1; s++) {
    if (s < nkb) issue(s);
    cp_async_commit();
  }
  for (int kb = 0; kb < nkb; kb++) {
    cp_async_wait<STAGES - 2>();
    __syncthreads();
    if (kb + STAGES - 1 < nkb) issue(kb + STAGES - 1);
    cp_async_commit();
    const double* as = As + (kb % STAGES) * Cfg::A_ELEMS;
    const double* bs = Bs + (kb % STAGES) * Cfg::B_ELEMS;
#pragma unroll
    for (int k4 = 0; k4 < BK; k4 += 4) {
      double af[MT], bf[NTL];
#pragma unroll
      for (int a = 0; a < MT; a++) {
        const int i = wm0 + a * 8 + gid, k = k4 + tig;  // A[i][k]
        af[a] = TA ? as[k * LDA + i] : as[i * LDA + k];
      }
#pragma unroll
      for (int b = 0; b < NTL; b++) {
        const int j = wn0 + b * 8 + gid, k = k4 + tig;  // B[k][j]
        bf[b] = TB ? bs[j * LDB + k] : bs[k * LDB + j];
      }
#pragma unroll
      for (int a = 0; a < MT; a++)
#pragma unroll
        for (int b = 0; b < NTL; b++) dmma884(acc[a][b][0], acc[a][b][1], af[a], bf[b]);
    }
  }
  cp_async_wait<0>();

  // ---- epilogue: C fragment (row = gid, cols = 2*tig, 2*tig+1) ----
  double* outp = g.out + (size_t)blockIdx.y * (size_t)g.M * (size_t)g.N;
#pragma unroll
  for (int a = 0; a < MT; a++) {
    const i64 i = i0 + wm0 + a * 8 + gid;
    if (i >= g.M) continue;
#pragma unroll
    for (int b = 0; b < NTL; b++) {
      const i64 j = j0 + wn0 + b * 8 + 2 * tig;
      if (j < g.N) outp[i * g.N + j] = acc[a][b][0];
      if (j + 1 < g.N) outp[i * g.N + j + 1] = acc[a][b][1];
    }
  }
}

// ------------------------------------------------------------------------------------------------
// Warp-specialised variant: one producer warp feeds the ring with 1-D bulk copies (cp.async.bulk, the
// TMA engine, SASS UBLKCP) -- one copy per contiguous row segment, straight into the padded layout --
// and signals mbarriers; the consumer warps do nothing but LDS + DMMA.  No __syncthreads and no address
// arithmetic in the consumers' loop, so the DMMA pipe never waits for the loaders.
// Requirements (checked on the host, otherwise the cp.async kernel above is used): 16-byte aligned
// bases, even leading dimensions, even M and N.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void bulk_g2s_plain(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst_smem)),
               "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

// rows of the logical [R][CCOLS] tile are contiguous in global memory; copies `rvalid` rows of `cvalid`
// columns (cvalid even) and, when zero_tail is set, clears what is not copied.  Returns bytes issued by
// the whole warp (identical on every lane).
template <int R, int CCOLS, int LD>
__device__ __forceinline__ uint32_t tile_bytes(i64 rvalid, i64 cvalid) {
  const i64 rv = rvalid < R ? (rvalid < 0 ? 0 : rvalid) : R;
  const i64 cv = cvalid < CCOLS ? (cvalid < 0 ? 0 : cvalid) : CCOLS;
  return (uint32_t)(rv * cv * 8);
}
template <int R, int CCOLS, int LD>
__device__ __forceinline__ void tile_zero_tail(double* s, i64 rvalid, i64 cvalid, int lane) {
  const int rv = (int)(rvalid < R ? (rvalid < 0 ? 0 : rvalid) : R);
  const int cv = (int)(cvalid < CCOLS ? (cvalid < 0 ? 0 : cvalid) : CCOLS);
  if (rv == R && cv == CCOLS) return;
  for (int c = lane; c < R * CCOLS; c += 32) {
    const int r = c / CCOLS, cc = c - r * CCOLS;
    if (r >= rv || cc >= cv) s[r * LD + cc] = 0.0;
  }
}
template <int R, int CCOLS, int LD>
__device__ __forceinline__ void tile_issue(double* s, const double* src, i64 ld, i64 rvalid, i64 cvalid, uint64_t* bar, int lane) {
  const int rv = (int)(rvalid < R ? (rvalid < 0 ? 0 : rvalid) : R);
  const int cv = (int)(cvalid < CCOLS ? (cvalid < 0 ? 0 : cvalid) : CCOLS);
  if (cv <= 0) return;
  for (int r = lane; r < rv; r += 32) bulk_g2s_plain(s + r * LD, src + (i64)r * ld, (uint32_t)cv * 8u, bar);
}

// K-contiguous operands (rows of 16 doubles = 128 B): per-row bulk copies would be 128-byte requests,
// which the TMA engine retires far too slowly; the producer warp uses 16-byte cp.async (LDGSTS) instead,
// 8 lanes per row segment, zero-filling through the src-size operand, and lets the copies arrive on the
// stage's mbarrier (cp.async.mbarrier.arrive.noinc).
template <int R, int LD, int KC>  // KC doubles (KC/2 16-byte chunks) per row
__device__ __forceinline__ void tile_issue_cpasync(double* s, const double* src, i64 ld, i64 rvalid, i64 kvalid, int lane, int pw, int npw) {
  if constexpr (KC == 16) {
    // 8 chunks per row, 4 rows per warp instruction
    const int r0 = pw * 4 + (lane >> 3), cc = (lane & 7) * 2;  // producer warp pw takes rows pw*4.., step 4*npw
    i64 nb = (kvalid - cc) * 8;
    nb = nb < 0 ? 0 : (nb > 16 ? 16 : nb);
    const double* p = src + (i64)r0 * ld + cc;
    double* d = s + r0 * LD + cc;
    const i64 step = (i64)4 * npw * ld;
    const int dstep = 4 * npw * LD;
#pragma unroll 4
    for (int r = r0; r < R; r += 4 * npw) {
      const int bytes = (r < rvalid) ? (int)nb : 0;
      cp_async16(d, bytes ? (const void*)p : (const void*)src, bytes);
      p += step;
      d += dstep;
    }
  } else {
    // KC == 64: 32 chunks per row, one row (512 contiguous bytes) per warp instruction
    static_assert(KC == 64, "K-contiguous tiles are 16 or 64 deep");
    const int cc = lane * 2;
    i64 nb = (kvalid - cc) * 8;
    nb = nb < 0 ? 0 : (nb > 16 ? 16 : nb);
    const double* p = src + (i64)pw * ld + cc;
    double* d = s + pw * LD + cc;
#pragma unroll 4
    for (int r = pw; r < R; r += npw) {
      const int bytes = (r < rvalid) ? (int)nb : 0;
      cp_async16(d, bytes ? (const void*)p : (const void*)src, bytes);
      p += (i64)npw * ld;
      d += npw * LD;
    }
  }
}
// rows-long tile [R][CCOLS] with SHORT rows (< 1 KB): same idea, 16-byte chunks spread over the warp
template <int R, int CCOLS, int LD>
__device__ __forceinline__ void tile_issue_cpasync_rows(double* s, const double* src, i64 ld, i64 rvalid, i64 cvalid, int lane, int pw, int npw) {
  constexpr int CH = CCOLS / 2;
#pragma unroll 2
  for (int c = pw * 32 + lane; c < R * CH; c += 32 * npw) {
    const int r = c / CH, cc = (c - r * CH) * 2;
    i64 nb = (r < rvalid) ? (cvalid - cc) * 8 : 0;
    nb = nb < 0 ? 0 : (nb > 16 ? 16 : nb);
    cp_async16(s + r * LD + cc, nb ? (const void*)(src + (i64)r * ld + cc) : (const void*)src, (int)nb);
  }
}
__device__ __forceinline__ void cp_async_mbar_arrive(uint64_t* bar) {
  asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// ---- strip mapping of one diagonal 128x128 tile (see dgemm_ws_kernel): warp w owns fragment rows w and 15-w ----
template <bool TA, bool TB, class Cfg, int BKT, int NST>
struct StripCtx {
  const double* As;
  const double* Bs;
  uint64_t* full;
  uint64_t* empty;
  double* outp;
  i64 M, N, i0, j0;
  int nkb, lane, warp;
};

// LO0 / LO1: first fragment column of the two fragment rows.  Consumes nkb stages of the ring, then writes the
// fragments (the tile lies inside the matrix).
template <int LO0, int LO1, bool TA, bool TB, class Cfg, int BKT, int NST>
__device__ __forceinline__ void strip_tile(const StripCtx<TA, TB, Cfg, BKT, NST>& c, uint32_t& it) {
  constexpr int LDA = Cfg::LDA, LDB = Cfg::LDB;
  constexpr int A_K4 = TA ? 4 * LDA : 4;
  constexpr int B_NT = TB ? 8 * LDB : 8;
  constexpr int B_K4 = TB ? 4 : 4 * LDB;
  const int gid = c.lane >> 2, tig = c.lane & 3;
  const int a_off0 = TA ? (tig * LDA + LO0 * 8 + gid) : ((LO0 * 8 + gid) * LDA + tig);
  const int a_off1 = TA ? (tig * LDA + LO1 * 8 + gid) : ((LO1 * 8 + gid) * LDA + tig);
  const int b_off = TB ? (gid * LDB + tig) : (tig * LDB + gid);
  double acc0[16][2], acc1[16][2];  // entries below LO0 / LO1 are never touched and cost no registers
#pragma unroll
  for (int b = 0; b < 16; b++) acc0[b][0] = acc0[b][1] = acc1[b][0] = acc1[b][1] = 0.0;
  for (int kb = 0; kb < c.nkb; kb++, it++) {
    const int s = it % NST;
    const uint32_t ph = (it / NST) & 1u;
    mbar_wait(&c.full[s], ph);
    const double* as = c.As + s * Cfg::A_ELEMS;
    const double* bs = c.Bs + s * Cfg::B_ELEMS + b_off;
#pragma unroll
    for (int k4 = 0; k4 < BKT / 4; k4++) {
      const double a0 = as[k4 * A_K4 + a_off0], a1 = as[k4 * A_K4 + a_off1];
      double bf[16];
#pragma unroll
      for (int b = LO0; b < 16; b++) bf[b] = bs[k4 * B_K4 + b * B_NT];
#pragma unroll
      for (int b = LO0; b < 16; b++) {
        dmma884(acc0[b][0], acc0[b][1], a0, bf[b]);
        if (b >= LO1) dmma884(acc1[b][0], acc1[b][1], a1, bf[b]);
      }
    }
    __syncwarp();
    if (c.lane == 0) mbar_arrive(&c.empty[s]);
  }
  const i64 r0 = c.i0 + LO0 * 8 + gid, r1 = c.i0 + LO1 * 8 + gid;
#pragma unroll
  for (int b = LO0; b < 16; b++) {
    const i64 j = c.j0 + b * 8 + 2 * tig;
    *reinterpret_cast<double2*>(c.outp + r0 * c.N + j) = make_double2(acc0[b][0], acc0[b][1]);
    if (b >= LO1) *reinterpret_cast<double2*>(c.outp + r1 * c.N + j) = make_double2(acc1[b][0], acc1[b][1]);
  }
}

// Ragged DIAGONAL tile of a symmetric product (the last one of m = 1000, or the only one of m <= 128): NF x NF
// fragments exist and only those on or right of the diagonal are needed.  Fragment rows are paired (w, NF-1-w):
// NF+1 fragments per pair, a middle row alone when NF is odd, warps beyond the pairs only keep the ring going.
// R0 / R1 / NF are compile-time (R1 < 0: single row, R0 < 0: idle warp).
template <int R0, int R1, int NF, bool TA, bool TB, class Cfg, int BKT, int NST>
__device__ __noinline__ void strip_tri_tile(const StripCtx<TA, TB, Cfg, BKT, NST>& c, uint32_t& it) {
  constexpr int LDA = Cfg::LDA, LDB = Cfg::LDB;
  constexpr int A_K4 = TA ? 4 * LDA : 4;
  constexpr int B_NT = TB ? 8 * LDB : 8;
  constexpr int B_K4 = TB ? 4 : 4 * LDB;
  constexpr int LO = R0 < 0 ? 0 : R0, L1 = R1 < 0 ? 0 : R1;
  const int gid = c.lane >> 2, tig = c.lane & 3;
  const int a_off0 = TA ? (tig * LDA + LO * 8 + gid) : ((LO * 8 + gid) * LDA + tig);
  const int a_off1 = TA ? (tig * LDA + L1 * 8 + gid) : ((L1 * 8 + gid) * LDA + tig);
  const int b_off = TB ? (gid * LDB + tig) : (tig * LDB + gid);
  double acc0[16][2], acc1[16][2];
#pragma unroll
  for (int b = 0; b < 16; b++) acc0[b][0] = acc0[b][1] = acc1[b][0] = acc1[b][1] = 0.0;
  for (int kb = 0; kb < c.nkb; kb++, it++) {
    const int s = it % NST;
    const uint32_t ph = (it / NST) & 1u;
    mbar_wait(&c.full[s], ph);
    if constexpr (R0 >= 0) {
      const double* as = c.As + s * Cfg::A_ELEMS;
      const double* bs = c.Bs + s * Cfg::B_ELEMS + b_off;
#pragma unroll
      for (int k4 = 0; k4 < BKT / 4; k4++) {
        const double a0 = as[k4 * A_K4 + a_off0];
        double a1 = 0.0;
        if constexpr (R1 >= 0) a1 = as[k4 * A_K4 + a_off1];
        double bf[16];
#pragma unroll
        for (int b = R0; b < NF; b++) bf[b] = bs[k4 * B_K4 + b * B_NT];
#pragma unroll
        for (int b = R0; b < NF; b++) {
          dmma884(acc0[b][0], acc0[b][1], a0, bf[b]);
          if constexpr (R1 >= 0)
            if (b >= R1) dmma884(acc1[b][0], acc1[b][1], a1, bf[b]);
        }
      }
    }
    __syncwarp();
    if (c.lane == 0) mbar_arrive(&c.empty[s]);
  }
  if constexpr (R0 >= 0) {
    const i64 r0 = c.i0 + R0 * 8 + gid, r1 = c.i0 + L1 * 8 + gid;
#pragma unroll
    for (int b = R0; b < NF; b++) {
      const i64 j = c.j0 + b * 8 + 2 * tig;
      if (r0 < c.M) {
        if (j + 1 < c.N) *reinterpret_cast<double2*>(c.outp + r0 * c.N + j) = make_double2(acc0[b][0], acc0[b][1]);
        else if (j < c.N) c.outp[r0 * c.N + j] = acc0[b][0];
      }
      if constexpr (R1 >= 0) {
        if (b >= R1 && r1 < c.M) {
          if (j + 1 < c.N) *reinterpret_cast<double2*>(c.outp + r1 * c.N + j) = make_double2(acc1[b][0], acc1[b][1]);
          else if (j < c.N) c.outp[r1 * c.N + j] = acc1[b][0];
        }
      }
    }
  }
}
template <int W, int NF, bool TA, bool TB, class Cfg, int BKT, int NST>
__device__ __forceinline__ void strip_tri_warp(const StripCtx<TA, TB, Cfg, BKT, NST>& c, uint32_t& it) {
  constexpr int P = NF - 1 - W;  // partner row
  if constexpr (W < P) strip_tri_tile<W, P, NF>(c, it);
  else if constexpr (W == P) strip_tri_tile<W, -1, NF>(c, it);
  else strip_tri_tile<-1, -1, NF>(c, it);
}
template <int NF, bool TA, bool TB, class Cfg, int BKT, int NST>
__device__ __forceinline__ void strip_tri_dispatch(const StripCtx<TA, TB, Cfg, BKT, NST>& c, uint32_t& it) {
  switch (c.warp) {
    case 0: strip_tri_warp<0, NF>(c, it); break;
    case 1: strip_tri_warp<1, NF>(c, it); break;
    case 2: strip_tri_warp<2, NF>(c, it); break;
    case 3: strip_tri_warp<3, NF>(c, it); break;
    case 4: strip_tri_warp<4, NF>(c, it); break;
    case 5: strip_tri_warp<5, NF>(c, it); break;
    case 6: strip_tri_warp<6, NF>(c, it); break;
    default: strip_tri_warp<7, NF>(c, it); break;
  }
}

// Off-diagonal tile that hangs over the right edge of the matrix: only NF of the 16 fragment columns exist
// (m = 1000: 13; m = 5000: 1).  Same strip mapping -- warp w owns fragment rows w and 15-w -- with the column
// count a compile-time constant, so the missing fragments cost neither DMMAs nor registers.
template <int NF, bool TA, bool TB, class Cfg, int BKT, int NST>
__device__ __forceinline__ void strip_edge_tile(const StripCtx<TA, TB, Cfg, BKT, NST>& c, uint32_t& it) {
  constexpr int LDA = Cfg::LDA, LDB = Cfg::LDB;
  constexpr int A_K4 = TA ? 4 * LDA : 4;
  constexpr int B_NT = TB ? 8 * LDB : 8;
  constexpr int B_K4 = TB ? 4 : 4 * LDB;
  const int gid = c.lane >> 2, tig = c.lane & 3;
  const int fa0 = c.warp, fa1 = 15 - c.warp;
  const int a_off0 = TA ? (tig * LDA + fa0 * 8 + gid) : ((fa0 * 8 + gid) * LDA + tig);
  const int a_off1 = TA ? (tig * LDA + fa1 * 8 + gid) : ((fa1 * 8 + gid) * LDA + tig);
  const int b_off = TB ? (gid * LDB + tig) : (tig * LDB + gid);
  double acc0[NF][2], acc1[NF][2];
#pragma unroll
  for (int b = 0; b < NF; b++) acc0[b][0] = acc0[b][1] = acc1[b][0] = acc1[b][1] = 0.0;
  for (int kb = 0; kb < c.nkb; kb++, it++) {
    const int s = it % NST;
    const uint32_t ph = (it / NST) & 1u;
    mbar_wait(&c.full[s], ph);
    const double* as = c.As + s * Cfg::A_ELEMS;
    const double* bs = c.Bs + s * Cfg::B_ELEMS + b_off;
#pragma unroll
    for (int k4 = 0; k4 < BKT / 4; k4++) {
      const double a0 = as[k4 * A_K4 + a_off0], a1 = as[k4 * A_K4 + a_off1];
      double bf[NF];
#pragma unroll
      for (int b = 0; b < NF; b++) bf[b] = bs[k4 * B_K4 + b * B_NT];
#pragma unroll
      for (int b = 0; b < NF; b++) {
        dmma884(acc0[b][0], acc0[b][1], a0, bf[b]);
        dmma884(acc1[b][0], acc1[b][1], a1, bf[b]);
      }
    }
    __syncwarp();
    if (c.lane == 0) mbar_arrive(&c.empty[s]);
  }
  const i64 r0 = c.i0 + fa0 * 8 + gid, r1 = c.i0 + fa1 * 8 + gid;
#pragma unroll
  for (int b = 0; b < NF; b++) {
    const i64 j = c.j0 + b * 8 + 2 * tig;
    if (j + 1 < c.N) {
      *reinterpret_cast<double2*>(c.outp + r0 * c.N + j) = make_double2(acc0[b][0], acc0[b][1]);
      *reinterpret_cast<double2*>(c.outp + r1 * c.N + j) = make_double2(acc1[b][0], acc1[b][1]);
    } else if (j < c.N) {
      c.outp[r0 * c.N + j] = acc0[b][0];
      c.outp[r1 * c.N + j] = acc1[b][0];
    }
  }
}

// Small output (M <= 64, N <= 128: t(P) %*% X of Kmeans, 50 x 100): the whole product is one 64 x 128 tile and
// every CTA a slice of the contraction.  Warp w owns fragment row w and the NF fragment columns that exist
// (compile time); warps whose rows lie past M only keep the ring protocol going.  On a 50 x 100 output that is
// 13 DMMAs per k-step and warp instead of the 16 of a 2 x 4 warp grid that cannot skip its padding.
template <int NF, bool TA, bool TB, class Cfg, int BKT, int NST>
__device__ __forceinline__ void strip_small_tile(const StripCtx<TA, TB, Cfg, BKT, NST>& c, uint32_t& it) {
  constexpr int LDA = Cfg::LDA, LDB = Cfg::LDB;
  constexpr int A_K4 = TA ? 4 * LDA : 4;
  constexpr int B_NT = TB ? 8 * LDB : 8;
  constexpr int B_K4 = TB ? 4 : 4 * LDB;
  const int gid = c.lane >> 2, tig = c.lane & 3;
  const int fa = c.warp;
  const bool active = c.i0 + fa * 8 < c.M;
  const int a_off = TA ? (tig * LDA + fa * 8 + gid) : ((fa * 8 + gid) * LDA + tig);
  const int b_off = TB ? (gid * LDB + tig) : (tig * LDB + gid);
  double acc[NF][2];
#pragma unroll
  for (int b = 0; b < NF; b++) acc[b][0] = acc[b][1] = 0.0;
  for (int kb = 0; kb < c.nkb; kb++, it++) {
    const int s = it % NST;
    const uint32_t ph = (it / NST) & 1u;
    mbar_wait(&c.full[s], ph);
    if (active) {
      const double* as = c.As + s * Cfg::A_ELEMS + a_off;
      const double* bs = c.Bs + s * Cfg::B_ELEMS + b_off;
#pragma unroll
      for (int k4 = 0; k4 < BKT / 4; k4++) {
        const double a0 = as[k4 * A_K4];
        double bf[NF];
#pragma unroll
        for (int b = 0; b < NF; b++) bf[b] = bs[k4 * B_K4 + b * B_NT];
#pragma unroll
        for (int b = 0; b < NF; b++) dmma884(acc[b][0], acc[b][1], a0, bf[b]);
      }
    }
    __syncwarp();
    if (c.lane == 0) mbar_arrive(&c.empty[s]);
  }
  const i64 r = c.i0 + fa * 8 + gid;
  if (r < c.M) {
#pragma unroll
    for (int b = 0; b < NF; b++) {
      const i64 j = c.j0 + b * 8 + 2 * tig;
      if (j + 1 < c.N) *reinterpret_cast<double2*>(c.outp + r * c.N + j) = make_double2(acc[b][0], acc[b][1]);
      else if (j < c.N) c.outp[r * c.N + j] = acc[b][0];
    }
  }
}

__device__ __forceinline__ void tile_coords(const GemmArgs& g, int t, int& ti, int& tj) {
  if (g.sym && g.diag_last) {  // strictly upper tiles row by row, then the diagonal (its ragged last tile first)
    const int T = g.tiles_n, noff = T * (T - 1) / 2;
    if (t >= noff) {
      const int d = t - noff;
      ti = tj = g.edge_diag ? (d == 0 ? T - 1 : d - 1) : d;
      return;
    }
    ti = 0;
    while (t >= T - 1 - ti) {
      t -= T - 1 - ti;
      ti++;
    }
    tj = ti + 1 + t;
  } else if (g.sym) {  // upper-triangular tile list, row by row
    ti = 0;
    while (t >= g.tiles_n - ti) {
      t -= g.tiles_n - ti;
      ti++;
    }
    tj = ti + t;
  } else {
    ti = t / g.tiles_n;
    tj = t - ti * g.tiles_n;
  }
}

template <bool TA, bool TB, int BM, int BN, int WARPS_M, int WARPS_N, int BKT, int NST, int NPW, bool STRIP>
__global__ void __launch_bounds__((WARPS_M * WARPS_N + NPW) * 32,
                                   // an explicit minimum: without one ptxas has at times settled on 32 registers and
                                   // kilobytes of spills for this kernel.  Two CTAs per SM where the stages leave room.
                                   (2 * (GemmCfg<TA, TB, BM, BN, WARPS_M, WARPS_N, BKT, NST>::SMEM + 1152) <= 228 * 1024) ? 2 : 1)
    dgemm_ws_kernel(const GemmArgs g) {
  using Cfg = GemmCfg<TA, TB, BM, BN, WARPS_M, WARPS_N, BKT, NST>;
  constexpr int NCW = WARPS_M * WARPS_N;  // consumer warps
  constexpr int MT = Cfg::MT, NTL = Cfg::NTL, LDA = Cfg::LDA, LDB = Cfg::LDB;
  constexpr bool A_BULK = TA && (BM * 8 >= 1024);
  constexpr bool B_BULK = !TB && (BN * 8 >= 1024);
  constexpr bool ANY_CPASYNC = !A_BULK || !B_BULK;
  extern __shared__ __align__(16) double smem[];
  uint64_t* full = reinterpret_cast<uint64_t*>(smem);
  uint64_t* empty = full + NST;
  double* As = smem + 16;  // 128 bytes reserved for the barriers
  double* Bs = As + NST * Cfg::A_ELEMS;

  // output tiles are walked with a grid stride: with gridDim.x == ntiles this is one tile per CTA; the direct
  // (no split-K) launches use one CTA per SM slot so that the copy pipeline never drains between tiles
  const int ntiles = g.sym ? g.tiles_n * (g.tiles_n + 1) / 2 : g.tiles_m * g.tiles_n;
  // diag_last launches are one-dimensional: first every (split, off-diagonal tile) with the tile index fastest --
  // concurrent CTAs then read the same k range and share its panels through L2 -- and only then the cheaper
  // (split, diagonal tile) CTAs, so that the short jobs fill the tail of the schedule
  int split = blockIdx.y, tile0 = blockIdx.x, tstep = gridDim.x;
  if (g.diag_last) {
    // expensive tiles: the strictly upper ones and a ragged last diagonal tile (it takes the generic path)
    const int T = g.tiles_n, n1 = T * (T - 1) / 2 + g.edge_diag, n2 = T - g.edge_diag, nsplit = g.nsplits;
    int lin = blockIdx.x;
    if (lin < n1 * nsplit) {
      split = lin / n1;
      tile0 = lin - split * n1;
    } else {
      lin -= n1 * nsplit;
      split = lin / n2;
      tile0 = n1 + (lin - split * n2);
    }
    tstep = ntiles;  // one tile per CTA
  }
  const i64 kbeg = (i64)split * g.k_per_split;
  const i64 kend = (kbeg + g.k_per_split < g.K) ? kbeg + g.k_per_split : g.K;
  const int nkb = (int)((kend - kbeg + BKT - 1) / BKT);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

  // clear all tiles once: padding columns/rows that no copy ever writes must read as zero
  for (int c = tid; c < NST * (Cfg::A_ELEMS + Cfg::B_ELEMS); c += (NCW + NPW) * 32) As[c] = 0.0;
  if (tid == 0) {
    for (int s = 0; s < NST; s++) {
      mbar_init(&full[s], ANY_CPASYNC ? 1 + 32 * NPW : 1);  // lane 0's expect_tx arrival (+ one cp.async arrival per producer lane)
      mbar_init(&empty[s], NCW);
    }
    mbar_fence_init();
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncthreads();

  if (warp >= NCW) {
    // ===================== producer warps (warp NCW does the bulk copies; all NPW share the cp.async rows) =====================
    const int pw = warp - NCW;
    uint32_t it = 0;  // k blocks issued so far, over all tiles of this CTA
    for (int tile = tile0; tile < ntiles; tile += tstep) {
    int ti, tj;
    tile_coords(g, tile, ti, tj);
    const i64 i0 = (i64)ti * BM, j0 = (i64)tj * BN;
    for (int kb = 0; kb < nkb; kb++, it++) {
      const int s = it % NST;
      const uint32_t ph = (it / NST) & 1u;
      mbar_wait(&empty[s], ph ^ 1u);
      const i64 k0 = kbeg + (i64)kb * BKT;
      const i64 kvalid = kend - k0;
      double* as = As + s * Cfg::A_ELEMS;
      double* bs = Bs + s * Cfg::B_ELEMS;
      // bulk copies only for operands whose tile rows are >= 1 KB; 16-byte cp.async for everything else
      if (pw == 0) {
        if (kvalid < BKT) {  // ragged last k block: clear what the bulk copies below will not overwrite
          if (A_BULK) tile_zero_tail<BKT, BM, LDA>(as, kvalid, g.M - i0, lane);
          if (B_BULK) tile_zero_tail<BKT, BN, LDB>(bs, kvalid, g.N - j0, lane);
          asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        }
        uint32_t bytes = 0;
        if (A_BULK) bytes += tile_bytes<BKT, BM, LDA>(kvalid, g.M - i0);
        if (B_BULK) bytes += tile_bytes<BKT, BN, LDB>(kvalid, g.N - j0);
        __syncwarp();
        if (lane == 0) mbar_expect_tx(&full[s], bytes);
        __syncwarp();
        if (A_BULK) tile_issue<BKT, BM, LDA>(as, g.A + k0 * g.lda + i0, g.lda, kvalid, g.M - i0, &full[s], lane);
        if (B_BULK) tile_issue<BKT, BN, LDB>(bs, g.B + k0 * g.ldb + j0, g.ldb, kvalid, g.N - j0, &full[s], lane);
      }
      if (!A_BULK) {
        if (TA) tile_issue_cpasync_rows<BKT, BM, LDA>(as, g.A + k0 * g.lda + i0, g.lda, kvalid, g.M - i0, lane, pw, NPW);
        else    tile_issue_cpasync<BM, LDA, BKT>(as, g.A + i0 * g.lda + k0, g.lda, g.M - i0, kvalid, lane, pw, NPW);
      }
      if (!B_BULK) {
        if (!TB) tile_issue_cpasync_rows<BKT, BN, LDB>(bs, g.B + k0 * g.ldb + j0, g.ldb, kvalid, g.N - j0, lane, pw, NPW);
        else     tile_issue_cpasync<BN, LDB, BKT>(bs, g.B + j0 * g.ldb + k0, g.ldb, g.N - j0, kvalid, lane, pw, NPW);
      }
      if (ANY_CPASYNC) cp_async_mbar_arrive(&full[s]);
    }
    }
    if (ANY_CPASYNC) cp_async_wait<0>();
    return;
  }

  // ===================== consumer warps =====================
  const int wm0 = (warp / WARPS_N) * Cfg::WM, wn0 = (warp % WARPS_N) * Cfg::WN;
  const int gid = lane >> 2, tig = lane & 3;
  // per-thread fragment base offsets (constant over the whole loop)
  const int a_off = TA ? (tig * LDA + wm0 + gid) : ((wm0 + gid) * LDA + tig);
  const int b_off = TB ? ((wn0 + gid) * LDB + tig) : (tig * LDB + wn0 + gid);
  constexpr int A_MT = TA ? 8 : 8 * LDA;       // step between m-tiles
  constexpr int A_K4 = TA ? 4 * LDA : 4;       // step between k4 blocks
  constexpr int B_NT = TB ? 8 * LDB : 8;
  constexpr int B_K4 = TB ? 4 : 4 * LDB;

  uint32_t it = 0;
  for (int tile = tile0; tile < ntiles; tile += tstep) {
  int ti, tj;
  tile_coords(g, tile, ti, tj);
  const i64 i0 = (i64)ti * BM, j0 = (i64)tj * BN;
  if constexpr (STRIP && BM == 64) {
    static_assert(!(STRIP && BM == 64) || (BN == 128 && WARPS_M * WARPS_N == 8), "small-output strip mapping: 64x128 tile, 8 warps");
    double* outd = g.out + (size_t)split * (size_t)g.M * (size_t)g.N;
    const StripCtx<TA, TB, Cfg, BKT, NST> sc{As, Bs, full, empty, outd, g.M, g.N, i0, j0, nkb, lane, warp};
    const i64 ncols = (g.N - j0 < BN) ? g.N - j0 : BN;
    switch ((int)((ncols + 7) / 8)) {
      case 1: strip_small_tile<1>(sc, it); break;
      case 2: strip_small_tile<2>(sc, it); break;
      case 3: strip_small_tile<3>(sc, it); break;
      case 4: strip_small_tile<4>(sc, it); break;
      case 5: strip_small_tile<5>(sc, it); break;
      case 6: strip_small_tile<6>(sc, it); break;
      case 7: strip_small_tile<7>(sc, it); break;
      case 8: strip_small_tile<8>(sc, it); break;
      case 9: strip_small_tile<9>(sc, it); break;
      case 10: strip_small_tile<10>(sc, it); break;
      case 11: strip_small_tile<11>(sc, it); break;
      case 12: strip_small_tile<12>(sc, it); break;
      case 13: strip_small_tile<13>(sc, it); break;
      case 14: strip_small_tile<14>(sc, it); break;
      case 15: strip_small_tile<15>(sc, it); break;
      default: strip_small_tile<16>(sc, it); break;
    }
    continue;
  }
  if constexpr (STRIP && BM == 128) {
    // Diagonal tiles of a symmetric product: only fragments on or right of the diagonal are needed.  The tile is
    // 16 x 16 fragments of 8x8; warp w takes fragment rows w and 15-w, which need 16-w and w+1 fragment columns:
    // 17 DMMAs per k-step for every warp instead of 32.  The column ranges are compile-time (one instantiation
    // per warp index): a predicated-off DMMA still occupies the tensor pipe (measured), and ptxas if-converts
    // short branches, so the skipped fragments must not exist in the code at all.
    static_assert(!(STRIP && BM == 128) || (BN == 128 && WARPS_M * WARPS_N == 8), "strip mapping is written for 128x128 tiles and 8 warps");
    if (g.sym && ti == tj && i0 + BM <= g.M) {
      double* outd = g.out + (size_t)split * (size_t)g.M * (size_t)g.N;
      const StripCtx<TA, TB, Cfg, BKT, NST> sc{As, Bs, full, empty, outd, g.M, g.N, i0, j0, nkb, lane, warp};
      switch (warp) {
        case 0: strip_tile<0, 15>(sc, it); break;
        case 1: strip_tile<1, 14>(sc, it); break;
        case 2: strip_tile<2, 13>(sc, it); break;
        case 3: strip_tile<3, 12>(sc, it); break;
        case 4: strip_tile<4, 11>(sc, it); break;
        case 5: strip_tile<5, 10>(sc, it); break;
        case 6: strip_tile<6, 9>(sc, it); break;
        default: strip_tile<7, 8>(sc, it); break;
      }
      continue;
    }
    if (g.sym && ti == tj && i0 + BM > g.M) {  // ragged diagonal tile
      double* outd = g.out + (size_t)split * (size_t)g.M * (size_t)g.N;
      const StripCtx<TA, TB, Cfg, BKT, NST> sc{As, Bs, full, empty, outd, g.M, g.N, i0, j0, nkb, lane, warp};
      switch ((int)((g.N - j0 + 7) / 8)) {
        case 1: strip_tri_dispatch<1>(sc, it); break;
        case 2: strip_tri_dispatch<2>(sc, it); break;
        case 3: strip_tri_dispatch<3>(sc, it); break;
        case 4: strip_tri_dispatch<4>(sc, it); break;
        case 5: strip_tri_dispatch<5>(sc, it); break;
        case 6: strip_tri_dispatch<6>(sc, it); break;
        case 7: strip_tri_dispatch<7>(sc, it); break;
        case 8: strip_tri_dispatch<8>(sc, it); break;
        case 9: strip_tri_dispatch<9>(sc, it); break;
        case 10: strip_tri_dispatch<10>(sc, it); break;
        case 11: strip_tri_dispatch<11>(sc, it); break;
        case 12: strip_tri_dispatch<12>(sc, it); break;
        case 13: strip_tri_dispatch<13>(sc, it); break;
        case 14: strip_tri_dispatch<14>(sc, it); break;
        default: strip_tri_dispatch<15>(sc, it); break;
      }
      continue;
    }
    if (g.sym && ti != tj && i0 + BM <= g.M && j0 + BN > g.N) {  // last tile column of a ragged matrix
      double* outd = g.out + (size_t)split * (size_t)g.M * (size_t)g.N;
      const StripCtx<TA, TB, Cfg, BKT, NST> sc{As, Bs, full, empty, outd, g.M, g.N, i0, j0, nkb, lane, warp};
      switch ((int)((g.N - j0 + 7) / 8)) {
        case 1: strip_edge_tile<1>(sc, it); break;
        case 2: strip_edge_tile<2>(sc, it); break;
        case 3: strip_edge_tile<3>(sc, it); break;
        case 4: strip_edge_tile<4>(sc, it); break;
        case 5: strip_edge_tile<5>(sc, it); break;
        case 6: strip_edge_tile<6>(sc, it); break;
        case 7: strip_edge_tile<7>(sc, it); break;
        case 8: strip_edge_tile<8>(sc, it); break;
        case 9: strip_edge_tile<9>(sc, it); break;
        case 10: strip_edge_tile<10>(sc, it); break;
        case 11: strip_edge_tile<11>(sc, it); break;
        case 12: strip_edge_tile<12>(sc, it); break;
        case 13: strip_edge_tile<13>(sc, it); break;
        case 14: strip_edge_tile<14>(sc, it); break;
        case 15: strip_edge_tile<15>(sc, it); break;
        default: strip_edge_tile<16>(sc, it); break;
      }
      continue;
    }
  }
  double acc[MT][NTL][2];
#pragma unroll
  for (int a = 0; a < MT; a++)
#pragma unroll
    for (int b = 0; b < NTL; b++) acc[a][b][0] = acc[a][b][1] = 0.0;
  // one k block: wait for its stage, run n4 k4-steps (all BKT/4 of them unrolled, or only those of a ragged last
  // block that hold data -- K = 100: 1 of 4 -- the rest of the stage is zeros), release the stage
  auto k_block = [&](auto full_block, int n4) {
    const int s = it % NST;
    const uint32_t ph = (it / NST) & 1u;
    mbar_wait(&full[s], ph);
    const double* as = As + s * Cfg::A_ELEMS + a_off;
    const double* bs = Bs + s * Cfg::B_ELEMS + b_off;
    auto k4_step = [&](int k4) {
      double af[MT], bf[NTL];
#pragma unroll
      for (int a = 0; a < MT; a++) af[a] = as[k4 * A_K4 + a * A_MT];
#pragma unroll
      for (int b = 0; b < NTL; b++) bf[b] = bs[k4 * B_K4 + b * B_NT];
#pragma unroll
      for (int a = 0; a < MT; a++)
#pragma unroll
        for (int b = 0; b < NTL; b++) dmma884(acc[a][b][0], acc[a][b][1], af[a], bf[b]);
    };
    if constexpr (decltype(full_block)::value) {
#pragma unroll
      for (int k4 = 0; k4 < BKT / 4; k4++) k4_step(k4);
    } else {
#pragma unroll 1
      for (int k4 = 0; k4 < n4; k4++) k4_step(k4);
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(&empty[s]);
    it++;
  };
  const int nfull = (int)((kend - kbeg) / BKT);
  if (nfull == nkb) {
    // the common case keeps the plain loop (measured: routing it through the helper above costs tsmm 0.4 % and
    // t(X)%*%P 4 %)
    for (int kb = 0; kb < nkb; kb++, it++) {
      const int s = it % NST;
      const uint32_t ph = (it / NST) & 1u;
      mbar_wait(&full[s], ph);
      const double* as = As + s * Cfg::A_ELEMS + a_off;
      const double* bs = Bs + s * Cfg::B_ELEMS + b_off;
#pragma unroll
      for (int k4 = 0; k4 < BKT / 4; k4++) {
        double af[MT], bf[NTL];
#pragma unroll
        for (int a = 0; a < MT; a++) af[a] = as[k4 * A_K4 + a * A_MT];
#pragma unroll
        for (int b = 0; b < NTL; b++) bf[b] = bs[k4 * B_K4 + b * B_NT];
#pragma unroll
        for (int a = 0; a < MT; a++)
#pragma unroll
          for (int b = 0; b < NTL; b++) dmma884(acc[a][b][0], acc[a][b][1], af[a], bf[b]);
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty[s]);
    }
  } else {
    for (int kb = 0; kb < nfull; kb++) k_block(std::true_type{}, 0);
    k_block(std::false_type{}, (int)((kend - kbeg - (i64)nfull * BKT + 3) / 4));
  }

  double* outp = g.out + (size_t)split * (size_t)g.M * (size_t)g.N;
#pragma unroll
  for (int a = 0; a < MT; a++) {
    const i64 i = i0 + wm0 + a * 8 + gid;
    if (i >= g.M) continue;
#pragma unroll
    for (int b = 0; b < NTL; b++) {
      const i64 j = j0 + wn0 + b * 8 + 2 * tig;
      if (j + 1 < g.N) {
        *reinterpret_cast<double2*>(outp + i * g.N + j) = make_double2(acc[a][b][0], acc[a][b][1]);
      } else if (j < g.N) {
        outp[i * g.N + j] = acc[a][b][0];
      }
    }
  }
  }
}

// C[idx] = sum over splits (fixed order, compensated)
__global__ void reduce_partials_kernel(const double* __restrict__ partial, int splits, i64 len, double* __restrict__ C) {
  const i64 stride = (i64)gridDim.x * blockDim.x;
  for (i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x; i < len; i += stride) {
    CompSum a;
    for (int s = 0; s < splits; s++) a.add(ldg_stream(partial + (size_t)s * len + i));
    C[i] = a.value();
  }
}

// tsmm epilogue: sum partials over the upper triangle, write both triangles, count non-zeros
// (diagonal counts 1, off-diagonal 2: LibMatrixMult.java:3651-3668).
__global__ void tsmm_finish_kernel(const double* __restrict__ partial, int splits, i64 n, double* __restrict__ C,
                                   unsigned long long* __restrict__ nnz) {
  const i64 len = n * n;
  const i64 stride = (i64)gridDim.x * blockDim.x;
  unsigned long long cnt = 0;
  i64 idx = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx < len) {
    i64 i = idx / n, j = idx - i * n;
    const i64 di = stride / n, dj = stride - di * n;
    for (; idx < len; idx += stride) {
      if (j >= i) {
        CompSum a;
        for (int s = 0; s < splits; s++) a.add(ldg_stream(partial + (size_t)s * len + idx));
        const double v = a.value();
        C[idx] = v;
        if (j > i) C[j * n + i] = v;
        if (v != 0.0) cnt += (j == i) ? 1ull : 2ull;
      }
      i += di;
      j += dj;
      if (j >= n) {
        j -= n;
        i++;
      }
    }
  }
  for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
  if ((threadIdx.x & 31) == 0 && cnt) atomicAdd(nnz, cnt);
}

inline bool even_aligned(const double* p, i64 ld) { return ((reinterpret_cast<uintptr_t>(p) & 15) == 0) && ((ld & 1) == 0); }

struct Plan {
  int tiles_m, tiles_n, ntiles, splits;
  i64 k_per_split;
};
inline Plan make_plan(sysml_b200_ctx* ctx, i64 M, i64 N, i64 K, int BM, int BN, bool sym, int bk) {
  Plan p;
  p.tiles_m = (int)((M + BM - 1) / BM);
  p.tiles_n = (int)((N + BN - 1) / BN);
  p.ntiles = sym ? p.tiles_n * (p.tiles_n + 1) / 2 : p.tiles_m * p.tiles_n;
  const i64 kblocks = (K + bk - 1) / bk;
  // Aim at ~24 waves of CTAs (1 CTA per SM): the last, partly filled wave costs up to one CTA duration, 1/6 of
  // the run at 6 waves (measured on 1M x 5000: 820 whole-K CTAs = 5.5 waves ran as 6).  Keep >= 32 k-blocks (512
  // contraction steps) per CTA and at most 1 GiB of partial outputs (written and read once: < 0.5 ms).
  i64 want = ((i64)ctx->sm_count * 24 + p.ntiles - 1) / p.ntiles;
  i64 maxs = kblocks * bk / 512;
  const i64 by_bytes = ((i64)1 << 30) / (M * N * 8 > 0 ? M * N * 8 : 1);
  if (maxs > by_bytes) maxs = by_bytes;
  if (maxs < 1) maxs = 1;
  if (want > maxs) want = maxs;
  if (want < 1) want = 1;
  if (want > 1) {
    // round the CTA count to a multiple of the SM count when that costs little
    const i64 total = want * p.ntiles;
    const i64 rounded = (total / ctx->sm_count) * ctx->sm_count;
    if (rounded >= p.ntiles && rounded / p.ntiles >= 1) want = rounded / p.ntiles > 0 ? rounded / p.ntiles : want;
  }
  if (want > 65535) want = 65535;
  i64 kbs = (kblocks + want - 1) / want;
  p.k_per_split = kbs * bk;
  p.splits = (int)((K + p.k_per_split - 1) / p.k_per_split);
  if (p.splits < 1) p.splits = 1;
  return p;
}

template <bool TA, bool TB, int BM, int BN, int WARPS_M, int WARPS_N, int BKT = BK, int NST = STAGES, int NPW = 1, bool STRIP = false>
int run_gemm(sysml_b200_ctx* ctx, const double* A, i64 lda, const double* B, i64 ldb, double* C, i64 M, i64 N, i64 K, bool sym,
             int64_t* h_nnz) {
  using Cfg = GemmCfg<TA, TB, BM, BN, WARPS_M, WARPS_N>;           // legacy cp.async kernel (BK = 16, 4 stages)
  using WsCfg = GemmCfg<TA, TB, BM, BN, WARPS_M, WARPS_N, BKT, NST>;  // warp-specialised kernel
  auto kern = dgemm_kernel<TA, TB, BM, BN, WARPS_M, WARPS_N>;
  SYSML_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Cfg::SMEM));
  const Plan p = make_plan(ctx, M, N, K, BM, BN, sym, BKT);
  const bool direct = (p.splits == 1) && !sym;
  GemmArgs g{};
  g.A = A; g.B = B; g.M = M; g.N = N; g.K = K; g.lda = lda; g.ldb = ldb;
  g.k_per_split = p.k_per_split; g.tiles_m = p.tiles_m; g.tiles_n = p.tiles_n; g.sym = sym ? 1 : 0;
  g.vecA = even_aligned(A, lda) ? 1 : 0;
  g.vecB = even_aligned(B, ldb) ? 1 : 0;
  double* partial = nullptr;
  unsigned long long* d_nnz = nullptr;
  if (!direct) {
    void* ws = nullptr;
    int rc = sysml_workspace(ctx, 256 + (size_t)p.splits * M * N * sizeof(double), &ws);
    if (rc) return rc;
    d_nnz = (unsigned long long*)ws;
    partial = (double*)((char*)ws + 256);
    g.out = partial;
  } else {
    g.out = C;
  }
  // the warp-specialised bulk-copy kernel needs aligned, even-sized operands (see its header comment)
  const bool ws_ok = g.vecA && g.vecB && (M % 2 == 0) && (N % 2 == 0) &&
                     ((reinterpret_cast<uintptr_t>(g.out) & 15) == 0) && getenv("SYSML_B200_GEMM_NO_WS") == nullptr;
  if (ws_ok) {
    auto kws = dgemm_ws_kernel<TA, TB, BM, BN, WARPS_M, WARPS_N, BKT, NST, NPW, STRIP>;
    const size_t smem_ws = WsCfg::SMEM + 128;
    SYSML_CUDA(ctx, cudaFuncSetAttribute(kws, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_ws));
    int gx = p.ntiles;
    if (direct) {  // persistent over tiles: one CTA per resident slot
      int per_sm = 1;
      SYSML_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kws, WsCfg::NT + 32 * NPW, smem_ws));
      if (per_sm < 1) per_sm = 1;
      if (gx > per_sm * ctx->sm_count) gx = per_sm * ctx->sm_count;
    }
    if (sym && STRIP && (i64)p.ntiles * p.splits < (1ll << 31)) {
      g.diag_last = 1;
      g.nsplits = p.splits;
      g.edge_diag = (M % BM != 0) ? 1 : 0;
      kws<<<dim3((unsigned)(p.ntiles * p.splits)), WsCfg::NT + 32 * NPW, smem_ws, ctx->stream>>>(g);
    } else {
      kws<<<dim3((unsigned)gx, (unsigned)p.splits), WsCfg::NT + 32 * NPW, smem_ws, ctx->stream>>>(g);
    }
    SYSML_LAUNCH_CHECK(ctx, "dgemm_ws_kernel");
  } else {
    dgemm_kernel<TA, TB, BM, BN, WARPS_M, WARPS_N><<<dim3((unsigned)p.ntiles, (unsigned)p.splits), Cfg::NT, Cfg::SMEM, ctx->stream>>>(g);
    SYSML_LAUNCH_CHECK(ctx, "dgemm_kernel");
  }
  if (direct) return SYSML_OK;
  const i64 len = M * N;
  i64 blocks = (len + 255) / 256;
  const i64 cap = (i64)ctx->sm_count * 8;
  if (blocks > cap) blocks = cap;
  if (sym) {
    SYSML_CUDA(ctx, cudaMemsetAsync(d_nnz, 0, sizeof(unsigned long long), ctx->stream));
    tsmm_finish_kernel<<<(unsigned)blocks, 256, 0, ctx->stream>>>(partial, p.splits, N, C, d_nnz);
    SYSML_LAUNCH_CHECK(ctx, "tsmm_finish_kernel");
    if (h_nnz) {
      unsigned long long h = 0;
      SYSML_CUDA(ctx, cudaMemcpyAsync(&h, d_nnz, sizeof(h), cudaMemcpyDeviceToHost, ctx->stream));
      SYSML_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
      *h_nnz = (int64_t)h;
    }
  } else {
    reduce_partials_kernel<<<(unsigned)blocks, 256, 0, ctx->stream>>>(partial, p.splits, len, C);
    SYSML_LAUNCH_CHECK(ctx, "reduce_partials_kernel");
  }
  return SYSML_OK;
}

// picks the tile shape from N: square 128x128 for wide outputs, 256 x (8..64) for skinny ones
template <bool TA, bool TB>
int dispatch_gemm(sysml_b200_ctx* ctx, const double* A, i64 lda, const double* B, i64 ldb, double* C, i64 M, i64 N, i64 K, bool sym,
                  int64_t* h_nnz) {
  if (!sym && N > 64 && M <= 64) return run_gemm<TA, TB, 64, 128, 2, 4, BK, STAGES, TA ? 1 : 4, true>(ctx, A, lda, B, ldb, C, M, N, K, false, nullptr);
  if (sym) {
    // tsmm is t(X)%*%X (TA, !TB) or X%*%t(X) (!TA, TB): only those two get the strip-mapped kernel (its per-warp,
    // per-fragment-count instantiations are the bulk of this file's compile time)
    if constexpr (TA != TB) return run_gemm<TA, TB, 128, 128, 2, 4, BK, STAGES, 1, true>(ctx, A, lda, B, ldb, C, M, N, K, true, h_nnz);
    else return run_gemm<TA, TB, 128, 128, 2, 4>(ctx, A, lda, B, ldb, C, M, N, K, true, h_nnz);
  }
  if (N > 64) return run_gemm<TA, TB, 128, 128, 2, 4>(ctx, A, lda, B, ldb, C, M, N, K, false, h_nnz);
  if (M <= 128 && N > 32) return run_gemm<TA, TB, 128, 64, 4, 2>(ctx, A, lda, B, ldb, C, M, N, K, false, nullptr);
  // K-contiguous A (no bulk copies possible): one producer warp cannot keep enough cp.async in flight
  // (measured ~28 KB per SM), so four producer warps share the rows of every stage
  constexpr int NP = TA ? 1 : 4;
  // t(X) %*% P with a narrow P: three stages instead of four let two CTAs share an SM (16 consumer warps keep the
  // tensor pipe busier while both it and HBM are near their limits)
  constexpr int SKST = TA ? 3 : STAGES;
  if (N <= 8) return run_gemm<TA, TB, 256, 8, 8, 1, BK, SKST, NP>(ctx, A, lda, B, ldb, C, M, N, K, false, nullptr);
  if (N <= 16) return run_gemm<TA, TB, 256, 16, 8, 1, BK, SKST, NP>(ctx, A, lda, B, ldb, C, M, N, K, false, nullptr);
  if (N <= 24) return run_gemm<TA, TB, 256, 24, 8, 1, BK, SKST, NP>(ctx, A, lda, B, ldb, C, M, N, K, false, nullptr);
  if (N <= 32) return run_gemm<TA, TB, 256, 32, 8, 1, BK, STAGES, NP>(ctx, A, lda, B, ldb, C, M, N, K, false, nullptr);
  if (N <= 40) return run_gemm<TA, TB, 256, 40, 8, 1, BK, STAGES, NP>(ctx, A, lda, B, ldb, C, M, N, K, false, nullptr);
  if (N <= 48) return run_gemm<TA, TB, 256, 48, 8, 1, BK, STAGES, NP>(ctx, A, lda, B, ldb, C, M, N, K, false, nullptr);
  if (N <= 56) return run_gemm<TA, TB, 256, 56, 8, 1, BK, STAGES, NP>(ctx, A, lda, B, ldb, C, M, N, K, false, nullptr);
  return run_gemm<TA, TB, 256, 64, 8, 1, BK, STAGES, NP>(ctx, A, lda, B, ldb, C, M, N, K, false, nullptr);
}

// ------------------------------------------------------------------------------------------------
// Odd shapes.  The warp-specialised kernel moves 16-byte pieces: it wants even leading dimensions, even M and N and
// 16-byte aligned bases.  An operand that is not like that (t(X) %*% X with 567 or 999 columns, an odd number of rows
// of a skinny product) is copied once into a zero-padded, even-sized image in the context's scratch block -- zeros add
// nothing to any sum -- the product runs on the images at full speed, and the m x n corner of the padded result is
// copied out.  The copies cost one pass over the operands (2 x 8 bytes per cell) next to a contraction that reads each
// cell many times.  Only when the images do not fit in half of the free device memory does the call fall back to the
// 8-byte cp.async kernel.
// ------------------------------------------------------------------------------------------------
__global__ void pad_copy_kernel(const double* __restrict__ src, i64 rs, i64 cs, double* __restrict__ dst, i64 rd, i64 cd) {
  const i64 total = rd * cd;
  const i64 stride = (i64)gridDim.x * blockDim.x;
  for (i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    const i64 r = i / cd, c = i - r * cd;
    dst[i] = (r < rs && c < cs) ? ldg_stream(src + r * cs + c) : 0.0;
  }
}
inline bool ws_friendly(const double* p, i64 ld) { return ((reinterpret_cast<uintptr_t>(p) & 15) == 0) && (ld % 2 == 0); }
inline i64 even_up(i64 x) { return x + (x & 1); }

template <bool TA, bool TB>
int gemm_any_shape(sysml_b200_ctx* ctx, const double* A, i64 a_rows, i64 a_cols, const double* B, i64 b_rows, i64 b_cols, double* C, i64 M, i64 N,
                   i64 K, bool sym, int64_t* h_nnz) {
  const bool ok = ws_friendly(A, a_cols) && ws_friendly(B, b_cols) && (M % 2 == 0) && (N % 2 == 0) && ((reinterpret_cast<uintptr_t>(C) & 15) == 0);
  if (ok || getenv("SYSML_B200_GEMM_NO_PAD")) return dispatch_gemm<TA, TB>(ctx, A, a_cols, B, b_cols, C, M, N, K, sym, h_nnz);
  // K is padded only where it is a leading dimension (the columns of a stored operand); an odd number of contraction
  // ROWS needs nothing, so t(X) %*% X of an odd-row X is not copied
  const bool kpad = (K & 1) && (!TA || TB);
  const i64 Mp = even_up(M), Np = even_up(N), Kp = kpad ? K + 1 : K;
  const i64 ar = TA ? Kp : Mp, ac = TA ? Mp : Kp;   // image of A
  const i64 br = TB ? Np : Kp, bc = TB ? Kp : Np;   // image of B
  const bool padA = (reinterpret_cast<uintptr_t>(A) & 15) != 0 || ar != a_rows || ac != a_cols;
  const bool padB = !sym && ((reinterpret_cast<uintptr_t>(B) & 15) != 0 || br != b_rows || bc != b_cols);
  const bool padC = Mp != M || Np != N || (reinterpret_cast<uintptr_t>(C) & 15) != 0;
  auto al = [](size_t b) { return (b + 255) & ~(size_t)255; };
  const size_t bytesA = padA ? al((size_t)ar * ac * 8) : 0, bytesB = padB ? al((size_t)br * bc * 8) : 0, bytesC = padC ? al((size_t)Mp * Np * 8) : 0;
  const size_t prefix = bytesA + bytesB + bytesC;
  const size_t nested = ((size_t)1 << 30) + ((size_t)1 << 20);  // run_gemm asks for at most 1 GiB + 256 bytes of split-K partials
  size_t free_b = 0, total_b = 0;
  cudaMemGetInfo(&free_b, &total_b);
  if (ctx->ws_prefix != 0 || (prefix + nested > ctx->ws_bytes && prefix + nested > (free_b + (ctx->ws_external ? 0 : ctx->ws_bytes)) / 2))
    return dispatch_gemm<TA, TB>(ctx, A, a_cols, B, b_cols, C, M, N, K, sym, h_nnz);  // no room for the images: the 8-byte kernel
  void* ws = nullptr;
  int rc = sysml_workspace(ctx, prefix + nested, &ws);
  if (rc) return rc;
  char* base = (char*)ws;
  const double* Ap = A;
  const double* Bp = B;
  double* Cp = C;
  auto grid = [&](i64 n) {
    i64 b = (n + 255) / 256;
    const i64 cap = (i64)ctx->sm_count * 16;
    return (unsigned)(b > cap ? cap : (b < 1 ? 1 : b));
  };
  if (padA) {
    pad_copy_kernel<<<grid(ar * ac), 256, 0, ctx->stream>>>(A, a_rows, a_cols, (double*)base, ar, ac);
    SYSML_LAUNCH_CHECK(ctx, "pad_copy_kernel");
    Ap = (const double*)base;
  }
  if (sym) Bp = Ap;
  if (padB) {
    pad_copy_kernel<<<grid(br * bc), 256, 0, ctx->stream>>>(B, b_rows, b_cols, (double*)(base + bytesA), br, bc);
    SYSML_LAUNCH_CHECK(ctx, "pad_copy_kernel");
    Bp = (const double*)(base + bytesA);
  }
  if (padC) Cp = (double*)(base + bytesA + bytesB);
  ctx->ws_prefix = prefix;
  rc = dispatch_gemm<TA, TB>(ctx, Ap, ac, Bp, sym ? ac : bc, Cp, Mp, Np, Kp, sym, h_nnz);
  ctx->ws_prefix = 0;
  if (rc) return rc;
  if (padC) SYSML_CUDA(ctx, cudaMemcpy2DAsync(C, (size_t)N * 8, Cp, (size_t)Np * 8, (size_t)N * 8, (size_t)M, cudaMemcpyDeviceToDevice, ctx->stream));
  return SYSML_OK;
}

}  // namespace

extern "C" int sysml_b200_tsmm_dense(sysml_b200_ctx* ctx, const double* X, double* out, int64_t rows, int64_t cols, int left,
                                     int64_t* h_nnz_out) {
  SYSML_ENTER(ctx);
  if (rows < 0 || cols < 0) return sysml_set_error(ctx, SYSML_E_DIM, "tsmm: negative dimension");
  if (h_nnz_out) *h_nnz_out = 0;
  const i64 dim = left ? cols : rows;
  const i64 K = left ? rows : cols;
  if (dim == 0) return SYSML_OK;
  if (!out) return sysml_set_error(ctx, SYSML_E_INVALID, "tsmm: out is null");
  if (K == 0) {
    SYSML_CUDA(ctx, cudaMemsetAsync(out, 0, (size_t)(dim * dim) * sizeof(double), ctx->stream));
    return SYSML_OK;
  }
  if (!X) return sysml_set_error(ctx, SYSML_E_INVALID, "tsmm: X is null");
  if (left) return gemm_any_shape<true, false>(ctx, X, rows, cols, X, rows, cols, out, dim, dim, K, true, h_nnz_out);
  return gemm_any_shape<false, true>(ctx, X, rows, cols, X, rows, cols, out, dim, dim, K, true, h_nnz_out);
}

extern "C" int sysml_b200_matmult_dense(sysml_b200_ctx* ctx, const double* A, const double* B, double* C, int64_t a_rows,
                                        int64_t a_cols, int64_t b_rows, int64_t b_cols, int trans_a, int trans_b) {
  SYSML_ENTER(ctx);
  if (a_rows < 0 || a_cols < 0 || b_rows < 0 || b_cols < 0) return sysml_set_error(ctx, SYSML_E_DIM, "matmult: negative dimension");
  const i64 M = trans_a ? a_cols : a_rows, K = trans_a ? a_rows : a_cols;
  const i64 K2 = trans_b ? b_cols : b_rows, N = trans_b ? b_rows : b_cols;
  if (K != K2) return sysml_set_error(ctx, SYSML_E_DIM, "matmult: inner dimensions %lld vs %lld", (long long)K, (long long)K2);
  if (M * N == 0) return SYSML_OK;
  if (!C) return sysml_set_error(ctx, SYSML_E_INVALID, "matmult: C is null");
  if (K == 0) {
    SYSML_CUDA(ctx, cudaMemsetAsync(C, 0, (size_t)(M * N) * sizeof(double), ctx->stream));
    return SYSML_OK;
  }
  if (!A || !B) return sysml_set_error(ctx, SYSML_E_INVALID, "matmult: null operand");
  // vector shapes -> the HBM-bound row-streaming kernels of mmchain.cu (cublasDgemv / Ddot twins)
  if (N == 1) return trans_a ? sysml_dense_vm(ctx, A, B, C, K, M) : sysml_dense_mv(ctx, A, B, C, M, K);
  if (M == 1) return trans_b ? sysml_dense_mv(ctx, B, A, C, N, K) : sysml_dense_vm(ctx, B, A, C, K, N);
  if (!trans_a && !trans_b) return gemm_any_shape<false, false>(ctx, A, a_rows, a_cols, B, b_rows, b_cols, C, M, N, K, false, nullptr);
  if (trans_a && !trans_b) return gemm_any_shape<true, false>(ctx, A, a_rows, a_cols, B, b_rows, b_cols, C, M, N, K, false, nullptr);
  if (!trans_a && trans_b) return gemm_any_shape<false, true>(ctx, A, a_rows, a_cols, B, b_rows, b_cols, C, M, N, K, false, nullptr);
  return gemm_any_shape<true, true>(ctx, A, a_rows, a_cols, B, b_rows, b_cols, C, M, N, K, false, nullptr);
}
