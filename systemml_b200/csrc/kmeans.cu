// kmeans.cu -- one Lloyd step of Kmeans.dml in ONE pass over X.
//
// The script's inner loop (scripts/algorithms/Kmeans.dml:146-166):
//     D     = -2 * (X %*% t(C)) + t(rowSums(C^2))      n x k
//     minD  = rowMins(D);  wcss = sumXsq + sum(minD)
//     P     = D <= minD;   P = P / rowSums(P)           ties share a record equally
//     P_denom = colSums(P);  C_new = (t(P) %*% X) / t(P_denom)
// is, in the reference, ba+* / r' / uarsqk+ / * / + / uarmin / uak+ / <= / uark+ / / / uack+ / ba+* / / : a dozen
// instructions, of which the two products read X (8 GB at C5) and the others read or write n x k matrices
// (LibMatrixMult.matrixMultDenseDense LibMatrixMult.java:880-1103, LibMatrixAgg.java:1624-1927,
// LibMatrixBincell.java:87-339; on the reference GPU: cublasDgemm twice plus seven kernels).
//
// Here a CTA streams 64-row tiles of X through shared memory (bulk copies, two stages).  Per tile:
//   phase 1  the tile times t(C) on the FP64 tensor cores (mma.sync m8n8k4 / SASS DMMA, C resident in shared memory,
//            one warp per 8 rows, all k centroids), -2*xc + c2 with the script's two roundings, the row minimum, the set
//            of centroids that attain it (a bit mask per row) and its size;
//   phase 2  every row is added -- times 1/|set|, which is exactly 1 when there is no tie -- to the rows of a k x m
//            accumulator in shared memory that belong to its centroids.  Warp w owns centroids c = w (mod 8), so no two
//            warps ever touch the same accumulator row: no atomics, and a fixed order of additions inside a CTA.
// D, minD and P never exist in memory.  Per-CTA results (k x m sums, k counts, sum of minD) are combined by a
// fixed-order compensated reduction.  Work: 2*n*m*k flops on the tensor pipe (the bound at m = 100, k = 50: 12.5 flop
// per byte), X read once: 8*n*m bytes.
//
// NaN in X or C: the script propagates NaN through rowMins into every centroid; this kernel leaves such records out of
// the sums (their mask is empty).  Callers that must reproduce that use the operator-by-operator path.
#include <stdlib.h>

#include "common.cuh"

namespace {

constexpr int KM_ROWS = 64;    // rows per tile = 8 warps x 8 rows
constexpr int KM_THREADS = 256;

__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

struct KmArgs {
  const double* X;
  const double* C;
  const double* c2;
  double* partial;  // [gridDim.x][k*m + k + 1]
  i64 n, tiles;
  int m, k, ld;     // ld: leading dimension of the shared-memory tiles, = 4 (mod 16) doubles -> conflict-free fragment loads
};

// NT: 8-centroid tiles (k <= 8*NT)
template <int NT>
__global__ void __launch_bounds__(KM_THREADS) kmeans_step_kernel(const KmArgs a) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const int m = a.m, k = a.k, ld = a.ld;
  const int m4 = (m + 3) & ~3;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem_raw);                 // [2]
  unsigned long long* mask_s = reinterpret_cast<unsigned long long*>(smem_raw + 16);   // [64]
  double* inv_s = reinterpret_cast<double*>(smem_raw + 16 + 64 * 8);     // [64]  1 / (number of closest centroids)
  double* c2_s = inv_s + 64;                                             // [NT*8]
  double* den_s = c2_s + NT * 8;                                         // [NT*8]
  double* red_s = den_s + NT * 8;                                        // [2*KM_THREADS] final wcss reduction
  double* Cs = red_s + 2 * KM_THREADS;                                   // [NT*8][ld]
  double* acc_s = Cs + (size_t)NT * 8 * ld;                              // [k][m]
  double* Xs = acc_s + (size_t)k * m;                                    // [2][KM_ROWS][ld]
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

  if (tid == 0) {
    mbar_init(&full[0], 1);
    mbar_init(&full[1], 1);
    mbar_fence_init();
  }
  // centroids (zero-padded to NT*8 rows and ld columns), c2 (+inf for the padding: never the minimum), accumulators
  for (int i = tid; i < NT * 8 * ld; i += KM_THREADS) {
    const int c = i / ld, j = i - c * ld;
    Cs[i] = (c < k && j < m) ? a.C[(size_t)c * m + j] : 0.0;
  }
  for (int i = tid; i < NT * 8; i += KM_THREADS) {
    c2_s[i] = (i < k) ? a.c2[i] : __longlong_as_double(0x7ff0000000000000LL);
    den_s[i] = 0.0;
  }
  for (int i = tid; i < k * m; i += KM_THREADS) acc_s[i] = 0.0;
  for (int i = tid; i < 2 * KM_ROWS * ld; i += KM_THREADS) Xs[i] = 0.0;  // padding columns stay zero for good
  __syncthreads();

  const uint64_t pol = l2_policy_evict_first();
  auto issue = [&](i64 tile, int s) {  // thread 0 only
    const i64 r0 = tile * KM_ROWS;
    const int rows = (int)((a.n - r0 < KM_ROWS) ? (a.n - r0) : KM_ROWS);
    double* dst = Xs + (size_t)s * KM_ROWS * ld;
    mbar_expect_tx(&full[s], (uint32_t)((size_t)rows * m * sizeof(double)));
    if (ld == m) {
      bulk_g2s(dst, a.X + r0 * m, (uint32_t)((size_t)rows * m * sizeof(double)), &full[s], pol);
    } else {
      for (int r = 0; r < rows; r++) bulk_g2s(dst + (size_t)r * ld, a.X + (r0 + r) * m, (uint32_t)(m * sizeof(double)), &full[s], pol);
    }
  };
  if (tid == 0) {
    // generic-proxy writes (the zero fill above) before async-proxy writes to the same buffers
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    if ((i64)blockIdx.x < a.tiles) issue(blockIdx.x, 0);
    if ((i64)blockIdx.x + gridDim.x < a.tiles) issue((i64)blockIdx.x + gridDim.x, 1);
  }

  CompSum wsum;  // sum of minD over the rows this thread reports (lanes 0, 4, 8, ... of every warp)
  int it = 0;
  for (i64 tile = blockIdx.x; tile < a.tiles; tile += gridDim.x, ++it) {
    const int s = it & 1;
    const uint32_t ph = (uint32_t)(it >> 1) & 1u;
    const i64 r0 = tile * KM_ROWS;
    const int rows = (int)((a.n - r0 < KM_ROWS) ? (a.n - r0) : KM_ROWS);
    mbar_wait(&full[s], ph);
    const double* Xt = Xs + (size_t)s * KM_ROWS * ld;

    // ---------------- phase 1: distances of this warp's 8 rows to all centroids ----------------
    {
      const double* A = Xt + (size_t)(warp * 8 + (lane >> 2)) * ld + (lane & 3);
      const double* B = Cs + (size_t)(lane >> 2) * ld + (lane & 3);
      double acc[NT][2];
#pragma unroll
      for (int t = 0; t < NT; t++) acc[t][0] = acc[t][1] = 0.0;
#pragma unroll 5
      for (int k0 = 0; k0 < m4; k0 += 4) {
        const double av = A[k0];
#pragma unroll
        for (int t = 0; t < NT; t++) dmma884(acc[t][0], acc[t][1], av, B[(size_t)t * 8 * ld + k0]);
      }
      // D = -2 * xc + c2: two roundings, as the script's two instructions
      const int cb = 2 * (lane & 3);
      double mn = __longlong_as_double(0x7ff0000000000000LL);
#pragma unroll
      for (int t = 0; t < NT; t++) {
        acc[t][0] = __dadd_rn(__dmul_rn(-2.0, acc[t][0]), c2_s[t * 8 + cb]);
        acc[t][1] = __dadd_rn(__dmul_rn(-2.0, acc[t][1]), c2_s[t * 8 + cb + 1]);
        mn = fmin(mn, fmin(acc[t][0], acc[t][1]));
      }
      mn = fmin(mn, __shfl_xor_sync(0xffffffffu, mn, 1));
      mn = fmin(mn, __shfl_xor_sync(0xffffffffu, mn, 2));
      unsigned long long msk = 0ull;
#pragma unroll
      for (int t = 0; t < NT; t++) {
        if (acc[t][0] <= mn) msk |= 1ull << (t * 8 + cb);
        if (acc[t][1] <= mn) msk |= 1ull << (t * 8 + cb + 1);
      }
      msk |= __shfl_xor_sync(0xffffffffu, msk, 1);
      msk |= __shfl_xor_sync(0xffffffffu, msk, 2);
      const int row = warp * 8 + (lane >> 2);
      if ((lane & 3) == 0) {
        const bool valid = row < rows && mn == mn;  // rows past the end of X, and NaN rows, contribute nothing
        const int cnt = __popcll(msk);
        mask_s[row] = (valid && cnt > 0) ? msk : 0ull;
        inv_s[row] = (cnt > 0) ? __ddiv_rn(1.0, (double)cnt) : 0.0;   // P / rowSums(P)
        if (valid && cnt > 0) wsum.add(mn);
      }
    }
    __syncthreads();

    // ---------------- phase 2: rows into the accumulator rows of their centroids (warp w owns c = w mod 8) ----------------
    {
      unsigned long long own = 0ull;
#pragma unroll
      for (int t = 0; t < NT; t++) own |= 1ull << (t * 8 + warp);
      // which of the 64 rows concern this warp: two rows per lane, two ballots; then the rows are visited in row order,
      // so the order of the additions into an accumulator row is fixed
      const unsigned lo_rows = __ballot_sync(0xffffffffu, (mask_s[lane] & own) != 0ull);
      const unsigned hi_rows = __ballot_sync(0xffffffffu, (mask_s[32 + lane] & own) != 0ull);
      unsigned long long todo = ((unsigned long long)hi_rows << 32) | lo_rows;
      while (todo) {
        const int r = __ffsll((long long)todo) - 1;
        todo &= todo - 1;
        unsigned long long mm = mask_s[r] & own;   // warp-uniform; more than one bit only when this warp's centroids tie
        const double wgt = inv_s[r];
        const double* xr = Xt + (size_t)r * ld;
        // the row's contribution P[i,c] * X[i,j], four columns per lane
        double px[4];
#pragma unroll
        for (int q = 0; q < 4; q++) {
          const int j = lane + 32 * q;
          px[q] = (j < m) ? __dmul_rn(wgt, xr[j]) : 0.0;
        }
        while (mm) {
          const int c = __ffsll((long long)mm) - 1;
          mm &= mm - 1;
          double* ar = acc_s + (size_t)c * m;
          if (m <= 128) {
            double av[4];
#pragma unroll
            for (int q = 0; q < 4; q++) av[q] = (lane + 32 * q < m) ? ar[lane + 32 * q] : 0.0;
#pragma unroll
            for (int q = 0; q < 4; q++)
              if (lane + 32 * q < m) ar[lane + 32 * q] = __dadd_rn(av[q], px[q]);
          } else {
            for (int j = lane; j < m; j += 32) ar[j] = __dadd_rn(ar[j], __dmul_rn(wgt, xr[j]));
          }
          if (lane == 0) den_s[c] = __dadd_rn(den_s[c], wgt);
        }
      }
    }
    __syncthreads();  // the stage is free again, the masks may be overwritten
    if (tid == 0) {
      const i64 nxt = tile + 2 * (i64)gridDim.x;
      if (nxt < a.tiles) issue(nxt, s);
    }
  }

  // ---------------- per-CTA results -> scratch; the finishing kernel adds the CTAs in a fixed order ----------------
  const size_t L = (size_t)k * m + k + 1;
  double* prow = a.partial + (size_t)blockIdx.x * L;
  for (int i = tid; i < k * m; i += KM_THREADS) prow[i] = acc_s[i];
  for (int i = tid; i < k; i += KM_THREADS) prow[(size_t)k * m + i] = den_s[i];
  red_s[tid] = wsum.hi;
  red_s[KM_THREADS + tid] = wsum.lo;
  __syncthreads();
  if (tid == 0) {
    CompSum t;
    for (int q = 0; q < KM_THREADS; q++) {
      CompSum u;
      u.hi = red_s[q];
      u.lo = red_s[KM_THREADS + q];
      t.add(u);
    }
    prow[(size_t)k * m + k] = t.value();
  }
}

// ------------------------------------------------------------------------------------------------------------------
// Warp-specialised form for m <= 128 (m a multiple of 4): the tensor pipe never waits for the accumulation.
//   warps 0..7       distances: 8 rows each on the tensor cores, then the row minimum, the tie mask and 1/|mask| per row
//                    into the stage's assignment slots; one arrival per warp on the stage's `asg` barrier
//   warps 8..15      accumulation: warp 8+u owns centroids c = u (mod 8) and keeps THEIR sums in registers (NT rows x 4
//                    columns per lane, lane l holding columns 4l..4l+3): it waits for the assignments of the stage, picks
//                    the rows of one owned centroid at a time with two ballots, adds them in row order, and releases the
//                    stage
//   warp 16          producer: streams 64-row tiles through a 3-stage ring (bulk copies, full / empty mbarriers)
// (Tried and dropped: two distance groups on alternate tiles with 16 rows per warp -- 7.3 ms instead of 4.8 ms at C5; one
// warp per SM sub-partition does not keep the tensor pipe fed across its fragment loads.  A head start of warps 0..3
// over 4..7, to stagger their epilogues, changes nothing: 4.76 ms for 0 ... 6 us.)
// No shared-memory accumulator, no __syncthreads in the loop; each centroid's sum is built by one warp in a fixed order.
// FP64 adds share their pipe with the tensor-core instructions, so the accumulation issues nothing it does not need.
// ------------------------------------------------------------------------------------------------------------------
constexpr int KW_STAGES = 3;
constexpr int KW_THREADS = 17 * 32;

template <int NT>
__global__ void __launch_bounds__(KW_THREADS, 1) kmeans_step_ws_kernel(const KmArgs a) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const int m = a.m, k = a.k, ld = a.ld;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem_raw);                 // [S]  tile landed
  uint64_t* asg = full + KW_STAGES;                                      // [S]  assignments of the tile written (8 arrivals)
  uint64_t* empty = asg + KW_STAGES;                                     // [S]  tile consumed (8 arrivals)
  unsigned long long* mask_s = reinterpret_cast<unsigned long long*>(smem_raw + 128);   // [S][64]
  double* inv_s = reinterpret_cast<double*>(smem_raw + 128 + KW_STAGES * 64 * 8);       // [S][64]
  double* c2_s = inv_s + KW_STAGES * 64;                                 // [NT*8]
  double* red_s = c2_s + NT * 8;                                         // [2*256] final sum(minD) reduction
  double* Cs = red_s + 512;                                              // [NT*8][ld]
  double* Xs = Cs + (size_t)NT * 8 * ld;                                 // [S][64][ld]
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

  if (tid == 0) {
    for (int s = 0; s < KW_STAGES; s++) {
      mbar_init(&full[s], 1);
      mbar_init(&asg[s], 8);
      mbar_init(&empty[s], 8);
    }
    mbar_fence_init();
  }
  for (int i = tid; i < NT * 8 * ld; i += KW_THREADS) {
    const int c = i / ld, j = i - c * ld;
    Cs[i] = (c < k && j < m) ? a.C[(size_t)c * m + j] : 0.0;
  }
  for (int i = tid; i < NT * 8; i += KW_THREADS) c2_s[i] = (i < k) ? a.c2[i] : __longlong_as_double(0x7ff0000000000000LL);
  for (int i = tid; i < KW_STAGES * KM_ROWS * ld; i += KW_THREADS) Xs[i] = 0.0;
  __syncthreads();

  const size_t L = (size_t)k * m + k + 1;
  double* prow = a.partial + (size_t)blockIdx.x * L;

  if (warp < 8) {
    // ===================== distance warps =====================
    const int m4 = (m + 3) & ~3;
    CompSum wsum;
    int it = 0;
    for (i64 tile = blockIdx.x; tile < a.tiles; tile += gridDim.x, ++it) {
      const int s = it % KW_STAGES;
      const uint32_t ph = (uint32_t)(it / KW_STAGES) & 1u;
      const i64 r0 = tile * KM_ROWS;
      const int rows = (int)((a.n - r0 < KM_ROWS) ? (a.n - r0) : KM_ROWS);
      mbar_wait(&full[s], ph);
      const double* Xt = Xs + (size_t)s * KM_ROWS * ld;
      const double* A = Xt + (size_t)(warp * 8 + (lane >> 2)) * ld + (lane & 3);
      const double* B = Cs + (size_t)(lane >> 2) * ld + (lane & 3);
      double acc[NT][2];
#pragma unroll
      for (int t = 0; t < NT; t++) acc[t][0] = acc[t][1] = 0.0;
#pragma unroll 5
      for (int k0 = 0; k0 < m4; k0 += 4) {
        const double av = A[k0];
#pragma unroll
        for (int t = 0; t < NT; t++) dmma884(acc[t][0], acc[t][1], av, B[(size_t)t * 8 * ld + k0]);
      }
      const int cb = 2 * (lane & 3);
      double mn = __longlong_as_double(0x7ff0000000000000LL);
#pragma unroll
      for (int t = 0; t < NT; t++) {
        acc[t][0] = __dadd_rn(__dmul_rn(-2.0, acc[t][0]), c2_s[t * 8 + cb]);
        acc[t][1] = __dadd_rn(__dmul_rn(-2.0, acc[t][1]), c2_s[t * 8 + cb + 1]);
        mn = fmin(mn, fmin(acc[t][0], acc[t][1]));
      }
      mn = fmin(mn, __shfl_xor_sync(0xffffffffu, mn, 1));
      mn = fmin(mn, __shfl_xor_sync(0xffffffffu, mn, 2));
      unsigned long long msk = 0ull;
#pragma unroll
      for (int t = 0; t < NT; t++) {
        if (acc[t][0] <= mn) msk |= 1ull << (t * 8 + cb);
        if (acc[t][1] <= mn) msk |= 1ull << (t * 8 + cb + 1);
      }
      msk |= __shfl_xor_sync(0xffffffffu, msk, 1);
      msk |= __shfl_xor_sync(0xffffffffu, msk, 2);
      const int row = warp * 8 + (lane >> 2);
      if ((lane & 3) == 0) {
        const bool valid = row < rows && mn == mn;
        const int cnt = __popcll(msk);
        mask_s[s * 64 + row] = (valid && cnt > 0) ? msk : 0ull;
        inv_s[s * 64 + row] = (cnt > 0) ? __ddiv_rn(1.0, (double)cnt) : 0.0;
        if (valid && cnt > 0) wsum.add(mn);
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&asg[s]);  // release: this warp's slots of the stage are written, its reads of the tile are done
    }
    red_s[tid] = wsum.hi;
    red_s[256 + tid] = wsum.lo;
    named_bar_sync(1, 256);
    if (tid == 0) {
      CompSum t;
      for (int q = 0; q < 256; q++) {
        CompSum u;
        u.hi = red_s[q];
        u.lo = red_s[256 + q];
        t.add(u);
      }
      prow[(size_t)k * m + k] = t.value();
    }
    return;
  }

  if (warp == 16) {
    // ===================== producer =====================
    if (lane == 0) {
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // the zero fill above, then bulk copies into the same buffers
      const uint64_t pol = l2_policy_evict_first();
      int it = 0;
      for (i64 tile = blockIdx.x; tile < a.tiles; tile += gridDim.x, ++it) {
        const int s = it % KW_STAGES;
        const uint32_t ph = (uint32_t)(it / KW_STAGES) & 1u;
        mbar_wait(&empty[s], ph ^ 1u);
        const i64 r0 = tile * KM_ROWS;
        const int rows = (int)((a.n - r0 < KM_ROWS) ? (a.n - r0) : KM_ROWS);
        double* dst = Xs + (size_t)s * KM_ROWS * ld;
        mbar_expect_tx(&full[s], (uint32_t)((size_t)rows * m * sizeof(double)));
        if (ld == m) {
          bulk_g2s(dst, a.X + r0 * m, (uint32_t)((size_t)rows * m * sizeof(double)), &full[s], pol);
        } else {
          for (int r = 0; r < rows; r++) bulk_g2s(dst + (size_t)r * ld, a.X + (r0 + r) * m, (uint32_t)(m * sizeof(double)), &full[s], pol);
        }
      }
    }
    return;
  }

  // ===================== accumulation warps =====================
  {
    const int u = warp - 8;
    double sum[NT][4];
    double den[NT];
#pragma unroll
    for (int t = 0; t < NT; t++) {
      den[t] = 0.0;
#pragma unroll
      for (int q = 0; q < 4; q++) sum[t][q] = 0.0;
    }
    const bool act = 4 * lane < m;  // m is a multiple of 4: a lane holds four whole columns or none
    int it = 0;
    for (i64 tile = blockIdx.x; tile < a.tiles; tile += gridDim.x, ++it) {
      const int s = it % KW_STAGES;
      const uint32_t ph = (uint32_t)(it / KW_STAGES) & 1u;
      mbar_wait(&full[s], ph);   // the tile itself (bulk-copy writes become visible through this barrier)
      mbar_wait(&asg[s], ph);    // and what the distance warps decided about its rows
      const double* Xt = Xs + (size_t)s * KM_ROWS * ld;
      const unsigned long long* mk = mask_s + s * 64;
      const unsigned long long m0 = mk[lane], m1 = mk[32 + lane];
      // One owned centroid at a time (static index: the sums stay in registers and every addition issued is a real
      // one), its rows in row order: the order of the additions into a centroid is fixed.
#pragma unroll
      for (int t = 0; t < NT; t++) {
        const int bit = t * 8 + u;
        const unsigned lo_rows = __ballot_sync(0xffffffffu, ((m0 >> bit) & 1ull) != 0ull);
        const unsigned hi_rows = __ballot_sync(0xffffffffu, ((m1 >> bit) & 1ull) != 0ull);
        unsigned long long todo = ((unsigned long long)hi_rows << 32) | lo_rows;
        while (todo) {
          const int r = __ffsll((long long)todo) - 1;
          todo &= todo - 1;
          const double wgt = inv_s[s * 64 + r];
          if (act) {
            const double2 x01 = *reinterpret_cast<const double2*>(Xt + (size_t)r * ld + 4 * lane);
            const double2 x23 = *reinterpret_cast<const double2*>(Xt + (size_t)r * ld + 4 * lane + 2);
            if (wgt == 1.0) {  // no tie: P[i,c] = 1 and the product is the value itself
              sum[t][0] = __dadd_rn(sum[t][0], x01.x);
              sum[t][1] = __dadd_rn(sum[t][1], x01.y);
              sum[t][2] = __dadd_rn(sum[t][2], x23.x);
              sum[t][3] = __dadd_rn(sum[t][3], x23.y);
            } else {
              sum[t][0] = __dadd_rn(sum[t][0], __dmul_rn(wgt, x01.x));
              sum[t][1] = __dadd_rn(sum[t][1], __dmul_rn(wgt, x01.y));
              sum[t][2] = __dadd_rn(sum[t][2], __dmul_rn(wgt, x23.x));
              sum[t][3] = __dadd_rn(sum[t][3], __dmul_rn(wgt, x23.y));
            }
          }
          den[t] = __dadd_rn(den[t], wgt);
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty[s]);
    }
#pragma unroll
    for (int t = 0; t < NT; t++) {
      const int c = t * 8 + u;
      if (c < k) {
        if (act) {
#pragma unroll
          for (int q = 0; q < 4; q++) prow[(size_t)c * m + 4 * lane + q] = sum[t][q];
        }
        if (lane == 0) prow[(size_t)k * m + c] = den[t];
      }
    }
  }
}

__global__ void kmeans_finish_kernel(const double* __restrict__ partial, int P, i64 L, i64 km, int k, double* __restrict__ sums,
                                     double* __restrict__ counts, double* __restrict__ sum_min) {
  const i64 j = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= L) return;
  CompSum acc;
  for (int p = 0; p < P; p++) acc.add(__ldcg(partial + (size_t)p * L + j));
  const double v = acc.value();
  if (j < km) sums[j] = v;
  else if (j < km + k) counts[j - km] = v;
  else *sum_min = v;
}

template <int NT>
int launch_kmeans_ws(sysml_b200_ctx* ctx, KmArgs a, double* sums, double* counts, double* sum_min) {
  const size_t smem = 128 + (size_t)KW_STAGES * 64 * 16 + (size_t)NT * 8 * 8 + 512 * 8 + (size_t)NT * 8 * a.ld * 8 + (size_t)KW_STAGES * KM_ROWS * a.ld * 8 + 128;
  if (smem > (size_t)ctx->max_smem_optin) return SYSML_E_UNSUPPORTED;
  auto kern = kmeans_step_ws_kernel<NT>;
  SYSML_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  i64 grid = ctx->sm_count;
  if (grid > a.tiles) grid = a.tiles;
  const size_t L = (size_t)a.k * a.m + a.k + 1;
  void* ws = nullptr;
  int rc = sysml_workspace(ctx, (size_t)grid * L * sizeof(double), &ws);
  if (rc) return rc;
  a.partial = (double*)ws;
  kern<<<(unsigned)grid, KW_THREADS, smem, ctx->stream>>>(a);
  SYSML_LAUNCH_CHECK(ctx, "kmeans_step_ws_kernel");
  kmeans_finish_kernel<<<(unsigned)((L + 255) / 256), 256, 0, ctx->stream>>>(a.partial, (int)grid, (i64)L, (i64)a.k * a.m, a.k, sums, counts, sum_min);
  SYSML_LAUNCH_CHECK(ctx, "kmeans_finish_kernel");
  return SYSML_OK;
}

template <int NT>
int launch_kmeans(sysml_b200_ctx* ctx, KmArgs a, double* sums, double* counts, double* sum_min) {
  if (a.m <= 128 && (a.m & 3) == 0 && !getenv("SYSML_B200_KMEANS_SIMPLE")) {
    int rc = launch_kmeans_ws<NT>(ctx, a, sums, counts, sum_min);
    if (rc != SYSML_E_UNSUPPORTED) return rc;
  }
  const size_t smem = 16 + 64 * 8 + 64 * 8 + (size_t)NT * 8 * 8 * 2 + 2 * KM_THREADS * 8 + (size_t)NT * 8 * a.ld * 8 + (size_t)a.k * a.m * 8 +
                      (size_t)2 * KM_ROWS * a.ld * 8 + 128;
  if (smem > (size_t)ctx->max_smem_optin) return SYSML_E_UNSUPPORTED;
  auto kern = kmeans_step_kernel<NT>;
  SYSML_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  i64 grid = ctx->sm_count;
  if (grid > a.tiles) grid = a.tiles;
  const size_t L = (size_t)a.k * a.m + a.k + 1;
  void* ws = nullptr;
  int rc = sysml_workspace(ctx, (size_t)grid * L * sizeof(double), &ws);
  if (rc) return rc;
  a.partial = (double*)ws;
  kern<<<(unsigned)grid, KM_THREADS, smem, ctx->stream>>>(a);
  SYSML_LAUNCH_CHECK(ctx, "kmeans_step_kernel");
  kmeans_finish_kernel<<<(unsigned)((L + 255) / 256), 256, 0, ctx->stream>>>(a.partial, (int)grid, (i64)L, (i64)a.k * a.m, a.k, sums, counts, sum_min);
  SYSML_LAUNCH_CHECK(ctx, "kmeans_finish_kernel");
  return SYSML_OK;
}

}  // namespace

extern "C" int sysml_b200_kmeans_step(sysml_b200_ctx* ctx, const double* X, const double* C, const double* c2, int64_t n, int64_t m, int64_t k,
                                      double* sums, double* counts, double* sum_min_dist) {
  SYSML_ENTER(ctx);
  if (n < 0 || m < 1 || k < 1) return sysml_set_error(ctx, SYSML_E_DIM, "kmeans_step: bad dimensions %lld x %lld, k = %lld", (long long)n, (long long)m, (long long)k);
  if (!C || !c2 || !sums || !counts || !sum_min_dist || (n > 0 && !X)) return sysml_set_error(ctx, SYSML_E_INVALID, "kmeans_step: null pointer");
  if (k > 64) return sysml_set_error(ctx, SYSML_E_UNSUPPORTED, "kmeans_step: more than 64 centroids; run the step operator by operator");
  if ((m & 1) || (reinterpret_cast<uintptr_t>(X) & 15)) return sysml_set_error(ctx, SYSML_E_UNSUPPORTED, "kmeans_step: rows of X must be 16-byte aligned (even number of columns)");
  if (n == 0) {
    SYSML_CUDA(ctx, cudaMemsetAsync(sums, 0, (size_t)(k * m) * sizeof(double), ctx->stream));
    SYSML_CUDA(ctx, cudaMemsetAsync(counts, 0, (size_t)k * sizeof(double), ctx->stream));
    SYSML_CUDA(ctx, cudaMemsetAsync(sum_min_dist, 0, sizeof(double), ctx->stream));
    return SYSML_OK;
  }
  KmArgs a{};
  a.X = X; a.C = C; a.c2 = c2; a.n = n; a.m = (int)m; a.k = (int)k;
  a.tiles = (n + KM_ROWS - 1) / KM_ROWS;
  int ld = (int)((m + 3) & ~3LL);
  while ((ld & 15) != 4) ld += 4;  // = 4 (mod 16): the 8 rows x 4 columns of a fragment fall into 16 distinct 8-byte banks per half warp
  a.ld = ld;
  int rc;
  const int nt = (int)((k + 7) / 8);
  switch (nt) {
    case 1: rc = launch_kmeans<1>(ctx, a, sums, counts, sum_min_dist); break;
    case 2: rc = launch_kmeans<2>(ctx, a, sums, counts, sum_min_dist); break;
    case 3: rc = launch_kmeans<3>(ctx, a, sums, counts, sum_min_dist); break;
    case 4: rc = launch_kmeans<4>(ctx, a, sums, counts, sum_min_dist); break;
    case 5: rc = launch_kmeans<5>(ctx, a, sums, counts, sum_min_dist); break;
    case 6: rc = launch_kmeans<6>(ctx, a, sums, counts, sum_min_dist); break;
    case 7: rc = launch_kmeans<7>(ctx, a, sums, counts, sum_min_dist); break;
    default: rc = launch_kmeans<8>(ctx, a, sums, counts, sum_min_dist); break;
  }
  if (rc == SYSML_E_UNSUPPORTED) return sysml_set_error(ctx, SYSML_E_UNSUPPORTED, "kmeans_step: %lld columns x %lld centroids do not fit in shared memory; run the step operator by operator", (long long)m, (long long)k);
  return rc;
}
