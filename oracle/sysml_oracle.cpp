// sysml_oracle.cpp -- CPU restatement of SystemML's LibMatrixMult / LibMatrixAgg / LibMatrixBincell
// hot path.  TEST INFRASTRUCTURE ONLY: this file is the parity checker and the timed CPU baseline.
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load
// it.  The product (systemml_b200/) never links, imports or calls anything in this directory.
//
// Parity status: PINNED BY SOURCE.  The reference is Java 8 (no JVM, javac, mvn or R in this image,
// SURVEY.md section 8c) and holds no golden vectors for numeric ops (its tests regenerate inputs with
// java.util.Random and compare against R at run time).  Every function below cites the reference
// file:line whose arithmetic ORDER it follows (Java: IEEE double, no FMA contraction, left-to-right
// evaluation) and is cross-checked in tests/ against an independent extended-precision computation
// at the reference tests' own shapes, seeds and tolerances.
//
// Build: g++ -O2 -ffp-contract=off -fno-fast-math -fopenmp -shared -fPIC (see oracle/Makefile).
//
// Path shorthand: LMM = src/main/java/org/apache/sysml/runtime/matrix/data/LibMatrixMult.java
//                 LMA = .../matrix/data/LibMatrixAgg.java
//                 LMB = .../matrix/data/LibMatrixBincell.java
//                 FO  = src/main/java/org/apache/sysml/runtime/functionobjects/
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

using i64 = int64_t;

// ---------------------------------------------------------------------------------------------
// java.util.Random (JDK definition; used by TestUtils.generateTestMatrix,
// src/test/java/org/apache/sysml/test/utils/TestUtils.java:1285-1297)
// ---------------------------------------------------------------------------------------------
struct JavaRandom {
  uint64_t s;
  explicit JavaRandom(i64 seed) { s = ((uint64_t)seed ^ 0x5DEECE66DULL) & ((1ULL << 48) - 1); }
  int32_t next(int bits) {
    s = (s * 0x5DEECE66DULL + 0xBULL) & ((1ULL << 48) - 1);
    return (int32_t)((i64)s >> (48 - bits));
  }
  double nextDouble() {
    i64 hi = (i64)next(26) << 27;
    i64 lo = next(27);
    return (double)(hi + lo) * 0x1.0p-53;
  }
};

// UtilFunctions.getBalancedBlockSizesDefault (src/main/java/org/apache/sysml/runtime/util/UtilFunctions.java:156-180)
inline int round_to_next(int val, int factor) {  // UtilFunctions.roundToNext
  int pval = std::max(val, factor);
  return ((pval + factor) / factor) * factor - ((pval % factor == 0) ? factor : 0);
}
std::vector<int> balanced_blocks(int len, int k, bool constK) {
  int nk = constK ? k : round_to_next(std::min(8 * k, len / 32), k);
  std::vector<int> ret;
  int base = len / nk, rest = len % nk;
  for (int i = 0; i < nk; i++) {
    int v = base + (i < rest ? 1 : 0);
    if (v > 0) ret.push_back(v);
  }
  return ret;
}

// ---------------------------------------------------------------------------------------------
// Vector primitives, LMM:2996-3482.  All follow the "len%8 rest first, then blocks of 8" order.
// The 8-block of dotProduct is ONE expression: val += p0+p1+...+p7 (left to right), LMM:3039-3046.
// ---------------------------------------------------------------------------------------------
inline double dot_dense(const double* a, const double* b, i64 len) {  // LMM:3025-3051
  double val = 0;
  const i64 bn = len % 8;
  for (i64 i = 0; i < bn; i++) val += a[i] * b[i];
  for (i64 i = bn; i < len; i += 8) {
    val += a[i + 0] * b[i + 0] + a[i + 1] * b[i + 1] + a[i + 2] * b[i + 2] + a[i + 3] * b[i + 3] +
           a[i + 4] * b[i + 4] + a[i + 5] * b[i + 5] + a[i + 6] * b[i + 6] + a[i + 7] * b[i + 7];
  }
  return val;
}
// sparse-dense dot: a[ai..ai+len) values, aix their column ids, b dense.  LMM:3054-3081
inline double dot_sparse(const double* a, const double* b, const int* aix, i64 ai, i64 len) {
  double val = 0;
  const i64 bn = len % 8;
  for (i64 i = ai; i < ai + bn; i++) val += a[i] * b[aix[i]];
  for (i64 i = ai + bn; i < ai + len; i += 8) {
    val += a[i + 0] * b[aix[i + 0]] + a[i + 1] * b[aix[i + 1]] + a[i + 2] * b[aix[i + 2]] +
           a[i + 3] * b[aix[i + 3]] + a[i + 4] * b[aix[i + 4]] + a[i + 5] * b[aix[i + 5]] +
           a[i + 6] * b[aix[i + 6]] + a[i + 7] * b[aix[i + 7]];
  }
  return val;
}
// c += aval*b.  LMM:3084-3108 (element-wise, so the 8-unrolling does not change any rounding)
inline void vmadd1(double av, const double* b, double* c, i64 len) {
  for (i64 j = 0; j < len; j++) c[j] += av * b[j];
}
inline void vmadd2(double a1, double a2, const double* b1, const double* b2, double* c, i64 len) {  // LMM:3110-3133
  for (i64 j = 0; j < len; j++) c[j] += a1 * b1[j] + a2 * b2[j];
}
inline void vmadd3(double a1, double a2, double a3, const double* b1, const double* b2, const double* b3,
                   double* c, i64 len) {  // LMM:3135-3158
  for (i64 j = 0; j < len; j++) c[j] += a1 * b1[j] + a2 * b2[j] + a3 * b3[j];
}
inline void vmadd4(double a1, double a2, double a3, double a4, const double* b1, const double* b2,
                   const double* b3, const double* b4, double* c, i64 len) {  // LMM:3160-3183
  for (i64 j = 0; j < len; j++) c[j] += a1 * b1[j] + a2 * b2[j] + a3 * b3[j] + a4 * b4[j];
}
// scatter variant c[bix[j]] += aval*b[j], j in [bi, bi+len).  LMM:3213-3237
inline void vmadd_scatter(double av, const double* b, double* c, const int* bix, i64 bi, i64 len) {
  for (i64 j = bi; j < bi + len; j++) c[bix[j]] += av * b[j];
}
// vectAddAll over task partials, LMM:3435-3443 (+vectAdd :3367-3389, vectAdd4 :3411-3433)
void vadd_all(const std::vector<std::vector<double>>& a, double* c, i64 len) {
  size_t bi = a.size() % 4;
  for (size_t i = 0; i < bi; i++)
    for (i64 j = 0; j < len; j++) c[j] += a[i][j];
  for (size_t i = bi; i < a.size(); i += 4)
    for (i64 j = 0; j < len; j++) c[j] += a[i][j] + a[i + 1][j] + a[i + 2][j] + a[i + 3][j];
}

const i64 L2_CACHESIZE = 256 * 1024;             // LMM:64
const i64 MEM_OVERHEAD_THRESHOLD = 2LL * 1024 * 1024;  // LMM:61
const i64 PAR_MINFLOP_THRESHOLD2 = 128LL * 1024;  // LMM:63 (shared thread pool: k == #vcores)

// DenseBlockFactory.createDenseBlock: one contiguous double[] only below 2^31-1 cells
// (src/main/java/org/apache/sysml/runtime/matrix/data/DenseBlockFactory.java:25-29)
inline bool drb_contiguous(i64 rows, i64 cols) { return rows * cols < (i64)std::numeric_limits<int32_t>::max(); }

// ---------------------------------------------------------------------------------------------
// mmchain.  LMM:1594-1653 (dense), LMM:1655-1684 (sparse).  ct: 0=XtXv 1=XtwXv 2=XtXvy
// ---------------------------------------------------------------------------------------------
void mmchain_dense_range(const double* X, const double* b, const double* w, double* c, i64 rows_total, i64 cd,
                         int ct, i64 rl, i64 ru) {
  const bool weights = (ct == 1), weights2 = (ct == 2);
  const int blocksizeI = 24;
  const i64 blocksizeJ = 1024;
  double tmp[24];
  const i64 bn = (ru - rl) % blocksizeI;
  const bool contiguous = drb_contiguous(rows_total, cd);
  for (i64 i = rl; i < rl + bn; i++) {  // LMM:1612-1619
    const double* arow = X + i * cd;
    double val = dot_dense(arow, b, cd);
    val *= weights ? w[i] : 1;
    val -= weights2 ? w[i] : 0;
    vmadd1(val, arow, c, cd);
  }
  for (i64 bi = rl + bn; bi < ru; bi += blocksizeI) {  // LMM:1622-1652
    for (int i = 0; i < blocksizeI; i++) tmp[i] = 0;
    for (i64 bj = 0; bj < cd; bj += blocksizeJ) {
      i64 bjmin = std::min(cd - bj, blocksizeJ);
      for (int i = 0; i < blocksizeI; i++) tmp[i] += dot_dense(X + (bi + i) * cd + bj, b + bj, bjmin);
    }
    if (weights)
      for (int i = 0; i < blocksizeI; i++) tmp[i] *= w[bi + i];
    else if (weights2)
      for (int i = 0; i < blocksizeI; i++) tmp[i] -= w[bi + i];
    for (i64 bj = 0; bj < cd; bj += blocksizeJ) {
      i64 bjmin = std::min(cd - bj, blocksizeJ);
      if (contiguous) {
        for (int i = 0; i < blocksizeI; i += 4) {
          const double* r0 = X + (bi + i) * cd + bj;
          vmadd4(tmp[i], tmp[i + 1], tmp[i + 2], tmp[i + 3], r0, r0 + cd, r0 + 2 * cd, r0 + 3 * cd, c + bj, bjmin);
        }
      } else {  // LDRB: per-row update, LMM:1647-1650
        for (int i = 0; i < blocksizeI; i++) vmadd1(tmp[i], X + (bi + i) * cd + bj, c + bj, bjmin);
      }
    }
  }
}

void mmchain_csr_range(const i64* rp, const int* ci, const double* va, const double* b, const double* w, double* c,
                       int ct, i64 rl, i64 ru) {
  const bool weights = (ct == 1), weights2 = (ct == 2);
  for (i64 i = rl; i < ru; i++) {
    i64 apos = rp[i], alen = rp[i + 1] - rp[i];
    if (alen == 0 || (weights && w[i] == 0)) continue;
    double val = dot_sparse(va, b, ci, apos, alen);
    val *= weights ? w[i] : 1;
    val -= weights2 ? w[i] : 0;
    if (val != 0) vmadd_scatter(val, va, c, ci, apos, alen);
  }
}

template <class F>
void run_mmchain(i64 n, i64 m, bool sparse, int k, double* out, F range_fn) {
  // threading rule: LMM:284-329 with satisfiesMultiThreadingConstraints(mX,true,true,sparse?2:4,k) LMM:3737-3743
  std::fill(out, out + m, 0.0);
  bool par = k > 1 && (8LL * m * k < MEM_OVERHEAD_THRESHOLD) && ((sparse ? 2 : 4) * n * m > PAR_MINFLOP_THRESHOLD2);
  if (!par) {
    range_fn(out, (i64)0, n);
    return;
  }
  std::vector<int> blk = balanced_blocks((int)std::min<i64>(n, INT32_MAX), k, true);
  // note: the reference indexes rows with int; n < 2^31 always holds for a MatrixBlock
  std::vector<i64> lb(blk.size() + 1, 0);
  for (size_t i = 0; i < blk.size(); i++) lb[i + 1] = lb[i] + blk[i];
  std::vector<std::vector<double>> part(blk.size(), std::vector<double>(m, 0.0));
#pragma omp parallel for schedule(dynamic, 1) num_threads(k)
  for (int t = 0; t < (int)blk.size(); t++) range_fn(part[t].data(), lb[t], lb[t + 1]);
  vadd_all(part, out, m);
}

// ---------------------------------------------------------------------------------------------
// tsmm.  LMM:1686-1837 (dense), LMM:1839-1960 (sparse), copyUpperToLowerTriangle LMM:3633-3672
// ---------------------------------------------------------------------------------------------
void tsmm_left_dense_range(const double* A, double* C, i64 m /*rows of X*/, i64 n /*cols of X*/, i64 rl, i64 ru) {
  if (n == 1) {  // LMM:1698-1702
    C[0] = dot_dense(A, A, m);
    return;
  }
  const i64 blocksizeI = 32, blocksizeK = 24, blocksizeJ = 1024;
  double ta[24];
  const double* tb[24];
  for (i64 bi = rl; bi < ru; bi += blocksizeI) {
    i64 bimin = std::min(ru, bi + blocksizeI);
    for (i64 bk = 0; bk < m; bk += blocksizeK) {
      i64 bkmin = std::min(m, bk + blocksizeK);
      for (i64 bj = bi; bj < n; bj += blocksizeJ) {
        i64 bklen = bkmin - bk;
        i64 bjlen = std::min(n, bj + blocksizeJ) - bj;
        for (i64 i = bi; i < bimin; i++) {
          double* crow = C + i * n + bj;
          // copyNonZeroElements (stride-n variant) LMM:3805-3816
          int knnz = 0;
          for (i64 k = 0; k < bklen; k++) {
            double av = A[(bk + k) * n + i];
            if (av != 0) {
              ta[knnz] = av;
              tb[knnz] = A + (bk + k) * n + bj;
              knnz++;
            }
          }
          const int bn = knnz % 4;
          switch (bn) {  // LMM:1743-1747
            case 1: vmadd1(ta[0], tb[0], crow, bjlen); break;
            case 2: vmadd2(ta[0], ta[1], tb[0], tb[1], crow, bjlen); break;
            case 3: vmadd3(ta[0], ta[1], ta[2], tb[0], tb[1], tb[2], crow, bjlen); break;
          }
          for (int k = bn; k < knnz; k += 4)
            vmadd4(ta[k], ta[k + 1], ta[k + 2], ta[k + 3], tb[k], tb[k + 1], tb[k + 2], tb[k + 3], crow, bjlen);
        }
      }
    }
  }
}

void tsmm_right_dense_range(const double* A, double* C, i64 m, i64 n, i64 rl, i64 ru) {  // LMM:1787-1820
  if (m == 1) {
    C[0] = dot_dense(A, A, n);
    return;
  }
  const i64 blocksizeK = 1024;
  const i64 blocksizeIJ = L2_CACHESIZE / 8 / blocksizeK / 2 - 1;  // 15
  for (i64 bi = rl; bi < ru; bi += blocksizeIJ) {
    i64 bimin = std::min(ru, bi + blocksizeIJ);
    for (i64 bk = 0; bk < n; bk += blocksizeK) {
      i64 bklen = std::min(blocksizeK, n - bk);
      for (i64 bj = bi; bj < m; bj += blocksizeIJ) {
        i64 bjmin = std::min(m, bj + blocksizeIJ);
        for (i64 i = bi; i < bimin; i++) {
          i64 bjmax = std::max(i, bj);
          for (i64 j = bjmax; j < bjmin; j++) C[i * m + j] += dot_dense(A + i * n + bk, A + j * n + bk, bklen);
        }
      }
    }
  }
}

// position of first column index >= col in row r, or -1 (SparseBlockCSR.posFIndexGTE)
inline i64 pos_f_index_gte(const i64* rp, const int* ci, i64 r, i64 col) {
  const int* b = ci + rp[r];
  const int* e = ci + rp[r + 1];
  const int* p = std::lower_bound(b, e, (int)col);
  return (p == e) ? -1 : (p - b);
}

void tsmm_left_csr_range(const i64* rp, const int* ci, const double* va, double* C, i64 rows, i64 n, i64 rl, i64 ru) {
  // LMM:1851-1879
  for (i64 r = 0; r < rows; r++) {
    i64 apos = rp[r], alen = rp[r + 1] - rp[r];
    if (alen == 0) continue;
    const double* avals = va + apos;
    if (alen == n) {  // dense row
      for (i64 i = rl; i < ru; i++) vmadd1(avals[i], avals + i, C + i * n + i, n - i);
    } else {
      i64 rlix = (rl == 0) ? 0 : pos_f_index_gte(rp, ci, r, rl);
      rlix = (rlix >= 0) ? apos + rlix : apos + alen;
      i64 len = apos + alen;
      for (i64 i = rlix; i < len && ci[i] < ru; i++) {
        double val = va[i];
        if (val != 0) vmadd_scatter(val, va, C + (i64)ci[i] * n, ci, i, len - i);
      }
    }
  }
}

i64 copy_upper_to_lower(double* c, i64 n) {  // LMM:3633-3672; nnz: diagonal counts 1, upper off-diagonal counts 2
  i64 nnz = 0;
  for (i64 i = 0; i < n; i++) {
    nnz += (c[i * n + i] != 0) ? 1 : 0;
    for (i64 j = i + 1; j < n; j++) {
      double v = c[i * n + j];
      c[j * n + i] = v;
      nnz += (v != 0) ? 2 : 0;
    }
  }
  return nnz;
}

// ---------------------------------------------------------------------------------------------
// general matmult.  LMM:86-229 dispatch, dense-dense LMM:880-1103, sparse-dense LMM:1170-1395
// ---------------------------------------------------------------------------------------------
struct DD {  // dense-dense problem: C[m x n] = A[m x cd] * B[cd x n]
  const double* A;
  const double* B;  // when tm2: B holds t(B) as n x cd
  double* C;
  i64 m, cd, n;
  bool tm2, pm2;
};

void mm_dd_mm(const DD& p, i64 rl, i64 ru, i64 cl, i64 cu) {  // matrixMultDenseDenseMM LMM:1047-1103
  const i64 blocksizeI = 32, blocksizeK = 24, blocksizeJ = 1024;
  double ta[24];
  const double* tb[24];
  for (i64 bi = rl; bi < ru; bi += blocksizeI) {
    i64 bimin = std::min(ru, bi + blocksizeI);
    for (i64 bk = 0; bk < p.cd; bk += blocksizeK) {
      i64 bkmin = std::min(p.cd, bk + blocksizeK);
      for (i64 bj = cl; bj < cu; bj += blocksizeJ) {
        i64 bklen = bkmin - bk;
        i64 bjlen = std::min(cu, bj + blocksizeJ) - bj;
        for (i64 i = bi; i < bimin; i++) {
          const double* arow = p.A + i * p.cd + bk;
          double* crow = p.C + i * p.n + bj;
          int knnz = 0;  // copyNonZeroElements LMM:3792-3803
          for (i64 k = 0; k < bklen; k++)
            if (arow[k] != 0) {
              ta[knnz] = arow[k];
              tb[knnz] = p.B + (bk + k) * p.n + bj;
              knnz++;
            }
          const int bn = knnz % 4;
          switch (bn) {
            case 1: vmadd1(ta[0], tb[0], crow, bjlen); break;
            case 2: vmadd2(ta[0], ta[1], tb[0], tb[1], crow, bjlen); break;
            case 3: vmadd3(ta[0], ta[1], ta[2], tb[0], tb[1], tb[2], crow, bjlen); break;
          }
          for (int k = bn; k < knnz; k += 4)
            vmadd4(ta[k], ta[k + 1], ta[k + 2], ta[k + 3], tb[k], tb[k + 1], tb[k + 2], tb[k + 3], crow, bjlen);
        }
      }
    }
  }
}

// rl/ru: row range of A (or of B when pm2), cl/cu: column range of B
void mm_dense_dense(const DD& p, i64 rl, i64 ru, i64 cl, i64 cu) {  // LMM:880-948
  const i64 m = p.m, n = p.n, cd = p.cd;
  if (m == 1 && n == 1) {  // dot product
    p.C[0] = dot_dense(p.A, p.B, cd);
  } else if (n > 1 && cd == 1) {  // outer product LMM:894-907
    for (i64 i = rl; i < ru; i++) {
      double* crow = p.C + i * n;
      if (p.A[i] == 1)
        std::memcpy(crow, p.B, n * sizeof(double));
      else if (p.A[i] != 0)
        for (i64 j = 0; j < n; j++) crow[j] = p.A[i] * p.B[j];
      else
        std::fill(crow, crow + n, 0.0);
    }
  } else if (n == 1 && cd == 1) {  // vector-scalar
    for (i64 i = rl; i < ru; i++) p.C[i] = p.B[0] * p.A[i];
  } else if (n == 1 && cd <= 2 * 1024) {  // MV short rhs LMM:950-955
    for (i64 i = rl; i < ru; i++) p.C[i] = dot_dense(p.A + i * cd, p.B, cd);
  } else if (n == 1) {  // MV tall rhs LMM:957-970
    const i64 blocksizeI = 32, blocksizeK = 2 * 1024;
    for (i64 bi = rl; bi < ru; bi += blocksizeI) {
      i64 bimin = std::min(bi + blocksizeI, ru);
      for (i64 bk = 0; bk < cd; bk += blocksizeK) {
        i64 bkmin = std::min(bk + blocksizeK, cd);
        for (i64 i = bi; i < bimin; i++) p.C[i] += dot_dense(p.A + i * cd + bk, p.B + bk, bkmin - bk);
      }
    }
  } else if (p.pm2 && m == 1) {  // VM LMM:972-993 (rows rl..ru of B)
    const bool contiguous = drb_contiguous(cd, n);
    const i64 kn = contiguous ? rl + (ru - rl) % 2 : ru;
    for (i64 k = rl; k < kn; k++)
      if (p.A[k] != 0) vmadd1(p.A[k], p.B + k * n, p.C, n);
    for (i64 k = kn; k < ru; k += 2) {
      const double* b0 = p.B + k * n;
      if (p.A[k] != 0 && p.A[k + 1] != 0)
        vmadd2(p.A[k], p.A[k + 1], b0, b0 + n, p.C, n);
      else if (p.A[k] != 0)
        vmadd1(p.A[k], b0, p.C, n);
      else if (p.A[k + 1] != 0)
        vmadd1(p.A[k + 1], b0 + n, p.C, n);
    }
  } else if (p.pm2 && m <= 16) {  // MM short lhs LMM:995-1034 (rows rl..ru of B)
    const i64 kn = (ru - rl) % 4;
    for (i64 i = 0; i < m; i++)
      for (i64 k = rl; k < rl + kn; k++)
        if (p.A[i * cd + k] != 0) vmadd1(p.A[i * cd + k], p.B + k * n, p.C + i * n, n);
    const i64 blocksizeK = 48, blocksizeJ = 1024;
    for (i64 bk = rl + kn; bk < ru; bk += blocksizeK) {
      i64 bkmin = std::min(ru, bk + blocksizeK);
      for (i64 bj = 0; bj < n; bj += blocksizeJ) {
        i64 bjlen = std::min(n, bj + blocksizeJ) - bj;
        for (i64 i = 0; i < m; i++) {
          const double* arow = p.A + i * cd;
          double* crow = p.C + i * n + bj;
          for (i64 k = bk; k < bkmin; k += 4) {
            const double* b0 = p.B + k * n + bj;
            vmadd4(arow[k], arow[k + 1], arow[k + 2], arow[k + 3], b0, b0 + n, b0 + 2 * n, b0 + 3 * n, crow, bjlen);
          }
        }
      }
    }
  } else if (p.tm2) {  // MM skinny rhs, B pre-transposed LMM:1036-1045
    for (i64 i = rl; i < ru; i++)
      for (i64 j = 0; j < n; j++) p.C[i * n + j] = dot_dense(p.A + i * cd, p.B + j * cd, cd);
  } else {
    mm_dd_mm(p, rl, ru, cl, cu);
  }
}

struct SD {  // sparse(CSR)-dense: C[m x n] = A[m x cd] * B[cd x n]
  const i64* rp;
  const int* ci;
  const double* va;
  const double* B;
  double* C;
  i64 m, cd, n, nnz;
  bool pm2;
};

void mm_sparse_dense(const SD& p, i64 rl, i64 ru) {  // LMM:1170-1221
  const i64 m = p.m, n = p.n, cd = p.cd;
  const i64 xsp = (i64)m * cd / std::max<i64>(p.nnz, 1);
  const bool bcontig = drb_contiguous(cd, n);
  if (m == 1 && n == 1) {
    if (p.rp[1] > p.rp[0]) p.C[0] = dot_sparse(p.va, p.B, p.ci, p.rp[0], p.rp[1] - p.rp[0]);
  } else if (n == 1 && cd <= 2 * 1024) {  // MV short rhs LMM:1223-1234
    for (i64 i = rl; i < ru; i++) {
      i64 apos = p.rp[i], alen = p.rp[i + 1] - apos;
      if (alen == 0) continue;
      p.C[i] = (alen == cd) ? dot_dense(p.va + apos, p.B, cd) : dot_sparse(p.va, p.B, p.ci, apos, alen);
    }
  } else if (n == 1) {  // MV tall rhs LMM:1236-1260
    const i64 blocksizeI = 32;
    const i64 blocksizeK = std::max<i64>(2 * 1024, 2 * 1024 * xsp / 32);
    i64 curk[32];
    for (i64 bi = rl; bi < ru; bi += blocksizeI) {
      std::fill(curk, curk + 32, 0);
      i64 bimin = std::min(ru, bi + blocksizeI);
      for (i64 bk = 0; bk < cd; bk += blocksizeK) {
        i64 bkmin = std::min(bk + blocksizeK, cd);
        for (i64 i = bi; i < bimin; i++) {
          i64 apos = p.rp[i], alen = p.rp[i + 1] - apos;
          if (alen == 0) continue;
          i64 k = curk[i - bi] + apos;
          for (; k < apos + alen && p.ci[k] < bkmin; k++) p.C[i] += p.va[k] * p.B[p.ci[k]];
          curk[i - bi] = k - apos;
        }
      }
    }
  } else if (p.pm2 && m == 1) {  // VM LMM:1262-1286 (rows rl..ru of B)
    i64 alen = p.rp[1] - p.rp[0];
    if (alen == 0) return;
    i64 rlix = (rl == 0) ? 0 : pos_f_index_gte(p.rp, p.ci, 0, rl);
    rlix = (rlix >= 0) ? rlix : alen;
    if (bcontig) {
      for (i64 k = rlix; k < alen && p.ci[k] < ru; k++) {
        if (k + 1 < alen && p.ci[k + 1] < ru) {
          vmadd2(p.va[k], p.va[k + 1], p.B + (i64)p.ci[k] * n, p.B + (i64)p.ci[k + 1] * n, p.C, n);
          k++;
        } else
          vmadd1(p.va[k], p.B + (i64)p.ci[k] * n, p.C, n);
      }
    } else {
      for (i64 k = rlix; k < alen && p.ci[k] < ru; k++) vmadd1(p.va[k], p.B + (i64)p.ci[k] * n, p.C, n);
    }
  } else if (p.pm2 && m <= 16) {  // MM short lhs LMM:1288-1326
    for (i64 i = 0; i < m; i++) {
      i64 apos = p.rp[i], alen = p.rp[i + 1] - apos;
      if (alen == 0) continue;
      double* crow = p.C + i * n;
      i64 k1 = (rl == 0) ? 0 : pos_f_index_gte(p.rp, p.ci, i, rl);
      k1 = (k1 >= 0) ? apos + k1 : apos + alen;
      i64 k2 = (ru == cd) ? alen : pos_f_index_gte(p.rp, p.ci, i, ru);
      k2 = (k2 >= 0) ? apos + k2 : apos + alen;
      const int bn = (int)((k2 - k1) % 4);
      auto brow = [&](i64 k) { return p.B + (i64)p.ci[k] * n; };
      switch (bn) {
        case 1: vmadd1(p.va[k1], brow(k1), crow, n); break;
        case 2: vmadd2(p.va[k1], p.va[k1 + 1], brow(k1), brow(k1 + 1), crow, n); break;
        case 3: vmadd3(p.va[k1], p.va[k1 + 1], p.va[k1 + 2], brow(k1), brow(k1 + 1), brow(k1 + 2), crow, n); break;
      }
      for (i64 k = k1 + bn; k < k2; k += 4)
        vmadd4(p.va[k], p.va[k + 1], p.va[k + 2], p.va[k + 3], brow(k), brow(k + 1), brow(k + 2), brow(k + 3), crow, n);
    }
  } else if (n <= 64) {  // MM skinny rhs LMM:1328-1347
    for (i64 i = rl; i < ru; i++) {
      i64 apos = p.rp[i], alen = p.rp[i + 1] - apos;
      if (alen == 0) continue;
      double* crow = p.C + i * n;
      i64 bn = bcontig ? alen % 4 : alen;
      for (i64 k = apos; k < apos + bn; k++) vmadd1(p.va[k], p.B + (i64)p.ci[k] * n, crow, n);
      for (i64 k = apos + bn; k < apos + alen; k += 4)
        vmadd4(p.va[k], p.va[k + 1], p.va[k + 2], p.va[k + 3], p.B + (i64)p.ci[k] * n, p.B + (i64)p.ci[k + 1] * n,
               p.B + (i64)p.ci[k + 2] * n, p.B + (i64)p.ci[k + 3] * n, crow, n);
    }
  } else {  // general MM LMM:1349-1395
    const i64 blocksizeI = std::max<i64>(1, 8 * xsp), blocksizeK = std::max<i64>(1, 8 * xsp), blocksizeJ = 1024;
    std::vector<i64> curk(std::min(blocksizeI, ru - rl));
    for (i64 bi = rl; bi < ru; bi += blocksizeI) {
      std::fill(curk.begin(), curk.end(), 0);
      i64 bimin = std::min(ru, bi + blocksizeI);
      for (i64 bk = 0; bk < cd; bk += blocksizeK) {
        i64 bkmin = std::min(cd, bk + blocksizeK);
        for (i64 bj = 0; bj < n; bj += blocksizeJ) {
          i64 bjlen = std::min(n, bj + blocksizeJ) - bj;
          for (i64 i = bi; i < bimin; i++) {
            i64 apos = p.rp[i], alen = p.rp[i + 1] - apos;
            if (alen == 0) continue;
            double* crow = p.C + i * n + bj;
            i64 k = curk[i - bi] + apos;
            i64 bn = bcontig ? alen % 4 : alen;
            for (; k < apos + bn && p.ci[k] < bkmin; k++) vmadd1(p.va[k], p.B + (i64)p.ci[k] * n + bj, crow, bjlen);
            for (; k < apos + alen && p.ci[k] < bkmin; k += 4)
              vmadd4(p.va[k], p.va[k + 1], p.va[k + 2], p.va[k + 3], p.B + (i64)p.ci[k] * n + bj,
                     p.B + (i64)p.ci[k + 1] * n + bj, p.B + (i64)p.ci[k + 2] * n + bj, p.B + (i64)p.ci[k + 3] * n + bj,
                     crow, bjlen);
            if (bj + bjlen == n) curk[i - bi] = k - apos;
          }
        }
      }
    }
  }
}

inline i64 count_nnz(const double* a, i64 len) {
  i64 c = 0;
  for (i64 i = 0; i < len; i++) c += (a[i] != 0) ? 1 : 0;
  return c;
}

// ---------------------------------------------------------------------------------------------
// Kahan addition.  FO/KahanPlus.java:63-102 (execute with 2 and 3 operands)
// ---------------------------------------------------------------------------------------------
struct Kahan {
  double sum = 0, corr = 0;
};
inline void kahan_add(Kahan& k, double in) {  // execute2, KahanPlus.java:89-102
  if (std::isinf(k.sum) || std::isinf(in)) {
    k.sum = std::isinf(in) ? in : k.sum;
    k.corr = 0;
    return;
  }
  double correction = in + k.corr;
  double sum = k.sum + correction;
  k.corr = correction - (sum - k.sum);
  k.sum = sum;
}
inline void kahan_add3(Kahan& k, double in2, double in3) {  // execute(Data,double,double), KahanPlus.java:63-78
  if (std::isinf(k.sum) || std::isinf(in2)) {
    k.sum = std::isinf(in2) ? in2 : k.sum;
    k.corr = 0;
    return;
  }
  double correction = in2 + (k.corr + in3);
  double sum = k.sum + correction;
  k.corr = correction - (sum - k.sum);
  k.sum = sum;
}
// Math.max / Math.min semantics (NaN propagates; -0.0 < +0.0).  FO/Builtin.java:356-361
inline double jmax(double a, double b) {
  if (a != a) return a;
  if (b != b) return b;
  if (a == 0.0 && b == 0.0) return std::signbit(a) ? b : a;
  return (a >= b) ? a : b;
}
inline double jmin(double a, double b) {
  if (a != a) return a;
  if (b != b) return b;
  if (a == 0.0 && b == 0.0) return std::signbit(a) ? a : b;
  return (a <= b) ? a : b;
}

}  // namespace

// =============================================================================================
// C entry points (ctypes).  Row pointers are int64 on this side so the same arrays can feed the
// product's int64-extent ABI; the reference's SparseBlockCSR uses int[] (SparseBlockCSR.java:50-53).
// =============================================================================================
extern "C" {

int orc_num_threads() {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

// Launchers such as torchrun export OMP_NUM_THREADS=1 to every rank; the CPU arm of bench.py asks for the
// cores this process may run on (sched_getaffinity) instead, as the reference's k = #vcores rule does
// (InfrastructureAnalyzer.getLocalParallelism -> LibMatrixMult.matrixMultChain(.., k), LibMatrixMult.java:284).
void orc_set_num_threads(int k) {
#ifdef _OPENMP
  if (k >= 1) omp_set_num_threads(k);
#else
  (void)k;
#endif
}

// TestUtils.generateTestMatrix(rows, cols, min, max, sparsity, seed) -> row-major out (pre-zeroed here)
void orc_java_test_matrix(int rows, int cols, double min, double max, double sparsity, i64 seed, double* out) {
  JavaRandom rnd(seed);
  for (i64 i = 0; i < (i64)rows * cols; i++) {
    out[i] = 0;
    if (rnd.nextDouble() > sparsity) continue;
    out[i] = rnd.nextDouble() * (max - min) + min;
  }
}

// first n values of new java.util.Random(seed).nextDouble() (known-answer pin for JavaRandom)
void orc_java_next_doubles(i64 seed, int n, double* out) {
  JavaRandom rnd(seed);
  for (int i = 0; i < n; i++) out[i] = rnd.nextDouble();
}

int orc_balanced_blocks(int len, int k, int constK, int* out) {
  std::vector<int> b = balanced_blocks(len, k, constK != 0);
  for (size_t i = 0; i < b.size(); i++) out[i] = b[i];
  return (int)b.size();
}

// LibMatrixMult.matrixMultChain (LMM:243-268 sequential when k<=1, LMM:284-329 with k threads)
void orc_mmchain_dense(const double* X, const double* v, const double* w, double* out, i64 n, i64 m, int ct, int k) {
  run_mmchain(n, m, false, k, out,
              [&](double* c, i64 rl, i64 ru) { mmchain_dense_range(X, v, w, c, n, m, ct, rl, ru); });
}
void orc_mmchain_csr(const i64* rp, const int* ci, const double* va, const double* v, const double* w, double* out,
                     i64 n, i64 m, int ct, int k) {
  run_mmchain(n, m, true, k, out,
              [&](double* c, i64 rl, i64 ru) { mmchain_csr_range(rp, ci, va, v, w, c, ct, rl, ru); });
}

// LibMatrixMult.matrixMultTransposeSelf (LMM:331-355, 357-399).  out is dim x dim, returns nnz.
i64 orc_tsmm_dense(const double* X, double* out, i64 rows, i64 cols, int left, int k) {
  const i64 dim = left ? cols : rows;
  std::fill(out, out + dim * dim, 0.0);
  // satisfiesMultiThreadingConstraintsTSMM LMM:3752-3759
  bool par = k > 1 && dim != 1 &&
             (double)(left ? rows * cols * cols : cols * rows * rows) > (double)PAR_MINFLOP_THRESHOLD2;
  auto range = [&](i64 rl, i64 ru) {
    if (left)
      tsmm_left_dense_range(X, out, rows, cols, rl, ru);
    else
      tsmm_right_dense_range(X, out, rows, cols, rl, ru);
  };
  if (!par) {
    range(0, dim);
  } else {
    i64 blklen = (i64)std::ceil((double)dim / (2 * k));  // LMM:380-383
    int ntasks = 0;
    for (int i = 0; i < 2 * k && i * blklen < dim; i++) ntasks++;
#pragma omp parallel for schedule(dynamic, 1) num_threads(k)
    for (int i = 0; i < ntasks; i++) range(i * blklen, std::min((i + 1) * blklen, dim));
  }
  return copy_upper_to_lower(out, dim);
}

// t(X) %*% X for CSR X (LEFT only; RIGHT transposes first, LMM:3679-3685)
i64 orc_tsmm_csr_left(const i64* rp, const int* ci, const double* va, double* out, i64 rows, i64 cols, int k) {
  std::fill(out, out + cols * cols, 0.0);
  bool par = k > 1 && cols != 1 && (double)(rows * cols * cols) > (double)PAR_MINFLOP_THRESHOLD2;
  if (!par) {
    tsmm_left_csr_range(rp, ci, va, out, rows, cols, 0, cols);
  } else {
    i64 blklen = (i64)std::ceil((double)cols / (2 * k));
    int ntasks = 0;
    for (int i = 0; i < 2 * k && i * blklen < cols; i++) ntasks++;
#pragma omp parallel for schedule(dynamic, 1) num_threads(k)
    for (int i = 0; i < ntasks; i++)
      tsmm_left_csr_range(rp, ci, va, out, rows, cols, i * blklen, std::min((i + 1) * blklen, cols));
  }
  return copy_upper_to_lower(out, cols);
}

// LibMatrixMult.matrixMult dense x dense (LMM:86-154 seq, 165-229 par).  Returns nnz(C).
i64 orc_matmult_dense(const double* A, const double* B, double* C, i64 m, i64 cd, i64 n, int k) {
  std::fill(C, C + m * n, 0.0);
  // checkPrepMatrixMultRightInput / isSkinnyRightHandSide LMM:3706-3716
  bool tm2 = m > n && cd > n && n > 1 && n < 64 && 8 * cd * n < L2_CACHESIZE;
  std::vector<double> Bt;
  const double* Buse = B;
  if (tm2) {
    Bt.resize(cd * n);
    for (i64 i = 0; i < cd; i++)
      for (i64 j = 0; j < n; j++) Bt[j * cd + i] = B[i * n + j];
    Buse = Bt.data();
  }
  // note: after the transpose prep the reference evaluates m2.rlen/m2.clen on the TRANSPOSED block
  const i64 m2r = tm2 ? n : cd, m2c = tm2 ? cd : n;
  bool par = k > 1 && (m != 1 || 8LL * n * k < MEM_OVERHEAD_THRESHOLD) && (2.0 * m * cd * n > (double)PAR_MINFLOP_THRESHOLD2);
  DD p{A, Buse, C, m, cd, n, tm2, false};
  if (!par) {
    // LMM:129-131: pm2 = checkParMatrixMultRightInputRows(m1, m2, Integer.MAX_VALUE)
    bool pm2 = (m == 1 && m2c > 1) || (m <= 16 && m2c > 1 && m2r > m && (double)INT32_MAX * 8 * m * m2c < (double)MEM_OVERHEAD_THRESHOLD);
    p.pm2 = pm2;
    i64 ru2 = pm2 ? m2r : m;
    mm_dense_dense(p, 0, ru2, 0, m2c);
    return count_nnz(C, m * n);
  }
  bool pm2r = (m == 1 && m2c > 1) || (m <= 16 && m2c > 1 && m2r > m && (i64)k * 8 * m * m2c < MEM_OVERHEAD_THRESHOLD);
  bool pm2c = !tm2 && m2c > (i64)k * 1024 && m < (i64)k * 32 && !pm2r && 8 * m * cd < 256 * 1024;  // LMM:3726-3731
  i64 num = pm2r ? m2r : pm2c ? m2c : m;
  std::vector<int> blk = balanced_blocks((int)num, k, pm2r || pm2c);
  std::vector<i64> lb(blk.size() + 1, 0);
  for (size_t i = 0; i < blk.size(); i++) lb[i + 1] = lb[i] + blk[i];
  p.pm2 = pm2r;
  if (pm2r) {  // private outputs per task, added in task order with vectAdd (LMM:211-214)
    std::vector<std::vector<double>> part(blk.size(), std::vector<double>(m * n, 0.0));
#pragma omp parallel for schedule(dynamic, 1) num_threads(k)
    for (int t = 0; t < (int)blk.size(); t++) {
      DD q = p;
      q.C = part[t].data();
      mm_dense_dense(q, lb[t], lb[t + 1], 0, m2c);
    }
    for (size_t t = 0; t < blk.size(); t++)
      for (i64 j = 0; j < m * n; j++) C[j] += part[t][j];
  } else {
#pragma omp parallel for schedule(dynamic, 1) num_threads(k)
    for (int t = 0; t < (int)blk.size(); t++) {
      if (pm2c)
        mm_dense_dense(p, 0, m, lb[t], lb[t + 1]);
      else
        mm_dense_dense(p, lb[t], lb[t + 1], 0, m2c);
    }
  }
  return count_nnz(C, m * n);
}

// LibMatrixMult.matrixMult CSR x dense.  Returns nnz(C).
i64 orc_matmult_csr_dense(const i64* rp, const int* ci, const double* va, const double* B, double* C, i64 m, i64 cd,
                          i64 n, int k) {
  std::fill(C, C + m * n, 0.0);
  const i64 nnz = rp[m];
  if (nnz == 0) return 0;
  bool par = k > 1 && (m != 1 || 8LL * n * k < MEM_OVERHEAD_THRESHOLD) && (2.0 * m * cd * n > (double)PAR_MINFLOP_THRESHOLD2);
  SD p{rp, ci, va, B, C, m, cd, n, nnz, false};
  if (!par) {
    bool pm2 = (m == 1 && n > 1) || (m <= 16 && n > 1 && cd > m && (double)INT32_MAX * 8 * m * n < (double)MEM_OVERHEAD_THRESHOLD);
    p.pm2 = pm2;
    mm_sparse_dense(p, 0, pm2 ? cd : m);
    return count_nnz(C, m * n);
  }
  bool pm2r = (m == 1 && n > 1) || (m <= 16 && n > 1 && cd > m && (i64)k * 8 * m * n < MEM_OVERHEAD_THRESHOLD);
  i64 num = pm2r ? cd : m;
  std::vector<int> blk = balanced_blocks((int)num, k, pm2r);
  std::vector<i64> lb(blk.size() + 1, 0);
  for (size_t i = 0; i < blk.size(); i++) lb[i + 1] = lb[i] + blk[i];
  p.pm2 = pm2r;
  if (pm2r) {
    std::vector<std::vector<double>> part(blk.size(), std::vector<double>(m * n, 0.0));
#pragma omp parallel for schedule(dynamic, 1) num_threads(k)
    for (int t = 0; t < (int)blk.size(); t++) {
      SD q = p;
      q.C = part[t].data();
      mm_sparse_dense(q, lb[t], lb[t + 1]);
    }
    for (size_t t = 0; t < blk.size(); t++)
      for (i64 j = 0; j < m * n; j++) C[j] += part[t][j];
  } else {
#pragma omp parallel for schedule(dynamic, 1) num_threads(k)
    for (int t = 0; t < (int)blk.size(); t++) mm_sparse_dense(p, lb[t], lb[t + 1]);
  }
  return count_nnz(C, m * n);
}

// ---------------------------------------------------------------------------------------------
// Aggregate-unary.  LibMatrixAgg.aggregateUnaryMatrix LMA:194-282.
// op : 0 SUM(k+) 1 SUMSQ(sqk+) 2 MAX 3 MIN 4 MAXINDEX 5 MININDEX 6 MEAN 7 PROD
// dir: 0 ALL (ua*)  1 ROW-wise = per-row result (uar*, ReduceCol)  2 COL-wise = per-column result (uac*, ReduceRow)
// out: values only (the Kahan correction cells the reference keeps beside them go to `corr` when
//      non-null, same length as out).  MAXINDEX/MININDEX: out = 1-based index, corr = the value.
// ---------------------------------------------------------------------------------------------

}  // extern "C"

static void agg_dense_range(const double* a, i64 n, int op, int dir, double* out, double* corr, i64 rl, i64 ru) {
  // LMA:1260-1363 dispatch; kernels LMA:1624-1669 (k+), 1685-1750 (sqk+), 1843-1889 (min/max), 1902-1927 (index)
  switch (op) {
    case 0:
    case 1: {
      const bool sq = (op == 1);
      if (dir == 0) {
        Kahan kb;
        for (i64 i = rl * n; i < ru * n; i++) kahan_add(kb, sq ? a[i] * a[i] : a[i]);
        out[0] = kb.sum;
        corr[0] = kb.corr;
      } else if (dir == 1) {
        for (i64 i = rl; i < ru; i++) {
          Kahan kb;
          for (i64 j = 0; j < n; j++) kahan_add(kb, sq ? a[i * n + j] * a[i * n + j] : a[i * n + j]);
          out[i] = kb.sum;
          corr[i] = kb.corr;
        }
      } else {
        for (i64 i = rl; i < ru; i++)
          for (i64 j = 0; j < n; j++) {  // sumAgg LMA:2900-2912
            Kahan kb{out[j], corr[j]};
            kahan_add(kb, sq ? a[i * n + j] * a[i * n + j] : a[i * n + j]);
            out[j] = kb.sum;
            corr[j] = kb.corr;
          }
      }
      break;
    }
    case 2:
    case 3: {
      const bool mx = (op == 2);
      const double init = mx ? -INFINITY : INFINITY;
      if (dir == 0) {
        double t = init;
        for (i64 i = rl * n; i < ru * n; i++) t = mx ? jmax(t, a[i]) : jmin(t, a[i]);
        out[0] = t;
      } else if (dir == 1) {
        for (i64 i = rl; i < ru; i++) {
          double t = init;
          for (i64 j = 0; j < n; j++) t = mx ? jmax(t, a[i * n + j]) : jmin(t, a[i * n + j]);
          out[i] = t;
        }
      } else {
        for (i64 j = 0; j < n; j++) out[j] = init;
        for (i64 i = rl; i < ru; i++)
          for (i64 j = 0; j < n; j++) out[j] = mx ? jmax(out[j], a[i * n + j]) : jmin(out[j], a[i * n + j]);
      }
      break;
    }
    case 4:
    case 5: {  // indexmax/indexmin LMA:3081-3099: >= / <= so the LAST extremum wins; NaN never wins
      const bool mx = (op == 4);
      for (i64 i = rl; i < ru; i++) {
        double best = mx ? -INFINITY : INFINITY;
        i64 idx = -1;
        for (i64 j = 0; j < n; j++) {
          double x = a[i * n + j];
          bool take = mx ? (x >= best) : (x <= best);
          idx = take ? j : idx;
          best = take ? x : best;
        }
        out[i] = (double)idx + 1;
        if (corr) corr[i] = (idx >= 0) ? a[i * n + idx] : NAN;  // reference reads a[i,-1] here (undefined)
      }
      break;
    }
    case 7: {  // product LMA:2929-2950 (NAN_AWARENESS=false: early abort at 0)
      auto prod = [](const double* p, i64 len, i64 stride) {
        double v = 1;
        for (i64 i = 0; i < len; i++) {
          v *= p[i * stride];
          if (v == 0) break;
        }
        return v;
      };
      if (dir == 0)
        out[0] = prod(a + rl * n, (ru - rl) * n, 1);
      else if (dir == 1)
        for (i64 i = rl; i < ru; i++) out[i] = prod(a + i * n, n, 1);
      else {
        for (i64 j = 0; j < n; j++) out[j] = 1;  // d_uacm LMA:2121-2135
        for (i64 i = rl; i < ru; i++)
          for (i64 j = 0; j < n; j++) out[j] *= a[i * n + j];
      }
      break;
    }
  }
}

static void agg_csr_range(const i64* rp, const int* ci, const double* va, i64 n, int op, int dir, double* out,
                          double* corr, i64 rl, i64 ru) {
  // LMA:1365-1460 dispatch; s_uakp/s_uarkp/s_uackp LMA:2137-2195; s_uamxx.. LMA:2426-2534; s_uarimxx LMA:2536-2614
  switch (op) {
    case 0:
    case 1: {
      const bool sq = (op == 1);
      if (dir == 0) {
        Kahan kb;
        for (i64 k = rp[rl]; k < rp[ru]; k++) kahan_add(kb, sq ? va[k] * va[k] : va[k]);
        out[0] = kb.sum;
        corr[0] = kb.corr;
      } else if (dir == 1) {
        for (i64 i = rl; i < ru; i++) {
          if (rp[i + 1] == rp[i]) continue;
          Kahan kb;
          for (i64 k = rp[i]; k < rp[i + 1]; k++) kahan_add(kb, sq ? va[k] * va[k] : va[k]);
          out[i] = kb.sum;
          corr[i] = kb.corr;
        }
      } else {
        for (i64 k = rp[rl]; k < rp[ru]; k++) {
          Kahan kb{out[ci[k]], corr[ci[k]]};
          kahan_add(kb, sq ? va[k] * va[k] : va[k]);
          out[ci[k]] = kb.sum;
          corr[ci[k]] = kb.corr;
        }
      }
      break;
    }
    case 2:
    case 3: {
      const bool mx = (op == 2);
      const double init = mx ? -INFINITY : INFINITY;
      auto f = [&](double x, double y) { return mx ? jmax(x, y) : jmin(x, y); };
      if (dir == 0) {
        double val = init;
        i64 alen = rp[ru] - rp[rl];
        for (i64 k = rp[rl]; k < rp[ru]; k++) val = f(val, va[k]);
        double ret = f(init, val);
        if (alen < (ru - rl) * n) ret = f(ret, 0);
        out[0] = ret;
      } else if (dir == 1) {
        for (i64 i = rl; i < ru; i++) {
          double t = init;
          for (i64 k = rp[i]; k < rp[i + 1]; k++) t = f(t, va[k]);
          if (rp[i + 1] - rp[i] < n) t = f(t, 0);
          out[i] = t;
        }
      } else {
        std::vector<i64> cnt(n, 0);
        for (i64 j = 0; j < n; j++) out[j] = init;
        for (i64 k = rp[rl]; k < rp[ru]; k++) {
          out[ci[k]] = f(out[ci[k]], va[k]);
          cnt[ci[k]]++;
        }
        for (i64 j = 0; j < n; j++)
          if (cnt[j] < ru - rl) out[j] = f(out[j], 0);
      }
      break;
    }
    case 4:
    case 5: {
      const bool mx = (op == 4);
      for (i64 i = rl; i < ru; i++) {
        i64 apos = rp[i], alen = rp[i + 1] - apos;
        if (alen == 0) {  // LMA:2559-2563
          out[i] = (double)n;
          if (corr) corr[i] = 0;
          continue;
        }
        double best = mx ? -INFINITY : INFINITY;
        i64 idx = -1;
        for (i64 k = 0; k < alen; k++) {
          double x = va[apos + k];
          bool take = mx ? (x >= best) : (x <= best);
          idx = take ? k : idx;
          best = take ? x : best;
        }
        double bv = (idx >= 0) ? va[apos + idx] : NAN;
        out[i] = (idx >= 0) ? (double)ci[apos + idx] + 1 : 0;
        if (corr) corr[i] = bv;
        // Builtin.execute(0, bv)==1 <=> 0 > bv (max) / 0 < bv (min), strictly.  FO/Builtin.java:374-393
        bool zero_wins = mx ? (0 > bv) : (0 < bv);
        if (alen < n && zero_wins) {
          i64 ix = n - 1;
          for (i64 j = apos + alen - 1; j >= apos; j--, ix--)
            if (ci[j] != ix) break;
          out[i] = (double)(ix + 1);
          if (corr) corr[i] = 0;
        }
      }
      break;
    }
  }
}

// aggregateFinalResult for partial results, LMA:648-662 + MatrixBlock.incrementalAggregate
// (matrix/data/MatrixBlock.java:3092-3180) -> aggregateBinaryMatrixLast{Row,Col}DenseGeneric LMA:1126-1218
static void agg_merge(int op, i64 len, double* out, double* corr, const double* pout, const double* pcorr) {
  if (op == 0 || op == 1) {
    for (i64 j = 0; j < len; j++) {
      Kahan kb{out[j], corr[j]};
      kahan_add3(kb, pout[j], pcorr[j]);
      out[j] = kb.sum;
      corr[j] = kb.corr;
    }
  } else if (op == 2) {
    for (i64 j = 0; j < len; j++) out[j] = jmax(out[j], pout[j]);
  } else if (op == 3) {
    for (i64 j = 0; j < len; j++) out[j] = jmin(out[j], pout[j]);
  } else if (op == 7) {
    for (i64 j = 0; j < len; j++) out[j] = out[j] * pout[j];
  }
}

template <class F>
static void agg_driver(i64 rows, i64 cols, i64 nnz, int op, int dir, double* out, double* corr, int k, F range_fn) {
  const i64 outlen = (dir == 0) ? 1 : (dir == 1 ? rows : cols);
  std::vector<double> corr_local;
  if (!corr) {
    corr_local.assign(outlen, 0.0);
    corr = corr_local.data();
  }
  std::fill(out, out + outlen, 0.0);
  std::fill(corr, corr + outlen, 0.0);
  // satisfiesMultiThreadingConstraints LMA:541-546 (sharedTP)
  bool par = k > 1 && rows > k / 8 && (dir == 1 || (dir == 0 ? 2 : cols) * 8 * k < 2 * 1024 * 1024) && nnz > 16 * 1024;
  if (!par) {
    range_fn(out, corr, (i64)0, rows);
    return;
  }
  std::vector<int> blk = balanced_blocks((int)rows, k, dir == 2);  // static partitioning for col*(), LMA:257-258
  std::vector<i64> lb(blk.size() + 1, 0);
  for (size_t i = 0; i < blk.size(); i++) lb[i + 1] = lb[i] + blk[i];
  if (dir == 1) {  // RowAggTask: disjoint rows of the shared output
#pragma omp parallel for schedule(dynamic, 1) num_threads(k)
    for (int t = 0; t < (int)blk.size(); t++) range_fn(out, corr, lb[t], lb[t + 1]);
    return;
  }
  std::vector<std::vector<double>> po(blk.size(), std::vector<double>(outlen, 0.0));
  std::vector<std::vector<double>> pc(blk.size(), std::vector<double>(outlen, 0.0));
#pragma omp parallel for schedule(dynamic, 1) num_threads(k)
  for (int t = 0; t < (int)blk.size(); t++) range_fn(po[t].data(), pc[t].data(), lb[t], lb[t + 1]);
  std::copy(po[0].begin(), po[0].end(), out);  // out.copy(first partial), LMA:266
  std::copy(pc[0].begin(), pc[0].end(), corr);
  for (size_t t = 1; t < blk.size(); t++) agg_merge(op, outlen, out, corr, po[t].data(), pc[t].data());
}

extern "C" {

// VARIANCE: Welford-style update with Kahan-compensated mean and m2 (FO/CM.java:166-176), result
// m2/(w-1), 0 for a single value (instructions/cp/CM_COV_Object.java:169-170).  Sequential form.
static double var_seq(const double* p, i64 len, i64 stride) {
  Kahan mean, m2;
  double w = 0;
  for (i64 i = 0; i < len; i++) {
    const double x = p[i * stride];
    const double w1 = w + 1;
    const double d = x - mean.sum;
    kahan_add(mean, d / w1);
    const double t1 = w / w1 * d;
    kahan_add(m2, t1 * d);
    w = w1;
  }
  return w == 1.0 ? 0.0 : m2.sum / (w - 1);
}

void orc_agg_unary_dense(const double* a, i64 rows, i64 cols, int op, int dir, double* out, double* corr, int k) {
  if (op == 8) {
    if (dir == 0)
      out[0] = var_seq(a, rows * cols, 1);
    else if (dir == 1)
      for (i64 i = 0; i < rows; i++) out[i] = var_seq(a + i * cols, cols, 1);
    else
      for (i64 j = 0; j < cols; j++) out[j] = var_seq(a + j, rows, cols);
    return;
  }
  if (op == 6) {  // MEAN: incremental Kahan mean, d_uamean/d_uarmean/d_uacmean LMA:1940-2010 with
                  // mean() LMA:2962-2966 and FO/Mean.java:60-64.  Sequential form only (k ignored).
    auto mean_seq = [](const double* p, i64 len, i64 stride) {  // Mean.execute: FO/Mean.java
      Kahan kb;
      double cnt = 0;
      for (i64 i = 0; i < len; i++) {
        cnt += 1;
        double delta = (p[i * stride] - kb.sum) / cnt;  // mean' = mean + (x-mean)/n with Kahan
        kahan_add(kb, delta);
      }
      return kb.sum;
    };
    if (dir == 0)
      out[0] = mean_seq(a, rows * cols, 1);
    else if (dir == 1)
      for (i64 i = 0; i < rows; i++) out[i] = mean_seq(a + i * cols, cols, 1);
    else
      for (i64 j = 0; j < cols; j++) out[j] = mean_seq(a + j, rows, cols);
    return;
  }
  agg_driver(rows, cols, rows * cols, op, dir, out, corr, k,
             [&](double* o, double* c, i64 rl, i64 ru) { agg_dense_range(a, cols, op, dir, o, c, rl, ru); });
}

void orc_agg_unary_csr(const i64* rp, const int* ci, const double* va, i64 rows, i64 cols, int op, int dir,
                       double* out, double* corr, int k) {
  agg_driver(rows, cols, rp[rows], op, dir, out, corr, k,
             [&](double* o, double* c, i64 rl, i64 ru) { agg_csr_range(rp, ci, va, cols, op, dir, o, c, rl, ru); });
}

// ---------------------------------------------------------------------------------------------
// Cellwise binary.  LibMatrixBincell.bincellOp LMB:87-132, 195-339, 938-1081 with the value functions
// in FO/{Plus,Minus,Multiply,Divide,Power,LessThan,...,Builtin,And,Or,Minus1Multiply,MinusNz,Modulus,
// IntegerDivide}.java.  Opcode numbering = LibMatrixCUDA.getBinaryOp (LibMatrixCUDA.java:1551-1575).
// ---------------------------------------------------------------------------------------------
static inline double intdiv(double a, double b) {  // FO/IntegerDivide.java:62-73, UtilFunctions.toLong :327-329
  double r = a / b;
  if (std::isnan(r) || std::isinf(r)) return r;
  return (double)(i64)std::floor(r + 0x1.0p-53);
}
double orc_binary_scalar(double x, double y, int op) {
  switch (op) {
    case 0: return x + y;
    case 1: return x - y;
    case 2: return x * y;
    case 3: return x / y;
    case 4: return std::pow(x, y);
    case 5: return (x < y) ? 1.0 : 0.0;
    case 6: return (x <= y) ? 1.0 : 0.0;
    case 7: return (x > y) ? 1.0 : 0.0;
    case 8: return (x >= y) ? 1.0 : 0.0;
    case 9: return (x == y) ? 1.0 : 0.0;
    case 10: return (x != y) ? 1.0 : 0.0;
    case 11: return jmin(x, y);
    case 12: return jmax(x, y);
    case 13: return ((x != 0) && (y != 0)) ? 1.0 : 0.0;  // FO/And.java:46-48
    case 14: return ((x != 0) || (y != 0)) ? 1.0 : 0.0;  // FO/Or.java:46-48
    case 15: return 1 - x * y;                           // FO/Minus1Multiply.java:41-43
    case 16: return (x != 0) ? x - y : 0;                // FO/MinusNz.java:42-44
    case 17: {                                           // FO/Modulus.java:49-53
      if (y == 0.0) return NAN;
      return x - intdiv(x, y) * y;
    }
    case 18: return intdiv(x, y);
  }
  return NAN;
}
// Builtin ROUND is Math.round(double) converted back to double (FO/Builtin.java:314).  java.lang.Math.round, JDK 8 (a
// dependency outside the reference tree; its specification: "the closest long to the argument, with ties rounding to
// positive infinity"; NaN -> 0; values beyond the long range saturate).  Restated from the JDK 8 source: with the
// unbiased shift s = 1074 - biasedExponent, for 0 <= s < 64 the result is ((+-significand >> s) + 1) >> 1, otherwise the
// Java narrowing cast (long) a.
double orc_java_round(double a) {
  uint64_t bits;
  memcpy(&bits, &a, 8);
  const int64_t biasedExp = (int64_t)((bits & 0x7ff0000000000000ULL) >> 52);
  const int64_t shift = (52 - 1 + 1023) - biasedExp;
  if ((shift & -64) == 0) {
    int64_t r = (int64_t)((bits & 0x000fffffffffffffULL) | 0x0010000000000000ULL);
    if ((int64_t)bits < 0) r = -r;
    return (double)(((r >> shift) + 1) >> 1);
  }
  if (a != a) return 0.0;                                   // (long) NaN == 0
  if (a >= 9223372036854775807.0) return 9223372036854775807.0;   // saturation: Long.MAX_VALUE, as a double 2^63
  if (a <= -9223372036854775808.0) return -9223372036854775808.0;
  return (double)(int64_t)a;                                // |a| < 0.5 (-> 0) or an integer of magnitude >= 2^52 ... 2^63
}
// Builtin SIGN is Math.signum (FO/Builtin.java:311): zero and NaN are returned as they are (the sign of a zero survives)
double orc_java_signum(double a) { return (a > 0.0) ? 1.0 : ((a < 0.0) ? -1.0 : a); }
void orc_unary_round(const double* A, double* C, i64 len) {
  for (i64 i = 0; i < len; i++) C[i] = orc_java_round(A[i]);
}
void orc_unary_signum(const double* A, double* C, i64 len) {
  for (i64 i = 0; i < len; i++) C[i] = orc_java_signum(A[i]);
}
// bcast: 0 = B is rows x cols, 1 = B is a column vector (rows x 1), 2 = B is a row vector (1 x cols)
void orc_binary_mm(const double* A, const double* B, double* C, i64 rows, i64 cols, int bcast, int op) {
#pragma omp parallel for schedule(static)
  for (i64 i = 0; i < rows; i++)
    for (i64 j = 0; j < cols; j++) {
      // bcast 3: OUTER_VECTOR_VECTOR, A is rows x 1 and B is 1 x cols (safeBinaryVVGeneric, LibMatrixBincell.java:565-589)
      double b = bcast == 0 ? B[i * cols + j] : (bcast == 1 ? B[i] : B[j]);
      double a = bcast == 3 ? A[i] : A[i * cols + j];
      C[i * cols + j] = orc_binary_scalar(a, b, op);
    }
}
void orc_binary_ms(const double* A, double s, double* C, i64 len, int op, int scalar_left) {
#pragma omp parallel for schedule(static)
  for (i64 i = 0; i < len; i++) C[i] = scalar_left ? orc_binary_scalar(s, A[i], op) : orc_binary_scalar(A[i], s, op);
}
// +* / -* (axpy): C = A + alpha*B with B broadcast like above.  LibMatrixCUDA.axpy :2295-2350,
// CP twin PlusMultiply/MinusMultiply (FO/PlusMultiply.java): in1 + cnt*in2
void orc_axpy(const double* A, const double* B, double alpha, double* C, i64 rows, i64 cols, int bcast) {
#pragma omp parallel for schedule(static)
  for (i64 i = 0; i < rows; i++)
    for (i64 j = 0; j < cols; j++) {
      double b = bcast == 0 ? B[i * cols + j] : (bcast == 1 ? B[i] : B[j]);
      C[i * cols + j] = A[i * cols + j] + alpha * b;
    }
}

i64 orc_count_nnz(const double* a, i64 len) { return count_nnz(a, len); }

// dense -> CSR in row order (MatrixBlock.denseToSparse: matrix/data/MatrixBlock.java:1017-1080), nnz = (!= 0)
i64 orc_dense_to_csr(const double* a, i64 rows, i64 cols, i64* rp, int* ci, double* va) {
  i64 k = 0;
  rp[0] = 0;
  for (i64 i = 0; i < rows; i++) {
    for (i64 j = 0; j < cols; j++)
      if (a[i * cols + j] != 0) {
        if (ci) {
          ci[k] = (int)j;
          va[k] = a[i * cols + j];
        }
        k++;
      }
    rp[i + 1] = k;
  }
  return k;
}

// ---------------------------------------------------------------------------------------------
// Synthetic-data generator shared (bit-exactly) with the product's device generator
// (systemml_b200/csrc/rand.cu): counter-based splitmix64 hash of (seed, global cell index).
// Not part of the reference; it exists so a row shard on any rank equals the same rows of the
// global matrix and so the CPU baseline sees the very same inputs as the GPU.
// ---------------------------------------------------------------------------------------------
static inline uint64_t splitmix64(uint64_t x) {
  x += 0x9E3779B97F4A7C15ULL;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ULL;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBULL;
  return x ^ (x >> 31);
}
void orc_rand_uniform(double* out, i64 len, uint64_t seed, i64 offset, double lo, double hi) {
  const uint64_t s = splitmix64(seed);
#pragma omp parallel for schedule(static)
  for (i64 i = 0; i < len; i++) {
    uint64_t h = splitmix64(s ^ (uint64_t)(offset + i));
    double u = (double)(h >> 11) * 0x1.0p-53;  // [0,1)
    out[i] = lo + (hi - lo) * u;
  }
}


// CPU twin of sysml_b200_rand_csr (systemml_b200/csrc/rand.cu): same hashes, same separately rounded recurrences.
static inline int orc_csr_row_count(uint64_t hr, double mean, double p0, i64 cols) {
  const uint64_t h = splitmix64(hr);
  const double u0 = (double)(h >> 11) * 0x1.0p-53;
  double pk = p0, cdf = p0;
  int k = 0;
  const int cap = (int)(cols < 4096 ? cols : 4096);
  while (cdf <= u0 && k < cap) {
    k++;
    pk = (pk * mean) / (double)k;
    cdf = cdf + pk;
    if (pk < 1e-300) break;
  }
  return k;
}
// rp: rows+1 int64 (filled here); returns nnz.  Call with ci == NULL first to size the arrays.
i64 orc_rand_csr(i64 rows, i64 cols, double mean, uint64_t seed, i64 row_offset, i64* rp, int* ci, double* va) {
  const uint64_t s = splitmix64(seed ^ 0xC5A1ull);
  const double p0 = std::exp(-mean);
  rp[0] = 0;
  for (i64 r = 0; r < rows; r++) rp[r + 1] = rp[r] + orc_csr_row_count(splitmix64(s ^ (uint64_t)(row_offset + r)), mean, p0, cols);
  if (!ci || !va) return rp[rows];
#pragma omp parallel for schedule(static)
  for (i64 r = 0; r < rows; r++) {
    const uint64_t hr = splitmix64(s ^ (uint64_t)(row_offset + r));
    const i64 lo = rp[r];
    const int c = (int)(rp[r + 1] - lo);
    if (c <= 0) continue;
    const uint64_t W = (uint64_t)(cols / c);
    for (int j = 0; j < c; j++) {
      const uint64_t hc = splitmix64(hr + 2ull * (uint64_t)j + 1ull), hv = splitmix64(hr + 2ull * (uint64_t)j + 2ull);
      ci[lo + j] = (int)((uint64_t)j * W + hc % W);
      const double u = (double)((hv >> 11) | 1ull) * 0x1.0p-53;
      va[lo + j] = -1.0 + 2.0 * u;
    }
  }
  return rp[rows];
}

// ---------------------------------------------------------------------------------------------
// Sparse-safe cellwise operations with a sparse result.  Results are appended through SparseRow.append /
// MatrixBlock.appendValue, both of which drop zeros.  Two-call use: ci == NULL sizes the result (rp is filled).
// ---------------------------------------------------------------------------------------------
struct CsrOut {
  i64* rp;
  int* ci;
  double* va;
  i64 k = 0;
  void append(int col, double v) {
    if (v == 0) return;
    if (ci) {
      ci[k] = col;
      va[k] = v;
    }
    k++;
  }
};
static double orc_unary_scalar(double x, int op) {
  switch (op) {
    case 1: return std::sqrt(x);
    case 2: return orc_java_round(x);
    case 3: return std::fabs(x);
    case 5: return std::floor(x);
    case 6: return std::ceil(x);
    case 7: return orc_java_signum(x);
    case 9: return std::sin(x);
    case 11: return std::tan(x);
    case 12: return std::tanh(x);
  }
  return NAN;  // exp, log, sigmoid, cos are not sparse-safe
}
// LibMatrixBincell.safeBinaryScalar, SPARSE <- SPARSE (LMB:952-995); unary = MatrixBlock.sparseUnaryOperations (MB:2640-2672)
// mode 0: a op s, 1: s op a, 2: unary
i64 orc_csr_scalar_op(const i64* rp, const int* ci, const double* va, i64 rows, double s, int op, int mode, i64* crp, int* cci,
                      double* cva) {
  CsrOut c{crp, cci, cva};
  crp[0] = 0;
  for (i64 r = 0; r < rows; r++) {
    for (i64 j = rp[r]; j < rp[r + 1]; j++) {
      const double val = mode == 2 ? orc_unary_scalar(va[j], op) : (mode == 1 ? orc_binary_scalar(s, va[j], op) : orc_binary_scalar(va[j], s, op));
      c.append(ci[j], val);
    }
    crp[r + 1] = c.k;
  }
  return c.k;
}
// sparse op dense.  bcast 1: safeBinaryMVSparse MATRIX_COL_VECTOR (LMB:359-405), bcast 2: MATRIX_ROW_VECTOR (LMB:406-440),
// bcast 0: safeBinaryMMSparseDenseSkip (LMB:765-789; Multiply / sparse-safe Divide).  The vector branches evaluate the
// operator at the cells a row does NOT store as well ("empty left"), except where Multiply lets them skip.
i64 orc_csr_dense_op(const i64* rp, const int* ci, const double* va, i64 rows, i64 cols, const double* B, int bcast, int op, i64* crp,
                     int* cci, double* cva) {
  const bool isMultiply = (op == 2), skipEmpty = isMultiply;
  CsrOut c{crp, cci, cva};
  crp[0] = 0;
  for (i64 i = 0; i < rows; i++) {
    const bool empty = rp[i] == rp[i + 1];
    if (bcast == 0) {
      for (i64 k = rp[i]; k < rp[i + 1]; k++) {
        const double in2 = B[i * cols + ci[k]];
        if (in2 == 0) continue;
        c.append(ci[k], orc_binary_scalar(va[k], in2, op));
      }
    } else if (bcast == 1) {
      const double v2 = B[i];
      if (!((skipEmpty && (empty || v2 == 0)) || (empty && v2 == 0))) {
        if (isMultiply && v2 == 1) {  // ROW COPY
          for (i64 k = rp[i]; k < rp[i + 1]; k++) c.append(ci[k], va[k]);
        } else {
          // "empty left": every cell the row does not store gets op(0, v2) -- one value per row, so it is computed once and
          // the gap loops are skipped when it is zero (appendValue would drop each of them)
          const double gv = orc_binary_scalar(0, v2, op);
          i64 lastIx = -1;
          for (i64 j = rp[i]; j < rp[i + 1]; j++) {
            for (i64 k = lastIx + 1; gv != 0 && k < ci[j]; k++) c.append((int)k, gv);
            c.append(ci[j], orc_binary_scalar(va[j], v2, op));
            lastIx = ci[j];
          }
          for (i64 k = lastIx + 1; gv != 0 && k < cols; k++) c.append((int)k, gv);
        }
      }
    } else {
      if (!(skipEmpty && empty)) {
        i64 lastIx = -1;
        for (i64 j = rp[i]; j < rp[i + 1]; j++) {
          for (i64 k = lastIx + 1; !skipEmpty && k < ci[j]; k++) c.append((int)k, orc_binary_scalar(0, B[k], op));
          c.append(ci[j], orc_binary_scalar(va[j], B[ci[j]], op));
          lastIx = ci[j];
        }
        for (i64 k = lastIx + 1; !skipEmpty && k < cols; k++) c.append((int)k, orc_binary_scalar(0, B[k], op));
      }
    }
    crp[i + 1] = c.k;
  }
  return c.k;
}
// sparse op sparse: safeBinaryMMSparseSparse (LMB:597-660) -> mergeForSparseBinary (LMB:1279-1322): inner join under
// Multiply, otherwise outer join, then appendLeft / appendRight for the remainders (LMB:1324-1340)
i64 orc_csr_csr_op(const i64* arp, const int* aci, const double* ava, const i64* brp, const int* bci, const double* bva, i64 rows, int op,
                   i64* crp, int* cci, double* cva) {
  CsrOut c{crp, cci, cva};
  crp[0] = 0;
  for (i64 r = 0; r < rows; r++) {
    i64 p1 = arp[r], p2 = brp[r];
    const i64 e1 = arp[r + 1], e2 = brp[r + 1];
    if (p1 < e1 && p2 < e2 && op == 2) {
      while (p1 < e1 && p2 < e2) {
        const int c1 = aci[p1], c2 = bci[p2];
        if (c1 == c2) c.append(c1, orc_binary_scalar(ava[p1], bva[p2], op));
        p1 += (c1 <= c2) ? 1 : 0;
        p2 += (c1 >= c2) ? 1 : 0;
      }
    } else {
      while (p1 < e1 && p2 < e2) {
        if (aci[p1] < bci[p2]) {
          c.append(aci[p1], orc_binary_scalar(ava[p1], 0, op));
          p1++;
        } else if (aci[p1] == bci[p2]) {
          c.append(aci[p1], orc_binary_scalar(ava[p1], bva[p2], op));
          p1++;
          p2++;
        } else {
          c.append(bci[p2], orc_binary_scalar(0, bva[p2], op));
          p2++;
        }
      }
      for (; p1 < e1; p1++) c.append(aci[p1], orc_binary_scalar(ava[p1], 0, op));
      for (; p2 < e2; p2++) c.append(bci[p2], orc_binary_scalar(0, bva[p2], op));
    }
    crp[r + 1] = c.k;
  }
  return c.k;
}

// LibMatrixMult.matrixMultUltraSparse, left operand ultra-sparse, right operand dense (LMM:1510-1566); the result is sparse.
// alen == 1: row selection, values appended where b != 0 (recomputeNonZeros / c.append), then scaled in place by aval when
// it is not 1 (vectMultiplyInPlace); general case: c(i,j) = c(i,j) + aval*b for every non-zero product, k ascending
// (quickGetValue / quickSetValue: a value that becomes 0 is removed).
i64 orc_matmult_ultra_sparse_left(const i64* arp, const int* aci, const double* ava, const double* B, i64 m, i64 n, i64* crp, int* cci,
                                  double* cva) {
  CsrOut c{crp, cci, cva};
  crp[0] = 0;
  std::vector<double> row((size_t)n);
  for (i64 i = 0; i < m; i++) {
    const i64 apos = arp[i], alen = arp[i + 1] - arp[i];
    if (alen == 1) {
      const double aval = ava[apos];
      const double* b = B + (i64)aci[apos] * n;
      for (i64 j = 0; j < n; j++)
        if (b[j] != 0) c.append((int)j, aval != 1 ? b[j] * aval : b[j]);
    } else if (alen > 1) {
      std::fill(row.begin(), row.end(), 0.0);
      for (i64 k = apos; k < apos + alen; k++) {
        const double aval = ava[k];
        const double* b = B + (i64)aci[k] * n;
        for (i64 j = 0; j < n; j++) {
          const double cvald = aval * b[j];
          if (cvald != 0) row[j] = row[j] + cvald;
        }
      }
      for (i64 j = 0; j < n; j++) c.append((int)j, row[j]);
    }
    crp[i + 1] = c.k;
  }
  return c.k;
}

}  // extern "C"
