// cg_tail.cuh -- the m-vector tail of one conjugate-gradient iteration as a device function, shared by the
// stand-alone kernel (cellwise.cu) and by the finishing kernel of the fused CG step (comm.cu), so that both
// produce the same bits.  See cellwise.cu for the reference citation (LinearRegCG.dml:183-195, GLM.dml:1011-1030).
#pragma once
#include "common.cuh"

// sum over the CTA, compensated, every thread returns the same value (any blockDim.x that is a multiple of 32, <= 1024)
__device__ __forceinline__ double block_sum_1024(double v, double* sh) {
  CompSum s;
  s.add(v);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    CompSum t;
    t.hi = __shfl_xor_sync(0xffffffffu, s.hi, o);
    t.lo = __shfl_xor_sync(0xffffffffu, s.lo, o);
    s.add(t);
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __syncthreads();
  if (lane == 0) {
    sh[warp] = s.hi;
    sh[32 + warp] = s.lo;
  }
  __syncthreads();
  CompSum t;
  const int nw = blockDim.x >> 5;
  for (int q = 0; q < nw; q++) {  // every thread adds the warp sums in the same order: identical result
    CompSum u;
    u.hi = sh[q];
    u.lo = sh[32 + q];
    t.add(u);
  }
  return t.value();
}

// q += lambda*p; a = nr2/sum(p*q); beta += a*p; r += a*q; nr2' = sum(r^2); p = -r + (nr2'/nr2)*p; nr2 = nr2'.
// One CTA, all of its threads; sh: 64 doubles of shared memory.  q is read with L2 loads (other CTAs may have
// just written it).  The per-thread partial sums depend on blockDim.x: callers that must agree bit for bit use 1024.
__device__ __forceinline__ void cg_vector_update_body(double* __restrict__ p, double* q, double* __restrict__ r,
                                                      double* __restrict__ beta, double* __restrict__ nr2, i64 m, double lambda,
                                                      double* sh) {
  const double nr2_old = *nr2;
  double part = 0.0;
  for (i64 j = threadIdx.x; j < m; j += blockDim.x) {
    const double pj = p[j];
    const double qj = __dadd_rn(__ldcg(q + j), __dmul_rn(lambda, pj));
    q[j] = qj;
    part = fma(pj, qj, part);
  }
  const double pq = block_sum_1024(part, sh);
  const double a = __ddiv_rn(nr2_old, pq);
  part = 0.0;
  for (i64 j = threadIdx.x; j < m; j += blockDim.x) {
    const double pj = p[j];
    beta[j] = __dadd_rn(beta[j], __dmul_rn(a, pj));
    const double rj = __dadd_rn(r[j], __dmul_rn(a, q[j]));
    r[j] = rj;
    part = fma(rj, rj, part);
  }
  const double nr2_new = block_sum_1024(part, sh);
  const double b = __ddiv_rn(nr2_new, nr2_old);
  for (i64 j = threadIdx.x; j < m; j += blockDim.x) p[j] = __dadd_rn(-r[j], __dmul_rn(b, p[j]));
  __syncthreads();
  if (threadIdx.x == 0) *nr2 = nr2_new;
}
