// rand.cu -- counter-based synthetic data generator (see include/sysml_b200.h, "Synthetic data").
// cell i of the global array = lo + (hi-lo) * u,  u = (splitmix64(splitmix64(seed) ^ (offset+i)) >> 11) * 2^-53.
// Every step is exact integer arithmetic or a single correctly rounded FP64 operation WITHOUT
// contraction, so the CPU twin (oracle/sysml_oracle.cpp: orc_rand_uniform) is bit-identical.
#include "common.cuh"

namespace {
__host__ __device__ __forceinline__ uint64_t splitmix64(uint64_t x) {
  x += 0x9E3779B97F4A7C15ULL;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ULL;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBULL;
  return x ^ (x >> 31);
}

__global__ void rand_uniform_kernel(double* __restrict__ out, i64 len, uint64_t s, i64 offset, double lo, double span) {
  const i64 stride = (i64)gridDim.x * blockDim.x;
  for (i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x; i < len; i += stride) {
    const uint64_t h = splitmix64(s ^ (uint64_t)(offset + i));
    const double u = __dmul_rn((double)(h >> 11), 0x1.0p-53);
    out[i] = __dadd_rn(lo, __dmul_rn(span, u));
  }
}
}  // namespace

extern "C" int sysml_b200_rand_uniform(sysml_b200_ctx* ctx, double* out, int64_t len, uint64_t seed, int64_t offset,
                                       double lo, double hi) {
  SYSML_ENTER(ctx);
  if (len < 0 || (len > 0 && !out)) return sysml_set_error(ctx, SYSML_E_INVALID, "rand_uniform: bad arguments");
  if (len == 0) return SYSML_OK;
  i64 blocks = (len + 255) / 256;
  const i64 maxb = (i64)ctx->sm_count * 16;
  if (blocks > maxb) blocks = maxb;
  rand_uniform_kernel<<<(unsigned)blocks, 256, 0, ctx->stream>>>(out, len, splitmix64(seed), offset, lo, hi - lo);
  SYSML_LAUNCH_CHECK(ctx, "rand_uniform_kernel");
  return SYSML_OK;
}
