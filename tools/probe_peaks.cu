// probe_peaks.cu -- box calibration (NOT part of the library): FP64 DMMA issue peak, FP64 DFMA peak,
// and read-only streaming bandwidth.  BASELINE.md section 2 asks for the FP64 tensor peak to be measured
// on the box before any tsmm fraction is quoted.  Build: nvcc -arch=sm_100a -O3 tools/probe_peaks.cu
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdint.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)

__device__ __forceinline__ void dmma_16x8x8(double (&c)[4], const double (&a)[4], const double (&b)[2]) {
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+d"(c[0]), "+d"(c[1]), "+d"(c[2]), "+d"(c[3])
               : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(b[0]), "d"(b[1]));
}
__device__ __forceinline__ void dmma_8x8x4(double (&c)[2], double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(c[0]), "+d"(c[1]) : "d"(a), "d"(b));
}
__device__ __forceinline__ void dmma_16x8x16(double (&c)[4], const double (&a)[8], const double (&b)[4]) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7,%8,%9,%10,%11}, {%12,%13,%14,%15}, {%0,%1,%2,%3};"
               : "+d"(c[0]), "+d"(c[1]), "+d"(c[2]), "+d"(c[3])
               : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(a[4]), "d"(a[5]), "d"(a[6]), "d"(a[7]), "d"(b[0]), "d"(b[1]), "d"(b[2]), "d"(b[3]));
}

template <int ILP>
__global__ void k_dmma_16x8x8(double* out, int iters) {
  double c[ILP][4]; double a[4], b[2];
  for (int i = 0; i < 4; i++) a[i] = threadIdx.x * 1e-3 + i;
  b[0] = 1.0 + threadIdx.x * 1e-6; b[1] = 0.5;
  for (int j = 0; j < ILP; j++) for (int i = 0; i < 4; i++) c[j][i] = 0;
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int j = 0; j < ILP; j++) dmma_16x8x8(c[j], a, b);
  }
  double s = 0; for (int j = 0; j < ILP; j++) for (int i = 0; i < 4; i++) s += c[j][i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int ILP>
__global__ void k_dmma_16x8x16(double* out, int iters) {
  double c[ILP][4]; double a[8], b[4];
  for (int i = 0; i < 8; i++) a[i] = threadIdx.x * 1e-3 + i;
  for (int i = 0; i < 4; i++) b[i] = 1.0 + threadIdx.x * 1e-6 + i;
  for (int j = 0; j < ILP; j++) for (int i = 0; i < 4; i++) c[j][i] = 0;
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int j = 0; j < ILP; j++) dmma_16x8x16(c[j], a, b);
  }
  double s = 0; for (int j = 0; j < ILP; j++) for (int i = 0; i < 4; i++) s += c[j][i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int ILP>
__global__ void k_dmma_8x8x4(double* out, int iters) {
  double c[ILP][2]; double a = threadIdx.x * 1e-3, b = 1.0 + threadIdx.x * 1e-6;
  for (int j = 0; j < ILP; j++) { c[j][0] = 0; c[j][1] = 0; }
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int j = 0; j < ILP; j++) dmma_8x8x4(c[j], a, b);
  }
  double s = 0; for (int j = 0; j < ILP; j++) s += c[j][0] + c[j][1];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int ILP>
__global__ void k_dfma(double* out, int iters) {
  double c[ILP]; double a = 1.0 + threadIdx.x * 1e-9, b = 1e-9;
  for (int j = 0; j < ILP; j++) c[j] = j;
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int j = 0; j < ILP; j++) c[j] = fma(c[j], a, b);
  }
  double s = 0; for (int j = 0; j < ILP; j++) s += c[j];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void k_read(const double2* __restrict__ in, double* out, long long n2) {
  long long stride = (long long)gridDim.x * blockDim.x;
  double s = 0;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n2; i += stride) {
    double2 v;
    asm volatile("ld.global.nc.L1::no_allocate.v2.f64 {%0,%1}, [%2];" : "=d"(v.x), "=d"(v.y) : "l"(in + i));
    s += v.x + v.y;
  }
  if (s == 12345.678) out[0] = s;
}
__global__ void k_copy(const double2* __restrict__ in, double2* __restrict__ out, long long n2) {
  long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n2; i += stride) out[i] = in[i];
}

template <class F> float time_ms(F f, int reps) {
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  f(); cudaDeviceSynchronize();
  float best = 1e30f;
  for (int r = 0; r < reps; r++) {
    cudaEventRecord(e0); f(); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
  }
  return best;
}

int main() {
  cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
  int sms = p.multiProcessorCount;
  printf("{\"gpu\": \"%s\", \"sms\": %d, \"cc\": \"%d.%d\"", p.name, sms, p.major, p.minor);
  double* out; CK(cudaMalloc(&out, sizeof(double) * sms * 16 * 1024));
  const int iters = 4096;
  for (int warps = 4; warps <= 16; warps *= 2) {
    int threads = warps * 32, blocks = sms * 2;
    double nw = (double)blocks * warps;
    float ms;
    ms = time_ms([&] { k_dmma_16x8x8<8><<<blocks, threads>>>(out, iters); }, 5);
    printf(", \"dmma_m16n8k8_w%d_tflops\": %.2f", warps * 2, nw * iters * 8 * (2.0 * 16 * 8 * 8) / (ms * 1e-3) / 1e12);
    ms = time_ms([&] { k_dmma_16x8x16<8><<<blocks, threads>>>(out, iters); }, 5);
    printf(", \"dmma_m16n8k16_w%d_tflops\": %.2f", warps * 2, nw * iters * 8 * (2.0 * 16 * 8 * 16) / (ms * 1e-3) / 1e12);
    ms = time_ms([&] { k_dmma_8x8x4<8><<<blocks, threads>>>(out, iters); }, 5);
    printf(", \"dmma_m8n8k4_w%d_tflops\": %.2f", warps * 2, nw * iters * 8 * (2.0 * 8 * 8 * 4) / (ms * 1e-3) / 1e12);
    ms = time_ms([&] { k_dfma<8><<<blocks, threads>>>(out, iters * 4); }, 5);
    printf(", \"dfma_w%d_tflops\": %.2f", warps * 2, nw * 32 * iters * 4.0 * 8 * 2.0 / (ms * 1e-3) / 1e12);
  }
  CK(cudaGetLastError());
  // sustained DMMA (2 s) to see the power-capped figure
  {
    int threads = 256, blocks = sms * 2; double nw = blocks * 8.0;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    int launches = 0; float ms = 0;
    do { for (int i = 0; i < 20; i++) { k_dmma_16x8x8<8><<<blocks, threads>>>(out, iters * 4); launches++; }
         cudaEventRecord(e1); cudaEventSynchronize(e1); cudaEventElapsedTime(&ms, e0, e1); } while (ms < 2000.f);
    printf(", \"dmma_m16n8k8_sustained_tflops\": %.2f", nw * launches * iters * 4.0 * 8 * (2.0 * 16 * 8 * 8) / (ms * 1e-3) / 1e12);
  }
  // streaming bandwidth: 8 GiB read, and 4 GiB copy
  {
    long long n2 = (8LL << 30) / 16;
    double2* buf; CK(cudaMalloc(&buf, n2 * 16)); CK(cudaMemset(buf, 0, n2 * 16));
    for (int mult = 4; mult <= 16; mult *= 2) {
      float ms = time_ms([&] { k_read<<<sms * mult, 512>>>(buf, out, n2); }, 5);
      printf(", \"read_gbs_x%d\": %.1f", mult, n2 * 16.0 / (ms * 1e-3) / 1e9);
    }
    float ms = time_ms([&] { k_copy<<<sms * 8, 512>>>(buf, buf + n2 / 2, n2 / 2); }, 5);
    printf(", \"copy_gbs\": %.1f", n2 * 16.0 / (ms * 1e-3) / 1e9);
    ms = time_ms([&] { cudaMemcpyAsync(buf + n2 / 2, buf, n2 * 8, cudaMemcpyDeviceToDevice); }, 5);
    printf(", \"memcpy_d2d_gbs\": %.1f", n2 * 16.0 / (ms * 1e-3) / 1e9);
    cudaFree(buf);
  }
  CK(cudaGetLastError());
  printf("}\n");
  return 0;
}
