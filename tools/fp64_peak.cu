// fp64_peak.cu -- measures the fp64 roofline denominators MEASURED_PEAKS.json lacks (SURVEY.md section 7,
// "no fp64 peak -- measure one"): register-resident DFMA and mma.sync f64 (DMMA) throughput on one GPU.
// Prints one JSON object. Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp64_peak fp64_peak.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>

__global__ void dfma_kernel(double* out, int iters, double a, double b) {
  double acc[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) acc[i] = threadIdx.x * 1e-3 + i;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] = fma(acc[i], a, b);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += acc[i];
  if (s == 1.2345) out[0] = s;
}

__device__ __forceinline__ void dmma884(double& d0, double& d1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
               : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}
__global__ void dmma884_kernel(double* out, int iters, double a, double b) {
  double c[8][2];
#pragma unroll
  for (int i = 0; i < 8; ++i) { c[i][0] = threadIdx.x; c[i][1] = i; }
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) dmma884(c[i][0], c[i][1], a, b);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += c[i][0] + c[i][1];
  if (s == 1.2345) out[0] = s;
}

__device__ __forceinline__ void dmma16816(double (&d)[4], const double (&a)[8], const double (&b)[4]) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7,%8,%9,%10,%11}, {%12,%13,%14,%15}, {%0,%1,%2,%3};\n"
               : "+d"(d[0]), "+d"(d[1]), "+d"(d[2]), "+d"(d[3])
               : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(a[4]), "d"(a[5]), "d"(a[6]), "d"(a[7]),
                 "d"(b[0]), "d"(b[1]), "d"(b[2]), "d"(b[3]));
}
__global__ void dmma16816_kernel(double* out, int iters, double av, double bv) {
  double c[4][4];
  double a[8], b[4];
#pragma unroll
  for (int i = 0; i < 8; ++i) a[i] = av + i;
#pragma unroll
  for (int i = 0; i < 4; ++i) b[i] = bv + i;
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) c[i][j] = threadIdx.x + i + j;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 4; ++i) dmma16816(c[i], a, b);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) s += c[i][j];
  if (s == 1.2345) out[0] = s;
}

// dependent-chain latencies (one warp, one chain), in clocks per op
__global__ void lat_kernel(long long* out, double a, double b) {
  double x = threadIdx.x;
  long long t0 = clock64();
#pragma unroll 1
  for (int i = 0; i < 1024; ++i) { x = fma(x, a, b); x = fma(x, a, b); x = fma(x, a, b); x = fma(x, a, b); }
  long long t1 = clock64();
  double y = x + 2.0;
#pragma unroll 1
  for (int i = 0; i < 256; ++i) { y = rsqrt(y) + 1.5; y = rsqrt(y) + 1.5; y = rsqrt(y) + 1.5; y = rsqrt(y) + 1.5; }
  long long t2 = clock64();
  double z = y + 2.0;
#pragma unroll 1
  for (int i = 0; i < 256; ++i) { z = sqrt(z) + 1.5; z = sqrt(z) + 1.5; z = sqrt(z) + 1.5; z = sqrt(z) + 1.5; }
  long long t3 = clock64();
  double w = z + 2.0;
#pragma unroll 1
  for (int i = 0; i < 256; ++i) { w = 1.5 / w + 1.0; w = 1.5 / w + 1.0; w = 1.5 / w + 1.0; w = 1.5 / w + 1.0; }
  long long t4 = clock64();
  double c0 = w, c1 = x;
#pragma unroll 1
  for (int i = 0; i < 1024; ++i) { dmma884(c0, c1, a, b); dmma884(c0, c1, a, b); dmma884(c0, c1, a, b); dmma884(c0, c1, a, b); }
  long long t5 = clock64();
  double v = c0 + c1;
#pragma unroll 1
  for (int i = 0; i < 1024; ++i) { v = __shfl_sync(0xffffffffu, v, (threadIdx.x + 1) & 31) + 1.0; v = __shfl_sync(0xffffffffu, v, (threadIdx.x + 1) & 31) + 1.0; }
  long long t6 = clock64();
  if (threadIdx.x == 0) {
    out[0] = (t1 - t0); out[1] = (t2 - t1); out[2] = (t3 - t2); out[3] = (t4 - t3); out[4] = (t5 - t4); out[5] = (t6 - t5);
    out[6] = (long long)(v);
  }
}

template <typename F>
static double time_ms(F launch, int reps) {
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  launch(); launch();
  cudaDeviceSynchronize();
  double best = 1e30;
  for (int r = 0; r < reps; ++r) {
    cudaEventRecord(e0);
    launch();
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    if (ms < best) best = ms;
  }
  return best;
}

int main() {
  cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
  const int sms = p.multiProcessorCount;
  double* out; cudaMalloc(&out, 64);
  long long* lout; cudaMalloc(&lout, 64);
  const int iters = 1 << 14;
  const int blocks = sms * 8, threads = 256;
  const double nthreads = (double)blocks * threads, nwarps = nthreads / 32;
  double t_dfma = time_ms([&] { dfma_kernel<<<blocks, threads>>>(out, iters, 1.0000001, 1e-9); }, 5);
  double t_884 = time_ms([&] { dmma884_kernel<<<blocks, threads>>>(out, iters, 1.0000001, 1e-9); }, 5);
  double t_16816 = time_ms([&] { dmma16816_kernel<<<blocks, threads>>>(out, iters / 4, 1.0000001, 1e-9); }, 5);
  cudaError_t e = cudaDeviceSynchronize();
  lat_kernel<<<1, 32>>>(lout, 1.0000001, 1e-9);
  long long h[8]; cudaMemcpy(h, lout, 56, cudaMemcpyDeviceToHost);
  double f_dfma = nthreads * iters * 8 * 2 / (t_dfma * 1e-3) / 1e12;
  double f_884 = nwarps * iters * 8 * 512.0 / (t_884 * 1e-3) / 1e12;
  double f_16816 = nwarps * (iters / 4) * 4 * 4096.0 / (t_16816 * 1e-3) / 1e12;
  printf("{\"gpu\": \"%s\", \"sms\": %d, \"sm_clock_khz\": %d, \"dfma_tflops\": %.3f, \"dmma_m8n8k4_tflops\": %.3f, "
         "\"dmma_m16n8k16_tflops\": %.3f, \"lat_clk\": {\"dfma\": %.1f, \"rsqrt\": %.1f, \"sqrt\": %.1f, \"div\": %.1f, "
         "\"dmma884\": %.1f, \"shfl_f64_plus_add\": %.1f}, \"cuda_status\": \"%s\"}\n",
         p.name, sms, p.clockRate, f_dfma, f_884, f_16816, h[0] / 4096.0, h[1] / 1024.0, h[2] / 1024.0, h[3] / 1024.0,
         h[4] / 4096.0, h[5] / 2048.0, cudaGetErrorString(e));
  return 0;
}
