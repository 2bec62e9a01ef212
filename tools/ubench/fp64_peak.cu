// microbenchmark: DFMA vs DMMA (m8n8k4, m16n8k8) throughput and LDS bandwidth on sm_100a
#include <cstdio>
#include <cuda_runtime.h>
#define ITERS 4096
__global__ void k_dfma(double *out, double a, double b)
{
  double x[16];
  for (int i = 0; i < 16; i++) x[i] = threadIdx.x + i;
  for (int it = 0; it < ITERS; it++) {
#pragma unroll
    for (int i = 0; i < 16; i++) x[i] = fma(x[i], a, b);
  }
  double s = 0; for (int i = 0; i < 16; i++) s += x[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void k_dmma884(double *out, double a, double b)
{
  double c[8][2];
  for (int i = 0; i < 8; i++) { c[i][0] = threadIdx.x; c[i][1] = i; }
  for (int it = 0; it < ITERS; it++) {
#pragma unroll
    for (int i = 0; i < 8; i++)
      asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                   : "+d"(c[i][0]), "+d"(c[i][1]) : "d"(a), "d"(b));
  }
  double s = 0; for (int i = 0; i < 8; i++) s += c[i][0] + c[i][1];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void k_dmma1688(double *out, double a, double b)
{
  double c[4][4];
  for (int i = 0; i < 4; i++) for (int j = 0; j < 4; j++) c[i][j] = threadIdx.x + i + j;
  for (int it = 0; it < ITERS; it++) {
#pragma unroll
    for (int i = 0; i < 4; i++)
      asm volatile("mma.sync.aligned.m16n8k8.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                   : "+d"(c[i][0]), "+d"(c[i][1]), "+d"(c[i][2]), "+d"(c[i][3])
                   : "d"(a), "d"(b), "d"(a), "d"(b), "d"(a), "d"(b));
  }
  double s = 0; for (int i = 0; i < 4; i++) for (int j = 0; j < 4; j++) s += c[i][j];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void k_lds(double *out, int stride)
{
  extern __shared__ double sm[];
  for (int i = threadIdx.x; i < 4096; i += blockDim.x) sm[i] = i;
  __syncthreads();
  double s = 0;
  int base = (threadIdx.x * stride) & 2047;
  for (int it = 0; it < ITERS; it++) {
#pragma unroll
    for (int i = 0; i < 8; i++) {
      double2 v = *reinterpret_cast<double2 *>(&sm[(base + 2 * i * 32 + (it & 1) * 2) & 4094]);
      s += v.x + v.y;
    }
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <class F> float timeit(F f)
{
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  f(); cudaDeviceSynchronize();
  cudaEventRecord(a); f(); cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b); return ms;
}
int main()
{
  double *out; cudaMalloc(&out, 148 * 8 * 1024 * sizeof(double));
  int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
  for (int warps = 4; warps <= 32; warps *= 2) {
    int th = warps * 32; int grid = 148;
    float t1 = timeit([&] { k_dfma<<<grid, th>>>(out, 1.0000001, 1e-9); });
    float t2 = timeit([&] { k_dmma884<<<grid, th>>>(out, 1.0000001, 1e-9); });
    float t3 = timeit([&] { k_dmma1688<<<grid, th>>>(out, 1.0000001, 1e-9); });
    double f1 = 2.0 * 16 * ITERS * th * grid / (t1 * 1e-3) / 1e12;
    double f2 = 2.0 * 8 * 256 * ITERS * warps * grid / (t2 * 1e-3) / 1e12;
    double f3 = 2.0 * 4 * 1024 * ITERS * warps * grid / (t3 * 1e-3) / 1e12;
    printf("warps/SM %2d: DFMA %.2f TF  DMMA884 %.2f TF  DMMA1688 %.2f TF\n", warps, f1, f2, f3);
  }
  cudaFuncSetAttribute(k_lds, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
  for (int stride = 2; stride <= 2; stride++) {
    int th = 512; float t = timeit([&] { k_lds<<<148, th, 32768>>>(out, stride); });
    double bw = 16.0 * 8 * ITERS * th * 148 / (t * 1e-3) / 1e12;
    printf("LDS.128 stride %d: %.2f TB/s aggregate (%.1f B/clk/SM at %d MHz nominal)\n", stride, bw, bw * 1e12 / 148 / (clk * 1e3), clk / 1000);
  }
  return 0;
}
