// mma.cu -- issue rate of the legacy mma.sync tensor path on sm_100a (tf32 m16n8k8, bf16 m16n8k16, s8 m16n8k32) and of FFMA.
// nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o mma mma.cu && ./mma
#include <cstdio>
#include <cuda_runtime.h>

constexpr int kIters = 2048, kTiles = 8;

__global__ void k_tf32(float *out, unsigned seed, long long *cyc) {
  float c[kTiles][4];
  unsigned a[4] = {seed ^ threadIdx.x, seed + 1, seed + 2, seed + 3}, b[2] = {seed + 5, seed + 7};
  for (int t = 0; t < kTiles; t++) for (int i = 0; i < 4; i++) c[t][i] = 0.f;
  long long t0 = clock64();
  for (int it = 0; it < kIters; it++) {
#pragma unroll
    for (int t = 0; t < kTiles; t++)
      asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
                   : "+f"(c[t][0]), "+f"(c[t][1]), "+f"(c[t][2]), "+f"(c[t][3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
  }
  long long t1 = clock64();
  float s = 0; for (int t = 0; t < kTiles; t++) for (int i = 0; i < 4; i++) s += c[t][i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
__global__ void k_bf16(float *out, unsigned seed, long long *cyc) {
  float c[kTiles][4];
  unsigned a[4] = {seed ^ threadIdx.x, seed + 1, seed + 2, seed + 3}, b[2] = {seed + 5, seed + 7};
  for (int t = 0; t < kTiles; t++) for (int i = 0; i < 4; i++) c[t][i] = 0.f;
  long long t0 = clock64();
  for (int it = 0; it < kIters; it++) {
#pragma unroll
    for (int t = 0; t < kTiles; t++)
      asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
                   : "+f"(c[t][0]), "+f"(c[t][1]), "+f"(c[t][2]), "+f"(c[t][3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
  }
  long long t1 = clock64();
  float s = 0; for (int t = 0; t < kTiles; t++) for (int i = 0; i < 4; i++) s += c[t][i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
__global__ void k_s8(float *out, unsigned seed, long long *cyc) {
  int c[kTiles][4];
  unsigned a[4] = {seed ^ threadIdx.x, seed + 1, seed + 2, seed + 3}, b[2] = {seed + 5, seed + 7};
  for (int t = 0; t < kTiles; t++) for (int i = 0; i < 4; i++) c[t][i] = 0;
  long long t0 = clock64();
  for (int it = 0; it < kIters; it++) {
#pragma unroll
    for (int t = 0; t < kTiles; t++)
      asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
                   : "+r"(c[t][0]), "+r"(c[t][1]), "+r"(c[t][2]), "+r"(c[t][3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
  }
  long long t1 = clock64();
  int s = 0; for (int t = 0; t < kTiles; t++) for (int i = 0; i < 4; i++) s += c[t][i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = (float)s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
__global__ void k_ffma(float *out, unsigned seed, long long *cyc) {
  float c[kTiles * 4];
  float a = __uint_as_float(0x3f800000u | (seed & 0xff)), b = 1e-3f * threadIdx.x;
  for (int t = 0; t < kTiles * 4; t++) c[t] = t;
  long long t0 = clock64();
  for (int it = 0; it < kIters; it++) {
#pragma unroll
    for (int t = 0; t < kTiles * 4; t++) c[t] = fmaf(c[t], a, b);
  }
  long long t1 = clock64();
  float s = 0; for (int t = 0; t < kTiles * 4; t++) s += c[t];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

int main() {
  cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
  float *out; long long *cyc; cudaMalloc(&out, 148 * 1024 * 4); cudaMallocManaged(&cyc, 8);
  struct { const char *name; void (*k)(float *, unsigned, long long *); int per_iter; double macs; } es[] = {
    {"mma.sync m16n8k8 tf32", k_tf32, kTiles, 1024}, {"mma.sync m16n8k16 bf16", k_bf16, kTiles, 2048},
    {"mma.sync m16n8k32 s8", k_s8, kTiles, 4096}, {"FFMA", k_ffma, kTiles * 4, 32}};
  for (int nw : {4, 8, 16}) {
    printf("--- %d warps/SM\n", nw);
    for (auto &e : es) {
      e.k<<<p.multiProcessorCount, nw * 32>>>(out, 12345, cyc); cudaDeviceSynchronize();
      e.k<<<p.multiProcessorCount, nw * 32>>>(out, 12345, cyc); cudaDeviceSynchronize();
      double inst_per_clk = (double)kIters * e.per_iter * nw / (double)*cyc;
      printf("%-26s %8lld cyc  %.3f warp-inst/clk/SM  %.0f MAC/clk/SM  -> %.1f TFLOP/s at 1.965 GHz x %d SMs\n", e.name, *cyc, inst_per_clk,
             inst_per_clk * e.macs, inst_per_clk * e.macs * 2 * 1.965e9 * p.multiProcessorCount / 1e12, p.multiProcessorCount);
    }
  }
  return 0;
}
