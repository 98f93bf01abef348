// lds.cu -- shared-memory data-pipe cost of LDS.32/64/128 under different broadcast patterns on sm_100a.
// nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o lds lds.cu && ./lds
#include <cstdio>
#include <cuda_runtime.h>
constexpr int kIters = 4096;

template <int W> __device__ inline float lds(unsigned addr);
template <> __device__ inline float lds<1>(unsigned a) { float x; asm volatile("ld.shared.f32 %0, [%1];" : "=f"(x) : "r"(a)); return x; }
template <> __device__ inline float lds<2>(unsigned a) { float x, y; asm volatile("ld.shared.v2.f32 {%0,%1}, [%2];" : "=f"(x), "=f"(y) : "r"(a)); return x + y; }
template <> __device__ inline float lds<4>(unsigned a) {
  float x, y, z, w; asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(x), "=f"(y), "=f"(z), "=f"(w) : "r"(a)); return x + y + z + w;
}

// pattern: byte offset of a lane = (lane / div) * stride_bytes, or (lane % -div) * stride_bytes when div < 0
template <int W>
__global__ void k(float *out, int div, int stride_bytes, long long *cyc, int zero) {
  extern __shared__ __align__(16) unsigned char sm[];
  for (int i = threadIdx.x; i < 8192; i += blockDim.x) reinterpret_cast<float *>(sm)[i] = i;
  __syncthreads();
  const int lane = threadIdx.x & 31;
  unsigned p = (unsigned)__cvta_generic_to_shared(sm) + (div > 0 ? lane / div : lane % (-div)) * stride_bytes;
  float acc = 0.f;
  long long t0 = clock64();
  for (int it = 0; it < kIters; it++) {
#pragma unroll
    for (int u = 0; u < 8; u++) {
      acc += lds<W>(p + u * 2048 + (it & 1) * 1024);
    }
    p += __float_as_int(acc) & zero;      // a dependence the compiler cannot fold: the loads stay in the loop
  }
  long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

// table-driven pattern: lane l reads 16 B at offs[l]
__global__ void k_tab(float *out, const int *offs, long long *cyc, int zero) {
  extern __shared__ __align__(16) unsigned char sm[];
  for (int i = threadIdx.x; i < 8192; i += blockDim.x) reinterpret_cast<float *>(sm)[i] = i;
  __syncthreads();
  unsigned p = (unsigned)__cvta_generic_to_shared(sm) + offs[threadIdx.x & 31];
  float acc = 0.f;
  long long t0 = clock64();
  for (int it = 0; it < kIters; it++) {
#pragma unroll
    for (int u = 0; u < 8; u++) acc += lds<4>(p + u * 2048 + (it & 1) * 1024);
    p += __float_as_int(acc) & zero;
  }
  long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

int main() {
  cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
  float *out; long long *cyc; cudaMalloc(&out, 148 * 1024 * 4); cudaMallocManaged(&cyc, 8);
  struct { const char *name; int w, div, stride; } es[] = {
    {"LDS.32  distinct (4 B/lane)", 1, 1, 4}, {"LDS.32  all lanes one address", 1, 32, 0}, {"LDS.32  8 lanes per address, 4 addresses", 1, 8, 4},
    {"LDS.64  distinct (8 B/lane)", 2, 1, 8}, {"LDS.64  all lanes one address", 2, 32, 0}, {"LDS.64  4 addresses x 8 lanes", 2, 8, 8},
    {"LDS.128 distinct (16 B/lane)", 4, 1, 16}, {"LDS.128 all lanes one address", 4, 32, 0},
    {"LDS.128 quarter-warps one address each (4 x 16 B)", 4, 8, 16},
    {"LDS.128 lane%8 -> 8 addresses (128 B), repeated by quarters", 4, -8, 16},
    {"LDS.128 pairs of lanes share (16 addresses, 256 B)", 4, 2, 16},
    {"LDS.128 4 lanes share (8 addresses, 128 B)", 4, 4, 16},
    {"LDS.128 lane%4 -> 4 addresses (64 B), repeated 8x", 4, -4, 16},
    {"LDS.128 lane%2 -> 2 addresses (32 B), repeated 16x", 4, -2, 16},
  };
  for (auto &e : es) {
    for (int nw : {8, 16}) {
      auto run = [&](auto kern) {
        kern<<<p.multiProcessorCount, nw * 32, 32768>>>(out, e.div, e.stride, cyc, 0); cudaDeviceSynchronize();
        kern<<<p.multiProcessorCount, nw * 32, 32768>>>(out, e.div, e.stride, cyc, 0); cudaDeviceSynchronize();
      };
      if (e.w == 1) run(k<1>); else if (e.w == 2) run(k<2>); else run(k<4>);
      printf("%-60s %2d warps: %.2f clk per warp-LDS (SM-wide)\n", e.name, nw, (double)*cyc / ((double)kIters * 8 * nw));
    }
  }
  struct { const char *name; int (*f)(int); } ts[] = {
    {"K4 weights, g = (lane>>1)&7 (pairs share, repeated by half warps)", [](int l) { return ((l >> 1) & 7) * 16; }},
    {"K4 pixels, (lane&1) + 2*(lane>>4): A,B,A,B.. | C,D,C,D..", [](int l) { return ((l & 1) + 2 * (l >> 4)) * 80; }},
    {"K4 weights, g = lane&7 (today)", [](int l) { return (l & 7) * 16; }},
    {"K4 pixels, lane>>3 (today)", [](int l) { return (l >> 3) * 80; }},
    {"g = (lane>>2)&7 (4 lanes share), 8 addresses", [](int l) { return ((l >> 2) & 7) * 16; }},
    {"pixels lane&3 : A,B,C,D repeated", [](int l) { return (l & 3) * 80; }},
    {"g = (lane>>1)&3 | pairs share, 4 addresses x4", [](int l) { return ((l >> 1) & 3) * 16; }},
    {"pixels (lane&1) + 2*(lane>>3): 8 pixels", [](int l) { return ((l & 1) + 2 * (l >> 3)) * 80; }},
    {"g = (lane>>1)&15 pairs share, 16 addresses (256 B)", [](int l) { return ((l >> 1) & 15) * 16; }},
    {"pixels lane&1: A,B x16", [](int l) { return (l & 1) * 80; }},
  };
  int *offs; cudaMallocManaged(&offs, 128);
  for (auto &t : ts) {
    for (int l = 0; l < 32; l++) offs[l] = t.f(l);
    k_tab<<<p.multiProcessorCount, 512, 32768>>>(out, offs, cyc, 0); cudaDeviceSynchronize();
    k_tab<<<p.multiProcessorCount, 512, 32768>>>(out, offs, cyc, 0); cudaDeviceSynchronize();
    printf("LDS.128 %-72s %.2f clk\n", t.name, (double)*cyc / ((double)kIters * 8 * 16));
  }
  return 0;
}
