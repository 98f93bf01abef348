// Integer-pipe throughput microbenchmark for sm_100a (B200).  DEVELOPMENT TOOL, not product code.
// For each candidate SASS op of the bit-exact q15 front-end it measures warp-instructions per
// cycle per SM.  Every chain is a pair (a, b) updated as a = op(a, b); b = op(b, a) so ptxas cannot
// fold consecutive ops; 8 independent chains per thread give ILP; the outer loop is not unrolled
// so the static SASS count of the loop body is the dynamic count per iteration (check with
// cuobjdump -sass).  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipes pipes.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define ITERS 2048
#define CHAINS 8
#define REP 4

#define DEFK(name, DECL, BODY)                                                         \
  __global__ void __launch_bounds__(1024) name(int *out, int seed, long long *cyc) {    \
    int a[CHAINS], b[CHAINS];                                                          \
    _Pragma("unroll") for (int i = 0; i < CHAINS; i++) { a[i] = seed * (i + 3) + threadIdx.x; b[i] = seed ^ (i * 77 + threadIdx.x); } \
    DECL;                                                                              \
    long long t0 = clock64();                                                          \
    _Pragma("unroll 1") for (int it = 0; it < ITERS; it++) {                           \
      _Pragma("unroll") for (int r = 0; r < REP; r++) {                                \
        _Pragma("unroll") for (int i = 0; i < CHAINS; i++) { BODY; }                   \
      }                                                                                \
    }                                                                                  \
    long long t1 = clock64();                                                          \
    int acc = 0;                                                                       \
    _Pragma("unroll") for (int i = 0; i < CHAINS; i++) acc ^= a[i] ^ b[i];             \
    if (acc == 0x12345678) out[0] = acc;                                               \
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;                                   \
  }

DEFK(k_iadd3, , a[i] += b[i]; b[i] += a[i])
DEFK(k_lop3, , asm("lop3.b32 %0, %0, %1, %2, 0xCA;" : "+r"(a[i]) : "r"(b[i]), "r"(it)); asm("lop3.b32 %0, %0, %1, %2, 0xCA;" : "+r"(b[i]) : "r"(a[i]), "r"(it)))
DEFK(k_shf, , asm("shf.r.clamp.b32 %0, %0, %1, 3;" : "+r"(a[i]) : "r"(b[i])); asm("shf.r.clamp.b32 %0, %0, %1, 5;" : "+r"(b[i]) : "r"(a[i])))
DEFK(k_sar_add, , a[i] = (b[i] >> 1); b[i] = a[i] + it)
DEFK(k_prmt, , a[i] = __byte_perm(a[i], b[i], 0x7632); b[i] = __byte_perm(b[i], a[i], 0x1054))
DEFK(k_prmt_sext, , a[i] = __byte_perm(b[i], 0, 0x9910); b[i] = a[i] + it)
DEFK(k_imad, , a[i] = a[i] * b[i] + it; b[i] = b[i] * a[i] + it)
DEFK(k_imadhi, , a[i] = __mulhi(a[i], b[i]) + it; b[i] = __mulhi(b[i], a[i]) + it)
DEFK(k_imadwide_hi, , asm("{ .reg .s64 w; .reg .s32 lo; mul.wide.s32 w, %0, %1; mad.wide.s32 w, %1, %2, w; mov.b64 {lo, %0}, w; }" : "+r"(a[i]) : "r"(b[i]), "r"(it)); asm("{ .reg .s64 w; .reg .s32 lo; mul.wide.s32 w, %0, %1; mad.wide.s32 w, %1, %2, w; mov.b64 {lo, %0}, w; }" : "+r"(b[i]) : "r"(a[i]), "r"(it)))
DEFK(k_tw_lo, , a[i] = (a[i] * b[i] + b[i] * it) >> 16; b[i] = (b[i] * a[i] + a[i] * it) >> 16)
DEFK(k_viadd16, , a[i] = __vadd2(a[i], b[i]); b[i] = __vsub2(b[i], a[i]))
DEFK(k_vimnmx16, , a[i] = __vmaxs2(a[i], b[i]) ^ it; b[i] = __vmins2(b[i], a[i]) ^ it)
DEFK(k_vimnmx, , a[i] = max(a[i], b[i]) ^ it; b[i] = min(b[i], a[i]) ^ it)
DEFK(k_sat, , a[i] = min(__viaddmax_s32(a[i], b[i], -32768), 32767); b[i] = min(__viaddmax_s32(b[i], a[i], -32768), 32767) ^ it)
DEFK(k_i2f_f2i, float f = 1.0001f, { float t = (float)a[i]; t = t * f; b[i] ^= __float2int_rz(t); a[i] += b[i]; })
DEFK(k_mufu, , { float t = __int_as_float((a[i] & 0x007fffff) | 0x3f800000); asm("sqrt.approx.f32 %0, %0;" : "+f"(t)); b[i] ^= __float_as_int(t); a[i] += b[i]; })
DEFK(k_shfl, , a[i] = __shfl_xor_sync(0xffffffffu, a[i], 1) + b[i]; b[i] = __shfl_xor_sync(0xffffffffu, b[i], 2) + a[i])
DEFK(k_dp4a, , a[i] = __dp4a(a[i], b[i], it); b[i] = __dp4a(b[i], a[i], it))
DEFK(k_mix_add_imad, , a[i] += b[i]; b[i] = b[i] * a[i] + it)
DEFK(k_mix_shf_imad, , asm("shf.r.clamp.b32 %0, %0, %1, 3;" : "+r"(a[i]) : "r"(b[i])); b[i] = b[i] * a[i] + it)
DEFK(k_mix_viadd_imad, , a[i] = __vadd2(a[i], b[i]); b[i] = b[i] * a[i] + it)
DEFK(k_mix3, , a[i] += b[i]; b[i] = b[i] * a[i] + it; asm("shf.r.clamp.b32 %0, %0, %1, 3;" : "+r"(a[i]) : "r"(b[i])); b[i] = b[i] * a[i] + a[i])
DEFK(k_clz, , a[i] = __clz(a[i]) + b[i]; b[i] = __clz(b[i]) ^ a[i])
DEFK(k_lds, extern __shared__ int sm[]; sm[threadIdx.x] = seed; __syncthreads(), a[i] += sm[(a[i] + threadIdx.x) & 1023]; b[i] ^= a[i])

typedef void (*kern_t)(int *, int, long long *);
struct Entry { const char *name; kern_t k; int ops_per_body; };

int main() {
  cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
  printf("device %s SMs %d clock %d kHz\n", p.name, p.multiProcessorCount, p.clockRate);
  int *out; long long *cyc; cudaMalloc(&out, 64); cudaMalloc(&cyc, sizeof(long long) * 4096);
  Entry es[] = {
    {"IADD3 x2", k_iadd3, 2}, {"LOP3 x2", k_lop3, 2}, {"SHF x2", k_shf, 2}, {"SHF.S32+IADD", k_sar_add, 2}, {"PRMT x2", k_prmt, 2}, {"PRMT.sext+IADD", k_prmt_sext, 2},
    {"IMAD x2", k_imad, 2}, {"IMAD.HI x2", k_imadhi, 2}, {"(IMAD.WIDE+IMAD.HI) x2", k_imadwide_hi, 4},
    {"(IMAD,IMAD,SHF) x2", k_tw_lo, 6},
    {"VIADD.16x2 x2", k_viadd16, 2}, {"(VIMNMX.S16x2+LOP) x2", k_vimnmx16, 4}, {"(VIMNMX+LOP) x2", k_vimnmx, 4}, {"(VIADDMNMX+VIMNMX) x2 +LOP", k_sat, 5},
    {"I2F+FMUL+F2I+LOP+IADD", k_i2f_f2i, 5}, {"MUFU.SQRT +4 alu", k_mufu, 5}, {"(SHFL+IADD) x2", k_shfl, 4}, {"IDP4A x2", k_dp4a, 2},
    {"IADD+IMAD", k_mix_add_imad, 2}, {"SHF+IMAD", k_mix_shf_imad, 2}, {"VIADD16+IMAD", k_mix_viadd_imad, 2}, {"IADD,IMAD,SHF,IMAD", k_mix3, 4},
    {"(FLO+IADD/LOP) x2", k_clz, 4}, {"LDS random + 3 alu", k_lds, 4},
  };
  for (int nw : {4, 8, 16, 32}) {
    printf("--- %d warps/SM (1 CTA/SM)\n", nw);
    for (auto &e : es) {
      int grid = p.multiProcessorCount;
      e.k<<<grid, nw * 32, 4096>>>(out, 12345, cyc);
      cudaDeviceSynchronize();
      cudaEvent_t ea, eb; cudaEventCreate(&ea); cudaEventCreate(&eb);
      cudaEventRecord(ea);
      e.k<<<grid, nw * 32, 4096>>>(out, 12345, cyc);
      cudaEventRecord(eb); cudaEventSynchronize(eb);
      float ms; cudaEventElapsedTime(&ms, ea, eb);
      static long long h[4096]; cudaMemcpy(h, cyc, sizeof(long long) * grid, cudaMemcpyDeviceToHost);
      double mean = 0; for (int i = 0; i < grid; i++) mean += h[i]; mean /= grid;
      double winst = (double)ITERS * CHAINS * REP * e.ops_per_body * nw;
      printf("%-34s %9.0f cyc  %8.2f cyc/iter  %6.3f nominal-warp-inst/cyc/SM  (%.3f ms, %.0f MHz eff, err=%d)\n", e.name, mean, mean / ITERS, winst / mean, ms, mean / ms / 1e3, (int)cudaGetLastError());
    }
  }
  return 0;
}
