// umma_probe.cu -- bring-up probe and rate microbenchmark for the Blackwell tensor path K2 uses:
//   tcgen05.mma.cta_group::1.kind::i8 (s8 x s8 -> s32, M = 128) with both operands read from shared memory
//   through NO-SWIZZLE K-major matrix descriptors, accumulators in TMEM, tcgen05.ld 32x32b read-back.
// Part 1 (correctness): a list of cases; each builds a shared-memory image on the host, issues a list of MMAs described
//   by (start, LBO, SBO) triples and compares TMEM with a host emulation of the descriptor addressing
//       A[m][k] = smem[a.start + (m / 8) * a.sbo + (m % 8) * 16 + (k / 16) * a.lbo + (k % 16)]      (B likewise over n)
//   -- the "implicit GEMM by descriptor strides" K2 relies on (conv taps = shifted start addresses, two taps per K = 32
//   through LBO, clip-minor activations so that 8 consecutive rows are 8 consecutive 16-byte units).
// Part 2 (rates): cycles per MMA at N = 16..256, tcgen05.ld throughput, and the whole-chip int8 rate (all SMs issuing
//   M128 N256 K32 MMAs back to back) = the measured int8 tensor peak that bench.py's cnn_roofline uses.
// nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o umma_probe umma_probe.cu && ./umma_probe [out.json]
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(2); } } while (0)

struct Mma {
  uint32_t a_start, a_lbo, a_sbo, b_start, b_lbo, b_sbo;   // bytes, relative to the image
  uint32_t n, tmem_col, accumulate, pad;
};
constexpr int kMaxMma = 64;
struct Case {
  uint32_t image_bytes, n_mma, out_cols, pad;
  Mma mma[kMaxMma];
};

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__host__ __device__ inline uint64_t make_desc(uint32_t addr, uint32_t lbo, uint32_t sbo) {
  // cute::UMMA::SmemDescriptor: start >> 4 [0,14), LBO >> 4 [16,30), SBO >> 4 [32,46), version 1 [46,48), layout_type 0 (no swizzle)
  return (uint64_t)((addr >> 4) & 0x3fff) | ((uint64_t)((lbo >> 4) & 0x3fff) << 16) | ((uint64_t)((sbo >> 4) & 0x3fff) << 32) |
         ((uint64_t)1 << 46);
}
__host__ __device__ inline uint32_t make_idesc_i8(uint32_t m, uint32_t n) {
  // cute::UMMA::InstrDescriptor: c_format S32 = 2 [4,6), a_format INT8 = 1 [7,10), b_format INT8 = 1 [10,13), K-major both,
  // n >> 3 [17,23), m >> 4 [24,29)
  return (2u << 4) | (1u << 7) | (1u << 10) | ((n >> 3) << 17) | ((m >> 4) << 24);
}

__device__ __forceinline__ void mma_i8(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t acc) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
               "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}\n"
               :: "r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" :: "r"(bar) : "memory");
}
__device__ __forceinline__ bool mbar_wait(uint32_t bar, uint32_t parity, long long max_spins = 1 << 22) {
  for (long long i = 0; i < max_spins; i++) {
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}\n"
                 : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    if (ok) return true;
  }
  return false;
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
                 "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
               : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// ---- part 1 -----------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128, 1) probe_kernel(const uint8_t *image, const Case *cs, int32_t *out, int *status) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t tmem_base_s;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const Case &c = *cs;
  for (uint32_t i = tid; i < c.image_bytes / 16; i += 128) reinterpret_cast<int4 *>(smem)[i] = reinterpret_cast<const int4 *>(image)[i];
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(&bar)), "r"(1));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" :: "r"(smem_u32(&tmem_base_s)), "r"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy smem writes -> visible to the MMA (async proxy)
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = tmem_base_s;
  if (tid == 0) {
    const uint32_t base = smem_u32(smem);
    for (uint32_t i = 0; i < c.n_mma; i++) {
      const Mma &m = c.mma[i];
      mma_i8(tmem + m.tmem_col, make_desc(base + m.a_start, m.a_lbo, m.a_sbo), make_desc(base + m.b_start, m.b_lbo, m.b_sbo),
             make_idesc_i8(128, m.n), m.accumulate);
    }
    umma_commit(smem_u32(&bar));
  }
  __syncwarp();
  const bool ok = mbar_wait(smem_u32(&bar), 0);
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  if (!ok) { if (tid == 0) *status = 1; }
  else {
    for (uint32_t col = 0; col < c.out_cols; col += 16) {
      uint32_t v[16];
      tmem_ld16(tmem + ((uint32_t)(warp * 32) << 16) + col, v);
      tmem_ld_wait();
      for (int j = 0; j < 16; j++) out[(warp * 32 + lane) * 512 + col + j] = (int32_t)v[j];
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(tmem), "r"(512));
}

// ---- part 2 -----------------------------------------------------------------------------------------------------------
// mode 0: one thread issues `iters` MMAs (M128, N = n, K32) into alternating TMEM halves, commit, wait; cycles of the whole train
// mode 1: four warps x `iters` tcgen05.ld 32x32b.x16 + wait
__global__ void __launch_bounds__(128, 1) rate_kernel(int mode, int n, int iters, long long *cycles, int *status) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t tmem_base_s;
  const int tid = threadIdx.x, warp = tid >> 5;
  for (int i = tid; i < 48 * 1024 / 4; i += 128) reinterpret_cast<uint32_t *>(smem)[i] = 0x01010101u * (uint32_t)(i & 3);
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(&bar)), "r"(1));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" :: "r"(smem_u32(&tmem_base_s)), "r"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = tmem_base_s;
  long long t0 = 0, t1 = 0;
  bool ok = true;
  const int chains = mode >> 8;
  mode &= 0xff;
  if (mode == 0) {
    if (tid == 0) {
      const uint32_t base = smem_u32(smem);
      const uint64_t da = make_desc(base, 2048, 128), db = make_desc(base + 8192, 4096, 128);
      const uint32_t idesc = make_idesc_i8(128, n);
      t0 = clock64();
      // `chains` independent accumulators (different TMEM columns), round robin: consecutive MMAs of one chain depend on each other
      const int stride = 512 / chains >= n ? 512 / chains : 0;   // chains * n <= 512, else all on one accumulator
      int ch = 0;
      for (int i = 0; i < iters; i++) {
        mma_i8(tmem + ch * stride, da, db, idesc, i >= chains);
        ch = ch + 1 == chains ? 0 : ch + 1;
      }
      umma_commit(smem_u32(&bar));
      ok = mbar_wait(smem_u32(&bar), 0, 1ll << 26);
      t1 = clock64();
    }
  } else {
    uint32_t acc = 0;
    __syncthreads();
    t0 = clock64();
    for (int i = 0; i < iters; i++) {
      uint32_t v[16];
      tmem_ld16(tmem + ((uint32_t)(warp * 32) << 16) + (i & 15) * 16, v);
      tmem_ld_wait();
      acc += v[0] ^ v[15];
    }
    t1 = clock64();
    if (acc == 0x12345678u) *status = 7;
  }
  if (tid == 0) {
    if (blockIdx.x == 0) *cycles = t1 - t0;
    if (!ok) *status = 1;
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(tmem), "r"(512));
}

// ---- host -------------------------------------------------------------------------------------------------------------
static uint32_t rng_state = 12345;
static int8_t rnd8() { rng_state = rng_state * 1664525u + 1013904223u; return (int8_t)(rng_state >> 24); }

static void emulate(const std::vector<uint8_t> &img, const Case &c, std::vector<int32_t> &D, std::vector<uint8_t> &written, bool swap_lbo_sbo) {
  D.assign(128 * 512, 0);
  written.assign(512, 0);
  for (uint32_t i = 0; i < c.n_mma; i++) {
    const Mma &m = c.mma[i];
    const uint32_t al = swap_lbo_sbo ? m.a_sbo : m.a_lbo, as = swap_lbo_sbo ? m.a_lbo : m.a_sbo;
    const uint32_t bl = swap_lbo_sbo ? m.b_sbo : m.b_lbo, bs = swap_lbo_sbo ? m.b_lbo : m.b_sbo;
    for (int r = 0; r < 128; r++)
      for (uint32_t n = 0; n < m.n; n++) {
        int32_t s = m.accumulate ? D[r * 512 + m.tmem_col + n] : 0;
        for (int k = 0; k < 32; k++) {
          const size_t aa = m.a_start + (r / 8) * as + (r % 8) * 16 + (k / 16) * al + (k % 16);
          const size_t ba = m.b_start + (n / 8) * bs + (n % 8) * 16 + (k / 16) * bl + (k % 16);
          const int a = aa < img.size() ? (int8_t)img[aa] : 0, b = ba < img.size() ? (int8_t)img[ba] : 0;
          s += a * b;
        }
        D[r * 512 + m.tmem_col + n] = s;
        written[m.tmem_col + n] = 1;
      }
  }
}

static int run_case(const char *name, const std::vector<uint8_t> &img, const Case &c, FILE *js) {
  uint8_t *d_img; Case *d_c; int32_t *d_out; int *d_status;
  CK(cudaMalloc(&d_img, img.size())); CK(cudaMalloc(&d_c, sizeof(Case))); CK(cudaMalloc(&d_out, 128 * 512 * 4)); CK(cudaMalloc(&d_status, 4));
  CK(cudaMemcpy(d_img, img.data(), img.size(), cudaMemcpyHostToDevice));
  CK(cudaMemcpy(d_c, &c, sizeof(Case), cudaMemcpyHostToDevice));
  CK(cudaMemset(d_out, 0xff, 128 * 512 * 4)); CK(cudaMemset(d_status, 0, 4));
  probe_kernel<<<1, 128, img.size()>>>(d_img, d_c, d_out, d_status);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("case %-28s CUDA ERROR %s\n", name, cudaGetErrorString(e)); if (js) fprintf(js, "\"%s\": \"cuda error\",\n", name); return -1; }
  std::vector<int32_t> out(128 * 512); int status;
  CK(cudaMemcpy(out.data(), d_out, out.size() * 4, cudaMemcpyDeviceToHost)); CK(cudaMemcpy(&status, d_status, 4, cudaMemcpyDeviceToHost));
  std::vector<int32_t> exp, exps;
  std::vector<uint8_t> wr;
  emulate(img, c, exp, wr, false); emulate(img, c, exps, wr, true);
  long bad = 0, bads = 0, nz = 0;
  for (int r = 0; r < 128; r++)
    for (uint32_t n = 0; n < c.out_cols; n++) {
      if (!wr[n]) continue;
      bad += out[r * 512 + n] != exp[r * 512 + n];
      bads += out[r * 512 + n] != exps[r * 512 + n];
      nz += exp[r * 512 + n] != 0;
    }
  printf("case %-28s status %d  mismatches %ld / %u (as documented)   %ld (LBO/SBO swapped)   nonzero expected %ld\n", name, status, bad,
         128 * c.out_cols, bads, nz);
  if (bad && bads) {
    for (int r : {0, 1, 8, 9, 33, 127}) {
      printf("   row %3d got:", r); for (int n = 0; n < 8; n++) printf(" %7d", out[r * 512 + n]);
      printf("\n           exp:"); for (int n = 0; n < 8; n++) printf(" %7d", exp[r * 512 + n]);
      printf("\n");
    }
  }
  if (js) fprintf(js, "\"%s\": {\"status\": %d, \"mismatch\": %ld, \"mismatch_swapped\": %ld, \"cells\": %u},\n", name, status, bad, bads, 128 * c.out_cols);
  cudaFree(d_img); cudaFree(d_c); cudaFree(d_out); cudaFree(d_status);
  return (int)bad;
}

int main(int argc, char **argv) {
  FILE *js = argc > 1 ? fopen(argv[1], "w") : nullptr;
  if (js) fprintf(js, "{\n");
  cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
  printf("device %s sm_%d%d, %d SMs\n", p.name, p.major, p.minor, p.multiProcessorCount);
  CK(cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  CK(cudaFuncSetAttribute(rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
  int total_bad = 0;
  {  // case 1: one dense MMA, A = [chunk][row][16 B], B = [chunk][n][16 B], N = 32
    std::vector<uint8_t> img(16384); for (auto &b : img) b = (uint8_t)rnd8();
    Case c{}; c.image_bytes = 16384; c.n_mma = 1; c.out_cols = 32;
    c.mma[0] = {0, 2048, 128, 8192, 512, 128, 32, 0, 0, 0};
    total_bad += run_case("dense_n32", img, c, js) != 0;
  }
  {  // case 2: same, N = 144 / 64 / 16 / 256 at different TMEM columns, then an accumulating second MMA on the N = 64 tile
    std::vector<uint8_t> img(32768); for (auto &b : img) b = (uint8_t)rnd8();
    Case c{}; c.image_bytes = 32768; c.n_mma = 5; c.out_cols = 512;
    c.mma[0] = {0, 2048, 128, 8192, 256 * 16, 128, 144, 0, 0, 0};
    c.mma[1] = {4096, 2048, 128, 8192 + 64, 256 * 16, 128, 64, 144, 0, 0};
    c.mma[2] = {0, 2048, 128, 16384, 256, 128, 16, 208, 0, 0};
    c.mma[3] = {0, 2048, 128, 8192, 256 * 16, 128, 256, 256, 0, 0};
    c.mma[4] = {1024, 2048, 128, 20480, 1024, 128, 64, 144, 1, 0};
    total_bad += run_case("n144_n64_n16_n256_acc", img, c, js) != 0;
  }
  {  // case 3: conv2-like.  activations [pixel][32 clips][16 B] (pixel stride 512 B), image width 9; one output tile = 4 pixels x 32 clips;
     // 9 taps in 5 MMAs: (ky, 0) + (ky, 1) via LBO = 512, ky = 0..2; then (0, 2) + (1, 2) via LBO = 9 * 512; then (2, 2) + zero weights
    const uint32_t act = 0, act_bytes = 48 * 512, wts = act_bytes, wts_bytes = 5 * 1024;
    std::vector<uint8_t> img(act_bytes + wts_bytes); for (auto &b : img) b = (uint8_t)rnd8();
    for (uint32_t i = 0; i < 512; i++) img[wts + 4 * 1024 + 512 + i] = 0;   // second chunk of the 5th weight block = zero
    Case c{}; c.image_bytes = (uint32_t)img.size(); c.out_cols = 32;
    const uint32_t q0 = 9 + 4;   // output pixels 13..16
    int k = 0;
    for (int ky = 0; ky < 3; ky++) c.mma[k] = {act + (q0 + ky * 9) * 512, 512, 128, wts + k * 1024, 512, 128, 32, 0, (uint32_t)(k > 0), 0}, k++;
    c.mma[k] = {act + (q0 + 2) * 512, 9 * 512, 128, wts + k * 1024, 512, 128, 32, 0, 1, 0}; k++;
    c.mma[k] = {act + (q0 + 18 + 2) * 512, 512, 128, wts + k * 1024, 512, 128, 32, 0, 1, 0}; k++;
    c.n_mma = k;
    total_bad += run_case("conv2_like_taps", img, c, js) != 0;
  }
  {  // case 4: conv1-like.  A start shifted by multiples of 512, LBO = 9216 (other-parity half), N = 144 twice sharing A
    std::vector<uint8_t> img(18432 + 6 * 4608); for (auto &b : img) b = (uint8_t)rnd8();
    Case c{}; c.image_bytes = (uint32_t)img.size(); c.out_cols = 288;
    for (int j = 0; j < 3; j++) {
      c.mma[2 * j] = {(uint32_t)(4 + j) * 512, 9216, 128, 18432u + (2 * j) * 4608, 144 * 16, 128, 144, 0, (uint32_t)(j > 0), 0};
      c.mma[2 * j + 1] = {(uint32_t)(4 + j) * 512, 9216, 128, 18432u + (2 * j + 1) * 4608, 144 * 16, 128, 144, 144, (uint32_t)(j > 0), 0};
    }
    c.n_mma = 6;
    total_bad += run_case("conv1_like_n144x2", img, c, js) != 0;
  }
  {  // case 5: starts that are 16-byte but not 128-byte aligned, odd SBO (144 B: image rows of 9 pixels x 16 B as one core matrix each)
    std::vector<uint8_t> img(32768); for (auto &b : img) b = (uint8_t)rnd8();
    Case c{}; c.image_bytes = 32768; c.n_mma = 2; c.out_cols = 32;
    c.mma[0] = {48, 16, 144, 16384 + 16, 512, 128, 32, 0, 0, 0};
    c.mma[1] = {48 + 32, 16 * 9, 144, 16384 + 1024 + 16, 512, 128, 32, 0, 1, 0};
    total_bad += run_case("unaligned16_sbo144", img, c, js) != 0;
  }
  // ---- rates
  long long *cyc; int *status; CK(cudaMallocManaged(&cyc, 8)); CK(cudaMallocManaged(&status, 4)); *status = 0;
  const int iters = 4096;
  for (int n : {16, 32, 48, 64, 80, 144, 256})
    for (int chains : {1, 2, 4, 8}) {
      if (chains * n > 512) continue;
      rate_kernel<<<1, 128, 48 * 1024>>>(0 | (chains << 8), n, iters, cyc, status); CK(cudaDeviceSynchronize());
      printf("mma i8 M128 N%-3d K32, %d independent accumulators: %.1f cycles/MMA (1 CTA)  status %d\n", n, chains, (double)*cyc / iters, *status);
      if (js) fprintf(js, "\"cyc_per_mma_n%d_chains%d\": %.2f,\n", n, chains, (double)*cyc / iters);
    }
  rate_kernel<<<1, 128, 48 * 1024>>>(1 | (1 << 8), 0, iters, cyc, status); CK(cudaDeviceSynchronize());
  printf("tcgen05.ld 32x32b.x16 + wait, 4 warps: %.1f cycles per load per warp\n", (double)*cyc / iters);
  if (js) fprintf(js, "\"cyc_per_ld_x16\": %.2f,\n", (double)*cyc / iters);
  {  // whole chip, N = 256
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    const int it2 = 1 << 16;
    rate_kernel<<<p.multiProcessorCount, 128, 48 * 1024>>>(0 | (2 << 8), 256, 1024, cyc, status); CK(cudaDeviceSynchronize());
    CK(cudaEventRecord(e0));
    rate_kernel<<<p.multiProcessorCount, 128, 48 * 1024>>>(0 | (2 << 8), 256, it2, cyc, status);
    CK(cudaEventRecord(e1)); CK(cudaDeviceSynchronize());
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
    const double ops = 2.0 * 128 * 256 * 32 * (double)it2 * p.multiProcessorCount;
    printf("whole chip int8 (tcgen05 cta_group::1, M128 N256 K32 from smem): %.1f TOP/s  (%.3f ms, %.1f cycles/MMA on CTA 0)  status %d\n",
           ops / ms / 1e9, ms, (double)*cyc / it2, *status);
    if (js) fprintf(js, "\"int8_tops_measured\": %.1f,\n", ops / ms / 1e9);
  }
  if (js) { fprintf(js, "\"cases_failed\": %d\n}\n", total_bad); fclose(js); }
  printf("%s\n", total_bad ? "PROBE: FAILURES" : "PROBE: all cases match the documented descriptor addressing");
  return 0;
}
