// fp2.cu -- issue throughput of Blackwell's packed fp32 instructions (FADD2 / FFMA2, PTX *.f32x2) alone and mixed with the
// integer instructions of the fixed-point chain (sm_100a).  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp2 fp2.cu
// Prints warp-instructions per clock per SM (8 independent chains per thread, 512 threads x 4 CTAs per SM).
#include <cstdio>
#include <cuda_runtime.h>

#define ITERS 4096
#define REP8(X) X(0) X(1) X(2) X(3) X(4) X(5) X(6) X(7)
typedef unsigned long long u64;

template <int KIND>
__global__ void __launch_bounds__(512) k(int *out, int seed, int mul, float fs) {
  int v[8], w[8];
  u64 p[8], q[8];
  float f[8], g[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    v[i] = seed + threadIdx.x * (i + 1); w[i] = seed ^ (i * 77 + threadIdx.x);
    f[i] = fs * (i + 1 + threadIdx.x); g[i] = fs * 0.5f;
    asm("mov.b64 %0, {%1, %2};" : "=l"(p[i]) : "f"(f[i]), "f"(g[i]));
    asm("mov.b64 %0, {%1, %2};" : "=l"(q[i]) : "f"(g[i]), "f"(fs));
  }
#pragma unroll 1
  for (int it = 0; it < ITERS; ++it) {
#define IMAD_(i) asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(v[i]) : "r"(mul), "r"(w[i]));
#define PRMT_(i) asm volatile("prmt.b32 %0, %0, %1, 0x9910;" : "+r"(v[i]) : "r"(w[i]));
#define SHF_(i) asm volatile("shr.s32 %0, %0, %1;" : "+r"(v[i]) : "r"(mul));
#define FADD_(i) asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(f[i]) : "f"(g[i]));
#define FFMA_(i) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(f[i]) : "f"(g[i]), "f"(fs));
#define FFMARP_(i) asm volatile("fma.rp.f32 %0, %0, %1, %2;" : "+f"(f[i]) : "f"(g[i]), "f"(fs));
#define FADD2_(i) asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(p[i]) : "l"(q[i]));
#define FFMA2_(i) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p[i]) : "l"(q[i]), "l"(q[(i + 1) & 7]));
#define FFMA2RP_(i) asm volatile("fma.rp.f32x2 %0, %0, %1, %2;" : "+l"(p[i]) : "l"(q[i]), "l"(q[(i + 1) & 7]));
    if (KIND == 0) { REP8(FADD_) }
    if (KIND == 1) { REP8(FFMA_) }
    if (KIND == 2) { REP8(FFMARP_) }
    if (KIND == 3) { REP8(FADD2_) }
    if (KIND == 4) { REP8(FFMA2_) }
    if (KIND == 5) { REP8(FFMA2RP_) }
    if (KIND == 6) { REP8(FADD2_) REP8(PRMT_) }
    if (KIND == 7) { REP8(FADD2_) REP8(IMAD_) }
    if (KIND == 8) { REP8(FADD_) REP8(IMAD_) }
    if (KIND == 9) { REP8(FADD2_) REP8(PRMT_) REP8(FFMA2RP_) REP8(SHF_) }
    if (KIND == 10) { REP8(FADD2_) REP8(IMAD_) REP8(PRMT_) }
    if (KIND == 11) { REP8(FADD_) REP8(PRMT_) }
    if (KIND == 12) { REP8(IMAD_) }
    if (KIND == 13) { REP8(FADD2_) REP8(FFMA2RP_) REP8(FADD2_) REP8(IMAD_) REP8(PRMT_) REP8(SHF_) }
  }
  int acc = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) { float a, b; asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(p[i])); acc += v[i] + (int)f[i] + (int)a + (int)b; }
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

template <int KIND> void run(const char *name, int per_iter, int *d, int sms, int clk_khz) {
  const int grid = sms * 4;
  k<KIND><<<grid, 512>>>(d, 1, 3, 1.0f);
  cudaDeviceSynchronize();
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  cudaEventRecord(a);
  k<KIND><<<grid, 512>>>(d, 1, 3, 1.0f);
  cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b);
  const double warp_insts = (double)grid * 16 * ITERS * per_iter;
  const double clocks = ms * 1e-3 * clk_khz * 1e3;
  printf("%-52s %8.3f ms  %6.3f warp-inst/clk/SM\n", name, ms, warp_insts / clocks / sms);
}

int main() {
  cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
  int clk = 0; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
  printf("%s, %d SMs, %d kHz (nominal; results assume this clock)\n", p.name, p.multiProcessorCount, clk);
  int *d; cudaMalloc(&d, 148 * 8 * 512 * 4);
  const int s = p.multiProcessorCount;
  run<0>("add.f32 (FADD)", 8, d, s, clk);
  run<1>("fma.rn.f32 (FFMA)", 8, d, s, clk);
  run<2>("fma.rp.f32 (FFMA.RP)", 8, d, s, clk);
  run<3>("add.f32x2 (FADD2)", 8, d, s, clk);
  run<4>("fma.rn.f32x2 (FFMA2)", 8, d, s, clk);
  run<5>("fma.rp.f32x2 (FFMA2.RP)", 8, d, s, clk);
  run<12>("mad.lo (IMAD)", 8, d, s, clk);
  run<6>("FADD2 + PRMT 1:1", 16, d, s, clk);
  run<11>("FADD + PRMT 1:1", 16, d, s, clk);
  run<7>("FADD2 + IMAD 1:1", 16, d, s, clk);
  run<8>("FADD + IMAD 1:1", 16, d, s, clk);
  run<10>("FADD2 + IMAD + PRMT 1:1:1", 24, d, s, clk);
  run<9>("FADD2 + PRMT + FFMA2.RP + SHF", 32, d, s, clk);
  run<13>("FADD2 FFMA2.RP FADD2 IMAD PRMT SHF", 48, d, s, clk);
  return 0;
}
