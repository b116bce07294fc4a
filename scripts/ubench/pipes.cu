// pipes.cu -- issue-throughput microbenchmarks of the integer instructions the fixed-point chain is made of (sm_100a).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipes pipes.cu ; run on the GPU box: ./pipes
// Prints warp-instructions per clock per SM for each instruction mix (8 independent chains per thread, 1024 threads x 2 CTAs/SM).
#include <cstdio>
#include <cuda_runtime.h>

#define ITERS 4096
#define REP8(X) X(0) X(1) X(2) X(3) X(4) X(5) X(6) X(7)

template <int KIND>
__global__ void __launch_bounds__(512) k(int *out, int seed, int mul, int sh) {
  int v[8], w[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) { v[i] = seed + threadIdx.x * (i + 1); w[i] = seed ^ (i * 77 + threadIdx.x); }
  __shared__ int sm[2048];
  sm[threadIdx.x] = seed; sm[threadIdx.x + 512] = seed; sm[threadIdx.x + 1024] = seed; sm[threadIdx.x + 1536] = seed;
  __syncthreads();
  const int sa = (int)__cvta_generic_to_shared(&sm[threadIdx.x]);
#pragma unroll 1
  for (int it = 0; it < ITERS; ++it) {
#define IADD3_(i) asm volatile("add.s32 %0, %0, %1;" : "+r"(v[i]) : "r"(w[i]));
#define IMAD_(i) asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(v[i]) : "r"(mul), "r"(w[i]));
#define IMADHI_(i) asm volatile("mad.hi.s32 %0, %0, %1, %2;" : "+r"(v[i]) : "r"(mul), "r"(w[i]));
#define MULHI_(i) asm volatile("mul.hi.s32 %0, %0, %1;" : "+r"(v[i]) : "r"(mul));
#define SHF_(i) asm volatile("shr.s32 %0, %0, %1;" : "+r"(v[i]) : "r"(sh));
#define PRMT_(i) asm volatile("prmt.b32 %0, %0, %1, 0x9910;" : "+r"(v[i]) : "r"(w[i]));
#define LOP_(i) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(v[i]) : "r"(w[i]), "r"(mul));
#define VADD2_(i) asm volatile("{.reg .b32 t; vadd2.s32.s32.s32 t, %0, %1, %2; mov.b32 %0, t;}" : "+r"(v[i]) : "r"(w[i]), "r"(0));
#define MNMX_(i) asm volatile("max.s32 %0, %0, %1;" : "+r"(v[i]) : "r"(w[i]));
#define LDS_(i) asm volatile("ld.shared.b32 %0, [%1+" #i "*128];" : "=r"(v[i]) : "r"(sa));
#define LDS16_(i) asm volatile("ld.shared.s16 %0, [%1+" #i "*128];" : "=r"(v[i]) : "r"(sa));
#define STS_(i) asm volatile("st.shared.b32 [%1+" #i "*128], %0;" :: "r"(v[i]), "r"(sa));
#define STS16_(i) asm volatile("st.shared.b16 [%1+" #i "*128], %0;" :: "h"((short)v[i]), "r"(sa));
#define VAMR_(i) v[i] = __viaddmin_s16x2_relu((unsigned)v[i], (unsigned)w[i], 0x00010001u);
#define VMX2_(i) v[i] = __vmaxs2((unsigned)v[i], (unsigned)w[i]);
#define LOPP_(i) asm volatile("{.reg .pred p, q; .reg .b32 d; setp.ne.u32 q, %0, %0; lop3.or.b32 d|p, %0, %1, 0xffff, 0x28, q; @!p st.shared.u16 [%2+" #i "*128], %3;}" :: "r"(v[i]), "r"(w[i]), "r"(sa), "h"((short)mul) : "memory");
    if (KIND == 19) { REP8(VAMR_) }
    if (KIND == 20) { REP8(VMX2_) }
    if (KIND == 21) { REP8(VAMR_) REP8(IMAD_) }
    if (KIND == 22) { REP8(LOPP_) REP8(IMAD_) }                                   // test + (almost never taken) predicated store, + mad
    if (KIND == 23) { REP8(LOPP_) REP8(VAMR_) REP8(IMAD_) REP8(VAMR_) REP8(IMAD_) } // the packed OS-CFAR step: 2 alu + 2 alu... per pair
    if (KIND == 0) { REP8(IADD3_) }
    if (KIND == 1) { REP8(IMAD_) }
    if (KIND == 2) { REP8(IMADHI_) }
    if (KIND == 3) { REP8(SHF_) }
    if (KIND == 4) { REP8(PRMT_) }
    if (KIND == 5) { REP8(LOP_) }
    if (KIND == 6) { REP8(IADD3_) REP8(IMAD_) }          // 1:1 alu:fma
    if (KIND == 7) { REP8(SHF_) REP8(IMAD_) }
    if (KIND == 8) { REP8(PRMT_) REP8(IMAD_) REP8(SHF_) REP8(IMAD_) }
    if (KIND == 9) { REP8(MULHI_) }
    if (KIND == 10) { REP8(MNMX_) }
    if (KIND == 11) { REP8(LDS_) }
    if (KIND == 12) { REP8(LDS16_) }
    if (KIND == 13) { REP8(STS_) }
    if (KIND == 14) { REP8(STS16_) }
    if (KIND == 15) { REP8(LDS_) REP8(IADD3_) REP8(IMAD_) REP8(SHF_) REP8(IMAD_) }   // 1 lsu : 2 alu : 2 fma
    if (KIND == 16) { REP8(VADD2_) }
    if (KIND == 17) { REP8(SHF_) REP8(IMADHI_) }
    if (KIND == 18) { REP8(LDS_) REP8(STS_) REP8(IADD3_) REP8(IMAD_) REP8(SHF_) REP8(IMAD_) REP8(PRMT_) REP8(IMAD_) }  // 2 lsu : 3 alu : 3 fma
  }
  int acc = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) acc += v[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

template <int KIND> void run(const char *name, int per_iter, int *d, int sms, int clk_khz) {
  const int grid = sms * 4;
  k<KIND><<<grid, 512>>>(d, 1, 3, 1);
  cudaDeviceSynchronize();
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  cudaEventRecord(a);
  k<KIND><<<grid, 512>>>(d, 1, 3, 1);
  cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b);
  const double warp_insts = (double)grid * 16 * ITERS * per_iter;
  const double clocks = ms * 1e-3 * clk_khz * 1e3;
  printf("%-44s %8.3f ms  %6.3f warp-inst/clk/SM\n", name, ms, warp_insts / clocks / sms);
}

int main() {
  cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
  int clk = 0; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
  printf("%s, %d SMs, %d kHz (nominal; results assume this clock)\n", p.name, p.multiProcessorCount, clk);
  int *d; cudaMalloc(&d, 148 * 8 * 512 * 4);
  const int s = p.multiProcessorCount;
  run<0>("add (IADD3)", 8, d, s, clk);
  run<1>("mad.lo (IMAD)", 8, d, s, clk);
  run<2>("mad.hi (IMAD.HI)", 8, d, s, clk);
  run<9>("mul.hi", 8, d, s, clk);
  run<3>("shr (SHF)", 8, d, s, clk);
  run<4>("prmt (PRMT)", 8, d, s, clk);
  run<5>("lop3 (LOP3)", 8, d, s, clk);
  run<10>("max (VIMNMX)", 8, d, s, clk);
  run<16>("vadd2 (VIADD.16x2)", 8, d, s, clk);
  run<6>("add + mad.lo 1:1", 16, d, s, clk);
  run<7>("shr + mad.lo 1:1", 16, d, s, clk);
  run<8>("prmt + mad + shr + mad", 32, d, s, clk);
  run<17>("shr + mad.hi 1:1", 16, d, s, clk);
  run<11>("ld.shared.b32", 8, d, s, clk);
  run<12>("ld.shared.s16", 8, d, s, clk);
  run<13>("st.shared.b32", 8, d, s, clk);
  run<14>("st.shared.b16", 8, d, s, clk);
  run<15>("lds + add + mad + shr + mad (1:2:2)", 40, d, s, clk);
  run<18>("lds sts + 3 alu + 3 fma (2:3:3)", 64, d, s, clk);
  run<19>("viaddmin.s16x2.relu (VIADDMNMX.S16x2.RELU)", 8, d, s, clk);
  run<20>("vimax.s16x2 (VIMNMX.S16x2)", 8, d, s, clk);
  run<21>("viaddmin.s16x2.relu + mad.lo 1:1", 16, d, s, clk);
  run<22>("lop3 -> pred + @p sts.u16, + mad.lo", 24, d, s, clk);
  run<23>("lop3/sts + 2 x (viaddmnmx + mad)", 48, d, s, clk);
  return 0;
}
