// ubench_pipes.cu -- measure sm_100a issue rates of the integer instructions this engine can use,
// alone and co-issued (ALU pipe: LOP3/SHF/IADD3/PRMT; FMA pipe: IMAD, IMAD.HI, IMAD.WIDE).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ubench_pipes ubench_pipes.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

constexpr int CH = 8;       // independent chains per op kind
constexpr int INNER = 32;   // unrolled inner repetitions

template <int NL, int NS, int NM, int NH, int NW, int NP, int NA>
__global__ void __launch_bounds__(1024) k(uint32_t* sink, unsigned iters, uint32_t m1, uint32_t m2) {
  uint32_t a[CH], s[CH], m[CH], h[CH], p[CH], d[CH];
  uint64_t w[CH];
#pragma unroll
  for (int i = 0; i < CH; i++) {
    a[i] = threadIdx.x * 2654435761u + i; s[i] = a[i] ^ 0x1234567u; m[i] = a[i] + 77u * i; h[i] = ~a[i];
    w[i] = ((uint64_t)a[i] << 32) | s[i]; p[i] = a[i] * 3u; d[i] = a[i] * 5u;
  }
#pragma unroll 1
  for (unsigned it = 0; it < iters; it++) {
#pragma unroll
    for (int j = 0; j < INNER; j++) {
#pragma unroll
      for (int i = 0; i < CH; i++) {
        if (i < NL) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[i]) : "r"(a[(i + 1) % CH]), "r"(m1));
        if (i < NS) asm volatile("shf.l.wrap.b32 %0, %0, %1, 1;" : "+r"(s[i]) : "r"(s[(i + 1) % CH]));
        if (i < NM) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(m[i]) : "r"(m1), "r"(m[(i + 1) % CH]));
        if (i < NH) asm volatile("mad.hi.u32 %0, %0, %1, %2;" : "+r"(h[i]) : "r"(m2), "r"(h[(i + 1) % CH]));
        if (i < NW) asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(w[i]) : "r"((uint32_t)w[(i + 1) % CH]), "r"(m2));
        if (i < NP) asm volatile("prmt.b32 %0, %0, %1, 0x3210;" : "+r"(p[i]) : "r"(p[(i + 1) % CH]));
        if (i < NA) asm volatile("add.u32 %0, %0, %1;" : "+r"(d[i]) : "r"(d[(i + 1) % CH]));
      }
    }
  }
  uint32_t acc = 0;
#pragma unroll
  for (int i = 0; i < CH; i++) acc ^= a[i] ^ s[i] ^ m[i] ^ h[i] ^ (uint32_t)w[i] ^ (uint32_t)(w[i] >> 32) ^ p[i] ^ d[i];
  if (acc == 0x12345u) sink[0] = acc;
}

template <int NL, int NS, int NM, int NH, int NW, int NP, int NA>
void run(const char* name, int sms, double mhz) {
  uint32_t* sink;
  cudaMalloc(&sink, 4);
  const unsigned iters = 512;
  const int blocks = sms * 2, threads = 1024;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  float best = 1e30f;
  for (int r = 0; r < 4; r++) {
    cudaEventRecord(e0);
    k<NL, NS, NM, NH, NW, NP, NA><<<blocks, threads>>>(sink, iters, 2u, 0x80000000u);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    if (r && ms < best) best = ms;
  }
  double per_kind = (double)blocks * threads / 32.0 * iters * INNER;  // warp-instr per chain slot
  double total = per_kind * (NL + NS + NM + NH + NW + NP + NA);
  double cyc = best * 1e-3 * mhz * 1e6;  // SM cycles at the assumed clock
  printf("%-34s %8.3f ms  warp-instr/clk/SM total %.3f  (lop3 %.2f shf %.2f imad %.2f imad.hi %.2f imad.wide %.2f prmt %.2f iadd %.2f)\n",
         name, best, total / cyc / sms, per_kind * NL / cyc / sms, per_kind * NS / cyc / sms, per_kind * NM / cyc / sms,
         per_kind * NH / cyc / sms, per_kind * NW / cyc / sms, per_kind * NP / cyc / sms, per_kind * NA / cyc / sms);
  cudaFree(sink);
}

int main() {
  cudaDeviceProp prop;
  cudaGetDeviceProperties(&prop, 0);
  int sms = prop.multiProcessorCount;
  double mhz = 1965.0;
  printf("%s, %d SMs, assuming %.0f MHz (check clocks)\n", prop.name, sms, mhz);
  //   NL NS NM NH NW NP NA
  run<8, 0, 0, 0, 0, 0, 0>("lop3 x8", sms, mhz);
  run<0, 8, 0, 0, 0, 0, 0>("shf x8", sms, mhz);
  run<0, 0, 8, 0, 0, 0, 0>("imad x8", sms, mhz);
  run<0, 0, 0, 8, 0, 0, 0>("imad.hi x8", sms, mhz);
  run<0, 0, 0, 0, 8, 0, 0>("imad.wide x8", sms, mhz);
  run<0, 0, 0, 0, 0, 8, 0>("prmt x8", sms, mhz);
  run<0, 0, 0, 0, 0, 0, 8>("iadd x8", sms, mhz);
  run<8, 0, 8, 0, 0, 0, 0>("lop3 x8 + imad x8", sms, mhz);
  run<8, 0, 4, 0, 0, 0, 0>("lop3 x8 + imad x4", sms, mhz);
  run<8, 0, 2, 0, 0, 0, 0>("lop3 x8 + imad x2", sms, mhz);
  run<8, 0, 0, 4, 0, 0, 0>("lop3 x8 + imad.hi x4", sms, mhz);
  run<8, 0, 0, 2, 0, 0, 0>("lop3 x8 + imad.hi x2", sms, mhz);
  run<8, 0, 0, 0, 4, 0, 0>("lop3 x8 + imad.wide x4", sms, mhz);
  run<8, 0, 0, 0, 2, 0, 0>("lop3 x8 + imad.wide x2", sms, mhz);
  run<8, 0, 2, 2, 0, 0, 0>("lop3 x8 + imad x2 + imad.hi x2", sms, mhz);
  run<8, 4, 0, 0, 0, 0, 0>("lop3 x8 + shf x4", sms, mhz);
  run<8, 0, 0, 0, 0, 0, 4>("lop3 x8 + iadd x4", sms, mhz);
  return 0;
}
