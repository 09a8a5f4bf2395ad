// ubench_fma.cu -- does moving gates of the step circuit from LOP3 (ALU pipe) to IMAD (FMA pipe) pay in the
// real loop?  (Not product code.)  TileX<8>::step with V of the 15 LUTs of a column pair replaced by
// `mad.lo.u32 d, a, one, b` (one = 1 from a kernel argument, so ptxas must keep an IMAD).  Variants > 0 are
// NOT Life any more -- this measures issue / pipe behaviour only, before a constrained-add circuit
// (tools/lut_synth_pair_add_cnf.py) is searched for.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I lifeapi_b200/csrc -o tools/ubench_fma tools/ubench_fma.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#include "life_multi.cuh"

using namespace lifeb200;

__device__ __forceinline__ uint32_t imad(uint32_t a, uint32_t one, uint32_t b) {
  uint32_t d;
  asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(one), "r"(b));
  return d;
}

// V bit 0: tail gate t2 as IMAD (2 per pair); bit 1: pair gate p2 as IMAD (1 per pair);
// bit 2: tail gate t0 as IMAD (2 per pair); bit 3: the last tail gate as IMAD (2 per pair)
template <int V>
struct TX : TileX<8> {
  using Col = typename TileX<8>::Col;
  __device__ __forceinline__ static uint32_t sel(const PairCode& p, uint32_t a0, uint32_t b0, uint32_t c0, uint32_t c1,
                                                 uint32_t cell, uint32_t one) {
    const uint32_t t0 = (V & 4) ? imad(c0, one, p.p2) : lop3<kMaj>(c0, p.p0, p.p2);
    const uint32_t t1 = lop3<0x16>(c1, p.p1, t0);
    const uint32_t t2 = (V & 1) ? imad(c0, one, a0) : lop3<kXor3>(c0, a0, b0);
    return (V & 8) ? imad(t1, one, t2) : lop3<0xC8>(cell, t1, t2);
  }
  __device__ __forceinline__ void step_v(int sub, uint32_t one) {
    const Col first = this->count(0), last = this->count(C - 1);
    Col prev = TileX<8>::halo(last, sub + 7);
    const Col after = TileX<8>::halo(first, sub + 1);
    Col c0 = first;
#pragma unroll
    for (int c = 0; c < C; c += 2) {
      const Col c1 = (c + 1 == C - 1) ? last : this->count(c + 1);
      const Col c2 = (c + 2 == C) ? after : (c + 2 == C - 1) ? last : this->count(c + 2);
#pragma unroll
      for (int k = 0; k < 8; k++) {
        PairCode p;
        p.p0 = lop3<0xB2>(c0.b1[k], c1.b0[k], c1.b1[k]);
        p.p1 = lop3<0xD4>(c0.b1[k], c1.b1[k], p.p0);
        p.p2 = (V & 2) ? imad(c0.b0[k], one, p.p0) : lop3<kMaj>(c0.b0[k], c1.b0[k], p.p0);
        r[c][k] = sel(p, c0.b0[k], c1.b0[k], prev.b0[k], prev.b1[k], r[c][k], one);
        r[c + 1][k] = sel(p, c0.b0[k], c1.b0[k], c2.b0[k], c2.b1[k], r[c + 1][k], one);
      }
      prev = c1;
      c0 = c2;
    }
  }
};

template <int V>
__global__ void __launch_bounds__(128, 1) k(const uint64_t* __restrict__ in, uint64_t* __restrict__ out, size_t n,
                                            unsigned gens, uint32_t one) {
  const size_t gid = (size_t)blockIdx.x * 128 + threadIdx.x;
  const size_t grp = gid >> 3;
  const int sub = threadIdx.x & 7;
  TX<V> t;
#pragma unroll
  for (int c = 0; c < 8; c++) {
    uint64_t w[4];
#pragma unroll
    for (int s = 0; s < 4; s++) {
      const size_t p = grp * 4 + s;
      w[s] = p < n ? in[p * 64 + sub * 8 + c] : 0;
    }
    t.pack(c, w);
  }
#pragma unroll 1
  for (unsigned g = 0; g < gens; g++) t.step_v(sub, one);
#pragma unroll
  for (int c = 0; c < 8; c++) {
    uint64_t w[4];
    t.unpack(c, w);
#pragma unroll
    for (int s = 0; s < 4; s++) {
      const size_t p = grp * 4 + s;
      if (p < n) out[p * 64 + sub * 8 + c] = w[s];
    }
  }
}

__global__ void fill(uint64_t* p, size_t words) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < words) p[i] = splitmix64(0x4C696665415049ULL + i);
}

static uint64_t *d_in, *d_out;
static const size_t N = 1 << 20;
static const unsigned GENS = 1024;

template <int V>
void run(const char* what, double lut_per_reg) {
  const unsigned grid = (unsigned)(N / 4 * 8 / 128);
  cudaFuncAttributes fa;
  cudaFuncGetAttributes(&fa, k<V>);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  float best = 1e30f;
  for (int r = 0; r < 4; r++) {
    cudaEventRecord(e0);
    k<V><<<grid, 128>>>(d_in, d_out, N, GENS, 1u);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    if (r && ms < best) best = ms;
  }
  const double cells = (double)N * 4096.0 * GENS / (best * 1e-3);
  const double p_int = 148 * 64 * 1.965e9;
  printf("V %2d %-44s regs %3d | %7.3f ms %8.1f Gcell/s | ALU instr/reg %.2f util %.4f %s\n", V, what, fa.numRegs, best,
         cells / 1e9, lut_per_reg, cells * (lut_per_reg / 32.0) / p_int, cudaGetLastError() == cudaSuccess ? "" : "ERR");
  fflush(stdout);
}

int main() {
  cudaMalloc(&d_in, N * 512);
  cudaMalloc(&d_out, N * 512);
  fill<<<(unsigned)((N * 64 + 255) / 256), 256>>>(d_in, N * 64);
  cudaDeviceSynchronize();
  run<0>("product circuit (15 LOP3 per pair)", 7.75);
  run<2>("p2 on FMA (14 LOP3 + 1 IMAD per pair)", 7.25);
  run<1>("t2 on FMA (13 LOP3 + 2 IMAD)", 6.75);
  run<3>("p2, t2 on FMA (12 LOP3 + 3 IMAD)", 6.25);
  run<5>("t0, t2 on FMA (11 LOP3 + 4 IMAD)", 5.75);
  run<8>("out on FMA (13 LOP3 + 2 IMAD)", 6.75);
  run<13>("t0, t2, out on FMA (9 LOP3 + 6 IMAD)", 4.75);
  return 0;
}
