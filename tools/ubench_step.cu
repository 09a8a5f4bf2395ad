// ubench_step.cu -- scheduling experiments on the step kernel (not product code): the same Tile<LPP>::step
// under different lanes-per-pattern, block sizes, generations unrolled per loop trip and patterns
// interleaved per lane group.  Prints Tcell/s and the executed int-pipe utilisation per variant.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I lifeapi_b200/csrc -o tools/ubench_step tools/ubench_step.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#include "life_core.cuh"

using namespace lifeb200;

template <int LPP, int BLOCK, int UNROLL, int ILP, int MINB>
__global__ void __launch_bounds__(BLOCK, MINB) k(const uint64_t* __restrict__ in, uint64_t* __restrict__ out, size_t n,
                                                 unsigned gens) {
  const size_t gid = (size_t)blockIdx.x * BLOCK + threadIdx.x;
  const size_t grp = gid / LPP;
  const int sub = threadIdx.x & (LPP - 1);
  Tile<LPP> t[ILP];
#pragma unroll
  for (int i = 0; i < ILP; i++) {
    const size_t p = grp * ILP + i;
    if (p < n) t[i].load(in + p * kN, sub);
    else {
#pragma unroll
      for (int w = 0; w < Tile<LPP>::W; w++) t[i].lo[w] = t[i].hi[w] = 0;
    }
  }
#pragma unroll 1
  for (unsigned g = 0; g < gens; g += UNROLL) {
#pragma unroll
    for (int u = 0; u < UNROLL; u++) {
#pragma unroll
      for (int i = 0; i < ILP; i++) t[i].step(sub);
    }
  }
#pragma unroll
  for (int i = 0; i < ILP; i++) {
    const size_t p = grp * ILP + i;
    if (p < n) t[i].store(out + p * kN, sub);
  }
}

__global__ void fill(uint64_t* p, size_t words) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < words) p[i] = splitmix64(0x4C696665415049ULL + i);
}
__global__ void fold(const uint64_t* p, size_t words, unsigned long long* acc) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  unsigned long long v = i < words ? splitmix64(p[i] + i) : 0;
  for (int o = 16; o > 0; o >>= 1) v ^= __shfl_xor_sync(0xffffffffu, v, o);
  if ((threadIdx.x & 31) == 0) atomicXor(acc, v);
}

static uint64_t *d_in, *d_out;
static unsigned long long* d_acc;
static const size_t N = 1 << 20;
static const unsigned GENS = 1024;
static double p_int = 148 * 64 * 1.965e9;

template <int LPP, int BLOCK, int UNROLL, int ILP, int MINB>
void run() {
  const size_t threads = (N + ILP - 1) / ILP * LPP;
  const unsigned grid = (unsigned)((threads + BLOCK - 1) / BLOCK);
  cudaFuncAttributes fa;
  cudaFuncGetAttributes(&fa, k<LPP, BLOCK, UNROLL, ILP, MINB>);
  int occ = 0;
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k<LPP, BLOCK, UNROLL, ILP, MINB>, BLOCK, 0);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  float best = 1e30f;
  for (int r = 0; r < 4; r++) {
    cudaEventRecord(e0);
    k<LPP, BLOCK, UNROLL, ILP, MINB><<<grid, BLOCK>>>(d_in, d_out, N, GENS);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    if (r && ms < best) best = ms;
  }
  cudaMemset(d_acc, 0, 8);
  fold<<<(unsigned)((N * 64 + 255) / 256), 256>>>(d_out, N * 64, d_acc);
  unsigned long long h = 0;
  cudaMemcpy(&h, d_acc, 8, cudaMemcpyDeviceToHost);
  const double cells = (double)N * 4096.0 * GENS / (best * 1e-3);
  printf("lpp %2d block %4d unroll %d ilp %d minb %d | regs %3d warps/SM %2d | %7.3f ms %8.1f Gcell/s util %.4f | fold %016llx %s\n", LPP,
         BLOCK, UNROLL, ILP, MINB, fa.numRegs, occ * BLOCK / 32, best, cells / 1e9, cells * (9.5 / 32.0) / p_int, h,
         cudaGetLastError() == cudaSuccess ? "" : "ERR");
  fflush(stdout);
}

int main() {
  cudaMalloc(&d_in, N * 512);
  cudaMalloc(&d_out, N * 512);
  cudaMalloc(&d_acc, 8);
  fill<<<(unsigned)((N * 64 + 255) / 256), 256>>>(d_in, N * 64);
  cudaDeviceSynchronize();
  //   LPP BLOCK UNROLL ILP MINB
  run<4, 256, 1, 1, 1>();
  run<16, 256, 1, 1, 1>();
  run<16, 128, 1, 1, 1>();
  run<16, 512, 1, 1, 1>();
  run<16, 1024, 1, 1, 1>();
  run<16, 64, 1, 1, 1>();
  run<16, 256, 2, 1, 1>();
  run<16, 256, 4, 1, 1>();
  run<16, 256, 1, 2, 1>();
  run<16, 256, 2, 2, 1>();
  run<32, 256, 1, 1, 1>();
  run<32, 256, 2, 1, 1>();
  run<32, 256, 1, 2, 1>();
  run<32, 256, 1, 4, 1>();
  run<32, 512, 1, 2, 1>();
  run<8, 256, 1, 1, 1>();
  run<8, 256, 2, 1, 1>();
  run<8, 128, 1, 1, 1>();
  run<8, 512, 1, 1, 1>();
  run<8, 256, 1, 2, 1>();
  run<4, 256, 2, 1, 1>();
  run<4, 128, 1, 1, 1>();
  run<4, 512, 1, 1, 1>();
  run<4, 256, 1, 1, 5>();
  run<4, 256, 1, 1, 6>();
  run<2, 256, 1, 1, 1>();
  run<2, 128, 1, 1, 1>();
  return 0;
}
