// ubench_stepx.cu -- scheduling lottery on the layout-8 step kernel (not product code): the product's
// step_kernel_x<8> body under different block sizes, generations unrolled per loop trip, launch bounds and
// orders of the (bit-position-wise, independent) register loops.  ptxas' schedule of the 531-instruction
// loop body moves the ALU-pipe utilisation by about a point between equivalent sources; this prints ms,
// Tcell/s and the fold of the output (all variants must agree) for each.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -I lifeapi_b200/csrc -o tools/ubench_stepx tools/ubench_stepx.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#include "life_multi.cuh"

using namespace lifeb200;

template <int ORDER>
__device__ __forceinline__ void stepv(TileX<8>& t, int sub) {
  using TX = TileX<8>;
  constexpr int M = 8, C = 8;
  const TX::Col first = t.count(0), last = t.count(C - 1);
  TX::Col prev = TX::halo(last, sub + 7);
  const TX::Col after = TX::halo(first, sub + 1);
  TX::Col c0 = first;
#pragma unroll
  for (int c = 0; c < C; c += 2) {
    const TX::Col c1 = (c + 1 == C - 1) ? last : t.count(c + 1);
    const TX::Col c2 = (c + 2 == C) ? after : (c + 2 == C - 1) ? last : t.count(c + 2);
#pragma unroll
    for (int kk = 0; kk < M; kk++) {
      const int k = (ORDER & 1) ? M - 1 - kk : kk;
      const PairCode p = pair_code(c0.b0[k], c0.b1[k], c1.b0[k], c1.b1[k]);
      if (ORDER & 2) {
        t.r[c + 1][k] = pair_select(p, c0.b0[k], c1.b0[k], c2.b0[k], c2.b1[k], t.r[c + 1][k]);
        t.r[c][k] = pair_select(p, c0.b0[k], c1.b0[k], prev.b0[k], prev.b1[k], t.r[c][k]);
      } else {
        t.r[c][k] = pair_select(p, c0.b0[k], c1.b0[k], prev.b0[k], prev.b1[k], t.r[c][k]);
        t.r[c + 1][k] = pair_select(p, c0.b0[k], c1.b0[k], c2.b0[k], c2.b1[k], t.r[c + 1][k]);
      }
    }
    prev = c1;
    c0 = c2;
  }
}

template <int BLOCK, int MINB, int UNROLL, int ORDER>
__global__ void __launch_bounds__(BLOCK, MINB) kx(const uint64_t* in, uint64_t* out, size_t n, unsigned gens) {
  constexpr int P = 4;
  const size_t gid = (size_t)blockIdx.x * BLOCK + threadIdx.x;
  const size_t p0 = (gid >> 3) * P;
  const int sub = threadIdx.x & 7;
  TileX<8> t;
#pragma unroll
  for (int c = 0; c < 8; c += 2) {
    uint64_t wa[P], wb[P];
#pragma unroll
    for (int s = 0; s < P; s++) {
      uint4 v = make_uint4(0, 0, 0, 0);
      if (p0 + s < n) v = *reinterpret_cast<const uint4*>(in + (p0 + s) * kN + sub * 8 + c);
      wa[s] = ((uint64_t)v.y << 32) | v.x;
      wb[s] = ((uint64_t)v.w << 32) | v.z;
    }
    t.pack(c, wa);
    t.pack(c + 1, wb);
  }
  if (ORDER == 4) {
#pragma unroll 1
    for (unsigned g = 0; g < gens; g++) t.step(sub);  // the product's own member
  } else {
#pragma unroll 1
    for (unsigned g = 0; g < gens; g += UNROLL) {
#pragma unroll
      for (int u = 0; u < UNROLL; u++) stepv<ORDER>(t, sub);
    }
  }
#pragma unroll
  for (int c = 0; c < 8; c += 2) {
    uint64_t wa[P], wb[P];
    t.unpack(c, wa);
    t.unpack(c + 1, wb);
#pragma unroll
    for (int s = 0; s < P; s++)
      if (p0 + s < n)
        *reinterpret_cast<uint4*>(out + (p0 + s) * kN + sub * 8 + c) =
            make_uint4((uint32_t)wa[s], (uint32_t)(wa[s] >> 32), (uint32_t)wb[s], (uint32_t)(wb[s] >> 32));
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

template <int BLOCK, int MINB, int UNROLL, int ORDER>
void run() {
  const size_t threads = (N + 3) / 4 * 8;
  const unsigned grid = (unsigned)((threads + BLOCK - 1) / BLOCK);
  cudaFuncAttributes fa;
  cudaFuncGetAttributes(&fa, kx<BLOCK, MINB, UNROLL, ORDER>);
  int occ = 0;
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kx<BLOCK, MINB, UNROLL, ORDER>, BLOCK, 0);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  float best = 1e30f, sum = 0;
  for (int r = 0; r < 5; r++) {
    cudaEventRecord(e0);
    kx<BLOCK, MINB, UNROLL, ORDER><<<grid, BLOCK>>>(d_in, d_out, N, GENS);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    if (r) { sum += ms; if (ms < best) best = ms; }
  }
  cudaMemset(d_acc, 0, 8);
  fold<<<(unsigned)((N * 64 + 255) / 256), 256>>>(d_out, N * 64, d_acc);
  unsigned long long h = 0;
  cudaMemcpy(&h, d_acc, 8, cudaMemcpyDeviceToHost);
  const double cells = (double)N * 4096.0 * GENS / (sum / 4 * 1e-3);
  printf("block %4d minb %d unroll %d order %d | regs %3d spill %zu warps/SM %2d | best %7.3f mean %7.3f ms %8.1f Gcell/s | fold %016llx %s\n",
         BLOCK, MINB, UNROLL, ORDER, fa.numRegs, (size_t)fa.localSizeBytes, occ * BLOCK / 32, best, sum / 4, cells / 1e9, h,
         cudaGetLastError() == cudaSuccess ? "" : "ERR");
  fflush(stdout);
}

int main() {
  cudaMalloc(&d_in, N * 512);
  cudaMalloc(&d_out, N * 512);
  cudaMalloc(&d_acc, 8);
  fill<<<(unsigned)((N * 64 + 255) / 256), 256>>>(d_in, N * 64);
  cudaDeviceSynchronize();
  //  BLOCK MINB UNROLL ORDER
  run<128, 1, 1, 4>();
  run<128, 1, 1, 0>();
  run<128, 1, 1, 1>();
  run<128, 1, 1, 2>();
  run<128, 1, 1, 3>();
  run<128, 2, 1, 0>();
  run<64, 1, 1, 0>();
  run<64, 4, 1, 0>();
  run<256, 1, 1, 0>();
  run<128, 1, 2, 0>();
  run<128, 1, 2, 1>();
  run<128, 1, 4, 0>();
  run<64, 1, 2, 0>();
  run<256, 1, 2, 0>();
  run<128, 3, 1, 0>();
  return 0;
}
