// ubench_rot.cu -- the last non-LOP3 integer-pipe instructions of the step loop are the 16 SHF rotates per lane and
// generation (layout 8: one rotate by 4 per column and direction).  Can they move to the FMA pipe?  A 32-bit rotate
// is exact as  w = x * 2^s (64-bit product: lo = x << s, hi = x >> (32 - s));  rot = hi + lo  -- an IMAD.WIDE and an
// IMAD, both on the FMA pipe, operands 2^s and 1 passed as kernel arguments so ptxas keeps the multiplies.
// V0 = product loop, V1 = both rotates as IMAD.WIDE + IMAD.  The two outputs are compared (this variant IS Life).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I lifeapi_b200/csrc -o tools/ubench_rot tools/ubench_rot.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#include "life_multi.cuh"

using namespace lifeb200;

__device__ __forceinline__ uint32_t rot_fma(uint32_t x, uint32_t pow2, uint32_t one) {
  uint32_t lo, hi, d;
  asm("{ .reg .u64 w; mul.wide.u32 w, %2, %3; mov.b64 {%0, %1}, w; }" : "=r"(lo), "=r"(hi) : "r"(x), "r"(pow2));
  asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(hi), "r"(one), "r"(lo));
  return d;
}

template <int V>
struct TX : TileX<8> {
  using Col = typename TileX<8>::Col;
  __device__ __forceinline__ Col count_v(int c, uint32_t one, uint32_t p4, uint32_t p28) const {
    Col o;
#pragma unroll
    for (int k = 0; k < 8; k++) {
      uint32_t up, dn;
      if (k == 0) up = V ? rot_fma(r[c][7], p4, one) : __funnelshift_l(r[c][7], r[c][7], 4);
      else up = r[c][k - 1];
      if (k == 7) dn = V ? rot_fma(r[c][0], p28, one) : __funnelshift_r(r[c][0], r[c][0], 4);
      else dn = r[c][k + 1];
      o.b0[k] = lop3<kXor3>(up, r[c][k], dn);
      o.b1[k] = lop3<kMaj>(up, r[c][k], dn);
    }
    return o;
  }
  __device__ __forceinline__ void step_v(int sub, uint32_t one, uint32_t p4, uint32_t p28) {
    const Col first = count_v(0, one, p4, p28), last = count_v(C - 1, one, p4, p28);
    Col prev = TileX<8>::halo(last, sub + 7);
    const Col after = TileX<8>::halo(first, sub + 1);
    Col c0 = first;
#pragma unroll
    for (int c = 0; c < C; c += 2) {
      const Col c1 = (c + 1 == C - 1) ? last : count_v(c + 1, one, p4, p28);
      const Col c2 = (c + 2 == C) ? after : (c + 2 == C - 1) ? last : count_v(c + 2, one, p4, p28);
#pragma unroll
      for (int k = 0; k < 8; k++) {
        const PairCode p = pair_code(c0.b0[k], c0.b1[k], c1.b0[k], c1.b1[k]);
        r[c][k] = pair_select(p, c0.b0[k], c1.b0[k], prev.b0[k], prev.b1[k], r[c][k]);
        r[c + 1][k] = pair_select(p, c0.b0[k], c1.b0[k], c2.b0[k], c2.b1[k], r[c + 1][k]);
      }
      prev = c1;
      c0 = c2;
    }
  }
};

template <int V>
__global__ void __launch_bounds__(128, 1) k(const uint64_t* __restrict__ in, uint64_t* __restrict__ out, size_t n,
                                            unsigned gens, uint32_t one, uint32_t p4, uint32_t p28) {
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
  for (unsigned g = 0; g < gens; g++) t.step_v(sub, one, p4, p28);
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
__global__ void diff(const uint64_t* a, const uint64_t* b, size_t words, unsigned long long* bad) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < words && a[i] != b[i]) atomicAdd(bad, 1ull);
}

static uint64_t *d_in, *d_out[2];
static const size_t N = 1 << 20;
static const unsigned GENS = 1024;

template <int V>
void run(const char* what) {
  const unsigned grid = (unsigned)(N / 4 * 8 / 128);
  cudaFuncAttributes fa;
  cudaFuncGetAttributes(&fa, k<V>);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  float best = 1e30f;
  for (int r = 0; r < 4; r++) {
    cudaEventRecord(e0);
    k<V><<<grid, 128>>>(d_in, d_out[V], N, GENS, 1u, 16u, 1u << 28);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    if (r && ms < best) best = ms;
  }
  const double cells = (double)N * 4096.0 * GENS / (best * 1e-3);
  printf("V %d %-52s regs %3d | %7.3f ms %8.1f Gcell/s %s\n", V, what, fa.numRegs, best, cells / 1e9,
         cudaGetLastError() == cudaSuccess ? "" : "ERR");
  fflush(stdout);
}

int main() {
  cudaMalloc(&d_in, N * 512);
  cudaMalloc(&d_out[0], N * 512);
  cudaMalloc(&d_out[1], N * 512);
  fill<<<(unsigned)((N * 64 + 255) / 256), 256>>>(d_in, N * 64);
  cudaDeviceSynchronize();
  run<0>("product loop (480 LOP3 + 16 SHF per lane-generation)");
  run<1>("rotates as IMAD.WIDE + IMAD (480 LOP3 + 32 FMA-pipe)");
  run<0>("product loop again");
  unsigned long long *d_bad, bad = 0;
  cudaMalloc(&d_bad, 8);
  cudaMemset(d_bad, 0, 8);
  diff<<<(unsigned)((N * 64 + 255) / 256), 256>>>(d_out[0], d_out[1], N * 64, d_bad);
  cudaMemcpy(&bad, d_bad, 8, cudaMemcpyDeviceToHost);
  printf("words that differ between the two variants after %u generations: %llu\n", GENS, bad);
  return bad ? 1 : 0;
}
