// life_multi.cuh -- the row-interleaved layout of life_core.cuh taken further: P patterns per lane group.
//
// A column of 64 rows is spread over M registers by row class: register k holds the rows r with
// r mod M == k.  That is 64/M rows per register, so P = M/2 patterns share the register, bit-interleaved:
//     register k, bit j*P + s  =  row M*j + k of pattern s          (j < 64/M, s < P)
// Row neighbours then live in the NEXT / PREVIOUS register at the SAME bit, except across the class
// wrap: the row above class 0 is class M-1 one field down, i.e. that register rotated left by P bits
// (one 32-bit rotate moves the fields of all P patterns at once, and its wrap is the torus wrap
// 63 -> 0); the row below class M-1 is class 0 rotated right by P.  So a generation needs 2 SHF per M
// registers instead of 2 per register (reference layout) or 1 per register (M = 2, life_core.cuh):
//     M = 2: 8.5   M = 4: 8.0   M = 8: 7.75   integer-pipe instructions per 32 cells and generation.
// Everything after the column count (pair code, select) is bit-position-wise and is the same code.
// A group of 8 lanes owns P patterns; each lane keeps 8 columns of all of them (8*M registers).
#pragma once
#include "life_core.cuh"

namespace lifeb200 {

// Exchange a register-index bit with a position-index bit (a delta swap across a register pair): the bits of
// `a` at positions with that index bit set trade places with the bits of `b` at positions with it clear
// (m = positions with the bit clear, d = its weight).  2 SHF + 3 LOP3.
__device__ __forceinline__ void exchange_bits(uint32_t& a, uint32_t& b, int d, uint32_t m) {
  const uint32_t t = ((a >> d) ^ b) & m;
  b ^= t;
  a ^= t << d;
}
template <int M>
struct TileX {
  static_assert(M == 4 || M == 8, "row classes per column");
  static constexpr int C = 8;      // columns per lane (8 lanes per group)
  static constexpr int P = M / 2;  // patterns per group
  uint32_t r[C][M];

  struct Col {
    uint32_t b0[M], b1[M];
  };
  __device__ __forceinline__ Col count(int c) const {
    Col o;
#pragma unroll
    for (int k = 0; k < M; k++) {
      const uint32_t up = k == 0 ? __funnelshift_l(r[c][M - 1], r[c][M - 1], P) : r[c][k - 1];
      const uint32_t dn = k == M - 1 ? __funnelshift_r(r[c][0], r[c][0], P) : r[c][k + 1];
      o.b0[k] = lop3<kXor3>(up, r[c][k], dn);
      o.b1[k] = lop3<kMaj>(up, r[c][k], dn);
    }
    return o;
  }
  __device__ __forceinline__ static Col halo(const Col& c, int src) {
    Col o;
#pragma unroll
    for (int k = 0; k < M; k++) {
      o.b0[k] = __shfl_sync(0xffffffffu, c.b0[k], src, 8);
      o.b1[k] = __shfl_sync(0xffffffffu, c.b1[k], src, 8);
    }
    return o;
  }

  // One generation of all P patterns in place (== LifeState::Step() on each, LifeAPI.hpp:1196-1216).
  __device__ __forceinline__ void step(int sub) {
    const Col first = count(0), last = count(C - 1);
    Col prev = halo(last, sub + 7);   // column to the left of this lane's first (mod 8 by the shuffle width)
    const Col after = halo(first, sub + 1);
    Col c0 = first;
#pragma unroll
    for (int c = 0; c < C; c += 2) {
      const Col c1 = (c + 1 == C - 1) ? last : count(c + 1);
      const Col c2 = (c + 2 == C) ? after : (c + 2 == C - 1) ? last : count(c + 2);
#pragma unroll
      for (int k = 0; k < M; k++) {
        const PairCode p = pair_code(c0.b0[k], c0.b1[k], c1.b0[k], c1.b1[k]);
        r[c][k] = pair_select(p, c0.b0[k], c1.b0[k], prev.b0[k], prev.b1[k], r[c][k]);
        r[c + 1][k] = pair_select(p, c0.b0[k], c1.b0[k], c2.b0[k], c2.b1[k], r[c + 1][k]);
      }
      prev = c1;
      c0 = c2;
    }
  }

  // words w[s] = column c of pattern s in the reference layout  ->  r[c][0..M-1]
  __device__ __forceinline__ void pack(int c, const uint64_t (&w)[P]) {
    if constexpr (M == 8) {
      // A permutation of index bits: the 8 input registers are numbered (s1 s0 y5) and hold position (y4 .. y0);
      // the output registers are numbered (y2 y1 y0) and hold position (y5 y4 y3 s1 s0).  Four register <->
      // position exchanges (s1 <-> y1, s0 <-> y0, then the register bit that starts as y5 trades places with
      // position bit 3 and then with position bit 2: it walks y5 -> y3 -> y2 while the position becomes
      // (y4 y5 y3 s1 s0)) and one byte swap (position bits 4 <-> 3) per register.  88 integer instructions per
      // column (11 per register; 100 with three exchanges + a nibble swap + the byte swap, round 1) where the
      // 8x8 transposes + byte gathers + two perfect shuffles per register took about 400.
      uint32_t x[P][2];
#pragma unroll
      for (int s = 0; s < P; s++) { x[s][0] = (uint32_t)w[s]; x[s][1] = (uint32_t)(w[s] >> 32); }
#pragma unroll
      for (int h = 0; h < 2; h++) {
        exchange_bits(x[0][h], x[2][h], 2, 0x33333333u);  // s1 <-> y1
        exchange_bits(x[1][h], x[3][h], 2, 0x33333333u);
        exchange_bits(x[0][h], x[1][h], 1, 0x55555555u);  // s0 <-> y0
        exchange_bits(x[2][h], x[3][h], 1, 0x55555555u);
      }
#pragma unroll
      for (int s = 0; s < P; s++) {
        exchange_bits(x[s][0], x[s][1], 8, 0x00FF00FFu);  // y5 <-> y3: position (y4 y5 y2 s1 s0), register bit h = y3
        exchange_bits(x[s][0], x[s][1], 4, 0x0F0F0F0Fu);  // y3 <-> y2: position (y4 y5 y3 s1 s0), register bit h = y2
      }
#pragma unroll
      for (int h = 0; h < 2; h++)
#pragma unroll
        for (int s = 0; s < P; s++) r[c][4 * h + s] = __byte_perm(x[s][h], 0, 0x3120);  // bytes 1 <-> 2: (y5 y4 y3 s1 s0)
    } else {
      // M = 4, P = 2: rows 4j+k.  Per pattern: unshuffle twice brings bit 4j+k of each 32-bit half to
      // bit 8k + j (byte k); byte k of the low half = rows 4j+k for j < 8, of the high half for j >= 8.
      uint32_t ul[P], uh[P];
#pragma unroll
      for (int s = 0; s < P; s++) {
        ul[s] = unshuffle32(unshuffle32((uint32_t)w[s]));
        uh[s] = unshuffle32(unshuffle32((uint32_t)(w[s] >> 32)));
      }
#pragma unroll
      for (int k = 0; k < M; k++) {
        // 16 rows of pattern s: low byte from ul, high byte from uh -> halfword; patterns side by side
        const uint32_t sel = (uint32_t)k | ((uint32_t)(4 + k) << 4);
        const uint32_t a = __byte_perm(ul[0], uh[0], sel);  // pattern 0: bit j = row 4j+k, j < 16
        const uint32_t b = __byte_perm(ul[1], uh[1], sel);
        r[c][k] = shuffle32(__byte_perm(a, b, 0x5410));     // bit 16s + j -> bit 2j + s
      }
    }
  }
  __device__ __forceinline__ void unpack(int c, uint64_t (&w)[P]) const {
    if constexpr (M == 8) {
      uint32_t x[P][2];  // pack() run backwards
#pragma unroll
      for (int h = 0; h < 2; h++)
#pragma unroll
        for (int s = 0; s < P; s++) x[s][h] = __byte_perm(r[c][4 * h + s], 0, 0x3120);
#pragma unroll
      for (int s = 0; s < P; s++) {
        exchange_bits(x[s][0], x[s][1], 4, 0x0F0F0F0Fu);
        exchange_bits(x[s][0], x[s][1], 8, 0x00FF00FFu);
      }
#pragma unroll
      for (int h = 0; h < 2; h++) {
        exchange_bits(x[0][h], x[1][h], 1, 0x55555555u);
        exchange_bits(x[2][h], x[3][h], 1, 0x55555555u);
        exchange_bits(x[0][h], x[2][h], 2, 0x33333333u);
        exchange_bits(x[1][h], x[3][h], 2, 0x33333333u);
      }
#pragma unroll
      for (int s = 0; s < P; s++) w[s] = ((uint64_t)x[s][1] << 32) | x[s][0];
    } else {
      uint32_t x[M];
#pragma unroll
      for (int k = 0; k < M; k++) x[k] = unshuffle32(r[c][k]);  // halfword s = pattern s, bit j = row 4j+k
#pragma unroll
      for (int s = 0; s < P; s++) {
        // byte k of ul = low byte of halfword s of x[k]; byte k of uh = its high byte
        const uint32_t bl = 2 * s, bh = 2 * s + 1;
        const uint32_t ul = __byte_perm(__byte_perm(x[0], x[1], bl | ((4 + bl) << 4)), __byte_perm(x[2], x[3], bl | ((4 + bl) << 4)), 0x5410);
        const uint32_t uh = __byte_perm(__byte_perm(x[0], x[1], bh | ((4 + bh) << 4)), __byte_perm(x[2], x[3], bh | ((4 + bh) << 4)), 0x5410);
        w[s] = ((uint64_t)shuffle32(shuffle32(uh)) << 32) | shuffle32(shuffle32(ul));
      }
    }
  }
};

}  // namespace lifeb200
