// widen.cu -- two more callers of the stepping path's building blocks, batched (VERDICT r01 "widening"):
//   NeighbourCount::Add / operator+ / Subtract / operator- / WithExactly   (NeighbourCount.hpp:71-102)
//   LifeWeld::InteractionCounts                                             (LifeWeld.hpp:193-215)
// Both are single passes over whole records, bound by HBM: the ripple adder is 8 LOP3 per 32 cells.
#include <algorithm>

#include "ctx.h"
#include "host_pipeline.cuh"
#include "life_core.cuh"

namespace lifeb200 {

constexpr int kBlockW = 256;

// A NeighbourCount record = bit3, bit2, bit1, bit0 planes (NeighbourCount.hpp:7-11) = 4 x 64 words.  One thread
// owns two consecutive words (16 bytes) of every plane of one record; a warp covers one whole record, so every
// plane access of a warp is one contiguous 512-byte request.
// mode 0: out = a.Add(b, incarry)  (incarry == nullptr: the empty plane, i.e. operator+)
// mode 1: out = a.Subtract(b) = a.Add(~b, all-ones)
__global__ void __launch_bounds__(kBlockW) nc_add_kernel(const uint64_t* __restrict__ a, const uint64_t* __restrict__ b,
                                                         const uint64_t* __restrict__ incarry, uint64_t* __restrict__ out,
                                                         size_t n, int mode) {
  const size_t gid = (size_t)blockIdx.x * kBlockW + threadIdx.x;
  const size_t p = gid >> 5;
  const int lane = threadIdx.x & 31;
  if (p >= n) return;
  const ulonglong2* pa = reinterpret_cast<const ulonglong2*>(a + p * 4 * kN) + lane;
  const ulonglong2* pb = reinterpret_cast<const ulonglong2*>(b + p * 4 * kN) + lane;
  ulonglong2* po = reinterpret_cast<ulonglong2*>(out + p * 4 * kN) + lane;
  ulonglong2 carry = make_ulonglong2(0, 0);
  if (mode == 1) carry = make_ulonglong2(~0ull, ~0ull);
  else if (incarry) carry = reinterpret_cast<const ulonglong2*>(incarry + p * kN)[lane];
#pragma unroll
  for (int plane = 3; plane >= 0; plane--) {  // bit0 is the LAST plane of the record: ripple from it upwards
    const ulonglong2 x = pa[plane * 32];
    ulonglong2 y = pb[plane * 32];
    if (mode == 1) { y.x = ~y.x; y.y = ~y.y; }
    ulonglong2 s;
    s.x = x.x ^ y.x ^ carry.x;  // LifeState::FullAdd (LifeAPI.hpp:850-864): sum, then carry = majority
    s.y = x.y ^ y.y ^ carry.y;
    carry.x = (x.x & y.x) | (carry.x & (x.x ^ y.x));
    carry.y = (x.y & y.y) | (carry.y & (x.y ^ y.y));
    po[plane * 32] = s;
  }
}

// out[p] = counts[p].WithExactly(k): AND over the four planes of (bit or ~bit) by the bits of k
__global__ void __launch_bounds__(kBlockW) nc_with_exactly_kernel(const uint64_t* __restrict__ a, unsigned k,
                                                                  uint64_t* __restrict__ out, size_t n) {
  const size_t gid = (size_t)blockIdx.x * kBlockW + threadIdx.x;
  const size_t p = gid >> 5;
  const int lane = threadIdx.x & 31;
  if (p >= n) return;
  const ulonglong2* pa = reinterpret_cast<const ulonglong2*>(a + p * 4 * kN) + lane;
  ulonglong2 r = make_ulonglong2(~0ull, ~0ull);
#pragma unroll
  for (int bit = 0; bit < 4; bit++) {
    const ulonglong2 x = pa[(3 - bit) * 32];
    const unsigned long long flip = (k >> bit) & 1u ? 0ull : ~0ull;
    r.x &= x.x ^ flip;
    r.y &= x.y ^ flip;
  }
  reinterpret_cast<ulonglong2*>(out + p * kN)[lane] = r;
}

// LifeWeld::InteractionCounts: LifeState::InteractionCounts of the state, every plane masked by the zone of influence
// of the NON-frozen cells (state & ~frozen2 & ~frozen1 & ~frozen0).ZOI().  A warp per record; out = out1, out2, outMore.
__global__ void __launch_bounds__(kBlockW) weld_interaction_kernel(const uint64_t* __restrict__ records,
                                                                   uint64_t* __restrict__ out, size_t n) {
  constexpr int LPP = 32, W = Tile<LPP>::W;
  const size_t p = ((size_t)blockIdx.x * kBlockW + threadIdx.x) >> 5;
  const int sub = threadIdx.x & 31;
  const bool valid = p < n;  // invalid warps still shuffle with a full mask (whole warps only: p is warp-uniform)
  Tile<LPP> s, f2, f1, f0;
  if (valid) {
    const uint64_t* r = records + p * 4 * kN;
    s.load(r, sub); f2.load(r + kN, sub); f1.load(r + 2 * kN, sub); f0.load(r + 3 * kN, sub);
  } else {
#pragma unroll
    for (int w = 0; w < W; w++) s.lo[w] = s.hi[w] = f2.lo[w] = f2.hi[w] = f1.lo[w] = f1.hi[w] = f0.lo[w] = f0.hi[w] = 0;
  }
  // the reference's formulas on the column counts (LifeAPI.hpp:956-995), as csrc/next.cu does for the unfrozen form
  Tile<LPP> o1, o2, om;
  {
    const int up_src = sub + LPP - 1, dn_src = sub + 1;
    Cnt first = count_rows(s.lo[0], s.hi[0]);
    Cnt last = count_rows(s.lo[W - 1], s.hi[W - 1]);
    Cnt prev = halo<LPP>(last, up_src);
    Cnt after = halo<LPP>(first, dn_src);
    Cnt cur = first;
#pragma unroll
    for (int w = 0; w < W; w++) {
      Cnt next = (w == W - 1) ? after : (w + 1 == W - 1) ? last : count_rows(s.lo[w + 1], s.hi[w + 1]);
#pragma unroll
      for (int h = 0; h < 2; h++) {
        const uint32_t a = h ? s.hi[w] : s.lo[w];
        const uint32_t u0 = h ? prev.b0hi : prev.b0lo, u1 = h ? prev.b1hi : prev.b1lo;
        const uint32_t c0 = h ? cur.b0hi : cur.b0lo, c1 = h ? cur.b1hi : cur.b1lo;
        const uint32_t l0 = h ? next.b0hi : next.b0lo, l1 = h ? next.b1hi : next.b1lo;
        const uint32_t fs = lop3<kXor3>(u0, c0, l0), fc = lop3<kMaj>(u0, c0, l0);
        const uint32_t cs = lop3<kXor3>(u1, c1, l1), cc = lop3<kMaj>(u1, c1, l1);
        const uint32_t r1 = ~a & ~cc & fs & ~cs & ~fc;
        const uint32_t r2 = ~a & ~cc & ~fs & (cs ^ fc);
        const uint32_t rm = ~a & ~r2 & (fc | cs | cc);
        if (h) { o1.hi[w] = r1; o2.hi[w] = r2; om.hi[w] = rm; }
        else { o1.lo[w] = r1; o2.lo[w] = r2; om.lo[w] = rm; }
      }
      prev = cur; cur = next;
    }
  }
  Tile<LPP> nf;
#pragma unroll
  for (int w = 0; w < W; w++) {
    nf.lo[w] = s.lo[w] & ~(f2.lo[w] | f1.lo[w] | f0.lo[w]);
    nf.hi[w] = s.hi[w] & ~(f2.hi[w] | f1.hi[w] | f0.hi[w]);
  }
  const Tile<LPP> zoi = dilate<LPP>(nf, sub, false);
  if (!valid) return;
#pragma unroll
  for (int w = 0; w < W; w++) {
    o1.lo[w] &= zoi.lo[w]; o1.hi[w] &= zoi.hi[w];
    o2.lo[w] &= zoi.lo[w]; o2.hi[w] &= zoi.hi[w];
    om.lo[w] &= zoi.lo[w]; om.hi[w] &= zoi.hi[w];
  }
  uint64_t* o = out + p * 3 * kN;
  o1.store(o, sub); o2.store(o + kN, sub); om.store(o + 2 * kN, sub);
}

static inline unsigned grid_w(size_t n) { return (unsigned)((n * 32 + kBlockW - 1) / kBlockW); }

}  // namespace lifeb200

using namespace lifeb200;

extern "C" {

int lifeb200_nc_add_dev(lifeb200_ctx* ctx, const uint64_t* a, const uint64_t* b, const uint64_t* incarry, uint64_t* out, size_t n) {
  LIFEB200_CHECK_CTX(ctx);
  if (n == 0) return LIFEB200_OK;
  if (!a || !b || !out || !aligned16(a) || !aligned16(b) || !aligned16(out) || !aligned16(incarry))
    return set_error(LIFEB200_E_ARG, "null or misaligned pointer");
  nc_add_kernel<<<grid_w(n), kBlockW, 0, ctx->stream>>>(a, b, incarry, out, n, 0);
  LIFEB200_LAUNCHED(ctx);
  return LIFEB200_OK;
}

int lifeb200_nc_subtract_dev(lifeb200_ctx* ctx, const uint64_t* a, const uint64_t* b, uint64_t* out, size_t n) {
  LIFEB200_CHECK_CTX(ctx);
  if (n == 0) return LIFEB200_OK;
  if (!a || !b || !out || !aligned16(a) || !aligned16(b) || !aligned16(out)) return set_error(LIFEB200_E_ARG, "null or misaligned pointer");
  nc_add_kernel<<<grid_w(n), kBlockW, 0, ctx->stream>>>(a, b, nullptr, out, n, 1);
  LIFEB200_LAUNCHED(ctx);
  return LIFEB200_OK;
}

int lifeb200_nc_with_exactly_dev(lifeb200_ctx* ctx, const uint64_t* counts, unsigned k, uint64_t* out, size_t n) {
  LIFEB200_CHECK_CTX(ctx);
  if (n == 0) return LIFEB200_OK;
  if (!counts || !out || !aligned16(counts) || !aligned16(out)) return set_error(LIFEB200_E_ARG, "null or misaligned pointer");
  nc_with_exactly_kernel<<<grid_w(n), kBlockW, 0, ctx->stream>>>(counts, k, out, n);
  LIFEB200_LAUNCHED(ctx);
  return LIFEB200_OK;
}

int lifeb200_weld_interaction_counts_dev(lifeb200_ctx* ctx, const uint64_t* records, uint64_t* planes, size_t n) {
  LIFEB200_CHECK_CTX(ctx);
  if (n == 0) return LIFEB200_OK;
  if (!records || !planes || !aligned16(records) || !aligned16(planes)) return set_error(LIFEB200_E_ARG, "null or misaligned pointer");
  weld_interaction_kernel<<<grid_w(n), kBlockW, 0, ctx->stream>>>(records, planes, n);
  LIFEB200_LAUNCHED(ctx);
  return LIFEB200_OK;
}

int lifeb200_weld_interaction_counts_host(lifeb200_ctx* ctx, const uint64_t* records, uint64_t* planes, size_t n) {
  LIFEB200_CHECK_CTX(ctx);
  if (n == 0) return LIFEB200_OK;
  if (!records || !planes) return set_error(LIFEB200_E_ARG, "null pointer");
  return host_pipeline_impl(ctx, records, planes, n, 4 * LIFEB200_PATTERN_BYTES, 3 * LIFEB200_PATTERN_BYTES, true, (size_t)1 << 14,
                            [&](cudaStream_t st, const uint64_t* din, uint64_t* dout, size_t cnt) -> int {
                              weld_interaction_kernel<<<grid_w(cnt), kBlockW, 0, st>>>(din, dout, cnt);
                              LIFEB200_LAUNCHED(ctx);
                              return LIFEB200_OK;
                            });
}

}  // extern "C"
