/*
 * next_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU restatement of the SURVEY.md 8(f) "next" rows: LifeWeld::Step (LifeWeld.hpp:169-186),
 * InteractionCountsAndNext (LifeAPI.hpp:997-1040) and GetHash = XXH3_64bits over the 512-byte state
 * (LifeAPI.hpp:373; algorithm restated from the published xxHash v0.8.3 specification of
 * XXH3_hashLong_64b with the default secret -- the reference vendors it as a submodule).
 * Pinned against the unmodified reference (oracle/_ref: ref_weld_step, ref_interaction_counts,
 * ref_hash) in tests/test_oracle.py and against the survey's known XXH3 answers.
 */
#include <stdint.h>
#include <stddef.h>
#include <string.h>

#define LN 64
void lo_neighbour_count(const uint64_t *s, uint64_t *planes);
void lo_count_rows(const uint64_t *s, uint64_t *bit0, uint64_t *bit1);

/* record = state, frozen2, frozen1, frozen0 */
void lo_weld_step(uint64_t *rec) {
  uint64_t nc[4 * LN];
  lo_neighbour_count(rec, nc); /* bit3, bit2, bit1, bit0; bit3 is ignored by the reference */
  for (int i = 0; i < LN; i++) {
    uint64_t b2 = nc[LN + i], b1 = nc[2 * LN + i], b0 = nc[3 * LN + i];
    uint64_t f2 = rec[LN + i], f1 = rec[2 * LN + i], f0 = rec[3 * LN + i];
    uint64_t s0 = b0 ^ f0, c0 = b0 & f0;
    uint64_t s1 = b1 ^ f1 ^ c0, c1 = (b1 & f1) | (c0 & (b1 ^ f1));
    uint64_t s2 = b2 ^ f2 ^ c1;
    rec[i] = (s0 ^ s2) & (s1 ^ s2) & (rec[i] | s0);
  }
}
void lo_weld_step_n(uint64_t *rec, unsigned gens) { for (unsigned g = 0; g < gens; g++) lo_weld_step(rec); }

/* out = out1, out2, outMore, next */
void lo_interaction_counts(const uint64_t *s, uint64_t *out) {
  uint64_t col0[LN], col1[LN];
  lo_count_rows(s, col0, col1);
  for (int i = 0; i < LN; i++) {
    int u = (i + LN - 1) & 63, b = (i + 1) & 63;
    uint64_t fs = col0[u] ^ col0[i] ^ col0[b], fc = (col0[u] & col0[i]) | (col0[b] & (col0[u] ^ col0[i]));
    uint64_t cs = col1[u] ^ col1[i] ^ col1[b], cc = (col1[u] & col1[i]) | (col1[b] & (col1[u] ^ col1[i]));
    uint64_t o1 = ~s[i] & ~cc & fs & ~cs & ~fc;
    uint64_t o2 = ~s[i] & ~cc & ~fs & (cs ^ fc);
    uint64_t om = ~s[i] & ~o2 & (fc | cs | cc);
    cc ^= cs & fc;
    out[i] = o1; out[LN + i] = o2; out[2 * LN + i] = om;
    out[3 * LN + i] = (fs ^ cc) & (fc ^ cs ^ cc) & (s[i] | fs);
  }
}

/* ---- XXH3_64bits for len = 512, default secret, seed 0 ---- */
static const uint8_t kSecret[192] = {
  0xb8, 0xfe, 0x6c, 0x39, 0x23, 0xa4, 0x4b, 0xbe, 0x7c, 0x01, 0x81, 0x2c, 0xf7, 0x21, 0xad, 0x1c,
  0xde, 0xd4, 0x6d, 0xe9, 0x83, 0x90, 0x97, 0xdb, 0x72, 0x40, 0xa4, 0xa4, 0xb7, 0xb3, 0x67, 0x1f,
  0xcb, 0x79, 0xe6, 0x4e, 0xcc, 0xc0, 0xe5, 0x78, 0x82, 0x5a, 0xd0, 0x7d, 0xcc, 0xff, 0x72, 0x21,
  0xb8, 0x08, 0x46, 0x74, 0xf7, 0x43, 0x24, 0x8e, 0xe0, 0x35, 0x90, 0xe6, 0x81, 0x3a, 0x26, 0x4c,
  0x3c, 0x28, 0x52, 0xbb, 0x91, 0xc3, 0x00, 0xcb, 0x88, 0xd0, 0x65, 0x8b, 0x1b, 0x53, 0x2e, 0xa3,
  0x71, 0x64, 0x48, 0x97, 0xa2, 0x0d, 0xf9, 0x4e, 0x38, 0x19, 0xef, 0x46, 0xa9, 0xde, 0xac, 0xd8,
  0xa8, 0xfa, 0x76, 0x3f, 0xe3, 0x9c, 0x34, 0x3f, 0xf9, 0xdc, 0xbb, 0xc7, 0xc7, 0x0b, 0x4f, 0x1d,
  0x8a, 0x51, 0xe0, 0x4b, 0xcd, 0xb4, 0x59, 0x31, 0xc8, 0x9f, 0x7e, 0xc9, 0xd9, 0x78, 0x73, 0x64,
  0xea, 0xc5, 0xac, 0x83, 0x34, 0xd3, 0xeb, 0xc3, 0xc5, 0x81, 0xa0, 0xff, 0xfa, 0x13, 0x63, 0xeb,
  0x17, 0x0d, 0xdd, 0x51, 0xb7, 0xf0, 0xda, 0x49, 0xd3, 0x16, 0x55, 0x26, 0x29, 0xd4, 0x68, 0x9e,
  0x2b, 0x16, 0xbe, 0x58, 0x7d, 0x47, 0xa1, 0xfc, 0x8f, 0xf8, 0xb8, 0xd1, 0x7a, 0xd0, 0x31, 0xce,
  0x45, 0xcb, 0x3a, 0x8f, 0x95, 0x16, 0x04, 0x28, 0xaf, 0xd7, 0xfb, 0xca, 0xbb, 0x4b, 0x40, 0x7e,
};
static inline uint64_t rd64(const uint8_t *p) { uint64_t v; memcpy(&v, p, 8); return v; } /* little-endian hosts */

static void accumulate_stripe(uint64_t acc[8], const uint8_t *in, const uint8_t *sec) {
  for (int i = 0; i < 8; i++) {
    uint64_t d = rd64(in + 8 * i), k = d ^ rd64(sec + 8 * i);
    acc[i ^ 1] += d;
    acc[i] += (k & 0xffffffffULL) * (k >> 32);
  }
}

uint64_t lo_xxh3_512(const uint64_t *state) {
  const uint8_t *in = (const uint8_t *)state;
  const size_t len = 512;
  uint64_t acc[8] = { 0xC2B2AE3DULL, 0x9E3779B185EBCA87ULL, 0xC2B2AE3D27D4EB4FULL, 0x165667B19E3779F9ULL,
                      0x85EBCA77C2B2AE63ULL, 0x85EBCA77ULL, 0x27D4EB2F165667C5ULL, 0x9E3779B1ULL };
  size_t stripes = (len - 1) / 64; /* 7 full stripes: fewer than one 16-stripe block, so no scramble */
  for (size_t s = 0; s < stripes; s++) accumulate_stripe(acc, in + 64 * s, kSecret + 8 * s);
  accumulate_stripe(acc, in + len - 64, kSecret + 192 - 64 - 7); /* last stripe */
  uint64_t r = (uint64_t)len * 0x9E3779B185EBCA87ULL;
  for (int k = 0; k < 4; k++) {
    unsigned __int128 m = (unsigned __int128)(acc[2 * k] ^ rd64(kSecret + 11 + 16 * k)) *
                          (acc[2 * k + 1] ^ rd64(kSecret + 11 + 16 * k + 8));
    r += (uint64_t)m ^ (uint64_t)(m >> 64);
  }
  r ^= r >> 37;
  r *= 0x165667919E3779F9ULL;
  r ^= r >> 32;
  return r;
}
