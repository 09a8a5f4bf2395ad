/*
 * convolve_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU restatement of the SURVEY.md 8(f) rank-4 rows, the step AFTER the stepping path:
 *   LifeState::Mirrored            LifeAPI.hpp:789-795 (FlipX :764, FlipY :753-755, Move(1,1))
 *   LifeState::Convolve            LifeAPI.hpp:1293-1370 (+ Bits.hpp:132-143 convolve_uint64_t)
 *   MatchLive / MatchLiveAndDead   LifeAPI.hpp:427-435
 *   Match(LifeState)               LifeAPI.hpp:440-442 (dead = live.GetBoundary(), :538)
 *   Match(LifeTarget)              LifeTarget.hpp:53-54
 *   InteractionOffsets             LifeAPI.hpp:1066-1095
 *   LifeState::RLE / LifeHistory::RLE / GenericRLE   Parsing.hpp:8-66,198-202; LifeHistory.hpp:62-68
 *   LifeState::Parse / LifeHistory::Parse            Parsing.hpp:141-196; LifeHistory.hpp:70-93
 * The reference's Convolve walks runs of set bits and doubles them; that is an optimisation of
 * the definition used here (the torus Minkowski sum, one shifted copy per ON cell), so the two
 * agree on every input.  Pinned against the unmodified reference (oracle/_ref: ref_convolve,
 * ref_match_*, ref_interaction_offsets, ref_rle, ref_history_rle) in tests/test_oracle.py.
 */
#include <stdint.h>
#include <stddef.h>
#include <stdio.h>
#include <string.h>

#define LN 64
void lo_neighbour_count(const uint64_t *s, uint64_t *planes);
void lo_zoi(const uint64_t *s, uint64_t *out);

static inline uint64_t rotl64(uint64_t v, unsigned s) { s &= 63; return s ? (v << s) | (v >> (64 - s)) : v; }
static inline uint64_t brev64(uint64_t v) {
  uint64_t r = 0;
  for (int i = 0; i < 64; i++) r |= ((v >> i) & 1ULL) << (63 - i);
  return r;
}

/* (x, y) -> (-x, -y) on the torus */
void lo_mirrored(const uint64_t *s, uint64_t *out) {
  uint64_t t[LN];
  for (int i = 0; i < LN; i++) t[i] = brev64(s[LN - 1 - i]);         /* FlipX (bits) then FlipY (words) */
  for (int i = 0; i < LN; i++) out[(i + 1) & 63] = rotl64(t[i], 1);   /* Move(1, 1) */
}

/* result = OR over ON cells (j, k) of a of b.Moved(j, k) */
void lo_convolve(const uint64_t *a, const uint64_t *b, uint64_t *out) {
  uint64_t r[LN];
  memset(r, 0, sizeof r);
  for (int j = 0; j < LN; j++)
    for (uint64_t x = a[j]; x; x &= x - 1) {
      unsigned k = (unsigned)__builtin_ctzll(x);
      for (int i = 0; i < LN; i++) r[(i + j) & 63] |= rotl64(b[i], k);
    }
  memcpy(out, r, sizeof r);
}

void lo_match_live_and_dead(const uint64_t *s, const uint64_t *live, const uint64_t *dead, uint64_t *out) {
  uint64_t inv[LN], ml[LN], md[LN], c1[LN], c2[LN];
  for (int i = 0; i < LN; i++) inv[i] = ~s[i];
  lo_mirrored(live, ml);
  lo_mirrored(dead, md);
  lo_convolve(inv, ml, c1);
  lo_convolve(s, md, c2);
  for (int i = 0; i < LN; i++) out[i] = ~c1[i] & ~c2[i];
}
void lo_match_live(const uint64_t *s, const uint64_t *live, uint64_t *out) {
  uint64_t inv[LN], ml[LN], c1[LN];
  for (int i = 0; i < LN; i++) inv[i] = ~s[i];
  lo_mirrored(live, ml);
  lo_convolve(inv, ml, c1);
  for (int i = 0; i < LN; i++) out[i] = ~c1[i];
}
void lo_match_state(const uint64_t *s, const uint64_t *live, uint64_t *out) {
  uint64_t dead[LN];
  lo_zoi(live, dead);
  for (int i = 0; i < LN; i++) dead[i] &= ~live[i]; /* GetBoundary */
  lo_match_live_and_dead(s, live, dead, out);
}
/* target = wanted[64], unwanted[64] */
void lo_match_target(const uint64_t *s, const uint64_t *target, uint64_t *out) {
  lo_match_live_and_dead(s, target, target + LN, out);
}

void lo_interaction_offsets(const uint64_t *a, const uint64_t *other, uint64_t *out) {
  uint64_t b[LN], an[4 * LN], bn[4 * LN], x[LN], y[LN], c[LN], r[LN];
  lo_mirrored(other, b);
  lo_neighbour_count(a, an); /* bit3, bit2, bit1, bit0 */
  lo_neighbour_count(b, bn);
  const uint64_t *a3 = an, *a2 = an + LN, *a1 = an + 2 * LN, *a0 = an + 3 * LN;
  const uint64_t *b3 = bn, *b2 = bn + LN, *b1 = bn + 2 * LN, *b0 = bn + 3 * LN;
  lo_convolve(a, b, r); /* overlaps */
#define TERM(XEXPR, YEXPR)                              \
  do {                                                  \
    for (int i = 0; i < LN; i++) { x[i] = (XEXPR); y[i] = (YEXPR); } \
    lo_convolve(x, y, c);                               \
    for (int i = 0; i < LN; i++) r[i] |= c[i];          \
  } while (0)
  /* births: exactly-1 next to exactly-2 */
  TERM(~a3[i] & ~a2[i] & ~a1[i] & a0[i] & ~a[i], ~b3[i] & ~b2[i] & b1[i] & ~b0[i] & ~b[i]);
  TERM(~b3[i] & ~b2[i] & ~b1[i] & b0[i] & ~b[i], ~a3[i] & ~a2[i] & a1[i] & ~a0[i] & ~a[i]);
  /* overcrowding */
  TERM(~a3[i] & ~a2[i] & a1[i] & a0[i] & a[i], (b3[i] | b2[i] | b1[i]) & ~b[i]);
  TERM((a2[i] | a3[i]) & a[i], (b3[i] | b2[i] | b1[i] | b0[i]) & ~b[i]);
  TERM(~b3[i] & ~b2[i] & b1[i] & b0[i] & b[i], (a3[i] | a2[i] | a1[i]) & ~a[i]);
  TERM((b2[i] | b3[i]) & b[i], (a3[i] | a2[i] | a1[i] | a0[i]) & ~a[i]);
#undef TERM
  memcpy(out, r, sizeof r);
}

/* ---- GenericRLE (Parsing.hpp:8-66): rows j = 32..63,0..31, columns i = 32..63,0..31 ---- */
typedef char (*cell_fn)(const void *ctx, int x, int y);

static size_t put_count(char *buf, size_t cap, size_t len, unsigned n) {
  char tmp[16];
  int k = snprintf(tmp, sizeof tmp, "%u", n);
  for (int i = 0; i < k; i++) { if (len < cap) buf[len] = tmp[i]; len++; }
  return len;
}
static size_t put_char(char *buf, size_t cap, size_t len, char c) {
  if (len < cap) buf[len] = c;
  return len + 1;
}

static size_t generic_rle(cell_fn cellchar, const void *ctx, char *buf, size_t cap) {
  size_t len = 0;
  unsigned eol_count = 0;
  for (unsigned j = 0; j < 64; j++) {
    char last_val = cellchar(ctx, 32, (int)((j + 32) & 63));
    unsigned run_count = 0;
    for (unsigned i = 0; i < LN; i++) {
      char val = cellchar(ctx, (int)((i + 32) & 63), (int)((j + 32) & 63));
      if (val != '.' && val != 'b' && eol_count > 0) {
        if (eol_count > 1) len = put_count(buf, cap, len, eol_count);
        len = put_char(buf, cap, len, '$');
        eol_count = 0;
      }
      if (val != last_val) {
        if (run_count > 1) len = put_count(buf, cap, len, run_count);
        len = put_char(buf, cap, len, last_val);
        run_count = 0;
      }
      run_count++;
      last_val = val;
    }
    if (last_val != '.' && last_val != 'b') {
      if (run_count > 1) len = put_count(buf, cap, len, run_count);
      len = put_char(buf, cap, len, last_val);
    }
    eol_count++;
  }
  len = put_char(buf, cap, len, '!');
  if (len < cap) buf[len] = 0;
  return len; /* length without the terminator; truncated if >= cap */
}

static inline int getbit(const uint64_t *p, int x, int y) { return (int)((p[x & 63] >> (y & 63)) & 1); }
static char state_char(const void *ctx, int x, int y) { return getbit((const uint64_t *)ctx, x, y) ? 'o' : 'b'; }
/* LifeHistory::StateToChar (LifeHistory.hpp:25-44): state + 2 history + 4 marked + 8 original */
static char history_char(const void *ctx, int x, int y) {
  const uint64_t *r = (const uint64_t *)ctx; /* state, history, marked, original */
  unsigned v = (unsigned)getbit(r, x, y) + ((unsigned)getbit(r + LN, x, y) << 1) + ((unsigned)getbit(r + 2 * LN, x, y) << 2) +
               ((unsigned)getbit(r + 3 * LN, x, y) << 3);
  switch (v) {
    case 0: return '.';
    case 1: return 'A';
    case 2: return 'B';
    case 5: return 'C';
    case 4: return 'D';
    case 9: return 'E';
    default: return 'F';
  }
}
size_t lo_rle(const uint64_t *state, char *buf, size_t cap) { return generic_rle(state_char, state, buf, cap); }
size_t lo_history_rle(const uint64_t *rec, char *buf, size_t cap) { return generic_rle(history_char, rec, buf, cap); }

/* GenericParse (Parsing.hpp:141-190) with the LifeState cell interpreter (:192-198): header lines
 * (first char 'x') dropped, digits accumulate a count, '$' moves down, '!' stops, CR/LF/space skipped,
 * any other character advances x by the count and sets the cell only when it is 'o'. */
void lo_parse(const char *rle, uint64_t *out) {
  memset(out, 0, LN * sizeof(uint64_t));
  int cnt = 0, x = 0, y = 0;
  const char *p = rle;
  while (*p) {
    /* one line */
    const char *e = strchr(p, '\n');
    size_t n = e ? (size_t)(e - p) : strlen(p);
    if (!(n > 0 && p[0] == 'x')) {
      for (size_t k = 0; k < n; k++) {
        char ch = p[k];
        if (ch >= '0' && ch <= '9') { cnt = cnt * 10 + (ch - '0'); }
        else if (ch == '$') { if (cnt == 0) cnt = 1; if (cnt == 129) return; y += cnt; x = 0; cnt = 0; }
        else if (ch == '!') return;
        else if (ch == '\r' || ch == '\n' || ch == ' ') continue;
        else {
          if (cnt == 0) cnt = 1;
          for (int j = 0; j < cnt; j++) { if (ch == 'o') out[x & 63] |= 1ULL << (y & 63); x++; }
          cnt = 0;
        }
      }
    }
    if (!e) break;
    p = e + 1;
  }
}
