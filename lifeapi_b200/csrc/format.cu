// format.cu -- the data formats either side of the stepping path: RLE in and out.
//   LifeState::RLE / LifeHistory::RLE    Parsing.hpp:8-66,198-202; LifeHistory.hpp:25-44,62-68
//   LifeState::Parse / LifeHistory::Parse  Parsing.hpp:141-196; LifeHistory.hpp:70-93
// Host code (strings are a host format): hit lists come back from the device as states / ids and are
// serialised here.  The encoder works on whole rows: the 64x64 bit matrix is transposed once
// (6 butterfly stages) so that a display row is one 64-bit word, and runs are taken with
// count-trailing-zeros instead of one cell at a time.
#include <cstring>
#include <string>

#include "ctx.h"

namespace lifeb200 {

// out[y] bit x = in[x] bit y
static void transpose64(const uint64_t* in, uint64_t* out) {
  uint64_t a[64];
  std::memcpy(a, in, sizeof a);
  // swap the off-diagonal s x s blocks of every 2s x 2s block, s = 32, 16, ..., 1
  for (int s = 32; s >= 1; s >>= 1) {
    uint64_t keep = 0;  // bit positions whose index has bit s clear
    for (int b = 0; b < 64; b++)
      if (!(b & s)) keep |= 1ULL << b;
    for (int k = 0; k < 64; k++) {
      if (k & s) continue;
      const uint64_t lo = a[k], hi = a[k + s];
      a[k] = (lo & keep) | ((hi & keep) << s);
      a[k + s] = ((lo >> s) & keep) | (hi & ~keep);
    }
  }
  std::memcpy(out, a, sizeof a);
}

static inline uint64_t rotr64(uint64_t v, int s) { return (v >> (s & 63)) | (v << ((64 - s) & 63)); }

struct Writer {
  char* buf;
  size_t cap, len = 0;
  void ch(char c) { if (len < cap) buf[len] = c; len++; }
  void num(unsigned n) {
    char tmp[12];
    int k = 0;
    do { tmp[k++] = (char)('0' + n % 10); n /= 10; } while (n);
    while (k) ch(tmp[--k]);
  }
  void run(unsigned count, char c) { if (count > 1) num(count); ch(c); }
};

// planes: n_planes transposed-and-recentred row words per display row; cls(row, i) gives the character
// class mask lookups.  Generic over the two alphabets through a small table of (mask planes -> char).
template <int NCH>
static size_t encode_rows(const uint64_t (*rows)[64], const char* chars, char dead, char* buf, size_t cap) {
  // rows[c][j] = mask of display row j whose cells print chars[c]; the classes partition the row
  Writer w{buf, cap};
  unsigned eol = 0;
  for (int j = 0; j < 64; j++) {
    uint64_t live = 0;
    for (int c = 0; c < NCH; c++)
      if (chars[c] != dead) live |= rows[c][j];
    if (live) {
      if (eol > 0) { if (eol > 1) w.num(eol); w.ch('$'); eol = 0; }
      const int last = 63 - __builtin_clzll(live);  // trailing dead cells are dropped
      int i = 0;
      while (i <= last) {
        int c = 0;
        while (!((rows[c][j] >> i) & 1ULL)) c++;
        const uint64_t rest = ~(rows[c][j] >> i);  // first cell of another class ends the run
        int r = rest ? __builtin_ctzll(rest) : 64 - i;
        if (r > 64 - i) r = 64 - i;
        w.run((unsigned)r, chars[c]);
        i += r;
      }
    }
    eol++;
  }
  w.ch('!');
  if (w.len < cap) buf[w.len] = 0;
  else if (cap) buf[cap - 1] = 0;
  return w.len;
}

// display row j, display column i  <->  cell x = (i + 32) & 63, y = (j + 32) & 63   (Parsing.hpp:13-18)
static void display_rows(const uint64_t* plane, uint64_t* rows) {
  uint64_t t[64];
  transpose64(plane, t);  // t[y] bit x
  for (int j = 0; j < 64; j++) rows[j] = rotr64(t[(j + 32) & 63], 32);
}

// GenericParse (Parsing.hpp:141-190): lines starting with 'x' are dropped; digits build a count; '$' moves
// down `count` rows (a count of 129 aborts, as in the reference); '!' stops; CR / LF / space are skipped;
// any other character occupies `count` cells, `cell(ch, x, y)` deciding what to set.
template <class CellFn>
static void parse_generic(const char* rle, CellFn cell) {
  int cnt = 0, x = 0, y = 0;
  bool line_start = true, skip_line = false;
  for (const char* p = rle; *p; p++) {
    const char ch = *p;
    if (ch == '\n') { line_start = true; skip_line = false; continue; }
    if (line_start) { line_start = false; skip_line = (ch == 'x'); }
    if (skip_line) continue;
    if (ch >= '0' && ch <= '9') cnt = cnt * 10 + (ch - '0');
    else if (ch == '$') {
      if (cnt == 0) cnt = 1;
      if (cnt == 129) return;
      y += cnt; x = 0; cnt = 0;
    } else if (ch == '!') return;
    else if (ch == '\r' || ch == ' ') continue;
    else {
      if (cnt == 0) cnt = 1;
      for (int k = 0; k < cnt; k++, x++) cell(ch, x, y);
      cnt = 0;
    }
  }
}

static inline void set_cell(uint64_t* plane, int x, int y) { plane[x & 63] |= 1ULL << (y & 63); }

}  // namespace lifeb200

using namespace lifeb200;

extern "C" {

int lifeb200_rle(const uint64_t* state, char* buf, size_t cap, size_t* len) {
  if (!state || (!buf && cap) || !len) return set_error(LIFEB200_E_ARG, "null pointer");
  uint64_t rows[2][64];
  display_rows(state, rows[0]);
  for (int j = 0; j < 64; j++) rows[1][j] = ~rows[0][j];
  static const char chars[2] = {'o', 'b'};
  *len = encode_rows<2>(rows, chars, 'b', buf, cap);
  if (*len >= cap) return set_error(LIFEB200_E_OVERFLOW, "RLE needs %zu bytes + terminator, buffer has %zu", *len, cap);
  return LIFEB200_OK;
}

int lifeb200_history_rle(const uint64_t* record, char* buf, size_t cap, size_t* len) {
  if (!record || (!buf && cap) || !len) return set_error(LIFEB200_E_ARG, "null pointer");
  uint64_t s[64], h[64], m[64], o[64];
  display_rows(record, s);
  display_rows(record + 64, h);
  display_rows(record + 128, m);
  display_rows(record + 192, o);
  // StateToChar (LifeHistory.hpp:25-44) on state + 2 history + 4 marked + 8 original
  uint64_t rows[7][64];
  static const char chars[7] = {'.', 'A', 'B', 'C', 'D', 'E', 'F'};
  for (int j = 0; j < 64; j++) {
    const uint64_t S = s[j], H = h[j], M = m[j], O = o[j];
    rows[0][j] = ~S & ~H & ~M & ~O;  // 0b0000
    rows[1][j] = S & ~H & ~M & ~O;   // 0b0001
    rows[2][j] = ~S & H & ~M & ~O;   // 0b0010
    rows[3][j] = S & ~H & M & ~O;    // 0b0101
    rows[4][j] = ~S & ~H & M & ~O;   // 0b0100
    rows[5][j] = S & ~H & ~M & O;    // 0b1001
    rows[6][j] = ~(rows[0][j] | rows[1][j] | rows[2][j] | rows[3][j] | rows[4][j] | rows[5][j]);
  }
  *len = encode_rows<7>(rows, chars, '.', buf, cap);
  if (*len >= cap) return set_error(LIFEB200_E_OVERFLOW, "RLE needs %zu bytes + terminator, buffer has %zu", *len, cap);
  return LIFEB200_OK;
}

int lifeb200_parse_rle(const char* rle, uint64_t* state) {
  if (!rle || !state) return set_error(LIFEB200_E_ARG, "null pointer");
  std::memset(state, 0, LIFEB200_PATTERN_BYTES);
  parse_generic(rle, [&](char ch, int x, int y) { if (ch == 'o') set_cell(state, x, y); });
  return LIFEB200_OK;
}

int lifeb200_history_parse_rle(const char* rle, uint64_t* record) {
  if (!rle || !record) return set_error(LIFEB200_E_ARG, "null pointer");
  std::memset(record, 0, 4 * LIFEB200_PATTERN_BYTES);
  uint64_t *s = record, *h = record + 64, *m = record + 128, *o = record + 192;
  parse_generic(rle, [&](char ch, int x, int y) {
    switch (ch) {  // LifeHistory.hpp:70-93
      case 'A': set_cell(s, x, y); break;
      case 'B': set_cell(h, x, y); break;
      case 'C': set_cell(s, x, y); set_cell(m, x, y); break;
      case 'D': set_cell(m, x, y); break;
      case 'E': set_cell(s, x, y); set_cell(o, x, y); break;
      default: break;
    }
  });
  return LIFEB200_OK;
}

}  // extern "C"
