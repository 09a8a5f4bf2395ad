// Search: can the Life rule tail be done in 3 LUT3 gates after 4 "adder-like" gates?
#include <stdio.h>
#include <stdint.h>
#include <stdlib.h>
// inputs: a, xa, xb, ya, yb (5) -> rows 0..31; bit r of input i = (r>>i)&1
static uint32_t IN[5];
static uint32_t F, CARE;
static inline int realizable(uint32_t s1, uint32_t s2, uint32_t s3, uint32_t F, uint32_t care, int *lut) {
  int l = 0;
  for (int k = 0; k < 8; k++) {
    uint32_t m = ((k & 1) ? s1 : ~s1) & ((k & 2) ? s2 : ~s2) & ((k & 4) ? s3 : ~s3) & care;
    uint32_t f = F & m;
    if (f != 0 && f != m) return 0;
    if (f) l |= 1 << k;
  }
  *lut = l;
  return 1;
}
static uint32_t apply(int lut, uint32_t s1, uint32_t s2, uint32_t s3) {
  uint32_t r = 0;
  for (int k = 0; k < 8; k++) if (lut >> k & 1) r |= ((k & 1) ? s1 : ~s1) & ((k & 2) ? s2 : ~s2) & ((k & 4) ? s3 : ~s3);
  return r;
}
int main() {
  for (int i = 0; i < 5; i++) { IN[i] = 0; for (int r = 0; r < 32; r++) if (r >> i & 1) IN[i] |= 1u << r; }
  int perms[3][4] = {{0,1,2,3},{0,1,3,2},{0,3,1,2}}; // code for value v: 3 classes up to flips/swap
  // value->code: class0: v -> v (binary); class1: 0->0,1->1,2->3,3->2 ; class2: 0->0,1->3,2->1,3->2
  int found = 0;
  for (int ex = 0; ex < 3; ex++) for (int ey = 0; ey < 3; ey++) {
    // build F over rows: decode X from (xa,xb) code, Y from (ya,yb)
    F = 0; CARE = 0xffffffffu;
    for (int r = 0; r < 32; r++) {
      int a = r & 1, cx = (r >> 1) & 3, cy = (r >> 3) & 3;
      int X = -1, Y = -1;
      for (int v = 0; v < 4; v++) { if (perms[ex][v] == cx) X = v; if (perms[ey][v] == cy) Y = v; }
      int S = X + 2 * Y;
      int out = (S == 3) || (a && S == 4);
      if (out) F |= 1u << r;
    }
    uint32_t sig[7];
    for (int i = 0; i < 5; i++) sig[i] = IN[i];
    for (int i1 = 0; i1 < 5; i1++) for (int j1 = i1 + 1; j1 < 5; j1++) for (int k1 = j1 + 1; k1 < 5; k1++)
      for (int l1 = 0; l1 < 256; l1++) {
        sig[5] = apply(l1, sig[i1], sig[j1], sig[k1]);
        for (int i2 = 0; i2 < 6; i2++) for (int j2 = i2 + 1; j2 < 6; j2++) for (int k2 = j2 + 1; k2 < 6; k2++)
          for (int l2 = 0; l2 < 256; l2++) {
            sig[6] = apply(l2, sig[i2], sig[j2], sig[k2]);
            // final gate must use sig[6] (else gate2 useless)
            for (int i3 = 0; i3 < 6; i3++) for (int j3 = i3 + 1; j3 < 6; j3++) {
              int lut;
              if (realizable(sig[i3], sig[j3], sig[6], F, CARE, &lut)) {
                printf("FOUND ex=%d ey=%d g1=(%d,%d,%d)lut%02x g2=(%d,%d,%d)lut%02x g3=(%d,%d,6)lut%02x\n", ex, ey, i1, j1, k1, l1, i2, j2, k2, l2, i3, j3, lut);
                found++;
                if (found > 20) return 0;
              }
            }
          }
      }
    printf("done ex=%d ey=%d found=%d\n", ex, ey, found); fflush(stdout);
  }
  return 0;
}
