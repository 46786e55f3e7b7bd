/* dc_oracle_bytewise.c -- CPU ORACLE, TEST INFRASTRUCTURE ONLY (see dc_oracle.h).
 *
 * Restatement of the byte-wise codec CT=1 (myCompress impl/dataCompression.c:3977-4118, myDecompress :3942-3975): four
 * open-loop predictors on the ORIGINAL history, a hit is one char 'a'..'d' plus its 1-based position, everything else is
 * the raw float.  -1.0f marks an empty history slot (the same quirk as SURVEY F7).  PINNED against the compiled reference
 * by tests/test_oracle_bytewise.py. */
#include <math.h>
#include "dc_oracle.h"

/* -> number of raw floats nf; hits = n - nf.  af[nf], ac[n-nf], ad[n-nf] (capacity n each). */
size_t dco_bytewise_compress(const float* x, size_t n, double eb, float* af, char* ac, int* ad) {
  float b1 = -1, b2 = -1, b3 = -1, b4 = -1;
  size_t nf = 0, nc = 0;
  for (size_t i = 0; i < n; i++) {
    const float v = x[i];
    if (b4 == -1 || b3 == -1 || b2 == -1 || b1 == -1) {                 /* :3996-4027 */
      af[nf++] = v;
      if (b4 == -1) b4 = v; else if (b3 == -1) b3 = v; else if (b2 == -1) b2 = v; else b1 = v;
      continue;
    }
    const float p1 = b1, p2 = 2 * b1 - b2, p3 = 3 * b1 - 3 * b2 + b3, p4 = 4 * b1 - 6 * b2 + 4 * b3 - b4;   /* :4030-4033 */
    const float d1 = (float)fabs(p1 - v), d2 = (float)fabs(p2 - v), d3 = (float)fabs(p3 - v), d4 = (float)fabs(p4 - v);
    float dm = d1; char t = 'a';
    if (d2 < dm) { dm = d2; t = 'b'; }
    if (d3 < dm) { dm = d3; t = 'c'; }
    if (d4 < dm) { dm = d4; t = 'd'; }
    b4 = b3; b3 = b2; b2 = b1; b1 = v;                                   /* :4063-4066: the ORIGINAL value */
    if (dm <= eb) { ac[nc] = t; ad[nc] = (int)(nf + nc + 1); nc++; }     /* :4068-4079 */
    else af[nf++] = v;
  }
  return nf;
}

/* myDecompress :3942-3975; nc = number of hits (the reference reads array_char_displacement[nc] past its end) */
void dco_bytewise_decompress(const float* af, const char* ac, const int* ad, size_t nc, size_t n, float* out) {
  size_t fp = 0, cp = 0;
  for (size_t i = 0; i < n; i++) {
    if (cp < nc && (size_t)(ad[cp] - 1) == i) {
      const char t = ac[cp++];
      if (t == 'a') out[i] = out[i - 1];
      else if (t == 'b') out[i] = 2 * out[i - 1] - out[i - 2];
      else if (t == 'c') out[i] = 3 * out[i - 1] - 3 * out[i - 2] + out[i - 3];
      else if (t == 'd') out[i] = 4 * out[i - 1] - 6 * out[i - 2] + 4 * out[i - 3] - out[i - 4];
    } else {
      out[i] = af[fp++];
    }
  }
}

/* float64 twin (myCompress_double :3815, myDecompress_double :3778) -> number of raw doubles nf; hits = n - nf.  af[nf], ac[n-nf], ad[n-nf] (capacity n each). */
size_t dco_bytewise_compress_f64(const double* x, size_t n, double eb, double* af, char* ac, int* ad) {
  double b1 = -1, b2 = -1, b3 = -1, b4 = -1;
  size_t nf = 0, nc = 0;
  for (size_t i = 0; i < n; i++) {
    const double v = x[i];
    if (b4 == -1 || b3 == -1 || b2 == -1 || b1 == -1) {                 /* :3996-4027 */
      af[nf++] = v;
      if (b4 == -1) b4 = v; else if (b3 == -1) b3 = v; else if (b2 == -1) b2 = v; else b1 = v;
      continue;
    }
    const double p1 = b1, p2 = 2 * b1 - b2, p3 = 3 * b1 - 3 * b2 + b3, p4 = 4 * b1 - 6 * b2 + 4 * b3 - b4;   /* :4030-4033 */
    const double d1 = fabs(p1 - v), d2 = fabs(p2 - v), d3 = fabs(p3 - v), d4 = fabs(p4 - v);
    double dm = d1; char t = 'a';
    if (d2 < dm) { dm = d2; t = 'b'; }
    if (d3 < dm) { dm = d3; t = 'c'; }
    if (d4 < dm) { dm = d4; t = 'd'; }
    b4 = b3; b3 = b2; b2 = b1; b1 = v;                                   /* :4063-4066: the ORIGINAL value */
    if (dm <= eb) { ac[nc] = t; ad[nc] = (int)(nf + nc + 1); nc++; }     /* :4068-4079 */
    else af[nf++] = v;
  }
  return nf;
}

/* myDecompress :3942-3975; nc = number of hits (the reference reads array_char_displacement[nc] past its end) */
void dco_bytewise_decompress_f64(const double* af, const char* ac, const int* ad, size_t nc, size_t n, double* out) {
  size_t fp = 0, cp = 0;
  for (size_t i = 0; i < n; i++) {
    if (cp < nc && (size_t)(ad[cp] - 1) == i) {
      const char t = ac[cp++];
      if (t == 'a') out[i] = out[i - 1];
      else if (t == 'b') out[i] = 2 * out[i - 1] - out[i - 2];
      else if (t == 'c') out[i] = 3 * out[i - 1] - 3 * out[i - 2] + out[i - 3];
      else if (t == 'd') out[i] = 4 * out[i - 1] - 6 * out[i - 2] + 4 * out[i - 3] - out[i - 4];
    } else {
      out[i] = af[fp++];
    }
  }
}
