/* dc_oracle_f64.c -- CPU ORACLE, TEST INFRASTRUCTURE ONLY (see dc_oracle.h).
 *
 * The float64 variants of the bit-wise codecs of /root/reference/impl/dataCompression.c: the
 * same algorithm as dc_oracle.c on the IEEE double layout -- raw code = sign + 11 exponent bits + m mantissa bits,
 * m = clamp(n_eb + (E-1023), 0, 52); CT=7 mask = the first 1+11+8 = 20 bits of the chunk mean; history in double.
 * Every function cites the reference lines it follows.  Pinned by tests/test_oracle_f64.py against the unmodified
 * reference compiled into oracle/_ref (the reference has no golden vectors for the double codecs).
 */
#include "dc_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

static inline uint64_t d2u(double d) { uint64_t u; memcpy(&u, &d, 8); return u; }
static inline double u2d(uint64_t u) { double d; memcpy(&d, &u, 8); return d; }

/* ---- toSmallDataset_double, :3522-3541 ------------------------------------- */
double dco_min_subtract_f64(const double* x, size_t n, double* y)
{
  double mn = x[0];
  for (size_t i = 1; i < n; i++) if (x[i] < mn) mn = x[i];
  for (size_t i = 0; i < n; i++) y[i] = x[i] - mn;
  return mn;
}

/* ---- med_dataset_double, :3564-3591: double accumulator in index order; thresholds 2^(add-1023),
 * add = 1024, 1536, 1792, ... (sum of 2^10 .. 2^i)  ->  type = 11 - i = 1..10 */
double dco_med_f64(const double* x, size_t n, int* type)
{
  double total = 0, mx = x[0];
  for (size_t i = 0; i < n; i++) {
    total += x[i];
    if (x[i] > mx) mx = x[i];
  }
  int add = 0;
  for (int i = 10; i > 0; i--) {
    add += 1 << i;
    if (mx < ldexp(1.0, add - 1023)) { *type = 11 - i; break; }
  }
  return total / (double)(int)n;
}

/* mask[1+11+8] of impl/k-means.c:527-530 = first 20 chars of doubletostr(medium) */
uint32_t dco_mask20(double medium) { return (uint32_t)(d2u(medium) >> 44); }

/* ---- MSB-first bit appender (add_bit_to_bytes :5456-5489), 64-bit codes ---- */
typedef struct { uint8_t* p; size_t bytes; uint64_t acc; int nacc; } bitw;   /* nacc < 8 between calls */
static inline void bw_put(bitw* w, uint64_t code, int len) /* len 0..64, code right-aligned */
{
  while (len > 0) {
    int take = len > 32 ? 32 : len;
    uint64_t part = (code >> (len - take)) & ((take == 64) ? ~0ull : ((1ull << take) - 1ull));
    w->acc = (w->acc << take) | part;
    w->nacc += take;
    while (w->nacc >= 8) { w->nacc -= 8; w->p[w->bytes++] = (uint8_t)(w->acc >> w->nacc); }
    len -= take;
  }
}
static inline int bw_finish(bitw* w)
{
  int rem = w->nacc;
  if (rem) { w->p[w->bytes++] = (uint8_t)(w->acc << (8 - rem)); w->nacc = 0; }
  return rem ? 8 - rem : 8;
}

/* m(x) = clamp(n_eb + (E-1023), 0, 52), compress_bitwise_double :3452-3470 */
static inline int mant_bits(uint64_t u, int n_eb)
{
  int m = n_eb + (int)((u >> 52) & 0x7FF) - 1023;
  return m > 52 ? 52 : (m < 0 ? 0 : m);
}

/* raw code: compress_bitwise_double :3446-3477 -- top 12+m bits, pure truncation */
static inline void put_raw(bitw* w, uint64_t u, int n_eb)
{
  int len = 12 + mant_bits(u, n_eb);
  bw_put(w, u >> (64 - len), len);
}

/* compress_bitwise_double_mask :1493-1588 */
static inline void put_mask(bitw* w, uint64_t u, int n_eb, int type, uint32_t mask20)
{
  int m = mant_bits(u, n_eb);
  if ((u >> 52) != (uint64_t)(mask20 >> 8)) { put_raw(w, u, n_eb); return; }     /* :1518-1527, :1581-1587 */
  uint64_t prefix = (1ull << type) - 1ull;                                       /* 0 then `type` ones */
  if (((u >> 44) & 0xFF) == (mask20 & 0xFF)) {                                   /* :1543-1556 */
    bw_put(w, prefix, 1 + type);
    bw_put(w, 0, 1);
    int k = m > 8 ? m - 8 : 0;                                                   /* chars 20..11+m */
    bw_put(w, k ? (u >> (52 - m)) & ((1ull << k) - 1ull) : 0, k);
  } else {                                                                       /* :1557-1571 */
    bw_put(w, prefix, 1 + type);
    bw_put(w, 1, 1);
    bw_put(w, m ? (u >> (52 - m)) & ((1ull << m) - 1ull) : 0, m);                /* chars 12..11+m */
  }
}

static inline void hist_push(double* b3, double* b2, double* b1, double v)
{
  if (*b3 == -1) *b3 = v;
  else if (*b2 == -1) *b2 = v;
  else if (*b1 == -1) *b1 = v;
  else { *b3 = *b2; *b2 = *b1; *b1 = v; }
}

/* myCompress_bitwise_double :3189-3307 (CT=5), _np :2633-2643 (CT=6), _op :355-474 (CT=11), _mask :1590-1701 (CT=7) */
size_t dco_compress_f64(int ct, const double* x, size_t n, double eb, int type, uint32_t mask20,
                        uint8_t* out, int* pos_out)
{
  bitw w = { out, 0, 0, 0 };
  int n_eb = dco_n_eb(eb);
  double b1 = -1, b2 = -1, b3 = -1;
  for (size_t i = 0; i < n; i++) {
    double v = x[i];
    uint64_t u = d2u(v);
    if (ct == DCO_CT_NP) { put_raw(&w, u, n_eb); continue; }
    int code = 0;
    if (b3 == -1 || b2 == -1 || b1 == -1) {                 /* :3204 history not full */
      if (fabs(v) < eb) code = 4;                           /* :3207 '100' */
      hist_push(&b3, &b2, &b1, v);
    } else {
      volatile double p2 = 2 * b1 - b2;                     /* :3237, every operation rounded to double */
      volatile double t1 = 3 * b1, t2 = 3 * b2;
      volatile double t3 = t1 - t2;
      volatile double p3 = t3 + b3;                         /* :3238 */
      double d1 = fabs(b1 - v), d2 = fabs(p2 - v), d3 = fabs(p3 - v);
      double dmin = d1; int sel = 5;
      if (d2 < dmin) { dmin = d2; sel = 6; }
      if (d3 < dmin) { dmin = d3; sel = 7; }
      b3 = b2; b2 = b1; b1 = v;                             /* :3260-3262 */
      if (fabs(v) < eb) code = 4;                           /* :3264 */
      else if (dmin <= eb) code = sel;                      /* :3271 */
    }
    if (code) bw_put(&w, (uint64_t)code, 3);
    else if (ct == DCO_CT_BITWISE) put_raw(&w, u, n_eb);
    else if (ct == DCO_CT_OP) bw_put(&w, u, 64);            /* :380-383, :467-470 full word */
    else put_mask(&w, u, n_eb, type, mask20);
  }
  int pos = bw_finish(&w);
  if (pos_out) *pos_out = pos;
  return w.bytes;
}

/* ---- bit reader: bits past the end read as 0 ---- */
typedef struct { const uint8_t* p; size_t bytes; uint64_t pos; } bitr;
static inline uint64_t br_peek64(const bitr* r) /* next 64 bits, left aligned */
{
  size_t b = (size_t)(r->pos >> 3);
  int sh = (int)(r->pos & 7);
  uint64_t hi = 0;
  for (int k = 0; k < 8; k++) hi = (hi << 8) | (b + k < r->bytes ? r->p[b + k] : 0);
  uint64_t nx = (b + 8 < r->bytes) ? r->p[b + 8] : 0;
  return sh ? (hi << sh) | (nx >> (8 - sh)) : hi;
}

/* transmitted top `have` bits, then a single '1', then zeros; exact when have == 64
 * (decompress_bitwise_double :2906-2920, _np :2438-2457) */
static inline double midpoint(uint64_t top, int have)
{
  if (have >= 64) return u2d(top);
  return u2d(top | (0x8000000000000000ull >> have));
}

/* myDecompress_bitwise_double :2656-2869 (CT=5), _np :2286-2436 (CT=6), _op :476-575 (CT=11),
 * _mask :1199-1394 + decompress_bitwise_double_mask :1396-1491 (CT=7).  Code-by-code, exactly n codes; the same
 * divergences in the reference's undefined corners as the float32 oracle (dc_oracle.c), with 9 -> 12 and 17 -> 20:
 * 3-bit codes always win (reference mis-parses them when type == 1, bits_num == 1+type+1, :1355); a flag-0 escape
 * carries max(m-8,0) payload bits (reference: offset_bits <= 0 for 0 < m <= 8, :1290-1296); codes of total length
 * 12 are emitted (reference blocks bits_num == 1+11, :1355). */
int dco_decompress_f64(int ct, const uint8_t* s, size_t bytes, size_t n, double eb, int type, uint32_t mask20,
                       double* out)
{
  bitr r = { s, bytes, 0 };
  int n_eb = dco_n_eb(eb);
  double b1 = -1, b2 = -1, b3 = -1;
  for (size_t i = 0; i < n; i++) {
    uint64_t w = br_peek64(&r);
    double v;
    int len;
    if (w >> 63) {
      if (ct == DCO_CT_NP) return -1;
      len = 3;
      switch ((w >> 61) & 3) {                       /* decompress_bitwise_double :2873-2892 */
        case 0: v = 0.0; break;
        case 1: v = b1; break;
        case 2: { volatile double p = 2 * b1 - b2; v = p; } break;
        default: { volatile double t1 = 3 * b1, t2 = 3 * b2; volatile double t3 = t1 - t2;
                   volatile double p = t3 + b3; v = p; } break;
      }
    } else if (ct == DCO_CT_OP) {
      len = 64; v = u2d(w);
    } else {
      int escape = 0;
      if (ct == DCO_CT_MASK) {
        uint64_t ones = (1ull << type) - 1ull;
        escape = (((w >> (63 - type)) & ones) == ones);
      }
      if (!escape) {
        int m = mant_bits(w, n_eb);
        len = 12 + m;
        v = midpoint(len == 64 ? w : (w & ~(~0ull >> len)), len);
      } else {
        int flag = (int)((w >> (62 - type)) & 1);
        int m = mant_bits((uint64_t)mask20 << 44, n_eb);       /* :1276-1281 exponent of the MASK */
        int k = flag ? m : (m > 8 ? m - 8 : 0);
        int hdr = type + 2;
        uint64_t payload = k ? ((w << hdr) >> (64 - k)) : 0;
        len = hdr + k;
        if (flag) {                                            /* :1461-1474 mask's 12 bits + mantissa 1..m */
          uint64_t top = ((uint64_t)(mask20 >> 8) << 52) | (k ? payload << (52 - k) : 0);
          v = midpoint(top, 12 + k);
        } else {                                               /* :1446-1460 mask's 20 bits + mantissa 9..m */
          uint64_t top = ((uint64_t)mask20 << 44) | (k ? payload << (44 - k) : 0);
          v = midpoint(top, 20 + k);
        }
      }
    }
    r.pos += (uint64_t)len;
    out[i] = v;
    if (ct != DCO_CT_NP) hist_push(&b3, &b2, &b1, v);
  }
  return 0;
}

/* ---- chunk drivers: the reference call sequence of impl/k-means.c:520-533 per chunk ---- */
size_t dco_compress_chunked_f64(int ct, const double* x, size_t n, size_t chunk_elems, double eb, int subtract_min,
                                int fixed_type, uint32_t fixed_mask20,
                                uint8_t* out, uint64_t* offsets, double* chunk_min, double* chunk_medium,
                                int32_t* chunk_type, uint32_t* chunk_crc)
{
  size_t K = n ? (n + chunk_elems - 1) / chunk_elems : 0, total = 0;
  size_t cap = chunk_elems < n ? chunk_elems : n;
  double* tmp = (double*)malloc(sizeof(double) * (cap ? cap : 1));
  offsets[0] = 0;
  for (size_t c = 0; c < K; c++) {
    size_t lo = c * chunk_elems, cn = (lo + chunk_elems <= n) ? chunk_elems : n - lo;
    const double* src = x + lo;
    double mn = 0.0;
    if (subtract_min) { mn = dco_min_subtract_f64(src, cn, tmp); src = tmp; }
    if (chunk_min) chunk_min[c] = mn;
    int type = fixed_type; uint32_t mask20 = fixed_mask20;
    if (ct == DCO_CT_MASK && fixed_type < 0) {
      type = 0;                                   /* caller's initial value, impl/k-means.c:518 */
      double med = dco_med_f64(src, cn, &type);
      mask20 = dco_mask20(med);
      if (chunk_medium) chunk_medium[c] = med;
      if (chunk_type) chunk_type[c] = type;
    }
    size_t b = dco_compress_f64(ct, src, cn, eb, type, mask20, out + total, NULL);
    if (chunk_crc) chunk_crc[c] = dco_crc32(out + total, b);
    total += b;
    offsets[c + 1] = total;
  }
  free(tmp);
  return total;
}

int dco_decompress_chunked_f64(int ct, const uint8_t* s, const uint64_t* offsets, size_t n, size_t chunk_elems,
                               double eb, int add_min, const double* chunk_min, int fixed_type,
                               uint32_t fixed_mask20, const uint32_t* chunk_mask20, const int32_t* chunk_type,
                               double* out)
{
  size_t K = n ? (n + chunk_elems - 1) / chunk_elems : 0;
  for (size_t c = 0; c < K; c++) {
    size_t lo = c * chunk_elems, cn = (lo + chunk_elems <= n) ? chunk_elems : n - lo;
    int type = fixed_type; uint32_t mask20 = fixed_mask20;
    if (ct == DCO_CT_MASK && fixed_type < 0) { type = chunk_type[c]; mask20 = chunk_mask20[c]; }
    int rc = dco_decompress_f64(ct, s + offsets[c], (size_t)(offsets[c + 1] - offsets[c]), cn, eb, type, mask20, out + lo);
    if (rc) return rc;
    if (add_min) for (size_t i = 0; i < cn; i++) out[lo + i] = out[lo + i] + chunk_min[c];  /* impl/k-means.c:597-598 */
  }
  return 0;
}
