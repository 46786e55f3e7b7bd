/* dc_oracle.c -- CPU ORACLE, TEST INFRASTRUCTURE ONLY (see dc_oracle.h).
 *
 * Restates, function by function, the float32 hot path of
 * /root/reference/impl/dataCompression.c.  Every function cites the reference
 * lines it follows.  Written from the behaviour (bit-exact specification in
 * SURVEY.md section 3.5), not from the text: the reference appends one bit per
 * call to a realloc'd byte array and parses one bit per loop trip; this file
 * keeps a 64-bit accumulator and parses one code per trip.
 *
 * Build: make -C oracle   (gcc -O2 -ffp-contract=off: the predictors must not
 * be fused into FMAs, dataCompression.c:3362-3364 evaluates them op by op).
 */
#include "dc_oracle.h"
#include <math.h>
#include <string.h>

static inline uint32_t f2u(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
static inline float u2f(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }

/* ---- to_absErrorBound_binary, dataCompression.c:5512-5522 ------------------ */
int dco_n_eb(double eb)
{
  for (int n = 0; n < 100; n++)
    if (eb >= pow(2.0, -n)) return n;
  return -1; /* reference: falls off the end of a non-void function (undefined) */
}

/* ---- toSmallDataset_float, :3543-3562 -------------------------------------- */
float dco_min_subtract(const float* x, size_t n, float* y)
{
  float mn = x[0];
  for (size_t i = 1; i < n; i++) if (x[i] < mn) mn = x[i];
  for (size_t i = 0; i < n; i++) y[i] = x[i] - mn;
  return mn;
}

/* ---- med_dataset_float, :3593-3620 ------------------------------------------
 * float32 accumulator summed in index order; max scanned with '>';
 * thresholds 2^(add-127), add = 128,192,224,240,248,252,254  ->  type 1..7. */
float dco_med(const float* x, size_t n, int* type)
{
  volatile float total = 0; /* volatile: keep every partial sum rounded to float32 */
  float mx = x[0];
  for (size_t i = 0; i < n; i++) {
    total = total + x[i];
    if (x[i] > mx) mx = x[i];
  }
  static const int expo[7] = { 1, 65, 97, 113, 121, 125, 127 };
  for (int t = 0; t < 7; t++)
    if ((double)mx < ldexp(1.0, expo[t])) { *type = t + 1; break; }
  return total / (float)(int)n; /* :3615 float / int */
}

/* mask[17] of impl/pingpong.c:202-206 = first 17 chars of floattostr(medium) */
uint32_t dco_mask17(float medium) { return f2u(medium) >> 15; }

/* ---- do_crc32, :5524-5534: zlib crc32 one byte at a time == CRC-32/ISO-HDLC - */
uint32_t dco_crc32(const uint8_t* p, size_t bytes)
{
  static uint32_t tab[256];
  static int init = 0;
  if (!init) {
    for (uint32_t i = 0; i < 256; i++) {
      uint32_t c = i;
      for (int k = 0; k < 8; k++) c = (c & 1) ? 0xEDB88320u ^ (c >> 1) : (c >> 1);
      tab[i] = c;
    }
    init = 1;
  }
  uint32_t c = 0xFFFFFFFFu;
  for (size_t i = 0; i < bytes; i++) c = tab[(c ^ p[i]) & 0xFF] ^ (c >> 8);
  return c ^ 0xFFFFFFFFu;
}

/* ---- MSB-first bit appender: add_bit_to_bytes :5456-5489 + bit_set :5492-5510.
 * Byte k holds stream bits 8k..8k+7, bit 8k at mask 0x80; last byte zero padded. */
typedef struct { uint8_t* p; size_t bytes; uint64_t acc; int nacc; } bitw;
static inline void bw_put(bitw* w, uint32_t code, int len) /* len 0..32, code right-aligned */
{
  if (len == 0) return;
  w->acc = (w->acc << len) | (uint64_t)(len == 32 ? code : (code & ((1u << len) - 1u)));
  w->nacc += len;
  while (w->nacc >= 8) { w->nacc -= 8; w->p[w->bytes++] = (uint8_t)(w->acc >> w->nacc); }
}
static inline int bw_finish(bitw* w) /* returns the reference's pos: 8 when aligned, else 8 - (bits mod 8) */
{
  int rem = w->nacc;
  if (rem) { w->p[w->bytes++] = (uint8_t)(w->acc << (8 - rem)); w->nacc = 0; }
  return rem ? 8 - rem : 8;
}

/* m(x) = clamp(n_eb + (E-127), 0, 23), :3485-3503 */
static inline int mant_bits(uint32_t u, int n_eb)
{
  int m = n_eb + (int)((u >> 23) & 0xFF) - 127;
  return m > 23 ? 23 : (m < 0 ? 0 : m);
}

/* raw code R(x): compress_bitwise_float :3479-3520 -- top 9+m bits, pure truncation */
static inline void put_raw(bitw* w, uint32_t u, int n_eb)
{
  int len = 9 + mant_bits(u, n_eb);
  bw_put(w, u >> (32 - len), len);
}

/* M(x): compress_bitwise_float_mask :2143-2284 */
static inline void put_mask(bitw* w, uint32_t u, int n_eb, int type, uint32_t mask17)
{
  int m = mant_bits(u, n_eb);
  if ((u >> 23) != (mask17 >> 8)) { put_raw(w, u, n_eb); return; }   /* :2172-2180, :2277-2283 */
  uint32_t prefix = (1u << type) - 1u;                               /* 0 then `type` ones */
  if (((u >> 15) & 0xFF) == (mask17 & 0xFF)) {                       /* :2208-2221 */
    bw_put(w, prefix, 1 + type);
    bw_put(w, 0, 1);
    int k = m > 8 ? m - 8 : 0;                                       /* chars 17..8+m */
    bw_put(w, (u >> (23 - m)) & ((1u << k) - 1u), k);
  } else {                                                           /* :2222-2266 */
    bw_put(w, prefix, 1 + type);
    bw_put(w, 1, 1);
    bw_put(w, (u >> (23 - m)) & ((1u << m) - 1u), m);                /* chars 9..8+m */
  }
}

/* history update, identical in encoder (:3347-3358, :3386-3388) and decoder (:3110-3127) */
static inline void hist_push(float* b3, float* b2, float* b1, float v)
{
  if (*b3 == -1) *b3 = v;
  else if (*b2 == -1) *b2 = v;
  else if (*b1 == -1) *b1 = v;
  else { *b3 = *b2; *b2 = *b1; *b1 = v; }
}

/* myCompress_bitwise :3310-3444 (CT=5), _np :2645-2654 (CT=6), _op :577-696 (CT=11), _mask :2030-2141 (CT=7) */
size_t dco_compress(int ct, const float* x, size_t n, double eb, int type, uint32_t mask17,
                    uint8_t* out, int* pos_out)
{
  bitw w = { out, 0, 0, 0 };
  int n_eb = dco_n_eb(eb);
  float b1 = -1, b2 = -1, b3 = -1;
  for (size_t i = 0; i < n; i++) {
    float v = x[i];
    uint32_t u = f2u(v);
    if (ct == DCO_CT_NP) { put_raw(&w, u, n_eb); continue; }
    int code = 0; /* 0 = fallback, else the 3-bit code */
    if (b3 == -1 || b2 == -1 || b1 == -1) {                 /* :3330 history not full */
      if (fabs((double)v) < eb) code = 4;                   /* :3333 '100' */
      hist_push(&b3, &b2, &b1, v);                          /* (only the fill branches can fire) */
    } else {
      volatile float p2 = 2 * b1 - b2;                      /* :3363, each op rounded to float32 */
      volatile float t1 = 3 * b1, t2 = 3 * b2;
      volatile float t3 = t1 - t2;
      volatile float p3 = t3 + b3;                          /* :3364 */
      float d1 = fabsf(b1 - v), d2 = fabsf(p2 - v), d3 = fabsf(p3 - v);
      float dmin = d1; int sel = 5;                         /* 'a' 101 */
      if (d2 < dmin) { dmin = d2; sel = 6; }                /* 'b' 110, strict < */
      if (d3 < dmin) { dmin = d3; sel = 7; }                /* 'c' 111 */
      b3 = b2; b2 = b1; b1 = v;                             /* :3386-3388 */
      if (fabs((double)v) < eb) code = 4;                   /* :3390 */
      else if ((double)dmin <= eb) code = sel;              /* :3399 */
    }
    if (code) bw_put(&w, (uint32_t)code, 3);
    else if (ct == DCO_CT_BITWISE) put_raw(&w, u, n_eb);
    else if (ct == DCO_CT_OP) bw_put(&w, u, 32);            /* :689-692 full word */
    else put_mask(&w, u, n_eb, type, mask17);
  }
  int pos = bw_finish(&w);
  if (pos_out) *pos_out = pos;
  return w.bytes;
}

/* ---- bit reader: stream bit k = (s[k/8] >> (7 - k%8)) & 1; bits past the end read as 0 ---- */
typedef struct { const uint8_t* p; size_t bytes; uint64_t pos; } bitr;
static inline uint32_t br_peek32(const bitr* r) /* next 32 bits, left aligned */
{
  uint64_t v = 0;
  size_t b = (size_t)(r->pos >> 3);
  for (int k = 0; k < 5; k++) v = (v << 8) | (b + k < r->bytes ? r->p[b + k] : 0);
  return (uint32_t)(v >> (8 - (r->pos & 7)));
}

/* raw / escape reconstruction: the transmitted top `have` bits, then a single '1', then zeros
 * (midpoint); exact when have == 32.  decompress_bitwise_float :3166-3185, _np :2611-2630. */
static inline float midpoint(uint32_t top, int have)
{
  if (have >= 32) return u2f(top);
  return u2f(top | (0x80000000u >> have));
}

/* myDecompress_bitwise :2922-3135 (CT=5), _np :2459-2609 (CT=6), _op :698-797 (CT=11),
 * _mask :1703-1898 + decompress_bitwise_float_mask :1900-2027 (CT=7).
 *
 * The reference walks the stream bit by bit; this walks it code by code and decodes exactly n codes.
 * Divergences, all in corners where the reference is undefined:
 *  - a 9-bit / zero-payload code that is the LAST code of a byte-aligned stream is decoded here; the
 *    reference only emits it when the next bit arrives (:3022-3025) and leaves out[n-1] uninitialised;
 *  - CT=7: a 3-bit code always wins (reference mis-parses every 3-bit code when type==1, :1866);
 *    a flag-0 escape carries max(m-8,0) payload bits (reference: offset_bits goes <= 0 for 0<m<=8, :1797-1802);
 *    an escape whose total length is 9 bits is emitted (reference blocks it with bits_num != 1+8, :1866);
 *    type==7 escapes are parsed as escapes (reference takes bits_num==9 for a raw code, :1765). */
int dco_decompress(int ct, const uint8_t* s, size_t bytes, size_t n, double eb, int type, uint32_t mask17,
                   float* out)
{
  bitr r = { s, bytes, 0 };
  int n_eb = dco_n_eb(eb);
  float b1 = -1, b2 = -1, b3 = -1;
  for (size_t i = 0; i < n; i++) {
    uint32_t w = br_peek32(&r);
    float v;
    int len;
    if (w >> 31) {                                   /* leading 1 */
      if (ct == DCO_CT_NP) return -1;                /* :2501-2505 exit(1) */
      len = 3;
      switch ((w >> 29) & 3) {                       /* decompress_bitwise_float :3139-3158 */
        case 0: v = 0.0f; break;
        case 1: v = b1; break;
        case 2: { volatile float p = 2 * b1 - b2; v = p; } break;
        default: { volatile float t1 = 3 * b1, t2 = 3 * b2; volatile float t3 = t1 - t2;
                   volatile float p = t3 + b3; v = p; } break;
      }
    } else if (ct == DCO_CT_OP) {                    /* :718-721 leading 0 -> 32 bits */
      len = 32; v = u2f(w);
    } else {
      int escape = 0;
      if (ct == DCO_CT_MASK) {                       /* :1739-1758: bits 1..type all ones -> escape */
        uint32_t ones = (1u << type) - 1u;
        escape = (((w >> (31 - type)) & ones) == ones);
      }
      if (!escape) {                                 /* raw: 9 bits, m from the exponent just read, m more */
        int m = mant_bits(w, n_eb);
        len = 9 + m;
        v = midpoint(w & (len == 32 ? 0xFFFFFFFFu : ~(0xFFFFFFFFu >> len)), len);
      } else {
        int flag = (w >> (30 - type)) & 1;
        int m = mant_bits(mask17 << 15, n_eb);       /* :1769-1771 exponent of the MASK */
        int k = flag ? m : (m > 8 ? m - 8 : 0);
        int hdr = type + 2;
        uint32_t payload = k ? ((uint32_t)((((uint64_t)w << hdr) & 0xFFFFFFFFu)) >> (32 - k)) : 0;
        len = hdr + k;
        if (flag) {                                  /* :1963-2010 mask's 9 bits + mantissa 1..m */
          uint32_t top = ((mask17 >> 8) << 23) | (k ? payload << (23 - k) : 0);
          v = midpoint(top, 9 + k);
        } else {                                     /* :1947-1962 mask's 17 bits + mantissa 9..m */
          uint32_t top = (mask17 << 15) | (k ? payload << (15 - k) : 0);
          v = midpoint(top, 17 + k);
        }
      }
    }
    r.pos += (uint64_t)len;
    out[i] = v;
    if (ct != DCO_CT_NP) hist_push(&b3, &b2, &b1, v);  /* :3110-3127 */
  }
  return 0;
}

/* ---- chunk drivers: "the reference routine run chunk-by-chunk"  ----------
 * per chunk: toSmallDataset_float (impl/pingpong.c:148) -> [med_dataset_float + mask, :199-206]
 *            -> myCompress_bitwise* from (NULL,0,8) -> [do_crc32, impl/pingpong.c:263]. */
#include <stdlib.h>
size_t dco_compress_chunked(int ct, const float* x, size_t n, size_t chunk_elems, double eb, int subtract_min,
                            int fixed_type, uint32_t fixed_mask17,
                            uint8_t* out, uint64_t* offsets, float* chunk_min, float* chunk_medium,
                            int32_t* chunk_type, uint32_t* chunk_crc)
{
  size_t K = n ? (n + chunk_elems - 1) / chunk_elems : 0, total = 0;
  size_t cap = chunk_elems < n ? chunk_elems : n;
  float* tmp = (float*)malloc(sizeof(float) * (cap ? cap : 1));
  offsets[0] = 0;
  for (size_t c = 0; c < K; c++) {
    size_t lo = c * chunk_elems, cn = (lo + chunk_elems <= n) ? chunk_elems : n - lo;
    const float* src = x + lo;
    float mn = 0.0f;
    if (subtract_min) { mn = dco_min_subtract(src, cn, tmp); src = tmp; }
    if (chunk_min) chunk_min[c] = mn;
    int type = fixed_type; uint32_t mask17 = fixed_mask17;
    if (ct == DCO_CT_MASK && fixed_type < 0) {
      type = 0;                                   /* caller's initial value, impl/pingpong.c:199 */
      float med = dco_med(src, cn, &type);
      mask17 = dco_mask17(med);
      if (chunk_medium) chunk_medium[c] = med;
      if (chunk_type) chunk_type[c] = type;
    }
    size_t b = dco_compress(ct, src, cn, eb, type, mask17, out + total, NULL);
    if (chunk_crc) chunk_crc[c] = dco_crc32(out + total, b);
    total += b;
    offsets[c + 1] = total;
  }
  free(tmp);
  return total;
}

int dco_decompress_chunked(int ct, const uint8_t* s, const uint64_t* offsets, size_t n, size_t chunk_elems,
                           double eb, int add_min, const float* chunk_min, int fixed_type,
                           uint32_t fixed_mask17, const float* chunk_medium, const int32_t* chunk_type,
                           float* out)
{
  size_t K = n ? (n + chunk_elems - 1) / chunk_elems : 0;
  for (size_t c = 0; c < K; c++) {
    size_t lo = c * chunk_elems, cn = (lo + chunk_elems <= n) ? chunk_elems : n - lo;
    int type = fixed_type; uint32_t mask17 = fixed_mask17;
    if (ct == DCO_CT_MASK && fixed_type < 0) { type = chunk_type[c]; mask17 = dco_mask17(chunk_medium[c]); }
    int rc = dco_decompress(ct, s + offsets[c], (size_t)(offsets[c + 1] - offsets[c]), cn, eb, type, mask17, out + lo);
    if (rc) return rc;
    if (add_min) for (size_t i = 0; i < cn; i++) out[lo + i] = out[lo + i] + chunk_min[c];  /* impl/pingpong.c:562 */
  }
  return 0;
}
