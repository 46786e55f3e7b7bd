// dcb_f64.cuh -- float64 variants of the bit-wise codecs: the reference's *_double functions
// (impl/dataCompression.c: myCompress_bitwise_double :3189, _np :2633, _op :355, _mask :1590; myDecompress_bitwise_double
// :2656, _np :2286, _op :476, _mask :1199; toSmallDataset_double :3522; med_dataset_double :3564) behind dcb_compress_f64 /
// dcb_decompress_f64.  Same stream rules as float32 with the IEEE double layout: raw code = 12 + m bits,
// m = clamp(n_eb + E - 1023, 0, 52); CT=7 mask = the first 1+11+8 = 20 bits of the chunk mean; predictors in double, every
// operation rounded on its own.
//
// First version of this row: correct and parallel, not yet tuned like the float32 path.
//   stats  : one block per chunk -- min (first occurrence wins, as the reference's '<' scan), then max / tree sum of
//            x - min; the 20 mask bits come from the tree sum when its error bracket cannot change them, else from the exact
//            sequential sum (one thread).
//   length : one block per 4096-element tile -- classify every value, sum the code lengths.
//   scan   : one block -- chunk byte offsets and every tile's bit offset.
//   pack   : one block per tile -- classify again, bit offsets by a block scan, codes OR-ed into a shared-memory image of
//            the tile's words, image copied out (the two boundary words with atomicOr into the zeroed stream).
//   decode : one thread per chunk (a chunk is sequential by construction: predictors read the decoded history).
#pragma once
#include "dcb_common.cuh"

#define F64_TILE 4096u
#define F64_NT 256
#define F64_PER 16
#define F64_PITCH 17          // doubles per thread slot in the staged tile (16 + 1: conflict-free column reads)

struct F64Const { double eb; int n_eb; };

struct F64Args {
  const double* in; uint64_t n; uint64_t chunk;        // chunk = elements per chunk (n for a single chunk)
  uint32_t num_tiles; uint32_t tiles_per_chunk;        // every chunk has tiles_per_chunk tiles (the last one may have fewer)
  uint64_t num_chunks;
  uint32_t flags; int fixed_type; uint32_t fixed_mask20;
  F64Const k;
  uint8_t* out; uint64_t* offsets; double* cmin; uint32_t* cmask; int32_t* ctype;
  unsigned long long* tile_bits;                       // [num_tiles]: code bits of the tile, then (after the scan) its bit offset
  unsigned long long* last_bits;                       // code bits of the LAST chunk (the legacy callers' `pos`)
};

__device__ __forceinline__ int f64_mant_bits(uint64_t u, int n_eb) {
  const int m = n_eb + (int)((u >> 52) & 0x7FFu) - 1023;
  return min(max(m, 0), 52);
}

// compress_bitwise_double :3446-3477
__device__ __forceinline__ void f64_raw(uint64_t u, int n_eb, uint64_t& code, int& len) {
  len = 12 + f64_mant_bits(u, n_eb);
  code = u >> (64 - len);
}

// compress_bitwise_double_mask :1493-1588
__device__ __forceinline__ void f64_mask(uint64_t u, int n_eb, int type, uint32_t mask20, uint64_t& code, int& len) {
  if ((uint32_t)(u >> 52) != (mask20 >> 8)) { f64_raw(u, n_eb, code, len); return; }
  const int m = f64_mant_bits(u, n_eb);
  const uint64_t prefix = (1ull << type) - 1ull;
  const bool same8 = ((uint32_t)(u >> 44) & 0xFFu) == (mask20 & 0xFFu);
  const int k = same8 ? max(m - 8, 0) : m;
  const uint64_t payload = k ? ((u >> (52 - m)) & ((1ull << k) - 1ull)) : 0ull;
  code = (((prefix << 1) | (same8 ? 0ull : 1ull)) << k) | payload;
  len = type + 2 + k;
}

// one element of myCompress_bitwise_double{,_np,_op,_mask}: the exact -1 "empty slot" state machine (:3204-3222, :3260-3262)
template <int CT>
__device__ __forceinline__ void f64_classify(double v, double& b3, double& b2, double& b1, const F64Const& k, int type,
                                             uint32_t mask20, uint64_t& code, int& len) {
  const uint64_t u = (uint64_t)__double_as_longlong(v);
  if (CT == CT_NP) { f64_raw(u, k.n_eb, code, len); return; }
  int c3 = 0;
  if (b3 == -1.0 || b2 == -1.0 || b1 == -1.0) {
    if (fabs(v) < k.eb) c3 = 4;
    if (b3 == -1.0) b3 = v; else if (b2 == -1.0) b2 = v; else if (b1 == -1.0) b1 = v;
  } else {
    const double p2 = __dsub_rn(__dmul_rn(2.0, b1), b2);
    const double p3 = __dadd_rn(__dsub_rn(__dmul_rn(3.0, b1), __dmul_rn(3.0, b2)), b3);
    const double d1 = fabs(__dsub_rn(b1, v)), d2 = fabs(__dsub_rn(p2, v)), d3 = fabs(__dsub_rn(p3, v));
    double dmin = d1; int sel = 5;
    if (d2 < dmin) { dmin = d2; sel = 6; }
    if (d3 < dmin) { dmin = d3; sel = 7; }
    b3 = b2; b2 = b1; b1 = v;
    if (fabs(v) < k.eb) c3 = 4;
    else if (dmin <= k.eb) c3 = sel;
  }
  if (c3) { code = (uint64_t)c3; len = 3; }
  else if (CT == CT_BITWISE) f64_raw(u, k.n_eb, code, len);
  else if (CT == CT_OP) { code = u; len = 64; }
  else f64_mask(u, k.n_eb, type, mask20, code, len);
}

// tile geometry: tile t -> chunk, first element, count
__device__ __forceinline__ void f64_tile_geom(const F64Args& a, uint32_t t, uint64_t& c, uint64_t& chunk_lo, uint64_t& e0,
                                              uint32_t& cnt) {
  c = t / a.tiles_per_chunk;
  const uint32_t j = t - (uint32_t)c * a.tiles_per_chunk;
  chunk_lo = c * a.chunk;
  const uint64_t chunk_hi = min(a.n, chunk_lo + a.chunk);
  e0 = chunk_lo + (uint64_t)j * F64_TILE;
  cnt = (uint32_t)min((uint64_t)F64_TILE, chunk_hi - e0);
}

// the history a thread starts from at chunk-local position e0 - chunk_lo: the reference's state after the elements before
// it = the last three values that are not the -1 marker, with b1 = -1 if the value just before was the marker
__device__ __forceinline__ void f64_entry_state(const double* __restrict__ x, uint64_t chunk_lo, uint64_t g, double mn,
                                                bool sub, double& b3, double& b2, double& b1) {
  double r[3]; int cnt = 0; bool last_sent = false;
  for (uint64_t i = g; i > chunk_lo && cnt < 3; ) {
    --i;
    const double v = sub ? __dsub_rn(x[i], mn) : x[i];
    if (i + 1 == g) last_sent = (v == -1.0);
    if (v != -1.0) r[cnt++] = v;
  }
  b3 = b2 = b1 = -1.0;
  if (cnt == 3) { if (last_sent) { b3 = r[1]; b2 = r[0]; } else { b3 = r[2]; b2 = r[1]; b1 = r[0]; } }
  else if (cnt == 2) { b3 = r[1]; b2 = r[0]; }
  else if (cnt == 1) { b3 = r[0]; }
}

// ---- stats: toSmallDataset_double (:3522) + med_dataset_double (:3564) per chunk ------------------------------------
struct F64StatsArgs {
  const double* in; uint64_t n; uint64_t chunk; int subtract_min; int want_mask;
  double* cmin; uint32_t* cmask; int32_t* ctype; double* cmedium;   // cmedium may be NULL
};

__device__ __forceinline__ bool f64_min_better(double av, uint64_t ai, double bv, uint64_t bi) {
  // sequential scan with '<' from data[0]: a NaN never replaces, an equal value (also -0 == +0) never replaces
  if (av != av) return false;
  if (bv != bv) return true;
  return (av < bv) || (av == bv && ai < bi);
}

__global__ void __launch_bounds__(F64_NT) dcb_f64_stats_kernel(F64StatsArgs s) {
  __shared__ double sh_v[F64_NT / 32]; __shared__ unsigned long long sh_i[F64_NT / 32];
  __shared__ double sh_a[F64_NT / 32], sh_b[F64_NT / 32]; __shared__ double sh_mn;
  const uint64_t c = blockIdx.x, lo = c * s.chunk, hi = min(s.n, lo + s.chunk);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const double* x = s.in;
  double mn = 0.0;
  if (s.subtract_min) {
    double bv = __longlong_as_double(0x7FF8000000000000ll); uint64_t bi = ~0ull;
    for (uint64_t i = lo + 1 + tid; i < hi; i += F64_NT) { const double v = x[i]; if (f64_min_better(v, i, bv, bi)) { bv = v; bi = i; } }
    for (int o = 16; o; o >>= 1) {
      const double ov = __shfl_xor_sync(0xffffffffu, bv, o); const uint64_t oi = __shfl_xor_sync(0xffffffffu, bi, o);
      if (f64_min_better(ov, oi, bv, bi)) { bv = ov; bi = oi; }
    }
    if (lane == 0) { sh_v[warp] = bv; sh_i[warp] = bi; }
    __syncthreads();
    if (tid == 0) {
      double m = x[lo];                                   // data[0]; a NaN here stays (every '<' is false)
      if (m == m) {
        uint64_t mi = lo;
        for (int w = 0; w < F64_NT / 32; w++) if (f64_min_better(sh_v[w], sh_i[w], m, mi)) { m = sh_v[w]; mi = sh_i[w]; }
      }
      sh_mn = m; s.cmin[c] = m;
    }
    __syncthreads();
    mn = sh_mn;
  }
  if (!s.want_mask) return;
  const bool sub = s.subtract_min != 0;
  // max with '>' from data[0] (NaN never replaces; a NaN data[0] stays), tree sum and sum of magnitudes
  double mx = __longlong_as_double(0x7FF8000000000000ll), sum = 0.0, mag = 0.0;
  for (uint64_t i = lo + tid; i < hi; i += F64_NT) {
    const double y = sub ? __dsub_rn(x[i], mn) : x[i];
    sum = __dadd_rn(sum, y); mag = __dadd_rn(mag, fabs(y));
    if (y == y && (mx != mx || y > mx)) mx = y;
  }
  for (int o = 16; o; o >>= 1) {
    sum = __dadd_rn(sum, __shfl_xor_sync(0xffffffffu, sum, o)); mag = __dadd_rn(mag, __shfl_xor_sync(0xffffffffu, mag, o));
    const double om = __shfl_xor_sync(0xffffffffu, mx, o);
    if (om == om && (mx != mx || om > mx)) mx = om;
  }
  if (lane == 0) { sh_v[warp] = mx; sh_a[warp] = sum; sh_b[warp] = mag; }
  __syncthreads();
  if (tid == 0) {
    double M = sh_v[0], S = sh_a[0], G = sh_b[0];
    for (int w = 1; w < F64_NT / 32; w++) {
      S = __dadd_rn(S, sh_a[w]); G = __dadd_rn(G, sh_b[w]);
      if (sh_v[w] == sh_v[w] && (M != M || sh_v[w] > M)) M = sh_v[w];
    }
    const double y0 = sub ? __dsub_rn(x[lo], mn) : x[lo];
    if (y0 != y0) M = y0;                                 // max = data[0] = NaN: every later '>' is false
    int type = 0, add = 0;                                // :3576-3585; 0 = "left at the caller's initial value"
    for (int i = 10; i > 0; i--) {
      add += 1 << i;
      if (M < __longlong_as_double((long long)add << 52)) { type = 11 - i; break; }
    }
    s.ctype[c] = type;
    const double cnt = (double)(int)(hi - lo);
    const double E = __dmul_rn(__dmul_rn(cnt, 2.220446049250313e-16), G);      // n * 2^-52 * sum|y| >= |seq - tree|
    const double mlo = __ddiv_rn(__dsub_rn(S, E), cnt), mhi = __ddiv_rn(__dadd_rn(S, E), cnt);
    const uint32_t klo = (uint32_t)((uint64_t)__double_as_longlong(mlo) >> 44), khi = (uint32_t)((uint64_t)__double_as_longlong(mhi) >> 44);
    const bool decided = (klo == khi) && (G == G) && (G < 1.0e300) && s.cmedium == nullptr;
    if (decided) s.cmask[c] = klo;
    if (!decided) {                                        // exact: total += data[i] in index order (:3568-3571)
      double total = 0.0;
      for (uint64_t i = lo; i < hi; i++) total = __dadd_rn(total, sub ? __dsub_rn(x[i], mn) : x[i]);
      const double med = __ddiv_rn(total, cnt);
      s.cmask[c] = (uint32_t)((uint64_t)__double_as_longlong(med) >> 44);
      if (s.cmedium) s.cmedium[c] = med;
    }
  }
}

// ---- classification of one tile: thread t owns elements [16t, 16t+16) of the tile -------------------------------------
// stage: the tile's values (min already subtracted) in shared memory, slot pitch 17 doubles
template <int CT>
__device__ __forceinline__ uint32_t f64_tile_codes(const F64Args& a, double* stage, uint64_t chunk_lo, uint64_t e0, uint32_t cnt,
                                                   double mn, int type, uint32_t mask20, uint64_t (&code)[F64_PER],
                                                   int (&len)[F64_PER]) {
  const int tid = threadIdx.x;
  const bool sub = (a.flags & DCB_F_SUBTRACT_MIN) != 0;
  for (uint32_t e = tid; e < cnt; e += F64_NT) {                                 // coalesced
    const double v = a.in[e0 + e];
    stage[(e >> 4) * F64_PITCH + (e & 15)] = sub ? __dsub_rn(v, mn) : v;
  }
  __syncthreads();
  const uint32_t my0 = (uint32_t)tid * F64_PER;
  uint32_t bits = 0;
  double b3 = -1.0, b2 = -1.0, b1 = -1.0;
  if (CT != CT_NP && my0 < cnt) f64_entry_state(a.in, chunk_lo, e0 + my0, mn, sub, b3, b2, b1);
  #pragma unroll
  for (int j = 0; j < F64_PER; j++) {
    code[j] = 0; len[j] = 0;
    if (my0 + j < cnt) {
      f64_classify<CT>(stage[tid * F64_PITCH + j], b3, b2, b1, a.k, type, mask20, code[j], len[j]);
      bits += (uint32_t)len[j];
    }
  }
  return bits;
}

__device__ __forceinline__ void f64_chunk_params(const F64Args& a, uint64_t c, double& mn, int& type, uint32_t& mask20) {
  mn = (a.flags & DCB_F_SUBTRACT_MIN) ? a.cmin[c] : 0.0;
  type = a.fixed_type; mask20 = a.fixed_mask20;
  if (a.cmask && !(a.flags & DCB_F_FIXED_MASK)) { type = a.ctype[c]; mask20 = a.cmask[c]; }
}

template <int CT>
__global__ void __launch_bounds__(F64_NT) dcb_f64_length_kernel(F64Args a) {
  __shared__ double stage[F64_NT * F64_PITCH];
  __shared__ uint32_t part[F64_NT / 32];
  uint64_t c, chunk_lo, e0; uint32_t cnt;
  f64_tile_geom(a, blockIdx.x, c, chunk_lo, e0, cnt);
  double mn; int type; uint32_t mask20;
  f64_chunk_params(a, c, mn, type, mask20);
  uint64_t code[F64_PER]; int len[F64_PER];
  uint32_t bits = f64_tile_codes<CT>(a, stage, chunk_lo, e0, cnt, mn, type, mask20, code, len);
  bits = __reduce_add_sync(0xffffffffu, bits);
  if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = bits;
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t t = 0;
    for (int w = 0; w < F64_NT / 32; w++) t += part[w];
    a.tile_bits[blockIdx.x] = t;
  }
}

// ---- scan: chunk byte offsets and tile bit offsets; one block, thread i owns a contiguous run of chunks ----------------
__global__ void __launch_bounds__(1024) dcb_f64_scan_kernel(F64Args a) {
  __shared__ unsigned long long base[1024];
  const uint64_t K = a.num_chunks, per = (K + 1023) / 1024;
  const uint64_t c0 = min(K, (uint64_t)threadIdx.x * per), c1 = min(K, c0 + per);
  unsigned long long mine = 0;
  for (uint64_t c = c0; c < c1; c++) {
    unsigned long long b = 0;
    const uint64_t t0 = c * a.tiles_per_chunk, t1 = min((uint64_t)a.num_tiles, t0 + a.tiles_per_chunk);
    for (uint64_t t = t0; t < t1; t++) b += a.tile_bits[t];
    mine += (b + 7) >> 3;
  }
  base[threadIdx.x] = mine;
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned long long run = 0;
    for (int i = 0; i < 1024; i++) { const unsigned long long v = base[i]; base[i] = run; run += v; }
    a.offsets[K] = run;
  }
  __syncthreads();
  unsigned long long off = base[threadIdx.x];
  for (uint64_t c = c0; c < c1; c++) {
    a.offsets[c] = off;
    unsigned long long bit = off * 8ull;
    const uint64_t t0 = c * a.tiles_per_chunk, t1 = min((uint64_t)a.num_tiles, t0 + a.tiles_per_chunk);
    for (uint64_t t = t0; t < t1; t++) { const unsigned long long b = a.tile_bits[t]; a.tile_bits[t] = bit; bit += b; }
    if (c + 1 == K) *a.last_bits = bit - off * 8ull;
    off += (bit - off * 8ull + 7) >> 3;
  }
}

// zero the words the pack kernel ORs into: [0, total + 8) rounded to words
__global__ void dcb_f64_zero_kernel(uint32_t* __restrict__ out, const uint64_t* __restrict__ total_bytes) {
  const uint64_t words = (*total_bytes + 11) >> 2;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < words; i += (uint64_t)gridDim.x * blockDim.x) out[i] = 0u;
}

// ---- pack --------------------------------------------------------------------------------------------------------------
#define F64_IMG_WORDS (F64_TILE * 2 + 8)
template <int CT>
__global__ void __launch_bounds__(F64_NT) dcb_f64_pack_kernel(F64Args a) {
  __shared__ __align__(16) unsigned char smem_raw[F64_NT * F64_PITCH * 8];          // staged tile, then the word image
  __shared__ uint32_t part[F64_NT / 32];
  static_assert(F64_IMG_WORDS * 4 <= F64_NT * F64_PITCH * 8, "image must fit the staging buffer");
  double* stage = reinterpret_cast<double*>(smem_raw);
  uint32_t* img = reinterpret_cast<uint32_t*>(smem_raw);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  uint64_t c, chunk_lo, e0; uint32_t cnt;
  f64_tile_geom(a, blockIdx.x, c, chunk_lo, e0, cnt);
  double mn; int type; uint32_t mask20;
  f64_chunk_params(a, c, mn, type, mask20);
  uint64_t code[F64_PER]; int len[F64_PER];
  const uint32_t bits = f64_tile_codes<CT>(a, stage, chunk_lo, e0, cnt, mn, type, mask20, code, len);
  uint32_t incl = bits;
  #pragma unroll
  for (int o = 1; o < 32; o <<= 1) { const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += v; }
  if (lane == 31) part[warp] = incl;
  __syncthreads();                                                                 // also: every thread has read its stage slot
  uint32_t wbase = 0, tile_bits = 0;
  #pragma unroll
  for (int w = 0; w < F64_NT / 32; w++) { const uint32_t v = part[w]; if (w < warp) wbase += v; tile_bits += v; }
  const unsigned long long bit0 = a.tile_bits[blockIdx.x];                         // global bit offset of the tile
  const uint32_t phase = (uint32_t)(bit0 & 31ull);
  const uint32_t nwords = (phase + tile_bits + 31u) >> 5;
  for (uint32_t w = tid; w < nwords; w += F64_NT) img[w] = 0u;
  __syncthreads();
  uint32_t pos = phase + wbase + incl - bits;
  #pragma unroll
  for (int j = 0; j < F64_PER; j++) {
    if (len[j]) {
      const uint64_t left = code[j] << (64 - len[j]);                              // left aligned
      const uint32_t o = pos & 31u, wi = pos >> 5;
      const uint64_t hi = left >> o;
      const uint32_t lo = o ? (uint32_t)((left << (64 - o)) >> 32) : 0u;
      const uint32_t w0 = (uint32_t)(hi >> 32), w1 = (uint32_t)hi;
      if (w0) atomicOr(&img[wi], w0);
      if (w1) atomicOr(&img[wi + 1], w1);
      if (lo) atomicOr(&img[wi + 2], lo);
      pos += (uint32_t)len[j];
    }
  }
  __syncthreads();
  uint32_t* gw = reinterpret_cast<uint32_t*>(a.out) + (bit0 >> 5);
  for (uint32_t w = tid; w < nwords; w += F64_NT) {
    const uint32_t v = __byte_perm(img[w], 0u, 0x0123);                            // stream bytes are MSB first
    if (w == 0 || w + 1 == nwords) { if (v) atomicOr(gw + w, v); }
    else gw[w] = v;
  }
}

// ---- fused encoder for "one chunk == one 4096-element tile" (r2) --------------------------------------
// stats + length + pack read the input three times (r1: 0.21 + 0.41 + 0.74 ms for 2^27 doubles, plus a one-block scan).  A chunk
// of 4096 doubles is 32 KB: ONE block keeps it in shared memory from the first load to the packed image -- minimum, [max, mean
// bracket, exact mean], classification, bit offsets, packing -- and writes the image to a fixed-pitch scratch slot; the chunk
// table and the final placement are the float32 path's dcb_compact_kernel (which also computes the CRC while it copies).
#define F64_SLOT_PITCH 32768u        // bytes per scratch slot: 4096 codes of at most 64 bits

struct F64FusedArgs {
  F64Args a;
  uint8_t* slots; uint32_t* sizes; uint32_t* gsum; uint32_t* last_tile_bits;
  double* cmedium;                   // exact mean wanted (DCB_F_EXACT_MEDIUM), else NULL
  int want_mask;
};

#define F64_AT(stage, e) (stage)[((e) >> 4) * F64_PITCH + ((e) & 15)]

template <int CT>
__global__ void __launch_bounds__(F64_NT) dcb_f64_fused_kernel(const F64FusedArgs f) {
  __shared__ __align__(16) unsigned char smem_raw[F64_NT * F64_PITCH * 8];          // the chunk, then the word image
  __shared__ double sh_v[F64_NT / 32]; __shared__ unsigned long long sh_i[F64_NT / 32];
  __shared__ double sh_a[F64_NT / 32], sh_b[F64_NT / 32]; __shared__ double sh_mn;
  __shared__ uint32_t sh_mask; __shared__ int sh_type;
  __shared__ uint32_t part[F64_NT / 32];
  const F64Args& a = f.a;
  double* stage = reinterpret_cast<double*>(smem_raw);
  uint32_t* img = reinterpret_cast<uint32_t*>(smem_raw);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint64_t c = blockIdx.x, lo = c * a.chunk;
  const uint32_t cnt = (uint32_t)min((uint64_t)F64_TILE, a.n - lo);
  const bool sub = (a.flags & DCB_F_SUBTRACT_MIN) != 0;
  for (uint32_t e = tid; e < cnt; e += F64_NT) F64_AT(stage, e) = a.in[lo + e];       // the only read of the input
  __syncthreads();
  // ---- toSmallDataset_double (:3522): first minimum with '<' from data[0] (dcb_f64_stats_kernel's rules) --------------
  double mn = 0.0;
  if (sub) {
    // plain minimum first (fmin ignores NaN: a NaN never replaces, :3524-3530); the reference's "first minimal element wins" only
    // shows when the minimum is a zero (+0 == -0): then the first zero's sign is looked up
    double bv = __longlong_as_double(0x7FF8000000000000ll);
    for (uint32_t e = 1 + tid; e < cnt; e += F64_NT) bv = fmin(bv, F64_AT(stage, e));
    for (int o = 16; o; o >>= 1) bv = fmin(bv, __shfl_xor_sync(0xffffffffu, bv, o));
    if (lane == 0) sh_v[warp] = bv;
    __syncthreads();
    double m1 = sh_v[0];
    #pragma unroll
    for (int w = 1; w < F64_NT / 32; w++) m1 = fmin(m1, sh_v[w]);
    if (m1 == 0.0) {                                      // rare: which zero comes first?
      unsigned long long fi = ~0ull;
      for (uint32_t e = 1 + tid; e < cnt; e += F64_NT) if (F64_AT(stage, e) == 0.0) { fi = e; break; }
      for (int o = 16; o; o >>= 1) fi = min(fi, __shfl_xor_sync(0xffffffffu, fi, o));
      __syncthreads();
      if (lane == 0) sh_i[warp] = fi;
      __syncthreads();
      fi = sh_i[0];
      #pragma unroll
      for (int w = 1; w < F64_NT / 32; w++) fi = min(fi, sh_i[w]);
      m1 = F64_AT(stage, (uint32_t)fi);
    }
    if (tid == 0) {
      double m = F64_AT(stage, 0);                        // data[0]; a NaN here stays (every '<' is false)
      if (m == m && m1 == m1 && m1 < m) m = m1;           // strict '<': an equal value (also -0 == +0) never replaces
      sh_mn = m; a.cmin[c] = m;
    }
    __syncthreads();
    mn = sh_mn;
    for (uint32_t e = tid; e < cnt; e += F64_NT) F64_AT(stage, e) = __dsub_rn(F64_AT(stage, e), mn);
    __syncthreads();
  }
  // ---- med_dataset_double (:3564): type from max, 20 mask bits of the sequential mean ------------------------------------
  int type = a.fixed_type; uint32_t mask20 = a.fixed_mask20;
  if (f.want_mask) {
    double mx = __longlong_as_double(0x7FF8000000000000ll), sum = 0.0, mag = 0.0;
    for (uint32_t e = tid; e < cnt; e += F64_NT) {
      const double y = F64_AT(stage, e);
      sum = __dadd_rn(sum, y); mag = __dadd_rn(mag, fabs(y));
      if (y == y && (mx != mx || y > mx)) mx = y;
    }
    for (int o = 16; o; o >>= 1) {
      sum = __dadd_rn(sum, __shfl_xor_sync(0xffffffffu, sum, o)); mag = __dadd_rn(mag, __shfl_xor_sync(0xffffffffu, mag, o));
      const double om = __shfl_xor_sync(0xffffffffu, mx, o);
      if (om == om && (mx != mx || om > mx)) mx = om;
    }
    if (lane == 0) { sh_v[warp] = mx; sh_a[warp] = sum; sh_b[warp] = mag; }
    __syncthreads();
    if (tid == 0) {
      double M = sh_v[0], S = sh_a[0], G = sh_b[0];
      for (int w = 1; w < F64_NT / 32; w++) {
        S = __dadd_rn(S, sh_a[w]); G = __dadd_rn(G, sh_b[w]);
        if (sh_v[w] == sh_v[w] && (M != M || sh_v[w] > M)) M = sh_v[w];
      }
      const double y0 = F64_AT(stage, 0);
      if (y0 != y0) M = y0;                               // max = data[0] = NaN: every later '>' is false
      int ty = 0, add = 0;                                // :3576-3585; 0 = "left at the caller's initial value"
      for (int i = 10; i > 0; i--) {
        add += 1 << i;
        if (M < __longlong_as_double((long long)add << 52)) { ty = 11 - i; break; }
      }
      const double fc = (double)(int)cnt;
      const double E = __dmul_rn(__dmul_rn(fc, 2.220446049250313e-16), G);       // n * 2^-52 * sum|y| >= |seq - tree|
      const double mlo = __ddiv_rn(__dsub_rn(S, E), fc), mhi = __ddiv_rn(__dadd_rn(S, E), fc);
      const uint32_t klo = (uint32_t)((uint64_t)__double_as_longlong(mlo) >> 44), khi = (uint32_t)((uint64_t)__double_as_longlong(mhi) >> 44);
      uint32_t mk = klo;
      if (!((klo == khi) && (G == G) && (G < 1.0e300) && f.cmedium == nullptr)) {   // exact: total += data[i] in index order
        double total = 0.0;
        for (uint32_t e = 0; e < cnt; e++) total = __dadd_rn(total, F64_AT(stage, e));
        const double med = __ddiv_rn(total, fc);
        mk = (uint32_t)((uint64_t)__double_as_longlong(med) >> 44);
        if (f.cmedium) f.cmedium[c] = med;
      }
      sh_mask = mk; sh_type = ty;
      a.cmask[c] = mk; a.ctype[c] = ty;
    }
    __syncthreads();
    type = sh_type; mask20 = sh_mask;
  }
  // ---- classification: thread t owns elements [16t, 16t+16) of the chunk (f64_tile_codes without its load) ---------------
  uint64_t code[F64_PER]; int len[F64_PER];
  const uint32_t my0 = (uint32_t)tid * F64_PER;
  uint32_t bits = 0;
  // CLEAN chunk: every value >= 0 and no NaN -- none can be the -1.0 "empty slot" marker, so the reference's history state
  // machine (:3204-3222) is "the three values before this one" and is read straight from the staged chunk
  bool dirty = false;
  for (uint32_t e = tid; e < cnt; e += F64_NT) dirty = dirty || !(F64_AT(stage, e) >= 0.0);
  const bool clean = (CT != CT_NP) && !__syncthreads_or(dirty);
  if (clean) {
    double b3 = 0.0, b2 = 0.0, b1 = 0.0;
    if (tid > 0) { b3 = stage[(tid - 1) * F64_PITCH + 13]; b2 = stage[(tid - 1) * F64_PITCH + 14]; b1 = stage[(tid - 1) * F64_PITCH + 15]; }
    #pragma unroll
    for (int j = 0; j < F64_PER; j++) {
      code[j] = 0; len[j] = 0;
      if (my0 + j < cnt) {
        const double v = stage[tid * F64_PITCH + j];
        const uint64_t u = (uint64_t)__double_as_longlong(v);
        const double p2 = __dsub_rn(__dmul_rn(2.0, b1), b2);
        const double p3 = __dadd_rn(__dsub_rn(__dmul_rn(3.0, b1), __dmul_rn(3.0, b2)), b3);
        const double d1 = fabs(__dsub_rn(b1, v)), d2 = fabs(__dsub_rn(p2, v)), d3 = fabs(__dsub_rn(p3, v));
        double dmin = d1; int sel = 5;
        if (d2 < dmin) { dmin = d2; sel = 6; }
        if (d3 < dmin) { dmin = d3; sel = 7; }
        int c3 = 0;
        if (fabs(v) < a.k.eb) c3 = 4;
        else if (my0 + j >= 3u && dmin <= a.k.eb) c3 = sel;            // elements 0..2: the history is not full yet (:3204)
        uint64_t cd; int ln;
        if (CT == CT_BITWISE) f64_raw(u, a.k.n_eb, cd, ln);
        else if (CT == CT_OP) { cd = u; ln = 64; }
        else f64_mask(u, a.k.n_eb, type, mask20, cd, ln);
        code[j] = c3 ? (uint64_t)c3 : cd;
        len[j] = c3 ? 3 : ln;
        bits += (uint32_t)len[j];
        b3 = b2; b2 = b1; b1 = v;
      }
    }
  } else {
    double b3 = -1.0, b2 = -1.0, b1 = -1.0;
    if (CT != CT_NP && my0 < cnt) f64_entry_state(a.in, lo, lo + my0, mn, sub, b3, b2, b1);
    #pragma unroll
    for (int j = 0; j < F64_PER; j++) {
      code[j] = 0; len[j] = 0;
      if (my0 + j < cnt) {
        f64_classify<CT>(stage[tid * F64_PITCH + j], b3, b2, b1, a.k, type, mask20, code[j], len[j]);
        bits += (uint32_t)len[j];
      }
    }
  }
  uint32_t incl = bits;
  #pragma unroll
  for (int o = 1; o < 32; o <<= 1) { const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += v; }
  if (lane == 31) part[warp] = incl;
  __syncthreads();                                                                 // also: every thread has read its stage slot
  uint32_t wbase = 0, chunk_bits = 0;
  #pragma unroll
  for (int w = 0; w < F64_NT / 32; w++) { const uint32_t v = part[w]; if (w < warp) wbase += v; chunk_bits += v; }
  const uint32_t nwords = (chunk_bits + 31u) >> 5;
  for (uint32_t w = tid; w < nwords + 4u; w += F64_NT) img[w] = 0u;
  __syncthreads();
  uint32_t pos = wbase + incl - bits;                                              // a slot starts its chunk: no phase
  #pragma unroll
  for (int j = 0; j < F64_PER; j++) {
    if (len[j]) {
      const uint64_t left = code[j] << (64 - len[j]);                              // left aligned
      const uint32_t o = pos & 31u, wi = pos >> 5;
      const uint64_t hi = left >> o;
      const uint32_t lw = o ? (uint32_t)((left << (64 - o)) >> 32) : 0u;
      const uint32_t w0 = (uint32_t)(hi >> 32), w1 = (uint32_t)hi;
      if (w0) atomicOr(&img[wi], w0);
      if (w1) atomicOr(&img[wi + 1], w1);
      if (lw) atomicOr(&img[wi + 2], lw);
      pos += (uint32_t)len[j];
    }
  }
  __syncthreads();
  // ---- the image leaves for the chunk's slot: MSB-first bytes, aligned 128-bit stores -------------------------------------
  uint4* o4 = reinterpret_cast<uint4*>(f.slots + (uint64_t)c * F64_SLOT_PITCH);
  const uint32_t units = (nwords + 3u) >> 2;
  for (uint32_t u = tid; u < units; u += F64_NT) {
    uint4 v = *reinterpret_cast<const uint4*>(img + 4u * u);
    v.x = __byte_perm(v.x, 0u, 0x0123); v.y = __byte_perm(v.y, 0u, 0x0123);
    v.z = __byte_perm(v.z, 0u, 0x0123); v.w = __byte_perm(v.w, 0u, 0x0123);
    o4[u] = v;
  }
  if (tid == 0) {
    const uint32_t bytes = (chunk_bits + 7u) >> 3;
    f.sizes[c] = bytes;
    atomicAdd(&f.gsum[c / 64u], bytes);
    if (c + 1 == a.num_chunks) *f.last_tile_bits = chunk_bits;
  }
}

#ifndef F64_DEC_PREFETCH
#define F64_DEC_PREFETCH 32      // stream words (0: off)
#endif
#ifndef F64_DEC_LANES
#define F64_DEC_LANES 16     // measured 2^27 doubles smooth / noisy: 32 lanes 0.431 / 1.978 ms, 16: 0.403 / 1.741, 8: 0.491 / 2.185
#endif
// ---- decode: one thread per chunk ---------------------------------------------------------------------------------------
struct F64DecArgs {
  const uint8_t* stream; const uint64_t* offsets; uint64_t n; uint64_t chunk; uint64_t num_chunks;
  uint32_t flags; int fixed_type; uint32_t fixed_mask20; F64Const k;
  const double* cmin; const uint32_t* cmask; const int32_t* ctype;
  double* out; int32_t* status;
};

template <int CT>
__global__ void __launch_bounds__(128) dcb_f64_decode_kernel(F64DecArgs a) {
  // F64_DEC_LANES lanes of a warp decode (one chunk each), the others leave at once: there are only num_chunks independent
  // streams, and with all 32 lanes busy a 2^27-element array fills 1.7 warps per scheduler, every latency exposed
  if ((threadIdx.x & 31) >= F64_DEC_LANES) return;
  const uint64_t c = ((uint64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * F64_DEC_LANES + (threadIdx.x & 31);
  if (c >= a.num_chunks) return;
  const uint64_t lo = c * a.chunk, cn = min(a.chunk, a.n - lo);
  const uint32_t* __restrict__ s32 = reinterpret_cast<const uint32_t*>(a.stream);
  const uint64_t end_word = (a.offsets[a.num_chunks] + 3) >> 2;                    // words past the stream read as 0
  uint64_t pos = a.offsets[c] * 8ull;
  const bool addmin = (a.flags & DCB_F_SUBTRACT_MIN) != 0;
  const double mn = addmin ? a.cmin[c] : 0.0;
  int type = a.fixed_type; uint32_t mask20 = a.fixed_mask20;
  if (CT == CT_MASK && !(a.flags & DCB_F_FIXED_MASK)) { type = a.ctype[c]; mask20 = a.cmask[c]; }
  const int m_mask = f64_mant_bits((uint64_t)mask20 << 44, a.k.n_eb);
  double b3 = -1.0, b2 = -1.0, b1 = -1.0;
  // Only 32,768 lanes decode a 2^27-element array (one per chunk), so every latency is exposed (r2 profile: issue active 25 %,
  // long-scoreboard + wait stalls, 79 instructions per element).  The reader therefore keeps a sliding window of three stream
  // words plus the next TWO already requested (a code advances the window by at most two words), the position is a word index
  // and a bit offset (no 64-bit shifts per element), the history is register renaming unless a value is the -1.0 marker
  // (:2797-2814, rare: exact update), and values leave in pairs as 16-byte stores.
  uint64_t wi = pos >> 5;
  uint32_t o = (uint32_t)(pos & 31ull);
  #define F64_LDW(IDX) ((IDX) < end_word ? s32[(IDX)] : 0u)
  uint32_t w0 = __byte_perm(F64_LDW(wi), 0u, 0x0123), w1 = __byte_perm(F64_LDW(wi + 1), 0u, 0x0123),
           w2 = __byte_perm(F64_LDW(wi + 2), 0u, 0x0123);
  uint32_t n3 = F64_LDW(wi + 3), n4 = F64_LDW(wi + 4);
  bool full = false;                                               // all three history slots hold a value
  const bool vec = ((reinterpret_cast<uintptr_t>(a.out + lo) & 15) == 0);
  double pend = 0.0;
  for (uint64_t i = 0; i < cn; i++) {
    // ---- eight 3-bit codes at once (the float32 decoder's fast round): no parse, the code bits only select among the
    //      predictors; taken when the history is full, and abandoned (nothing committed) if a value comes out negative,
    //      which the -1.0 marker is -- the code-by-code path below then handles these codes exactly
    if (CT != CT_NP) {
      const uint32_t t32 = __funnelshift_l(w1, w0, o);
      if (full && !(i & 1ull) && i + 8 <= cn && (t32 & 0x92492400u) == 0x92492400u) {
        double c1 = b1, c2 = b2, c3 = b3;
        double t1 = __dmul_rn(3.0, c1), t2 = __dmul_rn(3.0, c2);
        double x8[8];
        uint32_t neg = 0u;
        #pragma unroll
        for (int e = 0; e < 8; e++) {
          const double p2 = __dsub_rn(__dadd_rn(c1, c1), c2);          // 2.0 * b1 is exact
          const double p3 = __dadd_rn(__dsub_rn(t1, t2), c3);
          const bool bb = (t32 & (0x20000000u >> (3 * e))) != 0u;
          const bool ba = (t32 & (0x40000000u >> (3 * e))) != 0u;
          const double x = ba ? (bb ? p3 : p2) : (bb ? c1 : 0.0);
          x8[e] = x;
          neg |= (uint32_t)((uint64_t)__double_as_longlong(x) >> 32);
          c3 = c2; c2 = c1; c1 = x; t2 = t1; t1 = __dmul_rn(3.0, x);
        }
        bool ok = !(neg >> 31);                                      // a negative value is only a reason to look for the marker
        if (!ok) {
          bool mk = false;
          #pragma unroll
          for (int e = 0; e < 8; e++) mk = mk || (x8[e] == -1.0);
          ok = !mk;
        }
        if (ok) {
          b1 = c1; b2 = c2; b3 = c3;
          o += 24u;
          if (o >> 5) {                                              // at most one word
            o &= 31u; wi += 1;
            w0 = w1; w1 = w2; w2 = __byte_perm(n3, 0u, 0x0123);
            n3 = n4; n4 = F64_LDW(wi + 4);
          }
          if (addmin) {
            #pragma unroll
            for (int e = 0; e < 8; e++) x8[e] = __dadd_rn(x8[e], mn);
          }
          if (vec) {
            #pragma unroll
            for (int e = 0; e < 8; e += 2) *reinterpret_cast<double2*>(a.out + lo + i + e) = make_double2(x8[e], x8[e + 1]);
          } else {
            #pragma unroll
            for (int e = 0; e < 8; e++) a.out[lo + i + e] = x8[e];
          }
          i += 7;                                                    // (+1 by the loop)
          continue;
        }
      }
    }
    const uint64_t w = ((uint64_t)__funnelshift_l(w1, w0, o) << 32) | (uint64_t)__funnelshift_l(w2, w1, o);
    double v; int len;
    if (w >> 63) {
      if (CT == CT_NP) { if (a.status) atomicExch(a.status, DCB_EFORMAT); return; }
      len = 3;
      const uint32_t sel = (uint32_t)(w >> 61) & 3u;
      if (sel == 0) v = 0.0;
      else if (sel == 1) v = b1;
      else if (sel == 2) v = __dsub_rn(__dmul_rn(2.0, b1), b2);
      else v = __dadd_rn(__dsub_rn(__dmul_rn(3.0, b1), __dmul_rn(3.0, b2)), b3);
    } else if (CT == CT_OP) {
      len = 64; v = __longlong_as_double((long long)w);
    } else {
      bool escape = false;
      if (CT == CT_MASK) { const uint64_t ones = (1ull << type) - 1ull; escape = ((w >> (63 - type)) & ones) == ones; }
      uint64_t top; int have;
      if (!escape) {
        have = 12 + f64_mant_bits(w, a.k.n_eb);
        top = have == 64 ? w : (w & ~(~0ull >> have));
        len = have;
      } else {
        const int flag = (int)((w >> (62 - type)) & 1ull);
        const int k = flag ? m_mask : max(m_mask - 8, 0);
        const int hdr = type + 2;
        const uint64_t payload = k ? ((w << hdr) >> (64 - k)) : 0ull;
        len = hdr + k;
        if (flag) { top = ((uint64_t)(mask20 >> 8) << 52) | (k ? payload << (52 - k) : 0ull); have = 12 + k; }
        else { top = ((uint64_t)mask20 << 44) | (k ? payload << (44 - k) : 0ull); have = 20 + k; }
      }
      v = __longlong_as_double((long long)(have >= 64 ? top : (top | (0x8000000000000000ull >> have))));
    }
    o += (uint32_t)len;
    const uint32_t adv = o >> 5;                                   // 0, 1 or 2 words
    o &= 31u;
    if (adv) {
      wi += adv;
#if F64_DEC_PREFETCH
      // the words requested "ahead" are moved at the very next advance (a data-dependent rotation), so their loads are waited
      // for almost at once (r2 profile: 36 % of the samples); the sector F64_DEC_PREFETCH words ahead is asked into L1 instead
      if (((uint32_t)wi & 7u) < adv && wi + F64_DEC_PREFETCH < end_word) asm volatile("prefetch.global.L1 [%0];" :: "l"(s32 + wi + F64_DEC_PREFETCH));
#endif
      if (adv == 1u) {
        w0 = w1; w1 = w2; w2 = __byte_perm(n3, 0u, 0x0123);
        n3 = n4; n4 = F64_LDW(wi + 4);
      } else {
        w0 = w2; w1 = __byte_perm(n3, 0u, 0x0123); w2 = __byte_perm(n4, 0u, 0x0123);
        n3 = F64_LDW(wi + 3); n4 = F64_LDW(wi + 4);
      }
    }
    const double r = addmin ? __dadd_rn(v, mn) : v;
    if (vec) {
      if (i & 1ull) *reinterpret_cast<double2*>(a.out + lo + i - 1) = make_double2(pend, r);
      else if (i + 1 == cn) a.out[lo + i] = r;
      else pend = r;
    } else {
      a.out[lo + i] = r;
    }
    if (CT != CT_NP) {
      if (full && v != -1.0) { b3 = b2; b2 = b1; b1 = v; }
      else {
        if (b3 == -1.0) b3 = v; else if (b2 == -1.0) b2 = v; else if (b1 == -1.0) b1 = v;
        else { b3 = b2; b2 = b1; b1 = v; }
        full = b3 != -1.0 && b2 != -1.0 && b1 != -1.0;
      }
    }
  }
  #undef F64_LDW
}
