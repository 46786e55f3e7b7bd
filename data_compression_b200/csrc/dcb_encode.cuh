// dcb_encode.cuh -- single-pass encoder for CT=5/6/7/11 (replaces myCompress_bitwise{,_np,_mask,_op},
// dataCompression.c:3310, :2645, :2030, :577, and their per-bit add_bit_to_bytes :5456).
//
// PERSISTENT, SOFTWARE-PIPELINED kernel: 3 thread blocks per SM stay resident and pull TILES of 4096
// consecutive elements from a ticket counter (16 elements per thread, blocked, so a thread concatenates its
// own codes in registers and the 3-element look-back needs no shuffles).  v2 (one tile per block) was
// latency bound -- ticket atomic, DRAM load, five barriers and the look-back round trip are a serial chain
// per block and only 4-5 blocks fit an SM -- so every long-latency step is now issued one tile ahead:
//   * the NEXT tile's 16 KB are fetched with cp.async (16 B per thread x 4) into a padded shared-memory
//     slot per thread (80-byte pitch: conflict-free 128-bit reads) while the current tile is classified
//   * the ticket of the tile after next is fetched by thread 0 and parked in shared memory
//   * the decoupled look-back of tile t is resolved during iteration t+1: its status words are loaded at the
//     top of the iteration and consumed after the classification, when every predecessor has long published
// Per tile:
//   1. [fused toSmallDataset_float / med_dataset_float when the chunk is one tile] one block reduction
//   2. classification.  FAST tile (every value finite and >= 0, i.e. no -1.0f sentinel, no NaN/Inf):
//        the 3-bit code of every element is built with static shifts into a 48-bit string; the
//        fallback code (raw / escape / full word) is only computed in warps that actually contain one.
//      SLOW tile (anything else): the reference's exact sentinel state machine, element by element.
//   3. block scan of the per-thread bit counts -> bit offset inside the tile; the tile's aggregate is
//      published for the look-back at once
//   4. the PREVIOUS tile is finished: look-back (chunk ends round up to a byte), then its staged words are
//      funnel-shifted to the output alignment, byte-swapped to the reference's MSB-first byte order and
//      written as coalesced 32-bit words
//   5. packing of the current tile into the staging buffer: a warp whose threads all hold sixteen 3-bit
//      codes places its 48-bit strings with three funnel shifts; other warps run a branch-free shift/flush loop
#pragma once
#include "dcb_common.cuh"

#define DCB_ENC_STATS_READY 0x100u   // EncArgs.flags: cmin / cmed / ctype were filled by the stats pre-pass
#define DCB_ENC_ATOMIC_OUT  0x200u   // tile boundaries are not byte aligned: partial words go out with atomicOr
#define DCB_ENC_IN_ALIGNED  0x400u   // input pointer is 16-byte aligned (128-bit loads)

// ------------------------------------------------------------------------------------------------
// fallback code of one element, TOP-ALIGNED in 32 bits (stream bit 0 of the code = bit 31), plus its length
//   CT=5/6 raw code R(x): compress_bitwise_float :3479-3520 -- top 9+m bits of the word, truncation
//   CT=11: the full word (:689-692);  CT=7: M(x), compress_bitwise_float_mask :2143-2284
template <int CT, bool NONNEG>
__device__ __forceinline__ void fallback_top(uint32_t u, const CodecConst& k, int type, uint32_t mask17,
                                             uint32_t& top, int& len) {
  if (CT == CT_OP) { top = u; len = 32; return; }
  const int E = NONNEG ? (int)(u >> 23) : (int)((u >> 23) & 0xFFu);
  const int sh = min(max(150 - k.n_eb - E, 0), 23);        // 32 - (9 + m),  m = clamp(n_eb + E - 127, 0, 23)
  top = u & (0xFFFFFFFFu << sh);
  len = 32 - sh;
  if (CT == CT_MASK) {
    if ((u >> 23) == (mask17 >> 8)) {                       // sign + exponent equal the mask's (:2172-2180)
      const int m = 23 - sh;
      const bool low8 = ((u >> 15) & 0xFFu) == (mask17 & 0xFFu);
      const int kbits = low8 ? max(m - 8, 0) : m;          // mantissa bits 9..m (:2208-2221) or 1..m (:2222-2266)
      const uint32_t prefix = (((1u << type) - 1u) << 1) | (low8 ? 0u : 1u);   // 0, type x 1, flag
      const int hdr = type + 2;
      // payload = mantissa bits (m-kbits+1 .. m), placed right after the header
      const uint32_t pay = (u << (9 + m - kbits)) & ~__funnelshift_rc(0xFFFFFFFFu, 0u, kbits);   // top-aligned kbits bits
      top = (prefix << (32 - hdr)) | (pay >> hdr);
      len = hdr + kbits;
    }
  }
}

// slow path: exact sentinel state machine of the reference, history carried in (b3,b2,b1)
template <int CT>
__device__ __forceinline__ void classify_sentinel(float x, float& b3, float& b2, float& b1,
                                                  const CodecConst& k, int type, uint32_t mask17,
                                                  uint32_t& top, int& len) {
  const uint32_t u = f2u(x);
  if (CT == CT_NP) { fallback_top<CT_NP, false>(u, k, type, mask17, top, len); return; }
  const bool zero = fabsf(x) <= k.tz;                                    // fabs(x) < absErrorBound in double, :3390
  bool hit = false; uint32_t sel = 5u;
  if (b3 == -1.0f || b2 == -1.0f || b1 == -1.0f) {                       // :3330
    hist_push(b3, b2, b1, x);
  } else {
    const float p2 = pred2(b1, b2);
    const float p3 = pred3(b1, b2, b3);
    const float d1 = fabsf(__fsub_rn(b1, x));
    const float d2 = fabsf(__fsub_rn(p2, x));
    const float d3 = fabsf(__fsub_rn(p3, x));
    float dmin = d1;
    if (d2 < dmin) { dmin = d2; sel = 6u; }                              // 'b' 110 (strict <, :3373)
    if (d3 < dmin) { dmin = d3; sel = 7u; }                              // 'c' 111
    hit = dmin <= k.tp;                                                  // diff_min <= absErrorBound, :3399
    b3 = b2; b2 = b1; b1 = x;                                            // :3386-3388
  }
  uint32_t ft; int fl;
  fallback_top<CT, false>(u, k, type, mask17, ft, fl);
  top = zero ? (4u << 29) : (hit ? (sel << 29) : ft);
  len = (zero || hit) ? 3 : fl;
}

// ------------------------------------------------------------------------------------------------
// decoupled look-back over tiles (Merrill & Garland style), in BITS, with byte round-up at chunk ends.
// Called by warp 0.  Returns the absolute bit offset of tile `tile`'s first bit.
// Walking backwards from tile-1:  open = bits of the visited tiles up to and including the first
// (stream-order) chunk end among them, tail = everything after that chunk end (already well defined,
// it starts on a byte boundary).  See DESIGN.md "look-back with chunk ends".
// `w0` = status[tile-1-lane] as loaded earlier by the caller (stale INVALID entries are simply re-read).
__device__ __forceinline__ unsigned long long lookback_bits(const volatile unsigned long long* status,
                                                            uint32_t tile, int lane, unsigned long long w0) {
  unsigned long long open = 0, tail = 0;
  bool have_end = false, first = true;
  int64_t j0 = (int64_t)tile - 1;                 // nearest predecessor handled by lane 0
  while (true) {
    const int64_t j = j0 - lane;
    // tiles before 0 read as "inclusive prefix 0"; only real tiles can still be INVALID
    unsigned long long w = first ? w0 : ((j >= 0) ? status[j] : (ST_INCL << ST_SHIFT));
    first = false;
    while (__any_sync(0xffffffffu, (w >> ST_SHIFT) == ST_INVALID)) {
      __nanosleep(40);
      if ((w >> ST_SHIFT) == ST_INVALID) w = status[j];
    }
    const unsigned incl_mask = __ballot_sync(0xffffffffu, (w >> ST_SHIFT) == ST_INCL);
    const int first_incl = incl_mask ? (__ffs(incl_mask) - 1) : 32;
    // lanes 0 .. first_incl-1 hold aggregates, nearest first; fold them serially (uniform across the warp)
    for (int l = 0; l < first_incl; l++) {
      const unsigned long long wl = __shfl_sync(0xffffffffu, w, l);
      const unsigned long long bits = wl & ST_VAL_MASK;
      if (wl & ST_END_BIT) {
        tail += have_end ? ((open + 7ull) & ~7ull) : open;
        open = bits;
        have_end = true;
      } else {
        open += bits;
      }
    }
    if (first_incl < 32) {
      const unsigned long long X = __shfl_sync(0xffffffffu, w, first_incl) & ST_VAL_MASK;
      unsigned long long pos = X + open;
      if (have_end) pos = (pos + 7ull) & ~7ull;
      return pos + tail;
    }
    j0 -= 32;
  }
}


// ------------------------------------------------------------------------------------------------
#define DCB_XPITCH 20            // floats per thread slot of the input staging buffer (16 + 4 pad)
#define DCB_ENC_BLOCKS_PER_SM 3

struct EncSmem {
  float X[2][DCB_NT * DCB_XPITCH];   // double-buffered input tiles, one 80-byte slot per thread
  uint32_t W[2][DCB_TILE + 8];       // packed tiles (double buffered): W[b][0] = 0; stream word k of the tile lives at W[b][1 + k]
  float red_mn[2][DCB_NT / 32];      // partials of the block reductions, double buffered (see block_stats)
  float red_mx[2][DCB_NT / 32];
  float red_am[2][DCB_NT / 32];
  float red_s[2][DCB_NT / 32];
  uint32_t warp_tot[2][DCB_NT / 32];
  uint32_t xchg[DCB_NT / 32];        // tail word of every warp's last lane (completes the next warp's first word)
  float hx[2][DCB_NT / 32][4];       // last three RAW values of every warp (look-back of the next warp's lane 0)
  uint32_t tkt[2];                   // prefetched tickets
  float red_w[2][DCB_NT / 32];       // CT=7: centre and half-width of the bracket of the sequential sum
  float red_e[2][DCB_NT / 32];
  uint32_t walkmask[DCB_NT / 32];    // exact-sum fallback: threads where the serial walk has to stop
  unsigned long long P;              // absolute bit offset of the tile being copied out
  uint32_t pend_tile[2], pend_bits[2], pend_c[2], pend_last[2];   // tiles packed in the previous two iterations (each is finished two iterations later)
  float mn, med;
  int type;
  uint32_t fi;
  uint32_t role;
};

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
  const uint32_t d = (uint32_t)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(d), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory"); }

// Fetch tile `tile` into the staging buffer: each thread brings its own 16 consecutive floats.
// Tile prefetch.  A warp owns 512 consecutive floats (2 KB); it fetches them with four fully coalesced copies (lane L of copy q
// moves the 16 bytes at offset (32 q + L) * 16), each landing in the slot of the thread that will consume those four values:
// thread t's sixteen floats sit at Xbuf[t * DCB_XPITCH .. + 15].  (Copies in the per-thread order -- lane stride 64 B -- touch
// 32 sectors per instruction and kept the load/store unit busy for ~0.5 us per tile.)  A thread's slot is written by four
// different lanes of its own warp: readers need cp.async.wait_group AND __syncwarp.
__device__ __forceinline__ void issue_tile_load(const EncArgs& a, uint32_t tile, float* Xbuf, int tid) {
  const int lane = tid & 31;
  const uint64_t w0 = (uint64_t)tile * DCB_TILE + (uint64_t)(tid & ~31) * DCB_E;      // first element of this warp
  if (w0 + 32 * DCB_E <= a.n && (a.flags & DCB_ENC_IN_ALIGNED)) {                      // warp uniform
    const float* g = a.in + w0;
    float* sb = Xbuf + (tid & ~31) * DCB_XPITCH;
    #pragma unroll
    for (int q = 0; q < 4; q++) {
      const int u = q * 32 + lane;                      // 16-byte unit inside the warp's 2 KB
      cp_async16(sb + (u >> 2) * DCB_XPITCH + (u & 3) * 4, g + u * 4);
    }
  } else {                                              // ragged last tile / unaligned input: plain loads, zero padding
    const uint64_t my0 = (uint64_t)tile * DCB_TILE + (uint64_t)tid * DCB_E;
    float* slot = Xbuf + tid * DCB_XPITCH;
    #pragma unroll
    for (int e = 0; e < DCB_E; e++) slot[e] = (my0 + e < a.n) ? __ldg(a.in + my0 + e) : 0.0f;
  }
}

// max that PROPAGATES NaN (fmaxf drops it): one instruction (FMNMX.NAN), used to spot NaN/Inf/huge values
__device__ __forceinline__ float fmax_nan(float a, float b) {
  float r;
  asm("max.NaN.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b));
  return r;
}

// One barrier: block-wide min, [max, CT=7 only] and NaN-propagating max of |x| of the raw tile.  `par` alternates
// between consecutive calls so that the partial arrays need no trailing barrier.
template <bool WANT_MAX>
__device__ __forceinline__ void block_stats(float& mn, float& mx, float& am, EncSmem& sm, int par, int lane, int warp) {
  #pragma unroll
  for (int o = 16; o; o >>= 1) {
    mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o));
    if (WANT_MAX) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    am = fmax_nan(am, __shfl_xor_sync(0xffffffffu, am, o));
  }
  if (lane == 0) { sm.red_mn[par][warp] = mn; if (WANT_MAX) sm.red_mx[par][warp] = mx; sm.red_am[par][warp] = am; }
  __syncthreads();
  mn = sm.red_mn[par][0]; if (WANT_MAX) mx = sm.red_mx[par][0]; am = sm.red_am[par][0];
  #pragma unroll
  for (int w = 1; w < DCB_NT / 32; w++) {
    mn = fminf(mn, sm.red_mn[par][w]); if (WANT_MAX) mx = fmaxf(mx, sm.red_mx[par][w]); am = fmax_nan(am, sm.red_am[par][w]);
  }
}
__device__ __forceinline__ float block_sum_any_order(float v, EncSmem& sm, int par, int lane, int warp) {
  #pragma unroll
  for (int o = 16; o; o >>= 1) v = __fadd_rn(v, __shfl_xor_sync(0xffffffffu, v, o));
  if (lane == 0) sm.red_s[par][warp] = v;
  __syncthreads();
  float r = sm.red_s[par][0];
  #pragma unroll
  for (int w = 1; w < DCB_NT / 32; w++) r = __fadd_rn(r, sm.red_s[par][w]);
  return r;
}

// 3-bit field of element e inside the 48-bit string (chi:clo, element 0 at the top)
#define C48_FIELD(chi, clo, e) ((e) <= 9 ? (((chi) >> (29 - 3 * (e))) & 7u) : \
                                ((e) == 10 ? ((((chi) << 1) & 6u) | ((clo) >> 31)) : (((clo) >> (61 - 3 * (e))) & 7u)))

// copy-out: shift the packed tile to the output bit alignment, MSB-first byte order, coalesced 32-bit words
__device__ __forceinline__ void copy_out_tile(const EncArgs& a, const uint32_t* W, unsigned long long P,
                                              uint32_t tile_bits, bool last_in_chunk, int tid) {
  unsigned long long Eabs = P + tile_bits;
  if (last_in_chunk) Eabs = (Eabs + 7ull) & ~7ull;          // zero padding of the chunk's last byte (:5467)
  if (Eabs == P) return;
  const unsigned long long g0 = P >> 5, g1 = (Eabs - 1ull) >> 5;
  const uint32_t sh = (uint32_t)(P & 31ull), tb = (uint32_t)(Eabs & 31ull);
  uint32_t* out32 = reinterpret_cast<uint32_t*>(a.out);
  const bool atomic_out = (a.flags & DCB_ENC_ATOMIC_OUT) != 0;
  const uint32_t nw = (uint32_t)(g1 - g0) + 1u;            // <= DCB_TILE + 1 words
  uint32_t* const o32 = out32 + g0;
  #pragma unroll 4
  for (uint32_t k = (uint32_t)tid; k < nw; k += DCB_NT) {    // independent iterations: the loads of four go out together
    uint32_t v = __funnelshift_r(W[1 + k], W[k], sh);        // (W[k] : W[k+1]) >> sh; sh == 0 gives W[k+1]
    const bool head = (k == 0u) && sh, tail = (k == nw - 1u) && tb;
    if (!head && !tail) {
      o32[k] = __byte_perm(v, 0u, 0x0123);
    } else {
      const uint32_t lo_bit = head ? sh : 0u, hi_bit = tail ? tb : 32u;   // valid bits [lo_bit, hi_bit), MSB first
      uint32_t mask = 0xFFFFFFFFu >> lo_bit;
      if (hi_bit < 32u) mask &= ~(0xFFFFFFFFu >> hi_bit);
      v &= mask;
      if (atomic_out) {
        atomicOr(&o32[k], __byte_perm(v, 0u, 0x0123));
      } else {
        uint8_t* ob = reinterpret_cast<uint8_t*>(o32 + k);
        #pragma unroll
        for (uint32_t b = 0; b < 4; b++)
          if (8u * b >= lo_bit && 8u * b + 8u <= hi_bit) ob[b] = (uint8_t)(v >> (24u - 8u * b));
      }
    }
  }
}


// ------------------------------------------------------------------------------------------------
// CENTRAL SCANNER.  A look-back window of 32 tiles advances the inclusive-prefix frontier by at most 32 tiles
// per L2 round trip, which capped every earlier version at ~45 tiles/us (1.4 ms for 2^28 elements) no matter
// how cheap the rest was (profiles/r1_summary.md).  Instead the first block that starts running becomes
// the scanner: it streams over the status array 256 tiles per hop, turns aggregates into inclusive prefixes
// with a block-wide scan, and writes them back.  A worker needs exactly one word, the inclusive prefix of
// the tile before its own, one iteration after it published its aggregate.
//
// Status word of tile j:  [63:62] state (0 invalid, 1 aggregate, 2 inclusive)  [61] the tile ends a chunk
//                         [55:0] aggregate: bits of the tile / inclusive: E(j), the absolute bit offset after it
// E(j) = ends_chunk(j) ? round_up_8(E(j-1) + bits(j)) : E(j-1) + bits(j)        (chunks are byte aligned)
// The scan operator on (flag, B1, B2a, B2o) -- "no chunk end inside: +B1" or "round_up_8(p + B1) + B2a + B2o",
// B2a the byte-aligned part after the first chunk end, B2o the open bits after the last one -- is associative.
struct ScanEl { uint32_t f1; uint32_t a; uint32_t o; };   // f1 = flag << 31 | B1
__device__ __forceinline__ ScanEl scan_combine(const ScanEl L, const ScanEl R) {   // L earlier in the stream
  const uint32_t lf = L.f1 >> 31, rf = R.f1 >> 31, lb = L.f1 & 0x7FFFFFFFu, rb = R.f1 & 0x7FFFFFFFu;
  ScanEl x;
  if (!rf) {
    if (!lf) { x.f1 = lb + rb; x.a = 0u; x.o = 0u; }
    else { x.f1 = L.f1; x.a = L.a; x.o = L.o + rb; }
  } else {
    if (!lf) { x.f1 = 0x80000000u | (lb + rb); x.a = R.a; x.o = R.o; }
    else { x.f1 = L.f1; x.a = L.a + ((L.o + rb + 7u) & ~7u) + R.a; x.o = R.o; }
  }
  return x;
}
__device__ __forceinline__ unsigned long long scan_apply(unsigned long long p, const ScanEl x) {
  const unsigned long long b1 = x.f1 & 0x7FFFFFFFu;
  return (x.f1 >> 31) ? (((p + b1 + 7ull) & ~7ull) + x.a + x.o) : (p + b1);
}

#ifdef DCB_DEBUG_COUNTERS
// profiling aid (never in the shipped build): clock64 timelines of lane 0 of warps 0 and 7 of four worker blocks
#define DBG_BLOCKS 4
#define DBG_ITERS 40
#define DBG_MARKS 16
__device__ long long g_dbg_ts[DBG_BLOCKS][2][DBG_ITERS][DBG_MARKS];
#define DBG_ADD(i, v) do {} while (0)
#define DBG_CLK() clock64()
#else
#define DBG_ADD(i, v) do {} while (0)
#define DBG_CLK() 0ll
#endif
// Per-tile duties with a global-memory round trip are spread over the first lanes of different warps: the warp that
// carries one is late at the next barrier, and everybody waits for the latest warp.
#define DCB_TKT_TID 32                                  // claims tile tickets
#define DCB_RES_TID 64                                  // fetches the pending tile's prefix, publishes chunk offsets
#define DCB_SCAN_PER 4                                   // status words per scanner thread and hop (1024 tiles per hop)
struct ScanSmem {
  ScanEl wtot[DCB_NT / 32];
  uint32_t first_invalid[DCB_NT / 32];
  unsigned long long P;
};

// One hop costs an L2 round trip for the loads, a block scan and two barriers whatever its width; with 256 tiles per hop
// the scanner topped out near 70 tiles/us and every worker ended up polling for its prefix (12-25 polls per tile in the
// r1 profiles) -- the encoder needs 90+.  Four words per thread quadruple the width at the same hop latency.
__device__ __noinline__ void scanner_role(const EncArgs& a, ScanSmem& ss, int tid) {
  volatile unsigned long long* status = a.status;
  const int lane = tid & 31, warp = tid >> 5;
  const uint32_t limit = a.num_tiles - 1u;             // nobody needs the inclusive prefix of the last tile
  uint32_t q = 0u;
  unsigned long long P = 0ull;
  while (q < limit) {
    const uint32_t j0 = q + (uint32_t)tid * DCB_SCAN_PER;
    unsigned long long w[DCB_SCAN_PER];
    #pragma unroll
    for (int k = 0; k < DCB_SCAN_PER; k++) w[k] = (j0 + k < limit) ? status[j0 + k] : 0ull;
    uint32_t nv = 0u;                                   // leading valid words of this thread
    bool open = true;
    #pragma unroll
    for (int k = 0; k < DCB_SCAN_PER; k++) { open = open && ((w[k] >> ST_SHIFT) != ST_INVALID); nv += open ? 1u : 0u; }
    const unsigned inval = __ballot_sync(0xffffffffu, nv < DCB_SCAN_PER);
    const int fl = inval ? (__ffs(inval) - 1) : 0;
    const uint32_t nv_first = __shfl_sync(0xffffffffu, nv, fl);
    if (lane == 0) ss.first_invalid[warp] = inval ? ((uint32_t)fl * DCB_SCAN_PER + nv_first) : 32u * DCB_SCAN_PER;
    ScanEl xi[DCB_SCAN_PER];                            // inclusive over this thread's words
    #pragma unroll
    for (int k = 0; k < DCB_SCAN_PER; k++) {
      ScanEl e;
      e.f1 = ((w[k] & ST_END_BIT) ? 0x80000000u : 0u) | (uint32_t)(w[k] & 0x7FFFFFFFull);
      e.a = 0u; e.o = 0u;
      xi[k] = (k == 0) ? e : scan_combine(xi[k - 1], e);
    }
    ScanEl x = xi[DCB_SCAN_PER - 1];
    #pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      ScanEl y;
      y.f1 = __shfl_up_sync(0xffffffffu, x.f1, o);
      y.a = __shfl_up_sync(0xffffffffu, x.a, o);
      y.o = __shfl_up_sync(0xffffffffu, x.o, o);
      if (lane >= o) x = scan_combine(y, x);
    }
    if (lane == 31) ss.wtot[warp] = x;
    ScanEl pre;                                         // exclusive prefix of this thread inside its warp
    pre.f1 = __shfl_up_sync(0xffffffffu, x.f1, 1);
    pre.a = __shfl_up_sync(0xffffffffu, x.a, 1);
    pre.o = __shfl_up_sync(0xffffffffu, x.o, 1);
    if (lane == 0) { pre.f1 = 0u; pre.a = 0u; pre.o = 0u; }
    __syncthreads();
    uint32_t m = 0u;                                    // leading valid entries of this hop
    #pragma unroll
    for (int wq = 0; wq < DCB_NT / 32; wq++) {
      const uint32_t fi = ss.first_invalid[wq];
      if (m == 32u * DCB_SCAN_PER * wq) m += fi;
    }
    if (m == 0u) { __syncthreads(); __nanosleep(100); continue; }
    if ((uint32_t)tid * DCB_SCAN_PER < m) {
      ScanEl wpre; wpre.f1 = 0u; wpre.a = 0u; wpre.o = 0u;
      for (int wq = 0; wq < warp; wq++) wpre = scan_combine(wpre, ss.wtot[wq]);
      const ScanEl tpre = scan_combine(wpre, pre);
      #pragma unroll
      for (int k = 0; k < DCB_SCAN_PER; k++) {
        const uint32_t idx = (uint32_t)tid * DCB_SCAN_PER + k;
        if (idx < m) {
          const unsigned long long E = scan_apply(P, scan_combine(tpre, xi[k]));
          status[j0 + k] = (ST_INCL << ST_SHIFT) | E;
          if (idx == m - 1u) ss.P = E;
        }
      }
    }
    __syncthreads();
    P = ss.P;
    q += m;
  }
}

// one thread of a worker (DCB_RES_TID): wait for the inclusive prefix of tile-1, publish the chunk offset of the pending tile
__device__ __forceinline__ void resolve_pending(const EncArgs& a, EncSmem& sm, int b, unsigned long long w0) {
  volatile unsigned long long* status = a.status;
  const uint32_t tile = sm.pend_tile[b];
  unsigned long long P = 0ull;
  if (tile > 0u) {
    const long long dbg_t0 = DBG_CLK();
    (void)dbg_t0;                                       // only read in -DDCB_DEBUG_COUNTERS builds
    while ((w0 >> ST_SHIFT) != ST_INCL) { __nanosleep(20); w0 = status[tile - 1u]; DBG_ADD(2, 1); }
    DBG_ADD(3, DBG_CLK() - dbg_t0);
    P = w0 & ST_VAL_MASK;
  }
  unsigned long long E = P + sm.pend_bits[b];
  if (tile == a.num_tiles - 1u && a.last_bits) *a.last_bits = E;
  sm.P = P;
  if (tile == 0u) a.offsets[0] = 0ull;
  if (sm.pend_last[b]) a.offsets[(uint64_t)sm.pend_c[b] + 1ull] = (E + 7ull) >> 3;
}

template <int CT>
__global__ void __launch_bounds__(DCB_NT, DCB_ENC_BLOCKS_PER_SM) dcb_encode_kernel(const EncArgs a) {
  extern __shared__ __align__(16) unsigned char dcb_enc_smem[];
  EncSmem& sm = *reinterpret_cast<EncSmem*>(dcb_enc_smem);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  volatile unsigned long long* status = a.status;
  // The first block that starts running becomes the scanner (claimed at run time, so it is resident by
  // construction); every other block is a worker.  Tile ids are tickets for the same reason: a tile is only
  // ever claimed by a running block, so nobody waits on an unscheduled one.
  if (tid == 0) {
    const uint32_t role = (a.num_tiles > 1u && atomicCAS(a.scanner_claim, 0u, 1u) == 0u) ? 1u : 0u;
    sm.role = role;
    if (!role) { sm.tkt[0] = atomicAdd(a.ticket, 1u); sm.tkt[1] = atomicAdd(a.ticket, 1u); }
    sm.W[0][0] = 0u; sm.W[1][0] = 0u;
  }
  __syncthreads();
  if (sm.role) { scanner_role(a, *reinterpret_cast<ScanSmem*>(sm.X[0]), tid); return; }
  uint32_t cur = sm.tkt[0], nxt = sm.tkt[1];
  if (cur >= a.num_tiles) return;
  issue_tile_load(a, cur, sm.X[0], tid);
  cp_async_commit();

  const bool submin = (a.flags & 1u) != 0;            // DCB_F_SUBTRACT_MIN
  const bool fixed_mask = (a.flags & 2u) != 0;        // DCB_F_FIXED_MASK
  const bool stats_ready = (a.flags & DCB_ENC_STATS_READY) != 0;

  uint32_t fix_hw = 0; bool fix_need = false;         // lane 0 of warps > 0: first word of the pending tile still lacks the previous warp's tail
  int par = 0;

  uint32_t it = 0;
  for (;; it++) {
    const int buf = (int)(it & 1u);
    float* slot = &sm.X[buf][tid * DCB_XPITCH];
#ifdef DCB_DEBUG_COUNTERS
    const int dbg_blk = (int)blockIdx.x - 200;           // four blocks from the middle of the grid
    const bool dbg_on = dbg_blk >= 0 && dbg_blk < DBG_BLOCKS && it >= 20u && it < 20u + DBG_ITERS && lane == 0 && (warp == 0 || warp == 7);
    #define DBG_MARK(i) do { if (dbg_on) g_dbg_ts[dbg_blk][warp ? 1 : 0][it - 20u][i] = clock64(); } while (0)
    DBG_MARK(0);
#else
    #define DBG_MARK(i) do {} while (0)
#endif
    // ---- issue everything with a long latency for LATER use -------------------------------------------
    if (nxt < a.num_tiles) issue_tile_load(a, nxt, sm.X[buf ^ 1], tid);
    cp_async_commit();
    DBG_MARK(1);
    uint32_t tk = 0u;
    if (tid == DCB_TKT_TID) tk = atomicAdd(a.ticket, 1u);               // ticket of the tile after next
    unsigned long long w0 = 0ull;                                       // inclusive prefix of the tile before the pending one
    if (tid == DCB_RES_TID && it > 1u && sm.pend_tile[buf] > 0u) w0 = status[sm.pend_tile[buf] - 1u];   // tile packed two iterations ago
    DBG_MARK(2);
    uint32_t* const W = sm.W[buf];                                      // staging buffer of the current tile (and of the one two iterations ago)
    if (fix_need) sm.W[buf ^ 1][fix_hw] |= sm.xchg[warp - 1];           // previous tile (ordered before its copy-out by the barriers in between)
    fix_need = false;

    DBG_MARK(3);
    // ---- chunk geometry of the current tile ---------------------------------------------------------------
    const uint32_t tile = cur;
    const uint64_t base = (uint64_t)tile * DCB_TILE;
    const bool full = base + DCB_TILE <= a.n;          // block-uniform: no ragged tail in this tile
    uint32_t c = 0u, tic = 0u;                         // chunk index, tile index inside the chunk
    bool last_in_chunk = (tile == a.num_tiles - 1);
    if (a.tiles_per_chunk == 1u) {
      c = tile; last_in_chunk = true;
    } else if (a.tiles_per_chunk) {
      c = tile / a.tiles_per_chunk;
      tic = tile - c * a.tiles_per_chunk;
      last_in_chunk = last_in_chunk || (tic == a.tiles_per_chunk - 1);
    } else {
      tic = tile;
    }
    const uint64_t my0 = base + (uint64_t)tid * DCB_E;
    const int vcnt = full ? DCB_E : ((my0 >= a.n) ? 0 : (int)min((uint64_t)DCB_E, a.n - my0));
    const bool first = (tic == 0u && tid == 0);        // chunk start: elements 0..2 have no full history (:3330)
    float hr1 = 0.0f, hr2 = 0.0f, hr3 = 0.0f;          // RAW look-back values of thread 0 (previous tile of the same chunk)
    if (tid == 0 && tic > 0u && CT != CT_NP) { hr1 = __ldg(a.in + my0 - 1); hr2 = __ldg(a.in + my0 - 2); hr3 = __ldg(a.in + my0 - 3); }

    // ---- 1. the tile's 16 floats of this thread (landed while the previous tile was processed) ---------------
    DBG_MARK(4);
    cp_async_wait<1>();
    __syncwarp();                                       // the slot was filled by four lanes of this warp
    DBG_MARK(5);
    float x[DCB_E];
    #pragma unroll
    for (int q = 0; q < 4; q++) {
      const float4 v = *reinterpret_cast<const float4*>(slot + 4 * q);
      x[4 * q] = v.x; x[4 * q + 1] = v.y; x[4 * q + 2] = v.z; x[4 * q + 3] = v.w;
    }

    // ---- 2. one block reduction: min (toSmallDataset_float :3543-3562), max (type), largest |x| word (FAST test) --
    float lmn = __uint_as_float(0x7f800000u), lmx = __uint_as_float(0xff800000u);
    float am = 0.0f;                                   // NaN-propagating max |x|: NaN, Inf and |x| >= 2^127 all end up >= 2^127 or NaN
    bool negative = false;                             // without min-subtract a negative value (sign bit set, -0.0f included) makes the tile SLOW
    if (full) {
      #pragma unroll
      for (int e = 0; e < DCB_E; e++) {
        lmn = fminf(lmn, x[e]);
        if (CT == CT_MASK) lmx = fmaxf(lmx, x[e]);
        am = fmax_nan(am, fabsf(x[e]));
      }
    } else {
      #pragma unroll
      for (int e = 0; e < DCB_E; e++) {
        if (e < vcnt) { lmn = fminf(lmn, x[e]); if (CT == CT_MASK) lmx = fmaxf(lmx, x[e]); }   // NaN never wins a '<' (:3550)
        am = fmax_nan(am, fabsf(x[e]));                                                          // padding is 0.0f
      }
    }
    if (!submin) {
      uint32_t sg = 0u;
      #pragma unroll
      for (int e = 0; e < DCB_E; e++) sg |= f2u(x[e]);
      if (tid == 0) sg |= f2u(hr1) | f2u(hr2) | f2u(hr3);
      negative = (sg >> 31) != 0u;
      if (negative) am = __uint_as_float(0x7fc00000u);                              // fold into the same reduction
    }
    if (tid == 0) am = fmax_nan(am, fmax_nan(fabsf(hr1), fmax_nan(fabsf(hr2), fabsf(hr3))));
    if (lane == 31) { sm.hx[par][warp][0] = x[DCB_E - 1]; sm.hx[par][warp][1] = x[DCB_E - 2]; sm.hx[par][warp][2] = x[DCB_E - 3]; }
    block_stats<CT == CT_MASK>(lmn, lmx, am, sm, par, lane, warp);
    DBG_MARK(6);
    float mn = 0.0f;
    if (submin) mn = stats_ready ? a.cmin[c] : lmn;
    // FAST tile: every |x| (and |min|) below 2^127, so x - min is finite and >= 0; no NaN/Inf; no negative value
    // without min-subtract.  Negative values include the -1.0f "empty history slot" sentinel.
    bool slow = !(am < 1.7014118e38f) || !(fabsf(mn) < 1.7014118e38f);     // 2^127; NaN fails both
    if (submin && !stats_ready) {
      if (mn == 0.0f) {
        // +0 / -0 compare equal: the reference keeps the FIRST minimal element; recover its sign
        uint32_t fi = 0xFFFFFFFFu;
        float fz = 0.0f;
        #pragma unroll
        for (int e = DCB_E - 1; e >= 0; e--)
          if (e < vcnt && x[e] == 0.0f) { fi = (uint32_t)tid * DCB_E + e; fz = x[e]; }
        if (tid == 0) sm.fi = 0xFFFFFFFFu;
        __syncthreads();
        atomicMin(&sm.fi, fi);
        __syncthreads();
        if (fi == sm.fi && fi != 0xFFFFFFFFu) sm.mn = fz;
        // x - (+0) keeps a -0.0f input negative: not a FAST tile (the raw code would carry its sign bit)
        bool negz = false;
        #pragma unroll
        for (int e = 0; e < DCB_E; e++) negz |= (f2u(x[e]) == 0x80000000u);
        slow = __syncthreads_or(negz) || slow;
        mn = sm.mn;
      }
      if (slow) {                                       // data[0] NaN: every '<' fails and min stays NaN (:3546-3553)
        __syncthreads();
        if (tid == 0) sm.mn = (x[0] != x[0]) ? x[0] : mn;
        __syncthreads();
        mn = sm.mn;
      }
      if (tid == 0) a.cmin[c] = mn;
    } else if (!submin && a.cmin && tid == 0 && last_in_chunk) {
      a.cmin[c] = 0.0f;
    }
    if (submin) {
      if (full) {
        #pragma unroll
        for (int e = 0; e < DCB_E; e++) x[e] = __fsub_rn(x[e], mn);
      } else {
        #pragma unroll
        for (int e = 0; e < DCB_E; e++) x[e] = (e < vcnt) ? __fsub_rn(x[e], mn) : 0.0f;
      }
    }

    // ---- look-back values x[-1], x[-2], x[-3] of the thread's first element ------------------------------
    float h1 = __shfl_up_sync(0xffffffffu, x[DCB_E - 1], 1);
    float h2 = __shfl_up_sync(0xffffffffu, x[DCB_E - 2], 1);
    float h3 = __shfl_up_sync(0xffffffffu, x[DCB_E - 3], 1);
    if (lane == 0) {
      if (warp > 0) { h1 = sm.hx[par][warp - 1][0]; h2 = sm.hx[par][warp - 1][1]; h3 = sm.hx[par][warp - 1][2]; }
      else { h1 = hr1; h2 = hr2; h3 = hr3; }
      if (submin) { h1 = __fsub_rn(h1, mn); h2 = __fsub_rn(h2, mn); h3 = __fsub_rn(h3, mn); }
    }

    // ---- 2b. med_dataset_float (:3593-3620): SEQUENTIAL float32 sum (order matters), max -> type -----------
    int type = a.fixed_type;
    uint32_t mask17 = a.fixed_mask17;
    if (CT == CT_MASK && !fixed_mask) {
      if (!stats_ready) {
        const int cnt = (int)min((uint64_t)DCB_TILE, a.n - base);
        // Only the top 17 bits of medium reach the stream (mask = first 17 chars of floattostr(medium),
        // impl/pingpong.c:202-206), so the reference's SEQUENTIAL float32 sum S_seq only has to be bracketed tightly
        // enough to pin those bits.  While the running sum stays inside one binade [2^e, 2^(e+1)) it lives on the grid
        // q = 2^(e-23), and S -> fl(S + y) commutes with shifts of S by 2q (round-to-nearest-even only looks at the position
        // relative to the grid and at the parity of S/q).  So a thread whose 16 values provably neither start nor end outside
        // one binade is a 2-state transducer: S_out = S_in + d[parity(S_in)], with d[0], d[1] from two 16-step chains started
        // at 2^e and 2^e + q -- all threads in parallel, and every d is EXACT.  Since S_seq is the sum of the 256 per-thread
        // increments,
        //     S_seq = sum_t d_t[0]  +-  sum_t |d_t[1] - d_t[0]|          (certain threads; the two usually coincide)
        //                           +-  sum_t 24 ulp(hi_t)               (threads that may straddle a binade: plain sum, 16 roundings
        //                                                                 of <= ulp(hi_t)/2 each, plus the plain sum's own)
        // which is ~5e-6 * S for ordinary data -- 25x tighter than any a-priori bound (cnt * u * S), so that only ~0.5 % of
        // the tiles (instead of 25 %) fall through to the exact walk below.
        bool decided = false;
        float ps = 0.0f;
        float d0 = -1.0f, d1 = -1.0f;                      // < 0: not a transducer (the walk chains this thread's values)
        float lo_keep = 0.0f;
        float inc = 0.0f;
        if (!slow) {
          #pragma unroll
          for (int e = 0; e < DCB_E; e++) ps = __fadd_rn(ps, x[e]);
          inc = ps;
          #pragma unroll
          for (int o = 1; o < 32; o <<= 1) {
            const float t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc = __fadd_rn(inc, t);
          }
          if (lane == 31) sm.red_s[par][warp] = inc;
          __syncthreads();
          float wb = 0.0f;
          for (int w = 0; w < warp; w++) wb = __fadd_rn(wb, sm.red_s[par][w]);
          const float pex = __fadd_rn(wb, __fsub_rn(inc, ps));             // estimate of the running sum before this thread
          // |reference running sum - pex| <= (16 tid [its roundings] + 30 [ours]) u * sum; 96 and 1.02 leave slack
          const float relt = (float)(16 * tid + 96) * (1.02f / 16777216.0f);
          const float lo = __fmul_rd(pex, __fsub_rd(1.0f, relt));
          float hi = __fmul_ru(__fadd_ru(pex, ps), __fadd_ru(1.0f, relt));
          hi = __fmul_ru(hi, 1.00001f);                                     // the 16 round-ups inside the thread and the +q start
          const uint32_t elo = f2u(lo) >> 23, ehi = f2u(hi) >> 23;
          lo_keep = lo;
          float ctr = ps, err = __fmul_ru(hi, 24.0f / 8388608.0f);          // 24 ulp(hi) <= 24 * 2^-23 * hi
          if (elo == ehi && elo > 0u && elo < 255u) {
            const float r0 = u2f(elo << 23), r1 = u2f((elo << 23) | 1u);
            float c0 = r0, c1 = r1;
            #pragma unroll
            for (int e = 0; e < DCB_E; e++) { c0 = __fadd_rn(c0, x[e]); c1 = __fadd_rn(c1, x[e]); }
            d0 = __fsub_rn(c0, r0); d1 = __fsub_rn(c1, r1);                 // exact: both ends lie on the grid of one binade
            ctr = d0; err = fabsf(__fsub_rn(d1, d0));                       // exact too (a few grid steps)
          }
          if (!(a.flags & 8u)) {                                            // DCB_F_EXACT_MEDIUM skips the shortcut
            #pragma unroll
            for (int o = 16; o; o >>= 1) {
              ctr = __fadd_rn(ctr, __shfl_xor_sync(0xffffffffu, ctr, o));
              err = __fadd_ru(err, __shfl_xor_sync(0xffffffffu, err, o));
            }
            if (lane == 0) { sm.red_w[par][warp] = ctr; sm.red_e[par][warp] = err; }
            __syncthreads();
            float Cs = sm.red_w[par][0], Es = sm.red_e[par][0];
            #pragma unroll
            for (int w = 1; w < DCB_NT / 32; w++) { Cs = __fadd_rn(Cs, sm.red_w[par][w]); Es = __fadd_ru(Es, sm.red_e[par][w]); }
            // the reduction tree of the centres (13 roundings of <= u * Cs each): 32 u * Cs with slack
            const float dlt = __fadd_ru(__fmul_ru(Es, 1.001f), __fmul_ru(Cs, 32.0f * 1.02f / 16777216.0f));
            const float blo = fmaxf(__fsub_rd(Cs, dlt), 0.0f), bhi = __fadd_ru(Cs, dlt);
            const float fc = (float)cnt;
            uint32_t mlo, mhi;
            if ((cnt & (cnt - 1)) == 0) {                  // power of two (the usual 4096): an exact reciprocal, same rounding
              const float rc = __frcp_rn(fc);
              mlo = f2u(__fmul_rn(blo, rc)) >> 15; mhi = f2u(__fmul_rn(bhi, rc)) >> 15;
            } else {
              mlo = f2u(__fdiv_rn(blo, fc)) >> 15; mhi = f2u(__fdiv_rn(bhi, fc)) >> 15;
            }
            decided = (mlo == mhi) && (f2u(bhi) < 0x7F800000u);
            if (decided) {
              mask17 = mlo;
              type = type_from_max(submin ? __fsub_rn(lmx, mn) : lmx);   // max(x - min) == max(x) - min (rounding is monotone)
              if (tid == 0) { a.cmed[c] = u2f(mask17 << 15); a.ctype[c] = type; }   // only the 17 mask bits are defined
            }
          }
        }
        DBG_MARK(15);
        if (!decided) {                                    // block-uniform
          // Exact sequential sum: thread 0 walks the 256 transducers (one exact add each, runs of them pre-composed) and
          // chains the few threads that straddle a binade element by element.  SLOW tiles (negative / NaN / Inf) keep the
          // plain 4096-step chain.
          if (d0 < 0.0f) {                                 // only threads that get chained element by element stage their values
            #pragma unroll
            for (int e = 0; e < DCB_E; e++) slot[e] = x[e];
          }
          if (slow) {                                      // the reference's max with its NaN rule; FAST tiles use max(x) - min
            float ymx = __uint_as_float(0xff800000u);
            #pragma unroll
            for (int e = 0; e < DCB_E; e++) if (e < vcnt) ymx = fmaxf(ymx, x[e]);
            #pragma unroll
            for (int o = 16; o; o >>= 1) ymx = fmaxf(ymx, __shfl_xor_sync(0xffffffffu, ymx, o));
            __syncthreads();                               // every read of red_mx[par] by the stats reduction of this tile is done
            if (lane == 0) sm.red_mx[par][warp] = ymx;
          }
          // Runs of consecutive transducer threads in the same binade compose (associatively) into one transducer:
          // segmented warp scan.  (E then O)_p = E_p + O_{p ^ parity(E_p / q)}; every sum is exact (< 2^e, on the grid).
          {
            const bool certain = d0 >= 0.0f;
            const unsigned cm = __ballot_sync(0xffffffffu, certain);
            const uint32_t ebits = certain ? (f2u(lo_keep) >> 23) : 0u;
            const uint32_t eprev = __shfl_up_sync(0xffffffffu, ebits, 1);
            const bool head = certain && (lane == 0 || !((cm >> (lane - 1)) & 1u) || eprev != ebits);
            const unsigned hm = __ballot_sync(0xffffffffu, head);
            const int rs = 31 - __clz((int)(hm & (0xFFFFFFFFu >> (31 - lane))));    // first lane of this lane's run
            const float r0 = u2f(ebits << 23);
            #pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
              const float e0 = __shfl_up_sync(0xffffffffu, d0, o), e1 = __shfl_up_sync(0xffffffffu, d1, o);
              if (certain && lane - o >= rs) {
                const bool q0 = (f2u(__fadd_rn(r0, e0)) & 1u) != 0u, q1 = (f2u(__fadd_rn(r0, e1)) & 1u) != 0u;
                const float n0 = __fadd_rn(e0, q0 ? d1 : d0), n1 = __fadd_rn(e1, q1 ? d0 : d1);
                d0 = n0; d1 = n1;
              }
            }
            const bool tail = certain && (lane == 31 || ((hm >> (lane + 1)) & 1u) || !((cm >> (lane + 1)) & 1u));
            const unsigned wp = __ballot_sync(0xffffffffu, !certain || tail);       // where thread 0 has to stop
            if (lane == 0) sm.walkmask[warp] = wp;
          }
          slot[16] = d0; slot[17] = d1;
          __syncthreads();
          if (tid == 0) {
            float mx = submin ? __fsub_rn(lmx, mn) : lmx;  // FAST tile: max(x - min) == max(x) - min (rounding is monotone)
            if (slow) {
              mx = sm.red_mx[par][0];
              #pragma unroll
              for (int w = 1; w < DCB_NT / 32; w++) mx = fmaxf(mx, sm.red_mx[par][w]);
            }
            float total = 0.0f;
            const float* Y = sm.X[buf];
            // The stop list does not depend on the running sum: the next stop's transducer is fetched while the current
            // one is applied, so only the add (and an occasional 16-value chain) stays on the dependent path.
            int w = 0;
            unsigned m = sm.walkmask[0];
            int t = -1;
            float2 d = make_float2(0.0f, 0.0f);
            #define DCB_NEXT_STOP(T_, D_) do {                                                             \
                while (m == 0u && w + 1 < DCB_NT / 32) m = sm.walkmask[++w];                                \
                if (m) { T_ = w * 32 + (__ffs((int)m) - 1); m &= m - 1u;                                    \
                         D_ = *reinterpret_cast<const float2*>(Y + T_ * DCB_XPITCH + 16); }                 \
                else T_ = -1;                                                                               \
              } while (0)
            DCB_NEXT_STOP(t, d);
            while (t >= 0) {
              int tn; float2 dn = make_float2(0.0f, 0.0f);
              DCB_NEXT_STOP(tn, dn);
              if (d.x >= 0.0f) {
                total = __fadd_rn(total, (f2u(total) & 1u) ? d.y : d.x);      // exact
              } else {
                const float* yp = Y + t * DCB_XPITCH;
                const float4 v0 = *reinterpret_cast<const float4*>(yp), v1 = *reinterpret_cast<const float4*>(yp + 4);
                const float4 v2 = *reinterpret_cast<const float4*>(yp + 8), v3 = *reinterpret_cast<const float4*>(yp + 12);
                total = __fadd_rn(total, v0.x); total = __fadd_rn(total, v0.y); total = __fadd_rn(total, v0.z); total = __fadd_rn(total, v0.w);
                total = __fadd_rn(total, v1.x); total = __fadd_rn(total, v1.y); total = __fadd_rn(total, v1.z); total = __fadd_rn(total, v1.w);
                total = __fadd_rn(total, v2.x); total = __fadd_rn(total, v2.y); total = __fadd_rn(total, v2.z); total = __fadd_rn(total, v2.w);
                total = __fadd_rn(total, v3.x); total = __fadd_rn(total, v3.y); total = __fadd_rn(total, v3.z); total = __fadd_rn(total, v3.w);
              }
              t = tn; d = dn;
            }
            #undef DCB_NEXT_STOP
            const float med = __fdiv_rn(total, (float)cnt);
            const int ty = (Y[0] != Y[0]) ? 0 : type_from_max(mx);   // max starts at data[0]: NaN there defeats every '<'
            sm.med = med; sm.type = ty;
            a.cmed[c] = med; a.ctype[c] = ty;
          }
          __syncthreads();
          type = sm.type;
          mask17 = f2u(sm.med) >> 15;
        }
      } else {
        type = a.ctype[c];
        mask17 = f2u(a.cmed[c]) >> 15;
      }
    }

    DBG_MARK(7);
    // ---- 3. classify --------------------------------------------------------------------------------------
    uint32_t code[DCB_E];            // general path: top-aligned codes
    int len[DCB_E];
    uint32_t chi = 0u, clo = 0u;     // fast path: sixteen 3-bit fields, element 0 in the top bits of chi (0 = needs a fallback code)
    uint32_t total = 0;
    bool all3 = false;               // this thread holds sixteen 3-bit codes (48 bits)
    if (!slow) {
      if (CT != CT_NP) {
        float b1 = h1, b2 = h2, b3 = h3;
        float t1 = __fmul_rn(3.0f, b1), t2 = __fmul_rn(3.0f, b2);
        #pragma unroll
        for (int e = 0; e < DCB_E; e++) {
          const float xe = x[e];
          const float p2 = __fsub_rn(__fadd_rn(b1, b1), b2);                 // 2*b1 is exact: == __fmul_rn(2, b1)
          const float p3 = __fadd_rn(__fsub_rn(t1, t2), b3);
          const float d1 = fabsf(__fsub_rn(b1, xe));
          const float d2 = fabsf(__fsub_rn(p2, xe));
          const float d3 = fabsf(__fsub_rn(p3, xe));
          const float dmin = fminf(fminf(d1, d2), d3);                       // no NaN/Inf here: plain min == the '<' chain
          bool hit = dmin <= a.k.tp;                                         // diff_min <= absErrorBound, :3399
          if (e < 3) hit = hit && !first;
          const bool zero = xe <= a.k.tz;                                    // fabs(x) < absErrorBound (x >= 0), :3390
          const uint32_t sel = (d1 == dmin) ? 5u : ((d2 == dmin) ? 6u : 7u);  // first minimum wins == strict '<' updates
          const uint32_t f = zero ? 4u : (hit ? sel : 0u);
          if (e <= 9) chi |= f << (29 - 3 * e);
          else if (e == 10) { chi |= f >> 1; clo |= f << 31; }
          else clo |= f << (61 - 3 * e);
          b3 = b2; b2 = b1; b1 = xe;
          t2 = t1; t1 = __fmul_rn(3.0f, xe);
        }
        all3 = (vcnt == DCB_E) && ((chi & 0x92492492u) == 0x92492492u) && ((clo & 0x49240000u) == 0x49240000u);
      }
      if (__all_sync(0xffffffffu, all3)) {
        total = 48u;
      } else {
        all3 = false;
        #pragma unroll
        for (int e = 0; e < DCB_E; e++) {
          const uint32_t f = (CT == CT_NP) ? 0u : C48_FIELD(chi, clo, e);
          uint32_t ft; int fl;
          fallback_top<CT, true>(f2u(x[e]), a.k, type, mask17, ft, fl);
          code[e] = f ? (f << 29) : ft;
          len[e] = f ? 3 : fl;
          if (e >= vcnt) { code[e] = 0u; len[e] = 0; }
          total += (uint32_t)len[e];
        }
      }
    } else {
      // -1.0f is the reference's "empty history slot" marker: rebuild the exact state this thread
      // starts from by walking back over the chunk until three non-sentinel values are found.
      float b3 = -1.0f, b2 = -1.0f, b1 = -1.0f;
      if (vcnt > 0 && CT != CT_NP) {
        const uint64_t chunk_start = base - (uint64_t)tic * DCB_TILE;
        float n0 = 0.0f, n1 = 0.0f, n2 = 0.0f;     // newest first
        int found = 0; bool last_sent = false;
        for (uint64_t idx = my0; idx > chunk_start && found < 3; idx--) {
          float v = __ldg(a.in + idx - 1);
          if (submin) v = __fsub_rn(v, mn);
          if (idx == my0) last_sent = (v == -1.0f);
          if (v != -1.0f) {
            if (found == 0) n0 = v; else if (found == 1) n1 = v; else n2 = v;
            found++;
          }
        }
        if (found == 1) { b3 = n0; }
        else if (found == 2) { b3 = n1; b2 = n0; }
        else if (found == 3) {
          if (!last_sent) { b3 = n2; b2 = n1; b1 = n0; }
          else { b3 = n1; b2 = n0; b1 = -1.0f; }
        }
      }
      #pragma unroll
      for (int e = 0; e < DCB_E; e++) {
        len[e] = 0; code[e] = 0u;
        if (e < vcnt) classify_sentinel<CT>(x[e], b3, b2, b1, a.k, type, mask17, code[e], len[e]);
        total += (uint32_t)len[e];
      }
    }

    // ---- 4. block scan of the per-thread bit counts ---------------------------------------------------
    uint32_t incl = total;
    #pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += t;
    }
    if (lane == 31) sm.warp_tot[par][warp] = incl;
    DBG_MARK(14);
    // word-owner packing needs every thread that has bits to span a 32-bit word boundary (>= 32 bits)
    const bool fast_pack = __syncthreads_and(total >= 32u || total == 0u);
    uint32_t warp_base = 0, tile_bits = 0;
    #pragma unroll
    for (int w = 0; w < DCB_NT / 32; w++) {
      const uint32_t t = sm.warp_tot[par][w];
      if (w < warp) warp_base += t;
      tile_bits += t;
    }
    const uint32_t s = warp_base + incl - total;        // first bit of this thread inside the tile

    // publish the aggregate at once: the scanner turns it into an inclusive prefix while this block packs.
    // One tile per chunk: every tile ends byte aligned, so its contribution is simply its bits rounded up to bytes.
    if (tid == 0) {
      const bool single = (a.tiles_per_chunk == 1u);
      const unsigned long long v = single ? (unsigned long long)((tile_bits + 7u) & ~7u) : (unsigned long long)tile_bits;
      status[tile] = (ST_AGG << ST_SHIFT) | ((last_in_chunk && !single) ? ST_END_BIT : 0ull) | v;
    }

    DBG_MARK(8);
    // ---- 5. finish the tile packed TWO iterations ago: its start offset (status word loaded at the top), copy-out of its staging buffer ----
    if (it > 1u) {
      if (tid == DCB_RES_TID) resolve_pending(a, sm, buf, w0);
      __syncthreads();
      DBG_MARK(12);
      copy_out_tile(a, W, sm.P, sm.pend_bits[buf], sm.pend_last[buf] != 0u, tid);
      DBG_MARK(13);
      __syncthreads();
    }

    DBG_MARK(9);
    // ---- 6. pack the current tile into the staging buffer -------------------------------------------------
    uint32_t hi = 0u;                                   // on exit: the thread's unfinished last word (its tail)
    const uint32_t f0 = s & 31u;
    const uint32_t hw = 1u + (s >> 5);
    if (fast_pack) {
      uint32_t widx = hw;
      if (all3) {
        // 48-bit string chi:clo placed at bit f0 of word hw: three funnel shifts, at most two finished words
        const uint32_t w0_ = chi >> f0;
        const uint32_t w1_ = __funnelshift_r(clo, chi, f0);
        const uint32_t w2_ = __funnelshift_r(0u, clo, f0);
        W[hw] = w0_;
        if (f0 >= 16u) { W[hw + 1] = w1_; hi = w2_; widx = hw + 2; }
        else { hi = w1_; widx = hw + 1; }
      } else {
        uint32_t fill = f0;
        #pragma unroll
        for (int e = 0; e < DCB_E; e++) {
          const uint32_t cl = code[e];
          hi |= cl >> fill;
          const uint32_t lo = __funnelshift_r(0u, cl, fill);
          fill += (uint32_t)len[e];
          const bool fl = fill >= 32u;
          if (fl) W[widx] = hi;
          widx += fl ? 1u : 0u;
          hi = fl ? lo : hi;
          fill &= 31u;
        }
      }
      // the word straddling two threads is stored by the later one; it ORs in the earlier thread's tail
      const uint32_t prevT = __shfl_up_sync(0xffffffffu, hi, 1);
      if (lane > 0 && total > 0u && f0) W[hw] |= prevT;
      if (lane == 31) sm.xchg[warp] = hi;
      if (total > 0u && s + total == tile_bits) {          // last thread that has bits: flush its tail, zero what follows
        if ((s + total) & 31u) W[widx++] = hi;
        W[widx] = 0u; W[widx + 1] = 0u;
      }
      fix_need = (lane == 0 && warp > 0 && total > 0u && f0 != 0u);   // completed at the top of the next iteration
      fix_hw = hw;
    } else {
      // ragged tile (fewer than 11 valid elements in some thread): zero the buffer and OR every word in
      for (int i = tid + 1; i < DCB_TILE + 8; i += DCB_NT) W[i] = 0u;
      __syncthreads();
      uint32_t fill = f0, widx = hw;
      if (all3) {                                          // (cannot happen: all3 implies 48 bits; kept for safety)
        #pragma unroll
        for (int e = 0; e < DCB_E; e++) { code[e] = C48_FIELD(chi, clo, e) << 29; len[e] = 3; }
      }
      #pragma unroll
      for (int e = 0; e < DCB_E; e++) {
        const uint32_t L = (uint32_t)len[e];
        if (L) {
          const uint32_t cl = code[e];
          hi |= cl >> fill;
          const uint32_t lo = __funnelshift_r(0u, cl, fill);
          fill += L;
          if (fill >= 32u) { atomicOr(&W[widx++], hi); hi = lo; fill -= 32u; }
        }
      }
      if (fill && total > 0u) atomicOr(&W[widx], hi);
    }

    // ---- 7. rotate the pipeline --------------------------------------------------------------------------
    if (tid == 0) {
      sm.pend_tile[buf] = tile; sm.pend_bits[buf] = tile_bits; sm.pend_c[buf] = c; sm.pend_last[buf] = last_in_chunk ? 1u : 0u;
    }
    if (tid == DCB_TKT_TID) sm.tkt[buf] = tk;            // slot of the ticket just consumed
    par ^= 1;
    cur = nxt;
    DBG_MARK(10);
    __syncthreads();
    DBG_MARK(11);
    nxt = sm.tkt[buf];
    if (cur >= a.num_tiles) break;
  }

  // ---- epilogue: finish the last two tiles of this block (older first) --------------------------------------
  // here `it` is the index of the last iteration executed; its tile sits in W[it & 1], the one before in W[(it & 1) ^ 1]
  {
    const int lastb = (int)(it & 1u);
    if (fix_need) sm.W[lastb][fix_hw] |= sm.xchg[warp - 1];
    __syncthreads();
    for (int k = (it > 0u ? 0 : 1); k < 2; k++) {
      const int b = (k == 0) ? (lastb ^ 1) : lastb;
      if (tid == DCB_RES_TID) resolve_pending(a, sm, b, 0ull);
      __syncthreads();
      copy_out_tile(a, sm.W[b], sm.P, sm.pend_bits[b], sm.pend_last[b] != 0u, tid);
      __syncthreads();
    }
  }
}
