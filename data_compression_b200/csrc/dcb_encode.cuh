// dcb_encode.cuh -- single-pass encoder for CT=5/6/7/11 (replaces myCompress_bitwise{,_np,_mask,_op},
// dataCompression.c:3310, :2645, :2030, :577, and their per-bit add_bit_to_bytes :5456).
//
// One thread block = one TILE of 4096 consecutive elements, 16 per thread (blocked arrangement so a
// thread can concatenate its own codes in registers).  Steps:
//   1. 128-bit loads of the thread's 16 floats (+3-element look-back from the neighbouring lane)
//   2. [fused toSmallDataset_float / med_dataset_float when chunk == tile]
//   3. per element: zero test, best-of-three predictor on the ORIGINAL history (open loop),
//      else raw / escape / full word  ->  (code, len<=32)
//   4. block scan of the per-thread bit counts -> bit offset inside the tile
//   5. each thread shifts its codes into a 64-bit accumulator and stores finished 32-bit words into the
//      shared-memory staging buffer; the word straddling two threads is completed with one shuffle
//   6. decoupled look-back over the per-tile bit counts (chunk ends round up to a byte) -> absolute bit
//      offset of the tile in the output stream; chunk-end tiles write chunk_offsets[c+1]
//   7. staged words are funnel-shifted to the output alignment, byte-swapped to the reference's
//      MSB-first byte order and written as coalesced 32-bit words
#pragma once
#include "dcb_common.cuh"

// (code, len) of the fallback path for one element
template <int CT>
__device__ __forceinline__ void fallback_code(uint32_t u, const CodecConst& k, int type, uint32_t mask17,
                                              uint32_t& code, int& len) {
  if (CT == CT_OP) { code = u; len = 32; return; }                       // :689-692 full word
  const int lr = raw_len(u, k.n_eb_m118);
  if (CT == CT_MASK) {                                                   // compress_bitwise_float_mask :2143-2284
    if ((u >> 23) == (mask17 >> 8)) {
      const int m = lr - 9;
      const bool low8 = ((u >> 15) & 0xFFu) == (mask17 & 0xFFu);
      const int kbits = low8 ? max(m - 8, 0) : m;
      const uint32_t prefix = (((1u << type) - 1u) << 1) | (low8 ? 0u : 1u);   // 0, type x 1, flag
      const uint32_t pay = (u >> (23 - m)) & ((1u << kbits) - 1u);
      code = (prefix << kbits) | pay;
      len = type + 2 + kbits;
      return;
    }
  }
  code = (lr == 32) ? u : (u >> (32 - lr));                              // compress_bitwise_float :3505-3514
  len = lr;
}

// fast path: history = the three previous inputs, all known not to be the -1.0f sentinel
template <int CT>
__device__ __forceinline__ void classify(float x, float b1, float b2, float b3, bool predictable,
                                         const CodecConst& k, int type, uint32_t mask17,
                                         uint32_t& code, int& len) {
  const uint32_t u = f2u(x);
  if (CT == CT_NP) { fallback_code<CT_NP>(u, k, type, mask17, code, len); return; }
  const float p2 = pred2(b1, b2);
  const float p3 = pred3(b1, b2, b3);
  const float d1 = fabsf(__fsub_rn(b1, x));
  const float d2 = fabsf(__fsub_rn(p2, x));
  const float d3 = fabsf(__fsub_rn(p3, x));
  float dmin = d1; uint32_t sel = 5u;                                    // 'a' 101
  if (d2 < dmin) { dmin = d2; sel = 6u; }                                // 'b' 110 (strict <, :3373)
  if (d3 < dmin) { dmin = d3; sel = 7u; }                                // 'c' 111
  const bool zero = fabsf(x) <= k.tz;                                    // fabs(x) < absErrorBound in double, :3390
  const bool hit = predictable && (dmin <= k.tp);                        // diff_min <= absErrorBound, :3399
  uint32_t fc; int fl;
  fallback_code<CT>(u, k, type, mask17, fc, fl);
  code = zero ? 4u : (hit ? sel : fc);
  len = (zero || hit) ? 3 : fl;
}

// slow path: exact sentinel state machine of the reference, history carried in (b3,b2,b1)
template <int CT>
__device__ __forceinline__ void classify_sentinel(float x, float& b3, float& b2, float& b1,
                                                  const CodecConst& k, int type, uint32_t mask17,
                                                  uint32_t& code, int& len) {
  const uint32_t u = f2u(x);
  if (CT == CT_NP) { fallback_code<CT_NP>(u, k, type, mask17, code, len); return; }
  const bool zero = fabsf(x) <= k.tz;
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
    if (d2 < dmin) { dmin = d2; sel = 6u; }
    if (d3 < dmin) { dmin = d3; sel = 7u; }
    hit = dmin <= k.tp;
    b3 = b2; b2 = b1; b1 = x;                                            // :3386-3388
  }
  uint32_t fc; int fl;
  fallback_code<CT>(u, k, type, mask17, fc, fl);
  code = zero ? 4u : (hit ? sel : fc);
  len = (zero || hit) ? 3 : fl;
}

// ------------------------------------------------------------------------------------------------
// decoupled look-back over tiles (Merrill & Garland style), in BITS, with byte round-up at chunk ends.
// Called by warp 0.  Returns the absolute bit offset of tile `tile`'s first bit.
// Walking backwards from tile-1:  open = bits of the visited tiles up to and including the first
// (stream-order) chunk end among them, tail = everything after that chunk end (already well defined,
// it starts on a byte boundary).  See DESIGN.md "look-back with chunk ends".
__device__ __forceinline__ unsigned long long lookback_bits(const volatile unsigned long long* status,
                                                            uint32_t tile, int lane) {
  unsigned long long open = 0, tail = 0;
  bool have_end = false;
  int64_t j0 = (int64_t)tile - 1;                 // nearest predecessor handled by lane 0
  while (true) {
    const int64_t j = j0 - lane;
    // tiles before 0 read as "inclusive prefix 0"; only real tiles can still be INVALID
    unsigned long long w = (j >= 0) ? status[j] : (ST_INCL << ST_SHIFT);
    while (__any_sync(0xffffffffu, (w >> ST_SHIFT) == ST_INVALID)) {
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
struct EncSmem {
  uint32_t W[DCB_TILE + 8];      // W[0] = 0; stream word k of the tile lives at W[1 + k]
  uint32_t warp_tot[DCB_NT / 32];
  uint32_t xchg[DCB_NT / 32];
  float redf[DCB_NT / 32];
  uint32_t redu[DCB_NT / 32];
  unsigned long long P;
  uint32_t tile;
  float mn, med;
  int type;
};

__device__ __forceinline__ float block_min(float v, EncSmem& sm, int lane, int warp) {
  #pragma unroll
  for (int o = 16; o; o >>= 1) v = fminf(v, __shfl_xor_sync(0xffffffffu, v, o));
  if (lane == 0) sm.redf[warp] = v;
  __syncthreads();
  float r = sm.redf[0];
  #pragma unroll
  for (int w = 1; w < DCB_NT / 32; w++) r = fminf(r, sm.redf[w]);
  __syncthreads();
  return r;
}
__device__ __forceinline__ float block_max(float v, EncSmem& sm, int lane, int warp) {
  #pragma unroll
  for (int o = 16; o; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  if (lane == 0) sm.redf[warp] = v;
  __syncthreads();
  float r = sm.redf[0];
  #pragma unroll
  for (int w = 1; w < DCB_NT / 32; w++) r = fmaxf(r, sm.redf[w]);
  __syncthreads();
  return r;
}
__device__ __forceinline__ uint32_t block_min_u32(uint32_t v, EncSmem& sm, int lane, int warp) {
  #pragma unroll
  for (int o = 16; o; o >>= 1) v = min(v, __shfl_xor_sync(0xffffffffu, v, o));
  if (lane == 0) sm.redu[warp] = v;
  __syncthreads();
  uint32_t r = sm.redu[0];
  #pragma unroll
  for (int w = 1; w < DCB_NT / 32; w++) r = min(r, sm.redu[w]);
  __syncthreads();
  return r;
}

#define DCB_ENC_STATS_READY 0x100u   // EncArgs.flags: cmin / cmed / ctype were filled by the stats pre-pass
#define DCB_ENC_ATOMIC_OUT  0x200u   // tile boundaries are not byte aligned: partial words go out with atomicOr
#define DCB_ENC_IN_ALIGNED  0x400u   // input pointer is 16-byte aligned (128-bit loads)

template <int CT>
__global__ void __launch_bounds__(DCB_NT) dcb_encode_kernel(const EncArgs a) {
  __shared__ __align__(16) EncSmem sm;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) { sm.tile = atomicAdd(a.ticket, 1u); sm.W[0] = 0u; }   // dynamic tile id: look-back only waits on started tiles
  __syncthreads();
  const uint32_t tile = sm.tile;
  const uint64_t base = (uint64_t)tile * DCB_TILE;

  // ---- chunk geometry --------------------------------------------------------------------------
  uint64_t c = 0, chunk_start = 0;
  bool last_in_chunk = (tile == a.num_tiles - 1);
  if (a.tiles_per_chunk) {
    c = tile / a.tiles_per_chunk;
    chunk_start = c * a.chunk;
    last_in_chunk = last_in_chunk || ((tile % a.tiles_per_chunk) == a.tiles_per_chunk - 1);
  }
  const bool submin = (a.flags & 1u) != 0;            // DCB_F_SUBTRACT_MIN
  const bool fixed_mask = (a.flags & 2u) != 0;        // DCB_F_FIXED_MASK
  const bool stats_ready = (a.flags & DCB_ENC_STATS_READY) != 0;

  // ---- 1. load 16 consecutive floats per thread ---------------------------------------------------
  const uint64_t my0 = base + (uint64_t)tid * DCB_E;
  const int vcnt = (my0 >= a.n) ? 0 : (int)min((uint64_t)DCB_E, a.n - my0);
  const uint32_t ci = (uint32_t)min((uint64_t)0xFFFFFFF0u, my0 - chunk_start);  // index in chunk (only compared with 3)
  float x[DCB_E];
  if (vcnt == DCB_E && (a.flags & DCB_ENC_IN_ALIGNED)) {
    const float4* p = reinterpret_cast<const float4*>(a.in + my0);
    #pragma unroll
    for (int q = 0; q < 4; q++) {
      const float4 v = __ldg(p + q);
      x[4 * q] = v.x; x[4 * q + 1] = v.y; x[4 * q + 2] = v.z; x[4 * q + 3] = v.w;
    }
  } else {
    #pragma unroll
    for (int e = 0; e < DCB_E; e++) x[e] = (e < vcnt) ? __ldg(a.in + my0 + e) : 0.0f;
  }

  // ---- 2a. toSmallDataset_float (dataCompression.c:3543-3562), fused when the tile is the whole chunk --
  float mn = 0.0f;
  if (submin) {
    if (!stats_ready) {
      float lm = __uint_as_float(0x7f800000u);
      #pragma unroll
      for (int e = 0; e < DCB_E; e++) if (e < vcnt) lm = fminf(lm, x[e]);        // NaN never wins a '<' (:3550)
      mn = block_min(lm, sm, lane, warp);
      if (mn == 0.0f) {
        // +0 / -0 compare equal: the reference keeps the FIRST minimal element; recover its sign
        uint32_t fi = 0xFFFFFFFFu;
        float fz = 0.0f;
        #pragma unroll
        for (int e = DCB_E - 1; e >= 0; e--)
          if (e < vcnt && x[e] == 0.0f) { fi = (uint32_t)tid * DCB_E + e; fz = x[e]; }
        const uint32_t fmin = block_min_u32(fi, sm, lane, warp);
        if (fi == fmin) sm.mn = fz;
        __syncthreads();
        mn = sm.mn;
        __syncthreads();
      }
      if (tid == 0) { if (x[0] != x[0]) sm.mn = x[0]; else sm.mn = mn; }           // data[0] NaN stays the min
      __syncthreads();
      mn = sm.mn;
      if (tid == 0) a.cmin[c] = mn;
    } else {
      mn = a.cmin[c];
    }
    #pragma unroll
    for (int e = 0; e < DCB_E; e++) x[e] = __fsub_rn(x[e], mn);
  } else if (a.cmin && tid == 0 && last_in_chunk) {
    a.cmin[c] = 0.0f;
  }

  // ---- look-back values x[-1], x[-2], x[-3] of the thread's first element ------------------------------
  float h1 = __shfl_up_sync(0xffffffffu, x[DCB_E - 1], 1);
  float h2 = __shfl_up_sync(0xffffffffu, x[DCB_E - 2], 1);
  float h3 = __shfl_up_sync(0xffffffffu, x[DCB_E - 3], 1);
  if (lane == 0) {
    h1 = h2 = h3 = 0.0f;
    if (vcnt > 0 && CT != CT_NP) {
      if (ci >= 1) h1 = __ldg(a.in + my0 - 1);
      if (ci >= 2) h2 = __ldg(a.in + my0 - 2);
      if (ci >= 3) h3 = __ldg(a.in + my0 - 3);
      if (submin) { h1 = __fsub_rn(h1, mn); h2 = __fsub_rn(h2, mn); h3 = __fsub_rn(h3, mn); }
    }
  }

  // ---- 2b. med_dataset_float (:3593-3620): SEQUENTIAL float32 sum (order matters), max -> type -----------
  int type = a.fixed_type;
  uint32_t mask17 = a.fixed_mask17;
  if (CT == CT_MASK && !fixed_mask) {
    if (!stats_ready) {
      float* Wf = reinterpret_cast<float*>(sm.W + 4);   // 16-byte aligned alias of the staging buffer
      #pragma unroll
      for (int e = 0; e < DCB_E; e++) Wf[tid * DCB_E + e] = x[e];
      float lmx = __uint_as_float(0xff800000u);
      #pragma unroll
      for (int e = 0; e < DCB_E; e++) if (e < vcnt) lmx = fmaxf(lmx, x[e]);
      const float mx = block_max(lmx, sm, lane, warp);       // (barriers inside also publish Wf)
      if (tid == 0) {
        const int cnt = (int)min((uint64_t)DCB_TILE, a.n - base);
        float total = 0.0f;
        int i = 0;
        for (; i + 4 <= cnt; i += 4) {
          const float4 v = *reinterpret_cast<const float4*>(Wf + i);
          total = __fadd_rn(total, v.x); total = __fadd_rn(total, v.y);
          total = __fadd_rn(total, v.z); total = __fadd_rn(total, v.w);
        }
        for (; i < cnt; i++) total = __fadd_rn(total, Wf[i]);
        const float med = __fdiv_rn(total, (float)cnt);
        const int ty = (Wf[0] != Wf[0]) ? 0 : type_from_max(mx);   // max starts at data[0]: NaN there defeats every '<'
        sm.med = med; sm.type = ty;
        a.cmed[c] = med; a.ctype[c] = ty;
      }
      __syncthreads();
      type = sm.type;
      mask17 = f2u(sm.med) >> 15;
      __syncthreads();
    } else {
      type = a.ctype[c];
      mask17 = f2u(a.cmed[c]) >> 15;
    }
  }

  // ---- 3. classify: (code, len) per element ---------------------------------------------------------
  uint32_t code[DCB_E];
  uint32_t lens[DCB_E / 4] = {0u, 0u, 0u, 0u};
  uint32_t total = 0;
  bool sentinel = false;
  if (CT != CT_NP) {
    #pragma unroll
    for (int e = 0; e < DCB_E; e++) sentinel |= (e < vcnt) && (x[e] == -1.0f);
    sentinel |= (vcnt > 0) && ((ci >= 1 && h1 == -1.0f) || (ci >= 2 && h2 == -1.0f) || (ci >= 3 && h3 == -1.0f));
    sentinel = __syncthreads_or(sentinel);
  }
  if (!sentinel) {
    #pragma unroll
    for (int e = 0; e < DCB_E; e++) {
      const float b1 = (e >= 1) ? x[e - 1] : h1;
      const float b2 = (e >= 2) ? x[e - 2] : ((e == 1) ? h1 : h2);
      const float b3 = (e >= 3) ? x[e - 3] : ((e == 2) ? h1 : ((e == 1) ? h2 : h3));
      int len;
      classify<CT>(x[e], b1, b2, b3, ci + e >= 3u, a.k, type, mask17, code[e], len);
      if (e >= vcnt) len = 0;
      lens[e >> 2] |= (uint32_t)len << (8 * (e & 3));
      total += len;
    }
  } else {
    // -1.0f is the reference's "empty history slot" marker: rebuild the exact state this thread
    // starts from by walking back over the chunk until three non-sentinel values are found.
    float b3 = -1.0f, b2 = -1.0f, b1 = -1.0f;
    if (vcnt > 0) {
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
      int len = 0; code[e] = 0;
      if (e < vcnt) classify_sentinel<CT>(x[e], b3, b2, b1, a.k, type, mask17, code[e], len);
      lens[e >> 2] |= (uint32_t)len << (8 * (e & 3));
      total += len;
    }
  }

  // ---- 4. block scan of the per-thread bit counts ---------------------------------------------------
  uint32_t incl = total;
  #pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  if (lane == 31) sm.warp_tot[warp] = incl;
  // fast packing needs every thread that has bits to span a 32-bit word boundary (>= 32 bits)
  const bool fast_pack = __syncthreads_and(total >= 32u || total == 0u);
  uint32_t warp_base = 0, tile_bits = 0;
  #pragma unroll
  for (int w = 0; w < DCB_NT / 32; w++) {
    const uint32_t t = sm.warp_tot[w];
    if (w < warp) warp_base += t;
    tile_bits += t;
  }
  const uint32_t s = warp_base + incl - total;        // first bit of this thread inside the tile

  // ---- 6a. publish the aggregate, then warp 0 looks back while the other warps pack ----------------------
  volatile unsigned long long* status = a.status;
  if (warp == 0) {
    unsigned long long P = 0;
    if (tile > 0) {
      if (lane == 0)
        status[tile] = (ST_AGG << ST_SHIFT) | (last_in_chunk ? ST_END_BIT : 0ull) | (unsigned long long)tile_bits;
      P = lookback_bits(status, tile, lane);
    }
    if (lane == 0) {
      unsigned long long E = P + tile_bits;
      if (last_in_chunk) E = (E + 7ull) & ~7ull;
      status[tile] = (ST_INCL << ST_SHIFT) | E;
      sm.P = P;
      if (tile == 0) a.offsets[0] = 0ull;
      if (last_in_chunk) a.offsets[c + 1] = E >> 3;
    }
  }

  // ---- 5. pack ---------------------------------------------------------------------------------------
  uint32_t hi = 0, lo = 0;
  uint32_t fill = s & 31u;
  uint32_t widx = 1u + (s >> 5);
  const uint32_t hw = widx;
  if (fast_pack) {
    #pragma unroll
    for (int e = 0; e < DCB_E; e++) {
      const uint32_t L = (lens[e >> 2] >> (8 * (e & 3))) & 0xFFu;
      if (L) {
        const uint32_t cl = code[e] << (32u - L);
        hi |= cl >> fill;
        lo = __funnelshift_r(0u, cl, fill);
        fill += L;
        if (fill >= 32u) { sm.W[widx++] = hi; hi = lo; lo = 0u; fill -= 32u; }
      }
    }
    // the word straddling two threads is stored by the later one; it ORs in the earlier thread's tail
    const uint32_t prevT = __shfl_up_sync(0xffffffffu, hi, 1);
    if (lane > 0 && total > 0u && (s & 31u)) sm.W[hw] |= prevT;
    if (lane == 31) sm.xchg[warp] = hi;
    if (total > 0u && s + total == tile_bits) {          // last thread that has bits: flush its tail, zero what follows
      if (fill) sm.W[widx++] = hi;
      sm.W[widx] = 0u; sm.W[widx + 1] = 0u;
    }
    __syncthreads();
    if (lane == 0 && warp > 0 && total > 0u && (s & 31u)) sm.W[hw] |= sm.xchg[warp - 1];
    __syncthreads();
  } else {
    // ragged tile (fewer than 11 valid elements in some thread): zero the buffer and OR every word in
    for (int i = tid; i < DCB_TILE + 8; i += DCB_NT) sm.W[i] = 0u;
    __syncthreads();
    #pragma unroll
    for (int e = 0; e < DCB_E; e++) {
      const uint32_t L = (lens[e >> 2] >> (8 * (e & 3))) & 0xFFu;
      if (L) {
        const uint32_t cl = code[e] << (32u - L);
        hi |= cl >> fill;
        lo = __funnelshift_r(0u, cl, fill);
        fill += L;
        if (fill >= 32u) { atomicOr(&sm.W[widx++], hi); hi = lo; lo = 0u; fill -= 32u; }
      }
    }
    if (fill && total > 0u) atomicOr(&sm.W[widx], hi);
    __syncthreads();
  }

  // ---- 7. copy out: shift to the output bit alignment, MSB-first byte order, coalesced words ---------------
  const unsigned long long P = sm.P;
  unsigned long long Eabs = P + tile_bits;
  if (last_in_chunk) Eabs = (Eabs + 7ull) & ~7ull;          // zero padding of the chunk's last byte (:5467)
  if (Eabs == P) return;
  const unsigned long long g0 = P >> 5, g1 = (Eabs - 1ull) >> 5;
  const uint32_t sh = (uint32_t)(P & 31ull), tb = (uint32_t)(Eabs & 31ull);
  uint32_t* out32 = reinterpret_cast<uint32_t*>(a.out);
  const bool atomic_out = (a.flags & DCB_ENC_ATOMIC_OUT) != 0;
  for (unsigned long long g = g0 + tid; g <= g1; g += DCB_NT) {
    const uint32_t k = (uint32_t)(g - g0);
    uint32_t v = sh ? __funnelshift_r(sm.W[1 + k], sm.W[k], sh) : sm.W[1 + k];
    const bool head = (g == g0) && sh, tail = (g == g1) && tb;
    if (!head && !tail) {
      out32[g] = __byte_perm(v, 0u, 0x0123);
    } else {
      const uint32_t lo_bit = head ? sh : 0u, hi_bit = tail ? tb : 32u;   // valid bits [lo_bit, hi_bit), MSB first
      uint32_t mask = 0xFFFFFFFFu >> lo_bit;
      if (hi_bit < 32u) mask &= ~(0xFFFFFFFFu >> hi_bit);
      v &= mask;
      if (atomic_out) {
        atomicOr(&out32[g], __byte_perm(v, 0u, 0x0123));
      } else {
        #pragma unroll
        for (uint32_t b = 0; b < 4; b++)
          if (8u * b >= lo_bit && 8u * b + 8u <= hi_bit) a.out[g * 4ull + b] = (uint8_t)(v >> (24u - 8u * b));
      }
    }
  }
}
