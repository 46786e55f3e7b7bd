// dcb_encode_slots.cuh -- encoder for the common shape "one chunk == one 4096-element tile" (CT=5/6/7/11).
// Replaces myCompress_bitwise{,_np,_mask,_op} (dataCompression.c:3310, :2645, :2030, :577) + toSmallDataset_float (:3543)
// + med_dataset_float (:3593) per chunk, like dcb_encode.cuh, but WITHOUT any coupling between thread blocks:
//
//   pass 1  dcb_encode_slots_kernel : every chunk is classified, packed and written to a FIXED-PITCH slot (16 KB, the worst
//                                     case of a chunk) of a scratch area; its byte count goes to sizes[] and is added to the
//                                     sum of its group of 64 chunks.
//   pass 2  dcb_compact_kernel      : one block per group: prefix of the group sums + local scan = chunk offsets (the chunk
//                                     table), then the chunks are copied to their final, byte-granular place in the stream.
//
// Why: in the single-pass encoder (dcb_encode.cuh) a tile can only be written once the sizes of all tiles before it are
// known; tile offsets therefore retire IN ORDER, and the r1 profiles show what that costs -- a ticket, a status word, a
// scanner block, a copy-out deferred by two iterations behind five barriers, and every slow tile (CT=7's exact walk)
// stalls the blocks behind it (mean iteration 7.2 us against 5.6 us of work).  Here a block never waits for another one:
// two barriers per tile (statistics, scan), tiles from a ticket counter (pure load balancing), the packed tile leaves as aligned 128-bit stores
// without a funnel shift.  The price is pass 2: the compressed bytes are written and read once more (2B of extra traffic,
// B = 4n/r; 0.2 GB for the smooth field, most of it L2 hits), which is far less than what the coupling cost.
#pragma once
#include "dcb_encode.cuh"
#include "dcb_crc.cuh"

#define DCB_SLOT_PITCH 16384u      // bytes per scratch slot: 4096 codes of at most 32 bits
#define DCB_GROUP 64u              // chunks per compaction block
#define DCB_SLOT_BLOCKS_PER_SM 3

#ifdef DCB_DEBUG_MED
__device__ unsigned long long g_dbg_walk[8];   // cycles: whole level 2, serial loop; counts; stops
#endif
struct SlotSmem {
  float X[2][DCB_NT * DCB_XPITCH];   // double-buffered input tiles, one 80-byte slot per thread
  uint32_t W[2][DCB_TILE + 8];       // packed tiles (double buffered): stream word k of the tile lives at W[b][k]
  float red_mn[2][DCB_NT / 32];
  float red_mx[2][DCB_NT / 32];
  float red_am[2][DCB_NT / 32];
  float red_t[2][DCB_NT / 32];       // CT=7 level 0: tree sum of the raw values
  float red_s[2][DCB_NT / 32];
  float red_w[2][DCB_NT / 32];
  float red_e[2][DCB_NT / 32];
  uint32_t warp_tot[2][DCB_NT / 32];
  uint32_t xchg[DCB_NT / 32];        // tail word of every warp's last lane (completes the next warp's first word)
  float hx[2][DCB_NT / 32][4];       // last three RAW values of every warp
  uint32_t walkmask[DCB_NT / 32];
  float mn, med;
  int type;
  uint32_t fi;
  uint32_t tkt[2];                   // tile tickets: prologue pair, then the ticket of the tile after next (by parity)
};

// packed tile -> its scratch slot: MSB-first byte order, aligned 128-bit stores, no shift (a slot starts a chunk)
__device__ __forceinline__ void copy_out_slot(uint8_t* slot, const uint32_t* W, uint32_t bits, int tid) {
  const uint32_t units = (bits + 127u) >> 7;
  uint4* o = reinterpret_cast<uint4*>(slot);
  #pragma unroll 2
  for (uint32_t u = (uint32_t)tid; u < units; u += DCB_NT) {
    uint4 v = *reinterpret_cast<const uint4*>(W + 4u * u);
    v.x = __byte_perm(v.x, 0u, 0x0123); v.y = __byte_perm(v.y, 0u, 0x0123);
    v.z = __byte_perm(v.z, 0u, 0x0123); v.w = __byte_perm(v.w, 0u, 0x0123);
    o[u] = v;
  }
}

// One element of the FAST classification.  The 3-bit field (preset to 100 = "zero") is raised to 101 / 110 / 111 by up to
// three predicated multiply-adds, a miss sets the element's bit in `miss`.  Written in PTX so that the compares chain
// through predicates (FSETP ... .AND P) and the field updates are IMADs: the ALU pipe (FSETP, SEL, FMNMX, LOP3) is the busy
// one in this kernel, the FMA pipe (FADD, FMUL, IMAD) has room (profiles/r1_summary.md) -- 6 ALU-pipe instructions per
// element instead of 10.  zero: x <= tz; hit: min |e_i| <= tp; the FIRST minimum wins (strict '<' updates, :3373).
template <uint32_t FIELD_ONE>
__device__ __forceinline__ void classify_pred(uint32_t& fields, uint32_t& miss, float xe, float e1, float e2, float e3,
                                              float tz, float tp, uint32_t one, uint32_t miss_bit) {
  asm("{\n\t"
      ".reg .pred pnz, ph, pa, pb, pm;\n\t"
      ".reg .f32 a1, a2, a3, dm;\n\t"
      "abs.f32 a1, %3;\n\t"
      "abs.f32 a2, %4;\n\t"
      "abs.f32 a3, %5;\n\t"
      "min.f32 dm, a1, a2, a3;\n\t"                /* sm_100: one FMNMX3 */
      "setp.gt.f32 pnz, %2, %6;\n\t"
      "setp.le.and.f32 ph|pm, dm, %7, pnz;\n\t"      /* pm = !(dm <= tp) && pnz: one FSETP, two predicates */
      "setp.neu.and.f32 pa, a1, dm, ph;\n\t"
      "setp.neu.and.f32 pb, a2, dm, pa;\n\t"
      "@ph mad.lo.u32 %0, %8, %9, %0;\n\t"
      "@pa mad.lo.u32 %0, %8, %9, %0;\n\t"
      "@pb mad.lo.u32 %0, %8, %9, %0;\n\t"
      "@pm mad.lo.u32 %1, %8, %10, %1;\n\t"
      "}"
      : "+r"(fields), "+r"(miss)
      : "f"(xe), "f"(e1), "f"(e2), "f"(e3), "f"(tz), "f"(tp), "r"(one), "n"(FIELD_ONE), "r"(miss_bit));
}

template <int CT>
__global__ void __launch_bounds__(DCB_NT, DCB_SLOT_BLOCKS_PER_SM) dcb_encode_slots_kernel(const EncArgs a) {
  extern __shared__ __align__(16) unsigned char dcb_slot_smem[];
  SlotSmem& sm = *reinterpret_cast<SlotSmem*>(dcb_slot_smem);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  // Tiles are handed out by a ticket counter: their cost differs by 2x (CT=7: a chunk whose mask needs the exact walk, a
  // noisy chunk next to a constant one), and with a static stride the periodic bench field put every expensive chunk on a
  // quarter of the blocks (444 = 12 mod 16).  No block ever waits for another one, so a ticket is only load balancing.
  if (tid == 0) { sm.tkt[0] = atomicAdd(a.ticket, 1u); sm.tkt[1] = atomicAdd(a.ticket, 1u); }
  __syncthreads();
  uint32_t cur = sm.tkt[0], nxt = sm.tkt[1];
  if (cur >= a.num_tiles) return;
  issue_tile_load(a, cur, sm.X[0], tid);
  cp_async_commit();

  const bool submin = (a.flags & 1u) != 0;            // DCB_F_SUBTRACT_MIN
  const bool fixed_mask = (a.flags & 2u) != 0;        // DCB_F_FIXED_MASK
  const uint32_t one = (a.flags >> 31) + 1u;          // 1, but opaque to the compiler (keeps the field updates on IMAD)

  uint32_t fix_hw = 0; bool fix_need = false;         // lane 0 of warps > 0: first word of the pending tile still lacks the previous warp's tail
  uint32_t pend_tile = 0u, pend_bits = 0u;            // tile packed in the previous iteration (copied out in this one)
  int par = 0;

  uint32_t it = 0;
  for (;; it++) {
    const int buf = (int)(it & 1u);
    float* slot = &sm.X[buf][tid * DCB_XPITCH];
    if (nxt < a.num_tiles) issue_tile_load(a, nxt, sm.X[buf ^ 1], tid);
    cp_async_commit();
    uint32_t tk = 0u;
    if (tid == DCB_TKT_TID) tk = atomicAdd(a.ticket, 1u);               // ticket of the tile after next (lands during this tile)
    uint32_t* const W = sm.W[buf];

    const uint32_t tile = cur, c = cur;
    const uint64_t base = (uint64_t)tile * DCB_TILE;
    const bool full = base + DCB_TILE <= a.n;
    const uint64_t my0 = base + (uint64_t)tid * DCB_E;
    const int vcnt = full ? DCB_E : ((my0 >= a.n) ? 0 : (int)min((uint64_t)DCB_E, a.n - my0));
    const bool first = (tid == 0);                     // chunk start: elements 0..2 have no full history (:3330)
    const int cnt = (int)min((uint64_t)DCB_TILE, a.n - base);

    // ---- 1. this thread's 16 floats ----------------------------------------------------------------------
    cp_async_wait<1>();
    __syncwarp();
    float x[DCB_E];
    #pragma unroll
    for (int q = 0; q < 4; q++) {
      const float4 v = *reinterpret_cast<const float4*>(slot + 4 * q);
      x[4 * q] = v.x; x[4 * q + 1] = v.y; x[4 * q + 2] = v.z; x[4 * q + 3] = v.w;
    }

    // ---- 2. one block reduction: min, [CT=7: max, prefix sums of the raw values]; FAST test by a barrier vote -------
    // nacc turns Inf/NaN iff the thread holds a NaN, an Inf or |x| >= 2^127 (2|x| overflows) -- or, harmlessly, when the
    // sum of eight huge values overflows: such tiles take the exact SLOW path.  FFMA chain: the FMA pipe has room, the
    // ALU pipe (FMNMX) is the busy one (profiles/r1_summary.md).
    float lmn = __uint_as_float(0x7f800000u), lmx = __uint_as_float(0xff800000u);
    float nacc0 = 0.0f, nacc1 = 0.0f;
    if (full) {
      #pragma unroll
      for (int e = 0; e < DCB_E; e++) {
        lmn = fminf(lmn, x[e]);
        if (CT == CT_MASK) lmx = fmaxf(lmx, x[e]);
      }
    } else {
      #pragma unroll
      for (int e = 0; e < DCB_E; e++)
        if (e < vcnt) { lmn = fminf(lmn, x[e]); if (CT == CT_MASK) lmx = fmaxf(lmx, x[e]); }   // NaN never wins a '<' (:3550)
    }
    #pragma unroll
    for (int e = 0; e < DCB_E; e += 2) {
      nacc0 = __fmaf_rn(fabsf(x[e]), 2.0f, nacc0);
      nacc1 = __fmaf_rn(fabsf(x[e + 1]), 2.0f, nacc1);
    }
    bool odd = !(fabsf(__fadd_rn(nacc0, nacc1)) < __uint_as_float(0x7f800000u));
    if (!submin) {                                      // without min-subtract a negative value (sign bit, -0.0f included) makes the tile SLOW
      uint32_t sg = 0u;
      #pragma unroll
      for (int e = 0; e < DCB_E; e++) sg |= f2u(x[e]);
      odd = odd || ((sg >> 31) != 0u);
    }
    float px = 0.0f, inc = 0.0f;                        // CT=7: tree sum of this thread's raw values, inclusive scan over the warp
    if (CT == CT_MASK) {                                // pairwise: 4 levels (padding of a ragged tile is 0.0f)
      const float s0 = __fadd_rn(__fadd_rn(x[0], x[1]), __fadd_rn(x[2], x[3]));
      const float s1 = __fadd_rn(__fadd_rn(x[4], x[5]), __fadd_rn(x[6], x[7]));
      const float s2 = __fadd_rn(__fadd_rn(x[8], x[9]), __fadd_rn(x[10], x[11]));
      const float s3 = __fadd_rn(__fadd_rn(x[12], x[13]), __fadd_rn(x[14], x[15]));
      px = __fadd_rn(__fadd_rn(s0, s1), __fadd_rn(s2, s3));
      inc = px;
      #pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const float t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc = __fadd_rn(inc, t);
      }
    }
    if (lane == 31) {
      sm.hx[par][warp][0] = x[DCB_E - 1]; sm.hx[par][warp][1] = x[DCB_E - 2]; sm.hx[par][warp][2] = x[DCB_E - 3];
      if (CT == CT_MASK) sm.red_t[par][warp] = inc;
    }
    #pragma unroll
    for (int o = 16; o; o >>= 1) {
      lmn = fminf(lmn, __shfl_xor_sync(0xffffffffu, lmn, o));
      if (CT == CT_MASK) lmx = fmaxf(lmx, __shfl_xor_sync(0xffffffffu, lmx, o));
    }
    if (lane == 0) {
      sm.red_mn[par][warp] = lmn;
      if (CT == CT_MASK) sm.red_mx[par][warp] = lmx;
    }
    bool slow = __syncthreads_or(odd);                                    // barrier 1
    lmn = sm.red_mn[par][0];
    float wb = 0.0f, tsum = 0.0f;                       // CT=7: raw sum of the warps before this one, of the whole tile
    if (CT == CT_MASK) { lmx = sm.red_mx[par][0]; tsum = sm.red_t[par][0]; }
    #pragma unroll
    for (int w = 1; w < DCB_NT / 32; w++) {
      lmn = fminf(lmn, sm.red_mn[par][w]);
      if (CT == CT_MASK) {
        lmx = fmaxf(lmx, sm.red_mx[par][w]);
        if (w == warp) wb = tsum;
        tsum = __fadd_rn(tsum, sm.red_t[par][w]);
      }
    }
    // the previous tile's warp-boundary words (its tails were parked in xchg before this barrier)
    if (fix_need) sm.W[buf ^ 1][fix_hw] |= sm.xchg[warp - 1];
    fix_need = false;

    float mn = submin ? lmn : 0.0f;
    if (submin) {
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
      if (full) {
        #pragma unroll
        for (int e = 0; e < DCB_E; e++) x[e] = __fsub_rn(x[e], mn);
      } else {
        #pragma unroll
        for (int e = 0; e < DCB_E; e++) x[e] = (e < vcnt) ? __fsub_rn(x[e], mn) : 0.0f;
      }
    } else if (a.cmin && tid == 0) {
      a.cmin[c] = 0.0f;
    }

    // ---- look-back values x[-1], x[-2], x[-3] of the thread's first element ------------------------------
    float h1 = __shfl_up_sync(0xffffffffu, x[DCB_E - 1], 1);
    float h2 = __shfl_up_sync(0xffffffffu, x[DCB_E - 2], 1);
    float h3 = __shfl_up_sync(0xffffffffu, x[DCB_E - 3], 1);
    if (lane == 0) {
      if (warp > 0) {
        h1 = sm.hx[par][warp - 1][0]; h2 = sm.hx[par][warp - 1][1]; h3 = sm.hx[par][warp - 1][2];
        if (submin) { h1 = __fsub_rn(h1, mn); h2 = __fsub_rn(h2, mn); h3 = __fsub_rn(h3, mn); }
      } else { h1 = 0.0f; h2 = 0.0f; h3 = 0.0f; }      // chunk start: never used (hits of elements 0..2 are suppressed)
    }

    // ---- 2b. med_dataset_float (:3593-3620): SEQUENTIAL float32 sum (order matters), max -> type -----------
    int type = a.fixed_type;
    uint32_t mask17 = a.fixed_mask17;
    if (CT == CT_MASK && !fixed_mask) {
      bool decided = false;
      const float u24 = 1.0f / 16777216.0f;
      const float fc = (float)cnt;
      const float amb = fmaxf(fabsf(lmn), fabsf(lmx));   // largest |x| of the tile (FAST tiles: finite)
      // LEVEL 0 (no extra barrier).  Only the top 17 bits of medium reach the stream.  With y_i = fl(x_i - min) >= 0:
      //   sum_i y_i = (X - cnt*min)(1 +- u),  X = the real sum of the x_i = tsum +- 16 u sum|x_i|   (tree of depth 16),
      //   S_seq = sum_i y_i (1 +- gamma_{cnt-1})                                                    (recursive summation),
      // so S_seq lies in est +- E with est = fl(tsum - fl(cnt*min)), E1 = 24 u cnt max|x| (tree, product, difference),
      // E = E1 + (cnt + 4) * 1.01 u (est + E1); u = 2^-24.  If both ends give the same 17 bits, that is the mask.
      if (!slow && !(a.flags & 8u)) {
        float est = __fsub_rn(tsum, __fmul_rn(fc, mn));
        float E1 = __fadd_ru(__fmul_ru(__fmul_ru(fc, amb), 24.0f * u24), 1e-37f);   // + floor: roundings of denormal results are absolute
        if (submin && lmx == lmn) { est = 0.0f; E1 = 0.0f; }      // constant chunk: every y is +0 and so is the sum
        const float top = fmaxf(__fadd_ru(est, E1), 0.0f);
        const float E = __fadd_ru(E1, __fmul_ru(__fmul_ru(__fadd_ru(fc, 4.0f), 1.01f * u24), top));
        const float flo = fmaxf(__fsub_rd(est, E), 0.0f), fhi = __fadd_ru(est, E);
        uint32_t mlo, mhi;
        if (full) { mlo = f2u(__fmul_rn(flo, 1.0f / 4096.0f)) >> 15; mhi = f2u(__fmul_rn(fhi, 1.0f / 4096.0f)) >> 15; }   // == division
        else { mlo = f2u(__fdiv_rn(flo, fc)) >> 15; mhi = f2u(__fdiv_rn(fhi, fc)) >> 15; }
        decided = (mlo == mhi) && (f2u(fhi) < 0x7F800000u);
#ifdef DCB_DEBUG_MED
        if (a.flags & (1u << 17)) decided = true;
#endif
        if (decided) {
          mask17 = mlo;
          type = type_from_max(submin ? __fsub_rn(lmx, mn) : lmx);
          if (tid == 0) { a.cmed[c] = u2f(mask17 << 15); a.ctype[c] = type; }
        }
      }
      if (!decided) {
        // LEVEL 1: transducer bracket, one barrier; LEVEL 2: exact walk.
        // The running sum of the reference before / after this thread's 16 values is bracketed from the raw prefix sums of
        // barrier 1:  S_t in (PX_t - 16 t min -+ A)(1 -+ r),  A = 24 u 16 (t + 2) max|x|  (tree, scan, product, difference),
        // r = (16 t + 34) 1.02 u  (y_i = (x_i - min)(1 +- u), recursive summation of 16 (t + 1) terms); the sum is monotone (y >= 0).
        float d0 = -1.0f, d1 = -1.0f;                      // < 0: not a transducer (the walk chains this thread's values)
        float lo_keep = 0.0f;
        if (!slow) {
          const float t16 = (float)(16 * tid);
          const float PX = __fadd_rn(wb, __fsub_rn(inc, px));
          const float base0 = __fsub_rn(PX, __fmul_rn(t16, mn));
          const float base1 = __fsub_rn(__fadd_rn(PX, px), __fmul_rn(t16 + 16.0f, mn));
          const float A = __fadd_ru(__fmul_ru((t16 + 32.0f) * (24.0f * u24), amb), 1e-37f);
          const float r = (t16 + 34.0f) * (1.02f * u24);
          const float lo = __fmul_rd(fmaxf(__fsub_rd(base0, A), 0.0f), __fsub_rd(1.0f, r));
          float hi = __fmul_ru(fmaxf(__fadd_ru(base1, A), 0.0f), __fadd_ru(1.0f, r));
          hi = __fmul_ru(hi, 1.00001f);                                     // the 16 round-ups inside the thread and the +q start
          const uint32_t elo = f2u(lo) >> 23, ehi = f2u(hi) >> 23;
          lo_keep = lo;
          // plain (tree) sum of this thread's y: within 4 ulp(hi) of the real sum, the reference's 16 adds within 8 ulp(hi)
          const float q0 = __fadd_rn(__fadd_rn(x[0], x[1]), __fadd_rn(x[2], x[3]));
          const float q1 = __fadd_rn(__fadd_rn(x[4], x[5]), __fadd_rn(x[6], x[7]));
          const float q2 = __fadd_rn(__fadd_rn(x[8], x[9]), __fadd_rn(x[10], x[11]));
          const float q3 = __fadd_rn(__fadd_rn(x[12], x[13]), __fadd_rn(x[14], x[15]));
          float ctr = __fadd_rn(__fadd_rn(q0, q1), __fadd_rn(q2, q3));
          float err = __fmul_ru(hi, 24.0f / 8388608.0f);                    // 24 ulp(hi) <= 24 * 2^-23 * hi
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
            uint32_t mlo, mhi;
            if (full) { mlo = f2u(__fmul_rn(blo, 1.0f / 4096.0f)) >> 15; mhi = f2u(__fmul_rn(bhi, 1.0f / 4096.0f)) >> 15; }
            else { mlo = f2u(__fdiv_rn(blo, fc)) >> 15; mhi = f2u(__fdiv_rn(bhi, fc)) >> 15; }
            decided = (mlo == mhi) && (f2u(bhi) < 0x7F800000u);
#ifdef DCB_DEBUG_MED
            if (a.flags & (1u << 16)) decided = true;
#endif
            if (decided) {
              mask17 = mlo;
              type = type_from_max(submin ? __fsub_rn(lmx, mn) : lmx);
              if (tid == 0) { a.cmed[c] = u2f(mask17 << 15); a.ctype[c] = type; }
            }
          }
        }
        if (!decided) {                                    // block-uniform
          if (d0 < 0.0f) {
            #pragma unroll
            for (int e = 0; e < DCB_E; e++) slot[e] = x[e];
          }
          if (slow) {
            float ymx = __uint_as_float(0xff800000u);
            #pragma unroll
            for (int e = 0; e < DCB_E; e++) if (e < vcnt) ymx = fmaxf(ymx, x[e]);
            #pragma unroll
            for (int o = 16; o; o >>= 1) ymx = fmaxf(ymx, __shfl_xor_sync(0xffffffffu, ymx, o));
            __syncthreads();
            if (lane == 0) sm.red_mx[par][warp] = ymx;
          }
#ifdef DCB_DEBUG_MED
          const long long dbg_l2_t0 = clock64();
#endif
          // Runs of consecutive transducer threads in the same binade compose (associatively) into one transducer:
          // segmented warp scan.  (E then O)_p = E_p + O_{p ^ parity(E_p / q)}; every sum is exact (< 2^e, on the grid).
          const bool certain = d0 >= 0.0f;
          const unsigned cm = __ballot_sync(0xffffffffu, certain);
          const uint32_t ebits = certain ? (f2u(lo_keep) >> 23) : 0u;
          const uint32_t eprev = __shfl_up_sync(0xffffffffu, ebits, 1);
          const bool head = certain && (lane == 0 || !((cm >> (lane - 1)) & 1u) || eprev != ebits);
          const unsigned hm = __ballot_sync(0xffffffffu, head);
          const int rs = 31 - __clz((int)(hm & (0xFFFFFFFFu >> (31 - lane))));
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
          const bool stop = !certain || tail;             // where the sequential walk has to stop
          const unsigned wp = __ballot_sync(0xffffffffu, stop);
          if (lane == 0) sm.walkmask[warp] = wp;
          __syncthreads();
          // The stops are compacted into one list (run transducers, or "chain thread t" markers) so that the one
          // sequential pass reads consecutive entries: the walker's dependent chain is parity -> select -> add and nothing
          // else (a walk that looked up masks, thread ids and slots on the way took 15 us per tile, profiles/r1_summary.md).
          float2* const stopv = reinterpret_cast<float2*>(sm.W[buf]);      // free until this tile is packed
          uint32_t before = 0u, ns = 0u;
          #pragma unroll
          for (int w = 0; w < DCB_NT / 32; w++) {
            const uint32_t pc = (uint32_t)__popc(sm.walkmask[w]);
            if (w < warp) before += pc;
            ns += pc;
          }
          if (stop) {
            const uint32_t rank = before + (uint32_t)__popc(wp & ((1u << lane) - 1u));
            stopv[rank] = certain ? make_float2(d0, d1) : make_float2(-1.0f, __int_as_float(tid));
          }
          __syncthreads();
          if (tid == 0) {
            float mx = submin ? __fsub_rn(lmx, mn) : lmx;
            if (slow) {
              mx = sm.red_mx[par][0];
              #pragma unroll
              for (int w = 1; w < DCB_NT / 32; w++) mx = fmaxf(mx, sm.red_mx[par][w]);
            }
            float total = 0.0f;
            const float* Y = sm.X[buf];
#ifdef DCB_DEBUG_MED
            const long long dbg_w_t0 = clock64();
#endif
            float2 d = stopv[0];
            for (uint32_t k = 0; k < ns; k++) {
              const float2 dn = stopv[k + 1u];             // (one entry past the end is still inside the buffer)
              if (d.x >= 0.0f) {
                total = __fadd_rn(total, (f2u(total) & 1u) ? d.y : d.x);      // exact
              } else {
                const float* yp = Y + __float_as_int(d.y) * DCB_XPITCH;
                const float4 v0 = *reinterpret_cast<const float4*>(yp), v1 = *reinterpret_cast<const float4*>(yp + 4);
                const float4 v2 = *reinterpret_cast<const float4*>(yp + 8), v3 = *reinterpret_cast<const float4*>(yp + 12);
                total = __fadd_rn(total, v0.x); total = __fadd_rn(total, v0.y); total = __fadd_rn(total, v0.z); total = __fadd_rn(total, v0.w);
                total = __fadd_rn(total, v1.x); total = __fadd_rn(total, v1.y); total = __fadd_rn(total, v1.z); total = __fadd_rn(total, v1.w);
                total = __fadd_rn(total, v2.x); total = __fadd_rn(total, v2.y); total = __fadd_rn(total, v2.z); total = __fadd_rn(total, v2.w);
                total = __fadd_rn(total, v3.x); total = __fadd_rn(total, v3.y); total = __fadd_rn(total, v3.z); total = __fadd_rn(total, v3.w);
              }
              d = dn;
            }
#ifdef DCB_DEBUG_MED
            { const long long t1 = clock64(); atomicAdd(&g_dbg_walk[0], (unsigned long long)(t1 - dbg_l2_t0)); atomicAdd(&g_dbg_walk[1], (unsigned long long)(t1 - dbg_w_t0)); atomicAdd(&g_dbg_walk[2], 1ull); atomicAdd(&g_dbg_walk[3], (unsigned long long)ns); }
#endif
            const float med = __fdiv_rn(total, (float)cnt);
            const int ty = (Y[0] != Y[0]) ? 0 : type_from_max(mx);
            sm.med = med; sm.type = ty;
            a.cmed[c] = med; a.ctype[c] = ty;
          }
          __syncthreads();
          type = sm.type;
          mask17 = f2u(sm.med) >> 15;
        }
      }
    }

    // ---- 3. classify --------------------------------------------------------------------------------------
    uint32_t code[DCB_E];
    int len[DCB_E];
    uint32_t chi = 0u, clo = 0u;
    uint32_t total = 0;
    bool all3 = false;
    int np = DCB_E;                  // pieces the packing loop has to place (warp-uniform)
    if (!slow) {
      uint32_t miss = 0u;                              // bit e: element e needs a fallback code
      if (CT != CT_NP) {
        float b1 = h1, b2 = h2, b3 = h3;
        float t1 = __fmul_rn(3.0f, b1), t2 = __fmul_rn(3.0f, b2);
        uint32_t fa = 0x92492490u, fb = 0x92490000u;   // 3-bit fields of elements 0..9 / 10..15, all preset to 100 (zero)
        const float tp3 = first ? -1.0f : a.k.tp;      // chunk start: elements 0..2 never hit
        #define DCB_CLS(E, ACC, POS, TP) do {                                                                   \
            const float xe = x[E];                                                                               \
            const float p2 = __fsub_rn(__fadd_rn(b1, b1), b2);             /* 2*b1 is exact */                   \
            const float p3 = __fadd_rn(__fsub_rn(t1, t2), b3);                                                   \
            classify_pred<(1u << (POS))>(ACC, miss, xe, __fsub_rn(b1, xe), __fsub_rn(p2, xe), __fsub_rn(p3, xe), \
                                         a.k.tz, TP, one, 1u << (E));                                            \
            b3 = b2; b2 = b1; b1 = xe; t2 = t1; t1 = __fmul_rn(3.0f, xe);                                        \
          } while (0)
        DCB_CLS(0, fa, 29, tp3); DCB_CLS(1, fa, 26, tp3); DCB_CLS(2, fa, 23, tp3); DCB_CLS(3, fa, 20, a.k.tp);
        DCB_CLS(4, fa, 17, a.k.tp); DCB_CLS(5, fa, 14, a.k.tp); DCB_CLS(6, fa, 11, a.k.tp); DCB_CLS(7, fa, 8, a.k.tp);
        DCB_CLS(8, fa, 5, a.k.tp); DCB_CLS(9, fa, 2, a.k.tp);
        DCB_CLS(10, fb, 29, a.k.tp); DCB_CLS(11, fb, 26, a.k.tp); DCB_CLS(12, fb, 23, a.k.tp); DCB_CLS(13, fb, 20, a.k.tp);
        DCB_CLS(14, fb, 17, a.k.tp); DCB_CLS(15, fb, 14, a.k.tp);
        #undef DCB_CLS
        chi = fa | (fb >> 30); clo = fb << 2;          // the 48-bit string: 30 bits of fa, 18 bits of fb
        all3 = (vcnt == DCB_E) && (miss == 0u);
      }
      // The first three elements of a chunk never hit (:3330), so the chunk's first lane always needs fallback codes: when that
      // is the only reason the warp is not pure 3-bit codes, only that lane leaves the fast path, with five pieces to pack
      // (three codes, then its remaining thirteen 3-bit codes as 32 + 7 bits) instead of sixteen.
      const bool warp_all3 = __all_sync(0xffffffffu, all3);
      const bool head_ok = all3 || (CT != CT_NP && first && vcnt == DCB_E && (miss & ~7u) == 0u);
      const bool headcase = !warp_all3 && __all_sync(0xffffffffu, head_ok);
      if (warp_all3) {
        total = 48u;
      } else if (headcase) {
        np = 5;
        if (all3) {
          total = 48u;
        } else {
          #pragma unroll
          for (int e = 0; e < 3; e++) {
            const uint32_t f = ((miss >> e) & 1u) ? 0u : C48_FIELD(chi, clo, e);
            uint32_t ft; int fl;
            fallback_top<CT, true>(f2u(x[e]), a.k, type, mask17, ft, fl);
            code[e] = f ? (f << 29) : ft;
            len[e] = f ? 3 : fl;
          }
          code[3] = (chi << 9) | (clo >> 23); len[3] = 32;
          code[4] = clo << 9; len[4] = 7;
          total = (uint32_t)(len[0] + len[1] + len[2]) + 39u;
        }
      } else {
        all3 = false;
        #pragma unroll
        for (int e = 0; e < DCB_E; e++) {
          uint32_t ft; int fl;
          fallback_top<CT, true>(f2u(x[e]), a.k, type, mask17, ft, fl);
          // the element's 3-bit field, top aligned, straight from the 48-bit string
          const uint32_t f3 = (e < 10 ? (chi << (3 * e)) : (e == 10 ? ((chi << 30) | (clo >> 2)) : (clo << (3 * e - 32)))) & 0xE0000000u;
          const bool ms = (CT == CT_NP) || ((miss >> e) & 1u);
          code[e] = ms ? ft : f3;
          len[e] = ms ? fl : 3;
          total += (uint32_t)len[e];
        }
        if (!full) {                                        // block-uniform: a ragged tile, padding carries no code
          #pragma unroll
          for (int e = 0; e < DCB_E; e++)
            if (e >= vcnt) { total -= (uint32_t)len[e]; code[e] = 0u; len[e] = 0; }
        }
      }
    } else {
      // -1.0f is the reference's "empty history slot" marker: rebuild the exact state this thread
      // starts from by walking back over the chunk until three non-sentinel values are found.
      float b3 = -1.0f, b2 = -1.0f, b1 = -1.0f;
      if (vcnt > 0 && CT != CT_NP) {
        float n0 = 0.0f, n1 = 0.0f, n2 = 0.0f;
        int found = 0; bool last_sent = false;
        for (uint64_t idx = my0; idx > base && found < 3; idx--) {
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
    if (tid == DCB_TKT_TID) sm.tkt[par] = tk;
    const bool fast_pack = __syncthreads_and(total >= 32u || total == 0u);      // barrier 2
    const uint32_t nn = sm.tkt[par];
    uint32_t warp_base = 0, tile_bits = 0;
    #pragma unroll
    for (int w = 0; w < DCB_NT / 32; w++) {
      const uint32_t t = sm.warp_tot[par][w];
      if (w < warp) warp_base += t;
      tile_bits += t;
    }
    const uint32_t s = warp_base + incl - total;        // first bit of this thread inside the tile
    if (tid == 0) {
      const uint32_t bytes = (tile_bits + 7u) >> 3;
      a.sizes[tile] = bytes;
      atomicAdd(&a.gsum[tile / DCB_GROUP], bytes);
      if (tile == a.num_tiles - 1u) *a.last_tile_bits = tile_bits;
    }

    // ---- 5. the tile packed in the previous iteration leaves for its slot ------------------------------------
    if (it > 0u) copy_out_slot(a.slots + (uint64_t)pend_tile * DCB_SLOT_PITCH, sm.W[buf ^ 1], pend_bits, tid);

    // ---- 6. pack the current tile into the staging buffer -------------------------------------------------
    uint32_t hi = 0u;
    const uint32_t f0 = s & 31u;
    const uint32_t hw = s >> 5;
    if (fast_pack) {
      uint32_t widx = hw;
      if (all3) {
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
          if (e >= np) break;
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
      const uint32_t prevT = __shfl_up_sync(0xffffffffu, hi, 1);
      if (lane > 0 && total > 0u && f0) W[hw] |= prevT;
      if (lane == 31) sm.xchg[warp] = hi;
      if (total > 0u && s + total == tile_bits) {          // last thread that has bits: flush its tail, zero the rest of the 16-byte unit
        if ((s + total) & 31u) W[widx++] = hi;
        W[widx] = 0u; W[widx + 1] = 0u; W[widx + 2] = 0u;
      }
      fix_need = (lane == 0 && warp > 0 && total > 0u && f0 != 0u);
      fix_hw = hw;
    } else {
      // ragged tile (fewer than 11 valid elements in some thread): zero the buffer and OR every word in.
      // The buffer was last read by the copy-out of iteration it-1, which every thread left before barrier 1.
      for (int i = tid; i < DCB_TILE + 8; i += DCB_NT) W[i] = 0u;
      __syncthreads();
      uint32_t fill = f0, widx = hw;
      if (all3) {
        #pragma unroll
        for (int e = 0; e < DCB_E; e++) { code[e] = C48_FIELD(chi, clo, e) << 29; len[e] = 3; }
      }
      #pragma unroll
      for (int e = 0; e < DCB_E; e++) {
        if (e >= np) break;
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

    // ---- 7. next tile ------------------------------------------------------------------------------------------
    pend_tile = tile; pend_bits = tile_bits;
    par ^= 1;
    cur = nxt; nxt = nn;
    if (cur >= a.num_tiles) break;
  }

  // ---- epilogue: the last tile of this block -------------------------------------------------------------------
  {
    const int lastb = (int)(it & 1u);
    __syncthreads();                                     // xchg and every word of W[lastb] are written
    if (fix_need) sm.W[lastb][fix_hw] |= sm.xchg[warp - 1];
    __syncthreads();
    copy_out_slot(a.slots + (uint64_t)pend_tile * DCB_SLOT_PITCH, sm.W[lastb], pend_bits, tid);
  }
}

// ------------------------------------------------------------------------------------------------
// pass 2: chunk table + compaction.  Block g owns chunks [64 g, 64 g + 64).
struct CompactArgs {
  const uint8_t* slots;
  const uint32_t* sizes;
  const uint32_t* gsum;
  const uint32_t* last_tile_bits;
  uint64_t K;
  uint8_t* out;
  uint64_t* offsets;                 // [K+1]
  unsigned long long* last_bits;     // bit length of the whole stream before the final byte round-up (legacy `pos`)
  uint32_t* crc;                     // [K] or NULL: do_crc32 of every chunk's stream, computed on the bytes as they are copied
  uint32_t pitch;                    // bytes per slot: DCB_SLOT_PITCH (float32), 32768 (the float64 codecs' fused encoder)
};

struct CrcTabs {                     // shared-memory copies of the tables of dcb_crc.cuh
  uint32_t tab[256];
  uint32_t t4[4][256];
  uint32_t c2[257];
};

__device__ __forceinline__ uint4 ldg_u4(const uint8_t* p) { return __ldg(reinterpret_cast<const uint4*>(p)); }

// one warp moves one chunk: n bytes from the 16-byte aligned slot S to the byte address D.
// CRC = true: do_crc32 (dataCompression.c:5524) of the chunk rides on the copy's own loads (r1 ran dcb_crc_kernel over the
// finished stream: a third pass over B bytes, 0.07-0.24 ms).  The slot is 16-byte aligned and starts the chunk, so source unit
// j = chunk bytes [16 j, 16 j + 16); lane L loads units L, L + 32, ..: four Horner accumulators of stride 128 words,
// U_k <- U_k * x^4096 + w_k (four table lookups per word); at the end a lane folds its four (three times x^32 through the
// byte table), weighs the result with x^(32 (R - i_last)) (one bit-serial multiply) and the warp XORs; the <= 31 bytes behind
// the last unit it loaded go through the byte table.  Returns the CRC (warp uniform).
template <bool CRC>
__device__ __forceinline__ uint32_t warp_copy_chunk(const uint8_t* __restrict__ S, uint8_t* __restrict__ D, uint32_t n, int lane,
                                                    const CrcTabs* __restrict__ T) {
  const uint32_t a = (uint32_t)(reinterpret_cast<uintptr_t>(D) & 15u);
  const uint32_t head = min(n, (16u - a) & 15u);
  if ((uint32_t)lane < head) D[lane] = S[lane];
  const uint32_t nb = (n - head) >> 4;                   // whole 16-byte units of the destination
  const uint32_t q = head >> 2, b8 = (head & 3u) * 8u;   // warp-uniform source misalignment
  uint4* Dv = reinterpret_cast<uint4*>(D + head);
  uint32_t U0 = 0u, U1 = 0u, U2 = 0u, U3 = 0u;
  for (uint32_t j = (uint32_t)lane; j < nb; j += 32u) {
    const uint4 A = ldg_u4(S + 16u * j);
    if (CRC) {
      const uint32_t w0 = (j == 0u) ? (A.x ^ 0xFFFFFFFFu) : A.x;      // the initial state enters the first word (CRC is linear)
      U0 = T->t4[0][U0 & 0xFFu] ^ T->t4[1][(U0 >> 8) & 0xFFu] ^ T->t4[2][(U0 >> 16) & 0xFFu] ^ T->t4[3][U0 >> 24] ^ w0;
      U1 = T->t4[0][U1 & 0xFFu] ^ T->t4[1][(U1 >> 8) & 0xFFu] ^ T->t4[2][(U1 >> 16) & 0xFFu] ^ T->t4[3][U1 >> 24] ^ A.y;
      U2 = T->t4[0][U2 & 0xFFu] ^ T->t4[1][(U2 >> 8) & 0xFFu] ^ T->t4[2][(U2 >> 16) & 0xFFu] ^ T->t4[3][U2 >> 24] ^ A.z;
      U3 = T->t4[0][U3 & 0xFFu] ^ T->t4[1][(U3 >> 8) & 0xFFu] ^ T->t4[2][(U3 >> 16) & 0xFFu] ^ T->t4[3][U3 >> 24] ^ A.w;
    }
    uint4 r;
    if (head == 0u) {
      r = A;
    } else {
      const uint4 B = ldg_u4(S + 16u * j + 16u);         // inside the slot: head + 16 j + 16 <= n <= pitch
      uint32_t w0, w1, w2, w3, w4;
      switch (q) {
        case 0: w0 = A.x; w1 = A.y; w2 = A.z; w3 = A.w; w4 = B.x; break;
        case 1: w0 = A.y; w1 = A.z; w2 = A.w; w3 = B.x; w4 = B.y; break;
        case 2: w0 = A.z; w1 = A.w; w2 = B.x; w3 = B.y; w4 = B.z; break;
        default: w0 = A.w; w1 = B.x; w2 = B.y; w3 = B.z; w4 = B.w; break;
      }
      r.x = __funnelshift_r(w0, w1, b8); r.y = __funnelshift_r(w1, w2, b8);
      r.z = __funnelshift_r(w2, w3, b8); r.w = __funnelshift_r(w3, w4, b8);
    }
    Dv[j] = r;
  }
  const uint32_t done = head + (nb << 4);
  const uint32_t t = n - done;
  if ((uint32_t)lane < t) D[done + lane] = S[done + lane];
  if (!CRC) return 0u;
  uint32_t st = 0xFFFFFFFFu;
  if (nb) {
    uint32_t v = 0u;
    if ((uint32_t)lane < nb) {
      // V = ((U0 x^32 + U1) x^32 + U2) x^32 + U3; a multiplication by x^32 is four zero bytes through the byte table
      uint32_t V = U0;
      #pragma unroll
      for (int k = 0; k < 3; k++) {
        #pragma unroll
        for (int z = 0; z < 4; z++) V = T->tab[V & 0xFFu] ^ (V >> 8);
        V ^= (k == 0) ? U1 : ((k == 1) ? U2 : U3);
      }
      const uint32_t jlast = (uint32_t)lane + 32u * ((nb - 1u - (uint32_t)lane) >> 5);   // the last unit this lane loaded
      v = gf2_mulmod(V, T->c2[4u * (nb - jlast) - 3u]);          // x^(32 (R - i_last)), R = 4 nb words, i_last = 4 jlast + 3
    }
    #pragma unroll
    for (int o = 16; o; o >>= 1) v ^= __shfl_xor_sync(0xffffffffu, v, o);
    st = v;
  }
  for (uint32_t i = nb << 4; i < n; i++) st = T->tab[(st ^ S[i]) & 0xFFu] ^ (st >> 8);   // < 32 bytes, warp uniform
  return st ^ 0xFFFFFFFFu;
}

#define DCB_COMPACT_NT 256
__global__ void __launch_bounds__(DCB_COMPACT_NT) dcb_compact_kernel(const CompactArgs a) {
  __shared__ unsigned long long s_part[DCB_COMPACT_NT / 32];
  __shared__ unsigned long long s_off[DCB_GROUP + 1];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t g = blockIdx.x;
  const uint64_t c0 = (uint64_t)g * DCB_GROUP;
  const uint32_t cnt = (uint32_t)min((uint64_t)DCB_GROUP, a.K - c0);
  // bytes of all the groups before this one
  unsigned long long part = 0ull;
  for (uint32_t i = (uint32_t)tid; i < g; i += DCB_COMPACT_NT) part += __ldg(a.gsum + i);
  #pragma unroll
  for (int o = 16; o; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
  if (lane == 0) s_part[warp] = part;
  // local exclusive scan of the group's sizes (warp 0, two chunks per lane)
  uint32_t v0 = 0u, v1 = 0u;
  if (warp == 0) {
    if (2u * lane < cnt) v0 = __ldg(a.sizes + c0 + 2u * lane);
    if (2u * lane + 1u < cnt) v1 = __ldg(a.sizes + c0 + 2u * lane + 1u);
  }
  __syncthreads();
  if (warp == 0) {
    unsigned long long basev = 0ull;
    #pragma unroll
    for (int w = 0; w < DCB_COMPACT_NT / 32; w++) basev += s_part[w];
    uint32_t incl = v0 + v1;
    #pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += t;
    }
    const uint32_t ex = incl - v0 - v1;
    s_off[2 * lane] = basev + ex;
    s_off[2 * lane + 1] = basev + ex + v0;
    if (lane == 31) s_off[DCB_GROUP] = basev + incl;
  }
  __syncthreads();
  if ((uint32_t)tid < cnt) a.offsets[c0 + tid + 1u] = s_off[tid + 1];
  if (g == 0 && tid == 0) a.offsets[0] = 0ull;
  if (c0 + cnt == a.K && tid == 0 && a.last_bits) *a.last_bits = s_off[cnt - 1u] * 8ull + (unsigned long long)__ldg(a.last_tile_bits);
  if (a.crc) {
    __shared__ CrcTabs T;
    for (int i = tid; i < 256; i += DCB_COMPACT_NT) {
      T.tab[i] = g_crc_tab[i];
      #pragma unroll
      for (int k = 0; k < 4; k++) T.t4[k][i] = g_crc_t4[k][i];
    }
    for (int i = tid; i < 257; i += DCB_COMPACT_NT) T.c2[i] = g_crc_c2[i];
    __syncthreads();
    for (uint32_t k = (uint32_t)warp; k < cnt; k += DCB_COMPACT_NT / 32) {
      const uint32_t nbytes = (uint32_t)(s_off[k + 1u] - s_off[k]);
      const uint32_t crc = warp_copy_chunk<true>(a.slots + (c0 + k) * (uint64_t)a.pitch, a.out + s_off[k], nbytes, lane, &T);
      if (lane == 0) a.crc[c0 + k] = crc;
    }
    return;
  }
  for (uint32_t k = (uint32_t)warp; k < cnt; k += DCB_COMPACT_NT / 32) {
    const uint32_t nbytes = (uint32_t)(s_off[k + 1u] - s_off[k]);
    warp_copy_chunk<false>(a.slots + (c0 + k) * (uint64_t)a.pitch, a.out + s_off[k], nbytes, lane, nullptr);
  }
}
