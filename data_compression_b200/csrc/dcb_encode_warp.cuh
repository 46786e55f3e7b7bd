// dcb_encode_warp.cuh -- WARP-AUTONOMOUS encoder for "one chunk == one 4096-element tile" (CT=5/6/7/11), pass 1 of the slot
// path (dcb_encode_slots.cuh; pass 2 = dcb_compact_kernel).  Replaces, per chunk, toSmallDataset_float (dataCompression.c:3543)
// -> [med_dataset_float (:3593)] -> myCompress_bitwise{,_np,_mask,_op} (:3310, :2645, :2030, :577).
//
// One warp owns one chunk from the first load to the last store; nothing is shared between warps, so there is no
// __syncthreads, no shared-memory reduction and no cross-warp fix-up at all (the block-per-chunk kernel spent 18-27 % of its
// warp time at its two barriers and a third of its instructions on per-tile bookkeeping, profiles/r1_summary.md).
//   * the chunk is consumed in 8 ROUNDS of 512 elements, 16 consecutive elements per lane (a lane concatenates its own codes
//     in registers; the 3-element look-back comes from the neighbour lane by shuffle, from the previous round by three
//     carried registers);
//   * statistics (min, [max, raw sum], NaN/Inf/huge test) of the NEXT chunk are folded in one round at a time from a 2 KB
//     cp.async staging buffer -- that is also what pulls the next chunk from HBM into L2, so the transposing cp.async of
//     the classification pass hits L2; every load has a whole round (~2 us) to land;
//   * a round's codes are packed into a 2 KB warp buffer at the bit position where the previous round stopped, whole words
//     leave for the chunk's scratch slot at once (MSB-first bytes, coalesced), the open word is carried in a register.
#pragma once
#include "dcb_encode_slots.cuh"

#ifndef DCB_WARP_BLOCKS_CT7
#define DCB_WARP_BLOCKS_CT7 2        // CT=7 carries more live state: 2 blocks per SM (up to 128 registers) instead of spilling at 80
#endif
#ifndef DCB_WNT
#define DCB_WNT 256                 // threads per block of the warp kernel (its warps are independent)
#endif
#ifndef DCB_WARP_BLOCKS
#define DCB_WARP_BLOCKS 2          // 16 warps per SM with up to 128 registers beat 24 warps at 80 (0.47 -> 0.43 ms, CT=5 smooth)
#endif
#define DCB_WR 512                  // elements per round (32 lanes x 16)
#define DCB_WWORDS 520              // words of the round buffer: 512 codes of <= 32 bits + carry word + slack

#ifdef DCB_DEBUG_MED
__device__ unsigned long long g_dbg_exact[4];   // cycles in exact_medium, calls, cycles per whole chunk, chunks
#endif
#ifndef DCB_PACKED
#define DCB_PACKED 1                // classification arithmetic on sm_100 packed fp32 pairs (FADD2): two elements per instruction
#endif
// Packed fp32 (PTX add/sub/mul.rn.f32x2, sm_100+; SASS FADD2 / FMUL2): every half is an IEEE round-to-nearest operation of
// its own, so the reference's per-operation float32 rounding (dataCompression.c:3362-3364) is kept bit for bit
// (tools/probe_f32x2.cu checks that on the device against __fadd_rn / __fmul_rn: 2^28 quadruples, 0 mismatches).
__device__ __forceinline__ uint64_t pk2(float lo, float hi) { uint64_t r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ void upk2(uint64_t r, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(r)); }
__device__ __forceinline__ uint64_t add2(uint64_t a, uint64_t b) { uint64_t r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ uint64_t sub2(uint64_t a, uint64_t b) { uint64_t r; asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }

struct WarpBuf {
  float X[2][32 * DCB_XPITCH];      // classification pass: double-buffered rounds, one 80-byte slot per lane
  float4 P[128];                    // statistics pass of the next chunk: one round, linear
  uint32_t W[DCB_WWORDS];           // packed round
  unsigned long long mbar[2];       // [0]: mbarrier of the bulk copy into P (16 bytes keep the array of WarpBufs 16-byte aligned)
};

// ---- TMA bulk copy of one statistics round (2 KB, linear) ----------------------------------------------------------------------
// The statistics pass reads the NEXT chunk round by round in plain order, so a round is one contiguous, 16-byte aligned 2 KB
// block: ONE cp.async.bulk (SASS UBLKCP) issued by lane 0 with an mbarrier for completion can replace four LDGSTS per lane
// (128 per warp and round).  The classification pass keeps cp.async: its 16 consecutive elements per lane sit at an 80-byte
// pitch (conflict-free 128-bit reads), which a linear bulk copy cannot produce.
// MEASURED (round 2, 2^28 smooth, one B200, two A/B runs each): SLOWER than the four LDGSTS -- encoder kernel CT=7 0.597 ms
// against 0.566, CT=5 compress 0.432 against 0.414 (0.588 / 0.426 without the proxy fence): a 2 KB copy per warp and round is
// too small for the bulk engine's per-copy cost plus an mbarrier wait per round.  So the default build keeps cp.async
// (DCB_BULK_P = 0, no UBLKCP in the shipped binary); -DDCB_BULK_P=1 builds this path (parity tests pass on it).
#ifndef DCB_BULK_P
#define DCB_BULK_P 0
#endif
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long* bar) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;\n" ::"r"(smem_u32(bar)) : "memory");
  asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
  asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
}
__device__ __forceinline__ void bulk_load_2k(void* smem_dst, const void* gsrc, unsigned long long* bar) {
  // the buffer was read by ordinary loads in the previous round: order them before the asynchronous proxy's writes
  asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], 2048;\n" ::"r"(smem_u32(bar)) : "memory");
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], 2048, [%2];\n"
               ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, uint32_t phase) {
  uint32_t done;
  do {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(done) : "r"(smem_u32(bar)), "r"(phase) : "memory");
  } while (!done);
}

struct Pass1 {                      // per-lane partial statistics of a chunk
  float mn, mx, s0, s1, ws, n0, n1;
  uint32_t sg;
};
template <int CT>
__device__ __forceinline__ void p1_reset(Pass1& p) {
  p.mn = __uint_as_float(0x7f800000u); p.mx = __uint_as_float(0xff800000u);
  p.s0 = 0.0f; p.s1 = 0.0f; p.ws = 0.0f; p.n0 = 0.0f; p.n1 = 0.0f; p.sg = 0u;
}
// four consecutive raw values; w = (elements of the chunk from the first of them on), the weight of the CT=7 bound.
// nacc turns Inf/NaN iff one of them is NaN, Inf or >= 2^127 in magnitude (or, harmlessly, when the running sum of huge
// values overflows: such chunks take the exact SLOW path).
template <int CT, bool SG>
__device__ __forceinline__ void p1_add4(Pass1& p, const float4 v, int which, float w) {
  p.mn = fminf(fminf(p.mn, v.x), v.y); p.mn = fminf(fminf(p.mn, v.z), v.w);
  if (CT == CT_MASK) {
    p.mx = fmaxf(fmaxf(p.mx, v.x), v.y); p.mx = fmaxf(fmaxf(p.mx, v.z), v.w);
    const float t = __fadd_rn(__fadd_rn(v.x, v.y), __fadd_rn(v.z, v.w));
    if (which & 1) p.s1 = __fadd_rn(p.s1, t); else p.s0 = __fadd_rn(p.s0, t);
    p.ws = __fmaf_rn(t, w, p.ws);
  }
  p.n0 = __fmaf_rn(fabsf(v.x), 2.0f, p.n0); p.n1 = __fmaf_rn(fabsf(v.y), 2.0f, p.n1);
  p.n0 = __fmaf_rn(fabsf(v.z), 2.0f, p.n0); p.n1 = __fmaf_rn(fabsf(v.w), 2.0f, p.n1);
  if (SG) p.sg |= f2u(v.x) | f2u(v.y) | f2u(v.z) | f2u(v.w);
}
template <int CT>
__device__ __forceinline__ void p1_add1(Pass1& p, float v, bool submin) {      // ragged rounds, unaligned input
  p.mn = fminf(p.mn, v);
  if (CT == CT_MASK) { p.mx = fmaxf(p.mx, v); p.s0 = __fadd_rn(p.s0, v); }
  p.n0 = __fmaf_rn(fabsf(v), 2.0f, p.n0);
  if (!submin) p.sg |= f2u(v);
}

__device__ __forceinline__ bool round_is_bulk(const EncArgs& a, uint64_t e0, uint64_t cend) {
  return e0 + DCB_WR <= cend && (a.flags & DCB_ENC_IN_ALIGNED);
}
// classification pass: round starting at element e0 -> one 80-byte slot per lane (16 consecutive elements each)
__device__ __forceinline__ void warp_issue_x(const EncArgs& a, uint64_t e0, uint64_t cend, float* Xb, int lane, bool bulk) {
  if (bulk) {
    const float* g = a.in + e0;
    #pragma unroll
    for (int q = 0; q < 4; q++) {
      const int u = q * 32 + lane;
      cp_async16(Xb + (u >> 2) * DCB_XPITCH + (u & 3) * 4, g + u * 4);
    }
  } else {
    const uint64_t my0 = e0 + (uint64_t)lane * DCB_E;
    float* slot = Xb + lane * DCB_XPITCH;
    #pragma unroll
    for (int e = 0; e < DCB_E; e++) slot[e] = (my0 + e < cend) ? __ldg(a.in + my0 + e) : 0.0f;
  }
}
// statistics pass: the same round of the NEXT chunk, linear
__device__ __forceinline__ void warp_issue_p(const EncArgs& a, uint64_t e0, float4* P, int lane, bool bulk,
                                             unsigned long long* bar) {
  if (bulk) {
    const float* g = a.in + e0;
#if DCB_BULK_P
    if (lane == 0) bulk_load_2k(P, g, bar);
#else
    #pragma unroll
    for (int q = 0; q < 4; q++) cp_async16(&P[q * 32 + lane], g + (q * 32 + lane) * 4);
#endif
  }
}
// j = round inside its chunk (weights: 4096 - index of the float4's first element; only used for whole bulk chunks)
template <int CT>
__device__ __forceinline__ void warp_consume_p(const EncArgs& a, uint64_t e0, uint64_t cend, const float4* P, int lane,
                                               Pass1& p, bool submin, bool bulk, int j) {
  if (bulk) {
    const float w0 = (float)(DCB_TILE - j * DCB_WR - lane * 4);
    if (submin) {
      #pragma unroll
      for (int q = 0; q < 4; q++) p1_add4<CT, false>(p, P[q * 32 + lane], q, w0 - (float)(q * 128));
    } else {
      #pragma unroll
      for (int q = 0; q < 4; q++) p1_add4<CT, true>(p, P[q * 32 + lane], q, w0 - (float)(q * 128));
    }
  } else {
    for (int k = 0; k < DCB_E; k++) {
      const uint64_t idx = e0 + (uint64_t)k * 32 + lane;
      if (idx < cend) p1_add1<CT>(p, __ldg(a.in + idx), submin);
    }
  }
}
// whole chunk at once (a warp's first chunk): direct loads
template <int CT>
__device__ __forceinline__ void pass1_direct(const EncArgs& a, uint64_t cbase, uint64_t cend, int lane, Pass1& p, bool submin) {
  int j = 0;
  for (uint64_t e0 = cbase; e0 < cend; e0 += 2 * DCB_WR, j += 2) {
    if (round_is_bulk(a, e0 + DCB_WR, cend)) {           // two rounds: eight 128-bit loads in flight
      const float4* g = reinterpret_cast<const float4*>(a.in + e0);
      float4 v[8];
      #pragma unroll
      for (int q = 0; q < 8; q++) v[q] = __ldg(g + q * 32 + lane);
      const float w0 = (float)(DCB_TILE - j * DCB_WR - lane * 4);
      #pragma unroll
      for (int q = 0; q < 8; q++) p1_add4<CT, true>(p, v[q], q, w0 - (float)(q * 128));
    } else {
      for (uint64_t idx = e0 + lane; idx < min(e0 + 2 * DCB_WR, cend); idx += 32) p1_add1<CT>(p, __ldg(a.in + idx), submin);
    }
  }
}

// med_dataset_float (:3593-3620), exactly, for the chunks whose mask bits the bound cannot pin (and SLOW chunks): one more
// pass over the chunk, 32 consecutive values per lane and step (1024 per step).  While the running sum S stays inside one binade
// [2^e, 2^(e+1)) it lives on the grid q = 2^(e-23), and S -> fl(S + y) commutes with shifts of S by 2q, so a lane whose 32
// values provably keep S inside one binade is a 2-state transducer S_out = S_in + d[parity(S_in / q)], d[0], d[1] from two
// 32-step chains started at 2^e and 2^e + q -- all lanes in parallel, every d exact.  The binade is pinned by S at the start
// of the step (exact, carried) plus a warp scan of the lanes' plain sums: the reference's running sum before / after lane L
// lies within (32 L + 80) 1.02 u of it (32 L + 32 adds of the reference, 11 roundings of ours).  Then S is handed from run to
// run of lanes (see "General step" below), or through the 32 values of a lane that straddles a binade (all lanes of a SLOW
// chunk).
// Measured (round 2): one call took 27.9k cycles of its warp with 16 values per lane, eight steps and the input staged through
// shared memory one step ahead -- a chain of dependent latencies (load, tree, scan, two chains, ballots, butterfly) per step,
// not an instruction count (a cheaper combine stage changed nothing).  Hence half as many, fatter steps, values straight into
// registers, and the lines of the step after next requested from L1 (prefetch) while this one is worked on.
#define DCB_XE 32                    // values per lane and step of the exact pass
__device__ __noinline__ void exact_medium(const EncArgs& a, WarpBuf& sm, uint64_t cbase, uint64_t cend, float mn, bool submin,
                                          bool slow, bool cbulk, int lane, float& total_out, float& ymx_out, float& y0_out) {
  (void)sm; (void)cbulk;
  const float u24 = 1.0f / 16777216.0f;
  constexpr uint64_t XR = 32ull * DCB_XE;                            // values per step
  const uint64_t nr = (cend - cbase + XR - 1) / XR;                  // 64 bit: dcb_stats_kernel calls this for chunks of any length
  const bool aligned = (a.flags & DCB_ENC_IN_ALIGNED) != 0;
  float S = 0.0f, ymx = __uint_as_float(0xff800000u), y0 = 0.0f;
  auto prefetch = [&](uint64_t j) {
    const uint64_t my0 = cbase + j * XR + (uint64_t)lane * DCB_XE;
    if (my0 < cend) asm volatile("prefetch.global.L1 [%0];" :: "l"(a.in + my0));
  };
  prefetch(0);
  if (nr > 1) prefetch(1);
  for (uint64_t j = 0; j < nr; j++) {
    if (j + 2 < nr) prefetch(j + 2);
    const uint64_t e0 = cbase + j * XR;
    const uint64_t my0 = e0 + (uint64_t)lane * DCB_XE;
    const int vcnt = (my0 >= cend) ? 0 : (int)min((uint64_t)DCB_XE, cend - my0);
    float y[DCB_XE];
    if (aligned && e0 + XR <= cend) {
      const float4* g = reinterpret_cast<const float4*>(a.in + my0);
      #pragma unroll
      for (int q = 0; q < DCB_XE / 4; q++) {
        const float4 v = __ldg(g + q);
        y[4 * q] = v.x; y[4 * q + 1] = v.y; y[4 * q + 2] = v.z; y[4 * q + 3] = v.w;
      }
    } else {
      #pragma unroll
      for (int e = 0; e < DCB_XE; e++) y[e] = (e < vcnt) ? __ldg(a.in + my0 + e) : 0.0f;
    }
    if (submin) {
      #pragma unroll
      for (int e = 0; e < DCB_XE; e++) y[e] = (e < vcnt) ? __fsub_rn(y[e], mn) : 0.0f;   // + 0.0f leaves every sum as it is
    }
    if (j == 0) y0 = __shfl_sync(0xffffffffu, y[0], 0);
    float d0 = -1.0f, d1 = -1.0f;                        // < 0: not a transducer
    uint32_t ebits = 0u;
    if (!slow) {
      float t8[8];
      #pragma unroll
      for (int q = 0; q < 8; q++)
        t8[q] = __fadd_rn(__fadd_rn(y[4 * q], y[4 * q + 1]), __fadd_rn(y[4 * q + 2], y[4 * q + 3]));
      const float py = __fadd_rn(__fadd_rn(__fadd_rn(t8[0], t8[1]), __fadd_rn(t8[2], t8[3])),
                                 __fadd_rn(__fadd_rn(t8[4], t8[5]), __fadd_rn(t8[6], t8[7])));
      float inc = py;
      #pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const float t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc = __fadd_rn(inc, t);
      }
      const float before = __fadd_rn(S, __fsub_rn(inc, py)), after = __fadd_rn(S, inc);
      const float r = (float)(DCB_XE * lane + 80) * (1.02f * u24);
      const float lo = __fmul_rd(before, __fsub_rd(1.0f, r));
      const float hi = __fmul_ru(__fmul_ru(after, __fadd_ru(1.0f, r)), 1.00001f);
      const uint32_t elo = f2u(lo) >> 23, ehi = f2u(hi) >> 23;
      if (elo == ehi && elo > 0u && elo < 255u) {
        const float r0 = u2f(elo << 23), r1 = u2f((elo << 23) | 1u);
        float c0 = r0, c1 = r1;
        #pragma unroll
        for (int e = 0; e < DCB_XE; e++) { c0 = __fadd_rn(c0, y[e]); c1 = __fadd_rn(c1, y[e]); }
        d0 = __fsub_rn(c0, r0); d1 = __fsub_rn(c1, r1);  // exact: both ends lie on the grid of one binade
        ebits = elo;
      }
    } else {
      #pragma unroll
      for (int e = 0; e < DCB_XE; e++) if (e < vcnt) ymx = fmaxf(ymx, y[e]);
    }
    const bool certain = d0 >= 0.0f;
    const unsigned cm = __ballot_sync(0xffffffffu, certain);
    // The common step: every lane is a transducer of the SAME binade and none depends on the parity of S (no tie in its 32
    // additions).  Then S_out = S_in + sum of the d0, and that sum is exact in any order (multiples of the binade's grid below
    // 2^24 grid steps): one butterfly.
    const uint32_t ebits0 = __shfl_sync(0xffffffffu, ebits, 0);
    if (cm == 0xffffffffu && __all_sync(0xffffffffu, d0 == d1 && ebits == ebits0)) {
      float t = d0;
      #pragma unroll
      for (int o = 16; o; o >>= 1) t = __fadd_rn(t, __shfl_xor_sync(0xffffffffu, t, o));
      S = __fadd_rn(S, t);
      continue;
    }
    // General step.  A transducer lane maps the parity of S / q to the parity after it: p -> p ^ parity(d_p / q), one of the
    // four maps of {0, 1} (identity, flip, constant 0, constant 1).  Two ballots hold f(0) and f(1) of every lane, and the
    // parity that ENTERS lane L of a run is read off them with bit operations: the value of the nearest constant lane before L
    // (or the run's entry parity), flipped once per flip lane in between.  Every lane then knows its increment, and the run's
    // total is one butterfly (exact: multiples of q, every partial sum below 2^23 q).  So S visits one stop per run of
    // same-binade transducer lanes and one per straddling lane (its 32 additions done by every lane on its own values, the
    // straddler's result broadcast) -- no shuffle scan of pairs, no stop per tie.
    const float rb = u2f(ebits << 23);
    const unsigned F0 = __ballot_sync(0xffffffffu, certain && (f2u(__fadd_rn(rb, d0)) & 1u) != 0u);   // f(0) = parity(d0 / q)
    const unsigned F1 = __ballot_sync(0xffffffffu, certain && (f2u(__fadd_rn(rb, d1)) & 1u) == 0u);   // f(1) = 1 ^ parity(d1 / q)
    const uint32_t eprev = __shfl_up_sync(0xffffffffu, ebits, 1);
    const bool head = certain && (lane == 0 || !((cm >> (lane - 1)) & 1u) || eprev != ebits);
    const unsigned hm = __ballot_sync(0xffffffffu, head);
    const unsigned cst = ~(F0 ^ F1), flp = F0 & ~F1;
    const unsigned lt = (1u << lane) - 1u;
    unsigned m = hm | ~cm;                                 // stops: heads of transducer runs, and every other lane
    while (m) {                                            // warp-uniform: S is the same in every lane
      const int k = __ffs((int)m) - 1;
      m &= m - 1u;
      if ((cm >> k) & 1u) {
        const unsigned upto = m ? ((1u << (__ffs((int)m) - 1)) - 1u) : 0xffffffffu;
        const unsigned seg = upto & ~((1u << k) - 1u);     // lanes k .. next stop - 1
        const unsigned before = seg & lt;                  // lanes of the run before this one
        const unsigned cb = cst & before;
        uint32_t pin;
        if (cb) {
          const int jc = 31 - __clz((int)cb);
          pin = ((F0 >> jc) & 1u) ^ ((uint32_t)__popc(flp & before & ~((2u << jc) - 1u)) & 1u);
        } else {
          pin = (f2u(S) & 1u) ^ ((uint32_t)__popc(flp & before) & 1u);
        }
        float t = ((seg >> lane) & 1u) ? (pin ? d1 : d0) : 0.0f;
        #pragma unroll
        for (int o = 16; o; o >>= 1) t = __fadd_rn(t, __shfl_xor_sync(0xffffffffu, t, o));
        S = __fadd_rn(S, t);                               // exact
      } else {
        float c = S;
        #pragma unroll
        for (int e = 0; e < DCB_XE; e++) c = __fadd_rn(c, y[e]);
        S = __shfl_sync(0xffffffffu, c, k);
      }
    }
  }
  if (slow) {
    #pragma unroll
    for (int o = 16; o; o >>= 1) ymx = fmaxf(ymx, __shfl_xor_sync(0xffffffffu, ymx, o));
  }
  __syncwarp();
  total_out = S; ymx_out = ymx; y0_out = y0;
}

template <int CT>
__global__ void __launch_bounds__(DCB_WNT, (CT == CT_MASK ? DCB_WARP_BLOCKS_CT7 : DCB_WARP_BLOCKS)) dcb_encode_warp_kernel(const EncArgs a) {
  extern __shared__ __align__(16) unsigned char dcb_warp_smem[];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  WarpBuf& sm = reinterpret_cast<WarpBuf*>(dcb_warp_smem)[warp];
  const uint32_t K = a.num_tiles;
  // chunks are handed out by a ticket counter, one chunk ahead (the next chunk's statistics are folded in during this one)
  uint32_t c = 0u, cn = 0u;
  if (lane == 0) { c = atomicAdd(a.ticket, 1u); cn = atomicAdd(a.ticket, 1u); }
  c = __shfl_sync(0xffffffffu, c, 0); cn = __shfl_sync(0xffffffffu, cn, 0);
  if (c >= K) return;

  const bool submin = (a.flags & 1u) != 0;            // DCB_F_SUBTRACT_MIN
  const bool fixed_mask = (a.flags & 2u) != 0;        // DCB_F_FIXED_MASK
  const float u24 = 1.0f / 16777216.0f;
  uint32_t* const W = sm.W;
  const uint32_t one = (a.flags >> 31) + 1u;          // 1, but opaque to the compiler (keeps the field updates on IMAD)

#if DCB_BULK_P
  if (lane == 0) mbar_init(&sm.mbar[0]);
  __syncwarp();
  uint32_t pphase = 0u;                              // parity of the statistics buffer's mbarrier
#endif
  Pass1 acc;
  p1_reset<CT>(acc);
  pass1_direct<CT>(a, (uint64_t)c * DCB_TILE, min((uint64_t)c * DCB_TILE + DCB_TILE, a.n), lane, acc, submin);

  for (;;) {
    const uint32_t cw = c;
    const uint64_t cbase = (uint64_t)c * DCB_TILE;
    const uint64_t cend = min(cbase + DCB_TILE, a.n);
    const int cnt = (int)(cend - cbase);
    const int nr = (cnt + DCB_WR - 1) / DCB_WR;
    const bool cbulk = cnt == DCB_TILE && (a.flags & DCB_ENC_IN_ALIGNED);      // every round of this chunk is a bulk round
    const bool fold_next = true;

    // ---- statistics of this chunk (folded in during the previous chunk) ------------------------------------------
    float lmn = acc.mn, lmx = acc.mx, tsum = __fadd_rn(acc.s0, acc.s1), wsum = acc.ws;
    #pragma unroll
    for (int o = 16; o; o >>= 1) {
      lmn = fminf(lmn, __shfl_xor_sync(0xffffffffu, lmn, o));
      if (CT == CT_MASK) {
        lmx = fmaxf(lmx, __shfl_xor_sync(0xffffffffu, lmx, o));
        tsum = __fadd_rn(tsum, __shfl_xor_sync(0xffffffffu, tsum, o));
        wsum = __fadd_rn(wsum, __shfl_xor_sync(0xffffffffu, wsum, o));
      }
    }
    bool odd = !(fabsf(__fadd_rn(acc.n0, acc.n1)) < __uint_as_float(0x7f800000u));
    if (!submin) odd = odd || ((acc.sg >> 31) != 0u);   // a negative value (sign bit, -0.0f included) without min-subtract
    bool slow = __any_sync(0xffffffffu, odd);
    float mn = submin ? lmn : 0.0f;
    if (submin) {
      if (mn == 0.0f) {
        // +0 / -0 compare equal: the reference keeps the FIRST minimal element; recover its sign.  A -0.0f anywhere
        // makes the chunk SLOW (x - (+0) keeps it negative and the raw code would carry its sign bit).
        uint32_t fi = 0xFFFFFFFFu; bool negz = false;
        for (int i = lane; i < cnt; i += 32) {
          const float v = __ldg(a.in + cbase + i);
          if (v == 0.0f && fi == 0xFFFFFFFFu) fi = (uint32_t)i;
          negz = negz || (f2u(v) == 0x80000000u);
        }
        #pragma unroll
        for (int o = 16; o; o >>= 1) fi = min(fi, __shfl_xor_sync(0xffffffffu, fi, o));
        if (fi != 0xFFFFFFFFu) mn = __ldg(a.in + cbase + fi);
        slow = slow || __any_sync(0xffffffffu, negz);
      }
      if (slow) {                                       // data[0] NaN: every '<' fails and min stays NaN (:3546-3553)
        const float x0 = __ldg(a.in + cbase);
        if (x0 != x0) mn = x0;
      }
      if (lane == 0) a.cmin[c] = mn;
    } else if (a.cmin && lane == 0) {
      a.cmin[c] = 0.0f;
    }

    // ---- med_dataset_float: only the top 17 bits of the sequential float32 mean reach the stream -----------------
    int type = a.fixed_type;
    uint32_t mask17 = a.fixed_mask17;
    if (CT == CT_MASK && !fixed_mask) {
      bool decided = false;
      const float fc = (float)cnt;
      if (!slow && !(a.flags & 8u)) {
        // S_seq in est +- E (see dcb_encode_slots.cuh, LEVEL 0); the raw sum here is 2 + 16 + 1 + 5 roundings deep
        const float amb = fmaxf(fabsf(lmn), fabsf(lmx));
        float est = __fsub_rn(tsum, __fmul_rn(fc, mn));
        // (ragged chunk / unaligned input: the per-lane sums are plain chains of up to 128 adds)
        const float ke = cbulk ? 32.0f * u24 : 160.0f * u24;
        float E1 = __fadd_ru(__fmul_ru(__fmul_ru(fc, amb), ke), 1e-37f);
        if (submin && lmx == lmn) { est = 0.0f; E1 = 0.0f; }
        const float top = fmaxf(__fadd_ru(est, E1), 0.0f);
        float E = __fadd_ru(E1, __fmul_ru(__fmul_ru(__fadd_ru(fc, 4.0f), 1.01f * u24), top));
        if (cbulk) {
          // tighter: S_seq - sum y = sum_k delta_k with |delta_k| <= u (s_{k-1} + y_k) <= u (1 + gamma) P_k, P_k the exact partial
          // sums, and sum_k P_k = sum_i (n - i) y_i <= (1 + u) (WX - min SW): WX = sum_i w_i x_i with w_i = n - (index of the
          // first element of i's group of four) >= n - i, SW = sum_i w_i = n (n + 4) / 2.  wsum is WX to 40 roundings.
          const float SW = 0.5f * (float)DCB_TILE * (float)(DCB_TILE + 4);                  // 8396800, exact
          const float wup = fmaxf(__fadd_ru(__fsub_ru(wsum, __fmul_rn(mn, SW)), __fmul_ru(__fmul_ru(SW, amb), 48.0f * u24)), 0.0f);
          const float E2 = __fadd_ru(E1, __fadd_ru(__fmul_ru(top, 4.0f * 1.01f * u24), __fmul_ru(wup, 1.002f * u24)));
          E = fminf(E, E2);
        }
        const float flo = fmaxf(__fsub_rd(est, E), 0.0f), fhi = __fadd_ru(est, E);
        uint32_t mlo, mhi;
        if (cnt == DCB_TILE) { mlo = f2u(__fmul_rn(flo, 1.0f / 4096.0f)) >> 15; mhi = f2u(__fmul_rn(fhi, 1.0f / 4096.0f)) >> 15; }
        else { mlo = f2u(__fdiv_rn(flo, fc)) >> 15; mhi = f2u(__fdiv_rn(fhi, fc)) >> 15; }
        decided = (mlo == mhi) && (f2u(fhi) < 0x7F800000u);
        if (decided) {
          mask17 = mlo;
          type = type_from_max(submin ? __fsub_rn(lmx, mn) : lmx);
          if (lane == 0) { a.cmed[c] = u2f(mask17 << 15); a.ctype[c] = type; }
        }
      }
      if (!decided) {
        float total, ymx, y0;
#ifdef DCB_DEBUG_MED
        const long long dbg_t0 = clock64();
#endif
        exact_medium(a, sm, cbase, cend, mn, submin, slow, cbulk, lane, total, ymx, y0);
#ifdef DCB_DEBUG_MED
        if (lane == 0) { atomicAdd(&g_dbg_exact[0], (unsigned long long)(clock64() - dbg_t0)); atomicAdd(&g_dbg_exact[1], 1ull); }
#endif
        const float med = __fdiv_rn(total, fc);
        type = slow ? ((y0 != y0) ? 0 : type_from_max(ymx)) : type_from_max(submin ? __fsub_rn(lmx, mn) : lmx);
        mask17 = f2u(med) >> 15;
        if (lane == 0) { a.cmed[c] = med; a.ctype[c] = type; }
      }
    }

    // ---- the rounds ----------------------------------------------------------------------------------------------
    const bool has_next = fold_next && cn < K;
    uint32_t tk = 0u;
    if (lane == 0) tk = atomicAdd(a.ticket, 1u);        // the chunk after next (lands while this chunk is encoded)
    const uint64_t nbase = (uint64_t)cn * DCB_TILE;
    const uint64_t nend = has_next ? min(nbase + DCB_TILE, a.n) : 0ull;
    const int nrn = has_next ? (int)((nend - nbase + DCB_WR - 1) / DCB_WR) : 0;
    const bool nbulk = has_next && (nend - nbase) == DCB_TILE && (a.flags & DCB_ENC_IN_ALIGNED);   // every round of the next chunk is a bulk round
    if (fold_next) p1_reset<CT>(acc);
    warp_issue_x(a, cbase, cend, sm.X[0], lane, cbulk || round_is_bulk(a, cbase, cend));
    if (nrn > 0) warp_issue_p(a, nbase, sm.P, lane, nbulk || round_is_bulk(a, nbase, nend), &sm.mbar[0]);
    cp_async_commit();

    uint32_t* const out32 = reinterpret_cast<uint32_t*>(a.slots + (uint64_t)cw * DCB_SLOT_PITCH);
    uint32_t wbase = 0u, carry_bits = 0u, carry = 0u;  // words already in the slot, bits of the open word, the open word
    float c1 = 0.0f, c2 = 0.0f, c3 = 0.0f;              // last three values of the previous round (after min-subtract)

    for (int j = 0; j < nr; j++) {
      cp_async_wait<0>();
      __syncwarp();
      if (j < nrn) {
        const bool pb = nbulk || round_is_bulk(a, nbase + (uint64_t)j * DCB_WR, nend);
#if DCB_BULK_P
        if (pb) { mbar_wait(&sm.mbar[0], pphase); pphase ^= 1u; }   // the round's 2 KB have landed (complete_tx)
#endif
        warp_consume_p<CT>(a, nbase + (uint64_t)j * DCB_WR, nend, sm.P, lane, acc, submin, pb, j);
#if DCB_BULK_P
        __syncwarp();                                  // every lane has read P before lane 0 lets the next copy overwrite it
#endif
      }
      if (j + 1 < nr) warp_issue_x(a, cbase + (uint64_t)(j + 1) * DCB_WR, cend, sm.X[(j + 1) & 1], lane,
                                   cbulk || round_is_bulk(a, cbase + (uint64_t)(j + 1) * DCB_WR, cend));
      if (j + 1 < nrn) warp_issue_p(a, nbase + (uint64_t)(j + 1) * DCB_WR, sm.P, lane,
                                    nbulk || round_is_bulk(a, nbase + (uint64_t)(j + 1) * DCB_WR, nend), &sm.mbar[0]);
      cp_async_commit();

      const uint64_t my0 = cbase + (uint64_t)j * DCB_WR + (uint64_t)lane * DCB_E;
      const int vcnt = (my0 >= cend) ? 0 : (int)min((uint64_t)DCB_E, cend - my0);
      const bool first = (j == 0 && lane == 0);        // chunk start: elements 0..2 have no full history (:3330)
      const float* slot = &sm.X[j & 1][lane * DCB_XPITCH];
      float x[DCB_E];
      #pragma unroll
      for (int q = 0; q < 4; q++) {
        const float4 v = *reinterpret_cast<const float4*>(slot + 4 * q);
        x[4 * q] = v.x; x[4 * q + 1] = v.y; x[4 * q + 2] = v.z; x[4 * q + 3] = v.w;
      }
#if DCB_PACKED
      // element e and element e + 8 as the two halves of one 64-bit register: the min-subtract writes the pair (a scalar FADD per
      // half, inside one asm block so that ptxas allocates the two results as the pair instead of moving them there)
      uint64_t XP[DCB_E / 2];
      if (cbase + (uint64_t)(j + 1) * DCB_WR > cend) {     // warp-uniform: a ragged round, padding becomes 0.0f
        #pragma unroll
        for (int e = 0; e < DCB_E; e++) x[e] = (e < vcnt) ? x[e] : mn;
      }
      if (slow && !submin) {                               // rare: NaN payloads must survive, so no arithmetic at all
        #pragma unroll
        for (int e = 0; e < DCB_E / 2; e++) XP[e] = pk2(x[e], x[e + 8]);
      } else {
        #pragma unroll
        for (int e = 0; e < DCB_E / 2; e++)
          asm("{\n\t.reg .f32 lo, hi;\n\tsub.rn.f32 lo, %1, %3;\n\tsub.rn.f32 hi, %2, %3;\n\tmov.b64 %0, {lo, hi};\n\t}"
              : "=l"(XP[e]) : "f"(x[e]), "f"(x[e + 8]), "f"(mn));              // mn = 0.0f without min-subtract: x - 0 is x
      }
      #pragma unroll
      for (int e = 0; e < DCB_E / 2; e++) upk2(XP[e], x[e], x[e + 8]);
#else
      if (submin) {
        if (cbase + (uint64_t)(j + 1) * DCB_WR <= cend) {   // warp-uniform: a full round
          #pragma unroll
          for (int e = 0; e < DCB_E; e++) x[e] = __fsub_rn(x[e], mn);
        } else {
          #pragma unroll
          for (int e = 0; e < DCB_E; e++) x[e] = (e < vcnt) ? __fsub_rn(x[e], mn) : 0.0f;
        }
      }
#endif
      float h1 = __shfl_up_sync(0xffffffffu, x[DCB_E - 1], 1);
      float h2 = __shfl_up_sync(0xffffffffu, x[DCB_E - 2], 1);
      float h3 = __shfl_up_sync(0xffffffffu, x[DCB_E - 3], 1);
      if (lane == 0) { h1 = c1; h2 = c2; h3 = c3; }

      // ---- classify ----------------------------------------------------------------------------------------------
      uint32_t code[DCB_E];
      int len[DCB_E];
      uint32_t chi = 0u, clo = 0u;
      uint32_t total = 0;
      bool all3 = false;
      int np = DCB_E;                                  // pieces the packing loop has to place (warp-uniform)
      bool uni48 = false;                              // warp-uniform: every lane holds sixteen 3-bit codes (48 bits)
      if (!slow) {
        uint32_t miss = 0u;                              // bit e: element e needs a fallback code
        if (CT != CT_NP) {
          uint32_t fa = 0x92492490u, fb = 0x92490000u;   // 3-bit fields of elements 0..9 / 10..15, all preset to 100 (zero)
          const float tp3 = first ? -1.0f : a.k.tp;      // chunk start: elements 0..2 never hit
#if DCB_PACKED
          // The encoder is open loop (history = the INPUT values, :3386-3388), so elements are independent: element e and
          // element e + 8 of this lane share every instruction of the predictor arithmetic as the two halves of a packed
          // pair.  3*b is fl(2b + b): 2b is exact, so the sum rounds once, exactly like the reference's 3*b (and ptxas cannot
          // contract two additions into an FFMA2, which it does with mul.rn.f32x2 followed by sub.rn.f32x2).
          uint64_t B1 = pk2(h1, x[7]), B2 = pk2(h2, x[6]), B3 = pk2(h3, x[5]);
          uint64_t W1 = add2(B1, B1), W2 = add2(B2, B2);                      // 2*b1, 2*b2
          uint64_t T1 = add2(W1, B1), T2 = add2(W2, B2);                      // 3*b1, 3*b2
          #define DCB_CLS2(E, ACCL, POSL, TPL, ACCH, POSH) do {                                                    \
              const uint64_t X = XP[E];                                                                            \
              const uint64_t D1 = sub2(B1, X);                                                                     \
              const uint64_t D2 = sub2(sub2(W1, B2), X);                                                           \
              const uint64_t D3 = sub2(add2(sub2(T1, T2), B3), X);                                                 \
              float d1l, d1h, d2l, d2h, d3l, d3h;                                                                  \
              upk2(D1, d1l, d1h); upk2(D2, d2l, d2h); upk2(D3, d3l, d3h);                                          \
              classify_pred<(1u << (POSL))>(ACCL, miss, x[E], d1l, d2l, d3l, a.k.tz, TPL, one, 1u << (E));         \
              classify_pred<(1u << (POSH))>(ACCH, miss, x[(E) + 8], d1h, d2h, d3h, a.k.tz, a.k.tp, one, 1u << ((E) + 8)); \
              B3 = B2; B2 = B1; B1 = X; T2 = T1; W1 = add2(X, X); T1 = add2(W1, X);                                \
            } while (0)
          DCB_CLS2(0, fa, 29, tp3, fa, 5); DCB_CLS2(1, fa, 26, tp3, fa, 2); DCB_CLS2(2, fa, 23, tp3, fb, 29);
          DCB_CLS2(3, fa, 20, a.k.tp, fb, 26); DCB_CLS2(4, fa, 17, a.k.tp, fb, 23); DCB_CLS2(5, fa, 14, a.k.tp, fb, 20);
          DCB_CLS2(6, fa, 11, a.k.tp, fb, 17); DCB_CLS2(7, fa, 8, a.k.tp, fb, 14);
          #undef DCB_CLS2
#else
          float b1 = h1, b2 = h2, b3 = h3;
          float t1 = __fmul_rn(3.0f, b1), t2 = __fmul_rn(3.0f, b2);
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
#endif
          chi = fa | (fb >> 30); clo = fb << 2;          // the 48-bit string: 30 bits of fa, 18 bits of fb
          all3 = (vcnt == DCB_E) && (miss == 0u);
        }
        // The first three elements of a chunk never hit (:3330), so the chunk's first lane always needs fallback codes: when that
        // is the only reason the warp is not pure 3-bit codes, only that lane leaves the fast path, with five pieces to pack
        // (three codes, then its remaining thirteen 3-bit codes as 32 + 7 bits) instead of sixteen.
        const bool warp_all3 = __all_sync(0xffffffffu, all3);
        const bool head_ok = all3 || (CT != CT_NP && first && vcnt == DCB_E && (miss & ~7u) == 0u);
        const bool headcase = !warp_all3 && __all_sync(0xffffffffu, head_ok);
        uni48 = warp_all3;
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
          if (cbase + (uint64_t)(j + 1) * DCB_WR > cend) {     // warp-uniform: a ragged round, padding carries no code
            #pragma unroll
            for (int e = 0; e < DCB_E; e++)
              if (e >= vcnt) { total -= (uint32_t)len[e]; code[e] = 0u; len[e] = 0; }
          }
        }
      } else {
        // -1.0f is the reference's "empty history slot" marker: rebuild the exact state this lane
        // starts from by walking back over the chunk until three non-sentinel values are found.
        float b3 = -1.0f, b2 = -1.0f, b1 = -1.0f;
        if (vcnt > 0 && CT != CT_NP) {
          float n0 = 0.0f, n1 = 0.0f, n2 = 0.0f;
          int found = 0; bool last_sent = false;
          for (uint64_t idx = my0; idx > cbase && found < 3; idx--) {
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
      c1 = __shfl_sync(0xffffffffu, x[DCB_E - 1], 31);
      c2 = __shfl_sync(0xffffffffu, x[DCB_E - 2], 31);
      c3 = __shfl_sync(0xffffffffu, x[DCB_E - 3], 31);

      // ---- warp scan of the bit counts; the round starts at bit carry_bits of the buffer -----------------------------
      uint32_t incl, R;
      unsigned havebits = 0xffffffffu;
      bool fast_pack = true;
      if (uni48) {                                                              // the common round of compressible data: no scan
        incl = 48u * (uint32_t)(lane + 1);
        R = carry_bits + 32u * 48u;
      } else {
        incl = total;
        #pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
          if (lane >= o) incl += t;
        }
        R = carry_bits + __shfl_sync(0xffffffffu, incl, 31);                    // bits in the buffer after this round
        havebits = __ballot_sync(0xffffffffu, total > 0u);
        fast_pack = __all_sync(0xffffffffu, total >= 32u || total == 0u);
      }
      const uint32_t s = carry_bits + incl - total;                             // first bit of this lane

      // ---- pack ----------------------------------------------------------------------------------------------------
      uint32_t hi = 0u;
      const uint32_t f0 = s & 31u;
      const uint32_t hw = s >> 5;
      uint32_t newcarry;
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
        // the word straddling two lanes is stored by the later one; it ORs in the earlier lane's tail (lane 0: the carry)
        uint32_t prevT = __shfl_up_sync(0xffffffffu, hi, 1);
        if (lane == 0) prevT = carry;
        if (total > 0u && f0) W[hw] |= prevT;
        const int lastl = 31 - __clz((int)havebits);                            // last lane that has bits (a round has >= 1 element)
        newcarry = __shfl_sync(0xffffffffu, hi, lastl);
        __syncwarp();
      } else {
        // ragged round (a lane with fewer than 11 valid elements): zero the buffer and OR every word in
        for (int i = lane; i < DCB_WWORDS; i += 32) W[i] = 0u;
        __syncwarp();
        if (lane == 0 && carry_bits) atomicOr(&W[0], carry);
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
        __syncwarp();
        newcarry = W[R >> 5];
      }

      // ---- whole words leave for the slot ----------------------------------------------------------------------------
      const uint32_t fullw = R >> 5;
      {
        uint32_t* const o = out32 + wbase;                                      // a round of 3-bit codes is 48 or 49 words
        if ((uint32_t)lane < fullw) o[lane] = __byte_perm(W[lane], 0u, 0x0123);
        if ((uint32_t)lane + 32u < fullw) o[lane + 32] = __byte_perm(W[lane + 32], 0u, 0x0123);
        if (fullw > 64u)
          for (uint32_t k = (uint32_t)lane + 64u; k < fullw; k += 32u) o[k] = __byte_perm(W[k], 0u, 0x0123);
      }
      wbase += fullw;
      carry_bits = R & 31u;
      carry = carry_bits ? newcarry : 0u;
    }

    // ---- chunk end: the open word (zero padded, :5467), the size ----------------------------------------------------
    if (lane == 0) {
      if (carry_bits) out32[wbase] = __byte_perm(carry, 0u, 0x0123);
      const uint32_t bits = wbase * 32u + carry_bits;
      const uint32_t bytes = (bits + 7u) >> 3;
      a.sizes[cw] = bytes;
      atomicAdd(&a.gsum[cw / DCB_GROUP], bytes);
      if (cw == K - 1u) *a.last_tile_bits = bits;
    }
    if (!has_next) break;
    c = cn;
    cn = __shfl_sync(0xffffffffu, tk, 0);
  }
}
