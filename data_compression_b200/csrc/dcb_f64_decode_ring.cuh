// dcb_f64_decode_ring.cuh -- chunked float64 decoder in the shape of the float32 one (dcb_decode.cuh): one lane per chunk, the
// lane's stream staged by cp.async into a lane-private shared-memory ring, rounds of eight codes that are either parse-free
// (eight 3-bit codes) or branch-free (any mix of raw / escape / full-word / 3-bit codes), the history as register renaming with
// an exact redo when the -1.0 empty-slot marker shows up (:2797-2814).
// Why (round-2 profile of dcb_f64_decode_kernel, CT=6, 2^27 doubles, 1.44 ms): 85 instructions per code behind data-dependent
// branches (code kind, 0 / 1 / 2 words of advance, store pairing) and 36 % of the samples waiting for the stream word that was
// requested "two words ahead" but is rotated into the window -- hence waited for -- at the very next advance.  Here a code
// costs straight-line selects, and the only loads on the lane's path are shared-memory reads issued one code before their use.
#pragma once
#include "dcb_decode.cuh"             // dec_cp_async16, dec_cp_commit, lds32
#include "dcb_decode_single_f64.cuh"  // f64_decode_fallback

#define F64R_NT 128
#define F64R_RND 8
#define F64R_RING_BLOCKS 16           // 16-byte copies per lane ring: a round consumes up to 64 bytes
#define F64R_RING_WORDS (4 * F64R_RING_BLOCKS)
#define F64R_PITCH (F64R_RING_WORDS + 4)
#ifndef F64R_LANES
#define F64R_LANES 32                 // lanes of a warp that decode (one chunk each)
#endif

// Per-lane MSB-first bit reader with a THREE-word window (a code has up to 64 bits and starts at any bit of w0) and the two
// words behind it already read from the ring (an advance is 0, 1 or 2 words).
struct LaneReader64 {
  uint32_t pos;           // position of the next stream bit, counted from ring word 0 of block 0
  uint32_t w0, w1, w2;    // stream words at pos >> 5 and after it, MSB-first
  uint32_t w3raw, w4raw;  // the next two words as they lie in memory
  uint32_t ring, fetched, nblocks;
  const uint8_t* gbase;

  __device__ __forceinline__ void fetch_if(bool want) {
    if (want && fetched < nblocks) dec_cp_async16(ring + ((fetched & (F64R_RING_BLOCKS - 1u)) << 4), gbase + ((size_t)fetched << 4));
    fetched += want ? 1u : 0u;
  }
  // once per round (<= 8 codes = <= 64 bytes since the last call): blocks cur .. cur + 15 requested; what this round can touch
  // (<= cur + 6) was requested at least two calls ago, so "all but the two newest groups" covers it
  __device__ __forceinline__ void top_up() {
    const uint32_t want_until = (pos >> 7) + F64R_RING_BLOCKS;
    #pragma unroll
    for (int q = 0; q < 5; q++) fetch_if(fetched < want_until);
    dec_cp_commit();
    asm volatile("cp.async.wait_group 2;\n" ::: "memory");
  }
  __device__ __forceinline__ uint32_t ring_word(uint32_t word) const {
    return lds32(ring + ((word & (F64R_RING_WORDS - 1u)) << 2));
  }
  __device__ __forceinline__ uint64_t peek() const {                  // next 64 bits
    return ((uint64_t)__funnelshift_l(w1, w0, pos) << 32) | (uint64_t)__funnelshift_l(w2, w1, pos);
  }
  __device__ __forceinline__ void consume(int L) {                    // L in 0..64
    const uint32_t np = pos + (uint32_t)L;
    const uint32_t adv = (np >> 5) - (pos >> 5);                      // 0, 1 or 2
    const uint32_t s3 = __byte_perm(w3raw, 0u, 0x0123), s4 = __byte_perm(w4raw, 0u, 0x0123);
    const uint32_t a0 = adv == 1u ? w1 : w2, a1 = adv == 1u ? w2 : s3, a2 = adv == 1u ? s3 : s4;
    w0 = adv ? a0 : w0; w1 = adv ? a1 : w1; w2 = adv ? a2 : w2;
    pos = np;
    const uint32_t wi = np >> 5;
    w3raw = ring_word(wi + 3u); w4raw = ring_word(wi + 4u);
  }
  __device__ __forceinline__ void init(const uint8_t* stream, uint64_t byte_off, uint64_t byte_end, uint32_t ring_addr) {
    ring = ring_addr;
    gbase = stream + (byte_off & ~15ull);
    nblocks = (uint32_t)((((byte_end + 15ull) & ~15ull) - (byte_off & ~15ull)) >> 4);
    fetched = 0u;
    pos = (uint32_t)(byte_off & 15ull) * 8u;
    #pragma unroll
    for (int b = 0; b < F64R_RING_BLOCKS; b++) fetch_if(true);
    dec_cp_commit();
    asm volatile("cp.async.wait_group 0;\n" ::: "memory");
    const uint32_t wi = pos >> 5;
    w0 = __byte_perm(ring_word(wi), 0u, 0x0123);
    w1 = __byte_perm(ring_word(wi + 1u), 0u, 0x0123);
    w2 = __byte_perm(ring_word(wi + 2u), 0u, 0x0123);
    w3raw = ring_word(wi + 3u); w4raw = ring_word(wi + 4u);
  }
};

// the reference's history update (:2797-2814): fill b3, b2, b1 while a slot is empty, else shift
__device__ __forceinline__ void f64_hist_push(double& b3, double& b2, double& b1, double v) {
  const bool full = b3 != -1.0 && b2 != -1.0 && b1 != -1.0;
  if (full && v != -1.0) { b3 = b2; b2 = b1; b1 = v; }
  else if (b3 == -1.0) b3 = v;
  else if (b2 == -1.0) b2 = v;
  else if (b1 == -1.0) b1 = v;
  else { b3 = b2; b2 = b1; b1 = v; }
}

// one code with the exact state machine (chunk head, tail, and rounds that met the marker)
template <int CT>
__device__ __forceinline__ double f64r_one_exact(LaneReader64& br, const F64Const& k, int type, uint32_t mask20, int m_mask,
                                                 double& b3, double& b2, double& b1, bool& bad) {
  const uint64_t w = br.peek();
  double v; int len;
  if (CT != CT_NP && (w >> 63)) {
    len = 3;
    const uint32_t sel = (uint32_t)(w >> 61) & 3u;
    if (sel == 0) v = 0.0;
    else if (sel == 1) v = b1;
    else if (sel == 2) v = __dsub_rn(__dmul_rn(2.0, b1), b2);
    else v = __dadd_rn(__dsub_rn(__dmul_rn(3.0, b1), __dmul_rn(3.0, b2)), b3);
  } else {
    uint64_t vb;
    f64_decode_fallback<CT>(w, k, type, mask20, m_mask, vb, len);
    if (CT == CT_NP) bad |= (w >> 63) != 0ull;
    v = __longlong_as_double((long long)vb);
  }
  if (CT != CT_NP) f64_hist_push(b3, b2, b1, v);
  br.consume(len);
  return v;
}

// one round: eight codes of one lane -> v[] (before the minimum is added back); needs b3, b2, b1 all non-marker on entry
template <int CT>
__device__ __forceinline__ void f64r_round(LaneReader64& br, const F64Const& k, int type, uint32_t mask20, int m_mask,
                                           double& b3, double& b2, double& b1, bool& bad, double (&v)[F64R_RND]) {
  bool done = false;
  br.top_up();
  if (CT != CT_NP) {
    const uint32_t t32 = (uint32_t)(br.peek() >> 32);
    if ((t32 & 0x92492400u) == 0x92492400u) {               // eight 3-bit codes: no parse
      double c1 = b1, c2 = b2, c3 = b3;
      double t1 = __dmul_rn(3.0, c1), t2 = __dmul_rn(3.0, c2);
      uint32_t neg = 0u;
      #pragma unroll
      for (int e = 0; e < F64R_RND; e++) {
        const double p2 = __dsub_rn(__dadd_rn(c1, c1), c2);          // 2.0 * b1 is exact
        const double p3 = __dadd_rn(__dsub_rn(t1, t2), c3);
        const bool bb = (t32 & (0x20000000u >> (3 * e))) != 0u;
        const bool ba = (t32 & (0x40000000u >> (3 * e))) != 0u;
        const double x = ba ? (bb ? p3 : p2) : (bb ? c1 : 0.0);
        v[e] = x;
        neg |= (uint32_t)__double2hiint(x);
        c3 = c2; c2 = c1; c1 = x; t2 = t1; t1 = __dmul_rn(3.0, x);
      }
      bool ok = !(neg >> 31);                                        // a negative value is only a reason to look for the marker
      if (!ok) {
        bool mk = false;
        #pragma unroll
        for (int e = 0; e < F64R_RND; e++) mk = mk || (v[e] == -1.0);
        ok = !mk;
      }
      if (ok) { b1 = c1; b2 = c2; b3 = c3; br.consume(3 * F64R_RND); done = true; }
    }
  }
  if (!done) {
    // general round, branch free; the history is renamed assuming no value is the marker, else rewound and redone exactly
    const LaneReader64 br0 = br;
    const double s1 = b1, s2 = b2, s3 = b3;
    bool marker = false;
    #pragma unroll
    for (int e = 0; e < F64R_RND; e++) {
      const uint64_t w = br.peek();
      uint64_t vb; int len;
      f64_decode_fallback<CT>(w, k, type, mask20, m_mask, vb, len);
      double x = __longlong_as_double((long long)vb);
      if (CT == CT_NP) {
        bad |= (w >> 63) != 0ull;
      } else {
        const bool is3 = (w >> 63) != 0ull;
        const double p2 = __dsub_rn(__dadd_rn(b1, b1), b2);
        const double p3 = __dadd_rn(__dsub_rn(__dmul_rn(3.0, b1), __dmul_rn(3.0, b2)), b3);
        const bool bb = ((uint32_t)(w >> 61) & 1u) != 0u, ba = ((uint32_t)(w >> 62) & 1u) != 0u;
        const double x3 = ba ? (bb ? p3 : p2) : (bb ? b1 : 0.0);
        x = is3 ? x3 : x;
        len = is3 ? 3 : len;
        marker = marker || (x == -1.0);
        b3 = b2; b2 = b1; b1 = x;
      }
      v[e] = x;
      br.consume(len);
    }
    if (CT != CT_NP && marker) {                                     // rare: redo with the exact state machine
      br = br0; b1 = s1; b2 = s2; b3 = s3;
      #pragma unroll 1
      for (int e = 0; e < F64R_RND; e++) {
        const double x = f64r_one_exact<CT>(br, k, type, mask20, m_mask, b3, b2, b1, bad);
        #pragma unroll
        for (int j = 0; j < F64R_RND; j++) v[j] = (j == e) ? x : v[j];
      }
    }
  }
}

template <int CT>
__global__ void __launch_bounds__(F64R_NT) dcb_f64_decode_ring_kernel(const F64DecArgs a) {
  __shared__ __align__(16) uint32_t rings[F64R_NT * F64R_PITCH];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint64_t c_raw = ((uint64_t)blockIdx.x * (F64R_NT / 32) + warp) * F64R_LANES + lane;
  const bool active = lane < F64R_LANES && c_raw < a.num_chunks;
  if (!active) return;                                               // nothing below is warp-collective
  const uint64_t c = c_raw, lo = c * a.chunk;
  const uint32_t cn = (uint32_t)min(a.chunk, a.n - lo);
  const bool addmin = (a.flags & DCB_F_SUBTRACT_MIN) != 0;
  const double mn = addmin ? a.cmin[c] : 0.0;
  int type = a.fixed_type; uint32_t mask20 = a.fixed_mask20;
  if (CT == CT_MASK && !(a.flags & DCB_F_FIXED_MASK)) { type = a.ctype[c]; mask20 = a.cmask[c]; }
  const int m_mask = f64_mant_bits((uint64_t)mask20 << 44, a.k.n_eb);
  LaneReader64 br;
  br.init(a.stream, a.offsets[c], a.offsets[c + 1], (uint32_t)__cvta_generic_to_shared(rings + threadIdx.x * F64R_PITCH));
  double* out = a.out + lo;
  const bool vec = (reinterpret_cast<uintptr_t>(out) & 15) == 0;
  double b3 = -1.0, b2 = -1.0, b1 = -1.0;
  bool bad = false;
  uint32_t i = 0;
  // head: the state machine verbatim until all three history slots hold values (and the position is even for paired stores)
  if (CT != CT_NP) {
    while (i < cn && ((i & 1u) != 0u || b3 == -1.0 || b2 == -1.0 || b1 == -1.0)) {
      if ((i & 7u) == 0u) br.top_up();
      const double x = f64r_one_exact<CT>(br, a.k, type, mask20, m_mask, b3, b2, b1, bad);
      out[i] = addmin ? __dadd_rn(x, mn) : x;
      i++;
    }
  }
  for (; i + F64R_RND <= cn; i += F64R_RND) {
    double v[F64R_RND];
    if (CT != CT_NP && (b3 == -1.0 || b2 == -1.0 || b1 == -1.0)) {    // a marker emptied a slot: exact until the history is full again
      br.top_up();
      #pragma unroll 1
      for (int e = 0; e < F64R_RND; e++) {
        const double x = f64r_one_exact<CT>(br, a.k, type, mask20, m_mask, b3, b2, b1, bad);
        out[i + e] = addmin ? __dadd_rn(x, mn) : x;
      }
      continue;
    }
    f64r_round<CT>(br, a.k, type, mask20, m_mask, b3, b2, b1, bad, v);
    if (addmin) {
      #pragma unroll
      for (int e = 0; e < F64R_RND; e++) v[e] = __dadd_rn(v[e], mn);
    }
    if (vec) {
      #pragma unroll
      for (int e = 0; e < F64R_RND; e += 2) *reinterpret_cast<double2*>(out + i + e) = make_double2(v[e], v[e + 1]);
    } else {
      #pragma unroll
      for (int e = 0; e < F64R_RND; e++) out[i + e] = v[e];
    }
  }
  if (i < cn) br.top_up();
  for (; i < cn; i++) {                                               // ragged tail
    const double x = f64r_one_exact<CT>(br, a.k, type, mask20, m_mask, b3, b2, b1, bad);
    out[i] = addmin ? __dadd_rn(x, mn) : x;
  }
  if (bad && a.status) atomicExch(a.status, DCB_EFORMAT);
}
