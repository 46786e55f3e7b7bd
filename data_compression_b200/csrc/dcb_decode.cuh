// dcb_decode.cuh -- decoder for CT=5/6/7/11 (replaces myDecompress_bitwise{,_np,_mask,_op},
// dataCompression.c:2922, :2459, :1703, :698 and decompress_bitwise_float{,_np,_mask} :3137, :2611, :1900).
//
// Decoding has two serial dependencies per chunk: code boundaries (a code's length is known
// only after its first bit and exponent are read) and the float recurrence on predicted values (the decoder
// predicts from RECONSTRUCTED values; float rounding makes the recurrence non-associative).  Chunks are
// independent, so ONE LANE walks ONE CHUNK and the parallelism is across chunks.  v2 (profiles/r1_v1_summary.md
// showed v1's cooperative staging cost 2/3 of the time):
//   - every lane streams its own chunk with 16-byte loads into a 3-deep register queue (two blocks always in
//     flight), and writes its own output as full 32-byte sectors: no shared memory, no shuffles, no barriers
//   - a round is 8 elements; when the next 8 codes are all 3-bit codes (one mask test) the round runs a
//     parse-free path (static bit tests + the float recurrence), otherwise the general per-code path
//   - the reference's -1.0f "empty history slot" state machine is exact in the general path; the fast path
//     is only taken with a full history and re-runs the round generally if any value could be the sentinel
#pragma once
#include "dcb_common.cuh"

#define DEC_NT 128
#define DEC_RND 8

struct BitReader {
  const uint4* p;       // next 16-byte block to fetch
  const uint4* pend;    // first block that starts at/after the end of this chunk's stream
  uint4 cur, n1, n2;    // word queue (memory order); cur.x is the next word
  int cnt;              // words left in cur
  uint32_t hi, lo;      // bit buffer: hi = next 32 stream bits (MSB first), lo = the bits after them
  int nlo;              // valid bits in lo (0..32)

  __device__ __forceinline__ uint4 fetch() {
    uint4 v = make_uint4(0u, 0u, 0u, 0u);               // past the chunk: zeros (never dereferenced)
    if (p < pend) v = __ldg(p);
    p++;
    return v;
  }
  __device__ __forceinline__ uint32_t next_word() {
    const uint32_t w = cur.x;
    cur.x = cur.y; cur.y = cur.z; cur.z = cur.w;
    if (--cnt == 0) { cur = n1; n1 = n2; n2 = fetch(); cnt = 4; }
    return __byte_perm(w, 0u, 0x0123);                   // memory order -> MSB-first word
  }
  __device__ __forceinline__ void init(const uint8_t* stream, uint64_t byte_off, uint64_t byte_end) {
    p = reinterpret_cast<const uint4*>(stream + (byte_off & ~15ull));
    pend = reinterpret_cast<const uint4*>(stream + ((byte_end + 15ull) & ~15ull));
    cur = fetch(); n1 = fetch(); n2 = fetch();
    cnt = 4;
    const int skipw = (int)((byte_off & 15ull) >> 2);
    for (int i = 0; i < skipw; i++) (void)next_word();
    hi = next_word(); lo = next_word(); nlo = 32;
    const int skipb = (int)(byte_off & 3ull) * 8;
    if (skipb) consume(skipb);
  }
  // drop L (1..32) bits
  __device__ __forceinline__ void consume(int L) {
    hi = __funnelshift_lc(lo, hi, L);
    lo = __funnelshift_lc(0u, lo, L);
    nlo -= L;
    if (nlo < 0) {                                       // fewer than 32 valid bits left in hi
      const int valid = 32 + nlo;                        // 0..31
      const uint32_t w = next_word();
      hi |= __funnelshift_rc(w, 0u, valid);              // w >> valid  (valid = 0 -> w)
      lo = __funnelshift_lc(0u, w, 32 - valid);          // w << (32 - valid)  (valid = 0 -> 0)
      nlo = valid;
    }
  }
};

// one code of the general path: value + length from the next 32 bits `w`
template <int CT>
__device__ __forceinline__ void decode_one(uint32_t w, const CodecConst& k, int type, uint32_t ones, uint32_t mask17,
                                           int m_mask, float b1, float b2, float b3, float& v, int& len, bool& bad) {
  if ((w >> 31) && CT != CT_NP) {                                // 3-bit code (:2965-2967)
    len = 3;
    const float p2 = pred2(b1, b2), p3 = pred3(b1, b2, b3);     // decompress_bitwise_float :3139-3158
    const float lo2 = (w & 0x20000000u) ? b1 : 0.0f;            // 100 -> 0, 101 -> b1
    const float hi2 = (w & 0x20000000u) ? p3 : p2;              // 110 -> 2*b1-b2, 111 -> 3*b1-3*b2+b3
    v = (w & 0x40000000u) ? hi2 : lo2;
    return;
  }
  if (CT == CT_OP) { len = 32; v = u2f(w); return; }             // :718-721
  if (CT == CT_NP && (w >> 31)) bad = true;                      // :2501-2505 "Error leading bit 1"
  bool escape = false;
  if (CT == CT_MASK) escape = (((w >> (31 - type)) & ones) == ones);   // :1739-1758
  if (!escape) {                                                 // raw: 9 bits, m from the exponent just read
    len = raw_len(w, k.n_eb_m118);
    const uint32_t keep = ~__funnelshift_rc(0xFFFFFFFFu, 0u, len);     // top `len` bits (len = 32 -> all)
    v = u2f((w & keep) | __funnelshift_rc(0x80000000u, 0u, len));       // + midpoint bit, :3168-3184
  } else {
    const uint32_t flag = (w >> (30 - type)) & 1u;
    const uint32_t kb = flag ? (uint32_t)m_mask : (uint32_t)max(m_mask - 8, 0);
    const uint32_t hdr = (uint32_t)type + 2u;
    const uint32_t pay = kb ? ((w << hdr) >> (32u - kb)) : 0u;
    len = (int)(hdr + kb);
    const uint32_t have = (flag ? 9u : 17u) + kb;
    uint32_t top = flag ? (((mask17 >> 8) << 23) | (kb ? pay << (23u - kb) : 0u))   // :1963-2010
                        : ((mask17 << 15) | (kb ? pay << (15u - kb) : 0u));          // :1947-1962
    top |= __funnelshift_rc(0x80000000u, 0u, have);
    v = u2f(top);
  }
}

template <int CT>
__global__ void __launch_bounds__(DEC_NT) dcb_decode_kernel(const DecArgs a) {
  const uint64_t c = (uint64_t)blockIdx.x * DEC_NT + threadIdx.x;
  if (c >= a.num_chunks) return;
  const uint64_t elem0 = c * a.chunk;
  const uint32_t cn = (uint32_t)min(a.chunk, a.n - elem0);
  const bool addmin = (a.flags & 1u) != 0;
  const float mn = (addmin && a.cmin) ? a.cmin[c] : 0.0f;
  int type = a.fixed_type;
  uint32_t mask17 = a.fixed_mask17;
  if (CT == CT_MASK && !(a.flags & 2u)) { type = a.ctype[c]; mask17 = f2u(a.cmed[c]) >> 15; }
  type = min(max(type, 0), 7);
  const int m_mask = raw_len(mask17 << 15, a.k.n_eb_m118) - 9;     // escapes take m from the MASK's exponent (:1769)
  const uint32_t ones = (1u << type) - 1u;

  BitReader br;
  br.init(a.stream, a.offsets[c], a.offsets[c + 1]);
  float* out = a.out + elem0;
  const bool vec_ok = ((reinterpret_cast<uintptr_t>(out) & 31) == 0);

  float b3 = -1.0f, b2 = -1.0f, b1 = -1.0f;                         // :2930
  bool bad = false;
  uint32_t i = 0;
  for (; i + DEC_RND <= cn; i += DEC_RND) {
    float v[DEC_RND];
    bool done = false;
    if (CT != CT_NP) {
      // fast round: the next 8 codes all start with 1 (3-bit codes) and the history is full
      const bool clean = (b1 != -1.0f) && (b2 != -1.0f) && (b3 != -1.0f);
      if (clean && (br.hi & 0x92492400u) == 0x92492400u) {
        const uint32_t w = br.hi;
        float c1 = b1, c2 = b2, c3 = b3;
        float t1 = __fmul_rn(3.0f, c1), t2 = __fmul_rn(3.0f, c2);
        uint32_t neg = 0u;
        #pragma unroll
        for (int e = 0; e < DEC_RND; e++) {
          const float p2 = __fsub_rn(__fadd_rn(c1, c1), c2);          // 2*b1 is exact: same as __fmul_rn(2, b1)
          const float p3 = __fadd_rn(__fsub_rn(t1, t2), c3);
          const bool bb = (w & (0x20000000u >> (3 * e))) != 0u;
          const bool ba = (w & (0x40000000u >> (3 * e))) != 0u;
          const float lo2 = bb ? c1 : 0.0f;
          const float hi2 = bb ? p3 : p2;
          const float x = ba ? hi2 : lo2;
          v[e] = x;
          neg |= f2u(x);
          c3 = c2; c2 = c1; c1 = x;
          t2 = t1; t1 = __fmul_rn(3.0f, x);
        }
        if (!(neg >> 31)) {                  // no negative value -> none is the -1.0f sentinel: commit
          b1 = c1; b2 = c2; b3 = c3;
          br.consume(3 * DEC_RND);
          done = true;
        }
      }
    }
    if (!done) {
      #pragma unroll
      for (int e = 0; e < DEC_RND; e++) {
        int len;
        decode_one<CT>(br.hi, a.k, type, ones, mask17, m_mask, b1, b2, b3, v[e], len, bad);
        if (CT != CT_NP) hist_push(b3, b2, b1, v[e]);               // :3110-3127
        br.consume(len);
      }
    }
    if (addmin) {
      #pragma unroll
      for (int e = 0; e < DEC_RND; e++) v[e] = __fadd_rn(v[e], mn);  // decompressed[i] + min, impl/pingpong.c:562
    }
    if (vec_ok) {
      float4* o4 = reinterpret_cast<float4*>(out + i);
      o4[0] = make_float4(v[0], v[1], v[2], v[3]);
      o4[1] = make_float4(v[4], v[5], v[6], v[7]);
    } else {
      #pragma unroll
      for (int e = 0; e < DEC_RND; e++) out[i + e] = v[e];
    }
  }
  for (; i < cn; i++) {                                              // ragged tail of the last chunk
    float x; int len;
    decode_one<CT>(br.hi, a.k, type, ones, mask17, m_mask, b1, b2, b3, x, len, bad);
    if (CT != CT_NP) hist_push(b3, b2, b1, x);
    br.consume(len);
    out[i] = addmin ? __fadd_rn(x, mn) : x;
  }
  if (bad && a.status) atomicExch(a.status, -5);
}
