// dcb_decode.cuh -- decoder for CT=5/6/7/11 (replaces myDecompress_bitwise{,_np,_mask,_op},
// dataCompression.c:2922, :2459, :1703, :698 and decompress_bitwise_float{,_np,_mask} :3137, :2611, :1900).
//
// Decoding has two serial dependencies per chunk: code boundaries (a code's length is known
// only after its first bit and exponent are read) and the float recurrence on predicted values (the decoder
// predicts from RECONSTRUCTED values; float rounding makes the recurrence non-associative).  Chunks are
// independent, so ONE LANE walks ONE CHUNK and the parallelism is across chunks.  With 65,536 chunks in
// flight (2^28 elements) there are only ~3.5 warps per scheduler, so what matters is the length of the
// dependent instruction chain per element, not occupancy.  v3 (profiles/r1_summary.md: v2 spent 83 issued
// instructions per element at 20/32 lane efficiency):
//   - the compressed stream of every lane is staged by cp.async (16 B per lane per copy, 8 copies in flight)
//     into a lane-private shared-memory ring; the bit buffer is refilled 32 bits at a time from a word that was
//     read from the ring one refill earlier -- no global-memory latency and no register queue on the chain
//   - general round (8 elements): branch-free decode of any code -- length, raw/escape reconstruction with the
//     midpoint bit, the three predictors, selection -- and a branch-free history update that is EXACT for the
//     reference's -1.0f "empty slot" state machine once three values are in (only b1 can be -1.0f afterwards)
//   - rounds of eight 3-bit codes take a parse-free path (one mask test, static bit tests, one 24-bit consume)
//   - the first values of a chunk (history slots still empty) run the reference's state machine verbatim
//   - every lane writes its own output as full 32-byte sectors
#pragma once
#include "dcb_common.cuh"

#define DEC_NT 128
#define DEC_RND 8
#define DEC_RING_WORDS 32          // 8 copies of 16 B per lane
#define DEC_PITCH 36               // ring pitch in words (144 B: 16-byte aligned, lanes start in different banks)

__device__ __forceinline__ void dec_cp_async16(uint32_t smem_dst, const void* gsrc) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 16;\n" ::"r"(smem_dst), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void dec_cp_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
__device__ __forceinline__ void dec_cp_wait7() { asm volatile("cp.async.wait_group 7;\n" ::: "memory"); }
__device__ __forceinline__ uint32_t lds32(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.u32 %0, [%1];\n" : "=r"(v) : "r"(addr) : "memory");
  return v;
}

// Per-lane MSB-first bit reader over a shared-memory ring fed by cp.async.
// Invariant between codes: 32 <= avail <= 63, `hi` = the next 32 stream bits.
struct LaneReader {
  uint32_t hi, lo;        // bit window, top aligned
  int avail;              // valid bits in hi:lo
  uint32_t nextw;         // ring word `widx`, already read (memory byte order)
  uint32_t widx;          // running word index (word 0 = first word of the first 16-byte block fetched)
  uint32_t ring;          // shared-memory byte address of this lane's ring
  const uint8_t* gnext;   // next 16-byte block to fetch
  const uint8_t* gend;    // first block at/after the end of this chunk's stream

  __device__ __forceinline__ void fetch_block(uint32_t blk) {        // block index -> ring slot blk & 7
    if (gnext < gend) dec_cp_async16(ring + ((blk & 7u) << 4), gnext);
    gnext += 16;
    dec_cp_commit();                                                  // (an empty group keeps the count uniform)
  }
  // make `nextw` the word after the current one; crossing into a new 16-byte block recycles the block just left
  __device__ __forceinline__ void advance_word() {
    widx++;
    if ((widx & 3u) == 0u) {
      fetch_block((widx >> 2) + 7u);
      dec_cp_wait7();                                                 // all but the 7 newest copies done: block widx/4 landed
    }
    nextw = lds32(ring + ((widx & (DEC_RING_WORDS - 1u)) << 2));
  }
  __device__ __forceinline__ void refill() {
    if (avail < 32) {
      const uint32_t w = __byte_perm(nextw, 0u, 0x0123);              // memory order -> MSB-first word
      hi |= w >> avail;
      lo = __funnelshift_r(0u, w, avail);                            // (w:0) >> avail, low half
      avail += 32;
      advance_word();
    }
  }
  __device__ __forceinline__ void consume(int L) {                    // drop L (0..32) bits, then restore the invariant
    hi = __funnelshift_lc(lo, hi, L);
    lo = __funnelshift_lc(0u, lo, L);
    avail -= L;
    refill();
  }
  __device__ __forceinline__ void init(const uint8_t* stream, uint64_t byte_off, uint64_t byte_end, uint32_t ring_addr) {
    ring = ring_addr;
    gnext = stream + (byte_off & ~15ull);
    gend = stream + ((byte_end + 15ull) & ~15ull);
    #pragma unroll
    for (uint32_t b = 0; b < 8u; b++) fetch_block(b);
    dec_cp_wait7();
    widx = (uint32_t)(byte_off & 15ull) >> 2;
    nextw = lds32(ring + (widx << 2));
    hi = 0u; lo = 0u; avail = 0;
    refill();                                                         // 32 bits
    consume((int)(byte_off & 3ull) * 8);                              // sub-word start (chunks are byte aligned)
  }
};

// raw / escape code starting at the top of `w`: value (with the midpoint bit, :3168-3184) and length
template <int CT>
__device__ __forceinline__ void decode_fallback(uint32_t w, const CodecConst& k, int type, uint32_t ones, uint32_t mask17,
                                                int m_mask, uint32_t& vbits, int& len) {
  if (CT == CT_OP) { len = 32; vbits = w; return; }                 // :718-721
  const int sh = min(max(150 - k.n_eb - (int)(w >> 23), 0), 23);    // leading bit 0: w >> 23 is the exponent field
  len = 32 - sh;
  vbits = (w & (0xFFFFFFFFu << sh)) | ((1u << sh) >> 1);            // top `len` bits, then a single 1, then zeros
  if (CT == CT_MASK) {
    if (((w >> (31 - type)) & ones) == ones) {                      // 0, `type` ones: escape (:1739-1758)
      const uint32_t flag = (w >> (30 - type)) & 1u;
      const uint32_t kb = flag ? (uint32_t)m_mask : (uint32_t)max(m_mask - 8, 0);
      const uint32_t hdr = (uint32_t)type + 2u;
      const uint32_t pay = kb ? ((w << hdr) >> (32u - kb)) : 0u;
      len = (int)(hdr + kb);
      const uint32_t have = (flag ? 9u : 17u) + kb;
      uint32_t top = flag ? (((mask17 >> 8) << 23) | (kb ? pay << (23u - kb) : 0u))   // :1963-2010
                          : ((mask17 << 15) | (kb ? pay << (15u - kb) : 0u));          // :1947-1962
      top |= __funnelshift_rc(0x80000000u, 0u, have);
      vbits = top;
    }
  }
}

// one code with the reference's full history state machine (used while slots are still empty)
template <int CT>
__device__ __forceinline__ float decode_one_exact(LaneReader& br, const CodecConst& k, int type, uint32_t ones,
                                                  uint32_t mask17, int m_mask, float& b3, float& b2, float& b1, bool& bad) {
  const uint32_t w = br.hi;
  float v; int len;
  if ((w >> 31) && CT != CT_NP) {                                   // 3-bit code (:2965-2967)
    len = 3;
    const float p2 = pred2(b1, b2), p3 = pred3(b1, b2, b3);        // decompress_bitwise_float :3139-3158
    const float lo2 = (w & 0x20000000u) ? b1 : 0.0f;               // 100 -> 0, 101 -> b1
    const float hi2 = (w & 0x20000000u) ? p3 : p2;                 // 110 -> 2*b1-b2, 111 -> 3*b1-3*b2+b3
    v = (w & 0x40000000u) ? hi2 : lo2;
  } else {
    if (CT == CT_NP && (w >> 31)) bad = true;                       // :2501-2505 "Error leading bit 1"
    uint32_t vb;
    decode_fallback<CT>(CT == CT_NP ? (w & 0x7FFFFFFFu) : w, k, type, ones, mask17, m_mask, vb, len);
    if (CT == CT_NP) vb |= (w & 0x80000000u);
    v = u2f(vb);
  }
  if (CT != CT_NP) hist_push(b3, b2, b1, v);                        // :3110-3127
  br.consume(len);
  return v;
}

template <int CT>
__global__ void __launch_bounds__(DEC_NT) dcb_decode_kernel(const DecArgs a) {
  __shared__ __align__(16) uint32_t rings[DEC_NT * DEC_PITCH];
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

  LaneReader br;
  br.init(a.stream, a.offsets[c], a.offsets[c + 1],
          (uint32_t)__cvta_generic_to_shared(rings + threadIdx.x * DEC_PITCH));
  float* out = a.out + elem0;
  const bool vec_ok = ((reinterpret_cast<uintptr_t>(out) & 31) == 0);

  float b3 = -1.0f, b2 = -1.0f, b1 = -1.0f;                         // :2930
  bool bad = false;
  uint32_t i = 0;
  // ---- head: the reference's state machine verbatim until two history slots hold values (and a whole
  //      output sector is written), so that afterwards only b1 can ever be the -1.0f "empty" marker
  if (CT != CT_NP) {
    while (i < cn && ((i & (DEC_RND - 1)) != 0u || i == 0u || b3 == -1.0f || b2 == -1.0f)) {
      const float x = decode_one_exact<CT>(br, a.k, type, ones, mask17, m_mask, b3, b2, b1, bad);
      out[i] = addmin ? __fadd_rn(x, mn) : x;
      i++;
    }
  }
  for (; i + DEC_RND <= cn; i += DEC_RND) {
    float v[DEC_RND];
    bool done = false;
    if (CT != CT_NP) {
      // fast round: the next 8 codes all start with 1 (3-bit codes) and b1 holds a value
      if (b1 != -1.0f && (br.hi & 0x92492400u) == 0x92492400u) {
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
        if (!(neg >> 31)) {                  // no negative value -> none is the -1.0f marker: commit
          b1 = c1; b2 = c2; b3 = c3;
          br.consume(3 * DEC_RND);
          done = true;
        }
      }
    }
    if (!done) {
      // general round, branch free.  History update == hist_push (:3110-3127) given b3, b2 != -1.0f:
      // a -1.0f in b1 is overwritten instead of shifted out.
      #pragma unroll
      for (int e = 0; e < DEC_RND; e++) {
        const uint32_t w = br.hi;
        uint32_t vb; int len;
        decode_fallback<CT>(w, a.k, type, ones, mask17, m_mask, vb, len);
        float x = u2f(vb);
        if (CT == CT_NP) {
          bad |= (w >> 31) != 0u;                                     // :2501-2505 "Error leading bit 1"
        } else {
          const bool is3 = (w >> 31) != 0u;
          const float p2 = __fsub_rn(__fadd_rn(b1, b1), b2);
          const float p3 = __fadd_rn(__fsub_rn(__fmul_rn(3.0f, b1), __fmul_rn(3.0f, b2)), b3);
          const bool bb = (w & 0x20000000u) != 0u, ba = (w & 0x40000000u) != 0u;
          const float x3 = ba ? (bb ? p3 : p2) : (bb ? b1 : 0.0f);    // decompress_bitwise_float :3139-3158
          x = is3 ? x3 : x;
          len = is3 ? 3 : len;
          const bool shift_ok = (b1 != -1.0f);
          b3 = shift_ok ? b2 : b3;
          b2 = shift_ok ? b1 : b2;
          b1 = x;
        }
        v[e] = x;
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
    const float x = decode_one_exact<CT>(br, a.k, type, ones, mask17, m_mask, b3, b2, b1, bad);
    out[i] = addmin ? __fadd_rn(x, mn) : x;
  }
  if (bad && a.status) atomicExch(a.status, -5);
}
