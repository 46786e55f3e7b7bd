// dcb_hamming.cuh -- Hamming SECDED over blocks of the finished stream (CT=10).
// Replaces hamming_encode / hamming_decode (impl/dataCompression.c:5742-5779; hamming_code :5545, hamming_verify_bit :5782,
// error_info :5632, hamming_rectify_bit :5822), called once per block by impl/pingpong.c:296-311 and :487-505.
//
// The reference evaluates each of the r check bits with a loop over all r + k codeword positions.  Check bit i is the
// parity of the data bits whose (1-based) codeword position has bit i set, so all r of them together are the XOR of the
// codeword positions of the set data bits: one pass, a popcount and an XOR reduction.  One thread block per Hamming
// block (125,000 bytes at the reference's BER = 1e-6), 32-bit words read coalesced, set bits visited with CLZ.
#pragma once
#include "dcb_common.cuh"

#define DCB_HM_NT 256

// codeword position (1-based; powers of two hold check bits, :5551-5554) of data bit d (0-based)
__device__ __forceinline__ uint32_t hm_pos(uint32_t d) {
  const uint32_t m = d + 1u;
  const uint32_t t = 32u - (uint32_t)__clz((int)m);      // floor(log2 m) + 1
  uint32_t j = m + t;
  if (j >> t) j++;                                       // one more check position (2^t) lies below
  return j;
}
__host__ __device__ __forceinline__ int hm_length(long long k) {   // hmLength, :5587-5599
  int r = 0;
  while (((1LL << r) - 1 - r - k) < 0) r++;
  return r;
}

struct HmArgs {
  uint8_t* bits;                // the stream (decode rectifies in place)
  uint64_t total_bytes;
  uint32_t block_bytes;
  uint32_t* check;              // [blocks]  bit i (i < r) = c[i], bit r = c[r] (parity over data and check bits)
  int32_t* r_out;               // [blocks]  encode: r of every block (may be null)
  int32_t* error_type;          // [blocks]  decode: 0 none, 1 two-bit, 2 parity bit, 3 one bit (rectified)
  int decode;
};

__global__ void __launch_bounds__(DCB_HM_NT) dcb_hamming_kernel(const HmArgs a) {
  __shared__ uint32_t s_x[DCB_HM_NT / 32], s_n[DCB_HM_NT / 32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint64_t b0 = (uint64_t)blockIdx.x * a.block_bytes;
  const uint32_t nbytes = (uint32_t)min((uint64_t)a.block_bytes, a.total_bytes - b0);
  const uint8_t* p = a.bits + b0;
  uint32_t sx = 0u, n1 = 0u;
  // leading bytes up to a 4-byte boundary, whole words, trailing bytes
  const uint32_t head = min(nbytes, (uint32_t)((4u - (uint32_t)(reinterpret_cast<uintptr_t>(p) & 3u)) & 3u));
  const uint32_t nwords = (nbytes - head) >> 2;
  const uint32_t tail0 = head + (nwords << 2);
  const bool wordpar = head == 0u && nbytes >= 64u && nbytes < (1u << 26);   // (shift t + 2 stays below 32)                          // the word-parallel pass below covers every byte
  for (uint32_t i = (uint32_t)tid; !wordpar && i < head + (nbytes - tail0); i += DCB_HM_NT) {
    const uint32_t q = (i < head) ? i : (tail0 + (i - head));
    uint32_t v = (uint32_t)p[q] << 24;                   // data bit 8 q + t is bit 7 - t of byte q (cast_bits_to_char, :5712-5722)
    while (v) { const int t = __clz((int)v); v &= ~(0x80000000u >> t); sx ^= hm_pos(8u * q + (uint32_t)t); n1++; }
  }
  const uint32_t* pw = reinterpret_cast<const uint32_t*>(p + head);
  if (wordpar) {
    // Word-parallel pass.  Codeword word q = positions 32q .. 32q + 31; for q >= 2 they lie in one region between check
    // positions, 2^t <= 32q < 2^(t+1), where position = data bit + t + 2: the word is the data bit string shifted by
    // sh = t + 2 < 32 bits, one funnel shift of data words q - 1 and q (bit j of a data word = data bit 32w + j, i.e. the
    // stream's bytes bit-reversed).  The XOR of the positions of its set bits splits into (bit index within the word) -- the
    // XOR over all words of the words themselves, A, reduced to five parities at the very end -- and q for every word of odd
    // parity.  Position 2^t itself (bit 0 of word q = 2^(t-5)) holds a check bit, not data: those <= 27 words are corrected
    // afterwards.  Positions below 64 (data bits 0 .. 56) are visited bit by bit.
    const uint32_t nbits = 8u * nbytes;
    const uint32_t rr = (uint32_t)hm_length((long long)nbits);
    const uint32_t Q = ((nbits + rr) >> 5) + 1u;                            // words 0 .. Q-1 hold every position
    const uint32_t rem = nbytes & 3u;
    uint32_t lastw = 0u;                                                    // the ragged last data word, zero filled
    for (uint32_t k = 0; k < rem; k++) lastw |= (uint32_t)p[tail0 + k] << (8u * k);
    auto load_u = [&](uint32_t w) -> uint32_t {
      uint32_t x = __ldg(pw + min(w, nwords - 1u));                          // branch free: the loads of a step issue together
      x = (w < nwords) ? x : ((w == nwords) ? lastw : 0u);
      return __brev(__byte_perm(x, 0u, 0x0123));
    };
    auto cw_of = [&](uint32_t q, uint32_t ulo, uint32_t uhi) -> uint32_t {   // ulo = data word q-1, uhi = data word q
      const uint32_t sh = (uint32_t)(31 - __clz((int)q)) + 7u;               // t + 2, t = floor(log2 q) + 5
      return __funnelshift_r(ulo, uhi, 32u - sh);
    };
    uint32_t A = 0u, hx = 0u;
    // a warp takes 128 consecutive words per step, four loads per lane in flight (one load per step left the pass latency
    // bound: 1.6 TB/s)
    for (uint32_t q0 = 2u + 128u * (uint32_t)warp; q0 < Q; q0 += 4u * DCB_HM_NT) {
      uint32_t uhi[4], ul0 = 0u;
      #pragma unroll
      for (int k = 0; k < 4; k++) {
        const uint32_t q = q0 + 32u * (uint32_t)k + (uint32_t)lane;
        uhi[k] = load_u(q);                                                  // 0 beyond the data
      }
      if (lane == 0) ul0 = load_u(q0 - 1u);
      #pragma unroll
      for (int k = 0; k < 4; k++) {
        const uint32_t q = q0 + 32u * (uint32_t)k + (uint32_t)lane;
        uint32_t ulo = __shfl_up_sync(0xffffffffu, uhi[k], 1);
        const uint32_t prev = __shfl_sync(0xffffffffu, uhi[(k + 3) & 3], 31);   // lane 0: the last word of the 32 before
        if (lane == 0) ulo = k ? prev : ul0;
        if (q < Q) {
          const uint32_t cw = cw_of(q, ulo, uhi[k]);
          const uint32_t pc = (uint32_t)__popc(cw);
          A ^= cw; n1 += pc;
          hx ^= q & (0u - (pc & 1u));
        }
      }
    }
    if (tid < 27) {                                                          // take the check positions 2^t (t = 6 ..) out again
      const uint32_t q = 1u << (tid + 1);
      if (q >= 2u && q < Q && (cw_of(q, load_u(q - 1u), load_u(q)) & 1u)) { hx ^= q; n1 -= 1u; }
    }
    if (tid == 0) {
      for (uint32_t d = 0; d < nbits; d++) {
        const uint32_t ps = hm_pos(d);
        if (ps >= 64u) break;
        if ((p[d >> 3] >> (7u - (d & 7u))) & 1u) { sx ^= ps; n1++; }
      }
    }
    #pragma unroll
    for (int o = 16; o; o >>= 1) { A ^= __shfl_xor_sync(0xffffffffu, A, o); hx ^= __shfl_xor_sync(0xffffffffu, hx, o); }
    if (lane == 0) {
      const uint32_t lo5 = ((uint32_t)__popc(A & 0xAAAAAAAAu) & 1u) | (((uint32_t)__popc(A & 0xCCCCCCCCu) & 1u) << 1) |
                           (((uint32_t)__popc(A & 0xF0F0F0F0u) & 1u) << 2) | (((uint32_t)__popc(A & 0xFF00FF00u) & 1u) << 3) |
                           (((uint32_t)__popc(A & 0xFFFF0000u) & 1u) << 4);
      sx ^= (hx << 5) | lo5;
    }
  } else
  for (uint32_t w = (uint32_t)tid; w < nwords; w += DCB_HM_NT) {
    uint32_t v = __byte_perm(pw[w], 0u, 0x0123);         // MSB = the first data bit of the word
    const uint32_t d0 = 8u * (head + 4u * w);
    n1 += (uint32_t)__popc(v);
    while (v) { const int t = __clz((int)v); v &= ~(0x80000000u >> t); sx ^= hm_pos(d0 + (uint32_t)t); }
  }
  #pragma unroll
  for (int o = 16; o; o >>= 1) { sx ^= __shfl_xor_sync(0xffffffffu, sx, o); n1 += __shfl_xor_sync(0xffffffffu, n1, o); }
  if (lane == 0) { s_x[warp] = sx; s_n[warp] = n1; }
  __syncthreads();
  if (tid != 0) return;
  sx = 0u; n1 = 0u;
  for (int w = 0; w < DCB_HM_NT / 32; w++) { sx ^= s_x[w]; n1 += s_n[w]; }
  const int r = hm_length(8ll * nbytes);
  if (!a.decode) {
    const uint32_t par = (n1 + (uint32_t)__popc(sx)) & 1u;                 // :5568-5578
    a.check[blockIdx.x] = sx | (par << r);
    if (a.r_out) a.r_out[blockIdx.x] = r;
    return;
  }
  uint32_t c = a.check[blockIdx.x];
  const uint32_t cbits = c & ((1u << r) - 1u);
  const uint32_t pos = sx ^ cbits;                                          // error_info: sum v[i] 2^i
  const bool vr = ((n1 + (uint32_t)__popc(cbits)) & 1u) != ((c >> r) & 1u);
  int et = 0;
  if (pos > 0u && !vr) et = 1;
  else if (pos == 0u && vr) { et = 2; c ^= 1u << r; }
  else if (pos > 0u && vr) {
    et = 3;
    if ((pos & (pos - 1u)) == 0u) {                                         // a check bit
      const int cn = 31 - __clz((int)pos);
      if (cn < r) c ^= 1u << cn;
    } else {
      const uint32_t t = 32u - (uint32_t)__clz((int)pos);                   // powers of two below pos
      const uint32_t d = pos - 1u - t;
      if (d < 8u * nbytes) a.bits[b0 + d / 8u] ^= (uint8_t)(1u << (7u - d % 8u));
    }
  }
  a.check[blockIdx.x] = c;
  if (a.error_type) a.error_type[blockIdx.x] = et;
}
