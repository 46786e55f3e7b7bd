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
  for (uint32_t i = (uint32_t)tid; i < head + (nbytes - tail0); i += DCB_HM_NT) {
    const uint32_t q = (i < head) ? i : (tail0 + (i - head));
    uint32_t v = (uint32_t)p[q] << 24;                   // data bit 8 q + t is bit 7 - t of byte q (cast_bits_to_char, :5712-5722)
    while (v) { const int t = __clz((int)v); v &= ~(0x80000000u >> t); sx ^= hm_pos(8u * q + (uint32_t)t); n1++; }
  }
  const uint32_t* pw = reinterpret_cast<const uint32_t*>(p + head);
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
