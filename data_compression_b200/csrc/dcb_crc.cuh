// dcb_crc.cuh -- do_crc32 (dataCompression.c:5524-5534: zlib crc32, CRC-32/ISO-HDLC, reflected polynomial
// 0xEDB88320, init/xorout 0xFFFFFFFF) of every chunk's stream.  One warp per chunk, coalesced word loads, a
// strided Horner recurrence per lane and a GF(2) recombination -- CRC is linear in its state (details at the kernel).
#pragma once
#include "dcb_common.cuh"

__device__ __forceinline__ uint32_t gf2_mulmod(uint32_t a, uint32_t b) {   // reflected representation, mod 0x1EDB88320
  uint32_t p = 0;
  #pragma unroll 4
  for (int i = 0; i < 32; i++) {
    if (a & 0x80000000u) p ^= b;                 // bit 31 of a is x^0
    a <<= 1;
    b = (b >> 1) ^ ((b & 1u) ? 0xEDB88320u : 0u);
  }
  return p;
}

// state after `nbytes` zero bytes == state * x^(8*nbytes) mod P
__device__ __forceinline__ uint32_t crc_shift(uint32_t state, uint64_t nbytes, const uint32_t* x8pow) {
  for (int k = 0; nbytes; k++, nbytes >>= 1)
    if (nbytes & 1ull) state = gf2_mulmod(state, x8pow[k]);
  return state;
}

// Tables, built once per device by dcb_crc_init_kernel:
//   g_crc_tab[i]    byte-wise table of the reflected polynomial (head / last bytes)
//   g_crc_ts[k][i]  (i << 8k) * x^1024 mod P: one Horner step of a lane that owns every 32nd word
//   g_crc_c[j]      x^(32 j) mod P, j = 0..32: weight of a word that is followed by j - 1 more words, times x^32
__device__ uint32_t g_crc_tab[256];
__device__ uint32_t g_crc_ts[4][256];
__device__ uint32_t g_crc_c[33];
//   g_crc_t4[k][i]  (i << 8k) * x^4096 mod P: one Horner step of an accumulator that owns every 128th word (fused CRC of
//                   dcb_compact_kernel: a lane reads 16-byte units, i.e. four accumulators of stride 128 words)
//   g_crc_c2[j]     x^(32 j) mod P, j = 0..256
__device__ uint32_t g_crc_t4[4][256];
__device__ uint32_t g_crc_c2[257];

__global__ void dcb_crc_init_kernel() {
  __shared__ uint32_t x1024, x4096;
  const int t = threadIdx.x;                       // 256 threads
  if (t == 0) {
    uint32_t p = 0x00800000u;                      // x^8 (x^0 = bit 31)
    for (int k = 0; k < 7; k++) p = gf2_mulmod(p, p);          // x^(8 * 128) = x^1024
    x1024 = p;
    uint32_t x32 = 0x00800000u;
    for (int k = 0; k < 2; k++) x32 = gf2_mulmod(x32, x32);    // x^32
    uint32_t c = 0x80000000u;                      // x^0
    for (int j = 0; j <= 256; j++) { if (j <= 32) g_crc_c[j] = c; g_crc_c2[j] = c; c = gf2_mulmod(c, x32); }
    x4096 = gf2_mulmod(gf2_mulmod(x1024, x1024), gf2_mulmod(x1024, x1024));
  }
  __syncthreads();
  uint32_t c = (uint32_t)t;
  #pragma unroll
  for (int k = 0; k < 8; k++) c = (c & 1u) ? (0xEDB88320u ^ (c >> 1)) : (c >> 1);
  g_crc_tab[t] = c;
  for (int k = 0; k < 4; k++) g_crc_ts[k][t] = gf2_mulmod((uint32_t)t << (8 * k), x1024);
  for (int k = 0; k < 4; k++) g_crc_t4[k][t] = gf2_mulmod((uint32_t)t << (8 * k), x4096);
}

// One warp per chunk.  The chunk's 32-bit words are read row by row (32 consecutive words = one coalesced 128-byte load);
// lane L keeps U_L = sum_j w[L + 32 j] * x^(1024 (R-1-j)) by Horner: U <- U * x^1024 + w, four table lookups per word and
// nothing else on the dependent chain.  The state after R rows is sum_L U_L * x^(32 (32 - L)) (one bit-serial multiply per
// lane and chunk, XOR reduction), a last partial row uses the same weights shifted, and the <= 3 bytes before the first
// aligned word / after the last whole word go through the byte table.  CRC is linear: the state that enters a row is
// simply XORed into its first word.
__global__ void __launch_bounds__(128) dcb_crc_kernel(const uint8_t* __restrict__ stream,
                                                      const uint64_t* __restrict__ offsets, uint64_t num_chunks,
                                                      uint32_t* __restrict__ crc_out) {
  __shared__ uint32_t tab[256];
  __shared__ uint32_t ts[4][256];
  __shared__ uint32_t cw[33];
  for (int i = threadIdx.x; i < 256; i += blockDim.x) {
    tab[i] = g_crc_tab[i];
    #pragma unroll
    for (int k = 0; k < 4; k++) ts[k][i] = g_crc_ts[k][i];
  }
  if (threadIdx.x < 33) cw[threadIdx.x] = g_crc_c[threadIdx.x];
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const uint64_t c = (uint64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (c >= num_chunks) return;
  const uint64_t lo = offsets[c], hi = offsets[c + 1];
  const uint8_t* p = stream + lo;
  uint64_t B = hi - lo;
  uint32_t st = 0xFFFFFFFFu;                        // warp uniform
  uint32_t head = (uint32_t)((4ull - (reinterpret_cast<uintptr_t>(p) & 3ull)) & 3ull);   // bytes before the first aligned word
  if (head > B) head = (uint32_t)B;
  for (uint32_t i = 0; i < head; i++) st = tab[(st ^ p[i]) & 0xFFu] ^ (st >> 8);
  p += head; B -= head;
  const uint32_t* w32 = reinterpret_cast<const uint32_t*>(p);
  const uint64_t nwords = B >> 2;
  const uint64_t rows = nwords >> 5;
  if (rows) {
    uint32_t U = 0u;
    uint64_t r = 0;
    for (; r + 4 <= rows; r += 4) {                 // four loads in flight
      uint32_t w0 = __ldg(w32 + (r << 5) + lane), w1 = __ldg(w32 + ((r + 1) << 5) + lane);
      uint32_t w2 = __ldg(w32 + ((r + 2) << 5) + lane), w3 = __ldg(w32 + ((r + 3) << 5) + lane);
      if (r == 0 && lane == 0) w0 ^= st;
      U = ts[0][U & 0xFFu] ^ ts[1][(U >> 8) & 0xFFu] ^ ts[2][(U >> 16) & 0xFFu] ^ ts[3][U >> 24] ^ w0;
      U = ts[0][U & 0xFFu] ^ ts[1][(U >> 8) & 0xFFu] ^ ts[2][(U >> 16) & 0xFFu] ^ ts[3][U >> 24] ^ w1;
      U = ts[0][U & 0xFFu] ^ ts[1][(U >> 8) & 0xFFu] ^ ts[2][(U >> 16) & 0xFFu] ^ ts[3][U >> 24] ^ w2;
      U = ts[0][U & 0xFFu] ^ ts[1][(U >> 8) & 0xFFu] ^ ts[2][(U >> 16) & 0xFFu] ^ ts[3][U >> 24] ^ w3;
    }
    for (; r < rows; r++) {
      uint32_t w0 = __ldg(w32 + (r << 5) + lane);
      if (r == 0 && lane == 0) w0 ^= st;
      U = ts[0][U & 0xFFu] ^ ts[1][(U >> 8) & 0xFFu] ^ ts[2][(U >> 16) & 0xFFu] ^ ts[3][U >> 24] ^ w0;
    }
    st = gf2_mulmod(U, cw[32 - lane]);
    #pragma unroll
    for (int o = 16; o; o >>= 1) st ^= __shfl_xor_sync(0xffffffffu, st, o);
  }
  const uint32_t nt = (uint32_t)(nwords & 31ull);   // words of the last partial row
  if (nt) {
    uint32_t v = 0u;
    if ((uint32_t)lane < nt) {
      uint32_t w0 = __ldg(w32 + (rows << 5) + lane);
      if (lane == 0) w0 ^= st;
      v = gf2_mulmod(w0, cw[nt - lane]);
    }
    #pragma unroll
    for (int o = 16; o; o >>= 1) v ^= __shfl_xor_sync(0xffffffffu, v, o);
    st = v;
  }
  const uint8_t* q = p + (nwords << 2);
  for (uint32_t i = 0; i < (uint32_t)(B & 3ull); i++) st = tab[(st ^ q[i]) & 0xFFu] ^ (st >> 8);
  if (lane == 0) crc_out[c] = st ^ 0xFFFFFFFFu;
}

// bit_flip test hook (dataCompression.c:5858-5865 flips random bits; here the caller names them)
__global__ void dcb_flip_kernel(uint8_t* stream, const uint64_t* idx, uint64_t count) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < count) {
    const uint64_t b = idx[i];
    uint32_t* w = reinterpret_cast<uint32_t*>(stream + ((b >> 3) & ~3ull));
    atomicXor(w, (0x80u >> (b & 7ull)) << (8u * ((b >> 3) & 3ull)));
  }
}
