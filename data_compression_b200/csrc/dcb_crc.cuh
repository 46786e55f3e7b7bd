// dcb_crc.cuh -- do_crc32 (dataCompression.c:5524-5534: zlib crc32, CRC-32/ISO-HDLC, reflected polynomial
// 0xEDB88320, init/xorout 0xFFFFFFFF) of every chunk's stream.  One warp per chunk: each lane runs the
// byte-wise table recurrence over its own contiguous segment starting from state 0 (lane 0: 0xFFFFFFFF),
// then the 32 partial states are aligned with x^(8*bytes_after) mod P (GF(2) multiply, square-and-multiply
// over precomputed x^(8*2^k)) and XOR-reduced -- CRC is linear in its state.
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

__global__ void __launch_bounds__(128) dcb_crc_kernel(const uint8_t* __restrict__ stream,
                                                      const uint64_t* __restrict__ offsets, uint64_t num_chunks,
                                                      uint32_t* __restrict__ crc_out) {
  __shared__ uint32_t tab[256];
  __shared__ uint32_t x8pow[48];
  for (int i = threadIdx.x; i < 256; i += blockDim.x) {
    uint32_t c = (uint32_t)i;
    #pragma unroll
    for (int k = 0; k < 8; k++) c = (c & 1u) ? (0xEDB88320u ^ (c >> 1)) : (c >> 1);
    tab[i] = c;
  }
  if (threadIdx.x == 0) {
    uint32_t p = 0x00800000u;                    // x^8 in the reflected representation (x^0 = bit 31)
    for (int k = 0; k < 48; k++) { x8pow[k] = p; p = gf2_mulmod(p, p); }
  }
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const uint64_t c = (uint64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (c >= num_chunks) return;
  const uint64_t lo = offsets[c], hi = offsets[c + 1], B = hi - lo;
  const uint64_t seg = (B + 31ull) / 32ull;
  const uint64_t s0 = min(B, seg * lane), s1 = min(B, s0 + seg);
  uint32_t st = (lane == 0) ? 0xFFFFFFFFu : 0u;
  for (uint64_t i = s0; i < s1; i++) st = tab[(st ^ stream[lo + i]) & 0xFFu] ^ (st >> 8);
  st = crc_shift(st, B - s1, x8pow);
  #pragma unroll
  for (int o = 16; o; o >>= 1) st ^= __shfl_xor_sync(0xffffffffu, st, o);
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
