// dcb_decode.cuh -- decoder for CT=5/6/7/11 (replaces myDecompress_bitwise{,_np,_mask,_op},
// dataCompression.c:2922, :2459, :1703, :698 and decompress_bitwise_float{,_np,_mask} :3137, :2611, :1900).
//
// Decoding has two serial dependencies per chunk: code boundaries (a code's length is known
// only after its first bit and exponent are read) and the float recurrence on predicted values (the decoder
// predicts from RECONSTRUCTED values, and float rounding makes the recurrence non-associative).  Chunks are
// independent, so ONE LANE walks ONE CHUNK and the parallelism is across the ~n/4096 chunks:
//   - a warp owns 32 consecutive chunks and advances them in rounds of 32 elements per lane
//   - per round the warp cooperatively stages, for every lane, the <=136 stream bytes that lane can consume
//     (coalesced 128-byte row loads, byte-swapped to MSB-first words, odd row stride -> no bank conflicts)
//   - each lane parses its codes from a 64-bit bit buffer and runs the history state machine in registers
//   - decoded values go to a [32][33] staging tile and leave as 32 coalesced 128-byte rows (+min fused)
#pragma once
#include "dcb_common.cuh"

#define DEC_WARPS 4
#define DEC_R 32                 // elements per lane per round
#define DEC_IW 36                // staged stream words per lane per round (32 max consumed + alignment + look-ahead)
#define DEC_IS 37                // row stride (odd)

struct DecSmem {
  uint32_t in[DEC_WARPS][32][DEC_IS];
  float out[DEC_WARPS][32][DEC_R + 1];
};

template <int CT>
__global__ void __launch_bounds__(DEC_WARPS * 32) dcb_decode_kernel(const DecArgs a) {
  __shared__ DecSmem sm;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint64_t c = ((uint64_t)blockIdx.x * DEC_WARPS + warp) * 32ull + lane;
  const bool live = c < a.num_chunks;
  const uint64_t c0 = c - lane;                                  // first chunk of the warp
  if (c0 >= a.num_chunks) return;

  // per-lane chunk description
  uint64_t bit0 = 0, bit_end = 0;                                // absolute bit range of the chunk's stream
  uint32_t cn = 0;                                               // elements in this chunk
  uint64_t elem0 = 0;
  float mn = 0.0f;
  int type = a.fixed_type;
  uint32_t mask17 = a.fixed_mask17;
  if (live) {
    bit0 = a.offsets[c] * 8ull;
    bit_end = a.offsets[c + 1] * 8ull;
    elem0 = c * a.chunk;
    cn = (uint32_t)min(a.chunk, a.n - elem0);
    if ((a.flags & 1u) && a.cmin) mn = a.cmin[c];
    if (CT == CT_MASK && !(a.flags & 2u)) { type = a.ctype[c]; mask17 = f2u(a.cmed[c]) >> 15; }
  }
  type = min(max(type, 0), 7);
  // CT=7 escape constants: m comes from the MASK's exponent (dataCompression.c:1769-1771)
  const int m_mask = raw_len(mask17 << 15, a.k.n_eb_m118) - 9;
  const uint32_t ones = (1u << type) - 1u;

  float b3 = -1.0f, b2 = -1.0f, b1 = -1.0f;                      // :2930
  uint64_t bp = bit0;                                            // absolute bit position of the next code
  uint32_t done = 0;
  const uint32_t cn_max = __reduce_max_sync(0xffffffffu, cn);
  const uint32_t* s32 = reinterpret_cast<const uint32_t*>(a.stream);
  bool bad = false;

  for (uint32_t r0 = 0; r0 < cn_max; r0 += DEC_R) {
    // ---- stage every lane's window: words [bp>>5, bp>>5 + DEC_IW) ----------------------------------
    #pragma unroll 4
    for (int l = 0; l < 32; l++) {
      const uint64_t w0 = __shfl_sync(0xffffffffu, bp, l) >> 5;
      const uint64_t bend = __shfl_sync(0xffffffffu, bit_end, l);    // bits past the chunk's last byte read as 0
      #pragma unroll
      for (int h = 0; h < 2; h++) {
        if (h == 0 || lane < DEC_IW - 32) {
          const uint64_t w = w0 + lane + 32 * h;
          uint32_t v = 0u;
          if (w * 32ull < bend) {
            v = __byte_perm(__ldg(s32 + w), 0u, 0x0123);
            const uint64_t rem = bend - w * 32ull;
            if (rem < 32ull) v &= ~(0xFFFFFFFFu >> (uint32_t)rem);
          }
          sm.in[warp][l][lane + 32 * h] = v;
        }
      }
    }
    __syncwarp();

    // ---- decode up to DEC_R codes of this lane's chunk ------------------------------------------------
    const uint32_t* row = sm.in[warp][lane];
    const uint32_t boff = (uint32_t)(bp & 31ull);
    unsigned long long buf = (((unsigned long long)row[0] << 32) | row[1]) << boff;
    int avail = 64 - (int)boff;
    uint32_t widx = 2;
    if (avail < 32) { buf |= (unsigned long long)row[widx++] << (32 - avail); avail += 32; }
    uint32_t used = 0;
    const uint32_t todo = (done < cn) ? min((uint32_t)DEC_R, cn - done) : 0u;
    #pragma unroll 1
    for (uint32_t e = 0; e < DEC_R; e++) {
      if (e < todo) {
        const uint32_t w = (uint32_t)(buf >> 32);
        float v;
        uint32_t len;
        if ((w >> 31) && CT != CT_NP) {                              // 3-bit code (:2965-2967)
          len = 3u;
          const uint32_t sel = (w >> 29) & 3u;                      // decompress_bitwise_float :3139-3158
          const float p2 = pred2(b1, b2), p3 = pred3(b1, b2, b3);
          v = (sel == 0u) ? 0.0f : ((sel == 1u) ? b1 : ((sel == 2u) ? p2 : p3));
        } else if (CT == CT_OP) {                                    // :718-721
          len = 32u; v = u2f(w);
        } else {
          if (CT == CT_NP && (w >> 31)) bad = true;                  // :2501-2505 "Error leading bit 1"
          bool escape = false;
          if (CT == CT_MASK) escape = (((w >> (31 - type)) & ones) == ones);   // :1739-1758
          if (!escape) {                                             // raw: 9 bits, then m from the exponent read
            len = (uint32_t)raw_len(w, a.k.n_eb_m118);
            const uint32_t top = (len == 32u) ? w : (w & ~(0xFFFFFFFFu >> len));
            v = u2f((len == 32u) ? top : (top | (0x80000000u >> len)));        // midpoint, :3168-3184
          } else {
            const uint32_t flag = (w >> (30 - type)) & 1u;
            const uint32_t kb = flag ? (uint32_t)m_mask : (uint32_t)max(m_mask - 8, 0);
            const uint32_t hdr = (uint32_t)type + 2u;
            const uint32_t pay = kb ? ((w << hdr) >> (32u - kb)) : 0u;
            len = hdr + kb;
            const uint32_t have = (flag ? 9u : 17u) + kb;
            uint32_t top = flag ? (((mask17 >> 8) << 23) | (kb ? pay << (23u - kb) : 0u))   // :1963-2010
                                : ((mask17 << 15) | (kb ? pay << (15u - kb) : 0u));          // :1947-1962
            if (have < 32u) top |= (0x80000000u >> have);
            v = u2f(top);
          }
        }
        sm.out[warp][lane][e] = v;
        if (CT != CT_NP) hist_push(b3, b2, b1, v);                   // :3110-3127
        buf <<= len;                                                 // len <= 32 < 64
        avail -= (int)len;
        used += len;
        if (avail < 32) { buf |= (unsigned long long)row[widx++] << (32 - avail); avail += 32; }
      }
    }
    bp += used;
    done += todo;
    __syncwarp();

    // ---- write out: lane t takes element t of every lane's row --------------------------------------------
    #pragma unroll 4
    for (int l = 0; l < 32; l++) {
      const uint32_t cn_l = __shfl_sync(0xffffffffu, cn, l);
      const uint64_t e0_l = __shfl_sync(0xffffffffu, elem0, l);
      const float mn_l = __shfl_sync(0xffffffffu, mn, l);
      const uint32_t idx = r0 + lane;
      if (idx < cn_l) {
        float v = sm.out[warp][l][lane];
        if (a.flags & 1u) v = __fadd_rn(v, mn_l);                    // decompressed[i] + min, impl/pingpong.c:562
        a.out[e0_l + idx] = v;
      }
    }
    __syncwarp();
  }
  if (bad && a.status) atomicExch(a.status, -5);
}
