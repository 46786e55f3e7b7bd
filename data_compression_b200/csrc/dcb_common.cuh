// dcb_common.cuh -- shared definitions for the sm_100a codec kernels.
//
// Reference being replaced: /root/reference/impl/dataCompression.c (float32 bit-wise codecs).
// Bit-exact specification: SURVEY.md section 3.5.  This is NOT a translation of the reference's
// bit-at-a-time realloc loop: codes are computed per element in registers, placed with a block scan,
// packed through shared memory and written contiguously at an offset found by decoupled look-back.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define DCB_NT 256            // threads per encoder block
#define DCB_E 16              // consecutive elements owned by one thread
#define DCB_TILE (DCB_NT * DCB_E)   // 4096 elements per tile

enum { CT_BITWISE = 5, CT_NP = 6, CT_MASK = 7, CT_OP = 11 };

// look-back status word (one u64 per tile):  [63:62] state  [61] tile ends a chunk  [55:0] bits
#define ST_INVALID 0ull
#define ST_AGG 1ull
#define ST_INCL 2ull
#define ST_SHIFT 62
#define ST_END_BIT (1ull << 61)
#define ST_VAL_MASK ((1ull << 56) - 1)

struct CodecConst {
  int n_eb_m118;      // n_eb - 118:  len_raw = clamp(E + n_eb - 118, 9, 32) = 9 + clamp(n_eb + E - 127, 0, 23)
  float tz;           // largest float f with (double)f <  eb   (|x| <  eb  <=> |x| <= tz), dataCompression.c:3390
  float tp;           // largest float f with (double)f <= eb   (d   <= eb  <=> d   <= tp), dataCompression.c:3399
  int n_eb;
};

struct EncArgs {
  const float* in;
  uint64_t n;
  uint64_t chunk;            // elements per chunk (multiple of DCB_TILE) or >= n for a single chunk
  uint32_t tiles_per_chunk;  // chunk / DCB_TILE, or 0 = one chunk spans every tile
  uint32_t num_tiles;
  uint32_t flags;            // DCB_F_*
  int fixed_type;
  uint32_t fixed_mask17;
  CodecConst k;
  uint8_t* out;
  uint64_t* offsets;         // [K+1]
  float* cmin;               // [K]
  float* cmed;               // [K]
  int32_t* ctype;            // [K]
  unsigned long long* status;  // [num_tiles]
  unsigned int* ticket;
  unsigned int* scanner_claim;    // 0 -> 1 by the first block that runs: it becomes the prefix scanner
  unsigned long long* last_bits;  // bit length of the whole stream before the final byte round-up (legacy `pos`)
  // slot path (dcb_encode_slots.cuh): one chunk == one tile, blocks are independent
  uint8_t* slots;                 // [K] scratch slots of DCB_SLOT_PITCH bytes
  uint32_t* sizes;                // [K] bytes of every chunk
  uint32_t* gsum;                 // [ceil(K / DCB_GROUP)] bytes of every group of chunks (zeroed before the launch)
  uint32_t* last_tile_bits;       // code bits of the last chunk
};

struct DecArgs {
  const uint8_t* stream;
  const uint64_t* offsets;
  uint64_t n;
  uint64_t chunk;
  uint64_t num_chunks;
  uint32_t flags;
  int fixed_type;
  uint32_t fixed_mask17;
  CodecConst k;
  const float* cmin;
  const float* cmed;
  const int32_t* ctype;
  float* out;
  int32_t* status;
};

__device__ __forceinline__ uint32_t f2u(float f) { return __float_as_uint(f); }
__device__ __forceinline__ float u2f(uint32_t u) { return __uint_as_float(u); }

// 9 + m(x), m = clamp(n_eb + E - 127, 0, 23)   (compress_bitwise_float, dataCompression.c:3485-3505)
__device__ __forceinline__ int raw_len(uint32_t u, int n_eb_m118) {
  int l = (int)((u >> 23) & 0xFFu) + n_eb_m118;
  return min(max(l, 9), 32);
}

// history update shared by encoder (:3347-3358, :3386-3388) and decoder (:3110-3127); -1.0f marks an empty slot
__device__ __forceinline__ void hist_push(float& b3, float& b2, float& b1, float v) {
  if (b3 == -1.0f) b3 = v;
  else if (b2 == -1.0f) b2 = v;
  else if (b1 == -1.0f) b1 = v;
  else { b3 = b2; b2 = b1; b1 = v; }
}

// predictors, each operation rounded to float32, never contracted (:3362-3364)
__device__ __forceinline__ float pred2(float b1, float b2) { return __fsub_rn(__fmul_rn(2.0f, b1), b2); }
__device__ __forceinline__ float pred3(float b1, float b2, float b3) {
  return __fadd_rn(__fsub_rn(__fmul_rn(3.0f, b1), __fmul_rn(3.0f, b2)), b3);
}

// med_dataset_float's type from max (:3605-3614); 0 = "left at the caller's initial value"
__device__ __forceinline__ int type_from_max(float mx) {
  // thresholds 2^1, 2^65, 2^97, 2^113, 2^121, 2^125, 2^127 are exact in float32 (built from the exponent field)
  #pragma unroll
  for (int t = 0; t < 7; t++) {
    const int ex[7] = {1, 65, 97, 113, 121, 125, 127};
    if (mx < __uint_as_float((uint32_t)(127 + ex[t]) << 23)) return t + 1;
  }
  return 0;
}
