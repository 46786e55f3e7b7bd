// dcb_decode_single.cuh -- decoder for ONE LONG CHUNK (the reference's own call shape: myDecompress_bitwise{,_np,_op,_mask}
// on the whole array, dataCompression.c:2922, :2459, :698, :1703; impl/pingpong.c:553-566, impl/himenoBMTxps.c:696).
// dcb_decode_kernel walks a chunk with one lane; that is the right shape for 65,536 chunks and the wrong one for a single
// stream of 131,072 codes (4.3 ms measured).  Here the two serial dependencies of the format are taken apart:
//
//  (1) code boundaries.  A code is 3..32 bits long, so a code that starts in stream word w ends in w or w + 1, and the
//      offset at which word w + 1 is entered is a function f_w of the offset at which word w was entered.  Every lane builds
//      f_w (and the number of codes) of its word for all 32 entry offsets by a backward sweep over the 32 bit positions;
//      function composition is associative, so warps, blocks and the grid compose them: 32 lanes follow the 32 candidate
//      entries through the tables of a warp's words, then of the block's warps.
//        dcb_parse_kernel<CT, false> : per block, the composite (exit offset, code count) for each of the 32 entries
//        dcb_parse_chain_kernel      : one block chains the block composites -> true entry offset and first element index
//        dcb_parse_kernel<CT, true>  : same tables again, now followed from the true entry: every lane walks the codes of
//                                      its word and writes each element's kind (zero / predictor a, b, c / raw) and, for
//                                      raw and escape codes, the reconstructed value (midpoint bit, :3168-3184).
//  (2) the float recurrence of predicted values (the decoder predicts from RECONSTRUCTED values, :3139-3158; every operation
//      rounds, so it is not associative).  A predicted value needs at most the three values before it: wherever three
//      consecutive elements are known (raw / escape / zero codes, none of them the -1.0f empty-slot marker of :3110-3127)
//      the history is (x[i-3], x[i-2], x[i-1]) whatever came before, and a new RUN starts.
//        dcb_resolve_kernel : a thread owns the run starts inside its 32 elements and walks each run to the next start.
//      Noisy and constant data fall into short runs (all parallel); a stream of predicted codes only (a smooth field) is one
//      run and costs its dependent chain, 8..12 cycles per element -- the format's own limit for one reference call.
//  (3) dcb_addmin_kernel : decompressed[i] + min (impl/pingpong.c:562), only with DCB_F_SUBTRACT_MIN.
#pragma once
#include "dcb_decode.cuh"

#define DSP_NT 256                    // threads per parse block = stream words per tile
#define DSP_TPITCH 34                 // u16 entries per lane table (32 + pad: lanes start in different banks)
#define DS_OP_ZERO 0                  // code 100
#define DS_OP_A 1                     // code 101: b1
#define DS_OP_B 2                     // code 110: 2*b1 - b2
#define DS_OP_C 3                     // code 111: 3*b1 - 3*b2 + b3
#define DS_OP_KNOWN 4                 // raw / escape / full word: out[i] holds the value

struct ParseArgs {
  const uint8_t* stream;
  const uint64_t* offsets;            // [2]: the chunk's byte range
  uint64_t n;                         // elements to decode
  uint32_t tiles_per_block;
  uint32_t num_blocks;
  uint32_t flags;
  int fixed_type;
  uint32_t fixed_mask17;
  CodecConst k;
  const float* cmed;
  const int32_t* ctype;
  unsigned long long* fblk;           // [num_blocks][32]: exit offset | codes << 8, for every candidate entry offset
  uint32_t* blk_entry;                // [num_blocks]
  unsigned long long* blk_base;       // [num_blocks]: index of the first element whose code starts in the block
  uint8_t* ops;                       // [n]
  float* out;                         // [n]
  int32_t* status;
};

// big-endian stream word w (bits 32w .. 32w+31 of the chunk), zero past the chunk's last byte
__device__ __forceinline__ uint32_t ds_word(const uint8_t* stream, uint64_t off0, uint64_t total, uint64_t w) {
  const uint64_t b = 4ull * w;
  if (b >= total) return 0u;
  const uint64_t a = off0 + b;
  const uint32_t* p = reinterpret_cast<const uint32_t*>(stream + (a & ~3ull));
  const uint32_t sh = (uint32_t)(a & 3ull) * 8u;
  const uint32_t lo = __byte_perm(__ldg(p), 0u, 0x0123);
  uint32_t v = lo;
  if (sh) v = __funnelshift_l(__byte_perm(__ldg(p + 1), 0u, 0x0123), lo, sh);
  const uint64_t left = total - b;                                   // valid bytes of this word (>= 1)
  if (left < 4ull) v &= 0xFFFFFFFFu << (8u * (uint32_t)(4ull - left));
  return v;
}

template <int CT>
__device__ __forceinline__ int ds_code_len(uint32_t w, const CodecConst& k, int type, uint32_t ones, uint32_t mask17, int m_mask) {
  if (CT != CT_NP && (w >> 31)) return 3;                            // :2965-2967
  uint32_t vb; int len;
  decode_fallback<CT>(CT == CT_NP ? (w & 0x7FFFFFFFu) : w, k, type, ones, mask17, m_mask, vb, len);
  return len;
}

// EMIT = false: block composites.  EMIT = true: element kinds and known values.
template <int CT, bool EMIT>
__global__ void __launch_bounds__(DSP_NT) dcb_parse_kernel(const ParseArgs a) {
  __shared__ uint16_t T[DSP_NT * DSP_TPITCH];     // per word, per entry offset: exit offset | codes << 8
  __shared__ uint32_t WC[DSP_NT / 32][32];         // per warp, per entry offset: exit offset | codes << 8
  __shared__ uint32_t BC[32];                      // the tile
  __shared__ unsigned long long RC[32];            // composite of this block's tiles so far (EMIT = false)
  __shared__ uint32_t s_entry;                     // EMIT: entry offset / first element index of the current tile
  __shared__ unsigned long long s_base;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint64_t off0 = a.offsets[0], total = a.offsets[1] - a.offsets[0];
  int type = a.fixed_type;
  uint32_t mask17 = a.fixed_mask17;
  if (CT == CT_MASK && !(a.flags & 2u)) { type = a.ctype[0]; mask17 = f2u(a.cmed[0]) >> 15; }
  type = min(max(type, 0), 7);
  const int m_mask = raw_len(mask17 << 15, a.k.n_eb_m118) - 9;
  const uint32_t ones = (1u << type) - 1u;
  if (tid < 32) RC[tid] = (unsigned long long)tid;                 // identity: exit = entry, 0 codes
  if (EMIT && tid == 0) { s_entry = a.blk_entry[blockIdx.x]; s_base = a.blk_base[blockIdx.x]; }
  bool bad = false;
  uint16_t* const myT = T + tid * DSP_TPITCH;
  // a block that lies wholly behind the chunk's last byte holds no code: identity composite, nothing to emit
  if ((uint64_t)blockIdx.x * a.tiles_per_block * DSP_NT * 4ull >= total + 4ull) {
    if (!EMIT && tid < 32) a.fblk[(uint64_t)blockIdx.x * 32 + tid] = (unsigned long long)tid;
    return;
  }

  for (uint32_t t = 0; t < a.tiles_per_block; t++) {
    const uint64_t w = ((uint64_t)blockIdx.x * a.tiles_per_block + t) * DSP_NT + tid;
    const uint32_t w0 = ds_word(a.stream, off0, total, w), w1 = ds_word(a.stream, off0, total, w + 1);
    __syncthreads();                                              // previous tile's tables are no longer read
    // ---- backward sweep: the table of this word ------------------------------------------------------------------
    #pragma unroll 4
    for (int p = 31; p >= 0; p--) {
      const uint32_t bits = __funnelshift_l(w1, w0, p);
      const int np = p + ds_code_len<CT>(bits, a.k, type, ones, mask17, m_mask);
      uint32_t e;
      if (np >= 32) e = (uint32_t)(np - 32) | (1u << 8);
      else e = (uint32_t)myT[np] + (1u << 8);
      myT[p] = (uint16_t)e;
    }
    __syncthreads();
    // ---- warp composite: lane = candidate entry offset, followed through the warp's 32 words ---------------------
    {
      uint32_t cur = (uint32_t)lane, cnt = 0u;
      const uint16_t* wt = T + warp * 32 * DSP_TPITCH;
      #pragma unroll 8
      for (int k = 0; k < 32; k++) {
        const uint32_t e = wt[k * DSP_TPITCH + cur];
        cur = e & 31u; cnt += e >> 8;
      }
      WC[warp][lane] = cur | (cnt << 8);
    }
    __syncthreads();
    if (warp == 0) {
      uint32_t cur = (uint32_t)lane, cnt = 0u;
      #pragma unroll
      for (int k = 0; k < DSP_NT / 32; k++) {
        const uint32_t e = WC[k][cur];
        cur = e & 31u; cnt += e >> 8;
      }
      BC[lane] = cur | (cnt << 8);
      if (!EMIT) {                                                // running composite of the block's tiles
        const unsigned long long r = RC[lane];
        const uint32_t e = __shfl_sync(0xffffffffu, cur | (cnt << 8), (int)(r & 31ull));
        RC[lane] = (unsigned long long)(e & 31u) | (((r >> 8) + (unsigned long long)(e >> 8)) << 8);
      }
    }
    if (!EMIT) continue;
    __syncthreads();
    // ---- EMIT: follow the true entry down to this lane's word ------------------------------------------------------
    uint32_t cur = s_entry;
    unsigned long long idx = s_base;
    for (int k = 0; k < warp; k++) { const uint32_t e = WC[k][cur]; cur = e & 31u; idx += e >> 8; }
    {
      const uint16_t* wt = T + warp * 32 * DSP_TPITCH;
      for (int k = 0; k < lane; k++) { const uint32_t e = wt[k * DSP_TPITCH + cur]; cur = e & 31u; idx += e >> 8; }
    }
    // ---- the codes that start in this word ---------------------------------------------------------------------------
    int p = (int)cur;
    while (p < 32) {
      const uint32_t bits = __funnelshift_l(w1, w0, p);
      int len; uint32_t op, vb = 0u;
      if (CT != CT_NP && (bits >> 31)) {
        len = 3; op = (bits >> 29) & 3u;                          // 100 / 101 / 110 / 111
      } else {
        decode_fallback<CT>(CT == CT_NP ? (bits & 0x7FFFFFFFu) : bits, a.k, type, ones, mask17, m_mask, vb, len);
        if (CT == CT_NP) { bad |= (bits >> 31) != 0u; vb |= bits & 0x80000000u; }   // :2501-2505
        op = DS_OP_KNOWN;
      }
      if (idx < a.n) { a.ops[idx] = (uint8_t)op; a.out[idx] = u2f(vb); }
      idx++;
      p += len;
    }
    __syncthreads();
    if (tid == 0) { const uint32_t e = BC[s_entry]; s_entry = e & 31u; s_base += e >> 8; }
  }
  if (!EMIT) {
    __syncthreads();
    if (tid < 32) a.fblk[(uint64_t)blockIdx.x * 32 + tid] = RC[tid];
  }
  if (EMIT && bad && a.status) atomicExch(a.status, -5);
}

// One block: entry offset and first element index of every parse block (the chain of block composites).
__global__ void __launch_bounds__(256) dcb_parse_chain_kernel(const ParseArgs a) {
  __shared__ unsigned long long S[128 * 32];
  __shared__ uint32_t s_cur;
  __shared__ unsigned long long s_idx;
  if (threadIdx.x == 0) { s_cur = 0u; s_idx = 0ull; }             // word 0 starts at the chunk's first bit
  // blocks behind the chunk's last byte hold no code and emit nothing (they never read their entry): stop at the last one
  const uint64_t total = a.offsets[1] - a.offsets[0];
  const uint64_t words_per_block = (uint64_t)a.tiles_per_block * DSP_NT;
  const uint64_t used64 = (total / 4ull + 2ull) / words_per_block + 1ull;
  const uint32_t used = used64 < (uint64_t)a.num_blocks ? (uint32_t)used64 : a.num_blocks;
  for (uint32_t b0 = 0; b0 < used; b0 += 128) {
    const uint32_t nb = min(128u, used - b0);
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < nb * 32; i += blockDim.x) S[i] = a.fblk[(uint64_t)b0 * 32 + i];
    __syncthreads();
    if (threadIdx.x == 0) {
      uint32_t cur = s_cur; unsigned long long idx = s_idx;
      for (uint32_t b = 0; b < nb; b++) {
        a.blk_entry[b0 + b] = cur; a.blk_base[b0 + b] = idx;
        const unsigned long long e = S[b * 32 + cur];
        cur = (uint32_t)(e & 31ull); idx += e >> 8;
      }
      s_cur = cur; s_idx = idx;
    }
  }
}

// ---- (2) predicted values ----------------------------------------------------------------------------------------------
#define DSR_G 32                      // elements per thread (run starts it owns)
struct ResolveArgs {
  const uint8_t* ops;
  uint4* masks;                       // per group of 32 elements: .x = not predicted (zero / known), .y = 110, .z = 111
  float* out;
  uint64_t n;
};

// kinds of 32 elements -> three bit masks (bit e = element 32 g + e); groups past n read "not predicted"
__global__ void dcb_opmask_kernel(const ResolveArgs a, uint64_t groups) {
  const uint64_t g = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= groups) return;
  uint32_t mk = 0u, mb = 0u, mc = 0u;
  const uint4* src = reinterpret_cast<const uint4*>(a.ops + 32ull * g);
  #pragma unroll
  for (int h = 0; h < 2; h++) {
    const uint4 v = (32ull * g + 16ull * h < a.n) ? src[h] : make_uint4(0u, 0u, 0u, 0u);
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
    #pragma unroll
    for (int i = 0; i < 16; i++) {
      const uint32_t op = (w[i >> 2] >> (8 * (i & 3))) & 7u;
      const uint64_t idx = 32ull * g + 16ull * h + i;
      const uint32_t bit = 1u << (16 * h + i);
      if (idx >= a.n || op == DS_OP_ZERO || op == DS_OP_KNOWN) mk |= bit;
      else if (op == DS_OP_B) mb |= bit;
      else if (op == DS_OP_C) mc |= bit;
    }
  }
  a.masks[g] = make_uint4(mk, mb, mc, 0u);
}

__device__ __forceinline__ bool ds_known(const uint8_t* ops, const float* out, uint64_t i) {
  const uint32_t op = ops[i];
  return (op == DS_OP_ZERO || op == DS_OP_KNOWN) && out[i] != -1.0f;
}

// one element with the reference's exact history state machine; returns the value
__device__ __forceinline__ float ds_step_exact(uint32_t op, float known, float& b3, float& b2, float& b1) {
  float v;
  if (op == DS_OP_ZERO || op == DS_OP_KNOWN) v = known;
  else if (op == DS_OP_A) v = b1;                                   // decompress_bitwise_float :3139-3158
  else if (op == DS_OP_B) v = __fsub_rn(__fadd_rn(b1, b1), b2);     // 2*b1 is exact, but may overflow: no FMA
  else v = __fadd_rn(__fsub_rn(__fmul_rn(3.0f, b1), __fmul_rn(3.0f, b2)), b3);
  hist_push(b3, b2, b1, v);                                         // :3110-3127
  return v;
}

// the 32 predicted values of a group once more, with exact compares against the -1.0f marker
__device__ __noinline__ bool ds_group_has_marker(float c3, float c2, float c1, uint32_t mb, uint32_t mc) {
  bool mk = false;
  float u1 = __fmul_rn(3.0f, c1), u2 = __fmul_rn(3.0f, c2);
  #pragma unroll 8
  for (int e = 0; e < 32; e++) {
    const float w = __fadd_rn(c1, c1);
    const float d = __fsub_rn(u1, u2);
    float x = c1;
    if (mb & (1u << e)) x = __fsub_rn(w, c2);
    if (mc & (1u << e)) x = __fadd_rn(d, c3);
    mk = mk || (x == -1.0f);
    c3 = c2; c2 = c1; c1 = x; u2 = u1; u1 = __fmul_rn(3.0f, x);
  }
  return mk;
}

// A thread walks every run that STARTS inside its DSR_G elements, each to the next run start (wherever that lies).  Known
// elements are never written, so a neighbouring walker that reads the three values before its own start sees them unchanged.
// Long runs (a smooth field is ONE run) are bound by the dependent float chain, so the walk keeps everything else off it:
// the kinds of 32 elements arrive as three bit masks (dcb_opmask_kernel), two groups ahead; 32 predicted codes in a row are
// resolved with the history as pure register renaming, and only if a value came out negative (the -1.0f marker is) is the
// group redone with the exact state machine.
__global__ void __launch_bounds__(128) dcb_resolve_kernel(const ResolveArgs a) {
  const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t lo = t * DSR_G, hi = min(lo + DSR_G, a.n);
  if (lo >= a.n) return;
  const uint8_t* __restrict__ ops = a.ops;
  const uint4* __restrict__ gm = a.masks;                          // padded: three groups past n
  float* __restrict__ out = a.out;
  // consecutive known elements (no -1.0f among them) just before `lo`, capped at 3
  int kn = 0;
  for (int b = 1; b <= 3 && lo >= (uint64_t)b; b++) { if (ds_known(ops, out, lo - b)) kn++; else break; }
  for (uint64_t s = lo; s < hi; s++) {
    if (!(s == 0 || kn >= 3)) { kn = ds_known(ops, out, s) ? min(kn + 1, 3) : 0; continue; }
    // ---- walk the run that starts at s ----
    float b3 = -1.0f, b2 = -1.0f, b1 = -1.0f;                       // :2930
    int k = 0;                                                      // known non-marker elements in a row up to here
    if (s != 0) { b3 = out[s - 3]; b2 = out[s - 2]; b1 = out[s - 1]; k = 3; }
    uint64_t j = s;
    bool full = b3 != -1.0f && b2 != -1.0f && b1 != -1.0f;
    for (;;) {
      if ((j & 31ull) == 0ull && full) {
        // Groups of 32 predicted codes, as long as they last.  One lane walks a long run alone in its warp and nothing hides
        // a latency, so this is straight-line code: the kind of an element only predicates its additions (a bit test of the
        // group's masks), the history is register renaming, and the -1.0f marker is looked for afterwards (it is negative:
        // one OR per value, four interleaved accumulators).  (Round-2 profile: the arithmetic runs at the ~15 cycles per
        // element its dependent chain allows; the rest was the mask load and per-group bookkeeping.)
        // returns 0: group done, go on; 1: not 32 predicted codes (nothing done); 2: done, but the history lost a slot
        auto group = [&](const uint4& m) -> int {
          if (m.x != 0u) return 1;
          const float s3 = b3, s2 = b2, s1 = b1;
          const uint32_t mb = m.y, mc = m.z;
          k = 0;                                                    // no known element in this group
          float v[4];
          uint32_t neg[4] = {0u, 0u, 0u, 0u};
          float t1 = __fmul_rn(3.0f, b1), t2 = __fmul_rn(3.0f, b2);
          #pragma unroll
          for (int e = 0; e < 32; e++) {
            const float w = __fadd_rn(b1, b1);                      // 2*b1 is exact (may overflow: no FMA)
            const float d = __fsub_rn(t1, t2);
            float x = b1;                                           // 101
            if (mb & (1u << e)) x = __fsub_rn(w, b2);               // 110
            if (mc & (1u << e)) x = __fadd_rn(d, b3);               // 111
            v[e & 3] = x; neg[e & 3] |= f2u(x);
            b3 = b2; b2 = b1; b1 = x; t2 = t1; t1 = __fmul_rn(3.0f, x);
            // stored as soon as four are there (a redo below rewrites the whole group): no burst of stores, and no wait
            // for their operands, at the end of the group
            if ((e & 3) == 3) *reinterpret_cast<float4*>(out + j + e - 3) = make_float4(v[0], v[1], v[2], v[3]);
          }
          int r = 0;
          // a negative value: look whether one IS the marker (closed-loop predictions drift below zero around a minimum, the
          // marker itself is rare); out of line, so that the group's own code stays as it was
          const bool mk = (((neg[0] | neg[1]) | (neg[2] | neg[3])) >> 31) && ds_group_has_marker(s3, s2, s1, mb, mc);
          if (mk) {                                                 // the marker -> exact state machine
            b3 = s3; b2 = s2; b1 = s1;
            #pragma unroll 1
            for (int e = 0; e < 32; e++) {
              const uint32_t op = ((mb >> e) & 1u) ? DS_OP_B : (((mc >> e) & 1u) ? DS_OP_C : DS_OP_A);
              out[j + e] = ds_step_exact(op, 0.0f, b3, b2, b1);
            }
            full = b3 != -1.0f && b2 != -1.0f && b1 != -1.0f;
            if (!full) r = 2;
          }
          j += 32;
          return r;
        };
        // the masks of the next group are requested before this group's chain starts, into a variable of their own: two
        // copies of the group in the loop body, no rotation (a rotation's register moves were scheduled right behind the load
        // and waited for it: 17-22 % of the kernel's samples)
        uint64_t g = j >> 5;                                        // gm is padded: g + 1 is always readable
        uint4 mA = gm[g], mB;                                       // (pinning these loads with volatile asm was slower: 0.985 vs 0.903 ms)
        while (j + 32 <= a.n) {
          mB = gm[g + 1];
          if (group(mA)) break;
          g++;
          if (j + 32 > a.n) break;
          mA = gm[g + 1];
          if (group(mB)) break;
          g++;
        }
        if (j >= a.n) break;
      }
      {
        const uint32_t op = ops[j];
        const bool kno = (op == DS_OP_ZERO || op == DS_OP_KNOWN);
        const float kv = kno ? out[j] : 0.0f;
        const float v = ds_step_exact(op, kv, b3, b2, b1);
        if (!kno) out[j] = v;
        k = (kno && v != -1.0f) ? k + 1 : 0;
        full = b3 != -1.0f && b2 != -1.0f && b1 != -1.0f;
        j++;
        if (j >= a.n || k >= 3) break;                              // element j starts the next run
      }
    }
    // the elements s+1 .. j-1 start no run; continue the ownership scan at j
    if (j >= hi) return;
    s = j - 1; kn = 3;
  }
}

__global__ void dcb_addmin_kernel(float* out, uint64_t n, const float* cmin) {
  const float mn = cmin[0];
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x)
    out[i] = __fadd_rn(out[i], mn);                                 // decompressed[i] + min, impl/pingpong.c:562
}

// ---- host side -----------------------------------------------------------------------------------------------------------
static inline uint64_t ds_num_words(uint64_t n) { return n + 2; }  // a code has at most 32 bits: n codes end within n words (+ slack)
static inline void ds_geometry(uint64_t n, uint32_t* tiles_per_block, uint32_t* num_blocks) {
  const uint64_t tiles = (ds_num_words(n) + DSP_NT - 1) / DSP_NT;
  uint64_t tpb = (tiles + 4095) / 4096;                             // at most 4096 parse blocks
  if (tpb == 0) tpb = 1;
  *tiles_per_block = (uint32_t)tpb;
  *num_blocks = (uint32_t)((tiles + tpb - 1) / tpb);
}
// scratch: block composites [nb][32] u64 | blk_base [nb] u64 | blk_entry [nb] u32 | ops [n] u8 | masks [n/32 + 3] uint4
static inline uint64_t ds_ops_bytes(uint64_t n) { return (n + 64 + 255) & ~(uint64_t)255; }
static inline uint64_t ds_workspace_bytes(uint64_t n) {
  uint32_t tpb, nb; ds_geometry(n, &tpb, &nb);
  return (uint64_t)nb * (32 * 8 + 8 + 4) + 256 + ds_ops_bytes(n) + ((n + 31) / 32 + 3) * 16 + 256;
}

// launches the five kernels on `st`; d_ws: ds_workspace_bytes(n) bytes, 256-byte aligned
template <int CT>
static cudaError_t ds_launch(const DecArgs& d, void* d_ws, cudaStream_t st) {
  uint32_t tpb, nb; ds_geometry(d.n, &tpb, &nb);
  char* ws = (char*)d_ws;
  ParseArgs a;
  a.stream = d.stream; a.offsets = d.offsets; a.n = d.n; a.tiles_per_block = tpb; a.num_blocks = nb;
  a.flags = d.flags; a.fixed_type = d.fixed_type; a.fixed_mask17 = d.fixed_mask17; a.k = d.k; a.cmed = d.cmed; a.ctype = d.ctype;
  a.fblk = (unsigned long long*)ws;
  a.blk_base = (unsigned long long*)(ws + (uint64_t)nb * 256);
  a.blk_entry = (uint32_t*)(ws + (uint64_t)nb * 264);
  a.ops = (uint8_t*)(ws + (((uint64_t)nb * 268 + 255) & ~(uint64_t)255));
  a.out = d.out; a.status = d.status;
  dcb_parse_kernel<CT, false><<<nb, DSP_NT, 0, st>>>(a);
  dcb_parse_chain_kernel<<<1, 256, 0, st>>>(a);
  dcb_parse_kernel<CT, true><<<nb, DSP_NT, 0, st>>>(a);
  if (CT != CT_NP) {                                                // CT=6 has no predicted codes
    ResolveArgs r; r.ops = a.ops; r.out = d.out; r.n = d.n;
    r.masks = (uint4*)(a.ops + ds_ops_bytes(d.n));
    const uint64_t groups = (d.n + 31) / 32 + 3;                    // three "not predicted" groups behind the last element
    dcb_opmask_kernel<<<(unsigned)((groups + 127) / 128), 128, 0, st>>>(r, groups);
    const uint64_t threads = (d.n + DSR_G - 1) / DSR_G;
    dcb_resolve_kernel<<<(unsigned)((threads + 127) / 128), 128, 0, st>>>(r);
  }
  if ((d.flags & 1u) && d.cmin) {
    const uint64_t blocks = (d.n + 1023) / 1024;
    dcb_addmin_kernel<<<(unsigned)(blocks < 4096 ? blocks : 4096), 256, 0, st>>>(d.out, d.n, d.cmin);
  }
  return cudaGetLastError();
}
