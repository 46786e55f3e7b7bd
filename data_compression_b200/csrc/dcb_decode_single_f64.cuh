// dcb_decode_single_f64.cuh -- float64 twin of dcb_decode_single.cuh: decoder for ONE LONG CHUNK of the *_double codecs, the
// call shape of the reference's k-means / MM / LU drivers (myDecompress_bitwise_double{,_np,_op,_mask} on a whole array,
// impl/dataCompression.c:2656, :2286, :476, :1199; callers impl/k-means.c:520-590, impl/mm.c:404-690, impl/lu.c:199-345).
// dcb_f64_decode_kernel walks a chunk with one thread -- right for 32,768 chunks, hopeless for one stream of a million codes.
//
// Same two steps as the float32 file, with the float64 geometry:
//  (1) code boundaries.  A code is 3..64 bits long, so the unit is a 64-BIT stream word: a code that starts in word W ends
//      in W or W + 1, and the offset at which W + 1 is entered is a function f_W of the 64 possible entry offsets of W.
//      Lanes build f_W (and the code count) by a backward sweep over the 64 bit positions; warps, blocks and the grid compose
//      them (every lane follows two of the 64 candidate entries).
//        dcb_parse64_kernel<CT, false> -> dcb_parse64_chain_kernel -> dcb_parse64_kernel<CT, true> (element kinds + known values)
//  (2) predicted values: the decoder predicts from RECONSTRUCTED doubles (:2797-2814), every operation rounded; wherever three
//      consecutive elements are known and none is the -1.0 empty-slot marker, the history is those three whatever came before
//      (the float64 fill order b3, b2, b1 differs from float32's, the property does not), and a new run starts.
//        dcb_resolve64_kernel : a thread owns the run starts inside its 32 elements and walks each run to the next start.
//  (3) dcb_addmin64_kernel : decompressed[i] + min (impl/k-means.c:584), only with DCB_F_SUBTRACT_MIN.
#pragma once
#include "dcb_f64.cuh"
#include "dcb_decode_single.cuh"      // ds_word, DS_OP_*, ResolveArgs + dcb_opmask_kernel (kinds -> bit masks; shared)

#define DSP64_NT 256                  // threads per parse block = 64-bit stream words per tile
#define DSP64_TPITCH 66               // u16 entries per lane table (64 + 2: 33 banks apart)

struct Parse64Args {
  const uint8_t* stream;
  const uint64_t* offsets;            // [2]: the chunk's byte range
  uint64_t n;                         // elements to decode
  uint32_t tiles_per_block;
  uint32_t num_blocks;
  uint32_t flags;
  int fixed_type;
  uint32_t fixed_mask20;
  F64Const k;
  const uint32_t* cmask;
  const int32_t* ctype;
  unsigned long long* fblk;           // [num_blocks][64]: exit offset | codes << 8, for every candidate entry offset
  uint32_t* blk_entry;                // [num_blocks]
  unsigned long long* blk_base;       // [num_blocks]: index of the first element whose code starts in the block
  uint8_t* ops;                       // [n]
  double* out;                        // [n]
  int32_t* status;
};

// big-endian 64-bit stream word W (bits 64W .. 64W+63 of the chunk), zero past the chunk's last byte
__device__ __forceinline__ uint64_t ds64_word(const uint8_t* stream, uint64_t off0, uint64_t total, uint64_t W) {
  return ((uint64_t)ds_word(stream, off0, total, 2ull * W) << 32) | (uint64_t)ds_word(stream, off0, total, 2ull * W + 1ull);
}
__device__ __forceinline__ uint64_t ds64_window(uint64_t w0, uint64_t w1, int p) {
  return p ? ((w0 << p) | (w1 >> (64 - p))) : w0;
}

// A code that does not start with 1 (or any code of CT=6), at the top of the window w: reconstructed value bits and length.
// The same arithmetic as the non-3-bit branch of dcb_f64_decode_kernel (decompress_bitwise_double :2816-2862, _mask :1340-1491).
template <int CT>
__device__ __forceinline__ void f64_decode_fallback(uint64_t w, const F64Const& k, int type, uint32_t mask20, int m_mask,
                                                    uint64_t& vb, int& len) {
  if (CT == CT_OP) { len = 64; vb = w; return; }
  bool escape = false;
  if (CT == CT_MASK) { const uint64_t ones = (1ull << type) - 1ull; escape = ((w >> (63 - type)) & ones) == ones; }
  uint64_t top; int have;
  if (!escape) {
    have = 12 + f64_mant_bits(w, k.n_eb);
    top = have == 64 ? w : (w & ~(~0ull >> have));
    len = have;
  } else {
    const int flag = (int)((w >> (62 - type)) & 1ull);
    const int kk = flag ? m_mask : max(m_mask - 8, 0);
    const int hdr = type + 2;
    const uint64_t payload = kk ? ((w << hdr) >> (64 - kk)) : 0ull;
    len = hdr + kk;
    if (flag) { top = ((uint64_t)(mask20 >> 8) << 52) | (kk ? payload << (52 - kk) : 0ull); have = 12 + kk; }
    else { top = ((uint64_t)mask20 << 44) | (kk ? payload << (44 - kk) : 0ull); have = 20 + kk; }
  }
  vb = have >= 64 ? top : (top | (0x8000000000000000ull >> have));
}
template <int CT>
__device__ __forceinline__ int ds64_code_len(uint64_t w, const F64Const& k, int type, uint32_t mask20, int m_mask) {
  if (CT != CT_NP && (w >> 63)) return 3;
  uint64_t vb; int len;
  f64_decode_fallback<CT>(w, k, type, mask20, m_mask, vb, len);
  return len;
}

// EMIT = false: block composites.  EMIT = true: element kinds and known values.
template <int CT, bool EMIT>
__global__ void __launch_bounds__(DSP64_NT) dcb_parse64_kernel(const Parse64Args a) {
  __shared__ uint16_t T[DSP64_NT * DSP64_TPITCH];  // per word, per entry offset: exit offset | codes << 8
  __shared__ uint32_t WC[DSP64_NT / 32][64];        // per warp, per entry offset: exit offset | codes << 8
  __shared__ uint32_t BC[64];                       // the tile
  __shared__ unsigned long long RC[64];             // composite of this block's tiles so far (EMIT = false)
  __shared__ uint32_t s_entry;                      // EMIT: entry offset / first element index of the current tile
  __shared__ unsigned long long s_base;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint64_t off0 = a.offsets[0], total = a.offsets[1] - a.offsets[0];
  int type = a.fixed_type;
  uint32_t mask20 = a.fixed_mask20;
  if (CT == CT_MASK && !(a.flags & DCB_F_FIXED_MASK)) { type = a.ctype[0]; mask20 = a.cmask[0]; }
  const int m_mask = f64_mant_bits((uint64_t)mask20 << 44, a.k.n_eb);
  if (tid < 64) RC[tid] = (unsigned long long)tid;                 // identity: exit = entry, 0 codes
  if (EMIT && tid == 0) { s_entry = a.blk_entry[blockIdx.x]; s_base = a.blk_base[blockIdx.x]; }
  bool bad = false;
  uint16_t* const myT = T + tid * DSP64_TPITCH;
  // a block that lies wholly behind the chunk's last byte holds no code: identity composite, nothing to emit
  if ((uint64_t)blockIdx.x * a.tiles_per_block * DSP64_NT * 8ull >= total + 8ull) {
    if (!EMIT && tid < 64) a.fblk[(uint64_t)blockIdx.x * 64 + tid] = (unsigned long long)tid;
    return;
  }

  for (uint32_t t = 0; t < a.tiles_per_block; t++) {
    const uint64_t W = ((uint64_t)blockIdx.x * a.tiles_per_block + t) * DSP64_NT + tid;
    const uint64_t w0 = ds64_word(a.stream, off0, total, W), w1 = ds64_word(a.stream, off0, total, W + 1);
    __syncthreads();                                              // previous tile's tables are no longer read
    // ---- backward sweep: the table of this word ------------------------------------------------------------------
    #pragma unroll 4
    for (int p = 63; p >= 0; p--) {
      const int np = p + ds64_code_len<CT>(ds64_window(w0, w1, p), a.k, type, mask20, m_mask);
      uint32_t e;
      if (np >= 64) e = (uint32_t)(np - 64) | (1u << 8);
      else e = (uint32_t)myT[np] + (1u << 8);
      myT[p] = (uint16_t)e;
    }
    __syncthreads();
    // ---- warp composite: two candidate entry offsets per lane, followed through the warp's 32 words --------------
    {
      const uint16_t* wt = T + warp * 32 * DSP64_TPITCH;
      uint32_t cur0 = (uint32_t)lane, cur1 = (uint32_t)lane + 32u, cnt0 = 0u, cnt1 = 0u;
      #pragma unroll 8
      for (int k = 0; k < 32; k++) {
        const uint32_t e0 = wt[k * DSP64_TPITCH + cur0], e1 = wt[k * DSP64_TPITCH + cur1];
        cur0 = e0 & 63u; cnt0 += e0 >> 8;
        cur1 = e1 & 63u; cnt1 += e1 >> 8;
      }
      WC[warp][lane] = cur0 | (cnt0 << 8);
      WC[warp][lane + 32] = cur1 | (cnt1 << 8);
    }
    __syncthreads();
    if (warp == 0) {
      #pragma unroll
      for (int h = 0; h < 2; h++) {
        uint32_t cur = (uint32_t)lane + 32u * h, cnt = 0u;
        #pragma unroll
        for (int k = 0; k < DSP64_NT / 32; k++) {
          const uint32_t e = WC[k][cur];
          cur = e & 63u; cnt += e >> 8;
        }
        BC[lane + 32 * h] = cur | (cnt << 8);
      }
      __syncwarp();
      if (!EMIT) {                                                // running composite of the block's tiles
        #pragma unroll
        for (int h = 0; h < 2; h++) {
          const unsigned long long r = RC[lane + 32 * h];
          const uint32_t e = BC[(uint32_t)(r & 63ull)];
          RC[lane + 32 * h] = (unsigned long long)(e & 63u) | (((r >> 8) + (unsigned long long)(e >> 8)) << 8);
        }
      }
    }
    if (!EMIT) continue;
    __syncthreads();
    // ---- EMIT: follow the true entry down to this lane's word ------------------------------------------------------
    uint32_t cur = s_entry;
    unsigned long long idx = s_base;
    for (int k = 0; k < warp; k++) { const uint32_t e = WC[k][cur]; cur = e & 63u; idx += e >> 8; }
    {
      const uint16_t* wt = T + warp * 32 * DSP64_TPITCH;
      for (int k = 0; k < lane; k++) { const uint32_t e = wt[k * DSP64_TPITCH + cur]; cur = e & 63u; idx += e >> 8; }
    }
    // ---- the codes that start in this word ---------------------------------------------------------------------------
    int p = (int)cur;
    while (p < 64) {
      const uint64_t bits = ds64_window(w0, w1, p);
      int len; uint32_t op; uint64_t vb = 0ull;
      if (CT != CT_NP && (bits >> 63)) {
        len = 3; op = (uint32_t)(bits >> 61) & 3u;                // 100 / 101 / 110 / 111
      } else {
        f64_decode_fallback<CT>(bits, a.k, type, mask20, m_mask, vb, len);
        if (CT == CT_NP) bad |= (bits >> 63) != 0ull;             // "Error leading bit 1" (:2330-2334)
        op = DS_OP_KNOWN;
      }
      if (idx < a.n) { a.ops[idx] = (uint8_t)op; a.out[idx] = __longlong_as_double((long long)vb); }
      idx++;
      p += len;
    }
    __syncthreads();
    if (tid == 0) { const uint32_t e = BC[s_entry]; s_entry = e & 63u; s_base += e >> 8; }
  }
  if (!EMIT) {
    __syncthreads();
    if (tid < 64) a.fblk[(uint64_t)blockIdx.x * 64 + tid] = RC[tid];
  }
  if (EMIT && bad && a.status) atomicExch(a.status, DCB_EFORMAT);
}

// One block: entry offset and first element index of every parse block (the chain of block composites).
__global__ void __launch_bounds__(256) dcb_parse64_chain_kernel(const Parse64Args a) {
  __shared__ unsigned long long S[64 * 64];
  __shared__ uint32_t s_cur;
  __shared__ unsigned long long s_idx;
  if (threadIdx.x == 0) { s_cur = 0u; s_idx = 0ull; }             // word 0 starts at the chunk's first bit
  const uint64_t total = a.offsets[1] - a.offsets[0];
  const uint64_t words_per_block = (uint64_t)a.tiles_per_block * DSP64_NT;
  const uint64_t used64 = (total / 8ull + 2ull) / words_per_block + 1ull;    // blocks behind the last byte emit nothing
  const uint32_t used = used64 < (uint64_t)a.num_blocks ? (uint32_t)used64 : a.num_blocks;
  for (uint32_t b0 = 0; b0 < used; b0 += 64) {
    const uint32_t nb = min(64u, used - b0);
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < nb * 64; i += blockDim.x) S[i] = a.fblk[(uint64_t)b0 * 64 + i];
    __syncthreads();
    if (threadIdx.x == 0) {
      uint32_t cur = s_cur; unsigned long long idx = s_idx;
      for (uint32_t b = 0; b < nb; b++) {
        a.blk_entry[b0 + b] = cur; a.blk_base[b0 + b] = idx;
        const unsigned long long e = S[b * 64 + cur];
        cur = (uint32_t)(e & 63ull); idx += e >> 8;
      }
      s_cur = cur; s_idx = idx;
    }
  }
}

// ---- (2) predicted values ----------------------------------------------------------------------------------------------
struct Resolve64Args {
  const uint8_t* ops;
  const uint4* masks;                 // per group of 32 elements: .x = not predicted (zero / known), .y = 110, .z = 111
  double* out;
  uint64_t n;
};

__device__ __forceinline__ bool ds64_known(const uint8_t* ops, const double* out, uint64_t i) {
  const uint32_t op = ops[i];
  return (op == DS_OP_ZERO || op == DS_OP_KNOWN) && out[i] != -1.0;
}
// one element with the reference's exact history state machine (:2797-2814, fill order b3, b2, b1); returns the value
__device__ __forceinline__ double ds64_step_exact(uint32_t op, double known, double& b3, double& b2, double& b1) {
  double v;
  if (op == DS_OP_ZERO || op == DS_OP_KNOWN) v = known;
  else if (op == DS_OP_A) v = b1;
  else if (op == DS_OP_B) v = __dsub_rn(__dmul_rn(2.0, b1), b2);
  else v = __dadd_rn(__dsub_rn(__dmul_rn(3.0, b1), __dmul_rn(3.0, b2)), b3);
  const bool full = b3 != -1.0 && b2 != -1.0 && b1 != -1.0;
  if (full && v != -1.0) { b3 = b2; b2 = b1; b1 = v; }
  else if (b3 == -1.0) b3 = v;
  else if (b2 == -1.0) b2 = v;
  else if (b1 == -1.0) b1 = v;
  else { b3 = b2; b2 = b1; b1 = v; }
  return v;
}

// the 32 predicted values of a group once more, with exact compares against the -1.0 marker
__device__ __noinline__ bool ds64_group_has_marker(double c3, double c2, double c1, uint32_t mb, uint32_t mc) {
  bool mk = false;
  double u1 = __dmul_rn(3.0, c1), u2 = __dmul_rn(3.0, c2);
  #pragma unroll 8
  for (int e = 0; e < 32; e++) {
    const double w = __dadd_rn(c1, c1);
    const double d = __dsub_rn(u1, u2);
    double x = c1;
    if (mb & (1u << e)) x = __dsub_rn(w, c2);
    if (mc & (1u << e)) x = __dadd_rn(d, c3);
    mk = mk || (x == -1.0);
    c3 = c2; c2 = c1; c1 = x; u2 = u1; u1 = __dmul_rn(3.0, x);
  }
  return mk;
}

// As dcb_resolve_kernel: a thread walks every run that STARTS inside its 32 elements, each to the next run start.  Known
// elements are never written, so a neighbouring walker that reads the three values before its own start sees them unchanged.
__global__ void __launch_bounds__(128) dcb_resolve64_kernel(const Resolve64Args a) {
  const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint64_t lo = t * 32ull, hi = (lo + 32ull < a.n) ? lo + 32ull : a.n;
  if (lo >= a.n) return;
  const uint8_t* __restrict__ ops = a.ops;
  const uint4* __restrict__ gm = a.masks;                          // padded: three groups past n
  double* __restrict__ out = a.out;
  int kn = 0;                                                      // consecutive known non-marker elements just before `lo`
  for (int b = 1; b <= 3 && lo >= (uint64_t)b; b++) { if (ds64_known(ops, out, lo - b)) kn++; else break; }
  for (uint64_t s = lo; s < hi; s++) {
    if (!(s == 0 || kn >= 3)) { kn = ds64_known(ops, out, s) ? min(kn + 1, 3) : 0; continue; }
    // ---- walk the run that starts at s ----
    double b3 = -1.0, b2 = -1.0, b1 = -1.0;
    int k = 0;
    if (s != 0) { b3 = out[s - 3]; b2 = out[s - 2]; b1 = out[s - 1]; k = 3; }
    uint64_t j = s;
    bool full = b3 != -1.0 && b2 != -1.0 && b1 != -1.0;
    for (;;) {
      if ((j & 31ull) == 0ull && full) {
        // groups of 32 predicted codes: straight-line, history by renaming, the marker (negative) looked for afterwards
        auto group = [&](const uint4& m) -> int {
          if (m.x != 0u) return 1;
          const double s3 = b3, s2 = b2, s1 = b1;
          const uint32_t mb = m.y, mc = m.z;
          k = 0;
          double v[2];
          uint32_t neg = 0u;
          double t1 = __dmul_rn(3.0, b1), t2 = __dmul_rn(3.0, b2);
          #pragma unroll
          for (int e = 0; e < 32; e++) {
            const double w = __dadd_rn(b1, b1);                     // = 2.0 * b1, exact
            const double d = __dsub_rn(t1, t2);
            double x = b1;
            if (mb & (1u << e)) x = __dsub_rn(w, b2);
            if (mc & (1u << e)) x = __dadd_rn(d, b3);
            v[e & 1] = x; neg |= (uint32_t)__double2hiint(x);
            b3 = b2; b2 = b1; b1 = x; t2 = t1; t1 = __dmul_rn(3.0, x);
            if (e & 1) *reinterpret_cast<double2*>(out + j + e - 1) = make_double2(v[0], v[1]);
          }
          int r = 0;
          // a negative value: look whether one IS the marker (out of line: the group's own code stays as it was)
          const bool mk = (neg >> 31) && ds64_group_has_marker(s3, s2, s1, mb, mc);
          if (mk) {                                                 // the marker -> exact state machine
            b3 = s3; b2 = s2; b1 = s1;
            #pragma unroll 1
            for (int e = 0; e < 32; e++) {
              const uint32_t op = ((mb >> e) & 1u) ? DS_OP_B : (((mc >> e) & 1u) ? DS_OP_C : DS_OP_A);
              out[j + e] = ds64_step_exact(op, 0.0, b3, b2, b1);
            }
            full = b3 != -1.0 && b2 != -1.0 && b1 != -1.0;
            if (!full) r = 2;
          }
          j += 32;
          return r;
        };
        uint64_t g = j >> 5;                                        // gm is padded: g + 1 is always readable
        uint4 mA = gm[g], mB;
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
        const double kv = kno ? out[j] : 0.0;
        const double v = ds64_step_exact(op, kv, b3, b2, b1);
        if (!kno) out[j] = v;
        k = (kno && v != -1.0) ? k + 1 : 0;
        full = b3 != -1.0 && b2 != -1.0 && b1 != -1.0;
        j++;
        if (j >= a.n || k >= 3) break;                              // element j starts the next run
      }
    }
    if (j >= hi) return;
    s = j - 1; kn = 3;
  }
}

__global__ void dcb_addmin64_kernel(double* out, uint64_t n, const double* cmin) {
  const double mn = cmin[0];
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x)
    out[i] = __dadd_rn(out[i], mn);
}

// ---- host side -----------------------------------------------------------------------------------------------------------
static inline void ds64_geometry(uint64_t n, uint32_t* tiles_per_block, uint32_t* num_blocks) {
  const uint64_t tiles = (n + 2 + DSP64_NT - 1) / DSP64_NT;         // a code has at most 64 bits: n codes end within n words
  uint64_t tpb = (tiles + 4095) / 4096;
  if (tpb == 0) tpb = 1;
  *tiles_per_block = (uint32_t)tpb;
  *num_blocks = (uint32_t)((tiles + tpb - 1) / tpb);
}
// scratch: block composites [nb][64] u64 | blk_base [nb] u64 | blk_entry [nb] u32 | ops [n] u8 | masks [n/32 + 3] uint4
static inline uint64_t ds64_workspace_bytes(uint64_t n) {
  uint32_t tpb, nb; ds64_geometry(n, &tpb, &nb);
  return (uint64_t)nb * (64 * 8 + 8 + 4) + 256 + ds_ops_bytes(n) + ((n + 31) / 32 + 3) * 16 + 256;
}
template <int CT>
static cudaError_t ds64_launch(const F64DecArgs& d, void* d_ws, cudaStream_t st) {
  uint32_t tpb, nb; ds64_geometry(d.n, &tpb, &nb);
  char* ws = (char*)d_ws;
  Parse64Args a;
  a.stream = d.stream; a.offsets = d.offsets; a.n = d.n; a.tiles_per_block = tpb; a.num_blocks = nb;
  a.flags = d.flags; a.fixed_type = d.fixed_type; a.fixed_mask20 = d.fixed_mask20; a.k = d.k; a.cmask = d.cmask; a.ctype = d.ctype;
  a.fblk = (unsigned long long*)ws;
  a.blk_base = (unsigned long long*)(ws + (uint64_t)nb * 512);
  a.blk_entry = (uint32_t*)(ws + (uint64_t)nb * 520);
  a.ops = (uint8_t*)(ws + (((uint64_t)nb * 524 + 255) & ~(uint64_t)255));
  a.out = d.out; a.status = d.status;
  dcb_parse64_kernel<CT, false><<<nb, DSP64_NT, 0, st>>>(a);
  dcb_parse64_chain_kernel<<<1, 256, 0, st>>>(a);
  dcb_parse64_kernel<CT, true><<<nb, DSP64_NT, 0, st>>>(a);
  if (CT != CT_NP) {
    uint4* masks = (uint4*)(a.ops + ds_ops_bytes(d.n));
    ResolveArgs r; r.ops = a.ops; r.out = nullptr; r.n = d.n; r.masks = masks;
    const uint64_t groups = (d.n + 31) / 32 + 3;
    dcb_opmask_kernel<<<(unsigned)((groups + 127) / 128), 128, 0, st>>>(r, groups);
    Resolve64Args q; q.ops = a.ops; q.masks = masks; q.out = d.out; q.n = d.n;
    const uint64_t threads = (d.n + 31) / 32;
    dcb_resolve64_kernel<<<(unsigned)((threads + 127) / 128), 128, 0, st>>>(q);
  }
  if ((d.flags & DCB_F_SUBTRACT_MIN) && d.cmin) {
    const uint64_t blocks = (d.n + 1023) / 1024;
    dcb_addmin64_kernel<<<(unsigned)(blocks < 4096 ? blocks : 4096), 256, 0, st>>>(d.out, d.n, d.cmin);
  }
  return cudaGetLastError();
}
