// dcb_bytewise.cuh -- the byte-wise codec CT=1 (second half): myCompress (impl/dataCompression.c:3977-4118)
// and myDecompress (:3942-3975).  A value either hits one of four open-loop predictors on the ORIGINAL history (one char
// 'a'..'d' plus its 1-based position) or is kept as a raw float; one CHUNK == one reference call (the history restarts).
//
// The encoder is open loop, so every element is classified on its own from x[i-1..i-4] -- unless the chunk contains the
// value -1.0f, the reference's "empty history slot" marker (its state machine then stalls and refills slots):
// such chunks are redone by one thread with the state machine verbatim.  Three passes:
//   classify : one thread per element -> flag byte (0 raw, 'a'..'d'), hits per 256-element tile, "-1.0f seen" per chunk
//   offsets  : one block per chunk: exclusive scan of its tile counts, raw-float count of the chunk
//   scatter  : one thread per element: rank inside the tile by ballots -> array_float / array_char / displacement slots
// Outputs are fixed pitch: chunk c owns [c*chunk, c*chunk + n_c) of each of the three arrays.
// The decoder is closed loop (data[i] from the RECONSTRUCTED data[i-1..i-4]): one thread per chunk.
#pragma once
#include "dcb_common.cuh"

#define DCB_BW_NT 256

// every operation rounded on its own, never contracted; float32 (myCompress :3977) and float64 (myCompress_double :3815)
template <typename T> struct BwFp;
template <> struct BwFp<float> {
  static __device__ __forceinline__ float sub(float a, float b) { return __fsub_rn(a, b); }
  static __device__ __forceinline__ float add(float a, float b) { return __fadd_rn(a, b); }
  static __device__ __forceinline__ float mul(float a, float b) { return __fmul_rn(a, b); }
  static __device__ __forceinline__ float abs(float a) { return fabsf(a); }
};
template <> struct BwFp<double> {
  static __device__ __forceinline__ double sub(double a, double b) { return __dsub_rn(a, b); }
  static __device__ __forceinline__ double add(double a, double b) { return __dadd_rn(a, b); }
  static __device__ __forceinline__ double mul(double a, double b) { return __dmul_rn(a, b); }
  static __device__ __forceinline__ double abs(double a) { return fabs(a); }
};

template <typename T>
struct BwArgs {
  const T* in; uint64_t n; uint64_t chunk;              // chunk: elements per chunk (multiple of 256) or n for a single chunk
  T tp;                                                 // float32: largest float <= absErrorBound (the compare is in double, :4068); float64: the bound
  uint32_t tiles_per_chunk;                             // chunk / 256, or 0 for a single chunk (a tile never straddles chunks)
  uint8_t* flags;                                       // [n]
  uint32_t* tile_hits;                                  // [tiles]  hits per tile, then (offsets pass) exclusive prefix inside the chunk
  uint32_t* chunk_slow;                                 // [K]
  uint32_t* chunk_nfloat;                               // [K]
  T* af; uint8_t* ac; int32_t* ad;                      // [n] each, fixed pitch per chunk
};

template <typename T>
__device__ __forceinline__ uint8_t bw_classify(T b1, T b2, T b3, T b4, T v, T tp) {
  typedef BwFp<T> F;
  const T p2 = F::sub(F::mul((T)2, b1), b2);                                                // :4031
  const T p3 = F::add(F::sub(F::mul((T)3, b1), F::mul((T)3, b2)), b3);                       // :4032
  const T p4 = F::sub(F::add(F::sub(F::mul((T)4, b1), F::mul((T)6, b2)), F::mul((T)4, b3)), b4);   // :4033
  const T d1 = F::abs(F::sub(b1, v)), d2 = F::abs(F::sub(p2, v)), d3 = F::abs(F::sub(p3, v)), d4 = F::abs(F::sub(p4, v));
  T dm = d1; uint8_t t = 'a';
  if (d2 < dm) { dm = d2; t = 'b'; }                                                       // strict '<': NaN never wins (:4041-4058)
  if (d3 < dm) { dm = d3; t = 'c'; }
  if (d4 < dm) { dm = d4; t = 'd'; }
  return (dm <= tp) ? t : (uint8_t)0;
}

template <typename T>
__global__ void __launch_bounds__(DCB_BW_NT) dcb_bw_classify_kernel(const BwArgs<T> a) {
  const uint64_t i = (uint64_t)blockIdx.x * DCB_BW_NT + threadIdx.x;
  uint8_t f = 0; bool marker = false;
  uint64_t c = 0;
  if (i < a.n) {
    c = a.tiles_per_chunk ? blockIdx.x / a.tiles_per_chunk : 0u;             // (no 64-bit division per element)
    const uint64_t r = i - c * a.chunk;                 // index inside the chunk
    const T v = __ldg(a.in + i);
    marker = (v == (T)-1);
    if (r >= 4) f = bw_classify(__ldg(a.in + i - 1), __ldg(a.in + i - 2), __ldg(a.in + i - 3), __ldg(a.in + i - 4), v, a.tp);
    a.flags[i] = f;
  }
  const int hits = __syncthreads_count(f != 0);
  const int anym = __syncthreads_or(marker);
  if (threadIdx.x == 0) {
    a.tile_hits[blockIdx.x] = (uint32_t)hits;
    if (anym) atomicOr(&a.chunk_slow[a.tiles_per_chunk ? blockIdx.x / a.tiles_per_chunk : 0u], 1u);   // a tile lies inside one chunk
  }
}

// chunks that contain -1.0f: the reference's state machine, one thread per chunk (flags and tile counts rewritten)
template <typename T>
__global__ void dcb_bw_slow_kernel(const BwArgs<T> a, uint64_t K) {
  const uint64_t c = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= K || !a.chunk_slow[c]) return;
  const uint64_t base = c * a.chunk, nc = min(a.chunk, a.n - base);
  const T M = (T)-1;
  T b1 = M, b2 = M, b3 = M, b4 = M;
  uint32_t hits = 0;
  for (uint64_t r = 0; r < nc; r++) {
    const T v = a.in[base + r];
    uint8_t f = 0;
    if (b4 == M || b3 == M || b2 == M || b1 == M) {                                       // :3996-4027
      if (b4 == M) b4 = v; else if (b3 == M) b3 = v; else if (b2 == M) b2 = v; else b1 = v;
    } else {
      f = bw_classify(b1, b2, b3, b4, v, a.tp);
      b4 = b3; b3 = b2; b2 = b1; b1 = v;
    }
    a.flags[base + r] = f;
    hits += f != 0;
    if (((r + 1) % DCB_BW_NT) == 0 || r + 1 == nc) { a.tile_hits[(base + r) / DCB_BW_NT] = hits; hits = 0; }
  }
}

// one block per chunk: tile_hits -> exclusive prefix inside the chunk; raw-float count of the chunk
template <typename T>
__global__ void __launch_bounds__(DCB_BW_NT) dcb_bw_offsets_kernel(const BwArgs<T> a) {
  __shared__ uint32_t s_w[DCB_BW_NT / 32];
  __shared__ uint32_t s_run;
  const uint64_t c = blockIdx.x;
  const uint64_t base = c * a.chunk, nc = min(a.chunk, a.n - base);
  const uint64_t t0 = base / DCB_BW_NT, nt = (nc + DCB_BW_NT - 1) / DCB_BW_NT;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) s_run = 0u;
  __syncthreads();
  for (uint64_t k0 = 0; k0 < nt; k0 += DCB_BW_NT) {
    const uint64_t k = k0 + threadIdx.x;
    const uint32_t v = (k < nt) ? a.tile_hits[t0 + k] : 0u;
    uint32_t inc = v;
    #pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += t; }
    if (lane == 31) s_w[warp] = inc;
    __syncthreads();
    uint32_t wb = 0u, tot = 0u;
    for (int w = 0; w < DCB_BW_NT / 32; w++) { if (w < warp) wb += s_w[w]; tot += s_w[w]; }
    const uint32_t run = s_run;
    if (k < nt) a.tile_hits[t0 + k] = run + wb + inc - v;
    __syncthreads();
    if (threadIdx.x == 0) s_run = run + tot;
    __syncthreads();
  }
  if (threadIdx.x == 0) a.chunk_nfloat[c] = (uint32_t)nc - s_run;
}

template <typename T>
__global__ void __launch_bounds__(DCB_BW_NT) dcb_bw_scatter_kernel(const BwArgs<T> a) {
  __shared__ uint32_t s_w[DCB_BW_NT / 32];
  const uint64_t i = (uint64_t)blockIdx.x * DCB_BW_NT + threadIdx.x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint8_t f = (i < a.n) ? a.flags[i] : (uint8_t)0;
  const unsigned m = __ballot_sync(0xffffffffu, f != 0);
  if (lane == 0) s_w[warp] = (uint32_t)__popc(m);
  __syncthreads();
  if (i >= a.n) return;
  uint32_t before = a.tile_hits[blockIdx.x] + (uint32_t)__popc(m & ((1u << lane) - 1u));    // hits of the chunk before this element
  for (int w = 0; w < warp; w++) before += s_w[w];
  const uint64_t c = a.tiles_per_chunk ? blockIdx.x / a.tiles_per_chunk : 0u, base = c * a.chunk;
  const uint32_t r = (uint32_t)(i - base);
  if (f) { a.ac[base + before] = f; a.ad[base + before] = (int32_t)(r + 1u); }             // displacement: 1-based position (:4077)
  else a.af[base + (r - before)] = __ldg(a.in + i);
}

// Chunks of a few tiles (the default 4096 elements): ONE kernel, one block per chunk.  The block walks the chunk's 256-element
// tiles in order with the number of hits so far in a register, so an element's output slot is known when it is classified:
// no flag bytes, no tile counts, no second read of the input (the three passes above move 4n + n, then n + 4n + 5n bytes; this
// one 4n + 5n).  Up to the chunk's first -1.0f every element sees a clean history, so the parallel classification holds up to
// and including that element; thread 0 then finishes the chunk with the state machine (:3996-4027), appending in order.
template <typename T>
__global__ void __launch_bounds__(DCB_BW_NT) dcb_bw_fused_kernel(const BwArgs<T> a) {
  constexpr int NW = DCB_BW_NT / 32, TPG = 16;            // a group = 16 tiles = 4096 elements: all its loads in flight at once
  __shared__ uint32_t s_cnt[TPG * NW];                   // hits per (tile, warp), then their exclusive prefix
  __shared__ uint32_t s_tot, s_first;
  const uint64_t c = blockIdx.x, base = c * a.chunk;
  const uint32_t nc = (uint32_t)min(a.chunk, a.n - base);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const T* __restrict__ in = a.in + base;
  T* __restrict__ af = a.af + base; uint8_t* __restrict__ ac = a.ac + base; int32_t* __restrict__ ad = a.ad + base;
  const T M = (T)-1;
  uint32_t hits = 0u;                                    // hits of the chunk before the current group (block uniform)
  uint32_t g0 = 0u;
  if (tid == 0) s_first = 0xFFFFFFFFu;
  __syncthreads();
  for (; g0 < nc; g0 += TPG * DCB_BW_NT) {
    T v[TPG];
    #pragma unroll
    for (int k = 0; k < TPG; k++) {
      const uint32_t r = g0 + (uint32_t)(k * DCB_BW_NT + tid);
      v[k] = (r < nc) ? __ldg(in + r) : (T)0;
    }
    uint32_t fl[TPG / 4], rk[TPG / 4];                    // flag byte and rank inside the warp (hits before this lane), per tile
    #pragma unroll
    for (int k = 0; k < TPG / 4; k++) { fl[k] = 0u; rk[k] = 0u; }
    bool marker = false;
    #pragma unroll
    for (int k = 0; k < TPG; k++) {
      const uint32_t r = g0 + (uint32_t)(k * DCB_BW_NT + tid);
      const bool valid = r < nc;
      uint8_t f = 0;
      if (valid && r >= 4u) f = bw_classify(__ldg(in + r - 1), __ldg(in + r - 2), __ldg(in + r - 3), __ldg(in + r - 4), v[k], a.tp);
      marker = marker || (valid && v[k] == M);
      const unsigned hm = __ballot_sync(0xffffffffu, f != 0);
      if (lane == 0) s_cnt[k * NW + warp] = (uint32_t)__popc(hm);
      fl[k >> 2] |= (uint32_t)f << (8 * (k & 3));
      rk[k >> 2] |= (uint32_t)__popc(hm & ((1u << lane) - 1u)) << (8 * (k & 3));
    }
    if (marker) atomicMin(&s_first, g0);                  // this group (and the rest of the chunk) goes to the state machine
    __syncthreads();
    if (warp == 0) {                                      // exclusive prefix of the 128 counts: four per lane
      uint32_t c4[4], t = 0u;
      #pragma unroll
      for (int q = 0; q < 4; q++) { c4[q] = s_cnt[lane * 4 + q]; t += c4[q]; }
      uint32_t inc = t;
      #pragma unroll
      for (int o = 1; o < 32; o <<= 1) { const uint32_t u = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += u; }
      uint32_t run = inc - t;
      #pragma unroll
      for (int q = 0; q < 4; q++) { s_cnt[lane * 4 + q] = run; run += c4[q]; }
      if (lane == 31) s_tot = inc;
    }
    __syncthreads();
    if (s_first != 0xFFFFFFFFu) break;                    // block uniform
    #pragma unroll
    for (int k = 0; k < TPG; k++) {
      const uint32_t r = g0 + (uint32_t)(k * DCB_BW_NT + tid);
      const uint32_t f = (fl[k >> 2] >> (8 * (k & 3))) & 0xFFu;
      const uint32_t before = hits + s_cnt[k * NW + warp] + ((rk[k >> 2] >> (8 * (k & 3))) & 0xFFu);
      if (r < nc) {
        if (f) { ac[before] = (uint8_t)f; ad[before] = (int32_t)(r + 1u); }                // displacement: 1-based position (:4077)
        else af[r - before] = v[k];
      }
    }
    hits += s_tot;
    __syncthreads();                                      // s_cnt / s_tot are rewritten by the next group
  }
  if (g0 >= nc) {
    if (tid == 0) a.chunk_nfloat[c] = nc - hits;
    return;
  }
  if (tid != 0) return;
  // thread 0: the group with the chunk's first -1.0f and everything after it, by the state machine, appending in order.  The
  // history before the group is clean (no marker yet), so the machine starts in its steady state -- except in the first group,
  // where it starts empty as the reference does.
  T b1 = M, b2 = M, b3 = M, b4 = M;
  if (g0 >= 4u) { b1 = in[g0 - 1]; b2 = in[g0 - 2]; b3 = in[g0 - 3]; b4 = in[g0 - 4]; }
  for (uint32_t r = g0; r < nc; r++) {
    const T v = in[r];
    uint8_t f = 0;
    if (b4 == M || b3 == M || b2 == M || b1 == M) {
      if (b4 == M) b4 = v; else if (b3 == M) b3 = v; else if (b2 == M) b2 = v; else b1 = v;
    } else {
      f = bw_classify(b1, b2, b3, b4, v, a.tp);
      b4 = b3; b3 = b2; b2 = b1; b1 = v;
    }
    if (f) { ac[hits] = f; ad[hits] = (int32_t)(r + 1u); hits++; }
    else af[r - hits] = v;
  }
  a.chunk_nfloat[c] = nc - hits;
}

template <typename T>
struct BwDecArgs {
  const T* af; const uint8_t* ac; const int32_t* ad; const uint32_t* chunk_nfloat;
  uint64_t n; uint64_t chunk; uint64_t K; T* out;
};
// myDecompress :3942-3975: closed loop on the reconstructed values, one thread per chunk.
// A lane walks its own chunk, so every access of a warp touches 32 different sectors: with one scalar access per element and
// stream the pass was bound by those wavefronts (3.3 ms for 2^28 floats).  VEC: the three input streams are read through
// 16-byte register windows (4 displacements, 16 chars, 4 floats / 2 doubles per load) and the output leaves as 16-byte
// stores -- legal when the chunk's arrays are 16-byte aligned and its length is a multiple of 16 (every chunk of a chunked
// call; a ragged or single chunk takes the scalar loop).
// 16-byte read-only load that asks L2 for the whole 128-byte line: the next seven windows of the lane's stream then come
// from L2 instead of DRAM (the lane consumes them in order)
__device__ __forceinline__ uint4 bw_ld16(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L2::128B.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}
template <typename V> __device__ __forceinline__ auto bw_sel4(const V& w, uint32_t i) -> decltype(w.x) {
  return i == 0u ? w.x : (i == 1u ? w.y : (i == 2u ? w.z : w.w));
}
template <typename T, bool VEC>
__device__ __forceinline__ void bw_decode_chunk(const T* __restrict__ af, const uint8_t* __restrict__ ac,
                                                const int32_t* __restrict__ ad, T* __restrict__ out, uint32_t nc, uint32_t nhit) {
  typedef BwFp<T> F;
  constexpr uint32_t FW = 16 / sizeof(T);                 // floats per 16-byte window
  uint32_t fp = 0, cp = 0;
  T o1 = 0, o2 = 0, o3 = 0, o4 = 0;
  int4 adw = make_int4(0, 0, 0, 0); uint4 acw = make_uint4(0, 0, 0, 0);
  T afw[FW];
  #pragma unroll
  for (uint32_t q = 0; q < FW; q++) afw[q] = 0;
  auto disp = [&](uint32_t i) -> int32_t {                // ad[i], i == cp after an advance
    if (!VEC) return ad[i];
    if ((i & 3u) == 0u) { const uint4 w = bw_ld16(ad + i); adw = make_int4((int)w.x, (int)w.y, (int)w.z, (int)w.w); }
    return bw_sel4(adw, i & 3u);
  };
  int32_t next = nhit ? disp(0) - 1 : -1;
  for (uint32_t r4 = 0; r4 < nc; r4 += 4) {
    T ob[4];
    #pragma unroll
    for (uint32_t u = 0; u < 4; u++) {
      const uint32_t r = r4 + u;
      T v = 0;
      if (VEC || r < nc) {
        if ((int32_t)r == next) {
          uint32_t t;
          if (VEC) {
            if ((cp & 15u) == 0u) acw = bw_ld16(ac + cp);
            t = (bw_sel4(acw, (cp >> 2) & 3u) >> (8u * (cp & 3u))) & 0xFFu;
          } else {
            t = ac[cp];
          }
          cp++;
          if (t == 'a') v = o1;
          else if (t == 'b') v = F::sub(F::mul((T)2, o1), o2);
          else if (t == 'c') v = F::add(F::sub(F::mul((T)3, o1), F::mul((T)3, o2)), o3);
          else if (t == 'd') v = F::sub(F::add(F::sub(F::mul((T)4, o1), F::mul((T)6, o2)), F::mul((T)4, o3)), o4);
          next = (cp < nhit) ? disp(cp) - 1 : -1;
        } else {
          if (VEC) {
            if ((fp % FW) == 0u) {
              if (sizeof(T) == 4) {
                const uint4 w = bw_ld16(af + fp);
                afw[0] = (T)__uint_as_float(w.x); afw[1] = (T)__uint_as_float(w.y);
                afw[FW - 2] = (T)__uint_as_float(w.z); afw[FW - 1] = (T)__uint_as_float(w.w);
              } else {
                const uint4 w = bw_ld16(af + fp);
                afw[0] = (T)__hiloint2double((int)w.y, (int)w.x); afw[FW - 1] = (T)__hiloint2double((int)w.w, (int)w.z);
              }
            }
            if (FW == 4) v = (fp & 3u) == 0u ? afw[0] : ((fp & 3u) == 1u ? afw[1] : ((fp & 3u) == 2u ? afw[FW - 2] : afw[FW - 1]));
            else v = (fp & 1u) ? afw[FW - 1] : afw[0];
          } else {
            v = af[fp];
          }
          fp++;
        }
        if (!VEC) out[r] = v;
        o4 = o3; o3 = o2; o2 = o1; o1 = v;
      }
      ob[u] = v;
    }
    if (VEC) {
      if (sizeof(T) == 4) {
        *reinterpret_cast<float4*>(out + r4) = make_float4((float)ob[0], (float)ob[1], (float)ob[2], (float)ob[3]);
      } else {
        *reinterpret_cast<double2*>(out + r4) = make_double2((double)ob[0], (double)ob[1]);
        *reinterpret_cast<double2*>(out + r4 + 2) = make_double2((double)ob[2], (double)ob[3]);
      }
    }
  }
}
template <typename T>
__global__ void dcb_bw_decode_kernel(const BwDecArgs<T> a) {
  const uint64_t c = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= a.K) return;
  const uint64_t base = c * a.chunk, nc = min(a.chunk, a.n - base);
  const uint32_t nhit = (uint32_t)nc - a.chunk_nfloat[c];
  const T* af = a.af + base; const uint8_t* ac = a.ac + base; const int32_t* ad = a.ad + base;
  T* out = a.out + base;
  const bool vec = (nc & 15u) == 0u && ((reinterpret_cast<uintptr_t>(af) | reinterpret_cast<uintptr_t>(ac) |
                                        reinterpret_cast<uintptr_t>(ad) | reinterpret_cast<uintptr_t>(out)) & 15u) == 0u;
  if (vec) bw_decode_chunk<T, true>(af, ac, ad, out, (uint32_t)nc, nhit);
  else bw_decode_chunk<T, false>(af, ac, ad, out, (uint32_t)nc, nhit);
}
