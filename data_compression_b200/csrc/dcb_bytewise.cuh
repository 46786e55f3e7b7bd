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
    c = i / a.chunk;
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
    if (anym) atomicOr(&a.chunk_slow[(uint64_t)blockIdx.x * DCB_BW_NT / a.chunk], 1u);     // a tile lies inside one chunk
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
  const uint64_t c = i / a.chunk, base = c * a.chunk;
  const uint32_t r = (uint32_t)(i - base);
  if (f) { a.ac[base + before] = f; a.ad[base + before] = (int32_t)(r + 1u); }             // displacement: 1-based position (:4077)
  else a.af[base + (r - before)] = __ldg(a.in + i);
}

template <typename T>
struct BwDecArgs {
  const T* af; const uint8_t* ac; const int32_t* ad; const uint32_t* chunk_nfloat;
  uint64_t n; uint64_t chunk; uint64_t K; T* out;
};
// myDecompress :3942-3975: closed loop on the reconstructed values, one thread per chunk
template <typename T>
__global__ void dcb_bw_decode_kernel(const BwDecArgs<T> a) {
  typedef BwFp<T> F;
  const uint64_t c = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= a.K) return;
  const uint64_t base = c * a.chunk, nc = min(a.chunk, a.n - base);
  const uint32_t nhit = (uint32_t)nc - a.chunk_nfloat[c];
  const T* af = a.af + base; const uint8_t* ac = a.ac + base; const int32_t* ad = a.ad + base;
  T* out = a.out + base;
  uint32_t fp = 0, cp = 0;
  T o1 = 0, o2 = 0, o3 = 0, o4 = 0;
  int32_t next = nhit ? ad[0] - 1 : -1;
  for (uint32_t r = 0; r < (uint32_t)nc; r++) {
    T v;
    if ((int32_t)r == next) {
      const uint8_t t = ac[cp++];
      v = 0;
      if (t == 'a') v = o1;
      else if (t == 'b') v = F::sub(F::mul((T)2, o1), o2);
      else if (t == 'c') v = F::add(F::sub(F::mul((T)3, o1), F::mul((T)3, o2)), o3);
      else if (t == 'd') v = F::sub(F::add(F::sub(F::mul((T)4, o1), F::mul((T)6, o2)), F::mul((T)4, o3)), o4);
      next = (cp < nhit) ? ad[cp] - 1 : -1;
    } else {
      v = af[fp++];
    }
    out[r] = v;
    o4 = o3; o3 = o2; o2 = o1; o1 = v;
  }
}
