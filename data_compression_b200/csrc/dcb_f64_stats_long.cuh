// dcb_f64_stats_long.cuh -- toSmallDataset_double (:3522) / med_dataset_double (:3564) for ONE LONG chunk (the single-call shape
// of k-means / MM / LU).  dcb_f64_stats_kernel is one block per chunk: for a 2^20-element call that one block was 0.35 of the
// call's 0.41 ms, and its exact fallback (one thread re-adding from global memory) 22 ms.  Here the chunk is cut into parts:
//   min_part  : per part, the first minimum under the reference's '<' scan (value + index)
//   min_final : one block combines the parts and applies the data[0] rule -> cmin
//   sum_part  : per part, max / tree sum / sum of magnitudes of y = x - min
//   final     : one block: type from the max, the 20 mask bits from the bracket around the tree sum -- or, when the bracket
//               straddles a mask boundary (or the exact mean is asked for), the sequential sum with the values staged through
//               shared memory by the whole block and added by thread 0 in index order.
// Same rules, same results as dcb_f64_stats_kernel (the bracket n 2^-52 sum|y| holds for any summation order).
#pragma once
#include "dcb_f64.cuh"

#define F64L_MAX_PARTS 4096
#define F64L_STAGE 2048              // doubles staged per step of the exact sum

struct F64LongArgs {
  F64StatsArgs s;                    // chunk == n: one chunk
  uint32_t parts; uint64_t per;      // elements per part (a multiple of F64_NT)
  double* pv; unsigned long long* pi;              // [parts] partial minima
  double* pmx; double* psum; double* pmag;         // [parts]
};

__global__ void __launch_bounds__(F64_NT) dcb_f64l_min_part_kernel(F64LongArgs a) {
  __shared__ double sh_v[F64_NT / 32]; __shared__ unsigned long long sh_i[F64_NT / 32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint64_t lo = max((uint64_t)blockIdx.x * a.per, (uint64_t)1), hi = min(a.s.n, ((uint64_t)blockIdx.x + 1) * a.per);
  const double* __restrict__ x = a.s.in;
  double bv = __longlong_as_double(0x7FF8000000000000ll); uint64_t bi = ~0ull;
  for (uint64_t i = lo + tid; i < hi; i += F64_NT) { const double v = x[i]; if (f64_min_better(v, i, bv, bi)) { bv = v; bi = i; } }
  for (int o = 16; o; o >>= 1) {
    const double ov = __shfl_xor_sync(0xffffffffu, bv, o); const uint64_t oi = __shfl_xor_sync(0xffffffffu, bi, o);
    if (f64_min_better(ov, oi, bv, bi)) { bv = ov; bi = oi; }
  }
  if (lane == 0) { sh_v[warp] = bv; sh_i[warp] = bi; }
  __syncthreads();
  if (tid == 0) {
    for (int w = 1; w < F64_NT / 32; w++) if (f64_min_better(sh_v[w], sh_i[w], bv, bi)) { bv = sh_v[w]; bi = sh_i[w]; }
    a.pv[blockIdx.x] = bv; a.pi[blockIdx.x] = bi;
  }
}

__global__ void __launch_bounds__(F64_NT) dcb_f64l_min_final_kernel(F64LongArgs a) {
  __shared__ double sh_v[F64_NT / 32]; __shared__ unsigned long long sh_i[F64_NT / 32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  double bv = __longlong_as_double(0x7FF8000000000000ll); uint64_t bi = ~0ull;
  for (uint32_t p = tid; p < a.parts; p += F64_NT) if (f64_min_better(a.pv[p], a.pi[p], bv, bi)) { bv = a.pv[p]; bi = a.pi[p]; }
  for (int o = 16; o; o >>= 1) {
    const double ov = __shfl_xor_sync(0xffffffffu, bv, o); const uint64_t oi = __shfl_xor_sync(0xffffffffu, bi, o);
    if (f64_min_better(ov, oi, bv, bi)) { bv = ov; bi = oi; }
  }
  if (lane == 0) { sh_v[warp] = bv; sh_i[warp] = bi; }
  __syncthreads();
  if (tid == 0) {
    double m = a.s.in[0];                                 // data[0]; a NaN here stays (every '<' is false)
    if (m == m) {
      uint64_t mi = 0;
      for (int w = 0; w < F64_NT / 32; w++) if (f64_min_better(sh_v[w], sh_i[w], m, mi)) { m = sh_v[w]; mi = sh_i[w]; }
    }
    a.s.cmin[0] = m;
  }
}

__global__ void __launch_bounds__(F64_NT) dcb_f64l_sum_part_kernel(F64LongArgs a) {
  __shared__ double sh_v[F64_NT / 32], sh_a[F64_NT / 32], sh_b[F64_NT / 32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint64_t lo = (uint64_t)blockIdx.x * a.per, hi = min(a.s.n, lo + a.per);
  const double* __restrict__ x = a.s.in;
  const bool sub = a.s.subtract_min != 0;
  const double mn = sub ? a.s.cmin[0] : 0.0;
  double mx = __longlong_as_double(0x7FF8000000000000ll), sum = 0.0, mag = 0.0;
  for (uint64_t i = lo + tid; i < hi; i += F64_NT) {
    const double y = sub ? __dsub_rn(x[i], mn) : x[i];
    sum = __dadd_rn(sum, y); mag = __dadd_rn(mag, fabs(y));
    if (y == y && (mx != mx || y > mx)) mx = y;
  }
  for (int o = 16; o; o >>= 1) {
    sum = __dadd_rn(sum, __shfl_xor_sync(0xffffffffu, sum, o)); mag = __dadd_rn(mag, __shfl_xor_sync(0xffffffffu, mag, o));
    const double om = __shfl_xor_sync(0xffffffffu, mx, o);
    if (om == om && (mx != mx || om > mx)) mx = om;
  }
  if (lane == 0) { sh_v[warp] = mx; sh_a[warp] = sum; sh_b[warp] = mag; }
  __syncthreads();
  if (tid == 0) {
    double M = sh_v[0], S = sh_a[0], G = sh_b[0];
    for (int w = 1; w < F64_NT / 32; w++) {
      S = __dadd_rn(S, sh_a[w]); G = __dadd_rn(G, sh_b[w]);
      if (sh_v[w] == sh_v[w] && (M != M || sh_v[w] > M)) M = sh_v[w];
    }
    a.pmx[blockIdx.x] = M; a.psum[blockIdx.x] = S; a.pmag[blockIdx.x] = G;
  }
}

__global__ void __launch_bounds__(F64_NT) dcb_f64l_final_kernel(F64LongArgs a) {
  __shared__ double buf[F64L_STAGE];
  __shared__ int sh_exact;
  const int tid = threadIdx.x;
  const double* __restrict__ x = a.s.in;
  const bool sub = a.s.subtract_min != 0;
  const double mn = sub ? a.s.cmin[0] : 0.0;
  const double cnt = (double)(long long)a.s.n;
  if (tid == 0) {
    double M = a.pmx[0], S = a.psum[0], G = a.pmag[0];
    for (uint32_t p = 1; p < a.parts; p++) {
      S = __dadd_rn(S, a.psum[p]); G = __dadd_rn(G, a.pmag[p]);
      const double v = a.pmx[p];
      if (v == v && (M != M || v > M)) M = v;
    }
    const double y0 = sub ? __dsub_rn(x[0], mn) : x[0];
    if (y0 != y0) M = y0;                                 // max = data[0] = NaN: every later '>' is false
    int type = 0, add = 0;                                // :3576-3585
    for (int i = 10; i > 0; i--) {
      add += 1 << i;
      if (M < __longlong_as_double((long long)add << 52)) { type = 11 - i; break; }
    }
    a.s.ctype[0] = type;
    // (n + parts) 2^-52 sum|y| >= |sequential - this tree|: n - 1 roundings each, plus the partials' own
    const double E = __dmul_rn(__dmul_rn(__dadd_rn(cnt, (double)a.parts), 2.220446049250313e-16), G);
    const double mlo = __ddiv_rn(__dsub_rn(S, E), cnt), mhi = __ddiv_rn(__dadd_rn(S, E), cnt);
    const uint32_t klo = (uint32_t)((uint64_t)__double_as_longlong(mlo) >> 44), khi = (uint32_t)((uint64_t)__double_as_longlong(mhi) >> 44);
    const bool decided = (klo == khi) && (G == G) && (G < 1.0e300) && a.s.cmedium == nullptr;
    if (decided) a.s.cmask[0] = klo;
    sh_exact = decided ? 0 : 1;
  }
  __syncthreads();
  if (!sh_exact) return;
  // exact: total += data[i] in index order (:3568-3571); the block stages, thread 0 adds
  double total = 0.0;
  for (uint64_t t0 = 0; t0 < a.s.n; t0 += F64L_STAGE) {
    const uint32_t c = (uint32_t)min((uint64_t)F64L_STAGE, a.s.n - t0);
    for (uint32_t e = tid; e < c; e += F64_NT) buf[e] = sub ? __dsub_rn(x[t0 + e], mn) : x[t0 + e];
    __syncthreads();
    if (tid == 0) {
      uint32_t e = 0;
      for (; e + 8 <= c; e += 8) {
        const double v0 = buf[e], v1 = buf[e + 1], v2 = buf[e + 2], v3 = buf[e + 3], v4 = buf[e + 4], v5 = buf[e + 5], v6 = buf[e + 6], v7 = buf[e + 7];
        total = __dadd_rn(total, v0); total = __dadd_rn(total, v1); total = __dadd_rn(total, v2); total = __dadd_rn(total, v3);
        total = __dadd_rn(total, v4); total = __dadd_rn(total, v5); total = __dadd_rn(total, v6); total = __dadd_rn(total, v7);
      }
      for (; e < c; e++) total = __dadd_rn(total, buf[e]);
    }
    __syncthreads();
  }
  if (tid == 0) {
    const double med = __ddiv_rn(total, cnt);
    a.s.cmask[0] = (uint32_t)((uint64_t)__double_as_longlong(med) >> 44);
    if (a.s.cmedium) a.s.cmedium[0] = med;
  }
}

static inline uint64_t f64l_scratch_bytes() { return (uint64_t)F64L_MAX_PARTS * 5 * 8 + 256; }
// launches the passes on `st`; d_scratch: f64l_scratch_bytes() bytes
static cudaError_t f64l_launch(const F64StatsArgs& s, void* d_scratch, cudaStream_t st) {
  F64LongArgs a;
  a.s = s;
  uint64_t per = (s.n + F64L_MAX_PARTS - 1) / F64L_MAX_PARTS;
  per = (per + 8191) & ~(uint64_t)8191;                                 // at least 8192 elements per part
  a.per = per;
  a.parts = (uint32_t)((s.n + per - 1) / per);
  double* p = (double*)d_scratch;
  a.pv = p; a.pi = (unsigned long long*)(p + F64L_MAX_PARTS); a.pmx = p + 2 * F64L_MAX_PARTS; a.psum = p + 3 * F64L_MAX_PARTS;
  a.pmag = p + 4 * F64L_MAX_PARTS;
  if (s.subtract_min) {
    dcb_f64l_min_part_kernel<<<a.parts, F64_NT, 0, st>>>(a);
    dcb_f64l_min_final_kernel<<<1, F64_NT, 0, st>>>(a);
  }
  if (s.want_mask) {
    dcb_f64l_sum_part_kernel<<<a.parts, F64_NT, 0, st>>>(a);
    dcb_f64l_final_kernel<<<1, F64_NT, 0, st>>>(a);
  }
  return cudaGetLastError();
}
