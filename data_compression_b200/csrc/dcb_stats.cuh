// dcb_stats.cuh -- per-chunk toSmallDataset_float minimum (dataCompression.c:3543-3562) and
// med_dataset_float medium / type (:3593-3620) for chunks that span several tiles (the encoder fuses
// both when a chunk is one tile).  One block per chunk.
#pragma once
#include "dcb_common.cuh"
#include "dcb_encode_warp.cuh"          // exact_medium: the sequential float32 sum without the n-step dependent chain

struct StatsArgs {
  const float* in;
  uint64_t n;
  uint64_t chunk;        // elements per chunk (>= n: single chunk)
  int want_medium;
  int subtract_min;      // medium/max are taken over x - min (the caller compresses the min-subtracted data)
  float* cmin;
  float* cmed;
  int32_t* ctype;
};

__global__ void __launch_bounds__(256) dcb_stats_kernel(const StatsArgs a) {
  __shared__ float redf[8];
  __shared__ unsigned long long redi[8];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint64_t c = blockIdx.x;
  const uint64_t lo = c * a.chunk;
  const uint64_t cn = min(a.chunk, a.n - lo);
  const float* x = a.in + lo;

  float mn = 0.0f;
  if (a.subtract_min) {
    float lm = __uint_as_float(0x7f800000u);
    for (uint64_t i = tid; i < cn; i += 256) lm = fminf(lm, __ldg(x + i));
    #pragma unroll
    for (int o = 16; o; o >>= 1) lm = fminf(lm, __shfl_xor_sync(0xffffffffu, lm, o));
    if (lane == 0) redf[warp] = lm;
    __syncthreads();
    mn = redf[0];
    #pragma unroll
    for (int w = 1; w < 8; w++) mn = fminf(mn, redf[w]);
    __syncthreads();
    if (mn == 0.0f) {   // first minimal element decides between +0 and -0
      unsigned long long fi = ~0ull;
      for (uint64_t i = tid; i < cn; i += 256) if (__ldg(x + i) == 0.0f) { fi = i; break; }
      #pragma unroll
      for (int o = 16; o; o >>= 1) fi = min(fi, __shfl_xor_sync(0xffffffffu, fi, o));
      if (lane == 0) redi[warp] = fi;
      __syncthreads();
      fi = redi[0];
      #pragma unroll
      for (int w = 1; w < 8; w++) fi = min(fi, redi[w]);
      mn = __ldg(x + fi);
    }
    const float x0 = __ldg(x);
    if (x0 != x0) mn = x0;            // data[0] NaN: every '<' fails, min stays NaN (:3546-3553)
    if (tid == 0) a.cmin[c] = mn;
  } else if (a.cmin && tid == 0) {
    a.cmin[c] = 0.0f;
  }
  if (!a.want_medium) return;

  // max, and whether the chunk leaves the domain of the fast exact sum (a negative value, NaN, Inf, |v| >= 2^127)
  float lmx = __uint_as_float(0xff800000u);
  bool odd = false;
  for (uint64_t i = tid; i < cn; i += 256) {
    float v = __ldg(x + i);
    if (a.subtract_min) v = __fsub_rn(v, mn);
    lmx = fmaxf(lmx, v);
    odd = odd || !(fabsf(v) < __uint_as_float(0x7f000000u)) || (f2u(v) >> 31);
  }
  #pragma unroll
  for (int o = 16; o; o >>= 1) lmx = fmaxf(lmx, __shfl_xor_sync(0xffffffffu, lmx, o));
  const bool slow = __syncthreads_or(odd);
  if (lane == 0) redf[warp] = lmx;
  __syncthreads();
  if (warp == 0) {
    float mx = redf[0];
    #pragma unroll
    for (int w = 1; w < 8; w++) mx = fmaxf(mx, redf[w]);
    // The float32 sum must be the one taken in index order (total += data[i], :3599).  r1 had one thread add the n values
    // (3.2 ms for the 131,072-float pingpong array); the warp encoder's transducer sum (dcb_encode_warp.cuh, exact_medium:
    // lanes whose 16 values keep the running sum inside one binade compose exactly) gives the same bits a warp at a time.
    __shared__ WarpBuf wb;
    EncArgs ea;
    ea.in = a.in; ea.n = a.n;
    ea.flags = ((reinterpret_cast<uintptr_t>(a.in) & 15) == 0) ? DCB_ENC_IN_ALIGNED : 0u;
    float total, ymx, y0;
    exact_medium(ea, wb, lo, lo + cn, mn, a.subtract_min != 0, slow, false, lane, total, ymx, y0);
    if (lane == 0) {
      float first = __ldg(x);
      if (a.subtract_min) first = __fsub_rn(first, mn);
      a.cmed[c] = __fdiv_rn(total, (float)(int)cn);
      a.ctype[c] = (first != first) ? 0 : type_from_max(mx);
    }
  }
}
