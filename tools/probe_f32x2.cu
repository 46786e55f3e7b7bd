// probe_f32x2.cu -- evidence for the packed classification arithmetic of dcb_encode_warp.cuh:
//  (1) add/sub/mul.rn.f32x2 (SASS FADD2 / FMUL2) give, in each half, exactly the bits of __fadd_rn / __fsub_rn / __fmul_rn
//      on random bit patterns (denormals, infinities and NaN classes included) -- the reference's per-operation float32
//      rounding (impl/dataCompression.c:3362-3364) survives the packing;
//  (2) fl(2b + b) == fl(3 * b) (the kernel forms 3*b by two additions so that ptxas cannot contract into FFMA2);
//  (3) issue cost: cycles per iteration of FADD vs FADD2 streams, alone and interleaved with ALU-pipe work.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -fmad=false -o tools/probe_f32x2 tools/probe_f32x2.cu
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <cuda_runtime.h>

__device__ __forceinline__ uint64_t pk2(float lo, float hi) { uint64_t r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ void upk2(uint64_t r, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(r)); }
__device__ __forceinline__ uint64_t add2(uint64_t a, uint64_t b) { uint64_t r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ uint64_t sub2(uint64_t a, uint64_t b) { uint64_t r; asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ uint64_t mul2(uint64_t a, uint64_t b) { uint64_t r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }

__device__ uint32_t mix(uint32_t h) { h ^= h >> 16; h *= 0x85EBCA6Bu; h ^= h >> 13; h *= 0xC2B2AE35u; h ^= h >> 16; return h; }
// random float bit patterns: a quarter raw bits (NaN / Inf / denormals appear), a quarter denormals, the rest near each other
__device__ float pattern(uint32_t i, uint32_t salt) {
  const uint32_t h = mix(i * 0x9E3779B9u + salt), sel = mix(h + 77u) & 3u;
  if (sel == 0) return __uint_as_float(h);
  if (sel == 1) return __uint_as_float(h & 0x807FFFFFu);
  return __uint_as_float(0x3F000000u + (h & 0x03FFFFFFu));
}
__device__ bool same(float a, float b) {                       // bit-equal, or both NaN (payloads of generated NaNs are canonical)
  return __float_as_uint(a) == __float_as_uint(b) || (a != a && b != b);
}

__global__ void check_kernel(uint32_t n, unsigned long long* bad) {
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const float a0 = pattern(i, 1), a1 = pattern(i, 2), b0 = pattern(i, 3), b1 = pattern(i, 4);
    float r0, r1;
    unsigned long long e = 0;
    upk2(add2(pk2(a0, a1), pk2(b0, b1)), r0, r1);
    e += !same(r0, __fadd_rn(a0, b0)) + !same(r1, __fadd_rn(a1, b1));
    upk2(sub2(pk2(a0, a1), pk2(b0, b1)), r0, r1);
    e += !same(r0, __fsub_rn(a0, b0)) + !same(r1, __fsub_rn(a1, b1));
    upk2(mul2(pk2(a0, a1), pk2(b0, b1)), r0, r1);
    e += !same(r0, __fmul_rn(a0, b0)) + !same(r1, __fmul_rn(a1, b1));
    const uint64_t A = pk2(a0, a1);
    upk2(add2(add2(A, A), A), r0, r1);                          // 3*b as fl(2b + b)
    e += !same(r0, __fmul_rn(3.0f, a0)) + !same(r1, __fmul_rn(3.0f, a1));
    if (e) atomicAdd(bad, e);
  }
}

// ---- issue cost ------------------------------------------------------------------------------------------------
template <int MODE>
__global__ void rate_kernel(float* out, long long* cyc, int iters) {
  float a[8], c = out[threadIdx.x & 31];
  uint32_t u = threadIdx.x;
  #pragma unroll
  for (int k = 0; k < 8; k++) a[k] = c + (float)k;
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; it++) {
    if (MODE == 0 || MODE == 2) {                               // 8 scalar FADD
      #pragma unroll
      for (int k = 0; k < 8; k++) a[k] = __fadd_rn(a[k], c);
    }
    if (MODE == 1 || MODE == 3) {                               // 4 FADD2 = the same 8 element-operations
      #pragma unroll
      for (int k = 0; k < 8; k += 2) { uint64_t r = add2(pk2(a[k], a[k + 1]), pk2(c, c)); upk2(r, a[k], a[k + 1]); }
    }
    if (MODE >= 2) {                                            // plus ALU-pipe instructions
      #pragma unroll
      for (int k = 0; k < 8; k++) u = (u ^ (u >> 3)) + (u & 0x55u);
    }
  }
  const long long t1 = clock64();
  float s = 0.0f;
  #pragma unroll
  for (int k = 0; k < 8; k++) s += a[k];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s + (float)u;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

int main() {
  unsigned long long* bad; cudaMalloc(&bad, 8); cudaMemset(bad, 0, 8);
  const uint32_t n = 1u << 28;
  check_kernel<<<148 * 8, 256>>>(n, bad);
  unsigned long long hb = 1; cudaMemcpy(&hb, bad, 8, cudaMemcpyDeviceToHost);
  cudaError_t e = cudaDeviceSynchronize();
  printf("f32x2 vs scalar rn: %u quadruples x (add, sub, mul, 2b+b), mismatches %llu, cuda %s\n", n, hb, cudaGetErrorString(e));
  float* out; long long* cyc; cudaMalloc(&out, 148 * 1024 * 4 * 4); cudaMalloc(&cyc, 8);
  cudaMemset(out, 0, 148 * 1024 * 4 * 4);
  const int iters = 4096;
  const char* names[4] = {"8 FADD", "4 FADD2", "8 FADD + ALU", "4 FADD2 + ALU"};
  for (int warps = 4; warps <= 16; warps *= 2) {
    for (int m = 0; m < 4; m++) {
      for (int rep = 0; rep < 2; rep++) {
        if (m == 0) rate_kernel<0><<<148, warps * 32>>>(out, cyc, iters);
        if (m == 1) rate_kernel<1><<<148, warps * 32>>>(out, cyc, iters);
        if (m == 2) rate_kernel<2><<<148, warps * 32>>>(out, cyc, iters);
        if (m == 3) rate_kernel<3><<<148, warps * 32>>>(out, cyc, iters);
      }
      long long hc = 0; cudaMemcpy(&hc, cyc, 8, cudaMemcpyDeviceToHost);
      printf("warps/SM %2d  %-14s  %.2f cycles per iteration (%.3f per warp of a scheduler)\n", warps, names[m],
             (double)hc / iters, (double)hc / iters / (warps / 4.0));
    }
  }
  return hb != 0 || e != cudaSuccess;
}
