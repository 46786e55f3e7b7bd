// dcb_face.cuh -- the step either side of the halo compression in the Himeno driver:
//   gather  == transform_3d_array_to_1d_array (dataCompression.c:3741-3775): one face of p[MIMAX][MJMAX][MKMAX] -> 1-D array,
//              ijk = 1: data[v][a][b] (a < jmax, b < kmax), 2: data[a][v][b] (a < imax, b < kmax), 3: data[a][b][v] (a < imax, b < jmax)
//   scatter == the loops after decompression (himenoBMTxps.c:560-568, :640-648, ...): p[face] = decompressed[pointer++] + min
// With p resident on the GPU the whole halo path (gather, min-subtract, compress, NCCL, decompress, scatter) never touches
// the host.  k-faces (ijk = 3) are strided by MKMAX floats: one thread per element, consecutive threads walk b (stride MKMAX)
// -- each access is its own sector, which is inherent to the layout; i/j faces are contiguous runs of kmax floats.
#pragma once
#include "dcb_common.cuh"

struct FaceArgs {
  int MJ, MK;        // allocated extents of the two inner dimensions (strides MJ*MK, MK, 1)
  int ijk, v;        // face orientation and its fixed index
  int A, B;          // logical extents of the face (outer, inner)
};

__device__ __forceinline__ size_t face_index(const FaceArgs& f, int a, int b) {
  if (f.ijk == 1) return ((size_t)f.v * f.MJ + a) * f.MK + b;
  if (f.ijk == 2) return ((size_t)a * f.MJ + f.v) * f.MK + b;
  return ((size_t)a * f.MJ + b) * f.MK + f.v;
}

__global__ void __launch_bounds__(256) dcb_gather_face_kernel(const float* __restrict__ p, FaceArgs f, float* __restrict__ out) {
  const size_t n = (size_t)f.A * f.B;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    out[i] = p[face_index(f, (int)(i / f.B), (int)(i % f.B))];
}

__global__ void __launch_bounds__(256) dcb_scatter_face_kernel(float* __restrict__ p, FaceArgs f, const float* __restrict__ in,
                                                               float add) {
  const size_t n = (size_t)f.A * f.B;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    p[face_index(f, (int)(i / f.B), (int)(i % f.B))] = __fadd_rn(in[i], add);
}

// ---- Himeno Jacobi sweep (impl/himenoBMTxps.c:247-284) ------------------------------------------------------------------
// One thread per interior point, k fastest (coalesced).  Every product and sum is rounded on its own, in the order the C
// expression evaluates them (the reference is built without FMA contraction), so wrk2 -- and with it the next p -- is bit
// exact; gosa = sum ss^2 is accumulated in double per block and added atomically (the reference's float sum in i,j,k order
// is a diagnostic and is compared to tolerance).  14 coefficient/field arrays make the sweep HBM bound: ~60 B per point.
struct JacobiArgs {
  const float* p; const float* a; const float* b; const float* c; const float* bnd; const float* wrk1;
  float* wrk2;
  int MI, MJ, MK, imax, jmax, kmax;
  float omega;
  double* gosa;
};

__global__ void __launch_bounds__(256) dcb_himeno_jacobi_kernel(JacobiArgs q) {
  const size_t sj = (size_t)q.MK, si = (size_t)q.MJ * q.MK, sa = (size_t)q.MI * si;
  const int nk = q.kmax - 2, nj = q.jmax - 2, ni = q.imax - 2;
  const size_t total = (size_t)ni * nj * nk;
  double acc = 0.0;
  for (size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (size_t)gridDim.x * blockDim.x) {
    const int k = 1 + (int)(t % nk), j = 1 + (int)((t / nk) % nj), i = 1 + (int)(t / ((size_t)nk * nj));
    const size_t o = (size_t)i * si + (size_t)j * sj + k;
    const float* p = q.p;
    float s0 = __fmul_rn(q.a[o], p[o + si]);
    s0 = __fadd_rn(s0, __fmul_rn(q.a[sa + o], p[o + sj]));
    s0 = __fadd_rn(s0, __fmul_rn(q.a[2 * sa + o], p[o + 1]));
    s0 = __fadd_rn(s0, __fmul_rn(q.b[o], __fadd_rn(__fsub_rn(__fsub_rn(p[o + si + sj], p[o + si - sj]), p[o - si + sj]), p[o - si - sj])));
    s0 = __fadd_rn(s0, __fmul_rn(q.b[sa + o], __fadd_rn(__fsub_rn(__fsub_rn(p[o + sj + 1], p[o - sj + 1]), p[o + sj - 1]), p[o - sj - 1])));
    s0 = __fadd_rn(s0, __fmul_rn(q.b[2 * sa + o], __fadd_rn(__fsub_rn(__fsub_rn(p[o + si + 1], p[o - si + 1]), p[o + si - 1]), p[o - si - 1])));
    s0 = __fadd_rn(s0, __fmul_rn(q.c[o], p[o - si]));
    s0 = __fadd_rn(s0, __fmul_rn(q.c[sa + o], p[o - sj]));
    s0 = __fadd_rn(s0, __fmul_rn(q.c[2 * sa + o], p[o - 1]));
    s0 = __fadd_rn(s0, q.wrk1[o]);
    const float ss = __fmul_rn(__fsub_rn(__fmul_rn(s0, q.a[3 * sa + o]), p[o]), q.bnd[o]);
    acc += (double)__fmul_rn(ss, ss);
    q.wrk2[o] = __fadd_rn(p[o], __fmul_rn(q.omega, ss));
  }
  #pragma unroll
  for (int s = 16; s; s >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, s);
  __shared__ double part[8];
  if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double v = 0.0;
    for (int w = 0; w < 8; w++) v += part[w];
    atomicAdd(q.gosa, v);
  }
}

// p[i][j][k] = wrk2[i][j][k] on the interior (:279-282)
__global__ void __launch_bounds__(256) dcb_himeno_copy_kernel(float* __restrict__ p, const float* __restrict__ wrk2,
                                                              int MJ, int MK, int imax, int jmax, int kmax) {
  const size_t sj = (size_t)MK, si = (size_t)MJ * MK;
  const int nk = kmax - 2, nj = jmax - 2, ni = imax - 2;
  const size_t total = (size_t)ni * nj * nk;
  for (size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (size_t)gridDim.x * blockDim.x) {
    const int k = 1 + (int)(t % nk), j = 1 + (int)((t / nk) % nj), i = 1 + (int)(t / ((size_t)nk * nj));
    const size_t o = (size_t)i * si + (size_t)j * sj + k;
    p[o] = wrk2[o];
  }
}
