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
