// dcb_legacy.cu -- the reference's own float32 hot-path symbols (include/dataCompression_b200.h), as thin
// host-pointer wrappers over the device API of include/dc_b200.h: ONE chunk = the whole array, exactly one
// reference call.  Same names, argument meaning, ownership (malloc/realloc'd results the caller owns) and
// error behaviour (print + exit(1), dataCompression.c:2961-2963) as impl/dataCompression.c.
// Nothing here encodes or decodes on the CPU: without a CUDA device every codec call exits.
#include "../../include/dc_b200.h"
#include "../../include/dataCompression_b200.h"
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

static double g_abs_error_bound = -1.0;           // < 0: not set yet -> environment, then the header default

extern "C" void dc_set_abs_error_bound(double eb) { g_abs_error_bound = eb; }
extern "C" double dc_get_abs_error_bound(void) {
  if (g_abs_error_bound < 0.0) {
    const char* e = getenv("DC_ABS_ERROR_BOUND");
    g_abs_error_bound = (e && atof(e) > 0.0) ? atof(e) : 0.000001;     // impl/dataCompression.h:5
  }
  return g_abs_error_bound;
}

static void die(const char* what, int rc) {
  printf("dcb200: %s failed (rc=%d%s%s)\n", what, rc, rc == DCB_ECUDA ? ": " : "", rc == DCB_ECUDA ? dcb_last_cuda_error() : "");
  exit(1);
}

// mask[1+8+8] of '0'/'1' chars (impl/pingpong.c:202-206) -> the 17-bit integer the kernels take
static uint32_t mask17_from_chars(const char* mask) {
  uint32_t m = 0;
  for (int i = 0; i < 17; i++) m = (m << 1) | (uint32_t)(mask[i] == '1');
  return m;
}

// Appends `nbits` bits (MSB first in src) to the stream (*data_bits, *bytes, *pos) the way repeated
// add_bit_to_bytes calls would (dataCompression.c:5456-5489).  Every caller in the reference starts from
// (NULL, 0, 8), for which this is a realloc + memcpy; the shifted merge only runs for a caller-built prefix.
static void append_stream(unsigned char** data_bits, int* bytes, int* pos, const unsigned char* src, uint64_t nbits) {
  if (nbits == 0) return;
  const uint64_t src_bytes = (nbits + 7) / 8;
  int free_bits = (*pos == 8 || *bytes == 0) ? 0 : *pos;          // unused low bits of the current last byte
  const uint64_t old_bytes = (uint64_t)*bytes;
  const uint64_t total_bits = old_bytes * 8 - (uint64_t)free_bits + nbits;
  const uint64_t new_bytes = (total_bits + 7) / 8;
  unsigned char* p = (unsigned char*)realloc(*data_bits, (size_t)new_bytes);
  if (!p) { printf("ERROR: Out of memory\n"); exit(1); }          // dataCompression.c:5473-5477
  if (free_bits == 0) {
    memcpy(p + old_bytes, src, (size_t)src_bytes);
  } else {
    const int sh = 8 - free_bits;                                  // bits already used in the last byte
    memset(p + old_bytes, 0, (size_t)(new_bytes - old_bytes));
    for (uint64_t i = 0; i < src_bytes; i++) {
      p[old_bytes - 1 + i] |= (unsigned char)(src[i] >> sh);
      if (old_bytes + i < new_bytes) p[old_bytes + i] |= (unsigned char)(src[i] << free_bits);
    }
  }
  *data_bits = p;
  *bytes = (int)new_bytes;
  const int rem = (int)(total_bits % 8);
  *pos = rem ? 8 - rem : 8;
}

static void legacy_compress(int ct, float* data, int num, unsigned char** data_bits, int* bytes, int* pos, int type,
                            const char* mask) {
  if (num <= 0) return;
  const uint64_t n = (uint64_t)num;
  const uint64_t cap = dcb_max_stream_bytes(n, 0);
  unsigned char* tmp = (unsigned char*)malloc((size_t)cap);
  if (!tmp) { printf("ERROR: Out of memory\n"); exit(1); }
  uint64_t offs[2] = {0, 0};
  uint32_t flags = 0, m17 = 0;
  if (ct == DCB_CT_BITWISE_MASK) { flags |= DCB_F_FIXED_MASK; m17 = mask17_from_chars(mask); }
  const int rc = dcb_compress_host(ct, data, n, 0, dc_get_abs_error_bound(), flags, type, m17, tmp, offs, NULL, NULL,
                                   NULL, NULL);
  if (rc != DCB_OK) die("dcb_compress_host", rc);
  append_stream(data_bits, bytes, pos, tmp, dcb_last_stream_bits());
  free(tmp);
}

static float* legacy_decompress(int ct, unsigned char* data_bits, int bytes, int num, int type, const char* mask) {
  float* out = (float*)malloc(sizeof(float) * (size_t)(num > 0 ? num : 1));
  if (!out) { printf("ERROR: Out of memory\n"); exit(1); }
  if (num <= 0) return out;
  uint64_t offs[2] = {0, (uint64_t)(bytes > 0 ? bytes : 0)};
  uint32_t flags = 0, m17 = 0;
  if (ct == DCB_CT_BITWISE_MASK) { flags |= DCB_F_FIXED_MASK; m17 = mask17_from_chars(mask); }
  const int rc = dcb_decompress_host(ct, data_bits, offs, (uint64_t)num, 0, dc_get_abs_error_bound(), flags, type, m17,
                                     NULL, NULL, NULL, out);
  if (rc == DCB_EFORMAT) { printf("Error leading bit 1\n"); exit(1); }          // dataCompression.c:2501-2505
  if (rc != DCB_OK) die("dcb_decompress_host", rc);
  return out;
}

extern "C" void myCompress_bitwise(float data[], int num, unsigned char** data_bits, int* bytes, int* pos) {
  legacy_compress(DCB_CT_BITWISE, data, num, data_bits, bytes, pos, 0, NULL);
}
extern "C" float* myDecompress_bitwise(unsigned char* data_bits, int bytes, int num) {
  return legacy_decompress(DCB_CT_BITWISE, data_bits, bytes, num, 0, NULL);
}
extern "C" void myCompress_bitwise_np(float data[], int num, unsigned char** data_bits, int* bytes, int* pos) {
  legacy_compress(DCB_CT_BITWISE_NP, data, num, data_bits, bytes, pos, 0, NULL);
}
extern "C" float* myDecompress_bitwise_np(unsigned char* data_bits, int bytes, int num) {
  return legacy_decompress(DCB_CT_BITWISE_NP, data_bits, bytes, num, 0, NULL);
}
extern "C" void myCompress_bitwise_op(float data[], int num, unsigned char** data_bits, int* bytes, int* pos) {
  legacy_compress(DCB_CT_BITWISE_OP, data, num, data_bits, bytes, pos, 0, NULL);
}
extern "C" float* myDecompress_bitwise_op(unsigned char* data_bits, int bytes, int num) {
  return legacy_decompress(DCB_CT_BITWISE_OP, data_bits, bytes, num, 0, NULL);
}
extern "C" void myCompress_bitwise_mask(float data[], int num, unsigned char** data_bits, int* bytes, int* pos, int type,
                                        char mask[1 + 8 + 8]) {
  legacy_compress(DCB_CT_BITWISE_MASK, data, num, data_bits, bytes, pos, type < 0 ? 0 : (type > 7 ? 7 : type), mask);
}
extern "C" float* myDecompress_bitwise_mask(unsigned char* data_bits, int bytes, int num, int type, char mask[1 + 8 + 8]) {
  return legacy_decompress(DCB_CT_BITWISE_MASK, data_bits, bytes, num, type < 0 ? 0 : (type > 7 ? 7 : type), mask);
}

// ---- callers' front end ------------------------------------------------------------------------------------
__global__ void dcb_subtract_kernel(const float* __restrict__ x, const float* __restrict__ mn, float* __restrict__ y, uint64_t n) {
  const float m = *mn;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x)
    y[i] = __fsub_rn(x[i], m);
}

// every CUDA call of the front-end symbols is checked: a failed copy must not yield a silently wrong min / medium / CRC
#define LCK(call, what) do { if ((call) != cudaSuccess) die(what, DCB_ECUDA); } while (0)

struct DevBuf {
  void* p = nullptr;
  explicit DevBuf(size_t bytes) { if (cudaMalloc(&p, bytes ? bytes : 16) != cudaSuccess) p = nullptr; }
  ~DevBuf() { if (p) cudaFree(p); }
};

// toSmallDataset_float, dataCompression.c:3543-3562: min (first minimum wins), then data - min into a malloc'd array
extern "C" float toSmallDataset_float(float data[], float** data_small, int num) {
  *data_small = (float*)malloc(sizeof(float) * (size_t)(num > 0 ? num : 1));
  if (!*data_small) { printf("ERROR: Out of memory\n"); exit(1); }
  if (num <= 0) return 0.0f;
  if (dcb_device_count() == 0) die("toSmallDataset_float (no CUDA device)", DCB_ENODEVICE);
  const uint64_t n = (uint64_t)num;
  DevBuf dx(n * 4), dy(n * 4), dm(16);
  if (!dx.p || !dy.p || !dm.p) die("cudaMalloc", DCB_ECUDA);
  float mn = 0.0f;
  LCK(cudaMemcpy(dx.p, data, n * 4, cudaMemcpyHostToDevice), "toSmallDataset_float: copy in");
  int rc = dcb_chunk_stats((const float*)dx.p, n, 0, 0, (float*)dm.p, NULL, NULL, NULL);
  if (rc != DCB_OK) die("dcb_chunk_stats", rc);
  dcb_subtract_kernel<<<1184, 256>>>((const float*)dx.p, (const float*)dm.p, (float*)dy.p, n);
  LCK(cudaGetLastError(), "toSmallDataset_float: launch");
  LCK(cudaMemcpy(*data_small, dy.p, n * 4, cudaMemcpyDeviceToHost), "toSmallDataset_float: copy out");
  LCK(cudaMemcpy(&mn, dm.p, 4, cudaMemcpyDeviceToHost), "toSmallDataset_float: min");
  return mn;
}

// med_dataset_float, dataCompression.c:3593-3620: sequential float32 mean (exact: one thread walks the chain), max -> type
extern "C" float med_dataset_float(float* data, int num, int* type) {
  if (num <= 0) return 0.0f;
  if (dcb_device_count() == 0) die("med_dataset_float (no CUDA device)", DCB_ENODEVICE);
  const uint64_t n = (uint64_t)num;
  DevBuf dx(n * 4), dm(48);
  if (!dx.p || !dm.p) die("cudaMalloc", DCB_ECUDA);
  LCK(cudaMemcpy(dx.p, data, n * 4, cudaMemcpyHostToDevice), "med_dataset_float: copy in");
  float* d_min = (float*)dm.p; float* d_med = d_min + 4; int32_t* d_ty = (int32_t*)(d_min + 8);
  int rc = dcb_chunk_medium((const float*)dx.p, n, d_med, d_ty, NULL);
  if (rc != DCB_OK) die("dcb_chunk_medium", rc);
  float med = 0.0f; int32_t ty = 0;
  LCK(cudaMemcpy(&med, d_med, 4, cudaMemcpyDeviceToHost), "med_dataset_float: medium");
  LCK(cudaMemcpy(&ty, d_ty, 4, cudaMemcpyDeviceToHost), "med_dataset_float: type");
  if (ty != 0) *type = ty;                       // max >= 2^127: the reference leaves *type untouched (:3605-3614)
  return med;
}

// do_crc32, dataCompression.c:5524-5534 (zlib crc32 of the whole buffer)
extern "C" uint32_t do_crc32(unsigned char* data_bits, int bytes) {
  if (bytes <= 0) return 0u;
  if (dcb_device_count() == 0) die("do_crc32 (no CUDA device)", DCB_ENODEVICE);
  DevBuf ds((size_t)bytes + 16), dt(32);
  if (!ds.p || !dt.p) die("cudaMalloc", DCB_ECUDA);
  const uint64_t offs[2] = {0, (uint64_t)bytes};
  LCK(cudaMemcpy(ds.p, data_bits, (size_t)bytes, cudaMemcpyHostToDevice), "do_crc32: copy in");
  LCK(cudaMemcpy(dt.p, offs, 16, cudaMemcpyHostToDevice), "do_crc32: offsets");
  uint32_t* d_crc = (uint32_t*)((char*)dt.p + 16);
  int rc = dcb_crc32_chunks((const uint8_t*)ds.p, (const uint64_t*)dt.p, 1, d_crc, NULL);
  if (rc != DCB_OK) die("dcb_crc32_chunks", rc);
  uint32_t crc = 0;
  if (cudaMemcpy(&crc, d_crc, 4, cudaMemcpyDeviceToHost) != cudaSuccess) die("do_crc32", DCB_ECUDA);
  return crc;
}

// ---- bit-string helpers (pure formatting, dataCompression.c:5244-5276, :5512-5522) -----------------------------
extern "C" void floattostr(float* a, char* str) {
  uint32_t c;
  memcpy(&c, a, 4);
  for (int i = 0; i < 32; i++) str[i] = (char)('0' + ((c >> (31 - i)) & 1u));
  str[32] = '\0';
}
extern "C" float strtofloat(char* str) {
  uint32_t c = 0;
  for (int i = 0; i < 32; i++) c = (c << 1) + (uint32_t)(str[i] - '0');
  float f;
  memcpy(&f, &c, 4);
  return f;
}
extern "C" int to_absErrorBound_binary(double absErrBound) { return dcb_n_eb(absErrBound); }

// ================================ float64 symbols (k-means / MM / LU callers) ====================================
// mask[1+11+8] of '0'/'1' chars (impl/k-means.c:527-530) -> the 20-bit integer the kernels take
static uint32_t mask20_from_chars(const char* mask) {
  uint32_t m = 0;
  for (int i = 0; i < 20; i++) m = (m << 1) | (uint32_t)(mask[i] == '1');
  return m;
}

static void legacy_compress_f64(int ct, double* data, int num, unsigned char** data_bits, int* bytes, int* pos, int type,
                                const char* mask) {
  if (num <= 0) return;
  if (dcb_device_count() == 0) die("myCompress_bitwise_double* (no CUDA device)", DCB_ENODEVICE);
  const uint64_t n = (uint64_t)num;
  const uint64_t cap = dcb_max_stream_bytes_f64(n, 0), wsb = dcb_workspace_bytes_f64(n, 0);
  DevBuf dx(n * 8), ds(cap), doff(16), dws(wsb);
  if (!dx.p || !ds.p || !doff.p || !dws.p) die("cudaMalloc", DCB_ECUDA);
  uint32_t flags = 0, m20 = 0;
  if (ct == DCB_CT_BITWISE_MASK) { flags |= DCB_F_FIXED_MASK; m20 = mask20_from_chars(mask); }
  LCK(cudaMemcpy(dx.p, data, n * 8, cudaMemcpyHostToDevice), "legacy symbol: device copy");
  const int rc = dcb_compress_f64(ct, (const double*)dx.p, n, 0, dc_get_abs_error_bound(), flags, type, m20, (uint8_t*)ds.p, cap,
                                  (uint64_t*)doff.p, NULL, NULL, NULL, NULL, NULL, dws.p, wsb, NULL);
  if (rc != DCB_OK) die("dcb_compress_f64", rc);
  uint64_t offs[2] = {0, 0};
  unsigned long long nbits = 0;
  LCK(cudaMemcpy(offs, doff.p, 16, cudaMemcpyDeviceToHost), "legacy symbol: device copy");
  if (cudaMemcpy(&nbits, dws.p, 8, cudaMemcpyDeviceToHost) != cudaSuccess) die("myCompress_bitwise_double*", DCB_ECUDA);
  unsigned char* tmp = (unsigned char*)malloc((size_t)(offs[1] ? offs[1] : 1));
  if (!tmp) { printf("ERROR: Out of memory\n"); exit(1); }
  if (cudaMemcpy(tmp, ds.p, (size_t)offs[1], cudaMemcpyDeviceToHost) != cudaSuccess) die("myCompress_bitwise_double*", DCB_ECUDA);
  append_stream(data_bits, bytes, pos, tmp, nbits);
  free(tmp);
}

static double* legacy_decompress_f64(int ct, unsigned char* data_bits, int bytes, int num, int type, const char* mask) {
  double* out = (double*)malloc(sizeof(double) * (size_t)(num > 0 ? num : 1));
  if (!out) { printf("ERROR: Out of memory\n"); exit(1); }
  if (num <= 0) return out;
  if (dcb_device_count() == 0) die("myDecompress_bitwise_double* (no CUDA device)", DCB_ENODEVICE);
  const uint64_t n = (uint64_t)num, nb = (uint64_t)(bytes > 0 ? bytes : 0);
  DevBuf ds(nb + 16), doff(16), dout(n * 8), dst(16);
  if (!ds.p || !doff.p || !dout.p || !dst.p) die("cudaMalloc", DCB_ECUDA);
  const uint64_t offs[2] = {0, nb};
  uint32_t flags = 0, m20 = 0;
  if (ct == DCB_CT_BITWISE_MASK) { flags |= DCB_F_FIXED_MASK; m20 = mask20_from_chars(mask); }
  LCK(cudaMemset(ds.p, 0, (size_t)nb + 16), "legacy symbol: device copy");
  LCK(cudaMemcpy(ds.p, data_bits, (size_t)nb, cudaMemcpyHostToDevice), "legacy symbol: device copy");
  LCK(cudaMemcpy(doff.p, offs, 16, cudaMemcpyHostToDevice), "legacy symbol: device copy");
  const int rc = dcb_decompress_f64(ct, (const uint8_t*)ds.p, (const uint64_t*)doff.p, n, 0, dc_get_abs_error_bound(), flags, type,
                                    m20, NULL, NULL, NULL, (double*)dout.p, (int32_t*)dst.p, NULL);
  if (rc != DCB_OK) die("dcb_decompress_f64", rc);
  int32_t status = 0;
  LCK(cudaMemcpy(&status, dst.p, 4, cudaMemcpyDeviceToHost), "legacy symbol: device copy");
  if (cudaMemcpy(out, dout.p, n * 8, cudaMemcpyDeviceToHost) != cudaSuccess) die("myDecompress_bitwise_double*", DCB_ECUDA);
  if (status == DCB_EFORMAT) { printf("Error leading bit 1\n"); exit(1); }
  return out;
}

static int clamp_type_f64(int type) { return type < 0 ? 0 : (type > 10 ? 10 : type); }

extern "C" void myCompress_bitwise_double(double data[], int num, unsigned char** data_bits, int* bytes, int* pos) {
  legacy_compress_f64(DCB_CT_BITWISE, data, num, data_bits, bytes, pos, 0, NULL);
}
extern "C" double* myDecompress_bitwise_double(unsigned char* data_bits, int bytes, int num) {
  return legacy_decompress_f64(DCB_CT_BITWISE, data_bits, bytes, num, 0, NULL);
}
extern "C" void myCompress_bitwise_double_np(double data[], int num, unsigned char** data_bits, int* bytes, int* pos) {
  legacy_compress_f64(DCB_CT_BITWISE_NP, data, num, data_bits, bytes, pos, 0, NULL);
}
extern "C" double* myDecompress_bitwise_double_np(unsigned char* data_bits, int bytes, int num) {
  return legacy_decompress_f64(DCB_CT_BITWISE_NP, data_bits, bytes, num, 0, NULL);
}
extern "C" void myCompress_bitwise_double_op(double data[], int num, unsigned char** data_bits, int* bytes, int* pos) {
  legacy_compress_f64(DCB_CT_BITWISE_OP, data, num, data_bits, bytes, pos, 0, NULL);
}
extern "C" double* myDecompress_bitwise_double_op(unsigned char* data_bits, int bytes, int num) {
  return legacy_decompress_f64(DCB_CT_BITWISE_OP, data_bits, bytes, num, 0, NULL);
}
extern "C" void myCompress_bitwise_double_mask(double data[], int num, unsigned char** data_bits, int* bytes, int* pos, int type,
                                               char mask[1 + 11 + 8]) {
  legacy_compress_f64(DCB_CT_BITWISE_MASK, data, num, data_bits, bytes, pos, clamp_type_f64(type), mask);
}
extern "C" double* myDecompress_bitwise_double_mask(unsigned char* data_bits, int bytes, int num, int type, char mask[1 + 11 + 8]) {
  return legacy_decompress_f64(DCB_CT_BITWISE_MASK, data_bits, bytes, num, clamp_type_f64(type), mask);
}

__global__ void dcb_subtract_f64_kernel(const double* __restrict__ x, const double* __restrict__ mn, double* __restrict__ y, uint64_t n) {
  const double m = *mn;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x)
    y[i] = __dsub_rn(x[i], m);
}

// toSmallDataset_double, dataCompression.c:3522-3541
extern "C" double toSmallDataset_double(double data[], double** data_small, int num) {
  *data_small = (double*)malloc(sizeof(double) * (size_t)(num > 0 ? num : 1));
  if (!*data_small) { printf("ERROR: Out of memory\n"); exit(1); }
  if (num <= 0) return 0.0;
  if (dcb_device_count() == 0) die("toSmallDataset_double (no CUDA device)", DCB_ENODEVICE);
  const uint64_t n = (uint64_t)num;
  DevBuf dx(n * 8), dy(n * 8), dm(16);
  if (!dx.p || !dy.p || !dm.p) die("cudaMalloc", DCB_ECUDA);
  double mn = 0.0;
  LCK(cudaMemcpy(dx.p, data, n * 8, cudaMemcpyHostToDevice), "legacy symbol: device copy");
  int rc = dcb_chunk_stats_f64((const double*)dx.p, n, 0, 1, 0, (double*)dm.p, NULL, NULL, NULL, NULL);
  if (rc != DCB_OK) die("dcb_chunk_stats_f64", rc);
  dcb_subtract_f64_kernel<<<1184, 256>>>((const double*)dx.p, (const double*)dm.p, (double*)dy.p, n);
  LCK(cudaMemcpy(*data_small, dy.p, n * 8, cudaMemcpyDeviceToHost), "legacy symbol: device copy");
  if (cudaMemcpy(&mn, dm.p, 8, cudaMemcpyDeviceToHost) != cudaSuccess) die("toSmallDataset_double", DCB_ECUDA);
  return mn;
}

// med_dataset_double, dataCompression.c:3564-3591: exact sequential mean, max -> type (untouched when max >= 2^1023)
extern "C" double med_dataset_double(double* data, int num, int* type) {
  if (num <= 0) return 0.0;
  if (dcb_device_count() == 0) die("med_dataset_double (no CUDA device)", DCB_ENODEVICE);
  const uint64_t n = (uint64_t)num;
  DevBuf dx(n * 8), dm(64);
  if (!dx.p || !dm.p) die("cudaMalloc", DCB_ECUDA);
  LCK(cudaMemcpy(dx.p, data, n * 8, cudaMemcpyHostToDevice), "legacy symbol: device copy");
  double* d_med = (double*)dm.p; uint32_t* d_mask = (uint32_t*)((char*)dm.p + 16); int32_t* d_ty = (int32_t*)((char*)dm.p + 32);
  int rc = dcb_chunk_stats_f64((const double*)dx.p, n, 0, 0, 1, NULL, d_mask, d_ty, d_med, NULL);
  if (rc != DCB_OK) die("dcb_chunk_stats_f64", rc);
  double med = 0.0; int32_t ty = 0;
  LCK(cudaMemcpy(&med, d_med, 8, cudaMemcpyDeviceToHost), "legacy symbol: device copy");
  if (cudaMemcpy(&ty, d_ty, 4, cudaMemcpyDeviceToHost) != cudaSuccess) die("med_dataset_double", DCB_ECUDA);
  if (ty != 0) *type = ty;
  return med;
}

// doubletostr / strtodbl, dataCompression.c (bit-string helpers the drivers use to build the mask, impl/k-means.c:526-530)
extern "C" void doubletostr(double* a, char* str) {
  uint64_t c;
  memcpy(&c, a, 8);
  for (int i = 0; i < 64; i++) str[i] = (char)('0' + ((c >> (63 - i)) & 1ull));
  str[64] = '\0';
}
extern "C" double strtodbl(char* str) {
  uint64_t c = 0;
  for (int i = 0; i < 64; i++) c = (c << 1) + (uint64_t)(str[i] - '0');
  double d;
  memcpy(&d, &c, 8);
  return d;
}


// ---- CT=10: Hamming SECDED blocks of the finished stream (dataCompression.h:144-158) -------------------------------------
static double g_ber = -1.0;                        // < 0: not set yet -> environment, then the header default (dataCompression.h:4)
extern "C" void dc_set_ber(double ber) { g_ber = ber; }
extern "C" double dc_get_ber(void) {
  if (g_ber < 0.0) {
    const char* e = getenv("DC_BER");
    g_ber = (e && atof(e) > 0.0) ? atof(e) : 1e-6;
  }
  return g_ber;
}
extern "C" int hmLength(int k) { return dcb_hm_length(k); }                                   // dataCompression.c:5587
extern "C" int block_size(int data_bytes) {                                                    // :5868-5878
  const double ber = dc_get_ber();
  const uint64_t bits = (uint64_t)(1 / ber);
  const uint64_t bytes = bits / 8;
  int bs = data_bytes;
  if ((uint64_t)bs > bytes) bs = (int)bytes;
  return bs;
}
extern "C" uint64_t get_random_int(uint64_t from, uint64_t to) { return (uint64_t)rand() % (to - from + 1) + from; }   // :5536
extern "C" void bit_flip(unsigned char* bits, int bytes) {                                     // :5860-5866 (fault injection, host memory)
  const int num = rand() % (bytes * 8);
  bits[num / 8] ^= (unsigned char)(1 << (7 - num % 8));
}
extern "C" void cast_bits_to_char(unsigned char* bits, char* data, int bytes) {                // :5712-5722
  for (int i = 0; i < bytes; i++)
    for (int j = 0; j < 8; j++) data[i * 8 + j] = (char)(((bits[i] >> (7 - j)) & 1) + '0');
}

// one block on the device: stage the bytes, run dcb_hamming_{encode,decode}, read the packed check word back
static uint32_t hm_pos_host(uint32_t d) {             // codeword position of data bit d (powers of two hold check bits)
  uint32_t m = d + 1, t = 0;
  while ((1u << t) <= m) t++;
  uint32_t j = m + t;
  if (j >> t) j++;
  return j;
}
static uint32_t hamming_device(unsigned char* bits, int bytes, uint32_t check_in, int decode, int* error_type, uint32_t* err_pos) {
  uint8_t* d_b = nullptr; uint32_t* d_c = nullptr; int32_t* d_e = nullptr;
  uint32_t check = check_in; int32_t et = 0;
  if (cudaMalloc(&d_b, (size_t)bytes) != cudaSuccess || cudaMalloc(&d_c, 8) != cudaSuccess || cudaMalloc(&d_e, 8) != cudaSuccess)
    die("hamming: cudaMalloc", DCB_ECUDA);
  LCK(cudaMemcpy(d_b, bits, (size_t)bytes, cudaMemcpyHostToDevice), "legacy symbol: device copy");
  LCK(cudaMemcpy(d_c, &check, 4, cudaMemcpyHostToDevice), "legacy symbol: device copy");
  const int rc = decode ? dcb_hamming_decode(d_b, (uint64_t)bytes, (uint32_t)bytes, d_c, d_e, nullptr)
                        : dcb_hamming_encode(d_b, (uint64_t)bytes, (uint32_t)bytes, d_c, nullptr, nullptr);
  if (rc) die(decode ? "hamming_decode" : "hamming_encode", rc);
  if (cudaMemcpy(&check, d_c, 4, cudaMemcpyDeviceToHost) != cudaSuccess) die("hamming: copy back", DCB_ECUDA);
  if (decode) {
    LCK(cudaMemcpy(&et, d_e, 4, cudaMemcpyDeviceToHost), "legacy symbol: device copy");
    if (et == 3) {                                  // a rectified data bit: find it (for the reference's message), take it over
      unsigned char* fixed = (unsigned char*)malloc((size_t)bytes);
      if (!fixed) { printf("ERROR: Out of memory\n"); exit(1); }
      LCK(cudaMemcpy(fixed, d_b, (size_t)bytes, cudaMemcpyDeviceToHost), "legacy symbol: device copy");
      for (int i = 0; i < bytes; i++)
        if (fixed[i] != bits[i]) {
          const unsigned x = (unsigned)(fixed[i] ^ bits[i]);
          int b = 0; while (!((x >> (7 - b)) & 1u)) b++;
          *err_pos = hm_pos_host((uint32_t)(8 * i + b));
          break;
        }
      memcpy(bits, fixed, (size_t)bytes);
      free(fixed);
    }
    *error_type = et;
  }
  cudaFree(d_b); cudaFree(d_c); cudaFree(d_e);
  return check;
}
extern "C" void hamming_encode(unsigned char* bits, char** c, int bytes, int* r) {             // :5742-5750
  if (dcb_device_count() == 0) die("hamming_encode", DCB_ENODEVICE);
  *r = dcb_hm_length(bytes * 8);
  *c = (char*)malloc(sizeof(char) * (size_t)(*r + 1));
  if (bytes <= 0) return;
  const uint32_t check = hamming_device(bits, bytes, 0u, 0, nullptr, nullptr);
  for (int i = 0; i <= *r; i++) (*c)[i] = (char)('0' + ((check >> i) & 1u));
}
extern "C" int hamming_decode(unsigned char* bits, char* c, int bytes, int r) {               // :5752-5779
  if (dcb_device_count() == 0) die("hamming_decode", DCB_ENODEVICE);
  if (bytes <= 0) return 0;
  uint32_t check = 0u;
  for (int i = 0; i <= r; i++) check |= (uint32_t)(c[i] == '1') << i;
  int et = 0;
  uint32_t pos = 0u;
  const uint32_t fixed = hamming_device(bits, bytes, check, 1, &et, &pos);
  switch (et) {                                                                               // the reference's own messages
    case 0: break;
    case 1: printf("two-bit error\n"); break;
    case 2: printf("parity error\n"); break;
    case 3:
      if ((fixed ^ check) & ((1u << r) - 1u)) pos = (fixed ^ check) & ((1u << r) - 1u);   // a rectified check bit c[i] sits at position 2^i
      printf("one bit error: pos = %d\n", (int)pos);
      break;
    default: printf("ERROR");
  }
  for (int i = 0; i <= r; i++) c[i] = (char)('0' + ((fixed >> i) & 1u));
  return et;
}


// ---- CT=1: the byte-wise codec (dataCompression.h: myCompress / myDecompress, float) ------------------------------------------
#define BW_CK(call) do { if ((call) != cudaSuccess) die("bytewise: " #call, DCB_ECUDA); } while (0)
extern "C" int myCompress(float data[], float** array_float, char** array_char, int** array_char_displacement, int num) {   // :3977
  if (dcb_device_count() == 0) die("myCompress", DCB_ENODEVICE);
  if (num <= 0) return 0;
  const uint64_t n = (uint64_t)num;
  float *d_in, *d_f; uint8_t* d_c; int32_t* d_d; uint32_t* d_nf; void* d_ws;
  const uint64_t wsb = dcb_bytewise_workspace_bytes(n, 0);
  BW_CK(cudaMalloc(&d_in, n * 4)); BW_CK(cudaMalloc(&d_f, n * 4)); BW_CK(cudaMalloc(&d_c, n)); BW_CK(cudaMalloc(&d_d, n * 4));
  BW_CK(cudaMalloc(&d_nf, 4)); BW_CK(cudaMalloc(&d_ws, wsb));
  BW_CK(cudaMemcpy(d_in, data, n * 4, cudaMemcpyHostToDevice));
  const int rc = dcb_bytewise_compress(d_in, n, 0, dc_get_abs_error_bound(), d_f, d_c, d_d, d_nf, d_ws, wsb, nullptr);
  if (rc) die("myCompress", rc);
  uint32_t nf = 0;
  BW_CK(cudaMemcpy(&nf, d_nf, 4, cudaMemcpyDeviceToHost));
  const uint32_t nc = (uint32_t)num - nf;
  // the reference grows the caller's arrays with realloc (callers start from NULL) and leaves them untouched when empty
  if (nf) {
    float* f = (float*)realloc(*array_float, sizeof(float) * nf);
    if (!f) { free(*array_float); printf("Error (re)allocating memory\n"); exit(1); }                 // :4001-4005
    BW_CK(cudaMemcpy(f, d_f, (size_t)nf * 4, cudaMemcpyDeviceToHost));
    *array_float = f;
  }
  if (nc) {
    char* c = (char*)realloc(*array_char, sizeof(char) * nc);
    int* d = (int*)realloc(*array_char_displacement, sizeof(int) * nc);
    if (!c || !d) { printf("Error (re)allocating memory\n"); exit(1); }                                // :4081-4086
    BW_CK(cudaMemcpy(c, d_c, nc, cudaMemcpyDeviceToHost));
    BW_CK(cudaMemcpy(d, d_d, (size_t)nc * 4, cudaMemcpyDeviceToHost));
    *array_char = c; *array_char_displacement = d;
  }
  cudaFree(d_in); cudaFree(d_f); cudaFree(d_c); cudaFree(d_d); cudaFree(d_nf); cudaFree(d_ws);
  return (int)nf;
}
extern "C" float* myDecompress(float array_float[], char array_char[], int array_char_displacement[], int num) {           // :3942
  if (dcb_device_count() == 0) die("myDecompress", DCB_ENODEVICE);
  float* out = (float*)malloc(sizeof(float) * (size_t)(num > 0 ? num : 1));
  if (num <= 0) return out;
  // The call does not say how many hits there are: the reference finds out by comparing displacement[p] - 1 with the running
  // index (and reads displacement[p] one past its end once the last hit is consumed, :3948).  The same walk, the same reads.
  uint32_t nc = 0;
  if (array_char_displacement)
    for (int i = 0; i < num; i++)
      if (array_char_displacement[nc] - 1 == i) nc++;
  const uint32_t nf = (uint32_t)num - nc;
  const uint64_t n = (uint64_t)num;
  float *d_f, *d_o; uint8_t* d_c; int32_t* d_d; uint32_t* d_nf;
  BW_CK(cudaMalloc(&d_f, n * 4)); BW_CK(cudaMalloc(&d_o, n * 4)); BW_CK(cudaMalloc(&d_c, n)); BW_CK(cudaMalloc(&d_d, n * 4)); BW_CK(cudaMalloc(&d_nf, 4));
  if (nf) BW_CK(cudaMemcpy(d_f, array_float, (size_t)nf * 4, cudaMemcpyHostToDevice));
  if (nc) { BW_CK(cudaMemcpy(d_c, array_char, nc, cudaMemcpyHostToDevice)); BW_CK(cudaMemcpy(d_d, array_char_displacement, (size_t)nc * 4, cudaMemcpyHostToDevice)); }
  BW_CK(cudaMemcpy(d_nf, &nf, 4, cudaMemcpyHostToDevice));
  const int rc = dcb_bytewise_decompress(d_f, d_c, d_d, d_nf, n, 0, d_o, nullptr);
  if (rc) die("myDecompress", rc);
  BW_CK(cudaMemcpy(out, d_o, n * 4, cudaMemcpyDeviceToHost));
  cudaFree(d_f); cudaFree(d_o); cudaFree(d_c); cudaFree(d_d); cudaFree(d_nf);
  return out;
}


// ---- CT=1, float64: myCompress_double / myDecompress_double (the k-means / MM / LU callers) -----------------------------------
extern "C" int myCompress_double(double data[], double** array_double, char** array_char, int** array_char_displacement, int num) {   // :3815
  if (dcb_device_count() == 0) die("myCompress_double", DCB_ENODEVICE);
  if (num <= 0) return 0;
  const uint64_t n = (uint64_t)num;
  double *d_in, *d_f; uint8_t* d_c; int32_t* d_d; uint32_t* d_nf; void* d_ws;
  const uint64_t wsb = dcb_bytewise_workspace_bytes(n, 0);
  BW_CK(cudaMalloc(&d_in, n * 8)); BW_CK(cudaMalloc(&d_f, n * 8)); BW_CK(cudaMalloc(&d_c, n)); BW_CK(cudaMalloc(&d_d, n * 4));
  BW_CK(cudaMalloc(&d_nf, 4)); BW_CK(cudaMalloc(&d_ws, wsb));
  BW_CK(cudaMemcpy(d_in, data, n * 8, cudaMemcpyHostToDevice));
  const int rc = dcb_bytewise_compress_f64(d_in, n, 0, dc_get_abs_error_bound(), d_f, d_c, d_d, d_nf, d_ws, wsb, nullptr);
  if (rc) die("myCompress_double", rc);
  uint32_t nf = 0;
  BW_CK(cudaMemcpy(&nf, d_nf, 4, cudaMemcpyDeviceToHost));
  const uint32_t nc = (uint32_t)num - nf;
  if (nf) {
    double* f = (double*)realloc(*array_double, sizeof(double) * nf);
    if (!f) { free(*array_double); printf("Error (re)allocating memory\n"); exit(1); }
    BW_CK(cudaMemcpy(f, d_f, (size_t)nf * 8, cudaMemcpyDeviceToHost));
    *array_double = f;
  }
  if (nc) {
    char* c = (char*)realloc(*array_char, sizeof(char) * nc);
    int* d = (int*)realloc(*array_char_displacement, sizeof(int) * nc);
    if (!c || !d) { printf("Error (re)allocating memory\n"); exit(1); }
    BW_CK(cudaMemcpy(c, d_c, nc, cudaMemcpyDeviceToHost));
    BW_CK(cudaMemcpy(d, d_d, (size_t)nc * 4, cudaMemcpyDeviceToHost));
    *array_char = c; *array_char_displacement = d;
  }
  cudaFree(d_in); cudaFree(d_f); cudaFree(d_c); cudaFree(d_d); cudaFree(d_nf); cudaFree(d_ws);
  return (int)nf;
}
extern "C" double* myDecompress_double(double array_double[], char array_char[], int array_char_displacement[], int num) {           // :3778
  if (dcb_device_count() == 0) die("myDecompress_double", DCB_ENODEVICE);
  double* out = (double*)malloc(sizeof(double) * (size_t)(num > 0 ? num : 1));
  if (num <= 0) return out;
  uint32_t nc = 0;                                     // the reference's own walk over the displacements (see myDecompress)
  if (array_char_displacement)
    for (int i = 0; i < num; i++)
      if (array_char_displacement[nc] - 1 == i) nc++;
  const uint32_t nf = (uint32_t)num - nc;
  const uint64_t n = (uint64_t)num;
  double *d_f, *d_o; uint8_t* d_c; int32_t* d_d; uint32_t* d_nf;
  BW_CK(cudaMalloc(&d_f, n * 8)); BW_CK(cudaMalloc(&d_o, n * 8)); BW_CK(cudaMalloc(&d_c, n)); BW_CK(cudaMalloc(&d_d, n * 4)); BW_CK(cudaMalloc(&d_nf, 4));
  if (nf) BW_CK(cudaMemcpy(d_f, array_double, (size_t)nf * 8, cudaMemcpyHostToDevice));
  if (nc) { BW_CK(cudaMemcpy(d_c, array_char, nc, cudaMemcpyHostToDevice)); BW_CK(cudaMemcpy(d_d, array_char_displacement, (size_t)nc * 4, cudaMemcpyHostToDevice)); }
  BW_CK(cudaMemcpy(d_nf, &nf, 4, cudaMemcpyHostToDevice));
  const int rc = dcb_bytewise_decompress_f64(d_f, d_c, d_d, d_nf, n, 0, d_o, nullptr);
  if (rc) die("myDecompress_double", rc);
  BW_CK(cudaMemcpy(out, d_o, n * 8, cudaMemcpyDeviceToHost));
  cudaFree(d_f); cudaFree(d_o); cudaFree(d_c); cudaFree(d_d); cudaFree(d_nf);
  return out;
}
