/* dataCompression_b200.h -- the reference's own float32 (and float64) hot-path symbols, served by libdcb200.so.
 *
 * Same names, same signatures, same ownership rules as impl/dataCompression.h of
 * smallcat9603/data-compression (the line each prototype replaces is cited), so that the MPI drivers
 * (impl/pingpong.c, impl/himenoBMTxps.c) link libdcb200.so in place of dataCompression.o for these
 * symbols without a source change (INTEGRATION.md shows the link line).
 *
 * All pointers are HOST pointers, exactly as in the reference: every call stages the data to the GPU,
 * runs the sm_100a kernels behind include/dc_b200.h with ONE chunk = the whole array, and copies the
 * result back.  There is no CPU codec: without a CUDA device the calls print the reason and exit(1)
 * (the reference's own error behaviour, dataCompression.c:2961-2963, :2503-2504).
 *
 * Differences from the reference, all deliberate:
 *   - absErrorBound is a compile-time macro there (dataCompression.h:5, rewritten by set-parameter.sh);
 *     here it is a run-time value, default 0.000001 = the header's value, changed with
 *     dc_set_abs_error_bound() or the environment variable DC_ABS_ERROR_BOUND.
 *   - the decoders return defined values in the corners where the reference reads uninitialised memory
 *     or mis-parses (listed in DESIGN.md "Reference-undefined corners").
 */
#ifndef DATACOMPRESSION_B200_H
#define DATACOMPRESSION_B200_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* run-time replacement of `#define absErrorBound` (dataCompression.h:5) and `double absErrBound` (dataCompression.c:21) */
void   dc_set_abs_error_bound(double abs_error_bound);
double dc_get_abs_error_bound(void);

/* CT=5  bit-wise with prediction */
void   myCompress_bitwise(float data[], int num, unsigned char** data_bits, int* bytes, int* pos);            /* dataCompression.h:90 */
float* myDecompress_bitwise(unsigned char* data_bits, int bytes, int num);                                    /* dataCompression.h:86 */
/* CT=6  bit-wise, no prediction */
void   myCompress_bitwise_np(float data[], int num, unsigned char** data_bits, int* bytes, int* pos);         /* dataCompression.h:77 */
float* myDecompress_bitwise_np(unsigned char* data_bits, int bytes, int num);                                 /* dataCompression.h:74 */
/* CT=11 bit-wise, prediction only */
void   myCompress_bitwise_op(float data[], int num, unsigned char** data_bits, int* bytes, int* pos);         /* dataCompression.h:81 */
float* myDecompress_bitwise_op(unsigned char* data_bits, int bytes, int num);                                 /* dataCompression.h:82 */
/* CT=7  bitmask-based bit-wise */
void   myCompress_bitwise_mask(float data[], int num, unsigned char** data_bits, int* bytes, int* pos,
                               int type, char mask[1 + 8 + 8]);                                               /* dataCompression.h:70 */
float* myDecompress_bitwise_mask(unsigned char* data_bits, int bytes, int num, int type,
                                 char mask[1 + 8 + 8]);                                                       /* dataCompression.h:67 */

/* callers' front end: min-subtract (malloc's *data_small, returns min) and mean/type for the mask */
float  toSmallDataset_float(float data[], float** data_small, int num);                                       /* dataCompression.h:96 */
float  med_dataset_float(float* data, int num, int* type);                                                    /* dataCompression.h:99 */

/* CT=8/9: zlib-compatible CRC-32 of a finished stream */
uint32_t do_crc32(unsigned char* data_bits, int bytes);                                                       /* dataCompression.h:143 */

/* bit-string helpers the drivers use to build the mask (impl/pingpong.c:202-206) */
void   floattostr(float* a, char* str);                                                                       /* dataCompression.h:125 */
float  strtofloat(char* str);                                                                                 /* dataCompression.h:127 */
int    to_absErrorBound_binary(double absErrBound);                                                           /* dataCompression.h:141 */


/* ---- float64 variants (the k-means / MM / LU callers, impl/k-means.c:510-780, impl/mm.c:404-690, impl/lu.c:199-345) ---- */
void    myCompress_bitwise_double(double data[], int num, unsigned char** data_bits, int* bytes, int* pos);   /* dataCompression.h:89 */
double* myDecompress_bitwise_double(unsigned char* data_bits, int bytes, int num);                            /* dataCompression.h:84 */
void    myCompress_bitwise_double_np(double data[], int num, unsigned char** data_bits, int* bytes, int* pos);/* dataCompression.h:76 */
double* myDecompress_bitwise_double_np(unsigned char* data_bits, int bytes, int num);                         /* dataCompression.h:72 */
void    myCompress_bitwise_double_op(double data[], int num, unsigned char** data_bits, int* bytes, int* pos);/* dataCompression.h:79 */
double* myDecompress_bitwise_double_op(unsigned char* data_bits, int bytes, int num);                         /* dataCompression.h:80 */
void    myCompress_bitwise_double_mask(double data[], int num, unsigned char** data_bits, int* bytes, int* pos,
                                       int type, char mask[1 + 11 + 8]);                                      /* dataCompression.h:66 */
double* myDecompress_bitwise_double_mask(unsigned char* data_bits, int bytes, int num, int type,
                                         char mask[1 + 11 + 8]);                                              /* dataCompression.h:63 */
double  toSmallDataset_double(double data[], double** data_small, int num);                                   /* dataCompression.h:95 */
double  med_dataset_double(double* data, int num, int* type);                                                 /* dataCompression.h:98 */
void    doubletostr(double* a, char* str);                                                                    /* dataCompression.h:126 */
double  strtodbl(char* str);                                                                                  /* dataCompression.h:128 */


/* ---- CT=1: the byte-wise codec of pingpong / Himeno (impl/pingpong.c:148-161, impl/himenoBMTxps.c) ---- */
int    myCompress(float data[], float** array_float, char** array_char, int** array_char_displacement, int num);   /* dataCompression.h:120 */
float* myDecompress(float array_float[], char array_char[], int array_char_displacement[], int num);               /* dataCompression.h:121 */
int     myCompress_double(double data[], double** array_double, char** array_char, int** array_char_displacement, int num);   /* impl/dataCompression.c:3815 */
double* myDecompress_double(double array_double[], char array_char[], int array_char_displacement[], int num);             /* impl/dataCompression.c:3778 */

/* ---- CT=10: Hamming SECDED over blocks of the finished stream (impl/pingpong.c:212-222, :296-311, :487-505) ----
 * BER is a compile-time macro in the reference (dataCompression.h:4); here dc_set_ber() / DC_BER, default 1e-6. */
void     dc_set_ber(double ber);
double   dc_get_ber(void);
uint64_t get_random_int(uint64_t from, uint64_t to);                                                          /* dataCompression.h:144 */
int      hmLength(int k);                                                                                     /* dataCompression.h:146 */
void     cast_bits_to_char(unsigned char* bits, char* data, int bytes);                                       /* dataCompression.h:152 */
void     hamming_encode(unsigned char* bits, char** c, int bytes, int* r);                                    /* dataCompression.h:153 */
int      hamming_decode(unsigned char* bits, char* c, int bytes, int r);                                      /* dataCompression.h:154 */
void     bit_flip(unsigned char* bits, int bytes);                                                            /* dataCompression.h:157 */
int      block_size(int data_bytes);                                                                          /* dataCompression.h:158 */

#ifdef __cplusplus
}
#endif
#endif
