/* dc_oracle.h -- CPU ORACLE, TEST INFRASTRUCTURE ONLY.
 *
 * A plain-C restatement of the float32 bit-wise codecs of
 * smallcat9603/data-compression (impl/dataCompression.c), used as the checker
 * by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg.  The
 * product (data_compression_b200/csrc) never includes, links or calls this.
 *
 * Parity is PINNED: tests/test_oracle_*.py check this restatement against
 *   - the reference's golden bit dumps impl/dataset/bit.txt + bitcomp.txt,
 *   - the known-answer ratios impl/pingpong.csv:36-38,
 *   - the unmodified reference compiled by oracle/build_ref.sh (oracle/_ref),
 *   - the committed fixtures tests/golden/ (generated from oracle/_ref).
 */
#ifndef DC_ORACLE_H
#define DC_ORACLE_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* compress types (impl/dataCompression.h:10-23) */
enum { DCO_CT_BITWISE = 5, DCO_CT_NP = 6, DCO_CT_MASK = 7, DCO_CT_OP = 11 };

/* to_absErrorBound_binary, dataCompression.c:5512-5522 (returns -1 where the reference falls off the end) */
int dco_n_eb(double eb);
/* toSmallDataset_float, :3543-3562.  y may alias x.  Returns min. */
float dco_min_subtract(const float* x, size_t n, float* y);
/* med_dataset_float, :3593-3620.  *type is left untouched when max >= 2^127 (as the reference does). */
float dco_med(const float* x, size_t n, int* type);
/* top 17 bits (sign, exponent, 8 mantissa bits) of medium: the char mask[17] of impl/pingpong.c:202-206 as an integer */
uint32_t dco_mask17(float medium);
/* do_crc32, :5524-5534 (zlib crc32, CRC-32/ISO-HDLC), own table implementation */
uint32_t dco_crc32(const uint8_t* p, size_t bytes);

/* One reference call myCompress_bitwise{,_np,_op,_mask}(x, n, (NULL,0,8)).
 * out must hold 4*n+8 bytes.  Returns the byte count; *pos_out (may be NULL) gets the reference's final pos. */
size_t dco_compress(int ct, const float* x, size_t n, double eb, int type, uint32_t mask17,
                    uint8_t* out, int* pos_out);
/* One reference call myDecompress_bitwise{,_np,_op,_mask}(s, bytes, n): decodes exactly n codes
 * (bits past the end of the buffer read as 0).  Returns 0, or -1 on a malformed code (CT=6 leading 1). */
int dco_decompress(int ct, const uint8_t* s, size_t bytes, size_t n, double eb, int type, uint32_t mask17,
                   float* out);

/* Chunk driver = "the reference routine run chunk-by-chunk": for each chunk of chunk_elems values
 * [min-subtract] -> [med/type/mask for CT=7] -> compress from (NULL,0,8); streams concatenated.
 * offsets[K+1] byte offsets; chunk_min/medium/type[K] (may be NULL); returns total bytes. */
size_t dco_compress_chunked(int ct, const float* x, size_t n, size_t chunk_elems, double eb, int subtract_min,
                            int fixed_type, uint32_t fixed_mask17, /* used when fixed_type >= 0 */
                            uint8_t* out, uint64_t* offsets, float* chunk_min, float* chunk_medium,
                            int32_t* chunk_type, uint32_t* chunk_crc);
int dco_decompress_chunked(int ct, const uint8_t* s, const uint64_t* offsets, size_t n, size_t chunk_elems,
                           double eb, int add_min, const float* chunk_min, int fixed_type,
                           uint32_t fixed_mask17, const float* chunk_medium, const int32_t* chunk_type,
                           float* out);

/* ---- float64 variants (dc_oracle_f64.c).  Same contracts with double data, 8*n+16 byte worst case,
 * and the CT=7 mask as the first 1+11+8 = 20 bits of the chunk mean (impl/k-means.c:527-530). ---- */
double dco_min_subtract_f64(const double* x, size_t n, double* y);           /* toSmallDataset_double :3522-3541 */
double dco_med_f64(const double* x, size_t n, int* type);                    /* med_dataset_double :3564-3591 */
uint32_t dco_mask20(double medium);
size_t dco_compress_f64(int ct, const double* x, size_t n, double eb, int type, uint32_t mask20,
                        uint8_t* out, int* pos_out);
int dco_decompress_f64(int ct, const uint8_t* s, size_t bytes, size_t n, double eb, int type, uint32_t mask20,
                       double* out);
size_t dco_compress_chunked_f64(int ct, const double* x, size_t n, size_t chunk_elems, double eb, int subtract_min,
                                int fixed_type, uint32_t fixed_mask20,
                                uint8_t* out, uint64_t* offsets, double* chunk_min, double* chunk_medium,
                                int32_t* chunk_type, uint32_t* chunk_crc);
int dco_decompress_chunked_f64(int ct, const uint8_t* s, const uint64_t* offsets, size_t n, size_t chunk_elems,
                               double eb, int add_min, const double* chunk_min, int fixed_type,
                               uint32_t fixed_mask20, const uint32_t* chunk_mask20, const int32_t* chunk_type,
                               double* out);
/* ---- Hamming SECDED of CT=10 (dc_oracle_hamming.c): hmLength :5587, hamming_encode :5742, hamming_decode :5752 ---- */
int dco_hm_length(int k);
int dco_hamming_encode(const uint8_t* bits, int bytes, char* c);          /* c: r + 1 chars; returns r */
int dco_hamming_decode(uint8_t* bits, char* c, int bytes, int r);         /* returns the reference's error_type 0..3 */

/* ---- byte-wise codec CT=1 (dc_oracle_bytewise.c): myCompress :3977, myDecompress :3942 ---- */
size_t dco_bytewise_compress(const float* x, size_t n, double eb, float* af, char* ac, int* ad);   /* returns the raw-float count */
void dco_bytewise_decompress(const float* af, const char* ac, const int* ad, size_t nc, size_t n, float* out);
size_t dco_bytewise_compress_f64(const double* x, size_t n, double eb, double* af, char* ac, int* ad);   /* myCompress_double :3815 */
void dco_bytewise_decompress_f64(const double* af, const char* ac, const int* ad, size_t nc, size_t n, double* out);   /* :3778 */

#ifdef __cplusplus
}
#endif
#endif
