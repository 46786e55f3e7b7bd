/* dc_b200.h -- C ABI of the B200-native float32 bit-wise codec (libdcb200.so).
 *
 * Drop-in boundary for the per-element lossy float compression hot path of
 * smallcat9603/data-compression, impl/dataCompression.c (CT=5, 6, 11, 7 and the CRC-protected
 * CT=8/9).  Plain pointers and sizes only; no torch / CUDA types in the signatures (a CUDA stream
 * is passed as void*, NULL = the legacy default stream).
 *
 * Two layers:
 *   (1) dcb_*  : device-resident, chunked, 64-bit-sized entry points.  One CHUNK == one call of the
 *               reference routine myCompress_bitwise*(chunk, n, (NULL,0,8)): the history restarts and the
 *               stream is byte aligned at every chunk boundary.  The byte stream of every chunk is
 *               bit-identical to the reference's; the chunk table (offsets / min / medium / type / crc)
 *               lives beside the stream, never inside it.
 *   (2) the reference's own symbols (dataCompression.h:67-99, 125-143) with unchanged signatures and
 *               host pointers, declared in dataCompression_b200.h; thin wrappers over (1).
 *
 * Every function returns DCB_OK (0) or a negative DCB_E* code; nothing exits the process.
 */
#ifndef DC_B200_H
#define DC_B200_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* compress types: impl/dataCompression.h:10-23 */
#define DCB_CT_BITWISE      5   /* myCompress_bitwise       dataCompression.c:3310 */
#define DCB_CT_BITWISE_NP   6   /* myCompress_bitwise_np    :2645 */
#define DCB_CT_BITWISE_MASK 7   /* myCompress_bitwise_mask  :2030 */
#define DCB_CT_BITWISE_CRC  8   /* CT=5 stream + do_crc32   impl/pingpong.c:260-274 */
#define DCB_CT_MASK_CRC     9   /* CT=7 stream + do_crc32   impl/pingpong.c:275-289 */
#define DCB_CT_BITWISE_OP  11   /* myCompress_bitwise_op    :577 */

/* flags */
#define DCB_F_SUBTRACT_MIN 1u   /* fuse toSmallDataset_float (:3543) per chunk on compress; add min back on decompress (impl/pingpong.c:562) */
#define DCB_F_FIXED_MASK   2u   /* CT=7/9: use (fixed_type, fixed_mask17) for every chunk instead of med_dataset_float (:3593) per chunk */
#define DCB_F_CRC          4u   /* also write do_crc32 (:5524) of every chunk's stream (implied by CT=8/9) */
#define DCB_F_EXACT_MEDIUM 8u   /* CT=7/9: d_chunk_medium holds med_dataset_float's exact float (default: only its 17 mask bits are defined) */

#define DCB_OK            0
#define DCB_EINVAL       -1     /* bad argument (ct, error bound, alignment, chunk size) */
#define DCB_ECAPACITY    -2     /* output / workspace too small */
#define DCB_ECUDA        -3     /* CUDA runtime error (dcb_last_cuda_error()) */
#define DCB_ENODEVICE    -4     /* no CUDA device: there is no CPU fallback */
#define DCB_EFORMAT      -5     /* malformed stream (CT=6 code with leading 1, dataCompression.c:2501) */

#define DCB_TILE_ELEMS 4096u    /* elements one thread block encodes per step */

const char* dcb_version(void);
const char* dcb_last_cuda_error(void);
int      dcb_device_count(void);

/* to_absErrorBound_binary (dataCompression.c:5512): smallest n in [0,99] with eb >= 2^-n, else -1 */
int      dcb_n_eb(double abs_error_bound);
/* chunk_elems must be 0 (one chunk = whole array, the legacy call) or a multiple of DCB_TILE_ELEMS; dcb_decompress takes chunked
 * streams with chunk_elems <= 2^26 (a decoding lane's bit position is 32 bits wide); one chunk = whole array has no such limit
 * below 2^32 elements (it is decoded in parallel inside the chunk). */
uint64_t dcb_num_chunks(uint64_t n, uint64_t chunk_elems);
/* worst case of the stream: every code is <= 32 bits */
uint64_t dcb_max_stream_bytes(uint64_t n, uint64_t chunk_elems);
uint64_t dcb_workspace_bytes(uint64_t n, uint64_t chunk_elems);

/* Compress n float32 at d_in (device) into d_stream (device, 16-byte aligned, capacity >= dcb_max_stream_bytes).
 *   d_chunk_offsets[K+1] : byte offset of every chunk's stream; [K] = total bytes          (always written)
 *   d_chunk_min[K]       : per-chunk minimum           (required with DCB_F_SUBTRACT_MIN, else may be NULL)
 *   d_chunk_medium[K], d_chunk_type[K] : med_dataset_float results (CT=7/9 without DCB_F_FIXED_MASK)
 *   d_chunk_crc[K]       : CRC-32 of every chunk's stream                     (CT=8/9 or DCB_F_CRC)
 * Asynchronous on cuda_stream. */
int dcb_compress(int ct, const float* d_in, uint64_t n, uint64_t chunk_elems, double abs_error_bound,
                 uint32_t flags, int fixed_type, uint32_t fixed_mask17,
                 uint8_t* d_stream, uint64_t stream_capacity, uint64_t* d_chunk_offsets,
                 float* d_chunk_min, float* d_chunk_medium, int32_t* d_chunk_type, uint32_t* d_chunk_crc,
                 void* d_workspace, uint64_t workspace_bytes, void* cuda_stream);

/* dcb_compress without caller-provided scratch and without any allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, serving loops).  The workspace is
 * a static device array inside the library (DCB_STATIC_WS_BYTES on every GPU that loads it); n is limited to
 * DCB_STATIC_MAX_ELEMS elements per call (larger n: DCB_ECAPACITY, text in dcb_last_cuda_error).  All calls on one device
 * share that array, so calls that overlap on the GPU race: issue them on one stream or order them with events.
 * Results are those of dcb_compress, bit for bit. */
#define DCB_STATIC_MAX_ELEMS (1ull << 20)
#define DCB_STATIC_WS_BYTES ((1ull << 22) + (1ull << 16))   /* 256 slots of 16 KB + tables of 2^20 elements, rounded up */
uint64_t dcb_static_max_elems(void);
int dcb_compress_static(int ct, const float* d_in, uint64_t n, uint64_t chunk_elems, double abs_error_bound,
                        uint32_t flags, int fixed_type, uint32_t fixed_mask17,
                        uint8_t* d_stream, uint64_t stream_capacity, uint64_t* d_chunk_offsets,
                        float* d_chunk_min, float* d_chunk_medium, int32_t* d_chunk_type, uint32_t* d_chunk_crc,
                        void* cuda_stream);

/* Mirror of dcb_compress: decodes exactly n values (chunk c holds min(chunk_elems, n - c*chunk_elems)).
 * d_stream must be readable up to the next multiple of 16 bytes past d_chunk_offsets[K] (the decoder fetches 16-byte blocks;
 * a stream of dcb_max_stream_bytes capacity always is).  The chunk offsets are TRUSTED: they must be monotonic and inside the
 * stream buffer (the entry point does not know the buffer's capacity); a caller that receives them over an unprotected link
 * checks them first (transport.recv_compressed does, against n).
 * d_status (device int32, may be NULL) receives DCB_OK / DCB_EFORMAT. */
int dcb_decompress(int ct, const uint8_t* d_stream, const uint64_t* d_chunk_offsets, uint64_t n,
                   uint64_t chunk_elems, double abs_error_bound, uint32_t flags, int fixed_type,
                   uint32_t fixed_mask17, const float* d_chunk_min, const float* d_chunk_medium,
                   const int32_t* d_chunk_type, float* d_out, int32_t* d_status, void* cuda_stream);

/* Measurement aid (bench.py): with dcb_profile_enable(1) the calling thread's dcb_compress calls that take the slot path
 * (one chunk == one tile) record CUDA events around the encoder kernel and the compaction kernel; dcb_profile_last_ms
 * waits for the last such call and returns the two durations.  No effect on results. */
int dcb_profile_enable(int on);
int dcb_profile_last_ms(float* encode_kernel_ms, float* compact_kernel_ms);

/* do_crc32 (dataCompression.c:5524) of each chunk's stream, as a stand-alone pass */
int dcb_crc32_chunks(const uint8_t* d_stream, const uint64_t* d_chunk_offsets, uint64_t num_chunks,
                     uint32_t* d_chunk_crc, void* cuda_stream);

/* toSmallDataset_float (:3543) and med_dataset_float (:3593) on device, one result per chunk */
int dcb_chunk_stats(const float* d_in, uint64_t n, uint64_t chunk_elems, int want_medium,
                    float* d_chunk_min, float* d_chunk_medium, int32_t* d_chunk_type, void* cuda_stream);

/* med_dataset_float (:3593) of the whole array as it is (no min-subtract): exact sequential float32 mean, and the
 * type from max (0 = "left at the caller's initial value") */
int dcb_chunk_medium(const float* d_in, uint64_t n, float* d_medium, int32_t* d_type, void* cuda_stream);

/* bit_flip (dataCompression.c:5858) test hook: flips bit (7 - bit_index%8) of byte bit_index/8 for each index */
int dcb_flip_bits(uint8_t* d_stream, const uint64_t* d_bit_indices, uint64_t count, void* cuda_stream);

/* ---- byte-wise codec CT=1: myCompress (dataCompression.c:3977) / myDecompress (:3942), one reference call
 * per chunk.  Fixed-pitch outputs: chunk c owns elements [c*chunk_elems, ...) of d_float (its raw floats, d_chunk_nfloat[c] of
 * them), d_char (hit types 'a'..'d') and d_disp (their 1-based positions inside the chunk); hits = chunk length - raw floats.
 * chunk_elems = 0: one chunk = the whole array (the legacy call). */
uint64_t dcb_bytewise_workspace_bytes(uint64_t n, uint64_t chunk_elems);
int dcb_bytewise_compress(const float* d_in, uint64_t n, uint64_t chunk_elems, double abs_error_bound, float* d_float,
                          uint8_t* d_char, int32_t* d_disp, uint32_t* d_chunk_nfloat, void* d_workspace,
                          uint64_t workspace_bytes, void* cuda_stream);
int dcb_bytewise_decompress(const float* d_float, const uint8_t* d_char, const int32_t* d_disp, const uint32_t* d_chunk_nfloat,
                            uint64_t n, uint64_t chunk_elems, float* d_out, void* cuda_stream);
/* float64: myCompress_double (:3815) / myDecompress_double (:3778), the k-means / MM / LU callers; same layout and workspace */
int dcb_bytewise_compress_f64(const double* d_in, uint64_t n, uint64_t chunk_elems, double abs_error_bound, double* d_double,
                              uint8_t* d_char, int32_t* d_disp, uint32_t* d_chunk_ndouble, void* d_workspace,
                              uint64_t workspace_bytes, void* cuda_stream);
int dcb_bytewise_decompress_f64(const double* d_double, const uint8_t* d_char, const int32_t* d_disp,
                                const uint32_t* d_chunk_ndouble, uint64_t n, uint64_t chunk_elems, double* d_out, void* cuda_stream);

/* ---- Hamming SECDED of CT=10: hamming_encode / hamming_decode (dataCompression.c:5742, :5752), one
 * call per block of the finished stream as impl/pingpong.c:296-311 / :487-505 make them.  Block b covers bytes
 * [b*block_bytes, min(total_bytes, (b+1)*block_bytes)); its r = dcb_hm_length(8 * its bytes) check bits c[0..r-1] and the
 * parity bit c[r] are packed into d_check[b] (bit i = c[i]).  Decode recomputes the syndrome, classifies (0 no error,
 * 1 two-bit error, 2 parity bit, 3 one bit) and rectifies single errors in d_bits / d_check in place. */
int dcb_hm_length(int k);                                                    /* hmLength, dataCompression.c:5587 */
int dcb_hamming_encode(const uint8_t* d_bits, uint64_t total_bytes, uint32_t block_bytes, uint32_t* d_check,
                       int32_t* d_r, void* cuda_stream);
int dcb_hamming_decode(uint8_t* d_bits, uint64_t total_bytes, uint32_t block_bytes, uint32_t* d_check,
                       int32_t* d_error_type, void* cuda_stream);

/* ---- float64 codecs: the reference's *_double entry points ------------------------------------
 * dcb_compress_f64 replaces, per chunk, toSmallDataset_double (dataCompression.c:3522) -> [med_dataset_double (:3564) +
 * the mask chars, impl/k-means.c:525-530] -> myCompress_bitwise_double (:3189) / _np (:2633) / _mask (:1590) / _op (:355)
 * from (NULL,0,8) [-> do_crc32]; dcb_decompress_f64 replaces myDecompress_bitwise_double (:2656) / _np (:2286) / _mask
 * (:1199) / _op (:476) and the "+ min" loop (impl/k-means.c:597-598).  Same arguments as the float32 calls except:
 *   d_chunk_mask20[K] : the first 1+11+8 bits of the chunk mean (what the reference cuts out of doubletostr(medium));
 *   d_chunk_medium[K] : the exact mean, written only with DCB_F_EXACT_MEDIUM (may be NULL otherwise);
 *   fixed_type 0..10, fixed_mask20 < 2^20; worst case stream 8*n + 16 bytes. */
uint64_t dcb_max_stream_bytes_f64(uint64_t n, uint64_t chunk_elems);
uint64_t dcb_workspace_bytes_f64(uint64_t n, uint64_t chunk_elems);
int dcb_compress_f64(int ct, const double* d_in, uint64_t n, uint64_t chunk_elems, double abs_error_bound,
                     uint32_t flags, int fixed_type, uint32_t fixed_mask20,
                     uint8_t* d_stream, uint64_t stream_capacity, uint64_t* d_chunk_offsets,
                     double* d_chunk_min, uint32_t* d_chunk_mask20, int32_t* d_chunk_type, double* d_chunk_medium,
                     uint32_t* d_chunk_crc, void* d_workspace, uint64_t workspace_bytes, void* cuda_stream);
/* toSmallDataset_double (:3522) / med_dataset_double (:3564) per chunk as a stand-alone pass; after dcb_compress_f64 the first
 * 8 bytes of d_workspace hold the code bits of the last chunk (the legacy `pos`). */
int dcb_chunk_stats_f64(const double* d_in, uint64_t n, uint64_t chunk_elems, int subtract_min, int want_medium,
                        double* d_chunk_min, uint32_t* d_chunk_mask20, int32_t* d_chunk_type, double* d_chunk_medium,
                        void* cuda_stream);
/* d_stream: 4-byte aligned at least.  When it is 16-byte aligned the decoder stages it in 16-byte blocks, and, as for float32, it
 * must then be readable up to the next multiple of 16 bytes past d_chunk_offsets[K] (dcb_max_stream_bytes_f64 leaves that room).
 * One chunk of more than 4096 elements (chunk_elems = 0 on a long array: the reference's own call shape) is decoded in parallel
 * inside the chunk; its scratch comes from the library's stream-ordered memory pool. */
int dcb_decompress_f64(int ct, const uint8_t* d_stream, const uint64_t* d_chunk_offsets, uint64_t n,
                       uint64_t chunk_elems, double abs_error_bound, uint32_t flags, int fixed_type,
                       uint32_t fixed_mask20, const double* d_chunk_min, const uint32_t* d_chunk_mask20,
                       const int32_t* d_chunk_type, double* d_out, int32_t* d_status, void* cuda_stream);

/* Himeno halo faces, device resident.  dcb_gather_face replaces transform_3d_array_to_1d_array
 * (impl/dataCompression.c:3741-3775; p is float[MI][MJ][MK], ijk = 1/2/3 fixes i/j/k = v, the face has jmax*kmax /
 * imax*kmax / imax*jmax elements in the reference's loop order); dcb_scatter_face replaces the loops that write a
 * decompressed face back, p[face] = d_in[pointer++] + add (impl/himenoBMTxps.c:560-568, :640-648). */
int dcb_gather_face(const float* d_p, int MI, int MJ, int MK, int ijk, int v, int imax, int jmax, int kmax,
                    float* d_out, void* cuda_stream);
int dcb_scatter_face(float* d_p, int MI, int MJ, int MK, int ijk, int v, int imax, int jmax, int kmax,
                     const float* d_in, float add, void* cuda_stream);

/* Himeno Jacobi sweep on the device (impl/himenoBMTxps.c:247-284): arrays as in the driver -- p, bnd, wrk1, wrk2 are
 * float[MI][MJ][MK], a is float[4][MI][MJ][MK], b and c float[3][MI][MJ][MK].  dcb_himeno_jacobi writes wrk2 on the interior
 * (bit exact: every operation rounded in the C expression's order) and *d_gosa = sum ss^2 (double); dcb_himeno_copy_interior
 * is the p = wrk2 loop (:279-282).  The halo exchange between the two is halo.exchange_axis / dcb_gather_face etc. */
int dcb_himeno_jacobi(const float* d_p, const float* d_a, const float* d_b, const float* d_c, const float* d_bnd,
                      const float* d_wrk1, float* d_wrk2, int MI, int MJ, int MK, int imax, int jmax, int kmax,
                      float omega, double* d_gosa, void* cuda_stream);
int dcb_himeno_copy_interior(float* d_p, const float* d_wrk2, int MI, int MJ, int MK, int imax, int jmax, int kmax,
                             void* cuda_stream);

/* Frees the CALLING THREAD's cached device buffers, pinned staging and stream of the host-pointer layer (dcb_*_host and the
 * reference's symbols of dataCompression_b200.h keep them between calls, about 12-16 bytes per element).  They are also freed
 * when the thread exits. */
int dcb_host_release(void);

/* ---- host-pointer convenience (what the legacy symbols and the end-to-end bench call) ----------------
 * Pinned staging + H2D + kernels + D2H inside.  h_stream needs dcb_max_stream_bytes(); tables need K(+1).
 * Blocking and thread safe: every calling thread has its own stream and device buffers.  Bulk copies go out in slices of
 * DCB_COPY_SLICE_MB megabytes (environment, default 128) under a FIFO lock per direction, so that concurrent callers keep
 * both copy engines busy: measured on one B200 with pinned buffers, one caller 24.9 GB/s of input (compress + decompress),
 * four callers 43.5 GB/s = 0.95 of what plain concurrent cudaMemcpyAsync in both directions reaches on that box.
 * PAGEABLE buffers (malloc, what the reference's drivers pass) of 8 MB or more are copied by DCB_STAGED_THREADS helpers (default 6)
 * through pinned staging: 17 GB/s for one caller instead of the driver's 6.3. */
int dcb_compress_host(int ct, const float* h_in, uint64_t n, uint64_t chunk_elems, double abs_error_bound,
                      uint32_t flags, int fixed_type, uint32_t fixed_mask17,
                      uint8_t* h_stream, uint64_t* h_chunk_offsets, float* h_chunk_min, float* h_chunk_medium,
                      int32_t* h_chunk_type, uint32_t* h_chunk_crc);
int dcb_decompress_host(int ct, const uint8_t* h_stream, const uint64_t* h_chunk_offsets, uint64_t n,
                        uint64_t chunk_elems, double abs_error_bound, uint32_t flags, int fixed_type,
                        uint32_t fixed_mask17, const float* h_chunk_min, const float* h_chunk_medium,
                        const int32_t* h_chunk_type, float* h_out);

/* Bit length (before the final byte round-up) of the whole stream produced by this thread's most recent
 * dcb_compress_host call: the legacy `pos` output is 8 - bits%8 (8 when aligned), dataCompression.c:5456-5489. */
uint64_t dcb_last_stream_bits(void);

#ifdef __cplusplus
}
#endif
#endif
