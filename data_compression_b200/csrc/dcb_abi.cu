// dcb_abi.cu -- extern "C" entry points of libdcb200.so (include/dc_b200.h).
// Host logic only: argument checks, error-bound constants, launch configuration.  No CPU codec lives
// here: without a CUDA device every entry point returns DCB_ENODEVICE.
#include <atomic>
#include <thread>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <condition_variable>
#include "../../include/dc_b200.h"
#include "dcb_common.cuh"
#include "dcb_encode.cuh"
#include "dcb_encode_slots.cuh"
#include "dcb_encode_warp.cuh"
#include "dcb_decode.cuh"
#include "dcb_decode_single.cuh"
#include "dcb_stats.cuh"
#include "dcb_crc.cuh"
#include "dcb_face.cuh"
#include "dcb_f64.cuh"
#include "dcb_decode_single_f64.cuh"
#include "dcb_f64_stats_long.cuh"
#include "dcb_f64_decode_ring.cuh"
#include "dcb_hamming.cuh"
#include "dcb_bytewise.cuh"
static int ensure_crc_tables(cudaStream_t st);
#include <math.h>
#include <stdio.h>
#include <string.h>

static thread_local char g_err[256] = "";
static int cuda_fail(cudaError_t e, const char* what) {
  snprintf(g_err, sizeof g_err, "%s: %s", what, cudaGetErrorString(e));
  return DCB_ECUDA;
}
#define CK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return cuda_fail(e_, #call); } while (0)

extern "C" const char* dcb_version(void) { return "dcb200 0.1 (sm_100a)"; }
extern "C" const char* dcb_last_cuda_error(void) { return g_err; }
extern "C" int dcb_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}

// to_absErrorBound_binary, dataCompression.c:5512-5522 (2^-n is exact in double, no pow() needed)
extern "C" int dcb_n_eb(double eb) {
  for (int n = 0; n < 100; n++)
    if (eb >= ldexp(1.0, -n)) return n;
  return -1;
}

static int make_const(double eb, CodecConst* k) {
  const int n_eb = dcb_n_eb(eb);
  if (!(eb > 0.0) || n_eb < 0) return DCB_EINVAL;      // reference: undefined (falls off the end of the function)
  k->n_eb = n_eb;
  k->n_eb_m118 = n_eb - 118;
  // the reference compares float32 values with the double literal absErrorBound in DOUBLE precision
  // (:3390 fabs(x) < eb, :3399 diff_min <= eb); the same predicates on float32 thresholds:
  float t = (float)eb;                                  // may round up
  if (isinf(t)) t = 3.402823466e38f;
  while ((double)t > eb) t = nextafterf(t, -INFINITY);
  k->tp = t;                                            // largest float <= eb
  while ((double)t >= eb) t = nextafterf(t, -INFINITY);
  k->tz = t;                                            // largest float <  eb
  return DCB_OK;
}

static int base_ct(int ct, uint32_t* flags) {
  switch (ct) {
    case DCB_CT_BITWISE: return CT_BITWISE;
    case DCB_CT_BITWISE_NP: return CT_NP;
    case DCB_CT_BITWISE_MASK: return CT_MASK;
    case DCB_CT_BITWISE_OP: return CT_OP;
    case DCB_CT_BITWISE_CRC: *flags |= DCB_F_CRC; return CT_BITWISE;   // impl/pingpong.c:260-274
    case DCB_CT_MASK_CRC: *flags |= DCB_F_CRC; return CT_MASK;         // impl/pingpong.c:275-289
    default: return -1;
  }
}

static bool single_chunk(uint64_t n, uint64_t chunk) { return chunk == 0 || chunk >= n; }

extern "C" uint64_t dcb_num_chunks(uint64_t n, uint64_t chunk) {
  if (n == 0) return 0;
  return single_chunk(n, chunk) ? 1 : (n + chunk - 1) / chunk;
}
extern "C" uint64_t dcb_max_stream_bytes(uint64_t n, uint64_t chunk) { (void)chunk; return 4 * n + 16; }
static uint64_t num_tiles(uint64_t n) { return (n + DCB_TILE - 1) / DCB_TILE; }
// "one chunk == one tile": min / medium are fused in the encoder and blocks are independent (slot path)
static bool chunk_is_tile(uint64_t n, uint64_t chunk) {
  return num_tiles(n) == 1 || (!single_chunk(n, chunk) && chunk == DCB_TILE);
}
// workspace layout: [0,256) header (ticket, scanner claim, last_bits @8, last_tile_bits @16) | status[tiles] u64
//   slot path only:  | gsum[G] u32 | sizes[K] u32 | pad to 256 | K slots of DCB_SLOT_PITCH bytes
static uint64_t ws_head_bytes(uint64_t n) { return 256 + 8 * num_tiles(n); }
static uint64_t ws_groups(uint64_t K) { return (K + DCB_GROUP - 1) / DCB_GROUP; }
static uint64_t ws_slots_at(uint64_t n) {
  const uint64_t K = num_tiles(n);
  return (ws_head_bytes(n) + 4 * ws_groups(K) + 4 * K + 255) & ~(uint64_t)255;
}
// Encoder for one-tile chunks.  Default: the warp-autonomous kernel (fastest for every CT: 0.42 / 0.61 ms against 0.47 / 0.73 ms
// of the block-per-chunk slot kernel and 0.65 / 0.96 ms of the single-pass kernel, CT=5 / CT=7 on the smooth 2^28 field).
// DCB_ENCODER = warp | slots | single forces one kernel (A/B runs; the parity tests cover all three).
static int encoder_choice() {
  static const int v = [] {
    const char* e = getenv("DCB_ENCODER");
    if (getenv("DCB_SINGLE_PASS") || (e && !strcmp(e, "single"))) return 0;
    if (e && !strcmp(e, "slots")) return 1;
    return 2;
  }();
  return v;
}
extern "C" uint64_t dcb_workspace_bytes(uint64_t n, uint64_t chunk) {
  if (n && chunk_is_tile(n, chunk)) return ws_slots_at(n) + num_tiles(n) * (uint64_t)DCB_SLOT_PITCH;
  return ws_head_bytes(n);
}

static int num_sms() {                                 // per device (a process may drive several GPUs)
  static std::atomic<int> cache[64];
  int dev = 0, n = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  n = cache[dev].load(std::memory_order_relaxed);
  if (n == 0) {
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cache[dev].store(n, std::memory_order_relaxed);
  }
  return n;
}

// persistent launch: DCB_ENC_BLOCKS_PER_SM resident blocks per SM pull tiles from the ticket counter
template <int CT>
static cudaError_t launch_encode(const EncArgs& a, cudaStream_t st) {
  {                                                   // per device; cheap enough to set on every launch
    cudaError_t e = cudaFuncSetAttribute(dcb_encode_kernel<CT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(EncSmem));
    if (e != cudaSuccess) return e;
  }
  // one block becomes the prefix scanner, the others are workers
  const unsigned grid = 1u + (unsigned)min((uint64_t)a.num_tiles, (uint64_t)num_sms() * DCB_ENC_BLOCKS_PER_SM - 1u);
  dcb_encode_kernel<CT><<<grid, DCB_NT, sizeof(EncSmem), st>>>(a);
  return cudaGetLastError();
}

// Optional per-kernel timing of the slot path (bench.py's roofline line): when enabled, dcb_compress records CUDA events on its
// stream around the encoder kernel and around the compaction kernel; dcb_profile_last_ms reads them back.
static thread_local bool g_prof_on = false;
static thread_local cudaEvent_t g_prof_ev[3] = {nullptr, nullptr, nullptr};
extern "C" int dcb_profile_enable(int on) {
  if (on && !g_prof_ev[0]) for (int i = 0; i < 3; i++) CK(cudaEventCreate(&g_prof_ev[i]));
  g_prof_on = on != 0;
  return DCB_OK;
}
extern "C" int dcb_profile_last_ms(float* encode_ms, float* compact_ms) {
  if (!g_prof_ev[0]) return DCB_EINVAL;
  CK(cudaEventSynchronize(g_prof_ev[2]));
  if (encode_ms) CK(cudaEventElapsedTime(encode_ms, g_prof_ev[0], g_prof_ev[1]));
  if (compact_ms) CK(cudaEventElapsedTime(compact_ms, g_prof_ev[1], g_prof_ev[2]));
  return DCB_OK;
}
#define PROF_MARK(i) do { if (g_prof_on) { cudaError_t pe_ = cudaEventRecord(g_prof_ev[i], st); if (pe_ != cudaSuccess) return pe_; } } while (0)

template <int CT>
static cudaError_t launch_encode_warp(const EncArgs& a, const CompactArgs& ca, cudaStream_t st) {
  const int smem = (int)(sizeof(WarpBuf) * (DCB_WNT / 32));
  {
    cudaError_t e = cudaFuncSetAttribute(dcb_encode_warp_kernel<CT>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return e;
  }
  const uint64_t blocks = ((uint64_t)a.num_tiles + DCB_WNT / 32 - 1) / (DCB_WNT / 32);
  const unsigned grid = (unsigned)min(blocks, (uint64_t)num_sms() * (CT == CT_MASK ? DCB_WARP_BLOCKS_CT7 : DCB_WARP_BLOCKS));
  PROF_MARK(0);
  dcb_encode_warp_kernel<CT><<<grid, DCB_WNT, smem, st>>>(a);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  PROF_MARK(1);
  dcb_compact_kernel<<<(unsigned)ws_groups(ca.K), DCB_COMPACT_NT, 0, st>>>(ca);
  PROF_MARK(2);
  return cudaGetLastError();
}

template <int CT>
static cudaError_t launch_encode_slots(const EncArgs& a, const CompactArgs& ca, cudaStream_t st) {
  {                                                   // per device; cheap enough to set on every launch
    cudaError_t e = cudaFuncSetAttribute(dcb_encode_slots_kernel<CT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(SlotSmem));
    if (e != cudaSuccess) return e;
  }
  const unsigned grid = (unsigned)min((uint64_t)a.num_tiles, (uint64_t)num_sms() * DCB_SLOT_BLOCKS_PER_SM);
  PROF_MARK(0);
  dcb_encode_slots_kernel<CT><<<grid, DCB_NT, sizeof(SlotSmem), st>>>(a);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  PROF_MARK(1);
  dcb_compact_kernel<<<(unsigned)ws_groups(ca.K), DCB_COMPACT_NT, 0, st>>>(ca);
  PROF_MARK(2);
  return cudaGetLastError();
}

extern "C" int dcb_chunk_stats(const float* d_in, uint64_t n, uint64_t chunk, int want_medium,
                               float* d_chunk_min, float* d_chunk_medium, int32_t* d_chunk_type, void* stream) {
  if (dcb_device_count() == 0) return DCB_ENODEVICE;
  if (n == 0) return DCB_OK;
  if (!d_in || !d_chunk_min || (want_medium && (!d_chunk_medium || !d_chunk_type))) return DCB_EINVAL;
  StatsArgs s;
  s.in = d_in; s.n = n; s.chunk = single_chunk(n, chunk) ? n : chunk;
  s.want_medium = want_medium; s.subtract_min = 1;
  s.cmin = d_chunk_min; s.cmed = d_chunk_medium; s.ctype = d_chunk_type;
  dcb_stats_kernel<<<(unsigned)dcb_num_chunks(n, chunk), 256, 0, (cudaStream_t)stream>>>(s);
  CK(cudaGetLastError());
  return DCB_OK;
}

extern "C" int dcb_chunk_medium(const float* d_in, uint64_t n, float* d_medium, int32_t* d_type, void* stream) {
  if (dcb_device_count() == 0) return DCB_ENODEVICE;
  if (n == 0) return DCB_OK;
  if (!d_in || !d_medium || !d_type) return DCB_EINVAL;
  StatsArgs s;
  s.in = d_in; s.n = n; s.chunk = n;
  s.want_medium = 1; s.subtract_min = 0;
  s.cmin = nullptr; s.cmed = d_medium; s.ctype = d_type;
  dcb_stats_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(s);
  CK(cudaGetLastError());
  return DCB_OK;
}

extern "C" int dcb_compress(int ct, const float* d_in, uint64_t n, uint64_t chunk, double eb, uint32_t flags,
                            int fixed_type, uint32_t fixed_mask17, uint8_t* d_stream, uint64_t capacity,
                            uint64_t* d_offsets, float* d_cmin, float* d_cmed, int32_t* d_ctype, uint32_t* d_ccrc,
                            void* d_ws, uint64_t ws_bytes, void* stream) {
  if (dcb_device_count() == 0) return DCB_ENODEVICE;
  cudaStream_t st = (cudaStream_t)stream;
  const int bct = base_ct(ct, &flags);
  EncArgs a;
  memset(&a, 0, sizeof a);
  if (bct < 0 || make_const(eb, &a.k) != DCB_OK || !d_offsets) return DCB_EINVAL;
  if (n == 0) { CK(cudaMemsetAsync(d_offsets, 0, sizeof(uint64_t), st)); return DCB_OK; }
  const bool single = single_chunk(n, chunk);
  if (!single && (chunk % DCB_TILE) != 0) return DCB_EINVAL;
  if (!d_in || !d_stream || ((uintptr_t)d_stream & 15) || ((uintptr_t)d_in & 3)) return DCB_EINVAL;
  if (n >= (1ull << 40) || (!single && chunk >= (1ull << 32))) return DCB_EINVAL;
  const uint64_t K = dcb_num_chunks(n, chunk);
  const bool per_chunk_mask = (bct == CT_MASK) && !(flags & DCB_F_FIXED_MASK);
  if ((flags & DCB_F_SUBTRACT_MIN) && !d_cmin) return DCB_EINVAL;
  if (per_chunk_mask && (!d_cmed || !d_ctype)) return DCB_EINVAL;
  if ((flags & DCB_F_FIXED_MASK) && (fixed_type < 0 || fixed_type > 7 || fixed_mask17 >= (1u << 17))) return DCB_EINVAL;
  if ((flags & DCB_F_CRC) && !d_ccrc) return DCB_EINVAL;
  if (capacity < dcb_max_stream_bytes(n, chunk)) return DCB_ECAPACITY;
  if (!d_ws || ws_bytes < dcb_workspace_bytes(n, chunk)) return DCB_ECAPACITY;

  a.in = d_in; a.n = n;
  a.chunk = single ? n : chunk;
  a.num_tiles = (uint32_t)num_tiles(n);
  a.tiles_per_chunk = single ? 0u : (uint32_t)(chunk / DCB_TILE);
  a.flags = flags & 15u;
  a.fixed_type = fixed_type; a.fixed_mask17 = fixed_mask17;
  a.out = d_stream; a.offsets = d_offsets; a.cmin = d_cmin; a.cmed = d_cmed; a.ctype = d_ctype;
  a.ticket = (unsigned int*)d_ws;
  a.scanner_claim = (unsigned int*)((char*)d_ws + 4);
  a.last_bits = (unsigned long long*)((char*)d_ws + 8);
  a.status = (unsigned long long*)((char*)d_ws + 256);
  if (((uintptr_t)d_in & 15) == 0) a.flags |= DCB_ENC_IN_ALIGNED;
#ifdef DCB_DEBUG_MED
  if (const char* dm = getenv("DCB_DEBUG_MED")) a.flags |= (uint32_t)atoi(dm) << 16;
#endif

  // a chunk is "one tile" when every chunk fits a single thread block: min / medium are fused in the encoder
  const bool is_tile = chunk_is_tile(n, chunk);
  if (is_tile && encoder_choice() != 0) {
    // slot path: independent blocks write fixed-pitch slots, a second kernel builds the chunk table and compacts
    char* ws = (char*)d_ws;
    a.last_tile_bits = (uint32_t*)(ws + 16);
    a.gsum = (uint32_t*)(ws + ws_head_bytes(n));
    a.sizes = a.gsum + ws_groups(K);
    a.slots = (uint8_t*)(ws + ws_slots_at(n));
    CK(cudaMemsetAsync(d_ws, 0, (size_t)(ws_head_bytes(n) + 4 * ws_groups(K)), st));
    CompactArgs ca;
    ca.slots = a.slots; ca.sizes = a.sizes; ca.gsum = a.gsum; ca.last_tile_bits = a.last_tile_bits; ca.K = K;
    ca.out = d_stream; ca.offsets = d_offsets; ca.last_bits = a.last_bits; ca.pitch = DCB_SLOT_PITCH;
    // CT=8/9: the CRC of every chunk is computed by the compaction kernel while it copies the chunk (no third pass)
    static const bool crc_separate = getenv("DCB_CRC_SEPARATE") != nullptr;      // A/B: the r1 stand-alone kernel
    ca.crc = ((flags & DCB_F_CRC) && !crc_separate) ? d_ccrc : nullptr;
    if (ca.crc) { int rc_ = ensure_crc_tables(st); if (rc_) return rc_; }
    cudaError_t es;
    if (encoder_choice() == 2) {
      switch (bct) {
        case CT_BITWISE: es = launch_encode_warp<CT_BITWISE>(a, ca, st); break;
        case CT_NP: es = launch_encode_warp<CT_NP>(a, ca, st); break;
        case CT_MASK: es = launch_encode_warp<CT_MASK>(a, ca, st); break;
        default: es = launch_encode_warp<CT_OP>(a, ca, st); break;
      }
    } else {
      switch (bct) {
        case CT_BITWISE: es = launch_encode_slots<CT_BITWISE>(a, ca, st); break;
        case CT_NP: es = launch_encode_slots<CT_NP>(a, ca, st); break;
        case CT_MASK: es = launch_encode_slots<CT_MASK>(a, ca, st); break;
        default: es = launch_encode_slots<CT_OP>(a, ca, st); break;
      }
    }
    CK(es);
    if ((flags & DCB_F_CRC) && !ca.crc) {
      { int rc_ = ensure_crc_tables(st); if (rc_) return rc_; }
      dcb_crc_kernel<<<(unsigned)((K + 3) / 4), 128, 0, st>>>(d_stream, d_offsets, K, d_ccrc);
      CK(cudaGetLastError());
    }
    return DCB_OK;
  }
  if (!is_tile) {
    // tile boundaries inside a chunk fall on arbitrary BITS: boundary words are OR-ed into a zeroed stream
    a.flags |= DCB_ENC_ATOMIC_OUT;
    CK(cudaMemsetAsync(d_stream, 0, (size_t)(4 * n + 16), st));
    if ((flags & DCB_F_SUBTRACT_MIN) || per_chunk_mask) {
      StatsArgs s;
      s.in = d_in; s.n = n; s.chunk = a.chunk; s.want_medium = per_chunk_mask ? 1 : 0;
      s.subtract_min = (flags & DCB_F_SUBTRACT_MIN) ? 1 : 0;
      s.cmin = d_cmin; s.cmed = d_cmed; s.ctype = d_ctype;
      dcb_stats_kernel<<<(unsigned)K, 256, 0, st>>>(s);
      CK(cudaGetLastError());
      a.flags |= DCB_ENC_STATS_READY;
    }
  }
  CK(cudaMemsetAsync(d_ws, 0, (size_t)ws_head_bytes(n), st));
  cudaError_t e;
  switch (bct) {
    case CT_BITWISE: e = launch_encode<CT_BITWISE>(a, st); break;
    case CT_NP: e = launch_encode<CT_NP>(a, st); break;
    case CT_MASK: e = launch_encode<CT_MASK>(a, st); break;
    default: e = launch_encode<CT_OP>(a, st); break;
  }
  CK(e);
  if (flags & DCB_F_CRC) {
    { int rc_ = ensure_crc_tables(st); if (rc_) return rc_; }
    dcb_crc_kernel<<<(unsigned)((K + 3) / 4), 128, 0, st>>>(d_stream, d_offsets, K, d_ccrc);
    CK(cudaGetLastError());
  }
  return DCB_OK;
}

// ---- dcb_compress_static: dcb_compress for callers that may not allocate ------------------------------------------------
// No counterpart in the reference; added (callers under CUDA graph capture, tight
// serving loops, harnesses that forbid cudaMalloc).  The workspace dcb_compress needs (ticket, group sums, chunk sizes, one
// 16 KB slot per chunk) lives in a file-scope device array instead of caller memory: every GPU that loads the library holds
// DCB_STATIC_WS_BYTES until exit.  In HOST code the array's name is only a stand-in, not a device address (a kernel given
// it faults, or silently uses the wrong memory where the GPU can reach host memory): the address is taken with
// cudaGetSymbolAddress, once per device, with that device current.  Every call shares the array and it keeps the last
// call's contents; what must start at zero (header, group sums) is cleared by dcb_compress itself on the caller's stream.
// Calls that overlap on the GPU (different streams or host threads, same device) RACE on it: serialise them.
static __device__ __align__(256) unsigned char g_static_ws[DCB_STATIC_WS_BYTES];

static int static_ws_ptr(void** out) {
  static std::atomic<void*> cache[64];
  int dev = 0;
  CK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) { snprintf(g_err, sizeof g_err, "dcb_compress_static: device ordinal %d out of range", dev); return DCB_EINVAL; }
  void* p = cache[dev].load(std::memory_order_acquire);
  if (!p) {
    CK(cudaGetSymbolAddress(&p, g_static_ws));          // the current device's instance of the array
    cache[dev].store(p, std::memory_order_release);
  }
  *out = p;
  return DCB_OK;
}

extern "C" uint64_t dcb_static_max_elems(void) { return DCB_STATIC_MAX_ELEMS; }

extern "C" int dcb_compress_static(int ct, const float* d_in, uint64_t n, uint64_t chunk, double eb, uint32_t flags,
                                   int fixed_type, uint32_t fixed_mask17, uint8_t* d_stream, uint64_t capacity,
                                   uint64_t* d_offsets, float* d_cmin, float* d_cmed, int32_t* d_ctype, uint32_t* d_ccrc,
                                   void* stream) {
  if (dcb_device_count() == 0) return DCB_ENODEVICE;
  const uint64_t need = dcb_workspace_bytes(n, chunk);
  if (n > DCB_STATIC_MAX_ELEMS || need > DCB_STATIC_WS_BYTES) {
    snprintf(g_err, sizeof g_err, "dcb_compress_static: n = %llu elements needs %llu bytes of scratch; the static array holds %llu "
             "(at most %llu elements per call) -- use dcb_compress with caller memory", (unsigned long long)n,
             (unsigned long long)need, (unsigned long long)DCB_STATIC_WS_BYTES, (unsigned long long)DCB_STATIC_MAX_ELEMS);
    return DCB_ECAPACITY;
  }
  void* ws = nullptr;
  { int rc_ = static_ws_ptr(&ws); if (rc_) return rc_; }
  return dcb_compress(ct, d_in, n, chunk, eb, flags, fixed_type, fixed_mask17, d_stream, capacity, d_offsets, d_cmin, d_cmed,
                      d_ctype, d_ccrc, ws, DCB_STATIC_WS_BYTES, stream);
}

// Stream-ordered scratch of the single-chunk decoder: the library's own memory pool per device, which keeps what it has
// been given (release threshold = max) -- with the default pool's threshold of 0 every call paid a real allocation.
static int scratch_pool(cudaMemPool_t* out) {
  static std::atomic<cudaMemPool_t> pools[64];
  static std::mutex mu;
  int dev = 0;
  CK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) return DCB_EINVAL;
  cudaMemPool_t p = pools[dev].load(std::memory_order_acquire);
  if (!p) {
    std::lock_guard<std::mutex> lk(mu);
    p = pools[dev].load(std::memory_order_acquire);
    if (!p) {
      cudaMemPoolProps props;
      memset(&props, 0, sizeof props);
      props.allocType = cudaMemAllocationTypePinned;
      props.handleTypes = cudaMemHandleTypeNone;
      props.location.type = cudaMemLocationTypeDevice;
      props.location.id = dev;
      CK(cudaMemPoolCreate(&p, &props));
      uint64_t keep = ~0ull;
      CK(cudaMemPoolSetAttribute(p, cudaMemPoolAttrReleaseThreshold, &keep));
      pools[dev].store(p, std::memory_order_release);
    }
  }
  *out = p;
  return DCB_OK;
}

template <int CT>
static cudaError_t launch_decode(const DecArgs& a, cudaStream_t st) {
  // (the attribute is per device, and setting it is cheap: no process-wide "already set" flag)
  cudaError_t e = cudaFuncSetAttribute(dcb_decode_kernel<CT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)DEC_SMEM_BYTES);
  if (e != cudaSuccess) return e;
  dcb_decode_kernel<CT><<<(unsigned)((a.num_chunks + DEC_NT - 1) / DEC_NT), DEC_NT, DEC_SMEM_BYTES, st>>>(a);
  return cudaGetLastError();
}

extern "C" int dcb_decompress(int ct, const uint8_t* d_stream, const uint64_t* d_offsets, uint64_t n, uint64_t chunk,
                              double eb, uint32_t flags, int fixed_type, uint32_t fixed_mask17,
                              const float* d_cmin, const float* d_cmed, const int32_t* d_ctype, float* d_out,
                              int32_t* d_status, void* stream) {
  if (dcb_device_count() == 0) return DCB_ENODEVICE;
  cudaStream_t st = (cudaStream_t)stream;
  const int bct = base_ct(ct, &flags);
  DecArgs a;
  memset(&a, 0, sizeof a);
  if (bct < 0 || make_const(eb, &a.k) != DCB_OK) return DCB_EINVAL;
  if (d_status) CK(cudaMemsetAsync(d_status, 0, sizeof(int32_t), st));
  if (n == 0) return DCB_OK;
  const bool single = single_chunk(n, chunk);
  if (!single && (chunk % DCB_TILE) != 0) return DCB_EINVAL;
  if (!single && chunk > (1ull << 26)) return DCB_EINVAL;            // a lane's bit position is 32 bits wide: <= 2^31 bits per chunk
  if (!d_stream || !d_offsets || !d_out || ((uintptr_t)d_stream & 15)) return DCB_EINVAL;
  if (n >= (1ull << 32) && single) return DCB_EINVAL;
  const bool per_chunk_mask = (bct == CT_MASK) && !(flags & DCB_F_FIXED_MASK);
  if ((flags & DCB_F_SUBTRACT_MIN) && !d_cmin) return DCB_EINVAL;
  if (per_chunk_mask && (!d_cmed || !d_ctype)) return DCB_EINVAL;
  if ((flags & DCB_F_FIXED_MASK) && (fixed_type < 0 || fixed_type > 7 || fixed_mask17 >= (1u << 17))) return DCB_EINVAL;
  a.stream = d_stream; a.offsets = d_offsets; a.n = n; a.chunk = single ? n : chunk;
  a.num_chunks = dcb_num_chunks(n, chunk);
  a.flags = flags & 7u; a.fixed_type = fixed_type; a.fixed_mask17 = fixed_mask17;
  a.cmin = d_cmin; a.cmed = d_cmed; a.ctype = d_ctype; a.out = d_out; a.status = d_status;
  cudaError_t e;
  // ONE chunk of more than a tile (the reference's own call shape): codes are located and resolved in parallel inside the
  // chunk (dcb_decode_single.cuh) instead of by one lane.  Its scratch (a byte per element + tables) is stream-ordered
  // memory: allocated and freed on `st`, no host synchronisation.  DCB_DECODER=lane keeps the lane-per-chunk kernel.
  static const bool lane_only = [] { const char* v = getenv("DCB_DECODER"); return v && !strcmp(v, "lane"); }();
  if (a.num_chunks == 1 && n > DCB_TILE && !lane_only) {
    void* ws = nullptr;
    cudaMemPool_t pool = nullptr;
    { int rc_ = scratch_pool(&pool); if (rc_) return rc_; }
    CK(cudaMallocFromPoolAsync(&ws, (size_t)ds_workspace_bytes(n), pool, st));
    switch (bct) {
      case CT_BITWISE: e = ds_launch<CT_BITWISE>(a, ws, st); break;
      case CT_NP: e = ds_launch<CT_NP>(a, ws, st); break;
      case CT_MASK: e = ds_launch<CT_MASK>(a, ws, st); break;
      default: e = ds_launch<CT_OP>(a, ws, st); break;
    }
    cudaError_t ef = cudaFreeAsync(ws, st);
    CK(e);
    CK(ef);
    return DCB_OK;
  }
  switch (bct) {
    case CT_BITWISE: e = launch_decode<CT_BITWISE>(a, st); break;
    case CT_NP: e = launch_decode<CT_NP>(a, st); break;
    case CT_MASK: e = launch_decode<CT_MASK>(a, st); break;
    default: e = launch_decode<CT_OP>(a, st); break;
  }
  CK(e);
  return DCB_OK;
}

// CRC tables live in device globals and are built by one tiny kernel the first time a device needs them
static int ensure_crc_tables(cudaStream_t st) {
  static std::atomic<int> ready[64];
  int dev = 0;
  CK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) return DCB_EINVAL;
  if (ready[dev].load(std::memory_order_acquire) == 0) {
    static std::mutex mu;
    std::lock_guard<std::mutex> lk(mu);
    if (ready[dev].load(std::memory_order_acquire) == 0) {
      dcb_crc_init_kernel<<<1, 256, 0, st>>>();
      CK(cudaGetLastError());
      CK(cudaStreamSynchronize(st));               // other streams of this device may use the tables next
      ready[dev].store(1, std::memory_order_release);
    }
  }
  return DCB_OK;
}

extern "C" int dcb_crc32_chunks(const uint8_t* d_stream, const uint64_t* d_offsets, uint64_t K, uint32_t* d_crc,
                                void* stream) {
  if (dcb_device_count() == 0) return DCB_ENODEVICE;
  if (K == 0) return DCB_OK;
  if (!d_stream || !d_offsets || !d_crc) return DCB_EINVAL;
  { int rc_ = ensure_crc_tables((cudaStream_t)stream); if (rc_) return rc_; }
  dcb_crc_kernel<<<(unsigned)((K + 3) / 4), 128, 0, (cudaStream_t)stream>>>(d_stream, d_offsets, K, d_crc);
  CK(cudaGetLastError());
  return DCB_OK;
}

extern "C" int dcb_flip_bits(uint8_t* d_stream, const uint64_t* d_idx, uint64_t count, void* stream) {
  if (dcb_device_count() == 0) return DCB_ENODEVICE;
  if (count == 0) return DCB_OK;
  if (!d_stream || !d_idx || ((uintptr_t)d_stream & 3)) return DCB_EINVAL;
  dcb_flip_kernel<<<(unsigned)((count + 255) / 256), 256, 0, (cudaStream_t)stream>>>(d_stream, d_idx, count);
  CK(cudaGetLastError());
  return DCB_OK;
}

static int face_args(int MI, int MJ, int MK, int ijk, int v, int imax, int jmax, int kmax, FaceArgs* f) {
  if (MI <= 0 || MJ <= 0 || MK <= 0 || imax <= 0 || jmax <= 0 || kmax <= 0 || imax > MI || jmax > MJ || kmax > MK) return DCB_EINVAL;
  f->MJ = MJ; f->MK = MK; f->ijk = ijk; f->v = v;
  if (ijk == 1) { f->A = jmax; f->B = kmax; if (v < 0 || v >= MI) return DCB_EINVAL; }
  else if (ijk == 2) { f->A = imax; f->B = kmax; if (v < 0 || v >= MJ) return DCB_EINVAL; }
  else if (ijk == 3) { f->A = imax; f->B = jmax; if (v < 0 || v >= MK) return DCB_EINVAL; }
  else return DCB_EINVAL;                      // the reference leaves A, B uninitialised here
  return DCB_OK;
}

extern "C" int dcb_gather_face(const float* d_p, int MI, int MJ, int MK, int ijk, int v, int imax, int jmax, int kmax,
                               float* d_out, void* stream) {
  if (dcb_device_count() == 0) return DCB_ENODEVICE;
  FaceArgs f;
  if (!d_p || !d_out || face_args(MI, MJ, MK, ijk, v, imax, jmax, kmax, &f)) return DCB_EINVAL;
  const size_t n = (size_t)f.A * f.B;
  dcb_gather_face_kernel<<<(unsigned)((n + 255) / 256 < 4096 ? (n + 255) / 256 : 4096), 256, 0, (cudaStream_t)stream>>>(d_p, f, d_out);
  CK(cudaGetLastError());
  return DCB_OK;
}

extern "C" int dcb_scatter_face(float* d_p, int MI, int MJ, int MK, int ijk, int v, int imax, int jmax, int kmax,
                                const float* d_in, float add, void* stream) {
  if (dcb_device_count() == 0) return DCB_ENODEVICE;
  FaceArgs f;
  if (!d_p || !d_in || face_args(MI, MJ, MK, ijk, v, imax, jmax, kmax, &f)) return DCB_EINVAL;
  const size_t n = (size_t)f.A * f.B;
  dcb_scatter_face_kernel<<<(unsigned)((n + 255) / 256 < 4096 ? (n + 255) / 256 : 4096), 256, 0, (cudaStream_t)stream>>>(d_p, f, d_in, add);
  CK(cudaGetLastError());
  return DCB_OK;
}

// ---- float64 codecs -----------------------------------------------------------------------------
extern "C" uint64_t dcb_max_stream_bytes_f64(uint64_t n, uint64_t chunk) { (void)chunk; return 8 * n + 16; }
// float64, one chunk == one tile: the fused encoder's scratch = header | gsum[G] | sizes[K] | K slots of 32 KB
static bool f64_fused_on() {
  static const bool on = [] { const char* v = getenv("DCB_F64_FUSED"); return !(v && v[0] == '0'); }();
  return on;
}
static uint64_t ws64_slots_at(uint64_t K) { return (256 + 4 * ws_groups(K) + 4 * K + 255) & ~(uint64_t)255; }
extern "C" uint64_t dcb_workspace_bytes_f64(uint64_t n, uint64_t chunk) {
  const uint64_t plain = 256 + 8 * num_tiles(n);
  if (n && chunk_is_tile(n, chunk)) {
    const uint64_t fused = ws64_slots_at(num_tiles(n)) + num_tiles(n) * (uint64_t)F64_SLOT_PITCH;
    return fused > plain ? fused : plain;
  }
  return plain;
}

static int f64_setup(int ct, double eb, uint32_t* flags, int* bct, F64Const* k) {
  *bct = base_ct(ct, flags);
  k->eb = eb; k->n_eb = dcb_n_eb(eb);
  if (*bct < 0 || !(eb > 0.0) || k->n_eb < 0) return DCB_EINVAL;
  return DCB_OK;
}

// per-chunk statistics of the float64 codecs: one block per chunk, except ONE long chunk (the single-call shape), which is
// cut into parts (dcb_f64_stats_long.cuh; scratch from the library's stream-ordered pool)
static int f64_stats_run(const F64StatsArgs& s, uint64_t K, cudaStream_t st) {
  static const bool one_block = [] { const char* v = getenv("DCB_F64_STATS"); return v && !strcmp(v, "block"); }();
  if (K == 1 && s.n >= 65536 && !one_block) {
    void* ws = nullptr;
    cudaMemPool_t pool = nullptr;
    { int rc_ = scratch_pool(&pool); if (rc_) return rc_; }
    CK(cudaMallocFromPoolAsync(&ws, (size_t)f64l_scratch_bytes(), pool, st));
    cudaError_t e = f64l_launch(s, ws, st);
    cudaError_t ef = cudaFreeAsync(ws, st);
    CK(e);
    CK(ef);
    return DCB_OK;
  }
  dcb_f64_stats_kernel<<<(unsigned)K, F64_NT, 0, st>>>(s);
  CK(cudaGetLastError());
  return DCB_OK;
}

extern "C" int dcb_compress_f64(int ct, const double* d_in, uint64_t n, uint64_t chunk, double eb, uint32_t flags,
                                int fixed_type, uint32_t fixed_mask20, uint8_t* d_stream, uint64_t capacity,
                                uint64_t* d_offsets, double* d_cmin, uint32_t* d_cmask, int32_t* d_ctype, double* d_cmedium,
                                uint32_t* d_ccrc, void* d_ws, uint64_t ws_bytes, void* stream) {
  if (dcb_device_count() == 0) return DCB_ENODEVICE;
  cudaStream_t st = (cudaStream_t)stream;
  F64Args a;
  memset(&a, 0, sizeof a);
  int bct;
  if (f64_setup(ct, eb, &flags, &bct, &a.k) != DCB_OK || !d_offsets) return DCB_EINVAL;
  if (n == 0) { CK(cudaMemsetAsync(d_offsets, 0, sizeof(uint64_t), st)); return DCB_OK; }
  const bool single = single_chunk(n, chunk);
  if (!single && (chunk % DCB_TILE) != 0) return DCB_EINVAL;
  if (!d_in || !d_stream || ((uintptr_t)d_stream & 15) || ((uintptr_t)d_in & 7)) return DCB_EINVAL;
  if (n >= (1ull << 40) || (!single && chunk >= (1ull << 32))) return DCB_EINVAL;
  const uint64_t K = dcb_num_chunks(n, chunk);
  const bool per_chunk_mask = (bct == CT_MASK) && !(flags & DCB_F_FIXED_MASK);
  if ((flags & DCB_F_SUBTRACT_MIN) && !d_cmin) return DCB_EINVAL;
  if (per_chunk_mask && (!d_cmask || !d_ctype)) return DCB_EINVAL;
  if ((flags & DCB_F_EXACT_MEDIUM) && per_chunk_mask && !d_cmedium) return DCB_EINVAL;
  if ((flags & DCB_F_FIXED_MASK) && (fixed_type < 0 || fixed_type > 10 || fixed_mask20 >= (1u << 20))) return DCB_EINVAL;
  if ((flags & DCB_F_CRC) && !d_ccrc) return DCB_EINVAL;
  if (capacity < dcb_max_stream_bytes_f64(n, chunk)) return DCB_ECAPACITY;
  if (!d_ws || ws_bytes < dcb_workspace_bytes_f64(n, chunk)) return DCB_ECAPACITY;

  a.in = d_in; a.n = n; a.chunk = single ? n : chunk;
  a.num_tiles = (uint32_t)(single ? num_tiles(n) : 0);
  a.tiles_per_chunk = single ? (uint32_t)num_tiles(n) : (uint32_t)(chunk / DCB_TILE);
  if (!single) a.num_tiles = (uint32_t)((K - 1) * a.tiles_per_chunk + num_tiles(n - (K - 1) * chunk));
  a.num_chunks = K;
  a.flags = flags & 15u; a.fixed_type = fixed_type; a.fixed_mask20 = fixed_mask20;
  a.out = d_stream; a.offsets = d_offsets; a.cmin = d_cmin; a.cmask = per_chunk_mask ? d_cmask : nullptr; a.ctype = d_ctype;
  a.tile_bits = (unsigned long long*)((char*)d_ws + 256);
  a.last_bits = (unsigned long long*)d_ws;

  if (chunk_is_tile(n, chunk) && f64_fused_on()) {
    // one block per chunk keeps the chunk in shared memory from the first load to the packed image (dcb_f64_fused_kernel);
    // the chunk table, the final placement and the CRC are the float32 path's compaction kernel
    char* ws = (char*)d_ws;
    F64FusedArgs f;
    f.a = a;
    f.a.chunk = single ? (uint64_t)F64_TILE : chunk;               // blocks index chunks of one tile
    f.gsum = (uint32_t*)(ws + 256);
    f.sizes = f.gsum + ws_groups(K);
    f.slots = (uint8_t*)(ws + ws64_slots_at(K));
    f.last_tile_bits = (uint32_t*)(ws + 16);
    f.want_mask = per_chunk_mask ? 1 : 0;
    f.cmedium = (per_chunk_mask && (flags & DCB_F_EXACT_MEDIUM)) ? d_cmedium : nullptr;
    CK(cudaMemsetAsync(d_ws, 0, (size_t)(256 + 4 * ws_groups(K)), st));
    switch (bct) {
      case CT_BITWISE: dcb_f64_fused_kernel<CT_BITWISE><<<(unsigned)K, F64_NT, 0, st>>>(f); break;
      case CT_NP: dcb_f64_fused_kernel<CT_NP><<<(unsigned)K, F64_NT, 0, st>>>(f); break;
      case CT_MASK: dcb_f64_fused_kernel<CT_MASK><<<(unsigned)K, F64_NT, 0, st>>>(f); break;
      default: dcb_f64_fused_kernel<CT_OP><<<(unsigned)K, F64_NT, 0, st>>>(f); break;
    }
    CK(cudaGetLastError());
    CompactArgs ca;
    ca.slots = f.slots; ca.sizes = f.sizes; ca.gsum = f.gsum; ca.last_tile_bits = f.last_tile_bits; ca.K = K;
    ca.out = d_stream; ca.offsets = d_offsets; ca.last_bits = a.last_bits; ca.pitch = F64_SLOT_PITCH;
    ca.crc = (flags & DCB_F_CRC) ? d_ccrc : nullptr;
    if (ca.crc) { int rc_ = ensure_crc_tables(st); if (rc_) return rc_; }
    dcb_compact_kernel<<<(unsigned)ws_groups(K), DCB_COMPACT_NT, 0, st>>>(ca);
    CK(cudaGetLastError());
    return DCB_OK;
  }
  if ((flags & DCB_F_SUBTRACT_MIN) || per_chunk_mask) {
    F64StatsArgs s;
    s.in = d_in; s.n = n; s.chunk = a.chunk; s.subtract_min = (flags & DCB_F_SUBTRACT_MIN) ? 1 : 0;
    s.want_mask = per_chunk_mask ? 1 : 0;
    s.cmin = d_cmin; s.cmask = d_cmask; s.ctype = d_ctype;
    s.cmedium = (per_chunk_mask && (flags & DCB_F_EXACT_MEDIUM)) ? d_cmedium : nullptr;
    { int rc_ = f64_stats_run(s, K, st); if (rc_) return rc_; }
  }
  switch (bct) {
    case CT_BITWISE: dcb_f64_length_kernel<CT_BITWISE><<<a.num_tiles, F64_NT, 0, st>>>(a); break;
    case CT_NP: dcb_f64_length_kernel<CT_NP><<<a.num_tiles, F64_NT, 0, st>>>(a); break;
    case CT_MASK: dcb_f64_length_kernel<CT_MASK><<<a.num_tiles, F64_NT, 0, st>>>(a); break;
    default: dcb_f64_length_kernel<CT_OP><<<a.num_tiles, F64_NT, 0, st>>>(a); break;
  }
  CK(cudaGetLastError());
  dcb_f64_scan_kernel<<<1, 1024, 0, st>>>(a);
  CK(cudaGetLastError());
  dcb_f64_zero_kernel<<<(unsigned)num_sms() * 8, 256, 0, st>>>((uint32_t*)d_stream, d_offsets + K);
  CK(cudaGetLastError());
  switch (bct) {
    case CT_BITWISE: dcb_f64_pack_kernel<CT_BITWISE><<<a.num_tiles, F64_NT, 0, st>>>(a); break;
    case CT_NP: dcb_f64_pack_kernel<CT_NP><<<a.num_tiles, F64_NT, 0, st>>>(a); break;
    case CT_MASK: dcb_f64_pack_kernel<CT_MASK><<<a.num_tiles, F64_NT, 0, st>>>(a); break;
    default: dcb_f64_pack_kernel<CT_OP><<<a.num_tiles, F64_NT, 0, st>>>(a); break;
  }
  CK(cudaGetLastError());
  if (flags & DCB_F_CRC) {
    { int rc_ = ensure_crc_tables(st); if (rc_) return rc_; }
    dcb_crc_kernel<<<(unsigned)((K + 3) / 4), 128, 0, st>>>(d_stream, d_offsets, K, d_ccrc);
    CK(cudaGetLastError());
  }
  return DCB_OK;
}

// toSmallDataset_double (:3522) / med_dataset_double (:3564) per chunk as a stand-alone pass.  want_medium: also mask20, type
// and (d_chunk_medium != NULL) the exact sequential mean; subtract_min = 0 takes the values as they are.
extern "C" int dcb_chunk_stats_f64(const double* d_in, uint64_t n, uint64_t chunk, int subtract_min, int want_medium,
                                   double* d_cmin, uint32_t* d_cmask, int32_t* d_ctype, double* d_cmedium, void* stream) {
  if (dcb_device_count() == 0) return DCB_ENODEVICE;
  if (!d_in || n == 0 || (subtract_min && !d_cmin) || (want_medium && (!d_cmask || !d_ctype))) return DCB_EINVAL;
  const bool single = single_chunk(n, chunk);
  F64StatsArgs s;
  s.in = d_in; s.n = n; s.chunk = single ? n : chunk; s.subtract_min = subtract_min ? 1 : 0; s.want_mask = want_medium ? 1 : 0;
  s.cmin = d_cmin; s.cmask = d_cmask; s.ctype = d_ctype; s.cmedium = d_cmedium;
  return f64_stats_run(s, dcb_num_chunks(n, chunk), (cudaStream_t)stream);
}

extern "C" int dcb_decompress_f64(int ct, const uint8_t* d_stream, const uint64_t* d_offsets, uint64_t n, uint64_t chunk,
                                  double eb, uint32_t flags, int fixed_type, uint32_t fixed_mask20, const double* d_cmin,
                                  const uint32_t* d_cmask, const int32_t* d_ctype, double* d_out, int32_t* d_status,
                                  void* stream) {
  if (dcb_device_count() == 0) return DCB_ENODEVICE;
  cudaStream_t st = (cudaStream_t)stream;
  F64DecArgs a;
  memset(&a, 0, sizeof a);
  int bct;
  if (f64_setup(ct, eb, &flags, &bct, &a.k) != DCB_OK) return DCB_EINVAL;
  if (d_status) CK(cudaMemsetAsync(d_status, 0, sizeof(int32_t), st));
  if (n == 0) return DCB_OK;
  const bool single = single_chunk(n, chunk);
  if (!d_stream || !d_offsets || !d_out || ((uintptr_t)d_stream & 3)) return DCB_EINVAL;
  const bool per_chunk_mask = (bct == CT_MASK) && !(flags & DCB_F_FIXED_MASK);
  if ((flags & DCB_F_SUBTRACT_MIN) && !d_cmin) return DCB_EINVAL;
  if (per_chunk_mask && (!d_cmask || !d_ctype)) return DCB_EINVAL;
  if ((flags & DCB_F_FIXED_MASK) && (fixed_type < 0 || fixed_type > 10 || fixed_mask20 >= (1u << 20))) return DCB_EINVAL;
  a.stream = d_stream; a.offsets = d_offsets; a.n = n; a.chunk = single ? n : chunk; a.num_chunks = dcb_num_chunks(n, chunk);
  a.flags = flags & 15u; a.fixed_type = fixed_type; a.fixed_mask20 = fixed_mask20;
  a.cmin = d_cmin; a.cmask = d_cmask; a.ctype = d_ctype; a.out = d_out; a.status = d_status;
  // ONE chunk of more than a tile (the *_double call shape of k-means / MM / LU): codes located and resolved in parallel
  // inside the chunk (dcb_decode_single_f64.cuh); scratch from the library's stream-ordered pool.  DCB_DECODER=lane: one thread.
  static const bool lane_only = [] { const char* v = getenv("DCB_DECODER"); return v && !strcmp(v, "lane"); }();
  if (a.num_chunks == 1 && n > F64_TILE && n < (1ull << 32) && !lane_only) {
    void* ws = nullptr;
    cudaMemPool_t pool = nullptr;
    { int rc_ = scratch_pool(&pool); if (rc_) return rc_; }
    CK(cudaMallocFromPoolAsync(&ws, (size_t)ds64_workspace_bytes(n), pool, st));
    cudaError_t e;
    switch (bct) {
      case CT_BITWISE: e = ds64_launch<CT_BITWISE>(a, ws, st); break;
      case CT_NP: e = ds64_launch<CT_NP>(a, ws, st); break;
      case CT_MASK: e = ds64_launch<CT_MASK>(a, ws, st); break;
      default: e = ds64_launch<CT_OP>(a, ws, st); break;
    }
    cudaError_t ef = cudaFreeAsync(ws, st);
    CK(e);
    CK(ef);
    return DCB_OK;
  }
  // chunked streams: the ring decoder (dcb_f64_decode_ring.cuh; cp.async needs a 16-byte aligned stream).  DCB_F64_DECODER=window
  // keeps the register-window kernel.
  static const bool window_only = [] { const char* v = getenv("DCB_F64_DECODER"); return v && !strcmp(v, "window"); }();
  if (!window_only && ((uintptr_t)d_stream & 15) == 0 && a.chunk <= (1ull << 25)) {   // (the ring reader's bit position is 32 bits wide)
    const unsigned per_block = (F64R_NT / 32) * F64R_LANES;
    const unsigned rgrid = (unsigned)((a.num_chunks + per_block - 1) / per_block);
    switch (bct) {
      case CT_BITWISE: dcb_f64_decode_ring_kernel<CT_BITWISE><<<rgrid, F64R_NT, 0, st>>>(a); break;
      case CT_NP: dcb_f64_decode_ring_kernel<CT_NP><<<rgrid, F64R_NT, 0, st>>>(a); break;
      case CT_MASK: dcb_f64_decode_ring_kernel<CT_MASK><<<rgrid, F64R_NT, 0, st>>>(a); break;
      default: dcb_f64_decode_ring_kernel<CT_OP><<<rgrid, F64R_NT, 0, st>>>(a); break;
    }
    CK(cudaGetLastError());
    return DCB_OK;
  }
  const unsigned grid = (unsigned)((a.num_chunks + 4 * F64_DEC_LANES - 1) / (4 * F64_DEC_LANES));   // 4 warps per block
  switch (bct) {
    case CT_BITWISE: dcb_f64_decode_kernel<CT_BITWISE><<<grid, 128, 0, st>>>(a); break;
    case CT_NP: dcb_f64_decode_kernel<CT_NP><<<grid, 128, 0, st>>>(a); break;
    case CT_MASK: dcb_f64_decode_kernel<CT_MASK><<<grid, 128, 0, st>>>(a); break;
    default: dcb_f64_decode_kernel<CT_OP><<<grid, 128, 0, st>>>(a); break;
  }
  CK(cudaGetLastError());
  return DCB_OK;
}

extern "C" int dcb_himeno_jacobi(const float* d_p, const float* d_a, const float* d_b, const float* d_c, const float* d_bnd,
                                 const float* d_wrk1, float* d_wrk2, int MI, int MJ, int MK, int imax, int jmax, int kmax,
                                 float omega, double* d_gosa, void* stream) {
  if (dcb_device_count() == 0) return DCB_ENODEVICE;
  if (!d_p || !d_a || !d_b || !d_c || !d_bnd || !d_wrk1 || !d_wrk2 || !d_gosa) return DCB_EINVAL;
  if (imax < 3 || jmax < 3 || kmax < 3 || imax > MI || jmax > MJ || kmax > MK) return DCB_EINVAL;
  cudaStream_t st = (cudaStream_t)stream;
  JacobiArgs q;
  q.p = d_p; q.a = d_a; q.b = d_b; q.c = d_c; q.bnd = d_bnd; q.wrk1 = d_wrk1; q.wrk2 = d_wrk2;
  q.MI = MI; q.MJ = MJ; q.MK = MK; q.imax = imax; q.jmax = jmax; q.kmax = kmax; q.omega = omega; q.gosa = d_gosa;
  CK(cudaMemsetAsync(d_gosa, 0, sizeof(double), st));
  const size_t total = (size_t)(imax - 2) * (jmax - 2) * (kmax - 2);
  const unsigned grid = (unsigned)((total + 255) / 256 < (size_t)num_sms() * 16 ? (total + 255) / 256 : (size_t)num_sms() * 16);
  dcb_himeno_jacobi_kernel<<<grid, 256, 0, st>>>(q);
  CK(cudaGetLastError());
  return DCB_OK;
}

extern "C" int dcb_himeno_copy_interior(float* d_p, const float* d_wrk2, int MI, int MJ, int MK, int imax, int jmax, int kmax,
                                        void* stream) {
  if (dcb_device_count() == 0) return DCB_ENODEVICE;
  if (!d_p || !d_wrk2 || imax < 3 || jmax < 3 || kmax < 3 || imax > MI || jmax > MJ || kmax > MK) return DCB_EINVAL;
  const size_t total = (size_t)(imax - 2) * (jmax - 2) * (kmax - 2);
  const unsigned grid = (unsigned)((total + 255) / 256 < (size_t)num_sms() * 16 ? (total + 255) / 256 : (size_t)num_sms() * 16);
  dcb_himeno_copy_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(d_p, d_wrk2, MJ, MK, imax, jmax, kmax);
  CK(cudaGetLastError());
  return DCB_OK;
}

// ---- host-pointer layer: cached device buffers (grow only), one stream ---------------------------------------
struct HostCtx {
  cudaStream_t st = nullptr;
  void* buf[8] = {nullptr};
  size_t cap[8] = {0};
  char* pin = nullptr;            // pinned staging for the chunk tables: the caller's table pointers may be pageable, and the
  size_t pin_cap = 0;             // driver's pageable-copy path serialises against the bulk copies of other host threads
  char* stage[16] = {nullptr};    // pinned staging + one stream per helper of the pageable bulk path (copy_staged)
  cudaStream_t sst[16] = {nullptr};
  int dev = -1;                   // device the buffers live on
  void release() {                // everything back to the driver (a thread that exits must not leak 12-16 n bytes of HBM)
    if (dev >= 0) {
      int cur = -1;
      const bool sw = cudaGetDevice(&cur) == cudaSuccess && cur != dev && cudaSetDevice(dev) == cudaSuccess;
      for (int i = 0; i < 8; i++) { if (buf[i]) cudaFree(buf[i]); buf[i] = nullptr; cap[i] = 0; }
      if (pin) cudaFreeHost(pin);
      pin = nullptr; pin_cap = 0;
      for (int i = 0; i < 16; i++) {
        if (stage[i]) cudaFreeHost(stage[i]);
        if (sst[i]) cudaStreamDestroy(sst[i]);
        stage[i] = nullptr; sst[i] = nullptr;
      }
      if (st) cudaStreamDestroy(st);
      st = nullptr;
      if (sw) cudaSetDevice(cur);
      dev = -1;
      cudaGetLastError();         // at process exit the context may already be gone: nothing to report
    }
  }
  ~HostCtx() { release(); }
};
static thread_local HostCtx g_h;
// frees the calling thread's cached device buffers, pinned staging and stream (they are also freed when the thread exits)
extern "C" int dcb_host_release(void) { g_h.release(); return DCB_OK; }
static int host_ctx_stream(cudaStream_t* out) {
  int dev = 0;
  CK(cudaGetDevice(&dev));
  if (g_h.dev != dev) {           // first use, or the thread moved to another device: the cached buffers belong to the old one
    g_h.release();
    g_h.dev = dev;
  }
  if (!g_h.st) CK(cudaStreamCreateWithFlags(&g_h.st, cudaStreamNonBlocking));
  *out = g_h.st;
  return DCB_OK;
}
static thread_local uint64_t g_last_bits = 0;
extern "C" uint64_t dcb_last_stream_bits(void) { return g_last_bits; }
static int hbuf(int i, size_t bytes, void** p) {
  bytes = (bytes + 255) & ~(size_t)255;
  if (bytes == 0) bytes = 256;
  if (g_h.cap[i] < bytes) {
    if (g_h.buf[i]) cudaFree(g_h.buf[i]);
    g_h.buf[i] = nullptr; g_h.cap[i] = 0;
    CK(cudaMalloc(&g_h.buf[i], bytes));
    g_h.cap[i] = bytes;
  }
  *p = g_h.buf[i];
  return DCB_OK;
}
// Host <-> device transfers of the host-pointer layer.  A copy engine serves the streams that feed it one after the other
// (measured: a 100 MB copy queued behind another host thread's 1 GB copy in the same direction waits for all of it, even
// when that copy is issued in small slices), which leaves the opposite direction idle whenever two host threads overlap
// a compress with a decompress.  So every transfer goes out in slices, and a slice holds a FAIR per-direction lock
// until it has completed: concurrent callers alternate slice by slice, and both engines stay busy.
// Slice size (DCB_COPY_SLICE_MB, default 128): measured with 4 caller threads on one B200, 2^28 floats per call -- 32 MB 0.84-0.90
// of the box's concurrent-copy ceiling, 128 MB 0.96-0.97 (43.5 GB/s of input), 256 MB 0.95-0.97, 1024 MB (= no slicing) 0.85.
struct TicketLock {               // FIFO; waiters sleep on a condition variable (r1 spun on yield(): 16 spinners on 32 cores)
  std::mutex m;
  std::condition_variable cv;
  uint64_t next = 0, serving = 0;
  void lock() { std::unique_lock<std::mutex> lk(m); const uint64_t t = next++; cv.wait(lk, [&] { return serving == t; }); }
  void unlock() { { std::lock_guard<std::mutex> lk(m); serving++; } cv.notify_all(); }
};
static TicketLock g_h2d_mu, g_d2h_mu;
static const bool g_trace = getenv("DCB_TRACE") != nullptr;
static double now_ms() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
#define TRACE(what) do { if (g_trace) fprintf(stderr, "[dcb %p] %-28s %.2f\n", (void*)&g_h, what, now_ms()); } while (0)
// Bulk copies from / to PAGEABLE host memory (what the reference's drivers pass: malloc'd arrays).  cudaMemcpyAsync on pageable
// memory goes through the driver's own small staging buffer on one thread: 6.6 GB/s of input for the bench's round trip.  Here
// DCB_STAGED_THREADS helpers (default 6: the caller + five threads) take 4 MB slices off a shared counter, each through its OWN
// pinned staging buffer and stream: memcpy into staging + DMA for host-to-device, DMA + memcpy out for the other direction; while
// one helper copies in host memory another one's DMA runs.  Measured, compress + decompress of 2^26 floats from numpy arrays, one
// caller: driver path 6.3 GB/s; 2 / 4 / 6 / 8 / 16 helpers 8.0 / 14.3 / 17.0 / 16.8 / 17.4 GB/s (pinned buffers: 24.9).
// Only for copies of DCB_STAGED_MIN_MB megabytes or more (default 8: thread start-up is ~0.1 ms).
#define DCB_STAGE_SLICE ((size_t)4 << 20)
static bool host_ptr_is_pageable(const void* p) {
  cudaPointerAttributes at;
  if (cudaPointerGetAttributes(&at, p) != cudaSuccess) { cudaGetLastError(); return false; }
  return at.type == cudaMemoryTypeUnregistered;
}
static cudaError_t copy_staged(void* dst, const void* src, size_t bytes, cudaMemcpyKind kind, cudaStream_t st) {
  cudaError_t e = cudaStreamSynchronize(st);           // whatever produces the source has finished
  if (e != cudaSuccess) return e;
  static const int T = [] { const char* v = getenv("DCB_STAGED_THREADS"); int t = v ? atoi(v) : 6; return t < 1 ? 1 : (t > 16 ? 16 : t); }();
  for (int i = 0; i < T; i++) {
    if (!g_h.stage[i] && (e = cudaHostAlloc((void**)&g_h.stage[i], DCB_STAGE_SLICE, cudaHostAllocDefault)) != cudaSuccess) return e;
    if (!g_h.sst[i] && (e = cudaStreamCreateWithFlags(&g_h.sst[i], cudaStreamNonBlocking)) != cudaSuccess) return e;
  }
  const int dev = g_h.dev;
  char* const* stage = g_h.stage;
  cudaStream_t const* sst = g_h.sst;
  std::atomic<size_t> next{0};
  std::atomic<int> err{(int)cudaSuccess};
  auto work = [&](int t, bool set_device) {
    if (set_device && cudaSetDevice(dev) != cudaSuccess) { err.store((int)cudaErrorInvalidDevice); return; }
    for (;;) {
      const size_t o = next.fetch_add(DCB_STAGE_SLICE);
      if (o >= bytes || err.load() != (int)cudaSuccess) break;
      const size_t len = bytes - o < DCB_STAGE_SLICE ? bytes - o : DCB_STAGE_SLICE;
      cudaError_t r;
      if (kind == cudaMemcpyHostToDevice) {
        memcpy(stage[t], (const char*)src + o, len);
        r = cudaMemcpyAsync((char*)dst + o, stage[t], len, kind, sst[t]);
        if (r == cudaSuccess) r = cudaStreamSynchronize(sst[t]);
      } else {
        r = cudaMemcpyAsync(stage[t], (const char*)src + o, len, kind, sst[t]);
        if (r == cudaSuccess) r = cudaStreamSynchronize(sst[t]);
        if (r == cudaSuccess) memcpy((char*)dst + o, stage[t], len);
      }
      if (r != cudaSuccess) { err.store((int)r); break; }
    }
  };
  std::thread helpers[15];
  for (int t = 1; t < T; t++) helpers[t - 1] = std::thread(work, t, true);
  work(0, false);
  for (int t = 1; t < T; t++) helpers[t - 1].join();
  return (cudaError_t)err.load();
}
static cudaError_t copy_sliced(TicketLock& mu, void* dst, const void* src, size_t bytes, cudaMemcpyKind kind, cudaStream_t st) {
  static const size_t STAGED_MIN = [] { const char* e = getenv("DCB_STAGED_MIN_MB"); long v = e ? atol(e) : 8; return v > 0 ? (size_t)v << 20 : ~(size_t)0 >> 1; }();
  if (bytes >= STAGED_MIN && host_ptr_is_pageable(kind == cudaMemcpyHostToDevice ? src : dst)) return copy_staged(dst, src, bytes, kind, st);
  static const size_t SL = [] { const char* e = getenv("DCB_COPY_SLICE_MB"); long v = e ? atol(e) : 128; return v > 0 ? (size_t)v << 20 : ~(size_t)0 >> 1; }();
  cudaError_t e = cudaStreamSynchronize(st);           // whatever produces the source runs outside the lock
  for (size_t o = 0; o < bytes && e == cudaSuccess; o += SL) {
    mu.lock();
    e = cudaMemcpyAsync((char*)dst + o, (const char*)src + o, bytes - o < SL ? bytes - o : SL, kind, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    mu.unlock();
  }
  return e;
}
static int pinbuf(size_t bytes, char** p) {
  if (g_h.pin_cap < bytes) {
    if (g_h.pin) cudaFreeHost(g_h.pin);
    g_h.pin = nullptr; g_h.pin_cap = 0;
    CK(cudaHostAlloc((void**)&g_h.pin, bytes, cudaHostAllocDefault));
    g_h.pin_cap = bytes;
  }
  *p = g_h.pin;
  return DCB_OK;
}
#define HB(i, bytes, ptr) do { int rc_ = hbuf(i, bytes, (void**)&(ptr)); if (rc_) return rc_; } while (0)

extern "C" int dcb_compress_host(int ct, const float* h_in, uint64_t n, uint64_t chunk, double eb, uint32_t flags,
                                 int fixed_type, uint32_t fixed_mask17, uint8_t* h_stream, uint64_t* h_offsets,
                                 float* h_cmin, float* h_cmed, int32_t* h_ctype, uint32_t* h_ccrc) {
  if (dcb_device_count() == 0) return DCB_ENODEVICE;
  if (!h_offsets) return DCB_EINVAL;
  if (n == 0) { h_offsets[0] = 0; return DCB_OK; }
  if (!h_in || !h_stream) return DCB_EINVAL;
  cudaStream_t st = nullptr;
  { int rc_ = host_ctx_stream(&st); if (rc_) return rc_; }
  const uint64_t K = dcb_num_chunks(n, chunk);
  float* d_in; uint8_t* d_s; uint64_t* d_off; float* d_min; float* d_med; int32_t* d_ty; uint32_t* d_crc; void* d_ws;
  HB(0, n * 4, d_in); HB(1, dcb_max_stream_bytes(n, chunk), d_s); HB(2, (K + 1) * 8, d_off);
  HB(3, K * 4, d_min); HB(4, K * 4, d_med); HB(5, K * 4, d_ty); HB(6, K * 4, d_crc);
  HB(7, dcb_workspace_bytes(n, chunk), d_ws);
  TRACE("C: begin");
  CK(copy_sliced(g_h2d_mu, d_in, h_in, n * 4, cudaMemcpyHostToDevice, st));
  TRACE("C: input in");
  uint32_t f2 = flags;
  const int bct = base_ct(ct, &f2);
  int rc = dcb_compress(ct, d_in, n, chunk, eb, flags, fixed_type, fixed_mask17, d_s, dcb_max_stream_bytes(n, chunk),
                        d_off, d_min, d_med, d_ty, d_crc, d_ws, dcb_workspace_bytes(n, chunk), st);
  if (rc) return rc;
  char* pin; { int rc_ = pinbuf((K + 1) * 8 + 4 * K * 4 + 64, &pin); if (rc_) return rc_; }
  uint64_t* p_off = (uint64_t*)pin;
  char* p_tab = pin + (K + 1) * 8;                          // min | medium | type | crc | last_bits
  const bool want_mask = (bct == CT_MASK && !(flags & DCB_F_FIXED_MASK));
  const bool want_crc = (f2 & DCB_F_CRC) && h_ccrc;
  CK(cudaMemcpyAsync(p_off, d_off, (K + 1) * 8, cudaMemcpyDeviceToHost, st));
  if (h_cmin) CK(cudaMemcpyAsync(p_tab, d_min, K * 4, cudaMemcpyDeviceToHost, st));
  if (want_mask && h_cmed) CK(cudaMemcpyAsync(p_tab + K * 4, d_med, K * 4, cudaMemcpyDeviceToHost, st));
  if (want_mask && h_ctype) CK(cudaMemcpyAsync(p_tab + 2 * K * 4, d_ty, K * 4, cudaMemcpyDeviceToHost, st));
  if (want_crc) CK(cudaMemcpyAsync(p_tab + 3 * K * 4, d_crc, K * 4, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(p_tab + 4 * K * 4, (char*)d_ws + 8, 8, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));                       // the stream length is only known now
  TRACE("C: kernel+tables done");
  memcpy(h_offsets, p_off, (K + 1) * 8);
  if (h_cmin) memcpy(h_cmin, p_tab, K * 4);
  if (want_mask && h_cmed) memcpy(h_cmed, p_tab + K * 4, K * 4);
  if (want_mask && h_ctype) memcpy(h_ctype, p_tab + 2 * K * 4, K * 4);
  if (want_crc) memcpy(h_ccrc, p_tab + 3 * K * 4, K * 4);
  memcpy(&g_last_bits, p_tab + 4 * K * 4, 8);
  const uint64_t total = h_offsets[K];
  if (total) CK(copy_sliced(g_d2h_mu, h_stream, d_s, total, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  TRACE("C: stream out done");
  return DCB_OK;
}

extern "C" int dcb_decompress_host(int ct, const uint8_t* h_stream, const uint64_t* h_offsets, uint64_t n,
                                   uint64_t chunk, double eb, uint32_t flags, int fixed_type, uint32_t fixed_mask17,
                                   const float* h_cmin, const float* h_cmed, const int32_t* h_ctype, float* h_out) {
  if (dcb_device_count() == 0) return DCB_ENODEVICE;
  if (n == 0) return DCB_OK;
  if (!h_stream || !h_offsets || !h_out) return DCB_EINVAL;
  cudaStream_t st = nullptr;
  { int rc_ = host_ctx_stream(&st); if (rc_) return rc_; }
  const uint64_t K = dcb_num_chunks(n, chunk);
  const uint64_t total = h_offsets[K];
  if (total > 4 * n + 16) return DCB_EINVAL;
  float* d_out; uint8_t* d_s; uint64_t* d_off; float* d_min; float* d_med; int32_t* d_ty; int32_t* d_status; 
  HB(0, n * 4, d_out); HB(1, dcb_max_stream_bytes(n, chunk), d_s); HB(2, (K + 1) * 8, d_off);
  HB(3, K * 4, d_min); HB(4, K * 4, d_med); HB(5, K * 4, d_ty); HB(6, 256, d_status);
  TRACE("D: begin");
  CK(copy_sliced(g_h2d_mu, d_s, h_stream, total, cudaMemcpyHostToDevice, st));
  char* pin; { int rc_ = pinbuf((K + 1) * 8 + 4 * K * 4 + 64, &pin); if (rc_) return rc_; }
  char* p_tab = pin + (K + 1) * 8;
  memcpy(pin, h_offsets, (K + 1) * 8);
  CK(cudaMemcpyAsync(d_off, pin, (K + 1) * 8, cudaMemcpyHostToDevice, st));
  if (h_cmin) { memcpy(p_tab, h_cmin, K * 4); CK(cudaMemcpyAsync(d_min, p_tab, K * 4, cudaMemcpyHostToDevice, st)); }
  if (h_cmed) { memcpy(p_tab + K * 4, h_cmed, K * 4); CK(cudaMemcpyAsync(d_med, p_tab + K * 4, K * 4, cudaMemcpyHostToDevice, st)); }
  if (h_ctype) { memcpy(p_tab + 2 * K * 4, h_ctype, K * 4); CK(cudaMemcpyAsync(d_ty, p_tab + 2 * K * 4, K * 4, cudaMemcpyHostToDevice, st)); }
  int rc = dcb_decompress(ct, d_s, d_off, n, chunk, eb, flags, fixed_type, fixed_mask17,
                          h_cmin ? d_min : nullptr, h_cmed ? d_med : nullptr, h_ctype ? d_ty : nullptr,
                          d_out, d_status, st);
  if (rc) return rc;
  if (g_trace) { cudaStreamSynchronize(st); TRACE("D: stream in + kernel done"); }
  int32_t* p_status = (int32_t*)(p_tab + 4 * K * 4);
  CK(cudaMemcpyAsync(p_status, d_status, 4, cudaMemcpyDeviceToHost, st));
  CK(copy_sliced(g_d2h_mu, h_out, d_out, n * 4, cudaMemcpyDeviceToHost, st));
  TRACE("D: output out");
  return *p_status ? DCB_EFORMAT : DCB_OK;
}

#ifdef DCB_DEBUG_COUNTERS
extern "C" int dcb_debug_timeline(long long* host_dst) {
  CK(cudaDeviceSynchronize());
  CK(cudaMemcpyFromSymbol(host_dst, g_dbg_ts, sizeof(g_dbg_ts)));
  return DCB_OK;
}
#endif

// ---- byte-wise codec CT=1 (myCompress / myDecompress) ---------------------------------------------------------------------
static uint64_t bw_tiles(uint64_t n) { return (n + DCB_BW_NT - 1) / DCB_BW_NT; }
static uint64_t bw_flags_bytes(uint64_t n) { return (n + 255) & ~(uint64_t)255; }
extern "C" uint64_t dcb_bytewise_workspace_bytes(uint64_t n, uint64_t chunk) {
  return bw_flags_bytes(n) + 4 * bw_tiles(n) + 4 * dcb_num_chunks(n, chunk) + 256;
}
static bool bw_three_pass() {                            // DCB_BW_FUSED=0: the three-pass encoder for every shape (A/B, tests)
  static const bool v = [] { const char* e = getenv("DCB_BW_FUSED"); return e && e[0] == '0'; }();
  return v;
}
template <typename T>
static int bytewise_compress_t(const T* d_in, uint64_t n, uint64_t chunk, double eb, T* d_val, uint8_t* d_char, int32_t* d_disp,
                               uint32_t* d_chunk_nval, void* d_ws, uint64_t ws_bytes, void* stream) {
  if (dcb_device_count() == 0) return DCB_ENODEVICE;
  cudaStream_t st = (cudaStream_t)stream;
  CodecConst k;
  if (make_const(eb, &k) != DCB_OK) return DCB_EINVAL;
  if (n == 0) return DCB_OK;
  const bool single = single_chunk(n, chunk);
  if (!single && (chunk % DCB_TILE) != 0) return DCB_EINVAL;
  if (!d_in || !d_val || !d_char || !d_disp || !d_chunk_nval || n >= (1ull << 40) || (single && n >= (1ull << 31))) return DCB_EINVAL;
  if (!d_ws || ws_bytes < dcb_bytewise_workspace_bytes(n, chunk)) return DCB_ECAPACITY;
  const uint64_t K = dcb_num_chunks(n, chunk);
  BwArgs<T> a;
  a.in = d_in; a.n = n; a.chunk = single ? n : chunk;
  a.tp = sizeof(T) == 4 ? (T)k.tp : (T)eb;              // float32: diff_min (float) <= absErrorBound (double), folded into a float threshold
  a.flags = (uint8_t*)d_ws;
  a.tile_hits = (uint32_t*)((char*)d_ws + bw_flags_bytes(n));
  a.chunk_slow = a.tile_hits + bw_tiles(n);
  a.chunk_nfloat = d_chunk_nval; a.af = d_val; a.ac = d_char; a.ad = d_disp;
  a.tiles_per_chunk = single ? 0u : (uint32_t)min(chunk / DCB_BW_NT, (uint64_t)0xFFFFFFFFu);
  if (!single && chunk <= 65536 && !bw_three_pass()) {     // short chunks: one kernel, one block per chunk
    dcb_bw_fused_kernel<T><<<(unsigned)K, DCB_BW_NT, 0, st>>>(a);
    CK(cudaGetLastError());
    return DCB_OK;
  }
  CK(cudaMemsetAsync(a.chunk_slow, 0, 4 * K, st));
  dcb_bw_classify_kernel<T><<<(unsigned)bw_tiles(n), DCB_BW_NT, 0, st>>>(a);
  CK(cudaGetLastError());
  dcb_bw_slow_kernel<T><<<(unsigned)((K + 63) / 64), 64, 0, st>>>(a, K);
  CK(cudaGetLastError());
  dcb_bw_offsets_kernel<T><<<(unsigned)K, DCB_BW_NT, 0, st>>>(a);
  CK(cudaGetLastError());
  dcb_bw_scatter_kernel<T><<<(unsigned)bw_tiles(n), DCB_BW_NT, 0, st>>>(a);
  CK(cudaGetLastError());
  return DCB_OK;
}
template <typename T>
static int bytewise_decompress_t(const T* d_val, const uint8_t* d_char, const int32_t* d_disp, const uint32_t* d_chunk_nval,
                                 uint64_t n, uint64_t chunk, T* d_out, void* stream) {
  if (dcb_device_count() == 0) return DCB_ENODEVICE;
  if (n == 0) return DCB_OK;
  const bool single = single_chunk(n, chunk);
  if (!single && (chunk % DCB_TILE) != 0) return DCB_EINVAL;
  if (!d_val || !d_char || !d_disp || !d_chunk_nval || !d_out || (single && n >= (1ull << 31))) return DCB_EINVAL;
  BwDecArgs<T> a;
  a.af = d_val; a.ac = d_char; a.ad = d_disp; a.chunk_nfloat = d_chunk_nval; a.n = n; a.chunk = single ? n : chunk;
  a.K = dcb_num_chunks(n, chunk); a.out = d_out;
  dcb_bw_decode_kernel<T><<<(unsigned)((a.K + 63) / 64), 64, 0, (cudaStream_t)stream>>>(a);
  CK(cudaGetLastError());
  return DCB_OK;
}
extern "C" int dcb_bytewise_compress(const float* d_in, uint64_t n, uint64_t chunk, double eb, float* d_float, uint8_t* d_char,
                                     int32_t* d_disp, uint32_t* d_chunk_nfloat, void* d_ws, uint64_t ws_bytes, void* stream) {
  return bytewise_compress_t<float>(d_in, n, chunk, eb, d_float, d_char, d_disp, d_chunk_nfloat, d_ws, ws_bytes, stream);
}
extern "C" int dcb_bytewise_decompress(const float* d_float, const uint8_t* d_char, const int32_t* d_disp,
                                       const uint32_t* d_chunk_nfloat, uint64_t n, uint64_t chunk, float* d_out, void* stream) {
  return bytewise_decompress_t<float>(d_float, d_char, d_disp, d_chunk_nfloat, n, chunk, d_out, stream);
}
extern "C" int dcb_bytewise_compress_f64(const double* d_in, uint64_t n, uint64_t chunk, double eb, double* d_double, uint8_t* d_char,
                                         int32_t* d_disp, uint32_t* d_chunk_ndouble, void* d_ws, uint64_t ws_bytes, void* stream) {
  return bytewise_compress_t<double>(d_in, n, chunk, eb, d_double, d_char, d_disp, d_chunk_ndouble, d_ws, ws_bytes, stream);
}
extern "C" int dcb_bytewise_decompress_f64(const double* d_double, const uint8_t* d_char, const int32_t* d_disp,
                                           const uint32_t* d_chunk_ndouble, uint64_t n, uint64_t chunk, double* d_out, void* stream) {
  return bytewise_decompress_t<double>(d_double, d_char, d_disp, d_chunk_ndouble, n, chunk, d_out, stream);
}

// ---- Hamming SECDED over blocks of the finished stream (CT=10) ------------------------------------------------------------
extern "C" int dcb_hm_length(int k) { return k < 0 ? -1 : hm_length(k); }
static int hamming_launch(uint8_t* d_bits, uint64_t total, uint32_t block_bytes, uint32_t* d_check, int32_t* d_r,
                          int32_t* d_et, int decode, void* stream) {
  if (dcb_device_count() == 0) return DCB_ENODEVICE;
  if (total == 0) return DCB_OK;
  if (!d_bits || !d_check || block_bytes == 0 || block_bytes > (1u << 26)) return DCB_EINVAL;   // r <= 30 fits the packed word
  const uint64_t blocks = (total + block_bytes - 1) / block_bytes;
  if (blocks > 0x7FFFFFFFull) return DCB_EINVAL;
  HmArgs a;
  a.bits = d_bits; a.total_bytes = total; a.block_bytes = block_bytes; a.check = d_check; a.r_out = d_r; a.error_type = d_et;
  a.decode = decode;
  dcb_hamming_kernel<<<(unsigned)blocks, DCB_HM_NT, 0, (cudaStream_t)stream>>>(a);
  CK(cudaGetLastError());
  return DCB_OK;
}
extern "C" int dcb_hamming_encode(const uint8_t* d_bits, uint64_t total_bytes, uint32_t block_bytes, uint32_t* d_check,
                                  int32_t* d_r, void* stream) {
  return hamming_launch(const_cast<uint8_t*>(d_bits), total_bytes, block_bytes, d_check, d_r, nullptr, 0, stream);
}
extern "C" int dcb_hamming_decode(uint8_t* d_bits, uint64_t total_bytes, uint32_t block_bytes, uint32_t* d_check,
                                  int32_t* d_error_type, void* stream) {
  return hamming_launch(d_bits, total_bytes, block_bytes, d_check, nullptr, d_error_type, 1, stream);
}

#ifdef DCB_DEBUG_MED
extern "C" int dcb_debug_exact(unsigned long long* host_dst) {
  CK(cudaDeviceSynchronize());
  CK(cudaMemcpyFromSymbol(host_dst, g_dbg_exact, sizeof(g_dbg_exact)));
  return DCB_OK;
}
extern "C" int dcb_debug_walk(unsigned long long* host_dst) {
  CK(cudaDeviceSynchronize());
  CK(cudaMemcpyFromSymbol(host_dst, g_dbg_walk, sizeof(g_dbg_walk)));
  return DCB_OK;
}
#endif
