/* dc_oracle_hamming.c -- CPU ORACLE, TEST INFRASTRUCTURE ONLY (see dc_oracle.h).
 *
 * Restatement of the Hamming SECDED protection of CT=10 (impl/dataCompression.c:5544-5878, called per block of the
 * finished stream by impl/pingpong.c:296-311 and :487-505).  The reference evaluates every check bit with a loop over
 * all codeword positions (O(r k)); here the r check bits are the XOR of the codeword positions of the set data bits
 * (the same parity sums, :5545-5566), which is also how the CUDA kernel computes them.  PINNED against the compiled
 * reference (oracle/_ref) by tests/test_oracle_hamming.py. */
#include "dc_oracle.h"

/* hmLength, :5587-5599: smallest r with 2^r - 1 - r - k >= 0 */
int dco_hm_length(int k) {
  int r = 0;
  while (((1LL << r) - 1 - r - k) < 0) r++;
  return r;
}

/* codeword position (1-based; powers of two hold check bits, :5551-5554) of data bit d (0-based) */
static uint32_t hm_pos(uint32_t d) {
  uint32_t m = d + 1, t = 0;
  while ((1u << t) <= m) t++;                 /* t = floor(log2 m) + 1 */
  uint32_t j = m + t;
  if (j >> t) j++;                            /* one more check position (2^t) lies below */
  return j;
}

static void hm_sums(const uint8_t* bits, int bytes, uint32_t* syndrome, uint32_t* ones) {
  uint32_t s = 0, n1 = 0;
  for (int i = 0; i < bytes; i++)
    for (int b = 0; b < 8; b++)               /* data bit 8 i + b is bit 7 - b of byte i (cast_bits_to_char, :5712-5722) */
      if ((bits[i] >> (7 - b)) & 1) { s ^= hm_pos((uint32_t)(8 * i + b)); n1++; }
  *syndrome = s; *ones = n1;
}

/* hamming_encode, :5742-5750 (hamming_code :5545-5585).  c gets r + 1 chars '0'/'1'; returns r. */
int dco_hamming_encode(const uint8_t* bits, int bytes, char* c) {
  const int r = dco_hm_length(bytes * 8);
  uint32_t s, n1;
  hm_sums(bits, bytes, &s, &n1);
  uint32_t sum = n1;
  for (int i = 0; i < r; i++) { c[i] = (char)('0' + ((s >> i) & 1)); sum += (s >> i) & 1; }
  c[r] = (char)('0' + (sum & 1));             /* parity over data and check bits (secded), :5568-5578 */
  return r;
}

/* hamming_decode, :5752-5779 (hamming_verify_bit :5782-5819, error_info :5632-5655, hamming_rectify_bit :5822-5858):
 * returns 0 no error, 1 two-bit error, 2 parity bit error (c[r] flipped back), 3 one-bit error (rectified in bits or c). */
int dco_hamming_decode(uint8_t* bits, char* c, int bytes, int r) {
  uint32_t s, n1;
  hm_sums(bits, bytes, &s, &n1);
  uint32_t pos = 0, sum = n1;
  for (int i = 0; i < r; i++) {
    const uint32_t ci = (uint32_t)(c[i] - '0');
    if (((s >> i) & 1) != ci) pos |= 1u << i;
    sum += ci;
  }
  const int vr = ((sum & 1) != (uint32_t)(c[r] - '0'));
  if (pos > 0 && !vr) return 1;
  if (pos == 0 && vr) { c[r] = (c[r] == '0') ? '1' : '0'; return 2; }
  if (pos > 0 && vr) {
    if ((pos & (pos - 1)) == 0) {             /* a check-bit position */
      int cn = 0; while ((1u << cn) < pos) cn++;
      if (cn < r) c[cn] = (c[cn] == '0') ? '1' : '0';
    } else {
      uint32_t t = 0; while ((1u << t) <= pos) t++;      /* powers of two below pos */
      const uint32_t d = pos - 1 - t;
      if (d < (uint32_t)bytes * 8u) bits[d / 8] ^= (uint8_t)(1u << (7 - d % 8));
    }
    return 3;
  }
  return 0;
}
