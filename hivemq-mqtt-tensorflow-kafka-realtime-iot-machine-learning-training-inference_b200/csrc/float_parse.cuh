// Correctly rounded decimal → binary32 for CSV float fields (host + device, so the same code is fuzzed on the CPU).
//
// Replaces the per-field `safe_strtof` inside tf.io.decode_csv (reference call site:
// python-scripts/autoencoder-anomaly-detection/Sensor-Kafka-Consumer-and-TensorFlow-Model-Training.py:32-34).
// "Bit-exact to tf.io.decode_csv" means correctly rounded (round-to-nearest-even) binary32, which is also what
// glibc strtof produces — the oracle.  A decimal→double→float shortcut double-rounds and is NOT used.
//
// Method: scan digits into a 64-bit significand w (≤ 19 digits) and a decimal exponent q, then the Eisel-Lemire
// product w × 5^q against a 128-bit table (pow5_table.inc) with the Mushtak-Lemire proof that no fallback is needed
// when w is exact.  Tokens with more than 19 significant digits are evaluated at w and w+1; if both round to the
// same float that is the answer, otherwise the record gets B2_REC_TOO_LONG (never seen for ≤ 17-digit producers
// such as Java Float/Double.toString or Python repr).
#pragma once
#include <stdint.h>

#if defined(__CUDA_ARCH__)
#define B2_POW5_DECL static __device__ const
#define B2_HD __host__ __device__ __forceinline__
#elif defined(__CUDACC__)
#define B2_POW5_DECL static const
#define B2_HD __host__ __device__ __forceinline__
#else
#define B2_POW5_DECL static const
#define B2_HD static inline
#endif

// one table per compilation pass: device code reads the __device__ copy, host code the host copy
#if defined(__CUDACC__)
namespace b2dev {
#undef B2_POW5_DECL
#define B2_POW5_DECL static __device__ const
#include "pow5_table.inc"
}  // namespace b2dev
namespace b2host {
#undef B2_POW5_DECL
#define B2_POW5_DECL static const
#include "pow5_table.inc"
}  // namespace b2host
#else
namespace b2host {
#include "pow5_table.inc"
}
#endif

enum { B2_NUM_OK = 0, B2_NUM_BAD = 1, B2_NUM_TOO_LONG = 2 };

B2_HD void b2_mul64(uint64_t a, uint64_t b, uint64_t& hi, uint64_t& lo) {
#if defined(__CUDA_ARCH__)
  lo = a * b;
  hi = __umul64hi(a, b);
#else
  unsigned __int128 p = (unsigned __int128)a * b;
  lo = (uint64_t)p;
  hi = (uint64_t)(p >> 64);
#endif
}

B2_HD int b2_clz64(uint64_t x) {
#if defined(__CUDA_ARCH__)
  return __clzll((long long)x);
#else
  return __builtin_clzll(x);
#endif
}

B2_HD uint64_t b2_pow5(int q, int half) {
#if defined(__CUDA_ARCH__)
  return b2dev::b2_pow5_128[q - B2_POW5_MIN_Q][half];
#else
  return b2host::b2_pow5_128[q - B2_POW5_MIN_Q][half];
#endif
}

// w × 10^q → IEEE binary32 bits (sign excluded), round-to-nearest-even.  w != 0 is exact.
B2_HD uint32_t b2_eisel_lemire32(uint64_t w, int q) {
  if (w == 0 || q < -64) return 0u;            // below half the smallest subnormal for any 64-bit w
  if (q > B2_POW5_MAX_Q) return 0x7F800000u;   // overflow → inf
  const int lz = b2_clz64(w);
  w <<= lz;
  uint64_t hi, lo;
  b2_mul64(w, b2_pow5(q, 0), hi, lo);
  const uint64_t precision_mask = 0xFFFFFFFFFFFFFFFFull >> 26;  // mantissa bits (23) + 3
  if ((hi & precision_mask) == precision_mask) {
    uint64_t hi2, lo2;
    b2_mul64(w, b2_pow5(q, 1), hi2, lo2);
    lo += hi2;
    if (hi2 > lo) ++hi;
  }
  const int upperbit = (int)(hi >> 63);
  const int shift = upperbit + 64 - 23 - 3;
  uint64_t m = hi >> shift;
  int p2 = (((152170 + 65536) * q) >> 16) + 63 + upperbit - lz + 127;
  if (p2 <= 0) {  // subnormal
    if (-p2 + 1 >= 64) return 0u;
    m >>= -p2 + 1;
    m += (m & 1);
    m >>= 1;
    p2 = (m < (1ull << 23)) ? 0 : 1;
    return ((uint32_t)p2 << 23) | (uint32_t)(m & 0x7FFFFFu);
  }
  if (lo <= 1 && q >= -17 && q <= 10 && (m & 3) == 1) {  // exactly halfway and even: do not round up
    if ((m << shift) == hi) m &= ~1ull;
  }
  m += (m & 1);
  m >>= 1;
  if (m >= (2ull << 23)) {
    m = 1ull << 23;
    ++p2;
  }
  m &= ~(1ull << 23);
  if (p2 >= 0xFF) return 0x7F800000u;
  return ((uint32_t)p2 << 23) | (uint32_t)m;
}

B2_HD bool b2_is_digit(uint8_t c) { return (uint8_t)(c - '0') < 10; }
B2_HD uint8_t b2_lower(uint8_t c) { return (c >= 'A' && c <= 'Z') ? (uint8_t)(c + 32) : c; }

// Parse s[0..n) as one float field.  Grammar (what both safe_strtof and strtof accept for decimal input):
//   ws* [+-]? ( digits [. digits*] | . digits+ ) ( [eE] [+-]? digits+ )? ws*   |   ws* [+-]? (inf|infinity|nan) ws*
// Hex floats and nan(payload) are rejected.  Returns B2_NUM_*; *out receives the IEEE bits.
template <typename BytePtr>
B2_HD int b2_parse_float(BytePtr s, int n, uint32_t* out) {
  int i = 0;
  while (i < n && (s[i] == ' ' || s[i] == '\t')) ++i;
  while (n > i && (s[n - 1] == ' ' || s[n - 1] == '\t')) --n;
  if (i >= n) return B2_NUM_BAD;
  uint32_t sign = 0;
  if (s[i] == '-') {
    sign = 0x80000000u;
    ++i;
  } else if (s[i] == '+') {
    ++i;
  }
  if (i >= n) return B2_NUM_BAD;
  if (!b2_is_digit(s[i]) && s[i] != '.') {  // inf / infinity / nan
    const int len = n - i;
    uint8_t c0 = b2_lower(s[i]);
    if (len == 3 && c0 == 'n' && b2_lower(s[i + 1]) == 'a' && b2_lower(s[i + 2]) == 'n') {
      *out = sign | 0x7FC00000u;
      return B2_NUM_OK;
    }
    if ((len == 3 || len == 8) && c0 == 'i' && b2_lower(s[i + 1]) == 'n' && b2_lower(s[i + 2]) == 'f') {
      if (len == 8) {
        const char* rest = "inity";
        for (int k = 0; k < 5; ++k)
          if (b2_lower(s[i + 3 + k]) != (uint8_t)rest[k]) return B2_NUM_BAD;
      }
      *out = sign | 0x7F800000u;
      return B2_NUM_OK;
    }
    return B2_NUM_BAD;
  }
  uint64_t w = 0;
  int ndig = 0;        // significant digits accumulated into w (≤ 19)
  int dropped = 0;     // significant digits beyond the 19th (value only matters through `truncated`)
  bool truncated = false, any = false;
  int q = 0;
  // integer part
  while (i < n && b2_is_digit(s[i])) {
    const uint32_t d = s[i] - '0';
    any = true;
    if (ndig < 19) {
      if (ndig || d) {
        w = w * 10 + d;
        ++ndig;
      }
    } else {
      ++dropped;
      truncated |= d != 0;
    }
    ++i;
  }
  q = dropped;
  if (i < n && s[i] == '.') {
    ++i;
    while (i < n && b2_is_digit(s[i])) {
      const uint32_t d = s[i] - '0';
      any = true;
      if (ndig < 19) {
        if (ndig || d) {
          w = w * 10 + d;
          ++ndig;
        }
        --q;
      } else {
        truncated |= d != 0;
      }
      ++i;
    }
  }
  if (!any) return B2_NUM_BAD;
  if (i < n && (s[i] == 'e' || s[i] == 'E')) {
    ++i;
    bool eneg = false;
    if (i < n && (s[i] == '-' || s[i] == '+')) {
      eneg = s[i] == '-';
      ++i;
    }
    if (i >= n || !b2_is_digit(s[i])) return B2_NUM_BAD;
    int e = 0;
    while (i < n && b2_is_digit(s[i])) {
      if (e < 100000) e = e * 10 + (s[i] - '0');
      ++i;
    }
    q += eneg ? -e : e;
  }
  if (i != n) return B2_NUM_BAD;
  if (w == 0) {
    *out = sign;
    return B2_NUM_OK;
  }
  if (q < -400) q = -400;
  if (q > 400) q = 400;
  const uint32_t a = b2_eisel_lemire32(w, q);
  if (truncated) {
    const uint32_t b = b2_eisel_lemire32(w + 1, q);
    if (a != b) return B2_NUM_TOO_LONG;
  }
  *out = sign | a;
  return B2_NUM_OK;
}
