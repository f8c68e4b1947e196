// Correctly rounded decimal → binary32 for CSV float fields (host + device, so the same code is fuzzed on the CPU).
//
// Replaces the per-field `safe_strtof` inside tf.io.decode_csv (reference call site:
// python-scripts/autoencoder-anomaly-detection/Sensor-Kafka-Consumer-and-TensorFlow-Model-Training.py:32-34).
// "Bit-exact to tf.io.decode_csv" means correctly rounded (round-to-nearest-even) binary32, which is also what
// glibc strtof produces — the oracle.  A plain decimal→double→float shortcut double-rounds and is NOT used; the
// double-precision step below (b2_dec32) is taken only when the double lands provably clear of every binary32 rounding
// midpoint, and hands everything else to the Eisel-Lemire product.
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
// 10^q as binary64, q = -22..22 (exact for q >= 0, correctly rounded reciprocals below)
B2_POW5_DECL double b2_p10d[45] = {1e-22, 1e-21, 1e-20, 1e-19, 1e-18, 1e-17, 1e-16, 1e-15, 1e-14, 1e-13, 1e-12, 1e-11, 1e-10, 1e-9, 1e-8,
                                   1e-7,  1e-6,  1e-5,  1e-4,  1e-3,  1e-2,  1e-1,  1e0,   1e1,   1e2,   1e3,   1e4,   1e5,   1e6,  1e7,
                                   1e8,   1e9,   1e10,  1e11,  1e12,  1e13,  1e14,  1e15,  1e16,  1e17,  1e18,  1e19,  1e20,  1e21, 1e22};
}  // namespace b2dev
namespace b2host {
#undef B2_POW5_DECL
#define B2_POW5_DECL static const
#include "pow5_table.inc"
// 10^q as binary64, q = -22..22 (exact for q >= 0, correctly rounded reciprocals below)
B2_POW5_DECL double b2_p10d[45] = {1e-22, 1e-21, 1e-20, 1e-19, 1e-18, 1e-17, 1e-16, 1e-15, 1e-14, 1e-13, 1e-12, 1e-11, 1e-10, 1e-9, 1e-8,
                                   1e-7,  1e-6,  1e-5,  1e-4,  1e-3,  1e-2,  1e-1,  1e0,   1e1,   1e2,   1e3,   1e4,   1e5,   1e6,  1e7,
                                   1e8,   1e9,   1e10,  1e11,  1e12,  1e13,  1e14,  1e15,  1e16,  1e17,  1e18,  1e19,  1e20,  1e21, 1e22};
}  // namespace b2host
#else
namespace b2host {
#include "pow5_table.inc"
// 10^q as binary64, q = -22..22 (exact for q >= 0, correctly rounded reciprocals below)
B2_POW5_DECL double b2_p10d[45] = {1e-22, 1e-21, 1e-20, 1e-19, 1e-18, 1e-17, 1e-16, 1e-15, 1e-14, 1e-13, 1e-12, 1e-11, 1e-10, 1e-9, 1e-8,
                                   1e-7,  1e-6,  1e-5,  1e-4,  1e-3,  1e-2,  1e-1,  1e0,   1e1,   1e2,   1e3,   1e4,   1e5,   1e6,  1e7,
                                   1e8,   1e9,   1e10,  1e11,  1e12,  1e13,  1e14,  1e15,  1e16,  1e17,  1e18,  1e19,  1e20,  1e21, 1e22};
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
  if (q > 38) return 0x7F800000u;              // overflow → inf
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

// w × 10^q → binary32 bits, correctly rounded, with a binary64 short cut for the common case (w < 2^53, |q| ≤ 22):
// d = RN(w · RN(10^q)) is within 2 ulp53 of the exact value v (one rounding in the table entry for q < 0, one in the product),
// so RN32(d) = RN32(v) unless a binary32 midpoint lies within 2 ulp53 of d.  Midpoints in d's binade are the doubles whose low
// 29 fraction bits read 0x10000000; when d's low bits are more than 4 away from that, the conversion of d is the answer.  With
// w ≥ 1 and |q| ≤ 22 the result is a normal binary32 (1e-22 ≤ v < 2^53·1e22 < 2^127), so no range cases arise.
B2_HD uint32_t b2_dec32(uint64_t w, int q) {
  if (w < (1ull << 53) && q >= -22 && q <= 22) {
#if defined(__CUDA_ARCH__)
    const double d = __dmul_rn((double)w, b2dev::b2_p10d[q + 22]);
    const uint32_t low = (uint32_t)__double2loint(d) & 0x1FFFFFFFu;
    if (low - 0x0FFFFFFCu > 8u) return __float_as_uint(__double2float_rn(d));
#else
    volatile double dv = (double)w * b2host::b2_p10d[q + 22];
    const double d = dv;
    uint64_t bits;
    __builtin_memcpy(&bits, &d, 8);
    const uint32_t low = (uint32_t)bits & 0x1FFFFFFFu;
    if (low - 0x0FFFFFFCu > 8u) {
      const float f = (float)d;
      uint32_t fb;
      __builtin_memcpy(&fb, &f, 4);
      return fb;
    }
#endif
  }
  return b2_eisel_lemire32(w, q);
}

// ---- decimal → binary64 (the JSON wire format: KSQL's DOUBLE columns, then one more rounding to the model's float32) -------
// Same Eisel-Lemire step with binary64 parameters; exact for any 64-bit w (Mushtak & Lemire, "Fast number parsing without
// fallback", 2023).  Clinger's short cut in front: w < 2^53 and |q| ≤ 22 are exactly representable operands of ONE correctly
// rounded multiplication or division.
B2_HD uint64_t b2_eisel_lemire64(uint64_t w, int q) {
  if (w == 0 || q < -342) return 0ull;
  if (q > 308) return 0x7FF0000000000000ull;
  const int lz = b2_clz64(w);
  w <<= lz;
  uint64_t hi, lo;
  b2_mul64(w, b2_pow5(q, 0), hi, lo);
  const uint64_t precision_mask = 0xFFFFFFFFFFFFFFFFull >> 55;  // mantissa bits (52) + 3
  if ((hi & precision_mask) == precision_mask) {
    uint64_t hi2, lo2;
    b2_mul64(w, b2_pow5(q, 1), hi2, lo2);
    lo += hi2;
    if (hi2 > lo) ++hi;
  }
  const int upperbit = (int)(hi >> 63);
  const int shift = upperbit + 64 - 52 - 3;
  uint64_t m = hi >> shift;
  int p2 = (((152170 + 65536) * q) >> 16) + 63 + upperbit - lz + 1023;
  if (p2 <= 0) {  // subnormal
    if (-p2 + 1 >= 64) return 0ull;
    m >>= -p2 + 1;
    m += (m & 1);
    m >>= 1;
    p2 = (m < (1ull << 52)) ? 0 : 1;
    return ((uint64_t)p2 << 52) | (m & 0xFFFFFFFFFFFFFull);
  }
  if (lo <= 1 && q >= -4 && q <= 23 && (m & 3) == 1) {  // exactly halfway and even: do not round up
    if ((m << shift) == hi) m &= ~1ull;
  }
  m += (m & 1);
  m >>= 1;
  if (m >= (2ull << 52)) {
    m = 1ull << 52;
    ++p2;
  }
  m &= ~(1ull << 52);
  if (p2 >= 0x7FF) return 0x7FF0000000000000ull;
  return ((uint64_t)p2 << 52) | m;
}

B2_HD uint64_t b2_dec64(uint64_t w, int q) {
  if (w < (1ull << 53) && q >= -22 && q <= 22) {
    uint64_t bits;
#if defined(__CUDA_ARCH__)
    const double p = b2dev::b2_p10d[(q < 0 ? -q : q) + 22];
    const double d = q < 0 ? __ddiv_rn((double)w, p) : __dmul_rn((double)w, p);
    bits = (uint64_t)__double_as_longlong(d);
#else
    const double p = b2host::b2_p10d[(q < 0 ? -q : q) + 22];
    volatile double dv = q < 0 ? (double)w / p : (double)w * p;
    const double d = dv;
    __builtin_memcpy(&bits, &d, 8);
#endif
    return bits;
  }
  return b2_eisel_lemire64(w, q);
}

B2_HD uint32_t b2_f64bits_to_f32bits(uint64_t bits) {  // round-to-nearest-even narrowing, like tf.cast / np.float32(double)
#if defined(__CUDA_ARCH__)
  return __float_as_uint(__double2float_rn(__longlong_as_double((long long)bits)));
#else
  double d;
  __builtin_memcpy(&d, &bits, 8);
  const float f = (float)d;
  uint32_t fb;
  __builtin_memcpy(&fb, &f, 4);
  return fb;
#endif
}

B2_HD bool b2_is_digit(uint8_t c) { return (uint8_t)(c - '0') < 10; }
B2_HD uint8_t b2_lower(uint8_t c) { return (c >= 'A' && c <= 'Z') ? (uint8_t)(c + 32) : c; }

// Parse s[0..n) as one float field.  Grammar (what both safe_strtof and strtof accept for decimal input):
//   ws* [+-]? ( digits [. digits*] | . digits+ ) ( [eE] [+-]? digits+ )? ws*   |   ws* [+-]? (inf|infinity|nan) ws*
// Hex floats and nan(payload) are rejected.  Returns B2_NUM_*; *out receives the IEEE bits.
// WIDE = false: *out32 = binary32 bits (one rounding);  WIDE = true: *out64 = binary64 bits
template <bool WIDE, typename BytePtr>
B2_HD int b2_parse_number(BytePtr s, int n, uint32_t* out32, uint64_t* out64) {
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
      if (WIDE) *out64 = ((uint64_t)sign << 32) | 0x7FF8000000000000ull;
      else *out32 = sign | 0x7FC00000u;
      return B2_NUM_OK;
    }
    if ((len == 3 || len == 8) && c0 == 'i' && b2_lower(s[i + 1]) == 'n' && b2_lower(s[i + 2]) == 'f') {
      if (len == 8) {
        const char* rest = "inity";
        for (int k = 0; k < 5; ++k)
          if (b2_lower(s[i + 3 + k]) != (uint8_t)rest[k]) return B2_NUM_BAD;
      }
      if (WIDE) *out64 = ((uint64_t)sign << 32) | 0x7FF0000000000000ull;
      else *out32 = sign | 0x7F800000u;
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
    if (WIDE) *out64 = (uint64_t)sign << 32;
    else *out32 = sign;
    return B2_NUM_OK;
  }
  if (q < -400) q = -400;
  if (q > 400) q = 400;
  if (WIDE) {
    const uint64_t a = truncated ? b2_eisel_lemire64(w, q) : b2_dec64(w, q);
    if (truncated && a != b2_eisel_lemire64(w + 1, q)) return B2_NUM_TOO_LONG;
    *out64 = ((uint64_t)sign << 32) | a;
    return B2_NUM_OK;
  }
  const uint32_t a = truncated ? b2_eisel_lemire32(w, q) : b2_dec32(w, q);
  if (truncated) {
    const uint32_t b = b2_eisel_lemire32(w + 1, q);
    if (a != b) return B2_NUM_TOO_LONG;
  }
  *out32 = sign | a;
  return B2_NUM_OK;
}

template <typename BytePtr>
B2_HD int b2_parse_float(BytePtr s, int n, uint32_t* out) {
  return b2_parse_number<false>(s, n, out, (uint64_t*)nullptr);
}
template <typename BytePtr>
B2_HD int b2_parse_double(BytePtr s, int n, uint64_t* out) {
  return b2_parse_number<true>(s, n, (uint32_t*)nullptr, out);
}


// ---- SWAR fast path for short tokens -----------------------------------------------------------------------------------
// Handles  [-] digits [. digits] [ E|e [-] d [d] ]  with at most 12 mantissa characters, i.e. everything a sensor producer's
// Float.toString emits ("39.395103", "0.034893095", "5.723834E-4", "32", "-0.5").  The token's first 16 bytes arrive in four
// little-endian 32-bit words (byte k of the token = byte k of the 128-bit value; bytes at or beyond L are don't-care) and its
// last four bytes in `tail`.  No loops, no per-byte branches: the dot is squeezed out with one byte shift + masks, the
// mantissa is zero-padded on the RIGHT to 12 digits (w × 10^k with the exponent lowered by k is the same number, and the
// Eisel-Lemire step is exact for any 64-bit w), validated and converted four digits per word.  Returns false for anything
// else (caller falls back to b2_parse_float); when it returns true the result is bit-identical to b2_parse_float's.
B2_HD uint32_t b2_fsr(uint32_t lo, uint32_t hi, int s) {  // (hi:lo) >> s, s in [0, 32]
#if defined(__CUDA_ARCH__)
  return __funnelshift_rc(lo, hi, s);
#else
  return s >= 32 ? hi : (s == 0 ? lo : (lo >> s) | (hi << (32 - s)));
#endif
}
B2_HD uint32_t b2_bytemask(int n) {  // low n bytes set, n clamped to [0, 4]
  n = n < 0 ? 0 : (n > 4 ? 4 : n);
  return b2_fsr(0xFFFFFFFFu, 0u, 32 - 8 * n);
}
B2_HD uint32_t b2_nondigit(uint32_t x) {  // x = chars ^ 0x30303030: high bit of each byte set iff that byte is not 0..9
  return ((x + 0x76767676u) | x) & 0x80808080u;
}
B2_HD uint32_t b2_eqbyte(uint32_t w, uint32_t splat) {  // 0x80 in every byte of w equal to the splatted byte (exact, no carries)
  const uint32_t x = w ^ splat;
  return ~(((x & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | x) & 0x80808080u;
}
B2_HD int b2_first_byte(uint32_t m) {  // index of the lowest byte whose 0x80 bit is set (m != 0)
#if defined(__CUDA_ARCH__)
  return (__ffs((int)m) - 1) >> 3;
#else
  return (__builtin_ffs((int)m) - 1) >> 3;
#endif
}
B2_HD uint32_t b2_four_digits(uint32_t x) {  // x = 4 digit values, most significant in the low byte → 0..9999
  const uint32_t y = x * 10u + (x >> 8);     // byte0 = 10 d0 + d1, byte2 = 10 d2 + d3 (no inter-byte carry: ≤ 99)
  return (y & 0xFFu) * 100u + ((y >> 16) & 0xFFu);
}

B2_HD bool b2_fast_token16(uint32_t t0, uint32_t t1, uint32_t t2, uint32_t t3, uint32_t tail, int L, uint32_t* out) {
  if (L < 1 || L > 16) return false;
  // sign
  const bool neg = (t0 & 0xFFu) == '-';
  const int s8 = neg ? 8 : 0;
  t0 = b2_fsr(t0, t1, s8);
  t1 = b2_fsr(t1, t2, s8);
  t2 = b2_fsr(t2, t3, s8);
  L -= neg ? 1 : 0;
  // exponent: E d | E - d | E d d | E - d d at the very end of the token
  const uint32_t c0 = tail & 0xFFu, c1 = (tail >> 8) & 0xFFu, c2 = (tail >> 16) & 0xFFu, c3 = tail >> 24;
  const uint32_t d2 = c2 - '0', d3 = c3 - '0';
  const bool isd2 = d2 < 10u, isd3 = d3 < 10u;
  const bool E0 = (c0 | 0x20u) == 'e', E1 = (c1 | 0x20u) == 'e', E2 = (c2 | 0x20u) == 'e';
  const bool e1 = E2 && isd3 && L >= 3;
  const bool e2 = E1 && c2 == '-' && isd3 && L >= 4;
  const bool e3 = E1 && isd2 && isd3 && L >= 4;
  const bool e4 = E0 && c1 == '-' && isd2 && isd3 && L >= 5;
  const int elen = e1 ? 2 : ((e2 || e3) ? 3 : (e4 ? 4 : 0));
  int ex = (int)d3 + ((e3 || e4) ? 10 * (int)d2 : 0);
  ex = elen ? ((e2 || e4) ? -ex : ex) : 0;
  const int Lm = L - elen;
  if (Lm < 1 || Lm > 12) return false;
  // first '.' inside the mantissa
  const uint32_t m0 = b2_eqbyte(t0, 0x2E2E2E2Eu), m1 = b2_eqbyte(t1, 0x2E2E2E2Eu), m2 = b2_eqbyte(t2, 0x2E2E2E2Eu);
  int pd = m0 ? b2_first_byte(m0) : (m1 ? 4 + b2_first_byte(m1) : (m2 ? 8 + b2_first_byte(m2) : 12));
  pd = pd < Lm ? pd : Lm;
  const int has_dot = pd < Lm ? 1 : 0;
  const int nd = Lm - has_dot;  // digit characters
  if (nd < 1) return false;
  // squeeze the dot out: bytes before it stay, bytes after it move down by one; pad with '0' from byte nd on
  const uint32_t u0 = b2_fsr(t0, t1, 8), u1 = b2_fsr(t1, t2, 8), u2 = t2 >> 8;
  const uint32_t a0 = b2_bytemask(pd), a1 = b2_bytemask(pd - 4), a2 = b2_bytemask(pd - 8);
  const uint32_t k0 = b2_bytemask(nd), k1 = b2_bytemask(nd - 4), k2 = b2_bytemask(nd - 8);
  uint32_t r0 = (t0 & a0) | (u0 & ~a0), r1 = (t1 & a1) | (u1 & ~a1), r2 = (t2 & a2) | (u2 & ~a2);
  r0 = (r0 & k0) | (0x30303030u & ~k0);
  r1 = (r1 & k1) | (0x30303030u & ~k1);
  r2 = (r2 & k2) | (0x30303030u & ~k2);
  const uint32_t x0 = r0 ^ 0x30303030u, x1 = r1 ^ 0x30303030u, x2 = r2 ^ 0x30303030u;
  if (b2_nondigit(x0) | b2_nondigit(x1) | b2_nondigit(x2)) return false;  // anything but digits (second dot, '+', blanks, …)
  const uint32_t hi8 = b2_four_digits(x0) * 10000u + b2_four_digits(x1);  // ≤ 99 999 999
  const uint64_t w = (uint64_t)hi8 * 10000u + b2_four_digits(x2);         // 12 digits, < 2^40
  const int q = ex - (Lm - has_dot - pd) - (12 - nd);                     // fraction digits and the padding zeros
  *out = (neg ? 0x80000000u : 0u) | b2_dec32(w, q);
  return true;
}

// ---- CSV numeric field straight from the byte stream (the decode kernel's lock-step field walk) ---------------------------
// The 16 bytes starting at a field's first byte arrive as five aligned little-endian words + the byte shift `sh` ((pos & 3) * 8);
// `rem` is the number of bytes left in the record from that position, `delim` the field delimiter.  One pass classifies every
// byte as digit / not-a-digit (a 16-bit map), so the field's end, the decimal point and an exponent are found with bit tricks
// instead of a separate delimiter scan, and only a strict grammar is accepted:
//      [-] digits [. digits] [ E|e [-] d [d] ]   followed by the delimiter or the record's end, at most 12 mantissa characters.
// Returns the number of characters in the field (sign included; the caller steps over them and the delimiter) with *bits set;
// 0 for an empty field; -1 for anything else ('+', blanks, quotes, inf/nan, long tokens, …) — the caller then re-does the whole
// record on the careful path, which applies the full decode_csv grammar.  When it accepts, the value is bit-identical to
// b2_parse_float's (same digits → same b2_dec32).
B2_HD uint32_t b2_byte16(uint32_t t0, uint32_t t1, uint32_t t2, uint32_t t3, int k) {  // byte k (0..15) of t3:t2:t1:t0
#if defined(__CUDA_ARCH__)
  const uint32_t lo = (k & 8) ? t2 : t0, hi = (k & 8) ? t3 : t1;
  return __byte_perm(lo, hi, (uint32_t)(k & 7)) & 0xFFu;
#else
  const uint32_t w = k < 4 ? t0 : (k < 8 ? t1 : (k < 12 ? t2 : t3));
  return (w >> ((k & 3) * 8)) & 0xFFu;
#endif
}
B2_HD uint32_t b2_flagmap16(uint32_t f0, uint32_t f1, uint32_t f2, uint32_t f3) {  // 0x80-per-byte flags of 4 words → 16-bit map
#if defined(__CUDA_ARCH__)
  const uint32_t lo = __dp4a(f1 >> 7, 0x80402010u, __dp4a(f0 >> 7, 0x08040201u, 0u));
  const uint32_t hi = __dp4a(f3 >> 7, 0x80402010u, __dp4a(f2 >> 7, 0x08040201u, 0u));
  return lo | (hi << 8);
#else
  const uint32_t M = 0x00204081u;
  return ((((f0 >> 7) * M) >> 21) & 0xFu) | ((((f1 >> 7) * M) >> 17) & 0xF0u) | ((((f2 >> 7) * M) >> 13) & 0xF00u) |
         ((((f3 >> 7) * M) >> 9) & 0xF000u);
#endif
}
B2_HD int b2_ffs32(uint32_t m) {  // 1-based index of the lowest set bit, 0 for m == 0
#if defined(__CUDA_ARCH__)
  return __ffs((int)m);
#else
  return __builtin_ffs((int)m);
#endif
}
B2_HD uint32_t b2_four_digits_fast(uint32_t x) {  // like b2_four_digits; two byte-dot-products on the device
#if defined(__CUDA_ARCH__)
  return __dp4a(x, 0x0000010Au, 0u) * 100u + __dp4a(x, 0x010A0000u, 0u);
#else
  return b2_four_digits(x);
#endif
}

// Shared core: MODE 0 = CSV field (terminator `delim` or the record's end; result binary32, one rounding);
// MODE 1 = JSON number (terminator ',' or '}'; RFC 8259 shape checks; result binary32 by way of binary64, two roundings).
template <int MODE>
B2_HD int b2_num16(uint32_t w0, uint32_t w1, uint32_t w2, uint32_t w3, uint32_t w4, int sh, int rem, uint32_t delim, uint32_t* out) {
  const int neg = (((w0 >> sh) & 0xFFu) == '-' && rem > 0) ? 1 : 0;
  const int s = sh + 8 * neg;  // ≤ 32: the clamped funnel shift returns the upper word for 32
  const uint32_t t0 = b2_fsr(w0, w1, s), t1 = b2_fsr(w1, w2, s), t2 = b2_fsr(w2, w3, s), t3 = b2_fsr(w3, w4, s);
  rem -= neg;
  const uint32_t x0 = t0 ^ 0x30303030u, x1 = t1 ^ 0x30303030u, x2 = t2 ^ 0x30303030u, x3 = t3 ^ 0x30303030u;
  // not-a-digit flags; a byte ≥ 0xBA may carry a false flag into its successor, which only ever sends a record to the careful path
  uint32_t B = b2_flagmap16(b2_nondigit(x0), b2_nondigit(x1), b2_nondigit(x2), b2_nondigit(x3)) | 0x10000u;
  if (rem < 16) B |= 0xFFFFFFFFu << rem;  // everything from the record's end on terminates the field
  const int p1 = b2_ffs32(B) - 1;
  if (p1 >= 16) return -1;
  const uint32_t c1 = b2_byte16(t0, t1, t2, t3, p1);
  int Lm, pd, has_dot, ex = 0, L;
  if (MODE == 0 ? (p1 >= rem || c1 == delim) : (c1 == ',' || c1 == '}')) {
    if (p1 == 0) return neg ? -1 : 0;  // empty field ("-" alone is not a number)
    if (p1 <= 8) {
      // an integer token of at most 8 digits (tyre pressures, firmware version, time stamps' low part …): it fits 32 bits, so ONE
      // integer → binary32 conversion is the correctly rounded value (exact below 2^24).  The 8 bytes are shifted up so that the
      // last digit sits in byte 7: what follows the token falls off the top, the bytes below read as leading zeros.
      if (MODE == 1 && (p1 > rem || (p1 > 1 && (t0 & 0xFFu) == '0'))) return -1;
      const uint64_t X = (((uint64_t)x1 << 32) | x0) << (8 * (8 - p1));
      const uint32_t v = b2_four_digits_fast((uint32_t)X) * 10000u + b2_four_digits_fast((uint32_t)(X >> 32));
      float f = (float)v;
      uint32_t fb;
#if defined(__CUDA_ARCH__)
      fb = __float_as_uint(f);
#else
      __builtin_memcpy(&fb, &f, 4);
#endif
      *out = (neg ? 0x80000000u : 0u) | fb;
      return p1 + neg;
    }
    Lm = pd = L = p1;
    has_dot = 0;
  } else {
    int pe = -1;
    if (c1 == '.') {
      B &= B - 1;
      const int p2 = b2_ffs32(B) - 1;
      if (p2 >= 16) return -1;
      const uint32_t c2 = b2_byte16(t0, t1, t2, t3, p2);
      pd = p1;
      has_dot = 1;
      Lm = L = p2;
      if (!(MODE == 0 ? (p2 >= rem || c2 == delim) : (c2 == ',' || c2 == '}'))) {
        if ((c2 | 0x20u) != 'e') return -1;
        pe = p2;
      }
    } else if ((c1 | 0x20u) == 'e') {
      pe = Lm = pd = p1;
      has_dot = 0;
    } else {
      return -1;
    }
    if (pe >= 0) {  // exponent (rare for sensor values: Float.toString switches to E notation below 1e-3 and from 1e7)
      uint32_t Be = B & (B - 1);  // flags after the 'E'
      int es = pe + 1;
      if (es >= 16) return -1;
      const bool eneg = es < rem && b2_byte16(t0, t1, t2, t3, es) == '-';
      if (eneg) {
        Be &= Be - 1;
        ++es;
      }
      const int p3 = b2_ffs32(Be) - 1;
      const int ne = p3 - es;
      if (p3 >= 16 || ne < 1 || ne > 2) return -1;
      {
        const uint32_t c3 = b2_byte16(t0, t1, t2, t3, p3 & 15);
        if (!(MODE == 0 ? (p3 >= rem || c3 == delim) : (p3 < rem && (c3 == ',' || c3 == '}')))) return -1;
      }
      ex = (int)b2_byte16(t0, t1, t2, t3, es) - '0';
      if (ne == 2) ex = ex * 10 + (int)b2_byte16(t0, t1, t2, t3, es + 1) - '0';
      if (eneg) ex = -ex;
      L = p3;
    }
  }
  const int nd = Lm - has_dot;  // digit characters
  if (nd < 1 || Lm > 12) return -1;
  if (MODE == 1) {  // JSON: at least one digit on each side of a point, no leading zeros (the careful walk gives the verdict)
    if (L > rem || pd < 1 || (has_dot && Lm - pd - 1 < 1) || (pd > 1 && (t0 & 0xFFu) == '0')) return -1;
  }
  // squeeze the dot out (bytes after it move down by one) and zero everything from digit nd on: 12 digits, zero-padded on the right
  const uint32_t u0 = b2_fsr(x0, x1, 8), u1 = b2_fsr(x1, x2, 8), u2 = x2 >> 8;
  const uint32_t a0 = b2_bytemask(pd), a1 = b2_bytemask(pd - 4), a2 = b2_bytemask(pd - 8);
  const uint32_t k0 = b2_bytemask(nd), k1 = b2_bytemask(nd - 4), k2 = b2_bytemask(nd - 8);
  const uint32_t r0 = ((x0 & a0) | (u0 & ~a0)) & k0, r1 = ((x1 & a1) | (u1 & ~a1)) & k1, r2 = ((x2 & a2) | (u2 & ~a2)) & k2;
  const uint32_t hi8 = b2_four_digits_fast(r0) * 10000u + b2_four_digits_fast(r1);
  const uint64_t w = (uint64_t)hi8 * 10000u + b2_four_digits_fast(r2);
  const int q = ex - (nd - pd) - (12 - nd);  // fraction digits and the padding zeros
  if (MODE == 1) *out = b2_f64bits_to_f32bits((neg ? 0x8000000000000000ull : 0ull) | b2_dec64(w, q));
  else *out = (neg ? 0x80000000u : 0u) | b2_dec32(w, q);
  return L + neg;
}

B2_HD int b2_csv_num16(uint32_t w0, uint32_t w1, uint32_t w2, uint32_t w3, uint32_t w4, int sh, int rem, uint32_t delim, uint32_t* out) {
  return b2_num16<0>(w0, w1, w2, w3, w4, sh, rem, delim, out);
}
// a JSON number followed by ',' or '}': returns its length (≥ 1) or -1 (not the plain short form: let the careful walk decide)
B2_HD int b2_json_num16(uint32_t w0, uint32_t w1, uint32_t w2, uint32_t w3, uint32_t w4, int sh, int rem, uint32_t* out) {
  return b2_num16<1>(w0, w1, w2, w3, w4, sh, rem, 0u, out);
}
