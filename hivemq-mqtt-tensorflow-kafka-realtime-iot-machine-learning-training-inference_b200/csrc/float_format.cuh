// np.array2string(row) for a float32 row, as one host+device function (so the exact code the CUDA kernel runs is checked on
// the CPU against NumPy itself).
//
// Replaces the reference's result sink formatting: OutputCallback.on_predict_batch_end does
//     message = np.array2string(output)
// for every reconstructed row before handing it to the Kafka output sequence
// (python-scripts/autoencoder-anomaly-detection/cardata-v1.py:218-224, same in LSTM-…/cardata-v1.py:207-214).
// The arithmetic lives in NumPy (numpy/_core/arrayprint.py FloatingFormat + the Dragon4 printer); this file restates the
// behaviour for 1-D float32 input with NumPy's default print options (precision 8, floatmode 'maxprec', suppress False,
// linewidth 75, separator ' ', sign '-', threshold 1000 > row length):
//   * digits per element: the shortest decimal that identifies the float32 ("unique"), cut to 8 fractional digits by a
//     correctly rounded (half-even) cut when it is longer (positional mode only: a float32 never needs more than 9
//     significant digits, so the scientific mantissa is never cut);
//   * the row switches to scientific notation when max|x| ≥ 1e8, min|x|≠0 < 1e-4 or max/min > 1000 (float32 arithmetic);
//   * every element is padded to the row's widest integer part on the left and widest fraction on the right (scientific:
//     zero-filled to the longest mantissa, two exponent digits); trailing zeros are trimmed, the point always stays ("0.");
//   * elements are joined with one blank, lines wrap before an element that would pass column 74, continuation lines are
//     indented by one blank and lose their trailing blanks, and the whole is wrapped in [ ].
// Shortest digits come from the Schubfach construction (R. Giulietti, "The Schubfach way to render doubles", 2020) specialised
// to binary32: one 64x32-bit multiplication per interval bound against a 77-entry table of 10^e rounded up to 64 bits.
#pragma once
#include <stdint.h>

#include "float_parse.cuh"  // B2_HD, B2_POW5_DECL and the per-pass table namespaces

#if defined(__CUDACC__)
namespace b2dev {
#undef B2_POW5_DECL
#define B2_POW5_DECL static __device__ const
#include "pow10_g_table.inc"
}  // namespace b2dev
namespace b2host {
#undef B2_POW5_DECL
#define B2_POW5_DECL static const
#include "pow10_g_table.inc"
}  // namespace b2host
#else
namespace b2host {
#include "pow10_g_table.inc"
}
#endif

B2_HD uint64_t b2_pow10_g_at(int e) {
#if defined(__CUDA_ARCH__)
  return b2dev::b2_pow10_g[e - B2_POW10G_MIN_E];
#else
  return b2host::b2_pow10_g[e - B2_POW10G_MIN_E];
#endif
}

// g * cp / 2^64, rounded to odd (sticky low bit), g 64-bit, cp 32-bit
B2_HD uint32_t b2_round_to_odd(uint64_t g, uint32_t cp) {
  const uint64_t lo = (g & 0xFFFFFFFFull) * cp;
  const uint64_t hi = (g >> 32) * cp + (lo >> 32);  // bits 32..95 of the 96-bit product
  const uint32_t y1 = (uint32_t)(hi >> 32), y0 = (uint32_t)hi;
  return y1 | (y0 > 1u ? 1u : 0u);
}

// Shortest decimal of a finite, non-zero binary32 magnitude: value = digits × 10^k10, `digits` without trailing zeros.
struct B2Dec32 {
  uint32_t digits;
  int k10;
  int nd;  // number of decimal digits in `digits`
};

B2_HD int b2_count_digits(uint32_t v) {  // v ≥ 1
  return v < 10u ? 1 : v < 100u ? 2 : v < 1000u ? 3 : v < 10000u ? 4 : v < 100000u ? 5 : v < 1000000u ? 6 : v < 10000000u ? 7 : v < 100000000u ? 8 : v < 1000000000u ? 9 : 10;
}

B2_HD B2Dec32 b2_shortest32(uint32_t abs_bits) {
  const uint32_t frac = abs_bits & 0x7FFFFFu, be = abs_bits >> 23;
  uint32_t c, s;
  int q, k;
  bool done = false;
  if (be != 0) {
    c = 0x800000u | frac;
    q = (int)be - 150;
    if (q <= 0 && q > -24 && (c & ((1u << -q) - 1u)) == 0) {  // small integer: exact
      s = c >> -q;
      k = 0;
      done = true;
    }
  } else {
    c = frac;
    q = -149;
  }
  if (!done) {
    const bool even = (c & 1u) == 0;
    const bool closer = frac == 0 && be > 1;  // lower neighbour is half as far away
    const uint32_t cbl = 4u * c - 2u + (closer ? 1u : 0u), cb = 4u * c, cbr = 4u * c + 2u;
    k = (q * 1262611 - (closer ? 524031 : 0)) >> 22;  // floor(log10(2^q)) resp. floor(log10(3/4 · 2^q))
    const int h = q + ((-k * 1741647) >> 19) + 1;     // q + floor(log2(10^-k)) + 1, in 1..4
    const uint64_t g = b2_pow10_g_at(-k);
    const uint32_t vbl = b2_round_to_odd(g, cbl << h), vb = b2_round_to_odd(g, cb << h), vbr = b2_round_to_odd(g, cbr << h);
    const uint32_t lower = vbl + (even ? 0u : 1u), upper = vbr - (even ? 0u : 1u);
    s = vb >> 2;
    bool picked = false;
    if (s >= 10u) {
      const uint32_t sp = s / 10u;
      const bool up_in = lower <= 40u * sp, wp_in = 40u * sp + 40u <= upper;
      if (up_in != wp_in) {
        s = sp + (wp_in ? 1u : 0u);
        k += 1;
        picked = true;
      }
    }
    if (!picked) {
      const bool u_in = lower <= 4u * s, w_in = 4u * s + 4u <= upper;
      if (u_in != w_in) {
        s += w_in ? 1u : 0u;
      } else {
        const uint32_t mid = 4u * s + 2u;
        s += (vb > mid || (vb == mid && (s & 1u))) ? 1u : 0u;
      }
    }
  }
  while (s % 10u == 0u) {  // at most 9 rounds
    s /= 10u;
    ++k;
  }
  B2Dec32 r;
  r.digits = s;
  r.k10 = k;
  r.nd = b2_count_digits(s);
  return r;
}

// round-half-even(|x| · 10^8) for a finite binary32 below 1e8 (fits 54 bits)
B2_HD uint64_t b2_scaled_1e8(uint32_t abs_bits) {
  const uint32_t frac = abs_bits & 0x7FFFFFu, be = abs_bits >> 23;
  const uint64_t c = be ? (0x800000u | frac) : frac;
  const int q = be ? (int)be - 150 : -149;
  const uint64_t n = c * 100000000ull;  // < 2^51
  if (q >= 0) return n << q;            // integers below 1e8: q ≤ 3
  const int sh = -q;
  if (sh >= 64) return 0;
  const uint64_t r = n >> sh, rem = n & ((1ull << sh) - 1ull), half = 1ull << (sh - 1);
  return r + ((rem > half || (rem == half && (r & 1ull))) ? 1ull : 0ull);
}

B2_HD uint32_t b2_pow10_u32(int n) {  // 10^n, n in 0..9
  uint32_t p = 1;
  for (int i = 0; i < n; ++i) p *= 10u;
  return p;
}

// |x| rendered with exactly P mantissa digits after the point (P ≤ 8): D = round-half-even(|x| · 10^(P - e10)) with
// 10^P ≤ D < 10^(P+1), exact.  This is what NumPy ≥ 1.21 prints when a row's precision is longer than an element's own shortest
// form (Dragon4 with min_digits keeps producing true digits; past the shortest form both margin tests hold and its final
// rounding is the plain comparison with one half).  Five 32-bit limbs carry c · 5^t (≤ 148 bits) or c · 2^s.
struct B2Big {
  uint32_t w[5];
};
B2_HD void b2_big_mul(B2Big& a, uint32_t m) {
  uint64_t carry = 0;
  for (int i = 0; i < 5; ++i) {
    const uint64_t t = (uint64_t)a.w[i] * m + carry;
    a.w[i] = (uint32_t)t;
    carry = t >> 32;
  }
}
B2_HD uint32_t b2_big_div(B2Big& a, uint32_t d) {  // a /= d, returns the remainder
  uint64_t rem = 0;
  for (int i = 4; i >= 0; --i) {
    const uint64_t t = (rem << 32) | a.w[i];
    a.w[i] = (uint32_t)(t / d);
    rem = t % d;
  }
  return (uint32_t)rem;
}
B2_HD void b2_big_shl(B2Big& a, int s) {  // s in 0..159
  const int ws = s >> 5, bs = s & 31;
  for (int i = 4; i >= 0; --i) {
    const uint32_t lo = i - ws >= 0 ? a.w[i - ws] : 0u, lo2 = i - ws - 1 >= 0 ? a.w[i - ws - 1] : 0u;
    a.w[i] = bs ? ((lo << bs) | (lo2 >> (32 - bs))) : lo;
  }
}
B2_HD bool b2_big_shr_sticky(B2Big& a, int s) {  // a >>= s (s ≥ 0), returns whether any 1 bit was shifted out
  bool sticky = false;
  if (s >= 160) {
    for (int i = 0; i < 5; ++i) sticky |= a.w[i] != 0, a.w[i] = 0;
    return sticky;
  }
  const int ws = s >> 5, bs = s & 31;
  for (int i = 0; i < ws; ++i) sticky |= a.w[i] != 0;
  if (bs) sticky |= (a.w[ws] & ((1u << bs) - 1u)) != 0;
  for (int i = 0; i < 5; ++i) {
    const uint32_t lo = i + ws < 5 ? a.w[i + ws] : 0u, hi = i + ws + 1 < 5 ? a.w[i + ws + 1] : 0u;
    a.w[i] = bs ? ((lo >> bs) | (hi << (32 - bs))) : lo;
  }
  return sticky;
}
// floor(2 · c · 2^q · 10^t) and whether that was exact; the result must fit 64 bits (it is < 2·10^10 for every caller)
B2_HD uint64_t b2_twice_scaled(uint32_t c, int q, int t, bool* exact) {
  B2Big n;
  n.w[0] = c;
  n.w[1] = n.w[2] = n.w[3] = n.w[4] = 0;
  bool ex = true;
  int sh = q + 1 + t;  // power of two still to apply (10^t = 5^t · 2^t, and the factor 2)
  if (t >= 0) {
    int r = t;
    while (r >= 13) b2_big_mul(n, 1220703125u), r -= 13;  // 5^13
    if (r > 0) {
      uint32_t m = 1;
      for (int i = 0; i < r; ++i) m *= 5u;
      b2_big_mul(n, m);
    }
    if (sh >= 0) b2_big_shl(n, sh);
    else ex = !b2_big_shr_sticky(n, -sh);
  } else {
    // dividing by 5^-t: first bring in as many of the positive powers of two as fit (keeps the division exact-or-sticky)
    if (sh > 0) {
      b2_big_shl(n, sh);  // c·2^sh < 2^(24+104+1): fits
      sh = 0;
    }
    int r = -t;
    while (r >= 13) ex &= b2_big_div(n, 1220703125u) == 0, r -= 13;
    if (r > 0) {
      uint32_t m = 1;
      for (int i = 0; i < r; ++i) m *= 5u;
      ex &= b2_big_div(n, m) == 0;
    }
    if (sh < 0) ex &= !b2_big_shr_sticky(n, -sh);
  }
  *exact = ex;
  return ((uint64_t)n.w[1] << 32) | n.w[0];
}
B2_HD uint32_t b2_round_scaled(uint32_t c, int q, int t) {  // round-half-even(c · 2^q · 10^t)
  bool exact;
  const uint64_t Q = b2_twice_scaled(c, q, t, &exact);
  uint64_t D = Q >> 1;
  if (Q & 1ull) D += (!exact || (D & 1ull)) ? 1ull : 0ull;
  return (uint32_t)D;
}
// mantissa digits D (P+1 of them) and decimal exponent of |x| = abs_bits (finite, non-zero), given the exponent e10s of its
// shortest form (which is floor(log10|x|), or one more when the shortest form rounded up to a power of ten)
B2_HD uint32_t b2_sci_digits_exact(uint32_t abs_bits, int P, int e10s, int* e10_out) {
  const uint32_t frac = abs_bits & 0x7FFFFFu, be = abs_bits >> 23;
  const uint32_t c = be ? (0x800000u | frac) : frac;
  const int q = be ? (int)be - 150 : -149;
  const uint32_t lo = b2_pow10_u32(P);
  int e10 = e10s;
  uint32_t D = b2_round_scaled(c, q, P - e10);
  if (D < lo) {  // the shortest form had crossed a power of ten: the true leading digit sits one place lower
    --e10;
    D = b2_round_scaled(c, q, P - e10);
  }
  if (D >= lo * 10u) {  // rounding carried into a new digit (…999.5 → 1000…)
    D /= 10u;
    ++e10;
  }
  *e10_out = e10;
  return D;
}

// One element, resolved to what the row layout needs.  Positional: ipart / fpart are the integer and fraction digit strings
// as numbers (fpart has exactly flen digits, leading zeros included).  Scientific: ipart = first digit, fpart = the other
// nd-1 mantissa digits, e10 = decimal exponent.
struct B2Elem {
  uint32_t ipart, fpart;
  int8_t ilen, flen;  // ilen includes the '-' sign
  int8_t cls;         // 0 finite, 1 nan, 2 inf
  int8_t neg;
  int16_t e10;
};

B2_HD B2Elem b2_elem_positional(uint32_t bits, bool true_int_digits) {
  B2Elem e;
  e.neg = (int8_t)(bits >> 31);
  e.e10 = 0;
  const uint32_t a = bits & 0x7FFFFFFFu;
  e.cls = a > 0x7F800000u ? 1 : (a == 0x7F800000u ? 2 : 0);
  e.ipart = e.fpart = 0;
  e.flen = 0;
  e.ilen = (int8_t)(1 + e.neg);
  if (e.cls != 0 || a == 0) {
    if (e.cls == 1) e.ilen = 3, e.neg = 0;  // "nan" (the sign of a NaN is not printed)
    if (e.cls == 2) e.ilen = (int8_t)(3 + e.neg);
    return e;
  }
  const B2Dec32 d = b2_shortest32(a);
  if (d.k10 > 0 && true_int_digits && (a >> 23) >= 150u) {
    // the shortest form stops left of the units digit (integers from 2^24 on): NumPy ≥ 1.21 prints the true digits up to the point
    e.ipart = (0x800000u | (a & 0x7FFFFFu)) << ((a >> 23) - 150u);
  } else if (d.k10 >= 0) {  // integer: digits then zeros (below 1e8 in positional rows)
    e.ipart = d.digits * b2_pow10_u32(d.k10);
  } else if (-d.k10 <= 8) {
    const int f = -d.k10;  // fraction digits
    if (d.nd > f) {
      const uint32_t p = b2_pow10_u32(f);
      e.ipart = d.digits / p;
      e.fpart = d.digits - e.ipart * p;
    } else {
      e.fpart = d.digits;
    }
    e.flen = (int8_t)f;
  } else {  // longer than 8 fraction digits: correctly rounded cut, then trim
    const uint64_t r = b2_scaled_1e8(a);
    e.ipart = (uint32_t)(r / 100000000ull);
    uint32_t fp = (uint32_t)(r - (uint64_t)e.ipart * 100000000ull);
    int f = 8;
    while (f > 0 && fp % 10u == 0u) {
      fp /= 10u;
      --f;
    }
    e.fpart = fp;
    e.flen = (int8_t)f;
  }
  e.ilen = (int8_t)(b2_count_digits(e.ipart ? e.ipart : 1u) + e.neg);
  return e;
}

B2_HD B2Elem b2_elem_scientific(uint32_t bits) {
  B2Elem e;
  e.neg = (int8_t)(bits >> 31);
  const uint32_t a = bits & 0x7FFFFFFFu;
  e.cls = a > 0x7F800000u ? 1 : (a == 0x7F800000u ? 2 : 0);
  e.ipart = e.fpart = 0;
  e.flen = 0;
  e.e10 = 0;
  e.ilen = (int8_t)(1 + e.neg);
  if (e.cls != 0 || a == 0) {
    if (e.cls == 1) e.ilen = 3, e.neg = 0;
    if (e.cls == 2) e.ilen = (int8_t)(3 + e.neg);
    return e;
  }
  const B2Dec32 d = b2_shortest32(a);
  const uint32_t p = b2_pow10_u32(d.nd - 1);
  e.ipart = d.digits / p;
  e.fpart = d.digits - e.ipart * p;
  e.flen = (int8_t)(d.nd - 1);
  e.e10 = (int16_t)(d.k10 + d.nd - 1);
  return e;
}

// writes `len` decimal digits of v (leading zeros included), most significant first
template <typename Out>
B2_HD void b2_put_digits(Out& out, int pos, uint32_t v, int len) {
  for (int i = len - 1; i >= 0; --i) {
    out[pos + i] = (uint8_t)('0' + v % 10u);
    v /= 10u;
  }
}

// Formats x[0..n) (n ≥ 1, n ≤ B2_FMT_MAX_N) into out[]; returns the text length.  `out` needs room for b2_format_row_bound(n).
#define B2_FMT_MAX_N 64
B2_HD int b2_format_row_bound(int n) { return n * 20 + 2; }  // widest element: 2+1+8+4 = 15 chars, + separator / newline + indent

// `compat` selects the NumPy generation whose text is reproduced (the rules changed twice):
//   0  NumPy ≥ 2.3: scientific from max|x| ≥ 1e6 for float32 (10^min(8, finfo.precision)); elements shorter than the row's
//      precision are printed with their TRUE further digits (Dragon4 min_digits, correctly rounded)
//   1  NumPy 1.21 … 2.2: threshold 1e8, true further digits
//   2  NumPy 1.14 … 1.20 (what the reference's tensorflow:2.0.0 images shipped): threshold 1e8, shorter elements zero-filled
enum { B2_NPY_23 = 0, B2_NPY_121 = 1, B2_NPY_114 = 2 };

template <typename In, typename Out>
B2_HD int b2_format_row_f32(const In& x, int n, Out& out, int compat) {
  // ---- notation: row-wide, from the finite non-zero magnitudes ----
  float mx = 0.f, mn = 0.f;
  bool any = false, any_nonfinite = false, neg_inf = false;
  for (int i = 0; i < n; ++i) {
#if defined(__CUDA_ARCH__)
    const uint32_t b = __float_as_uint(x[i]);
#else
    uint32_t b;
    {
      const float xv = x[i];
      __builtin_memcpy(&b, &xv, 4);
    }
#endif
    const uint32_t a = b & 0x7FFFFFFFu;
    if (a >= 0x7F800000u) {
      any_nonfinite = true;
      if (a == 0x7F800000u && (b >> 31)) neg_inf = true;
      continue;
    }
    if (a == 0) continue;
#if defined(__CUDA_ARCH__)
    const float av = __uint_as_float(a);
#else
    float av;
    __builtin_memcpy(&av, &a, 4);
#endif
    if (!any || av > mx) mx = av;
    if (!any || av < mn) mn = av;
    any = true;
  }
  bool sci = false;
  if (any) {
#if defined(__CUDA_ARCH__)
    const float ratio = __fdiv_rn(mx, mn);
#else
    volatile float ratio_v = mx / mn;
    const float ratio = ratio_v;
#endif
    // 0.0001f is the float32 just below 1e-4: today's NumPy compares in float32 (min < 0.0001f); the 1.14-1.20 generation compared
    // the float32 scalar with the Python float in binary64, which makes that one value count as "below" too
    const bool tiny = compat == B2_NPY_114 ? mn <= 0.0001f : mn < 0.0001f;
    sci = mx >= (compat == B2_NPY_23 ? 1.0e6f : 1.0e8f) || tiny || ratio > 1000.0f;
  }
  // ---- widths ----
  B2Elem el[B2_FMT_MAX_N];
  int pad_left = 0, pad_right = 0;
  bool any_finite = false;
  for (int i = 0; i < n; ++i) {
#if defined(__CUDA_ARCH__)
    const uint32_t b = __float_as_uint(x[i]);
#else
    uint32_t b;
    {
      const float xv = x[i];
      __builtin_memcpy(&b, &xv, 4);
    }
#endif
    el[i] = sci ? b2_elem_scientific(b) : b2_elem_positional(b, compat != B2_NPY_114);
    if (el[i].cls == 0) {
      any_finite = true;
      if (el[i].ilen > pad_left) pad_left = el[i].ilen;
      if (el[i].flen > pad_right) pad_right = el[i].flen;
    }
  }
  const int prec = pad_right;           // scientific: mantissa digits after the point
  if (sci) pad_right = 2 + 2 + prec;    // exponent digits (always 2 for binary32) + 'e' and its sign + mantissa fraction
  if (!any_finite) pad_left = pad_right = 0;
  if (any_nonfinite) {
    const int offset = pad_right + 1;
    const int need_nan = 3 - offset, need_inf = 3 + (neg_inf ? 1 : 0) - offset;
    bool has_nan = false, has_inf = false;
    for (int i = 0; i < n; ++i) {
      has_nan |= el[i].cls == 1;
      has_inf |= el[i].cls == 2;
    }
    if (has_nan && need_nan > pad_left) pad_left = need_nan;
    // NumPy applies the inf width whenever the row has a non-finite value (nan alone included)
    if (need_inf > pad_left) pad_left = need_inf;
    (void)has_inf;
  }
  const int W = pad_left + 1 + pad_right;
  // ---- layout ----
  int pos = 0, line_len = 1;  // the first line's indent blank becomes the '['
  out[pos++] = (uint8_t)'[';
  for (int i = 0; i < n; ++i) {
    if (line_len + W > 74 && line_len > 1) {  // wrap: drop the line's trailing blanks, newline, one blank of indent
      while (pos > 0 && out[pos - 1] == (uint8_t)' ') --pos;
      out[pos++] = (uint8_t)'\n';
      out[pos++] = (uint8_t)' ';
      line_len = 1;
    }
    const B2Elem& e = el[i];
    if (e.cls != 0) {
      const int len = e.cls == 1 ? 3 : 3 + e.neg;
      for (int j = 0; j < W - len; ++j) out[pos++] = (uint8_t)' ';
      if (e.cls == 2 && e.neg) out[pos++] = (uint8_t)'-';
      const char* t = e.cls == 1 ? "nan" : "inf";
      out[pos++] = (uint8_t)t[0];
      out[pos++] = (uint8_t)t[1];
      out[pos++] = (uint8_t)t[2];
    } else {
      for (int j = 0; j < pad_left - e.ilen; ++j) out[pos++] = (uint8_t)' ';
      if (e.neg) out[pos++] = (uint8_t)'-';
      const int il = e.ilen - e.neg;
      if (!sci) {
        b2_put_digits(out, pos, e.ipart, il);
        pos += il;
        out[pos++] = (uint8_t)'.';
      }
      if (sci) {
        uint32_t ip = e.ipart, fp = e.fpart;
        int fl = e.flen, ex = e.e10;
        if (fl < prec && compat != B2_NPY_114 && (ip | fp) != 0) {  // true digits up to the row's precision
#if defined(__CUDA_ARCH__)
          const uint32_t ab = __float_as_uint(x[i]) & 0x7FFFFFFFu;
#else
          uint32_t ab;
          {
            const float xv = x[i];
            __builtin_memcpy(&ab, &xv, 4);
            ab &= 0x7FFFFFFFu;
          }
#endif
          const uint32_t D = b2_sci_digits_exact(ab, prec, ex, &ex);
          const uint32_t p10 = b2_pow10_u32(prec);
          ip = D / p10;
          fp = D - ip * p10;
          fl = prec;
        }
        b2_put_digits(out, pos, ip, 1);
        pos += 1;
        out[pos++] = (uint8_t)'.';
        b2_put_digits(out, pos, fp, fl);
        pos += fl;
        for (int j = fl; j < prec; ++j) out[pos++] = (uint8_t)'0';
        out[pos++] = (uint8_t)'e';
        out[pos++] = (uint8_t)(ex < 0 ? '-' : '+');
        b2_put_digits(out, pos, (uint32_t)(ex < 0 ? -ex : ex), 2);
        pos += 2;
      } else {
        b2_put_digits(out, pos, e.fpart, e.flen);
        pos += e.flen;
        for (int j = e.flen; j < pad_right; ++j) out[pos++] = (uint8_t)' ';
      }
    }
    line_len += W;
    if (i + 1 < n) {
      out[pos++] = (uint8_t)' ';
      line_len += 1;
    }
  }
  out[pos++] = (uint8_t)']';
  return pos;
}
