// One flat JSON object per record → the wanted numeric members as float32 (+ a true/false label), host+device.
//
// Upstream wire format of the path: the car simulator publishes JSON objects to the `sensor-data` topic
// (fixture: python-scripts/LSTM-TensorFlow-IO-Kafka/cardata-v1.json), KSQL reads them with value_format='JSON' as DOUBLE /
// INT / STRING columns (infrastructure/confluent/01_installConfluentPlatform.sh:235) and re-publishes Avro, which the scripts
// then cast to float32 (cardata-v1.py:16-37).  This walker takes the JSON directly: number text → binary64 (correctly
// rounded, what a JSON parser yields) → binary32 (round-to-nearest-even, the cast) — two roundings, on purpose.
//
// Grammar: RFC 8259 for one object whose wanted members are numbers, strings, true/false/null; members the spec does not
// list are skipped whatever they are (nested containers included).  A wanted member that is absent or null yields 0.0 (KSQL
// NULL → Avro null → 0, like the Avro walk); duplicate names: the last one wins.  Anything malformed gives the record a
// status (B2_REC_BAD_JSON / B2_REC_BAD_NUMBER) and never stops the batch.
#pragma once
#include "float_parse.cuh"

#ifndef B2_REC_OK
#define B2_REC_OK 0
#define B2_REC_BAD_NUMBER 2
#define B2_REC_TOO_LONG 7
#endif
#ifndef B2_REC_BAD_JSON
#define B2_REC_BAD_JSON 8
#endif

#define B2_JSON_MAX_KEYS 64
#define B2_JSON_NAME_BYTES 32

struct JsonP {
  int32_t n_keys;
  int32_t n_out;
  uint32_t sig[B2_JSON_MAX_KEYS];                          // length | first << 8 | last << 16 | middle << 24
  int8_t out[B2_JSON_MAX_KEYS];                            // output column, -2 = label, -1 = matched but ignored
  uint8_t len[B2_JSON_MAX_KEYS];
  uint32_t name[B2_JSON_MAX_KEYS][B2_JSON_NAME_BYTES / 4];  // zero-padded, little-endian words
  // the member as a producer without white space writes it, separator included:  {"name":  for the first,  ,"name":  after
  uint8_t plen[B2_JSON_MAX_KEYS];
  uint32_t pat[B2_JSON_MAX_KEYS][(B2_JSON_NAME_BYTES + 4 + 3) / 4];
};

// fills sig / plen / pat from name / len / out (host side, once per launch)
inline void b2_json_finish_spec(JsonP& p) {
  for (int k = 0; k < p.n_keys; ++k) {
    const int len = p.len[k];
    const uint8_t* nm = reinterpret_cast<const uint8_t*>(p.name[k]);
    p.sig[k] = (uint32_t)len | ((uint32_t)nm[0] << 8) | ((uint32_t)nm[len - 1] << 16) | ((uint32_t)nm[len >> 1] << 24);
    uint8_t buf[B2_JSON_NAME_BYTES + 8] = {0};
    buf[0] = k == 0 ? '{' : ',';
    buf[1] = '"';
    for (int i = 0; i < len; ++i) buf[2 + i] = nm[i];
    buf[2 + len] = '"';
    buf[3 + len] = ':';
    p.plen[k] = (uint8_t)(len + 4);
    for (int w = 0; w < (B2_JSON_NAME_BYTES + 4 + 3) / 4; ++w)
      p.pat[k][w] = (uint32_t)buf[4 * w] | ((uint32_t)buf[4 * w + 1] << 8) | ((uint32_t)buf[4 * w + 2] << 16) | ((uint32_t)buf[4 * w + 3] << 24);
  }
}

B2_HD bool b2_json_ws(uint32_t c) { return c == ' ' || c == '\t' || c == '\n' || c == '\r'; }

// RFC 8259 number:  -? (0 | [1-9][0-9]*) (\.[0-9]+)? ([eE][+-]?[0-9]+)?
template <typename BytePtr>
B2_HD bool b2_json_number_valid(BytePtr s, int n) {
  int i = 0;
  if (i < n && s[i] == '-') ++i;
  if (i >= n) return false;
  if (s[i] == '0') {
    ++i;
  } else if (s[i] >= '1' && s[i] <= '9') {
    while (i < n && b2_is_digit(s[i])) ++i;
  } else {
    return false;
  }
  if (i < n && s[i] == '.') {
    ++i;
    if (i >= n || !b2_is_digit(s[i])) return false;
    while (i < n && b2_is_digit(s[i])) ++i;
  }
  if (i < n && (s[i] == 'e' || s[i] == 'E')) {
    ++i;
    if (i < n && (s[i] == '+' || s[i] == '-')) ++i;
    if (i >= n || !b2_is_digit(s[i])) return false;
    while (i < n && b2_is_digit(s[i])) ++i;
  }
  return i == n;
}

// skips a string whose opening quote is at s[i]; returns the index after the closing quote or -1
template <typename BytePtr>
B2_HD int b2_json_skip_string(BytePtr s, int i, int e, bool* has_escape = nullptr) {
  ++i;
  if (has_escape) *has_escape = false;
  while (i < e) {
    const uint32_t c = s[i];
    if (c == '"') return i + 1;
    if (c < 0x20) return -1;  // raw control characters are not allowed inside strings
    if (c == '\\') {
      if (has_escape) *has_escape = true;
      if (i + 1 >= e) return -1;
      const uint32_t x = s[i + 1];
      if (x == 'u') {
        if (i + 5 >= e) return -1;
        for (int k = 2; k < 6; ++k) {
          const uint32_t h = s[i + k] | 0x20u;
          if (!((h >= '0' && h <= '9') || (h >= 'a' && h <= 'f'))) return -1;
        }
        i += 6;
        continue;
      }
      if (!(x == '"' || x == '\\' || x == '/' || x == 'b' || x == 'f' || x == 'n' || x == 'r' || x == 't')) return -1;
      i += 2;
      continue;
    }
    ++i;
  }
  return -1;
}

// a scalar that is not a string, starting at s[i]: number / true / false / null; returns the index after it or -1
template <typename BytePtr>
B2_HD int b2_json_skip_scalar(BytePtr s, int i, int e) {
  int j = i;
  while (j < e) {
    const uint32_t c = s[j];
    if (c == ',' || c == '}' || c == ']' || b2_json_ws(c)) break;
    ++j;
  }
  const int n = j - i;
  if (n == 4 && s[i] == 't' && s[i + 1] == 'r' && s[i + 2] == 'u' && s[i + 3] == 'e') return j;
  if (n == 4 && s[i] == 'n' && s[i + 1] == 'u' && s[i + 2] == 'l' && s[i + 3] == 'l') return j;
  if (n == 5 && s[i] == 'f' && s[i + 1] == 'a' && s[i + 2] == 'l' && s[i + 3] == 's' && s[i + 4] == 'e') return j;
  return b2_json_number_valid(s + i, n) ? j : -1;
}

// skips any value starting at s[i] (first byte already known not to be white space); returns the index after it or -1.
// Containers are validated, not just bracket-matched: a 64-level stack of "object or array" bits and a three-state machine.
template <typename BytePtr>
B2_HD int b2_json_skip_value(BytePtr s, int i, int e) {
  const uint32_t c0 = s[i];
  if (c0 == '"') return b2_json_skip_string(s, i, e);
  if (c0 != '{' && c0 != '[') return b2_json_skip_scalar(s, i, e);
  uint64_t stack = 0;  // bit d-1: container at depth d is an object
  int depth = 0;
  enum { VALUE, VALUE_OR_CLOSE, AFTER_VALUE, KEY, KEY_OR_CLOSE, COLON };
  int st = VALUE;
  while (i < e) {
    const uint32_t c = s[i];
    if (b2_json_ws(c)) {
      ++i;
      continue;
    }
    if (st == VALUE || st == VALUE_OR_CLOSE) {
      if (c == ']' && st == VALUE_OR_CLOSE) {
        --depth;
        ++i;
        if (depth == 0) return i;
        st = AFTER_VALUE;
      } else if (c == '{' || c == '[') {
        if (depth >= 64) return -1;
        stack = c == '{' ? (stack | (1ull << depth)) : (stack & ~(1ull << depth));
        ++depth;
        ++i;
        st = c == '{' ? KEY_OR_CLOSE : VALUE_OR_CLOSE;
      } else {
        i = c == '"' ? b2_json_skip_string(s, i, e) : b2_json_skip_scalar(s, i, e);
        if (i < 0) return -1;
        st = AFTER_VALUE;
      }
    } else if (st == KEY || st == KEY_OR_CLOSE) {
      if (c == '}' && st == KEY_OR_CLOSE) {
        --depth;
        ++i;
        if (depth == 0) return i;
        st = AFTER_VALUE;
      } else if (c == '"') {
        i = b2_json_skip_string(s, i, e);
        if (i < 0) return -1;
        st = COLON;
      } else {
        return -1;
      }
    } else if (st == COLON) {
      if (c != ':') return -1;
      ++i;
      st = VALUE;
    } else {  // AFTER_VALUE: ',' or the close of the current container
      const bool in_obj = (stack >> (depth - 1)) & 1ull;
      if (c == ',') {
        ++i;
        st = in_obj ? KEY : VALUE;
      } else if (c == (in_obj ? '}' : ']')) {
        --depth;
        ++i;
        if (depth == 0) return i;
      } else {
        return -1;
      }
    }
  }
  return -1;
}

template <typename BytePtr>
B2_HD uint32_t b2_json_load_u32(BytePtr s, int i) {  // four bytes at any alignment (may read up to 3 bytes past a token)
  return (uint32_t)s[i] | ((uint32_t)s[i + 1] << 8) | ((uint32_t)s[i + 2] << 16) | ((uint32_t)s[i + 3] << 24);
}

// A member name written with escapes ("speed"): unescape on the fly and compare code point by code point (names in the
// spec are Latin-1 bytes; a code point above U+00FF cannot match).  Rare, so it is a plain loop over every wanted name.
template <typename BytePtr>
B2_HD int b2_json_match_escaped_key(BytePtr s, int kb, int ke, const JsonP& p) {
  for (int k = 0; k < p.n_keys; ++k) {
    int i = kb, j = 0;
    const int len = p.len[k];
    bool same = true;
    while (i < ke && same) {
      uint32_t cp = s[i];
      if (cp == '\\') {
        const uint32_t x = s[i + 1];
        if (x == 'u') {
          cp = 0;
          for (int t = 2; t < 6; ++t) {
            const uint32_t h = s[i + t] | 0x20u;
            cp = cp * 16 + (h <= '9' ? h - '0' : h - 'a' + 10);
          }
          i += 6;
        } else {
          cp = x == 'b' ? 8 : x == 'f' ? 12 : x == 'n' ? 10 : x == 'r' ? 13 : x == 't' ? 9 : x;
          i += 2;
        }
      } else {
        ++i;
      }
      same = j < len && cp == ((p.name[k][j >> 2] >> (8 * (j & 3))) & 0xFFu);
      ++j;
    }
    if (same && j == len) return k;
  }
  return -1;
}

// Index of the wanted key equal to s[kb, ke), or -1.  `guess` is tried first (members normally arrive in the order of the
// spec, so all lanes of a warp ask for the same entry and the table read is a uniform one).
template <typename BytePtr>
B2_HD int b2_json_match_key(BytePtr s, int kb, int ke, const JsonP& p, int guess) {
  const int len = ke - kb;
  if (len < 1 || len > B2_JSON_NAME_BYTES) return -1;
  const uint32_t sig = (uint32_t)len | ((uint32_t)s[kb] << 8) | ((uint32_t)s[ke - 1] << 16) | ((uint32_t)s[kb + (len >> 1)] << 24);
  for (int t = -1; t < p.n_keys; ++t) {
    const int k = t < 0 ? guess : t;
    if (k >= p.n_keys || p.sig[k] != sig) continue;
    bool same = true;
    for (int w = 0; w * 4 < len; ++w) {
      const int left = len - w * 4;
      const uint32_t have = b2_json_load_u32(s, kb + w * 4) & (left >= 4 ? 0xFFFFFFFFu : ((1u << (8 * left)) - 1u));
      same = same && have == p.name[k][w];
    }
    if (same) return k;
  }
  return -1;
}

B2_HD uint32_t b2_ld32(const uint8_t* s, int i) {  // four bytes at any alignment; s itself is 4-byte aligned on the device
#if defined(__CUDA_ARCH__)
  const uint32_t* wp = reinterpret_cast<const uint32_t*>(s + (i & ~3));
  return __funnelshift_r(wp[0], wp[1], (i & 3) * 8);
#else
  uint32_t v;
  __builtin_memcpy(&v, s + i, 4);
  return v;
#endif
}

// Fast prefix: as long as the object reads exactly  {"name0":number,"name1":number,…  in the order of the spec and without
// white space (what the simulator and every compact serialiser emit), a member costs one word-wise comparison against its
// pre-built pattern and one SWAR number parse — no byte loops.  Returns how many leading members were taken this way and
// the position right behind the last value; the careful walk continues from there (or from the start when 0).
B2_HD int b2_json_fast_prefix(const uint8_t* s, int b, int e, const JsonP& p, float* xrow, uint8_t* label, int* resume) {
  int pos = b, k = 0;
  for (; k < p.n_keys; ++k) {
    const int plen = p.plen[k];
    if (pos + plen + 2 > e) break;
    bool same = true;
    for (int w = 0; w * 4 < plen; ++w) {
      const int left = plen - 4 * w;
      uint32_t have = b2_ld32(s, pos + 4 * w);
      if (left < 4) have &= (1u << (8 * left)) - 1u;
      same = same && have == p.pat[k][w];
    }
    if (!same) break;
    int v = pos + plen;
    const int col = p.out[k];
    if (col >= 0) {
#if defined(__CUDA_ARCH__)
      const uint32_t* wp = reinterpret_cast<const uint32_t*>(s + (v & ~3));
      const uint32_t w0 = wp[0], w1 = wp[1], w2 = wp[2], w3 = wp[3], w4 = wp[4];
      const int sh = (v & 3) * 8;
#else
      const uint32_t w0 = b2_ld32(s, v), w1 = b2_ld32(s, v + 4), w2 = b2_ld32(s, v + 8), w3 = b2_ld32(s, v + 12), w4 = b2_ld32(s, v + 16);
      const int sh = 0;
#endif
      uint32_t bits = 0;
      const int r = b2_json_num16(w0, w1, w2, w3, w4, sh, e - v, &bits);
      if (r < 1) break;
#if defined(__CUDA_ARCH__)
      xrow[col] = __uint_as_float(bits);
#else
      __builtin_memcpy(&xrow[col], &bits, 4);
#endif
      v += r;
    } else if (col == -2) {
      const uint32_t a = b2_ld32(s, v), c = b2_ld32(s, v + 4);
      if (a == 0x75727422u && (c & 0xFFFFu) == 0x2265u && v + 6 < e) {  // "true"
        *label = 1;
        v += 6;
      } else if (a == 0x6C616622u && (c & 0xFFFFFFu) == 0x226573u && v + 7 < e) {  // "false"
        *label = 0;
        v += 7;
      } else {
        break;
      }
    } else {
      break;
    }
    pos = v;
  }
  *resume = pos;
  return k;
}

// The record walk.  xrow[0..n_out) must be zero-filled by the caller (absent members stay 0).  Returns B2_REC_*.
template <typename BytePtr>
B2_HD int b2_json_record(BytePtr s, int b, int e, const JsonP& p, float* xrow, uint8_t* label, int resume = -1, int guess0 = 0,
                         int lab0 = 2) {
  int i = b;
  uint8_t lab = (uint8_t)lab0;
  uint64_t bad_mask = 0, long_mask = 0;
  bool open = true;          // still inside the object
  bool want_member = true;   // next token must be a member name (else: ',' or '}')
  if (resume >= 0) {         // continue right behind a value the fast prefix has taken
    i = resume;
    want_member = false;
  } else {
    while (i < e && b2_json_ws(s[i])) ++i;
    if (i >= e || s[i] != '{') return B2_REC_BAD_JSON;
    ++i;
    while (i < e && b2_json_ws(s[i])) ++i;
    if (i < e && s[i] == '}') {
      ++i;
      open = false;
    }
  }
  if (open) {
    int guess = guess0;
    for (;;) {
      if (!want_member) {
        while (i < e && b2_json_ws(s[i])) ++i;
        if (i >= e) return B2_REC_BAD_JSON;
        if (s[i] == '}') {
          ++i;
          break;
        }
        if (s[i] != ',') return B2_REC_BAD_JSON;
        ++i;
      }
      want_member = false;
      while (i < e && b2_json_ws(s[i])) ++i;
      if (i >= e || s[i] != '"') return B2_REC_BAD_JSON;
      const int kb = i + 1;
      bool key_escaped;
      i = b2_json_skip_string(s, i, e, &key_escaped);
      if (i < 0) return B2_REC_BAD_JSON;
      const int ke = i - 1;
      while (i < e && b2_json_ws(s[i])) ++i;
      if (i >= e || s[i] != ':') return B2_REC_BAD_JSON;
      ++i;
      while (i < e && b2_json_ws(s[i])) ++i;
      if (i >= e) return B2_REC_BAD_JSON;
      const int k = key_escaped ? b2_json_match_escaped_key(s, kb, ke, p) : b2_json_match_key(s, kb, ke, p, guess);
      const int vb = i;
      i = b2_json_skip_value(s, i, e);
      if (i < 0) return B2_REC_BAD_JSON;
      if (k >= 0) {
        guess = k + 1;
        const int col = p.out[k];
        const uint32_t c0 = s[vb];
        if (col >= 0) {
          float v = 0.0f;
          const uint64_t bit = 1ull << k;  // a later member of the same name replaces the value AND its verdict (last one wins)
          bad_mask &= ~bit;
          long_mask &= ~bit;
          if (c0 == '-' || b2_is_digit((uint8_t)c0)) {
            uint64_t bits = 0;
            const int ps = b2_parse_double(s + vb, i - vb, &bits);  // grammar already checked by skip_value
            if (ps == B2_NUM_TOO_LONG) long_mask |= bit;
            else if (ps != B2_NUM_OK) bad_mask |= bit;
#if defined(__CUDA_ARCH__)
            v = __uint_as_float(b2_f64bits_to_f32bits(bits));
#else
            {
              const uint32_t fb = b2_f64bits_to_f32bits(bits);
              __builtin_memcpy(&v, &fb, 4);
            }
#endif
          } else if (c0 != 'n') {  // a string, a boolean or a container where a number is wanted
            bad_mask |= bit;
          }
          xrow[col] = v;
        } else if (col == -2) {
          // "true" / "false" as string (the simulator's failure_occurred) or as literal; anything else: 3, null: 2
          const int q = c0 == '"' ? 1 : 0;
          const int n = i - vb - 2 * q;
          const int t = vb + q;
          if (n == 4 && s[t] == 't' && s[t + 1] == 'r' && s[t + 2] == 'u' && s[t + 3] == 'e') lab = 1;
          else if (n == 5 && s[t] == 'f' && s[t + 1] == 'a' && s[t + 2] == 'l' && s[t + 3] == 's' && s[t + 4] == 'e') lab = 0;
          else if (!q && c0 == 'n') lab = 2;
          else lab = (q && n == 0) ? 2 : 3;
        }
      }
    }
  }
  while (i < e && b2_json_ws(s[i])) ++i;
  if (i != e) return B2_REC_BAD_JSON;
  *label = lab;
  return bad_mask ? B2_REC_BAD_NUMBER : (long_mask ? B2_REC_TOO_LONG : B2_REC_OK);
}

// fast prefix, then the careful walk for whatever is left (everything, when the first member already deviates)
B2_HD int b2_json_record_auto(const uint8_t* s, int b, int e, const JsonP& p, float* xrow, uint8_t* label) {
  uint8_t lab = 2;
  int resume = b;
  const int k = b2_json_fast_prefix(s, b, e, p, xrow, &lab, &resume);
  return b2_json_record(s, b, e, p, xrow, label, k > 0 ? resume : -1, k, (int)lab);
}
