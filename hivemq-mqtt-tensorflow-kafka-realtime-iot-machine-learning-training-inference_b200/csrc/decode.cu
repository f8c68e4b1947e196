// K2 — record decode (CSV split + correctly rounded float parse; Avro zig-zag varint walk) and the fused
// decode → normalise → score kernel (the whole streaming scoring loop of cardata-v1.py:176-231 in one launch).
//
// Replaces  tf.io.decode_csv + tf.stack            Sensor-Kafka-Consumer-and-TensorFlow-Model-Training.py:32-39
//           tf.strings.substr(e, 5, -1)            AUTOENCODER-TensorFlow-IO-Kafka/cardata-v1.py:13
//           kafka_io.decode_avro(e, schema, dtype) cardata-v1.py:16-37  (+ int→float casts :61-66)
//
// HBM layout: a Kafka partition batch is `uint8 data[]` (message payloads back to back, no separators) plus
// `int32 offsets[n+1]`.  Every warp of a persistent grid is an independent worker over chunks of 32 consecutive records
// (lane i ↔ record i).  The bytes of a chunk are one contiguous range, fetched by ONE 1-D bulk async copy (TMA engine) into
// the warp's own shared-memory stage and signalled on the warp's own mbarrier; a chunk whose bytes exceed the stage is taken
// in several items.  Per item:
//   decode   lane-per-record   CSV: lock-step field walk, numbers parsed straight from the byte stream;  Avro: the schema
//                              walk (uniform field type across the warp, unaligned 32-bit loads) → the lane's f32 row
//   (then the next item is planned and its copy issued, so that it flies under the score phase)
//   score    lane-per-record   fused normalise → Dense stack → MSE → flag (ae_math.cuh), or a coalesced write-out of the rows
// Algorithmic bytes/record: payload + 4 (offset) read, 5 written (fused) or 4*F + 4 written (decode only).
#include "ae_mma.cuh"
#include "ae_tc.cuh"
#include "float_parse.cuh"
#include "json_walk.cuh"

constexpr int DEC_NT = 256;
constexpr int DEC_WARPS = DEC_NT / 32;
// two-records-per-lane kernels run as CTAs of at most 4 warps; the MINB template parameter (CTAs per SM the registers must
// allow: 5 → at most 102 registers, 4 → 128) is a launch-time choice
constexpr int DEC_NT_R2 = 128;
constexpr int STAGE_CAP_MAX = 24 * 1024;  // most bytes of payload one warp stages per item (the launch normally chooses far less)

enum { FMT_CSV = 0, FMT_AVRO = 1, FMT_JSON = 2 };
enum { AV_INT = 1, AV_LONG, AV_FLOAT, AV_DOUBLE, AV_STRING, AV_BOOLEAN, AV_BYTES, AV_NULL };

struct CsvP {
  int32_t n_fields, n_float, n_str, delim, use_quote;
  uint8_t kind[B2_MAX_FIELDS];       // 0 skip, 1 float, 2 string
  uint8_t has_default[B2_MAX_FIELDS];
  uint8_t col[B2_MAX_FIELDS];        // output column (float) or span slot (string) of field i
  uint8_t nov[B2_MAX_FIELDS];        // float field whose value is not needed (the normalisation forces its column to zero): only validated
  float dflt[B2_MAX_FIELDS];
};
struct AvroRun {
  int8_t type, nullable, count, out0;  // out0: first output column, -1 skip, -2 label
};
struct AvroP {
  int32_t n_fields, n_out, skip_prefix, all_routed;  // all_routed: every output column is written by some field
  int32_t n_runs, known, pad[2];  // known: 1 / 2 = the program is the reference's schema A / B (straight-line walk below)
  b2_avro_field f[B2_MAX_FIELDS];
  AvroRun run[B2_MAX_FIELDS];
};

template <int D>
struct DecodeArgs {
  const uint8_t* bytes;
  const int32_t* offs;
  const int32_t* ends;  // NULL: record i is [offs[i], offs[i+1]); else [offs[i], ends[i]) — ascending ranges with gaps (Kafka framing)
  long long n;
  int fmt;
  int n_out;            // floats per record produced by the decode (== D when scoring)
  int stage_cap;        // bytes of each warp's shared-memory byte stage chosen for this launch (multiple of 16)
  int warp_bytes;       // shared memory per warp: apron + stage + float tile + mbarrier
  CsvP csv;
  AvroP avro;
  JsonP json;
  // decode-only outputs
  float* out;
  int32_t* span;
  // common outputs
  int32_t* status;
  uint8_t* label;
  // fused score
  AeW<D> w;
  NormP<D> nm;
  int has_norm;
  float threshold;
  float* mse;
  uint8_t* flag;
  unsigned long long* counters;
  // records that do not fit a warp's stage: the main launch appends their row numbers to over_list ([0] = count, then rows);
  // the retry launch (list_mode = 1: one CTA, one record per chunk, STAGE_CAP_MAX bytes of stage) reads them back
  int32_t* over_list;
  int list_mode;
  const float* panels;  // tcgen05 score phase (decode_kernel_tc): packed (hi, lo) weight panels, see ae_tc.cuh
  // float64 normalisation order (b2_norm.order = B2_NORM_ORDER_F64, cardata-v3.py:53-168; decode_kernel_n64): the table as doubles
  double n64_lo[D], n64_rg[D];
  int norm64;
};

// ---- CSV, careful path step 1: field boundaries of one record (RFC-4180 quotes) ---------------------------------------------
// fb[f] = offset (relative to the stage base) of the first payload byte of field f, fb[n_fields] = end + 1, so that
// field f spans [fb[f], fb[f+1] - 1 - quoted(f)).  qmask bit f = field f was quoted.
__device__ __forceinline__ int csv_scan_slow(const uint8_t* base, int b, int e, const CsvP& c, uint16_t* fb, unsigned long long* qmask) {
  int i = b, field = 0;
  unsigned long long qm = 0;
  for (;;) {
    if (field >= c.n_fields) return B2_REC_FIELD_COUNT;
    if (c.use_quote && i < e && base[i] == '"') {
      ++i;
      fb[field] = (uint16_t)i;
      bool closed = false;
      while (i < e) {
        if (base[i] == '"') {
          if (i + 1 < e && base[i + 1] == '"') {
            i += 2;
            continue;
          }
          closed = true;
          break;
        }
        ++i;
      }
      if (!closed) return B2_REC_BAD_QUOTE;
      ++i;  // past the closing quote
      if (i < e && base[i] != c.delim) return B2_REC_BAD_QUOTE;
      qm |= 1ull << field;
    } else {
      fb[field] = (uint16_t)i;
      while (i < e && base[i] != c.delim) {
        const uint8_t ch = base[i];
        if ((c.use_quote && ch == '"') || ch == '\n' || ch == '\r') return B2_REC_BAD_QUOTE;
        ++i;
      }
    }
    ++field;
    if (i >= e) break;
    ++i;  // delimiter
    if (i == e) {  // trailing delimiter → one more empty field
      if (field >= c.n_fields) return B2_REC_FIELD_COUNT;
      fb[field++] = (uint16_t)i;
      fb[field] = (uint16_t)(i + 1);
      *qmask = qm;
      return field == c.n_fields ? B2_REC_OK : B2_REC_FIELD_COUNT;
    }
  }
  fb[field] = (uint16_t)(e + 1);
  *qmask = qm;
  return field == c.n_fields ? B2_REC_OK : B2_REC_FIELD_COUNT;
}

__device__ __forceinline__ int csv_scan(const uint8_t* base, int b, int e, const CsvP& c, uint16_t* fb, unsigned long long* qmask) {
  const uint32_t d4 = (uint32_t)c.delim * 0x01010101u;
  int nf = 0;
  uint32_t special = 0;
  fb[0] = (uint16_t)b;
  *qmask = 0;
  // 32-bit SWAR: exact byte-equality for the delimiter; "any byte below ','" as a cheap superset of quote / CR / LF (blanks
  // and '+' also trip it — they only cost the careful path, they do not change the result)
  // first and last word carry a byte mask for the record's edges; the words in between run the bare loop
  auto word = [&](int p, uint32_t keep) {
    const uint32_t wv = *reinterpret_cast<const uint32_t*>(base + p);
    const uint32_t x = wv ^ d4;
    uint32_t m = ~(((x & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | x) & keep;  // 0x80 where byte == delim
    special |= (wv - 0x2C2C2C2Cu) & ~wv & keep;                     // 0x80 where byte < 0x2C (may also flag a byte after one)
    while (m) {
      const int k = (__ffs((int)m) - 1) >> 3;
      ++nf;
      if (nf < c.n_fields) fb[nf] = (uint16_t)(p + k + 1);
      m &= m - 1;
    }
  };
  int p = b & ~3;
  if (p < e) {
    uint32_t keep = 0x80808080u << ((b - p) * 8);
    if (p + 4 > e) keep &= 0x80808080u >> ((p + 4 - e) * 8);
    word(p, keep);
    p += 4;
  }
  const int e4 = e & ~3;
#pragma unroll 1
  for (; p < e4; p += 4) word(p, 0x80808080u);
  if (p < e) word(p, 0x80808080u >> ((p + 4 - e) * 8));
  if (special || c.delim < 0x2C) return csv_scan_slow(base, b, e, c, fb, qmask);  // quotes / CR / LF possible: careful sequential scan
  if (nf + 1 != c.n_fields) return B2_REC_FIELD_COUNT;
  fb[c.n_fields] = (uint16_t)(e + 1);
  return B2_REC_OK;
}

// ---- CSV, careful path step 2: one float token, registers only -----------------------------------------------------------
// The token's first 16 bytes are pulled into registers with 5 aligned 32-bit loads + funnel shifts, its last 4 bytes with 2
// more, and handed to the branch-free SWAR parser b2_fast_token16() (float_parse.cuh; fuzzed on the CPU against strtof).
// Anything that parser declines (long mantissas, '+', blanks, inf/nan, junk) takes the general b2_parse_float().
__device__ __forceinline__ uint32_t lds_u32_unaligned(const uint8_t* __restrict__ st, int off) {
  const uint32_t* wp = reinterpret_cast<const uint32_t*>(st + (off & ~3));
  return __funnelshift_r(wp[0], wp[1], (off & 3) * 8);
}
__device__ __forceinline__ bool csv_fast_token(const uint8_t* __restrict__ st, int b, int L, uint32_t* bits) {
  const uint32_t* wp = reinterpret_cast<const uint32_t*>(st + (b & ~3));
  const int sh = (b & 3) * 8;
  const uint32_t w0 = wp[0], w1 = wp[1], w2 = wp[2], w3 = wp[3], w4 = wp[4];
  const uint32_t tail = lds_u32_unaligned(st, b + L - 4);  // may reach ≤ 3 bytes in front of the token: the stage has a 16-byte apron
  return b2_fast_token16(__funnelshift_r(w0, w1, sh), __funnelshift_r(w1, w2, sh), __funnelshift_r(w2, w3, sh), __funnelshift_r(w3, w4, sh),
                         tail, L, bits);
}

// ---- CSV, common case: lock-step field walk --------------------------------------------------------------------------------
// One record per thread; all lanes of a warp are at field f of their own record, so the field's kind / column / default are
// uniform constant-bank reads.  A numeric field is parsed straight from the 16 bytes at the cursor (b2_csv_num16: the same byte
// classification finds the field's end, so there is no separate delimiter scan); a skipped / string field is stepped over by
// a delimiter search.  (Round 2 tried a plain digit loop — LDS.U8, subtract, compare, branch, multiply-add per character, one
// 32-bit significand — in its place: 156 instead of 169 executed instructions per field, but a chain of dependent shared-memory
// loads; 5.55e9 against 5.8e9 records/s.  profiles/round2_notes.md §7.)  Returns B2_REC_OK, or -1 as soon as anything is not the plain grammar (quotes, blanks, '+', long tokens,
// missing or surplus fields, an empty field without default): the caller then runs csv_record_careful on the record.
template <int XS>
__device__ __forceinline__ int csv_walk(const uint8_t* __restrict__ stage, int b, int e, const CsvP& c, float* __restrict__ xrow,
                                        int32_t* __restrict__ span, int rel) {
  if (c.delim < 0x2C) return -1;  // the "byte below ','" test for quotes / CR / LF below needs the delimiter above them
  const uint32_t delim = (uint32_t)c.delim, d4 = delim * 0x01010101u;
  int pos = b;
  for (int f = 0; f < c.n_fields; ++f) {
    int rem = e - pos;
    if (rem < 0) return -1;  // the record ended with the previous field
    const int kind = c.kind[f];
    if (kind == 1 && c.nov[f]) {
      // the same token grammar, value unused: the digit folding and the rounding drop out of this copy of the parser
      const uint32_t* wp = reinterpret_cast<const uint32_t*>(stage + (pos & ~3));
      uint32_t unused = 0;
      const int r = b2_csv_num16(wp[0], wp[1], wp[2], wp[3], wp[4], (pos & 3) * 8, rem, delim, &unused);
      if (r < 0 || (r == 0 && !c.has_default[f])) return -1;
      pos += r + 1;
    } else if (kind == 1) {
      const uint32_t* wp = reinterpret_cast<const uint32_t*>(stage + (pos & ~3));
      uint32_t bits = 0;
      const int r = b2_csv_num16(wp[0], wp[1], wp[2], wp[3], wp[4], (pos & 3) * 8, rem, delim, &bits);
      if (r < 0) return -1;
      float v = __uint_as_float(bits);
      if (r == 0) {
        if (!c.has_default[f]) return -1;
        v = c.dflt[f];
      }
      xrow[c.col[f] * XS] = v;
      pos += r + 1;
    } else {
      const int beg = pos;
      for (;;) {
        const uint32_t* wp = reinterpret_cast<const uint32_t*>(stage + (pos & ~3));
        const int sh = (pos & 3) * 8;
        const uint32_t w0 = wp[0], w1 = wp[1], w2 = wp[2], w3 = wp[3], w4 = wp[4];
        const uint32_t t0 = __funnelshift_r(w0, w1, sh), t1 = __funnelshift_r(w1, w2, sh), t2 = __funnelshift_r(w2, w3, sh),
                       t3 = __funnelshift_r(w3, w4, sh);
        uint32_t m = b2_flagmap16(b2_eqbyte(t0, d4), b2_eqbyte(t1, d4), b2_eqbyte(t2, d4), b2_eqbyte(t3, d4)) | 0x10000u;
        if (rem < 16) m |= 0xFFFFFFFFu << rem;
        const int pz = __ffs((int)m) - 1;  // delimiter, record end, or 16 = keep looking
        // bytes below ',' in front of it (quote, CR, LF, blank, …; a borrow may flag one byte too many — that only costs time)
        const uint32_t sp = b2_flagmap16((t0 - 0x2C2C2C2Cu) & ~t0 & 0x80808080u, (t1 - 0x2C2C2C2Cu) & ~t1 & 0x80808080u,
                                         (t2 - 0x2C2C2C2Cu) & ~t2 & 0x80808080u, (t3 - 0x2C2C2C2Cu) & ~t3 & 0x80808080u);
        if (sp & ((1u << pz) - 1u)) return -1;
        pos += pz;
        if (pz < 16) break;
        rem -= 16;
      }
      if (kind == 2 && span) {
        span[2 * c.col[f]] = beg - rel;
        span[2 * c.col[f] + 1] = pos - rel;
      }
      pos += 1;
    }
  }
  return pos == e + 1 ? B2_REC_OK : -1;
}

// ---- CSV, everything else: full decode_csv grammar for one record ------------------------------------------------------------
// Delimiter scan with RFC-4180 quotes (csv_scan), then field by field through the SWAR token parser or the general one.
__device__ __noinline__ int csv_record_careful(const uint8_t* __restrict__ stage, int b, int e, const CsvP& c, float* __restrict__ xrow,
                                               int xs, int32_t* __restrict__ span, int rel) {
  uint16_t fb[B2_MAX_FIELDS + 1];
  unsigned long long qm = 0;
  int st = csv_scan(stage, b, e, c, fb, &qm);
  if (st != B2_REC_OK) return st;
  for (int f = 0; f < c.n_fields; ++f) {
    if (c.kind[f] == 0) continue;
    // a quoted field's payload sits between its quotes: drop its closing quote and, when the NEXT field is quoted too, that
    // field's opening quote (fb[] holds payload starts)
    const int quoted = (int)((qm >> f) & 1ull) + (int)((qm >> (f + 1)) & 1ull);
    const int fbeg = fb[f], fend = (int)fb[f + 1] - 1 - quoted;
    if (c.kind[f] == 1) {
      float v = 0.f;
      if (fend <= fbeg) {
        if (c.has_default[f]) v = c.dflt[f];
        else st = max(st, (int)B2_REC_EMPTY_FIELD);
      } else {
        uint32_t bits = 0;
        if (!csv_fast_token(stage, fbeg, fend - fbeg, &bits)) {
          const int ps = b2_parse_float(stage + fbeg, fend - fbeg, &bits);
          if (ps == B2_NUM_BAD) st = max(st, (int)B2_REC_BAD_NUMBER);
          else if (ps == B2_NUM_TOO_LONG) st = max(st, (int)B2_REC_TOO_LONG);
        }
        v = __uint_as_float(bits);
      }
      xrow[c.col[f] * xs] = v;
    } else if (span) {
      span[2 * c.col[f]] = fbeg - rel;
      span[2 * c.col[f] + 1] = fend - rel;
    }
  }
  return st;
}

// ---- Avro: one record per thread, whole walk -------------------------------------------------------------------------------
// Every thread of a warp is at the same field of its own record, so the field descriptor is a uniform constant-bank read and
// the type switch does not diverge; values are fetched with two aligned 32-bit loads + a funnel shift instead of byte loads.
// zig-zag varint of more than two bytes (rare: |value| ≥ 8192)
__device__ __noinline__ int av_varint_long(const uint8_t* base, int p, int e, int max_bytes, long long* v) {
  unsigned long long u = 0;
  int shift = 0;
  for (int k = 0; k < max_bytes; ++k) {
    if (p >= e) return -1;
    const uint8_t b = base[p++];
    u |= (unsigned long long)(b & 0x7F) << shift;
    shift += 7;
    if (!(b & 0x80)) {
      *v = (long long)(u >> 1) ^ -(long long)(u & 1);
      return p;
    }
  }
  return -1;
}
// returns the new position or -1; *v32 receives the value when it fits two bytes (the common case), else *v64 / wide = true
__device__ __forceinline__ int av_varint(const uint8_t* base, int p, int e, int max_bytes, int* v32, long long* v64, bool* wide) {
  *wide = false;
  const uint32_t w = lds_u32_unaligned(base, p);
  const uint32_t b0 = w & 0xFFu, b1 = (w >> 8) & 0xFFu;
  if (b0 < 0x80u) {
    *v32 = (int)(b0 >> 1) ^ -(int)(b0 & 1u);
    return p + 1 <= e ? p + 1 : -1;
  }
  if (b1 < 0x80u) {
    const uint32_t u = (b0 & 0x7Fu) | (b1 << 7);
    *v32 = (int)(u >> 1) ^ -(int)(u & 1u);
    return p + 2 <= e ? p + 2 : -1;
  }
  *wide = true;
  return av_varint_long(base, p, e, max_bytes, v64);
}

// one field of any type (strings, booleans, wide varints, nullable anything): the general, slower step
__device__ __forceinline__ int avro_field(const uint8_t* __restrict__ base, int p, int e, int type, int nullable, float* val,
                                          uint8_t* lab, int* status) {
  bool present = true, wide;
  int v32 = 0;
  long long v64 = 0;
  *val = 0.0f;
  *lab = 2;
  if (nullable) {
    p = av_varint(base, p, e, 10, &v32, &v64, &wide);
    if (p < 0) return *status = B2_REC_TRUNCATED, -1;
    if (wide || (v32 != 0 && v32 != 1)) return *status = B2_REC_BAD_BRANCH, -1;
    present = v32 != (nullable == 1 ? 0 : 1);
  }
  if (!present) return p;
  if (type == AV_FLOAT) {
    *val = __uint_as_float(lds_u32_unaligned(base, p));
    return p + 4;
  }
  if (type == AV_DOUBLE) {
    const unsigned long long bits = (unsigned long long)lds_u32_unaligned(base, p) | ((unsigned long long)lds_u32_unaligned(base, p + 4) << 32);
    *val = __double2float_rn(__longlong_as_double((long long)bits));  // one rounding, like Keras' float32 input cast
    return p + 8;
  }
  if (type == AV_INT || type == AV_LONG) {
    p = av_varint(base, p, e, type == AV_INT ? 5 : 10, &v32, &v64, &wide);
    if (p < 0) return *status = B2_REC_TRUNCATED, -1;
    // tf.cast(int32 → float32), cardata-v1.py:61-66
    *val = !wide ? __int2float_rn(v32) : (type == AV_INT ? __int2float_rn((int)v64) : __ll2float_rn(v64));
    return p;
  }
  if (type == AV_BOOLEAN) {
    *val = base[p] ? 1.0f : 0.0f;
    return p + 1;
  }
  if (type == AV_STRING || type == AV_BYTES) {
    p = av_varint(base, p, e, 10, &v32, &v64, &wide);
    if (p < 0) return *status = B2_REC_TRUNCATED, -1;
    const long long L = wide ? v64 : (long long)v32;
    if (L < 0 || (long long)p + L > e) return *status = B2_REC_TRUNCATED, -1;
    const uint32_t c4 = lds_u32_unaligned(base, p);
    if (L == 5 && c4 == 0x736C6166u && base[p + 4] == 'e') *lab = 0;  // "fals" + "e"
    else if (L == 4 && c4 == 0x65757274u) *lab = 1;                   // "true"
    else *lab = L == 0 ? 2 : 3;
    return p + (int)L;
  }
  return p;  // AV_NULL: no bytes
}

// The schema program is pre-compiled on the host into RUNS of consecutive fields that share (type, nullable) and go to
// consecutive output columns; plain float / double / small-int runs get tight type-specialised loops, everything else
// the general step above.  Fixed-size fields are not bounds-checked one by one: the walk can overshoot `e` by at most a
// few hundred bytes, which stays inside this CTA's shared memory, and the final check rejects the record.
__device__ __forceinline__ int avro_decode_record(const uint8_t* __restrict__ base, int p, int e, const AvroP& a, float* __restrict__ xrow,
                                                  int xs, uint8_t* label) {
  uint8_t lab_out = 2;
  p += a.skip_prefix;
  if (p > e) return B2_REC_TRUNCATED;
  for (int ri = 0; ri < a.n_runs; ++ri) {
    const int type = a.run[ri].type, nullable = a.run[ri].nullable, count = a.run[ri].count, out0 = a.run[ri].out0;
    if (!nullable && type == AV_FLOAT && out0 >= 0) {
      // consecutive floats share their unaligned words: count + 1 aligned loads, one funnel shift per value
      const uint32_t* wp = reinterpret_cast<const uint32_t*>(base + (p & ~3));
      const int sh = (p & 3) * 8;
      uint32_t w_prev = wp[0];
#pragma unroll 4
      for (int j = 0; j < count; ++j) {
        const uint32_t w_next = wp[j + 1];
        xrow[(out0 + j) * xs] = __uint_as_float(__funnelshift_r(w_prev, w_next, sh));
        w_prev = w_next;
      }
      p += 4 * count;
    } else if (!nullable && (type == AV_INT || type == AV_LONG) && out0 >= 0) {
      int j = 0;
      for (; count - j >= 4; j += 4) {  // four one-byte varints (tyre pressures) in one word: no per-value loads or branches
        const uint32_t w4 = lds_u32_unaligned(base, p);
        if (w4 & 0x80808080u) break;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const uint32_t u = (w4 >> (8 * q)) & 0xFFu;
          xrow[(out0 + j + q) * xs] = __int2float_rn((int)(u >> 1) ^ -(int)(u & 1u));  // tf.cast(int32 → float32), cardata-v1.py:61-66
        }
        p += 4;
      }
      for (; j < count; ++j) {  // one- and two-byte varints inline (firmware), the rest in the general step
        const uint32_t w = lds_u32_unaligned(base, p);
        const uint32_t b0 = w & 0xFFu, b1 = (w >> 8) & 0xFFu;
        if (b0 < 0x80u) {
          xrow[(out0 + j) * xs] = __int2float_rn((int)(b0 >> 1) ^ -(int)(b0 & 1u));
          p += 1;
        } else if (b1 < 0x80u) {
          const uint32_t u = (b0 & 0x7Fu) | (b1 << 7);
          xrow[(out0 + j) * xs] = __int2float_rn((int)(u >> 1) ^ -(int)(u & 1u));
          p += 2;
        } else {
          float val;
          uint8_t lab;
          int st = B2_REC_OK;
          p = avro_field(base, p, e, type, 0, &val, &lab, &st);
          if (p < 0) return st;
          xrow[(out0 + j) * xs] = val;
        }
      }
    } else if (!nullable && type == AV_DOUBLE && out0 >= 0) {
#pragma unroll 2
      for (int j = 0; j < count; ++j) {
        const unsigned long long bits =
            (unsigned long long)lds_u32_unaligned(base, p + 8 * j) | ((unsigned long long)lds_u32_unaligned(base, p + 8 * j + 4) << 32);
        xrow[(out0 + j) * xs] = __double2float_rn(__longlong_as_double((long long)bits));
      }
      p += 8 * count;
    } else if (nullable == 1 && type == AV_DOUBLE && out0 >= 0) {  // ["null", "double"]: tag 0x00 = null, 0x02 = value
      for (int j = 0; j < count; ++j) {
        const uint32_t tag = base[p];
        float val = 0.0f;
        if (tag == 2u) {
          const unsigned long long bits =
              (unsigned long long)lds_u32_unaligned(base, p + 1) | ((unsigned long long)lds_u32_unaligned(base, p + 5) << 32);
          val = __double2float_rn(__longlong_as_double((long long)bits));
          p += 9;
        } else if (tag == 0u) {
          p += 1;
        } else {  // non-minimal or invalid branch encoding: let the general step decide
          uint8_t lab;
          int st = B2_REC_OK;
          p = avro_field(base, p, e, type, nullable, &val, &lab, &st);
          if (p < 0) return st;
        }
        xrow[(out0 + j) * xs] = val;
      }
    } else {
      for (int j = 0; j < count; ++j) {
        float val;
        uint8_t lab;
        int st = B2_REC_OK;
        p = avro_field(base, p, e, type, nullable, &val, &lab, &st);
        if (p < 0) return st;
        if (out0 >= 0) xrow[(out0 + j) * xs] = val;
        else if (out0 == -2) lab_out = lab;
      }
    }
  }
  if (p > e) return B2_REC_TRUNCATED;
  *label = lab_out;
  return B2_REC_OK;
}

// ---- Avro, the reference's own two schemas: straight-line walks ---------------------------------------------------------------
// The generic walk above interprets runs with run-time counts; for the two schemas the reference commits — the simulator's
// (testdata/cardata-v1.avsc: 9 float, 4 int, 4 float, 1 int) and KSQL's (python-scripts/*/cardata-v1.avsc: the same 18 as
// ["null", double|int] unions + a ["null", string] label) — the run list is known at compile time, so the same steps are
// emitted with constant counts and columns: no loops, no descriptor reads.  A record that is not in the common shape (an int
// of more than two bytes, a union tag other than 0/2, a long label) returns -1 and takes the generic walk; results are
// identical by construction and checked against each other in tests/test_decode_gpu.py.
// ZM: bit i set = column i is overwritten with zero by the normalisation anyway (see ae_forward_rows): its value is not extracted
template <int COUNT, int OUT0, int XS, unsigned ZM = 0u>
__device__ __forceinline__ int avs_floats(const uint8_t* __restrict__ base, int p, float* __restrict__ xrow) {
  const uint32_t* wp = reinterpret_cast<const uint32_t*>(base + (p & ~3));
  const int sh = (p & 3) * 8;
  uint32_t w_prev = wp[0];
#pragma unroll
  for (int j = 0; j < COUNT; ++j) {
    const uint32_t w_next = wp[j + 1];
    if (!((ZM >> (OUT0 + j)) & 1u)) xrow[(OUT0 + j) * XS] = __uint_as_float(__funnelshift_r(w_prev, w_next, sh));
    w_prev = w_next;
  }
  return p + 4 * COUNT;
}
__device__ __forceinline__ float av_zigzag_f32(uint32_t u) { return __int2float_rn((int)(u >> 1) ^ -(int)(u & 1u)); }
// one zig-zag varint of one or two bytes → float; returns bytes used, 0 = longer (caller bails out)
__device__ __forceinline__ int avs_small_int(uint32_t w, float* v) {
  const uint32_t b0 = w & 0xFFu, b1 = (w >> 8) & 0xFFu;
  const bool one = b0 < 0x80u;
  *v = av_zigzag_f32(one ? b0 : ((b0 & 0x7Fu) | (b1 << 7)));
  return one ? 1 : (b1 < 0x80u ? 2 : 0);
}
template <int COUNT, int OUT0, int XS>
__device__ __forceinline__ int avs_ints(const uint8_t* __restrict__ base, int p, float* __restrict__ xrow) {
  if (COUNT == 4) {
    const uint32_t w4 = lds_u32_unaligned(base, p);
    if ((w4 & 0x80808080u) == 0) {  // four one-byte values (tyre pressures)
#pragma unroll
      for (int q = 0; q < 4; ++q) xrow[(OUT0 + q) * XS] = av_zigzag_f32((w4 >> (8 * q)) & 0xFFu);
      return p + 4;
    }
  }
#pragma unroll
  for (int j = 0; j < COUNT; ++j) {
    float v;
    const int used = avs_small_int(lds_u32_unaligned(base, p), &v);
    if (used == 0) return -1;
    xrow[(OUT0 + j) * XS] = v;
    p += used;
  }
  return p;
}
template <int XS, unsigned ZM = 0u>
__device__ __forceinline__ int avro_walk_schema_a(const uint8_t* __restrict__ base, int p, int e, int skip, float* __restrict__ xrow) {
  p += skip;
  p = avs_floats<9, 0, XS, ZM>(base, p, xrow);
  p = avs_ints<4, 9, XS>(base, p, xrow);
  if (p < 0) return -1;
  p = avs_floats<4, 13, XS, ZM>(base, p, xrow);
  p = avs_ints<1, 17, XS>(base, p, xrow);
  if (p < 0) return -1;
  return p > e ? B2_REC_TRUNCATED : B2_REC_OK;
}

// ["null", double]: tag 0x00 = null (→ 0), 0x02 = value; anything else bails out
template <int COUNT, int OUT0, int XS, unsigned ZM = 0u>
__device__ __forceinline__ int avs_nullable_doubles(const uint8_t* __restrict__ base, int p, float* __restrict__ xrow) {
  bool ok = true;
#pragma unroll
  for (int j = 0; j < COUNT; ++j) {
    if ((ZM >> (OUT0 + j)) & 1u) {  // a column the normalisation overwrites with zero: only the union tag is looked at
      const uint32_t tag = base[p];
      ok = ok && (tag == 0u || tag == 2u);
      p += tag == 2u ? 9 : 1;
      continue;
    }
    // tag + 8 value bytes always sit inside three aligned words
    const uint32_t* wp = reinterpret_cast<const uint32_t*>(base + (p & ~3));
    const uint32_t w0 = wp[0], w1 = wp[1], w2 = wp[2];
    const int sh = (p & 3) * 8;
    const uint32_t tag = (w0 >> sh) & 0xFFu;
    const unsigned long long bits =
        (unsigned long long)__funnelshift_rc(w0, w1, sh + 8) | ((unsigned long long)__funnelshift_rc(w1, w2, sh + 8) << 32);
    const float v = __double2float_rn(__longlong_as_double((long long)bits));
    ok = ok && (tag == 0u || tag == 2u);
    xrow[(OUT0 + j) * XS] = tag == 2u ? v : 0.0f;
    p += tag == 2u ? 9 : 1;
  }
  return ok ? p : -1;
}
template <int COUNT, int OUT0, int XS>
__device__ __forceinline__ int avs_nullable_ints(const uint8_t* __restrict__ base, int p, float* __restrict__ xrow) {
  bool ok = true;
#pragma unroll
  for (int j = 0; j < COUNT; ++j) {
    const uint32_t w = lds_u32_unaligned(base, p);
    const uint32_t tag = w & 0xFFu;
    float v;
    const int used = avs_small_int(w >> 8, &v);  // bytes 1..3 of the window: a one- or two-byte varint fits
    ok = ok && (tag == 0u || (tag == 2u && used != 0));
    xrow[(OUT0 + j) * XS] = tag == 2u ? v : 0.0f;
    p += tag == 2u ? 1 + used : 1;
  }
  return ok ? p : -1;
}
template <int XS, unsigned ZM = 0u>
__device__ __forceinline__ int avro_walk_schema_b(const uint8_t* __restrict__ base, int p, int e, int skip, float* __restrict__ xrow,
                                                  uint8_t* label) {
  p += skip;
  p = avs_nullable_doubles<9, 0, XS, ZM>(base, p, xrow);
  if (p < 0) return -1;
  p = avs_nullable_ints<4, 9, XS>(base, p, xrow);
  if (p < 0) return -1;
  p = avs_nullable_doubles<4, 13, XS>(base, p, xrow);
  if (p < 0) return -1;
  p = avs_nullable_ints<1, 17, XS>(base, p, xrow);
  if (p < 0) return -1;
  // ["null", string] label: tag, one-byte length, bytes
  const uint32_t w = lds_u32_unaligned(base, p);
  const uint32_t tag = w & 0xFFu, lz = (w >> 8) & 0xFFu;
  uint8_t lab = 2;
  if (tag == 2u) {
    if (lz >= 0x80u || (lz & 1u)) return -1;  // long or negative length: generic walk
    const int L = (int)(lz >> 1);
    if (p + 2 + L > e) return B2_REC_TRUNCATED;
    const uint32_t c4 = lds_u32_unaligned(base, p + 2);
    if (L == 5 && c4 == 0x736C6166u && base[p + 6] == 'e') lab = 0;  // "fals" + "e"
    else if (L == 4 && c4 == 0x65757274u) lab = 1;                   // "true"
    else lab = L == 0 ? 2 : 3;
    p += 2 + L;
  } else if (tag == 0u) {
    p += 1;
  } else {
    return -1;
  }
  if (p > e) return B2_REC_TRUNCATED;
  *label = lab;
  return B2_REC_OK;
}

// ---- Avro, float64 order (cardata-v3.py) ------------------------------------------------------------------------------------
// cardata-v3.py decodes every numeric field as float64 (:53-74; ints via tf.cast(..., tf.float64), :98-103) and runs normalize_fn
// on float64 (:78-168) — (v - lo) / (hi - lo) * 2.0 - 1.0 with every step rounded to binary64 — before Keras rounds the model
// input to float32.  This walk keeps a field's value in binary64 until that last rounding: plain field-by-field loop over the
// schema program (no run specialisation: the mode is opt-in), one IEEE double division per scaled feature.
template <int D>
__device__ __forceinline__ float norm64_store(const DecodeArgs<D>& a, int col, double v) {
  const int kind = a.nm.kind[col];
  if (kind == B2_NORM_SCALE) v = __fma_rn(__ddiv_rn(__dsub_rn(v, a.n64_lo[col]), a.n64_rg[col]), 2.0, -1.0);  // 2q is exact: one rounding, like (q * 2.0) - 1.0
  else if (kind == B2_NORM_ZERO) v = 0.0;
  return __double2float_rn(v);
}
template <int D>
__device__ __noinline__ int avro_decode_record_f64(const uint8_t* __restrict__ base, int p, int e, const DecodeArgs<D>& a, float* __restrict__ xrow,
                                                   int xs, uint8_t* label) {
  const AvroP& av = a.avro;
  uint8_t lab_out = 2;
  p += av.skip_prefix;
  if (p > e) return B2_REC_TRUNCATED;
  for (int i = 0; i < av.n_fields; ++i) {
    const int type = av.f[i].type, nullable = av.f[i].nullable, out = av.f[i].out;
    bool present = true, wide;
    int v32 = 0;
    long long v64 = 0;
    double val = 0.0;  // a null numeric field decodes as 0 ([upstream] tfio default) and is then scaled like any other value
    uint8_t lab = 2;
    if (nullable) {
      p = av_varint(base, p, e, 10, &v32, &v64, &wide);
      if (p < 0) return B2_REC_TRUNCATED;
      if (wide || (v32 != 0 && v32 != 1)) return B2_REC_BAD_BRANCH;
      present = v32 != (nullable == 1 ? 0 : 1);
    }
    if (present) {
      if (type == AV_FLOAT) {
        if (p + 4 > e) return B2_REC_TRUNCATED;
        val = (double)__uint_as_float(lds_u32_unaligned(base, p));
        p += 4;
      } else if (type == AV_DOUBLE) {
        if (p + 8 > e) return B2_REC_TRUNCATED;
        const unsigned long long bits = (unsigned long long)lds_u32_unaligned(base, p) | ((unsigned long long)lds_u32_unaligned(base, p + 4) << 32);
        val = __longlong_as_double((long long)bits);
        p += 8;
      } else if (type == AV_INT || type == AV_LONG) {
        p = av_varint(base, p, e, type == AV_INT ? 5 : 10, &v32, &v64, &wide);
        if (p < 0) return B2_REC_TRUNCATED;
        val = !wide ? (double)v32 : (type == AV_INT ? (double)(int)v64 : __ll2double_rn(v64));
      } else if (type == AV_BOOLEAN) {
        if (p + 1 > e) return B2_REC_TRUNCATED;
        val = base[p] ? 1.0 : 0.0;
        p += 1;
      } else if (type == AV_STRING || type == AV_BYTES) {
        p = av_varint(base, p, e, 10, &v32, &v64, &wide);
        if (p < 0) return B2_REC_TRUNCATED;
        const long long L = wide ? v64 : (long long)v32;
        if (L < 0 || (long long)p + L > e) return B2_REC_TRUNCATED;
        const uint32_t c4 = lds_u32_unaligned(base, p);
        if (L == 5 && c4 == 0x736C6166u && base[p + 4] == 'e') lab = 0;
        else if (L == 4 && c4 == 0x65757274u) lab = 1;
        else lab = L == 0 ? 2 : 3;
        p += (int)L;
      }
    }
    if (out >= 0) xrow[out * xs] = norm64_store<D>(a, out, val);
    else if (out == -2) lab_out = lab;
  }
  if (p > e) return B2_REC_TRUNCATED;
  *label = lab_out;
  return B2_REC_OK;
}

// ---- the kernel --------------------------------------------------------------------------------------------------
// MODE 0: decode only (runtime n_out);  MODE 1: fused decode + score for the D-14-7-7-D autoencoder.
// Every WARP is an independent worker: it owns a byte stage, a float tile and an mbarrier in shared memory and walks chunks of
// 32·R consecutive records (chunk = global warp index, + total warps, …).  Lane i owns records i, i+32, … of the chunk from their
// offsets to their results; there is no block-wide barrier anywhere, so a warp that hits slow records, a wrap of the line or
// the end of its range never holds up the other warps of the SM.  Per item (the records of the chunk whose bytes fit the stage —
// normally all of them): wait for the bulk copy, decode, then — the stage being free again — plan the next item and issue its
// copy so that it flies under this item's score phase, then score and store.
// R = 2 (fused Avro): a lane decodes two records into an element-interleaved row pair and scores both in one pass over the
// weights (ae_forward_pair): a quarter of the weight fetches per record, half the planning work per record.
// TC (fused mode, R = 1, CTAs of exactly 4 warps): the score phase runs on the tensor cores (ae_tc.cuh) — the CTA's four warps
// still fetch, plan and decode independently, but score their 4 x 32 records together as one 128-row tile, so the item loop runs
// in lock-step across the CTA (a warp that has run out of chunks contributes empty rows until all four are done).
template <int D, int MODE, bool FT, int FMT, int R, bool TC = false, bool WM = false, unsigned ZM = 0u, bool N64 = false>
__device__ __forceinline__ void decode_body(const DecodeArgs<D>& a) {
  static_assert(!N64 || (MODE == 1 && R == 1 && !TC && !WM && ZM == 0u && FMT == FMT_AVRO), "float64 order: fused Avro, FFMA2 score phase");
  static_assert(ZM == 0u || (MODE == 1 && R == 1 && !TC && !WM), "zero-set specialisation: fused mode, FFMA2 score phase");
  static_assert(R == 1 || (R == 2 && MODE == 1), "two records per lane only in the fused mode");
  static_assert(!TC || (MODE == 1 && R == 1), "tensor-core scoring: fused mode, one record per lane");
  static_assert(!WM || (MODE == 1 && R == 1 && !TC && FMT != FMT_JSON), "warp-MMA scoring: fused mode, one record per lane, CSV / Avro");
  constexpr int XS = WM ? 2 : R;  // distance between consecutive elements of a record in the tile (WM: rows interleaved in pairs)
  extern __shared__ __align__(128) uint8_t smem[];
  constexpr unsigned FULL = 0xffffffffu;
  constexpr int CH = 32 * R;  // records per chunk
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  // floats of the tile per lane: MODE 0 one row of n_out; MODE 1 R = 1 one row of D; R = 2 D + 1 interleaved pairs (the odd
  // pair count keeps the lanes' 64-bit accesses on distinct banks)
  const int xt_stride = MODE == 1 ? (R == 2 ? 2 * (D + 1) : D) : a.n_out;
  // per warp: [16-byte apron | stage (stage_cap) | 16 spare | float tile | mbarrier]
  uint8_t* wbase = smem + (size_t)warp * a.warp_bytes;
  uint8_t* stage = wbase + 16;  // apron in front: the token-tail load may start up to 3 bytes before the first token
  float* xt = reinterpret_cast<float*>(wbase + 32 + a.stage_cap);
  uint64_t* full = reinterpret_cast<uint64_t*>(wbase + a.warp_bytes - 16);
  if (lane == 0) {
    mbar_init(full, 1);
    mbar_fence_init();
  }
  __syncwarp();
  uint32_t phase = 0;
  unsigned flagged = 0, n_ok = 0, n_err = 0;
  unsigned long long n_bytes = 0;
  __shared__ uint32_t s_tmem;
  __shared__ __align__(8) uint64_t mma_bar;
  TcCtx tc;
  if (TC) tcs_setup<D>(tc, a.panels, reinterpret_cast<float*>(smem + (size_t)(blockDim.x >> 5) * a.warp_bytes), &s_tmem, &mma_bar, tid);
  // WM: the CTA's table of weight fragments behind the warps' regions (the one block-wide barrier of the kernel)
  const uint8_t* const wm_tab = smem + (size_t)(blockDim.x >> 5) * a.warp_bytes;
  if (WM) {
    mma_fill_table<D>(a.w, smem + (size_t)(blockDim.x >> 5) * a.warp_bytes, tid, blockDim.x);
    __syncthreads();
  }
  const bool list_mode = a.list_mode != 0;
  const long long n_chunks = list_mode ? (long long)min(a.over_list[0], B2_OVERSIZE_CAP) : (a.n + CH - 1) / CH;
  const int warps_per_cta = blockDim.x >> 5;  // the launch picks 8, 4 or 2 warps per CTA: whatever packs most warps into an SM
  const long long n_workers = (long long)gridDim.x * warps_per_cta;

  // a single record larger than the stage (lane 0): mark it refused and hand it to the retry launch, whose stage holds
  // STAGE_CAP_MAX bytes; what does not fit even there stays refused
  auto oversize = [&](long long row) {
    if (list_mode) return;  // already marked and counted by the main launch
    if (a.status) a.status[row] = B2_REC_TOO_LONG;
    if (MODE == 1) {
      if (a.mse) a.mse[row] = __int_as_float(0x7FC00000);
      if (a.flag) a.flag[row] = 0;
    } else {
      for (int j = 0; j < a.n_out; ++j) a.out[row * a.n_out + j] = __int_as_float(0x7FC00000);
    }
    if (a.label) a.label[row] = 2;
    ++n_err;
    if (a.over_list) {
      const int slot = atomicAdd(a.over_list, 1);
      if (slot < B2_OVERSIZE_CAP) a.over_list[1 + slot] = (int32_t)row;
    }
  };

  // cursor over this warp's chunks: lane i holds the byte ranges [ob[s], oe[s]) of records i + 32 s of the current chunk
  long long chunk = (long long)blockIdx.x * warps_per_cta + warp - n_workers, row0 = 0;
  int rows = 0, sub = 0, ob[R], oe[R];
#pragma unroll
  for (int s = 0; s < R; ++s) ob[s] = oe[s] = 0;
  // record index q (0 .. CH-1) of the chunk → value held by lane q & 31 in slot q >> 5
  auto pick = [&](const int (&v)[R], int q) -> int {
    int r = __shfl_sync(FULL, v[0], q & 31);
    if (R == 2) {
      const int r1 = __shfl_sync(FULL, v[R - 1], q & 31);
      if (q >= 32) r = r1;
    }
    return r;
  };
  // plans the next item (records [sub, sub+cnt) of the cursor's chunk), issues its bulk copy and advances the cursor;
  // it_b / it_e receive the lane's stage offsets of its records.  (Fetching a chunk's offsets one chunk ahead was tried: the
  // load then stays in flight across the decode phase, whose shared-memory loads end up waiting on the same scoreboard:
  // 2.71e10 -> 2.52e10 records/s on Avro-A.)
  auto next_item = [&](long long& it_row0, int& it_sub, int& it_cnt, int (&it_b)[R], int (&it_e)[R]) -> bool {
    for (;;) {
      if (sub >= rows) {
        chunk += n_workers;
        if (chunk >= n_chunks) return false;
        row0 = list_mode ? (long long)a.over_list[1 + chunk] : chunk * CH;
        rows = list_mode ? 1 : (int)min((long long)CH, a.n - row0);
        sub = 0;
        if (a.ends) {
          // ranges with gaps between them (the values inside a Kafka segment): an item's bulk copy simply brings the bytes
          // in between along
#pragma unroll
          for (int s = 0; s < R; ++s) {
            const int q = min(lane + 32 * s, rows - 1);
            ob[s] = a.offs[row0 + q];
            oe[s] = a.ends[row0 + q];
            if (lane + 32 * s >= rows) ob[s] = oe[s];  // a record that does not exist: empty, at the end of the chunk
          }
        } else {
#pragma unroll
          for (int s = 0; s < R; ++s) ob[s] = a.offs[row0 + min(lane + 32 * s, rows)];
#pragma unroll
          for (int s = 0; s < R; ++s) {
            oe[s] = __shfl_down_sync(FULL, ob[s], 1);  // lane 31 keeps its own value here
            if (s + 1 < R) {
              const int wrap = __shfl_sync(FULL, ob[s + 1 < R ? s + 1 : s], 0);
              if (lane == 31) oe[s] = wrap;
            }
          }
          // end of the chunk's last possible record (a record that does not exist has ob == oe == end of the chunk)
          if (lane == 31) oe[R - 1] = a.offs[row0 + rows];
        }
      }
      const int b_abs = pick(ob, sub);
      const uintptr_t p0 = reinterpret_cast<uintptr_t>(a.bytes) + (uintptr_t)(long long)b_abs;
      const uintptr_t a0 = p0 & ~(uintptr_t)15;
      const int lead = (int)(p0 - a0);
      int cnt = 0;
#pragma unroll
      for (int s = 0; s < R; ++s) {
        const int q = lane + 32 * s;
        const bool fit = q >= sub && q < rows && (long long)oe[s] - b_abs + lead <= (long long)a.stage_cap;
        cnt += __popc(__ballot_sync(FULL, fit));  // record ends are monotone: the records that fit are sub .. sub+cnt-1
      }
      if (cnt == 0) {
        if (lane == 0) oversize(row0 + sub);
        sub += 1;
        continue;
      }
      const int e_abs = pick(oe, sub + cnt - 1);
      if (lane == 0) {
        const uint32_t nbytes = (uint32_t)(((e_abs - b_abs + lead) + 15) & ~15);
        mbar_expect_tx(full, nbytes);
        bulk_g2s(stage, reinterpret_cast<const void*>(a0), nbytes, full);
        n_bytes += (unsigned long long)(e_abs - b_abs);
      }
      const int rel = lead - b_abs;  // stage offset of absolute byte x is x + rel
      it_row0 = row0;
      it_sub = sub;
      it_cnt = cnt;
#pragma unroll
      for (int s = 0; s < R; ++s) {
        it_b[s] = ob[s] + rel;
        it_e[s] = oe[s] + rel;
      }
      sub += cnt;
      return true;
    }
  };

  long long c_row0 = 0, n_row0 = 0;
  int c_sub = 0, c_cnt = 0, c_b[R], c_e[R], n_sub = 0, n_cnt = 0, n_b[R], n_e[R];
  bool have = next_item(c_row0, c_sub, c_cnt, c_b, c_e);
  float* const xlane = WM ? xt + mma_row_offset<D>(lane) : xt + (size_t)lane * xt_stride;
  bool cur = have;  // TC: this warp holds a current item (the loop itself runs while ANY warp of the CTA does)
  while (TC ? __syncthreads_or(cur) != 0 : have) {
    if (!TC || cur) {
      mbar_wait(full, phase);
      phase ^= 1u;
    }
    int st[R];
    uint8_t lab[R];
    bool active[R];
#pragma unroll
    for (int s = 0; s < R; ++s) {
      const int q = lane + 32 * s;
      active[s] = (!TC || cur) && q >= c_sub && q < c_sub + c_cnt;
      st[s] = B2_REC_OK;
      lab[s] = 2;
    }
    // one record of the item: slot s of this lane, stage bytes [cb, ce), absolute offset ab of its first byte
    auto decode_one = [&](int s, int cb, int ce, int ab, int& st_s, uint8_t& lab_s) {
      const int q = lane + 32 * s;
      float* xrow = xlane + s;  // element j of this record sits at xrow[j * XS]
      const long long gr = c_row0 + q;
      if (FMT == FMT_CSV) {
        int32_t* sp = (MODE == 0 && a.span) ? a.span + gr * a.csv.n_str * 2 : nullptr;
        const int rel = cb - ab;  // stage offset minus absolute offset (the cursor still sits on this item's chunk)
        st_s = csv_walk<XS>(stage, cb, ce, a.csv, xrow, sp, rel);
        if (st_s < 0) st_s = csv_record_careful(stage, cb, ce, a.csv, xrow, XS, sp, rel);
      } else if (FMT == FMT_JSON) {
        for (int j = 0; j < xt_stride; ++j) xrow[j] = 0.0f;  // members that are absent or null
        st_s = b2_json_record_auto(stage, cb, ce, a.json, xrow, &lab_s);
      } else if (N64) {
        // columns no field routes to hold a decoded 0.0, scaled like any other value
        if (!a.avro.all_routed)
          for (int j = 0; j < D; ++j) xrow[j * XS] = norm64_store<D>(a, j, 0.0);
        st_s = avro_decode_record_f64<D>(stage, cb, ce, a, xrow, XS, &lab_s);
      } else {
        if (!a.avro.all_routed)
          for (int j = 0; j < (MODE == 1 ? D : a.n_out); ++j) xrow[j * XS] = 0.0f;  // columns no field routes to
        st_s = -1;
        if (a.avro.known == 1) st_s = avro_walk_schema_a<XS, ZM>(stage, cb, ce, a.avro.skip_prefix, xrow);
        else if (a.avro.known == 2) st_s = avro_walk_schema_b<XS, ZM>(stage, cb, ce, a.avro.skip_prefix, xrow, &lab_s);
        if (st_s < 0) {
          lab_s = 2;
          st_s = avro_decode_record(stage, cb, ce, a.avro, xrow, XS, &lab_s);
        }
      }
    };
    if (R == 1) {
      if (active[0]) decode_one(0, c_b[0], c_e[0], ob[0], st[0], lab[0]);
    } else {
      // the two records run through ONE copy of the decoder (a rolled loop): the straight-line code of the item loop has to stay
      // inside the 32 KB instruction cache, and two inlined decoders + the pair forward pass do not
#pragma unroll 1
      for (int s = 0; s < 2; ++s) {
        const bool act = s ? active[R - 1] : active[0];
        int st_s = B2_REC_OK;
        uint8_t lab_s = 2;
        if (act) decode_one(s, s ? c_b[R - 1] : c_b[0], s ? c_e[R - 1] : c_e[0], s ? ob[R - 1] : ob[0], st_s, lab_s);
        if (s) st[R - 1] = st_s, lab[R - 1] = lab_s;
        else st[0] = st_s, lab[0] = lab_s;
      }
    }
    __syncwarp();  // every lane is done with the stage
    if (!TC || cur) have = next_item(n_row0, n_sub, n_cnt, n_b, n_e);
    if (MODE == 1) {
      float mm[R];
      if (TC) {
        // the lane's row → registers (normalised), then the 128-row tile of the whole CTA through the tensor cores
        float x[D];
#pragma unroll
        for (int i = 0; i < D; i += 2) {
          float2 v = active[0] ? *reinterpret_cast<const float2*>(xlane + i) : make_float2(0.f, 0.f);
          if (a.has_norm) v = norm_pair_fast<D>(v, a.nm, i);
          x[i] = v.x;
          x[i + 1] = v.y;
        }
        mm[0] = tcs_score_row<D, FT, true>(tc, a.w, x, tid);
      } else if (WM) {
        // every lane normalises its own row in place (uniform constants), then the warp's 32 rows run through the MMA chain
        if (a.has_norm && active[0]) {
#pragma unroll
          for (int i = 0; i < D; i += 2) {
            const float2 v = norm_pair_fast<D>(make_float2(xlane[2 * i], xlane[2 * i + 2]), a.nm, i);
            xlane[2 * i] = v.x;
            xlane[2 * i + 2] = v.y;
          }
        }
        __syncwarp();
        mm[0] = mma_score_warp<D, FT>(xt, wm_tab, lane);
      } else if (R == 2) {
        if (active[0] || active[R - 1]) {
          float2* rowp = reinterpret_cast<float2*>(xlane);
          float m2[2] = {0.f, 0.f};
          if (a.has_norm) ae_forward_pair<D, true, FT>(a.w, a.nm, rowp, m2);
          else ae_forward_pair<D, false, FT>(a.w, a.nm, rowp, m2);
          mm[0] = m2[0];
          mm[R - 1] = m2[1];
        }
      } else if (active[0]) {
        float* const rowp[1] = {xlane};
        float m1[1] = {0.f};
        if (!N64 && (ZM != 0u || a.has_norm)) ae_forward_rows<D, 1, true, false, FT, ZM>(a.w, a.nm, rowp, m1);  // (a ZM kernel is only launched with a table)
        else ae_forward_rows<D, 1, false, false, FT>(a.w, a.nm, rowp, m1);  // (N64: the decode has normalised the row already)
        mm[0] = m1[0];
      }
#pragma unroll
      for (int s = 0; s < R; ++s) {
        if (active[s]) {
          const long long gr = c_row0 + lane + 32 * s;
          float m = mm[s];
          if (st[s] != B2_REC_OK) m = __int_as_float(0x7FC00000);
          const bool f = m > a.threshold;
          if (a.mse) a.mse[gr] = m;
          if (a.flag) a.flag[gr] = f ? 1 : 0;
          if (a.status) a.status[gr] = st[s];
          if (a.label) a.label[gr] = FMT != FMT_CSV && st[s] == B2_REC_OK ? lab[s] : 2;
          flagged += f;
          n_ok += st[s] == B2_REC_OK;
          n_err += st[s] != B2_REC_OK;
        }
      }
      if (R == 2 || WM) __syncwarp();  // the tile is rewritten by the next item's decode (WM: the lanes have read each other's rows)
    } else {
      // coalesced write-out of the item's rows (they are contiguous in xt); a bad record's row becomes NaN
      const int total = c_cnt * xt_stride;
      const float* src = xt + (size_t)c_sub * xt_stride;
      float* dst = a.out + (c_row0 + c_sub) * xt_stride;
      for (int i0 = 0; i0 < total; i0 += 32) {
        const int i = i0 + lane;
        const int r = min(i, total - 1) / xt_stride;
        const int st_r = __shfl_sync(FULL, st[0], c_sub + r);
        if (i < total) dst[i] = st_r == B2_REC_OK ? src[i] : __int_as_float(0x7FC00000);
      }
      if (active[0]) {
        const long long gr = c_row0 + lane;
        if (a.status) a.status[gr] = st[0];
        if (a.label) a.label[gr] = FMT != FMT_CSV && st[0] == B2_REC_OK ? lab[0] : 2;
        n_ok += st[0] == B2_REC_OK;
        n_err += st[0] != B2_REC_OK;
      }
      __syncwarp();  // xt is rewritten by the next item's decode
    }
    c_row0 = n_row0, c_sub = n_sub, c_cnt = n_cnt;
#pragma unroll
    for (int s = 0; s < R; ++s) c_b[s] = n_b[s], c_e[s] = n_e[s];
    if (TC) cur = cur && have;
  }
  if (TC) tcs_teardown<D>(tc, tid);
  if (a.counters) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      flagged += __shfl_xor_sync(FULL, flagged, o);
      n_ok += __shfl_xor_sync(FULL, n_ok, o);
      n_err += __shfl_xor_sync(FULL, n_err, o);
    }
    if (lane == 0) {
      if (n_ok) atomicAdd(&a.counters[0], (unsigned long long)n_ok);
      if (flagged) atomicAdd(&a.counters[1], (unsigned long long)flagged);
      // retry launch: every record it was handed had been counted as an error; the ones that decoded now are taken back
      // (a record that fails again, with whatever status, stays counted once)
      if (list_mode) atomicAdd(&a.counters[2], 0ull - (unsigned long long)n_ok);
      else if (n_err) atomicAdd(&a.counters[2], (unsigned long long)n_err);
      if (n_bytes) atomicAdd(&a.counters[3], n_bytes);
    }
  }
  if (list_mode) {  // one CTA: everybody has read the count long ago; leave the list empty for the stream's next launch
    __syncthreads();
    if (tid == 0) a.over_list[0] = 0;
  }
}

// One record per lane: CTAs of up to 8 warps, registers left to ptxas (64: four CTAs per SM).  Two records per lane: CTAs of up
// to 4 warps, MINB of them per SM bound the registers (6 → 85, 5 → 102).
template <int D, int MODE, bool FT, int FMT, int R, int MINB>
struct DecodeKernel;
template <int D, int MODE, bool FT, int FMT>
__global__ void __launch_bounds__(DEC_NT) decode_kernel(const __grid_constant__ DecodeArgs<D> a) {
  decode_body<D, MODE, FT, FMT, 1>(a);
}
template <int D, int MODE, bool FT, int FMT, int MINB>
__global__ void __launch_bounds__(DEC_NT_R2, MINB) decode_kernel_r2(const __grid_constant__ DecodeArgs<D> a) {
  decode_body<D, MODE, FT, FMT, (MODE == 1 ? 2 : 1)>(a);
}
// tensor-core score phase: CTAs of exactly 4 warps = one 128-row TMEM tile; 512 / TMEM columns of them fit an SM (registers: 5)
template <int D, bool FT, int FMT>
__global__ void __launch_bounds__(TCS_ROWS, 512 / TcShape<D>::TMEM_COLS > 5 ? 5 : 512 / TcShape<D>::TMEM_COLS) decode_kernel_tc(const __grid_constant__ DecodeArgs<D> a) {
  decode_body<D, 1, FT, FMT, 1, true>(a);
}
// the instantiation with the reference table's zero set compiled in (REF_ZERO_MASK, ae_math.cuh), launched when the caller's table
// has exactly that zero set
template <int D, bool FT, int FMT>
__global__ void __launch_bounds__(DEC_NT, FMT == FMT_AVRO ? 4 : 3) decode_kernel_zs(const __grid_constant__ DecodeArgs<D> a) {
  decode_body<D, 1, FT, FMT, 1, false, false, REF_ZERO_MASK>(a);
}
// float64 normalisation order (cardata-v3.py): the walk that keeps values in binary64 until the model's input cast
template <int D, bool FT>
__global__ void __launch_bounds__(DEC_NT, 3) decode_kernel_n64(const __grid_constant__ DecodeArgs<D> a) {
  decode_body<D, 1, FT, FMT_AVRO, 1, false, false, 0u, true>(a);
}
// warp-MMA score phase (ae_mma.cuh): CTA shape as for decode_kernel; registers bounded for three CTAs of 8 warps per SM
template <int D, bool FT, int FMT>
__global__ void __launch_bounds__(DEC_NT, 3) decode_kernel_wm(const __grid_constant__ DecodeArgs<D> a) {
  decode_body<D, 1, FT, FMT, 1, false, true>(a);
}
template <int D, int MODE, bool FT, int FMT, int MINB>
struct DecodeKernel<D, MODE, FT, FMT, 1, MINB> {
  static auto get() { return decode_kernel<D, MODE, FT, FMT>; }
};
template <int D, int MODE, bool FT, int FMT, int MINB>
struct DecodeKernel<D, MODE, FT, FMT, 2, MINB> {
  static auto get() { return decode_kernel_r2<D, MODE, FT, FMT, MINB>; }
};

// ---- host side ----------------------------------------------------------------------------------------------------
static int fill_csv(const b2_csv_spec* s, CsvP& c) {
  B2_REQUIRE(s->n_fields >= 1 && s->n_fields <= 63, "csv: n_fields=%d not in 1..63", s->n_fields);
  memset(&c, 0, sizeof(c));
  c.n_fields = s->n_fields;
  c.delim = s->delim ? s->delim : ',';
  c.use_quote = s->use_quote_delim;
  for (int i = 0; i < s->n_fields; ++i) {
    B2_REQUIRE(s->kind[i] >= 0 && s->kind[i] <= 2, "csv: kind[%d]=%d unknown", i, s->kind[i]);
    c.kind[i] = (uint8_t)s->kind[i];
    c.has_default[i] = s->has_default[i] ? 1 : 0;
    c.dflt[i] = s->dflt[i];
    if (s->kind[i] == 1) c.col[i] = (uint8_t)c.n_float++;
    else if (s->kind[i] == 2) c.col[i] = (uint8_t)c.n_str++;
  }
  B2_REQUIRE(c.n_float >= 1, "csv: no float field selected");
  return B2_OK;
}

static int fill_avro(const b2_ctx* ctx, const b2_avro_prog* p, int skip_prefix, AvroP& a) {
  B2_REQUIRE(p->n_fields >= 1 && p->n_fields <= B2_MAX_FIELDS, "avro: n_fields=%d out of range", p->n_fields);
  B2_REQUIRE(p->n_out >= 1 && p->n_out <= B2_MAX_WIDTH, "avro: n_out=%d out of range", p->n_out);
  B2_REQUIRE(skip_prefix >= 0 && skip_prefix <= 64, "avro: skip_prefix=%d out of range", skip_prefix);
  memset(&a, 0, sizeof(a));
  a.n_fields = p->n_fields;
  a.n_out = p->n_out;
  a.skip_prefix = skip_prefix;
  for (int i = 0; i < p->n_fields; ++i) {
    B2_REQUIRE(p->f[i].type >= AV_INT && p->f[i].type <= AV_NULL, "avro: field %d has unknown type %d", i, p->f[i].type);
    B2_REQUIRE(p->f[i].out < p->n_out, "avro: field %d routed to column %d >= n_out", i, p->f[i].out);
    a.f[i] = p->f[i];
  }
  // runs of consecutive fields with the same (type, nullable) and consecutive output columns (or all skipped)
  for (int i = 0; i < p->n_fields; ++i) {
    const b2_avro_field& f = p->f[i];
    AvroRun* last = a.n_runs ? &a.run[a.n_runs - 1] : nullptr;
    const bool extend = last && last->type == f.type && last->nullable == f.nullable && last->count < 100 &&
                        ((last->out0 >= 0 && f.out == last->out0 + last->count) || (last->out0 == -1 && f.out == -1));
    if (extend) ++last->count;
    else a.run[a.n_runs++] = AvroRun{f.type, f.nullable, 1, f.out};
  }
  unsigned long long routed = 0;
  for (int i = 0; i < p->n_fields; ++i)
    if (p->f[i].out >= 0) routed |= 1ull << p->f[i].out;
  a.all_routed = routed == (p->n_out >= 64 ? ~0ull : ((1ull << p->n_out) - 1));
  // the reference's own schemas (see avro_walk_schema_a / _b); B2_AVRO_GENERIC=1 keeps the generic walk (tests compare the two)
  {
    static const int8_t is_int[18] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 0, 0, 0, 0, 1};
    bool is_a = p->n_fields == 18 && p->n_out == 18, is_b = p->n_fields == 19 && p->n_out == 18;
    for (int i = 0; i < 18 && (is_a || is_b); ++i) {
      const b2_avro_field& f = p->f[i];
      is_a = is_a && f.out == i && f.nullable == 0 && f.type == (is_int[i] ? AV_INT : AV_FLOAT);
      is_b = is_b && f.out == i && f.nullable == 1 && f.type == (is_int[i] ? AV_INT : AV_DOUBLE);
    }
    is_b = is_b && p->f[18].type == AV_STRING && p->f[18].nullable == 1 && p->f[18].out == -2;
    a.known = ctx->knobs.avro_generic ? 0 : (is_a ? 1 : (is_b ? 2 : 0));
  }
  return B2_OK;
}

// Upper bound of an Avro record's size when the schema has no strings/bytes (else 0 = unbounded)
static int avro_max_record_bytes(const AvroP& a) {
  int n = a.skip_prefix;
  for (int i = 0; i < a.n_fields; ++i) {
    if (a.f[i].nullable) n += 1;
    switch (a.f[i].type) {
      case AV_FLOAT: n += 4; break;
      case AV_DOUBLE: n += 8; break;
      case AV_BOOLEAN: n += 1; break;
      case AV_INT: n += 5; break;
      case AV_LONG: n += 10; break;
      case AV_NULL: break;
      default: return 0;
    }
  }
  return n;
}

static int align16i(int v) { return (v + 15) & ~15; }

// Bytes one warp stages per item.  Avro without strings: 32 records of the schema's maximum size.  Otherwise an estimate of 32
// typical records (a chunk whose bytes do not fit is simply split into several items, at the price of idle lanes):
// `hint` = the caller's average record size if known, else ~10 bytes per CSV field / the fixed part + 16 bytes per Avro string.
static int choose_stage_cap(int fmt, const CsvP& csv, const AvroP& avro, const JsonP& json, int hint, int ch /* records per chunk */) {
  int per_rec;
  if (fmt == FMT_JSON) {
    per_rec = 2;
    for (int k = 0; k < json.n_keys; ++k) per_rec += json.len[k] + 14;  // "name":value,
  } else if (fmt == FMT_AVRO) {
    const int mx = avro_max_record_bytes(avro);
    if (mx > 0) return min(STAGE_CAP_MAX, align16i(mx * ch + 32));
    per_rec = avro.skip_prefix;
    for (int i = 0; i < avro.n_fields; ++i) {
      per_rec += avro.f[i].nullable ? 1 : 0;
      switch (avro.f[i].type) {
        case AV_FLOAT: per_rec += 4; break;
        case AV_DOUBLE: per_rec += 8; break;
        case AV_BOOLEAN: per_rec += 1; break;
        case AV_INT: per_rec += 3; break;
        case AV_LONG: per_rec += 5; break;
        case AV_STRING: case AV_BYTES: per_rec += 17; break;
        default: break;
      }
    }
  } else {
    per_rec = csv.n_fields * 10;
  }
  if (hint > 0) per_rec = hint;
  // a known average gets 6 % head-room (32 records rarely sum to more), an estimate none: the stage also sets how many CTAs fit
  return max(2048, min(STAGE_CAP_MAX, align16i(per_rec * ch + (hint > 0 ? per_rec * (ch / 16) : 0) + 48)));
}

// the kernel of a launch: VAR = 1 selects the warp-MMA score phase (fused mode, one record per lane)
template <int D, int MODE, bool FT, int FMT, int R, int MINB, int VAR>
static auto decode_kernel_of() {
  if constexpr (VAR == 1 && FMT != FMT_JSON) return decode_kernel_wm<D, FT, FMT>;
  else if constexpr (VAR == 2) return decode_kernel_zs<D, FT, FMT>;
  else if constexpr (VAR == 3 && FMT == FMT_AVRO) return decode_kernel_n64<D, FT>;
  else return DecodeKernel<D, MODE, FT, FMT, R, MINB>::get();
}

template <int D, int MODE, bool FT, int FMT, int R, int MINB, int VAR = 0>
static int launch_decode_r(b2_ctx* ctx, DecodeArgs<D>& args, cudaStream_t stream) {
  static_assert(VAR == 0 || (MODE == 1 && R == 1 && (VAR == 2 || FMT != FMT_JSON) && (VAR != 3 || FMT == FMT_AVRO)), "kernel variants: fused mode, one record per lane");
  const int xt_stride = MODE == 1 ? (R == 2 ? 2 * (D + 1) : D) : args.n_out;
  args.stage_cap = choose_stage_cap(args.fmt, args.csv, args.avro, args.json, ctx->record_bytes_hint, 32 * R);
  args.warp_bytes = 16 + args.stage_cap + 16 + align16i(32 * xt_stride * 4) + 16;
  auto kern = decode_kernel_of<D, MODE, FT, FMT, R, MINB, VAR>();
  constexpr size_t cta_extra = VAR == 1 ? (size_t)MmaShape<D>::BYTES : 0;  // per-CTA table of weight fragments
  // no block-wide synchronisation in the kernel, so the CTA size is free: take the one that packs most warps into an SM.
  // The choice depends only on (instantiation, device, bytes per warp): remembered, a launch is then one driver call.
  struct Cfg {
    int device, warp_bytes, threads, occ;
  };
  static thread_local Cfg cached = {-1, 0, 0, 0};
  int threads = R == 2 ? DEC_NT_R2 : DEC_NT, occ = 0;
  if (cached.device == ctx->device && cached.warp_bytes == args.warp_bytes) {
    threads = cached.threads;
    occ = cached.occ;
  } else {
    int best_warps = 0;
    B2_REQUIRE((size_t)args.warp_bytes * 2 + cta_extra <= (size_t)ctx->smem_optin, "decode: a warp's stage needs %d bytes of shared memory (max %d per CTA)",
               args.warp_bytes, (int)(ctx->smem_optin - cta_extra) / 2);
    B2_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, ctx->smem_optin));
    for (int w = (R == 2 ? DEC_NT_R2 / 32 : DEC_WARPS); w >= 2; w >>= 1) {
      const size_t sm = (size_t)args.warp_bytes * w + cta_extra;
      if (sm > (size_t)ctx->smem_optin) continue;
      int o = 0;
      B2_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o, kern, w * 32, sm));
      if (o * w > best_warps) best_warps = o * w, threads = w * 32, occ = o;
    }
    if (occ >= 1) cached = Cfg{ctx->device, args.warp_bytes, threads, occ};
  }
  const size_t smem = (size_t)args.warp_bytes * (threads / 32) + cta_extra;
  B2_REQUIRE(occ >= 1, "decode: kernel does not fit on an SM");
  const int wpc = threads / 32;
  const long long n_chunks = (args.n + 32 * R - 1) / (32 * R);
  long long grid = (long long)ctx->sm_count * occ;
  const long long need = (n_chunks + wpc - 1) / wpc;
  if (grid > need) grid = need;
  // records longer than the stage can only occur when the stage was sized from an average (CSV, JSON, Avro with strings) and
  // is below the maximum: those launches get a retry pass (below) for the few records they had to refuse
  const bool may_overflow = args.stage_cap < STAGE_CAP_MAX && !(args.fmt == FMT_AVRO && avro_max_record_bytes(args.avro) > 0);
  args.list_mode = 0;
  args.over_list = may_overflow ? b2_ctx_oversize_slot(ctx, stream) : nullptr;
  kern<<<(unsigned)grid, threads, smem, stream>>>(args);
  B2_CUDA(cudaGetLastError());
  if (may_overflow) {
    // same kernel, list mode: ONE CTA whose warps each take one listed record at a time into a STAGE_CAP_MAX-byte stage.  The
    // list is normally empty and the launch returns at once (a few microseconds on the stream); the hard limit of a record
    // is therefore STAGE_CAP_MAX bytes whatever the average was, and B2_OVERSIZE_CAP such records per launch.
    DecodeArgs<D> r = args;
    r.list_mode = 1;
    r.stage_cap = STAGE_CAP_MAX;
    r.warp_bytes = 16 + r.stage_cap + 16 + align16i(32 * xt_stride * 4) + 16;
    int rw = 4;
    while (rw > 1 && (size_t)r.warp_bytes * rw + cta_extra > (size_t)ctx->smem_optin) rw >>= 1;
    kern<<<1, rw * 32, (size_t)r.warp_bytes * rw + cta_extra, stream>>>(r);
    B2_CUDA(cudaGetLastError());
  }
  return B2_OK;
}

// Fused decode + score with the score phase on the tensor cores (decode_kernel_tc): CTAs of exactly 4 warps, as many per SM as
// TMEM columns, registers and shared memory allow; the oversize retry pass runs through the FFMA2 kernel.
template <int D, bool FT, int FMT>
static int launch_decode_tc(b2_ctx* ctx, DecodeArgs<D>& args, cudaStream_t stream) {
  args.stage_cap = choose_stage_cap(args.fmt, args.csv, args.avro, args.json, ctx->record_bytes_hint, 32);
  args.warp_bytes = 16 + args.stage_cap + 16 + align16i(32 * D * 4) + 16;
  auto kern = decode_kernel_tc<D, FT, FMT>;
  auto retry = decode_kernel<D, 1, FT, FMT>;
  const size_t smem = (size_t)args.warp_bytes * 4 + sizeof(float) * TcShape<D>::TOTAL;
  const int dyn_max = ctx->smem_optin - 1024;  // the kernel also has a few static words (TMEM base, MMA mbarrier)
  B2_REQUIRE(smem <= (size_t)dyn_max, "decode (tensor-core score): a CTA needs %zu bytes of shared memory", smem);
  static thread_local int regs[16] = {0};
  if (regs[ctx->device & 15] == 0) {
    B2_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn_max));
    B2_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    B2_CUDA(cudaFuncSetAttribute(retry, cudaFuncAttributeMaxDynamicSharedMemorySize, ctx->smem_optin));
    cudaFuncAttributes fa;
    B2_CUDA(cudaFuncGetAttributes(&fa, kern));
    regs[ctx->device & 15] = fa.numRegs > 0 ? fa.numRegs : 128;
  }
  int ctas = 512 / TcShape<D>::TMEM_COLS;
  ctas = min(ctas, 65536 / (((regs[ctx->device & 15] + 7) & ~7) * TCS_ROWS));
  ctas = min(ctas, (int)((size_t)(228 * 1024 - 1024) / (smem + 1024)));
  if (ctx->knobs.tc_ctas > 0) ctas = min(ctas, ctx->knobs.tc_ctas);
  B2_REQUIRE(ctas >= 1, "decode (tensor-core score): kernel does not fit on an SM");
  const long long n_chunks = (args.n + 31) / 32;
  long long grid = (long long)ctx->sm_count * ctas;
  const long long need = (n_chunks + 3) / 4;
  if (grid > need) grid = need;
  const bool may_overflow = args.stage_cap < STAGE_CAP_MAX && !(args.fmt == FMT_AVRO && avro_max_record_bytes(args.avro) > 0);
  args.list_mode = 0;
  args.over_list = may_overflow ? b2_ctx_oversize_slot(ctx, stream) : nullptr;
  kern<<<(unsigned)grid, TCS_ROWS, smem, stream>>>(args);
  B2_CUDA(cudaGetLastError());
  if (may_overflow) {
    DecodeArgs<D> r = args;
    r.list_mode = 1;
    r.stage_cap = STAGE_CAP_MAX;
    r.warp_bytes = 16 + r.stage_cap + 16 + align16i(32 * D * 4) + 16;
    int rw = 4;
    while (rw > 1 && (size_t)r.warp_bytes * rw > (size_t)ctx->smem_optin) rw >>= 1;
    retry<<<1, rw * 32, (size_t)r.warp_bytes * rw, stream>>>(r);
    B2_CUDA(cudaGetLastError());
  }
  return B2_OK;
}

// Records per lane and item.  Default: ONE.  Two records per lane (row-packed FFMA2, ae_forward_pair) were built to cut the weight
// fetches — 12.5 % fewer executed instructions on Avro-A — and measured SLOWER on B200 (2.3e10 vs 2.6e10 records/s): the loop body
// no longer fits the instruction cache (ncu: no_instruction becomes the top stall, issue slots 52 % busy at 19 warps/SM;
// profiles/round2_notes.md).  The variant stays selectable for fused Avro (results are bit-identical, tests compare them):
// context option "decode_r2" = 2 (registers capped at 85: 6 CTAs of 4 warps per SM) or 3 (102 registers, 5 CTAs); 0 / -1 = one.
template <int D, int MODE, bool FT, int FMT>
static int launch_decode_fmt(b2_ctx* ctx, DecodeArgs<D>& args, cudaStream_t stream) {
  if constexpr (MODE == 1 && FMT == FMT_AVRO) {  // the float64 normalisation order has one kernel; the tuning options below do not apply
    if (args.norm64) return launch_decode_r<D, MODE, FT, FMT, 1, 1, 3>(ctx, args, stream);
  }
  if (MODE == 1 && FT && args.panels != nullptr) {  // context option decode_tc: score phase on tcgen05
    if (MODE == 1) return launch_decode_tc<D, true, FMT>(ctx, args, stream);  // (only the fast-tanh form is instantiated)
  }
  if (MODE == 1 && FMT == FMT_AVRO) {
    constexpr int R2 = (MODE == 1 && FMT == FMT_AVRO) ? 2 : 1;  // keeps the other instantiations free of R = 2 code
    const int k = ctx->knobs.decode_r2;
    if (k == 3) return launch_decode_r<D, MODE, FT, FMT, R2, 5>(ctx, args, stream);
    if (k == 2) return launch_decode_r<D, MODE, FT, FMT, R2, 6>(ctx, args, stream);
  }
  if constexpr (MODE == 1 && FMT != FMT_JSON) {
    if (ctx->knobs.decode_mma != 0) return launch_decode_r<D, MODE, FT, FMT, 1, 1, 1>(ctx, args, stream);
  }
  if constexpr (MODE == 1 && D == 18 && FT) {
    // the table's zero set is the reference's and the layer-1 rows it skips are finite (+0 · w must be exactly nothing)
    if (ctx->knobs.zero_skip != 0 && args.has_norm && norm_zero_set<D>(args.nm, args.w) == REF_ZERO_MASK)
      return launch_decode_r<D, MODE, FT, FMT, 1, 1, 2>(ctx, args, stream);
  }
  return launch_decode_r<D, MODE, FT, FMT, 1, 1>(ctx, args, stream);
}

// the record format is a template parameter: each instantiation carries only its own decoder (fewer registers, smaller code)
template <int D, int MODE, bool FT>
static int launch_decode(b2_ctx* ctx, DecodeArgs<D>& args, cudaStream_t stream) {
  if (args.fmt == FMT_CSV) return launch_decode_fmt<D, MODE, FT, FMT_CSV>(ctx, args, stream);
  if (args.fmt == FMT_JSON) return launch_decode_fmt<D, MODE, FT, FMT_JSON>(ctx, args, stream);
  return launch_decode_fmt<D, MODE, FT, FMT_AVRO>(ctx, args, stream);
}

extern "C" int b2_decode_csv(b2_ctx* ctx, const uint8_t* bytes_dev, const int32_t* offs_dev, int64_t n, const b2_csv_spec* spec,
                             float* out_dev, int32_t* span_dev, int32_t* status_dev, void* stream) {
  B2_REQUIRE(ctx && spec, "b2_decode_csv: NULL argument");
  B2_REQUIRE(n >= 0, "b2_decode_csv: n < 0");
  if (n == 0) return B2_OK;
  B2_REQUIRE(bytes_dev && offs_dev && out_dev, "b2_decode_csv: NULL device pointer");
  DeviceGuard g(ctx->device);
  DecodeArgs<18> args;  // D is irrelevant in MODE 0
  memset(&args, 0, sizeof(args));
  int rc = fill_csv(spec, args.csv);
  if (rc) return rc;
  args.bytes = bytes_dev;
  args.offs = offs_dev;
  args.n = n;
  args.fmt = FMT_CSV;
  args.n_out = args.csv.n_float;
  args.out = out_dev;
  args.span = span_dev;
  args.status = status_dev;
  return launch_decode<18, 0, false>(ctx, args, (cudaStream_t)stream);
}

extern "C" int b2_decode_avro(b2_ctx* ctx, const uint8_t* bytes_dev, const int32_t* offs_dev, int64_t n, int skip_prefix,
                              const b2_avro_prog* prog, float* out_dev, uint8_t* label_dev, int32_t* status_dev, void* stream) {
  B2_REQUIRE(ctx && prog, "b2_decode_avro: NULL argument");
  B2_REQUIRE(n >= 0, "b2_decode_avro: n < 0");
  if (n == 0) return B2_OK;
  B2_REQUIRE(bytes_dev && offs_dev && out_dev, "b2_decode_avro: NULL device pointer");
  DeviceGuard g(ctx->device);
  DecodeArgs<18> args;
  memset(&args, 0, sizeof(args));
  int rc = fill_avro(ctx, prog, skip_prefix, args.avro);
  if (rc) return rc;
  args.bytes = bytes_dev;
  args.offs = offs_dev;
  args.n = n;
  args.fmt = FMT_AVRO;
  args.n_out = prog->n_out;
  args.out = out_dev;
  args.status = status_dev;
  args.label = label_dev;
  return launch_decode<18, 0, false>(ctx, args, (cudaStream_t)stream);
}

static int fill_json(const b2_json_spec* sp, JsonP& j) {
  B2_REQUIRE(sp->n_keys >= 1 && sp->n_keys <= B2_JSON_MAX_KEYS, "json: n_keys=%d out of range", sp->n_keys);
  B2_REQUIRE(sp->n_out >= 1 && sp->n_out <= B2_MAX_WIDTH, "json: n_out=%d out of range", sp->n_out);
  memset(&j, 0, sizeof(j));
  j.n_keys = sp->n_keys;
  j.n_out = sp->n_out;
  for (int k = 0; k < sp->n_keys; ++k) {
    const int len = sp->name_len[k];
    B2_REQUIRE(len >= 1 && len <= B2_JSON_NAME_BYTES, "json: name %d has length %d (1..%d)", k, len, B2_JSON_NAME_BYTES);
    B2_REQUIRE(sp->out[k] >= -2 && sp->out[k] < sp->n_out, "json: name %d routed to column %d", k, sp->out[k]);
    const uint8_t* nm = reinterpret_cast<const uint8_t*>(sp->name[k]);
    memcpy(j.name[k], nm, len);
    j.len[k] = (uint8_t)len;
    j.out[k] = sp->out[k];
  }
  b2_json_finish_spec(j);
  return B2_OK;
}

extern "C" int b2_decode_json(b2_ctx* ctx, const uint8_t* bytes_dev, const int32_t* offs_dev, int64_t n, const b2_json_spec* spec,
                              float* out_dev, uint8_t* label_dev, int32_t* status_dev, void* stream) {
  B2_REQUIRE(ctx && spec, "b2_decode_json: NULL argument");
  B2_REQUIRE(n >= 0, "b2_decode_json: n < 0");
  if (n == 0) return B2_OK;
  B2_REQUIRE(bytes_dev && offs_dev && out_dev, "b2_decode_json: NULL device pointer");
  DeviceGuard g(ctx->device);
  DecodeArgs<18> args;
  memset(&args, 0, sizeof(args));
  int rc = fill_json(spec, args.json);
  if (rc) return rc;
  args.bytes = bytes_dev;
  args.offs = offs_dev;
  args.n = n;
  args.fmt = FMT_JSON;
  args.n_out = spec->n_out;
  args.out = out_dev;
  args.status = status_dev;
  args.label = label_dev;
  return launch_decode<18, 0, false>(ctx, args, (cudaStream_t)stream);
}

// Decode only, records given as byte ranges (see b2_decode_score_spans): the three decoders behind one entry point.
extern "C" int b2_decode_spans(b2_ctx* ctx, const uint8_t* bytes_dev, const int32_t* begin_dev, const int32_t* end_dev, int64_t n,
                               const b2_csv_spec* csv, const b2_avro_prog* avro, const b2_json_spec* json, int skip_prefix, float* out_dev,
                               uint8_t* label_dev, int32_t* span_dev, int32_t* status_dev, void* stream) {
  B2_REQUIRE(ctx, "b2_decode_spans: NULL context");
  B2_REQUIRE((csv != nullptr) + (avro != nullptr) + (json != nullptr) == 1, "b2_decode_spans: exactly one of csv / avro / json must be given");
  B2_REQUIRE(n >= 0, "b2_decode_spans: n < 0");
  if (n == 0) return B2_OK;
  B2_REQUIRE(bytes_dev && begin_dev && end_dev && out_dev, "b2_decode_spans: NULL device pointer");
  DeviceGuard g(ctx->device);
  DecodeArgs<18> args;  // D is irrelevant in MODE 0
  memset(&args, 0, sizeof(args));
  int rc;
  if (csv) {
    rc = fill_csv(csv, args.csv);
    args.fmt = FMT_CSV;
    args.n_out = args.csv.n_float;
    args.span = span_dev;
  } else if (avro) {
    rc = fill_avro(ctx, avro, skip_prefix, args.avro);
    args.fmt = FMT_AVRO;
    args.n_out = avro->n_out;
    args.label = label_dev;
  } else {
    rc = fill_json(json, args.json);
    args.fmt = FMT_JSON;
    args.n_out = json->n_out;
    args.label = label_dev;
  }
  if (rc) return rc;
  args.bytes = bytes_dev;
  args.offs = begin_dev;
  args.ends = end_dev;
  args.n = n;
  args.out = out_dev;
  args.status = status_dev;
  return launch_decode<18, 0, false>(ctx, args, (cudaStream_t)stream);
}

template <int D>
static int decode_score_spec(b2_ae* ae, const uint8_t* bytes, const int32_t* offs, const int32_t* ends, int64_t n, const b2_csv_spec* csv,
                             const b2_avro_prog* avro, const b2_json_spec* json, int skip_prefix, const b2_norm* norm, float threshold, float* mse, uint8_t* flag,
                             uint8_t* label, int32_t* status, unsigned long long* counters, cudaStream_t stream) {
  DecodeArgs<D> args;
  memset(&args, 0, sizeof(args));
  int rc;
  if (csv) {
    rc = fill_csv(csv, args.csv);
    if (rc) return rc;
    B2_REQUIRE(args.csv.n_float == D, "b2_decode_score: CSV spec yields %d float columns, model input is %d", args.csv.n_float, D);
    args.fmt = FMT_CSV;
  } else if (json) {
    rc = fill_json(json, args.json);
    if (rc) return rc;
    B2_REQUIRE(json->n_out == D, "b2_decode_score_json: spec yields %d columns, model input is %d", json->n_out, D);
    args.fmt = FMT_JSON;
  } else {
    rc = fill_avro(ae->ctx, avro, skip_prefix, args.avro);
    if (rc) return rc;
    B2_REQUIRE(avro->n_out == D, "b2_decode_score: Avro program yields %d columns, model input is %d", avro->n_out, D);
    args.fmt = FMT_AVRO;
  }
  args.bytes = bytes;
  args.offs = offs;
  args.ends = ends;
  args.n = n;
  args.n_out = D;
  args.status = status;
  args.label = label;
  ae_pack_weights<D>(ae->h_w.data(), args.w);
  args.has_norm = norm_fill<D>(norm, args.nm);
  if (args.has_norm && norm->order != B2_NORM_ORDER_F32) {
    B2_REQUIRE(norm->order == B2_NORM_ORDER_F64, "b2_decode_score: b2_norm.order=%d unknown", norm->order);
    if (!avro) {
      b2_set_error("b2_decode_score: the float64 normalisation order is defined for Avro records (cardata-v3.py decodes Avro as float64); "
                   "CSV and JSON values are float32 after tf.io.decode_csv / this library's JSON route");
      return B2_EUNSUPPORTED;
    }
    args.norm64 = 1;
    for (int i = 0; i < D; ++i) args.n64_lo[i] = (double)norm->lo[i], args.n64_rg[i] = (double)norm->range[i];
  }
  if (csv && args.has_norm && ae->ctx->knobs.zero_skip != 0) {
    // CSV float fields whose column the table forces to zero: the field walk checks the token and skips its conversion
    for (int f = 0; f < args.csv.n_fields; ++f)
      if (args.csv.kind[f] == 1 && args.nm.kind[args.csv.col[f]] == B2_NORM_ZERO) args.csv.nov[f] = 1;
  }
  args.threshold = threshold;
  args.mse = mse;
  args.flag = flag;
  args.counters = counters;
  if (ae->ctx->knobs.decode_tc && ae->ctx->knobs.fast_tanh) {
    rc = b2_ae_tc_panels(ae, &args.panels);
    if (rc) return rc;
  }
  if (ae->ctx->knobs.fast_tanh) return launch_decode<D, 1, true>(ae->ctx, args, stream);
  return launch_decode<D, 1, false>(ae->ctx, args, stream);
}

static int decode_score_any(const char* who, b2_ae* ae, const uint8_t* bytes_dev, const int32_t* offs_dev, const int32_t* ends_dev, int64_t n,
                            const b2_csv_spec* csv,
                            const b2_avro_prog* avro, const b2_json_spec* json, int skip_prefix, const b2_norm* norm, float threshold,
                            float* mse_dev, uint8_t* flag_dev, uint8_t* label_dev, int32_t* status_dev, unsigned long long* counters_dev,
                            void* stream) {
  B2_REQUIRE(ae != nullptr, "%s: ae is NULL", who);
  B2_REQUIRE(n >= 0, "%s: n < 0", who);
  if (n == 0) return B2_OK;
  B2_REQUIRE(bytes_dev && offs_dev, "%s: NULL device pointer", who);
  B2_REQUIRE(!norm || norm->n == 0 || norm->n == ae->desc.widths[0], "%s: norm table has %d features, model has %d", who,
             norm ? norm->n : 0, ae->desc.widths[0]);
  DeviceGuard g(ae->ctx->device);
  int rc = b2_ae_sync_host(ae);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  if (ae->spec_d == 18)
    return decode_score_spec<18>(ae, bytes_dev, offs_dev, ends_dev, n, csv, avro, json, skip_prefix, norm, threshold, mse_dev, flag_dev, label_dev,
                                 status_dev, counters_dev, st);
  if (ae->spec_d == 30)
    return decode_score_spec<30>(ae, bytes_dev, offs_dev, ends_dev, n, csv, avro, json, skip_prefix, norm, threshold, mse_dev, flag_dev, label_dev,
                                 status_dev, counters_dev, st);
  b2_set_error("%s: fused path needs the D-14-7-7-D autoencoder with D in {18, 30}; decode and score separately", who);
  return B2_EUNSUPPORTED;
}

extern "C" int b2_decode_score(b2_ae* ae, const uint8_t* bytes_dev, const int32_t* offs_dev, int64_t n, const b2_csv_spec* csv,
                               const b2_avro_prog* avro, int skip_prefix, const b2_norm* norm, float threshold, float* mse_dev,
                               uint8_t* flag_dev, uint8_t* label_dev, int32_t* status_dev, unsigned long long* counters_dev,
                               void* stream) {
  B2_REQUIRE((csv != nullptr) != (avro != nullptr), "b2_decode_score: exactly one of csv / avro must be given");
  return decode_score_any("b2_decode_score", ae, bytes_dev, offs_dev, nullptr, n, csv, avro, nullptr, skip_prefix, norm, threshold, mse_dev, flag_dev,
                          label_dev, status_dev, counters_dev, stream);
}

extern "C" int b2_decode_score_json(b2_ae* ae, const uint8_t* bytes_dev, const int32_t* offs_dev, int64_t n, const b2_json_spec* json,
                                    const b2_norm* norm, float threshold, float* mse_dev, uint8_t* flag_dev, uint8_t* label_dev,
                                    int32_t* status_dev, unsigned long long* counters_dev, void* stream) {
  B2_REQUIRE(json != nullptr, "b2_decode_score_json: spec is NULL");
  return decode_score_any("b2_decode_score_json", ae, bytes_dev, offs_dev, nullptr, n, nullptr, nullptr, json, 0, norm, threshold, mse_dev, flag_dev,
                          label_dev, status_dev, counters_dev, stream);
}

// The same fused pass over records given as byte RANGES [begin[i], end[i]) of bytes_dev: ascending, end[i] >= begin[i], gaps
// allowed — what b2_kafka_index_records writes, so that a Kafka segment is decoded and scored where it lies (no packed copy of the
// values).  Exactly one of csv / avro / json.
extern "C" int b2_decode_score_spans(b2_ae* ae, const uint8_t* bytes_dev, const int32_t* begin_dev, const int32_t* end_dev, int64_t n,
                                     const b2_csv_spec* csv, const b2_avro_prog* avro, const b2_json_spec* json, int skip_prefix,
                                     const b2_norm* norm, float threshold, float* mse_dev, uint8_t* flag_dev, uint8_t* label_dev,
                                     int32_t* status_dev, unsigned long long* counters_dev, void* stream) {
  B2_REQUIRE((csv != nullptr) + (avro != nullptr) + (json != nullptr) == 1, "b2_decode_score_spans: exactly one of csv / avro / json must be given");
  B2_REQUIRE(n == 0 || end_dev != nullptr, "b2_decode_score_spans: end_dev is NULL");
  return decode_score_any("b2_decode_score_spans", ae, bytes_dev, begin_dev, end_dev, n, csv, avro, json, json ? 0 : skip_prefix, norm,
                          threshold, mse_dev, flag_dev, label_dev, status_dev, counters_dev, stream);
}
