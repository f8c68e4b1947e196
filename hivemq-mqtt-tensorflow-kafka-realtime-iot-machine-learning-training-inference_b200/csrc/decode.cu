// K2 — record decode (CSV split + correctly rounded float parse; Avro zig-zag varint walk) and the fused
// decode → normalise → score kernel (the whole streaming scoring loop of cardata-v1.py:176-231 in one launch).
//
// Replaces  tf.io.decode_csv + tf.stack            Sensor-Kafka-Consumer-and-TensorFlow-Model-Training.py:32-39
//           tf.strings.substr(e, 5, -1)            AUTOENCODER-TensorFlow-IO-Kafka/cardata-v1.py:13
//           kafka_io.decode_avro(e, schema, dtype) cardata-v1.py:16-37  (+ int→float casts :61-66)
//
// HBM layout: a Kafka partition batch is `uint8 data[]` (message payloads back to back, no separators) plus
// `int32 offsets[n+1]`.  A persistent CTA walks tiles of TILE = 2*NT consecutive records.  The bytes of (a prefix of) the
// tile are one contiguous range, fetched by ONE 1-D bulk async copy (TMA engine) into shared memory — sub-tiled when
// the records are long (credit-card CSV ≈ 550 B/record) so that a sub-tile never exceeds STAGE_CAP bytes.
//   phase A  thread-per-record   CSV: SWAR delimiter scan → field-begin table (u16) in smem;  Avro: the whole varint walk
//   phase B  thread-per-field    CSV only: Eisel-Lemire parse of one token → f32 tile in smem
//   phase C  thread-per-2-records  fused normalise → Dense stack → MSE → flag (ae_math.cuh), or coalesced tile write-out
// Algorithmic bytes/record: payload + 4 (offset) read, 5 written (fused) or 4*F + 4 written (decode only).
#include "ae_math.cuh"
#include "float_parse.cuh"

constexpr int DEC_NT = 128;
constexpr int DEC_TILE = 2 * DEC_NT;
constexpr int STAGE_CAP = 48 * 1024;  // bytes of payload staged per sub-tile

enum { FMT_CSV = 0, FMT_AVRO = 1 };
enum { AV_INT = 1, AV_LONG, AV_FLOAT, AV_DOUBLE, AV_STRING, AV_BOOLEAN, AV_BYTES, AV_NULL };

struct CsvP {
  int32_t n_fields, n_float, n_str, delim, use_quote;
  uint8_t kind[B2_MAX_FIELDS];       // 0 skip, 1 float, 2 string
  uint8_t has_default[B2_MAX_FIELDS];
  uint8_t col[B2_MAX_FIELDS];        // output column (float) or span slot (string) of field i
  uint8_t fidx[B2_MAX_FIELDS];       // k-th non-skipped field → field index
  float dflt[B2_MAX_FIELDS];
};
struct AvroP {
  int32_t n_fields, n_out, skip_prefix, pad;
  b2_avro_field f[B2_MAX_FIELDS];
};

template <int D>
struct DecodeArgs {
  const uint8_t* bytes;
  const int32_t* offs;
  long long n;
  int fmt;
  int n_out;            // floats per record produced by the decode (== D when scoring)
  CsvP csv;
  AvroP avro;
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
};

// ---- CSV phase A: field boundaries of one record -------------------------------------------------------------
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
  for (int p = b & ~3; p < e; p += 4) {
    const uint32_t wv = *reinterpret_cast<const uint32_t*>(base + p);
    uint32_t keep = 0xFFFFFFFFu;
    if (p < b) keep &= 0xFFFFFFFFu << ((b - p) * 8);
    if (p + 4 > e) keep &= 0xFFFFFFFFu >> ((p + 4 - e) * 8);
    uint32_t m = __vcmpeq4(wv, d4) & keep;
    special |= (__vcmpeq4(wv, 0x22222222u) | __vcmpeq4(wv, 0x0A0A0A0Au) | __vcmpeq4(wv, 0x0D0D0D0Du)) & keep;
    while (m) {
      const int k = (__ffs(m) - 1) >> 3;
      ++nf;
      if (nf < c.n_fields) fb[nf] = (uint16_t)(p + k + 1);
      m &= ~(0xFFu << (k * 8));
    }
  }
  if (special) return csv_scan_slow(base, b, e, c, fb, qmask);  // quotes / CR / LF present: careful sequential scan
  if (nf + 1 != c.n_fields) return B2_REC_FIELD_COUNT;
  fb[c.n_fields] = (uint16_t)(e + 1);
  return B2_REC_OK;
}

// ---- Avro: one record, thread-sequential -----------------------------------------------------------------------
__device__ __forceinline__ bool av_varint(const uint8_t* base, int& p, int e, int max_bytes, long long* v) {
  unsigned long long u = 0;
  int shift = 0;
  for (int k = 0; k < max_bytes; ++k) {
    if (p >= e) return false;
    const uint8_t b = base[p++];
    u |= (unsigned long long)(b & 0x7F) << shift;
    shift += 7;
    if (!(b & 0x80)) {
      *v = (long long)(u >> 1) ^ -(long long)(u & 1);
      return true;
    }
  }
  return false;
}

__device__ __forceinline__ int avro_decode_record(const uint8_t* base, int p, int e, const AvroP& a, float* xrow, uint8_t* label) {
  uint8_t lab_out = 2;
  p += a.skip_prefix;
  if (p > e) return B2_REC_TRUNCATED;
  for (int i = 0; i < a.n_fields; ++i) {
    const int type = a.f[i].type, nullable = a.f[i].nullable, oc = a.f[i].out;
    bool present = true;
    long long v;
    if (nullable) {
      if (!av_varint(base, p, e, 10, &v)) return B2_REC_TRUNCATED;
      if (v != 0 && v != 1) return B2_REC_BAD_BRANCH;
      present = v != (nullable == 1 ? 0 : 1);
    }
    float val = 0.0f;
    uint8_t lab = 2;
    if (present) {
      if (type == AV_FLOAT) {
        if (p + 4 > e) return B2_REC_TRUNCATED;
        const uint32_t bits = (uint32_t)base[p] | ((uint32_t)base[p + 1] << 8) | ((uint32_t)base[p + 2] << 16) | ((uint32_t)base[p + 3] << 24);
        val = __uint_as_float(bits);
        p += 4;
      } else if (type == AV_DOUBLE) {
        if (p + 8 > e) return B2_REC_TRUNCATED;
        unsigned long long bits = 0;
#pragma unroll
        for (int k = 0; k < 8; ++k) bits |= (unsigned long long)base[p + k] << (8 * k);
        val = __double2float_rn(__longlong_as_double((long long)bits));  // one rounding, like Keras' float32 input cast
        p += 8;
      } else if (type == AV_INT) {
        if (!av_varint(base, p, e, 5, &v)) return B2_REC_TRUNCATED;
        val = __int2float_rn((int)v);  // tf.cast(int32 → float32), cardata-v1.py:61-66
      } else if (type == AV_LONG) {
        if (!av_varint(base, p, e, 10, &v)) return B2_REC_TRUNCATED;
        val = __ll2float_rn(v);
      } else if (type == AV_BOOLEAN) {
        if (p + 1 > e) return B2_REC_TRUNCATED;
        val = base[p++] ? 1.0f : 0.0f;
      } else if (type == AV_STRING || type == AV_BYTES) {
        if (!av_varint(base, p, e, 10, &v) || v < 0 || (long long)p + v > e) return B2_REC_TRUNCATED;
        const int L = (int)v;
        if (L == 5 && base[p] == 'f' && base[p + 1] == 'a' && base[p + 2] == 'l' && base[p + 3] == 's' && base[p + 4] == 'e') lab = 0;
        else if (L == 4 && base[p] == 't' && base[p + 1] == 'r' && base[p + 2] == 'u' && base[p + 3] == 'e') lab = 1;
        else lab = L == 0 ? 2 : 3;
        p += L;
      }
    }
    if (oc >= 0) xrow[oc] = val;
    else if (oc == -2) lab_out = lab;
  }
  *label = lab_out;
  return B2_REC_OK;
}

// ---- the kernel --------------------------------------------------------------------------------------------------
// MODE 0: decode only (runtime n_out);  MODE 1: fused decode + score for the D-14-7-7-D autoencoder
template <int D, int MODE, bool FT>
__global__ void __launch_bounds__(DEC_NT) decode_kernel(const __grid_constant__ DecodeArgs<D> a) {
  extern __shared__ __align__(128) uint8_t smem[];
  // carve: [stage bytes | offsets (TILE+1) | status (TILE) | label (TILE) | qmask (TILE) | fb (TILE*(nf+1)) | xt (TILE*stride)]
  uint8_t* stage = smem;
  int32_t* s_offs = reinterpret_cast<int32_t*>(smem + STAGE_CAP + 64);
  int32_t* s_stat = s_offs + DEC_TILE + 4;
  unsigned long long* s_qm = reinterpret_cast<unsigned long long*>(s_stat + DEC_TILE);
  uint8_t* s_lab = reinterpret_cast<uint8_t*>(s_qm + DEC_TILE);
  uint64_t* full = reinterpret_cast<uint64_t*>(s_lab + DEC_TILE);
  const int fb_stride = a.fmt == FMT_CSV ? a.csv.n_fields + 1 : 0;
  uint16_t* s_fb = reinterpret_cast<uint16_t*>(full + 2);
  const int xt_stride = MODE == 1 ? D : a.n_out;
  float* s_xt = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(s_fb) + (((size_t)DEC_TILE * fb_stride * 2 + 15) & ~(size_t)15));

  const int tid = threadIdx.x;
  if (tid == 0) {
    mbar_init(&full[0], 1);
    mbar_fence_init();
  }
  __syncthreads();
  uint32_t phase = 0;
  unsigned flagged = 0, n_ok = 0, n_err = 0;
  unsigned long long n_bytes = 0;
  const long long n_tiles = (a.n + DEC_TILE - 1) / DEC_TILE;

  for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const long long row0 = tile * DEC_TILE;
    const int rows = (int)min((long long)DEC_TILE, a.n - row0);
    for (int i = tid; i <= rows; i += DEC_NT) s_offs[i] = a.offs[row0 + i];
    __syncthreads();
    int sub = 0;
    while (sub < rows) {
      // how many records starting at `sub` fit in the stage (byte range is 16-byte aligned outwards)
      const long long b_abs = s_offs[sub];
      const uintptr_t a0 = (reinterpret_cast<uintptr_t>(a.bytes) + (uintptr_t)b_abs) & ~(uintptr_t)15;
      const int lead = (int)((reinterpret_cast<uintptr_t>(a.bytes) + (uintptr_t)b_abs) - a0);
      int fit = 0;
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int r = sub + tid + h * DEC_NT;
        if (r < rows) fit += (long long)s_offs[r + 1] - b_abs + lead <= (long long)STAGE_CAP;
      }
      const int cnt = __syncthreads_count(fit) + __syncthreads_count(fit > 1);  // records are monotone: count of fitting ones
      if (cnt == 0) {  // a single record larger than the stage: refuse it, keep going
        if (tid == 0) {
          const long long r = row0 + sub;
          if (a.status) a.status[r] = B2_REC_TOO_LONG;
          if (MODE == 1) {
            if (a.mse) a.mse[r] = __int_as_float(0x7FC00000);
            if (a.flag) a.flag[r] = 0;
          } else {
            for (int j = 0; j < a.n_out; ++j) a.out[r * a.n_out + j] = __int_as_float(0x7FC00000);
          }
          if (a.label) a.label[r] = 2;
          ++n_err;
        }
        sub += 1;
        __syncthreads();
        continue;
      }
      const long long e_abs = s_offs[sub + cnt];
      const uint32_t nbytes = (uint32_t)(((e_abs - b_abs + lead) + 15) & ~15ll);
      if (tid == 0) {
        mbar_expect_tx(&full[0], nbytes);
        bulk_g2s(stage, reinterpret_cast<const void*>(a0), nbytes, &full[0]);
        n_bytes += (unsigned long long)(e_abs - b_abs);
      }
      mbar_wait(&full[0], phase);
      phase ^= 1u;
      const int rel = lead - (int)b_abs;  // stage offset of absolute byte x is x + rel

      // ---- phase A ----
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int r = tid + h * DEC_NT;  // index inside the sub-tile
        if (r < cnt) {
          const int b = s_offs[sub + r] + rel, e = s_offs[sub + r + 1] + rel;
          int st;
          if (a.fmt == FMT_CSV) {
            st = csv_scan(stage, b, e, a.csv, s_fb + (size_t)r * fb_stride, &s_qm[r]);
          } else {
            float* xrow = s_xt + (size_t)r * xt_stride;
            for (int j = 0; j < xt_stride; ++j) xrow[j] = 0.0f;
            uint8_t lab = 2;
            st = avro_decode_record(stage, b, e, a.avro, xrow, &lab);
            s_lab[r] = lab;
          }
          s_stat[r] = st;
        }
      }
      __syncthreads();
      // ---- phase B (CSV): one task per (record, float/string field) ----
      if (a.fmt == FMT_CSV) {
        const int per_rec = a.csv.n_float + a.csv.n_str;
        for (int t = tid; t < cnt * per_rec; t += DEC_NT) {
          const int r = t / per_rec, k = t - r * per_rec;
          if (s_stat[r] != B2_REC_OK) continue;
          const int f = a.csv.fidx[k];
          const uint16_t* fb = s_fb + (size_t)r * fb_stride;
          // a quoted field's payload sits between its quotes: drop its closing quote and, when the NEXT field is quoted
          // too, that field's opening quote (fb[] holds payload starts)
          const int quoted = (int)((s_qm[r] >> f) & 1ull) + (int)((s_qm[r] >> (f + 1)) & 1ull);
          const int fbeg = fb[f], fend = (int)fb[f + 1] - 1 - quoted;
          if (a.csv.kind[f] == 1) {
            float v;
            int st = B2_REC_OK;
            if (fend <= fbeg) {
              if (a.csv.has_default[f]) v = a.csv.dflt[f];
              else st = B2_REC_EMPTY_FIELD, v = 0.f;
            } else {
              uint32_t bits = 0;
              const int ps = b2_parse_float(stage + fbeg, fend - fbeg, &bits);
              v = __uint_as_float(bits);
              if (ps == B2_NUM_BAD) st = B2_REC_BAD_NUMBER;
              else if (ps == B2_NUM_TOO_LONG) st = B2_REC_TOO_LONG;
            }
            if (st != B2_REC_OK) atomicMax(&s_stat[r], st);
            s_xt[(size_t)r * xt_stride + a.csv.col[f]] = v;
          } else if (MODE == 0 && a.span) {
            int32_t* sp = a.span + ((row0 + sub + r) * a.csv.n_str + a.csv.col[f]) * 2;
            sp[0] = fbeg - rel;
            sp[1] = fend - rel;
          }
        }
        __syncthreads();
      }
      // ---- phase C ----
      if (MODE == 1) {
        const bool va = tid < cnt, vb = tid + DEC_NT < cnt;
        float* const rowp[2] = {s_xt + (size_t)tid * D, s_xt + (size_t)(tid + DEC_NT) * D};
        float mm[2] = {0.f, 0.f};
        if (va) {
          if (a.has_norm) ae_forward_rows<D, 2, true, false, FT>(a.w, a.nm, rowp, mm);
          else ae_forward_rows<D, 2, false, false, FT>(a.w, a.nm, rowp, mm);
        }
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const int r = tid + h * DEC_NT;
          if (r < cnt) {
            const long long gr = row0 + sub + r;
            const int st = s_stat[r];
            float m = mm[h];
            if (st != B2_REC_OK) m = __int_as_float(0x7FC00000);
            const bool f = m > a.threshold;
            if (a.mse) a.mse[gr] = m;
            if (a.flag) a.flag[gr] = f ? 1 : 0;
            if (a.status) a.status[gr] = st;
            if (a.label) a.label[gr] = a.fmt == FMT_AVRO && st == B2_REC_OK ? s_lab[r] : 2;
            flagged += f;
            n_ok += st == B2_REC_OK;
            n_err += st != B2_REC_OK;
          }
        }
      } else {
        for (int i = tid; i < cnt * xt_stride; i += DEC_NT) {
          const int r = i / xt_stride;
          a.out[(row0 + sub) * xt_stride + i] = s_stat[r] == B2_REC_OK ? s_xt[i] : __int_as_float(0x7FC00000);
        }
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const int r = tid + h * DEC_NT;
          if (r < cnt) {
            const long long gr = row0 + sub + r;
            const int st = s_stat[r];
            if (a.status) a.status[gr] = st;
            if (a.label) a.label[gr] = a.fmt == FMT_AVRO && st == B2_REC_OK ? s_lab[r] : 2;
            n_ok += st == B2_REC_OK;
            n_err += st != B2_REC_OK;
          }
        }
      }
      __syncthreads();  // stage, tables and tile are free again
      sub += cnt;
    }
  }
  if (a.counters) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      flagged += __shfl_xor_sync(0xffffffffu, flagged, o);
      n_ok += __shfl_xor_sync(0xffffffffu, n_ok, o);
      n_err += __shfl_xor_sync(0xffffffffu, n_err, o);
    }
    if ((tid & 31) == 0) {
      if (n_ok) atomicAdd(&a.counters[0], (unsigned long long)n_ok);
      if (flagged) atomicAdd(&a.counters[1], (unsigned long long)flagged);
      if (n_err) atomicAdd(&a.counters[2], (unsigned long long)n_err);
    }
    if (tid == 0 && n_bytes) atomicAdd(&a.counters[3], n_bytes);
  }
}

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
    if (s->kind[i] != 0) c.fidx[c.n_float + c.n_str] = (uint8_t)i;
    if (s->kind[i] == 1) c.col[i] = (uint8_t)c.n_float++;
    else if (s->kind[i] == 2) c.col[i] = (uint8_t)c.n_str++;
  }
  B2_REQUIRE(c.n_float >= 1, "csv: no float field selected");
  return B2_OK;
}

static int fill_avro(const b2_avro_prog* p, int skip_prefix, AvroP& a) {
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
  return B2_OK;
}

static size_t decode_smem_bytes(int fb_stride, int xt_stride) {
  size_t s = STAGE_CAP + 64;
  s += sizeof(int32_t) * (DEC_TILE + 4);           // offsets
  s += sizeof(int32_t) * DEC_TILE;                 // status
  s += sizeof(unsigned long long) * DEC_TILE;      // qmask
  s += DEC_TILE;                                   // label
  s += 2 * sizeof(uint64_t);                       // mbarrier
  s += ((size_t)DEC_TILE * fb_stride * 2 + 15) & ~(size_t)15;
  s += (size_t)DEC_TILE * xt_stride * 4;
  return s + 64;
}

template <int D, int MODE, bool FT>
static int launch_decode(b2_ctx* ctx, const DecodeArgs<D>& args, cudaStream_t stream) {
  const int fb_stride = args.fmt == FMT_CSV ? args.csv.n_fields + 1 : 0;
  const int xt_stride = MODE == 1 ? D : args.n_out;
  const size_t smem = decode_smem_bytes(fb_stride, xt_stride);
  B2_REQUIRE(smem <= (size_t)ctx->smem_optin, "decode: tile needs %zu bytes of shared memory (max %d)", smem, ctx->smem_optin);
  auto kern = decode_kernel<D, MODE, FT>;
  B2_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int occ = 0;
  B2_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, DEC_NT, smem));
  B2_REQUIRE(occ >= 1, "decode: kernel does not fit on an SM");
  const long long n_tiles = (args.n + DEC_TILE - 1) / DEC_TILE;
  long long grid = (long long)ctx->sm_count * occ;
  if (grid > n_tiles) grid = n_tiles;
  kern<<<(unsigned)grid, DEC_NT, smem, stream>>>(args);
  B2_CUDA(cudaGetLastError());
  return B2_OK;
}

static int env_int(const char* name, int dflt) {
  const char* v = getenv(name);
  return v ? atoi(v) : dflt;
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
  int rc = fill_avro(prog, skip_prefix, args.avro);
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

template <int D>
static int decode_score_spec(b2_ae* ae, const uint8_t* bytes, const int32_t* offs, int64_t n, const b2_csv_spec* csv,
                             const b2_avro_prog* avro, int skip_prefix, const b2_norm* norm, float threshold, float* mse, uint8_t* flag,
                             uint8_t* label, int32_t* status, unsigned long long* counters, cudaStream_t stream) {
  DecodeArgs<D> args;
  memset(&args, 0, sizeof(args));
  int rc;
  if (csv) {
    rc = fill_csv(csv, args.csv);
    if (rc) return rc;
    B2_REQUIRE(args.csv.n_float == D, "b2_decode_score: CSV spec yields %d float columns, model input is %d", args.csv.n_float, D);
    args.fmt = FMT_CSV;
  } else {
    rc = fill_avro(avro, skip_prefix, args.avro);
    if (rc) return rc;
    B2_REQUIRE(avro->n_out == D, "b2_decode_score: Avro program yields %d columns, model input is %d", avro->n_out, D);
    args.fmt = FMT_AVRO;
  }
  args.bytes = bytes;
  args.offs = offs;
  args.n = n;
  args.n_out = D;
  args.status = status;
  args.label = label;
  ae_pack_weights<D>(ae->h_w.data(), args.w);
  args.has_norm = norm && norm->n > 0;
  if (args.has_norm) {
    for (int j = 0; j < D; ++j) {
      args.nm.kind[j] = norm->kind[j];
      args.nm.lo[j] = norm->lo[j];
      args.nm.range[j] = norm->range[j];
      args.nm.rinv[j] = 1.0f / norm->range[j];
    }
  }
  args.threshold = threshold;
  args.mse = mse;
  args.flag = flag;
  args.counters = counters;
  if (env_int("B2_FAST_TANH", 1)) return launch_decode<D, 1, true>(ae->ctx, args, stream);
  return launch_decode<D, 1, false>(ae->ctx, args, stream);
}

extern "C" int b2_decode_score(b2_ae* ae, const uint8_t* bytes_dev, const int32_t* offs_dev, int64_t n, const b2_csv_spec* csv,
                               const b2_avro_prog* avro, int skip_prefix, const b2_norm* norm, float threshold, float* mse_dev,
                               uint8_t* flag_dev, uint8_t* label_dev, int32_t* status_dev, unsigned long long* counters_dev,
                               void* stream) {
  B2_REQUIRE(ae != nullptr, "b2_decode_score: ae is NULL");
  B2_REQUIRE((csv != nullptr) != (avro != nullptr), "b2_decode_score: exactly one of csv / avro must be given");
  B2_REQUIRE(n >= 0, "b2_decode_score: n < 0");
  if (n == 0) return B2_OK;
  B2_REQUIRE(bytes_dev && offs_dev, "b2_decode_score: NULL device pointer");
  B2_REQUIRE(!norm || norm->n == 0 || norm->n == ae->desc.widths[0], "b2_decode_score: norm table has %d features, model has %d",
             norm ? norm->n : 0, ae->desc.widths[0]);
  DeviceGuard g(ae->ctx->device);
  int rc = b2_ae_sync_host(ae);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  if (ae->spec_d == 18)
    return decode_score_spec<18>(ae, bytes_dev, offs_dev, n, csv, avro, skip_prefix, norm, threshold, mse_dev, flag_dev, label_dev,
                                 status_dev, counters_dev, st);
  if (ae->spec_d == 30)
    return decode_score_spec<30>(ae, bytes_dev, offs_dev, n, csv, avro, skip_prefix, norm, threshold, mse_dev, flag_dev, label_dev,
                                 status_dev, counters_dev, st);
  b2_set_error("b2_decode_score: fused path needs the D-14-7-7-D autoencoder with D in {18, 30}; decode and score separately");
  return B2_EUNSUPPORTED;
}
