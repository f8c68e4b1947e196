/*
 * CPU ORACLE (C restatement) — test infrastructure and the timed CPU baseline; never linked into the product.
 *
 * Restates the reference's streaming scoring loop in plain C so that it can be (a) compared with the CUDA path and
 * (b) timed on the GPU box's host cores with every thread (OpenMP) as `bench.py --impl reference` / cpu_baseline.
 * PARITY UNPINNED: the reference publishes no golden outputs for this path and TensorFlow is not installable here;
 * this file is cross-checked against oracle/cpu_oracle.py (NumPy), glibc strtof and pyarrow instead.
 *
 * Reference lines restated (paths relative to the reference root):
 *   CSV split + float parse   tf.io.decode_csv   python-scripts/autoencoder-anomaly-detection/
 *                                                Sensor-Kafka-Consumer-and-TensorFlow-Model-Training.py:32-34
 *   Avro decode               substr(e,5,-1) + kafka_io.decode_avro   python-scripts/AUTOENCODER-TensorFlow-IO-Kafka/cardata-v1.py:12-37
 *   normalize_fn              cardata-v1.py:40-131   (scale_fn :68-69)
 *   Dense forward             cardata-v1.py:161-167  (y = act(x @ K + b), K row-major [in,out])
 *   reconstruction error      nb2.0:1015  mse = mean((x - pred)^2, axis=1)
 *   flag                      nb2.0:1163  mse > threshold (strict), threshold_fixed = 5 (nb2.0:1129)
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define MAXW 64
enum { ACT_LINEAR = 0, ACT_TANH = 1, ACT_RELU = 2 };
enum { NORM_IDENTITY = 0, NORM_ZERO = 1, NORM_SCALE = 2 };
enum { REC_OK = 0, REC_FIELD_COUNT, REC_BAD_NUMBER, REC_EMPTY_FIELD, REC_BAD_QUOTE, REC_TRUNCATED, REC_BAD_BRANCH, REC_TOO_LONG };

int orc_max_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}
void orc_set_threads(int n) {
#ifdef _OPENMP
  omp_set_num_threads(n);
#else
  (void)n;
#endif
}

/* normalize_fn: (v - lo) / (hi - lo) * 2.0 - 1.0 in float32, operation by operation (cardata-v1.py:69) */
static inline float norm_feature(float v, int kind, float lo, float range) {
  if (kind == NORM_SCALE) {
    volatile float t = v - lo;
    t = t / range;
    t = t * 2.0f;
    t = t - 1.0f;
    return t;
  }
  if (kind == NORM_ZERO) return 0.0f;
  return v;
}

/* one record through the Dense stack; returns mse; y (recon) optional */
static float forward_one(const float* x, int n_layers, const int* widths, const int* acts, const float* w, float* y_out) {
  float h[MAXW], t[MAXW];
  const int D = widths[0];
  for (int j = 0; j < D; ++j) h[j] = x[j];
  const float* p = w;
  for (int l = 0; l < n_layers; ++l) {
    const int in = widths[l], out = widths[l + 1];
    const float* k = p;
    const float* b = p + (size_t)in * out;
    for (int o = 0; o < out; ++o) t[o] = 0.0f;
    for (int i = 0; i < in; ++i) {
      const float hi = h[i];
      for (int o = 0; o < out; ++o) t[o] += hi * k[i * out + o];
    }
    for (int o = 0; o < out; ++o) {
      float z = t[o] + b[o];
      if (acts[l] == ACT_TANH) z = tanhf(z);
      else if (acts[l] == ACT_RELU) z = z > 0.0f ? z : 0.0f;
      h[o] = z;
    }
    p += (size_t)in * out + out;
  }
  float se = 0.0f;
  for (int j = 0; j < D; ++j) {
    const float d = x[j] - h[j];
    se += d * d;
    if (y_out) y_out[j] = h[j];
  }
  return se / (float)D;
}

/* normalise → forward → mse → flag over n rows; returns the number of flagged rows */
long long orc_score(const float* x, long long n, int n_layers, const int* widths, const int* acts, const float* w, int norm_n,
                    const int* nkind, const float* nlo, const float* nrange, float threshold, float* mse, uint8_t* flag,
                    float* recon) {
  const int D = widths[0];
  long long flagged = 0;
#pragma omp parallel for schedule(static) reduction(+ : flagged)
  for (long long r = 0; r < n; ++r) {
    float xn[MAXW];
    for (int j = 0; j < D; ++j) xn[j] = norm_n ? norm_feature(x[r * D + j], nkind[j], nlo[j], nrange[j]) : x[r * D + j];
    const float m = forward_one(xn, n_layers, widths, acts, w, recon ? recon + r * D : 0);
    const int f = m > threshold;
    if (mse) mse[r] = m;
    if (flag) flag[r] = (uint8_t)f;
    flagged += f;
  }
  return flagged;
}

/* ---- CSV: split on delim honouring RFC-4180 quotes, strtof per float field -------------------------------- */
static int parse_float_field(const uint8_t* s, int len, float* out) {
  char buf[80];
  while (len > 0 && (*s == ' ' || *s == '\t')) ++s, --len;
  while (len > 0 && (s[len - 1] == ' ' || s[len - 1] == '\t')) --len;
  if (len <= 0 || len >= (int)sizeof(buf)) return 0;
  memcpy(buf, s, (size_t)len);
  buf[len] = 0;
  { /* hex floats and nan(...) payloads are rejected (the CUDA path does the same) */
    const char* q = buf;
    if (*q == '+' || *q == '-') ++q;
    if (q[0] == '0' && (q[1] == 'x' || q[1] == 'X')) return 0;
    if ((q[0] == 'n' || q[0] == 'N') && len > 3 + (int)(q - buf)) return 0;
  }
  char* end = 0;
  *out = strtof(buf, &end);
  return end == buf + len;
}

static int decode_csv_record(const uint8_t* rec, int n, int n_fields, const int* kind, const int* has_default, const float* dflt,
                             int delim, int use_quote, float* out, int32_t* span, int32_t base) {
  int i = 0, field = 0, oc = 0, sc = 0;
  for (;;) {
    int b, e; /* payload [b, e) */
    if (use_quote && i < n && rec[i] == '"') {
      b = ++i;
      int closed = 0;
      while (i < n) {
        if (rec[i] == '"') {
          if (i + 1 < n && rec[i + 1] == '"') { i += 2; continue; }
          closed = 1;
          break;
        }
        ++i;
      }
      if (!closed) return REC_BAD_QUOTE;
      e = i++;
      if (i < n && rec[i] != delim) return REC_BAD_QUOTE;
    } else {
      b = i;
      while (i < n && rec[i] != delim) {
        if ((use_quote && rec[i] == '"') || rec[i] == '\n' || rec[i] == '\r') return REC_BAD_QUOTE;
        ++i;
      }
      e = i;
    }
    if (field >= n_fields) return REC_FIELD_COUNT;
    if (kind[field] == 1) {
      if (e == b) {
        if (!has_default[field]) return REC_EMPTY_FIELD;
        out[oc++] = dflt[field];
      } else if (!parse_float_field(rec + b, e - b, &out[oc++])) {
        return REC_BAD_NUMBER;
      }
    } else if (kind[field] == 2) {
      if (span) { span[2 * sc] = base + b; span[2 * sc + 1] = base + e; }
      ++sc;
    }
    ++field;
    if (i >= n) break;
    ++i; /* delimiter */
    if (i == n) { /* trailing delimiter: one more empty field */
      if (field >= n_fields) return REC_FIELD_COUNT;
      if (kind[field] == 1) {
        if (!has_default[field]) return REC_EMPTY_FIELD;
        out[oc++] = dflt[field];
      } else if (kind[field] == 2) {
        if (span) { span[2 * sc] = base + i; span[2 * sc + 1] = base + i; }
        ++sc;
      }
      ++field;
      break;
    }
  }
  return field == n_fields ? REC_OK : REC_FIELD_COUNT;
}

void orc_decode_csv(const uint8_t* data, const int32_t* offs, long long n, int n_fields, const int* kind, const int* has_default,
                    const float* dflt, int delim, int use_quote, float* out, int32_t* span, int32_t* status) {
  int nf = 0, ns = 0;
  for (int i = 0; i < n_fields; ++i) { nf += kind[i] == 1; ns += kind[i] == 2; }
#pragma omp parallel for schedule(static)
  for (long long r = 0; r < n; ++r) {
    float tmp[MAXW];
    const int st = decode_csv_record(data + offs[r], offs[r + 1] - offs[r], n_fields, kind, has_default, dflt, delim, use_quote, tmp,
                                     span ? span + r * ns * 2 : 0, offs[r]);
    for (int j = 0; j < nf; ++j) out[r * nf + j] = st == REC_OK ? tmp[j] : NAN;
    if (status) status[r] = st;
  }
}

/* ---- Avro ---------------------------------------------------------------------------------------------------- */
typedef struct { int8_t type, nullable, out, pad; } orc_avro_field;
enum { T_INT = 1, T_LONG, T_FLOAT, T_DOUBLE, T_STRING, T_BOOLEAN, T_BYTES, T_NULL };

static int read_varint(const uint8_t* p, const uint8_t* end, int max_bytes, int64_t* v, const uint8_t** np_) {
  uint64_t u = 0;
  int shift = 0;
  for (int k = 0; k < max_bytes; ++k) {
    if (p >= end) return 0;
    const uint8_t b = *p++;
    u |= (uint64_t)(b & 0x7F) << shift;
    shift += 7;
    if (!(b & 0x80)) {
      *v = (int64_t)(u >> 1) ^ -(int64_t)(u & 1);
      *np_ = p;
      return 1;
    }
  }
  return 0;
}

static int decode_avro_record(const uint8_t* p, const uint8_t* end, int n_fields, const orc_avro_field* f, float* out, uint8_t* label) {
  *label = 2;
  for (int i = 0; i < n_fields; ++i) {
    int present = 1;
    int64_t v;
    if (f[i].nullable) {
      if (!read_varint(p, end, 10, &v, &p)) return REC_TRUNCATED;
      if (v != 0 && v != 1) return REC_BAD_BRANCH;
      present = v != (f[i].nullable == 1 ? 0 : 1);
    }
    float val = 0.0f;
    uint8_t lab = 2;
    if (present) {
      switch (f[i].type) {
        case T_INT:
          if (!read_varint(p, end, 5, &v, &p)) return REC_TRUNCATED;
          val = (float)(int32_t)v;
          break;
        case T_LONG:
          if (!read_varint(p, end, 10, &v, &p)) return REC_TRUNCATED;
          val = (float)v;
          break;
        case T_FLOAT:
          if (p + 4 > end) return REC_TRUNCATED;
          memcpy(&val, p, 4);
          p += 4;
          break;
        case T_DOUBLE: {
          if (p + 8 > end) return REC_TRUNCATED;
          double d;
          memcpy(&d, p, 8);
          val = (float)d;
          p += 8;
          break;
        }
        case T_BOOLEAN:
          if (p + 1 > end) return REC_TRUNCATED;
          val = *p++ ? 1.0f : 0.0f;
          break;
        case T_STRING:
        case T_BYTES:
          if (!read_varint(p, end, 10, &v, &p) || v < 0 || p + v > end) return REC_TRUNCATED;
          lab = (v == 5 && !memcmp(p, "false", 5)) ? 0 : (v == 4 && !memcmp(p, "true", 4)) ? 1 : v == 0 ? 2 : 3;
          p += v;
          break;
        default: break;
      }
    }
    if (f[i].out >= 0) out[f[i].out] = val;
    else if (f[i].out == -2) *label = lab;
  }
  return REC_OK;
}

void orc_decode_avro(const uint8_t* data, const int32_t* offs, long long n, int skip_prefix, int n_fields, const orc_avro_field* f,
                     int n_out, float* out, uint8_t* label, int32_t* status) {
#pragma omp parallel for schedule(static)
  for (long long r = 0; r < n; ++r) {
    float tmp[MAXW];
    uint8_t lab = 2;
    for (int j = 0; j < n_out; ++j) tmp[j] = 0.0f;
    const uint8_t* p = data + offs[r] + skip_prefix;
    const uint8_t* end = data + offs[r + 1];
    const int st = p > end ? REC_TRUNCATED : decode_avro_record(p, end, n_fields, f, tmp, &lab);
    for (int j = 0; j < n_out; ++j) out[r * n_out + j] = st == REC_OK ? tmp[j] : NAN;
    if (label) label[r] = st == REC_OK ? lab : 2;
    if (status) status[r] = st;
  }
}

/* ---- the whole streaming loop on the CPU: decode → normalise → forward → mse → flag ----------------------------- */
long long orc_decode_score(const uint8_t* data, const int32_t* offs, long long n, int is_avro, int skip_prefix, int n_fields,
                           const int* csv_kind, const int* csv_has_default, const float* csv_dflt, int delim, int use_quote,
                           const orc_avro_field* avro, int n_layers, const int* widths, const int* acts, const float* w, int norm_n,
                           const int* nkind, const float* nlo, const float* nrange, float threshold, float* mse, uint8_t* flag,
                           int32_t* status) {
  const int D = widths[0];
  long long flagged = 0;
#pragma omp parallel for schedule(static) reduction(+ : flagged)
  for (long long r = 0; r < n; ++r) {
    float x[MAXW], xn[MAXW];
    uint8_t lab;
    int st;
    for (int j = 0; j < D; ++j) x[j] = 0.0f;
    if (is_avro) {
      const uint8_t* p = data + offs[r] + skip_prefix;
      const uint8_t* end = data + offs[r + 1];
      st = p > end ? REC_TRUNCATED : decode_avro_record(p, end, n_fields, avro, x, &lab);
    } else {
      st = decode_csv_record(data + offs[r], offs[r + 1] - offs[r], n_fields, csv_kind, csv_has_default, csv_dflt, delim, use_quote, x,
                             0, 0);
    }
    float m = NAN;
    int f = 0;
    if (st == REC_OK) {
      for (int j = 0; j < D; ++j) xn[j] = norm_n ? norm_feature(x[j], nkind[j], nlo[j], nrange[j]) : x[j];
      m = forward_one(xn, n_layers, widths, acts, w, 0);
      f = m > threshold;
    }
    if (mse) mse[r] = m;
    if (flag) flag[r] = (uint8_t)f;
    if (status) status[r] = st;
    flagged += f;
  }
  return flagged;
}
