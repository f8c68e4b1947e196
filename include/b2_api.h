/*
 * b2_api.h — C ABI of the B200 sensor-scoring hot path.
 *
 * The reference (uurl/hivemq-mqtt-tensorflow-kafka-realtime-iot-machine-learning-training-inference) has no
 * FFI of its own: its hot path sits directly behind TensorFlow / tensorflow-io Python calls.  Each entry point
 * below therefore names the Python call site(s) it replaces (paths relative to the reference root).  All
 * pointers suffixed _dev are device pointers owned by the caller (e.g. torch.Tensor.data_ptr()); `stream` is
 * a cudaStream_t passed as void* (NULL = legacy default stream).  Every function returns 0 on success or a
 * negative B2_E* code; b2_strerror() / b2_last_error() give text.  Functions that take a stream only enqueue
 * work; nothing here allocates device memory on the hot path (handles own their small parameter blocks).
 */
#ifndef B2_API_H
#define B2_API_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B2_OK            0
#define B2_EINVAL       -1   /* bad argument (null pointer, size out of range, unsupported layer stack) */
#define B2_ECUDA        -2   /* a CUDA runtime call failed; b2_last_error() has the CUDA message      */
#define B2_EUNSUPPORTED -3   /* valid request this build does not implement                             */
#define B2_ENOMEM       -4

#define B2_MAX_WIDTH    64   /* widest Dense layer / input the score + train kernels accept */
#define B2_MAX_FIELDS   64   /* CSV / Avro fields per record */

/* activations (Keras names: cardata-v1.py:163-166) */
#define B2_ACT_LINEAR 0
#define B2_ACT_TANH   1
#define B2_ACT_RELU   2

/* per-feature normalisation kinds (normalize_fn, cardata-v1.py:68-111) */
#define B2_NORM_IDENTITY 0
#define B2_NORM_ZERO     1   /* feature forced to 0.0 (:72,:78,:84,:87) */
#define B2_NORM_SCALE    2   /* (v - lo) / range * 2.0 - 1.0 with IEEE division (:69) */

/* per-record decode status */
#define B2_REC_OK          0
#define B2_REC_FIELD_COUNT 1  /* CSV: wrong number of fields ("Expect N fields but have M") */
#define B2_REC_BAD_NUMBER  2  /* CSV: "Field i in record j is not a valid float" */
#define B2_REC_EMPTY_FIELD 3  /* CSV: empty field and the column has no default ("Field i is required") */
#define B2_REC_BAD_QUOTE   4  /* CSV: quote inside unquoted field / unterminated quoted field */
#define B2_REC_TRUNCATED   5  /* Avro: record ends inside a field */
#define B2_REC_BAD_BRANCH  6  /* Avro: union branch index out of range */
#define B2_REC_TOO_LONG    7  /* CSV: number with > 19 significant digits that cannot be rounded unambiguously; any format: record > 24 KB */
#define B2_REC_BAD_JSON    8  /* JSON: not one well-formed RFC 8259 object */

typedef struct b2_ctx b2_ctx;   /* one per device / rank */
typedef struct b2_ae b2_ae;     /* Dense autoencoder: weights (+ Adam state once fit is called) */
typedef struct b2_lstm b2_lstm; /* LSTM stack of LSTM-TensorFlow-IO-Kafka/cardata-v1.py:165-178 */

/* ---- context ------------------------------------------------------------------------------------------ */
int  b2_ctx_create(int device, b2_ctx** out);
void b2_ctx_destroy(b2_ctx* ctx);
int  b2_ctx_sm_count(const b2_ctx* ctx);
/* Optional sizing hint for the decode calls that follow on this context: the average record size (bytes) of the batches to
 * come, e.g. total payload bytes / N (0 = unknown, the default: estimated from the schema).  Each warp stages 32 records at a
 * time; a good hint makes a 32-record chunk fit its stage in one piece and lets more CTAs share an SM. */
int  b2_ctx_set_record_bytes_hint(b2_ctx* ctx, int avg_record_bytes);
/* Tuning / debugging switches: "fast_tanh", "score_cfg", "no_bulk", "force_generic", "avro_generic", "lstm_tc", "score_tc",
 * "tc_ctas", "decode_r2", "decode_tc", "decode_mma" (score phase of the fused kernel as warp-level TF32 MMAs) and "zero_skip"
 * (1, the default: a normalisation table whose zero set is the reference's — features 0, 2, 4, 5, cardata-v1.py:72-87 — runs the
 * kernels compiled for that set; same bits as 0).
 * Their defaults come from the environment (B2_FAST_TANH, …) ONCE, when the context is created; no launch reads the
 * environment afterwards. */
int  b2_ctx_set_option(b2_ctx* ctx, const char* name, int value);
int  b2_ctx_get_option(const b2_ctx* ctx, const char* name, int* value);
const char* b2_strerror(int code);
const char* b2_last_error(void);      /* thread-local detail of the last failure */
const char* b2_version(void);

/* ---- Dense autoencoder -------------------------------------------------------------------------------- */
typedef struct {
  int32_t n_layers;                 /* number of Dense layers (4 for the reference models) */
  int32_t widths[9];                /* widths[0] = input dim D, widths[i+1] = units of layer i */
  int32_t act[8];                   /* B2_ACT_* per layer */
  float   l1_activity;              /* L1 activity regulariser on layer 0's output (1e-7; cardata-v1.py:163) */
} b2_ae_desc;

typedef struct {
  int32_t n;                        /* number of features (must equal widths[0]) or 0 = no normalisation */
  int32_t kind[B2_MAX_WIDTH];
  float   lo[B2_MAX_WIDTH];
  float   range[B2_MAX_WIDTH];      /* hi - lo */
  int32_t order;                    /* B2_NORM_ORDER_F32 (0): values are rounded to float32 first, then scaled in float32 —
                                       cardata-v1.py:16-37, :61-69.  B2_NORM_ORDER_F64 (1): decoded and scaled in float64, rounded
                                       to float32 last (Keras' input cast) — cardata-v3.py:53-74, :98-168.  The float64 order needs
                                       float64 values, i.e. it exists for the fused Avro calls (b2_decode_score /
                                       b2_decode_score_spans with an Avro program); every other call refuses it. */
} b2_norm;
#define B2_NORM_ORDER_F32 0
#define B2_NORM_ORDER_F64 1

/* Replaces tf.keras.models.load_model + model construction (cardata-v1.py:153-167,202; nb2.0:888).
 * host_weights = kernel0[in*out] (row-major [in,out], Keras layout), bias0[out], kernel1, bias1, ... */
int  b2_ae_create(b2_ctx* ctx, const b2_ae_desc* desc, const float* host_weights, b2_ae** out);
void b2_ae_destroy(b2_ae* ae);
int  b2_ae_param_count(const b2_ae* ae);
int  b2_ae_get_weights(b2_ae* ae, float* host_weights);                  /* after fit: read back */
int  b2_ae_set_weights(b2_ae* ae, const float* host_weights);

/* Fused normalise → Dense stack → per-record MSE → flag = mse > threshold.
 * Replaces normalize_fn (cardata-v1.py:40-131) + Model.predict (:231, nb2.0:1014) + np.mean(np.power(x - pred, 2),
 * axis=1) (nb2.0:1015) + `e > threshold_fixed` (nb2.0:1129,1163).
 * x_dev [N,D] row-major f32 (16-byte aligned); mse_dev [N] f32; flag_dev [N] u8 (either may be NULL);
 * recon_dev [N,D] or NULL (predict output); counters_dev u64[4] += {records, flagged, 0, 0} or NULL. */
int  b2_ae_score(b2_ae* ae, const float* x_dev, int64_t n, const b2_norm* norm, float threshold,
                 float* mse_dev, uint8_t* flag_dev, float* recon_dev, unsigned long long* counters_dev,
                 void* stream);

/* normalize_fn as a stand-alone operator (cardata-v1.py:40-131; LSTM-TensorFlow-IO-Kafka/cardata-v1.py:53-143):
 * out[r][j] = (x[r][j] - lo[j]) / range[j] * 2.0 - 1.0 (B2_NORM_SCALE), 0.0 (B2_NORM_ZERO) or x[r][j], x_dev / out_dev f32[n, d]
 * row-major, in binary32 with the reference's operation order — the device function the fused kernels apply, bit-identical
 * to an IEEE division for every input as long as 2^-64 <= |range[j]| <= 2^64 (the reference's table uses 1 … 7500; outside
 * that span the correction step's residual can leave the normal range and the last bit may differ).  b2_selftest_normalize proves that on the device: for every float32 bit pattern in
 * [first_bits, first_bits + count) and every feature of the table it compares that function with sub → IEEE div → fma and
 * adds the number of differences to out_dev[0] (u64[2], zeroed by the caller; [1] = feature << 32 | bits of one of them). */
int  b2_normalize(b2_ctx* ctx, const float* x_dev, int64_t n, int d, const b2_norm* norm, float* out_dev, void* stream);
int  b2_selftest_normalize(b2_ctx* ctx, const b2_norm* norm, int d, uint64_t first_bits, uint64_t count,
                           unsigned long long* out_dev, void* stream);

/* ---- record decode --------------------------------------------------------------------------------------
 * Input contract shared by every decode entry point: record i is bytes_dev[offs_dev[i] .. offs_dev[i+1]).  The kernels fetch
 * a chunk of records with 16-byte-granular bulk copies, so they READ (never write) the whole 16-byte lines that hold the
 * first and the last payload byte: up to 15 bytes before bytes_dev + offs_dev[0] and up to 15 bytes past
 * bytes_dev + offs_dev[n] must be readable.  Any cudaMalloc'ed (256-byte aligned, 256-byte granular) or torch buffer is;
 * a caller that sub-allocates with byte granularity must pad its arena by 16 bytes at both ends.
 * Record length: each warp stages 32 records in shared memory, sized from the schema or from b2_ctx_set_record_bytes_hint (a pure
 * performance hint).  A record that does not fit is decoded by a second, small launch on the same stream with a 24 KB stage,
 * so the only hard limit is 24 KB per record (beyond: B2_REC_TOO_LONG), and at most 1024 such records per call; launches on
 * more than 32 distinct streams per context share that retry list and must not overlap. */
typedef struct {
  int32_t n_fields;                 /* fields per record (20 sensor CSV, 31 credit-card CSV) */
  int32_t kind[B2_MAX_FIELDS];      /* 0 skip, 1 float32 → next output column, 2 string (span reported) */
  int32_t has_default[B2_MAX_FIELDS];
  float   dflt[B2_MAX_FIELDS];      /* value for an empty field when has_default (record_defaults) */
  int32_t delim;                    /* ',' */
  int32_t use_quote_delim;          /* 1: RFC-4180 double quotes honoured (tf.io.decode_csv default) */
} b2_csv_spec;

/* Replaces tf.io.decode_csv + tf.stack (Sensor-Kafka-Consumer-and-TensorFlow-Model-Training.py:32-39).
 * bytes_dev: concatenated message payloads; offs_dev int32[N+1]; out_dev f32[N, n_float] row-major;
 * span_dev int32[N, n_string, 2] (begin,end byte offsets of string fields, NULL to skip);
 * status_dev int32[N] (B2_REC_*, NULL to skip).  Parsed floats are correctly rounded binary32. */
int  b2_decode_csv(b2_ctx* ctx, const uint8_t* bytes_dev, const int32_t* offs_dev, int64_t n,
                   const b2_csv_spec* spec, float* out_dev, int32_t* span_dev, int32_t* status_dev, void* stream);

typedef struct { int8_t type, nullable, out, pad; } b2_avro_field;  /* see avro_schema.py for codes */
typedef struct {
  int32_t n_fields;
  int32_t n_out;
  b2_avro_field f[B2_MAX_FIELDS];
} b2_avro_prog;

/* Replaces tf.strings.substr(e, 5, -1) + kafka_io.decode_avro + the int→float casts of normalize_fn
 * (cardata-v1.py:13-37,61-66).  label_dev u8[N]: 0 "false", 1 "true", 2 null/empty, 3 other (NULL to skip). */
int  b2_decode_avro(b2_ctx* ctx, const uint8_t* bytes_dev, const int32_t* offs_dev, int64_t n, int skip_prefix,
                    const b2_avro_prog* prog, float* out_dev, uint8_t* label_dev, int32_t* status_dev, void* stream);

/* ---- JSON records (the raw `sensor-data` topic) ----------------------------------------------
 * One flat JSON object per message, as the car simulator publishes it and KSQL reads it with value_format='JSON'
 * (01_installConfluentPlatform.sh:235; fixture python-scripts/LSTM-TensorFlow-IO-Kafka/cardata-v1.json).  The members named
 * in the spec are routed to output columns; numbers go text → binary64 (correctly rounded) → binary32 (the scripts' cast);
 * absent or null members give 0, the last of several equal names wins, other members are skipped (validated as RFC 8259).
 * out[k] = column, -2 = the label member ("true"/"false" as string or literal → label_dev 1/0, null/"" 2, else 3), -1 ignore. */
#define B2_JSON_NAME_MAX 32
typedef struct {
  int32_t n_keys;
  int32_t n_out;
  char    name[B2_MAX_FIELDS][B2_JSON_NAME_MAX];   /* member names, raw bytes (no escapes), not NUL-terminated */
  uint8_t name_len[B2_MAX_FIELDS];
  int8_t  out[B2_MAX_FIELDS];
} b2_json_spec;
int  b2_decode_json(b2_ctx* ctx, const uint8_t* bytes_dev, const int32_t* offs_dev, int64_t n, const b2_json_spec* spec,
                    float* out_dev, uint8_t* label_dev, int32_t* status_dev, void* stream);
/* b2_decode_score for JSON records (same outputs and counters). */
int  b2_decode_score_json(b2_ae* ae, const uint8_t* bytes_dev, const int32_t* offs_dev, int64_t n, const b2_json_spec* json,
                          const b2_norm* norm, float threshold, float* mse_dev, uint8_t* flag_dev, uint8_t* label_dev,
                          int32_t* status_dev, unsigned long long* counters_dev, void* stream);
/* The same fused pass over records given as byte RANGES of bytes_dev: record i = [begin_dev[i], end_dev[i]), ranges ascending,
 * end >= begin, gaps between them allowed (a warp's bulk copy brings the bytes in between along; counters_dev[3] counts them).
 * That is what b2_kafka_index_records writes: the reference's KafkaDataset hands the scripts one string per message
 * (cardata-v1.py:10-16); here the segment's bytes are decoded and scored where they lie, with no packed copy of the values.
 * Exactly one of csv / avro / json; skip_prefix as in b2_decode_score (ignored for JSON).  Same input contract as offs_dev
 * otherwise (16 readable bytes behind the last record, a record of more than 24 KB is refused). */
/* b2_decode_spans: decode only (b2_decode_csv / _avro / _json by the spec given), same ranges; span_dev as in b2_decode_csv (string
 * field ranges, absolute offsets into bytes_dev), label_dev as in b2_decode_avro / _json. */
int  b2_decode_spans(b2_ctx* ctx, const uint8_t* bytes_dev, const int32_t* begin_dev, const int32_t* end_dev, int64_t n,
                     const b2_csv_spec* csv, const b2_avro_prog* avro, const b2_json_spec* json, int skip_prefix, float* out_dev,
                     uint8_t* label_dev, int32_t* span_dev, int32_t* status_dev, void* stream);
int  b2_decode_score_spans(b2_ae* ae, const uint8_t* bytes_dev, const int32_t* begin_dev, const int32_t* end_dev, int64_t n,
                           const b2_csv_spec* csv, const b2_avro_prog* avro, const b2_json_spec* json, int skip_prefix,
                           const b2_norm* norm, float threshold, float* mse_dev, uint8_t* flag_dev, uint8_t* label_dev,
                           int32_t* status_dev, unsigned long long* counters_dev, void* stream);

/* ---- Avro encoder (inverse of b2_decode_avro) -------------------------------------------------------------------
 * Rows x_dev f32[n, prog->n_out] (+ label_dev u8[n]: 0 "false", 1 "true", else null; NULL = all null) → Avro binary records
 * under the same field program: the KSQL projection of the reference (CREATE STREAM … WITH (VALUE_FORMAT='AVRO') AS SELECT *,
 * 01_installConfluentPlatform.sh:242) is b2_decode_json followed by this; also the result sink as Avro, the TODO of
 * python-scripts/README.md:108.  int/long = the value rounded to nearest, double = the float widened, a NaN in a nullable
 * column = null, prefix != 0 puts the Confluent framing (0x00 + big-endian schema_id) in front of every record.
 * Two passes: b2_avro_encode_sizes fills size_dev int32[n]; the caller scans them into offs_dev int32[n+1] (offs[0] = 0) and
 * provides bytes_dev of offs[n] bytes; b2_avro_encode writes the records back to back in row order. */
int  b2_avro_encode_sizes(b2_ctx* ctx, const float* x_dev, const uint8_t* label_dev, int64_t n, const b2_avro_prog* prog, int prefix,
                          int32_t* size_dev, void* stream);
int  b2_avro_encode(b2_ctx* ctx, const float* x_dev, const uint8_t* label_dev, int64_t n, const b2_avro_prog* prog, int prefix,
                    int schema_id, const int32_t* offs_dev, uint8_t* bytes_dev, void* stream);

/* Fused decode → normalise → score: the streaming scoring loop of cardata-v1.py:176-231 in one kernel.
 * Exactly one of csv / avro is non-NULL.  Records whose status != 0 get mse = NaN, flag = 0 and are counted
 * in counters[2]; counters_dev u64[4] += {records ok, flagged, decode errors, payload bytes}. */
int  b2_decode_score(b2_ae* ae, const uint8_t* bytes_dev, const int32_t* offs_dev, int64_t n,
                     const b2_csv_spec* csv, const b2_avro_prog* avro, int skip_prefix, const b2_norm* norm,
                     float threshold, float* mse_dev, uint8_t* flag_dev, uint8_t* label_dev, int32_t* status_dev,
                     unsigned long long* counters_dev, void* stream);

/* ---- training ------------------------------------------------------------------------------------------- */
typedef struct {
  float lr, beta1, beta2, eps;      /* Adam(1e-3, .9, .999, 1e-7) from the .h5 training_config */
  int64_t iter;                     /* optimizer step counter (`iter:0`) — in/out */
} b2_adam;

/* Model.fit hot loop (cardata-v1.py:192-194, cardata-v3.py:220-222, nb2.0:871-877): `steps` consecutive
 * mini-batch Adam steps on MSE(x, AE(x)) + l1·Σ|h1|/B inside ONE kernel launch.  Step s uses rows
 * order_dev[s*batch .. s*batch+batch_s) of x_dev (order_dev NULL = sequential), the last step of an epoch may be
 * short: rows_total rows are consumed in ceil(rows_total/batch) steps.  loss_dev f32[steps] and acc_dev
 * f32[steps] receive the per-step loss / "accuracy" metric (NULL to skip).  Single CTA, latency-bound path
 * for the reference's batch sizes (32/100/128). */
int  b2_ae_fit_steps(b2_ae* ae, const float* x_dev, const int32_t* order_dev, int64_t rows_total, int batch,
                     const b2_norm* norm, b2_adam* adam, float* loss_dev, float* acc_dev, void* stream);

/* Large-batch / data-parallel step, split so that a gradient all-reduce can sit between the halves:
 *   b2_ae_grad:  grads_dev f32[P+3] = Σ_rows d(loss)/dθ (un-normalised), then Σ loss, Σ correct, row count.
 *                scratch_dev must hold b2_ae_grad_scratch_bytes(ae, n) bytes.  n <= 2^24 per call: the row count is a float of
 *                the gradient vector, and the all-reduced count must stay below 2^24 too to remain exact.
 *   b2_ae_apply: θ ← Adam(θ, grads / total_rows) using grads_dev after the caller's all-reduce. */
int64_t b2_ae_grad_scratch_bytes(const b2_ae* ae, int64_t n);
int  b2_ae_grad(b2_ae* ae, const float* x_dev, int64_t n, const b2_norm* norm, float* grads_dev,
                void* scratch_dev, int64_t scratch_bytes, void* stream);
int  b2_ae_apply(b2_ae* ae, const float* grads_dev, b2_adam* adam, float* loss_out_dev, void* stream);

/* The same step as ONE launch (model.fit of cardata-v3.py:220-222 run data-parallel): gradient sums of this rank's n rows, the
 * cross-GPU sum and the Adam update all happen inside the gradient kernel — its last CTA folds the partial sums, writes the
 * 840-float vector into every peer's exchange buffer with plain peer stores over NVLink, waits for the peers' flags and applies
 * the update from the rank-ordered sum (bit-identical weights on every GPU).  total_rows = rows of the global batch.
 * world == 1: no exchange (peer_bufs may be NULL).  Otherwise peer_bufs[world] are the device addresses, as mapped in THIS
 * process, of every rank's exchange buffer of b2_ae_exchange_bytes() bytes (symmetric memory, zero-filled once), and seq is the
 * step's sequence number — 1, 2, 3, … on all ranks alike.  scratch_dev: b2_ae_step_scratch_bytes(ae, n) bytes, zero-filled
 * once before the first step.  stats_dev f32[2] (or NULL): mean loss and accuracy of the global batch; NaN if a peer failed
 * to arrive within ~4 s (the step is then skipped rather than hanging the GPU).  Every rank must launch the step. */
int64_t b2_ae_step_scratch_bytes(const b2_ae* ae, int64_t n);
int64_t b2_ae_exchange_bytes(void);
int  b2_ae_step_fused(b2_ae* ae, const float* x_dev, int64_t n, const b2_norm* norm, b2_adam* adam, void* scratch_dev,
                      int64_t scratch_bytes, int64_t total_rows, const uint64_t* peer_bufs, int world, int rank, uint32_t seq,
                      float* stats_dev, void* stream);

/* No counterpart in the reference.  Same result as b2_ae_grad for callers that may not allocate device memory while
 * running (pre-budgeted memory pools, real-time loops that must stay off the allocator lock, hosts whose cudaMalloc is
 * owned by another component): the per-CTA partial sums go to a fixed buffer inside the library image instead of a
 * caller-supplied one, and the call itself performs no allocation of any kind.
 *   - Limit: the launch may need at most B2_NOALLOC_SCRATCH_FLOATS floats (CTAs used x (parameters + 3)); anything larger
 *     is refused with B2_EINVAL and a message — fall back to b2_ae_grad.  Every GPU that loads the library holds those
 *     B2_NOALLOC_SCRATCH_FLOATS * 4 bytes until exit.
 *   - The first call on each device resolves the buffer's device address (cudaGetSymbolAddress); make that first call
 *     before starting any stream capture.
 *   - All calls on a device share the one buffer and it keeps whatever the previous call wrote.  Nothing needs clearing
 *     (each launch overwrites every slot it later reads), but two calls that overlap on the GPU — different streams, or
 *     different b2_ae handles — race on it: serialise them. */
#define B2_NOALLOC_SCRATCH_FLOATS (160 * 840)
int  b2_ae_grad_noalloc(b2_ae* ae, const float* x_dev, int64_t n, const b2_norm* norm, float* grads_dev, void* stream);

/* Adam slots (for .h5 save / resume; layout like the weights) */
int  b2_ae_get_adam(b2_ae* ae, float* host_m, float* host_v);
int  b2_ae_set_adam(b2_ae* ae, const float* host_m, const float* host_v);

/* ---- LSTM ---------------------------------------------------------------------------------------------- */
typedef struct {
  int32_t features;                 /* 18 */
  int32_t units[4];                 /* 32, 16, 16, 32 (LSTM-TensorFlow-IO-Kafka/cardata-v1.py:170-175) */
} b2_lstm_desc;

/* host_weights: per LSTM layer kernel[in,4u], recurrent_kernel[u,4u], bias[4u] (gate order i,f,c,o), then the
 * TimeDistributed Dense kernel[32,features], bias[features]. */
int  b2_lstm_create(b2_ctx* ctx, const b2_lstm_desc* desc, const float* host_weights, b2_lstm** out);
void b2_lstm_destroy(b2_lstm* m);
int  b2_lstm_param_count(const b2_lstm* m);
/* Model.predict of the LSTM stack (LSTM-…/cardata-v1.py:226): x_dev f32[B,T,features] → y_dev f32[B,T,features]. */
int  b2_lstm_forward(b2_lstm* m, const float* x_dev, int64_t batch, int timesteps, float* y_dev, void* stream);
/* The script's input pipeline fused into the forward pass (LSTM-…/cardata-v1.py:184-185:
 * dataset.window(look_back, shift=1, drop_remainder=True).flat_map(lambda w: w.batch(look_back))): window i is rows
 * [i*shift, i*shift + look_back) of rows_dev f32[n_rows, features]; the kernel reads them in place, nothing is materialised.
 * y_dev f32[b2_lstm_window_count(...), look_back, features]. */
int64_t b2_lstm_window_count(int64_t n_rows, int look_back, int shift);
int  b2_lstm_forward_windows(b2_lstm* m, const float* rows_dev, int64_t n_rows, int look_back, int shift, float* y_dev, void* stream);

/* Training of the same stack (LSTM-TensorFlow-IO-Kafka/cardata-v1.py:178 compile(mean_squared_error, adam, ['accuracy']) and
 * :191-194 fit(zip((dataset_x, dataset_y)).batch(batch_size).take(1000), epochs=5)): consecutive batches of `batch` sequences,
 * one Adam step each, all inside ONE launch (one persistent CTA: the reference's batch_size 1 x look_back 1 makes this a chain of
 * dependent steps).  Sequence s, step t reads x_dev + (s * x_seq_stride + t) * F and is compared with y_dev + (s * y_seq_stride
 * + t) * F (strides in rows: `timesteps` for dense [n, T, F] tensors; 1 / 1 with y_dev = rows + F for the script's sliding
 * windows of look_back 1).  w_dev / m_dev / v_dev f32[P] (Keras flat weight order: kernel, recurrent_kernel, bias per LSTM, then
 * the Dense kernel and bias) are updated in place, adam->iter advances by the number of steps; loss_dev / acc_dev f32[steps]
 * (NULL to skip) receive each step's loss and accuracy.  scratch_dev: b2_lstm_fit_scratch_bytes(desc, timesteps, batch) bytes.
 * Afterwards build a new b2_lstm from the trained weights to run the forward kernels. */
int64_t b2_lstm_fit_scratch_bytes(const b2_lstm_desc* desc, int timesteps, int batch);
int  b2_lstm_fit_steps(b2_ctx* ctx, const b2_lstm_desc* desc, float* w_dev, float* m_dev, float* v_dev, const float* x_dev,
                       int64_t x_seq_stride, const float* y_dev, int64_t y_seq_stride, int64_t n_seq, int timesteps, int batch,
                       b2_adam* adam, float* loss_dev, float* acc_dev, void* scratch_dev, int64_t scratch_bytes, void* stream);

/* ---- result sink: rows → text ------------------------------------------------------------------------------
 * Replaces the formatting half of OutputCallback.on_predict_batch_end (cardata-v1.py:218-224:
 * `message = np.array2string(output)` per reconstructed row, then KafkaOutputSequence.setitem(index, message)).
 * x_dev f32[n, n_cols] (n_cols ≤ 64) → the exact text NumPy's array2string gives for each row with default print options.
 * numpy_compat picks the NumPy generation to reproduce: 0 = NumPy ≥ 2.3, 1 = NumPy 1.21 … 2.2, 2 = NumPy 1.14 … 1.20 (the
 * generation the reference's tensorflow:2.0.0 images carried; differs only in rows printed in scientific notation and for
 * integers ≥ 2^24).  Texts are packed back to back into text_dev (capacity text_cap bytes; b2_format_rows_bound(n_cols) per
 * row always suffices) in tile order, NOT row order: row i is text_dev[offs_dev[i] : offs_dev[i] + len_dev[i]].
 * cursor_dev: uint64[2], zeroed by the caller; [0] receives the bytes used, [1] the number of rows dropped because the arena
 * was full (their len_dev is -1). */
int64_t b2_format_rows_bound(int n_cols);
int  b2_format_rows(b2_ctx* ctx, const float* x_dev, int64_t n, int n_cols, int numpy_compat, uint8_t* text_dev, int64_t text_cap,
                    int64_t* offs_dev, int32_t* len_dev, unsigned long long* cursor_dev, void* stream);

/* ---- Kafka record-batch framing (message format v2) ------------------------------------------
 * What tensorflow-io's KafkaDataset / librdkafka does in front of the reference's scripts (cardata-v1.py:10): fetch-response or
 * log-segment bytes → message values.  The caller walks the batch chain on the host (baseOffset i64, batchLength i32: the next
 * batch starts 12 + batchLength bytes on; recordCount i32 at byte 57 — b2_kafka_scan_batches does it on a host copy of the bytes:
 * returns the number of batches, writes the first `cap` positions / record counts, so cap = 0 sizes the arrays; a trailing partial
 * batch is dropped like consumers do) and passes batch_pos_dev int64[n_batches] and rec_base_dev int64[n_batches] (records in the
 * batches before each one).  One warp per batch checks magic == 2, the CRC-32C (verify_crc) and walks the records: begin_dev /
 * end_dev int32[n_records] receive each value's byte range inside bytes_dev, status_dev int32[n_batches] a B2_FRAME_* code (every
 * record a refused batch's header announces gets an EMPTY range placed at the batch's position, so the ranges of a segment whose
 * batch positions ascend are in ascending order; record, key and value lengths are untrusted input and are checked against the
 * bytes that remain).
 * Those ranges go straight into b2_decode_score_spans (below): the segment itself is the decode input and no value is copied.
 * b2_kafka_gather_values instead packs the values back to back at out_offs_dev int32[n] (the caller's scan of end - begin),
 * giving the (bytes, offsets) batch every other decode entry point takes. */
#define B2_FRAME_OK         0
#define B2_FRAME_BAD_MAGIC  1  /* not message format v2 */
#define B2_FRAME_BAD_CRC    2
#define B2_FRAME_TRUNCATED  3  /* batch or record runs past its container */
#define B2_FRAME_COMPRESSED 4  /* compressed batch handed to b2_kafka_index_records (inflate it first), or a codec this library has no decoder for */
#define B2_FRAME_CONTROL    5  /* control batch (attributes bit 5: transaction markers); consumers never see its records */
int64_t b2_kafka_scan_batches(const uint8_t* bytes_host, int64_t n_bytes, int64_t* batch_pos, int64_t* rec_count, int64_t cap);
int  b2_kafka_index_records(b2_ctx* ctx, const uint8_t* bytes_dev, int64_t n_bytes, const int64_t* batch_pos_dev, const int64_t* rec_base_dev,
                            int64_t n_batches, int verify_crc, int32_t* begin_dev, int32_t* end_dev, int32_t* status_dev, void* stream);
int  b2_kafka_gather_values(b2_ctx* ctx, const uint8_t* bytes_dev, const int32_t* begin_dev, const int32_t* end_dev, int64_t n,
                            const int32_t* out_offs_dev, uint8_t* out_bytes_dev, void* stream);
/* Compressed batches (attributes & 7: 1 gzip, 2 Snappy — raw or the Java client's xerial framing —, 3 LZ4 frame, 4 Zstandard
 * frames without dictionary): what librdkafka
 * inflates underneath KafkaDataset whatever the producer's compression.type was (cardata-v1.py:10).  Two passes, one thread per
 * batch.  b2_kafka_inflate_sizes: status_dev int32[n_batches] (header checks, CRC-32C over the bytes as stored when verify_crc;
 * an unknown codec → B2_FRAME_COMPRESSED, a damaged stream → B2_FRAME_TRUNCATED) and sizes_dev int64[n_batches] = bytes of each batch's
 * records section once inflated (as stored for an uncompressed batch, 0 for a refused one).  The caller lays the rebuilt segment
 * out: out_pos[b+1] = out_pos[b] + 61 + sizes[b].  b2_kafka_inflate writes every batch there as an UNCOMPRESSED batch (header
 * copied, codec bits cleared, batchLength adjusted, records inflated); the result goes through b2_kafka_index_records with
 * verify_crc = 0 (the stored CRC belongs to the compressed bytes and was checked in the first pass) — blank the records of the
 * batches whose first-pass status is not B2_FRAME_OK. */
int  b2_kafka_inflate_sizes(b2_ctx* ctx, const uint8_t* bytes_dev, int64_t n_bytes, const int64_t* batch_pos_dev, int64_t n_batches,
                            int verify_crc, int64_t* sizes_dev, int32_t* status_dev, void* stream);
int  b2_kafka_inflate(b2_ctx* ctx, const uint8_t* bytes_dev, int64_t n_bytes, const int64_t* batch_pos_dev, int64_t n_batches,
                      const int64_t* sizes_dev, const int64_t* out_pos_dev, int32_t* status_dev, uint8_t* out_dev, int64_t out_bytes,
                      void* stream);

/* ---- evaluation metrics ----------------------------------------------------------------------
 * The threshold sweep behind sklearn.metrics.roc_curve / precision_recall_curve as the notebooks call them on the
 * reconstruction error (nb2.0 cells 23-24): scores sorted in decreasing order, one threshold per distinct score,
 * tps[r] / fps[r] = positives / negatives with score >= thresholds[r] (scikit-learn's confusion_matrix_at_thresholds).
 * score_dev f32[n], y_dev u8[n] (non-zero = positive class); outputs sized n; n_thresholds_dev int64[2]: [0] number of
 * thresholds written, [1] number of NaN scores (the result is meaningless then; scikit-learn refuses NaN input).
 * scratch_dev: b2_threshold_sweep_scratch_bytes(n) bytes, caller-owned.  Exact integer work (bitonic sort + scan). */
int64_t b2_threshold_sweep_scratch_bytes(int64_t n);
int  b2_threshold_sweep(b2_ctx* ctx, const float* score_dev, const uint8_t* y_dev, int64_t n, float* thresholds_dev, int64_t* tps_dev,
                        int64_t* fps_dev, int64_t* n_thresholds_dev, void* scratch_dev, int64_t scratch_bytes, void* stream);
/* sklearn.metrics.confusion_matrix(y_true, pred_y) of nb2.0 cell 27 for two classes: counts_dev int64[4] = {tn, fp, fn, tp}. */
int  b2_confusion_matrix(b2_ctx* ctx, const uint8_t* y_true_dev, const uint8_t* y_pred_dev, int64_t n, int64_t* counts_dev, void* stream);

/* ---- per-GPU partition driver --------------------------------------------------------------------------- */
typedef struct {
  const uint8_t* bytes_dev;         /* device-resident message payloads of one Kafka partition batch */
  const int32_t* offs_dev;          /* int32[n+1] */
  int64_t        n;
  float*         mse_dev;           /* [n] */
  uint8_t*       flag_dev;          /* [n] */
} b2_partition;

typedef struct {
  unsigned long long records, flagged, decode_errors, bytes;
  float elapsed_ms;                 /* CUDA-event time of the whole run on this GPU */
} b2_run_stats;

/* One process per Kafka partition set (cardata-v1.py:10 reads "<topic>:0:<offset>"; 10 partitions per topic,
 * infrastructure/confluent/01_installConfluentPlatform.sh:180): runs decode+score over this GPU's partitions on
 * `n_streams` CUDA streams and returns the counters.  The cross-GPU count reduction is the caller's
 * (torch.distributed / NCCL) because the communicator lives there. counters_dev u64[4] is accumulated too. */
int  b2_partition_run(b2_ae* ae, const b2_partition* parts, int n_parts, const b2_csv_spec* csv,
                      const b2_avro_prog* avro, int skip_prefix, const b2_norm* norm, float threshold,
                      int n_streams, unsigned long long* counters_dev, b2_run_stats* host_out);
/* The same run `repeat` times over the partition list inside one call (a consumer that keeps polling the same partitions;
 * what bench.py times: no host work between launches but the launches themselves).  Streams and events belong to the
 * context: created on first use, reused by every run. */
int  b2_partition_run_repeat(b2_ae* ae, const b2_partition* parts, int n_parts, const b2_csv_spec* csv,
                             const b2_avro_prog* avro, int skip_prefix, const b2_norm* norm, float threshold,
                             int n_streams, int repeat, unsigned long long* counters_dev, b2_run_stats* host_out);

#ifdef __cplusplus
}
#endif
#endif /* B2_API_H */
