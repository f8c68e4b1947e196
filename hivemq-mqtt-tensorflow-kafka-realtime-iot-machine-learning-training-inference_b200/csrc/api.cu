// Context, error reporting and the autoencoder handle (weights upload / read-back).
#include <stdarg.h>
#include <stdlib.h>

#include "common.cuh"

static thread_local char g_err[512] = "";

void b2_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

extern "C" const char* b2_last_error(void) { return g_err; }

extern "C" const char* b2_strerror(int code) {
  switch (code) {
    case B2_OK: return "ok";
    case B2_EINVAL: return "invalid argument";
    case B2_ECUDA: return "CUDA runtime error";
    case B2_EUNSUPPORTED: return "unsupported";
    case B2_ENOMEM: return "out of memory";
    default: return "unknown error";
  }
}

extern "C" const char* b2_version(void) { return "b2sensor 0.1 (sm_100a)"; }

namespace {
struct KnobName {
  const char* name;
  const char* env;
  int b2_knobs::*field;
};
const KnobName kKnobs[] = {
    {"fast_tanh", "B2_FAST_TANH", &b2_knobs::fast_tanh},       {"score_cfg", "B2_SCORE_CFG", &b2_knobs::score_cfg},
    {"no_bulk", "B2_NO_BULK", &b2_knobs::no_bulk},             {"force_generic", "B2_FORCE_GENERIC", &b2_knobs::force_generic},
    {"avro_generic", "B2_AVRO_GENERIC", &b2_knobs::avro_generic}, {"lstm_tc", "B2_LSTM_TC", &b2_knobs::lstm_tc},
    {"score_tc", "B2_SCORE_TC", &b2_knobs::score_tc},          {"tc_ctas", "B2_TC_CTAS", &b2_knobs::tc_ctas},
    {"decode_r2", "B2_DECODE_R2", &b2_knobs::decode_r2},       {"decode_tc", "B2_DECODE_TC", &b2_knobs::decode_tc},
    {"decode_mma", "B2_DECODE_MMA", &b2_knobs::decode_mma},   {"zero_skip", "B2_ZERO_SKIP", &b2_knobs::zero_skip},
};
}  // namespace

extern "C" int b2_ctx_create(int device, b2_ctx** out) {
  B2_REQUIRE(out != nullptr, "b2_ctx_create: out is NULL");
  int count = 0;
  B2_CUDA(cudaGetDeviceCount(&count));
  B2_REQUIRE(device >= 0 && device < count, "b2_ctx_create: device %d out of range (have %d)", device, count);
  cudaDeviceProp prop;
  B2_CUDA(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10) {
    b2_set_error("b2_ctx_create: device %d is sm_%d%d; this library only carries sm_100a code (B200)", device, prop.major,
                 prop.minor);
    return B2_EUNSUPPORTED;
  }
  b2_ctx* c = new b2_ctx();
  c->device = device;
  c->sm_count = prop.multiProcessorCount;
  c->smem_optin = (int)prop.sharedMemPerBlockOptin;
  for (const KnobName& k : kKnobs) {
    const char* v = getenv(k.env);
    if (v && *v) c->knobs.*(k.field) = atoi(v);
  }
  {
    DeviceGuard g(device);
    const size_t bytes = sizeof(int32_t) * (size_t)B2_MAX_STREAM_SLOTS * (1 + B2_OVERSIZE_CAP);
    cudaError_t e = cudaMalloc(&c->oversize_pool, bytes);
    if (e == cudaSuccess) e = cudaMemset(c->oversize_pool, 0, bytes);
    if (e != cudaSuccess) {
      b2_set_error("b2_ctx_create: %s", cudaGetErrorString(e));
      if (c->oversize_pool) cudaFree(c->oversize_pool);
      delete c;
      return B2_ECUDA;
    }
  }
  *out = c;
  return B2_OK;
}

extern "C" void b2_ctx_destroy(b2_ctx* ctx) {
  if (!ctx) return;
  DeviceGuard g(ctx->device);
  for (int i = 0; i < ctx->n_run_streams; ++i) {
    cudaStreamDestroy(ctx->run_streams[i]);
    cudaEventDestroy(ctx->run_done[i]);
  }
  if (ctx->run_start) cudaEventDestroy(ctx->run_start);
  if (ctx->run_stop) cudaEventDestroy(ctx->run_stop);
  if (ctx->oversize_pool) cudaFree(ctx->oversize_pool);
  delete ctx;
}

int32_t* b2_ctx_oversize_slot(b2_ctx* ctx, cudaStream_t stream) {
  int i = 0;
  for (; i < ctx->n_stream_slots; ++i)
    if (ctx->slot_stream[i] == stream) break;
  if (i == ctx->n_stream_slots) {
    if (ctx->n_stream_slots < B2_MAX_STREAM_SLOTS) ctx->slot_stream[ctx->n_stream_slots++] = stream;
    else i = B2_MAX_STREAM_SLOTS - 1;  // more distinct streams than slots: the surplus share the last one (see b2_api.h)
  }
  return ctx->oversize_pool + (size_t)i * (1 + B2_OVERSIZE_CAP);
}

extern "C" int b2_ctx_set_option(b2_ctx* ctx, const char* name, int value) {
  B2_REQUIRE(ctx && name, "b2_ctx_set_option: NULL argument");
  for (const KnobName& k : kKnobs)
    if (!strcmp(name, k.name)) {
      ctx->knobs.*(k.field) = value;
      return B2_OK;
    }
  b2_set_error("b2_ctx_set_option: unknown option '%s'", name);
  return B2_EINVAL;
}

extern "C" int b2_ctx_get_option(const b2_ctx* ctx, const char* name, int* value) {
  B2_REQUIRE(ctx && name && value, "b2_ctx_get_option: NULL argument");
  for (const KnobName& k : kKnobs)
    if (!strcmp(name, k.name)) {
      *value = ctx->knobs.*(k.field);
      return B2_OK;
    }
  b2_set_error("b2_ctx_get_option: unknown option '%s'", name);
  return B2_EINVAL;
}
extern "C" int b2_ctx_sm_count(const b2_ctx* ctx) { return ctx ? ctx->sm_count : 0; }

extern "C" int b2_ctx_set_record_bytes_hint(b2_ctx* ctx, int avg_record_bytes) {
  B2_REQUIRE(ctx, "b2_ctx_set_record_bytes_hint: NULL context");
  B2_REQUIRE(avg_record_bytes >= 0, "b2_ctx_set_record_bytes_hint: negative size");
  ctx->record_bytes_hint = avg_record_bytes;
  return B2_OK;
}

// ---- autoencoder handle ------------------------------------------------------------------------------------
static int count_params(const b2_ae_desc& d) {
  int n = 0;
  for (int l = 0; l < d.n_layers; ++l) n += d.widths[l] * d.widths[l + 1] + d.widths[l + 1];
  return n;
}

extern "C" int b2_ae_create(b2_ctx* ctx, const b2_ae_desc* desc, const float* host_weights, b2_ae** out) {
  B2_REQUIRE(ctx && desc && host_weights && out, "b2_ae_create: NULL argument");
  B2_REQUIRE(desc->n_layers >= 1 && desc->n_layers <= 8, "b2_ae_create: n_layers=%d not in 1..8", desc->n_layers);
  for (int l = 0; l <= desc->n_layers; ++l)
    B2_REQUIRE(desc->widths[l] >= 1 && desc->widths[l] <= B2_MAX_WIDTH, "b2_ae_create: width[%d]=%d not in 1..%d", l,
               desc->widths[l], B2_MAX_WIDTH);
  for (int l = 0; l < desc->n_layers; ++l)
    B2_REQUIRE(desc->act[l] >= B2_ACT_LINEAR && desc->act[l] <= B2_ACT_RELU, "b2_ae_create: activation %d of layer %d unknown",
               desc->act[l], l);
  B2_REQUIRE(desc->widths[desc->n_layers] == desc->widths[0], "b2_ae_create: autoencoder output width %d != input width %d",
             desc->widths[desc->n_layers], desc->widths[0]);
  DeviceGuard g(ctx->device);
  b2_ae* ae = new b2_ae();
  ae->ctx = ctx;
  ae->desc = *desc;
  ae->n_params = count_params(*desc);
  ae->h_w.assign(host_weights, host_weights + ae->n_params);
  ae->d_w = ae->d_m = ae->d_v = nullptr;
  ae->host_stale = false;
  const b2_ae_desc& d = *desc;
  const bool shape_ok = d.n_layers == 4 && d.widths[1] == 14 && d.widths[2] == 7 && d.widths[3] == 7 &&
                        d.act[0] == B2_ACT_TANH && d.act[1] == B2_ACT_RELU && d.act[2] == B2_ACT_TANH && d.act[3] == B2_ACT_RELU;
  ae->spec_d = (shape_ok && (d.widths[0] == 18 || d.widths[0] == 30)) ? d.widths[0] : 0;
  cudaError_t e = cudaMalloc(&ae->d_w, sizeof(float) * ae->n_params);
  if (e == cudaSuccess) e = cudaMemcpy(ae->d_w, host_weights, sizeof(float) * ae->n_params, cudaMemcpyHostToDevice);
  if (e != cudaSuccess) {
    b2_set_error("b2_ae_create: weight upload failed: %s", cudaGetErrorString(e));
    if (ae->d_w) cudaFree(ae->d_w);
    delete ae;
    return B2_ECUDA;
  }
  *out = ae;
  return B2_OK;
}

extern "C" void b2_ae_destroy(b2_ae* ae) {
  if (!ae) return;
  DeviceGuard g(ae->ctx->device);
  if (ae->d_w) cudaFree(ae->d_w);
  if (ae->d_m) cudaFree(ae->d_m);
  if (ae->d_v) cudaFree(ae->d_v);
  if (ae->d_tc) cudaFree(ae->d_tc);
  delete ae;
}

extern "C" int b2_ae_param_count(const b2_ae* ae) { return ae ? ae->n_params : 0; }

int b2_ae_sync_host(b2_ae* ae) {
  if (!ae->host_stale) return B2_OK;
  DeviceGuard g(ae->ctx->device);
  B2_CUDA(cudaDeviceSynchronize());
  B2_CUDA(cudaMemcpy(ae->h_w.data(), ae->d_w, sizeof(float) * ae->n_params, cudaMemcpyDeviceToHost));
  ae->host_stale = false;
  return B2_OK;
}

extern "C" int b2_ae_get_weights(b2_ae* ae, float* host_weights) {
  B2_REQUIRE(ae && host_weights, "b2_ae_get_weights: NULL argument");
  int rc = b2_ae_sync_host(ae);
  if (rc) return rc;
  memcpy(host_weights, ae->h_w.data(), sizeof(float) * ae->n_params);
  return B2_OK;
}

extern "C" int b2_ae_set_weights(b2_ae* ae, const float* host_weights) {
  B2_REQUIRE(ae && host_weights, "b2_ae_set_weights: NULL argument");
  DeviceGuard g(ae->ctx->device);
  ae->h_w.assign(host_weights, host_weights + ae->n_params);
  B2_CUDA(cudaMemcpy(ae->d_w, host_weights, sizeof(float) * ae->n_params, cudaMemcpyHostToDevice));
  ae->host_stale = false;
  ae->tc_valid = false;
  return B2_OK;
}

static int ensure_adam(b2_ae* ae) {
  if (ae->d_m) return B2_OK;
  B2_CUDA(cudaMalloc(&ae->d_m, sizeof(float) * ae->n_params));
  B2_CUDA(cudaMalloc(&ae->d_v, sizeof(float) * ae->n_params));
  B2_CUDA(cudaMemset(ae->d_m, 0, sizeof(float) * ae->n_params));
  B2_CUDA(cudaMemset(ae->d_v, 0, sizeof(float) * ae->n_params));
  return B2_OK;
}
int b2_ae_ensure_adam(b2_ae* ae) {
  DeviceGuard g(ae->ctx->device);
  return ensure_adam(ae);
}

extern "C" int b2_ae_get_adam(b2_ae* ae, float* host_m, float* host_v) {
  B2_REQUIRE(ae && host_m && host_v, "b2_ae_get_adam: NULL argument");
  DeviceGuard g(ae->ctx->device);
  int rc = ensure_adam(ae);
  if (rc) return rc;
  B2_CUDA(cudaDeviceSynchronize());
  B2_CUDA(cudaMemcpy(host_m, ae->d_m, sizeof(float) * ae->n_params, cudaMemcpyDeviceToHost));
  B2_CUDA(cudaMemcpy(host_v, ae->d_v, sizeof(float) * ae->n_params, cudaMemcpyDeviceToHost));
  return B2_OK;
}

extern "C" int b2_ae_set_adam(b2_ae* ae, const float* host_m, const float* host_v) {
  B2_REQUIRE(ae && host_m && host_v, "b2_ae_set_adam: NULL argument");
  DeviceGuard g(ae->ctx->device);
  int rc = ensure_adam(ae);
  if (rc) return rc;
  B2_CUDA(cudaMemcpy(ae->d_m, host_m, sizeof(float) * ae->n_params, cudaMemcpyHostToDevice));
  B2_CUDA(cudaMemcpy(ae->d_v, host_v, sizeof(float) * ae->n_params, cudaMemcpyHostToDevice));
  return B2_OK;
}
