// K4 — forward pass of the reference LSTM stack
//   LSTM(32, relu, seq) → LSTM(16, relu) → RepeatVector(T) → LSTM(16, relu, seq) → LSTM(32, relu, seq) → TimeDistributed(Dense(F))
//   (python-scripts/LSTM-TensorFlow-IO-Kafka/cardata-v1.py:169-176; Model.predict :226)
// with tf.keras LSTM semantics: gate order i,f,c,o; z = x_t·W + h·U + b; c = σ(z_f)·c + σ(z_i)·relu(z_c);
// h = σ(z_o)·relu(c); h0 = c0 = 0.
//
// This is a genuine dense contraction (36 480 FLOP per record-step), so it is organised as a shared-memory tiled GEMM
// that never leaves the SM: a CTA owns 128 sequences; all 18 642 weights (re-laid-out as one [K, 4u] matrix per layer with
// the four gates of a unit adjacent) and the activation rows [x_t | h1 | h2 | h3 | h4] (k-major, 128 sequences wide) stay in
// shared memory for the whole time loop; each of the 256 threads keeps an 8-sequence × (4u/16)-column accumulator tile
// and the matching cell states in registers; the gate non-linearities are the GEMM epilogue.  Because consecutive layers'
// inputs are adjacent activation rows, every layer's GEMM reads ONE contiguous [K, 128] panel.
// Algorithmic bytes per record-step: 4F read + 4F written (weights are read once per CTA).
#include "lstm_tc.h"

constexpr int LS_SEQ = 128;   // sequences per CTA
constexpr int LS_NT = 256;    // threads: 16 (sequence groups of 8) x 16 (column groups)
constexpr int LS_MAXF = 32;   // feature count limit of this kernel (reference: 18)

struct b2_lstm {
  b2_ctx* ctx;
  b2_lstm_desc desc;
  int n_params;
  float* d_packed;  // device: per layer [K][N] (gate-interleaved columns) + bias[N], then dense [K][32] + bias[32]
  int packed_floats;
  // tensor-core path (lstm_tc.cu): canonical-layout (hi, lo) weights; tc_floats == 0 when the stack does not fit it
  float* d_packed_tc;
  int tc_floats;
  TcLayer tc_layers[5];
};

struct LstmArgs {
  const float* x;
  float* y;
  long long batch;
  long long seq_stride;  // rows between sequence starts (T, or the window shift)
  int T, F;
  int u1, u2, u3, u4;
  const float* packed;
  int packed_floats;
};

__device__ __forceinline__ float sigmoidf_fast(float x) {
  float t;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(t) : "f"(x * -1.4426950408889634f));
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(1.0f + t));
  return r;
}

// acc[s][c] = bias[col] + Σ_k As[k][seq] * Ws[k][col]   for this thread's 8 sequences × TN columns
template <int TN>
__device__ __forceinline__ void tile_gemm(const float* __restrict__ As, const float* __restrict__ Ws, const float* __restrict__ bias, int K,
                                          int N, int ty, int tx, float (&acc)[8][TN]) {
#pragma unroll
  for (int c = 0; c < TN; ++c) {
    const float b = bias[tx * TN + c];
#pragma unroll
    for (int s = 0; s < 8; ++s) acc[s][c] = b;
  }
  const float* ap = As + ty * 8;
  const float* wp = Ws + tx * TN;
#pragma unroll 2
  for (int k = 0; k < K; ++k) {
    const float4 a0 = *reinterpret_cast<const float4*>(ap + k * LS_SEQ);
    const float4 a1 = *reinterpret_cast<const float4*>(ap + k * LS_SEQ + 4);
    const float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
    float w[TN];
    if (TN % 4 == 0) {
#pragma unroll
      for (int c = 0; c < TN; c += 4) {
        const float4 wv = *reinterpret_cast<const float4*>(wp + k * N + c);
        w[c] = wv.x, w[c + 1] = wv.y, w[c + 2] = wv.z, w[c + 3] = wv.w;
      }
    } else {
#pragma unroll
      for (int c = 0; c < TN; ++c) w[c] = wp[k * N + c];
    }
#pragma unroll
    for (int s = 0; s < 8; ++s)
#pragma unroll
      for (int c = 0; c < TN; ++c) acc[s][c] = fmaf(a[s], w[c], acc[s][c]);
  }
}

// LSTM epilogue: this thread's TN/4 units × 8 sequences; writes h into the activation rows h_rows[unit][seq]
template <int TN>
__device__ __forceinline__ void lstm_gates(const float (&acc)[8][TN], float (&c)[8][TN / 4], float* __restrict__ h_rows, int ty, int tx) {
  constexpr int UPT = TN / 4;
#pragma unroll
  for (int q = 0; q < UPT; ++q) {
    const int unit = tx * UPT + q;
    float hv[8];
#pragma unroll
    for (int s = 0; s < 8; ++s) {
      const float ig = sigmoidf_fast(acc[s][4 * q]);
      const float fg = sigmoidf_fast(acc[s][4 * q + 1]);
      const float gg = fmaxf(acc[s][4 * q + 2], 0.f);
      const float og = sigmoidf_fast(acc[s][4 * q + 3]);
      c[s][q] = fmaf(fg, c[s][q], ig * gg);
      hv[s] = og * fmaxf(c[s][q], 0.f);
    }
    float* dst = h_rows + unit * LS_SEQ + ty * 8;
    *reinterpret_cast<float4*>(dst) = make_float4(hv[0], hv[1], hv[2], hv[3]);
    *reinterpret_cast<float4*>(dst + 4) = make_float4(hv[4], hv[5], hv[6], hv[7]);
  }
}

// Specialised for the reference's unit counts (32, 16, 16, 32): thread tiles 8x8, 8x4, 8x4, 8x8, dense 8x2.
__global__ void __launch_bounds__(LS_NT, 1) lstm_forward_kernel(const __grid_constant__ LstmArgs a) {
  extern __shared__ __align__(16) float lsm[];
  const int F = a.F;
  const int K1 = F + 32, K2 = 32 + 16, K3 = 16 + 16, K4 = 16 + 32, K5 = 32;
  // packed weights: W1[K1][128] b1[128] W2[K2][64] b2[64] W3[K3][64] b3[64] W4[K4][128] b4[128] W5[K5][32] b5[32]
  float* W1 = lsm;
  float* B1 = W1 + K1 * 128;
  float* W2 = B1 + 128;
  float* B2 = W2 + K2 * 64;
  float* W3 = B2 + 64;
  float* B3 = W3 + K3 * 64;
  float* W4 = B3 + 64;
  float* B4 = W4 + K4 * 128;
  float* W5 = B4 + 128;
  float* B5 = W5 + K5 * 32;
  float* act = B5 + 32;                 // activation rows, k-major: [F + 32 + 16 + 16 + 32][128]
  float* xr = act;                      // x_t        rows [0, F)
  float* h1 = xr + F * LS_SEQ;          // h1         32 rows
  float* h2 = h1 + 32 * LS_SEQ;         // h2         16 rows
  float* h3 = h2 + 16 * LS_SEQ;         // h3         16 rows
  float* h4 = h3 + 16 * LS_SEQ;         // h4         32 rows
  const int tid = threadIdx.x, ty = tid >> 4, tx = tid & 15;

  for (int i = tid; i < a.packed_floats; i += LS_NT) lsm[i] = a.packed[i];
  const long long n_tiles = (a.batch + LS_SEQ - 1) / LS_SEQ;
  for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const long long b0 = tile * LS_SEQ;
    const int nseq = (int)min((long long)LS_SEQ, a.batch - b0);
    __syncthreads();
    for (int i = tid; i < (32 + 16 + 16 + 32) * LS_SEQ; i += LS_NT) h1[i] = 0.f;  // h0 = 0 for every layer
    float c1[8][2], c2[8][1], c3[8][1], c4[8][2];
#pragma unroll
    for (int s = 0; s < 8; ++s) c1[s][0] = c1[s][1] = c2[s][0] = c3[s][0] = c4[s][0] = c4[s][1] = 0.f;

    auto load_x = [&](int t) {
      for (int i = tid; i < LS_SEQ * F; i += LS_NT) {
        const int seq = i / F, f = i - seq * F;
        xr[f * LS_SEQ + seq] = seq < nseq ? __ldg(a.x + ((b0 + seq) * a.seq_stride + t) * F + f) : 0.f;
      }
    };
    // ---- phase 1: encoder LSTMs over the whole sequence ----
    for (int t = 0; t < a.T; ++t) {
      load_x(t);
      __syncthreads();
      {
        float acc[8][8];
        tile_gemm<8>(xr, W1, B1, K1, 128, ty, tx, acc);
        __syncthreads();  // all reads of h1(t-1) done
        lstm_gates<8>(acc, c1, h1, ty, tx);
      }
      __syncthreads();
      {
        float acc[8][4];
        tile_gemm<4>(h1, W2, B2, K2, 64, ty, tx, acc);
        __syncthreads();
        lstm_gates<4>(acc, c2, h2, ty, tx);
      }
      __syncthreads();
    }
    // ---- phase 2: RepeatVector(h2_T) → decoder LSTMs → TimeDistributed(Dense) ----
    for (int t = 0; t < a.T; ++t) {
      {
        float acc[8][4];
        tile_gemm<4>(h2, W3, B3, K3, 64, ty, tx, acc);  // rows [h2 | h3] are adjacent
        __syncthreads();
        lstm_gates<4>(acc, c3, h3, ty, tx);
      }
      __syncthreads();
      {
        float acc[8][8];
        tile_gemm<8>(h3, W4, B4, K4, 128, ty, tx, acc);  // rows [h3 | h4]
        __syncthreads();
        lstm_gates<8>(acc, c4, h4, ty, tx);
      }
      __syncthreads();
      {
        float acc[8][2];
        tile_gemm<2>(h4, W5, B5, K5, 32, ty, tx, acc);
#pragma unroll
        for (int s = 0; s < 8; ++s) {
          const int seq = ty * 8 + s;
#pragma unroll
          for (int c = 0; c < 2; ++c) {
            const int col = tx * 2 + c;
            if (seq < nseq && col < F) a.y[((b0 + seq) * a.T + t) * F + col] = acc[s][c];
          }
        }
      }
      // next iteration's first GEMM reads h3 (written above, synced) and writes h3 only after its own barrier
    }
  }
}

// ---- host side -------------------------------------------------------------------------------------------------
static int lstm_param_count(const b2_lstm_desc& d) {
  int n = 0, in = d.features;
  for (int l = 0; l < 4; ++l) {
    n += in * 4 * d.units[l] + d.units[l] * 4 * d.units[l] + 4 * d.units[l];
    in = d.units[l];
  }
  return n + d.units[3] * d.features + d.features;
}

extern "C" int b2_lstm_create(b2_ctx* ctx, const b2_lstm_desc* desc, const float* host_weights, b2_lstm** out) {
  B2_REQUIRE(ctx && desc && host_weights && out, "b2_lstm_create: NULL argument");
  B2_REQUIRE(desc->features >= 1 && desc->features <= LS_MAXF, "b2_lstm_create: features=%d not in 1..%d", desc->features, LS_MAXF);
  if (!(desc->units[0] == 32 && desc->units[1] == 16 && desc->units[2] == 16 && desc->units[3] == 32)) {
    b2_set_error("b2_lstm_create: only the reference stack LSTM(32)-LSTM(16)-LSTM(16)-LSTM(32) is implemented");
    return B2_EUNSUPPORTED;
  }
  DeviceGuard g(ctx->device);
  const int F = desc->features;
  const int in_dim[4] = {F, 32, 16, 16};
  const int u[4] = {32, 16, 16, 32};
  std::vector<float> packed;
  const float* p = host_weights;
  for (int l = 0; l < 4; ++l) {
    const int K = in_dim[l] + u[l], N = 4 * u[l];
    const float* W = p;                       // [in][4u], column g*u + unit
    const float* U = W + in_dim[l] * N;       // [u][4u]
    const float* b = U + u[l] * N;            // [4u]
    p = b + N;
    std::vector<float> M((size_t)K * N + N);
    for (int k = 0; k < K; ++k) {
      const float* row = k < in_dim[l] ? W + (size_t)k * N : U + (size_t)(k - in_dim[l]) * N;
      for (int unit = 0; unit < u[l]; ++unit)
        for (int gate = 0; gate < 4; ++gate) M[(size_t)k * N + unit * 4 + gate] = row[gate * u[l] + unit];
    }
    for (int unit = 0; unit < u[l]; ++unit)
      for (int gate = 0; gate < 4; ++gate) M[(size_t)K * N + unit * 4 + gate] = b[gate * u[l] + unit];
    packed.insert(packed.end(), M.begin(), M.end());
  }
  {  // TimeDistributed(Dense(F)): [32][F] → [32][32] zero padded, bias[32]
    std::vector<float> M(32 * 32 + 32, 0.f);
    for (int k = 0; k < 32; ++k)
      for (int j = 0; j < F; ++j) M[k * 32 + j] = p[k * F + j];
    for (int j = 0; j < F; ++j) M[32 * 32 + j] = p[32 * F + j];
    packed.insert(packed.end(), M.begin(), M.end());
  }
  b2_lstm* m = new b2_lstm();
  m->ctx = ctx;
  m->desc = *desc;
  m->n_params = lstm_param_count(*desc);
  m->packed_floats = (int)packed.size();
  m->d_packed = nullptr;
  m->d_packed_tc = nullptr;
  m->tc_floats = 0;
  cudaError_t e = cudaMalloc(&m->d_packed, sizeof(float) * packed.size());
  if (e == cudaSuccess) e = cudaMemcpy(m->d_packed, packed.data(), sizeof(float) * packed.size(), cudaMemcpyHostToDevice);
  std::vector<float> tc_packed;
  if (e == cudaSuccess && b2_lstm_tc_pack(desc, host_weights, tc_packed, m->tc_layers) == 0) {
    e = cudaMalloc(&m->d_packed_tc, sizeof(float) * tc_packed.size());
    if (e == cudaSuccess) e = cudaMemcpy(m->d_packed_tc, tc_packed.data(), sizeof(float) * tc_packed.size(), cudaMemcpyHostToDevice);
    m->tc_floats = (int)tc_packed.size();
  }
  if (e != cudaSuccess) {
    b2_set_error("b2_lstm_create: weight upload failed: %s", cudaGetErrorString(e));
    if (m->d_packed) cudaFree(m->d_packed);
    if (m->d_packed_tc) cudaFree(m->d_packed_tc);
    delete m;
    return B2_ECUDA;
  }
  *out = m;
  return B2_OK;
}

extern "C" void b2_lstm_destroy(b2_lstm* m) {
  if (!m) return;
  DeviceGuard g(m->ctx->device);
  if (m->d_packed) cudaFree(m->d_packed);
  if (m->d_packed_tc) cudaFree(m->d_packed_tc);
  delete m;
}

extern "C" int b2_lstm_param_count(const b2_lstm* m) { return m ? m->n_params : 0; }

static int lstm_launch(b2_lstm* m, const float* x_dev, int64_t batch, int timesteps, int64_t seq_stride, float* y_dev, void* stream) {
  DeviceGuard g(m->ctx->device);
  {
    // default: tensor-core kernel (tcgen05, split-precision TF32, within the 1e-4 parity budget); B2_LSTM_TC=0 → FP32 FMA kernel
    const bool want_tc = m->ctx->knobs.lstm_tc != 0;
    if (want_tc && m->tc_floats > 0)
      return b2_lstm_tc_launch(m->ctx, m->d_packed_tc, m->tc_floats, m->tc_layers, m->desc.features, x_dev, batch, timesteps, seq_stride,
                               y_dev, (cudaStream_t)stream);
  }
  LstmArgs a;
  a.x = x_dev;
  a.y = y_dev;
  a.batch = batch;
  a.seq_stride = seq_stride;
  a.T = timesteps;
  a.F = m->desc.features;
  a.u1 = 32, a.u2 = 16, a.u3 = 16, a.u4 = 32;
  a.packed = m->d_packed;
  a.packed_floats = m->packed_floats;
  const size_t smem = sizeof(float) * ((size_t)m->packed_floats + (size_t)(a.F + 96) * LS_SEQ) + 64;
  B2_REQUIRE(smem <= (size_t)m->ctx->smem_optin, "b2_lstm_forward: needs %zu bytes of shared memory", smem);
  B2_CUDA(cudaFuncSetAttribute(lstm_forward_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  long long tiles = (batch + LS_SEQ - 1) / LS_SEQ;
  long long grid = m->ctx->sm_count;
  if (grid > tiles) grid = tiles;
  lstm_forward_kernel<<<(unsigned)grid, LS_NT, smem, (cudaStream_t)stream>>>(a);
  B2_CUDA(cudaGetLastError());
  return B2_OK;
}

extern "C" int b2_lstm_forward(b2_lstm* m, const float* x_dev, int64_t batch, int timesteps, float* y_dev, void* stream) {
  B2_REQUIRE(m && x_dev && y_dev, "b2_lstm_forward: NULL argument");
  B2_REQUIRE(batch >= 0 && timesteps >= 1, "b2_lstm_forward: batch=%lld timesteps=%d", (long long)batch, timesteps);
  if (batch == 0) return B2_OK;
  return lstm_launch(m, x_dev, batch, timesteps, timesteps, y_dev, stream);
}

extern "C" int64_t b2_lstm_window_count(int64_t n_rows, int look_back, int shift) {
  if (look_back < 1 || shift < 1 || n_rows < look_back) return 0;
  return (n_rows - look_back) / shift + 1;
}

extern "C" int b2_lstm_forward_windows(b2_lstm* m, const float* rows_dev, int64_t n_rows, int look_back, int shift, float* y_dev,
                                       void* stream) {
  B2_REQUIRE(m, "b2_lstm_forward_windows: NULL model");
  B2_REQUIRE(n_rows >= 0 && look_back >= 1 && shift >= 1, "b2_lstm_forward_windows: n_rows=%lld look_back=%d shift=%d", (long long)n_rows,
             look_back, shift);
  const int64_t nw = b2_lstm_window_count(n_rows, look_back, shift);
  if (nw == 0) return B2_OK;  // fewer rows than one window (drop_remainder)
  B2_REQUIRE(rows_dev && y_dev, "b2_lstm_forward_windows: NULL device pointer");
  return lstm_launch(m, rows_dev, nw, look_back, shift, y_dev, stream);
}
