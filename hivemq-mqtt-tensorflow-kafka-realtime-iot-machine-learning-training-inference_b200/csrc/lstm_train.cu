// K4b — training of the reference LSTM stack: forward, back-propagation through time, Adam, one persistent CTA per epoch.
//
// Replaces  model.compile(metrics=['accuracy'], loss='mean_squared_error', optimizer='adam')   LSTM-TensorFlow-IO-Kafka/cardata-v1.py:178
//           model.fit(zip((dataset_x, dataset_y)).batch(batch_size).take(1000), epochs=5)        :191-194  (same in cardata-v2.py)
// for the stack  LSTM(32, relu, seq) → LSTM(16, relu) → RepeatVector(T) → LSTM(16, relu, seq) → LSTM(32, relu, seq) →
// TimeDistributed(Dense(F))  (:169-176) with tf.keras semantics: gate order i,f,c,o; sigmoid gates; relu candidate and output;
// h0 = c0 = 0; loss = mean over batch, time and features of (ŷ - y)²; Adam with ε outside the bias correction (Keras 2.2.4-tf);
// 'accuracy' = argmax over the feature axis matches (what Keras resolves for this compile call).
//
// The reference trains with batch_size 1 and look_back 1 (:160-161): a chain of 5 000 dependent steps on 18 642 parameters —
// latency-bound, like the autoencoder's fit (ae_train.cu).  So ONE CTA of 1024 threads runs every step of a launch: the
// weights live in shared memory (74.6 KB) for the whole launch, the Adam slots stay in global memory (each thread owns a fixed
// set of parameters), the per-step activations / gate gradients go to a small scratch area that stays in L2.
// Per step:  forward (time-major, layer by layer)  →  loss, ∂loss/∂ŷ  →  BPTT layer 4 … 1 (gate gradients kept per row)  →
// one pass with a thread per parameter: gradient = Σ_rows input·∂z, immediately followed by that parameter's Adam update.
// Bound: latency (one CTA, ≈ 25 block-wide barriers per step at T = 1).  Parity: tests/test_lstm_train_gpu.py against the
// float64 BPTT of oracle/cpu_oracle.py (itself checked against torch autograd).
#include "common.cuh"

namespace {

constexpr int LT_NT = 1024;
constexpr int LT_LAYERS = 4;

struct LstmTrainArgs {
  const float* x;            // sequence s, step t: x + (s * x_stride + t) * F
  const float* y;            // target   s, step t: y + (s * y_stride + t) * F
  long long n_seq, x_stride, y_stride;
  int T, F, batch;
  int in_dim[LT_LAYERS], u[LT_LAYERS];
  int offW[LT_LAYERS], offU[LT_LAYERS], offB[LT_LAYERS], offWd, offBd, P;  // float offsets inside the flat Keras weight vector
  float lr, beta1, beta2, eps;
  long long iter0;
  float* w;                  // [P] device, updated in place
  float* m;
  float* v;
  float* scratch;
  float* loss;               // [steps] or NULL
  float* acc;
};

__device__ __forceinline__ float sigmoid_acc(float x) { return 1.0f / (1.0f + __expf(-x)); }

__global__ void __launch_bounds__(LT_NT, 1) lstm_fit_kernel(const __grid_constant__ LstmTrainArgs a) {
  extern __shared__ __align__(16) float sm[];
  float* th = sm;                 // weights [P]
  float* red = sm + ((a.P + 3) & ~3);  // [64] reduction scratch
  const int tid = threadIdx.x;
  const int F = a.F, T = a.T;
  for (int i = tid; i < a.P; i += LT_NT) th[i] = a.w[i];
  __syncthreads();

  // scratch layout for up to batch * T rows (row r = t * B + b): per layer the gates / cell / output [6u], the gate gradients
  // [4u] and the gradient arriving at the layer's outputs [u]; then ŷ and ∂loss/∂ŷ [F each]; then the two recurrent carries
  const int Rmax = a.batch * T;
  float* act[LT_LAYERS];
  float* dz[LT_LAYERS];
  float* dhx[LT_LAYERS];
  float* sp = a.scratch;
  for (int l = 0; l < LT_LAYERS; ++l) {
    act[l] = sp, sp += (size_t)Rmax * 6 * a.u[l];
    dz[l] = sp, sp += (size_t)Rmax * 4 * a.u[l];
    dhx[l] = sp, sp += (size_t)Rmax * a.u[l];
  }
  float* out = sp;
  float* dout = out + (size_t)Rmax * F;
  float* dhn = dout + (size_t)Rmax * F;
  float* dcn = dhn + (size_t)a.batch * 32;

  const long long n_steps = (a.n_seq + a.batch - 1) / a.batch;
  for (long long step = 0; step < n_steps; ++step) {
    const long long s0 = step * a.batch;
    const int B = (int)min((long long)a.batch, a.n_seq - s0);
    const int R = B * T;

    // ---------------- forward ----------------
    for (int l = 0; l < LT_LAYERS; ++l) {
      const int u = a.u[l], n4 = 4 * u, in = a.in_dim[l];
      const float* W = th + a.offW[l];
      const float* U = th + a.offU[l];
      const float* bias = th + a.offB[l];
      for (int t = 0; t < T; ++t) {
        // pre-activations z[b][col] → act slots (raw)
        for (int idx = tid; idx < B * n4; idx += LT_NT) {
          const int b = idx / n4, col = idx - b * n4;
          const int r = t * B + b;
          const float* inp;
          if (l == 0) inp = a.x + ((s0 + b) * a.x_stride + t) * F;
          else if (l == 2) inp = act[1] + (size_t)((T - 1) * B + b) * 6 * a.u[1] + 5 * a.u[1];  // RepeatVector: h2 of the last step
          else inp = act[l - 1] + (size_t)r * 6 * a.u[l - 1] + 5 * a.u[l - 1];
          float z = bias[col];
          for (int k = 0; k < in; ++k) z = fmaf(inp[k], W[k * n4 + col], z);
          if (t > 0) {
            const float* hp = act[l] + (size_t)((t - 1) * B + b) * 6 * u + 5 * u;
            for (int j = 0; j < u; ++j) z = fmaf(hp[j], U[j * n4 + col], z);
          }
          act[l][(size_t)r * 6 * u + col] = z;
        }
        __syncthreads();
        // gates, cell, output
        for (int idx = tid; idx < B * u; idx += LT_NT) {
          const int b = idx / u, j = idx - b * u;
          float* ar = act[l] + (size_t)(t * B + b) * 6 * u;
          const float ig = sigmoid_acc(ar[j]), fg = sigmoid_acc(ar[u + j]), gg = fmaxf(ar[2 * u + j], 0.f), og = sigmoid_acc(ar[3 * u + j]);
          const float cp = t > 0 ? act[l][(size_t)((t - 1) * B + b) * 6 * u + 4 * u + j] : 0.f;
          const float c = fmaf(fg, cp, ig * gg);
          ar[j] = ig, ar[u + j] = fg, ar[2 * u + j] = gg, ar[3 * u + j] = og;
          ar[4 * u + j] = c;
          ar[5 * u + j] = og * fmaxf(c, 0.f);
        }
        __syncthreads();
      }
    }
    // TimeDistributed(Dense): out[r][f]; ∂loss/∂out; loss and accuracy
    {
      const float* Wd = th + a.offWd;
      const float* bd = th + a.offBd;
      const int u4 = a.u[3];
      const float scale = 2.0f / (float)(R * F);
      float se = 0.f;
      for (int idx = tid; idx < R * F; idx += LT_NT) {
        const int r = idx / F, f = idx - r * F;
        const int t = r / B, b = r - t * B;
        const float* h4 = act[3] + (size_t)r * 6 * u4 + 5 * u4;
        float o = bd[f];
        for (int j = 0; j < u4; ++j) o = fmaf(h4[j], Wd[j * F + f], o);
        const float d = o - a.y[((s0 + b) * a.y_stride + t) * F + f];
        out[idx] = o;
        dout[idx] = d * scale;
        se = fmaf(d, d, se);
      }
      // block sum of se
      for (int o = 16; o > 0; o >>= 1) se += __shfl_xor_sync(0xffffffffu, se, o);
      if ((tid & 31) == 0) red[tid >> 5] = se;
      __syncthreads();
      if (tid < 32) {
        float s = red[tid];
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        if (tid == 0 && a.loss) a.loss[step] = s / (float)(R * F);
      }
      // accuracy: argmax over features of out vs y (first maximum wins, like np.argmax)
      float hits = 0.f;
      for (int r = tid; r < R; r += LT_NT) {
        const int t = r / B, b = r - t * B;
        const float* yr = a.y + ((s0 + b) * a.y_stride + t) * F;
        int am_o = 0, am_y = 0;
        for (int f = 1; f < F; ++f) {
          if (out[r * F + f] > out[r * F + am_o]) am_o = f;
          if (yr[f] > yr[am_y]) am_y = f;
        }
        hits += am_o == am_y ? 1.f : 0.f;
      }
      __syncthreads();  // red is reused
      for (int o = 16; o > 0; o >>= 1) hits += __shfl_xor_sync(0xffffffffu, hits, o);
      if ((tid & 31) == 0) red[tid >> 5] = hits;
      __syncthreads();
      if (tid < 32) {
        float s = red[tid];
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        if (tid == 0 && a.acc) a.acc[step] = s / (float)R;
      }
      // gradient w.r.t. the top LSTM's outputs: dhx[3][r][j] = Σ_f dout[r][f] · Wd[j][f]
      for (int idx = tid; idx < R * u4; idx += LT_NT) {
        const int r = idx / u4, j = idx - r * u4;
        float g = 0.f;
        for (int f = 0; f < F; ++f) g = fmaf(dout[r * F + f], Wd[j * F + f], g);
        dhx[3][idx] = g;
      }
      // the encoder's second layer receives gradient only through its last step (filled by layer 3's backward pass)
      for (int idx = tid; idx < R * a.u[1]; idx += LT_NT) dhx[1][idx] = 0.f;
      __syncthreads();
    }

    // ---------------- back-propagation through time ----------------
    for (int l = LT_LAYERS - 1; l >= 0; --l) {
      const int u = a.u[l], n4 = 4 * u, in = a.in_dim[l];
      const float* W = th + a.offW[l];
      const float* U = th + a.offU[l];
      for (int idx = tid; idx < B * u; idx += LT_NT) dhn[idx] = 0.f, dcn[idx] = 0.f;
      __syncthreads();
      for (int t = T - 1; t >= 0; --t) {
        for (int idx = tid; idx < B * u; idx += LT_NT) {
          const int b = idx / u, j = idx - b * u;
          const int r = t * B + b;
          const float* ar = act[l] + (size_t)r * 6 * u;
          const float ig = ar[j], fg = ar[u + j], gg = ar[2 * u + j], og = ar[3 * u + j], c = ar[4 * u + j];
          const float cp = t > 0 ? act[l][(size_t)((t - 1) * B + b) * 6 * u + 4 * u + j] : 0.f;
          const float dh = dhx[l][(size_t)r * u + j] + dhn[idx];
          const float d_o = dh * fmaxf(c, 0.f);
          const float dc = (c > 0.f ? dh * og : 0.f) + dcn[idx];
          float* dr = dz[l] + (size_t)r * n4;
          dr[j] = dc * gg * ig * (1.f - ig);
          dr[u + j] = dc * cp * fg * (1.f - fg);
          dr[2 * u + j] = gg > 0.f ? dc * ig : 0.f;
          dr[3 * u + j] = d_o * og * (1.f - og);
          dcn[idx] = dc * fg;
        }
        __syncthreads();
        // recurrent carry dhn[b][j] = Σ_col dz[r][col] · U[j][col]; the column loop starts at j so that the threads of a warp
        // (consecutive j) hit distinct shared-memory banks (rows of U are 4u = a multiple of 32 words apart)
        for (int idx = tid; idx < B * u; idx += LT_NT) {
          const int b = idx / u, j = idx - b * u;
          const float* dr = dz[l] + (size_t)(t * B + b) * n4;
          float g = 0.f;
          for (int cc = 0; cc < n4; ++cc) {
            const int col = (cc + j) & (n4 - 1);
            g = fmaf(dr[col], U[j * n4 + col], g);
          }
          dhn[idx] = g;
        }
        // gradient for the layer below: din[b][k] = Σ_col dz[r][col] · W[k][col]
        if (l > 0) {
          for (int idx = tid; idx < B * in; idx += LT_NT) {
            const int b = idx / in, k = idx - b * in;
            const int r = t * B + b;
            const float* dr = dz[l] + (size_t)r * n4;
            float g = 0.f;
            for (int cc = 0; cc < n4; ++cc) {
              const int col = (cc + k) & (n4 - 1);
              g = fmaf(dr[col], W[k * n4 + col], g);
            }
            if (l == 2) {  // RepeatVector: every decoder step's input gradient lands on h2 of the LAST encoder step
              float* dst = dhx[1] + (size_t)((T - 1) * B + b) * in + k;
              *dst += g;  // this thread owns (b, k) for every t: no race
            } else {
              dhx[l - 1][(size_t)r * in + k] = g;
            }
          }
        }
        __syncthreads();
      }
    }

    // ---------------- gradients per parameter + Adam ----------------
    {
      const double tt = (double)(a.iter0 + step + 1);
      const float alpha = (float)((double)a.lr * sqrt(1.0 - pow((double)a.beta2, tt)) / (1.0 - pow((double)a.beta1, tt)));
      for (int p = tid; p < a.P; p += LT_NT) {
        float g = 0.f;
        if (p >= a.offWd) {
          const int u4 = a.u[3];
          if (p >= a.offBd) {
            const int f = p - a.offBd;
            for (int r = 0; r < R; ++r) g += dout[r * F + f];
          } else {
            const int j = (p - a.offWd) / F, f = (p - a.offWd) - j * F;
            for (int r = 0; r < R; ++r) g = fmaf(act[3][(size_t)r * 6 * u4 + 5 * u4 + j], dout[r * F + f], g);
          }
        } else {
          int l = 0;
          while (l + 1 < LT_LAYERS && p >= a.offW[l + 1]) ++l;
          const int u = a.u[l], n4 = 4 * u;
          if (p >= a.offB[l]) {
            const int col = p - a.offB[l];
            for (int r = 0; r < R; ++r) g += dz[l][(size_t)r * n4 + col];
          } else if (p >= a.offU[l]) {
            const int j = (p - a.offU[l]) / n4, col = (p - a.offU[l]) - j * n4;
            for (int r = B; r < R; ++r)  // rows of step t >= 1; the previous step's h of the same sequence is B rows earlier
              g = fmaf(act[l][(size_t)(r - B) * 6 * u + 5 * u + j], dz[l][(size_t)r * n4 + col], g);
          } else {
            const int k = (p - a.offW[l]) / n4, col = (p - a.offW[l]) - k * n4;
            for (int r = 0; r < R; ++r) {
              const int t = r / B, b = r - t * B;
              float xin;
              if (l == 0) xin = a.x[((s0 + b) * a.x_stride + t) * F + k];
              else if (l == 2) xin = act[1][(size_t)((T - 1) * B + b) * 6 * a.u[1] + 5 * a.u[1] + k];
              else xin = act[l - 1][(size_t)r * 6 * a.u[l - 1] + 5 * a.u[l - 1] + k];
              g = fmaf(xin, dz[l][(size_t)r * n4 + col], g);
            }
          }
        }
        float mm = a.m[p], vv = a.v[p];
        mm = fmaf(1.0f - a.beta1, g - mm, mm);
        vv = fmaf(1.0f - a.beta2, g * g - vv, vv);
        a.m[p] = mm;
        a.v[p] = vv;
        th[p] = th[p] - alpha * mm / (sqrtf(vv) + a.eps);
      }
      __syncthreads();
    }
  }
  for (int i = tid; i < a.P; i += LT_NT) a.w[i] = th[i];
}

int64_t scratch_floats(const b2_lstm_desc* d, int T, int batch) {
  int64_t per_row = 2 * d->features;
  for (int l = 0; l < LT_LAYERS; ++l) per_row += 11 * d->units[l];
  return per_row * (int64_t)T * batch + 64 * (int64_t)batch + 64;
}

}  // namespace

extern "C" int64_t b2_lstm_fit_scratch_bytes(const b2_lstm_desc* desc, int timesteps, int batch) {
  if (!desc || timesteps < 1 || batch < 1) return 0;
  return (int64_t)sizeof(float) * scratch_floats(desc, timesteps, batch);
}

extern "C" int b2_lstm_fit_steps(b2_ctx* ctx, const b2_lstm_desc* desc, float* w_dev, float* m_dev, float* v_dev, const float* x_dev,
                                 int64_t x_seq_stride, const float* y_dev, int64_t y_seq_stride, int64_t n_seq, int timesteps, int batch,
                                 b2_adam* adam, float* loss_dev, float* acc_dev, void* scratch_dev, int64_t scratch_bytes, void* stream) {
  B2_REQUIRE(ctx && desc && w_dev && m_dev && v_dev && x_dev && y_dev && adam && scratch_dev, "b2_lstm_fit_steps: NULL argument");
  B2_REQUIRE(n_seq >= 1 && timesteps >= 1 && batch >= 1, "b2_lstm_fit_steps: n_seq=%lld timesteps=%d batch=%d", (long long)n_seq, timesteps, batch);
  B2_REQUIRE((int64_t)timesteps * batch <= 8192, "b2_lstm_fit_steps: batch x timesteps = %lld rows per step (limit 8192)",
             (long long)timesteps * batch);
  B2_REQUIRE(x_seq_stride >= 1 && y_seq_stride >= 1, "b2_lstm_fit_steps: sequence strides must be >= 1 row");
  B2_REQUIRE(scratch_bytes >= b2_lstm_fit_scratch_bytes(desc, timesteps, batch), "b2_lstm_fit_steps: scratch of %lld bytes, %lld needed",
             (long long)scratch_bytes, (long long)b2_lstm_fit_scratch_bytes(desc, timesteps, batch));
  for (int l = 0; l < LT_LAYERS; ++l) {
    const int u = desc->units[l];
    B2_REQUIRE(u >= 8 && u <= 32 && (u & (u - 1)) == 0, "b2_lstm_fit_steps: units[%d]=%d (8, 16 or 32)", l, u);
  }
  B2_REQUIRE(desc->features >= 1 && desc->features <= 32, "b2_lstm_fit_steps: features=%d not in 1..32", desc->features);
  DeviceGuard g(ctx->device);
  LstmTrainArgs a;
  memset(&a, 0, sizeof(a));
  a.x = x_dev, a.y = y_dev;
  a.n_seq = n_seq, a.x_stride = x_seq_stride, a.y_stride = y_seq_stride;
  a.T = timesteps, a.F = desc->features, a.batch = batch;
  int p = 0, in = desc->features;
  for (int l = 0; l < LT_LAYERS; ++l) {
    a.in_dim[l] = in, a.u[l] = desc->units[l];
    const int n4 = 4 * a.u[l];
    a.offW[l] = p, p += in * n4;
    a.offU[l] = p, p += a.u[l] * n4;
    a.offB[l] = p, p += n4;
    in = a.u[l];
  }
  a.offWd = p, p += in * a.F;
  a.offBd = p, p += a.F;
  a.P = p;
  a.lr = adam->lr, a.beta1 = adam->beta1, a.beta2 = adam->beta2, a.eps = adam->eps;
  a.iter0 = adam->iter;
  a.w = w_dev, a.m = m_dev, a.v = v_dev;
  a.scratch = (float*)scratch_dev;
  a.loss = loss_dev, a.acc = acc_dev;
  const size_t smem = sizeof(float) * (((size_t)a.P + 3) & ~(size_t)3) + sizeof(float) * 64;
  B2_REQUIRE(smem <= (size_t)ctx->smem_optin, "b2_lstm_fit_steps: %d parameters do not fit in shared memory", a.P);
  B2_CUDA(cudaFuncSetAttribute(lstm_fit_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  lstm_fit_kernel<<<1, LT_NT, smem, (cudaStream_t)stream>>>(a);
  B2_CUDA(cudaGetLastError());
  adam->iter += (n_seq + batch - 1) / batch;
  return B2_OK;
}
