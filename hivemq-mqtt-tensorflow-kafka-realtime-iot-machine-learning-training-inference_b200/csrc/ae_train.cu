// K3 — fused forward / backward / Adam training of the Dense autoencoder.
//
// Replaces Model.compile(optimizer='adam', loss='mean_squared_error', metrics=['accuracy']) + Model.fit
//   (AUTOENCODER-TensorFlow-IO-Kafka/cardata-v1.py:169-171,192-194; cardata-v3.py:203-205,220-222; nb2.0:858-877)
// whose per-step op list is pinned by the committed profiler trace:
//   L = mean_i mse_i + (l1/B)·Σ|h1|;  dŷ = 2/(B·D)·(ŷ-x)⊙[ŷ>0];  dh1 += (l1/B)·sign(h1) before TanhGrad;
//   Adam: α_t = lr·√(1-β2^t)/(1-β1^t), m += (1-β1)(g-m), v += (1-β2)(g²-v), θ -= α_t·m/(√v+ε)   [ResourceApplyAdam]
//
// Two shapes of the same math:
//  * ae_fit_kernel  — the reference's batch sizes (32/100/128, ≤ 256): training is a chain of dependent steps on 835
//    parameters, i.e. latency-bound.  ONE persistent CTA runs a whole epoch: θ (padded AeW layout), all activations and
//    deltas of the batch live in shared memory, every thread owns ≤ 4 parameters with their Adam slots in registers,
//    the next batch's rows are prefetched into registers while the current step computes.  No global traffic per step
//    besides the batch rows (which stay L2-resident: 227 468 × 120 B = 27 MB).
//  * ae_grad_kernel + ae_reduce_kernel + ae_apply_kernel — large batches / data parallel: a persistent grid computes
//    per-CTA partial gradient sums (scratch[cta][P+3] in global memory), a second kernel folds the partials, the caller
//    all-reduces the P+3 floats across ranks (NCCL) and ae_apply_kernel does the Adam step.
#include <mutex>

#include "ae_math.cuh"

constexpr int FIT_NT = 256;     // threads of the fit CTA = max batch rows
constexpr int PPT = 4;          // parameters per thread (4 * 256 >= 835)

template <int D>
struct TrainLayout {
  static constexpr int PD = AeW<D>::PD;
  static constexpr int P = D * 14 + 14 + 98 + 7 + 49 + 7 + 7 * D + D;
  static constexpr int WPAD = sizeof(AeW<D>) / 4;
};

// flat Keras parameter index → offset (in floats) inside the padded AeW<D> struct
template <int D>
__host__ __device__ inline int padded_offset(int p) {
  constexpr int PD = (D + 3) & ~3;
  int off = 0;
  if (p < D * 14) return off + (p / 14) * AE_P1 + p % 14;
  p -= D * 14, off += D * AE_P1;
  if (p < 14) return off + p;
  p -= 14, off += AE_P1;
  if (p < 98) return off + (p / 7) * AE_P2 + p % 7;
  p -= 98, off += AE_H1 * AE_P2;
  if (p < 7) return off + p;
  p -= 7, off += AE_P2;
  if (p < 49) return off + (p / 7) * AE_P3 + p % 7;
  p -= 49, off += AE_H2 * AE_P3;
  if (p < 7) return off + p;
  p -= 7, off += AE_P3;
  // layer 4 is stored per output pair: (b[2q], b[2q+1], k[0][2q], k[0][2q+1], ..., k[6][2q], k[6][2q+1]), see AeW::o4
  constexpr int PAIR = 2 + 2 * AE_H3;
  if (p < 7 * D) return off + ((p % D) >> 1) * PAIR + 2 + 2 * (p / D) + ((p % D) & 1);
  p -= 7 * D;
  return off + (p >> 1) * PAIR + (p & 1);
}

// gradient of one parameter = Σ_b A[b*sa + ia] * B[b*sb + ib]   (A == nullptr → bias: Σ_b B[...])
struct ParamSrc {
  int a_off, a_stride, b_off, b_stride;  // offsets in floats into the activation arena; a_off < 0 → bias
};

template <int D>
struct Arena {  // per-batch-row activations and deltas (floats), row-major per array
  static constexpr int X = 0;                // [B][D]   normalised inputs
  static constexpr int H1 = X + FIT_NT * D;  // [B][14]
  static constexpr int H2 = H1 + FIT_NT * 14;
  static constexpr int H3 = H2 + FIT_NT * 7;
  static constexpr int Z1 = H3 + FIT_NT * 7;  // dL/dz1 [B][14]
  static constexpr int Z2 = Z1 + FIT_NT * 14;
  static constexpr int Z3 = Z2 + FIT_NT * 7;
  static constexpr int Z4 = Z3 + FIT_NT * 7;  // [B][D]
  static constexpr int END = Z4 + FIT_NT * D;
};

template <int D>
__device__ inline ParamSrc param_src(int p) {
  using A = Arena<D>;
  ParamSrc s;
  if (p < D * 14) return s = {A::X + p / 14, D, A::Z1 + p % 14, 14}, s;
  p -= D * 14;
  if (p < 14) return s = {-1, 0, A::Z1 + p, 14}, s;
  p -= 14;
  if (p < 98) return s = {A::H1 + p / 7, 14, A::Z2 + p % 7, 7}, s;
  p -= 98;
  if (p < 7) return s = {-1, 0, A::Z2 + p, 7}, s;
  p -= 7;
  if (p < 49) return s = {A::H2 + p / 7, 7, A::Z3 + p % 7, 7}, s;
  p -= 49;
  if (p < 7) return s = {-1, 0, A::Z3 + p, 7}, s;
  p -= 7;
  if (p < 7 * D) return s = {A::H3 + p / D, 7, A::Z4 + p % D, D}, s;
  p -= 7 * D;
  return s = {-1, 0, A::Z4 + p, D}, s;
}

// Forward + backward of ONE row (thread-private), weights read from the padded AeW copy in shared memory.
// c_mse = 2/(B*D) (or 2/D for un-normalised sums), c_l1 = l1/B (or l1).  Writes h1,h2,h3 and the four deltas into
// the arena; returns the row's squared-error sum, Σ|h1| and argmax agreement.
template <int D>
__device__ __forceinline__ void row_fwd_bwd(const AeW<D>& w, float* __restrict__ arena, int t, float c_mse, float c_l1, float& se_out,
                                            float& l1_out, int& correct_out) {
  using A = Arena<D>;
  constexpr int PD = AeW<D>::PD;
  const float* x = arena + A::X + t * D;
  float h1[AE_H1], h2[AE_H2], h3[AE_H3];
#pragma unroll
  for (int o = 0; o < AE_H1; ++o) h1[o] = w.b1[o];
#pragma unroll 2
  for (int i = 0; i < D; ++i) {
    const float xi = x[i];
#pragma unroll
    for (int o = 0; o < AE_H1; ++o) h1[o] = fmaf(xi, w.k1[i][o], h1[o]);
  }
  float l1s = 0.f;
#pragma unroll
  for (int o = 0; o < AE_H1; ++o) {
    h1[o] = tanh1(h1[o]);
    l1s += fabsf(h1[o]);
    arena[A::H1 + t * 14 + o] = h1[o];
  }
#pragma unroll
  for (int o = 0; o < AE_H2; ++o) h2[o] = w.b2[o];
#pragma unroll
  for (int i = 0; i < AE_H1; ++i)
#pragma unroll
    for (int o = 0; o < AE_H2; ++o) h2[o] = fmaf(h1[i], w.k2[i][o], h2[o]);
#pragma unroll
  for (int o = 0; o < AE_H2; ++o) {
    h2[o] = fmaxf(h2[o], 0.f);
    arena[A::H2 + t * 7 + o] = h2[o];
  }
#pragma unroll
  for (int o = 0; o < AE_H3; ++o) h3[o] = w.b3[o];
#pragma unroll
  for (int i = 0; i < AE_H2; ++i)
#pragma unroll
    for (int o = 0; o < AE_H3; ++o) h3[o] = fmaf(h2[i], w.k3[i][o], h3[o]);
#pragma unroll
  for (int o = 0; o < AE_H3; ++o) {
    h3[o] = tanh1(h3[o]);
    arena[A::H3 + t * 7 + o] = h3[o];
  }
  // output layer, loss, dL/dz4 and dL/dh3 in one sweep over the outputs
  float dh3[AE_H3];
#pragma unroll
  for (int k = 0; k < AE_H3; ++k) dh3[k] = 0.f;
  float se = 0.f, ymax = -1.f, xmax = 0.f;
  int yarg = 0, xarg = 0;
#pragma unroll 2
  for (int j = 0; j < D; ++j) {
    float y = w.b4(j);
#pragma unroll
    for (int k = 0; k < AE_H3; ++k) y = fmaf(h3[k], w.k4(k, j), y);
    y = fmaxf(y, 0.f);
    const float xj = x[j];
    const float d = y - xj;
    se = fmaf(d, d, se);
    if (y > ymax) ymax = y, yarg = j;  // relu output ≥ 0 > -1: the first maximum wins, like ArgMax
    if (j == 0 || xj > xmax) xmax = xj, xarg = j;
    const float dz = y > 0.f ? c_mse * d : 0.f;
    arena[A::Z4 + t * D + j] = dz;
#pragma unroll
    for (int k = 0; k < AE_H3; ++k) dh3[k] = fmaf(dz, w.k4(k, j), dh3[k]);
  }
  (void)PD;
  float dz3[AE_H3], dz2[AE_H2];
#pragma unroll
  for (int k = 0; k < AE_H3; ++k) {
    dz3[k] = dh3[k] * fmaf(-h3[k], h3[k], 1.0f);
    arena[A::Z3 + t * 7 + k] = dz3[k];
  }
#pragma unroll
  for (int i = 0; i < AE_H2; ++i) {
    float acc = 0.f;
#pragma unroll
    for (int k = 0; k < AE_H3; ++k) acc = fmaf(dz3[k], w.k3[i][k], acc);
    dz2[i] = h2[i] > 0.f ? acc : 0.f;
    arena[A::Z2 + t * 7 + i] = dz2[i];
  }
#pragma unroll
  for (int i = 0; i < AE_H1; ++i) {
    float acc = 0.f;
#pragma unroll
    for (int k = 0; k < AE_H2; ++k) acc = fmaf(dz2[k], w.k2[i][k], acc);
    const float sg = h1[i] > 0.f ? 1.f : (h1[i] < 0.f ? -1.f : 0.f);  // Abs_grad/Sign
    acc = fmaf(c_l1, sg, acc);                                        // AddN before TanhGrad
    arena[A::Z1 + t * 14 + i] = acc * fmaf(-h1[i], h1[i], 1.0f);
  }
  se_out = se;
  l1_out = l1s;
  correct_out = yarg == xarg;
}

__device__ __forceinline__ float block_sum(float v, float* red, int tid, int nt) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  if ((tid & 31) == 0) red[tid >> 5] = v;
  __syncthreads();
  float s = 0.f;
  for (int i = 0; i < nt / 32; ++i) s += red[i];
  __syncthreads();
  return s;
}

template <int D>
struct FitArgs {
  const float* x;
  const int32_t* order;
  long long rows_total;
  int batch;
  NormP<D> nm;
  int has_norm;
  float l1;
  float lr, beta1, beta2, eps;
  long long iter0;
  float* w;  // device, Keras flat order (in/out)
  float* m;
  float* v;
  float* loss;
  float* acc;
};

template <int D>
__device__ __forceinline__ void load_row(const FitArgs<D>& a, long long r, float (&buf)[D]) {
  const float2* src = reinterpret_cast<const float2*>(a.x + r * D);
#pragma unroll
  for (int i = 0; i < D / 2; ++i) {
    const float2 v = __ldg(src + i);
    buf[2 * i] = v.x;
    buf[2 * i + 1] = v.y;
  }
}

template <int D>
__global__ void __launch_bounds__(FIT_NT, 1) ae_fit_kernel(const __grid_constant__ FitArgs<D> a) {
  using L = TrainLayout<D>;
  using A = Arena<D>;
  extern __shared__ __align__(16) float fsm[];
  AeW<D>& w = *reinterpret_cast<AeW<D>*>(fsm);
  float* arena = fsm + L::WPAD;
  float* red = arena + A::END;
  const int tid = threadIdx.x;

  // parameter ownership: thread tid owns flat parameters tid, tid+256, ... with their Adam slots in registers
  float th[PPT], mm[PPT], vv[PPT];
  int poff[PPT];
  ParamSrc src[PPT];
  for (int i = tid; i < L::WPAD; i += FIT_NT) fsm[i] = 0.f;
  __syncthreads();
#pragma unroll
  for (int q = 0; q < PPT; ++q) {
    const int p = tid + q * FIT_NT;
    if (p < L::P) {
      th[q] = a.w[p];
      mm[q] = a.m[p];
      vv[q] = a.v[p];
      poff[q] = padded_offset<D>(p);
      src[q] = param_src<D>(p);
      fsm[poff[q]] = th[q];
    } else {
      th[q] = mm[q] = vv[q] = 0.f;
      poff[q] = -1;
      src[q] = {-1, 0, 0, 0};
    }
  }
  __syncthreads();

  const int B = a.batch;
  const long long steps = (a.rows_total + B - 1) / B;
  double b1t = pow((double)a.beta1, (double)a.iter0), b2t = pow((double)a.beta2, (double)a.iter0);
  float nxt[D];
  auto row_of = [&](long long s, int t) -> long long {
    const long long k = s * B + t;
    return a.order ? (long long)a.order[k] : k;
  };
  {
    const long long rows0 = min((long long)B, a.rows_total);
    if (tid < rows0) load_row<D>(a, row_of(0, tid), nxt);
  }
  for (long long s = 0; s < steps; ++s) {
    const int rows = (int)min((long long)B, a.rows_total - s * B);
    // stage this step's rows (normalised) and prefetch the next step's
    if (tid < rows) {
      float* xr = arena + A::X + tid * D;
#pragma unroll
      for (int i = 0; i < D; i += 2) {
        float2 v = make_float2(nxt[i], nxt[i + 1]);
        if (a.has_norm) {
          v.x = norm_feature(v.x, a.nm.kind[i], a.nm.lo[i], a.nm.range[i]);
          v.y = norm_feature(v.y, a.nm.kind[i + 1], a.nm.lo[i + 1], a.nm.range[i + 1]);
        }
        xr[i] = v.x;
        xr[i + 1] = v.y;
      }
    }
    if (s + 1 < steps) {
      const int nrows = (int)min((long long)B, a.rows_total - (s + 1) * B);
      if (tid < nrows) load_row<D>(a, row_of(s + 1, tid), nxt);
    }
    float se = 0.f, l1s = 0.f;
    int correct = 0;
    if (tid < rows) row_fwd_bwd<D>(w, arena, tid, 2.0f / ((float)rows * (float)D), a.l1 / (float)rows, se, l1s, correct);
    __syncthreads();
    // loss / accuracy of this step (Keras: mean over the batch of per-sample mse + activity penalty / batch)
    const float sse = block_sum(se, red, tid, FIT_NT);
    const float sl1 = block_sum(l1s, red, tid, FIT_NT);
    const float scor = block_sum((float)correct, red, tid, FIT_NT);
    if (tid == 0) {
      if (a.loss) a.loss[s] = sse / ((float)rows * (float)D) + a.l1 * sl1 / (float)rows;
      if (a.acc) a.acc[s] = scor / (float)rows;
    }
    // weight gradients + Adam, one thread per parameter
    b1t *= (double)a.beta1;
    b2t *= (double)a.beta2;
    const float alpha = (float)((double)a.lr * sqrt(1.0 - b2t) / (1.0 - b1t));
    float g[PPT] = {0.f, 0.f, 0.f, 0.f};
    for (int b = 0; b < rows; ++b) {
#pragma unroll
      for (int q = 0; q < PPT; ++q) {
        const float bv = arena[src[q].b_off + b * src[q].b_stride];
        const float av = src[q].a_off >= 0 ? arena[src[q].a_off + b * src[q].a_stride] : 1.0f;
        g[q] = fmaf(av, bv, g[q]);
      }
    }
#pragma unroll
    for (int q = 0; q < PPT; ++q) {
      if (poff[q] >= 0) {
        mm[q] = fmaf(1.0f - a.beta1, g[q] - mm[q], mm[q]);
        vv[q] = fmaf(1.0f - a.beta2, g[q] * g[q] - vv[q], vv[q]);
        th[q] = th[q] - alpha * mm[q] / (sqrtf(vv[q]) + a.eps);
      }
    }
    __syncthreads();  // everyone is done reading the old weights / arena
#pragma unroll
    for (int q = 0; q < PPT; ++q)
      if (poff[q] >= 0) fsm[poff[q]] = th[q];
    __syncthreads();
  }
#pragma unroll
  for (int q = 0; q < PPT; ++q) {
    const int p = tid + q * FIT_NT;
    if (p < L::P) {
      a.w[p] = th[q];
      a.m[p] = mm[q];
      a.v[p] = vv[q];
    }
  }
}

// ---- large batch / data parallel -----------------------------------------------------------------------------------
// One data-parallel step as ONE kernel (b2_ae_step_fused).  The CTA that finishes its tiles last folds the per-CTA partial sums
// (fixed order), then — with more than one GPU — stores its 840-float vector straight into every peer's exchange buffer over
// NVLink (plain peer stores into symmetric memory: the cudaMalloc'ed buffer of each rank mapped into every process), raises
// a flag there, waits for the peers' flags in its own buffer, adds the `world` vectors in rank order (so every GPU forms the
// identical sum) and applies the Adam update.  No NCCL call, no extra launch, no host round trip per step.
constexpr int EXCH_W = 840;              // floats per exchange slot (>= P + 3 for D = 30)
constexpr int EXCH_MAX_WORLD = 8;
constexpr int EXCH_FLAGS_OFF = 2 * EXCH_MAX_WORLD * EXCH_W;   // uint32 flags[world] behind the two parities of slots
constexpr int EXCH_FLOATS = EXCH_FLAGS_OFF + 32;

struct StepTail {
  int enabled;
  int* ticket;            // device counter, zero before the first launch; the last CTA resets it
  float* w;               // same buffer as GradArgs::w, writable
  float* m;
  float* v;
  float lr, beta1, beta2, eps;
  long long iter;         // value before this step
  long long total_rows;   // rows of the GLOBAL batch (all ranks)
  float* stats;           // [2] mean loss, accuracy of the global batch, or NULL
  int world, rank;
  unsigned seq;           // exchange sequence number of this step (1, 2, 3, … on every rank alike)
  float* peer[EXCH_MAX_WORLD];  // exchange buffers of all ranks as mapped here; peer[rank] is this GPU's own
};

__device__ __forceinline__ void st_release_sys_u32(unsigned* p, unsigned v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned ld_acquire_sys_u32(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float ld_relaxed_sys_f32(const float* p) {
  float v;
  asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
  return v;
}

// executed by the last CTA of the launch; gl = at least EXCH_W floats of shared memory
__device__ void step_tail(const StepTail& t, const float* scratch, int n_parts, int P, float* gl, int tid) {
  const int W = P + 3;
  __threadfence();
  for (int j = tid; j < W; j += FIT_NT) {
    float s = 0.f;
    for (int c = 0; c < n_parts; ++c) s += __ldcg(scratch + (size_t)c * W + j);
    gl[j] = s;
  }
  __syncthreads();
  bool ok = true;
  if (t.world > 1) {
    const int par = (int)(t.seq & 1u);
    for (int p = 0; p < t.world; ++p) {
      float* dst = t.peer[p] + (size_t)(par * t.world + t.rank) * EXCH_W;
      for (int j = tid; j < W; j += FIT_NT) dst[j] = gl[j];
    }
    __threadfence_system();
    __syncthreads();
    if (tid < t.world) {
      st_release_sys_u32(reinterpret_cast<unsigned*>(t.peer[tid] + EXCH_FLAGS_OFF) + t.rank, t.seq);
      const unsigned* mine = reinterpret_cast<const unsigned*>(t.peer[t.rank] + EXCH_FLAGS_OFF) + tid;
      const long long t0 = clock64();
      while ((int)(ld_acquire_sys_u32(mine) - t.seq) < 0) {
        if (clock64() - t0 > 8000000000ll) {  // a peer never showed up (~4 s): give up instead of hanging the GPU
          ok = false;
          break;
        }
      }
    }
    ok = __syncthreads_and(ok);
    const float* my = t.peer[t.rank] + (size_t)(par * t.world) * EXCH_W;
    for (int j = tid; j < W; j += FIT_NT) {
      float s = 0.f;
      for (int q = 0; q < t.world; ++q) s += ld_relaxed_sys_f32(my + (size_t)q * EXCH_W + j);  // rank order: same sum everywhere
      gl[j] = s;
    }
    __syncthreads();
  }
  const float rows = (float)t.total_rows;
  if (tid == 0 && t.stats) {
    t.stats[0] = ok ? gl[P] / rows : __int_as_float(0x7FC00000);
    t.stats[1] = ok ? gl[P + 1] / rows : __int_as_float(0x7FC00000);
  }
  if (ok) {
    const double tt = (double)(t.iter + 1);
    const float alpha = (float)((double)t.lr * sqrt(1.0 - pow((double)t.beta2, tt)) / (1.0 - pow((double)t.beta1, tt)));
    for (int p = tid; p < P; p += FIT_NT) {
      const float g = gl[p] / rows;
      float m = t.m[p], v = t.v[p];
      m = fmaf(1.0f - t.beta1, g - m, m);
      v = fmaf(1.0f - t.beta2, g * g - v, v);
      t.m[p] = m;
      t.v[p] = v;
      t.w[p] = t.w[p] - alpha * m / (sqrtf(v) + t.eps);
    }
  }
  if (tid == 0) *t.ticket = 0;
}

template <int D>
struct GradArgs {
  const float* x;
  long long n;
  NormP<D> nm;
  int has_norm;
  float l1;
  const float* w;   // device, Keras flat order
  float* scratch;   // [gridDim.x][P+3] partial sums
  StepTail tail;    // fused step (b2_ae_step_fused): fold + cross-GPU exchange + Adam in the CTA that finishes last
};

template <int D>
__global__ void __launch_bounds__(FIT_NT, 1) ae_grad_kernel(const __grid_constant__ GradArgs<D> a) {
  using L = TrainLayout<D>;
  using A = Arena<D>;
  extern __shared__ __align__(16) float fsm[];
  AeW<D>& w = *reinterpret_cast<AeW<D>*>(fsm);
  float* arena = fsm + L::WPAD;
  float* red = arena + A::END;
  const int tid = threadIdx.x;
  for (int i = tid; i < L::WPAD; i += FIT_NT) fsm[i] = 0.f;
  __syncthreads();
  ParamSrc src[PPT];
  bool own[PPT];
#pragma unroll
  for (int q = 0; q < PPT; ++q) {
    const int p = tid + q * FIT_NT;
    own[q] = p < L::P;
    if (own[q]) {
      fsm[padded_offset<D>(p)] = a.w[p];
      src[q] = param_src<D>(p);
    } else {
      src[q] = {-1, 0, 0, 0};
    }
  }
  __syncthreads();
  float g[PPT] = {0.f, 0.f, 0.f, 0.f};
  float se_t = 0.f, l1_t = 0.f, cor_t = 0.f;
  const long long n_tiles = (a.n + FIT_NT - 1) / FIT_NT;
  for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const long long r0 = tile * FIT_NT;
    const int rows = (int)min((long long)FIT_NT, a.n - r0);
    if (tid < rows) {
      const float2* srcx = reinterpret_cast<const float2*>(a.x + (r0 + tid) * D);
      float* xr = arena + A::X + tid * D;
#pragma unroll
      for (int i = 0; i < D; i += 2) {
        float2 v = __ldg(srcx + i / 2);
        if (a.has_norm) {
          v.x = norm_feature(v.x, a.nm.kind[i], a.nm.lo[i], a.nm.range[i]);
          v.y = norm_feature(v.y, a.nm.kind[i + 1], a.nm.lo[i + 1], a.nm.range[i + 1]);
        }
        xr[i] = v.x;
        xr[i + 1] = v.y;
      }
      float se, l1s;
      int c;
      row_fwd_bwd<D>(w, arena, tid, 2.0f / (float)D, a.l1, se, l1s, c);  // un-normalised per-row gradients
      se_t += se;
      l1_t += l1s;
      cor_t += (float)c;
    }
    __syncthreads();
    for (int b = 0; b < rows; ++b) {
#pragma unroll
      for (int q = 0; q < PPT; ++q) {
        const float bv = arena[src[q].b_off + b * src[q].b_stride];
        const float av = src[q].a_off >= 0 ? arena[src[q].a_off + b * src[q].a_stride] : 1.0f;
        g[q] = fmaf(av, bv, g[q]);
      }
    }
    __syncthreads();
  }
  float* out = a.scratch + (size_t)blockIdx.x * (L::P + 3);
#pragma unroll
  for (int q = 0; q < PPT; ++q)
    if (own[q]) out[tid + q * FIT_NT] = g[q];
  const float sse = block_sum(se_t, red, tid, FIT_NT);
  const float sl1 = block_sum(l1_t, red, tid, FIT_NT);
  const float scor = block_sum(cor_t, red, tid, FIT_NT);
  if (tid == 0) {
    out[L::P] = sse / (float)D + a.l1 * sl1;  // Σ_rows (mse_row + l1·Σ|h1_row|)
    out[L::P + 1] = scor;
    long long mine = 0;
    for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) mine += min((long long)FIT_NT, a.n - tile * FIT_NT);
    out[L::P + 2] = (float)mine;
  }
  if (a.tail.enabled) {  // fused step: whoever finishes last does the fold, the exchange and the update
    __shared__ int s_last;
    __threadfence();
    __syncthreads();
    if (tid == 0) s_last = atomicAdd(a.tail.ticket, 1) == (int)gridDim.x - 1;
    __syncthreads();
    if (s_last) step_tail(a.tail, a.scratch, (int)gridDim.x, L::P, arena, tid);
  }
}

// fold the per-CTA partials: out[j] = Σ_c scratch[c][j]  (fixed order → deterministic)
__global__ void ae_reduce_kernel(const float* __restrict__ scratch, int n_parts, int width, float* __restrict__ out) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= width) return;
  float s = 0.f;
  for (int c = 0; c < n_parts; ++c) s += scratch[(size_t)c * width + j];
  out[j] = s;
}

struct ApplyArgs {
  const float* grads;  // [P+3]: Σ gradients, Σ loss, Σ correct, Σ rows (already all-reduced by the caller)
  int P;
  float lr, beta1, beta2, eps;
  long long iter;      // value BEFORE this step
  float* w;
  float* m;
  float* v;
  float* loss_out;     // [2]: mean loss, accuracy (may be NULL)
};

__global__ void ae_apply_kernel(const ApplyArgs a) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  const float rows = a.grads[a.P + 2];
  if (p == 0 && a.loss_out) {
    a.loss_out[0] = a.grads[a.P] / rows;
    a.loss_out[1] = a.grads[a.P + 1] / rows;
  }
  if (p >= a.P) return;
  const double t = (double)(a.iter + 1);
  const float alpha = (float)((double)a.lr * sqrt(1.0 - pow((double)a.beta2, t)) / (1.0 - pow((double)a.beta1, t)));
  const float g = a.grads[p] / rows;
  float m = a.m[p], v = a.v[p];
  m = fmaf(1.0f - a.beta1, g - m, m);
  v = fmaf(1.0f - a.beta2, g * g - v, v);
  a.m[p] = m;
  a.v[p] = v;
  a.w[p] = a.w[p] - alpha * m / (sqrtf(v) + a.eps);
}

// ---- host side -------------------------------------------------------------------------------------------------------
template <int D>
static void fill_norm(const b2_norm* norm, NormP<D>& nm, int& has_norm) {
  has_norm = norm_fill<D>(norm, nm);
}

template <int D>
static size_t train_smem() {
  return sizeof(float) * (TrainLayout<D>::WPAD + Arena<D>::END + 64);
}

template <int D>
static int fit_spec(b2_ae* ae, const float* x, const int32_t* order, int64_t rows_total, int batch, const b2_norm* norm, b2_adam* adam,
                    float* loss, float* acc, cudaStream_t stream) {
  FitArgs<D> a;
  a.x = x;
  a.order = order;
  a.rows_total = rows_total;
  a.batch = batch;
  fill_norm<D>(norm, a.nm, a.has_norm);
  a.l1 = ae->desc.l1_activity;
  a.lr = adam->lr, a.beta1 = adam->beta1, a.beta2 = adam->beta2, a.eps = adam->eps;
  a.iter0 = adam->iter;
  a.w = ae->d_w, a.m = ae->d_m, a.v = ae->d_v;
  a.loss = loss, a.acc = acc;
  const size_t smem = train_smem<D>();
  B2_CUDA(cudaFuncSetAttribute(ae_fit_kernel<D>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  ae_fit_kernel<D><<<1, FIT_NT, smem, stream>>>(a);
  B2_CUDA(cudaGetLastError());
  adam->iter += (rows_total + batch - 1) / batch;
  ae->host_stale = true;
  ae->tc_valid = false;
  return B2_OK;
}

extern "C" int b2_ae_fit_steps(b2_ae* ae, const float* x_dev, const int32_t* order_dev, int64_t rows_total, int batch,
                               const b2_norm* norm, b2_adam* adam, float* loss_dev, float* acc_dev, void* stream) {
  B2_NORM_F32_ONLY(norm, "b2_ae_fit_steps");
  B2_REQUIRE(ae && x_dev && adam, "b2_ae_fit_steps: NULL argument");
  B2_REQUIRE(rows_total >= 1, "b2_ae_fit_steps: rows_total=%lld < 1", (long long)rows_total);
  B2_REQUIRE(batch >= 1 && batch <= FIT_NT, "b2_ae_fit_steps: batch=%d not in 1..%d (use b2_ae_grad/b2_ae_apply for larger batches)", batch,
             FIT_NT);
  B2_REQUIRE(((uintptr_t)x_dev & 7) == 0, "b2_ae_fit_steps: x_dev must be 8-byte aligned");
  B2_REQUIRE(!norm || norm->n == 0 || norm->n == ae->desc.widths[0], "b2_ae_fit_steps: norm table size mismatch");
  if (ae->spec_d == 0) {
    b2_set_error("b2_ae_fit_steps: training is implemented for the D-14-7-7-D tanh/relu/tanh/relu autoencoder with D in {18, 30}");
    return B2_EUNSUPPORTED;
  }
  DeviceGuard g(ae->ctx->device);
  int rc = b2_ae_ensure_adam(ae);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  if (ae->spec_d == 18) return fit_spec<18>(ae, x_dev, order_dev, rows_total, batch, norm, adam, loss_dev, acc_dev, st);
  return fit_spec<30>(ae, x_dev, order_dev, rows_total, batch, norm, adam, loss_dev, acc_dev, st);
}

static int grad_grid(const b2_ae* ae, int64_t n) {
  long long tiles = (n + FIT_NT - 1) / FIT_NT;
  long long g = ae->ctx->sm_count;
  if (g > tiles) g = tiles;
  return (int)(g < 1 ? 1 : g);
}

extern "C" int64_t b2_ae_grad_scratch_bytes(const b2_ae* ae, int64_t n) {
  if (!ae || n < 0) return 0;
  return (int64_t)sizeof(float) * (int64_t)grad_grid(ae, n) * (ae->n_params + 3);
}

template <int D>
static int grad_spec(b2_ae* ae, const float* x, int64_t n, const b2_norm* norm, float* grads, float* scratch, cudaStream_t stream) {
  GradArgs<D> a;
  a.x = x;
  a.n = n;
  fill_norm<D>(norm, a.nm, a.has_norm);
  a.l1 = ae->desc.l1_activity;
  a.w = ae->d_w;
  a.scratch = scratch;
  memset(&a.tail, 0, sizeof(a.tail));
  const size_t smem = train_smem<D>();
  const int grid = grad_grid(ae, n);
  B2_CUDA(cudaFuncSetAttribute(ae_grad_kernel<D>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  ae_grad_kernel<D><<<grid, FIT_NT, smem, stream>>>(a);
  B2_CUDA(cudaGetLastError());
  const int width = ae->n_params + 3;
  ae_reduce_kernel<<<(width + 127) / 128, 128, 0, stream>>>(scratch, grid, width, grads);
  B2_CUDA(cudaGetLastError());
  return B2_OK;
}

extern "C" int b2_ae_grad(b2_ae* ae, const float* x_dev, int64_t n, const b2_norm* norm, float* grads_dev, void* scratch_dev,
                          int64_t scratch_bytes, void* stream) {
  B2_NORM_F32_ONLY(norm, "b2_ae_grad");
  B2_REQUIRE(ae && x_dev && grads_dev && scratch_dev, "b2_ae_grad: NULL argument");
  B2_REQUIRE(n >= 1, "b2_ae_grad: n=%lld < 1", (long long)n);
  B2_REQUIRE(n <= (1ll << 24), "b2_ae_grad: n=%lld rows in one step (the row count travels with the gradient as a float: exact to 2^24)", (long long)n);
  B2_REQUIRE(((uintptr_t)x_dev & 7) == 0, "b2_ae_grad: x_dev must be 8-byte aligned");
  B2_REQUIRE(scratch_bytes >= b2_ae_grad_scratch_bytes(ae, n), "b2_ae_grad: scratch of %lld bytes is smaller than the %lld needed",
             (long long)scratch_bytes, (long long)b2_ae_grad_scratch_bytes(ae, n));
  B2_REQUIRE(!norm || norm->n == 0 || norm->n == ae->desc.widths[0], "b2_ae_grad: norm table size mismatch");
  if (ae->spec_d == 0) {
    b2_set_error("b2_ae_grad: training is implemented for the D-14-7-7-D autoencoder with D in {18, 30}");
    return B2_EUNSUPPORTED;
  }
  DeviceGuard g(ae->ctx->device);
  cudaStream_t st = (cudaStream_t)stream;
  if (ae->spec_d == 18) return grad_spec<18>(ae, x_dev, n, norm, grads_dev, (float*)scratch_dev, st);
  return grad_spec<30>(ae, x_dev, n, norm, grads_dev, (float*)scratch_dev, st);
}

extern "C" int b2_ae_apply(b2_ae* ae, const float* grads_dev, b2_adam* adam, float* loss_out_dev, void* stream) {
  B2_REQUIRE(ae && grads_dev && adam, "b2_ae_apply: NULL argument");
  DeviceGuard g(ae->ctx->device);
  int rc = b2_ae_ensure_adam(ae);
  if (rc) return rc;
  ApplyArgs a;
  a.grads = grads_dev;
  a.P = ae->n_params;
  a.lr = adam->lr, a.beta1 = adam->beta1, a.beta2 = adam->beta2, a.eps = adam->eps;
  a.iter = adam->iter;
  a.w = ae->d_w, a.m = ae->d_m, a.v = ae->d_v;
  a.loss_out = loss_out_dev;
  ae_apply_kernel<<<(a.P + 127) / 128, 128, 0, (cudaStream_t)stream>>>(a);
  B2_CUDA(cudaGetLastError());
  adam->iter += 1;
  ae->host_stale = true;
  ae->tc_valid = false;
  return B2_OK;
}

// ---- one data-parallel step in one launch ---------------------------------------------------------------------------------
template <int D>
static int step_fused_spec(b2_ae* ae, const float* x, int64_t n, const b2_norm* norm, b2_adam* adam, float* scratch, int64_t total_rows,
                           const uint64_t* peers, int world, int rank, unsigned seq, float* stats, cudaStream_t stream) {
  GradArgs<D> a;
  a.x = x;
  a.n = n;
  fill_norm<D>(norm, a.nm, a.has_norm);
  a.l1 = ae->desc.l1_activity;
  a.w = ae->d_w;
  a.scratch = scratch;
  const int grid = grad_grid(ae, n);
  memset(&a.tail, 0, sizeof(a.tail));
  a.tail.enabled = 1;
  a.tail.ticket = reinterpret_cast<int*>(scratch + (size_t)grid * (ae->n_params + 3));
  a.tail.w = ae->d_w, a.tail.m = ae->d_m, a.tail.v = ae->d_v;
  a.tail.lr = adam->lr, a.tail.beta1 = adam->beta1, a.tail.beta2 = adam->beta2, a.tail.eps = adam->eps;
  a.tail.iter = adam->iter;
  a.tail.total_rows = total_rows;
  a.tail.stats = stats;
  a.tail.world = world, a.tail.rank = rank, a.tail.seq = seq;
  for (int p = 0; p < world && peers; ++p) a.tail.peer[p] = reinterpret_cast<float*>((uintptr_t)peers[p]);
  const size_t smem = train_smem<D>();
  B2_CUDA(cudaFuncSetAttribute(ae_grad_kernel<D>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  ae_grad_kernel<D><<<grid, FIT_NT, smem, stream>>>(a);
  B2_CUDA(cudaGetLastError());
  adam->iter += 1;
  ae->host_stale = true;
  ae->tc_valid = false;
  return B2_OK;
}

extern "C" int64_t b2_ae_step_scratch_bytes(const b2_ae* ae, int64_t n) {
  if (!ae || n < 0) return 0;
  return b2_ae_grad_scratch_bytes(ae, n) + 64;  // + the ticket counter
}
extern "C" int64_t b2_ae_exchange_bytes(void) { return (int64_t)sizeof(float) * EXCH_FLOATS; }

extern "C" int b2_ae_step_fused(b2_ae* ae, const float* x_dev, int64_t n, const b2_norm* norm, b2_adam* adam, void* scratch_dev,
                                int64_t scratch_bytes, int64_t total_rows, const uint64_t* peer_bufs, int world, int rank, uint32_t seq,
                                float* stats_dev, void* stream) {
  B2_NORM_F32_ONLY(norm, "b2_ae_step_fused");
  B2_REQUIRE(ae && x_dev && adam && scratch_dev, "b2_ae_step_fused: NULL argument");
  B2_REQUIRE(n >= 1 && total_rows >= n, "b2_ae_step_fused: n=%lld total_rows=%lld", (long long)n, (long long)total_rows);
  B2_REQUIRE(((uintptr_t)x_dev & 7) == 0, "b2_ae_step_fused: x_dev must be 8-byte aligned");
  B2_REQUIRE(scratch_bytes >= b2_ae_step_scratch_bytes(ae, n), "b2_ae_step_fused: scratch of %lld bytes is smaller than the %lld needed",
             (long long)scratch_bytes, (long long)b2_ae_step_scratch_bytes(ae, n));
  B2_REQUIRE(world >= 1 && world <= EXCH_MAX_WORLD && rank >= 0 && rank < world, "b2_ae_step_fused: world=%d rank=%d", world, rank);
  B2_REQUIRE(world == 1 || (peer_bufs != nullptr && seq != 0), "b2_ae_step_fused: %d ranks need the exchange buffers and a sequence number >= 1", world);
  B2_REQUIRE(!norm || norm->n == 0 || norm->n == ae->desc.widths[0], "b2_ae_step_fused: norm table size mismatch");
  if (ae->spec_d == 0) {
    b2_set_error("b2_ae_step_fused: training is implemented for the D-14-7-7-D autoencoder with D in {18, 30}");
    return B2_EUNSUPPORTED;
  }
  DeviceGuard g(ae->ctx->device);
  int rc = b2_ae_ensure_adam(ae);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  if (ae->spec_d == 18)
    return step_fused_spec<18>(ae, x_dev, n, norm, adam, (float*)scratch_dev, total_rows, peer_bufs, world, rank, seq, stats_dev, st);
  return step_fused_spec<30>(ae, x_dev, n, norm, adam, (float*)scratch_dev, total_rows, peer_bufs, world, rank, seq, stats_dev, st);
}

// ---- gradient sums without any allocation ------------------------------------------------------------------------------
// Some hosts of this library are not allowed to allocate device memory while they run: a consumer loop that has handed
// its whole memory budget to a pool at start-up, a real-time thread that must not take the driver's allocator lock, a
// caller sitting inside someone else's memory manager whose cudaMalloc is simply off limits.  b2_ae_grad() needs a scratch
// buffer from them; this variant brings its own: a fixed array that lives in the library's device image, so every GPU
// that loads the library carries these B2_NOALLOC_SCRATCH_FLOATS * 4 bytes until the process exits.
//
// `g_noalloc_scratch` as spelled in host code is only a host-side handle for the symbol — handing it to a kernel would
// make the kernel fault (or scribble over host memory where the GPU can reach it).  The real device address is looked up
// once per device with cudaGetSymbolAddress while that device is current, and that address is what the kernel receives.
static __device__ __align__(256) float g_noalloc_scratch[B2_NOALLOC_SCRATCH_FLOATS];

static float* g_noalloc_ptr[64];
static bool g_noalloc_ready[64];
static std::mutex g_noalloc_mu;  // host threads of different devices may arrive at the same time

static int noalloc_scratch_for_current_device(int device, float** out) {
  B2_REQUIRE(device >= 0 && device < 64, "b2_ae_grad_noalloc: device index %d not supported", device);
  std::lock_guard<std::mutex> lock(g_noalloc_mu);
  if (!g_noalloc_ready[device]) {
    float* p = nullptr;
    B2_CUDA(cudaGetSymbolAddress((void**)&p, g_noalloc_scratch));  // device `device` is current here (DeviceGuard)
    B2_REQUIRE(p != nullptr, "b2_ae_grad_noalloc: scratch symbol resolved to NULL");
    g_noalloc_ptr[device] = p;
    g_noalloc_ready[device] = true;
  }
  *out = g_noalloc_ptr[device];
  return B2_OK;
}

extern "C" int b2_ae_grad_noalloc(b2_ae* ae, const float* x_dev, int64_t n, const b2_norm* norm, float* grads_dev, void* stream) {
  B2_NORM_F32_ONLY(norm, "b2_ae_grad_noalloc");
  B2_REQUIRE(ae && x_dev && grads_dev, "b2_ae_grad_noalloc: NULL argument");
  B2_REQUIRE(n >= 1, "b2_ae_grad_noalloc: n=%lld < 1", (long long)n);
  B2_REQUIRE(((uintptr_t)x_dev & 7) == 0, "b2_ae_grad_noalloc: x_dev must be 8-byte aligned");
  B2_REQUIRE(!norm || norm->n == 0 || norm->n == ae->desc.widths[0], "b2_ae_grad_noalloc: norm table size mismatch");
  if (ae->spec_d == 0) {
    b2_set_error("b2_ae_grad_noalloc: training is implemented for the D-14-7-7-D autoencoder with D in {18, 30}");
    return B2_EUNSUPPORTED;
  }
  const long long need = (long long)grad_grid(ae, n) * (ae->n_params + 3);
  B2_REQUIRE(need <= (long long)B2_NOALLOC_SCRATCH_FLOATS,
             "b2_ae_grad_noalloc: this launch needs %lld scratch floats (%d CTAs x %d), the built-in buffer holds %d; "
             "use b2_ae_grad with caller-provided scratch",
             need, grad_grid(ae, n), ae->n_params + 3, (int)B2_NOALLOC_SCRATCH_FLOATS);
  DeviceGuard g(ae->ctx->device);
  float* scratch = nullptr;
  int rc = noalloc_scratch_for_current_device(ae->ctx->device, &scratch);
  if (rc) return rc;
  // Nothing in the buffer has to start at zero: every CTA of this launch overwrites its whole row before the fold reads
  // exactly those rows, so whatever an earlier call left behind is never looked at.
  cudaStream_t st = (cudaStream_t)stream;
  if (ae->spec_d == 18) return grad_spec<18>(ae, x_dev, n, norm, grads_dev, scratch, st);
  return grad_spec<30>(ae, x_dev, n, norm, grads_dev, scratch, st);
}
