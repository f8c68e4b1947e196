// Device math of the Dense autoencoder D→14→7→7→D (tanh, relu, tanh, relu).
//
// Restates the forward pass built at
//   python-scripts/AUTOENCODER-TensorFlow-IO-Kafka/cardata-v1.py:161-167 (18-wide)
//   python-scripts/autoencoder-anomaly-detection/Sensor-Kafka-Consumer-and-TensorFlow-Model-Training.py:16-24 (30-wide)
// the per-feature scaling of normalize_fn (cardata-v1.py:68-111) and the reconstruction error
//   mse = np.mean(np.power(x - pred, 2), axis=1)   (nb2.0:1015).
//
// B200 specifics: the FP32 pipe only reaches its peak with the packed FFMA2 instruction (two FMAs per lane and
// issue slot).  Accumulators are packed along the output dimension (h[2p], h[2p+1]); the weight pair arrives
// through the kernel parameter (constant bank → uniform registers, LDCU.128) and the input is broadcast by
// FFMA2's .F32 operand form — no shared-memory or register traffic for weights at all.
#pragma once
#include "common.cuh"

constexpr int AE_H1 = 14, AE_H2 = 7, AE_H3 = 7;
constexpr int AE_P1 = 16, AE_P2 = 8, AE_P3 = 8;  // padded row lengths → 16-byte aligned LDCU.128

template <int D>
struct AeW {  // padded, kernel-parameter resident copy of the weights
  static constexpr int PD = (D + 3) & ~3;
  float k1[D][AE_P1];      // [in][out]
  float b1[AE_P1];
  float k2[AE_H1][AE_P2];  // [in][out]
  float b2[AE_P2];
  float k3[AE_H2][AE_P3];
  float b3[AE_P3];
  // layer 4 per OUTPUT PAIR, the order the second pass consumes it in: (b[2p], b[2p+1], k[0][2p], k[0][2p+1], ... k[6][2p], k[6][2p+1])
  // = 16 floats = four LDCU.128 per pair (as [in][out] rows every one of the 7 + 1 operand pairs was an LDCU.64 of its own)
  float o4[PD / 2][2 + 2 * AE_H3];
  __host__ __device__ float k4(int k, int j) const { return o4[j >> 1][2 + 2 * k + (j & 1)]; }
  __host__ __device__ float b4(int j) const { return o4[j >> 1][j & 1]; }
};

template <int D>
struct NormP {             // normalize_fn table; kind: B2_NORM_*
  int32_t kind[D];
  float lo[D];
  float range[D];
  float rinv[D];           // RN(1 / range), filled by the host
  // Branch-free form for the packed paths (norm_pair_fast / norm_rows_fast): the three kinds become constants of ONE instruction
  // sequence, so a feature costs no compare / select on its kind —  out = fma(q, m2, a2) with q = (v & mask - flo) / frange:
  //   SCALE     mask ~0, flo lo, frange range, frinv = fcorr = 1/range, m2 2, a2 -1
  //   IDENTITY  mask ~0, flo 0,  frange 1,     frinv 1, fcorr -0 (keeps the sign of a zero), m2 1, a2 -0      → v
  //   ZERO      mask 0,  flo 0,  frange 1,     frinv 1, fcorr 1,  m2 0, a2 +0                                 → +0
  uint32_t mask[D];
  float flo[D], frange[D], frinv[D], fcorr[D], m2[D], a2[D];
  // the same constants once more, grouped per feature PAIR in the order norm_pair_fast consumes them — (flo, frinv | frange, fcorr |
  // m2, a2 | mask), two values each: 16 words = four 16-byte constant loads per pair instead of seven 8-byte ones
  float pk[(D + 1) / 2][16];
};

// b2_norm (C ABI) → NormP<D>; returns whether a table was given
template <int D>
static inline int norm_fill(const b2_norm* norm, NormP<D>& nm, int n = D) {
  memset(&nm, 0, sizeof(nm));
  if (!norm || norm->n <= 0) return 0;
  for (int j = 0; j < n; ++j) {
    const int k = norm->kind[j];
    nm.kind[j] = k;
    nm.lo[j] = norm->lo[j];
    nm.range[j] = norm->range[j];
    nm.rinv[j] = 1.0f / norm->range[j];
    const bool sc = k == B2_NORM_SCALE, zero = k == B2_NORM_ZERO;
    nm.mask[j] = zero ? 0u : 0xFFFFFFFFu;
    nm.flo[j] = sc ? nm.lo[j] : 0.0f;
    nm.frange[j] = sc ? nm.range[j] : 1.0f;
    nm.frinv[j] = sc ? nm.rinv[j] : 1.0f;
    nm.fcorr[j] = sc ? nm.rinv[j] : (zero ? 1.0f : -0.0f);
    nm.m2[j] = sc ? 2.0f : (zero ? 0.0f : 1.0f);
    nm.a2[j] = sc ? -1.0f : (zero ? 0.0f : -0.0f);
  }
  for (int j = 0; j < n; ++j) {
    float* q = nm.pk[j >> 1];
    const int h = j & 1;
    q[0 + h] = nm.flo[j], q[2 + h] = nm.frinv[j], q[4 + h] = nm.frange[j], q[6 + h] = nm.fcorr[j], q[8 + h] = nm.m2[j], q[10 + h] = nm.a2[j];
    memcpy(&q[12 + h], &nm.mask[j], 4);
  }
  return 1;
}

// The reference's normalisation table forces features 0, 2, 4 and 5 to zero (cardata-v1.py:72, :78, :84, :87): kernels carry one
// instantiation with that zero set compiled in (ZM of ae_forward_rows), launched when the caller's table has exactly this set.
constexpr unsigned REF_ZERO_MASK = 0x35u;
// bit i = feature i is of kind ZERO — or ~0u when the table is not of the shape the ZM instantiations assume: every feature
// outside the zero set of kind SCALE (as in the reference's table: the instantiation then folds "* 2 - 1" into immediates), and
// the layer-1 weight rows it skips finite (+0 · w must be exactly nothing for the skip to be invisible)
template <int D, class W, class N>
static inline unsigned norm_zero_set(const N& nm, const W& w) {
  unsigned zm = 0;
  for (int i = 0; i < D && i < 32; ++i) {
    if (nm.kind[i] == B2_NORM_SCALE) continue;
    if (nm.kind[i] != B2_NORM_ZERO) return ~0u;
    zm |= 1u << i;
    for (int j = 0; j < 14; ++j)
      if (!(w.k1[i][j] - w.k1[i][j] == 0.0f)) return ~0u;
  }
  return zm;
}

// Pack Keras flat order (kernel0[in,out], bias0, kernel1, ...) into AeW<D>.
template <int D>
static inline void ae_pack_weights(const float* w, AeW<D>& o) {
  memset(&o, 0, sizeof(o));
  const float* p = w;
  for (int i = 0; i < D; ++i)
    for (int j = 0; j < AE_H1; ++j) o.k1[i][j] = *p++;
  for (int j = 0; j < AE_H1; ++j) o.b1[j] = *p++;
  for (int i = 0; i < AE_H1; ++i)
    for (int j = 0; j < AE_H2; ++j) o.k2[i][j] = *p++;
  for (int j = 0; j < AE_H2; ++j) o.b2[j] = *p++;
  for (int i = 0; i < AE_H2; ++i)
    for (int j = 0; j < AE_H3; ++j) o.k3[i][j] = *p++;
  for (int j = 0; j < AE_H3; ++j) o.b3[j] = *p++;
  for (int i = 0; i < AE_H3; ++i)
    for (int j = 0; j < D; ++j) o.o4[j >> 1][2 + 2 * i + (j & 1)] = *p++;
  for (int j = 0; j < D; ++j) o.o4[j >> 1][j & 1] = *p++;
}

#ifdef __CUDACC__
__device__ __forceinline__ float2 splat(float v) { return make_float2(v, v); }
__device__ __forceinline__ float2 fma2(float2 a, float2 b, float2 c) { return __ffma2_rn(a, b, c); }
__device__ __forceinline__ float2 mul2(float2 a, float2 b) { return __fmul2_rn(a, b); }
__device__ __forceinline__ float2 add2(float2 a, float2 b) { return __fadd2_rn(a, b); }
__device__ __forceinline__ float2 relu2(float2 a) { return make_float2(fmaxf(a.x, 0.f), fmaxf(a.y, 0.f)); }

__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float rcp_approx(float x) {
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// tanh to ≤ ~4 ulp (max abs error ≈ 1.5e-7): odd minimax polynomial below 0.6, 1 - 2/(e^{2|x|}+1) above.
// tanh.approx.f32 (MUFU.TANH, 2^-11) is far too coarse for the 1e-5 relative MSE contract.
__device__ __forceinline__ float2 tanh2(float2 x) {
  const float2 t = mul2(x, x);
  float2 p = fma2(t, splat(-6.276387590e-03f), splat(2.111621396e-02f));
  p = fma2(p, t, splat(-5.387513819e-02f));
  p = fma2(p, t, splat(1.333292401e-01f));
  p = fma2(p, t, splat(-3.333333043e-01f));
  p = mul2(p, t);
  const float2 small = fma2(p, x, x);
  const float ax = fabsf(x.x), ay = fabsf(x.y);
  const float ex = ex2_approx(ax * 2.8853900817779268f);  // e^{2|x|}
  const float ey = ex2_approx(ay * 2.8853900817779268f);
  const float bx = fmaf(-2.f, rcp_approx(ex + 1.f), 1.f);
  const float by = fmaf(-2.f, rcp_approx(ey + 1.f), 1.f);
  float2 r;
  r.x = ax < 0.6f ? small.x : copysignf(bx, x.x);
  r.y = ay < 0.6f ? small.y : copysignf(by, x.y);
  return r;
}
__device__ __forceinline__ float tanh1(float x) { return tanh2(make_float2(x, x)).x; }

// Cheaper variant: tanh(x) = 1 - 2/(2^(2x·log2 e) + 1) for every x (3 packed FP32 ops + 4 MUFU per pair, no select).
// Saturates correctly (2^y → 0 / inf), absolute error ≤ ~2.5e-7 near 0 where 1 - 2r cancels (relative error is
// unbounded for tiny |x|, which a following Dense layer does not care about).
__device__ __forceinline__ float2 tanh2_fast(float2 x) {
  const float2 y = mul2(x, splat(2.8853900817779268f));
  float2 t = make_float2(ex2_approx(y.x), ex2_approx(y.y));
  t = add2(t, splat(1.0f));
  const float2 r = make_float2(rcp_approx(t.x), rcp_approx(t.y));
  return fma2(r, splat(-2.0f), splat(1.0f));
}
template <bool FAST>
__device__ __forceinline__ float2 tanh2_sel(float2 x) {
  return FAST ? tanh2_fast(x) : tanh2(x);
}

// scale_fn(v, lo, hi) = (v - lo) / (hi - lo) * 2.0 - 1.0 with the reference's operation order and a true IEEE
// division (cardata-v1.py:68-69); `*2.0 - 1.0` folds into one FMA because the product is exact.
__device__ __forceinline__ float norm_feature(float v, int kind, float lo, float range) {
  if (kind == B2_NORM_SCALE) return fmaf(__fdiv_rn(__fsub_rn(v, lo), range), 2.0f, -1.0f);
  if (kind == B2_NORM_ZERO) return 0.0f;
  return v;
}

// Packed path for two adjacent features (x_i, x_i+1) of one row, 5 packed FP32 ops instead of two IEEE divisions:
// with rinv = RN(1/range), q0 = RN(t*rinv), r = RN(t - q0*range) (exact by FMA) and q = RN(q0 + r*rinv) is the
// correctly rounded quotient t/range (Markstein's theorem) for every finite t whose quotient neither overflows nor
// lands in the subnormal range.  A subnormal-range slip is absorbed by the final `*2 - 1` (|q| < 2^-25 gives -1 either
// way); an infinite t or an overflowing quotient shows up as q0 = ±inf, which is then the IEEE result and is selected
// explicitly (the correction step would turn it into NaN).  NaN propagates.  The other two kinds run through the same
// instructions with neutral constants (see NormP).  Hence bit-identical to norm_feature() for
// 2^-64 <= |range| <= 2^64 (beyond that the residual r itself can go subnormal); checked exhaustively on the device over all
// 2^32 inputs by b2_selftest_normalize (tests/test_normalize_gpu.py).
// MASKED = false: the caller knows that both features are of kind SCALE: no masks, the correction factor is the reciprocal itself
// and the last step is fma(q, 2, -1) with immediates — four constants per feature instead of seven
template <int D, bool MASKED = true>
__device__ __forceinline__ float2 norm_pair_fast(float2 v, const NormP<D>& nm, int i) {
  const float* __restrict__ c = nm.pk[i >> 1];  // i is even
  const float2 vm = MASKED ? make_float2(__uint_as_float(__float_as_uint(v.x) & __float_as_uint(c[12])), __uint_as_float(__float_as_uint(v.y) & __float_as_uint(c[13])))
                           : v;
  const float2 t = fma2(make_float2(c[0], c[1]), splat(-1.0f), vm);  // v - lo (exactly one rounding, like the reference's subtraction)
  const float2 q0 = mul2(t, make_float2(c[2], c[3]));
  const float2 r = fma2(make_float2(-q0.x, -q0.y), make_float2(c[4], c[5]), t);
  float2 q = fma2(r, MASKED ? make_float2(c[6], c[7]) : make_float2(c[2], c[3]), q0);
  if (fabsf(q0.x) == __int_as_float(0x7F800000)) q.x = q0.x;
  if (fabsf(q0.y) == __int_as_float(0x7F800000)) q.y = q0.y;
  if (!MASKED) return fma2(q, splat(2.0f), splat(-1.0f));
  return fma2(q, make_float2(c[8], c[9]), make_float2(c[10], c[11]));
}

// Forward pass for R records living in shared-memory rows row[r] (D floats each, 8-byte aligned).
//
// FFMA2 packing is along the OUTPUT dimension: accumulator pair (h[2p], h[2p+1]) += x_i (broadcast, .F32 operand
// form) * (k[i][2p], k[i][2p+1]) (one 64-bit uniform-register pair).  The input pair (x_i, x_i+1) of one LDS.64 is
// consumed as-is, so no register shuffling is needed; R = 2 rows per thread share every LDCU.128 weight fetch.
//   pass 1 streams x (normalising it in place when NORM) into the layer-1 accumulators,
//   pass 2 produces one output pair at a time, accumulates (x - y)^2 and (RECON) overwrites x with y.
// ZM (only with NORM): bit i set = the normalisation table forces feature i to +0 (kind ZERO — four of the reference's eighteen,
// cardata-v1.py:72-87).  Such an input adds exactly nothing to layer 1 (+0 · w for finite w), so its multiply-adds, its weight
// fetches and its share of the normalisation arithmetic are left out at compile time; the launch selects a ZM instantiation only
// when the table's zero set matches and the skipped weights are finite, and the results are bit-identical to ZM = 0.
template <int D, int R, bool NORM, bool RECON, bool FAST_TANH, unsigned ZM = 0u>
__device__ __forceinline__ void ae_forward_rows(const AeW<D>& w, const NormP<D>& nm, float* const (&row)[R], float (&mse)[R]) {
  static_assert(D % 2 == 0, "feature count must be even (8-byte row alignment)");
  static_assert(ZM == 0u || NORM, "a zero set is a property of the normalisation table");
  constexpr int Q1 = AE_H1 / 2, Q2 = AE_P2 / 2, Q3 = AE_P3 / 2;
  float2 h1[R][Q1];
#pragma unroll
  for (int r = 0; r < R; ++r)
#pragma unroll
    for (int q = 0; q < Q1; ++q) h1[r][q] = make_float2(w.b1[2 * q], w.b1[2 * q + 1]);
#pragma unroll
  for (int i = 0; i < D; i += 2) {
#pragma unroll
    for (int r = 0; r < R; ++r) {
      if (((ZM >> i) & 3u) == 3u) {  // both features forced to zero: nothing to compute, the row just receives the zeros
        *reinterpret_cast<float2*>(row[r] + i) = make_float2(0.f, 0.f);
        continue;
      }
      float2 xv = *reinterpret_cast<const float2*>(row[r] + i);
      if (NORM) {
        // with a compiled-in zero set (launched for tables whose other features are all of kind SCALE), a pair outside it needs
        // neither masks nor the per-kind constants
        xv = (ZM != 0u && ((ZM >> i) & 3u) == 0u) ? norm_pair_fast<D, false>(xv, nm, i) : norm_pair_fast<D, true>(xv, nm, i);
        *reinterpret_cast<float2*>(row[r] + i) = xv;
      }
      if (!((ZM >> i) & 1u)) {
#pragma unroll
        for (int q = 0; q < Q1; ++q) h1[r][q] = fma2(splat(xv.x), make_float2(w.k1[i][2 * q], w.k1[i][2 * q + 1]), h1[r][q]);
      }
      if (!((ZM >> (i + 1)) & 1u)) {
#pragma unroll
        for (int q = 0; q < Q1; ++q)
          h1[r][q] = fma2(splat(xv.y), make_float2(w.k1[i + 1][2 * q], w.k1[i + 1][2 * q + 1]), h1[r][q]);
      }
    }
  }
#pragma unroll
  for (int r = 0; r < R; ++r)
#pragma unroll
    for (int q = 0; q < Q1; ++q) h1[r][q] = tanh2_sel<FAST_TANH>(h1[r][q]);

  float2 h2[R][Q2];  // 7 units padded to 8 (pad lane stays 0: zero weights, zero bias, relu)
#pragma unroll
  for (int r = 0; r < R; ++r)
#pragma unroll
    for (int q = 0; q < Q2; ++q) h2[r][q] = make_float2(w.b2[2 * q], w.b2[2 * q + 1]);
#pragma unroll
  for (int i = 0; i < AE_H1; ++i)
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const float xi = (i & 1) ? h1[r][i / 2].y : h1[r][i / 2].x;
#pragma unroll
      for (int q = 0; q < Q2; ++q) h2[r][q] = fma2(splat(xi), make_float2(w.k2[i][2 * q], w.k2[i][2 * q + 1]), h2[r][q]);
    }
#pragma unroll
  for (int r = 0; r < R; ++r)
#pragma unroll
    for (int q = 0; q < Q2; ++q) h2[r][q] = relu2(h2[r][q]);

  float2 h3[R][Q3];
#pragma unroll
  for (int r = 0; r < R; ++r)
#pragma unroll
    for (int q = 0; q < Q3; ++q) h3[r][q] = make_float2(w.b3[2 * q], w.b3[2 * q + 1]);
#pragma unroll
  for (int i = 0; i < AE_H2; ++i)
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const float xi = (i & 1) ? h2[r][i / 2].y : h2[r][i / 2].x;
#pragma unroll
      for (int q = 0; q < Q3; ++q) h3[r][q] = fma2(splat(xi), make_float2(w.k3[i][2 * q], w.k3[i][2 * q + 1]), h3[r][q]);
    }
#pragma unroll
  for (int r = 0; r < R; ++r)
#pragma unroll
    for (int q = 0; q < Q3; ++q) h3[r][q] = tanh2_sel<FAST_TANH>(h3[r][q]);

  // compiler barrier: re-read x from shared memory in pass 2 instead of keeping D values per row live in registers
  asm volatile("" ::: "memory");
  float2 se[R];
#pragma unroll
  for (int r = 0; r < R; ++r) se[r] = make_float2(0.f, 0.f);
#pragma unroll
  for (int j = 0; j < D; j += 2) {
#pragma unroll
    for (int r = 0; r < R; ++r) {
      float2 y = make_float2(w.o4[j / 2][0], w.o4[j / 2][1]);
#pragma unroll
      for (int k = 0; k < AE_H3; ++k) {
        const float hk = (k & 1) ? h3[r][k / 2].y : h3[r][k / 2].x;
        y = fma2(splat(hk), make_float2(w.o4[j / 2][2 + 2 * k], w.o4[j / 2][3 + 2 * k]), y);
      }
      y = relu2(y);
      const float2 xv = *reinterpret_cast<const float2*>(row[r] + j);
      const float2 d = fma2(y, splat(-1.0f), xv);  // x - y, one packed op
      se[r] = fma2(d, d, se[r]);
      if (RECON) *reinterpret_cast<float2*>(row[r] + j) = y;
    }
  }
#pragma unroll
  for (int r = 0; r < R; ++r) mse[r] = __fdiv_rn(se[r].x + se[r].y, (float)D);
}
// ---- two rows per thread, packed along the ROW dimension -------------------------------------------------------------------
// row[i] = (x_i of record A, x_i of record B): the lane's two records interleaved element by element, so one LDS.64 delivers an
// operand pair as it is.  Every FFMA2 multiplies that pair by ONE weight (broadcast .F32 operand straight from the constant
// bank), i.e. a weight is fetched once per two records and an LDCU.128 feeds four FFMA2 — a quarter of the weight fetches per
// record of the output-packed form above, which is what the fused decode kernel (issue-bound, one record per lane before) needs.
// Per record the products are accumulated in the same order as in ae_forward_rows, and the squared error keeps its even / odd
// partial sums, so both forms give the same bits.
template <int D>
__device__ __forceinline__ float2 norm_rows_fast(float2 v, const NormP<D>& nm, int i) {  // feature i of two records
  const uint32_t mk = nm.mask[i];
  const float2 vm = make_float2(__uint_as_float(__float_as_uint(v.x) & mk), __uint_as_float(__float_as_uint(v.y) & mk));
  const float2 t = fma2(splat(nm.flo[i]), splat(-1.0f), vm);  // v - lo, one rounding
  const float2 q0 = mul2(t, splat(nm.frinv[i]));
  const float2 r = fma2(make_float2(-q0.x, -q0.y), splat(nm.frange[i]), t);
  float2 q = fma2(r, splat(nm.fcorr[i]), q0);
  if (fabsf(q0.x) == __int_as_float(0x7F800000)) q.x = q0.x;
  if (fabsf(q0.y) == __int_as_float(0x7F800000)) q.y = q0.y;
  return fma2(q, splat(nm.m2[i]), splat(nm.a2[i]));
}

template <int D, bool NORM, bool FAST_TANH>
__device__ __forceinline__ void ae_forward_pair(const AeW<D>& w, const NormP<D>& nm, float2* __restrict__ row, float (&mse)[2]) {
  // Layer 1 is taken in two halves of 7 units, each folded into layer 2's accumulators as soon as it has gone through tanh: the
  // live state is 14 + 14 + 2 registers instead of 28 + 14 + 2 (x is read from shared memory once per half).  That is what lets
  // the fused kernel keep two records per lane within 80 registers, i.e. 24 warps per SM.
  constexpr int HALF = AE_H1 / 2;
  float2 h2[AE_H2];
#pragma unroll
  for (int j = 0; j < AE_H2; ++j) h2[j] = splat(w.b2[j]);
#pragma unroll
  for (int half = 0; half < 2; ++half) {
    float2 h1[HALF];
#pragma unroll
    for (int j = 0; j < HALF; ++j) h1[j] = splat(w.b1[half * HALF + j]);
#pragma unroll
    for (int i = 0; i < D; ++i) {
      float2 xv = row[i];
      if (NORM && half == 0) {
        xv = norm_rows_fast<D>(xv, nm, i);
        row[i] = xv;
      }
#pragma unroll
      for (int j = 0; j < HALF; ++j) h1[j] = fma2(xv, splat(w.k1[i][half * HALF + j]), h1[j]);
    }
#pragma unroll
    for (int j = 0; j < HALF; ++j) h1[j] = tanh2_sel<FAST_TANH>(h1[j]);
    // layer 2 accumulates over its inputs in index order: half 0 contributes inputs 0..6 first, half 1 inputs 7..13 — the same
    // order of additions as the one-pass form
#pragma unroll
    for (int i = 0; i < HALF; ++i)
#pragma unroll
      for (int j = 0; j < AE_H2; ++j) h2[j] = fma2(h1[i], splat(w.k2[half * HALF + i][j]), h2[j]);
    if (half == 0) asm volatile("" ::: "memory");  // read x again for the second half instead of keeping it in registers
  }
#pragma unroll
  for (int j = 0; j < AE_H2; ++j) h2[j] = relu2(h2[j]);

  float2 h3[AE_H3];
#pragma unroll
  for (int j = 0; j < AE_H3; ++j) h3[j] = splat(w.b3[j]);
#pragma unroll
  for (int i = 0; i < AE_H2; ++i)
#pragma unroll
    for (int j = 0; j < AE_H3; ++j) h3[j] = fma2(h2[i], splat(w.k3[i][j]), h3[j]);
#pragma unroll
  for (int j = 0; j < AE_H3; ++j) h3[j] = tanh2_sel<FAST_TANH>(h3[j]);

  asm volatile("" ::: "memory");  // re-read x from shared memory below instead of keeping D pairs live
  float2 se_even = make_float2(0.f, 0.f), se_odd = make_float2(0.f, 0.f);
#pragma unroll
  for (int j = 0; j < D; ++j) {
    float2 y = splat(w.b4(j));
#pragma unroll
    for (int k = 0; k < AE_H3; ++k) y = fma2(h3[k], splat(w.k4(k, j)), y);
    y = relu2(y);
    const float2 d = fma2(y, splat(-1.0f), row[j]);
    if (j & 1) se_odd = fma2(d, d, se_odd);
    else se_even = fma2(d, d, se_even);
  }
  mse[0] = __fdiv_rn(se_even.x + se_odd.x, (float)D);
  mse[1] = __fdiv_rn(se_even.y + se_odd.y, (float)D);
}
#endif
