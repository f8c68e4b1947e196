// K1w — the Dense autoencoder D-14-7-7-D on warp-level tensor-core MMAs (mma.sync.m16n8k8 TF32, accumulators in registers).
//
// Same math as ae_math.cuh (cardata-v1.py:161-167; nb2.0:1015).  Why a third form next to FFMA2 (ae_math.cuh) and tcgen05 (ae_tc.cuh):
// the fused decode kernel is ISSUE bound (ncu: 83 % of the issue slots, 3/4 of the instructions are the score phase at one record
// per lane), and its warps are independent workers.  The tcgen05 form removes the multiply-adds from the issue stream but needs a
// CTA-wide 128-row tile and a ≈ 950-cycle TMEM round trip per layer, which costs the warps their independence (measured: slower).
// A warp-level MMA keeps the warp as the unit: its 32 decoded rows (shared-memory tile, row stride D) are two 16-row M tiles, and
// the whole stack is a chain of register-to-register MMAs — 18 + 6 + 3 + 9 per M tile at D = 18 instead of 263 FFMA2 + weight
// fetches per ROW.  The accumulator layout of one layer IS the A layout of the next once the K index is permuted:
//   C fragment: c0, c1 = (row g, cols 2t, 2t+1), c2, c3 = (row g+8, same cols)         g = lane >> 2, t = lane & 3
//   A fragment: a0 = (row g, k t), a1 = (row g+8, k t), a2 = (row g, k t+4), a3 = (row g+8, k t+4)
// so with MMA k index t ↔ unit 8s+2t and t+4 ↔ unit 8s+2t+1 (the rows of B permuted to match, done once per CTA) a0..a3 =
// c0, c2, c1, c3: the activation writes its results in that order and nothing else moves.  Layer 1 takes the same elements from the
// tile, which therefore holds the warp's rows INTERLEAVED in pairs (rows g and g+8 of an M tile element by element:
// tile[pair][col][slot], see mma_row_offset) so that ONE 16-byte load is an A fragment; those x values are also exactly the elements
// the lane needs for (x - y)^2 at the end.
// Precision: 3-pass split TF32, D = A_lo·B_hi + A_hi·B_lo + A_hi·B_hi with FP32 accumulation (A split by truncation on the fly, B
// split with round-to-nearest once per CTA); biases are the initial accumulator, so they stay full FP32.
#pragma once
#include "ae_math.cuh"

template <int D>
struct MmaShape {
  static constexpr int KS1 = (D + 7) / 8, NT1 = 2;  // D → 14 (16)
  static constexpr int KS2 = 2, NT2 = 1;            // 14 (16) → 7 (8)
  static constexpr int KS3 = 1, NT3 = 1;            // 7 (8) → 7 (8)
  static constexpr int KS4 = 1, NT4 = (D + 7) / 8;  // 7 (8) → D
  static constexpr int F1 = 0, F2 = F1 + KS1 * NT1, F3 = F2 + KS2 * NT2, F4 = F3 + KS3 * NT3;
  static constexpr int NFRAG = F4 + KS4 * NT4;      // B fragments: 12 (D = 18), 15 (D = 30)
  static constexpr int B1 = 0, B2 = B1 + NT1, B3 = B2 + NT2, B4 = B3 + NT3;
  static constexpr int NBIAS = B4 + NT4;            // bias pairs per t: 7 / 8
  // shared-memory table: float4 frag[NFRAG][32] (b0_hi, b1_hi, b0_lo, b1_lo of the lane), then float4 bias[NBIAS][4] = the
  // initial accumulator (b[2t], b[2t+1], b[2t], b[2t+1]) of the lanes with that t
  static constexpr int BYTES = NFRAG * 32 * 16 + NBIAS * 4 * 16;
};

// Tile layout: row r (0..31) of the warp's chunk lives in pair (r >> 4) * 8 + (r & 7), slot (r >> 3) & 1; element j of it at float
// offset pair * 2 D + 2 j + slot.  Returns the float offset of element 0; consecutive elements are 2 floats apart.
template <int D>
__host__ __device__ constexpr int mma_row_offset(int r) {
  return ((r >> 4) * 8 + (r & 7)) * 2 * D + ((r >> 3) & 1);
}

#ifdef __CUDACC__
__device__ __forceinline__ float tf32_rna(float w) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(w));
  return __uint_as_float(r);
}

// D (+)= A·B, m16n8k8, TF32 inputs (the low 13 mantissa bits of an operand register are not used), FP32 accumulate
__device__ __forceinline__ void mma_tf32(float (&d)[4], const float (&a)[4], float b0, float b1, const float (&c)[4]) {
  asm("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%10,%11,%12,%13};"
      : "=f"(d[0]), "=f"(d[1]), "=f"(d[2]), "=f"(d[3])
      : "r"(__float_as_uint(a[0])), "r"(__float_as_uint(a[1])), "r"(__float_as_uint(a[2])), "r"(__float_as_uint(a[3])),
        "r"(__float_as_uint(b0)), "r"(__float_as_uint(b1)), "f"(c[0]), "f"(c[1]), "f"(c[2]), "f"(c[3]));
}

// Fills the CTA's fragment table from the kernel-parameter copy of the weights (all threads of the CTA; the caller synchronises).
// Entry (f, lane): the B fragment of k-step s / n-tile j of its layer: b0 = W[unit(s, t)][8j + g], b1 = W[unit(s, t + 4)][8j + g]
// with unit(s, k) = 8s + 2k for k < 4 and 8s + 2(k - 4) + 1 otherwise; zero outside the layer's real shape.
template <int D>
__device__ __forceinline__ void mma_fill_table(const AeW<D>& w, uint8_t* tab, int tid, int nthreads) {
  using S = MmaShape<D>;
  float4* frag = reinterpret_cast<float4*>(tab);
  float4* bias = reinterpret_cast<float4*>(tab + S::NFRAG * 32 * 16);
  for (int e = tid; e < S::NFRAG * 32; e += nthreads) {
    const int f = e >> 5, l = e & 31, g = l >> 2, t = l & 3;
    int s, j, layer;
    if (f < S::F2) layer = 1, s = (f - S::F1) / S::NT1, j = (f - S::F1) % S::NT1;
    else if (f < S::F3) layer = 2, s = (f - S::F2) / S::NT2, j = (f - S::F2) % S::NT2;
    else if (f < S::F4) layer = 3, s = 0, j = 0;
    else layer = 4, s = 0, j = f - S::F4;
    const int u0 = 8 * s + 2 * t, u1 = u0 + 1, n = 8 * j + g;
    float v0 = 0.f, v1 = 0.f;
    if (layer == 1) {
      if (n < AE_H1) {
        if (u0 < D) v0 = w.k1[u0][n];
        if (u1 < D) v1 = w.k1[u1][n];
      }
    } else if (layer == 2) {
      if (n < AE_H2) {
        if (u0 < AE_H1) v0 = w.k2[u0][n];
        if (u1 < AE_H1) v1 = w.k2[u1][n];
      }
    } else if (layer == 3) {
      if (n < AE_H3) {
        if (u0 < AE_H2) v0 = w.k3[u0][n];
        if (u1 < AE_H2) v1 = w.k3[u1][n];
      }
    } else {
      if (n < D) {
        if (u0 < AE_H3) v0 = w.k4(u0, n);
        if (u1 < AE_H3) v1 = w.k4(u1, n);
      }
    }
    const float h0 = tf32_rna(v0), h1 = tf32_rna(v1);
    frag[e] = make_float4(h0, h1, v0 - h0, v1 - h1);  // the remainders are exact; the MMA reads their leading 11 bits
  }
  for (int e = tid; e < S::NBIAS * 4; e += nthreads) {
    const int b = e >> 2, t = e & 3;
    float v0 = 0.f, v1 = 0.f;
    if (b < S::B2) {
      const int n = 8 * (b - S::B1) + 2 * t;
      if (n < AE_H1) v0 = w.b1[n];
      if (n + 1 < AE_H1) v1 = w.b1[n + 1];
    } else if (b < S::B3) {
      const int n = 2 * t;
      if (n < AE_H2) v0 = w.b2[n];
      if (n + 1 < AE_H2) v1 = w.b2[n + 1];
    } else if (b < S::B4) {
      const int n = 2 * t;
      if (n < AE_H3) v0 = w.b3[n];
      if (n + 1 < AE_H3) v1 = w.b3[n + 1];
    } else {
      const int n = 8 * (b - S::B4) + 2 * t;
      if (n < D) v0 = w.b4(n);
      if (n + 1 < D) v1 = w.b4(n + 1);
    }
    bias[e] = make_float4(v0, v1, v0, v1);
  }
}

// head / remainder of an activation for the split product (truncation: hi keeps the leading 11 bits, lo = a - hi exactly)
__device__ __forceinline__ void mma_split(const float (&a)[4], float (&hi)[4], float (&lo)[4]) {
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    hi[i] = __uint_as_float(__float_as_uint(a[i]) & 0xFFFFE000u);
    lo[i] = a[i] - hi[i];
  }
}

// One layer for the warp's two M tiles: acc[m][j] = bias + Σ_s A[m][s] · B[s][j]  (KS k-steps, NT n-tiles).
template <int KS, int NT>
__device__ __forceinline__ void mma_layer(const float4* __restrict__ frag, const float4* __restrict__ bias, int lane, const float (&a)[2][KS][4],
                                          float (&acc)[2][NT][4]) {
  const int t = lane & 3;
#pragma unroll
  for (int s = 0; s < KS; ++s) {
    float hi[2][4], lo[2][4];
#pragma unroll
    for (int m = 0; m < 2; ++m) mma_split(a[m][s], hi[m], lo[m]);
#pragma unroll
    for (int j = 0; j < NT; ++j) {
      const float4 w = frag[(s * NT + j) * 32 + lane];
      float c0[4];
      if (s == 0) {
        const float4 b = bias[j * 4 + t];
        c0[0] = b.x, c0[1] = b.y, c0[2] = b.z, c0[3] = b.w;
      }
#pragma unroll
      for (int m = 0; m < 2; ++m) {
        if (s == 0) mma_tf32(acc[m][j], lo[m], w.x, w.y, c0);  // small terms first; the bias is the initial accumulator
        else mma_tf32(acc[m][j], lo[m], w.x, w.y, acc[m][j]);
        mma_tf32(acc[m][j], hi[m], w.z, w.w, acc[m][j]);
        mma_tf32(acc[m][j], hi[m], w.x, w.y, acc[m][j]);
      }
    }
  }
}

// tanh of an accumulator tile, written in the A-fragment order of the next layer (c0, c2, c1, c3).  The last step of the fast
// form is done per element so that each result can land in its own register.
template <bool FT>
__device__ __forceinline__ void mma_tanh_c_to_a(const float (&c)[4], float (&a)[4]) {
  if (FT) {
    const float2 y0 = mul2(make_float2(c[0], c[1]), splat(2.8853900817779268f)), y1 = mul2(make_float2(c[2], c[3]), splat(2.8853900817779268f));
    const float2 e0 = add2(make_float2(ex2_approx(y0.x), ex2_approx(y0.y)), splat(1.0f));
    const float2 e1 = add2(make_float2(ex2_approx(y1.x), ex2_approx(y1.y)), splat(1.0f));
    a[0] = fmaf(rcp_approx(e0.x), -2.0f, 1.0f);
    a[2] = fmaf(rcp_approx(e0.y), -2.0f, 1.0f);
    a[1] = fmaf(rcp_approx(e1.x), -2.0f, 1.0f);
    a[3] = fmaf(rcp_approx(e1.y), -2.0f, 1.0f);
  } else {
    const float2 p = tanh2(make_float2(c[0], c[1])), q = tanh2(make_float2(c[2], c[3]));
    a[0] = p.x, a[1] = q.x, a[2] = p.y, a[3] = q.y;
  }
}

// Scores the warp's 32 rows (shared memory, interleaved pair layout, already normalised).  Returns, in EVERY lane, the mean
// squared reconstruction error of row `lane` (the lane that decoded it).  Rows are independent inside the MMAs, so stale rows of
// lanes without a record cannot disturb the others.
template <int D, bool FT>
__device__ __forceinline__ float mma_score_warp(const float* __restrict__ xt, const uint8_t* __restrict__ tab, int lane) {
  using S = MmaShape<D>;
  static_assert(D % 2 == 0 && D <= 32, "feature count");
  const float4* frag = reinterpret_cast<const float4*>(tab);
  const float4* bias = reinterpret_cast<const float4*>(tab + S::NFRAG * 32 * 16);
  const int g = lane >> 2, t = lane & 3;
  // layer-1 A fragments: (row g, col c), (row g+8, col c), (row g, col c+1), (row g+8, col c+1) with c = 8s + 2t are 16 contiguous
  // bytes of the pair's line; columns beyond D are zero (never read: they belong to the next pair, which could hold NaN, and
  // NaN · 0 would leak into this row)
  float x[2][S::KS1][4];
#pragma unroll
  for (int m = 0; m < 2; ++m)
#pragma unroll
    for (int s = 0; s < S::KS1; ++s) {
      const int col = 8 * s + 2 * t;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (8 * s + 8 <= D || col < D) v = *reinterpret_cast<const float4*>(xt + (8 * m + g) * 2 * D + 2 * col);
      x[m][s][0] = v.x, x[m][s][1] = v.y, x[m][s][2] = v.z, x[m][s][3] = v.w;
    }
  float h1[2][S::NT1][4];
  mma_layer<S::KS1, S::NT1>(frag + S::F1 * 32, bias + S::B1 * 4, lane, x, h1);
  float a2[2][S::KS2][4];
#pragma unroll
  for (int m = 0; m < 2; ++m)
#pragma unroll
    for (int j = 0; j < S::NT1; ++j) mma_tanh_c_to_a<FT>(h1[m][j], a2[m][j]);
  float h2[2][S::NT2][4];
  mma_layer<S::KS2, S::NT2>(frag + S::F2 * 32, bias + S::B2 * 4, lane, a2, h2);
  float a3[2][S::KS3][4];
#pragma unroll
  for (int m = 0; m < 2; ++m) {
    a3[m][0][0] = fmaxf(h2[m][0][0], 0.f);
    a3[m][0][1] = fmaxf(h2[m][0][2], 0.f);
    a3[m][0][2] = fmaxf(h2[m][0][1], 0.f);
    a3[m][0][3] = fmaxf(h2[m][0][3], 0.f);
  }
  float h3[2][S::NT3][4];
  mma_layer<S::KS3, S::NT3>(frag + S::F3 * 32, bias + S::B3 * 4, lane, a3, h3);
  float a4[2][S::KS4][4];
#pragma unroll
  for (int m = 0; m < 2; ++m) mma_tanh_c_to_a<FT>(h3[m][0], a4[m][0]);
  float y[2][S::NT4][4];
  mma_layer<S::KS4, S::NT4>(frag + S::F4 * 32, bias + S::B4 * 4, lane, a4, y);
  // squared error: y tile j = (row g: cols c, c+1 | row g+8: cols c, c+1), x fragment j = (g, c), (g+8, c), (g, c+1), (g+8, c+1);
  // padded columns give relu(0) - 0
  float p[4];  // partial sums over this lane's columns for rows g, g+8, g+16, g+24
#pragma unroll
  for (int m = 0; m < 2; ++m) {
    float2 s0 = make_float2(0.f, 0.f), s1 = s0;
#pragma unroll
    for (int j = 0; j < S::NT4; ++j) {
      float2 d0, d1;
      d0.x = x[m][j][0] - fmaxf(y[m][j][0], 0.f);
      d0.y = x[m][j][2] - fmaxf(y[m][j][1], 0.f);
      d1.x = x[m][j][1] - fmaxf(y[m][j][2], 0.f);
      d1.y = x[m][j][3] - fmaxf(y[m][j][3], 0.f);
      s0 = fma2(d0, d0, s0);
      s1 = fma2(d1, d1, s1);
    }
    p[2 * m] = s0.x + s0.y;
    p[2 * m + 1] = s1.x + s1.y;
  }
  // sum over the four lanes of a quad, transposing on the way: lane t ends up with the total of p[t], i.e. of row g + 8t
  constexpr unsigned FULL = 0xffffffffu;
  const bool o1 = t & 1, o2 = t & 2;
  float ka = o1 ? p[1] : p[0], ga = o1 ? p[0] : p[1];
  float kb = o1 ? p[3] : p[2], gb = o1 ? p[2] : p[3];
  ka += __shfl_xor_sync(FULL, ga, 1);
  kb += __shfl_xor_sync(FULL, gb, 1);
  float kc = o2 ? kb : ka;
  const float gc = o2 ? ka : kb;
  kc += __shfl_xor_sync(FULL, gc, 2);
  const float mse = __fdiv_rn(kc, (float)D);
  // row r sits in lane 4 (r & 7) + (r >> 3)
  return __shfl_sync(FULL, mse, 4 * (lane & 7) + (lane >> 3));
}
#endif
