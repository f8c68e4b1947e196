// K1t — the Dense autoencoder D-14-7-7-D on the 5th-generation tensor cores (tcgen05.mma kind::tf32, accumulators in TMEM).
//
// Same math as ae_math.cuh (cardata-v1.py:161-167; nb2.0:1015).  BASELINE.json keeps a tcgen05 variant "only if ncu shows the FMA
// pipe rather than HBM is the limiter": it does (profiles/: fma pipe 79 % busy, math_pipe_throttle the top stall), and the FFMA2
// form needs ≈ 2.4 issue slots per multiply-add pair once the weight fetches (LDCU) are counted.  Here a thread still owns a
// row, but the four matrix products of a 128-row tile are four short chains of tcgen05.mma issued by ONE thread:
//   * the A operand of each layer — [activations | 1 | 0…] — lives in TMEM, written by the row-owning threads with tcgen05.st as a
//     (hi, lo) TF32 pair (cvt.rna + remainder): lane = row, no shared-memory staging and no transposes;
//   * the B operands (weights with the bias as one more K row) sit in shared memory for the life of the CTA in the canonical
//     K-major no-swizzle layout, again as (hi, lo);
//   * D = A_hi·B_hi + A_lo·B_hi + A_hi·B_lo (3 × K/8 MMAs, M = 128, N = 16 or 32) → an FP32 accumulator in TMEM, read back by the
//     row-owning threads with tcgen05.ld for the activation / the squared error.
// TMEM per CTA: D = 30 → 128 columns (A panel hi [0,32) + lo [32,64) + accumulator [64,96)): four CTAs share an SM's 512 columns;
// D = 18 → 64 columns (hi [0,24) + lo [24,48) + a 16-column accumulator at 48; the last layer's 32-column accumulator sits at
// [32,64), free by then because its A panel is only [0,8) + [24,32)): up to eight CTAs per SM.
#pragma once
#include "ae_math.cuh"

constexpr int TCS_ROWS = 128;      // rows per tile = TMEM lanes = threads per CTA

// GEMM shapes: K includes the bias row and zero padding to a multiple of 8; N is padded to 16 / 32 (M = 128 needs N % 16 == 0)
template <int D>
struct TcShape {
  static constexpr int K1 = (D + 1 + 7) & ~7, N1 = 16;    // 30 → 32, 18 → 24
  static constexpr int K2 = 16, N2 = 16;                   // 14 + 1 → 16
  static constexpr int K3 = 8, N3 = 16;                    // 7 + 1
  static constexpr int K4 = 8, N4 = (D + 15) & ~15;        // 30 → 32, 18 → 32
  static constexpr int OFF1 = 0;                           // float offsets of (hi, lo) panels inside the packed block
  static constexpr int OFF2 = OFF1 + 2 * N1 * K1;
  static constexpr int OFF3 = OFF2 + 2 * N2 * K2;
  static constexpr int OFF4 = OFF3 + 2 * N3 * K3;
  static constexpr int TOTAL = OFF4 + 2 * N4 * K4;
  // TMEM columns
  static constexpr int ALO = K1;                           // lo copy of the A panel right behind the hi copy
  static constexpr int ACC = 2 * K1;                       // accumulator of layers 1-3 (16 columns)
  static constexpr int ACC4 = (2 * K1 + 16 <= 64) ? 32 : 2 * K1;  // accumulator of layer 4 (32 columns)
  static constexpr int TMEM_COLS = (2 * K1 + 16 <= 64) ? 64 : 128;
  static_assert(K1 <= 32 && N4 <= 32, "panel wider than the TMEM budget");
  static_assert(ACC + 16 <= TMEM_COLS && ACC4 + N4 <= TMEM_COLS && ACC4 >= ALO + 8, "TMEM layout overlaps");
};

template <int D>
struct TcPanels {  // host image of the packed panels; a device copy hangs off the b2_ae handle and is loaded into shared memory per CTA
  float p[TcShape<D>::TOTAL];
};

// B[n][k] → canonical K-major no-swizzle tile, split into TF32 head and remainder: the 16-byte chunk (4 k's) of row n sits at
// (k/4)·(N·16) + n·16 bytes; hi block first, lo block behind it
static inline void tc_pack_panel(const float* B, int N, int K, float* out) {
  for (int n = 0; n < N; ++n)
    for (int k = 0; k < K; ++k) {
      const float w = B[(size_t)n * K + k];
      uint32_t bits;
      memcpy(&bits, &w, 4);
      uint32_t hb = (bits + 0x1000u) & 0xFFFFE000u;  // round to nearest on the 13 dropped bits; lo carries the rest exactly
      float h;
      memcpy(&h, &hb, 4);
      if (!(h == h) || h - h != 0.0f) h = w;         // never round into inf / NaN
      const size_t off = (size_t)(k >> 2) * (N * 4) + (size_t)n * 4 + (k & 3);
      out[off] = h;
      out[(size_t)N * K + off] = w - h;
    }
}

// Keras flat weights (kernel0[in,out], bias0, …) → the four (hi, lo) B panels; bias = K row `in`, padding rows / columns zero
template <int D>
static inline void tc_pack_weights(const float* w, TcPanels<D>& o) {
  using S = TcShape<D>;
  const int in[4] = {D, AE_H1, AE_H2, AE_H3}, out[4] = {AE_H1, AE_H2, AE_H3, D};
  const int K[4] = {S::K1, S::K2, S::K3, S::K4}, N[4] = {S::N1, S::N2, S::N3, S::N4}, off[4] = {S::OFF1, S::OFF2, S::OFF3, S::OFF4};
  const float* p = w;
  for (int l = 0; l < 4; ++l) {
    float B[32 * 32];
    memset(B, 0, sizeof(B));
    for (int i = 0; i < in[l]; ++i)
      for (int j = 0; j < out[l]; ++j) B[j * K[l] + i] = p[i * out[l] + j];
    p += in[l] * out[l];
    for (int j = 0; j < out[l]; ++j) B[j * K[l] + in[l]] = p[j];
    p += out[l];
    tc_pack_panel(B, N[l], K[l], o.p + off[l]);
  }
}

#ifdef __CUDACC__
// ---- tcgen05 wrappers (lane = row: 32x32b shape, each thread touches its own TMEM lane) ------------------------------------
#define TCS_R4(a, i) "r"(a[i]), "r"(a[i + 1]), "r"(a[i + 2]), "r"(a[i + 3])
#define TCS_W4(a, i) "=r"(a[i]), "=r"(a[i + 1]), "=r"(a[i + 2]), "=r"(a[i + 3])
__device__ __forceinline__ void tcs_st8(uint32_t t, const uint32_t* r) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"r"(t), TCS_R4(r, 0), TCS_R4(r, 4) : "memory");
}
__device__ __forceinline__ void tcs_st16(uint32_t t, const uint32_t* r) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};" ::"r"(t),
               TCS_R4(r, 0), TCS_R4(r, 4), TCS_R4(r, 8), TCS_R4(r, 12)
               : "memory");
}
__device__ __forceinline__ void tcs_ld8(uint32_t t, uint32_t* r) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];" : TCS_W4(r, 0), TCS_W4(r, 4) : "r"(t) : "memory");
}
__device__ __forceinline__ void tcs_ld16(uint32_t t, uint32_t* r) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
               : TCS_W4(r, 0), TCS_W4(r, 4), TCS_W4(r, 8), TCS_W4(r, 12)
               : "r"(t)
               : "memory");
}
__device__ __forceinline__ void tcs_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tcs_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tcs_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tcs_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ uint64_t tcs_bdesc(uint32_t saddr, int lbo_bytes, int sbo_bytes) {
  return (uint64_t)((saddr >> 4) & 0x3FFF) | ((uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16) | ((uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32) |
         (1ull << 46);  // descriptor version 1 (sm_100), SWIZZLE_NONE
}
__device__ __forceinline__ void tcs_mma(uint32_t d_tmem, uint32_t a_tmem, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d_tmem),
      "r"(a_tmem), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void tcs_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// TF32 head (round to nearest on the 13 dropped mantissa bits) + exact remainder.  cvt.rna.tf32.f32 expands to five SASS
// instructions on sm_100a (ncu: LOP3 + FSETP + VIADD + SEL + FADD per value, 40 % of the first version of this kernel); the
// integer form below is three.  It differs only where the rounding carries into the exponent field of the largest binade
// (|v| > 3.4e38 · (1 - 2^-12)): hi becomes inf there and the record's error NaN — such a value squared overflows binary32 anyway.
__device__ __forceinline__ void tcs_split(float v, uint32_t& hi, uint32_t& lo) {
  hi = (__float_as_uint(v) + 0x1000u) & 0xFFFFE000u;
  lo = __float_as_uint(v - __uint_as_float(hi));
}

// store KP (8, 16, 24 or 32) A-panel values of this thread's row as (hi, lo); every store is aligned to its own width
template <int KP, int ALO>
__device__ __forceinline__ void tcs_a_store(uint32_t lane_base, const float* v) {
  uint32_t hi[KP], lo[KP];
#pragma unroll
  for (int i = 0; i < KP; ++i) tcs_split(v[i], hi[i], lo[i]);
#pragma unroll
  for (int c = 0; c + 16 <= KP; c += 16) tcs_st16(lane_base + c, hi + c);
  if (KP % 16 == 8) tcs_st8(lane_base + (KP - 8), hi + (KP - 8));
  if (ALO % 16 == 0) {
#pragma unroll
    for (int c = 0; c + 16 <= KP; c += 16) tcs_st16(lane_base + ALO + c, lo + c);
    if (KP % 16 == 8) tcs_st8(lane_base + ALO + (KP - 8), lo + (KP - 8));
  } else {  // ALO % 16 == 8: one 8-wide store brings the cursor to a 16-column boundary
    tcs_st8(lane_base + ALO, lo);
#pragma unroll
    for (int c = 8; c + 16 <= KP; c += 16) tcs_st16(lane_base + ALO + c, lo + c);
    if (KP % 16 == 0) tcs_st8(lane_base + ALO + (KP - 8), lo + (KP - 8));
  }
}

// one thread: D[128 x N] = A_hi·B_hi + A_lo·B_hi + A_hi·B_lo with A = TMEM columns [0, K) / [TCS_ALO, TCS_ALO + K), then commit
template <int N, int K, int ALO, int ACC>
__device__ __forceinline__ void tcs_issue(uint32_t tbase, uint32_t b_hi_saddr, uint64_t* bar) {
  tcs_fence_after();
  constexpr uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | ((128u >> 4) << 24);
  constexpr int chunk = N * 16;  // bytes between consecutive 16-byte K-chunks (LBO); rows 8 apart are 128 B apart (SBO)
  const uint32_t b_lo_saddr = b_hi_saddr + (uint32_t)(N * K * 4);
  uint32_t acc = 0;
#pragma unroll
  for (int pass = 0; pass < 3; ++pass) {
    const uint32_t a0 = tbase + (pass == 1 ? ALO : 0);
    const uint32_t b0 = pass == 2 ? b_lo_saddr : b_hi_saddr;
#pragma unroll
    for (int j = 0; j < K / 8; ++j) {
      tcs_mma(tbase + ACC, a0 + 8 * j, tcs_bdesc(b0 + j * 2 * chunk, chunk, 128), idesc, acc);
      acc = 1;
    }
  }
  tcs_commit(bar);
}

// Per-CTA state of the tensor-core score path: TMEM base, the weights' shared-memory address, the MMA-done mbarrier.
struct TcCtx {
  uint32_t tbase;       // TMEM base address (lane 0, column 0 of this CTA's allocation)
  uint32_t lane_base;   // … of this warp's 32 lanes
  uint32_t w_saddr;     // shared-memory address of the packed panels
  uint64_t* bar;
  uint32_t parity;
};

// All 128 threads: A-panel stores are complete → one thread issues layer (N, K) → everybody waits for its accumulator.
template <int N, int K, int ALO, int ACC>
__device__ __forceinline__ void tcs_run_layer(TcCtx& c, int w_off_floats, int tid) {
  tcs_wait_st();
  tcs_fence_before();
  __syncthreads();
  if (tid == 0) tcs_issue<N, K, ALO, ACC>(c.tbase, c.w_saddr + (uint32_t)w_off_floats * 4u, c.bar);
  mbar_wait(c.bar, c.parity);
  c.parity ^= 1u;
  tcs_fence_after();
}

// Forward pass + reconstruction error of ONE row per thread, 128 rows per CTA in lock-step (every thread of the CTA must call
// this together; rows that do not exist pass zeros).  x[] holds the (already normalised) input row and is left unchanged.
// MID_FP32: the two small middle layers (14→7→7: 147 of the 525 multiply-adds) run as FFMA2 in registers, weights from the
// kernel parameters (`w`), so a tile makes TWO tensor-core round trips (layer 1, layer 4) instead of four.  The round trip —
// tcgen05.st, fence, barrier, issue, commit, mbarrier wait, tcgen05.ld — costs ≈ 950 cycles and is what bounds this kernel,
// not the instruction count (ncu, profiles/round2_notes.md).
template <int D, bool FAST_TANH, bool MID_FP32>
__device__ __forceinline__ float tcs_score_row(TcCtx& c, const AeW<D>& w, const float (&x)[D], int tid) {
  using S = TcShape<D>;
  {  // layer 1 panel: [x | 1 | 0…]
    float a[S::K1];
#pragma unroll
    for (int i = 0; i < S::K1; ++i) a[i] = i < D ? x[i] : (i == D ? 1.0f : 0.0f);
    tcs_a_store<S::K1, S::ALO>(c.lane_base, a);
  }
  tcs_run_layer<S::N1, S::K1, S::ALO, S::ACC>(c, S::OFF1, tid);
  if (MID_FP32) {
    uint32_t z[16];
    tcs_ld16(c.lane_base + S::ACC, z);
    tcs_wait_ld();
    float2 h1[AE_H1 / 2];
#pragma unroll
    for (int q = 0; q < AE_H1 / 2; ++q)
      h1[q] = tanh2_sel<FAST_TANH>(make_float2(__uint_as_float(z[2 * q]), __uint_as_float(z[2 * q + 1])));
    constexpr int Q2 = AE_P2 / 2, Q3 = AE_P3 / 2;
    float2 h2[Q2];  // 7 units padded to 8 (pad lane stays 0: zero weights, zero bias, relu)
#pragma unroll
    for (int q = 0; q < Q2; ++q) h2[q] = make_float2(w.b2[2 * q], w.b2[2 * q + 1]);
#pragma unroll
    for (int i = 0; i < AE_H1; ++i) {
      const float xi = (i & 1) ? h1[i / 2].y : h1[i / 2].x;
#pragma unroll
      for (int q = 0; q < Q2; ++q) h2[q] = fma2(splat(xi), make_float2(w.k2[i][2 * q], w.k2[i][2 * q + 1]), h2[q]);
    }
#pragma unroll
    for (int q = 0; q < Q2; ++q) h2[q] = relu2(h2[q]);
    float2 h3[Q3];
#pragma unroll
    for (int q = 0; q < Q3; ++q) h3[q] = make_float2(w.b3[2 * q], w.b3[2 * q + 1]);
#pragma unroll
    for (int i = 0; i < AE_H2; ++i) {
      const float xi = (i & 1) ? h2[i / 2].y : h2[i / 2].x;
#pragma unroll
      for (int q = 0; q < Q3; ++q) h3[q] = fma2(splat(xi), make_float2(w.k3[i][2 * q], w.k3[i][2 * q + 1]), h3[q]);
    }
    float a[8];
#pragma unroll
    for (int q = 0; q < Q3; ++q) {
      const float2 t = tanh2_sel<FAST_TANH>(h3[q]);
      a[2 * q] = t.x;
      a[2 * q + 1] = t.y;
    }
    a[7] = 1.0f;  // (the pad unit's tanh(0) = 0 is replaced by the bias-1 column)
    tcs_a_store<8, S::ALO>(c.lane_base, a);
  } else {
    {  // h1 = tanh(z1) → layer 2 panel [h1(14) | 1 | 0]
      uint32_t z[16];
      tcs_ld16(c.lane_base + S::ACC, z);
      tcs_wait_ld();
      float a[16];
#pragma unroll
      for (int q = 0; q < AE_H1 / 2; ++q) {
        const float2 t = tanh2_sel<FAST_TANH>(make_float2(__uint_as_float(z[2 * q]), __uint_as_float(z[2 * q + 1])));
        a[2 * q] = t.x;
        a[2 * q + 1] = t.y;
      }
      a[14] = 1.0f;
      a[15] = 0.0f;
      tcs_a_store<16, S::ALO>(c.lane_base, a);
    }
    tcs_run_layer<S::N2, S::K2, S::ALO, S::ACC>(c, S::OFF2, tid);
    {  // h2 = relu(z2) → layer 3 panel [h2(7) | 1]
      uint32_t z[8];
      tcs_ld8(c.lane_base + S::ACC, z);
      tcs_wait_ld();
      float a[8];
#pragma unroll
      for (int i = 0; i < 7; ++i) a[i] = fmaxf(__uint_as_float(z[i]), 0.0f);
      a[7] = 1.0f;
      tcs_a_store<8, S::ALO>(c.lane_base, a);
    }
    tcs_run_layer<S::N3, S::K3, S::ALO, S::ACC>(c, S::OFF3, tid);
    {  // h3 = tanh(z3) → layer 4 panel [h3(7) | 1]
      uint32_t z[8];
      tcs_ld8(c.lane_base + S::ACC, z);
      tcs_wait_ld();
      float a[8];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float2 t = tanh2_sel<FAST_TANH>(make_float2(__uint_as_float(z[2 * q]), __uint_as_float(z[2 * q + 1])));
        a[2 * q] = t.x;
        a[2 * q + 1] = t.y;
      }
      a[7] = 1.0f;
      tcs_a_store<8, S::ALO>(c.lane_base, a);
    }
  }
  tcs_run_layer<S::N4, S::K4, S::ALO, S::ACC4>(c, S::OFF4, tid);
  float2 se = make_float2(0.f, 0.f);
  {  // ŷ = relu(z4); squared error against x, 16 columns at a time
#pragma unroll
    for (int c0 = 0; c0 < D; c0 += 16) {
      uint32_t z[16];
      if (D - c0 > 8) tcs_ld16(c.lane_base + S::ACC4 + c0, z);  // the accumulator is N4 = 32 columns wide
      else tcs_ld8(c.lane_base + S::ACC4 + c0, z);               // D = 18: columns 16..23
      tcs_wait_ld();
#pragma unroll
      for (int j = 0; j < 16; j += 2) {
        if (c0 + j < D) {
          const float2 y = relu2(make_float2(__uint_as_float(z[j]), __uint_as_float(z[j + 1])));
          const float2 d = fma2(y, splat(-1.0f), make_float2(x[c0 + j], x[c0 + j + 1]));
          se = fma2(d, d, se);
        }
      }
    }
  }
  return __fdiv_rn(se.x + se.y, (float)D);
}

// CTA set-up / tear-down: TMEM allocation (warp 0), mbarrier, weights → shared memory.  `wsm` must be 16-byte aligned and hold
// TcShape<D>::TOTAL floats; `s_tmem` / `bar` are __shared__ variables of the caller.
template <int D>
__device__ __forceinline__ void tcs_setup(TcCtx& c, const float* __restrict__ panels, float* wsm, uint32_t* s_tmem, uint64_t* bar, int tid) {
  const int warp = tid >> 5;
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(s_tmem)), "r"(TcShape<D>::TMEM_COLS));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  if (tid == 0) {
    mbar_init(bar, 1);
    mbar_fence_init();
  }
  for (int i = tid; i < TcShape<D>::TOTAL; i += TCS_ROWS) wsm[i] = __ldg(panels + i);
  fence_proxy_async();  // generic-proxy writes of the weights → visible to the tensor core's async-proxy reads
  tcs_fence_before();
  __syncthreads();
  tcs_fence_after();
  c.tbase = *s_tmem;
  c.lane_base = c.tbase + ((uint32_t)((warp & 3) * 32) << 16);
  c.w_saddr = smem_u32(wsm);
  c.bar = bar;
  c.parity = 0;
}
template <int D>
__device__ __forceinline__ void tcs_teardown(TcCtx& c, int tid) {
  tcs_fence_before();
  __syncthreads();
  if ((tid >> 5) == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(c.tbase), "r"(TcShape<D>::TMEM_COLS));
}
#endif
