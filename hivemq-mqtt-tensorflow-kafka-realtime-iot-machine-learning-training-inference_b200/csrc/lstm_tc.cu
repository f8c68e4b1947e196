// K4t — the LSTM stack forward on the 5th-generation tensor cores (tcgen05.mma kind::tf32, accumulators in TMEM).
//
// Same math as lstm.cu (LSTM-TensorFlow-IO-Kafka/cardata-v1.py:169-176; gate order i,f,c,o; relu activation, sigmoid gates;
// SURVEY.md App. F) — this is the one genuinely GEMM-shaped piece of the hot path (36 480 FLOP per record-step), and its parity
// budget (1e-4) leaves room for split-precision TF32, which the 1e-5 budget of the Dense autoencoder does not.
//
// One CTA = 128 sequences = the 128 TMEM lanes, two threads per sequence (t and t + 128: warps w and w + 4 address the same 32
// lanes) that split every layer's units between them — twice the warps to hide the epilogue's MUFU latency.  For the whole time loop:
//   * the A operand of every layer-step GEMM, [bias-1 | x_t | h_prev ...], lives in TMEM and is written by the owning thread
//     with tcgen05.st as a (hi, lo) TF32 pair — no shared-memory staging, no transposes; consecutive layers overlap their
//     column ranges so that h1 written once is both layer 1's recurrent input and layer 2's input (ditto h3, h4);
//   * the weights of all five GEMMs sit in shared memory for the life of the CTA in the canonical K-major no-swizzle UMMA
//     layout, again as (hi, lo); bias is a weight row multiplied by a constant-1 column of A;
//   * thread 0 issues the 3-pass product  A_hi·B_hi + A_lo·B_hi + A_hi·B_lo  (K/8 MMAs per pass, M = 128, N = 4u) into one TMEM
//     accumulator and commits to an mbarrier; every thread then pulls its row of gate pre-activations back with tcgen05.ld, 32
//     columns (= 8 units × 4 gates, columns are gate-interleaved) at a time, applies the gates with c in registers, and
//     writes h_new straight back into the TMEM A columns.
// Algorithmic bytes per record-step: 4F read + 4F written; weights are read once per CTA.
//
// Two tiles per CTA (G = 2, the default): ncu shows the threads of the one-tile kernel parked on the MMA-done mbarrier for ≈ 45 %
// of their time while the tensor pipe is 20 % busy — the per-layer chain issue → MMA → commit → read-back → gates is serial.  The
// weights fill 174 KB of shared memory, so a second CTA cannot share the SM; instead one CTA of 512 threads runs TWO 128-sequence
// tiles, one per group of 256 threads, each with its own A panel in TMEM (2 x 160 columns) and ONE accumulator (128 columns)
// that the groups take in turns: a group pulls its whole accumulator into registers right after its MMAs complete and hands the
// accumulator to the other group, whose MMAs then run under this group's gate arithmetic.  TMEM: 320 + 128 = 448 of 512 columns.
#include <type_traits>

#include "lstm_tc.h"

constexpr int TC_NT = 128;      // sequences per tile = TMEM lanes
constexpr int TC_THREADS = 256; // threads per tile: thread t and t + 128 share TMEM lane t and split the units between them
constexpr int TC_ALO = 80;    // TMEM column of the "lo" copy of a tile's A panel (hi copy at 0); the panel is 80 columns wide
constexpr int TC_PANEL = 160; // TMEM columns of one tile's A panel (hi + lo); tile g's panel starts at column g * TC_PANEL
constexpr int TC_D = 320;     // TMEM column of the accumulator (≤ 128 columns), shared by the tiles of a CTA
constexpr int TC_TMEM_COLS = 512;

struct LstmTcArgs {
  const float* x;
  float* y;
  long long batch;
  long long seq_stride;  // rows between the starts of consecutive sequences: T for [B,T,F] input, `shift` for sliding windows
  int T, F;
  const float* packed;  // [Σ 2·N·K] canonical-layout weights (hi, lo) for L1, L2, L3, L4, Dense
  int packed_floats;
  TcLayer layer[5];
};

// ---- tcgen05 wrappers ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void tmem_st2(uint32_t taddr, const uint32_t* r) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x2.b32 [%0], {%1,%2};" ::"r"(taddr), "r"(r[0]), "r"(r[1]) : "memory");
}
__device__ __forceinline__ void tmem_st8(uint32_t taddr, const uint32_t* r) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]),
               "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7])
               : "memory");
}
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t* r) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};" ::"r"(taddr),
               "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]), "r"(r[10]),
               "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
               : "memory");
}

__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,"
      "%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]),
        "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]),
        "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]),
        "=r"(r[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tc_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tc_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

__device__ __forceinline__ uint64_t tc_bdesc(uint32_t saddr, int lbo_bytes, int sbo_bytes) {
  return (uint64_t)((saddr >> 4) & 0x3FFF) | ((uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16) | ((uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32) |
         (1ull << 46);  // descriptor version 1 (sm_100), SWIZZLE_NONE
}
__device__ __forceinline__ void tc_mma(uint32_t d_tmem, uint32_t a_tmem, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d_tmem),
      "r"(a_tmem), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void tc_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// split one fp32 into a TF32-exact head (round to nearest) and the remainder
// (integer rounding: cvt.rna.tf32.f32 expands to five SASS instructions on sm_100a, this is three; it differs only where the
// rounding carries out of the largest binade, |v| > 3.4e38 · (1 - 2^-12))
__device__ __forceinline__ void tf32_split(float v, uint32_t& hi, uint32_t& lo) {
  hi = (__float_as_uint(v) + 0x1000u) & 0xFFFFE000u;
  lo = __float_as_uint(v - __uint_as_float(hi));
}
// store n (8 or 16) consecutive A-panel values of this thread's lane as (hi, lo)
template <int N>
__device__ __forceinline__ void a_store(uint32_t lane_base, int col, const float* v) {
  uint32_t hi[N], lo[N];
#pragma unroll
  for (int i = 0; i < N; ++i) tf32_split(v[i], hi[i], lo[i]);
  if (N == 16) {
    tmem_st16(lane_base + col, hi);
    tmem_st16(lane_base + TC_ALO + col, lo);
  } else if (N == 8) {
    tmem_st8(lane_base + col, hi);
    tmem_st8(lane_base + TC_ALO + col, lo);
  } else {
    tmem_st2(lane_base + col, hi);
    tmem_st2(lane_base + TC_ALO + col, lo);
  }
}

__device__ __forceinline__ float sigmoid_fast(float x) {
  float t, r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(t) : "f"(x * -1.4426950408889634f));
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(1.0f + t));
  return r;
}

// The five GEMMs of the reference stack (the only one b2_lstm_tc_pack accepts): shapes, panel columns and weight offsets are
// compile-time constants, so the issuing thread's loop is fully unrolled and a descriptor is "base + constant".
template <int L> struct TcL;
template <> struct TcL<0> { static constexpr int N = 128, K = 56, A = 0,  W = 0; };
template <> struct TcL<1> { static constexpr int N = 64,  K = 56, A = 24, W = TcL<0>::W + 2 * TcL<0>::N * TcL<0>::K; };
template <> struct TcL<2> { static constexpr int N = 64,  K = 40, A = 0,  W = TcL<1>::W + 2 * TcL<1>::N * TcL<1>::K; };
template <> struct TcL<3> { static constexpr int N = 128, K = 56, A = 24, W = TcL<2>::W + 2 * TcL<2>::N * TcL<2>::K; };
template <> struct TcL<4> { static constexpr int N = 32,  K = 40, A = 40, W = TcL<3>::W + 2 * TcL<3>::N * TcL<3>::K; };

// one thread: D = A_hi·B_hi + A_lo·B_hi + A_hi·B_lo for layer L, then commit (pbase = TMEM address of the tile's panel, dacc = of
// the accumulator, wbase = shared-memory address of the packed weights)
template <int L>
__device__ __forceinline__ void issue_layer(uint32_t pbase, uint32_t dacc, uint32_t wbase, uint64_t* bar) {
  using T = TcL<L>;
  tc_fence_after();
  constexpr uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(T::N >> 3) << 17) | ((128u >> 4) << 24);
  constexpr int chunk = T::N * 16;  // bytes between consecutive 16-byte K-chunks (LBO); rows 8 apart are 128 B apart (SBO)
  const uint32_t b_hi = wbase + (uint32_t)T::W * 4u, b_lo = b_hi + (uint32_t)(T::N * T::K * 4);
  // descriptor of (b_hi, chunk 0); the start-address field counts 16-byte units, so later descriptors are this one plus a constant
  const uint64_t d_hi = tc_bdesc(b_hi, chunk, 128), d_lo = tc_bdesc(b_lo, chunk, 128);
#pragma unroll
  for (int pass = 0; pass < 3; ++pass) {
    const uint32_t a0 = pbase + T::A + (pass == 1 ? TC_ALO : 0);
#pragma unroll
    for (int j = 0; j < T::K / 8; ++j) {
      const uint64_t bd = (pass == 2 ? d_lo : d_hi) + (uint64_t)((j * 2 * chunk) >> 4);
      tc_mma(dacc, a0 + 8 * j, bd, idesc, (pass | j) != 0 ? 1u : 0u);
    }
  }
  tc_commit(bar);
}

// LSTM epilogue for this thread's UH units (half of the layer's; the other half belongs to the thread 128 above / below on the
// same TMEM lane): gate pre-activations of 8 units at a time, update c (registers), write h into the A panel.
// PULLED: the thread's 4·UH accumulator columns are already in z_all (the accumulator itself may belong to the other tile by now);
// otherwise they are read from TMEM chunk by chunk.
template <int UH, bool PULLED>
__device__ __forceinline__ void lstm_epilogue(uint32_t lane_panel, uint32_t lane_acc, int half, int h_col, float (&c)[UH], float* h_out /* UH or nullptr */,
                                              const uint32_t* z_all) {
#pragma unroll
  for (int ch = 0; ch < UH / 8; ++ch) {
    uint32_t zl[32];
    if (!PULLED) {
      tmem_ld32(lane_acc + (half * (UH / 8) + ch) * 32, zl);
      tc_wait_ld();
    }
    const uint32_t* z = PULLED ? z_all + ch * 32 : zl;
    float h[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      const float ig = sigmoid_fast(__uint_as_float(z[4 * q]));
      const float fg = sigmoid_fast(__uint_as_float(z[4 * q + 1]));
      const float gg = fmaxf(__uint_as_float(z[4 * q + 2]), 0.f);
      const float og = sigmoid_fast(__uint_as_float(z[4 * q + 3]));
      const int u = ch * 8 + q;
      c[u] = fmaf(fg, c[u], ig * gg);
      h[q] = og * fmaxf(c[u], 0.f);
      if (h_out) h_out[u] = h[q];
    }
    a_store<8>(lane_panel, h_col + half * UH + ch * 8, h);
  }
}

template <int NCOL>
__device__ __forceinline__ void pull_cols(uint32_t taddr, uint32_t (&z)[NCOL]) {  // NCOL (32 or 64) accumulator columns of this lane → registers
#pragma unroll
  for (int c0 = 0; c0 < NCOL; c0 += 32) {
    uint32_t t32[32];
    tmem_ld32(taddr + c0, t32);
#pragma unroll
    for (int i = 0; i < 32; ++i) z[c0 + i] = t32[i];
  }
  tc_wait_ld();
}

__device__ __forceinline__ void group_barrier(int g, int G) {  // all 256 threads of tile g (named barrier; the whole CTA when G == 1)
  if (G == 1) __syncthreads();
  else asm volatile("bar.sync %0, %1;" ::"r"(1 + g), "r"(TC_THREADS) : "memory");
}

template <int G>
__global__ void __launch_bounds__(TC_THREADS * G, 1) lstm_forward_tc_kernel(const __grid_constant__ LstmTcArgs a) {
  extern __shared__ __align__(128) float wsm[];
  __shared__ uint32_t s_tmem;
  __shared__ __align__(8) uint64_t bar[2];   // MMA-done, one per tile
  __shared__ __align__(8) uint64_t tok[2];   // G == 2: "the accumulator is yours", one per tile
  const int tid_cta = threadIdx.x;
  const int g = G == 1 ? 0 : tid_cta / TC_THREADS;   // tile (group) of this thread
  const int tid = tid_cta - g * TC_THREADS, warp = tid >> 5;
  const int F = a.F;
  const bool vec18 = F == 18 && ((reinterpret_cast<uintptr_t>(a.x) | reinterpret_cast<uintptr_t>(a.y)) & 7) == 0;  // 8-byte row accesses
  if (tid_cta < 32) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s_tmem)), "r"(TC_TMEM_COLS));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  if (tid_cta == 0) {
    mbar_init(&bar[0], 1);
    mbar_init(&bar[1], 1);
    mbar_init(&tok[0], 1);
    mbar_init(&tok[1], 1);
    mbar_fence_init();
  }
  for (int i = tid_cta; i < a.packed_floats; i += TC_THREADS * G) wsm[i] = a.packed[i];
  fence_proxy_async();  // generic-proxy writes of the weights → visible to the tensor core's async-proxy reads
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tbase = s_tmem;
  const int half = tid >> 7, row = tid & (TC_NT - 1);  // warps w and w + 4 of a group address the same 32 TMEM lanes
  const uint32_t lane_off = (uint32_t)((warp & 3) * 32) << 16;
  const uint32_t pbase = tbase + (uint32_t)(g * TC_PANEL);     // this tile's A panel
  const uint32_t lane_base = pbase + lane_off;
  const uint32_t dacc = tbase + TC_D, lane_acc = dacc + lane_off;
  uint32_t parity = 0, tok_parity = 0;
  // Token protocol (G == 2): tile 0 owns the accumulator first.  Before issuing a layer the tile's issuing thread waits for its
  // token; after the tile has pulled its accumulator into registers one thread passes the token on.  Both tiles run the same
  // number of layers (a tile beyond the end of the batch computes on zeros), so the hand-over never stalls for good.
  if (G == 2 && tid_cta == 0) mbar_arrive(&tok[0]);

  // all threads of the tile: previous A-panel writes are complete → issue → wait for the accumulator
  const uint32_t wbase = smem_u32(wsm);
  auto run_layer = [&](auto layer_tag) {
    constexpr int L = decltype(layer_tag)::value;
    tc_wait_st();
    tc_fence_before();
    group_barrier(g, G);
    if (tid == 0) {
      if (G == 2) {
        mbar_wait(&tok[g], tok_parity);
        tc_fence_after();
      }
      issue_layer<L>(pbase, dacc, wbase, &bar[g]);
    }
    if (G == 2) tok_parity ^= 1u;
    mbar_wait(&bar[g], parity);
    parity ^= 1u;
    tc_fence_after();
  };
  // G == 2: once every thread of the tile has its accumulator columns in registers, the accumulator goes to the other tile
  auto release_acc = [&]() {
    tc_fence_before();
    group_barrier(g, G);
    if (tid == 0) mbar_arrive(&tok[1 - g]);
  };

  const long long n_tiles = (a.batch + TC_NT - 1) / TC_NT;
  const long long n_rounds = (n_tiles + (long long)gridDim.x * G - 1) / ((long long)gridDim.x * G);
  for (long long r = 0; r < n_rounds; ++r) {
    const long long tile = (r * gridDim.x + blockIdx.x) * G + g;
    const long long seq = tile * TC_NT + row;
    const bool live = tile < n_tiles && seq < a.batch;
    float c1[16], c2[8], c3[8], c4[16], h2[8];  // this thread's half of every layer's units
#pragma unroll
    for (int i = 0; i < 16; ++i) c1[i] = c4[i] = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) c2[i] = c3[i] = h2[i] = 0.f;
    // ---- phase 1 panel: [1, 0×5, x(18), h1(32), h2(16), 1, 0×7] ----
    if (half == 0) {
      float z16[16];
#pragma unroll
      for (int i = 0; i < 16; ++i) z16[i] = 0.f;
      float first[16];
#pragma unroll
      for (int i = 0; i < 16; ++i) first[i] = 0.f;
      first[0] = 1.0f;
      a_store<16>(lane_base, 0, first);   // cols 0..15 (bias-1 at 0; x fills 6.. later)
      a_store<8>(lane_base, 24, z16);     // 24..31 (cols 16..23 belong to the other half of the thread pair: it writes x there every
                                          // step, and nothing orders its store against one made here)
      a_store<16>(lane_base, 32, z16);    // 32..47
      a_store<16>(lane_base, 48, z16);    // 48..63
      float last[16];
#pragma unroll
      for (int i = 0; i < 16; ++i) last[i] = 0.f;
      last[8] = 1.0f;                     // col 72
      a_store<16>(lane_base, 64, last);   // 64..79
    }
    for (int t = 0; t < a.T; ++t) {
      // cols 0..23 of the panel: [1, 0, 0, 0, 0, 0, x0..x17]; half 0 writes cols 0..15, half 1 cols 16..23.  The reference's width
      // (F = 18, rows 72 bytes apart: aligned 8-byte accesses) has its own branch with its own arrays so that those stay in
      // registers; the generic branch indexes at run time and lives in local memory.
      const float* xp = a.x + (seq * a.seq_stride + t) * F;
      if (vec18) {
        if (half == 0) {
          float xv[16];
#pragma unroll
          for (int i = 0; i < 16; ++i) xv[i] = 0.f;
          xv[0] = 1.0f;
          if (live) {
#pragma unroll
            for (int f = 0; f < 10; f += 2) {
              const float2 v = __ldg(reinterpret_cast<const float2*>(xp + f));
              xv[6 + f] = v.x;
              xv[7 + f] = v.y;
            }
          }
          a_store<16>(lane_base, 0, xv);
        } else {
          float xv[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) xv[i] = 0.f;
          if (live) {
#pragma unroll
            for (int f = 0; f < 8; f += 2) {
              const float2 v = __ldg(reinterpret_cast<const float2*>(xp + 10 + f));
              xv[f] = v.x;
              xv[f + 1] = v.y;
            }
          }
          a_store<8>(lane_base, 16, xv);
        }
      } else if (half == 0) {
        float xg[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) xg[i] = 0.f;
        xg[0] = 1.0f;
        if (live)
          for (int f = 0; f < min(F, 10); ++f) xg[6 + f] = __ldg(xp + f);
        a_store<16>(lane_base, 0, xg);
      } else {
        float xg[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) xg[i] = 0.f;
        if (live)
          for (int f = 10; f < F; ++f) xg[f - 10] = __ldg(xp + f);
        a_store<8>(lane_base, 16, xg);
      }
      run_layer(std::integral_constant<int, 0>{});                                        // L1: cols [0,56) → 128 gate columns
      if (G == 2) {
        uint32_t z[64];
        pull_cols<64>(lane_acc + half * 64, z);
        release_acc();
        lstm_epilogue<16, true>(lane_base, lane_acc, half, 24, c1, nullptr, z);
      } else {
        lstm_epilogue<16, false>(lane_base, lane_acc, half, 24, c1, nullptr, nullptr);  // h1 → cols 24..55
      }
      run_layer(std::integral_constant<int, 1>{});                                        // L2: cols [24,80) → 64 gate columns
      if (G == 2) {
        uint32_t z[32];
        pull_cols<32>(lane_acc + half * 32, z);
        release_acc();
        lstm_epilogue<8, true>(lane_base, lane_acc, half, 56, c2, h2, z);
      } else {
        lstm_epilogue<8, false>(lane_base, lane_acc, half, 56, c2, h2, nullptr);        // h2 → cols 56..71 (kept in registers for phase 2)
      }
    }
    // ---- phase 2 panel: [1, 0×7, h2_final(16), h3(16), h4(32), 1, 0×7] ----
    {
      a_store<8>(lane_base, 8 + half * 8, h2);  // RepeatVector: the same h2_T feeds every decoder step (each half its 8 units)
      float z16[16];
#pragma unroll
      for (int i = 0; i < 16; ++i) z16[i] = 0.f;
      if (half == 0) {
        float head[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) head[i] = 0.f;
        head[0] = 1.0f;
        a_store<8>(lane_base, 0, head);
        a_store<16>(lane_base, 24, z16);    // h3 = 0
      } else {
        a_store<16>(lane_base, 40, z16);    // h4 = 0
        a_store<16>(lane_base, 56, z16);
      }
      // cols 72..79 still hold [1, 0×7] from phase 1
    }
    for (int t = 0; t < a.T; ++t) {
      run_layer(std::integral_constant<int, 2>{});                                        // L3: cols [0,40)
      if (G == 2) {
        uint32_t z[32];
        pull_cols<32>(lane_acc + half * 32, z);
        release_acc();
        lstm_epilogue<8, true>(lane_base, lane_acc, half, 24, c3, nullptr, z);
      } else {
        lstm_epilogue<8, false>(lane_base, lane_acc, half, 24, c3, nullptr, nullptr);   // h3 → cols 24..39
      }
      run_layer(std::integral_constant<int, 3>{});                                        // L4: cols [24,80)
      if (G == 2) {
        uint32_t z[64];
        pull_cols<64>(lane_acc + half * 64, z);
        release_acc();
        lstm_epilogue<16, true>(lane_base, lane_acc, half, 40, c4, nullptr, z);
      } else {
        lstm_epilogue<16, false>(lane_base, lane_acc, half, 40, c4, nullptr, nullptr);  // h4 → cols 40..71
      }
      run_layer(std::integral_constant<int, 4>{});                                        // TimeDistributed(Dense): cols [40,80) → 32 columns
      {
        uint32_t z[32];
        if (G == 2) {
          pull_cols<32>(lane_acc, z);  // both halves read the row (only half 0 writes it): the hand-over needs every thread's arrival
          release_acc();
        } else if (half == 0) {
          tmem_ld32(lane_acc, z);
          tc_wait_ld();
        }
        if (half == 0 && live) {  // the 18 output columns are one 32-column read: one thread per sequence writes the row
          float* yp = a.y + (seq * a.T + t) * F;
          if (vec18) {
#pragma unroll
            for (int f = 0; f < 18; f += 2)
              *reinterpret_cast<float2*>(yp + f) = make_float2(__uint_as_float(z[f]), __uint_as_float(z[f + 1]));
          } else {
#pragma unroll
            for (int f = 0; f < 32; ++f)   // static register indices, predicated stores
              if (f < F) yp[f] = __uint_as_float(z[f]);
          }
        }
      }
    }
    if (G == 1) {
      tc_fence_before();
      __syncthreads();  // accumulator reads done before the next tile's first MMA overwrites it
    }
  }
  tc_fence_before();
  __syncthreads();
  if (tid_cta < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tbase), "r"(TC_TMEM_COLS));
}

// ---- host side ----------------------------------------------------------------------------------------------------
// B[n][k] → canonical K-major no-swizzle tile: the 16-byte chunk (4 k's) of row n sits at (k/4)*(N*16) + n*16 bytes
static void pack_canonical(const std::vector<float>& B, int N, int K, std::vector<float>& out) {
  std::vector<float> hi((size_t)N * K), lo((size_t)N * K);
  for (int n = 0; n < N; ++n)
    for (int k = 0; k < K; ++k) {
      const float w = B[(size_t)n * K + k];
      uint32_t bits;
      memcpy(&bits, &w, 4);
      // round-to-nearest-even on the 13 dropped bits (the device uses cvt.rna; either is fine: lo carries the remainder)
      uint32_t hb = (bits + 0x1000u) & 0xFFFFE000u;
      float h;
      memcpy(&h, &hb, 4);
      if (!(fabsf(h) <= 3.0e38f)) h = w;  // do not round into inf
      const size_t off = (size_t)(k >> 2) * (N * 4) + (size_t)n * 4 + (k & 3);
      hi[off] = h;
      lo[off] = w - h;
    }
  out.insert(out.end(), hi.begin(), hi.end());
  out.insert(out.end(), lo.begin(), lo.end());
}

// Keras weights → the five B matrices ([N][K], gate-interleaved columns n = unit*4 + gate, k ordered like the A panels)
int b2_lstm_tc_pack(const b2_lstm_desc* desc, const float* w, std::vector<float>& packed, TcLayer (&layers)[5]) {
  const int F = desc->features;
  if (F > 18 || !(desc->units[0] == 32 && desc->units[1] == 16 && desc->units[2] == 16 && desc->units[3] == 32)) return -1;
  const int in_dim[4] = {F, 32, 16, 16}, u[4] = {32, 16, 16, 32};
  const int K[5] = {56, 56, 40, 56, 40}, N[5] = {128, 64, 64, 128, 32}, a_col[5] = {0, 24, 0, 24, 40};
  const float* p = w;
  packed.clear();
  for (int l = 0; l < 4; ++l) {
    const int Nk = 4 * u[l];
    const float* W = p;
    const float* U = W + (size_t)in_dim[l] * Nk;
    const float* b = U + (size_t)u[l] * Nk;
    p = b + Nk;
    std::vector<float> B((size_t)N[l] * K[l], 0.f);
    auto col = [&](int unit, int gate) { return gate * u[l] + unit; };  // Keras column of (unit, gate)
    for (int unit = 0; unit < u[l]; ++unit)
      for (int gate = 0; gate < 4; ++gate) {
        const int n = unit * 4 + gate;
        float* row = &B[(size_t)n * K[l]];
        int k_in, k_h, k_b;  // panel positions of this layer's input, recurrent input and bias-1 column
        if (l == 0) k_b = 0, k_in = 6, k_h = 24;          // [1, 0×5, x, h1]
        else if (l == 1) k_in = 0, k_h = 32, k_b = 48;    // [h1, h2, 1]
        else if (l == 2) k_b = 0, k_in = 8, k_h = 24;     // [1, 0×7, h2f, h3]
        else k_in = 0, k_h = 16, k_b = 48;                // [h3, h4, 1]
        for (int i = 0; i < in_dim[l]; ++i) row[k_in + i] = W[(size_t)i * Nk + col(unit, gate)];
        for (int j = 0; j < u[l]; ++j) row[k_h + j] = U[(size_t)j * Nk + col(unit, gate)];
        row[k_b] = b[col(unit, gate)];
      }
    layers[l] = TcLayer{N[l], K[l], a_col[l], (int)packed.size()};
    pack_canonical(B, N[l], K[l], packed);
  }
  {  // Dense: panel [h4(32), 1, 0×7]
    std::vector<float> B((size_t)N[4] * K[4], 0.f);
    for (int n = 0; n < F; ++n) {
      for (int k = 0; k < 32; ++k) B[(size_t)n * K[4] + k] = p[(size_t)k * F + n];
      B[(size_t)n * K[4] + 32] = p[(size_t)32 * F + n];
    }
    layers[4] = TcLayer{N[4], K[4], a_col[4], (int)packed.size()};
    pack_canonical(B, N[4], K[4], packed);
  }
  return 0;
}

int b2_lstm_tc_launch(b2_ctx* ctx, const float* d_packed, int packed_floats, const TcLayer (&layers)[5], int F, const float* x, int64_t batch,
                      int T, int64_t seq_stride, float* y, cudaStream_t stream) {
  LstmTcArgs a;
  a.x = x;
  a.y = y;
  a.batch = batch;
  a.seq_stride = seq_stride;
  a.T = T;
  a.F = F;
  a.packed = d_packed;
  a.packed_floats = packed_floats;
  for (int l = 0; l < 5; ++l) a.layer[l] = layers[l];
  const size_t smem = sizeof(float) * (size_t)packed_floats + 128;
  B2_REQUIRE(smem <= (size_t)ctx->smem_optin, "lstm_tc: weights need %zu bytes of shared memory", smem);
  const long long tiles = (batch + TC_NT - 1) / TC_NT;
  // context option lstm_tc: 2 (default) two tiles per CTA once there is more than one tile per SM, 3 always two, 1 always one
  const bool two = ctx->knobs.lstm_tc == 3 || (ctx->knobs.lstm_tc == 2 && tiles > ctx->sm_count);
  if (two) {
    B2_CUDA(cudaFuncSetAttribute(lstm_forward_tc_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    long long grid = ctx->sm_count;
    if (grid > (tiles + 1) / 2) grid = (tiles + 1) / 2;
    lstm_forward_tc_kernel<2><<<(unsigned)grid, TC_THREADS * 2, smem, stream>>>(a);
  } else {
    B2_CUDA(cudaFuncSetAttribute(lstm_forward_tc_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    long long grid = ctx->sm_count;
    if (grid > tiles) grid = tiles;
    lstm_forward_tc_kernel<1><<<(unsigned)grid, TC_THREADS, smem, stream>>>(a);
  }
  B2_CUDA(cudaGetLastError());
  return B2_OK;
}
