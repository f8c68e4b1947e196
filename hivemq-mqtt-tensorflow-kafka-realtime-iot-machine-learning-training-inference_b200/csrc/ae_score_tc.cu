// K1t — fused normalise → Dense stack (tcgen05) → per-record MSE → flag for pre-decoded rows; see ae_tc.cuh for the tensor-core
// formulation and ae_score.cu for the FP32 kernel it stands beside (same inputs, outputs and HBM layout).
//
// Persistent CTAs of 128 threads (= 128 rows = the TMEM lanes), four per SM; a 128-row tile of x arrives by one bulk async copy
// into a two-deep shared-memory ring; each thread normalises its row in registers, the tile's four layer products run on the
// tensor cores, and the thread folds its row of the last accumulator into the squared error.
// Algorithmic bytes/record: 4*D read + 5 written.
#include "ae_tc.cuh"

template <int D>
struct ScoreTcArgs {
  AeW<D> w;             // FP32 copy of the weights: the two middle layers run in registers (MID_FP32)
  NormP<D> nm;
  const float* panels;  // device: TcShape<D>::TOTAL floats
  const float* x;
  float* mse;
  uint8_t* flag;
  unsigned long long* counters;
  long long n;
  float threshold;
  int use_bulk;
};

template <int D, bool NORM, bool FT, bool MID>
__global__ void __launch_bounds__(TCS_ROWS, 512 / TcShape<D>::TMEM_COLS > 6 ? 6 : 512 / TcShape<D>::TMEM_COLS) ae_score_tc_kernel(const __grid_constant__ ScoreTcArgs<D> a) {
  constexpr int STAGES = 2;
  constexpr uint32_t TILE_BYTES = TCS_ROWS * D * 4;
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ uint32_t s_tmem;
  __shared__ __align__(8) uint64_t mma_bar;
  __shared__ __align__(8) uint64_t full[STAGES];
  float* wsm = reinterpret_cast<float*>(smem + (size_t)STAGES * TILE_BYTES);
  const int tid = threadIdx.x;
  const long long n_tiles = (a.n + TCS_ROWS - 1) / TCS_ROWS;

  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < STAGES; ++s) mbar_init(&full[s], 1);
  }
  TcCtx c;
  tcs_setup<D>(c, a.panels, wsm, &s_tmem, &mma_bar, tid);  // includes the mbarrier-init fence and a __syncthreads

  auto issue = [&](long long tile, int s) {  // thread 0 only
    if (tile < n_tiles && a.use_bulk && (tile + 1) * TCS_ROWS <= a.n) {
      mbar_expect_tx(&full[s], TILE_BYTES);
      bulk_g2s(smem + (size_t)s * TILE_BYTES, a.x + tile * (TCS_ROWS * D), TILE_BYTES, &full[s]);
    }
  };
  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < STAGES; ++s) issue((long long)blockIdx.x + (long long)s * gridDim.x, s);
  }

  uint32_t phase = 0;
  unsigned flagged = 0;
  int s = 0;
  for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    float* st = reinterpret_cast<float*>(smem + (size_t)s * TILE_BYTES);
    const long long row0 = tile * TCS_ROWS;
    const int rows = (int)min((long long)TCS_ROWS, a.n - row0);
    const bool bulk = a.use_bulk && rows == TCS_ROWS;
    if (bulk) {
      mbar_wait(&full[s], (phase >> s) & 1u);
      phase ^= 1u << s;
    } else {  // ragged last tile or unaligned input: plain coalesced loads
      const float* src = a.x + row0 * D;
      for (int i = tid; i < rows * D; i += TCS_ROWS) st[i] = __ldg(src + i);
      __syncthreads();
    }
    float x[D];
    const bool live = tid < rows;
#pragma unroll
    for (int i = 0; i < D; i += 2) {
      float2 v = live ? *reinterpret_cast<const float2*>(st + (size_t)tid * D + i) : make_float2(0.f, 0.f);
      if (NORM) v = norm_pair_fast<D>(v, a.nm, i);
      x[i] = v.x;
      x[i + 1] = v.y;
    }
    // the stage is free once every thread has its row in registers: the first barrier inside tcs_score_row orders that, and
    // thread 0 refills it right after issuing layer 1 … simpler and just as good: refill after the row's score
    const float m = tcs_score_row<D, FT, MID>(c, a.w, x, tid);
    if (tid == 0) issue(tile + (long long)STAGES * gridDim.x, s);
    const bool f = m > a.threshold;  // strict, nb2.0:1163
    if (live) {
      if (a.mse) a.mse[row0 + tid] = m;
      if (a.flag) a.flag[row0 + tid] = f ? 1 : 0;
      flagged += f;
    }
    s = (s + 1 == STAGES) ? 0 : s + 1;
  }
  tcs_teardown<D>(c, tid);

  if (a.counters) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) flagged += __shfl_xor_sync(0xffffffffu, flagged, o);
    if ((tid & 31) == 0 && flagged) atomicAdd(&a.counters[1], (unsigned long long)flagged);
    if (blockIdx.x == 0 && tid == 0) atomicAdd(&a.counters[0], (unsigned long long)a.n);
  }
}

// ---- host side ----------------------------------------------------------------------------------------------
int b2_ae_tc_panels(b2_ae* ae, const float** out) {
  B2_REQUIRE(ae->spec_d == 18 || ae->spec_d == 30, "tensor-core score path needs the D-14-7-7-D autoencoder with D in {18, 30}");
  if (!ae->tc_valid) {
    int rc = b2_ae_sync_host(ae);
    if (rc) return rc;
    const size_t bytes = sizeof(float) * (ae->spec_d == 18 ? TcShape<18>::TOTAL : TcShape<30>::TOTAL);
    if (!ae->d_tc) B2_CUDA(cudaMalloc(&ae->d_tc, sizeof(float) * TcShape<30>::TOTAL));
    if (ae->spec_d == 18) {
      TcPanels<18> h;
      tc_pack_weights<18>(ae->h_w.data(), h);
      B2_CUDA(cudaMemcpy(ae->d_tc, h.p, bytes, cudaMemcpyHostToDevice));
    } else {
      TcPanels<30> h;
      tc_pack_weights<30>(ae->h_w.data(), h);
      B2_CUDA(cudaMemcpy(ae->d_tc, h.p, bytes, cudaMemcpyHostToDevice));
    }
    ae->tc_valid = true;
  }
  *out = ae->d_tc;
  return B2_OK;
}

template <int D, bool NORM, bool FT, bool MID>
static int launch_score_tc(b2_ae* ae, const ScoreTcArgs<D>& args, cudaStream_t stream) {
  constexpr size_t smem = (size_t)2 * TCS_ROWS * D * 4 + sizeof(float) * TcShape<D>::TOTAL;
  auto kern = ae_score_tc_kernel<D, NORM, FT, MID>;
  static bool attr_set[16] = {false};
  const int dev = ae->ctx->device;
  if (!attr_set[dev & 15]) {
    B2_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    B2_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    attr_set[dev & 15] = true;
  }
  // CTAs per SM by construction: 512 TMEM columns / this kernel's allocation (4 for D = 30, up to 8 for D = 18; registers
  // allow 6 there), ~40 / ~28 KB of shared memory each; the context option "tc_ctas" overrides it for experiments.
  // (cudaOccupancyMaxActiveBlocksPerMultiprocessor answers 1 for this kernel; tile striding is static, so a CTA that had to
  // wait for a slot would only cost balance, never correctness.)
  const int want = ae->ctx->knobs.tc_ctas > 0 ? ae->ctx->knobs.tc_ctas : 512 / TcShape<D>::TMEM_COLS;
  const int CTAS_PER_SM = want < 1 ? 1 : (want > 512 / TcShape<D>::TMEM_COLS ? 512 / TcShape<D>::TMEM_COLS : want);
  const long long n_tiles = (args.n + TCS_ROWS - 1) / TCS_ROWS;
  long long grid = (long long)ae->ctx->sm_count * CTAS_PER_SM;
  if (grid > n_tiles) grid = n_tiles;
  if (grid < 1) grid = 1;
  kern<<<(unsigned)grid, TCS_ROWS, smem, stream>>>(args);
  B2_CUDA(cudaGetLastError());
  return B2_OK;
}

template <int D>
static int score_tc_spec(b2_ae* ae, const float* x, int64_t n, const b2_norm* norm, float threshold, float* mse, uint8_t* flag,
                         unsigned long long* counters, cudaStream_t stream) {
  ScoreTcArgs<D> args;
  const bool has_norm = norm_fill<D>(norm, args.nm) != 0;
  int rc = b2_ae_tc_panels(ae, &args.panels);
  if (rc) return rc;
  args.x = x;
  args.mse = mse;
  args.flag = flag;
  args.counters = counters;
  args.n = n;
  args.threshold = threshold;
  args.use_bulk = (((uintptr_t)x & 15) == 0) && !ae->ctx->knobs.no_bulk;
  ae_pack_weights<D>(ae->h_w.data(), args.w);
  const bool ft = ae->ctx->knobs.fast_tanh != 0;
  const bool mid = ae->ctx->knobs.score_tc != 2;  // score_tc = 2: all four layers on the tensor cores (four round trips per tile)
  if (!ft) {  // the ≤ 4-ulp tanh is a debugging aid: one configuration is enough
    if (has_norm) return launch_score_tc<D, true, false, true>(ae, args, stream);
    return launch_score_tc<D, false, false, true>(ae, args, stream);
  }
  if (has_norm) {
    if (mid) return launch_score_tc<D, true, true, true>(ae, args, stream);
    return launch_score_tc<D, true, true, false>(ae, args, stream);
  }
  if (mid) return launch_score_tc<D, false, true, true>(ae, args, stream);
  return launch_score_tc<D, false, true, false>(ae, args, stream);
}

// called by b2_ae_score when the context option "score_tc" is set and no reconstruction rows are requested
int b2_ae_score_tc(b2_ae* ae, const float* x_dev, int64_t n, const b2_norm* norm, float threshold, float* mse_dev, uint8_t* flag_dev,
                   unsigned long long* counters_dev, cudaStream_t stream) {
  if (ae->spec_d == 30) return score_tc_spec<30>(ae, x_dev, n, norm, threshold, mse_dev, flag_dev, counters_dev, stream);
  return score_tc_spec<18>(ae, x_dev, n, norm, threshold, mse_dev, flag_dev, counters_dev, stream);
}
