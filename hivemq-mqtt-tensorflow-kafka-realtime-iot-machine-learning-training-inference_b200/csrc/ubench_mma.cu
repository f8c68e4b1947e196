// Micro-benchmark (standalone binary, not part of libb2sensor.so): the warp-level MMA form of the Dense stack (ae_mma.cuh).
// Answers, before the fused decode kernel is touched: (1) what mma.sync.m16n8k8 TF32 (SASS HMMA.1688.F32.TF32) sustains per SM on
// sm_100a, (2) the error of the 3-pass split product against a float64 evaluation of the same rows, (3) rows/s of the score phase
// alone on L2-resident and on HBM-resident rows.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 --expt-relaxed-constexpr -I../../include -o ubench_mma ubench_mma.cu
#include <cuda_runtime.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <vector>

#include "ae_mma.cuh"

#define CK(x)                                                                      \
  do {                                                                             \
    cudaError_t e_ = (x);                                                          \
    if (e_ != cudaSuccess) {                                                       \
      fprintf(stderr, "%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e_)); \
      exit(1);                                                                     \
    }                                                                              \
  } while (0)

// (1) raw HMMA rate: every warp runs CH independent accumulator chains
template <int CH>
__global__ void __launch_bounds__(256) hmma_rate(float* out, int iters, long long* cyc) {
  float acc[CH][4];
  float a[4], b0 = threadIdx.x * 1e-3f, b1 = 0.5f;
#pragma unroll
  for (int i = 0; i < 4; ++i) a[i] = 1.0f + i + threadIdx.x;
#pragma unroll
  for (int c = 0; c < CH; ++c)
#pragma unroll
    for (int i = 0; i < 4; ++i) acc[c][i] = c + i;
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int c = 0; c < CH; ++c) mma_tf32(acc[c], a, b0, b1, acc[c]);
  }
  const long long t1 = clock64();
  float s = 0.f;
#pragma unroll
  for (int c = 0; c < CH; ++c) s += acc[c][0] + acc[c][1] + acc[c][2] + acc[c][3];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

// (1b) does an HMMA hold the issue port?  Per iteration CH HMMAs on independent accumulators plus NF independent scalar FFMAs:
// if the two streams overlap, the time is the maximum of the two alone, otherwise their sum
template <int CH, int NF>
__global__ void __launch_bounds__(256) hmma_mix(float* out, int iters, long long* cyc) {
  float acc[CH > 0 ? CH : 1][4], f[NF > 0 ? NF : 1];
  float a[4], b0 = threadIdx.x * 1e-3f, b1 = 0.5f;
#pragma unroll
  for (int i = 0; i < 4; ++i) a[i] = 1.0f + i + threadIdx.x;
#pragma unroll
  for (int c = 0; c < (CH > 0 ? CH : 1); ++c)
#pragma unroll
    for (int i = 0; i < 4; ++i) acc[c][i] = c + i;
#pragma unroll
  for (int i = 0; i < (NF > 0 ? NF : 1); ++i) f[i] = i * 0.25f + threadIdx.x;
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int c = 0; c < CH; ++c) {
      mma_tf32(acc[c], a, b0, b1, acc[c]);
#pragma unroll
      for (int i = c * NF / (CH > 0 ? CH : 1); i < (c + 1) * NF / (CH > 0 ? CH : 1); ++i) f[i] = fmaf(f[i], b1, b0);
    }
    if (CH == 0) {
#pragma unroll
      for (int i = 0; i < NF; ++i) f[i] = fmaf(f[i], b1, b0);
    }
  }
  const long long t1 = clock64();
  float s = 0.f;
#pragma unroll
  for (int c = 0; c < (CH > 0 ? CH : 1); ++c) s += acc[c][0] + acc[c][1] + acc[c][2] + acc[c][3];
#pragma unroll
  for (int i = 0; i < (NF > 0 ? NF : 1); ++i) s += f[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

// (2), (3) the score phase on pre-decoded rows: every warp walks chunks of 32 rows, stages them into its interleaved tile, scores
template <int D, bool FT>
__global__ void __launch_bounds__(256) score_rows(const __grid_constant__ AeW<D> w, const float* __restrict__ x, float* __restrict__ out, long long n,
                                                  int passes) {
  extern __shared__ __align__(128) uint8_t smem[];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, wpc = blockDim.x >> 5;
  uint8_t* tab = smem;
  float* xt = reinterpret_cast<float*>(smem + MmaShape<D>::BYTES) + warp * 32 * D;
  mma_fill_table<D>(w, tab, tid, blockDim.x);
  __syncthreads();
  for (int p = 0; p < passes; ++p)
    for (long long c = (long long)blockIdx.x * wpc + warp; c * 32 < n; c += (long long)gridDim.x * wpc) {
      const float2* src = reinterpret_cast<const float2*>(x + c * 32 * D);
      for (int i = lane; i < 16 * D; i += 32) {  // pairs of elements of one row
        const int r = i / (D / 2), j = 2 * (i % (D / 2));
        const float2 v = src[i];
        float* dst = xt + mma_row_offset<D>(r) + 2 * j;
        dst[0] = v.x;
        dst[2] = v.y;
      }
      __syncwarp();
      const float m = mma_score_warp<D, FT>(xt, tab, lane);
      out[c * 32 + lane] = m;
      __syncwarp();
    }
}

static double ref_row(const float* w, int D, const float* x) {
  const int widths[5] = {D, 14, 7, 7, D};
  double h[32], o[32];
  for (int i = 0; i < D; ++i) h[i] = x[i];
  const float* p = w;
  for (int l = 0; l < 4; ++l) {
    const int in = widths[l], out = widths[l + 1];
    const float* b = p + in * out;
    for (int j = 0; j < out; ++j) {
      double s = b[j];
      for (int i = 0; i < in; ++i) s += h[i] * (double)p[i * out + j];
      o[j] = (l % 2 == 0) ? tanh(s) : (s > 0 ? s : 0);
    }
    for (int j = 0; j < out; ++j) h[j] = o[j];
    p = b + out;
  }
  double se = 0;
  for (int i = 0; i < D; ++i) se += (x[i] - h[i]) * (x[i] - h[i]);
  return se / D;
}

template <int D>
static void run_score(int sms) {
  const int nw = D * 14 + 14 + 14 * 7 + 7 + 7 * 7 + 7 + 7 * D + D;
  std::vector<float> w(nw);
  srand(314);
  auto uni = [](float lim) { return lim * (2.0f * rand() / (float)RAND_MAX - 1.0f); };
  {
    const int widths[5] = {D, 14, 7, 7, D};
    float* p = w.data();
    for (int l = 0; l < 4; ++l) {
      const int in = widths[l], out = widths[l + 1];
      const float lim = sqrtf(6.0f / (in + out));
      for (int i = 0; i < in * out; ++i) *p++ = uni(lim);
      for (int j = 0; j < out; ++j) *p++ = uni(0.1f);
    }
  }
  static AeW<D> packed;
  ae_pack_weights<D>(w.data(), packed);
  const long long n_big = 1ll << 24, n_small = 1ll << 16;
  std::vector<float> hx((size_t)n_small * D);
  for (auto& v : hx) v = uni(1.0f);
  for (int r = 0; r < 64; ++r) hx[(size_t)r * D + (r % D)] = uni(8.0f);  // a few outliers
  float *dx, *dout;
  CK(cudaMalloc(&dx, (size_t)n_big * D * 4));
  CK(cudaMalloc(&dout, (size_t)n_big * 4));
  for (long long off = 0; off < n_big; off += n_small) CK(cudaMemcpy(dx + off * D, hx.data(), (size_t)n_small * D * 4, cudaMemcpyHostToDevice));
  const size_t smem = MmaShape<D>::BYTES + 8 * 32 * D * 4;
  auto kern = score_rows<D, true>;
  CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int occ = 0;
  CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, 256, smem));
  cudaFuncAttributes fa;
  CK(cudaFuncGetAttributes(&fa, kern));
  printf("D=%d: %d registers, %d CTAs of 8 warps per SM, %zu B shared per CTA\n", D, fa.numRegs, occ, smem);
  // error
  kern<<<sms * occ, 256, smem>>>(packed, dx, dout, n_small, 1);
  CK(cudaDeviceSynchronize());
  std::vector<float> ho(n_small);
  CK(cudaMemcpy(ho.data(), dout, n_small * 4, cudaMemcpyDeviceToHost));
  double max_rel = 0, sum_rel = 0;
  for (long long r = 0; r < n_small; ++r) {
    const double ref = ref_row(w.data(), D, &hx[(size_t)r * D]);
    const double rel = fabs(ho[r] - ref) / fmax(ref, 1e-2);
    max_rel = fmax(max_rel, rel);
    sum_rel += rel;
  }
  printf("D=%d: max relative error of the score vs float64 on %lld rows: %.3g (mean %.3g; budget 1e-5)\n", D, n_small, max_rel, sum_rel / n_small);
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  for (int which = 0; which < 2; ++which) {
    const long long n = which ? n_big : n_small;
    const int passes = which ? 1 : 256;
    for (int o = occ; o >= 1; o >>= 1) {
      kern<<<sms * o, 256, smem>>>(packed, dx, dout, n, passes);
      CK(cudaEventRecord(e0));
      for (int it = 0; it < 5; ++it) kern<<<sms * o, 256, smem>>>(packed, dx, dout, n, passes);
      CK(cudaEventRecord(e1));
      CK(cudaEventSynchronize(e1));
      float ms;
      CK(cudaEventElapsedTime(&ms, e0, e1));
      const double rate = 5.0 * n * passes / (ms * 1e-3);
      printf("D=%d %s rows, %d CTAs/SM: %.4g rows/s (%.3g B/row -> %.0f GB/s)\n", D, which ? "HBM-resident 2^24" : "L2-resident 2^16 x 256 passes", o, rate,
             4.0 * D + 4, rate * (4.0 * D + 4) / 1e9);
    }
  }
  CK(cudaFree(dx));
  CK(cudaFree(dout));
}

int main() {
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, 0));
  const int sms = prop.multiProcessorCount;
  printf("%s, %d SMs\n", prop.name, sms);
  float* out;
  long long* cyc;
  CK(cudaMalloc(&out, sizeof(float) * sms * 8 * 1024));
  CK(cudaMalloc(&cyc, sizeof(long long) * sms * 8));
  const int iters = 4096;
  auto report = [&](const char* name, int ch, int threads, int ctas) {
    std::vector<long long> h(sms * ctas);
    CK(cudaMemcpy(h.data(), cyc, sizeof(long long) * sms * ctas, cudaMemcpyDeviceToHost));
    double mean = 0;
    for (auto v : h) mean += (double)v;
    mean /= h.size();
    const double mmas_per_sm = (double)iters * ch * (threads / 32) * ctas;
    printf("%s: %d warps/SM x %d chains: %.2f cycles per HMMA.1688.TF32 per SM (= %.0f TF32 MAC/clk/SM)\n", name, threads / 32 * ctas, ch,
           mean / mmas_per_sm, 1024.0 * mmas_per_sm / mean);
  };
  for (int ctas = 1; ctas <= 4; ctas *= 2) {
    hmma_rate<4><<<sms * ctas, 256>>>(out, iters, cyc);
    CK(cudaDeviceSynchronize());
    hmma_rate<4><<<sms * ctas, 256>>>(out, iters, cyc);
    CK(cudaDeviceSynchronize());
    report("hmma_rate<4>", 4, 256, ctas);
  }
  hmma_rate<1><<<sms, 32>>>(out, iters, cyc);
  CK(cudaDeviceSynchronize());
  report("hmma_rate<1> (latency, one warp per SM)", 1, 32, 1);
  hmma_rate<8><<<sms, 128>>>(out, iters, cyc);
  CK(cudaDeviceSynchronize());
  report("hmma_rate<8> (one warp per scheduler)", 8, 128, 1);
  auto mix_report = [&](const char* name, int ctas) {
    std::vector<long long> h(sms * ctas);
    CK(cudaMemcpy(h.data(), cyc, sizeof(long long) * sms * ctas, cudaMemcpyDeviceToHost));
    double mean = 0;
    for (auto v : h) mean += (double)v;
    mean /= h.size();
    printf("%s, %d warps/SM: %.1f cycles per iteration per warp\n", name, 8 * ctas, mean / iters);
  };
  for (int ctas = 2; ctas <= 4; ctas *= 2) {
    hmma_mix<4, 0><<<sms * ctas, 256>>>(out, iters, cyc);
    CK(cudaDeviceSynchronize());
    mix_report("4 HMMA", ctas);
    hmma_mix<0, 32><<<sms * ctas, 256>>>(out, iters, cyc);
    CK(cudaDeviceSynchronize());
    mix_report("32 FFMA", ctas);
    hmma_mix<4, 32><<<sms * ctas, 256>>>(out, iters, cyc);
    CK(cudaDeviceSynchronize());
    mix_report("4 HMMA + 32 FFMA interleaved", ctas);
    hmma_mix<4, 16><<<sms * ctas, 256>>>(out, iters, cyc);
    CK(cudaDeviceSynchronize());
    mix_report("4 HMMA + 16 FFMA interleaved", ctas);
  }
  run_score<18>(sms);
  run_score<30>(sms);
  return 0;
}
