// Micro-benchmark (standalone binary, not part of libb2sensor.so): FP32 FMA issue rates on sm_100a.
// Answers the design question behind ae_math.cuh: how many FMA lanes/clk/SM do FFMA (3-reg, uniform-reg operand)
// and the packed FFMA2 forms sustain, alone and with LDCU weight fetches in the instruction stream.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ubench_fma ubench_fma.cu && ./ubench_fma
#include <cuda_runtime.h>
#include <stdio.h>

struct W { float w[1024]; };
constexpr int ITERS = 64;

template <int MODE>
__global__ void __launch_bounds__(256) k(const __grid_constant__ W p, float* out, long long* cyc) {
  float2 a[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) a[i] = make_float2(threadIdx.x * 1e-3f + i, i * 0.5f);
  float2 x = make_float2(out[threadIdx.x], out[threadIdx.x + 256]);
  float r0 = out[threadIdx.x + 512], r1 = out[threadIdx.x + 768];
  __syncthreads();
  long long t0 = clock64();
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int u = 0; u < 16; ++u) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        if (MODE == 0) {  // FFMA reg,reg,reg (weights pre-loaded in regs)
          a[j].x = fmaf(x.x, r0, a[j].x);
          a[j].y = fmaf(x.y, r1, a[j].y);
        } else if (MODE == 1) {  // FFMA reg, UR (constant → uniform reg), reg
          a[j].x = fmaf(x.x, p.w[u * 16 + 2 * j], a[j].x);
          a[j].y = fmaf(x.y, p.w[u * 16 + 2 * j + 1], a[j].y);
        } else if (MODE == 2) {  // FFMA2 row-packed: (xA,xB) * splat(w)
          a[j] = __ffma2_rn(x, make_float2(p.w[u * 8 + j], p.w[u * 8 + j]), a[j]);
        } else if (MODE == 3) {  // FFMA2 output-packed: splat(x) * (w0,w1)
          a[j] = __ffma2_rn(make_float2(x.x, x.x), make_float2(p.w[u * 16 + 2 * j], p.w[u * 16 + 2 * j + 1]), a[j]);
        } else if (MODE == 4) {  // FFMA2 all-register operands
          a[j] = __ffma2_rn(x, make_float2(r0, r1), a[j]);
        }
      }
    }
    x.x += 1e-7f;
  }
  long long t1 = clock64();
  float s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += a[i].x + a[i].y;
  out[blockIdx.x * 256 + threadIdx.x] = s;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int MODE>
static void run(const char* name, int ctas_per_sm, int fma_per_inst) {
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  int grid = sms * ctas_per_sm;
  float* out;
  long long* cyc;
  cudaMalloc(&out, sizeof(float) * 256 * (grid + 8));
  cudaMemset(out, 0, sizeof(float) * 256 * (grid + 8));
  cudaMalloc(&cyc, sizeof(long long) * grid);
  W w;
  for (int i = 0; i < 1024; ++i) w.w[i] = 1e-3f * i;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  k<MODE><<<grid, 256>>>(w, out, cyc);
  cudaEventRecord(e0);
  k<MODE><<<grid, 256>>>(w, out, cyc);
  cudaEventRecord(e1);
  cudaDeviceSynchronize();
  float ms;
  cudaEventElapsedTime(&ms, e0, e1);
  long long* h = new long long[grid];
  cudaMemcpy(h, cyc, sizeof(long long) * grid, cudaMemcpyDeviceToHost);
  double avg = 0;
  for (int i = 0; i < grid; ++i) avg += h[i];
  avg /= grid;
  const double inst_per_thread = (double)ITERS * 16 * 8 * (MODE <= 1 ? 2 : 1);
  const double fma_per_thread = (double)ITERS * 16 * 8 * 2;
  const double lanes = ctas_per_sm * 256.0 * fma_per_thread / avg;          // FMA lanes / clk / SM
  const double ipc = ctas_per_sm * 8.0 * inst_per_thread / avg / 4.0;        // warp-inst / clk / SMSP
  printf("%-34s ctas/SM=%d  %.0f cyc  FMA/clk/SM=%6.1f  FMA-inst/clk/SMSP=%.3f  %.2f TFLOP/s (event-timed)\n", name, ctas_per_sm,
         avg, lanes, ipc, 2.0 * grid * 256.0 * fma_per_thread / (ms * 1e-3) / 1e12);
  (void)fma_per_inst;
  cudaFree(out);
  cudaFree(cyc);
  delete[] h;
}

int main() {
  for (int c : {1, 2, 4}) {
    run<0>("FFMA  reg*reg+reg", c, 1);
    run<1>("FFMA  reg*UR+reg", c, 1);
    run<2>("FFMA2 pair*UR.F32+pair (row-packed)", c, 2);
    run<3>("FFMA2 R.F32*URpair+pair (out-packed)", c, 2);
    run<4>("FFMA2 pair*pair+pair (all reg)", c, 2);
  }
  return 0;
}
