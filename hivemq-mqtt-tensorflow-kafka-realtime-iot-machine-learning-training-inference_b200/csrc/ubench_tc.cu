// Stand-alone check of the tcgen05 building block used by the tensor-core score variant (not part of libb2sensor.so):
//   D[128 x N] = A[128 x K] * B[N x K]^T in "3xTF32" (A_hi*B_hi + A_lo*B_hi + A_hi*B_lo), A written to TMEM by the
//   threads that own the rows (tcgen05.st, thread t = TMEM lane t), B in shared memory (K-major, no swizzle), D read
//   back with tcgen05.ld.  Prints the max error against a float64 CPU product.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ubench_tc ubench_tc.cu && ./ubench_tc
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,"
      "%25,%26,%27,%28,%29,%30,%31,%32};" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]),
      "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]), "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]),
      "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
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

__device__ __forceinline__ uint64_t make_bdesc(uint32_t saddr, int lbo_bytes, int sbo_bytes) {
  return (uint64_t)((saddr >> 4) & 0x3FFF) | ((uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16) | ((uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32) |
         (1ull << 46);
}
__device__ __forceinline__ void mma_tf32_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(d_tmem), "r"(a_tmem), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}

// B[n][k] (n < N, k < K) → canonical K-major no-swizzle tile: 16-byte chunk (4 k's) of row n at  (k/4)*(N*16) + n*16
__global__ void __launch_bounds__(128) tc_test(const float* A, const float* B, float* D, int N, int K) {
  __shared__ __align__(128) float bhi[32 * 32], bmid[32 * 32], blo[32 * 32];
  __shared__ uint32_t s_tmem;
  __shared__ __align__(8) uint64_t mbar;
  const int tid = threadIdx.x, warp = tid >> 5;
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s_tmem)), "r"(128));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&mbar)));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  for (int i = tid; i < N * K; i += 128) {
    const int n = i / K, k = i % K;
    const float w = B[n * K + k];
    // exact 3-way split of the weight: 11 + 11 + 2 significant bits, each part representable in TF32
    const float hi = __uint_as_float(__float_as_uint(w) & 0xFFFFE000u);
    const float r1 = w - hi;
    const float mid = __uint_as_float(__float_as_uint(r1) & 0xFFFFE000u);
    const int off = (k >> 2) * (N * 4) + n * 4 + (k & 3);
    bhi[off] = hi;
    bmid[off] = mid;
    blo[off] = r1 - mid;
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tbase = s_tmem;
  const uint32_t lane_base = tbase + ((uint32_t)(warp * 32) << 16);
  uint32_t hi[32], lo[32];
  for (int k = 0; k < 32; ++k) {
    const float x = k < K ? A[tid * K + k] : 0.f;
    uint32_t hb;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(hb) : "f"(x));  // round-to-nearest split: |lo| <= 2^-12 |x|
    hi[k] = hb;
    lo[k] = __float_as_uint(x - __uint_as_float(hb));
  }
  tmem_st32(lane_base + 0, hi);
  tmem_st32(lane_base + 32, lo);
  asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (tid == 0) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | ((128u >> 4) << 24);
    const uint32_t d = tbase + 64;
    uint32_t acc = 0;
    for (int pass = 0; pass < 5; ++pass) {  // A_hi*(B_hi + B_mid + B_lo) + A_lo*(B_hi + B_mid)
      const uint32_t a_col = pass >= 3 ? 32 : 0;
      const float* bsm = (pass == 0 || pass == 3) ? bhi : ((pass == 1 || pass == 4) ? bmid : blo);
      for (int j = 0; j < K / 8; ++j) {
        const uint64_t bd = make_bdesc(smem_u32(bsm) + j * 2 * (N * 16), N * 16, 128);
        mma_tf32_ts(d, tbase + a_col + 8 * j, bd, idesc, acc);
        acc = 1;
      }
    }
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&mbar)) : "memory");
  }
  {
    uint32_t ok = 0;
    while (!ok) {
      asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                   : "=r"(ok)
                   : "r"(smem_u32(&mbar)), "r"(0)
                   : "memory");
    }
  }
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  uint32_t out[32];
  tmem_ld32(lane_base + 64, out);
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
  for (int n = 0; n < N; ++n) D[tid * N + n] = __uint_as_float(out[n]);
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tbase), "r"(128));
}

static int run(int N, int K) {
  float *hA = (float*)malloc(128 * K * 4), *hB = (float*)malloc(N * K * 4), *hD = (float*)malloc(128 * N * 4);
  srand(314 + N + K);
  for (int i = 0; i < 128 * K; ++i) hA[i] = (float)rand() / RAND_MAX * 4.f - 2.f;
  for (int i = 0; i < N * K; ++i) hB[i] = (float)rand() / RAND_MAX * 2.f - 1.f;
  float *dA, *dB, *dD;
  cudaMalloc(&dA, 128 * K * 4);
  cudaMalloc(&dB, N * K * 4);
  cudaMalloc(&dD, 128 * N * 4);
  cudaMemcpy(dA, hA, 128 * K * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(dB, hB, N * K * 4, cudaMemcpyHostToDevice);
  cudaMemset(dD, 0xFF, 128 * N * 4);
  tc_test<<<1, 128>>>(dA, dB, dD, N, K);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) {
    printf("N=%d K=%d CUDA error: %s\n", N, K, cudaGetErrorString(e));
    return 1;
  }
  cudaMemcpy(hD, dD, 128 * N * 4, cudaMemcpyDeviceToHost);
  double max_err = 0, max_ref = 0, max_err32 = 0, sum_err = 0, sum_err32 = 0;
  for (int m = 0; m < 128; ++m)
    for (int n = 0; n < N; ++n) {
      double ref = 0;
      float f32 = 0.f;
      for (int k = 0; k < K; ++k) {
        ref += (double)hA[m * K + k] * (double)hB[n * K + k];
        f32 = fmaf(hA[m * K + k], hB[n * K + k], f32);
      }
      max_err = fmax(max_err, fabs(ref - (double)hD[m * N + n]));
      max_err32 = fmax(max_err32, fabs(ref - (double)f32));
      sum_err += fabs(ref - (double)hD[m * N + n]);
      sum_err32 += fabs(ref - (double)f32);
      max_ref = fmax(max_ref, fabs(ref));
    }
  printf("   fp32 FMA chain on the CPU: max |err| = %.3g, mean %.3g;  tensor path mean |err| = %.3g\n", max_err32, sum_err32 / (128 * N), sum_err / (128 * N));
  printf("N=%d K=%d  max |err| = %.3g  (max |ref| = %.3g)  D[0][0..3] = %g %g %g %g\n", N, K, max_err, max_ref, hD[0], hD[1], hD[2], hD[3]);
  return max_err < 1e-4 ? 0 : 2;
}

int main() {
  int rc = 0;
  rc |= run(16, 32);
  rc |= run(32, 8);
  rc |= run(16, 16);
  return rc;
}
