// K7 evaluation metrics of the notebooks on the GPU: the sorted-threshold sweep behind
//   roc_curve / auc / precision_recall_curve   nb2.0 cells 23-24 (sklearn.metrics, on error_df.Reconstruction_error)
//   confusion_matrix                           nb2.0 cell 27 (pred_y = mse > threshold_fixed)
// The arithmetic owner is scikit-learn (confusion_matrix_at_thresholds: stable sort of the scores in decreasing order, one
// threshold per distinct score, cumulative true / false positives); it is installed, so the oracle for this row is the
// library itself.
//
//   pack      score → 64-bit key: (order-reversing map of the float bits) << 32 | label; padding keys sort last
//   sort      bitonic network on the keys: strides ≥ 2048 in global memory, the rest of every stage in shared memory
//   sweep     one pass over the sorted keys: inclusive scan of (is positive, is last of its score) as one 64-bit add,
//             three kernels (block sums, scan of the sums, apply) — at every "last of its score" position r the outputs are
//             thresholds[r] = score, tps[r] = positives so far, fps[r] = 1 + index − tps[r]
// Everything is integer / comparison work: results are exact, not approximations of the library's.
#include "common.cuh"

namespace {

constexpr int MT_NT = 1024;          // threads per block in the shared-memory sort / scan kernels
constexpr int MT_TILE = 2 * MT_NT;   // keys per block in the shared-memory sort

__device__ __forceinline__ uint32_t desc_key(uint32_t bits) {
  if (bits == 0x80000000u) bits = 0u;  // -0.0 and +0.0 are one threshold
  const uint32_t asc = bits ^ ((bits >> 31) ? 0xFFFFFFFFu : 0x80000000u);
  return ~asc;
}
__device__ __forceinline__ float score_of(uint32_t dk) {
  const uint32_t asc = ~dk;
  return __uint_as_float(asc ^ ((asc >> 31) ? 0x80000000u : 0xFFFFFFFFu));
}

__global__ void mt_pack(const float* __restrict__ score, const uint8_t* __restrict__ y, long long n, long long n2,
                        unsigned long long* __restrict__ key, unsigned long long* __restrict__ nan_count) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n2) return;
  if (i >= n) {
    key[i] = ~0ull;
    return;
  }
  const uint32_t b = __float_as_uint(score[i]);
  if ((b & 0x7FFFFFFFu) > 0x7F800000u) atomicAdd(nan_count, 1ull);
  key[i] = ((unsigned long long)desc_key(b) << 32) | (y[i] ? 1ull : 0ull);
}

__device__ __forceinline__ void cmp_swap(unsigned long long& a, unsigned long long& b, bool up) {
  if ((a > b) == up) {
    const unsigned long long t = a;
    a = b;
    b = t;
  }
}

// one compare-exchange step (stride j of stage k) in global memory
__global__ void mt_bitonic_global(unsigned long long* __restrict__ key, long long n2, long long j, long long k) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;  // one thread per pair
  if (t >= n2 / 2) return;
  const long long i = 2 * t - (t & (j - 1));  // insert a zero bit at position log2(j)
  const bool up = (i & k) == 0;
  unsigned long long a = key[i], b = key[i + j];
  if ((a > b) == up) {
    key[i] = b;
    key[i + j] = a;
  }
}

// all steps with stride ≤ MT_NT of stage k (FULL: every stage up to k = MT_TILE) on one tile in shared memory
template <bool FULL>
__global__ void __launch_bounds__(MT_NT) mt_bitonic_shared(unsigned long long* __restrict__ key, long long k_stage) {
  __shared__ unsigned long long s[MT_TILE];
  const long long base = (long long)blockIdx.x * MT_TILE;
  const int t = threadIdx.x;
  s[t] = key[base + t];
  s[t + MT_NT] = key[base + t + MT_NT];
  __syncthreads();
  for (long long k = FULL ? 2 : k_stage; k <= (FULL ? (long long)MT_TILE : k_stage); k <<= 1) {
    for (int j = (int)min((long long)MT_NT, k >> 1); j >= 1; j >>= 1) {
      const int i = 2 * t - (t & (j - 1));
      const bool up = ((base + i) & k) == 0;
      cmp_swap(s[i], s[i + j], up);
      __syncthreads();
    }
  }
  key[base + t] = s[t];
  key[base + t + MT_NT] = s[t + MT_NT];
}


// (is positive, is last of its score) of sorted position i as one 64-bit addend: low word counts positives, high word thresholds
__device__ __forceinline__ unsigned long long sweep_term(const unsigned long long* __restrict__ key, long long i, long long n) {
  if (i >= n) return 0ull;
  const unsigned long long a = key[i];
  const bool last = i == n - 1 || (uint32_t)(key[i + 1] >> 32) != (uint32_t)(a >> 32);
  return (a & 1ull) | ((unsigned long long)(last ? 1 : 0) << 32);
}

__device__ __forceinline__ unsigned long long block_scan_incl(unsigned long long v, unsigned long long* s_warp, unsigned long long* total) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const unsigned long long u = __shfl_up_sync(0xffffffffu, v, o);
    if (lane >= o) v += u;
  }
  if (lane == 31) s_warp[warp] = v;
  __syncthreads();
  if (warp == 0) {
    unsigned long long w = s_warp[lane];  // MT_NT / 32 == 32 warps
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const unsigned long long u = __shfl_up_sync(0xffffffffu, w, o);
      if (lane >= o) w += u;
    }
    s_warp[lane] = w;
  }
  __syncthreads();
  if (total) *total = s_warp[31];
  const unsigned long long before = warp ? s_warp[warp - 1] : 0ull;
  __syncthreads();
  return v + before;
}

__global__ void __launch_bounds__(MT_NT) mt_block_sums(const unsigned long long* __restrict__ key, long long n, unsigned long long* __restrict__ sums) {
  __shared__ unsigned long long s_warp[32];
  const long long i = ((long long)blockIdx.x * MT_NT + threadIdx.x) * 2;
  unsigned long long total;
  block_scan_incl(sweep_term(key, i, n) + sweep_term(key, i + 1, n), s_warp, &total);
  if (threadIdx.x == 0) sums[blockIdx.x] = total;
}

// exclusive scan of the block sums in place (one block; chunks of MT_NT)
__global__ void __launch_bounds__(MT_NT) mt_scan_sums(unsigned long long* __restrict__ sums, long long nb, unsigned long long* __restrict__ grand) {
  __shared__ unsigned long long s_warp[32];
  unsigned long long carry = 0;
  for (long long c = 0; c < nb; c += MT_NT) {
    const long long i = c + threadIdx.x;
    const unsigned long long v = i < nb ? sums[i] : 0ull;
    unsigned long long total;
    const unsigned long long incl = block_scan_incl(v, s_warp, &total);
    if (i < nb) sums[i] = carry + incl - v;
    carry += total;
  }
  if (threadIdx.x == 0) *grand = carry;
}

__global__ void __launch_bounds__(MT_NT) mt_sweep(const unsigned long long* __restrict__ key, long long n, const unsigned long long* __restrict__ sums,
                                                   float* __restrict__ thr, long long* __restrict__ tps, long long* __restrict__ fps) {
  __shared__ unsigned long long s_warp[32];
  const long long i = ((long long)blockIdx.x * MT_NT + threadIdx.x) * 2;
  const unsigned long long t0 = sweep_term(key, i, n), t1 = sweep_term(key, i + 1, n);
  const unsigned long long incl1 = block_scan_incl(t0 + t1, s_warp, nullptr) + sums[blockIdx.x];
  const unsigned long long incl0 = incl1 - t1;
  if (t0 >> 32) {
    const long long r = (long long)(incl0 >> 32) - 1, tp = (long long)(uint32_t)incl0;
    thr[r] = score_of((uint32_t)(key[i] >> 32));
    tps[r] = tp;
    fps[r] = 1 + i - tp;
  }
  if (t1 >> 32) {
    const long long r = (long long)(incl1 >> 32) - 1, tp = (long long)(uint32_t)incl1;
    thr[r] = score_of((uint32_t)(key[i + 1] >> 32));
    tps[r] = tp;
    fps[r] = 2 + i - tp;
  }
}

__global__ void mt_confusion(const uint8_t* __restrict__ yt, const uint8_t* __restrict__ yp, long long n, unsigned long long* __restrict__ cm) {
  unsigned c[4] = {0, 0, 0, 0};
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const int k = (yt[i] ? 2 : 0) + (yp[i] ? 1 : 0);
#pragma unroll
    for (int q = 0; q < 4; ++q) c[q] += k == q;
  }
#pragma unroll
  for (int q = 0; q < 4; ++q) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) c[q] += __shfl_xor_sync(0xffffffffu, c[q], o);
    if ((threadIdx.x & 31) == 0 && c[q]) atomicAdd(&cm[q], (unsigned long long)c[q]);
  }
}

long long pow2_at_least(long long n) {
  long long p = MT_TILE;
  while (p < n) p <<= 1;
  return p;
}

}  // namespace

extern "C" int64_t b2_threshold_sweep_scratch_bytes(int64_t n) {
  if (n < 0 || n > (1ll << 31) - 1) return 0;
  const long long n2 = pow2_at_least(n);
  return n2 * 8 + (n2 / MT_TILE + 2) * 8 + 16;
}

extern "C" int b2_threshold_sweep(b2_ctx* ctx, const float* score_dev, const uint8_t* y_dev, int64_t n, float* thresholds_dev, int64_t* tps_dev,
                                  int64_t* fps_dev, int64_t* n_thresholds_dev, void* scratch_dev, int64_t scratch_bytes, void* stream) {
  B2_REQUIRE(ctx, "b2_threshold_sweep: NULL context");
  B2_REQUIRE(n >= 1 && n <= (1ll << 31) - 1, "b2_threshold_sweep: n=%lld out of range", (long long)n);
  B2_REQUIRE(score_dev && y_dev && thresholds_dev && tps_dev && fps_dev && n_thresholds_dev && scratch_dev, "b2_threshold_sweep: NULL device pointer");
  B2_REQUIRE(scratch_bytes >= b2_threshold_sweep_scratch_bytes(n), "b2_threshold_sweep: scratch of %lld bytes, need %lld", (long long)scratch_bytes,
             (long long)b2_threshold_sweep_scratch_bytes(n));
  DeviceGuard g(ctx->device);
  cudaStream_t st = (cudaStream_t)stream;
  const long long n2 = pow2_at_least(n), nb = n2 / MT_TILE;
  unsigned long long* key = reinterpret_cast<unsigned long long*>(scratch_dev);
  unsigned long long* sums = key + n2;
  unsigned long long* nan_count = sums + nb;      // [1]
  unsigned long long* grand = nan_count + 1;      // [1]: low word positives, high word thresholds
  B2_CUDA(cudaMemsetAsync(nan_count, 0, 16, st));
  mt_pack<<<(unsigned)((n2 + 255) / 256), 256, 0, st>>>(score_dev, y_dev, n, n2, key, nan_count);
  mt_bitonic_shared<true><<<(unsigned)nb, MT_NT, 0, st>>>(key, 0);
  for (long long k = 2 * MT_TILE; k <= n2; k <<= 1) {
    for (long long j = k >> 1; j >= MT_TILE; j >>= 1) mt_bitonic_global<<<(unsigned)((n2 / 2 + 255) / 256), 256, 0, st>>>(key, n2, j, k);
    mt_bitonic_shared<false><<<(unsigned)nb, MT_NT, 0, st>>>(key, k);
  }
  mt_block_sums<<<(unsigned)nb, MT_NT, 0, st>>>(key, n, sums);
  mt_scan_sums<<<1, MT_NT, 0, st>>>(sums, nb, grand);
  mt_sweep<<<(unsigned)nb, MT_NT, 0, st>>>(key, n, sums, thresholds_dev, reinterpret_cast<long long*>(tps_dev), reinterpret_cast<long long*>(fps_dev));
  // n_thresholds_dev[0] = number of thresholds, [1] = NaN scores seen (the caller refuses those like scikit-learn does)
  B2_CUDA(cudaMemcpyAsync(n_thresholds_dev, reinterpret_cast<uint32_t*>(grand) + 1, 4, cudaMemcpyDeviceToDevice, st));
  B2_CUDA(cudaMemsetAsync(reinterpret_cast<uint32_t*>(n_thresholds_dev) + 1, 0, 4, st));
  B2_CUDA(cudaMemcpyAsync(n_thresholds_dev + 1, nan_count, 8, cudaMemcpyDeviceToDevice, st));
  B2_CUDA(cudaGetLastError());
  return B2_OK;
}

extern "C" int b2_confusion_matrix(b2_ctx* ctx, const uint8_t* y_true_dev, const uint8_t* y_pred_dev, int64_t n, int64_t* counts_dev, void* stream) {
  B2_REQUIRE(ctx, "b2_confusion_matrix: NULL context");
  B2_REQUIRE(n >= 0, "b2_confusion_matrix: n < 0");
  B2_REQUIRE(counts_dev, "b2_confusion_matrix: NULL counts");
  DeviceGuard g(ctx->device);
  B2_CUDA(cudaMemsetAsync(counts_dev, 0, 32, (cudaStream_t)stream));
  if (n == 0) return B2_OK;
  B2_REQUIRE(y_true_dev && y_pred_dev, "b2_confusion_matrix: NULL device pointer");
  const long long blocks = min((long long)ctx->sm_count * 8, (long long)((n + 255) / 256));
  mt_confusion<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(y_true_dev, y_pred_dev, n, reinterpret_cast<unsigned long long*>(counts_dev));
  B2_CUDA(cudaGetLastError());
  return B2_OK;
}
