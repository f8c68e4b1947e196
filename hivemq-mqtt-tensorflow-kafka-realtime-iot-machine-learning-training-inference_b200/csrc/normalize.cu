// normalize_fn as a stand-alone operator, and the exhaustive check of the division-free form the fused kernels use.
//
// Replaces  normalize_fn / scale_fn   AUTOENCODER-TensorFlow-IO-Kafka/cardata-v1.py:40-131 (float32 order: the columns are float32
//                                     when scale_fn runs), LSTM-TensorFlow-IO-Kafka/cardata-v1.py:53-143
// The fused score / decode / fit kernels never materialise the normalised row; this entry point does, with the very same
// device function (norm_pair_fast, ae_math.cuh), so that its bits can be compared with the reference's arithmetic
// `(v - lo) / (hi - lo) * 2.0 - 1.0` evaluated in IEEE binary32.  Purely HBM-bound: 4 B read + 4 B written per value.
#include "ae_math.cuh"

namespace {

using NormMax = NormP<B2_MAX_WIDTH>;

struct NormalizeArgs {
  NormMax nm;
  const float* x;
  float* out;
  long long n;  // rows
  int d;        // features per row
};

// one thread per pair of adjacent features; an odd last feature is paired with a dummy
__global__ void __launch_bounds__(256) normalize_kernel(const __grid_constant__ NormalizeArgs a) {
  const int pairs = (a.d + 1) / 2;
  const long long total = a.n * pairs;
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
    const long long r = t / pairs;
    const int i = (int)(t - r * pairs) * 2;
    const bool two = i + 1 < a.d;
    const float* src = a.x + r * a.d + i;
    const float2 v = make_float2(__ldg(src), two ? __ldg(src + 1) : 0.0f);
    const float2 o = norm_pair_fast<B2_MAX_WIDTH>(v, a.nm, i);
    float* dst = a.out + r * a.d + i;
    dst[0] = o.x;
    if (two) dst[1] = o.y;
  }
}

struct SelfTestArgs {
  NormMax nm;
  int d;
  unsigned long long first, count;  // float32 bit patterns [first, first + count)
  unsigned long long* out;          // [0] mismatches, [1] feature << 32 | bits of one of them
};

// every bit pattern × every feature: the packed Markstein form against the plain IEEE sequence sub → div → fma
__global__ void __launch_bounds__(256) normalize_selftest_kernel(const __grid_constant__ SelfTestArgs a) {
  unsigned long long bad = 0, where = 0;
  for (unsigned long long t = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; t < a.count;
       t += (unsigned long long)gridDim.x * blockDim.x) {
    const uint32_t bits = (uint32_t)(a.first + t);
    const float v = __uint_as_float(bits);
    for (int i = 0; i < a.d; i += 2) {
      const float2 fast = norm_pair_fast<B2_MAX_WIDTH>(make_float2(v, v), a.nm, i);
      const float r0 = norm_feature(v, a.nm.kind[i], a.nm.lo[i], a.nm.range[i]);
      const float r1 = i + 1 < a.d ? norm_feature(v, a.nm.kind[i + 1], a.nm.lo[i + 1], a.nm.range[i + 1]) : fast.y;
      const bool ok0 = __float_as_uint(fast.x) == __float_as_uint(r0) || (fast.x != fast.x && r0 != r0);
      const bool ok1 = __float_as_uint(fast.y) == __float_as_uint(r1) || (fast.y != fast.y && r1 != r1);
      if (!ok0) ++bad, where = ((unsigned long long)i << 32) | bits;
      if (!ok1) ++bad, where = ((unsigned long long)(i + 1) << 32) | bits;
    }
  }
  if (bad) {
    atomicAdd(&a.out[0], bad);
    a.out[1] = where;
  }
}

int fill(const b2_norm* norm, int d, NormMax& nm) {
  B2_REQUIRE(norm && norm->n == d && d >= 1 && d <= B2_MAX_WIDTH, "normalize: table has %d features, rows have %d", norm ? norm->n : 0, d);
  for (int j = 0; j < d; ++j)
    B2_REQUIRE(norm->kind[j] >= B2_NORM_IDENTITY && norm->kind[j] <= B2_NORM_SCALE, "normalize: kind[%d]=%d unknown", j, norm->kind[j]);
  norm_fill<B2_MAX_WIDTH>(norm, nm, d);
  return B2_OK;
}

}  // namespace

extern "C" int b2_normalize(b2_ctx* ctx, const float* x_dev, int64_t n, int d, const b2_norm* norm, float* out_dev, void* stream) {
  B2_NORM_F32_ONLY(norm, "b2_normalize");
  B2_REQUIRE(ctx, "b2_normalize: NULL context");
  B2_REQUIRE(n >= 0, "b2_normalize: n < 0");
  NormalizeArgs a;
  int rc = fill(norm, d, a.nm);
  if (rc) return rc;
  if (n == 0) return B2_OK;
  B2_REQUIRE(x_dev && out_dev, "b2_normalize: NULL device pointer");
  DeviceGuard g(ctx->device);
  a.x = x_dev;
  a.out = out_dev;
  a.n = n;
  a.d = d;
  const long long total = n * ((d + 1) / 2);
  long long blocks = (total + 255) / 256;
  const long long cap = (long long)ctx->sm_count * 8;
  if (blocks > cap) blocks = cap;
  normalize_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(a);
  B2_CUDA(cudaGetLastError());
  return B2_OK;
}

extern "C" int b2_selftest_normalize(b2_ctx* ctx, const b2_norm* norm, int d, uint64_t first_bits, uint64_t count, unsigned long long* out_dev,
                                     void* stream) {
  B2_REQUIRE(ctx && out_dev, "b2_selftest_normalize: NULL argument");
  B2_REQUIRE(first_bits + count <= (1ull << 32), "b2_selftest_normalize: range runs past 2^32");
  SelfTestArgs a;
  int rc = fill(norm, d, a.nm);
  if (rc) return rc;
  if (count == 0) return B2_OK;
  DeviceGuard g(ctx->device);
  a.d = d;
  a.first = first_bits;
  a.count = count;
  a.out = out_dev;
  normalize_selftest_kernel<<<(unsigned)(ctx->sm_count * 8), 256, 0, (cudaStream_t)stream>>>(a);
  B2_CUDA(cudaGetLastError());
  return B2_OK;
}
