// K6 result sink: np.array2string(row) for every row of a float32 matrix, on the device.
//
// Reference: OutputCallback.on_predict_batch_end formats each reconstructed row with np.array2string and hands the text to
// the Kafka output sequence keyed by the running row index (cardata-v1.py:212-226; LSTM-…/cardata-v1.py:201-216) — a Python
// double loop per predict batch.  Here one thread formats one row with the host+device routine in float_format.cuh (checked on
// the CPU against NumPy), into its own shared-memory line buffer; the CTA then packs its rows' texts back to back into the
// caller's byte arena (one atomicAdd on the arena cursor per tile) and records (offset, length) per row — the same
// (bytes, offsets) shape the decode kernels consume, ready to be cut into produce requests.
//
// Layout: x tile [rows][n_cols + 1] floats in shared memory (coalesced load, the +1 keeps per-thread row reads conflict-free);
// line buffers at a stride of an odd number of words so that byte stores of the 32 lanes fall into different banks.
// Traffic per row: 4·n_cols bytes in, ≈ 12·n_cols bytes of text out + 12 bytes of index; the kernel is instruction-bound
// (digit generation), not HBM-bound.
#include "common.cuh"
#include "float_format.cuh"

namespace {

constexpr int FMT_NT = 128;

struct FormatArgs {
  const float* x;
  long long n;
  int n_cols, compat;
  int line_stride;  // bytes per line buffer (multiple of 4, odd number of words)
  uint8_t* text;
  long long text_cap;
  long long* offs;
  int32_t* len;
  unsigned long long* cursor;  // [0] bytes used, [1] rows dropped because the arena was full
};

__global__ void __launch_bounds__(FMT_NT) format_rows_kernel(const FormatArgs a) {
  extern __shared__ __align__(16) uint8_t smem[];
  float* s_x = reinterpret_cast<float*>(smem);
  const int xs = a.n_cols + 1;
  uint8_t* s_line = smem + (((size_t)FMT_NT * xs * 4 + 15) & ~(size_t)15);
  __shared__ int s_warp_sum[FMT_NT / 32];
  __shared__ int s_off[FMT_NT], s_len[FMT_NT];
  __shared__ long long s_base;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const long long n_tiles = (a.n + FMT_NT - 1) / FMT_NT;
  for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const long long row0 = tile * FMT_NT;
    const int rows = (int)min((long long)FMT_NT, a.n - row0);
    // coalesced tile load
    const float* src = a.x + row0 * a.n_cols;
    for (int i = tid; i < rows * a.n_cols; i += FMT_NT) {
      const int r = i / a.n_cols, c = i - r * a.n_cols;
      s_x[r * xs + c] = src[i];
    }
    __syncthreads();
    int len = 0;
    uint8_t* line = s_line + (size_t)tid * a.line_stride;
    if (tid < rows) {
      const float* xr = s_x + tid * xs;
      len = b2_format_row_f32(xr, a.n_cols, line, a.compat);
    }
    // exclusive prefix of the lengths over the tile
    int incl = len;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int v = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += v;
    }
    if (lane == 31) s_warp_sum[warp] = incl;
    __syncthreads();
    int before = 0, total = 0;
#pragma unroll
    for (int w = 0; w < FMT_NT / 32; ++w) {
      if (w < warp) before += s_warp_sum[w];
      total += s_warp_sum[w];
    }
    if (tid == 0) {
      const unsigned long long b = atomicAdd(a.cursor, (unsigned long long)total);
      long long base = (long long)b;
      if (base + total > a.text_cap) {  // arena full: the tile's rows are reported with length -1
        atomicAdd(a.cursor + 1, (unsigned long long)rows);
        base = -1;
      }
      s_base = base;
    }
    __syncthreads();
    const long long base = s_base;
    if (tid < rows) {
      a.offs[row0 + tid] = base < 0 ? -1 : base + before + incl - len;
      a.len[row0 + tid] = base < 0 ? -1 : len;
    }
    if (base >= 0) {
      // pack: warp w copies rows w, w+4, …; a row's start offset travels by shuffle-free recomputation from shared memory
      s_off[tid] = before + incl - len;
      s_len[tid] = len;
      __syncthreads();
      for (int r = warp; r < rows; r += FMT_NT / 32) {
        const uint8_t* lp = s_line + (size_t)r * a.line_stride;
        uint8_t* dst = a.text + base + s_off[r];
        const int l = s_len[r];
        for (int j = lane; j < l; j += 32) dst[j] = lp[j];
      }
    }
    __syncthreads();
  }
}

int line_stride_for(int n_cols) {
  int words = (b2_format_row_bound(n_cols) + 3) / 4;
  if ((words & 1) == 0) ++words;
  return words * 4;
}

}  // namespace

extern "C" int64_t b2_format_rows_bound(int n_cols) { return n_cols >= 1 && n_cols <= B2_FMT_MAX_N ? b2_format_row_bound(n_cols) : 0; }

extern "C" int b2_format_rows(b2_ctx* ctx, const float* x_dev, int64_t n, int n_cols, int numpy_compat, uint8_t* text_dev, int64_t text_cap,
                              int64_t* offs_dev, int32_t* len_dev, unsigned long long* cursor_dev, void* stream) {
  B2_REQUIRE(ctx, "b2_format_rows: NULL context");
  B2_REQUIRE(n >= 0, "b2_format_rows: n < 0");
  B2_REQUIRE(n_cols >= 1 && n_cols <= B2_FMT_MAX_N, "b2_format_rows: n_cols=%d out of range (1..%d)", n_cols, B2_FMT_MAX_N);
  B2_REQUIRE(numpy_compat >= B2_NPY_23 && numpy_compat <= B2_NPY_114, "b2_format_rows: numpy_compat=%d unknown", numpy_compat);
  if (n == 0) return B2_OK;
  B2_REQUIRE(x_dev && text_dev && offs_dev && len_dev && cursor_dev, "b2_format_rows: NULL device pointer");
  B2_REQUIRE(text_cap >= 0, "b2_format_rows: text_cap < 0");
  DeviceGuard g(ctx->device);
  FormatArgs a;
  a.x = x_dev;
  a.n = n;
  a.n_cols = n_cols;
  a.compat = numpy_compat;
  a.line_stride = line_stride_for(n_cols);
  a.text = text_dev;
  a.text_cap = text_cap;
  a.offs = reinterpret_cast<long long*>(offs_dev);
  a.len = len_dev;
  a.cursor = cursor_dev;
  const size_t smem = (((size_t)FMT_NT * (n_cols + 1) * 4 + 15) & ~(size_t)15) + (size_t)FMT_NT * a.line_stride;
  B2_REQUIRE(smem <= (size_t)ctx->smem_optin, "b2_format_rows: tile needs %zu bytes of shared memory (max %d)", smem, ctx->smem_optin);
  B2_CUDA(cudaFuncSetAttribute(format_rows_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int occ = 0;
  B2_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, format_rows_kernel, FMT_NT, smem));
  B2_REQUIRE(occ >= 1, "b2_format_rows: kernel does not fit on an SM");
  const long long n_tiles = (n + FMT_NT - 1) / FMT_NT;
  long long grid = (long long)ctx->sm_count * occ;
  if (grid > n_tiles) grid = n_tiles;
  format_rows_kernel<<<(unsigned)grid, FMT_NT, smem, (cudaStream_t)stream>>>(a);
  B2_CUDA(cudaGetLastError());
  return B2_OK;
}
