// K8 Avro encoder: float rows (+ label) → Avro binary records, the inverse of the decode walk.
//
// Two uses on the reference's path:
//   * the KSQL projection  CREATE STREAM SENSOR_DATA_S_AVRO WITH (VALUE_FORMAT='AVRO') AS SELECT * FROM SENSOR_DATA_S
//     (infrastructure/confluent/01_installConfluentPlatform.sh:242): JSON topic → Avro topic = b2_decode_json + this kernel;
//   * the result sink as Avro instead of np.array2string text, the TODO of python-scripts/README.md:108.
// Encoding rules are Avro 1.9's binary encoding, field by field of the same host-compiled program the decoder interprets:
// int / long = zig-zag varint of the value rounded to the nearest integer, float = 4 bytes LE, double = the float widened to
// 8 bytes LE, ["null", T] unions = one branch byte (null when the value is NaN), the label member = "false" / "true" / null,
// optional Confluent framing (0x00 + big-endian schema id) in front.
//
// Two passes so that the output is a proper record batch in row order (bytes back to back + int32 offsets[n+1], what the decode
// kernels consume): sizes (b2_avro_encode_sizes) → exclusive scan by the caller → encode.  In the encode pass a warp stages its
// 32 records in shared memory (their bytes are one contiguous range of the output) and copies the range out coalesced.
#include "common.cuh"

namespace {

enum { AV_INT = 1, AV_LONG, AV_FLOAT, AV_DOUBLE, AV_STRING, AV_BOOLEAN, AV_BYTES, AV_NULL };
constexpr int ENC_NT = 256;

struct EncP {
  int32_t n_fields, n_out, prefix, schema_id;
  b2_avro_field f[B2_MAX_FIELDS];
};

struct EncArgs {
  const float* x;
  const uint8_t* label;
  long long n;
  EncP p;
  int32_t* size;         // sizes pass
  const int32_t* offs;   // encode pass: int32[n+1]
  uint8_t* bytes;
  int warp_cap;          // bytes of shared memory per warp (encode pass)
};

__device__ __forceinline__ int varint_len(unsigned long long u) {
  int l = 1;
  while (u >= 0x80ull) {
    u >>= 7;
    ++l;
  }
  return l;
}
__device__ __forceinline__ unsigned long long zigzag_of(long long v) { return ((unsigned long long)v << 1) ^ (unsigned long long)(v >> 63); }

// One record.  WRITE = false: returns the size only.  `out` may be shared or global memory.
template <bool WRITE>
__device__ __forceinline__ int encode_record(const EncP& p, const float* __restrict__ xrow, int lab, uint8_t* __restrict__ out) {
  int pos = 0;
  if (p.prefix) {
    if (WRITE) {
      out[0] = 0;
      out[1] = (uint8_t)(p.schema_id >> 24);
      out[2] = (uint8_t)(p.schema_id >> 16);
      out[3] = (uint8_t)(p.schema_id >> 8);
      out[4] = (uint8_t)p.schema_id;
    }
    pos = 5;
  }
  for (int i = 0; i < p.n_fields; ++i) {
    const int type = p.f[i].type, nullable = p.f[i].nullable, col = p.f[i].out;
    float v = 0.f;
    bool present = true;
    if (col >= 0) {
      v = xrow[col];
      present = !(nullable && v != v);  // NaN in a nullable column = null
    } else if (col == -2) {
      present = lab == 0 || lab == 1 || !nullable;
    } else {
      present = !nullable;  // a member the rows do not carry: null where the schema allows it, else the type's zero
    }
    if (nullable) {
      const int null_branch = nullable == 1 ? 0 : 1;
      const int branch = present ? 1 - null_branch : null_branch;
      if (WRITE) out[pos] = (uint8_t)(branch << 1);  // zig-zag of 0 / 1
      ++pos;
    }
    if (!present) continue;
    if (type == AV_FLOAT) {
      if (WRITE) {
        const uint32_t b = __float_as_uint(v);
#pragma unroll
        for (int k = 0; k < 4; ++k) out[pos + k] = (uint8_t)(b >> (8 * k));
      }
      pos += 4;
    } else if (type == AV_DOUBLE) {
      if (WRITE) {
        const unsigned long long b = (unsigned long long)__double_as_longlong((double)v);
#pragma unroll
        for (int k = 0; k < 8; ++k) out[pos + k] = (uint8_t)(b >> (8 * k));
      }
      pos += 8;
    } else if (type == AV_INT || type == AV_LONG) {
      const long long iv = type == AV_INT ? (long long)__float2int_rn(v) : __float2ll_rn(v);
      unsigned long long u = zigzag_of(iv);
      if (WRITE) {
        while (u >= 0x80ull) {
          out[pos++] = (uint8_t)(u | 0x80ull);
          u >>= 7;
        }
        out[pos++] = (uint8_t)u;
      } else {
        pos += varint_len(u);
      }
    } else if (type == AV_BOOLEAN) {
      if (WRITE) out[pos] = v != 0.f ? 1 : 0;
      ++pos;
    } else if (type == AV_STRING || type == AV_BYTES) {
      // only the label member carries text: "false" / "true"; anything else is the empty string
      const int L = col == -2 ? (lab == 0 ? 5 : (lab == 1 ? 4 : 0)) : 0;
      if (WRITE) {
        out[pos] = (uint8_t)(L << 1);
        const char* t = lab == 0 ? "false" : "true";
        for (int k = 0; k < L; ++k) out[pos + 1 + k] = (uint8_t)t[k];
      }
      pos += 1 + L;
    }
  }
  return pos;
}

__global__ void __launch_bounds__(ENC_NT) avro_sizes_kernel(const EncArgs a) {
  for (long long i = (long long)blockIdx.x * ENC_NT + threadIdx.x; i < a.n; i += (long long)gridDim.x * ENC_NT)
    a.size[i] = encode_record<false>(a.p, a.x + i * a.p.n_out, a.label ? a.label[i] : 2, nullptr);
}

__global__ void __launch_bounds__(ENC_NT) avro_encode_kernel(const EncArgs a) {
  extern __shared__ __align__(16) uint8_t smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint8_t* buf = smem + (size_t)warp * a.warp_cap;
  const long long n_chunks = (a.n + 31) / 32;
  for (long long chunk = (long long)blockIdx.x * (ENC_NT / 32) + warp; chunk < n_chunks; chunk += (long long)gridDim.x * (ENC_NT / 32)) {
    const long long row0 = chunk * 32;
    const int rows = (int)min(32ll, a.n - row0);
    const int o = a.offs[row0 + min(lane, rows)];
    const int base = __shfl_sync(0xffffffffu, o, 0);
    const int end = __shfl_sync(0xffffffffu, o, rows);  // lane `rows` holds offs[row0 + rows] (rows == 32: fetched below)
    const int total = (rows == 32 ? a.offs[row0 + 32] : end) - base;
    if (total <= a.warp_cap) {
      if (lane < rows) encode_record<true>(a.p, a.x + (row0 + lane) * a.p.n_out, a.label ? a.label[row0 + lane] : 2, buf + (o - base));
      __syncwarp();
      for (int j = lane; j < total; j += 32) a.bytes[(long long)base + j] = buf[j];
      __syncwarp();
    } else if (lane < rows) {  // cannot happen with warp_cap = 32 maximal records; kept as the safe path
      encode_record<true>(a.p, a.x + (row0 + lane) * a.p.n_out, a.label ? a.label[row0 + lane] : 2, a.bytes + o);
    }
  }
}

}  // namespace

static int fill_enc(const b2_avro_prog* prog, int schema_id, int prefix, EncP& p) {
  B2_REQUIRE(prog->n_fields >= 1 && prog->n_fields <= B2_MAX_FIELDS, "avro encode: n_fields=%d out of range", prog->n_fields);
  B2_REQUIRE(prog->n_out >= 1 && prog->n_out <= B2_MAX_WIDTH, "avro encode: n_out=%d out of range", prog->n_out);
  memset(&p, 0, sizeof(p));
  p.n_fields = prog->n_fields;
  p.n_out = prog->n_out;
  p.prefix = prefix ? 1 : 0;
  p.schema_id = schema_id;
  for (int i = 0; i < prog->n_fields; ++i) {
    B2_REQUIRE(prog->f[i].type >= AV_INT && prog->f[i].type <= AV_NULL, "avro encode: field %d has unknown type %d", i, prog->f[i].type);
    B2_REQUIRE(prog->f[i].out >= -2 && prog->f[i].out < prog->n_out, "avro encode: field %d routed to column %d", i, prog->f[i].out);
    p.f[i] = prog->f[i];
  }
  return B2_OK;
}

extern "C" int b2_avro_encode_sizes(b2_ctx* ctx, const float* x_dev, const uint8_t* label_dev, int64_t n, const b2_avro_prog* prog, int prefix,
                                    int32_t* size_dev, void* stream) {
  B2_REQUIRE(ctx && prog, "b2_avro_encode_sizes: NULL argument");
  B2_REQUIRE(n >= 0, "b2_avro_encode_sizes: n < 0");
  if (n == 0) return B2_OK;
  B2_REQUIRE(x_dev && size_dev, "b2_avro_encode_sizes: NULL device pointer");
  DeviceGuard g(ctx->device);
  EncArgs a;
  memset(&a, 0, sizeof(a));
  int rc = fill_enc(prog, 0, prefix, a.p);
  if (rc) return rc;
  a.x = x_dev;
  a.label = label_dev;
  a.n = n;
  a.size = size_dev;
  const long long blocks = min((long long)ctx->sm_count * 8, (long long)((n + ENC_NT - 1) / ENC_NT));
  avro_sizes_kernel<<<(unsigned)blocks, ENC_NT, 0, (cudaStream_t)stream>>>(a);
  B2_CUDA(cudaGetLastError());
  return B2_OK;
}

extern "C" int b2_avro_encode(b2_ctx* ctx, const float* x_dev, const uint8_t* label_dev, int64_t n, const b2_avro_prog* prog, int prefix,
                              int schema_id, const int32_t* offs_dev, uint8_t* bytes_dev, void* stream) {
  B2_REQUIRE(ctx && prog, "b2_avro_encode: NULL argument");
  B2_REQUIRE(n >= 0, "b2_avro_encode: n < 0");
  if (n == 0) return B2_OK;
  B2_REQUIRE(x_dev && offs_dev && bytes_dev, "b2_avro_encode: NULL device pointer");
  DeviceGuard g(ctx->device);
  EncArgs a;
  memset(&a, 0, sizeof(a));
  int rc = fill_enc(prog, schema_id, prefix, a.p);
  if (rc) return rc;
  a.x = x_dev;
  a.label = label_dev;
  a.n = n;
  a.offs = offs_dev;
  a.bytes = bytes_dev;
  int max_rec = prefix ? 5 : 0;  // upper bound of one record under this program
  for (int i = 0; i < prog->n_fields; ++i) {
    max_rec += prog->f[i].nullable ? 1 : 0;
    switch (prog->f[i].type) {
      case AV_FLOAT: max_rec += 4; break;
      case AV_DOUBLE: max_rec += 8; break;
      case AV_INT: max_rec += 5; break;
      case AV_LONG: max_rec += 10; break;
      case AV_BOOLEAN: max_rec += 1; break;
      case AV_STRING: case AV_BYTES: max_rec += 6; break;
      default: break;
    }
  }
  a.warp_cap = (32 * max_rec + 15) & ~15;
  const size_t smem = (size_t)a.warp_cap * (ENC_NT / 32);
  B2_REQUIRE(smem <= (size_t)ctx->smem_optin, "b2_avro_encode: needs %zu bytes of shared memory", smem);
  B2_CUDA(cudaFuncSetAttribute(avro_encode_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int occ = 0;
  B2_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, avro_encode_kernel, ENC_NT, smem));
  B2_REQUIRE(occ >= 1, "b2_avro_encode: kernel does not fit on an SM");
  const long long need = ((n + 31) / 32 + ENC_NT / 32 - 1) / (ENC_NT / 32);
  const long long blocks = min((long long)ctx->sm_count * occ, need);
  avro_encode_kernel<<<(unsigned)blocks, ENC_NT, smem, (cudaStream_t)stream>>>(a);
  B2_CUDA(cudaGetLastError());
  return B2_OK;
}
