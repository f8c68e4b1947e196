// K9 Kafka record-batch (message format v2, magic 2) framing: broker bytes → the record values the decode kernels take.
//
// Upstream of the path: the reference reads its topic through tensorflow-io's KafkaDataset
// (cardata-v1.py:10), i.e. librdkafka parses fetch responses / log segments into message values.  Layout (Kafka protocol guide,
// "RecordBatch"): baseOffset i64 | batchLength i32 | partitionLeaderEpoch i32 | magic i8 (= 2) | crc u32 (CRC-32C of everything
// after it) | attributes i16 | lastOffsetDelta i32 | firstTimestamp i64 | maxTimestamp i64 | producerId i64 | producerEpoch i16 |
// baseSequence i32 | recordCount i32 | records…, all big-endian; a record = length varint | attributes i8 | timestampDelta
// varlong | offsetDelta varint | keyLength varint | key | valueLength varint | value | headerCount varint | headers.
//
// The caller walks the batch chain on the host (one length field per batch) and hands over the batch positions and the running
// record count; here one thread per batch verifies the CRC-32C (slicing-by-4 tables in shared memory) and walks its records,
// writing each value's [begin, end) — then a warp-per-32-records gather packs the values back to back so that the result is the
// (bytes, offsets) batch every decode entry point consumes.
// Compressed batches (attributes & 7: gzip, Snappy, LZ4) are inflated first — b2_kafka_inflate_sizes / b2_kafka_inflate, one thread
// per batch running the sequential decoders of inflate.cuh — into a rebuilt, uncompressed copy of the segment that then takes
// the path above; zstd batches are reported (B2_FRAME_COMPRESSED) and skipped.
#include "common.cuh"
#include "inflate.cuh"

namespace {

constexpr int KF_NT = 128;
constexpr int KF_HEADER = 61;  // bytes in front of the first record

__device__ __forceinline__ uint32_t be32(const uint8_t* p) { return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3]; }

__device__ __forceinline__ bool read_varint(const uint8_t* __restrict__ s, long long& p, long long e, long long* v) {
  unsigned long long u = 0;
  int shift = 0;
  for (int k = 0; k < 10; ++k) {
    if (p >= e) return false;
    const uint32_t b = s[p++];
    u |= (unsigned long long)(b & 0x7Fu) << shift;
    if (b < 0x80u) {
      *v = (long long)(u >> 1) ^ -(long long)(u & 1ull);
      return true;
    }
    shift += 7;
  }
  return false;
}

struct FrameArgs {
  const uint8_t* bytes;
  long long n_bytes;
  const long long* batch_pos;  // [nb] byte offset of each batch
  const long long* rec_base;   // [nb] records in the batches before this one
  long long nb;
  int verify_crc;
  int32_t* begin;              // [n] value byte ranges
  int32_t* end;
  int32_t* status;             // [nb]
};

__global__ void __launch_bounds__(KF_NT) kafka_index_kernel(const FrameArgs a) {
  __shared__ uint32_t tab[4][256];
  for (int i = threadIdx.x; i < 256; i += KF_NT) {  // CRC-32C (Castagnoli, reflected 0x82F63B78), slicing-by-4
    uint32_t c = (uint32_t)i;
    for (int k = 0; k < 8; ++k) c = (c >> 1) ^ ((c & 1u) ? 0x82F63B78u : 0u);
    tab[0][i] = c;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < 256; i += KF_NT) {
    uint32_t c = tab[0][i];
    for (int t = 1; t < 4; ++t) {
      c = tab[0][c & 0xFFu] ^ (c >> 8);
      tab[t][i] = c;
    }
  }
  __syncthreads();
  const long long b = (long long)blockIdx.x * KF_NT + threadIdx.x;
  if (b >= a.nb) return;
  const long long pos = a.batch_pos[b];
  const uint8_t* s = a.bytes;
  int st = B2_FRAME_OK;
  long long n_rec = 0, endb = 0;
  if (pos < 0 || pos + KF_HEADER > a.n_bytes) {
    st = B2_FRAME_TRUNCATED;
  } else {
    const long long blen = (long long)(int32_t)be32(s + pos + 8);
    endb = pos + 12 + blen;
    n_rec = (long long)(int32_t)be32(s + pos + 57);
    if (s[pos + 16] != 2) st = B2_FRAME_BAD_MAGIC;
    else if (blen < KF_HEADER - 12 || endb > a.n_bytes || n_rec < 0) st = B2_FRAME_TRUNCATED;
    else if ((s[pos + 22] & 7u) != 0) st = B2_FRAME_COMPRESSED;  // low 3 bits of attributes = codec
  }
  if (st == B2_FRAME_OK && a.verify_crc) {
    uint32_t c = 0xFFFFFFFFu;
    long long p = pos + 21;
    for (; p + 4 <= endb; p += 4) {
      c ^= (uint32_t)s[p] | ((uint32_t)s[p + 1] << 8) | ((uint32_t)s[p + 2] << 16) | ((uint32_t)s[p + 3] << 24);
      c = tab[3][c & 0xFFu] ^ tab[2][(c >> 8) & 0xFFu] ^ tab[1][(c >> 16) & 0xFFu] ^ tab[0][c >> 24];
    }
    for (; p < endb; ++p) c = tab[0][(c ^ s[p]) & 0xFFu] ^ (c >> 8);
    if ((c ^ 0xFFFFFFFFu) != be32(s + pos + 17)) st = B2_FRAME_BAD_CRC;
  }
  const long long r0 = a.rec_base[b];
  if (st == B2_FRAME_OK && (s[pos + 22] & 0x20u)) st = B2_FRAME_CONTROL;  // attributes bit 5: transaction markers, never delivered
  if (st == B2_FRAME_OK) {
    long long p = pos + KF_HEADER;
    for (long long r = 0; r < n_rec; ++r) {
      // every length is untrusted: compare it with the bytes that are left instead of adding it to the cursor (a crafted
      // varint near 2^63 would wrap the sum negative and slip past a `>` test)
      long long len, v, klen, vlen;
      if (!read_varint(s, p, endb, &len) || len < 0 || len > endb - p) {
        st = B2_FRAME_TRUNCATED;
        break;
      }
      const long long rec_end = p + len;
      p += 1;                                                     // attributes
      if (p > rec_end || !read_varint(s, p, rec_end, &v) || !read_varint(s, p, rec_end, &v) || !read_varint(s, p, rec_end, &klen) ||
          klen < -1 || klen > rec_end - p) {
        st = B2_FRAME_TRUNCATED;
        break;
      }
      if (klen > 0) p += klen;
      if (!read_varint(s, p, rec_end, &vlen) || vlen < -1 || vlen > rec_end - p) {
        st = B2_FRAME_TRUNCATED;
        break;
      }
      a.begin[r0 + r] = (int32_t)p;                               // a null value (length -1) is an empty record
      a.end[r0 + r] = (int32_t)(p + (vlen > 0 ? vlen : 0));
      p = rec_end;                                                // headers are skipped with the record length
    }
  }
  // a refused batch yields the empty range [0, 0) for each of the records its header announces (the caller sized begin / end
  // from those same counts) — also when the record walk itself ran into a truncation: ranges written before the break are withdrawn
  if (st != B2_FRAME_OK)
    for (long long r = 0; r < n_rec; ++r) a.begin[r0 + r] = a.end[r0 + r] = 0;
  a.status[b] = st;
}

// values [begin[i], end[i]) → packed at out_offs[i] (the caller scanned the lengths); one warp per 32 records
__global__ void __launch_bounds__(256) kafka_gather_kernel(const uint8_t* __restrict__ bytes, const int32_t* __restrict__ begin,
                                                           const int32_t* __restrict__ end, long long n, const int32_t* __restrict__ out_offs,
                                                           uint8_t* __restrict__ out) {
  const int lane = threadIdx.x & 31;
  const long long warp = ((long long)blockIdx.x * 256 + threadIdx.x) >> 5, n_warps = ((long long)gridDim.x * 256) >> 5;
  for (long long c = warp; c * 32 < n; c += n_warps) {
    const long long i = c * 32 + lane;
    const int b = i < n ? begin[i] : 0, len = i < n ? end[i] - b : 0, o = i < n ? out_offs[i] : 0;
    for (int r = 0; r < 32; ++r) {
      const int rb = __shfl_sync(0xffffffffu, b, r), rl = __shfl_sync(0xffffffffu, len, r), ro = __shfl_sync(0xffffffffu, o, r);
      for (int j = lane; j < rl; j += 32) out[(long long)ro + j] = bytes[(long long)rb + j];
    }
  }
}

// CRC-32C tables (slicing-by-4) in shared memory, shared by the index and the inflate-size kernels
__device__ __forceinline__ void crc_tables(uint32_t (*tab)[256], int nt) {
  for (int i = threadIdx.x; i < 256; i += nt) {
    uint32_t c = (uint32_t)i;
    for (int k = 0; k < 8; ++k) c = (c >> 1) ^ ((c & 1u) ? 0x82F63B78u : 0u);
    tab[0][i] = c;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < 256; i += nt) {
    uint32_t c = tab[0][i];
    for (int t = 1; t < 4; ++t) {
      c = tab[0][c & 0xFFu] ^ (c >> 8);
      tab[t][i] = c;
    }
  }
  __syncthreads();
}
__device__ __forceinline__ uint32_t crc32c_dev(const uint32_t (*tab)[256], const uint8_t* s, long long p, long long e) {
  uint32_t c = 0xFFFFFFFFu;
  for (; p + 4 <= e; p += 4) {
    c ^= (uint32_t)s[p] | ((uint32_t)s[p + 1] << 8) | ((uint32_t)s[p + 2] << 16) | ((uint32_t)s[p + 3] << 24);
    c = tab[3][c & 0xFFu] ^ tab[2][(c >> 8) & 0xFFu] ^ tab[1][(c >> 16) & 0xFFu] ^ tab[0][c >> 24];
  }
  for (; p < e; ++p) c = tab[0][(c ^ s[p]) & 0xFFu] ^ (c >> 8);
  return c ^ 0xFFFFFFFFu;
}

struct InflateArgs {
  const uint8_t* bytes;
  long long n_bytes;
  const long long* batch_pos;
  long long nb;
  int verify_crc;
  long long* sizes;          // [nb] bytes of the records section once inflated (= as stored for an uncompressed batch); 0 when refused
  int32_t* status;           // [nb] B2_FRAME_* of the first pass
  const long long* out_pos;  // [nb] (second pass) where each rebuilt batch starts in `out`
  uint8_t* out;
  long long out_bytes;
};

// first pass: header checks, CRC-32C over the bytes as stored, size of the inflated records section
__global__ void __launch_bounds__(KF_NT) kafka_inflate_sizes_kernel(const InflateArgs a) {
  __shared__ uint32_t tab[4][256];
  crc_tables(tab, KF_NT);
  const long long b = (long long)blockIdx.x * KF_NT + threadIdx.x;
  if (b >= a.nb) return;
  const long long pos = a.batch_pos[b];
  const uint8_t* s = a.bytes;
  int st = B2_FRAME_OK;
  long long size = 0;
  if (pos < 0 || pos + KF_HEADER > a.n_bytes) {
    st = B2_FRAME_TRUNCATED;
  } else {
    const long long blen = (long long)(int32_t)be32(s + pos + 8), endb = pos + 12 + blen;
    if (s[pos + 16] != 2) st = B2_FRAME_BAD_MAGIC;
    else if (blen < KF_HEADER - 12 || endb > a.n_bytes) st = B2_FRAME_TRUNCATED;
    else if (a.verify_crc && crc32c_dev(tab, s, pos + 21, endb) != be32(s + pos + 17)) st = B2_FRAME_BAD_CRC;
    else {
      const int codec = s[pos + 22] & 7;
      if (codec == 0) {
        size = endb - (pos + KF_HEADER);
      } else {
        const long long r = b2_inflate_codec(codec, s + pos + KF_HEADER, endb - (pos + KF_HEADER), nullptr, 0);
        if (r == B2_INFL_UNSUPPORTED) st = B2_FRAME_COMPRESSED;   // zstd
        else if (r < 0) st = B2_FRAME_TRUNCATED;                  // the compressed stream is damaged
        else size = r;
      }
    }
  }
  a.sizes[b] = st == B2_FRAME_OK ? size : 0;
  a.status[b] = st;
}

// second pass: rebuilt batch = the 61 header bytes (codec bits cleared, batchLength adjusted) + the inflated records; a batch
// the first pass refused becomes an empty shell that announces its records but is marked by status (the caller blanks them)
__global__ void __launch_bounds__(KF_NT) kafka_inflate_kernel(const InflateArgs a) {
  const long long b = (long long)blockIdx.x * KF_NT + threadIdx.x;
  if (b >= a.nb) return;
  const long long pos = a.batch_pos[b], op = a.out_pos[b], size = a.sizes[b];
  const uint8_t* s = a.bytes;
  uint8_t* o = a.out + op;
  if (pos < 0 || pos + KF_HEADER > a.n_bytes || op + KF_HEADER + size > a.out_bytes) return;
  for (int i = 0; i < KF_HEADER; ++i) o[i] = s[pos + i];
  const long long blen = KF_HEADER - 12 + size;
  o[8] = (uint8_t)(blen >> 24), o[9] = (uint8_t)(blen >> 16), o[10] = (uint8_t)(blen >> 8), o[11] = (uint8_t)blen;
  if (a.status[b] != B2_FRAME_OK) return;
  const int codec = s[pos + 22] & 7;
  o[22] = (uint8_t)(s[pos + 22] & ~7);
  const long long src_len = (long long)(int32_t)be32(s + pos + 8) + 12 - KF_HEADER;
  if (codec == 0) {
    for (long long i = 0; i < size; ++i) o[KF_HEADER + i] = s[pos + KF_HEADER + i];
  } else {
    const long long r = b2_inflate_codec(codec, s + pos + KF_HEADER, src_len, o + KF_HEADER, size);
    if (r != size) a.status[b] = B2_FRAME_TRUNCATED;
  }
}

}  // namespace

extern "C" int b2_kafka_inflate_sizes(b2_ctx* ctx, const uint8_t* bytes_dev, int64_t n_bytes, const int64_t* batch_pos_dev, int64_t n_batches,
                                      int verify_crc, int64_t* sizes_dev, int32_t* status_dev, void* stream) {
  B2_REQUIRE(ctx, "b2_kafka_inflate_sizes: NULL context");
  B2_REQUIRE(n_batches >= 0 && n_bytes >= 0, "b2_kafka_inflate_sizes: n_batches=%lld n_bytes=%lld", (long long)n_batches, (long long)n_bytes);
  if (n_batches == 0) return B2_OK;
  B2_REQUIRE(bytes_dev && batch_pos_dev && sizes_dev && status_dev, "b2_kafka_inflate_sizes: NULL device pointer");
  DeviceGuard g(ctx->device);
  InflateArgs a;
  memset(&a, 0, sizeof(a));
  a.bytes = bytes_dev, a.n_bytes = n_bytes;
  a.batch_pos = reinterpret_cast<const long long*>(batch_pos_dev);
  a.nb = n_batches, a.verify_crc = verify_crc;
  a.sizes = reinterpret_cast<long long*>(sizes_dev), a.status = status_dev;
  kafka_inflate_sizes_kernel<<<(unsigned)((n_batches + KF_NT - 1) / KF_NT), KF_NT, 0, (cudaStream_t)stream>>>(a);
  B2_CUDA(cudaGetLastError());
  return B2_OK;
}

extern "C" int b2_kafka_inflate(b2_ctx* ctx, const uint8_t* bytes_dev, int64_t n_bytes, const int64_t* batch_pos_dev, int64_t n_batches,
                                const int64_t* sizes_dev, const int64_t* out_pos_dev, int32_t* status_dev, uint8_t* out_dev, int64_t out_bytes,
                                void* stream) {
  B2_REQUIRE(ctx, "b2_kafka_inflate: NULL context");
  B2_REQUIRE(n_batches >= 0 && n_bytes >= 0 && out_bytes >= 0, "b2_kafka_inflate: negative size");
  if (n_batches == 0) return B2_OK;
  B2_REQUIRE(bytes_dev && batch_pos_dev && sizes_dev && out_pos_dev && status_dev && out_dev, "b2_kafka_inflate: NULL device pointer");
  DeviceGuard g(ctx->device);
  InflateArgs a;
  memset(&a, 0, sizeof(a));
  a.bytes = bytes_dev, a.n_bytes = n_bytes;
  a.batch_pos = reinterpret_cast<const long long*>(batch_pos_dev);
  a.nb = n_batches;
  a.sizes = const_cast<long long*>(reinterpret_cast<const long long*>(sizes_dev));
  a.status = status_dev;
  a.out_pos = reinterpret_cast<const long long*>(out_pos_dev);
  a.out = out_dev, a.out_bytes = out_bytes;
  kafka_inflate_kernel<<<(unsigned)((n_batches + KF_NT - 1) / KF_NT), KF_NT, 0, (cudaStream_t)stream>>>(a);
  B2_CUDA(cudaGetLastError());
  return B2_OK;
}

extern "C" int b2_kafka_index_records(b2_ctx* ctx, const uint8_t* bytes_dev, int64_t n_bytes, const int64_t* batch_pos_dev,
                                      const int64_t* rec_base_dev, int64_t n_batches, int verify_crc, int32_t* begin_dev, int32_t* end_dev,
                                      int32_t* status_dev, void* stream) {
  B2_REQUIRE(ctx, "b2_kafka_index_records: NULL context");
  B2_REQUIRE(n_batches >= 0 && n_bytes >= 0 && n_bytes < (1ll << 31), "b2_kafka_index_records: n_batches=%lld n_bytes=%lld", (long long)n_batches,
             (long long)n_bytes);
  if (n_batches == 0) return B2_OK;
  B2_REQUIRE(bytes_dev && batch_pos_dev && rec_base_dev && begin_dev && end_dev && status_dev, "b2_kafka_index_records: NULL device pointer");
  DeviceGuard g(ctx->device);
  FrameArgs a;
  a.bytes = bytes_dev;
  a.n_bytes = n_bytes;
  a.batch_pos = reinterpret_cast<const long long*>(batch_pos_dev);
  a.rec_base = reinterpret_cast<const long long*>(rec_base_dev);
  a.nb = n_batches;
  a.verify_crc = verify_crc;
  a.begin = begin_dev;
  a.end = end_dev;
  a.status = status_dev;
  kafka_index_kernel<<<(unsigned)((n_batches + KF_NT - 1) / KF_NT), KF_NT, 0, (cudaStream_t)stream>>>(a);
  B2_CUDA(cudaGetLastError());
  return B2_OK;
}

extern "C" int b2_kafka_gather_values(b2_ctx* ctx, const uint8_t* bytes_dev, const int32_t* begin_dev, const int32_t* end_dev, int64_t n,
                                      const int32_t* out_offs_dev, uint8_t* out_bytes_dev, void* stream) {
  B2_REQUIRE(ctx, "b2_kafka_gather_values: NULL context");
  B2_REQUIRE(n >= 0, "b2_kafka_gather_values: n < 0");
  if (n == 0) return B2_OK;
  B2_REQUIRE(bytes_dev && begin_dev && end_dev && out_offs_dev && out_bytes_dev, "b2_kafka_gather_values: NULL device pointer");
  DeviceGuard g(ctx->device);
  const long long warps = (n + 31) / 32;
  const long long blocks = min((long long)ctx->sm_count * 8, (warps + 7) / 8);
  kafka_gather_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(bytes_dev, begin_dev, end_dev, n, out_offs_dev, out_bytes_dev);
  B2_CUDA(cudaGetLastError());
  return B2_OK;
}
