// K9 Kafka record-batch (message format v2, magic 2) framing: broker bytes → the record values the decode kernels take.
//
// Upstream of the path: the reference reads its topic through tensorflow-io's KafkaDataset
// (cardata-v1.py:10), i.e. librdkafka parses fetch responses / log segments into message values.  Layout (Kafka protocol guide,
// "RecordBatch"): baseOffset i64 | batchLength i32 | partitionLeaderEpoch i32 | magic i8 (= 2) | crc u32 (CRC-32C of everything
// after it) | attributes i16 | lastOffsetDelta i32 | firstTimestamp i64 | maxTimestamp i64 | producerId i64 | producerEpoch i16 |
// baseSequence i32 | recordCount i32 | records…, all big-endian; a record = length varint | attributes i8 | timestampDelta
// varlong | offsetDelta varint | keyLength varint | key | valueLength varint | value | headerCount varint | headers.
//
// The caller walks the batch chain on the host (one length field per batch: b2_kafka_scan_batches) and hands over the batch
// positions and the running record count; here one WARP per batch verifies the CRC-32C (coalesced, 32 interleaved lane states
// folded at the end) and walks its records (length chain in lock-step, record fields one record per lane), writing each value's
// [begin, end).  The decode entry points take those ranges as they are (b2_decode_score_spans: the segment is the input, nothing
// is copied); b2_kafka_gather_values packs the values back to back for callers that want a (bytes, offsets) batch.
// Compressed batches (attributes & 7: gzip, Snappy, LZ4, zstd) are inflated first — b2_kafka_inflate_sizes / b2_kafka_inflate, one thread
// per batch running the sequential decoders of inflate.cuh — into a rebuilt, uncompressed copy of the segment that then takes
// the path above.
#include <mutex>

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

// The same, for the index kernel: the four bytes at p arrive as ONE round trip (two aligned words, funnel shift) instead of a chain
// of dependent byte loads; a varint of up to four bytes (every length below 2^27) is taken from the register.  Near the end of
// its container and for longer varints the byte loop decides.
__device__ __forceinline__ bool read_varint_w(const uint8_t* __restrict__ s, long long& p, long long e, long long* v) {
  if (p + 8 <= e) {
    const uintptr_t at = reinterpret_cast<uintptr_t>(s + p);
    const uint32_t* wp = reinterpret_cast<const uint32_t*>(at & ~(uintptr_t)3);
    uint32_t x = __funnelshift_r(wp[0], wp[1], (int)(at & 3u) * 8);
    const uint32_t stop = ~x & 0x80808080u;  // the first byte without a continuation bit ends the varint
    if (stop) {
      const int nb = __ffs((int)stop) >> 3;  // 1..4 bytes
      if (nb < 4) x &= (1u << (8 * nb)) - 1u;
      const uint32_t u = (x & 0x7Fu) | ((x >> 1) & 0x3F80u) | ((x >> 2) & 0x1FC000u) | ((x >> 3) & 0xFE00000u);
      *v = (long long)(u >> 1) ^ -(long long)(u & 1u);
      p += nb;
      return true;
    }
  }
  return read_varint(s, p, e, v);
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

// CRC-32C tables: tab[k] = the slicing-by-4 set (advance the state over one byte / four bytes); far[k] = the same state advanced
// over 128 zero bytes (a 32 x 32 matrix over GF(2) laid out as four 256-entry tables, one per state byte).  Built once on the
// host, uploaded once per device.
struct CrcTabs {
  uint32_t tab[4][256];
  uint32_t far[4][256];
};
__device__ CrcTabs g_crc_tabs;

const CrcTabs& host_crc_tabs() {
  static CrcTabs t;
  static bool built = false;
  if (!built) {
    for (uint32_t i = 0; i < 256; ++i) {
      uint32_t c = i;
      for (int k = 0; k < 8; ++k) c = (c >> 1) ^ ((c & 1u) ? 0x82F63B78u : 0u);
      t.tab[0][i] = c;
    }
    for (uint32_t i = 0; i < 256; ++i) {
      uint32_t c = t.tab[0][i];
      for (int k = 1; k < 4; ++k) {
        c = t.tab[0][c & 0xFFu] ^ (c >> 8);
        t.tab[k][i] = c;
      }
    }
    for (int k = 0; k < 4; ++k)
      for (uint32_t i = 0; i < 256; ++i) {
        uint32_t c = i << (8 * k);
        for (int z = 0; z < 128; ++z) c = t.tab[0][c & 0xFFu] ^ (c >> 8);
        t.far[k][i] = c;
      }
    built = true;
  }
  return t;
}

std::mutex g_crc_mutex;
bool g_crc_uploaded[64];
cudaError_t upload_crc_tabs(int device, cudaStream_t stream) {
  std::lock_guard<std::mutex> lock(g_crc_mutex);
  if (device >= 0 && device < 64 && g_crc_uploaded[device]) return cudaSuccess;
  const CrcTabs& t = host_crc_tabs();
  // stream-ordered in front of the first kernel that reads it; the host copy is static, so the async copy's source stays valid
  cudaError_t e = cudaMemcpyToSymbolAsync(g_crc_tabs, &t, sizeof(CrcTabs), 0, cudaMemcpyHostToDevice, stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(stream);
  if (e == cudaSuccess && device >= 0 && device < 64) g_crc_uploaded[device] = true;
  return e;
}

constexpr int KF_WARPS = 8;        // batches in flight per CTA: one warp each
constexpr int KF_TILE_ROWS = 32;   // CRC rows (of 128 bytes) between two stretches of the record walk: 4 KB stays in L1

// One WARP per record batch.
//  * CRC-32C: the checked bytes are taken as rows of 32 little-endian words, lane l owning word l of every row — coalesced loads —
//    with the lane's running state advanced over the 128 bytes of a row by the `far` tables (c = far(c) ^ word: the four lookups
//    of slicing-by-4, different tables).  The stream is padded IN FRONT with zero bytes to a whole number of rows (leading zeros
//    leave a zero state alone), the initial 0xFFFFFFFF becomes an XOR onto the first four real bytes, and the 32 lane states are
//    folded by one more slicing pass over them.
//  * Record walk: the chain of record lengths is followed by all lanes in lock-step (same address in every lane: one broadcast
//    load), 32 records at a time, lane k keeping record k's bounds; then the lanes read their own record's varints in parallel and
//    store 32 value ranges coalesced.  The walk runs in stretches right behind the CRC rows of the same 4 KB, so that it finds
//    its bytes in L1.
__global__ void __launch_bounds__(KF_WARPS * 32) kafka_index_kernel(const FrameArgs a) {
  __shared__ CrcTabs t;
  {
    const uint32_t* src = reinterpret_cast<const uint32_t*>(&g_crc_tabs);
    uint32_t* dst = reinterpret_cast<uint32_t*>(&t);
    for (int i = threadIdx.x; i < (int)(sizeof(CrcTabs) / 4); i += KF_WARPS * 32) dst[i] = src[i];
  }
  __syncthreads();
  constexpr unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const long long n_warps = (long long)gridDim.x * KF_WARPS;
  const uint8_t* __restrict__ s = a.bytes;
  for (long long b = (long long)blockIdx.x * KF_WARPS + (threadIdx.x >> 5); b < a.nb; b += n_warps) {
    const long long pos = a.batch_pos[b];
    int st = B2_FRAME_OK;
    long long n_rec = 0, endb = 0;
    bool control = false;
    if (pos < 0 || pos + KF_HEADER > a.n_bytes) {
      st = B2_FRAME_TRUNCATED;
    } else {
      const long long blen = (long long)(int32_t)be32(s + pos + 8);
      endb = pos + 12 + blen;
      n_rec = (long long)(int32_t)be32(s + pos + 57);
      if (s[pos + 16] != 2) st = B2_FRAME_BAD_MAGIC;
      else if (blen < KF_HEADER - 12 || endb > a.n_bytes || n_rec < 0) st = B2_FRAME_TRUNCATED;
      else if ((s[pos + 22] & 7u) != 0) st = B2_FRAME_COMPRESSED;  // low 3 bits of attributes = codec
      else control = (s[pos + 22] & 0x20u) != 0;                   // attributes bit 5: transaction markers, never delivered
    }
    const long long r0 = a.rec_base[b];
    if (n_rec < 0) n_rec = 0;
    if (st == B2_FRAME_OK) {
      // the checked stream [S, endb), n bytes, padded in front to whole rows
      const long long S = pos + 21, n = endb - S;
      const int pad = (int)((128 - (n & 127)) & 127);
      const long long rows = (n + pad) >> 7;
      const int sh = (int)((reinterpret_cast<uintptr_t>(s) + (uintptr_t)(S - pad + 4ll * lane)) & 3u) * 8;  // same in every lane
      uint32_t c = 0;
      long long p = pos + KF_HEADER, r = 0;  // record walk: next record's first byte, records done
      bool bad = false, guess = true;
      const bool walk = !control;
      for (long long row0 = 0; row0 < rows; row0 += KF_TILE_ROWS) {
        const long long row1 = min(rows, row0 + KF_TILE_ROWS);
        // the padding, the initial value (rows 0 and 1) and the stream's last bytes (its last row) byte by byte; every other row
        // with aligned word loads, eight rows' loads in flight in front of the eight dependent table steps
        auto slow_word = [&](long long row) -> uint32_t {
          const long long o = (row << 7) + 4 * lane - pad;  // real offset of this lane's word
          uint32_t w = 0;
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const long long ro = o + j;
            if (ro >= 0 && ro < n) w |= ((uint32_t)s[S + ro] ^ (ro < 4 ? 0xFFu : 0u)) << (8 * j);
          }
          return w;
        };
        auto fast_word = [&](long long row) -> uint32_t {
          const long long o = (row << 7) + 4 * lane - pad;
          const uint32_t* wp = reinterpret_cast<const uint32_t*>(reinterpret_cast<uintptr_t>(s + S + o) & ~(uintptr_t)3);
          uint32_t w = wp[0];
          if (sh) w = __funnelshift_r(w, wp[1], sh);
          return w;
        };
        auto step = [&](uint32_t w) { c = t.far[0][c & 0xFFu] ^ t.far[1][(c >> 8) & 0xFFu] ^ t.far[2][(c >> 16) & 0xFFu] ^ t.far[3][c >> 24] ^ w; };
        long long row = row0;
        for (; row < row1 && row < 2; ++row) step(slow_word(row));
        const long long fast_end = min(row1, rows - 1);
        for (; row + 8 <= fast_end; row += 8) {
          uint32_t w[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) w[j] = fast_word(row + j);
#pragma unroll
          for (int j = 0; j < 8; ++j) step(w[j]);
        }
        for (; row < fast_end; ++row) step(fast_word(row));
        if (row < row1) step(slow_word(row));  // the stream's last row
        if (!walk || bad) continue;
        const long long g_hi = row1 == rows ? endb : S + (row1 << 7) - pad;  // bytes below g_hi have been read
        while (r < n_rec && p < g_hi && !bad) {
          // every length is untrusted: compare it with the bytes that are left instead of adding it to the cursor (a crafted
          // varint near 2^63 would wrap the sum negative and slip past a `>` test)
          long long ms = 0, me = 0;
          int cnt = 0;
          const int want = (int)min(32ll, n_rec - r);
          if (guess) {
            // Sensor records mostly have ONE length: take the stride from the record at p, let lane l read the length at
            // p + l * stride; the leading lanes whose record has exactly that length prove the next lane's start — 32 record
            // starts for the price of two, however long the chain.
            long long q0 = p, len0;
            if (!read_varint_w(s, q0, endb, &len0) || len0 < 0 || len0 > endb - q0) {
              bad = true;
              break;
            }
            const long long stride = q0 - p + len0;  // >= 1
            long long tot = -1;
            if (lane < want && (lane == 0 || stride <= (endb - p - 1) / lane)) {  // the guessed start lies inside the batch
              long long q = p + lane * stride, len;
              const long long at = q;
              if (read_varint_w(s, q, endb, &len) && len >= 0 && len <= endb - q) ms = q, me = q + len, tot = me - at;
            }
            const unsigned eq = __ballot_sync(FULL, tot == stride);  // lane 0 is always in
            const int m = eq == FULL ? 32 : __ffs((int)~eq) - 1;      // lanes 0 .. m-1 have the stride: lanes 0 .. m start a record
            const long long tot_m = __shfl_sync(FULL, tot, m & 31);
            cnt = m;
            if (m < want) {
              if (tot_m >= 0) cnt = m + 1;  // the first record of another length: known too
              else bad = true;              // a record that must exist cannot be read
            }
            p = __shfl_sync(FULL, me, cnt - 1);  // cnt >= 1: end of the last record taken
            guess = cnt >= 4;                    // lengths vary (text values): follow the chain one by one for a round
          } else {
            // the chain one record at a time, in 32-bit offsets from the batch's first byte: per record one round trip for the
            // four bytes at the cursor and a dozen dependent integer instructions
            const uint8_t* sb = s + pos;
            const uint32_t endr = (uint32_t)(endb - pos);
            uint32_t pp = (uint32_t)(p - pos), ms32 = 0, me32 = 0;
            for (int k = 0; k < want; ++k) {
              uint32_t nb = 0, len = 0;
              if (pp + 8u <= endr) {
                const uintptr_t at = reinterpret_cast<uintptr_t>(sb + pp);
                const uint32_t* wp = reinterpret_cast<const uint32_t*>(at & ~(uintptr_t)3);
                uint32_t x = __funnelshift_r(wp[0], wp[1], (int)(at & 3u) * 8);
                const uint32_t stop = ~x & 0x80808080u;
                if (stop) {
                  nb = (uint32_t)__ffs((int)stop) >> 3;
                  x &= 0xFFFFFFFFu >> (32u - 8u * nb);
                  const uint32_t u = (x & 0x7Fu) | ((x >> 1) & 0x3F80u) | ((x >> 2) & 0x1FC000u) | ((x >> 3) & 0xFE00000u);
                  if (u & 1u) {  // a negative length
                    bad = true;
                    break;
                  }
                  len = u >> 1;
                }
              }
              if (nb == 0) {  // the batch's last bytes, or a length of 2^27 and more: the byte loop decides
                long long q = pos + pp, l64;
                if (!read_varint(s, q, endb, &l64) || l64 < 0 || l64 > endb - q) {
                  bad = true;
                  break;
                }
                nb = (uint32_t)(q - pos) - pp;
                len = (uint32_t)l64;
              } else if (len > endr - pp - nb) {
                bad = true;
                break;
              }
              if (k == lane) ms32 = pp + nb, me32 = pp + nb + len;
              pp += nb + len;
              ++cnt;
            }
            p = pos + pp;
            ms = pos + ms32;
            me = pos + me32;
            guess = true;
          }
          int32_t vb = 0, ve = 0;
          bool ok = true;
          if (lane < cnt) {
            bool done = false;
            if (me - ms >= 20) {
              // the common record — attributes, three short varints, a null key, the value length — sits in its first 16 bytes:
              // one round trip (five aligned words), then the varints come out of registers
              const uintptr_t at = reinterpret_cast<uintptr_t>(s + ms);
              const uint32_t* wp = reinterpret_cast<const uint32_t*>(at & ~(uintptr_t)3);
              const int sh0 = (int)(at & 3u) * 8;
              const uint32_t a0 = wp[0], a1 = wp[1], a2 = wp[2], a3 = wp[3], a4 = wp[4];
              uint32_t w0 = __funnelshift_r(a0, a1, sh0), w1 = __funnelshift_r(a1, a2, sh0), w2 = __funnelshift_r(a2, a3, sh0),
                       w3 = __funnelshift_r(a3, a4, sh0);
              int used = 0;
              bool short4 = true;
              auto drop = [&](int nbytes) {  // the window moves on by 1..4 bytes
                const int b = nbytes * 8;
                w0 = __funnelshift_rc(w0, w1, b), w1 = __funnelshift_rc(w1, w2, b), w2 = __funnelshift_rc(w2, w3, b), w3 = __funnelshift_rc(w3, 0u, b);
                used += nbytes;
              };
              auto take = [&]() -> int {  // zig-zag varint of up to four bytes at the window's front
                const uint32_t stop = ~w0 & 0x80808080u;
                if (!stop) {
                  short4 = false;
                  return 0;
                }
                const int nb = __ffs((int)stop) >> 3;
                const uint32_t x = nb < 4 ? w0 & ((1u << (8 * nb)) - 1u) : w0;
                const uint32_t u = (x & 0x7Fu) | ((x >> 1) & 0x3F80u) | ((x >> 2) & 0x1FC000u) | ((x >> 3) & 0xFE00000u);
                drop(nb);
                return (int)(u >> 1) ^ -(int)(u & 1u);
              };
              drop(1);  // attributes
              take();   // timestampDelta
              take();   // offsetDelta
              const int klen = take();
              if (short4 && klen <= 0 && klen >= -1) {
                const int vlen = take();
                if (short4 && used <= 16) {  // everything read lay inside the 16 bytes
                  const long long q = ms + used;
                  ok = vlen >= -1 && vlen <= me - q;
                  vb = (int32_t)q;
                  ve = (int32_t)(q + (vlen > 0 ? vlen : 0));
                  done = true;
                }
              }
            }
            if (!done) {  // keys, long varints, records of a few bytes: field by field
              long long q = ms + 1, v, klen, vlen;  // attributes
              ok = q <= me && read_varint_w(s, q, me, &v) && read_varint_w(s, q, me, &v) && read_varint_w(s, q, me, &klen) && klen >= -1 &&
                   klen <= me - q;
              if (ok) {
                if (klen > 0) q += klen;
                ok = read_varint_w(s, q, me, &vlen) && vlen >= -1 && vlen <= me - q;
                vb = (int32_t)q;                               // a null value (length -1) is an empty record
                ve = (int32_t)(q + (vlen > 0 ? vlen : 0));     // headers are skipped with the record length
              }
            }
          }
          if (__any_sync(FULL, !ok)) bad = true;
          if (lane < cnt && !bad) {
            a.begin[r0 + r + lane] = vb;
            a.end[r0 + r + lane] = ve;
          }
          r += cnt;
        }
      }
      if (a.verify_crc) {
        // fold the lane states: lane l's words sit 4 (31 - l) bytes in front of the stream's end
        uint32_t f = 0;
        for (int l = 0; l < 32; ++l) {
          f ^= __shfl_sync(FULL, c, l);
          f = t.tab[3][f & 0xFFu] ^ t.tab[2][(f >> 8) & 0xFFu] ^ t.tab[1][(f >> 16) & 0xFFu] ^ t.tab[0][f >> 24];
        }
        if ((f ^ 0xFFFFFFFFu) != be32(s + pos + 17)) st = B2_FRAME_BAD_CRC;
      }
      if (st == B2_FRAME_OK) {
        if (control) st = B2_FRAME_CONTROL;
        else if (bad || r < n_rec) st = B2_FRAME_TRUNCATED;
      }
    }
    // a refused batch yields an EMPTY range for each of the records its header announces (the caller sized begin / end from
    // those same counts) — also when the record walk itself ran into a truncation: ranges written before are withdrawn.  The
    // empty range sits at the batch's position, so that the ranges of a segment stay in ascending order.
    if (st != B2_FRAME_OK) {
      const int32_t at = (int32_t)max(0ll, min(pos, a.n_bytes));
      for (long long i = lane; i < n_rec; i += 32) a.begin[r0 + i] = a.end[r0 + i] = at;
    }
    if (lane == 0) a.status[b] = st;
    __syncwarp();
  }
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
        if (r == B2_INFL_UNSUPPORTED) st = B2_FRAME_COMPRESSED;   // a codec number beyond zstd, a zstd dictionary
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
  B2_CUDA(upload_crc_tabs(ctx->device, (cudaStream_t)stream));
  long long grid = (n_batches + KF_WARPS - 1) / KF_WARPS;
  if (grid > (long long)ctx->sm_count * 4) grid = (long long)ctx->sm_count * 4;
  kafka_index_kernel<<<(unsigned)grid, KF_WARPS * 32, 0, (cudaStream_t)stream>>>(a);
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

// Host walk of the batch chain of a log segment / fetch response held in HOST memory: batch b starts at batch_pos[b] and announces
// rec_count[b] records (clamped at 0).  Stops at the first header that does not fit or announces a batch running past the end (a
// fetch response may end in a partial batch; consumers drop it).  Returns the number of batches found; only the first `cap` are
// written, so a first call with cap = 0 sizes the arrays.
extern "C" int64_t b2_kafka_scan_batches(const uint8_t* bytes_host, int64_t n_bytes, int64_t* batch_pos, int64_t* rec_count, int64_t cap) {
  if (!bytes_host || n_bytes < 0 || (cap > 0 && (!batch_pos || !rec_count))) return -1;
  auto be32h = [&](int64_t at) -> int64_t {
    return (int64_t)(int32_t)(((uint32_t)bytes_host[at] << 24) | ((uint32_t)bytes_host[at + 1] << 16) | ((uint32_t)bytes_host[at + 2] << 8) |
                              (uint32_t)bytes_host[at + 3]);
  };
  int64_t pos = 0, nb = 0;
  while (pos + KF_HEADER <= n_bytes) {
    const int64_t blen = be32h(pos + 8);
    if (blen < KF_HEADER - 12 || blen > n_bytes - pos - 12) break;
    if (nb < cap) {
      const int64_t cnt = be32h(pos + 57);
      batch_pos[nb] = pos;
      rec_count[nb] = cnt > 0 ? cnt : 0;
    }
    ++nb;
    pos += 12 + blen;
  }
  return nb;
}
