// Decompressors for compressed Kafka record batches (message format v2: the records section of a batch is compressed as a whole
// with the codec named by attributes bits 0-2): Snappy (raw block and the xerial "SnappyOutputStream" framing the Java client
// writes), LZ4 (frame format) and gzip (RFC 1952 / DEFLATE RFC 1951).  In the reference this is librdkafka's job underneath
// tensorflow-io's KafkaDataset (cardata-v1.py:10): the consumer accepts whatever compression.type the producer chose.
//
// One routine per codec, written as plain sequential byte code that compiles for the host and for the device (one thread
// inflates one batch; a fetch response holds many batches).  Every routine bounds-checks each read and write — the input is
// untrusted — and runs in two modes: dst == nullptr only counts the output bytes (the size pass), otherwise it writes at most
// `cap` bytes.  Return value: bytes produced, or a negative B2_INFL_* code.
#pragma once
#include <stdint.h>

#ifndef B2_HD
#if defined(__CUDACC__)
#define B2_HD __host__ __device__ inline
#else
#define B2_HD inline
#endif
#endif

#define B2_INFL_TRUNCATED (-1)   /* input ends inside a token / block */
#define B2_INFL_CORRUPT   (-2)   /* impossible offset, bad header, bad Huffman code, length mismatch */
#define B2_INFL_OVERFLOW  (-3)   /* output does not fit `cap` */
#define B2_INFL_UNSUPPORTED (-4) /* a feature this decoder does not implement (LZ4 dictionary / linked-block options are fine; zstd is not here) */

// ---- output cursor shared by the three decoders -------------------------------------------------------------------------
struct B2Out {
  uint8_t* dst;      // nullptr: count only
  long long cap, n;  // capacity, bytes produced so far
};
B2_HD bool b2_out_lit(B2Out& o, const uint8_t* src, long long len) {
  if (o.dst) {
    if (o.n + len > o.cap) return false;
    for (long long i = 0; i < len; ++i) o.dst[o.n + i] = src[i];
  }
  o.n += len;
  return true;
}
// copy `len` bytes from `dist` bytes back (may overlap forwards: byte by byte); `base` = first output index dist may reach back to
B2_HD int b2_out_match(B2Out& o, long long dist, long long len, long long base) {
  if (dist <= 0 || dist > o.n - base) return B2_INFL_CORRUPT;
  if (o.dst) {
    if (o.n + len > o.cap) return B2_INFL_OVERFLOW;
    for (long long i = 0; i < len; ++i) o.dst[o.n + i] = o.dst[o.n + i - dist];
  }
  o.n += len;
  return 0;
}

// ---- Snappy ----------------------------------------------------------------------------------------------------------------
// raw block: varint uncompressed length, then elements: tag & 3 = 0 literal, 1 copy (1-byte offset), 2 copy (2-byte), 3 copy (4-byte)
B2_HD long long b2_snappy_block(const uint8_t* s, long long n, B2Out& o) {
  long long p = 0;
  unsigned long long ulen = 0;
  int shift = 0;
  for (;;) {
    if (p >= n) return B2_INFL_TRUNCATED;
    const uint8_t b = s[p++];
    ulen |= (unsigned long long)(b & 0x7F) << shift;
    if (!(b & 0x80)) break;
    shift += 7;
    if (shift > 35) return B2_INFL_CORRUPT;
  }
  const long long start = o.n;
  while (p < n) {
    const uint32_t tag = s[p++];
    if ((tag & 3u) == 0) {
      long long len = (tag >> 2) + 1;
      if (len > 60) {
        const int nb = (int)len - 60;  // 1..4 length bytes follow
        if (p + nb > n) return B2_INFL_TRUNCATED;
        len = 0;
        for (int k = 0; k < nb; ++k) len |= (long long)s[p + k] << (8 * k);
        len += 1;
        p += nb;
      }
      if (p + len > n) return B2_INFL_TRUNCATED;
      if (!b2_out_lit(o, s + p, len)) return B2_INFL_OVERFLOW;
      p += len;
    } else {
      long long len, off;
      if ((tag & 3u) == 1) {
        if (p + 1 > n) return B2_INFL_TRUNCATED;
        len = 4 + ((tag >> 2) & 7);
        off = ((long long)(tag >> 5) << 8) | s[p];
        p += 1;
      } else if ((tag & 3u) == 2) {
        if (p + 2 > n) return B2_INFL_TRUNCATED;
        len = (tag >> 2) + 1;
        off = (long long)s[p] | ((long long)s[p + 1] << 8);
        p += 2;
      } else {
        if (p + 4 > n) return B2_INFL_TRUNCATED;
        len = (tag >> 2) + 1;
        off = (long long)s[p] | ((long long)s[p + 1] << 8) | ((long long)s[p + 2] << 16) | ((long long)s[p + 3] << 24);
        p += 4;
      }
      const int rc = b2_out_match(o, off, len, start);
      if (rc) return rc;
    }
  }
  if ((unsigned long long)(o.n - start) != ulen) return B2_INFL_CORRUPT;
  return o.n - start;
}
// raw block, or xerial framing: "\x82SNAPPY\0", version (4 bytes BE), compatible version (4 bytes BE), then [length (4 bytes BE)][raw block] …
B2_HD long long b2_snappy(const uint8_t* s, long long n, uint8_t* dst, long long cap) {
  B2Out o{dst, cap, 0};
  const bool xerial = n >= 16 && s[0] == 0x82 && s[1] == 'S' && s[2] == 'N' && s[3] == 'A' && s[4] == 'P' && s[5] == 'P' && s[6] == 'Y' && s[7] == 0;
  if (!xerial) {
    const long long r = b2_snappy_block(s, n, o);
    return r < 0 ? r : o.n;
  }
  long long p = 16;
  while (p < n) {
    if (p + 4 > n) return B2_INFL_TRUNCATED;
    const long long bl = ((long long)s[p] << 24) | ((long long)s[p + 1] << 16) | ((long long)s[p + 2] << 8) | (long long)s[p + 3];
    p += 4;
    if (bl < 0 || p + bl > n) return B2_INFL_TRUNCATED;
    const long long r = b2_snappy_block(s + p, bl, o);
    if (r < 0) return r;
    p += bl;
  }
  return o.n;
}

// ---- LZ4 frame -------------------------------------------------------------------------------------------------------------
B2_HD long long b2_lz4_block(const uint8_t* s, long long n, B2Out& o, long long base) {
  long long p = 0;
  while (p < n) {
    const uint32_t tok = s[p++];
    long long lit = tok >> 4;
    if (lit == 15) {
      for (;;) {
        if (p >= n) return B2_INFL_TRUNCATED;
        const uint8_t b = s[p++];
        lit += b;
        if (b != 255) break;
      }
    }
    if (p + lit > n) return B2_INFL_TRUNCATED;
    if (!b2_out_lit(o, s + p, lit)) return B2_INFL_OVERFLOW;
    p += lit;
    if (p >= n) break;  // the last sequence has no match part
    if (p + 2 > n) return B2_INFL_TRUNCATED;
    const long long off = (long long)s[p] | ((long long)s[p + 1] << 8);
    p += 2;
    long long ml = tok & 15u;
    if (ml == 15) {
      for (;;) {
        if (p >= n) return B2_INFL_TRUNCATED;
        const uint8_t b = s[p++];
        ml += b;
        if (b != 255) break;
      }
    }
    ml += 4;
    const int rc = b2_out_match(o, off, ml, base);
    if (rc) return rc;
  }
  return 0;
}
B2_HD long long b2_lz4_frame(const uint8_t* s, long long n, uint8_t* dst, long long cap) {
  B2Out o{dst, cap, 0};
  if (n < 7) return B2_INFL_TRUNCATED;
  if (!(s[0] == 0x04 && s[1] == 0x22 && s[2] == 0x4D && s[3] == 0x18)) return B2_INFL_CORRUPT;
  const uint32_t flg = s[4];
  if (((flg >> 6) & 3u) != 1u) return B2_INFL_CORRUPT;  // version 01
  const bool block_indep = (flg >> 5) & 1u, block_checksum = (flg >> 4) & 1u, content_size = (flg >> 3) & 1u, content_checksum = (flg >> 2) & 1u,
             dict_id = flg & 1u;
  long long p = 6;  // FLG, BD
  if (content_size) p += 8;
  if (dict_id) p += 4;
  p += 1;  // header checksum (not verified: Kafka's own batch CRC covers these bytes, and old clients wrote a wrong one)
  if (p > n) return B2_INFL_TRUNCATED;
  for (;;) {
    if (p + 4 > n) return B2_INFL_TRUNCATED;
    const uint32_t bs = (uint32_t)s[p] | ((uint32_t)s[p + 1] << 8) | ((uint32_t)s[p + 2] << 16) | ((uint32_t)s[p + 3] << 24);
    p += 4;
    if (bs == 0) break;  // EndMark
    const long long len = bs & 0x7FFFFFFFu;
    if (p + len > n) return B2_INFL_TRUNCATED;
    if (bs & 0x80000000u) {
      if (!b2_out_lit(o, s + p, len)) return B2_INFL_OVERFLOW;
    } else {
      // linked blocks may reach back into earlier blocks (up to 64 KB); independent blocks only into themselves
      const long long rc = b2_lz4_block(s + p, len, o, block_indep ? o.n : 0);
      if (rc < 0) return rc;
    }
    p += len;
    if (block_checksum) p += 4;
  }
  (void)content_checksum;  // trailing xxHash32 of the content: not verified (see above)
  return o.n;
}

// ---- gzip / DEFLATE ------------------------------------------------------------------------------------------------------------
struct B2Bits {
  const uint8_t* s;
  long long n, p;
  uint32_t buf;
  int cnt;
};
B2_HD int b2_bits(B2Bits& b, int need) {  // `need` ≤ 16 bits, LSB first; -1 when the input is exhausted
  while (b.cnt < need) {
    if (b.p >= b.n) return -1;
    b.buf |= (uint32_t)b.s[b.p++] << b.cnt;
    b.cnt += 8;
  }
  const int v = (int)(b.buf & ((1u << need) - 1u));
  b.buf >>= need;
  b.cnt -= need;
  return v;
}
struct B2Huff {
  uint16_t count[16];
  uint16_t symbol[288];
};
// canonical Huffman table from code lengths; returns 0 complete, > 0 incomplete (allowed only for a single-code distance table), < 0 over-subscribed
B2_HD int b2_huff_build(B2Huff& h, const uint8_t* len, int n) {
  for (int i = 0; i < 16; ++i) h.count[i] = 0;
  for (int i = 0; i < n; ++i) h.count[len[i]]++;
  if (h.count[0] == n) return 0;
  int left = 1;
  for (int l = 1; l < 16; ++l) {
    left <<= 1;
    left -= h.count[l];
    if (left < 0) return left;
  }
  uint16_t offs[16];
  offs[1] = 0;
  for (int l = 1; l < 15; ++l) offs[l + 1] = offs[l] + h.count[l];
  for (int i = 0; i < n; ++i)
    if (len[i]) h.symbol[offs[len[i]]++] = (uint16_t)i;
  return left;
}
B2_HD int b2_huff_decode(B2Bits& b, const B2Huff& h) {
  int code = 0, first = 0, index = 0;
  for (int l = 1; l < 16; ++l) {
    const int bit = b2_bits(b, 1);
    if (bit < 0) return -1;
    code |= bit;
    const int count = h.count[l];
    if (code - count < first) return h.symbol[index + (code - first)];
    index += count;
    first += count;
    first <<= 1;
    code <<= 1;
  }
  return -2;
}
B2_HD long long b2_inflate(const uint8_t* s, long long n, B2Out& o, long long* consumed) {
  const uint16_t lbase[29] = {3, 4, 5, 6, 7, 8, 9, 10, 11, 13, 15, 17, 19, 23, 27, 31, 35, 43, 51, 59, 67, 83, 99, 115, 131, 163, 195, 227, 258};
  const uint8_t lext[29] = {0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 0};
  const uint16_t dbase[30] = {1, 2, 3, 4, 5, 7, 9, 13, 17, 25, 33, 49, 65, 97, 129, 193, 257, 385, 513, 769, 1025, 1537, 2049, 3073, 4097, 6145, 8193, 12289, 16385, 24577};
  const uint8_t dext[30] = {0, 0, 0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 5, 5, 6, 6, 7, 7, 8, 8, 9, 9, 10, 10, 11, 11, 12, 12, 13, 13};
  const uint8_t order[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};
  B2Bits b{s, n, 0, 0u, 0};
  B2Huff hl, hd;
  uint8_t lengths[320];
  int last;
  do {
    last = b2_bits(b, 1);
    const int type = b2_bits(b, 2);
    if (last < 0 || type < 0) return B2_INFL_TRUNCATED;
    if (type == 0) {  // stored
      b.buf = 0, b.cnt = 0;
      if (b.p + 4 > n) return B2_INFL_TRUNCATED;
      const long long len = (long long)s[b.p] | ((long long)s[b.p + 1] << 8), nlen = (long long)s[b.p + 2] | ((long long)s[b.p + 3] << 8);
      if ((len ^ 0xFFFF) != nlen) return B2_INFL_CORRUPT;
      b.p += 4;
      if (b.p + len > n) return B2_INFL_TRUNCATED;
      if (!b2_out_lit(o, s + b.p, len)) return B2_INFL_OVERFLOW;
      b.p += len;
      continue;
    }
    if (type == 3) return B2_INFL_CORRUPT;
    if (type == 1) {  // fixed codes
      for (int i = 0; i < 144; ++i) lengths[i] = 8;
      for (int i = 144; i < 256; ++i) lengths[i] = 9;
      for (int i = 256; i < 280; ++i) lengths[i] = 7;
      for (int i = 280; i < 288; ++i) lengths[i] = 8;
      b2_huff_build(hl, lengths, 288);
      for (int i = 0; i < 30; ++i) lengths[i] = 5;
      b2_huff_build(hd, lengths, 30);
    } else {  // dynamic codes
      const int nlen = b2_bits(b, 5), ndist = b2_bits(b, 5), ncode = b2_bits(b, 4);
      if (nlen < 0 || ndist < 0 || ncode < 0) return B2_INFL_TRUNCATED;
      const int NL = nlen + 257, ND = ndist + 1, NC = ncode + 4;
      if (NL > 286 || ND > 30) return B2_INFL_CORRUPT;
      for (int i = 0; i < 19; ++i) lengths[i] = 0;
      for (int i = 0; i < NC; ++i) {
        const int v = b2_bits(b, 3);
        if (v < 0) return B2_INFL_TRUNCATED;
        lengths[order[i]] = (uint8_t)v;
      }
      if (b2_huff_build(hl, lengths, 19) != 0) return B2_INFL_CORRUPT;
      int idx = 0;
      while (idx < NL + ND) {
        int sym = b2_huff_decode(b, hl);
        if (sym < 0) return sym == -1 ? B2_INFL_TRUNCATED : B2_INFL_CORRUPT;
        if (sym < 16) {
          lengths[idx++] = (uint8_t)sym;
        } else {
          int rep, val = 0;
          if (sym == 16) {
            if (idx == 0) return B2_INFL_CORRUPT;
            val = lengths[idx - 1];
            rep = b2_bits(b, 2);
            if (rep < 0) return B2_INFL_TRUNCATED;
            rep += 3;
          } else if (sym == 17) {
            rep = b2_bits(b, 3);
            if (rep < 0) return B2_INFL_TRUNCATED;
            rep += 3;
          } else {
            rep = b2_bits(b, 7);
            if (rep < 0) return B2_INFL_TRUNCATED;
            rep += 11;
          }
          if (idx + rep > NL + ND) return B2_INFL_CORRUPT;
          while (rep--) lengths[idx++] = (uint8_t)val;
        }
      }
      if (lengths[256] == 0) return B2_INFL_CORRUPT;  // no end-of-block code
      int rc = b2_huff_build(hl, lengths, NL);
      if (rc < 0 || (rc > 0 && NL - hl.count[0] != 1)) return B2_INFL_CORRUPT;
      rc = b2_huff_build(hd, lengths + NL, ND);
      if (rc < 0 || (rc > 0 && ND - hd.count[0] != 1)) return B2_INFL_CORRUPT;
    }
    for (;;) {  // the block's symbols
      int sym = b2_huff_decode(b, hl);
      if (sym < 0) return sym == -1 ? B2_INFL_TRUNCATED : B2_INFL_CORRUPT;
      if (sym < 256) {
        const uint8_t c = (uint8_t)sym;
        if (!b2_out_lit(o, &c, 1)) return B2_INFL_OVERFLOW;
      } else if (sym == 256) {
        break;
      } else {
        sym -= 257;
        if (sym >= 29) return B2_INFL_CORRUPT;
        const int eb = b2_bits(b, lext[sym]);
        if (eb < 0) return B2_INFL_TRUNCATED;
        const int len = lbase[sym] + eb;
        const int ds = b2_huff_decode(b, hd);
        if (ds < 0) return ds == -1 ? B2_INFL_TRUNCATED : B2_INFL_CORRUPT;
        if (ds >= 30) return B2_INFL_CORRUPT;
        const int de = b2_bits(b, dext[ds]);
        if (de < 0) return B2_INFL_TRUNCATED;
        const int rc = b2_out_match(o, dbase[ds] + de, len, 0);
        if (rc) return rc;
      }
    }
  } while (!last);
  if (consumed) *consumed = b.p;
  return o.n;
}
// gzip member: 10-byte header (+ optional FEXTRA / FNAME / FCOMMENT / FHCRC), DEFLATE stream, CRC-32 and ISIZE (ISIZE is checked)
B2_HD long long b2_gzip(const uint8_t* s, long long n, uint8_t* dst, long long cap) {
  if (n < 18) return B2_INFL_TRUNCATED;
  if (s[0] != 0x1F || s[1] != 0x8B || s[2] != 8) return B2_INFL_CORRUPT;
  const uint32_t flg = s[3];
  long long p = 10;
  if (flg & 4u) {
    if (p + 2 > n) return B2_INFL_TRUNCATED;
    p += 2 + ((long long)s[p] | ((long long)s[p + 1] << 8));
  }
  if (flg & 8u) {
    while (p < n && s[p]) ++p;
    ++p;
  }
  if (flg & 16u) {
    while (p < n && s[p]) ++p;
    ++p;
  }
  if (flg & 2u) p += 2;
  if (p >= n) return B2_INFL_TRUNCATED;
  B2Out o{dst, cap, 0};
  long long used = 0;
  const long long r = b2_inflate(s + p, n - p, o, &used);
  if (r < 0) return r;
  p += used;
  if (p + 8 > n) return B2_INFL_TRUNCATED;
  const uint32_t isize = (uint32_t)s[p + 4] | ((uint32_t)s[p + 5] << 8) | ((uint32_t)s[p + 6] << 16) | ((uint32_t)s[p + 7] << 24);
  if (isize != (uint32_t)o.n) return B2_INFL_CORRUPT;
  return o.n;
}

// codec = attributes & 7 of the batch header: 1 gzip, 2 snappy, 3 lz4, 4 zstd (not implemented)
B2_HD long long b2_inflate_codec(int codec, const uint8_t* s, long long n, uint8_t* dst, long long cap) {
  if (codec == 1) return b2_gzip(s, n, dst, cap);
  if (codec == 2) return b2_snappy(s, n, dst, cap);
  if (codec == 3) return b2_lz4_frame(s, n, dst, cap);
  return B2_INFL_UNSUPPORTED;
}
