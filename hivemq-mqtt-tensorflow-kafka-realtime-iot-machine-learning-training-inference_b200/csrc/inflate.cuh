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
#define B2_INFL_UNSUPPORTED (-4) /* a feature this decoder does not implement (a zstd dictionary) */

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

// ---- Zstandard (RFC 8878) -------------------------------------------------------------------------------------------------
// Frames of raw / RLE / compressed blocks; a compressed block = literals section (raw, RLE or Huffman-coded, 1 or 4 streams) +
// sequences section (literal length / offset / match length codes, three interleaved FSE states over one backward bit stream).
// No dictionaries (Kafka's producers use none).  Everything a block needs lives in B2Zstd (≈ 10 KB: three FSE tables, one Huffman
// table) — on the GPU that is the thread's local memory; the literals are decoded lazily, as the sequences consume them, so no
// literal buffer exists.  The content checksum (xxHash64) is not verified: the Kafka batch's CRC-32C covers the compressed bytes.
struct B2ZEntry {
  uint16_t base;  // FSE: new state before the extra bits
  uint8_t sym, nb;
};
struct B2ZFse {
  B2ZEntry e[512];
  int log;
};
// forward bit reader (LSB first) for the FSE table descriptions
struct B2ZFwd {
  const uint8_t* p;
  long long n, bit;
};
B2_HD uint32_t b2z_peek_fwd(const B2ZFwd& r, int k) {  // k ≤ 24; bits beyond the end read as zero
  uint32_t v = 0;
  const long long by = r.bit >> 3;
  for (int i = 0; i < 4; ++i)
    if (by + i < r.n) v |= (uint32_t)r.p[by + i] << (8 * i);
  return (v >> (r.bit & 7)) & ((1u << k) - 1u);
}
// backward bit stream: the last byte carries a 1 marker above the payload; bits are taken from the top down
struct B2ZBack {
  const uint8_t* p;
  long long bitpos;  // bits not yet consumed; negative once more was asked for than there is
};
B2_HD int b2z_highbit(uint32_t v) {
  int h = 0;
  while (v >>= 1) ++h;
  return h;
}
B2_HD bool b2z_back_init(B2ZBack& b, const uint8_t* p, long long n) {
  if (n < 1 || p[n - 1] == 0) return false;
  b.p = p;
  b.bitpos = 8 * (n - 1) + b2z_highbit(p[n - 1]);
  return true;
}
// bits [pos, pos + k) of the stream as a little-endian number, k ≤ 32; positions below zero read as zero
B2_HD uint32_t b2z_bits_at(const uint8_t* p, long long pos, int k) {
  int up = 0;  // bits of the request that lie below position zero
  if (pos < 0) {
    if (pos + k <= 0) return 0;
    up = (int)(-pos);
    k -= up;
    pos = 0;
  }
  if (k == 0) return 0;
  uint64_t v = 0;
  const long long by = pos >> 3;
  const int nbytes = (int)(((pos & 7) + k + 7) >> 3);
  for (int i = 0; i < nbytes; ++i) v |= (uint64_t)p[by + i] << (8 * i);
  return (uint32_t)((v >> (pos & 7)) & (k == 32 ? 0xFFFFFFFFull : ((1ull << k) - 1ull))) << up;
}
B2_HD uint32_t b2z_read_back(B2ZBack& b, int k) {
  b.bitpos -= k;
  return b2z_bits_at(b.p, b.bitpos, k);
}

// normalised counts of an FSE table description → *consumed bytes; returns the accuracy log or a negative error
B2_HD int b2z_read_ncount(const uint8_t* s, long long n, int max_log, int max_sym, short* norm, int* n_sym, long long* consumed) {
  B2ZFwd r{s, n, 0};
  const int log = (int)b2z_peek_fwd(r, 4) + 5;
  r.bit += 4;
  if (log > max_log) return B2_INFL_CORRUPT;
  int remaining = (1 << log) + 1, threshold = 1 << log, nb = log + 1, sym = 0;
  bool prev0 = false;
  while (remaining > 1 && sym <= max_sym) {
    if (prev0) {
      int n0 = sym;
      for (;;) {
        const uint32_t two = b2z_peek_fwd(r, 2);
        r.bit += 2;
        n0 += (int)two;
        if (two != 3) break;
        if (r.bit > 8 * n) return B2_INFL_TRUNCATED;
      }
      if (n0 > max_sym + 1) return B2_INFL_CORRUPT;
      while (sym < n0) norm[sym++] = 0;
      if (sym > max_sym) break;
    }
    const int mx = (2 * threshold - 1) - remaining;
    const uint32_t bs = b2z_peek_fwd(r, nb);
    int count;
    if ((int)(bs & (uint32_t)(threshold - 1)) < mx) {
      count = (int)(bs & (uint32_t)(threshold - 1));
      r.bit += nb - 1;
    } else {
      count = (int)(bs & (uint32_t)(2 * threshold - 1));
      if (count >= threshold) count -= mx;
      r.bit += nb;
    }
    --count;  // -1 = "less than one": a single cell at the top of the table
    remaining -= count < 0 ? -count : count;
    norm[sym++] = (short)count;
    prev0 = count == 0;
    while (remaining < threshold) {
      --nb;
      threshold >>= 1;
    }
    if (r.bit > 8 * n) return B2_INFL_TRUNCATED;
  }
  if (remaining != 1 || sym > max_sym + 1) return B2_INFL_CORRUPT;
  *n_sym = sym;
  *consumed = (r.bit + 7) >> 3;
  return log;
}
B2_HD int b2z_build_fse(B2ZFse& t, const short* norm, int n_sym, int log) {
  const int size = 1 << log;
  uint16_t next[64];
  int high = size - 1;
  for (int s = 0; s < n_sym; ++s) {
    if (norm[s] == -1) {
      t.e[high--].sym = (uint8_t)s;
      next[s] = 1;
    } else {
      next[s] = (uint16_t)norm[s];
    }
  }
  const int step = (size >> 1) + (size >> 3) + 3, mask = size - 1;
  int pos = 0;
  for (int s = 0; s < n_sym; ++s)
    for (int i = 0; i < norm[s]; ++i) {
      t.e[pos].sym = (uint8_t)s;
      do pos = (pos + step) & mask;
      while (pos > high);
    }
  if (pos != 0) return B2_INFL_CORRUPT;
  for (int u = 0; u < size; ++u) {
    const int sy = t.e[u].sym;
    const uint32_t ns = next[sy]++;
    const int nb = log - b2z_highbit(ns);
    t.e[u].nb = (uint8_t)nb;
    t.e[u].base = (uint16_t)((ns << nb) - (uint32_t)size);
  }
  t.log = log;
  return 0;
}
B2_HD void b2z_rle_fse(B2ZFse& t, int sym) {
  t.e[0].sym = (uint8_t)sym;
  t.e[0].nb = 0;
  t.e[0].base = 0;
  t.log = 0;
}

struct B2ZHuf {
  uint8_t sym[2048], nb[2048];
  int log;  // 0 = no table yet
};
// Huffman tree description → table; returns bytes consumed or a negative error
B2_HD long long b2z_read_huf(B2ZHuf& h, const uint8_t* s, long long n) {
  if (n < 1) return B2_INFL_TRUNCATED;
  uint8_t w[256];
  int nw = 0;
  long long used;
  const int hb = s[0];
  if (hb >= 128) {  // weights stored directly, 4 bits each
    nw = hb - 127;
    used = 1 + (nw + 1) / 2;
    if (used > n) return B2_INFL_TRUNCATED;
    for (int i = 0; i < nw; ++i) w[i] = (i & 1) ? (s[1 + i / 2] & 15) : (s[1 + i / 2] >> 4);
  } else {  // FSE-compressed weights: two interleaved states over a backward stream
    used = 1 + hb;
    if (hb < 2 || used > n) return hb < 2 ? B2_INFL_CORRUPT : B2_INFL_TRUNCATED;
    short norm[16];
    int ns = 0;
    long long c = 0;
    const int log = b2z_read_ncount(s + 1, hb, 6, 12, norm, &ns, &c);
    if (log < 0) return log;
    B2ZFse t;
    if (b2z_build_fse(t, norm, ns, log) < 0) return B2_INFL_CORRUPT;
    B2ZBack b;
    if (c >= hb || !b2z_back_init(b, s + 1 + c, hb - c)) return B2_INFL_CORRUPT;
    uint32_t s1 = b2z_read_back(b, log), s2 = b2z_read_back(b, log);
    if (b.bitpos < 0) return B2_INFL_CORRUPT;
    for (;;) {
      if (nw > 253) return B2_INFL_CORRUPT;
      w[nw++] = t.e[s1].sym;
      s1 = t.e[s1].base + b2z_read_back(b, t.e[s1].nb);
      if (b.bitpos < 0) {
        w[nw++] = t.e[s2].sym;
        break;
      }
      w[nw++] = t.e[s2].sym;
      s2 = t.e[s2].base + b2z_read_back(b, t.e[s2].nb);
      if (b.bitpos < 0) {
        w[nw++] = t.e[s1].sym;
        break;
      }
    }
  }
  // the last weight completes the sum of 2^(w-1) to a power of two
  uint32_t sum = 0;
  for (int i = 0; i < nw; ++i) {
    if (w[i] > 11) return B2_INFL_CORRUPT;
    if (w[i]) sum += 1u << (w[i] - 1);
  }
  if (sum == 0) return B2_INFL_CORRUPT;
  const int log = b2z_highbit(sum) + 1;
  if (log > 11) return B2_INFL_CORRUPT;
  const uint32_t rest = (1u << log) - sum;
  if (rest == 0 || (rest & (rest - 1)) != 0) return B2_INFL_CORRUPT;
  w[nw++] = (uint8_t)(b2z_highbit(rest) + 1);
  // cells in order of rising weight (longest codes first), symbols of one weight in rising order
  int pos = 0;
  for (int wt = 1; wt <= log; ++wt)
    for (int sy = 0; sy < nw; ++sy)
      if (w[sy] == wt) {
        const int cells = 1 << (wt - 1);
        for (int i = 0; i < cells; ++i) {
          h.sym[pos + i] = (uint8_t)sy;
          h.nb[pos + i] = (uint8_t)(log + 1 - wt);
        }
        pos += cells;
      }
  if (pos != (1 << log)) return B2_INFL_CORRUPT;
  h.log = log;
  return used;
}

// the literals of one block, handed out in order
struct B2ZLit {
  int type;             // 0 raw, 1 RLE, 2 Huffman
  const uint8_t* raw;   // raw: next byte; RLE: the byte
  long long left;       // literals not yet handed out
  // Huffman: up to four streams, consumed one after the other
  B2ZBack st[4];
  long long st_left[4];
  int cur, n_st;
};
B2_HD int b2z_lit_take(B2ZLit& l, const B2ZHuf& h, B2Out& o, long long len) {
  if (len > l.left) return B2_INFL_CORRUPT;
  l.left -= len;
  if (o.dst && o.n + len > o.cap) return B2_INFL_OVERFLOW;
  if (l.type == 0) {
    if (o.dst)
      for (long long i = 0; i < len; ++i) o.dst[o.n + i] = l.raw[i];
    l.raw += len;
  } else if (l.type == 1) {
    if (o.dst)
      for (long long i = 0; i < len; ++i) o.dst[o.n + i] = l.raw[0];
  } else {
    for (long long i = 0; i < len; ++i) {
      while (l.cur < l.n_st && l.st_left[l.cur] == 0) {
        if (l.st[l.cur].bitpos != 0) return B2_INFL_CORRUPT;  // a stream must end exactly on its first bit
        ++l.cur;
      }
      if (l.cur >= l.n_st) return B2_INFL_CORRUPT;
      B2ZBack& b = l.st[l.cur];
      const uint32_t idx = b2z_bits_at(b.p, b.bitpos - h.log, h.log);
      b.bitpos -= h.nb[idx];
      if (b.bitpos < 0) return B2_INFL_CORRUPT;
      --l.st_left[l.cur];
      if (o.dst) o.dst[o.n + i] = h.sym[idx];
    }
  }
  o.n += len;
  return 0;
}

struct B2Zstd {
  B2ZFse ll, of, ml;
  B2ZHuf huf;
  bool have_ll, have_of, have_ml;
  uint32_t rep[3];
};
// one table of the sequences section: mode 0 predefined, 1 RLE, 2 described, 3 the previous block's
B2_HD long long b2z_seq_table(B2ZFse& t, bool& have, int mode, const uint8_t* s, long long n, int max_log, int max_sym, const short* dflt,
                              int dflt_n, int dflt_log) {
  if (mode == 0) {
    if (b2z_build_fse(t, dflt, dflt_n, dflt_log) < 0) return B2_INFL_CORRUPT;
    have = true;
    return 0;
  }
  if (mode == 1) {
    if (n < 1) return B2_INFL_TRUNCATED;
    if (s[0] > max_sym) return B2_INFL_CORRUPT;
    b2z_rle_fse(t, s[0]);
    have = true;
    return 1;
  }
  if (mode == 2) {
    short norm[64];
    int ns = 0;
    long long c = 0;
    const int log = b2z_read_ncount(s, n, max_log, max_sym, norm, &ns, &c);
    if (log < 0) return log;
    if (b2z_build_fse(t, norm, ns, log) < 0) return B2_INFL_CORRUPT;
    have = true;
    return c;
  }
  return have ? 0 : B2_INFL_CORRUPT;
}

B2_HD int b2z_block(B2Zstd& z, const uint8_t* s, long long n, B2Out& o, long long base) {
  // ---- literals section
  if (n < 1) return B2_INFL_TRUNCATED;
  B2ZLit lit;
  lit.cur = 0, lit.n_st = 0, lit.raw = s;
  const int ltype = s[0] & 3, sf = (s[0] >> 2) & 3;
  long long p = 0, regen = 0;
  if (ltype < 2) {
    if (sf == 0 || sf == 2) regen = s[0] >> 3, p = 1;
    else if (sf == 1) {
      if (n < 2) return B2_INFL_TRUNCATED;
      regen = (s[0] >> 4) | ((long long)s[1] << 4), p = 2;
    } else {
      if (n < 3) return B2_INFL_TRUNCATED;
      regen = (s[0] >> 4) | ((long long)s[1] << 4) | ((long long)s[2] << 12), p = 3;
    }
    lit.type = ltype;
    lit.raw = s + p;
    const long long stored = ltype == 0 ? regen : 1;
    if (p + stored > n) return B2_INFL_TRUNCATED;
    p += stored;
  } else {
    long long comp;
    const int hl = sf == 3 ? 5 : (sf == 2 ? 4 : 3);
    if (n < hl) return B2_INFL_TRUNCATED;
    unsigned long long hdr = 0;
    for (int i = 0; i < hl; ++i) hdr |= (unsigned long long)s[i] << (8 * i);
    if (sf < 2) regen = (hdr >> 4) & 0x3FF, comp = (hdr >> 14) & 0x3FF;
    else if (sf == 2) regen = (hdr >> 4) & 0x3FFF, comp = (hdr >> 18) & 0x3FFF;
    else regen = (hdr >> 4) & 0x3FFFF, comp = (hdr >> 22) & 0x3FFFF;
    p = hl;
    if (p + comp > n) return B2_INFL_TRUNCATED;
    const uint8_t* c = s + p;
    long long cn = comp;
    if (ltype == 2) {
      const long long u = b2z_read_huf(z.huf, c, cn);
      if (u < 0) return (int)u;
      c += u, cn -= u;
    } else if (z.huf.log == 0) {
      return B2_INFL_CORRUPT;  // "treeless" with no tree from an earlier block
    }
    lit.type = 2;
    if (sf == 0) {
      lit.n_st = 1;
      lit.st_left[0] = regen;
      if (!b2z_back_init(lit.st[0], c, cn)) return B2_INFL_CORRUPT;
    } else {
      if (cn < 6) return B2_INFL_CORRUPT;
      const long long z1 = c[0] | (c[1] << 8), z2 = c[2] | (c[3] << 8), z3 = c[4] | (c[5] << 8), z4 = cn - 6 - z1 - z2 - z3;
      if (z4 < 1) return B2_INFL_CORRUPT;
      const long long per = (regen + 3) / 4;
      if (regen < 3 * per) return B2_INFL_CORRUPT;
      const uint8_t* q = c + 6;
      const long long zs[4] = {z1, z2, z3, z4};
      lit.n_st = 4;
      for (int i = 0; i < 4; ++i) {
        lit.st_left[i] = i < 3 ? per : regen - 3 * per;
        if (!b2z_back_init(lit.st[i], q, zs[i])) return B2_INFL_CORRUPT;
        q += zs[i];
      }
    }
    p += comp;
  }
  lit.left = regen;
  // ---- sequences section
  if (p >= n) return B2_INFL_TRUNCATED;
  long long nseq = s[p++];
  if (nseq >= 128) {
    if (nseq == 255) {
      if (p + 2 > n) return B2_INFL_TRUNCATED;
      nseq = s[p] + ((long long)s[p + 1] << 8) + 0x7F00;
      p += 2;
    } else {
      if (p + 1 > n) return B2_INFL_TRUNCATED;
      nseq = ((nseq - 128) << 8) + s[p];
      p += 1;
    }
  }
  if (nseq > 0) {
    if (p >= n) return B2_INFL_TRUNCATED;
    const int modes = s[p++];
    if (modes & 3) return B2_INFL_CORRUPT;
    const short ll_d[36] = {4, 3, 2, 2, 2, 2, 2, 2, 2, 2, 2, 2, 2, 1, 1, 1, 2, 2, 2, 2, 2, 2, 2, 2, 2, 3, 2, 1, 1, 1, 1, 1, -1, -1, -1, -1};
    const short of_d[29] = {1, 1, 1, 1, 1, 1, 2, 2, 2, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, -1, -1, -1, -1, -1};
    const short ml_d[53] = {1, 4, 3, 2, 2, 2, 2, 2, 2, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1,
                            1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, -1, -1, -1, -1, -1, -1, -1};
    long long u = b2z_seq_table(z.ll, z.have_ll, (modes >> 6) & 3, s + p, n - p, 9, 35, ll_d, 36, 6);
    if (u < 0) return (int)u;
    p += u;
    u = b2z_seq_table(z.of, z.have_of, (modes >> 4) & 3, s + p, n - p, 8, 31, of_d, 29, 5);
    if (u < 0) return (int)u;
    p += u;
    u = b2z_seq_table(z.ml, z.have_ml, (modes >> 2) & 3, s + p, n - p, 9, 52, ml_d, 53, 6);
    if (u < 0) return (int)u;
    p += u;
    B2ZBack b;
    if (p >= n || !b2z_back_init(b, s + p, n - p)) return B2_INFL_CORRUPT;
    uint32_t sl = b2z_read_back(b, z.ll.log), so = b2z_read_back(b, z.of.log), sm = b2z_read_back(b, z.ml.log);
    if (b.bitpos < 0) return B2_INFL_CORRUPT;
    const uint32_t ll_base[36] = {0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15, 16, 18, 20, 22, 24, 28, 32, 40, 48, 64, 128, 256, 512,
                                  1024, 2048, 4096, 8192, 16384, 32768, 65536};
    const uint8_t ll_bits[36] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 3, 3, 4, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15, 16};
    const uint32_t ml_base[53] = {3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15, 16, 17, 18, 19, 20, 21, 22, 23, 24, 25, 26, 27, 28, 29, 30, 31, 32, 33,
                                  34, 35, 37, 39, 41, 43, 47, 51, 59, 67, 83, 99, 131, 259, 515, 1027, 2051, 4099, 8195, 16387, 32771, 65539};
    const uint8_t ml_bits[53] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0,
                                 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 3, 3, 4, 4, 5, 7, 8, 9, 10, 11, 12, 13, 14, 15, 16};
    for (long long q = 0; q < nseq; ++q) {
      const int oc = z.of.e[so].sym, mc = z.ml.e[sm].sym, lc = z.ll.e[sl].sym;
      if (oc > 31 || mc > 52 || lc > 35) return B2_INFL_CORRUPT;
      const uint32_t ov = (1u << oc) + b2z_read_back(b, oc);
      const uint32_t mlen = ml_base[mc] + b2z_read_back(b, ml_bits[mc]);
      const uint32_t llen = ll_base[lc] + b2z_read_back(b, ll_bits[lc]);
      if (q + 1 < nseq) {  // the states are refreshed in the order literal length, match length, offset
        sl = z.ll.e[sl].base + b2z_read_back(b, z.ll.e[sl].nb);
        sm = z.ml.e[sm].base + b2z_read_back(b, z.ml.e[sm].nb);
        so = z.of.e[so].base + b2z_read_back(b, z.of.e[so].nb);
      }
      if (b.bitpos < 0) return B2_INFL_CORRUPT;
      uint32_t off;
      if (ov > 3) {
        off = ov - 3;
        z.rep[2] = z.rep[1], z.rep[1] = z.rep[0], z.rep[0] = off;
      } else {
        const uint32_t idx = ov - 1 + (llen == 0 ? 1u : 0u);  // 0..3: repeat offsets 1..3, or (idx 3) the first one minus one
        if (idx == 0) {
          off = z.rep[0];
        } else {
          off = idx == 3 ? z.rep[0] - 1 : z.rep[idx];
          if (off == 0) return B2_INFL_CORRUPT;
          if (idx != 1) z.rep[2] = z.rep[1];
          z.rep[1] = z.rep[0];
          z.rep[0] = off;
        }
      }
      int rc = b2z_lit_take(lit, z.huf, o, llen);
      if (rc) return rc;
      rc = b2_out_match(o, off, mlen, base);
      if (rc) return rc;
    }
    if (b.bitpos != 0) return B2_INFL_CORRUPT;
  } else if (p != n) {
    return B2_INFL_CORRUPT;
  }
  return b2z_lit_take(lit, z.huf, o, lit.left);  // what the sequences left over
}

// a concatenation of frames (and skippable frames); returns the regenerated size or a negative error
B2_HD long long b2_zstd(const uint8_t* s, long long n, uint8_t* dst, long long cap) {
  B2Out o{dst, cap, 0};
  B2Zstd z;
  long long p = 0;
  if (n < 4) return B2_INFL_TRUNCATED;
  while (p < n) {
    if (p + 4 > n) return B2_INFL_TRUNCATED;
    const uint32_t magic = s[p] | (s[p + 1] << 8) | (s[p + 2] << 16) | ((uint32_t)s[p + 3] << 24);
    p += 4;
    if ((magic & 0xFFFFFFF0u) == 0x184D2A50u) {  // skippable frame
      if (p + 4 > n) return B2_INFL_TRUNCATED;
      const long long sz = s[p] | (s[p + 1] << 8) | (s[p + 2] << 16) | ((long long)s[p + 3] << 24);
      p += 4 + sz;
      if (p > n) return B2_INFL_TRUNCATED;
      continue;
    }
    if (magic != 0xFD2FB528u) return B2_INFL_CORRUPT;
    if (p >= n) return B2_INFL_TRUNCATED;
    const int fhd = s[p++];
    if (fhd & 0x08) return B2_INFL_CORRUPT;  // reserved bit
    const int fcs_flag = fhd >> 6, single = (fhd >> 5) & 1, checksum = (fhd >> 2) & 1, did_flag = fhd & 3;
    const int did_bytes = did_flag == 3 ? 4 : did_flag;
    const int fcs_bytes = fcs_flag == 0 ? single : (fcs_flag == 1 ? 2 : (fcs_flag == 2 ? 4 : 8));
    if (p + (single ? 0 : 1) + did_bytes + fcs_bytes > n) return B2_INFL_TRUNCATED;
    if (!single) ++p;  // window descriptor: the whole output is addressable here
    unsigned long long did = 0;
    for (int i = 0; i < did_bytes; ++i) did |= (unsigned long long)s[p + i] << (8 * i);
    if (did != 0) return B2_INFL_UNSUPPORTED;
    p += did_bytes + fcs_bytes;
    const long long base = o.n;
    z.huf.log = 0;
    z.have_ll = z.have_of = z.have_ml = false;
    z.rep[0] = 1, z.rep[1] = 4, z.rep[2] = 8;
    for (;;) {
      if (p + 3 > n) return B2_INFL_TRUNCATED;
      const uint32_t bh = s[p] | (s[p + 1] << 8) | (s[p + 2] << 16);
      p += 3;
      const int last = bh & 1, type = (bh >> 1) & 3;
      const long long bs = bh >> 3;
      if (type == 0) {
        if (p + bs > n) return B2_INFL_TRUNCATED;
        if (!b2_out_lit(o, s + p, bs)) return B2_INFL_OVERFLOW;
        p += bs;
      } else if (type == 1) {
        if (p + 1 > n) return B2_INFL_TRUNCATED;
        if (o.dst) {
          if (o.n + bs > o.cap) return B2_INFL_OVERFLOW;
          for (long long i = 0; i < bs; ++i) o.dst[o.n + i] = s[p];
        }
        o.n += bs;
        p += 1;
      } else if (type == 2) {
        if (p + bs > n) return B2_INFL_TRUNCATED;
        const int rc = b2z_block(z, s + p, bs, o, base);
        if (rc) return rc;
        p += bs;
      } else {
        return B2_INFL_CORRUPT;
      }
      if (last) break;
    }
    if (checksum) {
      if (p + 4 > n) return B2_INFL_TRUNCATED;
      p += 4;
    }
  }
  return o.n;
}

// codec = attributes & 7 of the batch header: 1 gzip, 2 snappy, 3 lz4, 4 zstd
B2_HD long long b2_inflate_codec(int codec, const uint8_t* s, long long n, uint8_t* dst, long long cap) {
  if (codec == 1) return b2_gzip(s, n, dst, cap);
  if (codec == 2) return b2_snappy(s, n, dst, cap);
  if (codec == 3) return b2_lz4_frame(s, n, dst, cap);
  if (codec == 4) return b2_zstd(s, n, dst, cap);
  return B2_INFL_UNSUPPORTED;
}
