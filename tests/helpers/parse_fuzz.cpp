// CPU fuzz of csrc/float_parse.cuh (the exact code the CUDA decode kernel runs) against glibc strtof.
//   g++ -O2 -std=c++17 -o parse_fuzz parse_fuzz.cpp && ./parse_fuzz <n_random> [token_file]
// Prints "mismatches=<k> checked=<n>"; exit code 1 when k > 0.
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <random>
#include <string>

#include "../../hivemq-mqtt-tensorflow-kafka-realtime-iot-machine-learning-training-inference_b200/csrc/float_parse.cuh"

static long long bad = 0, checked = 0, fast_hits = 0, checked64 = 0;

// the SWAR fast path sees the token the way the kernel hands it over: 16 bytes + the last 4 bytes, garbage around it
static void check_fast(const std::string& tok, uint32_t ref_bits, bool ref_ok) {
  if (tok.empty() || tok.size() > 16) return;
  unsigned char buf[48];
  for (int i = 0; i < 48; ++i) buf[i] = (unsigned char)(",x\"9E-.5 "[i % 9]);  // hostile neighbours
  const int b = 8;
  memcpy(buf + b, tok.data(), tok.size());
  uint32_t t[4], tail;
  memcpy(t, buf + b, 16);
  memcpy(&tail, buf + b + tok.size() - 4, 4);
  uint32_t got = 0;
  if (b2_fast_token16(t[0], t[1], t[2], t[3], tail, (int)tok.size(), &got)) {
    ++fast_hits;
    if (!ref_ok || got != ref_bits) {
      if (bad < 20) fprintf(stderr, "FAST MISMATCH '%s' got 0x%08x ref ok=%d 0x%08x\n", tok.c_str(), got, (int)ref_ok, ref_bits);
      ++bad;
    }
  }
}

// the lock-step field walk's parser sees the raw byte stream: token at any alignment, then either the delimiter and more
// bytes, or the record's end with foreign bytes behind it
static long long num16_hits = 0;
static void check_num16(const std::string& tok, uint32_t ref_bits, bool ref_ok) {
  if (tok.find(',') != std::string::npos) return;  // a delimiter inside the token is a different field split, not a parse question
  static unsigned variant = 0;
  ++variant;
  alignas(4) unsigned char buf[96];
  for (int i = 0; i < 96; ++i) buf[i] = (unsigned char)("7,-\"9E-.5 \xC3"[(i + variant) % 11]);  // hostile neighbours
  const int b = 4 + (int)(variant % 4);
  memcpy(buf + b, tok.data(), tok.size());
  const bool at_end = (variant / 4) % 2;
  if (!at_end) buf[b + tok.size()] = ',';
  const int rem = at_end ? (int)tok.size() : (int)tok.size() + 1 + (int)(variant % 7);
  uint32_t w[5];
  memcpy(w, buf + (b & ~3), 20);
  uint32_t got = 0;
  const int r = b2_csv_num16(w[0], w[1], w[2], w[3], w[4], (b & 3) * 8, rem, ',', &got);
  if (r < 0) return;
  ++num16_hits;
  const bool good = tok.empty() ? r == 0 : (r == (int)tok.size() && ref_ok && got == ref_bits);
  if (!good) {
    if (bad < 20) fprintf(stderr, "NUM16 MISMATCH '%s' r=%d got 0x%08x ref ok=%d 0x%08x\n", tok.c_str(), r, got, (int)ref_ok, ref_bits);
    ++bad;
  }
}

static void check(const std::string& tok) {
  uint32_t got = 0;
  const int st = b2_parse_float((const uint8_t*)tok.data(), (int)tok.size(), &got);
  // reference: glibc strtof on the token with trailing blanks removed (decode_csv allows them); empty and hex tokens are
  // invalid by our (and decode_csv's decimal) grammar
  std::string t = tok;
  while (!t.empty() && (t.back() == ' ' || t.back() == '\t')) t.pop_back();
  char* end = nullptr;
  const float f = strtof(t.c_str(), &end);
  uint32_t ref;
  memcpy(&ref, &f, 4);
  ++checked;
  if (st == B2_NUM_TOO_LONG) return;  // allowed refusal (counted separately by the caller if needed)
  const bool hex = t.find('x') != std::string::npos || t.find('X') != std::string::npos;
  const bool ref_ok = !t.empty() && !hex && end == t.c_str() + t.size() && t.find_first_not_of(" \t") != std::string::npos;
  {  // binary64 parse (the JSON path) against strtod
    uint64_t got64 = 0;
    const int st64 = b2_parse_double((const uint8_t*)tok.data(), (int)tok.size(), &got64);
    char* end64 = nullptr;
    const double dref = strtod(t.c_str(), &end64);
    uint64_t ref64;
    memcpy(&ref64, &dref, 8);
    if (st64 != B2_NUM_TOO_LONG) {
      ++checked64;
      const bool nan64 = isnan(dref) && (got64 & 0x7FFFFFFFFFFFFFFFull) > 0x7FF0000000000000ull;
      if ((st64 == B2_NUM_OK) != ref_ok || (ref_ok && got64 != ref64 && !nan64)) {
        if (bad < 20) fprintf(stderr, "F64 MISMATCH '%s' got st=%d 0x%016llx ref ok=%d 0x%016llx\n", tok.c_str(), st64, (unsigned long long)got64, (int)ref_ok, (unsigned long long)ref64);
        ++bad;
      }
    }
  }
  check_fast(tok, ref, ref_ok && !isnan(f));
  check_num16(tok, ref, ref_ok && !isnan(f));
  if ((st == B2_NUM_OK) != ref_ok || (ref_ok && got != ref && !(isnan(f) && (got & 0x7FFFFFFF) > 0x7F800000))) {
    if (bad < 20) fprintf(stderr, "MISMATCH '%s' got st=%d 0x%08x ref ok=%d 0x%08x\n", tok.c_str(), st, got, (int)ref_ok, ref);
    ++bad;
  }
}

int main(int argc, char** argv) {
  const long long n = argc > 1 ? atoll(argv[1]) : 1000000;
  std::mt19937_64 g(314);
  char buf[128];
  // 1. every float that is a rounding boundary neighbour: print random floats with 9 and 17 digits, and midpoints
  for (long long i = 0; i < n; ++i) {
    uint32_t bits = (uint32_t)g();
    if ((bits & 0x7F800000u) == 0x7F800000u) bits &= ~0x00800000u;  // finite
    float f;
    memcpy(&f, &bits, 4);
    snprintf(buf, sizeof buf, "%.9g", (double)f);
    check(buf);
    // exact midpoint between f and its successor, printed with many digits (hard round-to-even cases)
    uint32_t b2 = bits + 1;
    float f2;
    memcpy(&f2, &b2, 4);
    if (isfinite(f2)) {
      const double mid = ((double)f + (double)f2) / 2;
      snprintf(buf, sizeof buf, "%.19g", mid);
      check(buf);
      snprintf(buf, sizeof buf, "%.17g", nextafter(mid, 1e300));
      check(buf);
      snprintf(buf, sizeof buf, "%.17g", nextafter(mid, -1e300));
      check(buf);
    }
  }
  // 1b. random doubles at 17 and 16 digits, and Float.toString-like 9-digit values across the exponent range
  for (long long i = 0; i < n; ++i) {
    uint64_t b64 = g();
    if ((b64 & 0x7FF0000000000000ull) == 0x7FF0000000000000ull) b64 &= ~0x0010000000000000ull;
    double dv;
    memcpy(&dv, &b64, 8);
    snprintf(buf, sizeof buf, "%.17g", dv);
    check(buf);
    snprintf(buf, sizeof buf, "%.16g", dv);
    check(buf);
    snprintf(buf, sizeof buf, "%.15g", dv);
    check(buf);
  }
  // 2. random decimal strings: d digits, random point position, optional exponent
  for (long long i = 0; i < n; ++i) {
    std::string s;
    if (g() & 1) s += '-';
    const int nd = 1 + (int)(g() % 19);
    const int point = (int)(g() % (nd + 1));
    for (int k = 0; k < nd; ++k) {
      if (k == point) s += '.';
      s += (char)('0' + g() % 10);
    }
    const int mode = (int)(g() % 4);
    if (mode == 1) s += "E" + std::to_string((int)(g() % 12) - 6);
    else if (mode == 2) s += "e" + std::to_string((int)(g() % 100) - 50);
    else if (mode == 3) s += "E" + std::to_string((int)(g() % 700) - 350);
    check(s);
  }
  // 2b. short tokens that ARE binary32 rounding midpoints (odd integers above 2^24, x.5 above 2^23, …) and their neighbours:
  //     the binary64 short cut must hand exactly these to the Eisel-Lemire step
  for (long long i = 0; i < n; ++i) {
    const int e = 24 + (int)(g() % 10);                    // floats in [2^e, 2^(e+1)) are spaced 2^(e-23) apart
    const uint64_t step = 1ull << (e - 23);
    const uint64_t base = (1ull << e) + (g() % (1ull << 23)) * step;
    const uint64_t mid = base + step / 2;
    for (int dlt = -1; dlt <= 1; ++dlt) {
      snprintf(buf, sizeof buf, "%llu", (unsigned long long)(mid + dlt));
      check(buf);
      snprintf(buf, sizeof buf, "%llu.0", (unsigned long long)(mid + dlt));
      check(buf);
    }
    const uint64_t k = (1ull << 23) + (g() % (1ull << 23));  // 2^23 ≤ k < 2^24: spacing 1, k + 0.5 is a midpoint
    snprintf(buf, sizeof buf, "%llu.5", (unsigned long long)k);
    check(buf);
    snprintf(buf, sizeof buf, "%llu.25", (unsigned long long)(k >> 1));  // spacing 0.5 → .25 is a midpoint
    check(buf);
    snprintf(buf, sizeof buf, "%llu.5E-1", (unsigned long long)(k * 10 + g() % 10));
    check(buf);
    snprintf(buf, sizeof buf, "%llu.125", (unsigned long long)(k >> 2));
    check(buf);
  }
  // 3. long digit strings (> 19 significant digits): must either match or be refused, never be wrong
  for (long long i = 0; i < n / 10; ++i) {
    std::string s;
    const int nd = 20 + (int)(g() % 30);
    const int point = (int)(g() % (nd + 1));
    for (int k = 0; k < nd; ++k) {
      if (k == point) s += '.';
      s += (char)('0' + g() % 10);
    }
    check(s);
  }
  // 4. syntax corner cases
  const char* cases[] = {"0", "-0", "0.0", "-0.0", ".5", "5.", "+1.5", "1e5", "1E+5", "1e-5", " 1.5", "1.5 ", "\t2\t", "inf", "-inf",
                         "INF", "Infinity", "nan", "NaN", "-nan", "", " ", "-", "+", ".", "e5", "1e", "1e+", "1.5x", "x1.5", "1..5",
                         "1.5.2", "--1", "0x10", "1,5", "1 5", "3.4028235e38", "3.4028236e38", "3.4028235677973366e38", "1e39", "1e-46",
                         "1.4e-45", "7e-46", "7.0064923216240854e-46", "7.0064923216240853e-46", "1.17549435e-38", "1.17549421e-38",
                         "000000000000000000000001.5", "0.000000000000000000000000000000000000000000001", "123456789012345678901234567890",
                         "5.723834E-4", "1567606196", "0.034893095", "2493.487", "1e400", "1e-400", "0e999999999", "1e99999999999"};
  for (const char* c : cases) check(c);
  if (argc > 2) {
    FILE* fp = fopen(argv[2], "r");
    while (fp && fgets(buf, sizeof buf, fp)) {
      size_t l = strlen(buf);
      while (l && (buf[l - 1] == '\n' || buf[l - 1] == '\r')) buf[--l] = 0;
      check(buf);
    }
    if (fp) fclose(fp);
  }
  // shapes the fast path is meant for: d.dddd with 1..11 digits, optional E[-]d[d], optional sign
  for (long long i = 0; i < n; ++i) {
    std::string s;
    if (g() % 4 == 0) s += '-';
    const int nd = 1 + (int)(g() % 11);
    const int point = (int)(g() % (nd + 2)) - 1;  // -1: no dot
    for (int k = 0; k < nd; ++k) {
      if (k == point) s += '.';
      s += (char)('0' + g() % 10);
    }
    if (point == nd) s += '.';
    const int mode = (int)(g() % 6);
    if (mode == 1) s += "E" + std::to_string((int)(g() % 10));
    else if (mode == 2) s += "E-" + std::to_string((int)(g() % 10));
    else if (mode == 3) s += "E" + std::to_string(10 + (int)(g() % 30));
    else if (mode == 4) s += "e-" + std::to_string(10 + (int)(g() % 36));
    check(s);
  }
  printf("mismatches=%lld checked=%lld checked64=%lld fast_path_hits=%lld num16_hits=%lld\n", bad, checked, checked64, fast_hits, num16_hits);
  return bad ? 1 : 0;
}
