// Integer tokens ("26", "-7", "00012", "99999999", nine digits for the general path) through the CSV field parser against glibc strtof:
// the at-most-8-digit shortcut of b2_num16 (one integer -> binary32 conversion) must give the same bits.  Tokens end at a delimiter or at
// the end of the record with arbitrary bytes behind it.
//   g++ -O2 -std=c++17 -o int_token_fuzz int_token_fuzz.cpp && ./int_token_fuzz 1000000
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cstdint>
#include <random>
#include "../../hivemq-mqtt-tensorflow-kafka-realtime-iot-machine-learning-training-inference_b200/csrc/float_parse.cuh"
int main(int argc, char** argv) {
  std::mt19937_64 g(7);
  long bad = 0, n = 0;
  for (int it = 0; it < (argc > 1 ? atol(argv[1]) : 1000000); ++it) {
    char tok[32];
    int nd = 1 + g() % 9;  // up to 9 digits: 9 takes the general path
    int p = 0;
    if (g() % 4 == 0) tok[p++] = '-';
    for (int i = 0; i < nd; ++i) tok[p++] = '0' + ((i == 0 && g() % 3) ? 1 + g() % 9 : g() % 10);
    int L = p;
    char buf[40];
    memset(buf, ',', sizeof buf);
    int sh = g() % 4;
    memcpy(buf + sh, tok, L);
    // terminator: ',' or end of record
    bool at_end = g() % 2;
    uint32_t w[6];
    memcpy(w, buf, 24);
    if (at_end) { for (int i = sh + L; i < 24; ++i) ((char*)w)[i] = (char)(g() % 256); }
    uint32_t out = 0;
    int r = b2_csv_num16(w[0], w[1], w[2], w[3], w[4], sh * 8, at_end ? L : L + 5, ',', &out);
    tok[L] = 0;
    float ref = strtof(tok, nullptr);
    uint32_t rb; memcpy(&rb, &ref, 4);
    ++n;
    if (r != L || rb != out) { if (bad < 10) printf("MISMATCH tok=%s r=%d out=%08x ref=%08x\n", tok, r, out, rb); ++bad; }
  }
  printf("%ld tokens, %ld mismatches\n", n, bad);
  return bad != 0;
}
