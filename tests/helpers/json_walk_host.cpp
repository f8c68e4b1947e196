// Host build of csrc/json_walk.cuh (the exact record walk the CUDA decode kernel runs), for comparison with Python's json.
//   g++ -O2 -std=c++17 -o json_walk_host json_walk_host.cpp
//   ./json_walk_host spec.bin records.bin out.bin
//   spec.bin:    int32 n_keys, int32 n_out, then per key: int8 out, uint8 len, char name[32]
//   records.bin: int32 n, int32 offsets[n+1], bytes
//   out.bin:     float32[n*n_out], int32 status[n], uint8 label[n]
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <vector>

#include "../../hivemq-mqtt-tensorflow-kafka-realtime-iot-machine-learning-training-inference_b200/csrc/json_walk.cuh"

int main(int argc, char** argv) {
  if (argc < 4) return 2;
  FILE* f = fopen(argv[1], "rb");
  if (!f) return 2;
  static JsonP p;
  memset(&p, 0, sizeof p);
  if (fread(&p.n_keys, 4, 1, f) != 1 || fread(&p.n_out, 4, 1, f) != 1) return 2;
  for (int k = 0; k < p.n_keys; ++k) {
    char name[B2_JSON_NAME_BYTES];
    if (fread(&p.out[k], 1, 1, f) != 1 || fread(&p.len[k], 1, 1, f) != 1 || fread(name, 1, B2_JSON_NAME_BYTES, f) != B2_JSON_NAME_BYTES) return 2;
    memcpy(p.name[k], name, B2_JSON_NAME_BYTES);
  }
  b2_json_finish_spec(p);
  fclose(f);
  f = fopen(argv[2], "rb");
  if (!f) return 2;
  int32_t n = 0;
  if (fread(&n, 4, 1, f) != 1) return 2;
  std::vector<int32_t> offs(n + 1);
  if (fread(offs.data(), 4, n + 1, f) != (size_t)(n + 1)) return 2;
  std::vector<uint8_t> data(offs[n] + 64, (uint8_t)'9');  // hostile padding behind the last record
  if (offs[n] && fread(data.data(), 1, offs[n], f) != (size_t)offs[n]) return 2;
  fclose(f);
  std::vector<float> out((size_t)n * p.n_out, 0.f);
  std::vector<int32_t> st(n);
  std::vector<uint8_t> lab(n, 2);
  for (int i = 0; i < n; ++i) st[i] = b2_json_record_auto(data.data(), offs[i], offs[i + 1], p, out.data() + (size_t)i * p.n_out, &lab[i]);
  f = fopen(argv[3], "wb");
  fwrite(out.data(), 4, out.size(), f);
  fwrite(st.data(), 4, n, f);
  fwrite(lab.data(), 1, n, f);
  fclose(f);
  return 0;
}
