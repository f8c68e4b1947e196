// Host build of csrc/float_format.cuh (the exact row formatter the CUDA kernel runs), for comparison against NumPy.
//   g++ -O2 -std=c++17 -o format_rows_host format_rows_host.cpp
//   ./format_rows_host in.bin out.bin [compat]      in: int32 n, int32 d, float32[n*d]   out: per row int32 len + bytes
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include <vector>

#include "../../hivemq-mqtt-tensorflow-kafka-realtime-iot-machine-learning-training-inference_b200/csrc/float_format.cuh"

int main(int argc, char** argv) {
  if (argc < 3) return 2;
  FILE* f = fopen(argv[1], "rb");
  if (!f) return 2;
  int32_t n = 0, d = 0;
  if (fread(&n, 4, 1, f) != 1 || fread(&d, 4, 1, f) != 1) return 2;
  std::vector<float> x((size_t)n * d);
  if (fread(x.data(), 4, x.size(), f) != x.size()) return 2;
  fclose(f);
  const int compat = argc > 3 ? atoi(argv[3]) : 0;
  FILE* o = fopen(argv[2], "wb");
  std::vector<uint8_t> buf(b2_format_row_bound(d) + 16);
  for (int i = 0; i < n; ++i) {
    const float* row = x.data() + (size_t)i * d;
    uint8_t* out = buf.data();
    const int32_t len = b2_format_row_f32(row, d, out, compat);
    if (len > b2_format_row_bound(d)) {
      fprintf(stderr, "row %d: length %d exceeds the bound %d\n", i, len, b2_format_row_bound(d));
      return 3;
    }
    fwrite(&len, 4, 1, o);
    fwrite(buf.data(), 1, len, o);
  }
  fclose(o);
  return 0;
}
