// Host build of csrc/inflate.cuh (the byte code the CUDA inflate kernels run) as a shared library for ctypes:
//   g++ -O2 -std=c++17 -shared -fPIC -o libinflate_host.so inflate_host.cpp
#include "../../hivemq-mqtt-tensorflow-kafka-realtime-iot-machine-learning-training-inference_b200/csrc/inflate.cuh"

extern "C" long long inflate_codec(int codec, const uint8_t* src, long long n, uint8_t* dst, long long cap) {
  return b2_inflate_codec(codec, src, n, dst, cap);
}
