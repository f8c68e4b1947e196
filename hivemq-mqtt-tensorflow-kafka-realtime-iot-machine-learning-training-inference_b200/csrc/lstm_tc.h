// Interface between lstm.cu (handle, FP32 kernel) and lstm_tc.cu (tcgen05 kernel).
#pragma once
#include <vector>

#include "common.cuh"

struct TcLayer {
  int N, K;   // GEMM shape (M = 128)
  int a_col;  // first TMEM column of this layer's A panel
  int w_off;  // float offset of B_hi inside the packed weight block; B_lo follows at w_off + N*K
};

// returns 0 and fills `packed` / `layers` when the stack fits the tensor-core kernel (features ≤ 18, units 32-16-16-32)
int b2_lstm_tc_pack(const b2_lstm_desc* desc, const float* host_weights, std::vector<float>& packed, TcLayer (&layers)[5]);
int b2_lstm_tc_launch(b2_ctx* ctx, const float* d_packed, int packed_floats, const TcLayer (&layers)[5], int F, const float* x, int64_t batch,
                      int T, int64_t seq_stride, float* y, cudaStream_t stream);
