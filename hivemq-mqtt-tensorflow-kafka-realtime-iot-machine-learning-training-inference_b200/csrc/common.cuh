// Shared host/device helpers for the b2 sensor-scoring library (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <vector>

#include "../../include/b2_api.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "this library is written for sm_100a (B200) only"
#endif

// ---- error plumbing ---------------------------------------------------------------------------------------
void b2_set_error(const char* fmt, ...);

#define B2_CUDA(call)                                                                              \
  do {                                                                                             \
    cudaError_t e__ = (call);                                                                      \
    if (e__ != cudaSuccess) {                                                                      \
      b2_set_error("%s:%d: %s failed: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__));    \
      return B2_ECUDA;                                                                             \
    }                                                                                              \
  } while (0)

#define B2_REQUIRE(cond, ...)        \
  do {                               \
    if (!(cond)) {                   \
      b2_set_error(__VA_ARGS__);     \
      return B2_EINVAL;              \
    }                                \
  } while (0)

// b2_norm.order = B2_NORM_ORDER_F64 needs float64 values: only the fused Avro decode has them (decode.cu)
#define B2_NORM_F32_ONLY(norm, who)                                                                                        \
  B2_REQUIRE(!(norm) || (norm)->n <= 0 || (norm)->order == B2_NORM_ORDER_F32,                                              \
             "%s: the float64 normalisation order (cardata-v3.py) needs float64 values; only b2_decode_score / "             \
             "b2_decode_score_spans with an Avro program take it", who)

// Tuning / debugging knobs.  Read ONCE from the environment (B2_FAST_TANH, B2_SCORE_CFG, …) when the context is created — a
// launch never calls getenv — and changed afterwards with b2_ctx_set_option().
struct b2_knobs {
  int fast_tanh = 1;      // B2_FAST_TANH     1: exp2-based tanh (7 instr / pair), 0: the ≤ 4-ulp variant
  int score_cfg = 1;      // B2_SCORE_CFG     ae_score launch shape: 0 = 128 thr x 2 stages, 1 = 128 x 1, 2 = 256 x 1
  int no_bulk = 0;        // B2_NO_BULK       1: plain loads instead of bulk async copies in ae_score
  int force_generic = 0;  // B2_FORCE_GENERIC 1: generic Dense-stack kernel even for D-14-7-7-D
  int avro_generic = 0;   // B2_AVRO_GENERIC  1: interpreted Avro walk even for the two known schemas
  int lstm_tc = 2;        // B2_LSTM_TC       LSTM forward: 2 tcgen05 kernel, two tiles per CTA when the batch is large (default), 3 always two, 1 one tile, 0 FP32 FMA kernel
  int score_tc = -1;      // B2_SCORE_TC      pre-decoded scoring: -1 auto (tcgen05 kernel K1t for D = 30, FFMA2 for D = 18), 0 FFMA2, 1 K1t, 2 K1t with all four layers on tcgen05
  int decode_r2 = 0;      // B2_DECODE_R2     fused Avro decode: 0 one record per lane (default), 2 / 3 two (85 / 102 registers)
  int decode_tc = 0;      // B2_DECODE_TC     fused decode + score: 1 = score phase on tcgen05 (decode_kernel_tc)
  int decode_mma = 0;     // B2_DECODE_MMA    fused decode + score: 1 = score phase as warp-level TF32 MMAs (decode_kernel_wm, ae_mma.cuh)
  int zero_skip = 1;      // B2_ZERO_SKIP     fused decode + score: 1 = the instantiation that leaves out the features the reference's table forces to zero (when the table matches), 0 = always the generic one
  int tc_ctas = 0;        // B2_TC_CTAS       CTAs per SM of the tcgen05 score kernels (0 = as many as TMEM allows)
};

constexpr int B2_MAX_STREAM_SLOTS = 32;
constexpr int B2_OVERSIZE_CAP = 1024;  // oversize records one decode launch can hand to its retry pass

struct b2_ctx {
  int device;
  int sm_count;
  int smem_optin;  // max dynamic shared memory per block (bytes)
  int record_bytes_hint = 0;  // b2_ctx_set_record_bytes_hint: average record size of the batches to come, 0 = unknown
  b2_knobs knobs;
  // b2_partition_run: streams and events created on first use and kept for the life of the context
  int n_run_streams = 0;
  cudaStream_t run_streams[16] = {};
  cudaEvent_t run_done[16] = {};
  cudaEvent_t run_start = nullptr, run_stop = nullptr;
  // decode retry lists: one slot per stream that has launched a decode on this context (same-stream launches are ordered, so
  // a slot is never shared by two launches in flight).  Slot = int32[1 + B2_OVERSIZE_CAP] in one device allocation made at
  // context creation.
  int32_t* oversize_pool = nullptr;
  int n_stream_slots = 0;
  cudaStream_t slot_stream[B2_MAX_STREAM_SLOTS] = {};
};
int32_t* b2_ctx_oversize_slot(b2_ctx* ctx, cudaStream_t stream);  // never NULL once the pool exists; shares the last slot when full

// RAII "make this device current for the call" guard (callers may sit on another device)
struct DeviceGuard {
  int prev = -1;
  explicit DeviceGuard(int dev) {
    cudaGetDevice(&prev);
    if (prev != dev) cudaSetDevice(dev);
    else prev = -1;
  }
  ~DeviceGuard() {
    if (prev >= 0) cudaSetDevice(prev);
  }
};

// ---- the Dense autoencoder handle (shared by score / decode / train translation units) ---------------------
struct b2_ae {
  b2_ctx* ctx;
  b2_ae_desc desc;
  int n_params;
  int spec_d;                  // 18 / 30 when the stack is D-14-7-7-D tanh/relu/tanh/relu (fast kernels), else 0
  std::vector<float> h_w;      // Keras flat order: kernel0, bias0, kernel1, bias1, ...
  float* d_w;                  // device copy, same order
  float* d_m;                  // Adam slots (allocated on first fit)
  float* d_v;
  bool host_stale;             // device weights changed by fit; h_w must be refreshed before param-passing
  float* d_tc = nullptr;       // tensor-core path: packed (hi, lo) weight panels (ae_tc.cuh), built on first use
  bool tc_valid = false;       // … and rebuilt after every change of the weights
};
int b2_ae_tc_panels(b2_ae* ae, const float** out);  // device pointer to the current panels (packs + uploads when stale)
int b2_ae_sync_host(b2_ae* ae);    // refresh h_w from d_w if stale
int b2_ae_ensure_adam(b2_ae* ae);  // allocate + zero the Adam slots on first use

// ---- device-side PTX wrappers (mbarrier + 1-D bulk async copy = the TMA engine without a tensor map) -------
#ifdef __CUDACC__
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  while (!mbar_try_wait(bar, parity)) {
  }
}
// global → shared bulk copy; completion is signalled on `bar` (complete_tx). src/dst 16-byte aligned, bytes % 16 == 0.
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(dst_smem)),
               "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
// shared → global bulk copy (bulk async-group completion)
__device__ __forceinline__ void bulk_s2g(void* dst_gmem, const void* src_smem, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst_gmem), "r"(smem_u32(src_smem)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read_all() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
#endif
