// K5 — per-GPU partition driver.
//
// The reference runs one consumer process per Kafka partition ("<topic>:0:<offset>", cardata-v1.py:10; the topic has 10
// partitions, infrastructure/confluent/01_installConfluentPlatform.sh:180) and scores batch after batch inside
// Model.predict.  Here one process per GPU owns a set of partitions (p mod world_size): their device-resident record
// batches are pushed through the fused decode → normalise → score kernel round-robin over a few CUDA streams so that the
// tail of one partition overlaps the head of the next, and the per-GPU counters {records ok, flagged, decode errors,
// payload bytes} accumulate on the device.  The only cross-GPU exchange of the whole path — summing those four counters
// — is left to the caller, which owns the NCCL communicator (torch.distributed).
#include "common.cuh"

extern "C" int b2_partition_run(b2_ae* ae, const b2_partition* parts, int n_parts, const b2_csv_spec* csv, const b2_avro_prog* avro,
                                int skip_prefix, const b2_norm* norm, float threshold, int n_streams, unsigned long long* counters_dev,
                                b2_run_stats* host_out) {
  return b2_partition_run_repeat(ae, parts, n_parts, csv, avro, skip_prefix, norm, threshold, n_streams, 1, counters_dev, host_out);
}

extern "C" int b2_partition_run_repeat(b2_ae* ae, const b2_partition* parts, int n_parts, const b2_csv_spec* csv, const b2_avro_prog* avro,
                                       int skip_prefix, const b2_norm* norm, float threshold, int n_streams, int repeat,
                                       unsigned long long* counters_dev, b2_run_stats* host_out) {
  B2_REQUIRE(repeat >= 1, "b2_partition_run_repeat: repeat=%d < 1", repeat);
  B2_REQUIRE(ae != nullptr, "b2_partition_run: ae is NULL");
  B2_REQUIRE(n_parts >= 0 && (n_parts == 0 || parts != nullptr), "b2_partition_run: bad partition list");
  B2_REQUIRE(counters_dev != nullptr, "b2_partition_run: counters_dev is NULL");
  B2_REQUIRE(n_streams >= 1 && n_streams <= 16, "b2_partition_run: n_streams=%d not in 1..16", n_streams);
  DeviceGuard g(ae->ctx->device);
  // streams and events live on the context: created the first time they are needed, reused by every later run, released by
  // b2_ctx_destroy — a run neither creates nor destroys anything, and an early error return leaks nothing
  b2_ctx* ctx = ae->ctx;
  if (!ctx->run_start) B2_CUDA(cudaEventCreate(&ctx->run_start));
  if (!ctx->run_stop) B2_CUDA(cudaEventCreate(&ctx->run_stop));
  while (ctx->n_run_streams < n_streams) {
    const int i = ctx->n_run_streams;
    B2_CUDA(cudaStreamCreateWithFlags(&ctx->run_streams[i], cudaStreamNonBlocking));
    cudaError_t e = cudaEventCreateWithFlags(&ctx->run_done[i], cudaEventDisableTiming);
    if (e != cudaSuccess) {
      cudaStreamDestroy(ctx->run_streams[i]);
      b2_set_error("b2_partition_run: %s", cudaGetErrorString(e));
      return B2_ECUDA;
    }
    ctx->n_run_streams = i + 1;
  }
  cudaStream_t* streams = ctx->run_streams;
  cudaEvent_t ev_start = ctx->run_start, ev_stop = ctx->run_stop, *ev_done = ctx->run_done;
  int rc = B2_OK;
  B2_CUDA(cudaEventRecord(ev_start, streams[0]));
  for (int i = 1; i < n_streams; ++i) B2_CUDA(cudaStreamWaitEvent(streams[i], ev_start, 0));
  for (int rep = 0; rep < repeat && rc == B2_OK; ++rep)
    for (int p = 0; p < n_parts && rc == B2_OK; ++p) {
      const b2_partition& pt = parts[p];
      if (pt.n == 0) continue;
      rc = b2_decode_score(ae, pt.bytes_dev, pt.offs_dev, pt.n, csv, avro, skip_prefix, norm, threshold, pt.mse_dev, pt.flag_dev, nullptr,
                           nullptr, counters_dev, streams[p % n_streams]);
    }
  for (int i = 1; i < n_streams; ++i) {
    cudaEventRecord(ev_done[i], streams[i]);
    cudaStreamWaitEvent(streams[0], ev_done[i], 0);
  }
  cudaEventRecord(ev_stop, streams[0]);
  cudaError_t e = cudaStreamSynchronize(streams[0]);  // also on the error path: nothing of this run is left in flight
  float ms = 0.f;
  if (e == cudaSuccess) cudaEventElapsedTime(&ms, ev_start, ev_stop);
  unsigned long long h[4] = {0, 0, 0, 0};
  if (e == cudaSuccess) e = cudaMemcpy(h, counters_dev, sizeof(h), cudaMemcpyDeviceToHost);
  if (rc != B2_OK) return rc;
  if (e != cudaSuccess) {
    b2_set_error("b2_partition_run: %s", cudaGetErrorString(e));
    return B2_ECUDA;
  }
  if (host_out) {
    host_out->records = h[0];
    host_out->flagged = h[1];
    host_out->decode_errors = h[2];
    host_out->bytes = h[3];
    host_out->elapsed_ms = ms;
  }
  return B2_OK;
}
