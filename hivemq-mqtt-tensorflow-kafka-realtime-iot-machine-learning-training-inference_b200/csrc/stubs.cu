// Entry points not implemented yet in this build (each is replaced by its own translation unit as it lands).
#include "common.cuh"
#define STUB(name) b2_set_error(name ": not implemented in this build"); return B2_EUNSUPPORTED
extern "C" int b2_lstm_create(b2_ctx*, const b2_lstm_desc*, const float*, b2_lstm**) { STUB("b2_lstm_create"); }
extern "C" void b2_lstm_destroy(b2_lstm*) {}
extern "C" int b2_lstm_param_count(const b2_lstm*) { return 0; }
extern "C" int b2_lstm_forward(b2_lstm*, const float*, int64_t, int, float*, void*) { STUB("b2_lstm_forward"); }
extern "C" int b2_partition_run(b2_ae*, const b2_partition*, int, const b2_csv_spec*, const b2_avro_prog*, int, const b2_norm*, float, int, unsigned long long*, b2_run_stats*) { STUB("b2_partition_run"); }
