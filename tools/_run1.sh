mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_decode_gpu.py tests/test_kafka_gpu.py tests/test_partition_gpu.py tests/test_oversize_gpu.py -q -m gpu -x 2>&1 | tail -5
timeout 900 python tools/mma_probe.py avro_a avro_b csv 1 > gpurun_out/mma_probe6.txt 2>&1; echo "probe rc $?"
grep "decode_mma=0.*records/s" gpurun_out/mma_probe6.txt
