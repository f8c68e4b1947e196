mkdir -p gpurun_out
timeout 600 python tools/e2e_probe.py avro_a > gpurun_out/e2e_probe2.txt 2>&1; echo "rc $?"; cat gpurun_out/e2e_probe2.txt | tail -9
timeout 600 python -m pytest tests/test_decode_gpu.py -q -m gpu -x -k "host_end_to_end" 2>&1 | tail -2
