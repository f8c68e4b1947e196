mkdir -p gpurun_out
timeout 600 python tools/e2e_probe.py avro_a > gpurun_out/e2e_probe.txt 2>&1; echo "rc $?"; cat gpurun_out/e2e_probe.txt | tail -12
