set -x
mkdir -p gpurun_out
timeout 900 python tools/mma_probe.py avro_a avro_b csv 1 > gpurun_out/mma_probe5.txt 2>&1; echo "probe rc $?"
grep "records/s" gpurun_out/mma_probe5.txt
timeout 1500 python -m pytest tests -q -m gpu -x 2>&1 | tail -5
