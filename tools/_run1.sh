mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_score_gpu.py tests/test_decode_gpu.py tests/test_normalize_gpu.py tests/test_train_gpu.py -q -m gpu -x 2>&1 | tail -4
timeout 900 python tools/mma_probe.py avro_b 1 > gpurun_out/mma_probe7.txt 2>&1; echo "probe rc $?"
grep "decode_mma=0.*records/s" gpurun_out/mma_probe7.txt
timeout 600 python bench.py --workload score18 --steps 5 --warmup 3 --no-cpu-baseline --no-e2e --also none 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('score18', d['value'], d['roofline']['frac'])"
