#!/bin/bash
# decode attention with L2 evict-first streaming of the cross-attention memory cache: tests, then A/B of the greedy decoder bench.
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_kernels_gpu.py tests/test_decoder_gpu.py tests/test_greedy_parity_gpu.py -x -q -m gpu -k "decode or greedy or decoder" > gpurun_out/pytest_decode_evict.log 2>&1
echo "pytest rc=$?" | tee -a gpurun_out/pytest_decode_evict.log; tail -4 gpurun_out/pytest_decode_evict.log
for v in 0 1 0 1; do
  M3AE_DECODE_EVICT_FIRST=$v timeout 200 python bench.py --workload decoder --no-cpu-baseline --steps 30 --warmup 5 > gpurun_out/bench_decoder_evict$v.json 2> gpurun_out/bench_decoder_evict$v.err
  python - "$v" <<'PY'
import json, sys
d = json.loads(open(f"gpurun_out/bench_decoder_evict{sys.argv[1]}.json").read().strip().splitlines()[-1])
print("evict_first", sys.argv[1], d["value"], d["unit"], d["ms_per_step"], "roofline", d["roofline"]["frac"], "e2e", d["e2e"]["value"])
PY
done 2>&1 | tee gpurun_out/decode_evict_ab.txt
