#!/bin/bash
# CUDA-graph classifier forward: its test (+ the other full-model tests), then the bench line (config 1 eager and graphed).
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_full_model_gpu.py -x -q -m gpu -s > gpurun_out/pytest_graphed.log 2>&1
echo "pytest rc=$?" | tee -a gpurun_out/pytest_graphed.log
tail -25 gpurun_out/pytest_graphed.log
timeout 600 python bench.py > gpurun_out/bench_r2_graph.json 2> gpurun_out/bench_r2_graph.err; echo "bench rc=$?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/bench_r2_graph.json").read().strip().splitlines()[-1])
print(json.dumps(d.get("classifier_forward_b8"))[:900])
print(d["value"], d["ms_per_step"], d["e2e"]["value"], d["whole_step"]["ms_per_step"], d["decoder"]["value"])
PY
tail -5 gpurun_out/bench_r2_graph.err
