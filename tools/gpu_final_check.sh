#!/bin/bash
# Round-end sequence on the final tree: smoke, the whole GPU suite, the default bench line, the reference arm (short).
mkdir -p gpurun_out
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_final.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/smoke_final.log; tail -3 gpurun_out/smoke_final.log
timeout 600 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu_final.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu_final.log; tail -4 gpurun_out/pytest_gpu_final.log
timeout 600 python bench.py > gpurun_out/bench_final.json 2> gpurun_out/bench_final.err; echo "bench rc=$?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/bench_final.json").read().strip().splitlines()[-1])
print("fusion", d["value"], d["ms_per_step"], "e2e", d["e2e"]["value"], "roofline", d["roofline"]["frac"], "launches", d["gpu_launches"], "clocks", d.get("clocks"))
print("decoder", d["decoder"]["value"], d["decoder"]["ms_per_step"], d["decoder"]["roofline"]["frac"], "e2e", d["decoder"]["e2e"]["value"])
print("whole", d["whole_step"]["ms_per_step"], "cfg1", d["classifier_forward_b8"]["ms_per_forward"], d["classifier_forward_b8"]["cuda_graph"], "t5", d["t5_head"]["ms_per_step"])
PY
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference_final.json 2> gpurun_out/bench_reference_final.err; echo "reference rc=$?"; tail -c 600 gpurun_out/bench_reference_final.json
