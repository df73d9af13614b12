#!/bin/bash
# Standard GPU-box sequence: parity tests, bench line, per-kernel microbenches.  Usage (under gpurun): bash tools/gpu_check.sh [tag]
tag=${1:-run}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -8 | tee gpurun_out/pytest_$tag.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err; tail -c 3000 gpurun_out/bench_$tag.json; tail -5 gpurun_out/bench_$tag.err
timeout 120 python tools/gemm_bench.py > gpurun_out/gemm_bench_$tag.txt 2>&1
timeout 120 python tools/attn_bench.py > gpurun_out/attn_bench_$tag.txt 2>&1; cat gpurun_out/attn_bench_$tag.txt
