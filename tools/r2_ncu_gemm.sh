#!/bin/bash
# ncu --set full of the epilogue-heavy GEMM stages (one launch each after 3 warm-ups), image-stream shapes at batch 64
set -x
for only in "[i] fwd FFN-up N=3072 bias+gelu+aux=gelu'" "[i] fwd O-proj N=768 bias+drop+res" "[i] fwd FFN-up N=3072 plain" "[i] dgrad FFN-down (B MN) *aux" "[i] wgrad W1"; do
  tag=$(echo "$only" | tr -c 'A-Za-z0-9' '_' | cut -c1-40)
  ONLY="$only" REPS=1 ncu --set full --clock-control none --import-source on -k regex:gemm_tcgen05 -s 3 -c 1 -f -o gpurun_out/ncu_gemm_$tag python tools/gemm_bench.py > gpurun_out/ncu_gemm_$tag.log 2>&1
done
