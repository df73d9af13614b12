#!/bin/bash
# ncu --set full of the tcgen05 attention kernels at the fusion shapes (image self, both cross directions), dropout on.
# Usage (under gpurun): bash tools/r2_ncu_attn.sh [tag]
tag=${1:-r2}
mkdir -p gpurun_out
export REPS=1 SHAPES=325x325,325x32,32x325 DROPS=0.1
python tools/attn_bench.py > gpurun_out/attn_plain_$tag.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'attn_(fwd|bwd)_tc' -c 12 -f -o gpurun_out/ncu_attn_$tag python tools/attn_bench.py > gpurun_out/ncu_attn_$tag.log 2>&1
echo "rc=$?"; tail -3 gpurun_out/ncu_attn_$tag.log
REPS=20 DROPS=0.0,0.1 SHAPES=325x325,325x32,32x325,32x32 python tools/attn_bench.py | tee gpurun_out/attn_bench_$tag.txt
