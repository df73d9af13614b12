#!/bin/bash
# 1. the allocation-free entry point's test alone in a fresh process; 2. the whole GPU suite (-x) with it collected last;
# 3. if (2) stopped early: the suite again with the CLIP blocks on the composable nodes and the fused-block test deselected, so that
#    the no-alloc test is still seen after the other tests; 4. whole-model step, fused ViT blocks off / on; 5. bench line.
mkdir -p gpurun_out
python -m pytest tests/test_zz_no_alloc.py -q -m gpu > gpurun_out/noalloc_alone.log 2>&1
echo "alone rc=$?" | tee -a gpurun_out/noalloc_alone.log
tail -4 gpurun_out/noalloc_alone.log
timeout 500 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu_r2_noalloc.log 2>&1
rc=$?
echo "suite rc=$rc" | tee -a gpurun_out/pytest_gpu_r2_noalloc.log
tail -30 gpurun_out/pytest_gpu_r2_noalloc.log
if [ $rc -ne 0 ]; then
  M3AE_VIT_FUSED=0 timeout 500 python -m pytest tests -x -q -m gpu \
    --deselect "tests/test_dropin_gpu.py::test_vit_block_fused_node_matches_torch_block_and_composable_nodes" > gpurun_out/pytest_gpu_r2_noalloc_vit_off.log 2>&1
  echo "suite (ViT blocks composable) rc=$?" | tee -a gpurun_out/pytest_gpu_r2_noalloc_vit_off.log
  tail -5 gpurun_out/pytest_gpu_r2_noalloc_vit_off.log
fi
M3AE_VIT_FUSED=0 timeout 200 python tools/whole_step_probe.py > gpurun_out/whole_step_vit_composable.txt 2>&1; tail -6 gpurun_out/whole_step_vit_composable.txt
M3AE_VIT_FUSED=1 timeout 200 python tools/whole_step_probe.py > gpurun_out/whole_step_vit_fused.txt 2>&1; tail -6 gpurun_out/whole_step_vit_fused.txt
if [ $rc -eq 0 ]; then
  timeout 600 python bench.py > gpurun_out/bench_r2_vit.json 2> gpurun_out/bench_r2_vit.err; echo "bench rc=$?"; tail -c 3000 gpurun_out/bench_r2_vit.json
fi
