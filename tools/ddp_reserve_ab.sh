# A/B at N=2 on one box: SMs the persistent GEMM leaves to NCCL (usage under gpurun --gpus 2: bash tools/ddp_reserve_ab.sh)
run() { tag=$1; shift; env "$@" timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $((29520 + RANDOM % 200)) bench.py --gpus 2 --steps 20 --warmup 5 --watchdog 150 --no-cpu-baseline 2>gpurun_out/ddp_res_$tag.err | python -c "import sys,json; d=json.loads([l for l in sys.stdin.read().splitlines() if l.startswith(chr(123))][-1]); print('$tag', d['ms_per_step'], d['value'], d['e2e']['value'], d['roofline']['achieved'], d['host_issue_ms_per_step'], d['grad_check']['rel_err_vs_whole_batch'])"; }
echo "# tag ms/step samples/s e2e in-step-GEMM-TF/s host-issue-ms grad_check"
timeout 200 python bench.py --steps 20 --warmup 5 --fusion-only --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('single_gpu', d['ms_per_step'], d['value'], d['e2e']['value'], d['roofline']['achieved'], d['host_issue_ms_per_step'])"
run reserve0 M3AE_DDP_RESERVE_SMS=0
run reserve4 M3AE_DDP_RESERVE_SMS=4
run reserve8 M3AE_DDP_RESERVE_SMS=8
run reserve16 M3AE_DDP_RESERVE_SMS=16
run reserve8_cta8 M3AE_DDP_RESERVE_SMS=8 NCCL_MAX_CTAS=8
run reserve0_again M3AE_DDP_RESERVE_SMS=0
