# Bucket-size sweep of the gradient exchange at N=2 (usage under gpurun --gpus 2: bash tools/ddp_bucket_ab.sh)
run() { tag=$1; shift; env "$@" timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $((29520 + RANDOM % 200)) bench.py --gpus 2 --steps 20 --warmup 5 --watchdog 100 --no-cpu-baseline 2>gpurun_out/ddp_ab_$tag.err | python -c "import sys,json; d=json.loads([l for l in sys.stdin.read().splitlines() if l.startswith(chr(123))][-1]); print('$tag', d['ms_per_step'], d['value'], d['e2e']['value'], d['roofline']['achieved'])"; }
run bucket16 M3AE_DDP_BUCKET_MB=16
run bucket32 M3AE_DDP_BUCKET_MB=32
run bucket64 M3AE_DDP_BUCKET_MB=64
run bucket128 M3AE_DDP_BUCKET_MB=128
