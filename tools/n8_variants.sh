run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $1 bench.py --gpus 8 --steps 200 --warmup 10 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$2', d['value'], d['ms_per_step'], d['e2e']['ms_per_step'])"; }
NCCL_MAX_NCHANNELS=1 NCCL_MIN_NCHANNELS=1 run 29601 "pipe+1ch"
ORYX_BENCH_PIPELINE=0 ORYX_BENCH_RESERVED_SMS=0 run 29602 "nopipe"
ORYX_BENCH_RESERVED_SMS=4 run 29603 "pipe+res4"
ORYX_BENCH_RESERVED_SMS=4 NCCL_MAX_NCHANNELS=1 NCCL_MIN_NCHANNELS=1 run 29604 "pipe+res4+1ch"
