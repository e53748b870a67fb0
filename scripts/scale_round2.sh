#!/bin/bash
# Runs ON THE GPU BOX under `gpurun --gpus N`: the multi-GPU evidence of round 2 for N ranks (one per GPU, torchrun, NCCL).
#   cfg2 weak scaling (the contract workload), cfg4 weak scaling (BASELINE's designated DDP config), cfg2 strong scaling
#   (global batch 64 fixed), cfg3 sampling (batch-sharded, no collective), and -- at N = 2 -- the NCCL correctness tests.
# Outputs: gpurun_out/scale/<name>_n<N>.json (one bench JSON line each).
set -u
N=${1:-2}
O=gpurun_out/scale; mkdir -p $O
PORT=29611
run() {  # name, script, args...
  local name=$1; shift
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT "$@" \
      > $O/${name}_n$N.json 2> $O/${name}_n$N.err
  echo "$name n=$N rc=$? $(head -c 260 $O/${name}_n$N.json)"
  PORT=$((PORT + 1))
}
if [ "$N" = 2 ]; then
  python -m pytest tests/test_gpu_nccl.py -m gpu -q --timeout 600 > $O/nccl_tests.log 2>&1; echo "nccl tests rc=$? $(tail -1 $O/nccl_tests.log)"
  grep -h "nccl ok" $O/nccl_tests.log
fi
run cfg2_weak bench.py --gpus $N --steps 20 --warmup 5 --sampling-structures 0
run cfg4_weak bench.py --gpus $N --steps 10 --warmup 3 --workload cfg4 --sampling-structures 0
run cfg2_strong bench.py --gpus $N --steps 20 --warmup 5 --global-batch 64 --sampling-structures 0
run cfg3_sampling bench_sampling.py --gpus $N --structures 128
