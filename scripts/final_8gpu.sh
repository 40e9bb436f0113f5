#!/bin/bash
# Round-end evidence on an N-GPU box in one gpurun call: bench.py, BASELINE.json configs 2-4 through the on-box CLI, and the
# reference's own schedule (5 rounds x 10 epochs x 388 steps at 128^2, validation every epoch) with one client per GPU.
N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29615 bench.py --gpus $N --steps 20 --warmup 5 2>/dev/null | tail -1 | tee gpurun_out/bench_${N}gpu.json
timeout 600 bash scripts/onbox_configs.sh $N
timeout 400 $TR --master-port 29704 -m crackfl.cli.onbox --rounds 5 --eval-global --fast --train.tensorboard=false \
   --client.log_root=gpurun_out/onbox_logs > gpurun_out/onbox_ref_schedule_5x10_n$N.log 2>&1
tail -1 gpurun_out/onbox_ref_schedule_5x10_n$N.log
rm -rf gpurun_out/onbox_logs
