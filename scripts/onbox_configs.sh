#!/bin/bash
# BASELINE.json configs 2-4 on an N-GPU box (scaled-down epochs so they finish in a gpurun call).
# usage: scripts/onbox_configs.sh NGPU
N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
COMMON="--fast --train.tensorboard=false --train.validate=false --client.log_root=gpurun_out/onbox_logs"
# config 2: N clients (1 per GPU) FedAvg, U-Net 256x256 bf16, 5 local epochs/round (epoch = 32 steps here)
$TR --master-port 29701 -m crackfl.cli.onbox --rounds 2 --clients-per-gpu 1 --strategy fedavg $COMMON \
   --train.image_size=256x256 --train.local_epochs=5 --data.synthetic_train=512 > gpurun_out/onbox_cfg2_n$N.log 2>&1
tail -2 gpurun_out/onbox_cfg2_n$N.log
# config 3: 8 clients per GPU, sample-count weighting, 256x256
$TR --master-port 29702 -m crackfl.cli.onbox --rounds 2 --clients-per-gpu 8 --strategy fedavg_weighted --payload bf16_delta $COMMON \
   --train.image_size=256x256 --train.local_epochs=1 --data.synthetic_train=128 > gpurun_out/onbox_cfg3_n$N.log 2>&1
tail -2 gpurun_out/onbox_cfg3_n$N.log
# config 4: FedAdam server optimiser, wide U-Net, 448x448
$TR --master-port 29703 -m crackfl.cli.onbox --rounds 2 --clients-per-gpu 1 --strategy fedadam $COMMON \
   --train.image_size=448x448 --train.width_mult=2 --train.local_epochs=1 --data.synthetic_train=128 > gpurun_out/onbox_cfg4_n$N.log 2>&1
tail -2 gpurun_out/onbox_cfg4_n$N.log
